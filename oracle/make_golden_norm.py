"""Golden outputs and autograd gradients of the UNMODIFIED reference's InstanceNorm++ modules
(ncsnv2/models/normalization.py InstanceNorm2dPlus; ncsn/models/cond_refinenet_dilated.py ConditionalInstanceNorm2dPlus),
alone and followed by F.elu -> tests/golden/inorm_plus.npz.  Runs in the build container only (/root/reference is not on
the GPU box).  TEST INFRASTRUCTURE.
    python -m oracle.make_golden_norm"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

from .make_golden import REF

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "inorm_plus.npz")
CASES = {            # name: (N, C, H, W, conditional, bias)
    "plain_8x8": (3, 6, 8, 8, False, True),
    "plain_nobias_5x7": (2, 5, 5, 7, False, False),
    "plain_16x16": (2, 16, 16, 16, False, True),
    "cond_8x8": (4, 6, 8, 8, True, True),
    "cond_nobias_4x4": (3, 8, 4, 4, True, False),
}


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref_plain = _load(f"{REF}/ncsnv2/models/normalization.py", "ref_ncsnv2_normalization")
    # the ncsn file imports its sibling packages at module level; only the norm class is needed
    src = open(f"{REF}/ncsn/models/cond_refinenet_dilated.py").read()
    start = src.index("class ConditionalInstanceNorm2dPlus")
    end = src.index("class CondRefineNetDilated")
    ref_cond = types.ModuleType("ref_ncsn_cond_norm")
    exec(compile("import torch\nimport torch.nn as nn\n" + src[start:end], f"{REF}/ncsn/models/cond_refinenet_dilated.py", "exec"),
         ref_cond.__dict__)
    out = {}
    for name, (N, C, H, W, cond, bias) in CASES.items():
        rs = np.random.RandomState(sum(map(ord, name)))
        x = rs.randn(N, C, H, W) * rs.uniform(0.5, 2.0, size=(1, C, 1, 1)) + rs.randn(N, C, 1, 1)
        gout = rs.randn(N, C, H, W)
        nrow = 7 if cond else 1
        table = np.concatenate([rs.normal(1, 0.2, (nrow, 2 * C)), rs.normal(0, 0.3, (nrow, C))], axis=1)
        labels = rs.randint(0, nrow, size=N)
        for act in (False, True):
            for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
                xt = torch.tensor(x, dtype=dt, requires_grad=True)
                if cond:
                    mod = ref_cond.ConditionalInstanceNorm2dPlus(C, nrow, bias=bias).to(dt)
                    with torch.no_grad():
                        mod.embed.weight.copy_(torch.tensor(table[:, :(3 if bias else 2) * C], dtype=dt))
                    emb = mod.embed.weight
                    y = mod(xt, torch.tensor(labels))
                else:
                    mod = ref_plain.InstanceNorm2dPlus(C, bias=bias).to(dt)
                    with torch.no_grad():
                        mod.gamma.copy_(torch.tensor(table[0, :C], dtype=dt))
                        mod.alpha.copy_(torch.tensor(table[0, C:2 * C], dtype=dt))
                        if bias:
                            mod.beta.copy_(torch.tensor(table[0, 2 * C:], dtype=dt))
                    y = mod(xt)
                o = F.elu(y) if act else y
                o.backward(torch.tensor(gout, dtype=dt))
                k = f"{name}/{'elu' if act else 'id'}/{tag}/"
                out[k + "out"] = o.detach().numpy()
                out[k + "dx"] = xt.grad.numpy()
                if cond:
                    out[k + "dtable"] = emb.grad.numpy()
                else:
                    out[k + "dgamma"] = mod.gamma.grad.numpy()
                    out[k + "dalpha"] = mod.alpha.grad.numpy()
                    if bias:
                        out[k + "dbeta"] = mod.beta.grad.numpy()
        out[f"{name}/x"], out[f"{name}/gout"], out[f"{name}/table"], out[f"{name}/labels"] = x, gout, table, labels
    np.savez_compressed(OUT, **out)
    print(f"{OUT}: {len(out)} arrays, {os.path.getsize(OUT)} bytes")


if __name__ == "__main__":
    sys.exit(main())
