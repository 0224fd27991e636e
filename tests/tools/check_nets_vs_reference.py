"""Build-container check (needs /root/reference): the three conv networks of fdsm_b200.nets against the reference's
modules with ONE state dict loaded into both.  Run once per sub-repo (their top-level packages collide):
    python tests/tools/check_nets_vs_reference.py ncsnv2|ncsn|deepebm"""
import importlib.util
import os
import sys
import types
from argparse import Namespace as NS

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fdsm_b200 import nets  # noqa: E402

REF = "/root/reference"


def compare(ref, mine, *inputs):
    sd = ref.state_dict()
    diff = set(sd) ^ set(mine.state_dict())
    assert not diff, sorted(diff)[:8]
    mine.load_state_dict(sd)
    a, b = ref(*inputs), mine(*inputs)
    err = float((a - b).abs().max() / a.abs().max())
    print(f"  {type(mine).__name__}: {len(sd)} tensors, {sum(p.numel() for p in mine.parameters())} params, rel max diff {err:.2e}")
    assert err < 1e-5
    return a


def ref_model(which, ngf=8, n_classes=12):
    if which == "ncsnv2":
        sys.path.insert(0, f"{REF}/ncsnv2")
        from models.ncsnv2 import NCSNv2 as R
        cfg = NS(device="cpu", data=NS(logit_transform=False, rescaled=False, channels=3, image_size=32),
                 model=NS(ngf=ngf, num_classes=n_classes, nonlinearity="elu", normalization="InstanceNorm++",
                          sigma_begin=50.0, sigma_end=0.01, sigma_dist="geometric"))
        return R(cfg)
    if which == "ncsn":
        sys.path.insert(0, f"{REF}/ncsn")
        from models.cond_refinenet_dilated import CondRefineNetDilated as R
        cfg = NS(data=NS(logit_transform=False, channels=3, image_size=32), model=NS(ngf=ngf, num_classes=n_classes))
        return R(cfg)
    flags = types.ModuleType("Utils.flags")
    flags.FLAGS = NS(batch_size=4, esm_eps=0.05, loss_coe=1e-4)
    sys.modules["Utils"], sys.modules["Utils.flags"], sys.modules["Torture"] = types.ModuleType("Utils"), flags, types.ModuleType("Torture")
    spec = importlib.util.spec_from_file_location("ref_models", f"{REF}/DeepEBM/ESM/models.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.Res18_Quadratic(1, ngf, 32, normalize=False, AF=torch.nn.ELU())


if __name__ == "__main__":
    which = sys.argv[1]
    torch.manual_seed(0)
    x, y = torch.rand(3, 3, 32, 32), torch.tensor([0, 5, 11])
    if which == "ncsnv2":
        compare(ref_model(which), nets.NCSNv2(3, 8, 12), x, y)
    elif which == "ncsn":
        compare(ref_model(which), nets.CondRefineNetDilated(3, 8, 12), x, y)
    else:
        compare(ref_model(which), nets.Res18Quadratic(1, 8, 32), torch.rand(3, 1, 32, 32))
