"""Generate tests/golden/*.npz by executing the UNMODIFIED reference functions from
/root/reference (build container only -- the GPU box has no /root/reference).

    python -m oracle.make_golden

The reference draws its randomness with torch.randn_like / torch.randint; we replay
the case's numpy-seeded draws through those two entry points in the documented RNG
order, so reference, oracle and CUDA path see identical numbers.
TEST INFRASTRUCTURE.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

from . import cases, run_case

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _load(path, name, extra_path=None, stubs=None):
    saved = {k: sys.modules.get(k) for k in (stubs or {})}
    for k, v in (stubs or {}).items():
        sys.modules[k] = v
    if extra_path:
        sys.path.insert(0, extra_path)
    try:
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if extra_path:
            sys.path.remove(extra_path)
            for k in [k for k in sys.modules if k == "models" or k.startswith("models.") or k == "losses"
                      or k.startswith("losses.")]:
                del sys.modules[k]
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod


class _Flags:
    batch_size = 0
    esm_eps = 0.05
    loss_coe = 1e-4


def load_reference():
    flags_mod = types.ModuleType("Utils.flags")
    flags_mod.FLAGS = _Flags()
    utils_mod = types.ModuleType("Utils")
    utils_mod.flags = flags_mod
    stubs = {"Utils": utils_mod, "Utils.flags": flags_mod, "Torture": types.ModuleType("Torture")}
    ref = types.SimpleNamespace()
    ref.FLAGS = flags_mod.FLAGS
    ref.esm = _load(f"{REF}/SSM/losses/efficient_sm.py", "ref_ssm_efficient_sm")
    ref.sliced = _load(f"{REF}/SSM/losses/sliced_sm.py", "ref_ssm_sliced_sm")
    ref.score_matching = _load(f"{REF}/SSM/losses/score_matching.py", "ref_ssm_score_matching")
    ref.ssm_dsm = _load(f"{REF}/SSM/losses/dsm.py", "ref_ssm_dsm", extra_path=f"{REF}/SSM")
    ref.deepebm = _load(f"{REF}/DeepEBM/ESM/loss_train.py", "ref_deepebm_loss_train", stubs=stubs)
    ref.deepebm_models = _load(f"{REF}/DeepEBM/ESM/models.py", "ref_deepebm_models", stubs=stubs)
    ref.ncsn = _load(f"{REF}/ncsn/losses/sliced_sm.py", "ref_ncsn_sliced_sm")
    ref.ncsnv2 = _load(f"{REF}/ncsnv2/losses/dsm.py", "ref_ncsnv2_dsm")
    return ref


class Replay:
    """Patch torch.randn_like / torch.randint to pop pre-drawn tensors."""

    def __init__(self, normals, ints=()):
        self.normals, self.ints = list(normals), list(ints)

    def __enter__(self):
        self._rl, self._ri = torch.randn_like, torch.randint

        def randn_like(t, *a, **k):
            d = self.normals.pop(0)
            assert d.shape == t.shape, (d.shape, t.shape)
            return d.to(t.dtype).clone()

        def randint(*a, **k):
            return self.ints.pop(0).clone()

        torch.randn_like, torch.randint = randn_like, randint
        return self

    def __exit__(self, *exc):
        torch.randn_like, torch.randint = self._rl, self._ri
        assert not self.normals and not self.ints, "unused draws: RNG order mismatch"


def reference_loss(ref, c):
    fn, net, x, eps = c["fn"], c["net"], c["x"].clone(), c["eps"]
    u, n = c["u"], c["n"]
    ref.FLAGS.batch_size, ref.FLAGS.esm_eps = c["B"], eps
    ref.FLAGS.loss_coe = c.get("loss_coe", 1e-4)
    if fn == "a1":   # exercise both the drawn path and (below) the noise= path
        with Replay([u]):
            return ref.esm.efficient_score_matching_conjugate(net, x, eps=eps)
    if fn == "a7":
        with Replay([u.reshape(c["B"], -1)]):
            return ref.deepebm.ssm_fd(net, x)
    if fn == "a8":
        with Replay([u.reshape(c["B"], -1)]):
            return ref.deepebm.ssm_fd_nop(net, x)
    if fn == "a2":
        with Replay([u]):
            return ref.esm.single_efficient_score_matching(net, x, eps=eps, noise_type=c.get("noise_type", "sphere"))
    if fn == "a5":
        with Replay([u]):
            return ref.esm.MLE_efficient_score_matching_conjugate(net, x, eps=eps, mle_ratio=c["mle_ratio"])
    if fn == "a3":
        with Replay([n, u]):
            return ref.esm.dsm_fd(net, x, sigma=c["sigma"], eps=eps)
    if fn == "a6":
        with Replay([n, u]):
            return ref.ssm_dsm.dsm_tracetrick_FD(net, x, sigma=c["sigma"], eps=eps)
    if fn == "a4":
        with Replay([u]):
            return ref.esm.esm_scorenet_VR(net, x, eps=eps)
    if fn in ("a9", "a10"):
        f = ref.deepebm.mdsm_fd if fn == "a9" else ref.deepebm.mdsm_fd_nop
        with Replay([n, u]):
            return f(net, x, c["sig_b"], c["sigma02"], c["D"])
    if fn == "a11":
        with Replay([n, u]):
            return ref.ncsn.anneal_ESM_scorenet_VR(net, x, c["labels"], c["sigmas"], 1, eps)
    if fn == "a12":
        with Replay([n, u]):
            return ref.ncsn.anneal_ESM_scorenet(net, x, c["labels"], c["sigmas"], 1)
    if fn == "a13":   # labels=None -> drawn inside with torch.randint (ncsnv2/losses/dsm.py:88-91)
        with Replay([n, u], [c["labels"]]):
            return ref.ncsnv2.anneal_ESM_scorenet_VR(net, x, c["sigmas"], None, c["anneal_power"], None, 1, eps)
    if fn == "a17_anneal_vr":   # RNG order: randint labels, noise, projections (ncsnv2/losses/dsm.py:36-53)
        with Replay([n, u], [c["labels"]]):
            return ref.ncsnv2.anneal_sliced_score_estimation_vr(net, x, c["sigmas"], None, c["anneal_power"], None,
                                                                c["n_particles"])
    if fn == "a17_exact":   # the reference accumulates into torch.zeros(B) of the DEFAULT dtype (score_matching.py:52)
        saved = torch.get_default_dtype()
        torch.set_default_dtype(x.dtype)
        try:
            return ref.score_matching.exact_score_matching(net, x, train=True).mean()
        finally:
            torch.set_default_dtype(saved)
    if fn == "a17_mdsm_baseline":
        with Replay([n]):
            return ref.deepebm.mdsm_baseline(net, x, c["sig_b"], c["sigma02"], c["D"])
    if fn == "a17_mdsm_tracetrick":
        with Replay([n, u]):
            return ref.deepebm.mdsm_tracetrick(net, x, c["sig_b"], c["sigma02"], c["D"])
    if fn == "a17_exact_deepebm":   # returns a detached mean (loss_train.py:20-26): value only, the gradient entry stays zero
        saved = torch.get_default_dtype()
        torch.set_default_dtype(x.dtype)
        try:
            return ref.deepebm.exact_score_matching(net, x, c["D"])
        finally:
            torch.set_default_dtype(saved)
    if fn == "a17_ssm":
        with Replay([u.reshape(c["B"], -1)]):
            return ref.deepebm.ssm(net, x)
    if fn == "a17_ssm_vr":
        with Replay([u.reshape(c["B"], -1)]):
            return ref.deepebm.ssm_vr(net, x)
    if fn == "a17_sliced":      # default noise_type='radermacher': sign of the replayed normal draw
        with Replay([u]):
            return ref.sliced.single_sliced_score_matching(net, x)[0]
    if fn == "a17_sliced_vr":
        with Replay([u]):
            return ref.sliced.sliced_VR_score_matching(net, x)[0]
    raise KeyError(fn)


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = load_reference()
    out = {}
    for name in cases.CASES:
        rec = {}
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            c = cases.build(name, dt)
            for p in c["net"].parameters():
                p.grad = None
            loss = reference_loss(ref, c)
            if loss.requires_grad:
                loss.backward()
            rec[f"loss_{tag}"] = loss.detach().double().numpy()
            rec[f"grad_{tag}"] = cases.flat_grad(c["net"]).double().numpy()
            # restatement check at generation time too (the CPU test suite repeats it)
            ol, og = run_case.loss_and_grad(cases.build(name, dt))
            rel = abs(float(ol) - float(loss)) / max(abs(float(loss)), 1e-30)
            print(f"{name:18s} {tag} ref={float(loss):+.9e} oracle={float(ol):+.9e} rel={rel:.2e}")
        for k, v in rec.items():
            out[f"{name}/{k}"] = v

    # network restatement: reference MLP_Quadratic_MNIST vs oracle.nets.QuadraticMLP
    from . import nets
    for act in ("nsoftplus", "elu"):
        af = ref.deepebm_models.NormalizedSoftplus() if act == "nsoftplus" else torch.nn.ELU()
        m = ref.deepebm_models.MLP_Quadratic_MNIST(1, 0, 28, AF=af)
        nets.init_from_numpy(m, 777)
        x = nets.np_rand(778, 8, 784).float()
        out[f"net_mlpq_mnist_{act}/energy"] = m(x).detach().double().numpy()
        # a7 on the real model, fp64, tiny batch: loss only (grad is 238k floats -> keep a digest)
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            m2 = ref.deepebm_models.MLP_Quadratic_MNIST(1, 0, 28, AF=af)
            nets.init_from_numpy(m2, 777)
            m2 = m2.to(dt)
            ref.FLAGS.batch_size, ref.FLAGS.esm_eps = 8, 0.05
            xu = nets.np_rand(778, 8, 1, 28, 28).to(dt)
            with Replay([nets.np_randn(779, 8, 784).to(dt)]):
                l = ref.deepebm.ssm_fd(m2, xu)
            l.backward()
            g = cases.flat_grad(m2).double()
            out[f"net_mlpq_mnist_{act}/ssm_fd_loss_{tag}"] = l.detach().double().numpy()
            out[f"net_mlpq_mnist_{act}/ssm_fd_gradnorm_{tag}"] = g.norm().numpy()
            out[f"net_mlpq_mnist_{act}/ssm_fd_grad_head_{tag}"] = g[-603:].numpy()  # ln1, ln2, lq
    # N3 / N4: the reference's annealed Langevin sampler and EMA helper on a tiny conditional score net
    sys.path.insert(0, f"{REF}/ncsnv2")
    for k in [k for k in sys.modules if k == "models" or k.startswith("models.") or k == "losses" or k.startswith("losses.")]:
        del sys.modules[k]
    import models as ref_models          # ncsnv2/models/__init__.py
    from models.ema import EMAHelper as RefEMA
    from . import sampling as osamp
    sig = cases.geometric_sigmas(1.0, 0.1, 3)
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        net = nets.TinyCondScore((3, 8, 8), 40, 3, sigmas=sig)
        nets.init_from_numpy(net, 4242)
        net = net.to(dt)
        x0 = nets.np_rand(4243, 6, 3, 8, 8).to(dt)
        zs = [nets.np_randn(4300 + i, 6, 3, 8, 8).to(dt) for i in range(6)]
        with Replay([z.clone() for z in zs]):
            xs = ref_models.anneal_Langevin_dynamics(x0.clone(), net, sig.numpy(), n_steps_each=2, step_lr=2e-3, final_only=True)
        out[f"n3_langevin/x_final_{tag}"] = xs[0].double().numpy()
        mine = osamp.anneal_langevin(x0, net, sig.numpy(), 2, 2e-3, zs)
        print(f"n3_langevin        {tag} |ref - oracle| max = {float((xs[0] - mine).abs().max()):.2e}")
        # EMA: three updates with drifting parameters
        lin = torch.nn.Linear(7, 5).to(dt)
        nets.init_from_numpy(lin, 4400)
        h = RefEMA(mu=0.9)
        h.register(lin)
        for i in range(3):
            with torch.no_grad():
                for j, p in enumerate(lin.parameters()):
                    p.add_(nets.np_randn(4410 + 10 * i + j, *p.shape).to(dt) * 0.1)
            h.update(lin)
        out[f"n4_ema/shadow_{tag}"] = torch.cat([h.shadow[n].reshape(-1) for n, _ in lin.named_parameters()]).double().numpy()
    np.savez_compressed(os.path.join(OUT, "reference_outputs.npz"), **out)
    print("wrote", os.path.join(OUT, "reference_outputs.npz"), os.path.getsize(os.path.join(OUT, "reference_outputs.npz")))


if __name__ == "__main__":
    main()
