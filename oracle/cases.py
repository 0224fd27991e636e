"""Fixture case table shared by oracle/make_golden.py (which runs the REFERENCE on
each case in the build container) and tests/ (which run the oracle restatement and
the CUDA path on the same case).  TEST INFRASTRUCTURE.

Every input is regenerated from numpy MT19937 seeds, so a fixture stores only the
reference's outputs (loss in fp32 and fp64, flat parameter gradient in both).
"""
import math

import torch

from . import nets

# name -> dict(fn=reference function id, net spec, sizes, hyper-parameters)
CASES = {
    # a1 efficient_score_matching_conjugate, config-1 shape (D=11 -> scalar tail path)
    "a1_gmm11_h128": dict(fn="a1", net=("softplus_mlp", 11, 128), B=64, D=11, eps=0.1, data="gmm"),
    "a1_gmm11_h30": dict(fn="a1", net=("softplus_mlp", 11, 30), B=100, D=11, eps=0.1, data="gmm"),
    "a1_quad_d64": dict(fn="a1", net=("quadratic_mlp", 64, 48, 3, "nsoftplus"), B=32, D=64, eps=0.1, data="unit"),
    # a7 ssm_fd / a8 ssm_fd_nop (DeepEBM; FLAGS.esm_eps=0.05 default.yaml:28)
    "a7_quad_d64": dict(fn="a7", net=("quadratic_mlp", 64, 48, 3, "nsoftplus"), B=32, D=64, eps=0.05, data="unit"),
    "a7_quad_elu_img": dict(fn="a7", net=("quadratic_mlp", 100, 40, 3, "elu"), B=16, D=100, eps=0.05, data="unit",
                            view=(1, 10, 10)),
    "a8_quad_d64": dict(fn="a8", net=("quadratic_mlp", 64, 48, 3, "nsoftplus"), B=32, D=64, eps=0.05, data="unit"),
    # a2 one-sided, a5 MLE
    "a2_gmm11": dict(fn="a2", net=("softplus_mlp", 11, 30), B=48, D=11, eps=0.01, data="gmm"),
    "a5_gmm11": dict(fn="a5", net=("softplus_mlp", 11, 30), B=48, D=11, eps=0.01, mle_ratio=10, data="gmm"),
    # a3 dsm_fd, a6 dsm_tracetrick_FD
    "a3_d20": dict(fn="a3", net=("softplus_mlp", 20, 32), B=40, D=20, eps=0.001, sigma=0.7, data="gmm"),
    "a6_d20": dict(fn="a6", net=("softplus_mlp", 20, 32), B=40, D=20, eps=0.1, sigma=0.7, data="gmm"),
    # a4 esm_scorenet_VR (VAE latent z in R^32)
    "a4_z32": dict(fn="a4", net=("score", (32,), 24, 1, False), B=24, D=32, eps=0.1, data="randn"),
    # a9 mdsm_fd / a10 nop (sigma per batch position, train_dsm.py:50-57)
    "a9_d48": dict(fn="a9", net=("quadratic_mlp", 48, 32, 3, "elu"), B=20, D=48, eps=0.05, sigma02=0.01,
                   loss_coe=1e-4, min_noise=0.05, max_noise=1.2, data="unit"),
    "a10_d48": dict(fn="a10", net=("quadratic_mlp", 48, 32, 3, "elu"), B=20, D=48, eps=0.05, sigma02=0.01,
                    loss_coe=1e-4, min_noise=0.05, max_noise=1.2, data="unit"),
    # a11/a12 ncsn (10 sigmas 1 -> 0.01), a13 ncsnv2 (sigma^anneal_power, hook, labels inside)
    "a11_img": dict(fn="a11", net=("score", (3, 8, 8), 40, 10, False), B=12, D=192, eps=0.1, L=10, smax=1.0,
                    smin=0.01, data="unit"),
    "a12_img": dict(fn="a12", net=("score", (3, 8, 8), 40, 10, False), B=12, D=192, L=10, smax=1.0, smin=0.01,
                    eps=0.1, data="unit"),
    "a13_img": dict(fn="a13", net=("score", (3, 8, 8), 40, 12, True), B=12, D=192, eps=0.1, L=12, smax=50.0,
                    smin=0.01, anneal_power=2.0, data="unit"),
    # a17 exact (Hessian-vector-product) SSM comparison losses: DeepEBM ssm / ssm_vr (Gaussian projections), SSM-repo
    # single_sliced_score_matching (sign projections) and sliced_VR_score_matching
    "a17_ssm_d64": dict(fn="a17_ssm", net=("quadratic_mlp", 64, 48, 3, "nsoftplus"), B=32, D=64, eps=0.0, data="unit"),
    "a17_ssmvr_d64": dict(fn="a17_ssm_vr", net=("quadratic_mlp", 64, 48, 3, "nsoftplus"), B=32, D=64, eps=0.0, data="unit"),
    "a17_sliced_gmm11": dict(fn="a17_sliced", net=("softplus_mlp", 11, 30), B=48, D=11, eps=0.0, data="gmm"),
    "a17_slicedvr_gmm11": dict(fn="a17_sliced_vr", net=("softplus_mlp", 11, 30), B=48, D=11, eps=0.0, data="gmm"),
    "a13_img_p1": dict(fn="a13", net=("score", (3, 8, 8), 40, 12, True), B=10, D=192, eps=0.1, L=12, smax=50.0,
                       smin=0.01, anneal_power=1.0, data="unit"),
}

# Cases added after the first GPU parity run.  Their GPU tests live in tests/test_zz_late_cases_gpu.py (collected last),
# the CPU oracle tests and make_golden treat them like every other case.
LATE = {
    # a2 with noise_type='radermacher' (the reference's spelling): v = sign(u) * eps  (efficient_sm.py:10-11)
    "a2_rad_gmm11": dict(fn="a2", net=("softplus_mlp", 11, 30), B=48, D=11, eps=0.01, data="gmm",
                         noise_type="radermacher"),
    # a17 ncsnv2 anneal_sliced_score_estimation_vr (ncsnv2/losses/dsm.py:33-74): labels drawn inside, Gaussian projections
    "a17_anneal_vr_img": dict(fn="a17_anneal_vr", net=("score", (3, 8, 8), 40, 12, True), B=10, D=192, eps=0.0, L=12,
                              smax=50.0, smin=0.01, anneal_power=2.0, n_particles=1, data="unit"),
    "a17_anneal_vr_p2": dict(fn="a17_anneal_vr", net=("score", (3, 8, 8), 40, 12, True), B=6, D=192, eps=0.0, L=12,
                             smax=50.0, smin=0.01, anneal_power=1.5, n_particles=2, data="unit"),
    # a17 exact_score_matching(train=True) (SSM/losses/score_matching.py:48-73): 0.5 ||grad logp||^2 + tr(Hessian logp)
    "a17_exact_gmm11": dict(fn="a17_exact", net=("softplus_mlp", 11, 30), B=24, D=11, eps=0.0, data="gmm"),
    # the rest of DeepEBM's loss table (ESM/inputs.py:96-136): the non-FD multiscale DSM losses a9 is compared with (and
    # evaluated by: get_eval_loss) and the offline exact-SM evaluator of test_exact_sm.py
    "a17_mdsm_baseline_d48": dict(fn="a17_mdsm_baseline", net=("quadratic_mlp", 48, 32, 3, "elu"), B=20, D=48, eps=0.0,
                                  sigma02=0.01, loss_coe=1e-4, min_noise=0.05, max_noise=1.2, data="unit"),
    "a17_mdsm_tracetrick_d48": dict(fn="a17_mdsm_tracetrick", net=("quadratic_mlp", 48, 32, 3, "elu"), B=20, D=48, eps=0.0,
                                    sigma02=0.01, loss_coe=1e-4, min_noise=0.05, max_noise=1.2, data="unit"),
    "a17_exact_deepebm_d16": dict(fn="a17_exact_deepebm", net=("quadratic_mlp", 16, 24, 2, "nsoftplus"), B=10, D=16, eps=0.0,
                                  data="unit", view=(1, 4, 4)),
}
CASES.update(LATE)


def geometric_sigmas(smax, smin, L):
    """ncsn anneal_runner / ncsnv2 models/__init__.py:4-17: exp(linspace(log smax, log smin, L))."""
    import numpy as np
    return torch.tensor(np.exp(np.linspace(np.log(smax), np.log(smin), L))).float()


def build(name, dtype=torch.float32):
    """Materialise a case: returns dict(net, x, draws..., labels, sigmas, hyper-params)."""
    c = dict(CASES[name])
    seed = abs(hash_name(name)) % 100000
    B, D = c["B"], c["D"]
    spec = c["net"]
    sig = None
    if "L" in c:
        sig = geometric_sigmas(c["smax"], c["smin"], c["L"])
    if spec[0] == "softplus_mlp":
        net = nets.SoftplusMLP(spec[1], spec[2])
    elif spec[0] == "quadratic_mlp":
        net = nets.QuadraticMLP(spec[1], spec[2], spec[3], spec[4])
    else:
        net = nets.TinyCondScore(spec[1], spec[2], spec[3], sigmas=sig if spec[4] else None)
    nets.init_from_numpy(net, seed)
    net = net.to(dtype)
    if c["data"] == "gmm":
        g = torch.Generator().manual_seed(0)  # only the mixture index; values from numpy
        sign = (nets.np_rand(seed + 1, B, 1) < 0.8).double() * 2 - 1
        x = nets.np_randn(seed + 2, B, D) + 5.0 * sign
    elif c["data"] == "unit":
        x = nets.np_rand(seed + 2, B, D)
    else:
        x = nets.np_randn(seed + 2, B, D)
    shape = c.get("view") or (spec[1] if spec[0] == "score" else None)
    if shape is not None and len(shape) > 1:
        x = x.reshape(B, *shape)
    c["net"] = net
    c["x"] = x.to(dtype)
    P = c.get("n_particles", 1)
    # direction draw (one per particle: the exact-SSM annealed loss duplicates the noised batch, ncsnv2/losses/dsm.py:43-56)
    c["u"] = nets.np_randn(seed + 3, P * x.shape[0], *x.shape[1:]).to(dtype)
    c["n"] = nets.np_randn(seed + 4, *x.shape).to(dtype)   # DSM noise draw
    if "L" in c:
        c["sigmas"] = sig.to(dtype)
        c["labels"] = nets.np_randint(seed + 5, c["L"], B)
    if c["fn"] in ("a9", "a10", "a17_mdsm_baseline", "a17_mdsm_tracetrick"):
        import numpy as np
        c["sig_b"] = torch.tensor(np.linspace(c["min_noise"], c["max_noise"], B)).view(-1, 1).to(dtype)
    return c


def hash_name(name):
    h = 0
    for ch in name:
        h = (h * 131 + ord(ch)) % 1000003
    return h


def flat_grad(net):
    return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                      for p in net.parameters()])
