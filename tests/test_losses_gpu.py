"""GPU parity, end to end: every reference-signature adapter (generic mode: our kernels (1),(2),(4)
around the caller's torch network) vs the reference's own outputs (tests/golden) and the CPU oracle
on identical injected draws.

Tolerance (SURVEY.md §7 hard part 1): 1e-4 relative vs the reference's fp32 result on
the well-conditioned forms; every form must satisfy err(ours, fp64) <= max(1e-4, 3*err(ref_fp32, fp64))
for the loss and for the flat parameter gradient."""
import numpy as np
import pytest
import torch

from oracle import cases, run_case

pytestmark = pytest.mark.gpu

WELL_CONDITIONED = {"a4", "a5", "a6", "a9", "a10", "a11", "a12", "a13", "a17_ssm", "a17_ssm_vr", "a17_sliced", "a17_sliced_vr"}


def _rel(a, b):
    return abs(float(a) - float(b)) / max(abs(float(b)), 1e-30)


def _relvec(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def run_adapter(fdsm, c, dev):
    fn, net, eps = c["fn"], c["net"].to(dev), c["eps"]
    x, u, n = c["x"].to(dev), c["u"].to(dev), c["n"].to(dev)
    L = fdsm.losses
    if fn == "a17_ssm":
        return L.ssm(net, x, u=u.reshape(c["B"], -1))
    if fn == "a17_ssm_vr":
        return L.ssm_vr(net, x, u=u.reshape(c["B"], -1))
    if fn == "a17_sliced":
        return L.single_sliced_score_matching(net, x, u=u)[0]
    if fn == "a17_sliced_vr":
        return L.sliced_VR_score_matching(net, x, u=u)[0]
    if fn == "a1":
        return L.efficient_score_matching_conjugate(net, x, eps=eps, u=u)
    if fn == "a7":
        return L.ssm_fd(net, x, eps=eps, u=u)
    if fn == "a8":
        return L.ssm_fd_nop(net, x, eps=eps, u=u)
    if fn == "a2":
        return L.single_efficient_score_matching(net, x, eps=eps, u=u)
    if fn == "a5":
        return L.MLE_efficient_score_matching_conjugate(net, x, eps=eps, mle_ratio=c["mle_ratio"], u=u)
    if fn == "a3":
        return L.dsm_fd(net, x, sigma=c["sigma"], eps=eps, u=u, dsm_noise=n)
    if fn == "a6":
        return L.dsm_tracetrick_FD(net, x, sigma=c["sigma"], eps=eps, u=u, dsm_noise=n)
    if fn == "a4":
        return L.esm_scorenet_VR(net, x, eps=eps, u=u)
    if fn in ("a9", "a10"):
        f = L.mdsm_fd if fn == "a9" else L.mdsm_fd_nop
        return f(net, x, c["sig_b"].to(dev), c["sigma02"], c["D"], eps=eps, loss_coe=c["loss_coe"], u=u, dsm_noise=n)
    labels, sigmas = c["labels"].to(dev), c["sigmas"].to(dev)
    if getattr(net, "sigmas", None) is not None:
        net.sigmas = net.sigmas.to(dev)
    if fn == "a11":
        return L.ncsn_anneal_ESM_scorenet_VR(net, x, labels, sigmas, 1, eps, u=u, dsm_noise=n)
    if fn == "a12":
        return L.ncsn_anneal_ESM_scorenet(net, x, labels, sigmas, 1, u=u, dsm_noise=n)
    if fn == "a13":
        return L.anneal_ESM_scorenet_VR(net, x, sigmas, labels, c["anneal_power"], None, 1, eps, u=u, dsm_noise=n)
    raise KeyError(fn)


@pytest.mark.parametrize("name", [n for n in cases.CASES if n not in cases.LATE])
def test_adapter_matches_reference(fdsm, golden, name):
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    c = cases.build(name, torch.float32)
    loss = run_adapter(fdsm, c, dev)
    loss.backward()
    grad = cases.flat_grad(c["net"]).cpu().numpy()
    l32, l64 = golden[f"{name}/loss_f32"], golden[f"{name}/loss_f64"]
    g32, g64 = golden[f"{name}/grad_f32"], golden[f"{name}/grad_f64"]
    if c["fn"] in WELL_CONDITIONED:
        assert _rel(loss, l32) < 1e-4, (float(loss), float(l32))
        assert _relvec(grad, g32) < 1e-4
    # fp32 noise floor of the second difference -E2-E3+2E1 (O(eps^2) signal from O(|E|) numbers): a few ulp
    # of |E| per term, times the loss scale.  Only the 3-row energy forms need it (a1/a7/a8).
    floor = 0.0
    if c["fn"] in ("a1", "a7", "a8"):
        with torch.no_grad():
            emag = float(c["net"](c["x"].to(dev).reshape(c["B"], -1)).abs().mean())
        scale = (c["D"] if c["fn"] == "a1" else 1.0) / c["eps"] ** 2
        floor = 16 * 2.0 ** -24 * emag * scale / abs(float(l64))
    assert _rel(loss, l64) <= max(1e-4, 3 * _rel(l32, l64), floor), (float(loss), float(l64), float(l32), floor)
    assert _relvec(grad, g64) <= max(1e-4, 3 * _relvec(g32, g64))


def test_detach_and_noise_hook(fdsm):
    """a1: the reference's own `noise=` injection and `detach=True` (value only)."""
    dev = torch.device("cuda:0")
    c = cases.build("a1_gmm11_h128", torch.float32)
    net, x = c["net"].to(dev), c["x"].to(dev)
    v = run_case.directions(c).to(dev)
    l1 = fdsm.efficient_score_matching_conjugate(net, x, noise=v, eps=c["eps"])
    l2 = fdsm.efficient_score_matching_conjugate(net, x, eps=c["eps"], u=c["u"].to(dev))
    assert _rel(l1, l2) < 5e-3   # same direction up to 1 ulp; second difference amplifies it
    ld = fdsm.efficient_score_matching_conjugate(net, x, noise=v, eps=c["eps"], detach=True)
    assert not ld.requires_grad and _rel(ld, l1) < 1e-6


def test_hook_receives_per_sample_loss(fdsm):
    dev = torch.device("cuda:0")
    c = cases.build("a13_img", torch.float32)
    seen = {}
    net = c["net"].to(dev)
    net.sigmas = net.sigmas.to(dev)
    loss = fdsm.anneal_ESM_scorenet_VR(net, c["x"].to(dev), c["sigmas"].to(dev), c["labels"].to(dev), 2.0,
                                       lambda per, lab: seen.update(per=per, lab=lab), u=c["u"].to(dev),
                                       dsm_noise=c["n"].to(dev))
    assert seen["per"].shape == (c["B"],) and torch.equal(seen["lab"], c["labels"].to(dev))
    assert _rel(seen["per"].mean(), loss) < 1e-5


def test_philox_mode_is_a_valid_estimator(fdsm):
    """No injection: Philox directions.  FD-SSM on an analytic quadratic energy is exact for any eps, so
    the production-mode loss must equal the closed form for the v the kernel drew."""
    dev = torch.device("cuda:0")
    D, B, eps = 16, 4096, 0.1
    A = torch.randn(D, D, device=dev, dtype=torch.float64)
    A = A + A.T
    energy = lambda z: (0.5 * ((z.double() @ A) * z.double()).sum(-1)).float()
    x = torch.randn(B, D, device=dev)
    fdsm.manual_seed(123)
    loss = fdsm.efficient_score_matching_conjugate(energy, x, eps=eps)
    v = fdsm.ops.direction(B, D, 1.0, device=dev, draw=(123, 0, 0)).double()
    want = (0.5 * ((x.double() @ A) * v).sum(-1) ** 2 - ((v @ A) * v).sum(-1)).mean() * D
    assert _rel(loss, want) < 2e-2   # fp32 energies, eps=0.1 second difference
    # and it is deterministic under the same seed, different under another
    fdsm.manual_seed(123)
    assert float(fdsm.efficient_score_matching_conjugate(energy, x, eps=eps)) == float(loss)


def test_gan_fd_gradient_penalty_matches_reference_expression(fdsm):
    """N4: GAN_FD/library/trainer_fd.py:57-73 restated inline (randn_like -> l2norm over dims [1,2,3] -> noise = u/||u|| *
    fd_eps * sqrt(D); cat[x, x - noise]; reg = mean(((d1 - d2)/fd_eps)^2)) on injected draws."""
    dev = torch.device("cuda:0")
    B, shape, fd_eps, reg_param = 16, (3, 8, 8), 0.1, 10.0
    x = torch.rand(B, *shape, device=dev)
    y = torch.randint(0, 10, (B,), device=dev)
    u = torch.randn(B, *shape, device=dev)
    emb = torch.nn.Embedding(10, 8).to(dev)
    body = torch.nn.Sequential(torch.nn.Linear(3 * 8 * 8 + 8, 32), torch.nn.ELU(), torch.nn.Linear(32, 1)).to(dev)
    disc = lambda inp, lab: body(torch.cat([inp.reshape(inp.shape[0], -1), emb(lab)], 1)).view(-1)
    d_real, reg = fdsm.fd_gradient_penalty(disc, x, y, fd_eps, reg_param, u=u)
    l2norm = torch.pow(torch.sum(u ** 2, dim=[1, 2, 3]), 0.5).view(B, 1, 1, 1)
    noise = u / l2norm * fd_eps * np.sqrt(3 * 8 * 8)
    cat_d = disc(torch.cat([x, x - noise], 0), torch.cat([y, y], 0))
    want_reg = reg_param * (((cat_d[:B] - cat_d[B:]) / fd_eps) ** 2).mean()
    assert torch.allclose(d_real, cat_d[:B], rtol=0, atol=0)
    assert _rel(reg, want_reg) < 2e-4          # (u/||u||)*c vs u/||u||*eps*sqrt(D): one rounding apart, amplified by 1/eps
    (d_real.mean() + reg).backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in body.parameters())
