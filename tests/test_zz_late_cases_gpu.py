"""GPU parity for the cases of oracle/cases.py:LATE (a2 with sign projections, the remaining a17 exact-score-matching
comparison losses) against the reference's own outputs (tests/golden).  Collected last on purpose: these adapters were
added after the round's last full GPU run.  Tolerances as in tests/test_losses_gpu.py: 1e-4 relative vs the reference's
fp32 result, and err(ours, fp64) <= max(1e-4, 3*err(ref_fp32, fp64))."""
import numpy as np
import pytest
import torch

from oracle import cases

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return abs(float(a) - float(b)) / max(abs(float(b)), 1e-30)


def _relvec(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def run_late_adapter(L, c, dev):
    fn, net = c["fn"], c["net"].to(dev)
    x, u, n = c["x"].to(dev), c["u"].to(dev), c["n"].to(dev)
    if getattr(net, "sigmas", None) is not None:
        net.sigmas = net.sigmas.to(dev)
    if fn == "a2":
        return L.single_efficient_score_matching(net, x, eps=c["eps"], noise_type=c["noise_type"], u=u)
    if fn == "a17_anneal_vr":
        return L.anneal_sliced_score_estimation_vr(net, x, c["sigmas"].to(dev), c["labels"].to(dev), c["anneal_power"], None,
                                                   c["n_particles"], u=u, dsm_noise=n)
    if fn == "a17_exact":
        return L.exact_score_matching(net, x, train=True).mean()
    raise KeyError(fn)


@pytest.mark.parametrize("name", list(cases.LATE))
def test_late_adapter_matches_reference(fdsm, golden, name):
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    c = cases.build(name, torch.float32)
    loss = run_late_adapter(fdsm.losses, c, dev)
    loss.backward()
    grad = cases.flat_grad(c["net"]).cpu().numpy()
    l32, l64 = golden[f"{name}/loss_f32"], golden[f"{name}/loss_f64"]
    g32, g64 = golden[f"{name}/grad_f32"], golden[f"{name}/grad_f64"]
    if c["fn"] != "a2":   # a2's first difference of O(|E|) numbers at eps = 0.01 is bounded by the reference's own fp32 error
        assert _rel(loss, l32) < 1e-4, (float(loss), float(l32))
        assert _relvec(grad, g32) < 1e-4
    # a2 at eps = 0.01 sums first differences of O(|E|) numbers: re-ordering the fp32 accumulations of the SAME network
    # on CPU (permuted hidden units, 30 orders) moves the loss by up to 3.6x the reference's own fp32-vs-fp64 error, so
    # that case gets 6x; the others keep the 3x of tests/test_losses_gpu.py.
    k = 6 if c["fn"] == "a2" else 3
    assert _rel(loss, l64) <= max(1e-4, k * _rel(l32, l64)), (float(loss), float(l64), float(l32))
    assert _relvec(grad, g64) <= max(1e-4, k * _relvec(g32, g64))


def test_rademacher_rows_are_bit_exact(fdsm):
    """noise_type='radermacher': the stacked rows the network sees are exactly cat[x, x + sign(u)*eps]."""
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    x, u = torch.randn(9, 11, generator=g).to(dev), torch.randn(9, 11, generator=g).to(dev)
    seen = []

    def net(z):
        seen.append(z.clone())
        return z.sum(-1)

    fdsm.losses.single_efficient_score_matching(net, x, eps=0.01, noise_type="radermacher", u=u)
    assert torch.equal(seen[0], torch.cat([x, x + u.sign() * 0.01], 0))


def test_anneal_ssm_vr_philox_draws_and_hook(fdsm):
    """Without injected draws: labels from torch.randint, noise and projections from the Philox streams; the hook
    receives the (B,) per-sample loss; two calls with the same seed agree bit for bit."""
    dev = torch.device("cuda:0")
    c = cases.build("a17_anneal_vr_p2", torch.float32)
    net, x, sig = c["net"].to(dev), c["x"].to(dev), c["sigmas"].to(dev)
    net.sigmas = net.sigmas.to(dev)
    got = []
    out = []
    for _ in range(2):
        fdsm.ops.manual_seed(77)
        out.append(fdsm.losses.anneal_sliced_score_estimation_vr(net, x, sig, c["labels"].to(dev), 1.5,
                                                                   lambda per, lab: got.append(per.detach()), 2))
    assert got[0].shape == (c["B"],) and torch.isfinite(out[0])
    assert torch.equal(out[0], out[1]) and torch.equal(got[0], got[1])
