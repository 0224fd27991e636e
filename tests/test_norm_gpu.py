"""GPU parity of the fused InstanceNorm++ (+ ELU) kernels (first piece) through ops.inorm_plus /
the C ABI: against the reference modules' golden outputs and autograd gradients (tests/golden/inorm_plus.npz), against the
float64 oracle (oracle/norm.py) on seeded inputs over every code path (plane sizes -> group widths 4..32, scalar and 128-bit
paths, register-held and long planes, shared and per-sample parameter rows), and at the config-4/5 activation shape against
the ATen op sequence.  Bound: 2e-5 of the tensor's largest magnitude (fp32 rounding of sums over up to 4096 terms; the
reference's own fp32 run is held to the same bound against float64 in tests/test_norm_cpu.py)."""
import numpy as np
import pytest
import torch

from oracle import norm as onorm
from test_norm_cpu import CASES, GOLD, load_case, param_grads

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 2e-5


def close(got, want, what, tol=TOL):
    got = got.detach().cpu().double().numpy() if isinstance(got, torch.Tensor) else np.asarray(got, dtype=np.float64)
    err = np.abs(got - want).max() if want.size else 0.0
    assert err <= tol * max(1.0, np.abs(want).max()), f"{what}: max abs err {err:.3e}"


def run_kernel(fdsm, x, gamma, alpha, beta, act, gout):
    t = lambda a: None if a is None else torch.tensor(np.ascontiguousarray(a), dtype=torch.float32, device=DEV)
    xt, g, a = t(x).requires_grad_(), t(gamma).requires_grad_(), t(alpha).requires_grad_()
    b = t(beta).requires_grad_() if beta is not None else None
    out = fdsm.ops.inorm_plus(xt, g, a, b, act)
    out.backward(t(gout))
    return out, xt.grad, g.grad, a.grad, (b.grad if b is not None else None)


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("act", [False, True])
def test_kernels_match_reference_modules(fdsm, name, act):
    gold = np.load(GOLD)
    x, gout, table, labels, gamma, alpha, beta, cond, bias = load_case(gold, name)
    out, dx, dg, da, db = run_kernel(fdsm, x, gamma, alpha, beta, act, gout)
    key = f"{name}/{'elu' if act else 'id'}/f64/"
    close(out, gold[key + "out"], "out")
    close(dx, gold[key + "dx"], "dx")
    cpu = lambda t: None if t is None else t.cpu().double().numpy()
    N = x.shape[0]
    rows = lambda t: None if t is None else (cpu(t) if cond else np.concatenate([cpu(t)] + [np.zeros_like(cpu(t))] * (N - 1)))
    for got, want in param_grads(gold, key, rows(dg), rows(da), rows(db) if bias else None, table, labels, cond, bias):
        close(got, want, "parameter gradient")


SHAPES = [(4, 128, 32, 32), (3, 256, 16, 16), (2, 64, 8, 8), (2, 16, 4, 4), (2, 8, 6, 6), (2, 5, 5, 7), (1, 3, 64, 64), (2, 4, 40, 40),
          (5, 2, 1, 4), (0, 4, 8, 8)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("per_sample", [False, True])
@pytest.mark.parametrize("act", [False, True])
def test_kernels_match_oracle(fdsm, shape, per_sample, act):
    N, C, H, W = shape
    rs = np.random.RandomState(N * 1000 + C * 10 + H)
    x = rs.randn(N, C, H, W) * rs.uniform(0.5, 2.0, size=(1, C, 1, 1)) + rs.randn(N, C, 1, 1)
    gout = rs.randn(N, C, H, W)
    rows = N if per_sample else 1
    gamma, alpha, beta = rs.normal(1, 0.2, (rows, C)), rs.normal(1, 0.2, (rows, C)), rs.normal(0, 0.3, (rows, C))
    out, dx, dg, da, db = run_kernel(fdsm, x, gamma, alpha, beta, act, gout)
    if N == 0:
        assert out.shape == (0, C, H, W) and dx.shape == (0, C, H, W)
        return
    want, cache = onorm.forward(x.astype(np.float32), gamma.astype(np.float32), alpha.astype(np.float32), beta.astype(np.float32), act)
    wdx, wdg, wda, wdb = onorm.backward(gout.astype(np.float32), cache)
    red = (lambda a: a) if per_sample else (lambda a: a.sum(0, keepdims=True))
    close(out, want, "out")
    close(dx, wdx, "dx")
    close(dg, red(wdg), "dgamma", 5e-5)
    close(da, red(wda), "dalpha", 5e-5)
    close(db, red(wdb), "dbeta", 5e-5)


def test_embedding_chunks_and_unaligned_base(fdsm):
    """Parameter rows as the conditional module hands them over (views into one (N, 3C) embedding lookup: row stride 3C),
    and an activation whose base is not 16-byte aligned (scalar path)."""
    N, C, H, W = 6, 12, 8, 8
    torch.manual_seed(3)
    emb = torch.nn.Embedding(5, 3 * C).to(DEV)
    y = torch.randint(0, 5, (N,), device=DEV)
    big = torch.randn(N * C * H * W + 1, device=DEV)
    for x in (big[:-1].view(N, C, H, W), big[1:].view(N, C, H, W)):
        gout = torch.randn(N, C, H, W, device=DEV)
        res = []
        for kernel in (True, False):
            xt = x.clone().requires_grad_() if kernel else x.double().requires_grad_()
            emb.zero_grad()
            g, a, b = (emb(y) if kernel else emb(y).double()).chunk(3, dim=1)
            out = fdsm.ops.inorm_plus(xt, g, a, b, True) if kernel else fdsm.ops.inorm_plus_reference(xt, g, a, b, True)
            out.backward(gout if kernel else gout.double())
            res.append((out.detach().double(), xt.grad.double(), emb.weight.grad.double().clone()))
        for got, want, what in zip(res[0], res[1], ("out", "dx", "dtable")):
            close(got, want.cpu().numpy(), what, 5e-5)
    # the unaligned view itself, not a clone of it
    xu = big[1:].view(N, C, H, W)
    g, a, b = emb(y).detach().chunk(3, dim=1)
    close(fdsm.ops.inorm_plus(xu, g, a, b, False), fdsm.ops.inorm_plus_reference(xu.double(), g.double(), a.double(), b.double()).cpu().numpy(),
          "out (unaligned base)")


def test_gradient_of_gradient_takes_the_op_sequence(fdsm):
    """create_graph=True (the exact-SSM comparison losses differentiate the score): same second-order result as the op sequence."""
    N, C, H, W = 3, 8, 8, 8
    x = torch.randn(N, C, H, W, device=DEV)
    gamma, alpha, beta = (torch.randn(1, C, device=DEV) * 0.2 + 1 for _ in range(3))
    v = torch.randn_like(x)
    res = []
    for fn in (fdsm.ops.inorm_plus, fdsm.ops.inorm_plus_reference):
        xt, g = x.clone().requires_grad_(), gamma.clone().requires_grad_()
        out = fn(xt, g, alpha, beta, True)
        (gx,) = torch.autograd.grad(out.sum(), xt, create_graph=True)
        (gx * v).sum().backward()
        res.append((gx.detach(), xt.grad, g.grad))
    for got, want in zip(*res):
        close(got, want.double().cpu().numpy(), "second order", 1e-4)


def test_config_shape_against_aten(fdsm):
    """(256, 128, 32, 32): the activation of configs 4 / 5 at B = 128 (two stacked rows), 134 MB."""
    N, C, H, W = 256, 128, 32, 32
    x = torch.randn(N, C, H, W, device=DEV) * 1.5 + 0.3
    gout = torch.randn(N, C, H, W, device=DEV)
    gamma, alpha, beta = (torch.randn(1, C, device=DEV) * 0.1 + 1 for _ in range(3))
    res = []
    for fn in (fdsm.ops.inorm_plus, fdsm.ops.inorm_plus_reference):
        xt, g, a, b = (t.clone().requires_grad_() for t in (x, gamma, alpha, beta))
        out = fn(xt, g, a, b, True)
        out.backward(gout)
        res.append((out.detach(), xt.grad, g.grad, a.grad, b.grad))
        del out, xt
    for got, want, what in zip(res[0], res[1], ("out", "dx", "dgamma", "dalpha", "dbeta")):
        scale = max(1.0, float(want.abs().max()))
        tol = 2e-5 if what in ("out", "dx") else 2e-4      # parameter sums over 268 M terms: ATen's own fp32 summation order
        assert float((got - want).abs().max()) <= tol * scale, what
    # determinism: fixed-order reductions
    out2 = fdsm.ops.inorm_plus(x, gamma, alpha, beta, True)
    assert torch.equal(out2, res[0][0])


def test_score_nets_same_output_with_and_without_the_kernels(fdsm, monkeypatch):
    """NCSNv2 / CondRefineNetDilated with our normaliser kernels against the same network on the ATen op sequence.  Thirty
    normalised layers amplify fp32 rounding (the two fp32 runs differ by ~1e-3 of the output), so both are held against the
    float64 run of the same network: ours may be no further from it than three times the ATen fp32 run."""
    import copy
    torch.manual_seed(11)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        for net, ylim in ((fdsm.nets.NCSNv2(ngf=16, num_classes=7), 7), (fdsm.nets.CondRefineNetDilated(ngf=16, num_classes=5), 5)):
            net = net.to(DEV)
            x = torch.rand(4, 3, 32, 32, device=DEV)
            y = torch.randint(0, ylim, (4,), device=DEV)
            net64 = copy.deepcopy(net).double()
            out64 = net64(x.double(), y)
            out64.square().mean().backward()
            want = (out64.detach(), torch.cat([p.grad.reshape(-1) for p in net64.parameters()]))
            errs = []
            for flag in ("1", "0"):
                monkeypatch.setenv("FDSM_NORM_KERNEL", flag)
                net.zero_grad()
                launches0 = fdsm.ops.LAUNCHES
                out = net(x, y)
                out.square().mean().backward()
                assert (fdsm.ops.LAUNCHES > launches0) == (flag == "1")
                got = (out.detach(), torch.cat([p.grad.reshape(-1) for p in net.parameters()]))
                errs.append([float((g.double() - w).abs().max() / w.abs().max()) for g, w in zip(got, want)])
            for e_ours, e_aten, what in zip(errs[0], errs[1], ("output", "parameter gradient")):
                assert e_ours <= max(3.0 * e_aten, 1e-5), f"{type(net).__name__} {what}: ours {e_ours:.2e}, ATen fp32 {e_aten:.2e} (relative to float64)"
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
