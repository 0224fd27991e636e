"""GPU parity, stage level: kernels (1), (2), (4) through the C ABI vs the torch op sequence of the
reference / the numpy Philox oracle, on identical inputs.  Edge cases: empty batch, D=1, D=11
(scalar tail path), misaligned bases, padded leading dimensions, rows longer than the register
cache."""
import math

import numpy as np
import pytest
import torch

from oracle import fd_losses, philox

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


@pytest.mark.parametrize("B,D", [(1, 1), (7, 11), (64, 784), (5, 3072), (3, 4100), (2, 20000), (33, 6)])
def test_philox_normals_match_numpy_oracle(fdsm, B, D):
    got = fdsm.ops.philox_normal(B, D, 0, seed=0x1234ABCD5678, offset=5, row0=3).cpu().numpy()
    want = philox.normals(B, D, 0, 0x1234ABCD5678, 5, 3)
    # uniforms are bit-exact (integer Philox); the Box-Muller transform runs on the MUFU pipe (lg2/sin/cos.approx,
    # log1p series near u = 1): within 1e-5 absolute of the float64 transform for |n| up to 5.8
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-5)


def test_philox_statistics(fdsm):
    n = fdsm.ops.philox_normal(4096, 784, 1, seed=7, offset=0).double()
    assert abs(float(n.mean())) < 2e-3 and abs(float(n.std()) - 1) < 2e-3
    from scipy import stats
    assert stats.kstest(n.flatten()[:50000].cpu().numpy(), "norm").pvalue > 1e-3
    cov = (n[:, :16].T @ n[:, :16]) / n.shape[0]
    assert float((cov - torch.eye(16, device=n.device, dtype=n.dtype)).abs().max()) < 0.08  # isotropy


@pytest.mark.parametrize("B,D", [(1, 1), (9, 11), (128, 784), (16, 3072), (4, 4100), (3, 16388)])
def test_direction_unit_norm_and_sharding(fdsm, B, D):
    eps = 0.05
    v = fdsm.ops.direction(B, D, eps, device=_dev(), draw=(11, 2, 0))
    nrm = v.double().norm(dim=-1)
    assert float((nrm - eps).abs().max()) < 1e-6 * eps * 10
    want = fd_losses.sphere_direction(torch.from_numpy(philox.normals(B, D, 0, 11, 2)), eps)
    np.testing.assert_allclose(v.cpu().numpy(), want.numpy(), rtol=2e-5, atol=1e-8)
    if B >= 2:  # a shard starting at row0 draws what the full batch draws for those rows
        h = B // 2
        v2 = fdsm.ops.direction(B - h, D, eps, device=_dev(), draw=(11, 2, h))
        assert torch.equal(v2, v[h:])


@pytest.mark.parametrize("B,D", [(0, 8), (5, 0), (1, 1), (13, 11), (64, 784), (8, 3072), (3, 4100), (2, 17001)])
@pytest.mark.parametrize("mask", [7, 6, 3])
def test_stack_injected_direction_bit_exact(fdsm, B, D, mask):
    """Already-normalised v (the reference's noise= hook): rows must equal torch's x, x+v, x-v bit for bit."""
    g = torch.Generator().manual_seed(B * 1000 + D)
    x = torch.randn(B, D, generator=g).to(_dev())
    v = (torch.randn(B, D, generator=g) * 0.01).to(_dev())
    out, _, _ = fdsm.ops.stack(x, row_mask=mask, scale=1.0, direction=v, normalize=False)
    rows = [r for bit, r in ((1, x), (2, x + v), (4, x - v)) if mask & bit]
    want = torch.cat(rows, 0) if rows else None
    assert out.shape == want.shape
    assert torch.equal(out, want)


@pytest.mark.parametrize("B,D", [(13, 11), (64, 784), (8, 3072), (3, 4100)])
def test_stack_normalises_like_torch(fdsm, B, D):
    g = torch.Generator().manual_seed(5)
    x = torch.rand(B, D, generator=g).to(_dev())
    u = torch.randn(B, D, generator=g).to(_dev())
    eps = 0.05
    out, v, _ = fdsm.ops.stack(x, row_mask=7, scale=eps, direction=u, normalize=True, want_v=True)
    vt = u / torch.norm(u, p=2, dim=-1, keepdim=True) * eps      # loss_train.py:218
    # norm summation order differs -> <= 2 ulp on v; rows follow
    assert float(((v - vt).abs() / vt.abs().clamp_min(1e-30)).max()) < 3e-7
    want = torch.cat([x, x + v, x - v], 0)
    assert torch.equal(out, want)


def test_stack_misaligned_and_padded(fdsm):
    """Base not 16-byte aligned / ld not a multiple of 4 -> scalar path, same numbers."""
    B, D = 6, 64
    big = torch.randn(B, D + 3, device=_dev())
    x = big[:, 1:D + 1]                       # ld = D+3, base offset 4 bytes
    v = torch.randn(B, D, device=_dev()) * 0.1
    out, _, _ = fdsm.ops.stack(x, row_mask=7, scale=1.0, direction=v, normalize=False)
    assert torch.equal(out, torch.cat([x, x + v, x - v], 0))
    outp = torch.full((3 * B, D + 4), -1.0, device=_dev())   # padded output leading dimension
    fdsm.ops.stack(x, row_mask=7, scale=1.0, direction=v, normalize=False, out=outp[:, :D])
    assert torch.equal(outp[:, :D], torch.cat([x, x + v, x - v], 0))
    assert bool((outp[:, D:] == -1).all())


@pytest.mark.parametrize("B,D", [(10, 20), (16, 768), (4, 3072)])
@pytest.mark.parametrize("dot_mode", [0, 1])
def test_stack_noised_matches_torch(fdsm, B, D, dot_mode):
    g = torch.Generator().manual_seed(3)
    x = torch.rand(B, D, generator=g).to(_dev())
    n = torch.randn(B, D, generator=g).to(_dev())
    u = torch.randn(B, D, generator=g).to(_dev())
    sig = torch.linspace(0.05, 1.2, B).to(_dev())
    eps = 0.05
    out, v, dot = fdsm.ops.stack(x, row_mask=6, scale=eps, direction=u, normalize=True, noise=n, sigma=sig,
                                 perturb=True, dot_mode=dot_mode, want_v=True, want_dot=True)
    xn = x + sig.view(-1, 1) * n                                   # loss_train.py:101-102
    assert torch.equal(out, torch.cat([xn + v, xn - v], 0))
    want = (v * (xn - x)).sum(-1) if dot_mode == 1 else (v * (sig.view(-1, 1) * n)).sum(-1)
    np.testing.assert_allclose(dot.cpu().numpy(), want.cpu().numpy(), rtol=2e-5, atol=1e-7)
    # label-indexed sigma table (ncsn / ncsnv2)
    table = torch.exp(torch.linspace(math.log(1.0), math.log(0.01), 10)).to(_dev())
    labels = torch.randint(0, 10, (B,), generator=g).to(_dev())
    out2, v2, _ = fdsm.ops.stack(x, row_mask=6, scale=eps, direction=u, normalize=True, noise=n, sigma=table,
                                 labels=labels, want_v=True)
    xn2 = x + n * table[labels].view(-1, 1)
    assert torch.equal(out2, torch.cat([xn2 + v2, xn2 - v2], 0))


def test_stack_philox_fused_equals_two_step(fdsm):
    """Philox-mode fused build == standalone Philox draws fed through the injected path, up to the normalisation's last
    bit: injected directions divide per element (the reference's rounding sequence), in-kernel draws multiply by one
    per-row reciprocal (no reference bit pattern exists for them)."""
    B, D = 32, 784
    x = torch.rand(B, D, device=_dev())
    draw = (99, 4, 100)
    sig = torch.full((B,), 0.3, device=_dev())
    out, v, dot = fdsm.ops.stack(x, row_mask=6, scale=0.1, sigma=sig, want_v=True, want_dot=True, draw=draw)
    u = fdsm.ops.philox_normal(B, D, 0, *draw)
    n = fdsm.ops.philox_normal(B, D, 1, *draw)
    out2, v2, dot2 = fdsm.ops.stack(x, row_mask=6, scale=0.1, direction=u, noise=n, sigma=sig, want_v=True, want_dot=True)
    torch.testing.assert_close(v, v2, rtol=3e-7, atol=0)
    torch.testing.assert_close(out, out2, rtol=0, atol=1.2e-7)        # rows in [0, 1.1]: at most one ulp
    torch.testing.assert_close(dot, dot2, rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("B", [1, 5, 1024, 4099])
@pytest.mark.parametrize("mode", [0, 1])
def test_combine_energy_fwd_bwd(fdsm, B, mode):
    rows = 3 if mode == 0 else 2
    E = torch.randn(rows * B, device=_dev(), requires_grad=True)
    scale, c1 = 784 / 0.05 ** 2, 2.0
    loss, per = fdsm.ops.CombineEnergy.apply(E, B, mode, c1, scale, True, True)
    Ed = E.detach().double().requires_grad_(True)
    if mode == 0:
        want, wper = fd_losses.combine_energy(Ed, B, scale, c1)
    else:
        d = Ed[:B] - Ed[B:]
        wper = (d * d + 4 * d) * scale
        want = wper.mean()
    assert abs(float(loss) - float(want)) <= 1e-6 * max(abs(float(want)), float(wper.abs().mean()))
    np.testing.assert_allclose(per.cpu().numpy(), wper.detach().cpu().numpy(), rtol=1e-5, atol=1e-5 * float(wper.abs().mean()))
    (loss * 3.0).backward()
    (want * 3.0).backward()
    np.testing.assert_allclose(E.grad.cpu().numpy(), Ed.grad.cpu().numpy(), rtol=1e-5, atol=1e-6 * float(Ed.grad.abs().max()))


@pytest.mark.parametrize("B", [1, 20, 4099])
@pytest.mark.parametrize("weighted", [False, True])
def test_combine_dsm_fwd_bwd(fdsm, B, weighted):
    E = torch.randn(2 * B, device=_dev(), requires_grad=True)
    dot = torch.randn(B, device=_dev())
    sig = torch.linspace(0.05, 1.2, B).to(_dev()) if weighted else None
    alpha, beta, gamma = 0.5, 100.0, 1e-4 * 784 / 0.05 ** 2
    loss = fdsm.ops.CombineDsm.apply(E, dot, sig, B, alpha, beta, gamma)
    Ed = E.detach().double().requires_grad_(True)
    r = alpha * (Ed[B:] - Ed[:B]) + beta * dot.double()
    per = r * r / (sig.double() ** 2 if weighted else 1.0)
    want = per.mean() * gamma
    assert abs(float(loss) - float(want)) <= 2e-6 * abs(float(want))
    (loss * 0.5).backward()
    (want * 0.5).backward()
    np.testing.assert_allclose(E.grad.cpu().numpy(), Ed.grad.cpu().numpy(), rtol=1e-5, atol=1e-6 * float(Ed.grad.abs().max()))


@pytest.mark.parametrize("B,D", [(1, 1), (6, 11), (12, 192), (128, 3072), (3, 4100), (2, 20000)])
@pytest.mark.parametrize("projected", [False, True])
def test_combine_score_fwd_bwd_and_fused(fdsm, B, D, projected):
    g = torch.Generator().manual_seed(B + D)
    s = torch.randn(2 * B, D, generator=g).to(_dev()).requires_grad_(True)
    v = (torch.randn(B, D, generator=g) * 0.01).to(_dev())
    table = torch.exp(torch.linspace(math.log(50.0), math.log(0.01), 12)).to(_dev())
    labels = torch.randint(0, 12, (B,), generator=g).to(_dev())
    k = D / (2 * 0.1 ** 2)
    for power in (2.0, 1.0, 1.5):
        s.grad = None
        loss, per = fdsm.ops.CombineScore.apply(s, v, table, labels, power, B, D, k, projected, True)
        sd = s.detach().double().requires_grad_(True)
        w = table.double()[labels] ** power
        want, wper = fd_losses.combine_score(sd, v.double(), B, k, w, projected)
        tol = 2e-6 * float(wper.abs().mean() + abs(float(want)))
        assert abs(float(loss) - float(want)) <= tol
        np.testing.assert_allclose(per.cpu().numpy(), wper.detach().cpu().numpy(), rtol=2e-5, atol=tol)
        (loss * 2.0).backward()
        (want * 2.0).backward()
        gmax = float(sd.grad.abs().max())
        np.testing.assert_allclose(s.grad.cpu().numpy(), sd.grad.cpu().numpy(), rtol=1e-5, atol=1e-6 * gmax)
        l2, p2, ds = fdsm.ops.combine_score_fused(s.detach(), v, k=k, sigma=table, labels=labels, power=power,
                                                  projected=projected)
        assert torch.equal(l2, loss.detach()) and torch.equal(p2, per)
        np.testing.assert_allclose(ds.cpu().numpy() * 2.0, s.grad.cpu().numpy(), rtol=1e-6, atol=1e-7 * gmax)


def test_combine_rejects_bad_arguments(fdsm):
    E = torch.randn(10, device=_dev())
    with pytest.raises(ValueError):
        fdsm.ops.CombineEnergy.apply(E, 4, 0, 2.0, 1.0, True, False)   # 10 != 3*4
    with pytest.raises(TypeError):
        fdsm.ops.stack(torch.randn(4, 8, device=_dev()).double(), row_mask=7, scale=0.1)
    lib = fdsm._lib.load()
    rc = lib.fdsm_combine_energy_fwd(None, 4, 5, 2.0, 1.0, 1, None, None, None)
    assert rc == -1 and b"mode" in lib.fdsm_last_error_string()


def test_peer_allreduce_across_gpus():
    """fdsm_peer_allreduce (NVLink peer-memory all-reduce + squared norm + step counter) against NCCL, across ranks and
    under graph replay -- scripts/peer_allreduce_check.py under torchrun on every visible GPU (needs >= 2)."""
    import os
    import subprocess
    import sys
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
                        "127.0.0.1", "--master-port", "29533", os.path.join(root, "scripts", "peer_allreduce_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
