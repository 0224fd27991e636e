"""GPU parity of kernel (3)'s building block: the tcgen05/TMEM GEMM (through the C ABI) against an fp64 matmul,
for every operand-majorness combination the MLP uses (fwd: K/MN, dgrad: K/K, wgrad: MN/MN), plain TF32 and 3xTF32,
split-K, ragged M/N/K (TMA zero fill)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = {1: 2e-3, 3: 3e-6}   # max |D - ref| / (|a_m| |b_n|): TF32 (10-bit mantissa) vs 3xTF32 (~fp32)


def _run(fdsm, M, N, K, am, bm, passes, splits=1, tile_n=0, seed=0, raw=False, want_out=False):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(seed)
    A = torch.randn(M, K, device=dev, generator=g)
    B = torch.randn(N, K, device=dev, generator=g)
    # raw=True: plain fp32 operands (padded copy), hi/lo split inside the kernel
    ah, al = fdsm.ops.tf32_split(A if am == 0 else A.t().contiguous(), raw=raw)
    bh, bl = fdsm.ops.tf32_split(B if bm == 0 else B.t().contiguous(), raw=raw)
    out = fdsm.ops.tc_gemm(ah, al, bh, bl, M, N, K, a_major=am, b_major=bm, passes=passes, splits=splits, tile_n=tile_n)
    if want_out:
        return out
    D = out.double().sum(0)
    ref = A.double() @ B.double().t()
    scale = A.double().norm(dim=1)[:, None] * B.double().norm(dim=1)[None, :]
    return float(((D - ref).abs() / scale).max())


@pytest.mark.parametrize("am,bm", [(0, 0), (0, 1), (1, 1), (1, 0)])
@pytest.mark.parametrize("passes", [1, 3])
@pytest.mark.parametrize("M,N,K", [(128, 16, 32), (128, 208, 64), (300, 200, 224), (77, 100, 45 * 4)])
def test_gemm_matches_fp64(fdsm, am, bm, passes, M, N, K):
    assert _run(fdsm, M, N, K, am, bm, passes) < TOL[passes]


@pytest.mark.parametrize("am,bm", [(0, 0), (0, 1), (1, 1), (1, 0)])
@pytest.mark.parametrize("M,N,K", [(128, 16, 32), (128, 208, 64), (300, 200, 224), (77, 100, 45 * 4), (256, 256, 32 * 9)])
def test_raw_operand_gemm_matches_fp64_and_presplit(fdsm, am, bm, M, N, K):
    """In-kernel hi/lo split (shared-memory converter warps): fp32-grade vs fp64, and equal to the pre-split kernel up to
    fp32 rounding (same MMA order; hi = truncation in-kernel, round-to-nearest in the pre-split format)."""
    assert _run(fdsm, M, N, K, am, bm, 3, raw=True) < TOL[3]
    a = _run(fdsm, M, N, K, am, bm, 3, raw=True, want_out=True)
    b = _run(fdsm, M, N, K, am, bm, 3, raw=False, want_out=True)
    assert float((a - b).abs().max()) <= 4e-6 * (K ** 0.5) * 4      # |a_m||b_n| ~ K for N(0,1) operands; TOL-grade


def test_raw_operand_gemm_config_shapes(fdsm):
    assert _run(fdsm, 12288, 200, 800, 0, 1, 3, raw=True) < TOL[3]
    assert _run(fdsm, 12288, 200, 208, 0, 0, 3, raw=True) < TOL[3]
    assert _run(fdsm, 800, 200, 12288, 1, 1, 3, splits=21, raw=True) < TOL[3]
    assert _run(fdsm, 201, 200, 12288, 1, 1, 3, splits=74, raw=True) < TOL[3]


def test_gemm_split_k_and_config_shapes(fdsm):
    # cfg2 forward layer 1 (augmented K = 800), dgrad (K = 208) and wgrad (contraction over 3B rows) shapes
    assert _run(fdsm, 12288, 200, 800, 0, 1, 3) < TOL[3]
    assert _run(fdsm, 12288, 200, 208, 0, 0, 3) < TOL[3]
    assert _run(fdsm, 800, 200, 12288, 1, 1, 3, splits=21) < TOL[3]
    assert _run(fdsm, 300, 100, 200, 0, 0, 3, splits=3) < TOL[3]


def test_tf32_split_is_exact_and_padded(fdsm):
    dev = torch.device("cuda:0")
    x = torch.randn(37, 11, device=dev)
    hi, lo = fdsm.ops.tf32_split(x, out_cols=16, ones_col=True)
    assert hi.shape == (37, 16)
    # hi + lo reproduces x to 2^-21 relative; both halves have the low 13 mantissa bits clear
    assert float(((hi[:, :11] + lo[:, :11]) - x).abs().max() / x.abs().max()) < 2 ** -20
    assert int((hi.view(torch.int32) & 0x1FFF).abs().max()) == 0 and int((lo.view(torch.int32) & 0x1FFF).abs().max()) == 0
    assert torch.all(hi[:, 11] == 1) and torch.all(hi[:, 12:] == 0) and torch.all(lo[:, 11:] == 0)


def test_gemm_argument_errors(fdsm):
    dev = torch.device("cuda:0")
    a = torch.zeros(128, 32, device=dev)
    with pytest.raises(fdsm._lib.FdsmError):
        fdsm.ops.tc_gemm(a, a, a, None, 128, 128, 32, passes=3)             # lo halves: both operands or neither
    with pytest.raises(fdsm._lib.FdsmError):
        fdsm.ops.tc_gemm(a, a, a, a, 128, 128, 32, passes=3, splits=5)     # more splits than k-blocks


def test_cta_pair_kernel_in_a_subprocess():
    """The opt-in cta_group::2 flavour of the raw-operand GEMM (FDSM_GEMM_PAIR=1 is read once per process): same parity
    bars on the config shapes and the ragged ones, odd tile counts included (padded pair)."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, FDSM_GEMM_PAIR="1")
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", __file__, "-k", "raw_operand"], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
