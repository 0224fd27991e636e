"""Bring-up aid for the tcgen05 GEMM: runs a few shapes/layouts and prints where the result is wrong."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)


def run(M, N, K, am, bm, passes, splits=1, tile_n=0):
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev)
    Kp = (K + 3) // 4 * 4
    Ast = A if am == 0 else A.t().contiguous()
    Bst = B if bm == 0 else B.t().contiguous()
    ah, al = ops.tf32_split(Ast)
    bh, bl = ops.tf32_split(Bst)
    out = ops.tc_gemm(ah, al, bh, bl, M, N, K, a_major=am, b_major=bm, passes=passes, splits=splits, tile_n=tile_n)
    torch.cuda.synchronize()
    D = out.double().sum(0)
    ref = A.double() @ B.double().t()
    scale = (A.double().norm(dim=1)[:, None] * B.double().norm(dim=1)[None, :])
    err = ((D - ref).abs() / scale)
    tag = f"M={M} N={N} K={K} a_major={am} b_major={bm} passes={passes} splits={splits}"
    print(f"{tag:70s} max rel err {float(err.max()):.3e}  mean {float(err.mean()):.3e}")
    if float(err.max()) > (1e-5 if passes == 3 else 5e-3):
        bad = err > (1e-5 if passes == 3 else 5e-3)
        rows = bad.any(1).nonzero().flatten()
        cols = bad.any(0).nonzero().flatten()
        print("   bad rows:", rows[:16].tolist(), "... n=", rows.numel(), " bad cols:", cols[:16].tolist(), "... n=", cols.numel())
        print("   D[0,:4]", D[0, :4].tolist(), " ref[0,:4]", ref[0, :4].tolist())
    return float(err.max())


which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("kk", "all"):
    run(128, 16, 32, 0, 0, 1)
    run(128, 208, 32, 0, 0, 1)
    run(128, 208, 64, 0, 0, 3)
    run(256, 200, 800, 0, 0, 3)
    run(300, 100, 200, 0, 0, 3, splits=3)
if which in ("kmn", "all"):
    run(128, 32, 32, 0, 1, 1)
    run(128, 208, 64, 0, 1, 3)
    run(256, 200, 224, 0, 1, 3)
if which in ("mnmn", "all"):
    run(128, 32, 32, 1, 1, 1)
    run(128, 208, 64, 1, 1, 3)
    run(800, 200, 12288, 1, 1, 3, splits=8)
if which in ("mnk", "all"):
    run(128, 32, 32, 1, 0, 1)
    run(200, 208, 96, 1, 0, 3)
print("done")
