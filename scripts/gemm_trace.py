"""Per-CTA phase timeline of the raw-operand GEMM and of the MLP step's GEMM launches (debug hook, %globaltimer)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.load()
lib.fdsm_debug_set_gemm_trace.argtypes = [C.c_void_p]
lib.fdsm_debug_set_gemm_trace.restype = None
NAMES = ["start->setup", "setup->first conv", "first->last conv", "last conv->acc done", "acc->epi done", "epi->all done"]


def trace(tag, M, N, K, am, bm, splits=1, tile_n=208):
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev)
    ah = ops.tf32_split(A if am == 0 else A.t().contiguous(), raw=True)[0]
    bh = ops.tf32_split(B if bm == 0 else B.t().contiguous(), raw=True)[0]
    nct = ((M + 127) // 128) * ((N + tile_n - 1) // tile_n) * splits
    buf = torch.zeros(nct * 8, dtype=torch.int64, device=dev)
    for it in range(3):
        flush = torch.empty(64 << 20, device=dev).fill_(1.0)   # 256 MB: push the operands out of L2
        if it == 2:
            lib.fdsm_debug_set_gemm_trace(C.c_void_p(buf.data_ptr()))
        ops.tc_gemm(ah, None, bh, None, M, N, K, a_major=am, b_major=bm, passes=3, splits=splits, tile_n=tile_n)
        torch.cuda.synchronize()
    lib.fdsm_debug_set_gemm_trace(None)
    t = buf.view(nct, 8).double().cpu()
    t0 = t[:, 0].min()
    d = t[:, 1:7] - t[:, 0:6]
    print(f"{tag}: {nct} CTAs, kernel span {float(t[:, 6].max() - t0) / 1e3:.2f} us; CTA start spread {float(t[:, 0].max() - t0) / 1e3:.2f} us")
    for i, n in enumerate(NAMES):
        print(f"    {n:22s} mean {float(d[:, i].mean()) / 1e3:7.2f} us   max {float(d[:, i].max()) / 1e3:7.2f}")


trace("fwd L1 (cold L2)", 12288, 200, 785, 0, 1)
trace("fwd L2", 12288, 200, 201, 0, 1)
trace("dgrad", 12288, 200, 208, 0, 0)
trace("wgrad L1", 785, 208, 12288, 1, 1, splits=21)
trace("wgrad L2", 201, 208, 12288, 1, 1, splits=74)
