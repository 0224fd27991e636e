"""Times fdsm_tc_gemm on the config-2 step's GEMM shapes (graph-replayed launches, CUDA events)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
REPS = 20


def bench(tag, M, N, K, am, bm, passes, splits=1, tile_n=0, raw=False):
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev)
    Ast = A if am == 0 else A.t().contiguous()
    Bst = B if bm == 0 else B.t().contiguous()
    ah, al = ops.tf32_split(Ast)
    bh, bl = ops.tf32_split(Bst)
    if raw:   # raw fp32 operands, lo halves generated in the kernel
        ah, al, bh, bl = ops.tf32_split(Ast, want_lo=False)[0], None, ops.tf32_split(Bst, want_lo=False)[0], None
        ah.copy_(torch.nn.functional.pad(Ast, (0, ah.shape[1] - Ast.shape[1])))
        bh.copy_(torch.nn.functional.pad(Bst, (0, bh.shape[1] - Bst.shape[1])))
    f = lambda: ops.tc_gemm(ah, al, bh, bl, M, N, K, a_major=am, b_major=bm, passes=passes, splits=splits, tile_n=tile_n)
    out = f()
    torch.cuda.synchronize()
    ref = A.double() @ B.double().t()
    scale = (A.double().norm(dim=1)[:, None] * B.double().norm(dim=1)[None, :])
    err = float(((out.double().sum(0) - ref).abs() / scale).max())
    # keep `out` allocations out of the graph: call the C entry directly into a fixed buffer
    lib = _lib.load()
    g = _lib.GemmArgs()
    p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
    g.a_hi, g.a_lo, g.lda, g.a_major = p(ah), p(al), ah.stride(0), am
    g.b_hi, g.b_lo, g.ldb, g.b_major = p(bh), p(bl), bh.stride(0), bm
    g.M, g.N, g.K, g.passes, g.splits, g.tile_n = M, N, K, passes, splits, tile_n
    g.out, g.ldo, g.split_stride = p(out), N, M * N
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(REPS):
                _lib.check(lib.fdsm_tc_gemm(C.byref(g), C.c_void_p(torch.cuda.current_stream().cuda_stream)), "gemm")
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (5 * REPS)
    tf = 2.0 * M * N * K / us / 1e6
    print(f"{tag:28s} M={M:6d} N={N:4d} K={K:6d} maj={am}{bm} passes={passes} splits={splits:3d} tile_n={tile_n:3d} raw={int(raw)} "
          f"{us:8.2f} us  {tf:7.1f} TF/s alg  {tf * passes:7.1f} exec   err {err:.2e}", flush=True)


RAW = [False] + ([True] if os.environ.get("FDSM_RAW", "0") == "1" else [])
for raw in RAW:
    for passes in (3, 1):
        bench("fwd L1", 12288, 200, 785, 0, 1, passes, tile_n=208, raw=raw)
        bench("fwd L2", 12288, 200, 201, 0, 1, passes, tile_n=208, raw=raw)
        bench("dgrad", 12288, 200, 208, 0, 0, passes, tile_n=208, raw=raw)
        bench("wgrad L1", 785, 208, 12288, 1, 1, passes, splits=21, tile_n=208, raw=raw)
        bench("wgrad L2", 201, 208, 12288, 1, 1, passes, splits=74, tile_n=208, raw=raw)
        bench("wgrad L2 s37", 201, 208, 12288, 1, 1, passes, splits=37, tile_n=208, raw=raw)
    bench("fwd L1 tile112", 12288, 200, 785, 0, 1, 3, tile_n=112, raw=raw)
    bench("fwd L2 tile112", 12288, 200, 201, 0, 1, 3, tile_n=112, raw=raw)
    bench("square 4096", 4096, 4096, 4096, 0, 0, 3, tile_n=256, raw=raw)
    bench("square 4096 p1", 4096, 4096, 4096, 0, 0, 1, tile_n=256, raw=raw)
