"""Run each FD kernel a few times at config and saturating sizes (for ncu and for CUDA-graph timing)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
graph_timing = "--graph" in sys.argv


def timed(name, fn, alg_bytes, iters=20):
    """Replay `iters` launches inside one CUDA graph between two events: launch-gap free kernel time."""
    fn()
    torch.cuda.synchronize()
    if not graph_timing:
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        return
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(iters):
                fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    print(f"{name:55s} {us:9.2f} us  {alg_bytes / us / 1e3:8.1f} GB/s (algorithmic)")


for B, D, tag in ((4096, 784, "cfg2"), (65536, 784, "sat"), (128, 3072, "cfg5"), (16384, 3072, "sat")):
    # rotate over enough distinct inputs to exceed L2 (126 MB) at the saturating sizes
    nbuf = max(2, min(8, int(300e6 // (B * D * 4)) + 1))
    xs = [torch.rand(B, D, device=dev) for _ in range(nbuf)]
    out3 = [torch.empty(3 * B, D, device=dev) for _ in range(2)]
    out2 = [torch.empty(2 * B, D, device=dev) for _ in range(2)]
    it = [0]

    def stack3():
        it[0] += 1
        ops.stack(xs[it[0] % nbuf], row_mask=7, scale=0.05, out=out3[it[0] % 2], draw=(1, it[0], 0))

    timed(f"stack3 philox   B={B} D={D} ({tag})", stack3, 16.0 * B * D)
    table = torch.exp(torch.linspace(3.9, -4.6, 232)).to(dev)
    labels = torch.randint(0, 232, (B,), device=dev)

    def stack2():
        it[0] += 1
        ops.stack(xs[it[0] % nbuf], row_mask=6, scale=0.1, sigma=table, labels=labels, out=out2[it[0] % 2],
                  draw=(1, it[0], 0))

    timed(f"stack2 noised philox B={B} D={D} ({tag})", stack2, 12.0 * B * D)
    v = torch.randn(B, D, device=dev) * 0.01
    ss = [torch.randn(2 * B, D, device=dev) for _ in range(max(2, nbuf // 2))]

    def comb():
        it[0] += 1
        ops.combine_score_fused(ss[it[0] % len(ss)], v, k=D / 0.02, sigma=table, labels=labels, power=2.0)

    timed(f"combine_score fused fwd+bwd B={B} D={D} ({tag})", comb, 20.0 * B * D)
