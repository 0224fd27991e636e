"""Per-phase timeline of the chain kernel (csrc/fd_chain.cuh) from its in-kernel %globaltimer stamps, at config 2.
    python scripts/chain_trace.py [B] [passes]      (env FDSM_JET_TILE = samples per tile)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = "cuda:0"
net = fdsm.mlp.FusedEnergyMLP(784, 200, 3, act="nsoftplus", head="quadratic", device=dev, seed=1, formulation="jet", jet_passes=passes)
x = torch.rand(B, 784, device=dev)
for _ in range(3):
    net.fd_ssm_loss_and_grad(x, 0.05, 400.0)
st = net._step_state
tiles, L = st["offsets"][7], 3
ops.CHAIN_TRACE = torch.zeros(tiles, 38, dtype=torch.int64, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
net.fd_ssm_loss_and_grad(x, 0.05, 400.0)
e1.record()
torch.cuda.synchronize()
t = ops.CHAIN_TRACE.cpu().double()
ops.CHAIN_TRACE = None
t0 = t[:, 0].min()
rel = (t - t0) / 1e3
print(f"B={B} passes={passes} S={st['offsets'][0]} tiles={tiles} cluster={st['offsets'][9]}  loss+grad call {e0.elapsed_time(e1) * 1e3:.1f} us")
print(f"CTA start  mean {rel[:, 0].mean():6.2f} max {rel[:, 0].max():6.2f}   setup done mean {rel[:, 1].mean():6.2f}")
names = [f"fwd{g}" for g in range(L)] + [f"dgrad{2 * L - 1 - g}" for g in range(L, 2 * L - 1)]
prev = rel[:, 1]
print("phase     acc-ready(us)  since prev epilogue end   dump    epilogue(w3)   mma-issued(us)   | mean over CTAs; max of epilogue end")
for g, n in enumerate(names):
    a, d, j, m = rel[:, 2 + 4 * g], rel[:, 3 + 4 * g], rel[:, 4 + 4 * g], rel[:, 5 + 4 * g]
    print(f"{n:8s} {a.mean():10.2f} {(a - prev).mean():16.2f} {(d - a).mean():14.2f} {(j - d).mean():10.2f} {m.mean():16.2f}   | end max {j.max():.2f}")
    prev = j
print(f"kernel span (first CTA start -> last epilogue done): {rel[:, 4 + 4 * (2 * L - 2)].max():.2f} us")
mode = int(os.environ.get("FDSM_CHAIN_TRACE_MODE", "0"))
if L == 3 and mode == 0:
    a0 = rel[:, 2]          # fwd0 accumulator ready
    print("fwd1, MMA thread, per k-block (us after fwd0's accumulator was ready; mean over CTAs):  operand block published | its weights landed")
    print("   " + "  ".join(f"kb{kb}: {(rel[:, 22 + 2 * kb] - a0).mean():5.2f} | {(rel[:, 23 + 2 * kb] - a0).mean():5.2f}" for kb in range(7)))
    print(f"   MMAs issued {(rel[:, 5 + 4] - a0).mean():5.2f}   accumulator ready {(rel[:, 2 + 4] - a0).mean():5.2f}")
elif L == 3 and mode == 3:
    s1 = rel[:, 1]          # set-up done
    print("fwd0, MMA thread: operands of k-blocks 0, 2, 4, ... landed (us after set-up; mean over CTAs)")
    print("   " + "  ".join(f"{(rel[:, 22 + u] - s1).mean():5.2f}" for u in range(13)))
elif L == 3:
    d0 = rel[:, 3]          # fwd0's accumulator dumped: the epilogue warps start together
    what = "whole epilogue" if mode == 1 else "block 1 published"
    print(f"fwd1's epilogue, per warp: {what} (us after the dump; mean / max over CTAs)")
    print("   " + "  ".join(f"w{w}: {(rel[:, 22 + w] - d0).mean():4.2f}/{(rel[:, 22 + w] - d0).max():4.2f}" for w in range(16)))
