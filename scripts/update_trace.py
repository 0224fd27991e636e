"""Timeline of the fused reduce + clip + Adam launch (jet_update_kernel, csrc/fd_jet_host.inc) from its in-kernel
%globaltimer stamps, at config 2, inside an un-graphed training step.   FDSM_CHAIN_TRACE_MODE=9 python scripts/update_trace.py"""
import os
import sys

import torch

os.environ["FDSM_CHAIN_TRACE_MODE"] = "9"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = "cuda:0"
net = fdsm.mlp.FusedEnergyMLP(784, 200, 3, act="nsoftplus", head="quadratic", device=dev, seed=1, formulation="jet")
net.configure_optimizer(lr=1e-5, betas=(0.9, 0.95), clip=0.1, seed=1, use_graph=False)
xs = [torch.rand(B, 784, device=dev) for _ in range(4)]
for i in range(8):
    net.fd_ssm_step(xs[i % 4], eps=0.05)
tiles = net._step_state["offsets"][7]
buf = torch.zeros(max(tiles * 38, 1024 * 6), dtype=torch.int64, device=dev)
ops.CHAIN_TRACE = buf
for i in range(3):          # the last one is read
    buf.zero_()
    net.fd_ssm_step(xs[i % 4], eps=0.05)
torch.cuda.synchronize()
ops.CHAIN_TRACE = None
t = buf.cpu().double().view(-1, 6)
t = t[t[:, 0] > 0]
t0 = t[:, 1].min()          # first block past griddepcontrol.wait
rel = (t - t0) / 1e3
names = ["block resident", "past griddepcontrol.wait", "partials reduced, gradient written", "past the grid barrier",
         "clip coefficient known", "Adam done"]
print(f"B={B}: {t.shape[0]} blocks; us after the first block left griddepcontrol.wait (mean / min / max over blocks)")
for i, n in enumerate(names):
    print(f"   {n:38s} {rel[:, i].mean():7.2f} / {rel[:, i].min():7.2f} / {rel[:, i].max():7.2f}")
