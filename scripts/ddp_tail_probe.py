"""Timeline of the multi-GPU tail of the step (reduce -> peer exchange -> Adam) from the kernels' own %globaltimer stamps, inside
the step's CUDA graph.   FDSM_CHAIN_TRACE_MODE=9 torchrun --nproc-per-node 2 scripts/ddp_tail_probe.py"""
import os
import sys

import torch
import torch.distributed as dist

os.environ["FDSM_CHAIN_TRACE_MODE"] = "9"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import ops  # noqa: E402
from fdsm_b200.train import FDSSMTrainer  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
net = fdsm.mlp.FusedEnergyMLP(784, 200, 3, act="nsoftplus", head="quadratic", device=dev, seed=1234, jet_passes=4)
tr = FDSSMTrainer(net, eps=0.05, lr=1e-5, betas=(0.9, 0.95), clip=0.1, seed=1234)
x = torch.rand(4096, 784, device=dev)
buf = torch.zeros(1024 * 6, dtype=torch.int64, device=dev)
ops.CHAIN_TRACE = buf                 # baked into the step's graph at capture: every replay stamps
for i in range(60):
    tr.step(x, borrow=True)
torch.cuda.synchronize()
dist.barrier()
for i in range(20):
    tr.step(x, borrow=True)
torch.cuda.synchronize()
t = buf.cpu().double().view(-1, 6)
t = t[t[:, 0] > 0]
ex = tr._allreduce.scratch[192:200].view(torch.int64).cpu().double()
t0 = t[:, 1].min()
if rank == 0:
    r = lambda v: (v - t0) / 1e3
    print(f"reduce kernel ({t.shape[0]} blocks), us after its first block left griddepcontrol.wait: gradient written mean {r(t[:, 2]).mean():.2f} max {r(t[:, 2]).max():.2f}")
    print(f"exchange kernel, block 0: start {r(ex[0]):.2f}  pushed {r(ex[1]):.2f}  reduced {r(ex[2]):.2f}  done {r(ex[3]):.2f}")
    print("(the reduce kernel's later stamps are not taken in its reduce-only mode; Adam follows the exchange)")
dist.destroy_process_group()
