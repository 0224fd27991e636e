"""Is the config-2 step GPU- or host-bound?  Times (a) FDSSMTrainer.step, (b) bare graph replays, (c) host time per call."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200.train import FDSSMTrainer  # noqa: E402

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
net = fdsm.mlp.FusedEnergyMLP(784, 200, 3, act="nsoftplus", head="quadratic", device=dev, seed=1234)
tr = FDSSMTrainer(net, eps=0.05, lr=1e-5, betas=(0.9, 0.95), clip=0.1, seed=1234)
pool = [torch.rand(B, 784, device=dev) for _ in range(16)]
for i in range(20):
    tr.step(pool[i % 16])
torch.cuda.synchronize()
N = 400
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for i in range(N):
    tr.step(pool[i % 16])
t1 = time.perf_counter(); e1.record(); torch.cuda.synchronize()
print(f"B={B}: tr.step        GPU {e0.elapsed_time(e1) * 1e3 / N:7.1f} us/step   host issue {1e6 * (t1 - t0) / N:7.1f} us/step")
g1 = list(net._graphs.values())[0][0]
e0.record(); t0 = time.perf_counter()
for i in range(N):
    g1.replay()
t1 = time.perf_counter(); e1.record(); torch.cuda.synchronize()
print(f"B={B}: graph replay   GPU {e0.elapsed_time(e1) * 1e3 / N:7.1f} us/step   host issue {1e6 * (t1 - t0) / N:7.1f} us/step")
x = net._step_state["x"]
e0.record()
for i in range(N):
    x.copy_(pool[i % 16])
e1.record(); torch.cuda.synchronize()
print(f"B={B}: x.copy_ alone  GPU {e0.elapsed_time(e1) * 1e3 / N:7.1f} us")
