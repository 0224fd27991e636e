"""In-graph timeline of the GEMM launches of one config-2 training step (debug %globaltimer stamps; nsys substitute)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from fdsm_b200 import _lib  # noqa: E402
from fdsm_b200.train import FDSSMTrainer  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.load()
lib.fdsm_debug_set_gemm_trace.argtypes = [C.c_void_p]
lib.fdsm_debug_set_gemm_trace.restype = None
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
net = fdsm.mlp.FusedEnergyMLP(784, 200, 3, act="nsoftplus", head="quadratic", device=dev, seed=1234)
tr = FDSSMTrainer(net, eps=0.05, lr=1e-5, betas=(0.9, 0.95), clip=0.1, seed=1234)
pool = [torch.rand(B, 784, device=dev) for _ in range(16)]
net._opt["use_graph"] = False
for i in range(3):
    tr.step(pool[i])          # warm-up, un-graphed
net._opt["use_graph"] = True
NL = 8
buf = torch.zeros(NL * 8 * 160 * 2, dtype=torch.int64, device=dev)
torch.cuda.synchronize()
lib.fdsm_debug_set_gemm_trace(C.c_void_p(buf.data_ptr()))
tr.step(pool[3])              # warm-up launch inside fd_ssm_step (slots 0..10) + capture (slots 11..21) + first replay
lib.fdsm_debug_set_gemm_trace(None)
for i in range(20):
    tr.step(pool[i % 16])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(100):
    tr.step(pool[i % 16])
e1.record()
torch.cuda.synchronize()
print(f"step {e0.elapsed_time(e1) * 10:.1f} us")
t = buf.view(2 * NL, 160, 8)[NL:].double().cpu()   # the captured launches: stamps of the LAST replay
rows = []
for i in range(NL):
    m = t[i][:, 0] > 0
    if not m.any():
        continue
    ld = m & (t[i][:, 2] > 0)      # CTA pairs: only the leader's MMA thread stamps slots 2 and 3
    rows.append((float(t[i][m][:, 0].min()), float(t[i][m][:, 6].max()), int(m.sum()),
                 float((t[i][ld][:, 2] - t[i][ld][:, 0]).mean()), float((t[i][ld][:, 4] - t[i][ld][:, 2]).mean()),
                 float((t[i][m][:, 5] - t[i][m][:, 4]).mean()), float((t[i][m][:, 6] - t[i][m][:, 5]).mean())))
rows.sort()
T0 = rows[0][0]
prev_end = T0
print("launch  CTAs   start     end     dur   gap_before | per-CTA: to-first-MMA  mainloop  epi-math epi-store (us; * = side stream)")
for i, (a, b, n, f, ml, ep, st) in enumerate(rows):
    print(f"{i:3d} {n:6d} {(a - T0) / 1e3:8.2f} {(b - T0) / 1e3:8.2f} {(b - a) / 1e3:7.2f} {(a - prev_end) / 1e3:9.2f}   | "
          f"{f / 1e3:10.2f} {ml / 1e3:9.2f} {ep / 1e3:9.2f} {st / 1e3:9.2f}")
    prev_end = max(prev_end, b)
