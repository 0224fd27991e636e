"""Print, for every fixture case, err(ours, fp64) next to err(reference fp32, fp64) (GPU box)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import fdsm_b200 as fdsm  # noqa: E402
from oracle import cases  # noqa: E402
from test_losses_gpu import _rel, _relvec, run_adapter  # noqa: E402

golden = np.load(os.path.join(ROOT, "tests", "golden", "reference_outputs.npz"))
dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = False
print(f"{'case':18s} {'loss':>13s} {'ours-vs-f64':>11s} {'ref32-vs-f64':>12s} {'ours-vs-r32':>11s} | "
      f"{'g ours/f64':>10s} {'g r32/f64':>10s} {'g ours/r32':>10s}")
for name in [n for n in cases.CASES if n not in cases.LATE]:   # the late cases: tests/test_zz_late_cases_gpu.py
    c = cases.build(name, torch.float32)
    loss = run_adapter(fdsm, c, dev)
    loss.backward()
    g = cases.flat_grad(c["net"]).cpu().numpy()
    l32, l64 = golden[f"{name}/loss_f32"], golden[f"{name}/loss_f64"]
    g32, g64 = golden[f"{name}/grad_f32"], golden[f"{name}/grad_f64"]
    print(f"{name:18s} {float(loss):+.6e} {_rel(loss, l64):11.2e} {_rel(l32, l64):12.2e} {_rel(loss, l32):11.2e} | "
          f"{_relvec(g, g64):10.2e} {_relvec(g32, g64):10.2e} {_relvec(g, g32):10.2e}")
