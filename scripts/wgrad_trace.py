"""Per-CTA phase timeline of the three weight-gradient GEMMs of one chain-path step (tc::gemm_wide_kernel's %globaltimer stamps),
at config 2.    python scripts/wgrad_trace.py"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200 as fdsm  # noqa: E402

dev = "cuda:0"
lib = fdsm._lib.load()
lib.fdsm_debug_set_gemm_trace.argtypes = [C.c_void_p]
lib.fdsm_debug_set_gemm_trace.restype = None
net = fdsm.mlp.FusedEnergyMLP(784, 200, 3, act="nsoftplus", head="quadratic", device=dev, seed=1, formulation="jet")
x = torch.rand(4096, 784, device=dev)
for _ in range(3):
    net.fd_ssm_loss_and_grad(x, 0.05, 400.0)
buf = torch.zeros(3, 160 * 8, dtype=torch.int64, device=dev)
torch.cuda.synchronize()
lib.fdsm_debug_set_gemm_trace(C.c_void_p(buf.data_ptr()))
net.fd_ssm_loss_and_grad(x, 0.05, 400.0)
torch.cuda.synchronize()
lib.fdsm_debug_set_gemm_trace(None)
t = buf.cpu().double().view(3, 160, 8)
t0 = t[t > 0].min()
names = ["CTA start", "set-up done", "first stage landed", "last stage landed", "accumulator ready", "epilogue (warp 4) done", "CTA done"]
for l in range(3):
    tl = t[l]
    live = tl[:, 0] > 0
    r = (tl[live] - t0) / 1e3
    print(f"wgrad launch {l}: {int(live.sum())} CTAs; us after the first CTA of the three launches started (mean / max over CTAs)")
    for i, n in enumerate(names):
        print(f"   {n:26s} {r[:, i].mean():7.2f} / {r[:, i].max():7.2f}")
