"""FD-DSM (DeepEBM train_dsm.py, loss mdsm_fd) training-step throughput on one B200 at the config-2 shape
(B = 4096, D = 784, MLP_Quadratic_MNIST 784-200-200-200) and at DeepEBM's default batch (B = 128):
  (a) fused path (ours): noised 2-row stack -> tcgen05 MLP fwd -> DSM combine -> MLP bwd -> clip -> Adam, one CUDA-graph replay
  (b) generic path: our FD kernels (stack2-noised, DSM combine) around a torch nn.Linear MLP (cuBLAS fp32), torch clip + Adam
Inputs rotate over a pool larger than L2; CUDA events on the launching stream.
Activation = bench.py's FDSM_BENCH_ACT (nsoftplus default, elu = configs/default.yaml).
    python scripts/bench_fused_dsm.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200 as fdsm  # noqa: E402
from bench import TorchQuadraticMLP, synth_batch, ACT, D, HID, NHID, EPS, LR, BETAS, CLIP  # noqa: E402
from fdsm_b200.train import FDSSMTrainer  # noqa: E402

dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = False


def timed(step, pool, n=100, warm=10):
    for i in range(warm):
        step(pool[i % len(pool)])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        step(pool[i % len(pool)])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for B in (4096, 128):
    gen = torch.Generator(device=dev).manual_seed(1234)
    pool = [synth_batch(gen, B, dev) for _ in range(16 if B == 4096 else 64)]
    # train_dsm.py:50-57, noise_distribution 'exp' between min_noise and max_noise
    sig = torch.tensor(np.logspace(np.log10(0.05), np.log10(1.2), B), dtype=torch.float32, device=dev).view(-1, 1)
    fused = fdsm.mlp.FusedEnergyMLP(D, HID, NHID, act=ACT, head="quadratic", device=dev, seed=1234)
    tr = FDSSMTrainer(fused, eps=EPS, lr=LR, betas=BETAS, clip=CLIP, seed=1234, loss="mdsm_fd", sigmas=sig)
    t_fused = timed(tr.step, pool)
    net = TorchQuadraticMLP(D, HID, NHID).to(dev)
    tg = FDSSMTrainer(net, eps=EPS, lr=LR, betas=BETAS, clip=CLIP, seed=1234, loss="mdsm_fd", sigmas=sig)
    t_gen = timed(tg.step, pool)
    print(f"FD-DSM step B={B} D={D}: fused (ours) {t_fused * 1e3:8.1f} us = {B / t_fused / 1e3:7.2f} M samples/s | "
          f"generic (torch MLP between our FD kernels) {t_gen * 1e3:8.1f} us = {B / t_gen / 1e3:7.2f} M samples/s | "
          f"speed-up {t_gen / t_fused:.2f}x")
