"""Fused-MLP FD-SSM step at config 2 vs the fp64 oracle, with torch fp32 (cuBLAS SGEMM) as the yardstick (GPU box)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from oracle import fd_losses, nets  # noqa: E402

DEV = "cuda:0"
torch.backends.cuda.matmul.allow_tf32 = False


def relv(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


for act in ("nsoftplus", "softplus", "elu"):
    for seed in (0, 1):
        B, D, eps = 4096, 784, 0.05
        ref = nets.init_from_numpy(nets.QuadraticMLP(D, 200, 3, act), seed)
        net = fdsm.mlp.FusedEnergyMLP(D, 200, 3, act=act, head="quadratic", device=DEV)
        net.load_reference_state_dict([p.detach().float() for p in ref.parameters()])
        ref = ref.double()
        x, u = nets.np_rand(5 + seed, B, D), nets.np_randn(6 + seed, B, D)
        lr = fd_losses.fd_ssm(ref, x, fd_losses.sphere_direction(u, eps), 1 / eps ** 2)
        lr.backward()
        gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
        loss, grad = net.fd_ssm_loss_and_grad(x.float().to(DEV), eps, 1 / eps ** 2, u=u.float().to(DEV))
        r32 = nets.init_from_numpy(nets.QuadraticMLP(D, 200, 3, act), seed).to(DEV)
        l32 = fd_losses.fd_ssm(r32, x.float().to(DEV), fd_losses.sphere_direction(u.float().to(DEV), eps), 1 / eps ** 2)
        l32.backward()
        g32 = torch.cat([p.grad.reshape(-1) for p in r32.parameters()])
        E = net(x.float().to(DEV))
        print(f"{act:10s} seed {seed}: loss {float(loss):+.6e}  ours/f64 {abs(float(loss) - float(lr)) / abs(float(lr)):.2e}  "
              f"torch32/f64 {abs(float(l32) - float(lr)) / abs(float(lr)):.2e} | grad ours/f64 {relv(grad, gr):.2e}  "
              f"torch32/f64 {relv(g32, gr):.2e} | E ours/f64 {relv(E, ref(x)):.2e} torch32/f64 {relv(r32(x.float().to(DEV)), ref(x)):.2e}",
              flush=True)
