"""Bring-up aid for the fused MLP: forward/backward vs an fp64 torch copy, and the fused FD-SSM step vs the oracle."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402
from oracle import fd_losses, nets  # noqa: E402

dev = torch.device("cuda:0")


def relv(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300))


def check(D, H, L, act, head, rows, seed=0):
    if head == "quadratic":
        ref = nets.QuadraticMLP(D, H, L, act)
    else:
        ref = nets.SoftplusMLP(D, H)
    nets.init_from_numpy(ref, seed)
    ref = ref.double()
    net = fdsm.mlp.FusedEnergyMLP(D, H, L, act=act, head=head, device=dev)
    net.load_reference_state_dict([p.detach().float() for p in ref.parameters()])
    x = nets.np_rand(seed + 1, rows, D)
    E = net(x.float().to(dev))
    Er = ref(x)
    g = nets.np_randn(seed + 2, rows)
    E.backward(g.float().to(dev))
    Er.backward(g)
    gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
    print(f"D={D} H={H} L={L} {act} {head} rows={rows}:  E rel {relv(E.cpu(), Er):.2e}   grad rel {relv(net.flat.grad.cpu(), gr):.2e}")
    views = net.named_views(net.flat.grad)
    o = 0
    for (name, shape), p in zip(net.param_shapes(), ref.parameters()):
        print(f"    {name:12s} {relv(views[name].cpu().reshape(-1), p.grad.reshape(-1)):.2e}", end="")
    print()
    return net, ref


check(64, 48, 3, "nsoftplus", "quadratic", 96)
check(11, 128, 2, "softplus", "linear", 192)
check(100, 40, 3, "elu", "quadratic", 48)
net, ref = check(784, 200, 3, "nsoftplus", "quadratic", 12288)

# fused FD-SSM step vs the fp64 oracle (same injected direction)
B, D, eps = 4096, 784, 0.05
x = nets.np_rand(5, B, D)
u = nets.np_randn(6, B, D)
for p in ref.parameters():
    p.grad = None
v = fd_losses.sphere_direction(u, eps)
lr = fd_losses.fd_ssm(ref, x, v, 1 / eps ** 2)
lr.backward()
gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
loss, grad = net.fd_ssm_loss_and_grad(x.float().to(dev), eps, 1 / eps ** 2, u=u.float().to(dev))
print(f"fused FD-SSM cfg2: loss {float(loss):.6e} vs fp64 {float(lr):.6e} rel {abs(float(loss) - float(lr)) / abs(float(lr)):.2e};"
      f" grad rel {relv(grad.cpu(), gr):.2e}")
# the torch fp32 path on the same inputs for scale
ref32 = nets.init_from_numpy(nets.QuadraticMLP(784, 200, 3, "nsoftplus"), 0).to(dev)
torch.backends.cuda.matmul.allow_tf32 = False
l32 = fd_losses.fd_ssm(ref32, x.float().to(dev), fd_losses.sphere_direction(u.float().to(dev), eps), 1 / eps ** 2)
l32.backward()
g32 = torch.cat([p.grad.reshape(-1) for p in ref32.parameters()])
print(f"torch fp32 (cuBLAS) : loss {float(l32):.6e} rel {abs(float(l32) - float(lr)) / abs(float(lr)):.2e}; grad rel {relv(g32.cpu(), gr):.2e}")
print("done")
