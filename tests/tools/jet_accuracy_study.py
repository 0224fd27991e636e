"""Accuracy of the difference-propagating evaluation (oracle/jet.py) at the cfg2 shape: fp64 stacked reference vs
fp32 stacked (what the reference runs) vs fp32 jets vs fp32 jets with TF32-rounded GEMM operands (what the chain kernel
runs).  CPU only.   python tests/tools/jet_accuracy_study.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import fd_losses, jet, nets  # noqa: E402


def rel(a, b):
    return float((a - b).norm() / b.norm())


def flat_grad(net):
    return torch.cat([p.grad.reshape(-1).double() for p in net.parameters()])


def run(B, D, H, L, act, eps, scale, kind="quad", seed=1234):
    def mk(dtype):
        net = nets.QuadraticMLP(D, H, L, act) if kind == "quad" else nets.SoftplusMLP(D, H)
        return nets.init_from_numpy(net, seed).to(dtype)
    x = nets.np_rand(seed + 2, B, D) if kind == "quad" else nets.np_randn(seed + 2, B, D) + 5.0
    u = nets.np_randn(seed + 3, B, D)
    out = {}
    for name, dtype, fn in [("f64 stacked", torch.float64, lambda n, x, v: fd_losses.fd_ssm(n, x, v, scale)),
                            ("f64 jet", torch.float64, lambda n, x, v: jet.fd_ssm_jet(n, x, v, scale)),
                            ("f32 stacked", torch.float32, lambda n, x, v: fd_losses.fd_ssm(n, x, v, scale)),
                            ("f32 jet", torch.float32, lambda n, x, v: jet.fd_ssm_jet(n, x, v, scale)),
                            ("f32 jet tf32", torch.float32, lambda n, x, v: jet.fd_ssm_jet(n, x, v, scale, tf32=True))]:
        net = mk(dtype)
        xx, uu = x.to(dtype), u.to(dtype)
        v = fd_losses.sphere_direction(uu, eps)
        loss = fn(net, xx, v)
        loss.backward()
        out[name] = (float(loss), flat_grad(net))
    l64, g64 = out["f64 stacked"]
    print(f"B={B} D={D} H={H} L={L} act={act} eps={eps}: loss(f64) = {l64:.8e}")
    for name, (l, g) in out.items():
        print(f"  {name:14s} loss rel err {abs(l - l64) / abs(l64):9.2e}   grad rel err {rel(g, g64):9.2e}")


if __name__ == "__main__":
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    run(B, 784, 200, 3, "nsoftplus", 0.05, 1 / 0.05 ** 2)
    run(B, 784, 200, 3, "elu", 0.05, 1 / 0.05 ** 2)
    run(min(B, 1024), 11, 128, 2, "softplus", 0.1, 11 / 0.1 ** 2, kind="sp")
    run(32, 64, 48, 3, "nsoftplus", 0.05, 1 / 0.05 ** 2)
