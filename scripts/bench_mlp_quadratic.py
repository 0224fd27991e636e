"""DeepEBM's large energy MLP as the tensor-core stress shape: MLP_Quadratic 1024 -> 5000 x 5
(DeepEBM/ESM/models.py:129-160), FD-SSM loss + gradient at B = 128 (384 stacked rows), fused path.
Prints ms per loss+gradient evaluation and the algorithmic / executed TFLOP/s of the GEMMs."""
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200 as fdsm  # noqa: E402

dev = torch.device("cuda:0")
B, D, H, L = (int(a) for a in (sys.argv[1:5] + ["128", "1024", "5000", "5"][len(sys.argv) - 1:]))
net = fdsm.mlp.FusedEnergyMLP(D, H, L, act="nsoftplus", head="quadratic", device=dev, seed=1)
xs = [torch.rand(B, D, device=dev) for _ in range(4)]
for i in range(3):
    loss, grad = net.fd_ssm_loss_and_grad(xs[i % 4], 0.05, 1 / 0.05 ** 2)
torch.cuda.synchronize()
assert torch.isfinite(loss) and torch.isfinite(grad).all()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
N = 20
e0.record()
for i in range(N):
    net.fd_ssm_loss_and_grad(xs[i % 4], 0.05, 1 / 0.05 ** 2)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / N
rows = 3 * B
fwd = 2.0 * rows * (D * H + (L - 1) * H * H)
total = 3 * fwd - 2.0 * rows * D * H          # fwd + wgrad (= fwd) + dgrad (= fwd - first layer)
print(f"MLP_Quadratic {D}->{H}x{L}, B={B} ({rows} rows, {net.flat.numel() / 1e6:.1f} M params): {ms:.3f} ms per loss+grad, "
      f"{total / ms / 1e9:.1f} TFLOP/s algorithmic, {3 * total / ms / 1e9:.1f} executed (3xTF32), "
      f"{B / ms * 1e3:.0f} samples/s; loss {float(loss):+.4e}")
