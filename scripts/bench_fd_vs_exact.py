"""Config 2 as BASELINE.json words it: "FD-SSM vs exact SSM on DeepEBM MLP energy, synthetic 784-dim MNIST-shaped data,
batch 4096" -- training-step throughput (loss, backward, clip, Adam) on one B200 for
  (a) FD-SSM, fused path (ours: kernels (1)-(4) + tcgen05 MLP, one CUDA-graph replay)
  (b) FD-SSM, generic path (our FD kernels around a torch nn.Linear MLP: cuBLAS fp32)
  (c) exact SSM  `ssm`     (DeepEBM/ESM/loss_train.py:173-189: two autograd.grad(create_graph=True) passes, torch MLP)
  (d) exact SSM-VR `ssm_vr` (loss_train.py:192-209)
and the loss values of the four on the same batch (ssm_fd uses a unit direction and no D factor, ssm Gaussian projections:
exact = D x FD in expectation)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fdsm_b200 as fdsm  # noqa: E402
from bench import TorchQuadraticMLP, synth_batch, B_PER_GPU, D, HID, NHID, EPS, LR, BETAS, CLIP  # noqa: E402
from fdsm_b200.train import FDSSMTrainer  # noqa: E402

dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = False
gen = torch.Generator(device=dev).manual_seed(1234)
pool = [synth_batch(gen, B_PER_GPU, dev) for _ in range(16)]


def timed(step, n=100, warm=10):
    for i in range(warm):
        step(pool[i % 16])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        step(pool[i % 16])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


fused = fdsm.mlp.FusedEnergyMLP(D, HID, NHID, act="nsoftplus", head="quadratic", device=dev, seed=1234)
tr = FDSSMTrainer(fused, eps=EPS, lr=LR, betas=BETAS, clip=CLIP, seed=1234)
res = {"FD-SSM fused (ours)": timed(tr.step)}
# exact SSM on the same fused net, in forward mode inside the same chain kernel
fused_x = fdsm.mlp.FusedEnergyMLP(D, HID, NHID, act="nsoftplus", head="quadratic", device=dev, seed=1234)
fused_x.configure_optimizer(lr=LR, betas=BETAS, clip=CLIP, seed=1234)
res["exact SSM fused, forward mode (ours)"] = timed(lambda x: fused_x.fd_ssm_step(x, eps=None))

net = TorchQuadraticMLP(D, HID, NHID).to(dev)
with torch.no_grad():   # same weights as the fused net
    views = fused.named_views()
    for l, fc in enumerate(net.fcs):
        fc.weight.copy_(views[f"fc{l + 1}.weight"]); fc.bias.copy_(views[f"fc{l + 1}.bias"])
    for name in ("ln1", "ln2", "lq"):
        getattr(net, name).weight.copy_(views[f"{name}.weight"]); getattr(net, name).bias.copy_(views[f"{name}.bias"])
opt = torch.optim.Adam(net.parameters(), lr=LR, betas=BETAS, fused=True)


def torch_step(loss_fn):
    def step(x):
        opt.zero_grad(set_to_none=True)
        loss = loss_fn(net, x)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(net.parameters(), CLIP)
        opt.step()
    return step


with torch.no_grad():
    pass
x0 = pool[0]
ux = torch.randn(B_PER_GPU, D, device=dev, generator=gen)
fused_chk = fdsm.mlp.FusedEnergyMLP(D, HID, NHID, act="nsoftplus", head="quadratic", device=dev, seed=1234)
vals = {"FD-SSM fused (ours)": float(fused.fd_ssm_loss_and_grad(x0, EPS, 1 / EPS ** 2)[0]),
        "exact SSM fused, forward mode (ours)": float(fused_chk.exact_ssm_loss_and_grad(x0, u=ux)[0]),
        "FD-SSM generic (torch MLP)": float(fdsm.ssm_fd(net, x0, eps=EPS)),
        "exact SSM (torch double backward)": float(fdsm.ssm(net, x0.clone())),
        "exact SSM-VR (torch double backward)": float(fdsm.ssm_vr(net, x0.clone()))}
res["FD-SSM generic (torch MLP)"] = timed(torch_step(lambda n, x: fdsm.ssm_fd(n, x, eps=EPS)), n=50)
res["exact SSM (torch double backward)"] = timed(torch_step(lambda n, x: fdsm.ssm(n, x.clone())), n=50)
res["exact SSM-VR (torch double backward)"] = timed(torch_step(lambda n, x: fdsm.ssm_vr(n, x.clone())), n=50)
# same projections through the reference's double-backward form on the torch twin of the UNTRAINED net: the two exact losses
with torch.no_grad():
    v0 = fused_chk.named_views()
    twin = TorchQuadraticMLP(D, HID, NHID).to(dev)
    for l, fc in enumerate(twin.fcs):
        fc.weight.copy_(v0[f"fc{l + 1}.weight"]); fc.bias.copy_(v0[f"fc{l + 1}.bias"])
    for name in ("ln1", "ln2", "lq"):
        getattr(twin, name).weight.copy_(v0[f"{name}.weight"]); getattr(twin, name).bias.copy_(v0[f"{name}.bias"])
ref_exact = float(fdsm.ssm(twin, x0.clone(), u=ux))
print(f"exact SSM on the same batch and projections, untrained net: forward mode (ours) {vals['exact SSM fused, forward mode (ours)']:+.6e}  "
      f"torch double backward {ref_exact:+.6e}")
print(f"config 2: MLP_Quadratic_MNIST(nsoftplus) {D}-{HID}x{NHID}, B={B_PER_GPU}, one B200, training step (loss+backward+clip+Adam)")
for k, ms in res.items():
    print(f"  {k:40s} {ms * 1e3:9.1f} us/step  {B_PER_GPU / ms * 1e3 / 1e6:7.2f} M samples/s   loss on batch 0: {vals[k]:+.6e}")
