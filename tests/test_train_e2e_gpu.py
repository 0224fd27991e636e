"""GPU tests of the end-to-end training call (`FDSSMTrainer.step_from_host`: pinned host batch -> H2D on the copy
streams -> the step -> loss into pinned host memory by `fdsm_publish_scalars`): it must give bit for bit what the
device-resident `step` gives (DeepEBM/train_ssm.py:60-70 is one loop, whichever side the batch starts on)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_publish_scalars_writes_pinned_host_memory(fdsm):
    for n in (1, 37, 4096):
        src = torch.randn(n, device=DEV)
        dst = torch.zeros(n).pin_memory()
        fdsm.ops.publish_scalars(src, dst)
        torch.cuda.synchronize()
        assert torch.equal(dst, src.cpu())
    with pytest.raises(ValueError):
        fdsm.ops.publish_scalars(torch.randn(4, device=DEV), torch.zeros(4))          # pageable destination
    lib = fdsm._lib.load()
    src = torch.randn(4, device=DEV)
    pageable = np.zeros(4, dtype=np.float32)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.fdsm_publish_scalars(C.c_void_p(src.data_ptr()), C.c_void_p(pageable.ctypes.data), 4, stream)
    assert rc == -1 and b"pinned" in lib.fdsm_last_error_string()
    rc = lib.fdsm_publish_scalars(C.c_void_p(src.data_ptr()), C.c_void_p(src.data_ptr()), 4, stream)   # device memory
    assert rc == -1
    rc = lib.fdsm_publish_scalars(C.c_void_p(src.data_ptr()), C.c_void_p(pageable.ctypes.data), 5000, stream)
    assert rc == -1
    torch.cuda.synchronize()


def _trainer(fdsm, D, H):
    net = fdsm.mlp.FusedEnergyMLP(D, H, 3, act="nsoftplus", head="quadratic", device=torch.device(DEV), seed=7)
    from fdsm_b200.train import FDSSMTrainer
    return FDSSMTrainer(net, eps=0.05, lr=1e-3, betas=(0.9, 0.95), clip=0.1, seed=99)


@pytest.mark.parametrize("B,D,H", [(256, 784, 200), (96, 64, 32)])
def test_step_from_host_equals_device_step(fdsm, B, D, H):
    batches = [torch.rand(B, D) for _ in range(3)]
    tr_dev, tr_host = _trainer(fdsm, D, H), _trainer(fdsm, D, H)
    host = [b.pin_memory() for b in batches]
    loss_host = torch.zeros((), pin_memory=True)
    for i in range(7):
        want = float(tr_dev.step(batches[i % 3].to(DEV)))
        tr_host.step_from_host(host[i % 3], loss_host)
        torch.cuda.synchronize()
        assert float(loss_host) == want, f"step {i}"
    assert torch.equal(tr_dev.flat.flat, tr_host.flat.flat)
    # pageable loss destination: torch's copy path, same values
    loss_pageable = torch.zeros(())
    tr_dev.step(batches[0].to(DEV))
    tr_host.step_from_host(host[0], loss_pageable)
    torch.cuda.synchronize()
    assert torch.equal(tr_dev.flat.flat, tr_host.flat.flat)


def test_step_from_host_u8_equals_device_u8_step(fdsm):
    B, D, H = 128, 784, 200
    if _trainer(fdsm, D, H).net.formulation != "jet":
        pytest.skip("uint8 pixel batches need the chain kernel")
    pix = [torch.randint(0, 256, (B, D), dtype=torch.uint8) for _ in range(2)]
    tr_dev, tr_host = _trainer(fdsm, D, H), _trainer(fdsm, D, H)
    host = [p.pin_memory() for p in pix]
    dev = [p.to(DEV) for p in pix]
    loss_host = torch.zeros((), pin_memory=True)
    for i in range(5):
        want = float(tr_dev.step(dev[i % 2], borrow=True, x_noise=0.005))
        tr_host.step_from_host_u8(host[i % 2], loss_host, 0.005)
        torch.cuda.synchronize()
        assert float(loss_host) == want, f"step {i}"
    assert torch.equal(tr_dev.flat.flat, tr_host.flat.flat)


def test_fp16_input_layer_on_pixel_batches(fdsm):
    """jet_passes = 4 (input layer in kind::f16) through both input kernels -- fp32 rows and uint8 pixels + noise: the gradient
    stays within TF32 rounding of the jet_passes = 2 gradient (same draws), and a training step runs through it."""
    B, D, H = 256, 784, 200
    nets_ = [fdsm.mlp.FusedEnergyMLP(D, H, 3, act="nsoftplus", head="quadratic", device=torch.device(DEV), seed=7, jet_passes=p)
             for p in (2, 4)]
    if nets_[0].formulation != "jet":
        pytest.skip("needs the chain kernel")
    pix = torch.randint(0, 256, (B, D), dtype=torch.uint8, device=DEV)
    xf = pix.float() / 255.0
    for x, kw in ((xf, {}), (pix, {"x_noise": 0.005})):
        grads = []
        for net in nets_:
            _, g = net._jet_loss_and_grad(x, 0.05, 400.0, None, None, (5, 3, 0), 2.0, True, None, **kw)
            grads.append(g.clone())
        rel = float((grads[1] - grads[0]).norm() / grads[0].norm())
        assert rel <= 5e-4, rel
    from fdsm_b200.train import FDSSMTrainer
    tr = FDSSMTrainer(nets_[1], eps=0.05, lr=1e-3, betas=(0.9, 0.95), clip=0.1, seed=99)
    l0 = float(tr.step(xf))
    l1 = float(tr.step(pix, borrow=True, x_noise=0.005))
    assert np.isfinite(l0) and np.isfinite(l1)
