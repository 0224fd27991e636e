"""Host logic of the a17 comparison adapters that contain no kernel of ours once the draws are injected
(anneal_sliced_score_estimation_vr, exact_score_matching): they must reproduce the reference's outputs (tests/golden)
on CPU tensors.  The FD adapters themselves cannot run here: their kernels refuse CPU tensors (checked below)."""
import numpy as np
import pytest
import torch

from oracle import cases


def _rel(a, b):
    return abs(float(a) - float(b)) / max(abs(float(b)), 1e-30)


def _relvec(a, b):
    return float(np.linalg.norm(np.asarray(a, dtype=np.float64) - b) / max(np.linalg.norm(b), 1e-30))


@pytest.mark.parametrize("name", ["a17_anneal_vr_img", "a17_anneal_vr_p2", "a17_exact_gmm11", "a17_mdsm_baseline_d48",
                                  "a17_mdsm_tracetrick_d48", "a17_exact_deepebm_d16"])
@pytest.mark.parametrize("tag,dtype,tol", [("f64", torch.float64, 1e-9), ("f32", torch.float32, 1e-4)])
def test_torch_only_adapters_match_reference(fdsm, golden, name, tag, dtype, tol):
    c = cases.build(name, dtype)
    L = fdsm.losses
    if c["fn"] == "a17_exact":
        loss = L.exact_score_matching(c["net"], c["x"], train=True).mean()
    elif c["fn"] == "a17_exact_deepebm":
        loss = L.deepebm_exact_score_matching(c["net"], c["x"], c["D"])
        assert not loss.requires_grad
    elif c["fn"] == "a17_mdsm_baseline":
        loss = L.mdsm_baseline(c["net"], c["x"], c["sig_b"], c["sigma02"], c["D"], loss_coe=c["loss_coe"], dsm_noise=c["n"])
    elif c["fn"] == "a17_mdsm_tracetrick":
        loss = L.mdsm_tracetrick(c["net"], c["x"], c["sig_b"], c["sigma02"], c["D"], loss_coe=c["loss_coe"], u=c["u"],
                                 dsm_noise=c["n"])
    else:
        loss = L.anneal_sliced_score_estimation_vr(c["net"], c["x"], c["sigmas"], c["labels"], c["anneal_power"], None,
                                                   c["n_particles"], u=c["u"], dsm_noise=c["n"])
    if loss.requires_grad:
        loss.backward()
    assert _rel(loss, golden[f"{name}/loss_{tag}"]) < tol
    assert _relvec(cases.flat_grad(c["net"]).numpy(), golden[f"{name}/grad_{tag}"]) < tol * 10


def test_exact_score_matching_eval_mode_is_detached(fdsm):
    c = cases.build("a17_exact_gmm11", torch.float64)
    per = fdsm.losses.exact_score_matching(c["net"], c["x"])
    assert per.shape == (c["B"],) and not per.requires_grad
    want = fdsm.losses.exact_score_matching(c["net"], c["x"], train=True)
    assert torch.allclose(per, want.detach(), rtol=1e-12, atol=0)


def test_fd_adapters_refuse_cpu_tensors(fdsm):
    """No CPU fallback: every FD adapter reaches a kernel, and the kernels refuse host tensors."""
    c = cases.build("a2_rad_gmm11", torch.float32)
    for nt in ("sphere", "radermacher"):
        with pytest.raises(fdsm._lib.FdsmError):
            fdsm.losses.single_efficient_score_matching(c["net"], c["x"], eps=0.01, noise_type=nt, u=c["u"])


def test_read_loss_to_host_takes_the_copy_path_for_pageable_destinations(fdsm):
    """train.read_loss_to_host: only a pinned fp32 destination on a CUDA loss goes through fdsm_publish_scalars; everything
    else (here: CPU tensors) is torch's copy -- same value either way."""
    import torch
    from fdsm_b200.train import read_loss_to_host
    loss, dst = torch.tensor(1.25), torch.zeros(())
    assert read_loss_to_host(loss, dst) is dst and float(dst) == 1.25
    dst64 = torch.zeros((), dtype=torch.float64)
    read_loss_to_host(torch.tensor(-3.5), dst64)
    assert float(dst64) == -3.5


def test_inorm_plus_op_sequence_second_order_on_cpu(fdsm):
    """The op sequence InormPlus.backward differentiates when the caller asks for the gradient of the gradient
    (create_graph=True) is an ordinary differentiable torch function: gradcheck in float64, first and second order."""
    import torch
    torch.manual_seed(0)
    x = torch.randn(2, 3, 4, 4, dtype=torch.float64, requires_grad=True)
    g, a, b = (torch.randn(1, 3, dtype=torch.float64, requires_grad=True) for _ in range(3))
    f = lambda x, g, a, b: fdsm.ops.inorm_plus_reference(x, g, a, b, act=True)
    assert torch.autograd.gradcheck(f, (x, g, a, b), eps=1e-6, atol=1e-6)
    assert torch.autograd.gradgradcheck(f, (x, g, a, b), eps=1e-6, atol=1e-5)
