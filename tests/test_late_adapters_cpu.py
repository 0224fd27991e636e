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
