"""GPU parity of the difference-propagating FD-SSM path (csrc/fd_jet.cuh, fd_chain.cuh; fdsm_jet_fd_ssm through the C ABI)
against the float64 oracle -- the stacked evaluation the reference performs (oracle/fd_losses.fd_ssm on an fp64 copy of the
net) -- and against the reference's own golden outputs at the BASELINE shapes.

Tolerances (relative, vs float64): jet_passes = 3 (all weights as TF32 hi + lo; activations TF32): loss 2e-4, gradient 2e-4
on the softplus family (measured 5e-6 / 4e-5 at config 2), 5e-2 on ELU nets, whose FD objective is sensitive to the kink of
elu'' (the reference's own fp32 is off by 9e-2 there: tests/tools/jet_accuracy_study.py); jet_passes = 2 (the default: the
input layer's weights rounded to TF32): loss 2e-4, gradient 5e-4 (2e-5 / 9e-5 at config 2; narrow input layers average the
rounding over fewer terms: 2.3e-4 at D = 64); jet_passes = 1: 1e-3.
For comparison the reference's fp32 stacked arithmetic is off by 1e-2 .. 1 on the same inputs (golden loss_f32 vs loss_f64)."""
import os

import numpy as np
import pytest
import torch

from oracle import fd_losses, jet, nets
from oracle.make_golden_baseline import SEED

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _relv(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def _pair(fdsm, D, H, L, act, head, seed=0, passes=2):
    ref = nets.QuadraticMLP(D, H, L, act) if head == "quadratic" else nets.SoftplusMLP(D, H)
    nets.init_from_numpy(ref, seed)
    net = fdsm.mlp.FusedEnergyMLP(D, H, L, act=act, head=head, device=DEV, formulation="jet", jet_passes=passes)
    net.load_reference_state_dict([p.detach().float() for p in ref.parameters()])
    return net, ref.double()


def _oracle(ref, x, v, scale, c1=2.0, reduce="mean"):
    for p in ref.parameters():
        p.grad = None
    loss = fd_losses.fd_ssm(ref, x, v, scale, c1=c1, reduce=reduce)
    loss.backward()
    return float(loss), torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in ref.parameters()])


CASES = [
    # D, H, L, act, head, B, eps, data
    (64, 48, 3, "nsoftplus", "quadratic", 37, 0.05, "unit"),
    (64, 48, 3, "softplus", "quadratic", 32, 0.1, "unit"),
    (11, 128, 2, "softplus", "linear", 1024, 0.1, "gmm"),
    (11, 30, 2, "softplus", "linear", 100, 0.1, "gmm"),
    (100, 40, 3, "elu", "quadratic", 16, 0.05, "unit"),
    (784, 200, 3, "nsoftplus", "quadratic", 4096, 0.05, "unit"),
    (784, 200, 3, "elu", "quadratic", 1000, 0.05, "unit"),
    (20, 223, 1, "softplus", "quadratic", 9, 0.05, "unit"),
    (33, 16, 5, "nsoftplus", "quadratic", 300, 0.05, "unit"),
]


@pytest.mark.parametrize("D,H,L,act,head,B,eps,data", CASES)
@pytest.mark.parametrize("passes", [3, 2, 1, 4])
def test_jet_loss_and_grad_match_fp64(fdsm, D, H, L, act, head, B, eps, data, passes):
    net, ref = _pair(fdsm, D, H, L, act, head, passes=passes)
    x = nets.np_rand(1, B, D) if data == "unit" else nets.np_randn(1, B, D) + 5.0 * ((nets.np_rand(3, B, 1) < 0.8).double() * 2 - 1)
    u = nets.np_randn(2, B, D)
    scale = (D if head == "linear" else 1.0) / eps ** 2
    want, gwant = _oracle(ref, x, fd_losses.sphere_direction(u, eps), scale)
    loss, grad = net.fd_ssm_loss_and_grad(x.float().to(DEV), eps, scale, u=u.float().to(DEV))
    tol_l, tol_g = {3: (2e-4, 2e-4), 2: (2e-4, 5e-4), 1: (1e-3, 1e-3), 4: (2e-4, 5e-4)}[passes]     # 4: input layer in fp16 = TF32's 11 bits
    if act == "elu":
        tol_g = 5e-2
    # the loss is a signed mean: measure its error against the mean |per-sample term| when the mean itself nearly cancels
    with torch.no_grad():
        E0, dE, sE = jet.energy_jet(ref, x, fd_losses.sphere_direction(u, eps))
        mag = float((dE * dE / 8 - sE).abs().mean() * scale)
    assert abs(float(loss) - want) <= tol_l * max(abs(want), mag), (float(loss), want, mag)
    assert _relv(grad, gwant) <= tol_g, _relv(grad, gwant)


def test_jet_variants_c1_sum_injected_v(fdsm):
    """a5 (c1 = mle_ratio, summed, no /eps^2) and an injected, already-scaled direction (the reference's noise= hook)."""
    net, ref = _pair(fdsm, 11, 30, 2, "softplus", "linear")
    B, eps = 48, 0.01
    x, u = nets.np_randn(1, B, 11) + 5.0, nets.np_randn(2, B, 11)
    v = fd_losses.sphere_direction(u, eps)
    want, gwant = _oracle(ref, x, v, 1.0, c1=10.0, reduce="sum")
    loss, grad = net.fd_ssm_loss_and_grad(x.float().to(DEV), eps, 1.0, v=v.float().to(DEV), c1=10.0, mean=False)
    assert abs(float(loss) - want) <= 2e-4 * abs(want)
    assert _relv(grad, gwant) <= 2e-4


def test_jet_philox_direction_is_the_stack_kernels(fdsm):
    """Production mode draws v from the same Philox stream, keyed the same way, as fdsm_stack: identical directions."""
    net, _ = _pair(fdsm, 64, 48, 3, "nsoftplus", "quadratic")
    x = nets.np_rand(1, 50, 64).float().to(DEV)
    net.fd_ssm_loss_and_grad(x, 0.05, 400.0, draw=(99, 7, 1000))
    st = net._step_state
    o = st["offsets"]
    XV = st["ws"][o[5]:o[5] + 2 * o[1] * o[3]].view(2, o[1], o[3])
    v = fdsm.ops.direction(50, 64, 0.05, device=DEV, draw=(99, 7, 1000))
    # stored rounded to TF32 (round-to-nearest): 2^-11 relative
    torch.testing.assert_close(XV[1, :50, :64], v, rtol=2.0 ** -11, atol=0)
    assert float((XV[1, :50, :64].view(torch.int32) & 0x1FFF).abs().max()) == 0


def test_jet_inputs_specialisation_is_bit_identical_to_the_general_kernel(fdsm):
    """jet_inputs_kernel<1> (fp32 rows, 16-byte aligned, ld % 4 == 0: the production configuration) against the general
    instantiation, reached with the same samples in a buffer whose row stride is not a multiple of 4: same XV bits."""
    net, _ = _pair(fdsm, 64, 48, 3, "nsoftplus", "quadratic")
    x = nets.np_rand(1, 50, 64).float().to(DEV)
    def xv(xin):
        net.fd_ssm_loss_and_grad(xin, 0.05, 400.0, draw=(99, 7, 1000))
        o = net._step_state["offsets"]
        return net._step_state["ws"][o[5]:o[5] + 2 * o[1] * o[3]].clone()
    fast = xv(x)
    wide = torch.zeros(50, 65, device=DEV)
    wide[:, :64] = x
    slow = xv(wide[:, :64])            # row stride 65 floats
    assert wide[:, :64].stride(0) == 65
    assert torch.equal(fast, slow)


def test_jet_is_deterministic_and_shard_invariant(fdsm):
    """Same inputs -> bit-identical loss and gradient; a batch split in two halves (row0 offsets) gives the mean of the
    halves' gradients up to fp32 summation order."""
    net, _ = _pair(fdsm, 64, 48, 3, "nsoftplus", "quadratic")
    x = nets.np_rand(1, 64, 64).float().to(DEV)
    l1, g1 = net.fd_ssm_loss_and_grad(x, 0.05, 400.0, draw=(5, 3, 0))
    l1, g1 = l1.clone(), g1.clone()
    l2, g2 = net.fd_ssm_loss_and_grad(x, 0.05, 400.0, draw=(5, 3, 0))
    assert torch.equal(l1, l2) and torch.equal(g1, g2)
    la, ga = net.fd_ssm_loss_and_grad(x[:32], 0.05, 400.0, draw=(5, 3, 0))
    la, ga = la.clone(), ga.clone()
    lb, gb = net.fd_ssm_loss_and_grad(x[32:].contiguous(), 0.05, 400.0, draw=(5, 3, 32))
    torch.testing.assert_close(0.5 * (la + lb), l1, rtol=1e-5, atol=0)
    assert _relv(0.5 * (ga + gb), g1) < 1e-5


@pytest.mark.parametrize("act", ["nsoftplus", "elu"])
def test_jet_cfg2_against_reference_golden(fdsm, act):
    """Config 2 at full size against the UNMODIFIED reference's float64 output (tests/golden/baseline_cfg2.npz), with the
    reference's own fp32 error printed next to ours."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "baseline_cfg2.npz"))
    B, D, eps = 4096, 784, 0.05
    ref = nets.init_from_numpy(nets.QuadraticMLP(D, 200, 3, act), SEED)
    net = fdsm.mlp.FusedEnergyMLP(D, 200, 3, act=act, head="quadratic", device=DEV, formulation="jet")
    net.load_reference_state_dict([p.detach().float() for p in ref.parameters()])
    x = (nets.np_rand(SEED + 2, B, 1, 28, 28) + 0.005 * nets.np_randn(SEED + 4, B, 1, 28, 28)).float().reshape(B, D)
    u = nets.np_randn(SEED + 3, B, D).float()
    loss, grad = net.fd_ssm_loss_and_grad(x.to(DEV), eps, 1 / eps ** 2, u=u.to(DEV))
    l64, l32 = float(g[f"{act}/loss_f64"]), float(g[f"{act}/loss_f32"])
    g64, g32 = g[f"{act}/grad_f64"].astype(np.float64), g[f"{act}/grad_f32"].astype(np.float64)
    ours_l = abs(float(loss) - l64) / abs(l64)
    ours_g = float(np.linalg.norm(grad.cpu().double().numpy() - g64) / np.linalg.norm(g64))
    ref_l, ref_g = abs(l32 - l64) / abs(l64), float(np.linalg.norm(g32 - g64) / np.linalg.norm(g64))
    print(f"cfg2 {act}: loss err ours {ours_l:.2e} / reference fp32 {ref_l:.2e}; grad err ours {ours_g:.2e} / reference fp32 {ref_g:.2e}")
    assert ours_l <= 2e-4
    assert ours_g <= (2e-4 if act == "nsoftplus" else max(5e-2, ref_g))


def test_jet_training_step_graph_replay(fdsm):
    """fd_ssm_step on the jet path: the captured graph reproduces the eager sequence bit for bit and the parameters move."""
    net, _ = _pair(fdsm, 64, 48, 3, "nsoftplus", "quadratic")
    twin, _ = _pair(fdsm, 64, 48, 3, "nsoftplus", "quadratic")
    net.configure_optimizer(lr=1e-3, betas=(0.9, 0.95), clip=0.1, use_graph=True, seed=3)
    twin.configure_optimizer(lr=1e-3, betas=(0.9, 0.95), clip=0.1, use_graph=False, seed=3)
    xs = [nets.np_rand(10 + i, 96, 64).float().to(DEV) for i in range(4)]
    for x in xs:
        la = net.fd_ssm_step(x, eps=0.05).clone()
        lb = twin.fd_ssm_step(x, eps=0.05).clone()
        assert torch.equal(la, lb)
    assert torch.equal(net.flat.data, twin.flat.data)
    assert int(net._opt["step"]) == 4


def test_jet_fused_update_matches_the_separate_kernels(fdsm, monkeypatch):
    """The fused reduce + clip + Adam launch (fdsm_jet_update) against reduce / finish / fdsm_grad_sumsq / fdsm_adam_step: same
    gradient (the head sums and the squared norm are added in a different fixed order: 1e-6) and the same parameters."""
    xs = [nets.np_rand(20 + i, 200, 64).float().to(DEV) for i in range(5)]
    out = {}
    for fused in ("1", "0"):
        monkeypatch.setenv("FDSM_JET_FUSED_UPDATE", fused)
        net, _ = _pair(fdsm, 64, 48, 3, "nsoftplus", "quadratic")
        net.configure_optimizer(lr=1e-3, betas=(0.9, 0.95), clip=0.1, use_graph=True, seed=3)
        losses = [float(net.fd_ssm_step(x, eps=0.05)) for x in xs]
        out[fused] = (losses, net.flat.data.clone(), net._step_state["grad"].clone(), int(net._opt["step"]))
    (la, pa, ga, sa), (lb, pb, gb, sb) = out["1"], out["0"]
    assert sa == sb == 5
    assert np.allclose(la, lb, rtol=1e-5, atol=0)
    assert float((ga - gb).norm() / gb.norm()) <= 1e-5
    assert float((pa - pb).norm() / pb.norm()) <= 1e-6


def test_jet_uint8_pixels_equal_the_explicit_float_batch(fdsm):
    """x_u8 / 255 + ssm_eps * N(0,1) built inside the input kernel (fdsm_jet_args.x_u8; DeepEBM/train_ssm.py:61-62 on uint8
    pixels) against the same batch built explicitly in fp32 from the same Philox stream: same loss and gradient up to the
    one fused multiply-add per element."""
    from fdsm_b200 import ops
    B, D = 200, 64
    net, _ = _pair(fdsm, D, 48, 3, "nsoftplus", "quadratic")
    x8 = torch.randint(0, 256, (B, D), dtype=torch.uint8, device=DEV)
    draw = (77, 5, 1000)
    n = ops.philox_normal(B, D, 1, draw[0], draw[1], row0=draw[2], device=DEV)
    xf = x8.float() * (1.0 / 255.0) + 0.005 * n
    la, ga = net._jet_loss_and_grad(x8, 0.05, 400.0, None, None, draw, 2.0, True, None, x_noise=0.005)
    la, ga = la.clone(), ga.clone()
    lb, gb = net._jet_loss_and_grad(xf, 0.05, 400.0, None, None, draw, 2.0, True, None)
    assert abs(float(la) - float(lb)) <= 1e-5 * max(abs(float(lb)), 1e-12) + 1e-9
    assert _relv(ga, gb) <= 1e-4
    # and through the training step (graph replay, borrowed staging buffer)
    net.configure_optimizer(lr=1e-3, betas=(0.9, 0.95), clip=0.1, use_graph=True, seed=3)
    l1 = float(net.fd_ssm_step(x8, eps=0.05, borrow=True, x_noise=0.005))
    l2 = float(net.fd_ssm_step(x8, eps=0.05, borrow=True, x_noise=0.005))
    assert np.isfinite(l1) and np.isfinite(l2) and int(net._opt["step"]) == 2


@pytest.mark.parametrize("name", ["a17_ssm_d64", "a17_sliced_gmm11"])
def test_exact_ssm_forward_mode_matches_reference_golden(fdsm, golden, name):
    """exact sliced score matching on the fused MLP in FORWARD mode inside the chain kernel
    (fdsm_jet_args.mode = 1) against the goldens generated from the unmodified reference's double-backward code
    (DeepEBM/ESM/loss_train.py:173-189 ssm on MLP_Quadratic; SSM/losses/sliced_sm.py:7-34 on the Softplus MLP).
    Tolerance against the reference's fp64 values: 2e-4 on the loss, 5e-4 on the gradient (activations and gradients travel
    as TF32; at B = 32 / 48 their rounding averages over few terms: measured 2.5e-4; at config 2 the loss agrees with the
    double-backward form to 4e-5, profiles/r02_fd_vs_exact_ssm.txt)."""
    from oracle import cases
    c = cases.build(name, torch.float64)
    spec = cases.CASES[name]["net"]
    if spec[0] == "quadratic_mlp":
        net = fdsm.mlp.FusedEnergyMLP(spec[1], spec[2], spec[3], act=spec[4], head="quadratic", device=DEV, jet_passes=3)
    else:
        net = fdsm.mlp.FusedEnergyMLP(spec[1], spec[2], 2, act="softplus", head="linear", device=DEV, jet_passes=3)
    net.load_reference_state_dict([p.detach().float() for p in c["net"].parameters()])
    u = c["u"].reshape(c["B"], -1)
    if c["fn"] == "a17_sliced":
        u = u.sign()          # SSM/losses/sliced_sm.py:12 default noise_type
    x = c["x"].reshape(c["B"], -1).float().to(DEV)
    loss, grad = net.exact_ssm_loss_and_grad(x, u=u.float().to(DEV))
    l64, g64 = float(golden[f"{name}/loss_f64"]), golden[f"{name}/grad_f64"]
    assert abs(float(loss) - l64) <= 2e-4 * abs(l64), (float(loss), l64)
    assert _relv(grad, torch.from_numpy(np.asarray(g64))) <= 5e-4
    # the autograd route of the adapter (losses.ssm on a fused net) gives the same numbers
    if c["fn"] == "a17_ssm":
        net.flat.grad = None
        l2 = fdsm.losses.ssm(net, x, batch_size=c["B"], u=u.float().to(DEV))
        l2.backward()
        assert float(l2) == float(loss) and torch.equal(net.flat.grad, grad)


def test_exact_ssm_philox_projections_and_step(fdsm):
    """In-kernel Gaussian projections (unnormalised Philox rows) = the same rows injected; the exact-SSM training step
    (fd_ssm_step(eps=None)) replays from a graph and moves the parameters."""
    from fdsm_b200 import ops
    B, D = 96, 64
    net, _ = _pair(fdsm, D, 48, 3, "nsoftplus", "quadratic")
    x = nets.np_rand(3, B, D).float().to(DEV)
    draw = (11, 2, 500)
    la, ga = net.exact_ssm_loss_and_grad(x, draw=draw)
    la, ga = la.clone(), ga.clone()
    u = ops.philox_normal(B, D, 0, draw[0], draw[1], row0=draw[2], device=DEV)
    lb, gb = net.exact_ssm_loss_and_grad(x, u=u)
    assert abs(float(la) - float(lb)) <= 1e-6 * abs(float(lb)) and _relv(ga, gb) <= 1e-5
    net.configure_optimizer(lr=1e-3, betas=(0.9, 0.95), clip=0.1, use_graph=True, seed=3)
    p0 = net.flat.data.clone()
    for _ in range(3):
        l = net.fd_ssm_step(x, eps=None)
    assert np.isfinite(float(l)) and int(net._opt["step"]) == 3 and not torch.equal(p0, net.flat.data)
