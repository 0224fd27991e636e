"""GPU parity of kernel (3) assembled: the fused energy MLP (tcgen05 GEMMs through the C ABI) against an fp64 copy of
the oracle's torch restatement of the reference models, the fused FD-SSM step against the reference's golden outputs,
and the fused clip + Adam update against torch.optim.Adam."""
import numpy as np
import pytest
import torch

from oracle import cases, fd_losses, nets, run_case

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _relv(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def _pair(fdsm, D, H, L, act, head, seed=0, presplit=False):
    ref = nets.QuadraticMLP(D, H, L, act) if head == "quadratic" else nets.SoftplusMLP(D, H)
    nets.init_from_numpy(ref, seed)
    net = fdsm.mlp.FusedEnergyMLP(D, H, L, act=act, head=head, device=DEV, presplit=presplit)
    net.load_reference_state_dict([p.detach().float() for p in ref.parameters()])
    return net, ref.double()


@pytest.mark.parametrize("D,H,L,act,head,rows", [
    (64, 48, 3, "nsoftplus", "quadratic", 96), (11, 128, 2, "softplus", "linear", 192), (11, 30, 2, "softplus", "linear", 7),
    (100, 40, 3, "elu", "quadratic", 48), (784, 200, 3, "nsoftplus", "quadratic", 3 * 256), (784, 200, 3, "elu", "quadratic", 130),
    (20, 256, 1, "softplus", "quadratic", 33),
    # H > 256: several N tiles per GEMM, per-tile head partials, column-per-thread head backward (MLP_Quadratic is 5000 wide)
    (64, 300, 2, "nsoftplus", "quadratic", 200), (20, 520, 2, "softplus", "linear", 130), (48, 272, 2, "softplus", "quadratic", 7)])
@pytest.mark.parametrize("presplit", [False, True])
def test_forward_backward_match_fp64(fdsm, D, H, L, act, head, rows, presplit):
    """3xTF32 GEMMs + fused epilogues vs fp64 autograd: energies 5e-6, flat parameter gradient 1e-5 (fp32 SGEMM grade).
    presplit=False: fp32 activations in HBM, hi/lo split inside the GEMMs; True: hi/lo halves materialised in HBM."""
    net, ref = _pair(fdsm, D, H, L, act, head, presplit=presplit)
    x = nets.np_rand(1, rows, D)
    g = nets.np_randn(2, rows)
    E = net(x.float().to(DEV))
    E.backward(g.float().to(DEV))
    Er = ref(x)
    Er.backward(g)
    gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
    assert _relv(E, Er) < 5e-6
    assert _relv(net.flat.grad, gr) < 1e-5
    for (name, _), p in zip(net.param_shapes(), ref.parameters()):     # every tensor, not only the big ones
        assert _relv(net.named_views(net.flat.grad)[name].reshape(-1), p.grad.reshape(-1)) < 5e-5, name


def test_module_is_a_drop_in_for_the_loss_adapters(fdsm, golden):
    """FusedEnergyMLP as `energy_net` of the reference-signature adapters (generic mode): a1 / a7 / a9 fixtures, same
    acceptance rule as test_losses_gpu (err vs fp64 reference <= max(1e-4, 3 x the reference's own fp32 error))."""
    for name in ("a1_quad_d64", "a7_quad_d64", "a7_quad_elu_img", "a9_d48", "a1_gmm11_h128"):
        c = cases.build(name, torch.float32)
        spec = cases.CASES[name]["net"]
        if spec[0] == "quadratic_mlp":
            net = fdsm.mlp.FusedEnergyMLP(spec[1], spec[2], spec[3], act=spec[4], head="quadratic", device=DEV)
        else:
            net = fdsm.mlp.FusedEnergyMLP(spec[1], spec[2], 2, act="softplus", head="linear", device=DEV)
        net.load_reference_state_dict([p.detach() for p in c["net"].parameters()])
        x, u, eps = c["x"].to(DEV), c["u"].to(DEV), c["eps"]
        if c["fn"] == "a1":
            loss = fdsm.efficient_score_matching_conjugate(net, x, eps=eps, u=u)
        elif c["fn"] == "a7":
            loss = fdsm.ssm_fd(net, x, eps=eps, u=u)
        else:
            loss = fdsm.mdsm_fd(net, x, c["sig_b"].to(DEV), c["sigma02"], c["D"], eps=eps, loss_coe=c["loss_coe"], u=u,
                                dsm_noise=c["n"].to(DEV))
        loss.backward()
        l32, l64 = golden[f"{name}/loss_f32"], golden[f"{name}/loss_f64"]
        g32, g64 = golden[f"{name}/grad_f32"], golden[f"{name}/grad_f64"]
        rel = lambda a, b: abs(float(a) - float(b)) / abs(float(b))
        relv = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
        # fp32 noise floor of the second difference (same rule as test_losses_gpu): a few ulp of |E| per term x loss scale
        floor = 0.0
        if c["fn"] in ("a1", "a7"):
            with torch.no_grad():
                emag = float(net(x.reshape(c["B"], -1)).abs().mean())
            floor = 16 * 2.0 ** -24 * emag * ((c["D"] if c["fn"] == "a1" else 1.0) / eps ** 2) / abs(float(l64))
        assert rel(loss, l64) <= max(1e-4, 3 * rel(l32, l64), floor), name
        assert relv(net.flat.grad.cpu().double().numpy(), g64) <= max(1e-4, 3 * relv(g32, g64)), name


def test_fused_fd_ssm_step_matches_oracle_cfg2(fdsm):
    """Config 2 (B=4096, D=784, MLP_Quadratic_MNIST, eps 0.05): the fused path vs the fp64 oracle, with the torch fp32
    path (cuBLAS SGEMM, the reference's arithmetic) on the same inputs as the yardstick."""
    B, D, eps = 4096, 784, 0.05
    net, ref = _pair(fdsm, D, 200, 3, "nsoftplus", "quadratic")
    x, u = nets.np_rand(5, B, D), nets.np_randn(6, B, D)
    lr = fd_losses.fd_ssm(ref, x, fd_losses.sphere_direction(u, eps), 1 / eps ** 2)
    lr.backward()
    gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
    loss, grad = net.fd_ssm_loss_and_grad(x.float().to(DEV), eps, 1 / eps ** 2, u=u.float().to(DEV))
    torch.backends.cuda.matmul.allow_tf32 = False
    r32 = nets.init_from_numpy(nets.QuadraticMLP(D, 200, 3, "nsoftplus"), 0).to(DEV)
    l32 = fd_losses.fd_ssm(r32, x.float().to(DEV), fd_losses.sphere_direction(u.float().to(DEV), eps), 1 / eps ** 2)
    l32.backward()
    g32 = torch.cat([p.grad.reshape(-1) for p in r32.parameters()])
    e_loss, e_loss32 = abs(float(loss) - float(lr)) / abs(float(lr)), abs(float(l32) - float(lr)) / abs(float(lr))
    assert e_loss <= max(1e-4, 3 * e_loss32)
    assert _relv(grad, gr) <= max(1e-4, 3 * _relv(g32, gr))


def test_fused_step_on_exact_quadratic_energy(fdsm):
    """Size-independent property at full config size: for E(x) = 0.5 x'Ax + b'x FD-SSM is exact for any eps.  A 1-hidden-
    layer ELU net with tiny weights is affine in its active region, so the quadratic head gives exactly such an E."""
    B, D, H = 4096, 784, 64
    net = fdsm.mlp.FusedEnergyMLP(D, H, 1, act="elu", head="quadratic", device=DEV, seed=3)
    v = net.named_views()
    v["fc1.weight"].mul_(0.5); v["fc1.bias"].fill_(2.0)           # pre-activations stay in 2 +- 0.8 > 0: ELU is the identity
    x = torch.rand(B, D, device=DEV)
    u = torch.randn(B, D, device=DEV)
    losses = [float(net.fd_ssm_loss_and_grad(x, eps, D / eps ** 2, u=u)[0]) for eps in (0.2, 0.1)]
    W, bb = v["fc1.weight"].double(), v["fc1.bias"].double()
    w1, w2, wq = (v[k].double().view(-1) for k in ("ln1.weight", "ln2.weight", "lq.weight"))
    A = W.t() @ (torch.outer(w1, w2) + torch.outer(w2, w1) + 2 * torch.diag(wq)) @ W       # Hessian of E
    h = x.double() @ W.t() + bb
    gE = ((h @ w1 + v["ln1.bias"].double()) [:, None] * w2 + (h @ w2 + v["ln2.bias"].double())[:, None] * w1 + 2 * wq * h) @ W
    vh = u.double() / u.double().norm(dim=1, keepdim=True)
    want = float((0.5 * (gE * vh).sum(1) ** 2 - ((vh @ A) * vh).sum(1)).mean() * D)
    for got in losses:
        assert abs(got - want) <= 2e-2 * abs(want) + 1e-3


def test_clip_and_adam_match_torch(fdsm):
    net = fdsm.mlp.FusedEnergyMLP(64, 48, 3, device=DEV, seed=1)
    net.configure_optimizer(lr=1e-3, betas=(0.9, 0.95), clip=0.1, use_graph=False)
    p = net.flat.data.clone().requires_grad_(True)
    opt = torch.optim.Adam([p], lr=1e-3, betas=(0.9, 0.95))
    x = torch.rand(32, 64, device=DEV)
    for it in range(3):
        net.fd_ssm_step(x, eps=0.05)
        g = net._step_state["grad"].clone()
        p.grad = g
        torch.nn.utils.clip_grad_norm_([p], 0.1)
        opt.step()
        assert _relv(net.flat.data, p.data) < 1e-6, it
        with torch.no_grad():
            p.copy_(net.flat.data)           # keep the two trajectories on the same weights


def test_graph_replay_equals_eager(fdsm):
    a = fdsm.mlp.FusedEnergyMLP(64, 48, 3, device=DEV, seed=2)
    b = fdsm.mlp.FusedEnergyMLP(64, 48, 3, device=DEV, seed=2)
    a.configure_optimizer(lr=1e-3, use_graph=True, seed=9)
    b.configure_optimizer(lr=1e-3, use_graph=False, seed=9)
    xs = [torch.rand(48, 64, device=DEV) for _ in range(4)]
    for x in xs:
        la, lb = a.fd_ssm_step(x, eps=0.05), b.fd_ssm_step(x, eps=0.05)
        assert float(la) == float(lb)            # same Philox draws (device step counter), same kernels
    assert torch.equal(a.flat.data, b.flat.data)
    assert int(a._opt["step"]) == 4


@pytest.mark.parametrize("D,H,L,act,head,B", [(64, 48, 3, "nsoftplus", "quadratic", 100), (11, 30, 2, "softplus", "linear", 37),
                                              (784, 200, 3, "nsoftplus", "quadratic", 4096)])
def test_fused_combine_backward_equals_explicit_sequence(fdsm, D, H, L, act, head, B):
    """fdsm_mlp_backward_fd_ssm (combine forward + analytic dE + MLP backward in one entry point) against the explicit
    sequence combine_energy_fwd -> combine_energy_bwd -> fdsm_mlp_backward(dE): the loss is bit-identical (same block
    reduction), the gradient is bit-identical too (same dE arithmetic, fixed-order sums) -- and twice in a row."""
    import ctypes as C
    from fdsm_b200 import _lib, ops
    from fdsm_b200.mlp import _ptr, _stream
    lib = _lib.load()
    net = fdsm.mlp.FusedEnergyMLP(D, H, L, act=act, head=head, device=DEV, seed=5, formulation="stacked")
    x = torch.rand(B, D, device=DEV)
    u = torch.randn(B, D, device=DEV)
    eps, scale = 0.05, 1.0 / 0.05 ** 2
    loss, grad = net.fd_ssm_loss_and_grad(x, eps, scale, u=u)
    loss1, grad1 = loss.clone(), grad.clone()
    st = net._step_state
    # explicit sequence on the same workspace / energies
    l2 = torch.zeros((), device=DEV)
    dE = torch.empty(3 * B, device=DEV)
    g2 = torch.empty_like(grad)
    s = _stream()
    _lib.check(lib.fdsm_combine_energy_fwd(_ptr(st["E"]), B, 0, 2.0, scale, 1, None, _ptr(l2), s), "fwd")
    _lib.check(lib.fdsm_combine_energy_bwd(_ptr(st["E"]), B, 0, 2.0, scale, 1, None, _ptr(dE), s), "bwd")
    _lib.check(lib.fdsm_mlp_backward(C.byref(net.desc), _ptr(st["xh"]), _ptr(st["xl"]), 3 * B, _ptr(net.flat.data), _ptr(st["ws"]),
                                     _ptr(dE), _ptr(g2), 0, s), "mlp_bwd")
    torch.cuda.synchronize()
    assert float(loss1) == float(l2)
    assert torch.equal(grad1, g2)
    loss3, grad3 = net.fd_ssm_loss_and_grad(x, eps, scale, u=u)
    assert float(loss3) == float(loss1) and torch.equal(grad3, grad1)      # deterministic: no atomics anywhere
