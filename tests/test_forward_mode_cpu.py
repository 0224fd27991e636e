"""SURVEY.md §8f row N2, arithmetic specification: the forward-mode (order-2 Taylor) form of exact SSM for the MLP energies
(oracle/forward_mode.py) equals the reference's double-backward form -- against the golden vectors generated from the
unmodified reference (a17 cases) and against the oracle's HVP restatement, value and parameter gradient, in fp64."""
import numpy as np
import pytest
import torch

from oracle import cases, fd_losses, forward_mode, nets


def _rel(a, b):
    return abs(float(a) - float(b)) / max(abs(float(b)), 1e-30)


@pytest.mark.parametrize("name", ["a17_ssm_d64", "a17_sliced_gmm11"])
def test_forward_mode_matches_reference_golden(golden, name):
    c = cases.build(name, torch.float64)
    u = c["u"].reshape(c["B"], -1)
    if c["fn"] == "a17_sliced":
        u = u.sign()          # SSM/losses/sliced_sm.py:12 default noise_type
    loss = forward_mode.ssm_forward_mode(c["net"], c["x"], u)
    loss.backward()
    assert _rel(loss.detach(), golden[f"{name}/loss_f64"]) < 1e-10
    g = cases.flat_grad(c["net"]).numpy()
    want = golden[f"{name}/grad_f64"]
    assert np.linalg.norm(g - want) / np.linalg.norm(want) < 1e-9


@pytest.mark.parametrize("act", ["nsoftplus", "softplus", "elu"])
def test_taylor_streams_match_autograd(act):
    """d1 = u . grad E and d2 = u' Hess E u per sample, against two autograd passes; ELU included (a'' jumps at 0, which
    random pre-activations never hit exactly)."""
    net = nets.init_from_numpy(nets.QuadraticMLP(20, 16, 3, act), 11).double()
    x = nets.np_randn(12, 9, 20).double().requires_grad_(True)
    u = nets.np_randn(13, 9, 20).double()
    E, d1, d2 = forward_mode.energy_taylor2(net, x, u)
    E_ref = net(x)
    g = torch.autograd.grad(E_ref.sum(), x, create_graph=True)[0]
    hv = torch.autograd.grad((g * u).sum(), x)[0]
    torch.testing.assert_close(E, E_ref, rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(d1, (g * u).sum(-1), rtol=1e-10, atol=1e-12)
    torch.testing.assert_close(d2, (hv * u).sum(-1), rtol=1e-9, atol=1e-11)


def test_forward_mode_is_the_eps_to_zero_limit_of_fd_ssm():
    """FD-SSM (a1 with v = eps * vhat) -> forward-mode exact SSM with u = sqrt(D) * vhat as eps -> 0."""
    c = cases.build("a1_gmm11_h30", torch.float64)
    D = c["D"]
    vhat = fd_losses.sphere_direction(c["u"], 1.0)
    want = forward_mode.ssm_forward_mode(c["net"], c["x"], vhat * D ** 0.5)
    errs = [_rel(fd_losses.fd_ssm(c["net"], c["x"], vhat * eps, D / eps ** 2), want) for eps in (0.1, 0.01)]
    assert errs[1] < errs[0] / 10 and errs[1] < 1e-4
