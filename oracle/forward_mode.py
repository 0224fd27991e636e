"""Forward-mode (Taylor-mode, order 2) restatement of the exact sliced score matching objective for the MLP energies.
TEST INFRASTRUCTURE (see oracle/__init__.py) -- the arithmetic specification of SURVEY.md §8f row N2, checked against the
double-backward form of the reference (DeepEBM/ESM/loss_train.py:173-189 `ssm`, SSM/losses/sliced_sm.py:7-34).

For a direction u the reference needs  d1 = u . grad E(x)  and  d2 = u' Hess E(x) u  (logp = -E):
    per-sample loss = 0.5 * d1^2 - d2            (loss_train.py:181-186 with grad1 = -grad E)
Both are derivatives of the scalar function  tau -> E(x + tau u)  at 0, so they can be carried FORWARD through the
network next to the activations, with no backward pass over the input:

    layer       z  = W h  + b        z'  = W h'               z'' = W h''
    activation  h+ = a(z)            h'+ = a'(z) z'           h''+ = a''(z) z'^2 + a'(z) z''
    start       h  = x               h'  = u                  h''  = 0
    head        E  = l1 l2 + lq.(h^2) + bq,  l_i = w_i . h + b_i
                d1 = l1' l2 + l1 l2' + 2 lq.(h h')            with l_i' = w_i . h',  l_i'' = w_i . h''
                d2 = l1'' l2 + 2 l1' l2' + l1 l2'' + 2 lq.(h'^2 + h h'')

The three streams (h | h' | h'') pass through the SAME weight matrix in every layer: stacked as 3B rows they are one GEMM
of exactly the shape the FD-SSM step already runs on [x | x+v | x-v]; the bias (ones column) applies to the first block
only.  That is the kernel plan for N2: the FD step's GEMMs with a three-stream activation epilogue.
"""
import math

import torch
import torch.nn.functional as F

LN2 = math.log(2.0)


def act_derivatives(name, z):
    """(a(z), a'(z), a''(z)) for softplus / nsoftplus (softplus - ln 2) / elu (alpha = 1)."""
    if name in ("softplus", "nsoftplus"):
        s = torch.sigmoid(z)
        a = F.softplus(z) - (LN2 if name == "nsoftplus" else 0.0)
        return a, s, s * (1 - s)
    if name == "elu":
        e = torch.exp(torch.clamp(z, max=0.0))
        neg = z <= 0
        return torch.where(neg, e - 1, z), torch.where(neg, e, torch.ones_like(z)), torch.where(neg, e, torch.zeros_like(z))
    raise ValueError(name)


def mlp_taylor2(layers, act, x, u):
    """Carry (h, h', h'') through [(W, b), ...] with activation `act` after every layer."""
    h, h1, h2 = x, u, torch.zeros_like(x)
    for W, b in layers:
        z, z1, z2 = h @ W.t() + b, h1 @ W.t(), h2 @ W.t()
        a, da, dda = act_derivatives(act, z)
        h, h1, h2 = a, da * z1, dda * z1 * z1 + da * z2
    return h, h1, h2


def quadratic_head_taylor2(net, h, h1, h2):
    """(E, d1, d2) of E = ln1(h) ln2(h) + lq(h^2) (DeepEBM/ESM/models.py:154-158, 949-953)."""
    w1, w2, wq = net.ln1.weight.view(-1), net.ln2.weight.view(-1), net.lq.weight.view(-1)
    l1, l2 = h @ w1 + net.ln1.bias, h @ w2 + net.ln2.bias
    l11, l21, l12, l22 = h1 @ w1, h1 @ w2, h2 @ w1, h2 @ w2
    E = l1 * l2 + (h * h) @ wq + net.lq.bias
    d1 = l11 * l2 + l1 * l21 + 2 * ((h * h1) @ wq)
    d2 = l12 * l2 + 2 * l11 * l21 + l1 * l22 + 2 * ((h1 * h1 + h * h2) @ wq)
    return E, d1, d2


def energy_taylor2(net, x, u):
    """(E, u.grad E, u' Hess E u) for oracle.nets.QuadraticMLP / SoftplusMLP, forward mode only."""
    x, u = x.reshape(x.shape[0], -1), u.reshape(u.shape[0], -1)
    if hasattr(net, "ln1"):
        layers = [(getattr(net, f"fc{i + 1}").weight, getattr(net, f"fc{i + 1}").bias) for i in range(net.n_hidden)]
        return quadratic_head_taylor2(net, *mlp_taylor2(layers, net.act, x, u))
    lins = [m for m in net.net if isinstance(m, torch.nn.Linear)]
    h, h1, h2 = mlp_taylor2([(m.weight, m.bias) for m in lins[:-1]], "softplus", x, u)
    w, b = lins[-1].weight.view(-1), lins[-1].bias
    return h @ w + b, h1 @ w, h2 @ w


def ssm_forward_mode(net, x, u):
    """mean_b [0.5 (u . grad logp)^2 + u' Hess logp u] with logp = -E: the value (and, through torch autograd over the
    forward-mode expression, the parameter gradient) of loss_train.ssm / single_sliced_score_matching."""
    _, d1, d2 = energy_taylor2(net, x, u)
    return (0.5 * d1 * d1 - d2).mean()
