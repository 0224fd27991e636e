"""Difference-propagating ("jet") restatement of the FD-SSM objective for the MLP energies.  TEST INFRASTRUCTURE
(see oracle/__init__.py): the arithmetic specification of the fused chain kernel (csrc/fd_chain.cu), in torch CPU ops,
any float dtype.

The reference evaluates the network on the stacked batch [x, x+v, x-v] and subtracts O(|E|) numbers to get O(eps^2)
differences (DeepEBM/ESM/loss_train.py:212-234, SSM/losses/efficient_sm.py:47-73).  The same three function values can
be carried as (h0, a, b) with

    h+ = h0 + b + a          a = (h+ - h-)/2   (odd part,  O(eps))
    h- = h0 + b - a          b = (h+ + h-)/2 - h0   (even part, O(eps^2))

    layer       z0 = W h0 + bias      P = W a            Q = W b          (one GEMM over the three streams)
    activation  h0' = s(z0)           a' = (s(z0+Q+P) - s(z0+Q-P))/2      b' = (s(z0+Q+P) + s(z0+Q-P))/2 - s(z0)
    start       h0 = x                a  = v             b  = 0

with a', b' evaluated in cancellation-free closed forms.  Mathematically identical to the stacked evaluation; rounding
is now relative to the DIFFERENCES, so single-pass TF32 tensor-core operands perturb the loss at the 1e-4 level instead
of swamping it (hard part 2(i)).

    E+ - E-        =: dE      E+ + E- - 2 E0 =: sE
    per-sample loss = dE^2/8 - sE + (c1 - 2) E0     (c1 = 2: the reference's 2*E1 term)
"""
import math

import torch
import torch.nn.functional as F

LN2 = math.log(2.0)


def tf32_round(t):
    """Round-to-nearest (ties away) to TF32 (10 mantissa bits), as cvt.rna.tf32.f32; identity for float64 inputs is NOT
    applied -- a float64 tensor is rounded through float32."""
    x = t.detach().to(torch.float32).contiguous()
    bits = x.view(torch.int32)
    r = ((bits + 0x1000) & ~0x1FFF).view(torch.float32)
    return r.to(t.dtype)


class _Tf32Linear(torch.autograd.Function):
    """y = rn(x) rn(W)^T (+ bias in full precision); both backward GEMMs with TF32-rounded operands as well."""

    @staticmethod
    def forward(ctx, x, W, bias):
        xr, Wr = tf32_round(x), tf32_round(W)
        ctx.save_for_backward(xr, Wr)
        ctx.has_bias = bias is not None
        y = xr @ Wr.t()
        return y + bias if bias is not None else y

    @staticmethod
    def backward(ctx, g):
        xr, Wr = ctx.saved_tensors
        gr = tf32_round(g)
        return gr @ Wr, gr.t() @ xr, (g.sum(0) if ctx.has_bias else None)


def linear(x, W, bias, tf32):
    if tf32:
        return _Tf32Linear.apply(x, W, bias)
    y = x @ W.t()
    return y + bias if bias is not None else y


def _coshm1(p):
    s = torch.sinh(0.5 * p)
    return 2.0 * s * s


class _SoftplusJet(torch.autograd.Function):
    """(z0, P, Q) -> (h0, a, b) for s = softplus (offset = 0) or softplus - ln 2 (the offset only moves h0).

    forward, with s0 = sigmoid(z0), d+ = Q + P, d- = Q - P:
        (1 + e^{z0 + d}) / (1 + e^{z0}) = 1 + s0 expm1(d)
        a = 1/2 log[(1 + s0 expm1(d+)) / (1 + s0 expm1(d-))] = 1/2 log1p(2 s0 e^Q sinh P / (1 + s0 expm1(d-)))
        b = 1/2 log[(1 + s0 expm1(d+))(1 + s0 expm1(d-))]
          = 1/2 log1p(s0 (expm1(d+) + expm1(d-)) + s0^2 expm1(d+) expm1(d-)),
            expm1(d+) + expm1(d-) = 2 (expm1(Q) cosh P + (cosh P - 1))
    backward, from the OUTPUTS (softplus(z+-) = sp0 + b +- a, sigmoid = 1 - exp(-softplus)), e0 = 1 - s0:
        S1 = (s'(z+) - s'(z-))/2 = e0 e^{-b} sinh a
        Sm = (s'(z+) + s'(z-))/2 = 1 - e0 e^{-b} cosh a          S2 = Sm - s0 = -e0 (expm1(-b) cosh a + (cosh a - 1))
        dz0 = s0 g0 + S2 gb + S1 ga     dP = Sm ga + S1 gb     dQ = Sm gb + S1 ga
    """

    @staticmethod
    def forward(ctx, z0, P, Q, offset):
        s0 = torch.sigmoid(z0)
        h0 = F.softplus(z0) - offset
        em = torch.expm1(Q - P)
        ep = torch.expm1(Q + P)
        a = 0.5 * torch.log1p(2.0 * s0 * torch.exp(Q) * torch.sinh(P) / (1.0 + s0 * em))
        sum_e = 2.0 * (torch.expm1(Q) * torch.cosh(P) + _coshm1(P))
        b = 0.5 * torch.log1p(s0 * sum_e + s0 * s0 * ep * em)
        ctx.save_for_backward(s0, a, b)
        return h0, a, b

    @staticmethod
    def backward(ctx, g0, ga, gb):
        s0, a, b = ctx.saved_tensors
        e0 = 1.0 - s0
        eb = torch.exp(-b)
        S1 = e0 * eb * torch.sinh(a)
        Sm = 1.0 - e0 * eb * torch.cosh(a)
        S2 = -e0 * (torch.expm1(-b) * torch.cosh(a) + _coshm1(a))
        return s0 * g0 + S2 * gb + S1 * ga, Sm * ga + S1 * gb, Sm * gb + S1 * ga, None


def _elu_delta(u, d):
    """elu(u + d) - elu(u) without cancellation."""
    t = u + d
    both_pos = (u > 0) & (t > 0)
    both_neg = (u <= 0) & (t <= 0)
    up = (u <= 0) & (t > 0)
    out = torch.where(both_pos, d, torch.exp(torch.clamp(u, max=0.0)) * torch.expm1(d))
    out = torch.where(up, t - torch.expm1(torch.clamp(u, max=0.0)), out)
    out = torch.where(~both_pos & ~both_neg & ~up, torch.expm1(torch.clamp(t, max=0.0)) - u, out)
    return out


class _EluJet(torch.autograd.Function):
    """(z0, P, Q) -> (h0, a, b) for elu (alpha = 1).  All three arguments in the linear region: a = P, b = Q exactly;
    all three in the exponential region: a = e^{z0} e^Q sinh P, b = e^{z0}(expm1(Q) cosh P + cosh P - 1); otherwise the
    two one-sided differences elu(z0 + d+-) - elu(z0) are combined directly.
    backward from the outputs: s'(z) = 1 (h > 0) or h + 1 (h <= 0) with h+- = h0 + b +- a."""

    @staticmethod
    def forward(ctx, z0, P, Q):
        zp, zm = z0 + Q + P, z0 + Q - P
        h0 = F.elu(z0)
        pos = (z0 > 0) & (zp > 0) & (zm > 0)
        neg = (z0 <= 0) & (zp <= 0) & (zm <= 0)
        dp, dm = _elu_delta(z0, Q + P), _elu_delta(z0, Q - P)
        e = torch.exp(torch.clamp(z0, max=0.0))
        a = torch.where(pos, P, torch.where(neg, e * torch.exp(Q) * torch.sinh(P), 0.5 * (dp - dm)))
        b = torch.where(pos, Q, torch.where(neg, e * (torch.expm1(Q) * torch.cosh(P) + _coshm1(P)), 0.5 * (dp + dm)))
        ctx.save_for_backward(h0, a, b)
        return h0, a, b

    @staticmethod
    def backward(ctx, g0, ga, gb):
        h0, a, b = ctx.saved_tensors
        hp, hm = h0 + b + a, h0 + b - a
        d0 = torch.where(h0 > 0, torch.ones_like(h0), h0 + 1.0)
        dp = torch.where(hp > 0, torch.ones_like(h0), hp + 1.0)
        dm = torch.where(hm > 0, torch.ones_like(h0), hm + 1.0)
        allneg = (h0 <= 0) & (hp <= 0) & (hm <= 0)
        S1 = torch.where(allneg, a, 0.5 * (dp - dm))
        S2 = torch.where(allneg, b, 0.5 * (dp + dm) - d0)
        Sm = S2 + d0
        return d0 * g0 + S2 * gb + S1 * ga, Sm * ga + S1 * gb, Sm * gb + S1 * ga


def act_jet(name, z0, P, Q):
    if name == "softplus":
        return _SoftplusJet.apply(z0, P, Q, 0.0)
    if name == "nsoftplus":
        return _SoftplusJet.apply(z0, P, Q, LN2)
    if name == "elu":
        return _EluJet.apply(z0, P, Q)
    raise ValueError(name)


def mlp_jet(layers, act, x, v, tf32=False):
    """Carry (h0, a, b) through [(W, bias), ...] with activation `act` after every layer.  Layer 1 has b = 0 (Q = 0)."""
    h0, a, b = x, v, None
    for W, bias in layers:
        z0 = linear(h0, W, bias, tf32)
        P = linear(a, W, None, tf32)
        Q = linear(b, W, None, tf32) if b is not None else torch.zeros_like(P)
        h0, a, b = act_jet(act, z0, P, Q)
    return h0, a, b


def quadratic_head_jet(w1, b1, w2, b2, wq, bq, h0, a, b):
    """(E0, dE, sE) of E = ln1(h) ln2(h) + lq(h^2) (DeepEBM/ESM/models.py:154-158, 949-953) at h+- = h0 + b +- a:
        dE = E+ - E- = 2 [A1 (l2 + B2) + A2 (l1 + B1)] + 4 wq.(a (b + h0))
        sE = E+ + E- - 2 E0 = 2 [l1 B2 + B1 l2 + B1 B2 + A1 A2] + 2 wq.(b^2 + a^2 + 2 h0 b)
    with l_i = w_i.h0 + b_i, A_i = w_i.a, B_i = w_i.b."""
    l1, l2 = h0 @ w1 + b1, h0 @ w2 + b2
    A1, A2, B1, B2 = a @ w1, a @ w2, b @ w1, b @ w2
    E0 = l1 * l2 + (h0 * h0) @ wq + bq
    dE = 2.0 * (A1 * (l2 + B2) + A2 * (l1 + B1)) + 4.0 * ((a * (b + h0)) @ wq)
    sE = 2.0 * (l1 * B2 + B1 * l2 + B1 * B2 + A1 * A2) + 2.0 * ((b * b + a * a + 2.0 * h0 * b) @ wq)
    return E0, dE, sE


def energy_jet(net, x, v, tf32=False):
    """(E0, dE, sE) for oracle.nets.QuadraticMLP / SoftplusMLP."""
    x, v = x.reshape(x.shape[0], -1), v.reshape(v.shape[0], -1)
    if hasattr(net, "ln1"):
        layers = [(getattr(net, f"fc{i + 1}").weight, getattr(net, f"fc{i + 1}").bias) for i in range(net.n_hidden)]
        h0, a, b = mlp_jet(layers, net.act, x, v, tf32)
        return quadratic_head_jet(net.ln1.weight.view(-1), net.ln1.bias, net.ln2.weight.view(-1), net.ln2.bias,
                                  net.lq.weight.view(-1), net.lq.bias, h0, a, b)
    lins = [m for m in net.net if isinstance(m, torch.nn.Linear)]
    h0, a, b = mlp_jet([(m.weight, m.bias) for m in lins[:-1]], "softplus", x, v, tf32)
    w, bias = lins[-1].weight.view(-1), lins[-1].bias
    return h0 @ w + bias, 2.0 * (a @ w), 2.0 * (b @ w)


def fd_ssm_jet(net, x, v, scale, c1=2.0, reduce="mean", tf32=False):
    """The value of oracle.fd_losses.fd_ssm (a1 / a7 / a5) through the difference-propagating evaluation."""
    E0, dE, sE = energy_jet(net, x, v, tf32)
    per = dE * dE / 8.0 - sE + (c1 - 2.0) * E0
    return (per.mean() if reduce == "mean" else per.sum()) * scale


def mlp_jet_trace(layers, act, x, v, tf32=False):
    """mlp_jet that also returns, per layer, the pre-activation streams (z0, P, Q) -- with retain_grad() set, so that after
    a backward their .grad are the (dz0, dP, dQ) the chain kernel stores -- and the outputs (h0, a, b)."""
    h0, a, b = x, v, None
    pre, post = [], []
    for W, bias in layers:
        z0 = linear(h0, W, bias, tf32)
        P = linear(a, W, None, tf32)
        Q = linear(b, W, None, tf32) if b is not None else torch.zeros_like(P, requires_grad=True)
        for t in (z0, P, Q):
            t.retain_grad()
        h0, a, b = act_jet(act, z0, P, Q)
        pre.append((z0, P, Q))
        post.append((h0, a, b))
    return pre, post
