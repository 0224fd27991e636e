"""Run the oracle restatement (oracle/fd_losses.py) on a built case (oracle/cases.py).
TEST INFRASTRUCTURE."""
import math

import torch

from . import fd_losses as F


def directions(c):
    """Per-function direction scaling (see fd_losses.sphere_direction)."""
    fn, eps, u = c["fn"], c["eps"], c["u"]
    if fn == "a3":
        return F.sphere_direction(u, eps, math.sqrt(c["D"]))
    if c.get("noise_type") == "radermacher":
        return u.sign() * eps              # SSM/losses/efficient_sm.py:10-11
    if fn == "a12":
        return F.sphere_direction(u, 0.1)  # eps hard-coded, ncsn/losses/sliced_sm.py:220
    return F.sphere_direction(u, eps)


def oracle_loss(c, net=None):
    net = net or c["net"]
    fn, x, eps, D, B = c["fn"], c["x"], c["eps"], c["D"], c["B"]
    if fn == "a17_anneal_vr":
        return F.anneal_ssm_vr(net, x, c["n"], c["u"], c["labels"], c["sigmas"], c["anneal_power"], c["n_particles"])
    if fn == "a17_exact":
        return F.exact_sm(net, x).mean()
    if fn == "a17_exact_deepebm":      # loss_train.py:11-26: the same quantity on samples.view(-1, dim), mean
        return F.exact_sm(net, x.reshape(B, -1)).mean().detach()   # an evaluator: value only
    if fn == "a17_mdsm_baseline":
        return F.mdsm_exact(net, x, c["n"], c["sig_b"], c["sigma02"], c["loss_coe"])
    if fn == "a17_mdsm_tracetrick":    # RNG order: noise, then projection (loss_train.py:49-50)
        return F.mdsm_exact(net, x, c["n"], c["sig_b"], c["sigma02"], c["loss_coe"], projection=c["u"])
    if fn.startswith("a17"):   # exact SSM (HVP) comparison losses: Gaussian (DeepEBM) or sign (SSM repo) projections
        xf, uf = x.reshape(B, -1), c["u"].reshape(B, -1)
        if fn in ("a17_sliced", "a17_sliced_vr"):
            uf = uf.sign()
        return F.ssm_hvp(net, xf, uf, vr=fn in ("a17_ssm_vr", "a17_sliced_vr"))
    v = directions(c)
    if fn == "a1":
        return F.fd_ssm(net, x, v, D / eps ** 2)
    if fn == "a7":
        xf = x.reshape(B, -1)
        return F.fd_ssm(net, xf, v.reshape(B, -1), 1.0 / eps ** 2)
    if fn == "a8":
        return F.fd_ssm_nop(net, x.reshape(B, -1), v.reshape(B, -1), 1.0 / eps ** 2)
    if fn == "a2":
        return F.fd_one_sided(net, x, v)
    if fn == "a5":
        return F.fd_ssm(net, x, v, 1.0, c1=c["mle_ratio"], reduce="sum")
    if fn == "a3":
        return F.fd_dsm(net, x, c["n"] * c["sigma"], v, c["sigma"], eps)
    if fn == "a6":
        return F.dsm_tracetrick_fd(net, x, c["n"] * c["sigma"], v, c["sigma"], eps)
    if fn == "a4":
        return F.fd_ssmvr_score(net, x, v, eps)
    if fn in ("a9", "a10"):
        return F.mdsm_fd(net, x, c["n"], v, c["sig_b"], c["sigma02"], eps, c["loss_coe"], nop=fn == "a10")
    if fn == "a11":
        return F.anneal_fd_ssmvr(net, x, c["n"], v, c["labels"], c["sigmas"], eps, 32. * 32. * 3., 2.0)
    if fn == "a12":
        return F.anneal_fd_ssmvr(net, x, c["n"], v, c["labels"], c["sigmas"], 0.1, 0.0, 2.0, projected=True)
    if fn == "a13":
        return F.anneal_fd_ssmvr(net, x, c["n"], v, c["labels"], c["sigmas"], eps, float(D), c["anneal_power"])
    raise KeyError(fn)


def loss_and_grad(c, net=None):
    net = net or c["net"]
    for p in net.parameters():
        p.grad = None
    loss = oracle_loss(c, net)
    if loss.requires_grad:
        loss.backward()
    g = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                   for p in net.parameters()])
    return loss.detach(), g
