"""CPU restatement (torch CPU ops, any float dtype) of the reference's FD losses.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Every function takes the random
draws as explicit tensors (``u`` = un-normalised Gaussian direction, ``n`` =
unit Gaussian DSM noise) in the order the reference draws them, so that the
oracle, the reference and the CUDA path can be fed identical numbers.  File:line
citations are relative to /root/reference.
"""
import math

import torch


def sphere_direction(u, eps, scale=1.0):
    """v = u / ||u||_2 (over all non-batch dims) * eps * scale.

    SSM/losses/efficient_sm.py:53-56, DeepEBM/ESM/loss_train.py:217-218,
    ncsn/losses/sliced_sm.py:183-184, ncsnv2/losses/dsm.py:107-110.
    """
    flat = u.reshape(u.shape[0], -1)
    nrm = torch.sqrt(torch.sum(flat * flat, dim=-1))
    nrm = nrm.reshape(-1, *([1] * (u.dim() - 1)))
    return u / nrm * (eps * scale)


def combine_energy(E, B, scale, c1=2.0, reduce="mean"):
    """Central-difference combine, energy form.

    E = energy([x, x+v, x-v]) of shape (3B,).  per-sample
    (E2-E3)^2/8 - E2 - E3 + c1*E1; SSM/losses/efficient_sm.py:61-71 (c1=2, mean,
    *D/eps^2), DeepEBM/ESM/loss_train.py:225-232 (c1=2, mean, /eps^2),
    efficient_sm.py:113-121 (c1=mle_ratio, sum, no scale).
    """
    e1, e2, e3 = E[:B], E[B:2 * B], E[2 * B:]
    d = e2 - e3
    per = d * d / 8 - e2 - e3 + c1 * e1
    if reduce == "mean":
        return per.mean() * scale, per * scale
    return per.sum() * scale, per * scale


def fd_ssm(energy_net, x, v, scale, c1=2.0, reduce="mean"):
    """a1 efficient_score_matching_conjugate (efficient_sm.py:47-73, scale=D/eps^2),
    a7 ssm_fd (loss_train.py:212-234, scale=1/eps^2), a5 MLE variant (:101-123)."""
    B = x.shape[0]
    cat = torch.cat([x, x + v, x - v], 0)
    return combine_energy(energy_net(cat), B, scale, c1, reduce)[0]


def fd_ssm_nop(energy_net, x, v, scale):
    """a8 ssm_fd_nop: three separate forwards (loss_train.py:237-255)."""
    e1, e2, e3 = energy_net(x), energy_net(x + v), energy_net(x - v)
    d = e2 - e3
    return (d * d / 8 - e2 - e3 + 2 * e1).mean() * scale


def fd_one_sided(energy_net, x, v):
    """a2 single_efficient_score_matching (efficient_sm.py:7-24): d=E(x)-E(x+v),
    sum(d^2 + 4d)."""
    B = x.shape[0]
    out = energy_net(torch.cat([x, x + v], 0))
    d = out[:B] - out[B:]
    return torch.sum(d * d + 4 * d)


def fd_dsm(energy_net, x, n_scaled, v, sigma, eps):
    """a3 dsm_fd (efficient_sm.py:27-44).  n_scaled = sigma*N(0,I) ("vector"),
    v = eps*sqrt(D)*u/||u|| ("noise")."""
    B = x.shape[0]
    xt = x + n_scaled
    logp = -energy_net(torch.cat([xt + v, xt - v]))
    dlogpv = sigma ** 2 * (logp[:B] - logp[B:]) * 0.5
    kernel = torch.sum(n_scaled * v, dim=-1)
    return ((dlogpv + kernel) ** 2).mean() / (2.0 * eps ** 2)


def dsm_tracetrick_fd(energy_net, x, n_scaled, v, sigma, eps):
    """a6 dsm_tracetrick_FD (SSM/losses/dsm.py:122-141).  NB the reference perturbs
    ``samples`` (not the noised input) at :131 -- restated as is."""
    B = x.shape[0]
    out = -energy_net(torch.cat([x + v, x - v], 0))
    dlogp = sigma ** 2 * (out[:B] - out[B:])
    kernel = 2 * torch.sum(n_scaled * v, dim=-1)
    return ((dlogp + kernel) ** 2).mean() * v.shape[-1] / (8 * eps * eps)


def combine_score(s, v, B, k, weights=None, projected=False):
    """Score-net combine.  s = scorenet([x+v, x-v]) reshaped (2B, D); per-sample
    sum_d[(s1+s2)^2/8 + (s1-s2)*v*k] * w_b  (efficient_sm.py:86-96,
    ncsn/losses/sliced_sm.py:192-205, ncsnv2/losses/dsm.py:118-131); the non-VR
    variant (projected=True, ncsn/losses/sliced_sm.py:234-238) is
    sum_d[((s1+s2)*v)^2/8 + (s1-s2)*v/2] * w_b."""
    s = s.reshape(2 * B, -1)
    v = v.reshape(B, -1)
    s1, s2 = s[:B], s[B:]
    if projected:
        per = torch.sum(((s1 + s2) * v) ** 2 / 8.0 + (s1 - s2) * v / 2.0, dim=-1)
    else:
        g1, g2 = s1 + s2, s1 - s2
        per = torch.sum(g1 * g1 / 8.0 + g2 * v * k, dim=-1)
    if weights is not None:
        per = per * weights
    return per.mean(dim=0), per


def fd_ssmvr_score(scorenet, x, v, eps):
    """a4 esm_scorenet_VR (efficient_sm.py:76-98); dim = samples.shape[-1]."""
    B, dim = x.shape[0], x.shape[-1]
    s = scorenet(torch.cat([x + v, x - v], 0))
    return combine_score(s, v, B, dim / (2 * eps * eps))[0]


def mdsm_fd(energy_net, x, n, v, sigmas, sigma02, eps, loss_coe, nop=False):
    """a9 mdsm_fd / a10 mdsm_fd_nop (loss_train.py:96-118, :151-170).
    x (B,D); n unit Gaussian; sigmas (B,1); v norm eps."""
    B, dim = x.shape
    xn = x + sigmas * n
    if nop:
        logp1, logp2 = -energy_net(xn + v), -energy_net(xn - v)
    else:
        logp = -energy_net(torch.cat([xn + v, xn - v], 0))
        logp1, logp2 = logp[:B], logp[B:]
    first = (logp1 - logp2) * 0.5
    second = torch.sum(v * (xn - x), dim=-1) / sigma02
    per = (first + second) ** 2 / (sigmas.view(-1) ** 2)
    return (per * (loss_coe / eps ** 2 * dim)).mean()


def anneal_fd_ssmvr(scorenet, x, n, v, labels, sigmas, eps, data_dim, anneal_power=2.0,
                    projected=False, hook=None):
    """a11 ncsn anneal_ESM_scorenet_VR (ncsn/losses/sliced_sm.py:174-207;
    data_dim=3072 hard-coded at :202, power 2), a12 anneal_ESM_scorenet (:210-240,
    projected=True), a13 ncsnv2 anneal_ESM_scorenet_VR (ncsnv2/losses/dsm.py:77-136;
    data_dim=prod(shape[1:]), sigma**anneal_power, hook).  n_particles=1."""
    B = x.shape[0]
    used = sigmas[labels].view(B, *([1] * (x.dim() - 1)))
    xt = x + n * used
    s = scorenet(torch.cat([xt + v, xt - v], 0), torch.cat([labels, labels], 0))
    w = used.reshape(B) ** anneal_power
    loss, per = combine_score(s, v, B, data_dim / (2.0 * eps * eps), w, projected)
    if hook is not None:
        hook(per, labels)
    return loss


# --- comparison baselines ("exact SSM" side of config 2) -------------------------------

def ssm_hvp(energy_net, x, u, vr=False):
    """a17 loss_train.ssm / ssm_vr (loss_train.py:173-209): Hessian-vector-product SSM."""
    x = x.detach().clone().requires_grad_(True)
    logp = -energy_net(x)
    score = torch.autograd.grad(logp.sum(), x, create_graph=True)[0]
    if vr:
        loss1 = (score ** 2).sum(dim=-1) * 0.5
    else:
        loss1 = ((score * u).sum(dim=-1) ** 2) * 0.5
    hvp = torch.autograd.grad((score * u).sum(), x, create_graph=True)[0]
    return (loss1 + (hvp * u).sum(dim=-1)).mean()


def anneal_ssm_vr(scorenet, x, n, u, labels, sigmas, anneal_power=2.0, n_particles=1, hook=None):
    """a17 ncsnv2 anneal_sliced_score_estimation_vr (ncsnv2/losses/dsm.py:33-74): x~ = x + n*sigma[labels], duplicated
    n_particles times; per row 0.5 ||s||^2 + u . d(u . s)/dx with GAUSSIAN u (the "use Rademacher" comment at :52 is not
    what :53 does); particle mean, weight sigma**anneal_power, batch mean."""
    B = x.shape[0]
    used = sigmas[labels].view(B, *([1] * (x.dim() - 1)))
    xt = x + n * used
    dup = xt.unsqueeze(0).expand(n_particles, *x.shape).contiguous().view(-1, *x.shape[1:]).detach().requires_grad_(True)
    dup_labels = labels.unsqueeze(0).expand(n_particles, *labels.shape).contiguous().view(-1)
    s = scorenet(dup, dup_labels)
    grad2 = torch.autograd.grad(torch.sum(s * u), dup, create_graph=True)[0]
    loss1 = (s.reshape(dup.shape[0], -1) ** 2).sum(-1) / 2.0
    loss2 = (u * grad2).reshape(dup.shape[0], -1).sum(-1)
    per = (loss1.view(n_particles, -1).mean(0) + loss2.view(n_particles, -1).mean(0)) * used.reshape(B) ** anneal_power
    if hook is not None:
        hook(per, labels)
    return per.mean()


def exact_sm(energy_net, x):
    """a17 exact_score_matching(train=True) (SSM/losses/score_matching.py:48-73): per sample
    0.5 ||grad logp||^2 + sum_i d^2 logp / dx_i^2 (one backward pass per input dimension)."""
    x = x.detach().clone().requires_grad_(True)
    score = torch.autograd.grad(-energy_net(x).sum(), x, create_graph=True)[0]
    per = (score ** 2).sum(-1) / 2.0
    for i in range(x.shape[1]):
        per = per + torch.autograd.grad(score[:, i].sum(), x, create_graph=True)[0][:, i]
    return per


def mdsm_exact(energy_net, x, n, sigmas, sigma02, loss_coe, projection=None):
    """a17 DeepEBM mdsm_baseline (loss_train.py:29-44) and, with `projection`, mdsm_tracetrick (:47-63): the multiscale
    denoising score matching objective that mdsm_fd approximates -- ((x - xt)/sigma0^2 + grad E(xt)) squared (or its
    squared projection), weighted loss_coe / sigma_b^2, batch mean."""
    B = x.shape[0]
    xf = x.reshape(B, -1)
    xt = (xf + sigmas * n.reshape(B, -1)).detach().requires_grad_(True)
    grad_x = torch.autograd.grad(energy_net(xt).sum(), xt, create_graph=True)[0]
    r = (xf - xt) / sigma02 + grad_x
    per = (r ** 2).sum(-1) if projection is None else ((r * projection.reshape(B, -1)).sum(-1)) ** 2
    return (per * (loss_coe / sigmas.view(-1) ** 2)).mean()


def quadratic_energy_fd_ssm_exact(A, b, x, vhat):
    """Analytic fixture: E(x)=0.5 x^T A x + b^T x (A symmetric).  For v=eps*vhat the
    a1 loss / D == mean(0.5*(vhat^T(Ax+b))^2 - vhat^T A vhat) for ANY eps
    (SURVEY.md §8c iii)."""
    g = x @ A + b
    proj = (g * vhat).sum(-1)
    curv = ((vhat @ A) * vhat).sum(-1)
    return (0.5 * proj ** 2 - curv).mean()
