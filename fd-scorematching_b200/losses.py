"""Drop-in FD score-matching losses: the reference's call signatures over the sm_100a kernels.

Each adapter keeps the reference function's name, argument order, defaults and error behaviour
(file:line cited per function, relative to the reference repo).  Generic mode: kernels (1)+(2)
build the stacked batch, the caller's network (any callable) runs in PyTorch, kernel (4) turns the
outputs into the loss and feeds the analytic gradient back into autograd.  When the network is one
of our fused MLP energies (``mlp.FusedEnergyMLP``) its forward/backward are the tcgen05 GEMMs.

Extra keyword-only arguments (``u=``, ``dsm_noise=``) inject the random draws the reference takes
from the global generator, in its RNG order, for parity tests; without them the draws come from
the in-kernel Philox streams (``ops.manual_seed``).
"""
import math

import torch

from . import ops
from ._lib import ROW_MINUS, ROW_PLUS, ROW_X


class Flags:
    """Stand-in for DeepEBM's process-global ``Utils.flags.FLAGS`` (Utils/flags.py:27-77), which
    the reference's DeepEBM losses read directly (loss_train.py:6,97-98,213-215)."""
    batch_size = None   # None -> samples.shape[0]
    esm_eps = 0.05      # configs/default.yaml:28
    loss_coe = 1e-4     # configs/default.yaml:24


FLAGS = Flags()


def _flat(samples):
    if samples.dim() < 2:
        raise ValueError("samples must have a batch dimension and at least one data dimension")
    B = samples.shape[0]
    return samples.reshape(B, -1), B, int(samples[0].numel()) if B > 0 else int(math.prod(samples.shape[1:]))


def _check_particles(n_particles, what):
    if n_particles != 1:
        # the reference multiplies a (n_particles*B,) loss by a (B,) sigma vector: shape error for != 1
        raise RuntimeError(f"{what}: only n_particles=1 is shape-correct in the reference "
                           "(ncsn/losses/sliced_sm.py:205, ncsnv2/losses/dsm.py:131)")


# ------------------------------------------------------------------------------------------------
# energy form
# ------------------------------------------------------------------------------------------------
def _fd_energy(energy_net, x2, *, v=None, u=None, eps, scale, c1=2.0, mean=True, rows3=True, detach=False,
               mode=0, want_per_sample=False):
    B, D = x2.shape
    mask = (ROW_X | ROW_PLUS | ROW_MINUS) if rows3 else (ROW_X | ROW_PLUS)
    if v is not None:
        stacked, _, _ = ops.stack(x2, row_mask=mask, scale=1.0, direction=v, normalize=False)
    else:
        stacked, _, _ = ops.stack(x2, row_mask=mask, scale=eps, direction=u, normalize=True)
    if detach:
        with torch.no_grad():
            E = energy_net(stacked)
            return ops.CombineEnergy.apply(E, B, mode, float(c1), float(scale), mean, want_per_sample)
    E = energy_net(stacked)
    return ops.CombineEnergy.apply(E, B, mode, float(c1), float(scale), mean, want_per_sample)


def efficient_score_matching_conjugate(energy_net, samples, noise=None, eps=0.1, noise_type="sphere", detach=False,
                                       *, u=None):
    """FD-SSM, energy form -- SSM/losses/efficient_sm.py:47-73.  ``noise`` (already-normalised v) is
    the reference's own injection hook; ``noise_type`` is ignored there and here."""
    dim = samples.shape[1]
    return _fd_energy(energy_net, samples, v=noise, u=u, eps=eps, scale=dim / eps ** 2, detach=detach)[0]


def single_efficient_score_matching(energy_net, samples, eps=0.01, noise_type="sphere", *, u=None):
    """One-sided 2-point FD -- SSM/losses/efficient_sm.py:7-24 (sum of d^2+4d, no /eps^2).  'sphere': eps*u/||u||
    inside the stack kernel; 'radermacher' (the reference's spelling, :10-11): v = sign(u)*eps, handed to the stack
    kernel as an already-scaled direction; any other value leaves the Gaussian draw unscaled, as the reference does."""
    x2 = samples.reshape(samples.shape[0], -1)
    if noise_type == "sphere":
        return _fd_energy(energy_net, x2, u=u, eps=eps, scale=1.0, mean=False, rows3=False, mode=1)[0]
    if u is not None:
        u = u.reshape(x2.shape)
    v = _projection_vectors(x2, "radermacher" if noise_type == "radermacher" else "gaussian", u)
    if noise_type == "radermacher":
        v = v * eps
    return _fd_energy(energy_net, x2, v=v, eps=eps, scale=1.0, mean=False, rows3=False, mode=1)[0]


def MLE_efficient_score_matching_conjugate(energy_net, samples, eps=0.01, mle_ratio=10, noise_type="sphere", *, u=None):
    """SSM/losses/efficient_sm.py:101-123: c1 = mle_ratio, summed, no /eps^2.  noise_type as in
    single_efficient_score_matching (:105-110): 'sphere' = eps*u/||u||, 'radermacher' = sign(u)*eps, anything else
    leaves the Gaussian draw unscaled."""
    x2 = samples.reshape(samples.shape[0], -1)
    if noise_type == "sphere":
        return _fd_energy(energy_net, x2, u=u, eps=eps, scale=1.0, c1=mle_ratio, mean=False)[0]
    if u is not None:
        u = u.reshape(x2.shape)
    v = _projection_vectors(x2, "radermacher" if noise_type == "radermacher" else "gaussian", u)
    if noise_type == "radermacher":
        v = v * eps
    return _fd_energy(energy_net, x2, v=v, eps=eps, scale=1.0, c1=mle_ratio, mean=False)[0]


def ssm_fd(energy, samples, *, eps=None, batch_size=None, u=None):
    """DeepEBM FD-SSM -- DeepEBM/ESM/loss_train.py:212-234 (no *D factor; eps=FLAGS.esm_eps;
    samples.view(FLAGS.batch_size, dim))."""
    eps = FLAGS.esm_eps if eps is None else eps
    bs = batch_size or FLAGS.batch_size or samples.shape[0]
    x2 = samples.reshape(bs, -1)
    if u is not None:
        u = u.reshape(bs, -1)
    return _fd_energy(energy, x2, u=u, eps=eps, scale=1.0 / eps ** 2)[0]


def ssm_fd_nop(energy, samples, *, eps=None, batch_size=None, u=None):
    """Timing-ablation twin with three separate forwards -- loss_train.py:237-255.  Same stacked
    build; the three row blocks are handed to the network one at a time."""
    eps = FLAGS.esm_eps if eps is None else eps
    bs = batch_size or FLAGS.batch_size or samples.shape[0]
    x2 = samples.reshape(bs, -1)
    if u is not None:
        u = u.reshape(bs, -1)
    stacked, _, _ = ops.stack(x2, row_mask=7, scale=eps, direction=u, normalize=True)
    E = torch.cat([energy(stacked[:bs]), energy(stacked[bs:2 * bs]), energy(stacked[2 * bs:])], 0)
    return ops.CombineEnergy.apply(E, bs, 0, 2.0, 1.0 / eps ** 2, True, False)[0]


# ------------------------------------------------------------------------------------------------
# DSM energy form
# ------------------------------------------------------------------------------------------------
def _fd_dsm(energy_net, x2, sigma_b, *, eps_scale, alpha, beta, gamma, weight_by_sigma, perturb, dot_mode, u, dsm_noise,
            separate=False):
    B, D = x2.shape
    stacked, _, dot = ops.stack(x2, row_mask=ROW_PLUS | ROW_MINUS, scale=eps_scale, direction=u, normalize=True,
                                noise=dsm_noise, sigma=sigma_b, perturb=perturb, dot_mode=dot_mode, want_dot=True)
    if separate:
        E = torch.cat([energy_net(stacked[:B]), energy_net(stacked[B:])], 0)
    else:
        E = energy_net(stacked)
    return ops.CombineDsm.apply(E, dot, sigma_b if weight_by_sigma else None, B, float(alpha), float(beta), float(gamma))


def dsm_fd(energy_net, samples, sigma=1, eps=0.001, *, u=None, dsm_noise=None):
    """FD-DSM -- SSM/losses/efficient_sm.py:27-44.  RNG order: DSM noise, then direction."""
    B, dim = samples.shape[0], samples.shape[1]
    sig = torch.full((B,), float(sigma), device=samples.device, dtype=torch.float32)
    return _fd_dsm(energy_net, samples, sig, eps_scale=eps * dim ** 0.5, alpha=sigma ** 2 * 0.5, beta=1.0,
                   gamma=1.0 / (2.0 * eps ** 2), weight_by_sigma=False, perturb=True, dot_mode=0, u=u, dsm_noise=dsm_noise)


def dsm_tracetrick_FD(energy_net, samples, sigma=1, eps=0.1, *, u=None, dsm_noise=None):
    """SSM/losses/dsm.py:122-141.  As in the reference the stacked input perturbs ``samples``, not the
    noised input (:131)."""
    B, dim = samples.shape[0], samples.shape[-1]
    sig = torch.full((B,), float(sigma), device=samples.device, dtype=torch.float32)
    return _fd_dsm(energy_net, samples, sig, eps_scale=eps, alpha=sigma ** 2, beta=2.0, gamma=dim / (8.0 * eps * eps),
                   weight_by_sigma=False, perturb=False, dot_mode=0, u=u, dsm_noise=dsm_noise)


def mdsm_fd(energy, x_real, sigmas, sigma02, dim, *, eps=None, batch_size=None, loss_coe=None, u=None, dsm_noise=None,
            _separate=False):
    """FD multiscale DSM -- DeepEBM/ESM/loss_train.py:96-118; sigmas is the (B,1) per-batch-position
    schedule of train_dsm.py:50-57."""
    eps = FLAGS.esm_eps if eps is None else eps
    coe = FLAGS.loss_coe if loss_coe is None else loss_coe
    bs = batch_size or FLAGS.batch_size or x_real.shape[0]
    x2 = x_real.reshape(bs, dim)
    sig = sigmas.reshape(-1).to(torch.float32)
    if sig.numel() != bs:
        raise RuntimeError(f"sigmas has {sig.numel()} entries for batch size {bs}")
    if u is not None:
        u = u.reshape(bs, dim)
    if dsm_noise is not None:
        dsm_noise = dsm_noise.reshape(bs, dim)
    return _fd_dsm(energy, x2, sig, eps_scale=eps, alpha=0.5, beta=1.0 / sigma02, gamma=coe / eps ** 2 * dim,
                   weight_by_sigma=True, perturb=True, dot_mode=1, u=u, dsm_noise=dsm_noise, separate=_separate)


def mdsm_fd_nop(energy, x_real, sigmas, sigma02, dim, **kw):
    """loss_train.py:151-170: two separate forwards."""
    return mdsm_fd(energy, x_real, sigmas, sigma02, dim, _separate=True, **kw)


# ------------------------------------------------------------------------------------------------
# score-net form
# ------------------------------------------------------------------------------------------------
def _fd_score(scorenet, samples, *, eps_dir, k, sigma_table=None, labels=None, power=2.0, projected=False, u=None,
              dsm_noise=None, hook=None):
    x2, B, D = _flat(samples)
    if u is not None:
        u = u.reshape(B, D)
    if dsm_noise is not None:
        dsm_noise = dsm_noise.reshape(B, D)
    sig = sigma_table.to(torch.float32).contiguous() if sigma_table is not None else None
    stacked, v, _ = ops.stack(x2, row_mask=ROW_PLUS | ROW_MINUS, scale=eps_dir, direction=u, normalize=True,
                              noise=dsm_noise, sigma=sig, labels=labels, perturb=True, want_v=True)
    cat_input = stacked.view(2 * B, *samples.shape[1:])
    if labels is not None:
        out = scorenet(cat_input, torch.cat([labels, labels], 0))
    else:
        out = scorenet(cat_input)
    loss, per = ops.CombineScore.apply(out, v, sig, labels, float(power), B, D, float(k), projected, True)
    if hook is not None:
        hook(per, labels)
    return loss


def esm_scorenet_VR(scorenet, samples, eps=0.1, *, u=None):
    """FD-SSMVR for score nets -- SSM/losses/efficient_sm.py:76-98 (dim = samples.shape[-1])."""
    dim = samples.shape[-1]
    return _fd_score(scorenet, samples, eps_dir=eps, k=dim / (2 * eps * eps), u=u)


def ncsn_anneal_ESM_scorenet_VR(scorenet, samples, labels, sigmas, n_particles=1, eps=0.1, *, u=None, dsm_noise=None):
    """ncsn annealed FD-SSMVR -- ncsn/losses/sliced_sm.py:174-207 (constant 32*32*3 hard-coded at
    :202, sigma^2 weighting :205)."""
    _check_particles(n_particles, "anneal_ESM_scorenet_VR")
    return _fd_score(scorenet, samples, eps_dir=eps, k=(32. * 32. * 3.) / (2. * eps * eps), sigma_table=sigmas,
                     labels=labels, power=2.0, u=u, dsm_noise=dsm_noise)


def ncsn_anneal_ESM_scorenet(scorenet, samples, labels, sigmas, n_particles=1, *, u=None, dsm_noise=None):
    """ncsn annealed FD-SSM (non-VR) -- ncsn/losses/sliced_sm.py:210-240 (eps fixed 0.1, no /eps^2)."""
    _check_particles(n_particles, "anneal_ESM_scorenet")
    return _fd_score(scorenet, samples, eps_dir=0.1, k=0.0, sigma_table=sigmas, labels=labels, power=2.0,
                     projected=True, u=u, dsm_noise=dsm_noise)


def anneal_ESM_scorenet_VR(scorenet, samples, sigmas, labels=None, anneal_power=2.0, hook=None, n_particles=1, eps=0.1,
                           *, u=None, dsm_noise=None):
    """ncsnv2 annealed FD-SSMVR -- ncsnv2/losses/dsm.py:77-136 (data_dim = prod(shape[1:]), labels
    drawn inside when None, sigma**anneal_power, hook(per_sample, labels))."""
    _check_particles(n_particles, "anneal_ESM_scorenet_VR")
    data_dim = int(math.prod(samples.shape[1:]))
    if labels is None:
        labels = torch.randint(0, len(sigmas), (samples.shape[0],), device=samples.device)
    return _fd_score(scorenet, samples, eps_dir=eps, k=data_dim / (2.0 * eps * eps), sigma_table=sigmas, labels=labels,
                     power=anneal_power, u=u, dsm_noise=dsm_noise, hook=hook)


# ------------------------------------------------------------------------------------------------
# a17: the exact (Hessian-vector-product) sliced score matching losses the FD objectives are compared against
# (BASELINE.json config 2 "FD-SSM vs exact SSM").  Generic mode only: the two autograd.grad(create_graph=True) passes
# run through the caller's torch network; for our fused MLP energies `ssm` runs in forward mode inside the chain kernel
# (mlp.FusedEnergyMLP.exact_ssm_loss_and_grad, no double backward);
# the projection vectors come from kernel (1) (Philox normals; sign / sphere variants derived from the same draw).
# ------------------------------------------------------------------------------------------------
def _projection_vectors(x2, noise_type, u):
    B, D = x2.shape
    if u is None:
        seed, offset, row0 = ops.next_draw()
        u = ops.philox_normal(B, D, 0, seed, offset, row0, device=x2.device)
    if noise_type == "radermacher":      # the reference's spelling (SSM/losses/sliced_sm.py:12)
        return u.sign()
    if noise_type == "sphere":
        return ops.direction(B, D, math.sqrt(D), u=u)
    if noise_type == "gaussian":
        return u
    raise ValueError("Noise type not implemented")


def _ssm_hvp(energy_net, x, vectors, vr, detach=False):
    x = x.requires_grad_(True)
    logp = -energy_net(x).sum()
    grad1 = torch.autograd.grad(logp, x, create_graph=True)[0]
    gradv = torch.sum(grad1 * vectors)
    loss1 = (torch.norm(grad1, dim=-1) ** 2 if vr else torch.sum(grad1 * vectors, dim=-1) ** 2) * 0.5
    grad2 = torch.autograd.grad(gradv, x, create_graph=True)[0]
    loss2 = torch.sum(vectors * grad2, dim=-1)
    if detach:
        loss1, loss2 = loss1.detach(), loss2.detach()
    return loss1 + loss2, grad1, grad2


def single_sliced_score_matching(energy_net, samples, noise=None, detach=False, noise_type="radermacher", *, u=None):
    """SSM/losses/sliced_sm.py:7-34: 0.5 (v.grad logp)^2 + v.grad(v.grad logp), one projection; returns
    (loss, grad1, grad2) like the reference."""
    vectors = noise if noise is not None else _projection_vectors(samples.reshape(samples.shape[0], -1), noise_type, u).view_as(samples)
    per, g1, g2 = _ssm_hvp(energy_net, samples, vectors, vr=False, detach=detach)
    return per.mean(), g1, g2


def sliced_VR_score_matching(energy_net, samples, noise=None, detach=False, noise_type="radermacher", *, u=None):
    """SSM/losses/sliced_sm.py:37-62 (variance-reduced: 0.5 ||grad logp||^2 replaces the squared projection); the reference
    accepts 'radermacher' and 'gaussian' here."""
    if noise is None and noise_type not in ("radermacher", "gaussian"):
        raise ValueError("Noise type not implemented")
    vectors = noise if noise is not None else _projection_vectors(samples.reshape(samples.shape[0], -1), noise_type, u).view_as(samples)
    per, g1, g2 = _ssm_hvp(energy_net, samples, vectors, vr=True, detach=detach)
    return per.mean(), g1, g2


def ssm(energy, x, *, batch_size=None, u=None):
    """DeepEBM/ESM/loss_train.py:173-189 (Gaussian projections, x viewed as (FLAGS.batch_size, dim))."""
    bs = batch_size if batch_size is not None else (FLAGS.batch_size or x.shape[0])
    x2 = x.reshape(bs, -1)
    if getattr(energy, "formulation", None) == "jet":      # mlp.FusedEnergyMLP: forward mode in the chain kernel (N2)
        return energy.exact_ssm_loss(x2, u=u)
    per, _, _ = _ssm_hvp(energy, x2, _projection_vectors(x2, "gaussian", u), vr=False)
    return per.mean()


def ssm_vr(energy, x, mean=True, *, batch_size=None, u=None):
    """DeepEBM/ESM/loss_train.py:192-209."""
    bs = batch_size if batch_size is not None else (FLAGS.batch_size or x.shape[0])
    x2 = x.reshape(bs, -1)
    per, _, _ = _ssm_hvp(energy, x2, _projection_vectors(x2, "gaussian", u), vr=True)
    return per.mean() if mean is True else per


def anneal_sliced_score_estimation_vr(scorenet, samples, sigmas, labels=None, anneal_power=2.0, hook=None, n_particles=1,
                                      *, u=None, dsm_noise=None):
    """ncsnv2/losses/dsm.py:33-74: the exact (double-backward) annealed SSM-VR objective that a13 replaces.  The noise
    and the Gaussian projections (:53; one per particle) come from kernel (1)'s Philox streams; the score net and its
    vector-Jacobian product run in torch.  Unlike the FD forms this one is shape-correct for n_particles > 1
    (per-particle means at :64-65)."""
    x2, B, D = _flat(samples)
    if labels is None:
        labels = torch.randint(0, len(sigmas), (B,), device=samples.device)
    if dsm_noise is None:
        seed, offset, row0 = ops.next_draw()
        dsm_noise = ops.philox_normal(B, D, 1, seed, offset, row0, device=samples.device)
    if u is None:
        seed, offset, row0 = ops.next_draw()
        u = ops.philox_normal(n_particles * B, D, 0, seed, offset, row0, device=samples.device)
    used = sigmas[labels].view(B, 1).to(x2.dtype)
    xt = x2 + dsm_noise.reshape(B, D) * used
    dup = xt.unsqueeze(0).expand(n_particles, B, D).reshape(n_particles * B, *samples.shape[1:]).detach().requires_grad_(True)
    dup_labels = labels.unsqueeze(0).expand(n_particles, B).reshape(-1)
    vectors = u.reshape(dup.shape)
    grad1 = scorenet(dup, dup_labels)
    grad2 = torch.autograd.grad(torch.sum(grad1 * vectors), dup, create_graph=True)[0]
    loss1 = torch.sum(grad1.reshape(n_particles * B, -1) ** 2, dim=-1) / 2.0
    loss2 = torch.sum((vectors * grad2).reshape(n_particles * B, -1), dim=-1)
    per = (loss1.view(n_particles, B).mean(0) + loss2.view(n_particles, B).mean(0)) * used.view(B) ** anneal_power
    if hook is not None:
        hook(per, labels)
    return per.mean(dim=0)


def exact_score_matching(energy_net, samples, train=False):
    """SSM/losses/score_matching.py:48-73: per-sample 0.5 ||grad logp||^2 + tr(Hessian logp) by one backward pass per
    input dimension.  Evaluation metric of the DKEF runner (dkef_runner.py:373-389); no random draws and no kernel of
    ours: D passes through the caller's torch network.  Returns the per-sample vector (detached unless train)."""
    samples.requires_grad_(True)
    logp = -energy_net(samples).sum()
    grad1 = torch.autograd.grad(logp, samples, create_graph=True)[0]
    loss = torch.norm(grad1, dim=-1) ** 2 / 2.0
    for i in range(samples.shape[1]):
        g = torch.autograd.grad(grad1[:, i].sum(), samples, create_graph=train, retain_graph=True)[0][:, i]
        loss = loss + (g if train else g.detach())
    return loss if train else loss.detach()


def _mdsm_exact(energy, x_real, sigmas, sigma02, dim, mean, projected, batch_size, loss_coe, u, dsm_noise):
    coe = FLAGS.loss_coe if loss_coe is None else loss_coe
    bs = batch_size or FLAGS.batch_size or x_real.shape[0]
    x2 = x_real.reshape(bs, dim)
    if dsm_noise is None:
        seed, offset, row0 = ops.next_draw()
        dsm_noise = ops.philox_normal(bs, dim, 1, seed, offset, row0, device=x2.device)
        if projected and u is None:       # same draw, the other Philox stream
            u = ops.philox_normal(bs, dim, 0, seed, offset, row0, device=x2.device)
    x_noisy = (x2 + sigmas * dsm_noise.reshape(bs, dim)).detach().requires_grad_()
    grad_x = torch.autograd.grad(energy(x_noisy).sum(), x_noisy, create_graph=True)[0]
    r = (x2 - x_noisy) / sigma02 + grad_x
    if projected:
        if u is None:
            u = _projection_vectors(x2, "gaussian", None)
        per = torch.sum(r * u.reshape(bs, dim).detach(), dim=-1) ** 2
    else:
        per = torch.sum(r ** 2, dim=-1)
    per = per * (coe / (sigmas.view(-1) ** 2))
    return per.mean() if mean is True else per


def mdsm_baseline(energy, x_real, sigmas, sigma02, dim, mean=True, *, batch_size=None, loss_coe=None, dsm_noise=None):
    """DeepEBM/ESM/loss_train.py:29-44: the multiscale DSM objective mdsm_fd approximates, and train_dsm.py's evaluation
    loss (inputs.py:126-134).  One autograd pass over the input through the caller's network; the noise comes from
    kernel (1)'s Philox noise stream."""
    return _mdsm_exact(energy, x_real, sigmas, sigma02, dim, mean, False, batch_size, loss_coe, None, dsm_noise)


def mdsm_tracetrick(energy, x_real, sigmas, sigma02, dim, mean=True, *, batch_size=None, loss_coe=None, u=None, dsm_noise=None):
    """DeepEBM/ESM/loss_train.py:47-63: mdsm_baseline with the residual projected on one Gaussian direction before
    squaring.  RNG order: noise, then projection."""
    return _mdsm_exact(energy, x_real, sigmas, sigma02, dim, mean, True, batch_size, loss_coe, u, dsm_noise)


def deepebm_exact_score_matching(energy_net, samples, dim):
    """DeepEBM/ESM/loss_train.py:11-26 (test_exact_sm.py's evaluator): mean over the batch of exact_score_matching on
    samples.view(-1, dim); a detached scalar."""
    return exact_score_matching(energy_net, samples.reshape(-1, dim), train=False).mean()


# ------------------------------------------------------------------------------------------------
# N4: GAN_FD's one-sided finite-difference gradient penalty (GAN_FD/library/trainer_fd.py:57-73 real data, :82-100 fake)
# ------------------------------------------------------------------------------------------------
def fd_gradient_penalty(discriminator, x, y, fd_eps, reg_param=1.0, *, u=None):
    """d = D(cat[x, x - noise], cat[y, y]) with noise = fd_eps * sqrt(D) * u/||u|| (one launch of kernels (1)+(2), rows
    X | MINUS); returns (d_real = d[:B], reg = reg_param * mean(((d[:B] - d[B:]) / fd_eps)**2)) as the reference computes
    them; the caller adds its own GAN loss on d_real and calls backward()."""
    x2, B, D = _flat(x)
    stacked, _, _ = ops.stack(x2, row_mask=ROW_X | ROW_MINUS, scale=float(fd_eps) * math.sqrt(D),
                              direction=None if u is None else u.reshape(B, -1), normalize=True)
    cat_y = torch.cat([y, y], 0) if y is not None else None
    d = discriminator(stacked.view(2 * B, *x.shape[1:]), cat_y) if cat_y is not None else discriminator(stacked.view(2 * B, *x.shape[1:]))
    d1, d2 = d[:B], d[B:]
    reg = reg_param * (((d1 - d2) / fd_eps) ** 2).mean()
    return d1, reg


# ------------------------------------------------------------------------------------------------
# unified entry point (BASELINE.json: "(model, samples, n_particles, eps, sigma)")
# ------------------------------------------------------------------------------------------------
def fd_loss(model, samples, n_particles=1, eps=0.1, sigma=None, *, mode="ssm", labels=None, sigmas=None, noise=None,
            u=None, dsm_noise=None, anneal_power=2.0, hook=None, sigma02=0.01, loss_coe=1e-4):
    """mode: 'ssm' (a1), 'ssm_deepebm' (a7), 'dsm' (a3, needs sigma), 'mdsm' (a9, sigma = per-sample
    schedule), 'ssmvr_score' (a4), 'anneal_ssmvr' (a13, needs sigmas)."""
    _check_particles(n_particles, f"fd_loss[{mode}]")
    if mode == "ssm":
        return efficient_score_matching_conjugate(model, samples.reshape(samples.shape[0], -1), noise=noise, eps=eps, u=u)
    if mode == "ssm_deepebm":
        return ssm_fd(model, samples, eps=eps, u=u)
    if mode == "dsm":
        return dsm_fd(model, samples, sigma=1 if sigma is None else sigma, eps=eps, u=u, dsm_noise=dsm_noise)
    if mode == "mdsm":
        return mdsm_fd(model, samples, sigma, sigma02, int(math.prod(samples.shape[1:])), eps=eps, loss_coe=loss_coe,
                       u=u, dsm_noise=dsm_noise)
    if mode == "ssmvr_score":
        return esm_scorenet_VR(model, samples, eps=eps, u=u)
    if mode == "anneal_ssmvr":
        return anneal_ESM_scorenet_VR(model, samples, sigmas, labels, anneal_power, hook, 1, eps, u=u, dsm_noise=dsm_noise)
    raise ValueError(f"unknown mode {mode!r}")
