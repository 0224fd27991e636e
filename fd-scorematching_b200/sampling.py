"""Callers either side of the hot path, with the reference's signatures:
annealed Langevin dynamics (ncsnv2/models/__init__.py:19-56) and DeepEBM's energy-model samplers
(DeepEBM/ESM/sampling_mdsm.py:6-96) on the fused update kernel, and the EMA helper (ncsnv2/models/ema.py:4-47) on the
fused EMA kernel.  The score / energy network is the caller's torch module (generic mode)."""
import numpy as np
import torch

from . import ops


@torch.no_grad()
def anneal_Langevin_dynamics(x_mod, scorenet, sigmas, n_steps_each=200, step_lr=0.000008, final_only=False, verbose=False,
                             *, noises=None):
    """ncsnv2/models/__init__.py:19-56.  Per step: score = scorenet(x, labels); x <- x + a*score + sqrt(2a)*z with
    a = step_lr * (sigma / sigmas[-1])**2 -- the update is ONE kernel (z from the in-kernel Philox noise stream, or taken
    from the iterator `noises` for parity runs).  Returns the reference's list of CPU tensors."""
    images = []
    sig = sigmas.detach().cpu().numpy() if isinstance(sigmas, torch.Tensor) else np.asarray(sigmas)
    x_mod = x_mod.clone()
    it = iter(noises) if noises is not None else None
    for c, sigma in enumerate(sig):
        labels = (torch.ones(x_mod.shape[0], device=x_mod.device) * c).long()
        step_size = step_lr * (sigma / sig[-1]) ** 2        # numpy scalar arithmetic, as in the reference (:29)
        for _ in range(n_steps_each):
            grad = scorenet(x_mod, labels)
            z = next(it) if it is not None else None
            if verbose:   # the reference's monitors need the noise tensor: draw it explicitly on this path
                z = torch.randn_like(x_mod) if z is None else z
                grad_norm = torch.norm(grad.view(grad.shape[0], -1), dim=-1).mean()
                noise_norm = torch.norm(z.view(z.shape[0], -1), dim=-1).mean()
            ops.langevin_step(x_mod, grad, float(step_size), noise=z, noise_scale=float(np.sqrt(step_size * 2)), out=x_mod)
            if not final_only:
                images.append(x_mod.to("cpu"))
            if verbose:
                image_norm = torch.norm(x_mod.view(x_mod.shape[0], -1), dim=-1).mean()
                snr = np.sqrt(step_size / 2.) * grad_norm / noise_norm
                grad_mean_norm = torch.norm(grad.mean(dim=0).view(-1)) ** 2 * sigma ** 2
                print("level: {}, step_size: {}, grad_norm: {}, image_norm: {}, snr: {}, grad_mean_norm: {}".format(
                    c, step_size, grad_norm.item(), image_norm.item(), snr.item(), grad_mean_norm.item()))
    return [x_mod.to("cpu")] if final_only else images


# ---- DeepEBM's samplers for energy models (DeepEBM/ESM/sampling_mdsm.py) -------------------------------------------------------
# Every step is  x <- x - sigma^2/2 * dE/dx + c * sigma * z  (c = 1, or sqrt(T_i) for the annealed sampler): the gradient is the
# caller's network through torch autograd, the update is ONE kernel (fdsm_langevin_step with step_size = -sigma^2/2 and
# noise_scale = c * sigma: the reference's rounding order), z from the in-kernel Philox noise stream or from `noises`.
def _energy_grad(netE, x):
    x = x.detach().requires_grad_()
    E_values = netE(x)
    grad = torch.autograd.grad(E_values.sum(), x)[0]
    return E_values, grad


def _mdsm_update(x, grad, sigma2, noise_scale, z):
    if noise_scale is None:       # Langevin_E(noise=False): the drift alone
        return x.detach() - 0.5 * sigma2 * grad
    return ops.langevin_step(x.detach(), grad, -0.5 * sigma2, noise=z, noise_scale=float(noise_scale))


def Langevin(netE, init_x, K, sigma, *, noises=None):
    """DeepEBM/ESM/sampling_mdsm.py:6-29: K Langevin steps at temperature sigma; returns the final x."""
    sigma2 = sigma ** 2
    netE.eval()
    it = iter(noises) if noises is not None else None
    x = init_x
    for _ in range(K):
        _, grad = _energy_grad(netE, x)
        x = _mdsm_update(x, grad, sigma2, sigma, next(it) if it is not None else None)
    netE.train()
    return x


def Langevin_E(netE, init_x, K, sigma, noise=True, *, noises=None):
    """DeepEBM/ESM/sampling_mdsm.py:32-60: as Langevin, also returning the (K, B) matrix of energies along the way."""
    sigma2 = sigma ** 2
    netE.eval()
    it = iter(noises) if noises is not None else None
    x = init_x
    E_list = []
    for _ in range(K):
        E_values, grad = _energy_grad(netE, x)
        E_list.append(E_values.squeeze().detach().cpu().numpy())
        x = _mdsm_update(x, grad, sigma2, sigma if noise else None, next(it) if (it is not None and noise) else None)
    netE.train()
    return x.detach(), np.vstack(E_list)


def Annealed_Langevin_E(netE, init_x, sigma, T_vect, Sample_every, *, noises=None):
    """DeepEBM/ESM/sampling_mdsm.py:63-96: annealed Langevin sampling, noise scaled by sqrt(T_vect[i]); returns (list of CPU
    samples taken every `Sample_every` steps, (B, K) energy matrix)."""
    sigma2 = sigma ** 2
    T_vect = np.sqrt(T_vect)
    K = T_vect.shape[0]
    netE.eval()
    it = iter(noises) if noises is not None else None
    E_list, x_list = [], []
    x = init_x
    for i in range(K):
        E_values, grad = _energy_grad(netE, x)
        E_list.append(E_values.squeeze().detach())
        x = _mdsm_update(x, grad, sigma2, T_vect[i] * sigma, next(it) if it is not None else None)
        if (i + 1) % Sample_every == 0:
            x_list.append(x.detach().cpu())
    netE.train()
    return x_list, torch.stack(E_list, 1).cpu().numpy()


class EMAHelper(object):
    """ncsnv2/models/ema.py:4-47: shadow = (1 - mu) * param + mu * shadow per trainable parameter."""

    def __init__(self, mu=0.999):
        self.mu = mu
        self.shadow = {}

    @staticmethod
    def _inner(module):
        return module.module if isinstance(module, torch.nn.DataParallel) else module

    def register(self, module):
        for name, param in self._inner(module).named_parameters():
            if param.requires_grad:
                self.shadow[name] = param.data.clone().contiguous()

    def update(self, module):
        for name, param in self._inner(module).named_parameters():
            if param.requires_grad:
                ops.ema_update(self.shadow[name], param.data.contiguous(), self.mu)

    def ema(self, module):
        for name, param in self._inner(module).named_parameters():
            if param.requires_grad:
                param.data.copy_(self.shadow[name])

    def state_dict(self):
        return self.shadow

    def load_state_dict(self, state_dict):
        self.shadow = state_dict
