"""CPU restatement of the callers either side of the hot path.  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's cpu baseline may import this package."""
import numpy as np
import torch


@torch.no_grad()
def anneal_langevin(x, scorenet, sigmas, n_steps_each, step_lr, noises):
    """ncsnv2/models/__init__.py:19-56 with the normal draws taken from `noises` in the reference's order
    (one randn_like per step).  Returns the final x."""
    it = iter(noises)
    x = x.clone()
    for c, sigma in enumerate(sigmas):                               # numpy array: numpy scalar arithmetic as in the reference
        labels = (torch.ones(x.shape[0]) * c).long()
        step_size = step_lr * (sigma / sigmas[-1]) ** 2              # :29
        for _ in range(n_steps_each):
            grad = scorenet(x, labels)                               # :31
            x = x + step_size * grad + next(it) * np.sqrt(step_size * 2)   # :39
    return x


def ema_update(shadow, param, mu):
    """ncsnv2/models/ema.py:21."""
    return (1. - mu) * param + mu * shadow


def mdsm_annealed_langevin(netE, x, sigma, T_vect, noises):
    """DeepEBM/ESM/sampling_mdsm.py:63-96 (Annealed_Langevin_E) with the normal draws taken from `noises`; Langevin / Langevin_E
    (:6-60) are the case T_vect = 1.  Returns (final x, (B, K) energies)."""
    sigma2 = sigma ** 2
    T = np.sqrt(np.asarray(T_vect, dtype=np.float64))
    it = iter(noises)
    E_list = []
    for i in range(T.shape[0]):
        x = x.detach().requires_grad_()
        E_values = netE(x)                                            # :82
        E_list.append(E_values.squeeze().detach())
        grad = torch.autograd.grad(E_values.sum(), x)[0]              # :84-85
        x = x.detach() - 0.5 * sigma2 * grad + T[i] * sigma * next(it)      # :86
    return x.detach(), torch.stack(E_list, 1)
