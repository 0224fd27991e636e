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
