"""CPU restatement of the networks the FD losses drive (TEST INFRASTRUCTURE).

Plain torch modules with the reference's parameter names so that reference
state_dicts load unchanged.  Citations relative to /root/reference.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

LN2 = math.log(2.0)


def activation(name):
    """softplus (nn.Softplus, beta=1, threshold=20), nsoftplus = softplus - ln2
    (DeepEBM/ESM/models.py:19-28), elu (default.yaml:12)."""
    if name == "softplus":
        return F.softplus
    if name == "nsoftplus":
        return lambda z: F.softplus(z) - LN2
    if name == "elu":
        return F.elu
    raise ValueError(name)


class QuadraticMLP(nn.Module):
    """MLP_Quadratic_MNIST (models.py:932-955: 784-200-200-200) and MLP_Quadratic
    (:129-160: D-5000x5): AF(fc_k(.)) then E = ln1(h)*ln2(h) + lq(h^2)."""

    def __init__(self, in_dim, hidden, n_hidden, act="nsoftplus"):
        super().__init__()
        dims = [in_dim] + [hidden] * n_hidden
        for i in range(n_hidden):
            setattr(self, f"fc{i + 1}", nn.Linear(dims[i], dims[i + 1]))
        self.n_hidden = n_hidden
        self.ln1 = nn.Linear(hidden, 1)
        self.ln2 = nn.Linear(hidden, 1)
        self.lq = nn.Linear(hidden, 1)
        self.act = act

    def forward(self, x):
        af = activation(self.act)
        h = x.reshape(x.shape[0], -1)
        for i in range(self.n_hidden):
            h = af(getattr(self, f"fc{i + 1}")(h))
        return (self.ln1(h) * self.ln2(h) + self.lq(h * h)).view(-1)


class SoftplusMLP(nn.Module):
    """Config-1 energy: Linear-Softplus-Linear-Softplus-Linear(->1), the Phi_w /
    toy-runner shape (SSM/models/dkef.py:442-461, ncsn/runners/toy_runner.py:197-203)
    with out_features=1."""

    def __init__(self, in_dim, hidden, act="softplus"):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(in_dim, hidden), nn.Softplus(),
                                 nn.Linear(hidden, hidden), nn.Softplus(),
                                 nn.Linear(hidden, 1))
        assert act == "softplus"

    def forward(self, x):
        return self.net(x.reshape(x.shape[0], -1)).view(-1)


class TinyCondScore(nn.Module):
    """Small label-conditioned score net for fixtures: s(x, y) = W2 elu(W1 x + emb[y]),
    output divided by sigma[y] when ``sigmas`` is given (the NCSNv2 convention,
    ncsnv2/models/ncsnv2.py:97-99)."""

    def __init__(self, shape, hidden, n_classes, sigmas=None):
        super().__init__()
        self.shape = tuple(shape)
        D = 1
        for s in shape:
            D *= s
        self.l1 = nn.Linear(D, hidden)
        self.emb = nn.Embedding(n_classes, hidden)
        self.l2 = nn.Linear(hidden, D)
        self.sigmas = sigmas

    def forward(self, x, y=None):
        h = self.l1(x.reshape(x.shape[0], -1))
        if y is not None:
            h = h + self.emb(y)
        out = self.l2(F.elu(h)).reshape(x.shape[0], *self.shape)
        if self.sigmas is not None and y is not None:
            out = out / self.sigmas.to(out.dtype)[y].view(-1, *([1] * len(self.shape)))
        return out


def gmm_sample(n, dim, gen):
    """GMMDist(dim).sample (ncsn/models/gmm.py:58-73): means +-5, weights .8/.2, sigma 1."""
    idx = torch.multinomial(torch.tensor([0.8, 0.2]), n, replacement=True, generator=gen)
    means = torch.stack([5 * torch.ones(dim), -5 * torch.ones(dim)], 0)[idx]
    return torch.randn(n, dim, generator=gen) + means


def init_from_numpy(module, seed):
    """Deterministic, platform-independent init: every parameter drawn from numpy's
    legacy MT19937 RandomState as U(-1/sqrt(fan_in), 1/sqrt(fan_in)) (the nn.Linear
    default range), in ``named_parameters`` order.  Fixtures store only the seed."""
    import numpy as np
    rs = np.random.RandomState(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            fan_in = p.shape[1] if p.dim() > 1 else p.shape[0]
            if name.endswith("bias"):
                # bias bound uses the fan_in of its weight; approximate with own size for
                # embeddings / heads -- exact value is irrelevant, determinism is the point
                w = dict(module.named_parameters()).get(name[:-4] + "weight")
                if w is not None and w.dim() > 1:
                    fan_in = w.shape[1]
            bound = 1.0 / math.sqrt(max(fan_in, 1))
            vals = rs.uniform(-bound, bound, size=tuple(p.shape))
            p.copy_(torch.from_numpy(vals).to(p.dtype))
    return module


def np_randn(seed, *shape):
    import numpy as np
    return torch.from_numpy(np.random.RandomState(seed).standard_normal(shape))


def np_rand(seed, *shape):
    import numpy as np
    return torch.from_numpy(np.random.RandomState(seed).uniform(0.0, 1.0, shape))


def np_randint(seed, high, n):
    import numpy as np
    return torch.from_numpy(np.random.RandomState(seed).randint(0, high, size=(n,))).long()


def init_by_name(module, seed):
    """Deterministic init that does not depend on parameter ORDER: every floating-point parameter is drawn from a numpy
    MT19937 stream keyed by (seed, crc32(name)).  Used where the same state must be given to the reference's module and
    to fdsm_b200.nets (whose registration order differs).  Ranges: weights with >= 2 dims U(+-1/sqrt(fan_in)) (torch's
    conv / linear default bound); normaliser scales (alpha, gamma, LayerNorm weight, the first two thirds of an
    InstanceNorm++ embedding row) U(0.9, 1.1); every other vector U(+-0.05)."""
    import zlib
    import numpy as np
    with torch.no_grad():
        for name, p in sorted(module.named_parameters()):
            rs = np.random.RandomState((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 31))
            leaf = name.rsplit(".", 1)[-1]
            if name.endswith("embed.weight"):
                c = p.shape[1] // 3
                vals = np.concatenate([rs.uniform(0.9, 1.1, (p.shape[0], 2 * c)), rs.uniform(-0.05, 0.05, (p.shape[0], c))], 1)
            elif leaf in ("alpha", "gamma") or (leaf == "weight" and p.dim() == 3):   # LayerNorm([C, H, W]).weight
                vals = rs.uniform(0.9, 1.1, tuple(p.shape))
            elif p.dim() >= 2:
                bound = 1.0 / math.sqrt(max(int(np.prod(p.shape[1:])), 1))
                vals = rs.uniform(-bound, bound, tuple(p.shape))
            else:
                vals = rs.uniform(-0.05, 0.05, tuple(p.shape))
            p.copy_(torch.from_numpy(vals).to(p.dtype))
    return module
