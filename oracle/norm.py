"""CPU restatement (numpy, float64 unless told otherwise) of InstanceNorm++ followed by an optional ELU, forward and backward.
TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.

Follows InstanceNorm2dPlus.forward (ncsnv2/models/normalization.py:163-176) and ConditionalInstanceNorm2dPlus.forward
(ncsn/models/cond_refinenet_dilated.py:318-333; gamma / alpha / beta rows per sample); F.elu as applied after `normalize`
in ncsnv2/models/layers.py:443-448.  Pinned to outputs AND autograd gradients of the unmodified reference modules:
tests/golden/inorm_plus.npz (oracle/make_golden_norm.py), checked by tests/test_norm_cpu.py.
"""
import numpy as np

EPS = 1e-5


def _rows(p, N):
    p = np.asarray(p)
    return np.broadcast_to(p, (N, p.shape[-1])) if p.ndim == 2 else np.broadcast_to(p[None], (N, p.shape[0]))


def forward(x, gamma, alpha, beta=None, act=False, dtype=np.float64):
    """x (N, C, H, W); gamma / alpha / beta (C,), (1, C) or (N, C).  Returns (out, cache)."""
    x = np.asarray(x, dtype=dtype)
    N, C = x.shape[:2]
    g, a = _rows(gamma, N).astype(dtype), _rows(alpha, N).astype(dtype)
    b = _rows(beta, N).astype(dtype) if beta is not None else np.zeros((N, C), dtype=dtype)
    mu = x.mean(axis=(2, 3))                                               # normalization.py:164
    m = mu.mean(axis=1, keepdims=True)                                     # :165
    v = mu.var(axis=1, keepdims=True, ddof=1)                              # :166 (torch.var: unbiased)
    s = 1.0 / np.sqrt(v + EPS)
    mustd = (mu - m) * s                                                   # :167
    var = x.var(axis=(2, 3))                                               # nn.InstanceNorm2d: biased, eps 1e-5 (:168)
    rstd = 1.0 / np.sqrt(var + EPS)
    xhat = (x - mu[:, :, None, None]) * rstd[:, :, None, None]
    h = xhat + (mustd * a)[:, :, None, None]                               # :171
    y = g[:, :, None, None] * h + b[:, :, None, None]                      # :172
    out = np.where(y > 0, y, np.expm1(np.minimum(y, 0))) if act else y
    return out, dict(xhat=xhat, rstd=rstd, mustd=mustd, s=s, g=g, a=a, y=y, act=act)


def backward(gout, cache):
    """Analytic gradients of sum(out * gout): dx (N,C,H,W) and PER-SAMPLE dgamma, dalpha, dbeta (N, C) (sum over N for
    a shared parameter row)."""
    xhat, rstd, mustd, s, g, a, y = (cache[k] for k in ("xhat", "rstd", "mustd", "s", "g", "a", "y"))
    gy = np.asarray(gout, dtype=xhat.dtype)
    if cache["act"]:
        gy = gy * np.where(y > 0, 1.0, np.exp(np.minimum(y, 0)))
    N, C, H, W = xhat.shape
    HW = H * W
    S1 = gy.sum(axis=(2, 3))
    S2 = (gy * xhat).sum(axis=(2, 3))
    dbeta = S1
    dgamma = S2 + mustd * a * S1
    dalpha = g * mustd * S1
    d = g * a * S1                                                           # dL/dmustd
    # mustd_k = (mu_k - m) s,  s = (v + eps)^-1/2,  v = sum_k (mu_k - m)^2 / (C - 1)
    dmu = s * (d - d.mean(axis=1, keepdims=True) - mustd * (d * mustd).sum(axis=1, keepdims=True) / (C - 1))
    k1 = (g * rstd)[:, :, None, None]
    dx = k1 * (gy - (S1 / HW)[:, :, None, None] - xhat * (S2 / HW)[:, :, None, None]) + (dmu / HW)[:, :, None, None]
    return dx, dgamma, dalpha, dbeta
