"""fp32 emulation (numpy) of the closed forms of csrc/fd_jet.cuh against fp64 evaluation of the definitions: how accurate
are a', b' (forward) and S1, S2 (backward) in the regime the main path covers (max(|P|, |Q|) <= 0.25)?
MUFU ex2 / lg2 / rcp are modelled as correctly rounded fp32 functions (their documented error is <= 2 ulp).
    python tests/tools/jet_math_check.py"""
import numpy as np

f = np.float32
rng = np.random.default_rng(0)


def expm1_s(x):
    p = f(1.984127e-4)
    for c in (1.3888889e-3, 8.3333333e-3, 4.1666667e-2, 1.6666667e-1, 0.5):
        p = f(p * x + f(c))
    return f(f(p * x) * x + x)


def rcp(x):
    return (f(1) / x).astype(f)


def log1p_ratio(n, d):
    s = f(n * rcp(f(f(2) * d + n)))
    u = f(s * s)
    q = f(0.076923077)
    for c in (0.090909091, 0.11111111, 0.14285715, 0.2, 0.33333334):
        q = f(q * u + f(c))
    s2 = f(s + s)
    return f(f(s2 * u) * q + s2)


def fwd32(z0, P, Q):
    t = np.exp(-np.abs(z0)).astype(f)
    r = rcp(f(1) + t)
    s0 = np.where(z0 >= 0, r, f(t * r)).astype(f)
    e0 = np.where(z0 >= 0, f(t * r), r).astype(f)
    h0 = f(np.maximum(z0, f(0)) + np.log1p(t.astype(np.float64)).astype(f))
    emP = expm1_s(P)
    inv = rcp(f(1) + emP)
    emn = f(-emP * inv)
    sinh = f(f(0.5) * f(emP - emn))
    coshm1 = f(f(f(0.5) * emP) * f(emP * inv))
    emq = expm1_s(Q)
    eq = f(f(1) + emq)
    em = f(f(emn + emq) + f(emn * emq))
    a = f(f(0.5) * log1p_ratio(f(f(f(2) * s0) * f(eq * sinh)), f(f(1) + f(s0 * em))))
    S = f(f(2) * f(emq * f(f(1) + coshm1) + coshm1))
    tb = f(s0 * f(e0 * S + f(f(s0 * emq) * f(f(2) + emq))))
    b = f(f(0.5) * log1p_ratio(tb, f(1)))
    return h0, a, b


def sp64(z):
    return np.logaddexp(0.0, z)


def bwd32(h0, a, b):
    e0 = np.exp(-h0).astype(f)
    s0 = f(f(1) - e0)
    ema = expm1_s(a)
    inv = rcp(f(1) + ema)
    emn = f(-ema * inv)
    sinh = f(f(0.5) * f(ema - emn))
    coshm1 = f(f(f(0.5) * ema) * f(ema * inv))
    emb = expm1_s(f(-b))
    S1 = f(f(e0 * f(f(1) + emb)) * sinh)
    S2 = f(-e0 * f(emb * f(f(1) + coshm1) + coshm1))
    return s0, S1, S2


def report(name, got, want, floor=0.0):
    err = np.abs(got.astype(np.float64) - want) / np.maximum(np.abs(want), floor + 1e-300)
    print(f"  {name:4s} rel err  median {np.median(err):.2e}  p99.9 {np.quantile(err, 0.999):.2e}  max {err.max():.2e}")


for label, ps, qs in (("typical (|P| ~ 3e-2, |Q| ~ 1e-3)", 3e-2, 1e-3), ("edge (|P|, |Q| up to 0.25)", None, None)):
    n = 400000
    z0 = (rng.standard_normal(n) * 3).astype(f)
    if ps is None:
        P = rng.uniform(-0.25, 0.25, n).astype(f)
        Q = rng.uniform(-0.25, 0.25, n).astype(f)
    else:
        P = (rng.standard_normal(n) * ps).astype(f)
        Q = (rng.standard_normal(n) * qs).astype(f)
    z, p, q = z0.astype(np.float64), P.astype(np.float64), Q.astype(np.float64)
    # fp64 definitions, cancellation-free enough at these magnitudes when done in extended steps
    sp0, spp, spm = sp64(z), sp64(z + q + p), sp64(z + q - p)
    a64 = 0.5 * (spp - spm)
    b64 = 0.5 * ((spp - sp0) + (spm - sp0))
    h0, a, b = fwd32(z0, P, Q)
    print(label)
    report("h0", h0, sp0)
    report("a", a, a64, 1e-12)
    report("b", b, b64, 1e-9)
    sig = lambda t: 1.0 / (1.0 + np.exp(-t))
    S1_64 = 0.5 * (sig(z + q + p) - sig(z + q - p))
    S2_64 = 0.5 * (sig(z + q + p) + sig(z + q - p)) - sig(z)
    s0, S1, S2 = bwd32(sp0.astype(f), a64.astype(f), b64.astype(f))
    report("s0", s0, sig(z), 1e-3)
    report("S1", S1, S1_64, 1e-12)
    report("S2", S2, S2_64, 1e-9)
