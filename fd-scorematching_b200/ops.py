"""Tensor-level wrappers of the libfdsm kernels and the autograd Functions around kernel (4).

PyTorch is plumbing here: it owns device memory and streams; every computation below is a call
through the C ABI (``_lib``).  No fallback path exists.
"""
import ctypes as C
import math
import threading

import torch

from . import _lib
from ._lib import ROW_MINUS, ROW_PLUS, ROW_X, StackArgs, check


LAUNCHES = 0          # kernels launched through the C ABI by this process (bench.py "gpu_launches")
PROFILE_HOOK = None   # bench.py: list collecting (start, end) CUDA events around the dominant kernel
PROFILE_HOOK_STACK = None   # same, around the fused direction + stack-build kernel
CHAIN_TRACE = None          # scripts/chain_trace.py: int64 CUDA tensor receiving the chain kernel's per-CTA phase stamps


def _count(n):
    global LAUNCHES
    LAUNCHES += n


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _req(t, name, dtype=torch.float32):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise _lib.FdsmError(f"{name} must be a CUDA tensor (no CPU fallback); got "
                             f"{type(t).__name__}{'' if not isinstance(t, torch.Tensor) else ' on ' + str(t.device)}")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype} (the hot path computes in fp32)")
    return t


def _rows(t, name):
    """(B, D) row-major view with unit inner stride; returns (tensor, ld)."""
    if t.dim() != 2:
        raise ValueError(f"{name} must be 2-D, got shape {tuple(t.shape)}")
    if t.shape[0] > 1 and t.shape[1] > 0 and (t.stride(1) != 1 or t.stride(0) < t.shape[1]):
        t = t.contiguous()
    if t.shape[1] > 0 and t.stride(1) != 1:
        t = t.contiguous()
    ld = t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1)
    return t, ld


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


# --------------------------------------------------------------------------------------------
# Philox state: (seed, offset).  Every production-mode draw consumes one offset.
# --------------------------------------------------------------------------------------------
class _Rng(threading.local):
    seed = 0x5EED_F05D
    offset = 0
    row0 = 0


_rng = _Rng()


def manual_seed(seed, offset=0):
    """Seed the Philox streams used when no noise is injected."""
    _rng.seed, _rng.offset = int(seed) & (2 ** 64 - 1), int(offset)


def set_row_offset(row0):
    """Global index of this rank's first sample (rank*B_local), so a sharded batch draws exactly
    what the un-sharded batch would."""
    _rng.row0 = int(row0)


def next_draw():
    o = _rng.offset
    _rng.offset += 1
    return _rng.seed, o, _rng.row0


# --------------------------------------------------------------------------------------------
# kernels (1)+(2)
# --------------------------------------------------------------------------------------------
def stack(x, *, row_mask, scale, direction=None, normalize=True, noise=None, sigma=None, labels=None,
          perturb=True, dot_mode=0, want_v=False, want_dot=False, draw=None, out=None, out_lo=None, offset_dev=None,
          pad_ones=False, dot_out=None):
    """Fused direction sampling + normalisation + stacked-batch build (fdsm_stack).

    x: (B, D) fp32 CUDA.  Returns (stacked (rows*B, D) | None, v (B, D) | None, dot (B,) | None)."""
    lib = _lib.load()
    x = _req(x, "samples")
    x, ldx = _rows(x, "samples")
    B, D = x.shape
    a = StackArgs()
    a.x, a.ldx = _ptr(x), ldx
    keep = [x]
    if direction is not None:
        direction, ldd = _rows(_req(direction, "direction"), "direction")
        if direction.shape != x.shape:
            raise ValueError(f"direction shape {tuple(direction.shape)} != samples {tuple(x.shape)}")
        a.dir, a.lddir = _ptr(direction), ldd
        keep.append(direction)
    a.normalize_dir, a.dir_scale = int(bool(normalize)), float(scale)
    if sigma is not None:
        sigma = _req(sigma, "sigma").contiguous().view(-1)
        a.sigma = _ptr(sigma)
        keep.append(sigma)
        if labels is not None:
            labels = _req(labels, "labels", torch.int64).contiguous()
            if labels.numel() != B:
                raise ValueError("labels must have one entry per sample")
            a.labels = _ptr(labels)
            keep.append(labels)
        elif sigma.numel() != B:
            raise ValueError(f"per-sample sigma must have B={B} entries, got {sigma.numel()}")
        if noise is not None:
            noise, ldn = _rows(_req(noise, "noise"), "noise")
            if noise.shape != x.shape:
                raise ValueError("noise shape != samples shape")
            a.noise, a.ldnoise = _ptr(noise), ldn
            keep.append(noise)
    a.perturb_input, a.dot_mode = int(bool(perturb)), int(dot_mode)
    rows = bin(row_mask & 7).count("1")
    if rows:
        if out is None:
            out = torch.empty((rows * B, D), device=x.device, dtype=torch.float32)
        a.out, a.ldout = _ptr(out), (out.stride(0) if out.shape[0] > 1 else max(out.shape[1], 1))
        if out_lo is not None:
            if out_lo.shape != out.shape or out_lo.stride() != out.stride():
                raise ValueError("out_lo must have the shape and strides of out")
            a.out_lo = _ptr(out_lo)
        a.pad_ones = int(bool(pad_ones))
    a.row_mask = row_mask
    v = dot = None
    if want_v:
        v = torch.empty((B, D), device=x.device, dtype=torch.float32)
        a.v_out, a.ldv = _ptr(v), max(D, 1)
    if dot_out is not None:     # caller-owned (persistent) buffer: CUDA-graph replays write the same address
        dot = _req(dot_out, "dot_out")
        if dot.numel() != B or not dot.is_contiguous():
            raise ValueError("dot_out must be a contiguous (B,) tensor")
        a.dot_out = _ptr(dot)
    elif want_dot:
        dot = torch.empty((B,), device=x.device, dtype=torch.float32)
        a.dot_out = _ptr(dot)
    seed, offset, row0 = draw if draw is not None else next_draw()
    a.seed, a.offset, a.row0 = seed, offset, row0
    if offset_dev is not None:
        a.offset_dev = _ptr(_req(offset_dev, "offset_dev", torch.int64))
    a.B, a.D = B, D
    _count(1); check(lib.fdsm_stack(C.byref(a), _stream()), "fdsm_stack")
    return out, v, dot


def direction(B, D, scale, *, u=None, device=None, draw=None):
    """Kernel (1) alone: v = scale*u/||u|| with u injected or Philox."""
    lib = _lib.load()
    if u is not None:
        u, ldu = _rows(_req(u, "u"), "u")
        device = u.device
    v = torch.empty((B, D), device=device, dtype=torch.float32)
    seed, offset, row0 = draw if draw is not None else next_draw()
    _count(1); check(lib.fdsm_direction(_ptr(u), ldu if u is not None else 0, _ptr(v), max(D, 1), B, D, float(scale), seed,
                             offset, row0, _stream()), "fdsm_direction")
    return v


def philox_normal(B, D, stream_id, seed, offset, row0=0, device="cuda"):
    lib = _lib.load()
    if torch.device(device).type != "cuda":
        raise _lib.FdsmError(f"philox_normal draws on the GPU (no CPU fallback); got device {device}")
    out = torch.empty((B, D), device=device, dtype=torch.float32)
    _count(1); check(lib.fdsm_philox_normal(_ptr(out), max(D, 1), B, D, stream_id, seed, offset, row0, _stream()),
          "fdsm_philox_normal")
    return out


# --------------------------------------------------------------------------------------------
# kernel (4): autograd Functions
# --------------------------------------------------------------------------------------------
class CombineEnergy(torch.autograd.Function):
    """loss = scale * reduce_b[(E2-E3)^2/8 - E2 - E3 + c1*E1]  (mode 0) / sum(d^2+4d) (mode 1)."""

    @staticmethod
    def forward(ctx, E, B, mode, c1, scale, mean, want_per_sample):
        lib = _lib.load()
        ctx.e_shape = tuple(E.shape)      # (N,) or (N, 1): the backward returns the gradient in the caller's shape
        E = _req(E, "energy output").contiguous().view(-1)
        rows = 3 if mode == 0 else 2
        if E.numel() != rows * B:
            raise ValueError(f"energy net returned {E.numel()} values for a stacked batch of {rows * B} rows")
        loss = torch.empty((), device=E.device, dtype=torch.float32)
        per = torch.empty((B,), device=E.device, dtype=torch.float32) if want_per_sample else None
        _count(1); check(lib.fdsm_combine_energy_fwd(_ptr(E), B, mode, c1, scale, int(mean), _ptr(per), _ptr(loss), _stream()),
              "fdsm_combine_energy_fwd")
        ctx.save_for_backward(E)
        ctx.cfg = (B, mode, c1, scale, int(mean))
        if want_per_sample:
            ctx.mark_non_differentiable(per)
            return loss, per
        return loss, None

    @staticmethod
    def backward(ctx, gout, _gper):
        lib = _lib.load()
        (E,) = ctx.saved_tensors
        B, mode, c1, scale, mean = ctx.cfg
        gout = gout.contiguous().float()
        dE = torch.empty_like(E)
        _count(1); check(lib.fdsm_combine_energy_bwd(_ptr(E), B, mode, c1, scale, mean, _ptr(gout), _ptr(dE), _stream()),
              "fdsm_combine_energy_bwd")
        return dE.view(ctx.e_shape), None, None, None, None, None, None


class CombineDsm(torch.autograd.Function):
    """loss = gamma*mean_b[(alpha*(E2-E1) + beta*dot_b)^2 / sigma_b^2]."""

    @staticmethod
    def forward(ctx, E, dot, sigma_b, B, alpha, beta, gamma):
        lib = _lib.load()
        ctx.e_shape = tuple(E.shape)
        E = _req(E, "energy output").contiguous().view(-1)
        if E.numel() != 2 * B:
            raise ValueError(f"energy net returned {E.numel()} values for a stacked batch of {2 * B} rows")
        loss = torch.empty((), device=E.device, dtype=torch.float32)
        _count(1); check(lib.fdsm_combine_dsm_fwd(_ptr(E), _ptr(dot), _ptr(sigma_b), B, alpha, beta, gamma, None, _ptr(loss),
                                       _stream()), "fdsm_combine_dsm_fwd")
        ctx.save_for_backward(E, dot, sigma_b)
        ctx.cfg = (B, alpha, beta, gamma)
        return loss

    @staticmethod
    def backward(ctx, gout):
        lib = _lib.load()
        E, dot, sigma_b = ctx.saved_tensors
        B, alpha, beta, gamma = ctx.cfg
        gout = gout.contiguous().float()
        dE = torch.empty_like(E)
        _count(1); check(lib.fdsm_combine_dsm_bwd(_ptr(E), _ptr(dot), _ptr(sigma_b), B, alpha, beta, gamma, _ptr(gout), _ptr(dE),
                                       _stream()), "fdsm_combine_dsm_bwd")
        return dE.view(ctx.e_shape), None, None, None, None, None, None


class CombineScore(torch.autograd.Function):
    """loss = mean_b w_b * sum_d[(s1+s2)^2/8 + (s1-s2) v k]  (or the projected non-VR form).

    When the score output needs a gradient, the forward launch is the kernel's fused fwd+bwd pass (R s1, s2, v; W ds1,
    ds2 = 20*D bytes per sample): ds is written for an upstream gradient of 1 and the backward
    only rescales it when the upstream gradient differs from 1 (fdsm_scale_unless_one: a scalar test per block)."""

    @staticmethod
    def forward(ctx, s, v, sigma, labels, power, B, D, k, projected, want_per_sample):
        lib = _lib.load()
        s2d = _req(s, "score output").reshape(2 * B, D) if s.numel() == 2 * B * D else None
        if s2d is None:
            raise ValueError(f"score net returned shape {tuple(s.shape)}; expected 2B*D = {2 * B * D} elements")
        s2d, lds = _rows(s2d, "score output")
        loss = torch.empty((), device=s.device, dtype=torch.float32)
        per = torch.empty((B,), device=s.device, dtype=torch.float32)
        ds = torch.empty((2 * B, D), device=s.device, dtype=torch.float32) if ctx.needs_input_grad[0] else None
        _count(2); check(lib.fdsm_combine_score(_ptr(s2d), lds, _ptr(v), max(D, 1), _ptr(sigma), _ptr(labels), power, B, D, k,
                                     int(projected), None, _ptr(per), _ptr(loss), _ptr(ds), max(D, 1), _stream()),
              "fdsm_combine_score(fwd+bwd)" if ds is not None else "fdsm_combine_score(fwd)")
        ctx.ds, ctx.shape = ds, tuple(s.shape)
        ctx.mark_non_differentiable(per)
        return loss, per

    @staticmethod
    def backward(ctx, gout, _gper):
        lib = _lib.load()
        ds, ctx.ds = ctx.ds, None      # consumed: a second backward through the same graph is an error, as in torch
        if ds is None:
            raise RuntimeError("CombineScore.backward: the gradient buffer was already consumed (retain_graph is not "
                               "supported on the fused combine) or the score output did not require grad")
        gout = gout.contiguous().float()
        _count(1); check(lib.fdsm_scale_unless_one(_ptr(ds), ds.numel(), _ptr(gout), _stream()), "fdsm_scale_unless_one")
        return ds.view(ctx.shape), None, None, None, None, None, None, None, None, None


def combine_score_fused(s2d, v, *, k, sigma=None, labels=None, power=2.0, projected=False):
    """Single-pass fwd+bwd (root loss, upstream grad 1): returns (loss, per_sample, ds)."""
    lib = _lib.load()
    s2d, lds = _rows(_req(s2d, "score output"), "score output")
    B, D = v.shape
    loss = torch.empty((), device=s2d.device, dtype=torch.float32)
    per = torch.empty((B,), device=s2d.device, dtype=torch.float32)
    ds = torch.empty((2 * B, D), device=s2d.device, dtype=torch.float32)
    _count(2); check(lib.fdsm_combine_score(_ptr(s2d), lds, _ptr(v), max(D, 1), _ptr(sigma), _ptr(labels), float(power), B, D,
                                 float(k), int(projected), None, _ptr(per), _ptr(loss), _ptr(ds), max(D, 1),
                                 _stream()), "fdsm_combine_score(fused)")
    return loss, per, ds


# --------------------------------------------------------------------------------------------
# kernel (3) building blocks: TF32 hi/lo split and the tcgen05 GEMM
# --------------------------------------------------------------------------------------------
def tf32_split(x, out_cols=None, ones_col=False, want_lo=True, raw=False):
    """(hi, lo) with x = hi + lo, both TF32-representable; rows padded to `out_cols` columns (multiple of 4),
    column x.shape[1] set to 1 when `ones_col` (bias column of an augmented operand).  raw=True: (padded fp32 copy, None) --
    the operand format of the in-kernel-split GEMMs."""
    lib = _lib.load()
    x, ldx = _rows(_req(x, "x"), "x")
    R, Cc = x.shape
    oc = out_cols if out_cols is not None else (Cc + int(ones_col) + 3) // 4 * 4
    hi = torch.empty((R, oc), device=x.device, dtype=torch.float32)
    lo = torch.empty((R, oc), device=x.device, dtype=torch.float32) if (want_lo and not raw) else None
    _count(1); check(lib.fdsm_tf32_split(_ptr(x), ldx, R, Cc, _ptr(hi), _ptr(lo), oc, oc, int(ones_col) | (2 if raw else 0), _stream()),
                     "fdsm_tf32_split")
    return hi, lo


def tc_gemm(a_hi, a_lo, b_hi, b_lo, M, N, K, *, a_major=0, b_major=0, passes=3, splits=1, tile_n=0):
    """D[M,N] = sum_k A[m,k] B[n,k] on the tensor cores (fdsm_tc_gemm).  Operands are the pre-split hi/lo
    halves; major 0: stored [rows, K]; major 1: stored [K, rows].  Returns (splits, M, N) partials."""
    lib = _lib.load()
    out = torch.empty((splits, M, N), device=a_hi.device, dtype=torch.float32)
    g = _lib.GemmArgs()
    g.a_hi, g.a_lo, g.lda, g.a_major = _ptr(a_hi), _ptr(a_lo), a_hi.stride(0), a_major
    g.b_hi, g.b_lo, g.ldb, g.b_major = _ptr(b_hi), _ptr(b_lo), b_hi.stride(0), b_major
    g.M, g.N, g.K = M, N, K
    g.passes, g.splits, g.tile_n = passes, splits, tile_n
    g.out, g.ldo, g.split_stride = _ptr(out), N, M * N
    _count(1); check(lib.fdsm_tc_gemm(C.byref(g), _stream()), "fdsm_tc_gemm")
    return out


# --------------------------------------------------------------------------------------------
# Langevin update and EMA (callers either side of the hot path)
# --------------------------------------------------------------------------------------------
def langevin_step(x, score, step_size, *, noise=None, noise_scale=None, draw=None, out=None):
    """x + step_size*score + noise*sqrt(2*step_size) in one pass (fdsm_langevin_step); `noise` injected (parity) or drawn
    in-kernel from the Philox noise stream.  `out` may be `x` (in place)."""
    lib = _lib.load()
    shape = x.shape
    if x.dim() < 2:
        raise ValueError("x must have a batch dimension and at least one data dimension")
    Dflat = 1
    for d in shape[1:]:
        Dflat *= d
    x2, ldx = _rows(_req(x, "x").reshape(shape[0], Dflat), "x")
    s2, lds = _rows(_req(score, "score").reshape(shape[0], Dflat), "score")
    B, D = x2.shape
    n2, ldn = (None, 0)
    if noise is not None:
        n2, ldn = _rows(_req(noise, "noise").reshape(shape[0], Dflat), "noise")
    o = torch.empty_like(x2) if out is None else _req(out, "out").reshape(shape[0], Dflat)
    if o.dim() != 2 or o.shape != x2.shape or (o.numel() and o.stride(1) != 1):
        raise ValueError("out must be a row-major tensor of the shape of x")
    ldo = o.stride(0) if B > 1 else max(D, 1)
    seed, offset, row0 = draw if draw is not None else (next_draw() if noise is None else (0, 0, 0))
    scale = math.sqrt(2.0 * float(step_size)) if noise_scale is None else float(noise_scale)
    _count(1)
    check(lib.fdsm_langevin_step(_ptr(x2), ldx, _ptr(s2), lds, _ptr(n2), ldn, _ptr(o), ldo, B, D, float(step_size), scale, seed,
                                 offset, row0, _stream()), "fdsm_langevin_step")
    return o.view(shape)


def ema_update(shadow, params, mu):
    """shadow <- (1 - mu) * params + mu * shadow, in place (fdsm_ema_update)."""
    lib = _lib.load()
    _req(shadow, "shadow"); _req(params, "params")
    if shadow.numel() != params.numel() or not shadow.is_contiguous() or not params.is_contiguous():
        raise ValueError("shadow and params must be contiguous tensors of the same size")
    _count(1)
    check(lib.fdsm_ema_update(_ptr(shadow), _ptr(params), shadow.numel(), 1.0 - float(mu), float(mu), _stream()), "fdsm_ema_update")
    return shadow


def publish_scalars(src, host_dst):
    """host_dst (pinned fp32 host tensor) <- src (device fp32), stored by a kernel on the current stream
    (fdsm_publish_scalars): the host's read of a step's loss without a copy-engine operation between two graph replays.
    The host reads host_dst after synchronising with the stream, as after a non-blocking copy_."""
    lib = _lib.load()
    _req(src, "src")
    if host_dst.is_cuda or host_dst.dtype != torch.float32 or not host_dst.is_pinned() or not host_dst.is_contiguous():
        raise ValueError("host_dst must be a contiguous pinned fp32 host tensor")
    if src.numel() != host_dst.numel() or not src.is_contiguous():
        raise ValueError("src must be contiguous and as long as host_dst")
    _count(1)
    check(lib.fdsm_publish_scalars(_ptr(src), C.c_void_p(host_dst.data_ptr()), src.numel(), _stream()), "fdsm_publish_scalars")
    return host_dst


# --------------------------------------------------------------------------------------------
# SURVEY.md §8f row N1 (first piece): InstanceNorm++ (+ ELU), forward and backward in two launches each
# --------------------------------------------------------------------------------------------
def _param_rows(t, N, C, name):
    """(1 | N, C) fp32 rows with unit inner stride -> (tensor, row stride in floats; 0 = one shared row)."""
    _req(t, name)
    if t.dim() != 2 or t.shape[1] != C or t.shape[0] not in (1, N):
        raise ValueError(f"{name} must have shape (1, {C}) or ({N}, {C}), got {tuple(t.shape)}")
    if t.stride(1) != 1 and C > 1:
        t = t.contiguous()
    return t, (0 if t.shape[0] == 1 else (t.stride(0) if C > 1 or t.shape[0] > 1 else 1))


def inorm_plus_reference(x, gamma, alpha, beta, act=False):
    """The torch op sequence the kernel replaces (ncsnv2/models/normalization.py:163-176 + F.elu).  Used for CPU tensors
    (the networks' CPU goldens) and, inside InormPlus.backward, when the caller differentiates the gradient itself
    (create_graph=True: the exact-SSM comparison losses); never for a first-order CUDA step."""
    mu = x.mean(dim=(2, 3))
    centred = mu - mu.mean(dim=1, keepdim=True)
    mu_std = centred / torch.sqrt(mu.var(dim=1, keepdim=True) + 1e-5)
    h = torch.nn.functional.instance_norm(x) + (mu_std * alpha)[:, :, None, None]
    h = gamma[:, :, None, None] * h
    if beta is not None:
        h = h + beta[:, :, None, None]
    return torch.nn.functional.elu(h) if act else h


class InormPlus(torch.autograd.Function):
    """out = act(InstanceNorm++(x; gamma, alpha, beta)) on a contiguous (N, C, H, W) CUDA tensor."""

    @staticmethod
    def forward(ctx, x, gamma, alpha, beta, act):
        lib = _lib.load()
        _req(x, "x")
        if x.dim() != 4:
            raise ValueError(f"x must be (N, C, H, W), got {tuple(x.shape)}")
        if not x.is_contiguous():
            raise ValueError("x must be contiguous (ops.inorm_plus makes it so)")
        N, Cc, H, W = x.shape
        g, ps = _param_rows(gamma, N, Cc, "gamma")
        a, pa = _param_rows(alpha, N, Cc, "alpha")
        b, pb = _param_rows(beta, N, Cc, "beta") if beta is not None else (None, ps)
        if not (ps == pa == pb):         # one stride serves the three rows: the embedding's chunks or three plain vectors
            g, a = g.expand(N, Cc).contiguous(), a.expand(N, Cc).contiguous()
            b = b.expand(N, Cc).contiguous() if b is not None else None
            ps = Cc
        out = torch.empty_like(x)
        stats = torch.empty(max(int(lib.fdsm_inorm_plus_stats_floats(N, Cc)), 1), device=x.device, dtype=torch.float32)
        _count(int(lib.fdsm_inorm_plus_launches(N, Cc, H * W, 0, int(bool(act)), 1)))
        check(lib.fdsm_inorm_plus_fwd(_ptr(x), N, Cc, H * W, _ptr(g), _ptr(a), _ptr(b), ps, int(bool(act)), _ptr(out), _ptr(stats),
                                      _stream()), "fdsm_inorm_plus_fwd")
        ctx.save_for_backward(x, out if act else None, gamma, alpha, beta, stats, g, a)
        ctx.act, ctx.ps, ctx.shared = bool(act), ps, (gamma.shape[0] == 1, alpha.shape[0] == 1, beta is not None and beta.shape[0] == 1)
        return out

    @staticmethod
    def backward(ctx, gout):
        x, out, gamma, alpha, beta, stats, g, a = ctx.saved_tensors
        if torch.is_grad_enabled():
            # gradient of the gradient requested (create_graph=True): differentiate the torch op sequence instead
            with torch.enable_grad():
                y = inorm_plus_reference(x, gamma, alpha, beta, ctx.act)
                ins = [t for t in (x, gamma, alpha, beta) if t is not None and t.requires_grad]
                grads = iter(torch.autograd.grad(y, ins, gout, create_graph=True))
            res = [next(grads) if (t is not None and t.requires_grad) else None for t in (x, gamma, alpha, beta)]
            return res[0], res[1], res[2], res[3], None
        lib = _lib.load()
        N, Cc, H, W = x.shape
        gout = _req(gout, "grad_out").contiguous()
        need_dx = ctx.needs_input_grad[0]
        dx = torch.empty_like(x) if need_dx else None
        dpar = torch.empty(4, N, Cc, device=x.device, dtype=torch.float32)     # dgamma, dalpha, dbeta, scratch
        _count(int(lib.fdsm_inorm_plus_launches(N, Cc, H * W, 1, int(ctx.act), int(need_dx))))
        check(lib.fdsm_inorm_plus_bwd(_ptr(x), _ptr(out), _ptr(gout), N, Cc, H * W, _ptr(g), _ptr(a), ctx.ps, int(ctx.act), _ptr(stats),
                                      _ptr(dpar[3]), _ptr(dx), _ptr(dpar[0]), _ptr(dpar[1]), _ptr(dpar[2]), _stream()),
              "fdsm_inorm_plus_bwd")

        if any(ctx.shared):       # shared rows: ONE reduction over the samples for the three parameters
            summed = dpar[:3].sum(1, keepdim=True)

        def rows(i, shared):
            return summed[i] if shared else dpar[i]
        return (dx, rows(0, ctx.shared[0]) if ctx.needs_input_grad[1] else None,
                rows(1, ctx.shared[1]) if ctx.needs_input_grad[2] else None,
                rows(2, ctx.shared[2]) if (beta is not None and ctx.needs_input_grad[3]) else None, None)


def inorm_plus(x, gamma, alpha, beta=None, act=False):
    """InstanceNorm++ (+ ELU when act) of x (N, C, H, W) with parameter rows gamma / alpha / beta of shape (1 | N, C)."""
    return InormPlus.apply(x.contiguous(), gamma, alpha, beta, act)
