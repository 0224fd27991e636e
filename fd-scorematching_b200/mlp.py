"""Fused energy MLPs: the reference's ``MLP_Quadratic_MNIST`` / ``MLP_Quadratic`` (DeepEBM/ESM/models.py:129-160,
932-955) and the config-1 3-layer Softplus MLP, forward and backward on the tcgen05 GEMMs of libfdsm
(``fdsm_mlp_*``, include/fdsm.h).  No torch matmul, no fallback.

Two ways in:
  * ``net(x)`` -- an ``nn.Module`` drop-in for the reference's energy networks: works with every energy-form FD loss
    adapter in ``losses`` (generic mode; the a17 comparison losses differentiate w.r.t. x and need a torch module); parameters live in ONE flat ``nn.Parameter`` laid out in the reference's
    ``named_parameters()`` order, so ``load_reference_state_dict`` accepts the reference's checkpoints.
  * ``net.fd_ssm_loss_and_grad(x, eps, scale)`` -- the fused FD-SSM step: stack build (hi/lo TF32 operands written
    directly by kernel (1)+(2)), MLP forward, combine forward+backward (kernel (4)), MLP backward.
"""
import ctypes as C
import math
import os

import numpy as np
import torch
import torch.nn as nn

from . import _lib, ops
from ._lib import JetArgs, MlpDesc, check

ACTS = {"softplus": 1, "nsoftplus": 2, "elu": 3}


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _Sizes:
    def __init__(self, desc, rows):
        n, w, ld = C.c_int64(), C.c_int64(), C.c_int64()
        check(_lib.load().fdsm_mlp_sizes(C.byref(desc), rows, C.byref(n), C.byref(w), C.byref(ld)), "fdsm_mlp_sizes")
        self.n_params, self.workspace, self.in_ld = n.value, w.value, ld.value


class _MlpFunction(torch.autograd.Function):
    """E = MLP(x) with the flat parameter vector as the only differentiable input."""

    @staticmethod
    def forward(ctx, flat, x2, net):
        lib = _lib.load()
        if x2.requires_grad:
            # the backward emits parameter gradients only (the FD objectives never differentiate w.r.t. the samples):
            # a caller that needs dE/dx (the exact-SSM comparison losses, a GAN generator) must not get silence
            raise NotImplementedError("FusedEnergyMLP does not provide the gradient with respect to its input; use a torch "
                                      "module for losses that differentiate through x (ssm, ssm_vr, exact_score_matching)")
        rows = x2.shape[0]
        sz = net.sizes(rows)
        xh, xl = ops.tf32_split(x2, out_cols=sz.in_ld, ones_col=True, raw=not net.presplit)
        ws = torch.empty(sz.workspace, device=x2.device, dtype=torch.float32)
        E = torch.empty(rows, device=x2.device, dtype=torch.float32)
        d = net.desc
        ops._count(1); check(lib.fdsm_mlp_prepare(C.byref(d), rows, _ptr(flat), _ptr(ws), _stream()), "fdsm_mlp_prepare")
        ops._count(net.n_hidden); check(lib.fdsm_mlp_forward(C.byref(d), _ptr(xh), _ptr(xl), rows, _ptr(flat), _ptr(ws), _ptr(E),
                                                           _stream()), "fdsm_mlp_forward")
        ctx.save_for_backward(flat, xh, xl if xl is not None else xh.new_empty(0), ws)
        ctx.net, ctx.rows = net, rows
        return E

    @staticmethod
    def backward(ctx, dE):
        lib = _lib.load()
        flat, xh, xl, ws = ctx.saved_tensors
        xl = xl if xl.numel() else None
        net, rows = ctx.net, ctx.rows
        dE = dE.contiguous().float()
        g = torch.empty_like(flat)
        ops._count(3 * net.n_hidden + 1)
        check(lib.fdsm_mlp_backward(C.byref(net.desc), _ptr(xh), _ptr(xl), rows, _ptr(flat), _ptr(ws), _ptr(dE), _ptr(g), 0,
                                    _stream()), "fdsm_mlp_backward")
        return g, None, None


class _ExactSsmFunction(torch.autograd.Function):
    """loss = mean_b[(u.grad E)^2/2 - u' Hess E u] through FusedEnergyMLP.exact_ssm_loss_and_grad; the flat parameter vector
    is the only differentiable input (the objective's derivatives w.r.t. x are carried forward inside the kernel)."""

    @staticmethod
    def forward(ctx, flat, x, u, net, draw, scale, noise_type):
        loss, grad = net.exact_ssm_loss_and_grad(x, u=u, draw=draw, scale=scale, noise_type=noise_type)
        ctx.save_for_backward(grad.clone())
        return loss.clone()

    @staticmethod
    def backward(ctx, gout):
        (grad,) = ctx.saved_tensors
        return grad * gout, None, None, None, None, None, None


class FusedEnergyMLP(nn.Module):
    def __init__(self, in_dim, hidden, n_hidden=3, act="nsoftplus", head="quadratic", device="cuda", seed=None, passes=3,
                 presplit=False, formulation="auto", jet_passes=2):
        """formulation of the fused FD-SSM step (`fd_ssm_loss_and_grad` / `fd_ssm_step`): "jet" = the difference-
        propagating chain kernel (csrc/fd_chain.cuh; hidden <= 223, n_hidden <= 5), "stacked" = the reference's stacked batch
        [x | x+v | x-v] through the 3xTF32 GEMM launches, "auto" = jet where supported.  `jet_passes`: 3 = every GEMM with
        weights as TF32 hi + lo, 2 (default) = hi + lo for the hidden layers and round-to-nearest TF32 weights for the input
        layer, 1 = weights rounded to TF32 everywhere (include/fdsm.h quotes the measured errors); 4 = as 2 with the input layer
        in kind::f16 (fp16 copies of W_0 and of the rows [x | v]: the same 11 significant bits as TF32 in half the bytes, which is
        what phase 0 of the kernel is bound by; needs |x|, |W_0| < 65504, else the loss turns NaN).  `passes` / `presplit`
        configure the stacked path and `net(x)`."""
        super().__init__()
        if act not in ACTS:
            raise ValueError(f"act must be one of {sorted(ACTS)}")
        if head not in ("quadratic", "linear"):
            raise ValueError("head must be 'quadratic' or 'linear'")
        self.in_dim, self.hidden, self.n_hidden, self.act, self.head, self.passes = in_dim, hidden, n_hidden, act, head, passes
        # presplit=False: activations stay plain fp32 in HBM, the GEMMs split into TF32 hi/lo in shared memory;
        # presplit=True: hi/lo halves are materialised in HBM by the producing kernels (twice the traffic)
        self.presplit = bool(presplit) or passes != 3
        self.desc = MlpDesc(in_dim, hidden, n_hidden, ACTS[act], 1 if head == "quadratic" else 0, passes, int(self.presplit))
        if formulation not in ("auto", "jet", "stacked"):
            raise ValueError("formulation must be 'auto', 'jet' or 'stacked'")
        self.jet_desc = MlpDesc(in_dim, hidden, n_hidden, ACTS[act], 1 if head == "quadratic" else 0, int(jet_passes), 0)
        jet_ok = bool(_lib.load().fdsm_jet_supported(C.byref(self.jet_desc)))
        if formulation == "jet" and not jet_ok:
            raise ValueError("the chain kernel supports hidden <= 223 and n_hidden <= 5")
        self.formulation = "jet" if (formulation in ("auto", "jet") and jet_ok) else "stacked"
        self.jet_passes = int(jet_passes)
        self._sizes = {}
        n = self.sizes(0).n_params
        self.flat = nn.Parameter(torch.zeros(n, device=device, dtype=torch.float32))
        self.reset_parameters(seed)
        self._step_state = None     # the state of the most recent step (bench.py reads its sizes)
        self._states = {}
        self._prep_stream = None

    # ---- parameter bookkeeping (reference named_parameters order) ----------------------------------------------
    def param_shapes(self):
        names = []
        for l in range(self.n_hidden):
            k = self.in_dim if l == 0 else self.hidden
            names += [(f"fc{l + 1}.weight", (self.hidden, k)), (f"fc{l + 1}.bias", (self.hidden,))]
        heads = ("ln1", "ln2", "lq") if self.head == "quadratic" else ("out",)
        for h in heads:
            names += [(f"{h}.weight", (1, self.hidden)), (f"{h}.bias", (1,))]
        return names

    def named_views(self, flat=None):
        """name -> view into the flat vector (or into `flat`, e.g. the gradient)."""
        flat = self.flat.data if flat is None else flat
        out, o = {}, 0
        for name, shape in self.param_shapes():
            n = int(np.prod(shape))
            out[name] = flat[o:o + n].view(shape)
            o += n
        assert o == flat.numel()
        return out

    def reset_parameters(self, seed=None):
        """nn.Linear's default range U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weights and biases, drawn from numpy's
        MT19937 in named_parameters order (platform-independent given `seed`)."""
        rs = np.random.RandomState(seed)
        with torch.no_grad():
            views = self.named_views()
            for name, shape in self.param_shapes():
                wshape = shape if name.endswith("weight") else dict(self.param_shapes())[name[:-4] + "weight"]
                bound = 1.0 / math.sqrt(wshape[1])
                views[name].copy_(torch.from_numpy(rs.uniform(-bound, bound, size=shape)).float())

    @torch.no_grad()
    def load_reference_state_dict(self, sd):
        """Accepts a state_dict of the reference's MLP_Quadratic* (or an nn.Sequential-style 3-layer MLP given as an
        ordered list of tensors); 'module.' DataParallel prefixes are stripped."""
        views = self.named_views()
        if isinstance(sd, (list, tuple)):
            for (name, _), t in zip(self.param_shapes(), sd):
                views[name].copy_(t.reshape(views[name].shape))
            return self
        sd = {k[7:] if k.startswith("module.") else k: v for k, v in sd.items()}
        for name, _ in self.param_shapes():
            views[name].copy_(sd[name].reshape(views[name].shape))
        return self

    def sizes(self, rows):
        if rows not in self._sizes:
            self._sizes[rows] = _Sizes(self.desc, rows)
        return self._sizes[rows]

    # ---- nn.Module drop-in ---------------------------------------------------------------------------------------
    def forward(self, x):
        x2 = x.reshape(x.shape[0], -1)
        if x2.shape[1] != self.in_dim:
            raise ValueError(f"expected {self.in_dim} input features, got {x2.shape[1]}")
        return _MlpFunction.apply(self.flat, x2, self)

    # ---- fused FD-SSM (a1 / a7): kernels (1)+(2) -> (3) fwd -> (4) fwd+bwd -> (3) bwd ---------------------------
    def _state(self, B, device, nrow=3):
        """Persistent buffers for batch size B and `nrow` stacked rows per sample (3: [x, x+v, x-v]; 2: [xt+v, xt-v]):
        addresses stay fixed, so the sequence is CUDA-graph capturable.  One state per (B, nrow), kept for the life of
        the module: a captured graph holds raw pointers into its state, so a state is never dropped when the batch size
        changes (last, shorter batch of an epoch) and comes back."""
        key = (B, nrow, str(device))
        st = self._states.get(key)
        if st is None:
            rows = nrow * B
            sz = self.sizes(rows)
            f = dict(device=device, dtype=torch.float32)
            st = dict(B=B, nrow=nrow, rows=rows, sz=sz,
                      xh=torch.empty(rows, sz.in_ld, **f), xl=torch.empty(rows, sz.in_ld, **f) if self.presplit else None,
                      ws=torch.empty(sz.workspace, **f), E=torch.empty(rows, **f), dE=torch.empty(rows, **f),
                      dot=torch.empty(B, **f) if nrow == 2 else None,
                      loss=torch.zeros((), **f), grad=torch.zeros(sz.n_params, **f))
            self._states[key] = st
        self._step_state = st
        return st

    def _jet_state(self, B, device):
        key = ("jet", B, str(device))
        st = self._states.get(key)
        if st is None:
            n, w = C.c_int64(), C.c_int64()
            offs = (C.c_int64 * (8 + 4 * self.n_hidden))()
            check(_lib.load().fdsm_jet_sizes(C.byref(self.jet_desc), B, C.byref(n), C.byref(w), offs), "fdsm_jet_sizes")
            f = dict(device=device, dtype=torch.float32)
            st = dict(B=B, nrow=3, rows=3 * B, jet=True, offsets=list(offs), ws=torch.zeros(w.value, **f), loss=torch.zeros((), **f),
                      grad=torch.zeros(n.value, **f), xh=None, xl=None)
            self._states[key] = st
        self._step_state = st
        return st

    def _jet_loss_and_grad(self, x2, eps, scale, u, v, draw, c1, mean, offset_dev, update=False, x_noise=0.0, mode=0,
                           normalize=None):
        """The same loss and gradient through the difference-propagating chain kernel (fdsm_jet_fd_ssm); `update`: the clipped
        Adam step of configure_optimizer() in the same call (fdsm_jet_update: one fused launch for reduce + clip + Adam)."""
        lib = _lib.load()
        B, D = x2.shape
        st = self._jet_state(B, x2.device)
        a = JetArgs()
        if x2.dtype == torch.uint8:      # pixels: x = x_u8 / 255 + x_noise * N(0,1), built inside the input kernel
            if not (x2.is_cuda and x2.is_contiguous()):
                raise _lib.FdsmError("uint8 samples must be a contiguous CUDA tensor")
            a.x_u8, a.ldx, a.B = C.c_void_p(x2.data_ptr()), D, B
            a.x_u8_scale, a.x_noise_std = 1.0 / 255.0, float(x_noise)
        else:
            x2 = ops._req(x2, "samples")
            x2, ldx = ops._rows(x2, "samples")
            a.x, a.ldx, a.B = _ptr(x2), ldx, B
        keep = [x2]
        d = v if v is not None else u
        if d is not None:
            d, ldd = ops._rows(ops._req(d.reshape(B, -1), "direction"), "direction")
            a.dir, a.lddir = _ptr(d), ldd
            keep.append(d)
        norm = (v is None) if normalize is None else bool(normalize)
        a.normalize_dir, a.dir_scale, a.mode = int(norm), float(eps if norm else 1.0), int(mode)
        seed, offset, row0 = draw if draw is not None else ops.next_draw()
        a.seed, a.offset, a.row0 = seed, offset, row0
        if offset_dev is not None:
            a.offset_dev = _ptr(offset_dev)
        a.params, a.workspace = _ptr(self.flat.data), _ptr(st["ws"])
        a.c1, a.scale, a.mean = float(c1), float(scale), int(mean)
        a.loss, a.grads, a.accumulate = _ptr(st["loss"]), _ptr(st["grad"]), 0
        if ops.CHAIN_TRACE is not None:      # scripts/chain_trace.py: int64 tensor (tiles, 38) of %globaltimer stamps
            a.trace = _ptr(ops.CHAIN_TRACE)
        hook = ops.PROFILE_HOOK
        if hook is not None:   # bench.py: CUDA events around the chain kernel, recorded by the library on the launching stream
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); e1.record()        # creates the underlying cudaEvent_t handles
            a.ev_chain_start, a.ev_chain_stop = C.c_void_p(e0.cuda_event), C.c_void_p(e1.cuda_event)
        if update:
            o = self._opt
            if "update_scratch" not in o:
                o["update_scratch"] = torch.zeros(1024, device=self.flat.device)
            ju = _lib.JetUpdate()
            ju.exp_avg, ju.exp_avg_sq = _ptr(o["m"]), _ptr(o["v"])
            ju.lr, ju.lr_dev = o["lr"], _ptr(o["lr_dev"])
            ju.beta1, ju.beta2, ju.eps, ju.max_norm = o["betas"][0], o["betas"][1], o["eps"], o["clip"]
            ju.step_counter, ju.scratch = C.c_void_p(o["step"].data_ptr()), _ptr(o["update_scratch"])
            a.update = C.cast(C.pointer(ju), C.c_void_p)
            keep.append(ju)
        # launches of the default configuration: front (weight prep + input build), chain, batched weight-gradient GEMMs, and the
        # fixed-order reduce (+ clip + Adam when `update`)
        ops._count(4)
        check(lib.fdsm_jet_fd_ssm(C.byref(self.jet_desc), C.byref(a), _stream()), "fdsm_jet_fd_ssm")
        if hook is not None:
            hook.append((e0, e1))
        st["_keep"] = keep
        return st["loss"], st["grad"]

    def exact_ssm_loss_and_grad(self, x, *, u=None, draw=None, scale=1.0, mean=True, noise_type="gaussian", offset_dev=None):
        """EXACT sliced score matching, loss = scale * mean_b[(u.grad E)^2 / 2 - u' Hess E u] (SSM/losses/sliced_sm.py:7-34
        single_sliced_score_matching; DeepEBM/ESM/loss_train.py:173-189 ssm), and its gradient w.r.t. the flat parameters --
        in FORWARD mode through the chain kernel (fdsm_jet_args.mode = 1: the streams carry h, dh/dtau, 1/2 d2h/dtau2 of
        tau -> h(x + tau u)), no double backward.  `u`: the projection vectors (B, D) as given, or None: Philox N(0,1) rows
        ("gaussian"), scaled to norm sqrt(D) for noise_type "sphere" (sliced_sm.py:14-15)."""
        if self.formulation != "jet":
            raise NotImplementedError("exact SSM in forward mode needs the chain kernel (hidden <= 223, n_hidden <= 5)")
        x2 = x.reshape(x.shape[0], -1)
        if noise_type not in ("gaussian", "sphere"):
            raise ValueError("noise_type must be 'gaussian' or 'sphere' (pass sign vectors as `u` for 'radermacher')")
        sphere = u is None and noise_type == "sphere"
        return self._jet_loss_and_grad(x2, math.sqrt(x2.shape[1]) if sphere else 1.0, scale, None, u, draw, 2.0, mean, offset_dev,
                                       mode=1, normalize=sphere)

    def exact_ssm_loss(self, x, *, u=None, draw=None, scale=1.0, noise_type="gaussian"):
        """The exact-SSM loss as an autograd scalar (its backward feeds `self.flat.grad`): what losses.ssm returns for a
        fused net."""
        return _ExactSsmFunction.apply(self.flat, x, u, self, draw, float(scale), noise_type)

    def fd_ssm_loss_and_grad(self, x, eps, scale, *, u=None, v=None, draw=None, c1=2.0, mean=True, offset_dev=None):
        """loss = scale * mean_b[(E2-E3)^2/8 - E2 - E3 + c1 E1] on the stacked batch [x, x+v, x-v] and its gradient
        w.r.t. the flat parameters (SSM/losses/efficient_sm.py:47-73 with scale = D/eps^2; DeepEBM/ESM/loss_train.py:
        212-234 with scale = 1/eps^2).  Returns (loss 0-dim tensor, flat gradient); both are persistent buffers."""
        if self.formulation == "jet":
            return self._jet_loss_and_grad(x.reshape(x.shape[0], -1), eps, scale, u, v, draw, c1, mean, offset_dev)
        lib = _lib.load()
        x2 = x.reshape(x.shape[0], -1)
        B = x2.shape[0]
        st = self._state(B, x2.device)
        rows, d = st["rows"], self.desc
        s = _stream()
        flat = self.flat.data
        # weight preparation is independent of the stack build: fork it onto a side stream (capturable fork/join)
        cur = torch.cuda.current_stream()
        if self._prep_stream is None:
            self._prep_stream = torch.cuda.Stream()
        self._prep_stream.wait_stream(cur)
        with torch.cuda.stream(self._prep_stream):
            ops._count(1); check(lib.fdsm_mlp_prepare(C.byref(d), rows, _ptr(flat), _ptr(st["ws"]), _stream()), "fdsm_mlp_prepare")
        shook = ops.PROFILE_HOOK_STACK      # bench.py: CUDA events around the stack-build launch
        if shook is not None:
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
        if v is not None:
            ops.stack(x2, row_mask=7, scale=1.0, direction=v.reshape(B, -1), normalize=False, out=st["xh"], out_lo=st["xl"],
                      pad_ones=not self.presplit)
        else:
            ops.stack(x2, row_mask=7, scale=eps, direction=None if u is None else u.reshape(B, -1), normalize=True,
                      out=st["xh"], out_lo=st["xl"], draw=draw, offset_dev=offset_dev, pad_ones=not self.presplit)
        if shook is not None:
            s1.record()
            shook.append((s0, s1))
        cur.wait_stream(self._prep_stream)
        hook = ops.PROFILE_HOOK
        if hook is not None:   # bench.py: CUDA events around the forward GEMM launches, on the launching stream
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        ops._count(self.n_hidden); check(lib.fdsm_mlp_forward(C.byref(d), _ptr(st["xh"]), _ptr(st["xl"]), rows, _ptr(flat),
                                                            _ptr(st["ws"]), _ptr(st["E"]), s), "fdsm_mlp_forward")
        if hook is not None:
            e1.record()
            hook.append((e0, e1))
        # combine forward (loss) + combine backward (dE, analytic) + MLP backward: one entry point, dE never materialised
        ops._count(3 * self.n_hidden + 1)
        check(lib.fdsm_mlp_backward_fd_ssm(C.byref(d), _ptr(st["xh"]), _ptr(st["xl"]), B, _ptr(flat), _ptr(st["ws"]), _ptr(st["E"]),
                                           float(c1), float(scale), int(mean), _ptr(st["loss"]), _ptr(st["grad"]), 0, s),
              "fdsm_mlp_backward_fd_ssm")
        return st["loss"], st["grad"]

    # ---- fused FD-DSM (a9 mdsm_fd): kernels (1)+(2) noised 2-row form -> (3) fwd -> (4) folded into (3) bwd ------------
    def fd_dsm_loss_and_grad(self, x, sigma_b, *, eps, sigma02, loss_coe, u=None, dsm_noise=None, draw=None, offset_dev=None):
        """DeepEBM/ESM/loss_train.py:96-118 on the fused MLP: xt = x + sigma_b * n, v = eps * u/||u||, one forward on
        cat[xt + v, xt - v] (2B rows), loss = mean_b[((logp1 - logp2)/2 + v.(xt - x)/sigma0^2)^2 / sigma_b^2] * loss_coe *
        D / eps^2, and its gradient w.r.t. the flat parameters.  `sigma_b` (B,) is the per-batch-position noise schedule
        of train_dsm.py:50-57.  RNG order as in the reference: noise, then direction (injected for parity, else the two
        Philox streams of one draw).  Returns (loss 0-dim tensor, flat gradient); both are persistent buffers."""
        lib = _lib.load()
        x2 = x.reshape(x.shape[0], -1)
        B, D = x2.shape
        st = self._state(B, x2.device, nrow=2)
        rows, d = st["rows"], self.desc
        s = _stream()
        flat = self.flat.data
        sig = sigma_b.reshape(-1).to(torch.float32)
        if sig.numel() != B:
            raise RuntimeError(f"sigmas has {sig.numel()} entries for batch size {B}")
        alpha, beta, gamma = 0.5, 1.0 / float(sigma02), float(loss_coe) / float(eps) ** 2 * D
        cur = torch.cuda.current_stream()
        if self._prep_stream is None:
            self._prep_stream = torch.cuda.Stream()
        self._prep_stream.wait_stream(cur)
        with torch.cuda.stream(self._prep_stream):
            ops._count(1); check(lib.fdsm_mlp_prepare(C.byref(d), rows, _ptr(flat), _ptr(st["ws"]), _stream()), "fdsm_mlp_prepare")
        ops.stack(x2, row_mask=ops.ROW_PLUS | ops.ROW_MINUS, scale=eps, direction=None if u is None else u.reshape(B, -1),
                  normalize=True, noise=None if dsm_noise is None else dsm_noise.reshape(B, -1), sigma=sig, perturb=True,
                  dot_mode=1, dot_out=st["dot"], out=st["xh"], out_lo=st["xl"], draw=draw, offset_dev=offset_dev,
                  pad_ones=not self.presplit)
        cur.wait_stream(self._prep_stream)
        ops._count(self.n_hidden); check(lib.fdsm_mlp_forward(C.byref(d), _ptr(st["xh"]), _ptr(st["xl"]), rows, _ptr(flat),
                                                            _ptr(st["ws"]), _ptr(st["E"]), s), "fdsm_mlp_forward")
        # DSM combine forward (loss) + analytic dE + MLP backward: one entry point, dE never materialised
        ops._count(3 * self.n_hidden + 1)
        check(lib.fdsm_mlp_backward_fd_dsm(C.byref(d), _ptr(st["xh"]), _ptr(st["xl"]), B, _ptr(flat), _ptr(st["ws"]), _ptr(st["E"]),
                                           _ptr(st["dot"]), _ptr(sig), alpha, beta, gamma, _ptr(st["loss"]), _ptr(st["grad"]), 0, s),
              "fdsm_mlp_backward_fd_dsm")
        st["sig"] = sig      # keeps the schedule alive for graph replays
        return st["loss"], st["grad"]

    # ---- the training step: FD-SSM loss -> backward -> (all-reduce) -> clip -> Adam, graph-replayed ----------------
    def flat_params(self):
        return self

    def configure_optimizer(self, lr=1e-5, betas=(0.9, 0.95), clip=0.1, adam_eps=1e-8, use_graph=True, seed=0x5EED, rank=0):
        dev = self.flat.device
        self._opt = dict(lr=float(lr), betas=(float(betas[0]), float(betas[1])), clip=float(clip) if clip else 0.0,
                         eps=float(adam_eps), use_graph=bool(use_graph), seed=int(seed), rank=int(rank),
                         m=torch.zeros_like(self.flat.data), v=torch.zeros_like(self.flat.data),
                         sumsq=torch.zeros(32, device=dev), step=torch.zeros((), device=dev, dtype=torch.int64),
                         lr_dev=torch.full((), float(lr), device=dev))
        self._graphs = {}

    def set_lr(self, lr):
        """Learning-rate schedules (DeepEBM/ESM/inputs.py:75-90) write the device scalar the captured step reads."""
        self._opt["lr_dev"].fill_(float(lr))

    def _loss_and_grad_captured(self, st, kind, cfg, update=False):
        o = self._opt
        draw = (o["seed"], 0, o["rank"] * st["B"])      # offset = device step counter; rows keyed by the GLOBAL sample index
        if kind == "ssm":
            eps, scale = cfg
            if eps is None:        # exact SSM (ssm of loss_train.py:173-189): Gaussian projections, forward mode
                x2 = st["x"].reshape(st["x"].shape[0], -1)
                return self._jet_loss_and_grad(x2, 1.0, scale, None, None, draw, 2.0, True, o["step"], update=update, mode=1,
                                               normalize=False, x_noise=st.get("x_noise", 0.0))
            if update or st["x"].dtype == torch.uint8:
                x2 = st["x"].reshape(st["x"].shape[0], -1)
                return self._jet_loss_and_grad(x2, eps, scale, None, None, draw, 2.0, True, o["step"], update=update,
                                               x_noise=st.get("x_noise", 0.0))
            return self.fd_ssm_loss_and_grad(st["x"], eps, scale, draw=draw, offset_dev=o["step"])
        eps, sigma02, loss_coe = cfg
        return self.fd_dsm_loss_and_grad(st["x"], st["sigma_b"], eps=eps, sigma02=sigma02, loss_coe=loss_coe, draw=draw,
                                         offset_dev=o["step"])

    def _apply_update(self, st, have_sumsq=False):
        lib, o = _lib.load(), self._opt
        n, s = self.flat.numel(), _stream()
        ops._count(1 if have_sumsq else 2)
        if not have_sumsq:
            check(lib.fdsm_grad_sumsq(_ptr(st["grad"]), n, _ptr(o["sumsq"]), _ptr(o["step"]), s), "fdsm_grad_sumsq")
        check(lib.fdsm_adam_step(_ptr(self.flat.data), _ptr(st["grad"]), _ptr(o["m"]), _ptr(o["v"]), n, o["lr"], _ptr(o["lr_dev"]),
                                 o["betas"][0], o["betas"][1], o["eps"], o["clip"], _ptr(o["sumsq"]), _ptr(o["step"]), s),
              "fdsm_adam_step")

    def fd_ssm_step(self, x, eps, allreduce=None, scale=None, distributed=False, borrow=False, x_noise=0.0):
        """One optimisation step of DeepEBM/train_ssm.py:63-69 on a device batch x (B, ...).  Returns the loss (device
        scalar, persistent buffer).  The whole step is ONE CUDA-graph replay -- also for world > 1 when `allreduce` is the
        peer-memory kernel (ddp.PeerAllReduce); with the NCCL library collective it is two replays around the all-reduce.
        `borrow`: the step reads x where it lies instead of copying it into its own input buffer first -- for callers that
        hand in one of a few long-lived buffers (the trainer's H2D staging pair): one captured graph per buffer address.
        eps=None runs the EXACT sliced score matching objective instead (exact_ssm_loss_and_grad, Gaussian projections).
        A uint8 x is a batch of PIXELS: the input kernel builds x / 255 + x_noise * N(0,1) (torchvision ToTensor and the
        noise of DeepEBM/train_ssm.py:62) on the fly -- the host ships one byte per pixel."""
        x2 = x.reshape(x.shape[0], -1)
        st = self._jet_state(x2.shape[0], x2.device) if self.formulation == "jet" else self._state(x2.shape[0], x2.device)
        if eps is None and self.formulation != "jet":
            raise NotImplementedError("the exact-SSM step (eps=None) needs the chain kernel")
        scale = (1.0 if eps is None else 1.0 / eps ** 2) if scale is None else scale
        if x2.dtype == torch.uint8:
            if self.formulation != "jet":
                raise NotImplementedError("uint8 pixel batches need the difference-propagating path (formulation='jet')")
            if not borrow:
                raise ValueError("uint8 pixel batches are read where they lie: pass borrow=True")
            st["x"], st["x_noise"], xkey = x2, float(x_noise), (x2.data_ptr(), float(x_noise))
        elif borrow and x2.is_contiguous() and x2.dtype == torch.float32:
            st["x"], xkey = x2, x2.data_ptr()
        else:
            if st.get("x_own") is None:
                st["x_own"] = torch.empty_like(x2, memory_format=torch.contiguous_format)
            st["x_own"].copy_(x2)
            st["x"], xkey = st["x_own"], 0
        return self._train_step(st, "ssm", (None if eps is None else float(eps), float(scale)), allreduce, distributed, xkey)

    def mdsm_fd_step(self, x, sigmas, sigma02=0.01, eps=0.05, loss_coe=1e-4, allreduce=None, distributed=False):
        """One optimisation step of DeepEBM/train_dsm.py:62-71 with loss_func = mdsm_fd (loss_train.py:96-118): `sigmas`
        is the (B, 1) per-batch-position noise schedule of train_dsm.py:50-57 (a rank of a sharded batch passes its slice);
        defaults = configs/default.yaml (sigma0 0.1, esm_eps 0.05, loss_coe 1e-4).  Same graph-replayed sequence as
        fd_ssm_step with the noised 2-row stack and the DSM combine.  The reference's `tloss.item() > 1e8` skip
        (train_dsm.py:64-66) needs a host read of the loss every step and is left to the caller."""
        x2 = x.reshape(x.shape[0], -1)
        B = x2.shape[0]
        st = self._state(B, x2.device, nrow=2)
        if "x" not in st:
            st["x"] = torch.empty_like(x2)
            st["sigma_b"] = torch.empty(B, device=x2.device, dtype=torch.float32)
        if sigmas.numel() != B:
            raise RuntimeError(f"sigmas has {sigmas.numel()} entries for batch size {B}")
        st["x"].copy_(x2)
        st["sigma_b"].copy_(sigmas.reshape(-1))
        return self._train_step(st, "dsm", (float(eps), float(sigma02), float(loss_coe)), allreduce, distributed)

    def _train_step(self, st, kind, cfg, allreduce, distributed, xkey=0):
        self._last_graphs = None
        peer = distributed and getattr(allreduce, "fused_sumsq", False)    # ddp.PeerAllReduce: capturable, emits the sumsq
        key = (st["B"], kind, cfg, bool(distributed), bool(peer))

        # single GPU, difference-propagating path: the update rides in the same library call (one fused launch)
        fused = (self.formulation == "jet" and kind == "ssm" and not distributed and
                 os.environ.get("FDSM_JET_FUSED_UPDATE", "1") != "0")
        key = key + (fused, xkey)

        def exchange_and_update():
            if fused:
                return
            if peer:
                allreduce(st["grad"], sumsq=self._opt["sumsq"], step=self._opt["step"])
                ops._count(1)
                self._apply_update(st, have_sumsq=True)
            else:
                self._apply_update(st)

        if not self._opt["use_graph"]:
            self._loss_and_grad_captured(st, kind, cfg, update=fused)
            if distributed and allreduce is not None and not peer:
                allreduce(st["grad"])
            exchange_and_update()
            return st["loss"]
        if key not in self._graphs:
            # warm-up outside capture (module-level one-time setup inside the library), then capture
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._loss_and_grad_captured(st, kind, cfg)           # (no update: parameters and step counter untouched)
            torch.cuda.current_stream().wait_stream(side)
            launches0 = ops.LAUNCHES
            g1, g2 = torch.cuda.CUDAGraph(), None
            one_graph = not distributed or peer
            with torch.cuda.graph(g1):
                self._loss_and_grad_captured(st, kind, cfg, update=fused)
                if one_graph:
                    exchange_and_update()
            if not one_graph:
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2):
                    self._apply_update(st)
            self._graphs[key] = (g1, g2, ops.LAUNCHES - launches0)
            ops.LAUNCHES = launches0
        g1, g2, nlaunch = self._graphs[key]
        self._last_graphs = self._graphs[key]        # train.FDSSMTrainer's lean host path replays it directly next time
        g1.replay()
        if g2 is not None:
            if allreduce is not None:
                allreduce(st["grad"])
            g2.replay()
        ops._count(nlaunch)
        return st["loss"]

    def forward_flops(self, rows):
        """Algorithmic flops of the forward GEMMs (2*M*N*K per layer; bias column, padding and head dots not credited)."""
        return 2.0 * rows * (self.in_dim * self.hidden + (self.n_hidden - 1) * self.hidden * self.hidden)

    def roofline(self, events, peaks, peaks_kind):
        """bench.py: tensor-pipe roofline of the GEMM kernel from live CUDA-event timings of fdsm_mlp_forward
        (n_hidden launches of gemm_kernel back to back).  Peak = measured dense bf16 / 2 (TF32 runs at half the bf16
        rate); `achieved` counts ALGORITHMIC fp32 flops, the 3xTF32 scheme executes 3x that on the pipe."""
        if not events:
            return None
        ms = sorted(a.elapsed_time(b) for a, b in events)
        avg = sum(ms) / len(ms)
        if self.formulation == "jet":
            B, D, H, L = self._step_state["B"], self.in_dim, self.hidden, self.n_hidden
            fwd = 2.0 * (2 * B * D * H + 3 * B * H * H * (L - 1))          # layer 0 on the 2B rows [x | v], the rest on 3B
            dgrad = 2.0 * 3 * B * H * H * (L - 1)
            flops = fwd + dgrad
            fwd0 = 2.0 * 2 * B * D * H
            executed = {1: flops, 2: fwd0 + 2 * (flops - fwd0), 3: 2 * flops, 4: fwd0 + 2 * (flops - fwd0)}[self.jet_passes]
            ach = flops / (avg * 1e-3) / 1e12
            peak = peaks["bf16_tflops"] / 2.0
            wdesc = {1: "TF32 weights, 1 MMA pass", 2: "hidden-layer weights as TF32 hi+lo (2 MMA passes), input layer 1 pass",
                     3: "all weights as TF32 hi+lo, 2 MMA passes",
                     4: "hidden-layer weights as TF32 hi+lo (2 MMA passes), input layer in kind::f16 (fp16 operands, 1 pass)"}[self.jet_passes]
            return {"kernel": f"fdsm::jet::chain_kernel (tcgen05 kind::tf32, TMA-fed, {wdesc}; all {L} forward layers + head + combine + {L - 1} dgrad GEMMs in one launch)",
                    "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": None,
                    "peak_kind": f"{peaks_kind} dense bf16 burst / 2 (TF32 kind)", "passes": self.jet_passes,
                    "executed_tf32_tflops": ach * executed / flops, "executed_frac": ach * executed / flops / peak,
                    "algorithmic_flops_per_launch": flops, "avg_launch_us": avg * 1e3, "launches_timed": len(ms),
                    "stacked_formulation_flops": 2.0 * 3 * B * (D * H + (L - 1) * H * H) + dgrad,
                    "note": "achieved counts the 2*M*N*K of OUR formulation (layer 0 contracts the 2B rows [x | v]); the "
                            "reference's stacked evaluation of the same quantities needs stacked_formulation_flops"}
        rows = self._step_state["rows"]
        per_launch_flops = self.forward_flops(rows) / self.n_hidden
        per_launch_s = avg * 1e-3 / self.n_hidden
        ach = per_launch_flops / per_launch_s / 1e12
        peak = peaks["bf16_tflops"] / 2.0
        kname = "gemm_kernel" if self.presplit else "gemm_raw_kernel"
        return {"kernel": f"fdsm::tc::{kname} ({self.passes}-pass tcgen05 kind::tf32, TMA-fed; forward launches of the energy MLP)",
                "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": None,
                "peak_kind": f"{peaks_kind} dense bf16 burst / 2 (TF32 kind)", "passes": self.passes,
                "executed_tf32_tflops": ach * self.passes, "executed_frac": ach * self.passes / peak,
                "algorithmic_flops_per_launch": per_launch_flops, "avg_launch_us": per_launch_s * 1e6,
                "launches_timed": len(ms) * self.n_hidden}


def smoke(dev):
    """__graft_entry__.smoke(): one fused FD-SSM step on a tiny batch (finite loss, non-zero update)."""
    net = FusedEnergyMLP(64, 48, 3, act="nsoftplus", head="quadratic", device=dev, seed=7)
    net.configure_optimizer(lr=1e-3, betas=(0.9, 0.95), clip=0.1, use_graph=False)
    x = torch.rand(32, 64, device=dev)
    before = net.flat.data.clone()
    loss = net.fd_ssm_step(x, eps=0.05)
    assert torch.isfinite(loss).item(), "fused FD-SSM smoke: non-finite loss"
    assert float((net.flat.data - before).abs().max()) > 0, "fused FD-SSM smoke: parameters did not move"
