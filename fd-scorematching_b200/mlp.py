"""Fused energy MLPs: the reference's ``MLP_Quadratic_MNIST`` / ``MLP_Quadratic`` (DeepEBM/ESM/models.py:129-160,
932-955) and the config-1 3-layer Softplus MLP, forward and backward on the tcgen05 GEMMs of libfdsm
(``fdsm_mlp_*``, include/fdsm.h).  No torch matmul, no fallback.

Two ways in:
  * ``net(x)`` -- an ``nn.Module`` drop-in for the reference's energy networks: works with every energy-form loss
    adapter in ``losses`` (generic mode); parameters live in ONE flat ``nn.Parameter`` laid out in the reference's
    ``named_parameters()`` order, so ``load_reference_state_dict`` accepts the reference's checkpoints.
  * ``net.fd_ssm_loss_and_grad(x, eps, scale)`` -- the fused FD-SSM step: stack build (hi/lo TF32 operands written
    directly by kernel (1)+(2)), MLP forward, combine forward+backward (kernel (4)), MLP backward.
"""
import ctypes as C
import math

import numpy as np
import torch
import torch.nn as nn

from . import _lib, ops
from ._lib import MlpDesc, check

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
        rows = x2.shape[0]
        sz = net.sizes(rows)
        xh, xl = ops.tf32_split(x2, out_cols=sz.in_ld, ones_col=True)
        ws = torch.empty(sz.workspace, device=x2.device, dtype=torch.float32)
        E = torch.empty(rows, device=x2.device, dtype=torch.float32)
        d = net.desc
        ops._count(net.n_hidden); check(lib.fdsm_mlp_prepare(C.byref(d), rows, _ptr(flat), _ptr(ws), _stream()), "fdsm_mlp_prepare")
        ops._count(net.n_hidden); check(lib.fdsm_mlp_forward(C.byref(d), _ptr(xh), _ptr(xl), rows, _ptr(flat), _ptr(ws), _ptr(E),
                                                           _stream()), "fdsm_mlp_forward")
        ctx.save_for_backward(flat, xh, xl, ws)
        ctx.net, ctx.rows = net, rows
        return E

    @staticmethod
    def backward(ctx, dE):
        lib = _lib.load()
        flat, xh, xl, ws = ctx.saved_tensors
        net, rows = ctx.net, ctx.rows
        dE = dE.contiguous().float()
        g = torch.empty_like(flat)
        ops._count(3 * net.n_hidden + 1)
        check(lib.fdsm_mlp_backward(C.byref(net.desc), _ptr(xh), _ptr(xl), rows, _ptr(flat), _ptr(ws), _ptr(dE), _ptr(g), 0,
                                    _stream()), "fdsm_mlp_backward")
        return g, None, None


class FusedEnergyMLP(nn.Module):
    def __init__(self, in_dim, hidden, n_hidden=3, act="nsoftplus", head="quadratic", device="cuda", seed=None, passes=3):
        super().__init__()
        if act not in ACTS:
            raise ValueError(f"act must be one of {sorted(ACTS)}")
        if head not in ("quadratic", "linear"):
            raise ValueError("head must be 'quadratic' or 'linear'")
        self.in_dim, self.hidden, self.n_hidden, self.act, self.head, self.passes = in_dim, hidden, n_hidden, act, head, passes
        self.desc = MlpDesc(in_dim, hidden, n_hidden, ACTS[act], 1 if head == "quadratic" else 0, passes)
        self._sizes = {}
        n = self.sizes(0).n_params
        self.flat = nn.Parameter(torch.zeros(n, device=device, dtype=torch.float32))
        self.reset_parameters(seed)
        self._step_state = None

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
    def _state(self, B, device):
        """Persistent buffers for batch size B (rows = 3B): addresses stay fixed, so the sequence is CUDA-graph capturable."""
        st = self._step_state
        if st is None or st["B"] != B:
            rows = 3 * B
            sz = self.sizes(rows)
            f = dict(device=device, dtype=torch.float32)
            st = dict(B=B, rows=rows, sz=sz,
                      xh=torch.empty(rows, sz.in_ld, **f), xl=torch.empty(rows, sz.in_ld, **f),
                      ws=torch.empty(sz.workspace, **f), E=torch.empty(rows, **f), dE=torch.empty(rows, **f),
                      loss=torch.zeros((), **f), grad=torch.zeros(sz.n_params, **f))
            self._step_state = st
        return st

    def fd_ssm_loss_and_grad(self, x, eps, scale, *, u=None, v=None, draw=None, c1=2.0, mean=True):
        """loss = scale * mean_b[(E2-E3)^2/8 - E2 - E3 + c1 E1] on the stacked batch [x, x+v, x-v] and its gradient
        w.r.t. the flat parameters (SSM/losses/efficient_sm.py:47-73 with scale = D/eps^2; DeepEBM/ESM/loss_train.py:
        212-234 with scale = 1/eps^2).  Returns (loss 0-dim tensor, flat gradient); both are persistent buffers."""
        lib = _lib.load()
        x2 = x.reshape(x.shape[0], -1)
        B = x2.shape[0]
        st = self._state(B, x2.device)
        rows, d = st["rows"], self.desc
        if v is not None:
            ops.stack(x2, row_mask=7, scale=1.0, direction=v.reshape(B, -1), normalize=False, out=st["xh"], out_lo=st["xl"])
        else:
            ops.stack(x2, row_mask=7, scale=eps, direction=None if u is None else u.reshape(B, -1), normalize=True,
                      out=st["xh"], out_lo=st["xl"], draw=draw)
        s = _stream()
        flat = self.flat.data
        ops._count(self.n_hidden); check(lib.fdsm_mlp_prepare(C.byref(d), rows, _ptr(flat), _ptr(st["ws"]), s), "fdsm_mlp_prepare")
        ops._count(self.n_hidden); check(lib.fdsm_mlp_forward(C.byref(d), _ptr(st["xh"]), _ptr(st["xl"]), rows, _ptr(flat),
                                                            _ptr(st["ws"]), _ptr(st["E"]), s), "fdsm_mlp_forward")
        ops._count(2)
        check(lib.fdsm_combine_energy_fwd(_ptr(st["E"]), B, 0, float(c1), float(scale), int(mean), None, _ptr(st["loss"]), s),
              "fdsm_combine_energy_fwd")
        check(lib.fdsm_combine_energy_bwd(_ptr(st["E"]), B, 0, float(c1), float(scale), int(mean), None, _ptr(st["dE"]), s),
              "fdsm_combine_energy_bwd")
        ops._count(3 * self.n_hidden + 1)
        check(lib.fdsm_mlp_backward(C.byref(d), _ptr(st["xh"]), _ptr(st["xl"]), rows, _ptr(flat), _ptr(st["ws"]), _ptr(st["dE"]),
                                    _ptr(st["grad"]), 0, s), "fdsm_mlp_backward")
        return st["loss"], st["grad"]
