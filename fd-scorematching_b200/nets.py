"""The convolutional networks the FD losses drive in BASELINE configs 3-5, as
plain torch modules for the GENERIC mode of the loss adapters: our kernels build the stacked batch and combine the
outputs, the network itself runs on cuDNN ("library-backed; reported, not judged").

    Res18Quadratic        DeepEBM/ESM/models.py:236-342 (ResidualBlock :163-233, quadratic head :338)      config 3
    CondRefineNetDilated  ncsn/models/cond_refinenet_dilated.py:336-417                                    config 4
    NCSNv2                ncsnv2/models/ncsnv2.py:11-101 (layers.py, normalization.py:150-176)             config 5

Parameter and buffer names equal the reference's, so its checkpoints load with `load_state_dict` (strip a `module.`
DataParallel prefix first) and the forward goldens in tests/golden/ are produced by loading ONE state dict into both.
The two RefineNets share one implementation here: a normaliser either owns (alpha, gamma, beta) vectors (NCSNv2's
InstanceNorm++) or looks them up in a per-noise-level embedding table (ncsn's conditional variant).
"""
import math
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops


# ---------------------------------------------------------------------------------------------------------------------
# InstanceNorm++ : instance norm plus a term that re-injects the (standardised) per-channel means
# ---------------------------------------------------------------------------------------------------------------------
def _inorm_plus(x, gamma, alpha, beta, act=False):
    """x (N,C,H,W); gamma / alpha / beta (N or 1, C); act: the ELU that follows the normaliser in the residual blocks.
    ncsnv2/models/normalization.py:163-176.  CUDA fp32 tensors go through our fused kernels (ops.inorm_plus: two launches
    per direction); FDSM_NORM_KERNEL=0 keeps the ATen op sequence there too (A/B timing).
    CPU tensors (the forward goldens of tests/test_nets_cpu.py) take the op sequence."""
    if x.is_cuda and x.dtype == torch.float32 and os.environ.get("FDSM_NORM_KERNEL", "1") != "0":
        return ops.inorm_plus(x, gamma, alpha, beta, act)
    return ops.inorm_plus_reference(x, gamma, alpha, beta, act)


class InstanceNormPlus(nn.Module):
    """Unconditional InstanceNorm++ (parameters alpha, gamma ~ N(1, 0.02), beta = 0)."""

    def __init__(self, channels, n_classes=None):
        super().__init__()
        self.alpha = nn.Parameter(torch.empty(channels).normal_(1.0, 0.02))
        self.gamma = nn.Parameter(torch.empty(channels).normal_(1.0, 0.02))
        self.beta = nn.Parameter(torch.zeros(channels))

    def forward(self, x, y=None, act=False):
        return _inorm_plus(x, self.gamma[None], self.alpha[None], self.beta[None], act)


class CondInstanceNormPlus(nn.Module):
    """Per-class (= per-noise-level) InstanceNorm++: one embedding row [gamma | alpha | beta] per class
    (ncsn/models/cond_refinenet_dilated.py:304-333)."""

    def __init__(self, channels, n_classes):
        super().__init__()
        self.embed = nn.Embedding(n_classes, 3 * channels)
        with torch.no_grad():
            self.embed.weight[:, :2 * channels].normal_(1.0, 0.02)
            self.embed.weight[:, 2 * channels:].zero_()

    def forward(self, x, y, act=False):
        gamma, alpha, beta = self.embed(y).chunk(3, dim=1)
        return _inorm_plus(x, gamma, alpha, beta, act)


def _conv(cin, cout, k=3, dilation=1, bias=True):
    return nn.Conv2d(cin, cout, k, padding=dilation * (k // 2), dilation=dilation, bias=bias)


def _pool2(x):
    """2x2 mean pool written as the sum of the four strided slices / 4 (ncsnv2/models/layers.py:308-312)."""
    return (x[:, :, ::2, ::2] + x[:, :, 1::2, ::2] + x[:, :, ::2, 1::2] + x[:, :, 1::2, 1::2]) / 4.0


class _ConvThenPool(nn.Module):
    def __init__(self, cin, cout, k):
        super().__init__()
        self.conv = _conv(cin, cout, k)

    def forward(self, x):
        return _pool2(self.conv(x))


# ---------------------------------------------------------------------------------------------------------------------
# RefineNet pieces (shared by ncsn's conditional net and NCSNv2)
# ---------------------------------------------------------------------------------------------------------------------
class _ScoreResBlock(nn.Module):
    """norm-act-conv-norm-act-conv + shortcut; `down` without dilation halves the resolution by mean pooling, with
    dilation the resolution stays (ncsnv2/models/layers.py:401-456)."""

    def __init__(self, cin, cout, norm, n_classes, down=False, dilation=None):
        super().__init__()
        d = dilation or 1
        mid = cin if down else cout
        self.conv1 = _conv(cin, mid, 3, d)
        self.normalize2 = norm(mid, n_classes)
        if down and dilation is None:
            self.conv2 = _ConvThenPool(mid, cout, 3)
        else:
            self.conv2 = _conv(mid, cout, 3, d)
        self.has_shortcut = down or cin != cout
        if self.has_shortcut:
            if down and dilation is None:
                self.shortcut = _ConvThenPool(cin, cout, 1)
            elif dilation is not None:
                self.shortcut = _conv(cin, cout, 3, d)
            else:
                self.shortcut = _conv(cin, cout, 1)
        self.normalize1 = norm(cin, n_classes)

    def forward(self, x, y):
        h = self.conv1(self.normalize1(x, y, act=True))
        h = self.conv2(self.normalize2(h, y, act=True))
        return h + (self.shortcut(x) if self.has_shortcut else x)


class _RCU(nn.Module):
    """n_blocks residual units of two (norm-)act-conv stages; bias-free 3x3 convs named `{block}_{stage}_conv`."""

    def __init__(self, ch, n_blocks, norm, n_classes, cond):
        super().__init__()
        self.n_blocks, self.cond = n_blocks, cond
        for b in range(1, n_blocks + 1):
            for s in (1, 2):
                if cond:
                    setattr(self, f"{b}_{s}_norm", norm(ch, n_classes))
                setattr(self, f"{b}_{s}_conv", _conv(ch, ch, 3, bias=False))

    def forward(self, x, y):
        for b in range(1, self.n_blocks + 1):
            h = x
            for s in (1, 2):
                h = getattr(self, f"{b}_{s}_norm")(h, y, act=True) if self.cond else F.elu(h)
                h = getattr(self, f"{b}_{s}_conv")(h)
            x = x + h
        return x


class _CRP(nn.Module):
    """Chained residual pooling: 5x5 stride-1 pool -> conv, accumulated.  NCSNv2 pools with max, ncsn's conditional net
    normalises then pools with avg (layers.py:62-83 vs cond_refinenet_dilated.py:93-113)."""

    def __init__(self, ch, norm, n_classes, cond):
        super().__init__()
        self.cond = cond
        self.convs = nn.ModuleList(_conv(ch, ch, 3, bias=False) for _ in range(2))
        if cond:
            self.norms = nn.ModuleList(norm(ch, n_classes) for _ in range(2))

    def forward(self, x, y):
        x = F.elu(x)
        path = x
        for i, conv in enumerate(self.convs):
            if self.cond:
                path = F.avg_pool2d(self.norms[i](path, y), 5, stride=1, padding=2)
            else:
                path = F.max_pool2d(path, 5, stride=1, padding=2)
            path = conv(path)
            x = x + path
        return x


class _MSF(nn.Module):
    """Multi-resolution fusion: conv each input to `ch` channels, bilinear-resize (align_corners) to `size`, sum."""

    def __init__(self, in_chs, ch, norm, n_classes, cond):
        super().__init__()
        self.cond = cond
        self.convs = nn.ModuleList(_conv(c, ch, 3) for c in in_chs)
        if cond:
            self.norms = nn.ModuleList(norm(c, n_classes) for c in in_chs)

    def forward(self, xs, y, size):
        total = None
        for i, conv in enumerate(self.convs):
            h = self.norms[i](xs[i], y) if self.cond else xs[i]
            h = F.interpolate(conv(h), size=size, mode="bilinear", align_corners=True)
            total = h if total is None else total + h
        return total


class _Refine(nn.Module):
    def __init__(self, in_chs, ch, norm, n_classes, cond, start=False, end=False):
        super().__init__()
        self.adapt_convs = nn.ModuleList(_RCU(c, 2, norm, n_classes, cond) for c in in_chs)
        self.output_convs = _RCU(ch, 3 if end else 1, norm, n_classes, cond)
        if not start:
            self.msf = _MSF(in_chs, ch, norm, n_classes, cond)
        self.crp = _CRP(ch, norm, n_classes, cond)

    def forward(self, xs, y, size):
        hs = [m(x, y) for m, x in zip(self.adapt_convs, xs)]
        h = self.msf(hs, y, size) if len(hs) > 1 else hs[0]
        return self.output_convs(self.crp(h, y), y)


class _RefineNet(nn.Module):
    """RefineNet score network s(x, y) -> x.shape on (N, C, S, S) images; 4 residual stages (the last two dilated 2, 4),
    4 refine blocks, InstanceNorm++ everywhere, ELU."""

    def __init__(self, channels, ngf, n_classes, cond, image_size=32):
        super().__init__()
        if image_size == 28:
            raise NotImplementedError("the adjust_padding variant for 28x28 inputs is not part of the BASELINE configs")
        norm = CondInstanceNormPlus if cond else InstanceNormPlus
        nc = n_classes
        self.begin_conv = _conv(channels, ngf, 3)
        self.normalizer = norm(ngf, nc)
        self.end_conv = _conv(ngf, channels, 3)

        def stage(cin, cout, down, dilation=None):
            return nn.ModuleList([_ScoreResBlock(cin, cout, norm, nc, down=down, dilation=dilation),
                                  _ScoreResBlock(cout, cout, norm, nc, dilation=dilation)])
        self.res1 = stage(ngf, ngf, False)
        self.res2 = stage(ngf, 2 * ngf, True)
        self.res3 = stage(2 * ngf, 2 * ngf, True, 2)
        self.res4 = stage(2 * ngf, 2 * ngf, True, 4)
        self.refine1 = _Refine([2 * ngf], 2 * ngf, norm, nc, cond, start=True)
        self.refine2 = _Refine([2 * ngf, 2 * ngf], 2 * ngf, norm, nc, cond)
        self.refine3 = _Refine([2 * ngf, 2 * ngf], ngf, norm, nc, cond)
        self.refine4 = _Refine([ngf, ngf], ngf, norm, nc, cond, end=True)

    def _features(self, x, y):
        h = self.begin_conv(2.0 * x - 1.0)          # data in [0, 1] -> [-1, 1] (logit_transform / rescaled both false)
        feats = []
        for blocks in (self.res1, self.res2, self.res3, self.res4):
            for blk in blocks:
                h = blk(h, y)
            feats.append(h)
        f1, f2, f3, f4 = feats
        r = self.refine1([f4], y, f4.shape[2:])
        r = self.refine2([f3, r], y, f3.shape[2:])
        r = self.refine3([f2, r], y, f2.shape[2:])
        r = self.refine4([f1, r], y, f1.shape[2:])
        return self.end_conv(self.normalizer(r, y, act=True))


class CondRefineNetDilated(_RefineNet):
    """ncsn/models/cond_refinenet_dilated.py:336-417 (ngf 128, 10 noise levels: configs/anneal_fdssmvr.yml:33-38)."""

    def __init__(self, channels=3, ngf=128, num_classes=10, image_size=32):
        super().__init__(channels, ngf, num_classes, cond=True, image_size=image_size)

    def forward(self, x, y):
        return self._features(x, y)


def geometric_sigmas(sigma_begin, sigma_end, num_classes):
    """ncsnv2/models/__init__.py:4-17 (sigma_dist: geometric)."""
    return torch.tensor(np.exp(np.linspace(np.log(sigma_begin), np.log(sigma_end), num_classes))).float()


class NCSNv2(_RefineNet):
    """ncsnv2/models/ncsnv2.py:11-101: unconditional InstanceNorm++ RefineNet whose output is divided by sigma[y]
    (:97-99).  Defaults = configs/cifar10_esm.yml:51-60 (232 geometric noise levels 50 -> 0.01, ngf 128)."""

    def __init__(self, channels=3, ngf=128, num_classes=232, sigma_begin=50.0, sigma_end=0.01, image_size=32):
        super().__init__(channels, ngf, num_classes, cond=False, image_size=image_size)
        self.register_buffer("sigmas", geometric_sigmas(sigma_begin, sigma_end, num_classes))

    def forward(self, x, y):
        out = self._features(x, y)
        return out / self.sigmas[y].view(-1, *([1] * (x.dim() - 1)))


# ---------------------------------------------------------------------------------------------------------------------
# DeepEBM ResNet energy (config 3)
# ---------------------------------------------------------------------------------------------------------------------
class _Wrapped(nn.Module):
    """A conv held under the attribute name `conv` (the reference wraps every conv, so its parameter names carry it)."""

    def __init__(self, cin, cout, k, bias=True, pool=None):
        super().__init__()
        self.conv = _conv(cin, cout, k, bias=bias)
        self.pool_mode = pool

    def forward(self, x):
        if self.pool_mode == "before":
            x = F.avg_pool2d(x, 2)
        x = self.conv(x)
        return F.avg_pool2d(x, 2) if self.pool_mode == "after" else x


class _WrappedTwice(nn.Module):
    """conv_2 / conv_shortcut of a down block: `<name>.conv.conv.weight` in the reference's state dict."""

    def __init__(self, cin, cout, k, pool):
        super().__init__()
        self.conv = _Wrapped(cin, cout, k)
        self.pool_mode = pool

    def forward(self, x):
        if self.pool_mode == "before":
            return self.conv(F.avg_pool2d(x, 2))
        return F.avg_pool2d(self.conv(x), 2)


class _EnergyResBlock(nn.Module):
    """Pre-activation block act-conv-act-conv + 1x1 shortcut; `down` halves the resolution with 2x2 average pooling
    (after the second conv, before the shortcut conv).  DeepEBM/ESM/models.py:163-233.  The reference constructs two
    LayerNorms per block that its forward never calls when `normalization: None` (default.yaml:11); they are kept as
    (unused) parameters so that state dicts and parameter counts match (64.17 M for Res18, n_chan 128)."""

    def __init__(self, cin, cout, hw, down, act):
        super().__init__()
        self.act = act
        self.bn1 = nn.LayerNorm([cin, hw, hw])
        self.bn2 = nn.LayerNorm([cin, hw, hw])
        self.conv_1 = _Wrapped(cin, cin, 3, bias=False)
        if down:
            self.conv_shortcut = _WrappedTwice(cin, cout, 1, "before")
            self.conv_2 = _WrappedTwice(cin, cout, 3, "after")
        else:
            self.conv_shortcut = _Wrapped(cin, cout, 1)
            self.conv_2 = _Wrapped(cin, cout, 3)

    def forward(self, x):
        h = self.conv_2(self.act(self.conv_1(self.act(x))))
        return self.conv_shortcut(x) + h


def _activation(name):
    if name == "elu":
        return F.elu
    if name == "softplus":
        return F.softplus
    if name == "nsoftplus":
        return lambda z: F.softplus(z) - math.log(2.0)
    if name == "relu":
        return F.relu
    raise ValueError(name)


class Res18Quadratic(nn.Module):
    """DeepEBM/ESM/models.py:236-342: conv3x3 -> 9 residual blocks (3 of them downsampling, widths dim*2/4/8) -> flatten
    -> E = ln1(h) * ln2(h) + lq(h^2).  Input (N, inchan*hw*hw) or (N, inchan, hw, hw); output (N,)."""

    def __init__(self, inchan=1, dim=128, hw=32, act="elu"):
        super().__init__()
        self.inchan, self.dim, self.hw = inchan, dim, hw
        a = _activation(act)
        self.conv1 = _Wrapped(inchan, dim, 3)
        plan = [("rb1", dim, 2 * dim, hw, True), ("rbc1", 2 * dim, 2 * dim, hw // 2, False),
                ("rbc11", 2 * dim, 2 * dim, hw // 2, False), ("rb2", 2 * dim, 4 * dim, hw // 2, True),
                ("rbc2", 4 * dim, 4 * dim, hw // 4, False), ("rbc22", 4 * dim, 4 * dim, hw // 4, False),
                ("rb3", 4 * dim, 8 * dim, hw // 4, True), ("rbc3", 8 * dim, 8 * dim, hw // 8, False),
                ("rbc33", 8 * dim, 8 * dim, hw // 8, False)]
        self._order = [p[0] for p in plan]
        for name, cin, cout, res, down in plan:
            setattr(self, name, _EnergyResBlock(cin, cout, res, down, a))
        feat = (hw // 8) ** 2 * 8 * dim
        self.ln1, self.ln2, self.lq = nn.Linear(feat, 1), nn.Linear(feat, 1), nn.Linear(feat, 1)

    def forward(self, x):
        h = self.conv1(x.reshape(-1, self.inchan, self.hw, self.hw))
        for name in self._order:
            h = getattr(self, name)(h)
        h = h.flatten(1)
        return (self.ln1(h) * self.ln2(h) + self.lq(h * h)).view(-1)
