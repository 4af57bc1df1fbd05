"""Drop-in nn.Modules with the reference's constructors, forward signatures and
state-dict keys, backed by the sm_100a kernels.

    AdaIN_block            retinanet/AdaIN.py:3-26
    CrossAttentionBlock    retinanet/Attention.py:7-36
    REFusion               retinanet/fusion.py:7-36
    Conv1x1                nn.Conv2d(Cin, Cout, kernel_size=1) of the FPN laterals, retinanet/model.py:28-40

The nn.Conv2d submodules are parameter containers only (their weights keep the
reference's (out, in, 1, 1) shape so published checkpoints load unchanged,
SURVEY.md section 5 "Checkpoint / resume"); their forward is never called.
"""
from __future__ import annotations

import torch.nn as nn

from . import functional as F


def c2_xavier_fill(module: nn.Module) -> None:
    """fvcore.nn.weight_init.c2_xavier_fill as used at Attention.py:16-17:
    kaiming_uniform_(a=1) on the weight, zeros on the bias."""
    nn.init.kaiming_uniform_(module.weight, a=1)
    if module.bias is not None:
        nn.init.constant_(module.bias, 0)


class AdaIN_block(nn.Module):
    def __init__(self):
        super().__init__()

    def calc_mean_std(self, feat, eps=1e-5):
        size = feat.size()
        assert (len(size) == 4)                                   # AdaIN.py:10
        return F.mean_std(feat, eps)

    def forward(self, rgb, evt):
        assert (rgb.size()[:2] == evt.size()[:2])                 # AdaIN.py:20
        assert len(rgb.size()) == 4 and len(evt.size()) == 4      # AdaIN.py:10 (via calc_mean_std)
        return F.adain(rgb, evt)


def _conv_params(conv):
    return conv.weight, conv.bias


class CrossAttentionBlock(nn.Module):
    def __init__(self, in_channels):
        super().__init__()
        self.g = nn.Conv2d(in_channels, in_channels // 8, kernel_size=1, stride=1, padding=0)
        self.theta = nn.Conv2d(in_channels, in_channels // 8, kernel_size=1, stride=1, padding=0)
        self.phi = nn.Conv2d(in_channels, in_channels // 8, kernel_size=1, stride=1, padding=0)
        self.W = nn.Conv2d(in_channels // 8, in_channels, kernel_size=1, stride=1, padding=0)
        for layer in [self.g, self.theta, self.phi, self.W]:
            c2_xavier_fill(layer)
        self.adain = AdaIN_block()

    def kernel_params(self):
        """The 8 tensors in the order the C ABI's frn_cab_params expects."""
        return (*_conv_params(self.g), *_conv_params(self.theta), *_conv_params(self.phi),
                *_conv_params(self.W))

    def forward(self, x):
        x0, x1 = x[0], x[1]
        assert len(x0.size()) == 4
        assert x0.size() == x1.size()
        return F.cross_attention(x0, x1, *self.kernel_params())


class REFusion(nn.Module):
    def __init__(self, in_planes, out_planes):
        super().__init__()
        # fusion.py:12-17 builds CrossAttentionBlock(in_planes) on conv0 outputs of out_planes
        # channels: the reference constructs with any pair but its forward only works (and is only
        # called, model.py:248-251) with in_planes == out_planes.  Same here: forward checks.
        self.in_planes, self.out_planes = in_planes, out_planes
        self.rgb_cross_attention = CrossAttentionBlock(in_planes)
        self.event_cross_attention = CrossAttentionBlock(in_planes)
        self.conv0_rgb = nn.Conv2d(in_planes, out_planes, kernel_size=1, stride=1, padding=0)
        self.conv0_evt = nn.Conv2d(out_planes, out_planes, kernel_size=1, stride=1, padding=0)

    def forward(self, rgb, evt):
        assert len(rgb.size()) == 4
        assert rgb.size() == evt.size()
        if self.in_planes != self.out_planes or rgb.size(1) != self.in_planes:
            raise RuntimeError("REFusion: expected in_planes == out_planes == input channels "
                               f"(got {self.in_planes}, {self.out_planes}, {rgb.size(1)})")
        conv0 = (*_conv_params(self.conv0_rgb), *_conv_params(self.conv0_evt))
        return F.refusion(rgb, evt, conv0, self.rgb_cross_attention.kernel_params(),
                          self.event_cross_attention.kernel_params())


class Conv1x1(nn.Conv2d):
    """nn.Conv2d(in_channels, out_channels, kernel_size=1, stride=1, padding=0) on the library's GEMM
    engine: the lateral convolutions P{5,4,3,2}_1 that consume the fused maps (model.py:28,32,36,40;
    SURVEY.md 8.2 row f1).  Same constructor arguments, parameters and state-dict keys as nn.Conv2d;
    anything but a dense 1x1 stride-1 unpadded convolution is refused."""

    def __init__(self, in_channels, out_channels, kernel_size=1, stride=1, padding=0, bias=True):
        as2 = lambda v: tuple(v) if isinstance(v, (tuple, list)) else (v, v)
        if as2(kernel_size) != (1, 1) or as2(stride) != (1, 1) or as2(padding) != (0, 0):
            raise ValueError("frn_b200.Conv1x1: kernel_size=1, stride=1, padding=0 only")
        super().__init__(in_channels, out_channels, kernel_size=1, stride=1, padding=0, bias=bias)

    def forward(self, x):
        assert len(x.size()) == 4
        return F.conv1x1(x, self.weight, self.bias)


class REFusionLateral(nn.Module):
    """One FPN level's REFusion together with the lateral 1x1 conv that consumes its output (model.py:305-308 followed
    by P{5,4,3,2}_1 at model.py:59-76): forward(rgb, evt) == lateral(fusion(rgb, evt)), computed
    without writing or re-reading the (B, 2C, H, W) map -- the AdaIN affine is folded into per-sample scaled lateral
    weights and the GEMM reads the gated maps.  `fusion` and `lateral` are the model's own modules (parameters and
    state-dict keys shared, nothing copied).  return_map=True also returns the 2C map (level 4: P6 reads C5 too,
    model.py:44)."""

    def __init__(self, fusion: REFusion, lateral: nn.Conv2d, return_map: bool = False):
        super().__init__()
        if lateral.kernel_size != (1, 1) or lateral.stride != (1, 1) or lateral.padding != (0, 0) or lateral.groups != 1:
            raise ValueError("REFusionLateral: the lateral must be a dense 1x1 stride-1 unpadded convolution")
        if lateral.in_channels != 2 * fusion.out_planes:
            raise ValueError("REFusionLateral: lateral.in_channels must be 2 * fusion.out_planes")
        self.fusion, self.lateral, self.return_map = fusion, lateral, return_map

    def forward(self, rgb, evt):
        assert len(rgb.size()) == 4
        assert rgb.size() == evt.size()
        f = self.fusion
        conv0 = (*_conv_params(f.conv0_rgb), *_conv_params(f.conv0_evt))
        return F.refusion_lateral(rgb, evt, self.lateral.weight, self.lateral.bias, conv0,
                                  f.rgb_cross_attention.kernel_params(), f.event_cross_attention.kernel_params(),
                                  self.return_map)


def convert_fpn_laterals(model: nn.Module):
    """Replace the lateral 1x1 convs (attributes P5_1, P4_1, P3_1, P2_1 of a PyramidFeatures-like module,
    model.py:28-40) by Conv1x1 sharing the same Parameter objects.  Returns the rebound names."""
    done = []
    for mod_name, m in model.named_modules():
        for attr in ("P5_1", "P4_1", "P3_1", "P2_1"):
            conv = getattr(m, attr, None)
            if (isinstance(conv, nn.Conv2d) and not isinstance(conv, Conv1x1) and conv.kernel_size == (1, 1)
                    and conv.stride == (1, 1) and conv.padding == (0, 0) and conv.groups == 1):
                new = Conv1x1(conv.in_channels, conv.out_channels, bias=conv.bias is not None)
                new.weight = conv.weight
                if conv.bias is not None:
                    new.bias = conv.bias
                setattr(m, attr, new)
                done.append(f"{mod_name}.{attr}" if mod_name else attr)
    return done


def fusion_parameters(model: nn.Module):
    """Parameters that belong to REFusion modules of `model` (the allreduce set)."""
    seen = []
    for m in model.modules():
        if isinstance(m, REFusion):
            seen.extend(m.parameters())
    return seen
