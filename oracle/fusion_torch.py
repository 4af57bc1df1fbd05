"""CPU restatement of FRN's fusion block with the same ATen calls the reference makes.

TEST INFRASTRUCTURE ONLY (see oracle/fusion_numpy.py for the rules): used as
(a) the fp64 autograd cross-check of the NumPy oracle and (b) the timed CPU
baseline ("port") of bench.py, because /root/reference does not exist on the
GPU box.  It is written functionally (no nn.Module, parameters passed as a
dict keyed by the reference's state_dict names) but issues the same operator
sequence as the reference so the timing is the reference's CPU cost:

    conv2d 1x1 x2, mul, add, add                     retinanet/fusion.py:23-28
    conv2d 1x1 x3, view/permute, matmul, softmax,
      matmul, view (no permute back), conv2d 1x1     retinanet/Attention.py:24-34
    var(dim=2)+eps, sqrt, mean, sub, div, mul, add   retinanet/AdaIN.py:12-14,24-26
    cat(dim=1)                                       retinanet/fusion.py:34

Pinned against the unmodified reference module by oracle/make_golden.py
(bit-identical outputs in fp64; recorded in tests/golden/pin_report.json).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

EPS = 1e-5


def mean_std(feat, eps=EPS):
    """AdaIN.py:7-15."""
    assert feat.dim() == 4
    b, c = feat.shape[:2]
    var = feat.view(b, c, -1).var(dim=2) + eps
    return feat.view(b, c, -1).mean(dim=2).view(b, c, 1, 1), var.sqrt().view(b, c, 1, 1)


def adain(content, style):
    """AdaIN.py:17-26."""
    assert content.shape[:2] == style.shape[:2]
    size = content.size()
    s_mean, s_std = mean_std(style)
    c_mean, c_std = mean_std(content)
    normalized = (content - c_mean.expand(size)) / c_std.expand(size)
    return normalized * s_std.expand(size) + s_mean.expand(size)


def cross_attention(x0, x1, prm, blk):
    """Attention.py:21-36."""
    b, c, h, w = x0.shape
    g_x = F.conv2d(x0, prm[f"{blk}.g.weight"], prm[f"{blk}.g.bias"]).view(b, -1, h * w).permute(0, 2, 1)
    theta_x = F.conv2d(x1, prm[f"{blk}.theta.weight"], prm[f"{blk}.theta.bias"]).view(b, -1, h * w)
    phi_x = F.conv2d(x1, prm[f"{blk}.phi.weight"], prm[f"{blk}.phi.bias"]).view(b, -1, h * w).permute(0, 2, 1)
    f = torch.matmul(phi_x, theta_x)
    p = F.softmax(f, dim=-1)
    y = torch.matmul(p, g_x).view(b, c // 8, h, w)   # reinterpretation, Attention.py:32
    w_y = F.conv2d(y, prm[f"{blk}.W.weight"], prm[f"{blk}.W.bias"])
    return adain(x0, w_y)


def refusion(a, e, prm):
    """fusion.py:20-36.  prm: {reference state_dict key: tensor}."""
    r0 = F.conv2d(a, prm["conv0_rgb.weight"], prm["conv0_rgb.bias"])
    e0 = F.conv2d(e, prm["conv0_evt.weight"], prm["conv0_evt.bias"])
    m = r0.mul(e0)
    r1 = r0 + m
    e1 = e0 + m
    za = cross_attention(r1, e1, prm, "rgb_cross_attention")
    zb = cross_attention(e1, r1, prm, "event_cross_attention")
    return torch.cat([za, zb], dim=1)


def init_params(c, seed=0, dtype=torch.float32):
    """Parameters with the reference's shapes and init *distributions*.

    conv0_*: nn.Conv2d default (kaiming_uniform a=sqrt(5), bias U(+-1/sqrt(fan_in)),
    fusion.py:16-17); g/theta/phi/W: c2_xavier_fill = kaiming_uniform a=1, zero
    bias (Attention.py:16-17).  Not the reference's RNG stream: parity tests copy
    a state_dict instead of relying on seeds.
    """
    gen = torch.Generator().manual_seed(seed)
    d = c // 8

    def uni(shape, bound):
        return ((torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1) * bound).to(dtype)

    prm = {}
    for blk in ("rgb_cross_attention", "event_cross_attention"):
        for conv, (o, i) in (("g", (d, c)), ("theta", (d, c)), ("phi", (d, c)), ("W", (c, d))):
            prm[f"{blk}.{conv}.weight"] = uni((o, i, 1, 1), (3.0 / i) ** 0.5)
            prm[f"{blk}.{conv}.bias"] = torch.zeros(o, dtype=dtype)
    for conv in ("conv0_rgb", "conv0_evt"):
        prm[f"{conv}.weight"] = uni((c, c, 1, 1), (1.0 / c) ** 0.5)
        prm[f"{conv}.bias"] = uni((c,), (1.0 / c) ** 0.5)
    return prm


def refusion_fwd_bwd(a, e, prm, dout):
    """One fwd+bwd step; returns out, da, de, {key: grad}.  Used for timing and cross-checks."""
    a = a.detach().clone().requires_grad_(True)
    e = e.detach().clone().requires_grad_(True)
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in prm.items()}
    out = refusion(a, e, leaves)
    out.backward(dout)
    return out.detach(), a.grad, e.grad, {k: v.grad for k, v in leaves.items()}
