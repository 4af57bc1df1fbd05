"""torch.autograd.Functions over the C ABI (libfrn_b200.so).

PyTorch is used for device memory (caching allocator), streams and autograd
bookkeeping only; every arithmetic op of the fusion block runs in the
hand-written sm_100a kernels.  CPU tensors are rejected: there is no fallback.

Reference semantics:
    adain / mean_std   retinanet/AdaIN.py:7-26
    cross_attention    retinanet/Attention.py:21-36
    refusion           retinanet/fusion.py:20-36
    conv1x1            nn.Conv2d(Cin, Cout, 1) of the FPN laterals, retinanet/model.py:28-40 (row f1)
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

CAB_PARAM_ORDER = ("g_w", "g_b", "theta_w", "theta_b", "phi_w", "phi_b", "W_w", "W_b")


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _autocast_dtype():
    """dtype of the active CUDA autocast region or None (torch >= 2.4 and the older torch.cuda.amp API)."""
    try:
        if not torch.is_autocast_enabled("cuda"):
            return None
        return torch.get_autocast_dtype("cuda")
    except TypeError:          # torch < 2.4: is_autocast_enabled() takes no argument
        if not torch.is_autocast_enabled():
            return None
        return torch.get_autocast_gpu_dtype()


def _wants_grad(*tensors):
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


# SURVEY.md 8.2 b: behave under autocast (train_dsec_det_fast.py:240-241, torch.cuda.amp.autocast() = fp16 + GradScaler):
# the Functions run with autocast disabled on fp32 copies of their inputs; the public wrappers cast the result to the
# autocast dtype, which is what the reference's ATen ops return under the same context.
try:
    _custom_fwd = torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    _custom_bwd = torch.amp.custom_bwd(device_type="cuda")
except (AttributeError, TypeError):      # torch < 2.4
    _custom_fwd = torch.cuda.amp.custom_fwd(cast_inputs=torch.float32)
    _custom_bwd = torch.cuda.amp.custom_bwd


def _to_autocast(out):
    ac = _autocast_dtype()
    return out if ac is None or out.dtype == ac else out.to(ac)


def _check_cuda(*tensors):
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError("frn_b200: CUDA tensors required (there is no CPU fallback)")


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _bytes(n, device):
    return torch.empty(max(int(n), 16), dtype=torch.uint8, device=device)


def _cab_struct(tensors):
    s = _lib.CabParams()
    for name, t in zip(CAB_PARAM_ORDER, tensors):
        setattr(s, name, t.data_ptr())
    return s


def _fusion_struct(conv0, cab_a, cab_b):
    s = _lib.RefusionParams()
    s.conv0_rgb_w, s.conv0_rgb_b, s.conv0_evt_w, s.conv0_evt_b = (t.data_ptr() for t in conv0)
    s.rgb_ca = _cab_struct(cab_a)
    s.event_ca = _cab_struct(cab_b)
    return s


# --------------------------------------------------------------------------- AdaIN
def mean_std(feat, eps=1e-5):
    """AdaIN_block.calc_mean_std (AdaIN.py:7-15); not differentiable through this entry."""
    _check_cuda(feat)
    lib = _lib.load()
    b, c = feat.shape[:2]
    x = _f32c(feat).view(b, c, -1)
    n = x.shape[2]
    mean = torch.empty(b * c, dtype=torch.float32, device=x.device)
    std = torch.empty_like(mean)
    with torch.cuda.device(x.device):
        _lib.check(lib.frn_mean_std(_stream(), b, c, n, _ptr(x), eps, _ptr(mean), _ptr(std)), "mean_std")
    return mean.view(b, c, 1, 1).to(feat.dtype), std.view(b, c, 1, 1).to(feat.dtype)


class _AdaINFn(torch.autograd.Function):
    @staticmethod
    @_custom_fwd
    def forward(ctx, content, style):
        _check_cuda(content, style)
        lib = _lib.load()
        b, c = content.shape[:2]
        x, t = _f32c(content), _f32c(style)
        # the style statistics run over the style map's own H*W (AdaIN.py:12-14,22): the reference only asserts
        # that the two maps agree in (N, C) (AdaIN.py:20)
        n, ns = x[0, 0].numel(), t[0, 0].numel()
        out = torch.empty_like(x)
        stats = torch.empty(4 * b * c, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(lib.frn_adain_forward(_stream(), b, c, n, ns, _ptr(x), _ptr(t), _ptr(out), _ptr(stats)),
                       "adain_forward")
        ctx.save_for_backward(x, t, stats)
        ctx.dims = (b, c, n, ns)
        ctx.in_dtypes = (content.dtype, style.dtype)
        return out.to(content.dtype)

    @staticmethod
    @_custom_bwd
    def backward(ctx, dout):
        lib = _lib.load()
        x, t, stats = ctx.saved_tensors
        b, c, n, ns = ctx.dims
        g = _f32c(dout)
        dx, dt = torch.empty_like(x), torch.empty_like(t)
        with torch.cuda.device(x.device):
            ws_bytes = lib.frn_adain_workspace_bytes(b, c, n)
            ws = _bytes(ws_bytes, x.device)
            _lib.check(lib.frn_adain_backward(_stream(), b, c, n, ns, _ptr(x), _ptr(t), _ptr(stats), _ptr(g),
                                              _ptr(dx), _ptr(dt), _ptr(ws), ws_bytes), "adain_backward")
        return dx.to(ctx.in_dtypes[0]), dt.to(ctx.in_dtypes[1])


def adain(content, style):
    return _to_autocast(_AdaINFn.apply(content, style))


# --------------------------------------------------------------------------- CrossAttentionBlock
class _CabFn(torch.autograd.Function):
    @staticmethod
    @_custom_fwd
    def forward(ctx, need_grad, x0, x1, *params):
        _check_cuda(x0, x1, *params)
        lib = _lib.load()
        b, c, h, w = x0.shape
        n = h * w
        a, e = _f32c(x0), _f32c(x1)
        prm = [_f32c(p) for p in params]
        out = torch.empty_like(a)
        with torch.cuda.device(a.device):
            ws_bytes = lib.frn_cab_workspace_bytes(b, c, n)
            ws = _bytes(ws_bytes, a.device)
            saved, saved_bytes = None, 0
            if need_grad:
                saved_bytes = lib.frn_cab_saved_bytes(b, c, n)
                saved = _bytes(saved_bytes, a.device)
            st = _cab_struct(prm)
            _lib.check(lib.frn_cab_forward(_stream(), b, c, n, _ptr(a), _ptr(e), C.byref(st), _ptr(out), c * n,
                                           _ptr(saved) if need_grad else None, saved_bytes, _ptr(ws), ws_bytes),
                       "cab_forward")
        if need_grad:
            ctx.save_for_backward(a, e, saved, *prm)
        ctx.dims = (b, c, n)
        ctx.in_dtypes = (x0.dtype, x1.dtype) + tuple(p.dtype for p in params)
        return out.to(x0.dtype)

    @staticmethod
    @_custom_bwd
    def backward(ctx, dout):
        lib = _lib.load()
        a, e, saved, *prm = ctx.saved_tensors
        b, c, n = ctx.dims
        g = _f32c(dout)
        da, de = torch.empty_like(a), torch.empty_like(e)
        grads = [torch.empty_like(p) for p in prm]
        with torch.cuda.device(a.device):
            ws_bytes = lib.frn_cab_workspace_bytes(b, c, n)
            ws = _bytes(ws_bytes, a.device)
            st, gst = _cab_struct(prm), _cab_struct(grads)
            _lib.check(lib.frn_cab_backward(_stream(), b, c, n, _ptr(a), _ptr(e), C.byref(st), _ptr(saved),
                                            _ptr(g), c * n, _ptr(da), _ptr(de), 0, C.byref(gst), _ptr(ws),
                                            ws_bytes), "cab_backward")
        dts = ctx.in_dtypes
        return (None, da.to(dts[0]), de.to(dts[1])) + tuple(gr.to(dt) for gr, dt in zip(grads, dts[2:]))


def cross_attention(x0, x1, g_w, g_b, theta_w, theta_b, phi_w, phi_b, W_w, W_b):
    params = (g_w, g_b, theta_w, theta_b, phi_w, phi_b, W_w, W_b)
    return _to_autocast(_CabFn.apply(_wants_grad(x0, x1, *params), x0, x1, *params))


# --------------------------------------------------------------------------- REFusion
class _REFusionFn(torch.autograd.Function):
    """params: conv0_rgb (w, b), conv0_evt (w, b), rgb_cross_attention (8), event_cross_attention (8)."""

    @staticmethod
    @_custom_fwd
    def forward(ctx, need_grad, rgb, evt, *params):
        _check_cuda(rgb, evt, *params)
        lib = _lib.load()
        b, c, h, w = rgb.shape
        n = h * w
        a, e = _f32c(rgb), _f32c(evt)
        prm = [_f32c(p) for p in params]
        out = torch.empty((b, 2 * c, h, w), dtype=torch.float32, device=a.device)
        with torch.cuda.device(a.device):
            ws_bytes = lib.frn_refusion_workspace_bytes(b, c, n)
            ws = _bytes(ws_bytes, a.device)
            saved, saved_bytes = None, 0
            if need_grad:
                saved_bytes = lib.frn_refusion_saved_bytes(b, c, n)
                saved = _bytes(saved_bytes, a.device)
            st = _fusion_struct(prm[0:4], prm[4:12], prm[12:20])
            _lib.check(lib.frn_refusion_forward(_stream(), b, c, n, _ptr(a), _ptr(e), C.byref(st), _ptr(out),
                                                _ptr(saved) if need_grad else None, saved_bytes, _ptr(ws),
                                                ws_bytes), "refusion_forward")
        if need_grad:
            ctx.save_for_backward(a, e, saved, *prm)
        ctx.dims = (b, c, n)
        ctx.in_dtypes = (rgb.dtype, evt.dtype) + tuple(p.dtype for p in params)
        return out.to(rgb.dtype)

    @staticmethod
    @_custom_bwd
    def backward(ctx, dout):
        lib = _lib.load()
        a, e, saved, *prm = ctx.saved_tensors
        b, c, n = ctx.dims
        g = _f32c(dout)
        da, de = torch.empty_like(a), torch.empty_like(e)
        grads = [torch.empty_like(p) for p in prm]
        with torch.cuda.device(a.device):
            ws_bytes = lib.frn_refusion_workspace_bytes(b, c, n)
            ws = _bytes(ws_bytes, a.device)
            st = _fusion_struct(prm[0:4], prm[4:12], prm[12:20])
            gst = _fusion_struct(grads[0:4], grads[4:12], grads[12:20])
            _lib.check(lib.frn_refusion_backward(_stream(), b, c, n, _ptr(a), _ptr(e), C.byref(st), _ptr(saved),
                                                 _ptr(g), _ptr(da), _ptr(de), C.byref(gst), _ptr(ws), ws_bytes),
                       "refusion_backward")
        dts = ctx.in_dtypes
        return (None, da.to(dts[0]), de.to(dts[1])) + tuple(gr.to(dt) for gr, dt in zip(grads, dts[2:]))


def refusion_infer_bf16(rgb, evt, conv0, rgb_ca, event_ca):
    """Inference with bf16 activations (BASELINE.json config 4 "bf16 fusion kernels": the module under
    torch.autocast(dtype=torch.bfloat16) + no_grad.  The reference's own trainer uses the fp16 default of
    torch.cuda.amp.autocast() with a GradScaler, train_dsec_det_fast.py:240-265; that mode takes the fp32-parity
    kernels through _REFusionFn, see refusion()): bf16 in, bf16 out, fp32 parameters,
    fp32 statistics / softmax / accumulation, single-pass TF32 products.  No autograd graph."""
    params = (*conv0, *rgb_ca, *event_ca)
    _check_cuda(rgb, evt, *params)
    lib = _lib.load()
    b, c, h, w = rgb.shape
    n = h * w
    a = rgb.detach().to(torch.bfloat16).contiguous()
    e = evt.detach().to(torch.bfloat16).contiguous()
    prm = [_f32c(p) for p in params]
    out = torch.empty((b, 2 * c, h, w), dtype=torch.bfloat16, device=a.device)
    with torch.cuda.device(a.device):
        ws_bytes = lib.frn_refusion_infer_workspace_bytes(_lib.FRN_BF16, b, c, n)
        ws = _bytes(ws_bytes, a.device)
        st = _fusion_struct(prm[0:4], prm[4:12], prm[12:20])
        _lib.check(lib.frn_refusion_infer(_stream(), _lib.FRN_BF16, b, c, n, _ptr(a), _ptr(e), C.byref(st),
                                          _ptr(out), _ptr(ws), ws_bytes), "refusion_infer")
    return out


def refusion(rgb, evt, conv0, rgb_ca, event_ca):
    """conv0: (rgb_w, rgb_b, evt_w, evt_b); rgb_ca / event_ca: 8 tensors in CAB_PARAM_ORDER.

    bf16 activations (bf16 inputs, or fp32 inputs under torch.autocast(bfloat16)) with no gradient required
    take the bf16 inference path; everything else runs the fp32-parity kernels."""
    params = (*conv0, *rgb_ca, *event_ca)
    wants_grad = _wants_grad(rgb, evt, *params)
    autocast_bf16 = _autocast_dtype() == torch.bfloat16
    if not wants_grad and (autocast_bf16 or (rgb.dtype == torch.bfloat16 and evt.dtype == torch.bfloat16)):
        return refusion_infer_bf16(rgb, evt, conv0, rgb_ca, event_ca)
    # fp32, and fp16 autocast + GradScaler (train_dsec_det_fast.py:240-265): fp32-parity kernels on fp32 copies
    return _to_autocast(_REFusionFn.apply(wants_grad, rgb, evt, *params))


# --------------------------------------------------------------------------- REFusion + lateral conv (row f1)
class _REFusionLateralFn(torch.autograd.Function):
    """P_1(REFusion(rgb, evt)) without writing the (B, 2C, H, W) map (include/frn_b200.h, frn_refusion_lateral_*).
    params: lateral (w, b), then the 20 REFusion tensors in _REFusionFn order.  return_map: also return the map."""

    @staticmethod
    @_custom_fwd
    def forward(ctx, return_map, rgb, evt, lat_w, lat_b, *params):
        _check_cuda(rgb, evt, lat_w, *params)
        lib = _lib.load()
        b, c, h, w = rgb.shape
        n, cout = h * w, lat_w.shape[0]
        a, e = _f32c(rgb), _f32c(evt)
        lw = _f32c(lat_w).view(cout, 2 * c)
        lb = _f32c(lat_b) if lat_b is not None else None
        prm = [_f32c(p) for p in params]
        lat = torch.empty((b, cout, h, w), dtype=torch.float32, device=a.device)
        out = torch.empty((b, 2 * c, h, w), dtype=torch.float32, device=a.device) if return_map else None
        with torch.cuda.device(a.device):
            ws_bytes = lib.frn_refusion_lateral_workspace_bytes(b, c, n, cout)
            ws = _bytes(ws_bytes, a.device)
            saved_bytes = lib.frn_refusion_saved_bytes(b, c, n)
            saved = _bytes(saved_bytes, a.device)
            st = _fusion_struct(prm[0:4], prm[4:12], prm[12:20])
            _lib.check(lib.frn_refusion_lateral_forward(
                _stream(), b, c, n, cout, _ptr(a), _ptr(e), C.byref(st), _ptr(lw), _ptr(lb) if lb is not None else None,
                _ptr(lat), _ptr(out) if out is not None else None, _ptr(saved), saved_bytes, _ptr(ws), ws_bytes),
                "refusion_lateral_forward")
        ctx.save_for_backward(a, e, saved, lw, *prm)
        ctx.dims = (b, c, n, cout)
        ctx.has_bias = lb is not None
        ctx.lat_w_shape = tuple(lat_w.shape)
        ctx.in_dtypes = (rgb.dtype, evt.dtype, lat_w.dtype, lat_b.dtype if lat_b is not None else None) + \
            tuple(p.dtype for p in params)
        if return_map:
            return lat.to(rgb.dtype), out.to(rgb.dtype)
        return lat.to(rgb.dtype)

    @staticmethod
    @_custom_bwd
    def backward(ctx, dlat, dout=None):
        lib = _lib.load()
        a, e, saved, lw, *prm = ctx.saved_tensors
        b, c, n, cout = ctx.dims
        g = _f32c(dlat)
        gx = _f32c(dout) if dout is not None else None
        da, de = torch.empty_like(a), torch.empty_like(e)
        grads = [torch.empty_like(p) for p in prm]
        dlw = torch.empty_like(lw)
        dlb = torch.empty(cout, dtype=torch.float32, device=a.device) if ctx.has_bias else None
        with torch.cuda.device(a.device):
            ws_bytes = lib.frn_refusion_lateral_workspace_bytes(b, c, n, cout)
            ws = _bytes(ws_bytes, a.device)
            st = _fusion_struct(prm[0:4], prm[4:12], prm[12:20])
            gst = _fusion_struct(grads[0:4], grads[4:12], grads[12:20])
            _lib.check(lib.frn_refusion_lateral_backward(
                _stream(), b, c, n, cout, _ptr(a), _ptr(e), C.byref(st), _ptr(lw), _ptr(saved), _ptr(g),
                _ptr(gx) if gx is not None else None, _ptr(da), _ptr(de), C.byref(gst), _ptr(dlw),
                _ptr(dlb) if dlb is not None else None, _ptr(ws), ws_bytes), "refusion_lateral_backward")
        dts = ctx.in_dtypes
        return (None, da.to(dts[0]), de.to(dts[1]), dlw.view(ctx.lat_w_shape).to(dts[2]),
                dlb.to(dts[3]) if dlb is not None else None) + tuple(gr.to(dt) for gr, dt in zip(grads, dts[4:]))


def refusion_lateral(rgb, evt, lat_w, lat_b, conv0, rgb_ca, event_ca, return_map=False):
    """P_1(REFusion(rgb, evt)) for the lateral conv P_1 = (lat_w (Cout, 2C, 1, 1), lat_b) of model.py:28-40 with the
    AdaIN affine folded into the lateral weights; return_map=True also returns the (B, 2C, H, W) map."""
    res = _REFusionLateralFn.apply(bool(return_map), rgb, evt, lat_w, lat_b, *conv0, *rgb_ca, *event_ca)
    if return_map:
        return _to_autocast(res[0]), _to_autocast(res[1])
    return _to_autocast(res)


# --------------------------------------------------------------------------- 1x1 conv (FPN laterals, row f1)
class _Conv1x1Fn(torch.autograd.Function):
    @staticmethod
    @_custom_fwd
    def forward(ctx, x, weight, bias):
        _check_cuda(x, weight) if bias is None else _check_cuda(x, weight, bias)
        lib = _lib.load()
        b, cin, h, w = x.shape
        cout = weight.shape[0]
        n = h * w
        xf, wf = _f32c(x), _f32c(weight)
        bf = _f32c(bias) if bias is not None else None
        y = torch.empty((b, cout, h, w), dtype=torch.float32, device=xf.device)
        with torch.cuda.device(xf.device):
            ws_bytes = lib.frn_conv1x1_workspace_bytes(b, cin, cout, n)
            ws = _bytes(ws_bytes, xf.device)
            _lib.check(lib.frn_conv1x1_forward(_stream(), b, cin, cout, n, _ptr(xf), _ptr(wf),
                                               _ptr(bf) if bf is not None else None, _ptr(y), _ptr(ws), ws_bytes),
                       "conv1x1_forward")
        ctx.save_for_backward(xf, wf)
        ctx.dims = (b, cin, cout, n)
        ctx.has_bias = bias is not None
        ctx.in_dtypes = (x.dtype, weight.dtype, bias.dtype if bias is not None else None)
        return y.to(x.dtype)

    @staticmethod
    @_custom_bwd
    def backward(ctx, dy):
        lib = _lib.load()
        xf, wf = ctx.saved_tensors
        b, cin, cout, n = ctx.dims
        g = _f32c(dy)
        dx = torch.empty_like(xf) if ctx.needs_input_grad[0] else None
        dw = torch.empty_like(wf)
        db = torch.empty(cout, dtype=torch.float32, device=xf.device) if ctx.has_bias else None
        with torch.cuda.device(xf.device):
            ws_bytes = lib.frn_conv1x1_workspace_bytes(b, cin, cout, n)
            ws = _bytes(ws_bytes, xf.device)
            _lib.check(lib.frn_conv1x1_backward(_stream(), b, cin, cout, n, _ptr(xf), _ptr(wf), _ptr(g),
                                                _ptr(dx) if dx is not None else None, _ptr(dw),
                                                _ptr(db) if db is not None else None, _ptr(ws), ws_bytes),
                       "conv1x1_backward")
        dts = ctx.in_dtypes
        return (dx.to(dts[0]) if dx is not None else None, dw.to(dts[1]),
                db.to(dts[2]) if db is not None else None)


def conv1x1(x, weight, bias=None):
    """F.conv2d(x, weight, bias) for a (Cout, Cin, 1, 1) weight, stride 1, no padding (model.py:28-40)."""
    return _to_autocast(_Conv1x1Fn.apply(x, weight, bias))
