"""GPU: out-of-bounds WRITE detection without compute-sanitizer (it is closed on this GPU pool,
profiles/r2_n_compute_sanitizer_closed.txt).  Every buffer the C ABI writes -- outputs, gradients, `saved`, the
workspace -- is carved out of a larger allocation with 64 KB guard bands on both sides filled with a canary pattern;
after REFusion forward + backward, the fused lateral and the bf16 inference call (through the C ABI, exact byte sizes
from the *_bytes queries) every canary must be intact and every output finite.  Shapes: ragged tiles (N % 128 != 0),
N % 4 != 0 (cp.async GEMM path, padded bf16 maps), d <= 32, d = 64 and d > 64."""
import ctypes as C

import pytest
import torch

from frn_b200 import _lib
from frn_b200 import functional as F
from helpers import make_params

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GUARD = 64 * 1024
CANARY = 0xA5


class Guarded:
    """nbytes of device memory between two guard bands."""

    def __init__(self, nbytes):
        self.n = int(nbytes)
        pad = (-self.n) % 256
        self.raw = torch.full((GUARD + self.n + pad + GUARD,), CANARY, dtype=torch.uint8, device=DEV)
        self.body = self.raw[GUARD:GUARD + self.n]
        self.tail_from = GUARD + self.n

    @property
    def ptr(self):
        return C.c_void_p(self.body.data_ptr())

    def intact(self):
        return bool((self.raw[:GUARD] == CANARY).all()) and bool((self.raw[self.tail_from:] == CANARY).all())

    def as_f32(self, *shape):
        return self.body.view(torch.float32).view(*shape)


SHAPES = [(2, 64, 12, 10), (1, 256, 13, 17), (2, 512, 9, 11), (2, 1024, 6, 7)]


@pytest.mark.parametrize("b,c,h,w", SHAPES, ids=str)
def test_no_write_outside_any_buffer(b, c, h, w):
    lib = _lib.load()
    n, cout = h * w, 64
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator().manual_seed(c)
    a = torch.randn(b, c, h, w, generator=g).relu().to(DEV)
    e = torch.randn(b, c, h, w, generator=g).relu().to(DEV)
    prm = [v.float().to(DEV).contiguous() for v in make_params(c, seed=1).values()]
    names = list(make_params(c, seed=1).keys())
    by = dict(zip(names, prm))
    conv0 = [by["conv0_rgb.weight"], by["conv0_rgb.bias"], by["conv0_evt.weight"], by["conv0_evt.bias"]]
    cab = lambda blk: [by[f"{blk}.{m}.{k}"] for m in ("g", "theta", "phi", "W") for k in ("weight", "bias")]
    ps = F._fusion_struct(conv0, cab("rgb_cross_attention"), cab("event_cross_attention"))
    f4 = 4
    out = Guarded(b * 2 * c * n * f4)
    saved = Guarded(lib.frn_refusion_saved_bytes(b, c, n))
    ws = Guarded(lib.frn_refusion_workspace_bytes(b, c, n))
    _lib.check(lib.frn_refusion_forward(st, b, c, n, C.c_void_p(a.data_ptr()), C.c_void_p(e.data_ptr()), C.byref(ps),
                                        out.ptr, saved.ptr, saved.n, ws.ptr, ws.n), "forward")
    dout = torch.randn(b, 2 * c, h, w, generator=g).to(DEV)
    da, de = Guarded(b * c * n * f4), Guarded(b * c * n * f4)
    grads = [Guarded(p.numel() * f4) for p in prm]
    gby = dict(zip(names, [C.c_void_p(x.body.data_ptr()).value for x in grads]))

    class Fake:
        def __init__(self, v): self.v = v
        def data_ptr(self): return self.v
    fk = lambda keys: [Fake(gby[k]) for k in keys]
    gs = F._fusion_struct(fk(["conv0_rgb.weight", "conv0_rgb.bias", "conv0_evt.weight", "conv0_evt.bias"]),
                          fk([f"rgb_cross_attention.{m}.{k}" for m in ("g", "theta", "phi", "W") for k in ("weight", "bias")]),
                          fk([f"event_cross_attention.{m}.{k}" for m in ("g", "theta", "phi", "W") for k in ("weight", "bias")]))
    _lib.check(lib.frn_refusion_backward(st, b, c, n, C.c_void_p(a.data_ptr()), C.c_void_p(e.data_ptr()), C.byref(ps),
                                         saved.ptr, C.c_void_p(dout.data_ptr()), da.ptr, de.ptr, C.byref(gs), ws.ptr, ws.n),
               "backward")
    # fused lateral (row f1)
    lw = (torch.randn(cout, 2 * c, generator=g) / (2 * c) ** 0.5).to(DEV)
    lb = torch.randn(cout, generator=g).to(DEV)
    lat, ws2 = Guarded(b * cout * n * f4), Guarded(lib.frn_refusion_lateral_workspace_bytes(b, c, n, cout))
    _lib.check(lib.frn_refusion_lateral_forward(st, b, c, n, cout, C.c_void_p(a.data_ptr()), C.c_void_p(e.data_ptr()),
                                                C.byref(ps), C.c_void_p(lw.data_ptr()), C.c_void_p(lb.data_ptr()), lat.ptr,
                                                None, saved.ptr, saved.n, ws2.ptr, ws2.n), "lateral forward")
    dlat = torch.randn(b, cout, h, w, generator=g).to(DEV)
    dlw, dlb = Guarded(cout * 2 * c * f4), Guarded(cout * f4)
    _lib.check(lib.frn_refusion_lateral_backward(st, b, c, n, cout, C.c_void_p(a.data_ptr()), C.c_void_p(e.data_ptr()),
                                                 C.byref(ps), C.c_void_p(lw.data_ptr()), saved.ptr,
                                                 C.c_void_p(dlat.data_ptr()), None, da.ptr, de.ptr, C.byref(gs), dlw.ptr,
                                                 dlb.ptr, ws2.ptr, ws2.n), "lateral backward")
    # bf16 inference (padded maps)
    ab, eb = a.bfloat16().contiguous(), e.bfloat16().contiguous()
    ob = Guarded(b * 2 * c * n * 2)
    ws3 = Guarded(lib.frn_refusion_infer_workspace_bytes(_lib.FRN_BF16, b, c, n))
    _lib.check(lib.frn_refusion_infer(st, _lib.FRN_BF16, b, c, n, C.c_void_p(ab.data_ptr()), C.c_void_p(eb.data_ptr()),
                                      C.byref(ps), ob.ptr, ws3.ptr, ws3.n), "infer bf16")
    torch.cuda.synchronize()
    bufs = {"out": out, "saved": saved, "ws": ws, "da": da, "de": de, "lat": lat, "ws_lateral": ws2, "dlat_w": dlw,
            "dlat_b": dlb, "out_bf16": ob, "ws_infer": ws3}
    bufs.update({f"grad[{k}]": x for k, x in zip(names, grads)})
    broken = [k for k, x in bufs.items() if not x.intact()]
    assert not broken, f"guard bands overwritten around: {broken}"
    for k in ("out", "da", "de", "lat", "dlat_w"):
        assert torch.isfinite(bufs[k].body.view(torch.float32)).all(), k
    assert torch.isfinite(ob.body.view(torch.bfloat16).float()).all()
