"""GPU: bf16 inference path (BASELINE.json config 4: DDD17-shaped inference with bf16 fusion kernels;
frn_refusion_infer with FRN_BF16 through frn_b200.REFusion under no_grad).

Oracle: the fp64 restatement run on the SAME bf16-rounded inputs and fp32-representable parameters.
Stated tolerance (north_star: "bf16 stated separately"): l2-relative error of the bf16 output <= 2.5e-3
and <= 1.25 x the error of rounding the exact result to bf16 (measured on B200: 1.67e-3 ... 1.69e-3 against a
floor of 1.66e-3; the reference's ATen ops in bf16: 5.4e-3).
Rounding the exact result to bf16 alone gives ~1.7e-3 (8 significant bits); the rest is the single-pass
TF32 arithmetic (11 significant bits).  The reference's own ATen ops run in bf16 on the same inputs are the
yardstick on the small case: our error must not exceed theirs."""
import ctypes as C

import numpy as np
import pytest
import torch

import frn_b200
from frn_b200 import _lib
from frn_b200 import functional as F
from helpers import load_into, make_params, rel_l2
from oracle import fusion_numpy as onp
from oracle import fusion_torch as ot

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL_BF16 = 2.5e-3

SHAPES = [
    # DDD17 level shapes: ragged N (no 128-bit rows), all four d
    ("ddd17_L1", 2, 256, 65, 87),
    ("ddd17_L2", 2, 512, 33, 44),
    ("ddd17_L3", 2, 1024, 17, 22),
    ("ddd17_L4", 3, 2048, 9, 11),
    ("dsec_L2", 1, 512, 60, 80),
    ("small", 2, 64, 12, 10),
]


def _inputs(b, c, h, w, seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(b, c, h, w, generator=g).relu().bfloat16()
    e = torch.randn(b, c, h, w, generator=g).relu().bfloat16()
    return a, e


@pytest.mark.parametrize("name,b,c,h,w", SHAPES, ids=[s[0] for s in SHAPES])
def test_bf16_inference_matches_fp64_oracle(name, b, c, h, w):
    a, e = _inputs(b, c, h, w, seed=c + h)
    prm = make_params(c, seed=11)
    m = frn_b200.REFusion(c, c)
    load_into(m, {k: v.numpy() for k, v in prm.items()})
    m = m.to(DEV).eval()
    with torch.no_grad():
        out = m(a.to(DEV), e.to(DEV))
    torch.cuda.synchronize()
    assert out.dtype == torch.bfloat16 and out.shape == (b, 2 * c, h, w)
    ref, _ = onp.refusion_forward(a.double().numpy(), e.double().numpy(), {k: v.numpy() for k, v in prm.items()})
    err = rel_l2(out.float().cpu().numpy(), ref)
    floor = rel_l2(torch.tensor(ref).bfloat16().float().numpy(), ref)     # rounding the exact answer to bf16
    print(f"{name}: bf16 path {err:.2e}, bf16 rounding of the exact result {floor:.2e}")
    assert np.isfinite(out.float().cpu().numpy()).all()
    assert err <= TOL_BF16, (err, floor)
    assert err <= 1.25 * floor, (err, floor)


def test_not_worse_than_reference_ops_in_bf16():
    b, c, h, w = 2, 64, 12, 10
    a, e = _inputs(b, c, h, w, seed=3)
    prm = make_params(c, seed=5)
    m = frn_b200.REFusion(c, c)
    load_into(m, {k: v.numpy() for k, v in prm.items()})
    m = m.to(DEV)
    with torch.no_grad():
        out = m(a.to(DEV), e.to(DEV))
    ref, _ = onp.refusion_forward(a.double().numpy(), e.double().numpy(), {k: v.numpy() for k, v in prm.items()})
    with torch.no_grad():
        aten = ot.refusion(a, e, {k: v.bfloat16() for k, v in prm.items()})      # ATen ops in bf16 on the CPU
    ours, theirs = rel_l2(out.float().cpu().numpy(), ref), rel_l2(aten.float().numpy(), ref)
    print(f"bf16: ours {ours:.2e}, ATen bf16 {theirs:.2e}")
    assert ours <= theirs


def test_autocast_and_grad_dispatch():
    b, c, h, w = 2, 64, 6, 7
    g = torch.Generator().manual_seed(1)
    a = torch.randn(b, c, h, w, generator=g).relu().to(DEV)
    e = torch.randn(b, c, h, w, generator=g).relu().to(DEV)
    m = frn_b200.REFusion(c, c).to(DEV)
    with torch.no_grad():
        o32 = m(a, e)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            o16 = m(a, e)                                  # fp32 inputs under autocast -> bf16 path
    assert o32.dtype == torch.float32 and o16.dtype == torch.bfloat16
    assert rel_l2(o16.float().cpu().numpy(), o32.double().cpu().numpy()) <= 2 * TOL_BF16
    # gradients required -> the fp32-parity kernels, whatever the activation dtype
    ab = a.bfloat16().requires_grad_(True)
    out = m(ab, e.bfloat16())
    out.float().sum().backward()
    assert out.dtype == torch.bfloat16 and ab.grad is not None and ab.grad.dtype == torch.bfloat16


def test_infer_f32_is_the_training_forward():
    """frn_refusion_infer(FRN_F32) == frn_refusion_forward bit for bit (same kernels, no saved state)."""
    b, c, h, w = 2, 256, 9, 11
    n = h * w
    g = torch.Generator().manual_seed(2)
    a = torch.randn(b, c, h, w, generator=g).relu().to(DEV)
    e = torch.randn(b, c, h, w, generator=g).relu().to(DEV)
    m = frn_b200.REFusion(c, c).to(DEV)
    ar = a.clone().requires_grad_(True)
    ref = m(ar, e).detach()
    lib = _lib.load()
    prm = [p.detach().float().contiguous() for p in
           (m.conv0_rgb.weight, m.conv0_rgb.bias, m.conv0_evt.weight, m.conv0_evt.bias,
            *m.rgb_cross_attention.kernel_params(), *m.event_cross_attention.kernel_params())]
    st = F._fusion_struct(prm[0:4], prm[4:12], prm[12:20])
    out = torch.empty(b, 2 * c, h, w, device=DEV)
    nbytes = lib.frn_refusion_infer_workspace_bytes(_lib.FRN_F32, b, c, n)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
    rc = lib.frn_refusion_infer(C.c_void_p(torch.cuda.current_stream().cuda_stream), _lib.FRN_F32, b, c, n,
                                C.c_void_p(a.data_ptr()), C.c_void_p(e.data_ptr()), C.byref(st),
                                C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), nbytes)
    torch.cuda.synchronize()
    assert rc == 0
    assert torch.equal(out, ref)
    # error paths: unknown dtype, short workspace
    assert lib.frn_refusion_infer(None, 7, b, c, n, C.c_void_p(a.data_ptr()), C.c_void_p(e.data_ptr()), C.byref(st),
                                  C.c_void_p(out.data_ptr()), None, 0) == -4
    assert lib.frn_refusion_infer(None, _lib.FRN_BF16, b, c, n, C.c_void_p(a.data_ptr()), C.c_void_p(e.data_ptr()),
                                  C.byref(st), C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), 16) == -3


def test_config4_batch32_by_sample_decomposition():
    """BASELINE.json configs[3]: DDD17-shaped inference at batch 32 with bf16 kernels.  The B = 32 call must equal
    the 32 single-sample calls (no cross-sample coupling) and, for the first and last sample of every level, meet
    the stated bf16 tolerance against the fp64 oracle."""
    for name, c, h, w in (("ddd17_L1", 256, 65, 87), ("ddd17_L2", 512, 33, 44), ("ddd17_L3", 1024, 17, 22),
                          ("ddd17_L4", 2048, 9, 11)):
        b = 32
        a, e = _inputs(b, c, h, w, seed=4000 + c)
        prm = make_params(c, seed=11)
        m = frn_b200.REFusion(c, c)
        load_into(m, {k: v.numpy() for k, v in prm.items()})
        m = m.to(DEV).eval()
        with torch.no_grad():
            full = m(a.to(DEV), e.to(DEV))
            assert full.dtype == torch.bfloat16 and full.shape == (b, 2 * c, h, w)
            for i in range(b):
                one = m(a[i:i + 1].to(DEV), e[i:i + 1].to(DEV))
                # identical arithmetic per sample; allow one bf16 ulp (2^-8) on isolated elements
                assert rel_l2(full[i:i + 1].float().cpu().numpy(), one.double().cpu().numpy()) <= 1e-4, (name, i)
        for i in (0, b - 1):
            ref, _ = onp.refusion_forward(a[i:i + 1].double().numpy(), e[i:i + 1].double().numpy(),
                                          {k: v.numpy() for k, v in prm.items()})
            err = rel_l2(full[i:i + 1].float().cpu().numpy(), ref)
            floor = rel_l2(torch.tensor(ref).bfloat16().float().numpy(), ref)
            assert err <= TOL_BF16 and err <= 1.25 * floor, (name, i, err, floor)
