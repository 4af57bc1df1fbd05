"""Profiling driver: the tcgen05 GEMM on the block's dominant shapes (DSEC level 1, B = 8).
  conv   C (8, 256, 19200) = W (256, 256) @ x (8, 256, 19200)          retinanet/fusion.py:23
  wgrad  dW (256, 256) = sum_b dy_b (256, 19200) @ x_b^T (19200, 256)
  proj   V (8, 32, 19200) = Wg (32, 256) @ x                            retinanet/Attention.py:24

Instrumented builds (frn_b200/csrc/Makefile: GEMM_KNOBS / GEMM_TS / GEMM_TRACE), loaded with FRN_B200_LIB=...:
  SWEEP=1      the v2 kernel with parts switched off (FRN_GEMM_DBG bits), 20 back-to-back launches per setting
  TIMELINE=n   clock64 stamps of the first n stage hand-overs of CTA 0
"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from test_gemm_gpu import run_gemm, DEV

B, C_, N, d = 8, 256, 19200, 32
C = C_
g = torch.Generator().manual_seed(0)
W = torch.randn(1, C, C, generator=g).to(DEV)
Wg = torch.randn(1, d, C, generator=g).to(DEV)
x = torch.randn(B, C, N, generator=g).to(DEV)
dy = torch.randn(B, C, N, generator=g).to(DEV)
out = torch.empty(B, C, N, device=DEV)
outp = torch.empty(B, d, N, device=DEV)
dW = torch.empty(C, C, device=DEV)


def timeit(name, fn, flops, bytes_):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"{name}: {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s  {bytes_ / ms / 1e6:.0f} GB/s (algorithmic)")


eng = int(os.environ.get("ENGINE", "1"))
timeit("conv ", lambda: run_gemm(eng, W, x, out, C, N, C, B, (C, 1, 0), (N, 1, C * N), (N, 1, C * N)),
       2.0 * C * C * N * B, 2.0 * 4 * B * C * N)
timeit("wgrad", lambda: run_gemm(eng, dy, x, dW, C, C, N, B, (N, 1, C * N), (1, N, C * N), (C, 1, 0), reduce_batch=1),
       2.0 * C * C * N * B, 2.0 * 4 * B * C * N)
timeit("proj ", lambda: run_gemm(eng, Wg, x, outp, d, N, C, B, (C, 1, 0), (N, 1, C * N), (N, 1, d * N)),
       2.0 * d * C * N * B, 4.0 * B * (C + d) * N)

# K = d = 32 shapes of the backward pass (HBM-bound by arithmetic: one (B, C, N) tensor written or updated)
Wb = torch.randn(B, C_, d, generator=g).to(DEV)            # per-sample scaled W (style path)
Yc = torch.randn(B, d, N, generator=g).to(DEV)
dVp = torch.randn(B, N, d, generator=g).to(DEV)            # position-major projection gradient
timeit("dT   ", lambda: run_gemm(eng, Wb, Yc, out, C_, N, d, B, (d, 1, C_ * d), (N, 1, d * N), (N, 1, C_ * N)),
       2.0 * C_ * d * N * B, 4.0 * B * (C_ + d) * N)
timeit("dgrad", lambda: run_gemm(eng, Wg, dVp, out, C_, N, d, B, (1, C_, 0), (1, d, N * d), (N, 1, C_ * N), accumulate=1),
       2.0 * C_ * d * N * B, 4.0 * B * (2 * C_ + d) * N)
timeit("pwgrd", lambda: run_gemm(eng, dVp, x, outp[0], d, C_, N, B, (1, d, N * d), (1, N, C_ * N), (C_, 1, 0), reduce_batch=1),
       2.0 * C_ * d * N * B, 4.0 * B * (C_ + d) * N)

# role cycle counters (library built with -DFRN_GEMM_TRACE only)
import ctypes as C
from frn_b200 import _lib
lib = _lib.load()
if hasattr(lib, "frn_debug_gemm_trace"):
    names = ["ld cp.async wait", "ld bar", "ld wait empty", "ld convert", "mma wait acc_free", "mma wait full",
             "mma issue", "epi wait acc_full", "epi store", "cvt A", "cvt B", "fence+arrive", "issue"]
    if os.environ.get("FRN_GEMM_V2", "1") != "0":      # v2 kernel slots (CTA 0; split / epi slots are summed over 8 warps)
        names = ["tma wait empty", "split wait raw (x8)", "split work (x8)", "mma wait lo_full", "mma wait acc_free",
                 "mma wait raw_full", "mma issue", "epi wait acc_full (x8)", "epi store (x8)"]
    buf = (C.c_ulonglong * 16)()
    if lib.frn_debug_gemm_trace(buf):
        for nm, fn in (("conv", lambda: run_gemm(eng, W, x, out, C_, N, C_, B, (C_, 1, 0), (N, 1, C_ * N), (N, 1, C_ * N))),
                       ("dT", lambda: run_gemm(eng, Wb, Yc, out, C_, N, d, B, (d, 1, C_ * d), (N, 1, d * N), (N, 1, C_ * N))),
                       ("dgrad", lambda: run_gemm(eng, Wg, dVp, out, C_, N, d, B, (1, C_, 0), (1, d, N * d), (N, 1, C_ * N), accumulate=1))):
            lib.frn_debug_gemm_trace(buf)
            t0 = time.time(); fn(); dt = time.time() - t0
            lib.frn_debug_gemm_trace(buf)
            print(nm, f"{dt * 1e3:.3f} ms;", ", ".join(f"{n}: {buf[i] / 1e3:.0f}k" for i, n in enumerate(names)))
            if buf[14]:
                print(f"    CTA 0: {buf[13] / 1e3:.0f}k cycles in {buf[14] / 1e3:.1f} us = {buf[13] / buf[14]:.3f} GHz")

# timing experiments of a -DFRN_GEMM_KNOBS build (FRN_GEMM_DBG bits, no trace atomics; the results are wrong on
# purpose, only the time counts): what does the v2 kernel's time consist of?  20 back-to-back launches between events.
if os.environ.get("SWEEP"):
    x2 = torch.randn(B, 512, 4800, generator=g).to(DEV); W2 = torch.randn(1, 512, 512, generator=g).to(DEV)
    o2 = torch.empty(B, 512, 4800, device=DEV)
    x4 = torch.randn(B, 2048, 300, generator=g).to(DEV); W4 = torch.randn(1, 2048, 2048, generator=g).to(DEV)
    o4 = torch.empty(B, 2048, 300, device=DEV)
    R = 20
    kw = dict(repeat=R, sync=False)
    shapes = (("conv L1", lambda: run_gemm(eng, W, x, out, C_, N, C_, B, (C_, 1, 0), (N, 1, C_ * N), (N, 1, C_ * N), **kw)),
              ("conv L2", lambda: run_gemm(eng, W2, x2, o2, 512, 4800, 512, B, (512, 1, 0), (4800, 1, 512 * 4800), (4800, 1, 512 * 4800), **kw)),
              ("conv L4", lambda: run_gemm(eng, W4, x4, o4, 2048, 300, 2048, B, (2048, 1, 0), (300, 1, 2048 * 300), (300, 1, 2048 * 300), **kw)),
              ("wgrad L1", lambda: run_gemm(eng, dy, x, dW, C_, C_, N, B, (N, 1, C_ * N), (1, N, C_ * N), (C_, 1, 0), reduce_batch=1, **kw)))
    knobs = ((0, "as shipped (4 products)"), (2, "no lo.lo (3 products)"), (4, "[hh|hl] only (2 products, 1 instruction)"),
             (1, "no split work"), (16, "no epilogue stores"), (128, "no accumulator drain"), (1 | 16, "no split, no epilogue"),
             (1 | 4 | 16, "no split, [hh|hl] only, no epilogue"), (1 | 8 | 16, "TMA + barriers only (no MMA, no split, no epilogue)"),
             (1 | 8 | 16 | 128, "TMA + barriers, no accumulator drain either"),
             (1 | 8 | 16 | 32, "TMA of B only + barriers"), (1 | 8 | 16 | 64, "TMA of A only + barriers"),
             (1 | 8 | 16 | 32 | 64, "barriers only (no TMA at all)"), (8 | 16, "TMA + split (no MMA, no epilogue)"),
             (1 | 16 | 32 | 64, "MMA only (no TMA, no split, no epilogue)"), (16 | 32 | 64, "MMA + split (no TMA, no epilogue)"),
             (256, "no generic->async proxy fence in the split warps"), (1 | 8 | 16 | 32 | 64 | 256, "barriers only, no fence"),
             (1 | 256, "no split work, no fence"))
    for nm, fn in shapes:
        print(f"--- {nm}")
        for bits, what in knobs:
            os.environ["FRN_GEMM_DBG"] = str(bits)
            fn(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record(); torch.cuda.synchronize()
            print(f"    dbg={bits:3d} {e0.elapsed_time(e1) / R * 1e3:7.1f} us   {what}")
    os.environ["FRN_GEMM_DBG"] = "0"

# stage hand-over timeline of CTA 0 (a -DFRN_GEMM_TS -DFRN_GEMM_KNOBS build): one line per stage, cycles relative
# to the producer's first stamp
if hasattr(lib, "frn_debug_gemm_ts") and os.environ.get("TIMELINE"):
    ts = (C.c_longlong * (8 * 256))()
    conv = lambda: run_gemm(eng, W, x, out, C_, N, C_, B, (C_, 1, 0), (N, 1, C_ * N), (N, 1, C_ * N))
    for bits in (0, 1 | 8 | 16 | 32 | 64):
        os.environ["FRN_GEMM_DBG"] = str(bits)
        conv(); conv()
        if not lib.frn_debug_gemm_ts(ts):
            break
        t0 = ts[0]
        print(f"--- conv L1 timeline, dbg={bits}: stage | producer woke, TMA issued | split woke, arrived | MMA woke, left the issue block | MMAs issued, commits issued")
        for it in range(int(os.environ.get("TIMELINE"))):
            print(f"  {it:3d} | " + " ".join(f"{ts[r * 256 + it] - t0:7d}" for r in range(8)))
    os.environ["FRN_GEMM_DBG"] = "0"
