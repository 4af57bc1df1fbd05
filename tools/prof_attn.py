"""Profiling driver: one attention forward + backward on the tcgen05 engine at a DSEC level shape
(default level 1: B=8, d=32, N=19200).  Usage: python tools/prof_attn.py [B d N]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from test_attn_gpu import attn_bwd, attn_fwd, make

B, d, N = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (8, 32, 19200)
Qc, Kc, Vc, dY = make(B, d, N, seed=1, scale=0.3)
for it in range(2):
    torch.cuda.synchronize()
    t0 = time.time()
    Y, lse = attn_fwd(1, Qc, Kc, Vc)
    t1 = time.time()
    g = attn_bwd(1, Qc, Kc, Vc, Y, lse, dY)
    t2 = time.time()
    print(f"B={B} d={d} N={N}: fwd {1e3 * (t1 - t0):.2f} ms, bwd {1e3 * (t2 - t1):.2f} ms (host-timed, incl. pack)")

# role cycle counters of the dK/dV kernel (library built with -DFRN_ATTN_TRACE only); CTA (0, 0), lane 0 of each warp
import ctypes as C
from frn_b200 import _lib
lib = _lib.load()
if hasattr(lib, "frn_debug_attn_trace"):
    buf = (C.c_ulonglong * 16)()
    if lib.frn_debug_attn_trace(buf):
        lib.frn_debug_attn_trace(buf)
        g = attn_bwd(1, Qc, Kc, Vc, Y, lse, dY)
        lib.frn_debug_attn_trace(buf)
        names = ["mma: wait q_full", "mma: wait s_free", "mma: issue S/dP", "mma: wait ds_full", "mma: issue dV/dK",
                 "cmp: wait sdp_full (8 warps)", "cmp: ld + math (8)", "cmp: wait pv_done (8)", "cmp: st + arrive (8)"]
        steps = (N + 31) // 32
        nz = (d + 31) // 32
        for i, nm in enumerate(names):
            div = 8 if nm.startswith("cmp") else 1
            print(f"  {nm:32s} {buf[i] / steps / div / nz:8.0f} cycles / step / warp")
