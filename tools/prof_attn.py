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
