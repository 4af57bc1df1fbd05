"""Per-shape GEMM times of one DSEC fusion step (debug hook frn_debug_gemm_log; FRN_GEMM_LOG=1).

    FRN_GEMM_LOG=1 python tools/gemm_shape_log.py [batch] [steps]

Prints, per FPN level, every distinct GEMM shape with its call count and time per step, largest first.
flags: 1 = reduce over batch (weight gradient), 2 = accumulate into C, 4 = A is k-contiguous, 8 = B is
n-contiguous."""
import ctypes as C
import os
import sys

os.environ.setdefault("FRN_GEMM_LOG", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import frn_b200
from frn_b200 import _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 3
LEVELS = [(256, 120, 160), (512, 60, 80), (1024, 30, 40), (2048, 15, 20)]
dev = torch.device("cuda:0")
lib = _lib.load()
lib.frn_debug_gemm_log.restype = C.c_long
lib.frn_debug_gemm_log.argtypes = [C.c_char_p, C.c_long]
buf = C.create_string_buffer(1 << 20)
torch.manual_seed(0)
grand = 0.0
for (c, h, w) in LEVELS:
    m = frn_b200.REFusion(c, c).to(dev)
    a = torch.randn(B, c, h, w, device=dev).relu().requires_grad_(True)
    e = torch.randn(B, c, h, w, device=dev).relu().requires_grad_(True)
    go = torch.randn(B, 2 * c, h, w, device=dev)
    for _ in range(2):
        m(a, e).backward(go)
    torch.cuda.synchronize()
    lib.frn_timing_enable(1)
    lib.frn_debug_gemm_log(buf, len(buf))
    for _ in range(STEPS):
        m(a, e).backward(go)
    torch.cuda.synchronize()
    lib.frn_debug_gemm_log(buf, len(buf))
    lib.frn_timing_enable(0)
    for cls in range(5):
        lib.frn_timing_read(cls, C.byref(C.c_double()), C.byref(C.c_long()))
    rows = []
    for line in buf.value.decode().strip().splitlines():
        key, n, ms = [x.strip() for x in line.split("|")]
        rows.append((float(ms) / STEPS, int(n) // STEPS, key))
    rows.sort(reverse=True)
    tot = sum(r[0] for r in rows)
    grand += tot
    print(f"--- level C={c} {h}x{w} (N={h * w}, d={c // 8}), B={B}: GEMM total {tot:.3f} ms/step")
    for ms, n, key in rows:
        print(f"  {ms:8.4f} ms  x{n:<3d} {ms / max(n, 1) * 1e3:8.1f} us/call  {key}")
print(f"=== all levels: {grand:.3f} ms/step")
