"""The program compute-sanitizer wraps (one tool per gpurun call, B200_PROFILING.md):

    compute-sanitizer --tool memcheck  --error-exitcode 1 python tools/sanitize_run.py
    compute-sanitizer --tool racecheck --error-exitcode 1 python tools/sanitize_run.py small

Runs, in one process: the smoke shape (d = 8, ragged tiles), one DSEC level-1 sample (every d <= 32 tcgen05 attention
kernel, the v2 GEMM in all its modes, the bulk-copy gate_stats), one level-2 sample (the d = 64 kernels), one DDD17
level-3 sample (d = 128: materialised attention on the GEMMs, odd N: cp.async GEMM path), the fused lateral and the
bf16 inference path.  `small` keeps only shapes whose kernels finish in seconds under racecheck."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import frn_b200

small = len(sys.argv) > 1 and sys.argv[1] == "small"
dev = torch.device("cuda:0")
torch.manual_seed(0)
shapes = [(2, 64, 12, 10), (1, 256, 24, 32), (1, 512, 12, 16)] if small else \
    [(2, 64, 12, 10), (1, 256, 120, 160), (1, 512, 60, 80), (2, 1024, 17, 22)]
for b, c, h, w in shapes:
    m = frn_b200.REFusion(c, c).to(dev)
    a = torch.randn(b, c, h, w, device=dev).relu().requires_grad_(True)
    e = torch.randn(b, c, h, w, device=dev).relu().requires_grad_(True)
    out = m(a, e)
    out.backward(torch.randn_like(out))
    torch.cuda.synchronize()
    assert torch.isfinite(out).all() and torch.isfinite(a.grad).all()
    print(f"ok fwd+bwd {b}x{c}x{h}x{w}", flush=True)
b, c, h, w = (1, 256, 24, 32) if small else (1, 256, 60, 80)
fus = frn_b200.REFusion(c, c).to(dev)
lat = frn_b200.REFusionLateral(fus, torch.nn.Conv2d(2 * c, 256, 1).to(dev))
a = torch.randn(b, c, h, w, device=dev).relu().requires_grad_(True)
e = torch.randn(b, c, h, w, device=dev).relu().requires_grad_(True)
y = lat(a, e)
y.backward(torch.randn_like(y))
torch.cuda.synchronize()
print("ok fused lateral", flush=True)
with torch.no_grad():
    for hh, ww in ((17, 22), (24, 32)):
        o = fus(torch.randn(2, c, hh, ww, device=dev).bfloat16(), torch.randn(2, c, hh, ww, device=dev).bfloat16())
        torch.cuda.synchronize()
        assert o.dtype == torch.bfloat16 and torch.isfinite(o.float()).all()
print("ok bf16 inference", flush=True)
