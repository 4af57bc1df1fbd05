"""Experiment: is the tcgen05 accumulator truncation symmetric (toward zero) or a floor (toward -inf)?
Runs C1 = A.B and C2 = (-A).B on the tcgen05 GEMM engine; symmetric truncation gives C1 + C2 == 0 bitwise."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from test_gemm_gpu import run_gemm, DEV
for kind in ("randn", "relu"):
    for K in (32, 128, 1024, 8192):
        g = torch.Generator().manual_seed(K)
        A = torch.randn(1, 256, K, generator=g).to(DEV); B = torch.randn(1, K, 256, generator=g).to(DEV)
        if kind == "relu": A, B = A.relu(), B.relu()
        ref = A.double() @ B.double()
        o1 = torch.empty(1, 256, 256, device=DEV); o2 = torch.empty_like(o1)
        run_gemm(1, A, B, o1, 256, 256, K, 1, (K, 1, 0), (256, 1, 0), (256, 1, 0))
        run_gemm(1, (-A).contiguous(), B, o2, 256, 256, K, 1, (K, 1, 0), (256, 1, 0), (256, 1, 0))
        s = (o1 + o2).double()
        e1 = (o1.double() - ref); e2 = (o2.double() + ref)
        scale = ref.abs().mean()
        print(f"{kind} K={K:5d}: mean(C1+C2)/mean|C| {float(s.mean()/scale):+.2e}  nonzero frac {float((s != 0).float().mean()):.3f}"
              f" | mean err C1 {float(e1.mean()/scale):+.2e}  mean err C2 {float(e2.mean()/scale):+.2e}"
              f" | antisym part rel {float(((o1-o2).double()/2 - ref).norm()/ref.norm()):.2e}")
