"""Experiment: error of the tcgen05 3xTF32 GEMM vs contraction length (accumulator truncation)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from test_gemm_gpu import run_gemm, DEV
for kind in ("randn", "relu"):
    for K in (32, 64, 128, 256, 512, 1024, 2048, 4096, 16384):
        g = torch.Generator().manual_seed(K)
        A = torch.randn(1, 256, K, generator=g).to(DEV); B = torch.randn(1, K, 256, generator=g).to(DEV)
        if kind == "relu": A, B = A.relu(), B.relu()
        ref = A.double() @ B.double()
        row = [kind, K]
        for eng in (0, 1):
            out = torch.empty(1, 256, 256, device=DEV)
            run_gemm(eng, A, B, out, 256, 256, K, 1, (K, 1, 0), (256, 1, 0), (256, 1, 0))
            err = (out.double() - ref)
            row += [float(err.norm() / ref.norm()), float((err / ref.abs().clamp_min(1e-3)).mean())]
        print("%s K=%5d  ffma rel %.2e bias %+.2e | tc rel %.2e bias %+.2e" % tuple(row))
