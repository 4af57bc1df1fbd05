"""GPU: the strided batched GEMM (frn_gemm_f32) on both engines -- FFMA and tcgen05 3xTF32 --
against an fp64 matmul, over the operand orientations the fusion block uses (conv, dgrad, wgrad,
position-major projections) and ragged sizes.  Gate: l2-relative <= 2e-6 (both engines must sit well
inside the block-level 1e-5 budget)."""
import ctypes as C

import numpy as np
import pytest
import torch

from frn_b200 import _lib
from helpers import debug_engines_available

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def run_gemm(engine, A, B, Cout, M, N, K, batch, sA, sB, sC, bias_m=None, bias_n=None, accumulate=0,
             reduce_batch=0, repeat=1, sync=True):
    lib = _lib.load()
    d = _lib.GemmDesc()
    d.A, d.B, d.C = A.data_ptr(), B.data_ptr(), Cout.data_ptr()
    d.M, d.N, d.K, d.batch = M, N, K, batch
    d.sAm, d.sAk, d.sAb = sA
    d.sBk, d.sBn, d.sBb = sB
    d.sCm, d.sCn, d.sCb = sC
    d.bias_m = bias_m.data_ptr() if bias_m is not None else None
    d.bias_n = bias_n.data_ptr() if bias_n is not None else None
    d.accumulate, d.reduce_batch = accumulate, reduce_batch
    nbytes = lib.frn_gemm_workspace_bytes(C.byref(d))
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=DEV)
    for _ in range(repeat):
        rc = lib.frn_gemm_f32(C.c_void_p(torch.cuda.current_stream().cuda_stream), engine, C.byref(d),
                              C.c_void_p(ws.data_ptr()), nbytes)
    if sync:
        torch.cuda.synchronize()
    assert rc == 0, lib.frn_error_string(rc)


def rel(x, ref):
    x, ref = x.double().cpu(), ref.double().cpu()
    return float((x - ref).norm() / ref.norm())


CASES = [
    # name, batch, M, N, K, A layout, B layout, C layout  ("km" = k-major i.e. k contiguous)
    ("conv_L2", 2, 512, 1000, 512, "mk", "kn", "mn"),
    ("conv_ragged", 3, 200, 333, 77, "mk", "kn", "mn"),
    ("proj_small_M", 2, 32, 1452, 256, "mk", "kn", "mn"),
    ("proj_posmajor_out", 2, 64, 600, 512, "mk", "kn", "nm"),
    ("dgrad", 2, 256, 777, 32, "km", "nk", "mn"),
    ("scores", 2, 300, 300, 256, "km", "kn", "mn"),
    ("tiny", 1, 5, 7, 3, "mk", "kn", "mn"),
    ("k_not_mult_4", 2, 130, 129, 35, "mk", "nk", "mn"),
    # TMA-able strides (multiples of 4 floats): the v2 kernel, all four operand-major combinations, ragged M / N / K
    # edges (zero fill), every tile width (BN = 128 / 64 / 32) and accumulation chains longer than one chunk
    ("v2_kA_mnB_conv", 2, 256, 2400, 256, "mk", "kn", "mn"),
    ("v2_mnA_mnB", 2, 200, 480, 96, "km", "kn", "mn"),
    ("v2_kA_kB_ragged", 2, 384, 300, 200, "mk", "nk", "mn"),
    ("v2_mnA_kB", 3, 132, 260, 72, "km", "nk", "mn"),
    ("v2_bn64", 2, 512, 64, 256, "mk", "kn", "mn"),
    ("v2_bn32_posmajor", 2, 320, 32, 128, "km", "nk", "nm"),
    ("v2_long_k", 1, 128, 128, 4096, "mk", "nk", "mn"),
    ("v2_swapped_small_M", 2, 96, 1200, 512, "mk", "kn", "mn"),
    # persistent CTAs walking MANY tiles (more tiles than SMs) with 1, 3 and 5 stages per tile: the two MMA-issuing
    # warps own stages by the parity of the CTA's global stage counter, so with an odd stage count per tile the owner of
    # a tile's first stage alternates, chains of both issuers end inside one tile, and every ring slot changes hands
    ("v2_many_tiles_k32", 16, 128, 4096, 32, "mk", "kn", "mn"),
    ("v2_many_tiles_k96", 8, 256, 5120, 96, "mk", "kn", "mn"),
    ("v2_many_tiles_k160_kB", 4, 384, 3000, 160, "mk", "nk", "mn"),
    ("v2_many_tiles_bn32", 24, 1024, 32, 96, "mk", "kn", "mn"),
]


def make(layout, rows, cols, batch, g):
    """Tensor addressed as [b][r][c] with the given memory order; returns (tensor, (s_r, s_c, s_b), view)."""
    if layout[0] in "mk" and layout in ("mk", "kn", "mn"):      # row-major in the logical order
        t = torch.randn(batch, rows, cols, generator=g).to(DEV)
        return t, (cols, 1, rows * cols), t
    t = torch.randn(batch, cols, rows, generator=g).to(DEV)     # transposed storage
    return t, (1, rows, rows * cols), t.transpose(1, 2)


@pytest.mark.parametrize("engine", [0, 1], ids=["ffma", "tcgen05"])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gemm_orientations(engine, case):
    if engine == 0 and not debug_engines_available():
        pytest.skip("FFMA engine: -DFRN_DEBUG_ENGINES builds only")
    name, batch, M, N, K, la, lb, lc = case
    g = torch.Generator().manual_seed(hash(name) % 1000)
    A, sA, Av = make("mk" if la == "mk" else "T", M, K, batch, g)
    B, sB, Bv = make("kn" if lb == "kn" else "T", K, N, batch, g)
    bias_m = torch.randn(M, generator=g).to(DEV)
    bias_n = torch.randn(N, generator=g).to(DEV)
    ref = Av.double() @ Bv.double() + bias_m.double()[None, :, None] + bias_n.double()[None, None, :]
    if lc == "mn":
        Cout = torch.full((batch, M, N), 7.0, device=DEV)
        sC, Cv = (N, 1, M * N), Cout
    else:
        Cout = torch.full((batch, N, M), 7.0, device=DEV)
        sC, Cv = (1, M, M * N), Cout.transpose(1, 2)
    run_gemm(engine, A, B, Cout, M, N, K, batch, sA, sB, sC, bias_m, bias_n)
    assert rel(Cv, ref) < 2e-6, rel(Cv, ref)
    # accumulate into C
    run_gemm(engine, A, B, Cout, M, N, K, batch, sA, sB, sC, None, None, accumulate=1)
    ref2 = ref + Av.double() @ Bv.double()
    assert rel(Cv, ref2) < 2e-6


@pytest.mark.parametrize("engine", [0, 1], ids=["ffma", "tcgen05"])
@pytest.mark.parametrize("shape", [(8, 256, 256, 19200), (4, 32, 256, 5655), (2, 2048, 2048, 300), (3, 6, 48, 28)],
                         ids=str)
def test_gemm_wgrad_reduce_batch(engine, shape):
    """dW (M, K') = sum_b dy_b (M, n) x_b^T (n, K'): contraction over (batch, n), deterministic split-K."""
    if engine == 0 and not debug_engines_available():
        pytest.skip("FFMA engine: -DFRN_DEBUG_ENGINES builds only")
    batch, M, Kp, n = shape
    g = torch.Generator().manual_seed(n)
    dy = torch.randn(batch, M, n, generator=g).to(DEV)
    x = torch.randn(batch, Kp, n, generator=g).to(DEV)
    ref = torch.einsum("bmn,bkn->mk", dy.double(), x.double())
    out = torch.zeros(M, Kp, device=DEV)
    run_gemm(engine, dy, x, out, M, Kp, n, batch, (n, 1, M * n), (1, n, Kp * n), (Kp, 1, 0), reduce_batch=1)
    assert rel(out, ref) < 2e-6, rel(out, ref)
    out2 = out.clone()
    run_gemm(engine, dy, x, out2, M, Kp, n, batch, (n, 1, M * n), (1, n, Kp * n), (Kp, 1, 0), reduce_batch=1)
    assert torch.equal(out, out2)        # deterministic


def test_tcgen05_is_not_plain_tf32():
    """Single-pass TF32 would show ~5e-4 here; the split-precision path must be ~1e-6."""
    g = torch.Generator().manual_seed(0)
    A = torch.randn(1, 256, 512, generator=g).to(DEV)
    B = torch.randn(1, 512, 384, generator=g).to(DEV)
    out = torch.empty(1, 256, 384, device=DEV)
    run_gemm(1, A, B, out, 256, 384, 512, 1, (512, 1, 0), (384, 1, 0), (384, 1, 0))
    e = rel(out, A.double() @ B.double())
    print("tcgen05 3xTF32 rel err", e)
    assert e < 2e-6
