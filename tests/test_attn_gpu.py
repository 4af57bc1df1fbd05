"""GPU: the attention core (frn_attn_forward / frn_attn_backward, retinanet/Attention.py:29-32) on both
engines -- FFMA streaming kernels and the tcgen05 / TMEM kernels -- against an fp64 softmax(Q^T K) V
and its autograd gradients.  Logit scales follow the block: no 1/sqrt(d), |S| up to ~50.

Gates (l2-relative vs fp64): Y <= 2e-6; dQ, dK, dV <= 5e-6 (the block-level gate is 1e-5).  Also the
size-independent properties: rows of the implied softmax sum to one (Y of a constant V is that
constant), and the backward is deterministic on the tcgen05 engine (bit-identical reruns: no atomics).
"""
import ctypes as C

import pytest
import torch

from frn_b200 import _lib

from helpers import debug_engines_available

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def attn_fwd(engine, Qc, Kc, Vc):
    lib = _lib.load()
    B, d, N = Qc.shape
    Y = torch.empty(B, N, d, device=DEV)
    lse = torch.empty(B, N, device=DEV)
    nb = lib.frn_attn_workspace_bytes(B, N, d)
    ws = torch.empty(max(nb, 16), dtype=torch.uint8, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.frn_attn_forward(st, engine, B, N, d, _ptr(Qc), _ptr(Kc), _ptr(Vc), _ptr(Y), _ptr(lse), _ptr(ws), nb)
    torch.cuda.synchronize()
    assert rc == 0, lib.frn_error_string(rc)
    return Y, lse


def attn_bwd(engine, Qc, Kc, Vc, Y, lse, dY):
    lib = _lib.load()
    B, d, N = Qc.shape
    dQ, dK, dV = (torch.full((B, N, d), float("nan"), device=DEV) for _ in range(3))
    nb = lib.frn_attn_workspace_bytes(B, N, d)
    ws = torch.empty(max(nb, 16), dtype=torch.uint8, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.frn_attn_backward(st, engine, B, N, d, _ptr(Qc), _ptr(Kc), _ptr(Vc), _ptr(Y), _ptr(lse), _ptr(dY),
                               _ptr(dQ), _ptr(dK), _ptr(dV), _ptr(ws), nb)
    torch.cuda.synchronize()
    assert rc == 0, lib.frn_error_string(rc)
    return dQ, dK, dV


def reference(Qc, Kc, Vc, dY):
    """fp64 on the GPU: S[i][j] = sum_c Q[c][i] K[c][j]; Y = softmax_j(S) V^T; grads by autograd."""
    q = Qc.double().requires_grad_(True)
    k = Kc.double().requires_grad_(True)
    v = Vc.double().requires_grad_(True)
    S = q.transpose(1, 2) @ k
    Y = torch.softmax(S, dim=-1) @ v.transpose(1, 2)
    Y.backward(dY.double())
    # gradients in the library's (B, N, d) orientation
    return Y.detach(), q.grad.transpose(1, 2), k.grad.transpose(1, 2), v.grad.transpose(1, 2)


def rel(x, ref):
    return float((x.double() - ref).norm() / ref.norm())


def make(B, d, N, seed, scale):
    g = torch.Generator().manual_seed(seed)
    Qc = (torch.randn(B, d, N, generator=g) * scale).to(DEV)
    Kc = (torch.randn(B, d, N, generator=g) * scale).to(DEV)
    Vc = torch.randn(B, d, N, generator=g).to(DEV)
    dY = torch.randn(B, N, d, generator=g).to(DEV)
    return Qc, Kc, Vc, dY


CASES = [
    # name, B, d, N, logit scale (std of Q and K entries)
    ("d32_small", 2, 32, 200, 1.0),
    ("d32_ragged", 2, 32, 333, 1.5),
    ("d32_one_tile", 1, 32, 17, 1.0),
    ("d8_pad_channels", 2, 8, 120, 2.0),
    ("d2_tiny", 1, 2, 9, 1.0),
    ("d64_ragged", 2, 64, 301, 1.0),
    ("d48_pad", 1, 48, 257, 1.0),
    ("d32_ddd17_L1", 1, 32, 5655, 1.2),
    ("d64_ddd17_L2", 2, 64, 1452, 0.9),
    ("d32_multi_tile_pairs", 1, 32, 128 * 5 + 3, 1.0),
    ("d64_dsec_L2", 1, 64, 4800, 0.9),
    ("d33_first_odd_width_above_32", 2, 33, 100, 1.0),
    ("d32_two_key_tiles_one_query_step", 1, 32, 31, 1.0),
]


@pytest.mark.parametrize("engine", [0, 1], ids=["ffma", "tcgen05"])
@pytest.mark.parametrize("name,B,d,N,scale", CASES, ids=[c[0] for c in CASES])
def test_attention_core(engine, name, B, d, N, scale):
    if engine == 0 and not debug_engines_available():
        pytest.skip("FFMA engine: -DFRN_DEBUG_ENGINES builds only")
    Qc, Kc, Vc, dY = make(B, d, N, seed=N + d, scale=scale)
    rY, rdQ, rdK, rdV = reference(Qc, Kc, Vc, dY)
    Y, lse = attn_fwd(engine, Qc, Kc, Vc)
    Kd = Kc.double()
    if engine == 1:      # the tcgen05 engine centres the keys over positions (softmax is unchanged by it)
        Kd = Kd - Kd.mean(dim=2, keepdim=True)
    S = Qc.double().transpose(1, 2) @ Kd
    assert rel(Y, rY) <= 2e-6, rel(Y, rY)
    assert float((lse.double() - torch.logsumexp(S, dim=-1)).abs().max()) <= 2e-5
    dQ, dK, dV = attn_bwd(engine, Qc, Kc, Vc, Y, lse, dY)
    errs = {"dQ": rel(dQ, rdQ), "dK": rel(dK, rdK), "dV": rel(dV, rdV)}
    print(name, ["ffma", "tc"][engine], f"Y {rel(Y, rY):.2e}", {k: f"{v:.2e}" for k, v in errs.items()})
    assert all(v <= 5e-6 for v in errs.values()), errs


def test_rows_sum_to_one_and_determinism():
    Qc, Kc, Vc, dY = make(1, 32, 19200, seed=7, scale=1.0)      # DSEC level-1 size
    Vc.fill_(3.0)
    Y, lse = attn_fwd(1, Qc, Kc, Vc)
    assert float((Y - 3.0).abs().max()) <= 1e-5        # 19 200 fp32 terms per row
    Vc = torch.randn_like(Vc)
    Y, lse = attn_fwd(1, Qc, Kc, Vc)
    a = attn_bwd(1, Qc, Kc, Vc, Y, lse, dY)
    b = attn_bwd(1, Qc, Kc, Vc, Y, lse, dY)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    if not debug_engines_available():
        return
    # and the two engines agree at full size
    Yf, lsef = attn_fwd(0, Qc, Kc, Vc)            # each engine's backward takes its own forward's lse
    f = attn_bwd(0, Qc, Kc, Vc, Yf, lsef, dY)
    assert rel(Y, Yf.double()) <= 2e-6
    for x, y in zip(a, f):
        assert rel(x, y.double()) <= 5e-6
