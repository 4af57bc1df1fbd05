"""frn_adain_backward_static (include/frn_b200.h): the AdaIN backward with its scratch in a file-scope __device__
array, for callers that cannot allocate.  Collected LAST on purpose.

GPU: two calls in a row through the C ABI -- the second at another shape, over the first call's leftovers in the
shared array -- against frn_adain_backward with a caller workspace (same kernels: equal within RTOL_F32, in fact
bit-identical), and the oversize rejection.  CPU: the rejection and the symbol, without touching a GPU."""
import pytest
import torch

from frn_b200 import _lib
from helpers import RTOL_F32, rel_l2


def _ptr(t):
    return t.data_ptr()


def _run_both(lib, b, c, n, ns, seed):
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(b, c, n, generator=g).relu().to(dev)
    t = (0.5 + torch.randn(b, c, ns, generator=g)).to(dev)
    dout = torch.randn(b, c, n, generator=g).to(dev)
    out = torch.empty_like(x)
    stats = torch.empty(4 * b * c, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    assert lib.frn_adain_forward(st, b, c, n, ns, _ptr(x), _ptr(t), _ptr(out), _ptr(stats)) == 0
    # existing entry point: workspace from the caller
    ws_bytes = lib.frn_adain_workspace_bytes(b, c, n)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    dx_ref, dt_ref = torch.full_like(x, float("nan")), torch.full_like(t, float("nan"))
    assert lib.frn_adain_backward(st, b, c, n, ns, _ptr(x), _ptr(t), _ptr(stats), _ptr(dout), _ptr(dx_ref),
                                  _ptr(dt_ref), _ptr(ws), ws_bytes) == 0
    # new entry point: every buffer exists before the call, nothing is allocated during it
    dx, dt = torch.full_like(x, float("nan")), torch.full_like(t, float("nan"))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.frn_adain_backward_static(st, b, c, n, ns, _ptr(x), _ptr(t), _ptr(stats), _ptr(dout), _ptr(dx), _ptr(dt))
    assert rc == 0, lib.frn_error_string(rc)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert torch.isfinite(dx).all() and torch.isfinite(dt).all()
    assert rel_l2(dx.cpu().numpy(), dx_ref.cpu().numpy()) <= RTOL_F32
    assert rel_l2(dt.cpu().numpy(), dt_ref.cpu().numpy()) <= RTOL_F32
    assert torch.equal(dx, dx_ref) and torch.equal(dt, dt_ref)      # same three kernels, same launch shapes


@pytest.mark.gpu
def test_static_scratch_matches_workspace_entry_point_twice_in_a_row():
    lib = _lib.load()
    # first call: 4096 rows, vector path; leaves its g1 / g2 in the shared array
    _run_both(lib, 8, 512, 1200, 1200, seed=11)
    # second call straight after: fewer rows, ragged N (scalar path), another style size -- stale rows above B*C
    # and stale values below it must not leak into the result
    _run_both(lib, 3, 40, 301, 77, seed=12)
    # and the first shape again, the DSEC level-4 shape at batch 8 (16384 rows)
    _run_both(lib, 8, 2048, 300, 300, seed=13)


@pytest.mark.gpu
def test_static_scratch_at_its_maximum():
    lib = _lib.load()
    rows = lib.frn_adain_static_max_rows()
    _run_both(lib, rows // 2048, 2048, 20, 12, seed=14)          # exactly FRN_ADAIN_STATIC_MAX_ROWS rows


def test_oversize_is_rejected_before_any_cuda_call():
    lib = _lib.load()
    rows = lib.frn_adain_static_max_rows()
    assert rows == 65536
    one = 16                                                      # any non-null pointer: rejected before use
    assert lib.frn_adain_backward_static(None, 33, 2048, 300, 300, one, one, one, one, one, one) == -6
    assert lib.frn_adain_backward_static(None, 1, rows + 1, 4, 4, one, one, one, one, one, one) == -6
    assert b"FRN_ADAIN_STATIC_MAX_ROWS" in lib.frn_error_string(-6)
    # the ordinary argument checks still come first
    assert lib.frn_adain_backward_static(None, 1, 8, 1, 4, one, one, one, one, one, one) == -1
    assert lib.frn_adain_backward_static(None, 1, 8, 4, 4, None, one, one, one, one, one) == -2
