"""CPU: the C-ABI library loads, exports every symbol include/frn_b200.h declares, its size queries
behave, and the product path refuses CPU tensors (no fallback).  No compute calls."""
import os
import re

import pytest
import torch

from conftest import ROOT
from frn_b200 import _lib


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "frn_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(frn_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert sorted(_lib.EXPORTS) == declared_symbols()


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert lib.frn_abi_version() == 2
    inv = lib.frn_kernel_inventory()
    assert b"attn_fwd_tc_kernel" in inv and b"tc_gemm2_kernel" in inv and b"gate_stats_bulk_kernel" in inv


def test_size_queries():
    lib = _lib.load()
    # bad shapes -> 0
    assert lib.frn_refusion_saved_bytes(1, 30, 100) == 0          # C % 8 != 0
    assert lib.frn_refusion_workspace_bytes(0, 32, 100) == 0
    assert lib.frn_cab_saved_bytes(1, 32, 1) == 0                 # N < 2: variance undefined
    # DSEC level shapes at batch 8: monotone in B, and bounded (no N x N buffer
    # at level 1: 19200^2 * 4 B = 1.47 GB per sample would show up immediately)
    for c, n in ((256, 19200), (512, 4800), (1024, 1200), (2048, 300)):
        s1, s8 = lib.frn_refusion_saved_bytes(1, c, n), lib.frn_refusion_saved_bytes(8, c, n)
        w1, w8 = lib.frn_refusion_workspace_bytes(1, c, n), lib.frn_refusion_workspace_bytes(8, c, n)
        assert 0 < s1 < s8 and 0 < w1 <= w8
        feature_map = 8 * c * n * 4
        assert s8 < 6 * feature_map
        assert w8 < 8 * feature_map + (600 << 20)
    assert lib.frn_error_string(-3) and lib.frn_error_string(0) == b"ok"


def test_error_codes_without_touching_the_gpu():
    lib = _lib.load()
    # shape / null checks happen before any CUDA call
    assert lib.frn_refusion_forward(None, 1, 30, 10, None, None, None, None, None, 0, None, 0) == -1
    assert lib.frn_refusion_forward(None, 1, 32, 10, None, None, None, None, None, 0, None, 0) == -2
    assert lib.frn_mean_std(None, 1, 4, 1, None, 1e-5, None, None) == -1


def test_cpu_tensors_are_rejected():
    import frn_b200
    m = frn_b200.REFusion(16, 16)
    x = torch.randn(1, 16, 4, 4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(x, x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        frn_b200.AdaIN_block()(x, x)
    with pytest.raises(AssertionError):
        frn_b200.AdaIN_block().calc_mean_std(x[0])           # AdaIN.py:10
    with pytest.raises(AssertionError):
        frn_b200.AdaIN_block()(x, torch.randn(1, 8, 4, 4))   # AdaIN.py:20


def test_state_dict_keys_match_reference():
    import frn_b200
    from helpers import golden_files, load_golden
    _, params, _ = load_golden(golden_files("refusion_relu_c32")[0])
    m = frn_b200.REFusion(32, 32)
    assert list(m.state_dict().keys()) == list(params.keys())     # same names AND registration order
    for k, v in m.state_dict().items():
        assert tuple(v.shape) == params[k].shape, k
    # c2_xavier_fill zeroes the attention biases (Attention.py:16-17)
    assert float(m.rgb_cross_attention.g.bias.abs().max()) == 0.0
