"""GPU: parity of the tcgen05 engines (attention core and GEMM), selected through the C ABI, on the same
gates as test_parity_gpu.py.  The default engines are exercised by test_parity_gpu.py."""
import os

import numpy as np
import pytest
import torch

import frn_b200
from frn_b200 import _lib
from helpers import (check_against, debug_engines_available, golden_files, load_golden, load_into, make_inputs,
                     make_params, run_module)
from oracle import fusion_numpy as onp

pytestmark = pytest.mark.gpu

ENGINES = [(1, 0), (0, 1), (1, 1)]          # (attention, gemm): 1 = tcgen05


@pytest.fixture(params=ENGINES, ids=["attn_tc", "gemm_tc", "both_tc"])
def engines(request):
    if request.param != (1, 1) and not debug_engines_available():
        pytest.skip("FFMA engines are compiled into -DFRN_DEBUG_ENGINES builds only (make DEBUG_ENGINES=1)")
    lib = _lib.load()
    old = lib.frn_get_attn_engine(), lib.frn_get_gemm_engine()
    lib.frn_set_attn_engine(request.param[0])
    lib.frn_set_gemm_engine(request.param[1])
    yield request.param
    lib.frn_set_attn_engine(old[0])
    lib.frn_set_gemm_engine(old[1])


@pytest.mark.parametrize("path", [p for p in golden_files("refusion_") if "nearconst" not in p],
                         ids=os.path.basename)
def test_golden_with_tc_engines(engines, path):
    d, params, grads = load_golden(path)
    c = d["a"].shape[1]
    m = frn_b200.REFusion(c, c)
    load_into(m, params)
    got = run_module(m, d["a"], d["e"], d["dout"])
    check_against(os.path.basename(path), got, (d["out"], d["da"], d["de"], grads))


LEVELS = [
    ("ddd17_L2", 2, 512, 33, 44, "relu"),
    ("ddd17_L1", 1, 256, 65, 87, "relu"),
    ("dsec_L3", 1, 1024, 30, 40, "relu"),
    ("dsec_L2", 1, 512, 60, 80, "randn"),
    ("dsec_L1", 1, 256, 120, 160, "relu"),
]


@pytest.mark.parametrize("name,b,c,h,w,kind", LEVELS, ids=[l[0] for l in LEVELS])
def test_level_shapes_with_tc_engines(engines, name, b, c, h, w, kind):
    if name == "dsec_L1" and engines != (1, 1):
        pytest.skip("the 10 s fp64 oracle at N = 19 200 runs once, with both engines on")
    a, e, dout = make_inputs(kind, (b, c, h, w), seed=1000 + c)
    a, e, dout = a.float().double(), e.float().double(), dout.float().double()
    prm = make_params(c, seed=0)
    npp = {k: v.numpy() for k, v in prm.items()}
    ref = onp.refusion_forward_backward(a.numpy(), e.numpy(), npp, dout.numpy())
    m = frn_b200.REFusion(c, c)
    load_into(m, npp)
    got = run_module(m, a.numpy(), e.numpy(), dout.numpy())
    errs = check_against(name, got, ref)
    print(name, engines, {k: f"{v:.2e}" for k, v in errs.items() if k in ("out", "da", "de")},
          "worst grad", f"{max(v for k, v in errs.items() if k not in ('out', 'da', 'de')):.2e}")
