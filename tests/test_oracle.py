"""CPU: the oracle (oracle/fusion_numpy.py, oracle/fusion_torch.py) against the golden vectors that
oracle/make_golden.py produced by running the UNMODIFIED reference module in fp64."""
import json
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN, golden_files, load_golden, rel_l2
from oracle import fusion_numpy as onp
from oracle import fusion_torch as ot


def test_pin_report_is_clean():
    rep = json.load(open(os.path.join(GOLDEN, "pin_report.json")))
    assert rep["cases"], "empty pin report"
    for name, r in rep["cases"].items():
        for k, v in r.items():
            if k.startswith("torch_"):
                assert v == 0.0, (name, k, v)        # same ATen ops -> bit identical in fp64
            elif "grad_worst" in k:
                assert v < 1e-8, (name, k, v)
            else:
                assert v < 1e-11, (name, k, v)


@pytest.mark.parametrize("path", golden_files("refusion_"), ids=os.path.basename)
def test_numpy_oracle_matches_reference_refusion(path):
    d, params, grads = load_golden(path)
    out, da, de, g = onp.refusion_forward_backward(d["a"], d["e"], params, d["dout"])
    assert rel_l2(out, d["out"]) < 1e-12
    assert rel_l2(da, d["da"]) < 1e-11 and rel_l2(de, d["de"]) < 1e-11
    for k, ref in grads.items():
        if k.endswith("theta.bias"):
            assert np.abs(g[k]).max() < 1e-9 and np.abs(ref).max() < 1e-9
        else:
            assert rel_l2(g[k], ref) < 1e-8, k


@pytest.mark.parametrize("path", golden_files("refusion_"), ids=os.path.basename)
def test_torch_oracle_matches_reference_refusion(path):
    d, params, grads = load_golden(path)
    prm = {k: torch.from_numpy(v) for k, v in params.items()}
    out, da, de, g = ot.refusion_fwd_bwd(torch.from_numpy(d["a"]), torch.from_numpy(d["e"]), prm,
                                         torch.from_numpy(d["dout"]))
    assert np.array_equal(out.numpy(), d["out"])
    assert np.abs(da.numpy() - d["da"]).max() < 1e-12
    for k, ref in grads.items():
        assert np.abs(g[k].numpy() - ref).max() <= 1e-10 * max(1.0, np.abs(ref).max()), k


def test_adain_golden():
    d, _, _ = load_golden(golden_files("adain_")[0])
    out = onp.adain_forward(d["content"], d["style"])
    assert np.abs(out - d["out"]).max() < 1e-12
    mean, std = onp.calc_mean_std(d["content"])
    assert np.abs(mean - d["mean"]).max() < 1e-13 and np.abs(std - d["std"]).max() < 1e-13


def test_cab_golden_and_scramble():
    d, params, _ = load_golden(golden_files("cab_")[0])
    prm = {f"rgb_cross_attention.{k}": v for k, v in params.items()}
    out = onp.cross_attention_forward(d["x0"], d["x1"], prm)
    assert np.abs(out - d["out"]).max() < 1e-12
    # d does not divide N here (6 vs 28), a transpose instead of the view must differ
    b, c, h, w = d["x0"].shape
    assert (h * w) % (c // 8) != 0


def test_theta_bias_gradient_is_zero_in_reference():
    for path in golden_files("refusion_"):
        _, _, grads = load_golden(path)
        for k, v in grads.items():
            if k.endswith("theta.bias"):
                assert np.abs(v).max() < 1e-9
