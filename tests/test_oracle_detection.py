"""CPU: oracle/detection_numpy.py (anchors + focal loss) against the golden fixtures that
oracle/make_golden_detection.py produced from the UNMODIFIED reference (retinanet/anchors.py, retinanet/losses.py),
plus the host-side drop-in surface (constructor defaults, CPU tensors rejected, C-ABI argument checks)."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN
from oracle import detection_numpy as odet

FILES = sorted(glob.glob(os.path.join(GOLDEN, "detection_*.npz")))


def test_fixtures_present():
    assert len(FILES) == 3
    rep = json.load(open(os.path.join(GOLDEN, "detection_pin_report.json")))
    assert rep["anchors_480x640"] == {"count": 230220, "bit_identical": True}      # ~230 k
    assert all(v.get("assignment_identical", True) for v in rep.values())


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p)[:-4] for p in FILES])
def test_oracle_matches_reference_fixture(path):
    z = np.load(path)
    h, w = (int(v) for v in z["image_hw"])
    assert np.array_equal(odet.anchors_for_image(h, w), z["anchors"])               # bit-exact
    r = odet.focal_loss(z["cls"], z["reg"], z["anchors"], z["annotations"], upstream=tuple(z["upstream"]))
    assert np.array_equal(r["assignment"], z["assignment"])                         # index work: exact
    assert abs(r["cls_loss"] - float(z["cls_loss"])) <= 1e-13 * abs(float(z["cls_loss"]))
    assert abs(r["reg_loss"] - float(z["reg_loss"])) <= 1e-13 * max(abs(float(z["reg_loss"])), 1e-30)
    assert odet.rel_l2(r["dcls"], z["dcls"]) <= 1e-13
    assert odet.rel_l2(r["dreg"], z["dreg"]) <= 1e-13


def test_anchor_counts_and_layout():
    a = odet.anchors_for_image(480, 640)
    assert a.shape == (1, 230220, 4) and a.dtype == np.float32
    # first cell of level 2: centre (2, 2), nine boxes, ratio-major
    first = a[0, :9]
    assert np.allclose((first[:, 0] + first[:, 2]) / 2, 2.0) and np.allclose((first[:, 1] + first[:, 3]) / 2, 2.0)
    areas = (first[:, 2] - first[:, 0]) * (first[:, 3] - first[:, 1])
    assert np.allclose(areas[[0, 3, 6]], 16.0 ** 2, rtol=1e-6)


def test_gradient_of_oracle_by_finite_differences():
    z = np.load(FILES[0])
    cls, reg = z["cls"].astype(np.float64)[:1, :400], z["reg"].astype(np.float64)[:1, :400]
    anc, ann = z["anchors"][:, :400], z["annotations"][:1]
    up = (0.7, 1.3)
    base = odet.focal_loss(cls, reg, anc, ann, upstream=up)
    f = lambda c, r: (lambda o: up[0] * o["cls_loss"] + up[1] * o["reg_loss"])(odet.focal_loss(c, r, anc, ann))
    rng = np.random.default_rng(0)
    for _ in range(10):
        i = tuple(rng.integers(0, s) for s in cls.shape)
        if not (1e-3 < cls[i] < 1 - 1e-3):
            continue
        d = np.zeros_like(cls); d[i] = 1e-6
        assert abs((f(cls + d, reg) - f(cls - d, reg)) / 2e-6 - base["dcls"][i]) <= 1e-6 * max(1, abs(base["dcls"][i]))
    pos = np.argwhere(base["assignment"][0] >= 0)[:, 0]
    for a in pos[:5]:
        d = np.zeros_like(reg); d[0, a, 2] = 1e-6
        assert abs((f(cls, reg + d) - f(cls, reg - d)) / 2e-6 - base["dreg"][0, a, 2]) <= 1e-6


def test_dropin_surface_on_cpu():
    import frn_b200
    from frn_b200 import _lib
    a = frn_b200.Anchors()
    assert a.pyramid_levels == [2, 3, 4, 5, 6] and a.strides == [4, 8, 16, 32, 64] and a.sizes == [16, 32, 64, 128, 256]
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        a(torch.zeros(1, 3, 32, 32))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        frn_b200.FocalLoss()(torch.rand(1, 9, 2), torch.rand(1, 9, 4), torch.rand(1, 9, 4), torch.rand(1, 1, 5))
    lib = _lib.load()
    import ctypes as C
    lv = (C.c_int * 5)(2, 3, 4, 5, 6)
    assert lib.frn_anchors_count(480, 640, 5, lv, 9) == 230220
    assert lib.frn_anchors_count(260, 346, 5, lv, 9) == 68490
    assert lib.frn_focal_loss_saved_bytes(8, 230220) >= 8 * 230220 * 4 + 8 * 24
    assert lib.frn_focal_loss_forward(None, 0, 10, 3, 1, None, None, None, None, 0.25, None, None, 0, None, 0) == -1
    assert lib.frn_focal_loss_forward(None, 1, 10, 3, 1, None, None, None, None, 0.25, None, None, 0, None, 0) == -2
    assert lib.frn_focal_loss_forward(None, 1, 10, 3, 5000, None, None, None, None, 0.25, None, None, 0, None, 0) == -4


def test_patch_reference_rebinds_a_package(tmp_path, monkeypatch):
    """patch_reference on a stand-in package with the reference's module / attribute names (model.py:7,10,221,227)."""
    import sys
    import frn_b200
    pkg = tmp_path / "standin_retinanet"
    pkg.mkdir()
    (pkg / "__init__.py").write_text("")
    (pkg / "AdaIN.py").write_text("class AdaIN_block: pass\n")
    (pkg / "Attention.py").write_text("from .AdaIN import AdaIN_block\nclass CrossAttentionBlock: pass\n")
    (pkg / "fusion.py").write_text("from .Attention import CrossAttentionBlock\nclass REFusion: pass\n")
    (pkg / "anchors.py").write_text("class Anchors: pass\n")
    (pkg / "losses.py").write_text("class FocalLoss: pass\n")
    (pkg / "model.py").write_text("from .fusion import REFusion\nfrom .anchors import Anchors\nfrom . import losses\n")
    monkeypatch.syspath_prepend(str(tmp_path))
    done = frn_b200.patch_reference("standin_retinanet")
    m = sys.modules["standin_retinanet.model"]
    assert m.REFusion is frn_b200.REFusion and m.Anchors is frn_b200.Anchors
    assert m.losses.FocalLoss is frn_b200.FocalLoss
    assert sys.modules["standin_retinanet.Attention"].AdaIN_block is frn_b200.AdaIN_block
    assert ("standin_retinanet.losses", "FocalLoss") in done and ("standin_retinanet.model", "Anchors") in done
