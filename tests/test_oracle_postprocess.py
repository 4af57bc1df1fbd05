"""CPU: oracle/postprocess_numpy.py (box decoding + per-class NMS) against the fixtures made
with the reference's BBoxTransform / ClipBoxes and torchvision.ops.nms (oracle/make_golden_postprocess.py)."""
import glob
import json
import os

import numpy as np
import pytest

from helpers import GOLDEN
from oracle import postprocess_numpy as opp

FILES = sorted(glob.glob(os.path.join(GOLDEN, "postprocess_*.npz")))


def test_fixtures_present():
    assert len(FILES) == 3
    rep = json.load(open(os.path.join(GOLDEN, "postprocess_pin_report.json")))
    assert all(v["kept_identical"] for v in rep.values()) and rep["pin_260x346"]["candidates"] > 50000


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p)[:-4] for p in FILES])
def test_oracle_matches_reference_fixture(path):
    z = np.load(path)
    h, w = (int(v) for v in z["image_hw"])
    dec = opp.decode_clip(z["anchors"], z["reg"], h, w)
    assert np.abs(dec - z["decoded"]).max() <= 1e-6 * np.abs(z["decoded"]).max()     # exp differs by an ulp at most
    s, c, b, i = opp.postprocess(z["cls"], z["reg"], z["anchors"], h, w, boxes=z["decoded"])
    assert np.array_equal(i, z["index"]) and np.array_equal(c, z["classes"])          # index work: exact
    assert np.array_equal(s, z["scores"]) and np.array_equal(b, z["boxes"])


def test_nms_properties():
    rng = np.random.default_rng(0)
    xy = rng.uniform(0, 100, (400, 2)).astype(np.float32)
    boxes = np.concatenate([xy, xy + rng.uniform(5, 40, (400, 2)).astype(np.float32)], axis=1)
    scores = rng.random(400).astype(np.float32)
    keep = opp.nms(boxes, scores, 0.5)
    assert np.all(np.diff(scores[keep]) <= 0)                       # by decreasing score
    assert np.array_equal(opp.nms(boxes[keep], scores[keep], 0.5), np.arange(len(keep)))      # idempotent
    assert len(opp.nms(boxes, scores, 1.0)) == 400 and keep[0] == np.argmax(scores)


def test_host_side_checks():
    import torch
    import frn_b200
    from frn_b200 import _lib
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        frn_b200.nms(torch.zeros(3, 4), torch.zeros(3), 0.5)
    lib = _lib.load()
    assert lib.frn_detect_workspace_bytes(0, 3) == 0
    assert lib.frn_detect_workspace_bytes(230220, 3) > 230220 * 3 * (4 + 4 + 4 + 4 + 16)
    args = [None] * 4 + [480, 640, 0.05, 0.5, 10] + [None] * 6 + [0]
    assert lib.frn_detect_postprocess(None, 0, 3, *args) == -1
    assert lib.frn_detect_postprocess(None, 10, 3, *args) == -2
    assert lib.frn_decode_boxes(None, 10, None, None, 480, 640, None) == -2
