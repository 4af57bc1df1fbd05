"""GPU: inference post-processing kernels through frn_b200.detect_postprocess / nms /
decode_boxes and the C ABI.  Bars: decoded boxes within 1e-6 of the reference's (expf differs from the CPU exp by an
ulp at most); the NMS keep list (index work) identical to torchvision.ops.nms / the oracle given the same boxes."""
import glob
import os

import numpy as np
import pytest
import torch

import frn_b200
from helpers import GOLDEN
from oracle import detection_numpy as odet
from oracle import postprocess_numpy as opp

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
FILES = sorted(glob.glob(os.path.join(GOLDEN, "postprocess_*.npz")))


def _gpu(z_or_arrays):
    return {k: torch.tensor(np.asarray(v), device=DEV) for k, v in z_or_arrays.items()}


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p)[:-4] for p in FILES])
def test_fixture(path):
    z = np.load(path)
    h, w = (int(v) for v in z["image_hw"])
    t = _gpu({k: z[k] for k in ("cls", "reg", "anchors", "decoded")})
    img = torch.zeros(1, 3, h, w, device=DEV)
    dec = frn_b200.decode_boxes(t["anchors"], t["reg"][None], img)
    assert dec.shape == (1, z["anchors"].shape[1], 4)
    assert np.abs(dec[0].cpu().numpy() - z["decoded"]).max() <= 1e-6 * np.abs(z["decoded"]).max()
    s, c, b = frn_b200.detect_postprocess(t["cls"][None], t["reg"][None], t["anchors"], img)
    assert c.dtype == torch.int64 and s.shape == c.shape and b.shape == (s.shape[0], 4)
    assert np.array_equal(c.cpu().numpy(), z["classes"])
    assert np.array_equal(s.cpu().numpy(), z["scores"])                      # same anchors kept, same order
    assert np.abs(b.cpu().numpy() - z["boxes"]).max() <= 1e-6 * max(np.abs(z["boxes"]).max(), 1)
    # NMS alone on the reference's boxes: bit-identical keep list per class
    for k in range(z["cls"].shape[1]):
        cand = np.nonzero(z["cls"][:, k] > 0.05)[0]
        if cand.size == 0:
            continue
        keep = frn_b200.nms(t["decoded"][cand], t["cls"][cand, k], 0.5).cpu().numpy()
        assert np.array_equal(cand[keep], z["index"][z["classes"] == k])


@pytest.mark.parametrize("hw,K,hot", [((480, 640), 3, 0.002), ((260, 346), 2, 0.07), ((128, 160), 20, 0.01)],
                         ids=["dsec_k3", "ddd17_k2_many_chunks", "k20"])
def test_full_size_vs_oracle(hw, K, hot):
    anchors = odet.anchors_for_image(*hw)
    A = anchors.shape[1]
    cls, reg = opp.synthetic_heads(A, K, seed=A % 13, hot=hot)
    t = _gpu({"cls": cls, "reg": reg, "anchors": anchors})
    img = torch.zeros(1, 3, *hw, device=DEV)
    dec = frn_b200.decode_boxes(t["anchors"], t["reg"][None], img)[0]
    s, c, b = frn_b200.detect_postprocess(t["cls"][None], t["reg"][None], t["anchors"], img)
    ws, wc, wb, wi = opp.postprocess(cls, reg, anchors, *hw, boxes=dec.cpu().numpy())   # oracle NMS on the same boxes
    assert (cls > 0.05).sum() > 1024                                       # more than one chunk somewhere
    assert np.array_equal(c.cpu().numpy(), wc) and np.array_equal(s.cpu().numpy(), ws)
    assert np.array_equal(b.cpu().numpy(), wb)
    s2, c2, b2 = frn_b200.detect_postprocess(t["cls"][None], t["reg"][None], t["anchors"], img)
    assert torch.equal(s, s2) and torch.equal(c, c2) and torch.equal(b, b2)           # bit-reproducible


def test_nothing_above_threshold_and_nms_dropin():
    anchors = odet.anchors_for_image(64, 64)
    A = anchors.shape[1]
    t = _gpu({"cls": np.full((1, A, 3), 0.01, np.float32), "reg": np.zeros((1, A, 4), np.float32), "anchors": anchors})
    s, c, b = frn_b200.detect_postprocess(t["cls"], t["reg"], t["anchors"], torch.zeros(1, 3, 64, 64, device=DEV))
    assert s.numel() == 0 and c.numel() == 0 and b.shape == (0, 4)
    rng = np.random.default_rng(1)
    xy = rng.uniform(0, 300, (5000, 2)).astype(np.float32)
    boxes = np.concatenate([xy, xy + rng.uniform(5, 60, (5000, 2)).astype(np.float32)], axis=1)
    scores = rng.standard_normal(5000).astype(np.float32)                   # negative scores are fine for nms()
    keep = frn_b200.nms(torch.tensor(boxes, device=DEV), torch.tensor(scores, device=DEV), 0.3)
    assert keep.dtype == torch.int64 and np.array_equal(keep.cpu().numpy(), opp.nms(boxes, scores, 0.3))
    assert frn_b200.nms(torch.zeros(0, 4, device=DEV), torch.zeros(0, device=DEV), 0.5).numel() == 0
