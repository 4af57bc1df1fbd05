"""Generate tests/golden/postprocess_*.npz with the reference's own classes.  Build container only:

    python oracle/make_golden_postprocess.py

TEST INFRASTRUCTURE ONLY.  Imports the UNMODIFIED `retinanet.utils.BBoxTransform`, `retinanet.utils.ClipBoxes` and
`retinanet.anchors.Anchors` from /root/reference and `torchvision.ops.nms` (the reference's dependency,
model.py:5), and calls them in the order of retinanet/model.py:326-364 (that block is the body of
`ResNet.forward`'s eval branch; its ten lines of glue -- threshold, mask, nms, cat -- are repeated here because the
branch cannot be called without the whole network)."""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch
from torchvision.ops import nms

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REFERENCE = os.environ.get("FRN_REFERENCE", "/root/reference")
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, REFERENCE)
sys.path.insert(0, ROOT)

from retinanet.anchors import Anchors  # noqa: E402  (reference, unmodified)
from retinanet.utils import BBoxTransform, ClipBoxes  # noqa: E402

from oracle import postprocess_numpy as opp  # noqa: E402


def reference_postprocess(classification, regression, anchors, img):
    boxes = ClipBoxes()(BBoxTransform()(anchors, regression), img)          # model.py:326-327
    scores_out, classes_out, boxes_out, index_out = [], [], [], []
    for i in range(classification.shape[2]):                                # model.py:340-362
        scores = torch.squeeze(classification[:, :, i])
        over = scores > 0.05
        if over.sum() == 0:
            continue
        idx = torch.nonzero(over)[:, 0]
        keep = nms(torch.squeeze(boxes)[over], scores[over], 0.5)
        scores_out.append(scores[over][keep]); classes_out.append(torch.tensor([i] * keep.shape[0]))
        boxes_out.append(torch.squeeze(boxes)[over][keep]); index_out.append(idx[keep])
    cat = lambda xs, e: torch.cat(xs) if xs else e
    return (boxes[0], cat(scores_out, torch.zeros(0)), cat(classes_out, torch.zeros(0, dtype=torch.long)),
            cat(boxes_out, torch.zeros(0, 4)), cat(index_out, torch.zeros(0, dtype=torch.long)))


def main():
    report = {}
    for name, (h, w, K, seed, hot) in {"postprocess_k3_64x96": (64, 96, 3, 1, 0.07),
                                       "postprocess_k2_dense_48x48": (48, 48, 2, 2, 0.6),
                                       "postprocess_none_32x32": (32, 32, 2, 3, 1e-4)}.items():
        img = torch.zeros(1, 3, h, w)
        anchors = Anchors()(img)
        A = anchors.shape[1]
        cls, reg = opp.synthetic_heads(A, K, seed, hot)
        dec, s, c, b, idx = reference_postprocess(torch.tensor(cls)[None], torch.tensor(reg)[None], anchors, img)
        mine_dec = opp.decode_clip(anchors.numpy(), reg, h, w)
        ms, mc, mb, mi = opp.postprocess(cls, reg, anchors.numpy(), h, w, boxes=dec.numpy())
        rep = {"A": int(A), "candidates": int((cls > 0.05).sum()), "kept": int(s.numel()),
               "decode_max_rel": float(np.abs(mine_dec - dec.numpy()).max() / np.abs(dec.numpy()).max()),
               "kept_identical": bool(np.array_equal(mi, idx.numpy()) and np.array_equal(mc, c.numpy())
                                      and np.array_equal(ms, s.numpy()) and np.array_equal(mb, b.numpy()))}
        print(name, rep)
        assert rep["kept_identical"] and rep["decode_max_rel"] < 1e-6
        report[name] = rep
        np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), cls=cls, reg=reg, anchors=anchors.numpy(),
                            image_hw=np.array([h, w]), decoded=dec.numpy(), scores=s.numpy(), classes=c.numpy(),
                            boxes=b.numpy(), index=idx.numpy())
    # larger pin, not stored: DDD17-sized image, thousands of candidates per class (several 1024-chunks)
    img = torch.zeros(1, 3, 260, 346)
    anchors = Anchors()(img)
    cls, reg = opp.synthetic_heads(anchors.shape[1], 3, 7, 0.07)
    dec, s, c, b, idx = reference_postprocess(torch.tensor(cls)[None], torch.tensor(reg)[None], anchors, img)
    _, mc, _, mi = opp.postprocess(cls, reg, anchors.numpy(), 260, 346, boxes=dec.numpy())
    ok = bool(np.array_equal(mi, idx.numpy()) and np.array_equal(mc, c.numpy()))
    report["pin_260x346"] = {"candidates": int((cls > 0.05).sum()), "kept": int(s.numel()), "kept_identical": ok}
    print(report["pin_260x346"])
    assert ok
    json.dump(report, open(os.path.join(GOLDEN, "postprocess_pin_report.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
