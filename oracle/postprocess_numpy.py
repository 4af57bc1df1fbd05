"""TEST INFRASTRUCTURE ONLY -- CPU restatement (NumPy, fp32 step by step) of FRN's inference post-processing
(SURVEY.md 8.2 row f3).  Only tests/, __graft_entry__.smoke() and a bench's cpu_baseline leg may import this.

    decode_clip   retinanet/utils.py:103-126 (BBoxTransform.forward) + :135-145 (ClipBoxes.forward)
    nms           torchvision.ops.nms -- a third-party dependency of the reference (model.py:5; torchvision is
                  not vendored, README pins 0.15.2, this image has 0.26.0); restated from its published CPU
                  algorithm: sort by score descending, keep a box, suppress every later box whose
                  inter / (area_i + area_j - inter) exceeds the threshold
    postprocess   retinanet/model.py:326-364 (threshold 0.05, NMS 0.5, classes concatenated)

Pinned by oracle/make_golden_postprocess.py against the reference's own BBoxTransform / ClipBoxes classes and
torchvision.ops.nms executed in the build container (tests/golden/postprocess_*.npz, postprocess_pin_report.json)."""
from __future__ import annotations

import numpy as np

F = np.float32


def decode_clip(anchors, deltas, height, width):
    a, d = np.asarray(anchors, dtype=F).reshape(-1, 4), np.asarray(deltas, dtype=F).reshape(-1, 4)
    w, h = a[:, 2] - a[:, 0], a[:, 3] - a[:, 1]
    cx, cy = a[:, 0] + F(0.5) * w, a[:, 1] + F(0.5) * h
    dx, dy = d[:, 0] * F(0.1) + F(0), d[:, 1] * F(0.1) + F(0)
    dw, dh = d[:, 2] * F(0.2) + F(0), d[:, 3] * F(0.2) + F(0)
    pcx, pcy = cx + dx * w, cy + dy * h
    pw, ph = np.exp(dw) * w, np.exp(dh) * h
    out = np.stack([pcx - F(0.5) * pw, pcy - F(0.5) * ph, pcx + F(0.5) * pw, pcy + F(0.5) * ph], axis=1)
    out[:, 0] = np.maximum(out[:, 0], F(0)); out[:, 1] = np.maximum(out[:, 1], F(0))
    out[:, 2] = np.minimum(out[:, 2], F(width)); out[:, 3] = np.minimum(out[:, 3], F(height))
    return out.astype(F)


def nms(boxes, scores, iou_threshold):
    """Indices kept, by decreasing score (ties: lower index first)."""
    b, s = np.asarray(boxes, dtype=F), np.asarray(scores, dtype=F)
    order = np.argsort(-s, kind="stable")
    b = b[order]
    area = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    dead = np.zeros(len(order), dtype=bool)
    keep = []
    thr = F(iou_threshold)
    for i in range(len(order)):
        if dead[i]:
            continue
        keep.append(order[i])
        r = slice(i + 1, None)
        w = np.maximum(F(0), np.minimum(b[i, 2], b[r, 2]) - np.maximum(b[i, 0], b[r, 0]))
        h = np.maximum(F(0), np.minimum(b[i, 3], b[r, 3]) - np.maximum(b[i, 1], b[r, 1]))
        inter = w * h
        with np.errstate(divide="ignore", invalid="ignore"):
            ovr = inter / (area[i] + area[r] - inter)
        dead[r] |= ovr > thr
    return np.asarray(keep, dtype=np.int64)


def postprocess(classification, regression, anchors, height, width, score_threshold=0.05, iou_threshold=0.5,
                boxes=None):
    """classification (A, K) -> scores (n,), classes (n,) int64, boxes (n, 4), anchor index (n,)."""
    cls = np.asarray(classification, dtype=F)
    if boxes is None:
        boxes = decode_clip(anchors, regression, height, width)
    out = ([], [], [], [])
    for k in range(cls.shape[1]):
        cand = np.nonzero(cls[:, k] > F(score_threshold))[0]
        if cand.size == 0:
            continue
        keep = cand[nms(boxes[cand], cls[cand, k], iou_threshold)]
        out[0].append(cls[keep, k]); out[1].append(np.full(keep.size, k, dtype=np.int64))
        out[2].append(boxes[keep]); out[3].append(keep)
    if not out[0]:
        return np.zeros(0, F), np.zeros(0, np.int64), np.zeros((0, 4), F), np.zeros(0, np.int64)
    return tuple(np.concatenate(o) for o in out)


def synthetic_heads(A, K, seed, hot=0.07):
    """Head outputs with a few percent of the scores above 0.05 and overlapping boxes."""
    rng = np.random.default_rng(seed)
    cls = (1.0 / (1.0 + np.exp(-(2.0 * rng.standard_normal((A, K)) - 3.0 + np.log(hot / 0.07))))).astype(F)
    reg = (0.5 * rng.standard_normal((A, 4))).astype(F)
    return cls, reg
