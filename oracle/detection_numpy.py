"""TEST INFRASTRUCTURE ONLY -- CPU restatement (NumPy) of the anchor generator and the focal loss of
JinkaiLiu/FRN.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg
may import this; the product path (frn_b200/detection.py -> libfrn_b200.so) never does.

    anchors_for_image   retinanet/anchors.py:25-45 (Anchors.forward), :47-82 (generate_anchors), :120-147 (shift)
    iou_matrix          retinanet/losses.py:5-24  (calc_iou)
    focal_loss          retinanet/losses.py:28-169 (FocalLoss.forward) + hand-derived gradients

Pinned against the unmodified reference (`retinanet.anchors.Anchors`, `retinanet.losses.FocalLoss` and
`calc_iou`, imported from /root/reference by oracle/make_golden_detection.py): anchors bit-identical, assignment
identical, losses and autograd gradients <= 1e-13 relative in fp64 (tests/golden/detection_pin_report.json,
tests/test_oracle_detection.py on the committed fixtures)."""
from __future__ import annotations

import numpy as np

NEGATIVE, IGNORE = -1, -2          # assignment codes (>= 0: matched annotation row)


def anchors_for_image(height, width, pyramid_levels=(2, 3, 4, 5, 6), strides=None, sizes=None,
                      ratios=(0.5, 1.0, 2.0), scales=(2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0))):
    """(1, A, 4) float32 boxes (x1, y1, x2, y2); float64 arithmetic, one rounding at the end."""
    levels = list(pyramid_levels)
    strides = [2 ** p for p in levels] if strides is None else list(strides)
    sizes = [2 ** (p + 2) for p in levels] if sizes is None else list(sizes)
    ratios = np.asarray(ratios, dtype=np.float64)
    scales = np.asarray(scales, dtype=np.float64)
    per_ratio = np.repeat(ratios, len(scales))            # box i: ratio i // n_scales, scale i % n_scales
    per_scale = np.tile(scales, len(ratios))
    chunks = []
    for p, stride, size in zip(levels, strides, sizes):
        side = size * per_scale
        w = np.sqrt(side * side / per_ratio)
        h = w * per_ratio
        base = np.stack([0.0 - w * 0.5, 0.0 - h * 0.5, w - w * 0.5, h - h * 0.5], axis=1)       # (nA, 4)
        fh, fw = (height + 2 ** p - 1) // 2 ** p, (width + 2 ** p - 1) // 2 ** p
        cx = (np.arange(fw) + 0.5) * stride
        cy = (np.arange(fh) + 0.5) * stride
        shift = np.stack([np.tile(cx, fh), np.repeat(cy, fw), np.tile(cx, fh), np.repeat(cy, fw)], axis=1)
        chunks.append((shift[:, None, :] + base[None, :, :]).reshape(-1, 4))
    return np.concatenate(chunks, axis=0).astype(np.float32)[None]


def iou_matrix(a, b):
    """IoU of every box of a (A, 4) with every box of b (M, 4), in the arrays' dtype and the reference's
    operation order (losses.py:5-24)."""
    area_b = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    iw = np.minimum(a[:, 2:3], b[:, 2]) - np.maximum(a[:, 0:1], b[:, 0])
    ih = np.minimum(a[:, 3:4], b[:, 3]) - np.maximum(a[:, 1:2], b[:, 1])
    iw = np.maximum(iw, 0)
    ih = np.maximum(ih, 0)
    inter = iw * ih
    ua = ((a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1]))[:, None] + area_b - inter
    ua = np.maximum(ua, a.dtype.type(1e-8))
    return inter / ua


def assign(anchors, annotations, dtype=np.float32):
    """Per-anchor assignment (A,) int32 for one sample: matched annotation ROW (of the unfiltered annotation
    array) where IoU_max >= 0.5, NEGATIVE where < 0.4, IGNORE between (losses.py:81-96).  Thresholds are
    evaluated in `dtype` (the reference works in fp32)."""
    anchors = np.asarray(anchors, dtype=dtype)
    ann = np.asarray(annotations, dtype=dtype)
    rows = np.nonzero(ann[:, 4] != -1)[0] if ann.size else np.zeros(0, dtype=np.int64)
    out = np.full(anchors.shape[0], NEGATIVE, dtype=np.int32)
    if rows.size == 0:
        return out
    iou = iou_matrix(anchors, ann[rows, :4])
    arg = np.argmax(iou, axis=1)                      # first maximum
    best = iou[np.arange(iou.shape[0]), arg]
    out[best >= dtype(0.4)] = IGNORE
    pos = best >= dtype(0.5)
    out[pos] = rows[arg[pos]].astype(np.int32)
    return out


def _box_residual(anc, gt, reg):
    aw, ah = anc[:, 2] - anc[:, 0], anc[:, 3] - anc[:, 1]
    acx, acy = anc[:, 0] + 0.5 * aw, anc[:, 1] + 0.5 * ah
    gw, gh = gt[:, 2] - gt[:, 0], gt[:, 3] - gt[:, 1]
    gcx, gcy = gt[:, 0] + 0.5 * gw, gt[:, 1] + 0.5 * gh
    gw, gh = np.maximum(gw, 1.0), np.maximum(gh, 1.0)
    t = np.stack([(gcx - acx) / aw, (gcy - acy) / ah, np.log(gw / aw), np.log(gh / ah)], axis=1)
    t = t / np.array([[0.1, 0.1, 0.2, 0.2]], dtype=np.float32).astype(np.float64)     # losses.py:143-146: fp32 constants
    return t - reg


def focal_loss(classifications, regressions, anchors, annotations, alpha=0.25, upstream=(1.0, 1.0),
               assign_dtype=np.float32):
    """fp64 focal loss of the whole batch and its gradients.

    classifications (B, A, K), regressions (B, A, 4), anchors (1, A, 4) or (A, 4), annotations (B, M, 5).
    Returns dict: cls_loss, reg_loss (floats: the two batch means), assignment (B, A) int32,
    per_sample (B, 3) = [cls loss, box loss, positives], dcls, dreg = gradients of
    upstream[0] * cls_loss + upstream[1] * reg_loss."""
    cls_raw = np.asarray(classifications, dtype=np.float64)
    reg = np.asarray(regressions, dtype=np.float64)
    anc32 = np.asarray(anchors).reshape(-1, 4)
    anc = anc32.astype(np.float64)
    B, A, K = cls_raw.shape
    lo, hi = 1e-4, 1.0 - 1e-4
    c = np.clip(cls_raw, lo, hi)
    inside = (cls_raw >= lo) & (cls_raw <= hi)
    state = np.zeros((B, A), dtype=np.int32)
    per_sample = np.zeros((B, 3))
    dcls, dreg = np.zeros_like(cls_raw), np.zeros_like(reg)
    for b in range(B):
        ann = np.asarray(annotations[b]).reshape(-1, 5)
        st = assign(anc32, ann, assign_dtype) if ann.shape[0] else np.full(A, NEGATIVE, dtype=np.int32)
        state[b] = st
        pos = st >= 0
        npos = int(pos.sum())
        one = np.zeros((A, K), dtype=bool)
        if npos:
            one[np.nonzero(pos)[0], ann[st[pos], 4].astype(np.int64)] = True
        counted = (st != IGNORE)[:, None]
        cb = c[b]
        term = np.where(one, alpha * (1 - cb) ** 2 * -np.log(cb), (1 - alpha) * cb ** 2 * -np.log1p(-cb))
        dterm = np.where(one, alpha * (2 * (1 - cb) * np.log(cb) - (1 - cb) ** 2 / cb),
                         (1 - alpha) * (2 * cb * -np.log1p(-cb) + cb ** 2 / (1 - cb)))
        norm = max(npos, 1)
        per_sample[b, 0] = (term * counted).sum() / norm
        dcls[b] = dterm * counted * inside[b] * (upstream[0] / B / norm)
        per_sample[b, 2] = npos
        if npos:
            d = _box_residual(anc[pos], ann[st[pos], :4].astype(np.float64), reg[b, pos])
            v = np.abs(d)
            small = v <= 1.0 / 9.0
            per_sample[b, 1] = np.where(small, 4.5 * v * v, v - 0.5 / 9.0).mean()
            dreg[b, pos] = -np.sign(d) * np.where(small, 9.0 * v, 1.0) * (upstream[1] / B / (4.0 * npos))
    return {"cls_loss": per_sample[:, 0].mean(), "reg_loss": per_sample[:, 1].mean(), "assignment": state,
            "per_sample": per_sample, "dcls": dcls, "dreg": dreg}


def rel_l2(x, ref):
    x, ref = np.asarray(x, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    den = np.linalg.norm(ref.ravel())
    return float(np.linalg.norm((x - ref).ravel()) / den) if den > 0 else float(np.linalg.norm(x.ravel()))
