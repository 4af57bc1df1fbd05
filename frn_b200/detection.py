"""Drop-in replacements for the two host-bound hops after the heads:

* `Anchors`   -- retinanet/anchors.py:6-45.  The reference regenerates all boxes in NumPy and copies them
                 host -> device on every step (model.py:321); here a CUDA kernel writes them once per image
                 shape into device memory (bit-identical values) and the tensor is cached.
* `FocalLoss` -- retinanet/losses.py:26-169.  The reference loops over the batch in Python with boolean-mask
                 indexing and `if positive_indices.sum() > 0` (three host synchronisations and ~60 launches per
                 sample); here one forward and one backward kernel for the whole batch, no synchronisation.

Same constructors, call signatures and return values: `Anchors()(image) -> (1, A, 4)` fp32 on the image's
device, `FocalLoss()(classifications, regressions, anchors, annotations) -> (cls_loss (1,), reg_loss (1,))`.
CUDA tensors only: there is no CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .functional import _bytes, _check_cuda, _f32c, _ptr, _stream


def _int_array(values):
    return (C.c_int * len(values))(*[int(v) for v in values])


def _double_array(values):
    return (C.c_double * len(values))(*[float(v) for v in values])


class Anchors(nn.Module):
    def __init__(self, pyramid_levels=None, strides=None, sizes=None, ratios=None, scales=None):
        super().__init__()
        # anchors.py:10-22 (defaults).  The reference leaves an attribute unset when the argument is given
        # (an AttributeError later); here a given argument is simply used.
        self.pyramid_levels = [2, 3, 4, 5, 6] if pyramid_levels is None else list(pyramid_levels)
        self.strides = [2 ** x for x in self.pyramid_levels] if strides is None else list(strides)
        self.sizes = [2 ** (x + 2) for x in self.pyramid_levels] if sizes is None else list(sizes)
        self.ratios = np.array([0.5, 1, 2]) if ratios is None else np.asarray(ratios, dtype=np.float64)
        self.scales = (np.array([2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)]) if scales is None
                       else np.asarray(scales, dtype=np.float64))
        self._cache = {}

    def forward(self, image):
        h, w = int(image.shape[2]), int(image.shape[3])
        if not image.is_cuda:
            raise _lib.FrnError("frn_b200.Anchors: CUDA image expected (no CPU fallback)")
        key = (h, w, image.device.index)
        hit = self._cache.get(key)
        if hit is not None:
            return hit
        lib = _lib.load()
        levels = _int_array(self.pyramid_levels)
        per_cell = len(self.ratios) * len(self.scales)
        count = lib.frn_anchors_count(h, w, len(self.pyramid_levels), levels, per_cell)
        out = torch.empty((1, count, 4), dtype=torch.float32, device=image.device)
        with torch.cuda.device(image.device):
            _lib.check(lib.frn_anchors_generate(_stream(), h, w, len(self.pyramid_levels), levels,
                                                _int_array(self.strides), _int_array(self.sizes),
                                                len(self.ratios), _double_array(self.ratios),
                                                len(self.scales), _double_array(self.scales), _ptr(out)),
                       "anchors_generate")
        self._cache[key] = out
        return out


class _FocalLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, classifications, regressions, anchors, annotations, alpha):
        _check_cuda(classifications, regressions, anchors, annotations)
        lib = _lib.load()
        b, a, k = classifications.shape
        m = int(annotations.shape[1])
        cls, reg = _f32c(classifications), _f32c(regressions)
        anc, ann = _f32c(anchors[0]), _f32c(annotations)
        dev = cls.device
        losses = torch.empty(2, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            saved_bytes = lib.frn_focal_loss_saved_bytes(b, a)
            ws_bytes = lib.frn_focal_loss_workspace_bytes(b, a)
            saved, ws = _bytes(saved_bytes, dev), _bytes(ws_bytes, dev)
            _lib.check(lib.frn_focal_loss_forward(_stream(), b, a, k, m, _ptr(cls), _ptr(reg), _ptr(anc),
                                                  _ptr(ann) if m else None, alpha, _ptr(losses), _ptr(saved),
                                                  saved_bytes, _ptr(ws), ws_bytes), "focal_loss_forward")
        ctx.save_for_backward(cls, reg, anc, ann, saved)
        ctx.dims, ctx.alpha = (b, a, k, m), alpha
        ctx.in_dtypes = (classifications.dtype, regressions.dtype)
        return losses[0:1], losses[1:2]

    @staticmethod
    def backward(ctx, dcls_loss, dreg_loss):
        cls, reg, anc, ann, saved = ctx.saved_tensors
        b, a, k, m = ctx.dims
        lib = _lib.load()
        dl = torch.cat([dcls_loss.reshape(1), dreg_loss.reshape(1)]).float().contiguous()
        dcls, dreg = torch.empty_like(cls), torch.empty_like(reg)
        with torch.cuda.device(cls.device):
            _lib.check(lib.frn_focal_loss_backward(_stream(), b, a, k, m, _ptr(cls), _ptr(reg), _ptr(anc),
                                                   _ptr(ann) if m else None, ctx.alpha, _ptr(saved), _ptr(dl),
                                                   _ptr(dcls), _ptr(dreg)), "focal_loss_backward")
        return dcls.to(ctx.in_dtypes[0]), dreg.to(ctx.in_dtypes[1]), None, None, None


def focal_loss(classifications, regressions, anchors, annotations, alpha=0.25):
    return _FocalLossFn.apply(classifications, regressions, anchors, annotations, float(alpha))


def focal_assignment(classifications, regressions, anchors, annotations, alpha=0.25):
    """Debug / test hook: (losses (2,), assignment (B, A) int32, per_sample (B, 3) float64) of one forward."""
    _check_cuda(classifications, regressions, anchors, annotations)
    lib = _lib.load()
    b, a, k = classifications.shape
    m = int(annotations.shape[1])
    cls, reg, anc, ann = _f32c(classifications), _f32c(regressions), _f32c(anchors[0]), _f32c(annotations)
    dev = cls.device
    losses = torch.empty(2, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        saved_bytes = lib.frn_focal_loss_saved_bytes(b, a)
        ws_bytes = lib.frn_focal_loss_workspace_bytes(b, a)
        saved, ws = _bytes(saved_bytes, dev), _bytes(ws_bytes, dev)
        _lib.check(lib.frn_focal_loss_forward(_stream(), b, a, k, m, _ptr(cls), _ptr(reg), _ptr(anc),
                                              _ptr(ann) if m else None, float(alpha), _ptr(losses), _ptr(saved),
                                              saved_bytes, _ptr(ws), ws_bytes), "focal_loss_forward")
    assign = saved[: b * a * 4].view(torch.int32).reshape(b, a)
    off = (b * a * 4 + 255) // 256 * 256
    per_sample = saved[off: off + b * 3 * 8].view(torch.float64).reshape(b, 3)
    return losses, assign, per_sample


class FocalLoss(nn.Module):
    def forward(self, classifications, regressions, anchors, annotations):
        assert classifications.dim() == 3 and regressions.dim() == 3 and regressions.shape[2] == 4
        assert anchors.dim() == 3 and annotations.dim() == 3 and annotations.shape[2] == 5
        return focal_loss(classifications, regressions, anchors, annotations, alpha=0.25)


# --------------------------------------------------------------------------- row f3: inference post-processing
def _postprocess(classification, boxes_or_none, regression, anchors, img_hw, score_threshold, iou_threshold):
    lib = _lib.load()
    cls = _f32c(classification)
    a, k = cls.shape
    dev = cls.device
    cap = a * k
    scores = torch.empty(cap, dtype=torch.float32, device=dev)
    classes = torch.empty(cap, dtype=torch.int32, device=dev)
    boxes = torch.empty((cap, 4), dtype=torch.float32, device=dev)
    index = torch.empty(cap, dtype=torch.int32, device=dev)
    count = torch.empty(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        ws_bytes = lib.frn_detect_workspace_bytes(a, k)
        ws = _bytes(ws_bytes, dev)
        if boxes_or_none is not None:
            bx = _f32c(boxes_or_none)
            rc = lib.frn_detect_postprocess(_stream(), a, k, _ptr(cls), None, None, _ptr(bx), 0, 0, score_threshold,
                                            iou_threshold, cap, _ptr(scores), _ptr(classes), _ptr(boxes), _ptr(index),
                                            _ptr(count), _ptr(ws), ws_bytes)
        else:
            reg, anc = _f32c(regression), _f32c(anchors)
            rc = lib.frn_detect_postprocess(_stream(), a, k, _ptr(cls), _ptr(reg), _ptr(anc), None, int(img_hw[0]),
                                            int(img_hw[1]), score_threshold, iou_threshold, cap, _ptr(scores),
                                            _ptr(classes), _ptr(boxes), _ptr(index), _ptr(count), _ptr(ws), ws_bytes)
        _lib.check(rc, "detect_postprocess")
    n = int(count.item())                      # the one D2H read: the result tensors are sized by the data
    return scores[:n], classes[:n], boxes[:n], index[:n]


def detect_postprocess(classification, regression, anchors, img_batch, score_threshold=0.05, iou_threshold=0.5):
    """model.py:326-364 (eval branch of ResNet.forward) for the reference's batch of 1: classification (1, A, K),
    regression (1, A, 4), anchors (1, A, 4), img_batch (1, C, H, W) -> [scores, class indexes (int64), boxes]."""
    _check_cuda(classification, regression, anchors)
    assert classification.dim() == 3 and classification.shape[0] == 1, "batch of 1 (model.py:340 squeezes it)"
    h, w = int(img_batch.shape[2]), int(img_batch.shape[3])
    s, c, b, _ = _postprocess(classification[0], None, regression[0], anchors[0], (h, w), float(score_threshold),
                              float(iou_threshold))
    return [s, c.long(), b]


def decode_boxes(anchors, regression, img_batch):
    """regressBoxes + clipBoxes (model.py:326-327; utils.py:103-126, 135-145): (1, A, 4) -> (1, A, 4)."""
    _check_cuda(anchors, regression)
    lib = _lib.load()
    anc, reg = _f32c(anchors[0]), _f32c(regression[0])
    out = torch.empty_like(anc)
    with torch.cuda.device(anc.device):
        _lib.check(lib.frn_decode_boxes(_stream(), anc.shape[0], _ptr(anc), _ptr(reg), int(img_batch.shape[2]),
                                        int(img_batch.shape[3]), _ptr(out)), "decode_boxes")
    return out[None]


def nms(boxes, scores, iou_threshold):
    """torchvision.ops.nms(boxes (N, 4), scores (N), iou_threshold) -> int64 indices kept, by decreasing score."""
    _check_cuda(boxes, scores)
    if boxes.shape[0] == 0:
        return torch.empty(0, dtype=torch.int64, device=boxes.device)
    _, _, _, idx = _postprocess(scores.reshape(-1, 1), boxes, None, None, None, -3.0e38, float(iou_threshold))
    return idx.long()
