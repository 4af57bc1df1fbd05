"""Generate tests/golden/detection_*.npz by running the UNMODIFIED reference (retinanet/anchors.py,
retinanet/losses.py from /root/reference) on CPU.  Build container only:

    python oracle/make_golden_detection.py

TEST INFRASTRUCTURE ONLY.  Stores inputs (fp32), the reference's anchors, its losses in fp32 and fp64, the
fp64 autograd gradients of 0.7 * cls_loss + 1.3 * reg_loss, and the assignment derived from the reference's own
`calc_iou` + `torch.max` in fp32.  Also pins oracle/detection_numpy.py against the reference at a larger size
that is not stored (tests/golden/detection_pin_report.json)."""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REFERENCE = os.environ.get("FRN_REFERENCE", "/root/reference")
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, REFERENCE)
sys.path.insert(0, ROOT)

from retinanet.anchors import Anchors  # noqa: E402  (reference, unmodified)
from retinanet.losses import FocalLoss, calc_iou  # noqa: E402

from oracle import detection_numpy as odet  # noqa: E402

UP = (0.7, 1.3)


def reference_assignment(anchors, annotations):
    """Assignment codes from the reference's own fp32 ops (losses.py:45,81-90)."""
    out = []
    for ann in annotations:
        rows = torch.nonzero(ann[:, 4] != -1)[:, 0] if ann.shape[0] else torch.zeros(0, dtype=torch.long)
        st = torch.full((anchors.shape[1],), odet.NEGATIVE, dtype=torch.int32)
        if rows.numel():
            iou = calc_iou(anchors[0], ann[rows, :4])
            best, arg = torch.max(iou, dim=1)
            st[~torch.lt(best, 0.4)] = odet.IGNORE
            pos = torch.ge(best, 0.5)
            st[pos] = rows[arg[pos]].to(torch.int32)
        out.append(st)
    return torch.stack(out).numpy()


def make_case(seed, B, K, M, h, w, special=True):
    g = torch.Generator().manual_seed(seed)
    image = torch.zeros(B, 3, h, w)
    anchors = Anchors()(image)                                   # (1, A, 4) fp32, reference
    A = anchors.shape[1]
    cls = torch.sigmoid(2.0 * torch.randn(B, A, K, generator=g))
    flat = cls.view(-1)
    idx = torch.randperm(flat.numel(), generator=g)[:40]
    flat[idx[:20]] = 1e-6                                        # outside the clamp range (losses.py:46)
    flat[idx[20:]] = 1.0 - 1e-7
    reg = 0.5 * torch.randn(B, A, 4, generator=g)
    ann = torch.full((B, M, 5), -1.0)
    for b in range(B):
        if M == 0:
            break
        n_valid = int(torch.randint(1, M + 1, (1,), generator=g))
        if special and b == 1:
            continue                                             # every row padding: losses.py:48-79
        if special and b == 2:                                   # one tiny far-away box: no positive anchor
            ann[b, 0] = torch.tensor([1.0, 1.0, 2.5, 2.0, 1.0])
            continue
        for j in range(n_valid):
            src = anchors[0, int(torch.randint(0, A, (1,), generator=g))]
            jit = 0.08 * (src[2] - src[0]) * torch.randn(4, generator=g)
            ann[b, j, :4] = src + jit
            ann[b, j, 4] = float(torch.randint(0, K, (1,), generator=g))
        if special and b == 3 and n_valid >= 2:                  # duplicated box, different class: argmax tie
            ann[b, n_valid - 1, :4] = ann[b, 0, :4]
            ann[b, n_valid - 1, 4] = (ann[b, 0, 4] + 1) % K
        if M >= 3 and n_valid >= 3:                              # a padding row BEFORE valid rows
            ann[b, 1] = -1.0
    return image, anchors, cls, reg, ann


def run_reference(cls, reg, anchors, ann, dtype):
    c = cls.clone().to(dtype).requires_grad_(True)
    r = reg.clone().to(dtype).requires_grad_(True)
    lc, lr = FocalLoss()(c, r, anchors.to(dtype), ann.to(dtype))
    (UP[0] * lc + UP[1] * lr).sum().backward()
    zero = lambda t: torch.zeros_like(t).numpy()          # no annotation anywhere: the box loss is a constant
    return (float(lc.detach()), float(lr.detach()), c.grad.numpy() if c.grad is not None else zero(c),
            r.grad.numpy() if r.grad is not None else zero(r))


def compare(name, cls, reg, anchors, ann, image):
    lc64, lr64, dc64, dr64 = run_reference(cls, reg, anchors, ann, torch.float64)
    lc32, lr32, _, _ = run_reference(cls, reg, anchors, ann, torch.float32)
    st = reference_assignment(anchors, ann)
    mine = odet.focal_loss(cls.numpy(), reg.numpy(), anchors.numpy(), ann.numpy(), upstream=UP)
    my_anchors = odet.anchors_for_image(image.shape[2], image.shape[3])
    rep = {
        "anchors_bit_identical": bool(np.array_equal(my_anchors, anchors.numpy())),
        "assignment_identical": bool(np.array_equal(mine["assignment"], st)),
        "positives": int((st >= 0).sum()), "ignored": int((st == odet.IGNORE).sum()),
        "cls_loss_rel": abs(mine["cls_loss"] - lc64) / abs(lc64),
        "reg_loss_rel": abs(mine["reg_loss"] - lr64) / max(abs(lr64), 1e-300),
        "dcls_rel": odet.rel_l2(mine["dcls"], dc64), "dreg_rel": odet.rel_l2(mine["dreg"], dr64),
        "fp32_vs_fp64_cls_loss_rel": abs(lc32 - lc64) / abs(lc64),
    }
    print(name, rep)
    assert rep["anchors_bit_identical"] and rep["assignment_identical"], rep
    assert max(rep["cls_loss_rel"], rep["reg_loss_rel"], rep["dcls_rel"], rep["dreg_rel"]) < 1e-12, rep
    return rep, dict(cls=cls.numpy(), reg=reg.numpy(), anchors=anchors.numpy(), annotations=ann.numpy(),
                     image_hw=np.array(image.shape[2:]), upstream=np.array(UP),
                     cls_loss=np.float64(lc64), reg_loss=np.float64(lr64), cls_loss_f32=np.float32(lc32),
                     reg_loss_f32=np.float32(lr32), dcls=dc64, dreg=dr64, assignment=st)


def main():
    report = {}
    for name, args in (("detection_k3_40x56", dict(seed=1, B=4, K=3, M=6, h=40, w=56)),
                       ("detection_k8_33x47", dict(seed=2, B=2, K=8, M=3, h=33, w=47, special=False)),
                       ("detection_empty_32x32", dict(seed=3, B=2, K=2, M=0, h=32, w=32))):
        image, anchors, cls, reg, ann = make_case(**args)
        rep, arrays = compare(name, cls, reg, anchors, ann, image)
        report[name] = rep
        np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **arrays)
    # larger pin, not stored: DDD17-sized image, K = 3, 12 boxes per sample
    image, anchors, cls, reg, ann = make_case(seed=4, B=2, K=3, M=12, h=260, w=346, special=False)
    report["pin_260x346"], _ = compare("pin_260x346", cls, reg, anchors, ann, image)
    for hw in ((480, 640), (260, 346), (97, 131)):
        ref = Anchors()(torch.zeros(1, 3, *hw)).numpy()
        ok = bool(np.array_equal(odet.anchors_for_image(*hw), ref))
        report[f"anchors_{hw[0]}x{hw[1]}"] = {"count": int(ref.shape[1]), "bit_identical": ok}
        assert ok
    with open(os.path.join(GOLDEN, "detection_pin_report.json"), "w") as f:
        json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
