"""GPU: anchors + focal loss kernels through frn_b200.Anchors / frn_b200.FocalLoss and the
C ABI, against the golden fixtures of the unmodified reference and the fp64 oracle (oracle/detection_numpy.py).

Bars: anchors bit-identical (fp32 bits); assignment (integer work) identical; losses within 1e-6 relative and
gradients within 1e-5 l2-relative of fp64 (north_star: fp32 <= 1e-5 relative)."""
import glob
import os

import numpy as np
import pytest
import torch

import frn_b200
from frn_b200 import detection as det
from helpers import GOLDEN
from oracle import detection_numpy as odet

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
FILES = sorted(glob.glob(os.path.join(GOLDEN, "detection_*.npz")))


@pytest.mark.parametrize("hw", [(480, 640), (260, 346), (97, 131), (40, 56), (1, 1)])
def test_anchors_bit_identical(hw):
    mod = frn_b200.Anchors()
    out = mod(torch.zeros(2, 3, *hw, device=DEV))
    ref = odet.anchors_for_image(*hw)
    assert out.shape == ref.shape and out.dtype == torch.float32 and out.is_cuda
    assert np.array_equal(out.cpu().numpy().view(np.uint32), ref.view(np.uint32))
    assert mod(torch.zeros(1, 3, *hw, device=DEV)) is out          # cached per image shape: no regeneration


def test_anchors_custom_pyramid():
    mod = frn_b200.Anchors(pyramid_levels=[3, 4], ratios=[1.0, 3.0], scales=[1.0, 1.5, 2.0, 2.5])
    out = mod(torch.zeros(1, 3, 50, 70, device=DEV))
    ref = odet.anchors_for_image(50, 70, pyramid_levels=[3, 4], ratios=[1.0, 3.0], scales=[1.0, 1.5, 2.0, 2.5])
    assert np.array_equal(out.cpu().numpy(), ref)


def _run(cls, reg, anchors, ann, up):
    c = torch.tensor(cls, device=DEV).requires_grad_(True)
    r = torch.tensor(reg, device=DEV).requires_grad_(True)
    lc, lr = frn_b200.FocalLoss()(c, r, torch.tensor(anchors, device=DEV), torch.tensor(ann, device=DEV))
    assert lc.shape == (1,) and lr.shape == (1,)
    (up[0] * lc + up[1] * lr).sum().backward()
    return float(lc), float(lr), c.grad.cpu().numpy(), r.grad.cpu().numpy()


def _check(got, want, name):
    lc, lr, dc, dr = got
    assert abs(lc - want["cls_loss"]) <= 1e-6 * abs(want["cls_loss"]), name
    assert abs(lr - want["reg_loss"]) <= 1e-6 * abs(want["reg_loss"]) + 1e-30, name
    assert odet.rel_l2(dc, want["dcls"]) <= 1e-5, name
    if np.abs(want["dreg"]).max() > 0:
        assert odet.rel_l2(dr, want["dreg"]) <= 1e-5, name
    else:
        assert np.abs(dr).max() == 0.0
    assert ((want["dreg"] == 0) == (dr == 0)).all()            # gradient only on positive anchors
    assert ((want["dcls"] == 0) == (dc == 0)).all()            # zero on ignored anchors and outside the clamp


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p)[:-4] for p in FILES])
def test_focal_loss_matches_reference_fixture(path):
    z = np.load(path)
    up = tuple(float(v) for v in z["upstream"])
    got = _run(z["cls"], z["reg"], z["anchors"], z["annotations"], up)
    _check(got, {"cls_loss": float(z["cls_loss"]), "reg_loss": float(z["reg_loss"]), "dcls": z["dcls"],
                 "dreg": z["dreg"]}, path)
    _, assign, per_sample = det.focal_assignment(*(torch.tensor(z[k], device=DEV)
                                                   for k in ("cls", "reg", "anchors", "annotations")))
    assert np.array_equal(assign.cpu().numpy(), z["assignment"])
    assert np.array_equal(per_sample[:, 2].cpu().numpy(), (z["assignment"] >= 0).sum(1).astype(np.float64))


def _synthetic(B, K, M, hw, seed):
    g = torch.Generator().manual_seed(seed)
    anchors = odet.anchors_for_image(*hw)
    A = anchors.shape[1]
    cls = torch.sigmoid(torch.randn(B, A, K, generator=g) - 2.0).numpy()
    reg = (0.3 * torch.randn(B, A, 4, generator=g)).numpy()
    ann = np.full((B, M, 5), -1.0, dtype=np.float32)
    for b in range(B):
        n = int(torch.randint(0, M + 1, (1,), generator=g))
        for j in range(n):
            x1 = float(torch.rand(1, generator=g)) * hw[1] * 0.8
            y1 = float(torch.rand(1, generator=g)) * hw[0] * 0.8
            bw = 8 + float(torch.rand(1, generator=g)) * hw[1] * 0.3
            bh = 8 + float(torch.rand(1, generator=g)) * hw[0] * 0.3
            ann[b, j] = (x1, y1, x1 + bw, y1 + bh, float(torch.randint(0, K, (1,), generator=g)))
    return cls, reg, anchors, ann


@pytest.mark.parametrize("B,K,M,hw", [(8, 3, 16, (480, 640)), (4, 2, 7, (260, 346)), (3, 80, 5, (128, 160))],
                         ids=["dsec_b8_k3", "ddd17_b4_k2", "k80"])
def test_focal_loss_full_size_vs_oracle(B, K, M, hw):
    cls, reg, anchors, ann = _synthetic(B, K, M, hw, seed=B + K)
    up = (1.0, 1.0)
    want = odet.focal_loss(cls, reg, anchors, ann, upstream=up)
    _check(_run(cls, reg, anchors, ann, up), want, "synthetic")
    _, assign, _ = det.focal_assignment(*(torch.tensor(v, device=DEV) for v in (cls, reg, anchors, ann)))
    assert np.array_equal(assign.cpu().numpy(), want["assignment"])


def test_bit_reproducible_and_batch_independent():
    cls, reg, anchors, ann = _synthetic(4, 3, 6, (260, 346), seed=9)
    a = _run(cls, reg, anchors, ann, (1.0, 1.0))
    b = _run(cls, reg, anchors, ann, (1.0, 1.0))
    assert a[0] == b[0] and a[1] == b[1] and np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])
    # the loss of the batch is the mean of the per-sample losses (losses.py:169)
    singles = [_run(cls[i:i + 1], reg[i:i + 1], anchors, ann[i:i + 1], (1.0, 1.0)) for i in range(4)]
    assert abs(np.mean([s[0] for s in singles]) - a[0]) <= 1e-6 * abs(a[0])
    assert abs(np.mean([s[1] for s in singles]) - a[1]) <= 1e-6 * abs(a[1]) + 1e-12
