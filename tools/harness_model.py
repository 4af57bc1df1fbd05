"""BENCH / TEST HARNESS, not product: a network of the same shape as the reference's `ResNet` with
fusion_model='fpn_fusion' (retinanet/model.py:160-364) assembled from stock layers, so that BASELINE.json config 3
("end-to-end FRN RetinaNet training step ... with the new fusion modules, synthetic DSEC batch, random-init weights")
can run on the GPU box, where /root/reference does not exist.

    two torchvision ResNet-50 trunks (RGB: 3 channels, events: 5), BatchNorm frozen (model.py:245,270-274)
    -> four fusion blocks on (event, rgb) feature pairs at C = 256 / 512 / 1024 / 2048 (model.py:247-252,305-308)
    -> FPN over the 2C-channel fused maps, P2 ... P6 (model.py:21-85)
    -> regression / classification heads, 9 anchors per cell, prior 0.01 (model.py:88-158,237-243)
    -> anchors + focal loss (model.py:321-324)

`impl='ours'`: frn_b200.REFusion, frn_b200.Conv1x1 laterals, frn_b200.Anchors, frn_b200.FocalLoss.
`impl='eager'`: the same arithmetic issued as stock PyTorch ops on the GPU (what the unmodified reference does
when it is run on a GPU): the fusion block as conv2d / matmul / softmax / var / mean calls that materialise the
N x N attention matrix, NumPy anchors copied to the device every step, and the per-sample Python loss loop."""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
import torchvision


# ------------------------------------------------------------------ eager stand-ins (stock PyTorch ops on the GPU)
class EagerCrossAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.g, self.theta, self.phi = nn.Conv2d(c, c // 8, 1), nn.Conv2d(c, c // 8, 1), nn.Conv2d(c, c // 8, 1)
        self.W = nn.Conv2d(c // 8, c, 1)

    @staticmethod
    def stats(t, eps=1e-5):
        b, c = t.shape[:2]
        v = t.reshape(b, c, -1)
        return v.mean(2).view(b, c, 1, 1), (v.var(2) + eps).sqrt().view(b, c, 1, 1)

    def forward(self, x0, x1):
        b, c, h, w = x0.shape
        v = self.g(x0).view(b, c // 8, -1).permute(0, 2, 1)
        k = self.theta(x1).view(b, c // 8, -1)
        q = self.phi(x1).view(b, c // 8, -1).permute(0, 2, 1)
        y = torch.matmul(F.softmax(torch.matmul(q, k), dim=-1), v).contiguous().view(b, c // 8, h, w)
        t = self.W(y)
        mx, sx = self.stats(x0)
        mt, st = self.stats(t)
        return (x0 - mx) / sx * st + mt


class EagerREFusion(nn.Module):
    def __init__(self, c_in, c_out):
        super().__init__()
        self.rgb_cross_attention, self.event_cross_attention = EagerCrossAttention(c_in), EagerCrossAttention(c_in)
        self.conv0_rgb, self.conv0_evt = nn.Conv2d(c_in, c_out, 1), nn.Conv2d(c_out, c_out, 1)

    def forward(self, rgb, evt):
        r0, e0 = self.conv0_rgb(rgb), self.conv0_evt(evt)
        m = r0 * e0
        r1, e1 = r0 + m, e0 + m
        return torch.cat([self.rgb_cross_attention(r1, e1), self.event_cross_attention(e1, r1)], dim=1)


class EagerAnchors(nn.Module):
    """NumPy on the host + H2D copy every step, as the reference does."""

    def forward(self, image):
        h, w = image.shape[2:]
        ratios, scales = np.array([0.5, 1, 2]), np.array([2 ** 0, 2 ** (1 / 3), 2 ** (2 / 3)])
        out = []
        for p in (2, 3, 4, 5, 6):
            side = 2 ** (p + 2) * np.tile(scales, 3)
            bw = np.sqrt(side * side / np.repeat(ratios, 3))
            bh = bw * np.repeat(ratios, 3)
            base = np.stack([-bw / 2, -bh / 2, bw / 2, bh / 2], 1)
            fh, fw = (h + 2 ** p - 1) // 2 ** p, (w + 2 ** p - 1) // 2 ** p
            sx, sy = np.meshgrid((np.arange(fw) + 0.5) * 2 ** p, (np.arange(fh) + 0.5) * 2 ** p)
            shift = np.stack([sx.ravel(), sy.ravel(), sx.ravel(), sy.ravel()], 1)
            out.append((shift[:, None] + base[None]).reshape(-1, 4))
        return torch.from_numpy(np.concatenate(out).astype(np.float32)[None]).to(image.device)


class EagerFocalLoss(nn.Module):
    """Per-sample Python loop with boolean-mask indexing (host synchronisations), as the reference does."""

    def forward(self, cls, reg, anchors, ann):
        a = anchors[0]
        aw, ah = a[:, 2] - a[:, 0], a[:, 3] - a[:, 1]
        ax, ay = a[:, 0] + 0.5 * aw, a[:, 1] + 0.5 * ah
        cl, rl = [], []
        for j in range(cls.shape[0]):
            c = cls[j].clamp(1e-4, 1 - 1e-4)
            box = ann[j][ann[j][:, 4] != -1]
            if box.shape[0] == 0:
                cl.append((0.75 * c ** 2 * -torch.log(1 - c)).sum()); rl.append(c.new_zeros(()))
                continue
            area = (box[:, 2] - box[:, 0]) * (box[:, 3] - box[:, 1])
            iw = (torch.min(a[:, 2:3], box[:, 2]) - torch.max(a[:, 0:1], box[:, 0])).clamp(min=0)
            ih = (torch.min(a[:, 3:4], box[:, 3]) - torch.max(a[:, 1:2], box[:, 1])).clamp(min=0)
            iou = iw * ih / ((aw * ah)[:, None] + area - iw * ih).clamp(min=1e-8)
            best, arg = iou.max(1)
            tgt = -torch.ones_like(c)
            tgt[best < 0.4] = 0
            pos = best >= 0.5
            got = box[arg]
            tgt[pos] = 0
            tgt[pos, got[pos, 4].long()] = 1
            af = torch.where(tgt == 1, 0.25, 0.75)
            fw = af * torch.where(tgt == 1, 1 - c, c) ** 2
            l = fw * -(tgt * torch.log(c) + (1 - tgt) * torch.log(1 - c))
            cl.append(torch.where(tgt != -1, l, torch.zeros_like(l)).sum() / pos.sum().float().clamp(min=1))
            if pos.sum() > 0:
                g = got[pos]
                gw, gh = g[:, 2] - g[:, 0], g[:, 3] - g[:, 1]
                gx, gy = g[:, 0] + 0.5 * gw, g[:, 1] + 0.5 * gh
                gw, gh = gw.clamp(min=1), gh.clamp(min=1)
                t = torch.stack([(gx - ax[pos]) / aw[pos], (gy - ay[pos]) / ah[pos], torch.log(gw / aw[pos]),
                                 torch.log(gh / ah[pos])], 1) / c.new_tensor([[0.1, 0.1, 0.2, 0.2]])
                d = (t - reg[j][pos]).abs()
                rl.append(torch.where(d <= 1 / 9, 4.5 * d ** 2, d - 0.5 / 9).mean())
            else:
                rl.append(c.new_zeros(()))
        return torch.stack(cl).mean(0, keepdim=True), torch.stack(rl).mean(0, keepdim=True)


# ------------------------------------------------------------------------------------------------- the harness
class Head(nn.Module):
    def __init__(self, out_per_anchor, sigmoid, anchors=9, width=256):
        super().__init__()
        self.convs = nn.ModuleList([nn.Conv2d(width, width, 3, padding=1) for _ in range(4)])
        self.output = nn.Conv2d(width, anchors * out_per_anchor, 3, padding=1)
        self.k, self.sigmoid = out_per_anchor, sigmoid

    def forward(self, x):
        for conv in self.convs:
            x = F.relu(conv(x))
        x = self.output(x)
        x = torch.sigmoid(x) if self.sigmoid else x
        return x.permute(0, 2, 3, 1).contiguous().view(x.shape[0], -1, self.k)


class Pyramid(nn.Module):
    def __init__(self, sizes, lateral, width=256):
        super().__init__()
        c2, c3, c4, c5 = sizes
        self.P5_1, self.P4_1, self.P3_1, self.P2_1 = (lateral(c, width) for c in (c5, c4, c3, c2))
        self.P5_2, self.P4_2, self.P3_2, self.P2_2 = (nn.Conv2d(width, width, 3, padding=1) for _ in range(4))
        self.P6 = nn.Conv2d(c5, width, 3, stride=2, padding=1)

    def forward(self, feats):
        c2, c3, c4, c5 = feats
        up = lambda t, like: F.interpolate(t, size=like.shape[2:], mode="nearest")
        p5 = self.P5_1(c5)
        p4 = self.P4_1(c4) + up(p5, c4)
        p3 = self.P3_1(c3) + up(p4, c3)
        p2 = self.P2_1(c2) + up(p3, c2)
        return [self.P2_2(p2), self.P3_2(p3), self.P4_2(p4), self.P5_2(p5), self.P6(c5)]


class FusionDetector(nn.Module):
    def __init__(self, num_classes=3, impl="ours"):
        super().__init__()
        import frn_b200
        self.rgb = torchvision.models.resnet50(weights=None)
        self.evt = torchvision.models.resnet50(weights=None)
        self.evt.conv1 = nn.Conv2d(5, 64, 7, stride=2, padding=3, bias=False)
        for trunk in (self.rgb, self.evt):
            trunk.fc = nn.Identity()
        chans = (256, 512, 1024, 2048)
        ours = impl == "ours"
        fusion = frn_b200.REFusion if ours else EagerREFusion
        self.fus = nn.ModuleList([fusion(c, c) for c in chans])
        lateral = (lambda ci, co: frn_b200.Conv1x1(ci, co)) if ours else (lambda ci, co: nn.Conv2d(ci, co, 1))
        self.fpn = Pyramid([2 * c for c in chans], lateral)
        self.regressionModel, self.classificationModel = Head(4, False), Head(num_classes, True)
        self.anchors = frn_b200.Anchors() if ours else EagerAnchors()
        self.focalLoss = frn_b200.FocalLoss() if ours else EagerFocalLoss()
        prior = 0.01
        nn.init.zeros_(self.classificationModel.output.weight)
        nn.init.constant_(self.classificationModel.output.bias, -math.log((1 - prior) / prior))
        nn.init.zeros_(self.regressionModel.output.weight)
        nn.init.zeros_(self.regressionModel.output.bias)

    def train(self, mode=True):
        super().train(mode)
        for m in self.modules():                       # frozen BatchNorm (model.py:270-274)
            if isinstance(m, nn.BatchNorm2d):
                m.eval()
        return self

    @staticmethod
    def stem(trunk, x):
        x = trunk.maxpool(trunk.relu(trunk.bn1(trunk.conv1(x))))
        out = []
        for layer in (trunk.layer1, trunk.layer2, trunk.layer3, trunk.layer4):
            x = layer(x)
            out.append(x)
        return out

    def forward(self, img_rgb, img_evt, annotations=None):
        fr, fe = self.stem(self.rgb, img_rgb), self.stem(self.evt, img_evt)
        fused = [f(e, r) for f, e, r in zip(self.fus, fe, fr)]          # model.py:305-308: (event, rgb)
        feats = self.fpn(fused)
        regression = torch.cat([self.regressionModel(f) for f in feats], dim=1)
        classification = torch.cat([self.classificationModel(f) for f in feats], dim=1)
        anchors = self.anchors(img_rgb)
        if self.training:
            return self.focalLoss(classification, regression, anchors, annotations)
        return classification, regression, anchors


@torch.no_grad()
def calibrate_batchnorm(net, img_rgb, img_evt):
    """Random-init trunks with untouched running statistics (mean 0, var 1) let the activations grow by orders of
    magnitude through the 16 bottlenecks, which a trained network never shows and which puts the attention softmax
    into saturation.  One forward pass through the trunks with the BatchNorm layers in training mode and momentum 1
    sets the running statistics to those of the batch; the layers are frozen again afterwards."""
    bns = [m for m in net.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.momentum = 1.0
        m.train()
    net.stem(net.rgb, img_rgb)
    net.stem(net.evt, img_evt)
    for m in bns:
        m.momentum = 0.1
        m.eval()


def synthetic_batch(batch, device, height=480, width=640, boxes=2, classes=3, seed=0):
    g = torch.Generator().manual_seed(seed)
    rgb = torch.randn(batch, 3, height, width, generator=g)
    evt = torch.randn(batch, 5, height, width, generator=g)
    ann = torch.zeros(batch, boxes, 5)
    for b in range(batch):
        for j in range(boxes):
            x, y = float(torch.rand(1, generator=g)) * width * 0.6, float(torch.rand(1, generator=g)) * height * 0.6
            w, h = 30 + float(torch.rand(1, generator=g)) * 150, 30 + float(torch.rand(1, generator=g)) * 120
            ann[b, j] = torch.tensor([x, y, x + w, y + h, float(torch.randint(0, classes, (1,), generator=g))])
    return rgb.to(device), evt.to(device), ann.to(device)
