#!/usr/bin/env python
"""Row f2 measurement: anchors + focal loss fwd+bwd at the DSEC head shapes, one GPU.

    python tools/bench_detection.py [--steps K] [--warmup W] [--batch B] [--classes K] [--boxes M] [--no-cpu]

One step = FocalLoss forward + backward over a batch of synthetic head outputs for a 480x640 image
(A = 230 220 anchors).  Prints one JSON line: samples/s with inputs resident in HBM (`value`), end to end with
host buffers (`e2e`), the HBM roofline of the two kernels (algorithmic bytes / CUDA-event time), and the CPU port
(oracle/detection_numpy.py, the reference's per-sample loop restated in NumPy) timed on the host cores."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--classes", type=int, default=3)
    ap.add_argument("--boxes", type=int, default=16)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import frn_b200
    from frn_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    B, K, M, H, W = args.batch, args.classes, args.boxes, 480, 640
    image = torch.zeros(B, 3, H, W, device=dev)
    anchors_mod, loss_mod = frn_b200.Anchors(), frn_b200.FocalLoss()
    anchors = anchors_mod(image)
    A = anchors.shape[1]
    g = torch.Generator().manual_seed(0)
    cls_h = torch.sigmoid(torch.randn(B, A, K, generator=g) - 2.0).pin_memory()
    reg_h = (0.3 * torch.randn(B, A, 4, generator=g)).pin_memory()
    ann_h = torch.full((B, M, 5), -1.0)
    for b in range(B):
        for j in range(int(torch.randint(1, M + 1, (1,), generator=g))):
            xy = torch.rand(2, generator=g) * torch.tensor([W * 0.8, H * 0.8])
            wh = 8 + torch.rand(2, generator=g) * torch.tensor([W * 0.3, H * 0.3])
            ann_h[b, j] = torch.cat([xy, xy + wh, torch.randint(0, K, (1,), generator=g).float()])
    ann_h = ann_h.pin_memory()
    cls_d, reg_d, ann_d = cls_h.to(dev), reg_h.to(dev), ann_h.to(dev)
    res = torch.empty(2, dtype=torch.float32).pin_memory()

    def step(c, r, a):
        c = c.detach().requires_grad_(True); r = r.detach().requires_grad_(True)
        lc, lr = loss_mod(c, r, anchors_mod(image), a)
        (lc + lr).sum().backward()
        return lc, lr

    def step_e2e():
        lc, lr = step(cls_h.to(dev, non_blocking=True), reg_h.to(dev, non_blocking=True),
                      ann_h.to(dev, non_blocking=True))
        res.copy_(torch.cat([lc, lr]).detach(), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def timed(fn, n):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    for _ in range(max(args.warmup, 3)):
        step(cls_d, reg_d, ann_d)
    n0 = lib.frn_launch_count()
    ms = timed(lambda: step(cls_d, reg_d, ann_d), args.steps)
    launches = lib.frn_launch_count() - n0
    for _ in range(3):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)

    # kernel-only time: the library's row-kernel timing class brackets the focal kernels with CUDA events
    import ctypes as C
    lib.frn_timing_enable(1)
    t, n = C.c_double(), C.c_long()
    lib.frn_timing_read(3, C.byref(t), C.byref(n))
    for _ in range(args.steps):
        step(cls_d, reg_d, ann_d)
    torch.cuda.synchronize()
    lib.frn_timing_read(3, C.byref(t), C.byref(n))
    lib.frn_timing_enable(0)
    kernel_ms = t.value / args.steps
    # algorithmic bytes per step: fwd reads cls + reg, writes the assignment; bwd reads cls + reg + assignment and
    # writes both gradients (anchors / annotations: 3.7 MB, L2-resident across the batch)
    bytes_step = B * A * ((K + 4) * 4 * 3 + 8) + 2 * A * 16

    # uncached anchor generation (what the reference redoes every step in NumPy + H2D)
    def gen():
        anchors_mod._cache.clear()
        anchors_mod(image)
    for _ in range(3):
        gen()
    ms_anchors = timed(gen, 20)

    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak, src = peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json), copy bandwidth"
    except OSError:
        peak, src = 6650.0, "fallback (B200_PROFILING.md)"
    cpu = None
    if not args.no_cpu:
        from oracle import detection_numpy as odet
        cn, rn, an, nn_ = cls_h.numpy(), reg_h.numpy(), anchors.cpu().numpy(), ann_h.numpy()
        t0 = time.perf_counter()
        odet.anchors_for_image(H, W)
        t_anchor_cpu = time.perf_counter() - t0
        t0 = time.perf_counter()
        odet.focal_loss(cn[:2], rn[:2], an, nn_[:2])
        sec = (time.perf_counter() - t0) / 2
        cpu = {"value": 1.0 / sec, "unit": "samples/s", "cores": 1, "kind": "port",
               "sample": "2 samples of the same batch, fwd + gradients, oracle/detection_numpy.py (fp64 NumPy, 1 thread)",
               "anchors_ms": t_anchor_cpu * 1e3}
    ach = bytes_step / (kernel_ms / 1e3) / 1e9
    print(json.dumps({
        "metric": "focal loss fwd+bwd samples/s at DSEC 640x480 heads (row f2)", "value": B / (ms / 1e3),
        "unit": "samples/s", "ms_per_step": ms, "steps": args.steps, "warmup": max(args.warmup, 3), "dtype": "f32",
        "config": {"workload": f"FocalLoss fwd+bwd, batch={B}, A={A} anchors, K={K} classes, M={M} annotation rows, "
                               "anchors cached on the device", "data": "synthetic"},
        "gpu_launches": int(launches),
        "e2e": {"value": B / (ms_e2e / 1e3), "unit": "samples/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(cls_h.numel() + reg_h.numel() + ann_h.numel()) * 4, "d2h_bytes_per_step": 8},
        "roofline": {"bound": "hbm", "kernel": "focal_forward_kernel + focal_backward_kernel", "achieved": ach,
                     "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": src,
                     "bytes_per_step": bytes_step, "kernel_ms_per_step": kernel_ms},
        "anchors": {"count": int(A), "generate_ms_uncached": ms_anchors},
        "cpu_baseline": cpu,
    }))


if __name__ == "__main__":
    main()
