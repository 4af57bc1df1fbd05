#!/usr/bin/env python
"""Row f3 measurement: inference post-processing (decode + clip + per-class NMS), one GPU.

    python tools/bench_postprocess.py [--steps K] [--warmup W] [--hot F] [--no-cpu]

One step = one image (the reference evaluates at batch 1): A = 230 220 anchors of a 480x640 frame, K = 3 classes.
Prints one JSON line: images/s with head outputs resident in HBM (`value`; includes the 4-byte D2H read of the
detection count that sizes the result), `e2e` from pinned host arrays, and the CPU port (oracle NumPy)."""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--hot", type=float, default=0.0005, help="controls the fraction of scores above 0.05")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import frn_b200
    from frn_b200 import _lib
    from oracle import postprocess_numpy as opp
    lib = _lib.load()
    dev = torch.device("cuda:0")
    H, W, K = 480, 640, 3
    img = torch.zeros(1, 3, H, W, device=dev)
    anchors = frn_b200.Anchors()(img)
    A = anchors.shape[1]
    cls, reg = opp.synthetic_heads(A, K, seed=0, hot=args.hot)
    cls_h, reg_h = torch.from_numpy(cls)[None].pin_memory(), torch.from_numpy(reg)[None].pin_memory()
    cls_d, reg_d = cls_h.to(dev), reg_h.to(dev)

    def step():
        return frn_b200.detect_postprocess(cls_d, reg_d, anchors, img)

    def step_e2e():
        s, c, b = frn_b200.detect_postprocess(cls_h.to(dev, non_blocking=True), reg_h.to(dev, non_blocking=True), anchors, img)
        return s.cpu(), c.cpu(), b.cpu()

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
        out = step()
    n0 = lib.frn_launch_count()
    ms = timed(step, args.steps)
    launches = lib.frn_launch_count() - n0
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    lib.frn_timing_enable(1)
    t, n = C.c_double(), C.c_long()
    lib.frn_timing_read(3, C.byref(t), C.byref(n))
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize()
    lib.frn_timing_read(3, C.byref(t), C.byref(n))
    lib.frn_timing_enable(0)
    cpu = None
    if not args.no_cpu:
        t0 = time.perf_counter()
        opp.postprocess(cls, reg, anchors[0].cpu().numpy(), H, W)
        sec = time.perf_counter() - t0
        cpu = {"value": 1.0 / sec, "unit": "images/s", "cores": 1, "kind": "port",
               "sample": "one image, oracle/postprocess_numpy.py (fp32 NumPy decode + greedy NMS per class)"}
    print(json.dumps({
        "metric": "inference post-processing images/s at DSEC 640x480 (row f3)", "value": 1e3 / ms, "unit": "images/s",
        "ms_per_step": ms, "steps": args.steps, "warmup": max(args.warmup, 3), "dtype": "f32",
        "config": {"workload": f"decode + clip + per-class NMS, A={A}, K={K}, {int((cls > 0.05).sum())} scores above "
                               f"0.05, {int(out[0].numel())} detections kept", "data": "synthetic"},
        "gpu_launches": int(launches), "library_launches": "3 cub radix sorts per step (not counted)",
        "e2e": {"value": 1e3 / ms_e2e, "unit": "images/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(cls_h.numel() + reg_h.numel()) * 4,
                "d2h_bytes_per_step": int(out[0].numel()) * 28 + 4},
        "kernel_ms_per_step": t.value / args.steps,
        "cpu_baseline": cpu,
    }))


if __name__ == "__main__":
    main()
