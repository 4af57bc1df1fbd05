#!/usr/bin/env python
"""Row f1: the FPN lateral 1x1 conv that consumes the fused map (model.py:28,32,36,40), at the four
DSEC level shapes, batch 8:

  lateral alone  frn_b200.Conv1x1 (split TF32 on tcgen05, 1e-5 parity) vs cuDNN fp32 (TF32 off) and cuDNN TF32, fwd+bwd
  fused          frn_b200.REFusionLateral (the 2C map is neither written nor read) vs lateral(REFusion(...)), fwd+bwd

    python tools/bench_lateral.py [--steps K] [--mode unfused|fused]     # --mode: run ONE arm once (for ncu dram bytes)

One JSON line per level."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F

import frn_b200

DSEC = [(256, 120, 160), (512, 60, 80), (1024, 30, 40), (2048, 15, 20)]


def timed(fn, steps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--mode", default="")
    ap.add_argument("--level", type=int, default=-1)
    args = ap.parse_args()
    dev, B, cout = torch.device("cuda:0"), args.batch, 256
    for k, (c, h, w) in enumerate(DSEC):
        if args.level >= 0 and k != args.level:
            continue
        torch.manual_seed(k)
        fus = frn_b200.REFusion(c, c).to(dev)
        lat = frn_b200.Conv1x1(2 * c, cout).to(dev)
        fused = frn_b200.REFusionLateral(fus, lat)
        a = torch.randn(B, c, h, w, device=dev).relu().requires_grad_(True)
        e = torch.randn(B, c, h, w, device=dev).relu().requires_grad_(True)
        x = torch.randn(B, 2 * c, h, w, device=dev, requires_grad=True)
        go = torch.randn(B, cout, h, w, device=dev)

        def run_unfused():
            lat(fus(a, e)).backward(go)

        def run_fused():
            fused(a, e).backward(go)

        if args.mode:
            (run_fused if args.mode == "fused" else run_unfused)()
            torch.cuda.synchronize()
            print(f"ok {args.mode} level {k + 1}")
            continue

        def cudnn(tf32):
            torch.backends.cudnn.allow_tf32 = tf32
            torch.backends.cuda.matmul.allow_tf32 = tf32
            return timed(lambda: F.conv2d(x, lat.weight, lat.bias).backward(go), args.steps)

        row = {"level": k + 1, "shape": [B, c, h, w], "lateral_out_channels": cout,
               "lateral_fwd_bwd_ms": {"frn_b200.Conv1x1": timed(lambda: lat(x).backward(go), args.steps),
                                      "cudnn_fp32": cudnn(False), "cudnn_tf32": cudnn(True)},
               "fusion_plus_lateral_fwd_bwd_ms": {"unfused": timed(run_unfused, args.steps),
                                                  "fused": timed(run_fused, args.steps)},
               "map_bytes_not_moved_fwd": 2 * 2 * B * 2 * c * h * w * 4}
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
