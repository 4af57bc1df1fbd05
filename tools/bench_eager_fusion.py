#!/usr/bin/env python
"""The only pre-existing GPU implementation of the path: the reference's ATen call sequence run by PyTorch eager on
the B200 (cuDNN / cuBLAS kernels, N x N attention matrices materialised), timed on the headline workload
(BASELINE.json configs[1]: four DSEC levels, batch 8, fwd+bwd) next to frn_b200.

    python tools/bench_eager_fusion.py [--batch B] [--steps K] [--warmup W]

One JSON line per arm: eager fp32 (TF32 off: the parity-comparable arm), eager with TF32 allowed (1e-3 accuracy),
and frn_b200 (fp32 parity <= 1e-5).  tools/harness_model.EagerREFusion is the stock-op restatement."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch  # noqa: E402

DSEC = [(256, 120, 160), (512, 60, 80), (1024, 30, 40), (2048, 15, 20)]


def run(name, make, tf32, args):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.allow_tf32 = tf32
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    mods = [make(c, c).to(dev) for c, _, _ in DSEC]
    data = []
    for k, (c, h, w) in enumerate(DSEC):
        g = torch.Generator().manual_seed(1000 + k)
        a = torch.randn(args.batch, c, h, w, generator=g).relu().to(dev).requires_grad_(True)
        e = torch.randn(args.batch, c, h, w, generator=g).relu().to(dev).requires_grad_(True)
        data.append((a, e, torch.randn(args.batch, 2 * c, h, w, generator=g).to(dev)))

    def step():
        for m, (a, e, go) in zip(mods, data):
            a.grad = None; e.grad = None
            for p in m.parameters():
                p.grad = None
            m(a, e).backward(go)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    print(json.dumps({"arm": name, "metric": "fusion-block fwd+bwd samples/s at DSEC 640x480", "batch": args.batch,
                      "value": args.batch / (ms / 1e3), "unit": "samples/s", "ms_per_step": ms, "steps": args.steps,
                      "warmup": max(args.warmup, 3), "peak_memory_gb": torch.cuda.max_memory_allocated() / 1e9}), flush=True)
    del mods, data
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-ours", action="store_true", help="only the two eager arms")
    args = ap.parse_args()
    import frn_b200
    from harness_model import EagerREFusion
    run("PyTorch eager on the GPU, fp32 (TF32 off)", EagerREFusion, False, args)
    run("PyTorch eager on the GPU, TF32 allowed", EagerREFusion, True, args)
    if not args.no_ours:
        run("frn_b200 (fp32 parity, split TF32 on tcgen05)", frn_b200.REFusion, False, args)


if __name__ == "__main__":
    main()
