#!/usr/bin/env python
"""BASELINE.json config 3: end-to-end FRN-shaped RetinaNet training step (dual ResNet-50 trunks, four fusion blocks,
FPN, heads, anchors, focal loss, SGD step) on a synthetic DSEC 480x640 batch with random-init weights, one GPU.

    python tools/bench_train_step.py [--batch B] [--steps K] [--warmup W] [--impl ours|eager|both]

The network is the harness in tools/harness_model.py (the reference's own model.py cannot travel to the GPU box).
`ours`: frn_b200 modules at the fusion / lateral / anchors / loss positions.  `eager`: the same arithmetic as stock
PyTorch ops on the GPU (what the unmodified reference executes on a GPU; fp32, TF32 off).  One JSON line per impl:
samples/s, ms per step and the share of the step spent in the fusion blocks (CUDA events around them)."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch  # noqa: E402


def run(impl, args):
    from harness_model import FusionDetector, calibrate_batchnorm, synthetic_batch
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    net = FusionDetector(num_classes=3, impl=impl).to(dev).train()
    opt = torch.optim.SGD([p for p in net.parameters() if p.requires_grad], lr=1e-4, momentum=0.9)
    rgb, evt, ann = synthetic_batch(args.batch, dev)
    calibrate_batchnorm(net, rgb, evt)
    marks = []

    def pre(_m, _i):
        e = torch.cuda.Event(enable_timing=True); e.record(); marks.append(e)

    def post(_m, _i, _o):
        e = torch.cuda.Event(enable_timing=True); e.record(); marks.append(e)

    def step():
        opt.zero_grad(set_to_none=True)
        lc, lr = net(rgb, evt, ann)
        loss = lc.mean() + lr.mean()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(net.parameters(), 0.1)          # train_dsec.py:169
        opt.step()
        return loss

    for _ in range(max(args.warmup, 3)):
        loss = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    # forward share of the fusion blocks (hooks only in this extra pass)
    hooks = [h for f in net.fus for h in (f.register_forward_pre_hook(pre), f.register_forward_hook(post))]
    step()
    torch.cuda.synchronize()
    fus_fwd_ms = sum(marks[i].elapsed_time(marks[i + 1]) for i in range(0, len(marks), 2))
    for h in hooks:
        h.remove()
    print(json.dumps({
        "config": "BASELINE.json configs[2]: end-to-end train step, harness of the reference's ResNet-50 fpn_fusion "
                  "network, synthetic DSEC 480x640 batch, random init, fp32", "impl": impl, "batch": args.batch,
        "value": args.batch / (ms / 1e3), "unit": "samples/s", "ms_per_step": ms, "steps": args.steps,
        "warmup": max(args.warmup, 3), "loss": float(loss), "fusion_forward_ms": fus_fwd_ms,
        "peak_memory_gb": torch.cuda.max_memory_allocated() / 1e9,
        "parameters_m": sum(p.numel() for p in net.parameters()) / 1e6}), flush=True)
    del net, opt
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=2)       # train_dsec_det_fast.py:359 (the reference default is 1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="both", choices=["ours", "eager", "both"])
    args = ap.parse_args()
    for impl in (("ours", "eager") if args.impl == "both" else (args.impl,)):
        run(impl, args)


if __name__ == "__main__":
    main()
