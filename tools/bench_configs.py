#!/usr/bin/env python
"""Secondary bench lines for the BASELINE.json configs that bench.py's headline does not cover (one GPU):

  config 4: DDD17-shaped 346x260 inference, batch=32, bf16 fusion kernels (forward only, no_grad)
  ddd17-train: DDD17-shaped fp32 fwd+bwd, batch=8 (north_star: "DSEC-shaped and DDD17-shaped ... pyramids")
  config 1 on the GPU: one REFusion(256) block fwd+bwd, batch=1, DSEC level 1 (the reference's CPU case)

    python tools/bench_configs.py [--steps K] [--warmup W]

Prints one JSON line per config.  Timing: CUDA events on the current stream around K steps after W warm-up
steps; inputs resident in HBM (`value`) and from pinned host memory with the D2H of a checksum (`e2e`)."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

DSEC = [(256, 120, 160), (512, 60, 80), (1024, 30, 40), (2048, 15, 20)]
DDD17 = [(256, 65, 87), (512, 33, 44), (1024, 17, 22), (2048, 9, 11)]


def timed(fn, steps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def run(name, levels, B, dtype, train, steps, warmup, peaks):
    import frn_b200
    from frn_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    mods = [frn_b200.REFusion(c, c).to(dev) for c, _, _ in levels]
    if not train:
        for m in mods:
            m.eval()
    host, devin, douts = [], [], []
    for k, (c, h, w) in enumerate(levels):
        g = torch.Generator().manual_seed(1000 + k)
        a = torch.randn(B, c, h, w, generator=g).relu().to(dtype)
        e = torch.randn(B, c, h, w, generator=g).relu().to(dtype)
        host.append((a.pin_memory(), e.pin_memory()))
        devin.append((a.to(dev), e.to(dev)))
        douts.append(torch.randn(B, 2 * c, h, w, generator=g).to(dev) if train else None)
    res = torch.empty(1, dtype=torch.float32).pin_memory()

    streams = [torch.cuda.Stream() for _ in levels] if len(levels) > 1 else None

    def one(m, a, e, go):
        if train:
            a = a.detach().requires_grad_(True); e = e.detach().requires_grad_(True)
            m(a, e).backward(go)
            return a.grad
        with torch.no_grad():
            return m(a, e)

    def step(inputs):
        """The independent levels on their own streams (forked from / joined into the current one), as bench.py."""
        acc = None
        if streams is None:
            for m, (a, e), go in zip(mods, inputs, douts):
                acc = one(m, a, e, go)
            return acc
        main = torch.cuda.current_stream()
        for s, m, (a, e), go in zip(streams, mods, inputs, douts):
            s.wait_stream(main)
            with torch.cuda.stream(s):
                acc = one(m, a, e, go)
        for s in streams:
            main.wait_stream(s)
        return acc

    def step_resident():
        step(devin)

    def step_e2e():
        inputs = [(a.to(dev, non_blocking=True), e.to(dev, non_blocking=True)) for a, e in host]
        acc = step(inputs)
        res.copy_(acc.float().sum().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(max(warmup, 3)):
        step_resident()
    n0 = lib.frn_launch_count()
    ms = timed(step_resident, steps)
    launches = lib.frn_launch_count() - n0
    # the same step replayed from ONE captured CUDA graph (the library only enqueues: no allocation, no
    # synchronisation, side streams fork from / join into the capturing stream), i.e. without the ~1300 - 4600 launch
    # calls per step that bound these small shapes
    ms_graph, graph_err = None, None
    try:
        static_in = [(a.clone(), e.clone()) for a, e in devin]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                step(static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            step(static_in)
        graph.replay()
        ms_graph = timed(graph.replay, steps)
    except Exception as ex:      # reported, not hidden
        graph_err = f"{type(ex).__name__}: {ex}"[:200]
        torch.cuda.synchronize()
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, steps)
    esz = 2 if dtype == torch.bfloat16 else 4
    sum_nc = sum(c * h * w for c, h, w in levels)
    n2c = sum((h * w) ** 2 * c for c, h, w in levels)
    nc2 = sum(h * w * c * c for c, h, w in levels)
    flops = (6 * nc2 + n2c) + ((12 * nc2 + 2.5 * n2c) if train else 0)
    hbm = (10 if train else 4) * esz * sum_nc
    sps = B / (ms / 1e3)
    print(json.dumps({
        "config": name, "metric": ("fusion-block fwd+bwd samples/s" if train else "fusion-block forward samples/s"),
        "value": sps, "unit": "samples/s", "ms_per_step": ms, "steps": steps, "warmup": max(warmup, 3),
        "dtype": "bf16 I/O, single-pass tf32 products, fp32 accumulation" if esz == 2 else "f32 (split tf32)",
        "levels": [[B, c, h, w] for c, h, w in levels], "gpu_launches": int(launches),
        "cuda_graph": ({"value": B / (ms_graph / 1e3), "unit": "samples/s", "ms_per_step": ms_graph}
                       if ms_graph else {"failed": graph_err}),
        "e2e": {"value": B / (ms_e2e / 1e3), "unit": "samples/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": 2 * B * esz * sum_nc, "d2h_bytes_per_step": 4},
        "block_hbm_frac": hbm * sps / 1e9 / peaks["hbm_gbs"],
        "block_tensor_frac": flops * sps / 1e12 / peaks["bf16_tflops_sustained"],
        "algorithmic": {"hbm_bytes_per_sample": hbm, "flops_per_sample": flops},
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--only", type=int, default=-1, help="run only the i-th config (0-based)")
    args = ap.parse_args()
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        peaks = {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0}
    configs = [("config 4: DDD17 346x260 inference, batch=32, bf16", DDD17, 32, torch.bfloat16, False),
               ("DDD17 346x260 fp32 inference, batch=32", DDD17, 32, torch.float32, False),
               ("DDD17 346x260 fp32 fwd+bwd, batch=8", DDD17, 8, torch.float32, True),
               ("config 1 on the GPU: REFusion(256) fwd+bwd, batch=1, DSEC level 1", DSEC[:1], 1, torch.float32, True)]
    for i, (name, levels, batch, dtype, train) in enumerate(configs):
        if args.only < 0 or args.only == i:
            run(name, levels, batch, dtype, train, args.steps, args.warmup, peaks)


if __name__ == "__main__":
    main()
