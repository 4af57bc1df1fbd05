#!/usr/bin/env python
"""Row f4 measurement: event voxelisation, one GPU.

    python tools/bench_events.py [--events N] [--steps K] [--warmup W] [--no-cpu]

One step = one sample: N synthetic DSEC-like events -> (5, 480, 640) grid + tanh normalisation.  Prints one JSON
line: events/s with the event arrays resident in HBM (`value`), end to end from pinned host arrays with a D2H of
the grid's checksum (`e2e`), the HBM roofline of the pass (13 B per event read + the grid cleared, scattered
into and normalised), and the CPU port (oracle/events_numpy.py; np.add.at) on the host."""
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
    ap.add_argument("--events", type=int, default=1_000_000)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import frn_b200
    from frn_b200 import _lib
    from oracle import events_numpy as oev
    lib = _lib.load()
    ev = oev.synthetic_events(args.events, seed=0)
    host = {k: torch.from_numpy(v).pin_memory() for k, v in ev.items()}
    dev = {k: v.cuda() for k, v in host.items()}
    res = torch.empty(1, dtype=torch.float32).pin_memory()

    def step():
        return frn_b200.voxelize_events(dev, normalize="tanh")

    def step_e2e():
        g = frn_b200.voxelize_events(host, normalize="tanh")
        res.copy_(g.sum().reshape(1), non_blocking=True)
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
        step()
    n0 = lib.frn_launch_count()
    ms = timed(step, args.steps)
    launches = lib.frn_launch_count() - n0
    for _ in range(3):
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
    kernel_ms = t.value / args.steps
    grid_bytes = 5 * 480 * 640 * 4
    # bounds pass reads x, y (4 B); scatter reads x, y, t, p (13 B) and adds 4 B per event; the grid is cleared
    # once (write) and normalised once (read + write)
    bytes_step = args.events * (4 + 13 + 4) + 3 * grid_bytes
    try:
        peak, src = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except OSError:
        peak, src = 6650.0, "fallback (B200_PROFILING.md)"
    cpu = None
    if not args.no_cpu:
        t0 = time.perf_counter()
        oev.normalize_event_data(oev.voxelize(ev), "tanh")
        sec = time.perf_counter() - t0
        cpu = {"value": args.events / sec, "unit": "events/s", "cores": 1, "kind": "port",
               "sample": f"one sample of {args.events} events, oracle/events_numpy.py (np.add.at instead of the "
                         "reference's per-event Python loop, which is ~10x slower still)"}
    ach = bytes_step / (kernel_ms / 1e3) / 1e9
    print(json.dumps({
        "metric": "event voxelisation events/s (row f4)", "value": args.events / (ms / 1e3), "unit": "events/s",
        "samples_per_s": 1e3 / ms, "ms_per_step": ms, "steps": args.steps, "warmup": max(args.warmup, 3),
        "dtype": "u16/i64/u8 -> f32 counts", "gpu_launches": int(launches),
        "config": {"workload": f"{args.events} synthetic events -> (5, 480, 640) voxel grid + tanh", "data": "synthetic"},
        "e2e": {"value": args.events / (ms_e2e / 1e3), "unit": "events/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": args.events * 13, "d2h_bytes_per_step": 4},
        "roofline": {"bound": "hbm", "kernel": "event_bounds_kernel + event_scatter_kernel + voxel_normalize_kernel",
                     "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                     "peak_source": src, "bytes_per_step": bytes_step, "kernel_ms_per_step": kernel_ms},
        "cpu_baseline": cpu,
    }))


if __name__ == "__main__":
    main()
