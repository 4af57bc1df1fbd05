"""TEST INFRASTRUCTURE ONLY -- CPU restatement (NumPy) of FRN's event voxelisation.
Only tests/, __graft_entry__.smoke() and a bench's cpu_baseline leg may import this.

    voxelize              retinanet/data/dsec_data.py:347-387 (DSECDet.preprocess_events)
    normalize_event_data  retinanet/data/dsec_data.py:78-93

Pinned against the unmodified reference functions (their source is read from /root/reference and executed by
oracle/make_golden_events.py, because the module itself imports h5py / dsec_det, absent here): voxel grids
bit-identical on every fixture (tests/golden/events_*.npz, events_pin_report.json)."""
from __future__ import annotations

import numpy as np


def voxelize(events, C=5, H=480, W=640):
    """(C, H, W) float32.  The reference adds +-1 event by event; np.add.at performs the same unbuffered adds
    (sums of +-1 are exact in fp32, so the order is immaterial)."""
    x, y, t, p = (np.asarray(events[k]) for k in ("x", "y", "t", "p"))
    keep = (y < H) & (x < W)
    x, y, t, p = x[keep], y[keep], t[keep], p[keep]
    grid = np.zeros((C, H, W), dtype=np.float32)
    if t.size == 0:
        return grid
    window = t[-1] - t[0] + 1e-6
    tn = (t - t[0]) / window
    x, y = x.astype(np.int64), y.astype(np.int64)
    sign = np.where(p.astype(np.int64) > 0, np.float32(1), np.float32(-1))
    bins = np.clip((tn * (C - 1)).astype(np.int64), 0, C - 1)
    inside = (x >= 0) & (y >= 0)
    np.add.at(grid, (bins[inside], y[inside], x[inside]), sign[inside])
    return grid


def normalize_event_data(grid, method="tanh"):
    g = np.asarray(grid, dtype=np.float64)
    if method == "tanh":
        return np.tanh((np.asarray(grid, dtype=np.float32) / np.float32(5.0)).astype(np.float64))
    if method == "clip":
        return np.clip(g, -2, 2)
    raise ValueError(method)


def synthetic_events(n, seed, H=480, W=640, out_of_range=0.02, dtype=np.uint16):
    """DSEC-like synthetic events: sorted int64 microsecond timestamps, uint16 coordinates with a few outside
    the frame, uint8 polarity, a moving cluster so that some pixels collect many events."""
    rng = np.random.default_rng(seed)
    t = np.sort(rng.integers(1_000_000, 1_050_000, size=n)).astype(np.int64)
    cx = W * (0.3 + 0.4 * np.linspace(0, 1, n)) + rng.normal(0, 40, n)
    cy = H * 0.5 + rng.normal(0, 60, n)
    uniform = rng.random(n) < 0.5
    x = np.where(uniform, rng.integers(0, W, n), cx).round()
    y = np.where(uniform, rng.integers(0, H, n), cy).round()
    far = rng.random(n) < out_of_range
    x = np.where(far, W + rng.integers(0, 50, n), x)
    y = np.where(rng.random(n) < out_of_range, H + rng.integers(0, 50, n), y)
    lo = 0 if np.issubdtype(dtype, np.unsignedinteger) else -3
    return {"x": np.clip(x, lo, 65535).astype(dtype), "y": np.clip(y, lo, 65535).astype(dtype), "t": t,
            "p": (rng.random(n) < 0.5).astype(np.uint8) * 255 if dtype == np.uint16 else (rng.random(n) < 0.5).astype(np.uint8)}
