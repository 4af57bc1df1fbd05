"""GPU: event voxelisation kernels through frn_b200.voxelize_events / preprocess_events and
the C ABI.  Bar: the voxel grid is BIT-IDENTICAL to the reference's sequential loop (integer / index work) on the
fixtures of the unmodified reference and against the oracle at full size; tanh normalisation within 2e-7."""
import glob
import os

import numpy as np
import pytest
import torch

import frn_b200
from helpers import GOLDEN
from oracle import events_numpy as oev
from test_oracle_events import dense

pytestmark = pytest.mark.gpu
FILES = sorted(glob.glob(os.path.join(GOLDEN, "events_*.npz")))


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p)[:-4] for p in FILES])
def test_fixture_bit_identical(path):
    z = np.load(path)
    ev = {k: z[k] for k in "xytp"}
    out = frn_b200.preprocess_events(ev)["img"]
    assert out.shape == (5, 480, 640) and out.dtype == torch.float32 and out.is_cuda
    assert np.array_equal(out.cpu().numpy(), dense(z))
    t = frn_b200.voxelize_events(ev, normalize="tanh").cpu().numpy()
    assert np.abs(t - dense(z, "nz_tanh")).max() <= 2e-7
    assert np.abs(frn_b200.normalize_event_data(out, "tanh").cpu().numpy() - dense(z, "nz_tanh")).max() <= 2e-7
    assert np.array_equal(frn_b200.normalize_event_data(out, "clip").cpu().numpy(), np.clip(dense(z), -2, 2))


@pytest.mark.parametrize("n,dtype", [(1_000_000, np.uint16), (250_000, np.int32), (0, np.uint16)],
                         ids=["1M_u16", "250k_i32", "empty"])
def test_full_size_vs_oracle(n, dtype):
    ev = oev.synthetic_events(n, seed=n % 97, dtype=dtype) if n else {
        "x": np.zeros(0, np.uint16), "y": np.zeros(0, np.uint16), "t": np.zeros(0, np.int64), "p": np.zeros(0, np.uint8)}
    if dtype == np.int32 and n:
        ev["x"][0] = -5; ev["y"][-1] = -1
    want = oev.voxelize(ev)
    got = frn_b200.voxelize_events(ev).cpu().numpy()
    assert np.array_equal(got, want)
    # bit-reproducible, device-resident inputs give the same grid, other grid shapes work
    dev = {k: torch.from_numpy(v).cuda() for k, v in ev.items()}
    assert np.array_equal(frn_b200.voxelize_events(dev).cpu().numpy(), want)
    assert np.array_equal(frn_b200.voxelize_events(ev, C_bins=3, H=260, W=346).cpu().numpy(),
                          oev.voxelize(ev, C=3, H=260, W=346))


def test_conservation_at_full_size():
    ev = oev.synthetic_events(2_000_000, seed=5)
    g = frn_b200.voxelize_events(ev)
    keep = (ev["y"] < 480) & (ev["x"] < 640)
    assert float(g.double().sum()) == float(np.where(ev["p"][keep] > 0, 1, -1).sum())
