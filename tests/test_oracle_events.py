"""CPU: oracle/events_numpy.py (event voxelisation) against the fixtures produced from the
UNMODIFIED reference (oracle/make_golden_events.py), plus host-side argument checks of the drop-in."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN
from oracle import events_numpy as oev

FILES = sorted(glob.glob(os.path.join(GOLDEN, "events_*.npz")))


def dense(z, key="nz_value", shape=(5, 480, 640)):
    g = np.zeros(shape, dtype=z[key].dtype)
    g[tuple(z["nz_index"])] = z[key]
    return g


def test_fixtures_present():
    assert len(FILES) == 5
    rep = json.load(open(os.path.join(GOLDEN, "events_pin_report.json")))
    assert all(v["oracle_bit_identical"] for v in rep.values())


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p)[:-4] for p in FILES])
def test_oracle_bit_identical_to_reference_fixture(path):
    z = np.load(path)
    grid = oev.voxelize({k: z[k] for k in "xytp"})
    assert np.array_equal(grid, dense(z))
    assert np.abs(oev.normalize_event_data(grid, "tanh") - dense(z, "nz_tanh")).max() < 1e-7


def test_properties():
    ev = oev.synthetic_events(20000, seed=11)
    g = oev.voxelize(ev)
    keep = (ev["y"] < 480) & (ev["x"] < 640)
    sign = np.where(ev["p"][keep] > 0, 1, -1)
    assert g.sum() == sign.sum()                                   # every in-range event lands exactly once
    perm = np.random.default_rng(0).permutation(np.nonzero(keep)[0][1:-1])      # order inside the window is immaterial
    idx = np.concatenate([np.nonzero(keep)[0][:1], perm, np.nonzero(keep)[0][-1:]])
    assert np.array_equal(oev.voxelize({k: v[idx] for k, v in ev.items()}), g)


def test_host_side_checks():
    import frn_b200
    from frn_b200 import _lib
    ev = oev.synthetic_events(10, seed=1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        frn_b200.voxelize_events(ev, device="cpu")
    with pytest.raises(ValueError):
        frn_b200.voxelize_events(ev, normalize="minmax")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        frn_b200.normalize_event_data(torch.zeros(5, 4, 4))
    lib = _lib.load()
    assert lib.frn_voxelize_workspace_bytes() >= 8
    assert lib.frn_voxelize_events(None, -1, 2, None, None, None, None, 5, 480, 640, 0, None, None, 0) == -1
    assert lib.frn_voxelize_events(None, 10, 3, None, None, None, None, 5, 480, 640, 0, None, None, 0) == -4
    assert lib.frn_voxelize_events(None, 10, 2, None, None, None, None, 5, 480, 640, 0, None, None, 0) == -2
    assert lib.frn_voxel_normalize(None, 10, 3, None) == -4
