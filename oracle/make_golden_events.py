"""Generate tests/golden/events_*.npz from the UNMODIFIED reference voxelisation.  Build container only:

    python oracle/make_golden_events.py

TEST INFRASTRUCTURE ONLY.  retinanet/data/dsec_data.py cannot be imported here (h5py, dsec_det, torch_geometric
are absent), so the source text of `DSECDet.preprocess_events` and `normalize_event_data` is cut out of the
reference file with `ast` and executed as it stands (with numpy / torch in scope and a stand-in `self` whose
`debug` is False).  Nothing of it is stored in this repository -- only the resulting grids."""
from __future__ import annotations

import ast
import json
import os
import sys
import textwrap
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REFERENCE = os.environ.get("FRN_REFERENCE", "/root/reference")
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)
from oracle import events_numpy as oev  # noqa: E402


def reference_functions():
    path = os.path.join(REFERENCE, "retinanet", "data", "dsec_data.py")
    src = open(path).read()
    tree = ast.parse(src)
    ns = {"np": np, "torch": torch}
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in ("preprocess_events", "normalize_event_data"):
            exec(textwrap.dedent(ast.get_source_segment(src, node)), ns)
    self = types.SimpleNamespace(debug=False)
    return (lambda ev: ns["preprocess_events"](self, ev)["img"].numpy()), ns["normalize_event_data"]


def main():
    voxelize_ref, normalize_ref = reference_functions()
    report = {}
    cases = {
        "events_n3000_u16": oev.synthetic_events(3000, seed=1),
        "events_n2000_i32_negative": oev.synthetic_events(2000, seed=2, dtype=np.int32),
        "events_hot_pixel": None, "events_all_out_of_range": None, "events_single": None,
    }
    ev = cases["events_n2000_i32_negative"]                               # negative coordinates pass the mask (:353),
    ev["x"][0] = -2; ev["x"][-1] = -1; ev["y"][5] = -3                   # set the time window, and are skipped at :380
    ev = oev.synthetic_events(1500, seed=3)
    ev["x"][:] = 17; ev["y"][:] = 23                                     # every event on one pixel
    cases["events_hot_pixel"] = ev
    ev = oev.synthetic_events(50, seed=4)
    ev["x"][:] = 700                                                     # x >= W everywhere -> empty grid
    cases["events_all_out_of_range"] = ev
    ev = oev.synthetic_events(1, seed=5, out_of_range=0.0)               # t_max == t_min: window = 1e-6
    cases["events_single"] = ev
    for name, ev in cases.items():
        ref = voxelize_ref({k: v.copy() for k, v in ev.items()})
        mine = oev.voxelize(ev)
        tanh_ref = normalize_ref(torch.tensor(ref), "tanh").numpy()
        rep = {"n": int(ev["t"].size), "nonzero": int((ref != 0).sum()), "sum_abs": float(np.abs(ref).sum()),
               "oracle_bit_identical": bool(np.array_equal(mine, ref)),
               "tanh_max_abs_diff": float(np.abs(oev.normalize_event_data(mine, "tanh") - tanh_ref).max())}
        print(name, rep)
        assert rep["oracle_bit_identical"] and rep["tanh_max_abs_diff"] < 1e-7
        report[name] = rep
        idx = np.nonzero(ref)
        np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), x=ev["x"], y=ev["y"], t=ev["t"], p=ev["p"],
                            nz_index=np.stack(idx).astype(np.int32), nz_value=ref[idx],
                            nz_tanh=tanh_ref[idx])
    ev = oev.synthetic_events(200_000, seed=6)                            # larger pin, not stored
    ok = bool(np.array_equal(oev.voxelize(ev), voxelize_ref(ev)))
    report["pin_n200000"] = {"oracle_bit_identical": ok}
    assert ok
    json.dump(report, open(os.path.join(GOLDEN, "events_pin_report.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
