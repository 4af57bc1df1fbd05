import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def pytest_sessionfinish(session, exitstatus):
    """Measured errors of every block-level parity gate (tests/helpers.check_against) -> gpurun_out/parity_report.json,
    so that the margins under the 1e-5 gate are on record, not just pass / fail."""
    try:
        import json
        import helpers
        if not helpers.PARITY_LOG:
            return
        out_dir = os.path.join(ROOT, "gpurun_out")
        os.makedirs(out_dir, exist_ok=True)
        rows = [{"case": n, "worst": max(e.values()), "worst_quantity": max(e, key=e.get), "errors": e}
                for n, e in helpers.PARITY_LOG]
        rows.sort(key=lambda r: -r["worst"])
        with open(os.path.join(out_dir, "parity_report.json"), "w") as f:
            json.dump({"gate": 1e-5, "metric": "l2-relative error vs the fp64 oracle (theta.bias: scaled absolute bound)",
                       "cases": rows}, f, indent=1)
    except Exception:
        pass
