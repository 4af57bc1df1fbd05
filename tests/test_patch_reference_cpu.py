"""CPU: patch_reference() on the REAL reference package (skipped where /root/reference does not exist, i.e. on the
GPU box).  retinanet/model.py:10 imports REFusion by name and builds self.fus at model.py:247-252; after the patch
model.resnet50('dsec', 3, 'fpn_fusion') must expose exactly the reference's state-dict keys and shapes, in the same
order, so published checkpoints load unchanged."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"

_SCRIPT = r"""
import json, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {shim!r}); sys.path.insert(0, {ref!r})
import torch
patched = {patched}
if patched:
    import frn_b200
    done = frn_b200.patch_reference("retinanet")
from retinanet import model
torch.manual_seed(0)
net = model.resnet50('dsec', 3, 'fpn_fusion')
sd = net.state_dict()
info = dict(keys=[(k, list(v.shape), str(v.dtype)) for k, v in sd.items()],
            fus=[type(m).__module__ + '.' + type(m).__name__ for m in net.fus],
            anchors=type(net.anchors).__module__, loss=type(net.focalLoss).__module__)
if patched:
    # a checkpoint written by the unmodified reference loads strictly
    net.load_state_dict({{k: torch.zeros_like(v) for k, v in sd.items()}}, strict=True)
    info['rebound'] = done
print('JSON' + json.dumps(info))
"""


def _run(patched):
    code = _SCRIPT.format(root=ROOT, shim=os.path.join(ROOT, "oracle", "fvcore_shim"), ref=REF, patched=patched)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("JSON")][-1]
    return json.loads(line[4:])


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "retinanet")), reason="/root/reference is only present in the build container")
def test_patch_reference_on_the_real_package():
    ref, ours = _run(False), _run(True)
    assert len(ref["keys"]) > 700
    assert ours["keys"] == ref["keys"]                       # same keys, shapes, dtypes, same order
    assert all(t == "retinanet.fusion.REFusion" for t in ref["fus"])
    assert all(t == "frn_b200.modules.REFusion" for t in ours["fus"])
    assert ours["anchors"].startswith("frn_b200") and ours["loss"].startswith("frn_b200")
    assert ["retinanet.model", "REFusion"] in ours["rebound"]
