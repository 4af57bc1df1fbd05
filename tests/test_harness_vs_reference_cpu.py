"""CPU: BASELINE.json config 3 is run on the GPU box through tools/harness_model.py (a network assembled from stock
layers, because /root/reference does not exist there).  This test ties that harness to the reference's OWN
retinanet/model.py (skipped where /root/reference is absent): the same weights loaded into
model.resnet50('dsec', 3, 'fpn_fusion') (unpatched, model.py:170-364) and into FusionDetector(impl='eager') give the same
training losses and the same gradients on a small synthetic batch, so `tests/test_train_step_gpu.py` (harness eager
vs harness with the frn_b200 modules, on the B200) is a statement about the reference network, not about a look-alike."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"

_SCRIPT = r"""
import json, random, re, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {tools!r}); sys.path.insert(0, {shim!r}); sys.path.insert(0, {ref!r})
import torch
torch.manual_seed(0)
torch.set_num_threads(8)
from retinanet import model as ref_model
import harness_model as hm

H, W = 128, 192
rgb, evt, ann = hm.synthetic_batch(2, 'cpu', height=H, width=W, boxes=2, seed=3)
ann[:, :, :4] *= 0.3                                   # keep the boxes inside the small image
ann[:, :, 2:4] += 24

ours = hm.FusionDetector(3, impl='eager')
hm.calibrate_batchnorm(ours, rgb, evt)                 # sane activation scales through the random-init trunks
for name, p in ours.named_parameters():                # non-zero head outputs and attention biases
    if name.endswith('output.weight') or ('cross_attention' in name and name.endswith('bias')):
        torch.nn.init.normal_(p, std=0.01)

def to_ref(k):
    k = re.sub(r'^rgb\.(layer\d)\.', r'\1.', k)
    k = re.sub(r'^evt\.(layer\d)\.', r'\1_event.', k)
    k = re.sub(r'^rgb\.(conv1|bn1)\.', r'\1.', k)
    k = re.sub(r'^evt\.(conv1|bn1)\.', r'\1_event.', k)
    k = re.sub(r'^(regressionModel|classificationModel)\.convs\.(\d)\.', lambda m: f'{{m.group(1)}}.conv{{int(m.group(2)) + 1}}.', k)
    return k

ref = ref_model.resnet50('dsec', 3, 'fpn_fusion')
sd = {{to_ref(k): v.clone() for k, v in ours.state_dict().items()}}
ref.load_state_dict(sd, strict=True)                   # every reference key is covered, none is left over
ref.train(); ref.freeze_bn()                           # training.py order (train_dsec_det_fast.py:230-232)
ours.train()
ref_model.random.uniform = lambda a, b: 1.0            # no RGB blanking in this step (model.py:279-280)

def step(net, call):
    for p in net.parameters():
        p.grad = None
    cls, reg = call()
    (cls.mean() + reg.mean()).backward()
    return float(cls.mean()), float(reg.mean())

lr = step(ref, lambda: ref([rgb.clone(), evt.clone(), ann.clone()]))
lo = step(ours, lambda: ours(rgb.clone(), evt.clone(), ann.clone()))
gref = dict(ref.named_parameters())
worst, worst_key, n = 0.0, None, 0
for k, p in ours.named_parameters():
    g, r = p.grad, gref[to_ref(k)].grad
    if g is None and r is None:
        continue
    n += 1
    den = float(r.norm())
    err = float((g - r).norm()) / den if den > 0 else float(g.norm())
    if err > worst:
        worst, worst_key = err, k
print('JSON' + json.dumps(dict(ref=lr, ours=lo, worst=worst, worst_key=worst_key, compared=n,
                               fus_grad=float(gref['fus.0.conv0_rgb.weight'].grad.norm()))))
"""


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "retinanet")), reason="/root/reference is only present in the build container")
def test_harness_network_is_the_reference_network():
    code = _SCRIPT.format(root=ROOT, tools=os.path.join(ROOT, "tools"), shim=os.path.join(ROOT, "oracle", "fvcore_shim"), ref=REF)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-3000:]
    r = json.loads([l for l in out.stdout.splitlines() if l.startswith("JSON")][-1][4:])
    print(r)
    assert r["compared"] > 300 and r["fus_grad"] > 0            # the fusion blocks take part in the step
    for a, b in zip(r["ref"], r["ours"]):
        assert abs(a - b) <= 1e-5 * max(abs(a), 1e-3), r        # same losses (fp32, same op order up to the heads)
    assert r["worst"] <= 1e-4, r                                 # same gradients for every parameter (fp32 CPU conv noise)
