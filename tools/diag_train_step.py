import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
from harness_model import FusionDetector, calibrate_batchnorm, synthetic_batch
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
dev = torch.device("cuda:0")
torch.manual_seed(0)
ours = FusionDetector(3, "ours").to(dev).train()
eager = FusionDetector(3, "eager").to(dev).train()
rgb, evt, ann = synthetic_batch(2, dev, height=192, width=256, boxes=3)
calibrate_batchnorm(ours, rgb, evt)
eager.load_state_dict(ours.state_dict())
with torch.no_grad():
    for net in (ours, eager):
        g = torch.Generator().manual_seed(1)
        for head in (net.regressionModel, net.classificationModel):
            head.output.weight.copy_(0.01 * torch.randn(head.output.weight.shape, generator=g))
res = {}
for name, net in (("ours", ours), ("eager", eager)):
    keep = {}
    def hook(key):
        def f(m, i, o):
            o.retain_grad(); keep[key] = o
        return f
    hs = [net.fus[i].register_forward_hook(hook(f"fus{i}")) for i in range(4)]
    hs += [net.fpn.register_forward_hook(lambda m, i, o: [t.retain_grad() for t in o] and keep.update({f"P{j}": t for j, t in enumerate(o)}))]
    lc, lr = net(rgb, evt, ann)
    (lc + lr).sum().backward()
    res[name] = (lc.item(), lr.item(), {k: p.grad.clone() for k, p in net.named_parameters() if p.grad is not None},
                 {k: (v.detach().clone(), v.grad.clone()) for k, v in keep.items()})
    for h in hs: h.remove()
print("loss", res["ours"][:2], res["eager"][:2])
rel = lambda a, b: float((a - b).norm() / b.norm().clamp(min=1e-30))
for k in sorted(res["ours"][3]):
    o, e = res["ours"][3][k], res["eager"][3][k]
    print(f"{k:6s} out {rel(o[0], e[0]):.2e}  grad {rel(o[1], e[1]):.2e}  |grad| {float(e[1].norm()):.3e}")
for k in ["classificationModel.output.bias", "regressionModel.output.bias", "classificationModel.convs.0.weight",
          "fpn.P2_2.weight", "fpn.P2_1.weight", "fpn.P5_1.weight", "fpn.P6.weight", "fus.0.conv0_rgb.weight", "fus.3.conv0_rgb.weight",
          "fus.0.rgb_cross_attention.W.weight", "fus.3.rgb_cross_attention.W.weight", "fus.3.rgb_cross_attention.g.weight", "rgb.layer4.2.conv3.weight", "rgb.conv1.weight"]:
    print(f"{k:45s} {rel(res['ours'][2][k], res['eager'][2][k]):.2e}  |g| {float(res['eager'][2][k].norm()):.3e}")
