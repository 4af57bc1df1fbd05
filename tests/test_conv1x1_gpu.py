"""GPU: the FPN lateral 1x1 convolution (nn.Conv2d(2C, 256, 1) of
retinanet/model.py:28-40) through frn_b200.Conv1x1 and the C ABI (frn_conv1x1_*), against an fp64
F.conv2d with autograd, at the DSEC / DDD17 level shapes (K = 2C = 512 ... 4096) and ragged sizes.
Gate: l2-relative <= 2e-6 for y, dx, dw, db (the GEMM engine's own gate, tests/test_gemm_gpu.py)."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as TF

import frn_b200

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 2e-6

CASES = [
    # name, B, Cin, Cout, H, W
    ("dsec_P2", 1, 512, 256, 120, 160),
    ("dsec_P3", 2, 1024, 256, 60, 80),
    ("dsec_P4", 2, 2048, 256, 30, 40),
    ("dsec_P5", 2, 4096, 256, 15, 20),
    ("ddd17_P2", 2, 512, 256, 65, 87),
    ("ddd17_P5", 3, 4096, 256, 9, 11),
    ("ragged", 3, 77, 19, 5, 7),
]


def rel(x, ref):
    x, ref = x.double().cpu(), ref.double().cpu()
    return float((x - ref).norm() / ref.norm())


@pytest.mark.parametrize("name,b,cin,cout,h,w", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("bias", [True, False], ids=["bias", "nobias"])
def test_conv1x1_matches_fp64(name, b, cin, cout, h, w, bias):
    g = torch.Generator().manual_seed(len(name) * 131 + cin)
    x = torch.randn(b, cin, h, w, generator=g).relu()
    dy = torch.randn(b, cout, h, w, generator=g)
    m = frn_b200.Conv1x1(cin, cout, bias=bias)
    if bias:
        with torch.no_grad():
            m.bias.copy_(torch.randn(cout, generator=g))
    x64 = x.double().requires_grad_(True)
    w64 = m.weight.detach().double().requires_grad_(True)
    b64 = m.bias.detach().double().requires_grad_(True) if bias else None
    TF.conv2d(x64, w64, b64).backward(dy.double())
    with torch.no_grad():
        y64 = TF.conv2d(x64, w64, b64)

    m = m.to(DEV)
    xd = x.to(DEV).requires_grad_(True)
    y = m(xd)
    y.backward(dy.to(DEV))
    torch.cuda.synchronize()
    errs = {"y": rel(y.detach(), y64), "dx": rel(xd.grad, x64.grad), "dw": rel(m.weight.grad, w64.grad)}
    if bias:
        errs["db"] = rel(m.bias.grad, b64.grad)
    assert all(v <= TOL for v in errs.values()), errs


def test_no_input_grad_and_inference():
    g = torch.Generator().manual_seed(5)
    m = frn_b200.Conv1x1(64, 32).to(DEV)
    x = torch.randn(2, 64, 9, 11, generator=g).to(DEV)
    with torch.no_grad():
        y0 = m(x)
    y1 = m(x)                      # x does not require grad: dx is skipped inside the library
    y1.sum().backward()
    assert torch.equal(y0, y1)
    ref = x.double().sum(dim=(0, 2, 3))
    assert rel(m.weight.grad[7].flatten(), ref) <= TOL      # dy = 1: every row of dw is sum_bn x


def test_convert_fpn_laterals_shares_parameters():
    class Pyr(nn.Module):           # attribute names of PyramidFeatures, model.py:28-40
        def __init__(self):
            super().__init__()
            self.P5_1 = nn.Conv2d(64, 16, kernel_size=1, stride=1, padding=0)
            self.P5_2 = nn.Conv2d(16, 16, kernel_size=3, stride=1, padding=1)
            self.P4_1 = nn.Conv2d(32, 16, kernel_size=1, stride=1, padding=0)

    net = nn.Sequential(Pyr()).to(DEV)
    keys = list(net.state_dict().keys())
    w_before = net[0].P5_1.weight
    assert frn_b200.convert_fpn_laterals(net) == ["0.P5_1", "0.P4_1"]
    assert isinstance(net[0].P5_1, frn_b200.Conv1x1) and not isinstance(net[0].P5_2, frn_b200.Conv1x1)
    assert net[0].P5_1.weight is w_before and list(net.state_dict().keys()) == keys
    x = torch.randn(2, 64, 6, 5, device=DEV)
    ref = TF.conv2d(x.double(), w_before.double(), net[0].P5_1.bias.double())
    assert rel(net[0].P5_1(x), ref) <= TOL
