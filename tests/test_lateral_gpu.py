"""GPU: REFusion fused with the FPN lateral 1x1 conv that consumes it (model.py:305-308 followed
by P{5,4,3,2}_1 at model.py:59-76) against the fp64 oracle of lateral(REFusion(a, e)): lateral output, input
gradients, every fusion-parameter gradient and the lateral's own weight / bias gradients, l2-relative <= 1e-5."""
import numpy as np
import pytest
import torch

import frn_b200
from helpers import RTOL_F32, check_against, load_into, make_inputs, make_params, rel_l2
from oracle import fusion_numpy as onp

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

CASES = [("small", 2, 64, 9, 8, 32), ("ddd17_L3", 2, 1024, 17, 22, 256), ("dsec_L4", 3, 2048, 15, 20, 256),
         ("dsec_L2", 1, 512, 60, 80, 256)]


def _oracle(a, e, prm, wl, bl, dlat, dmap=None):
    """fp64: lat = Wl out + bl with out = REFusion(a, e); gradients by the chain rule through the fp64 oracle."""
    p = {k: v.numpy() for k, v in prm.items()}
    out, _ = onp.refusion_forward(a, e, p)
    b, c2, h, w = out.shape
    o2 = out.reshape(b, c2, h * w)
    lat = np.einsum("ok,bkn->bon", wl, o2) + bl[None, :, None]
    g = dlat.reshape(b, -1, h * w)
    dout = np.einsum("ok,bon->bkn", wl, g).reshape(out.shape)
    if dmap is not None:
        dout = dout + dmap
    _, da, de, grads = onp.refusion_forward_backward(a, e, p, dout)
    return out, lat.reshape(b, -1, h, w), da, de, grads, np.einsum("bon,bkn->ok", g, o2), g.sum(axis=(0, 2))


@pytest.mark.parametrize("name,b,c,h,w,cout", CASES, ids=[x[0] for x in CASES])
@pytest.mark.parametrize("with_map", [False, True], ids=["lat_only", "lat_and_map"])
def test_refusion_lateral_vs_oracle(name, b, c, h, w, cout, with_map):
    if with_map and name == "dsec_L2":
        pytest.skip("the oracle at N = 4800 runs once")
    a, e, _ = make_inputs("relu", (b, c, h, w), seed=500 + c)
    a, e = a.float().double(), e.float().double()
    prm = make_params(c, seed=2)
    g = torch.Generator().manual_seed(7)
    wl = (torch.randn(cout, 2 * c, generator=g, dtype=torch.float64) / (2 * c) ** 0.5).float().double()
    bl = (0.1 * torch.randn(cout, generator=g, dtype=torch.float64)).float().double()
    dlat = torch.randn(b, cout, h, w, generator=g, dtype=torch.float64).float().double()
    dmap = torch.randn(b, 2 * c, h, w, generator=g, dtype=torch.float64).float().double() if with_map else None
    out_r, lat_r, da_r, de_r, grads_r, dwl_r, dbl_r = _oracle(a.numpy(), e.numpy(), prm, wl.numpy(), bl.numpy(), dlat.numpy(),
                                                              dmap.numpy() if with_map else None)
    fus = frn_b200.REFusion(c, c)
    load_into(fus, {k: v.numpy() for k, v in prm.items()})
    lateral = torch.nn.Conv2d(2 * c, cout, kernel_size=1)
    with torch.no_grad():
        lateral.weight.copy_(wl.float().view(cout, 2 * c, 1, 1))
        lateral.bias.copy_(bl.float())
    m = frn_b200.REFusionLateral(fus, lateral, return_map=with_map).to(DEV)
    x = a.float().to(DEV).requires_grad_(True)
    y = e.float().to(DEV).requires_grad_(True)
    res = m(x, y)
    if with_map:
        lat, fmap = res
        assert rel_l2(fmap.detach().cpu().numpy(), out_r) < RTOL_F32
        torch.autograd.backward([lat, fmap], [dlat.float().to(DEV), dmap.float().to(DEV)])
    else:
        lat = res
        lat.backward(dlat.float().to(DEV))
    torch.cuda.synchronize()
    assert lat.shape == (b, cout, h, w)
    assert rel_l2(lat.detach().cpu().numpy(), lat_r) < RTOL_F32, rel_l2(lat.detach().cpu().numpy(), lat_r)
    grads = {k: p.grad.detach().cpu().numpy() for k, p in fus.named_parameters()}
    check_against(name, (out_r, x.grad.cpu().numpy(), y.grad.cpu().numpy(), grads), (out_r, da_r, de_r, grads_r))
    assert rel_l2(lateral.weight.grad.cpu().numpy().reshape(cout, 2 * c), dwl_r) < RTOL_F32
    assert rel_l2(lateral.bias.grad.cpu().numpy(), dbl_r) < RTOL_F32


def test_fused_equals_unfused_modules():
    """Same modules, same parameters: REFusionLateral == Conv1x1(REFusion(...)) to fp32 rounding, forward and backward."""
    b, c, h, w, cout = 2, 256, 12, 16, 256
    torch.manual_seed(3)
    fus = frn_b200.REFusion(c, c).to(DEV)
    lateral = frn_b200.Conv1x1(2 * c, cout).to(DEV)
    a = torch.randn(b, c, h, w, device=DEV).relu()
    e = torch.randn(b, c, h, w, device=DEV).relu()
    go = torch.randn(b, cout, h, w, device=DEV)
    x, y = a.clone().requires_grad_(True), e.clone().requires_grad_(True)
    lateral(fus(x, y)).backward(go)
    ref = {k: p.grad.clone() for k, p in list(fus.named_parameters()) + [("lat." + k, p) for k, p in lateral.named_parameters()]}
    ref_lat = lateral(fus(a, e)).detach()
    fus.zero_grad(); lateral.zero_grad()
    x2, y2 = a.clone().requires_grad_(True), e.clone().requires_grad_(True)
    fused = frn_b200.REFusionLateral(fus, lateral)
    lat = fused(x2, y2)
    lat.backward(go)
    assert rel_l2(lat.detach().cpu().numpy(), ref_lat.double().cpu().numpy()) < 5e-6
    assert rel_l2(x2.grad.cpu().numpy(), x.grad.double().cpu().numpy()) < 5e-6
    for k, p in list(fus.named_parameters()) + [("lat." + k, p) for k, p in lateral.named_parameters()]:
        if k.endswith("theta.bias"):
            continue
        assert rel_l2(p.grad.cpu().numpy(), ref[k].double().cpu().numpy()) < 2e-5, k
    assert set(fused.state_dict()) == {"fusion." + k for k in fus.state_dict()} | {"lateral." + k for k in lateral.state_dict()}
