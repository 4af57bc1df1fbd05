"""Generate tests/golden/*.npz by running the UNMODIFIED reference module in fp64.

Run in the build container only (it needs /root/reference):

    python oracle/make_golden.py

TEST INFRASTRUCTURE ONLY.  Imports retinanet.fusion / Attention / AdaIN from
/root/reference behind the fvcore shim (oracle/fvcore_shim), runs them on CPU in
float64 with fixed seeds, and stores inputs, parameters (reference state_dict
keys), outputs and autograd gradients.  It also pins both restatements
(oracle/fusion_numpy.py, oracle/fusion_torch.py) against the module at larger
shapes that are not stored, and writes the residuals to
tests/golden/pin_report.json.  Nothing on the GPU box reads /root/reference;
the tests read only the committed .npz / .json files.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REFERENCE = os.environ.get("FRN_REFERENCE", "/root/reference")
GOLDEN = os.path.join(ROOT, "tests", "golden")

sys.path.insert(0, os.path.join(HERE, "fvcore_shim"))
sys.path.insert(0, REFERENCE)
sys.path.insert(0, ROOT)

from retinanet.AdaIN import AdaIN_block  # noqa: E402  (reference, unmodified)
from retinanet.Attention import CrossAttentionBlock  # noqa: E402
from retinanet.fusion import REFusion  # noqa: E402

from oracle import fusion_numpy as onp  # noqa: E402
from oracle import fusion_torch as otorch  # noqa: E402


def make_inputs(kind, shape, seed):
    g = torch.Generator().manual_seed(seed)
    b, c, h, w = shape
    if kind == "randn":
        a = torch.randn(shape, generator=g, dtype=torch.float64)
        e = torch.randn(shape, generator=g, dtype=torch.float64)
    elif kind == "relu":          # post-ReLU backbone features (utils.py:77-78)
        a = torch.randn(shape, generator=g, dtype=torch.float64).relu()
        e = torch.randn(shape, generator=g, dtype=torch.float64).relu()
    elif kind == "nearconst":     # |mean| >> std: stresses the variance
        a = 1.0 + 1e-3 * torch.randn(shape, generator=g, dtype=torch.float64)
        e = 3.0 + 1e-3 * torch.randn(shape, generator=g, dtype=torch.float64)
    elif kind == "zero_second":   # model.py:279-280 blanks RGB; model.py:305 passes rgb second
        a = torch.randn(shape, generator=g, dtype=torch.float64).relu()
        e = torch.zeros(shape, dtype=torch.float64)
    elif kind == "zero_first":
        a = torch.zeros(shape, dtype=torch.float64)
        e = torch.randn(shape, generator=g, dtype=torch.float64).relu()
    else:
        raise ValueError(kind)
    dout = torch.randn((b, 2 * c, h, w), generator=g, dtype=torch.float64)
    return a, e, dout


def build_module(c, seed):
    torch.manual_seed(seed)
    mod = REFusion(c, c).double()
    # c2_xavier_fill zeroes the attention biases; randomise them so that the bias
    # gradients and the bias terms themselves are exercised.
    g = torch.Generator().manual_seed(seed + 77)
    with torch.no_grad():
        for name, p in mod.named_parameters():
            if name.endswith("bias") and "cross_attention" in name:
                p.copy_(0.1 * torch.randn(p.shape, generator=g, dtype=torch.float64))
    return mod


def run_reference(mod, a, e, dout):
    a = a.clone().requires_grad_(True)
    e = e.clone().requires_grad_(True)
    mod.zero_grad()
    out = mod(a, e)
    out.backward(dout)
    grads = {k: p.grad.detach().clone() for k, p in mod.named_parameters()}
    return out.detach(), a.grad.detach(), e.grad.detach(), grads


def maxabs(x, y):
    return float(np.max(np.abs(np.asarray(x) - np.asarray(y)))) if np.asarray(x).size else 0.0


def pin(mod, a, e, dout, ref):
    """Residuals of both restatements against the module."""
    out, da, de, grads = ref
    prm = {k: v.detach() for k, v in mod.state_dict().items()}
    prm_np = {k: v.numpy() for k, v in prm.items()}
    o2, da2, de2, g2 = onp.refusion_forward_backward(a.numpy(), e.numpy(), prm_np, dout.numpy())
    o3, da3, de3, g3 = otorch.refusion_fwd_bwd(a, e, prm, dout)
    rep = {
        "numpy_out_rel": onp.rel_l2(o2, out), "numpy_da_rel": onp.rel_l2(da2, da), "numpy_de_rel": onp.rel_l2(de2, de),
        "torch_out_maxabs": maxabs(o3, out), "torch_da_maxabs": maxabs(da3, da), "torch_de_maxabs": maxabs(de3, de),
        "numpy_grad_rel": {}, "torch_grad_maxabs": {},
    }
    for k in grads:
        ref_g = grads[k].numpy()
        if k.endswith("theta.bias"):   # analytically zero: absolute
            rep["numpy_grad_rel"][k] = maxabs(g2[k], ref_g)
        else:
            rep["numpy_grad_rel"][k] = onp.rel_l2(g2[k], ref_g)
        rep["torch_grad_maxabs"][k] = maxabs(g3[k], ref_g)
    return rep


def summarise(rep):
    return {
        "numpy_out_rel": rep["numpy_out_rel"], "numpy_da_rel": rep["numpy_da_rel"], "numpy_de_rel": rep["numpy_de_rel"],
        "numpy_grad_worst": max(rep["numpy_grad_rel"].values()),
        "torch_out_maxabs": rep["torch_out_maxabs"],
        "torch_grad_worst_maxabs": max(rep["torch_grad_maxabs"].values()),
    }


STORED = [  # name, kind, (B, C, H, W), seed  -- small enough to commit
    ("refusion_randn_c64_5x7", "randn", (2, 64, 5, 7), 11),
    ("refusion_relu_c32_6x8", "relu", (3, 32, 6, 8), 12),
    ("refusion_nearconst_c32_4x9", "nearconst", (2, 32, 4, 9), 13),
    ("refusion_zero_second_c32_5x5", "zero_second", (2, 32, 5, 5), 14),
    ("refusion_zero_first_c32_5x5", "zero_first", (1, 32, 5, 5), 15),
    ("refusion_relu_c128_9x11", "relu", (1, 128, 9, 11), 16),   # DDD17 L4 spatial size
    ("refusion_relu_c16_3x3", "relu", (1, 16, 3, 3), 17),       # d = 2, N = 9
]

PINNED_ONLY = [  # real level shapes, checked here, not stored
    ("ddd17_L3", "relu", (1, 1024, 17, 22), 21),
    ("ddd17_L4", "relu", (2, 2048, 9, 11), 22),
    ("dsec_L3", "relu", (1, 1024, 30, 40), 23),
    ("ddd17_L2", "relu", (1, 512, 33, 44), 24),
]


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    report = {"torch": torch.__version__, "reference": REFERENCE, "cases": {}}

    for name, kind, shape, seed in STORED:
        mod = build_module(shape[1], seed)
        a, e, dout = make_inputs(kind, shape, seed)
        ref = run_reference(mod, a, e, dout)
        out, da, de, grads = ref
        arrays = {"a": a.numpy(), "e": e.numpy(), "dout": dout.numpy(),
                  "out": out.numpy(), "da": da.numpy(), "de": de.numpy()}
        for k, v in mod.state_dict().items():
            arrays["param/" + k] = v.numpy()
        for k, v in grads.items():
            arrays["grad/" + k] = v.numpy()
        np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **arrays)
        report["cases"][name] = summarise(pin(mod, a, e, dout, ref))
        print(name, report["cases"][name])

    for name, kind, shape, seed in PINNED_ONLY:
        mod = build_module(shape[1], seed)
        a, e, dout = make_inputs(kind, shape, seed)
        ref = run_reference(mod, a, e, dout)
        report["cases"][name] = summarise(pin(mod, a, e, dout, ref))
        print(name, report["cases"][name])

    # AdaIN_block alone (AdaIN.py) and CrossAttentionBlock alone (Attention.py)
    g = torch.Generator().manual_seed(31)
    content = torch.randn((2, 6, 5, 7), generator=g, dtype=torch.float64) * 2 + 0.5
    style = torch.randn((2, 6, 5, 7), generator=g, dtype=torch.float64) * 0.3 - 1.0
    ad = AdaIN_block()
    mean, std = ad.calc_mean_std(content)
    z = ad(content, style)
    np.savez_compressed(os.path.join(GOLDEN, "adain_c6_5x7.npz"), content=content.numpy(), style=style.numpy(),
                        mean=mean.numpy(), std=std.numpy(), out=z.numpy())
    report["cases"]["adain_c6_5x7"] = {
        "numpy_out_maxabs": maxabs(onp.adain_forward(content.numpy(), style.numpy()), z.numpy()),
        "torch_out_maxabs": maxabs(otorch.adain(content, style), z),
    }

    torch.manual_seed(32)
    cab = CrossAttentionBlock(48).double()
    with torch.no_grad():
        for p in cab.parameters():
            if p.dim() == 1:
                p.copy_(0.1 * torch.randn(p.shape, generator=g, dtype=torch.float64))
    x0 = torch.randn((2, 48, 4, 7), generator=g, dtype=torch.float64)
    x1 = torch.randn((2, 48, 4, 7), generator=g, dtype=torch.float64)
    zc = cab([x0, x1]).detach()
    arrays = {"x0": x0.numpy(), "x1": x1.numpy(), "out": zc.numpy()}
    for k, v in cab.state_dict().items():
        arrays["param/" + k] = v.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "cab_c48_4x7.npz"), **arrays)
    prm_np = {"rgb_cross_attention." + k: v.numpy() for k, v in cab.state_dict().items()}
    report["cases"]["cab_c48_4x7"] = {
        "numpy_out_maxabs": maxabs(onp.cross_attention_forward(x0.numpy(), x1.numpy(), prm_np), zc.numpy()),
    }

    # state_dict contract: keys and shapes of REFusion(64, 64)
    ref_mod = REFusion(64, 64)
    report["state_dict_c64"] = {k: list(v.shape) for k, v in ref_mod.state_dict().items()}
    with open(os.path.join(GOLDEN, "pin_report.json"), "w") as fh:
        json.dump(report, fh, indent=1, sort_keys=True)
    print("wrote", GOLDEN)


if __name__ == "__main__":
    main()
