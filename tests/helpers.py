"""Shared test helpers: golden fixtures, seeded inputs, oracle calls, error metrics."""
import glob
import os

import numpy as np
import torch

from oracle import fusion_numpy as onp

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# Tolerances (BASELINE.json north_star: fp32 <= 1e-5 relative vs an fp64 run of the reference;
# l2-relative norm).  theta.bias has an analytically zero gradient -> absolute bound
# relative to the scale of the other gradients.
RTOL_F32 = 1e-5
ATOL_THETA_BIAS = 1e-4


def golden_files(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def load_golden(path):
    z = np.load(path)
    d = {k: z[k] for k in z.files}
    params = {k[len("param/"):]: v for k, v in d.items() if k.startswith("param/")}
    grads = {k[len("grad/"):]: v for k, v in d.items() if k.startswith("grad/")}
    return d, params, grads


def rel_l2(x, ref):
    return onp.rel_l2(x, ref)


def make_inputs(kind, shape, seed):
    """Same recipes as oracle/make_golden.py (fp64 on CPU)."""
    g = torch.Generator().manual_seed(seed)
    b, c, h, w = shape
    if kind == "randn":
        a = torch.randn(shape, generator=g, dtype=torch.float64)
        e = torch.randn(shape, generator=g, dtype=torch.float64)
    elif kind == "relu":
        a = torch.randn(shape, generator=g, dtype=torch.float64).relu()
        e = torch.randn(shape, generator=g, dtype=torch.float64).relu()
    elif kind == "nearconst":
        a = 1.0 + 1e-3 * torch.randn(shape, generator=g, dtype=torch.float64)
        e = 3.0 + 1e-3 * torch.randn(shape, generator=g, dtype=torch.float64)
    else:
        raise ValueError(kind)
    dout = torch.randn((b, 2 * c, h, w), generator=g, dtype=torch.float64)
    return a, e, dout


def make_params(c, seed, bias_scale=0.1):
    """Reference-shaped parameters (fp32-representable fp64 values) with non-zero attention biases."""
    from oracle import fusion_torch as ot
    prm = ot.init_params(c, seed=seed, dtype=torch.float64)
    g = torch.Generator().manual_seed(seed + 77)
    for k in prm:
        if k.endswith("bias") and "cross_attention" in k:
            prm[k] = bias_scale * torch.randn(prm[k].shape, generator=g, dtype=torch.float64)
        prm[k] = prm[k].float().double()
    return prm


def load_into(module, params):
    sd = {k: torch.as_tensor(np.asarray(v)).float() for k, v in params.items()}
    missing = module.load_state_dict(sd, strict=True)
    return missing


def run_module(module, a, e, dout):
    """fwd+bwd of a frn_b200 REFusion on cuda:0 -> numpy out, da, de, grads."""
    dev = torch.device("cuda:0")
    module = module.to(dev)
    a = torch.as_tensor(np.asarray(a)).float().to(dev).requires_grad_(True)
    e = torch.as_tensor(np.asarray(e)).float().to(dev).requires_grad_(True)
    module.zero_grad()
    out = module(a, e)
    out.backward(torch.as_tensor(np.asarray(dout)).float().to(dev))
    torch.cuda.synchronize()
    grads = {k: p.grad.detach().cpu().numpy() for k, p in module.named_parameters()}
    return out.detach().cpu().numpy(), a.grad.cpu().numpy(), e.grad.cpu().numpy(), grads


PARITY_LOG = []          # (name, {quantity: l2-relative error}) of every block-level gate of this session (conftest dumps it)


def check_against(name, got, ref, tol=RTOL_F32):
    out, da, de, grads = got
    rout, rda, rde, rgrads = ref
    errs = {"out": rel_l2(out, rout), "da": rel_l2(da, rda), "de": rel_l2(de, rde)}
    gscale = max(float(np.abs(np.asarray(v)).max()) for v in rgrads.values())
    for k, v in rgrads.items():
        if k.endswith("theta.bias"):
            # true gradient is exactly zero; the reference itself returns rounding noise
            # -> absolute bound: |g| <= ATOL_THETA_BIAS * (largest gradient entry of the block)
            errs[k] = float(np.abs(grads[k]).max()) / max(gscale, 1e-30) * (tol / ATOL_THETA_BIAS)
        else:
            errs[k] = rel_l2(grads[k], v)
    PARITY_LOG.append((name, {k: float(v) for k, v in errs.items()}))
    bad = {k: v for k, v in errs.items() if not (v <= tol)}
    assert not bad, f"{name}: over tolerance {tol}: {bad}\nall: {errs}"
    return errs


def debug_engines_available():
    """True when the library was built with -DFRN_DEBUG_ENGINES (make DEBUG_ENGINES=1): the FFMA engines of round 1
    are then selectable through frn_set_*_engine.  The product build is tcgen05-only and ignores the switches."""
    from frn_b200 import _lib
    lib = _lib.load()
    old = lib.frn_get_gemm_engine()
    lib.frn_set_gemm_engine(0)
    ok = lib.frn_get_gemm_engine() == 0
    lib.frn_set_gemm_engine(old)
    return ok
