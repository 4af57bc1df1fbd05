"""fp64 NumPy oracle for FRN's hierarchical cross-modal fusion block.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (frn_b200/) may import
this file: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs do, and only as the checker or the timed CPU baseline.

It restates, in plain NumPy float64 with a hand-derived backward pass (no
autograd), what the reference computes in

    retinanet/AdaIN.py:7-26      calc_mean_std + AdaIN_block.forward
    retinanet/Attention.py:21-36 CrossAttentionBlock.forward
    retinanet/fusion.py:20-36    REFusion.forward

The arithmetic itself lives in a third-party dependency that is not vendored
in /root/reference: PyTorch ATen (conv2d, matmul, softmax, var, mean, sqrt;
README.md:25,36 names torch 2.0.1, this image has torch 2.11.0).  The
reference owns no golden vectors or tests for this path,
so the oracle is pinned by executing the unmodified reference module here:
oracle/make_golden.py imports retinanet.fusion.REFusion from /root/reference,
runs it in fp64 and commits its inputs / outputs / autograd gradients under
tests/golden/; tests/test_oracle.py checks this file against those vectors.

Notation (per sample; feature maps are C x N matrices, N = H*W, row-major):
    A, E       the two inputs (first / second positional argument of forward)
    R0, E0     conv0_rgb(A), conv0_evt(E)                       fusion.py:23-24
    M          R0 * E0 ; R1 = R0 + M ; E1 = E0 + M              fusion.py:26-28
    block "A"  rgb_cross_attention([R1, E1])                    fusion.py:31
    block "B"  event_cross_attention([E1, R1])                  fusion.py:32
    out        concat(blockA, blockB) on channels               fusion.py:34
"""
from __future__ import annotations

import numpy as np

EPS = 1e-5  # AdaIN.py:7 (added to the variance before the sqrt, AdaIN.py:12-13)
_ROW_BLOCK = 1024  # query rows per softmax block; bounds memory, changes nothing

BLOCKS = ("rgb_cross_attention", "event_cross_attention")
CONVS = ("g", "theta", "phi", "W")


def param_names():
    """state_dict keys of the reference REFusion, in its registration order."""
    names = []
    for blk in BLOCKS:
        for conv in CONVS:
            names += [f"{blk}.{conv}.weight", f"{blk}.{conv}.bias"]
    names += ["conv0_rgb.weight", "conv0_rgb.bias", "conv0_evt.weight", "conv0_evt.bias"]
    return names


def _w2d(w):
    w = np.asarray(w, dtype=np.float64)
    return w.reshape(w.shape[0], w.shape[1])


def calc_mean_std(feat, eps=EPS):
    """AdaIN.py:7-15.  feat (B, C, H, W) -> mean, std each (B, C, 1, 1).

    Unbiased variance (torch .var default, divisor N-1), eps inside the sqrt.
    """
    feat = np.asarray(feat, dtype=np.float64)
    assert feat.ndim == 4  # AdaIN.py:10
    b, c = feat.shape[:2]
    flat = feat.reshape(b, c, -1)
    var = flat.var(axis=2, ddof=1) + eps
    return flat.mean(axis=2).reshape(b, c, 1, 1), np.sqrt(var).reshape(b, c, 1, 1)


def _stats(x):
    """x (B, C, N) -> mean, std (B, C)."""
    n = x.shape[2]
    mu = x.mean(axis=2)
    d = x - mu[:, :, None]
    var = (d * d).sum(axis=2) / (n - 1) + EPS
    return mu, np.sqrt(var)


def adain_forward(content, style):
    """AdaIN.py:17-26 on (B, C, H, W) arrays."""
    content = np.asarray(content, dtype=np.float64)
    style = np.asarray(style, dtype=np.float64)
    assert content.shape[:2] == style.shape[:2]  # AdaIN.py:20
    b, c = content.shape[:2]
    mu_c, sd_c = _stats(content.reshape(b, c, -1))
    mu_s, sd_s = _stats(style.reshape(b, c, -1))
    xhat = (content - mu_c[:, :, None, None]) / sd_c[:, :, None, None]
    return xhat * sd_s[:, :, None, None] + mu_s[:, :, None, None]


def _softmax_rows(s):
    m = s.max(axis=-1, keepdims=True)
    e = np.exp(s - m)
    l = e.sum(axis=-1, keepdims=True)
    return e / l, (m + np.log(l))[..., 0]


def _attention_forward(q, k, v):
    """q (N, d), k (d, N), v (N, d) -> y (N, d), lse (N).  Attention.py:29-32."""
    n = q.shape[0]
    y = np.empty_like(q)
    lse = np.empty(n)
    for r0 in range(0, n, _ROW_BLOCK):
        r1 = min(n, r0 + _ROW_BLOCK)
        p, lse[r0:r1] = _softmax_rows(q[r0:r1] @ k)
        y[r0:r1] = p @ v
    return y, lse


def _attention_backward(q, k, v, y, lse, dy):
    """Gradients of y = softmax(q k) v.  SURVEY.md 8.1."""
    n = q.shape[0]
    dq = np.empty_like(q)
    dk = np.zeros_like(k)
    dv = np.zeros_like(v)
    delta = (dy * y).sum(axis=1)
    for r0 in range(0, n, _ROW_BLOCK):
        r1 = min(n, r0 + _ROW_BLOCK)
        p = np.exp(q[r0:r1] @ k - lse[r0:r1, None])
        dv += p.T @ dy[r0:r1]
        dp = dy[r0:r1] @ v.T
        ds = p * (dp - delta[r0:r1, None])
        dq[r0:r1] = ds @ k.T
        dk += q[r0:r1].T @ ds
    return dq, dk, dv


def _cab_forward(x0, x1, prm, blk):
    """CrossAttentionBlock.forward([x0, x1]) on (B, C, N) arrays.  Attention.py:21-36."""
    b, c, n = x0.shape
    d = c // 8
    wg, bg = _w2d(prm[f"{blk}.g.weight"]), np.asarray(prm[f"{blk}.g.bias"], np.float64)
    wt, bt = _w2d(prm[f"{blk}.theta.weight"]), np.asarray(prm[f"{blk}.theta.bias"], np.float64)
    wp, bp = _w2d(prm[f"{blk}.phi.weight"]), np.asarray(prm[f"{blk}.phi.bias"], np.float64)
    ww, bw = _w2d(prm[f"{blk}.W.weight"]), np.asarray(prm[f"{blk}.W.bias"], np.float64)
    z = np.empty_like(x0)
    cache = dict(x0=x0, x1=x1, q=[], k=[], v=[], y=[], lse=[], yp=[], t=[])
    t_all = np.empty_like(x0)
    for i in range(b):
        v = (wg @ x0[i] + bg[:, None]).T          # g(x[0]) permuted      Attention.py:24
        k = wt @ x1[i] + bt[:, None]              # theta(x[1])           Attention.py:26
        q = (wp @ x1[i] + bp[:, None]).T          # phi(x[1]) permuted    Attention.py:27
        y, lse = _attention_forward(q, k, v)      #                       Attention.py:29-32
        # Attention.py:32: the (N, d) buffer is reinterpreted as (d, H, W) with no
        # permute back, i.e. Y'[k', n'] = Yflat[k' * N + n'].
        yp = np.ascontiguousarray(y).reshape(d, n)
        t_all[i] = ww @ yp + bw[:, None]          # W conv                Attention.py:34
        for name, val in (("q", q), ("k", k), ("v", v), ("y", y), ("lse", lse), ("yp", yp)):
            cache[name].append(val)
    mu_x, sd_x = _stats(x0)
    mu_t, sd_t = _stats(t_all)
    xhat = (x0 - mu_x[:, :, None]) / sd_x[:, :, None]
    z = xhat * sd_t[:, :, None] + mu_t[:, :, None]   # AdaIN.py:24-26
    cache.update(t=t_all, mu_x=mu_x, sd_x=sd_x, mu_t=mu_t, sd_t=sd_t, xhat=xhat)
    return z, cache


def _cab_backward(g, cache, prm, blk):
    """Returns dx0, dx1 and this block's parameter gradients.  SURVEY.md 8.1."""
    x0, x1 = cache["x0"], cache["x1"]
    b, c, n = x0.shape
    d = c // 8
    wg = _w2d(prm[f"{blk}.g.weight"])
    wt = _w2d(prm[f"{blk}.theta.weight"])
    wp = _w2d(prm[f"{blk}.phi.weight"])
    ww = _w2d(prm[f"{blk}.W.weight"])
    xhat, sd_x, sd_t, mu_t, t = cache["xhat"], cache["sd_x"], cache["sd_t"], cache["mu_t"], cache["t"]

    d_mu_t = g.sum(axis=2)
    d_sd_t = (g * xhat).sum(axis=2)
    dxhat = g * sd_t[:, :, None]
    dx0 = (dxhat - dxhat.mean(axis=2, keepdims=True)
           - xhat * (dxhat * xhat).sum(axis=2, keepdims=True) / (n - 1)) / sd_x[:, :, None]
    dt = d_mu_t[:, :, None] / n + d_sd_t[:, :, None] * (t - mu_t[:, :, None]) / ((n - 1) * sd_t[:, :, None])

    grads = {f"{blk}.{cv}.{wb}": 0.0 for cv in CONVS for wb in ("weight", "bias")}
    dx1 = np.zeros_like(x1)
    for i in range(b):
        q, k, v, y, lse, yp = (cache[s][i] for s in ("q", "k", "v", "y", "lse", "yp"))
        grads[f"{blk}.W.weight"] = grads[f"{blk}.W.weight"] + dt[i] @ yp.T
        grads[f"{blk}.W.bias"] = grads[f"{blk}.W.bias"] + dt[i].sum(axis=1)
        dyp = ww.T @ dt[i]                        # (d, N)
        dy = dyp.reshape(n, d)                    # undo the reinterpretation
        dq, dk, dv = _attention_backward(q, k, v, y, lse, dy)
        grads[f"{blk}.g.weight"] = grads[f"{blk}.g.weight"] + dv.T @ x0[i].T
        grads[f"{blk}.g.bias"] = grads[f"{blk}.g.bias"] + dv.sum(axis=0)
        grads[f"{blk}.theta.weight"] = grads[f"{blk}.theta.weight"] + dk @ x1[i].T
        grads[f"{blk}.theta.bias"] = grads[f"{blk}.theta.bias"] + dk.sum(axis=1)
        grads[f"{blk}.phi.weight"] = grads[f"{blk}.phi.weight"] + dq.T @ x1[i].T
        grads[f"{blk}.phi.bias"] = grads[f"{blk}.phi.bias"] + dq.sum(axis=0)
        dx0[i] += wg.T @ dv.T
        dx1[i] += wt.T @ dk + wp.T @ dq.T
    for key in list(grads):
        if key.endswith("weight"):
            grads[key] = grads[key].reshape(grads[key].shape + (1, 1))
    return dx0, dx1, grads


def refusion_forward(a, e, prm):
    """REFusion.forward(rgb=a, evt=e).  a, e (B, C, H, W) -> out (B, 2C, H, W), cache."""
    a = np.asarray(a, dtype=np.float64)
    e = np.asarray(e, dtype=np.float64)
    b, c, h, w = a.shape
    n = h * w
    a2, e2 = a.reshape(b, c, n), e.reshape(b, c, n)
    wr, br = _w2d(prm["conv0_rgb.weight"]), np.asarray(prm["conv0_rgb.bias"], np.float64)
    we, be = _w2d(prm["conv0_evt.weight"]), np.asarray(prm["conv0_evt.bias"], np.float64)
    r0 = np.einsum("oc,bcn->bon", wr, a2) + br[None, :, None]    # fusion.py:23
    e0 = np.einsum("oc,bcn->bon", we, e2) + be[None, :, None]    # fusion.py:24
    m = r0 * e0                                                  # fusion.py:26
    r1, e1 = r0 + m, e0 + m                                      # fusion.py:27-28
    za, cache_a = _cab_forward(r1, e1, prm, BLOCKS[0])           # fusion.py:31
    zb, cache_b = _cab_forward(e1, r1, prm, BLOCKS[1])           # fusion.py:32
    out = np.concatenate([za, zb], axis=1).reshape(b, 2 * c, h, w)  # fusion.py:34
    cache = dict(a=a2, e=e2, r0=r0, e0=e0, r1=r1, e1=e1, blk_a=cache_a, blk_b=cache_b, shape=(b, c, h, w))
    return out, cache


def refusion_backward(dout, cache, prm):
    """dout (B, 2C, H, W) -> da, de (B, C, H, W), grads {state_dict key: array}."""
    b, c, h, w = cache["shape"]
    n = h * w
    dout = np.asarray(dout, dtype=np.float64).reshape(b, 2 * c, n)
    ga, gb = dout[:, :c], dout[:, c:]
    dx0_a, dx1_a, grads_a = _cab_backward(ga, cache["blk_a"], prm, BLOCKS[0])
    dx0_b, dx1_b, grads_b = _cab_backward(gb, cache["blk_b"], prm, BLOCKS[1])
    dr1 = dx0_a + dx1_b
    de1 = dx0_b + dx1_a
    r0, e0 = cache["r0"], cache["e0"]
    dr0 = dr1 * (1.0 + e0) + de1 * e0
    de0 = de1 * (1.0 + r0) + dr1 * r0
    wr, we = _w2d(prm["conv0_rgb.weight"]), _w2d(prm["conv0_evt.weight"])
    grads = {}
    grads.update(grads_a)
    grads.update(grads_b)
    grads["conv0_rgb.weight"] = np.einsum("bon,bcn->oc", dr0, cache["a"]).reshape(c, c, 1, 1)
    grads["conv0_rgb.bias"] = dr0.sum(axis=(0, 2))
    grads["conv0_evt.weight"] = np.einsum("bon,bcn->oc", de0, cache["e"]).reshape(c, c, 1, 1)
    grads["conv0_evt.bias"] = de0.sum(axis=(0, 2))
    da = np.einsum("oc,bon->bcn", wr, dr0).reshape(b, c, h, w)
    de = np.einsum("oc,bon->bcn", we, de0).reshape(b, c, h, w)
    return da, de, grads


def refusion_forward_backward(a, e, prm, dout):
    out, cache = refusion_forward(a, e, prm)
    da, de, grads = refusion_backward(dout, cache, prm)
    return out, da, de, grads


def cross_attention_forward(x0, x1, prm, blk=BLOCKS[0]):
    """CrossAttentionBlock.forward([x0, x1]) on (B, C, H, W) arrays (params keyed '<blk>.<conv>.*')."""
    x0 = np.asarray(x0, dtype=np.float64)
    x1 = np.asarray(x1, dtype=np.float64)
    b, c, h, w = x0.shape
    z, _ = _cab_forward(x0.reshape(b, c, -1), x1.reshape(b, c, -1), prm, blk)
    return z.reshape(b, c, h, w)


def rel_l2(x, ref):
    """l2-relative error used by every parity gate."""
    x = np.asarray(x, dtype=np.float64).ravel()
    ref = np.asarray(ref, dtype=np.float64).ravel()
    den = np.linalg.norm(ref)
    return float(np.linalg.norm(x - ref) / den) if den > 0 else float(np.linalg.norm(x))
