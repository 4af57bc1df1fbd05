"""Experiment: error of the attention core on BLOCK-SHAPED data (Q, K, V and dY taken from an fp64 autograd
run of the fusion block), per engine, against fp64.  Prints l2-relative errors and the error's projection
on the reference (a systematic scale error shows up as |proj| ~ rel)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.nn.functional as F
from helpers import make_inputs, make_params
from test_attn_gpu import attn_fwd, attn_bwd

DEV = "cuda:0"


def stats(x, ref):
    e = x.double() - ref
    rel = float(e.norm() / ref.norm())
    proj = float((e * ref).sum() / (ref * ref).sum())
    return f"rel {rel:.2e} proj {proj:+.2e}"


def block_tensors(c, h, w, kind, blk):
    a, e, dout = (t.to(DEV) for t in make_inputs(kind, (1, c, h, w), seed=1000 + c))
    p = {k: v.to(DEV) for k, v in make_params(c, seed=0).items()}
    conv = lambda x, k: F.conv2d(x, p[k + ".weight"], p[k + ".bias"])
    r0, e0 = conv(a, "conv0_rgb"), conv(e, "conv0_evt")
    m = r0 * e0
    r1, e1 = r0 + m, e0 + m
    x0, x1 = (r1, e1) if blk == "rgb_cross_attention" else (e1, r1)
    n, d = h * w, c // 8
    V = conv(x0, blk + ".g").view(1, d, n).detach().requires_grad_(True)
    K = conv(x1, blk + ".theta").view(1, d, n).detach().requires_grad_(True)
    Q = conv(x1, blk + ".phi").view(1, d, n).detach().requires_grad_(True)
    S = Q.transpose(1, 2) @ K
    Y = torch.softmax(S, -1) @ V.transpose(1, 2)
    Y.retain_grad()
    T = conv(Y.contiguous().view(1, d, h, w), blk + ".W")
    mu_t, sd_t = T.mean((2, 3), keepdim=True), (T.var((2, 3), keepdim=True) + 1e-5).sqrt()
    mu_x, sd_x = x0.mean((2, 3), keepdim=True), (x0.var((2, 3), keepdim=True) + 1e-5).sqrt()
    z = (x0 - mu_x) / sd_x * sd_t + mu_t
    half = dout[:, :c] if blk == "rgb_cross_attention" else dout[:, c:]
    z.backward(half)
    return Q, K, V, Y, Y.grad, x0.detach(), x1.detach()


for (c, h, w, kind) in [(512, 60, 80, "randn"), (256, 120, 160, "relu"), (512, 33, 44, "relu")]:
    for blk in ("rgb_cross_attention", "event_cross_attention"):
        Q, K, V, Y, dY, x0, x1 = block_tensors(c, h, w, kind, blk)
        q32, k32, v32, dy32 = (t.detach().float().contiguous() for t in (Q, K, V, dY))
        # reference gradients for the fp32-rounded operands, in fp64
        q, k, v = (t.double().requires_grad_(True) for t in (q32, k32, v32))
        Yr = torch.softmax(q.transpose(1, 2) @ k, -1) @ v.transpose(1, 2)
        Yr.backward(dy32.double())
        ref = (q.grad.transpose(1, 2), k.grad.transpose(1, 2), v.grad.transpose(1, 2))
        for eng in (0, 1):
            y, lse = attn_fwd(eng, q32, k32, v32)
            g = attn_bwd(eng, q32, k32, v32, y, lse, dy32)
            print(f"C={c} {h}x{w} {kind} {blk[:3]} {['ffma', 'tc  '][eng]}: Y {stats(y, Yr.detach())} | "
                  f"dQ {stats(g[0], ref[0])} | dK {stats(g[1], ref[1])} | dV {stats(g[2], ref[2])}"
                  f" | sum_n dQ {stats(g[0].double().sum(1), ref[0].sum(1))}")
            # weight gradients formed in fp64 from the engine's fp32 dK / dQ (isolates the attention error)
            X1 = x1.view(1, c, -1).double()                      # (1, C, N)
            for nm, gi, ri in (("theta.w", g[1], ref[1]), ("phi.w", g[0], ref[0])):
                w_ref = ri.transpose(1, 2) @ X1.transpose(1, 2)     # (1, d, C)
                gd = gi.double()
                w_eng = gd.transpose(1, 2) @ X1.transpose(1, 2)
                gz = gd - gd.mean(1, keepdim=True) if nm == "theta.w" else gd
                w_prj = gz.transpose(1, 2) @ X1.transpose(1, 2)
                print(f"      {nm}: wgrad-from-engine-grad {stats(w_eng, w_ref)} | after zero-mean projection {stats(w_prj, w_ref)}"
                      f" | |sum_n g| / sum_n |g| = {float(gd.sum(1).abs().sum() / gd.abs().sum()):.2e}")
