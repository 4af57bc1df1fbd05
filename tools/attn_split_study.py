#!/usr/bin/env python
"""Error study on the CPU (numpy, no GPU): what would cheaper correction products cost the attention in accuracy?

The tcgen05 attention kernels form every fp32 product as three `kind::tf32` MMAs (hi.hi + hi.lo + lo.hi, hi = the
operand truncated to tf32, lo = the rest rounded to tf32).  VERDICT r1 item 4(i) proposes to issue the two 2^-11-sized
correction products as `kind::f16` MMAs on bf16 images (K = 16 per instruction at the cost of K = 8 in tf32: 2 units of
tensor time per product instead of 3).  This script emulates both schemes' OPERAND arithmetic exactly (products of tf32
/ bf16 values are exact in fp32; accumulation is done in float64 here, so what is measured is the operand rounding
alone, a lower bound for both) through the whole attention of one CrossAttentionBlock, forward and backward:

    S = Q K^T, P = softmax(S), Y = P V, dV = P^T dY, dP = dY V^T, dS = P (dP - rowsum(P dP)), dQ = dS K, dK = dS^T Q

with Q, K, V produced like the block does (1x1 projections of ReLU feature maps, Attention.py:24-27), and prints the
l2-relative error of Y, dQ, dK, dV against float64 for
    split3    hi.hi + hi.lo + lo.hi in tf32                      (what the kernels do)
    bf16corr  hi.hi in tf32 + bf16(a).bf16(lo_b) + bf16(lo_a).bf16(hi_b)   (the proposal)
    single    hi.hi only                                          (the bf16 inference path)

    python tools/attn_split_study.py [N] [d] [C] [seeds]
"""
import sys

import numpy as np


def trunc_tf32(x):
    u = np.asarray(x, np.float32).view(np.uint32) & np.uint32(0xFFFFE000)
    return u.view(np.float32)


def rn_tf32(x):
    u = np.asarray(x, np.float32).view(np.uint32)
    return ((u + np.uint32(0x1000)) & np.uint32(0xFFFFE000)).view(np.float32)


def rn_bf16(x):
    u = np.asarray(x, np.float32).view(np.uint32)
    r = u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))
    return (r & np.uint32(0xFFFF0000)).view(np.float32)


def split(x):
    x = np.asarray(x, np.float32)
    hi = trunc_tf32(x)
    return hi, x - hi                      # exact remainder (13 bits)


def mm(a, b, scheme):
    """a (m, k) @ b (k, n) with the operands of the given scheme, float64 accumulation."""
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    if scheme == "fp64":
        return a.astype(np.float64) @ b.astype(np.float64)
    ah, ar = split(a); bh, br = split(b)
    f = lambda t: t.astype(np.float64)
    out = f(ah) @ f(bh)
    if scheme == "single":
        return out
    if scheme == "split3":
        return out + f(ah) @ f(rn_tf32(br)) + f(rn_tf32(ar)) @ f(bh)
    if scheme == "bf16corr":
        # a.lo_b carries lo_a.lo_b along for free (bf16(a), not bf16(hi_a)); the other term uses hi_b so that it is
        # not counted twice
        return out + f(rn_bf16(a)) @ f(rn_bf16(br)) + f(rn_bf16(ar)) @ f(rn_bf16(bh))
    raise ValueError(scheme)


def attention(q, k, v, dy, scheme):
    """q, k, v, dy: (N, d).  Every matmul operand goes through the scheme, elementwise work is float64 -> float32
    where the kernels hold fp32 values (P, dS are re-split on chip)."""
    s = mm(q, k.T, scheme)
    s = s - s.max(1, keepdims=True)
    p = np.exp(s); p /= p.sum(1, keepdims=True)
    p32 = p.astype(np.float32)
    y = mm(p32, v, scheme)
    dv = mm(p32.T, dy, scheme)
    dp = mm(dy, v.T, scheme)
    delta = (p * dp).sum(1, keepdims=True)
    ds = (p * (dp - delta)).astype(np.float32)
    dq = mm(ds, k, scheme)
    dk = mm(ds.T, q, scheme)
    return y, dq, dk, dv


def rel(x, ref):
    return float(np.linalg.norm(x - ref) / np.linalg.norm(ref))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1200
    d = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    c = int(sys.argv[3]) if len(sys.argv) > 3 else 256
    seeds = int(sys.argv[4]) if len(sys.argv) > 4 else 3
    print(f"N = {n}, d = {d}, C = {c}; l2-relative error vs float64 (operand rounding only, float64 accumulation)")
    worst = {}
    for seed in range(seeds):
        g = np.random.default_rng(seed)
        x0 = np.maximum(g.standard_normal((c, n)), 0).astype(np.float32)        # ReLU feature maps (C, N)
        x1 = np.maximum(g.standard_normal((c, n)), 0).astype(np.float32)
        w = lambda: (g.standard_normal((d, c)) * np.sqrt(2.0 / c)).astype(np.float32)   # xavier-like, Attention.py:16
        bias = lambda: (0.1 * g.standard_normal((d, 1))).astype(np.float32)
        v = (w().astype(np.float64) @ x0 + bias()).T.astype(np.float32)         # g(x0)      (N, d)
        k = (w().astype(np.float64) @ x1 + bias()).T.astype(np.float32)         # theta(x1)
        q = (w().astype(np.float64) @ x1 + bias()).T.astype(np.float32)         # phi(x1)
        dy = g.standard_normal((n, d)).astype(np.float32)
        ref = attention(q, k, v, dy, "fp64")
        for scheme in ("split3", "bf16corr", "single"):
            got = attention(q, k, v, dy, scheme)
            errs = [rel(a, b) for a, b in zip(got, ref)]
            print(f"  seed {seed} {scheme:9s} Y {errs[0]:.2e}  dQ {errs[1]:.2e}  dK {errs[2]:.2e}  dV {errs[3]:.2e}")
            worst[scheme] = max(worst.get(scheme, 0.0), max(errs))
    print("worst:", ", ".join(f"{k} {v:.2e}" for k, v in worst.items()), " (gate for the whole block: 1e-5)")


if __name__ == "__main__":
    main()
