"""ORACLE -- TEST INFRASTRUCTURE ONLY.  CPU emulation of tensor-core operand precisions for the step.

    python -m oracle.precision_study [case] [fwd_scheme,bwd_scheme ...]

Runs the step of oracle/step_np.py with every GEMM operand rounded the way a tensor-core mode would
round it (fp32 accumulate, as TMEM does) and reports loss / gradient rel-L2 error against the fp64
oracle.  This is how the operand formats of the CUDA tiers were chosen (DESIGN.md "accuracy tiers")
without spending GPU time.  Schemes: f32, bf16, f16, tf32, and 3-term splits bf16x3 / f16x3
(x = hi + lo; products hi*hi + lo*hi + hi*lo).
"""
from __future__ import annotations

import sys

import numpy as np

from . import cases, step_np


def q_bf16(x):
    u = np.ascontiguousarray(x, np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def q_tf32(x):
    u = np.ascontiguousarray(x, np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0xFFF + ((u >> 13) & 1)) & 0xFFFFE000
    return u.astype(np.uint32).view(np.float32)


def q_f16(x):
    return np.asarray(x, np.float32).astype(np.float16).astype(np.float32)


Q = {"bf16": q_bf16, "f16": q_f16, "tf32": q_tf32}


def make_mm(scheme):
    """matmul(a, w) with operands rounded per `scheme`, fp32 accumulate."""
    if scheme == "f32":
        return lambda a, w: a.astype(np.float32) @ w.astype(np.float32)
    if scheme in Q:
        q = Q[scheme]
        return lambda a, w: q(a) @ q(w)
    if scheme.endswith("x3") or scheme.endswith("x2"):
        q = Q[scheme[:-2]]
        three = scheme.endswith("x3")

        def mm(a, w):
            a = a.astype(np.float32); w = w.astype(np.float32)
            ah = q(a); al = q(a - ah)
            wh = q(w)
            out = ah @ wh + al @ wh
            if three:
                out = out + ah @ q(w - wh)
            return out
        return mm
    raise KeyError(scheme)


def emulated_step(case, Pc, Ps, fwd, bwd, gram="f32"):
    """Same algebra as step_np.step, fp32 everywhere, GEMM operands rounded per scheme."""
    f32 = np.float32
    w = case["w"]
    mmf, mmb, mmg = make_mm(fwd), make_mm(bwd), make_mm(gram)
    x = case["x"].astype(f32)
    T, C = x.shape[0], 128
    NL = case["NL"]
    sh = step_np.shift_rows
    xs = (x / f32(128.0))[:, None]
    h = step_np.conv3(xs, w.W0.astype(f32), w.b0.astype(f32), 1)
    H, Z = [h], []
    for l in range(NL):
        d = step_np.dilation(l)
        a = np.maximum(h, 0)
        z = mmf(sh(a, -d), w.Wd[l][0]) + mmf(a, w.Wd[l][1]) + mmf(sh(a, d), w.Wd[l][2]) + w.bd[l].astype(f32)
        r = np.maximum(z, 0)
        h = h + (mmf(r, w.Wr[l]) + w.br[l].astype(f32))
        H.append(h); Z.append(z)
    ext = lambda i: H[min(i, 29) + 1]
    cnt = case["cnt"]
    Ec = np.concatenate([ext(i)[:, :cnt] for i in case["cont"]], axis=1)
    dEc = Ec.astype(np.float64) - Pc
    Lc = 10.0 * np.mean(dEc * dEc)
    gatys = case["gatys"]
    if gatys:
        G = np.stack([mmg(ext(i).T.copy(), ext(i)) for i in case["style"]]).astype(np.float64)
    else:
        S = np.stack([ext(i) for i in case["style"]]).transpose(2, 0, 1)
        G = (S @ S.transpose(0, 2, 1)).astype(np.float64)
    ss = np.sum(G * G, axis=(1, 2), keepdims=True)
    inv = 1.0 / np.sqrt(np.maximum(ss, 1e-12))
    Gn = G * inv
    nb = Gn.shape[0] if (gatys or case["nb"] >= 128) else case["nb"]
    dGs = Gn[:nb] - Ps
    Ls = 1e3 * np.mean(dGs * dGs)
    loss = Lc + case["lambd"] * Ls
    dGn = np.zeros_like(Gn)
    dGn[:nb] = case["lambd"] * 1e3 * 2.0 * dGs / dGs.size
    dot = np.sum(dGn * Gn, axis=(1, 2), keepdims=True)
    dG = (dGn - Gn * dot) * inv
    Ssym = (dG + dG.transpose(0, 2, 1)).astype(f32)
    dF = {}

    def acc(i, g):
        i = min(i, 29)
        dF[i] = dF[i] + g if i in dF else g
    for slot, i in enumerate(case["cont"]):
        g = np.zeros((T, C), f32)
        g[:, :cnt] = ((10.0 * 2.0 / dEc.size) * dEc[:, slot * cnt:(slot + 1) * cnt]).astype(f32)
        acc(i, g)
    if gatys:
        for li, i in enumerate(case["style"]):
            acc(i, mmg(ext(i), Ssym[li]))
    else:
        F = np.stack([ext(i) for i in case["style"]])
        dFs = np.einsum('clm,mtc->ltc', Ssym, F)
        for li, i in enumerate(case["style"]):
            acc(i, dFs[li])
    g = np.zeros((T, C), f32)
    for l in range(NL - 1, -1, -1):
        if l in dF:
            g = g + dF[l]
        d = step_np.dilation(l)
        dz = mmb(g, w.Wr[l].T.copy()) * (Z[l] > 0)
        da = mmb(sh(dz, d), w.Wd[l][0].T.copy()) + mmb(dz, w.Wd[l][1].T.copy()) + mmb(sh(dz, -d), w.Wd[l][2].T.copy())
        g = g + da * (H[l] > 0)
    dxs = step_np.conv3_T(g, w.W0.astype(f32), 1)[:, 0]
    return (loss, Lc, Ls), dxs / f32(128.0)


def main(argv):
    name = argv[0] if argv else "t4096_cfg1"
    schemes = [tuple(a.split(",")) for a in argv[1:]] or [
        ("f32", "f32"), ("bf16", "bf16"), ("f16", "bf16"), ("tf32", "bf16"), ("bf16x3", "bf16"),
        ("f16x3", "bf16"), ("f16x3", "bf16x3"), ("f16x2", "bf16"), ("f16x3", "f32"), ("f32", "bf16"), ("f32", "bf16x3")]
    case = cases.build_inputs(name)
    Pc, Ps = cases.oracle_targets(case)
    losses, dx = cases.oracle_step(case, Pc, Ps)
    print(f"case {name}: loss {losses[0]:.6g} |dx| {np.linalg.norm(dx):.4g}")
    for s in schemes:
        fwd, bwd = s[0], s[1]
        gram = s[2] if len(s) > 2 else "f32"
        lo, g = emulated_step(case, Pc, Ps, fwd, bwd, gram)
        print(f"  fwd {fwd:7s} bwd {bwd:7s} gram {gram:5s}: loss rel {abs(lo[0] - losses[0]) / abs(losses[0]):.2e}  "
              f"grad rel-L2 {np.linalg.norm(g - dx) / np.linalg.norm(dx):.2e}", flush=True)


if __name__ == "__main__":
    main(sys.argv[1:])
