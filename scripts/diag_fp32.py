"""Where does the fp32 tier's gradient error at cfg0 size come from?  (GPU box; test infrastructure, uses the oracle.)
Replaces, one at a time, the oracle's fp64 raw Gram / content extract by the GPU's fp32 ones and reports the gradient
rel-L2 that alone would cause."""
import sys
from pathlib import Path
import numpy as np
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from audio_style_transfer_b200.step import StyleStep
from oracle import cases, step_np

name = sys.argv[1] if len(sys.argv) > 1 else "t16384_cfg0"
prec = sys.argv[2] if len(sys.argv) > 2 else "fp32"
case = cases.build_inputs(name)
Pc, Ps = cases.oracle_targets(case)
losses, dx, aux = cases.oracle_step(case, Pc, Ps, return_aux=True)
rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / np.linalg.norm(b))
st = StyleStep(case["T"], case["cont"], case["style"], case["gatys"], precision=prec, device=0)
st.set_weights(case["w"])
st.set_targets(Pc[None], Ps[None], case["lambd"])
lo, g = st.step(case["x"][None])
torch.cuda.synchronize()
print(f"{name}[{prec}] grad rel-L2 {rel(g.cpu().numpy()[0], dx):.3e} loss rel {abs(float(lo[0,0]) - losses[0]) / abs(losses[0]):.2e}")
st.forward_partial(case["x"][None])
torch.cuda.synchronize()
part = st.partials().cpu().numpy().astype(np.float64)
G = aux["G"]
Gg = part[:G.size].reshape(G.shape)
print("raw Gram rel err (all)", rel(Gg, G), "max per matrix", max(rel(Gg[i], G[i]) for i in range(G.shape[0])))
cos = [float((aux["Gn"][i] * Ps[i]).sum()) for i in range(Ps.shape[0])]
print("cos(Gn, Ps): min %.6f median %.6f max %.8f" % (min(cos), np.median(cos), max(cos)))
for l in sorted(set([min(i, 29) for i in case["cont"] + case["style"]])):
    h = st.extract(l).cpu().numpy()[0]
    print(f"  extract {l}: rel err {rel(h, aux['H'][l + 1]):.3e}")
# oracle with the GPU's G substituted
import types
orig = step_np.raw_gram
step_np.raw_gram = lambda H, ids, gatys: Gg.copy()
_, dx_g = cases.oracle_step(case, Pc, Ps)
step_np.raw_gram = orig
print(f"oracle with the GPU's raw Gram: grad rel-L2 {rel(dx_g, dx):.3e}")
for lam in (0.0, 1e6):
    case2 = dict(case, lambd=lam)
    l2, dx2 = cases.oracle_step(case2, Pc, Ps)
    st.set_targets(Pc[None], Ps[None], lam)
    lo2, g2 = st.step(case["x"][None])
    torch.cuda.synchronize()
    g2 = g2.cpu().numpy()[0].astype(np.float64)
    e = g2 - dx2
    blk = np.sqrt((e.reshape(-1, 1024) ** 2).sum(1)) / np.linalg.norm(dx2)
    print(f"lambd {lam:g}: grad rel-L2 {rel(g2, dx2):.3e}; per-1024 block share max {blk.max():.2e} at {int(blk.argmax())}, median {np.median(blk):.2e}")
    top = np.argsort(-np.abs(e))[:8]
    print("   largest |err| at t =", sorted(top.tolist()), "values", np.abs(e)[top].max() / np.abs(dx2).max())
