"""Accurate tensor-core tier (AST_PREC_TC3) against the fp64 oracle: features, loss, raw and mask-imposed gradient distance."""
import sys
from pathlib import Path
import numpy as np
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from audio_style_transfer_b200.step import StyleStep
from oracle import cases

rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / np.linalg.norm(b))
for name in sys.argv[1:] or ["t1024_ours_s0", "t2048_gatys_s1", "t4096_cfg1", "t16384_cfg0", "t16384_cfg1"]:
    case = cases.build_inputs(name)
    Pc, Ps = cases.oracle_targets(case)
    losses, dx, aux = cases.oracle_step(case, Pc, Ps, return_aux=True)
    for prec in ("tc3", "bf16"):
        st = StyleStep(case["T"], case["cont"], case["style"], case["gatys"], case["nb"], case["cnt"], precision=prec)
        st.set_weights(case["w"])
        st.set_targets(Pc[None], Ps[None], case["lambd"])
        lo, g = st.step(case["x"][None])
        torch.cuda.synchronize()
        st.check()
        lo, g = lo.cpu().numpy()[0], g.cpu().numpy()[0]
        hm, zm = st.masks()
        nflip = sum(int((hm[l, 0] != (aux["H"][l] > 0)).sum() + (zm[l, 0] != (aux["Z"][l] > 0)).sum()) for l in range(case["NL"]))
        _, dxm = cases.oracle_step(case, Pc, Ps, masks=(hm[:, 0], zm[:, 0]), fwd=(aux["H"], aux["Z"]))
        ex = st.extract(case["NL"] - 1)[0].cpu().numpy()
        print(f"{name}[{prec}]: last extract rel {rel(ex, aux['H'][case['NL']]):.2e}  loss rel {abs(lo[0] - losses[0]) / abs(losses[0]):.2e}  "
              f"grad raw {rel(g, dx):.2e}  with device masks {rel(g, dxm):.2e}  ({nflip} sign bits differ)  launches {st.launches_per_step}", flush=True)
        st.close()
