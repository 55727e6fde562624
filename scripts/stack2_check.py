"""GPU bring-up of the two-pairs-per-CTA stack kernels: same B-pair context with AST_B200_DUO unset (default) and AST_B200_DUO=1.

    python scripts/stack2_check.py [T] [NL] [gatys 0/1] [B]        (run under `timeout`)
"""
import os
import sys
import time

import torch

sys.path.insert(0, '.')
from audio_style_transfer_b200 import synth  # noqa: E402
from audio_style_transfer_b200.step import StyleStep  # noqa: E402


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


T = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
NL = int(sys.argv[2]) if len(sys.argv) > 2 else 12
gatys = bool(int(sys.argv[3])) if len(sys.argv) > 3 else True
B = int(sys.argv[4]) if len(sys.argv) > 4 else 2
w = synth.make_weights(NL, 0, 0.1)
torch.manual_seed(0)
x = torch.randn(B, T, device='cuda') * 40
cont, style = [NL - 1], list(range(NL)) if gatys else list(range(min(NL, 10)))
res = {}
for mode in ("1", "0"):                      # "1": single-pair kernels (default), "0": duo
    os.environ["AST_B200_DUO"] = "0" if mode == "1" else "1"
    st = StyleStep(T, cont, style, gatys, precision='bf16', B=B)
    st.set_weights(w)
    Ec, Gn = st.embeds(x)
    torch.cuda.synchronize()
    ex = [st.extract(l).clone() for l in range(-1, NL)]
    if mode == "1":
        tgt = (Ec * 0.9, Gn.roll(1, 1))
    st.set_targets(tgt[0], tgt[1], 100.0)
    lo, g = st.step(x)
    torch.cuda.synchronize()
    lo2 = torch.empty_like(lo); g2 = torch.empty_like(g)
    for _ in range(3):
        st.step_graph(x, lo2, g2)
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(30):
        st.step_graph(x, lo2, g2)
    torch.cuda.synchronize()
    ms = (time.time() - t0) / 30 * 1e3
    res[mode] = dict(Ec=Ec.clone(), Gn=Gn.clone(), ex=ex, lo=lo.clone(), g=g.clone(), ms=ms, err=st.stack_error,
                     launches=st.launches_per_step, phases=st.step_timed(x, lo2, g2), same=bool(torch.equal(g2, g)))
    st.close()
a, b = res["1"], res["0"]
for i, l in enumerate(range(-1, NL)):
    e = rel(b["ex"][i], a["ex"][i])
    if e > 1e-4 or l in (-1, 0, NL - 1):
        print(f"layer {l:2d}: duo vs single rel {e:.3e}")
print('Ec rel', rel(b["Ec"], a["Ec"]), 'Gn rel', rel(b["Gn"], a["Gn"]))
print('losses single', a["lo"].tolist(), 'duo', b["lo"].tolist())
print('grad rel', rel(b["g"], a["g"]), 'stack_err single/duo', a["err"], b["err"], 'graph==eager', a["same"], b["same"])
print(f'single-pair kernels: {a["ms"]:.4f} ms/step ({B / a["ms"] * 1e3:.0f} pair-steps/s) launches {a["launches"]} phases {a["phases"]}')
print(f'duo kernels        : {b["ms"]:.4f} ms/step ({B / b["ms"] * 1e3:.0f} pair-steps/s) launches {b["launches"]} phases {b["phases"]}')
