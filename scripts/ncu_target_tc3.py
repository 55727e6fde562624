"""Small target for `ncu --set full`: 3 steps of cfg1 in the bf16 tier, then 3 in the accurate tier (AST_PREC_TC3: hi / lo Gatys heads)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from audio_style_transfer_b200 import synth  # noqa: E402
from audio_style_transfer_b200.step import StyleStep  # noqa: E402

T, NL = 16384, 30
w = synth.make_weights(NL, seed=0, bias_scale=0.0)
torch.manual_seed(0)
x = torch.randn(1, T, device="cuda") * 40
for prec in ("bf16", "tc3"):
    st = StyleStep(T, [NL - 1], list(range(NL)), True, precision=prec)
    st.set_weights(w)
    rng = np.random.default_rng(0)
    Pc = rng.standard_normal((1, T, 128)).astype(np.float32)
    Ps = rng.standard_normal(st.gn_shape).astype(np.float32)
    Ps /= np.sqrt((Ps ** 2).sum(axis=(-2, -1), keepdims=True))
    st.set_targets(Pc, Ps, 100.0)
    for _ in range(3):
        lo, g = st.step(x)
    torch.cuda.synchronize()
    st.check()
    print(prec, lo.cpu().numpy().tolist(), st.launches_per_step)
    st.close()
