"""Small target for `ncu --set full`: 3 steps of cfg1 (Gatys, whole-stack kernels) and 3 of the CLI-default workload
(all 30 layers channel-wise: gram_ours_tma / style_grad_ours_tma).  No oracle, no timing: see B200_PROFILING.md."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from audio_style_transfer_b200 import synth  # noqa: E402
from audio_style_transfer_b200.step import StyleStep  # noqa: E402

T, NL = 16384, 30
w = synth.make_weights(NL, seed=0, bias_scale=0.0)
x = torch.randn(1, T, device="cuda") * 40
for gatys in (True, False):
    st = StyleStep(T, [NL - 1], list(range(NL)), gatys, precision="bf16")
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
    print("gatys" if gatys else "ours", lo.cpu().numpy().tolist(), st.launches_per_step)
    st.close()
