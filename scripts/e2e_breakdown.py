"""Where the end-to-end evaluation (ast_step_host: host float64 in / out) spends its time next to the device-resident step."""
import sys, time
from pathlib import Path
import numpy as np
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from audio_style_transfer_b200 import synth
from audio_style_transfer_b200.step import StyleStep

T, NL = 16384, 30
w = synth.make_weights(NL, seed=0)
st = StyleStep(T, [NL - 1], list(range(NL)), True, precision="bf16")
st.set_weights(w)
rng = np.random.default_rng(0)
Pc = rng.standard_normal((1, T, 128)).astype(np.float32)
Ps = rng.standard_normal(st.gn_shape).astype(np.float32)
Ps /= np.sqrt((Ps ** 2).sum(axis=(-2, -1), keepdims=True))
st.set_targets(Pc, Ps, 100.0)
x = (rng.standard_normal((1, T)) * 40).astype(np.float64)
xd = torch.as_tensor(x, dtype=torch.float32, device="cuda")
lo = torch.empty((1, 4), device="cuda"); g = torch.empty((1, T), device="cuda")
N = 300
for _ in range(20):
    st.step_host(x); st.step_graph(xd, lo, g); torch.cuda.synchronize()   # one evaluation of a context at a time
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(N):
    st.step_host(x)
t_host = (time.perf_counter() - t0) / N * 1e6
t0 = time.perf_counter()
for _ in range(N):
    st.step_graph(xd, lo, g); torch.cuda.synchronize()
t_graph_sync = (time.perf_counter() - t0) / N * 1e6
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(N):
    st.step_graph(xd, lo, g)
e1.record(); torch.cuda.synchronize()
t_dev = e0.elapsed_time(e1) / N * 1e3
t0 = time.perf_counter()
for _ in range(N):
    xf = x.astype(np.float32); gd = np.empty((1, T), np.float64); gd[:] = xf
t_conv = (time.perf_counter() - t0) / N * 1e6
print(f"step_host {t_host:.1f} us | step_graph + sync per call {t_graph_sync:.1f} us | device back-to-back {t_dev:.1f} us | numpy f64<->f32 of T {t_conv:.1f} us")
