"""A few evaluations of the device-resident optimiser on cfg1 (for `ncu --metrics gpu__time_duration.sum`: the launch list
shows lbfgs_advance_kernel next to the step's kernels) and a CUDA-event timing of the advance kernel alone.

    python scripts/opt_profile.py [evals] [workload]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench                                                        # noqa: E402
from audio_style_transfer_b200.step import DeviceLBFGS, StyleStep   # noqa: E402

evals = int(sys.argv[1]) if len(sys.argv) > 1 else 24
name = sys.argv[2] if len(sys.argv) > 2 else "cfg1"
wl = bench.build_workload(name, 1)
dev = torch.device("cuda", 0)
st = StyleStep(wl["T"], wl["cont"], wl["style"], wl["gatys"], precision="bf16", device=0)
st.set_weights(wl["w"])
Pc, Ps = bench.gpu_targets(st, torch.as_tensor(wl["content"], device=dev), torch.as_tensor(wl["style_a"], device=dev))
st.set_targets(Pc, Ps, wl["lambd"])
opt = DeviceLBFGS(wl["T"], maxiter=100, epochs=1, min_evals=0, factr=10.0, max_log=evals + 8)
opt.reset(wl["x"].astype(np.float64))
n = opt.minimize(st, max_evals=evals)
status, fk = opt.status()
print(f"{name}: enqueued {n}, status {status[0].tolist()}, f {fk[0]:.6f}")

stamps = torch.zeros(16, dtype=torch.int64, device=dev)
import ctypes
st.lib.ast_opt_debug_stamps.argtypes = [ctypes.c_void_p]
st.lib.ast_opt_debug_stamps(ctypes.c_void_p(stamps.data_ptr()))
stamp_rows = []
# the advance kernel alone: hand-driven loop, events around every advance (the machine really moves: the step runs between)
a = [torch.cuda.Event(enable_timing=True) for _ in range(evals)]
b = [torch.cuda.Event(enable_timing=True) for _ in range(evals)]
opt.reset(wl["x"].astype(np.float64))
kinds = []
for k in range(evals):
    st.step_graph(opt.x, opt.losses, opt.grad)
    a[k].record()
    opt.advance()
    b[k].record()
    kinds.append(int(opt.status()[0][0, 7]))
    stamp_rows.append(stamps.cpu().numpy().copy())                          # line-search evaluations so far (1 = a new direction was just formed)
torch.cuda.synchronize()
ms = np.array([x.elapsed_time(y) for x, y in zip(a, b)])
new_dir = np.array(kinds) == 1
print(f"advance kernel: new direction (two-loop, history up to 10) median {np.median(ms[new_dir]) * 1e3:.1f} us over {new_dir.sum()}; "
      f"line-search continue median {np.median(ms[~new_dir]) * 1e3 if (~new_dir).any() else float('nan'):.1f} us over {(~new_dir).sum()}")
names = ["state", "prefetch+gd pass+combine", "dcsrch", "pass A", "pass B", "combine62", "gram upd", "two-loop", "pass C", "combine2", "ls start+trial", "writeback"]
for k in (20, 25):
    r = stamp_rows[k]
    print(f"advance {k} (kind {kinds[k]}): cycles per phase:", {n: int(r[i + 1] - r[i]) for i, n in enumerate(names)}, "total", int(r[12] - r[0]))
print("per advance (us):", np.round(ms * 1e3, 1).tolist())
