"""GPU bring-up of the whole-stack kernels: same context config through the per-layer kernels (kernel_path) and the stack kernels.

    python scripts/stack_check.py [T] [NL] [gatys 0/1] [B]
Run under `timeout`: a protocol bug in a persistent kernel shows up as a hang.
"""
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
B = int(sys.argv[4]) if len(sys.argv) > 4 else 1
REF_MODE = sys.argv[5] if len(sys.argv) > 5 else "per_layer"      # or "per_layer_bwd": stack forward + per-layer backward
w = synth.make_weights(NL, 0, 0.1)
torch.manual_seed(0)
x = torch.randn(B, T, device='cuda') * 40
cont, style = [NL - 1], list(range(NL)) if gatys else list(range(min(NL, 10)))


def make(no_stack):
    st = StyleStep(T, cont, style, gatys, precision='bf16', B=B, kernel_path=REF_MODE if no_stack else "auto")
    st.set_weights(w)
    return st


def chk(st, what):
    try:
        st.check()
        print(f'check after {what}: ok', flush=True)
    except Exception as e:  # noqa: BLE001
        print(f'check after {what}: {e}', flush=True)


ref, tc = make(True), make(False)
print('uses_stack ref/new:', ref.uses_stack_kernels, tc.uses_stack_kernels, flush=True)
Ec_r, Gn_r = ref.embeds(x)
Ec_t, Gn_t = tc.embeds(x)
chk(tc, 'embeds')
for l in range(-1, NL):
    a, b = tc.extract(l), ref.extract(l)
    print(f"layer {l:2d}: rel {rel(a, b):.3e}  max|ref| {float(b.abs().max()):.3g} nan={bool(torch.isnan(a).any())}")
print('Ec rel', rel(Ec_t, Ec_r), 'Gn rel', rel(Gn_t, Gn_r), flush=True)
for st in (ref, tc):
    st.set_targets(Ec_r * 0.9, Gn_r.roll(1, 1), 100.0)
lo_r, g_r = ref.step(x)
lo_t, g_t = tc.step(x)
chk(tc, 'step')
print('losses ref', lo_r.tolist(), 'new', lo_t.tolist())
print('grad rel', rel(g_t, g_r), 'launches ref/new', ref.launches_per_step, tc.launches_per_step)
lo = torch.empty(B, 4, device='cuda')
g = torch.empty(B, T, device='cuda')
for name, st in (('per-layer', ref), ('stack', tc)):
    for _ in range(3):
        st.step_graph(x, lo, g)
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(50):
        st.step_graph(x, lo, g)
    torch.cuda.synchronize()
    print(f'{name:10s} ms/step {(time.time() - t0) / 50 * 1e3:.4f}  phases {st.step_timed(x, lo, g)}', flush=True)
lo2, g2 = tc.step(x)
print('graph replay equals eager:', bool(torch.equal(g2, g)))
chk(tc, 'graph replays')
