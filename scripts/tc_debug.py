"""GPU bring-up of the tcgen05 path: bf16 context vs fp32 context, layer by layer (both through the C ABI)."""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
from audio_style_transfer_b200.step import StyleStep
from audio_style_transfer_b200 import synth

def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
NL = int(sys.argv[2]) if len(sys.argv) > 2 else 12
w = synth.make_weights(NL, 0, 0.1)
x = torch.randn(1, T, device='cuda') * 40
cont, style = [NL - 1], list(range(min(NL, 10)))
ref = StyleStep(T, cont, style, False, precision='fp32'); ref.set_weights(w)
tc = StyleStep(T, cont, style, False, precision='bf16'); tc.set_weights(w)
Ec_r, Gn_r = ref.embeds(x)
Ec_t, Gn_t = tc.embeds(x)
torch.cuda.synchronize()
for l in range(-1, NL):
    a, b = tc.extract(l), ref.extract(l)
    print(f"layer {l:2d}: rel {rel(a, b):.3e}  max|ref| {float(b.abs().max()):.3g} nan={bool(torch.isnan(a).any())}")
print('Ec rel', rel(Ec_t, Ec_r), 'Gn rel', rel(Gn_t, Gn_r))
for st in (ref, tc):
    st.set_targets(Ec_r * 0.9, Gn_r.roll(1, 1), 100.0)
lo_r, g_r = ref.step(x); lo_t, g_t = tc.step(x)
torch.cuda.synchronize()
print('losses ref', lo_r.tolist(), 'tc', lo_t.tolist())
print('grad rel', rel(g_t, g_r), 'launches', tc.launches_per_step)
import time
lo = torch.empty(1, 4, device='cuda'); g = torch.empty(1, T, device='cuda')
for _ in range(3): tc.step(x, lo, g)
torch.cuda.synchronize(); t0 = time.time()
for _ in range(20): tc.step(x, lo, g)
torch.cuda.synchronize(); print('tc ms/step', (time.time() - t0) / 20 * 1e3, 'phases', tc.step_timed(x, lo, g))
