"""Timeline of the whole-stack forward kernel (TL / TLD clock64 stamps, stack_fwd.cuh).

    python scripts/timeline_stack.py [layer] [gatys 0/1]
Coarse: per layer, when the epilogue saw GEMM1 complete and when the layer's epilogue 2 + publish finished.
Detailed: every event of ONE layer for a few CTAs.
"""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, '.')
from audio_style_transfer_b200 import _lib, synth  # noqa: E402
from audio_style_transfer_b200.step import StyleStep  # noqa: E402

layer = int(sys.argv[1]) if len(sys.argv) > 1 else 12
gatys = bool(int(sys.argv[2])) if len(sys.argv) > 2 else True
T, NL = 16384, 30
NT = T // 128
st = StyleStep(T, [NL - 1], list(range(NL)) if gatys else list(range(9, 29, 2)), gatys, precision='bf16', experiments=True)
st.set_weights(synth.make_weights(NL, 0, 0.0))
x = torch.randn(1, T, device='cuda') * 40
Ec, Gn = st.embeds(x)
st.set_targets(Ec * 0.9, Gn.roll(1, 1), 100.0)
buf = torch.zeros(4 * NT * 64, dtype=torch.int64, device='cuda')
lib = _lib.load(experiments=True)
for _ in range(3):
    st.step(x)
torch.cuda.synchronize()
lib.ast_debug_tl_layer(layer)
lib.ast_debug_timeline(C.c_void_p(buf.data_ptr()))
st.step(x)
torch.cuda.synchronize()
lib.ast_debug_timeline(C.c_void_p(0))
tl = buf.cpu().numpy().reshape(4, NT, 64)
co, de, bde, bco = tl[0], tl[1], tl[2], tl[3]
print('kernel duration per CTA (cycles): min %d median %d max %d' % tuple(
    f(co[:, 1] - co[:, 0]) for f in (np.min, np.median, np.max)))
for cta in (0, 64):
    t0 = co[cta, 0]
    acc1 = co[cta, 2:2 + NL] - t0
    fin = co[cta, 32:32 + NL] - t0
    print(f'--- CTA {cta}: layer: GEMM1-complete, layer-finished, (period)')
    prev = 0
    for l in range(NL):
        print(f'   l{l:2d} d={1 << (l % 10):3d}  acc1 {acc1[l]:8d}  fin {fin[l]:8d}  period {fin[l] - prev:6d}  (gemm1 wait {acc1[l] - prev:6d})')
        prev = fin[l]
names = {0: 'P: layer start', 2: 'P: far flags ok', 6: 'P: W2 issued', 8: 'M: ACT(0) ready', 9: 'M: ACT(1) ready',
         10: 'M: HALO(0) landed', 11: 'M: HALO(1) landed / far tap0 j0', 12: 'M: far tap0 j1', 13: 'M: far tap2 j0', 14: 'M: far tap2 j1',
         15: 'M: acc1 commit issued', 16: 'M: gemm2 j0 go', 17: 'M: gemm2 j1 go', 18: 'M: acc2 commit issued',
         20: 'H: flags(0) ok', 21: 'H: flags(1) ok', 30: 'E: acc1 complete', 32: 'E: epi1 done', 33: 'E: acc2 complete',
         36: 'E: published j0, ACT(0) arrive', 37: 'E: published j1, ACT(1) arrive', 41: 'E: layer done'}
for cta in (5, 64, 100):
    ev = de[cta]
    t0 = min(ev[k] for k in names if ev[k])
    print(f'--- CTA {cta}, layer {layer} (d={1 << (layer % 10)})')
    for k in sorted(names, key=lambda k: ev[k]):
        if ev[k]:
            print(f'  {names[k]:24s} {ev[k] - t0:8d}')

print('=== backward stack kernel')
for cta in (64,):
    done = bco[cta, 1:NL + 1]                 # stamp at boundary l = publish of dz_{l-1} done
    print(f'--- CTA {cta}: boundary period (cycles), boundaries {NL}..1')
    print('   ', [int(done[l - 1] - done[l]) for l in range(NL - 1, 0, -1)])
bnames = {0: 'P: boundary start', 3: 'P: far flags ok', 8: 'P: W2 issued', 10: 'M: DZ(0) ready', 11: 'M: DZ(1) ready',
          12: 'M: HALO(0) landed', 13: 'M: HALO(1) landed', 14: 'M: far tap0 j0', 15: 'M: far tap0 j1', 16: 'M: far tap2 j0',
          17: 'M: far tap2 j1', 18: 'M: acc1 commit issued', 19: 'M: gemm2 j0 go', 20: 'M: gemm2 j1 go', 21: 'M: acc2 commit issued',
          24: 'H: flags(0) ok', 25: 'H: flags(1) ok', 30: 'E: acc1 complete', 31: 'E: IO landed', 32: 'E: epi1 done',
          33: 'E: acc2 complete', 34: 'E: published j0, DZ(0) arrive', 35: 'E: published j1, DZ(1) arrive'}
for cta in (64, 100):
    ev = bde[cta]
    t0 = min(ev[k] for k in bnames if ev[k])
    print(f'--- CTA {cta}, boundary {layer} (d={1 << (layer % 10)})')
    for k in sorted(bnames, key=lambda k: ev[k]):
        if ev[k]:
            print(f'  {bnames[k]:28s} {ev[k] - t0:8d}')
