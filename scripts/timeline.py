"""Per-CTA phase timeline of the forward layer kernel (clock64 stamps written by the TL() hooks)."""
import ctypes as C
import sys
import numpy as np, torch
sys.path.insert(0, '.')
from audio_style_transfer_b200.step import StyleStep
from audio_style_transfer_b200 import synth, _lib

T, NL = 16384, 30
st = StyleStep(T, [NL - 1], list(range(NL)), True, precision='bf16')
st.set_weights(synth.make_weights(NL, 0, 0.0))
x = torch.randn(1, T, device='cuda') * 40
buf = torch.zeros(128 * 64, dtype=torch.int64, device='cuda')
lib = _lib.load()
for _ in range(3):
    st.embeds(x, style=False)
torch.cuda.synchronize()
lib.ast_debug_timeline(C.c_void_p(buf.data_ptr()))
if '--skipw' in sys.argv:          # timing experiment: forward producer skips the conv weight tiles (results garbage)
    lib.ast_debug_skipw(1)
st.embeds(x, style=False)          # the LAST layer's stamps survive
torch.cuda.synchronize()
lib.ast_debug_timeline(C.c_void_p(0))
lib.ast_debug_skipw(0)
tl = buf.cpu().numpy().reshape(128, 64)
names = {0: 'entry', 1: 'setup done', **{2 + i: f'tma issued it{i}' for i in range(6)}, 8: 'tma issued W2/H 0', 9: 'tma issued W2/H 1',
         **{10 + i: f'mma full it{i}' for i in range(6)}, 16: 'gemm2 half0 start', 19: 'gemm2 half1 start',
         18: 'acc1 commit', 20: 'acc2 commit', 24: 'epi: acc1 full', 25: 'epi1 done', 26: 'epi: acc2 full',
         27: 'epi: H landed', 40: 'epi2 tmem loaded', 41: 'epi2 H updated', 42: 'epi2 F staged+mask', 43: 'epi2 hi/lo staged', 44: 'epi2 hmask stored', 28: 'epi2 math done', 29: 'stores issued', 30: 'stores complete', 31: 'exit sync'}
for cta in (0, 5, 64, 127):
    t0 = tl[cta, 0]
    print(f'--- CTA {cta}')
    for k in sorted(names, key=lambda k: tl[cta, k]):
        if tl[cta, k]:
            print(f'  {names[k]:20s} {tl[cta, k] - t0:8d} cyc')
rel = tl[:, 31] - tl[:, 0]
print('CTA duration cycles: min %d median %d max %d' % (rel.min(), np.median(rel), rel.max()))
print('entry spread (cycles, unreliable across SMs):', tl[:, 0].max() - tl[:, 0].min())
