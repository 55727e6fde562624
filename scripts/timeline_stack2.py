"""Timeline of the two-pairs-per-CTA forward kernel (layer given on the command line)."""
import ctypes as C
import os
import sys
os.environ['AST_B200_DUO'] = '1'
import numpy as np
import torch
sys.path.insert(0, '.')
from audio_style_transfer_b200 import _lib, synth  # noqa: E402
from audio_style_transfer_b200.step import StyleStep  # noqa: E402

layer = int(sys.argv[1]) if len(sys.argv) > 1 else 12
T, NL, B = 16384, 30, 2
NT = T // 128
st = StyleStep(T, [NL - 1], list(range(NL)), True, precision='bf16', B=B)
st.set_weights(synth.make_weights(NL, 0, 0.0))
x = torch.randn(B, T, device='cuda') * 40
Ec, Gn = st.embeds(x)
st.set_targets(Ec * 0.9, Gn.roll(1, 1), 100.0)
buf = torch.zeros(4 * 2 * NT * 64, dtype=torch.int64, device='cuda')
lib = _lib.load()
for _ in range(3):
    st.step(x)
torch.cuda.synchronize()
lib.ast_debug_tl_layer(layer)
lib.ast_debug_timeline(C.c_void_p(buf.data_ptr()))
st.step(x)
torch.cuda.synchronize()
lib.ast_debug_timeline(C.c_void_p(0))
tl = buf.cpu().numpy()
de = tl[NT * 64: 2 * NT * 64].reshape(NT, 64)          # detailed block of the duo launch (gridDim = NT)
names = {}
for pa in range(2):
    for u in range(8):
        names[pa * 10 + u] = f'M: pair{pa} unit{u} go'
    names[30 + pa * 10] = f'E: pair{pa} acc1 complete'
    names[31 + pa * 10] = f'E: pair{pa} epi1 done'
    names[32 + pa * 10] = f'E: pair{pa} acc2 complete'
    names[33 + pa * 10] = f'E: pair{pa} h staged, DONE'
    names[34 + pa * 10] = f'E: pair{pa} publish + pass 2 done'
    names[50 + pa * 5] = f'P: pair{pa} item start'
    names[51 + pa * 5] = f'P: pair{pa} unit0 W slot armed'
    names[52 + pa * 5] = f'P: pair{pa} flags ok'
    names[53 + pa * 5] = f'P: pair{pa} unit0 A slot armed'
    names[54 + pa * 5] = f'P: pair{pa} all fills issued'
for cta in (64,):
    ev = de[cta]
    t0 = min(ev[k] for k in names if ev[k])
    print(f'--- CTA {cta}, layer {layer}')
    for k in sorted(names, key=lambda k: ev[k]):
        if ev[k]:
            print(f'  {names[k]:34s} {ev[k] - t0:8d}')
