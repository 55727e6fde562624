"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Named parity cases: seeded inputs + oracle answers.

Shared by tests/ and oracle/make_golden.py so that the committed golden vectors and the GPU parity
tests are built from the same definitions.  Inputs follow SURVEY.md 8(d).
"""
from __future__ import annotations

import numpy as np

from audio_style_transfer_b200 import synth
from audio_style_transfer_b200.mu_law import mu_law_numpy

from . import step_np

# name -> (T, cont_ids, stack, style_lyr_ids, gatys, nb_channels, cnt_channels, style_chunks)
CASES = {
    # small: every code path, seconds on CPU
    "t1024_ours_s0":    (1024, [9], 0, None, False, 128, 128, 2),
    "t1024_ours_sub":   (1024, [3, 7], 0, None, False, 64, 32, 2),     # --channels / --cnt_channels subsets
    "t1024_ours_ids":   (1024, [12], None, [2, 5, 11, 12], False, 128, 128, 1),   # explicit --style_lyrs
    "t2048_gatys_s1":   (2048, [19], 1, None, True, 128, 128, 1),
    "t4096_cfg0":       (4096, [25], 0, None, False, 128, 128, 5),     # BASELINE configs[0] layer set
    "t4096_cfg1":       (4096, [29], None, None, True, 128, 128, 2),   # BASELINE configs[1] layer set
    "t4096_alias30":    (4096, [30], 2, None, False, 128, 128, 1),     # extracts[30] aliases [29]
    "t1024_ours_l17":   (1024, [19], None, list(range(3, 20)), False, 96, 128, 1),   # 17 style layers (not a multiple of 8), 96 channels
    "t2048_ours_all30": (2048, [29], None, None, False, 128, 128, 1),  # the CLI default head: --stack unset = all 30 layers, channel-wise (cfg4's head)
    # BASELINE.json sizes
    "t16384_cfg0":      (16384, [25], 0, None, False, 128, 128, 1),
    "t16384_cfg1":      (16384, [29], None, None, True, 128, 128, 1),
    "t16384_cli":       (16384, [25], None, None, False, 128, 128, 1),  # methods.py CLI defaults at full size: all 30 layers channel-wise (bench --workload cli, cfg4's head)
    # BASELINE configs[2]: 4 s clip (time-sharded on the GPU), cfg0's layer set
    "t65536_cfg2":      (65536, [25], 0, None, False, 128, 128, 1),
}
# "<case>_pN": the same case with the step input of pair N of a batch (BASELINE configs[3]): x = mu_law(content) +
# N(0,1) drawn with seed 3 + N; targets unchanged.  Pair 0 is the plain case.
PAIR_SEED0 = 3
LAMBD = 100.0   # methods.py:256 default


def pair_x(case, n):
    """Step input of pair n of a batch built from `case` (n = 0: the case's own x)."""
    return synth.step_input(case["c_mu"], seed=PAIR_SEED0 + n).astype(np.float64)


def build_inputs(name, bias_scale=0.1):
    if "_p" in name and name.rsplit("_p", 1)[1].isdigit():
        base, n = name.rsplit("_p", 1)
        case = build_inputs(base, bias_scale)
        return dict(case, name=name, x=pair_x(case, int(n)))
    T, cont, stack, ids, gatys, nb, cnt, chunks = CASES[name]
    style = synth.resolve_style_ids(stack, ids)
    NL = synth.layers_needed(cont, style)
    w = synth.make_weights(NL, seed=0, bias_scale=bias_scale)
    content = synth.content_audio(T)
    style_a = synth.style_audio(T, chunks=chunks)
    c_mu = mu_law_numpy(content).astype(np.float64)
    s_mu = mu_law_numpy(style_a).astype(np.float64)
    x = synth.step_input(c_mu).astype(np.float64)
    return dict(name=name, T=T, cont=cont, style=style, gatys=gatys, nb=nb, cnt=cnt, NL=NL, w=w,
                c_mu=c_mu, s_chunks=[s_mu[i * T:(i + 1) * T] for i in range(chunks)], x=x, lambd=LAMBD)


def oracle_targets(case):
    return step_np.targets(case["c_mu"], case["s_chunks"], [case["c_mu"]], case["w"], case["cont"], case["style"],
                           case["gatys"], case["nb"], case["cnt"])


def oracle_step(case, Pc, Ps, return_aux=False, masks=None, fwd=None):
    return step_np.step(case["x"], case["w"], Pc, Ps, case["cont"], case["style"], case["gatys"], case["lambd"], 0.0,
                        case["nb"], case["cnt"], return_aux=return_aux, masks=masks, fwd=fwd)


def summary(case, Pc, Ps, losses, dx, aux):
    """What gets committed under tests/golden/: small, but enough to catch any drift."""
    sub = 1 if case["T"] <= 4096 else 16
    return dict(
        losses=np.array(losses[:3], np.float64),
        dx=dx[::sub].astype(np.float64), dx_stride=np.int64(sub), dx_norm=np.float64(np.linalg.norm(dx)),
        Gn_sum=np.float64(aux["Gn"].sum()), Gn_sq=np.float64((aux["Gn"] ** 2).sum()),
        Ps_sum=np.float64(Ps.sum()), Pc_sq=np.float64((Pc ** 2).sum()),
        H_sq=np.array([(h ** 2).sum() for h in aux["H"]], np.float64),
        x_sum=np.float64(case["x"].sum()),
        w_digest=np.array(case["w"].digest()),
    )
