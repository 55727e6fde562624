"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.

CPU (numpy, fp64 by default) restatement of the reference's fwd + bwd + Gram-loss step.  Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import it.

PARITY UNPINNED: the reference ships no tests, fixtures or golden vectors for this path
 and its arithmetic lives in TensorFlow 1.x, which cannot be installed or run
here (Python 3.12, no network).  This restatement is therefore pinned only by (a) a second, literal
restatement of masked.py's reshape/transpose/SAME-conv route (oracle/masked_literal.py, bit-exact on
integer data), (b) torch.autograd in fp64 (oracle/torch_step.py) and (c) central finite differences
(tests/test_oracle.py).  It has NOT been compared with output of the reference itself.

Every function cites the reference lines it follows (paths relative to /root/reference).
"""
from __future__ import annotations

import numpy as np

C = 128


def dilation(l: int) -> int:
    """model.py:98 -- 2 ** (num_layer % ae_num_stages), ae_num_stages = 10 (model.py:70)."""
    return 1 << (l % 10)


def shift_rows(a: np.ndarray, s: int) -> np.ndarray:
    """out[t] = a[t + s], rows outside [0, T) are zero.

    This is what 'SAME' padding of the K=3 conv amounts to after time_to_batch(d):
    one zero row either side in the [d, T/d] domain = d zero rows either side in time
    (masked.py:139,147-158; SURVEY.md 3.2 closed form)."""
    T = a.shape[0]
    out = np.zeros_like(a)
    if s == 0:
        out[:] = a
    elif s > 0:
        if s < T:
            out[:T - s] = a[s:]
    else:
        if -s < T:
            out[-s:] = a[:T + s]
    return out


def conv3(a: np.ndarray, W: np.ndarray, b: np.ndarray, d: int) -> np.ndarray:
    """masked.conv1d, filter_length=3, causal=False (masked.py:110-160 via model.py:88-93,100-106).

    a [T, Cin], W [3, Cin, Cout] (= TF's [1,3,Cin,Cout] without the leading 1), cross-correlation:
    y[t] = b + W[0]^T a[t-d] + W[1]^T a[t] + W[2]^T a[t+d]."""
    return shift_rows(a, -d) @ W[0] + a @ W[1] + shift_rows(a, d) @ W[2] + b


def conv3_T(dy: np.ndarray, W: np.ndarray, d: int) -> np.ndarray:
    """Adjoint of conv3 w.r.t. its input (TF's Conv2DBackpropInput for this geometry):
    da[t] = W[0] dy[t+d] + W[1] dy[t] + W[2] dy[t-d]   (SURVEY.md 8(a))."""
    return shift_rows(dy, d) @ W[0].T + dy @ W[1].T + shift_rows(dy, -d) @ W[2].T


def encoder_forward(x, w, n_layers):
    """model.py:80-116.  x [T] in mu-law units.  Returns (H, Z) with H[l] = residual stream entering
    layer l (H[0] = start conv output), H[l+1] = cfg.extracts[l] (model.py:116), Z[l] = pre-ReLU
    output of the dilated conv of layer l."""
    dt = x.dtype
    xs = (x / dt.type(128.0))[:, None]                        # model.py:82-83
    h = conv3(xs, w.W0.astype(dt), w.b0.astype(dt), 1)        # model.py:88-93 (ae_startconv)
    H, Z = [h], []
    for l in range(n_layers):
        a = np.maximum(h, 0)                                  # model.py:99
        z = conv3(a, w.Wd[l].astype(dt), w.bd[l].astype(dt), dilation(l))   # :100-106
        r = np.maximum(z, 0)                                  # :107
        h = h + (r @ w.Wr[l].astype(dt) + w.br[l].astype(dt))  # :109-114 (K=1 conv + residual)
        H.append(h)
        Z.append(z)
    return H, Z


def extract(H, i):
    """cfg.extracts[i] (model.py:116-119): i in 0..29 -> stream after layer i; 30 aliases 29."""
    return H[min(i, 29) + 1]


def content_embeds(H, cont_ids, cnt_channels=128):
    """methods.py:58 -- concat of extracts[i][:, :cnt_channels] along channels -> [T, n_c*cnt_ch]."""
    return np.concatenate([extract(H, i)[:, :cnt_channels] for i in cont_ids], axis=1)


def raw_gram(H, style_ids, gatys):
    """methods.py:62-73.  ours: [C, L, L] (per channel, across layers); gatys: [L, C, C]."""
    S = np.stack([extract(H, i) for i in style_ids], axis=0)      # [L, T, C]   (:62-66)
    if not gatys:
        S = S.transpose(2, 0, 1)                                  # [C, L, T]   (:69)
    else:
        S = S.transpose(0, 2, 1)                                  # [L, C, T]   (:71)
    return S @ S.transpose(0, 2, 1)                               # (:73)


def l2_normalize(G, eps=1e-12):
    """tf.nn.l2_normalize(G, axis=(1,2)) (methods.py:74): x * rsqrt(max(sum(x^2), eps))."""
    ss = np.sum(G * G, axis=(1, 2), keepdims=True)
    return G / np.sqrt(np.maximum(ss, eps))


def style_embeds(H, style_ids, gatys, nb_channels=128):
    """methods.py:68-76."""
    Gn = l2_normalize(raw_gram(H, style_ids, gatys))
    if nb_channels < 128 and not gatys:                           # :75-76
        Gn = Gn[:nb_channels]
    return Gn


def embeds(x, w, cont_ids, style_ids, gatys, nb_channels=128, cnt_channels=128, n_layers=None):
    """get_embeds for both heads (methods.py:86-95): forward only."""
    if n_layers is None:
        n_layers = min(max(list(cont_ids) + list(style_ids)), 29) + 1
    H, _ = encoder_forward(x, w, n_layers)
    return content_embeds(H, cont_ids, cnt_channels), style_embeds(H, style_ids, gatys, nb_channels)


def step(x, w, Pc, Ps, cont_ids, style_ids, gatys, lambd, gamma=0.0,
         nb_channels=128, cnt_channels=128, n_layers=None, return_aux=False, masks=None, fwd=None):
    """One (loss, dloss/dx) evaluation: define_loss (methods.py:113-125) + the gradient that
    ScipyOptimizerInterface asks TF for (methods.py:133-137), var_list = [x] only.

    x [T] (mu-law units); Pc [T, n_c*cnt_ch]; Ps [nb, L, L] (ours) or [L, C, C] (gatys).
    Returns (losses, dx) with losses = (total, content, style, regulariser).
    gamma != 0 adds the STFT regulariser (methods.py:121-123, oracle/stft_reg.py); with gamma == 0 the
    regulariser slot reports 0 (the reference evaluates it for logging only, methods.py:167).
    masks = (hmask, zmask), bool [n_layers, T, C] each, replaces the backward's own ReluGrad masks (H[l] > 0, Z[l] > 0)
    by given ones (tests impose the device's masks so that sign bits flipped by rounding do not dominate the comparison).
    fwd = (H, Z) of an earlier call on the same x (aux["H"], aux["Z"]) skips the forward.
    Backward formulas: SURVEY.md 8(a)."""
    dt = x.dtype
    cont_ids, style_ids = list(cont_ids), list(style_ids)
    if n_layers is None:
        n_layers = min(max(cont_ids + style_ids), 29) + 1
    H, Z = encoder_forward(x, w, n_layers) if fwd is None else fwd
    T = x.shape[0]

    # ---- heads and losses ------------------------------------------------------------------
    Ec = content_embeds(H, cont_ids, cnt_channels)
    dEc = Ec - Pc
    Lc = 10.0 * np.mean(dEc * dEc)                                   # methods.py:116-117
    G = raw_gram(H, style_ids, gatys)
    ss = np.sum(G * G, axis=(1, 2), keepdims=True)
    inv = 1.0 / np.sqrt(np.maximum(ss, 1e-12))
    Gn = G * inv
    nb = Gn.shape[0] if (gatys or nb_channels >= 128) else nb_channels
    dGs = Gn[:nb] - Ps
    Ls = 1e3 * np.mean(dGs * dGs)                                    # methods.py:118-119
    loss = Lc + lambd * Ls                                           # methods.py:125

    # ---- backward through the heads ---------------------------------------------------------
    dGn = np.zeros_like(Gn)
    dGn[:nb] = lambd * 1e3 * 2.0 * dGs / dGs.size
    # l2_normalize backward: y = G * rsqrt(max(ss, eps)).  In the clamped regime rsqrt is constant.
    dot = np.sum(dGn * Gn, axis=(1, 2), keepdims=True)
    dG = np.where(ss >= 1e-12, (dGn - Gn * dot) * inv, dGn * inv)
    Ssym = dG + dG.transpose(0, 2, 1)

    dF = {}   # extract index (0..29) -> gradient w.r.t. H[idx+1], [T, C]

    def acc(i, g):
        i = min(i, 29)
        dF[i] = dF[i] + g if i in dF else g

    for slot, i in enumerate(cont_ids):                              # content head
        g = np.zeros((T, C), dt)
        g[:, :cnt_channels] = (10.0 * 2.0 / dEc.size) * dEc[:, slot * cnt_channels:(slot + 1) * cnt_channels]
        acc(i, g)
    if gatys:                                                        # dF_l = F_l (dG_l + dG_l^T)
        for li, i in enumerate(style_ids):
            acc(i, extract(H, i) @ Ssym[li])
    else:                                                            # dF_l[t,c] = sum_m Ssym[c,l,m] F_m[t,c]
        F = np.stack([extract(H, i) for i in style_ids], axis=0)     # [L, T, C]
        dFs = np.einsum('clm,mtc->ltc', Ssym, F)
        for li, i in enumerate(style_ids):
            acc(i, dFs[li])

    # ---- backward through the encoder -------------------------------------------------------
    g = np.zeros((T, C), dt)
    for l in range(n_layers - 1, -1, -1):
        if l in dF:
            g = g + dF[l]
        dr = g @ w.Wr[l].astype(dt).T
        dz = dr * ((Z[l] > 0) if masks is None else masks[1][l])    # ReluGrad: mask = output > 0
        da = conv3_T(dz, w.Wd[l].astype(dt), dilation(l))
        g = g + da * ((H[l] > 0) if masks is None else masks[0][l])
    dxs = conv3_T(g, w.W0.astype(dt), 1)[:, 0]
    dx = dxs / dt.type(128.0)
    reg = 0.0
    if gamma != 0.0:                                                 # methods.py:121-125
        from . import stft_reg
        reg, dreg, _ = stft_reg.stft_l1(x)
        loss = loss + gamma * reg
        dx = dx + dt.type(gamma) * dreg.astype(dt)
    losses = (loss, Lc, Ls, reg)
    if return_aux:
        return losses, dx, dict(H=H, Z=Z, G=G, Gn=Gn, Ec=Ec)
    return losses, dx


def targets(content_mu, style_mu_chunks, content_mu_chunks, w, cont_ids, style_ids, gatys,
            nb_channels=128, cnt_channels=128):
    """GatysNet.run target construction (methods.py:192-212):
    Pc = Ec(content); phi = Gn(content) + mean_k Gn(target chunk k) - mean_k Gn(source chunk k),
    then l2_normalize over axes (1,2).  In the CLI source == content file (methods.py:240)."""
    Pc, phi = embeds(content_mu, w, cont_ids, style_ids, gatys, nb_channels, cnt_channels)
    phi_t = np.mean([embeds(c, w, cont_ids, style_ids, gatys, nb_channels, cnt_channels)[1]
                     for c in style_mu_chunks], axis=0)              # get_style_phi, :97-111
    phi_s = np.mean([embeds(c, w, cont_ids, style_ids, gatys, nb_channels, cnt_channels)[1]
                     for c in content_mu_chunks], axis=0)
    Ps = l2_normalize(phi + phi_t - phi_s)                           # :211-212
    return Pc, Ps
