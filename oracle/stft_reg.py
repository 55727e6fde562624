"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.

CPU restatement of the reference's STFT regulariser (the gamma term of define_loss,
methods.py:121-123, with utils.py:92-104) and of the gradient TensorFlow's autodiff gives for it.

    a   = utils.inv_mu_law(quantized_input[0])                                 methods.py:121
    Y   = tf.contrib.signal.stft(a, frame_length=1024, frame_step=512)         methods.py:122
    reg = reduce_mean(utils.abs(real(Y)) + utils.abs(imag(Y)))                 methods.py:123

tf.contrib.signal.stft defaults (TF 1.x, third-party, not vendored in /root/reference): fft_length = 1024 (the
enclosing power of two of frame_length), window_fn = hann_window(periodic=True), pad_end=False, so
frames = 1 + (T - 1024) // 512 rows of 513 rfft bins.

PARITY UNPINNED like the rest of oracle/ (see step_np.py): no TensorFlow here.  Pinned by (a) a second, literal
restatement through torch ops + torch.autograd in fp64 (`torch_literal`), (b) central finite differences
(tests/test_oracle.py).

The reference's abs / sign are not the textbook ones (utils.py:92-97):
    abs(v)  = maximum(v, 1e-12) + maximum(0, -v)
    sign(v) = where(|v| <= 1e-12, 0, v) / abs(v)
and TF's Maximum gradient sends the whole gradient to the first argument where first >= second, so
    abs'(v) = [v >= 1e-12] - [v < 0].
"""
from __future__ import annotations

import numpy as np

FRAME = 1024          # methods.py:122 frame_length (= fft_length)
HOP = 512             # methods.py:122 frame_step
BINS = FRAME // 2 + 1
EPS = 1e-12           # utils.py:93,96
MU = 255.0            # utils.py:99


def n_frames(T: int) -> int:
    """tf.contrib.signal.frame with pad_end=False."""
    return 0 if T < FRAME else 1 + (T - FRAME) // HOP


def ref_abs(v):
    """utils.py:92-93."""
    return np.maximum(v, EPS) + np.maximum(0.0, -v)


def ref_abs_grad(v):
    """TF MaximumGrad: gradient to the first operand where first >= second."""
    return (v >= EPS).astype(v.dtype) - (v < 0).astype(v.dtype)


def inv_mu_law(x):
    """utils.py:99-104 on float x (mu-law units).  Returns (a, da/dx)."""
    v = (x + 0.5) * 2.0 / (MU + 1.0)                                 # :101
    A = ref_abs(v)
    dA = ref_abs_grad(v)
    wv = np.where(np.abs(v) <= EPS, 0.0, v)                          # :96
    dw = (np.abs(v) > EPS).astype(x.dtype)
    s = wv / A                                                       # :97
    ds = (dw - s * dA) / A
    P = (1.0 + MU) ** A
    a = s / MU * (P - 1.0)                                           # :102
    da_dv = ds / MU * (P - 1.0) + s / MU * P * np.log(1.0 + MU) * dA
    da_dx = da_dv * (2.0 / (MU + 1.0))
    zero = (x == 0)                                                  # :103 where(equal(x, 0), x, out)
    return np.where(zero, x, a), np.where(zero, 1.0, da_dx)


def hann_periodic(n: int = FRAME, dtype=np.float64):
    """window_ops.hann_window(periodic=True)."""
    return (0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)).astype(dtype)


def stft_l1(x, frames=None, total_frames=None):
    """reg and d reg / d x for one clip x [T] (mu-law units).

    frames: optional iterable of frame indices to include (time sharding: the frames a rank owns);
    total_frames: the mean's frame count (defaults to n_frames(T)).
    Returns (reg, dreg_dx [T], raw_sum) with reg = raw_sum / (total_frames * 513)."""
    x = np.asarray(x)
    dt = x.dtype if x.dtype in (np.float32, np.float64) else np.float64
    x = x.astype(dt)
    T = x.shape[0]
    F = n_frames(T)
    if total_frames is None:
        total_frames = F
    if frames is None:
        frames = range(F)
    a, dadx = inv_mu_law(x)
    w = hann_periodic(FRAME, dt)
    raw = 0.0
    da = np.zeros(T, dt)
    for f in frames:
        seg = a[f * HOP:f * HOP + FRAME] * w
        Y = np.fft.rfft(seg.astype(np.float64)).astype(np.complex128 if dt == np.float64 else np.complex64)
        re, im = Y.real.astype(dt), Y.imag.astype(dt)
        raw += float(np.sum(ref_abs(re) + ref_abs(im), dtype=np.float64))
        gre, gim = ref_abs_grad(re), ref_abs_grad(im)
        # adjoint of the real -> half-complex DFT (513 independent bins, what TF's RFFT gradient computes):
        # d/d seg[m] = sum_k gre[k] cos(2 pi k m / N) - gim[k] sin(2 pi k m / N)
        #            = Re( sum_k (gre[k] + i gim[k]) e^{+2 pi i k m / N} ) over k = 0..512 only
        z = np.zeros(FRAME, np.complex128)
        z[:BINS] = gre.astype(np.float64) + 1j * gim.astype(np.float64)
        back = (np.fft.ifft(z) * FRAME).real
        da[f * HOP:f * HOP + FRAME] += (back * w).astype(dt)
    norm = float(total_frames * BINS)
    return raw / norm, da * dadx / dt.type(norm), raw


def torch_literal(x):
    """The same graph written op by op with torch (fp64) and differentiated by torch.autograd.
    torch.maximum splits the gradient at exact ties where TF does not; ties only occur at values whose
    gradient contribution is zero anyway (Im of bins 0 and 512)."""
    import torch
    xt = torch.tensor(np.asarray(x, np.float64), requires_grad=True)

    def tabs(v):
        return torch.maximum(v, torch.full_like(v, EPS)) + torch.maximum(torch.zeros_like(v), -v)

    v = (xt + 0.5) * 2.0 / (MU + 1.0)
    s = torch.where(v.abs() <= EPS, torch.zeros_like(v), v) / tabs(v)
    a = s / MU * ((1.0 + MU) ** tabs(v) - 1.0)
    a = torch.where(xt == 0, xt, a)
    win = torch.tensor(hann_periodic())
    fr = a.unfold(0, FRAME, HOP) * win
    Y = torch.fft.rfft(fr, dim=1)
    reg = (tabs(Y.real) + tabs(Y.imag)).mean()
    (g,) = torch.autograd.grad(reg, xt)
    return float(reg.detach()), g.numpy()
