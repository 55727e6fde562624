"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Second, LITERAL restatement of nsynth/wavenet/masked.py.

oracle/step_np.py uses the closed form y[t] = b + W0^T a[t-d] + W1^T a[t] + W2^T a[t+d].  This file
instead walks the reference's actual route step by step -- time_to_batch reshape/transpose
(masked.py:57-86), 'SAME' / 'VALID' stride-1 cross-correlation in the batched domain
(masked.py:139,147-155), batch_to_time (masked.py:89-107) -- so that the integer indexing of the
closed form (tap offsets, zero padding at both clip ends) is pinned by an independent derivation.
tests/test_oracle.py requires the two to agree BIT-EXACTLY on integer-valued data.

TF's conv2d SAME rule for stride 1 (not in /root/reference; TF 1.x semantics): total padding
K-1, pad_left = (K-1)//2, pad_right = K-1-pad_left; tf.nn.conv2d does not flip the taps.
"""
import numpy as np


def time_to_batch(x, block_size):
    """masked.py:57-86.  x [nb, k*block_size, n] -> [nb*block_size, k, n]."""
    nb, t, n = x.shape
    y = np.reshape(x, [nb, t // block_size, block_size, n])
    y = np.transpose(y, [0, 2, 1, 3])
    return np.reshape(y, [nb * block_size, t // block_size, n])


def batch_to_time(x, block_size):
    """masked.py:89-107.  x [nb*block_size, k, n] -> [nb, k*block_size, n]."""
    nbb, k, n = x.shape
    y = np.reshape(x, [nbb // block_size, block_size, k, n])
    y = np.transpose(y, [0, 2, 1, 3])
    return np.reshape(y, [nbb // block_size, k * block_size, n])


def _conv2d_1xK(x, W, padding):
    """tf.nn.conv2d on [nb, 1, width, Cin] with W [1, K, Cin, Cout], strides 1 (masked.py:151-154)."""
    nb, width, cin = x.shape
    K = W.shape[0]
    if padding == 'SAME':
        pl = (K - 1) // 2
        pr = K - 1 - pl
        x = np.pad(x, [(0, 0), (pl, pr), (0, 0)])
    out_w = x.shape[1] - K + 1
    y = np.zeros((nb, out_w, W.shape[2]), x.dtype)
    for k in range(K):
        y += x[:, k:k + out_w, :] @ W[k]
    return y


def conv1d(x, W, b, dilation=1, causal=True):
    """masked.py:110-160.  x [mb, T, Cin]; W [K, Cin, Cout]; returns [mb, T, Cout]."""
    mb, length, cin = x.shape
    assert length % dilation == 0                                  # masked.py:134
    K = W.shape[0]
    padding = 'VALID' if causal else 'SAME'                        # masked.py:139
    x_ttb = time_to_batch(x, dilation)                             # :147
    if K > 1 and causal:                                           # :148-149
        x_ttb = np.pad(x_ttb, [(0, 0), (K - 1, 0), (0, 0)])
    y = _conv2d_1xK(x_ttb, W, padding) + b                         # :151-155
    y = batch_to_time(y, dilation)                                 # :159
    assert y.shape == (mb, length, W.shape[2])                     # :160
    return y


def encoder_forward(x, w, n_layers):
    """model.py:80-116 through the literal conv1d.  x [T] -> list of extracts H[0..n_layers]."""
    dt = x.dtype
    xs = (x / dt.type(128.0))[None, :, None]
    enc = conv1d(xs, w.W0.astype(dt), w.b0.astype(dt), 1, causal=False)
    H = [enc[0]]
    for l in range(n_layers):
        d = 2 ** (l % 10)
        d_enc = np.maximum(enc, 0)
        d_enc = conv1d(d_enc, w.Wd[l].astype(dt), w.bd[l].astype(dt), d, causal=False)
        d_enc = np.maximum(d_enc, 0)
        d_enc = conv1d(d_enc, w.Wr[l].astype(dt)[None], w.br[l].astype(dt), 1, causal=True)
        enc = enc + d_enc
        H.append(enc[0])
    return H
