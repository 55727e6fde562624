"""Seeded synthetic inputs for the style-transfer step (no network: no NSynth checkpoint, no wavs).

Neutral test/bench data only -- no arithmetic of the hot path lives here.  Conventions follow
SURVEY.md section 8(d): seeds 0 (weights), 1 (content), 2 (style), 3 (x noise).

Weight layout mirrors the reference's TF variables (nsynth/wavenet/masked.py:136,141-145):
``W`` is ``[1, K, Cin, Cout]`` there; here the leading 1 is dropped:

    W0 [3, 1, 128]   b0 [128]              ae_startconv            (model.py:88-93)
    Wd [NL, 3, 128, 128]  bd [NL, 128]     ae_dilatedconv_{l+1}    (model.py:100-106)
    Wr [NL, 128, 128]     br [NL, 128]     ae_res_{l+1}  (K=1)     (model.py:109-113)
"""
from __future__ import annotations

import hashlib
from dataclasses import dataclass

import numpy as np

C = 128          # ae_width, model.py:73
K = 3            # ae_filter_length, model.py:72
NUM_LAYERS = 30  # ae_num_layers, model.py:71
NUM_STAGES = 10  # ae_num_stages, model.py:70


def dilation(layer: int) -> int:
    """model.py:98 -- dilation = 2 ** (num_layer % ae_num_stages)."""
    return 1 << (layer % NUM_STAGES)


@dataclass
class EncoderWeights:
    W0: np.ndarray  # [3, 1, C]  float32
    b0: np.ndarray  # [C]
    Wd: np.ndarray  # [NL, 3, C, C]   (k, cin, cout)
    bd: np.ndarray  # [NL, C]
    Wr: np.ndarray  # [NL, C, C]      (cin, cout)
    br: np.ndarray  # [NL, C]

    @property
    def n_layers(self) -> int:
        return int(self.Wd.shape[0])

    def astype(self, dt) -> "EncoderWeights":
        return EncoderWeights(*(np.ascontiguousarray(a, dtype=dt) for a in
                                (self.W0, self.b0, self.Wd, self.bd, self.Wr, self.br)))

    def digest(self) -> str:
        h = hashlib.sha256()
        for a in (self.W0, self.b0, self.Wd, self.bd, self.Wr, self.br):
            h.update(np.ascontiguousarray(a, dtype=np.float32).tobytes())
        return h.hexdigest()[:16]


def _unit_scaling(rng, shape, fan_in):
    # tf.uniform_unit_scaling_initializer(1.0) (masked.py:116): U(+-sqrt(3/fan_in)), fan_in = K*Cin.
    lim = np.sqrt(3.0 / fan_in)
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def make_weights(n_layers: int = NUM_LAYERS, seed: int = 0, bias_scale: float = 0.0,
                 gain: float = 1.0) -> EncoderWeights:
    """Random-init encoder.  ``bias_scale=0`` is the reference init (masked.py:117);
    parity tests use ``bias_scale=0.1`` so that the bias paths are exercised.
    ``gain`` scales the residual-branch 1x1 weights (1.0 = reference init)."""
    rng = np.random.default_rng(seed)
    W0 = _unit_scaling(rng, (K, 1, C), K * 1)
    Wd = _unit_scaling(rng, (n_layers, K, C, C), K * C)
    Wr = _unit_scaling(rng, (n_layers, C, C), C) * np.float32(gain)
    if bias_scale:
        b0 = rng.uniform(-bias_scale, bias_scale, C).astype(np.float32)
        bd = rng.uniform(-bias_scale, bias_scale, (n_layers, C)).astype(np.float32)
        br = rng.uniform(-bias_scale, bias_scale, (n_layers, C)).astype(np.float32)
    else:
        b0 = np.zeros(C, np.float32)
        bd = np.zeros((n_layers, C), np.float32)
        br = np.zeros((n_layers, C), np.float32)
    return EncoderWeights(W0, b0, Wd, bd, Wr, br)


def make_int_weights(n_layers: int, seed: int = 0) -> EncoderWeights:
    """Small-integer weights for the bit-exact index probes: every product and
    partial sum is an exactly representable integer in fp32/bf16, so any tap-offset, padding or
    tile-seam mistake changes the result while rounding cannot."""
    rng = np.random.default_rng(seed)
    W0 = rng.integers(-1, 2, (K, 1, C)).astype(np.float32) * 128.0  # cancels the /128 input scale
    Wd = np.zeros((n_layers, K, C, C), np.float32)
    Wr = np.zeros((n_layers, C, C), np.float32)
    for l in range(n_layers):
        # sparse +-1 taps: each output channel reads 2 input channels per tap
        for k in range(K):
            for co in range(C):
                ci = rng.integers(0, C, 2)
                Wd[l, k, ci, co] = rng.choice([-1.0, 1.0], 2)
        for co in range(C):
            ci = rng.integers(0, C, 1)
            Wr[l, ci, co] = rng.choice([-1.0, 1.0], 1)
    b0 = rng.integers(-1, 2, C).astype(np.float32)
    bd = rng.integers(-1, 2, (n_layers, C)).astype(np.float32)
    br = rng.integers(-1, 2, (n_layers, C)).astype(np.float32)
    return EncoderWeights(W0, b0, Wd, bd, Wr, br)


def make_audio(n: int, seed: int, f0: float, sr: int = 16000) -> np.ndarray:
    """two partials + noise, clipped to [-1, 1), float32."""
    rng = np.random.default_rng(seed)
    t = np.arange(n, dtype=np.float64)
    phi = rng.uniform(0, 2 * np.pi)
    a = (0.4 * np.sin(2 * np.pi * f0 * t / sr + phi)
         + 0.2 * np.sin(2 * np.pi * 3.01 * f0 * t / sr)
         + 0.05 * rng.standard_normal(n))
    return np.clip(a, -1.0, np.nextafter(1.0, 0.0)).astype(np.float32)


def content_audio(T: int, seed: int = 1) -> np.ndarray:
    return make_audio(T, seed, 220.0)


def style_audio(T: int, seed: int = 2, chunks: int = 5) -> np.ndarray:
    return make_audio(chunks * T, seed, 311.0)


def step_input(content_mu: np.ndarray, seed: int = 3) -> np.ndarray:
    """x = mu_law(content) + N(0,1): generic ReLU masks, non-trivial gradient."""
    rng = np.random.default_rng(seed)
    return (content_mu.astype(np.float64) + rng.standard_normal(content_mu.shape)).astype(np.float32)


def resolve_style_ids(stack, style_lyr_ids=None):
    """methods.py:60-66 -- explicit ids, else one 10-layer stack, else all 30 layers."""
    if style_lyr_ids is not None:
        assert isinstance(style_lyr_ids, (tuple, list)), "style_lyr_ids must be of type tuple or list!"
        return [int(i) for i in style_lyr_ids]
    if stack is not None:
        return list(range(stack * 10, stack * 10 + 10))
    return list(range(30))


def layers_needed(cont_ids, style_ids) -> int:
    """NL: TF prunes every encoder layer above the highest fetched extract.
    extracts[30] aliases extracts[29] (model.py:118-119)."""
    top = max(list(cont_ids) + list(style_ids))
    if top == 30:
        top = 29
    if top > 29 or min(list(cont_ids) + list(style_ids)) < 0:
        raise ValueError("extracts[31] (bottleneck) and negative ids are outside the hot path")
    return top + 1
