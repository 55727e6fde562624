"""Multi-GPU host logic: time sharding of one long clip, and data parallelism over pairs.

The reference is single-device (methods.py:48,143); both modes are new capabilities asked for by
BASELINE.json configs[2..4].

Time sharding = overlapped windows.  Rank r owns samples [own_b, own_e) and evaluates the ordinary
single-GPU step on the window [own_b - 2R, own_e + 2R) clipped to the clip, R = 1 + sum_l d_l the
encoder's one-sided receptive field.  Zero padding at an artificial window edge is wrong, but the
error travels at most R rows inwards in the forward and another R in the backward, so features are
exact on [own_b - R, own_e + R) and the gradient on the owned rows.  True clip ends keep the
reference's zero padding (masked.py:139).  The only exchange is ONE all-reduce (sum) per step of the
raw Gram partials + content squared-error partials (51 KB - 2 MB), after which every rank normalises,
forms the (identical) losses and back-propagates locally.  No per-layer halo traffic.
"""
from __future__ import annotations

from typing import Callable, List, Tuple

ALIGN = 128      # ast_config.T must be a multiple of 128


def receptive_field(n_layers: int) -> int:
    """One-sided receptive field of the start conv + n_layers dilated layers (model.py:88-106):
    1 + sum 2**(l % 10).  3070 for 30 layers, 2110 for 26."""
    return 1 + sum(1 << (l % 10) for l in range(n_layers))


STFT_HALO = 1024    # gamma != 0: every 1024-sample STFT frame (methods.py:122) touching an owned sample must lie in the window


def plan_windows(T: int, P: int, n_layers: int, min_halo: int = 0) -> List[Tuple[int, int, int, int]]:
    """[(win_begin, win_end, own_begin, own_end)] per rank; all multiples of 128, windows clipped to
    [0, T).  Owned ranges tile [0, T) exactly.  min_halo widens the overlap beyond the encoder's 2R
    (STFT_HALO when the regulariser is on)."""
    if T % ALIGN:
        raise ValueError(f"T={T} must be a multiple of {ALIGN}")
    blocks = T // ALIGN
    if P > blocks:
        raise ValueError(f"cannot split {blocks} blocks of {ALIGN} samples over {P} ranks")
    halo = -(-max(2 * receptive_field(n_layers), min_halo) // ALIGN) * ALIGN
    plan = []
    for r in range(P):
        ob = (blocks * r // P) * ALIGN
        oe = (blocks * (r + 1) // P) * ALIGN
        plan.append((max(0, ob - halo), min(T, oe + halo), ob, oe))
    return plan


def pairs_for_rank(n_pairs: int, rank: int, world: int) -> range:
    """Data parallelism over independent content/style pairs: pair p -> rank p % world
    (SURVEY.md 8(e)); no collective in the step."""
    return range(rank, n_pairs, world)


class TimeShardedStep:
    """Drives one rank's window context.  `make_ctx(T_win, window)` builds the per-rank evaluator (a
    StyleStep on the GPU; tests inject a CPU evaluator to check this logic under gloo) exposing
    forward_partial(x) / partials() / backward() -> (losses, grad_window)."""

    def __init__(self, T: int, n_layers: int, rank: int, world: int, make_ctx: Callable, all_reduce: Callable,
                 gamma: float = 0.0):
        self.T, self.rank, self.world = T, rank, world
        self.plan = plan_windows(T, world, n_layers, STFT_HALO if gamma != 0.0 else 0)
        self.wb, self.we, self.ob, self.oe = self.plan[rank]
        self.ctx = make_ctx(self.we - self.wb, (T, self.wb, self.ob, self.oe))
        self.all_reduce = all_reduce

    def window(self, full):
        """Slice a [.., T, ...] clip-length array (x, Pc) to this rank's window along axis 1."""
        return full[:, self.wb:self.we]

    def step(self, x_window):
        """Returns (losses [B,4] -- identical on every rank, grad of the OWNED samples [B, own_e-own_b])."""
        self.ctx.forward_partial(x_window)
        self.all_reduce(self.ctx.partials())          # in place, sum over ranks
        losses, grad = self.ctx.backward()
        return losses, grad[:, self.ob - self.wb:self.oe - self.wb]
