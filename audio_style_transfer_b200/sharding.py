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


HALO = 1024         # AST_HALO_ROWS: >= 1023 = sum of the dilations of one 10-layer stack (model.py:98)


def plan_windows_exchange(T: int, P: int) -> List[Tuple[int, int, int, int]]:
    """Windows for shards that EXCHANGE halo rows stack by stack (HaloShardedStep): owned ranges as in plan_windows,
    window = owned range +- 2 * HALO rows (clipped): HALO rows whose features must be exact for the backward of the
    owned rows, and HALO more that the forward of those needs and that are refreshed from the neighbour after every
    stack.  Every owned range must hold the 2 * HALO rows its neighbours fetch."""
    if T % ALIGN:
        raise ValueError(f"T={T} must be a multiple of {ALIGN}")
    blocks = T // ALIGN
    plan = []
    for r in range(P):
        ob = (blocks * r // P) * ALIGN
        oe = (blocks * (r + 1) // P) * ALIGN
        if P > 1 and oe - ob < 2 * HALO:
            raise ValueError(f"shard of {oe - ob} samples is shorter than the {2 * HALO} rows its neighbours fetch")
        plan.append((max(0, ob - 2 * HALO), min(T, oe + 2 * HALO), ob, oe))
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


class HaloShardedStep:
    """Time sharding with a receptive-field halo exchange (BASELINE.json north_star; SURVEY.md 8(e) option ii).

    The encoder is three stacks of ten layers with dilations 1..512 (model.py:96-106), so ONE stack moves information by
    at most 1023 rows.  Rank r evaluates its window stack by stack (StyleStep.forward_seg / backward_seg); after each
    forward stack it overwrites the outer HALO rows of its window (which its own zero-padded evaluation got wrong) with
    the neighbour's exact rows of the fp32 residual stream, after each backward stack the HALO rows next to its owned
    range with the neighbour's running gradient.  Per step and interior side: 2 + 2 messages of HALO x 128 fp32
    (512 KB), plus the one all-reduce of the raw Gram partials.  `exchange(pairs)` moves rows between ranks:
    pairs = [(send_tensor, recv_tensor, peer_rank), ...]; on NCCL it is one batch_isend_irecv, tests inject copies.

    attach_p2p() replaces those messages by peer-mapped memory: every rank maps its neighbours' buffers once (CUDA IPC) and
    the stack kernels read the halo rows over NVLink themselves, gated by completion words the neighbours' kernels write
    (include/ast_b200.h: ast_p2p_*).  A step is then forward, all-reduce, backward: no host-side exchange at all, and
    capture() turns it into one CUDA graph per rank."""

    def __init__(self, T: int, rank: int, world: int, make_ctx: Callable, all_reduce: Callable, exchange: Callable = None):
        self.T, self.rank, self.world = T, rank, world
        self.plan = plan_windows_exchange(T, world)
        self.wb, self.we, self.ob, self.oe = self.plan[rank]
        self.ctx = make_ctx(self.we - self.wb, (T, self.wb, self.ob, self.oe))
        self.all_reduce, self.exchange = all_reduce, exchange
        self.nseg = self.ctx.segments
        self.up = rank - 1 if rank > 0 else None             # neighbour that owns the samples before mine
        self.dn = rank + 1 if rank + 1 < world else None
        self.p2p = False

    def attach_p2p(self, all_gather_object: Callable):
        """all_gather_object(obj) -> [obj of rank 0, obj of rank 1, ...] (torch.distributed.all_gather_object).  Every rank
        must call this before any rank starts a step, and no rank may close its context while a neighbour still steps."""
        infos = all_gather_object((self.ctx.p2p_export(), self.wb, self.we - self.wb))
        for side, peer in ((0, self.up), (1, self.dn)):
            if peer is not None:
                handles, wb, Tw = infos[peer]
                self.ctx.p2p_attach(side, handles, wb, Tw)
        self.p2p = True

    def window(self, full):
        return full[:, self.wb:self.we]

    def capture(self, x_window, losses, grad, warmup: int = 2):
        """Peer-mapped halos only: captures forward -> all-reduce -> backward of this rank into ONE CUDA graph (the stack
        kernels, their flag memsets, the NCCL all-reduce) and returns `replay()`.  The evaluation number the completion words
        carry is counted in device memory on the stream (kernels_tc.cu: tc_peer_step_bump_kernel), so every replay is a new
        evaluation on every rank.  x_window / losses / grad are the graph's static buffers: write the next x into
        x_window in place.  Every rank must capture, and replay equally often."""
        import torch
        if not self.p2p:
            raise RuntimeError("capture() needs attach_p2p(): the NCCL send/recv route exchanges rows between host calls")
        for _ in range(max(1, warmup)):               # NCCL communicator and every lazy allocation exist before the capture
            self.step(x_window, losses, grad)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self.step(x_window, losses, grad)
        self._graph = g
        return g.replay

    def _swap(self, buf, send_up, recv_up, send_dn, recv_dn):
        """buf [B, T_win, 128]; the four (begin, end) ranges are global sample indices."""
        rows = lambda a: buf[:, a[0] - self.wb:a[1] - self.wb]
        pairs = []
        if self.up is not None:
            pairs.append((rows(send_up), rows(recv_up), self.up))
        if self.dn is not None:
            pairs.append((rows(send_dn), rows(recv_dn), self.dn))
        if pairs:
            self.exchange(pairs)

    def step(self, x_window, losses, grad):
        """Returns (losses [B,4] -- identical on every rank, grad of the OWNED samples [B, own_e-own_b])."""
        c, H, ob, oe = self.ctx, HALO, self.ob, self.oe
        if self.p2p:                                  # the kernels fetch the halo rows themselves
            c.forward_partial(x_window)
            self.all_reduce(c.partials())
            c.backward(losses, grad)
            return losses, grad[:, ob - self.wb:oe - self.wb]
        for k in range(self.nseg):
            c.forward_seg(x_window, k)
            if k + 1 < self.nseg:
                # my exact region is [ob-H, oe+H); the neighbour above needs [ob+H, ob+2H) of it, the one below [oe-2H, oe-H)
                self._swap(c.segment_buffer(False, k), (ob + H, ob + 2 * H), (ob - 2 * H, ob - H), (oe - 2 * H, oe - H), (oe + H, oe + 2 * H))
        self.all_reduce(c.partials())
        for k in range(self.nseg - 1, -1, -1):
            c.backward_seg(k, losses, grad)
            if k > 0:
                self._swap(c.segment_buffer(True, k), (ob, ob + H), (ob - H, ob), (oe - H, oe), (oe, oe + H))
        return losses, grad[:, ob - self.wb:oe - self.wb]
