"""Multi-process CPU tests (gloo, world_size 2) of the N>1 host logic: window planning, the two-phase
step with ONE all-reduce of raw Gram / content partials, owned-gradient slicing, pair distribution.
The per-rank evaluator injected here is a CPU stand-in with the StyleStep two-phase interface (built on
the oracle's torch restatement); on the GPU box the same TimeShardedStep drives a StyleStep."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from audio_style_transfer_b200 import sharding, synth
from audio_style_transfer_b200.mu_law import mu_law_numpy
from oracle import step_np
from oracle.torch_step import TorchStep


def test_plan_windows_tile_the_clip():
    for T, P, NL in [(65536, 2, 26), (65536, 8, 30), (479232, 8, 30), (16384, 4, 10)]:
        plan = sharding.plan_windows(T, P, NL)
        R = sharding.receptive_field(NL)
        assert plan[0][2] == 0 and plan[-1][3] == T
        for (wb, we, ob, oe), nxt in zip(plan, plan[1:] + [None]):
            assert wb % 128 == 0 and we % 128 == 0 and ob % 128 == 0 and oe % 128 == 0
            assert wb <= ob < oe <= we
            assert wb == 0 or ob - wb >= 2 * R          # forward exact on [ob-R, oe+R), gradient exact on owned rows
            assert we == T or we - oe >= 2 * R
            if nxt:
                assert oe == nxt[2]
    assert sharding.receptive_field(30) == 3070 and sharding.receptive_field(26) == 2110
    with pytest.raises(ValueError):
        sharding.plan_windows(1000, 2, 10)


def test_plan_windows_hold_the_stft_frames_of_the_owned_samples():
    """gamma != 0 (methods.py:121-123): a rank evaluates every 1024-sample / hop-512 STFT frame that touches one of
    its owned samples, so all of them must lie inside its window; frames that START in the owned range are each
    counted by exactly one rank."""
    for T, P, NL in [(8192, 3, 6), (65536, 8, 30), (16384, 4, 2), (16384 + 384, 4, 2)]:
        plan = sharding.plan_windows(T, P, NL, min_halo=sharding.STFT_HALO)
        F = 1 + (T - 1024) // 512
        counted = []
        for wb, we, ob, oe in plan:
            touching = [f for f in range(F) if 512 * f < oe and 512 * f + 1024 > ob]
            assert all(512 * f >= wb and 512 * f + 1024 <= we for f in touching), (T, P, NL, wb, we)
            counted += [f for f in range(F) if ob <= 512 * f < oe]
        assert sorted(counted) == list(range(F))
    short = sharding.plan_windows(8192, 3, 6)            # without the regulariser the halo is the encoder's 2R only
    assert short[1][2] - short[1][0] < 1024


def test_exchange_windows_and_the_capture_guard():
    """Halo-exchange windows: owned range +- 2 * 1024 rows, every owned range long enough for what the neighbours fetch;
    capture() (one CUDA graph per rank) is refused unless the halos are peer-mapped -- the NCCL send/recv route exchanges
    rows between host calls."""
    for T, P in [(65536, 2), (65536, 8), (479232, 8), (479232, 4)]:
        plan = sharding.plan_windows_exchange(T, P)
        assert plan[0][2] == 0 and plan[-1][3] == T
        for (wb, we, ob, oe), nxt in zip(plan, plan[1:] + [None]):
            assert wb == max(0, ob - 2 * sharding.HALO) and we == min(T, oe + 2 * sharding.HALO)
            assert oe - ob >= 2 * sharding.HALO and ob % 128 == 0 and oe % 128 == 0
            if nxt:
                assert oe == nxt[2]
    with pytest.raises(ValueError):
        sharding.plan_windows_exchange(16384, 16)          # 1024-sample shards cannot serve 2048 halo rows

    class Ctx:
        segments = 3
    hs = sharding.HaloShardedStep(65536, 1, 4, lambda Tw, window: Ctx(), lambda p: None, lambda pairs: None)
    assert (hs.up, hs.dn) == (0, 2) and hs.we - hs.wb == 16384 + 4 * sharding.HALO
    with pytest.raises(RuntimeError, match="attach_p2p"):
        hs.capture(None, None, None)


def test_pairs_for_rank_cover_every_pair_once():
    got = sorted(p for r in range(8) for p in sharding.pairs_for_rank(64, r, 8))
    assert got == list(range(64))


class CpuWindowStep:
    """StyleStep's two-phase interface on the CPU (fp64 torch autograd): forward_partial / partials / backward."""

    def __init__(self, w, cont, style, gatys, lambd, Pc_win, Ps, window):
        self.ts = TorchStep(w, cont, style, gatys, lambd, dtype=torch.float64)
        self.total_T, self.wb, self.ob, self.oe = window
        self.Pc, self.Ps, self.gatys, self.lambd = torch.tensor(Pc_win), torch.tensor(Ps), gatys, lambd

    def forward_partial(self, x):
        self.x = torch.tensor(x, dtype=torch.float64).requires_grad_(True)
        ex = self.ts.forward(self.x)
        lo, hi = self.ob - self.wb, self.oe - self.wb

        def heads(a, b):
            g = lambda i: ex[min(i, 29)][:, :, a:b]
            Ec = torch.cat([g(i) for i in self.ts.cont_ids], dim=1).transpose(1, 2)
            S = torch.stack([g(i) for i in self.ts.style_ids], dim=1)
            if not self.gatys:
                S = S.transpose(1, 2)
            return S @ S.transpose(2, 3), ((Ec - self.Pc[:, a:b]) ** 2).sum()
        G_own, sse_own = heads(lo, hi)                       # what this rank contributes to the all-reduce
        # the gradient flows through EVERY window row (the head gradient dF is formed for the whole window from
        # the global dG, as ast_step_backward does); rows farther than R from the owned range are inexact but
        # cannot reach the owned samples
        self.G, self.sse = heads(0, self.x.shape[1])
        self.buf = torch.cat([G_own.detach().reshape(-1), sse_own.detach().reshape(1)])

    def partials(self):
        return self.buf

    def backward(self):
        n = self.G.numel()
        G = self.G + (self.buf[:n].reshape(self.G.shape) - self.G.detach())          # total, gradient through the local part
        sse = self.sse + (self.buf[n] - self.sse.detach())
        ss = (G * G).sum(dim=(2, 3), keepdim=True)
        Gn = G * torch.rsqrt(torch.clamp(ss, min=1e-12))
        Ls = 1e3 * ((Gn - self.Ps) ** 2).mean()
        Lc = 10.0 * sse / (self.total_T * self.Pc.shape[2])
        tot = Lc + self.lambd * Ls
        (dx,) = torch.autograd.grad(tot, self.x)
        return torch.stack([tot, Lc, Ls, torch.zeros(())]).detach()[None], dx


def _worker(rank, world, port, T, NL, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    cont, style, gatys, lambd = [NL - 1], list(range(NL)), rank < 0, 100.0
    w = synth.make_weights(NL, 0, 0.1)
    c_mu = mu_law_numpy(synth.content_audio(T)).astype(np.float64)
    x = synth.step_input(c_mu).astype(np.float64)
    Pc, Gc = step_np.embeds(c_mu, w, cont, style, gatys)
    Ps = step_np.l2_normalize(np.roll(Gc, 1, axis=0))

    def make_ctx(Tw, window):
        total_T, wb, ob, oe = window
        return CpuWindowStep(w, cont, style, gatys, lambd, Pc[None, wb:wb + Tw], Ps[None], window)
    ts = sharding.TimeShardedStep(T, NL, rank, world, make_ctx, lambda p: dist.all_reduce(p))
    losses, g_own = ts.step(ts.window(x[None]))
    lo, dx = step_np.step(x, w, Pc, Ps, cont, style, gatys, lambd)
    err_l = abs(float(losses[0, 0]) - lo[0]) / abs(lo[0])
    err_g = np.linalg.norm(g_own[0].numpy() - dx[ts.ob:ts.oe]) / np.linalg.norm(dx[ts.ob:ts.oe])
    gathered = [None] * world
    dist.all_gather_object(gathered, (ts.ob, ts.oe, err_l, err_g))
    if rank == 0:
        ret.put(gathered)
    dist.destroy_process_group()


def test_time_sharded_step_world2_gloo():
    """T=8192 over 2 ranks, 4 layers (receptive field 16 -> halo 128): every rank sees the same loss as the
    unsharded oracle and its owned gradient slice matches to fp64 round-off."""
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 8192, 4, ret)) for r in range(2)]
    for p in procs:
        p.start()
    out = ret.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [o[:2] for o in out] == [(0, 4096), (4096, 8192)]
    for ob, oe, err_l, err_g in out:
        assert err_l < 1e-10 and err_g < 1e-9, out
