"""torchrun target: the methods.py CLI on ONE clip time-sharded over the ranks (--time_shards) against the same run on one
GPU.  Every rank runs SciPy's L-BFGS-B on identical (loss, gradient) pairs, so all ranks must end with the same waveform, and
it must track the single-GPU run (bf16 tier: the Gram sums are split differently, so equal to optimisation noise, not bits).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/sharded_cli_check.py
"""
import os
import sys
import tempfile
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from audio_style_transfer_b200 import methods  # noqa: E402

T = int(os.environ.get("CHECK_T", "32768"))
tmp = tempfile.mkdtemp()
common = ['c', 's', '--synthetic', '--epochs', '1', '--batch_size', str(T), '--stack', '0', '--cont_lyrs', '25',
          '--outdir', tmp + '/out', '--logdir', tmp + '/log']
seen = []
orig = methods.GatysNet.l_bfgs


def spy(self, sess, phi_c, phi_s, epochs, lambd, gamma, callback=None):
    return orig(self, sess, phi_c, phi_s, epochs, lambd, gamma, callback=lambda ep, i, lo: seen.append(float(lo[0])))


methods.GatysNet.l_bfgs = spy
x_sh = methods.piece_work(methods.build_parser().parse_args(common + ['--time_shards']))
loss_sh = list(seen)
rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
xs = [torch.empty(T, dtype=torch.float64, device=dev) for _ in range(world)]
dist.all_gather(xs, torch.as_tensor(np.asarray(x_sh, np.float64), device=dev))
same = all(torch.equal(xs[0], t) for t in xs)
if rank == 0:
    seen.clear()
    x_1 = methods.piece_work(methods.build_parser().parse_args(common))
    rel = float(np.linalg.norm(x_sh - x_1) / np.linalg.norm(x_1))
    print(f"time_shards x{world}, T={T}: {len(loss_sh)} evaluations, loss {loss_sh[0]:.4f} -> {loss_sh[-1]:.4f}; one GPU: "
          f"{len(seen)} evaluations, loss {seen[0]:.4f} -> {seen[-1]:.4f}; first-loss rel diff {abs(loss_sh[0] - seen[0]) / seen[0]:.2e}; "
          f"all ranks hold the same waveform: {same}; waveform rel-L2 sharded vs one GPU {rel:.2e}", flush=True)
    assert same and abs(loss_sh[0] - seen[0]) <= 1e-3 * seen[0] and loss_sh[-1] < 0.6 * loss_sh[0]
dist.barrier()
dist.destroy_process_group()
