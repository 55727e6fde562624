"""methods.py of the reference, on the B200 step: same class, method names, argument meaning and CLI.

Mirrors /root/reference/methods.py (cited per method).  What changes underneath: the TF1 graph +
session is replaced by one `StyleStep` context (hand-written sm_100a CUDA behind the C ABI of
include/ast_b200.h); `tf.contrib.opt.ScipyOptimizerInterface` is replaced by a direct call of
`scipy.optimize.minimize(method='L-BFGS-B')` on `StyleStep.step_host` (float64 host buffers each way,
one CUDA-graph replay per function evaluation).  `sess` arguments are kept for call compatibility and
ignored.  Outputs that need librosa / matplotlib (PNG plots) are skipped: those packages are not part
of this path; wavs are written with scipy.io.wavfile.

No CPU fallback: constructing GatysNet without an sm_100 GPU raises.
"""
from __future__ import annotations

import argparse
import os
import time

import numpy as np

from . import synth
from .mu_law import inv_mu_law_numpy, mu_law_numpy

abbrevs = {'length': 'l', 'layers': 'lyr', 'n_components': 'cpn', 'examples': 'ex', 'epochs': 'ep',
           'qualities': 'qult', 'lambd': 'lbd', 'batch_size': 'btch', 'stack': 'stk'}   # utils.py:14-15


def l2_normalize(G, eps=1e-12):
    """tf.nn.l2_normalize(G, axis=(1, 2)) (methods.py:74,212)."""
    ss = np.sum(G * G, axis=(-2, -1), keepdims=True)
    return G / np.sqrt(np.maximum(ss, eps))


def load_audio(fn, sr, audio_channel):
    """utils.load_audio (utils.py:260-265) without librosa: wav only, no resampling."""
    from scipy.io import wavfile
    file_sr, a = wavfile.read(fn)
    if file_sr != sr:
        raise ValueError(f"{fn}: sample rate {file_sr} != --sr {sr} (resampling needs librosa, which is not part of this path)")
    if a.dtype.kind == 'i':
        a = a.astype(np.float32) / float(np.iinfo(a.dtype).max + 1)
    a = a.astype(np.float32)
    if a.ndim > 1:
        return a[:, audio_channel], sr
    return a, sr


def write_wav(path, audio, sr):
    from scipy.io import wavfile
    wavfile.write(path, sr, np.asarray(audio, np.float32))


def gt_s_path(suppath, **kwargs):
    """utils.gt_s_path (utils.py:18-64): run directory named after the arguments."""
    path = ''
    for name, value in sorted(kwargs.items()):
        if name == 'cont_fn':
            path += '_cnt_{}_'.format(value)
        elif name == 'style_fn':
            path += '_style_{}_'.format(value)
        elif name == 'gatys':
            path = ('gatys_' if value else 'ours_') + path
        elif name == 'sr':
            path += '_sr{}kHz_'.format(value / 1000)
        elif not name.endswith(('dir', 'path', 'pieces')) and value is not None:
            if name in abbrevs:
                name = abbrevs[name]
            if isinstance(value, (list, tuple)):
                value = ''.join('-%d' % i for i in value)
            path += '_{}_{}_'.format(name, value)
    path = os.path.join(suppath, path)
    os.makedirs(path, exist_ok=True)
    return path


def crt_t_fol(suppath, hour=False):
    """utils.crt_t_fol (utils.py:67-76)."""
    dte = time.localtime()
    fol_n = os.path.join(suppath, ('{}{}{}{}' if hour else '{}{}').format(*([dte[1], dte[2], dte[3], dte[4]] if hour else [dte[1], dte[2]])))
    os.makedirs(fol_n, exist_ok=True)
    return fol_n


class GatysNet(object):
    """methods.py:19-216."""

    def __init__(self,
                 savepath='./data/out',
                 checkpoint_path='./nsynth/model/wavenet-ckpt/model.ckpt-200000',
                 logdir='./log',
                 figdir='./data/fig',
                 stack=0,
                 batch_size=16384,
                 sr=16000,
                 cont_lyr_ids=[29],
                 nb_channels=128,
                 cnt_channels=128,
                 gatys=False,
                 style_lyr_ids=None,
                 precision='bf16', seed=0, device=None, optimizer='scipy', random_init=False, time_shards=False):
        if optimizer not in ('scipy', 'device'):
            raise ValueError("optimizer must be 'scipy' (the reference's SciPy L-BFGS-B on host buffers) or 'device'")
        self.optimizer = optimizer
        self.logdir = logdir
        self.savepath = savepath
        self.checkpoint_path = checkpoint_path
        self.figdir = figdir
        self.batch_size = batch_size
        self.sr = sr
        self.late = (batch_size - (batch_size // 4096) * 4000) // 2          # methods.py:39
        self.cont_lyr_ids = cont_lyr_ids
        self.seed = seed
        self.random_init = bool(random_init)
        # time_shards: ONE clip evaluated by all ranks of the torch.distributed job (one process per GPU, launched with
        # torchrun): every rank owns a slice of the clip and exchanges halo rows with its neighbours (sharding.py).
        # The reference is single-device (methods.py:48,143); this is how --batch_size beyond one GPU's reach is run.
        self.time_shards = bool(time_shards)
        self.sharded = None
        if self.time_shards and optimizer != 'scipy':
            raise ValueError("time_shards runs SciPy's L-BFGS-B on every rank (the device optimiser's vectors are not sharded)")
        self.precision = precision
        self.graph, self.embeds_c, self.embeds_s, self.gatys = self.build(
            batch_size, cont_lyr_ids, stack, nb_channels, cnt_channels, gatys, style_lyr_ids,
            precision=precision, device=device)
        # the optimisation variable 'regenerated_wav' (methods.py:49-54): zeros + 1e-6, mu-law units
        self.x = np.zeros([1, batch_size], np.float64) + 1e-6

    @staticmethod
    def build(length, cont_lyr_ids, stack, nb_channels, cnt_channels=128, gatys=False, style_lyr_ids=None,
              precision='bf16', device=None):
        """methods.py:44-77: static shapes (length, layer sets, head) -> one StyleStep context.
        Returns (graph, cont_embeds, style_embeds, gatys) like the reference; `graph` is the StyleStep and the
        two embeds entries are the names of its heads."""
        from .step import StyleStep
        if length % 512:
            raise AssertionError("length must be divisible by every dilation (masked.py:134)")
        style_ids = synth.resolve_style_ids(stack, style_lyr_ids)            # methods.py:60-66
        graph = StyleStep(length, list(cont_lyr_ids), style_ids, bool(gatys), nb_channels, cnt_channels,
                          precision=precision, device=device)
        return graph, 'content', 'style', gatys

    def load_model(self, sess=None):
        """methods.py:79-84.  `checkpoint_path` is what the reference takes: the prefix of a TF V2 checkpoint
        (`<prefix>.index` + `<prefix>.data-*`, read by tf_checkpoint.py without TensorFlow; the plain variable names,
        as tf.train.Saver over tf.global_variables() restores them).  `.npz` weight files with the ast_set_weights
        layouts are accepted too.  A missing or unreadable checkpoint RAISES, as tf.train.Saver.restore does; a seeded
        random-init encoder (masked.py:116-117) is used only when asked for (`random_init=True`, CLI --random_init /
        --synthetic) or when `checkpoint_path` is None."""
        from . import tf_checkpoint
        path = str(self.checkpoint_path) if self.checkpoint_path else ''
        if path and tf_checkpoint.is_bundle(path):
            w = tf_checkpoint.load_encoder_weights(path, n_layers=self.graph.n_layers)
        elif path.endswith('.npz') and os.path.exists(path):
            d = np.load(path)
            w = synth.EncoderWeights(d['W0'], d['b0'], d['Wd'], d['bd'], d['Wr'], d['br'])
        elif self.random_init or not path:
            print(f'[ast_b200] random-init encoder (seed {self.seed}) on request: no checkpoint is read')
            w = synth.make_weights(self.graph.n_layers, seed=self.seed)
        else:
            raise FileNotFoundError(
                f'checkpoint {self.checkpoint_path!r} is neither a TF V2 bundle prefix (<prefix>.index + .data-*) nor an .npz '
                'weight file; pass --random_init (or --synthetic) to run with a seeded random-init encoder instead')
        self.weights = w
        self.graph.set_weights(w)

    def get_embeds(self, sess, aud, is_content=True):
        """methods.py:86-95."""
        aud = np.asarray(aud)
        if len(aud.shape) == 1:
            aud = aud[: self.batch_size]
            aud = np.reshape(aud, [1, self.batch_size])
        Ec, Gn = self.graph.embeds(mu_law_numpy(aud).astype(np.float32), content=is_content, style=not is_content)
        out = Ec if is_content else Gn
        return out[0].cpu().numpy()

    def _style_chunks(self, filename, max_examples=5):
        """The consecutive chunks get_style_phi averages over (methods.py:100-108)."""
        audio = filename if isinstance(filename, np.ndarray) else load_audio(filename, sr=self.sr, audio_channel=0)[0]
        I = []
        i = 0
        while i + self.batch_size <= min(len(audio), max_examples * self.batch_size):
            I.append(np.asarray(audio[i: i + self.batch_size]))
            i += self.batch_size
        return I

    def embeds_batched(self, chunks, content=False, style=True):
        """The embeddings of several equally long clips from ONE B = len(chunks) forward (ast_embeds with B > 1)
        instead of len(chunks) B = 1 forwards (the reference runs them one session.run at a time, methods.py:97-111,
        192-212).  Pairs of a batch are bit-identical to single-pair runs, so the result equals the
        looped one bit for bit.  Returns numpy (Ec [n, T, cols] | None, Gn [n, ...] | None)."""
        from .step import StyleStep
        n = len(chunks)
        if n == 0:
            raise ValueError('no full chunk of batch_size samples in the audio')
        g = self.graph
        x = mu_law_numpy(np.stack([np.asarray(c)[: self.batch_size] for c in chunks])).astype(np.float32)
        if n == 1:
            ctx = g
        else:
            ctx = StyleStep(g.T, g.cont_ids, g.style_ids, g.gatys, g.nb_channels, g.cnt_channels, B=n,
                            precision=g.precision, device=g.device)
            ctx.set_weights(self.weights)
        try:
            Ec, Gn = ctx.embeds(x, content=content, style=style)
            Ec = None if Ec is None else Ec.cpu().numpy()
            Gn = None if Gn is None else Gn.cpu().numpy()
        finally:
            if ctx is not g:
                ctx.close()
        return Ec, Gn

    def get_style_phi(self, sess, filename, max_examples=5, show_mat=True):
        """methods.py:97-111: mean style embedding over <= max_examples consecutive chunks -- one batched forward."""
        _, Gn = self.embeds_batched(self._style_chunks(filename, max_examples))
        return np.mean(list(Gn), axis=0)

    def precompute_targets(self, aud, source, target, max_examples=5):
        """Everything GatysNet.run needs before the loop (methods.py:192-212) -- phi_c, phi(content clip), the target's
        and the source's mean style embedding -- from ONE batched forward over <= 1 + 5 + 5 clips.
        Returns (phi_c, phi, phi_t, phi_s) with the values of the one-by-one calls, bit for bit."""
        ct, cs = self._style_chunks(target, max_examples), self._style_chunks(source, max_examples)
        Ec, Gn = self.embeds_batched([aud] + ct + cs, content=True, style=True)
        phi_t = np.mean(list(Gn[1: 1 + len(ct)]), axis=0)
        phi_s = np.mean(list(Gn[1 + len(ct):]), axis=0)
        return Ec[0], Gn[0], phi_t, phi_s

    def define_loss(self, name, stl_emb, cnt_emb, lambd, gamma, gpu=None):
        """methods.py:113-138: bakes the targets and loss weights; returns the evaluation closure that
        ScipyOptimizerInterface would have built: fun(x64) -> (loss, grad64) and the last loss parts."""
        last = {}
        if self.time_shards:
            return self._define_loss_sharded(stl_emb, cnt_emb, lambd, gamma, last), last
        self.graph.set_targets(np.asarray(cnt_emb)[None], np.asarray(stl_emb)[None], lambd, gamma)

        def fun(x64):
            losses, grad = self.graph.step_host(x64)
            last['losses'] = losses[0]
            return float(losses[0, 0]), grad[0]
        return fun, last

    def _define_loss_sharded(self, stl_emb, cnt_emb, lambd, gamma, last):
        """define_loss for a clip sharded over the ranks: fun(x64) evaluates this rank's window with halo exchange
        (HaloShardedStep), all-gathers the owned gradient slices and returns the SAME (loss, gradient) on every rank, so
        that every rank's SciPy takes the same iterations from the same numbers and no broadcast of x is needed."""
        import torch
        import torch.distributed as dist
        from .sharding import HaloShardedStep
        from .step import StyleStep
        if gamma != 0.0:
            raise ValueError('time_shards: the gamma term is supported by the overlapped-window path only (sharding.TimeShardedStep)')
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError('time_shards needs an initialised torch.distributed process group (launch with torchrun)')
        rank, world = dist.get_rank(), dist.get_world_size()
        g = self.graph

        def make_ctx(Tw, window):
            c = StyleStep(Tw, g.cont_ids, g.style_ids, g.gatys, g.nb_channels, g.cnt_channels, precision=g.precision,
                          device=g.device, window=window, halo_exchange=True)
            c.set_weights(self.weights)
            return c

        def exchange(pairs):
            ops = []
            for send, recv, peer in pairs:
                ops.append(dist.P2POp(dist.isend, send, peer))
                ops.append(dist.P2POp(dist.irecv, recv, peer))
            for r in dist.batch_isend_irecv(ops):
                r.wait()
        if self.sharded is not None:
            self.sharded.ctx.close()
        hs = HaloShardedStep(self.batch_size, rank, world, make_ctx, lambda p: dist.all_reduce(p), exchange)
        self.sharded = hs
        hs.ctx.set_targets(np.ascontiguousarray(np.asarray(cnt_emb)[None, hs.wb:hs.we]), np.asarray(stl_emb)[None], lambd, 0.0)
        dev = hs.ctx.dev
        losses = torch.empty((1, 4), device=dev)
        grad = torch.empty((1, hs.we - hs.wb), device=dev)
        sizes = [oe - ob for (_, _, ob, oe) in hs.plan]
        pad = max(sizes)
        mine = torch.zeros(pad, device=dev)
        parts = [torch.empty(pad, device=dev) for _ in range(world)]

        def fun(x64):
            xw = torch.as_tensor(np.ascontiguousarray(x64[hs.wb:hs.we], dtype=np.float32)[None], device=dev)
            _, own = hs.step(xw, losses, grad)
            mine[:own.shape[1]] = own[0]
            dist.all_gather(parts, mine)
            hs.ctx.check()
            full = torch.cat([p[:n] for p, n in zip(parts, sizes)]).double().cpu().numpy()
            lo = losses[0].double().cpu().numpy()
            last['losses'] = lo
            return float(lo[0]), full
        return fun

    def l_bfgs(self, sess, phi_c, phi_s, epochs, lambd, gamma, callback=None):
        """methods.py:140-181: `epochs` fresh L-BFGS-B runs of maxiter=100 from the variable's current value;
        stop when an epoch used < 50 evaluations."""
        from scipy.optimize import minimize
        fun, last = self.define_loss('loss', phi_s, phi_c, lambd, gamma)
        print('Saving file ... to fol {{{}}}'.format(self.savepath))
        since = time.time()
        if self.optimizer == 'device':
            return self._l_bfgs_device(epochs, since, callback)
        i_ = 0
        self.evaluations = 0
        for ep in range(epochs):
            i = 0

            def tracked(x64):
                nonlocal i
                f, g = fun(x64)
                lo = last['losses']
                if not i % 5:
                    print('Ep {0:}/{1:}-it {2:}({3:})-tlapse {4:.4f}s-loss{5:.4f}-{6:.4f}-{7:.4f}-{8:.4f}'.format(
                        ep + 1, epochs, i, i_, time.time() - since, lo[0], lo[1], lo[2], lo[3]), end='\r', flush=True)
                if callback is not None:
                    callback(ep, i, lo)
                i += 1
                return f, g
            res = minimize(tracked, self.x[0], jac=True, method='L-BFGS-B', options={'maxiter': 100})
            self.x = res.x[None].astype(np.float64)
            self.evaluations += i
            i_ = i
            audio = inv_mu_law_numpy(self.x)[0, self.late:-self.late]
            if self.savepath and (self.sharded is None or self.sharded.rank == 0):
                os.makedirs(self.savepath, exist_ok=True)
                write_wav(os.path.join(self.savepath, 'ep-{}.wav'.format(ep)), audio / max(np.max(audio), 1e-12), self.sr)
            if i_ < 50:
                break
        print()

    def _l_bfgs_device(self, epochs, since, callback):
        """The same loop with the optimiser resident on the device (csrc/lbfgs.cu): every epoch of every evaluation is
        enqueued without a host round trip; the per-evaluation losses and the per-epoch iterates the reference prints /
        writes (methods.py:147-157,169-179) are read back once at the end."""
        from .step import DeviceLBFGS
        opt = DeviceLBFGS(self.batch_size, B=1, maxiter=100, epochs=epochs, min_evals=50, max_log=min(epochs * 256, 1 << 18),
                          keep_epoch_x=True, device=self.graph.device)
        try:
            opt.reset(self.x)
            opt.minimize(self.graph)
            status, _ = opt.status()
            nfev, end, xs = opt.epochs_info()
            log = opt.log()[0]
            self.x = opt.result().astype(np.float64)
        finally:
            opt.close()
        done_epochs = int(status[0, 1])
        self.evaluations = int(status[0, 4])
        row = 0
        for ep in range(done_epochs):
            n = int(nfev[0, ep])
            for i in range(n):
                lo = log[row] if row < len(log) else np.full(4, np.nan, np.float32)
                if callback is not None:
                    callback(ep, i, lo)
                row += 1
            lo = log[min(row, len(log)) - 1]
            print('Ep {0:}/{1:}-it {2:}-tlapse {3:.4f}s-loss{4:.4f}-{5:.4f}-{6:.4f}-{7:.4f} ({8:})'.format(
                ep + 1, epochs, n, time.time() - since, lo[0], lo[1], lo[2], lo[3], DeviceLBFGS.END_REASONS[int(end[0, ep])]))
            audio = inv_mu_law_numpy(xs[0, ep].astype(np.float64)[None])[0, self.late:-self.late]
            if self.savepath:
                os.makedirs(self.savepath, exist_ok=True)
                write_wav(os.path.join(self.savepath, 'ep-{}.wav'.format(ep)), audio / max(np.max(audio), 1e-12), self.sr)

    def run(self, cont_file, source, target, epochs, lambd=0.1, gamma=0.1, audio_channel=0, start=1.0):
        """methods.py:183-216.  File names may also be numpy arrays of audio in [-1, 1)."""
        self.load_model(None)
        ld = lambda f: f if isinstance(f, np.ndarray) else load_audio(f, sr=self.sr, audio_channel=audio_channel)[0]
        aud = ld(cont_file)
        st = max(int(start * self.sr - self.late), 0)
        aud = aud[st: st + self.batch_size]
        if self.savepath and int(os.environ.get('RANK', '0')) == 0:
            os.makedirs(self.savepath, exist_ok=True)
            write_wav(os.path.join(self.savepath, 'ori.wav'), aud[self.late:-self.late], self.sr)
            write_wav(os.path.join(self.savepath, 'style.wav'), ld(target)[st: st + self.batch_size][self.late:-self.late], self.sr)
        phi_c, phi, phi_t, phi_s = self.precompute_targets(aud, ld(source), ld(target))
        phi = l2_normalize(phi + phi_t - phi_s)                               # methods.py:211-212
        self.l_bfgs(None, phi_c, phi, epochs=epochs, lambd=lambd, gamma=gamma)
        return inv_mu_law_numpy(self.x)[0]


def get_dir(dir, args):
    return gt_s_path(crt_t_fol(dir), **{k: v for k, v in vars(args).items() if k not in ('seed', 'precision', 'synthetic', 'optimizer', 'random_init', 'time_shards')})


def get_fpath(fn, args):
    return os.path.join(args.dir, fn) + '.wav'


def piece_work(args):
    """methods.py:227-240."""
    savepath, logdir = map(lambda dir: get_dir(dir, args), [args.outdir, args.logdir])
    figdir = os.path.join(savepath, 'fig')
    os.makedirs(figdir, exist_ok=True)
    if args.synthetic:
        content = synth.content_audio(args.batch_size + 2 * args.sr, seed=1)
        style = synth.style_audio(args.batch_size, seed=2, chunks=5)
    else:
        content, style = map(lambda name: get_fpath(name, args), [args.cont_fn, args.style_fn])
    device = None
    if args.time_shards:                       # one process per GPU (torchrun): rank r drives GPU LOCAL_RANK
        import torch
        import torch.distributed as dist
        device = int(os.environ.get('LOCAL_RANK', '0'))
        torch.cuda.set_device(device)
        if not dist.is_initialized():
            os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
            dist.init_process_group('nccl', device_id=torch.device('cuda', device))
    test = GatysNet(savepath, args.ckpt_path, logdir, figdir, args.stack, args.batch_size, args.sr, args.cont_lyrs,
                    args.channels, args.cnt_channels, args.gatys, args.style_lyrs, precision=args.precision, seed=args.seed,
                    optimizer=args.optimizer, random_init=args.random_init or args.synthetic, time_shards=args.time_shards,
                    device=device)
    return test.run(content, content, style, epochs=args.epochs, lambd=args.lambd, gamma=args.gamma, start=args.start)


def build_parser():
    """methods.py:243-267, flag for flag (incl. the `--gatys` type=bool quirk), plus --seed/--precision/--synthetic."""
    parser = argparse.ArgumentParser()
    parser.add_argument('cont_fn', help='relative content file name')
    parser.add_argument('style_fn', help='relative style file name')
    parser.add_argument('--epochs', help='number of epochs, each epoch contains 100 iterations of optimization',
                        nargs='?', type=int, default=100)
    parser.add_argument('--batch_size', help='length of output signal, must be divided by 4096', nargs='?', type=int, default=16384)
    parser.add_argument('--sr', help='sampling rate, default to 16kHz', nargs='?', type=int, default=16000)
    parser.add_argument('--stack', nargs='?', type=int, default=None,
                        help='stack of layers chosen for computing style loss (only if style_lyrs is None); None = all three stacks')
    parser.add_argument('--cont_lyrs', nargs='*', type=int, default=[29])
    parser.add_argument('--style_lyrs', nargs='*', type=int)
    parser.add_argument('--lambd', help='style loss scalar coefficient', nargs='?', type=float, default=100.0)
    parser.add_argument('--gamma', help='regularizer scalar coefficient', nargs='?', type=float, default=0.0)
    parser.add_argument('--channels', help='how many channels taken into account for style loss', nargs='?', type=int, default=128)
    parser.add_argument('--cnt_channels', help='how many channels taken into account for content loss', nargs='?', type=int, default=128)
    parser.add_argument('--start', nargs='?', type=float, default=1.0)
    parser.add_argument('--gatys', nargs='?', type=bool, default=False, const=True)
    parser.add_argument('--ckpt_path', nargs='?', default='./nsynth/model/wavenet-ckpt/model.ckpt-200000',
                        help="path to the pretrained model's checkpoint (TF checkpoint prefix, or .npz with W0,b0,Wd,bd,Wr,br)")
    parser.add_argument('--dir', help='path to source files', nargs='?', default='./data/src')
    parser.add_argument('--outdir', help='path to output', nargs='?', default='./data/out')
    parser.add_argument('--logdir', help='path to logs', nargs='?', default='./log')
    parser.add_argument('--cmt')
    # additions (not in the reference)
    parser.add_argument('--seed', type=int, default=0, help='seed of the random-init encoder (--random_init / --synthetic)')
    parser.add_argument('--random_init', action='store_true',
                        help='run with a seeded random-init encoder instead of a checkpoint (without it a missing --ckpt_path is an error, as in the reference)')
    parser.add_argument('--precision', default='bf16', choices=['bf16', 'fp32'])
    parser.add_argument('--optimizer', default='scipy', choices=['scipy', 'device'],
                        help="scipy: SciPy L-BFGS-B on host buffers, as the reference; device: the same algorithm resident on the GPU")
    parser.add_argument('--synthetic', action='store_true', help='ignore the file names, use seeded synthetic audio')
    parser.add_argument('--time_shards', action='store_true',
                        help='one clip over all ranks of the job (launch with torchrun, one process per GPU): each rank owns a '
                             'slice of the clip, halo rows are exchanged between neighbours, Gram partials all-reduced')
    return parser


def main(argv=None):
    piece_work(build_parser().parse_args(argv))


if __name__ == '__main__':
    main()
