"""StyleStep -- Python owner of one ast_ctx: the fwd + bwd + Gram-loss evaluation on one B200.

torch is used for device-memory ownership and streams only; all arithmetic happens in
libast_b200.so (hand-written sm_100a CUDA).  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .synth import EncoderWeights, layers_needed

PRECISIONS = {"fp32": _lib.AST_PREC_FP32, "bf16": _lib.AST_PREC_BF16, "tc3": _lib.AST_PREC_TC3}


class _DevMem:
    """__cuda_array_interface__ shim so torch can view library-owned device memory without a copy."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def _ptr(t):
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        assert t.is_contiguous()
        return C.c_void_p(t.data_ptr())
    if isinstance(t, np.ndarray):
        assert t.flags["C_CONTIGUOUS"]
        return C.c_void_p(t.ctypes.data)
    raise TypeError(type(t))


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class StyleStep:
    """One context = fixed (T, B, layer set, head, precision) -- the static shapes of GatysNet.build
    (methods.py:44-77).  `window=(total_T, win_begin, own_begin, own_end)` makes this context one
    time shard of a longer clip (multi-GPU)."""

    KERNEL_PATHS = {"auto": 0, "per_layer": 1, "per_layer_bwd": 2}

    def __init__(self, T, cont_ids, style_ids, gatys=False, nb_channels=128, cnt_channels=128, B=1,
                 precision="bf16", device=None, window=None, kernel_path="auto", experiments=False, halo_exchange=False):
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device: the B200 step has no CPU fallback")
        self.lib = _lib.load(experiments)
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.dev = torch.device("cuda", self.device)
        self.T, self.B = int(T), int(B)
        self.cont_ids, self.style_ids = [int(i) for i in cont_ids], [int(i) for i in style_ids]
        self.gatys = bool(gatys)
        self.nb_channels, self.cnt_channels = int(nb_channels), int(cnt_channels)
        self.n_layers = layers_needed(self.cont_ids, self.style_ids)
        self.L = len(self.style_ids)
        self.precision = precision
        cfg = _lib.AstConfig()
        cfg.T, cfg.B, cfg.n_layers = self.T, self.B, self.n_layers
        cfg.n_cont, cfg.n_style = len(self.cont_ids), self.L
        if cfg.n_cont > _lib.AST_MAX_IDS or cfg.n_style > _lib.AST_MAX_IDS:
            raise ValueError("too many layer ids")
        for i, v in enumerate(self.cont_ids):
            cfg.cont_ids[i] = v
        for i, v in enumerate(self.style_ids):
            cfg.style_ids[i] = v
        cfg.gatys, cfg.nb_channels, cfg.cnt_channels = int(self.gatys), self.nb_channels, self.cnt_channels
        cfg.precision = PRECISIONS[precision]
        if window is not None:
            cfg.total_T, cfg.win_begin, cfg.own_begin, cfg.own_end = [int(v) for v in window]
        cfg.device = self.device
        cfg.kernel_path = self.KERNEL_PATHS[kernel_path]
        cfg.halo_exchange = int(bool(halo_exchange))
        self._h = C.c_void_p()
        _lib.check(self.lib.ast_create(C.byref(cfg), C.byref(self._h)), self.lib)
        self.ec_cols = len(self.cont_ids) * self.cnt_channels
        if self.gatys:
            self.gn_shape = (self.B, self.L, 128, 128)
        else:
            self.gn_shape = (self.B, self.nb_channels, self.L, self.L)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.ast_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- setup ------------------------------------------------------------------------------
    def set_weights(self, w: EncoderWeights):
        """GatysNet.load_model (methods.py:79-84)."""
        if w.n_layers < self.n_layers:
            raise ValueError(f"need {self.n_layers} layers of weights, got {w.n_layers}")
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        arrs = [f(w.W0), f(w.b0), f(w.Wd[:self.n_layers]), f(w.bd[:self.n_layers]),
                f(w.Wr[:self.n_layers]), f(w.br[:self.n_layers])]
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_set_weights(self._h, *[_ptr(a) for a in arrs], _stream()), self.lib)

    def set_targets(self, Pc, Ps, lambd, gamma=0.0):
        """define_loss (methods.py:113-125): bake targets and loss weights."""
        Pc = self._as_f32(Pc, (self.B, self.T, self.ec_cols))
        Ps = self._as_f32(Ps, self.gn_shape)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_set_targets(self._h, _ptr(Pc), _ptr(Ps), float(lambd), float(gamma), _stream()), self.lib)

    def _as_f32(self, a, shape):
        if isinstance(a, torch.Tensor):
            a = a.to(torch.float32).contiguous()
            assert a.numel() == int(np.prod(shape)), (tuple(a.shape), shape)
            return a
        a = np.ascontiguousarray(a, dtype=np.float32)
        assert a.size == int(np.prod(shape)), (a.shape, shape)
        return a

    def _x(self, x):
        if not isinstance(x, torch.Tensor):
            x = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32), device=self.dev)
        x = x.to(device=self.dev, dtype=torch.float32).contiguous()
        assert x.numel() == self.B * self.T, (tuple(x.shape), self.B, self.T)
        return x

    # -- get_embeds (methods.py:86-95) --------------------------------------------------------
    def embeds(self, x, content=True, style=True):
        x = self._x(x)
        with torch.cuda.device(self.dev):
            Ec = torch.empty((self.B, self.T, self.ec_cols), dtype=torch.float32, device=self.dev) if content else None
            Gn = torch.empty(self.gn_shape, dtype=torch.float32, device=self.dev) if style else None
            _lib.check(self.lib.ast_embeds(self._h, _ptr(x), _ptr(Ec), _ptr(Gn), _stream()), self.lib)
        return Ec, Gn

    # -- the step -------------------------------------------------------------------------------
    def step(self, x, losses=None, grad=None):
        """Device in, device out, asynchronous on the current stream.  Returns (losses [B,4], grad [B,T])."""
        x = self._x(x)
        with torch.cuda.device(self.dev):
            if losses is None:
                losses = torch.empty((self.B, 4), dtype=torch.float32, device=self.dev)
            if grad is None:
                grad = torch.empty((self.B, self.T), dtype=torch.float32, device=self.dev)
            _lib.check(self.lib.ast_step(self._h, _ptr(x), _ptr(losses), _ptr(grad), _stream()), self.lib)
        return losses, grad

    def step_graph(self, x, losses, grad):
        """The same evaluation replayed from the library's captured CUDA graph (fixed device buffers)."""
        assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_step_graph(self._h, _ptr(x), _ptr(losses), _ptr(grad), _stream()), self.lib)
        return losses, grad

    def step_timed(self, x, losses, grad):
        """One step with CUDA events between phases; returns ms (encoder fwd, bwd, gram+heads, total)."""
        ms = (C.c_float * 4)()
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_step_timed(self._h, _ptr(x), _ptr(losses), _ptr(grad), _stream(), ms), self.lib)
        return list(ms)

    def step_host(self, x64: np.ndarray):
        """SciPy-facing evaluation: float64 host in, (losses [B,4] float64, grad [B,T] float64) out;
        H2D + captured CUDA graph + D2H inside (ast_step_host)."""
        x64 = np.ascontiguousarray(x64, dtype=np.float64)
        assert x64.size == self.B * self.T
        losses = np.empty((self.B, 4), np.float64)
        grad = np.empty((self.B, self.T), np.float64)
        _lib.check(self.lib.ast_step_host(self._h, _ptr(x64), _ptr(losses), _ptr(grad)), self.lib)
        return losses, grad

    # -- two-phase step for time sharding -------------------------------------------------------
    def forward_partial(self, x):
        x = self._x(x)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_step_forward(self._h, _ptr(x), _stream()), self.lib)

    def partials(self) -> torch.Tensor:
        """Zero-copy torch view of the raw Gram + content-SSE partial sums (all-reduce this)."""
        p, n = C.c_void_p(), C.c_int64()
        _lib.check(self.lib.ast_partials(self._h, C.byref(p), C.byref(n)), self.lib)
        return torch.as_tensor(_DevMem(p.value, n.value), device=self.dev)

    def backward(self, losses=None, grad=None):
        with torch.cuda.device(self.dev):
            if losses is None:
                losses = torch.empty((self.B, 4), dtype=torch.float32, device=self.dev)
            if grad is None:
                grad = torch.empty((self.B, self.T), dtype=torch.float32, device=self.dev)
            _lib.check(self.lib.ast_step_backward(self._h, _ptr(losses), _ptr(grad), _stream()), self.lib)
        return losses, grad

    # -- the same two phases stack by stack (time shards that exchange halo rows; include/ast_b200.h) ------------
    @property
    def segments(self) -> int:
        return int(self.lib.ast_segments(self._h))

    def forward_seg(self, x, seg: int):
        x = self._x(x)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_step_forward_seg(self._h, _ptr(x), int(seg), _stream()), self.lib)

    def backward_seg(self, seg: int, losses, grad):
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_step_backward_seg(self._h, int(seg), _ptr(losses), _ptr(grad), _stream()), self.lib)

    def segment_buffer(self, backward: bool, seg: int) -> torch.Tensor:
        """Zero-copy [B, T, 128] fp32 view of what neighbours exchange after segment `seg`: the residual stream at the
        segment's end (forward) or the running gradient (backward)."""
        p = C.c_void_p()
        _lib.check(self.lib.ast_segment_buffer(self._h, int(bool(backward)), int(seg), C.byref(p)), self.lib)
        return torch.as_tensor(_DevMem(p.value, self.B * self.T * 128), device=self.dev).view(self.B, self.T, 128)

    # -- peer-mapped halos (include/ast_b200.h: ast_p2p_*) ---------------------------------------------------------
    def p2p_export(self) -> bytes:
        buf = C.create_string_buffer(256)
        _lib.check(self.lib.ast_p2p_export(self._h, buf), self.lib)
        return buf.raw

    def p2p_attach(self, side: int, handles: bytes, peer_win_begin: int, peer_T: int):
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_p2p_attach(self._h, int(side), C.create_string_buffer(handles, 256), int(peer_win_begin), int(peer_T)), self.lib)

    def p2p_detach(self):
        _lib.check(self.lib.ast_p2p_detach(self._h), self.lib)

    def p2p_attach_local(self, side: int, peer: "StyleStep"):
        _lib.check(self.lib.ast_p2p_attach_local(self._h, int(side), peer._h), self.lib)

    def extract(self, layer: int) -> torch.Tensor:
        """cfg.extracts[layer] of the last forward (model.py:116); layer -1 = start-conv output."""
        out = torch.empty((self.B, self.T, 128), dtype=torch.float32, device=self.dev)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_get_extract(self._h, int(layer), _ptr(out), _stream()), self.lib)
        return out

    def masks(self):
        """ReLU sign masks of the last forward as bool arrays (hmask, zmask), each [n_layers, B, T, 128] (ast_get_masks)."""
        n = self.n_layers * self.B * self.T * 4
        hm = torch.empty(n, dtype=torch.int32, device=self.dev)
        zm = torch.empty(n, dtype=torch.int32, device=self.dev)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_get_masks(self._h, _ptr(hm), _ptr(zm), _stream()), self.lib)
        torch.cuda.synchronize(self.dev)
        unpack = lambda m: np.unpackbits(m.cpu().numpy().view(np.uint8), bitorder="little").astype(bool).reshape(
            self.n_layers, self.B, self.T, 128)
        return unpack(hm), unpack(zm)

    @property
    def uses_stack_kernels(self) -> bool:
        """True when the encoder forward / backward run as ONE cooperative kernel each (clip fits one tile per SM)."""
        return bool(self.lib.ast_uses_stack_kernels(self._h))

    def check(self):
        """ast_check: synchronise the current stream; raises AstError if a hand-off wait of the whole-stack kernels
        expired in any evaluation since the last check (those results are invalid) or CUDA reported an error."""
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_check(self._h, _stream()), self.lib)

    @property
    def launches_per_step(self) -> int:
        return int(self.lib.ast_launches_per_step(self._h))

    @property
    def workspace_bytes(self) -> int:
        return int(self.lib.ast_workspace_bytes(self._h))


class DeviceLBFGS:
    """Owner of one ast_opt: SciPy's L-BFGS-B loop of the reference (methods.py:133-137,164-181) as a state machine in
    device memory (csrc/lbfgs.cu).  `x`, `losses`, `grad` are zero-copy views of the machine's buffers: evaluate the
    step at `x` into (`losses`, `grad`), call `advance()`, repeat -- or let `minimize(step)` pump a StyleStep."""

    PHASE_DONE = 2
    END_REASONS = {0: "none", 1: "maxiter", 2: "pgtol", 3: "ftol", 4: "abnormal line search"}

    def __init__(self, T, B=1, m=10, maxiter=100, epochs=1, min_evals=50, maxls=20, max_log=0, keep_epoch_x=False,
                 factr=1e7, pgtol=1e-5, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device: the device-resident optimiser has no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.dev = torch.device("cuda", self.device)
        self.T, self.B, self.epochs, self.max_log, self.keep_epoch_x = int(T), int(B), int(epochs), int(max_log), bool(keep_epoch_x)
        cfg = _lib.AstOptConfig()
        cfg.T, cfg.B, cfg.m, cfg.maxiter, cfg.epochs, cfg.min_evals = self.T, self.B, int(m), int(maxiter), self.epochs, int(min_evals)
        cfg.maxls, cfg.max_log, cfg.keep_epoch_x, cfg.device = int(maxls), self.max_log, int(self.keep_epoch_x), self.device
        cfg.factr, cfg.pgtol = float(factr), float(pgtol)
        self._h = C.c_void_p()
        _lib.check(self.lib.ast_opt_create(C.byref(cfg), C.byref(self._h)), self.lib)
        px, pl, pg = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _lib.check(self.lib.ast_opt_buffers(self._h, C.byref(px), C.byref(pl), C.byref(pg)), self.lib)
        view = lambda p, n, shape: torch.as_tensor(_DevMem(p.value, n), device=self.dev).view(shape)
        self.x = view(px, self.B * self.T, (self.B, self.T))
        self.losses = view(pl, self.B * 4, (self.B, 4))
        self.grad = view(pg, self.B * self.T, (self.B, self.T))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.ast_opt_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self, x0):
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        assert x0.size == self.B * self.T
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_opt_reset(self._h, _ptr(x0), _stream()), self.lib)
            torch.cuda.current_stream().synchronize()       # x0 is pageable host memory

    def advance(self):
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_opt_advance(self._h, _stream()), self.lib)

    def status(self):
        """(int32 [B,8], float64 [B]): see ast_opt_status in include/ast_b200.h."""
        st = np.zeros((self.B, 8), np.int32)
        f = np.zeros(self.B, np.float64)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_opt_status(self._h, _ptr(st), _ptr(f), _stream()), self.lib)
        return st, f

    @property
    def done(self):
        return bool((self.status()[0][:, 0] == self.PHASE_DONE).all())

    def result(self):
        out = np.empty((self.B, self.T), np.float64)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_opt_result(self._h, _ptr(out), _stream()), self.lib)
        return out

    def epochs_info(self):
        """(evaluations [B,epochs], end reason [B,epochs], iterate per epoch [B,epochs,T] or None)."""
        nfev = np.zeros((self.B, self.epochs), np.int32)
        end = np.zeros((self.B, self.epochs), np.int32)
        xs = np.zeros((self.B, self.epochs, self.T), np.float32) if self.keep_epoch_x else None
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.ast_opt_epochs(self._h, _ptr(nfev), _ptr(end), _ptr(xs), _stream()), self.lib)
        return nfev, end, xs

    def log(self):
        out = np.zeros((self.B, max(self.max_log, 1), 4), np.float32)
        if self.max_log:
            with torch.cuda.device(self.dev):
                _lib.check(self.lib.ast_opt_log(self._h, _ptr(out), _stream()), self.lib)
        return out[:, :self.max_log]

    def minimize(self, step: "StyleStep", max_evals=0, chunk=4) -> int:
        """ast_minimize: [step, advance] graph replays until every pair has stopped.  Returns evaluations enqueued."""
        assert step.T == self.T and step.B == self.B and step.device == self.device
        n = C.c_int64()
        with torch.cuda.device(self.dev):
            torch.cuda.current_stream().synchronize()
            _lib.check(self.lib.ast_minimize(step._h, self._h, int(max_evals), int(chunk), C.byref(n)), self.lib)
        return int(n.value)
