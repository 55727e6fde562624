"""ctypes binding of libast_b200.so (include/ast_b200.h) and its in-tree nvcc build.

The product path has NO CPU fallback: if the shared library is missing or no sm_100 GPU is present,
`load()` / `Context()` raise.  PyTorch is used by callers only to own device memory and streams;
nothing torch-typed crosses this boundary (plain pointers and sizes).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
SO = PKG / "libast_b200.so"
SOURCES = ["csrc/api.cu", "csrc/kernels_f32.cu", "csrc/kernels_tc.cu", "csrc/gram_tc.cu", "csrc/stft_reg.cu", "csrc/lbfgs.cu", "csrc/gram_ours_tc.cu", "csrc/gram_ours_tma.cu"]
HEADERS = ["csrc/common.cuh", "csrc/tc.cuh", "csrc/ptx.cuh", "csrc/stack_fwd.cuh", "csrc/stack_bwd.cuh", "csrc/stack_bwd_split.cuh", "../include/ast_b200.h"]
SO_EXP = PKG / "libast_b200_exp.so"      # the same sources with -DAST_EXPERIMENTS: timeline stamps, fault injection, A/B switches

AST_PREC_FP32 = 0
AST_PREC_BF16 = 1
AST_PREC_TC3 = 2
AST_MAX_IDS = 32
AST_HALO_ROWS = 1024
ERRORS = {-1: "AST_EINVAL", -2: "AST_ECUDA", -3: "AST_EUNSUPPORTED", -4: "AST_ESTATE"}

SYMBOLS = [
    "ast_create", "ast_destroy", "ast_last_error", "ast_version", "ast_workspace_bytes",
    "ast_set_weights", "ast_set_targets", "ast_embeds", "ast_step", "ast_step_graph", "ast_step_host",
    "ast_step_forward", "ast_partials", "ast_step_backward", "ast_get_extract", "ast_get_masks",
    "ast_segments", "ast_step_forward_seg", "ast_step_backward_seg", "ast_segment_buffer",
    "ast_p2p_export", "ast_p2p_attach", "ast_p2p_attach_local", "ast_p2p_detach",
    "ast_launches_per_step", "ast_step_timed", "ast_check", "ast_uses_stack_kernels", "ast_stft_reg_noalloc",
    "ast_opt_create", "ast_opt_destroy", "ast_opt_workspace_bytes", "ast_opt_reset", "ast_opt_buffers", "ast_opt_advance",
    "ast_opt_status", "ast_opt_result", "ast_opt_epochs", "ast_opt_log", "ast_minimize",
]


class AstConfig(C.Structure):
    _fields_ = [
        ("T", C.c_int32), ("B", C.c_int32), ("n_layers", C.c_int32),
        ("n_cont", C.c_int32), ("cont_ids", C.c_int32 * AST_MAX_IDS),
        ("n_style", C.c_int32), ("style_ids", C.c_int32 * AST_MAX_IDS),
        ("gatys", C.c_int32), ("nb_channels", C.c_int32), ("cnt_channels", C.c_int32),
        ("precision", C.c_int32),
        ("total_T", C.c_int64), ("win_begin", C.c_int64), ("own_begin", C.c_int64), ("own_end", C.c_int64),
        ("device", C.c_int32), ("kernel_path", C.c_int32), ("halo_exchange", C.c_int32),
    ]


class AstOptConfig(C.Structure):
    _fields_ = [
        ("T", C.c_int32), ("B", C.c_int32), ("m", C.c_int32), ("maxiter", C.c_int32), ("epochs", C.c_int32),
        ("min_evals", C.c_int32), ("maxls", C.c_int32), ("max_log", C.c_int32), ("keep_epoch_x", C.c_int32),
        ("device", C.c_int32), ("factr", C.c_double), ("pgtol", C.c_double),
    ]


class AstError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERRORS.get(code, code)}: {msg}")
        self.code = code


def nvcc_command(out: Path = SO, experiments: bool = False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    return [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "--threads", "0",
            "-shared", "-Xcompiler", "-fPIC"] + (["-DAST_EXPERIMENTS"] if experiments else []) + \
           ["-o", str(out)] + [str(PKG / s) for s in SOURCES]


def needs_build(so: Path = SO) -> bool:
    if not so.exists():
        return True
    t = so.stat().st_mtime
    return any((PKG / p).resolve().stat().st_mtime > t for p in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every CUDA source for sm_100a into the in-tree libast_b200.so (nvcc cross-compiles without a GPU),
    and the same sources with -DAST_EXPERIMENTS into libast_b200_exp.so (timeline stamps, fault injection and the
    A/B switches: test / profiling infrastructure, never loaded by the product path)."""
    for so, exp in ((SO, False), (SO_EXP, True)):
        if force or needs_build(so):
            cmd = nvcc_command(so, exp)
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    return SO


_libs = {}


def load(experiments: bool = False) -> C.CDLL:
    """dlopen the in-tree library and declare every prototype of include/ast_b200.h.  experiments=True loads the
    -DAST_EXPERIMENTS build instead (tests/test_gpu_loud.py, scripts/timeline_stack.py)."""
    if experiments in _libs:
        return _libs[experiments]
    so = SO_EXP if experiments else SO
    if not so.exists():
        raise ImportError(f"{so} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback for the step)")
    lib = C.CDLL(str(so))
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    lib.ast_create.argtypes = [C.POINTER(AstConfig), C.POINTER(vp)]
    lib.ast_destroy.argtypes = [vp]
    lib.ast_last_error.restype = C.c_char_p
    lib.ast_version.restype = C.c_char_p
    lib.ast_workspace_bytes.argtypes = [vp]
    lib.ast_workspace_bytes.restype = i64
    lib.ast_set_weights.argtypes = [vp] + [vp] * 6 + [vp]
    lib.ast_set_targets.argtypes = [vp, vp, vp, dbl, dbl, vp]
    lib.ast_embeds.argtypes = [vp, vp, vp, vp, vp]
    lib.ast_step.argtypes = [vp, vp, vp, vp, vp]
    lib.ast_step_graph.argtypes = [vp, vp, vp, vp, vp]
    lib.ast_step_host.argtypes = [vp, vp, vp, vp]
    lib.ast_step_forward.argtypes = [vp, vp, vp]
    lib.ast_partials.argtypes = [vp, C.POINTER(vp), C.POINTER(i64)]
    lib.ast_step_backward.argtypes = [vp, vp, vp, vp]
    lib.ast_get_extract.argtypes = [vp, i32, vp, vp]
    lib.ast_get_masks.argtypes = [vp, vp, vp, vp]
    lib.ast_segments.argtypes = [vp]
    lib.ast_step_forward_seg.argtypes = [vp, vp, i32, vp]
    lib.ast_step_backward_seg.argtypes = [vp, i32, vp, vp, vp]
    lib.ast_segment_buffer.argtypes = [vp, i32, i32, C.POINTER(vp)]
    lib.ast_p2p_export.argtypes = [vp, vp]
    lib.ast_p2p_attach.argtypes = [vp, i32, vp, i64, i32]
    lib.ast_p2p_attach_local.argtypes = [vp, i32, vp]
    lib.ast_p2p_detach.argtypes = [vp]
    lib.ast_launches_per_step.argtypes = [vp]
    lib.ast_step_timed.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.ast_check.argtypes = [vp, vp]
    lib.ast_stft_reg_noalloc.argtypes = [vp, i32, vp, vp, vp]
    lib.ast_uses_stack_kernels.argtypes = [vp]
    if experiments:
        lib.ast_debug_timeline.argtypes = [vp]
        lib.ast_debug_tl_layer.argtypes = [i32]
        lib.ast_debug_skip_publish.argtypes = [i32]
        lib.ast_debug_group_tiles.argtypes = [i32]
    lib.ast_opt_create.argtypes = [C.POINTER(AstOptConfig), C.POINTER(vp)]
    lib.ast_opt_destroy.argtypes = [vp]
    lib.ast_opt_workspace_bytes.argtypes = [vp]
    lib.ast_opt_workspace_bytes.restype = i64
    lib.ast_opt_reset.argtypes = [vp, vp, vp]
    lib.ast_opt_buffers.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.ast_opt_advance.argtypes = [vp, vp]
    lib.ast_opt_status.argtypes = [vp, vp, vp, vp]
    lib.ast_opt_result.argtypes = [vp, vp, vp]
    lib.ast_opt_epochs.argtypes = [vp, vp, vp, vp, vp]
    lib.ast_opt_log.argtypes = [vp, vp, vp]
    lib.ast_minimize.argtypes = [vp, vp, i64, i32, C.POINTER(i64)]
    for name in SYMBOLS:
        f = getattr(lib, name)
        if f.restype is C.c_int:      # default restype; the int-returning entry points
            f.restype = C.c_int
    _libs[experiments] = lib
    return lib


def check(code: int, lib=None):
    """Raise AstError with the message of the library that returned `code` (the message is thread-local per .so)."""
    if code != 0:
        raise AstError(code, (lib or load()).ast_last_error().decode())
