"""CPU tests of the C-ABI boundary: the library builds, loads, exports every symbol that
include/ast_b200.h declares, and fails loudly (no CPU fallback) when no B200 is present."""
import ctypes as C
import re
from pathlib import Path

import pytest
import torch

from audio_style_transfer_b200 import _lib

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    _lib.build()
    return _lib.load()


def header_symbols():
    text = (ROOT / "include" / "ast_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ast_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert header_symbols() == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_version_string(lib):
    assert b"sm_100a" in lib.ast_version()


def test_config_struct_layout_matches_header():
    # 3 + 1 + 32 + 1 + 32 + 4 int32 = 73 int32 -> padded to 296 bytes, then 4 int64, then 3 int32 (+ 4 bytes of tail padding)
    assert C.sizeof(_lib.AstConfig) == 296 + 32 + 16
    assert _lib.AstConfig.halo_exchange.offset == 296 + 32 + 8
    assert _lib.AstConfig.total_T.offset == 296


def test_invalid_arguments_are_rejected_before_touching_cuda(lib):
    h = C.c_void_p()
    cfg = _lib.AstConfig()
    cfg.T, cfg.B, cfg.n_layers = 1000, 1, 10           # T not a multiple of 128
    assert lib.ast_create(C.byref(cfg), C.byref(h)) == -1
    assert b"multiple of 128" in lib.ast_last_error()
    assert lib.ast_create(None, C.byref(h)) == -1


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(lib):
    h = C.c_void_p()
    cfg = _lib.AstConfig()
    cfg.T, cfg.B, cfg.n_layers = 1024, 1, 10
    cfg.n_cont, cfg.n_style = 1, 10
    cfg.cont_ids[0] = 9
    for i in range(10):
        cfg.style_ids[i] = i
    cfg.nb_channels = cfg.cnt_channels = 128
    assert lib.ast_create(C.byref(cfg), C.byref(h)) == -2          # AST_ECUDA
    assert b"no CPU fallback" in lib.ast_last_error() or b"CUDA" in lib.ast_last_error()
    from audio_style_transfer_b200.step import StyleStep
    with pytest.raises(RuntimeError):
        StyleStep(1024, [9], list(range(10)))


def test_sass_is_sm100a():
    """The shipped binary carries sm_100a code only."""
    import subprocess
    _lib.build()
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", str(_lib.SO)], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out
