"""ast_stft_reg_noalloc (include/ast_b200.h): the gamma term of one clip through an entry point that allocates nothing
(scratch in a file-scope device array, csrc/stft_reg.cu).  No counterpart in the reference.
Checked through the C ABI against the existing entry points (ast_set_targets(gamma) + ast_step) at the
tolerances tests/test_gpu_parity.py::test_regulariser_matches_oracle uses (value 1e-5, gradient 1e-4), on two calls in a
row (the scratch keeps the previous call's contents), and with torch's allocator watched: the calls must not allocate.

Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest
import torch

from audio_style_transfer_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def reference_through_existing_entry_points(x, gamma=50.0):
    """reg = losses[3] and d reg / dx = (grad(gamma) - grad(0)) / gamma of ast_step on a small context."""
    from audio_style_transfer_b200.step import StyleStep
    T, NL = x.shape[1], 2
    st = StyleStep(T, [NL - 1], list(range(NL)), precision="fp32")
    st.set_weights(synth.make_weights(NL, 0, 0.1))
    rng = np.random.default_rng(3)
    Pc = rng.standard_normal((1, T, 128)); Ps = rng.standard_normal((1, 128, NL, NL)) * 0.1
    st.set_targets(Pc, Ps, 100.0, 0.0)
    lo0, g0 = (t.cpu().numpy()[0].astype(np.float64) for t in st.step(x))
    st.set_targets(Pc, Ps, 100.0, gamma)
    lo1, g1 = (t.cpu().numpy()[0].astype(np.float64) for t in st.step(x))
    st.close()
    return lo1[3], (g1 - g0) / gamma


def call_noalloc(lib, x_dev, reg, grad):
    rc = lib.ast_stft_reg_noalloc(C.c_void_p(x_dev.data_ptr()), x_dev.numel(), C.c_void_p(reg.data_ptr()),
                                  C.c_void_p(grad.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, lib)


@pytest.mark.parametrize("T", [4096, 16384])
def test_noalloc_matches_existing_entry_point_twice(T):
    lib = _lib.load()
    rng = np.random.default_rng(11)
    xs = [rng.normal(0, 40, (1, T)), np.floor(rng.normal(0, 25, (1, T)))]       # the second one hits x == 0 pass-throughs
    xs[1][0, ::97] = 0.0
    refs = [reference_through_existing_entry_points(x) for x in xs]
    dev = [torch.tensor(x[0], dtype=torch.float32, device="cuda") for x in xs]
    reg = torch.full((1,), float("nan"), device="cuda")
    grad = torch.full((T,), float("nan"), device="cuda")                          # must be overwritten, not accumulated into
    call_noalloc(lib, dev[0], reg, grad)                                          # first call on this device: one-time set-up
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"], torch.cuda.memory_allocated()
    free_before = torch.cuda.mem_get_info()[0]
    # two calls in a row on different inputs, then the first input again: the shared scratch carries nothing over
    for i in (0, 1, 0):
        call_noalloc(lib, dev[i], reg, grad)
        torch.cuda.synchronize()
        r_ref, g_ref = refs[i]
        r, g = float(reg.cpu()[0]), grad.cpu().numpy().astype(np.float64)
        assert abs(r - r_ref) <= 1e-5 * abs(r_ref), (i, r, r_ref)
        assert rel(g, g_ref) <= 1e-4, (i, rel(g, g_ref))
    assert (torch.cuda.memory_stats()["allocation.all.allocated"], torch.cuda.memory_allocated()) == before
    assert torch.cuda.mem_get_info()[0] == free_before                            # no cudaMalloc behind torch's back either


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    lib = _lib.load()
    x = torch.zeros(131072, device="cuda")
    reg = torch.zeros(1, device="cuda"); grad = torch.zeros(131072, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    assert lib.ast_stft_reg_noalloc(p(x), 131072, p(reg), p(grad), s) == -1
    assert b"AST_NOALLOC_MAX_T" in lib.ast_last_error()
    assert lib.ast_stft_reg_noalloc(p(x), 1000, p(reg), p(grad), s) == -1       # not a multiple of 128
    assert lib.ast_stft_reg_noalloc(p(x), 896, p(reg), p(grad), s) == -1        # shorter than one STFT frame
    assert lib.ast_stft_reg_noalloc(None, 4096, p(reg), p(grad), s) == -1
    assert lib.ast_stft_reg_noalloc(p(x), 65536, p(reg), p(grad), s) == 0       # the maximum itself is fine
    torch.cuda.synchronize()
