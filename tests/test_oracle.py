"""CPU tests of the oracle itself: it is pinned by three independent witnesses (no reference-made
fixtures exist -- see oracle/step_np.py header, PARITY UNPINNED) and against the committed goldens."""
import numpy as np
import pytest
import torch

from audio_style_transfer_b200 import synth
from oracle import cases, masked_literal, step_np
from oracle.torch_step import TorchStep


@pytest.mark.parametrize("d", [1, 2, 4, 64, 512])
@pytest.mark.parametrize("K,cin", [(3, 1), (3, 8), (1, 8)])
def test_conv_closed_form_equals_literal_masked_py(d, K, cin):
    """Integer data: the closed form must equal the reshape/transpose/SAME route bit for bit."""
    rng = np.random.default_rng(d * 10 + K + cin)
    T = 2048
    x = rng.integers(-8, 9, (1, T, cin)).astype(np.float64)
    W = rng.integers(-3, 4, (K, cin, 16)).astype(np.float64)
    b = rng.integers(-3, 4, 16).astype(np.float64)
    lit = masked_literal.conv1d(x, W, b, dilation=d, causal=(K == 1))[0]
    if K == 3:
        ref = step_np.conv3(x[0], W, b, d)
    else:
        ref = x[0] @ W[0] + b
    assert np.array_equal(lit, ref)


def test_encoder_closed_form_equals_literal_bit_exact():
    wi = synth.make_int_weights(12, 5)
    x = np.random.default_rng(7).integers(-128, 128, 4096).astype(np.float64)
    H1, _ = step_np.encoder_forward(x, wi, 12)
    H2 = masked_literal.encoder_forward(x, wi, 12)
    assert all(np.array_equal(a, b) for a, b in zip(H1, H2))


def test_conv3_adjoint():
    rng = np.random.default_rng(0)
    a, dy = rng.standard_normal((256, 8)), rng.standard_normal((256, 5))
    W = rng.standard_normal((3, 8, 5))
    for d in (1, 4, 64, 128, 256):
        lhs = np.sum(step_np.conv3(a, W, np.zeros(5), d) * dy)
        rhs = np.sum(a * step_np.conv3_T(dy, W, d))
        assert abs(lhs - rhs) < 1e-9 * max(1.0, abs(lhs))


@pytest.mark.parametrize("name", ["t1024_ours_s0", "t1024_ours_sub", "t1024_ours_ids", "t2048_gatys_s1"])
def test_backward_matches_torch_autograd_fp64(name):
    case = cases.build_inputs(name)
    Pc, Ps = cases.oracle_targets(case)
    losses, dx = cases.oracle_step(case, Pc, Ps)
    ts = TorchStep(case["w"], case["cont"], case["style"], case["gatys"], case["lambd"], case["nb"], case["cnt"],
                   dtype=torch.float64)
    l2, dx2 = ts.step(torch.tensor(case["x"])[None], torch.tensor(Pc)[None], torch.tensor(Ps)[None])
    assert np.allclose(l2[0].numpy(), losses[:3], rtol=1e-12)
    assert np.linalg.norm(dx - dx2[0].numpy()) <= 1e-12 * np.linalg.norm(dx)


def test_backward_matches_finite_differences():
    case = cases.build_inputs("t1024_ours_s0")
    Pc, Ps = cases.oracle_targets(case)
    _, dx = cases.oracle_step(case, Pc, Ps)
    rng = np.random.default_rng(11)
    f = lambda x: step_np.step(x, case["w"], Pc, Ps, case["cont"], case["style"], case["gatys"], case["lambd"])[0][0]
    for _ in range(3):
        v = rng.standard_normal(case["T"])
        eps = 1e-5    # small enough that no ReLU kink is crossed (the loss is piecewise smooth)
        fd = (f(case["x"] + eps * v) - f(case["x"] - eps * v)) / (2 * eps)
        assert abs(fd - dx @ v) <= 1e-5 * max(abs(fd), 1e-3), (fd, dx @ v)


def test_l2_normalize_clamp_branch():
    G = np.zeros((2, 3, 3)); G[1] = np.eye(3)
    Gn = step_np.l2_normalize(G)
    assert np.all(Gn[0] == 0) and np.allclose((Gn[1] ** 2).sum(), 1.0)


def _check_golden(name, golden_dir):
    gold = np.load(golden_dir / f"{name}.npz")
    case = cases.build_inputs(name)
    assert str(gold["w_digest"]) == case["w"].digest(), "seeded weight generator drifted"
    assert np.isclose(gold["x_sum"], case["x"].sum(), rtol=1e-12)
    Pc, Ps = cases.oracle_targets(case)
    losses, dx, aux = cases.oracle_step(case, Pc, Ps, return_aux=True)
    s = cases.summary(case, Pc, Ps, losses, dx, aux)
    assert np.allclose(s["losses"], gold["losses"], rtol=1e-9)
    assert np.linalg.norm(s["dx"] - gold["dx"]) <= 1e-8 * np.linalg.norm(gold["dx"])
    assert np.isclose(s["dx_norm"], gold["dx_norm"], rtol=1e-9)
    assert np.allclose(s["H_sq"], gold["H_sq"], rtol=1e-10)
    assert np.isclose(s["Gn_sq"], gold["Gn_sq"], rtol=1e-10)


@pytest.mark.parametrize("name", [n for n in cases.CASES if not n.startswith("t16384")])
def test_oracle_reproduces_golden(name, golden_dir):
    _check_golden(name, golden_dir)


def test_oracle_reproduces_golden_baseline_size(golden_dir):
    _check_golden("t16384_cfg0", golden_dir)


def test_fp32_oracle_noise_floor():
    """What an fp32 CPU run of the same step loses against fp64: the yardstick for the 1e-4 tier."""
    case = cases.build_inputs("t1024_ours_s0")
    Pc, Ps = cases.oracle_targets(case)
    l64, dx64 = cases.oracle_step(case, Pc, Ps)
    l32, dx32 = step_np.step(case["x"].astype(np.float32), case["w"], Pc.astype(np.float32), Ps.astype(np.float32),
                             case["cont"], case["style"], case["gatys"], case["lambd"])
    assert abs(l32[0] - l64[0]) <= 1e-4 * abs(l64[0])
    assert np.linalg.norm(dx32 - dx64) <= 1e-3 * np.linalg.norm(dx64)


def test_resolve_style_ids_and_layers_needed():
    assert synth.resolve_style_ids(0) == list(range(10))
    assert synth.resolve_style_ids(None) == list(range(30))
    assert synth.resolve_style_ids(1, [3, 4]) == [3, 4]
    assert synth.layers_needed([25], list(range(10))) == 26
    assert synth.layers_needed([30], [0]) == 30
    with pytest.raises(ValueError):
        synth.layers_needed([31], [0])
