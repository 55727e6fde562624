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


# ---- gamma term: STFT regulariser (methods.py:121-123, oracle/stft_reg.py) --------------------------------
def _reg_inputs(kind, T=4096, seed=3):
    rng = np.random.default_rng(seed)
    x = rng.normal(0, 40, T)
    if kind == "int":              # the first evaluation of a run sees integer mu-law codes, zeros included
        x = np.floor(x)
    x[5] = 0.0
    x[7] = -0.5                    # (x + 0.5) = 0: the reference's sign() is 0 there
    return x


@pytest.mark.parametrize("kind", ["float", "int"])
def test_regulariser_gradient_matches_torch_autograd_of_literal_graph(kind):
    from oracle import stft_reg
    x = _reg_inputs(kind)
    reg, g, raw = stft_reg.stft_l1(x)
    reg2, g2 = stft_reg.torch_literal(x)
    assert abs(reg - reg2) <= 1e-13 * abs(reg2)
    assert np.linalg.norm(g - g2) <= 1e-12 * np.linalg.norm(g2)
    assert stft_reg.n_frames(4096) == 7 and stft_reg.n_frames(16384) == 31 and stft_reg.n_frames(1023) == 0


def test_regulariser_finite_differences_and_frame_subsets():
    from oracle import stft_reg
    x = _reg_inputs("float")
    reg, g, _ = stft_reg.stft_l1(x)
    rng = np.random.default_rng(4)
    for _ in range(3):
        v = rng.standard_normal(x.shape[0])
        eps = 1e-6
        fd = (stft_reg.stft_l1(x + eps * v)[0] - stft_reg.stft_l1(x - eps * v)[0]) / (2 * eps)
        assert abs(fd - g @ v) <= 2e-4 * max(abs(fd), 1e-6), (fd, g @ v)     # |Re|, |Im| kinks: piecewise smooth only
    F = stft_reg.n_frames(x.shape[0])
    a = stft_reg.stft_l1(x, frames=range(0, 3), total_frames=F)
    b = stft_reg.stft_l1(x, frames=range(3, F), total_frames=F)
    assert abs(a[0] + b[0] - reg) <= 1e-14 and np.allclose(a[1] + b[1], g, rtol=0, atol=1e-18)


def test_reference_abs_and_inv_mu_law_definitions():
    from oracle import stft_reg
    from audio_style_transfer_b200.mu_law import inv_mu_law_numpy
    v = np.array([-2.0, -1e-13, 0.0, 5e-13, 1e-12, 3.0])
    assert np.allclose(stft_reg.ref_abs(v), [2.0 + 1e-12, 1e-13 + 1e-12, 1e-12, 1e-12, 1e-12, 3.0], rtol=1e-15)
    assert list(stft_reg.ref_abs_grad(v)) == [-1.0, -1.0, 0.0, 0.0, 1.0, 1.0]
    x = np.arange(-128, 128).astype(np.float64)
    a, d = stft_reg.inv_mu_law(x)
    assert np.allclose(a, inv_mu_law_numpy(x), rtol=1e-5, atol=1e-12)         # utils.py:85-90, the fp32 numpy twin
    assert a[128] == 0.0 and d[128] == 1.0                                     # x == 0 passes x through
    assert np.all(d > 0)


def test_step_with_gamma_adds_the_regulariser():
    from oracle import stft_reg
    case = cases.build_inputs("t1024_ours_s0")
    Pc, Ps = cases.oracle_targets(case)
    l0, dx0 = cases.oracle_step(case, Pc, Ps)
    lg, dxg = step_np.step(case["x"], case["w"], Pc, Ps, case["cont"], case["style"], case["gatys"], case["lambd"], 2.5,
                           case["nb"], case["cnt"])
    reg, g, _ = stft_reg.stft_l1(case["x"])
    assert l0[3] == 0.0 and abs(lg[3] - reg) == 0.0
    assert abs(lg[0] - (l0[0] + 2.5 * reg)) <= 1e-12 * abs(lg[0])
    assert np.allclose(dxg - dx0, 2.5 * g, rtol=0, atol=1e-15)


def test_regulariser_golden(golden_dir):
    from oracle import stft_reg
    gold = np.load(golden_dir / "reg_t16384.npz")
    x = cases.build_inputs("t16384_cfg0")["x"]
    assert float(x.sum()) == float(gold["x_sum"])
    reg, g, raw = stft_reg.stft_l1(x)
    assert abs(reg - float(gold["reg"])) <= 1e-12 * abs(reg)
    assert np.allclose(g[::16], gold["g"], rtol=1e-10, atol=1e-18)
    assert abs(np.linalg.norm(g) - float(gold["g_norm"])) <= 1e-10 * float(gold["g_norm"])
