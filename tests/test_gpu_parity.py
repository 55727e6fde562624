"""GPU parity tests proper: the CUDA path, called through the C ABI, against the fp64 oracle on the
same seeded inputs.  Tolerances are BASELINE.json's: rel-L2 <= 1e-4 on loss and gradient in the fp32
tier, <= 1e-2 in the bf16 tensor-core tier; integer indexing bit-exact."""
import numpy as np
import pytest
import torch

from audio_style_transfer_b200 import synth
from oracle import cases, step_np

pytestmark = pytest.mark.gpu

# tc3 = the accurate tensor-core tier (AST_PREC_TC3: tcgen05 convs with hi/lo bf16 operands both ways, fp32 heads).  Its
# forward is the bf16 tier's (features ~1e-5..2e-5 from fp64: the fp32 accumulation inside the tensor core, not the operand
# precision, is the floor -- fp16 hi/lo operands measured the same), its backward is 5-10x closer: <= 6e-4 with the device's
# masks at every case (bf16 tier: 1.6e-3..5.6e-3).  It does not reach the 1e-4 of the CUDA-core fp32 tier.
TOL = {"fp32": dict(loss=1e-4, grad=1e-4, feat=2e-5), "bf16": dict(loss=1e-2, grad=1e-2, feat=2e-3),
       "tc3": dict(loss=1e-4, grad=1e-3, feat=2e-4)}
PRECS = ["fp32", "bf16", "tc3"]


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def fp32_floor(case, Pc, Ps, dx64):
    """rel-L2 distance between an fp32 CPU run of the same step (what the reference's fp32 TensorFlow
    graph computes) and fp64 truth.  The 30-layer ReLU stack amplifies rounding (sign-mask flips), so
    for deep cases two correct fp32 implementations differ by more than 1e-4; the tolerance is then
    a small multiple of this measured floor instead ("accuracy tiers")."""
    _, dx32 = step_np.step(case["x"].astype(np.float32), case["w"], Pc.astype(np.float32), Ps.astype(np.float32),
                           case["cont"], case["style"], case["gatys"], case["lambd"], 0.0, case["nb"], case["cnt"])
    return rel(dx32, dx64)


def grad_tol(prec, floor):
    """Bound on the RAW gradient distance (each side with its own ReLU masks).  bf16 tier: BASELINE.json's flat 1e-2.
    fp32 tier: 1e-4, or ten times the distance of an fp32 CPU run of the same step from fp64 where that is larger.  The
    raw figure of a 20..30-layer case is decided by a handful of ReLU sign bits that two roundings of the same forward
    legitimately disagree on (measured on B200, t16384_cfg0: ONE flipped bit near t = 11328 is 3.0e-4 of the gradient,
    everything else 1.5e-7), so it is a loose secondary bound; the strict statement is check_with_device_masks."""
    return TOL["bf16"]["grad"] if prec in ("bf16", "tc3") else max(TOL["fp32"]["grad"], 10.0 * floor)


# a sign bit may differ from the oracle's only where the fp64 value is within rounding of zero: |v| <= FLIP_EPS * rms(v)
FLIP_EPS = {"fp32": 2e-5, "bf16": 2e-2, "tc3": 2e-2}


def check_with_device_masks(st, case, Pc, Ps, aux, g, prec, b=0):
    """The strict gradient statement: BASELINE.json's tolerance (1e-4 fp32 tier, 1e-2 bf16 tier), no cushion, against the
    oracle's backward evaluated with the DEVICE's ReLU sign masks (ast_get_masks), plus: every sign bit that differs from
    the oracle's own sits on a value within rounding of zero.  Returns (masked distance, differing bits)."""
    hm, zm = st.masks()
    hm, zm = hm[:, b], zm[:, b]
    nflip = 0
    for dev, ref in ((hm, aux["H"]), (zm, aux["Z"])):
        for l in range(case["NL"]):
            diff = dev[l] != (ref[l] > 0)
            n = int(diff.sum())
            if n:
                nflip += n
                worst = float(np.abs(ref[l][diff]).max()) / float(np.sqrt(np.mean(ref[l] ** 2)))
                assert worst <= FLIP_EPS[prec], (l, n, worst)
    _, dxm = cases.oracle_step(case, Pc, Ps, masks=(hm, zm), fwd=(aux["H"], aux["Z"]))
    e = rel(g, dxm)
    assert e <= TOL[prec]["grad"], (e, nflip)
    return e, nflip


def make_step(case, prec, B=1, window=None, T=None, kernel_path="auto"):
    from audio_style_transfer_b200.step import StyleStep
    st = StyleStep(T or case["T"], case["cont"], case["style"], case["gatys"], case["nb"], case["cnt"], B=B,
                   precision=prec, window=window, kernel_path=kernel_path)
    st.set_weights(case["w"])
    return st


SMALL = ["t1024_ours_s0", "t1024_ours_sub", "t1024_ours_ids", "t2048_gatys_s1", "t4096_cfg0", "t4096_cfg1",
         "t4096_alias30", "t2048_ours_all30", "t1024_ours_l17"]


@pytest.mark.parametrize("prec", PRECS)
@pytest.mark.parametrize("name", SMALL)
def test_step_matches_oracle(name, prec):
    case = cases.build_inputs(name)
    Pc, Ps = cases.oracle_targets(case)
    losses, dx, aux = cases.oracle_step(case, Pc, Ps, return_aux=True)
    st = make_step(case, prec)
    # forward features first (get_embeds, methods.py:86-95)
    Ec, Gn = st.embeds(case["x"][None])
    tol = TOL[prec]
    assert rel(Ec[0].cpu().numpy(), aux["Ec"]) <= tol["feat"]
    assert rel(Gn[0].cpu().numpy(), aux["Gn"][:Gn.shape[1]]) <= tol["feat"]
    for l in (-1, 0, case["NL"] - 1):
        assert rel(st.extract(l)[0].cpu().numpy(), aux["H"][l + 1]) <= tol["feat"], l
    st.set_targets(Pc[None], Ps[None], case["lambd"])
    lo, g = st.step(case["x"][None])
    lo, g = lo.cpu().numpy()[0], g.cpu().numpy()[0]
    assert abs(lo[0] - losses[0]) <= tol["loss"] * abs(losses[0])
    assert abs(lo[2] - losses[2]) <= tol["loss"] * abs(losses[2])
    # content loss is a difference of nearly equal features (x = content + noise): the absolute error
    # of the features shows up relative to a small number, so it is bounded relative to the total
    assert abs(lo[1] - losses[1]) <= tol["loss"] * max(abs(losses[1]), abs(losses[0]))
    assert lo[3] == 0.0
    floor = fp32_floor(case, Pc, Ps, dx)
    em, nflip = check_with_device_masks(st, case, Pc, Ps, aux, g, prec)
    print(f"{name}[{prec}]: grad rel-L2 {em:.2e} with the device's masks ({nflip} sign bits differ), raw {rel(g, dx):.2e} "
          f"(fp32 CPU floor {floor:.2e})")
    assert rel(g, dx) <= grad_tol(prec, floor), (rel(g, dx), floor)
    st.close()


@pytest.mark.parametrize("prec", PRECS)
@pytest.mark.parametrize("name", ["t16384_cfg0", "t16384_cfg1", "t16384_cli"])
def test_baseline_size_matches_oracle_and_golden(name, prec, golden_dir):
    """BASELINE.json configs[0]/[1] sizes and the methods.py CLI defaults (all 30 layers channel-wise: cfg4's head) at
    16384 samples: oracle computed here (tens of seconds) and cross-checked with the committed golden vectors, then
    the CUDA step against both."""
    case = cases.build_inputs(name)
    Pc, Ps = cases.oracle_targets(case)
    losses, dx, aux = cases.oracle_step(case, Pc, Ps, return_aux=True)
    gold = np.load(golden_dir / f"{name}.npz")
    assert np.allclose(losses[:3], gold["losses"], rtol=1e-9)
    assert rel(dx[::int(gold["dx_stride"])], gold["dx"]) <= 1e-8
    st = make_step(case, prec)
    st.set_targets(Pc[None], Ps[None], case["lambd"])
    lo, g = st.step(case["x"][None])
    lo, g = lo.cpu().numpy()[0], g.cpu().numpy()[0]
    tol = TOL[prec]
    assert abs(lo[0] - losses[0]) <= tol["loss"] * abs(losses[0])
    floor = fp32_floor(case, Pc, Ps, dx)
    em, nflip = check_with_device_masks(st, case, Pc, Ps, aux, g, prec)
    print(f"{name}[{prec}]: loss rel {abs(lo[0] - losses[0]) / abs(losses[0]):.2e} grad rel-L2 {em:.2e} with the device's "
          f"masks ({nflip} sign bits differ), raw {rel(g, dx):.2e} (fp32 CPU floor {floor:.2e})")
    assert rel(g, dx) <= grad_tol(prec, floor), (rel(g, dx), floor)
    # the SciPy-facing host entry point returns the same numbers (float64 each way, CUDA graph inside)
    lh, gh = st.step_host(case["x"])
    assert np.allclose(lh[0], lo, rtol=1e-6, atol=1e-7)
    assert rel(gh[0], g) <= 1e-6
    lh2, gh2 = st.step_host(case["x"])      # graph replay is deterministic
    assert np.array_equal(gh, gh2) and np.array_equal(lh, lh2)
    st.close()


@pytest.mark.parametrize("prec", PRECS)
def test_index_probe_bit_exact(prec):
    """Integer weights and inputs: every product and sum is exact in fp32 (and in the bf16 hi/lo split),
    so tap offsets t+-d, zero padding at both clip ends and tile seams must reproduce the oracle
    EXACTLY at every layer."""
    NL, T = 12, 2048
    w = synth.make_int_weights(NL, 5)
    x = np.random.default_rng(7).integers(-3, 4, T).astype(np.float64)
    H, _ = step_np.encoder_forward(x, w, NL)
    from audio_style_transfer_b200.step import StyleStep
    st = StyleStep(T, [NL - 1], list(range(NL)), precision=prec)
    st.set_weights(w)
    st.embeds(x[None], style=False)
    # Every value (and every conv output before the 1x1) stays far below 2^16, so the bf16 hi/lo split is exact too:
    # BOTH tiers must reproduce every layer bit for bit -- d = 1 .. 512 (layers 0..9, the far taps reach 4 tiles away)
    # and the 512 -> 1 wrap (layers 10, 11).
    assert max(np.abs(h).max() for h in H) < 2 ** 13
    for l in range(-1, NL):
        got = st.extract(l)[0].cpu().numpy().astype(np.float64)
        assert np.array_equal(got, H[l + 1]), f"layer {l} (d={1 << (max(l, 0) % 10)})"
    st.close()


@pytest.mark.parametrize("prec", PRECS)
def test_one_hot_impulse_response_positions(prec):
    """A single non-zero sample: the support of every layer's response must sit at exactly the
    oracle's rows (dilation offsets and both clip edges), here with the impulse near the end."""
    NL, T = 10, 1024
    w = synth.make_weights(NL, 3, bias_scale=0.0)
    for pos in (0, 517, T - 1):
        x = np.zeros(T); x[pos] = 64.0
        H, _ = step_np.encoder_forward(x, w, NL)
        from audio_style_transfer_b200.step import StyleStep
        st = StyleStep(T, [NL - 1], list(range(NL)), precision=prec)
        st.set_weights(w)
        st.embeds(x[None], style=False)
        for l in (0, 4, 9):
            got = st.extract(l)[0].cpu().numpy()
            assert np.array_equal(np.abs(got).sum(1) > 0, np.abs(H[l + 1]).sum(1) > 0), (pos, l)
        st.close()


@pytest.mark.parametrize("prec", PRECS)
def test_batch_of_pairs_equals_loop(prec):
    """B independent pairs (BASELINE configs[3], new capability): each row of the batch must equal the
    B=1 result for that pair bit for bit (no cross-pair leakage through the dilated taps)."""
    case = cases.build_inputs("t1024_ours_s0")
    Pc, Ps = cases.oracle_targets(case)
    rng = np.random.default_rng(5)
    B = 3
    xs = np.stack([case["x"] + rng.standard_normal(case["T"]) * (i > 0) for i in range(B)])
    stB = make_step(case, prec, B=B)
    stB.set_targets(np.repeat(Pc[None], B, 0), np.repeat(Ps[None], B, 0), case["lambd"])
    loB, gB = stB.step(xs)
    st1 = make_step(case, prec)
    st1.set_targets(Pc[None], Ps[None], case["lambd"])
    exB = stB.extract(case["NL"] - 1)
    for i in range(B):
        lo, g = st1.step(xs[i:i + 1])
        assert torch.equal(st1.extract(case["NL"] - 1)[0], exB[i]), i      # encoder: bit for bit
        # the Gram split does not depend on B (csrc/api.cu): a pair inside a batch IS the pair alone
        assert torch.equal(g[0], gB[i]) and torch.equal(lo[0], loB[i]), (i, rel(g[0].cpu().numpy(), gB[i].cpu().numpy()))
    stB.close(); st1.close()


@pytest.mark.parametrize("nl", [6, 14])       # 14 style layers: the 13..32-layer channel-wise head (tensor cores in the bf16 tier)
@pytest.mark.parametrize("prec", PRECS)
def test_time_sharded_windows_equal_unsharded(prec, nl):
    """Time sharding emulated on one GPU: P window contexts, partial sums added (= the all-reduce),
    owned gradient slices concatenated.  Must agree with the single-context step; on integer data the
    forward at the seams is bit-exact."""
    from audio_style_transfer_b200.sharding import plan_windows
    case = cases.build_inputs("t4096_cfg0")
    case = dict(case, cont=[nl - 1], style=list(range(nl)), NL=nl)
    case["w"] = synth.make_weights(nl, 0, 0.1)
    Pc, Ps = cases.oracle_targets(case)
    full = make_step(case, prec)
    full.set_targets(Pc[None], Ps[None], case["lambd"])
    lo_f, g_f = full.step(case["x"][None])
    P = 3 if nl == 6 else 2
    plan = plan_windows(case["T"], P, case["NL"])
    steps, parts = [], []
    for (wb, we, ob, oe) in plan:
        st = make_step(case, prec, window=(case["T"], wb, ob, oe), T=we - wb)
        st.set_targets(Pc[None, wb:we], Ps[None], case["lambd"])
        st.forward_partial(case["x"][None, wb:we])
        steps.append(st); parts.append(st.partials())
    total = torch.stack([p.clone() for p in parts]).sum(0)
    g = torch.empty(case["T"], device="cuda")
    for st, p, (wb, we, ob, oe) in zip(steps, parts, plan):
        p.copy_(total)
        lo, gw = st.backward()
        assert st.launches_per_step > 0          # the two-phase path must account for its kernels (bench gpu_launches)
        g[ob:oe] = gw[0, ob - wb:oe - wb]
        assert torch.allclose(lo, lo_f, rtol=2e-5 if prec == "fp32" else 2e-3)
    tol = 2e-5 if prec == "fp32" else 5e-3
    assert rel(g.cpu().numpy(), g_f[0].cpu().numpy()) <= tol
    for st in steps:
        st.close()
    full.close()


def test_errors_are_loud():
    from audio_style_transfer_b200._lib import AstError
    from audio_style_transfer_b200.step import StyleStep
    with pytest.raises(AstError):
        StyleStep(1000, [9], list(range(10)), precision="fp32")          # T % 128
    st = StyleStep(1024, [9], list(range(10)), precision="fp32")
    with pytest.raises(AstError):                                         # AST_ESTATE: no weights/targets
        st.step(np.zeros((1, 1024), np.float32))
    st.set_weights(synth.make_weights(10))
    short = StyleStep(896, [9], list(range(10)), precision="fp32")
    short.set_weights(synth.make_weights(10))
    with pytest.raises(AstError):                                         # gamma != 0 needs one 1024-sample STFT frame
        short.set_targets(np.zeros((1, 896, 128), np.float32), np.zeros((1, 128, 10, 10), np.float32), 100.0, 0.1)
    short.close()
    with pytest.raises(AstError):
        StyleStep(1024, [9], [3, 3], precision="fp32")                    # duplicate style layer
    st.close()


# ---- whole-stack kernels (one cooperative launch per direction) ---------------------------------------
def _step_with_path(case, Pc, Ps, mode):
    """mode "auto": whole-stack kernels; "per_layer": one kernel per layer; "per_layer_bwd": stack forward + per-layer
    backward (ast_config.kernel_path)."""
    st = make_step(case, "bf16", kernel_path=mode)
    st.set_targets(Pc[None], Ps[None], case["lambd"])
    lo, g = st.step(case["x"][None])
    return st, lo.cpu().numpy()[0], g.cpu().numpy()[0]


@pytest.mark.parametrize("name", ["t4096_cfg0", "t4096_cfg1", "t16384_cfg1"])
def test_stack_kernels_agree_with_per_layer_path(name):
    """The persistent whole-stack kernels and the per-layer kernels compute the same step.  The forward differs only in
    the order the three taps are accumulated (fp32): features agree to ~1e-5.  Gradients of the bf16 tier move by a few
    1e-3 under ANY reordering of the fp32 accumulation (bf16 rounding flips cascade through the masked backward; CPU
    emulation in DESIGN.md), so both are held to the oracle tolerance and to each other at the same level."""
    case = cases.build_inputs(name)
    Pc, Ps = cases.oracle_targets(case)
    losses, dx = cases.oracle_step(case, Pc, Ps)
    st_s, lo_s, g_s = _step_with_path(case, Pc, Ps, "auto")
    st_l, lo_l, g_l = _step_with_path(case, Pc, Ps, "per_layer")
    assert st_s.uses_stack_kernels and not st_l.uses_stack_kernels
    st_s.check()
    assert st_s.launches_per_step < st_l.launches_per_step
    last = case["NL"] - 1
    if last in [min(i, 29) for i in case["cont"]]:                 # retained by both paths
        assert rel(st_s.extract(last)[0].cpu().numpy(), st_l.extract(last)[0].cpu().numpy()) <= 1e-4
    assert abs(lo_s[0] - lo_l[0]) <= 1e-4 * abs(lo_l[0])
    tol = TOL["bf16"]["grad"]
    assert rel(g_s, dx) <= tol and rel(g_l, dx) <= tol
    assert rel(g_s, g_l) <= tol
    # identical forward (mode "per_layer_bwd"): the two backward implementations alone
    st_m, lo_m, g_m = _step_with_path(case, Pc, Ps, "per_layer_bwd")
    assert np.array_equal(lo_m, lo_s)
    assert rel(g_s, g_m) <= tol
    for st in (st_s, st_l, st_m):
        st.close()


def test_stack_step_is_reproducible_and_graph_equals_eager():
    case = cases.build_inputs("t16384_cfg1")
    Pc, Ps = cases.oracle_targets(case)
    st = make_step(case, "bf16")
    st.set_targets(Pc[None], Ps[None], case["lambd"])
    assert st.uses_stack_kernels
    x = torch.as_tensor(case["x"][None], dtype=torch.float32, device="cuda")
    lo0, g0 = st.step(x)
    lo0, g0 = lo0.clone(), g0.clone()
    lo = torch.empty_like(lo0); g = torch.empty_like(g0)
    for _ in range(5):
        st.step_graph(x, lo, g)
        torch.cuda.synchronize()
        assert torch.equal(g, g0) and torch.equal(lo, lo0)       # flag hand-off leaves no run-to-run variation
    st.check()
    st.close()


def test_step_keeps_only_head_extracts_in_stack_mode():
    """ast_step holds the residual stream in TMEM and writes only what the heads read; ast_embeds keeps every layer."""
    from audio_style_transfer_b200._lib import AstError
    case = cases.build_inputs("t4096_cfg1")
    Pc, Ps = cases.oracle_targets(case)
    st = make_step(case, "bf16")
    st.embeds(case["x"][None])
    st.extract(3)                                                 # fine after embeds
    st.set_targets(Pc[None], Ps[None], case["lambd"])
    st.step(case["x"][None])
    kept = sorted(set(min(i, 29) for i in case["cont"]))
    st.extract(kept[-1])
    dropped = [l for l in range(case["NL"]) if l not in kept]
    with pytest.raises(AstError, match="not retained"):
        st.extract(dropped[0])
    st.close()


# ---- gamma term: STFT regulariser (methods.py:121-123) ------------------------------------------------
@pytest.mark.parametrize("prec", PRECS)
@pytest.mark.parametrize("kind", ["float", "int"])
def test_regulariser_matches_oracle(kind, prec):
    """loss[3] = reg, loss[0] += gamma * reg, grad += gamma * d reg / dx.  The regulariser runs on fp32 CUDA cores in
    both tiers, so its own value and gradient (isolated as step(gamma) - step(0) of the same context) are held to
    the fp32 tolerance in both; integer x exercises the reference's x == 0 pass-through (utils.py:103)."""
    from oracle import stft_reg
    case = cases.build_inputs("t4096_cfg0")
    Pc, Ps = cases.oracle_targets(case)
    x = case["x"].copy()
    if kind == "int":
        x = np.floor(x)
        x[::97] = 0.0
    gamma = 50.0
    reg, dreg, _ = stft_reg.stft_l1(x)
    st = make_step(case, prec)
    st.set_targets(Pc[None], Ps[None], case["lambd"], 0.0)
    lo0, g0 = (t.cpu().numpy()[0].astype(np.float64) for t in st.step(x[None]))
    st.set_targets(Pc[None], Ps[None], case["lambd"], gamma)
    lo1, g1 = (t.cpu().numpy()[0].astype(np.float64) for t in st.step(x[None]))
    assert lo0[3] == 0.0
    assert abs(lo1[3] - reg) <= 1e-5 * reg
    assert abs((lo1[0] - lo0[0]) - gamma * reg) <= 1e-4 * gamma * reg
    assert np.array_equal(lo1[1:3], lo0[1:3])
    assert rel(g1 - g0, gamma * dreg) <= 1e-4
    # and the whole step against the oracle at the tier's tolerance
    losses, dx = step_np.step(x, case["w"], Pc, Ps, case["cont"], case["style"], case["gatys"], case["lambd"], gamma,
                              case["nb"], case["cnt"])
    assert abs(lo1[0] - losses[0]) <= TOL[prec]["loss"] * abs(losses[0])
    assert rel(g1, dx) <= grad_tol(prec, fp32_floor(dict(case, x=x), Pc, Ps, dx - gamma * dreg))
    # graph replay carries the regulariser kernels too
    lo = torch.empty(1, 4, device="cuda"); g = torch.empty(1, case["T"], device="cuda")
    st.step_graph(torch.tensor(x[None], dtype=torch.float32, device="cuda"), lo, g)
    torch.cuda.synchronize()
    assert np.allclose(g.cpu().numpy()[0], g1, rtol=0, atol=1e-6 * np.abs(g1).max())
    st.close()


def test_regulariser_time_sharded_equals_unsharded():
    """Each window context evaluates the STFT frames that touch its owned samples and counts the ones that START
    there; summed partials give the same reg, owned gradient slices the same gradient."""
    from audio_style_transfer_b200.sharding import plan_windows
    case = cases.build_inputs("t4096_cfg0")
    case = dict(case, cont=[5], style=[0, 1, 2, 3, 4, 5], NL=6, T=8192)
    case["w"] = synth.make_weights(6, 0, 0.1)
    rng = np.random.default_rng(2)
    x = rng.normal(0, 40, 8192)
    Pc = rng.standard_normal((8192, 128)); Ps = rng.standard_normal((128, 6, 6)) * 0.1
    gamma = 50.0
    full = make_step(case, "fp32")
    full.set_targets(Pc[None], Ps[None], case["lambd"], gamma)
    lo_f, g_f = full.step(x[None])
    plan = plan_windows(8192, 3, case["NL"], min_halo=1024)
    steps, parts = [], []
    for (wb, we, ob, oe) in plan:
        st = make_step(case, "fp32", window=(8192, wb, ob, oe), T=we - wb)
        st.set_targets(Pc[None, wb:we], Ps[None], case["lambd"], gamma)
        st.forward_partial(x[None, wb:we])
        steps.append(st); parts.append(st.partials())
    total = torch.stack([p.clone() for p in parts]).sum(0)
    g = torch.empty(8192, device="cuda")
    for st, p, (wb, we, ob, oe) in zip(steps, parts, plan):
        p.copy_(total)
        lo, gw = st.backward()
        g[ob:oe] = gw[0, ob - wb:oe - wb]
        assert torch.allclose(lo, lo_f, rtol=2e-5)
    assert float(lo_f[0, 3]) > 0
    assert rel(g.cpu().numpy(), g_f[0].cpu().numpy()) <= 2e-5
    for st in steps:
        st.close()
    full.close()


@pytest.mark.parametrize("L", [1, 2, 12, 13, 16, 17, 30])
def test_tma_fed_channelwise_head_agrees_with_the_second_implementation(L):
    """The channel-wise head of the bf16 tier streams transposed extract copies by TMA (gram_ours_tma.cu, 16- and 32-layer
    stages); the per-layer kernel path keeps the kernels that read the [t][c] extracts (kernels_f32.cu / gram_ours_tc.cu).
    Same Gram (TF32 / fp32 products, fp32 sums: 1e-4) for every stage geometry -- one layer, a full 16-layer stage, one past
    it, all 30 -- and, for the edge counts, the Gram against the fp64 oracle."""
    from audio_style_transfer_b200 import synth
    from audio_style_transfer_b200.mu_law import mu_law_numpy
    T = 1024
    style = list(range(30 - L, 30))                                 # the top L layers: every dilation class
    cont = [29]
    w = synth.make_weights(30, seed=0, bias_scale=0.1)
    c_mu = mu_law_numpy(synth.content_audio(T)).astype(np.float64)
    s_mu = mu_law_numpy(synth.style_audio(T, chunks=1)).astype(np.float64)
    case = dict(T=T, cont=cont, style=style, gatys=False, nb=128, cnt=128, NL=30, w=w, c_mu=c_mu, s_chunks=[s_mu],
                x=synth.step_input(c_mu).astype(np.float64), lambd=cases.LAMBD)
    st_a, st_b = make_step(case, "bf16"), make_step(case, "bf16", kernel_path="per_layer")
    assert st_a.uses_stack_kernels and not st_b.uses_stack_kernels
    Ec_a, Gn_a = st_a.embeds(s_mu[None])
    Ec_b, Gn_b = st_b.embeds(s_mu[None])
    assert rel(Gn_a.cpu().numpy(), Gn_b.cpu().numpy()) <= 1e-4, L
    if L in (1, 16, 17):
        Pc, Ps = cases.oracle_targets(case)
        assert rel(Gn_a[0].cpu().numpy(), Ps[:Gn_a.shape[1]]) <= TOL["bf16"]["feat"], L
    rng = np.random.default_rng(L)
    Pc = rng.standard_normal((1, T, 128)).astype(np.float32)
    Ps = rng.standard_normal(tuple(Gn_a.shape)).astype(np.float32)
    Ps /= np.sqrt((Ps ** 2).sum(axis=(-2, -1), keepdims=True))
    res = []
    for st in (st_a, st_b):
        st.set_targets(Pc, Ps, case["lambd"])
        lo, g = st.step(case["x"][None])
        torch.cuda.synchronize()
        st.check()
        res.append((lo.cpu().numpy()[0], g.cpu().numpy()[0]))
        st.close()
    (lo_a, g_a), (lo_b, g_b) = res
    assert abs(lo_a[0] - lo_b[0]) <= 1e-3 * abs(lo_b[0]), (L, lo_a, lo_b)
    assert rel(g_a, g_b) <= TOL["bf16"]["grad"], L                 # two accumulation orders of the bf16 tier (see above)


def test_channelwise_head_saturates_instead_of_overflowing():
    """The TMA-fed channel-wise head reads fp16 copies of the style extracts (gram_ours_tma.cu): values beyond +-65504 are
    clipped by a saturating conversion ("Differences ..."), never turned into inf / NaN.  Biases of 2e5 push
    every extract past the fp16 range: Gram, losses and gradient must stay finite, and the extract itself (fp32) is exact."""
    from audio_style_transfer_b200 import synth
    from audio_style_transfer_b200.mu_law import mu_law_numpy
    T = 1024
    w = synth.make_weights(10, seed=0, bias_scale=2e5)
    c_mu = mu_law_numpy(synth.content_audio(T)).astype(np.float64)
    case = dict(T=T, cont=[9], style=list(range(10)), gatys=False, nb=128, cnt=128, NL=10, w=w)
    st = make_step(case, "bf16")
    x = synth.step_input(c_mu)[None]
    Ec, Gn = st.embeds(x)
    assert float(st.extract(9).abs().max()) > 65504.0               # the premise: out of fp16 range
    assert bool(torch.isfinite(Gn).all())
    rng = np.random.default_rng(0)
    Pc = rng.standard_normal((1, T, 128)).astype(np.float32)
    st.set_targets(Pc, Gn.cpu().numpy() * 0.5, cases.LAMBD)
    lo, g = st.step(x)
    torch.cuda.synchronize()
    st.check()
    assert bool(torch.isfinite(lo).all()) and bool(torch.isfinite(g).all())
    st.close()
