"""BASELINE.json configs[2..4] at (or near) their real shapes: the paths the small parity cases never reach.

* configs[3] shape: B = 8 pairs of T = 16384 in one context -- the whole-stack kernels are launched pair by pair with
  tile0 > 0 (csrc/kernels_tc.cu: tc_forward_stack / tc_backward_stack);
* configs[2]: T = 65536 -- 512 tiles, more than one tile per SM: the whole-stack kernels run stack by stack on groups of
  <= 148 tiles (csrc/kernels_tc.cu: tc_plan_launches), checked against the golden vectors, the one-kernel-per-layer path
  and its 8 time-shard windows, both the overlapped ones (plan_windows) and the ones that exchange halo rows stack by
  stack (HaloShardedStep; here the "ranks" are contexts on one GPU and the exchange a device copy);
* configs[4] geometry: 30 layers, channel-wise head over all 30 (tensor-core head), two 516-tile windows.
Every result must equal the one-clip / one-pair evaluation (reference methods.py:50,58-76: one clip, one graph) and,
where the fp64 oracle is affordable, the committed golden vectors (oracle/make_golden.py)."""
import numpy as np
import pytest
import torch

from audio_style_transfer_b200 import synth
from audio_style_transfer_b200.sharding import plan_windows
from oracle import cases

pytestmark = pytest.mark.gpu
TOL_BF16 = 1e-2           # BASELINE.json: rel. L2 <= 1e-2 in bf16 mode (loss and gradient)


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def make_step(case, B=1, window=None, T=None, prec="bf16", **kw):
    from audio_style_transfer_b200.step import StyleStep
    st = StyleStep(T or case["T"], case["cont"], case["style"], case["gatys"], case["nb"], case["cnt"], B=B,
                   precision=prec, window=window, **kw)
    st.set_weights(case["w"])
    return st


def check_against_golden(gold, lo, g, what):
    stride = int(gold["dx_stride"])
    e_loss = abs(lo[0] - gold["losses"][0]) / abs(gold["losses"][0])
    e_grad = rel(g[::stride], gold["dx"])
    print(f"{what}: loss rel {e_loss:.2e}, gradient rel-L2 (every {stride}th sample) {e_grad:.2e}")
    assert e_loss <= TOL_BF16 and e_grad <= TOL_BF16, (what, e_loss, e_grad)


@pytest.mark.parametrize("name", ["t16384_cfg0", "t16384_cfg1"])
def test_eight_pairs_per_context_equal_single_pairs_and_oracle(name, golden_dir):
    B = 8
    case = cases.build_inputs(name)
    Pc, Ps = cases.oracle_targets(case)
    xs = np.stack([cases.pair_x(case, n) for n in range(B)])
    assert np.array_equal(xs[0], case["x"])
    stB = make_step(case, B=B)
    assert stB.uses_stack_kernels                     # 128 tiles per pair on 148 SMs: one launch per pair, tile0 = 128 n
    stB.set_targets(np.repeat(Pc[None], B, 0), np.repeat(Ps[None], B, 0), case["lambd"])
    loB, gB = stB.step(xs)
    stB.check()
    exB = stB.extract(case["NL"] - 1)
    st1 = make_step(case)
    st1.set_targets(Pc[None], Ps[None], case["lambd"])
    for i in range(B):
        lo, g = st1.step(xs[i:i + 1])
        assert torch.equal(st1.extract(case["NL"] - 1)[0], exB[i]), i          # encoder: bit for bit
        # the Gram split does not depend on B (csrc/api.cu): a pair inside a batch IS the pair alone
        assert torch.equal(g[0], gB[i]) and torch.equal(lo[0], loB[i]), (i, rel(g[0].cpu().numpy(), gB[i].cpu().numpy()))
    st1.check()
    loB, gB = loB.cpu().numpy(), gB.cpu().numpy()
    check_against_golden(np.load(golden_dir / f"{name}.npz"), loB[0], gB[0], f"{name} pair 0 of 8")
    check_against_golden(np.load(golden_dir / f"{name}_p7.npz"), loB[7], gB[7], f"{name} pair 7 of 8")
    stB.close(); st1.close()


def run_windows(case, Pc, Ps, plan, x):
    """P window contexts on one GPU; the sum of their partial buffers stands in for the all-reduce."""
    steps, parts = [], []
    for (wb, we, ob, oe) in plan:
        st = make_step(case, window=(case["T"], wb, ob, oe), T=we - wb)
        st.set_targets(Pc[None, wb:we], Ps[None], case["lambd"])
        st.forward_partial(x[None, wb:we])
        steps.append(st); parts.append(st.partials())
    total = torch.stack([p.clone() for p in parts]).sum(0)
    g = torch.empty(case["T"], device="cuda")
    los = []
    for st, p, (wb, we, ob, oe) in zip(steps, parts, plan):
        p.copy_(total)
        lo, gw = st.backward()
        st.check()
        g[ob:oe] = gw[0, ob - wb:oe - wb]
        los.append(lo[0].cpu().numpy())
    return steps, np.array(los), g.cpu().numpy()


def run_halo_windows(case, Pc, Ps, P, x, B=1, p2p=False):
    """HaloShardedStep for P "ranks" that are contexts on ONE GPU, driven in lockstep; the exchange is a device copy
    between the contexts' segment buffers and the all-reduce a sum of their partial buffers.  Returns (losses, gradient
    of the whole clip, last extract of the whole clip)."""
    from audio_style_transfer_b200.sharding import HaloShardedStep
    T = case["T"]
    mail = {}                                          # (src rank, dst rank, tag) -> rows in flight

    def make(rank):
        def exchange(pairs, rank=rank):
            for send, recv, peer in pairs:
                mail[(rank, peer)] = (send.clone(), recv)
        hs = HaloShardedStep(T, rank, P, lambda Tw, win: make_step(case, window=win, T=Tw, halo_exchange=True),
                             None, exchange)
        hs.ctx.set_targets(Pc[None, hs.wb:hs.we], Ps[None], case["lambd"])
        return hs
    shards = [make(r) for r in range(P)]
    if p2p:                                            # peer-mapped halos: the neighbours are contexts of this process
        for r, h in enumerate(shards):
            if r > 0:
                h.ctx.p2p_attach_local(0, shards[r - 1].ctx)
            if r + 1 < P:
                h.ctx.p2p_attach_local(1, shards[r + 1].ctx)
    nseg = shards[0].nseg
    assert nseg == 3 and all(h.ctx.uses_stack_kernels for h in shards)
    xw = [torch.as_tensor(x[None, h.wb:h.we], dtype=torch.float32, device="cuda").contiguous() for h in shards]
    los = [torch.empty((1, 4), device="cuda") for _ in shards]
    grads = [torch.empty((1, h.we - h.wb), device="cuda") for h in shards]

    def deliver():
        # every rank posted its sends and named its receive buffers: rank r's receive from peer q gets q's send to r
        for (src, dst), (rows, _) in list(mail.items()):
            mail[(dst, src)][1].copy_(rows)
        mail.clear()
    H = 1024
    for k in range(nseg):
        for h, xx in zip(shards, xw):
            h.ctx.forward_seg(xx, k)
        # (every context's segment k is enqueued before any segment k + 1: a kernel that resumes from a neighbour's rows
        # finds the neighbour's completion word already written -- ranks that WAIT on one another must not share a GPU)
        if k + 1 < nseg and not p2p:
            for h in shards:
                ob, oe = h.ob, h.oe
                h._swap(h.ctx.segment_buffer(False, k), (ob + H, ob + 2 * H), (ob - 2 * H, ob - H), (oe - 2 * H, oe - H), (oe + H, oe + 2 * H))
            deliver()
    parts = [h.ctx.partials() for h in shards]
    total = torch.stack([p.clone() for p in parts]).sum(0)
    for p in parts:
        p.copy_(total)
    for k in range(nseg - 1, -1, -1):
        for h, lo, g in zip(shards, los, grads):
            h.ctx.backward_seg(k, lo, g)
        if k > 0 and not p2p:
            for h in shards:
                ob, oe = h.ob, h.oe
                h._swap(h.ctx.segment_buffer(True, k), (ob, ob + H), (ob - H, ob), (oe - H, oe), (oe, oe + H))
            deliver()
    g = torch.empty(T, device="cuda")
    ex = torch.empty((T, 128), device="cuda")
    for h, gw in zip(shards, grads):
        h.ctx.check()
        g[h.ob:h.oe] = gw[0, h.ob - h.wb:h.oe - h.wb]
        ex[h.ob:h.oe] = h.ctx.extract(case["NL"] - 1)[0, h.ob - h.wb:h.oe - h.wb]
    lo0 = los[0].cpu().numpy()[0]
    for lo in los:
        assert np.allclose(lo.cpu().numpy()[0], lo0, rtol=1e-6)
    for h in shards:
        h.ctx.close()
    return lo0, g.cpu().numpy(), ex


def test_segments_and_groups_reproduce_the_one_launch_step_bit_for_bit():
    """Running the stack kernels stack by stack (the fp32 stream / running gradient pass through global memory between
    launches) and on groups of tiles with discarded halo tiles changes NOTHING: same bits as one launch per direction.
    The group planner is told the GPU has 40 SMs (-DAST_EXPERIMENTS build only) so that a 128-tile clip needs 6 groups."""
    from audio_style_transfer_b200 import _lib
    for name in ("t16384_cfg1", "t16384_cfg0"):
        case = cases.build_inputs(name)
        Pc, Ps = cases.oracle_targets(case)
        x = torch.as_tensor(case["x"][None], dtype=torch.float32, device="cuda")
        ref = make_step(case)
        ref.set_targets(Pc[None], Ps[None], case["lambd"])
        assert ref.segments == 1
        lo_r, g_r = ref.step(x)
        ex_r = ref.extract(case["NL"] - 1).clone()
        n_ref = ref.launches_per_step
        seg = make_step(case, halo_exchange=True)      # unsharded, but stack by stack
        seg.set_targets(Pc[None], Ps[None], case["lambd"])
        assert seg.segments == 3
        lo_s, g_s = seg.step(x)
        seg.check()
        assert seg.launches_per_step == n_ref + 4
        assert torch.equal(seg.extract(case["NL"] - 1), ex_r) and torch.equal(g_s, g_r) and torch.equal(lo_s, lo_r), name
        # ... and driven segment by segment through the C ABI: the two-phase step (ast_step_forward / ast_partials /
        # ast_step_backward, where an all-reduce may sit between the phases) taken apart further
        ref.forward_partial(x)
        lo_r, g_r = ref.backward()
        lo2, g2 = torch.empty_like(lo_r), torch.empty_like(g_r)
        for k in range(3):
            seg.forward_seg(x, k)
        for k in (2, 1, 0):
            seg.backward_seg(k, lo2, g2)
        assert torch.equal(g2, g_r) and torch.equal(lo2, lo_r), name
        seg.close()
        lo_r, g_r = ref.step(x)
        lib = _lib.load(experiments=True)
        lib.ast_debug_group_tiles(40)
        try:
            grp = make_step(case, experiments=True)
            grp.set_targets(Pc[None], Ps[None], case["lambd"])
            assert grp.segments == 3
            lo_g, g_g = grp.step(x)
            grp.check()
            assert grp.launches_per_step > n_ref + 4
            assert torch.equal(grp.extract(case["NL"] - 1), ex_r) and torch.equal(g_g, g_r) and torch.equal(lo_g, lo_r), name
            grp.close()
        finally:
            lib.ast_debug_group_tiles(0)
        ref.close()


def test_cfg2_clip_unsharded_and_eight_windows_match_oracle(golden_dir):
    case = cases.build_inputs("t65536_cfg2")
    gold = np.load(golden_dir / "t65536_cfg2.npz")
    Pc, Ps = cases.oracle_targets(case)
    assert abs(Ps.sum() - gold["Ps_sum"]) <= 1e-9 * abs(gold["Ps_sum"])
    full = make_step(case)
    assert full.uses_stack_kernels and full.segments == 3        # 512 tiles > 148 SMs: 3 stacks x 4 groups per direction
    full.set_targets(Pc[None], Ps[None], case["lambd"])
    lo_f, g_f = full.step(case["x"][None])
    full.check()
    assert full.launches_per_step <= 36
    ex_f = full.extract(case["NL"] - 1)[0].clone()
    lo_f, g_f = lo_f.cpu().numpy()[0], g_f.cpu().numpy()[0]
    check_against_golden(gold, lo_f, g_f, "T=65536 unsharded (whole-stack kernels, segments x groups)")
    full.close()
    pl = make_step(case, kernel_path="per_layer")
    assert not pl.uses_stack_kernels
    pl.set_targets(Pc[None], Ps[None], case["lambd"])
    lo_p, g_p = pl.step(case["x"][None])
    lo_p, g_p = lo_p.cpu().numpy()[0], g_p.cpu().numpy()[0]
    check_against_golden(gold, lo_p, g_p, "T=65536 unsharded (one kernel per layer)")
    assert rel(pl.extract(case["NL"] - 1)[0].cpu().numpy(), ex_f.cpu().numpy()) <= 1e-4
    assert rel(g_f, g_p) <= 5e-3
    pl.close()
    plan = plan_windows(case["T"], 8, case["NL"])
    steps, los, g = run_windows(case, Pc, Ps, plan, case["x"])
    assert all(st.uses_stack_kernels for st in steps)  # 130-tile (interior) / 97-tile (edge) windows
    assert np.allclose(los, los[0], rtol=1e-6)         # every rank forms the same losses from the summed partials
    check_against_golden(gold, los[0], g, "T=65536, 8 windows (whole-stack kernels)")
    print(f"8 windows vs unsharded: gradient rel-L2 {rel(g, g_f):.2e}")
    assert rel(g, g_f) <= 5e-3                         # the Gram sums are split differently (bf16 tier)
    for st in steps:
        st.close()
    # the same 8 shards with the receptive-field halo EXCHANGE: windows of own +- 2048 rows, stack by stack
    lo_h, g_h, ex_h = run_halo_windows(case, Pc, Ps, 8, case["x"])
    check_against_golden(gold, lo_h, g_h, "T=65536, 8 shards with halo exchange")
    assert torch.equal(ex_h, ex_f)                     # the seam: every shard's owned rows of the last extract, bit for bit
    print(f"8 halo-exchange shards vs unsharded: gradient rel-L2 {rel(g_h, g_f):.2e}")
    assert rel(g_h, g_f) <= 5e-3
    # ... and with the halo rows read by the resuming kernels from the neighbours' buffers (ast_p2p_attach_local), twice:
    # the completion words carry the evaluation number
    for rep_ in range(2):
        lo_p, g_p, ex_p = run_halo_windows(case, Pc, Ps, 8, case["x"], p2p=True)
        assert torch.equal(ex_p, ex_f)
        assert np.array_equal(g_p, g_h) and np.array_equal(lo_p, lo_h)      # the same rows by another route: same bits


def test_cfg4_geometry_two_windows_equal_unsharded():
    """30 layers, channel-wise Gram over all 30 layers, windows of 516 tiles (cfg4's are 564): no oracle at this size
    (it would take tens of minutes); the sharded evaluation must equal the unsharded one."""
    NL, T = 30, 2 * 59904
    w = synth.make_weights(NL, 0, 0.1)
    case = dict(T=T, cont=[NL - 1], style=list(range(NL)), gatys=False, nb=128, cnt=128, NL=NL, w=w, lambd=100.0)
    rng = np.random.default_rng(4)
    x = rng.normal(0, 40, T)
    full = make_step(case)
    Ec, Gn = full.embeds(x[None])
    Pc, Ps = (Ec * 0.9)[0].cpu().numpy(), Gn.roll(1, 1)[0].cpu().numpy()
    full.set_targets(Pc[None], Ps[None], case["lambd"])
    lo_f, g_f = full.step(x[None])
    lo_f, g_f = lo_f.cpu().numpy()[0], g_f.cpu().numpy()[0]
    full.close()
    plan = plan_windows(T, 2, NL)
    assert plan[0][1] - plan[0][0] == 516 * 128
    steps, los, g = run_windows(case, Pc, Ps, plan, x)
    assert all(st.uses_stack_kernels and st.segments == 3 for st in steps)
    print(f"cfg4 geometry, 2 windows vs unsharded: loss rel {abs(los[0][0] - lo_f[0]) / abs(lo_f[0]):.2e}, gradient rel-L2 {rel(g, g_f):.2e}")
    assert np.allclose(los[0], lo_f, rtol=2e-3)
    assert rel(g, g_f) <= 5e-3
    for st in steps:
        st.close()
