"""The device-resident optimiser (csrc/lbfgs.cu, C ABI ast_opt_* / ast_minimize) against oracle/lbfgs_np.py, the CPU
restatement of SciPy's L-BFGS-B loop the reference runs (methods.py:133-137,164-181)."""
import numpy as np
import pytest

from oracle import cases, lbfgs_np as L

pytestmark = pytest.mark.gpu


def rosen(x):
    f = np.sum(100 * (x[1:] - x[:-1] ** 2) ** 2 + (1 - x[:-1]) ** 2)
    g = np.zeros_like(x)
    g[:-1] += -400 * x[:-1] * (x[1:] - x[:-1] ** 2) - 2 * (1 - x[:-1])
    g[1:] += 200 * (x[1:] - x[:-1] ** 2)
    return f, g


def smooth_chain(x):
    """Cheap non-quadratic function of many variables (n up to 10^5) with O(1) curvature."""
    f = np.sum(np.log(np.cosh(x - 0.3))) + 0.5 * np.sum((x[1:] - x[:-1]) ** 2) + 0.1 * np.sum(np.sin(3 * x))
    g = np.tanh(x - 0.3) + 0.3 * np.cos(3 * x)
    g[1:] += x[1:] - x[:-1]
    g[:-1] -= x[1:] - x[:-1]
    return f, g


def lockstep(fn, x0, B=1, evals=60, **kw):
    """Drive the device machine and the oracle machine (fp32 storage) with the SAME (f, g), evaluated on the CPU at the
    DEVICE's point and rounded to fp32; every evaluation point and every counter must agree."""
    import torch
    from audio_style_transfer_b200.step import DeviceLBFGS
    T = x0.shape[-1]
    opt = DeviceLBFGS(T, B=B, max_log=evals, **kw)
    opt.reset(x0)
    machs = [L.LbfgsMachine(x0[b], store=np.float32, **kw) for b in range(B)]
    worst = 0.0
    for k in range(evals):
        xd = opt.x.cpu().numpy()
        st, _ = opt.status()
        for b in range(B):
            assert (st[b, 0] == 2) == machs[b].done, (k, b, st[b])
            if machs[b].done:
                continue
            xo = machs[b].x_trial
            err = np.max(np.abs(xd[b] - xo)) / max(1e-30, np.max(np.abs(xo)))
            worst = max(worst, err)
            assert err <= 2e-6, (k, b, err)
            assert st[b, 2] == machs[b].iter and st[b, 4] == machs[b].nfev_total, (k, b, st[b])
        if all(m.done for m in machs):
            break
        f32 = np.zeros((B, 4), np.float32)
        g32 = np.zeros((B, T), np.float32)
        for b in range(B):
            f, g = fn(xd[b].astype(np.float64))
            f32[b, 0], g32[b] = f, g
        opt.losses.copy_(torch.as_tensor(f32))
        opt.grad.copy_(torch.as_tensor(g32))
        opt.advance()
        for b in range(B):
            machs[b].advance(float(f32[b, 0]), g32[b])
    st, fk = opt.status()
    for b in range(B):
        m = machs[b]
        assert st[b, 1] == m.epoch and st[b, 4] == m.nfev_total and st[b, 6] == m.col, (st[b], m.epoch, m.nfev_total, m.col)
        if m.epoch_end:
            assert st[b, 5] == m.epoch_end[-1]
    xk = opt.result()
    for b in range(B):
        assert np.max(np.abs(xk[b] - machs[b].xk)) <= 2e-6 * max(1.0, np.max(np.abs(machs[b].xk)))
    nfev, end, _ = opt.epochs_info()
    for b in range(B):
        assert list(nfev[b, :machs[b].epoch]) == machs[b].epoch_nfev
    opt.close()
    return worst


def test_machine_lockstep_rosenbrock_small():
    x0 = np.random.default_rng(0).standard_normal((1, 700))         # one CTA (cluster of 1)
    lockstep(rosen, x0, evals=80, maxiter=100)


def test_machine_lockstep_cluster_of_8_two_pairs():
    x0 = np.random.default_rng(1).standard_normal((2, 16384)) * 2.0
    x0[1] *= 0.5
    lockstep(smooth_chain, x0, B=2, evals=70, maxiter=100)


def test_machine_lockstep_ragged_length_and_epochs():
    """T not a multiple of the cluster slice; several epochs with the < min_evals stop rule (methods.py:180-181)."""
    x0 = np.random.default_rng(2).standard_normal((1, 5001))
    lockstep(smooth_chain, x0, evals=150, maxiter=12, epochs=6, min_evals=14)


def test_minimize_on_the_step_follows_oracle_machine():
    """ast_minimize (graph of [step, advance], no host in the loop) vs the oracle machine pumping ast_step_host."""
    from audio_style_transfer_b200.step import DeviceLBFGS, StyleStep
    case = cases.build_inputs("t1024_ours_s0")
    Pc, Ps = cases.oracle_targets(case)
    st = StyleStep(case["T"], case["cont"], case["style"], case["gatys"], precision="fp32", device=0)
    st.set_weights(case["w"])
    st.set_targets(Pc[None], Ps[None], case["lambd"])
    x0 = case["x"][None].astype(np.float64)
    maxiter = 25
    opt = DeviceLBFGS(case["T"], maxiter=maxiter, epochs=1, min_evals=0, max_log=256)
    opt.reset(x0)
    n = opt.minimize(st, chunk=4)
    status, fk = opt.status()
    log = opt.log()[0]
    assert status[0, 0] == 2 and status[0, 2] == maxiter and status[0, 5] == 1
    nfev = int(status[0, 4])
    assert nfev <= n <= nfev + 8                      # at most 2 * chunk - 1 evaluations after the machine stopped

    mach = L.LbfgsMachine(x0[0], maxiter=maxiter, epochs=1, min_evals=0, store=np.float32)
    ref_log = []
    while not mach.done:
        lo, g = st.step_host(mach.x_trial.astype(np.float64)[None])
        ref_log.append(lo[0, 0])
        mach.advance(np.float32(lo[0, 0]), g[0].astype(np.float32))
    # identical f/g source and identical algorithm: the first evaluations agree closely, the end point in value
    k = min(12, nfev, len(ref_log))
    assert np.allclose(log[:k, 0], ref_log[:k], rtol=1e-4), (log[:k, 0], ref_log[:k])
    assert abs(nfev - mach.nfev_total) <= 3
    assert log[nfev - 1, 0] < 0.7 * log[0, 0]
    assert abs(fk[0] - mach.fk) <= 0.05 * abs(mach.fk)
    # the losses row is the step's own: total = content + lambd * style
    assert np.allclose(log[:nfev, 0], log[:nfev, 1] + case["lambd"] * log[:nfev, 2], rtol=1e-4)
    opt.close()
    st.close()


def test_pairs_advance_independently_and_equal_single_runs():
    from audio_style_transfer_b200.step import DeviceLBFGS, StyleStep
    case = cases.build_inputs("t1024_ours_s0")
    Pc, Ps = cases.oracle_targets(case)
    rng = np.random.default_rng(5)
    x0 = np.stack([case["x"], case["x"] + rng.standard_normal(case["T"])]).astype(np.float64)
    outs = []
    for B, xs in ((2, x0), (1, x0[:1]), (1, x0[1:])):
        st = StyleStep(case["T"], case["cont"], case["style"], case["gatys"], B=B, precision="bf16", device=0)
        st.set_weights(case["w"])
        st.set_targets(np.repeat(Pc[None], B, 0), np.repeat(Ps[None], B, 0), case["lambd"])
        opt = DeviceLBFGS(case["T"], B=B, maxiter=10, epochs=2, min_evals=5, max_log=64)
        opt.reset(xs)
        opt.minimize(st)
        outs.append((opt.result(), opt.status()[0], opt.log()))
        opt.close()
        st.close()
    (x2, s2, l2), (xa, sa, la), (xb, sb, lb) = outs
    assert np.array_equal(s2[0], sa[0]) and np.array_equal(s2[1], sb[0])
    assert np.array_equal(x2[0], xa[0]) and np.array_equal(x2[1], xb[0])         # B pairs == B single runs, bit for bit
    assert np.array_equal(l2[0], la[0]) and np.array_equal(l2[1], lb[0])


def test_gatysnet_device_optimizer(tmp_path):
    from audio_style_transfer_b200 import synth
    from audio_style_transfer_b200.methods import GatysNet, l2_normalize
    T = 4096
    res = {}
    for which in ("scipy", "device"):
        net = GatysNet(savepath=str(tmp_path / which), checkpoint_path=None, batch_size=T, stack=0, cont_lyr_ids=[9],
                       precision="fp32", seed=0, optimizer=which)
        net.load_model(None)
        aud = synth.content_audio(T, seed=1)
        phi_c = net.get_embeds(None, aud)
        phi = l2_normalize(net.get_embeds(None, aud, is_content=False) + net.get_style_phi(None, synth.style_audio(T, seed=2, chunks=2))
                           - net.get_style_phi(None, aud))
        seen = []
        net.l_bfgs(None, phi_c, phi, epochs=2, lambd=100.0, gamma=0.0, callback=lambda ep, i, l: seen.append(float(l[0])))
        res[which] = (seen, net.evaluations)
        assert (tmp_path / which / "ep-0.wav").exists() and (tmp_path / which / "ep-1.wav").exists()
    (ss, ns), (sd, nd) = res["scipy"], res["device"]
    # (how many evaluations the flat tail takes before ftol fires depends on the last bits: 153 vs 204 were both seen)
    assert len(sd) == nd and nd >= 50 and ns >= 50
    assert np.allclose(sd[:8], ss[:8], rtol=1e-3)
    assert sd[-1] <= 1.25 * ss[-1] and sd[-1] < 0.3 * sd[0]
