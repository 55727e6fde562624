"""The optimiser oracle (oracle/lbfgs_np.py) against the reference's own optimiser dependency: SciPy's L-BFGS-B as
`ScipyOptimizerInterface(method='L-BFGS-B', options={'maxiter': 100})` calls it (methods.py:133-137), and SciPy's
MINPACK-2 line search.  CPU only."""
import numpy as np
import pytest
from scipy.optimize import minimize

from oracle import lbfgs_np as L


def rosen(x):
    f = np.sum(100 * (x[1:] - x[:-1] ** 2) ** 2 + (1 - x[:-1]) ** 2)
    g = np.zeros_like(x)
    g[:-1] += -400 * x[:-1] * (x[1:] - x[:-1] ** 2) - 2 * (1 - x[:-1])
    g[1:] += 200 * (x[1:] - x[:-1] ** 2)
    return f, g


def make_quad(n, seed):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((n, n))
    A = A @ A.T + 0.1 * np.eye(n)
    b = rng.standard_normal(n)
    return lambda x: (0.5 * x @ A @ x - b @ x, A @ x - b)


def rough(x):
    f = np.sum(np.log(np.cosh(3 * x))) + 0.5 * np.sum((x[1:] - x[:-1]) ** 2) + np.sum(np.sin(5 * x))
    g = 3 * np.tanh(3 * x) + 5 * np.cos(5 * x)
    g[1:] += x[1:] - x[:-1]
    g[:-1] -= x[1:] - x[:-1]
    return f, g


CASES = {
    "rosen20": (rosen, 20, 1.0, 100),
    "rosen50_maxiter30": (rosen, 50, 1.0, 30),
    "quad40": (make_quad(40, 5), 40, 1.0, 100),
    "rough30": (rough, 30, 2.0, 100),
}


@pytest.mark.parametrize("name", list(CASES))
def test_machine_reproduces_scipy_lbfgsb(name):
    fn, n, scale, maxiter = CASES[name]
    x0 = np.random.default_rng(11).standard_normal(n) * scale
    pts = []

    def traced(x):
        pts.append(x.copy())
        return fn(x)
    ref = minimize(traced, x0, jac=True, method="L-BFGS-B", options={"maxiter": maxiter})
    x, mach, mine = L.minimize(fn, x0, maxiter=maxiter, epochs=1)
    # the same evaluation points, one by one, until rounding differences (two-loop vs compact form) are amplified by
    # the problem itself: the first 30 evaluations agree to 1e-9 on every case
    k = min(30, len(pts), len(mine))
    for i in range(k):
        assert np.linalg.norm(pts[i] - mine[i]) <= 1e-9 * max(1.0, np.linalg.norm(pts[i])), (name, i)
    if name != "rosen20":                       # rosen20 drifts apart after ~50 evaluations (chaotic), the others do not
        assert mach.nfev_total == ref.nfev and mach.iter == ref.nit
        assert np.linalg.norm(x - ref.x) <= 1e-6 * max(1.0, np.linalg.norm(ref.x))
    else:
        assert abs(mach.nfev_total - ref.nfev) <= 5 and mach.iter == ref.nit
    reason = mach.epoch_end[0]
    assert (reason == L.END_MAXITER) == ("ITERATIONS REACHED LIMIT" in ref.message)
    if "REL" in ref.message:
        assert reason == L.END_FTOL


def test_dcsrch_equals_scipy_minpack2_port():
    from scipy.optimize._dcsrch import DCSRCH
    rng = np.random.default_rng(3)
    for trial in range(40):
        c = rng.standard_normal(5)
        c[4] = abs(c[4]) + 0.1                              # quartic that goes up at infinity
        c[1] = -abs(c[1]) - 1.0                             # descent at 0 (beats the +0.9 of the sine term)
        phi = lambda a: float(c[0] + c[1] * a + c[2] * a * a + c[3] * a ** 3 + c[4] * a ** 4 + 0.3 * np.sin(3 * a))
        dphi = lambda a: float(c[1] + 2 * c[2] * a + 3 * c[3] * a * a + 4 * c[4] * a ** 3 + 0.9 * np.cos(3 * a))
        stp0 = float(rng.uniform(0.05, 3.0))
        visited = []

        def phi_t(a):
            visited.append(a)
            return phi(a)
        ref = DCSRCH(phi_t, dphi, L.LS_FTOL, L.LS_GTOL, L.LS_XTOL, L.LS_STPMIN, L.LS_STPMAX)
        stp_ref, _, _, task_ref = ref(stp0, phi0=phi(0.0), derphi0=dphi(0.0), maxiter=30)
        ls = L.Dcsrch()
        assert ls.start(phi(0.0), dphi(0.0), stp0) == L.T_FG
        mine = []
        for _ in range(30):
            mine.append(ls.stp)
            if ls.step(phi(ls.stp), dphi(ls.stp)) != L.T_FG:
                break
        assert mine == visited[:len(mine)], (trial, mine, visited)
        if stp_ref is not None:
            assert ls.stp == stp_ref


def test_epochs_and_early_stop_rule():
    """methods.py:164-181: `epochs` fresh minimisations from the current iterate; stop after an epoch with < 50 evaluations."""
    fn = make_quad(60, 9)
    x0 = np.random.default_rng(2).standard_normal(60)
    x, mach, pts = L.minimize(fn, x0, maxiter=30, epochs=5, min_evals=20)
    # what the reference's loop does with SciPy
    xs, used = x0.copy(), []
    for ep in range(5):
        cnt = [0]

        def f(x):
            cnt[0] += 1
            return fn(x)
        r = minimize(f, xs, jac=True, method="L-BFGS-B", options={"maxiter": 30})
        xs = r.x
        used.append(cnt[0])
        if cnt[0] < 20:
            break
    assert mach.epoch_nfev == used
    # ~150 evaluations on an ill-conditioned quadratic: rounding differences (two-loop vs compact form) have grown
    assert np.linalg.norm(x - xs) <= 1e-3 * np.linalg.norm(xs)
    assert abs(fn(x)[0] - fn(xs)[0]) <= 1e-6 * abs(fn(xs)[0])
    assert len(pts) == sum(used)


def test_float32_storage_tracks_float64():
    """The device keeps x_trial, g, d, S, Y in fp32: the same machine with store=float32 must follow the fp64 one."""
    fn = make_quad(40, 5)
    x0 = np.random.default_rng(11).standard_normal(40)
    x64, m64, _ = L.minimize(fn, x0, maxiter=40)
    x32, m32, _ = L.minimize(lambda x: fn(x.astype(np.float64)), x0, maxiter=40, store=np.float32)
    f64, f32 = fn(x64)[0], fn(x32)[0]
    assert abs(f32 - f64) <= 1e-4 * abs(f64)
    assert abs(m32.nfev_total - m64.nfev_total) <= 3
