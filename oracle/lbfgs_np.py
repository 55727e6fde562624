"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement of the optimiser that drives the step in the reference:
`tf.contrib.opt.ScipyOptimizerInterface(loss, method='L-BFGS-B', options={'maxiter': 100})`
(methods.py:133-137), minimised once per epoch from the variable's current value (methods.py:164-181),
stopping after an epoch that used fewer than 50 evaluations (methods.py:180-181).

The arithmetic lives in a third-party dependency that is not in /root/reference: SciPy's L-BFGS-B
(Zhu, Byrd, Lu, Nocedal, "Algorithm 778", v3.0) with the MINPACK-2 line search `dcsrch`/`dcstep`
(More' & Thuente 1994).  The reference passes NO bounds, so the generalized Cauchy point and the subspace
minimisation reduce to the quasi-Newton direction d = -H g with H the limited-memory BFGS inverse of
B0 = theta I; this file restates that published algorithm as an explicit state machine
(one `advance(f, g)` per function evaluation) because that is the shape the device code has
(audio_style_transfer_b200/csrc/lbfgs.cu): the host only pumps evaluations, all decisions are taken
where f and g already are.

Pinned against the dependency itself (tests/test_lbfgs_oracle.py): the sequence of evaluation points
and the result must equal `scipy.optimize.minimize(method='L-BFGS-B')` (SciPy 1.18 here) on quadratic,
Rosenbrock and a non-smooth-ish test function, and `dcsrch` must equal `scipy.optimize._dcsrch.DCSRCH`
step by step.

SciPy defaults used by the reference (it sets only maxiter): m = 10, factr = 1e7 (ftol 2.22e-9),
pgtol = 1e-5, maxls = 20, maxfun = 15000.
"""
from __future__ import annotations

import numpy as np

EPSMCH = np.finfo(np.float64).eps
# line-search constants of lnsrlb (L-BFGS-B 3.0)
LS_FTOL, LS_GTOL, LS_XTOL, LS_STPMIN, LS_STPMAX = 1e-3, 0.9, 0.1, 0.0, 1e10

# dcsrch tasks
T_FG, T_CONV, T_WARN, T_ERROR = 0, 1, 2, 3


def dcstep(stx, fx, dx, sty, fy, dy, stp, fp, dp, brackt, stpmin, stpmax):
    """One safeguarded cubic/quadratic step of More'-Thuente; returns the updated interval and trial step."""
    sgnd = dp * (dx / abs(dx))
    if fp > fx:                                               # case 1: higher value -> minimum bracketed
        theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp
        s = max(abs(theta), abs(dx), abs(dp))
        gamma = s * np.sqrt((theta / s) ** 2 - (dx / s) * (dp / s))
        if stp < stx:
            gamma = -gamma
        p = (gamma - dx) + theta
        q = ((gamma - dx) + gamma) + dp
        r = p / q
        stpc = stx + r * (stp - stx)
        stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx)
        stpf = stpc if abs(stpc - stx) < abs(stpq - stx) else stpc + (stpq - stpc) / 2.0
        brackt = True
    elif sgnd < 0.0:                                          # case 2: lower value, derivative changed sign
        theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp
        s = max(abs(theta), abs(dx), abs(dp))
        gamma = s * np.sqrt((theta / s) ** 2 - (dx / s) * (dp / s))
        if stp > stx:
            gamma = -gamma
        p = (gamma - dp) + theta
        q = ((gamma - dp) + gamma) + dx
        r = p / q
        stpc = stp + r * (stx - stp)
        stpq = stp + (dp / (dp - dx)) * (stx - stp)
        stpf = stpc if abs(stpc - stp) > abs(stpq - stp) else stpq
        brackt = True
    elif abs(dp) < abs(dx):                                   # case 3: lower value, same sign, |derivative| shrinks
        theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp
        s = max(abs(theta), abs(dx), abs(dp))
        gamma = s * np.sqrt(max(0.0, (theta / s) ** 2 - (dx / s) * (dp / s)))
        if stp > stx:
            gamma = -gamma
        p = (gamma - dp) + theta
        q = (gamma + (dx - dp)) + gamma
        r = p / q
        if r < 0.0 and gamma != 0.0:
            stpc = stp + r * (stx - stp)
        elif stp > stx:
            stpc = stpmax
        else:
            stpc = stpmin
        stpq = stp + (dp / (dp - dx)) * (stx - stp)
        if brackt:
            stpf = stpc if abs(stpc - stp) < abs(stpq - stp) else stpq
            if stp > stx:
                stpf = min(stp + 0.66 * (sty - stp), stpf)
            else:
                stpf = max(stp + 0.66 * (sty - stp), stpf)
        else:
            stpf = stpc if abs(stpc - stp) > abs(stpq - stp) else stpq
            stpf = min(stpmax, stpf)
            stpf = max(stpmin, stpf)
    else:                                                     # case 4: lower value, same sign, |derivative| grows
        if brackt:
            theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp
            s = max(abs(theta), abs(dy), abs(dp))
            gamma = s * np.sqrt((theta / s) ** 2 - (dy / s) * (dp / s))
            if stp > sty:
                gamma = -gamma
            p = (gamma - dp) + theta
            q = ((gamma - dp) + gamma) + dy
            r = p / q
            stpf = stp + r * (sty - stp)
        elif stp > stx:
            stpf = stpmax
        else:
            stpf = stpmin
    if fp > fx:
        sty, fy, dy = stp, fp, dp
    else:
        if sgnd < 0.0:
            sty, fy, dy = stx, fx, dx
        stx, fx, dx = stp, fp, dp
    return stx, fx, dx, sty, fy, dy, stpf, brackt


class Dcsrch:
    """MINPACK-2 dcsrch as a resumable object: start(f0, g0, stp0) then step(f, g) per evaluation."""

    def __init__(self, ftol=LS_FTOL, gtol=LS_GTOL, xtol=LS_XTOL, stpmin=LS_STPMIN, stpmax=LS_STPMAX):
        self.ftol, self.gtol, self.xtol, self.stpmin, self.stpmax = ftol, gtol, xtol, stpmin, stpmax

    def start(self, f, g, stp):
        if stp < self.stpmin or stp > self.stpmax or g >= 0.0:
            self.stp = stp
            return T_ERROR
        self.brackt, self.stage = False, 1
        self.finit, self.ginit = f, g
        self.gtest = self.ftol * g
        self.width = self.stpmax - self.stpmin
        self.width1 = self.width / 0.5
        self.stx, self.fx, self.gx = 0.0, f, g
        self.sty, self.fy, self.gy = 0.0, f, g
        self.stmin, self.stmax = 0.0, stp + 4.0 * stp
        self.stp = stp
        return T_FG

    def step(self, f, g):
        stp = self.stp
        ftest = self.finit + stp * self.gtest
        if self.stage == 1 and f <= ftest and g >= 0.0:
            self.stage = 2
        task = T_FG
        if self.brackt and (stp <= self.stmin or stp >= self.stmax):
            task = T_WARN
        if self.brackt and self.stmax - self.stmin <= self.xtol * self.stmax:
            task = T_WARN
        if stp == self.stpmax and f <= ftest and g <= self.gtest:
            task = T_WARN
        if stp == self.stpmin and (f > ftest or g >= self.gtest):
            task = T_WARN
        if f <= ftest and abs(g) <= self.gtol * (-self.ginit):
            task = T_CONV
        if task != T_FG:
            return task
        if self.stage == 1 and f <= self.fx and f > ftest:
            gt = self.gtest
            fm, fxm, fym = f - stp * gt, self.fx - self.stx * gt, self.fy - self.sty * gt
            gm, gxm, gym = g - gt, self.gx - gt, self.gy - gt
            self.stx, fxm, gxm, self.sty, fym, gym, stp, self.brackt = dcstep(
                self.stx, fxm, gxm, self.sty, fym, gym, stp, fm, gm, self.brackt, self.stmin, self.stmax)
            self.fx, self.fy = fxm + self.stx * gt, fym + self.sty * gt
            self.gx, self.gy = gxm + gt, gym + gt
        else:
            self.stx, self.fx, self.gx, self.sty, self.fy, self.gy, stp, self.brackt = dcstep(
                self.stx, self.fx, self.gx, self.sty, self.fy, self.gy, stp, f, g, self.brackt, self.stmin, self.stmax)
        if self.brackt:
            if abs(self.sty - self.stx) >= 0.66 * self.width1:
                stp = self.stx + 0.5 * (self.sty - self.stx)
            self.width1 = self.width
            self.width = abs(self.sty - self.stx)
        if self.brackt:
            self.stmin, self.stmax = min(self.stx, self.sty), max(self.stx, self.sty)
        else:
            self.stmin = stp + 1.1 * (stp - self.stx)
            self.stmax = stp + 4.0 * (stp - self.stx)
        stp = min(max(stp, self.stpmin), self.stpmax)
        if (self.brackt and (stp <= self.stmin or stp >= self.stmax)) or \
                (self.brackt and self.stmax - self.stmin <= self.xtol * self.stmax):
            stp = self.stx
        self.stp = stp
        return T_FG


# machine phases
PH_START, PH_LS, PH_DONE = 0, 1, 2
# why an epoch ended
END_NONE, END_MAXITER, END_PGTOL, END_FTOL, END_ABNORMAL = 0, 1, 2, 3, 4


class LbfgsMachine:
    """The optimiser as a state machine.  `x_trial` is where the next evaluation is wanted; feed its
    (f, g) to `advance`.  `store` is the dtype the vectors x-trial / g / d / S / Y are kept in (float32 on the
    device; float64 here reproduces SciPy), scalars and the iterate x_k are float64."""

    def __init__(self, x0, m=10, maxiter=100, epochs=1, min_evals=50, factr=1e7, pgtol=1e-5, maxls=20,
                 store=np.float64):
        self.n = int(np.size(x0))
        self.m, self.maxiter, self.epochs, self.min_evals = m, maxiter, epochs, min_evals
        self.ftol, self.pgtol, self.maxls = factr * EPSMCH, pgtol, maxls
        self.store = store
        self.xk = np.array(x0, np.float64).reshape(-1)
        self.x_trial = self.xk.astype(store)
        self.S, self.Y = np.zeros((m, self.n), store), np.zeros((m, self.n), store)
        self.sy = np.zeros(m)
        self.col, self.head, self.theta = 0, 0, 1.0
        self.phase, self.epoch, self.iter, self.nfev, self.nfev_total = PH_START, 0, 0, 0, 0
        self.ls = Dcsrch()
        self.epoch_x, self.epoch_nfev, self.epoch_end, self.log = [], [], [], []
        self.fk = np.inf

    @property
    def done(self):
        return self.phase == PH_DONE

    # -- d = -H g, H = limited-memory BFGS inverse with H0 = I / theta (two-loop recursion) --------
    def _direction(self, g):
        q = g.astype(np.float64).copy()
        idx = [(self.head + i) % self.m for i in range(self.col)]     # oldest ... newest
        a = {}
        for i in reversed(idx):
            a[i] = float(np.dot(self.S[i].astype(np.float64), q)) / self.sy[i]
            q -= a[i] * self.Y[i].astype(np.float64)
        q *= 1.0 / self.theta
        for i in idx:
            b = float(np.dot(self.Y[i].astype(np.float64), q)) / self.sy[i]
            q += (a[i] - b) * self.S[i].astype(np.float64)
        return (-q).astype(self.store)

    def _begin_line_search(self):
        """lnsrlb entry: returns False when d is not a descent direction (info = -4)."""
        d64 = self.d.astype(np.float64)
        self.dnorm = float(np.sqrt(np.dot(d64, d64)))
        self.stp = min(1.0 / self.dnorm, LS_STPMAX) if self.iter == 0 else 1.0     # not boxed: first step 1/|d|
        gd = float(np.dot(self.gk.astype(np.float64), d64))
        self.gdold = self.gd = gd
        self.ifun = 0
        if gd >= 0.0:
            return False
        self.ls.start(self.fk, gd, self.stp)
        self.ifun = 1
        self._set_trial()
        return True

    def _set_trial(self):
        self.x_trial = (self.xk + self.ls.stp * self.d.astype(np.float64)).astype(self.store)

    def _new_direction(self):
        """Direction + line-search start, with L-BFGS-B's 'refresh the memory and restart' on failure."""
        while True:
            self.d = self._direction(self.gk)
            if self._begin_line_search():
                return True
            if self.col == 0:
                return False                                   # ABNORMAL_TERMINATION_IN_LNSRCH
            self.col, self.head, self.theta = 0, 0, 1.0

    def _end_epoch(self, why):
        self.epoch_x.append(self.xk.copy())
        self.epoch_nfev.append(self.nfev)
        self.epoch_end.append(why)
        self.epoch += 1
        if self.epoch >= self.epochs or self.nfev < self.min_evals:      # methods.py:180-181
            self.phase = PH_DONE
            return
        self.phase, self.nfev = PH_START, 0
        self.x_trial = self.xk.astype(self.store)

    def advance(self, f, g, losses=None):
        """Consume the evaluation at `x_trial`; afterwards `x_trial` is the next point wanted (or `done`)."""
        if self.phase == PH_DONE:
            return
        f = float(f)
        g = np.asarray(g, self.store).reshape(-1)
        self.nfev += 1
        self.nfev_total += 1
        self.log.append((f,) if losses is None else tuple(losses))
        if self.phase == PH_START:                              # a fresh minimize() call
            self.col, self.head, self.theta, self.iter = 0, 0, 1.0, 0
            self.fk, self.gk = f, g.copy()
            if float(np.max(np.abs(g))) <= self.pgtol:
                return self._end_epoch(END_PGTOL)
            self.phase = PH_LS
            if not self._new_direction():
                return self._end_epoch(END_ABNORMAL)
            return
        # ---- inside a line search ---------------------------------------------------------------
        d64 = self.d.astype(np.float64)
        gd = float(np.dot(g.astype(np.float64), d64))
        self.gd = gd
        task = self.ls.step(f, gd)
        if task == T_FG:
            if self.ifun >= self.maxls:                         # iback >= maxls: restore, refresh memory, restart
                if self.col == 0:
                    return self._end_epoch(END_ABNORMAL)
                self.col, self.head, self.theta = 0, 0, 1.0
                if not self._new_direction():
                    return self._end_epoch(END_ABNORMAL)
                return
            self.ifun += 1
            self._set_trial()
            return
        # ---- line search finished: x_trial becomes the iterate -------------------------------
        stp = self.ls.stp
        self.xk = self.xk + stp * d64
        self.iter += 1
        fold = self.fk
        y = (g.astype(np.float64) - self.gk.astype(np.float64))
        self.fk, self.gk = f, g.copy()
        if float(np.max(np.abs(g))) <= self.pgtol:
            return self._end_epoch(END_PGTOL)
        if (fold - f) <= self.ftol * max(abs(fold), abs(f), 1.0):
            return self._end_epoch(END_FTOL)
        ys = y.astype(self.store).astype(np.float64)           # the device takes y.y of the STORED (rounded) pair
        rr = float(np.dot(ys, ys))
        if stp == 1.0:
            dr, ddum = gd - self.gdold, -self.gdold
        else:
            dr, ddum = (gd - self.gdold) * stp, -self.gdold * stp * stp
        if dr > EPSMCH * ddum:                                  # curvature ok: store the pair
            if self.col < self.m:
                slot = (self.head + self.col) % self.m
                self.col += 1
            else:
                slot = self.head
                self.head = (self.head + 1) % self.m
            self.S[slot] = (stp * d64).astype(self.store)
            self.Y[slot] = y.astype(self.store)
            self.sy[slot] = dr
            self.theta = rr / dr
        if self.iter >= self.maxiter:
            return self._end_epoch(END_MAXITER)
        if not self._new_direction():
            return self._end_epoch(END_ABNORMAL)


def minimize(fun, x0, **kw):
    """Pump: fun(x) -> (f, g) until the machine is done.  Returns (x, machine, list of evaluation points)."""
    mach = LbfgsMachine(x0, **kw)
    pts = []
    while not mach.done:
        x = mach.x_trial.astype(np.float64)
        pts.append(x.copy())
        f, g = fun(x)
        mach.advance(f, g)
    return mach.xk.copy(), mach, pts
