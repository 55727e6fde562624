// lbfgs.cu -- the optimiser that calls the step, resident on the device.
//
// Replaces, for runs that do not need SciPy in the loop, the reference's
//   tf.contrib.opt.ScipyOptimizerInterface(loss, method='L-BFGS-B', options={'maxiter': 100})   methods.py:133-137
//   for ep in range(epochs): optim.minimize(sess, ...) ... if i_ < 50: break                       methods.py:164-181
// i.e. SciPy's L-BFGS-B (Algorithm 778 v3.0; no bounds are passed, so the direction is the limited-memory BFGS
// quasi-Newton step with B0 = theta I) and its MINPACK-2 line search dcsrch / dcstep, with SciPy's defaults
// (m 10, factr 1e7, pgtol 1e-5, maxls 20).  The algorithm is restated as a state machine: ONE kernel launch
// (`lbfgs_advance_kernel`) consumes the (losses, gradient) the step just produced at `x_trial` and leaves the next
// evaluation point in `x_trial`, so an optimisation run is  [step graph, advance] x n  enqueued by the host with
// no read-back in between: every decision (accept / continue line search / restart / next epoch / stop) is taken on
// the device.  CPU restatement of the same machine: oracle/lbfgs_np.py (pinned against SciPy itself).
//
// One thread-block CLUSTER per pair.  Each CTA owns a contiguous slice of the T samples; a thread owns the same
// elements in every pass, so vector passes need no barrier of their own.  Dot products: warp shuffle -> CTA ->
// every CTA writes its partial into every peer's shared memory (DSMEM) -> cluster barrier -> fixed-order sum, so all
// CTAs hold bit-identical scalars and run the scalar part (dcsrch, tests, two-loop coefficients) redundantly in
// fp64 without any broadcast.  Vectors are fp32 (x_trial, g, d, S, Y) except the iterate x_k and the two-loop work
// vector q, which are fp64.
#include "../../include/ast_b200.h"
#include "common.cuh"

#include <cooperative_groups.h>
#include <cstdio>
#include <cstring>
#include <vector>

namespace cg = cooperative_groups;

#define OPT_MAXM 16
#define OPT_MAXCL 16
#define OPT_THREADS 512

enum { PH_START = 0, PH_LS = 1, PH_DONE = 2 };
enum { END_NONE = 0, END_MAXITER = 1, END_PGTOL = 2, END_FTOL = 3, END_ABNORMAL = 4 };
enum { T_FG = 0, T_CONV = 1, T_WARN = 2, T_ERROR = 3 };

struct OptState {
    int phase, epoch, iter, nfev, nfev_total, col, head, ifun, last_end;
    int brackt, stage, pad;
    double theta, fk, dnorm, stp, gd, gdold;
    double finit, ginit, gtest, width, width1, stx, fx, gx, sty, fy, gy, stmin, stmax;
    double sy[OPT_MAXM];
};

struct OptParams {
    int T, B, m, maxiter, epochs, min_evals, maxls, max_log, keep_epoch_x;
    double ftol, pgtol;
    OptState* st;          // [B]
    float* x_trial;        // [B][T]  where the next evaluation is wanted
    const float* losses;   // [B][4]  written by the step
    const float* g;        // [B][T]  written by the step
    double* xk;            // [B][T]  current iterate
    float* gk;             // [B][T]  gradient at x_k
    float* d;              // [B][T]  search direction
    double* q;             // [B][T]  two-loop work vector
    float* S;              // [B][m][T]
    float* Y;              // [B][m][T]
    float* epoch_x;        // [B][epochs][T] or NULL
    int* epoch_nfev;       // [B][epochs]
    int* epoch_end;        // [B][epochs]
    float* log;            // [B][max_log][4]
};

struct RedBuf {
    double warp[OPT_THREADS / 32][4];
    double slot[2][OPT_MAXCL][4];
};

// Cluster-wide reduction of N values (bit j of maxmask: max instead of sum).  Same result in every thread of the
// cluster.  `par` alternates the slot buffer so that a fast CTA cannot overwrite values a slow one still reads.
template <int N>
__device__ __forceinline__ void cluster_reduce(double (&v)[N], unsigned maxmask, RedBuf* rb, int& par) {
    cg::cluster_group cl = cg::this_cluster();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < N; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double u = __shfl_xor_sync(0xffffffffu, v[j], o);
            v[j] = ((maxmask >> j) & 1) ? fmax(v[j], u) : v[j] + u;
        }
        if (lane == 0) rb->warp[warp][j] = v[j];
    }
    __syncthreads();
    if (threadIdx.x < N) {
        const int j = threadIdx.x;
        double a = rb->warp[0][j];
        for (int w = 1; w < OPT_THREADS / 32; ++w) a = ((maxmask >> j) & 1) ? fmax(a, rb->warp[w][j]) : a + rb->warp[w][j];
        const unsigned me = cl.block_rank(), n = cl.num_blocks();
        for (unsigned r = 0; r < n; ++r) *cl.map_shared_rank(&rb->slot[par][me][j], r) = a;
    }
    cl.sync();
    const unsigned n = cl.num_blocks();
#pragma unroll
    for (int j = 0; j < N; ++j) {
        double a = rb->slot[par][0][j];
        for (unsigned r = 1; r < n; ++r) a = ((maxmask >> j) & 1) ? fmax(a, rb->slot[par][r][j]) : a + rb->slot[par][r][j];
        v[j] = a;
    }
    par ^= 1;
}

// ---- MINPACK-2 dcstep / dcsrch (More' & Thuente), as L-BFGS-B's lnsrlb calls them: ftol 1e-3, gtol 0.9, xtol 0.1,
// stpmin 0, stpmax 1e10 ------------------------------------------------------------------------------------------
#define LS_FTOL 1e-3
#define LS_GTOL 0.9
#define LS_XTOL 0.1
#define LS_STPMIN 0.0
#define LS_STPMAX 1e10

__device__ void dcstep(double& stx, double& fx, double& dx, double& sty, double& fy, double& dy, double& stp,
                       double fp, double dp, int& brackt, double stpmin, double stpmax) {
    const double sgnd = dp * (dx / fabs(dx));
    double stpf;
    if (fp > fx) {
        const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
        const double s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
        double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
        if (stp < stx) gamma = -gamma;
        const double p = (gamma - dx) + theta, q = ((gamma - dx) + gamma) + dp, r = p / q;
        const double stpc = stx + r * (stp - stx);
        const double stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx);
        stpf = fabs(stpc - stx) < fabs(stpq - stx) ? stpc : stpc + (stpq - stpc) / 2.0;
        brackt = 1;
    } else if (sgnd < 0.0) {
        const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
        const double s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
        double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
        if (stp > stx) gamma = -gamma;
        const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dx, r = p / q;
        const double stpc = stp + r * (stx - stp);
        const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
        stpf = fabs(stpc - stp) > fabs(stpq - stp) ? stpc : stpq;
        brackt = 1;
    } else if (fabs(dp) < fabs(dx)) {
        const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
        const double s = fmax(fabs(theta), fmax(fabs(dx), fabs(dp)));
        double gamma = s * sqrt(fmax(0.0, (theta / s) * (theta / s) - (dx / s) * (dp / s)));
        if (stp > stx) gamma = -gamma;
        const double p = (gamma - dp) + theta, q = (gamma + (dx - dp)) + gamma, r = p / q;
        double stpc;
        if (r < 0.0 && gamma != 0.0) stpc = stp + r * (stx - stp);
        else if (stp > stx) stpc = stpmax;
        else stpc = stpmin;
        const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
        if (brackt) {
            stpf = fabs(stpc - stp) < fabs(stpq - stp) ? stpc : stpq;
            if (stp > stx) stpf = fmin(stp + 0.66 * (sty - stp), stpf);
            else stpf = fmax(stp + 0.66 * (sty - stp), stpf);
        } else {
            stpf = fabs(stpc - stp) > fabs(stpq - stp) ? stpc : stpq;
            stpf = fmin(stpmax, stpf);
            stpf = fmax(stpmin, stpf);
        }
    } else {
        if (brackt) {
            const double theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp;
            const double s = fmax(fabs(theta), fmax(fabs(dy), fabs(dp)));
            double gamma = s * sqrt((theta / s) * (theta / s) - (dy / s) * (dp / s));
            if (stp > sty) gamma = -gamma;
            const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dy, r = p / q;
            stpf = stp + r * (sty - stp);
        } else if (stp > stx) stpf = stpmax;
        else stpf = stpmin;
    }
    if (fp > fx) { sty = stp; fy = fp; dy = dp; }
    else {
        if (sgnd < 0.0) { sty = stx; fy = fx; dy = dx; }
        stx = stp; fx = fp; dx = dp;
    }
    stp = stpf;
}

__device__ void dcsrch_start(OptState& s, double f, double g, double stp) {
    s.brackt = 0; s.stage = 1;
    s.finit = f; s.ginit = g; s.gtest = LS_FTOL * g;
    s.width = LS_STPMAX - LS_STPMIN; s.width1 = s.width / 0.5;
    s.stx = 0.0; s.fx = f; s.gx = g;
    s.sty = 0.0; s.fy = f; s.gy = g;
    s.stmin = 0.0; s.stmax = stp + 4.0 * stp;
    s.stp = stp;
}

__device__ int dcsrch_step(OptState& s, double f, double g) {
    double stp = s.stp;
    const double ftest = s.finit + stp * s.gtest;
    if (s.stage == 1 && f <= ftest && g >= 0.0) s.stage = 2;
    int task = T_FG;
    if (s.brackt && (stp <= s.stmin || stp >= s.stmax)) task = T_WARN;
    if (s.brackt && s.stmax - s.stmin <= LS_XTOL * s.stmax) task = T_WARN;
    if (stp == LS_STPMAX && f <= ftest && g <= s.gtest) task = T_WARN;
    if (stp == LS_STPMIN && (f > ftest || g >= s.gtest)) task = T_WARN;
    if (f <= ftest && fabs(g) <= LS_GTOL * (-s.ginit)) task = T_CONV;
    if (task != T_FG) return task;
    if (s.stage == 1 && f <= s.fx && f > ftest) {
        const double gt = s.gtest;
        double fxm = s.fx - s.stx * gt, fym = s.fy - s.sty * gt, gxm = s.gx - gt, gym = s.gy - gt;
        dcstep(s.stx, fxm, gxm, s.sty, fym, gym, stp, f - stp * gt, g - gt, s.brackt, s.stmin, s.stmax);
        s.fx = fxm + s.stx * gt; s.fy = fym + s.sty * gt;
        s.gx = gxm + gt; s.gy = gym + gt;
    } else {
        dcstep(s.stx, s.fx, s.gx, s.sty, s.fy, s.gy, stp, f, g, s.brackt, s.stmin, s.stmax);
    }
    if (s.brackt) {
        if (fabs(s.sty - s.stx) >= 0.66 * s.width1) stp = s.stx + 0.5 * (s.sty - s.stx);
        s.width1 = s.width;
        s.width = fabs(s.sty - s.stx);
    }
    if (s.brackt) { s.stmin = fmin(s.stx, s.sty); s.stmax = fmax(s.stx, s.sty); }
    else { s.stmin = stp + 1.1 * (stp - s.stx); s.stmax = stp + 4.0 * (stp - s.stx); }
    stp = fmin(fmax(stp, LS_STPMIN), LS_STPMAX);
    if ((s.brackt && (stp <= s.stmin || stp >= s.stmax)) || (s.brackt && s.stmax - s.stmin <= LS_XTOL * s.stmax)) stp = s.stx;
    s.stp = stp;
    return T_FG;
}

// ---- the machine ------------------------------------------------------------------------------------------------
struct Slice { int lo, hi; };     // this CTA's samples; a thread walks lo + tid, lo + tid + blockDim, ...

#define FOR_OWN(i) for (int i = sl.lo + (int)threadIdx.x; i < sl.hi; i += OPT_THREADS)

// d = -H g_k (two-loop recursion, H0 = I / theta), then the line-search start of lnsrlb.  On a non-descent direction
// L-BFGS-B drops the memory and retries from steepest descent; with no memory left it gives up.  Returns false = gave up.
__device__ bool new_direction(const OptParams& P, OptState& s, const Slice sl, int pair, RedBuf* rb, int& par) {
    const size_t base = (size_t)pair * P.T;
    const float* gk = P.gk + base;
    double* q = P.q + base;
    float* d = P.d + base;
    const float* S = P.S + (size_t)pair * P.m * P.T;
    const float* Y = P.Y + (size_t)pair * P.m * P.T;
    for (;;) {
        double alpha[OPT_MAXM];
        // first loop, newest -> oldest; pass j applies the previous update and takes the next dot
        {
            double acc[1] = {0.0};
            if (s.col > 0) {
                const float* Sn = S + (size_t)((s.head + s.col - 1) % P.m) * P.T;
                FOR_OWN(i) { const double v = (double)gk[i]; q[i] = v; acc[0] += (double)Sn[i] * v; }
            } else {
                FOR_OWN(i) q[i] = (double)gk[i];
            }
            for (int j = s.col - 1; j >= 0; --j) {
                const int slot = (s.head + j) % P.m;
                cluster_reduce<1>(acc, 0u, rb, par);
                const double a = acc[0] / s.sy[slot];
                alpha[j] = a;
                const float* Yj = Y + (size_t)slot * P.T;
                acc[0] = 0.0;
                if (j > 0) {
                    const float* Sp = S + (size_t)((s.head + j - 1) % P.m) * P.T;
                    FOR_OWN(i) { const double v = q[i] - a * (double)Yj[i]; q[i] = v; acc[0] += (double)Sp[i] * v; }
                } else {                                     // last update of loop 1, scale by H0, first dot of loop 2
                    const float* Y0 = Y + (size_t)(s.head % P.m) * P.T;
                    const double it = 1.0 / s.theta;
                    FOR_OWN(i) { const double v = (q[i] - a * (double)Yj[i]) * it; q[i] = v; acc[0] += (double)Y0[i] * v; }
                }
            }
            if (s.col == 0) {
                const double it = 1.0 / s.theta;
                FOR_OWN(i) q[i] *= it;
            }
            // second loop, oldest -> newest
            for (int j = 0; j < s.col; ++j) {
                const int slot = (s.head + j) % P.m;
                cluster_reduce<1>(acc, 0u, rb, par);
                const double c = alpha[j] - acc[0] / s.sy[slot];
                const float* Sj = S + (size_t)slot * P.T;
                acc[0] = 0.0;
                if (j + 1 < s.col) {
                    const float* Yn = Y + (size_t)((s.head + j + 1) % P.m) * P.T;
                    FOR_OWN(i) { const double v = q[i] + c * (double)Sj[i]; q[i] = v; acc[0] += (double)Yn[i] * v; }
                } else {
                    FOR_OWN(i) q[i] += c * (double)Sj[i];
                }
            }
        }
        // d = -q (fp32), |d|^2 and g_k . d of the stored direction
        double r2[2] = {0.0, 0.0};
        FOR_OWN(i) {
            const float dv = (float)(-q[i]);
            d[i] = dv;
            r2[0] += (double)dv * (double)dv;
            r2[1] += (double)gk[i] * (double)dv;
        }
        cluster_reduce<2>(r2, 0u, rb, par);
        s.dnorm = sqrt(r2[0]);
        s.stp = s.iter == 0 ? fmin(1.0 / s.dnorm, LS_STPMAX) : 1.0;
        s.gd = s.gdold = r2[1];
        s.ifun = 0;
        if (r2[1] < 0.0) {
            dcsrch_start(s, s.fk, r2[1], s.stp);
            s.ifun = 1;
            return true;
        }
        if (s.col == 0) return false;
        s.col = 0; s.head = 0; s.theta = 1.0;
    }
}

__device__ void set_trial(const OptParams& P, const OptState& s, const Slice sl, int pair) {
    const size_t base = (size_t)pair * P.T;
    const double stp = s.stp;
    FOR_OWN(i) P.x_trial[base + i] = (float)(P.xk[base + i] + stp * (double)P.d[base + i]);
}

__device__ void end_epoch(const OptParams& P, OptState& s, const Slice sl, int pair, int why, bool lead) {
    const size_t base = (size_t)pair * P.T;
    if (P.epoch_x && P.keep_epoch_x) {
        float* ex = P.epoch_x + ((size_t)pair * P.epochs + s.epoch) * P.T;
        FOR_OWN(i) ex[i] = (float)P.xk[base + i];
    }
    if (lead) {
        P.epoch_nfev[pair * P.epochs + s.epoch] = s.nfev;
        P.epoch_end[pair * P.epochs + s.epoch] = why;
    }
    s.last_end = why;
    s.epoch += 1;
    if (s.epoch >= P.epochs || s.nfev < P.min_evals) {        // methods.py:180-181
        s.phase = PH_DONE;
        return;
    }
    s.phase = PH_START;
    s.nfev = 0;
    FOR_OWN(i) P.x_trial[base + i] = (float)P.xk[base + i];   // a fresh minimize() evaluates its starting point
}

__global__ void __launch_bounds__(OPT_THREADS) lbfgs_advance_kernel(const OptParams P) {
    __shared__ RedBuf rb;
    cg::cluster_group cl = cg::this_cluster();
    const int ncl = (int)cl.num_blocks(), rank = (int)cl.block_rank();
    const int pair = blockIdx.x / ncl;
    const bool lead = rank == 0 && threadIdx.x == 0;
    OptState s = P.st[pair];
    cl.sync();                                   // everyone has read the state before the leader may rewrite it
    if (s.phase == PH_DONE) return;
    int par = 0;
    const int chunk = (P.T + ncl - 1) / ncl;
    Slice sl;
    sl.lo = rank * chunk;
    sl.hi = min(P.T, sl.lo + chunk);
    const size_t base = (size_t)pair * P.T;
    const float* g = P.g + base;
    const double f = (double)P.losses[pair * 4];
    if (lead && s.nfev_total < P.max_log) {
        float4 lo = *reinterpret_cast<const float4*>(P.losses + pair * 4);
        *reinterpret_cast<float4*>(P.log + ((size_t)pair * P.max_log + s.nfev_total) * 4) = lo;
    }
    s.nfev += 1;
    s.nfev_total += 1;

    if (s.phase == PH_START) {
        s.col = 0; s.head = 0; s.theta = 1.0; s.iter = 0;
        s.fk = f;
        double gm[1] = {0.0};
        FOR_OWN(i) { const float v = g[i]; P.gk[base + i] = v; gm[0] = fmax(gm[0], (double)fabsf(v)); }
        cluster_reduce<1>(gm, 1u, &rb, par);
        if (gm[0] <= P.pgtol) end_epoch(P, s, sl, pair, END_PGTOL, lead);
        else {
            s.phase = PH_LS;
            if (new_direction(P, s, sl, pair, &rb, par)) set_trial(P, s, sl, pair);
            else end_epoch(P, s, sl, pair, END_ABNORMAL, lead);
        }
    } else {
        double gd[1] = {0.0};
        FOR_OWN(i) gd[0] += (double)g[i] * (double)P.d[base + i];
        cluster_reduce<1>(gd, 0u, &rb, par);
        s.gd = gd[0];
        const int task = dcsrch_step(s, f, gd[0]);
        if (task == T_FG) {
            if (s.ifun >= P.maxls) {                          // lnsrlb: iback >= maxls -> restore, refresh memory, restart
                bool ok = s.col > 0;
                if (ok) {
                    s.col = 0; s.head = 0; s.theta = 1.0;
                    ok = new_direction(P, s, sl, pair, &rb, par);
                }
                if (ok) set_trial(P, s, sl, pair);
                else end_epoch(P, s, sl, pair, END_ABNORMAL, lead);
            } else {
                s.ifun += 1;
                set_trial(P, s, sl, pair);
            }
        } else {
            // line search finished: the trial point becomes the iterate
            const double stp = s.stp;
            double dr, ddum;
            if (stp == 1.0) { dr = s.gd - s.gdold; ddum = -s.gdold; }
            else { dr = (s.gd - s.gdold) * stp; ddum = -s.gdold * stp * stp; }
            const bool store = dr > 2.220446049250313e-16 * ddum;
            int slot = 0;
            if (store) slot = s.col < P.m ? (s.head + s.col) % P.m : s.head;
            float* Ss = P.S + ((size_t)pair * P.m + slot) * P.T;
            float* Ys = P.Y + ((size_t)pair * P.m + slot) * P.T;
            double r2[2] = {0.0, 0.0};                         // y.y, max |g|
            FOR_OWN(i) {
                const double sd = stp * (double)P.d[base + i];
                P.xk[base + i] += sd;
                const float gn = g[i];
                const double y = (double)gn - (double)P.gk[base + i];
                P.gk[base + i] = gn;
                if (store) { Ss[i] = (float)sd; Ys[i] = (float)y; }
                r2[0] += y * y;
                r2[1] = fmax(r2[1], (double)fabsf(gn));
            }
            cluster_reduce<2>(r2, 2u, &rb, par);
            s.iter += 1;
            const double fold = s.fk;
            s.fk = f;
            if (r2[1] <= P.pgtol) end_epoch(P, s, sl, pair, END_PGTOL, lead);
            else if ((fold - f) <= P.ftol * fmax(fmax(fabs(fold), fabs(f)), 1.0)) end_epoch(P, s, sl, pair, END_FTOL, lead);
            else {
                if (store) {
                    if (s.col < P.m) s.col += 1;
                    else s.head = (s.head + 1) % P.m;
                    s.sy[slot] = dr;
                    s.theta = r2[0] / dr;
                }
                if (s.iter >= P.maxiter) end_epoch(P, s, sl, pair, END_MAXITER, lead);
                else if (new_direction(P, s, sl, pair, &rb, par)) set_trial(P, s, sl, pair);
                else end_epoch(P, s, sl, pair, END_ABNORMAL, lead);
            }
        }
    }
    if (lead) P.st[pair] = s;
}

__global__ void lbfgs_reset_kernel(const OptParams P, const double* x0) {
    const size_t n = (size_t)P.B * P.T;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const double v = x0[i];
        P.xk[i] = v;
        P.x_trial[i] = (float)v;
    }
    if (blockIdx.x == 0 && threadIdx.x < P.B) {
        OptState s;
        memset(&s, 0, sizeof(s));
        s.phase = PH_START;
        s.theta = 1.0;
        P.st[threadIdx.x] = s;
    }
}

// ---- C ABI ----------------------------------------------------------------------------------------------------------
static thread_local char o_err[256] = "";
extern "C" const char* ast_last_error(void);
int ast_set_error(int code, const char* msg);      // api.cu: routes into ast_last_error()

#define OCK(call)                                                                              \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            snprintf(o_err, sizeof(o_err), "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return ast_set_error(AST_ECUDA, o_err);                                            \
        }                                                                                      \
    } while (0)

struct ast_opt {
    ast_opt_config cfg;
    OptParams P;
    int cluster;
    float *d_losses, *d_grad;
    double* d_x0;
    std::vector<void*> allocs;
    int64_t bytes;
    OptState* h_st;            // pinned, 2 slots of B states (status polling of the pump)
    cudaEvent_t ev[2];
    cudaGraphExec_t gexec;     // [step, advance] captured for (ctx, this)
    ast_ctx* gctx;
    cudaStream_t stream;
};

template <typename Tp>
static int oalloc(ast_opt* o, Tp** p, size_t n) {
    void* q = nullptr;
    const size_t b = n * sizeof(Tp);
    if (cudaMalloc(&q, b ? b : 16) != cudaSuccess) return ast_set_error(AST_ECUDA, "cudaMalloc failed (optimiser state)");
    cudaMemset(q, 0, b ? b : 16);
    o->allocs.push_back(q);
    o->bytes += (int64_t)b;
    *p = (Tp*)q;
    return AST_OK;
}

extern "C" int ast_opt_destroy(ast_opt* o) {
    if (!o) return AST_OK;
    cudaSetDevice(o->cfg.device);
    cudaDeviceSynchronize();
    if (o->gexec) cudaGraphExecDestroy(o->gexec);
    for (void* p : o->allocs) cudaFree(p);
    if (o->h_st) cudaFreeHost(o->h_st);
    for (auto& e : o->ev) if (e) cudaEventDestroy(e);
    if (o->stream) cudaStreamDestroy(o->stream);
    delete o;
    return AST_OK;
}

extern "C" int ast_opt_create(const ast_opt_config* cfg, ast_opt** out) {
    if (!cfg || !out) return ast_set_error(AST_EINVAL, "null argument");
    if (cfg->T <= 0 || cfg->B <= 0 || cfg->B > 1024) return ast_set_error(AST_EINVAL, "ast_opt_create: bad T or B");
    ast_opt_config c = *cfg;
    if (c.m <= 0) c.m = 10;
    if (c.m > OPT_MAXM) return ast_set_error(AST_EINVAL, "ast_opt_create: m > 16");
    if (c.maxiter <= 0) c.maxiter = 100;
    if (c.epochs <= 0) c.epochs = 1;
    if (c.min_evals < 0) c.min_evals = 0;
    if (c.maxls <= 0) c.maxls = 20;
    if (c.factr <= 0) c.factr = 1e7;
    if (c.pgtol < 0) c.pgtol = 1e-5;
    if (c.max_log < 0) c.max_log = 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || c.device < 0 || c.device >= ndev)
        return ast_set_error(AST_ECUDA, "ast_opt_create: no such CUDA device (there is no CPU fallback)");
    OCK(cudaSetDevice(c.device));
    ast_opt* o = new ast_opt();
    o->cfg = c;
    o->bytes = 0; o->h_st = nullptr; o->gexec = nullptr; o->gctx = nullptr; o->stream = nullptr;
    o->ev[0] = o->ev[1] = nullptr;
    const size_t n = (size_t)c.B * c.T;
    OptParams& P = o->P;
    memset(&P, 0, sizeof(P));
    P.T = c.T; P.B = c.B; P.m = c.m; P.maxiter = c.maxiter; P.epochs = c.epochs; P.min_evals = c.min_evals;
    P.maxls = c.maxls; P.max_log = c.max_log; P.keep_epoch_x = c.keep_epoch_x;
    P.ftol = c.factr * 2.220446049250313e-16; P.pgtol = c.pgtol;
    int r = AST_OK;
    float* losses = nullptr; float* grad = nullptr;
    if ((r = oalloc(o, &P.st, (size_t)c.B)) || (r = oalloc(o, &P.x_trial, n)) || (r = oalloc(o, &losses, (size_t)c.B * 4)) ||
        (r = oalloc(o, &grad, n)) || (r = oalloc(o, &P.xk, n)) || (r = oalloc(o, &P.gk, n)) || (r = oalloc(o, &P.d, n)) ||
        (r = oalloc(o, &P.q, n)) || (r = oalloc(o, &P.S, n * c.m)) || (r = oalloc(o, &P.Y, n * c.m)) ||
        (r = oalloc(o, &P.epoch_nfev, (size_t)c.B * c.epochs)) || (r = oalloc(o, &P.epoch_end, (size_t)c.B * c.epochs)) ||
        (r = oalloc(o, &P.log, (size_t)c.B * (c.max_log ? c.max_log : 1) * 4)) || (r = oalloc(o, &o->d_x0, n)) ||
        (c.keep_epoch_x && (r = oalloc(o, &P.epoch_x, n * c.epochs)))) {
        ast_opt_destroy(o);
        return r;
    }
    P.losses = losses; P.g = grad;
    o->d_losses = losses; o->d_grad = grad;
    // one cluster per pair: as many CTAs as give every thread work, at most the portable 8
    int cl = 8;
    while (cl > 1 && (c.T + cl - 1) / cl < OPT_THREADS) cl >>= 1;
    o->cluster = cl;
    if (cudaMallocHost((void**)&o->h_st, 2 * sizeof(OptState) * c.B) != cudaSuccess ||
        cudaEventCreateWithFlags(&o->ev[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&o->ev[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaStreamCreateWithFlags(&o->stream, cudaStreamNonBlocking) != cudaSuccess) {
        ast_opt_destroy(o);
        return ast_set_error(AST_ECUDA, "ast_opt_create: pinned memory / event / stream creation failed");
    }
    *out = o;
    return AST_OK;
}

extern "C" int64_t ast_opt_workspace_bytes(const ast_opt* o) { return o ? o->bytes : 0; }

extern "C" int ast_opt_reset(ast_opt* o, const double* p_x0, void* stream) {
    if (!o || !p_x0) return ast_set_error(AST_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    OCK(cudaSetDevice(o->cfg.device));
    const size_t n = (size_t)o->cfg.B * o->cfg.T;
    OCK(cudaMemcpyAsync(o->d_x0, p_x0, n * sizeof(double), cudaMemcpyDefault, s));
    lbfgs_reset_kernel<<<64, 256, 0, s>>>(o->P, o->d_x0);
    ++ast_launch_counter;
    OCK(cudaGetLastError());
    return AST_OK;
}

extern "C" int ast_opt_buffers(ast_opt* o, float** d_x, float** d_losses, float** d_grad) {
    if (!o) return ast_set_error(AST_EINVAL, "null argument");
    if (d_x) *d_x = o->P.x_trial;
    if (d_losses) *d_losses = o->d_losses;
    if (d_grad) *d_grad = o->d_grad;
    return AST_OK;
}

extern "C" int ast_opt_advance(ast_opt* o, void* stream) {
    if (!o) return ast_set_error(AST_EINVAL, "null argument");
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(o->cluster * o->cfg.B));
    cfg.blockDim = dim3(OPT_THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)o->cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    OCK(cudaLaunchKernelEx(&cfg, lbfgs_advance_kernel, o->P));
    ++ast_launch_counter;
    return AST_OK;
}

static void fill_status(const OptState* st, int B, int32_t* out) {
    for (int b = 0; b < B; ++b) {
        int32_t* r = out + 8 * b;
        r[0] = st[b].phase; r[1] = st[b].epoch; r[2] = st[b].iter; r[3] = st[b].nfev;
        r[4] = st[b].nfev_total; r[5] = st[b].last_end; r[6] = st[b].col; r[7] = st[b].ifun;
    }
}

extern "C" int ast_opt_status(ast_opt* o, int32_t* h_status, double* h_f, void* stream) {
    if (!o || !h_status) return ast_set_error(AST_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    OCK(cudaMemcpyAsync(o->h_st, o->P.st, sizeof(OptState) * o->cfg.B, cudaMemcpyDeviceToHost, s));
    OCK(cudaStreamSynchronize(s));
    fill_status(o->h_st, o->cfg.B, h_status);
    if (h_f) for (int b = 0; b < o->cfg.B; ++b) h_f[b] = o->h_st[b].fk;
    return AST_OK;
}

extern "C" int ast_opt_result(ast_opt* o, double* h_x, void* stream) {
    if (!o || !h_x) return ast_set_error(AST_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    OCK(cudaMemcpyAsync(h_x, o->P.xk, sizeof(double) * o->cfg.B * (size_t)o->cfg.T, cudaMemcpyDeviceToHost, s));
    OCK(cudaStreamSynchronize(s));
    return AST_OK;
}

extern "C" int ast_opt_epochs(ast_opt* o, int32_t* h_nfev, int32_t* h_end, float* h_x, void* stream) {
    if (!o) return ast_set_error(AST_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t ne = (size_t)o->cfg.B * o->cfg.epochs;
    if (h_nfev) OCK(cudaMemcpyAsync(h_nfev, o->P.epoch_nfev, ne * sizeof(int), cudaMemcpyDeviceToHost, s));
    if (h_end) OCK(cudaMemcpyAsync(h_end, o->P.epoch_end, ne * sizeof(int), cudaMemcpyDeviceToHost, s));
    if (h_x) {
        if (!o->P.epoch_x) return ast_set_error(AST_ESTATE, "ast_opt_epochs: created without keep_epoch_x");
        OCK(cudaMemcpyAsync(h_x, o->P.epoch_x, ne * o->cfg.T * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    OCK(cudaStreamSynchronize(s));
    return AST_OK;
}

extern "C" int ast_opt_log(ast_opt* o, float* h_log, void* stream) {
    if (!o || !h_log) return ast_set_error(AST_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    OCK(cudaMemcpyAsync(h_log, o->P.log, (size_t)o->cfg.B * o->cfg.max_log * 4 * sizeof(float), cudaMemcpyDeviceToHost, s));
    OCK(cudaStreamSynchronize(s));
    return AST_OK;
}

// The pump: [ast_step on x_trial -> (losses, grad)] + [advance] captured once as ONE graph; replayed `chunk` times
// between status read-backs.  The next chunk is enqueued BEFORE the previous chunk's status is waited for, so the GPU
// never idles on the host; at most 2 * chunk - 1 evaluations run after the machine has stopped (they are no-ops for
// the optimiser state).
extern "C" int ast_minimize(ast_ctx* ctx, ast_opt* o, int64_t max_evals, int32_t chunk, int64_t* evals_enqueued) {
    if (!ctx || !o) return ast_set_error(AST_EINVAL, "null argument");
    if (chunk <= 0) chunk = 4;
    if (max_evals <= 0) max_evals = (int64_t)o->cfg.epochs * 15000;
    OCK(cudaSetDevice(o->cfg.device));
    cudaStream_t s = o->stream;
    if (o->gexec && o->gctx != ctx) { cudaGraphExecDestroy(o->gexec); o->gexec = nullptr; }
    if (!o->gexec) {
        cudaGraph_t graph = nullptr;
        OCK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        int r = ast_step(ctx, o->P.x_trial, o->d_losses, o->d_grad, s);
        if (r == AST_OK) r = ast_opt_advance(o, s);
        const cudaError_t e = cudaStreamEndCapture(s, &graph);
        if (r != AST_OK) { if (graph) cudaGraphDestroy(graph); return r; }
        if (e != cudaSuccess) return ast_set_error(AST_ECUDA, "ast_minimize: graph capture failed");
        const cudaError_t e2 = cudaGraphInstantiate(&o->gexec, graph, 0);
        cudaGraphDestroy(graph);
        if (e2 != cudaSuccess) { o->gexec = nullptr; return ast_set_error(AST_ECUDA, "ast_minimize: graph instantiate failed"); }
        o->gctx = ctx;
    }
    int64_t n = 0;
    int slot = 0;
    bool pending = false, done = false;
    while (!done && n < max_evals) {
        for (int k = 0; k < chunk && n < max_evals; ++k, ++n) OCK(cudaGraphLaunch(o->gexec, s));
        OCK(cudaMemcpyAsync(o->h_st + (size_t)slot * o->cfg.B, o->P.st, sizeof(OptState) * o->cfg.B, cudaMemcpyDeviceToHost, s));
        OCK(cudaEventRecord(o->ev[slot], s));
        if (pending) {
            OCK(cudaEventSynchronize(o->ev[slot ^ 1]));
            done = true;
            for (int b = 0; b < o->cfg.B; ++b) done = done && o->h_st[(size_t)(slot ^ 1) * o->cfg.B + b].phase == PH_DONE;
        }
        pending = true;
        slot ^= 1;
    }
    OCK(cudaStreamSynchronize(s));
    if (evals_enqueued) *evals_enqueued = n;
    return AST_OK;
}
