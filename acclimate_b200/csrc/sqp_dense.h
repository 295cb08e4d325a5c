// sqp_dense.h -- sequential quadratic programming with a DENSE BFGS model, for small problems whose objective is not
// separable: the consumers' nested-CES utility (reference src/model/Consumer.cpp:170-253, at most one good per sector,
// n <= a few dozen) with box bounds and one general constraint.
//
// Companion of sqp_core.h (diagonal model, separable purchasing objective).  Same skeleton -- Kraft's SLSQP as NLopt
// drives it: B starts at the identity, one QP sub-problem per iteration, L1 merit function, Armijo line search with
// quadratic interpolation floored at 0.1 and at most 10 backtracks, at most 5 resets of B on a non-descent direction,
// NLopt's xtol_abs / maxeval rules and best-point bookkeeping -- with the model Kraft uses: a full matrix updated by
// Powell's damped BFGS formula on the gradient of the Lagrangian.  The QP (dense B, box, one linear constraint) is
// solved by an active-set method on the bounds: free variables from the KKT system through a Cholesky factor of B_FF,
// every bound violator is clamped and fixed, a fixed variable whose multiplier has the wrong sign is released (worst
// first), until nothing changes.
//
// Execution model.  The routines are written for a GROUP of threads (one warp on the device) that all run the same
// scalar control flow on the same values; `par` says how loops are shared out:
//   par.foreach(m, f)   f(0..m-1), iterations independent of each other, followed by a group barrier
//   par.single(f)       f() by one thread, fenced by group barriers (sequential recurrences)
// Memory is only ever written inside foreach (one thread per element) or single; everything a thread decides on is read
// after a barrier, so all threads decide alike.  Element-wise work (the BFGS update, matrix-vector products, the
// columns of the Cholesky factor, the bound multipliers) is spread over the group; every ELEMENT is still computed by
// one thread with the same operations in the same order as a plain sequential loop, so the host (SerialPar: oracle/
// nlopt_shim, tests) and the device produce identical bits.
#pragma once

#include "sqp_core.h"

namespace accsqp {

struct SerialPar {
    template <class F> SQP_HD void foreach(int m, F f) const { for (int i = 0; i < m; ++i) f(i); }
    template <class F> SQP_HD void single(F f) const { f(); }
};

struct DenseWork {
    int n;
    double *x, *x0, *g, *a, *d, *gl, *xbest, *t1, *t2, *t3;  // length n
    double *B, *L;                                           // n * n: model, Cholesky factor of its free block
    double* scal;                                            // 4 scalars handed from single() to the group (F, c, pivot)
    int* state;                                              // 0 free, -1 fixed at the lower step bound, +1 at the upper
    const double *xl, *xu, *xtol;
};
// doubles and ints a DenseWork needs
SQP_HD size_t dense_doubles(int n) { return (size_t)10 * n + (size_t)2 * n * n + 4; }

SQP_HD double dlo(const DenseWork& w, int i) { return w.xl[i] - w.x[i]; }
SQP_HD double dhi(const DenseWork& w, int i) { return w.xu[i] - w.x[i]; }

// Cholesky factor of the free block of B into L (lower triangle, free variables in index order); false if not PD.
// Column by column: the diagonal entry (one thread), then the entries below it side by side.  Each entry is
// B_pq - sum_{k<q} L_pk L_qk folded in increasing k, as in the row-by-row form.
template <class Par>
SQP_HD bool dense_factor(Par& par, DenseWork& w, const int* fr, int nf) {
    const int n = w.n;
    for (int q = 0; q < nf; ++q) {
        par.single([&] {
            double s = w.B[fr[q] * n + fr[q]];
            for (int k = 0; k < q; ++k) s -= w.L[q * n + k] * w.L[q * n + k];
            w.scal[2] = s;
            if (s > 0.0) w.L[q * n + q] = sqrt(s);
        });
        if (!(w.scal[2] > 0.0)) return false;
        par.foreach(nf - q - 1, [&](int t) {
            const int p = q + 1 + t;
            double s = w.B[fr[p] * n + fr[q]];
            for (int k = 0; k < q; ++k) s -= w.L[p * n + k] * w.L[q * n + k];
            w.L[p * n + q] = s / w.L[q * n + q];
        });
    }
    return true;
}
// solves (L L^T) y = rhs in place (rhs indexed by free position); a sequential recurrence: one thread
SQP_HD void dense_solve(const DenseWork& w, int nf, double* y) {
    const int n = w.n;
    for (int p = 0; p < nf; ++p) {
        double s = y[p];
        for (int k = 0; k < p; ++k) s -= w.L[p * n + k] * y[k];
        y[p] = s / w.L[p * n + p];
    }
    for (int p = nf - 1; p >= 0; --p) {
        double s = y[p];
        for (int k = p + 1; k < nf; ++k) s -= w.L[k * n + p] * y[k];
        y[p] = s / w.L[p * n + p];
    }
}

// QP sub-problem  min 1/2 d^T B d + g^T d  s.t.  a^T d + c (= | >=) 0,  lo <= d <= hi.  Returns the multiplier of the
// general constraint (B d + g = lambda a + z); w.d receives the step.  *ok = false if B lost positive definiteness.
template <class Par>
SQP_HD double qp_dense(Par& par, DenseWork& w, int ctype, double c, Stats& st, bool* ok) {
    const int n = w.n;
    *ok = true;
    int* fr = w.state + n;  // scratch: list of free variables (the state array has 2 n entries)
    bool use_con = ctype == CON_EQ;
    double lambda = 0.0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        par.foreach(n, [&](int i) { w.state[i] = (dlo(w, i) == dhi(w, i)) ? -2 : 0; w.d[i] = (w.state[i] == -2) ? dlo(w, i) : 0.0; });
        const int max_pass = 6 * n + 10;
        for (int pass = 0; pass < max_pass; ++pass) {
            ++st.qp_iters;
            int nf = 0;
            for (int i = 0; i < n; ++i) if (w.state[i] == 0) ++nf;
            par.single([&] { int m = 0; for (int i = 0; i < n; ++i) if (w.state[i] == 0) fr[m++] = i; });
            // right-hand sides over the free set: u = B_FF^-1 (g_F + B_FW d_W), v = B_FF^-1 a_F
            double cw = c;
            for (int i = 0; i < n; ++i) if (w.state[i] != 0) cw += w.a[i] * w.d[i];
            par.foreach(nf, [&](int p) {
                double s = w.g[fr[p]];
                for (int j = 0; j < n; ++j) if (w.state[j] != 0) s += w.B[fr[p] * n + j] * w.d[j];
                w.t1[p] = s;
                w.t2[p] = w.a[fr[p]];
            });
            if (nf > 0) {
                if (!dense_factor(par, w, fr, nf)) { *ok = false; return 0.0; }
                par.foreach(use_con ? 2 : 1, [&](int which) { dense_solve(w, nf, which == 0 ? w.t1 : w.t2); });  // two right-hand sides, two threads
            }
            lambda = 0.0;
            if (use_con) {
                double au = 0.0, av = 0.0;
                for (int p = 0; p < nf; ++p) { au += w.a[fr[p]] * w.t1[p]; av += w.a[fr[p]] * w.t2[p]; }
                if (av > 0.0) lambda = (au - cw) / av;
            }
            bool clamped = false;
            for (int p = 0; p < nf; ++p) {
                const int i = fr[p];
                const double di = (use_con ? lambda * w.t2[p] : 0.0) - w.t1[p];
                if (di < dlo(w, i) || di > dhi(w, i)) clamped = true;
            }
            par.foreach(nf, [&](int p) {
                const int i = fr[p];
                const double di = (use_con ? lambda * w.t2[p] : 0.0) - w.t1[p];
                if (di < dlo(w, i)) { w.d[i] = dlo(w, i); w.state[i] = -1; }
                else if (di > dhi(w, i)) { w.d[i] = dhi(w, i); w.state[i] = 1; }
                else w.d[i] = di;
            });
            if (clamped) continue;
            // multipliers of the fixed variables: z = B d + g - lambda a; at a lower bound z >= 0, at an upper bound z <= 0
            par.foreach(n, [&](int i) {
                double viol = 0.0;
                if (w.state[i] == -1 || w.state[i] == 1) {
                    double z = w.g[i] - lambda * w.a[i];
                    for (int j = 0; j < n; ++j) z += w.B[i * n + j] * w.d[j];
                    viol = w.state[i] < 0 ? -z : z;
                }
                w.t3[i] = viol;
            });
            int worst = -1;
            double worst_v = 0.0;
            for (int i = 0; i < n; ++i) {
                if (w.state[i] != -1 && w.state[i] != 1) continue;
                const double viol = w.t3[i];
                const double tol = 1e-14 * (1.0 + fabs(w.g[i]) + fabs(lambda * w.a[i]));
                if (viol > tol && viol > worst_v) { worst_v = viol; worst = i; }
            }
            if (worst < 0) break;
            par.single([&] { w.state[worst] = 0; });
        }
        if (ctype != CON_INEQ) break;
        // inequality c(x) >= 0: solved without it first; if the step violates its linearisation, or the multiplier of the
        // equality form is negative, switch form once
        double lin = c;
        for (int i = 0; i < n; ++i) lin += w.a[i] * w.d[i];
        if (!use_con && lin < 0.0) { use_con = true; continue; }
        if (use_con && lambda < 0.0) { use_con = false; lambda = 0.0; continue; }
        break;
    }
    return lambda;
}

// Eval: ev(par, x, g, a, scal) -- the group evaluates objective (to minimise) and constraint at x, fills the gradients
// g and a, and leaves F in scal[0] and c in scal[1], visible to the whole group on return.
template <class Par, class Eval>
SQP_HD int sqp_minimize_dense(Par& par, Eval& ev, DenseWork& w, const Params& p, Stats& st) {
    const int n = w.n;
    st.iters = 0; st.evals = 0; st.qp_iters = 0; st.qp_inexact = 0; st.resets = 0; st.minf = 0.0; st.feasible = 0;
    par.foreach(n, [&](int i) {
        double xi = w.x[i];
        if (xi < w.xl[i]) xi = w.xl[i];
        if (xi > w.xu[i]) xi = w.xu[i];
        w.x[i] = xi;
    });
    auto identity = [&] { par.foreach(n * n, [&](int k) { w.B[k] = (k / n == k % n) ? 1.0 : 0.0; }); };
    identity();
    ev(par, w.x, w.g, w.a, w.scal);
    double F = w.scal[0], c = w.scal[1];
    st.evals = 1;
    double viol = con_violation(p.ctype, c);
    bool feasible = viol <= p.ctol;
    double infeas = viol;
    st.minf = F;
    par.foreach(n, [&](int i) { w.xbest[i] = w.x[i]; });
    bool b_identity = true;
    double mu = 0.0;
    int ireset = 0;
    int stalled = 0;  // line searches in a row that ran out of trials without the sufficient decrease
    int ret = SQP_SUCCESS;
    while (ret == SQP_SUCCESS) {
        ++st.iters;
        if (p.max_iter > 0 && st.iters > p.max_iter) { ret = SQP_FAILURE; break; }
        bool ok = true;
        const double lambda = qp_dense(par, w, p.ctype, c, st, &ok);
        if (!ok) {  // the model lost positive definiteness (round-off): start it again
            ++ireset; ++st.resets;
            if (ireset > 5 || b_identity) { ret = SQP_ROUNDOFF_LIMITED; break; }
            identity(); b_identity = true;
            continue;
        }
        double ad = 0.0, gs = 0.0;
        bool large = false;
        for (int i = 0; i < n; ++i) {
            ad += w.a[i] * w.d[i];
            gs += w.g[i] * w.d[i];
            if (!(fabs(w.d[i]) < w.xtol[i])) large = true;
        }
        par.foreach(n, [&](int i) {
            w.gl[i] = w.g[i] - lambda * w.a[i];   // gradient of the Lagrangian at the old point, new multiplier
            w.x0[i] = w.x[i];
        });
        if (!large) { ret = SQP_XTOL_REACHED; break; }  // step test, as in sqp_core.h
        const double cres = (p.ctype == CON_NONE) ? 0.0 : c + ad;
        const double f0 = F;
        const double al = fabs(lambda);
        mu = (al > 0.5 * (mu + al)) ? al : 0.5 * (mu + al);
        const double viol0 = con_violation(p.ctype, c);
        const double viol_lin = con_violation(p.ctype, cres);
        const double t0 = f0 + mu * viol0;
        double h3 = gs - mu * (viol0 - viol_lin);
        if (!(h3 < 0.0)) {
            ++ireset; ++st.resets;
            if (ireset > 5 || b_identity) { ret = SQP_ROUNDOFF_LIMITED; break; }
            identity(); b_identity = true;
            continue;
        }
        int line = 0;
        double alpha = 1.0;
        for (;;) {
            ++line;
            h3 = alpha * h3;
            par.foreach(n, [&](int i) {
                const double si = alpha * w.d[i];
                w.d[i] = si;
                double xi = w.x0[i] + si;
                if (xi < w.xl[i]) xi = w.xl[i];
                if (xi > w.xu[i]) xi = w.xu[i];
                w.x[i] = xi;
            });
            ev(par, w.x, w.g, w.a, w.scal);
            F = w.scal[0]; c = w.scal[1];
            ++st.evals;
            {
                const double vi = con_violation(p.ctype, c);
                const bool fc = vi <= p.ctol;
                if ((F < st.minf && (fc || !feasible)) || (!feasible && vi < infeas)) {
                    feasible = fc; infeas = vi; st.minf = F;
                    par.foreach(n, [&](int i) { w.xbest[i] = w.x[i]; });
                }
            }
            if (p.maxeval > 0 && st.evals >= p.maxeval) { ret = SQP_MAXEVAL_REACHED; break; }
            const double tt = F + mu * con_violation(p.ctype, c);
            const double h1 = tt - t0;
            if (h1 <= h3 / 10.0) { stalled = 0; break; }
            if (line > 10) { ++stalled; break; }
            const double cand = h3 / (2.0 * (h3 - h1));
            alpha = cand > 0.1 ? cand : 0.1;
        }
        if (ret != SQP_SUCCESS) break;
        // Twenty line searches in a row without the sufficient decrease: the merit function moves in its last digits only (an
        // objective at its optimum to 1e-12 with steps of a few xtol left, the multiplier equal to the penalty weight) and the
        // iteration would crawl on until max_iter.  NLopt calls this "roundoff errors limited progress"; the best point stands.
        // (A handful of such searches in a row do occur in ordinary solves that then go on to meet xtol.)
        if (stalled >= 20) { ret = SQP_ROUNDOFF_LIMITED; break; }
        bool moved = false;
        for (int i = 0; i < n; ++i) if (!(fabs(w.x[i] - w.x0[i]) < w.xtol[i])) moved = true;
        if (!moved) { ret = SQP_XTOL_REACHED; break; }
        // Powell's damped BFGS update with s = x - x0, y = gradL(x) - gradL(x0)
        double sBs = 0.0, sy = 0.0;
        par.foreach(n, [&](int i) {
            w.t1[i] = w.x[i] - w.x0[i];                       // s
            w.t3[i] = (w.g[i] - lambda * w.a[i]) - w.gl[i];   // y
        });
        par.foreach(n, [&](int i) {
            double s = 0.0;
            for (int j = 0; j < n; ++j) s += w.B[i * n + j] * w.t1[j];
            w.t2[i] = s;                                      // B s
        });
        for (int i = 0; i < n; ++i) { sBs += w.t1[i] * w.t2[i]; sy += w.t1[i] * w.t3[i]; }
        if (sBs > 0.0) {
            if (sy < 0.2 * sBs) {
                const double theta = 0.8 * sBs / (sBs - sy);
                par.foreach(n, [&](int i) { w.t3[i] = theta * w.t3[i] + (1.0 - theta) * w.t2[i]; });
                sy = 0.0;
                for (int i = 0; i < n; ++i) sy += w.t1[i] * w.t3[i];
            }
            if (sy > 0.0) {
                par.foreach(n * n, [&](int k) {
                    const int i = k / n, j = k % n;
                    w.B[k] += w.t3[i] * w.t3[j] / sy - w.t2[i] * w.t2[j] / sBs;
                });
                b_identity = false;
            }
        }
    }
    st.feasible = feasible ? 1 : 0;
    return ret;
}

}  // namespace accsqp
