// sqp_core.h -- sequential quadratic programming core for problems with box bounds and
// at most ONE general constraint (equality or inequality).
//
// This is the solver that replaces the reference's per-agent NLopt(SLSQP) calls
// (reference include/optimization.h:123-190; call sites src/model/PurchasingManager.cpp:690-702
// and src/model/Consumer.cpp:381-433).  It keeps the skeleton of Kraft's SLSQP as NLopt drives
// it: Hessian model started at the identity, one QP sub-problem per iteration, L1 merit function
// with Kraft's inexact line search (Armijo factor 0.1, quadratic interpolation floored at 0.1, at
// most 10 backtracks), at most 5 Hessian resets on a non-descent direction (then "roundoff
// limited"), NLopt's stopping rules (xtol_abs between completed iterations, maxeval) and NLopt's
// best-point bookkeeping.  One thing is deliberately different: the Hessian model is DIAGONAL,
// updated by the per-coordinate secant  b_i = dgradL_i / dx_i.  The purchasing objective is
// separable (each term depends on its own D_r only, PurchasingManager.cpp:220-241) and its
// constraint is linear, so the Lagrangian's Hessian IS diagonal and the secant is exact for it;
// the QP sub-problem then collapses to a monotone one-dimensional root-find in the constraint
// multiplier, which a warp solves exactly out of shared memory.  The first iteration (identity
// Hessian) coincides with SLSQP's.
//
// Kinks.  The purchasing objective contains, per coordinate, a term that is piecewise linear with ONE kink
// (markup * |D_r - target_r|, PurchasingManager.cpp:388-396; or the linear transport penalty, :432-450).  A smooth
// model cannot represent it: at the baseline fixed point every coordinate sits exactly on its kink and line searches
// along a smooth-model direction fail.  The work set may therefore describe the kink of coordinate i: its position
// kpos(i), the side kside(i, x) a point lies on, and the slopes ksl(i) <= ksr(i) left and right of it (kmid(i) = what
// the caller's gradient contains AT the kink).  The solver then works with the SMOOTH part of the gradient
// (g - sigma_i(x_i)), keeps the kink term exactly inside the QP sub-problem -- whose per-coordinate solution stays
// closed-form: right branch, left branch, or "stay on the kink" -- and uses the model decrease including the kink term
// in the merit test.  With no kinks described everything reduces to the smooth algorithm.
//
// The file compiles as plain C++ (host, used by oracle/nlopt_shim so that the reference's own
// CPU run and the device run execute the same arithmetic) and as CUDA device code.  All
// reductions go through Ctx::sum / Ctx::vmin / Ctx::vmax; Ctx::sum is defined to accumulate
// element i into partial (i mod 32) in increasing i and then combine the 32 partials by an
// xor-butterfly, so host and device produce bit-identical results.
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define SQP_HD __host__ __device__ __forceinline__
#else
#define SQP_HD inline
#endif

// Shared limits (host shim and device kernel must agree so that both sides stop alike).
#ifndef ACCSQP_MAX_ITER
#define ACCSQP_MAX_ITER 200  // SQP iterations per call before "more than iter SQP iterations"
#endif
// Size of the thread group that runs one solve on the device.  Its reduction order (coop.cuh) is part of the
// arithmetic, so the oracle's host context replays exactly this group size.
#ifndef ACCSQP_GROUP
#define ACCSQP_GROUP 96
#endif
// Purchasing managers with more connections than two per thread of the base group get a larger group, chosen from the
// LARGEST manager of the network (one launch shape for all; the oracle's harness applies the same rule):
// two connections per thread, a multiple of the warp size, at most 512 threads.
#if defined(__CUDACC__)
__host__ __device__
#endif
inline int accsqp_group_for(int n_max) {
    if (n_max <= 2 * ACCSQP_GROUP) return ACCSQP_GROUP;
    const int g = ((n_max + 1) / 2 + 31) / 32 * 32;
    return g > 512 ? 512 : g;
}
#define ACCSQP_BMIN 1e-12
#define ACCSQP_BMAX 1e+15

namespace accsqp {

enum : int {
    SQP_SUCCESS = 1,
    SQP_XTOL_REACHED = 4,
    SQP_MAXEVAL_REACHED = 5,
    SQP_ROUNDOFF_LIMITED = -4,
    SQP_FAILURE = -1,
};

enum : int { CON_NONE = 0, CON_EQ = 1, CON_INEQ = 2 };  // CON_INEQ: c(x) >= 0 (Kraft's convention)

struct Params {
    int n;
    int ctype;
    double ctol;      // feasibility tolerance on the constraint (NLopt's per-constraint tol)
    int maxeval;      // <=0: unlimited
    int max_iter;     // cap on SQP iterations (<=0: unlimited)
};

struct Work {
    // length-n vectors
    double *x, *x0, *g, *a, *d, *b, *gl, *xbest;
    const double *xl, *xu, *xtol;
    // optional kink descriptors (null = smooth objective), see "Kinks" below
    const double *kscale = nullptr, *ktarget = nullptr, *ksl_ = nullptr, *ksr_ = nullptr, *kmid_ = nullptr;
    SQP_HD double lower(int i) const { return xl != nullptr ? xl[i] : 0.0; }
    SQP_HD double tol(int i) const { return xtol[i]; }
    SQP_HD bool kinked() const { return kscale != nullptr; }
    SQP_HD int kside(int i, double xi) const { const double D = xi * kscale[i]; return D < ktarget[i] ? -1 : (D > ktarget[i] ? 1 : 0); }
    SQP_HD double kpos(int i) const { return ktarget[i] / kscale[i]; }
    SQP_HD double ksl(int i) const { return ksl_[i]; }
    SQP_HD double ksr(int i) const { return ksr_[i]; }
    SQP_HD double kmid(int i) const { return kmid_[i]; }
};
#define ACCSQP_NVEC 8  // number of length-n work vectors above
// lower bounds may be omitted (null pointer = all zero); step bounds are derived where they are used
#define ACCSQP_XL(w, i) ((w).lower(i))
#define ACCSQP_LO(w, i) (ACCSQP_XL(w, i) - (w).x[i])
#define ACCSQP_HI(w, i) ((w).xu[i] - (w).x[i])

struct Stats {
    int iters;
    int evals;
    int qp_iters;
    int qp_inexact;
    int resets;
    double minf;
    int feasible;
};

// Reduction payloads (Ctx::reduce<T> folds them in the fixed order documented in coop.cuh).
struct Sum2 {
    double a, b;
    static SQP_HD Sum2 identity() { Sum2 r; r.a = 0.0; r.b = 0.0; return r; }
    static SQP_HD Sum2 combine(Sum2 x, Sum2 y) { Sum2 r; r.a = x.a + y.a; r.b = x.b + y.b; return r; }
};
struct Sum3 {
    double a, b, c;
    static SQP_HD Sum3 identity() { Sum3 r; r.a = 0.0; r.b = 0.0; r.c = 0.0; return r; }
    static SQP_HD Sum3 combine(Sum3 x, Sum3 y) { Sum3 r; r.a = x.a + y.a; r.b = x.b + y.b; r.c = x.c + y.c; return r; }
};
struct Sum4 {
    double a, b, c, d;
    static SQP_HD Sum4 identity() { Sum4 r; r.a = 0.0; r.b = 0.0; r.c = 0.0; r.d = 0.0; return r; }
    static SQP_HD Sum4 combine(Sum4 x, Sum4 y) { Sum4 r; r.a = x.a + y.a; r.b = x.b + y.b; r.c = x.c + y.c; r.d = x.d + y.d; return r; }
};
struct MinMax {
    double lo, hi;
    static SQP_HD MinMax identity() { MinMax r; r.lo = 1e308; r.hi = -1e308; return r; }
    static SQP_HD MinMax combine(MinMax x, MinMax y) { MinMax r; r.lo = y.lo < x.lo ? y.lo : x.lo; r.hi = y.hi > x.hi ? y.hi : x.hi; return r; }
};

// ---------------------------------------------------------------------------------------
// QP sub-problem:  min 1/2 d^T diag(1/h) d + g^T d   s.t.  a^T d + c (=|>=) 0,  lo <= d <= hi,
// with the INVERSE of the diagonal Hessian model in w.b (h_i = 1/b_i), so the sub-problem needs no division.
// Solution d_i(lambda) = clamp((-g_i + lambda a_i) h_i, lo_i, hi_i); psi(lambda) = a^T d + c is
// non-decreasing and piecewise linear, its root is found by Newton steps kept inside a bracket; psi and
// its slope come out of ONE pass.  Returns lambda (gradient condition  B d + g = lambda a + z).
// *d_ready tells the caller whether w.d already holds the step (vertex solutions) or is still to be formed
// from lambda (the caller fuses that pass with its own sums).
// ---------------------------------------------------------------------------------------
// slope of the kink term at xi (what the caller's gradient contains for it) and its value relative to the kink
template <class W>
SQP_HD double kink_sigma(const W& w, int i, double xi) {
    const int sd = w.kside(i, xi);
    return sd < 0 ? w.ksl(i) : (sd > 0 ? w.ksr(i) : w.kmid(i));
}
template <class W>
SQP_HD double kink_phi(const W& w, int i, double xi) {
    const double u = xi - w.kpos(i);
    return u < 0.0 ? w.ksl(i) * u : w.ksr(i) * u;
}

// unclamped minimiser of  d^2 / (2 h) + (g - lambda a) d + kink(x + d)  and whether it lies on a smooth piece
template <class W>
SQP_HD double qp_free_step(const W& w, int i, double lambda, bool kinked, bool* smooth) {
    const double r = -w.g[i] + lambda * w.a[i];
    *smooth = true;
    if (!kinked) return r * w.b[i];
    const double u = w.x[i] - w.kpos(i);
    const double dr = (r - w.ksr(i)) * w.b[i];
    if (u + dr > 0.0) return dr;
    const double dl = (r - w.ksl(i)) * w.b[i];
    if (u + dl < 0.0) return dl;
    *smooth = false;
    return -u;
}
template <class W>
SQP_HD double qp_step(const W& w, int i, double lambda, bool kinked) {
    bool smooth;
    const double t = qp_free_step(w, i, lambda, kinked, &smooth);
    return t < ACCSQP_LO(w, i) ? ACCSQP_LO(w, i) : (t > ACCSQP_HI(w, i) ? ACCSQP_HI(w, i) : t);
}

// contribution of coordinate i to the attainable range [sum of minima, sum of maxima] of a^T d over the box
template <class W>
SQP_HD Sum2 qp_range_term(const W& w, int i) {
    const double p = w.a[i] * ACCSQP_LO(w, i), q = w.a[i] * ACCSQP_HI(w, i);
    Sum2 r; r.a = p < q ? p : q; r.b = p > q ? p : q; return r;
}

// `known_range`: the attainable range of a^T d at the current x when the caller has it already (the evaluation pass
// of a per-coordinate objective computes it on the way, see sqp_eval_pass), else null.
// `lambda_start`: where the root-find begins -- the multiplier of the previous SQP iteration (psi is piecewise linear and
// Newton's steps exact between breakpoints, so starting next to the root saves passes; 0 for the first iteration).
// `step_sums` / `have_sums`: every pass of the root-find forms the whole step d(lambda) anyway, so it also stores it in w.d and
// returns the three sums the driver needs from the step (a.d, number of coordinates whose step reaches the x tolerance, model
// change of the objective g.d + kink terms).  When the root-find ends on a pass at the returned multiplier -- every exit but
// the exhausted iteration cap -- the driver needs no pass of its own to form the step.
template <class Ctx, class W>
SQP_HD double qp_solve(Ctx& cx, int n, int ctype, double c, W& w, Stats& st, bool* d_ready, const Sum2* known_range, double lambda_start,
                       Sum3* step_sums, bool* have_sums) {
    *d_ready = false;
    *have_sums = false;
    const bool kinked = w.kinked();
    if (ctype == CON_NONE) return 0.0;
    const double target = -c;  // want a^T d = target
    // attainable range of a^T d over the box
    const Sum2 range = known_range != nullptr ? *known_range : cx.template reduce<Sum2>(n, [&](int i) { return qp_range_term(w, i); });
    const double smin = range.a, smax = range.b;
    int forced = 0;  // -1: push a^T d to its minimum, +1: to its maximum
    if (ctype == CON_EQ && target <= smin) forced = -1;
    else if (target >= smax) forced = 1;
    if (forced != 0) {
        // vertex of the box; multiplier = the extreme ratio that keeps every bound multiplier sign-feasible
        cx.foreach(n, [&](int i) {
            const double p = w.a[i] * ACCSQP_LO(w, i), q = w.a[i] * ACCSQP_HI(w, i);
            w.d[i] = (forced > 0) ? (p >= q ? ACCSQP_LO(w, i) : ACCSQP_HI(w, i)) : (p <= q ? ACCSQP_LO(w, i) : ACCSQP_HI(w, i));
        });
        *d_ready = true;
        const double big = 1e300;
        // need  (d/h + g)_i - lambda a_i <= 0 where d sits at hi,  >= 0 where it sits at lo:
        // lmin = largest lower bound on lambda, lmax = smallest upper bound
        const MinMax lim = cx.template reduce<MinMax>(n, [&](int i) {
            MinMax r; r.lo = big; r.hi = -big;
            if (w.a[i] == 0.0 || ACCSQP_LO(w, i) == ACCSQP_HI(w, i)) return r;
            const bool at_hi = (w.d[i] == ACCSQP_HI(w, i));
            const double q = (w.d[i] / w.b[i] + w.g[i]) / w.a[i];
            if (at_hi == (w.a[i] > 0.0)) r.hi = q; else r.lo = q;
            return r;
        });
        const double lmin = lim.hi, lmax = lim.lo;
        double lambda;
        if (lmin > -big && lmax < big) lambda = (fabs(lmin) < fabs(lmax)) ? lmin : lmax;
        else if (lmin > -big) lambda = lmin;
        else if (lmax < big) lambda = lmax;
        else lambda = 0.0;
        if (ctype == CON_INEQ && lambda < 0.0) lambda = 0.0;
        return lambda;
    }

    // |psi| below the rounding noise of the n-term sums it is made of (c itself is such a sum) counts as zero:
    // n * 2^-52 relative to the magnitudes involved.  A tighter test only makes Newton chase that noise.
    const double psi_tol = 2.3e-16 * (n < 2 ? 2 : n) * (fabs(c) + (fabs(smin) > fabs(smax) ? fabs(smin) : fabs(smax)));
    double lambda = (ctype == CON_EQ) ? lambda_start : 0.0;  // the inequality form tests psi(0) first (constraint inactive)
    double lamL = 0.0, lamR = 0.0;
    bool haveL = false, haveR = false;
    double step = 1.0;
    double prev_lambda = 0.0;
    bool prev_valid = false;
    bool converged = false;
    int it = 0;
    Sum4 ps = Sum4::identity();
    for (; it < 200; ++it) {
        // psi and its slope at lambda; the step itself and the driver's sums on the way
        ps = cx.template reduce<Sum4>(n, [&](int i) {
            bool smooth;
            const double t = qp_free_step(w, i, lambda, kinked, &smooth);
            const double lo = ACCSQP_LO(w, i), hi = ACCSQP_HI(w, i);
            Sum4 r;
            double di;
            if (t <= lo) { di = lo; r.b = 0.0; }
            else if (t >= hi) { di = hi; r.b = 0.0; }
            else { di = t; r.b = smooth ? w.a[i] * w.a[i] * w.b[i] : 0.0; }
            r.a = w.a[i] * di;
            w.d[i] = di;
            r.c = (fabs(di) < w.tol(i)) ? 0.0 : 1.0;
            r.d = w.g[i] * di;
            if (kinked) r.d += kink_phi(w, i, w.x[i] + di) - kink_phi(w, i, w.x[i]);  // model decrease incl. the kink term
            return r;
        });
        const double psi = c + ps.a, slope = ps.b;
        if (it == 0 && ctype == CON_INEQ && psi >= 0.0) { converged = true; break; }  // constraint inactive
        if (fabs(psi) <= psi_tol) { converged = true; break; }
        if (psi < 0.0) { lamL = lambda; haveL = true; } else { lamR = lambda; haveR = true; }
        if (prev_valid && lambda == prev_lambda) { converged = true; break; }
        prev_lambda = lambda; prev_valid = true;
        double next;
        bool ok = false;
        if (slope > 0.0) {
            next = lambda - psi / slope;
            ok = (!haveL || next > lamL) && (!haveR || next < lamR);
        }
        if (!ok) {
            if (haveL && haveR) {
                next = 0.5 * (lamL + lamR);
                if (next == lamL || next == lamR) { converged = true; break; }
            } else {
                next = (psi < 0.0) ? lambda + step : lambda - step;
                step *= 4.0;
            }
        }
        lambda = next;
    }
    if (it < 200) {  // the last pass ran at the multiplier that is returned: w.d holds its step
        *d_ready = true;
        *have_sums = true;
        step_sums->a = ps.a; step_sums->b = ps.c; step_sums->c = ps.d;
    }
    st.qp_iters += it;
    if (!converged) st.qp_inexact += 1;
    return lambda;
}

// ---------------------------------------------------------------------------------------
// Main driver.  Eval is a callable:  ev(const double* x, bool want_grad, double* F, double* c)
// returning the objective to MINIMISE in *F and the constraint value in *c (Kraft's convention:
// equality c == 0, inequality c >= 0), and -- when want_grad -- filling w.g (objective gradient)
// and w.a (constraint gradient).  On return w.xbest holds the best point found (NLopt's rule:
// best feasible objective, else least infeasible) and st.minf its objective.
// ---------------------------------------------------------------------------------------
// An objective may declare itself separable: EvalTraits<Eval>::PER_COORD = 1 promises
//   ev.coord(i, xi)      evaluates coordinate i at xi (stores its summand, fills w.g[i]; touches nothing else), and
//   ev.finish(x, F, c)   completes the evaluation (the ordered sums over all coordinates).
// The solver then moves x, evaluates, strips the kink slope and gathers the QP's range in ONE pass over the coordinates
// instead of four.  Every coordinate sees the same operations in the same order as in the generic sequence, so the
// results are bit-identical (the oracle's NLopt shim runs the generic sequence: its objective is the reference's
// callback, a black box).
template <class Eval>
struct EvalTraits {
    enum { PER_COORD = 0 };
};

// contribution of coordinate i to the driver's first multiplier estimate (see sqp_minimize): a_i g_i and a_i^2 next to the range
template <class W>
SQP_HD Sum4 qp_start_term(const W& w, int i) {
    const Sum2 rg = qp_range_term(w, i);
    Sum4 r; r.a = rg.a; r.b = rg.b; r.c = w.a[i] * w.g[i]; r.d = w.a[i] * w.a[i]; return r;
}

// x_i <- newx(i) for every coordinate, objective / constraint / gradient there, smooth part of the gradient in w.g.
// Returns true when *range holds the attainable range of a^T d at the new x.  With `start` not null (the first evaluation
// of a run) the range comes with the two sums of the multiplier estimate, whichever way the objective is evaluated.
template <class Ctx, class Eval, class W, class NewX>
SQP_HD bool sqp_eval_pass(Ctx& cx, Eval& ev, W& w, int n, bool kinked, NewX newx, double* F, double* c, Sum2* range, Sum4* start = nullptr) {
    if constexpr (EvalTraits<Eval>::PER_COORD != 0) {
        if (start != nullptr) {
            *start = cx.template reduce<Sum4>(n, [&](int i) {
                const double xi = newx(i);
                w.x[i] = xi;
                ev.coord(i, xi);
                if (kinked) w.g[i] -= kink_sigma(w, i, xi);
                return qp_start_term(w, i);
            });
            range->a = start->a; range->b = start->b;
        } else {
            *range = cx.template reduce<Sum2>(n, [&](int i) {
                const double xi = newx(i);
                w.x[i] = xi;
                ev.coord(i, xi);
                if (kinked) w.g[i] -= kink_sigma(w, i, xi);
                return qp_range_term(w, i);
            });
        }
        ev.finish(w.x, F, c);
        return true;
    } else {
        cx.foreach(n, [&](int i) { w.x[i] = newx(i); });
        ev(w.x, true, F, c);
        if (kinked) cx.foreach(n, [&](int i) { w.g[i] -= kink_sigma(w, i, w.x[i]); });  // keep the smooth part only
        if (start != nullptr) {
            *start = cx.template reduce<Sum4>(n, [&](int i) { return qp_start_term(w, i); });
            range->a = start->a; range->b = start->b;
            return true;
        }
        return false;
    }
}

SQP_HD double con_violation(int ctype, double c) {
    if (ctype == CON_EQ) return fabs(c);
    if (ctype == CON_INEQ) return c < 0.0 ? -c : 0.0;
    return 0.0;
}

// W is any work-vector set with members x, x0, g, a, d, b, gl, xbest, xu indexable by [i], lower(i) and tol(i) (x tolerance):
// plain pointers (struct Work: host shim, emulation, shared-memory kernel) or per-lane register vectors (device).
template <class Ctx, class Eval, class W>
SQP_HD int sqp_minimize(Ctx& cx, Eval& ev, W& w, const Params& p, Stats& st) {
    const int n = p.n;
    st.iters = 0; st.evals = 0; st.qp_iters = 0; st.qp_inexact = 0; st.resets = 0; st.minf = 0.0; st.feasible = 0;

    const bool kinked = w.kinked();
    double F = 0.0, c = 0.0;
    Sum2 range;
    Sum4 start;
    // NLopt clips the start point into the box; the first point is the best one so far
    bool have_range = sqp_eval_pass(cx, ev, w, n, kinked, [&](int i) {
        double xi = w.x[i];
        if (xi < ACCSQP_XL(w, i)) xi = ACCSQP_XL(w, i);
        if (xi > w.xu[i]) xi = w.xu[i];
        w.b[i] = 1.0;
        w.xbest[i] = xi;
        return xi;
    }, &F, &c, &range, &start);
    st.evals = 1;
    double viol = con_violation(p.ctype, c);
    bool feasible = viol <= p.ctol;
    double infeas = viol;
    st.minf = F;

    bool b_identity = true;
    // Where the first root-find starts: with the identity model and no bound or kink in the way the QP's multiplier is
    // (a.g - c) / a.a exactly; the later ones start at their predecessor's multiplier.
    double lambda_prev = (p.ctype == CON_EQ && start.d > 0.0) ? (start.c - c) / start.d : 0.0;
    double mu = 0.0;      // merit penalty parameter
    int ireset = 0;
    int ret = SQP_SUCCESS;

    while (ret == SQP_SUCCESS) {
        ++st.iters;
        if (p.max_iter > 0 && st.iters > p.max_iter) { ret = SQP_FAILURE; break; }
        bool d_ready = false, have_sums = false;
        Sum3 sd;
        const double lambda = qp_solve(cx, n, p.ctype, c, w, st, &d_ready, have_range ? &range : nullptr, lambda_prev, &sd, &have_sums);
        lambda_prev = lambda;
        // The step, its linearised constraint a.d, the number of coordinates whose step reaches the x tolerance and the model's
        // first-order change of the objective g.d (+ change of the kink terms) come out of the root-find's last pass; only a
        // vertex solution or an exhausted root-find leaves them to a pass of its own.
        if (!have_sums) {
            sd = cx.template reduce<Sum3>(n, [&](int i) {
                const double di = d_ready ? w.d[i] : qp_step(w, i, lambda, kinked);
                w.d[i] = di;
                Sum3 r; r.a = w.a[i] * di; r.b = (fabs(di) < w.tol(i)) ? 0.0 : 1.0; r.c = w.g[i] * di;
                if (kinked) r.c += kink_phi(w, i, w.x[i] + di) - kink_phi(w, i, w.x[i]);  // model decrease incl. the kink term
                return r;
            });
        }
        const double cres = (p.ctype == CON_NONE) ? 0.0 : c + sd.a;
        // Step test (Kraft's convergence test is on the size of the QP step as well): when the full step is below the
        // x tolerance in every coordinate, no line search result can move x by a tolerance -- stop here instead of
        // spending evaluations on a step NLopt's xtol rule would discard at the end of the iteration.
        if (sd.b == 0.0) { ret = SQP_XTOL_REACHED; break; }
        const double f0 = F;
        const double gs = sd.c;
        const double al = fabs(lambda);
        mu = (al > 0.5 * (mu + al)) ? al : 0.5 * (mu + al);
        const double viol0 = con_violation(p.ctype, c);
        const double viol_lin = con_violation(p.ctype, cres);
        const double t0 = f0 + mu * viol0;
        double h3 = gs - mu * (viol0 - viol_lin);
#ifdef ACCSQP_DEBUG
        printf("iter %d lambda %g gs %g mu %g viol0 %g viol_lin %g h3 %g F %.12g c %g\n", st.iters, lambda, gs, mu, viol0, viol_lin, h3, F, c);
#endif
        if (!(h3 < 0.0)) {
            // not a descent direction for the merit function: reset the Hessian model (Kraft L110), at most 5 times
            ++ireset; ++st.resets;
            if (ireset > 5 || b_identity) { ret = SQP_ROUNDOFF_LIMITED; break; }
            cx.foreach(n, [&](int i) { w.b[i] = 1.0; });
            b_identity = true;
            continue;
        }
        // Kraft's inexact line search (Armijo factor 0.1, quadratic interpolation floored at 0.1, at most 10
        // backtracks).  Every trial is evaluated WITH gradients (nothing reads the old gradient any more: the
        // Lagrangian gradient at the old point sits in w.gl), so the accepted trial needs no second evaluation.
        int line = 0;
        double alpha = 1.0;
        for (;;) {
            ++line;
            h3 = alpha * h3;
            have_range = sqp_eval_pass(cx, ev, w, n, kinked, [&](int i) {
                if (line == 1) {  // the old point and the Lagrangian gradient there with the new multiplier (Kraft: v), before x and g move
                    w.gl[i] = w.g[i] - lambda * w.a[i];
                    w.x0[i] = w.x[i];
                }
                const double si = alpha * w.d[i];
                w.d[i] = si;
                double xi = w.x0[i] + si;
                if (xi < ACCSQP_XL(w, i)) xi = ACCSQP_XL(w, i);
                if (xi > w.xu[i]) xi = w.xu[i];
                return xi;
            }, &F, &c, &range);
            ++st.evals;
            {   // NLopt: keep the best point seen at every evaluation
                const double vi = con_violation(p.ctype, c);
                const bool fc = vi <= p.ctol;
                if ((F < st.minf && (fc || !feasible)) || (!feasible && vi < infeas)) {
                    feasible = fc; infeas = vi; st.minf = F;
                    cx.foreach(n, [&](int i) { w.xbest[i] = w.x[i]; });
                }
            }
            if (p.maxeval > 0 && st.evals >= p.maxeval) { ret = SQP_MAXEVAL_REACHED; break; }
            const double tt = F + mu * con_violation(p.ctype, c);
            const double h1 = tt - t0;
            if (h1 <= h3 / 10.0 || line > 10) break;
            const double cand = h3 / (2.0 * (h3 - h1));
            alpha = cand > 0.1 ? cand : 0.1;
        }
        if (ret != SQP_SUCCESS) break;
        // NLopt's xtol test between completed iterations, fused with the per-coordinate secant update of the
        // inverse diagonal Hessian model  h_i = dx_i / dgradL_i
        const double moved = cx.sum(n, [&](int i) {
            const double si = w.x[i] - w.x0[i];
            if (si != 0.0) {
                const double ui = (w.g[i] - lambda * w.a[i]) - w.gl[i];
                const double r = si / ui;
                if (r > 0.0 && r < 1e300) {
                    w.b[i] = r < 1.0 / ACCSQP_BMAX ? 1.0 / ACCSQP_BMAX : (r > 1.0 / ACCSQP_BMIN ? 1.0 / ACCSQP_BMIN : r);
                }
            }
            return (fabs(si) < w.tol(i)) ? 0.0 : 1.0;
        });
        if (moved == 0.0) { ret = SQP_XTOL_REACHED; break; }
        if (p.maxeval > 0 && st.evals >= p.maxeval) { ret = SQP_MAXEVAL_REACHED; break; }
        b_identity = false;
    }
    st.feasible = feasible ? 1 : 0;
    return ret;
}

}  // namespace accsqp
