// consumer_core.cuh -- the utilitarian consumer: Consumer::iterate_consumption_and_production with the nested-CES
// utility optimisation, for ONE consumer, executed by one thread group.
//
// Follows reference src/model/Consumer.cpp:
//   :170-187  autodiff_nested_CES_utility_function (value and forward-mode gradient, lib/cpp-library/autodiff.h:67-250)
//   :204-226  inequality_constraint / equality_constraint (the budget)
//   :234-253  max_objective
//   :312-364  start value, bounds, xtol of the optimisation
//   :381-433  optimiser set-up (local SLSQP; the global MLSL path is not provided)
//   :455-503  consumption_optimize: result codes -> events
//   :510-523  consume_optimisation_result
// The utility couples all goods of a consumer; objective, gradient and constraint follow the reference's own operation
// order (every intermediate of the autodiff chain is reproduced, which makes the host-side emulation bit-identical to the
// reference's autodiff; the CUDA pow() may differ from glibc's in the last bit).  The optimiser is csrc/sqp_dense.h (dense
// BFGS model: the utility is not separable).  The whole warp runs it: the same scalar control flow on every lane, the
// element-wise work -- BFGS update, Cholesky columns, matrix-vector products, bound multipliers, the pow() of every good --
// shared out over the lanes, recurrences in one lane; every ELEMENT is still computed by one thread with the operations
// of a sequential loop in the same order, so the bits equal those of the serial run of the same code in oracle/nlopt_shim
// (which uses it for every objective without described kinks).
#pragma once

#include "model_core.cuh"
#include "sqp_dense.h"

namespace acc {

struct ConsumerView {
    const int* cons_ids;                    // consumers this engine computes
    const int* st_off; const int* st_list;  // agent -> its storages in input_storages order
    // per storage (consumer storages only)
    const double* share;     // share_factors (normalised inside the basket), Consumer.cpp:138-160
    const double* eshare;    // exponent_share_factors
    const i64* base_q;       // baseline_consumption quantity (0 outside every basket)
    i64* prev_q; double* prev_v;  // previous_consumption
    // per agent
    const double* budget;    // to_float(consumption_budget)
    const int* order_cnt;    // number of storages inside baskets
    const int* order;        // [st_off[a] + k]: local index of the k-th in-basket storage, sorted by (basket, sector)
    const int* order_basket; // its basket
    const double* bshare; const double* ebshare;  // [agent * n_baskets + b] basket_share_factors, exponent_basket_share_factors
    double* utility;         // optimised utility
    // model
    const double* intra_exp; // [n_baskets] intra_basket_substitution_exponent
    double inter_exp;        // inter_basket_substitution_exponent
    int n_baskets;
    double xtol_c;           // FlowQuantity::precision * utility_optimization_precision_adjustment
    double ctol;             // FlowValue::precision
    int maxeval;             // utility_optimization_maxiter
};

ACC_HD size_t consumer_workspace_bytes(int n_max) {
    return (accsqp::dense_doubles(n_max) + (size_t)8 * n_max + 4) * sizeof(double) + (size_t)4 * n_max * sizeof(int);
}

struct ConsumerWork {
    accsqp::DenseWork w;
    double *xl, *xu, *xtol;
    double *price;   // consumption_prices as Price quanta (exact in f64)
    double *basef;   // to_float(baseline_consumption quantity)
    double *share, *eshare;
    double *chain;   // per-good derivative chain of the current evaluation
    int *ord, *bsk;
    ACC_HD void carve(unsigned char* mem, int n_max, int n) {
        double* p = reinterpret_cast<double*>(mem);
        w.n = n;
        w.x = p; p += n_max; w.x0 = p; p += n_max; w.g = p; p += n_max; w.a = p; p += n_max; w.d = p; p += n_max;
        w.gl = p; p += n_max; w.xbest = p; p += n_max; w.t1 = p; p += n_max; w.t2 = p; p += n_max; w.t3 = p; p += n_max;
        w.B = p; p += (size_t)n_max * n_max; w.L = p; p += (size_t)n_max * n_max; w.scal = p; p += 4;
        xl = p; p += n_max; xu = p; p += n_max; xtol = p; p += n_max;
        price = p; p += n_max; basef = p; p += n_max; share = p; p += n_max; eshare = p; p += n_max;
        chain = p; p += n_max + 4;
        w.xl = xl; w.xu = xu; w.xtol = xtol;
        int* q = reinterpret_cast<int*>(p);
        w.state = q; q += 2 * n_max;
        ord = q; q += n_max; bsk = q;
    }
};

// objective (to MINIMISE: -utility), its gradient, budget constraint and its gradient in sqp_core's convention; what
// oracle/nlopt_shim's CallbackEval hands the solver from the reference's max_objective / equality_constraint
struct ConsumerEval {
    ConsumerWork& k;
    const ConsumerView& cv;
    int agent, n, n_in;
    double budget;
    // Group convention of sqp_dense.h: the pow() of every good is taken side by side (the expensive part), the sums
    // over goods and baskets stay strictly sequential in one thread, in the reference's order.
    template <class Par>
    ACC_HD void operator()(Par& par, const double* x, double* g, double* ga, double* scal) {
        ConsumerWork& kk = k;
        const ConsumerView& v = cv;
        const int a = agent, nn = n, nin = n_in;
        const double B = budget;
        double* const powsum = kk.w.t1;  // per ordered good: pow(c, e_b) * exp-share, the summand of its basket (t1 is free during an evaluation)
        // ---- autodiff_nested_CES_utility_function, :170-187 ----
        par.foreach(nn, [&](int i) { kk.chain[i] = 0.0; });
        par.foreach(nin, [&](int q) {
            const int i = kk.ord[q];
            const double eb = v.intra_exp[kk.bsk[q]];
            const double cq = x[i] * kk.share[i];
            powsum[q] = pow(cq, eb) * kk.eshare[i];
            kk.chain[i] = (kk.share[i] * (eb * pow(cq, eb - 1))) * kk.eshare[i];
        });
        par.single([&] {
            double U = 0.0;
            int q = 0;
            while (q < nin) {
                const int b = kk.bsk[q];
                const double eb = v.intra_exp[b];
                const int q0 = q;
                double ub = 0.0;
                for (; q < nin && kk.bsk[q] == b; ++q) ub += powsum[q];
                const double inv = 1 / eb;
                const double bs = v.bshare[a * v.n_baskets + b], ebs = v.ebshare[a * v.n_baskets + b];
                const double ub2 = pow(ub, inv) * bs;
                const double term = pow(ub2, v.inter_exp) * ebs;
                const double f1 = inv * pow(ub, inv - 1);
                const double f2 = v.inter_exp * pow(ub2, v.inter_exp - 1);
                for (int r = q0; r < q; ++r) { const int i = kk.ord[r]; kk.chain[i] = (((kk.chain[i] * f1) * bs) * f2) * ebs; }
                U += term;
            }
            const double invE = 1 / v.inter_exp;
            const double util = pow(U, invE);
            scal[0] = -1.0 * util;
            const double f3 = invE * pow(U, invE - 1);
            for (int i = 0; i < nn; ++i) g[i] = -1.0 * (kk.chain[i] * f3);
            // ---- equality_constraint == inequality_constraint, :204-226 (not_spent_budget is 0 here, :521) ----
            double scaled_budget = B / B;
            for (int i = 0; i < nn; ++i) {
                const double price = x[i] * (kk.price[i] * ACC_PREC6);
                const double term = (x[i] * kk.basef[i]) * price / B;
                scaled_budget -= term;
                ga[i] = 1.0 * term;
            }
            scal[1] = 1.0 * (-scaled_budget);
        });
    }
};

// the group of a context as sqp_dense.h's executor
template <class Ctx>
struct CtxPar {
    Ctx& cx;
    // barrier first: the threads may still be reading (redundant scalar loops) what this loop overwrites
    template <class F> ACC_HD void foreach(int m, F f) const { cx.sync(); cx.foreach(m, f); }
    template <class F> ACC_HD void single(F f) const { cx.single(f); }
};

// What a solve leaves behind (consumer_one) for consumer_apply: the solve only READS model state, so it can run ahead of the step
// it belongs to -- as soon as the storages of the previous step are final -- while nothing the rest of that step reads has changed.
struct ConsumerStage {
    i64* q; double* v;   // per storage: optimised consumption
    double* utility;     // per agent
    int* ret;            // per agent: the solver's return code
};

template <class Ctx>
ACC_HD void consumer_one(Ctx& cx, int a, const ConsumerView& cv, const StorageView& st, const ModelP& mp, const ConsumerStage& out, unsigned char* smem,
                         int n_max) {
    const int s0 = cv.st_off[a];
    const int n = cv.st_off[a + 1] - s0;
    ConsumerWork k;
    k.carve(smem, n_max, n);
    const int n_in = cv.order_cnt[a];
    const double B = cv.budget[a];
    // start value, bounds, tolerances, :331-377
    cx.foreach(n, [&](int i) {
        const int s = cv.st_list[s0 + i];
        const Flow uhat = possible_use(st, s, mp.dt, false);
        const i64 poss_q = uhat.q;
        const i64 price = flow_price(uhat);
        const i64 base_q = cv.base_q[s];
        const Flow base = base_q != 0 ? mkflow(st.Istar_q[s], st.Istar_v[s]) : mkflow(0, 0.0);
        const Flow start = cv.prev_q[s] == 0 ? base : mkflow(cv.prev_q[s], cv.prev_v[s]);
        i64 sq = start.q < poss_q ? start.q : poss_q;
        sq = qmul(sq, pow(ratio(price, flow_price(start)), st.a[s]));
        const double ssv = base_q != 0 ? ratio(sq, base_q) : 0.0;
        double lb = 0.0, ub = 0.0;
        if (!(ssv <= 0.0)) {
            const i64 p1 = mkp(B / qf(poss_q));            // consumption_budget / possible_consumption_quantity -> Price
            const i64 p2 = mkp(pf(p1) * pf(price));        // ... * to_float(consumption_prices) -> Price
            const double affordable = pf(p2) * ratio(poss_q, base_q);
            lb = stdmin(0.5, ssv);
            ub = stdmin(1.5, affordable);
        }
        k.xl[i] = lb; k.xu[i] = ub;
        k.xtol[i] = cv.xtol_c / qf(base_q);
        k.w.x[i] = stdmin(ssv, ub);
        k.price[i] = (double)price;
        k.basef[i] = qf(base_q);
        k.share[i] = cv.share[s]; k.eshare[i] = cv.eshare[s];
        k.ord[i] = i < n_in ? cv.order[s0 + i] : -1;
        k.bsk[i] = i < n_in ? cv.order_basket[s0 + i] : -1;
    });
    // the whole group runs the solver (same control flow on every thread, loops shared out: sqp_dense.h)
    ConsumerEval ev{k, cv, a, n, n_in, B};
    accsqp::Params sp;
    sp.n = n; sp.ctype = accsqp::CON_EQ; sp.ctol = cv.ctol; sp.maxeval = cv.maxeval; sp.max_iter = ACCSQP_MAX_ITER;
    accsqp::Stats stt;
    CtxPar<Ctx> par{cx};
    const int ret = accsqp::sqp_minimize_dense(par, ev, k.w, sp, stt);
    if (cx.leader()) { out.ret[a] = ret; out.utility[a] = -stt.minf; }
    cx.foreach(n, [&](int i) {
        const int s = cv.st_list[s0 + i];
        const Flow cons = fcopy(flow_qp(qmul_l(k.w.xbest[i], cv.base_q[s]), (i64)k.price[i]));
        out.q[s] = cons.q; out.v[s] = cons.v;
    });
}

// The step's own part of Consumer::iterate_consumption_and_production, one group per consumer: events of consumption_optimize
// (:455-503) and consume_optimisation_result (:510-523): desired use = use = previous consumption = optimised consumption.
// Returns false when the optimiser failed outright (the reference throws, Consumer.cpp:497-502).
template <class Ctx>
ACC_HD bool consumer_apply(Ctx& cx, int a, const ConsumerView& cv, const StorageView& st, const ConsumerStage& in, const EventSink& events) {
    const int ret = in.ret[a];
    if (cx.leader()) {
        if (ret == accsqp::SQP_ROUNDOFF_LIMITED) events.push(EV_OPTIMIZER_ROUNDOFF_LIMITED, a, -1, (double)NAN);
        else if (ret == accsqp::SQP_MAXEVAL_REACHED) events.push(EV_OPTIMIZER_MAXITER, a, -1, (double)NAN);
        cv.utility[a] = in.utility[a];
    }
    const int s0 = cv.st_off[a];
    cx.foreach(cv.st_off[a + 1] - s0, [&](int i) {
        const int s = cv.st_list[s0 + i];
        st.Ut_q[s] = in.q[s]; st.Ut_v[s] = in.v[s];
        cv.prev_q[s] = in.q[s]; cv.prev_v[s] = in.v[s];
    });
    return ret != accsqp::SQP_FAILURE;
}

}  // namespace acc
