// purchase_core.cuh -- PurchasingManager::iterate_purchase for ONE purchasing manager, executed by
// one thread group (a small CTA) with the connection data staged in shared memory.
//
// Follows reference src/model/PurchasingManager.cpp:
//   :563-660  desired purchase, per-connection bounds / start value / xtol, filtering
//   :149-451  objective, gradient, equality constraint (price curve n_r, transport penalty)
//   :465-561  optimiser driver: restart on "more than iter SQP iterations" and on maxeval
//   :704-737  demand requests D_r -> Demand(quantity, round(n_r)), accumulation of demand / costs
// The optimiser is csrc/sqp_core.h (the same code the oracle's NLopt shim runs on the CPU).
// Objective and constraint are accumulated strictly left to right in connection order, exactly as
// the reference's loops do (:203-217, :221-240), so that the solver sees bit-identical values.
#pragma once

#include "model_core.cuh"
#include "sqp_core.h"

namespace acc {

// ---- transport penalty parameters of one purchasing manager ----
struct PMParams {
    double Uf;        // to_float(initial_used_flow_U_star): scale of objective and constraint (:194,:200)
    double markup;    // to_float(buyer sector initial_markup)
    double pen_large, pen_small;
    int quadratic, relative, limited_by_markup;
};

// The transport penalty's kink at D_r == target (:388-396 quadratic, :432-450 linear): slopes of the piecewise linear
// part left and right of the target and the value the reference's gradient contains AT the target, per unit of D_r.
// Shared with the oracle's harness (oracle/ref_harness), which describes the same kinks to the NLopt shim.
// STD = the reference's default penalty (quadratic, relative) known at compile time.
struct KinkSlopes {
    double sl, sr, mid;
};
template <bool STD = false>
ACC_HD KinkSlopes pm_kink_slopes(const PMParams& pp, double target) {
    KinkSlopes k;
    const bool quadratic = STD || pp.quadratic, relative = STD || pp.relative;
    if (quadratic) {
        if (relative && !(target > ACC_PREC3)) { k.sl = 0.0; k.sr = 0.0; k.mid = 0.0; return k; }  // D^2 form: treated as smooth
        k.sl = -pp.markup; k.sr = pp.markup; k.mid = 0.0;
    } else if (relative) {
        k.sl = -pp.pen_small / target; k.sr = pp.pen_large / target; k.mid = (pp.pen_large - pp.pen_small) / 2 / target;
    } else {
        k.sl = -pp.pen_small; k.sr = pp.pen_large; k.mid = (pp.pen_large - pp.pen_small) / 2;
    }
    return k;
}
// in the solver's scaled variables (x = D / Z*, objective / U*): slope * (Z* / U*)
ACC_HD double pm_scaled_slope(double slope, double z_over_u) { return slope * z_over_u; }

// ---- work vectors of one purchasing manager (shared memory) ----
// Slot-major: the V f64 values of connection slot r are contiguous, so one index computation per slot serves every
// vector (the vector is a compile-time offset) and one base register serves them all.  With an even V one f64 of
// padding follows every eight slots, which keeps the 16 eight-byte accesses of a half-warp on different banks.
// V = 20: 9 of the solver (x, x0, g, a, d, b, gl, xbest, xu) + 8 price-curve values + the two summands of the ordered
// sums + the x tolerance; one more with deviation_penalty (the penalty target is then the last demand request instead of
// Z*).  Connection ids are a separate int vector.
// DVAL = false drops the D_r summand (the ordered sum then recomputes x_r Z*_r on the way) and the x tolerance (divided
// out where it is tested): two vectors less, which is what lets two CTAs of the large solver groups (748 connections:
// 111 KB instead of 123 KB) share an SM.
// Number of f64 per slot for a given number of vectors: a multiple of four would put lanes i and i + 4 of a half-warp on the
// same banks whatever the padding (stride 8 i words mod 32), so it gets one unused f64 per slot.
ACC_HD constexpr int pm_slot_doubles(int vectors) { return vectors % 4 == 0 ? vectors + 1 : vectors; }
template <int V>
ACC_HD int pm_slot_off(int i) { return (V % 2 == 0) ? i * V + (i >> 3) : i * V; }
ACC_HD size_t pm_workspace_bytes(int n_max, bool separate_target, bool dval) {
    const int v = pm_slot_doubles(18 + (separate_target ? 1 : 0) + (dval ? 2 : 0));
    const size_t doubles = (size_t)n_max * v + (v % 2 == 0 ? (size_t)(n_max >> 3) + 1 : 0);
    return doubles * sizeof(double) + (size_t)n_max * sizeof(int);
}

template <bool SEP, bool STD, bool DVAL>
struct PMWorkT {
    enum { NVEC = 18 + (SEP ? 1 : 0) + (DVAL ? 2 : 0), V = pm_slot_doubles(NVEC), TGT = SEP ? NVEC - 1 : 9, COL_X = 0, COL_X0 = 1, COL_G = 2, COL_ZF = 9, COL_TERM = 17, COL_DVAL = 18,
           COL_TOL = 19, HAS_DVAL = DVAL ? 1 : 0 };
    template <int K>
    struct Col {
        double* base;
        ACC_HD double& operator[](int i) const { return base[pm_slot_off<V>(i) + K]; }
    };
    struct W {
        Col<0> x; Col<1> x0; Col<2> g; Col<3> a; Col<4> d; Col<5> b; Col<6> gl; Col<7> xbest;  // sqp_core work vectors
        Col<8> xu;                            // upper bounds (lower bounds are all zero, :620)
        Col<9> zf;                            // x tolerance = quantum of the demand request over the variable's scale (:640-644)
        Col<TGT> tgt;                         // kink position in D_r
        Col<COL_TOL> tolv;                    // (only with DVAL) the x tolerance, divided out once in the set-up pass
        double tol_c;
        PMParams pp;
        ACC_HD double lower(int) const { return 0.0; }
        ACC_HD double tol(int i) const { return DVAL ? tolv[i] : tol_c / zf[i]; }
        // kink of the transport penalty (sqp_core.h "Kinks"); a[i] = -(Z*_i / U*) is the constraint gradient
        ACC_HD bool kinked() const { return true; }
        ACC_HD int kside(int i, double xi) const { const double D = xi * zf[i], t = tgt[i]; return D < t ? -1 : (D > t ? 1 : 0); }
        ACC_HD double kpos(int i) const { return SEP ? tgt[i] / zf[i] : 1.0; }
        ACC_HD double ksl(int i) const { return pm_scaled_slope(pm_kink_slopes<STD>(pp, tgt[i]).sl, -a[i]); }
        ACC_HD double ksr(int i) const { return pm_scaled_slope(pm_kink_slopes<STD>(pp, tgt[i]).sr, -a[i]); }
        ACC_HD double kmid(int i) const { return pm_scaled_slope(pm_kink_slopes<STD>(pp, tgt[i]).mid, -a[i]); }
    } w;
    // price-curve constants per purchasing connection
    Col<9> zf;               // to_float(Z*): scale of the variable (:184-188)
    Col<TGT> tgt;            // transport penalty target (Z*, or the last demand request with deviation_penalty)
    Col<10> addX;            // expected_additional_production (:343-349)
    Col<11> Xhatf;           // seller constants (SellerPub); c1 = c2 / 2 exactly
    Col<12> lamXf;
    Col<13> c2;
    Col<14> base;            // n_bar [- pen(X_exp)/X_exp]: E_n_r without its X_new term (:255-271)
    Col<15> n_co;            // reservation price (:283-289); the kink position of the price curve D_r_min is max(0, lamX - addX)
    Col<16> slope;           // (n_bar_min - n_co) / D_r_min in the linear regime; NaN marks the cropped regime (n_co > n_bar_min)
    Col<17> term;            // per-connection summands of the current evaluation (cost term, D_r), read by the ordered sums
    Col<18> dval;            // (only with DVAL)
    int* eid;                // connection id of slot r
    ACC_HD double& at(int r, int col) const { return term.base[pm_slot_off<V>(r) + col]; }  // vector chosen at run time
    ACC_HD int& idx(int r) const { return *reinterpret_cast<int*>(&term[r]); }  // compacted position -> incoming position (set-up only)
    ACC_HD void carve(unsigned char* mem, int n_max) {
        double* p = reinterpret_cast<double*>(mem);
        w.x.base = p; w.x0.base = p; w.g.base = p; w.a.base = p; w.d.base = p; w.b.base = p; w.gl.base = p; w.xbest.base = p;
        w.xu.base = p; w.zf.base = p; w.tgt.base = p; w.tolv.base = p;
        zf.base = p; tgt.base = p; addX.base = p; Xhatf.base = p; lamXf.base = p; c2.base = p; base.base = p; n_co.base = p;
        slope.base = p; term.base = p; dval.base = p;
        eid = reinterpret_cast<int*>(mem + pm_workspace_bytes(n_max, SEP, DVAL) - (size_t)n_max * sizeof(int));
    }
};

// n_r, grad_n_r, transport_penalty, partial_D_r_transport_penalty for connection r at demand D.
// Same expressions as the reference, with two exact shortcuts: constant sub-expressions (dn/(2 lamX), dn/lamX,
// the linear-regime slope) are computed once per connection, and 0/X_new terms (no production extension) are not
// divided out (0.0 / x == 0.0 and base + 0.0 == base bit for bit).
struct CurvePoint {
    double n, gn, tp, tpp;
};
template <bool STD, class PW>
ACC_HD CurvePoint curve_at(const PW& k, const PMParams& pp, int r, double D) {
    const bool quadratic = STD || pp.quadratic, relative = STD || pp.relative;
    CurvePoint c;
    // expected_average_price_E_n_r (:243-275) and its gradient (:305-321)
    double Xn = D + k.addX[r];
    if (Xn > k.Xhatf[r]) Xn = k.Xhatf[r];
    double E = k.base[r];
    double gE = 0.0;
    const double lam = k.lamXf[r];
    if (Xn > lam && Xn > 0.0) {  // in production extension (:151,:175)
        const double dx = Xn - lam;
        const double pen = stdmax(0.0, 0.5 * k.c2[r] * dx * dx);
        E = k.base[r] + pen / Xn;
        gE = k.c2[r] * dx / Xn - pen / Xn / Xn;
    }
    // n_r (:277-303), grad_n_r (:351-379)
    const double sl = k.slope[r];
    if (sl == sl) {  // n_co <= n_bar_min: linear regime below D_r_min
        if (D < stdmax(0.0, lam - k.addX[r])) {  // D_r_min, :283
            c.n = k.n_co[r] + sl * D;
            c.gn = sl;
        } else {
            c.n = E;
            c.gn = gE;
        }
    } else if (E <= k.n_co[r]) {  // curve cropped from below with n_co
        c.n = k.n_co[r];
        c.gn = 0.0;
    } else {
        c.n = E;
        c.gn = gE;
    }
    // transport_penalty (:381-409), partial_D_r_transport_penalty (:411-451)
    const double t = k.tgt[r];
    if (quadratic) {
        double marg = 0.0;
        if (D < t) marg = -pp.markup; else if (D > t) marg = pp.markup;
        if (relative) {
            if (t > ACC_PREC3) {
                const double qd = (D - t) * pp.pen_large / (t * t);  // shared by both expressions of the reference
                c.tp = (D - t) * (qd / 2 + marg);
                c.tpp = qd + marg;
            } else {
                // to_float(transport_penalty_large / 2 + Price(marg_penalty)): typed Price arithmetic
                c.tp = D * D * pf(pdiv(mkp(pp.pen_large), 2.0) + mkp(marg));
                c.tpp = D * pp.pen_large + marg;
            }
        } else {
            c.tp = (D - t) * ((D - t) * pp.pen_large / 2 + marg);
            c.tpp = (D - t) * pp.pen_large + marg;
        }
    } else {
        if (relative) {
            if (D < t) c.tpp = -pp.pen_small / t;
            else if (D > t) c.tpp = pp.pen_large / t;
            else c.tpp = (pp.pen_large - pp.pen_small) / 2 / t;
            c.tp = c.tpp * (D - t) / t;
        } else {
            if (D < t) c.tpp = -pp.pen_small;
            else if (D > t) c.tpp = pp.pen_large;
            else c.tpp = (pp.pen_large - pp.pen_small) / 2;
            c.tp = c.tpp * (D - t);
        }
    }
    return c;
}

// Objective / constraint evaluation in the convention of sqp_core: minimise F = costs / U*,
// constraint c = (desired - sum D_r) / U* == 0 whose gradient -Z*_r / U* is constant (filled once).
// The objective is separable, so it is handed to the solver coordinate by coordinate (accsqp::EvalTraits).
template <bool STD, class Ctx, class PW>
struct PMEval {
    Ctx& cx;
    PW& k;
    const PMParams& pp;
    int n;
    double desired_f;
    // connection r at x_r = xi: its cost term (kept for the ordered sum) and the gradient entry
    ACC_HD void coord(int r, double xi) const {
        const double D = xi * k.zf[r];
        const CurvePoint cp = curve_at<STD>(k, pp, r, D);
        k.term[r] = cp.n * D + cp.tp;
        if (PW::HAS_DVAL) k.dval[r] = D;
        k.w.g[r] = (k.zf[r] * (cp.gn * D + cp.n + cp.tpp)) / pp.Uf;
    }
    // costs and purchase summed strictly left to right, like the reference's loops (:203-217, :221-240)
    template <class X>
    ACC_HD void finish(const X& x, double* F, double* c) const {
        const PW& kk = k;
        double s[2];
        // summand 0: the cost term; summand 1: D_r = x_r Z*_r (x is the solver's x vector), stored or recomputed
        if (PW::HAS_DVAL) {
            cx.template seqsum<2>(n, [&](int r, int j) { return kk.at(r, PW::COL_TERM + j); }, s);
        } else {
            cx.template seqsum<2>(n, [&](int r, int j) {
                const double a = kk.at(r, j == 0 ? PW::COL_TERM : PW::COL_X), b = kk.at(r, PW::COL_ZF);
                return j == 0 ? a : a * b;
            }, s);
        }
        *F = s[0] / pp.Uf;
        *c = (desired_f - s[1]) / pp.Uf;
    }
    template <class X>
    ACC_HD void operator()(const X& x, bool, double* F, double* c) const {
        cx.foreach(n, [&](int r) { coord(r, x[r]); });
        finish(x, F, c);
    }
};

struct CountSum {
    i64 n, s;
    static ACC_HD CountSum identity() { CountSum r; r.n = 0; r.s = 0; return r; }
    static ACC_HD CountSum combine(CountSum a, CountSum b) { CountSum r; r.n = a.n + b.n; r.s = a.s + b.s; return r; }
};

struct WriteBack {  // sums of the write-back pass: demand quantity (exact), expected costs, transport penalty, demand value
    i64 q;
    double costs, pen, value;
    static ACC_HD WriteBack identity() { WriteBack r; r.q = 0; r.costs = 0.0; r.pen = 0.0; r.value = 0.0; return r; }
    static ACC_HD WriteBack combine(WriteBack a, WriteBack b) {
        WriteBack r; r.q = a.q + b.q; r.costs = a.costs + b.costs; r.pen = a.pen + b.pen; r.value = a.value + b.value; return r;
    }
};

struct PMResult {
    Flow demand;           // demand_D_
    double expected_costs; // expected_costs_
    double optimized;      // optimized_value_
    double penalty;        // total_transport_penalty_
    int iters, evals;
};

// What the set-up pass reads per incoming connection through index chains (connection id -> delay -> conveyor slot).
// The caller's `gather(i)` resolves it; the group contexts call it for two connections of a thread before either is
// worked on (reduce_staged), so the dependent loads of both are in flight together.
struct PMGather {
    int e, seller;   // connection id, seller row (SellerTable) of the i-th incoming connection (seller ascending)
    i64 last_q;      // last shipment on the connection (:346)
    i64 Z_q;         // initial flow Z*
    i64 lastD_q;     // last demand request (read with deviation_penalty only)
    SellerPub sp;    // the seller's published supply parameters (SalesManager.h:38-43)
    double rr, npeE; // expected production / production and the extension penalty at the expected production, for minimum passage 1
};

// Inputs of one purchasing manager, gathered by the caller-provided accessors:
//   in_edge(i) -> connection id of the i-th incoming connection;  gather(i) -> PMGather of it
template <bool SEP, bool STD, class Ctx, class PW, class InEdge, class Gather, class MinPassage, class SendD>
ACC_HD PMResult purchase_one(Ctx& cx, int nin, InEdge in_edge, Gather gather, MinPassage min_passage, SendD send_D, PW& k, const ModelP& mp, const SectorP& sec, i64 Ustar_q,
                             i64 Sstar_q, i64 S_q, i64 Ut_q, i64 deficit_q, const EventSink& events, int ev_sector, int ev_agent) {
    PMResult res;
    res.demand = mkflow(0, 0.0); res.expected_costs = 0.0; res.optimized = 0.0; res.penalty = 0.0; res.iters = 0; res.evals = 0;
    // desired purchase, :587-593
    const i64 shortage = mkq(qf(deficit_q) * mp.dt) + Sstar_q - S_q;
    i64 desired = Ut_q + qdiv(shortage, shortage > 0 ? sec.refill : sec.withdraw);
    if (desired <= 0) {  // :595-600
        cx.foreach(nin, [&](int i) { send_D(in_edge(i), (i64)0, (i64)0); });
        return res;
    }
    PMParams pp;
    k.w.tol_c = ACC_PREC3 * mp.prec_adj;
    pp.Uf = qf(Ustar_q);
    pp.markup = pf(sec.markup);
    pp.pen_large = pf(mp.pen_large); pp.pen_small = pf(mp.pen_small);
    pp.quadratic = mp.quadratic_penalty; pp.relative = mp.relative_penalty; pp.limited_by_markup = mp.limited_by_markup;
    k.w.pp = pp;

    // One gather pass per connection: bounds (:609-650), price-curve constants (:279-289,:323-332) and start value,
    // written to slot r.  Returns D_r_max; <= 0 means the connection takes no part in the optimisation.
    auto fill = [&](const PMGather& in, int r) -> i64 {
        const int e = in.e;
        const SellerPub& sp = in.sp;
        k.eid[r] = e;
        if (sp.Xhat_q <= 0) return 0;                                           // :604-605
        const double nu = min_passage(e);
        const bool open_route = nu == 1.0;                                      // then nu * x == x bit for bit and the seller's own values apply
        const double Xe = nu * sp.Xexpf;                                        // expected_production, :334-341
        const double X = sp.Xf;
        const double Zl = qf(in.last_q);
        const double rr = open_route ? in.rr : spt_ratio(Xe, X);
        const double add = rr * (X - Zl);                                       // expected_additional_production, :343-349
        double Xmax = sp.Xhatf;
        Xmax *= nu;                                                             // :612-614
        const double upper = Xmax - add;
        const i64 dmax = mkq(upper);                                            // :622
        if (dmax <= 0) return 0;                                                // :646-648
        const double zf = qf(in.Z_q);
        const double lamXf = sp.lamXf;
        k.zf[r] = zf;
        if (SEP) k.tgt[r] = qf(in.lastD_q);
        k.addX[r] = add;
        k.Xhatf[r] = sp.Xhatf;
        k.lamXf[r] = lamXf;
        k.c2[r] = sp.c2;
        // n_r constants, :279-289 (penalty terms only divided out inside the production extension)
        const double Dmin = stdmax(0.0, lamXf - add);
        const double Xmin = Dmin + add;
        double npeMin = 0.0;
        const double npeE = open_route ? in.npeE : spt_extension_penalty(sp.c1, Xe, lamXf);
        if (Xmin > lamXf && Xmin > 0.0) npeMin = stdmax(0.0, sp.c1 * (Xmin - lamXf) * (Xmin - lamXf)) / Xmin;
        const double nbar_min = sp.nbarf - npeE + npeMin;
        // E_n_r's four cases collapse to  base [+ pen(X_new)/X_new]  with base = n_bar [- pen(X_exp)/X_exp]
        k.base[r] = (Xe > 0.0) ? sp.nbarf - npeE : sp.nbarf;
        double n_co = sp.n_co;                                                  // calc_n_co, :323-332
        if (pp.limited_by_markup) n_co = stdmax(n_co, nbar_min - pp.markup * Dmin);
        k.n_co[r] = n_co;
        k.slope[r] = (n_co <= nbar_min) ? (nbar_min - n_co) / Dmin : (double)NAN;
        // bounds, start value, xtol, :620-644; constraint gradient, :209
        k.w.xu[r] = upper / zf;
        k.w.x[r] = stdmin(upper, stdmax(0.0, Xe - add)) / zf;
        k.w.a[r] = -(zf / pp.Uf);
        if (PW::HAS_DVAL) k.w.tolv[r] = k.w.tol_c / zf;
        return dmax;
    };
    // the same pass counts the connections that take part and sums their D_r_max (integers: any order)
    const CountSum part = cx.template reduce_staged<CountSum>(nin, gather, [&](int i, const PMGather& in) {
        const i64 dmax = fill(in, i);
        k.w.d[i] = (double)dmax;  // exact below 2^53 quanta
        CountSum r; r.n = 0; r.s = 0;
        if (dmax <= 0) send_D(k.eid[i], (i64)0, (i64)0);
        else { r.n = 1; r.s = dmax; }
        return r;
    });
    const int n = (int)part.n;
    if (n == 0) return res;  // :652-655
    if (n != nin) {  // some connections dropped out: rebuild the slots in compacted order (slot r <= source idx[r])
        cx.compact(nin, [&](int i) { return k.w.d[i] > 0.0; }, [&](int m, int i) { k.idx(m) = i; });
        cx.foreach(n, [&](int r) { fill(gather(k.idx(r)), r); });
    }
    const i64 max_purchase = part.s;
    if (desired > max_purchase) desired = max_purchase;  // :658-660

    // local optimisation, :690-702 with the restart logic of :465-561
    PMEval<STD, Ctx, PW> ev{cx, k, pp, n, qf(desired)};
    accsqp::Params sp_;
    sp_.n = n; sp_.ctype = accsqp::CON_EQ; sp_.ctol = ACC_PREC3; sp_.maxeval = mp.maxeval; sp_.max_iter = ACCSQP_MAX_ITER;
    accsqp::Stats st;
    int restarts = 0;
    for (;;) {
        const int ret = accsqp::sqp_minimize(cx, ev, k.w, sp_, st);
        res.iters += st.iters; res.evals += st.evals;
        bool again = false;
        if (ret == accsqp::SQP_FAILURE) {                 // "bug: more than iter SQP iterations", :480-504
            again = ++restarts < 10;
        } else if (ret == accsqp::SQP_MAXEVAL_REACHED) {  // :526-546
            if (cx.leader()) events.push(EV_OPTIMIZER_MAXITER, ev_sector, ev_agent, 0.0);
            again = ++restarts < 10;
        }
        if (!again) break;
        cx.foreach(n, [&](int r) { k.w.x[r] = k.w.xbest[r]; });  // restart from the best point (the result is read from xbest below)
    }
    res.optimized = (-st.minf) * pp.Uf;  // unscaled_objective(opt.optimized_value()), :509

    // distribute demand requests, :704-737.  The three value sums of this pass (expected costs, transport penalty, value of
    // the demand) are observables only -- nothing of the model reads them back (include/model/Storage.h:108-140) -- so they are
    // folded in the group's fixed reduction order with the integer sum of the requests, not in a strictly sequential chain.
    const WriteBack wb = cx.template reduce<WriteBack>(n, [&](int r) {
        const double D = k.w.xbest[r] * k.zf[r];
        const i64 q = mkq(D);
        // with deviation_penalty the penalty target is last_demand_request_D, which the reference has already
        // overwritten with the new request when it accumulates the costs (:723 before :729-730)
        if (SEP) k.tgt[r] = qf(q);
        const CurvePoint cp = curve_at<STD>(k, pp, r, D);
        const i64 p = q > 0 ? mkp(cp.n) : 0;
        send_D(k.eid[r], q, p);
        WriteBack t;
        t.q = q; t.costs = cp.n * D + cp.tp; t.pen = cp.tp; t.value = q > 0 ? q_times_p(q, p) : 0.0;
        return t;
    });
    const i64 dq = wb.q;
    const double costs = wb.costs, pen = wb.pen, dv = wb.value;
    res.demand = mkflow(dq, dv);
    res.expected_costs = costs;
    res.penalty = pen;
    return res;
}

}  // namespace acc

namespace accsqp {
template <bool STD, class Ctx, class PW>
struct EvalTraits<acc::PMEval<STD, Ctx, PW>> {
    enum { PER_COORD = 1 };
};
}  // namespace accsqp
