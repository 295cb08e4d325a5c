// model_core.cuh -- the per-agent arithmetic of one Acclimate timestep, restated for cooperative
// thread groups (see coop.cuh).  Each routine cites the reference code it follows; every typed
// intermediate of the reference (Price, FlowQuantity, Quantity) is re-quantised at the same point.
//
// Scope of the restatement (what acc_create accepts, everything else is rejected with an error):
//   naive_expectations true or false (with or without always_extend_expected_demand_curve: seller_phase2 carries the
//   demand-curve extension and its bisection), respect_markup_in_production_extension = false (the reference never
//   assigns it, parameters.h:46), tax = 0 (Government is dead code at the pinned commit), local SLSQP purchasing
//   optimisation only, transport as one connection per region pair (const / csv / centroids) or as multi-hop routes
//   over geo entities ("network": per-hop delay, passage forcing and overflow, applied in the delivery pass).
#pragma once

#include "coop.cuh"
#include "stage.cuh"

namespace acc {

// reference include/ModelRun.h:39-55
enum EventType : int {
    EV_NO_CONSUMPTION = 0,
    EV_STORAGE_UNDERRUN,
    EV_STORAGE_OVERRUN,
    EV_NO_PRODUCTION_SUPPLY_SHORTAGE,
    EV_NO_PRODUCTION_DEMAND_QUANTITY_SHORTAGE,
    EV_NO_PRODUCTION_DEMAND_VALUE_SHORTAGE,
    EV_NO_PRODUCTION_HIGH_COSTS,
    EV_NO_EXP_PRODUCTION_SUPPLY_SHORTAGE,
    EV_NO_EXP_PRODUCTION_DEMAND_QUANTITY_SHORTAGE,
    EV_NO_EXP_PRODUCTION_DEMAND_VALUE_SHORTAGE,
    EV_NO_EXP_PRODUCTION_HIGH_COSTS,
    EV_DEMAND_FULFILL_HISTORY_UNDERFLOW,
    EV_OPTIMIZER_TIMEOUT,
    EV_OPTIMIZER_MAXITER,
    EV_OPTIMIZER_ROUNDOFF_LIMITED,
};

struct ModelP {  // reference include/parameters.h:33-84 (the keys this path reads) + delta_t
    double dt;               // to_float(delta_t)
    i64 pen_small, pen_large;  // transport_penalty_small / large (Price)
    double min_storage;
    int generic_width;       // cheapest_price_range_generic_size
    i64 width;               // cheapest_price_range_width (Price)
    int quadratic_penalty, relative_penalty, limited_by_markup, deviation_penalty, preserve_seller_price;
    int maxeval;             // optimization_maxiter
    int prec_adj;            // optimization_precision_adjustment (int, parameters.h:61)
    int utilitarian;         // consumers.<..>.bool_utilitarian (ModelInitializer.cpp:136)
    int naive_expectations, always_extend;  // naive_expectations, always_extend_expected_demand_curve (ModelInitializer.cpp:1024-1025)
};

struct SectorP {  // reference Sector.h + Parameters::SectorParameters
    double omega;       // upper_storage_limit_omega
    double psi;         // to_float(initial_storage_fill_factor_psi) (a Time)
    int transport_type; // 0 aviation, 1 immediate, 2 roadsea
    double eps;         // supply_elasticity
    i64 dn;             // price_increase_production_extension
    i64 dn_est;         // estimated_price_increase_production_extension
    i64 markup;         // initial_markup
    double refill;      // to_float(target_storage_refill_time)
    double withdraw;    // to_float(target_storage_withdraw_time)
};

struct Event {
    i64 step;
    int type, a, b;
    double value;
};

struct EventSink {
    Event* buf;
    int* count;
    int cap;
    i64 step;
    ACC_HD void push(int type, int a, int b, double value) const {
#if defined(__CUDA_ARCH__)
        const int k = atomicAdd(count, 1);
#else
        const int k = (*count)++;
#endif
        if (k < cap) { Event e; e.step = step; e.type = type; e.a = a; e.b = b; e.value = value; buf[k] = e; }
    }
};

// ------------------------------------------------------------------------------------------------
// SalesManager cost curves (reference src/model/SalesManager.cpp:645-702)
// ------------------------------------------------------------------------------------------------
// calc_marginal_production_extension_penalty, :674-683:  Price / FlowQuantity -> PriceGrad (f64), * FlowQuantity -> Price
ACC_HD i64 marginal_extension_penalty(i64 x, i64 lamX, i64 dn) {
    if (x <= lamX) return 0;
    return mkp((pf(dn) / qf(lamX)) * qf(x - lamX));
}
// calc_marginal_production_costs, :685-693
ACC_HD i64 marginal_costs(i64 x, i64 lamX, i64 dn, i64 n_c) {
    if (x <= lamX) return n_c;
    return n_c + marginal_extension_penalty(x, lamX, dn);
}
// calc_total_production_costs, :645-672
ACC_HD double total_costs(i64 x, i64 lamX, i64 dn, i64 n_c) {
    const double base = q_times_p(x, n_c);
    if (x <= lamX) return base;
    const i64 two_lam = mkq(2.0 * qf(lamX));                       // 2 * FlowQuantity -> FlowQuantity
    const i64 p1 = mkp((pf(dn) / qf(two_lam)) * qf(x - lamX));     // PriceGrad * FlowQuantity -> Price
    return base + pf(p1) * qf(x - lamX);                           // Price * FlowQuantity -> FlowValue
}
// analytic_solution_in_production_extension, :695-702
ACC_HD i64 analytic_extension(i64 lamX, i64 dn, i64 n_c, i64 price) { return qmul(lamX, 1.0 + ratio(price - n_c, dn)); }

// ------------------------------------------------------------------------------------------------
// One seller's demand book: totals and the first connection that cannot be served completely.
// Restates the sort of SalesManager::calc_production_X (:105-127) plus the scan loop shared by
// calc_supply_distribution_scenario (:301-316) and calc_expected_supply_distribution_scenario
// (:473-497) WITHOUT sorting: every decision of the scan depends only on int64 quantities and prices,
// the stop predicate is monotone along the sorted order, so the stop position is found by bisection
// over (price, quantity, index) thresholds with exact integer reductions.  Total order used:
// price desc, quantity desc, index asc (the reference's std::sort leaves ties unspecified).
// ------------------------------------------------------------------------------------------------
struct BookTotals {
    i64 sq;     // sum of quantities (all connections)
    double sv;  // sum of values over connections with q > 0
    i64 cnt;    // connections with q > 0
    i64 pmax;   // highest price among q > 0
    i64 pmin;   // lowest price among q > 0
    static ACC_HD BookTotals identity() { BookTotals r; r.sq = 0; r.sv = 0.0; r.cnt = 0; r.pmax = -1; r.pmin = 0x7fffffffffffffffLL; return r; }
    static ACC_HD BookTotals combine(BookTotals a, BookTotals b) {
        BookTotals r;
        r.sq = a.sq + b.sq; r.sv = a.sv + b.sv; r.cnt = a.cnt + b.cnt;
        r.pmax = a.pmax > b.pmax ? a.pmax : b.pmax;
        r.pmin = a.pmin < b.pmin ? a.pmin : b.pmin;
        return r;
    }
};

template <class Ctx>
ACC_HD BookTotals book_totals(Ctx& cx, int n, const i64* Dq, const i64* Dp) {
    return cx.template reduce<BookTotals>(n, [&](int i) {
        BookTotals r = BookTotals::identity();
        const i64 q = Dq[i];
        if (q > 0) { const i64 p = Dp[i]; r.sq = q; r.sv = q_times_p(q, p); r.cnt = 1; r.pmax = p; r.pmin = p; }
        return r;
    });
}

struct QV {
    i64 q; double v; i64 cnt;
    static ACC_HD QV identity() { QV r; r.q = 0; r.v = 0.0; r.cnt = 0; return r; }
    static ACC_HD QV combine(QV a, QV b) { QV r; r.q = a.q + b.q; r.v = a.v + b.v; r.cnt = a.cnt + b.cnt; return r; }
};
struct MaxI64 {
    i64 s;
    static ACC_HD MaxI64 identity() { MaxI64 r; r.s = -0x7fffffffffffffffLL; return r; }
    static ACC_HD MaxI64 combine(MaxI64 a, MaxI64 b) { return b.s > a.s ? b : a; }
};
struct MinI64 {
    i64 s;
    static ACC_HD MinI64 identity() { MinI64 r; r.s = 0x7fffffffffffffffLL; return r; }
    static ACC_HD MinI64 combine(MinI64 a, MinI64 b) { return b.s < a.s ? b : a; }
};

struct ScanResult {
    int all_served;   // no connection stopped the scan
    Flow cum;         // sum of the completely served demand requests (scan order prefix)
    i64 p_stop;       // price of the connection not served completely
    int has_prev;     // the stop connection is not the first one
    i64 p_prev;       // price of the cheapest completely served connection
};

// stop predicate of the scan for a cumulative quantity `cum` (including the candidate) and its price
ACC_HD bool scan_stops(i64 cum, i64 price, i64 xhat, i64 lamX, i64 dn, i64 n_c) {
    if (cum > xhat) return true;                              // :303 / :475
    return price < marginal_costs(cum, lamX, dn, n_c);        // :308-311 / :485-492
}

template <class Ctx>
ACC_HD ScanResult seller_scan(Ctx& cx, int n, const i64* Dq, const i64* Dp, const BookTotals& tot, i64 xhat, i64 lamX, i64 dn, i64 n_c) {
    ScanResult res;
    res.p_stop = 0; res.has_prev = 0; res.p_prev = 0;
    // the predicate is monotone, so the scan runs through iff the LAST connection passes
    if (!scan_stops(tot.sq, tot.pmin, xhat, lamX, dn, n_c)) {
        res.all_served = 1;
        res.cum = mkflow(tot.sq, tot.sv);
        res.has_prev = 1;
        res.p_prev = tot.pmin;
        return res;
    }
    res.all_served = 0;
    // level 1: highest price P with  stop(G(P), P),  G(P) = sum of q over connections with p >= P
    i64 lo = tot.pmin, hi = tot.pmax;  // invariant: stop true at lo
    while (lo < hi) {
        const i64 mid = lo + (hi - lo + 1) / 2;
        const i64 g = cx.isum(n, [&](int i) { return (Dq[i] > 0 && Dp[i] >= mid) ? Dq[i] : (i64)0; });
        if (scan_stops(g, mid, xhat, lamX, dn, n_c)) lo = mid; else hi = mid - 1;
    }
    // the price level that contains the stop: highest existing price <= lo
    const i64 L = cx.template reduce<MaxI64>(n, [&](int i) { MaxI64 r = MaxI64::identity(); if (Dq[i] > 0 && Dp[i] <= lo) r.s = Dp[i]; return r; }).s;
    const QV above = cx.template reduce<QV>(n, [&](int i) { QV r = QV::identity(); if (Dq[i] > 0 && Dp[i] > L) { r.q = Dq[i]; r.v = q_times_p(Dq[i], Dp[i]); r.cnt = 1; } return r; });
    // level 2: inside price level L, by quantity descending
    const MaxI64 qhi_ = cx.template reduce<MaxI64>(n, [&](int i) { MaxI64 r = MaxI64::identity(); if (Dq[i] > 0 && Dp[i] == L) r.s = Dq[i]; return r; });
    const MinI64 qlo_ = cx.template reduce<MinI64>(n, [&](int i) { MinI64 r = MinI64::identity(); if (Dq[i] > 0 && Dp[i] == L) r.s = Dq[i]; return r; });
    i64 qlo = qlo_.s, qhi = qhi_.s;  // stop true at the end of sub-level qlo (it ends the level)
    while (qlo < qhi) {
        const i64 mid = qlo + (qhi - qlo + 1) / 2;
        const i64 g = cx.isum(n, [&](int i) { return (Dp[i] == L && Dq[i] >= mid) ? Dq[i] : (i64)0; });
        if (scan_stops(above.q + g, L, xhat, lamX, dn, n_c)) qlo = mid; else qhi = mid - 1;
    }
    const i64 Qs = cx.template reduce<MaxI64>(n, [&](int i) { MaxI64 r = MaxI64::identity(); if (Dp[i] == L && Dq[i] > 0 && Dq[i] <= qlo) r.s = Dq[i]; return r; }).s;
    const QV bigger = cx.template reduce<QV>(n, [&](int i) { QV r = QV::identity(); if (Dp[i] == L && Dq[i] > Qs) { r.q = Dq[i]; r.v = q_times_p(Dq[i], L); r.cnt = 1; } return r; });
    // level 3: connections with identical (L, Qs): the j-th of them (index order) is the first that stops
    const i64 same = cx.isum(n, [&](int i) { return (Dp[i] == L && Dq[i] == Qs) ? (i64)1 : (i64)0; });
    i64 j = 1;
    while (j < same && !scan_stops(above.q + bigger.q + j * Qs, L, xhat, lamX, dn, n_c)) ++j;
    const i64 served_same = j - 1;  // completely served connections with exactly (L, Qs)
    // value of those: served_same equal terms folded in index order
    double v_same = 0.0;
    const double term = q_times_p(Qs, L);
    for (i64 k = 0; k < served_same; ++k) v_same += term;
    res.cum = mkflow(above.q + bigger.q + served_same * Qs, (above.v + bigger.v) + v_same);
    res.p_stop = L;
    const i64 n_before = above.cnt + bigger.cnt + served_same;
    res.has_prev = n_before > 0 ? 1 : 0;
    if (bigger.cnt + served_same > 0) {
        res.p_prev = L;
    } else if (above.cnt > 0) {
        res.p_prev = cx.template reduce<MinI64>(n, [&](int i) { MinI64 r = MinI64::identity(); if (Dq[i] > 0 && Dp[i] > L) r.s = Dp[i]; return r; }).s;
    }
    return res;
}

// ------------------------------------------------------------------------------------------------
// CapacityManager::get_possible_production_X_hat_intern (reference src/model/CapacityManager.cpp:72-102)
// for one firm: loops the firm's input storages in the reference's container order.
// ------------------------------------------------------------------------------------------------
struct StorageView {
    // state, indexed by storage
    i64* S_q; double* S_v;          // content_S
    i64* Ip_q; double* Ip_v;        // input_flow_I[other_register]: what arrived during the previous step
    i64* In_q; double* In_v;        // input_flow_I[current_register]: what arrives during this step
    i64* Il_q; double* Il_v;        // input_flow_I[2]: last input flow (output surface)
    i64* U_q; double* U_v;          // used_flow_U
    i64* Ut_q; double* Ut_v;        // desired_used_flow_U_tilde
    i64* T_q; double* T_v;          // in-transit flow on all incoming connections (queues + overflow)
    i64* dlv_q;                     // sum of this step's deliveries (for the flow deficit)
    i64* imp_q; double* imp_v;      // non-domestic part of this step's deliveries (region import)
    // constants
    const i64* Istar_q; const double* Istar_v;  // initial_input_flow_I_star
    const i64* Sstar_q; const double* Sstar_v;  // initial_content_S_star
    const double* a;                // technology coefficient (firms) / consumption_price_elasticity (consumers)
    const i64* defc;                // sum over incoming connections of Z* (tau + 1)
    const int* agent; const int* sector;
};

ACC_HD Flow possible_use(const StorageView& st, int s, double dt, bool estimate) {
    // Storage::get_possible_use_U_hat / estimate_possible_use_U_hat, Storage.cpp:51-59
    const Flow in = estimate ? mkflow(st.In_q[s], st.In_v[s]) : mkflow(st.Ip_q[s], st.Ip_v[s]);
    return fcopy(fadd(stock_over_time(mkflow(st.S_q[s], st.S_v[s]), dt), in));
}

// CapacityManager::get_possible_production_X_hat_intern (CapacityManager.cpp:72-102), one input storage per thread of
// the group: the unit commodity costs are a sum of int64 Price quanta and the production ratio is a minimum, so neither
// depends on the order in which the storages are visited -- the result is the sequential loop's, bit for bit.
struct CapacityTerm {
    i64 ucc;      // sum of the inputs' unit costs
    double pmin;  // smallest possible-use ratio seen
    static ACC_HD CapacityTerm identity() { CapacityTerm r; r.ucc = 0; r.pmin = 1e308; return r; }
    static ACC_HD CapacityTerm combine(CapacityTerm a, CapacityTerm b) {  // a NaN ratio (0 / 0) never wins, as in `if (tmp < p_hat)`
        CapacityTerm r; r.ucc = a.ucc + b.ucc; r.pmin = b.pmin < a.pmin ? b.pmin : a.pmin; return r;
    }
};
template <class Ctx>
ACC_HD Flow possible_production(Ctx& cx, const StorageView& st, const int* list, int n_st, Flow Xstar, double lambda, double beta, i64 var_costs,
                                double dt, bool estimate) {
    const CapacityTerm t = cx.template reduce<CapacityTerm>(n_st, [&](int k) {
        const int s = list[k];
        const Flow u = possible_use(st, s, dt, estimate);
        CapacityTerm r;
        if (estimate) {  // consider_transport_in_production_costs, CapacityManager.cpp:84-86
            const Flow tf = fcopy(mkflow(st.T_q[s], st.T_v[s]));
            r.ucc = pmul(flow_price(fcopy(fadd(u, tf))), st.a[s]);
        } else {
            r.ucc = pmul(flow_price(u), st.a[s]);
        }
        r.pmin = ratio(u.q, st.Istar_q[s]);
        return r;
    });
    double p_hat = lambda * beta;
    if (t.pmin < p_hat) p_hat = t.pmin;
    Flow result = fcopy(fmul(Xstar, p_hat));
    if (result.q > 0) result = set_price(result, t.ucc + var_costs);
    return result;
}

// per-seller record read by every buyer of that seller (SupplyParameters, SalesManager.h:38-43, plus
// the seller constants the price curve needs)
struct SellerPub {
    i64 Xhat_q;       // possible_production_X_hat quantity
    double Xhatf;     // ... as float
    double Xf;        // to_float(production_X quantity)
    double Xexpf;     // to_float(expected_production_X quantity)
    double nbarf;     // to_float(offer_price_n_bar)
    double lamXf;     // forced_initial_production_quantity_lambda_X_star_float (Firm.h:73)
    double c1, c2;    // est. price increase / (2 lamX) and / lamX: the factors of PurchasingManager.cpp:155-156,179-180
    double n_co;      // estimate_marginal_production_costs at the seller's production with n_c un-rounded (:323-326)
};

// The same record once more, field-major over the firms in SECTOR-MAJOR order ("seller rows": the firms of one sector are
// consecutive).  A purchasing manager's connections all lead to sellers of one sector, in ascending agent order, so a warp
// that gathers the records of consecutive connections reads consecutive rows: every field is one coalesced 8-byte read per
// lane instead of 32 strided 72-byte rows.  Two values every buyer would derive from the record alone are stored with it
// (expected_production / production and the extension penalty at the expected production, PurchasingManager.cpp:334-349,
// :279-289 -- both for an unforced route, minimum passage 1; a buyer behind a throttled route recomputes them).
enum { SPT_XHAT_Q = 0, SPT_XHATF, SPT_XF, SPT_XEXPF, SPT_NBARF, SPT_LAMXF, SPT_C1, SPT_C2, SPT_NCO, SPT_RR, SPT_NPEE, SPT_FIELDS };
struct SellerTable {
    double* f;        // [SPT_FIELDS][stride]; field SPT_XHAT_Q holds the int64's bits
    const int* row;   // agent -> seller row (-1: not a firm)
    i64 stride;
};
ACC_HD double spt_ratio(double Xe, double X) { return (X > 0.0) ? Xe / X : 0.0; }
ACC_HD double spt_extension_penalty(double c1, double Xe, double lamXf) {
    return (Xe > lamXf && Xe > 0.0) ? stdmax(0.0, c1 * (Xe - lamXf) * (Xe - lamXf)) / Xe : 0.0;
}
ACC_HD void publish_seller(const SellerTable& t, int agent, const SellerPub& pb) {
    if (t.f == nullptr) return;
    const int r = t.row[agent];
    if (r < 0) return;
    double* f = t.f + r;
    i64 bits = pb.Xhat_q;
    double xq;
    memcpy(&xq, &bits, sizeof xq);
    f[SPT_XHAT_Q * t.stride] = xq; f[SPT_XHATF * t.stride] = pb.Xhatf; f[SPT_XF * t.stride] = pb.Xf; f[SPT_XEXPF * t.stride] = pb.Xexpf;
    f[SPT_NBARF * t.stride] = pb.nbarf; f[SPT_LAMXF * t.stride] = pb.lamXf; f[SPT_C1 * t.stride] = pb.c1; f[SPT_C2 * t.stride] = pb.c2;
    f[SPT_NCO * t.stride] = pb.n_co;
    f[SPT_RR * t.stride] = spt_ratio(pb.Xexpf, pb.Xf);
    f[SPT_NPEE * t.stride] = spt_extension_penalty(pb.c1, pb.Xexpf, pb.lamXf);
}

// ------------------------------------------------------------------------------------------------
// Per-firm state
// ------------------------------------------------------------------------------------------------
struct PeerTable;
struct FirmView {
    SellerPub* pub;                 // published after the expectation phase, read by the purchasing phase
    // peer stores (null: none): the other parts' copies of both seller tables; an owned firm's record goes straight into them
    const PeerTable* peers; int n_parts, my_part;
    SellerTable pubT;               // ... and its field-major twin
    const i64* Xstar_q; const double* Xstar_v; const double* beta; const double* lambda; const int* sector; const int* region; const int* type;
    const i64* markup_eff; const i64* var_costs;  // SalesManager::get_initial_markup / get_initial_unit_variable_production_costs
    const int* st_off; const int* st_list;        // the agent's input storages, reference container order
    i64* cap_q; double* cap_v;      // CapacityManager::possible_production_X_hat_
    i64* Xhat_q; double* Xhat_v;    // communicated possible_production_X_hat (price = n_c)
    i64* est_q; double* est_v;      // estimated_possible_production_X_hat_
    i64* X_q; double* X_v;          // production_X
    i64* Xexp_q; double* Xexp_v;    // expected_production_X
    i64* nbar;                      // offer_price_n_bar
    i64* cutoff;                    // price_cheapest_buyer_accepted_in_optimization
    double* C; double* R;           // total_production_costs_C, total_revenue_R
    i64* desired_q;                 // desired_production_X_tilde (quantity)
    i64* sumD_q; double* sumD_v;    // sum_demand_requests_D at the start of the step
    i64* bk_sq; double* bk_sv; i64* bk_cnt; i64* bk_pmax; i64* bk_pmin;  // demand book totals, reused by the expectation phase
    i64* exp_q; double* exp_v;      // non-domestic shipments of this step (region export)
    int* sorted;                    // 1 where this step's pro-rata distribution took its sums in the reference's sorted order (diagnostic)
};

// Peer memory of a region-partitioned run (acc_comm_enable_peer_stores): the other parts' halo receive buffers, mapped into this
// device's address space over NVLink (CUDA IPC).  A kernel that produces a shipment or a demand request for a connection whose
// other end lives on another part stores the 16-byte entry straight into that part's receive buffer, at the entry's position
// in the halo order, next to its local store: no pack kernel and no collective -- the transfer rides under the producing
// kernel; the receiving part applies the buffer with its unpack kernel once a cross-part barrier says everything has landed.
#define ACC_MAX_PARTS 16
struct PeerTable {
    i64* flow_recv[ACC_MAX_PARTS];  // [n_recv_p][2] shipments (quantity, value bits) part p receives
    i64* dem_recv[ACC_MAX_PARTS];   // [n_send_p][2] demand requests (quantity, price) part p receives
    unsigned long long* flags[ACC_MAX_PARTS];  // [ACC_MAX_PARTS] arrival counters of part p, one per announcing part (cross-part barrier)
    SellerPub* pub[ACC_MAX_PARTS];  // part p's seller records (FirmView::pub) ...
    double* pubT[ACC_MAX_PARTS];    // ... and their field-major twin (SellerTable::f; rows and stride are the same on every part)
};
// an owned firm's new record, stored into every other part's tables (the purchasing managers there read it after the next barrier)
ACC_HD void peer_publish_seller(const FirmView& fv, int f, int p) {
    if (p == fv.my_part) return;
    const SellerPub pb = fv.pub[f];
    fv.peers->pub[p][f] = pb;
    SellerTable t = fv.pubT;
    t.f = fv.peers->pubT[p];
    publish_seller(t, f, pb);
}
struct alignas(16) HaloEntry {
    i64 a, b;
};

// The conveyor: every connection owns a ring of tau + 1 slots (the delay queue of its transport chain plus the slot of the
// shipment under way to it).  Connections of equal delay form a RING GROUP stored plane by plane -- [tau + 1][connections of
// the group] --, so that neighbouring connections of a seller (neighbouring buyers: same region, same delay) still sit side by
// side in every plane.  A connection's word `ring << 24 | rank` names its group and its position in the group's planes (a
// group holds at most 2^24 connections; larger ones are split).  Memory is 16 * sum(tau + 1) bytes instead of 16 * T * E for
// dense planes of the longest delay.  The table travels by value with the kernel arguments.
#ifndef ACC_RING_RANK_BITS
#define ACC_RING_RANK_BITS 24  // (a test build lowers it to run the split of large groups on a small network)
#endif
enum { ACC_MAX_RINGS = 64 };
ACC_HD unsigned ring_of(unsigned w) { return w >> ACC_RING_RANK_BITS; }
ACC_HD i64 rank_of(unsigned w) { return (i64)(w & ((1u << ACC_RING_RANK_BITS) - 1u)); }
struct ConveyorRings {
    // for the step the view was made for: the plane of what arrives now (time step) and of this step's shipment (time step + tau)
    i64 front[ACC_MAX_RINGS];
    i64 last[ACC_MAX_RINGS];
    i64 base[ACC_MAX_RINGS];             // first slot of the group
    int count[ACC_MAX_RINGS];            // connections in the group
    unsigned char len[ACC_MAX_RINGS];    // tau + 1
    unsigned char phase[ACC_MAX_RINGS];  // step mod (tau + 1)
};
ACC_HD int ring_tau(const ConveyorRings& rg, unsigned w) { return (int)rg.len[ring_of(w)] - 1; }
// slot of time step + k, 0 <= k <= tau (k = 0: what arrives now; k = tau: the shipment of this step)
ACC_HD i64 ring_slot(const ConveyorRings& rg, unsigned w, int k) {
    const unsigned r = ring_of(w);
    const int L = rg.len[r];
    int i = (int)rg.phase[r] + k;
    if (i >= L) i -= L;
    return rg.base[r] + (i64)i * rg.count[r] + rank_of(w);
}
ACC_HD i64 ring_slot_front(const ConveyorRings& rg, unsigned w) { return rg.front[ring_of(w)] + rank_of(w); }  // k = 0
ACC_HD i64 ring_slot_last(const ConveyorRings& rg, unsigned w) { return rg.last[ring_of(w)] + rank_of(w); }    // k = tau
// slot of an absolute time u (any sign): for reads outside the step's kernels
ACC_HD i64 ring_slot_time(const ConveyorRings& rg, unsigned w, i64 u) {
    const unsigned r = ring_of(w);
    const i64 L = rg.len[r];
    i64 i = u % L;
    if (i < 0) i += L;
    return rg.base[r] + i * rg.count[r] + rank_of(w);
}

// Connection arrays in CSR-by-seller order
struct EdgeView {
    const int* seller_off;          // [n_agents + 1]
    i64* D_q; i64* D_p;             // last_demand_request_D (quantity, price)
    const i64* Z_q;                 // initial_flow_Z_star
    const unsigned char* tau;       // transport delay: sum over the hops of the connection's route
    const unsigned char* domestic;  // buyer in the seller's region
    const int* route;               // route of the connection (-1: immediate / domestic, no geo entity)
    const int* route_off; const int* route_ent;  // route -> its geo entities in travel order (TransportChainLink chain)
    const unsigned char* ent_delay; // delay of each geo entity
    i64* conv_q; double* conv_v;    // conveyor slots (ConveyorRings): slot (t + tau) mod (tau + 1) of a connection's ring receives the shipment of step t
    const unsigned* slotw;          // ring << 24 | rank of every connection
    ConveyorRings rings;
    i64 n_edges; int T;             // T = longest delay + 1
    i64* ovf_q; double* ovf_v;      // overflow per (hop, connection): [hop][n_edges] (null until passage forcing is used)
    const double* nu;               // passage per geo entity (null until used; < 0: none)
    i64* ship_q; double* ship_v;    // last shipment (only kept apart when hops may rewrite its conveyor slot, else null)
    int* sort_scratch;              // [2 n_edges]: a seller's connections in the reference's sorted order, when a pass needs it (ordered_range_sums)
    // peer stores (null: none): part that holds the other end of each connection (-1: this one) and the entry's position in that
    // part's receive buffer
    const signed char* x_peer; const int* x_rid; const PeerTable* peers;
};
// the shipment on connection e also goes into the receive buffer of the part that owns the buyer (one 16-byte store)
ACC_HD void peer_store_shipment(const EdgeView& ev, i64 e, Flow z) {
    if (ev.x_peer == nullptr) return;
    const int p = ev.x_peer[e];
    if (p < 0) return;
    HaloEntry h;
    h.a = z.q; memcpy(&h.b, &z.v, sizeof h.b);
    reinterpret_cast<HaloEntry*>(ev.peers->flow_recv[p])[ev.x_rid[e]] = h;
}
ACC_HD void peer_store_demand(const EdgeView& ev, i64 e, i64 q, i64 pr) {
    if (ev.x_peer == nullptr) return;
    const int p = ev.x_peer[e];
    if (p < 0) return;
    HaloEntry h;
    h.a = q; h.b = pr;
    reinterpret_cast<HaloEntry*>(ev.peers->dem_recv[p])[ev.x_rid[e]] = h;
}

// ------------------------------------------------------------------------------------------------
// Phase 1, one seller: production X, offer price, cutoff price (calc_supply_distribution_scenario,
// SalesManager.cpp:253-423) and distribute (:140-251); shipments go straight onto the conveyor.
// ------------------------------------------------------------------------------------------------
struct DistTotals {
    i64 fq; double fv;   // served completely (price above the cheapest price range)
    i64 rq; double rv;   // in the cheapest price range
    i64 rcnt;
    static ACC_HD DistTotals identity() { DistTotals r; r.fq = 0; r.fv = 0.0; r.rq = 0; r.rv = 0.0; r.rcnt = 0; return r; }
    static ACC_HD DistTotals combine(DistTotals a, DistTotals b) {
        DistTotals r; r.fq = a.fq + b.fq; r.fv = a.fv + b.fv; r.rq = a.rq + b.rq; r.rv = a.rv + b.rv; r.rcnt = a.rcnt + b.rcnt; return r;
    }
};
struct ShipTotals {
    double rev; i64 eq; double ev;
    i64 doubt;  // shipments whose pro-rata quantity lies so close to a rounding boundary that the order of the sums behind it may decide
    static ACC_HD ShipTotals identity() { ShipTotals r; r.rev = 0.0; r.eq = 0; r.ev = 0.0; r.doubt = 0; return r; }
    static ACC_HD ShipTotals combine(ShipTotals a, ShipTotals b) {
        ShipTotals r; r.rev = a.rev + b.rev; r.eq = a.eq + b.eq; r.ev = a.ev + b.ev; r.doubt = a.doubt + b.doubt; return r;
    }
};
// Is f / precision (as f * inv) within `slack` (in quanta) of a half-integer, i.e. can a perturbation of that size change round()?
ACC_HD bool near_rounding_boundary(double f, double inv, double slack) {
    const double y = fabs(f * inv);
    const double fr = y - floor(y);
    return fabs(fr - 0.5) <= slack + 1e-9;
}

// cheapest_price_range_half_width, SalesManager.cpp:151-159 (each typed step re-quantised)
ACC_HD i64 half_width(const ModelP& mp, i64 dn, double beta) {
    if (mp.generic_width) return pdiv(pmul(pmul(pdiv(dn, 2.0), beta - 1), beta - 1), beta);
    return pdiv(mp.width, 2.0);
}

// Demand of the cheapest price range as the reference accumulates it (SalesManager.cpp:163-185): two DOUBLE sums -- quantity and
// value -- over the connections in the order calc_production_X has sorted them into (price descending, then quantity descending,
// :104-127; connections with equal price and quantity contribute equal summands, so their mutual order does not matter).  The
// engine does not sort (every decision of the scan is taken on int64 totals) and sums values in its fixed tree order,
// which agrees with the reference to ~1e-16 -- not enough where the pro-rata formula (:216-224) lands on an exact half quantum:
// at the onset of a shock from the baseline all prices are equal, a firm at half capacity ships q / 2 to every buyer, and for
// every odd q the last bits of the two sums decide which way round() goes (found by the three-year flood run: one shipment in
// 27 000 off by a quantum at step 10).  seller_phase1 therefore bounds how far its sums can lie from the reference's ((n + 16)
// unit roundoffs, both being sums of positive terms), carries the bound through the three roundings that depend on them (average
// price, price shift, every shipment) and, if any of them comes within the bound of a rounding boundary, takes the sums in the
// reference's order instead: the range's connections are sorted (a bitonic network over 16-bit indices in the CTA's shared memory
// -- over the seller's slice of a per-connection scratch array where the launch had no room for them --, connections outside the
// range as sentinels at the end; comparisons read the staged book), then added up left to right.  O(n log^2 n) for a seller in doubt -- every shocked seller at a shock's onset, about one in a thousand otherwise.
ACC_HD int ordered_range_slots(int n) { int N = 1; while (N < n) N <<= 1; return N; }  // < 2 n: the scratch holds two ints per connection
// where the permutation lives: the CTA's shared memory when the launch provided room for it (16-bit entries), else global scratch
struct PermGlobal {
    int* p;
    ACC_HD int get(int i) const { return p[i]; }
    ACC_HD void set(int i, int v) const { p[i] = v; }
};
struct PermShared {
    unsigned short* p;
    ACC_HD int get(int i) const { const unsigned short v = p[i]; return v == 0xffffu ? -1 : (int)v; }
    ACC_HD void set(int i, int v) const { p[i] = v < 0 ? (unsigned short)0xffffu : (unsigned short)v; }
};
template <class Ctx, class Perm>
ACC_HD void ordered_range_sums(Ctx& cx, int n, const i64* Dq, const i64* Dp, i64 cutoff, i64 h, int m, Perm perm, double (&rs)[2]) {
    const int N = ordered_range_slots(n);
    cx.foreach(N, [&](int i) {
        int v = -1;
        if (i < n) {
            const i64 d = Dp[i] - cutoff;
            if (Dq[i] > 0 && (d < 0 ? -d : d) < h) v = i;
        }
        perm.set(i, v);
    });
    // strict total order: sentinels last, equal keys by connection (any order of equal keys gives the same sums)
    // (no early exits: the four loads of a pair are unconditional, so those of the eight pairs of a turn can be in flight together)
    auto before = [&](int a, int b) {
        const int ia = a < 0 ? 0 : a, ib = b < 0 ? 0 : b;
        const i64 pa = Dp[ia], pb = Dp[ib], qa = Dq[ia], qb = Dq[ib];
        const bool first = pa > pb || (pa == pb && (qa > qb || (qa == qb && a < b)));
        return a >= 0 && (b < 0 || first);
    };
    for (int k = 2; k <= N; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            cx.sync();
            // eight pairs per turn, all of their loads first: a thread's stores to perm would otherwise order its next loads behind them
            enum { B = 8 };
            cx.foreach((N / 2 + B - 1) / B, [&](int g) {
                int a[B], b[B];
#pragma unroll
                for (int u = 0; u < B; ++u) {
                    const int t = g * B + u;
                    const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // the pair's lower index: bit j clear
                    a[u] = t < N / 2 ? perm.get(i) : -1;
                    b[u] = t < N / 2 ? perm.get(i | j) : -1;
                }
#pragma unroll
                for (int u = 0; u < B; ++u) {
                    const int t = g * B + u;
                    const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                    const bool ascending = (i & k) == 0;
                    if (t < N / 2 && (ascending ? before(b[u], a[u]) : before(a[u], b[u]))) { perm.set(i, b[u]); perm.set(i | j, a[u]); }
                }
            });
        }
    }
    cx.sync();
    cx.template seqsum<2>(m, [&](int r, int k) {
        const int i = perm.get(r);
        return k == 0 ? qf(Dq[i]) : q_times_p(Dq[i], Dp[i]);
    }, rs);
}

template <class Ctx>
ACC_HD void seller_phase1(Ctx& cx, int f, const FirmView& fv, const EdgeView& ev, const ModelP& mp, const SectorP* sectors, const EventSink& events, i64 step,
                          const SellerBook& book) {
    const int e0 = fv.type[f] == 1 ? ev.seller_off[f] : 0;
    const int n = fv.type[f] == 1 ? ev.seller_off[f + 1] - e0 : 0;
    if (n == 0) return;
    // the demand book: staged in shared memory by the caller (csrc/stage.cuh) or read in place
    const i64* Dq = book.q != nullptr ? book.q : ev.D_q + e0;
    const i64* Dp = book.q != nullptr ? book.p : ev.D_p + e0;
    const SectorP sp = sectors[fv.sector[f]];
    const double lambda = fv.lambda[f];
    const Flow Xstar = mkflow(fv.Xstar_q[f], fv.Xstar_v[f]);

    const BookTotals tot = book_totals(cx, n, Dq, Dp);
    // possible production: capacity manager's value, priced with last step's estimate when there is one
    // (SalesManager.cpp:129-133)
    Flow xhat = mkflow(fv.cap_q[f], fv.cap_v[f]);
    {
        const Flow est = mkflow(fv.est_q[f], fv.est_v[f]);
        if (est.q > 0) xhat = set_price(xhat, flow_price(est));
    }
    const i64 n_c = flow_price(xhat);                                 // minimal_production_price, :263
    const i64 offer_min = n_c + pdiv(fv.markup_eff[f], 1.0 - 0.0);    // minimal_offer_price, :264 (tax = 0)
    const i64 lamX = qmul(Xstar.q, lambda);                           // forced_initial_production_quantity_lambda_X_star
    const i64 desired = tot.sq;                                       // desired_production_X_tilde = sum of demand requests

    Flow X = mkflow(0, 0.0);
    i64 nbar = 0, cutoff = 0;
    double C = 0.0;
    if (xhat.q <= 0) {                                                // :266-278
        if (cx.leader() && !(lambda <= 0.0)) events.push(EV_NO_PRODUCTION_SUPPLY_SHORTAGE, f, -1, 0.0);
    } else if (tot.cnt == 0 || tot.pmax < n_c) {                      // :279-295
        if (cx.leader()) events.push(tot.cnt == 0 ? EV_NO_PRODUCTION_DEMAND_QUANTITY_SHORTAGE : EV_NO_PRODUCTION_DEMAND_VALUE_SHORTAGE, f, -1, 0.0);
    } else {
        const ScanResult sc = seller_scan(cx, n, Dq, Dp, tot, xhat.q, lamX, sp.dn, n_c);
        X = fcopy(sc.cum);                                            // :317
        cutoff = fv.cutoff[f];                                        // stale value survives some branches of the reference
        if (sc.all_served) {                                          // :320-350
            C = total_costs(X.q, lamX, sp.dn, n_c);
            if (desired < lamX && desired < xhat.q) {                 // :327-344
                const i64 cand = pmul(flow_price(X), 1 + sp.eps * ratio(X.q - lamX, lamX));
                nbar = cand > offer_min ? cand : offer_min;
            } else {
                nbar = flow_price(X);                                 // :346
            }
            cutoff = sc.p_prev;
        } else {                                                      // :351-422
            Flow extra = mkflow(0, 0.0);
            if (marginal_costs(X.q, lamX, sp.dn, n_c) - sc.p_stop < 0) {
                if (marginal_costs(xhat.q, lamX, sp.dn, n_c) - sc.p_stop < 0) {
                    extra = fcopy(flow_qp(xhat.q - X.q, sc.p_stop));  // :362-367
                    cutoff = sc.p_stop;
                } else {
                    const i64 total = analytic_extension(lamX, sp.dn, n_c, sc.p_stop);  // :370-371
                    if (total > X.q) {
                        extra = fcopy(flow_qp(total - X.q, sc.p_stop));
                        cutoff = sc.p_stop;
                    } else if (sc.has_prev) {
                        cutoff = sc.p_prev;                           // :388-391
                    }
                }
                X = fcopy(fadd(X, extra));                            // :395-396
            } else if (sc.has_prev) {
                cutoff = sc.p_prev;                                   // :400-405
            }
            C = total_costs(X.q, lamX, sp.dn, n_c);                   // :407
            if (X.q > 0) {
                nbar = flow_price(X);
            } else {                                                  // :413-422
                if (cx.leader()) events.push(EV_NO_PRODUCTION_HIGH_COSTS, f, -1, 0.0);
                cutoff = 0;
                nbar = n_c;
            }
        }
    }

    // ---- distribute, :140-251 ----
    const unsigned* slotw = ev.slotw + e0;
    const unsigned char* dom = book.q != nullptr ? book.dom : ev.domestic + e0;
    double R = 0.0;
    i64 exq = 0; double exv = 0.0;
    bool sorted_sums = false;
    if (X.q <= 0) {
        cx.foreach(n, [&](int i) {
            const i64 at = ring_slot_last(ev.rings, slotw[i]);
            ev.conv_q[at] = 0;
            if (ev.ship_q != nullptr) { ev.ship_q[e0 + i] = 0; ev.ship_v[e0 + i] = 0.0; }
            ev.conv_v[at] = 0.0;
            peer_store_shipment(ev, (i64)e0 + i, mkflow(0, 0.0));
        });
    } else {
        const i64 h = half_width(mp, sp.dn, fv.beta[f]);
        const DistTotals dt_ = cx.template reduce<DistTotals>(n, [&](int i) {
            DistTotals r = DistTotals::identity();
            const i64 q = Dq[i];
            if (q > 0) {
                const i64 d = Dp[i] - cutoff;
                if (d >= h) { r.fq = q; r.fv = q_times_p(q, Dp[i]); }
                else if ((d < 0 ? -d : d) < h) { r.rq = q; r.rv = q_times_p(q, Dp[i]); r.rcnt = 1; }
            }
            return r;
        });
        const i64 P = X.q - dt_.fq;                                   // production_cheapest_price_range, :197-198
        const bool prorata = dt_.rcnt > 0 && dt_.rq > P;              // :199
        // demand_cheapest_price_range / demand_value_cheapest_price_range (:176-177): the engine's sums first; `exact` once they
        // have been taken in the reference's order (see ordered_range_sums)
        double rq_f = qf(dt_.rq), rv_f = dt_.rv;
        bool exact = ev.sort_scratch == nullptr;
        const double kappa = (double)(n + 16) * 0x1p-53;              // relative distance of either sum from the reference's
        ShipTotals st_ = ShipTotals::identity();
        for (;;) {
            i64 shift = 0, seller_price = 0;
            double den = 1.0, ship_slack = 0.0;
            bool doubt = false;
            if (prorata) {
                const i64 pmax_r = cutoff + h, pmin_r = cutoff - h;
                const double ratio = rv_f / rq_f;
                const i64 avg = mkp(ratio);                           // :206
                const double num = q_times_p(P, pmax_r) - rq_f * pf(avg), den2 = rq_f - qf(P);
                const double c = num / den2;
                const i64 cand = mkp(c);                              // :207-209
                shift = cand > (0 - pmin_r) ? cand : (0 - pmin_r);
                seller_price = mkp((X.v - dt_.fv) / qf(P));           // :211-212
                den = rq_f * pf(shift) + rv_f;                        // :220
                if (!exact) {
                    const double dc = 2.0 * kappa * (rq_f * pf(avg) + fabs(c) * rq_f) / fabs(den2);
                    doubt = near_rounding_boundary(ratio, 1e6, 4.0 * kappa * ratio * 1e6) || near_rounding_boundary(c, 1e6, dc * 1e6);
                    ship_slack = 2.0 * kappa * (rq_f * fabs(pf(shift)) + rv_f) / den;  // relative, of every pro-rata quantity
                }
            }
            if (!doubt) {
                st_ = cx.template reduce<ShipTotals>(n, [&](int i) {
                    ShipTotals r = ShipTotals::identity();
                    const i64 q = Dq[i];
                    Flow z = mkflow(0, 0.0);
                    if (q > 0) {
                        const i64 p = Dp[i];
                        const i64 d = p - cutoff;
                        if (d >= h) {
                            z = flow_qp(q, p);                        // :170
                        } else if ((d < 0 ? -d : d) < h) {
                            if (prorata) {                            // :216-224
                                const double x = qf(P) * (q_times_p(q, shift) + q_times_p(q, p)) / den;
                                if (!exact && near_rounding_boundary(x, 1e3, ship_slack * x * 1e3)) r.doubt = 1;
                                const i64 zq = mkq(x);
                                z = fcopy(flow_qp(zq, mp.preserve_seller_price ? seller_price : p));
                                r.rev = z.v;
                            } else {
                                z = flow_qp(q, p);                    // :234
                            }
                        }
                    }
                    const i64 at = ring_slot_last(ev.rings, slotw[i]);
                    ev.conv_q[at] = z.q;                              // BusinessConnection::push_flow_Z -> first transport link
                    if (ev.ship_q != nullptr) { ev.ship_q[e0 + i] = z.q; ev.ship_v[e0 + i] = z.v; }  // last_shipment_Z, kept apart when hops rewrite the slot
                    ev.conv_v[at] = z.v;
                    peer_store_shipment(ev, (i64)e0 + i, z);
                    if (!dom[i]) { r.eq = z.q; r.ev = z.v; }          // Region::add_export_Z, BusinessConnection.cpp:91-93
                    return r;
                });
                doubt = st_.doubt != 0;
            }
            if (!doubt || exact) break;
            double rs[2];
            if (book.perm16 != nullptr && ordered_range_slots(n) <= book.perm_slots)
                ordered_range_sums(cx, n, Dq, Dp, cutoff, h, (int)dt_.rcnt, PermShared{book.perm16}, rs);
            else
                ordered_range_sums(cx, n, Dq, Dp, cutoff, h, (int)dt_.rcnt, PermGlobal{ev.sort_scratch + 2 * (i64)e0}, rs);
            rq_f = rs[0]; rv_f = rs[1];
            exact = true;                                             // ... and the shipments are written once more
            sorted_sums = true;
        }
        if (dt_.rcnt > 0 && prorata) R = dt_.fv + st_.rev;            // :214,225
        else R = X.v;                                                 // :231,241
        exq = st_.eq; exv = st_.ev;
    }
    if (cx.leader()) {
        fv.Xhat_q[f] = xhat.q; fv.Xhat_v[f] = xhat.v;
        fv.X_q[f] = X.q; fv.X_v[f] = X.v;
        fv.nbar[f] = nbar; fv.cutoff[f] = cutoff;
        fv.C[f] = C; fv.R[f] = X.q <= 0 ? fv.R[f] : R;
        fv.desired_q[f] = desired;
        fv.sumD_q[f] = tot.sq; fv.sumD_v[f] = tot.sv;
        fv.bk_sq[f] = tot.sq; fv.bk_sv[f] = tot.sv; fv.bk_cnt[f] = tot.cnt; fv.bk_pmax[f] = tot.pmax; fv.bk_pmin[f] = tot.pmin;
        fv.exp_q[f] = exq; fv.exp_v[f] = exv;
        if (fv.sorted != nullptr) fv.sorted[f] = sorted_sums ? 1 : 0;
    }
}

// ------------------------------------------------------------------------------------------------
// Phase 2, one seller: expected production and offer price
// (SalesManager::iterate_expectation :614-624 and calc_expected_supply_distribution_scenario :425-612
// with naive_expectations = true, respect_markup_in_production_extension = false, tax = 0).
// Needs the demand book only when the scan stops early; its totals were stored by phase 1.
// ------------------------------------------------------------------------------------------------
template <class Ctx>
ACC_HD void seller_phase2(Ctx& cx, int f, const FirmView& fv, const EdgeView& ev, const StorageView& st, const ModelP& mp, const SectorP* sectors,
                          const EventSink& events) {
    if (fv.type[f] != 1) return;
    const int e0 = ev.seller_off[f];
    const int n = ev.seller_off[f + 1] - e0;
    if (n == 0) return;
    const SectorP sp = sectors[fv.sector[f]];
    const double lambda = fv.lambda[f];
    const Flow Xstar = mkflow(fv.Xstar_q[f], fv.Xstar_v[f]);
    const int s0 = fv.st_off[f];
    Flow est = possible_production(cx, st, fv.st_list + s0, fv.st_off[f + 1] - s0, Xstar, lambda, fv.beta[f], fv.var_costs[f], mp.dt, true);
    if (est.q > 0) est = set_price(est, pdiv(flow_price(est), 1.0 - 0.0));  // :617-620
    BookTotals tot;
    tot.sq = fv.bk_sq[f]; tot.sv = fv.bk_sv[f]; tot.cnt = fv.bk_cnt[f]; tot.pmax = fv.bk_pmax[f]; tot.pmin = fv.bk_pmin[f];
    const i64 n_c = flow_price(est);
    const i64 offer_min = n_c + pdiv(fv.markup_eff[f], 1.0 - 0.0);
    const i64 lamX = qmul(Xstar.q, lambda);
    Flow X = mkflow(0, 0.0);
    i64 nbar = 0;
    if (est.q <= 0) {                                                 // :436-448
        if (cx.leader() && !(lambda <= 0.0)) events.push(EV_NO_EXP_PRODUCTION_SUPPLY_SHORTAGE, f, -1, 0.0);
    } else if (tot.cnt == 0 || tot.pmax < n_c) {                      // :449-467
        if (cx.leader()) events.push(tot.cnt == 0 ? EV_NO_EXP_PRODUCTION_DEMAND_QUANTITY_SHORTAGE : EV_NO_EXP_PRODUCTION_DEMAND_VALUE_SHORTAGE, f, -1, 0.0);
        nbar = n_c;
        X = flow_qp(lamX < est.q ? lamX : est.q, nbar);
    } else {
        const ScanResult sc = seller_scan(cx, n, ev.D_q + e0, ev.D_p + e0, tot, est.q, lamX, sp.dn, n_c);
        X = fcopy(sc.cum);                                            // :498
        if (sc.all_served && !mp.naive_expectations && (mp.always_extend || X.q < lamX)) {
            // Expectations: extend the demand curve beyond the sum of the demand requests with the sector's supply
            // elasticity and produce up to where marginal costs meet marginal revenue (:505-545, :704-779)
            const i64 nmin = sc.p_prev;                               // price of the cheapest served connection
            const i64 sumD = tot.sq;                                  // sum_demand_requests_D_
            const double eps = sp.eps;
            auto goal = [&](i64 xq) {                                 // goal_fkt_marginal_costs_minus_marginal_revenue, :737-744
                return marginal_costs(xq, lamX, sp.dn, offer_min) - pmul(nmin, pow(ratio(sumD, xq), eps));
            };
            auto add_rev = [&](i64 xq) {                              // calc_additional_revenue_expectation, :704-716
                const i64 t = qmul(xq, pow(ratio(sumD, xq), eps)) - sumD;
                return (pf(nmin) * qf(t)) / (1.0 - eps);
            };
            if (goal(X.q) < 0) {
                const i64 lam_beta_X = qmul(Xstar.q, fv.beta[f] * lambda);  // forced_maximal_production_quantity_lambda_beta_X_star
                const i64 max_add = (est.q < lam_beta_X ? est.q : lam_beta_X) - X.q;
                Flow add;
                if (goal(X.q + max_add) < 0) {
                    add = mkflow(max_add, add_rev(X.q + max_add));
                } else {                                              // search_root_bisec_expectation, :754-779
                    i64 left = 0, right = max_add;
                    const i64 prec = 1;                               // Price(Price::precision): one quantum
                    for (;;) {
                        const i64 middle = qdiv(left + right, 2.0);
                        auto mag = [&](i64 d) { const i64 g = goal(X.q + d); return g < 0 ? -g : g; };
                        if (left + 1 >= right) { const i64 d = mag(left) < mag(right) ? left : right; add = mkflow(d, add_rev(X.q + d)); break; }
                        if (mag(middle) < prec) { add = mkflow(middle, add_rev(X.q + middle)); break; }
                        if (mag(left) < prec) { add = mkflow(left, add_rev(X.q + left)); break; }
                        if (mag(right) < prec) { add = mkflow(right, add_rev(X.q + right)); break; }
                        const bool sm = goal(X.q + middle) >= 0, sr = goal(X.q + right) >= 0;  // same_sgn: (t >= 0) == (t >= 0)
                        if (sm != sr) left = middle; else right = middle;
                    }
                }
                X = fcopy(fadd(X, fcopy(add)));
            }
        }
        if (!sc.all_served) {                                         // :554-599
            Flow extra = mkflow(0, 0.0);
            if (marginal_costs(X.q, lamX, sp.dn, n_c) - sc.p_stop < 0) {
                if (marginal_costs(est.q, lamX, sp.dn, n_c) - sc.p_stop < 0) {
                    extra = fcopy(flow_qp(est.q - X.q, sc.p_stop));
                } else {
                    const i64 total = analytic_extension(lamX, sp.dn, n_c, sc.p_stop);
                    if (total > X.q) extra = fcopy(flow_qp(total - X.q, sc.p_stop));
                }
                X = fcopy(fadd(X, extra));
            }
        }
        if (X.q > 0) {                                                // :548-549, :600-605
            const i64 a = flow_price(X);
            const i64 b = mkp(total_costs(X.q, lamX, sp.dn, offer_min) / qf(X.q));
            nbar = a > b ? a : b;
        } else {                                                      // :607-611
            if (cx.leader()) events.push(EV_NO_EXP_PRODUCTION_HIGH_COSTS, f, -1, 0.0);
            nbar = offer_min;
        }
    }
    if (cx.leader()) {
        fv.est_q[f] = est.q; fv.est_v[f] = est.v;
        fv.Xexp_q[f] = X.q; fv.Xexp_v[f] = X.v;
        fv.nbar[f] = nbar;
        SellerPub pb;
        const Flow xh = mkflow(fv.Xhat_q[f], fv.Xhat_v[f]);
        pb.Xhat_q = xh.q; pb.Xhatf = qf(xh.q); pb.Xf = qf(fv.X_q[f]); pb.Xexpf = qf(X.q); pb.nbarf = pf(nbar);
        pb.lamXf = qf(lamX);
        const double dnf = pf(sp.dn_est);
        pb.c1 = dnf / (2 * pb.lamXf);
        pb.c2 = dnf / pb.lamXf;
        const double ncf = xh.v / qf(xh.q);  // get_price_float(); only read when Xhat_q > 0
        pb.n_co = (pb.Xf <= pb.lamXf) ? ncf : ncf + pb.c2 * (pb.Xf - pb.lamXf);
        fv.pub[f] = pb;
        publish_seller(fv.pubT, f, pb);
    }
    if (fv.peers != nullptr) {
        cx.sync();  // the record is in this part's table: one lane per other part copies it over
        cx.foreach(fv.n_parts, [&](int p) { peer_publish_seller(fv, f, p); });
    }
    // Firm::iterate_expectation, Firm.cpp:65-73: sum_demand_requests_D was zeroed two lines earlier
    // (SalesManager.cpp:623), so the desired production is the expected production
    const i64 desired = X.q > 0 ? X.q : 0;
    const int ns = fv.st_off[f + 1] - s0;
    cx.foreach(ns, [&](int k) {
        const int s = fv.st_list[s0 + k];
        const Flow ut = flow_q(qmul(desired, st.a[s]));
        st.Ut_q[s] = ut.q; st.Ut_v[s] = ut.v;
    });
}

// ------------------------------------------------------------------------------------------------
// Storage update (Storage::calc_content_S + iterate_consumption_and_production, Storage.cpp:38-91)
// ------------------------------------------------------------------------------------------------
ACC_HD void storage_update(const StorageView& st, int s, Flow used, const ModelP& mp, double omega, const EventSink& events) {
    st.U_q[s] = fcopy(used).q; st.U_v[s] = fcopy(used).v;             // use_content_S
    const Flow former = mkflow(st.S_q[s], st.S_v[s]);
    const Flow in = mkflow(st.Ip_q[s], st.Ip_v[s]);
    Flow S = fcopy(fadd(former, flow_times_time(fsub(in, fcopy(used)), mp.dt)));
    const i64 floor_q = qmul_l(mp.min_storage, st.Sstar_q[s]);
    if (S.q <= floor_q) {
        events.push(EV_STORAGE_UNDERRUN, st.sector[s], st.agent[s], 0.0);
        S = fcopy(flow_qp(floor_q, flow_price(former)));
    }
    const Flow max_stock = fmul(fmul(mkflow(st.Sstar_q[s], st.Sstar_v[s]), 1.0), omega);
    if (max_stock.q < S.q) {
        events.push(EV_STORAGE_OVERRUN, st.sector[s], st.agent[s], qf(S.q - max_stock.q));
        const i64 tmp = flow_price(S);
        S = set_price(fcopy(max_stock), tmp);
    }
    st.S_q[s] = S.q; st.S_v[s] = S.v;
    st.Il_q[s] = fcopy(in).q; st.Il_v[s] = fcopy(in).v;              // input_flow_I[2] = input_flow_I[other]
    st.Ip_q[s] = 0; st.Ip_v[s] = 0.0;                                 // input_flow_I[other] = Flow(0.0)
}

}  // namespace acc
