// observe.cuh -- the output surface on the device: every observable of the reference's observe<Observer, H>() visitors,
// gathered for a SELECTION of agents into a buffer shaped like ArrayOutput::Variable::data of one timestep.
//
// Reference: include/model/{EconomicAgent,Firm,Consumer,Storage,BusinessConnection}.h (the observable lists),
// src/output/ArrayOutput.cpp:115-158 (to_float, Flow / Stock split into <name>_quantity / <name>_value) and :479-646
// (ArrayOutput::iterate: firms / consumers [agents], storages [sectors][agents], flows [agent_from][agent_to]; cells that
// nothing writes stay NaN, :170).  One thread per output cell (agents, storages) or one CTA per selected agent (flows);
// nothing but the selected cells is touched, and only the finished buffer crosses to the host.
#pragma once

#include "model_core.cuh"

namespace acc {

#define ACC_OBS_NAN ((double)NAN)

// ---- variable ids (engine.cu maps the reference's names onto them) ----
enum AgentVar {
    // EconomicAgent::observe, include/model/EconomicAgent.h:70-96
    AV_BUSINESS_CONNECTIONS = 0, AV_CONSUMPTION_Q, AV_CONSUMPTION_V, AV_DEMAND_Q, AV_DEMAND_V, AV_INPUT_FLOW_Q, AV_INPUT_FLOW_V, AV_STORAGE_Q, AV_STORAGE_V,
    // Firm::observe, include/model/Firm.h:78-153
    FV_COMMUNICATED_POSSIBLE_PRODUCTION_Q, FV_COMMUNICATED_POSSIBLE_PRODUCTION_V, FV_DESIRED_PRODUCTION_CAPACITY, FV_DIRECT_LOSS_Q, FV_DIRECT_LOSS_V,
    FV_EXPECTED_OFFER_PRICE, FV_EXPECTED_PRODUCTION_Q, FV_EXPECTED_PRODUCTION_V, FV_FORCING, FV_INCOMING_DEMAND_Q, FV_INCOMING_DEMAND_V, FV_OFFER_PRICE,
    FV_PRODUCTION_Q, FV_PRODUCTION_V, FV_PRODUCTION_CAPACITY, FV_POSSIBLE_PRODUCTION_CAPACITY, FV_TAX, FV_TOTAL_LOSS_Q, FV_TOTAL_LOSS_V,
    FV_TOTAL_PRODUCTION_COSTS, FV_TOTAL_REVENUE, FV_TOTAL_VALUE_LOSS, FV_UNIT_PRODUCTION_COSTS,
    // Consumer::observe, include/model/Consumer.h:98-110
    CV_UTILITY, CV_LOCAL_OPTIMAL_UTILITY
};
enum StorageVar {  // Storage::observe, include/model/Storage.h:90-150
    SV_BUSINESS_CONNECTIONS = 0, SV_CONTENT_Q, SV_CONTENT_V, SV_DEMAND_Q, SV_DEMAND_V, SV_DESIRED_USED_FLOW_Q, SV_DESIRED_USED_FLOW_V, SV_EXPECTED_COSTS,
    SV_INPUT_FLOW_Q, SV_INPUT_FLOW_V, SV_OPTIMIZED_VALUE, SV_POSSIBLE_USE_Q, SV_POSSIBLE_USE_V, SV_PURCHASE_Q, SV_PURCHASE_V, SV_SHIPMENT_Q, SV_SHIPMENT_V,
    SV_STORAGE_DEMAND_Q, SV_STORAGE_DEMAND_V, SV_TOTAL_TRANSPORT_PENALTY, SV_USE_Q, SV_USE_V, SV_USED_FLOW_Q, SV_USED_FLOW_V
};
enum FlowVar {  // BusinessConnection::observe, include/model/BusinessConnection.h:79-115
    XV_INITIAL_FLOW_Q = 0, XV_INITIAL_FLOW_V, XV_DEMAND_REQUEST_Q, XV_DEMAND_REQUEST_V, XV_FLOW_DEFICIT, XV_FLOW_MEAN_Q, XV_FLOW_MEAN_V, XV_RECEIVED_FLOW_Q,
    XV_RECEIVED_FLOW_V, XV_SENT_FLOW_Q, XV_SENT_FLOW_V, XV_TOTAL_FLOW_Q, XV_TOTAL_FLOW_V, XV_MINIMUM_PASSAGE
};

struct ObsFill {  // ArrayOutput::resize_data: everything starts as NaN
    double* out;
    ACC_HD void operator()(int i) const { out[i] = ACC_OBS_NAN; }
};

// last shipment of connection e: the slot of time t + delay in its conveyor ring, t = step of the last production pass
ACC_HD Flow obs_last_shipment(const EdgeView& ev, i64 t, i64 e) {
    const i64 at = ring_slot_time(ev.rings, ev.slotw[e], t + ev.tau[e]);
    return ev.ship_q != nullptr ? mkflow(ev.ship_q[e], ev.ship_v[e]) : mkflow(ev.conv_q[at], ev.conv_v[at]);
}

// firms / consumers: out[i] for the i-th selected agent (idx == nullptr: agent i); agents of the other kind stay NaN
struct ObserveAgents {
    int var, want_type;
    const i64* idx;
    FirmView fv; StorageView st; ModelP mp;
    const i64* in_off;                       // incoming connections per storage
    const i64* dem_q; const double* dem_v;   // PurchasingManager::demand_D
    const i64* inc_q; const double* inc_v;   // per-firm sum of the demand requests it holds (FirmIncomingDemand), or null
    const double* utility;                   // consumers' utility, or null (price-elastic consumers)
    double* out;
    ACC_HD void operator()(int i) const {
        const int a = idx != nullptr ? (int)idx[i] : i;
        if (fv.type[a] != want_type) { out[i] = ACC_OBS_NAN; return; }
        double r = ACC_OBS_NAN;
        if (var <= AV_STORAGE_V) {
            // std::accumulate over input_storages in container order; Flow + Flow adds quantities and values (types.h:552-554)
            Flow sum = mkflow(0, 0.0);
            i64 count = 0;
            for (int k = fv.st_off[a]; k < fv.st_off[a + 1]; ++k) {
                const int s = fv.st_list[k];
                switch (var) {
                    case AV_BUSINESS_CONNECTIONS: count += in_off[s + 1] - in_off[s]; break;
                    case AV_CONSUMPTION_Q: case AV_CONSUMPTION_V: sum = fadd(sum, mkflow(st.U_q[s], st.U_v[s])); break;
                    case AV_DEMAND_Q: case AV_DEMAND_V: sum = fadd(sum, mkflow(dem_q[s], dem_v[s])); break;
                    case AV_INPUT_FLOW_Q: case AV_INPUT_FLOW_V: sum = fadd(sum, mkflow(st.Il_q[s], st.Il_v[s])); break;
                    default: sum = fadd(sum, mkflow(st.S_q[s], st.S_v[s])); break;
                }
            }
            if (var == AV_BUSINESS_CONNECTIONS) r = (double)count;
            else r = (var == AV_CONSUMPTION_Q || var == AV_DEMAND_Q || var == AV_INPUT_FLOW_Q || var == AV_STORAGE_Q) ? qf(sum.q) : sum.v;
            out[i] = r;
            return;
        }
        const Flow X = mkflow(fv.X_q[a], fv.X_v[a]);
        switch (var) {
            case FV_COMMUNICATED_POSSIBLE_PRODUCTION_Q: r = qf(fv.Xhat_q[a]); break;
            case FV_COMMUNICATED_POSSIBLE_PRODUCTION_V: r = fv.Xhat_v[a]; break;
            case FV_DESIRED_PRODUCTION_CAPACITY: r = ratio(fv.desired_q[a], fv.Xstar_q[a]); break;         // CapacityManager.cpp:40
            case FV_DIRECT_LOSS_Q: case FV_DIRECT_LOSS_V: {                                               // Firm.h:89-93
                const i64 q = qmul(fv.Xstar_q[a], 1.0 - fv.lambda[a]);
                r = var == FV_DIRECT_LOSS_Q ? qf(q) : q_times_p(q, X.q > 0 ? flow_price(X) : 0);
                break;
            }
            case FV_EXPECTED_OFFER_PRICE: r = pf(fv.nbar[a]); break;
            case FV_EXPECTED_PRODUCTION_Q: r = qf(fv.Xexp_q[a]); break;
            case FV_EXPECTED_PRODUCTION_V: r = fv.Xexp_v[a]; break;
            case FV_FORCING: r = fv.lambda[a]; break;
            case FV_INCOMING_DEMAND_Q: r = qf(inc_q[a]); break;
            case FV_INCOMING_DEMAND_V: r = inc_v[a]; break;
            case FV_OFFER_PRICE: r = pf(flow_price(X)); break;
            case FV_PRODUCTION_Q: r = qf(X.q); break;
            case FV_PRODUCTION_V: r = X.v; break;
            case FV_PRODUCTION_CAPACITY: r = ratio(X.q, fv.Xstar_q[a]); break;                            // CapacityManager.cpp:38
            case FV_POSSIBLE_PRODUCTION_CAPACITY: r = ratio(fv.cap_q[a], fv.Xstar_q[a]); break;           // CapacityManager.cpp:42
            case FV_TAX: r = 0.0; break;                                                                  // Government is dead code at the pinned commit
            case FV_TOTAL_LOSS_Q: case FV_TOTAL_LOSS_V: {                                                 // Firm.h:133-137
                const i64 q = fv.Xstar_q[a] - X.q;
                r = var == FV_TOTAL_LOSS_Q ? qf(q) : q_times_p(q, X.q > 0 ? flow_price(X) : 0);
                break;
            }
            case FV_TOTAL_PRODUCTION_COSTS: r = fv.C[a]; break;
            case FV_TOTAL_REVENUE: r = fv.R[a]; break;
            case FV_TOTAL_VALUE_LOSS: r = fv.Xstar_v[a] - X.v; break;
            case FV_UNIT_PRODUCTION_COSTS: r = pf(flow_price(mkflow(fv.Xhat_q[a], fv.Xhat_v[a]))); break;
            case CV_UTILITY: case CV_LOCAL_OPTIMAL_UTILITY: r = utility != nullptr ? utility[a] : ACC_OBS_NAN; break;  // Consumer.cpp:304: the two are equal
            default: break;
        }
        out[i] = r;
    }
};

// storages: out[sector * n_sel + i] for the i-th selected agent; (sector, agent) pairs without a storage stay NaN
struct ObserveStorages {
    int var, n_sel, n_agents;
    const i64* idx;
    const int* sto_of;                       // [n_sectors * n_agents] -> storage or -1
    StorageView st; EdgeView ev; ModelP mp; i64 t;
    const i64* in_off; const int* in_edge;
    const i64* dem_q; const double* dem_v; const double* exp_costs; const double* opt_val; const double* tpen;
    double* out;
    ACC_HD void operator()(int cell) const {
        const int sec = cell / n_sel, i = cell - sec * n_sel;
        const int a = idx != nullptr ? (int)idx[i] : i;
        const int s = sto_of[(i64)sec * n_agents + a];
        if (s < 0) { out[cell] = ACC_OBS_NAN; return; }
        double r = ACC_OBS_NAN;
        switch (var) {
            case SV_BUSINESS_CONNECTIONS: r = (double)(in_off[s + 1] - in_off[s]); break;
            case SV_CONTENT_Q: r = qf(st.S_q[s]); break;
            case SV_CONTENT_V: r = st.S_v[s]; break;
            // purchase_ = demand_D_ at the end of iterate_purchase (PurchasingManager.cpp:736; both zero on every early return),
            // and "use" returns demand_D as well (Storage.h:141-144)
            case SV_DEMAND_Q: case SV_PURCHASE_Q: case SV_USE_Q: r = qf(dem_q[s]); break;
            case SV_DEMAND_V: case SV_PURCHASE_V: case SV_USE_V: r = dem_v[s]; break;
            case SV_STORAGE_DEMAND_Q: r = qf(dem_q[s] - dem_q[s]); break;                                 // demand_D_ - purchase_, :125-128
            case SV_STORAGE_DEMAND_V: r = dem_v[s] - dem_v[s]; break;
            case SV_DESIRED_USED_FLOW_Q: r = qf(st.Ut_q[s]); break;
            case SV_DESIRED_USED_FLOW_V: r = st.Ut_v[s]; break;
            case SV_EXPECTED_COSTS: r = exp_costs[s]; break;
            case SV_INPUT_FLOW_Q: r = qf(st.Il_q[s]); break;
            case SV_INPUT_FLOW_V: r = st.Il_v[s]; break;
            case SV_OPTIMIZED_VALUE: r = opt_val[s]; break;
            case SV_POSSIBLE_USE_Q: case SV_POSSIBLE_USE_V: {                                             // Storage::last_possible_use_U_hat, Storage.cpp:46-49
                const Flow u = fadd(stock_over_time(mkflow(st.S_q[s], st.S_v[s]), mp.dt), mkflow(st.Il_q[s], st.Il_v[s]));
                r = var == SV_POSSIBLE_USE_Q ? qf(u.q) : u.v;
                break;
            }
            case SV_SHIPMENT_Q: case SV_SHIPMENT_V: {                                                     // get_sum_of_last_shipments, PurchasingManager.cpp:94-98
                Flow sum = mkflow(0, 0.0);
                for (i64 k = in_off[s]; k < in_off[s + 1]; ++k) sum = fadd(sum, obs_last_shipment(ev, t, in_edge[k]));
                r = var == SV_SHIPMENT_Q ? qf(sum.q) : sum.v;
                break;
            }
            case SV_TOTAL_TRANSPORT_PENALTY: r = tpen[s]; break;
            case SV_USED_FLOW_Q: r = qf(st.U_q[s]); break;
            case SV_USED_FLOW_V: r = st.U_v[s]; break;
            default: break;
        }
        out[cell] = r;
    }
};

// one connection's observable; t = step of the last production pass
ACC_HD double observe_connection(int var, const EdgeView& ev, i64 t, i64 e) {
    const int tau = ev.tau[e];
    const unsigned w = ev.slotw[e];
    const i64 at0 = ring_slot_time(ev.rings, w, t);  // what arrived at step t
    switch (var) {
        case XV_INITIAL_FLOW_Q: return qf(ev.Z_q[e]);
        case XV_INITIAL_FLOW_V: return flow_q(ev.Z_q[e]).v;
        case XV_DEMAND_REQUEST_Q: return qf(ev.D_q[e]);
        case XV_DEMAND_REQUEST_V: return ev.D_q[e] > 0 ? q_times_p(ev.D_q[e], ev.D_p[e]) : 0.0;
        case XV_SENT_FLOW_Q: return qf(obs_last_shipment(ev, t, e).q);
        case XV_SENT_FLOW_V: return obs_last_shipment(ev, t, e).v;
        case XV_RECEIVED_FLOW_Q: return qf(ev.conv_q[at0]);
        case XV_RECEIVED_FLOW_V: return ev.conv_v[at0];
        case XV_MINIMUM_PASSAGE: {  // BusinessConnection::get_minimum_passage, BusinessConnection.cpp:69-83
            double m = 1.0;
            if (ev.nu != nullptr && ev.route[e] >= 0)
                for (int h = ev.route_off[ev.route[e]]; h < ev.route_off[ev.route[e] + 1]; ++h) {
                    const double nu = ev.nu[ev.route_ent[h]];
                    if (nu >= 0.0 && nu < m) m = nu;
                }
            return m;
        }
        default: break;
    }
    // what is under way: the delay slots of planes t+1 .. t+delay plus the overflow of every hop (get_transport_flow, :162-171)
    const Flow dlv = mkflow(ev.conv_q[at0], ev.conv_v[at0]);
    Flow sum = mkflow(0, 0.0);
    for (int j = 1; j <= tau; ++j) {
        const i64 at = ring_slot_time(ev.rings, w, t + j);
        sum = fadd(sum, mkflow(ev.conv_q[at], ev.conv_v[at]));
    }
    if (ev.ovf_q != nullptr && ev.route[e] >= 0) {
        const int nh = ev.route_off[ev.route[e] + 1] - ev.route_off[ev.route[e]];
        for (int h = 0; h < nh; ++h) sum = fadd(sum, mkflow(ev.ovf_q[(i64)h * ev.n_edges + e], ev.ovf_v[(i64)h * ev.n_edges + e]));
    }
    switch (var) {
        case XV_FLOW_DEFICIT: return qf(ev.Z_q[e] * (tau + 1) - dlv.q - sum.q);                           // get_flow_deficit, :146-155
        case XV_TOTAL_FLOW_Q: return qf(fadd(sum, dlv).q);                                                // get_total_flow, :157-160
        case XV_TOTAL_FLOW_V: return fadd(fcopy(sum), dlv).v;
        case XV_FLOW_MEAN_Q: case XV_FLOW_MEAN_V: {                                                       // get_flow_mean, :133-144
            if (tau == 0) return ACC_OBS_NAN;  // the reference divides by a zero delay here
            const Flow tot = fadd(dlv, sum);
            const Flow mean = mkflow(mkq(qf(tot.q) / (double)tau), tot.v / (double)tau);
            return var == XV_FLOW_MEAN_Q ? qf(mean.q) : mean.v;
        }
        default: return ACC_OBS_NAN;
    }
}

// flows by seller: one CTA per selected selling agent i (idx == nullptr: agent i), out[i * n_to + buyer]
struct ObserveFlowsFrom {
    enum { THREADS = 256, MIN_BLOCKS = 1 };
    int var, n_to;
    const i64* idx;
    EdgeView ev; i64 t;
    const int* type; const int* edge_to;
    double* out;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int i) const {
        const int a = idx != nullptr ? (int)idx[i] : i;
        if (type[a] != 1) return;
        const int e0 = ev.seller_off[a], n = ev.seller_off[a + 1] - e0;
        cx.foreach(n, [&](int k) { out[(i64)i * n_to + edge_to[e0 + k]] = observe_connection(var, ev, t, (i64)e0 + k); });
    }
};

// flows by buyer: one CTA per (selected buying agent i, one of its storages), out[seller * n_sel + i]
struct ObserveFlowsTo {
    enum { THREADS = 128, MIN_BLOCKS = 1 };
    int var, n_sel, max_storages;
    const i64* idx;
    EdgeView ev; i64 t;
    const int* st_off; const int* st_list; const i64* in_off; const int* in_edge; const int* in_seller; const int* firm_of_row;
    double* out;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int item) const {
        const int i = item / max_storages, k = item - i * max_storages;
        const int a = (int)idx[i];
        if (st_off[a] + k >= st_off[a + 1]) return;
        const int s = st_list[st_off[a] + k];
        const i64 o = in_off[s];
        cx.foreach((int)(in_off[s + 1] - o), [&](int j) {
            const int seller = firm_of_row[(unsigned)in_seller[o + j] & 0xffffffu];  // in_seller holds the seller ROW | delay << 24
            out[(i64)seller * n_sel + i] = observe_connection(var, ev, t, in_edge[o + j]);
        });
    }
};

}  // namespace acc
