// ingest.cuh -- network ingest on the device: a candidate MRIO flow table becomes the connection list of a
// clean network.
//
// Reference: src/input/ModelInitializer.cpp
//   :832-834, :947-965  dense branch of build_agent_network_from_table: annual flows as float [agents][agents], a cell becomes
//                       a connection when  *d > threshold && flow > daily_threshold,  flow = round(FlowQuantity(*d) * time_factor)
//   :187-213            initialize_connection (no self-supply when model.no_self_supply)
//   :215-299            clean_network: sweeps over the agents in order; a firm without value added, without buyers or without
//                       inputs (a lone self-supply connection counts as none) is removed AT ONCE -- its buyers' storages and
//                       its suppliers' production shrink before the next agent is examined -- until a sweep removes nothing
//
// The table stays on the device as a dense int64 matrix Z[seller][buyer] of quantised flows (0 = no connection).  Per agent
// the device keeps what the removal test reads: production X* (row sum), total input (column sum), number of buyers, number
// of input storages (supplying sectors with at least one connection), self-supply flag.  The reference's sweep is sequential
// and order-dependent (a removal changes the value added of later agents), so it is replayed exactly: one kernel finds the
// FIRST removable agent at or after a cursor, one kernel takes it out (its row and its column in parallel), the cursor moves
// past it; a sweep ends when nothing is found, the fix point when a whole sweep found nothing.  The number of kernel pairs is
// the number of removed agents, not the number of agents.
#pragma once

#include "coop.cuh"

namespace acc {

struct IngestView {
    i64 nA; int nS;
    const int* type; const int* sector;
    i64* Z;                 // [nA][nA] quantised flows, 0 = none
    i64* X; i64* in_sum;    // production (row sum), input (column sum)
    int* n_out; int* n_st;  // buyers, input storages
    int* cnt;               // [nS][nA] connections per (supplying sector, buyer)
    unsigned char* alive; unsigned char* self;
};

// cell (i, j) of the float table -> quantised flow or 0
struct IngestFilter {
    IngestView v; const float* table; double threshold, time_factor; i64 daily_threshold; int no_self_supply;
    ACC_HD void operator()(int cell_) const {
        const i64 cell = cell_;
        const i64 i = cell / v.nA, j = cell - i * v.nA;
        const float d = table[cell];
        i64 z = 0;
        if (v.type[i] == 1 && !(no_self_supply && i == j)) {
            const i64 flow = qmul(mkq((double)d), time_factor);  // round(FlowQuantity(*d) * time_factor)
            if ((double)d > threshold && flow > daily_threshold) z = flow;
        }
        v.Z[cell] = z;
    }
};

struct RowCol {
    i64 x, in; int n_out, pad;
    static ACC_HD RowCol identity() { RowCol r; r.x = 0; r.in = 0; r.n_out = 0; r.pad = 0; return r; }
    static ACC_HD RowCol combine(RowCol p, RowCol q) { RowCol r; r.x = p.x + q.x; r.in = p.in + q.in; r.n_out = p.n_out + q.n_out; r.pad = 0; return r; }
};
// one CTA per agent: its row (what it sells) and its column (what it buys)
struct IngestStats {
    enum { THREADS = 256, MIN_BLOCKS = 1 };
    IngestView v;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int a) const {
        const i64 nA = v.nA;
        const RowCol t = cx.template reduce<RowCol>((int)nA, [&](int k) {
            RowCol r;
            const i64 zo = v.Z[(i64)a * nA + k], zi = v.Z[(i64)k * nA + a];
            r.x = zo; r.in = zi; r.n_out = zo > 0 ? 1 : 0; r.pad = 0;
            return r;
        });
        // connections per supplying sector into this buyer: the sellers of one sector are counted by the threads in turn
        cx.foreach(v.nS, [&](int s) {
            int c = 0;
            for (i64 k = 0; k < nA; ++k) if (v.sector[k] == s && v.Z[k * nA + a] > 0) ++c;
            v.cnt[(i64)s * nA + a] = c;
        });
        if (cx.leader()) {
            int n_st = 0;
            for (int s = 0; s < v.nS; ++s) n_st += v.cnt[(i64)s * nA + a] > 0 ? 1 : 0;
            v.X[a] = t.x; v.in_sum[a] = t.in; v.n_out[a] = t.n_out; v.n_st[a] = n_st;
            v.alive[a] = 1; v.self[a] = v.Z[(i64)a * nA + a] > 0 ? 1 : 0;
        }
    }
};

// the removal test of clean_network (:229-236, :279): true when agent a would be removed now
ACC_HD bool ingest_removable(const IngestView& v, i64 a) {
    if (!v.alive[a]) return false;
    if (v.type[a] == 1) {
        const bool self = v.self[a] != 0;
        return v.X[a] - v.in_sum[a] <= 0 || v.n_out[a] == 0 || (v.n_out[a] == 1 && self) || v.n_st[a] == 0 || (v.n_st[a] == 1 && self);
    }
    return v.n_st[a] == 0;
}
struct IngestFindFirst {  // first removable agent at or after the cursor -> *found (initialised to nA)
    IngestView v; i64 cursor; i64* found;
    ACC_HD void operator()(int k) const {
        const i64 a = cursor + k;
        if (!ingest_removable(v, a)) return;
#if defined(__CUDA_ARCH__)
        atomicMin(reinterpret_cast<long long*>(found), (long long)a);
#else
        if (a < *found) *found = a;
#endif
    }
};
// one CTA takes agent a out: its buyers lose the input (:253-260), its suppliers the buyer (:263-271)
struct IngestRemove {
    enum { THREADS = 256, MIN_BLOCKS = 1 };
    IngestView v; i64 a;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int) const {
        const i64 nA = v.nA;
        const int sa = v.sector[a];
        cx.foreach((int)nA, [&](int j) {
            const i64 zo = v.Z[a * nA + j];
            if (zo > 0) {  // buyer j: Storage::subtract_initial_flow_Z_star / PurchasingManager::remove_business_connection
                v.in_sum[j] -= zo;
                const int left = --v.cnt[(i64)sa * nA + j];
                if (left == 0) v.n_st[j] -= 1;
                v.Z[a * nA + j] = 0;
            }
            if (j != a) {
                const i64 zi = v.Z[(i64)j * nA + a];
                if (zi > 0) {  // supplier j: Firm::subtract_initial_production_X_star / SalesManager::remove_business_connection
                    v.X[j] -= zi;
                    v.n_out[j] -= 1;
                    v.Z[(i64)j * nA + a] = 0;
                }
            }
        });
        if (cx.leader()) { v.alive[a] = 0; v.self[a] = 0; v.n_out[a] = 0; v.n_st[a] = 0; v.X[a] = 0; v.in_sum[a] = 0; }
        cx.foreach(v.nS, [&](int s) { v.cnt[(i64)s * nA + a] = 0; });
    }
};

// connection list out of the cleaned matrix: seller-major, buyer ascending (the order acc_create expects)
struct IngestCount {
    enum { THREADS = 256, MIN_BLOCKS = 1 };
    IngestView v; i64* row_count;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int a) const {
        const i64 n = cx.isum((int)v.nA, [&](int j) { return (i64)(v.Z[(i64)a * v.nA + j] > 0 ? 1 : 0); });
        if (cx.leader()) row_count[a] = n;
    }
};
struct IngestEmit {
    enum { THREADS = 256, MIN_BLOCKS = 1 };
    IngestView v; const i64* row_off; i64* from; i64* to; i64* zq;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int a) const {
        const i64 base = row_off[a];
        cx.compact((int)v.nA, [&](int j) { return v.Z[(i64)a * v.nA + j] > 0; },
                   [&](int m, int j) { from[base + m] = a; to[base + m] = j; zq[base + m] = v.Z[(i64)a * v.nA + j]; });
    }
};

}  // namespace acc
