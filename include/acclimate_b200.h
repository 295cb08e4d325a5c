/* acclimate_b200.h -- C ABI of the B200 engine for Acclimate's per-timestep agent iteration.
 *
 * The reference has no plugin registry; the path sits behind the phase interface of its Model class.
 * Each entry point below names the reference interface it replaces (paths relative to the reference
 * repository root).  Conventions follow the reference's own (stale) C shim src/external.cpp:56-122:
 * every function returns 0 on success and -1 on error, the message is available from
 * acc_last_error(); no C++ exception crosses the boundary; the engine owns all device memory, host
 * buffers are caller-owned.  One engine = one CUDA device = one host thread at a time.
 *
 * There is no CPU implementation behind this interface: acc_create fails when no CUDA device is
 * available.
 */
#ifndef ACCLIMATE_B200_H
#define ACCLIMATE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct acc_engine acc_engine;

/* order of acc_params.model[] -- Parameters::ModelParameters, include/parameters.h:33-84, and
 * model.delta_t / model.no_self_supply, src/input/ModelInitializer.cpp:58-62,1015-1079 */
enum {
    ACC_MP_TRANSPORT_PENALTY_SMALL = 0,
    ACC_MP_TRANSPORT_PENALTY_LARGE,
    ACC_MP_MIN_STORAGE,
    ACC_MP_CHEAPEST_PRICE_RANGE_WIDTH, /* < 0: "auto" */
    ACC_MP_OPTIMIZATION_MAXITER,
    ACC_MP_OPTIMIZATION_PRECISION_ADJUSTMENT,
    ACC_MP_QUADRATIC_TRANSPORT_PENALTY,
    ACC_MP_RELATIVE_TRANSPORT_PENALTY,
    ACC_MP_MAX_DECREASE_RESERVATION_PRICE_LIMITED_BY_MARKUP,
    ACC_MP_DEVIATION_PENALTY,
    ACC_MP_CHEAPEST_PRICE_RANGE_PRESERVE_SELLER_PRICE,
    ACC_MP_NO_SELF_SUPPLY,
    ACC_MP_UTILITARIAN,
    ACC_MP_DELTA_T,
    ACC_MP_NAIVE_EXPECTATIONS,
    ACC_MP_ALWAYS_EXTEND_EXPECTED_DEMAND_CURVE,
    ACC_MP_COUNT
};

/* order of each row of acc_network.sector_params -- sectors.<name> keys,
 * src/input/ModelInitializer.cpp:159-185 */
enum {
    ACC_SP_UPPER_STORAGE_LIMIT = 0,
    ACC_SP_INITIAL_STORAGE_FILL_FACTOR,
    ACC_SP_TRANSPORT_TYPE, /* 0 aviation, 1 immediate, 2 roadsea */
    ACC_SP_SUPPLY_ELASTICITY,
    ACC_SP_PRICE_INCREASE_PRODUCTION_EXTENSION,
    ACC_SP_ESTIMATED_PRICE_INCREASE_PRODUCTION_EXTENSION,
    ACC_SP_INITIAL_MARKUP,
    ACC_SP_TARGET_STORAGE_REFILL_TIME,
    ACC_SP_TARGET_STORAGE_WITHDRAW_TIME,
    ACC_SP_COUNT
};

/* Flat snapshot of what the reference's ModelInitializer builds
 * (src/input/ModelInitializer.cpp:94-213 agents and connections, :689-700 one transport connection
 * per region pair, :1081-1085 post_initialize).  All arrays are caller-owned and copied. */
typedef struct {
    int64_t n_sectors;            /* including the consumer sector (last) */
    int64_t n_regions;
    int64_t n_agents;
    int64_t n_edges;
    const int64_t* agent_type;    /* 1 firm, 0 consumer */
    const int64_t* agent_sector;
    const int64_t* agent_region;
    const double* agent_beta;     /* firms.<..>.possible_overcapacity_ratio */
    const int64_t* edge_from;     /* seller agent; edges must be listed seller-major */
    const int64_t* edge_to;       /* buying agent */
    const int64_t* edge_zq;       /* initial flow Z* in 1e-3 quanta */
    const int64_t* delay;         /* [n_regions * n_regions] transport delay between regions, symmetric */
    const double* sector_params;  /* [n_sectors * ACC_SP_COUNT] */
    const double* consumer_elasticity; /* [n_sectors] consumers.<..>.consumption_price_elasticity by supplying sector */
    /* consumer baskets of the nested-CES utility (consumers.<..>.consumer_baskets / inter_basket_substitution_coefficient,
     * src/input/ModelInitializer.cpp:118-139); only read when model[ACC_MP_UTILITARIAN] != 0 */
    int64_t n_baskets;
    const int64_t* basket_of_sector;   /* [n_sectors] basket of each supplying sector, -1: in no basket; sectors of a
                                          basket are taken in ascending order */
    const double* basket_substitution; /* [n_baskets] intra-basket substitution coefficient */
    double inter_basket_substitution;  /* inter_basket_substitution_coefficient */
    /* multi-hop transport routes (the "network" transport, src/input/ModelInitializer.cpp:339-607): the geo entities
     * (connections, ports, sea areas) a shipment passes between two regions, each a TransportChainLink with its own
     * delay and passage forcing (src/model/TransportChainLink.cpp:40-97).  n_entities == 0: one connection per region
     * pair with the delays of `delay` above */
    int64_t n_entities;
    const int64_t* entity_delay;       /* [n_entities] steps */
    const int64_t* route_off;          /* [2 * n_regions * n_regions + 1]: route (k * n_regions + from) * n_regions + to, k = 0 aviation, 1 road/sea */
    const int64_t* route_entity;       /* entities of each route in travel order */
} acc_network;

typedef struct {
    double model[ACC_MP_COUNT];
} acc_params;

/* EventType values are the reference's (include/ModelRun.h:39-55) */
typedef struct {
    int64_t step;
    int32_t type;
    int32_t a;   /* sector index (storage events) or agent index */
    int32_t b;   /* agent index or -1 */
    double value;
} acc_event;

/* ---- network ingest: the step BEFORE the path, on the device ----
 * Replaces the dense branch of ModelInitializer::build_agent_network_from_table (src/input/ModelInitializer.cpp:832-834,
 * :947-965), initialize_connection's self-supply rule (:187-190) and clean_network (:215-299): flows[n_agents][n_agents] are
 * annual flows as float (row = selling agent), a cell becomes a connection when  flow > threshold  and its quantised
 * per-timestep flow  round(FlowQuantity(flow) * time_factor)  exceeds the quantised per-timestep threshold; then firms without
 * value added, buyers or inputs (and consumers without inputs) are removed in the reference's sweep order until nothing
 * changes.  agent_alive[n_agents] (may be NULL) tells who is left; the surviving connections come back seller-major, buyer
 * ascending, in ORIGINAL agent numbers (edge_* may be NULL: *n_edges then only reports how many there are).  The caller
 * renumbers the survivors and hands the result to acc_create (acclimate_b200/ingest.py does). */
int acc_ingest_table(int device, int64_t n_agents, int64_t n_sectors, const int64_t* agent_type, const int64_t* agent_sector, const float* flows,
                     double threshold, double time_factor, int no_self_supply, uint8_t* agent_alive, int64_t* n_edges, int64_t* edge_from,
                     int64_t* edge_to, int64_t* edge_zq, size_t cap_edges, int64_t* sweeps, int64_t* removed);

/* replaces Model::Model + ModelInitializer::initialize + Model::start
 * (include/model/Model.h:62,72; src/ModelRun.cpp:108-116,163-168) */
acc_engine* acc_create(const acc_network* net, const acc_params* params, int device);
/* replaces ~Model */
void acc_destroy(acc_engine* e);

/* ---- ensembles of scenarios on one network (BASELINE.json configs[4]; SURVEY 8e "replicate topology once per GPU, batch state
 * along a scenario axis") ----
 * No counterpart in the reference, which runs one scenario per process.  acc_fork makes an ensemble MEMBER: a new engine on the
 * parent's device that shares the parent's immutable tables -- the network itself: connections, Z*, delays, routes, both
 * buyer-side lists, agent and sector constants -- and owns only the state (what acc_save_state names), copied from the parent
 * as it is at the time of the call.  Parent and members are then forced, stepped and read independently, each on its own
 * stream; the shared tables are released when the last engine that uses them is destroyed, in whatever order. */
acc_engine* acc_fork(acc_engine* parent);

/* ---- region-partitioned run: one engine per device, each owning the agents of a block of regions ----
 * The reference parallelises the same loops over agents with OpenMP on shared memory (src/model/Model.cpp:59-118);
 * across devices the agents are split by region and the two per-connection messages of a step cross the wire:
 * shipments seller -> buyer after SalesManager::distribute (BusinessConnection::push_flow_Z,
 * src/model/BusinessConnection.cpp:87-94) and demand requests buyer -> seller after iterate_purchase
 * (BusinessConnection::send_demand_request_D, :127-131), plus the sellers' supply parameters every purchasing manager
 * reads (SalesManager::communicated_parameters, src/model/SalesManager.cpp:611-624).
 *
 * net lists ALL agents (global numbering) but only the connections with an owned seller or an owned buyer, in global
 * file order; agent_owned[n_agents] marks the agents this engine computes.  The engine never communicates itself: the
 * caller moves the exchange buffers between the stages (NCCL all-to-all / all-reduce in acclimate_b200/partition.py). */
acc_engine* acc_create_partition(const acc_network* net, const acc_params* params, int device, const uint8_t* agent_owned);
/* Halo lists and caller-provided exchange buffers (memory of the engine's device, valid for the engine's life):
 *   send_edges[n_send]: local ids of connections whose seller is owned and whose buyer is not, grouped by peer
 *   recv_edges[n_recv]: local ids of connections whose buyer is owned and whose seller is not, grouped by peer
 *   flow_send [n_send][2], flow_recv [n_recv][2]: shipment (quantity quanta, f64 value bits) as int64 pairs
 *   demand_send [n_recv][2], demand_recv [n_send][2]: demand request (quantity quanta, price quanta)
 *   seller_table [n_agents][ACC_SELLER_WORDS] int64: supply parameters, zero outside the owned rows when packed, so
 *   that an integer SUM all-reduce assembles the full table exactly */
#define ACC_SELLER_WORDS 9
int acc_set_halo(acc_engine* e, const int64_t* send_edges, size_t n_send, const int64_t* recv_edges, size_t n_recv, void* flow_send,
                 void* flow_recv, void* demand_send, void* demand_recv, void* seller_table);
/* one timestep in four stages; between them the caller exchanges, in order:
 *   after SELL:     flow_send -> peers' flow_recv            (all-to-all)
 *   after DELIVER:  seller_table                             (all-reduce SUM over int64)
 *   after PURCHASE: demand_send -> peers' demand_recv        (all-to-all)
 * CLOSE applies the received demand requests and ticks.  On an unpartitioned engine the four stages equal acc_step. */
enum { ACC_STAGE_SELL = 0, ACC_STAGE_DELIVER = 1, ACC_STAGE_PURCHASE = 2, ACC_STAGE_CLOSE = 3 };
int acc_stage(acc_engine* e, int stage);

/* ---- the same run with the engine exchanging its halos itself: NCCL from C++, no host code in the data plane ----
 * (no counterpart in the reference, which is shared-memory only)  One rank makes the 128-byte NCCL id and hands
 * it to the others by any means (MPI, a file, torch.distributed ...); every rank then joins with acc_comm_init, which also
 * says how the halo lists of acc_set_halo split over the peers (send_counts / recv_counts [world], entries grouped by peer
 * in rank order) and which part owns each agent (agent_part [n_agents], the same on every rank).  acc_set_halo may be
 * given five NULL buffers: the engine then allocates the exchange buffers itself.
 * acc_step_partitioned = one timestep: the four stages of acc_stage with, between them, on the engine's stream,
 *   shipments:       one ncclSend / ncclRecv group with the peers (16 bytes per cross connection),
 *   supply records:  ncclAllGather of the rows each part's own firms published (72 bytes per firm; not the zero-padded full
 *                    table of the acc_stage protocol),
 *   demand requests: one ncclSend / ncclRecv group.
 * NCCL is bound at run time (dlopen of libnccl.so.2, the copy already in the process first): an unpartitioned run does not
 * need it.  A P-part run is bit-identical to the one-engine run. */
#define ACC_COMM_ID_BYTES 128
int acc_comm_unique_id(void* id);
int acc_comm_init(acc_engine* e, const void* id, int rank, int world, const int32_t* agent_part, const int64_t* send_counts,
                  const int64_t* recv_counts);
int acc_step_partitioned(acc_engine* e);
/* Peer stores (call once after acc_comm_init, before the first step; all parts; needs the engine-owned exchange buffers of
 * acc_set_halo(..., NULL x 5)).  The parts map each other's halo RECEIVE buffers into their address spaces (CUDA IPC over
 * NVLink / NVSwitch); from then on the kernel that produces a shipment or a demand request for a connection whose other end
 * lives on another part stores the 16-byte entry straight into that part's receive buffer, at its position in the halo order,
 * next to its local store.  The pack kernels and the two NCCL send / receive groups of acc_step_partitioned disappear: the
 * transfer rides under the sellers' pass and under the purchasing kernel, and two one-word all-reduces per step order the
 * parts (all shipments landed / all demand requests landed) before the receiving part applies its buffer.  Results are
 * unchanged bit for bit.  Fails (and leaves the NCCL exchanges in place) where peer access is not possible. */
int acc_comm_enable_peer_stores(acc_engine* e);
/* All parts must run the same protocol: when the call failed on ANY part (the caller agrees on that, e.g. with a MIN all-reduce
 * of the return codes, which also orders the parts before the first step), every part calls this and the NCCL exchanges stay. */
int acc_comm_disable_peer_stores(acc_engine* e);
/* collectives (point-to-point groups and all-gathers) the engine has issued since acc_comm_init */
int64_t acc_collective_count(const acc_engine* e);

/* Threads that solve one purchasing manager's optimisation.  The order of the solver's reductions -- part of its
 * floating-point arithmetic -- follows from it, so all parts of a partitioned run must agree: each part derives the
 * value from its own largest manager, the caller takes the maximum over the parts and sets it everywhere before the
 * first step (acclimate_b200/partition.py does). */
int acc_solver_group(const acc_engine* e);
int acc_set_solver_group(acc_engine* e, int group);

/* replaces EconomicAgent::set_forcing for all agents at once (src/model/EconomicAgent.cpp:43-47);
 * lambda has n_agents entries */
int acc_set_agent_forcing(acc_engine* e, const double* lambda, size_t n);
/* replaces GeoEntity::set_forcing_nu (src/model/GeoEntity.cpp:30-34); nu has one entry per geo entity -- for a
 * snapshot without routes one per region pair (i, j < i) at index i*(i-1)/2 + j; nu < 0 means "no forcing" */
int acc_set_entity_passage(acc_engine* e, const double* nu, size_t n);

/* one timestep = Model::switch_registers + iterate_consumption_and_production + iterate_expectation
 * + iterate_purchase + iterate_investment + tick (src/ModelRun.cpp:178-190,226-227) */
int acc_step(acc_engine* e);
/* the same, split by phase as ModelRun::run calls them */
enum { ACC_PHASE_CONSUMPTION_AND_PRODUCTION = 0, ACC_PHASE_EXPECTATION = 1, ACC_PHASE_PURCHASE = 2, ACC_PHASE_TICK = 3 };
int acc_phase(acc_engine* e, int phase);
/* block until the device has finished all queued phases -- including the consumers' utility optimisation of the NEXT step, which
 * the engine starts on a side stream as soon as this step's storages are final (it only reads model state; its results are applied
 * at the start of the next step, and dropped if acc_load_state replaces the state in between).  Raises the error the reference
 * ends a run with when a consumer's optimisation failed outright (src/model/Consumer.cpp:505-512). */
int acc_sync(acc_engine* e);

/* replaces ModelRun::event (src/ModelRun.cpp:255-275): *n receives the number of pending events; up to cap of them
 * (ordered by step, type, indices) are copied to buf and removed, the rest stays pending.  buf == NULL with cap == 0
 * only counts.  The engine buffers 65 536 events between two drains (acc_set_event_capacity changes that): when more
 * were raised the call delivers what it has and then FAILS, naming the number of events that were dropped. */
int acc_drain_events(acc_engine* e, acc_event* buf, size_t cap, size_t* n);
int acc_set_event_capacity(acc_engine* e, size_t capacity);

/* Observable pull, replaces the observe<>() visitor sweep of ArrayOutput::iterate
 * (src/output/ArrayOutput.cpp:479-646): gathers one variable of one group ("firms", "consumers",
 * "sectors", "regions", "storages", "flows"; variable names of the reference, Flow/Stock
 * observables split into <name>_quantity / <name>_value as ArrayOutput.cpp:58-66 does) into dst
 * as doubles (to_float), shaped like ArrayOutput::Variable::data.  *n receives the element count. */
int acc_read(acc_engine* e, const char* group, const char* variable, double* dst, size_t cap, size_t* n);

/* The same pull with SELECTION and the reference's array shapes -- what ArrayOutput::iterate writes for one timestep
 * (src/output/ArrayOutput.cpp:479-646; the observable lists are the observe<>() visitors of include/model/EconomicAgent.h:70-96,
 * Firm.h:78-153, Consumer.h:98-110, Sector.h:73-90, Region.h, Storage.h:90-150, BusinessConnection.h:79-115):
 *   "firms", "consumers"   [n0 or n_agents]            idx0 = agents; agents of the other kind are NaN
 *   "sectors", "regions"   [n0 or n_sectors/n_regions] idx0 = sectors / regions
 *   "storages"             [n_sectors][n1 or n_agents] idx1 = agents (no selection by sector, as in the reference, :575)
 *   "flows"                [n0 or n_agents][n1 or n_agents]  idx0 = selling agents OR idx1 = buying agents (not both, :469-471)
 * Cells without an object (no storage of that sector, no connection between the two agents) are NaN (:170).  The cells are
 * gathered by kernels on the device -- only the selected ones are touched, nothing else of the state moves -- and arrive in
 * dst (caller-owned, cap doubles) in one copy; dst == NULL only reports the size in *n.  Valid after the purchase phase of a
 * step, before or after its tick. */
int acc_observe(acc_engine* e, const char* group, const char* variable, const int64_t* idx0, size_t n0, const int64_t* idx1, size_t n1,
                double* dst, size_t cap, size_t* n);

/* No counterpart in the reference: added for callers that cannot allocate device
 * memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same numbers as
 * acc_read(e, "firms", "incoming_demand_quantity" / "incoming_demand_value", ...) -- the sum of the demand requests each
 * firm holds, one entry per agent, zero for consumers -- but the per-firm sums are formed in a file-scope device array
 * instead of freshly allocated buffers: the call allocates and frees nothing on the device.  Limits: at most
 * ACC_STATIC_SCRATCH_AGENTS agents (more is an error); every GPU that loads the library holds the 1 MiB array until the
 * process exits; all calls share it, so calls that overlap on one GPU race.  out_quantity / out_value are caller-owned host
 * buffers of cap doubles, either may be NULL (both NULL: *n receives the count only). */
#define ACC_STATIC_SCRATCH_AGENTS 65536
int acc_firm_incoming_demand_static(acc_engine* e, double* out_quantity, double* out_value, size_t cap, size_t* n);

/* Raw state for parity checks: int64 / f64 arrays named like the oracle's per-step dump
 * (oracle/ref_harness/harness.cpp dump_step).  is_f64 tells which type was written to dst.
 * Three names have no counterpart in the dump (diagnostics): "sto_solver_iters" / "sto_solver_evals" (SQP iterations and objective
 * evaluations of every purchasing manager's last solve), "consumer_solver_ret" (return code of every consumer's last utility
 * optimisation, csrc/sqp_core.h) and "firm_sorted_sums" (1 for the firms whose last pro-rata distribution took its demand sums in
 * the reference's sorted order, csrc/model_core.cuh ordered_range_sums). */
int acc_read_state(acc_engine* e, const char* name, void* dst, size_t cap_bytes, size_t* n, int* is_f64);

/* counters */
int64_t acc_edge_count(const acc_engine* e);
int64_t acc_step_count(const acc_engine* e);
/* kernels launched since acc_create */
int64_t acc_launch_count(const acc_engine* e);
/* CUDA stream the engine launches on (cudaStream_t as void*) */
void* acc_stream(acc_engine* e);
/* per-kernel accumulated device time in ms since the last call with reset != 0 (needs acc_profile(e, 1)):
 * names[i] are static strings, ms[i] / launches[i] the totals; returns the number of kernels */
int acc_profile(acc_engine* e, int enable);
int acc_kernel_times(acc_engine* e, const char** names, double* ms, int64_t* launches, int cap, int reset);

/* ---- state checkpoint ----
 * Replaces the reference's process-level DMTCP checkpoint (reference src/checkpointing.cpp:40-68, the request handled in
 * the time loop at src/ModelRun.cpp:208-223): everything that changes from step to step -- storages, firms, demand
 * requests, the conveyor planes, overflows, forcing, the consumers' previous consumption, pending events, the step
 * counter -- as one flat little-endian blob in a caller-owned HOST buffer.  The topology is not part of it: a blob loads
 * into an engine created from the same snapshot (sizes are checked).  acc_state_bytes tells the size the blob has NOW
 * (it grows once when passage forcing is first used). */
int acc_state_bytes(acc_engine* e, size_t* bytes);
int acc_save_state(acc_engine* e, void* dst, size_t cap, size_t* written);
int acc_load_state(acc_engine* e, const void* src, size_t bytes);

const char* acc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
