// engine.cu -- device-resident Acclimate timestep engine and its C ABI (include/acclimate_b200.h).
//
// Data layout in HBM (DESIGN.md "Data layout"):
//   * BusinessConnections are stored once, CSR by seller (the file order of the snapshot).  Per
//     connection: Z* (int64), last demand request D as (quantity, price) int64 pair, delay tau (u8),
//     domestic flag (u8), geo entity (int32).
//   * TransportChainLink delay queues are a "conveyor": every connection owns a ring of tau + 1 slots (int64
//     quantity + f64 value); connections of equal delay form a ring group stored plane by plane (model_core.cuh
//     ConveyorRings).  The shipment of step t on a connection with delay tau is written to slot (t + tau) mod
//     (tau + 1) and arrives when slot t mod (tau + 1) is read -- every connection advances by exactly one slot per
//     step (reference TransportChainLink.cpp:69-71), so the ring shift is pure index arithmetic and nothing is ever
//     moved.
//   * Buyer side: storages are numbered sector-major / buyer-minor so that neighbouring storages read
//     neighbouring connections; incoming connections are listed per storage (CSC) twice: sliced-ELL
//     (k-major in slices of 32 storages) for the thread-per-storage delivery pass and storage-major
//     for the warp-per-purchasing-manager pass.
//
// Kernels of one step (reference phase they implement):
//   seller_production          CapacityManager::get_possible_production_X_hat (one warp, while the demand book is staged),
//                              SalesManager::calc_production_X + distribute              (phase 1)
//   consumer_utility           Consumer::iterate_consumption_and_production: the optimisation, on a side stream, started
//                              behind storage_delivery of the step BEFORE (it only reads state)       (phase 1)
//   consumer_apply             ... its results become the step's desired use                           (phase 1)
//   storage_delivery           Firm/Consumer input use, Storage::calc_content_S, TransportChainLink::push_flow_Z
//                              arrivals, Storage::push_flow_Z accumulation               (phase 1)
//   seller_expectation         SalesManager::iterate_expectation, Firm::iterate_expectation (phase 2)
//   purchase                   PurchasingManager::iterate_purchase                       (phase 3)
//
// Compiled twice from this one source: by nvcc into libacclimate_b200.so (the product; CUDA only, no
// CPU path) and -- with -DACC_HOST_EMU, by g++ -- into a test-only library in which every "launch" is
// a host loop over the same templated routines (tests/emu; lets the CPU-only test tier check the
// kernel logic against the oracle).  The emulation is never part of the product library.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <vector>

#include "../../include/acclimate_b200.h"
#include "purchase_core.cuh"
#include "consumer_core.cuh"
#include "observe.cuh"
#include "ingest.cuh"

#ifndef ACC_HOST_EMU
#include <dlfcn.h>
#endif

#ifndef ACC_PM_MAXREG_ANY
#define ACC_PM_MAXREG_ANY 80  // ... at any other group size: two CTAs of 384 threads (748 connections per manager) per SM; measured on a
                              // 748x3 network: 1.81 ms with 112 registers (one CTA per SM), 1.30 ms with 80
#endif
#ifndef ACC_PM_PREFETCH_AHEAD
#define ACC_PM_PREFETCH_AHEAD 444  // purchasing managers ahead whose connection lists are prefetched into the L2 (half a wave of 6 CTAs x 148 SMs)
#endif
#ifndef ACC_PM_MAXREG_FIXED
#define ACC_PM_MAXREG_FIXED 96  // registers per thread of the purchase kernel at the base group size (see struct Purchase)
#endif

namespace acc {

// ------------------------------------------------------------------------------------------------
// backend: memory + launches
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;
struct EngineError {
    std::string msg;
};
[[noreturn]] static void fail(const std::string& m) { throw EngineError{m}; }

#ifdef ACC_HOST_EMU
typedef void* stream_t;
static void* dev_alloc(size_t bytes) {
    void* p = std::calloc(1, bytes ? bytes : 1);
    if (!p) fail("out of host memory (emulation)");
    return p;
}
static void dev_free(void* p) { std::free(p); }
static void h2d(void* d, const void* h, size_t b, stream_t) { std::memcpy(d, h, b); }
static void d2h(void* h, const void* d, size_t b, stream_t) { std::memcpy(h, d, b); }
static void dev_zero(void* d, size_t b, stream_t) { std::memset(d, 0, b); }
static void dev_sync(stream_t) {}
template <class K>
static void launch_thread(stream_t, int n, K k) { for (int i = 0; i < n; ++i) k(i); }
template <class K>
static void launch_block(stream_t, int n, K k) { for (int i = 0; i < n; ++i) { HostCtx cx(K::THREADS); k(cx, i); } }
template <class K>
static void launch_seller(stream_t, int n, int, int, K k) {
    for (int i = 0; i < n; ++i) {
        HostCtx warp(32), cx(K::THREADS);
        k.prepare(warp, i);
        k(cx, i, SellerBook{nullptr, nullptr, nullptr, nullptr, 0});
    }
}
template <class K>
static void launch_warp(stream_t, int n, size_t smem, K k) {
    std::vector<unsigned char> sm(smem + 16);
    for (int i = 0; i < n; ++i) { HostCtx cx(32); k(cx, i, sm.data()); }
}
template <int G, class K>
static void launch_block_smem(stream_t, int n, int threads, size_t smem, K k) {
    std::vector<unsigned char> sm(smem + 16);
    for (int i = 0; i < n; ++i) { HostCtx cx(threads); k(cx, i, sm.data()); }
}
template <class K>
static void launch_warps(stream_t s, int n, size_t smem_per_warp, K k) { launch_warp(s, n, smem_per_warp, k); }
struct DeviceGuard { explicit DeviceGuard(int) {} };
#else
typedef cudaStream_t stream_t;
#define ACC_CUDA(x)                                                                                   \
    do {                                                                                              \
        cudaError_t err__ = (x);                                                                      \
        if (err__ != cudaSuccess) fail(std::string(#x) + ": " + cudaGetErrorString(err__));           \
    } while (0)
static void* dev_alloc(size_t bytes) {
    void* p = nullptr;
    ACC_CUDA(cudaMalloc(&p, bytes ? bytes : 1));
    return p;
}
static void dev_free(void* p) { cudaFree(p); }
static void h2d(void* d, const void* h, size_t b, stream_t s) { if (b) ACC_CUDA(cudaMemcpyAsync(d, h, b, cudaMemcpyHostToDevice, s)); }
static void d2h(void* h, const void* d, size_t b, stream_t s) {
    if (b) ACC_CUDA(cudaMemcpyAsync(h, d, b, cudaMemcpyDeviceToHost, s));
    ACC_CUDA(cudaStreamSynchronize(s));
}
static void dev_zero(void* d, size_t b, stream_t s) { if (b) ACC_CUDA(cudaMemsetAsync(d, 0, b, s)); }
static void dev_sync(stream_t s) { ACC_CUDA(cudaStreamSynchronize(s)); }
// Every C ABI entry makes the engine's device current for its duration (a caller may drive several engines, one per
// device, from one thread) and puts the caller's device back afterwards.
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != device) ACC_CUDA(cudaSetDevice(device));
        else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device (per-context) attribute: what has been configured is
// remembered per device, not per process.
static bool smem_needs_config(const void* kernel, size_t bytes) {
    static std::map<std::pair<int, const void*>, size_t> configured;
    int dev = 0;
    ACC_CUDA(cudaGetDevice(&dev));
    size_t& have = configured[std::make_pair(dev, kernel)];
    if (bytes <= have) return false;
    have = bytes;
    return true;
}

// Cross-part barrier over peer memory (acc_comm_enable_peer_stores): thread p announces this part's epoch in part p's counter
// array and waits until part p has announced the same epoch here.  The fence orders the peer stores of the kernels before it
// (same stream) ahead of the announcement; the waiting side's later kernels see them.  A wait that lasts seconds (a part died)
// ends with an error flag instead of hanging the GPU.
__global__ void cross_barrier_kernel(PeerTable pt, int rank, int world, unsigned long long epoch, int* timed_out) {
    const int p = (int)threadIdx.x;
    if (p >= world || p == rank) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(pt.flags[p] + rank), "l"(epoch) : "memory");
    const unsigned long long* mine = pt.flags[rank] + p;
    const long long t0 = clock64();
    for (;;) {
        unsigned long long seen;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(mine) : "memory");
        if (seen >= epoch) break;
        if (clock64() - t0 > 20000000000LL) { *timed_out = 1; break; }  // ~10 s
    }
}

template <class K>
__global__ void __launch_bounds__(128) thread_kernel(int n, K k) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) k(i);
}
template <class K>
__global__ void __launch_bounds__(K::THREADS, K::MIN_BLOCKS) block_kernel(K k) {
    __shared__ __align__(16) unsigned char scratch[ACC_BLOCK_SCRATCH_BYTES];
    BlockCtx cx(scratch);
    k(cx, (int)blockIdx.x);
}
// One CTA per seller; `cap` > 0: the seller's demand book is staged in dynamic shared memory first (csrc/stage.cuh),
// books larger than `cap` connections are read in place.
// `ahead` > 0: while this CTA waits for its own book, one thread asks the copy engine to bring the book of the seller `ahead`
// CTAs further on -- the one a CTA of the next wave will take on this SM -- into the L2, so that its bulk copy is fed from there.
template <class K>
__global__ void __launch_bounds__(K::THREADS, K::MIN_BLOCKS) seller_kernel(K k, int cap, int ahead, int perm_slots) {
    __shared__ __align__(16) unsigned char scratch[ACC_BLOCK_SCRATCH_BYTES];
    __shared__ __align__(8) unsigned long long bar;
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    BlockCtx cx(scratch);
    SellerBook book{nullptr, nullptr, nullptr, nullptr, 0};
    if (cap > 0) {
        if (threadIdx.x == 0) mbar_init(&bar, 1);
        __syncthreads();
    }
    // k.prepare: the seller's possible production, one input storage per lane of the second warp.  Its chain of dependent
    // loads (firm -> storage list -> storages) starts at once and runs beside the other warps' work: the seller's range,
    // handing the book to the copy engine and waiting for it.  The second warp joins them afterwards.
    if ((threadIdx.x >> 5) == 1) { WarpCtx warp; k.prepare(warp, (int)blockIdx.x); }
    int e0, n;
    k.range((int)blockIdx.x, &e0, &n);
    if (ahead > 0 && threadIdx.x == 64 && blockIdx.x + ahead < gridDim.x) {
        int e2, n2;
        k.range((int)blockIdx.x + ahead, &e2, &n2);
        if (n2 > 0 && n2 <= cap) {
            prefetch_l2(k.ev.D_q + e2, (size_t)n2 * sizeof(i64));
            prefetch_l2(k.ev.D_p + e2, (size_t)n2 * sizeof(i64));
            prefetch_l2(k.ev.slotw + e2, (size_t)n2 * sizeof(unsigned));
            prefetch_l2(k.ev.domestic + e2, (size_t)n2);
        }
    }
    if (n > 0 && n <= cap) book = stage_seller_book(dyn_smem, cap, &bar, k.ev.D_q, k.ev.D_p, k.ev.domestic, e0, n);
    if (perm_slots > 0) {  // room for a sort of the seller's connections behind the book (ordered_range_sums)
        book.perm16 = reinterpret_cast<unsigned short*>(dyn_smem + (cap > 0 ? seller_book_bytes(cap) : 0));
        book.perm_slots = perm_slots;
    }
    __syncthreads();
    k(cx, (int)blockIdx.x, book);
}
// G > 0: the CTA size is a compile-time constant (K::MAXREG_FIXED registers); G == 0: any CTA size (K::MAXREG)
template <class K, int G>
__global__ void __maxnreg__(G > 0 ? K::MAXREG_FIXED : K::MAXREG) block_smem_kernel(K k) {
    __shared__ __align__(16) unsigned char scratch[G > 0 ? ACC_BLOCK_SCRATCH_BYTES_FOR(G) : ACC_BLOCK_SCRATCH_BYTES];
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    BlockCtxT<G> cx(scratch);
    k(cx, (int)blockIdx.x, dyn_smem);
}
template <class K>
__global__ void __launch_bounds__(32) warp_kernel(K k) {
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    WarpCtx cx;
    k(cx, (int)blockIdx.x, dyn_smem);
}
// K::WARPS items per CTA, one warp each, every warp with its own slice of the dynamic shared memory
template <class K>
__global__ void __launch_bounds__(K::WARPS * 32) warps_kernel(K k, int n, unsigned smem_per_warp) {
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    const int w = threadIdx.x >> 5;
    const int item = (int)blockIdx.x * K::WARPS + w;
    if (item >= n) return;  // whole warps leave; the contexts of the others only use warp-level synchronisation
    WarpCtx cx;
    k(cx, item, dyn_smem + (size_t)w * smem_per_warp);
}
template <class K>
static void launch_warps(stream_t s, int n, size_t smem_per_warp, K k) {
    if (n <= 0) return;
    smem_per_warp = (smem_per_warp + 15) / 16 * 16;
    const size_t smem = smem_per_warp * K::WARPS;
    if (smem_needs_config((const void*)warps_kernel<K>, smem)) ACC_CUDA(cudaFuncSetAttribute(warps_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    warps_kernel<K><<<(n + K::WARPS - 1) / K::WARPS, K::WARPS * 32, smem, s>>>(k, n, (unsigned)smem_per_warp);
    ACC_CUDA(cudaGetLastError());
}
template <class K>
static void launch_thread(stream_t s, int n, K k) {
    if (n <= 0) return;
    thread_kernel<K><<<(n + 127) / 128, 128, 0, s>>>(n, k);
    ACC_CUDA(cudaGetLastError());
}
template <class K>
static void launch_block(stream_t s, int n, K k) {
    if (n <= 0) return;
    block_kernel<K><<<n, K::THREADS, 0, s>>>(k);
    ACC_CUDA(cudaGetLastError());
}
template <class K>
static void launch_seller(stream_t s, int n, int cap, int perm_slots, K k) {
    if (n <= 0) return;
    const size_t smem = (cap > 0 ? seller_book_bytes(cap) : 0) + (size_t)perm_slots * sizeof(unsigned short);
    if (smem_needs_config((const void*)seller_kernel<K>, smem)) ACC_CUDA(cudaFuncSetAttribute(seller_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int ahead = 0;
    if (cap > 0) {  // half a wave of CTAs ahead (measured on the EORA shape: none 0.521 ms, 148 ahead 0.504, 296 ahead 0.512, 592 ahead 0.539)
        static std::map<int, int> sms;
        int dev = 0;
        ACC_CUDA(cudaGetDevice(&dev));
        if (sms.find(dev) == sms.end()) { int v = 0; ACC_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev)); sms[dev] = v; }
        ahead = sms[dev] * K::MIN_BLOCKS / 2;
#ifdef ACC_SELLER_PREFETCH_AHEAD
        ahead = ACC_SELLER_PREFETCH_AHEAD;
#endif
    }
    seller_kernel<K><<<n, K::THREADS, smem, s>>>(k, cap, ahead, perm_slots);
    ACC_CUDA(cudaGetLastError());
}
template <class K>
static void launch_warp(stream_t s, int n, size_t smem, K k) {
    if (n <= 0) return;
    if (smem_needs_config((const void*)warp_kernel<K>, smem)) ACC_CUDA(cudaFuncSetAttribute(warp_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    warp_kernel<K><<<n, 32, smem, s>>>(k);
    ACC_CUDA(cudaGetLastError());
}
template <int G, class K>
static void launch_block_smem(stream_t s, int n, int threads, size_t smem, K k) {
    if (n <= 0) return;
    if (G > 0 && threads != G) fail("launch_block_smem: CTA size does not match the compiled group size");
    if (smem_needs_config((const void*)block_smem_kernel<K, G>, smem)) {
        ACC_CUDA(cudaFuncSetAttribute(block_smem_kernel<K, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ACC_CUDA(cudaFuncSetAttribute(block_smem_kernel<K, G>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    }
    block_smem_kernel<K, G><<<n, threads, smem, s>>>(k);
    ACC_CUDA(cudaGetLastError());
}
#endif

#ifndef ACC_HOST_EMU
// ---- NCCL, bound at run time ----
// The halo exchanges of a region-partitioned run are NCCL point-to-point groups and one all-gather issued by the engine
// itself on its own stream (acc_comm_init / acc_step_partitioned).  The library is looked up when the first communicator is
// made: the copy the process has already loaded (a host that also uses torch.distributed shares torch's), else the system
// one; an unpartitioned run never needs it.  Prototypes as in nccl.h (2.x: stable since 2.7).
struct NcclApi {
    typedef struct ncclComm* comm_t;
    struct unique_id { char internal[128]; };
    int (*GetUniqueId)(unique_id*) = nullptr;
    int (*CommInitRank)(comm_t*, int, unique_id, int) = nullptr;
    int (*CommDestroy)(comm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, comm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    enum { Int64 = 4 };
    static NcclApi& get() {
        static NcclApi api;
        if (api.Send != nullptr) return api;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (h == nullptr) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (h == nullptr) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (h == nullptr) fail(std::string("NCCL is not available: ") + dlerror());
        auto sym = [&](const char* name) {
            void* p = dlsym(h, name);
            if (p == nullptr) fail(std::string("NCCL symbol missing: ") + name);
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
        return api;
    }
};
#define ACC_NCCL(x)                                                                                                  \
    do {                                                                                                             \
        const int err__ = (x);                                                                                       \
        if (err__ != 0) fail(std::string(#x) + ": " + NcclApi::get().GetErrorString(err__));                         \
    } while (0)
#endif

template <class T>
struct DBuf {  // device array with its element count
    T* p = nullptr;
    size_t n = 0;
    bool borrowed = false;  // the array belongs to another engine (the immutable tables of an ensemble member, acc_fork)
    void alloc(size_t count) { free(); n = count; p = static_cast<T*>(dev_alloc(count * sizeof(T))); }
    void upload(const std::vector<T>& h, stream_t s) { alloc(h.size()); h2d(p, h.data(), h.size() * sizeof(T), s); }
    void zero(stream_t s) { dev_zero(p, n * sizeof(T), s); }
    std::vector<T> download(stream_t s) const { std::vector<T> h(n); d2h(h.data(), p, n * sizeof(T), s); return h; }
    void free() { if (p && !borrowed) dev_free(p); p = nullptr; n = 0; borrowed = false; }
};

// ------------------------------------------------------------------------------------------------
// kernels (functors over the routines of model_core.cuh / purchase_core.cuh)
// ------------------------------------------------------------------------------------------------
struct SellerProduction {  // CTA per firm
#ifndef ACC_SELLER_THREADS
#define ACC_SELLER_THREADS 256
#define ACC_SELLER_MIN_BLOCKS 2
#endif
#ifndef ACC_SELLER_BOOK_BUDGET
#define ACC_SELLER_BOOK_BUDGET (108 * 1024)  // bytes of dynamic shared memory per CTA: two CTAs per SM
#endif
    enum { THREADS = ACC_SELLER_THREADS, MIN_BLOCKS = ACC_SELLER_MIN_BLOCKS };
    FirmView fv; EdgeView ev; StorageView st; ModelP mp; const SectorP* sectors; EventSink events; i64 step; const int* firm_ids;
    // CapacityManager::get_possible_production_X_hat (CapacityManager.cpp:72-102) by one warp, before the sales manager's pass
    template <class Ctx>
    ACC_HD void prepare(Ctx& warp, int k) const {
        const int f = firm_ids[k];
        const int s0 = fv.st_off[f];
        const Flow cap = possible_production(warp, st, fv.st_list + s0, fv.st_off[f + 1] - s0, mkflow(fv.Xstar_q[f], fv.Xstar_v[f]), fv.lambda[f],
                                             fv.beta[f], fv.var_costs[f], mp.dt, false);
        if (warp.leader()) { fv.cap_q[f] = cap.q; fv.cap_v[f] = cap.v; }
    }
    ACC_HD void range(int k, int* e0, int* n) const {  // the seller's connections (none for a consumer)
        const int f = firm_ids[k];
        *e0 = ev.seller_off[f];
        *n = fv.type[f] == 1 ? ev.seller_off[f + 1] - *e0 : 0;
    }
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int k, const SellerBook& book) const { seller_phase1(cx, firm_ids[k], fv, ev, mp, sectors, events, step, book); }
};

struct SellerExpectation {  // warp per firm
    FirmView fv; EdgeView ev; StorageView st; ModelP mp; const SectorP* sectors; EventSink events; const int* firm_ids;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int k, unsigned char*) const { seller_phase2(cx, firm_ids[k], fv, ev, st, mp, sectors, events); }
};

template <bool FORCED>  // FORCED: passage forcing has been used, the chain of hops with its overflows is walked
struct StorageDelivery {  // thread per storage
    StorageView st; FirmView fv; EdgeView ev; ModelP mp; const SectorP* sectors; EventSink events; i64 step;
    const i64* ell_off; const int* ell_width; const int* ell_edge; const unsigned* ell_slotw; const unsigned char* owned;  // owned == nullptr: every agent
    ACC_HD void operator()(int s) const {
        const int a = st.agent[s];
        if (owned != nullptr && !owned[a]) return;
        const SectorP sp = sectors[st.sector[s]];
        // ---- input use and storage content ----
        const Flow uhat = possible_use(st, s, mp.dt, false);
        Flow used;
        if (fv.type[a] == 1) {  // Firm::iterate_consumption_and_production, Firm.cpp:54-61
            const Flow X = mkflow(fv.X_q[a], fv.X_v[a]);
            used = fcopy(fmul(X, st.a[s]));
            if (X.q > 0) used = set_price(used, flow_price(uhat));
        } else if (mp.utilitarian) {  // the consumer kernel has set the desired use to the optimised consumption, Consumer.cpp:510-517
            used = mkflow(st.Ut_q[s], st.Ut_v[s]);
        } else {  // price-elastic consumer, Consumer.cpp:276-302
            i64 rp;
            if (uhat.q > 0) {
                rp = flow_price(uhat);
            } else {
                rp = flow_price(mkflow(st.Ut_q[s], st.Ut_v[s]));
                events.push(EV_NO_CONSUMPTION, a, -1, pf(rp));
            }
            const i64 want = qmul(qmul(st.Istar_q[s], fv.lambda[a]), pow(ratio(rp, 1000000), st.a[s]));
            const Flow ut = flow_qp(want, rp);
            used = fcopy(flow_qp(want < uhat.q ? want : uhat.q, rp));
            st.Ut_q[s] = ut.q; st.Ut_v[s] = fcopy(ut).v;
        }
        storage_update(st, s, used, mp, sp.omega, events);
        // ---- arrivals of this step: TransportChainLink::push_flow_Z (TransportChainLink.cpp:65-97),
        //      BusinessConnection::deliver_flow_Z (BusinessConnection.cpp:106-113), Storage::push_flow_Z ----
        const int slice = s >> 5, lane = s & 31;
        const i64 base = ell_off[slice];
        const int width = ell_width[slice];
        const ConveyorRings& rg = ev.rings;
        Flow in = mkflow(0, 0.0);
        Flow imp = mkflow(0, 0.0);
        i64 tq = st.T_q[s];
        double tv = st.T_v[s];
        // The fold below is strictly sequential (the reference's delivery order: sellers ascending), so the loads of
        // eight connections are issued together first: one thread keeps 8 x 5 independent requests in flight instead
        // of walking a chain of dependent ones.
        enum { U = 8 };
        for (int k0 = 0; k0 < width; k0 += U) {
            int e_[U], tau_[U];
            unsigned w_[U];
            Flow front_[U], ship_[U];
            unsigned char dom_[U];
#pragma unroll
            for (int j = 0; j < U; ++j) {
                const bool have = k0 + j < width;
                e_[j] = have ? ell_edge[base + (i64)(k0 + j) * 32 + lane] : -1;
                w_[j] = ell_slotw[base + (i64)(have ? k0 + j : 0) * 32 + lane];  // padding cells hold connection -1 and the word of connection 0
            }
#pragma unroll
            for (int j = 0; j < U; ++j) {
                const int e = e_[j] < 0 ? 0 : e_[j];
                const unsigned w = w_[j];
                tau_[j] = ring_tau(rg, w);
                dom_[j] = ev.domestic[e];
                const i64 at = ring_slot_front(rg, w);
                front_[j] = mkflow(ev.conv_q[at], ev.conv_v[at]);
            }
#pragma unroll
            for (int j = 0; j < U; ++j) {
                ship_[j] = front_[j];
                if (tau_[j] != 0) {
                    const i64 at = ring_slot_last(rg, w_[j]);
                    ship_[j] = mkflow(ev.conv_q[at], ev.conv_v[at]);
                }
            }
#pragma unroll
            for (int j = 0; j < U; ++j) {
                const int e = e_[j];
                if (e < 0) continue;
                const Flow front = front_[j];
                const Flow ship = ship_[j];
                Flow out;
                const int rt = FORCED ? ev.route[e] : -1;
                if (rt < 0) {
                    out = fcopy(front);  // no passage forcing was ever set (overflow == 0), or no geo entity on the way
                } else {
                    // the chain of links, TransportChainLink.cpp:65-97: at this step hop h releases the shipment that
                    // still has rem_h steps to go, i.e. the conveyor slot (step + rem_h) mod T; what it lets pass goes
                    // back into that slot for the next hop (zero-delay hops follow at once), the last hop delivers
                    int rem = tau_[j];
                    const int h0 = ev.route_off[rt], h1 = ev.route_off[rt + 1];
                    out = mkflow(0, 0.0);
                    for (int h = h0; h < h1; ++h) {
                        const int ent = ev.route_ent[h];
                        rem -= ev.ent_delay[ent];
                        const i64 slot = ring_slot(rg, w_[j], rem);
                        const i64 oslot = (i64)(h - h0) * ev.n_edges + e;
                        const Flow fr = mkflow(ev.conv_q[slot], ev.conv_v[slot]);
                        const double nu = ev.nu[ent];
                        const Flow total = fadd(mkflow(ev.ovf_q[oslot], ev.ovf_v[oslot]), fr);
                        Flow o;
                        if (nu < 0) {
                            o = fcopy(total);
                        } else {
                            const Flow cap = flow_qp(qmul_l(nu, ev.Z_q[e]), flow_price(fr));
                            o = fcopy(cap.q < total.q ? cap : total);  // std::min by quantity
                        }
                        const Flow ovf = fcopy(fsub(total, o));
                        ev.ovf_q[oslot] = ovf.q; ev.ovf_v[oslot] = ovf.v;
                        // what passes goes back into the slot: for the next hop, or -- last hop -- as the connection's last delivery
                        // (the arrival slot is dead until its plane comes round again; read by the output surface)
                        ev.conv_q[slot] = o.q; ev.conv_v[slot] = o.v;
                        if (h + 1 == h1) out = o;
                    }
                }
                if (FORCED && ev.ship_q != nullptr) { ev.ship_q[e] = ship.q; ev.ship_v[e] = ship.v; }
                in = fadd(in, out);
                if (!dom_[j]) imp = fadd(imp, out);
                tq += ship.q - out.q;
                tv += ship.v - out.v;
            }
        }
        st.In_q[s] = in.q; st.In_v[s] = in.v;
        st.dlv_q[s] = in.q;
        st.imp_q[s] = imp.q; st.imp_v[s] = imp.v;
        st.T_q[s] = tq; st.T_v[s] = tv;
    }
};

// One warp per utilitarian consumer, eight consumers per CTA: the solve is sequential code run by the warp's leader, so
// the kernel needs residency, not width -- packed like this it takes two dozen SMs instead of one CTA slot on every SM
// next to seller_production, which runs at the same time.
#ifndef ACC_CONSUMER_WARPS
#define ACC_CONSUMER_WARPS 8
#endif
struct ConsumerUtility {
    enum { WARPS = ACC_CONSUMER_WARPS };
    ConsumerView cv; StorageView st; ModelP mp; ConsumerStage out; int n_max;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int item, unsigned char* smem) const { consumer_one(cx, cv.cons_ids[item], cv, st, mp, out, smem, n_max); }
};
struct ConsumerApply {  // warp per consumer: the solve's results become the step's desired use
    enum { WARPS = 8 };
    ConsumerView cv; StorageView st; ConsumerStage in; EventSink events; int* fatal;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int item, unsigned char*) const {
        if (!consumer_apply(cx, cv.cons_ids[item], cv, st, in, events) && cx.leader()) *fatal = 1;
    }
};

// One CTA of pm_group threads per purchasing manager.  SEP: deviation_penalty (separate penalty target); STD: the
// reference's default transport penalty (quadratic, relative) folded at compile time.
template <bool SEP, bool STD, bool DVAL>
struct Purchase {
    // any group size: 512 threads still fit; base group (96 threads, compile-time): 96 registers = 6 CTAs of 3 warps per
    // SM (5 warps on a scheduler: 16384 / (5 * 32) = 102).  80 registers would admit the 7 CTAs the shared memory allows
    // but spill: measured 3.22 ms against 3.10 ms on the EORA shape
    // (a compile-time LARGE group -- the 384 threads of the 748-connection managers of the 4x network -- keeps the 80 registers
    // that let two of its CTAs share an SM; it is told apart by the dropped D_r vector)
    enum { MAXREG = ACC_PM_MAXREG_ANY, MAXREG_FIXED = DVAL ? ACC_PM_MAXREG_FIXED : ACC_PM_MAXREG_ANY };
    StorageView st; EdgeView ev; ModelP mp; const SectorP* sectors; EventSink events; i64 step;
    // incoming connections per storage (seller ascending): connection id, seller | delay << 24, Z* (copies in this order:
    // one coalesced read instead of a gather through the connection id)
    // in_seller: seller ROW (SellerTable) | delay << 24
    const i64* in_off; const int* in_edge; const int* in_seller; const unsigned* in_slotw; const i64* in_Z; SellerTable pubT; int n_max; const int* pm_ids;  // pm_ids == nullptr: all
    i64* dem_q; double* dem_v; double* exp_costs; double* opt_val; double* tpen; int* iters; int* evals;
    int n_items, ahead;  // ahead > 0: the last thread asks for the connection lists of the manager `ahead` items further on (L2 prefetch)
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int item, unsigned char* smem) const {
#if defined(__CUDA_ARCH__)
        if (ahead > 0 && threadIdx.x == blockDim.x - 1 && item + ahead < n_items) {
            // the three static lists of a manager further down the grid are streamed from HBM once per step: by the time a CTA takes
            // that manager they wait in the L2
            const int s2 = pm_ids != nullptr ? pm_ids[item + ahead] : item + ahead;
            const i64 o2 = in_off[s2];
            const size_t n2 = (size_t)(in_off[s2 + 1] - o2);
            prefetch_l2(in_edge + o2, n2 * sizeof(int));
            prefetch_l2(in_seller + o2, n2 * sizeof(int));
            prefetch_l2(in_slotw + o2, n2 * sizeof(unsigned));
            prefetch_l2(in_Z + o2, n2 * sizeof(i64));
        }
#endif
        const int s = pm_ids != nullptr ? pm_ids[item] : item;
        PMWorkT<SEP, STD, DVAL> k;
        k.carve(smem, n_max);
        const i64 o = in_off[s];
        const int nin = (int)(in_off[s + 1] - o);
        const int* ie = in_edge + o;
        const int* is = in_seller + o;
        const unsigned* iw = in_slotw + o;
        const i64* iz = in_Z + o;
        const EdgeView& E = ev;
        const double* const tf = pubT.f;
        const i64 ts = pubT.stride;
        const i64 deficit = st.defc[s] - st.dlv_q[s] - st.T_q[s];
        const PMResult r = purchase_one<SEP, STD>(
            cx, nin, [&](int i) { return ie[i]; },
            [&](int i) {
                PMGather g;
                g.e = ie[i];
                const unsigned st_ = (unsigned)is[i];
                g.seller = (int)(st_ & 0xffffffu);
                const int e = g.e;
                // the last shipment: the slot of time step + delay in the connection's conveyor ring
                const i64 at = E.ship_q != nullptr ? (i64)e : ring_slot_last(E.rings, iw[i]);
                g.last_q = (E.ship_q != nullptr ? E.ship_q : E.conv_q)[at];  // a select, not a branch: loads of two gathers overlap
                g.Z_q = iz[i];
                g.lastD_q = SEP ? E.D_q[e] : 0;
                const double* f = tf + g.seller;  // consecutive connections -> consecutive rows: coalesced field by field
                double xq = f[SPT_XHAT_Q * ts];
                memcpy(&g.sp.Xhat_q, &xq, sizeof xq);
                g.sp.Xhatf = f[SPT_XHATF * ts]; g.sp.Xf = f[SPT_XF * ts]; g.sp.Xexpf = f[SPT_XEXPF * ts]; g.sp.nbarf = f[SPT_NBARF * ts];
                g.sp.lamXf = f[SPT_LAMXF * ts]; g.sp.c1 = f[SPT_C1 * ts]; g.sp.c2 = f[SPT_C2 * ts]; g.sp.n_co = f[SPT_NCO * ts];
                g.rr = f[SPT_RR * ts]; g.npeE = f[SPT_NPEE * ts];
                return g;
            },
            [&](int e) {  // BusinessConnection::get_minimum_passage, BusinessConnection.cpp:69-83
                double m = 1.0;
                if (E.nu != nullptr && E.route[e] >= 0) {
                    for (int h = E.route_off[E.route[e]]; h < E.route_off[E.route[e] + 1]; ++h) {
                        const double nu = E.nu[E.route_ent[h]];
                        if (nu >= 0.0 && nu < m) m = nu;
                    }
                }
                return m;
            },
            [&](int e, i64 q, i64 p) { E.D_q[e] = q; E.D_p[e] = p; peer_store_demand(E, e, q, p); }, k, mp, sectors[st.sector[s]], st.Istar_q[s], st.Sstar_q[s],
            st.S_q[s], st.Ut_q[s], deficit, events, st.sector[s], st.agent[s]);
        if (cx.leader()) {
            dem_q[s] = r.demand.q; dem_v[s] = r.demand.v; exp_costs[s] = r.expected_costs; opt_val[s] = r.optimized; tpen[s] = r.penalty;
            iters[s] = r.iters; evals[s] = r.evals;
        }
    }
};

struct FirmIncomingDemand {  // CTA per firm: sum of the demand requests it holds (output surface only)
    enum { THREADS = 256, MIN_BLOCKS = 1 };
    EdgeView ev; const int* firm_ids; i64* out_q; double* out_v;
    template <class Ctx>
    ACC_HD void operator()(Ctx& cx, int k) const {
        const int f = firm_ids[k];
        const int e0 = ev.seller_off[f];
        const BookTotals t = book_totals(cx, ev.seller_off[f + 1] - e0, ev.D_q + e0, ev.D_p + e0);
        if (cx.leader()) { out_q[f] = t.sq; out_v[f] = t.sv; }
    }
};

// Per-connection observables of the output surface (thread per connection), all derived from the conveyor:
//   last shipment   BusinessConnection::last_shipment_Z   plane (t + delay) mod T, t = step of the last production pass
//   last delivery   BusinessConnection::last_delivery_Z   plane t mod T (what arrived during step t stays there until the
//                                                         plane comes round again; a forced chain's last hop writes what
//                                                         it released back into that slot)
//   in transit      BusinessConnection::get_transport_flow (src/model/BusinessConnection.cpp:162-171): the delay slots
//                   still under way, planes t+1 .. t+delay, plus the overflow of every hop
//   flow deficit    get_flow_deficit (:146-155) = Z* (delay + 1) - delivered - in transit
// idx == nullptr: connections 0..n-1, else the selected ones.  Null outputs are skipped.
struct EdgeObservables {
    EdgeView ev; i64 t; const i64* idx;
    i64* ship_q; double* ship_v; i64* dlv_q; double* dlv_v; i64* tr_q; double* tr_v; i64* def_q;
    ACC_HD void operator()(int k) const {
        const i64 e = idx != nullptr ? idx[k] : (i64)k;
        const int tau = ev.tau[e];
        const unsigned w = ev.slotw[e];
        const i64 at0 = ring_slot_time(ev.rings, w, t);
        if (ship_q != nullptr || ship_v != nullptr) {
            const i64 at = ring_slot_time(ev.rings, w, t + tau);
            if (ship_q != nullptr) ship_q[k] = ev.ship_q != nullptr ? ev.ship_q[e] : ev.conv_q[at];
            if (ship_v != nullptr) ship_v[k] = ev.ship_v != nullptr ? ev.ship_v[e] : ev.conv_v[at];
        }
        if (dlv_q != nullptr || dlv_v != nullptr) {
            if (dlv_q != nullptr) dlv_q[k] = ev.conv_q[at0];
            if (dlv_v != nullptr) dlv_v[k] = ev.conv_v[at0];
        }
        if (tr_q != nullptr || tr_v != nullptr || def_q != nullptr) {
            Flow sum = mkflow(0, 0.0);
            for (int j = 1; j <= tau; ++j) {
                const i64 at = ring_slot_time(ev.rings, w, t + j);
                sum = fadd(sum, mkflow(ev.conv_q[at], ev.conv_v[at]));
            }
            if (ev.ovf_q != nullptr && ev.route[e] >= 0) {
                const int nh = ev.route_off[ev.route[e] + 1] - ev.route_off[ev.route[e]];
                for (int h = 0; h < nh; ++h) sum = fadd(sum, mkflow(ev.ovf_q[(i64)h * ev.n_edges + e], ev.ovf_v[(i64)h * ev.n_edges + e]));
            }
            sum = fcopy(sum);
            if (tr_q != nullptr) tr_q[k] = sum.q;
            if (tr_v != nullptr) tr_v[k] = sum.v;
            if (def_q != nullptr) def_q[k] = ev.Z_q[e] * (tau + 1) - ev.conv_q[at0] - sum.q;
        }
    }
};

// ---- exchange of a region-partitioned run (thread per halo entry / per agent) ----
ACC_HD i64 f64_bits(double v) { i64 b; memcpy(&b, &v, sizeof b); return b; }
ACC_HD double bits_f64(i64 b) { double v; memcpy(&v, &b, sizeof v); return v; }
struct PackFlow {  // shipment of this step: the conveyor slot it was written to
    EdgeView ev; const i64* ids; i64* buf;
    ACC_HD void operator()(int i) const {
        const i64 e = ids[i];
        const i64 at = ring_slot_last(ev.rings, ev.slotw[e]);
        buf[2 * (i64)i] = ev.conv_q[at]; buf[2 * (i64)i + 1] = f64_bits(ev.conv_v[at]);
    }
};
struct UnpackFlow {
    EdgeView ev; const i64* ids; const i64* buf;
    ACC_HD void operator()(int i) const {
        const i64 e = ids[i];
        const i64 at = ring_slot_last(ev.rings, ev.slotw[e]);
        ev.conv_q[at] = buf[2 * (i64)i]; ev.conv_v[at] = bits_f64(buf[2 * (i64)i + 1]);
    }
};
struct PackDemand {
    EdgeView ev; const i64* ids; i64* buf;
    ACC_HD void operator()(int i) const { const i64 e = ids[i]; buf[2 * (i64)i] = ev.D_q[e]; buf[2 * (i64)i + 1] = ev.D_p[e]; }
};
struct UnpackDemand {
    EdgeView ev; const i64* ids; const i64* buf;
    ACC_HD void operator()(int i) const { const i64 e = ids[i]; ev.D_q[e] = buf[2 * (i64)i]; ev.D_p[e] = buf[2 * (i64)i + 1]; }
};
static_assert(sizeof(SellerPub) == ACC_SELLER_WORDS * sizeof(i64), "seller table row size is part of the C ABI");
struct PackSellers {  // owned rows as they are, zero elsewhere: an integer SUM over the ranks then copies every row from its owner
    const SellerPub* pub; const unsigned char* owned; i64* table;
    ACC_HD void operator()(int a) const {
        const i64* src = reinterpret_cast<const i64*>(pub + a);
        for (int k = 0; k < ACC_SELLER_WORDS; ++k) table[(i64)a * ACC_SELLER_WORDS + k] = owned[a] ? src[k] : 0;
    }
};
struct UnpackSellers {
    SellerPub* pub; SellerTable pubT; const unsigned char* owned; const i64* table;
    ACC_HD void operator()(int a) const {
        if (owned[a]) return;
        i64* dst = reinterpret_cast<i64*>(pub + a);
        for (int k = 0; k < ACC_SELLER_WORDS; ++k) dst[k] = table[(i64)a * ACC_SELLER_WORDS + k];
        publish_seller(pubT, a, pub[a]);
    }
};
// the NCCL plane gathers only the rows their owners computed: owned firms in agent order, padded to the largest part
struct PackSellerRows {
    const SellerPub* pub; const int* firms; i64* rows;
    ACC_HD void operator()(int k) const {
        const i64* src = reinterpret_cast<const i64*>(pub + firms[k]);
        for (int w = 0; w < ACC_SELLER_WORDS; ++w) rows[(i64)k * ACC_SELLER_WORDS + w] = src[w];
    }
};
struct UnpackSellerRows {  // thread per (part, row): firms of every part in agent order; the own part's rows are skipped
    SellerPub* pub; SellerTable pubT; const int* part_firms; const int* part_off; int n_parts, rank, max_rows; const i64* rows;
    ACC_HD void operator()(int i) const {
        const int p = i / max_rows, k = i - p * max_rows;
        if (p == rank || k >= part_off[p + 1] - part_off[p]) return;
        const int a = part_firms[part_off[p] + k];
        i64* dst = reinterpret_cast<i64*>(pub + a);
        for (int w = 0; w < ACC_SELLER_WORDS; ++w) dst[w] = rows[((i64)p * max_rows + k) * ACC_SELLER_WORDS + w];
        publish_seller(pubT, a, pub[a]);
    }
};
struct RepublishSellers {  // the field-major twin of the seller records, rebuilt from them (after a checkpoint was loaded)
    const SellerPub* pub; SellerTable pubT;
    ACC_HD void operator()(int a) const { publish_seller(pubT, a, pub[a]); }
};

// ------------------------------------------------------------------------------------------------
// the engine
// ------------------------------------------------------------------------------------------------
struct KernelClock {
    const char* name;
    double ms = 0.0;
    i64 launches = 0;
#ifndef ACC_HOST_EMU
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
#endif
};

}  // namespace acc

using namespace acc;

struct acc_engine {
    int device = 0;
    stream_t stream = nullptr;
    stream_t side = nullptr;  // consumers' utility optimisation, concurrent with the sellers' pass
#ifndef ACC_HOST_EMU
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
#endif
    i64 step = 0;
    i64 ship_step = -1;  // step of the last production / distribution pass (its shipments sit on plane (ship_step + tau) mod T)
    i64 launches = 0;
    bool profile = false;
    std::vector<KernelClock> clocks;

    // sizes
    int nA = 0, nS = 0, nR = 0, nSt = 0, nF = 0, n_max = 0, T = 1;
    int pm_group = ACCSQP_GROUP;  // threads per purchasing manager
    int book_cap = 0;             // largest seller's number of connections when its demand book fits the shared-memory budget, else 0
    int perm_slots = 0;           // 16-bit sort slots behind the book (power of two >= the largest seller) when they fit as well, else 0
    i64 nE = 0;
    ModelP mp;
    std::vector<SectorP> h_sectors;
    // host copies of topology needed for reads
    std::vector<int> h_agent_type, h_agent_sector, h_agent_region, h_st_agent, h_st_sector, h_cid2s, h_firm_ids;
    std::vector<unsigned char> h_tau;
    std::vector<int> h_consumer_st;  // storages of consumers in (agent, container) order

    DBuf<SectorP> sectors;
    // agents / firms
    DBuf<int> a_type, a_sector, a_region, a_st_off, a_st_list, firm_ids;
    DBuf<double> a_beta, a_lambda;
    DBuf<i64> f_Xstar_q, f_markup_eff, f_var_costs;
    DBuf<double> f_Xstar_v;
    DBuf<i64> f_cap_q, f_Xhat_q, f_est_q, f_X_q, f_Xexp_q, f_nbar, f_cutoff, f_desired_q, f_sumD_q, f_bk_sq, f_bk_cnt, f_bk_pmax, f_bk_pmin, f_exp_q;
    DBuf<double> f_cap_v, f_Xhat_v, f_est_v, f_X_v, f_Xexp_v, f_C, f_R, f_sumD_v, f_bk_sv, f_exp_v;
    DBuf<SellerPub> pub;
    DBuf<double> pubT;         // field-major twin of pub over the seller rows (SellerTable)
    DBuf<int> seller_row;      // agent -> seller row (firms in sector-major order), -1 for consumers
    DBuf<int> row_firm;        // seller row -> agent
    // output surface (acc_observe): buyer of every connection, storage of every (sector, agent), staging buffers
    DBuf<int> e_to, sto_of;
    DBuf<double> obs_out;
    DBuf<i64> obs_idx, obs_inc_q;
    DBuf<double> obs_inc_v;
    int max_storages = 0;      // most input storages of one agent
    std::vector<int> h_row_firm;  // seller row -> agent
    // storages
    DBuf<i64> s_S_q, s_I0_q, s_I1_q, s_Il_q, s_U_q, s_Ut_q, s_T_q, s_dlv_q, s_imp_q, s_Istar_q, s_Sstar_q, s_defc, s_dem_q;
    DBuf<double> s_S_v, s_I0_v, s_I1_v, s_Il_v, s_U_v, s_Ut_v, s_T_v, s_imp_v, s_Istar_v, s_Sstar_v, s_a, s_dem_v, s_exp_costs, s_opt_val, s_tpen;
    DBuf<int> s_agent, s_sector, s_iters, s_evals;
    int reg = 0;  // which of I0/I1 is the "other" register (arrived during the previous step)
    // connections
    DBuf<int> e_seller_off, e_route, route_off, route_ent;
    DBuf<unsigned char> ent_delay;
    DBuf<i64> e_ship_q;
    DBuf<double> e_ship_v;
    int nEnt = 0, h_max = 1;
    bool hops_rewrite_shipment = false;  // some route starts with a zero-delay entity
    DBuf<i64> e_D_q, e_D_p, e_Z_q, conv_q, e_ovf_q;
    DBuf<double> conv_v, e_ovf_v, nu;
    DBuf<unsigned char> e_tau, e_dom;
    bool forcing_used = false;
    // buyer-side lists
    DBuf<i64> ell_off, in_off;
    DBuf<int> ell_width, ell_edge, in_edge, in_seller;
    DBuf<int> f_sorted;  // per firm: the last sellers' pass took the pro-rata sums in sorted order (diagnostic, read_state "firm_sorted_sums")
    DBuf<int> e_sort;  // scratch of the sellers' pass: a seller's connections in sorted order (model_core.cuh ordered_range_sums)
    DBuf<unsigned> e_slotw, ell_slotw, in_slotw;  // conveyor words (ring << 24 | rank) in connection order and in both buyer-side orders
    ConveyorRings rings;                          // groups of equal delay (phase filled in per view)
    int n_rings = 0;
    i64 conv_slots = 0;
    DBuf<i64> in_Z;
    // utilitarian consumers (nested-CES utility optimisation)
    int nC = 0, nB = 0, c_n_max = 0;
    double c_inter_exp = 0.0, c_ctol = 0.0;
    DBuf<int> cons_ids, c_order_cnt, c_order, c_order_basket, fatal;
    DBuf<double> c_share, c_eshare, c_prev_v, c_budget, c_bshare, c_ebshare, c_utility, c_intra_exp;
    DBuf<i64> c_base_q, c_prev_q;
    // results of the consumers' solve until consumer_apply takes them; ahead_step: the step the solve in flight (or done) belongs to
    DBuf<i64> c_nx_q; DBuf<double> c_nx_v, c_nx_util; DBuf<int> c_nx_ret;
    i64 ahead_step = -1;
    // region partition (empty = the engine owns every agent)
    bool partitioned = false;
    int n_pm = 0;  // purchasing managers of owned agents
    DBuf<unsigned char> owned;
    DBuf<int> pm_ids;
    DBuf<i64> halo_send, halo_recv;
    i64 *x_flow_send = nullptr, *x_flow_recv = nullptr, *x_dem_send = nullptr, *x_dem_recv = nullptr, *x_sellers = nullptr;  // caller-owned ...
    DBuf<i64> own_flow_send, own_flow_recv, own_dem_send, own_dem_recv, own_sellers;  // ... or the engine's own (acc_set_halo with null buffers)
    // NCCL plane (acc_comm_init): the engine exchanges the halos itself
    int comm_rank = -1, comm_world = 0, comm_max_rows = 0;
    std::vector<i64> comm_send_cnt, comm_recv_cnt, comm_send_off, comm_recv_off;
    DBuf<int> comm_part_firms, comm_part_off;
    DBuf<i64> comm_rows_send, comm_rows_recv;
    // peer stores (acc_comm_enable_peer_stores): the other parts' arrays mapped over NVLink, the connections' ids there
    bool peer_stores = false;
    DBuf<signed char> x_peer;
    DBuf<int> x_rid;
    DBuf<PeerTable> peer_table;
    DBuf<i64> barrier_word;
    DBuf<unsigned long long> barrier_flags;  // this part's arrival counters (written by the peers)
    DBuf<int> barrier_timeout;
    PeerTable h_peers;
    unsigned long long barrier_epoch = 0;
    bool flag_barrier = false;
    bool peer_sellers = false;  // the sellers' records travel by peer stores too (no collective in the step)
    std::vector<void*> ipc_opened;
#ifndef ACC_HOST_EMU
    NcclApi::comm_t comm = nullptr;
#endif
    // events
    DBuf<Event> ev_buf;
    DBuf<int> ev_count;
    int ev_cap = 1 << 16;

    // every array that changes from step to step, in the fixed order of the checkpoint blob (acc_save_state / acc_load_state)
    template <class V>
    void visit_state(V& v) {
        v(a_lambda); v(nu);
        v(f_cap_q); v(f_Xhat_q); v(f_est_q); v(f_X_q); v(f_Xexp_q); v(f_nbar); v(f_cutoff); v(f_desired_q); v(f_sumD_q); v(f_bk_sq); v(f_bk_cnt);
        v(f_bk_pmax); v(f_bk_pmin); v(f_exp_q);
        v(f_cap_v); v(f_Xhat_v); v(f_est_v); v(f_X_v); v(f_Xexp_v); v(f_C); v(f_R); v(f_sumD_v); v(f_bk_sv); v(f_exp_v);
        v(pub);
        v(s_S_q); v(s_I0_q); v(s_I1_q); v(s_Il_q); v(s_U_q); v(s_Ut_q); v(s_T_q); v(s_dlv_q); v(s_imp_q); v(s_dem_q);
        v(s_S_v); v(s_I0_v); v(s_I1_v); v(s_Il_v); v(s_U_v); v(s_Ut_v); v(s_T_v); v(s_imp_v); v(s_dem_v); v(s_exp_costs); v(s_opt_val); v(s_tpen);
        v(s_iters); v(s_evals);
        v(e_D_q); v(e_D_p); v(conv_q); v(conv_v);
        v(e_ovf_q); v(e_ovf_v); v(e_ship_q); v(e_ship_v);  // empty until passage forcing is used
        v(c_prev_q); v(c_prev_v); v(c_utility);
        v(ev_count); v(ev_buf);
    }
    // every device array of the engine (acc_destroy releases them, acc_fork shares the ones visit_state does not name)
    template <class V>
    void visit_all(V& v) {
        visit_state(v);
        v(sectors); v(a_type); v(a_sector); v(a_region); v(a_st_off); v(a_st_list); v(firm_ids); v(a_beta);
        v(f_Xstar_q); v(f_markup_eff); v(f_var_costs); v(f_Xstar_v);
        v(pubT); v(seller_row); v(row_firm); v(e_to); v(sto_of); v(obs_out); v(obs_idx); v(obs_inc_q); v(obs_inc_v);
        v(s_Istar_q); v(s_Sstar_q); v(s_defc); v(s_Istar_v); v(s_Sstar_v); v(s_a); v(s_agent); v(s_sector);
        v(e_seller_off); v(e_route); v(route_off); v(route_ent); v(ent_delay); v(e_Z_q); v(e_tau); v(e_dom);
        v(ell_off); v(in_off); v(ell_width); v(ell_edge); v(in_edge); v(in_seller); v(in_Z); v(e_slotw); v(ell_slotw); v(in_slotw); v(e_sort); v(f_sorted);
        v(cons_ids); v(c_order_cnt); v(c_order); v(c_order_basket); v(fatal); v(c_nx_q); v(c_nx_v); v(c_nx_util); v(c_nx_ret);
        v(c_share); v(c_eshare); v(c_budget); v(c_bshare); v(c_ebshare); v(c_intra_exp); v(c_base_q);
        v(owned); v(pm_ids); v(halo_send); v(halo_recv);
        v(own_flow_send); v(own_flow_recv); v(own_dem_send); v(own_dem_recv); v(own_sellers);
        v(comm_part_firms); v(comm_part_off); v(comm_rows_send); v(comm_rows_recv);
        v(x_peer); v(x_rid); v(peer_table); v(barrier_word); v(barrier_flags); v(barrier_timeout);
    }
    // ensemble members (acc_fork): a member shares its parent's immutable tables; the parent's memory lives until the last member is gone
    acc_engine* parent = nullptr;
    int members = 0;
    bool zombie = false;

    void ensure_forcing_state() {  // overflow per (hop, connection) exists from the first non-negative passage on
        if (forcing_used) return;
        e_ovf_q.alloc((size_t)nE * h_max); e_ovf_v.alloc((size_t)nE * h_max);
        e_ovf_q.zero(stream); e_ovf_v.zero(stream);
        if (hops_rewrite_shipment) { e_ship_q.alloc((size_t)nE); e_ship_q.zero(stream); e_ship_v.alloc((size_t)nE); e_ship_v.zero(stream); }
        forcing_used = true;
    }

    ~acc_engine() {
#ifndef ACC_HOST_EMU
        for (auto& c : clocks)
            for (auto& pr : c.pending) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
        for (void* q : ipc_opened) cudaIpcCloseMemHandle(q);
        if (comm != nullptr) NcclApi::get().CommDestroy(comm);
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (ev_join) cudaEventDestroy(ev_join);
        if (side) cudaStreamDestroy(side);
        if (stream) cudaStreamDestroy(stream);
#endif
    }

    StorageView storage_view() {
        StorageView v;
        v.S_q = s_S_q.p; v.S_v = s_S_v.p;
        v.Ip_q = reg == 0 ? s_I0_q.p : s_I1_q.p; v.Ip_v = reg == 0 ? s_I0_v.p : s_I1_v.p;
        v.In_q = reg == 0 ? s_I1_q.p : s_I0_q.p; v.In_v = reg == 0 ? s_I1_v.p : s_I0_v.p;
        v.Il_q = s_Il_q.p; v.Il_v = s_Il_v.p;
        v.U_q = s_U_q.p; v.U_v = s_U_v.p; v.Ut_q = s_Ut_q.p; v.Ut_v = s_Ut_v.p;
        v.T_q = s_T_q.p; v.T_v = s_T_v.p; v.dlv_q = s_dlv_q.p; v.imp_q = s_imp_q.p; v.imp_v = s_imp_v.p;
        v.Istar_q = s_Istar_q.p; v.Istar_v = s_Istar_v.p; v.Sstar_q = s_Sstar_q.p; v.Sstar_v = s_Sstar_v.p;
        v.a = s_a.p; v.defc = s_defc.p; v.agent = s_agent.p; v.sector = s_sector.p;
        return v;
    }
    FirmView firm_view() {
        FirmView v;
        v.pub = pub.p;
        v.pubT = seller_table();
        v.peers = (peer_stores && peer_sellers) ? peer_table.p : nullptr; v.n_parts = comm_world; v.my_part = comm_rank;
        v.Xstar_q = f_Xstar_q.p; v.Xstar_v = f_Xstar_v.p; v.beta = a_beta.p; v.lambda = a_lambda.p; v.sector = a_sector.p; v.region = a_region.p;
        v.type = a_type.p; v.markup_eff = f_markup_eff.p; v.var_costs = f_var_costs.p; v.st_off = a_st_off.p; v.st_list = a_st_list.p;
        v.cap_q = f_cap_q.p; v.cap_v = f_cap_v.p; v.Xhat_q = f_Xhat_q.p; v.Xhat_v = f_Xhat_v.p; v.est_q = f_est_q.p; v.est_v = f_est_v.p;
        v.X_q = f_X_q.p; v.X_v = f_X_v.p; v.Xexp_q = f_Xexp_q.p; v.Xexp_v = f_Xexp_v.p; v.nbar = f_nbar.p; v.cutoff = f_cutoff.p;
        v.C = f_C.p; v.R = f_R.p; v.desired_q = f_desired_q.p; v.sumD_q = f_sumD_q.p; v.sumD_v = f_sumD_v.p;
        v.bk_sq = f_bk_sq.p; v.bk_sv = f_bk_sv.p; v.bk_cnt = f_bk_cnt.p; v.bk_pmax = f_bk_pmax.p; v.bk_pmin = f_bk_pmin.p;
        v.exp_q = f_exp_q.p; v.exp_v = f_exp_v.p; v.sorted = f_sorted.p;
        return v;
    }
    SellerTable seller_table() {
        SellerTable t;
        t.f = pubT.p; t.row = seller_row.p; t.stride = (i64)h_row_firm.size();
        return t;
    }
    EdgeView edge_view() {
        EdgeView v;
        v.seller_off = e_seller_off.p; v.D_q = e_D_q.p; v.D_p = e_D_p.p; v.Z_q = e_Z_q.p; v.tau = e_tau.p; v.domestic = e_dom.p; v.route = e_route.p; v.route_off = route_off.p; v.route_ent = route_ent.p; v.ent_delay = ent_delay.p;
        v.conv_q = conv_q.p; v.conv_v = conv_v.p; v.n_edges = nE; v.T = T;
        v.slotw = e_slotw.p; v.rings = rings; v.sort_scratch = e_sort.p;
        for (int r = 0; r < n_rings; ++r) {
            const int L = rings.len[r], ph = (int)(step % (i64)L);
            v.rings.phase[r] = (unsigned char)ph;
            v.rings.front[r] = rings.base[r] + (i64)ph * rings.count[r];
            v.rings.last[r] = rings.base[r] + (i64)(ph == 0 ? L - 1 : ph - 1) * rings.count[r];
        }
        v.ovf_q = forcing_used ? e_ovf_q.p : nullptr; v.ovf_v = forcing_used ? e_ovf_v.p : nullptr; v.nu = forcing_used ? nu.p : nullptr;
        v.ship_q = e_ship_q.n ? e_ship_q.p : nullptr;
        v.ship_v = e_ship_v.n ? e_ship_v.p : nullptr;
        v.x_peer = peer_stores ? x_peer.p : nullptr; v.x_rid = peer_stores ? x_rid.p : nullptr; v.peers = peer_stores ? peer_table.p : nullptr;
        return v;
    }
    ConsumerView consumer_view() {
        ConsumerView v;
        v.cons_ids = cons_ids.p; v.st_off = a_st_off.p; v.st_list = a_st_list.p;
        v.share = c_share.p; v.eshare = c_eshare.p; v.base_q = c_base_q.p; v.prev_q = c_prev_q.p; v.prev_v = c_prev_v.p;
        v.budget = c_budget.p; v.order_cnt = c_order_cnt.p; v.order = c_order.p; v.order_basket = c_order_basket.p;
        v.bshare = c_bshare.p; v.ebshare = c_ebshare.p; v.utility = c_utility.p; v.intra_exp = c_intra_exp.p; v.inter_exp = c_inter_exp;
        v.n_baskets = nB; v.xtol_c = ACC_PREC3 * 1.0; v.ctol = c_ctol; v.maxeval = mp.maxeval;
        return v;
    }
    EventSink sink() {
        EventSink s;
        s.buf = ev_buf.p; s.count = ev_count.p; s.cap = ev_cap; s.step = step;
        return s;
    }

    int clock_id(const char* name) {
        for (size_t i = 0; i < clocks.size(); ++i) if (std::strcmp(clocks[i].name, name) == 0) return (int)i;
        KernelClock c;
        c.name = name;
        clocks.push_back(c);
        return (int)clocks.size() - 1;
    }
    template <class Fn>
    void timed(const char* name, Fn fn) { timed_on(stream, name, fn); }
    template <class Fn>
    void timed_comm(const char* name, Fn fn) { timed_on(stream, name, fn, false); }  // a collective: clocked, not counted as a kernel launch
    template <class Fn>
    void timed_on(stream_t on, const char* name, Fn fn, bool kernel = true) {
        if (kernel) ++launches;
#ifndef ACC_HOST_EMU
        if (profile) {
            const int id = clock_id(name);
            cudaEvent_t a, b;
            ACC_CUDA(cudaEventCreate(&a));
            ACC_CUDA(cudaEventCreate(&b));
            ACC_CUDA(cudaEventRecord(a, on));
            fn();
            ACC_CUDA(cudaEventRecord(b, on));
            clocks[id].pending.push_back(std::make_pair(a, b));
            clocks[id].launches += 1;
            return;
        }
#endif
        (void)name; (void)on;
        fn();
    }
    void collect_clocks() {
#ifndef ACC_HOST_EMU
        dev_sync(stream);
        for (auto& c : clocks) {
            for (auto& pr : c.pending) {
                float ms = 0.f;
                cudaEventElapsedTime(&ms, pr.first, pr.second);
                c.ms += ms;
                cudaEventDestroy(pr.first);
                cudaEventDestroy(pr.second);
            }
            c.pending.clear();
        }
#endif
    }

    // The consumers' utility optimisation reads storages and writes their desired use only, so it runs on a side
    // stream next to the sellers' production / distribution pass; the storage pass waits for both.
    void launch_delivery(const unsigned char* own) {
        if (forcing_used) {
            StorageDelivery<true> k{storage_view(), firm_view(), edge_view(), mp, sectors.p, sink(), step, ell_off.p, ell_width.p, ell_edge.p, ell_slotw.p, own};
            timed("storage_delivery", [&] { launch_thread(stream, nSt, k); });
        } else {
            StorageDelivery<false> k{storage_view(), firm_view(), edge_view(), mp, sectors.p, sink(), step, ell_off.p, ell_width.p, ell_edge.p, ell_slotw.p, own};
            timed("storage_delivery", [&] { launch_thread(stream, nSt, k); });
        }
        consumer_ahead();  // the storages are final for this step: the consumers' solve of the next step starts here
    }
    // The solve of step t reads the storages as the delivery pass of step t - 1 left them and nothing else that changes, and
    // writes only its own result arrays, so it is started right behind that delivery pass (`ahead`: with the input-flow registers
    // already seen the way step t will see them) and runs beside the sellers' expectation pass and the purchasing managers of
    // step t - 1.  consumer_join applies the results where Consumer::iterate_consumption_and_production would have computed
    // them.  Whatever rewrites state from outside (acc_load_state, forcing) drops a solve made ahead: it is repeated in place.
    ConsumerStage consumer_stage() { return ConsumerStage{c_nx_q.p, c_nx_v.p, c_nx_util.p, c_nx_ret.p}; }
    void consumer_fork(bool ahead = false) {
        if (!mp.utilitarian || nC == 0) return;
        if (!ahead && ahead_step == step) return;  // solved behind the previous step's delivery pass
        if (ahead) reg = 1 - reg;
        ConsumerUtility k{consumer_view(), storage_view(), mp, consumer_stage(), c_n_max};
        if (ahead) reg = 1 - reg;
#ifndef ACC_HOST_EMU
        ACC_CUDA(cudaEventRecord(ev_fork, stream));
        ACC_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
#endif
        timed_on(side, "consumer_utility", [&] { launch_warps(side, nC, consumer_workspace_bytes(c_n_max), k); });
#ifndef ACC_HOST_EMU
        ACC_CUDA(cudaEventRecord(ev_join, side));
#endif
        ahead_step = ahead ? step + 1 : step;
    }
    void consumer_join() {
        if (!mp.utilitarian || nC == 0) return;
#ifndef ACC_HOST_EMU
        ACC_CUDA(cudaStreamWaitEvent(stream, ev_join, 0));
#endif
        ConsumerApply k{consumer_view(), storage_view(), consumer_stage(), sink(), fatal.p};
        timed("consumer_apply", [&] { launch_warps(stream, nC, 0, k); });
    }
    void consumer_ahead() {
#ifndef ACC_NO_CONSUMER_AHEAD
        consumer_fork(true);
#endif
    }
    void stage_sell() {
        consumer_fork();
        {
            SellerProduction k{firm_view(), edge_view(), storage_view(), mp, sectors.p, sink(), step, firm_ids.p};
            timed("seller_production", [&] { launch_seller(stream, nF, book_cap, perm_slots, k); });
            ship_step = step;
        }
        if (halo_send.n && !peer_stores) {
            PackFlow k{edge_view(), halo_send.p, x_flow_send};
            timed("pack_flow", [&] { launch_thread(stream, (int)halo_send.n, k); });
        }
    }
    void stage_deliver() {
        if (halo_recv.n) {
            UnpackFlow k{edge_view(), halo_recv.p, x_flow_recv};
            timed("unpack_flow", [&] { launch_thread(stream, (int)halo_recv.n, k); });
        }
        consumer_join();
        {
            launch_delivery(partitioned ? owned.p : nullptr);
        }
        {
            SellerExpectation k{firm_view(), edge_view(), storage_view(), mp, sectors.p, sink(), firm_ids.p};
            timed("seller_expectation", [&] { launch_warp(stream, nF, 0, k); });
        }
        if (partitioned && x_sellers != nullptr) {
            PackSellers k{pub.p, owned.p, x_sellers};
            timed("pack_sellers", [&] { launch_thread(stream, nA, k); });
        }
    }
    template <bool SEP, bool STD, int G, bool DVAL = (G > 0)>  // the base group keeps the D_r summands; the large groups trade them for a second CTA per SM
    void launch_purchase() {
        Purchase<SEP, STD, DVAL> k{storage_view(), edge_view(), mp, sectors.p, sink(), step, in_off.p, in_edge.p, in_seller.p, in_slotw.p, in_Z.p, seller_table(), n_max,
                             partitioned ? pm_ids.p : nullptr, s_dem_q.p, s_dem_v.p, s_exp_costs.p, s_opt_val.p, s_tpen.p, s_iters.p, s_evals.p,
                             partitioned ? n_pm : nSt, ACC_PM_PREFETCH_AHEAD};
        timed("purchase", [&] { launch_block_smem<G>(stream, partitioned ? n_pm : nSt, pm_group, pm_workspace_bytes(n_max, SEP, DVAL), k); });
    }
    void stage_purchase() {
        if (partitioned && x_sellers != nullptr) {
            UnpackSellers k{pub.p, seller_table(), owned.p, x_sellers};
            timed("unpack_sellers", [&] { launch_thread(stream, nA, k); });
        }
        purchase_and_pack();
    }
    void purchase_and_pack() {
        // One CTA of pm_group threads per purchasing manager, all its vectors in shared memory.  The group size is part
        // of the arithmetic (reduction order): ACCSQP_GROUP, or accsqp_group_for(largest manager) -- the same rule the
        // oracle's host context follows.
        const bool dev = mp.deviation_penalty != 0;
        const bool std_pen = mp.quadratic_penalty != 0 && mp.relative_penalty != 0 && !dev;
        if (std_pen) {
            if (pm_group == ACCSQP_GROUP) launch_purchase<false, true, ACCSQP_GROUP>();
            else if (pm_group == 384) launch_purchase<false, true, 384, false>();  // 748 connections per manager (BASELINE.json configs[3]): 748x3 probe 1.30 -> 0.98 ms
            else launch_purchase<false, true, 0>();
        } else if (dev) {
            launch_purchase<true, false, 0>();
        } else {
            launch_purchase<false, false, 0>();
        }
        if (halo_recv.n && !peer_stores) {
            PackDemand pk{edge_view(), halo_recv.p, x_dem_send};
            timed("pack_demand", [&] { launch_thread(stream, (int)halo_recv.n, pk); });
        }
    }
    void stage_close() {
        if (halo_send.n) {
            UnpackDemand k{edge_view(), halo_send.p, x_dem_recv};
            timed("unpack_demand", [&] { launch_thread(stream, (int)halo_send.n, k); });
        }
        tick();
    }
#ifndef ACC_HOST_EMU
    // all-to-all of (2 x int64) entries grouped by peer: one NCCL point-to-point group on the engine's stream
    void exchange_halo(const i64* send, const std::vector<i64>& scnt, const std::vector<i64>& soff, i64* recv, const std::vector<i64>& rcnt,
                       const std::vector<i64>& roff) {
        NcclApi& nc = NcclApi::get();
        ACC_NCCL(nc.GroupStart());
        for (int p = 0; p < comm_world; ++p) {
            if (p == comm_rank) continue;
            if (scnt[p]) ACC_NCCL(nc.Send(send + 2 * soff[p], (size_t)(2 * scnt[p]), NcclApi::Int64, p, comm, stream));
            if (rcnt[p]) ACC_NCCL(nc.Recv(recv + 2 * roff[p], (size_t)(2 * rcnt[p]), NcclApi::Int64, p, comm, stream));
        }
        ACC_NCCL(nc.GroupEnd());
        ++collectives;
    }
#endif
#ifndef ACC_HOST_EMU
    // all parts have reached this point of their streams, and the peer stores of the kernels before it are visible: a
    // one-word all-reduce (kernel boundaries make the stores visible system-wide, the collective orders the parts)
    void cross_barrier() {
        if (flag_barrier) {  // counters in peer memory: a few microseconds instead of a collective's launch
            ++barrier_epoch;
            cross_barrier_kernel<<<1, ACC_MAX_PARTS, 0, stream>>>(h_peers, comm_rank, comm_world, barrier_epoch, barrier_timeout.p);
            ACC_CUDA(cudaGetLastError());
            return;
        }
        ACC_NCCL(NcclApi::get().AllReduce(barrier_word.p, barrier_word.p + 1, 1, NcclApi::Int64, 0 /* ncclSum */, comm, stream));
        ++collectives;
    }
#endif
    i64 collectives = 0;
    // One timestep of a region-partitioned run with the engine doing its own exchanges (NCCL on the engine's stream):
    //   sellers' pass -> shipments to the buyers' parts -> storages, deliveries, expectation -> all-gather of the owned
    //   sellers' records -> purchase -> demand requests to the sellers' parts -> tick
    void step_partitioned() {
#ifdef ACC_HOST_EMU
        fail("acc_step_partitioned: the emulation has no NCCL; drive the stages with acc_stage");
#else
        if (comm == nullptr) fail("acc_step_partitioned: call acc_comm_init first");
        NcclApi& nc = NcclApi::get();
        stage_sell();
        if (peer_stores) timed_comm("barrier_shipments", [&] { cross_barrier(); });  // every part's shipments have landed on its buyers' conveyors
        else timed_comm("exchange_flow", [&] { exchange_halo(x_flow_send, comm_send_cnt, comm_send_off, x_flow_recv, comm_recv_cnt, comm_recv_off); });
        {   // stage_deliver without the full-table pack
            if (halo_recv.n) {
                UnpackFlow k{edge_view(), halo_recv.p, x_flow_recv};
                timed("unpack_flow", [&] { launch_thread(stream, (int)halo_recv.n, k); });
            }
            consumer_join();
            launch_delivery(owned.p);
            SellerExpectation k{firm_view(), edge_view(), storage_view(), mp, sectors.p, sink(), firm_ids.p};
            timed("seller_expectation", [&] { launch_warp(stream, nF, 0, k); });
            if (!(peer_stores && peer_sellers)) {
                PackSellerRows pk{pub.p, firm_ids.p, comm_rows_send.p};
                timed("pack_sellers", [&] { launch_thread(stream, nF, pk); });
            }
        }
        if (peer_stores && peer_sellers) {
            // the expectation pass has stored every owned firm's record into the other parts' tables: all of them have landed
            timed_comm("barrier_sellers", [&] { cross_barrier(); });
        } else {
            timed_comm("exchange_sellers", [&] {
                ACC_NCCL(nc.AllGather(comm_rows_send.p, comm_rows_recv.p, (size_t)comm_max_rows * ACC_SELLER_WORDS, NcclApi::Int64, comm, stream));
                ++collectives;
            });
            UnpackSellerRows k{pub.p, seller_table(), comm_part_firms.p, comm_part_off.p, comm_world, comm_rank, comm_max_rows, comm_rows_recv.p};
            timed("unpack_sellers", [&] { launch_thread(stream, comm_world * comm_max_rows, k); });
        }
        purchase_and_pack();
        if (peer_stores) timed_comm("barrier_demand", [&] { cross_barrier(); });  // every part's demand requests have landed in its sellers' books
        else timed_comm("exchange_demand", [&] { exchange_halo(x_dem_send, comm_recv_cnt, comm_recv_off, x_dem_recv, comm_send_cnt, comm_send_off); });
        stage_close();
#endif
    }
    void require_single(const char* what) {
        if (partitioned) fail(std::string(what) + ": this engine is one part of a region-partitioned run; drive it with acc_stage and exchange the halos in between");
    }
    // the reference's phases (Model::iterate_*) on an unpartitioned engine
    void phase_cp() {
        require_single("acc_phase");
        consumer_fork();
        SellerProduction k2{firm_view(), edge_view(), storage_view(), mp, sectors.p, sink(), step, firm_ids.p};
        timed("seller_production", [&] { launch_seller(stream, nF, book_cap, perm_slots, k2); });
        ship_step = step;
        consumer_join();
        launch_delivery(nullptr);
    }
    void phase_exp() {
        require_single("acc_phase");
        SellerExpectation k{firm_view(), edge_view(), storage_view(), mp, sectors.p, sink(), firm_ids.p};
        timed("seller_expectation", [&] { launch_warp(stream, nF, 0, k); });
    }
    void phase_pur() {
        require_single("acc_phase");
        stage_purchase();
    }
    void tick() {
        reg = 1 - reg;  // this step's arrivals become the next step's "other" register (Model::switch_registers)
        ++step;
    }
};

// ------------------------------------------------------------------------------------------------
// snapshot -> tables (host, once)
// ------------------------------------------------------------------------------------------------
namespace acc {

static void build(acc_engine& E, const acc_network& net, const acc_params& par, const uint8_t* agent_owned) {
    stream_t s = E.stream;
    auto own = [&](i64 a) { return agent_owned == nullptr || agent_owned[a] != 0; };
    const i64 nS = net.n_sectors, nR = net.n_regions, nA = net.n_agents, nE = net.n_edges;
    if (nS <= 0 || nR <= 0 || nA <= 0 || nE <= 0) fail("empty network");
    if (nE >= (i64)0x7fffffff) fail("more than 2^31-1 connections on one device: partition the network");
    E.nS = (int)nS; E.nR = (int)nR; E.nA = (int)nA; E.nE = nE;
    const double* m = par.model;
    // ModelInitializer.cpp:58-62, 1015-1079
    ModelP& mp = E.mp;
    mp.dt = (double)mkt(m[ACC_MP_DELTA_T]);
    if (!(mp.dt > 0)) fail("model.delta_t must be positive");
    mp.pen_small = mkp(m[ACC_MP_TRANSPORT_PENALTY_SMALL]);
    mp.pen_large = mkp(m[ACC_MP_TRANSPORT_PENALTY_LARGE]);
    mp.min_storage = m[ACC_MP_MIN_STORAGE];
    mp.generic_width = m[ACC_MP_CHEAPEST_PRICE_RANGE_WIDTH] < 0.0 ? 1 : 0;
    mp.width = mp.generic_width ? 0 : mkp(m[ACC_MP_CHEAPEST_PRICE_RANGE_WIDTH]);
    mp.quadratic_penalty = m[ACC_MP_QUADRATIC_TRANSPORT_PENALTY] != 0.0;
    mp.relative_penalty = m[ACC_MP_RELATIVE_TRANSPORT_PENALTY] != 0.0;
    mp.limited_by_markup = m[ACC_MP_MAX_DECREASE_RESERVATION_PRICE_LIMITED_BY_MARKUP] != 0.0;
    mp.deviation_penalty = m[ACC_MP_DEVIATION_PENALTY] != 0.0;
    mp.preserve_seller_price = m[ACC_MP_CHEAPEST_PRICE_RANGE_PRESERVE_SELLER_PRICE] != 0.0;
    mp.maxeval = (int)m[ACC_MP_OPTIMIZATION_MAXITER];
    mp.prec_adj = (int)m[ACC_MP_OPTIMIZATION_PRECISION_ADJUSTMENT];
    mp.utilitarian = m[ACC_MP_UTILITARIAN] != 0.0;
    mp.naive_expectations = m[ACC_MP_NAIVE_EXPECTATIONS] != 0.0;
    mp.always_extend = m[ACC_MP_ALWAYS_EXTEND_EXPECTED_DEMAND_CURVE] != 0.0;
    const bool no_self_supply = m[ACC_MP_NO_SELF_SUPPLY] != 0.0;

    E.h_sectors.resize(nS);
    for (i64 k = 0; k < nS; ++k) {
        const double* q = net.sector_params + k * ACC_SP_COUNT;
        SectorP sp;
        sp.omega = q[ACC_SP_UPPER_STORAGE_LIMIT];
        sp.psi = (double)mkt(q[ACC_SP_INITIAL_STORAGE_FILL_FACTOR] * mp.dt);
        sp.transport_type = (int)q[ACC_SP_TRANSPORT_TYPE];
        sp.eps = q[ACC_SP_SUPPLY_ELASTICITY];
        sp.dn = mkp(q[ACC_SP_PRICE_INCREASE_PRODUCTION_EXTENSION]);
        sp.dn_est = mkp(q[ACC_SP_ESTIMATED_PRICE_INCREASE_PRODUCTION_EXTENSION]);
        sp.markup = mkp(q[ACC_SP_INITIAL_MARKUP]);
        sp.refill = (double)mkt(q[ACC_SP_TARGET_STORAGE_REFILL_TIME] * mp.dt);
        sp.withdraw = (double)mkt(q[ACC_SP_TARGET_STORAGE_WITHDRAW_TIME] * mp.dt);
        if (!(sp.refill > 0) || !(sp.withdraw > 0)) fail("target storage refill / withdraw times must be positive");
        E.h_sectors[k] = sp;
    }

    std::vector<int> a_type(nA), a_sector(nA), a_region(nA);
    std::vector<double> a_beta(nA), a_lambda(nA, 1.0);
    for (i64 a = 0; a < nA; ++a) {
        a_type[a] = (int)net.agent_type[a]; a_sector[a] = (int)net.agent_sector[a]; a_region[a] = (int)net.agent_region[a]; a_beta[a] = net.agent_beta[a];
        if (a_sector[a] < 0 || a_sector[a] >= nS || a_region[a] < 0 || a_region[a] >= nR) fail("agent sector / region index out of range");
    }

    // ---- storages in creation order (ModelInitializer::initialize_connection, :187-213) ----
    std::map<std::pair<int, int>, int> sto_of;  // (input sector, buying agent) -> creation id
    std::vector<int> c_agent, c_sector;
    std::vector<int> e_cid(nE);
    std::vector<int> seller_off(nA + 1, 0);
    for (i64 e = 0; e < nE; ++e) {
        const i64 f = net.edge_from[e], t = net.edge_to[e];
        if (f < 0 || f >= nA || t < 0 || t >= nA) fail("connection endpoint out of range");
        if (a_type[f] != 1) fail("a connection's seller must be a firm");
        if (e > 0 && net.edge_from[e - 1] > f) fail("connections must be listed seller-major (sorted by edge_from)");
        if (no_self_supply && f == t) fail("self-supply connection although no_self_supply is set");
        if (net.edge_zq[e] <= 0) fail("initial flows must be positive");
        seller_off[f + 1] += 1;
        const std::pair<int, int> key(a_sector[f], (int)t);
        auto it = sto_of.find(key);
        if (it == sto_of.end()) {
            it = sto_of.emplace(key, (int)c_agent.size()).first;
            c_agent.push_back((int)t);
            c_sector.push_back(a_sector[f]);
        }
        e_cid[e] = it->second;
    }
    for (i64 a = 0; a < nA; ++a) seller_off[a + 1] += seller_off[a];
    {   // the sellers' demand books are staged in shared memory when two CTAs with the largest book fit one SM
        int out_max = 0;
        for (i64 a = 0; a < nA; ++a) out_max = std::max(out_max, seller_off[a + 1] - seller_off[a]);
        E.book_cap = seller_book_bytes(out_max) <= ACC_SELLER_BOOK_BUDGET ? out_max : 0;
        {   // the sort scratch of ordered_range_sums: in shared memory when it fits next to the book (or alone, for books read in place)
            int N = 1;
            while (N < out_max) N <<= 1;
            const size_t with = (E.book_cap > 0 ? seller_book_bytes(E.book_cap) : 0) + (size_t)N * sizeof(unsigned short);
            E.perm_slots = (N <= 32768 && with <= ACC_SELLER_BOOK_BUDGET) ? N : 0;
        }
    }
    const int nSt = (int)c_agent.size();
    E.nSt = nSt;
    // internal order: sector-major, buyer-minor
    std::vector<int> s2cid(nSt);
    for (int k = 0; k < nSt; ++k) s2cid[k] = k;
    std::sort(s2cid.begin(), s2cid.end(), [&](int x, int y) {
        if (c_sector[x] != c_sector[y]) return c_sector[x] < c_sector[y];
        return c_agent[x] < c_agent[y];
    });
    std::vector<int> cid2s(nSt);
    for (int k = 0; k < nSt; ++k) cid2s[s2cid[k]] = k;
    std::vector<int> st_agent(nSt), st_sector(nSt);
    for (int k = 0; k < nSt; ++k) { st_agent[k] = c_agent[s2cid[k]]; st_sector[k] = c_sector[s2cid[k]]; }
    // per-agent storage lists in creation order
    std::vector<int> a_st_off(nA + 1, 0), a_st_list(nSt);
    for (int c = 0; c < nSt; ++c) a_st_off[c_agent[c] + 1] += 1;
    for (i64 a = 0; a < nA; ++a) a_st_off[a + 1] += a_st_off[a];
    {
        std::vector<int> fill(a_st_off.begin(), a_st_off.end() - 1);
        for (int c = 0; c < nSt; ++c) a_st_list[fill[c_agent[c]]++] = cid2s[c];
    }

    // ---- per-connection constants, transport (BusinessConnection.cpp:38-65, ModelInitializer.cpp:689-700) ----
    // Routes: the geo entities between two regions per kind of transport (k = 0 aviation, 1 road/sea), at index
    // (k R + from) R + to.  A snapshot without routes has one connection entity per unordered region pair
    // (create_simple_transport_connection, ModelInitializer.cpp:689-700), numbered i (i - 1) / 2 + j for j < i.
    std::vector<unsigned char> ent_delay;
    std::vector<int> route_off((size_t)2 * nR * nR + 1, 0), route_ent;
    if (net.n_entities > 0) {
        if (net.entity_delay == nullptr || net.route_off == nullptr || net.route_entity == nullptr) fail("routes: null table");
        ent_delay.resize(net.n_entities);
        for (i64 k = 0; k < net.n_entities; ++k) {
            if (net.entity_delay[k] < 0 || net.entity_delay[k] > 254) fail("geo entity delay out of range (0..254 steps)");
            ent_delay[k] = (unsigned char)net.entity_delay[k];
        }
        const i64 n_routes = 2 * nR * nR;
        route_ent.assign(net.route_entity, net.route_entity + net.route_off[n_routes]);
        for (i64 r = 0; r <= n_routes; ++r) route_off[r] = (int)net.route_off[r];
        for (int x : route_ent) if (x < 0 || x >= net.n_entities) fail("route names an unknown geo entity");
    } else {
        ent_delay.resize((size_t)nR * (nR - 1) / 2 + 1, 0);
        for (i64 k = 0; k < 2; ++k)
            for (i64 i = 0; i < nR; ++i)
                for (i64 j = 0; j < nR; ++j) {
                    if (i != j) {
                        const i64 hi = i > j ? i : j, lo = i > j ? j : i;
                        const i64 d = net.delay[hi * nR + lo];
                        if (d < 0 || d > 254) fail("transport delay out of range (0..254 steps)");
                        ent_delay[hi * (hi - 1) / 2 + lo] = (unsigned char)d;
                        route_ent.push_back((int)(hi * (hi - 1) / 2 + lo));
                    }
                    route_off[(k * nR + i) * nR + j + 1] = (int)route_ent.size();
                }
    }
    E.nEnt = net.n_entities > 0 ? (int)net.n_entities : (int)(nR * (nR - 1) / 2);
    std::vector<unsigned char> tau(nE), dom(nE);
    std::vector<int> route(nE);
    std::vector<i64> Zq(net.edge_zq, net.edge_zq + nE);
    int Tmax = 0;
    E.h_max = 1;
    for (i64 e = 0; e < nE; ++e) {
        const int f = (int)net.edge_from[e], t = (int)net.edge_to[e];
        const int rf = a_region[f], rt = a_region[t];
        dom[e] = rf == rt;
        const int kind = E.h_sectors[a_sector[f]].transport_type;
        if (kind == 1 || rf == rt) {  // immediate transport or domestic: one link without geo entity, BusinessConnection.cpp:49-50
            tau[e] = 0; route[e] = -1;
        } else {
            const i64 r = ((i64)(kind == 0 ? 0 : 1) * nR + rf) * nR + rt;
            const int h0 = route_off[r], h1 = route_off[r + 1];
            if (h1 <= h0) fail("no route between two regions that trade");
            int d = 0;
            for (int h = h0; h < h1; ++h) d += ent_delay[route_ent[h]];
            if (d > 254) fail("transport delay out of range (0..254 steps)");
            tau[e] = (unsigned char)d; route[e] = (int)r;
            if (h1 - h0 > E.h_max) E.h_max = h1 - h0;
            if (ent_delay[route_ent[h0]] == 0) E.hops_rewrite_shipment = true;  // a hop may rewrite the slot the shipment sits on
        }
        if (tau[e] > Tmax) Tmax = tau[e];
    }
    E.T = Tmax + 1;
    // ---- conveyor rings: connections of equal delay, in connection order; at most 2^24 per group ----
    std::vector<unsigned> slotw(nE);
    {
        std::vector<i64> per_delay(Tmax + 1, 0);
        for (i64 e = 0; e < nE; ++e) ++per_delay[tau[e]];
        const i64 cap = (i64)1 << ACC_RING_RANK_BITS;
        std::vector<int> first_ring(Tmax + 1, -1);
        int nr = 0;
        i64 slots = 0;
        std::memset(&E.rings, 0, sizeof E.rings);
        for (int d = 0; d <= Tmax; ++d) {
            if (per_delay[d] == 0) continue;
            first_ring[d] = nr;
            for (i64 left = per_delay[d]; left > 0; left -= cap) {
                if (nr >= ACC_MAX_RINGS) fail("more than " + std::to_string((int)ACC_MAX_RINGS) + " conveyor ring groups (distinct delays, 2^24 connections each): partition the network");
                E.rings.base[nr] = slots;
                E.rings.count[nr] = (int)std::min(left, cap);
                E.rings.len[nr] = (unsigned char)(d + 1);
                slots += (i64)(d + 1) * E.rings.count[nr];
                ++nr;
            }
        }
        E.n_rings = nr;
        E.conv_slots = slots;
        std::vector<i64> seen(Tmax + 1, 0);
        for (i64 e = 0; e < nE; ++e) {
            const i64 k = seen[tau[e]]++;
            slotw[e] = ((unsigned)(first_ring[tau[e]] + (int)(k >> ACC_RING_RANK_BITS)) << ACC_RING_RANK_BITS) | (unsigned)(k & (cap - 1));
        }
    }

    // ---- initial state (Storage::add_initial_flow_Z_star, Firm::add_initial_production_X_star) ----
    std::vector<i64> Xstar_q(nA, 0), Istar_q(nSt, 0), Sstar_q(nSt, 0), defc(nSt, 0), T_q(nSt, 0);
    std::vector<double> Xstar_v(nA, 0.0), Istar_v(nSt, 0.0), Sstar_v(nSt, 0.0), T_v(nSt, 0.0);
    std::vector<i64> in_cnt(nSt + 1, 0);
    for (i64 e = 0; e < nE; ++e) {
        const int f = (int)net.edge_from[e];
        const int st = cid2s[e_cid[e]];
        const Flow z = flow_qp(Zq[e], mkp(1.0));
        Xstar_q[f] += z.q; Xstar_v[f] += z.v;
        Istar_q[st] += z.q; Istar_v[st] += z.v;
        const Flow add = flow_times_time(z, E.h_sectors[st_sector[st]].psi);
        const Flow ns = fcopy(fadd(mkflow(Sstar_q[st], Sstar_v[st]), add));
        Sstar_q[st] = ns.q; Sstar_v[st] = ns.v;
        defc[st] += z.q * (tau[e] + 1);
        double tv = 0.0;
        for (int k = 0; k < tau[e]; ++k) tv += z.v;
        T_q[st] += z.q * tau[e]; T_v[st] += tau[e] ? tv : 0.0;
        in_cnt[st + 1] += 1;
    }
    // clean_network must be a no-op for a valid snapshot (ModelInitializer.cpp:215-299)
    for (i64 a = 0; a < nA; ++a) {
        if (!own(a)) continue;  // agents of other parts are present with only the connections that touch this part
        if (a_type[a] == 1) {
            i64 input = 0;
            for (int k = a_st_off[a]; k < a_st_off[a + 1]; ++k) input += Istar_q[a_st_list[k]];
            if (Xstar_q[a] - input <= 0 || seller_off[a + 1] == seller_off[a] || a_st_off[a + 1] == a_st_off[a])
                fail("network snapshot is not clean: a firm would be removed by clean_network (no value added, no buyers or no inputs)");
        } else if (a_st_off[a + 1] == a_st_off[a]) {
            fail("network snapshot is not clean: consumer without inputs");
        }
    }
    // technology coefficients / elasticities (Storage.cpp:68, ModelInitializer.cpp:195-200)
    std::vector<double> st_a(nSt);
    for (int k = 0; k < nSt; ++k) {
        const int a = st_agent[k];
        st_a[k] = !own(a) ? 0.0 : (a_type[a] == 1 ? ratio(Istar_q[k], Xstar_q[a]) : net.consumer_elasticity[st_sector[k]]);
    }
    // SalesManager::initialize, SalesManager.cpp:626-643
    std::vector<i64> markup_eff(nA, 0), var_costs(nA, 0), est_q(nA, 0), X_q(nA, 0);
    std::vector<double> est_v(nA, 0.0), X_v(nA, 0.0);
    std::vector<int> firm_ids, all_firms;
    for (i64 a = 0; a < nA; ++a) {
        if (a_type[a] != 1) continue;
        all_firms.push_back((int)a);
        if (!own(a)) continue;
        firm_ids.push_back((int)a);
        i64 c0 = 0;
        for (int k = a_st_off[a]; k < a_st_off[a + 1]; ++k) c0 += pmul(mkp(1.0), st_a[a_st_list[k]]);
        const i64 one = mkp(1.0);
        const i64 mk = std::min(one - c0, E.h_sectors[a_sector[a]].markup);
        const i64 var = std::max((i64)0, one - (c0 + mk));
        markup_eff[a] = mk; var_costs[a] = var;
        Flow est = fcopy(fmul(mkflow(Xstar_q[a], Xstar_v[a]), a_beta[a]));
        est = set_price(est, c0 + var);
        est_q[a] = est.q; est_v[a] = est.v;
        X_q[a] = Xstar_q[a]; X_v[a] = Xstar_v[a];
    }
    E.nF = (int)firm_ids.size();
    // seller rows: all firms (of every part: a buyer reads the record of any seller), sector-major, agent-minor
    std::vector<int> seller_row(nA, -1);
    {
        std::vector<int> by_sector(all_firms);
        std::stable_sort(by_sector.begin(), by_sector.end(), [&](int x, int y) { return a_sector[x] < a_sector[y]; });
        for (size_t r = 0; r < by_sector.size(); ++r) seller_row[by_sector[r]] = (int)r;
        E.h_row_firm = by_sector;
        if (by_sector.size() >= ((size_t)1 << 24)) fail("more than 2^24 firms");
    }

    // ---- buyer-side lists ----
    std::vector<i64> in_off(nSt + 1, 0);
    for (int k = 0; k < nSt; ++k) in_off[k + 1] = in_off[k] + in_cnt[k + 1];
    std::vector<int> in_edge(nE), in_seller(nE);  // in_seller: seller row | delay << 24
    std::vector<unsigned> in_slotw(nE);
    std::vector<i64> in_Z(nE);
    if (nA >= (1 << 24)) fail("more than 2^24 agents");
    {
        std::vector<i64> fill(in_off.begin(), in_off.end() - 1);
        for (i64 e = 0; e < nE; ++e) {  // file order == seller ascending
            const int st = cid2s[e_cid[e]];
            in_edge[fill[st]] = (int)e;
            in_seller[fill[st]] = (int)((unsigned)seller_row[net.edge_from[e]] | ((unsigned)tau[e] << 24));
            in_Z[fill[st]] = Zq[e];
            in_slotw[fill[st]] = slotw[e];
            ++fill[st];
        }
    }
    int n_max = 0;
    for (int k = 0; k < nSt; ++k) n_max = std::max<i64>(n_max, in_off[k + 1] - in_off[k]);
    E.n_max = (n_max + 7) / 8 * 8;
    E.pm_group = accsqp_group_for(n_max);
    const int n_slices = (nSt + 31) / 32;
    std::vector<i64> ell_off(n_slices + 1, 0);
    std::vector<int> ell_width(n_slices, 0);
    for (int sl = 0; sl < n_slices; ++sl) {
        int w = 0;
        for (int k = sl * 32; k < std::min(nSt, sl * 32 + 32); ++k) w = std::max<i64>(w, in_off[k + 1] - in_off[k]);
        ell_width[sl] = w;
        ell_off[sl + 1] = ell_off[sl] + (i64)w * 32;
    }
    std::vector<int> ell_edge(ell_off[n_slices], -1);
    std::vector<unsigned> ell_slotw(ell_off[n_slices], slotw[0]);  // padding cells: a valid word (the loads are issued before the cell is tested)
    for (int k = 0; k < nSt; ++k) {
        const int sl = k >> 5, lane = k & 31;
        for (i64 j = in_off[k]; j < in_off[k + 1]; ++j) {
            ell_edge[ell_off[sl] + (j - in_off[k]) * 32 + lane] = in_edge[j];
            ell_slotw[ell_off[sl] + (j - in_off[k]) * 32 + lane] = in_slotw[j];
        }
    }

    // ---- upload ----
    E.sectors.upload(E.h_sectors, s);
    E.a_type.upload(a_type, s); E.a_sector.upload(a_sector, s); E.a_region.upload(a_region, s);
    E.a_st_off.upload(a_st_off, s); E.a_st_list.upload(a_st_list, s); E.firm_ids.upload(firm_ids, s);
    E.a_beta.upload(a_beta, s); E.a_lambda.upload(a_lambda, s);
    E.f_Xstar_q.upload(Xstar_q, s); E.f_Xstar_v.upload(Xstar_v, s); E.f_markup_eff.upload(markup_eff, s); E.f_var_costs.upload(var_costs, s);
    E.f_est_q.upload(est_q, s); E.f_est_v.upload(est_v, s); E.f_X_q.upload(X_q, s); E.f_X_v.upload(X_v, s);
    for (DBuf<i64>* b : {&E.f_cap_q, &E.f_Xhat_q, &E.f_Xexp_q, &E.f_nbar, &E.f_cutoff, &E.f_desired_q, &E.f_sumD_q, &E.f_bk_sq, &E.f_bk_cnt, &E.f_bk_pmax,
                         &E.f_bk_pmin, &E.f_exp_q}) { b->alloc(nA); b->zero(s); }
    for (DBuf<double>* b : {&E.f_cap_v, &E.f_Xhat_v, &E.f_Xexp_v, &E.f_C, &E.f_R, &E.f_sumD_v, &E.f_bk_sv, &E.f_exp_v}) { b->alloc(nA); b->zero(s); }
    E.pub.alloc(nA); E.pub.zero(s);
    E.pubT.alloc((size_t)SPT_FIELDS * E.h_row_firm.size()); E.pubT.zero(s);
    E.seller_row.upload(seller_row, s);
    E.row_firm.upload(E.h_row_firm, s);
    {
        std::vector<int> e_to(nE);
        for (i64 e = 0; e < nE; ++e) e_to[e] = (int)net.edge_to[e];
        E.e_to.upload(e_to, s);
        std::vector<int> sto_of((size_t)nS * nA, -1);
        for (int k = 0; k < nSt; ++k) sto_of[(size_t)st_sector[k] * nA + st_agent[k]] = k;
        E.sto_of.upload(sto_of, s);
        for (i64 a = 0; a < nA; ++a) E.max_storages = std::max(E.max_storages, a_st_off[a + 1] - a_st_off[a]);
    }

    E.s_agent.upload(st_agent, s); E.s_sector.upload(st_sector, s);
    E.s_Istar_q.upload(Istar_q, s); E.s_Istar_v.upload(Istar_v, s); E.s_Sstar_q.upload(Sstar_q, s); E.s_Sstar_v.upload(Sstar_v, s);
    E.s_S_q.upload(Sstar_q, s); E.s_S_v.upload(Sstar_v, s);
    // input_flow_I_[1] = input_flow_I_[2] = sum of Z*, [0] = 0; the first step reads register 1 as "other"
    E.s_I0_q.upload(Istar_q, s); E.s_I0_v.upload(Istar_v, s);
    E.s_I1_q.alloc(nSt); E.s_I1_q.zero(s); E.s_I1_v.alloc(nSt); E.s_I1_v.zero(s);
    E.s_Il_q.upload(Istar_q, s); E.s_Il_v.upload(Istar_v, s);
    E.reg = 0;
    for (DBuf<i64>* b : {&E.s_U_q, &E.s_Ut_q, &E.s_dlv_q, &E.s_imp_q, &E.s_dem_q}) { b->alloc(nSt); b->zero(s); }
    for (DBuf<double>* b : {&E.s_U_v, &E.s_Ut_v, &E.s_imp_v, &E.s_dem_v, &E.s_exp_costs, &E.s_opt_val, &E.s_tpen}) { b->alloc(nSt); b->zero(s); }
    E.s_iters.alloc(nSt); E.s_iters.zero(s); E.s_evals.alloc(nSt); E.s_evals.zero(s);
    E.s_T_q.upload(T_q, s); E.s_T_v.upload(T_v, s); E.s_defc.upload(defc, s); E.s_a.upload(st_a, s);
    // PurchasingManager::add_initial_demand_D_star: demand_D = sum of Z*
    h2d(E.s_dem_q.p, Istar_q.data(), nSt * sizeof(i64), s);
    h2d(E.s_dem_v.p, Istar_v.data(), nSt * sizeof(double), s);
    h2d(E.s_exp_costs.p, Istar_v.data(), nSt * sizeof(double), s);

    E.e_seller_off.upload(seller_off, s); E.e_route.upload(route, s); E.route_off.upload(route_off, s); E.route_ent.upload(route_ent, s); E.ent_delay.upload(ent_delay, s); E.e_tau.upload(tau, s); E.e_dom.upload(dom, s); E.e_Z_q.upload(Zq, s);
    {   // last_demand_request_D = Z* at price 1 (BusinessConnection.cpp:42)
        std::vector<i64> p1(nE, mkp(1.0));
        E.e_D_q.upload(Zq, s); E.e_D_p.upload(p1, s);
    }
    {   // delay queues start full of Z* (TransportChainLink.cpp:49): the slots of times 0 .. tau - 1 of every ring, one ring group at a time
        E.conv_q.alloc((size_t)E.conv_slots); E.conv_v.alloc((size_t)E.conv_slots);
        E.conv_q.zero(s); E.conv_v.zero(s);
        std::vector<std::vector<i64>> members(E.n_rings);
        for (int r = 0; r < E.n_rings; ++r) members[r].reserve(E.rings.count[r]);
        for (i64 e = 0; e < nE; ++e) members[ring_of(slotw[e])].push_back(e);  // connection order == rank order
        std::vector<i64> pq;
        std::vector<double> pv;
        for (int r = 0; r < E.n_rings; ++r) {
            const i64 cnt = E.rings.count[r];
            const int d = (int)E.rings.len[r] - 1;
            if (d == 0) continue;
            pq.resize(cnt); pv.resize(cnt);
            for (i64 k = 0; k < cnt; ++k) { pq[k] = Zq[members[r][k]]; pv[k] = flow_qp(pq[k], mkp(1.0)).v; }
            for (int i = 0; i < d; ++i) {
                h2d(E.conv_q.p + E.rings.base[r] + (i64)i * cnt, pq.data(), cnt * sizeof(i64), s);
                h2d(E.conv_v.p + E.rings.base[r] + (i64)i * cnt, pv.data(), cnt * sizeof(double), s);
            }
            dev_sync(s);
            std::vector<i64>().swap(members[r]);
        }
    }
    if (mp.utilitarian) {
        // ---- Consumer::Consumer + Consumer::initialize, Consumer.cpp:43-163 (host, once) ----
        const i64 nB = net.n_baskets;
        if (nB <= 0 || net.basket_of_sector == nullptr || net.basket_substitution == nullptr) fail("utilitarian consumers need consumer baskets");
        std::vector<double> intra_exp(nB);
        for (i64 b = 0; b < nB; ++b) {
            const double c = net.basket_substitution[b];
            intra_exp[b] = (c - 1) / c;
        }
        const double inter = net.inter_basket_substitution;
        E.c_inter_exp = (inter - 1) / inter;
        E.nB = (int)nB;
        double prec8 = 1.0;
        for (int k = 0; k < 8; ++k) prec8 = 0.1 * prec8;  // FlowValue::precision, types.h:313-315
        E.c_ctol = prec8;
        std::vector<int> cons_ids, order_cnt(nA, 0), order(nSt, -1), order_basket(nSt, -1);
        std::vector<double> share(nSt, 0.0), eshare(nSt, 0.0), prev_v(nSt, 0.0), budget(nA, 0.0), bshare((size_t)nA * nB, 0.0), ebshare((size_t)nA * nB, 0.0);
        std::vector<i64> base_q(nSt, 0), prev_q(nSt, 0);
        int c_n_max = 0;
        for (i64 a = 0; a < nA; ++a) {
            if (a_type[a] == 1 || !own(a)) continue;
            cons_ids.push_back((int)a);
            const int o = a_st_off[a], n = a_st_off[a + 1] - o;
            c_n_max = std::max(c_n_max, n);
            auto sto = [&](int i) { return a_st_list[o + i]; };
            auto basket_of = [&](int i) { return (int)net.basket_of_sector[st_sector[sto(i)]]; };
            double B = 0.0;  // consumption_budget
            for (int i = 0; i < n; ++i) B += Istar_v[sto(i)];
            budget[a] = B;
            // visit order of every loop over (basket, sector of the basket, storage of that sector)
            std::vector<std::pair<int, int> > visit;  // (basket, local index)
            for (i64 b = 0; b < nB; ++b)
                for (i64 sec = 0; sec < nS; ++sec) {
                    if (net.basket_of_sector[sec] != b) continue;
                    for (int i = 0; i < n; ++i) if (st_sector[sto(i)] == sec) visit.emplace_back((int)b, i);
                }
            double* bs = &bshare[(size_t)a * nB];
            for (auto& v : visit) bs[v.first] += Istar_v[sto(v.second)] / B;
            for (i64 b = 0; b < nB; ++b)
                if (bs[b] == 0) fail("a consumer has no consumption in one of its baskets (the reference removes such baskets; not provided)");
            double Cr = 0.0;  // consumption_in_relevant_baskets
            for (auto& v : visit) Cr += Istar_v[sto(v.second)];
            for (auto& v : visit) bs[v.first] = bs[v.first] * (B / Cr);  // once per storage of the basket, as the reference does (:124-131)
            for (int i = 0; i < n; ++i) {
                const int k = sto(i);
                if (basket_of(i) >= 0) {
                    share[k] = (Istar_v[k] / B) * (B / Cr);
                    base_q[k] = Istar_q[k]; prev_q[k] = Istar_q[k]; prev_v[k] = Istar_v[k];
                }
            }
            int cnt = 0;
            for (auto& v : visit) {
                const int k = sto(v.second);
                order[o + cnt] = v.second; order_basket[o + cnt] = v.first; ++cnt;
                share[k] = share[k] / bs[v.first];
                eshare[k] = std::pow(share[k], 1 / net.basket_substitution[v.first]);
            }
            order_cnt[a] = cnt;
            for (i64 b = 0; b < nB; ++b) ebshare[(size_t)a * nB + b] = std::pow(bs[b], 1 / inter);
        }
        E.nC = (int)cons_ids.size();
        E.c_n_max = (c_n_max + 1) / 2 * 2;
        E.cons_ids.upload(cons_ids, s); E.c_order_cnt.upload(order_cnt, s); E.c_order.upload(order, s); E.c_order_basket.upload(order_basket, s);
        E.c_share.upload(share, s); E.c_eshare.upload(eshare, s); E.c_prev_v.upload(prev_v, s); E.c_budget.upload(budget, s);
        E.c_bshare.upload(bshare, s); E.c_ebshare.upload(ebshare, s); E.c_intra_exp.upload(intra_exp, s);
        E.c_base_q.upload(base_q, s); E.c_prev_q.upload(prev_q, s);
        E.c_utility.alloc(nA); E.c_utility.zero(s);
        E.c_nx_q.alloc(nSt); E.c_nx_v.alloc(nSt); E.c_nx_util.alloc(nA); E.c_nx_ret.alloc(nA);
        E.c_nx_q.zero(s); E.c_nx_v.zero(s); E.c_nx_util.zero(s); E.c_nx_ret.zero(s);
    }
    E.fatal.alloc(1); E.fatal.zero(s);
    E.in_off.upload(in_off, s); E.in_edge.upload(in_edge, s); E.in_seller.upload(in_seller, s); E.in_Z.upload(in_Z, s);
    E.ell_off.upload(ell_off, s); E.ell_width.upload(ell_width, s); E.ell_edge.upload(ell_edge, s);
    E.e_slotw.upload(slotw, s); E.ell_slotw.upload(ell_slotw, s); E.in_slotw.upload(in_slotw, s);
    E.e_sort.alloc(2 * (size_t)nE); E.e_sort.zero(s);
    E.f_sorted.alloc((size_t)nA); E.f_sorted.zero(s);
    E.ev_buf.alloc(E.ev_cap); E.ev_count.alloc(1); E.ev_count.zero(s);
    E.nu.alloc((size_t)E.nEnt + 1);
    dev_sync(s);

    E.h_agent_type = a_type; E.h_agent_sector = a_sector; E.h_agent_region = a_region; E.h_st_agent = st_agent; E.h_st_sector = st_sector;
    E.h_cid2s = cid2s; E.h_firm_ids = all_firms; E.h_tau = tau;
    if (agent_owned != nullptr) {
        E.partitioned = true;
        std::vector<unsigned char> ow(nA);
        for (i64 a = 0; a < nA; ++a) ow[a] = own(a) ? 1 : 0;
        std::vector<int> pm;
        for (int k = 0; k < nSt; ++k) if (ow[st_agent[k]]) pm.push_back(k);
        E.n_pm = (int)pm.size();
        E.owned.upload(ow, s); E.pm_ids.upload(pm, s);
        dev_sync(s);
    }
    for (i64 a = 0; a < nA; ++a)
        if (a_type[a] != 1)
            for (int k = a_st_off[a]; k < a_st_off[a + 1]; ++k) E.h_consumer_st.push_back(a_st_list[k]);
}

// ---- reads ----
struct StateArray {
    std::vector<i64> i;
    std::vector<double> f;
    bool is_f64 = false;
};

template <class T>
static std::vector<T> to_creation_order(const acc_engine& E, const std::vector<T>& dev_order) {
    std::vector<T> out(dev_order.size());
    for (size_t c = 0; c < out.size(); ++c) out[c] = dev_order[E.h_cid2s[c]];
    return out;
}

static bool read_state(acc_engine& E, const std::string& name, StateArray& out) {
    stream_t s = E.stream;
    auto I = [&](const DBuf<i64>& b) { out.i = b.download(s); out.is_f64 = false; return true; };
    auto F = [&](const DBuf<double>& b) { out.f = b.download(s); out.is_f64 = true; return true; };
    auto SI = [&](const DBuf<i64>& b) { out.i = to_creation_order(E, b.download(s)); out.is_f64 = false; return true; };
    auto SF = [&](const DBuf<double>& b) { out.f = to_creation_order(E, b.download(s)); out.is_f64 = true; return true; };
    // connections (file order)
    if (name == "edge_D_q") return I(E.e_D_q);
    if (name == "edge_D_p") return I(E.e_D_p);
    if (name == "edge_ship_q" || name == "edge_ship_v" || name == "edge_deliv_q" || name == "edge_deliv_v" || name == "edge_transit_q" ||
        name == "edge_transit_v" || name == "edge_deficit_q") {
        // derived from the conveyor by a gather kernel (EdgeObservables).  t = step of the last production pass, whether or
        // not the step has been ticked since (the reference's output sweep runs before Model::tick, src/ModelRun.cpp:198-226).
        // Before the first step everything is Z*.
        const bool q = name.back() == 'q';
        out.is_f64 = !q;
        const i64 t = E.ship_step;
        if (t < 0) {
            std::vector<i64> z = E.e_Z_q.download(s);
            if (q) out.i.assign(E.nE, 0); else out.f.assign(E.nE, 0.0);
            const bool transit = name.find("transit") != std::string::npos;
            for (i64 e = 0; e < E.nE; ++e) {
                const Flow one = flow_qp(z[e], mkp(1.0));
                Flow f = one;
                if (transit) { f = mkflow(0, 0.0); for (int k = 0; k < E.h_tau[e]; ++k) f = fadd(f, one); }
                if (name == "edge_deficit_q") out.i[e] = 0;
                else if (q) out.i[e] = f.q; else out.f[e] = f.v;
            }
            return true;
        }
        DBuf<i64> bq; DBuf<double> bv;
        if (q) bq.alloc(E.nE); else bv.alloc(E.nE);
        EdgeObservables k{E.edge_view(), t, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        if (name == "edge_ship_q") k.ship_q = bq.p;
        else if (name == "edge_ship_v") k.ship_v = bv.p;
        else if (name == "edge_deliv_q") k.dlv_q = bq.p;
        else if (name == "edge_deliv_v") k.dlv_v = bv.p;
        else if (name == "edge_transit_q") k.tr_q = bq.p;
        else if (name == "edge_transit_v") k.tr_v = bv.p;
        else k.def_q = bq.p;
        E.timed("edge_observables", [&] { launch_thread(s, (int)E.nE, k); });
        if (q) out.i = bq.download(s); else out.f = bv.download(s);
        bq.free(); bv.free();
        return true;
    }
    // storages (creation order)
    if (name == "sto_agent" || name == "sto_sector") {  // identity of each storage (creation order): buying agent, input sector
        const std::vector<int>& h = name == "sto_agent" ? E.h_st_agent : E.h_st_sector;
        std::vector<i64> dev(h.begin(), h.end());
        out.i = to_creation_order(E, dev); out.is_f64 = false;
        return true;
    }
    if (name == "sto_S_q") return SI(E.s_S_q);
    if (name == "sto_S_v") return SF(E.s_S_v);
    if (name == "sto_I_q") return SI(E.s_Il_q);
    if (name == "sto_I_v") return SF(E.s_Il_v);
    if (name == "sto_Inext_q") {  // Storage::current_input_flow_I() as the oracle dumps it before the tick: input_flow_I[other_register],
        // the register the storage pass has consumed and zeroed (Storage.cpp:61,90); after the tick it is the one that fills next
        const bool ticked = E.step != E.ship_step;
        const bool first = ticked ? E.reg != 0 : E.reg == 0;
        return SI(first ? E.s_I0_q : E.s_I1_q);
    }
    if (name == "sto_U_q") return SI(E.s_U_q);
    if (name == "sto_U_v") return SF(E.s_U_v);
    if (name == "sto_Ut_q") return SI(E.s_Ut_q);
    if (name == "sto_Ut_v") return SF(E.s_Ut_v);
    if (name == "sto_demand_q") return SI(E.s_dem_q);
    if (name == "sto_demand_v") return SF(E.s_dem_v);
    if (name == "sto_expected_costs") return SF(E.s_exp_costs);
    if (name == "sto_optimized_value") return SF(E.s_opt_val);
    if (name == "sto_transport_penalty") return SF(E.s_tpen);
    if (name == "sto_transit_q") return SI(E.s_T_q);
    if (name == "sto_transit_v") return SF(E.s_T_v);
    if (name == "sto_import_q") return SI(E.s_imp_q);
    if (name == "sto_import_v") return SF(E.s_imp_v);
    if (name == "sto_delivered_q") return SI(E.s_dlv_q);
    if (name == "sto_solver_iters" || name == "sto_solver_evals") {
        std::vector<int> h = (name == "sto_solver_iters" ? E.s_iters : E.s_evals).download(s);
        std::vector<i64> w(h.begin(), h.end());
        out.i = to_creation_order(E, w); out.is_f64 = false;
        return true;
    }
    if (name == "sto_deficit_q") {
        std::vector<i64> a = E.s_defc.download(s), b = E.s_dlv_q.download(s), c = E.s_T_q.download(s);
        for (size_t k = 0; k < a.size(); ++k) a[k] = a[k] - b[k] - c[k];
        out.i = to_creation_order(E, a); out.is_f64 = false;
        return true;
    }
    // agents
    if (name == "agent_forcing") return F(E.a_lambda);
    if (name == "consumer_utility" && E.mp.utilitarian) return F(E.c_utility);
    if (name == "consumer_solver_ret" && E.mp.utilitarian) {  // return code of the last utility optimisation per agent (csrc/sqp_core.h)
        std::vector<int> h = E.c_nx_ret.download(s);
        out.i.assign(h.begin(), h.end()); out.is_f64 = false;
        return true;
    }
    if (name == "firm_sorted_sums") {  // diagnostic: sellers whose last pro-rata distribution needed the reference's summation order
        std::vector<int> h = E.f_sorted.download(s);
        out.i.assign(h.begin(), h.end()); out.is_f64 = false;
        return true;
    }
    if (name == "firm_X_q") return I(E.f_X_q);
    if (name == "firm_X_v") return F(E.f_X_v);
    if (name == "firm_Xhat_q") return I(E.f_Xhat_q);
    if (name == "firm_Xhat_v") return F(E.f_Xhat_v);
    if (name == "firm_Xhat_p") {
        std::vector<i64> q = E.f_Xhat_q.download(s);
        std::vector<double> v = E.f_Xhat_v.download(s);
        out.i.resize(q.size()); out.is_f64 = false;
        for (size_t k = 0; k < q.size(); ++k) out.i[k] = flow_price(mkflow(q[k], v[k]));
        return true;
    }
    if (name == "firm_Xexp_q") return I(E.f_Xexp_q);
    if (name == "firm_Xexp_v") return F(E.f_Xexp_v);
    if (name == "firm_nbar") return I(E.f_nbar);
    if (name == "firm_cutoff") return I(E.f_cutoff);
    if (name == "firm_costs_C") return F(E.f_C);
    if (name == "firm_revenue_R") return F(E.f_R);
    if (name == "firm_desired_q") return I(E.f_desired_q);
    if (name == "firm_possible_q") return I(E.f_cap_q);
    if (name == "firm_export_q") return I(E.f_exp_q);
    if (name == "firm_export_v") return F(E.f_exp_v);
    if (name == "firm_incoming_demand_q" || name == "firm_incoming_demand_v") {
        DBuf<i64> q; DBuf<double> v;
        q.alloc(E.nA); v.alloc(E.nA); q.zero(s); v.zero(s);
        FirmIncomingDemand k{E.edge_view(), E.firm_ids.p, q.p, v.p};
        E.timed("firm_incoming_demand", [&] { launch_block(s, E.nF, k); });
        if (name == "firm_incoming_demand_q") I(q); else F(v);
        q.free(); v.free();
        return true;
    }
    // sector / region sums, folded in the reference's serial order (agent order)
    if (name == "sector_production_q" || name == "sector_production_v") {
        std::vector<i64> q = E.f_X_q.download(s);
        std::vector<double> v = E.f_X_v.download(s);
        out.is_f64 = name == "sector_production_v";
        out.i.assign(E.nS, 0); out.f.assign(E.nS, 0.0);
        for (int a = 0; a < E.nA; ++a)
            if (E.h_agent_type[a] == 1) { out.i[E.h_agent_sector[a]] += q[a]; out.f[E.h_agent_sector[a]] += v[a]; }
        return true;
    }
    if (name == "sector_demand_q" || name == "sector_demand_v") {
        // Sector::total_demand_D (src/model/Sector.cpp:36-39): every demand request of the step, by the SELLER's sector
        DBuf<i64> q; DBuf<double> v;
        q.alloc(E.nA); v.alloc(E.nA); q.zero(s); v.zero(s);
        FirmIncomingDemand k{E.edge_view(), E.firm_ids.p, q.p, v.p};
        E.timed("firm_incoming_demand", [&] { launch_block(s, E.nF, k); });
        const std::vector<i64> hq = q.download(s);
        const std::vector<double> hv = v.download(s);
        q.free(); v.free();
        out.is_f64 = name == "sector_demand_v";
        out.i.assign(E.nS, 0); out.f.assign(E.nS, 0.0);
        for (int a = 0; a < E.nA; ++a)
            if (E.h_agent_type[a] == 1) { out.i[E.h_agent_sector[a]] += hq[a]; out.f[E.h_agent_sector[a]] += hv[a]; }
        return true;
    }
    if (name == "region_export_q" || name == "region_export_v" || name == "region_import_q" || name == "region_import_v" || name == "region_consumption_q"
        || name == "region_consumption_v" || name == "region_gdp_v" || name == "region_gdp_q") {
        std::vector<i64> xq = E.f_exp_q.download(s), iq = E.s_imp_q.download(s), uq = E.s_U_q.download(s);
        std::vector<double> xv = E.f_exp_v.download(s), iv = E.s_imp_v.download(s), uv = E.s_U_v.download(s);
        std::vector<Flow> ex(E.nR, mkflow(0, 0.0)), im(E.nR, mkflow(0, 0.0)), co(E.nR, mkflow(0, 0.0));
        for (int a = 0; a < E.nA; ++a)
            if (E.h_agent_type[a] == 1) ex[E.h_agent_region[a]] = fadd(ex[E.h_agent_region[a]], mkflow(xq[a], xv[a]));
        for (int k = 0; k < E.nSt; ++k) { const int r = E.h_agent_region[E.h_st_agent[k]]; im[r] = fadd(im[r], mkflow(iq[k], iv[k])); }
        for (int k : E.h_consumer_st) { const int r = E.h_agent_region[E.h_st_agent[k]]; co[r] = fadd(co[r], mkflow(uq[k], uv[k])); }
        out.i.assign(E.nR, 0); out.f.assign(E.nR, 0.0);
        out.is_f64 = name.back() == 'v';
        for (int r = 0; r < E.nR; ++r) {
            Flow x;
            if (name.find("export") != std::string::npos) x = ex[r];
            else if (name.find("import") != std::string::npos) x = im[r];
            else if (name.find("consumption") != std::string::npos) x = co[r];
            else x = fsub(fadd(co[r], ex[r]), im[r]);  // Region::get_gdp, Region.cpp:52-54
            out.i[r] = x.q; out.f[r] = x.v;
        }
        return true;
    }
    return false;
}

}  // namespace acc

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
#define ACC_API_BEGIN try {
#define ACC_ON_DEVICE(e) acc::DeviceGuard device_guard__((e)->device)
#define ACC_API_END(failval)                                     \
    }                                                            \
    catch (const acc::EngineError& ex) {                         \
        acc::g_last_error = ex.msg;                              \
        return failval;                                          \
    }                                                            \
    catch (const std::exception& ex) {                           \
        acc::g_last_error = ex.what();                           \
        return failval;                                          \
    }


// ---- acc_observe: the output surface with selection (ArrayOutput::iterate, src/output/ArrayOutput.cpp:479-646) ----
namespace acc {
struct ObsName { const char* name; int id; };
static const ObsName k_agent_names[] = {
    {"business_connections", AV_BUSINESS_CONNECTIONS}, {"consumption_quantity", AV_CONSUMPTION_Q}, {"consumption_value", AV_CONSUMPTION_V},
    {"demand_quantity", AV_DEMAND_Q}, {"demand_value", AV_DEMAND_V}, {"input_flow_quantity", AV_INPUT_FLOW_Q}, {"input_flow_value", AV_INPUT_FLOW_V},
    {"storage_quantity", AV_STORAGE_Q}, {"storage_value", AV_STORAGE_V}};
static const ObsName k_firm_names[] = {
    {"communicated_possible_production_quantity", FV_COMMUNICATED_POSSIBLE_PRODUCTION_Q}, {"communicated_possible_production_value", FV_COMMUNICATED_POSSIBLE_PRODUCTION_V},
    {"desired_production_capacity", FV_DESIRED_PRODUCTION_CAPACITY}, {"direct_loss_quantity", FV_DIRECT_LOSS_Q}, {"direct_loss_value", FV_DIRECT_LOSS_V},
    {"expected_offer_price", FV_EXPECTED_OFFER_PRICE}, {"expected_production_quantity", FV_EXPECTED_PRODUCTION_Q}, {"expected_production_value", FV_EXPECTED_PRODUCTION_V},
    {"forcing", FV_FORCING}, {"incoming_demand_quantity", FV_INCOMING_DEMAND_Q}, {"incoming_demand_value", FV_INCOMING_DEMAND_V}, {"offer_price", FV_OFFER_PRICE},
    {"production_quantity", FV_PRODUCTION_Q}, {"production_value", FV_PRODUCTION_V}, {"production_capacity", FV_PRODUCTION_CAPACITY},
    {"possible_production_capacity", FV_POSSIBLE_PRODUCTION_CAPACITY}, {"tax", FV_TAX}, {"total_loss_quantity", FV_TOTAL_LOSS_Q}, {"total_loss_value", FV_TOTAL_LOSS_V},
    {"total_production_costs", FV_TOTAL_PRODUCTION_COSTS}, {"total_revenue", FV_TOTAL_REVENUE}, {"total_value_loss", FV_TOTAL_VALUE_LOSS},
    {"unit_production_costs", FV_UNIT_PRODUCTION_COSTS}};
static const ObsName k_consumer_names[] = {{"utility", CV_UTILITY}, {"local_optimal_utility", CV_LOCAL_OPTIMAL_UTILITY}};
static const ObsName k_storage_names[] = {
    {"business_connections", SV_BUSINESS_CONNECTIONS}, {"content_quantity", SV_CONTENT_Q}, {"content_value", SV_CONTENT_V}, {"demand_quantity", SV_DEMAND_Q},
    {"demand_value", SV_DEMAND_V}, {"desired_used_flow_quantity", SV_DESIRED_USED_FLOW_Q}, {"desired_used_flow_value", SV_DESIRED_USED_FLOW_V},
    {"expected_costs", SV_EXPECTED_COSTS}, {"input_flow_quantity", SV_INPUT_FLOW_Q}, {"input_flow_value", SV_INPUT_FLOW_V}, {"optimized_value", SV_OPTIMIZED_VALUE},
    {"possible_use_quantity", SV_POSSIBLE_USE_Q}, {"possible_use_value", SV_POSSIBLE_USE_V}, {"purchase_quantity", SV_PURCHASE_Q}, {"purchase_value", SV_PURCHASE_V},
    {"shipment_quantity", SV_SHIPMENT_Q}, {"shipment_value", SV_SHIPMENT_V}, {"storage_demand_quantity", SV_STORAGE_DEMAND_Q},
    {"storage_demand_value", SV_STORAGE_DEMAND_V}, {"total_transport_penalty", SV_TOTAL_TRANSPORT_PENALTY}, {"use_quantity", SV_USE_Q}, {"use_value", SV_USE_V},
    {"used_flow_quantity", SV_USED_FLOW_Q}, {"used_flow_value", SV_USED_FLOW_V}};
static const ObsName k_flow_names[] = {
    {"initial_flow_quantity", XV_INITIAL_FLOW_Q}, {"initial_flow_value", XV_INITIAL_FLOW_V}, {"demand_request_quantity", XV_DEMAND_REQUEST_Q},
    {"demand_request_value", XV_DEMAND_REQUEST_V}, {"flow_deficit", XV_FLOW_DEFICIT}, {"flow_mean_quantity", XV_FLOW_MEAN_Q}, {"flow_mean_value", XV_FLOW_MEAN_V},
    {"received_flow_quantity", XV_RECEIVED_FLOW_Q}, {"received_flow_value", XV_RECEIVED_FLOW_V}, {"sent_flow_quantity", XV_SENT_FLOW_Q},
    {"sent_flow_value", XV_SENT_FLOW_V}, {"total_flow_quantity", XV_TOTAL_FLOW_Q}, {"total_flow_value", XV_TOTAL_FLOW_V}, {"minimum_passage", XV_MINIMUM_PASSAGE}};
template <size_t N>
static int find_obs(const ObsName (&table)[N], const char* name) {
    for (size_t k = 0; k < N; ++k) if (std::strcmp(table[k].name, name) == 0) return table[k].id;
    return -1;
}

// copies a selection to the device (checked against `limit`); returns the device pointer or null for "all"
static const i64* stage_selection(acc_engine& E, const int64_t* idx, size_t n, i64 limit, size_t offset, const char* what) {
    if (n == 0) return nullptr;
    if (idx == nullptr) fail(std::string("acc_observe: null selection of ") + what);
    for (size_t k = 0; k < n; ++k) if (idx[k] < 0 || idx[k] >= limit) fail(std::string("acc_observe: selected ") + what + " out of range");
    if (E.obs_idx.n < offset + n) {
        DBuf<i64> bigger;
        bigger.alloc(std::max<size_t>(offset + n, 2 * E.obs_idx.n));
        E.obs_idx.free();
        E.obs_idx = bigger;
    }
    std::vector<i64> h(idx, idx + n);
    h2d(E.obs_idx.p + offset, h.data(), n * sizeof(i64), E.stream);
    dev_sync(E.stream);
    return E.obs_idx.p + offset;
}

static void observe(acc_engine& E, const std::string& group, const char* var, const int64_t* idx0, size_t n0, const int64_t* idx1, size_t n1, double* dst,
                    size_t cap, size_t* n) {
    stream_t s = E.stream;
    auto out_buffer = [&](size_t cells) {
        if (E.obs_out.n < cells) { E.obs_out.free(); E.obs_out.alloc(cells); }
        return E.obs_out.p;
    };
    auto deliver = [&](size_t cells) { d2h(dst, E.obs_out.p, cells * sizeof(double), s); };
    if (group == "firms" || group == "consumers") {
        const bool firms = group == "firms";
        int id = find_obs(k_agent_names, var);
        if (id < 0) id = firms ? find_obs(k_firm_names, var) : find_obs(k_consumer_names, var);
        if (id < 0) fail("acc_observe: unknown observable " + group + "/" + var);
        if (n1 != 0) fail("acc_observe: " + group + " take one selection (agents)");
        const size_t cells = n0 ? n0 : (size_t)E.nA;
        *n = cells;
        if (dst == nullptr) return;
        if (cap < cells) fail("acc_observe: destination too small");
        const i64* sel = stage_selection(E, idx0, n0, E.nA, 0, "agent");
        if (id == FV_INCOMING_DEMAND_Q || id == FV_INCOMING_DEMAND_V) {
            if (E.obs_inc_q.n == 0) { E.obs_inc_q.alloc(E.nA); E.obs_inc_v.alloc(E.nA); }
            E.obs_inc_q.zero(s); E.obs_inc_v.zero(s);
            FirmIncomingDemand k{E.edge_view(), E.firm_ids.p, E.obs_inc_q.p, E.obs_inc_v.p};
            E.timed("firm_incoming_demand", [&] { launch_block(s, E.nF, k); });
        }
        ObserveAgents k{id, firms ? 1 : 0, sel, E.firm_view(), E.storage_view(), E.mp, E.in_off.p, E.s_dem_q.p, E.s_dem_v.p, E.obs_inc_q.p, E.obs_inc_v.p,
                        E.mp.utilitarian ? E.c_utility.p : nullptr, out_buffer(cells)};
        E.timed("observe_agents", [&] { launch_thread(s, (int)cells, k); });
        deliver(cells);
        return;
    }
    if (group == "storages") {
        const int id = find_obs(k_storage_names, var);
        if (id < 0) fail("acc_observe: unknown observable storages/" + std::string(var));
        if (n0 != 0) fail("acc_observe: storages cannot be selected by sector (nor can the reference's output, ArrayOutput.cpp:575); select agents with idx1");
        const size_t n_sel = n1 ? n1 : (size_t)E.nA;
        const size_t cells = (size_t)E.nS * n_sel;
        *n = cells;
        if (dst == nullptr) return;
        if (cap < cells) fail("acc_observe: destination too small");
        if (cells >= (size_t)0x7fffffff) fail("acc_observe: selection too large");
        if ((id == SV_SHIPMENT_Q || id == SV_SHIPMENT_V) && E.ship_step < 0) fail("acc_observe: no timestep has run yet");
        const i64* sel = stage_selection(E, idx1, n1, E.nA, 0, "agent");
        ObserveStorages k{id, (int)n_sel, E.nA, sel, E.sto_of.p, E.storage_view(), E.edge_view(), E.mp, E.ship_step, E.in_off.p, E.in_edge.p, E.s_dem_q.p,
                          E.s_dem_v.p, E.s_exp_costs.p, E.s_opt_val.p, E.s_tpen.p, out_buffer(cells)};
        E.timed("observe_storages", [&] { launch_thread(s, (int)cells, k); });
        deliver(cells);
        return;
    }
    if (group == "flows") {
        const int id = find_obs(k_flow_names, var);
        if (id < 0) fail("acc_observe: unknown observable flows/" + std::string(var));
        if (n0 != 0 && n1 != 0) fail("acc_observe: selection on both source agent and target agent not supported (as in the reference, ArrayOutput.cpp:469-471)");
        if (E.partitioned) fail("acc_observe: flows of a region-partitioned engine are read per part with acc_read_state");
        const size_t rows = n0 ? n0 : (size_t)E.nA, cols = n1 ? n1 : (size_t)E.nA;
        const size_t cells = rows * cols;
        *n = cells;
        if (dst == nullptr) return;
        if (cap < cells) fail("acc_observe: destination too small");
        if (cells >= (size_t)0x7fffffff) fail("acc_observe: selection too large: select source or target agents");
        if (E.ship_step < 0) fail("acc_observe: no timestep has run yet");
        double* out = out_buffer(cells);
        ObsFill fill{out};
        E.timed("observe_fill", [&] { launch_thread(s, (int)cells, fill); });
        if (n1 == 0) {
            const i64* sel = stage_selection(E, idx0, n0, E.nA, 0, "agent");
            ObserveFlowsFrom k{id, (int)cols, sel, E.edge_view(), E.ship_step, E.a_type.p, E.e_to.p, out};
            E.timed("observe_flows", [&] { launch_block(s, (int)rows, k); });
        } else {
            const i64* sel = stage_selection(E, idx1, n1, E.nA, 0, "agent");
            ObserveFlowsTo k{id, (int)cols, E.max_storages, sel, E.edge_view(), E.ship_step, E.a_st_off.p, E.a_st_list.p, E.in_off.p, E.in_edge.p, E.in_seller.p,
                             E.row_firm.p, out};
            E.timed("observe_flows", [&] { launch_block(s, (int)(cols * (size_t)E.max_storages), k); });
        }
        deliver(cells);
        return;
    }
    if (group == "sectors" || group == "regions") {
        // a few dozen to a few hundred cells: folded on the host in the reference's serial order from per-agent / per-storage arrays
        const bool sectors = group == "sectors";
        const std::string v = var;
        std::string state;
        char kind = 'v';
        if (sectors) {
            if (v == "total_production_quantity") { state = "sector_production_q"; kind = 'q'; }
            else if (v == "total_production_value") state = "sector_production_v";
            else if (v == "total_demand_quantity") { state = "sector_demand_q"; kind = 'q'; }
            else if (v == "total_demand_value") state = "sector_demand_v";
            else if (v == "offer_price") { state = "sector_production_q"; kind = 'P'; }
        } else {
            static const char* names[] = {"import", "export", "consumption", "gdp"};
            for (const char* nm : names) {
                if (v == std::string(nm) + "_quantity") { state = std::string("region_") + nm + "_q"; kind = 'q'; }
                if (v == std::string(nm) + "_value") state = std::string("region_") + nm + "_v";
            }
        }
        // GeoEntity::observe on a region (include/model/GeoEntity.h:64-80): the transport chain links registered AT the region.  Routes
        // of a snapshot lead over connections, ports and sea areas, never through a region, so a region holds no link: Flow(0.0)
        const bool no_links = !sectors && (v == "total_flow_quantity" || v == "total_flow_value" || v == "total_outflow_quantity" || v == "total_outflow_value");
        if (state.empty() && !no_links) fail("acc_observe: unknown observable " + group + "/" + v);
        if (n1 != 0) fail("acc_observe: " + group + " take one selection");
        const i64 limit = sectors ? E.nS : E.nR;
        const size_t cells = n0 ? n0 : (size_t)limit;
        *n = cells;
        if (dst == nullptr) return;
        if (cap < cells) fail("acc_observe: destination too small");
        for (size_t k = 0; k < n0; ++k) if (idx0[k] < 0 || idx0[k] >= limit) fail("acc_observe: selection out of range");
        if (no_links) { for (size_t k = 0; k < cells; ++k) dst[k] = 0.0; return; }
        StateArray a, b;
        if (!read_state(E, state, a)) fail("acc_observe: internal mapping error");
        if (kind == 'P' && !read_state(E, "sector_production_v", b)) fail("acc_observe: internal mapping error");
        for (size_t k = 0; k < cells; ++k) {
            const size_t j = n0 ? (size_t)idx0[k] : k;
            // the consumers' sector produces and sells nothing: its sums are the empty Flow the reference reports too
            if (kind == 'q') dst[k] = qf(a.i[j]);
            else if (kind == 'P') dst[k] = pf(flow_price(mkflow(a.i[j], b.f[j])));  // Sector.h:78-80: total_production_X().get_price()
            else dst[k] = a.f[j];
        }
        return;
    }
    fail("acc_observe: unknown group '" + group + "' (firms, consumers, sectors, regions, storages, flows)");
}
}  // namespace acc

extern "C" int acc_observe(acc_engine* e, const char* group, const char* variable, const int64_t* idx0, size_t n0, const int64_t* idx1, size_t n1, double* dst,
                           size_t cap, size_t* n) {
    ACC_API_BEGIN
    if (e == nullptr || group == nullptr || variable == nullptr || n == nullptr) fail("acc_observe: null argument");
    ACC_ON_DEVICE(e);
    acc::observe(*e, group, variable, idx0, n0, idx1, n1, dst, cap, n);
    return 0;
    ACC_API_END(-1)
}


// ---- acc_ingest_table: threshold filter + clean_network fix point on the device (csrc/ingest.cuh) ----
extern "C" int acc_ingest_table(int device, int64_t n_agents, int64_t n_sectors, const int64_t* agent_type, const int64_t* agent_sector, const float* flows,
                                double threshold, double time_factor, int no_self_supply, uint8_t* agent_alive, int64_t* n_edges, int64_t* edge_from,
                                int64_t* edge_to, int64_t* edge_zq, size_t cap_edges, int64_t* sweeps, int64_t* removed) {
    ACC_API_BEGIN
    if (agent_type == nullptr || agent_sector == nullptr || flows == nullptr || n_edges == nullptr) fail("acc_ingest_table: null argument");
    if (n_agents <= 0 || n_sectors <= 0 || n_agents * n_agents >= (int64_t)0x7fffffff) fail("acc_ingest_table: table size out of range");
#ifndef ACC_HOST_EMU
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) fail("acc_ingest_table: no CUDA device available (this library has no CPU path)");
    if (device < 0 || device >= count) fail("acc_ingest_table: no such CUDA device");
#endif
    acc::DeviceGuard device_guard__(device);
    stream_t s = nullptr;
#ifndef ACC_HOST_EMU
    ACC_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    struct StreamOwner { stream_t s; ~StreamOwner() { cudaStreamDestroy(s); } } stream_owner{s};
#endif
    const i64 nA = n_agents;
    const int nS = (int)n_sectors;
    std::vector<int> h_type(nA), h_sector(nA);
    for (i64 a = 0; a < nA; ++a) {
        h_type[a] = (int)agent_type[a]; h_sector[a] = (int)agent_sector[a];
        if (h_sector[a] < 0 || h_sector[a] >= nS) fail("acc_ingest_table: agent sector out of range");
    }
    DBuf<int> d_type, d_sector, d_nout, d_nst, d_cnt;
    DBuf<float> d_table;
    DBuf<i64> d_Z, d_X, d_in, d_found, d_rows, d_off, d_from, d_to, d_zq;
    DBuf<unsigned char> d_alive, d_self;
    struct Release {
        std::vector<std::function<void()>> f;
        ~Release() { for (auto& g : f) g(); }
    } rel;
    auto own = [&](auto& b) { rel.f.push_back([&b] { b.free(); }); };
    own(d_type); own(d_sector); own(d_nout); own(d_nst); own(d_cnt); own(d_table); own(d_Z); own(d_X); own(d_in); own(d_found); own(d_rows); own(d_off);
    own(d_from); own(d_to); own(d_zq); own(d_alive); own(d_self);
    d_type.upload(h_type, s); d_sector.upload(h_sector, s);
    d_table.alloc((size_t)(nA * nA)); h2d(d_table.p, flows, (size_t)(nA * nA) * sizeof(float), s);
    d_Z.alloc((size_t)(nA * nA)); d_X.alloc(nA); d_in.alloc(nA); d_nout.alloc(nA); d_nst.alloc(nA); d_cnt.alloc((size_t)nS * nA); d_alive.alloc(nA); d_self.alloc(nA);
    d_found.alloc(1); d_rows.alloc(nA);
    IngestView v{nA, nS, d_type.p, d_sector.p, d_Z.p, d_X.p, d_in.p, d_nout.p, d_nst.p, d_cnt.p, d_alive.p, d_self.p};
    {
        IngestFilter k{v, d_table.p, threshold, time_factor, mkq(threshold * time_factor), no_self_supply};
        launch_thread(s, (int)(nA * nA), k);
        IngestStats st{v};
        launch_block(s, (int)nA, st);
    }
    // the reference's sweeps, replayed exactly: find the first removable agent at or after the cursor, take it out, go on behind it
    i64 n_sweeps = 0, n_removed = 0;
    for (;;) {
        ++n_sweeps;
        i64 cursor = 0, in_sweep = 0;
        while (cursor < nA) {
            i64 found = nA;
            h2d(d_found.p, &found, sizeof found, s);
            IngestFindFirst ff{v, cursor, d_found.p};
            launch_thread(s, (int)(nA - cursor), ff);
            d2h(&found, d_found.p, sizeof found, s);
            if (found >= nA) break;
            IngestRemove rm{v, found};
            launch_block(s, 1, rm);
            ++in_sweep;
            cursor = found + 1;
        }
        n_removed += in_sweep;
        if (in_sweep == 0) break;
    }
    {
        IngestCount k{v, d_rows.p};
        launch_block(s, (int)nA, k);
    }
    std::vector<i64> rows = d_rows.download(s), off(nA + 1, 0);
    for (i64 a = 0; a < nA; ++a) off[a + 1] = off[a] + rows[a];
    *n_edges = off[nA];
    if (sweeps) *sweeps = n_sweeps;
    if (removed) *removed = n_removed;
    if (agent_alive != nullptr) {
        std::vector<unsigned char> al = d_alive.download(s);
        std::memcpy(agent_alive, al.data(), (size_t)nA);
    }
    if (edge_from != nullptr || edge_to != nullptr || edge_zq != nullptr) {
        if (edge_from == nullptr || edge_to == nullptr || edge_zq == nullptr) fail("acc_ingest_table: pass all three connection arrays or none");
        if (cap_edges < (size_t)off[nA]) fail("acc_ingest_table: connection arrays too small (call with NULL arrays first to learn the count)");
        if (off[nA] == 0) fail("No economic agents present");  // ModelInitializer.cpp:296-298
        d_off.upload(off, s);
        d_from.alloc((size_t)off[nA]); d_to.alloc((size_t)off[nA]); d_zq.alloc((size_t)off[nA]);
        IngestEmit k{v, d_off.p, d_from.p, d_to.p, d_zq.p};
        launch_block(s, (int)nA, k);
        d2h(edge_from, d_from.p, (size_t)off[nA] * sizeof(i64), s);
        d2h(edge_to, d_to.p, (size_t)off[nA] * sizeof(i64), s);
        d2h(edge_zq, d_zq.p, (size_t)off[nA] * sizeof(i64), s);
    }
    return 0;
    ACC_API_END(-1)
}

// ---- state checkpoint: blob header and the visitors over acc_engine::visit_state ----
namespace {
struct StateHeader {
    char magic[8];
    int64_t n_agents, n_storages, n_edges, planes, step, reg, forcing_used, payload_bytes, ship_step, event_capacity;
};
struct SizeVisitor {
    size_t bytes = 0;
    template <class T> void operator()(acc::DBuf<T>& b) { bytes += b.n * sizeof(T); }
};
struct SaveVisitor {
    unsigned char* dst; acc::stream_t s;
    template <class T> void operator()(acc::DBuf<T>& b) { if (b.n) { acc::d2h(dst, b.p, b.n * sizeof(T), s); dst += b.n * sizeof(T); } }
};
struct LoadVisitor {
    const unsigned char* src; acc::stream_t s;
    template <class T> void operator()(acc::DBuf<T>& b) { if (b.n) { acc::h2d(b.p, src, b.n * sizeof(T), s); src += b.n * sizeof(T); } }
};
}  // namespace

namespace {
struct FreeVisitor {
    template <class T> void operator()(acc::DBuf<T>& b) { b.free(); }
};
struct BorrowVisitor {
    template <class T> void operator()(acc::DBuf<T>& b) { b.borrowed = true; }
};
struct CloneVisitor {  // the member gets its own copy of a state array, filled with the parent's current contents
    acc::stream_t s;
    template <class T> void operator()(acc::DBuf<T>& b) {
        const T* src = b.p;
        const size_t n = b.n;
        b.p = nullptr; b.n = 0; b.borrowed = false;
        if (n == 0) return;
        b.alloc(n);
#ifdef ACC_HOST_EMU
        std::memcpy(b.p, src, n * sizeof(T));
#else
        ACC_CUDA(cudaMemcpyAsync(b.p, src, n * sizeof(T), cudaMemcpyDeviceToDevice, s));
#endif
    }
};
void release_engine(acc_engine* e) {
    FreeVisitor fv;
    e->visit_all(fv);
    acc_engine* parent = e->parent;
    delete e;
    if (parent != nullptr && --parent->members == 0 && parent->zombie) release_engine(parent);
}
}  // namespace

extern "C" {

const char* acc_last_error(void) { return acc::g_last_error.c_str(); }

acc_engine* acc_create(const acc_network* net, const acc_params* params, int device) { return acc_create_partition(net, params, device, nullptr); }

acc_engine* acc_create_partition(const acc_network* net, const acc_params* params, int device, const uint8_t* agent_owned) {
    acc_engine* E = nullptr;
    try {
        if (net == nullptr || params == nullptr) fail("acc_create: null argument");
        E = new acc_engine;
        E->device = device;
#ifndef ACC_HOST_EMU
        int count = 0;
        if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) fail("acc_create: no CUDA device available (this library has no CPU path)");
        if (device < 0 || device >= count) fail("acc_create: no such CUDA device");
#endif
        acc::DeviceGuard device_guard__(device);
#ifndef ACC_HOST_EMU
        ACC_CUDA(cudaStreamCreateWithFlags(&E->stream, cudaStreamNonBlocking));
        {   // the consumers' side stream outranks the main one: their few CTAs of sequential work must become resident when
            // the first slots free up, not after the sellers' 4 862 CTAs launched beside them have drained
            int prio_low = 0, prio_high = 0;
            ACC_CUDA(cudaDeviceGetStreamPriorityRange(&prio_low, &prio_high));
            ACC_CUDA(cudaStreamCreateWithPriority(&E->side, cudaStreamNonBlocking, prio_high));
        }
        ACC_CUDA(cudaEventCreateWithFlags(&E->ev_fork, cudaEventDisableTiming));
        ACC_CUDA(cudaEventCreateWithFlags(&E->ev_join, cudaEventDisableTiming));
#endif
        build(*E, *net, *params, agent_owned);
        return E;
    } catch (const acc::EngineError& ex) {
        acc::g_last_error = ex.msg;
    } catch (const std::exception& ex) {
        acc::g_last_error = ex.what();
    }
    delete E;
    return nullptr;
}

void acc_destroy(acc_engine* e) {
    if (e == nullptr) return;
    acc::DeviceGuard* guard = nullptr;
    try { guard = new acc::DeviceGuard(e->device); } catch (...) {}
#ifndef ACC_HOST_EMU
    cudaStreamSynchronize(e->stream);
    if (e->side) cudaStreamSynchronize(e->side);
#endif
    if (e->members > 0) e->zombie = true;  // ensemble members still read this engine's tables: released with the last of them
    else release_engine(e);
    delete guard;
}

/* acc_fork: an ensemble member.  The new engine SHARES the parent's immutable tables (topology, Z*, delays, the buyer-side lists:
 * everything acc_save_state does not name) and owns a copy of the state, taken as the parent has it now.  Members and parent
 * are stepped, forced and read independently (each on its own stream); many scenarios of one network cost one topology. */
acc_engine* acc_fork(acc_engine* parent) {
    acc_engine* E = nullptr;
    try {
        if (parent == nullptr) fail("acc_fork: null engine");
        if (parent->partitioned) fail("acc_fork: a part of a region-partitioned run cannot be forked");
        acc::DeviceGuard device_guard__(parent->device);
        acc::dev_sync(parent->stream);
        if (parent->side) acc::dev_sync(parent->side);
        acc_engine* root = parent->parent != nullptr ? parent->parent : parent;  // members of a member share the same tables
        E = new acc_engine(*parent);
        E->stream = nullptr; E->side = nullptr;
#ifndef ACC_HOST_EMU
        E->ev_fork = nullptr; E->ev_join = nullptr; E->comm = nullptr;
#endif
        E->clocks.clear(); E->launches = 0; E->profile = false;
        E->parent = nullptr; E->members = 0; E->zombie = false;
        BorrowVisitor bv;
        E->visit_all(bv);
#ifndef ACC_HOST_EMU
        ACC_CUDA(cudaStreamCreateWithFlags(&E->stream, cudaStreamNonBlocking));
        int prio_low = 0, prio_high = 0;
        ACC_CUDA(cudaDeviceGetStreamPriorityRange(&prio_low, &prio_high));
        ACC_CUDA(cudaStreamCreateWithPriority(&E->side, cudaStreamNonBlocking, prio_high));
        ACC_CUDA(cudaEventCreateWithFlags(&E->ev_fork, cudaEventDisableTiming));
        ACC_CUDA(cudaEventCreateWithFlags(&E->ev_join, cudaEventDisableTiming));
#endif
        CloneVisitor cv{E->stream};
        E->visit_state(cv);
        cv(E->pubT); cv(E->fatal);  // derived / scratch state that changes with the steps
        cv(E->c_nx_q); cv(E->c_nx_v); cv(E->c_nx_util); cv(E->c_nx_ret); cv(E->e_sort); cv(E->f_sorted);
        E->ahead_step = -1;         // the member repeats the consumers' solve of its next step on its own copy of the state
        for (acc::DBuf<double>* b : {&E->obs_out, &E->obs_inc_v}) { b->p = nullptr; b->n = 0; b->borrowed = false; }
        for (acc::DBuf<i64>* b : {&E->obs_idx, &E->obs_inc_q}) { b->p = nullptr; b->n = 0; b->borrowed = false; }
        acc::dev_sync(E->stream);
        E->parent = root;
        root->members += 1;
        return E;
    } catch (const acc::EngineError& ex) {
        acc::g_last_error = ex.msg;
    } catch (const std::exception& ex) {
        acc::g_last_error = ex.what();
    }
    if (E != nullptr) {
        FreeVisitor fv;
        E->visit_all(fv);
        delete E;
    }
    return nullptr;
}

int acc_set_agent_forcing(acc_engine* e, const double* lambda, size_t n) {
    ACC_API_BEGIN
    if (e == nullptr || lambda == nullptr) fail("acc_set_agent_forcing: null argument");
    ACC_ON_DEVICE(e);
    if (n != (size_t)e->nA) fail("acc_set_agent_forcing: expected one entry per agent");
    h2d(e->a_lambda.p, lambda, n * sizeof(double), e->stream);
    return 0;
    ACC_API_END(-1)
}

int acc_set_entity_passage(acc_engine* e, const double* nu, size_t n) {
    ACC_API_BEGIN
    if (e == nullptr || nu == nullptr) fail("acc_set_entity_passage: null argument");
    ACC_ON_DEVICE(e);
    if (n != (size_t)e->nEnt) fail("acc_set_entity_passage: expected one entry per geo entity (per region pair for a snapshot without routes)");
    if (!e->forcing_used) {
        bool any = false;
        for (size_t k = 0; k < n; ++k) any = any || nu[k] >= 0.0;
        if (!any) return 0;  // nothing forced yet: overflow stays identically zero
        e->ensure_forcing_state();
    }
    h2d(e->nu.p, nu, n * sizeof(double), e->stream);
    return 0;
    ACC_API_END(-1)
}

int acc_phase(acc_engine* e, int phase) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_phase: null engine");
    ACC_ON_DEVICE(e);
    switch (phase) {
        case ACC_PHASE_CONSUMPTION_AND_PRODUCTION: e->phase_cp(); break;
        case ACC_PHASE_EXPECTATION: e->phase_exp(); break;
        case ACC_PHASE_PURCHASE: e->phase_pur(); break;
        case ACC_PHASE_TICK: e->tick(); break;
        default: fail("acc_phase: unknown phase");
    }
    return 0;
    ACC_API_END(-1)
}

int acc_set_halo(acc_engine* e, const int64_t* send_edges, size_t n_send, const int64_t* recv_edges, size_t n_recv, void* flow_send, void* flow_recv,
                 void* demand_send, void* demand_recv, void* seller_table) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_set_halo: null engine");
    ACC_ON_DEVICE(e);
    if (!e->partitioned) fail("acc_set_halo: the engine was not created with acc_create_partition");
    if ((n_send && send_edges == nullptr) || (n_recv && recv_edges == nullptr)) fail("acc_set_halo: null list");
    const bool own_buffers = flow_send == nullptr && flow_recv == nullptr && demand_send == nullptr && demand_recv == nullptr && seller_table == nullptr;
    if (own_buffers) {  // the engine keeps the exchange buffers itself (the NCCL plane of acc_step_partitioned needs no caller's)
        e->own_flow_send.alloc(2 * n_send); e->own_dem_recv.alloc(2 * n_send); e->own_flow_recv.alloc(2 * n_recv); e->own_dem_send.alloc(2 * n_recv);
        e->own_sellers.alloc((size_t)e->nA * ACC_SELLER_WORDS);
        for (DBuf<i64>* b : {&e->own_flow_send, &e->own_dem_recv, &e->own_flow_recv, &e->own_dem_send, &e->own_sellers}) b->zero(e->stream);
        flow_send = e->own_flow_send.p; flow_recv = e->own_flow_recv.p; demand_send = e->own_dem_send.p; demand_recv = e->own_dem_recv.p;
        seller_table = e->own_sellers.p;
    }
    if ((n_send && (flow_send == nullptr || demand_recv == nullptr)) || (n_recv && (flow_recv == nullptr || demand_send == nullptr)) || seller_table == nullptr)
        fail("acc_set_halo: null exchange buffer (pass all five, or none to let the engine allocate them)");
    if (n_send >= (size_t)0x7fffffff || n_recv >= (size_t)0x7fffffff) fail("acc_set_halo: halo too large");
    for (size_t k = 0; k < n_send; ++k) if (send_edges[k] < 0 || send_edges[k] >= e->nE) fail("acc_set_halo: send connection out of range");
    for (size_t k = 0; k < n_recv; ++k) if (recv_edges[k] < 0 || recv_edges[k] >= e->nE) fail("acc_set_halo: receive connection out of range");
    e->halo_send.upload(std::vector<i64>(send_edges, send_edges + n_send), e->stream);
    e->halo_recv.upload(std::vector<i64>(recv_edges, recv_edges + n_recv), e->stream);
    dev_sync(e->stream);
    e->x_flow_send = static_cast<i64*>(flow_send); e->x_flow_recv = static_cast<i64*>(flow_recv);
    e->x_dem_send = static_cast<i64*>(demand_send); e->x_dem_recv = static_cast<i64*>(demand_recv);
    e->x_sellers = static_cast<i64*>(seller_table);
    return 0;
    ACC_API_END(-1)
}

int acc_comm_unique_id(void* id) {
    ACC_API_BEGIN
    if (id == nullptr) fail("acc_comm_unique_id: null argument");
#ifdef ACC_HOST_EMU
    fail("acc_comm_unique_id: the emulation has no NCCL");
#else
    NcclApi::unique_id u;
    ACC_NCCL(NcclApi::get().GetUniqueId(&u));
    std::memcpy(id, &u, sizeof u);
#endif
    return 0;
    ACC_API_END(-1)
}

int acc_comm_init(acc_engine* e, const void* id, int rank, int world, const int32_t* agent_part, const int64_t* send_counts, const int64_t* recv_counts) {
    ACC_API_BEGIN
    if (e == nullptr || id == nullptr || agent_part == nullptr || send_counts == nullptr || recv_counts == nullptr) fail("acc_comm_init: null argument");
    ACC_ON_DEVICE(e);
    if (!e->partitioned || e->x_sellers == nullptr) fail("acc_comm_init: create the engine with acc_create_partition and call acc_set_halo first");
    if (world < 1 || rank < 0 || rank >= world) fail("acc_comm_init: rank / world out of range");
#ifdef ACC_HOST_EMU
    fail("acc_comm_init: the emulation has no NCCL; drive the stages with acc_stage");
#else
    if (e->comm != nullptr) fail("acc_comm_init: the engine has a communicator already");
    // how the halo lists split over the peers (they are grouped by peer, acc_set_halo)
    e->comm_send_cnt.assign(send_counts, send_counts + world); e->comm_recv_cnt.assign(recv_counts, recv_counts + world);
    e->comm_send_off.assign(world + 1, 0); e->comm_recv_off.assign(world + 1, 0);
    for (int p = 0; p < world; ++p) {
        if (send_counts[p] < 0 || recv_counts[p] < 0 || (p == rank && (send_counts[p] || recv_counts[p]))) fail("acc_comm_init: bad halo counts");
        e->comm_send_off[p + 1] = e->comm_send_off[p] + send_counts[p];
        e->comm_recv_off[p + 1] = e->comm_recv_off[p] + recv_counts[p];
    }
    if ((size_t)e->comm_send_off[world] != e->halo_send.n || (size_t)e->comm_recv_off[world] != e->halo_recv.n)
        fail("acc_comm_init: the per-peer counts do not add up to the halo lists of acc_set_halo");
    // the firms of every part, in agent order: rows of the sellers' all-gather
    std::vector<int> part_off(world + 1, 0), part_firms;
    for (int p = 0; p < world; ++p) {
        for (int a = 0; a < e->nA; ++a) {
            if (agent_part[a] < 0 || agent_part[a] >= world) fail("acc_comm_init: agent_part out of range");
            if (agent_part[a] == p && e->h_agent_type[a] == 1) part_firms.push_back(a);
        }
        part_off[p + 1] = (int)part_firms.size();
        e->comm_max_rows = std::max(e->comm_max_rows, part_off[p + 1] - part_off[p]);
    }
    if (part_off[rank + 1] - part_off[rank] != e->nF) fail("acc_comm_init: agent_part does not match the ownership mask the engine was created with");
    e->comm_part_firms.upload(part_firms, e->stream); e->comm_part_off.upload(part_off, e->stream);
    e->comm_rows_send.alloc((size_t)e->comm_max_rows * ACC_SELLER_WORDS); e->comm_rows_send.zero(e->stream);
    e->comm_rows_recv.alloc((size_t)world * e->comm_max_rows * ACC_SELLER_WORDS); e->comm_rows_recv.zero(e->stream);
    dev_sync(e->stream);
    NcclApi::unique_id u;
    std::memcpy(&u, id, sizeof u);
    e->comm_rank = rank; e->comm_world = world;
    ACC_NCCL(NcclApi::get().CommInitRank(&e->comm, world, u, rank));
#endif
    return 0;
    ACC_API_END(-1)
}

int acc_comm_enable_peer_stores(acc_engine* e) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_comm_enable_peer_stores: null engine");
    ACC_ON_DEVICE(e);
#ifdef ACC_HOST_EMU
    fail("acc_comm_enable_peer_stores: the emulation has no peer memory");
#else
    if (e->comm == nullptr) fail("acc_comm_enable_peer_stores: call acc_comm_init first");
    if (e->peer_stores) return 0;
    if (e->step != 0) fail("acc_comm_enable_peer_stores: only before the first step");
    if (e->x_flow_recv != e->own_flow_recv.p || e->x_dem_recv != e->own_dem_recv.p)
        fail("acc_comm_enable_peer_stores: the receive buffers must be the engine's own (pass NULL buffers to acc_set_halo)");
    const int world = e->comm_world, rank = e->comm_rank;
    if (world > ACC_MAX_PARTS) fail("acc_comm_enable_peer_stores: more parts than ACC_MAX_PARTS");
    NcclApi& nc = NcclApi::get();
    stream_t s = e->stream;
    // every part tells the others where its two receive buffers are (IPC handles) and how its halo lists split over the parts:
    // the k-th entry this part sends to part p is the k-th entry p receives from it (both lists are in global order), so its
    // position in p's buffer is p's offset for this part + k -- no per-connection id has to travel
    struct Hello { i64 send_cnt[ACC_MAX_PARTS], recv_cnt[ACC_MAX_PARTS]; cudaIpcMemHandle_t h[5]; };
    static_assert(sizeof(Hello) % 8 == 0, "exchanged as int64 words");
    Hello mine;
    std::memset(&mine, 0, sizeof mine);
    for (int p = 0; p < world; ++p) { mine.send_cnt[p] = e->comm_send_cnt[p]; mine.recv_cnt[p] = e->comm_recv_cnt[p]; }
    e->barrier_flags.alloc(ACC_MAX_PARTS); e->barrier_flags.zero(s);
    e->barrier_timeout.alloc(1); e->barrier_timeout.zero(s);
    dev_sync(s);
    void* arrays[5] = {e->own_flow_recv.p, e->own_dem_recv.p, e->barrier_flags.p, e->pub.p, e->pubT.p};
    for (int k = 0; k < 5; ++k) ACC_CUDA(cudaIpcGetMemHandle(&mine.h[k], arrays[k]));
    DBuf<i64> d_hello, d_all;
    const size_t words = sizeof(Hello) / 8;
    d_hello.alloc(words); d_all.alloc(words * world);
    h2d(d_hello.p, &mine, sizeof mine, s);
    ACC_NCCL(nc.AllGather(d_hello.p, d_all.p, words, NcclApi::Int64, e->comm, s));
    std::vector<Hello> all(world);
    d2h(all.data(), d_all.p, sizeof(Hello) * world, s);
    d_hello.free(); d_all.free();
    // (the only collective of this call is behind us: a part that cannot map its peers fails alone, nobody hangs)
    PeerTable pt;
    std::memset(&pt, 0, sizeof pt);
    for (int p = 0; p < world; ++p) {
        if (all[p].recv_cnt[rank] != e->comm_send_cnt[p] || all[p].send_cnt[rank] != e->comm_recv_cnt[p])
            fail("acc_comm_enable_peer_stores: the parts' halo lists do not match");
        void* ptr[5];
        for (int k = 0; k < 5; ++k) {
            if (p == rank) { ptr[k] = arrays[k]; continue; }
            ACC_CUDA(cudaIpcOpenMemHandle(&ptr[k], all[p].h[k], cudaIpcMemLazyEnablePeerAccess));
            e->ipc_opened.push_back(ptr[k]);
        }
        pt.flow_recv[p] = static_cast<i64*>(ptr[0]); pt.dem_recv[p] = static_cast<i64*>(ptr[1]);
        pt.flags[p] = static_cast<unsigned long long*>(ptr[2]);
        pt.pub[p] = static_cast<SellerPub*>(ptr[3]); pt.pubT[p] = static_cast<double*>(ptr[4]);
    }
    std::vector<i64> hs = e->halo_send.download(s), hr = e->halo_recv.download(s);
    std::vector<signed char> xp((size_t)e->nE, (signed char)-1);
    std::vector<int> xr((size_t)e->nE, 0);
    for (int p = 0; p < world; ++p) {
        if (p == rank) continue;
        i64 their_recv_off = 0, their_send_off = 0;  // where this part's entries start in p's receive buffers
        for (int q = 0; q < rank; ++q) { their_recv_off += all[p].recv_cnt[q]; their_send_off += all[p].send_cnt[q]; }
        for (i64 k = 0; k < e->comm_send_cnt[p]; ++k) {  // shipments to p: p's flow_recv, in p's receive order
            const i64 edge = hs[e->comm_send_off[p] + k];
            xp[edge] = (signed char)p; xr[edge] = (int)(their_recv_off + k);
        }
        for (i64 k = 0; k < e->comm_recv_cnt[p]; ++k) {  // demand requests to p: p's dem_recv, in p's send order
            const i64 edge = hr[e->comm_recv_off[p] + k];
            xp[edge] = (signed char)p; xr[edge] = (int)(their_send_off + k);
        }
    }
    e->x_peer.upload(xp, s); e->x_rid.upload(xr, s);
    e->peer_table.upload(std::vector<PeerTable>(1, pt), s);
    e->barrier_word.alloc(2); e->barrier_word.zero(s);
    dev_sync(s);
    e->h_peers = pt;
    e->flag_barrier = std::getenv("ACC_NCCL_BARRIER") == nullptr;  // ACC_NCCL_BARRIER=1: one-word all-reduces instead (for comparison)
    e->peer_sellers = std::getenv("ACC_NCCL_SELLERS") == nullptr;  // ACC_NCCL_SELLERS=1: the sellers' records by ncclAllGather instead (for comparison)
    e->peer_stores = true;
#endif
    return 0;
    ACC_API_END(-1)
}

int acc_comm_disable_peer_stores(acc_engine* e) {
    if (e == nullptr) return -1;
    e->peer_stores = false;
    return 0;
}

int acc_step_partitioned(acc_engine* e) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_step_partitioned: null engine");
    ACC_ON_DEVICE(e);
    e->step_partitioned();
    return 0;
    ACC_API_END(-1)
}

int64_t acc_collective_count(const acc_engine* e) { return e ? e->collectives : -1; }

int acc_stage(acc_engine* e, int stage) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_stage: null engine");
    ACC_ON_DEVICE(e);
    if (e->partitioned && e->x_sellers == nullptr) fail("acc_stage: call acc_set_halo first");
    switch (stage) {
        case ACC_STAGE_SELL: e->stage_sell(); break;
        case ACC_STAGE_DELIVER: e->stage_deliver(); break;
        case ACC_STAGE_PURCHASE: e->stage_purchase(); break;
        case ACC_STAGE_CLOSE: e->stage_close(); break;
        default: fail("acc_stage: unknown stage");
    }
    return 0;
    ACC_API_END(-1)
}

int acc_step(acc_engine* e) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_step: null engine");
    ACC_ON_DEVICE(e);
    e->require_single("acc_step");
    e->phase_cp();
    e->phase_exp();
    e->phase_pur();
    e->tick();
    return 0;
    ACC_API_END(-1)
}

int acc_sync(acc_engine* e) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_sync: null engine");
    ACC_ON_DEVICE(e);
    dev_sync(e->stream);
#ifndef ACC_HOST_EMU
    if (e->side) dev_sync(e->side);  // the consumers' solve of the next step, started behind this step's delivery pass
    if (e->flag_barrier && e->barrier_timeout.n) {
        int late = 0;
        d2h(&late, e->barrier_timeout.p, sizeof(int), e->stream);
        if (late) fail("acc_sync: a cross-part barrier waited for more than ten seconds: another part of the run has stopped");
    }
#endif
    if (e->mp.utilitarian) {
        int flag = 0;
        d2h(&flag, e->fatal.p, sizeof(int), e->stream);
        // the reference ends the run here: "optimization failed ... (for n consumption goods)", Consumer.cpp:497-502
        if (flag) fail("utility optimisation of a consumer failed (more than " + std::to_string(ACCSQP_MAX_ITER) + " SQP iterations)");
    }
    return 0;
    ACC_API_END(-1)
}

int acc_drain_events(acc_engine* e, acc_event* buf, size_t cap, size_t* n) {
    ACC_API_BEGIN
    if (e == nullptr || n == nullptr) fail("acc_drain_events: null argument");
    ACC_ON_DEVICE(e);
    int count = 0;
    d2h(&count, e->ev_count.p, sizeof(int), e->stream);
    const int raised = count;
    if (count > e->ev_cap) count = e->ev_cap;
    std::vector<Event> h((size_t)count);
    d2h(h.data(), e->ev_buf.p, (size_t)count * sizeof(Event), e->stream);
    // the device appends in arrival order; present a deterministic order
    std::sort(h.begin(), h.end(), [](const Event& x, const Event& y) {
        if (x.step != y.step) return x.step < y.step;
        if (x.type != y.type) return x.type < y.type;
        if (x.a != y.a) return x.a < y.a;
        return x.b < y.b;
    });
    *n = h.size();
    if (buf == nullptr && cap == 0 && raised <= e->ev_cap) return 0;  // a count query: nothing is taken
    const size_t m = buf != nullptr ? std::min(cap, h.size()) : 0;
    for (size_t k = 0; k < m; ++k) { buf[k].step = h[k].step; buf[k].type = h[k].type; buf[k].a = h[k].a; buf[k].b = h[k].b; buf[k].value = h[k].value; }
    // what did not fit the caller's buffer stays pending (sorted, at the front of the device buffer)
    const int rest = (int)(h.size() - m);
    if (rest > 0 && m > 0) h2d(e->ev_buf.p, h.data() + m, (size_t)rest * sizeof(Event), e->stream);
    h2d(e->ev_count.p, &rest, sizeof(int), e->stream);
    dev_sync(e->stream);
    if (raised > e->ev_cap)
        fail("acc_drain_events: " + std::to_string(raised - e->ev_cap) + " events were dropped: more than " + std::to_string(e->ev_cap) +
             " were raised between two drains (drain every step, or raise the capacity with acc_set_event_capacity)");
    return 0;
    ACC_API_END(-1)
}

int acc_set_event_capacity(acc_engine* e, size_t capacity) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_set_event_capacity: null engine");
    ACC_ON_DEVICE(e);
    if (capacity < 1 || capacity > ((size_t)1 << 28)) fail("acc_set_event_capacity: capacity out of range");
    int count = 0;
    d2h(&count, e->ev_count.p, sizeof(int), e->stream);
    if (count != 0) fail("acc_set_event_capacity: drain the pending events first");
    e->ev_cap = (int)capacity;
    e->ev_buf.alloc(capacity);
    return 0;
    ACC_API_END(-1)
}

int acc_read_state(acc_engine* e, const char* name, void* dst, size_t cap_bytes, size_t* n, int* is_f64) {
    ACC_API_BEGIN
    if (e == nullptr || name == nullptr || n == nullptr) fail("acc_read_state: null argument");
    ACC_ON_DEVICE(e);
    StateArray a;
    if (!read_state(*e, name, a)) fail(std::string("acc_read_state: unknown array '") + name + "'");
    const size_t cnt = a.is_f64 ? a.f.size() : a.i.size();
    *n = cnt;
    if (is_f64) *is_f64 = a.is_f64 ? 1 : 0;
    if (dst != nullptr) {
        if (cap_bytes < cnt * 8) fail("acc_read_state: destination too small");
        std::memcpy(dst, a.is_f64 ? (const void*)a.f.data() : (const void*)a.i.data(), cnt * 8);
    }
    return 0;
    ACC_API_END(-1)
}

int acc_read(acc_engine* e, const char* group, const char* variable, double* dst, size_t cap, size_t* n) {
    ACC_API_BEGIN
    if (e == nullptr || group == nullptr || variable == nullptr || n == nullptr) fail("acc_read: null argument");
    ACC_ON_DEVICE(e);
    // observable name -> (state array, kind): kind 'q' quantity (x 1e-3), 'p' price (x 1e-6), 'v' value / plain
    struct Map { const char* group; const char* var; const char* state; char kind; };
    static const Map maps[] = {
        {"firms", "production_quantity", "firm_X_q", 'q'}, {"firms", "production_value", "firm_X_v", 'v'},
        {"firms", "expected_production_quantity", "firm_Xexp_q", 'q'}, {"firms", "expected_production_value", "firm_Xexp_v", 'v'},
        {"firms", "communicated_possible_production_quantity", "firm_Xhat_q", 'q'}, {"firms", "communicated_possible_production_value", "firm_Xhat_v", 'v'},
        {"firms", "offer_price", "firm_nbar", 'p'}, {"firms", "forcing", "agent_forcing", 'v'},
        {"firms", "total_production_costs", "firm_costs_C", 'v'}, {"firms", "total_revenue", "firm_revenue_R", 'v'},
        {"firms", "incoming_demand_quantity", "firm_incoming_demand_q", 'q'}, {"firms", "incoming_demand_value", "firm_incoming_demand_v", 'v'},
        {"sectors", "total_production_quantity", "sector_production_q", 'q'}, {"sectors", "total_production_value", "sector_production_v", 'v'},
        {"regions", "import_quantity", "region_import_q", 'q'}, {"regions", "import_value", "region_import_v", 'v'},
        {"regions", "export_quantity", "region_export_q", 'q'}, {"regions", "export_value", "region_export_v", 'v'},
        {"regions", "consumption_quantity", "region_consumption_q", 'q'}, {"regions", "consumption_value", "region_consumption_v", 'v'},
        {"regions", "gdp_quantity", "region_gdp_q", 'q'}, {"regions", "gdp_value", "region_gdp_v", 'v'},
        {"storages", "content_quantity", "sto_S_q", 'q'}, {"storages", "content_value", "sto_S_v", 'v'},
        {"storages", "input_flow_quantity", "sto_I_q", 'q'}, {"storages", "input_flow_value", "sto_I_v", 'v'},
        {"storages", "used_flow_quantity", "sto_U_q", 'q'}, {"storages", "used_flow_value", "sto_U_v", 'v'},
        {"storages", "desired_used_flow_quantity", "sto_Ut_q", 'q'}, {"storages", "desired_used_flow_value", "sto_Ut_v", 'v'},
        {"storages", "demand_quantity", "sto_demand_q", 'q'}, {"storages", "demand_value", "sto_demand_v", 'v'},
        {"storages", "expected_costs", "sto_expected_costs", 'v'}, {"storages", "optimized_value", "sto_optimized_value", 'v'},
        {"storages", "total_transport_penalty", "sto_transport_penalty", 'v'},
        {"flows", "demand_request_quantity", "edge_D_q", 'q'}, {"flows", "sent_flow_quantity", "edge_ship_q", 'q'}, {"flows", "sent_flow_value", "edge_ship_v", 'v'},
    };
    for (const Map& mpp : maps) {
        if (std::strcmp(mpp.group, group) != 0 || std::strcmp(mpp.var, variable) != 0) continue;
        StateArray a;
        if (!read_state(*e, mpp.state, a)) fail("acc_read: internal mapping error");
        const size_t cnt = a.is_f64 ? a.f.size() : a.i.size();
        *n = cnt;
        if (dst == nullptr) return 0;
        if (cap < cnt) fail("acc_read: destination too small");
        for (size_t k = 0; k < cnt; ++k) dst[k] = a.is_f64 ? a.f[k] : (mpp.kind == 'q' ? qf(a.i[k]) : mpp.kind == 'p' ? pf(a.i[k]) : (double)a.i[k]);
        return 0;
    }
    fail(std::string("acc_read: unknown observable ") + group + "/" + variable);
    ACC_API_END(-1)
}

int acc_solver_group(const acc_engine* e) { return e ? e->pm_group : -1; }
int acc_set_solver_group(acc_engine* e, int group) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_set_solver_group: null engine");
    ACC_ON_DEVICE(e);
    if (group < 32 || group > 512 || group % 32 != 0) fail("acc_set_solver_group: the group must be a multiple of 32 in 32..512");
    if (e->step != 0) fail("acc_set_solver_group: only before the first step");
    e->pm_group = group;
    return 0;
    ACC_API_END(-1)
}
int64_t acc_edge_count(const acc_engine* e) { return e ? e->nE : -1; }
int64_t acc_step_count(const acc_engine* e) { return e ? e->step : -1; }
int64_t acc_launch_count(const acc_engine* e) { return e ? e->launches : -1; }
void* acc_stream(acc_engine* e) { return e ? (void*)e->stream : nullptr; }

// ---- state checkpoint ----

int acc_state_bytes(acc_engine* e, size_t* bytes) {
    ACC_API_BEGIN
    if (e == nullptr || bytes == nullptr) fail("acc_state_bytes: null argument");
    ACC_ON_DEVICE(e);
    SizeVisitor v;
    e->visit_state(v);
    *bytes = sizeof(StateHeader) + v.bytes;
    return 0;
    ACC_API_END(-1)
}

int acc_save_state(acc_engine* e, void* dst, size_t cap, size_t* written) {
    ACC_API_BEGIN
    if (e == nullptr || dst == nullptr) fail("acc_save_state: null argument");
    ACC_ON_DEVICE(e);
    SizeVisitor sz;
    e->visit_state(sz);
    const size_t total = sizeof(StateHeader) + sz.bytes;
    if (cap < total) fail("acc_save_state: buffer too small (see acc_state_bytes)");
    acc::dev_sync(e->stream);
    StateHeader h;
    std::memcpy(h.magic, "ACCSTAT2", 8);
    h.n_agents = e->nA; h.n_storages = e->nSt; h.n_edges = e->nE; h.planes = e->T; h.step = e->step; h.reg = e->reg;
    h.forcing_used = e->forcing_used ? 1 : 0; h.payload_bytes = (int64_t)sz.bytes; h.ship_step = e->ship_step; h.event_capacity = e->ev_cap;
    std::memcpy(dst, &h, sizeof h);
    SaveVisitor v{static_cast<unsigned char*>(dst) + sizeof h, e->stream};
    e->visit_state(v);
    acc::dev_sync(e->stream);
    if (written) *written = total;
    return 0;
    ACC_API_END(-1)
}

int acc_load_state(acc_engine* e, const void* src, size_t bytes) {
    ACC_API_BEGIN
    if (e == nullptr || src == nullptr) fail("acc_load_state: null argument");
    ACC_ON_DEVICE(e);
    if (bytes < sizeof(StateHeader)) fail("acc_load_state: not a state blob");
    StateHeader h;
    std::memcpy(&h, src, sizeof h);
    if (std::memcmp(h.magic, "ACCSTAT2", 8) != 0) fail("acc_load_state: not a state blob");
    if (h.n_agents != e->nA || h.n_storages != e->nSt || h.n_edges != e->nE || h.planes != e->T)
        fail("acc_load_state: the blob belongs to a different network snapshot");
    if (h.event_capacity != e->ev_cap) fail("acc_load_state: the blob was saved with another event capacity (acc_set_event_capacity)");
    if (h.forcing_used != 0) e->ensure_forcing_state();
    else if (e->forcing_used) fail("acc_load_state: this engine has used passage forcing, the blob's run had not; load into a fresh engine");
    SizeVisitor sz;
    e->visit_state(sz);
    if ((int64_t)sz.bytes != h.payload_bytes || bytes < sizeof h + sz.bytes) fail("acc_load_state: blob size does not match this engine");
    acc::dev_sync(e->stream);
    LoadVisitor v{static_cast<const unsigned char*>(src) + sizeof h, e->stream};
    e->visit_state(v);
    acc::dev_sync(e->stream);
    e->step = h.step; e->reg = (int)h.reg; e->ship_step = h.ship_step;
    e->ahead_step = -1;  // a consumers' solve made ahead belonged to the state that has just been replaced
    {   // the field-major twin of the seller records is derived state: rebuild it from the loaded records
        RepublishSellers k{e->pub.p, e->seller_table()};
        launch_thread(e->stream, e->nA, k);
        acc::dev_sync(e->stream);
    }
    return 0;
    ACC_API_END(-1)
}

// ---- acc_firm_incoming_demand_static: the firms' incoming demand without any device allocation ----
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops).  The per-firm sums that acc_read("firms", "incoming_demand_*") puts into
// freshly allocated device buffers go into a file-scope device array instead.  Every GPU that loads the library holds
// the array (ACC_STATIC_SCRATCH_AGENTS x 16 bytes = 1 MiB) until the process exits.  All calls share it and it keeps the
// last call's contents: each call clears the rows it uses first, and calls that overlap on one GPU (two engines on one
// device driven from two threads) race.
}  // extern "C"
#ifdef ACC_HOST_EMU
static i64 g_static_scratch_q[ACC_STATIC_SCRATCH_AGENTS];
static double g_static_scratch_v[ACC_STATIC_SCRATCH_AGENTS];
static void static_scratch(int, i64** q, double** v) { *q = g_static_scratch_q; *v = g_static_scratch_v; }
#else
static __device__ __align__(256) i64 g_static_scratch_q[ACC_STATIC_SCRATCH_AGENTS];
static __device__ __align__(256) double g_static_scratch_v[ACC_STATIC_SCRATCH_AGENTS];
// In host code the arrays' names are host-side stand-ins, not device addresses: the device addresses are asked for once
// per device, with that device current.
static void static_scratch(int device, i64** q, double** v) {
    struct Addr { i64* q; double* v; };
    static std::map<int, Addr> per_device;
    auto it = per_device.find(device);
    if (it == per_device.end()) {
        Addr a{nullptr, nullptr};
        ACC_CUDA(cudaGetSymbolAddress((void**)&a.q, g_static_scratch_q));
        ACC_CUDA(cudaGetSymbolAddress((void**)&a.v, g_static_scratch_v));
        it = per_device.emplace(device, a).first;
    }
    *q = it->second.q; *v = it->second.v;
}
#endif
extern "C" {

int acc_firm_incoming_demand_static(acc_engine* e, double* out_quantity, double* out_value, size_t cap, size_t* n) {
    ACC_API_BEGIN
    if (e == nullptr || n == nullptr) fail("acc_firm_incoming_demand_static: null argument");
    ACC_ON_DEVICE(e);
    const size_t cnt = (size_t)e->nA;
    *n = cnt;
    if (cnt > (size_t)ACC_STATIC_SCRATCH_AGENTS)
        fail("acc_firm_incoming_demand_static: " + std::to_string(cnt) + " agents exceed the static scratch of " + std::to_string(ACC_STATIC_SCRATCH_AGENTS) +
             " agents (ACC_STATIC_SCRATCH_AGENTS); use acc_read(\"firms\", \"incoming_demand_quantity\")");
    if (out_quantity == nullptr && out_value == nullptr) return 0;  // size query
    if (cap < cnt) fail("acc_firm_incoming_demand_static: destination too small");
    i64* sq = nullptr;
    double* sv = nullptr;
    static_scratch(e->device, &sq, &sv);
    stream_t s = e->stream;
    dev_zero(sq, cnt * sizeof(i64), s);  // rows of agents that are not (owned) firms must read zero, whatever the last call left
    dev_zero(sv, cnt * sizeof(double), s);
    FirmIncomingDemand k{e->edge_view(), e->firm_ids.p, sq, sv};
    e->timed("firm_incoming_demand", [&] { launch_block(s, e->nF, k); });
    // straight into the caller's buffers: the int64 quanta travel in the doubles' bytes and are converted in place
    if (out_value != nullptr) d2h(out_value, sv, cnt * sizeof(double), s);
    if (out_quantity != nullptr) {
        d2h(out_quantity, sq, cnt * sizeof(i64), s);
        for (size_t a = 0; a < cnt; ++a) {
            i64 q;
            std::memcpy(&q, &out_quantity[a], sizeof q);
            out_quantity[a] = qf(q);
        }
    } else {
        dev_sync(s);
    }
    return 0;
    ACC_API_END(-1)
}

#ifdef ACC_DEBUG_API
// Debugging aids, compiled into variant builds only (scripts/consumer_replay.py): what one consumer's utility optimisation reads
// from the model state, per input storage in input_storages order -- content (q, v), arrived input flow (q, v), previous
// consumption (q, v) -- can be taken from one engine and replayed on another (the host emulation) for that consumer alone.
int acc_debug_consumer_inputs(acc_engine* e, int agent, double* buf, size_t cap, size_t* n) {
    ACC_API_BEGIN
    ACC_ON_DEVICE(e);
    stream_t s = e->stream;
    dev_sync(s);
    if (e->side) dev_sync(e->side);
    std::vector<int> off = e->a_st_off.download(s), list = e->a_st_list.download(s);
    const int s0 = off[agent], cnt = off[agent + 1] - s0;
    *n = (size_t)cnt * 6;
    if (buf == nullptr) return 0;
    if (cap < *n) fail("acc_debug_consumer_inputs: buffer too small");
    const StorageView v = e->storage_view();
    std::vector<i64> Sq = e->s_S_q.download(s), pq = e->c_prev_q.download(s);
    std::vector<double> Sv = e->s_S_v.download(s), pv = e->c_prev_v.download(s);
    std::vector<i64> Iq = (v.Ip_q == e->s_I0_q.p ? e->s_I0_q : e->s_I1_q).download(s);
    std::vector<double> Iv = (v.Ip_v == e->s_I0_v.p ? e->s_I0_v : e->s_I1_v).download(s);
    for (int i = 0; i < cnt; ++i) {
        const int st = list[s0 + i];
        double* o = buf + 6 * (size_t)i;
        o[0] = (double)Sq[st]; o[1] = Sv[st]; o[2] = (double)Iq[st]; o[3] = Iv[st]; o[4] = (double)pq[st]; o[5] = pv[st];
    }
    return 0;
    ACC_API_END(-1)
}
int acc_debug_consumer_replay(acc_engine* e, int agent, const double* buf, size_t n, int* ret, double* x_out) {
    ACC_API_BEGIN
    ACC_ON_DEVICE(e);
    stream_t s = e->stream;
    dev_sync(s);
    if (e->side) dev_sync(e->side);
    std::vector<int> off = e->a_st_off.download(s), list = e->a_st_list.download(s), ids = e->cons_ids.download(s);
    const int s0 = off[agent], cnt = off[agent + 1] - s0;
    if (n != (size_t)cnt * 6) fail("acc_debug_consumer_replay: wrong number of values");
    int item = -1;
    for (size_t k = 0; k < ids.size(); ++k) if (ids[k] == agent) item = (int)k;
    if (item < 0) fail("acc_debug_consumer_replay: not a consumer of this engine");
    const StorageView v = e->storage_view();
    for (int i = 0; i < cnt; ++i) {
        const int st = list[s0 + i];
        const double* o = buf + 6 * (size_t)i;
        const i64 Sq = (i64)o[0], Iq = (i64)o[2], pq = (i64)o[4];
        h2d(e->s_S_q.p + st, &Sq, sizeof Sq, s); h2d(e->s_S_v.p + st, &o[1], sizeof(double), s);
        h2d(const_cast<i64*>(v.Ip_q) + st, &Iq, sizeof Iq, s); h2d(const_cast<double*>(v.Ip_v) + st, &o[3], sizeof(double), s);
        h2d(e->c_prev_q.p + st, &pq, sizeof pq, s); h2d(e->c_prev_v.p + st, &o[5], sizeof(double), s);
    }
    dev_sync(s);
    ConsumerView cv = e->consumer_view();
    cv.cons_ids = e->cons_ids.p + item;
    ConsumerUtility k{cv, e->storage_view(), e->mp, e->consumer_stage(), e->c_n_max};
    launch_warps(s, 1, consumer_workspace_bytes(e->c_n_max), k);
    dev_sync(s);
    std::vector<int> r = e->c_nx_ret.download(s);
    std::vector<i64> q = e->c_nx_q.download(s);
    *ret = r[agent];
    if (x_out != nullptr) for (int i = 0; i < cnt; ++i) x_out[i] = (double)q[list[s0 + i]];
    e->ahead_step = -1;
    return 0;
    ACC_API_END(-1)
}
#endif

int acc_profile(acc_engine* e, int enable) {
    if (e == nullptr) return -1;
    e->profile = enable != 0;
    return 0;
}

int acc_kernel_times(acc_engine* e, const char** names, double* ms, int64_t* launches, int cap, int reset) {
    ACC_API_BEGIN
    if (e == nullptr) fail("acc_kernel_times: null engine");
    ACC_ON_DEVICE(e);
    e->collect_clocks();
    const int cnt = (int)e->clocks.size();
    for (int k = 0; k < cnt && k < cap; ++k) {
        if (names) names[k] = e->clocks[k].name;
        if (ms) ms[k] = e->clocks[k].ms;
        if (launches) launches[k] = e->clocks[k].launches;
    }
    if (reset) for (auto& c : e->clocks) { c.ms = 0.0; c.launches = 0; }
    return cnt;
    ACC_API_END(-1)
}

}  // extern "C"
