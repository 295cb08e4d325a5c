// coop.cuh -- cooperative execution contexts.
//
// Every per-agent routine of the engine (sales manager, purchasing manager, ...) is written once as
// a template over a context `Ctx` that stands for the group of threads working on ONE agent:
//   BlockCtx : a whole CTA (one seller with thousands of outgoing connections)
//   WarpCtx  : one warp    (one purchasing manager, <= a few hundred connections)
// A routine only uses  cx.foreach(n, f)  (element-parallel loop followed by a group barrier),
// cx.reduce / cx.sum / ... (group reductions with a FIXED association order) and uniform scalar code
// that every thread of the group executes redundantly.
//
// The association order of a reduction over n elements by G threads is part of the contract:
//   thread t folds elements i = t, t+G, t+2G, ... in increasing i, starting from the identity;
//   inside each warp the 32 partials are combined by an xor butterfly (offsets 16,8,4,2,1);
//   warp results are then folded sequentially, warp 0 first.
// HostCtx reproduces exactly that order with a loop, so the same routine compiled for the host gives
// bit-identical f64 results.  HostCtx exists for two users only: oracle/nlopt_shim (the reference's CPU run
// must execute the same SQP arithmetic as the device) and the CPU-side unit tests of the kernel logic
// (tests/emu).  The product library never executes it: there is no CPU path in libacclimate_b200.so.
#pragma once

#include "numeric.cuh"

namespace acc {

struct SumF64 {
    double s;
    static ACC_HD SumF64 identity() { SumF64 r; r.s = 0.0; return r; }
    static ACC_HD SumF64 combine(SumF64 a, SumF64 b) { SumF64 r; r.s = a.s + b.s; return r; }
};
struct SumI64 {
    i64 s;
    static ACC_HD SumI64 identity() { SumI64 r; r.s = 0; return r; }
    static ACC_HD SumI64 combine(SumI64 a, SumI64 b) { SumI64 r; r.s = a.s + b.s; return r; }
};
struct MinF64 {
    double s;
    static ACC_HD MinF64 identity() { MinF64 r; r.s = 1e308; return r; }
    static ACC_HD MinF64 combine(MinF64 a, MinF64 b) { return b.s < a.s ? b : a; }
};
struct MaxF64 {
    double s;
    static ACC_HD MaxF64 identity() { MaxF64 r; r.s = -1e308; return r; }
    static ACC_HD MaxF64 combine(MaxF64 a, MaxF64 b) { return b.s > a.s ? b : a; }
};

// Convenience layer shared by all contexts (CRTP): sum / isum / vmin / vmax on top of reduce<T>().
template <class Derived>
struct CtxOps {
    template <class F>
    ACC_HD double sum(int n, F f) {
        return static_cast<Derived*>(this)->template reduce<SumF64>(n, [&](int i) { SumF64 r; r.s = f(i); return r; }).s;
    }
    template <class F>
    ACC_HD i64 isum(int n, F f) {
        return static_cast<Derived*>(this)->template reduce<SumI64>(n, [&](int i) { SumI64 r; r.s = f(i); return r; }).s;
    }
    template <class F>
    ACC_HD double vmin(int n, F f) {
        return static_cast<Derived*>(this)->template reduce<MinF64>(n, [&](int i) { MinF64 r; r.s = f(i); return r; }).s;
    }
    template <class F>
    ACC_HD double vmax(int n, F f) {
        return static_cast<Derived*>(this)->template reduce<MaxF64>(n, [&](int i) { MaxF64 r; r.s = f(i); return r; }).s;
    }
};

// ------------------------------------------------------------------------------------------------
// Host emulation of a group of G threads (G = 32 for a warp, the CTA size for a block).
// ------------------------------------------------------------------------------------------------
struct HostCtx : CtxOps<HostCtx> {
    int G;
    explicit HostCtx(int g = 32) : G(g) {}
    inline bool leader() const { return true; }
    inline void sync() const {}
    inline int lane() const { return 0; }
    template <class F>
    inline void foreach(int n, F f) const { for (int i = 0; i < n; ++i) f(i); }
    // code that only the group's leader executes, fenced by group barriers (results travel through the group's memory)
    template <class F>
    inline void single(F f) const { f(); }
    template <class T, class F>
    inline T reduce(int n, F f) const {
        const int W = G / 32;
        T total = T::identity();
        for (int w = 0; w < W; ++w) {
            T p[32];
            for (int l = 0; l < 32; ++l) {
                T acc_ = T::identity();
                for (int i = w * 32 + l; i < n; i += G) acc_ = T::combine(acc_, f(i));
                p[l] = acc_;
            }
            for (int off = 16; off >= 1; off >>= 1) {
                T q[32];
                for (int l = 0; l < 32; ++l) q[l] = T::combine(p[l], p[l ^ off]);
                for (int l = 0; l < 32; ++l) p[l] = q[l];
            }
            total = (w == 0) ? p[0] : T::combine(total, p[0]);
        }
        return total;
    }
    // K strictly sequential left-to-right folds (the reference's own loop order)
    // reduce over f(i, pre(i)); on the device pre() of a thread's next element is issued before f() of the current one
    template <class T, class Pre, class F>
    inline T reduce_staged(int n, Pre pre, F f) const {
        return reduce<T>(n, [&](int i) { return f(i, pre(i)); });
    }
    // K strictly sequential left-to-right sums (the reference's own loop order); elem(i, k) = k-th summand of element i
    template <int K, class Elem>
    inline void seqsum(int n, Elem elem, double (&out)[K]) const {
        for (int k = 0; k < K; ++k) out[k] = 0.0;
        for (int i = 0; i < n; ++i)
            for (int k = 0; k < K; ++k) out[k] += elem(i, k);
    }
    // order-preserving compaction: put(m, i) for ascending i with pred(i), m = 0, 1, ...; returns the count
    template <class Pred, class Put>
    inline int compact(int n, Pred pred, Put put) const {
        int m = 0;
        for (int i = 0; i < n; ++i) if (pred(i)) put(m++, i);
        return m;
    }
};

#if defined(__CUDACC__)
// ------------------------------------------------------------------------------------------------
// Device contexts
// ------------------------------------------------------------------------------------------------
template <class T>
__device__ __forceinline__ T shfl_xor_any(const T& v, int off) {
    static_assert(sizeof(T) % 4 == 0, "reduction payload must be a multiple of 4 bytes");
    T r;
    const unsigned* src = reinterpret_cast<const unsigned*>(&v);
    unsigned* dst = reinterpret_cast<unsigned*>(&r);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 4); ++k) dst[k] = __shfl_xor_sync(0xffffffffu, src[k], off);
    return r;
}
template <class T>
__device__ __forceinline__ T shfl_any(const T& v, int src_lane) {
    T r;
    const unsigned* src = reinterpret_cast<const unsigned*>(&v);
    unsigned* dst = reinterpret_cast<unsigned*>(&r);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 4); ++k) dst[k] = __shfl_sync(0xffffffffu, src[k], src_lane);
    return r;
}

// One strictly sequential left-to-right sum per lane: lane k < K folds the k-th summand elem(i, k) over i = 0..n-1.
// The K chains run side by side in the lanes of one warp (one instruction stream); each chain is bound by the latency
// of its dependent additions, so the summands of eight elements are fetched first (independent loads).
template <class Elem>
__device__ __forceinline__ double seqsum_lane(int n, Elem elem, int k) {
    double a = 0.0;
    const int nb = n >> 3;
    if (nb > 0) {
        double t[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = elem(j, k);
        for (int b = 0; b < nb; ++b) {
            // the next batch is fetched (the last one once more: no branch) while the additions of this one run
            const int nxt = (b + 1 < nb ? b + 1 : b) << 3;
            double u[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) u[j] = elem(nxt + j, k);
#pragma unroll
            for (int j = 0; j < 8; ++j) a += t[j];
#pragma unroll
            for (int j = 0; j < 8; ++j) t[j] = u[j];
        }
    }
    for (int i = nb << 3; i < n; ++i) a += elem(i, k);
    return a;
}

// One warp per item.
struct WarpCtx : CtxOps<WarpCtx> {
    int lane_;
    __device__ WarpCtx() : lane_(threadIdx.x & 31) {}
    __device__ __forceinline__ bool leader() const { return lane_ == 0; }
    __device__ __forceinline__ void sync() const { __syncwarp(); }
    __device__ __forceinline__ int lane() const { return lane_; }
    template <class F>
    __device__ __forceinline__ void foreach(int n, F f) const {
        for (int i = lane_; i < n; i += 32) f(i);
        __syncwarp();
    }
    template <class F>
    __device__ __forceinline__ void single(F f) const {
        __syncwarp();
        if (lane_ == 0) f();
        __syncwarp();
    }
    template <class T, class F>
    __device__ __forceinline__ T reduce(int n, F f) const {
        T acc_ = T::identity();
        for (int i = lane_; i < n; i += 32) acc_ = T::combine(acc_, f(i));
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) acc_ = T::combine(acc_, shfl_xor_any(acc_, off));
        return acc_;
    }
    template <class T, class Pre, class F>
    __device__ __forceinline__ T reduce_staged(int n, Pre pre, F f) const {
        return reduce<T>(n, [&](int i) { return f(i, pre(i)); });
    }
    template <int K, class Elem>
    __device__ __forceinline__ void seqsum(int n, Elem elem, double (&out)[K]) const {
        double a = 0.0;
        __syncwarp();
        if (lane_ < K) a = seqsum_lane(n, elem, lane_);
#pragma unroll
        for (int k = 0; k < K; ++k) out[k] = __shfl_sync(0xffffffffu, a, k);
    }
    template <class Pred, class Put>
    __device__ __forceinline__ int compact(int n, Pred pred, Put put) const {
        int m = 0;
        for (int base = 0; base < n; base += 32) {
            const int i = base + lane_;
            const bool keep = i < n && pred(i);
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            if (keep) put(m + __popc(mask & ((1u << lane_) - 1u)), i);
            m += __popc(mask);
        }
        __syncwarp();
        return m;
    }
};

// A CTA working on one item.  `scratch` is shared memory with room for one payload of the largest reduction type
// per warp (+1), see ACC_BLOCK_SCRATCH_BYTES.  G > 0 fixes the CTA size at compile time (strides and the fold over
// the warps unroll); G == 0 reads blockDim.x.
#define ACC_BLOCK_SCRATCH_BYTES (33 * 64)
#define ACC_BLOCK_SCRATCH_BYTES_FOR(G) ((((G) + 31) / 32 + 1) * 64)
template <int G>
struct BlockCtxT : CtxOps<BlockCtxT<G>> {
    unsigned char* scratch;
    __device__ explicit BlockCtxT(unsigned char* s) : scratch(s) {}
    __device__ __forceinline__ int size() const { return G > 0 ? G : (int)blockDim.x; }
    __device__ __forceinline__ bool leader() const { return threadIdx.x == 0; }
    __device__ __forceinline__ void sync() const { __syncthreads(); }
    __device__ __forceinline__ int lane() const { return threadIdx.x; }
    template <class F>
    __device__ __forceinline__ void foreach(int n, F f) const {
        for (int i = threadIdx.x; i < n; i += size()) f(i);
        __syncthreads();
    }
    template <class F>
    __device__ __forceinline__ void single(F f) const {
        __syncthreads();
        if (threadIdx.x == 0) f();
        __syncthreads();
    }
    template <class T, class F>
    __device__ __forceinline__ T reduce(int n, F f) const {
        T acc_ = T::identity();
        for (int i = threadIdx.x; i < n; i += size()) acc_ = T::combine(acc_, f(i));
        return fold_group(acc_);
    }
    // Same fold order as reduce(); the gather pre() of a thread's two next elements is issued (clamped index, no
    // branch) before either element is worked on, so their dependent loads overlap.
    template <class T, class Pre, class F>
    __device__ __forceinline__ T reduce_staged(int n, Pre pre, F f) const {
        T acc_ = T::identity();
        const int g = size();
        for (int i = threadIdx.x; i < n; i += 2 * g) {
            const int i1 = i + g;
            const auto p0 = pre(i);
            const auto p1 = pre(i1 < n ? i1 : n - 1);
            acc_ = T::combine(acc_, f(i, p0));
            if (i1 < n) acc_ = T::combine(acc_, f(i1, p1));
        }
        return fold_group(acc_);
    }
    template <class T>
    __device__ __forceinline__ T fold_group(T acc_) const {
        static_assert(sizeof(T) <= 64, "reduction payload too large for the block scratch");
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) acc_ = T::combine(acc_, shfl_xor_any(acc_, off));
        T* slots = reinterpret_cast<T*>(scratch);
        const int w = threadIdx.x >> 5, W = size() >> 5;
        __syncthreads();  // scratch may still be read by a previous reduction
        if ((threadIdx.x & 31) == 0) slots[w] = acc_;
        __syncthreads();
        T total = slots[0];
        if (G > 0) {
#pragma unroll
            for (int k = 1; k < (G > 0 ? G / 32 : 1); ++k) total = T::combine(total, slots[k]);
        } else {
            for (int k = 1; k < W; ++k) total = T::combine(total, slots[k]);
        }
        return total;
    }
    template <int K, class Elem>
    __device__ __forceinline__ void seqsum(int n, Elem elem, double (&out)[K]) const {
        double* slots = reinterpret_cast<double*>(scratch);
        __syncthreads();
        if (threadIdx.x < K) slots[threadIdx.x] = seqsum_lane(n, elem, (int)threadIdx.x);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; ++k) out[k] = slots[k];
    }
    template <class Pred, class Put>
    __device__ __forceinline__ int compact(int n, Pred pred, Put put) const {
        int* counts = reinterpret_cast<int*>(scratch);  // per-warp counts of the current chunk
        const int w = threadIdx.x >> 5, W = size() >> 5, lane_ = threadIdx.x & 31;
        int m = 0;
        for (int base = 0; base < n; base += size()) {
            const int i = base + threadIdx.x;
            const bool keep = i < n && pred(i);
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            __syncthreads();
            if (lane_ == 0) counts[w] = __popc(mask);
            __syncthreads();
            int before = 0, total = 0;
            if (G > 0) {
#pragma unroll
                for (int k = 0; k < (G > 0 ? G / 32 : 1); ++k) { if (k < w) before += counts[k]; total += counts[k]; }
            } else {
                for (int k = 0; k < W; ++k) { if (k < w) before += counts[k]; total += counts[k]; }
            }
            if (keep) put(m + before + __popc(mask & ((1u << lane_) - 1u)), i);
            m += total;
        }
        __syncthreads();
        return m;
    }
};
typedef BlockCtxT<0> BlockCtx;
#endif  // __CUDACC__

}  // namespace acc
