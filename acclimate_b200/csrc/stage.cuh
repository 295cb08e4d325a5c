// stage.cuh -- bulk staging of one seller's demand book into shared memory (sm_100a: cp.async.bulk + mbarrier).
//
// SalesManager::distribute_production and calc_supply_distribution_scenario (reference src/model/SalesManager.cpp:140-423)
// walk the seller's demand requests several times (totals, the price scan, the cheapest-price-range totals, the
// shipments).  A seller's connections are contiguous (CSR by seller), so ONE thread hands the whole book to the copy
// engine -- three 1-D bulk copies (quantity, price, domestic flag) that complete on an mbarrier -- and every pass
// then runs out of shared memory: HBM sees each demand request once, and the number of loads in flight no longer
// depends on how many warps are resident.
#pragma once

#include "numeric.cuh"

namespace acc {

// what seller_phase1 reads per connection; null pointers = read the global arrays
struct SellerBook {
    const i64* q; const i64* p; const unsigned char* dom;
    unsigned short* perm16; int perm_slots;  // shared-memory scratch for a sort of the seller's connections (model_core.cuh ordered_range_sums), or null
};

ACC_HD size_t seller_book_bytes(int cap) {
    // two i64 arrays with one element of slack (16-byte phase of the source), one byte array with 16 bytes of slack
    const size_t words = ((size_t)(cap + 2) * sizeof(i64) + 15) / 16 * 16;
    const size_t bytes = ((size_t)cap + 16 + 15) / 16 * 16;
    return 2 * words + bytes;
}

#if defined(__CUDACC__)
__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)), "l"(src), "r"(bytes),
                 "r"(smem_addr(bar))
                 : "memory");
}
// asks the copy engine to bring [p, p + bytes) into the L2 (no destination: a later bulk copy or load then finds it there)
__device__ __forceinline__ void prefetch_l2(const void* p, size_t bytes) {
    const size_t a = reinterpret_cast<size_t>(p) & ~(size_t)15;
    const size_t end = (reinterpret_cast<size_t>(p) + bytes + 15) & ~(size_t)15;
    if (end > a) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a), "r"((unsigned)(end - a)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "ACC_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra ACC_MBAR_DONE;\n"
        "bra ACC_MBAR_WAIT;\n"
        "ACC_MBAR_DONE:\n"
        "}" ::"r"(smem_addr(bar)),
        "r"(parity)
        : "memory");
}

// One array of the book: dst keeps the source's 16-byte phase, so that the 16-byte aligned interior goes through the
// copy engine; the (at most 15-byte) head and tail are moved by ordinary loads.  Returns the interior's size in bytes.
template <class T>
struct BookPart {
    const T* src; T* dst; int n; int head; unsigned body_bytes;
    __device__ __forceinline__ BookPart(const T* src_, unsigned char* slot, int n_) : src(src_), n(n_) {
        const size_t a = reinterpret_cast<size_t>(src_);
        dst = reinterpret_cast<T*>(slot + (a & 15));
        const int to_align = (int)(((16 - (a & 15)) & 15) / sizeof(T));
        head = to_align < n_ ? to_align : n_;
        body_bytes = (unsigned)(((size_t)(n_ - head) * sizeof(T)) & ~(size_t)15);
    }
    __device__ __forceinline__ void issue(unsigned long long* bar) const {
        if (body_bytes) bulk_g2s(dst + head, src + head, body_bytes, bar);
    }
    __device__ __forceinline__ void edges() const {  // first warp: head and tail elements (fewer than 32)
        const int body = (int)(body_bytes / sizeof(T));
        const int rest = n - body;  // head + tail elements
        for (int j = threadIdx.x; j < rest; j += 32) {
            const int i = j < head ? j : body + j;
            dst[i] = src[i];
        }
    }
};

// Stage the book of the connections [e0, e0 + n) of one seller.  `bar` is one mbarrier in static shared memory that this
// CTA uses once (phase 0); the caller has initialised it (mbar_init(bar, 1) by one thread, then a CTA barrier) and puts
// a CTA barrier behind this call.  Thread 0 hands the three arrays to the copy engine, the first warp moves the few
// unaligned head / tail elements, every caller waits for the copies.  A warp may arrive late (it works on something else
// first, see seller_kernel): it then only waits.
__device__ __forceinline__ SellerBook stage_seller_book(unsigned char* smem, int cap, unsigned long long* bar, const i64* D_q, const i64* D_p,
                                                        const unsigned char* dom, int e0, int n) {
    const size_t words = ((size_t)(cap + 2) * sizeof(i64) + 15) / 16 * 16;
    const BookPart<i64> q(D_q + e0, smem, n), p(D_p + e0, smem + words, n);
    const BookPart<unsigned char> d(dom + e0, smem + 2 * words, n);
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, q.body_bytes + p.body_bytes + d.body_bytes);
        q.issue(bar); p.issue(bar); d.issue(bar);
    }
    if (threadIdx.x < 32) { q.edges(); p.edges(); d.edges(); }
    mbar_wait(bar, 0);
    SellerBook b;
    b.q = q.dst; b.p = p.dst; b.dom = d.dst; b.perm16 = nullptr; b.perm_slots = 0;
    return b;
}
#endif

}  // namespace acc
