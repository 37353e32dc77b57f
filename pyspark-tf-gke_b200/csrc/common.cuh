// Shared device/host helpers for the sm_100a ETL kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

#include "../../include/etl_b200.h"

namespace etl {

// ---------------------------------------------------------------- host side

void set_error(const char* fmt, ...);

struct Handle {
    int device = -1;
    int sm_count = 0;
    void* scratch = nullptr;       // general arena (partials, tables); contents undefined between calls
    size_t scratch_bytes = 0;
    uint64_t* tile_state = nullptr;  // look-back words only; zeroed when (re)allocated or on epoch wrap
    size_t tile_words = 0;
    uint32_t epoch = 0;            // tag of the last scan launch; 0 is never a live tag
    unsigned int* counters = nullptr;  // 64 "last block done" tickets; zero between launches
    // staging for the *_host entry points
    void* stage = nullptr;
    size_t stage_bytes = 0;
    // CSV index state kept in the arena by etl_csv_index_rows (masks of every 64-byte window of that text);
    // cleared whenever the arena is reallocated
    const void* csv_text = nullptr;
    int64_t csv_nbytes = 0;
    // projection kernel launch state of this device (set on the first projection through this handle)
    int proj_stages = 0, proj_ctas_per_sm = 0, proj_hints = 3;
    cudaStream_t streams[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t events[12] = {};
};

// make the general arena at least `bytes` (cudaMalloc + device sync when it grows)
int arena_reserve(Handle* h, size_t bytes);
// make room for `words` tile-state words and hand out the tag for one scan launch
int tile_state_acquire(Handle* h, size_t words, cudaStream_t stream, uint32_t* epoch);

#define ETL_CUDA(call)                                                              \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess) {                                                   \
            etl::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                           __FILE__, __LINE__);                                     \
            return ETL_ECUDA;                                                       \
        }                                                                           \
    } while (0)

#define ETL_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            etl::set_error(__VA_ARGS__);  \
            return ETL_EINVAL;            \
        }                                 \
    } while (0)

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline bool aligned4(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 3) == 0; }
static inline size_t dtype_size(int dt) { return (dt == ETL_F64 || dt == ETL_I64) ? 8 : 4; }

// -------------------------------------------------------------- device side

#ifdef __CUDACC__

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// doubles as unsigned integers that sort like Spark orders doubles (for integer atomicMin / atomicMax)
__device__ __forceinline__ uint64_t dbl_sortable(double x) {
    if (x != x) return 0xfff8000000000000ull;  // every NaN -> one value above +inf (Spark: NaN is largest)
    const uint64_t b = (uint64_t)__double_as_longlong(x);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double dbl_from_sortable(uint64_t e) {
    const uint64_t b = (e >> 63) ? (e & 0x7fffffffffffffffull) : ~e;
    return __longlong_as_double((long long)b);
}

// streaming 128-bit loads: read once, keep out of L1
__device__ __forceinline__ int4 ld_stream_v4(const void* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ double2 ld_stream_d2(const void* p) {
    double2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}

// tile-state word traffic for the decoupled look-back scans (one 64-bit word
// carries tag + flag + value, so relaxed gpu-scope accesses are enough)
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- mbarrier + 1-D bulk async copy (TMA engine, no tensor map) ------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes % 16 == 0; completes on `bar`
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// the same copy with an L2 eviction-priority hint (createpolicy) - for data that is read exactly once
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_load_hint(void* smem_dst, const void* gmem_src, unsigned bytes, uint64_t* bar,
                                               uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
// streaming stores: written once, not read again by this kernel
__device__ __forceinline__ void st_stream_f64(double* p, double v) { asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory"); }
__device__ __forceinline__ void st_stream_s32(int32_t* p, int32_t v) { asm volatile("st.global.cs.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ETL_SMEM_CARVEOUT=<percent>: preferred shared-memory carve-out of the partition and merge kernels (unset = the
// driver's choice, the smallest configuration that holds the resident CTAs).  A larger carve-out leaves room for
// another kernel's shared memory on the same SM (the TMA slab mover beside the merge).
inline int smem_carveout_percent() {
    static const int v = [] {
        const char* e = getenv("ETL_SMEM_CARVEOUT");
        return e && *e ? atoi(e) : -1;
    }();
    return v;
}

// Tile-state word: [63:42] epoch (22 bits) | [41:40] flag | [39:0] value.
constexpr int TS_VALUE_BITS = 40;
constexpr uint64_t TS_VALUE_MASK = (1ull << TS_VALUE_BITS) - 1;
constexpr uint64_t TS_FLAG_AGG = 1ull << TS_VALUE_BITS;
constexpr uint64_t TS_FLAG_INCL = 2ull << TS_VALUE_BITS;
constexpr uint64_t TS_FLAG_MASK = 3ull << TS_VALUE_BITS;
constexpr int TS_EPOCH_SHIFT = 42;
constexpr uint32_t TS_EPOCH_MAX = (1u << 22) - 1;

// Decoupled look-back, run by ONE full warp of tile `tile`: publishes the
// tile's aggregate, walks back over predecessors' states until it meets an
// inclusive prefix, publishes its own inclusive prefix and returns the
// exclusive prefix (same value in every lane).  Tiles must be dispatched in
// index order (blockIdx order), as in any single-pass chained scan.
__device__ __forceinline__ uint64_t lookback_exclusive(uint64_t* state, uint32_t epoch, int64_t tile,
                                                       uint64_t aggregate) {
    const unsigned lane = threadIdx.x & 31;
    const uint64_t tag = (uint64_t)epoch << TS_EPOCH_SHIFT;
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(state, tag | TS_FLAG_INCL | aggregate);
        return 0;
    }
    if (lane == 0) st_relaxed_u64(state + tile, tag | TS_FLAG_AGG | aggregate);
    uint64_t excl = 0;
    int64_t j = tile - 1;
    while (true) {
        const int64_t idx = j - lane;
        uint64_t s;
        if (idx >= 0) {
            do {
                s = ld_relaxed_u64(state + idx);
            } while ((s >> TS_EPOCH_SHIFT) != epoch || (s & TS_FLAG_MASK) == 0);
        } else {
            s = TS_FLAG_INCL;  // virtual predecessor of tile 0: inclusive prefix 0
        }
        const unsigned incl = __ballot_sync(0xffffffffu, (s & TS_FLAG_MASK) == TS_FLAG_INCL);
        uint64_t v = s & TS_VALUE_MASK;
        if (incl) {
            const unsigned first = __ffs(incl) - 1;  // nearest predecessor holding a prefix
            if (lane > first) v = 0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        excl += v;
        if (incl) break;
        j -= 32;
    }
    if (lane == 0) st_relaxed_u64(state + tile, tag | TS_FLAG_INCL | (excl + aggregate));
    return excl;
}

#endif  // __CUDACC__

}  // namespace etl
