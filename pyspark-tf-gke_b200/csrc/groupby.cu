// groupBy(key).agg(count, sum, mean, min, max) - hash aggregation.
// North-star extension: the reference's only groupBy is the count hidden inside
// StringIndexer.fit (k_means.py:34,73); Spark semantics are restated in
// oracle/etl_oracle.py::groupby_agg.
//
// One global open-addressing table (array of fixed-stride entries: key, row
// count, per column {null cells, sum} [, {min, max}]; count(c) = rows - nulls,
// so a row without nulls costs one atomic per column) is the result AND the
// spill target.  Two kernels fill it:
//   groupby_smem_kernel    low cardinality: every CTA aggregates into its own
//        shared-memory open-addressing table (64-bit CAS on the key slot,
//        32-bit native shared atomics for counts, CAS-loop shared atomics for
//        fp64 sums) and flushes it once at the end; a row whose key finds no
//        shared slot within the probe limit spills straight to the global table;
//   groupby_global_kernel  high cardinality: every row goes to the global table
//        (one 32-byte entry = one sector for a single value column).
// Then groupby_compact_kernel turns used entries into dense columnar output.
// Hash = splitmix64 finaliser; the same function picks the destination rank of
// the exchange (hash_partition), so a key's rank and its slot are independent
// bits of one hash.
#include "common.cuh"
#include "scatter.cuh"

namespace etl {

constexpr uint64_t GB_EMPTY = 0x8000000000000000ull;  // INT64_MIN: reserved, cannot be a key
constexpr int GB_THREADS = 256;
constexpr int GB_MAX_PROBES_SMEM = 32;

struct GroupByParams {
    const void* keys;
    int32_t key_dtype;      // ETL_I32 / ETL_I64
    int64_t key_stride;     // in elements (1 = contiguous)
    const uint32_t* key_valid;
    const double* val[3];
    int64_t val_stride[3];
    const uint32_t* val_valid[3];
    int32_t ncols;
    int32_t minmax;
    int64_t n;
    uint64_t* table;        // [capacity + 1] entries of `stride` words; entry `capacity` = the null-key group
    uint64_t mask;          // capacity - 1
    int32_t stride;         // words per entry: 2 + 2 * ncols (+ 2 * ncols with min/max)
    int32_t smem_slots;     // power of two
    int32_t key_range;      // > 0: keys are known to lie in [0, key_range) -> direct-indexed shared table
    unsigned long long* status;  // [0] rows that found no global slot (table full), [1] rows with the reserved key
};

__device__ __forceinline__ int64_t gb_load_key(const GroupByParams& p, int64_t i) {
    if (p.key_dtype == ETL_I64) return static_cast<const int64_t*>(p.keys)[i * p.key_stride];
    return (int64_t)static_cast<const int32_t*>(p.keys)[i * p.key_stride];
}
__device__ __forceinline__ bool gb_bit(const uint32_t* bm, int64_t i) {
    return !bm || ((__ldg(bm + (i >> 5)) >> (i & 31)) & 1u);
}

// add one row's (or one flushed partial's) contribution to a global entry
__device__ __forceinline__ void gb_global_add(uint64_t* e, const GroupByParams& p, unsigned long long rows,
                                              const unsigned long long cnt[3], const double sum[3],
                                              const uint64_t mn[3], const uint64_t mx[3]) {
    atomicAdd(reinterpret_cast<unsigned long long*>(e + 1), rows);
    for (int c = 0; c < p.ncols; ++c) {
        if (cnt[c] != rows) atomicAdd(reinterpret_cast<unsigned long long*>(e + 2 + 2 * c), rows - cnt[c]);  // nulls
        if (!cnt[c]) continue;
        atomicAdd(reinterpret_cast<double*>(e + 3 + 2 * c), sum[c]);
        if (p.minmax) {
            atomicMin(reinterpret_cast<unsigned long long*>(e + 2 + 2 * p.ncols + 2 * c), (unsigned long long)mn[c]);
            atomicMax(reinterpret_cast<unsigned long long*>(e + 3 + 2 * p.ncols + 2 * c), (unsigned long long)mx[c]);
        }
    }
}

// entry of `key` in the global table (inserting it), or nullptr when the table is full
__device__ __forceinline__ uint64_t* gb_global_slot(const GroupByParams& p, uint64_t key) {
    // home slot = top bits of the LOW hash word (range-reduced), so the keys of bucket b of B
    // (etl_hash_partition's second level, same word) live in one contiguous 1/B of the table
    uint64_t slot = __umulhi((unsigned)mix64(key), (unsigned)(p.mask + 1));
    for (uint64_t probe = 0; probe <= p.mask; ++probe) {
        uint64_t* e = p.table + slot * p.stride;
        uint64_t cur = ld_relaxed_u64(e);
        if (cur == GB_EMPTY)
            cur = atomicCAS(reinterpret_cast<unsigned long long*>(e), (unsigned long long)GB_EMPTY, (unsigned long long)key);
        if (cur == GB_EMPTY || cur == key) return e;
        slot = (slot + 1) & p.mask;
    }
    return nullptr;
}

__device__ __forceinline__ void gb_row_to_global(const GroupByParams& p, uint64_t key, bool key_null, const bool ok[3],
                                                 const double v[3]) {
    uint64_t* e = key_null ? p.table + (p.mask + 1) * p.stride : gb_global_slot(p, key);
    if (!e) {
        atomicAdd(p.status, 1ull);
        return;
    }
    unsigned long long cnt[3];
    uint64_t mn[3], mx[3];
    double sum[3];
    for (int c = 0; c < 3; ++c) {
        cnt[c] = ok[c] ? 1 : 0;
        sum[c] = v[c];
        mn[c] = mx[c] = dbl_sortable(v[c]);
    }
    gb_global_add(e, p, 1, cnt, sum, mn, mx);
}

__global__ void __launch_bounds__(GB_THREADS) groupby_global_kernel(const GroupByParams p) {
    for (int64_t i = (int64_t)blockIdx.x * GB_THREADS + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * GB_THREADS) {
        const bool key_null = !gb_bit(p.key_valid, i);
        const uint64_t key = (uint64_t)gb_load_key(p, i);
        if (!key_null && key == GB_EMPTY) {
            atomicAdd(p.status + 1, 1ull);
            continue;
        }
        bool ok[3] = {false, false, false};
        double v[3] = {0.0, 0.0, 0.0};
        for (int c = 0; c < p.ncols; ++c) {
            ok[c] = gb_bit(p.val_valid[c], i);
            if (ok[c]) v[c] = p.val[c][i * p.val_stride[c]];
        }
        gb_row_to_global(p, key, key_null, ok, v);
    }
}

// shared table layout (dynamic smem): keys u64[S] | sum f64[3][S] | min u64[3][S] | max u64[3][S] (if minmax)
//                                     | rows u32[S] | cnt u32[3][S]
template <bool MINMAX>
__global__ void __launch_bounds__(GB_THREADS) groupby_smem_kernel(const GroupByParams p) {
    extern __shared__ __align__(16) unsigned char gb_smem[];
    const int S = p.smem_slots;
    const int V = p.ncols;
    uint64_t* skeys = reinterpret_cast<uint64_t*>(gb_smem);
    double* ssum = reinterpret_cast<double*>(skeys + S);
    uint64_t* smin = reinterpret_cast<uint64_t*>(ssum + (size_t)V * S);
    uint64_t* smax = smin + (MINMAX ? (size_t)V * S : 0);
    uint32_t* srows = reinterpret_cast<uint32_t*>(smax + (MINMAX ? (size_t)V * S : 0));
    uint32_t* scnt = srows + S;
    for (int i = threadIdx.x; i < S; i += GB_THREADS) {
        skeys[i] = GB_EMPTY;
        srows[i] = 0;
        for (int c = 0; c < V; ++c) {
            ssum[c * S + i] = 0.0;
            scnt[c * S + i] = 0;
            if (MINMAX) { smin[c * S + i] = ~0ull; smax[c * S + i] = 0ull; }
        }
    }
    __syncthreads();
    const uint32_t smask = (uint32_t)S - 1;
    for (int64_t i = (int64_t)blockIdx.x * GB_THREADS + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * GB_THREADS) {
        const bool key_null = !gb_bit(p.key_valid, i);
        const uint64_t key = (uint64_t)gb_load_key(p, i);
        if (!key_null && key == GB_EMPTY) {
            atomicAdd(p.status + 1, 1ull);
            continue;
        }
        bool ok[3] = {false, false, false};
        double v[3] = {0.0, 0.0, 0.0};
        for (int c = 0; c < V; ++c) {
            ok[c] = gb_bit(p.val_valid[c], i);
            if (ok[c]) v[c] = p.val[c][i * p.val_stride[c]];
        }
        int slot = -1;
        if (!key_null) {
            uint32_t s = (uint32_t)(mix64(key) >> 40) & smask;
            for (int probe = 0; probe < GB_MAX_PROBES_SMEM; ++probe) {
                uint64_t cur = skeys[s];
                if (cur == GB_EMPTY)
                    cur = atomicCAS(reinterpret_cast<unsigned long long*>(skeys + s), (unsigned long long)GB_EMPTY,
                                    (unsigned long long)key);
                if (cur == GB_EMPTY || cur == key) { slot = (int)s; break; }
                s = (s + 1) & smask;
            }
        }
        if (slot < 0) {  // null key, or no shared slot within the probe limit: spill to the global table
            gb_row_to_global(p, key, key_null, ok, v);
            continue;
        }
        atomicAdd(srows + slot, 1u);
        for (int c = 0; c < V; ++c) {
            if (!ok[c]) continue;
            atomicAdd(scnt + c * S + slot, 1u);
            atomicAdd(ssum + c * S + slot, v[c]);
            if (MINMAX) {
                const uint64_t e = dbl_sortable(v[c]);
                atomicMin(reinterpret_cast<unsigned long long*>(smin + c * S + slot), (unsigned long long)e);
                atomicMax(reinterpret_cast<unsigned long long*>(smax + c * S + slot), (unsigned long long)e);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < S; i += GB_THREADS) {
        const uint64_t key = skeys[i];
        if (key == GB_EMPTY) continue;
        uint64_t* e = gb_global_slot(p, key);
        if (!e) {
            atomicAdd(p.status, (unsigned long long)srows[i]);
            continue;
        }
        unsigned long long cnt[3] = {0, 0, 0};
        double sum[3] = {0, 0, 0};
        uint64_t mn[3] = {~0ull, ~0ull, ~0ull}, mx[3] = {0, 0, 0};
        for (int c = 0; c < V; ++c) {
            cnt[c] = scnt[c * S + i];
            sum[c] = ssum[c * S + i];
            if (MINMAX) { mn[c] = smin[c * S + i]; mx[c] = smax[c * S + i]; }
        }
        gb_global_add(e, p, srows[i], cnt, sum, mn, mx);
    }
}

// Low cardinality with a known small key range (dictionary codes, bucket ids,
// config 3's keys): the shared table is indexed by the key itself - no hashing,
// no probing, no key compare.  Per row: one native 32-bit shared atomic for the
// row count and, per value column, one fp64 shared atomic add (or, for a null
// cell, a native 32-bit add to the column's null counter: count = rows - nulls).
// 256 threads x 4 rows per tile with 128-bit loads; slot R collects null keys.
template <int NC>
__global__ void __launch_bounds__(GB_THREADS) groupby_dense_kernel(const GroupByParams p) {
    extern __shared__ __align__(16) unsigned char gb_smem[];
    const int R = p.key_range;          // slots 0..R-1 = keys, slot R = null key
    const int S = R + 1;
    double* ssum = reinterpret_cast<double*>(gb_smem);
    uint32_t* srows = reinterpret_cast<uint32_t*>(ssum + (size_t)NC * S);
    uint32_t* snull = srows + S;
    for (int i = threadIdx.x; i < S; i += GB_THREADS) {
        srows[i] = 0;
#pragma unroll
        for (int c = 0; c < NC; ++c) { ssum[c * S + i] = 0.0; snull[c * S + i] = 0; }
    }
    __syncthreads();
    const int32_t* keys = static_cast<const int32_t*>(p.keys);
    const int64_t tiles = (p.n + GB_THREADS * 4 - 1) / (GB_THREADS * 4);
    for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int64_t row = t * GB_THREADS * 4 + (int64_t)threadIdx.x * 4;
        if (row >= p.n) continue;
        const bool full = row + 4 <= p.n;
        int32_t k[4];
        double v[NC][4];
        unsigned vn[NC];
        if (full) {
            const int4 q = ld_stream_v4(keys + row);
            k[0] = q.x; k[1] = q.y; k[2] = q.z; k[3] = q.w;
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const double2 a = ld_stream_d2(p.val[c] + row), b = ld_stream_d2(p.val[c] + row + 2);
                v[c][0] = a.x; v[c][1] = a.y; v[c][2] = b.x; v[c][3] = b.y;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool in = row + j < p.n;
                k[j] = in ? keys[row + j] : 0;
#pragma unroll
                for (int c = 0; c < NC; ++c) v[c][j] = in ? p.val[c][row + j] : 0.0;
            }
        }
        const int left = p.n - row < 4 ? (int)(p.n - row) : 4;
        const unsigned in_mask = (1u << left) - 1;
        unsigned kn = 0xF;
        if (p.key_valid) kn = (__ldg(p.key_valid + (row >> 5)) >> (row & 31)) & 0xF;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            vn[c] = 0xF;
            if (p.val_valid[c]) vn[c] = (__ldg(p.val_valid[c] + (row >> 5)) >> (row & 31)) & 0xF;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (!((in_mask >> j) & 1u)) continue;
            const bool key_null = !((kn >> j) & 1u);
            if (!key_null && (k[j] < 0 || k[j] >= R)) {  // outside the promised range: straight to the global table
                bool ok[3] = {false, false, false};
                double x[3] = {0.0, 0.0, 0.0};
#pragma unroll
                for (int c = 0; c < NC; ++c) { ok[c] = (vn[c] >> j) & 1u; x[c] = v[c][j]; }
                gb_row_to_global(p, (uint64_t)(int64_t)k[j], false, ok, x);
                continue;
            }
            const int slot = key_null ? R : k[j];
            atomicAdd(srows + slot, 1u);
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                if ((vn[c] >> j) & 1u) atomicAdd(ssum + c * S + slot, v[c][j]);
                else atomicAdd(snull + c * S + slot, 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < S; i += GB_THREADS) {
        const uint32_t rows = srows[i];
        if (!rows) continue;
        uint64_t* e = i == R ? p.table + (p.mask + 1) * p.stride : gb_global_slot(p, (uint64_t)(int64_t)i);
        if (!e) {
            atomicAdd(p.status, (unsigned long long)rows);
            continue;
        }
        unsigned long long cnt[3] = {0, 0, 0};
        double sum[3] = {0, 0, 0};
        const uint64_t mn[3] = {~0ull, ~0ull, ~0ull}, mx[3] = {0, 0, 0};
#pragma unroll
        for (int c = 0; c < NC; ++c) { cnt[c] = rows - snull[c * S + i]; sum[c] = ssum[c * S + i]; }
        gb_global_add(e, p, rows, cnt, sum, mn, mx);
    }
}

// The table as a flat array of 16-byte vectors (every entry is an even number of words): vector v
// holds words 2v, 2v + 1 of entry v / (stride / 2) - fully coalesced stores.
__global__ void __launch_bounds__(256) groupby_init_kernel(uint64_t* table, int64_t entries, int32_t stride,
                                                           int32_t ncols, int32_t minmax) {
    const int vpe = stride / 2;  // vectors per entry
    const int64_t total = entries * vpe;
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < total; v += (int64_t)gridDim.x * blockDim.x) {
        const int f = (int)(v % vpe) * 2;  // first field of this vector
        ulonglong2 w = make_ulonglong2(0ull, 0ull);
        if (f == 0) w.x = GB_EMPTY;                                  // {key, rows}
        else if (minmax && f >= 2 + 2 * ncols) { w.x = ~0ull; w.y = 0ull; }  // {min, max} in sortable form
        reinterpret_cast<ulonglong2*>(table)[v] = w;
    }
}

// used entries -> dense columnar arrays (order = slot order; callers sort by key)
__global__ void __launch_bounds__(256) groupby_compact_kernel(const uint64_t* __restrict__ table, int64_t capacity,
                                                              int32_t stride, int32_t ncols, int32_t minmax,
                                                              int64_t* out_keys, int64_t* out_rows, int64_t* out_cnt,
                                                              double* out_sum, double* out_min, double* out_max,
                                                              int64_t out_capacity, unsigned long long* d_groups) {
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < capacity; s += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t* e = table + s * stride;
        if (e[0] == GB_EMPTY) continue;
        // warp-aggregated append
        const unsigned active = __activemask();
        const int leader = __ffs(active) - 1;
        unsigned long long base = 0;
        if ((threadIdx.x & 31) == leader) base = atomicAdd(d_groups, (unsigned long long)__popc(active));
        base = __shfl_sync(active, base, leader);
        const int64_t g = (int64_t)base + __popc(active & lanemask_lt());
        if (g >= out_capacity) continue;
        out_keys[g] = (int64_t)e[0];
        out_rows[g] = (int64_t)e[1];
        for (int c = 0; c < ncols; ++c) {
            out_cnt[c * out_capacity + g] = (int64_t)(e[1] - e[2 + 2 * c]);
            out_sum[c * out_capacity + g] = __longlong_as_double((long long)e[3 + 2 * c]);
            if (minmax) {
                const bool any = e[1] != e[2 + 2 * c];
                const double nanv = __longlong_as_double(0x7ff8000000000000ll);
                out_min[c * out_capacity + g] = any ? dbl_from_sortable(e[2 + 2 * ncols + 2 * c]) : nanv;
                out_max[c * out_capacity + g] = any ? dbl_from_sortable(e[3 + 2 * ncols + 2 * c]) : nanv;
            }
        }
    }
}

// ---------------------------------------------------------------- exchange
// Hash partition of (key, value) rows into P destination ranges of one output
// buffer of 16-byte records {key, value bits} - the send buffer of the
// all-to-all.  dest = (hi32(mix64(key)) * P) >> 32.
__global__ void __launch_bounds__(PART_THREADS) partition_count_kernel(const int64_t* __restrict__ keys,
                                                                       int64_t key_stride, int64_t n, int ranks,
                                                                       int buckets, unsigned long long* counts) {
    const int parts = ranks * buckets;
    __shared__ unsigned int s_cnt[PART_MAX];
    if (threadIdx.x < PART_MAX) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * PART_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * PART_THREADS) {
        const int d = part_dest((uint64_t)keys[i * key_stride], ranks, buckets);
        atomicAdd(&s_cnt[d & (PART_MAX - 1)], 1u);
    }
    __syncthreads();
    if (threadIdx.x < parts && s_cnt[threadIdx.x]) atomicAdd(counts + threadIdx.x, (unsigned long long)s_cnt[threadIdx.x]);
}

// exclusive scan of the destination counts
__global__ void partition_scan_kernel(const int64_t* counts, int64_t* base, int parts) {
    int64_t run = 0;
    for (int d = 0; d < parts; ++d) { base[d] = run; run += counts[d]; }
}

__global__ void __launch_bounds__(PART_THREADS, 4) partition_scatter_kernel(const int64_t* __restrict__ keys,
                                                                         int64_t key_stride,
                                                                         const double* __restrict__ vals,
                                                                         int64_t val_stride, int64_t n, int ranks,
                                                                         int buckets, const int64_t* __restrict__ part_base,
                                                                         unsigned long long* cursor, const PartDest dst) {
    constexpr int PER = PART_TILE / PART_THREADS;
    constexpr int THREADS = PART_THREADS;
    const int parts = ranks * buckets;
    __shared__ unsigned int s_cnt[PART_MAX], s_off[PART_MAX];
    __shared__ long long s_delta[PART_MAX];   // record index in dst.base[d] of staged position 0
    __shared__ int64_t s_key[PART_TILE];
    __shared__ double s_val[PART_TILE];
    __shared__ uint8_t s_dest[PART_TILE];
    const int64_t tiles = (n + PART_TILE - 1) / PART_TILE;
    for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        __syncthreads();
        if (threadIdx.x < PART_MAX) s_cnt[threadIdx.x] = 0;
        __syncthreads();
        const int64_t row0 = t * PART_TILE;
        const int rows = n - row0 < PART_TILE ? (int)(n - row0) : PART_TILE;
        // pass 1: destination and rank within (tile, destination)
        unsigned rank[PER];
        int dest[PER];
        int64_t k[PER];
        double v[PER];
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int r = j * THREADS + threadIdx.x;
            dest[j] = -1;
            if (r < rows) {
                k[j] = keys[(row0 + r) * key_stride];
                v[j] = vals[(row0 + r) * val_stride];
                dest[j] = part_dest((uint64_t)k[j], ranks, buckets);
                rank[j] = atomicAdd(&s_cnt[dest[j]], 1u);
            }
        }
        __syncthreads();
        // reserve this tile's run in every destination (atomics in flight while the offsets are scanned)
        unsigned long long gb = 0;
        if (threadIdx.x < parts && s_cnt[threadIdx.x])
            gb = atomicAdd(cursor + threadIdx.x, (unsigned long long)s_cnt[threadIdx.x]);
        if (parts <= 32) {
            if (threadIdx.x == 0) {
                unsigned run = 0;
                for (int d = 0; d < parts; ++d) { s_off[d] = run; run += s_cnt[d]; }
            }
        } else if (threadIdx.x < 32) {  // one warp, PART_MAX / 32 destinations per lane
            unsigned tot = 0;
#pragma unroll
            for (int q = 0; q < PART_MAX / 32; ++q) tot += s_cnt[threadIdx.x * (PART_MAX / 32) + q];
            unsigned incl = tot;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned up = __shfl_up_sync(0xffffffffu, incl, d);
                if ((int)threadIdx.x >= d) incl += up;
            }
            unsigned run = incl - tot;
#pragma unroll
            for (int q = 0; q < PART_MAX / 32; ++q) {
                s_off[threadIdx.x * (PART_MAX / 32) + q] = run;
                run += s_cnt[threadIdx.x * (PART_MAX / 32) + q];
            }
        }
        __syncthreads();
        if (threadIdx.x < parts)
            s_delta[threadIdx.x] = (long long)part_base[threadIdx.x] + (long long)gb - (long long)s_off[threadIdx.x];
        // pass 2: stage by destination, then write each destination's run contiguously
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            if (dest[j] < 0) continue;
            const unsigned pos = s_off[dest[j]] + rank[j];
            s_key[pos] = k[j];
            s_val[pos] = v[j];
            s_dest[pos] = (uint8_t)dest[j];
        }
        __syncthreads();
        for (int i = threadIdx.x; i < rows; i += THREADS) {
            const int d = s_dest[i];
            const int64_t g = (int64_t)s_delta[d] + i;
            longlong2 rec;
            rec.x = s_key[i];
            rec.y = __double_as_longlong(s_val[i]);
            *reinterpret_cast<longlong2*>(dst.base[d] + 2 * g) = rec;
        }
    }
}

}  // namespace etl

using etl::Handle;

// The scatter pass: input by bulk async copies (scatter.cuh) when the rows are two contiguous
// columns or {key, value} records; the register-staged kernel takes every other stride / alignment.
static cudaError_t launch_partition_scatter(const Handle* h, const int64_t* keys, int64_t key_stride, const double* vals,
                                            int64_t val_stride, int64_t n, int ranks, int buckets, const int64_t* part_base,
                                            unsigned long long* cursor, const etl::PartDest& dst, cudaStream_t stream) {
    static const int pipe = [] { const char* e = getenv("ETL_PART_PIPE"); return e ? atoi(e) : 1; }();
    const bool columns = key_stride == 1 && val_stride == 1 && etl::aligned16(keys) && etl::aligned16(vals);
    const bool records = key_stride == 2 && val_stride == 2 && etl::aligned16(keys) &&
                         reinterpret_cast<const int64_t*>(vals) == keys + 1;
    if (pipe && (columns || records)) {
        etl::ScatterJob job{};
        job.keys = keys; job.vals = vals; job.n = n; job.ranks = ranks; job.buckets = buckets;
        job.parts = ranks * buckets; job.dest_base = part_base; job.cursor = cursor;
        return columns ? etl::launch_scatter_pipelined<false>(h->sm_count, job, dst, 1, n, stream)
                       : etl::launch_scatter_pipelined<true>(h->sm_count, job, dst, 1, n, stream);
    }
    const int64_t tiles = (n + etl::PART_TILE - 1) / etl::PART_TILE;
    const int64_t g = tiles < (int64_t)h->sm_count * 4 ? tiles : (int64_t)h->sm_count * 4;
    etl::partition_scatter_kernel<<<(unsigned)g, etl::PART_THREADS, 0, stream>>>(keys, key_stride, vals, val_stride, n, ranks,
                                                                                buckets, part_base, cursor, dst);
    return cudaGetLastError();
}

static inline int gb_stride(int ncols, int minmax) { return 2 + 2 * ncols + (minmax ? 2 * ncols : 0); }

extern "C" int etl_groupby_table_words(int32_t ncols, int32_t minmax, int64_t capacity, int64_t* words) {
    ETL_REQUIRE(words && ncols >= 0 && ncols <= 3 && capacity >= 2 && (capacity & (capacity - 1)) == 0,
                "groupby_table_words: ncols in 0..3 and a power-of-two capacity required");
    *words = (capacity + 1) * gb_stride(ncols, minmax);
    return ETL_OK;
}

extern "C" int etl_groupby_init(etl_handle* hh, uint64_t* table, int64_t capacity, int32_t ncols, int32_t minmax,
                                void* stream_) {
    ETL_REQUIRE(hh && table, "groupby_init: NULL argument");
    ETL_REQUIRE(ncols >= 0 && ncols <= 3 && capacity >= 2 && (capacity & (capacity - 1)) == 0,
                "groupby_init: ncols in 0..3 and a power-of-two capacity required");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_REQUIRE(etl::aligned16(table), "groupby_init: table must be 16-byte aligned");
    int64_t blocks = ((capacity + 1) * (gb_stride(ncols, minmax) / 2) + 255) / 256;
    if (blocks > (int64_t)h->sm_count * 16) blocks = (int64_t)h->sm_count * 16;
    etl::groupby_init_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(table, capacity + 1,
                                                                                             gb_stride(ncols, minmax),
                                                                                             ncols, minmax);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_groupby_accumulate(etl_handle* hh, const void* keys, int32_t key_dtype, int64_t key_stride,
                                      const uint32_t* key_valid, int32_t ncols, const double* const val[],
                                      const int64_t val_stride[], const uint32_t* const val_valid[], int64_t n,
                                      uint64_t* table, int64_t capacity, int32_t minmax, int64_t groups_hint,
                                      int32_t key_range, int64_t* d_status, void* stream_) {
    ETL_REQUIRE(hh && table && d_status, "groupby_accumulate: NULL argument");
    ETL_REQUIRE(ncols >= 0 && ncols <= 3, "groupby_accumulate: ncols=%d not in 0..3", ncols);
    ETL_REQUIRE(key_dtype == ETL_I32 || key_dtype == ETL_I64, "groupby_accumulate: key_dtype=%d", key_dtype);
    ETL_REQUIRE(capacity >= 2 && (capacity & (capacity - 1)) == 0, "groupby_accumulate: capacity must be a power of two");
    ETL_REQUIRE(n >= 0 && key_stride >= 1, "groupby_accumulate: bad n / key_stride");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(keys != nullptr, "groupby_accumulate: keys is NULL");
    Handle* h = reinterpret_cast<Handle*>(hh);
    etl::GroupByParams p{};
    p.keys = keys; p.key_dtype = key_dtype; p.key_stride = key_stride; p.key_valid = key_valid;
    p.ncols = ncols; p.minmax = minmax ? 1 : 0; p.n = n; p.table = table; p.mask = (uint64_t)capacity - 1;
    p.stride = gb_stride(ncols, minmax);
    p.status = reinterpret_cast<unsigned long long*>(d_status);
    for (int c = 0; c < ncols; ++c) {
        ETL_REQUIRE(val && val[c], "groupby_accumulate: val[%d] is NULL", c);
        p.val[c] = val[c];
        p.val_stride[c] = val_stride ? val_stride[c] : 1;
        p.val_valid[c] = val_valid ? val_valid[c] : nullptr;
    }
    ETL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int64_t want = (n + etl::GB_THREADS - 1) / etl::GB_THREADS;
    // shared-memory path when the groups (at load <= 0.5) fit a per-CTA table of at most 4096 slots
    int slots = 256;
    const int64_t hint = groups_hint > 0 ? groups_hint : 1024;
    while (slots < 2 * hint && slots < 8192) slots *= 2;
    const size_t per_slot = 8 + 4 + (size_t)ncols * (8 + 4) + (minmax ? (size_t)ncols * 16 : 0);
    const bool use_smem = hint <= 2048 && slots * per_slot <= 200 * 1024;
    bool dense_ok = key_range > 0 && key_range <= 8192 && key_dtype == ETL_I32 && key_stride == 1 && !minmax && ncols >= 1 &&
                    etl::aligned16(keys);
    for (int c = 0; c < ncols && dense_ok; ++c) dense_ok = p.val_stride[c] == 1 && etl::aligned16(p.val[c]);
    const size_t dense_smem = ((size_t)key_range + 1) * (4 + (size_t)ncols * 12);
    if (dense_ok && dense_smem <= 100 * 1024) {
        p.key_range = key_range;
        auto kern = ncols == 1 ? etl::groupby_dense_kernel<1> : (ncols == 2 ? etl::groupby_dense_kernel<2> : etl::groupby_dense_kernel<3>);
        ETL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dense_smem));
        int occ = 0;
        ETL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, etl::GB_THREADS, dense_smem));
        if (occ < 1) occ = 1;
        const int64_t tiles = (n + etl::GB_THREADS * 4 - 1) / (etl::GB_THREADS * 4);
        int64_t blocks = (int64_t)h->sm_count * occ;
        if (blocks > tiles) blocks = tiles;
        kern<<<(unsigned)blocks, etl::GB_THREADS, dense_smem, stream>>>(p);
    } else if (use_smem) {
        p.smem_slots = slots;
        const size_t smem = slots * per_slot;
        auto kern = minmax ? etl::groupby_smem_kernel<true> : etl::groupby_smem_kernel<false>;
        ETL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        ETL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, etl::GB_THREADS, smem));
        if (occ < 1) occ = 1;
        int64_t blocks = (int64_t)h->sm_count * occ;
        if (blocks > want) blocks = want;
        kern<<<(unsigned)blocks, etl::GB_THREADS, smem, stream>>>(p);
    } else {
        int64_t blocks = (int64_t)h->sm_count * 8;
        if (blocks > want) blocks = want;
        etl::groupby_global_kernel<<<(unsigned)blocks, etl::GB_THREADS, 0, stream>>>(p);
    }
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_groupby_compact(etl_handle* hh, const uint64_t* table, int64_t capacity, int32_t ncols,
                                   int32_t minmax, int64_t out_capacity, int64_t* out_keys, int64_t* out_rows,
                                   int64_t* out_count, double* out_sum, double* out_min, double* out_max,
                                   int64_t* d_groups, void* stream_) {
    ETL_REQUIRE(hh && table && out_keys && out_rows && d_groups, "groupby_compact: NULL argument");
    ETL_REQUIRE(ncols == 0 || (out_count && out_sum), "groupby_compact: out_count/out_sum is NULL");
    ETL_REQUIRE(!minmax || ncols == 0 || (out_min && out_max), "groupby_compact: out_min/out_max is NULL");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_groups, 0, sizeof(int64_t), stream));
    int64_t blocks = (capacity + 255) / 256;
    if (blocks > (int64_t)h->sm_count * 16) blocks = (int64_t)h->sm_count * 16;
    etl::groupby_compact_kernel<<<(unsigned)blocks, 256, 0, stream>>>(
        table, capacity, gb_stride(ncols, minmax), ncols, minmax ? 1 : 0, out_keys, out_rows, out_count, out_sum, out_min,
        out_max, out_capacity, reinterpret_cast<unsigned long long*>(d_groups));
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_hash_partition(etl_handle* hh, const int64_t* keys, int64_t key_stride, const double* vals,
                                  int64_t val_stride, int64_t n, int32_t ranks, int32_t buckets, int64_t* d_part_counts,
                                  int64_t* d_part_base, int64_t* out_records, void* stream_) {
    ETL_REQUIRE(hh && d_part_counts && d_part_base, "hash_partition: NULL argument");
    ETL_REQUIRE(ranks >= 1 && buckets >= 1 && (int64_t)ranks * buckets <= etl::PART_MAX,
                "hash_partition: ranks=%d x buckets=%d must be in 1..%d", ranks, buckets, etl::PART_MAX);
    ETL_REQUIRE(n >= 0 && key_stride >= 1 && val_stride >= 1, "hash_partition: bad n / stride");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int parts = ranks * buckets;
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_part_counts, 0, sizeof(int64_t) * parts, stream));
    if (n == 0) {
        ETL_CUDA(cudaMemsetAsync(d_part_base, 0, sizeof(int64_t) * parts, stream));
        return ETL_OK;
    }
    ETL_REQUIRE(keys && vals && out_records && etl::aligned16(out_records), "hash_partition: NULL or misaligned buffer");
    int rc = etl::arena_reserve(h, sizeof(unsigned long long) * etl::PART_MAX);
    if (rc != ETL_OK) return rc;
    unsigned long long* cursor = static_cast<unsigned long long*>(h->scratch);
    ETL_CUDA(cudaMemsetAsync(cursor, 0, sizeof(unsigned long long) * etl::PART_MAX, stream));
    int64_t blocks = (n + etl::PART_THREADS * 8 - 1) / (etl::PART_THREADS * 8);
    if (blocks > (int64_t)h->sm_count * 8) blocks = (int64_t)h->sm_count * 8;
    etl::partition_count_kernel<<<(unsigned)blocks, etl::PART_THREADS, 0, stream>>>(
        keys, key_stride, n, ranks, buckets, reinterpret_cast<unsigned long long*>(d_part_counts));
    ETL_CUDA(cudaGetLastError());
    etl::partition_scan_kernel<<<1, 1, 0, stream>>>(d_part_counts, d_part_base, parts);
    ETL_CUDA(cudaGetLastError());
    etl::PartDest dst;
    for (int d = 0; d < etl::PART_MAX; ++d) dst.base[d] = out_records;
    ETL_CUDA(launch_partition_scatter(h, keys, key_stride, vals, val_stride, n, ranks, buckets, d_part_base, cursor, dst, stream));
    return ETL_OK;
}

extern "C" int etl_hash_partition_count(etl_handle* hh, const int64_t* keys, int64_t key_stride, int64_t n, int32_t ranks,
                                        int32_t buckets, int64_t* d_part_counts, void* stream_) {
    ETL_REQUIRE(hh && d_part_counts, "hash_partition_count: NULL argument");
    ETL_REQUIRE(ranks >= 1 && buckets >= 1 && (int64_t)ranks * buckets <= etl::PART_MAX,
                "hash_partition_count: ranks=%d x buckets=%d must be in 1..%d", ranks, buckets, etl::PART_MAX);
    ETL_REQUIRE(n >= 0 && key_stride >= 1 && (n == 0 || keys), "hash_partition_count: bad n / stride / keys");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_part_counts, 0, sizeof(int64_t) * ranks * buckets, stream));
    if (n == 0) return ETL_OK;
    int64_t blocks = (n + etl::PART_THREADS * 8 - 1) / (etl::PART_THREADS * 8);
    if (blocks > (int64_t)h->sm_count * 8) blocks = (int64_t)h->sm_count * 8;
    etl::partition_count_kernel<<<(unsigned)blocks, etl::PART_THREADS, 0, stream>>>(
        keys, key_stride, n, ranks, buckets, reinterpret_cast<unsigned long long*>(d_part_counts));
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_hash_partition_scatter_peers(etl_handle* hh, const int64_t* keys, int64_t key_stride,
                                                const double* vals, int64_t val_stride, int64_t n, int32_t ranks,
                                                int32_t buckets, int64_t* const* peer_records,
                                                const int64_t* d_peer_offset, void* stream_) {
    ETL_REQUIRE(hh && peer_records && d_peer_offset, "hash_partition_scatter_peers: NULL argument");
    ETL_REQUIRE(ranks >= 1 && buckets >= 1 && (int64_t)ranks * buckets <= etl::PART_MAX,
                "hash_partition_scatter_peers: ranks=%d x buckets=%d must be in 1..%d", ranks, buckets, etl::PART_MAX);
    ETL_REQUIRE(n >= 0 && key_stride >= 1 && val_stride >= 1, "hash_partition_scatter_peers: bad n / stride");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(keys && vals, "hash_partition_scatter_peers: keys / vals is NULL");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    etl::PartDest dst;
    for (int d = 0; d < etl::PART_MAX; ++d) dst.base[d] = nullptr;
    for (int d = 0; d < ranks; ++d) {
        ETL_REQUIRE(peer_records[d] && etl::aligned16(peer_records[d]), "hash_partition_scatter_peers: peer buffer %d NULL or misaligned", d);
        for (int b = 0; b < buckets; ++b) dst.base[d * buckets + b] = peer_records[d];
    }
    ETL_CUDA(cudaSetDevice(h->device));
    int rc = etl::arena_reserve(h, sizeof(unsigned long long) * etl::PART_MAX);
    if (rc != ETL_OK) return rc;
    unsigned long long* cursor = static_cast<unsigned long long*>(h->scratch);
    ETL_CUDA(cudaMemsetAsync(cursor, 0, sizeof(unsigned long long) * etl::PART_MAX, stream));
    ETL_CUDA(launch_partition_scatter(h, keys, key_stride, vals, val_stride, n, ranks, buckets, d_peer_offset, cursor, dst,
                                      stream));
    return ETL_OK;
}
