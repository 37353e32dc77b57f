// Final merge of the shuffle (config 4): {key, value} records -> groupBy(key).agg(count, sum)
// with every table update in SHARED memory.
//
// The table-in-HBM merge (groupby_global_kernel, bucket by bucket so the slice being updated is
// L2-resident) costs three or four L2 atomics per row and runs at ~15 G rows/s however the buckets
// are cut.  Here the records are partitioned once more, into "fine" buckets small enough that all
// groups of one bucket fit an open-addressing table in one CTA's shared memory:
//
//   level A   etl_hash_partition(ranks = 1, buckets = B)            (groupby.cu)
//   level B   etl_hash_subpartition: every level-A bucket into `sub` pieces, one launch for all
//             buckets (grid.y = bucket; the scatter pass is scatter.cuh's pipelined kernel); fine index = umulhi(lo32(mix64(key)), B * sub), so
//             fine / sub is the level-A bucket and the fine buckets of bucket b are the output
//             range level A gave b - the two levels compose into one ordering;
//   merge     etl_groupby_merge_partitioned: persistent CTAs, one fine bucket at a time: clear the
//             shared table, stream the bucket's records through it (64-bit CAS claims a key slot,
//             native 32-bit atomic counts the row, fp64 shared atomic adds the value), then append
//             the used slots to dense columnar output at a range reserved with ONE global atomic.
//
// HBM traffic per record: 16 read + 16 written per partition level, 16 read by the merge, 32 written
// per GROUP - all of it streaming, none of it atomic.  A record whose key finds no free slot within
// the probe limit (a bucket with more groups than the table holds: skew, or a groups hint far too
// low) is appended to a spill list, which the host aggregates with the ordinary global-table path;
// a key is either entirely in shared memory or entirely spilled (slots never empty again, so every
// row of a key ends the same probe sequence the same way), hence the two outputs are disjoint.
#include "common.cuh"
#include "scatter.cuh"

namespace etl {

constexpr uint64_t MG_EMPTY = 0x8000000000000000ull;  // same reserved key as the global table
constexpr int SUB_TILE = 2048;
constexpr int SUB_THREADS = 256;
constexpr int SUB_MAX = 256;         // pieces per level-A bucket
constexpr int MG_THREADS = 640;         // a launch that has the SMs to itself: 2 x 640 threads x 48 registers = 60 K of 64 K
constexpr int MG_THREADS_SHARED = 512;  // a launch beside the exchange's copy kernel: 48 K, room for the copy CTAs (slabs.cu)
constexpr int MG_MAX_PROBES = 48;
constexpr int MG_UNROLL = 4;

__device__ __forceinline__ longlong2 ld_record(const longlong2* p) {
    longlong2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.s64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
    return r;
}

// ---------------------------------------------------------------- level B
__global__ void __launch_bounds__(SUB_THREADS) subpartition_count_kernel(const longlong2* __restrict__ rec,
                                                                         const int64_t* __restrict__ bucket_base,
                                                                         const int64_t* __restrict__ bucket_counts,
                                                                         int sub, unsigned fine,
                                                                         unsigned long long* fine_counts) {
    __shared__ unsigned int s_cnt[SUB_MAX];
    const int b = blockIdx.y;
    const int64_t lo = bucket_base[b], cnt = bucket_counts[b];
    s_cnt[threadIdx.x] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * SUB_THREADS + threadIdx.x; i < cnt; i += (int64_t)gridDim.x * SUB_THREADS) {
        const unsigned d = fine_index((uint64_t)rec[lo + i].x, fine) - (unsigned)b * (unsigned)sub;
        atomicAdd(&s_cnt[d & (SUB_MAX - 1)], 1u);
    }
    __syncthreads();
    if (threadIdx.x < sub && s_cnt[threadIdx.x])
        atomicAdd(fine_counts + (size_t)b * sub + threadIdx.x, (unsigned long long)s_cnt[threadIdx.x]);
}

// exclusive scan of up to 65536 counts by one CTA of 1024 threads
__global__ void __launch_bounds__(1024) fine_scan_kernel(const int64_t* __restrict__ counts, int64_t* __restrict__ base, int n) {
    __shared__ int64_t s_warp[32];
    const int per = (n + 1023) / 1024;
    const int lo = threadIdx.x * per;
    int64_t mine = 0;
    for (int i = lo; i < lo + per && i < n; ++i) mine += counts[i];
    int64_t incl = mine;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int64_t up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int64_t w = s_warp[lane], wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int64_t up = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += up;
        }
        s_warp[lane] = wi - w;
    }
    __syncthreads();
    int64_t run = s_warp[warp] + incl - mine;
    for (int i = lo; i < lo + per && i < n; ++i) {
        base[i] = run;
        run += counts[i];
    }
}

// ---------------------------------------------------------------- merge
// shared table (dynamic smem): key u64[S] | sum f64[S] | [min u64[S] | max u64[S] (MINMAX, sortable form)] | rows u32[S]
struct MergeParams {
    const longlong2* rec;
    const int64_t* fine_base;
    const int64_t* fine_counts;
    int fine;
    int piece0;                // first piece of this launch (pieces piece0 .. piece0 + fine - 1)
    int64_t piece_cap;         // > 0: piece f is slab f (piece_cap records apart) holding min(fine_counts[f], piece_cap) records
    int slots;                 // S, a power of two
    int64_t* out_keys;
    int64_t* out_rows;
    int64_t* out_count;
    double* out_sum;
    double* out_min;              // MINMAX only
    double* out_max;
    int64_t out_capacity;
    unsigned long long* groups;   // [0] groups appended (may exceed out_capacity: the caller checks)
    longlong2* spill;
    int64_t spill_capacity;
    unsigned long long* spilled;  // [0] records appended to `spill` (may exceed spill_capacity)
    unsigned long long* status;   // [1] records with the reserved key
};

template <bool MINMAX>
__device__ __forceinline__ void merge_one(const MergeParams& p, uint64_t* skey, double* ssum, uint64_t* smin, uint64_t* smax,
                                          uint32_t* srows, uint32_t mask, const longlong2 r) {
    const uint64_t key = (uint64_t)r.x;
    if (key == MG_EMPTY) {
        atomicAdd(p.status + 1, 1ull);
        return;
    }
    // slot from the HIGH hash word: the low word's top bits are what the bucket was cut by
    uint32_t s = (uint32_t)(mix64(key) >> 32) & mask;
    for (int probe = 0; probe < MG_MAX_PROBES; ++probe) {
        uint64_t cur = skey[s];
        if (cur == MG_EMPTY)
            cur = atomicCAS(reinterpret_cast<unsigned long long*>(skey + s), (unsigned long long)MG_EMPTY,
                            (unsigned long long)key);
        if (cur == MG_EMPTY || cur == key) {
            atomicAdd(srows + s, 1u);
            atomicAdd(ssum + s, __longlong_as_double(r.y));
            if (MINMAX) {
                const unsigned long long e = dbl_sortable(__longlong_as_double(r.y));
                atomicMin(reinterpret_cast<unsigned long long*>(smin + s), e);
                atomicMax(reinterpret_cast<unsigned long long*>(smax + s), e);
            }
            return;
        }
        s = (s + 1) & mask;
    }
    const unsigned long long at = atomicAdd(p.spilled, 1ull);
    if ((int64_t)at < p.spill_capacity) p.spill[at] = r;
}

// 48 registers, two CTAs per SM.  THREADS = 512: 48 K of the SM's 64 K registers, which leaves room for the exchange's
// copy CTAs (slabs.cu) to sit beside the merge of the bucket groups that have already landed (measured at 2 GPUs:
// 12.5 ms, against 13.1 with the 640-thread CTAs that keep the copy kernel off the SM).  THREADS = 640: 40 warps per SM
// instead of 32 for the shared-memory atomics' latency - config 4 on one GPU 7.9 -> 7.4 ms.
template <bool MINMAX, int THREADS>
__global__ void __maxnreg__(48) groupby_merge_kernel(const MergeParams p) {
    extern __shared__ __align__(16) unsigned char mg_smem[];
    const int S = p.slots;
    uint64_t* skey = reinterpret_cast<uint64_t*>(mg_smem);
    double* ssum = reinterpret_cast<double*>(skey + S);
    uint64_t* smin = reinterpret_cast<uint64_t*>(ssum + S);
    uint64_t* smax = smin + (MINMAX ? S : 0);
    uint32_t* srows = reinterpret_cast<uint32_t*>(smax + (MINMAX ? S : 0));
    __shared__ unsigned s_warp[THREADS / 32];
    __shared__ unsigned long long s_base;
    const uint32_t mask = (uint32_t)S - 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int f = p.piece0 + blockIdx.x; f < p.piece0 + p.fine; f += gridDim.x) {
        int64_t lo, cnt;
        if (p.piece_cap > 0) {
            lo = (int64_t)f * p.piece_cap;
            cnt = p.fine_counts[f] < p.piece_cap ? p.fine_counts[f] : p.piece_cap;
        } else {
            lo = p.fine_base[f];
            cnt = p.fine_counts[f];
        }
        if (cnt == 0) continue;  // the same for every thread of the CTA
        for (int i = threadIdx.x; i < S; i += THREADS) {
            skey[i] = MG_EMPTY;
            ssum[i] = 0.0;
            srows[i] = 0u;
            if (MINMAX) { smin[i] = ~0ull; smax[i] = 0ull; }
        }
        __syncthreads();
        const longlong2* rec = p.rec + lo;
        int64_t i = threadIdx.x;
        for (; i + (MG_UNROLL - 1) * THREADS < cnt; i += MG_UNROLL * THREADS) {
            longlong2 r[MG_UNROLL];
#pragma unroll
            for (int j = 0; j < MG_UNROLL; ++j) r[j] = ld_record(rec + i + j * THREADS);
#pragma unroll
            for (int j = 0; j < MG_UNROLL; ++j) merge_one<MINMAX>(p, skey, ssum, smin, smax, srows, mask, r[j]);
        }
        for (; i < cnt; i += THREADS) merge_one<MINMAX>(p, skey, ssum, smin, smax, srows, mask, ld_record(rec + i));
        __syncthreads();
        // append the used slots: count per warp, one global atomic per CTA, then warp-ordered writes
        unsigned mine = 0;
        for (int s = threadIdx.x; s < S; s += THREADS) mine += skey[s] != MG_EMPTY;
        mine = __reduce_add_sync(0xffffffffu, mine);
        if (lane == 0) s_warp[warp] = mine;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned run = 0;
#pragma unroll
            for (int w = 0; w < THREADS / 32; ++w) { const unsigned c = s_warp[w]; s_warp[w] = run; run += c; }
            s_base = run ? atomicAdd(p.groups, (unsigned long long)run) : 0ull;
        }
        __syncthreads();
        int64_t run = (int64_t)s_base + s_warp[warp];
        for (int s = threadIdx.x; s < S; s += THREADS) {
            const uint64_t key = skey[s];
            const bool used = key != MG_EMPTY;
            const unsigned bal = __ballot_sync(0xffffffffu, used);
            const int64_t g = run + __popc(bal & lanemask_lt());
            if (used && g < p.out_capacity) {
                const int64_t rows = (int64_t)srows[s];
                p.out_keys[g] = (int64_t)key;
                p.out_rows[g] = rows;
                p.out_count[g] = rows;   // records carry no nulls: count(value) = rows
                p.out_sum[g] = ssum[s];
                if (MINMAX) {
                    p.out_min[g] = dbl_from_sortable(smin[s]);
                    p.out_max[g] = dbl_from_sortable(smax[s]);
                }
            }
            run += __popc(bal);
        }
        __syncthreads();  // the table is cleared again at the top
    }
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_hash_subpartition(etl_handle* hh, const int64_t* records, int32_t buckets, int32_t sub,
                                     const int64_t* d_bucket_counts, const int64_t* d_bucket_base,
                                     int64_t max_bucket_records, int64_t* d_fine_counts, int64_t* d_fine_base,
                                     int64_t* out_records, void* stream_) {
    ETL_REQUIRE(hh && d_bucket_counts && d_bucket_base && d_fine_counts && d_fine_base, "hash_subpartition: NULL argument");
    ETL_REQUIRE(buckets >= 1 && buckets <= 65535 && sub >= 1 && sub <= etl::SUB_MAX && (int64_t)buckets * sub <= 65536,
                "hash_subpartition: buckets=%d x sub=%d must be <= 65536 with sub <= %d", buckets, sub, etl::SUB_MAX);
    ETL_REQUIRE(max_bucket_records >= 0, "hash_subpartition: max_bucket_records < 0");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int fine = buckets * sub;
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_fine_counts, 0, sizeof(int64_t) * fine, stream));
    if (max_bucket_records == 0) {
        ETL_CUDA(cudaMemsetAsync(d_fine_base, 0, sizeof(int64_t) * fine, stream));
        return ETL_OK;
    }
    ETL_REQUIRE(records && out_records && etl::aligned16(records) && etl::aligned16(out_records),
                "hash_subpartition: NULL or misaligned record buffer");
    int rc = etl::arena_reserve(h, sizeof(unsigned long long) * 65536);
    if (rc != ETL_OK) return rc;
    unsigned long long* cursor = static_cast<unsigned long long*>(h->scratch);
    ETL_CUDA(cudaMemsetAsync(cursor, 0, sizeof(unsigned long long) * fine, stream));
    // grid.y = level-A bucket; grid.x = CTAs striding over that bucket's tiles (uniform hash: equal buckets)
    const int64_t tiles = (max_bucket_records + etl::SUB_TILE - 1) / etl::SUB_TILE;
    int64_t gx = ((int64_t)h->sm_count * 6 + buckets - 1) / buckets;
    if (gx > tiles) gx = tiles;
    if (gx < 1) gx = 1;
    const dim3 grid((unsigned)gx, (unsigned)buckets);
    const longlong2* rec = reinterpret_cast<const longlong2*>(records);
    etl::subpartition_count_kernel<<<grid, etl::SUB_THREADS, 0, stream>>>(
        rec, d_bucket_base, d_bucket_counts, sub, (unsigned)fine, reinterpret_cast<unsigned long long*>(d_fine_counts));
    ETL_CUDA(cudaGetLastError());
    etl::fine_scan_kernel<<<1, 1024, 0, stream>>>(d_fine_counts, d_fine_base, fine);
    ETL_CUDA(cudaGetLastError());
    etl::ScatterJob job{};
    job.keys = records; job.seg_base = d_bucket_base; job.seg_counts = d_bucket_counts; job.sub = sub;
    job.fine = (unsigned)fine; job.parts = sub; job.dest_base = d_fine_base; job.cursor = cursor;
    etl::PartDest dst;
    for (int d = 0; d < etl::PART_MAX; ++d) dst.base[d] = out_records;
    ETL_CUDA(etl::launch_scatter_pipelined<true>(h->sm_count, job, dst, buckets, max_bucket_records, stream));
    return ETL_OK;
}

static int merge_pieces(etl_handle* hh, const int64_t* records, int32_t piece0, int32_t fine, bool append, bool shared_sms,
                        const int64_t* d_fine_counts, const int64_t* d_fine_base, int64_t piece_cap, int32_t smem_slots,
                        int64_t out_capacity,
                        int64_t* out_keys, int64_t* out_rows, int64_t* out_count, double* out_sum, double* out_min,
                        double* out_max, int64_t* spill_records, int64_t spill_capacity, int64_t* d_counters, void* stream_) {
    ETL_REQUIRE(hh && d_fine_counts && (d_fine_base || piece_cap > 0) && d_counters, "groupby_merge_partitioned: NULL argument");
    ETL_REQUIRE(fine >= 1 && piece0 >= 0 && piece0 + fine <= 65536, "groupby_merge_partitioned: pieces %d..%d not in 0..65536",
                piece0, piece0 + fine);
    ETL_REQUIRE(smem_slots >= 64 && smem_slots <= 8192 && (smem_slots & (smem_slots - 1)) == 0,
                "groupby_merge_partitioned: smem_slots=%d must be a power of two in 64..8192", smem_slots);
    ETL_REQUIRE(out_capacity >= 0 && spill_capacity >= 0, "groupby_merge_partitioned: negative capacity");
    ETL_REQUIRE(out_capacity == 0 || (out_keys && out_rows && out_count && out_sum),
                "groupby_merge_partitioned: NULL output array");
    ETL_REQUIRE((out_min == nullptr) == (out_max == nullptr), "groupby_merge_partitioned: out_min and out_max go together");
    const bool minmax = out_min != nullptr;
    ETL_REQUIRE(!minmax || smem_slots <= 4096, "groupby_merge_partitioned: with min / max at most 4096 slots fit");
    ETL_REQUIRE(spill_capacity == 0 || (spill_records && etl::aligned16(spill_records)),
                "groupby_merge_partitioned: NULL or misaligned spill buffer");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    if (!append) ETL_CUDA(cudaMemsetAsync(d_counters, 0, 4 * sizeof(int64_t), stream));
    ETL_REQUIRE(records == nullptr || etl::aligned16(records), "groupby_merge_partitioned: misaligned records");
    if (!records) return ETL_OK;
    etl::MergeParams p{};
    p.rec = reinterpret_cast<const longlong2*>(records);
    p.fine_base = d_fine_base; p.fine_counts = d_fine_counts; p.fine = fine; p.piece0 = piece0; p.slots = smem_slots; p.piece_cap = piece_cap;
    p.out_keys = out_keys; p.out_rows = out_rows; p.out_count = out_count; p.out_sum = out_sum;
    p.out_min = out_min; p.out_max = out_max;
    p.out_capacity = out_capacity;
    unsigned long long* ctr = reinterpret_cast<unsigned long long*>(d_counters);
    p.groups = ctr; p.spilled = ctr + 1; p.status = ctr + 2;   // status[1] = ctr[3]: reserved-key records
    p.spill = reinterpret_cast<longlong2*>(spill_records); p.spill_capacity = spill_capacity;
    const size_t smem = (size_t)smem_slots * (minmax ? 36 : 20);
    static const int env_threads = [] { const char* e = getenv("ETL_MERGE_THREADS"); return e && *e ? atoi(e) : 0; }();
    const int threads = env_threads == etl::MG_THREADS || env_threads == etl::MG_THREADS_SHARED
                            ? env_threads : (shared_sms ? etl::MG_THREADS_SHARED : etl::MG_THREADS);
    const bool wide = threads == etl::MG_THREADS;
    auto kern = minmax ? (wide ? etl::groupby_merge_kernel<true, etl::MG_THREADS> : etl::groupby_merge_kernel<true, etl::MG_THREADS_SHARED>)
                       : (wide ? etl::groupby_merge_kernel<false, etl::MG_THREADS> : etl::groupby_merge_kernel<false, etl::MG_THREADS_SHARED>);
    ETL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (const int carve = etl::smem_carveout_percent(); carve >= 0)
        ETL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    int occ = 0;
    ETL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem));
    if (occ < 1) occ = 1;
    int64_t blocks = (int64_t)h->sm_count * occ;
    if (blocks > fine) blocks = fine;
    kern<<<(unsigned)blocks, threads, smem, stream>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_groupby_merge_partitioned(etl_handle* hh, const int64_t* records, int32_t fine,
                                             const int64_t* d_fine_counts, const int64_t* d_fine_base,
                                             int32_t smem_slots, int64_t out_capacity, int64_t* out_keys,
                                             int64_t* out_rows, int64_t* out_count, double* out_sum, double* out_min,
                                             double* out_max, int64_t* spill_records, int64_t spill_capacity,
                                             int64_t* d_counters, void* stream_) {
    ETL_REQUIRE(d_fine_base != nullptr, "groupby_merge_partitioned: NULL argument");
    return merge_pieces(hh, records, 0, fine, false, false, d_fine_counts, d_fine_base, 0, smem_slots, out_capacity, out_keys, out_rows,
                        out_count, out_sum, out_min, out_max, spill_records, spill_capacity, d_counters, stream_);
}

extern "C" int etl_groupby_merge_slabs(etl_handle* hh, const int64_t* records, int32_t piece0, int32_t pieces, int32_t append,
                                       const int64_t* d_cursor, int64_t piece_capacity, int32_t smem_slots, int64_t out_capacity,
                                       int64_t* out_keys, int64_t* out_rows, int64_t* out_count, double* out_sum,
                                       double* out_min, double* out_max, int64_t* spill_records, int64_t spill_capacity,
                                       int64_t* d_counters, void* stream_) {
    ETL_REQUIRE(piece_capacity > 0, "groupby_merge_slabs: piece_capacity must be > 0");
    return merge_pieces(hh, records, piece0, pieces, (append & 1) != 0, (append & 2) != 0, d_cursor, nullptr, piece_capacity, smem_slots, out_capacity, out_keys, out_rows,
                        out_count, out_sum, out_min, out_max, spill_records, spill_capacity, d_counters, stream_);
}
