// Hash-partition scatter, pipelined: the pass that moves every {key, value} row to its
// destination's contiguous range.  Used by both partition levels and by the fused exchange.
//
// One CTA works on tiles of PART_TILE rows of one "segment" (level 1: the whole input; level 2:
// one level-1 bucket per blockIdx.y).  The tile's input arrives in shared memory by 1-D bulk
// async copies (TMA engine, completion on an mbarrier), double buffered: the copy of the CTA's
// next tile is in flight while the current one is ranked, permuted and written, so the HBM read
// latency that stalled the register-staged version (ncu: 56 % of stall samples long_scoreboard,
// DRAM at 45 %) overlaps the shared-memory work.  Three block barriers per tile.  Per tile:
//   1. destination d of every row; rank of the row within (tile, d) by a shared atomic;
//   2. exclusive scan of the tile's per-destination counts; one global atomic per destination
//      reserves the tile's run in that destination's range;
//   3. the permutation "sorted position -> source row" (2 bytes per row) - the rows themselves
//      stay where the bulk copy put them;
//   4. rows are stored to global memory in sorted order: consecutive threads write consecutive
//      16-byte records of one destination's run.
// Two shapes: 4096-row tiles x 1024 threads (2 x 64 KB input + 8 KB permutation + 5 KB counters = 141 KB,
// one CTA per SM; the default - a (tile, destination) run is 256 B at 256 destinations) and 2048-row
// tiles x 256 / 512 threads (74 KB, three CTAs per SM).
//
// SLAB mode (template flag): no count pass in front.  Every destination owns a fixed-capacity slab
// (destination d's records start at slab_origin + d * slab_stride of its buffer, room for dest_cap
// records) and a cursor that persists across launches, so a table can be scattered chunk by chunk.
// A run that does not fit its slab is cut at the slab's end and the records beyond it are counted in
// job.overflow and dropped - the host then repeats the work on the counted path, which sizes every
// destination exactly.  Segments of a slab-mode input are slabs themselves: segment s holds the
// records [min(seg_prev[s], seg_cap), min(seg_counts[s], seg_cap)) of slab s.
#pragma once
#include "common.cuh"

namespace etl {

constexpr int PART_TILE = 2048;
constexpr int PART_THREADS = 256;
constexpr int PART_MAX = 256;

// destination = rank * buckets + bucket: rank from the high hash word, bucket from the low one
// (Lemire range reductions; the bucket matches the slot ranges of gb_global_slot)
__device__ __forceinline__ int part_dest(uint64_t key, int ranks, int buckets) {
    const uint64_t h = mix64(key);
    return (int)(__umulhi((unsigned)(h >> 32), (unsigned)ranks) * (unsigned)buckets + __umulhi((unsigned)h, (unsigned)buckets));
}
// level 2: index among `fine` = buckets * sub pieces; fine_index / sub == the level-1 bucket
__device__ __forceinline__ unsigned fine_index(uint64_t key, unsigned fine) { return __umulhi((unsigned)mix64(key), fine); }

// Where destination d's records go: base[d] + 2 * (offset[d] + running position).  For a local
// partition every base is the one output buffer and offset[] its exclusive scan; for the fused
// exchange base[d] is rank d's receive buffer mapped into this GPU's address space (NVLink peer
// memory) and offset[d] the first record this rank owns there.
struct PartDest {
    int64_t* base[PART_MAX];
};

struct ScatterJob {
    const int64_t* keys;        // RECORDS: {key, value bits} pairs; else the key column
    const double* vals;         // value column (unused for RECORDS)
    const int64_t* seg_base;    // per blockIdx.y: first row / row count of the segment (nullptr: one segment [0, n))
    const int64_t* seg_counts;
    int64_t n;
    int ranks, buckets;         // sub == 0: destination = part_dest(key, ranks, buckets)
    int sub;                    // sub > 0: destination = fine_index(key, fine) - segment * sub
    unsigned fine;
    int parts;                  // destinations per segment (<= PART_MAX)
    const int64_t* dest_base;   // [segments * parts] first record of every destination (counted mode)
    unsigned long long* cursor; // [segments * parts] records already claimed (counted mode: zero at launch)
    // ---- slab mode
    int64_t seg_cap;            // > 0: segment s = slab s of the input (seg_cap records apart), counts clamped to seg_cap
    const int64_t* seg_prev;    // optional [segments]: records of the segment consumed by earlier launches
    int shared_dest;            // 1: all segments scatter into the same `parts` destinations (cursor / slabs not per segment)
    // second level over what a bucketed exchange delivered: the input slabs are laid out [source][bucket]
    // (seg_bucket_stride buckets per source); grid.y walks (bucket, source) pairs: segment y is slab
    // (y % seg_sources) * seg_bucket_stride + seg_bucket0 + y / seg_sources and belongs to bucket
    // seg_bucket0 + y / seg_sources, whose `parts` pieces all its sources share.  seg_sources == 0: off.
    int seg_sources, seg_bucket0, seg_bucket_stride;
    int64_t slab_origin, slab_stride, dest_cap;
    unsigned long long* overflow;   // [0] += records that found no room in their slab
};

template <int TILE>
constexpr size_t scatter_smem_bytes() { return 2 * (size_t)TILE * 16 + (size_t)TILE * 2 + PART_MAX * (4 + 4 + 8 + 4) + 16; }
constexpr size_t SCATTER_SMEM = scatter_smem_bytes<PART_TILE>();

template <bool RECORDS, int THREADS, bool SLAB = false, int TILE = PART_TILE>
__global__ void __launch_bounds__(THREADS, TILE <= 2048 ? 3 : 1) scatter_pipelined_kernel(const ScatterJob job, const PartDest dst) {
    extern __shared__ __align__(128) unsigned char sc_smem[];
    constexpr int PER = TILE / THREADS;
    uint16_t* s_src = reinterpret_cast<uint16_t*>(sc_smem + 2 * TILE * 16);
    unsigned* s_cnt = reinterpret_cast<unsigned*>(s_src + TILE);
    unsigned* s_off = s_cnt + PART_MAX;
    unsigned* s_lim = s_off + PART_MAX;     // SLAB: sorted positions >= s_lim[d] of destination d found no room
    long long* s_delta = reinterpret_cast<long long*>(s_lim + PART_MAX);
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_delta + PART_MAX);
    int seg = blockIdx.y;       // input segment (slab) ...
    int group = seg;            // ... and the destination group it scatters into
    if (SLAB && job.seg_sources > 0) {
        group = job.seg_bucket0 + seg / job.seg_sources;
        seg = (seg % job.seg_sources) * job.seg_bucket_stride + group;
    }
    int64_t lo, cnt;
    if (SLAB && job.seg_cap > 0) {
        int64_t prev = job.seg_prev ? job.seg_prev[seg] : 0, cur = job.seg_counts[seg];
        prev = prev < job.seg_cap ? prev : job.seg_cap;
        cur = cur < job.seg_cap ? cur : job.seg_cap;
        lo = (int64_t)seg * job.seg_cap + prev;
        cnt = cur - prev;
    } else {
        lo = job.seg_base ? job.seg_base[seg] : 0;
        cnt = job.seg_counts ? job.seg_counts[seg] : job.n;
    }
    const int64_t tiles = (cnt + TILE - 1) / TILE;
    const size_t d0 = (SLAB && job.shared_dest) ? 0 : (size_t)group * job.parts;
    const unsigned sub0 = (unsigned)group * (unsigned)job.sub;
    const int parts = job.parts;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_init(bar + 1, 1);
        mbar_init_fence();
    }
    __syncthreads();

    // thread 0: start the copy of tile t into buffer `buf`
    auto issue = [&](int64_t t, int buf) {
        const int64_t left = cnt - t * TILE;
        const unsigned rows = left < TILE ? (unsigned)left : (unsigned)TILE;
        unsigned char* in = sc_smem + (size_t)buf * TILE * 16;
        if (RECORDS) {
            mbar_arrive_expect_tx(bar + buf, rows * 16);
            bulk_load(in, job.keys + 2 * (lo + t * TILE), rows * 16, bar + buf);
        } else {
            const int64_t* gk = job.keys + lo + t * TILE;
            const double* gv = job.vals + lo + t * TILE;
            const unsigned even = rows & ~1u;  // bulk copies move multiples of 16 bytes
            if (rows & 1u) {                   // an odd last row: by hand, published by the arrive below
                reinterpret_cast<int64_t*>(in)[even] = gk[even];
                reinterpret_cast<double*>(in + TILE * 8)[even] = gv[even];
            }
            mbar_arrive_expect_tx(bar + buf, even * 16);
            if (even) {
                bulk_load(in, gk, even * 8, bar + buf);
                bulk_load(in + TILE * 8, gv, even * 8, bar + buf);
            }
        }
    };
    auto dest_of = [&](uint64_t key) -> unsigned {
        if (job.sub) return (fine_index(key, job.fine) - sub0) & (PART_MAX - 1);
        return (unsigned)part_dest(key, job.ranks, job.buckets) & (PART_MAX - 1);
    };

    // Three block barriers per tile.  The counters are cleared for the NEXT tile right after the scan has
    // read them, and the copy of the next tile is issued after the first barrier of the current one - by
    // then every thread has left the previous tile's write loop, the last reader of that buffer.
    int64_t t = blockIdx.x;
    if (threadIdx.x == 0 && t < tiles) issue(t, 0);
    if (threadIdx.x < PART_MAX) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    for (int it = 0; t < tiles; t += gridDim.x, ++it) {
        const int buf = it & 1;
        mbar_wait(bar + buf, (unsigned)(it >> 1) & 1u);
        const int64_t left = cnt - t * TILE;
        const int rows = left < TILE ? (int)left : TILE;
        const unsigned char* in = sc_smem + (size_t)buf * TILE * 16;
        const int64_t* kin = reinterpret_cast<const int64_t*>(in);
        // 1. destination and rank within (tile, destination), packed: destination << 16 | rank
        unsigned dr[PER];
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int r = j * THREADS + threadIdx.x;
            dr[j] = 0xffffffffu;
            if (r < rows) {
                const unsigned d = dest_of((uint64_t)(RECORDS ? kin[2 * r] : kin[r]));
                dr[j] = (d << 16) | atomicAdd(&s_cnt[d], 1u);
            }
        }
        __syncthreads();   // counts complete; nobody is in the previous tile's write loop any more
        if (threadIdx.x == 0 && t + gridDim.x < tiles) issue(t + gridDim.x, buf ^ 1);
        // 2. reserve this tile's run in every destination (atomics in flight while the offsets are scanned)
        unsigned long long gb = 0;
        const unsigned mycnt = (int)threadIdx.x < parts ? s_cnt[threadIdx.x] : 0u;
        if (mycnt) gb = atomicAdd(job.cursor + d0 + threadIdx.x, (unsigned long long)mycnt);
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
        if (threadIdx.x < PART_MAX) s_cnt[threadIdx.x] = 0;   // for the next tile (read for the last time above)
        // 3. sorted position -> source row (the reservations' round trip to L2 overlaps this)
#pragma unroll
        for (int j = 0; j < PER; ++j)
            if (dr[j] != 0xffffffffu) s_src[s_off[dr[j] >> 16] + (dr[j] & 0xffffu)] = (uint16_t)(j * THREADS + threadIdx.x);
        if ((int)threadIdx.x < parts) {
            if (SLAB) {
                long long fit = (long long)job.dest_cap - (long long)gb;      // room left in the slab before this run
                fit = fit < 0 ? 0 : (fit > (long long)mycnt ? (long long)mycnt : fit);
                if (fit < (long long)mycnt) atomicAdd(job.overflow, (unsigned long long)((long long)mycnt - fit));
                s_lim[threadIdx.x] = s_off[threadIdx.x] + (unsigned)fit;
                s_delta[threadIdx.x] = job.slab_origin + (long long)(d0 + threadIdx.x) * job.slab_stride + (long long)gb -
                                       (long long)s_off[threadIdx.x];
            } else {
                s_delta[threadIdx.x] = (long long)job.dest_base[d0 + threadIdx.x] + (long long)gb - (long long)s_off[threadIdx.x];
            }
        }
        __syncthreads();
        // 4. write in sorted order
        for (int i = threadIdx.x; i < rows; i += THREADS) {
            const int src = s_src[i];
            longlong2 rec;
            if (RECORDS) {
                rec = *reinterpret_cast<const longlong2*>(in + (size_t)src * 16);
            } else {
                rec.x = kin[src];
                rec.y = kin[TILE + src];
            }
            const unsigned d = dest_of((uint64_t)rec.x);
            if (!SLAB || (unsigned)i < s_lim[d])
                *reinterpret_cast<longlong2*>(dst.base[d] + 2 * ((int64_t)s_delta[d] + i)) = rec;
        }
    }
}

// host side: launch over `segments` segments (grid.y) whose largest holds at most max_rows rows.
// max_ctas > 0 caps the grid (a link-bound exchange scatter leaves SMs to the kernels it overlaps with).
template <bool RECORDS, int THREADS, bool SLAB, int TILE>
static inline cudaError_t launch_scatter_variant(int sm_count, const ScatterJob& job, const PartDest& dst, int segments,
                                                 int64_t max_rows, int max_ctas, cudaStream_t stream) {
    constexpr size_t smem = scatter_smem_bytes<TILE>();
    constexpr int per_sm = TILE <= 2048 ? 3 : 1;
    const int64_t tiles = (max_rows + TILE - 1) / TILE;
    int64_t total = (int64_t)sm_count * per_sm;
    if (max_ctas > 0 && max_ctas < total) total = max_ctas;
    // CTAs per segment: with several segments (the second level: one per bucket) the grid is segments x gx CTAs
    // of equal work, so gx is chosen to fill whole waves of the `total` CTAs that are resident at a time - 256
    // segments x 1 CTA on 148 SMs would run as one full wave and one 73 % empty one
    int64_t gx = (total + segments - 1) / segments;
    if (segments > 1) {
        double best = 0.0;
        const int64_t g0 = gx;
        for (int64_t g = g0; g <= g0 + 80 && g <= tiles; ++g) {
            const int64_t ctas = g * segments, waves = (ctas + total - 1) / total;
            const double eff = (double)ctas / (double)(waves * total);
            if (eff > best + 1e-9) { best = eff; gx = g; }
            if (eff > 0.995) break;
        }
    }
    if (gx > tiles) gx = tiles;
    if (gx < 1) gx = 1;
    const dim3 grid((unsigned)gx, (unsigned)segments);
    auto kern = scatter_pipelined_kernel<RECORDS, THREADS, SLAB, TILE>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (const int carve = smem_carveout_percent(); carve >= 0) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        if (e != cudaSuccess) return e;
    }
    kern<<<grid, THREADS, smem, stream>>>(job, dst);
    return cudaGetLastError();
}

template <bool RECORDS, bool SLAB = false>
static inline cudaError_t launch_scatter_pipelined(int sm_count, const ScatterJob& job, const PartDest& dst, int segments,
                                                   int64_t max_rows, cudaStream_t stream, int max_ctas = 0) {
    // Measured on config 4 (1 B rows, 256 destinations, B200): tile 2048 x 256 / 512 threads 7.45 / 8.33 ms for
    // level 1 / level 2; tile 4096 x 512 threads 6.77 / 7.45; tile 4096 x 1024 threads 6.13 / 6.76 (twice the run
    // length per (tile, destination); one 141 KB CTA per SM).  A launch with a CTA cap (max_ctas > 0) is one that
    // shares the SMs with another scatter (the exchange overlapped with the receiver's first level): it takes the
    // 2048-row tile, whose 74 KB fit beside a 4096-row CTA.
    static const int env_threads = [] { const char* e = getenv("ETL_PART_THREADS"); return e ? atoi(e) : 0; }();
    static const int env_tile = [] { const char* e = getenv("ETL_PART_TILE"); return e ? atoi(e) : 0; }();
    const int tile = env_tile ? env_tile : (max_ctas > 0 ? 2048 : 4096);
    const int threads = env_threads ? env_threads : (tile == 4096 ? 1024 : (RECORDS ? 512 : 256));
    if (tile == 4096) {
        if (threads == 1024) return launch_scatter_variant<RECORDS, 1024, SLAB, 4096>(sm_count, job, dst, segments, max_rows, max_ctas, stream);
        return launch_scatter_variant<RECORDS, 512, SLAB, 4096>(sm_count, job, dst, segments, max_rows, max_ctas, stream);
    }
    if (threads == 512) return launch_scatter_variant<RECORDS, 512, SLAB, 2048>(sm_count, job, dst, segments, max_rows, max_ctas, stream);
    return launch_scatter_variant<RECORDS, 256, SLAB, 2048>(sm_count, job, dst, segments, max_rows, max_ctas, stream);
}

}  // namespace etl
