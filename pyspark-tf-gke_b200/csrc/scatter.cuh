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
// Shared memory: 2 x 32 KB input + 4 KB permutation + 4 KB counters = 73 KB -> 3 CTAs per SM.
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
    const int64_t* dest_base;   // [segments * parts] first record of every destination
    unsigned long long* cursor; // [segments * parts] records already claimed, zero at launch
};

constexpr size_t SCATTER_SMEM = 2 * PART_TILE * 16 + PART_TILE * 2 + PART_MAX * (4 + 4 + 8) + 16;

template <bool RECORDS, int THREADS>
__global__ void __launch_bounds__(THREADS, 3) scatter_pipelined_kernel(const ScatterJob job, const PartDest dst) {
    extern __shared__ __align__(128) unsigned char sc_smem[];
    constexpr int PER = PART_TILE / THREADS;
    uint16_t* s_src = reinterpret_cast<uint16_t*>(sc_smem + 2 * PART_TILE * 16);
    unsigned* s_cnt = reinterpret_cast<unsigned*>(s_src + PART_TILE);
    unsigned* s_off = s_cnt + PART_MAX;
    long long* s_delta = reinterpret_cast<long long*>(s_off + PART_MAX);
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_delta + PART_MAX);
    const int seg = blockIdx.y;
    const int64_t lo = job.seg_base ? job.seg_base[seg] : 0;
    const int64_t cnt = job.seg_counts ? job.seg_counts[seg] : job.n;
    const int64_t tiles = (cnt + PART_TILE - 1) / PART_TILE;
    const size_t d0 = (size_t)seg * job.parts;
    const unsigned sub0 = (unsigned)seg * (unsigned)job.sub;
    const int parts = job.parts;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_init(bar + 1, 1);
        mbar_init_fence();
    }
    __syncthreads();

    // thread 0: start the copy of tile t into buffer `buf`
    auto issue = [&](int64_t t, int buf) {
        const int64_t left = cnt - t * PART_TILE;
        const unsigned rows = left < PART_TILE ? (unsigned)left : (unsigned)PART_TILE;
        unsigned char* in = sc_smem + (size_t)buf * PART_TILE * 16;
        if (RECORDS) {
            mbar_arrive_expect_tx(bar + buf, rows * 16);
            bulk_load(in, job.keys + 2 * (lo + t * PART_TILE), rows * 16, bar + buf);
        } else {
            const int64_t* gk = job.keys + lo + t * PART_TILE;
            const double* gv = job.vals + lo + t * PART_TILE;
            const unsigned even = rows & ~1u;  // bulk copies move multiples of 16 bytes
            if (rows & 1u) {                   // an odd last row: by hand, published by the arrive below
                reinterpret_cast<int64_t*>(in)[even] = gk[even];
                reinterpret_cast<double*>(in + PART_TILE * 8)[even] = gv[even];
            }
            mbar_arrive_expect_tx(bar + buf, even * 16);
            if (even) {
                bulk_load(in, gk, even * 8, bar + buf);
                bulk_load(in + PART_TILE * 8, gv, even * 8, bar + buf);
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
        const int64_t left = cnt - t * PART_TILE;
        const int rows = left < PART_TILE ? (int)left : PART_TILE;
        const unsigned char* in = sc_smem + (size_t)buf * PART_TILE * 16;
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
        if (threadIdx.x < parts && s_cnt[threadIdx.x])
            gb = atomicAdd(job.cursor + d0 + threadIdx.x, (unsigned long long)s_cnt[threadIdx.x]);
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
        if (threadIdx.x < parts)
            s_delta[threadIdx.x] = (long long)job.dest_base[d0 + threadIdx.x] + (long long)gb - (long long)s_off[threadIdx.x];
        __syncthreads();
        // 4. write in sorted order
        for (int i = threadIdx.x; i < rows; i += THREADS) {
            const int src = s_src[i];
            longlong2 rec;
            if (RECORDS) {
                rec = *reinterpret_cast<const longlong2*>(in + (size_t)src * 16);
            } else {
                rec.x = kin[src];
                rec.y = kin[PART_TILE + src];
            }
            const unsigned d = dest_of((uint64_t)rec.x);
            *reinterpret_cast<longlong2*>(dst.base[d] + 2 * ((int64_t)s_delta[d] + i)) = rec;
        }
    }
}

// host side: launch over `segments` segments (grid.y) whose largest holds at most max_rows rows
template <bool RECORDS>
static inline cudaError_t launch_scatter_pipelined(int sm_count, const ScatterJob& job, const PartDest& dst, int segments,
                                                   int64_t max_rows, cudaStream_t stream) {
    // measured (1 B rows, 256 destinations): records 11.8 ms at 256 threads, 11.3 at 512; columns 8.9 / 9.1
    static const int threads = [] { const char* e = getenv("ETL_PART_THREADS"); return e ? atoi(e) : (RECORDS ? 512 : 256); }();
    const int64_t tiles = (max_rows + PART_TILE - 1) / PART_TILE;
    int64_t gx = ((int64_t)sm_count * 3 + segments - 1) / segments;
    if (gx > tiles) gx = tiles;
    if (gx < 1) gx = 1;
    const dim3 grid((unsigned)gx, (unsigned)segments);
    cudaError_t e;
    if (threads == 512) {
        e = cudaFuncSetAttribute(scatter_pipelined_kernel<RECORDS, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCATTER_SMEM);
        if (e != cudaSuccess) return e;
        scatter_pipelined_kernel<RECORDS, 512><<<grid, 512, SCATTER_SMEM, stream>>>(job, dst);
    } else {
        e = cudaFuncSetAttribute(scatter_pipelined_kernel<RECORDS, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCATTER_SMEM);
        if (e != cudaSuccess) return e;
        scatter_pipelined_kernel<RECORDS, 256><<<grid, 256, SCATTER_SMEM, stream>>>(job, dst);
    }
    return cudaGetLastError();
}

}  // namespace etl
