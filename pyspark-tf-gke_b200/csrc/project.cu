// Fused map-side projection: isNotNull filter + mean-impute + cast-to-double +
// order-preserving compaction, one pass over the rows (k_means.py:27-28,49-51).
//
// Persistent, warp-specialised kernel.  Rows are cut into 1024-row tiles and
// tiles into chunks (1 tile each unless the table is huge); CTAs claim chunks
// dynamically (first one static, then an atomic ticket) so no SM idles at the
// tail.  Three roles meet only on mbarriers, over a ring of shared-memory
// stages - no __syncthreads in the steady state:
//   loader warp   claims chunks; one lane issues each tile's column slices as
//                 1-D bulk async copies (TMA engine, cp.async.bulk) onto the
//                 stage's `full` mbarrier, PROJ_STAGES tiles ahead - the bytes
//                 in flight live in shared memory, not in registers;
//   scanner warp  the keep predicate depends on the key bitmap alone, so at start
//                 the CTA popcounts the bitmap of its share of the chunks into
//                 one tagged 64-bit word per chunk in scratch, and a chunk's
//                 first output row is just the sum of the lower chunks' words -
//                 no chained look-back; the scanner keeps that sum incrementally
//                 from chunk to chunk.  Per tile, once it has landed: popcount of its 32 key
//                 validity words (a validity bitmap is a stored warp ballot) +
//                 warp scan -> per-32-row segment offsets and the tile's base;
//   8 consumer warps  rank = base + segment offset + popc(word & lanemask_lt),
//                 impute / cast, and store: kept lanes of a warp write
//                 consecutive rows, so stores are coalesced without staging.
// HBM traffic = the algorithmic bytes: every input byte is read once, every
// output byte written once (plus 4 B of bitmap per 32 rows read twice).
#include "common.cuh"
#include <stdlib.h>
#include <vector>

namespace etl {

constexpr int PROJ_TILE = 1024;
constexpr int PROJ_CONSUMER_WARPS = 8;
constexpr int PROJ_THREADS = (PROJ_CONSUMER_WARPS + 2) * 32;  // + loader + scanner
constexpr int PROJ_MAX_STAGES = 6;
constexpr int64_t PROJ_MAX_CHUNKS = 1 << 16;

struct ProjParams {
    const int32_t* key_code;
    const uint32_t* key_valid;
    const void* val[3];
    const uint32_t* val_valid[3];
    int32_t val_dtype[3];
    double fill[3];
    int64_t n;
    int32_t* out_code;
    double* out_val[3];
    int64_t* out_rows;
    uint64_t* chunk_state;   // one tagged word per chunk: kept rows of the chunk
    unsigned int* ticket;    // [0] next chunk to claim, [1] CTAs finished; both 0 between launches
    uint32_t epoch;
    int stages;
    int chunk_tiles;
    int64_t chunks;
    int hints;  // bit 0: L2 evict-first on the bulk loads, bit 1: streaming (.cs) stores
};

struct __align__(128) ProjStage {
    double val[3][PROJ_TILE];  // raw column slices; 4-byte dtypes use the first half of a slot
    int32_t code[PROJ_TILE];
    uint32_t valid[4][32];     // key, val0, val1, val2 bitmap words of the tile
};

__device__ __forceinline__ uint32_t seg_mask(int seg, int rows) {
    const int left = rows - seg * 32;
    return left >= 32 ? 0xffffffffu : (left <= 0 ? 0u : ((1u << left) - 1));
}

__device__ __forceinline__ double stage_value(const ProjStage& st, int c, int dtype, int r) {
    if (dtype == ETL_F64) return st.val[c][r];
    if (dtype == ETL_F32) return (double)reinterpret_cast<const float*>(st.val[c])[r];
    return (double)reinterpret_cast<const int32_t*>(st.val[c])[r];
}

// tail tile only (rows < 1024, sizes not multiples of 16 bytes): plain copies by the loader warp
__device__ __forceinline__ void warp_copy_words(void* dst, const void* src, int words, int lane) {
    for (int i = lane; i < words; i += 32) static_cast<uint32_t*>(dst)[i] = static_cast<const uint32_t*>(src)[i];
}

__global__ void __launch_bounds__(PROJ_THREADS) project_filter_impute_kernel(const ProjParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ProjStage* stages = reinterpret_cast<ProjStage*>(smem_raw);
    __shared__ uint64_t full_bar[PROJ_MAX_STAGES], scan_bar[PROJ_MAX_STAGES], empty_bar[PROJ_MAX_STAGES];
    __shared__ uint32_t s_seg_off[PROJ_MAX_STAGES][32];
    __shared__ uint64_t s_base[PROJ_MAX_STAGES];
    __shared__ int64_t s_tile[PROJ_MAX_STAGES];  // tile in the stage, -1 = no more work

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = p.stages;
    const int64_t tiles = (p.n + PROJ_TILE - 1) / PROJ_TILE;
    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&scan_bar[s], 1);
            mbar_init(&empty_bar[s], PROJ_CONSUMER_WARPS);
        }
        mbar_init_fence();
    }
    __syncthreads();

    if (warp != PROJ_CONSUMER_WARPS && p.key_valid) {
        // Every warp but the loader: kept rows of this CTA's share of the chunks
        // (blockIdx.x, + gridDim.x, ...; lowest first - they are needed first),
        // published as tagged words.  Nobody waits here; readers spin on the tag.
        const uint64_t tag = (uint64_t)p.epoch << TS_EPOCH_SHIFT;
        const int64_t chunk_rows = (int64_t)p.chunk_tiles * PROJ_TILE;
        const int cw = warp < PROJ_CONSUMER_WARPS ? warp : PROJ_CONSUMER_WARPS;  // 0..8
        for (int64_t c = blockIdx.x + (int64_t)cw * gridDim.x; c < p.chunks;
             c += (int64_t)(PROJ_CONSUMER_WARPS + 1) * gridDim.x) {
            const int64_t r0 = c * chunk_rows;
            const int64_t r1 = r0 + chunk_rows < p.n ? r0 + chunk_rows : p.n;
            const int64_t w0 = r0 >> 5, w1 = (r1 + 31) >> 5;
            uint64_t mine = 0;
            for (int64_t w = w0 + lane; w < w1; w += 32 * 4) {
                uint32_t bits[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) bits[u] = w + 32 * u < w1 ? __ldg(p.key_valid + w + 32 * u) : 0u;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (w + 32 * u == w1 - 1 && (r1 & 31)) bits[u] &= (1u << (r1 & 31)) - 1;
                    mine += __popc(bits[u]);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
            if (lane == 0) st_relaxed_u64(p.chunk_state + c, tag | TS_FLAG_AGG | mine);
        }
    }

    if (warp == PROJ_CONSUMER_WARPS) {
        // ------------------------------------------------------------ loader
        int64_t chunk = blockIdx.x;  // first chunk is static, the rest come from the ticket
        int in_chunk = 0;
        for (int k = 0;; ++k) {
            const int s = k % S;
            if (k >= S) mbar_wait(&empty_bar[s], ((k / S) - 1) & 1);
            if (in_chunk == p.chunk_tiles || chunk * p.chunk_tiles + in_chunk >= tiles) {
                unsigned next = 0;
                if (lane == 0) next = atomicAdd(p.ticket, 1u);
                chunk = (int64_t)gridDim.x + __shfl_sync(0xffffffffu, next, 0);
                in_chunk = 0;
            }
            const int64_t t = chunk * p.chunk_tiles + in_chunk;
            ++in_chunk;
            if (chunk >= p.chunks) {  // out of work: tell the other roles and leave
                if (lane == 0) {
                    s_tile[s] = -1;
                    mbar_arrive(&full_bar[s]);
                }
                break;
            }
            ProjStage& st = stages[s];
            const int64_t row0 = t * PROJ_TILE;
            const int64_t left = p.n - row0;
            if (left >= PROJ_TILE) {
                if (lane == 0) {
                    s_tile[s] = t;
                    const bool hint = p.hints & 1;
                    const uint64_t pol = hint ? l2_policy_evict_first() : 0;
                    unsigned bytes = PROJ_TILE * 4;
#pragma unroll
                    for (int c = 0; c < 3; ++c) bytes += PROJ_TILE * (p.val_dtype[c] == ETL_F64 ? 8 : 4);
                    bytes += p.key_valid ? 128 : 0;
#pragma unroll
                    for (int c = 0; c < 3; ++c) bytes += p.val_valid[c] ? 128 : 0;
                    mbar_arrive_expect_tx(&full_bar[s], bytes);
                    if (p.key_valid) bulk_load(st.valid[0], p.key_valid + (row0 >> 5), 128, &full_bar[s]);
                    if (hint) bulk_load_hint(st.code, p.key_code + row0, PROJ_TILE * 4, &full_bar[s], pol);
                    else bulk_load(st.code, p.key_code + row0, PROJ_TILE * 4, &full_bar[s]);
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const unsigned es = p.val_dtype[c] == ETL_F64 ? 8 : 4;
                        if (hint)
                            bulk_load_hint(st.val[c], static_cast<const char*>(p.val[c]) + row0 * es, PROJ_TILE * es,
                                           &full_bar[s], pol);
                        else
                            bulk_load(st.val[c], static_cast<const char*>(p.val[c]) + row0 * es, PROJ_TILE * es,
                                      &full_bar[s]);
                        if (p.val_valid[c])
                            bulk_load(st.valid[1 + c], p.val_valid[c] + (row0 >> 5), 128, &full_bar[s]);
                    }
                }
            } else {
                const int rows = (int)left, words = (rows + 31) >> 5;
                warp_copy_words(st.code, p.key_code + row0, rows, lane);
                if (p.key_valid) warp_copy_words(st.valid[0], p.key_valid + (row0 >> 5), words, lane);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const int w = p.val_dtype[c] == ETL_F64 ? 2 : 1;
                    warp_copy_words(st.val[c], static_cast<const uint32_t*>(p.val[c]) + row0 * w, rows * w, lane);
                    if (p.val_valid[c]) warp_copy_words(st.valid[1 + c], p.val_valid[c] + (row0 >> 5), words, lane);
                }
                __syncwarp();
                if (lane == 0) {
                    s_tile[s] = t;
                    mbar_arrive(&full_bar[s]);
                }
            }
        }
        // last CTA out re-arms the ticket for the next launch
        if (lane == 0) {
            if (atomicAdd(p.ticket + 1, 1u) == gridDim.x - 1) {
                p.ticket[0] = 0;
                p.ticket[1] = 0;
            }
        }
    } else if (warp == PROJ_CONSUMER_WARPS + 1) {
        // ----------------------------------------------------------- scanner
        const uint64_t tag = (uint64_t)p.epoch << TS_EPOCH_SHIFT;
        const int64_t chunk_rows = (int64_t)p.chunk_tiles * PROJ_TILE;
        uint64_t running = 0;     // first output row of the next tile if it continues the current chunk
        int64_t next_chunk = 0;   // `running` = kept rows of chunks [0, next_chunk) when a chunk starts
        for (int k = 0;; ++k) {
            const int s = k % S;
            mbar_wait(&full_bar[s], (k / S) & 1);
            const int64_t t = s_tile[s];
            if (t < 0) break;
            const int64_t chunk = t / p.chunk_tiles;
            if (t == chunk * p.chunk_tiles) {  // a new chunk starts: add the chunks skipped since the last one
                if (p.key_valid) {
                    uint64_t add = 0;
                    for (int64_t c = next_chunk + lane; c < chunk; c += 32 * 8) {
                        uint64_t sw[8];  // 8 independent loads in flight per lane
#pragma unroll
                        for (int u = 0; u < 8; ++u)
                            sw[u] = c + 32 * u < chunk ? ld_relaxed_u64(p.chunk_state + c + 32 * u) : tag;
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            while ((sw[u] >> TS_EPOCH_SHIFT) != p.epoch) sw[u] = ld_relaxed_u64(p.chunk_state + c + 32 * u);
                            add += sw[u] & TS_VALUE_MASK;
                        }
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) add += __shfl_xor_sync(0xffffffffu, add, o);
                    running += add;
                } else {
                    running = (uint64_t)chunk * chunk_rows;
                }
                next_chunk = chunk;
            }
            const int64_t left = p.n - t * PROJ_TILE;
            const int rows = left >= PROJ_TILE ? PROJ_TILE : (int)left;
            const uint32_t kw = (p.key_valid ? stages[s].valid[0][lane] : 0xffffffffu) & seg_mask(lane, rows);
            const unsigned cnt = __popc(kw);
            unsigned incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += up;
            }
            s_seg_off[s][lane] = incl - cnt;
            if (lane == 0) s_base[s] = running;
            running += __shfl_sync(0xffffffffu, incl, 31);
            if (t == tiles - 1 && lane == 0) *p.out_rows = (int64_t)running;
            // `running` now covers chunks [0, chunk] once the chunk's last tile is done
            if (t + 1 == (chunk + 1) * p.chunk_tiles || t == tiles - 1) next_chunk = chunk + 1;
            __syncwarp();
            if (lane == 0) mbar_arrive(&scan_bar[s]);
        }
    } else {
        // --------------------------------------------------------- consumers
        const unsigned lt = lanemask_lt();
        for (int k = 0;; ++k) {
            const int s = k % S;
            const unsigned ph = (k / S) & 1;
            mbar_wait(&full_bar[s], ph);
            const int64_t t = s_tile[s];
            if (t < 0) break;
            mbar_wait(&scan_bar[s], ph);
            const ProjStage& st = stages[s];
            const int64_t left = p.n - t * PROJ_TILE;
            const int rows = left >= PROJ_TILE ? PROJ_TILE : (int)left;
            const uint64_t base = s_base[s];
            int32_t code[4];
            double x[3][4];
            uint64_t pos[4];
            bool keep[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int g = j * PROJ_CONSUMER_WARPS + warp;  // 32-row segment = bitmap word
                const int r = g * 32 + lane;
                const uint32_t m = seg_mask(g, rows);
                const uint32_t kw = (p.key_valid ? st.valid[0][g] : 0xffffffffu) & m;
                keep[j] = (kw >> lane) & 1u;
                pos[j] = base + s_seg_off[s][g] + __popc(kw & lt);
                code[j] = st.code[r];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t vw = p.val_valid[c] ? st.valid[1 + c][g] : 0xffffffffu;
                    const double v = stage_value(st, c, p.val_dtype[c], r);
                    const bool bad = !((vw >> lane) & 1u) || (v != v);
                    x[c][j] = bad ? p.fill[c] : v;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);  // stage may be refilled while we store
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (keep[j]) {
                    if (p.hints & 2) {
                        st_stream_s32(p.out_code + pos[j], code[j]);
#pragma unroll
                        for (int c = 0; c < 3; ++c) st_stream_f64(p.out_val[c] + pos[j], x[c][j]);
                    } else {
                        p.out_code[pos[j]] = code[j];
#pragma unroll
                        for (int c = 0; c < 3; ++c) p.out_val[c][pos[j]] = x[c][j];
                    }
                }
            }
        }
    }
}

static int check_project_args(const int32_t* key_code, const void* const val[3], const uint32_t* key_valid,
                              const uint32_t* const val_valid[3], const int32_t val_dtype[3], int64_t n,
                              const int32_t* out_code, double* const out_val[3], bool device_bitmaps = true) {
    ETL_REQUIRE(n >= 0 && n < ((int64_t)1 << 39), "project: n=%lld out of range", (long long)n);
    ETL_REQUIRE(val && val_valid && val_dtype && out_val, "project: NULL argument array");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(key_code && out_code, "project: key_code/out_code is NULL");
    // device bitmaps are fetched by 16-byte-granular bulk copies; host bitmaps get re-staged
    ETL_REQUIRE(aligned16(key_code) && (device_bitmaps ? aligned16(key_valid) : aligned4(key_valid)),
                "project: key buffers misaligned");
    for (int c = 0; c < 3; ++c) {
        ETL_REQUIRE(val[c] && out_val[c], "project: val[%d]/out_val[%d] is NULL", c, c);
        ETL_REQUIRE(val_dtype[c] == ETL_F64 || val_dtype[c] == ETL_F32 || val_dtype[c] == ETL_I32,
                    "project: val_dtype[%d]=%d unsupported", c, val_dtype[c]);
        ETL_REQUIRE(aligned16(val[c]) && (device_bitmaps ? aligned16(val_valid[c]) : aligned4(val_valid[c])),
                    "project: val[%d] buffers misaligned", c);
    }
    return ETL_OK;
}

int launch_project(Handle* h, const int32_t* key_code, const uint32_t* key_valid, const void* const val[3],
                   const uint32_t* const val_valid[3], const int32_t val_dtype[3], const double fill[3],
                   int64_t n, int32_t* out_code, double* const out_val[3], int64_t* d_out_rows,
                   cudaStream_t stream) {
    if (n == 0) {
        ETL_CUDA(cudaMemsetAsync(d_out_rows, 0, sizeof(int64_t), stream));
        return ETL_OK;
    }
    const int64_t tiles = (n + PROJ_TILE - 1) / PROJ_TILE;
    ProjParams p;
    p.key_code = key_code;
    p.key_valid = key_valid;
    for (int c = 0; c < 3; ++c) {
        p.val[c] = val[c];
        p.val_valid[c] = val_valid[c];
        p.val_dtype[c] = val_dtype[c];
        p.fill[c] = fill[c];
        p.out_val[c] = out_val[c];
    }
    p.n = n;
    p.out_code = out_code;
    p.out_rows = d_out_rows;
    // persistent grid: every CTA must be resident (a scanner waits for other CTAs' chunk counts).
    // The opt-in shared-memory limit and the occupancy belong to the DEVICE, so they are kept per handle
    // (one handle per GPU); the limit is always raised to the largest ring any path uses, so this path and
    // the allocation-free entry never undercut each other.
    if (h->proj_stages == 0) {
        int stages = 3;
        if (const char* e = getenv("ETL_PROJ_STAGES")) stages = atoi(e);
        if (stages < 2) stages = 2;
        if (stages > PROJ_MAX_STAGES) stages = PROJ_MAX_STAGES;
        const size_t dyn = (size_t)stages * sizeof(ProjStage);
        ETL_CUDA(cudaFuncSetAttribute(project_filter_impute_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(PROJ_MAX_STAGES * sizeof(ProjStage))));
        int occ = 0;
        ETL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, project_filter_impute_kernel, PROJ_THREADS, dyn));
        if (occ < 1) {
            set_error("project: kernel does not fit on an SM with %d stages", stages);
            return ETL_ECUDA;
        }
        if (const char* e = getenv("ETL_PROJ_CTAS_PER_SM")) {
            const int want = atoi(e);
            if (want >= 1 && want < occ) occ = want;
        }
        // default 3: evict-first bulk loads + streaming stores.  Measured on config 2: 108.5 us with
        // neither or only one of them, 98.5 us with both (the inputs stop evicting the dirty output
        // lines the L2 is still gathering into full sectors, and vice versa).
        const char* e = getenv("ETL_PROJ_HINTS");
        h->proj_hints = e ? atoi(e) : 3;
        h->proj_ctas_per_sm = occ;
        h->proj_stages = stages;
    }
    const int s_stages = h->proj_stages, s_ctas_per_sm = h->proj_ctas_per_sm;
    p.stages = s_stages;
    p.hints = h->proj_hints;
    int64_t grid = (int64_t)h->sm_count * s_ctas_per_sm;
    if (grid > tiles) grid = tiles;
    // chunk = unit of dynamic scheduling: ~8+ chunks per CTA, at most 4 tiles (measured best on
    // config 2: fewer claims and state sums per tile, still fine-grained at the tail), more only
    // when the table is so large that the chunk-state array would pass PROJ_MAX_CHUNKS words
    int64_t ct = tiles / (grid * 8);
    ct = ct < 1 ? 1 : (ct > 4 ? 4 : ct);
    const int64_t ct_min = (tiles + PROJ_MAX_CHUNKS - 1) / PROJ_MAX_CHUNKS;
    if (ct < ct_min) ct = ct_min;
    if (const char* e = getenv("ETL_PROJ_CHUNK_TILES")) {
        const int want = atoi(e);
        if (want >= ct_min) ct = want;
    }
    p.chunk_tiles = (int)ct;
    p.chunks = (tiles + p.chunk_tiles - 1) / p.chunk_tiles;
    if (grid > p.chunks) grid = p.chunks;
    int rc = tile_state_acquire(h, (size_t)p.chunks, stream, &p.epoch);
    if (rc != ETL_OK) return rc;
    p.chunk_state = h->tile_state;
    p.ticket = h->counters + 2;
    project_filter_impute_kernel<<<(unsigned)grid, PROJ_THREADS, (size_t)s_stages * sizeof(ProjStage), stream>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

// ---------------------------------------------------------------------------
// Host-buffer variant: 3-slot pipeline over row chunks.  The kept-row count of
// a chunk depends only on the key bitmap, so the host popcounts it up front and
// every copy-out has a known size and destination - no mid-pipeline sync.
// ---------------------------------------------------------------------------
constexpr int64_t HOST_CHUNK_ROWS = (int64_t)1 << 20;
constexpr int HOST_SLOTS = 3;

// Chunk sizes of the host pipeline: uniform 1 Mi-row chunks (a ramp of smaller
// chunks at the head and tail, to shorten the un-overlapped first copy-in and
// last copy-out, measured slightly slower: more, smaller copies).  Every chunk
// but the last is a multiple of 1024 rows.  Returns the largest size.
static int64_t host_chunk_schedule(int64_t n, std::vector<int64_t>* sizes) {
    int64_t big = HOST_CHUNK_ROWS;
    if (const char* e = getenv("ETL_HOST_CHUNK_ROWS")) {
        const long long want = atoll(e);
        if (want >= 1024) big = (want + 1023) & ~(long long)1023;
    }
    sizes->clear();
    for (int64_t left = n; left > 0; left -= big) sizes->push_back(left < big ? left : big);
    const int64_t mx = n < big ? n : big;
    return (mx + 1023) & ~(int64_t)1023;
}

static int64_t host_popcount_bits(const uint32_t* bitmap, int64_t row0, int64_t rows) {
    if (!bitmap) return rows;
    int64_t cnt = 0;
    const int64_t w0 = row0 >> 5, words = (rows + 31) >> 5;  // row0 % 32 == 0
    for (int64_t w = 0; w < words; ++w) {
        uint32_t bits = bitmap[w0 + w];
        if (w == words - 1 && (rows & 31)) bits &= (1u << (rows & 31)) - 1;
        cnt += __builtin_popcount(bits);
    }
    return cnt;
}

static int ensure_host_pipeline(Handle* h, size_t stage_bytes) {
    for (auto& s : h->streams)
        if (!s) ETL_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    for (auto& e : h->events)
        if (!e) ETL_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    if (stage_bytes > h->stage_bytes) {
        ETL_CUDA(cudaDeviceSynchronize());
        if (h->stage) ETL_CUDA(cudaFree(h->stage));
        h->stage = nullptr;
        h->stage_bytes = 0;
        cudaError_t e = cudaMalloc(&h->stage, stage_bytes);
        if (e != cudaSuccess) {
            set_error("host pipeline: cudaMalloc(%zu) failed: %s", stage_bytes, cudaGetErrorString(e));
            return ETL_ENOMEM;
        }
        h->stage_bytes = stage_bytes;
    }
    return ETL_OK;
}

struct StageSlot {
    int32_t* code;
    uint32_t* key_valid;
    void* val[3];
    uint32_t* val_valid[3];
    int32_t* out_code;
    double* out_val[3];
    int64_t* out_rows;
};

static size_t carve_slot(char* base, int64_t chunk, StageSlot* s) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char* p = base ? base + off : nullptr;
        off += (bytes + 255) & ~(size_t)255;
        return p;
    };
    StageSlot t;
    t.code = (int32_t*)take(chunk * 4);
    t.key_valid = nullptr;  // bitmaps are staged whole, once (they are 1/64 of the data)
    for (int c = 0; c < 3; ++c) {
        t.val[c] = take(chunk * 8);
        t.val_valid[c] = nullptr;
    }
    t.out_code = (int32_t*)take(chunk * 4);
    for (int c = 0; c < 3; ++c) t.out_val[c] = (double*)take(chunk * 8);
    t.out_rows = (int64_t*)take(8);
    if (s) *s = t;
    return off;
}

int project_host(Handle* h, const int32_t* key_code, const uint32_t* key_valid, const void* const val[3],
                 const uint32_t* const val_valid[3], const int32_t val_dtype[3], const double fill[3], int64_t n,
                 int32_t* out_code, double* const out_val[3], int64_t* out_rows) {
    *out_rows = 0;
    if (n == 0) return ETL_OK;
    std::vector<int64_t> sizes;
    const int64_t chunk = host_chunk_schedule(n, &sizes);
    const size_t slot_bytes = carve_slot(nullptr, chunk, nullptr);
    const size_t bitmap_bytes = ((size_t)((n + 31) >> 5) * 4 + 255) & ~(size_t)255;
    int rc = ensure_host_pipeline(h, slot_bytes * HOST_SLOTS + 4 * bitmap_bytes);
    if (rc != ETL_OK) return rc;
    StageSlot slot[HOST_SLOTS];
    for (int s = 0; s < HOST_SLOTS; ++s) carve_slot(static_cast<char*>(h->stage) + s * slot_bytes, chunk, &slot[s]);
    uint32_t* d_bitmap[4];
    for (int b = 0; b < 4; ++b)
        d_bitmap[b] = reinterpret_cast<uint32_t*>(static_cast<char*>(h->stage) + slot_bytes * HOST_SLOTS + b * bitmap_bytes);
    cudaStream_t s_in = h->streams[0], s_k = h->streams[1], s_out = h->streams[2];
    cudaEvent_t* in_ready = h->events;            // [slot]
    cudaEvent_t* k_done = h->events + HOST_SLOTS; // [slot]
    cudaEvent_t* out_done = h->events + 2 * HOST_SLOTS;

    // all four validity bitmaps first: 4 copies instead of 4 per chunk
    const uint32_t* h_bitmap[4] = {key_valid, val_valid[0], val_valid[1], val_valid[2]};
    for (int b = 0; b < 4; ++b)
        if (h_bitmap[b])
            ETL_CUDA(cudaMemcpyAsync(d_bitmap[b], h_bitmap[b], (size_t)((n + 31) >> 5) * 4, cudaMemcpyHostToDevice, s_in));
    int64_t written = 0, row0 = 0;
    const int64_t chunks = (int64_t)sizes.size();
    for (int64_t i = 0; i < chunks; row0 += sizes[i], ++i) {
        const int s = (int)(i % HOST_SLOTS);
        const int64_t rows = sizes[i];
        const int64_t kept = host_popcount_bits(key_valid, row0, rows);
        StageSlot& st = slot[s];
        // copy-in (after the kernel that last read this slot's inputs)
        if (i >= HOST_SLOTS) ETL_CUDA(cudaStreamWaitEvent(s_in, k_done[s], 0));
        ETL_CUDA(cudaMemcpyAsync(st.code, key_code + row0, rows * 4, cudaMemcpyHostToDevice, s_in));
        for (int c = 0; c < 3; ++c) {
            const size_t es = dtype_size(val_dtype[c]);
            ETL_CUDA(cudaMemcpyAsync(st.val[c], static_cast<const char*>(val[c]) + row0 * es, rows * es,
                                     cudaMemcpyHostToDevice, s_in));
        }
        ETL_CUDA(cudaEventRecord(in_ready[s], s_in));
        // kernel (after the copy-out that last read this slot's outputs)
        ETL_CUDA(cudaStreamWaitEvent(s_k, in_ready[s], 0));
        if (i >= HOST_SLOTS) ETL_CUDA(cudaStreamWaitEvent(s_k, out_done[s], 0));
        const void* dval[3] = {st.val[0], st.val[1], st.val[2]};
        const uint32_t* dvalid[3] = {val_valid[0] ? d_bitmap[1] + (row0 >> 5) : nullptr,
                                     val_valid[1] ? d_bitmap[2] + (row0 >> 5) : nullptr,
                                     val_valid[2] ? d_bitmap[3] + (row0 >> 5) : nullptr};
        rc = launch_project(h, st.code, key_valid ? d_bitmap[0] + (row0 >> 5) : nullptr, dval, dvalid, val_dtype, fill, rows,
                            st.out_code, st.out_val, st.out_rows, s_k);
        if (rc != ETL_OK) return rc;
        ETL_CUDA(cudaEventRecord(k_done[s], s_k));
        // copy-out
        ETL_CUDA(cudaStreamWaitEvent(s_out, k_done[s], 0));
        if (kept > 0) {
            ETL_CUDA(cudaMemcpyAsync(out_code + written, st.out_code, kept * 4, cudaMemcpyDeviceToHost, s_out));
            for (int c = 0; c < 3; ++c)
                ETL_CUDA(cudaMemcpyAsync(out_val[c] + written, st.out_val[c], kept * 8, cudaMemcpyDeviceToHost, s_out));
        }
        ETL_CUDA(cudaEventRecord(out_done[s], s_out));
        written += kept;
    }
    ETL_CUDA(cudaStreamSynchronize(s_out));
    ETL_CUDA(cudaStreamSynchronize(s_k));
    *out_rows = written;
    return ETL_OK;
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_project_filter_impute(etl_handle* hh, const int32_t* key_code, const uint32_t* key_valid,
                                         const void* const val[3], const uint32_t* const val_valid[3],
                                         const int32_t val_dtype[3], const double fill[3], int64_t n,
                                         int32_t* out_code, double* const out_val[3], int64_t* d_out_rows,
                                         void* stream) {
    ETL_REQUIRE(hh != nullptr, "project: handle is NULL");
    ETL_REQUIRE(fill != nullptr && d_out_rows != nullptr, "project: fill/d_out_rows is NULL");
    Handle* h = reinterpret_cast<Handle*>(hh);
    int rc = etl::check_project_args(key_code, val, key_valid, val_valid, val_dtype, n, out_code, out_val);
    if (rc != ETL_OK) return rc;
    ETL_CUDA(cudaSetDevice(h->device));
    return etl::launch_project(h, key_code, key_valid, val, val_valid, val_dtype, fill, n, out_code, out_val,
                               d_out_rows, static_cast<cudaStream_t>(stream));
}

extern "C" int etl_project_filter_impute_host(etl_handle* hh, const int32_t* key_code, const uint32_t* key_valid,
                                              const void* const val[3], const uint32_t* const val_valid[3],
                                              const int32_t val_dtype[3], const double fill[3], int64_t n,
                                              int32_t* out_code, double* const out_val[3], int64_t* out_rows) {
    ETL_REQUIRE(hh != nullptr, "project_host: handle is NULL");
    ETL_REQUIRE(fill != nullptr && out_rows != nullptr, "project_host: fill/out_rows is NULL");
    Handle* h = reinterpret_cast<Handle*>(hh);
    int rc = etl::check_project_args(key_code, val, key_valid, val_valid, val_dtype, n, out_code, out_val, false);
    if (rc != ETL_OK) return rc;
    ETL_CUDA(cudaSetDevice(h->device));
    return etl::project_host(h, key_code, key_valid, val, val_valid, val_dtype, fill, n, out_code, out_val, out_rows);
}

// ---------------------------------------------------------------------------
// Allocation-free entry point.
//
// Who needs it: a caller whose device memory is already spoken for.  Typical
// cases are a host process that gave the whole GPU to a pool allocator (a torch
// or RMM arena sized to the card: a raw cudaMalloc from inside this library
// would fail, or succeed only by taking memory the pool thinks it owns), and a
// latency-sensitive loop that cannot afford the implicit device-wide
// synchronisation of cudaMalloc / cudaFree.  Such callers cannot use
// etl_create(), because a handle allocates its scratch.
//
// What it does instead: the projection kernel's scratch (one tagged 64-bit word
// per chunk plus the two ticket words) lives in a file-scope __device__ array
// that every GPU loading this library carries until the process exits
// (NOALLOC_SCRATCH_BYTES = 131,328 bytes).  Nothing is allocated, here or in
// any library routine; inputs and outputs are still the caller's buffers.
//
// Rules that follow from the shared, fixed-size scratch:
//   * n is limited to ETL_NOALLOC_MAX_ROWS (2^31 rows); larger tables are
//     rejected with ETL_ECAPACITY before anything touches the GPU;
//   * the first call on each device looks the array's device address up
//     (cudaGetSymbolAddress - in host code the array's name is only a handle,
//     not a pointer a kernel can use), reads the SM count and sizes the grid.
//     Make that first call before starting any stream capture;
//   * every call on a device uses the same words.  Each call clears the chunk
//     words it is going to use and the ticket words on the caller's stream (and
//     tags the chunk words with a fresh epoch), so calls issued one after
//     another on one stream are independent;
//     two calls that overlap on the GPU (different streams, no ordering) race
//     on the scratch and produce garbage.
// ---------------------------------------------------------------------------
#include <mutex>

namespace etl {

constexpr int64_t NOALLOC_MAX_CHUNKS = 16384;
constexpr int64_t NOALLOC_TICKET_WORDS = 32;  // 256 bytes: keeps the ticket on its own lines
constexpr int64_t NOALLOC_MAX_ROWS = (int64_t)1 << 31;
constexpr size_t NOALLOC_SCRATCH_BYTES = (NOALLOC_MAX_CHUNKS + NOALLOC_TICKET_WORDS) * sizeof(uint64_t);
constexpr int NOALLOC_STAGES = 3;
constexpr int NOALLOC_MAX_DEVICES = 64;

static __device__ __align__(256) uint64_t g_noalloc_scratch[NOALLOC_MAX_CHUNKS + NOALLOC_TICKET_WORDS];
static_assert(sizeof(g_noalloc_scratch) == NOALLOC_SCRATCH_BYTES && NOALLOC_SCRATCH_BYTES == 131328,
              "the documented scratch size is part of the interface");

struct NoAllocDevice {
    bool ready = false;
    uint64_t* scratch = nullptr;  // device address of g_noalloc_scratch on this GPU
    int sm_count = 0;
    int ctas_per_sm = 0;
    uint32_t epoch = 0;
};
static NoAllocDevice g_noalloc_dev[NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mutex;

// one-time set-up for the current device (caller holds the mutex)
static int noalloc_setup(int dev, NoAllocDevice* d) {
    uint64_t* ptr = nullptr;
    ETL_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    if (ptr == nullptr) {
        set_error("project_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev);
        return ETL_ECUDA;
    }
    ETL_CUDA(cudaDeviceGetAttribute(&d->sm_count, cudaDevAttrMultiProcessorCount, dev));
    const size_t dyn = (size_t)NOALLOC_STAGES * sizeof(ProjStage);
    // raise the opt-in limit to the largest ring any path uses, so neither path undercuts the other
    ETL_CUDA(cudaFuncSetAttribute(project_filter_impute_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(PROJ_MAX_STAGES * sizeof(ProjStage))));
    int occ = 0;
    ETL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, project_filter_impute_kernel, PROJ_THREADS, dyn));
    if (occ < 1) {
        set_error("project_noalloc: kernel does not fit on an SM of device %d", dev);
        return ETL_ECUDA;
    }
    d->ctas_per_sm = occ;
    d->scratch = ptr;
    d->epoch = 0;  // the array is zero at module load; 0 is never a live tag
    d->ready = true;
    return ETL_OK;
}

}  // namespace etl

extern "C" int etl_project_filter_impute_noalloc(const int32_t* key_code, const uint32_t* key_valid,
                                                 const void* const val[3], const uint32_t* const val_valid[3],
                                                 const int32_t val_dtype[3], const double fill[3], int64_t n,
                                                 int32_t* out_code, double* const out_val[3], int64_t* d_out_rows,
                                                 void* stream_) {
    using namespace etl;
    ETL_REQUIRE(fill != nullptr && d_out_rows != nullptr, "project_noalloc: fill/d_out_rows is NULL");
    if (n > NOALLOC_MAX_ROWS) {
        set_error("project_noalloc: n=%lld exceeds the fixed scratch (max %lld rows); use etl_project_filter_impute",
                  (long long)n, (long long)NOALLOC_MAX_ROWS);
        return ETL_ECAPACITY;
    }
    int rc = check_project_args(key_code, val, key_valid, val_valid, val_dtype, n, out_code, out_val);
    if (rc != ETL_OK) return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    int dev = -1;
    ETL_CUDA(cudaGetDevice(&dev));
    ETL_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "project_noalloc: device ordinal %d unsupported", dev);

    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    NoAllocDevice* d = &g_noalloc_dev[dev];
    if (!d->ready) {
        rc = noalloc_setup(dev, d);
        if (rc != ETL_OK) return rc;
    }
    if (n == 0) {
        ETL_CUDA(cudaMemsetAsync(d_out_rows, 0, sizeof(int64_t), stream));
        return ETL_OK;
    }
    const int64_t tiles = (n + PROJ_TILE - 1) / PROJ_TILE;
    ProjParams p;
    p.key_code = key_code;
    p.key_valid = key_valid;
    for (int c = 0; c < 3; ++c) {
        p.val[c] = val[c];
        p.val_valid[c] = val_valid[c];
        p.val_dtype[c] = val_dtype[c];
        p.fill[c] = fill[c];
        p.out_val[c] = out_val[c];
    }
    p.n = n;
    p.out_code = out_code;
    p.out_rows = d_out_rows;
    p.stages = NOALLOC_STAGES;
    p.hints = 3;
    int64_t grid = (int64_t)d->sm_count * d->ctas_per_sm;
    if (grid > tiles) grid = tiles;
    int64_t ct = tiles / (grid * 8);
    ct = ct < 1 ? 1 : (ct > 4 ? 4 : ct);
    const int64_t ct_min = (tiles + NOALLOC_MAX_CHUNKS - 1) / NOALLOC_MAX_CHUNKS;
    if (ct < ct_min) ct = ct_min;
    p.chunk_tiles = (int)ct;
    p.chunks = (tiles + ct - 1) / ct;  // <= NOALLOC_MAX_CHUNKS by construction
    if (grid > p.chunks) grid = p.chunks;
    p.chunk_state = d->scratch;
    p.ticket = reinterpret_cast<unsigned int*>(d->scratch + NOALLOC_MAX_CHUNKS);
    // the array still holds the previous call's words: a fresh tag retires the chunk words
    // (wiped outright when the 22-bit tag wraps), and the ticket pair must start at zero
    if (d->epoch >= TS_EPOCH_MAX) {
        ETL_CUDA(cudaMemsetAsync(d->scratch, 0, NOALLOC_MAX_CHUNKS * sizeof(uint64_t), stream));
        d->epoch = 0;
    }
    p.epoch = ++d->epoch;
    // ... and the words this call will use are cleared on the caller's stream as well, so a call's result
    // never depends on what an earlier call left behind (at most 128 KiB, usually a few KiB)
    ETL_CUDA(cudaMemsetAsync(d->scratch, 0, (size_t)p.chunks * sizeof(uint64_t), stream));
    ETL_CUDA(cudaMemsetAsync(p.ticket, 0, 2 * sizeof(unsigned int), stream));
    project_filter_impute_kernel<<<(unsigned)grid, PROJ_THREADS, (size_t)NOALLOC_STAGES * sizeof(ProjStage), stream>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int64_t etl_project_filter_impute_host_launches(int64_t n) {
    if (n <= 0) return 0;
    std::vector<int64_t> sizes;
    etl::host_chunk_schedule(n, &sizes);
    return (int64_t)sizes.size();
}
