// Fused map-side projection: isNotNull filter + mean-impute + cast-to-double +
// order-preserving compaction, one pass over the rows (k_means.py:27-28,49-51).
//
// One CTA = one tile of 1024 rows, 256 threads x 4 consecutive rows.  Every
// thread issues all of its 128-bit loads first (1 x int4 of key codes and, for
// fp64 columns, 2 x double2 per column: 7 independent 16-byte requests in
// flight per thread), then the keep bits are ranked with warp ballots, the tile
// total goes through a single-pass decoupled look-back (one 64-bit state word
// per tile), kept rows are staged in shared memory in output order and leave as
// fully coalesced stores.  HBM traffic = the algorithmic bytes: each input byte
// is read once and each output byte written once.
#include "common.cuh"

namespace etl {

constexpr int PROJ_THREADS = 256;
constexpr int PROJ_RPT = 4;  // rows per thread
constexpr int PROJ_TILE = PROJ_THREADS * PROJ_RPT;

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
    uint64_t* tile_state;
    uint32_t epoch;
};

// 4 consecutive values of one column as doubles (the implicit cast of the
// when/otherwise expression).  Full tiles take the 128-bit path.
__device__ __forceinline__ void load4_full(const void* base, int dtype, int64_t row, double v[4]) {
    if (dtype == ETL_F64) {
        const double2 a = ld_stream_d2(static_cast<const double*>(base) + row);
        const double2 b = ld_stream_d2(static_cast<const double*>(base) + row + 2);
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
    } else {
        const int4 q = ld_stream_v4(static_cast<const int32_t*>(base) + row);
        if (dtype == ETL_F32) {
            v[0] = (double)__int_as_float(q.x); v[1] = (double)__int_as_float(q.y);
            v[2] = (double)__int_as_float(q.z); v[3] = (double)__int_as_float(q.w);
        } else {
            v[0] = (double)q.x; v[1] = (double)q.y; v[2] = (double)q.z; v[3] = (double)q.w;
        }
    }
}

__device__ __forceinline__ double load1(const void* base, int dtype, int64_t row) {
    if (dtype == ETL_F64) return static_cast<const double*>(base)[row];
    if (dtype == ETL_F32) return (double)static_cast<const float*>(base)[row];
    return (double)static_cast<const int32_t*>(base)[row];
}

// the 4 validity bits of rows [row, row+4), row % 4 == 0; rows >= n read as 0
__device__ __forceinline__ unsigned valid_nibble(const uint32_t* bitmap, int64_t row, int64_t n) {
    unsigned nib = 0xF;
    if (bitmap) nib = (__ldg(bitmap + (row >> 5)) >> (row & 31)) & 0xF;
    const int64_t left = n - row;
    if (left < 4) nib &= left <= 0 ? 0u : ((1u << left) - 1);
    return nib;
}

__global__ void __launch_bounds__(PROJ_THREADS) project_filter_impute_kernel(const ProjParams p) {
    __shared__ int32_t s_code[PROJ_TILE];
    __shared__ double s_val[3][PROJ_TILE];
    __shared__ uint32_t s_warp_count[PROJ_THREADS / 32];
    __shared__ uint64_t s_tile_base;

    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int64_t tile = blockIdx.x;
    const int64_t row = tile * PROJ_TILE + (int64_t)tid * PROJ_RPT;
    const bool full = (tile + 1) * PROJ_TILE <= p.n;

    // ---- loads first: everything below until the staging step depends only on
    // the validity nibbles, so the wide loads stay in flight across the scan
    int32_t code[4];
    double v[3][4];
    if (full) {
        const int4 q = ld_stream_v4(p.key_code + row);
        code[0] = q.x; code[1] = q.y; code[2] = q.z; code[3] = q.w;
#pragma unroll
        for (int c = 0; c < 3; ++c) load4_full(p.val[c], p.val_dtype[c], row, v[c]);
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool in = row + j < p.n;
            code[j] = in ? p.key_code[row + j] : 0;
#pragma unroll
            for (int c = 0; c < 3; ++c) v[c][j] = in ? load1(p.val[c], p.val_dtype[c], row + j) : 0.0;
        }
    }
    const unsigned keep = valid_nibble(p.key_valid, row, p.n);
    unsigned vnib[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) vnib[c] = valid_nibble(p.val_valid[c], row, p.n);

    // ---- rank kept rows: 4 ballots (one per row slot), rows ordered lane-major
    const unsigned lt = lanemask_lt();
    unsigned below = 0, warp_total = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const unsigned b = __ballot_sync(0xffffffffu, (keep >> j) & 1u);
        below += __popc(b & lt);
        warp_total += __popc(b);
    }
    if (lane == 0) s_warp_count[warp] = warp_total;
    __syncthreads();
    unsigned warp_base = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < PROJ_THREADS / 32; ++w) {
        const unsigned cnt = s_warp_count[w];
        if (w < warp) warp_base += cnt;
        tile_total += cnt;
    }
    if (warp == 0) {
        const uint64_t excl = lookback_exclusive(p.tile_state, p.epoch, tile, tile_total);
        if (lane == 0) {
            s_tile_base = excl;
            if (tile == gridDim.x - 1) *p.out_rows = (int64_t)(excl + tile_total);
        }
    }

    // ---- impute + stage in output order
    unsigned pos = warp_base + below;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if ((keep >> j) & 1u) {
            s_code[pos] = code[j];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const double x = v[c][j];
                const bool bad = !((vnib[c] >> j) & 1u) || (x != x);
                s_val[c][pos] = bad ? p.fill[c] : x;
            }
            ++pos;
        }
    }
    __syncthreads();

    // ---- coalesced copy-out of the tile's kept rows
    const uint64_t base = s_tile_base;
    for (unsigned i = tid; i < tile_total; i += PROJ_THREADS) p.out_code[base + i] = s_code[i];
#pragma unroll
    for (int c = 0; c < 3; ++c)
        for (unsigned i = tid; i < tile_total; i += PROJ_THREADS) p.out_val[c][base + i] = s_val[c][i];
}

static int check_project_args(const int32_t* key_code, const void* const val[3], const uint32_t* key_valid,
                              const uint32_t* const val_valid[3], const int32_t val_dtype[3], int64_t n,
                              const int32_t* out_code, double* const out_val[3]) {
    ETL_REQUIRE(n >= 0 && n < ((int64_t)1 << 39), "project: n=%lld out of range", (long long)n);
    ETL_REQUIRE(val && val_valid && val_dtype && out_val, "project: NULL argument array");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(key_code && out_code, "project: key_code/out_code is NULL");
    ETL_REQUIRE(aligned16(key_code) && aligned4(key_valid), "project: key buffers misaligned");
    for (int c = 0; c < 3; ++c) {
        ETL_REQUIRE(val[c] && out_val[c], "project: val[%d]/out_val[%d] is NULL", c, c);
        ETL_REQUIRE(val_dtype[c] == ETL_F64 || val_dtype[c] == ETL_F32 || val_dtype[c] == ETL_I32,
                    "project: val_dtype[%d]=%d unsupported", c, val_dtype[c]);
        ETL_REQUIRE(aligned16(val[c]) && aligned4(val_valid[c]), "project: val[%d] buffers misaligned", c);
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
    int rc = tile_state_acquire(h, (size_t)tiles, stream, &p.epoch);
    if (rc != ETL_OK) return rc;
    p.tile_state = h->tile_state;
    project_filter_impute_kernel<<<(unsigned)tiles, PROJ_THREADS, 0, stream>>>(p);
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
    const size_t words = (size_t)((chunk + 31) >> 5) * 4;
    StageSlot t;
    t.code = (int32_t*)take(chunk * 4);
    t.key_valid = (uint32_t*)take(words);
    for (int c = 0; c < 3; ++c) {
        t.val[c] = take(chunk * 8);
        t.val_valid[c] = (uint32_t*)take(words);
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
    const int64_t chunk = n < HOST_CHUNK_ROWS ? ((n + 1023) & ~(int64_t)1023) : HOST_CHUNK_ROWS;
    const size_t slot_bytes = carve_slot(nullptr, chunk, nullptr);
    int rc = ensure_host_pipeline(h, slot_bytes * HOST_SLOTS);
    if (rc != ETL_OK) return rc;
    StageSlot slot[HOST_SLOTS];
    for (int s = 0; s < HOST_SLOTS; ++s) carve_slot(static_cast<char*>(h->stage) + s * slot_bytes, chunk, &slot[s]);
    cudaStream_t s_in = h->streams[0], s_k = h->streams[1], s_out = h->streams[2];
    cudaEvent_t* in_ready = h->events;            // [slot]
    cudaEvent_t* k_done = h->events + HOST_SLOTS; // [slot]
    cudaEvent_t* out_done = h->events + 2 * HOST_SLOTS;

    int64_t written = 0;
    const int64_t chunks = (n + chunk - 1) / chunk;
    for (int64_t i = 0; i < chunks; ++i) {
        const int s = (int)(i % HOST_SLOTS);
        const int64_t row0 = i * chunk;
        const int64_t rows = n - row0 < chunk ? n - row0 : chunk;
        const size_t words = (size_t)((rows + 31) >> 5);
        const int64_t kept = host_popcount_bits(key_valid, row0, rows);
        StageSlot& st = slot[s];
        // copy-in (after the kernel that last read this slot's inputs)
        if (i >= HOST_SLOTS) ETL_CUDA(cudaStreamWaitEvent(s_in, k_done[s], 0));
        ETL_CUDA(cudaMemcpyAsync(st.code, key_code + row0, rows * 4, cudaMemcpyHostToDevice, s_in));
        if (key_valid)
            ETL_CUDA(cudaMemcpyAsync(st.key_valid, key_valid + (row0 >> 5), words * 4, cudaMemcpyHostToDevice, s_in));
        for (int c = 0; c < 3; ++c) {
            const size_t es = dtype_size(val_dtype[c]);
            ETL_CUDA(cudaMemcpyAsync(st.val[c], static_cast<const char*>(val[c]) + row0 * es, rows * es,
                                     cudaMemcpyHostToDevice, s_in));
            if (val_valid[c])
                ETL_CUDA(cudaMemcpyAsync(st.val_valid[c], val_valid[c] + (row0 >> 5), words * 4,
                                         cudaMemcpyHostToDevice, s_in));
        }
        ETL_CUDA(cudaEventRecord(in_ready[s], s_in));
        // kernel (after the copy-out that last read this slot's outputs)
        ETL_CUDA(cudaStreamWaitEvent(s_k, in_ready[s], 0));
        if (i >= HOST_SLOTS) ETL_CUDA(cudaStreamWaitEvent(s_k, out_done[s], 0));
        const void* dval[3] = {st.val[0], st.val[1], st.val[2]};
        const uint32_t* dvalid[3] = {val_valid[0] ? st.val_valid[0] : nullptr, val_valid[1] ? st.val_valid[1] : nullptr,
                                     val_valid[2] ? st.val_valid[2] : nullptr};
        rc = launch_project(h, st.code, key_valid ? st.key_valid : nullptr, dval, dvalid, val_dtype, fill, rows,
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
    int rc = etl::check_project_args(key_code, val, key_valid, val_valid, val_dtype, n, out_code, out_val);
    if (rc != ETL_OK) return rc;
    ETL_CUDA(cudaSetDevice(h->device));
    return etl::project_host(h, key_code, key_valid, val, val_valid, val_dtype, fill, n, out_code, out_val, out_rows);
}
