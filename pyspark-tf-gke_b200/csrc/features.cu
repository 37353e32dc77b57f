// Feature-side kernels of the ETL stage:
//   code_histogram        StringIndexer.fit's group-by-count (k_means.py:34,73)
//   assemble_dense/sparse OneHotEncoder + VectorAssembler([vec]*repeats + numerics)
//                         (k_means.py:38,64-68,74-77), optional standardisation
//   pack_tf               the X float32[N,3] / y int32[N] layout of load_csv
//                         (train_tf_ps.py:143-149) that from_tensor_slices reads
//   bucketize             Spark ML Bucketizer semantics (north-star extension)
// All are HBM-bound streaming kernels; the assemblers are write-bound
// (1224 B written per 28 B read at D = 153).
#include "common.cuh"

namespace etl {

// ------------------------------------------------------------- histogram
constexpr int HIST_THREADS = 256;
constexpr int HIST_MAX_SMEM_BINS = 8192;

__global__ void __launch_bounds__(HIST_THREADS) code_histogram_kernel(const int32_t* __restrict__ codes,
                                                                      const uint32_t* __restrict__ valid, int64_t n,
                                                                      int32_t bins, unsigned long long* __restrict__ out,
                                                                      unsigned long long* __restrict__ out_of_range) {
    extern __shared__ uint32_t s_hist[];
    const bool use_smem = bins <= HIST_MAX_SMEM_BINS;
    if (use_smem) {
        for (int i = threadIdx.x; i < bins; i += HIST_THREADS) s_hist[i] = 0;
        __syncthreads();
    }
    unsigned long long bad = 0;
    const int64_t tiles = (n + HIST_THREADS * 4 - 1) / (HIST_THREADS * 4);
    for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int64_t row = t * HIST_THREADS * 4 + (int64_t)threadIdx.x * 4;
        int32_t c[4] = {-1, -1, -1, -1};
        if (row + 4 <= n) {
            const int4 q = ld_stream_v4(codes + row);
            c[0] = q.x; c[1] = q.y; c[2] = q.z; c[3] = q.w;
        } else {
            for (int j = 0; j < 4; ++j)
                if (row + j < n) c[j] = codes[row + j];
        }
        unsigned nib = 0xF;
        if (valid && row < n) nib = (__ldg(valid + (row >> 5)) >> (row & 31)) & 0xF;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (row + j >= n || !((nib >> j) & 1u)) continue;
            if (c[j] < 0 || c[j] >= bins) { ++bad; continue; }
            if (use_smem) atomicAdd(&s_hist[c[j]], 1u);  // native 32-bit shared atomic
            else atomicAdd(out + c[j], 1ull);
        }
    }
    if (bad) atomicAdd(out_of_range, bad);
    if (use_smem) {
        __syncthreads();
        for (int i = threadIdx.x; i < bins; i += HIST_THREADS)
            if (s_hist[i]) atomicAdd(out + i, (unsigned long long)s_hist[i]);
    }
}

// ----------------------------------------------------- dense / sparse assembly
struct AssembleParams {
    const int32_t* index;     // final label index per row, in [0, num_labels]; num_labels = the all-zero vector
    const double* num[3];
    double shift[3], scale[3];  // numerics leave as (x - shift) * scale  (0 / 1 = plain copy)
    int64_t n;
    int32_t num_labels, repeats;
    void* out;                // [n, repeats * num_labels + 3] row-major
};

constexpr int ASM_THREADS = 256;
constexpr int ASM_MAXJ = 8;  // columns per lane held in registers: D <= 256 takes the fast kernel

// One warp per row: lane l writes columns l, l + 32, ... so every store
// instruction of the warp covers 32 consecutive elements of the row-major
// output.  A lane's columns are the same for every row, so what each of them
// means - which label a one-hot column stands for, or which numeric it carries -
// is worked out once per thread; per row and element only a compare, a select
// and the store remain.
template <typename T>
__global__ void __launch_bounds__(ASM_THREADS) assemble_dense_kernel(const AssembleParams p) {
    const int lane = threadIdx.x & 31;
    const int D = p.repeats * p.num_labels + 3;
    const int onehot = p.repeats * p.num_labels;
    int label_of[ASM_MAXJ];  // >= 0: one-hot column of that label; -1..-3: numeric 0..2; -4: past the row
#pragma unroll
    for (int j = 0; j < ASM_MAXJ; ++j) {
        const int col = lane + 32 * j;
        label_of[j] = col >= D ? -4 : (col >= onehot ? -1 - (col - onehot) : col % p.num_labels);
    }
    T* out = static_cast<T*>(p.out);
    const int64_t warps = (int64_t)gridDim.x * (ASM_THREADS / 32);
    for (int64_t r = (int64_t)blockIdx.x * (ASM_THREADS / 32) + (threadIdx.x >> 5); r < p.n; r += warps) {
        int32_t idx = __ldg(p.index + r);
        if (idx < 0 || idx >= p.num_labels) idx = -100;  // unknown label: matches no column
        const T x0 = (T)((__ldg(p.num[0] + r) - p.shift[0]) * p.scale[0]);
        const T x1 = (T)((__ldg(p.num[1] + r) - p.shift[1]) * p.scale[1]);
        const T x2 = (T)((__ldg(p.num[2] + r) - p.shift[2]) * p.scale[2]);
        T* row = out + r * D + lane;
#pragma unroll
        for (int j = 0; j < ASM_MAXJ; ++j) {
            const int l = label_of[j];
            if (l == -4) break;
            const T v = l >= 0 ? (T)(l == idx ? 1 : 0) : (l == -1 ? x0 : (l == -2 ? x1 : x2));
            row[32 * j] = v;
        }
    }
}

// any feature dimension (D > 256): same mapping, column meaning recomputed per element
template <typename T>
__global__ void __launch_bounds__(ASM_THREADS) assemble_dense_wide_kernel(const AssembleParams p) {
    const int lane = threadIdx.x & 31;
    const int D = p.repeats * p.num_labels + 3;
    const int onehot = p.repeats * p.num_labels;
    T* out = static_cast<T*>(p.out);
    const int64_t warps = (int64_t)gridDim.x * (ASM_THREADS / 32);
    for (int64_t r = (int64_t)blockIdx.x * (ASM_THREADS / 32) + (threadIdx.x >> 5); r < p.n; r += warps) {
        int32_t idx = __ldg(p.index + r);
        if (idx < 0 || idx >= p.num_labels) idx = -100;
        const T x0 = (T)((__ldg(p.num[0] + r) - p.shift[0]) * p.scale[0]);
        const T x1 = (T)((__ldg(p.num[1] + r) - p.shift[1]) * p.scale[1]);
        const T x2 = (T)((__ldg(p.num[2] + r) - p.shift[2]) * p.scale[2]);
        T* row = out + r * D;
        for (int col = lane; col < D; col += 32) {
            const int c = col - onehot;
            row[col] = c < 0 ? (T)(col % p.num_labels == idx ? 1 : 0) : (c == 0 ? x0 : (c == 1 ? x1 : x2));
        }
    }
}

// The same matrix through shared memory and the TMA engine.  The warp-per-row kernel above
// stores element by element - a 612-byte fp32 row (D = 153) is never 16-byte aligned - and
// spends 4x the store instructions per byte on fp32 (0.62 of copy peak against 0.89 for fp64).
// Here a CTA assembles ASM_CHUNK rows (a multiple of 4, so the chunk's bytes are a multiple of
// 16 for any D) in shared memory with the same lane-per-column mapping, then ONE bulk async
// copy (cp.async.bulk.global.shared) moves the chunk out; two buffers, so the copy of one chunk
// overlaps the assembly of the next.  The last, partial chunk leaves by plain stores.  The
// chunk's inputs (index + three numerics of 32 rows) are fetched one chunk ahead.

__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}

constexpr int ASM_CHUNK = 32;

template <typename T>
__global__ void __launch_bounds__(ASM_THREADS) assemble_dense_tma_kernel(const AssembleParams p) {
    extern __shared__ __align__(128) unsigned char asm_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = p.repeats * p.num_labels + 3;
    const int onehot = p.repeats * p.num_labels;
    int label_of[ASM_MAXJ];  // >= 0: one-hot column of that label; -1..-3: numeric 0..2; -4: past the row
#pragma unroll
    for (int j = 0; j < ASM_MAXJ; ++j) {
        const int col = lane + 32 * j;
        label_of[j] = col >= D ? -4 : (col >= onehot ? -1 - (col - onehot) : col % p.num_labels);
    }
    const size_t chunk_elems = (size_t)ASM_CHUNK * D;
    T* out = static_cast<T*>(p.out);
    const int64_t chunks = (p.n + ASM_CHUNK - 1) / ASM_CHUNK;
    // inputs of a chunk: threads 0..127 fetch one value each (array = warp: index, then the three
    // numerics; row = lane) one chunk AHEAD, so the loads' latency overlaps the assembly
    __shared__ double s_in[2][4][ASM_CHUNK];
    constexpr int FETCHERS = ASM_CHUNK / 8;   // warps that fetch: 4 arrays x ASM_CHUNK rows, one value per thread
    const int arr = threadIdx.x / ASM_CHUNK, frow = threadIdx.x % ASM_CHUNK;
    auto fetch = [&](int64_t c) -> double {
        const int64_t r = c * ASM_CHUNK + frow;
        if (r >= p.n) return 0.0;
        return arr == 0 ? (double)__ldg(p.index + r) : __ldg(p.num[arr - 1] + r);
    };
    double pre = 0.0;
    if (warp < FETCHERS && (int64_t)blockIdx.x < chunks) pre = fetch(blockIdx.x);
    int it = 0;
    for (int64_t c = blockIdx.x; c < chunks; c += gridDim.x, ++it) {
        T* sm = reinterpret_cast<T*>(asm_smem) + (it & 1) * chunk_elems;
        // the bulk copy that read this buffer two chunks ago must have finished reading it
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        if (warp < FETCHERS) s_in[it & 1][arr][frow] = pre;
        __syncthreads();
        if (warp < FETCHERS && c + gridDim.x < chunks) pre = fetch(c + gridDim.x);
        const int64_t row0 = c * ASM_CHUNK;
        const int rows = p.n - row0 < ASM_CHUNK ? (int)(p.n - row0) : ASM_CHUNK;
        for (int q = warp; q < rows; q += ASM_THREADS / 32) {
            int32_t idx = (int32_t)s_in[it & 1][0][q];
            if (idx < 0 || idx >= p.num_labels) idx = -100;  // unknown label: matches no column
            const T x0 = (T)((s_in[it & 1][1][q] - p.shift[0]) * p.scale[0]);
            const T x1 = (T)((s_in[it & 1][2][q] - p.shift[1]) * p.scale[1]);
            const T x2 = (T)((s_in[it & 1][3][q] - p.shift[2]) * p.scale[2]);
            T* row = sm + (size_t)q * D + lane;
#pragma unroll
            for (int j = 0; j < ASM_MAXJ; ++j) {
                const int l = label_of[j];
                if (l == -4) break;
                row[32 * j] = l >= 0 ? (T)(l == idx ? 1 : 0) : (l == -1 ? x0 : (l == -2 ? x1 : x2));
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the copy engine
        __syncthreads();
        T* dst = out + row0 * D;
        if (rows == ASM_CHUNK) {
            if (threadIdx.x == 0) {
                bulk_store(dst, sm, (unsigned)(chunk_elems * sizeof(T)));
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        } else {
            for (int i = threadIdx.x; i < rows * D; i += ASM_THREADS) dst[i] = sm[i];
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // smem must outlive the copies
}

// ELL form of Spark's SparseVector rows: repeats + 3 (index, value) slots per
// row in ascending index order; an unused slot (unknown label) has index -1.
__global__ void __launch_bounds__(256) assemble_sparse_kernel(const AssembleParams p, int32_t* __restrict__ out_idx,
                                                              double* __restrict__ out_val) {
    const int W = p.repeats + 3;
    const int64_t total = p.n * W;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / W;
        const int s = (int)(e - r * W);
        int32_t ix;
        double v;
        if (s < p.repeats) {
            const int32_t li = p.index[r];
            const bool known = li >= 0 && li < p.num_labels;
            ix = known ? s * p.num_labels + li : -1;
            v = known ? 1.0 : 0.0;
        } else {
            const int c = s - p.repeats;
            ix = p.repeats * p.num_labels + c;
            v = (p.num[c][r] - p.shift[c]) * p.scale[c];
        }
        out_idx[e] = ix;
        out_val[e] = v;
    }
}

// ------------------------------------------------------------------ TF layout
// keep = label non-null & every numeric non-null (the decoder, or
// etl_nan_to_null, has already turned NaN cells into nulls): skip rule of
// load_csv (train_tf_ps.py:120-131).  Bitmap-only predicate -> kept rows per
// 1024-row tile come from popcounts, a single CTA scans the tile counts, and the
// writer places every kept row directly.
constexpr int TF_TILE = 1024;

struct TfPackParams {
    const int32_t* label;        // dictionary slot per row
    const uint32_t* valid[4];    // label, num0, num1, num2 (NULL = no nulls)
    const double* num[3];
    double shift[3], scale[3];
    const int32_t* remap;        // slot -> class index (sorted-vocabulary position); may be NULL
    int32_t remap_size;
    int64_t n;
    float* X;                    // [kept, 3] row-major
    int32_t* y;                  // [kept]
    int64_t* tile_base;          // scratch: [tiles + 1]
    int64_t* out_rows;
};

__device__ __forceinline__ uint32_t tf_keep_word(const TfPackParams& p, int64_t w, int64_t n) {
    uint32_t k = 0xffffffffu;
#pragma unroll
    for (int c = 0; c < 4; ++c)
        if (p.valid[c]) k &= __ldg(p.valid[c] + w);
    const int64_t left = n - (w << 5);
    if (left < 32) k &= left <= 0 ? 0u : ((1u << left) - 1);
    return k;
}

__global__ void __launch_bounds__(256) tf_tile_count_kernel(const TfPackParams p) {
    // one warp per tile: 32 keep words
    const int lane = threadIdx.x & 31;
    const int64_t tiles = (p.n + TF_TILE - 1) / TF_TILE;
    const int64_t words = (p.n + 31) >> 5;
    for (int64_t t = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); t < tiles; t += (int64_t)gridDim.x * 8) {
        const int64_t w = t * 32 + lane;
        unsigned c = w < words ? __popc(tf_keep_word(p, w, p.n)) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0) p.tile_base[t] = c;
    }
}

__global__ void __launch_bounds__(1024) tf_tile_scan_kernel(int64_t* tile_base, int64_t tiles, int64_t* out_rows) {
    __shared__ int64_t s_sum[1024];
    const int tid = threadIdx.x;
    const int64_t per = (tiles + 1023) / 1024;
    const int64_t lo = tid * per, hi = lo + per < tiles ? lo + per : tiles;
    int64_t s = 0;
    for (int64_t i = lo; i < hi; ++i) s += tile_base[i];
    s_sum[tid] = s;
    __syncthreads();
    if (tid == 0) {
        int64_t run = 0;
        for (int t = 0; t < 1024; ++t) { const int64_t v = s_sum[t]; s_sum[t] = run; run += v; }
        *out_rows = run;
        tile_base[tiles] = run;
    }
    __syncthreads();
    int64_t run = s_sum[tid];
    for (int64_t i = lo; i < hi; ++i) { const int64_t v = tile_base[i]; tile_base[i] = run; run += v; }
}

__global__ void __launch_bounds__(256) tf_pack_kernel(const TfPackParams p) {
    __shared__ uint32_t s_word[32];
    __shared__ uint32_t s_off[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = lanemask_lt();
    const int64_t tiles = (p.n + TF_TILE - 1) / TF_TILE;
    const int64_t words = (p.n + 31) >> 5;
    for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        __syncthreads();
        if (warp == 0) {
            const int64_t w = t * 32 + lane;
            const uint32_t k = w < words ? tf_keep_word(p, w, p.n) : 0u;
            const unsigned cnt = __popc(k);
            unsigned incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += up;
            }
            s_word[lane] = k;
            s_off[lane] = incl - cnt;
        }
        __syncthreads();
        const int64_t base = p.tile_base[t];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int g = j * 8 + warp;
            const uint32_t k = s_word[g];
            if (!((k >> lane) & 1u)) continue;
            const int64_t r = t * TF_TILE + g * 32 + lane;
            const int64_t pos = base + s_off[g] + __popc(k & lt);
            const int32_t slot = p.label[r];
            p.y[pos] = p.remap ? ((slot >= 0 && slot < p.remap_size) ? __ldg(p.remap + slot) : -1) : slot;
#pragma unroll
            for (int c = 0; c < 3; ++c) p.X[pos * 3 + c] = (float)((p.num[c][r] - p.shift[c]) * p.scale[c]);
        }
    }
}

// isnan(value) -> null: clears validity bits (creating the bitmap content from
// all-ones if the column had none is the caller's job)
__global__ void __launch_bounds__(256) nan_to_null_kernel(const double* __restrict__ val, uint32_t* __restrict__ valid,
                                                          int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // grid covers n rounded up to 32
    const double x = i < n ? val[i] : 0.0;
    const unsigned notnan = __ballot_sync(0xffffffffu, x == x);
    if ((threadIdx.x & 31) == 0 && i < n) valid[i >> 5] &= notnan;
}

// ------------------------------------------------------------------ bucketize
__global__ void __launch_bounds__(256) bucketize_kernel(const double* __restrict__ val, const uint32_t* __restrict__ valid,
                                                        int64_t n, const double* __restrict__ splits, int32_t nsplits,
                                                        int32_t* __restrict__ out, unsigned long long* out_of_range) {
    extern __shared__ double s_splits[];
    for (int i = threadIdx.x; i < nsplits; i += blockDim.x) s_splits[i] = splits[i];
    __syncthreads();
    const int nb = nsplits - 1;
    unsigned long long bad = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const bool ok = !valid || ((__ldg(valid + (i >> 5)) >> (i & 31)) & 1u);
        const double x = val[i];
        int32_t b;
        if (!ok || x != x) {
            b = nb;  // handleInvalid="keep": the extra bucket
        } else if (x < s_splits[0] || x > s_splits[nb]) {
            b = -1;
            ++bad;
        } else if (x == s_splits[nb]) {
            b = nb - 1;
        } else {
            int lo = 0, hi = nb;  // largest lo with splits[lo] <= x
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (s_splits[mid] <= x) lo = mid; else hi = mid;
            }
            b = lo;
        }
        out[i] = b;
    }
    if (bad) atomicAdd(out_of_range, bad);
}

}  // namespace etl

using etl::Handle;

static inline int64_t capped_blocks(const Handle* h, int64_t want, int per_sm) {
    const int64_t cap = (int64_t)h->sm_count * per_sm;
    return want < 1 ? 1 : (want > cap ? cap : want);
}

extern "C" int etl_code_histogram(etl_handle* hh, const int32_t* codes, const uint32_t* valid, int64_t n, int32_t bins,
                                  int64_t* d_counts, int64_t* d_out_of_range, void* stream_) {
    ETL_REQUIRE(hh && d_counts && d_out_of_range, "code_histogram: NULL argument");
    ETL_REQUIRE(n >= 0 && bins >= 1, "code_histogram: n=%lld bins=%d", (long long)n, bins);
    ETL_REQUIRE(n == 0 || (codes && etl::aligned16(codes)), "code_histogram: codes NULL or misaligned");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_counts, 0, sizeof(int64_t) * bins, stream));
    ETL_CUDA(cudaMemsetAsync(d_out_of_range, 0, sizeof(int64_t), stream));
    if (n == 0) return ETL_OK;
    const int64_t tiles = (n + etl::HIST_THREADS * 4 - 1) / (etl::HIST_THREADS * 4);
    const size_t smem = bins <= etl::HIST_MAX_SMEM_BINS ? sizeof(uint32_t) * bins : 0;
    etl::code_histogram_kernel<<<(unsigned)capped_blocks(h, tiles, 8), etl::HIST_THREADS, smem, stream>>>(
        codes, valid, n, bins, reinterpret_cast<unsigned long long*>(d_counts),
        reinterpret_cast<unsigned long long*>(d_out_of_range));
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

static int fill_assemble(etl::AssembleParams* p, const int32_t* index, const double* const num[3], const double* shift,
                         const double* scale, int64_t n, int32_t num_labels, int32_t repeats) {
    ETL_REQUIRE(num, "assemble: num is NULL");
    ETL_REQUIRE(n >= 0 && num_labels >= 0 && repeats >= 1, "assemble: n=%lld num_labels=%d repeats=%d", (long long)n,
                num_labels, repeats);
    ETL_REQUIRE((int64_t)repeats * num_labels + 3 < (1 << 24), "assemble: feature dimension too large");
    p->index = index;
    for (int c = 0; c < 3; ++c) {
        ETL_REQUIRE(n == 0 || num[c], "assemble: num[%d] is NULL", c);
        p->num[c] = num[c];
        p->shift[c] = shift ? shift[c] : 0.0;
        p->scale[c] = scale ? scale[c] : 1.0;
    }
    p->n = n;
    p->num_labels = num_labels;
    p->repeats = repeats;
    return ETL_OK;
}

extern "C" int etl_assemble_features_dense(etl_handle* hh, const int32_t* index, const double* const num[3],
                                           const double* shift, const double* scale, int64_t n, int32_t num_labels,
                                           int32_t repeats, int32_t out_dtype, void* out, void* stream_) {
    ETL_REQUIRE(hh, "assemble_dense: NULL handle");
    etl::AssembleParams p{};
    int rc = fill_assemble(&p, index, num, shift, scale, n, num_labels, repeats);
    if (rc != ETL_OK) return rc;
    ETL_REQUIRE(out_dtype == ETL_F64 || out_dtype == ETL_F32, "assemble_dense: out_dtype=%d", out_dtype);
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(index && out && etl::aligned16(out), "assemble_dense: index/out NULL or out misaligned");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    p.out = out;
    const int64_t blocks = capped_blocks(h, (n + 7) / 8, 8);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool narrow = (int64_t)repeats * num_labels + 3 <= 32 * etl::ASM_MAXJ && num_labels > 0;
    // fp32 rows (612 bytes at D = 153) are never 16-byte aligned: they leave through shared memory and
    // bulk async stores (measured 1.58 -> 1.49 ms per 10 M rows; chunks of 8 / 16 / 64 / 128 rows are
    // all slower than 32).  fp64 rows keep the direct warp-per-row stores: 2.15 ms against 2.6 ms.
    static const int tma = [] { const char* e = getenv("ETL_ASM_TMA"); return e ? atoi(e) : 1; }();
    if (tma && narrow && out_dtype == ETL_F32) {
        const int D = repeats * num_labels + 3;
        const size_t smem = 2 * (size_t)etl::ASM_CHUNK * D * sizeof(float);
        const int64_t chunks = (n + etl::ASM_CHUNK - 1) / etl::ASM_CHUNK;
        int per_sm = (int)((200u << 10) / (smem + 3072));
        per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
        ETL_CUDA(cudaFuncSetAttribute(etl::assemble_dense_tma_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        etl::assemble_dense_tma_kernel<float><<<(unsigned)capped_blocks(h, chunks, per_sm), etl::ASM_THREADS, smem, stream>>>(p);
    } else if (out_dtype == ETL_F64) {
        if (narrow) etl::assemble_dense_kernel<double><<<(unsigned)blocks, etl::ASM_THREADS, 0, stream>>>(p);
        else etl::assemble_dense_wide_kernel<double><<<(unsigned)blocks, etl::ASM_THREADS, 0, stream>>>(p);
    } else {
        if (narrow) etl::assemble_dense_kernel<float><<<(unsigned)blocks, etl::ASM_THREADS, 0, stream>>>(p);
        else etl::assemble_dense_wide_kernel<float><<<(unsigned)blocks, etl::ASM_THREADS, 0, stream>>>(p);
    }
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_assemble_features_sparse(etl_handle* hh, const int32_t* index, const double* const num[3],
                                            const double* shift, const double* scale, int64_t n, int32_t num_labels,
                                            int32_t repeats, int32_t* out_index, double* out_value, void* stream_) {
    ETL_REQUIRE(hh, "assemble_sparse: NULL handle");
    etl::AssembleParams p{};
    int rc = fill_assemble(&p, index, num, shift, scale, n, num_labels, repeats);
    if (rc != ETL_OK) return rc;
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(index && out_index && out_value, "assemble_sparse: NULL argument");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    const int64_t total = n * (repeats + 3);
    etl::assemble_sparse_kernel<<<(unsigned)capped_blocks(h, (total + 255) / 256, 16), 256, 0,
                                  static_cast<cudaStream_t>(stream_)>>>(p, out_index, out_value);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_pack_tf_features(etl_handle* hh, const int32_t* label, const uint32_t* label_valid,
                                    const double* const num[3], const uint32_t* const num_valid[3], const double* shift,
                                    const double* scale, const int32_t* remap, int32_t remap_size, int64_t n, float* X,
                                    int32_t* y, int64_t* d_out_rows, void* stream_) {
    ETL_REQUIRE(hh && num && num_valid && d_out_rows, "pack_tf: NULL argument");
    ETL_REQUIRE(n >= 0, "pack_tf: n < 0");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    if (n == 0) {
        ETL_CUDA(cudaMemsetAsync(d_out_rows, 0, sizeof(int64_t), stream));
        return ETL_OK;
    }
    ETL_REQUIRE(label && X && y, "pack_tf: label/X/y is NULL");
    etl::TfPackParams p{};
    p.label = label;
    p.valid[0] = label_valid;
    for (int c = 0; c < 3; ++c) {
        ETL_REQUIRE(num[c], "pack_tf: num[%d] is NULL", c);
        p.num[c] = num[c];
        p.valid[1 + c] = num_valid[c];
        p.shift[c] = shift ? shift[c] : 0.0;
        p.scale[c] = scale ? scale[c] : 1.0;
    }
    p.remap = remap; p.remap_size = remap_size; p.n = n; p.X = X; p.y = y; p.out_rows = d_out_rows;
    const int64_t tiles = (n + etl::TF_TILE - 1) / etl::TF_TILE;
    int rc = etl::arena_reserve(h, (size_t)(tiles + 1) * sizeof(int64_t));
    if (rc != ETL_OK) return rc;
    p.tile_base = static_cast<int64_t*>(h->scratch);
    etl::tf_tile_count_kernel<<<(unsigned)capped_blocks(h, (tiles + 7) / 8, 8), 256, 0, stream>>>(p);
    etl::tf_tile_scan_kernel<<<1, 1024, 0, stream>>>(p.tile_base, tiles, d_out_rows);
    etl::tf_pack_kernel<<<(unsigned)capped_blocks(h, tiles, 8), 256, 0, stream>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_nan_to_null(etl_handle* hh, const double* val, uint32_t* valid, int64_t n, void* stream_) {
    ETL_REQUIRE(hh && (n == 0 || (val && valid)), "nan_to_null: NULL argument");
    ETL_REQUIRE(n >= 0, "nan_to_null: n < 0");
    if (n == 0) return ETL_OK;
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    const int64_t padded = (n + 31) & ~(int64_t)31;
    etl::nan_to_null_kernel<<<(unsigned)((padded + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(val, valid, n);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_bucketize(etl_handle* hh, const double* val, const uint32_t* valid, int64_t n, const double* d_splits,
                             int32_t nsplits, int32_t* out, int64_t* d_out_of_range, void* stream_) {
    ETL_REQUIRE(hh && d_splits && d_out_of_range, "bucketize: NULL argument");
    ETL_REQUIRE(n >= 0 && nsplits >= 3 && nsplits <= 4096, "bucketize: n=%lld nsplits=%d (need 3..4096 splits)",
                (long long)n, nsplits);
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_out_of_range, 0, sizeof(int64_t), stream));
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(val && out, "bucketize: val/out is NULL");
    etl::bucketize_kernel<<<(unsigned)capped_blocks(h, (n + 255) / 256, 16), 256, sizeof(double) * nsplits, stream>>>(
        val, valid, n, d_splits, nsplits, out, reinterpret_cast<unsigned long long*>(d_out_of_range));
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
