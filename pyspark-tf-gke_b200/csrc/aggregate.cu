// Scalar aggregates of the ETL stage: null count (k_means.py:23) and the
// (sum, count) partial state of Spark's Average (k_means.py:45-48).
//
// Both are single-pass streaming reductions.  The fp64 sum uses a reduction
// tree whose shape depends only on n: a fixed grid of SUM_GRID CTAs walks the
// 1024-row tiles round-robin, each thread keeps one running sum per column,
// then warp shuffles, a block tree, and the last CTA to finish folds the
// SUM_GRID partials in index order.  No floating-point atomics, so results are
// bit-reproducible, and exact whenever every partial sum is representable
// (integer-valued data below 2^53, the reference's default input).
//
// etl_column_moments (StandardScaler.fit, north-star item 5) runs the same kernel
// twice: pass 1 leaves count and mean = sum / count on the device, pass 2
// (CENTERED) sums (x - mean)^2 - the textbook two-pass variance, as accurate as a
// Welford / Chan merge and with the same fixed-shape tree; nothing visits the host
// in between.
#include "common.cuh"

namespace etl {

constexpr int SUM_THREADS = 256;
constexpr int SUM_TILE = SUM_THREADS * 4;
constexpr int SUM_GRID = 148 * 4;

struct SumParams {
    const void* val[3];
    const uint32_t* val_valid[3];
    int32_t val_dtype[3];
    const uint32_t* row_valid;
    int64_t n;
    double* partial_sum;     // [3][SUM_GRID]
    long long* partial_cnt;  // [3][SUM_GRID]
    double* out_sum;
    long long* out_cnt;
    double* out_mean;        // optional: sum / count (NaN for an empty column)
    const double* center;    // CENTERED: accumulate (x - center[c])^2 instead of x
    unsigned int* ticket;
};

__device__ __forceinline__ unsigned nibble_at(const uint32_t* bitmap, int64_t row, int64_t n) {
    unsigned nib = 0xF;
    if (bitmap) nib = (__ldg(bitmap + (row >> 5)) >> (row & 31)) & 0xF;
    const int64_t left = n - row;
    if (left < 4) nib &= left <= 0 ? 0u : ((1u << left) - 1);
    return nib;
}

template <int NC, bool CENTERED = false>
__global__ void __launch_bounds__(SUM_THREADS) column_sum_count_kernel(const SumParams p) {
    __shared__ double s_sum[NC][SUM_THREADS / 32];
    __shared__ long long s_cnt[NC][SUM_THREADS / 32];
    __shared__ bool s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double acc[NC], ctr[NC];
    long long cnt[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) { acc[c] = 0.0; cnt[c] = 0; ctr[c] = CENTERED ? p.center[c] : 0.0; }

    const int64_t tiles = (p.n + SUM_TILE - 1) / SUM_TILE;
    for (int64_t t = blockIdx.x; t < tiles; t += SUM_GRID) {
        const int64_t row = t * SUM_TILE + (int64_t)tid * 4;
        const bool full = (t + 1) * SUM_TILE <= p.n;
        const unsigned rnib = nibble_at(p.row_valid, row, p.n);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            double v[4];
            const int dt = p.val_dtype[c];
            if (full) {
                if (dt == ETL_F64) {
                    const double2 a = ld_stream_d2(static_cast<const double*>(p.val[c]) + row);
                    const double2 b = ld_stream_d2(static_cast<const double*>(p.val[c]) + row + 2);
                    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
                } else {
                    const int4 q = ld_stream_v4(static_cast<const int32_t*>(p.val[c]) + row);
                    if (dt == ETL_F32) {
                        v[0] = (double)__int_as_float(q.x); v[1] = (double)__int_as_float(q.y);
                        v[2] = (double)__int_as_float(q.z); v[3] = (double)__int_as_float(q.w);
                    } else {
                        v[0] = (double)q.x; v[1] = (double)q.y; v[2] = (double)q.z; v[3] = (double)q.w;
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int64_t r = row + j;
                    v[j] = 0.0;
                    if (r < p.n) {
                        if (dt == ETL_F64) v[j] = static_cast<const double*>(p.val[c])[r];
                        else if (dt == ETL_F32) v[j] = (double)static_cast<const float*>(p.val[c])[r];
                        else v[j] = (double)static_cast<const int32_t*>(p.val[c])[r];
                    }
                }
            }
            const unsigned nib = rnib & nibble_at(p.val_valid[c], row, p.n);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool ok = ((nib >> j) & 1u) && (v[j] == v[j]);
                if (ok) {
                    if (CENTERED) {
                        const double d = __dsub_rn(v[j], ctr[c]);
                        acc[c] = __dadd_rn(acc[c], __dmul_rn(d, d));
                    } else {
                        acc[c] += v[j];
                    }
                    ++cnt[c];
                }
            }
        }
    }
    // warp tree, then block tree (fixed order)
#pragma unroll
    for (int c = 0; c < NC; ++c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            acc[c] += __shfl_down_sync(0xffffffffu, acc[c], o);
            cnt[c] += __shfl_down_sync(0xffffffffu, cnt[c], o);
        }
        if (lane == 0) { s_sum[c][warp] = acc[c]; s_cnt[c][warp] = cnt[c]; }
    }
    __syncthreads();
    if (tid == 0) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            double s = 0.0;
            long long k = 0;
#pragma unroll
            for (int w = 0; w < SUM_THREADS / 32; ++w) { s += s_sum[c][w]; k += s_cnt[c][w]; }
            p.partial_sum[c * SUM_GRID + blockIdx.x] = s;
            p.partial_cnt[c * SUM_GRID + blockIdx.x] = k;
        }
        __threadfence();
        s_last = atomicAdd(p.ticket, 1u) == SUM_GRID - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // last CTA: warp c folds column c's partials in index order (strided per
    // lane, then a shuffle tree) - the same order every run
    if (warp < NC) {
        const int c = warp;
        double s = 0.0;
        long long k = 0;
        for (int i = lane; i < SUM_GRID; i += 32) {
            s += __ldcg(p.partial_sum + c * SUM_GRID + i);
            k += __ldcg(p.partial_cnt + c * SUM_GRID + i);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s += __shfl_down_sync(0xffffffffu, s, o);
            k += __shfl_down_sync(0xffffffffu, k, o);
        }
        if (lane == 0) {
            p.out_sum[c] = s;
            p.out_cnt[c] = k;
            if (p.out_mean) p.out_mean[c] = k ? s / (double)k : __longlong_as_double(0x7ff8000000000000ll);
        }
    }
    if (tid == 0) *p.ticket = 0;  // ready for the next launch
}

__global__ void __launch_bounds__(256) null_count_kernel(const uint32_t* __restrict__ valid, int64_t n,
                                                         unsigned long long* out) {
    const int64_t words = (n + 31) >> 5;
    unsigned long long zeros = 0;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words;
         w += (int64_t)gridDim.x * blockDim.x) {
        uint32_t bits = ~__ldg(valid + w);
        if (w == words - 1 && (n & 31)) bits &= (1u << (n & 31)) - 1;
        zeros += __popc(bits);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) zeros += __shfl_down_sync(0xffffffffu, zeros, o);
    if ((threadIdx.x & 31) == 0 && zeros) atomicAdd(out, zeros);
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_null_count(etl_handle* hh, const uint32_t* valid, int64_t n, int64_t* d_out, void* stream_) {
    ETL_REQUIRE(hh && d_out, "null_count: NULL handle or output");
    ETL_REQUIRE(n >= 0, "null_count: n < 0");
    ETL_REQUIRE(etl::aligned4(valid), "null_count: bitmap misaligned");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_out, 0, sizeof(int64_t), stream));
    if (!valid || n == 0) return ETL_OK;
    const int64_t words = (n + 31) >> 5;
    int64_t blocks = (words + 256 * 8 - 1) / (256 * 8);
    if (blocks > h->sm_count * 8) blocks = h->sm_count * 8;
    etl::null_count_kernel<<<(unsigned)blocks, 256, 0, stream>>>(valid, n, reinterpret_cast<unsigned long long*>(d_out));
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

static int launch_sum(etl_handle* hh, const char* who, bool centered, int32_t ncols, const void* const val[],
                      const uint32_t* const val_valid[], const int32_t val_dtype[], const uint32_t* row_valid, int64_t n,
                      double* d_sum, int64_t* d_count, double* d_mean, const double* d_center, void* stream_) {
    ETL_REQUIRE(hh && val && val_valid && val_dtype && d_sum && d_count, "%s: NULL argument", who);
    ETL_REQUIRE(ncols >= 1 && ncols <= 3, "%s: ncols=%d not in 1..3", who, ncols);
    ETL_REQUIRE(n >= 0, "%s: n < 0", who);
    ETL_REQUIRE(etl::aligned4(row_valid), "%s: row_valid misaligned", who);
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    etl::SumParams p{};
    for (int c = 0; c < ncols; ++c) {
        ETL_REQUIRE(val[c] != nullptr || n == 0, "%s: val[%d] is NULL", who, c);
        ETL_REQUIRE(val_dtype[c] == ETL_F64 || val_dtype[c] == ETL_F32 || val_dtype[c] == ETL_I32,
                    "%s: val_dtype[%d]=%d unsupported", who, c, val_dtype[c]);
        ETL_REQUIRE(etl::aligned16(val[c]) && etl::aligned4(val_valid[c]), "%s: val[%d] misaligned", who, c);
        p.val[c] = val[c];
        p.val_valid[c] = val_valid[c];
        p.val_dtype[c] = val_dtype[c];
    }
    ETL_CUDA(cudaSetDevice(h->device));
    const size_t need = 3 * etl::SUM_GRID * (sizeof(double) + sizeof(long long));
    int rc = etl::arena_reserve(h, need);
    if (rc != ETL_OK) return rc;
    p.row_valid = row_valid;
    p.n = n;
    p.partial_sum = static_cast<double*>(h->scratch);
    p.partial_cnt = reinterpret_cast<long long*>(p.partial_sum + 3 * etl::SUM_GRID);
    p.out_sum = d_sum;
    p.out_cnt = reinterpret_cast<long long*>(d_count);
    p.out_mean = d_mean;
    p.center = d_center;
    p.ticket = h->counters + 0;
    if (centered) {
        switch (ncols) {
            case 1: etl::column_sum_count_kernel<1, true><<<etl::SUM_GRID, etl::SUM_THREADS, 0, stream>>>(p); break;
            case 2: etl::column_sum_count_kernel<2, true><<<etl::SUM_GRID, etl::SUM_THREADS, 0, stream>>>(p); break;
            default: etl::column_sum_count_kernel<3, true><<<etl::SUM_GRID, etl::SUM_THREADS, 0, stream>>>(p); break;
        }
    } else {
        switch (ncols) {
            case 1: etl::column_sum_count_kernel<1><<<etl::SUM_GRID, etl::SUM_THREADS, 0, stream>>>(p); break;
            case 2: etl::column_sum_count_kernel<2><<<etl::SUM_GRID, etl::SUM_THREADS, 0, stream>>>(p); break;
            default: etl::column_sum_count_kernel<3><<<etl::SUM_GRID, etl::SUM_THREADS, 0, stream>>>(p); break;
        }
    }
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_column_sum_count(etl_handle* hh, int32_t ncols, const void* const val[],
                                    const uint32_t* const val_valid[], const int32_t val_dtype[],
                                    const uint32_t* row_valid, int64_t n, double* d_sum, int64_t* d_count,
                                    void* stream_) {
    return launch_sum(hh, "column_sum_count", false, ncols, val, val_valid, val_dtype, row_valid, n, d_sum, d_count, nullptr,
                      nullptr, stream_);
}

extern "C" int etl_column_moments(etl_handle* hh, int32_t ncols, const void* const val[], const uint32_t* const val_valid[],
                                  const int32_t val_dtype[], const uint32_t* row_valid, int64_t n, int64_t* d_count,
                                  double* d_mean, double* d_m2, void* stream_) {
    ETL_REQUIRE(d_mean && d_m2, "column_moments: NULL argument");
    // pass 1: count and mean (the sums land in d_m2 and are overwritten by pass 2)
    int rc = launch_sum(hh, "column_moments", false, ncols, val, val_valid, val_dtype, row_valid, n, d_m2, d_count, d_mean,
                        nullptr, stream_);
    if (rc != ETL_OK) return rc;
    // pass 2: sum of squared deviations from the mean pass 1 left on the device
    return launch_sum(hh, "column_moments", true, ncols, val, val_valid, val_dtype, row_valid, n, d_m2, d_count, nullptr,
                      d_mean, stream_);
}
