// Counter-based synthetic columns (bench / test infrastructure).
// bits(seed, stream, row) is the row-th output of the splitmix64 sequence keyed
// by (seed, stream), so any slice of any column can be produced independently
// on the host (pyspark_tf_gke_b200/synth.py, NumPy) and on the device.
#include "common.cuh"

namespace etl {

__device__ __forceinline__ uint64_t synth_bits(uint64_t key, uint64_t row) {
    return mix64(key + (row + 1) * 0x9E3779B97F4A7C15ull);
}
__device__ __forceinline__ double synth_unit(uint64_t bits) {
    return (double)(bits >> 11) * (1.0 / 9007199254740992.0);  // [0, 1), 53 bits
}

// One thread = 32 rows' validity word + those rows' values would serialise the
// stores; instead one thread = one row, the warp ballot builds the word.
__global__ void __launch_bounds__(256) synth_column_kernel(uint64_t key_val, uint64_t key_null, int kind,
                                                           uint64_t modulus, double null_rate, double nan_rate,
                                                           int64_t row0, int64_t n, void* out, uint32_t* valid) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // n is padded to a warp multiple
    const bool in = i < n;
    const uint64_t row = (uint64_t)(row0 + i);
    const uint64_t b = synth_bits(key_val, row);
    const double un = synth_unit(synth_bits(key_null, row));
    const bool is_valid = in && un >= null_rate;
    const bool is_nan = un >= null_rate && un < null_rate + nan_rate;
    if (in) {
        if (kind == 0) {
            double v = floor(__dmul_rn(synth_unit(b), 100001.0));
            static_cast<double*>(out)[i] = is_nan ? __longlong_as_double(0x7ff8000000000000ll) : v;
        } else if (kind == 1) {
            double v = __dmul_rn(synth_unit(b), 100000.0);
            static_cast<double*>(out)[i] = is_nan ? __longlong_as_double(0x7ff8000000000000ll) : v;
        } else if (kind == 2) {
            static_cast<int32_t*>(out)[i] = (int32_t)(b % modulus);
        } else {
            static_cast<int64_t*>(out)[i] = (int64_t)(b % modulus);
        }
    }
    if (valid) {
        const unsigned word = __ballot_sync(0xffffffffu, is_valid);
        if ((threadIdx.x & 31) == 0 && (i < n)) valid[i >> 5] = word;
    }
}

}  // namespace etl

extern "C" int etl_synth_column(etl_handle* hh, uint64_t seed, uint32_t stream_id, int32_t kind, uint64_t modulus,
                                double null_rate, double nan_rate, int64_t row0, int64_t n, void* out,
                                uint32_t* valid, void* stream_) {
    ETL_REQUIRE(hh && out, "synth_column: NULL handle or output");
    ETL_REQUIRE(kind >= 0 && kind <= 3, "synth_column: kind=%d", kind);
    ETL_REQUIRE(kind < 2 || modulus > 0, "synth_column: modulus must be > 0 for key columns");
    ETL_REQUIRE(n >= 0 && row0 >= 0 && (row0 & 31) == 0, "synth_column: n>=0, row0>=0 and row0 %% 32 == 0 required");
    if (n == 0) return ETL_OK;
    etl::Handle* h = reinterpret_cast<etl::Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    const uint64_t base = seed * 0x9E3779B97F4A7C15ull;
    const uint64_t key_val = base ^ ((uint64_t)(2 * stream_id) * 0xBF58476D1CE4E5B9ull);
    const uint64_t key_null = base ^ ((uint64_t)(2 * stream_id + 1) * 0xBF58476D1CE4E5B9ull);
    const int64_t blocks = (n + 255) / 256;
    etl::synth_column_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        key_val, key_null, kind, modulus, null_rate, nan_rate, row0, n, out, valid);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
