// A caller of the C ABI with no Python and no torch in the process: what a JVM / Go / C++ host
// would do through JNI / cgo / a plain link.  It builds a small table in host
// memory, runs the fused filter + impute projection (etl_project_filter_impute, k_means.py:27-28,
// 49-51) and the (sum, count) pass (etl_column_sum_count, k_means.py:45-48) on the GPU and checks
// both against straightforward host loops.  Exit code 0 = identical.
//
//   g++ -O2 -I include -I /usr/local/cuda/include tests/native/cabi_caller.cpp \
//       -L pyspark-tf-gke_b200 -letl_b200 -L /usr/local/cuda/lib64 -lcudart -o cabi_caller
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "etl_b200.h"

#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_)); return 2; } \
    } while (0)
#define ETL(call)                                                                        \
    do {                                                                                 \
        int rc_ = (call);                                                                \
        if (rc_ != ETL_OK) { std::fprintf(stderr, "%s -> %d: %s\n", #call, rc_, etl_last_error()); return 3; } \
    } while (0)

static uint64_t mix(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

int main(int argc, char** argv) {
    const int64_t n = argc > 1 ? std::atoll(argv[1]) : 1000003;
    const int64_t words = (n + 31) / 32;
    std::vector<int32_t> code(n);
    std::vector<uint32_t> key_valid(words + 4, 0), val_valid[3];
    std::vector<double> val[3];
    for (int c = 0; c < 3; ++c) { val[c].resize(n); val_valid[c].assign(words + 4, 0); }
    for (int64_t i = 0; i < n; ++i) {
        const uint64_t r = mix((uint64_t)i);
        code[i] = (int32_t)(r % 30);
        if (r % 100 != 0) key_valid[i >> 5] |= 1u << (i & 31);            // 1 % null keys
        for (int c = 0; c < 3; ++c) {
            const uint64_t q = mix(r + 1 + c);
            val[c][i] = (q % 50 == 1) ? std::nan("") : (double)(q % 100001);
            if (q % 14 != 0) val_valid[c][i >> 5] |= 1u << (i & 31);       // ~7 % null cells
        }
    }
    // device copies (bitmaps need 16-byte alignment for the projection entry: cudaMalloc gives 256)
    int32_t *d_code, *d_out_code;
    uint32_t *d_kv, *d_vv[3];
    double *d_val[3], *d_out[3], *d_sum;
    int64_t *d_rows, *d_cnt;
    CK(cudaMalloc(&d_code, n * 4)); CK(cudaMalloc(&d_out_code, n * 4)); CK(cudaMalloc(&d_kv, (words + 4) * 4));
    CK(cudaMalloc(&d_rows, 8)); CK(cudaMalloc(&d_sum, 3 * 8)); CK(cudaMalloc(&d_cnt, 3 * 8));
    CK(cudaMemcpy(d_code, code.data(), n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_kv, key_valid.data(), (words + 4) * 4, cudaMemcpyHostToDevice));
    for (int c = 0; c < 3; ++c) {
        CK(cudaMalloc(&d_val[c], n * 8)); CK(cudaMalloc(&d_out[c], n * 8)); CK(cudaMalloc(&d_vv[c], (words + 4) * 4));
        CK(cudaMemcpy(d_val[c], val[c].data(), n * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_vv[c], val_valid[c].data(), (words + 4) * 4, cudaMemcpyHostToDevice));
    }
    etl_handle* h = nullptr;
    ETL(etl_create(0, &h));
    // a4: (sum, count) over non-null, non-NaN cells -> the three means
    const void* vals[3] = {d_val[0], d_val[1], d_val[2]};
    const uint32_t* vvs[3] = {d_vv[0], d_vv[1], d_vv[2]};
    const int32_t dts[3] = {ETL_F64, ETL_F64, ETL_F64};
    ETL(etl_column_sum_count(h, 3, vals, vvs, dts, nullptr, n, d_sum, d_cnt, nullptr));
    double sum[3];
    int64_t cnt[3];
    CK(cudaMemcpy(sum, d_sum, 24, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cnt, d_cnt, 24, cudaMemcpyDeviceToHost));
    double fill[3];
    int bad = 0;
    for (int c = 0; c < 3; ++c) {
        double s = 0.0;   // integer-valued cells: any summation order gives the same double
        int64_t k = 0;
        for (int64_t i = 0; i < n; ++i)
            if (((val_valid[c][i >> 5] >> (i & 31)) & 1u) && !std::isnan(val[c][i])) { s += val[c][i]; ++k; }
        if (s != sum[c] || k != cnt[c]) { std::fprintf(stderr, "column %d: sum/count differ\n", c); ++bad; }
        fill[c] = sum[c] / (double)cnt[c];
    }
    // a3 + a5: keep rows with a key, impute null / NaN cells with the mean, compact
    double* outs[3] = {d_out[0], d_out[1], d_out[2]};
    ETL(etl_project_filter_impute(h, d_code, d_kv, vals, vvs, dts, fill, n, d_out_code, outs, d_rows, nullptr));
    int64_t kept = 0;
    CK(cudaMemcpy(&kept, d_rows, 8, cudaMemcpyDeviceToHost));
    std::vector<int32_t> got_code(n);
    std::vector<double> got[3];
    CK(cudaMemcpy(got_code.data(), d_out_code, n * 4, cudaMemcpyDeviceToHost));
    for (int c = 0; c < 3; ++c) { got[c].resize(n); CK(cudaMemcpy(got[c].data(), d_out[c], n * 8, cudaMemcpyDeviceToHost)); }
    int64_t w = 0;
    for (int64_t i = 0; i < n; ++i) {
        if (!((key_valid[i >> 5] >> (i & 31)) & 1u)) continue;
        if (w < kept) {
            if (got_code[w] != code[i]) ++bad;
            for (int c = 0; c < 3; ++c) {
                const bool ok = ((val_valid[c][i >> 5] >> (i & 31)) & 1u) && !std::isnan(val[c][i]);
                if (got[c][w] != (ok ? val[c][i] : fill[c])) ++bad;
            }
        }
        ++w;
    }
    if (w != kept) { std::fprintf(stderr, "kept rows: %lld, expected %lld\n", (long long)kept, (long long)w); ++bad; }
    // the allocation-free entry, with no torch caching allocator in the way: after its set-up call the
    // driver's free-memory figure must be the same before and after a call, and the result the same rows
    ETL(etl_project_filter_impute_noalloc(d_code, d_kv, vals, vvs, dts, fill, n, d_out_code, outs, d_rows, nullptr));
    CK(cudaDeviceSynchronize());
    for (int rep = 0; rep < 2; ++rep) {
        size_t free0 = 0, free1 = 0, total = 0;
        CK(cudaMemGetInfo(&free0, &total));
        CK(cudaMemset(d_rows, 0xff, 8));
        ETL(etl_project_filter_impute_noalloc(d_code, d_kv, vals, vvs, dts, fill, n, d_out_code, outs, d_rows, nullptr));
        CK(cudaDeviceSynchronize());
        CK(cudaMemGetInfo(&free1, &total));
        int64_t kept2 = -1;
        CK(cudaMemcpy(&kept2, d_rows, 8, cudaMemcpyDeviceToHost));
        if (free0 != free1) { std::fprintf(stderr, "noalloc call changed free device memory: %zu -> %zu\n", free0, free1); ++bad; }
        if (kept2 != kept) { std::fprintf(stderr, "noalloc kept rows %lld, expected %lld\n", (long long)kept2, (long long)kept); ++bad; }
    }
    std::vector<int32_t> again(n);
    CK(cudaMemcpy(again.data(), d_out_code, n * 4, cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < kept; ++i) if (again[i] != got_code[i]) { ++bad; break; }
    ETL(etl_destroy(h));
    std::printf("cabi_caller: abi %d, %lld rows -> %lld kept, means %.6f %.6f %.6f, mismatches %d\n", etl_abi_version(),
                (long long)n, (long long)kept, fill[0], fill[1], fill[2], bad);
    return bad ? 1 : 0;
}
