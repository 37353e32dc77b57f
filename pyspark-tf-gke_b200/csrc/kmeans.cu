// KMeans on the assembled `features` (k_means.py:83-87: KMeans().setK(25)
// .setSeed(1).setMaxIter(1000).fit(dataset)) - SURVEY.md §8f rank 2, the ETL
// stage's in-process consumer.
//
// The feature rows are never materialised for this: row i is
//   [onehot(index_i)] * repeats ++ [num0_i, num1_i, num2_i]
// so its squared distance to a centre c needs only
//   |x|^2 - 2 * (T[c][index_i] + sum_j num_j * c[onehot + j]) + |c|^2
// with T[c][l] = sum_rep c[rep * L + l] (0 for the unknown label).  T, the three
// numeric coordinates and |c|^2 of every centre are prepared once per iteration
// (kmeans_prepare_kernel) and staged in shared memory, so an assignment pass reads
// 28 B per row and does ~6 flops per (row, centre).
//   kmeans_assign_kernel      nearest centre + squared distance per row (optionally
//                             min-combined with a previous cost: the k-means|| step)
//   kmeans_accumulate_kernel  per-cluster row count, label histogram and numeric
//                             sums -> the next centres (shared-memory partials,
//                             native 32-bit atomics for counts, fp64 adds for sums)
#include "common.cuh"

namespace etl {

constexpr int KM_THREADS = 256;

struct KmPrepared {          // device arrays derived from the centres [k][D]
    const double* onehot_sum; // [k][L + 1]: T[c][l]; column L = 0 (unknown label)
    const double* num;        // [k][3]
    const double* norm2;      // [k]
};

__global__ void __launch_bounds__(256) kmeans_prepare_kernel(const double* __restrict__ centers, int k, int L, int repeats,
                                                             double* __restrict__ onehot_sum, double* __restrict__ num,
                                                             double* __restrict__ norm2) {
    const int D = repeats * L + 3;
    const int c = blockIdx.x;
    const double* row = centers + (int64_t)c * D;
    __shared__ double s_part[256];
    double acc = 0.0;
    for (int d = threadIdx.x; d < D; d += blockDim.x) acc = __dadd_rn(acc, __dmul_rn(row[d], row[d]));
    s_part[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {  // fixed tree: reproducible
        if (threadIdx.x < o) s_part[threadIdx.x] += s_part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) norm2[c] = s_part[0];
    for (int l = threadIdx.x; l <= L; l += blockDim.x) {
        double t = 0.0;
        if (l < L)
            for (int r = 0; r < repeats; ++r) t += row[r * L + l];
        onehot_sum[(int64_t)c * (L + 1) + l] = t;
    }
    if (threadIdx.x < 3) num[c * 3 + threadIdx.x] = row[repeats * L + threadIdx.x];
}

struct KmAssignParams {
    const int32_t* index;
    const double* num[3];
    int64_t n;
    int32_t k, L, repeats;
    KmPrepared prep;
    const double* prev_cost;  // optional: cost[i] = min(prev_cost[i], new); the cluster is only updated when new wins
    int32_t* out_cluster;     // optional
    double* out_cost;         // optional
    int32_t cluster_offset;   // added to the winning centre's id (k-means|| appends candidate batches)
};

__global__ void __launch_bounds__(KM_THREADS) kmeans_assign_kernel(const KmAssignParams p) {
    extern __shared__ double km_smem[];
    const int L1 = p.L + 1;
    double* s_T = km_smem;                         // [k][L + 1]
    double* s_num = s_T + (size_t)p.k * L1;        // [k][3]
    double* s_norm = s_num + (size_t)p.k * 3;      // [k]
    for (int i = threadIdx.x; i < p.k * L1; i += KM_THREADS) s_T[i] = p.prep.onehot_sum[i];
    for (int i = threadIdx.x; i < p.k * 3; i += KM_THREADS) s_num[i] = p.prep.num[i];
    for (int i = threadIdx.x; i < p.k; i += KM_THREADS) s_norm[i] = p.prep.norm2[i];
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * KM_THREADS + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * KM_THREADS) {
        int32_t idx = p.index[i];
        const bool known = idx >= 0 && idx < p.L;
        if (!known) idx = p.L;
        const double a = p.num[0][i], b = p.num[1][i], c = p.num[2][i];
        // explicit _rn ops: no FMA contraction, so the CPU oracle (NumPy, same order) gets the same bits
        const double x2 = __dadd_rn(__dadd_rn(__dadd_rn(known ? (double)p.repeats : 0.0, __dmul_rn(a, a)), __dmul_rn(b, b)),
                                    __dmul_rn(c, c));
        double best = 1.7976931348623157e308;
        int best_c = 0;
        for (int j = 0; j < p.k; ++j) {
            const double dot = __dadd_rn(__dadd_rn(__dadd_rn(s_T[j * L1 + idx], __dmul_rn(a, s_num[j * 3])),
                                                   __dmul_rn(b, s_num[j * 3 + 1])), __dmul_rn(c, s_num[j * 3 + 2]));
            double d2 = __dadd_rn(__dadd_rn(x2, -__dmul_rn(2.0, dot)), s_norm[j]);
            d2 = d2 < 0.0 ? 0.0 : d2;  // cancellation can leave a tiny negative
            if (d2 < best) { best = d2; best_c = j; }   // ties keep the lower centre id, as Spark's findClosest does
        }
        bool wins = true;
        if (p.prev_cost) {
            const double old = p.prev_cost[i];
            wins = best < old;
            if (!wins) best = old;
        }
        if (p.out_cost) p.out_cost[i] = best;
        if (p.out_cluster && wins) p.out_cluster[i] = best_c + p.cluster_offset;
    }
}

struct KmAccumParams {
    const int32_t* index;
    const double* num[3];
    const int32_t* cluster;
    int64_t n;
    int32_t k, L;
    unsigned long long* count;  // [k]
    unsigned long long* hist;   // [k][L + 1]
    double* numsum;             // [k][3]
};

__global__ void __launch_bounds__(KM_THREADS) kmeans_accumulate_kernel(const KmAccumParams p) {
    extern __shared__ double km_smem[];
    const int L1 = p.L + 1;
    double* s_sum = km_smem;                                             // [k][3]
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(s_sum + (size_t)p.k * 3);  // [k]
    uint32_t* s_hist = s_cnt + p.k;                                      // [k][L + 1]
    for (int i = threadIdx.x; i < p.k * 3; i += KM_THREADS) s_sum[i] = 0.0;
    for (int i = threadIdx.x; i < p.k; i += KM_THREADS) s_cnt[i] = 0;
    for (int i = threadIdx.x; i < p.k * L1; i += KM_THREADS) s_hist[i] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * KM_THREADS + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * KM_THREADS) {
        const int c = p.cluster[i];
        if (c < 0 || c >= p.k) continue;
        int32_t idx = p.index[i];
        if (idx < 0 || idx >= p.L) idx = p.L;
        atomicAdd(s_cnt + c, 1u);
        atomicAdd(s_hist + c * L1 + idx, 1u);
#pragma unroll
        for (int j = 0; j < 3; ++j) atomicAdd(s_sum + c * 3 + j, p.num[j][i]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < p.k; i += KM_THREADS)
        if (s_cnt[i]) atomicAdd(p.count + i, (unsigned long long)s_cnt[i]);
    for (int i = threadIdx.x; i < p.k * L1; i += KM_THREADS)
        if (s_hist[i]) atomicAdd(p.hist + i, (unsigned long long)s_hist[i]);
    for (int i = threadIdx.x; i < p.k * 3; i += KM_THREADS)
        if (s_sum[i] != 0.0) atomicAdd(p.numsum + i, s_sum[i]);
}

// k <= 32: the numeric sums without a single floating-point atomic.  fp64 shared atomics are CAS
// loops on sm_100a and 25 clusters x 3 sums are 75 hot addresses: the kernel above spends its time
// retrying.  Here lane c of every warp OWNS cluster c: per batch of 32 rows the warp stages the three
// numerics in shared memory, one ballot per cluster tells lane c which rows are its own, and it adds
// them into registers in row order (deterministic within a launch shape).  Registers live across the
// whole grid-stride loop; the CTA folds its warps at the end and issues one global add per (cluster,
// coordinate).  The label histogram keeps its native 32-bit shared atomics; the row count of a
// cluster is its histogram row's sum.
__global__ void __launch_bounds__(KM_THREADS) kmeans_accumulate_owner_kernel(const KmAccumParams p) {
    extern __shared__ double km_smem[];
    const int L1 = p.L + 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* s_stage = km_smem;                                               // [warps][3][32]
    double* s_fold = s_stage + (KM_THREADS / 32) * 96;                       // [warps][32][3]
    uint32_t* s_hist = reinterpret_cast<uint32_t*>(s_fold + (KM_THREADS / 32) * 96);   // [k][L + 1]
    for (int i = threadIdx.x; i < p.k * L1; i += KM_THREADS) s_hist[i] = 0;
    __syncthreads();
    double* mine = s_stage + warp * 96;
    double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0;
    const int64_t batches = (p.n + 31) / 32;
    for (int64_t b = (int64_t)blockIdx.x * (KM_THREADS / 32) + warp; b < batches; b += (int64_t)gridDim.x * (KM_THREADS / 32)) {
        const int64_t i = b * 32 + lane;
        int c = -1;
        if (i < p.n) {
            c = p.cluster[i];
            if (c < 0 || c >= p.k) c = -1;
        }
        if (c >= 0) {
            int32_t idx = p.index[i];
            if (idx < 0 || idx >= p.L) idx = p.L;
            atomicAdd(s_hist + c * L1 + idx, 1u);
            mine[lane] = p.num[0][i];
            mine[32 + lane] = p.num[1][i];
            mine[64 + lane] = p.num[2][i];
        }
        unsigned own = 0;   // rows of this batch that belong to cluster `lane`
        for (int q = 0; q < p.k; ++q) {
            const unsigned m = __ballot_sync(0xffffffffu, c == q);
            if (lane == q) own = m;
        }
        __syncwarp();
        while (own) {
            const int r = __ffs(own) - 1;
            own &= own - 1;
            acc0 += mine[r];
            acc1 += mine[32 + r];
            acc2 += mine[64 + r];
        }
        __syncwarp();
    }
    double* fold = s_fold + (warp * 32 + lane) * 3;
    fold[0] = acc0; fold[1] = acc1; fold[2] = acc2;
    __syncthreads();
    if (threadIdx.x < p.k * 3) {
        const int c = threadIdx.x / 3, j = threadIdx.x % 3;
        double t = 0.0;
        for (int w = 0; w < KM_THREADS / 32; ++w) t += s_fold[(w * 32 + c) * 3 + j];   // fixed order
        if (t != 0.0) atomicAdd(p.numsum + c * 3 + j, t);
    }
    for (int c = threadIdx.x; c < p.k; c += KM_THREADS) {
        unsigned long long rows = 0;
        for (int l = 0; l < L1; ++l) rows += s_hist[c * L1 + l];
        if (rows) atomicAdd(p.count + c, rows);
    }
    for (int i = threadIdx.x; i < p.k * L1; i += KM_THREADS)
        if (s_hist[i]) atomicAdd(p.hist + i, (unsigned long long)s_hist[i]);
}

// the centre update of one Lloyd iteration, on the device: centre c <- mean of its rows (unchanged when
// the cluster is empty); the divisions are the ones the host did before, so the centres keep their bits
__global__ void __launch_bounds__(256) kmeans_update_kernel(const unsigned long long* __restrict__ count,
                                                            const unsigned long long* __restrict__ hist,
                                                            const double* __restrict__ numsum, int L, int repeats,
                                                            double* __restrict__ centers) {
    const int c = blockIdx.x;
    const unsigned long long rows = count[c];
    if (rows == 0) return;
    const int D = repeats * L + 3;
    const double denom = (double)rows;
    double* row = centers + (int64_t)c * D;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        const int j = d - repeats * L;
        row[d] = j < 0 ? (double)hist[(int64_t)c * (L + 1) + d % L] / denom : numsum[c * 3 + j] / denom;
    }
}

// ---------------------------------------------------------------- silhouette
// ClusteringEvaluator (metric "silhouette", squared Euclidean distance) as Spark computes it, in
// O(n k): per cluster G the row count N_G, Psi_G = sum of |y|^2 and the vector sum Y_G; the mean
// squared distance from x to the rows of G is |x|^2 + Psi_G / N_G - 2 x.Y_G / N_G
// (spark_workload_to_cloud_k8s.py:139-144 -> org.apache.spark.ml.evaluation.SquaredEuclideanSilhouette).
// Pass 1 writes |x|^2 per row (its per-cluster sums come from kmeans_accumulate); pass 2 the
// coefficient of every row from the per-cluster tables.  Same explicit _rn arithmetic as the assign
// kernel, so the NumPy restatement gets the same bits.
__global__ void __launch_bounds__(KM_THREADS) kmeans_row_norms_kernel(const int32_t* __restrict__ index, const double* num0,
                                                                      const double* num1, const double* num2, int64_t n, int L,
                                                                      int repeats, double* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * KM_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * KM_THREADS) {
        const int32_t idx = index[i];
        const bool known = idx >= 0 && idx < L;
        const double a = num0[i], b = num1[i], c = num2[i];
        out[i] = __dadd_rn(__dadd_rn(__dadd_rn(known ? (double)repeats : 0.0, __dmul_rn(a, a)), __dmul_rn(b, b)), __dmul_rn(c, c));
    }
}

struct KmSilhouetteParams {
    const int32_t* index;
    const double* num[3];
    const int32_t* cluster;
    int64_t n;
    int32_t k, L, repeats;
    const double* label_sum;   // [k][L + 1]: repeats * (rows of the cluster with that label); column L = 0
    const double* num_sum;     // [k][3]
    const double* psi;         // [k]: sum of |y|^2 over the cluster
    const double* count;       // [k]: rows (as double); 0 = the cluster does not exist
    double* out;               // [n]
};

__global__ void __launch_bounds__(KM_THREADS) kmeans_silhouette_kernel(const KmSilhouetteParams p) {
    extern __shared__ double km_smem[];
    const int L1 = p.L + 1;
    double* s_T = km_smem;                         // [k][L + 1]
    double* s_num = s_T + (size_t)p.k * L1;        // [k][3]
    double* s_psi = s_num + (size_t)p.k * 3;       // [k]
    double* s_cnt = s_psi + p.k;                   // [k]
    for (int i = threadIdx.x; i < p.k * L1; i += KM_THREADS) s_T[i] = p.label_sum[i];
    for (int i = threadIdx.x; i < p.k * 3; i += KM_THREADS) s_num[i] = p.num_sum[i];
    for (int i = threadIdx.x; i < p.k; i += KM_THREADS) { s_psi[i] = p.psi[i]; s_cnt[i] = p.count[i]; }
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * KM_THREADS + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * KM_THREADS) {
        int32_t idx = p.index[i];
        const bool known = idx >= 0 && idx < p.L;
        if (!known) idx = p.L;
        const int own = p.cluster[i];
        const double a = p.num[0][i], b = p.num[1][i], c = p.num[2][i];
        const double x2 = __dadd_rn(__dadd_rn(__dadd_rn(known ? (double)p.repeats : 0.0, __dmul_rn(a, a)), __dmul_rn(b, b)),
                                    __dmul_rn(c, c));
        double neighbour = 1.7976931348623157e308, mine = 0.0;
        for (int j = 0; j < p.k; ++j) {
            const double nj = s_cnt[j];
            if (nj == 0.0) continue;
            const double dot = __dadd_rn(__dadd_rn(__dadd_rn(s_T[j * L1 + idx], __dmul_rn(a, s_num[j * 3])),
                                                   __dmul_rn(b, s_num[j * 3 + 1])), __dmul_rn(c, s_num[j * 3 + 2]));
            // pointDotSquaredNorm + squaredNormSum / numOfPoints - 2 * dot / numOfPoints
            const double d = __dadd_rn(__dadd_rn(x2, __ddiv_rn(s_psi[j], nj)), -__ddiv_rn(__dmul_rn(2.0, dot), nj));
            if (j == own) mine = d;
            else if (d < neighbour) neighbour = d;
        }
        double sil = 0.0;
        if (own >= 0 && own < p.k && s_cnt[own] > 1.0 && neighbour < 1.7976931348623157e308) {
            const double n_own = s_cnt[own];
            const double cur = __ddiv_rn(__dmul_rn(mine, n_own), __dadd_rn(n_own, -1.0));
            if (cur < neighbour) sil = __dadd_rn(1.0, -__ddiv_rn(cur, neighbour));
            else if (cur > neighbour) sil = __dadd_rn(__ddiv_rn(neighbour, cur), -1.0);
        }
        p.out[i] = sil;
    }
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_kmeans_prepare(etl_handle* hh, const double* centers, int32_t k, int32_t num_labels, int32_t repeats,
                                  double* d_onehot_sum, double* d_num, double* d_norm2, void* stream_) {
    ETL_REQUIRE(hh && centers && d_onehot_sum && d_num && d_norm2, "kmeans_prepare: NULL argument");
    ETL_REQUIRE(k >= 1 && num_labels >= 0 && repeats >= 1, "kmeans_prepare: k=%d num_labels=%d repeats=%d", k, num_labels, repeats);
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::kmeans_prepare_kernel<<<(unsigned)k, 256, 0, static_cast<cudaStream_t>(stream_)>>>(centers, k, num_labels, repeats,
                                                                                         d_onehot_sum, d_num, d_norm2);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

static size_t km_assign_smem(int k, int L) { return sizeof(double) * ((size_t)k * (L + 1) + (size_t)k * 4); }
static size_t km_accum_smem(int k, int L) { return sizeof(double) * (size_t)k * 3 + sizeof(uint32_t) * ((size_t)k * (L + 2)); }

extern "C" int etl_kmeans_assign(etl_handle* hh, const int32_t* index, const double* const num[3], int64_t n, int32_t k,
                                 int32_t num_labels, int32_t repeats, const double* d_onehot_sum, const double* d_num,
                                 const double* d_norm2, const double* d_prev_cost, int32_t cluster_offset,
                                 int32_t* out_cluster, double* out_cost, void* stream_) {
    ETL_REQUIRE(hh && num && d_onehot_sum && d_num && d_norm2, "kmeans_assign: NULL argument");
    ETL_REQUIRE(n >= 0 && k >= 1 && num_labels >= 0 && repeats >= 1, "kmeans_assign: bad sizes");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(index && num[0] && num[1] && num[2], "kmeans_assign: NULL column");
    const size_t smem = km_assign_smem(k, num_labels);
    ETL_REQUIRE(smem <= 200 * 1024, "kmeans_assign: k * (num_labels + 5) = %lld doubles exceed shared memory",
                (long long)(smem / 8));
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::KmAssignParams p{};
    p.index = index; p.n = n; p.k = k; p.L = num_labels; p.repeats = repeats;
    for (int j = 0; j < 3; ++j) p.num[j] = num[j];
    p.prep.onehot_sum = d_onehot_sum; p.prep.num = d_num; p.prep.norm2 = d_norm2;
    p.prev_cost = d_prev_cost; p.out_cluster = out_cluster; p.out_cost = out_cost; p.cluster_offset = cluster_offset;
    ETL_CUDA(cudaFuncSetAttribute(etl::kmeans_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = (n + etl::KM_THREADS - 1) / etl::KM_THREADS;
    const int per_sm = smem > 100 * 1024 ? 1 : (smem > 50 * 1024 ? 2 : 4);
    if (blocks > (int64_t)h->sm_count * per_sm) blocks = (int64_t)h->sm_count * per_sm;
    etl::kmeans_assign_kernel<<<(unsigned)blocks, etl::KM_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_kmeans_accumulate(etl_handle* hh, const int32_t* index, const double* const num[3],
                                     const int32_t* cluster, int64_t n, int32_t k, int32_t num_labels, int64_t* d_count,
                                     int64_t* d_hist, double* d_numsum, void* stream_) {
    ETL_REQUIRE(hh && num && d_count && d_hist && d_numsum, "kmeans_accumulate: NULL argument");
    ETL_REQUIRE(n >= 0 && k >= 1 && num_labels >= 0, "kmeans_accumulate: bad sizes");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    ETL_CUDA(cudaMemsetAsync(d_count, 0, sizeof(int64_t) * k, stream));
    ETL_CUDA(cudaMemsetAsync(d_hist, 0, sizeof(int64_t) * k * (num_labels + 1), stream));
    ETL_CUDA(cudaMemsetAsync(d_numsum, 0, sizeof(double) * k * 3, stream));
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(index && cluster && num[0] && num[1] && num[2], "kmeans_accumulate: NULL column");
    const size_t smem = km_accum_smem(k, num_labels);
    ETL_REQUIRE(smem <= 200 * 1024, "kmeans_accumulate: k * (num_labels + 2) too large for shared memory");
    etl::KmAccumParams p{};
    p.index = index; p.cluster = cluster; p.n = n; p.k = k; p.L = num_labels;
    for (int j = 0; j < 3; ++j) p.num[j] = num[j];
    p.count = reinterpret_cast<unsigned long long*>(d_count);
    p.hist = reinterpret_cast<unsigned long long*>(d_hist);
    p.numsum = d_numsum;
    static const int owner = [] { const char* e = getenv("ETL_KM_OWNER"); return e ? atoi(e) : 1; }();
    const size_t osmem = (size_t)(etl::KM_THREADS / 32) * 96 * 2 * sizeof(double) + (size_t)k * (num_labels + 1) * sizeof(uint32_t);
    if (owner && k <= 32 && osmem <= 200 * 1024) {   // lane c owns cluster c: no floating-point atomics in the row loop
        ETL_CUDA(cudaFuncSetAttribute(etl::kmeans_accumulate_owner_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)osmem));
        int64_t ob = (n + etl::KM_THREADS - 1) / etl::KM_THREADS;
        const int opsm = osmem > 100 * 1024 ? 1 : (osmem > 50 * 1024 ? 2 : 8);
        if (ob > (int64_t)h->sm_count * opsm) ob = (int64_t)h->sm_count * opsm;
        etl::kmeans_accumulate_owner_kernel<<<(unsigned)ob, etl::KM_THREADS, osmem, stream>>>(p);
        ETL_CUDA(cudaGetLastError());
        return ETL_OK;
    }
    ETL_CUDA(cudaFuncSetAttribute(etl::kmeans_accumulate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = (n + etl::KM_THREADS - 1) / etl::KM_THREADS;
    const int per_sm = smem > 100 * 1024 ? 1 : (smem > 50 * 1024 ? 2 : 4);
    if (blocks > (int64_t)h->sm_count * per_sm) blocks = (int64_t)h->sm_count * per_sm;
    etl::kmeans_accumulate_kernel<<<(unsigned)blocks, etl::KM_THREADS, smem, stream>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_kmeans_update(etl_handle* hh, const int64_t* d_count, const int64_t* d_hist, const double* d_numsum,
                                 int32_t k, int32_t num_labels, int32_t repeats, double* centers, void* stream_) {
    ETL_REQUIRE(hh && d_count && d_hist && d_numsum && centers, "kmeans_update: NULL argument");
    ETL_REQUIRE(k >= 1 && num_labels >= 0 && repeats >= 1, "kmeans_update: k=%d num_labels=%d repeats=%d", k, num_labels, repeats);
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::kmeans_update_kernel<<<(unsigned)k, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        reinterpret_cast<const unsigned long long*>(d_count), reinterpret_cast<const unsigned long long*>(d_hist), d_numsum,
        num_labels, repeats, centers);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_kmeans_row_norms(etl_handle* hh, const int32_t* index, const double* const num[3], int64_t n,
                                    int32_t num_labels, int32_t repeats, double* out, void* stream_) {
    ETL_REQUIRE(hh && num, "kmeans_row_norms: NULL argument");
    ETL_REQUIRE(n >= 0 && num_labels >= 0 && repeats >= 1, "kmeans_row_norms: bad n / num_labels / repeats");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(index && num[0] && num[1] && num[2] && out, "kmeans_row_norms: NULL column");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    int64_t blocks = (n + etl::KM_THREADS - 1) / etl::KM_THREADS;
    if (blocks > (int64_t)h->sm_count * 8) blocks = (int64_t)h->sm_count * 8;
    etl::kmeans_row_norms_kernel<<<(unsigned)blocks, etl::KM_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        index, num[0], num[1], num[2], n, num_labels, repeats, out);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_kmeans_silhouette(etl_handle* hh, const int32_t* index, const double* const num[3], const int32_t* cluster,
                                     int64_t n, int32_t k, int32_t num_labels, int32_t repeats, const double* d_label_sum,
                                     const double* d_num_sum, const double* d_psi, const double* d_count, double* out,
                                     void* stream_) {
    ETL_REQUIRE(hh && num && d_label_sum && d_num_sum && d_psi && d_count, "kmeans_silhouette: NULL argument");
    ETL_REQUIRE(n >= 0 && k >= 1 && num_labels >= 0 && repeats >= 1, "kmeans_silhouette: bad n / k / num_labels / repeats");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(index && cluster && num[0] && num[1] && num[2] && out, "kmeans_silhouette: NULL column");
    const size_t smem = ((size_t)k * (num_labels + 1) + (size_t)k * 5) * sizeof(double);
    ETL_REQUIRE(smem <= 200 * 1024, "kmeans_silhouette: k * (num_labels + 6) too large for shared memory");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::KmSilhouetteParams p{};
    p.index = index; p.cluster = cluster; p.n = n; p.k = k; p.L = num_labels; p.repeats = repeats;
    for (int j = 0; j < 3; ++j) p.num[j] = num[j];
    p.label_sum = d_label_sum; p.num_sum = d_num_sum; p.psi = d_psi; p.count = d_count; p.out = out;
    ETL_CUDA(cudaFuncSetAttribute(etl::kmeans_silhouette_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = (n + etl::KM_THREADS - 1) / etl::KM_THREADS;
    const int per_sm = smem > 100 * 1024 ? 1 : (smem > 50 * 1024 ? 2 : 4);
    if (blocks > (int64_t)h->sm_count * per_sm) blocks = (int64_t)h->sm_count * per_sm;
    etl::kmeans_silhouette_kernel<<<(unsigned)blocks, etl::KM_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
