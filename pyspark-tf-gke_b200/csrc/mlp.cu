// Device-resident hand-off to the trainer: the Keras model of
// workloads/raw-tf/train_tf_ps.py:328-343 - Dense(16, relu) -> Dense(32, relu) ->
// Dense(64, relu) -> Dense(num_classes, softmax), sparse categorical cross-entropy,
// Adam(1e-3) - trained straight from the X float32[N,3] / y int32[N] arrays the ETL
// left in HBM (train_tf_ps.py:596-601,664-672), batches chosen by an index array.
//
// The model has ~3.7 K parameters and a step is ~1.4 MFLOP at batch 64, so steps
// are launch-latency bound, not throughput bound: ONE launch runs a whole run of
// consecutive steps in a single CTA with the parameters and Adam moments resident
// in shared memory (no per-step launch, no parameter traffic to HBM).  Per step:
// gather the batch rows by index, forward layer by layer (one task per (sample,
// neuron)), softmax + loss, backward, per-parameter gradient sums over the batch,
// Adam update (Keras form: lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t),
// w -= lr_t * m / (sqrt(v) + eps)).  fp32 throughout, as in the reference.
// mlp_evaluate_kernel is the forward-only pass (validation loss / accuracy).
#include "common.cuh"

namespace etl {

constexpr int MLP_THREADS = 256;        // evaluate kernel
constexpr int MLP_TRAIN_THREADS = 1024; // training kernel: one CTA, as many warps as an SM takes (latency hiding)
constexpr int MLP_IN = 3, MLP_H1 = 16, MLP_H2 = 32, MLP_H3 = 64;
constexpr int MLP_MAX_CLASSES = 64;
constexpr int MLP_CHUNK = 64;  // samples whose activations are held in shared memory at a time

__host__ __device__ inline int mlp_param_count(int C) {
    return MLP_IN * MLP_H1 + MLP_H1 + MLP_H1 * MLP_H2 + MLP_H2 + MLP_H2 * MLP_H3 + MLP_H3 + MLP_H3 * C + C;
}
// offsets of (W1, b1, W2, b2, W3, b3, W4, b4); W is [in][out] row-major, the Keras kernel layout
struct MlpLayout {
    int w[4], b[4], in[4], out[4];
    __host__ __device__ explicit MlpLayout(int C) {
        const int ins[4] = {MLP_IN, MLP_H1, MLP_H2, MLP_H3}, outs[4] = {MLP_H1, MLP_H2, MLP_H3, C};
        int o = 0;
        for (int l = 0; l < 4; ++l) {
            in[l] = ins[l]; out[l] = outs[l];
            w[l] = o; o += ins[l] * outs[l];
            b[l] = o; o += outs[l];
        }
    }
};

struct MlpTrainParams {
    const float* X;        // [N][3]
    const int32_t* y;      // [N]
    const int32_t* order;  // [steps][batch] row indices
    const int32_t* step_rows;  // optional [steps]: rows of step s that count (1..batch): the short batch that ends a pass
    float* params;         // [P] in / out
    float* adam_m;         // [P] in / out
    float* adam_v;         // [P] in / out
    int32_t classes, batch, steps;
    int64_t step0;         // Adam time step of the first step minus one (0 for a fresh optimiser)
    float lr, beta1, beta2, eps;
    float* loss;           // [steps] mean loss of each step's batch
    float* acc;            // [steps] accuracy of each step's batch (before the update)
};

// activations of one chunk of samples, [sample][neuron]
struct MlpActs {
    float* x;   // [CHUNK][3]
    float* a1;  // [CHUNK][16]
    float* a2;  // [CHUNK][32]
    float* a3;  // [CHUNK][64]
    float* z;   // [CHUNK][C]  logits -> probabilities -> dlogits
    float* d3;  // [CHUNK][64]
    float* d2;  // [CHUNK][32]
    float* d1;  // [CHUNK][16]
    int32_t* label;
};

// NIN / NOUT > 0: compile-time layer shape (loops unroll, index division by a constant);
// 0: taken from the arguments (the output layer, whose width is the class count)
template <int NIN, int NOUT>
__device__ __forceinline__ void mlp_dense_forward(const float* w, const float* b, const float* in, float* out, int n, int nin_,
                                                  int nout_, bool relu) {
    const int nin = NIN ? NIN : nin_, nout = NOUT ? NOUT : nout_;
    for (int t = threadIdx.x; t < n * nout; t += blockDim.x) {
        const int s = t / nout, j = t - s * nout;
        float acc = b[j];
#pragma unroll 16
        for (int i = 0; i < nin; ++i) acc = fmaf(in[s * nin + i], w[i * nout + j], acc);
        out[t] = relu ? fmaxf(acc, 0.f) : acc;
    }
}
// din[s][i] = (ain[s][i] > 0) * sum_j dout[s][j] * w[i][j]
template <int NIN, int NOUT>
__device__ __forceinline__ void mlp_dense_backward(const float* w, const float* dout, const float* ain, float* din, int n,
                                                   int nin_, int nout_) {
    const int nin = NIN ? NIN : nin_, nout = NOUT ? NOUT : nout_;
    for (int t = threadIdx.x; t < n * nin; t += blockDim.x) {
        const int s = t / nin, i = t - s * nin;
        float acc = 0.f;
#pragma unroll 16
        for (int j = 0; j < nout; ++j) acc = fmaf(dout[s * nout + j], w[i * nout + j], acc);
        din[t] = ain[t] > 0.f ? acc : 0.f;
    }
}
// g[W] += in^T . delta, g[b] += column sums of delta, over the n samples of a chunk
template <int NIN, int NOUT>
__device__ __forceinline__ void mlp_dense_grad(float* gw, float* gb, const float* in, const float* delta, int n, int nin_,
                                               int nout_) {
    const int nin = NIN ? NIN : nin_, nout = NOUT ? NOUT : nout_;
    for (int t = threadIdx.x; t < nin * nout; t += blockDim.x) {
        const int i = t / nout, j = t - i * nout;
        float acc = 0.f;
#pragma unroll 16
        for (int s = 0; s < n; ++s) acc = fmaf(in[s * nin + i], delta[s * nout + j], acc);
        gw[t] += acc;
    }
    for (int j = threadIdx.x; j < nout; j += blockDim.x) {
        float acc = 0.f;
        for (int s = 0; s < n; ++s) acc += delta[s * nout + j];
        gb[j] += acc;
    }
}

__global__ void __launch_bounds__(MLP_TRAIN_THREADS) mlp_train_kernel(const MlpTrainParams p) {
    extern __shared__ __align__(16) float mlp_smem[];
    const int C = p.classes;
    const int P = mlp_param_count(C);
    const MlpLayout L(C);
    float* w = mlp_smem;        // parameters
    float* am = w + P;          // Adam first moment
    float* av = am + P;         // Adam second moment
    float* g = av + P;          // gradient of the current step
    MlpActs A;
    A.x = g + P;
    A.a1 = A.x + MLP_CHUNK * MLP_IN;
    A.a2 = A.a1 + MLP_CHUNK * MLP_H1;
    A.a3 = A.a2 + MLP_CHUNK * MLP_H2;
    A.z = A.a3 + MLP_CHUNK * MLP_H3;
    A.d3 = A.z + MLP_CHUNK * C;
    A.d2 = A.d3 + MLP_CHUNK * MLP_H3;
    A.d1 = A.d2 + MLP_CHUNK * MLP_H2;
    A.label = reinterpret_cast<int32_t*>(A.d1 + MLP_CHUNK * MLP_H1);
    __shared__ float s_loss[MLP_CHUNK];
    __shared__ int s_hit[MLP_CHUNK];
    for (int i = threadIdx.x; i < P; i += blockDim.x) { w[i] = p.params[i]; am[i] = p.adam_m[i]; av[i] = p.adam_v[i]; }
    __syncthreads();
    for (int step = 0; step < p.steps; ++step) {
        const int bs = p.step_rows ? p.step_rows[step] : p.batch;     // the same for every thread
        const float inv_b = 1.f / (float)bs;
        for (int i = threadIdx.x; i < P; i += blockDim.x) g[i] = 0.f;
        float loss_sum = 0.f;  // thread 0 only
        int hits = 0;
        for (int c0 = 0; c0 < bs; c0 += MLP_CHUNK) {
            const int n = bs - c0 < MLP_CHUNK ? bs - c0 : MLP_CHUNK;
            __syncthreads();
            for (int t = threadIdx.x; t < n; t += blockDim.x) {
                const int32_t r = p.order[(int64_t)step * p.batch + c0 + t];
                A.x[t * 3] = p.X[(int64_t)r * 3]; A.x[t * 3 + 1] = p.X[(int64_t)r * 3 + 1]; A.x[t * 3 + 2] = p.X[(int64_t)r * 3 + 2];
                A.label[t] = p.y[r];
            }
            __syncthreads();
            mlp_dense_forward<MLP_IN, MLP_H1>(w + L.w[0], w + L.b[0], A.x, A.a1, n, MLP_IN, MLP_H1, true);
            __syncthreads();
            mlp_dense_forward<MLP_H1, MLP_H2>(w + L.w[1], w + L.b[1], A.a1, A.a2, n, MLP_H1, MLP_H2, true);
            __syncthreads();
            mlp_dense_forward<MLP_H2, MLP_H3>(w + L.w[2], w + L.b[2], A.a2, A.a3, n, MLP_H2, MLP_H3, true);
            __syncthreads();
            mlp_dense_forward<MLP_H3, 0>(w + L.w[3], w + L.b[3], A.a3, A.z, n, MLP_H3, C, false);
            __syncthreads();
            // softmax, loss, dlogits = (p - onehot) / batch   (one thread per sample)
            for (int s = threadIdx.x; s < n; s += blockDim.x) {
                float* z = A.z + s * C;
                float mx = z[0];
                int arg = 0;
                for (int j = 1; j < C; ++j)
                    if (z[j] > mx) { mx = z[j]; arg = j; }
                float den = 0.f;
                for (int j = 0; j < C; ++j) { z[j] = expf(z[j] - mx); den += z[j]; }
                const int lab = A.label[s];
                const float pl = z[lab] / den;
                s_loss[s] = -logf(fmaxf(pl, 1e-7f));  // Keras clips probabilities at epsilon = 1e-7
                s_hit[s] = arg == lab;
                for (int j = 0; j < C; ++j) z[j] = (z[j] / den - (j == lab ? 1.f : 0.f)) * inv_b;
            }
            __syncthreads();
            if (threadIdx.x == 0)
                for (int s = 0; s < n; ++s) { loss_sum += s_loss[s]; hits += s_hit[s]; }
            mlp_dense_backward<MLP_H3, 0>(w + L.w[3], A.z, A.a3, A.d3, n, MLP_H3, C);
            __syncthreads();
            mlp_dense_backward<MLP_H2, MLP_H3>(w + L.w[2], A.d3, A.a2, A.d2, n, MLP_H2, MLP_H3);
            __syncthreads();
            mlp_dense_backward<MLP_H1, MLP_H2>(w + L.w[1], A.d2, A.a1, A.d1, n, MLP_H1, MLP_H2);
            __syncthreads();
            // gradients: one thread per parameter, summed over the chunk's samples
            mlp_dense_grad<MLP_IN, MLP_H1>(g + L.w[0], g + L.b[0], A.x, A.d1, n, MLP_IN, MLP_H1);
            mlp_dense_grad<MLP_H1, MLP_H2>(g + L.w[1], g + L.b[1], A.a1, A.d2, n, MLP_H1, MLP_H2);
            mlp_dense_grad<MLP_H2, MLP_H3>(g + L.w[2], g + L.b[2], A.a2, A.d3, n, MLP_H2, MLP_H3);
            mlp_dense_grad<MLP_H3, 0>(g + L.w[3], g + L.b[3], A.a3, A.z, n, MLP_H3, C);
        }
        __syncthreads();
        // Adam (Keras form)
        __shared__ float s_lr_t;
        if (threadIdx.x == 0) {
            const double t1 = (double)(p.step0 + step + 1);
            s_lr_t = p.lr * (float)(sqrt(1.0 - pow((double)p.beta2, t1)) / (1.0 - pow((double)p.beta1, t1)));
        }
        __syncthreads();
        const float lr_t = s_lr_t;
        for (int i = threadIdx.x; i < P; i += blockDim.x) {
            const float gi = g[i];
            const float m = p.beta1 * am[i] + (1.f - p.beta1) * gi;
            const float v = p.beta2 * av[i] + (1.f - p.beta2) * gi * gi;
            am[i] = m; av[i] = v;
            w[i] -= lr_t * m / (sqrtf(v) + p.eps);
        }
        if (threadIdx.x == 0) {
            p.loss[step] = loss_sum * inv_b;
            p.acc[step] = (float)hits * inv_b;
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < P; i += blockDim.x) { p.params[i] = w[i]; p.adam_m[i] = am[i]; p.adam_v[i] = av[i]; }
}

// forward only: every CTA takes 64-sample chunks (activations in shared memory, as in training)
// and writes the per-sample loss and predicted class
__global__ void __launch_bounds__(MLP_THREADS) mlp_evaluate_kernel(const float* __restrict__ X, const int32_t* __restrict__ y,
                                                                   int64_t n, const float* __restrict__ params, int C,
                                                                   float* __restrict__ loss, int32_t* __restrict__ pred) {
    extern __shared__ __align__(16) float mlp_smem[];
    const int P = mlp_param_count(C);
    const MlpLayout L(C);
    float* w = mlp_smem;
    float* x = w + P;
    float* a1 = x + MLP_CHUNK * MLP_IN;
    float* a2 = a1 + MLP_CHUNK * MLP_H1;
    float* a3 = a2 + MLP_CHUNK * MLP_H2;
    float* z = a3 + MLP_CHUNK * MLP_H3;
    for (int i = threadIdx.x; i < P; i += MLP_THREADS) w[i] = params[i];
    const int64_t chunks = (n + MLP_CHUNK - 1) / MLP_CHUNK;
    for (int64_t c = blockIdx.x; c < chunks; c += gridDim.x) {
        const int64_t s0 = c * MLP_CHUNK;
        const int m = n - s0 < MLP_CHUNK ? (int)(n - s0) : MLP_CHUNK;
        __syncthreads();
        for (int t = threadIdx.x; t < m * MLP_IN; t += MLP_THREADS) x[t] = X[s0 * MLP_IN + t];
        __syncthreads();
        mlp_dense_forward<MLP_IN, MLP_H1>(w + L.w[0], w + L.b[0], x, a1, m, MLP_IN, MLP_H1, true);
        __syncthreads();
        mlp_dense_forward<MLP_H1, MLP_H2>(w + L.w[1], w + L.b[1], a1, a2, m, MLP_H1, MLP_H2, true);
        __syncthreads();
        mlp_dense_forward<MLP_H2, MLP_H3>(w + L.w[2], w + L.b[2], a2, a3, m, MLP_H2, MLP_H3, true);
        __syncthreads();
        mlp_dense_forward<MLP_H3, 0>(w + L.w[3], w + L.b[3], a3, z, m, MLP_H3, C, false);
        __syncthreads();
        for (int s = threadIdx.x; s < m; s += MLP_THREADS) {
            const float* zs = z + s * C;
            float mx = zs[0];
            int arg = 0;
            for (int j = 1; j < C; ++j)
                if (zs[j] > mx) { mx = zs[j]; arg = j; }
            if (pred) pred[s0 + s] = arg;
            if (loss) {
                float den = 0.f;
                for (int j = 0; j < C; ++j) den += expf(zs[j] - mx);
                loss[s0 + s] = -logf(fmaxf(expf(zs[y[s0 + s]] - mx) / den, 1e-7f));
            }
        }
    }
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_mlp_param_count(int32_t classes) { return etl::mlp_param_count(classes); }

extern "C" int etl_mlp_train_steps(etl_handle* hh, const float* X, const int32_t* y, int64_t n, const int32_t* d_order,
                                   int32_t batch, int32_t steps, int32_t classes, float* d_params, float* d_adam_m,
                                   float* d_adam_v, int64_t step0, float lr, float beta1, float beta2, float eps,
                                   float* d_loss, float* d_acc, void* stream_) {
    return etl_mlp_train_steps_ragged(hh, X, y, n, d_order, nullptr, batch, steps, classes, d_params, d_adam_m, d_adam_v, step0,
                                      lr, beta1, beta2, eps, d_loss, d_acc, stream_);
}

extern "C" int etl_mlp_train_steps_ragged(etl_handle* hh, const float* X, const int32_t* y, int64_t n, const int32_t* d_order,
                                          const int32_t* d_step_rows, int32_t batch, int32_t steps, int32_t classes,
                                          float* d_params, float* d_adam_m, float* d_adam_v, int64_t step0, float lr,
                                          float beta1, float beta2, float eps, float* d_loss, float* d_acc, void* stream_) {
    ETL_REQUIRE(hh && X && y && d_order && d_params && d_adam_m && d_adam_v && d_loss && d_acc, "mlp_train_steps: NULL argument");
    ETL_REQUIRE(n >= 1 && batch >= 1 && steps >= 0 && classes >= 2 && classes <= etl::MLP_MAX_CLASSES,
                "mlp_train_steps: n=%lld batch=%d steps=%d classes=%d", (long long)n, batch, steps, classes);
    if (steps == 0) return ETL_OK;
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::MlpTrainParams p{};
    p.X = X; p.y = y; p.order = d_order; p.step_rows = d_step_rows; p.params = d_params; p.adam_m = d_adam_m; p.adam_v = d_adam_v;
    p.classes = classes; p.batch = batch; p.steps = steps; p.step0 = step0;
    p.lr = lr; p.beta1 = beta1; p.beta2 = beta2; p.eps = eps; p.loss = d_loss; p.acc = d_acc;
    const int P = etl::mlp_param_count(classes);
    const size_t smem = sizeof(float) * ((size_t)4 * P + (size_t)etl::MLP_CHUNK * (3 + 2 * (16 + 32 + 64) + classes)) +
                        sizeof(int32_t) * etl::MLP_CHUNK;
    ETL_CUDA(cudaFuncSetAttribute(etl::mlp_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    etl::mlp_train_kernel<<<1, etl::MLP_TRAIN_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_mlp_evaluate(etl_handle* hh, const float* X, const int32_t* y, int64_t n, const float* d_params,
                                int32_t classes, float* d_loss, int32_t* d_pred, void* stream_) {
    ETL_REQUIRE(hh && d_params, "mlp_evaluate: NULL argument");
    ETL_REQUIRE(n >= 0 && classes >= 2 && classes <= etl::MLP_MAX_CLASSES, "mlp_evaluate: n=%lld classes=%d", (long long)n, classes);
    ETL_REQUIRE(!d_loss || y, "mlp_evaluate: a loss needs labels");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(X != nullptr, "mlp_evaluate: X is NULL");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    const size_t smem = sizeof(float) * ((size_t)etl::mlp_param_count(classes) + (size_t)etl::MLP_CHUNK * (3 + 16 + 32 + 64 + classes));
    ETL_CUDA(cudaFuncSetAttribute(etl::mlp_evaluate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = (n + etl::MLP_CHUNK - 1) / etl::MLP_CHUNK;
    if (blocks > (int64_t)h->sm_count * 4) blocks = (int64_t)h->sm_count * 4;
    etl::mlp_evaluate_kernel<<<(unsigned)blocks, etl::MLP_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(
        X, y, n, d_params, classes, d_loss, d_pred);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
