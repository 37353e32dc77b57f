// Handle lifecycle, error reporting and scratch management of the C ABI.
#include "common.cuh"

namespace etl {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int arena_reserve(Handle* h, size_t bytes) {
    h->csv_text = nullptr;   // whoever reserves the arena is about to overwrite what the CSV index left in it
    if (bytes <= h->scratch_bytes) return ETL_OK;
    size_t want = h->scratch_bytes ? h->scratch_bytes : (size_t)1 << 20;
    while (want < bytes) want *= 2;
    ETL_CUDA(cudaDeviceSynchronize());
    if (h->scratch) ETL_CUDA(cudaFree(h->scratch));
    h->scratch = nullptr;
    h->scratch_bytes = 0;
    cudaError_t e = cudaMalloc(&h->scratch, want);
    if (e != cudaSuccess) {
        set_error("scratch arena: cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
        return ETL_ENOMEM;
    }
    h->scratch_bytes = want;
    return ETL_OK;
}

int tile_state_acquire(Handle* h, size_t words, cudaStream_t stream, uint32_t* epoch) {
    if (words > h->tile_words) {
        size_t want = h->tile_words ? h->tile_words : 4096;
        while (want < words) want *= 2;
        ETL_CUDA(cudaDeviceSynchronize());
        if (h->tile_state) ETL_CUDA(cudaFree(h->tile_state));
        h->tile_state = nullptr;
        h->tile_words = 0;
        cudaError_t e = cudaMalloc((void**)&h->tile_state, want * sizeof(uint64_t));
        if (e != cudaSuccess) {
            set_error("tile state: cudaMalloc(%zu) failed: %s", want * sizeof(uint64_t), cudaGetErrorString(e));
            return ETL_ENOMEM;
        }
        h->tile_words = want;
        ETL_CUDA(cudaMemset(h->tile_state, 0, want * sizeof(uint64_t)));
    }
    if (h->epoch >= TS_EPOCH_MAX) {  // tags would repeat: clear them (stream ordered)
        ETL_CUDA(cudaMemsetAsync(h->tile_state, 0, h->tile_words * sizeof(uint64_t), stream));
        h->epoch = 0;
    }
    *epoch = ++h->epoch;
    return ETL_OK;
}

}  // namespace etl

using etl::Handle;

extern "C" {

int etl_abi_version(void) { return 3; }

const char* etl_last_error(void) { return etl::g_err; }

int etl_create(int device, etl_handle** out) {
    ETL_REQUIRE(out != nullptr, "etl_create: out is NULL");
    *out = nullptr;
    int count = 0;
    ETL_CUDA(cudaGetDeviceCount(&count));
    ETL_REQUIRE(device >= 0 && device < count, "etl_create: device %d out of range (%d visible)", device, count);
    ETL_CUDA(cudaSetDevice(device));
    int sm_count = 0;
    ETL_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
    Handle* h = new Handle();
    h->device = device;
    h->sm_count = sm_count;
    int rc = etl::arena_reserve(h, (size_t)1 << 20);
    if (rc != ETL_OK) {
        delete h;
        return rc;
    }
    cudaError_t e = cudaMalloc((void**)&h->counters, 64 * sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMemset(h->counters, 0, 64 * sizeof(unsigned int));
    if (e != cudaSuccess) {
        etl::set_error("etl_create: counters: %s", cudaGetErrorString(e));
        cudaFree(h->scratch);
        delete h;
        return ETL_ECUDA;
    }
    *out = reinterpret_cast<etl_handle*>(h);
    return ETL_OK;
}

int etl_destroy(etl_handle* hh) {
    if (!hh) return ETL_OK;
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    if (h->scratch) cudaFree(h->scratch);
    if (h->tile_state) cudaFree(h->tile_state);
    if (h->counters) cudaFree(h->counters);
    if (h->stage) cudaFree(h->stage);
    for (auto& s : h->streams)
        if (s) cudaStreamDestroy(s);
    for (auto& e : h->events)
        if (e) cudaEventDestroy(e);
    delete h;
    return ETL_OK;
}

int etl_reserve(etl_handle* hh, size_t bytes) {
    ETL_REQUIRE(hh != nullptr, "etl_reserve: handle is NULL");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    return etl::arena_reserve(h, bytes);
}

}  // extern "C"
