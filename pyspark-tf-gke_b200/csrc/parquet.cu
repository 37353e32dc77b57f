// Parquet page decode (north-star item 1 "CSV/Parquet decode"; no reference site -
// the reference reads CSV / JDBC).  The host (parquet.py) parses footer and page
// headers and hands over a table of page descriptors; everything that scales
// with the number of rows happens here:
//   pq_hybrid_decode_kernel  one CTA per RLE / bit-packed hybrid stream
//        (definition levels, dictionary indices): one thread walks the run
//        headers into a run table, then the warps expand the runs in parallel;
//   pq_gather_kernel         one CTA per data page: definition levels -> validity
//        bits (ballot) and, through a block scan, each row's position among the
//        page's non-null values; PLAIN values are cast to double, dictionary
//        indices go through the column's lookup table (labels -> sorted codes, or
//        numeric dictionary values).
#include "common.cuh"

namespace etl {

constexpr int PQ_THREADS = 256;
constexpr int PQ_KIND_F64 = 0, PQ_KIND_F32 = 1, PQ_KIND_I64 = 3, PQ_KIND_DICT = 4;  // 2 = int32 (the default branch)

struct PqStream {        // one hybrid stream to expand
    int64_t src_offset;  // byte offset in the file buffer
    int32_t src_bytes;
    int32_t bit_width;
    int64_t out_offset;  // element offset in the int32 output
    int32_t max_values;  // stop after this many values (bit-packed groups are padded to 8)
    int32_t exact;       // 1: the stream must yield max_values (definition levels); 0: it may hold fewer (indices of
                         // the non-null values only - their number is not known before the levels are decoded)
    int64_t run_offset;  // element offset of this stream's run table
    int64_t run_capacity;
};
struct PqRun {
    int32_t out_start, count;
    int32_t src_pos;     // bit-packed: byte position of the first group; RLE: unused
    int32_t value;       // RLE: the repeated value; bit-packed: -1
};
struct PqPage {
    int64_t row_start;
    int32_t num_values;
    int32_t kind;        // PQ_KIND_*
    int64_t levels_offset;  // element offset of the page's definition levels, -1 = required column
    int64_t val_offset;     // PLAIN: byte offset of the values; DICT: element offset of the decoded indices
    int32_t dict_id;
    int32_t pad;
};

__device__ __forceinline__ uint32_t pq_extract(const uint8_t* p, int64_t bit, int bw) {
    const uint8_t* q = p + (bit >> 3);
    const int sh = (int)(bit & 7);
    uint64_t w = 0;
    const int need = (sh + bw + 7) >> 3;  // <= 5 bytes for bw <= 32
    for (int i = 0; i < need; ++i) w |= (uint64_t)q[i] << (8 * i);
    return (uint32_t)((w >> sh) & ((bw >= 32) ? 0xffffffffull : ((1ull << bw) - 1)));
}

__global__ void __launch_bounds__(PQ_THREADS) pq_hybrid_decode_kernel(const uint8_t* __restrict__ file,
                                                                      const PqStream* __restrict__ streams,
                                                                      PqRun* __restrict__ runs, int32_t* __restrict__ out,
                                                                      int32_t* status) {
    __shared__ int s_nruns;
    const PqStream st = streams[blockIdx.x];
    const uint8_t* src = file + st.src_offset;
    PqRun* my_runs = runs + st.run_offset;
    if (threadIdx.x == 0) {
        int pos = 0, produced = 0, n = 0;
        const int vbytes = (st.bit_width + 7) >> 3;
        while (pos < st.src_bytes && produced < st.max_values) {
            uint32_t h = 0;
            int shift = 0;
            while (pos < st.src_bytes) {  // ULEB128 run header
                const uint8_t c = src[pos++];
                h |= (uint32_t)(c & 0x7f) << shift;
                shift += 7;
                if (!(c & 0x80)) break;
            }
            PqRun r;
            r.out_start = produced;
            if (h & 1) {  // bit-packed groups of 8 values
                const int groups = (int)(h >> 1);
                r.count = groups * 8;
                r.src_pos = pos;
                r.value = -1;
                pos += groups * st.bit_width;
            } else {      // run-length: value repeated h >> 1 times
                r.count = (int)(h >> 1);
                r.src_pos = pos;
                uint32_t v = 0;
                for (int i = 0; i < vbytes && pos + i < st.src_bytes; ++i) v |= (uint32_t)src[pos + i] << (8 * i);
                r.value = (int32_t)v;
                pos += vbytes;
            }
            if (r.count <= 0) break;
            if (r.count > st.max_values - produced) r.count = st.max_values - produced;
            if (n >= st.run_capacity) { atomicAdd(status, 1); break; }
            my_runs[n++] = r;
            produced += r.count;
        }
        if (st.exact && produced < st.max_values) atomicAdd(status + 1, 1);  // stream ended early
        s_nruns = n;
    }
    __syncthreads();
    const int nruns = s_nruns;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t* dst = out + st.out_offset;
    for (int r = warp; r < nruns; r += PQ_THREADS / 32) {
        const PqRun run = my_runs[r];
        if (run.value >= 0 || st.bit_width == 0) {
            const int32_t v = st.bit_width == 0 ? 0 : run.value;
            for (int i = lane; i < run.count; i += 32) dst[run.out_start + i] = v;
        } else {
            const uint8_t* p = src + run.src_pos;
            for (int i = lane; i < run.count; i += 32)
                dst[run.out_start + i] = (int32_t)pq_extract(p, (int64_t)i * st.bit_width, st.bit_width);
        }
    }
}

struct PqGatherParams {
    const uint8_t* file;
    const PqPage* pages;
    const int32_t* levels;      // decoded definition levels (all optional pages, concatenated)
    const int32_t* indices;     // decoded dictionary indices (all dictionary pages, concatenated)
    const void* const* dict;    // per dictionary: double[] (numeric) or int32[] (string label codes)
    const int32_t* dict_size;
    int32_t string_column;      // 1: out is int32 codes; 0: out is double
    void* out;
    uint32_t* valid;            // zero-initialised bitmap, filled with atomicOr (pages need not start on a word)
    int32_t* status;            // [2] dictionary index out of range
};

__global__ void __launch_bounds__(PQ_THREADS) pq_gather_kernel(const PqGatherParams p) {
    __shared__ uint32_t s_warp[PQ_THREADS / 32];
    __shared__ uint32_t s_running;
    const PqPage pg = p.pages[blockIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_running = 0;
    __syncthreads();
    for (int base = 0; base < pg.num_values; base += PQ_THREADS) {
        const int i = base + threadIdx.x;
        const bool in = i < pg.num_values;
        const bool ok = in && (pg.levels_offset < 0 || p.levels[pg.levels_offset + i] > 0);
        const unsigned b = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) s_warp[warp] = __popc(b);
        __syncthreads();
        uint32_t before = s_running;
        uint32_t chunk = 0;
#pragma unroll
        for (int w = 0; w < PQ_THREADS / 32; ++w) {
            if (w < warp) before += s_warp[w];
            chunk += s_warp[w];
        }
        const uint32_t rank = before + __popc(b & lanemask_lt());  // position among the page's non-null values
        const int64_t row = pg.row_start + i;
        if (in) {
            double v = 0.0;
            int32_t code = -1;
            if (ok) {
                if (pg.kind == PQ_KIND_DICT) {
                    const int32_t ix = p.indices[pg.val_offset + rank];
                    if (ix < 0 || ix >= p.dict_size[pg.dict_id]) {
                        atomicAdd(p.status + 2, 1);
                    } else if (p.string_column) {
                        code = static_cast<const int32_t*>(p.dict[pg.dict_id])[ix];
                    } else {
                        v = static_cast<const double*>(p.dict[pg.dict_id])[ix];
                    }
                } else {
                    const uint8_t* src = p.file + pg.val_offset;
                    if (pg.kind == PQ_KIND_F64) {
                        uint64_t u = 0;  // pages are not 8-byte aligned in the file
                        for (int k = 0; k < 8; ++k) u |= (uint64_t)src[(int64_t)rank * 8 + k] << (8 * k);
                        v = __longlong_as_double((long long)u);
                    } else if (pg.kind == PQ_KIND_I64) {
                        uint64_t u = 0;
                        for (int k = 0; k < 8; ++k) u |= (uint64_t)src[(int64_t)rank * 8 + k] << (8 * k);
                        v = (double)(long long)u;
                    } else {
                        uint32_t u = 0;
                        for (int k = 0; k < 4; ++k) u |= (uint32_t)src[(int64_t)rank * 4 + k] << (8 * k);
                        v = pg.kind == PQ_KIND_F32 ? (double)__uint_as_float(u) : (double)(int32_t)u;
                    }
                }
            }
            if (p.string_column) static_cast<int32_t*>(p.out)[row] = code;
            else static_cast<double*>(p.out)[row] = v;
        }
        // validity bits: this warp's 32 rows may straddle two bitmap words
        if (lane == 0 && b) {
            const int64_t r0 = pg.row_start + base + warp * 32;
            const int sh = (int)(r0 & 31);
            atomicOr(p.valid + (r0 >> 5), b << sh);
            if (sh && (b >> (32 - sh))) atomicOr(p.valid + (r0 >> 5) + 1, b >> (32 - sh));
        }
        __syncthreads();
        if (threadIdx.x == 0) s_running += chunk;
        __syncthreads();
    }
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_parquet_expand_streams(etl_handle* hh, const uint8_t* file, const void* d_streams, int32_t nstreams,
                                          void* d_runs, int32_t* d_out, int32_t* d_status, void* stream_) {
    ETL_REQUIRE(hh && d_status, "parquet_expand_streams: NULL argument");
    ETL_REQUIRE(nstreams >= 0, "parquet_expand_streams: nstreams < 0");
    if (nstreams == 0) return ETL_OK;
    ETL_REQUIRE(file && d_streams && d_runs && d_out, "parquet_expand_streams: NULL buffer");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::pq_hybrid_decode_kernel<<<(unsigned)nstreams, etl::PQ_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        file, static_cast<const etl::PqStream*>(d_streams), static_cast<etl::PqRun*>(d_runs), d_out, d_status);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_parquet_gather_pages(etl_handle* hh, const uint8_t* file, const void* d_pages, int32_t npages,
                                        const int32_t* d_levels, const int32_t* d_indices, const void* const* d_dict,
                                        const int32_t* d_dict_size, int32_t string_column, void* out, uint32_t* valid,
                                        int32_t* d_status, void* stream_) {
    ETL_REQUIRE(hh && d_status, "parquet_gather_pages: NULL argument");
    ETL_REQUIRE(npages >= 0, "parquet_gather_pages: npages < 0");
    if (npages == 0) return ETL_OK;
    ETL_REQUIRE(file && d_pages && out && valid, "parquet_gather_pages: NULL buffer");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::PqGatherParams p{};
    p.file = file; p.pages = static_cast<const etl::PqPage*>(d_pages); p.levels = d_levels; p.indices = d_indices;
    p.dict = d_dict; p.dict_size = d_dict_size; p.string_column = string_column; p.out = out; p.valid = valid;
    p.status = d_status;
    etl::pq_gather_kernel<<<(unsigned)npages, etl::PQ_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(p);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
