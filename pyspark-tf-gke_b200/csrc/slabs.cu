// The partition passes of the high-cardinality groupBy WITHOUT count passes: every destination of a
// scatter owns a fixed-capacity slab and a persistent cursor (scatter.cuh, SLAB mode).
//
//   etl_slab_partition_rows      (key, value) rows - two columns or {key, value} records - into the slabs of
//                                `ranks x buckets` destinations.  ranks > 1 with one buffer pointer per rank is
//                                the exchange: rank r's pointer is r's receive buffer mapped into this GPU's
//                                address space (NVLink peer memory) and this rank's slab in it starts at
//                                slab_origin.  Cursors persist, so the rows can be sent chunk by chunk while
//                                the receivers already work on what has arrived.
//   etl_slab_partition_records   segments of a slab buffer -> slabs: the receiver's first level (all `segments`
//                                sources into one set of buckets, only the records that arrived since
//                                d_seg_prev) or the second level (segment = bucket, cut into `sub` pieces).
//   etl_peer_publish_counts      after a chunk: this rank's cursors -> one word in every receiver's count table.
//
// A record that finds its slab full is counted in *d_overflow and dropped: the caller must then discard the
// result and repeat the operation on the counted path (etl_hash_partition / etl_hash_subpartition), which
// sizes every destination exactly.  With capacities of mean + 6 sigma that happens for skewed keys only.
#include "common.cuh"
#include "scatter.cuh"

namespace etl {

__device__ __forceinline__ longlong2 ld_stream_rec(const longlong2* p) {
    longlong2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.s64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
    return r;
}

__global__ void publish_counts_kernel(const int64_t* __restrict__ cursor, int ranks, int per_rank, PartDest tables, int64_t slot) {
    const int i = threadIdx.x;            // cursor index = rank * per_rank + j  ->  tables[rank][slot + j]
    if (i < ranks * per_rank) {
        // system-scope release store: ordered after this stream's earlier peer stores by the kernel boundary
        asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(tables.base[i / per_rank] + slot + i % per_rank), "l"(cursor[i])
                     : "memory");
    }
}

// ---- the bucketed exchange's data movement: this rank's slabs of bucket group [bucket0, bucket0 + nbuckets) go to
//      every owner's receive buffer (layout [source][bucket]) with 16-byte stores from registers - measured at
//      ~700 GB/s per GPU over NVLink (tools/nvlink_probe.py), where stream-ordered peer copies through the copy
//      engines reached 330-400 GB/s at 8 GPUs (64 copies of 32-64 MB per step).  Only the filled part of a slab
//      moves (the cursor is read here), grid.y = (destination rank, bucket) pair, few registers and no shared
//      memory, so the CTAs sit beside the merge kernels of the groups that have already landed.
struct SlabCopyJob {
    const longlong2* send;
    const int64_t* cursor;        // [ranks * buckets] records in every slab of the send buffer
    int ranks, my_rank, buckets, bucket0, nbuckets;
    int64_t slab;
};
constexpr int COPY_THREADS = 128;
constexpr int COPY_UNROLL = 8;
// 128 threads x <= 56 registers: a CTA takes 7 K of an SM's 64 K registers and no shared memory, so it fits beside
// the second-level scatter (1024 threads x 40 registers) and beside two merge CTAs (512 x 48 each).  Few CTAs
// (half the SM count by default) with eight 16-byte loads in flight per thread: measured at 8 GPUs, 64 such CTAs
// move the slabs at ~480 GB/s per GPU and leave the merge kernels alone (0.3 ms left after the last group lands),
// where 300 CTAs reach 530 GB/s but delay the merge by 0.9 ms.
__global__ void __maxnreg__(56) peer_copy_slabs_kernel(const SlabCopyJob job, const PartDest peers) {
    const int r = blockIdx.y / job.nbuckets, b = job.bucket0 + blockIdx.y % job.nbuckets;
    int64_t n = job.cursor[r * job.buckets + b];
    n = n < job.slab ? n : job.slab;
    const longlong2* src = job.send + ((int64_t)r * job.buckets + b) * job.slab;
    longlong2* dst = reinterpret_cast<longlong2*>(peers.base[r]) + ((int64_t)job.my_rank * job.buckets + b) * job.slab;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (COPY_UNROLL - 1) * stride < n; i += COPY_UNROLL * stride) {
        longlong2 v[COPY_UNROLL];
#pragma unroll
        for (int k = 0; k < COPY_UNROLL; ++k) v[k] = ld_stream_rec(src + i + k * stride);
#pragma unroll
        for (int k = 0; k < COPY_UNROLL; ++k) dst[i + k * stride] = v[k];
    }
    for (; i < n; i += stride) dst[i] = ld_stream_rec(src + i);
}

// ---- NVLink write probe (tools/nvlink_probe.py): how fast can one GPU push a buffer into a peer's memory
//      with (0) 16-byte stores from registers - what the exchange scatter does - or (1) bulk async stores of
//      `chunk` bytes from shared memory (TMA engine, no tensor map, no multicast)?
__global__ void __launch_bounds__(256) peer_write_plain_kernel(const longlong2* __restrict__ src, longlong2* dst, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }

constexpr int PROBE_TILE = 32768;
__global__ void __launch_bounds__(32) peer_write_bulk_kernel(const unsigned char* __restrict__ src, unsigned char* dst, int64_t nbytes,
                                                             unsigned chunk) {
    extern __shared__ __align__(128) unsigned char pb_smem[];
    __shared__ uint64_t bar[2];
    if (threadIdx.x != 0) return;
    mbar_init(bar, 1);
    mbar_init(bar + 1, 1);
    mbar_init_fence();
    const int64_t tiles = (nbytes + PROBE_TILE - 1) / PROBE_TILE;
    auto issue = [&](int64_t t, int buf) {
        const int64_t left = nbytes - t * PROBE_TILE;
        const unsigned bytes = left < PROBE_TILE ? (unsigned)left : (unsigned)PROBE_TILE;
        mbar_arrive_expect_tx(bar + buf, bytes);
        bulk_load(pb_smem + buf * PROBE_TILE, src + t * PROBE_TILE, bytes, bar + buf);
    };
    int64_t t = blockIdx.x;
    if (t < tiles) issue(t, 0);
    for (int it = 0; t < tiles; t += gridDim.x, ++it) {
        const int buf = it & 1;
        // the other buffer is free once the stores issued from it two iterations ago have read it
        bulk_wait_read<0>();
        if (t + gridDim.x < tiles) issue(t + gridDim.x, buf ^ 1);
        mbar_wait(bar + buf, (unsigned)(it >> 1) & 1u);
        const int64_t left = nbytes - t * PROBE_TILE;
        const unsigned bytes = left < PROBE_TILE ? (unsigned)left : (unsigned)PROBE_TILE;
        for (unsigned o = 0; o < bytes; o += chunk)
            bulk_store(dst + t * PROBE_TILE + o, pb_smem + buf * PROBE_TILE + o, bytes - o < chunk ? bytes - o : chunk);
        bulk_commit();
    }
    bulk_wait_read<0>();
}

// ---- the same movement through the TMA engine: one thread per CTA streams its share of a slab through a ring of
//      shared-memory stages - bulk async load from this GPU's send buffer, bulk async store into the owner's receive
//      buffer (no tensor map, no multicast: TMA multicast loads here read only this GPU's memory, and this kernel
//      uses none).  The bytes in flight are the ring (48 KB per CTA) instead of eight registers per thread, so the
//      transfer keeps its rate when the partition and merge kernels load HBM beside it, and it takes no issue slots
//      from them.
constexpr int BULK_CHUNK = 8192;
constexpr int BULK_STAGES = 6;
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__global__ void __launch_bounds__(32) peer_copy_slabs_bulk_kernel(const SlabCopyJob job, const PartDest peers) {
    extern __shared__ __align__(128) unsigned char cb_smem[];
    __shared__ uint64_t bar[BULK_STAGES];
    if (threadIdx.x != 0) return;
    const int r = blockIdx.y / job.nbuckets, b = job.bucket0 + blockIdx.y % job.nbuckets;
    int64_t n = job.cursor[r * job.buckets + b];
    n = n < job.slab ? n : job.slab;
    const int64_t nbytes = n * 16;
    const unsigned char* src = reinterpret_cast<const unsigned char*>(job.send + ((int64_t)r * job.buckets + b) * job.slab);
    unsigned char* dst = reinterpret_cast<unsigned char*>(reinterpret_cast<longlong2*>(peers.base[r]) +
                                                          ((int64_t)job.my_rank * job.buckets + b) * job.slab);
    const int64_t chunks = (nbytes + BULK_CHUNK - 1) / BULK_CHUNK;
    const int64_t mine = chunks > blockIdx.x ? (chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;   // chunks x, x + grid, ...
    if (mine == 0) return;
    for (int s = 0; s < BULK_STAGES; ++s) mbar_init(bar + s, 1);
    mbar_init_fence();
    auto bytes_of = [&](int64_t c) {
        const int64_t left = nbytes - c * BULK_CHUNK;
        return left < BULK_CHUNK ? (unsigned)left : (unsigned)BULK_CHUNK;
    };
    auto issue = [&](int64_t j) {
        const int s = (int)(j % BULK_STAGES);
        const int64_t c = blockIdx.x + j * gridDim.x;
        const unsigned bytes = bytes_of(c);
        mbar_arrive_expect_tx(bar + s, bytes);
        bulk_load(cb_smem + s * BULK_CHUNK, src + c * BULK_CHUNK, bytes, bar + s);
    };
    for (int64_t j = 0; j < BULK_STAGES - 1 && j < mine; ++j) issue(j);
    for (int64_t j = 0; j < mine; ++j) {
        const int s = (int)(j % BULK_STAGES);
        const int64_t c = blockIdx.x + j * gridDim.x;
        mbar_wait(bar + s, (unsigned)(j / BULK_STAGES) & 1u);
        bulk_store(dst + c * BULK_CHUNK, cb_smem + s * BULK_CHUNK, bytes_of(c));
        bulk_commit();
        if (j + BULK_STAGES - 1 < mine) {
            bulk_wait_read<1>();      // every store but the one just issued has read its stage: the stage of chunk j - 1 is free
            issue(j + BULK_STAGES - 1);
        }
    }
    bulk_wait_all();
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_slab_partition_rows(etl_handle* hh, const int64_t* keys, int64_t key_stride, const double* vals,
                                       int64_t val_stride, int64_t n, int32_t ranks, int32_t buckets,
                                       int64_t* const* dest_records, int64_t slab_origin, int64_t slab_stride,
                                       int64_t slab_capacity, int64_t* d_cursor, int64_t* d_overflow, int32_t max_ctas,
                                       void* stream_) {
    ETL_REQUIRE(hh && dest_records && d_cursor && d_overflow, "slab_partition_rows: NULL argument");
    ETL_REQUIRE(ranks >= 1 && buckets >= 1 && (int64_t)ranks * buckets <= etl::PART_MAX,
                "slab_partition_rows: ranks=%d x buckets=%d must be in 1..%d", ranks, buckets, etl::PART_MAX);
    ETL_REQUIRE(n >= 0 && slab_origin >= 0 && slab_stride >= 0 && slab_capacity >= 1, "slab_partition_rows: bad n / slab layout");
    if (n == 0) return ETL_OK;
    ETL_REQUIRE(keys && vals, "slab_partition_rows: keys / vals is NULL");
    const bool columns = key_stride == 1 && val_stride == 1 && etl::aligned16(keys) && etl::aligned16(vals);
    const bool records = key_stride == 2 && val_stride == 2 && etl::aligned16(keys) &&
                         reinterpret_cast<const int64_t*>(vals) == keys + 1;
    ETL_REQUIRE(columns || records,
                "slab_partition_rows: rows must be two 16-byte aligned contiguous columns or {key, value} records");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    etl::PartDest dst;
    for (int d = 0; d < etl::PART_MAX; ++d) dst.base[d] = nullptr;
    for (int r = 0; r < ranks; ++r) {
        ETL_REQUIRE(dest_records[r] && etl::aligned16(dest_records[r]), "slab_partition_rows: buffer %d NULL or misaligned", r);
        for (int b = 0; b < buckets; ++b) dst.base[r * buckets + b] = dest_records[r];
    }
    ETL_CUDA(cudaSetDevice(h->device));
    etl::ScatterJob job{};
    job.keys = keys; job.vals = vals; job.n = n; job.ranks = ranks; job.buckets = buckets; job.parts = ranks * buckets;
    job.cursor = reinterpret_cast<unsigned long long*>(d_cursor);
    job.overflow = reinterpret_cast<unsigned long long*>(d_overflow);
    job.slab_origin = slab_origin; job.slab_stride = slab_stride; job.dest_cap = slab_capacity;
    ETL_CUDA(columns ? (etl::launch_scatter_pipelined<false, true>(h->sm_count, job, dst, 1, n, stream, max_ctas))
                     : (etl::launch_scatter_pipelined<true, true>(h->sm_count, job, dst, 1, n, stream, max_ctas)));
    return ETL_OK;
}

extern "C" int etl_slab_partition_records(etl_handle* hh, const int64_t* records, int32_t segments, int64_t seg_capacity,
                                          const int64_t* d_seg_counts, const int64_t* d_seg_prev, int64_t max_seg_records,
                                          int32_t buckets, int32_t sub, int32_t sources, int32_t bucket0,
                                          int64_t* out_records, int64_t slab_capacity, int64_t* d_cursor,
                                          int64_t* d_overflow, int32_t max_ctas, void* stream_) {
    ETL_REQUIRE(hh && d_seg_counts && d_cursor && d_overflow, "slab_partition_records: NULL argument");
    ETL_REQUIRE(segments >= 1 && segments <= 65535 && seg_capacity >= 1 && slab_capacity >= 1 && max_seg_records >= 0,
                "slab_partition_records: bad segments / capacities");
    ETL_REQUIRE(buckets >= 1 && sub >= 0 && sub <= etl::PART_MAX && (sub > 0 || buckets <= etl::PART_MAX),
                "slab_partition_records: buckets=%d sub=%d", buckets, sub);
    ETL_REQUIRE(sources >= 0 && bucket0 >= 0, "slab_partition_records: sources / bucket0 < 0");
    ETL_REQUIRE(sub == 0 || (int64_t)buckets * sub <= 65536, "slab_partition_records: buckets x sub must be <= 65536");
    ETL_REQUIRE(sub == 0 || (sources == 0 ? segments == buckets : (segments % sources == 0 && bucket0 + segments / sources <= buckets)),
                "slab_partition_records: the second level needs segments == buckets, or sources x (a range of buckets)");
    ETL_REQUIRE(sub > 0 || sources == 0, "slab_partition_records: a [source][bucket] input is a second-level input");
    if (max_seg_records == 0) return ETL_OK;
    ETL_REQUIRE(records && out_records && etl::aligned16(records) && etl::aligned16(out_records),
                "slab_partition_records: NULL or misaligned record buffer");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    etl::ScatterJob job{};
    job.keys = records; job.seg_counts = d_seg_counts; job.seg_prev = d_seg_prev; job.seg_cap = seg_capacity;
    job.cursor = reinterpret_cast<unsigned long long*>(d_cursor);
    job.overflow = reinterpret_cast<unsigned long long*>(d_overflow);
    job.slab_origin = 0; job.slab_stride = slab_capacity; job.dest_cap = slab_capacity;
    if (sub == 0) {
        job.ranks = 1; job.buckets = buckets; job.parts = buckets; job.shared_dest = 1;
    } else {
        job.sub = sub; job.fine = (unsigned)(buckets * sub); job.parts = sub;
        job.seg_sources = sources; job.seg_bucket0 = bucket0; job.seg_bucket_stride = buckets;
    }
    etl::PartDest dst;
    for (int d = 0; d < etl::PART_MAX; ++d) dst.base[d] = out_records;
    ETL_CUDA((etl::launch_scatter_pipelined<true, true>(h->sm_count, job, dst, segments, max_seg_records, stream, max_ctas)));
    return ETL_OK;
}

extern "C" int etl_peer_publish_counts(etl_handle* hh, const int64_t* d_cursor, int32_t ranks, int32_t per_rank,
                                       int64_t* const* peer_tables, int64_t slot, void* stream_) {
    ETL_REQUIRE(hh && d_cursor && peer_tables, "peer_publish_counts: NULL argument");
    ETL_REQUIRE(ranks >= 1 && per_rank >= 1 && (int64_t)ranks * per_rank <= etl::PART_MAX && slot >= 0,
                "peer_publish_counts: ranks=%d per_rank=%d slot=%lld", ranks, per_rank, (long long)slot);
    Handle* h = reinterpret_cast<Handle*>(hh);
    etl::PartDest t;
    for (int d = 0; d < etl::PART_MAX; ++d) t.base[d] = nullptr;
    for (int r = 0; r < ranks; ++r) {
        ETL_REQUIRE(peer_tables[r] != nullptr, "peer_publish_counts: table %d is NULL", r);
        t.base[r] = peer_tables[r];
    }
    ETL_CUDA(cudaSetDevice(h->device));
    etl::publish_counts_kernel<<<1, etl::PART_MAX, 0, static_cast<cudaStream_t>(stream_)>>>(d_cursor, ranks, per_rank, t, slot);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

// Diagnostic (no reference counterpart): push nbytes (a multiple of 16) from src (local) to dst (peer or local)
// with 16-byte register stores (mode 0) or bulk async stores of `chunk` bytes (mode 1; chunk a multiple of 16).
extern "C" int etl_peer_write_probe(etl_handle* hh, const void* src, void* dst, int64_t nbytes, int32_t mode, int32_t chunk,
                                    int32_t ctas, void* stream_) {
    ETL_REQUIRE(hh && src && dst && nbytes >= 0 && (nbytes & 15) == 0 && etl::aligned16(src) && etl::aligned16(dst),
                "peer_write_probe: bad buffers");
    ETL_REQUIRE(mode == 0 || (mode == 1 && chunk >= 16 && (chunk & 15) == 0), "peer_write_probe: mode=%d chunk=%d", mode, chunk);
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    if (nbytes == 0) return ETL_OK;
    if (ctas <= 0) ctas = h->sm_count * 4;
    if (mode == 0) {
        etl::peer_write_plain_kernel<<<ctas, 256, 0, stream>>>(static_cast<const longlong2*>(src), static_cast<longlong2*>(dst),
                                                                nbytes / 16);
    } else {
        ETL_CUDA(cudaFuncSetAttribute(etl::peer_write_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * etl::PROBE_TILE));
        etl::peer_write_bulk_kernel<<<ctas, 32, 2 * etl::PROBE_TILE, stream>>>(static_cast<const unsigned char*>(src),
                                                                               static_cast<unsigned char*>(dst), nbytes, (unsigned)chunk);
    }
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

// Stream-ordered copy (copy engine): how the bucketed exchange moves a rank's slabs into a peer's receive
// buffer without occupying SMs.  dst may be peer memory mapped into this process.
extern "C" int etl_copy_async(etl_handle* hh, void* dst, const void* src, int64_t nbytes, void* stream_) {
    ETL_REQUIRE(hh && nbytes >= 0 && (nbytes == 0 || (dst && src)), "copy_async: bad argument");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    if (nbytes) ETL_CUDA(cudaMemcpyAsync(dst, src, (size_t)nbytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream_)));
    return ETL_OK;
}

static int peer_copy_slabs(bool bulk, etl_handle* hh, const int64_t* send_records, int64_t* const* peer_records, int32_t ranks,
                           int32_t my_rank, int32_t buckets, int32_t bucket0, int32_t nbuckets, int64_t slab_capacity,
                           const int64_t* d_cursor, int32_t ctas_per_slab, void* stream_) {
    ETL_REQUIRE(hh && send_records && peer_records && d_cursor, "peer_copy_slabs: NULL argument");
    ETL_REQUIRE(ranks >= 1 && buckets >= 1 && (int64_t)ranks * buckets <= etl::PART_MAX && my_rank >= 0 && my_rank < ranks,
                "peer_copy_slabs: ranks=%d buckets=%d my_rank=%d", ranks, buckets, my_rank);
    ETL_REQUIRE(bucket0 >= 0 && nbuckets >= 1 && bucket0 + nbuckets <= buckets && slab_capacity >= 1,
                "peer_copy_slabs: bucket range %d..%d of %d", bucket0, bucket0 + nbuckets, buckets);
    ETL_REQUIRE(etl::aligned16(send_records), "peer_copy_slabs: send buffer misaligned");
    Handle* h = reinterpret_cast<Handle*>(hh);
    etl::PartDest peers;
    for (int d = 0; d < etl::PART_MAX; ++d) peers.base[d] = nullptr;
    for (int r = 0; r < ranks; ++r) {
        ETL_REQUIRE(peer_records[r] && etl::aligned16(peer_records[r]), "peer_copy_slabs: buffer %d NULL or misaligned", r);
        peers.base[r] = peer_records[r];
    }
    ETL_CUDA(cudaSetDevice(h->device));
    etl::SlabCopyJob job{};
    job.send = reinterpret_cast<const longlong2*>(send_records); job.cursor = d_cursor; job.ranks = ranks; job.my_rank = my_rank;
    job.buckets = buckets; job.bucket0 = bucket0; job.nbuckets = nbuckets; job.slab = slab_capacity;
    const int pairs = ranks * nbuckets;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (bulk) {
        int gx = ctas_per_slab > 0 ? ctas_per_slab : (h->sm_count + pairs - 1) / pairs;     // about one one-warp CTA per SM
        const int64_t want = (slab_capacity * 16 + etl::BULK_CHUNK - 1) / etl::BULK_CHUNK;
        if (gx > want) gx = (int)want;
        if (gx < 1) gx = 1;
        const int smem = etl::BULK_STAGES * etl::BULK_CHUNK;
        ETL_CUDA(cudaFuncSetAttribute(etl::peer_copy_slabs_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        etl::peer_copy_slabs_bulk_kernel<<<dim3((unsigned)gx, (unsigned)pairs), 32, smem, stream>>>(job, peers);
    } else {
        int gx = ctas_per_slab > 0 ? ctas_per_slab : (h->sm_count / 2 + pairs - 1) / pairs;     // about one small CTA per two SMs
        const int64_t want = (slab_capacity + etl::COPY_THREADS * etl::COPY_UNROLL - 1) / (etl::COPY_THREADS * etl::COPY_UNROLL);
        if (gx > want) gx = (int)want;
        if (gx < 1) gx = 1;
        etl::peer_copy_slabs_kernel<<<dim3((unsigned)gx, (unsigned)pairs), etl::COPY_THREADS, 0, stream>>>(job, peers);
    }
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_peer_copy_slabs(etl_handle* hh, const int64_t* send_records, int64_t* const* peer_records, int32_t ranks,
                                   int32_t my_rank, int32_t buckets, int32_t bucket0, int32_t nbuckets, int64_t slab_capacity,
                                   const int64_t* d_cursor, int32_t ctas_per_slab, void* stream_) {
    return peer_copy_slabs(false, hh, send_records, peer_records, ranks, my_rank, buckets, bucket0, nbuckets, slab_capacity, d_cursor,
                           ctas_per_slab, stream_);
}

extern "C" int etl_peer_copy_slabs_bulk(etl_handle* hh, const int64_t* send_records, int64_t* const* peer_records, int32_t ranks,
                                        int32_t my_rank, int32_t buckets, int32_t bucket0, int32_t nbuckets, int64_t slab_capacity,
                                        const int64_t* d_cursor, int32_t ctas_per_slab, void* stream_) {
    return peer_copy_slabs(true, hh, send_records, peer_records, ranks, my_rank, buckets, bucket0, nbuckets, slab_capacity, d_cursor,
                           ctas_per_slab, stream_);
}
