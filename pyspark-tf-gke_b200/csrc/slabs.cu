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

__global__ void publish_counts_kernel(const int64_t* __restrict__ cursor, int ranks, PartDest tables, int64_t slot) {
    const int r = threadIdx.x;
    if (r < ranks) {
        // system-scope release store: ordered after this stream's earlier peer stores by the kernel boundary
        asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(tables.base[r] + slot), "l"(cursor[r]) : "memory");
    }
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
                                          int32_t buckets, int32_t sub, int64_t* out_records, int64_t slab_capacity,
                                          int64_t* d_cursor, int64_t* d_overflow, int32_t max_ctas, void* stream_) {
    ETL_REQUIRE(hh && d_seg_counts && d_cursor && d_overflow, "slab_partition_records: NULL argument");
    ETL_REQUIRE(segments >= 1 && segments <= 65535 && seg_capacity >= 1 && slab_capacity >= 1 && max_seg_records >= 0,
                "slab_partition_records: bad segments / capacities");
    ETL_REQUIRE(buckets >= 1 && sub >= 0 && sub <= etl::PART_MAX && (sub > 0 || buckets <= etl::PART_MAX),
                "slab_partition_records: buckets=%d sub=%d", buckets, sub);
    ETL_REQUIRE(sub == 0 || (segments == buckets && (int64_t)buckets * sub <= 65536),
                "slab_partition_records: the second level needs segments == buckets and buckets x sub <= 65536");
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
    }
    etl::PartDest dst;
    for (int d = 0; d < etl::PART_MAX; ++d) dst.base[d] = out_records;
    ETL_CUDA((etl::launch_scatter_pipelined<true, true>(h->sm_count, job, dst, segments, max_seg_records, stream, max_ctas)));
    return ETL_OK;
}

extern "C" int etl_peer_publish_counts(etl_handle* hh, const int64_t* d_cursor, int32_t ranks, int64_t* const* peer_tables,
                                       int64_t slot, void* stream_) {
    ETL_REQUIRE(hh && d_cursor && peer_tables, "peer_publish_counts: NULL argument");
    ETL_REQUIRE(ranks >= 1 && ranks <= etl::PART_MAX && slot >= 0, "peer_publish_counts: ranks=%d slot=%lld", ranks, (long long)slot);
    Handle* h = reinterpret_cast<Handle*>(hh);
    etl::PartDest t;
    for (int d = 0; d < etl::PART_MAX; ++d) t.base[d] = nullptr;
    for (int r = 0; r < ranks; ++r) {
        ETL_REQUIRE(peer_tables[r] != nullptr, "peer_publish_counts: table %d is NULL", r);
        t.base[r] = peer_tables[r];
    }
    ETL_CUDA(cudaSetDevice(h->device));
    etl::publish_counts_kernel<<<1, etl::PART_MAX, 0, static_cast<cudaStream_t>(stream_)>>>(d_cursor, ranks, t, slot);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
