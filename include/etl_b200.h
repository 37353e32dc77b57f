/* etl_b200.h - C ABI of the B200-native raw-spark ETL hot path.
 *
 * The reference (greg-ogs/PySpark-TF-GKE) has no FFI for this path: every
 * operation below is a PySpark DataFrame / Spark ML call made from Python
 * (workloads/raw-spark/k_means.py:22-77) or a pure-Python loop
 * (workloads/raw-tf/train_tf_ps.py:75-149).  Each entry point cites the
 * reference call site it replaces (paths relative to the reference root).
 * INTEGRATION.md shows the ctypes binding a maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, a negative ETL_E* code otherwise;
 *     etl_last_error() gives the message of the calling thread's last failure;
 *     no C++ exception crosses this boundary.
 *   - the CALLER owns every data buffer (device pointers unless the name ends
 *     in _host); the library owns only the opaque per-GPU handle: a scratch
 *     arena that grows on demand (growth = cudaMalloc + stream sync; call
 *     etl_reserve() up front to avoid it on the hot path).
 *   - calls are asynchronous on `stream` (a cudaStream_t passed as void*);
 *     results written to device memory are ready when the stream reaches them.
 *     One handle per GPU; a handle is not thread-safe.
 *   - columns are Arrow-style: a data buffer plus an optional validity bitmap,
 *     bit i of the column in bit (i % 32) of uint32 word (i / 32), 1 = valid;
 *     a NULL bitmap means "no nulls".  Data pointers must be 16-byte aligned,
 *     bitmaps 4-byte aligned; a bitmap must hold ceil(n / 32) words.
 *   - value dtypes: ETL_F64, ETL_F32, ETL_I32, ETL_I64.
 */
#ifndef ETL_B200_H
#define ETL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ETL_OK 0
#define ETL_EINVAL (-1)   /* bad argument (null pointer, misalignment, size)   */
#define ETL_ECUDA (-2)    /* a CUDA runtime call failed                        */
#define ETL_ENOMEM (-3)   /* scratch arena could not grow                      */
#define ETL_EDOMAIN (-4)  /* input outside the documented domain               */
#define ETL_ECAPACITY (-5)/* a caller-sized table or buffer was too small      */

#define ETL_F64 0
#define ETL_F32 1
#define ETL_I32 2
#define ETL_I64 3

#if defined(__GNUC__)
#define ETL_API __attribute__((visibility("default")))
#else
#define ETL_API
#endif

typedef struct etl_handle etl_handle;

/* ---- lifecycle -------------------------------------------------------- */

/* ABI version of this build (bumped when a signature changes). */
ETL_API int etl_abi_version(void);
/* Message of the calling thread's most recent failure ("" if none). */
ETL_API const char* etl_last_error(void);
/* Replaces CreateSparkSession.new_spark_session (spark_session.py:37-88): the
 * per-executor state.  `device` = CUDA ordinal.  Allocates a small scratch
 * arena; etl_destroy frees it. */
ETL_API int etl_create(int device, etl_handle** out);
ETL_API int etl_destroy(etl_handle* h);
/* Grow the scratch arena to at least `bytes` now (synchronises the device). */
ETL_API int etl_reserve(etl_handle* h, size_t bytes);

/* ---- a2/a3/a5: fused filter + impute + cast + compaction -------------- */

/* Replaces, in one pass over the rows,
 *   input_df.filter(col("measure_name").isNotNull())            k_means.py:27
 *   .count()                                                    k_means.py:28
 *   withColumn(c, when(isNull(c) | isnan(c), mean).otherwise(c)) x3
 *                                                               k_means.py:49-51
 * keep row i iff key_valid bit i is set; for each kept row, in input order,
 * write the key code and the three numerics cast to double, a null or NaN cell
 * replaced by fill[c].  val_dtype[c] is the input dtype of column c (ETL_F64 /
 * ETL_F32 / ETL_I32; an integer column cannot hold NaN).  out_* must hold n
 * rows.  *d_out_rows (device int64) receives the kept-row count.
 * Scratch: one 8-byte word per chunk (<= 65536 chunks), from the handle. */
ETL_API int etl_project_filter_impute(etl_handle* h,
                              const int32_t* key_code, const uint32_t* key_valid,
                              const void* const val[3], const uint32_t* const val_valid[3],
                              const int32_t val_dtype[3], const double fill[3], int64_t n,
                              int32_t* out_code, double* const out_val[3],
                              int64_t* d_out_rows, void* stream);

/* Same operation with HOST buffers (pinned or pageable): copies the inputs to
 * device staging owned by the handle in row chunks, runs the kernel, copies the
 * kept rows back, on internal streams so copy-in / kernel / copy-out of
 * neighbouring chunks overlap; returns after everything has landed.
 * *out_rows = kept rows.  This is the call the end-to-end figure times. */
ETL_API int etl_project_filter_impute_host(etl_handle* h,
                                   const int32_t* key_code, const uint32_t* key_valid,
                                   const void* const val[3], const uint32_t* const val_valid[3],
                                   const int32_t val_dtype[3], const double fill[3], int64_t n,
                                   int32_t* out_code, double* const out_val[3], int64_t* out_rows);

/* No counterpart in the reference (it never manages device memory): the same
 * projection as etl_project_filter_impute for callers that must not trigger a
 * device allocation - e.g. a process whose GPU memory is wholly owned by a pool
 * allocator, or a latency-critical loop that cannot take cudaMalloc's device
 * synchronisation.  Needs no handle and calls no cudaMalloc / cudaFree or
 * allocating library routine: the kernel's scratch is a fixed 131,328-byte
 * __device__ array that each GPU loading the library holds until exit.
 *   - runs on the CURRENT device; buffers as for etl_project_filter_impute;
 *   - n > ETL_NOALLOC_MAX_ROWS returns ETL_ECAPACITY (nothing is launched);
 *   - the first call on a device resolves the array's device address and sizes
 *     the grid: make it before starting any stream capture;
 *   - all calls on a device share the scratch (it keeps the last call's words;
 *     each call re-arms what it needs on `stream`): calls ordered on one stream
 *     are fine, calls that overlap on the GPU race and corrupt each other. */
#define ETL_NOALLOC_MAX_ROWS (((int64_t)1) << 31)
ETL_API int etl_project_filter_impute_noalloc(const int32_t* key_code, const uint32_t* key_valid,
                                      const void* const val[3], const uint32_t* const val_valid[3],
                                      const int32_t val_dtype[3], const double fill[3], int64_t n,
                                      int32_t* out_code, double* const out_val[3],
                                      int64_t* d_out_rows, void* stream);

/* ---- a2: null count ---------------------------------------------------- */

/* Replaces input_df.filter(col("measure_name").isNull()).count()
 * (k_means.py:23): *d_out (device int64) = number of 0 bits among the first n
 * bits of `valid` (0 if valid == NULL). */
ETL_API int etl_null_count(etl_handle* h, const uint32_t* valid, int64_t n, int64_t* d_out, void* stream);

/* ---- a4: column sum / count (the Average partial state) ---------------- */

/* Replaces select(c).filter(~isnan(c) & c.isNotNull()).agg({c: "avg"})
 * (k_means.py:45-48) for up to 3 columns in one pass.  If row_valid != NULL
 * only rows whose bit is set take part (the lazy isNotNull filter of :27 that
 * Spark re-evaluates in this job).  d_sum[c] (device double) = fp64 sum,
 * d_count[c] (device int64) = contributing rows; the caller divides.  The
 * reduction tree has a fixed shape for a given n: results are reproducible
 * run to run, and exact when all partial sums are exactly representable. */
ETL_API int etl_column_sum_count(etl_handle* h, int32_t ncols,
                         const void* const val[], const uint32_t* const val_valid[],
                         const int32_t val_dtype[], const uint32_t* row_valid, int64_t n,
                         double* d_sum, int64_t* d_count, void* stream);

/* ---- synthetic tables (bench / test infrastructure) -------------------- */

/* Counter-based generator, identical to pyspark_tf_gke_b200.synth on the host:
 * bits(seed, stream, row) = splitmix64 state `row` of the stream keyed by
 * (seed, stream).  Fills one column and (if valid != NULL) its validity
 * bitmap.  kind: 0 = integer-valued double floor(u * 100001) ("exact"),
 * 1 = u * 1e5 ("real"), 2 = int32 key (bits % modulus), 3 = int64 key
 * (bits % modulus).  null_rate / nan_rate in [0, 1). */
ETL_API int etl_synth_column(etl_handle* h, uint64_t seed, uint32_t stream_id, int32_t kind,
                     uint64_t modulus, double null_rate, double nan_rate, int64_t row0, int64_t n,
                     void* out, uint32_t* valid, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ETL_B200_H */
