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
 * rows.  *d_out_rows (device int64) receives the kept-row count.  For this entry
 * the bitmaps, too, must be 16-byte aligned (they are fetched by bulk copies).
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

/* Number of kernel launches (= row chunks) the host-buffer entry makes for n rows. */
ETL_API int64_t etl_project_filter_impute_host_launches(int64_t n);

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

/* StandardScaler.fit's summary of up to 3 columns (north-star item 5 "feature
 * standardisation"; the reference's pipeline stops at VectorAssembler,
 * k_means.py:64-68, so there is no reference call site): over the same cells
 * etl_column_sum_count takes (row_valid & val_valid & ~isnan), d_count[c] = cells,
 * d_mean[c] = sum / count, d_m2[c] = sum of (x - mean)^2 - two passes of the same
 * fixed-shape tree, the mean never leaves the device.  Spark's unbiased std is
 * sqrt(m2 / (count - 1)); feed shift = mean (withMean) and scale = 1 / std to
 * etl_assemble_features_* / etl_pack_tf_features. */
ETL_API int etl_column_moments(etl_handle* h, int32_t ncols, const void* const val[], const uint32_t* const val_valid[],
                       const int32_t val_dtype[], const uint32_t* row_valid, int64_t n, int64_t* d_count,
                       double* d_mean, double* d_m2, void* stream);

/* ---- a1: CSV text -> columnar device buffers ----------------------------- */

/* Together these replace spark.read.csv(path, header=True, inferSchema=True)
 * (spark_checks/python_checks/spark_workload_to_cloud_k8s.py:48) and the
 * csv.DictReader loop of load_csv (workloads/raw-tf/train_tf_ps.py:114-137) for
 * the columns the ETL stage uses.  `text` is the whole file in device memory,
 * 16-byte aligned, with at least 16 readable bytes after nbytes.  RFC-4180
 * quoting (quoted fields may hold commas and newlines, "" is an escaped quote);
 * blank lines are dropped, as Spark and csv.reader do.
 *
 * etl_csv_index_rows:  *d_rows (device int64) = number of records, header line
 *                      included.  Leaves per-block state and the quote / line-start / comma / newline
 *                      masks of every 64-byte window (nbytes / 2 bytes) in the
 *                      handle's arena, so that the second pass reads those
 *                      instead of the text.
 * etl_csv_row_offsets: row_start[r] (device int64[rows]) = byte offset of record
 *                      r.  Must directly follow etl_csv_index_rows on the same
 *                      text (it reads that state). */
ETL_API int etl_csv_index_rows(etl_handle* h, const uint8_t* text, int64_t nbytes, int64_t* d_rows, void* stream);
ETL_API int etl_csv_row_offsets(etl_handle* h, const uint8_t* text, int64_t nbytes, int64_t* row_start, void* stream);

/* Decode up to 8 columns of records [first_row, first_row + rows).  field[s] =
 * position of column s in a record (0-based, < 256); kind[s]: 0 = numeric ->
 * out[s] is double[rows], 1 = string -> out[s] is int32[rows] holding the slot
 * of the string in the column's dictionary (-1 where null).  valid[s] =
 * ceil(rows/32) bitmap words, fully written.  An empty cell, a cell missing from
 * a short record and (numeric) an unparsable cell are null.  Numerals are
 * converted correctly rounded (as Double.parseDouble / float() do) for up to 60
 * significant digits and any exponent (overflow to inf, underflow through the
 * subnormals); a numeral with more significant digits and a non-zero tail is
 * never approximated: it is counted in d_status[0] and the caller must treat the
 * decode as failed.
 * Dictionaries (string columns): dict_hash[s] / dict_rep[s] = uint64[capacity]
 * device arrays zeroed by the caller before the first call (they persist across
 * calls, so several calls share one dictionary); capacity = power of two.  A slot
 * s is used iff dict_hash[s][slot] != 0; dict_rep[s][slot] = (byte offset << 24)
 * | (quoted << 23) | length of one occurrence of the string in `text`.
 * flags: ETL_CSV_TRIM_STRINGS strips ASCII whitespace around string cells
 * (load_csv, train_tf_ps.py:120); ETL_CSV_NAN_IS_NULL makes a NaN literal null
 * (load_csv drops such rows, train_tf_ps.py:129) instead of a NaN value (Spark).
 * d_status (device int32[4], zeroed by the caller): [0] numerals outside the
 * exact domain, [1] dictionary-full events, [2] bit s set = column s held a
 * non-integer numeral (inferSchema: double rather than int), [3] unparsable
 * numerals.  *d_first_bad_row (device int64, caller sets it to INT64_MAX) = the
 * lowest record with a [0] or [1] event. */
#define ETL_CSV_TRIM_STRINGS 1
#define ETL_CSV_NAN_IS_NULL 2
/* The caller vouches that `text` is byte for byte the text the handle's last etl_csv_index_rows call
 * indexed: the row scan then walks the quote / comma / newline masks that call left in the arena (no
 * byte compares).  Ignored - the text itself is scanned - when the arena no longer holds them. */
#define ETL_CSV_USE_INDEX_MASKS 4
ETL_API int etl_csv_parse_columns(etl_handle* h, const uint8_t* text, int64_t nbytes, const int64_t* row_start,
                          int64_t first_row, int64_t rows, int32_t ncols, const int32_t* field,
                          const int32_t* kind, void* const* out, uint32_t* const* valid,
                          uint64_t* const* dict_hash, uint64_t* const* dict_rep,
                          const int32_t* dict_capacity, int32_t flags, int32_t* d_status,
                          int64_t* d_first_bad_row, void* stream);

/* a7: StringIndexerModel.transform (k_means.py:74) on dictionary codes, in place:
 * codes[i] = table[codes[i]] if 0 <= codes[i] < table_size else `missing`
 * (handleInvalid="keep" maps unseen / null to numLabels). */
ETL_API int etl_remap_codes(etl_handle* h, int32_t* codes, int64_t n, const int32_t* table, int32_t table_size,
                    int32_t missing, void* stream);

/* ---- north-star extension: Parquet page decode (no reference site) -------- */

/* The reference reads CSV / JDBC only; Parquet is named by the north star.  The
 * host parses footer and page headers (Thrift compact, a few hundred bytes per
 * MB) and describes the work in two device tables; the row-proportional work is
 * here.  `file` = the whole file in device memory (plus, for compressed column
 * chunks, the region etl_parquet_inflate_pages fills).
 * etl_parquet_expand_streams: d_streams = nstreams records {int64 src_offset;
 *   int32 src_bytes, bit_width; int64 out_offset; int32 max_values, exact; int64
 *   run_offset, run_capacity} - one RLE / bit-packed hybrid stream each
 *   (definition levels: bit_width 1; dictionary indices: width from the page);
 *   values land in d_out[out_offset ...] (int32).  One warp per stream walks the
 *   run headers and expands each run as it is read; d_runs and the run_offset /
 *   run_capacity fields are unused since ABI 2 (pass NULL / 0).
 * etl_parquet_gather_pages: d_pages = npages records {int64 row_start; int32
 *   num_values, kind (0 f64, 1 f32, 2 i32, 3 i64 PLAIN; 4 dictionary); int64
 *   levels_offset (-1: required column), val_offset (PLAIN: byte offset in file;
 *   dictionary: element offset in d_indices); int32 dict_id, pad}.  Row r of a
 *   page is null iff its definition level is 0; non-null rows take the page's
 *   values in order.  out: double[rows] (numeric types are cast to double) or,
 *   string_column = 1, int32[rows] codes looked up in d_dict[dict_id] (-1 where
 *   null); valid: zeroed bitmap.  d_status (int32[3], zeroed): [0] run table
 *   overflow, [1] stream shorter than promised, [2] dictionary index out of range. */
ETL_API int etl_parquet_expand_streams(etl_handle* h, const uint8_t* file, const void* d_streams, int32_t nstreams,
                               void* d_runs, int32_t* d_out, int32_t* d_status, void* stream);
ETL_API int etl_parquet_gather_pages(etl_handle* h, const uint8_t* file, const void* d_pages, int32_t npages,
                             const int32_t* d_levels, const int32_t* d_indices, const void* const* d_dict,
                             const int32_t* d_dict_size, int32_t string_column, void* out, uint32_t* valid,
                             int32_t* d_status, void* stream);

/* Compressed pages and PLAIN strings (same north-star item; no reference site).
 * etl_parquet_inflate_pages: d_jobs = njobs records {int64 src, src_bytes, dst,
 *   dst_bytes} (byte offsets in `buffer`): Snappy block format, one warp per page;
 *   the caller sizes `buffer` as file bytes + a region for the inflated pages and
 *   describes inflated pages to the other entry points by offsets into that region.
 *   A bit_width < 0 in a stream record means "read the width byte in front of the
 *   indices on the device" (the host cannot see into a compressed page).
 * etl_parquet_plain_strings: d_pages = npages records {int64 val_offset; int32
 *   val_bytes, max_values; int64 out_offset}: PLAIN BYTE_ARRAY values ([u32
 *   length][bytes] ...) are entered into a device dictionary (dict_hash / dict_rep:
 *   zeroed uint64[dict_capacity], a power of two; rep = byte offset << 24 | length)
 *   and their slot numbers written to d_out[out_offset ...], where
 *   etl_parquet_gather_pages reads them as dictionary indices.
 * d_status int32[4]: additionally [0] dictionary full, [3] corrupt Snappy stream. */
ETL_API int etl_parquet_inflate_pages(etl_handle* h, uint8_t* buffer, const void* d_jobs, int32_t njobs, int32_t* d_status,
                              void* stream);
ETL_API int etl_parquet_plain_strings(etl_handle* h, const uint8_t* buffer, const void* d_pages, int32_t npages,
                              uint64_t* dict_hash, uint64_t* dict_rep, int32_t dict_capacity, int32_t* d_out,
                              int32_t* d_status, void* stream);

/* ---- a6: StringIndexer.fit's group-by-count ------------------------------ */

/* Replaces the aggregation inside StringIndexer(handleInvalid="keep").fit
 * (k_means.py:34,73): d_counts[b] (device int64[bins]) = number of valid rows
 * whose code is b; codes outside [0, bins) are counted in *d_out_of_range.  The
 * caller orders labels by (count desc, label asc) on the host - at most a few
 * thousand pairs - as Spark does on the driver. */
ETL_API int etl_code_histogram(etl_handle* h, const int32_t* codes, const uint32_t* valid, int64_t n, int32_t bins,
                       int64_t* d_counts, int64_t* d_out_of_range, void* stream);

/* ---- a8-a10: OneHotEncoder + VectorAssembler -> `features` ---------------- */

/* Replaces OneHotEncoder (dropLast) + VectorAssembler([vec] * repeats + numeric
 * cols) + select("features") (k_means.py:38,64-68,74-77).  index[i] in
 * [0, num_labels]: num_labels (or any value outside [0, num_labels)) is the
 * "unknown" category whose one-hot block is all zeros.  Row i of the output is
 * [onehot(index[i])] * repeats ++ [(num[c][i] - shift[c]) * scale[c], c = 0..2];
 * shift / scale may be NULL (0 / 1).  Dense: out = [n, repeats * num_labels + 3]
 * row-major, out_dtype ETL_F64 (Spark's DenseVector) or ETL_F32 (the trainer's
 * dtype).  Sparse: ELL form of Spark's SparseVector - out_index / out_value =
 * [n, repeats + 3], ascending indices, index -1 = unused slot. */
ETL_API int etl_assemble_features_dense(etl_handle* h, const int32_t* index, const double* const num[3],
                                const double* shift, const double* scale, int64_t n, int32_t num_labels,
                                int32_t repeats, int32_t out_dtype, void* out, void* stream);
ETL_API int etl_assemble_features_sparse(etl_handle* h, const int32_t* index, const double* const num[3],
                                 const double* shift, const double* scale, int64_t n, int32_t num_labels,
                                 int32_t repeats, int32_t* out_index, double* out_value, void* stream);

/* ---- a11/a12: the layout the Keras trainer reads -------------------------- */

/* Replaces the row filter and array building of load_csv
 * (workloads/raw-tf/train_tf_ps.py:120-149): keep a row iff its label and all
 * three numerics are non-null (NaN cells must already be null: decode with
 * ETL_CSV_NAN_IS_NULL or call etl_nan_to_null), in input order;
 * X[k] = float32((num[c] - shift[c]) * scale[c]), y[k] = remap[label] (remap NULL:
 * the label code itself).  X = float32[n * 3] row-major, y = int32[n]; the first
 * *d_out_rows rows are written - the arrays from_tensor_slices consumes
 * (train_tf_ps.py:597,664). */
ETL_API int etl_pack_tf_features(etl_handle* h, const int32_t* label, const uint32_t* label_valid,
                         const double* const num[3], const uint32_t* const num_valid[3], const double* shift,
                         const double* scale, const int32_t* remap, int32_t remap_size, int64_t n, float* X,
                         int32_t* y, int64_t* d_out_rows, void* stream);
/* valid[i] &= !isnan(val[i]) (the `v.lower() == "nan"` rule of load_csv for
 * columns that did not come from the CSV decoder). */
ETL_API int etl_nan_to_null(etl_handle* h, const double* val, uint32_t* valid, int64_t n, void* stream);

/* ---- north-star extension: explicit groupBy aggregates (no reference site) - */

/* df.groupBy(key).agg(count(*), count(c), sum(c), avg(c), min(c), max(c)) for up
 * to 3 double columns; the reference's only groupBy is the count inside
 * StringIndexer.fit (k_means.py:34,73).  Spark semantics: a null key is its own
 * group; count(c) counts non-null cells; sum/min/max skip nulls; NaN propagates
 * through sum and is the largest value for min/max.
 * The result lives in a caller-owned open-addressing table of
 * etl_groupby_table_words(...) uint64 words: `capacity` (power of two, >= 2 x
 * the number of groups) entries plus one for the null-key group; entry layout
 * {key, rows, (nulls_c, sum_c) x ncols [, (min_c, max_c) x ncols]}, count(c) =
 * rows - nulls_c.  Key value
 * INT64_MIN is reserved.  Usage: init once, accumulate any number of batches
 * (also partial tables received from other ranks), compact.
 * etl_groupby_accumulate: keys int32/int64 with element stride key_stride; val[c]
 * double with element stride val_stride[c] (NULL = 1) - strides let it read the
 * 16-byte {key, value} records of the exchange directly.  groups_hint <= 2048
 * (0 = unknown -> 1024) selects the shared-memory kernel (per-CTA table, one
 * flush), larger the direct global-table kernel.  key_range > 0 promises int32
 * keys in [0, key_range) (dictionary codes, bucket ids; <= 8192, contiguous
 * columns, no min/max): the shared table is then indexed by the key itself; a key
 * outside the promise is still aggregated correctly, through the global table.
 * d_status (device int64[2],
 * zeroed by the caller): [0] rows dropped because the table was full, [1] rows
 * carrying the reserved key - both must be 0 for a valid result.
 * etl_groupby_compact: used entries -> out_keys/out_rows int64[out_capacity],
 * out_count int64 / out_sum,out_min,out_max double [ncols][out_capacity]
 * (column-major), in table order (sort by key to compare); *d_groups = groups
 * found (may exceed out_capacity: then only the first out_capacity were
 * written).  An all-null group has count 0, sum 0 and NaN min/max. */
ETL_API int etl_groupby_table_words(int32_t ncols, int32_t minmax, int64_t capacity, int64_t* words);
ETL_API int etl_groupby_init(etl_handle* h, uint64_t* table, int64_t capacity, int32_t ncols, int32_t minmax,
                     void* stream);
ETL_API int etl_groupby_accumulate(etl_handle* h, const void* keys, int32_t key_dtype, int64_t key_stride,
                           const uint32_t* key_valid, int32_t ncols, const double* const val[],
                           const int64_t val_stride[], const uint32_t* const val_valid[], int64_t n,
                           uint64_t* table, int64_t capacity, int32_t minmax, int64_t groups_hint,
                           int32_t key_range, int64_t* d_status, void* stream);
ETL_API int etl_groupby_compact(etl_handle* h, const uint64_t* table, int64_t capacity, int32_t ncols, int32_t minmax,
                        int64_t out_capacity, int64_t* out_keys, int64_t* out_rows, int64_t* out_count,
                        double* out_sum, double* out_min, double* out_max, int64_t* d_groups, void* stream);

/* The map side of Spark's shuffle exchange (the Exchange hashpartitioning node
 * under every aggregate of k_means.py:23-48; north-star item 4): rows (key int64,
 * value double, element strides key_stride / val_stride so {key, value} records
 * can be re-partitioned in place) are grouped by destination into one buffer of
 * 16-byte records {key, value bits}.  Two levels from one hash h = mix64(key):
 * rank = (hi32(h) * ranks) >> 32 picks the GPU, bucket = (lo32(h) * buckets)
 * >> 32 a 1/buckets slice of that GPU's aggregation table (etl_groupby_* place
 * a key's home slot at (lo32(h) * capacity) >> 32).  Destination d = rank *
 * buckets + bucket occupies records [d_part_base[d], + d_part_counts[d]).  Rank
 * ranges are the send side of an all-to-all (NCCL, torch.distributed); feeding
 * the receiver's table bucket by bucket keeps the slice being updated resident
 * in L2.  out_records: int64[2 * n], 16-byte aligned; ranks * buckets <= 256. */
ETL_API int etl_hash_partition(etl_handle* h, const int64_t* keys, int64_t key_stride, const double* vals,
                       int64_t val_stride, int64_t n, int32_t ranks, int32_t buckets, int64_t* d_part_counts,
                       int64_t* d_part_base, int64_t* out_records, void* stream);

/* The exchange itself, fused into the partition: instead of building a send
 * buffer for a separate all-to-all, the scatter pass writes every record
 * straight into its destination rank's receive buffer over NVLink - peer memory
 * mapped into this process (CUDA IPC / torch symmetric memory).  Protocol:
 *   1. etl_hash_partition_count -> this rank's counts per destination
 *      (rank r, bucket b) = r * buckets + b;
 *   2. the caller all-gathers the counts (tiny) and derives, for every
 *      destination, the first record this rank owns in rank r's receive buffer
 *      -> d_peer_offset[ranks * buckets] on the device.  With buckets > 1 a
 *      receive buffer is laid out bucket by bucket (within a bucket: source rank
 *      by source rank), i.e. it arrives already partitioned for
 *      etl_hash_subpartition / a bucket-wise merge - the receiver's own first
 *      partition pass is folded into the exchange;
 *   3. etl_hash_partition_scatter_peers: peer_records[r] (host array of `ranks`
 *      device pointers) = rank r's receive buffer as mapped here (the local
 *      buffer for r == this rank); record i of destination d = r * buckets + b
 *      lands at peer_records[r] + 2 * (d_peer_offset[d] + position);
 *   4. a cross-rank barrier after the kernel, then every rank merges what landed.
 * No reference counterpart below the Spark shuffle. */
ETL_API int etl_hash_partition_count(etl_handle* h, const int64_t* keys, int64_t key_stride, int64_t n, int32_t ranks,
                             int32_t buckets, int64_t* d_part_counts, void* stream);
ETL_API int etl_hash_partition_scatter_peers(etl_handle* h, const int64_t* keys, int64_t key_stride, const double* vals,
                                     int64_t val_stride, int64_t n, int32_t ranks, int32_t buckets,
                                     int64_t* const* peer_records, const int64_t* d_peer_offset, void* stream);

/* The reduce side of the exchange with the aggregation table in SHARED memory
 * (the final HashAggregate above the Exchange node; north-star item 4).  The
 * records a rank owns after the exchange are cut into `fine` = buckets * sub
 * pieces by fine index (lo32(h) * fine) >> 32 - small enough that the groups of
 * one piece fit one CTA's shared-memory table - and merged piece by piece:
 *   1. etl_hash_partition(ranks = 1, buckets)   -> records grouped by bucket,
 *      d_part_counts / d_part_base on the device;
 *   2. etl_hash_subpartition (only when fine > buckets): every bucket into `sub`
 *      pieces in one launch; piece f = bucket * sub + s occupies records
 *      [d_fine_base[f], + d_fine_counts[f]) of out_records.  max_bucket_records:
 *      an upper bound of the largest bucket (n is always one) - it only sizes the
 *      grid.  buckets * sub <= 65536, sub <= 256;
 *   3. etl_groupby_merge_partitioned: one piece at a time per CTA; the groups are
 *      appended (in no particular order) to out_keys / out_rows / out_count /
 *      out_sum [out_capacity] and, when out_min / out_max are given (both or
 *      neither; min(value) / max(value) with Spark's ordering: NaN is largest), to
 *      those; smem_slots = table size (power of two, 64..8192; 20 bytes of shared
 *      memory each, 36 and at most 4096 slots with min / max).  A record whose key finds no slot within
 *      the probe limit goes to spill_records (int64[2 * spill_capacity], may be
 *      the level-1 buffer, which is dead by then); the caller aggregates those
 *      with etl_groupby_accumulate - a key is never in both outputs.
 *      d_counters int64[4]: [0] groups appended, [1] records spilled (either may
 *      exceed its capacity: then the surplus was dropped and the caller must
 *      retry with more room), [3] records carrying the reserved key INT64_MIN.
 * Records carry no nulls, so count(value) = rows.  No reference counterpart below
 * the Spark shuffle. */
ETL_API int etl_hash_subpartition(etl_handle* h, const int64_t* records, int32_t buckets, int32_t sub,
                          const int64_t* d_bucket_counts, const int64_t* d_bucket_base, int64_t max_bucket_records,
                          int64_t* d_fine_counts, int64_t* d_fine_base, int64_t* out_records, void* stream);
ETL_API int etl_groupby_merge_partitioned(etl_handle* h, const int64_t* records, int32_t fine,
                                  const int64_t* d_fine_counts, const int64_t* d_fine_base, int32_t smem_slots,
                                  int64_t out_capacity, int64_t* out_keys, int64_t* out_rows, int64_t* out_count,
                                  double* out_sum, double* out_min, double* out_max, int64_t* spill_records,
                                  int64_t spill_capacity, int64_t* d_counters, void* stream);

/* The same exchange and the same two partition levels WITHOUT their count passes
 * (the Exchange + final HashAggregate under k_means.py:23-48; north-star item 4;
 * no reference counterpart below the Spark shuffle).  Every destination of a
 * scatter owns a fixed-capacity SLAB of a record buffer and a cursor word that
 * persists across calls (caller-owned, zeroed by the caller before the first
 * call), so rows can be scattered chunk by chunk: a sender streams its table over
 * NVLink in chunks while the receivers already partition what has arrived.
 *   etl_slab_partition_rows: destination d = rank * buckets + bucket (as in
 *     etl_hash_partition) writes into dest_records[rank] at record index
 *     slab_origin + d * slab_stride + position, position < slab_capacity.  For a
 *     local first level: ranks = 1, one buffer, slab_stride = slab_capacity.  For
 *     the exchange: buckets = 1, dest_records[r] = rank r's receive buffer as
 *     mapped here, slab_origin = (this rank) * slab_capacity, slab_stride = 0.
 *     Rows: two contiguous 16-byte aligned columns (strides 1) or {key, value}
 *     records (strides 2, vals == keys + 1).  max_ctas > 0 caps the grid so a
 *     link-bound exchange leaves SMs to the kernels it overlaps with.
 *   etl_slab_partition_records: input = `segments` slabs of `records`, seg_capacity
 *     records apart; segment s holds the records [min(d_seg_prev[s], cap),
 *     min(d_seg_counts[s], cap)) (d_seg_prev NULL = from 0).  sub == 0: all segments
 *     scatter into the same `buckets` slabs of out_records (the receiver's first
 *     level over what each source has delivered since d_seg_prev).  sub > 0:
 *     segments == buckets, segment b is cut into pieces b * sub .. b * sub + sub - 1
 *     of `buckets * sub` (the second level); d_cursor has buckets * sub words.
 *     max_seg_records only sizes the grid.
 *     sources > 0 (with sub > 0): the input is what a BUCKETED exchange delivered -
 *     slabs laid out [source][bucket], `buckets` per source; the call cuts buckets
 *     bucket0 .. bucket0 + segments / sources - 1 (segments = that range x sources),
 *     all sources of a bucket scattering into that bucket's `sub` pieces.
 *   etl_groupby_merge_slabs: etl_groupby_merge_partitioned over slabs: piece f =
 *     records [f * piece_capacity, + min(d_cursor[f], piece_capacity)), for the
 *     pieces piece0 .. piece0 + pieces - 1; `append` is a flag word: bit 0 keeps
 *     d_counters and appends to the groups already in the output (bucket groups
 *     merged one after another while later ones are still crossing NVLink); bit 1
 *     says the launch shares the SMs with a concurrent transfer kernel
 *     (etl_peer_copy_slabs): it then runs 512-thread CTAs, which leave registers
 *     for the copy CTAs, instead of the 640-thread CTAs a launch on its own gets.
 *   etl_peer_publish_counts: peer_tables[r][slot + j] = d_cursor[r * per_rank + j]
 *     for every rank r and j < per_rank (system-scope release stores; follow with a
 *     cross-rank barrier) - how a receiver learns how much of each source's slabs is
 *     filled.  ranks * per_rank <= 256.
 *   etl_peer_copy_slabs: the bucketed exchange's data movement.  The send buffer
 *     holds ranks x buckets slabs, destination-rank major (etl_slab_partition_rows
 *     with ranks > 1 and ONE local buffer); the filled part of the slabs of
 *     buckets bucket0 .. bucket0 + nbuckets - 1 of every destination r is stored
 *     into peer_records[r] at slab (my_rank * buckets + bucket) - the receivers'
 *     [source][bucket] layout - with 16-byte stores over NVLink (~700 GB/s per GPU
 *     measured; no shared memory, so it runs beside the merge of earlier groups).
 *   etl_peer_copy_slabs_bulk: the same movement, same arguments, through the TMA
 *     engine: one thread per CTA streams its share of a slab through a 48 KB ring
 *     of shared memory (bulk async load from the send buffer, bulk async store into
 *     the owner's buffer; no tensor map).  ctas_per_slab 0 = about one CTA per SM.
 *   etl_copy_async: stream-ordered device-to-device copy through a copy engine
 *     (dst may be a peer's memory); an alternative mover for the same slabs
 *     (measured slower at 8 GPUs: 330-400 GB/s per GPU).
 * Overflow: a record that finds its slab full is counted in *d_overflow (int64,
 * caller-zeroed) and DROPPED.  A non-zero count means the result is incomplete:
 * repeat on the counted path (etl_hash_partition / etl_hash_subpartition), which
 * sizes every destination exactly.  With slabs of mean + 6 sigma records that
 * only happens for heavily skewed keys. */
ETL_API int etl_slab_partition_rows(etl_handle* h, const int64_t* keys, int64_t key_stride, const double* vals,
                            int64_t val_stride, int64_t n, int32_t ranks, int32_t buckets,
                            int64_t* const* dest_records, int64_t slab_origin, int64_t slab_stride,
                            int64_t slab_capacity, int64_t* d_cursor, int64_t* d_overflow, int32_t max_ctas,
                            void* stream);
ETL_API int etl_slab_partition_records(etl_handle* h, const int64_t* records, int32_t segments, int64_t seg_capacity,
                               const int64_t* d_seg_counts, const int64_t* d_seg_prev, int64_t max_seg_records,
                               int32_t buckets, int32_t sub, int32_t sources, int32_t bucket0, int64_t* out_records,
                               int64_t slab_capacity, int64_t* d_cursor, int64_t* d_overflow, int32_t max_ctas,
                               void* stream);
ETL_API int etl_groupby_merge_slabs(etl_handle* h, const int64_t* records, int32_t piece0, int32_t pieces, int32_t append,
                            const int64_t* d_cursor, int64_t piece_capacity, int32_t smem_slots, int64_t out_capacity,
                            int64_t* out_keys, int64_t* out_rows, int64_t* out_count, double* out_sum, double* out_min,
                            double* out_max, int64_t* spill_records, int64_t spill_capacity, int64_t* d_counters,
                            void* stream);
ETL_API int etl_peer_publish_counts(etl_handle* h, const int64_t* d_cursor, int32_t ranks, int32_t per_rank,
                            int64_t* const* peer_tables, int64_t slot, void* stream);
ETL_API int etl_copy_async(etl_handle* h, void* dst, const void* src, int64_t nbytes, void* stream);
ETL_API int etl_peer_copy_slabs(etl_handle* h, const int64_t* send_records, int64_t* const* peer_records, int32_t ranks,
                        int32_t my_rank, int32_t buckets, int32_t bucket0, int32_t nbuckets, int64_t slab_capacity,
                        const int64_t* d_cursor, int32_t ctas_per_slab, void* stream);
ETL_API int etl_peer_copy_slabs_bulk(etl_handle* h, const int64_t* send_records, int64_t* const* peer_records, int32_t ranks,
                             int32_t my_rank, int32_t buckets, int32_t bucket0, int32_t nbuckets, int64_t slab_capacity,
                             const int64_t* d_cursor, int32_t ctas_per_slab, void* stream);

/* Diagnostic, no reference counterpart (tools/nvlink_probe.py): copy nbytes (a
 * multiple of 16) from src to dst - dst may be a peer GPU's memory mapped here -
 * with 16-byte stores from registers (mode 0, what the exchange scatter issues) or
 * with bulk async stores of `chunk` bytes from shared memory (mode 1; TMA engine,
 * no multicast).  ctas <= 0: four CTAs per SM. */
ETL_API int etl_peer_write_probe(etl_handle* h, const void* src, void* dst, int64_t nbytes, int32_t mode, int32_t chunk,
                         int32_t ctas, void* stream);

/* ---- north-star extension: bucketing (no reference site) ------------------ */

/* Spark ML Bucketizer(handleInvalid="keep"): out[i] = b with splits[b] <= x <
 * splits[b+1] (the last bucket also takes x == splits[last]); null or NaN ->
 * nsplits - 1 (the extra bucket); x outside [splits[0], splits[last]] -> -1 and
 * counted in *d_out_of_range (Spark raises).  d_splits: device double[nsplits],
 * strictly increasing, 3 <= nsplits <= 4096. */
ETL_API int etl_bucketize(etl_handle* h, const double* val, const uint32_t* valid, int64_t n, const double* d_splits,
                  int32_t nsplits, int32_t* out, int64_t* d_out_of_range, void* stream);

/* ---- next row: KMeans on `features` ---------------- */

/* Replaces the distributed passes inside KMeans().setK(25).setSeed(1)
 * .setMaxIter(1000).fit(dataset) (k_means.py:83-87): the closest-centre search and
 * the per-cluster sums of one Lloyd iteration (and of the k-means|| init).  The
 * feature rows are not materialised: row i = [onehot(index[i])] * repeats ++
 * num[0..2][i], index in [0, num_labels] (num_labels = unknown label, all-zero
 * block).  centers: double[k][repeats * num_labels + 3].
 * etl_kmeans_prepare: per-centre tables -> d_onehot_sum[k][num_labels + 1],
 *   d_num[k][3], d_norm2[k].
 * etl_kmeans_assign: squared distance |x|^2 - 2 x.c + |c|^2 (clamped at 0) to
 *   every centre; out_cost[i] = the minimum (min-combined with d_prev_cost[i] if
 *   given), out_cluster[i] = cluster_offset + argmin (lowest id on ties; left
 *   untouched where the previous cost was not beaten).  Either output may be NULL.
 * etl_kmeans_accumulate: d_count[k], d_hist[k][num_labels + 1] (rows per cluster
 *   and label) and d_numsum[k][3]; new centre = hist / count per label (copied
 *   into each of the `repeats` blocks) ++ numsum / count. */
ETL_API int etl_kmeans_prepare(etl_handle* h, const double* centers, int32_t k, int32_t num_labels, int32_t repeats,
                       double* d_onehot_sum, double* d_num, double* d_norm2, void* stream);
ETL_API int etl_kmeans_assign(etl_handle* h, const int32_t* index, const double* const num[3], int64_t n, int32_t k,
                      int32_t num_labels, int32_t repeats, const double* d_onehot_sum, const double* d_num,
                      const double* d_norm2, const double* d_prev_cost, int32_t cluster_offset,
                      int32_t* out_cluster, double* out_cost, void* stream);
ETL_API int etl_kmeans_accumulate(etl_handle* h, const int32_t* index, const double* const num[3],
                          const int32_t* cluster, int64_t n, int32_t k, int32_t num_labels, int64_t* d_count,
                          int64_t* d_hist, double* d_numsum, void* stream);
/* The centre update that follows (Spark does it on the driver): centers[c] <- hist[c] / count[c] per label,
 * copied into each of the `repeats` blocks, ++ numsum[c] / count[c]; a cluster with count 0 keeps its
 * centre.  centers: double[k][repeats * num_labels + 3], updated in place on the device. */
ETL_API int etl_kmeans_update(etl_handle* h, const int64_t* d_count, const int64_t* d_hist, const double* d_numsum,
                      int32_t k, int32_t num_labels, int32_t repeats, double* centers, void* stream);

/* ClusteringEvaluator().evaluate(predictions) - silhouette with squared Euclidean distance
 * (spark_workload_to_cloud_k8s.py:139-144), Spark's O(n k) formulation: with N_G rows, Psi_G = sum of
 * |y|^2 and Y_G = vector sum of cluster G, the mean squared distance from x to G is
 * |x|^2 + Psi_G / N_G - 2 x.Y_G / N_G; a = that for x's own cluster times N / (N - 1), b = the minimum over
 * the other clusters, coefficient = 1 - a / b (a < b), b / a - 1 (a > b), 0 (equal, or a cluster of one).
 * etl_kmeans_row_norms: out[i] = |x_i|^2 (its per-cluster sums Psi come from etl_kmeans_accumulate).
 * etl_kmeans_silhouette: out[i] = the coefficient of row i; d_label_sum[k][num_labels + 1] = repeats * rows of
 *   the cluster with that label (column num_labels = 0), d_num_sum[k][3], d_psi[k], d_count[k] (rows as
 *   double, 0 = no such cluster).  The evaluator's value is the mean of out (etl_column_sum_count). */
ETL_API int etl_kmeans_row_norms(etl_handle* h, const int32_t* index, const double* const num[3], int64_t n,
                         int32_t num_labels, int32_t repeats, double* out, void* stream);
ETL_API int etl_kmeans_silhouette(etl_handle* h, const int32_t* index, const double* const num[3], const int32_t* cluster,
                          int64_t n, int32_t k, int32_t num_labels, int32_t repeats, const double* d_label_sum,
                          const double* d_num_sum, const double* d_psi, const double* d_count, double* out,
                          void* stream);

/* ---- next row: device-resident hand-off to the trainer */

/* Replaces model.fit / the custom train step on the tabular model of
 * workloads/raw-tf/train_tf_ps.py:328-343 (Dense 16-32-64 relu, softmax output,
 * sparse categorical cross-entropy, Adam) fed by from_tensor_slices((X, y))
 * .shuffle().batch().repeat() (train_tf_ps.py:596-601,664-672): X float32[n][3]
 * and y int32[n] stay where etl_pack_tf_features left them, in device memory.
 * Parameters: one float32 blob of etl_mlp_param_count(classes) values laid out
 * (W1[3][16], b1, W2[16][32], b2, W3[32][64], b3, W4[64][classes], b4), kernels
 * [in][out] row-major as Keras stores them.
 * etl_mlp_train_steps runs `steps` consecutive optimiser steps in ONE launch
 * (parameters and Adam moments live in shared memory for its duration): step s
 * trains on rows d_order[s * batch .. + batch); d_params / d_adam_m / d_adam_v are
 * updated in place; step0 = optimiser steps already taken (Adam bias correction);
 * d_loss[s] / d_acc[s] = mean loss / accuracy of step s's batch before its update.
 * etl_mlp_train_steps_ragged: the same with d_step_rows[s] (1..batch, device
 *   int32[steps]; NULL = batch) rows of step s counting - batch() before repeat()
 *   ends every pass over the data with one short batch (train_tf_ps.py:664-665),
 *   which model.fit consumes as an ordinary step; d_order stays [steps][batch].
 * etl_mlp_evaluate: forward only - per-sample loss (needs y) and / or arg-max class. */
ETL_API int etl_mlp_param_count(int32_t classes);
ETL_API int etl_mlp_train_steps(etl_handle* h, const float* X, const int32_t* y, int64_t n, const int32_t* d_order,
                        int32_t batch, int32_t steps, int32_t classes, float* d_params, float* d_adam_m,
                        float* d_adam_v, int64_t step0, float lr, float beta1, float beta2, float eps,
                        float* d_loss, float* d_acc, void* stream);
ETL_API int etl_mlp_train_steps_ragged(etl_handle* h, const float* X, const int32_t* y, int64_t n, const int32_t* d_order,
                               const int32_t* d_step_rows, int32_t batch, int32_t steps, int32_t classes, float* d_params,
                               float* d_adam_m, float* d_adam_v, int64_t step0, float lr, float beta1, float beta2,
                               float eps, float* d_loss, float* d_acc, void* stream);
ETL_API int etl_mlp_evaluate(etl_handle* h, const float* X, const int32_t* y, int64_t n, const float* d_params,
                     int32_t classes, float* d_loss, int32_t* d_pred, void* stream);

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
