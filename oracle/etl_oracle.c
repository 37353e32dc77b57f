/* CPU oracle, C restatement of the hot loops.  TEST INFRASTRUCTURE ONLY.
 *
 * Same semantics as oracle/etl_oracle.py (which cites the reference call
 * sites); this file exists so that the CPU baseline bench.py reports is compiled
 * code using every host core (pthreads; the image has no libgomp), not a NumPy
 * script.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  Built by oracle/Makefile into
 * oracle/liboracle.so.
 *
 * Parity status: the Spark-side loops here are "parity unpinned" (no PySpark to
 * run; see the header of etl_oracle.py); tests/test_oracle_c.py checks them
 * against the NumPy restatement on seeded inputs.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#define API __attribute__((visibility("default")))
#define MAX_THREADS 256

static int g_threads = 0;

API int oracle_threads(void) {
    if (g_threads <= 0) {
        long n = sysconf(_SC_NPROCESSORS_ONLN);
        g_threads = n < 1 ? 1 : (n > MAX_THREADS ? MAX_THREADS : (int)n);
    }
    return g_threads;
}
API void oracle_set_threads(int n) { g_threads = n < 1 ? 1 : (n > MAX_THREADS ? MAX_THREADS : n); }

typedef void (*body_fn)(int t, int nt, void* ctx);
typedef struct { body_fn fn; int t, nt; void* ctx; } task_t;
static void* trampoline(void* a) { task_t* k = (task_t*)a; k->fn(k->t, k->nt, k->ctx); return NULL; }

/* run fn(t, nt, ctx) for t in [0, nt) on nt threads (thread 0 = the caller) */
static void run_parallel(body_fn fn, void* ctx) {
    int nt = oracle_threads();
    pthread_t th[MAX_THREADS];
    task_t task[MAX_THREADS];
    for (int t = 0; t < nt; ++t) { task[t].fn = fn; task[t].t = t; task[t].nt = nt; task[t].ctx = ctx; }
    for (int t = 1; t < nt; ++t) pthread_create(&th[t], NULL, trampoline, &task[t]);
    fn(0, nt, ctx);
    for (int t = 1; t < nt; ++t) pthread_join(th[t], NULL);
}

static inline int bit(const uint32_t* bm, int64_t i) { return bm ? (bm[i >> 5] >> (i & 31)) & 1 : 1; }

static inline double load_as_double(const void* p, int dtype, int64_t i) {
    switch (dtype) {
        case 0: return ((const double*)p)[i];
        case 1: return (double)((const float*)p)[i];
        default: return (double)((const int32_t*)p)[i];
    }
}

/* rows [lo, hi) of thread t, split on 32-row boundaries so bitmap words are not shared */
static inline void row_range(int64_t n, int t, int nt, int64_t* lo, int64_t* hi) {
    int64_t words = (n + 31) >> 5;
    *lo = (words * t / nt) << 5;
    *hi = (words * (t + 1) / nt) << 5;
    if (*hi > n) *hi = n;
    if (*lo > n) *lo = n;
}

/* ---- k_means.py:23 ------------------------------------------------------ */
typedef struct { const uint32_t* valid; int64_t n; int64_t part[MAX_THREADS]; } nullcnt_t;
static void nullcnt_body(int t, int nt, void* c) {
    nullcnt_t* a = (nullcnt_t*)c;
    int64_t lo, hi, z = 0;
    row_range(a->n, t, nt, &lo, &hi);
    for (int64_t i = lo; i < hi; ++i) z += !bit(a->valid, i);
    a->part[t] = z;
}
API int64_t oracle_null_count(const uint32_t* valid, int64_t n) {
    if (!valid) return 0;
    nullcnt_t a;
    a.valid = valid; a.n = n;
    run_parallel(nullcnt_body, &a);
    int64_t z = 0;
    for (int t = 0; t < oracle_threads(); ++t) z += a.part[t];
    return z;
}

/* ---- k_means.py:45-48: (sum, count) over row_valid & ~isnan & isNotNull ----
 * Each thread sums a contiguous block with Neumaier compensation; blocks are
 * added in order, so the result is fixed for a given thread count. */
typedef struct {
    const void* val; const uint32_t* val_valid; int dtype; const uint32_t* row_valid; int64_t n;
    double ps[MAX_THREADS]; int64_t pc[MAX_THREADS];
} sum_t;
static void sum_body(int t, int nt, void* c) {
    sum_t* a = (sum_t*)c;
    int64_t lo, hi, k = 0;
    row_range(a->n, t, nt, &lo, &hi);
    double s = 0.0, comp = 0.0;
    for (int64_t i = lo; i < hi; ++i) {
        if (!bit(a->row_valid, i) || !bit(a->val_valid, i)) continue;
        double x = load_as_double(a->val, a->dtype, i);
        if (x != x) continue;
        double ts = s + x;
        comp += fabs(s) >= fabs(x) ? (s - ts) + x : (x - ts) + s;
        s = ts;
        ++k;
    }
    a->ps[t] = s + comp;
    a->pc[t] = k;
}
API void oracle_sum_count(const void* val, const uint32_t* val_valid, int dtype, const uint32_t* row_valid, int64_t n,
                          double* out_sum, int64_t* out_count) {
    sum_t a;
    a.val = val; a.val_valid = val_valid; a.dtype = dtype; a.row_valid = row_valid; a.n = n;
    run_parallel(sum_body, &a);
    double s = 0.0;
    int64_t k = 0;
    for (int t = 0; t < oracle_threads(); ++t) { s += a.ps[t]; k += a.pc[t]; }
    *out_sum = s;
    *out_count = k;
}

/* ---- k_means.py:27-28 + :49-51 ------------------------------------------
 * keep rows whose key is non-null, impute null/NaN numerics with fill[c], cast
 * to double, keep input order.  Pass 1: kept rows per thread block; pass 2:
 * each thread writes its block at its offset. */
typedef struct {
    const int32_t* key_code; const uint32_t* key_valid; const void* const* val; const uint32_t* const* val_valid;
    const int32_t* val_dtype; const double* fill; int64_t n; int32_t* out_code; double* const* out_val;
    int64_t off[MAX_THREADS + 1];
} proj_t;
static void proj_count_body(int t, int nt, void* c) {
    proj_t* a = (proj_t*)c;
    int64_t lo, hi, k = 0;
    row_range(a->n, t, nt, &lo, &hi);
    if (!a->key_valid) k = hi - lo;
    else for (int64_t i = lo; i < hi; ++i) k += bit(a->key_valid, i);
    a->off[t + 1] = k;
}
static void proj_write_body(int t, int nt, void* c) {
    proj_t* a = (proj_t*)c;
    int64_t lo, hi;
    row_range(a->n, t, nt, &lo, &hi);
    int64_t o = a->off[t];
    for (int64_t i = lo; i < hi; ++i) {
        if (!bit(a->key_valid, i)) continue;
        a->out_code[o] = a->key_code[i];
        for (int k = 0; k < 3; ++k) {
            double x = load_as_double(a->val[k], a->val_dtype[k], i);
            int bad = !bit(a->val_valid[k], i) || x != x;
            a->out_val[k][o] = bad ? a->fill[k] : x;
        }
        ++o;
    }
}
API int64_t oracle_project_filter_impute(const int32_t* key_code, const uint32_t* key_valid, const void* const val[3],
                                         const uint32_t* const val_valid[3], const int32_t val_dtype[3],
                                         const double fill[3], int64_t n, int32_t* out_code, double* const out_val[3]) {
    proj_t a;
    a.key_code = key_code; a.key_valid = key_valid; a.val = val; a.val_valid = val_valid; a.val_dtype = val_dtype;
    a.fill = fill; a.n = n; a.out_code = out_code; a.out_val = out_val;
    a.off[0] = 0;
    run_parallel(proj_count_body, &a);
    int nt = oracle_threads();
    for (int t = 0; t < nt; ++t) a.off[t + 1] += a.off[t];
    run_parallel(proj_write_body, &a);
    return a.off[nt];
}

/* ---- synthetic columns (pyspark_tf_gke_b200/synth.py, same definition) ----
 * bits(seed, stream, row) = mix64(key + (row + 1) * GOLDEN); kind 0 = integer-valued
 * double in [0, 1e5], kind 3 = int64 key in [0, modulus).  No nulls (config 4). */
static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
typedef struct { uint64_t key; int kind; uint64_t modulus; int64_t row0, n; void* out; } synth_t;
static void synth_body(int t, int nt, void* c) {
    synth_t* a = (synth_t*)c;
    int64_t lo = a->n * t / nt, hi = a->n * (t + 1) / nt;
    for (int64_t i = lo; i < hi; ++i) {
        uint64_t b = mix64(a->key + (uint64_t)(a->row0 + i + 1) * 0x9E3779B97F4A7C15ull);
        if (a->kind == 0) ((double*)a->out)[i] = floor((double)(b >> 11) * (1.0 / 9007199254740992.0) * 100001.0);
        else ((int64_t*)a->out)[i] = (int64_t)(b % a->modulus);
    }
}
API void oracle_synth_column(uint64_t seed, uint32_t stream_id, int kind, uint64_t modulus, int64_t row0, int64_t n,
                             void* out) {
    synth_t a;
    a.key = (seed * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)(2 * stream_id) * 0xBF58476D1CE4E5B9ull);
    a.kind = kind; a.modulus = modulus; a.row0 = row0; a.n = n; a.out = out;
    run_parallel(synth_body, &a);
}

/* ---- groupBy(key).agg(count, sum(value)) on (int64 key, double value) rows ----
 * Spark's HashAggregate restated (partial aggregation per partition is skipped:
 * with ~10 rows per group it only adds work).  The rows are cut into NPART
 * partitions by the high bits of mix64(key) - the Exchange - and every partition
 * is aggregated in an open-addressing table by one thread at a time.  Inside a
 * partition the rows keep their input order, so a group's sum is accumulated in
 * input order, like a single-partition Spark run.  Output order is unspecified
 * (as in Spark); callers sort.  No null keys / values (config 4 has none). */
#define GB_NPART 1024
typedef struct { int64_t key; double val; } gb_rec_t;
typedef struct {
    const int64_t* keys; const double* vals; int64_t n;
    int64_t* hist;            /* [nt][GB_NPART] */
    int64_t* base;            /* [GB_NPART + 1] */
    gb_rec_t* rec;            /* [n] partitioned */
    int64_t* out_keys; int64_t* out_count; double* out_sum;   /* [n]: partition p's groups start at base[p] */
    int64_t groups[GB_NPART];
    int next;                 /* dynamic partition counter */
    int failed;
} gb_t;
static inline int gb_part(int64_t key) { return (int)(mix64((uint64_t)key) >> 54); }   /* top 10 bits */
static void gb_hist_body(int t, int nt, void* c) {
    gb_t* a = (gb_t*)c;
    int64_t lo = a->n * t / nt, hi = a->n * (t + 1) / nt;
    int64_t* h = a->hist + (size_t)t * GB_NPART;
    memset(h, 0, sizeof(int64_t) * GB_NPART);
    for (int64_t i = lo; i < hi; ++i) h[gb_part(a->keys[i])]++;
}
static void gb_scatter_body(int t, int nt, void* c) {
    gb_t* a = (gb_t*)c;
    int64_t lo = a->n * t / nt, hi = a->n * (t + 1) / nt;
    int64_t* h = a->hist + (size_t)t * GB_NPART;   /* now: this thread's first slot in every partition */
    for (int64_t i = lo; i < hi; ++i) {
        gb_rec_t r = { a->keys[i], a->vals[i] };
        a->rec[h[gb_part(r.key)]++] = r;
    }
}
static void gb_agg_body(int t, int nt, void* c) {
    (void)t; (void)nt;
    gb_t* a = (gb_t*)c;
    int64_t cap = 0;
    int64_t* tk = NULL; int64_t* tslot = NULL;
    for (;;) {
        int p = __atomic_fetch_add(&a->next, 1, __ATOMIC_RELAXED);
        if (p >= GB_NPART) break;
        int64_t lo = a->base[p], cnt = a->base[p + 1] - lo;
        if (cnt == 0) { a->groups[p] = 0; continue; }
        int64_t want = 64;
        while (want < 2 * cnt) want <<= 1;
        if (want > cap) {
            free(tk); free(tslot);
            tk = (int64_t*)malloc(sizeof(int64_t) * want);
            tslot = (int64_t*)malloc(sizeof(int64_t) * want);
            cap = want;
            if (!tk || !tslot) { a->failed = 1; break; }
        }
        memset(tslot, 0xff, sizeof(int64_t) * want);   /* -1 = empty; tk holds the key of a used slot */
        int64_t g = 0;
        const uint64_t mask = (uint64_t)want - 1;
        for (int64_t i = 0; i < cnt; ++i) {
            const gb_rec_t r = a->rec[lo + i];
            uint64_t s = (mix64((uint64_t)r.key) >> 8) & mask;
            for (;;) {
                if (tslot[s] < 0) {
                    tslot[s] = g; tk[s] = r.key;
                    a->out_keys[lo + g] = r.key; a->out_count[lo + g] = 1; a->out_sum[lo + g] = r.val;
                    ++g;
                    break;
                }
                if (tk[s] == r.key) { int64_t q = lo + tslot[s]; a->out_count[q]++; a->out_sum[q] += r.val; break; }
                s = (s + 1) & mask;
            }
        }
        a->groups[p] = g;
    }
    free(tk); free(tslot);
}
/* -> number of groups, or -1 when memory ran out.  out_* need room for n entries
 * (groups are first written per partition, then compacted to the front). */
API int64_t oracle_groupby_sum(const int64_t* keys, const double* vals, int64_t n, int64_t* out_keys,
                               int64_t* out_count, double* out_sum) {
    if (n <= 0) return 0;
    gb_t a;
    int nt = oracle_threads();
    a.keys = keys; a.vals = vals; a.n = n; a.out_keys = out_keys; a.out_count = out_count; a.out_sum = out_sum;
    a.next = 0; a.failed = 0;
    a.hist = (int64_t*)malloc(sizeof(int64_t) * (size_t)nt * GB_NPART);
    a.base = (int64_t*)malloc(sizeof(int64_t) * (GB_NPART + 1));
    a.rec = (gb_rec_t*)malloc(sizeof(gb_rec_t) * (size_t)n);
    if (!a.hist || !a.base || !a.rec) { free(a.hist); free(a.base); free(a.rec); return -1; }
    run_parallel(gb_hist_body, &a);
    int64_t run = 0;
    for (int p = 0; p < GB_NPART; ++p) {
        a.base[p] = run;
        for (int t = 0; t < nt; ++t) { int64_t c = a.hist[(size_t)t * GB_NPART + p]; a.hist[(size_t)t * GB_NPART + p] = run; run += c; }
    }
    a.base[GB_NPART] = run;
    run_parallel(gb_scatter_body, &a);
    run_parallel(gb_agg_body, &a);
    int64_t g = -1;
    if (!a.failed) {
        g = 0;
        for (int p = 0; p < GB_NPART; ++p) {
            int64_t lo = a.base[p], c = a.groups[p];
            if (lo != g && c) {
                memmove(out_keys + g, out_keys + lo, sizeof(int64_t) * c);
                memmove(out_count + g, out_count + lo, sizeof(int64_t) * c);
                memmove(out_sum + g, out_sum + lo, sizeof(double) * c);
            }
            g += c;
        }
    }
    free(a.hist); free(a.base); free(a.rec);
    return g;
}
