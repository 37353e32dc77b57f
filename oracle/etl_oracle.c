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
