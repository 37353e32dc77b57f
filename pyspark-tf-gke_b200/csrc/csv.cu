// CSV text -> device-resident columnar buffers.
// Replaces spark.read.csv(path, header=True, inferSchema=True)
// (spark_workload_to_cloud_k8s.py:48) and the csv.DictReader loop of load_csv
// (train_tf_ps.py:114-137) for the columns the ETL stage uses.
//
// Quoting is RFC 4180 (quote ", doubled-quote escape; a quoted field may hold
// commas and newlines); backslash escapes are outside the documented domain.
//
// Four steps, all byte-parallel until the field parse:
//   1. csv_block_summary_kernel  per 16 KiB block (staged in shared memory by a bulk async copy):
//      every thread builds the quote / newline / comma masks of its 64-byte window; the block's
//      summary = parity of the quote count and the number of record starts assuming the block
//      begins outside / inside quotes (a record start = the byte after a newline that is outside
//      quotes, unless the line it opens is blank - Spark and csv.reader drop blank lines).  The
//      window masks (32 B per 64 B of text) are KEPT in the handle's arena for steps 3 and 4.
//   2. csv_span_reduce / _scan / _apply  compose the block summaries in order
//      (an associative operator on (parity, starts-outside, starts-inside))
//      -> every block's true starting quote state and first row number
//   3. csv_row_offsets_kernel    reads the kept masks (not the text), applies the true state and
//      writes row_start[row] (int64 byte offsets)
//   4. csv_parse_rows_kernel     one warp per 32 rows, lane r owns row r: (A) it walks the row over
//      the kept masks, a 64-byte window at a time (prefix xor for the quote state, ffs over the
//      unquoted-comma mask; without the masks it scans the text 8 bytes at a time with SWAR byte
//      compares) and leaves every requested field's bounds in shared memory; (B) it converts those
//      fields: numerals through dec_parse (correctly rounded), strings through an open-addressing
//      dictionary in global memory keyed by a 64-bit hash (8 bytes per step) with a byte-exact
//      collision check.  Each column leaves as one coalesced store plus one ballot-built validity
//      word per 32 rows.
#include "common.cuh"
#include "decimal.cuh"

namespace etl {

constexpr int CSV_THREADS = 256;
constexpr int CSV_BYTES_PER_THREAD = 64;
constexpr int CSV_BLOCK_BYTES = CSV_THREADS * CSV_BYTES_PER_THREAD;  // 16 KiB
constexpr int CSV_MAX_COLS = 8;      // requested columns per parse call
constexpr int CSV_MAX_FIELDS = 256;  // field positions that can be requested

struct CsvBlockSummary {
    uint32_t parity;  // quote count of the block & 1
    uint32_t pad;
    uint64_t starts_if_outside;  // record starts if the block begins outside quotes
    uint64_t starts_if_inside;
    uint64_t reserved;
};
struct CsvBlockStart {
    uint64_t first_row;
    uint32_t inside;  // 1 = block begins inside a quoted field
    uint32_t pad;
};

// 4-bit mask of the bytes of w equal to the byte replicated in `pat`
__device__ __forceinline__ unsigned eq4(uint32_t w, uint32_t pat) {
    const uint32_t t = w ^ pat;   // exact zero-byte flags (bit 7 of each matching byte), then one multiply gathers them
    const uint32_t f = ~(((t & 0x7f7f7f7fu) + 0x7f7f7f7fu) | t) & 0x80808080u;
    return ((f >> 7) * 0x01020408u) >> 24;
}
// does w hold a byte equal to the byte replicated in pat?  (bit 7 of some byte of the result is set)
__device__ __forceinline__ uint32_t has_byte(uint32_t w, uint32_t pat) {
    const uint32_t t = w ^ pat;
    return (t - 0x01010101u) & ~t;
}

// bit 7 of every byte of w that equals the byte replicated in pat (exact per byte)
__device__ __forceinline__ uint32_t eq_flags(uint32_t w, uint32_t pat) {
    const uint32_t t = w ^ pat;
    return ~(((t & 0x7f7f7f7fu) + 0x7f7f7f7fu) | t) & 0x80808080u;
}
// those four flags as a 4-bit mask: one multiply moves bits 0 / 8 / 16 / 24 to 24..27
__device__ __forceinline__ uint32_t flags_nibble(uint32_t f) { return ((f >> 7) * 0x01020408u) >> 24; }

// The index passes stage the text in shared memory: a CTA's block of CSV_BLOCK_BYTES (+16 bytes of
// look-ahead) arrives by one bulk async copy, double buffered, while the previous block is
// processed; every thread then reads ITS 64-byte window from shared memory.  (Read straight from
// global memory, the 64-byte-per-thread windows make every 16-byte load of a warp hit 32 different
// sectors: ncu showed 1.7x the text size in DRAM reads and 31 % L2 hits.)
constexpr int CSV_STAGE_BYTES = CSV_BLOCK_BYTES + 128;   // look-ahead + padding that keeps both buffers 128-byte aligned
constexpr size_t CSV_STAGE_SMEM = 2 * CSV_STAGE_BYTES + 16;

// thread 0: start the copy of block `blk` into `dst`; the < 16 bytes a bulk copy cannot move go by hand
__device__ __forceinline__ void csv_stage_issue(const uint8_t* __restrict__ text, int64_t n, int64_t blk, unsigned char* dst,
                                                uint64_t* bar) {
    const int64_t base = blk * CSV_BLOCK_BYTES;
    int64_t want = n - base;
    if (want > CSV_BLOCK_BYTES + 16) want = CSV_BLOCK_BYTES + 16;
    const unsigned bulk = (unsigned)(want & ~(int64_t)15);
    for (int64_t i = bulk; i < want; ++i) dst[i] = text[base + i];
    mbar_arrive_expect_tx(bar, bulk);
    if (bulk) bulk_load(dst, text + base, bulk, bar);
}

// csv_masks64 on a staged block: `win` = the thread's window in shared memory, `base` its text offset
__device__ __forceinline__ void csv_masks64_staged(const unsigned char* win, int64_t base, int64_t n, uint64_t* quotes,
                                                   uint64_t* newlines, uint64_t* opens, uint64_t* commas) {
    uint64_t q = 0, nl = 0, cr = 0, cm = 0;
    uint8_t a = 0, b = 0;
    if (base + 66 <= n) {
        uint32_t mq[4], mn[4], mc[4], any_cr = 0;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const uint4 x = *reinterpret_cast<const uint4*>(win + 16 * v);
            const uint32_t w[4] = {x.x, x.y, x.z, x.w};
            mq[v] = 0;
            mn[v] = 0;
            mc[v] = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                mq[v] |= flags_nibble(eq_flags(w[k], 0x22222222u)) << (4 * k);
                mn[v] |= flags_nibble(eq_flags(w[k], 0x0a0a0a0au)) << (4 * k);
                mc[v] |= flags_nibble(eq_flags(w[k], 0x2c2c2c2cu)) << (4 * k);
                const uint32_t t = w[k] ^ 0x0d0d0d0du;
                any_cr |= (t - 0x01010101u) & ~t;   // bit 7 of some byte set <=> the word holds a '\r'
            }
        }
        q = (uint64_t)(mq[0] | (mq[1] << 16)) | ((uint64_t)(mq[2] | (mq[3] << 16)) << 32);
        nl = (uint64_t)(mn[0] | (mn[1] << 16)) | ((uint64_t)(mn[2] | (mn[3] << 16)) << 32);
        cm = (uint64_t)(mc[0] | (mc[1] << 16)) | ((uint64_t)(mc[2] | (mc[3] << 16)) << 32);
        a = win[64];
        b = win[65];
        if (any_cr & 0x80808080u) {   // carriage returns are rare (none in LF files): their mask is built on demand
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const uint4 x = *reinterpret_cast<const uint4*>(win + 16 * v);
                const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) cr |= (uint64_t)flags_nibble(eq_flags(w[k], 0x0d0d0d0du)) << (16 * v + 4 * k);
            }
        }
    } else {
        for (int i = 0; i < 64 && base + i < n; ++i) {
            const uint8_t c = win[i];
            q |= (uint64_t)(c == '"') << i;
            nl |= (uint64_t)(c == '\n') << i;
            cr |= (uint64_t)(c == '\r') << i;
            cm |= (uint64_t)(c == ',') << i;
        }
        a = base + 64 < n ? win[64] : 0;
        b = base + 65 < n ? win[65] : 0;
    }
    const uint64_t nl1 = (nl >> 1) | ((uint64_t)(a == '\n') << 63);
    const uint64_t cr1 = (cr >> 1) | ((uint64_t)(a == '\r') << 63);
    const uint64_t nl2 = (nl >> 2) | ((uint64_t)(a == '\n') << 62) | ((uint64_t)(b == '\n') << 63);
    const int64_t left = n - base;
    const uint64_t in1 = left >= 65 ? ~0ull : (left <= 1 ? 0ull : ((1ull << (left - 1)) - 1));
    const uint64_t in2 = left >= 66 ? ~0ull : (left <= 2 ? 0ull : ((1ull << (left - 2)) - 1));
    *opens = in1 & ~nl1 & ~(cr1 & (~in2 | nl2));
    *quotes = q;
    *newlines = nl;
    *commas = cm;
}

// bit i = parity of the quotes strictly before position i (window-relative)
__device__ __forceinline__ uint64_t inside_mask(uint64_t q) {
    uint64_t x = q;
    x ^= x << 1; x ^= x << 2; x ^= x << 4; x ^= x << 8; x ^= x << 16; x ^= x << 32;
    return x << 1;
}

// does byte 0 of the text open a non-blank line?
__device__ __forceinline__ bool csv_first_line_nonblank(const uint8_t* text, int64_t n) {
    if (n <= 0) return false;
    const uint8_t c = text[0];
    if (c == '\n') return false;
    if (c == '\r' && (n == 1 || text[1] == '\n')) return false;
    return true;
}

__global__ void __launch_bounds__(CSV_THREADS) csv_block_summary_kernel(const uint8_t* __restrict__ text, int64_t n,
                                                                        int64_t blocks, CsvBlockSummary* __restrict__ out,
                                                                        ulonglong2* __restrict__ masks,
                                                                        ulonglong2* __restrict__ field_masks) {
    extern __shared__ __align__(128) unsigned char csv_smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(csv_smem + 2 * CSV_STAGE_BYTES);
    __shared__ uint32_t s_par[CSV_THREADS / 32];
    __shared__ uint32_t s_out[CSV_THREADS / 32], s_in[CSV_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(bar + 1, 1);
        mbar_init_fence();
    }
    __syncthreads();
    int64_t blk = blockIdx.x;
    if (tid == 0 && blk < blocks) csv_stage_issue(text, n, blk, csv_smem, bar);
    for (int it = 0; blk < blocks; blk += gridDim.x, ++it) {
        const int buf = it & 1;
        __syncthreads();  // the other buffer and the s_* words of the previous block are free
        if (tid == 0 && blk + gridDim.x < blocks)
            csv_stage_issue(text, n, blk + gridDim.x, csv_smem + (buf ^ 1) * CSV_STAGE_BYTES, bar + (buf ^ 1));
        mbar_wait(bar + buf, (unsigned)(it >> 1) & 1u);
        const int64_t base = blk * CSV_BLOCK_BYTES + (int64_t)tid * CSV_BYTES_PER_THREAD;
        uint64_t q, nl, opens, cm;
        csv_masks64_staged(csv_smem + buf * CSV_STAGE_BYTES + tid * CSV_BYTES_PER_THREAD, base, n, &q, &nl, &opens, &cm);
        // pass 2 needs exactly these two words per window: kept (16 B per 64 B of text) so that it neither
        // re-reads the text nor rebuilds the masks; the field parse walks rows over them plus the comma mask
        masks[blk * CSV_THREADS + tid] = make_ulonglong2(q, nl & opens);
        field_masks[blk * CSV_THREADS + tid] = make_ulonglong2(cm, nl);   // {commas, newlines}
        const uint64_t ins = inside_mask(q);
        unsigned c_out = __popcll(nl & ~ins & opens), c_in = __popcll(nl & ins & opens);  // thread starts outside / inside
        const unsigned tp = __popcll(q) & 1;
        // thread's starting state relative to the block start
        const unsigned wb = __ballot_sync(0xffffffffu, tp);
        const unsigned lane_state = __popc(wb & lanemask_lt()) & 1;
        if (lane == 0) s_par[warp] = __popc(wb) & 1;
        __syncthreads();
        unsigned warp_state = 0, block_par = 0;
#pragma unroll
        for (int w = 0; w < CSV_THREADS / 32; ++w) {
            if (w < warp) warp_state ^= s_par[w];
            block_par ^= s_par[w];
        }
        const unsigned st = lane_state ^ warp_state;
        unsigned a = st ? c_in : c_out;   // starts if the block begins outside quotes
        unsigned b = st ? c_out : c_in;   // ... inside quotes
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        if (lane == 0) { s_out[warp] = a; s_in[warp] = b; }
        __syncthreads();
        if (tid == 0) {
            uint64_t ta = 0, tb = 0;
#pragma unroll
            for (int w = 0; w < CSV_THREADS / 32; ++w) { ta += s_out[w]; tb += s_in[w]; }
            if (blk == 0 && csv_first_line_nonblank(text, n)) { ta += 1; tb += 1; }  // row 0 starts at byte 0
            CsvBlockSummary r;
            r.parity = block_par; r.pad = 0; r.starts_if_outside = ta; r.starts_if_inside = tb; r.reserved = 0;
            out[blk] = r;
        }
    }
}

// Ordered composition of the block summaries.  A span of blocks acts on the quote state like
// (parity, starts-if-it-begins-outside, starts-if-it-begins-inside); spans compose associatively:
//   (p1,a1,b1) . (p2,a2,b2) = (p1^p2, a1 + (p1 ? b2 : a2), b1 + (p1 ? a2 : b2))
// so the scan is three small kernels: reduce 256 summaries per CTA, scan the per-CTA spans in
// one CTA, then re-scan every CTA's 256 summaries from its known starting state.
struct CsvSpan {
    uint32_t par;
    uint64_t a, b;
};
__device__ __forceinline__ CsvSpan span_compose(const CsvSpan& x, const CsvSpan& y) {
    CsvSpan r;
    r.par = x.par ^ y.par;
    r.a = x.a + (x.par ? y.b : y.a);
    r.b = x.b + (x.par ? y.a : y.b);
    return r;
}
__device__ __forceinline__ CsvSpan span_shfl_up(const CsvSpan& x, int o) {
    CsvSpan r;
    r.par = __shfl_up_sync(0xffffffffu, x.par, o);
    r.a = __shfl_up_sync(0xffffffffu, x.a, o);
    r.b = __shfl_up_sync(0xffffffffu, x.b, o);
    return r;
}
// inclusive scan over the 256 threads of a CTA (thread order = block order); returns the thread's
// inclusive span, *total = the CTA's span
__device__ __forceinline__ CsvSpan span_block_scan(CsvSpan v, CsvSpan* s_warp, CsvSpan* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const CsvSpan up = span_shfl_up(v, o);
        if (lane >= o) v = span_compose(up, v);
    }
    if (lane == 31) s_warp[warp] = v;
    __syncthreads();
    CsvSpan pre{0, 0, 0}, all{0, 0, 0};
    for (int w = 0; w < CSV_THREADS / 32; ++w) {
        if (w < warp) pre = span_compose(pre, s_warp[w]);
        all = span_compose(all, s_warp[w]);
    }
    *total = all;
    __syncthreads();
    return span_compose(pre, v);
}

__global__ void __launch_bounds__(CSV_THREADS) csv_span_reduce_kernel(const CsvBlockSummary* __restrict__ sum,
                                                                      int64_t blocks, CsvBlockSummary* __restrict__ super) {
    __shared__ CsvSpan s_warp[CSV_THREADS / 32];
    const int64_t i = (int64_t)blockIdx.x * CSV_THREADS + threadIdx.x;
    CsvSpan v{0, 0, 0};
    if (i < blocks) { v.par = sum[i].parity; v.a = sum[i].starts_if_outside; v.b = sum[i].starts_if_inside; }
    CsvSpan total;
    span_block_scan(v, s_warp, &total);
    if (threadIdx.x == 0) {
        CsvBlockSummary r;
        r.parity = total.par; r.pad = 0; r.starts_if_outside = total.a; r.starts_if_inside = total.b; r.reserved = 0;
        super[blockIdx.x] = r;
    }
}

// one CTA: exclusive scan of the (few hundred to few thousand) per-CTA spans -> each CTA's start
__global__ void __launch_bounds__(CSV_THREADS) csv_span_scan_kernel(const CsvBlockSummary* __restrict__ super, int64_t supers,
                                                                    CsvBlockStart* __restrict__ super_start,
                                                                    int64_t* total_rows) {
    __shared__ CsvSpan s_warp[CSV_THREADS / 32];
    uint32_t state = 0;   // quote state at the start of the current chunk
    uint64_t row = 0;     // records before the current chunk
    for (int64_t base = 0; base < supers; base += CSV_THREADS) {
        const int64_t i = base + threadIdx.x;
        CsvSpan v{0, 0, 0};
        if (i < supers) { v.par = super[i].parity; v.a = super[i].starts_if_outside; v.b = super[i].starts_if_inside; }
        CsvSpan total;
        const CsvSpan incl = span_block_scan(v, s_warp, &total);
        // exclusive span of thread t = inclusive span of thread t-1 (the operator has no inverse)
        __shared__ CsvSpan s_incl[CSV_THREADS];
        s_incl[threadIdx.x] = incl;
        __syncthreads();
        CsvSpan ex{0, 0, 0};
        if (threadIdx.x > 0) ex = s_incl[threadIdx.x - 1];
        if (i < supers) {
            CsvBlockStart o;
            o.inside = state ^ ex.par;
            o.first_row = row + (state ? ex.b : ex.a);
            o.pad = 0;
            super_start[i] = o;
        }
        row += state ? total.b : total.a;
        state ^= total.par;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_rows = (int64_t)row;
}

__global__ void __launch_bounds__(CSV_THREADS) csv_span_apply_kernel(const CsvBlockSummary* __restrict__ sum, int64_t blocks,
                                                                     const CsvBlockStart* __restrict__ super_start,
                                                                     CsvBlockStart* __restrict__ start) {
    __shared__ CsvSpan s_warp[CSV_THREADS / 32];
    __shared__ CsvSpan s_incl[CSV_THREADS];
    const int64_t i = (int64_t)blockIdx.x * CSV_THREADS + threadIdx.x;
    CsvSpan v{0, 0, 0};
    if (i < blocks) { v.par = sum[i].parity; v.a = sum[i].starts_if_outside; v.b = sum[i].starts_if_inside; }
    CsvSpan total;
    s_incl[threadIdx.x] = span_block_scan(v, s_warp, &total);
    __syncthreads();
    CsvSpan ex{0, 0, 0};
    if (threadIdx.x > 0) ex = s_incl[threadIdx.x - 1];
    const CsvBlockStart ss = super_start[blockIdx.x];
    if (i < blocks) {
        CsvBlockStart o;
        o.inside = ss.inside ^ ex.par;
        o.first_row = ss.first_row + (ss.inside ? ex.b : ex.a);
        o.pad = 0;
        start[i] = o;
    }
}

__global__ void __launch_bounds__(CSV_THREADS) csv_row_offsets_kernel(const uint8_t* __restrict__ text, int64_t n,
                                                                      int64_t blocks, const CsvBlockStart* __restrict__ start,
                                                                      const ulonglong2* __restrict__ masks,
                                                                      int64_t* __restrict__ row_start) {
    __shared__ uint32_t s_par[CSV_THREADS / 32];
    __shared__ uint32_t s_cnt[CSV_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int64_t blk = blockIdx.x; blk < blocks; blk += gridDim.x) {
        const CsvBlockStart bs = start[blk];
        const ulonglong2 m = masks[blk * CSV_THREADS + tid];   // {quotes, newlines that open a non-blank line} of the window
        const uint64_t q = m.x;
        const int64_t base = blk * CSV_BLOCK_BYTES + (int64_t)tid * CSV_BYTES_PER_THREAD;
        const unsigned tp = __popcll(q) & 1;
        const unsigned wb = __ballot_sync(0xffffffffu, tp);
        const unsigned lane_state = __popc(wb & lanemask_lt()) & 1;
        __syncthreads();   // the previous block's s_par / s_cnt are no longer read
        if (lane == 0) s_par[warp] = __popc(wb) & 1;
        __syncthreads();
        unsigned st = lane_state ^ bs.inside;
#pragma unroll
        for (int w = 0; w < CSV_THREADS / 32; ++w)
            if (w < warp) st ^= s_par[w];
        uint64_t ins = inside_mask(q);
        if (st) ins = ~ins;
        uint64_t ev = m.y & ~ins;  // newlines outside quotes whose next byte opens a non-blank line
        const bool virt = blk == 0 && tid == 0 && csv_first_line_nonblank(text, n);
        unsigned cnt = __popcll(ev) + (virt ? 1 : 0);
        unsigned incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        if (lane == 31) s_cnt[warp] = incl;
        __syncthreads();
        uint64_t row = bs.first_row + (incl - cnt);
#pragma unroll
        for (int w = 0; w < CSV_THREADS / 32; ++w)
            if (w < warp) row += s_cnt[w];
        if (virt) row_start[row++] = 0;
        while (ev) {
            const int i = __ffsll((long long)ev) - 1;
            ev &= ev - 1;
            row_start[row++] = base + i + 1;
        }
    }
}

// ------------------------------------------------------------------ field parse

struct CsvParseParams {
    const uint8_t* text;
    int64_t n;
    const int64_t* row_start;
    int64_t first_row;   // first row to decode (1 skips the header)
    int64_t rows;        // rows to decode
    int ncols;
    int32_t field[CSV_MAX_COLS];   // field position of column s in a row
    int32_t kind[CSV_MAX_COLS];    // 0 = numeric -> double, 1 = string -> dictionary slot (int32)
    void* out[CSV_MAX_COLS];
    uint32_t* valid[CSV_MAX_COLS];
    uint64_t* dict_hash[CSV_MAX_COLS];  // string columns: capacity words, 0 = empty
    uint64_t* dict_rep[CSV_MAX_COLS];   // representative field of the slot (packed offset / length / quoted)
    uint32_t dict_mask[CSV_MAX_COLS];   // capacity - 1 (capacity is a power of two)
    int32_t last_field;                 // largest requested field position: the row scan stops once it is closed
    int flags;                          // CSV_TRIM_STRINGS | CSV_NAN_IS_NULL
    int32_t* status;                    // [0] numerals outside the exact domain, [1] dictionary overflows,
                                        // [2] bit s set: column s saw a non-integer numeral, [3] unparsable numerals
    int64_t* first_bad_row;             // smallest row (file order) with a status[0] / status[1] event
    const ulonglong2* quote_masks;      // MASKS kernel: per 64-byte window {quotes, -} and {commas, newlines}, as
    const ulonglong2* field_masks;      // etl_csv_index_rows left them in the arena
};
constexpr int CSV_TRIM_STRINGS = 1;  // load_csv strips the label cell (train_tf_ps.py:120)
constexpr int CSV_NAN_IS_NULL = 2;   // load_csv drops rows whose numeric text is "nan" (train_tf_ps.py:129)

constexpr uint64_t REP_LEN_BITS = 23;
constexpr uint64_t REP_QUOTED = 1ull << REP_LEN_BITS;
__device__ __forceinline__ uint64_t rep_pack(int64_t off, int len, bool quoted) {
    return ((uint64_t)off << (REP_LEN_BITS + 1)) | (quoted ? REP_QUOTED : 0) | (uint64_t)len;
}

// canonical bytes of a field: a quoted field's content with "" read as one "
struct FieldCursor {
    const uint8_t* p;
    int len, i;
    bool quoted;
    __device__ __forceinline__ FieldCursor(const uint8_t* p_, int len_, bool q) : p(p_), len(len_), i(0), quoted(q) {}
    __device__ __forceinline__ int next() {  // -1 at the end
        if (i >= len) return -1;
        const uint8_t c = p[i++];
        if (quoted && c == '"' && i < len && p[i] == '"') ++i;
        return c;
    }
};

__device__ __forceinline__ uint64_t field_hash(const uint8_t* p, int len, bool quoted) {
    uint64_t h = 0xcbf29ce484222325ull;
    FieldCursor cur(p, len, quoted);
    for (int c = cur.next(); c >= 0; c = cur.next()) h = (h ^ (uint64_t)c) * 0x100000001b3ull;
    h = mix64(h);
    return h ? h : 1;
}

__device__ __forceinline__ bool field_equal(const uint8_t* text, uint64_t ra, uint64_t rb) {
    FieldCursor a(text + (ra >> (REP_LEN_BITS + 1)), (int)(ra & (REP_QUOTED - 1)), (ra & REP_QUOTED) != 0);
    FieldCursor b(text + (rb >> (REP_LEN_BITS + 1)), (int)(rb & (REP_QUOTED - 1)), (rb & REP_QUOTED) != 0);
    while (true) {
        const int x = a.next(), y = b.next();
        if (x != y) return false;
        if (x < 0) return true;
    }
}

// returns the slot of the field's string, inserting it if new; -1 if the table is full
__device__ __forceinline__ int dict_find_or_insert(uint64_t* hash_tab, uint64_t* rep_tab, uint32_t mask,
                                                   const uint8_t* text, uint64_t h, uint64_t rep) {
    uint32_t slot = (uint32_t)(h >> 24) & mask;
    for (uint32_t probe = 0; probe <= mask; ++probe) {
        uint64_t cur = ld_relaxed_u64(hash_tab + slot);
        if (cur == 0) {
            cur = atomicCAS(reinterpret_cast<unsigned long long*>(hash_tab + slot), 0ull, (unsigned long long)h);
            if (cur == 0) {  // we own the slot: publish its representative
                st_relaxed_u64(rep_tab + slot, rep);
                return (int)slot;
            }
        }
        if (cur == h) {
            uint64_t r;
            do { r = ld_relaxed_u64(rep_tab + slot); } while (r == 0);
            if (field_equal(text, r, rep)) return (int)slot;
        }
        slot = (slot + 1) & mask;
    }
    return -1;
}

// 8-bit mask of the bytes of w equal to c
__device__ __forceinline__ unsigned eq8(uint64_t w, uint32_t pat) {
    return eq4((uint32_t)w, pat) | (eq4((uint32_t)(w >> 32), pat) << 4);
}

constexpr int PARSE_WARPS = 4;
constexpr uint32_t F_NONE = 0xffffffffu;

// 8 text bytes at an arbitrary address as a little-endian word (two aligned loads + funnel shift;
// the text buffer is readable up to the next multiple of 8 past its end)
__device__ __forceinline__ uint64_t load8_unaligned(const uint8_t* p) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint64_t* q = reinterpret_cast<const uint64_t*>(a & ~(uintptr_t)7);
    const unsigned sh = (unsigned)(a & 7) * 8;
    const uint64_t lo = q[0];
    if (sh == 0) return lo;
    return (lo >> sh) | (q[1] << (64 - sh));
}

// hash of a field's bytes, 8 at a time; `plain` = no doubled quotes inside (the common case)
__device__ __forceinline__ uint64_t field_hash_words(const uint8_t* p, int len) {
    uint64_t h = 0x9E3779B97F4A7C15ull ^ (uint64_t)len;
    int i = 0;
    for (; i + 8 <= len; i += 8) h = (h ^ load8_unaligned(p + i)) * 0xff51afd7ed558ccdull, h ^= h >> 32;
    if (i < len) {
        const uint64_t w = load8_unaligned(p + i) & (~0ull >> (8 * (8 - (len - i))));
        h = (h ^ w) * 0xff51afd7ed558ccdull, h ^= h >> 32;
    }
    h = mix64(h);
    return h ? h : 1;
}
__device__ __forceinline__ bool field_equal_words(const uint8_t* a, const uint8_t* b, int len) {
    int i = 0;
    for (; i + 8 <= len; i += 8)
        if (load8_unaligned(a + i) != load8_unaligned(b + i)) return false;
    if (i < len) {
        const uint64_t m = ~0ull >> (8 * (8 - (len - i)));
        if ((load8_unaligned(a + i) ^ load8_unaligned(b + i)) & m) return false;
    }
    return true;
}
// does the field hold a quote character (a doubled one inside a quoted field, or a bare one)?
__device__ __forceinline__ bool field_has_quote(const uint8_t* p, int len) {
    for (int i = 0; i < len; i += 8) {
        const int take = len - i < 8 ? len - i : 8;
        const unsigned m = eq8(load8_unaligned(p + i), 0x22222222u) & (0xffu >> (8 - take));
        if (m) return true;
    }
    return false;
}

// dictionary probe with the word-wise compare for plain fields (canonical cursor compare otherwise)
__device__ __forceinline__ int dict_find_or_insert_fast(uint64_t* hash_tab, uint64_t* rep_tab, uint32_t mask,
                                                        const uint8_t* text, uint64_t h, const uint8_t* f, int len,
                                                        bool quoted, bool plain) {
    const uint64_t rep = rep_pack(f - text, len, quoted);
    uint32_t slot = (uint32_t)(h >> 24) & mask;
    for (uint32_t probe = 0; probe <= mask; ++probe) {
        uint64_t cur = ld_relaxed_u64(hash_tab + slot);
        if (cur == 0) {
            cur = atomicCAS(reinterpret_cast<unsigned long long*>(hash_tab + slot), 0ull, (unsigned long long)h);
            if (cur == 0) {
                st_relaxed_u64(rep_tab + slot, rep);
                return (int)slot;
            }
        }
        if (cur == h) {
            uint64_t r;
            do { r = ld_relaxed_u64(rep_tab + slot); } while (r == 0);
            const bool r_quoted = (r & REP_QUOTED) != 0;
            const int r_len = (int)(r & (REP_QUOTED - 1));
            const uint8_t* rp = text + (r >> (REP_LEN_BITS + 1));
            // two plain fields are equal iff their bytes are; anything involving doubled quotes goes the slow way
            const bool r_plain = !field_has_quote(rp, r_len);
            (void)r_quoted;
            if (plain && r_plain ? (r_len == len && field_equal_words(rp, f, len)) : field_equal(text, r, rep))
                return (int)slot;
        }
        slot = (slot + 1) & mask;
    }
    return -1;
}

// [-]digits[.digits] with at most 15 significant digits: exact through one divide (Clinger);
// returns false when the cell needs the general parser
__device__ __forceinline__ bool dec_parse_simple(const uint8_t* p, int len, double* out, int* flags) {
    if (len <= 0 || len > 17) return false;
    int i = 0;
    const bool neg = p[0] == '-';
    if (neg) i = 1;
    uint64_t m = 0;
    int digits = 0, frac = 0;
    bool dot = false;
    for (; i < len; ++i) {
        const unsigned c = p[i];
        if (c - '0' <= 9u) {
            m = m * 10 + (c - '0');
            ++digits;
            frac += dot ? 1 : 0;
        } else if (c == '.' && !dot) {
            dot = true;
        } else {
            return false;
        }
    }
    if (digits == 0 || digits > 15) return false;
    double r = (double)m;
    if (frac) r /= dec_pow10(frac);
    *out = neg ? -r : r;
    *flags = dot ? DEC_FLAG_NONINT : 0;
    return true;
}

// One warp per group of 32 rows, lane r owns row r in both phases:
//   A  tokenise: walk the row 8 bytes at a time and leave each requested field's [start, end) in
//      shared memory;
//   B  convert the requested fields; the group's values leave as one coalesced store per column
//      plus one ballot-built validity word.
template <bool MASKS>
__global__ void __launch_bounds__(PARSE_WARPS * 32) csv_parse_rows_kernel(const CsvParseParams p) {
    __shared__ uint32_t s_fs[PARSE_WARPS][CSV_MAX_COLS][32], s_fe[PARSE_WARPS][CSV_MAX_COLS][32];  // relative to the group
    __shared__ uint32_t s_rend[PARSE_WARPS][32];
    // the requested columns in ascending field order: the row scan keeps two cursors into this list (the
    // next field to close, the next field to open), so a separator costs two register compares
    __shared__ int s_ord_field[CSV_MAX_COLS + 1], s_ord_slot[CSV_MAX_COLS + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        int nf = 0;
        for (int c = 0; c < p.ncols; ++c) {   // insertion sort of <= 8 entries
            int j = nf++;
            while (j > 0 && s_ord_field[j - 1] > p.field[c]) {
                s_ord_field[j] = s_ord_field[j - 1];
                s_ord_slot[j] = s_ord_slot[j - 1];
                --j;
            }
            s_ord_field[j] = p.field[c];
            s_ord_slot[j] = c;
        }
        s_ord_field[nf] = 0x7fffffff;   // sentinel: never reached
        s_ord_slot[nf] = 0;
    }
    __syncthreads();
    const unsigned lt = lanemask_lt();
    const int64_t groups = (p.rows + 31) / 32;
    for (int64_t g = (int64_t)blockIdx.x * PARSE_WARPS + warp; g < groups; g += (int64_t)gridDim.x * PARSE_WARPS) {
        const int64_t r0 = g * 32;
        const int in_group = p.rows - r0 < 32 ? (int)(p.rows - r0) : 32;
        const int64_t my_start = lane < in_group ? p.row_start[p.first_row + r0 + lane] : 0;
        const int64_t gbase = __shfl_sync(0xffffffffu, my_start, 0);  // offsets below are relative to this
        for (int s = 0; s < p.ncols; ++s) { s_fs[warp][s][lane] = F_NONE; s_fe[warp][s][lane] = F_NONE; }
        __syncwarp();
        // ---------------------------------------------------------------- phase A: lane = row
        // every lane walks its own row 8 bytes at a time: quote state by an in-word prefix xor plus a
        // carried parity bit, separators by ffs over the unquoted-comma mask
        if (MASKS && lane < in_group) {
            // the index pass kept, for every 64-byte window of the text, the positions of quotes, commas
            // and newlines: the row is walked a window at a time with no byte compares at all
            const int64_t start = my_start;
            int ci = 0, oi = 0;                       // cursors: next requested field to close / to open
            int cf = s_ord_field[0], cslot = s_ord_slot[0], of = cf, oslot = cslot;
            if (of == 0) {                            // field 0 opens at the row start
                s_fs[warp][oslot][lane] = (uint32_t)(start - gbase);
                ++oi; of = s_ord_field[oi]; oslot = s_ord_slot[oi];
            }
            int64_t win = start >> 6;
            const int64_t nwin = (p.n + 63) >> 6;
            uint64_t m = ~0ull << (start & 63);   // bits of the window that belong to the row
            unsigned inside = 0, fi = 0;
            int64_t row_end = p.n;
            while (win < nwin) {
                const uint64_t q = p.quote_masks[win].x & m;
                const ulonglong2 fm = p.field_masks[win];
                uint64_t ins = inside_mask(q);
                if (inside) ins = ~ins;
                const uint64_t e = fm.y & ~ins & m;                              // newlines outside quotes
                const uint64_t keep = e ? ((e & (0ull - e)) - 1) : ~0ull;        // bytes before the record end
                uint64_t oc = fm.x & ~ins & m & keep;
                const int64_t wbase = win << 6;
                while (oc) {
                    const int b = __ffsll((long long)oc) - 1;
                    oc &= oc - 1;
                    const uint32_t at = (uint32_t)(wbase + b - gbase);
                    if ((int)fi == cf) {              // separator fi closes field fi ...
                        s_fe[warp][cslot][lane] = at;
                        ++ci; cf = s_ord_field[ci]; cslot = s_ord_slot[ci];
                    }
                    ++fi;
                    if ((int)fi == of) {              // ... and opens field fi + 1
                        s_fs[warp][oslot][lane] = at + 1;
                        ++oi; of = s_ord_field[oi]; oslot = s_ord_slot[oi];
                    }
                }
                if (e) {
                    row_end = wbase + __ffsll((long long)e) - 1;
                    break;
                }
                if ((int)fi > p.last_field) {  // every requested field is closed: the rest of the row is not needed
                    row_end = wbase > start ? wbase : start;
                    break;
                }
                inside ^= __popcll(q) & 1;
                ++win;
                m = ~0ull;
            }
            s_rend[warp][lane] = (uint32_t)(row_end - gbase);
        }
        if (!MASKS && lane < in_group) {
            const int64_t start = my_start;
            int ci = 0, oi = 0;                       // cursors: next requested field to close / to open
            int cf = s_ord_field[0], cslot = s_ord_slot[0], of = cf, oslot = cslot;
            if (of == 0) {                            // field 0 opens at the row start
                s_fs[warp][oslot][lane] = (uint32_t)(start - gbase);
                ++oi; of = s_ord_field[oi]; oslot = s_ord_slot[oi];
            }
            int64_t pos = start & ~(int64_t)7;
            unsigned inside = 0, fi = 0;
            int64_t row_end = p.n;
            while (pos < p.n) {
                const uint64_t w = *reinterpret_cast<const uint64_t*>(p.text + pos);
                unsigned vm = 0xff;
                if (pos < start) vm = 0xffu << (start - pos);
                if (pos + 8 > p.n) vm &= 0xffu >> (8 - (p.n - pos));
                const unsigned c8 = eq8(w, 0x2c2c2c2cu) & vm;
                // a word with neither a quote nor a newline (and outside quotes) needs only its commas - the
                // usual case when the requested fields end before the row does (a false positive from the
                // bytes outside vm merely takes the exact path)
                const uint32_t lo = (uint32_t)w, hi = (uint32_t)(w >> 32);
                const uint32_t special = (has_byte(lo, 0x22222222u) | has_byte(hi, 0x22222222u) | has_byte(lo, 0x0a0a0a0au) |
                                          has_byte(hi, 0x0a0a0a0au)) & 0x80808080u;
                unsigned q8 = 0, end8 = 0, oc = c8;
                if (special | inside) {
                    q8 = eq8(w, 0x22222222u) & vm;
                    const unsigned n8 = eq8(w, 0x0a0a0a0au) & vm;
                    unsigned ins = q8;
                    ins ^= ins << 1; ins ^= ins << 2; ins ^= ins << 4;
                    ins = (ins << 1) & 0xff;
                    if (inside) ins ^= 0xff;
                    end8 = n8 & ~ins & 0xff;
                    const unsigned keep = end8 ? ((end8 & (0u - end8)) - 1) : 0xffu;  // bytes before the record end
                    oc = c8 & ~ins & keep;
                }
                while (oc) {
                    const int b = __ffs(oc) - 1;
                    oc &= oc - 1;
                    const uint32_t at = (uint32_t)(pos + b - gbase);
                    if ((int)fi == cf) {              // separator fi closes field fi ...
                        s_fe[warp][cslot][lane] = at;
                        ++ci; cf = s_ord_field[ci]; cslot = s_ord_slot[ci];
                    }
                    ++fi;
                    if ((int)fi == of) {              // ... and opens field fi + 1
                        s_fs[warp][oslot][lane] = at + 1;
                        ++oi; of = s_ord_field[oi]; oslot = s_ord_slot[oi];
                    }
                }
                if (end8) {
                    row_end = pos + __ffs(end8) - 1;
                    break;
                }
                if ((int)fi > p.last_field) {  // every requested field is closed: the rest of the row is not needed
                    row_end = pos;             // (unused then; kept inside the group's 32-bit offset range)
                    break;
                }
                inside ^= __popc(q8) & 1;
                pos += 8;
            }
            s_rend[warp][lane] = (uint32_t)(row_end - gbase);
        }
        __syncwarp();
        // ---------------------------------------------------------------- phase B: lane = row
        const int64_t row = p.first_row + r0 + lane;
        const uint32_t rend = lane < in_group ? s_rend[warp][lane] : 0;
        (void)lt;
        for (int s = 0; s < p.ncols; ++s) {
            bool ok = false;
            uint64_t bits = 0;
            const uint32_t fs = s_fs[warp][s][lane];
            if (lane < in_group && fs != F_NONE) {
                uint32_t fe = s_fe[warp][s][lane];
                if (fe == F_NONE) fe = rend;  // the row's last field
                const uint8_t* f = p.text + gbase + fs;
                int len = (int)(fe - fs);
                if (len > 0 && fe == rend && f[len - 1] == '\r') --len;  // CRLF line end
                const bool numeric = p.kind[s] == 0;
                const bool trim = numeric || (p.flags & CSV_TRIM_STRINGS);
                if (trim) {
                    while (len > 0 && dec_is_space(f[0])) { ++f; --len; }
                    while (len > 0 && dec_is_space(f[len - 1])) --len;
                }
                bool quoted = false;
                if (len >= 2 && f[0] == '"' && f[len - 1] == '"') {
                    quoted = true; ++f; len -= 2;
                    if (trim) {  // csv yields the inner text, then .strip()
                        while (len > 0 && dec_is_space(f[0])) { ++f; --len; }
                        while (len > 0 && dec_is_space(f[len - 1])) --len;
                    }
                }
                if (numeric) {
                    double v = 0.0;
                    int fl = 0;
                    int st = DEC_OK;
                    if (!dec_parse_simple(f, len, &v, &fl)) st = dec_parse(f, len, &v, &fl);
                    if (st == DEC_OK) {
                        ok = !((fl & DEC_FLAG_NAN) && (p.flags & CSV_NAN_IS_NULL));
                        bits = (uint64_t)__double_as_longlong(v);
                        if ((fl & DEC_FLAG_NONINT) && !(p.status[2] & (1 << s))) atomicOr(p.status + 2, 1 << s);
                    } else if (st == DEC_INEXACT) {
                        atomicAdd(p.status + 0, 1);
                        atomicMin(reinterpret_cast<long long*>(p.first_bad_row), (long long)row);
                    } else if (st == DEC_INVALID) {
                        atomicAdd(p.status + 3, 1);
                    }
                } else if (len > 0) {
                    // "plain" = no quote character at all: such strings always take the word-wise hash, strings
                    // holding a quote (doubled inside a quoted field, or bare) always the canonical byte-wise one
                    const bool plain = !field_has_quote(f, len);
                    const uint64_t h = plain ? field_hash_words(f, len) : field_hash(f, len, quoted);
                    const int slot = dict_find_or_insert_fast(p.dict_hash[s], p.dict_rep[s], p.dict_mask[s], p.text, h, f,
                                                              len, quoted, plain);
                    if (slot >= 0) {
                        ok = true;
                        bits = (uint64_t)(uint32_t)slot;
                    } else {
                        atomicAdd(p.status + 1, 1);
                        atomicMin(reinterpret_cast<long long*>(p.first_bad_row), (long long)row);
                    }
                }
            }
            const uint32_t vw = __ballot_sync(0xffffffffu, ok);
            if (lane < in_group) {
                if (p.kind[s] == 0) static_cast<double*>(p.out[s])[r0 + lane] = __longlong_as_double((long long)bits);
                else static_cast<int32_t*>(p.out[s])[r0 + lane] = ok ? (int32_t)bits : -1;
            }
            if (lane == 0) p.valid[s][g] = vw;
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256) remap_codes_kernel(int32_t* __restrict__ codes, int64_t n,
                                                          const int32_t* __restrict__ table, int32_t table_size,
                                                          int32_t missing) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int32_t c = codes[i];
        codes[i] = (c >= 0 && c < table_size) ? __ldg(table + c) : missing;
    }
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_csv_index_rows(etl_handle* hh, const uint8_t* text, int64_t nbytes, int64_t* d_rows,
                                  void* stream_) {
    ETL_REQUIRE(hh && d_rows, "csv_index_rows: NULL handle or output");
    ETL_REQUIRE(nbytes >= 0, "csv_index_rows: nbytes < 0");
    ETL_REQUIRE(nbytes == 0 || (text && etl::aligned16(text)), "csv_index_rows: text must be 16-byte aligned");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    if (nbytes == 0) {
        ETL_CUDA(cudaMemsetAsync(d_rows, 0, sizeof(int64_t), stream));
        return ETL_OK;
    }
    const int64_t blocks = (nbytes + etl::CSV_BLOCK_BYTES - 1) / etl::CSV_BLOCK_BYTES;
    const int64_t supers = (blocks + etl::CSV_THREADS - 1) / etl::CSV_THREADS;
    // scratch: block summaries | block starts | per-window masks (16 B per 64 B of text) | span arrays
    const size_t need = (size_t)blocks * (sizeof(etl::CsvBlockSummary) + sizeof(etl::CsvBlockStart) +
                                          etl::CSV_THREADS * 2 * sizeof(ulonglong2)) +
                        (size_t)supers * (sizeof(etl::CsvBlockSummary) + sizeof(etl::CsvBlockStart));
    int rc = etl::arena_reserve(h, need);
    if (rc != ETL_OK) return rc;
    auto* sum = static_cast<etl::CsvBlockSummary*>(h->scratch);
    auto* start = reinterpret_cast<etl::CsvBlockStart*>(sum + blocks);
    auto* masks = reinterpret_cast<ulonglong2*>(start + blocks);
    auto* field_masks = masks + blocks * etl::CSV_THREADS;
    auto* super = reinterpret_cast<etl::CsvBlockSummary*>(field_masks + blocks * etl::CSV_THREADS);
    h->csv_text = text;      // the masks in the arena describe this text (until the arena is reallocated)
    h->csv_nbytes = nbytes;
    auto* super_start = reinterpret_cast<etl::CsvBlockStart*>(super + supers);
    ETL_CUDA(cudaFuncSetAttribute(etl::csv_block_summary_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)etl::CSV_STAGE_SMEM));
    const int64_t grid = blocks < (int64_t)h->sm_count * 6 ? blocks : (int64_t)h->sm_count * 6;   // 33 KB each: 6 CTAs per SM
    etl::csv_block_summary_kernel<<<(unsigned)grid, etl::CSV_THREADS, etl::CSV_STAGE_SMEM, stream>>>(text, nbytes, blocks, sum, masks,
                                                                                                    field_masks);
    etl::csv_span_reduce_kernel<<<(unsigned)supers, etl::CSV_THREADS, 0, stream>>>(sum, blocks, super);
    etl::csv_span_scan_kernel<<<1, etl::CSV_THREADS, 0, stream>>>(super, supers, super_start, d_rows);
    etl::csv_span_apply_kernel<<<(unsigned)supers, etl::CSV_THREADS, 0, stream>>>(sum, blocks, super_start, start);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_csv_row_offsets(etl_handle* hh, const uint8_t* text, int64_t nbytes, int64_t* row_start,
                                   void* stream_) {
    ETL_REQUIRE(hh, "csv_row_offsets: NULL handle");
    ETL_REQUIRE(nbytes >= 0, "csv_row_offsets: nbytes < 0");
    if (nbytes == 0) return ETL_OK;
    ETL_REQUIRE(text && etl::aligned16(text) && row_start, "csv_row_offsets: NULL or misaligned argument");
    Handle* h = reinterpret_cast<Handle*>(hh);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ETL_CUDA(cudaSetDevice(h->device));
    const int64_t blocks = (nbytes + etl::CSV_BLOCK_BYTES - 1) / etl::CSV_BLOCK_BYTES;
    const size_t need = (size_t)blocks * (sizeof(etl::CsvBlockSummary) + sizeof(etl::CsvBlockStart) +
                                          etl::CSV_THREADS * sizeof(ulonglong2));
    // the masks pass 1 left in the arena must still be this text's: any call that reserves the arena in between
    // (sums, partitions, merges ...) clears csv_text
    ETL_REQUIRE(h->scratch_bytes >= need && h->csv_text == text && h->csv_nbytes == nbytes,
                "csv_row_offsets: call etl_csv_index_rows on the same text first (nothing that uses the handle's arena in between)");
    auto* sum = static_cast<etl::CsvBlockSummary*>(h->scratch);
    auto* start = reinterpret_cast<etl::CsvBlockStart*>(sum + blocks);
    auto* masks = reinterpret_cast<const ulonglong2*>(start + blocks);   // left there by etl_csv_index_rows
    const int64_t grid = blocks < (int64_t)h->sm_count * 8 ? blocks : (int64_t)h->sm_count * 8;
    etl::csv_row_offsets_kernel<<<(unsigned)grid, etl::CSV_THREADS, 0, stream>>>(text, nbytes, blocks, start, masks, row_start);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_csv_parse_columns(etl_handle* hh, const uint8_t* text, int64_t nbytes, const int64_t* row_start,
                                     int64_t first_row, int64_t rows, int32_t ncols, const int32_t* field,
                                     const int32_t* kind, void* const* out, uint32_t* const* valid,
                                     uint64_t* const* dict_hash, uint64_t* const* dict_rep,
                                     const int32_t* dict_capacity, int32_t flags, int32_t* d_status,
                                     int64_t* d_first_bad_row, void* stream_) {
    ETL_REQUIRE(hh && field && kind && out && valid && d_status && d_first_bad_row, "csv_parse_columns: NULL argument");
    ETL_REQUIRE(ncols >= 1 && ncols <= etl::CSV_MAX_COLS, "csv_parse_columns: ncols=%d not in 1..%d", ncols,
                etl::CSV_MAX_COLS);
    ETL_REQUIRE(rows >= 0 && first_row >= 0, "csv_parse_columns: negative row range");
    if (rows == 0) return ETL_OK;
    ETL_REQUIRE(text && row_start && nbytes > 0, "csv_parse_columns: NULL text / row_start");
    ETL_REQUIRE((reinterpret_cast<uintptr_t>(text) & 7) == 0, "csv_parse_columns: text must be 8-byte aligned");
    Handle* h = reinterpret_cast<Handle*>(hh);
    etl::CsvParseParams p{};
    p.text = text; p.n = nbytes; p.row_start = row_start; p.first_row = first_row; p.rows = rows; p.ncols = ncols;
    p.flags = flags; p.status = d_status; p.first_bad_row = d_first_bad_row;
    p.last_field = 0;
    for (int s = 0; s < ncols; ++s) p.last_field = field[s] > p.last_field ? field[s] : p.last_field;
    for (int s = 0; s < ncols; ++s) {
        ETL_REQUIRE(field[s] >= 0 && field[s] < etl::CSV_MAX_FIELDS, "csv_parse_columns: field[%d]=%d out of range", s,
                    field[s]);
        for (int t = 0; t < s; ++t) ETL_REQUIRE(field[t] != field[s], "csv_parse_columns: field %d requested twice", field[s]);
        ETL_REQUIRE(kind[s] == 0 || kind[s] == 1, "csv_parse_columns: kind[%d]=%d", s, kind[s]);
        ETL_REQUIRE(out[s] && valid[s], "csv_parse_columns: out[%d]/valid[%d] is NULL", s, s);
        p.field[s] = field[s]; p.kind[s] = kind[s]; p.out[s] = out[s]; p.valid[s] = valid[s];
        if (kind[s] == 1) {
            ETL_REQUIRE(dict_hash && dict_rep && dict_capacity && dict_hash[s] && dict_rep[s],
                        "csv_parse_columns: string column %d needs a dictionary", s);
            const int32_t cap = dict_capacity[s];
            ETL_REQUIRE(cap >= 2 && (cap & (cap - 1)) == 0, "csv_parse_columns: dict_capacity[%d]=%d is not a power of two",
                        s, cap);
            p.dict_hash[s] = dict_hash[s]; p.dict_rep[s] = dict_rep[s]; p.dict_mask[s] = (uint32_t)cap - 1;
        }
    }
    ETL_CUDA(cudaSetDevice(h->device));
    const int64_t groups = (rows + 31) / 32;
    int64_t blocks = (groups + etl::PARSE_WARPS - 1) / etl::PARSE_WARPS;
    const int64_t cap = (int64_t)h->sm_count * 16;
    if (blocks > cap) blocks = cap;
    // the masks etl_csv_index_rows left in the arena, if the caller vouches that this is still the text it indexed
    const int64_t csv_blocks = (nbytes + etl::CSV_BLOCK_BYTES - 1) / etl::CSV_BLOCK_BYTES;
    const size_t mask_end = (size_t)csv_blocks * (sizeof(etl::CsvBlockSummary) + sizeof(etl::CsvBlockStart) +
                                                  etl::CSV_THREADS * 2 * sizeof(ulonglong2));
    if ((flags & ETL_CSV_USE_INDEX_MASKS) && h->csv_text == text && h->csv_nbytes == nbytes && h->scratch_bytes >= mask_end) {
        auto* sum = static_cast<etl::CsvBlockSummary*>(h->scratch);
        auto* start = reinterpret_cast<etl::CsvBlockStart*>(sum + csv_blocks);
        p.quote_masks = reinterpret_cast<const ulonglong2*>(start + csv_blocks);
        p.field_masks = p.quote_masks + csv_blocks * etl::CSV_THREADS;
        etl::csv_parse_rows_kernel<true><<<(unsigned)blocks, etl::PARSE_WARPS * 32, 0, static_cast<cudaStream_t>(stream_)>>>(p);
    } else {
        etl::csv_parse_rows_kernel<false><<<(unsigned)blocks, etl::PARSE_WARPS * 32, 0, static_cast<cudaStream_t>(stream_)>>>(p);
    }
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_remap_codes(etl_handle* hh, int32_t* codes, int64_t n, const int32_t* table, int32_t table_size,
                               int32_t missing, void* stream_) {
    ETL_REQUIRE(hh && (n == 0 || (codes && table)), "remap_codes: NULL argument");
    ETL_REQUIRE(n >= 0 && table_size >= 0, "remap_codes: negative size");
    if (n == 0) return ETL_OK;
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    int64_t blocks = (n + 255) / 256;
    if (blocks > (int64_t)h->sm_count * 16) blocks = (int64_t)h->sm_count * 16;
    etl::remap_codes_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(codes, n, table, table_size,
                                                                                             missing);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
