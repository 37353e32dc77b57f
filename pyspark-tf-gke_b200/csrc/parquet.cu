// Parquet page decode (north-star item 1 "CSV/Parquet decode"; no reference site -
// the reference reads CSV / JDBC).  The host (parquet.py) parses footer and page
// headers and hands over a table of page descriptors; everything that scales
// with the number of rows happens here:
//   pq_hybrid_decode_kernel  one WARP per RLE / bit-packed hybrid stream
//        (definition levels, dictionary indices) of ANY column: the lanes decode a
//        run header speculatively at 32 consecutive byte offsets, a shuffle hop per
//        run settles which guesses lie on the real chain, and the runs are expanded
//        at once (no run table);
//   pq_snappy_kernel         one WARP per compressed page: the Snappy block format (what Spark and Arrow write
//        by default) is a chain of literal / copy elements that can only be found one after another; the
//        warp walks the chain in lockstep and all 32 lanes move each element's bytes (the most recent output
//        mirrored in shared memory, so a short-range copy costs no trip to L2), into a region behind the
//        file bytes, so every later kernel addresses inflated pages like any other part of the buffer;
//   pq_plain_strings_kernel  one WARP per PLAIN BYTE_ARRAY page (a writer's fallback when its dictionary grew
//        too large): lane 0 walks the [length][bytes] chain 32 values at a time, every lane hashes its value
//        into a device dictionary (as the CSV decoder does) and the slot numbers take the place of
//        dictionary indices;
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

// bw bits (<= 32) starting `bit` bits after p: two aligned 32-bit loads and a funnel shift.  Reads at
// most 7 bytes past the value - inside the file buffer, a stream is never the file's last bytes (the
// footer and the closing magic follow).
__device__ __forceinline__ uint32_t pq_extract(const uint8_t* p, int64_t bit, int bw) {
    const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
    bit += 8 * (int64_t)(addr & 3);
    const uint32_t lo = __ldg(w + (bit >> 5)), hi = __ldg(w + (bit >> 5) + 1);
    const uint32_t v = __funnelshift_r(lo, hi, (unsigned)(bit & 31));
    return bw >= 32 ? v : (v & ((1u << bw) - 1));
}
// 8 / 4 bytes at any alignment (PLAIN value sections start wherever the page header ended)
__device__ __forceinline__ uint64_t pq_load_u64(const uint8_t* p) {
    const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
    const uint64_t* w = reinterpret_cast<const uint64_t*>(addr & ~(uintptr_t)7);
    const unsigned sh = 8 * (unsigned)(addr & 7);
    const uint64_t lo = __ldg(w);
    if (sh == 0) return lo;
    return (lo >> sh) | (__ldg(w + 1) << (64 - sh));
}
__device__ __forceinline__ uint32_t pq_load_u32(const uint8_t* p) {
    const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
    const unsigned sh = 8 * (unsigned)(addr & 3);
    const uint32_t lo = __ldg(w);
    if (sh == 0) return lo;
    return __funnelshift_r(lo, __ldg(w + 1), sh);
}

constexpr int PQ_WINDOW = 2048;   // bytes of its stream a warp keeps in shared memory while it walks the run headers
constexpr int PQ_WARPS = PQ_THREADS / 32;

// One WARP per hybrid stream.  Run headers can only be found one after another, and writers emit
// short runs (Arrow switches between RLE and bit-packed runs every few values: ~1250 runs per
// 20 000-value page of definition levels), so the walk is a latency chain: the warp walks it in
// lockstep (every lane runs the same scalar code) over a window of the stream held in shared
// memory - a header byte costs a shared-memory access, not a dependent trip to L2 - and expands
// each run as soon as its header is read, all lanes storing.  One warp per stream (not one CTA)
// puts 64 streams in flight per SM; there is no run table.
__global__ void __launch_bounds__(PQ_THREADS) pq_hybrid_decode_kernel(const uint8_t* __restrict__ file,
                                                                      const PqStream* __restrict__ streams, int nstreams,
                                                                      int32_t* __restrict__ out, int32_t* status) {
    __shared__ __align__(16) uint8_t s_win_all[PQ_WARPS][PQ_WINDOW + 16];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sidx = blockIdx.x * PQ_WARPS + warp;
    if (sidx >= nstreams) return;
    uint8_t* s_win = s_win_all[warp];
    const PqStream st = streams[sidx];
    const uint8_t* src = file + st.src_offset;
    int32_t* dst = out + st.out_offset;
    const uintptr_t src_addr = reinterpret_cast<uintptr_t>(src);
    int64_t win_lo = 0, win_hi = 0;   // stream positions [win_lo, win_hi) are in s_win
    // make [pos, pos + need) resident (need <= PQ_WINDOW - 16)
    auto ensure = [&](int pos, int need) {
        if (pos >= win_lo && pos + need <= win_hi) return;
        __syncwarp();
        const uintptr_t g0 = (src_addr + (uintptr_t)pos) & ~(uintptr_t)15;   // 16-byte aligned global address
        const uint4* g = reinterpret_cast<const uint4*>(g0);
        // no further than the 16-byte chunk after the one holding the stream's last byte (what follows a
        // stream in the file - at the very least the footer - keeps those chunks inside the buffer)
        const int64_t chunks = (int64_t)((src_addr + (uintptr_t)st.src_bytes + 31 - g0) >> 4);
        for (int i = lane; i < PQ_WINDOW / 16 && i < chunks; i += 32) reinterpret_cast<uint4*>(s_win)[i] = __ldg(g + i);
        win_lo = (int64_t)(g0 - src_addr);
        win_hi = win_lo + PQ_WINDOW;
        __syncwarp();
    };
    // 8 stream bytes from position pos (resident), little endian
    auto peek64 = [&](int pos) -> uint64_t {
        const unsigned idx = (unsigned)(pos - win_lo);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(s_win) + (idx >> 2);
        const unsigned sh = 8 * (idx & 3);
        const uint32_t a = w[0], b = w[1], c = w[2];
        return (uint64_t)__funnelshift_r(a, b, sh) | ((uint64_t)__funnelshift_r(b, c, sh) << 32);
    };
    int64_t pos = 0;
    int produced = 0;
    // bit_width < 0: the page was compressed, so the host could not read the width byte in front of the indices
    const int bw = st.bit_width >= 0 ? st.bit_width : (int)src[-1];
    if (bw > 32) {
        if (lane == 0) atomicAdd(status + 1, 1);
        return;
    }
    const int vbytes = (bw + 7) >> 3;
    const uint32_t vmask = vbytes >= 4 ? 0xffffffffu : ((1u << (8 * vbytes)) - 1);
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(s_win);
    constexpr int SMALL = 64;        // runs up to this many values are expanded by the lane that decoded them
    constexpr int LOOKAHEAD = 32 + 5 + SMALL * 4 + 16;
    bool stop = false;
    while (!stop && pos < st.src_bytes && produced < st.max_values) {
        // Span = the next 32 stream bytes.  Lane j decodes a run header AS IF one started at pos + j;
        // which lanes guessed right is then settled by hopping along the chain (one shuffle per run
        // instead of a full serial decode per run), and the lanes on the chain expand their runs.
        ensure((int)pos, LOOKAHEAD);
        const int64_t my = pos + lane;
        const uint64_t w = peek64((int)my);
        uint32_t h = (uint32_t)w & 0x7f;
        int hl = 1;
        if (w & 0x80) {
            int shift = 7;
            while (hl < 5) {
                const uint32_t c = (uint32_t)(w >> (8 * hl)) & 0xff;
                h |= (c & 0x7f) << shift;
                shift += 7;
                ++hl;
                if (!(c & 0x80)) break;
            }
        }
        const bool lit = h & 1;
        const uint32_t units = h >> 1;                      // groups of 8 (bit-packed) or repetitions (RLE)
        const bool sane = units > 0 && units < (1u << 24) && my < st.src_bytes;
        const int cnt = sane ? (int)(lit ? units * 8 : units) : 0;
        const int dlen = lit ? (int)((sane ? units : 0) * (unsigned)bw) : vbytes;
        const int adv = hl + dlen;
        const uint32_t val = hl + vbytes <= 8 ? ((uint32_t)(w >> (8 * hl)) & vmask) : ((uint32_t)peek64((int)my + hl) & vmask);
        unsigned chain = 0;
        int64_t j = 0;
        while (j < 32) {
            const int a = __shfl_sync(0xffffffffu, adv, (int)j);
            const int ok = __shfl_sync(0xffffffffu, (int)sane, (int)j);
            if (!ok) { stop = true; break; }               // end of the stream (or a malformed header): nothing follows
            chain |= 1u << j;
            j += a;
        }
        const bool mine = (chain >> lane) & 1u;
        int c = mine ? cnt : 0;
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        int start = produced + incl - c;
        if (start >= st.max_values) c = 0;
        else if (c > st.max_values - start) c = st.max_values - start;
        const int data = (int)my + hl;
        if (c > 0 && c <= SMALL) {   // every lane on the chain writes its own short run
            if (!lit || bw == 0) {
                const int32_t v = lit ? 0 : (int32_t)val;
                for (int i = 0; i < c; ++i) dst[start + i] = v;
            } else {
                const unsigned base_bit = 8u * (unsigned)(data - win_lo);
                for (int i = 0; i < c; ++i) {
                    const unsigned bit = base_bit + (unsigned)i * (unsigned)bw;
                    const uint32_t v = __funnelshift_r(sw[bit >> 5], sw[(bit >> 5) + 1], bit & 31);
                    dst[start + i] = (int32_t)(bw >= 32 ? v : (v & ((1u << bw) - 1)));
                }
            }
        }
        unsigned big = __ballot_sync(0xffffffffu, c > SMALL);
        while (big) {                // long runs: the whole warp expands them, one after another
            const int f = __ffs(big) - 1;
            big &= big - 1;
            const int f_start = __shfl_sync(0xffffffffu, start, f), f_c = __shfl_sync(0xffffffffu, c, f);
            const int f_lit = __shfl_sync(0xffffffffu, (int)lit, f), f_data = __shfl_sync(0xffffffffu, data, f);
            const uint32_t f_val = __shfl_sync(0xffffffffu, val, f);
            if (!f_lit || bw == 0) {
                const int32_t v = f_lit ? 0 : (int32_t)f_val;
                for (int i = lane; i < f_c; i += 32) dst[f_start + i] = v;
            } else {
                const int nbytes = (int)(((int64_t)f_c * bw + 7) >> 3);
                if (nbytes + 8 <= PQ_WINDOW - 16) {
                    ensure(f_data, nbytes + 8);
                    const unsigned base_bit = 8u * (unsigned)(f_data - win_lo);
                    for (int i = lane; i < f_c; i += 32) {
                        const unsigned bit = base_bit + (unsigned)i * (unsigned)bw;
                        const uint32_t v = __funnelshift_r(sw[bit >> 5], sw[(bit >> 5) + 1], bit & 31);
                        dst[f_start + i] = (int32_t)(bw >= 32 ? v : (v & ((1u << bw) - 1)));
                    }
                } else {             // longer than the window: straight from global memory
                    const uint8_t* p = src + f_data;
                    for (int i = lane; i < f_c; i += 32) dst[f_start + i] = (int32_t)pq_extract(p, (int64_t)i * bw, bw);
                }
            }
        }
        produced = total > st.max_values - produced ? st.max_values : produced + total;
        pos += j;
    }
    if (lane == 0 && st.exact && produced < st.max_values) atomicAdd(status + 1, 1);  // stream ended early
}

// ---------------------------------------------------------------- Snappy
struct PqInflate { int64_t src, src_bytes, dst, dst_bytes; };   // byte offsets in the buffer

// One WARP per page (tried: one CTA per page with every thread walking the chain and all 256 threads moving the
// bytes - 3x slower on pages of integer-valued doubles, whose zero bytes give ~10-byte elements: the chain, not
// the copying, is the cost).  The chain is a latency chain, so what matters is the latency of one element: the
// last PQ_RING output bytes are mirrored in shared memory, and a copy element that reaches no further back than
// that (offsets of 8 or 16 in columns of doubles) reads the mirror instead of making a round trip to L2.
constexpr int PQ_RING = 2048;
__global__ void __launch_bounds__(PQ_THREADS) pq_snappy_kernel(uint8_t* buf, const PqInflate* __restrict__ jobs, int njobs,
                                                               int32_t* status) {
    __shared__ uint8_t s_ring_all[PQ_WARPS][PQ_RING];
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * PQ_WARPS + (threadIdx.x >> 5);
    if (j >= njobs) return;
    uint8_t* ring = s_ring_all[threadIdx.x >> 5];
    const PqInflate job = jobs[j];
    const uint8_t* in = buf + job.src;      // compressed bytes: read only
    uint8_t* out = buf + job.dst;           // this warp's output range: written, and read back by copy elements
    int64_t ip = 0, op = 0;
    // preamble: the uncompressed length as a varint (every lane runs the same scalar code)
    uint64_t n = 0;
    for (int shift = 0; ip < job.src_bytes && shift < 35; shift += 7) {
        const uint8_t c = __ldg(in + ip++);
        n |= (uint64_t)(c & 0x7f) << shift;
        if (!(c & 0x80)) break;
    }
    bool bad = (int64_t)n != job.dst_bytes;
    while (!bad && op < job.dst_bytes && ip < job.src_bytes) {
        const unsigned tag = __ldg(in + ip++);
        if ((tag & 3) == 0) {               // literal
            int64_t len = tag >> 2;
            if (len >= 60) {
                const int extra = (int)len - 59;
                len = 0;
                for (int k = 0; k < extra; ++k) len |= (int64_t)__ldg(in + ip + k) << (8 * k);
                ip += extra;
            }
            len += 1;
            if (ip + len > job.src_bytes || op + len > job.dst_bytes) { bad = true; break; }
            const uint8_t* s = in + ip;
            uint8_t* d = out + op;
            if (len <= 64) {                // the common short literal: one or two bytes per lane
                for (int i = lane; i < len; i += 32) {
                    const uint8_t c = __ldg(s + i);
                    d[i] = c;
                    ring[(op + i) & (PQ_RING - 1)] = c;
                }
            } else {
                // bytes up to 4-byte alignment of the destination, 4-byte words, tail (source words by funnel shift)
                const int head = (int)((4 - (reinterpret_cast<uintptr_t>(d) & 3)) & 3);
                const int64_t h = head < len ? head : len;
                if (lane < h) d[lane] = __ldg(s + lane);
                const int64_t words = (len - h) >> 2;
                uint32_t* dw = reinterpret_cast<uint32_t*>(d + h);
                const uint8_t* sw = s + h;
                int64_t w = lane;
                for (; w + 96 < words; w += 128) {
                    const uint32_t a = pq_load_u32(sw + 4 * w), b = pq_load_u32(sw + 4 * (w + 32)), c = pq_load_u32(sw + 4 * (w + 64)),
                                   e = pq_load_u32(sw + 4 * (w + 96));
                    dw[w] = a; dw[w + 32] = b; dw[w + 64] = c; dw[w + 96] = e;
                }
                for (; w < words; w += 32) dw[w] = pq_load_u32(sw + 4 * w);
                const int64_t done = h + 4 * words;
                if (done + lane < len) d[done + lane] = __ldg(s + done + lane);
                // mirror the literal's last PQ_RING bytes
                const int64_t m0 = len > PQ_RING ? len - PQ_RING : 0;
                for (int64_t i = m0 + lane; i < len; i += 32) ring[(op + i) & (PQ_RING - 1)] = __ldg(s + i);
            }
            ip += len;
            op += len;
        } else {                            // copy of `len` bytes that start `off` bytes back in the output
            int64_t len, off;
            if ((tag & 3) == 1) {
                len = 4 + ((tag >> 2) & 7);
                off = ((int64_t)(tag >> 5) << 8) | __ldg(in + ip);
                ip += 1;
            } else if ((tag & 3) == 2) {
                len = (tag >> 2) + 1;
                off = (int64_t)__ldg(in + ip) | ((int64_t)__ldg(in + ip + 1) << 8);
                ip += 2;
            } else {
                len = (tag >> 2) + 1;
                off = 0;
                for (int k = 0; k < 4; ++k) off |= (int64_t)__ldg(in + ip + k) << (8 * k);
                ip += 4;
            }
            if (off <= 0 || off > op || op + len > job.dst_bytes || ip > job.src_bytes) { bad = true; break; }
            // an overlapping copy (off < len) repeats the last `off` bytes: byte i comes from i mod off.  len <= 64.
            const int64_t from = op - off;
            uint8_t c0 = 0, c1 = 0;
            const bool two = lane + 32 < len;
            if (off <= PQ_RING) {
                if (lane < len) c0 = ring[(from + (lane % off)) & (PQ_RING - 1)];
                if (two) c1 = ring[(from + ((lane + 32) % off)) & (PQ_RING - 1)];
            } else {
                if (lane < len) c0 = __ldcg(out + from + (lane % off));
                if (two) c1 = __ldcg(out + from + ((lane + 32) % off));
            }
            __syncwarp();                   // every source byte is read before the mirror is overwritten
            if (lane < len) { out[op + lane] = c0; ring[(op + lane) & (PQ_RING - 1)] = c0; }
            if (two) { out[op + lane + 32] = c1; ring[(op + lane + 32) & (PQ_RING - 1)] = c1; }
            op += len;
        }
        __syncwarp();                       // the bytes just written may be the source of the next copy
    }
    if (lane == 0 && (bad || op != job.dst_bytes)) atomicAdd(status + 3, 1);
}

// ---------------------------------------------------------------- PLAIN BYTE_ARRAY pages
struct PqStringPage {
    int64_t val_offset;   // byte offset of the page's values: [u32 length][bytes] ...
    int32_t val_bytes;
    int32_t max_values;   // at most this many values (the non-null ones; their exact number is in the levels)
    int64_t out_offset;   // element offset of the page's slot numbers in the int32 output
};

__device__ __forceinline__ uint64_t pq_string_hash(const uint8_t* p, int len) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (int i = 0; i < len; ++i) h = (h ^ (uint64_t)p[i]) * 0x100000001b3ull;
    h = mix64(h ^ (uint64_t)len);
    return h ? h : 1;
}
// representative of a slot: offset << 24 | length (the CSV decoder's packing with the "quoted" bit clear, so the
// host reads the labels back the same way)
__device__ __forceinline__ int pq_dict_slot(uint64_t* hash_tab, uint64_t* rep_tab, uint32_t mask, const uint8_t* buf,
                                            const uint8_t* s, int len, uint64_t h) {
    const uint64_t rep = ((uint64_t)(s - buf) << 24) | (uint64_t)len;
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
            const int rl = (int)(r & 0x7fffff);
            const uint8_t* rp = buf + (r >> 24);
            bool same = rl == len;
            for (int i = 0; same && i < len; ++i) same = rp[i] == s[i];
            if (same) return (int)slot;
        }
        slot = (slot + 1) & mask;
    }
    return -1;
}

__global__ void __launch_bounds__(PQ_THREADS) pq_plain_strings_kernel(const uint8_t* __restrict__ buf,
                                                                      const PqStringPage* __restrict__ pages, int npages,
                                                                      uint64_t* hash_tab, uint64_t* rep_tab, uint32_t mask,
                                                                      int32_t* out, int32_t* status) {
    const int lane = threadIdx.x & 31;
    const int pi = blockIdx.x * PQ_WARPS + (threadIdx.x >> 5);
    if (pi >= npages) return;
    const PqStringPage pg = pages[pi];
    const uint8_t* base = buf + pg.val_offset;
    int64_t pos = 0;
    for (int v0 = 0; v0 < pg.max_values && pos + 4 <= pg.val_bytes; v0 += 32) {
        // lane 0 walks 32 links of the chain; the others receive their value's position and length
        int my_pos = -1, my_len = 0;
        if (lane == 0) my_pos = 0;
        int64_t p = pos;
        int count = 0;
        for (int k = 0; k < 32; ++k) {
            int len = -1;
            if (lane == 0 && v0 + k < pg.max_values && p + 4 <= pg.val_bytes) {
                len = (int)pq_load_u32(base + p);
                if (len < 0 || len >= (1 << 23) || p + 4 + len > pg.val_bytes) len = -2;
            }
            len = __shfl_sync(0xffffffffu, len, 0);
            if (len == -2) { if (lane == 0) atomicAdd(status + 1, 1); len = -1; }
            if (len < 0) break;
            const int64_t at = __shfl_sync(0xffffffffu, p, 0);
            if (lane == k) { my_pos = (int)(at + 4); my_len = len; }
            if (lane == 0) p += 4 + len;
            ++count;
        }
        pos = __shfl_sync(0xffffffffu, p, 0);
        if (lane < count) {
            const uint8_t* sp = base + my_pos;
            const int slot = pq_dict_slot(hash_tab, rep_tab, mask, buf, sp, my_len, pq_string_hash(sp, my_len));
            if (slot < 0) atomicAdd(status + 0, 1);     // dictionary full
            out[pg.out_offset + v0 + lane] = slot;
        }
        if (count < 32) break;
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

// One CTA per data page, in spans of PQ_SPAN values: (1) warp per 32 values: definition levels ->
// validity word + non-null count (no block barrier inside the loop); (2) one block scan of the
// span's counts; (3) warp per 32 values again: rank among the page's non-null values = scanned base
// + popcount below the lane, value fetched (PLAIN) or looked up (dictionary), row written.
constexpr int PQ_SPAN = 32 * 1024;   // values per span: 1024 groups of 32

__global__ void __launch_bounds__(PQ_THREADS) pq_gather_kernel(const PqGatherParams p) {
    __shared__ uint32_t s_bits[PQ_SPAN / 32];
    __shared__ uint32_t s_base[PQ_SPAN / 32];
    __shared__ uint32_t s_warp[PQ_WARPS];
    __shared__ uint32_t s_running;
    const PqPage pg = p.pages[blockIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_running = 0;
    for (int span0 = 0; span0 < pg.num_values; span0 += PQ_SPAN) {
        const int span_n = pg.num_values - span0 < PQ_SPAN ? pg.num_values - span0 : PQ_SPAN;
        const int groups = (span_n + 31) >> 5;
        __syncthreads();   // s_bits / s_base of the previous span are no longer read; s_running is visible
        for (int g = warp; g < groups; g += PQ_WARPS) {
            const int i = span0 + g * 32 + lane;
            const bool ok = i < pg.num_values && (pg.levels_offset < 0 || p.levels[pg.levels_offset + i] > 0);
            const unsigned b = __ballot_sync(0xffffffffu, ok);
            if (lane == 0) s_bits[g] = b;
        }
        __syncthreads();
        // exclusive scan of the groups' non-null counts: PQ_SPAN / 32 / PQ_THREADS consecutive groups per thread
        constexpr int PER = PQ_SPAN / 32 / PQ_THREADS;
        uint32_t c[PER], mine = 0;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int g = threadIdx.x * PER + k;
            c[k] = g < groups ? __popc(s_bits[g]) : 0;
            mine += c[k];
        }
        uint32_t incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t run = s_running + incl - mine;
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < PQ_WARPS; ++w) {
            if (w < warp) run += s_warp[w];
            total += s_warp[w];
        }
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int g = threadIdx.x * PER + k;
            if (g < groups) s_base[g] = run;
            run += c[k];
        }
        __syncthreads();
        if (threadIdx.x == 0) s_running += total;
        for (int g = warp; g < groups; g += PQ_WARPS) {
            const int i = span0 + g * 32 + lane;
            const bool in = i < pg.num_values;
            const unsigned b = s_bits[g];
            const bool ok = (b >> lane) & 1u;
            const uint32_t rank = s_base[g] + __popc(b & lanemask_lt());  // position among the page's non-null values
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
                        if (pg.kind == PQ_KIND_F64) {   // pages are not 8-byte aligned in the file
                            v = __longlong_as_double((long long)pq_load_u64(src + (int64_t)rank * 8));
                        } else if (pg.kind == PQ_KIND_I64) {
                            v = (double)(long long)pq_load_u64(src + (int64_t)rank * 8);
                        } else {
                            const uint32_t u = pq_load_u32(src + (int64_t)rank * 4);
                            v = pg.kind == PQ_KIND_F32 ? (double)__uint_as_float(u) : (double)(int32_t)u;
                        }
                    }
                }
                if (p.string_column) static_cast<int32_t*>(p.out)[row] = code;
                else static_cast<double*>(p.out)[row] = v;
            }
            // validity bits: this group's 32 rows may straddle two bitmap words
            if (lane == 0 && b) {
                const int64_t r0 = pg.row_start + span0 + g * 32;
                const int sh = (int)(r0 & 31);
                atomicOr(p.valid + (r0 >> 5), b << sh);
                if (sh && (b >> (32 - sh))) atomicOr(p.valid + (r0 >> 5) + 1, b >> (32 - sh));
            }
        }
    }
}

}  // namespace etl

using etl::Handle;

extern "C" int etl_parquet_expand_streams(etl_handle* hh, const uint8_t* file, const void* d_streams, int32_t nstreams,
                                          void* d_runs, int32_t* d_out, int32_t* d_status, void* stream_) {
    ETL_REQUIRE(hh && d_status, "parquet_expand_streams: NULL argument");
    ETL_REQUIRE(nstreams >= 0, "parquet_expand_streams: nstreams < 0");
    if (nstreams == 0) return ETL_OK;
    ETL_REQUIRE(file && d_streams && d_out, "parquet_expand_streams: NULL buffer");
    (void)d_runs;  // ABI 1 kept a run table here; runs are expanded as they are parsed now
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    const unsigned grid = (unsigned)((nstreams + etl::PQ_WARPS - 1) / etl::PQ_WARPS);
    etl::pq_hybrid_decode_kernel<<<grid, etl::PQ_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        file, static_cast<const etl::PqStream*>(d_streams), nstreams, d_out, d_status);
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

extern "C" int etl_parquet_inflate_pages(etl_handle* hh, uint8_t* buffer, const void* d_jobs, int32_t njobs, int32_t* d_status,
                                         void* stream_) {
    ETL_REQUIRE(hh && d_status, "parquet_inflate_pages: NULL argument");
    ETL_REQUIRE(njobs >= 0, "parquet_inflate_pages: njobs < 0");
    if (njobs == 0) return ETL_OK;
    ETL_REQUIRE(buffer && d_jobs, "parquet_inflate_pages: NULL buffer");
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    const unsigned grid = (unsigned)((njobs + etl::PQ_WARPS - 1) / etl::PQ_WARPS);
    etl::pq_snappy_kernel<<<grid, etl::PQ_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        buffer, static_cast<const etl::PqInflate*>(d_jobs), njobs, d_status);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}

extern "C" int etl_parquet_plain_strings(etl_handle* hh, const uint8_t* buffer, const void* d_pages, int32_t npages,
                                         uint64_t* dict_hash, uint64_t* dict_rep, int32_t dict_capacity, int32_t* d_out,
                                         int32_t* d_status, void* stream_) {
    ETL_REQUIRE(hh && d_status, "parquet_plain_strings: NULL argument");
    ETL_REQUIRE(npages >= 0, "parquet_plain_strings: npages < 0");
    if (npages == 0) return ETL_OK;
    ETL_REQUIRE(buffer && d_pages && dict_hash && dict_rep && d_out, "parquet_plain_strings: NULL buffer");
    ETL_REQUIRE(dict_capacity >= 2 && (dict_capacity & (dict_capacity - 1)) == 0,
                "parquet_plain_strings: dict_capacity=%d is not a power of two", dict_capacity);
    Handle* h = reinterpret_cast<Handle*>(hh);
    ETL_CUDA(cudaSetDevice(h->device));
    const unsigned grid = (unsigned)((npages + etl::PQ_WARPS - 1) / etl::PQ_WARPS);
    etl::pq_plain_strings_kernel<<<grid, etl::PQ_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        buffer, static_cast<const etl::PqStringPage*>(d_pages), npages, dict_hash, dict_rep, (uint32_t)dict_capacity - 1, d_out,
        d_status);
    ETL_CUDA(cudaGetLastError());
    return ETL_OK;
}
