// groupBy on a small known key range (config 3: 1 K groups) WITHOUT a floating-point atomic per cell.
//
// groupby_dense_kernel pays one fp64 shared-memory add per (row, value column), and fp64 has no native
// shared atomic on sm_100a: ncu puts its ATOMS.CAST.SPIN.64 at ~17 shared wavefronts per warp instruction
// and the kernel at 94 % of the shared-memory pipe with DRAM at 36 %.  Here a tile of rows is SORTED by key
// first, with the one atomic that is native (32-bit integer add -> the row's rank within its key):
//   1. tile (keys, value columns, validity words) -> shared memory by 1-D bulk async copies, double buffered;
//   2. rank of every row within (tile, key): native shared atomic; exclusive scan of the per-key counts;
//   3. one packed word {key, row, validity bits} per row stored at its sorted position;
//   4. in sorted order, 32 positions per warp: values gathered from the tile, a warp-level SEGMENTED scan over
//      the key runs (shuffles, no shared memory), and the last lane of every run adds the run's totals to the
//      CTA's table - with plain loads and stores when the run lies strictly inside the warp's 32 positions
//      (nobody else can hold that key in this tile), with atomics only for the runs that touch the warp's edge.
// With ~2 rows per (key, tile) that is ~0.2 fp64 atomics per row instead of 3.
#pragma once
#include "common.cuh"

namespace etl {

constexpr int DS_TILE = 2048;
constexpr int DS_THREADS = 1024;
constexpr int DS_SLOTS = 1024;        // keys 0 .. R-1 (R <= 1023) + slot R for the null key
constexpr unsigned DS_NONE = 0xffffffffu;

template <int NC>
struct DenseSortedSmem {
    // one input buffer: keys | NC value columns | (NC + 1) x 64 validity words (key bitmap last)
    static constexpr size_t kKeys = (size_t)DS_TILE * 4;
    static constexpr size_t kVals = (size_t)DS_TILE * 8;
    static constexpr size_t kBits = 64 * 4;
    static constexpr size_t kBuffer = kKeys + NC * kVals + (NC + 1) * kBits;
    static constexpr size_t kPerm = 2 * kBuffer;                       // u32[DS_TILE]
    static constexpr size_t kCnt = kPerm + (size_t)DS_TILE * 4;        // u32[DS_SLOTS]
    static constexpr size_t kOff = kCnt + DS_SLOTS * 4;                // u32[DS_SLOTS]
    static constexpr size_t kSum = kOff + DS_SLOTS * 4;                // f64[NC][DS_SLOTS]
    static constexpr size_t kRows = kSum + (size_t)NC * DS_SLOTS * 8;  // u32[DS_SLOTS]
    static constexpr size_t kNull = kRows + DS_SLOTS * 4;              // u32[NC][DS_SLOTS]
    static constexpr size_t kMisc = kNull + (size_t)NC * DS_SLOTS * 4; // u32[32] warp totals, u32 total, 2 mbarriers
    static constexpr size_t kTotal = kMisc + 32 * 4 + 16 + 16;
};

}  // namespace etl
