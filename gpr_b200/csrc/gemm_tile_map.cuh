// Linear tile index -> (ti, tj) of a GEMM grid, shared by the cp.async kernel (dgemm.cu) and the persistent TMA kernel
// (dgemm_tma.cu).
#pragma once
#include <math.h>
#include <stdint.h>

namespace gprb200 {

// linear block index -> (ti, tj), rasterised for L2 reuse: CTAs are issued in groups of GM tile-rows, column-major
// inside a group, so the ~300 co-resident CTAs share GM A-panels and a few dozen B-panels (a ~25 MB working set at
// K = 1024) instead of one A-panel and ~300 B-panels (150 MB > L2), which made the SYRK update re-read its panel
// 7x from DRAM in 128-byte pieces at a 256 KB stride (profiles/r01_ncu_summary.md).
// lower: group g holds tile-rows [g GM, (g+1) GM) and R GM (g+1) tile-columns (R = BM / BN); tiles right of the
// diagonal and rows past the edge exit at once.
constexpr int GM = 8;
template <int BM, int BN>
__device__ __forceinline__ bool map_tile(int lin, int lower, int tiles_m, int tiles_n, int& ti, int& tj) {
    if (!lower) {
        const int width = GM * tiles_n;
        const int group = lin / width;
        const int first = group * GM;
        const int gsize = min(tiles_m - first, GM);
        const int local = lin - group * width;
        ti = first + local % gsize;
        tj = local / gsize;
        return true;
    }
    constexpr int R = BM / BN;
    constexpr int64_t Q = (int64_t)GM * GM * R;  // tiles of group g: Q (g + 1); prefix(g) = Q g (g + 1) / 2
    int gidx = (int)((sqrt(1.0 + 8.0 * (double)lin / (double)Q) - 1.0) * 0.5);
    while (Q * (gidx + 1) * (gidx + 2) / 2 <= lin) ++gidx;
    while (Q * gidx * (gidx + 1) / 2 > lin) --gidx;
    const int local = lin - (int)(Q * gidx * (gidx + 1) / 2);
    ti = gidx * GM + local % GM;
    tj = local / GM;
    return ti < tiles_m && tj < tiles_n && (int64_t)tj * BN <= (int64_t)ti * BM + BM - 1;
}

// Compact enumeration of the lower-triangle tiles of a square grid of equal tiles (BM == BN), for persistent kernels
// that stride through the list: the same group order as map_tile (groups of GM tile-rows, column-major inside a group),
// but the tiles right of the diagonal are not in the list at all, so a static round-robin over v gives every CTA the
// same number of real tiles. Group g holds 64 g full tiles + the 36 lower tiles of its diagonal 8 x 8 block.
__host__ __device__ __forceinline__ int64_t lower_compact_count(int64_t tiles_m) {
    const int64_t ng = (tiles_m + GM - 1) / GM;
    return 32 * ng * ng + 4 * ng;   // sum_{h < ng} (64 h + 36); rows past tiles_m are filtered by the caller
}
__device__ __forceinline__ bool map_tile_lower_compact(int v, int tiles_m, int& ti, int& tj) {
    int g = (int)((sqrt(1.0 + 8.0 * (double)v) - 1.0) * 0.0625);   // root of 32 g^2 + 4 g = v, then corrected
    while (32 * (g + 1) * (g + 1) + 4 * (g + 1) <= v) ++g;
    while (32 * g * g + 4 * g > v) --g;
    int local = v - (32 * g * g + 4 * g);
    if (local < 64 * g) {
        ti = g * GM + (local & 7);
        tj = local >> 3;
    } else {
        local -= 64 * g;   // 0..35: lower triangle of the diagonal block, column-major (column c holds rows c..7)
        int c = 0;
        while (local >= 8 - c) {
            local -= 8 - c;
            ++c;
        }
        ti = g * GM + c + local;
        tj = g * GM + c;
    }
    return ti < tiles_m;
}

}  // namespace gprb200
