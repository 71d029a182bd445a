// Work split of the persistent decode kernel (decode_mega.cu), shared by the device code and the
// host-side planner (dl_decode_mega_plan): every CTA owns a contiguous row range of each matrix over
// the full K extent.  Row ranges are multiples of `align` rows (2: RoPE pairs / (gate, up) pairs;
// 32: the Q80 blocks of the wo / w2 partial sums) and are cut into pieces at the 128-row Q40T tile
// boundaries; a piece of nr rows packs kpack consecutive k-tiles into one 9216-byte pipeline stage.
// (reference: the per-thread row split SPLIT_THREADS, src/nn/nn-quants.hpp:82-86, and the node split
// sliceRowMatmul / sliceColMatmul, src/nn/nn-core.cpp:207-230 -- here one level, over 148 SMs.)
#pragma once

#if defined(__CUDACC__)
#define MG_HD __host__ __device__ __forceinline__
#else
#define MG_HD inline
#endif

MG_HD int mg_rows_per_cta(int d, int grid, int align) {
    const int per = (d + grid - 1) / grid;
    return (per + align - 1) / align * align;
}

// Matrices with no more row tiles than CTAs are cut tile by tile (grid / tiles CTAs share one 128-row tile), so no CTA
// straddles a tile boundary and pays the per-piece reduction + epilogue twice; bigger ones get equal contiguous shares.
MG_HD void mg_cta_rows(int d, int grid, int align, int cta, int& r0, int& r1) {
    const int tiles = (d + 127) / 128;
    if (tiles <= grid) {
        const int cpt = grid / tiles;                                  // CTAs per tile
        int rpc = (128 + cpt - 1) / cpt;
        rpc = (rpc + align - 1) / align * align;                       // rows per CTA inside a tile
        const int tile = cta / cpt, sub = cta - tile * cpt;
        if (tile >= tiles) { r0 = r1 = d; return; }
        const int tend = (tile * 128 + 128 < d) ? tile * 128 + 128 : d;
        const int a0 = tile * 128 + sub * rpc, a1 = a0 + rpc;
        r0 = a0 < tend ? a0 : tend;
        r1 = a1 < tend ? a1 : tend;
        return;
    }
    const long long rpc = mg_rows_per_cta(d, grid, align);
    const long long a0 = rpc * cta, a1 = a0 + rpc;
    r0 = (int)(a0 < d ? a0 : d);
    r1 = (int)(a1 < d ? a1 : d);
}

// k-tiles of a piece of nr rows that share one pipeline stage (nr * kpack <= 128 consumer threads of a group).  At most
// 4: each (group, k-tile) pair leaves a partial sum per row, and the per-piece reduction should stay short.
MG_HD int mg_kpack(int nr) {
    const int k = 128 / nr;
    return k < 1 ? 1 : (k > 4 ? 4 : k);
}
