// md5_strided_tma.cuh -- K1/K2 for equal-sized units laid out at a fixed stride, fed by TMA.
//
// This is the shape of both headline workloads: a packed workdir of equal small files (C2) and the
// parts of one S3 multipart object (C3) -- unit u is bytes [u*stride, u*stride + unit_len) of one
// buffer.  Seen as a 2-D byte tensor (row = unit, column = byte within the unit) the next 128 bytes
// of a warp's 32 units are one box of 32 rows x 128 columns, so ONE thread of the warp fetches the
// whole stage with a single tensor copy (cp.async.bulk.tensor.2d, SASS UTMALDG) that reports to an
// mbarrier; nobody computes addresses, predicates or copies 16 bytes at a time.  Per-row bulk copies
// were measured too slow (one TMA request per 128 bytes: 3.2 TB/s); one request per 4 KiB is not.
//
// The box lands densely (row pitch 128 B).  Thread t consumes row t, which would put all 32 threads
// on the same banks; the tensor map therefore asks for the 128-byte swizzle (16-byte chunk index
// XOR row mod 8), and the thread undoes it in its LDS.128 addresses: a quarter warp then touches
// eight different chunks = all 32 banks once.
//
// All units have the same length, so the whole warp runs the same number of blocks and the padding
// is warp-uniform.  Rows past the last unit and bytes past unit_len are out of bounds for the
// tensor and arrive as zeros without touching memory.
#pragma once
#include <cuda.h>

#include "md5_lanes.cuh"

namespace tpimd5 {

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tmap, uint32_t c0, uint32_t c1, uint32_t mbar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
        "l"(tmap), "r"(c0), "r"(c1), "r"(mbar)
        : "memory");
}

// The same for a 3-D view of the buffer -- (byte within a 128-byte chunk, unit, chunk index) -- whose box is 128 bytes x
// 32 units x TILES chunks: ONE request fetches a whole stage, landing as TILES consecutive swizzled tiles.
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *tmap, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t mbar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
        "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(mbar)
        : "memory");
}

// TILES: boxes per stage.  1 keeps shared memory small (20 warps/SM for throughput-bound launches);
// 2 or 4 halve / quarter the per-stage costs that sit on a lone chain's critical path (mbarrier wait,
// TMA issue, loop) for launches with too few chains to fill the SMs (S3 parts, large files).
template <int WARPS, int NSTAGE, int TILES = 1>
struct StridedCfg {
    static constexpr int TILE_BYTES = 32 * 128;                     // one box
    static constexpr int STAGE_BYTES = TILES * TILE_BYTES;
    static constexpr int WARP_BYTES = NSTAGE * STAGE_BYTES;
    static constexpr int DATA_BYTES = WARPS * WARP_BYTES;
    static constexpr int SMEM_BYTES = DATA_BYTES + WARPS * NSTAGE * 8 + 1024;   // + mbarriers + slack to align to 1 KiB
};

template <int WARPS, int MINB, int NSTAGE, int MODE, int TILES = 1>
__global__ void __launch_bounds__(WARPS * 32, MINB)
md5_strided_tma_kernel(const __grid_constant__ CUtensorMap tmap, uint64_t n, uint32_t unit_len, uint4 *__restrict__ table) {
    using Cfg = StridedCfg<WARPS, NSTAGE, TILES>;
    constexpr uint32_t SB = 128u * TILES;                     // bytes of one unit per stage
    constexpr uint32_t BPS = 2u * TILES;                      // MD5 blocks per stage
    extern __shared__ uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31u;
    // broadcast from lane 0 so that the compiler knows the warp index (and all that follows from it: tile
    // and mbarrier addresses, TMA coordinates) is warp-uniform and keeps it in uniform registers
    const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const uint64_t u0 = ((uint64_t)blockIdx.x * WARPS + warp) * 32u;         // first unit of this warp
    if (u0 >= n) return;                                                      // warp-uniform
    const uint64_t u = u0 + lane;

    // swizzled tiles must start on 1 KiB boundaries
    const uint32_t sbase = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;
    const uint32_t wbase = sbase + warp * Cfg::WARP_BYTES;
    const uint32_t mbar0 = sbase + Cfg::DATA_BYTES + warp * (NSTAGE * 8u);
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NSTAGE; ++i) mbar_init(mbar0 + 8u * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    const uint32_t nst = (unit_len + SB - 1u) / SB;           // stages that hold data
    const uint32_t full_st = unit_len / SB;                   // stages whose blocks are all whole
    const uint32_t bm = unit_len >> 6;                        // block that takes the 0x80 marker
    const uint32_t nb = ((unit_len + 8u) >> 6) + 1u;          // blocks incl. padding

    // Lane 0 only.  The buffer being refilled was last read by the warp before the __syncwarp that
    // precedes this call, and those reads have completed (their values were consumed by the MD5
    // steps), so the TMA write cannot overtake them; no proxy fence is needed for this direction.
    auto fetch = [&](uint32_t s) {
        const uint32_t buf = s % NSTAGE;
        mbar_expect_tx(mbar0 + 8u * buf, Cfg::STAGE_BYTES);
#pragma unroll
        for (int t = 0; t < TILES; ++t)                       // a box wholly past unit_len arrives as zeros
            tma_load_2d(wbase + buf * Cfg::STAGE_BYTES + t * Cfg::TILE_BYTES, &tmap, (s * TILES + t) << 7, (uint32_t)u0,
                        mbar0 + 8u * buf);
    };
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < NSTAGE - 1; ++s)
            if ((uint32_t)s < nst) fetch(s);
    }

    uint32_t st[4];
    md5_iv(st);
    // this thread's row, and the un-swizzle of its eight 16-byte chunks: chunk c sits at (c ^ (row & 7))
    const uint32_t rowoff = lane * 128u;
    const uint32_t swz = (lane & 7u) << 4;
    uint32_t xo[8];                                           // byte offset of chunk c inside a stage buffer
#pragma unroll
    for (int c = 0; c < 8; ++c) xo[c] = rowoff + (((uint32_t)c << 4) ^ swz);

    // ---- whole stages ---------------------------------------------------------------------------
#pragma unroll 1
    for (uint32_t s = 0; s < full_st; ++s) {
        if (lane == 0 && s + (NSTAGE - 1) < nst) fetch(s + (NSTAGE - 1));
        const uint32_t buf = s % NSTAGE;
        mbar_wait(mbar0 + 8u * buf, (s / NSTAGE) & 1u);
        // All 128 bytes of a row are read up front (8 x LDS.128), so the shared-memory latency is paid
        // once per box instead of once per load -- it is exposed when few chains share a scheduler
        // (S3 parts, large files), and costs only registers when many do.
        constexpr int GROUP = TILES >= 2 ? 2 : 1;             // boxes whose words are loaded before any is hashed
#pragma unroll
        for (int t0 = 0; t0 < TILES; t0 += GROUP) {
            uint32_t m[GROUP * 2][16];
#pragma unroll
            for (int g = 0; g < GROUP; ++g) {
                const uint32_t tile = wbase + buf * Cfg::STAGE_BYTES + (t0 + g) * Cfg::TILE_BYTES;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                                 : "=r"(m[2 * g + (q >> 2)][4 * (q & 3)]), "=r"(m[2 * g + (q >> 2)][4 * (q & 3) + 1]),
                                   "=r"(m[2 * g + (q >> 2)][4 * (q & 3) + 2]), "=r"(m[2 * g + (q >> 2)][4 * (q & 3) + 3])
                                 : "r"(tile + xo[q]));
            }
#pragma unroll
            for (int k = 0; k < GROUP * 2; ++k) md5_block<MODE>(st, m[k]);
        }
        __syncwarp();                                         // all 32 rows consumed before the buffer is refilled
    }

    // ---- tail: a partial stage (if unit_len is not a multiple of the stage) and the padding -----------
#pragma unroll 1
    for (uint32_t b = full_st * BPS; b < nb; ++b) {
        const uint32_t s = b / BPS, within = b % BPS, j = within & 1u;
        const uint32_t buf = s % NSTAGE;
        if (within == 0 && s < nst) {
            if (lane == 0 && s + (NSTAGE - 1) < nst) fetch(s + (NSTAGE - 1));
            mbar_wait(mbar0 + 8u * buf, (s / NSTAGE) & 1u);
        }
        uint32_t m[16];
        if (b <= bm && s < nst) {
            const uint32_t row = wbase + buf * Cfg::STAGE_BYTES + (within >> 1) * Cfg::TILE_BYTES + rowoff;
#pragma unroll
            for (int q = 0; q < 4; ++q)
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(m[4 * q]), "=r"(m[4 * q + 1]), "=r"(m[4 * q + 2]), "=r"(m[4 * q + 3])
                             : "r"(row + (((j * 4u + q) << 4) ^ swz)));
        } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) m[q] = 0;
        }
        if (b >= bm) md5_pad_block(m, unit_len & 63u, b > bm, (uint64_t)unit_len);
        md5_block<MODE>(st, m);
        if (within == BPS - 1) __syncwarp();
    }
    if (u < n) table[u] = make_uint4(st[0], st[1], st[2], st[3]);
}

}  // namespace tpimd5
