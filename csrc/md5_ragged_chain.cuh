// md5_ragged_chain.cuh -- the latency kernel for RAGGED batches: units of any length, few of them.
//
// What the host pipeline sends when a workdir's files differ in size and there are not enough of them to
// fill the SMs (BASELINE config 1: 1,000 files of 1 KiB..8 MiB; the large files of config 4; streams):
// each MD5 chain then runs at the latency of its three dependent ops per step, and the batch takes as long as
// its longest file's chain.  md5_lanes_kernel spends ~970 cycles per 64-byte block on such a chain (per-stage
// votes, per-block predicates, cp.async issue by the hashing threads); this kernel keeps the structure of
// md5_chain_tma.cuh (781) for units that are not equal-sized:
//   * a producer warp per CTA moves the data: lane r issues ONE bulk copy (cp.async.bulk, up to SB bytes) per
//     stage for row r of each hashing warp, completion on the stage's "full" mbarrier (every producer lane
//     arrives with its own byte count); rows are SB+16 bytes apart so that LDS.128 by row is conflict-free;
//   * hashing warps run the branch-free U-block loop of the equal-sized kernel for as many iterations as EVERY
//     lane that is still busy has whole blocks for (one warp reduction tells how many); lanes whose unit has
//     ended sit the loop out;
//   * whenever some lane's unit ends inside the next U blocks, ONE iteration takes the general path (partial
//     block, padding, per-lane predicates), then the fast loop resumes with the lanes that are left.  A warp of 32
//     units takes that path at most 32 + a few times, whatever the lengths.
// Units must start on 16-byte boundaries (the staging packer guarantees it; a bulk copy moves whole 16-byte
// granules, so up to 15 bytes past a unit's end are read -- inside the staging buffer).  Chaining across
// launches (prior bytes, CONT/FINAL) as in md5_lanes_kernel.
#pragma once
#include "md5_chain_tma.cuh"

namespace tpimd5 {

template <int WARPS, int NSTAGE, int SB>
struct RaggedCfg {
    static constexpr int PITCH = SB + 16;                           // row pitch: +16 B skews the banks
    static constexpr int STAGE_BYTES = 32 * PITCH;
    static constexpr int WARP_BYTES = NSTAGE * STAGE_BYTES;
    static constexpr int DATA_BYTES = WARPS * WARP_BYTES;
    static constexpr int SMEM_BYTES = DATA_BYTES + 2 * WARPS * NSTAGE * 8 + 128;
    static constexpr int THREADS = (WARPS + 1) * 32;
    static_assert((NSTAGE & (NSTAGE - 1)) == 0 && (SB == 256 || SB == 512), "ring arithmetic");
};

__device__ __forceinline__ void mbar_arrive_tx(uint32_t mbar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}

template <int WARPS, int MINB, int NSTAGE, int SB, int MODE, int U = 4>
__global__ void __launch_bounds__((WARPS + 1) * 32, MINB) md5_ragged_chain_kernel(const LaneArgs p) {
    using Cfg = RaggedCfg<WARPS, NSTAGE, SB>;
    constexpr uint32_t BPS = SB / 64u;                         // MD5 blocks per stage
    constexpr uint32_t LOG_BPS = SB == 256 ? 2 : 3;
    constexpr uint32_t LOG_NST = NSTAGE == 1 ? 0 : NSTAGE == 2 ? 1 : 2;
    static_assert((uint32_t)U <= BPS && (U == 2 || U == 4), "an iteration must not span stages");
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const uint64_t cta_u0 = (uint64_t)blockIdx.x * (WARPS * 32u);
    const uint32_t sbase = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 127u) & ~127u;
    const uint32_t bars = sbase + Cfg::DATA_BYTES;             // full[w][s] then empty[w][s]
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < WARPS * NSTAGE; ++i) {
            mbar_init(bars + 8u * i, 32);                                 // full: one arrival (with its byte count) per producer lane
            mbar_init(bars + 8u * (WARPS * NSTAGE + i), BPS / U);         // empty: one arrival per loop iteration
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ---- producer warp: lane r feeds row r of every hashing warp ------------------------------------
    if (warp == WARPS) {
        const uint8_t *src[WARPS];
        uint32_t l16[WARPS], nst[WARPS];
        uint32_t nst_max = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            const uint64_t u = cta_u0 + (uint32_t)w * 32u + lane;
            const bool valid = u < p.n;
            src[w] = p.base + (valid ? p.off[u] : 0);
            l16[w] = valid ? (uint32_t)((p.len[u] + 15u) & ~(uint64_t)15) : 0u;
            nst[w] = warp_max((l16[w] + SB - 1u) / SB);
            nst_max = nst[w] > nst_max ? nst[w] : nst_max;
        }
#pragma unroll 1
        for (uint32_t k = 0; k < nst_max; ++k) {
            const uint32_t buf = k & (NSTAGE - 1u);
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                if (k >= nst[w]) continue;                                // warp-uniform
                const uint32_t full = bars + 8u * ((uint32_t)w * NSTAGE + buf);
                const uint32_t empty = full + 8u * (WARPS * NSTAGE);
                if (k >= (uint32_t)NSTAGE) mbar_wait(empty, ((k / NSTAGE) - 1u) & 1u);   // the warp has read this buffer's previous stage
                const uint32_t done = k * SB;
                const uint32_t bytes = l16[w] > done ? (l16[w] - done < (uint32_t)SB ? l16[w] - done : (uint32_t)SB) : 0u;
                if (bytes) {
                    mbar_arrive_tx(full, bytes);
                    bulk_g2s(sbase + (uint32_t)w * Cfg::WARP_BYTES + buf * Cfg::STAGE_BYTES + lane * Cfg::PITCH, src[w] + done, bytes, full);
                } else {
                    mbar_arrive(full);
                }
            }
        }
        return;
    }

    // ---- hashing warps ------------------------------------------------------------------------------
    const uint64_t u = cta_u0 + warp * 32u + lane;
    if (cta_u0 + warp * 32u >= p.n) return;                    // warp-uniform
    const bool valid = u < p.n;
    uint64_t len = 0, prior = 0;
    uint32_t fl = LANE_FINAL, oidx = (uint32_t)u;
    if (valid) {
        len = p.len[u];
        if (p.prior) prior = p.prior[u];
        if (p.flags) fl = p.flags[u];
        if (p.out_idx) oidx = p.out_idx[u];
    }
    const bool fin = (fl & LANE_FINAL) != 0;
    uint32_t st[4];
    if (valid && (fl & LANE_CONT)) {
        const uint4 v = p.table[oidx];
        st[0] = v.x; st[1] = v.y; st[2] = v.z; st[3] = v.w;
    } else {
        md5_iv(st);
    }
    const uint32_t bm = (uint32_t)(len >> 6);                  // whole data blocks; block bm takes the 0x80 marker
    const uint32_t rem = (uint32_t)(len & 63u);
    const uint32_t nb = !valid ? 0u : (fin ? (uint32_t)((len + 8) >> 6) + 1u : bm);
    const uint32_t l16 = valid ? (uint32_t)((len + 15u) & ~(uint64_t)15) : 0u;
    const uint32_t warp_nst = warp_max((l16 + SB - 1u) / SB);  // stages the producer fills for this warp
    const uint32_t warp_nb = warp_max(nb);
    const uint32_t leader = (uint32_t)__ffs(__ballot_sync(0xffffffffu, nb == warp_nb)) - 1u;   // busy to the very end: does the barrier arrivals
    const uint64_t total = prior + len;

    const uint32_t full0 = bars + 8u * (warp * NSTAGE);
    const uint32_t empty0 = full0 + 8u * (WARPS * NSTAGE);
    const uint32_t row = sbase + warp * Cfg::WARP_BYTES + lane * Cfg::PITCH;
    auto boff = [&](uint32_t b) { return ((b >> LOG_BPS) & (NSTAGE - 1u)) * Cfg::STAGE_BYTES + ((b & (BPS - 1u)) << 6); };
    auto lds16 = [&](uint32_t (&m)[16], uint32_t addr) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(m[4 * q]), "=r"(m[4 * q + 1]), "=r"(m[4 * q + 2]), "=r"(m[4 * q + 3])
                         : "r"(addr + 16u * q)
                         : "memory");
    };
    auto full_bar = [&](uint32_t S) { return full0 + ((S & (NSTAGE - 1u)) << 3); };
    auto wait_stage = [&](uint32_t S) { mbar_wait(full_bar(S), (S >> LOG_NST) & 1u); };

    uint32_t b = 0;                                            // blocks done; the same for every lane still busy
#pragma unroll 1
    for (;;) {
        const bool alive = b < nb;
        const uint32_t alive_mask = __ballot_sync(0xffffffffu, alive);
        if (!alive_mask) break;
        // iterations the fast loop can run: every busy lane must have U whole blocks plus the prefetched one
        const uint32_t mine = alive ? (bm > b ? (bm - b - 1u) / U : 0u) : 0x7fffffffu;
        uint32_t iters = __reduce_min_sync(0xffffffffu, mine);
        if (iters && alive) {
            uint32_t mA[16], mB[16];
            wait_stage(b >> LOG_BPS);                          // returns at once for a stage already waited for
            lds16(mA, row + boff(b));
#pragma unroll 1
            do {
                const uint32_t s2 = (b + U) >> LOG_BPS;        // stage of block b+U
                const uint32_t landed = mbar_try_wait(full_bar(s2), (s2 >> LOG_NST) & 1u);
                const uint32_t at = row + boff(b), at_next = row + boff(b + U);
                --iters;
#pragma unroll
                for (int j = 0; j < U; j += 2) {
                    lds16(mB, at + (uint32_t)(j + 1) * 64u);
                    md5_block<MODE>(st, mA);
                    if (j + 2 < U) {
                        lds16(mA, at + (uint32_t)(j + 2) * 64u);
                    } else {
                        if (!landed) wait_stage(s2);
                        lds16(mA, at_next);
                    }
                    md5_block<MODE>(st, mB);
                }
                __syncwarp(alive_mask);                        // the busy lanes' rows of these blocks are in registers (and consumed)
                if (lane == leader) mbar_arrive(empty0 + (((b >> LOG_BPS) & (NSTAGE - 1u)) << 3));
                b += U;
            } while (iters);
        }
        b = __shfl_sync(0xffffffffu, b, leader);               // lanes that sat out catch up
        // ---- one general iteration: some unit ends in blocks [b, b+U) (or the fast loop had nothing to run) ----
        const uint32_t S = b >> LOG_BPS;
        if (S < warp_nst) wait_stage(S);
#pragma unroll 1
        for (uint32_t j = 0; j < (uint32_t)U; ++j) {
            const uint32_t blk = b + j;
            if (blk < nb) {
                uint32_t m[16];
                if ((uint64_t)blk * 64u < len) {
                    lds16(m, row + boff(blk));
                } else {
#pragma unroll
                    for (int q = 0; q < 16; ++q) m[q] = 0;
                }
                if (fin && blk >= bm) md5_pad_block(m, rem, blk > bm, total);
                md5_block<MODE>(st, m);
            }
        }
        __syncwarp();
        if (lane == leader) mbar_arrive(empty0 + ((S & (NSTAGE - 1u)) << 3));
        b += U;
    }
    if (valid) p.table[oidx] = make_uint4(st[0], st[1], st[2], st[3]);
}

}  // namespace tpimd5
