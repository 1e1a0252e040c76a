// md5_chain_tma.cuh -- K2 (and K1/K3 for large files) when chains are scarce: latency per chain is all that counts.
//
// A launch of S3 parts ([rclone] backend/s3 uploadMultipart: one md5.Sum per 8 MiB part of a 64 GiB
// object = 8,192 chains), of a few thousand multi-MiB files, or of the single digest-of-digests message
// cannot fill 148 SMs with independent MD5 chains: most schedulers hold one warp or none.  Each chain
// then advances at the latency of its three dependent integer ops per step (LOP3 -> IADD3 -> LEA.HI,
// 12 cycles, 768 per 64-byte block), and whatever else the hashing warp executes between two blocks is
// added to the job's run time one for one.  This kernel therefore takes everything off that warp:
//
//   * a PRODUCER warp per CTA owns the data movement.  It issues the tensor copies
//     (cp.async.bulk.tensor.2d, one box = the next 128 bytes of a warp's 32 units) for all hashing
//     warps of the CTA, NSTAGE stages ahead, and is the only one that ever blocks on a barrier.
//   * hashing warps wait on a "full" mbarrier per stage.  The wait for the next stage is issued one
//     block early (try_wait, result consumed a block later), so its ~90 cycles run under 64 MD5 steps.
//     When a stage has been read, lane 0 arrives on the stage's "empty" mbarrier (one instruction,
//     release semantics, never blocks); the producer waits on that before it refills the buffer.  The
//     write-after-read ordering is thus carried by the barrier pair, as in any TMA pipeline.
//   * the 16 message words of block b+1 are fetched from shared memory (4 x LDS.128, un-swizzled
//     addresses) before block b's 64 steps start: the LDS latency is hidden as well.
//   * the loop body is U = 2 or 4 blocks (9 or 18 KiB of SASS, inside the 32 KiB L1.5 instruction cache)
//     and has no branch but its own back edge; all per-stage bookkeeping is unconditional arithmetic.
//
// Layout, tensor map and swizzle are those of md5_strided_tma.cuh: unit u = bytes [u*stride,
// u*stride + unit_len) of one buffer.  Units of any length: a launch hashes at most 2^30 bytes of each
// (one tensor-map row) and long units continue from launch to launch through the digest table.
#pragma once
#include <cuda.h>

#include "md5_strided_tma.cuh"

namespace tpimd5 {

__device__ __forceinline__ void mbar_arrive(uint32_t mbar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar) : "memory");
}
// One poll (the hardware may sleep the warp for a bounded time): 1 = the phase with this parity is over.
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t mbar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(mbar), "r"(parity)
        : "memory");
    return done;
}

template <int WARPS, int NSTAGE, int TILES>
struct ChainCfg {
    static constexpr int TILE_BYTES = 32 * 128;                      // one box: 32 units x 128 bytes
    static constexpr int STAGE_BYTES = TILES * TILE_BYTES;
    static constexpr int WARP_BYTES = NSTAGE * STAGE_BYTES;
    static constexpr int DATA_BYTES = WARPS * WARP_BYTES;
    static constexpr int SMEM_BYTES = DATA_BYTES + 2 * WARPS * NSTAGE * 8 + 1024;   // + full/empty mbarriers + slack to align to 1 KiB
    static constexpr int THREADS = (WARPS + 1) * 32;                  // hashing warps + the producer warp
    static_assert((NSTAGE & (NSTAGE - 1)) == 0 && (TILES & (TILES - 1)) == 0, "ring arithmetic wants powers of two");
};

template <int WARPS, int MINB, int NSTAGE, int TILES, int MODE, int U = 2>
__global__ void __launch_bounds__((WARPS + 1) * 32, MINB)
md5_chain_tma_kernel(const __grid_constant__ CUtensorMap tmap_main, uint64_t n_main, uint32_t unit_len_main, uint4 *__restrict__ table_main,
                     uint64_t prior, uint32_t flags, const __grid_constant__ CUtensorMap tmap_tail, uint32_t tail_len) {
    // The buffer's end usually cuts the last unit short (the last part of an S3 object).  That unit cannot be a row
    // of the main tensor (its row would reach past the buffer), so it is a tensor of its own, hashed by one extra
    // CTA of the SAME launch (the last one): it runs next to the others instead of after them.
    const bool is_tail = tail_len != 0u && blockIdx.x == gridDim.x - 1u;
    const CUtensorMap *const tmap = is_tail ? &tmap_tail : &tmap_main;
    const uint64_t n = is_tail ? 1u : n_main;
    const uint32_t unit_len = is_tail ? tail_len : unit_len_main;
    uint4 *const table = is_tail ? table_main + n_main : table_main;
    using Cfg = ChainCfg<WARPS, NSTAGE, TILES>;
    constexpr uint32_t SB = 128u * TILES;                      // bytes of one unit per stage
    constexpr uint32_t BPS = 2u * TILES;                       // MD5 blocks per stage
    constexpr uint32_t RB = BPS * NSTAGE;                      // blocks in the whole ring of a warp
    static_assert((U == 2 || U == 4) && (uint32_t)U <= BPS, "an iteration must not span stages");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);     // warp-uniform by construction
    const uint64_t cta_u0 = is_tail ? 0u : (uint64_t)blockIdx.x * (WARPS * 32u);

    const uint32_t sbase = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;   // swizzled boxes: 1 KiB aligned
    const uint32_t bars = sbase + Cfg::DATA_BYTES;             // full[w][s] then empty[w][s]
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < WARPS * NSTAGE; ++i) {
            mbar_init(bars + 8u * i, 1);                                  // full: the producer's expect_tx arrival
            mbar_init(bars + 8u * (WARPS * NSTAGE + i), BPS / U);         // empty: one arrival per loop iteration
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t nst = (unit_len + SB - 1u) / SB;            // stages that hold data

    // ---- producer warp ----------------------------------------------------------------------------
    if (warp == WARPS) {
        if (lane != 0) return;
        const uint64_t left = n - cta_u0;                      // the launch never starts a CTA past n
        const uint32_t nw = left >= WARPS * 32u ? WARPS : (uint32_t)((left + 31u) >> 5);
#pragma unroll 1
        for (uint32_t k = 0; k < nst; ++k) {
            const uint32_t buf = k & (NSTAGE - 1u);
#pragma unroll 1
            for (uint32_t w = 0; w < nw; ++w) {
                const uint32_t full = bars + 8u * (w * NSTAGE + buf);
                const uint32_t empty = full + 8u * (WARPS * NSTAGE);
                if (k >= (uint32_t)NSTAGE) mbar_wait(empty, ((k / NSTAGE) - 1u) & 1u);   // the warp has read this buffer's previous stage
                mbar_expect_tx(full, Cfg::STAGE_BYTES);
                const uint32_t dst = sbase + w * Cfg::WARP_BYTES + buf * Cfg::STAGE_BYTES;
                if (TILES > 1 && (flags & LANE_BOX3D) && !is_tail) {
                    // unit_len is a multiple of 128: the buffer seen as (byte in chunk, unit, chunk) -- one request per stage
                    tma_load_3d(dst, tmap, 0u, (uint32_t)(cta_u0 + w * 32u), k * TILES, full);
                } else {
#pragma unroll
                    for (int t = 0; t < TILES; ++t)            // a box wholly past unit_len (or past n) arrives as zeros
                        tma_load_2d(dst + t * Cfg::TILE_BYTES, tmap, (k * TILES + t) << 7, (uint32_t)(cta_u0 + w * 32u), full);
                }
            }
        }
        return;
    }

    // ---- hashing warps ------------------------------------------------------------------------------
    const uint64_t u0 = cta_u0 + warp * 32u;
    if (u0 >= n) return;                                       // warp-uniform
    const uint64_t u = u0 + lane;
    const uint32_t full0 = bars + 8u * (warp * NSTAGE);
    const uint32_t empty0 = full0 + 8u * (WARPS * NSTAGE);
    const uint32_t wbase = sbase + warp * Cfg::WARP_BYTES;

    const uint32_t bm = unit_len >> 6;                         // whole data blocks; block bm takes the 0x80 marker
    const uint32_t nb = ((unit_len + 8u) >> 6) + 1u;           // blocks incl. padding
    const uint32_t rem = unit_len & 63u;

    // this thread's row inside a box, and the un-swizzle of its eight 16-byte chunks (chunk c sits at
    // c ^ (row & 7)): xa = first 64 bytes of the row (even blocks), xb = second 64 bytes (odd blocks)
    const uint32_t swz = (lane & 7u) << 4;
    uint32_t xa[4], xb[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        xa[q] = wbase + lane * 128u + (((uint32_t)q << 4) ^ swz);
        xb[q] = wbase + lane * 128u + (((uint32_t)(q + 4) << 4) ^ swz);
    }
    auto box_of = [&](uint32_t b) { return ((b & (RB - 1u)) >> 1) << 12; };       // byte offset of block b's box in the ring
    auto lds16 = [&](uint32_t (&m)[16], const uint32_t (&x)[4], uint32_t box) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(m[4 * q]), "=r"(m[4 * q + 1]), "=r"(m[4 * q + 2]), "=r"(m[4 * q + 3])
                         : "r"(x[q] + box)
                         : "memory");
    };
    auto full_bar = [&](uint32_t S) { return full0 + 8u * (S & (NSTAGE - 1u)); };
    auto wait_stage = [&](uint32_t S) { mbar_wait(full_bar(S), (S / NSTAGE) & 1u); };
    // A unit longer than one launch may carry (a tensor-map row is at most 2^31 bytes) is hashed as a
    // run of launches: all but the first start from the chaining value the previous one left in the
    // table (LANE_CONT, `prior` bytes already hashed), all but the last store the chaining value
    // instead of finalising (no LANE_FINAL; their unit_len is a multiple of 64).
    uint32_t st[4];
    if ((flags & LANE_CONT) && u < n) {
        const uint4 v = table[u];
        st[0] = v.x; st[1] = v.y; st[2] = v.z; st[3] = v.w;
    } else {
        md5_iv(st);
    }
    uint32_t mA[16], mB[16];
    uint32_t b = 0;
    if (bm >= 1u) {
        wait_stage(0);
        lds16(mA, xa, 0u);
    }
    // Main loop: U whole blocks per iteration (b .. b+U-1; b already in mA), while a further whole block
    // b+U exists to be prefetched.  Two register sets alternate, each block's words fetched under the
    // 64 steps of the block before it.  Nothing in the body is conditional except the (never taken)
    // spin when the next stage has not landed:
    //  - the "full" barrier of block b+U's stage is polled EVERY iteration -- within a stage it has long
    //    completed and the poll is free; at a stage boundary its latency runs under the iteration's steps;
    //  - lane 0 arrives on the "empty" barrier of the current stage EVERY iteration; the barrier counts
    //    BPS/U arrivals, so it completes exactly when the stage has been read.
    constexpr uint32_t LOG_BPS = TILES == 1 ? 1 : TILES == 2 ? 2 : TILES == 4 ? 3 : 4;
    constexpr uint32_t LOG_NST = NSTAGE == 1 ? 0 : NSTAGE == 2 ? 1 : NSTAGE == 4 ? 2 : 3;
    constexpr uint32_t RING_MASK = (RB / 2u) * 4096u - 1u;     // boxes of the ring, in bytes
    uint32_t box = 0;                                          // byte offset of block b's box in the ring
    if (bm > (uint32_t)U) {
        bool again;
        do {
            const uint32_t s2 = (b + U) >> LOG_BPS;            // stage of block b+U
            const uint32_t landed = mbar_try_wait(full0 + ((s2 & (NSTAGE - 1u)) << 3), (s2 >> LOG_NST) & 1u);
            again = b + 2u * U < bm;
            const uint32_t box_next = (box + (U / 2u) * 4096u) & RING_MASK;
#pragma unroll
            for (int j = 0; j < U; j += 2) {
                lds16(mB, xb, box + (uint32_t)(j / 2) * 4096u);             // block b+j+1: same box as b+j
                md5_block<MODE>(st, mA);                                    // block b+j
                if (j + 2 < U) {
                    lds16(mA, xa, box + (uint32_t)(j / 2 + 1) * 4096u);     // block b+j+2: next box of this stage
                } else {
                    if (!landed) wait_stage(s2);
                    lds16(mA, xa, box_next);                                // block b+U
                }
                md5_block<MODE>(st, mB);                                    // block b+j+1
            }
            __syncwarp();                                      // all 32 rows of these boxes are in registers (and consumed)
            if (lane == 0) mbar_arrive(empty0 + (((b >> LOG_BPS) & (NSTAGE - 1u)) << 3));
            b += U;
            box = box_next;
        } while (again);
    }
    // up to U whole blocks are left; b is in mA.  (xb[q] == xa[q] ^ 64: the other half of the row.)
    while (b < bm) {
        md5_block<MODE>(st, mA);
        ++b;
        if (b < bm) {
            wait_stage(b / BPS);                               // returns at once for a stage already waited for
            const uint32_t bx = box_of(b), half = (b & 1u) << 6;
#pragma unroll
            for (int q = 0; q < 4; ++q)
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(mA[4 * q]), "=r"(mA[4 * q + 1]), "=r"(mA[4 * q + 2]), "=r"(mA[4 * q + 3])
                             : "r"((xa[q] ^ half) + bx)
                             : "memory");
        }
    }

    // ---- the block that takes the 0x80 marker, and the length block if the 8 length bytes do not fit ----
    if (flags & LANE_FINAL) {
        const uint64_t total = prior + unit_len;
        uint32_t m[16];
        if (rem) {
            wait_stage(bm / BPS);                              // returns at once if this stage was waited for before
            if (bm & 1u) lds16(m, xb, box_of(bm));
            else lds16(m, xa, box_of(bm));
        } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) m[q] = 0;
        }
        md5_pad_block(m, rem, false, total);
        md5_block<MODE>(st, m);
        if (nb > bm + 1u) {
#pragma unroll
            for (int q = 0; q < 16; ++q) m[q] = 0;
            md5_pad_block(m, rem, true, total);
            md5_block<MODE>(st, m);
        }
    }
    if (u < n) table[u] = make_uint4(st[0], st[1], st[2], st[3]);
}

}  // namespace tpimd5
