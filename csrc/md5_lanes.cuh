// md5_lanes.cuh -- K1/K2/K3: one MD5 chain per thread, input staged through shared memory.
//
// A "unit" is whatever the caller wants one digest for: a whole small file (K1, replaces
// [rclone] backend/local Object.Hash per file), one S3 part (K2, replaces md5.Sum(buf) in
// [rclone] backend/s3 uploadMultipart) or the next slice of a long stream (K3, Go
// digest.Write).  Units never interact, so a warp simply carries 32 of them.
//
// Data movement.  A unit's bytes are contiguous in HBM but 32 units are 32 different places, so
// letting each thread read its own unit would touch 32 cache lines per load instruction.
// Instead the warp copies, for every one of its 32 units, the next CH bytes into a shared-memory
// row with 16-byte cp.async (LDGSTS) -- consecutive threads cover consecutive 16-byte granules
// of one row, i.e. whole 128-byte lines -- double buffered so the copy of stage s+1 runs under
// the 64-step arithmetic of stage s.  Thread t then reads row t.  Rows are CH+16 bytes apart, so
// the four LDS.128 of a block hit all 32 banks once per quarter warp.
//
// Ragged input.  Units may have any length.  The block loop runs to the warp's longest unit;
// finished threads sit out.  The padding block(s) are synthesised in registers.  Callers sort
// units by length so that a warp's 32 units end together (see tpimd5.cu).
#pragma once
#include "md5_device.cuh"

namespace tpimd5 {

constexpr uint32_t LANE_FINAL = 1u;  // pad and finalise: table[out] receives the digest
constexpr uint32_t LANE_CONT = 2u;   // start from the chaining value in table[out] instead of the IV
constexpr uint32_t LANE_BOX3D = 4u;  // (strided TMA kernels) the main tensor map is the 3-D view: one request per stage

struct LaneArgs {
    const uint8_t *base;       // device buffer holding every unit of this launch
    uint64_t base_bytes;       // its extent
    const uint64_t *off;       // unit -> byte offset in base, or nullptr: unit i starts at i*stride
    const uint64_t *len;       // unit -> bytes to hash in this launch (ignored when off == nullptr)
    uint64_t stride;
    const uint64_t *prior;     // unit -> bytes already hashed by earlier launches (nullptr: 0)
    const uint32_t *flags;     // unit -> LANE_* (nullptr: LANE_FINAL)
    const uint32_t *out_idx;   // unit -> row of `table` (nullptr: identity)
    uint4 *table;              // 16-byte chaining values / digests
    uint64_t n;                // units in this launch
};

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void *gptr) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ uint32_t warp_max(uint32_t v) { return __reduce_max_sync(0xffffffffu, v); }

// CH      bytes of one unit staged per pipeline stage (64, 128, 256 or 512)
// ALIGNED every unit starts on a 16-byte boundary of device memory (fast path: LDS.128, no shifts)
// WARPS   warps per CTA; each warp is independent (no CTA-wide barrier anywhere)
template <int CH, bool ALIGNED, int WARPS>
struct LaneCfg {
    static constexpr int SEGS = CH / 16;              // 16-byte granules per row per stage
    static constexpr int RPI = 32 / SEGS;             // rows covered by one warp-wide copy instruction
    static constexpr int BPS = CH / 64;               // MD5 blocks per stage
    static constexpr int ROWB = CH + 16;              // row pitch: +16 B skews banks (aligned) / holds the shifted tail
    static constexpr int STAGE_BYTES = 32 * ROWB;
    static constexpr int WARP_BYTES = 2 * STAGE_BYTES;
    static constexpr int SMEM_BYTES = WARPS * WARP_BYTES;
    static_assert(CH == 64 || CH == 128 || CH == 256 || CH == 512, "CH");
};

// The 16 message words of one block from this thread's shared-memory row.  Aligned units: four
// LDS.128.  Unaligned units: 17 words from the 4-byte boundary below, realigned by funnel shifts.
template <bool ALIGNED>
__device__ __forceinline__ void load_block_words(uint32_t src, uint32_t sh, uint32_t (&m)[16]) {
    if (ALIGNED) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(m[4 * q]), "=r"(m[4 * q + 1]), "=r"(m[4 * q + 2]), "=r"(m[4 * q + 3])
                         : "r"(src + 16u * q));
    } else {
        uint32_t w[17];
        const uint32_t a0 = src + (sh & ~3u);
#pragma unroll
        for (int q = 0; q < 17; ++q) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[q]) : "r"(a0 + 4u * q));
        const uint32_t bs = (sh & 3u) * 8u;
#pragma unroll
        for (int q = 0; q < 16; ++q) m[q] = __funnelshift_r(w[q], w[q + 1], bs);
    }
}

template <int CH, bool ALIGNED, int WARPS, int MINB, int MODE>
__global__ void __launch_bounds__(WARPS * 32, MINB) md5_lanes_kernel(const LaneArgs p) {
    using Cfg = LaneCfg<CH, ALIGNED, WARPS>;
    constexpr int SEGS = Cfg::SEGS, RPI = Cfg::RPI, BPS = Cfg::BPS, ROWB = Cfg::ROWB;
    extern __shared__ __align__(128) uint8_t smem_raw[];

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: stays in uniform registers
    const uint64_t u = (uint64_t)blockIdx.x * (WARPS * 32) + threadIdx.x;
    const bool valid = u < p.n;

    // ---- this thread's unit ------------------------------------------------------------------
    uint64_t off = 0, len = 0, prior = 0;
    uint32_t fl = LANE_FINAL, oidx = (uint32_t)u;
    if (valid) {
        if (p.off) {
            off = p.off[u];
            len = p.len[u];
        } else {
            off = u * p.stride;
            const uint64_t left = p.base_bytes > off ? p.base_bytes - off : 0;
            len = left < p.stride ? left : p.stride;
        }
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
    const uint64_t addr = (uint64_t)p.base + off;
    const uint32_t sh = ALIGNED ? 0u : (uint32_t)(addr & 15u);
    const uint64_t addr16 = addr & ~(uint64_t)15;
    // 16-byte granules that overlap this unit's bytes: only those are ever fetched, so nothing
    // outside the 16-byte-aligned hull of the caller's data is read.
    const uint32_t ngran = valid ? (uint32_t)((sh + len + 15) >> 4) : 0u;
    const uint32_t nst = (uint32_t)((len + CH - 1) / CH);                      // data stages
    const uint32_t bm = (uint32_t)(len >> 6);                                  // block that gets the 0x80 marker
    const uint32_t nb = !valid ? 0u : (fin ? (uint32_t)((len + 8) >> 6) + 1u : bm);

    const uint32_t warp_nb = warp_max(nb);
    const uint32_t warp_nst = warp_max(nst);

    // ---- copy plan: thread t moves granule (t % SEGS) of rows (t / SEGS) + k*RPI -----------------
    const uint32_t seg = lane % SEGS, rsub = lane / SEGS;
    const uint8_t *gsrc[SEGS];
    int32_t gleft[SEGS];                                  // granules of that row from `seg` on; > 0 means fetch
#pragma unroll
    for (int k = 0; k < SEGS; ++k) {
        const int row = k * RPI + rsub;
        const uint64_t ra = __shfl_sync(0xffffffffu, addr16, row);
        const uint32_t rg = __shfl_sync(0xffffffffu, ngran, row);
        gsrc[k] = (const uint8_t *)ra + seg * 16u;
        gleft[k] = (int32_t)rg - (int32_t)seg;
    }
    const uint32_t wbase = (uint32_t)__cvta_generic_to_shared(smem_raw) + warp * Cfg::WARP_BYTES;
    const uint32_t dst0 = wbase + rsub * ROWB + seg * 16u;       // (+ k*RPI*ROWB + stage*STAGE_BYTES)
    // unaligned units need one granule more per row and stage (the row's own thread fetches it)
    const uint8_t *xsrc = (const uint8_t *)addr16 + SEGS * 16u;
    int32_t xleft = (int32_t)ngran - SEGS;
    const uint32_t xdst0 = wbase + lane * ROWB + SEGS * 16u;

    auto issue = [&](uint32_t s) {
        const uint32_t sb = (s & 1u) * Cfg::STAGE_BYTES;
#pragma unroll
        for (int k = 0; k < SEGS; ++k) {
            if (gleft[k] > 0) cp_async16(dst0 + sb + k * (RPI * ROWB), gsrc[k]);
            gsrc[k] += CH;
            gleft[k] -= SEGS;
        }
        if (!ALIGNED) {
            if (xleft > 0) cp_async16(xdst0 + sb, xsrc);
            xsrc += CH;
            xleft -= SEGS;
        }
    };

    // the same copy without the per-row predicates, for stages in which every row is full
    auto issue_full = [&](uint32_t s) {
        const uint32_t sb = (s & 1u) * Cfg::STAGE_BYTES;
#pragma unroll
        for (int k = 0; k < SEGS; ++k) {
            cp_async16(dst0 + sb + k * (RPI * ROWB), gsrc[k]);
            gsrc[k] += CH;
            gleft[k] -= SEGS;
        }
        if (!ALIGNED) {
            cp_async16(xdst0 + sb, xsrc);
            xsrc += CH;
            xleft -= SEGS;
        }
    };

    if (warp_nst > 0) issue(0);
    cp_async_commit();

    const uint32_t myrow = wbase + lane * ROWB;
    const uint64_t total = prior + len;

    // ---- stage loop ---------------------------------------------------------------------------------
    // Per stage the warp votes: if every unit either has CH whole bytes left or is completely done
    // (padding included), the stage runs without any tail logic -- copy, wait, 4 x LDS.128 and the 64
    // steps per block, finished threads sitting out.  That is the whole run of a uniform batch and, in a
    // length-sorted batch, everything except the stage in which a unit ends.  Only such a stage takes
    // the per-block path that knows about partial blocks and padding.
    // (An unaligned unit needs CH+16 bytes for its shifted window, hence the margins below.)
    const uint64_t full64 = (ALIGNED ? len : (len > sh ? len - sh : 0)) / CH;
    const uint32_t full_st = full64 > 0xffffffffull ? 0xffffffffu : (uint32_t)full64;
    const uint32_t my_whole = ALIGNED ? full_st : (full_st ? full_st - 1 : 0);   // stages this unit runs without tail logic
    const uint32_t warp_fast = __reduce_min_sync(0xffffffffu, my_whole);          // ... and every unit of the warp does
    const uint32_t stages_total = (warp_nb + BPS - 1) / BPS;
#pragma unroll 1
    for (uint32_t s = 0; s < stages_total; ++s) {
        if (s < warp_nst) {
            if (s + 1 < warp_fast) issue_full(s + 1);
            else if (s + 1 < warp_nst) issue(s + 1);
            cp_async_commit();
            cp_async_wait<1>();      // everything but the newest group: stage s has landed
            __syncwarp();
        }
        const bool whole = s < my_whole;
        const bool over = (uint64_t)s * BPS >= nb;
        const uint32_t src0 = myrow + (s & 1u) * Cfg::STAGE_BYTES;
        if (__all_sync(0xffffffffu, whole || over)) {
            if (whole) {
#pragma unroll 1
                for (uint32_t j = 0; j < (uint32_t)BPS; ++j) {
                    uint32_t m[16];
                    load_block_words<ALIGNED>(src0 + j * 64u, sh, m);
                    md5_block<MODE>(st, m);
                }
            }
        } else {
#pragma unroll 1
            for (uint32_t j = 0; j < (uint32_t)BPS; ++j) {
                const uint32_t b = s * BPS + j;
                if (b < nb) {
                    uint32_t m[16];
                    if (b <= bm && b < nst * BPS) {
                        load_block_words<ALIGNED>(src0 + j * 64u, sh, m);
                    } else {
#pragma unroll
                        for (int q = 0; q < 16; ++q) m[q] = 0;
                    }
                    if (fin && b >= bm) md5_pad_block(m, (uint32_t)(len & 63u), b > bm, total);
                    md5_block<MODE>(st, m);
                }
            }
        }
        __syncwarp();                // every thread is done with this stage before it is refilled
    }
    cp_async_wait<0>();
    if (valid) p.table[oidx] = make_uint4(st[0], st[1], st[2], st[3]);
}

// ---------------------------------------------------------------------------------------------
// mbarrier / async-proxy helpers shared with the tensor-map kernel (md5_strided_tma.cuh).
// (A variant of the kernel above in which every thread fetched its own row with one
// cp.async.bulk per stage was measured and dropped: one TMA request per 64-256 bytes tops out at
// 2.3-3.8 TB/s on C2, see profiles/r1_sweep_bulk_per_row.jsonl.)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(mbar), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(mbar)
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// Register-only probe: the same 64-step body with no memory traffic at all.  Its throughput is the
// integer-issue ceiling the lanes kernel is measured against ("calibrate with
// a register-only microbenchmark of the same instruction mix").
// ---------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(128) int_probe_kernel(uint32_t iters, uint4 *__restrict__ out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t st[4], m[16];
    md5_iv(st);
    st[0] ^= t;
#pragma unroll
    for (int i = 0; i < 16; ++i) m[i] = t * 0x9E3779B9u + i;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; ++it) {
        md5_block<MODE>(st, m);
        m[0] ^= st[0];
    }
    out[t] = make_uint4(st[0], st[1], st[2], st[3]);
}

}  // namespace tpimd5
