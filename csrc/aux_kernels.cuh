// aux_kernels.cuh -- K4 manifest diff and the synthetic-workdir generator.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace tpimd5 {

// ---------------------------------------------------------------------------------------------
// K4: which units changed?  Replaces the per-pair `srcHash != dstHash` of [rclone]
// fs/operations checkHashes (reached from /root/reference/task/common/machine/storage.go:158)
// for two whole manifests of 16-byte digests.  Output: ascending indices of differing rows.
//
// Pass 1 compares (each manifest is read ONCE: 32 B per row, the algorithmic traffic), keeps the outcome
// as one bit per row and counts per 1024-row tile; pass 2 is an exclusive scan of the tile counts (one CTA,
// the array is n/1024 long); pass 3 scatters from the bitmap alone (n/8 bytes).  Within a tile the order
// comes from ballots: warp w owns rows [128w, 128w+128) as four coalesced 512-byte reads.
// HBM-bound; scratch per 1,024-row tile: 8 B of count + 128 B of bitmap (DIFF_SCRATCH_PER_TILE).
// ---------------------------------------------------------------------------------------------
constexpr int DIFF_THREADS = 256;
constexpr int DIFF_K = 4;
constexpr int DIFF_TILE = DIFF_THREADS * DIFF_K;
constexpr size_t DIFF_SCRATCH_PER_TILE = sizeof(uint64_t) + DIFF_TILE / 8;

// layout of the scratch: tile counts first, then the bitmap (one uint32 per 32 rows)
__host__ __device__ inline uint32_t *diff_bitmap(uint64_t *tile_counts, uint64_t tiles) { return (uint32_t *)(tile_counts + tiles); }

template <bool SCATTER>
__global__ void __launch_bounds__(DIFF_THREADS) diff_tile_kernel(const uint4 *__restrict__ a, const uint4 *__restrict__ b,
                                                                  uint64_t n, uint64_t *__restrict__ tile_counts,
                                                                  uint32_t *__restrict__ bitmap, uint32_t *__restrict__ idx_out) {
    __shared__ uint32_t wsum[DIFF_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint64_t row0 = (uint64_t)blockIdx.x * DIFF_TILE + warp * (32 * DIFF_K);
    uint32_t bal[DIFF_K];
    uint32_t cnt = 0;
    if (!SCATTER) {
#pragma unroll
        for (int k = 0; k < DIFF_K; ++k) {
            const uint64_t r = row0 + k * 32 + lane;
            bool ne = false;
            if (r < n) {
                const uint4 x = __ldg(a + r), y = __ldg(b + r);
                ne = ((x.x ^ y.x) | (x.y ^ y.y) | (x.z ^ y.z) | (x.w ^ y.w)) != 0u;
            }
            bal[k] = __ballot_sync(0xffffffffu, ne);
            cnt += __popc(bal[k]);
        }
        if (lane < DIFF_K) {                       // the warp's four ballots, one 16-byte store
            uint32_t v = bal[0];
#pragma unroll
            for (int k = 1; k < DIFF_K; ++k) v = lane == (uint32_t)k ? bal[k] : v;
            bitmap[(row0 >> 5) + lane] = v;
        }
    } else {
        const uint32_t mine = lane < DIFF_K ? bitmap[(row0 >> 5) + lane] : 0u;
#pragma unroll
        for (int k = 0; k < DIFF_K; ++k) {
            bal[k] = __shfl_sync(0xffffffffu, mine, k);
            cnt += __popc(bal[k]);
        }
    }
    if (lane == 0) wsum[warp] = cnt;
    __syncthreads();
    if (!SCATTER) {
        if (threadIdx.x == 0) {
            uint32_t t = 0;
#pragma unroll
            for (int w = 0; w < DIFF_THREADS / 32; ++w) t += wsum[w];
            tile_counts[blockIdx.x] = t;
        }
    } else {
        uint64_t pos = tile_counts[blockIdx.x];          // exclusive prefix after the scan
        for (uint32_t w = 0; w < warp; ++w) pos += wsum[w];
#pragma unroll
        for (int k = 0; k < DIFF_K; ++k) {
            if ((bal[k] >> lane) & 1u) idx_out[pos + __popc(bal[k] & ((1u << lane) - 1u))] = (uint32_t)(row0 + k * 32 + lane);
            pos += __popc(bal[k]);
        }
    }
}

// exclusive scan in place over `m` tile counts, total to *total_out.  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024) diff_scan_kernel(uint64_t *__restrict__ counts, uint64_t m, uint64_t *__restrict__ total_out) {
    __shared__ uint64_t wtot[32];
    __shared__ uint64_t carry_s;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint64_t base = 0; base < m; base += 1024) {
        const uint64_t i = base + threadIdx.x;
        const uint64_t v = i < m ? counts[i] : 0;
        uint64_t x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= (uint32_t)d) x += y;
        }
        if (lane == 31) wtot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint64_t t = wtot[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint64_t y = __shfl_up_sync(0xffffffffu, t, d);
                if (lane >= (uint32_t)d) t += y;
            }
            wtot[lane] = t;          // inclusive over warps
        }
        __syncthreads();
        const uint64_t carry = carry_s;
        const uint64_t before = carry + (warp ? wtot[warp - 1] : 0) + (x - v);
        if (i < m) counts[i] = before;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + wtot[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = carry_s;
}

// ---------------------------------------------------------------------------------------------
// Synthetic workdir content, generated straight into HBM.  8 bytes per counter, counter keyed by
// (seed, file index, byte offset / 8); must stay byte-identical with the host generators used for
// real files and by the test oracle (DESIGN.md "Synthetic content").
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t synth_word(uint64_t seed, uint64_t file_index, uint64_t q) {
    uint64_t z = (seed + 1) * 0x9E3779B97F4A7C15ull;
    z ^= (file_index + 0x632BE59BD9B4E019ull) * 0xD6E8FEB86659FD93ull;
    z += q * 0xA0761D6478BD642Full;
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

// file_len must be a multiple of 16; thread handles 16 bytes per iteration.
__global__ void __launch_bounds__(256) synth_fill_kernel(uint4 *__restrict__ out, uint64_t seed, uint64_t first_file,
                                                          uint64_t file_len, uint64_t n_files) {
    const uint64_t total16 = (file_len >> 4) * n_files;
    const uint64_t per_file16 = file_len >> 4;
    for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total16; g += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t f = g / per_file16;
        const uint64_t q = (g - f * per_file16) * 2;
        const uint64_t w0 = synth_word(seed, first_file + f, q), w1 = synth_word(seed, first_file + f, q + 1);
        out[g] = make_uint4((uint32_t)w0, (uint32_t)(w0 >> 32), (uint32_t)w1, (uint32_t)(w1 >> 32));
    }
}

// ---------------------------------------------------------------------------------------------
// Length bucketing for ragged device-resident units.  A warp runs until its longest unit is done,
// so 32 units of mixed length waste lanes.  Units are therefore ordered by block count (longest
// first: their chains bound the finish time) with a device radix sort; digests still land at the
// caller's index through out_idx.  These two kernels produce the keys and apply the permutation.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bucket_keys_kernel(const uint64_t *__restrict__ len, uint64_t n, uint32_t *__restrict__ keys,
                                                           uint32_t *__restrict__ idx) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t blocks = len[i] >> 6;
    keys[i] = ~(uint32_t)(blocks > 0xffffffffull ? 0xffffffffull : blocks);      // ascending sort = longest first
    idx[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) bucket_gather_kernel(const uint64_t *__restrict__ off, const uint64_t *__restrict__ len,
                                                             const uint32_t *__restrict__ perm, uint64_t n,
                                                             uint64_t *__restrict__ off_out, uint64_t *__restrict__ len_out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t j = perm[i];
    off_out[i] = off[j];
    len_out[i] = len[j];
}

}  // namespace tpimd5
