// md5_device.cuh -- the MD5 compression function as sm_100a integer code.
//
// One thread owns one MD5 chain (one "lane": a file, an S3 part or a stream).  The arithmetic is
// RFC 1321 section 3.4, which is what Go's crypto/md5 block function computes for
// [rclone] fs/hash StreamTypes (reached from /root/reference/task/common/machine/storage.go:158).
// Message words are little-endian, so bytes staged in shared memory are used as they are.
//
// Instruction budget per step: one LOP3 (F/G/H/I are single 3-input LUT ops), the rotate fused with
// the final addition into one LEA.HI, and the sum a + f + m + k (see the step modes below).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace tpimd5 {

__device__ __forceinline__ uint32_t rotl(uint32_t x, int s) { return __funnelshift_l(x, x, s); }

// lop3 with an explicit truth table keeps ptxas from splitting the boolean into two ops.
template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return r;
}
// truth tables for inputs (b, c, d) = (0xF0, 0xCC, 0xAA)
__device__ __forceinline__ uint32_t fF(uint32_t b, uint32_t c, uint32_t d) { return lop3<0xCA>(b, c, d); }  // (b&c)|(~b&d)
__device__ __forceinline__ uint32_t fG(uint32_t b, uint32_t c, uint32_t d) { return lop3<0xE4>(b, c, d); }  // (b&d)|(c&~d)
__device__ __forceinline__ uint32_t fH(uint32_t b, uint32_t c, uint32_t d) { return lop3<0x96>(b, c, d); }  // b^c^d
__device__ __forceinline__ uint32_t fI(uint32_t b, uint32_t c, uint32_t d) { return lop3<0x39>(b, c, d); }  // c^(b|~d)

// How the three additions of a step are issued.  sm_100a has two integer-capable pipes per SM
// sub-partition: alu (LOP3, IADD3, LEA, SHF; one warp instruction every other cycle) and fma (IMAD,
// and the reg+imm VIADD).  LOP3 and the fused rotate-add (LEA.HI) must go to alu; the additions can go
// either way, but ptxas, left alone, folds them into one IADD3 on alu and puts the constant add on the
// dependent chain.  The step modes below were measured with the register-only probe and in the kernels
// (profiles/r1_int_probe_modes0-6.jsonl):
//   MODE 0  plain C, ptxas decides: LOP3 -> IADD3 -> VIADD -> LEA.HI, a four-deep chain
//   MODE 4  m + k as a plain add hidden from the optimiser (VIADD, off the chain), then a + x + f as one
//           IADD3: the three-deep chain LOP3 -> IADD3 -> LEA.HI, all on alu.  Shortest latency per step:
//           used when chains are scarce.
//   MODE 5  m + k as in 4; a + x and y + f as two IMADs by a 1 held in the constant bank / a uniform
//           register (two register reads each) on the fma pipe: alu 2, fma 3 per step
//   MODE 6  steps alternate between 4 and 5: alu 2.5, fma 2 on average.  Highest throughput: default
//           when the SMs are full.
__constant__ uint32_t c_one = 1u;

__device__ __forceinline__ uint32_t add_imm(uint32_t a, uint32_t k) {
    uint32_t r;
    asm("add.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(k));
    return r;
}
__device__ __forceinline__ uint32_t madc(uint32_t a, uint32_t c) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(c_one), "r"(c));
    return r;
}

// does step i (0..63) of a mixed mode take the fma-pipe form?
template <int MODE>
__host__ __device__ constexpr int fma_step(int i) {
    return MODE == 5 ? 1 : MODE == 6 ? (i & 1) : 0;
}

// m + k as an IMAD by the constant-bank 1 with k as the immediate addend: the fma pipe, whatever ptxas
// would choose for a plain add
__device__ __forceinline__ uint32_t madk(uint32_t m, uint32_t k) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(m), "r"(c_one), "r"(k));
    return r;
}

template <int MODE, int FMA_FORM>
__device__ __forceinline__ uint32_t step_sum(uint32_t a, uint32_t f, uint32_t m, uint32_t k) {
    if (MODE == 0) {
        return (a + (m + k)) + f;
    } else if (MODE == 7) {
        // experimental: a + m + k as one three-input add off the chain (alu), then a two-input add on it
        uint32_t x;
        asm("{\n\t.reg .u32 t;\n\tadd.u32 t, %1, %2;\n\tadd.u32 %0, t, %3;\n\t}" : "=r"(x) : "r"(a), "r"(m), "r"(k));
        return x + f;
    } else if (MODE == 8) {
        // experimental: m + k on the fma pipe for certain (IMAD), the rest as mode 4
        const uint32_t x = madk(m, k);
        return a + x + f;
    } else if (!FMA_FORM) {
        const uint32_t x = add_imm(m, k);
        return a + x + f;
    } else {
        const uint32_t x = add_imm(m, k);
        const uint32_t y = madc(a, x);
        return madc(f, y);
    }
}

#define TPI_STEP(f, a, b, c, d, m, k, s, idx)                                    \
    {                                                                            \
        const uint32_t t_ = step_sum<MODE, fma_step<MODE>(idx)>((a), f((b), (c), (d)), (m), (uint32_t)(k)); \
        (a) = (b) + rotl(t_, (s));                                               \
    }

// One 64-byte block: st <- compress(st, m[0..15]).
template <int MODE>
__device__ __forceinline__ void md5_block(uint32_t (&st)[4], const uint32_t (&m)[16]) {
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
    // round 1: g = i
    TPI_STEP(fF, a, b, c, d, m[0], 0xd76aa478u, 7, 0)
    TPI_STEP(fF, d, a, b, c, m[1], 0xe8c7b756u, 12, 1)
    TPI_STEP(fF, c, d, a, b, m[2], 0x242070dbu, 17, 2)
    TPI_STEP(fF, b, c, d, a, m[3], 0xc1bdceeeu, 22, 3)
    TPI_STEP(fF, a, b, c, d, m[4], 0xf57c0fafu, 7, 4)
    TPI_STEP(fF, d, a, b, c, m[5], 0x4787c62au, 12, 5)
    TPI_STEP(fF, c, d, a, b, m[6], 0xa8304613u, 17, 6)
    TPI_STEP(fF, b, c, d, a, m[7], 0xfd469501u, 22, 7)
    TPI_STEP(fF, a, b, c, d, m[8], 0x698098d8u, 7, 8)
    TPI_STEP(fF, d, a, b, c, m[9], 0x8b44f7afu, 12, 9)
    TPI_STEP(fF, c, d, a, b, m[10], 0xffff5bb1u, 17, 10)
    TPI_STEP(fF, b, c, d, a, m[11], 0x895cd7beu, 22, 11)
    TPI_STEP(fF, a, b, c, d, m[12], 0x6b901122u, 7, 12)
    TPI_STEP(fF, d, a, b, c, m[13], 0xfd987193u, 12, 13)
    TPI_STEP(fF, c, d, a, b, m[14], 0xa679438eu, 17, 14)
    TPI_STEP(fF, b, c, d, a, m[15], 0x49b40821u, 22, 15)
    // round 2: g = (5i + 1) mod 16
    TPI_STEP(fG, a, b, c, d, m[1], 0xf61e2562u, 5, 16)
    TPI_STEP(fG, d, a, b, c, m[6], 0xc040b340u, 9, 17)
    TPI_STEP(fG, c, d, a, b, m[11], 0x265e5a51u, 14, 18)
    TPI_STEP(fG, b, c, d, a, m[0], 0xe9b6c7aau, 20, 19)
    TPI_STEP(fG, a, b, c, d, m[5], 0xd62f105du, 5, 20)
    TPI_STEP(fG, d, a, b, c, m[10], 0x02441453u, 9, 21)
    TPI_STEP(fG, c, d, a, b, m[15], 0xd8a1e681u, 14, 22)
    TPI_STEP(fG, b, c, d, a, m[4], 0xe7d3fbc8u, 20, 23)
    TPI_STEP(fG, a, b, c, d, m[9], 0x21e1cde6u, 5, 24)
    TPI_STEP(fG, d, a, b, c, m[14], 0xc33707d6u, 9, 25)
    TPI_STEP(fG, c, d, a, b, m[3], 0xf4d50d87u, 14, 26)
    TPI_STEP(fG, b, c, d, a, m[8], 0x455a14edu, 20, 27)
    TPI_STEP(fG, a, b, c, d, m[13], 0xa9e3e905u, 5, 28)
    TPI_STEP(fG, d, a, b, c, m[2], 0xfcefa3f8u, 9, 29)
    TPI_STEP(fG, c, d, a, b, m[7], 0x676f02d9u, 14, 30)
    TPI_STEP(fG, b, c, d, a, m[12], 0x8d2a4c8au, 20, 31)
    // round 3: g = (3i + 5) mod 16
    TPI_STEP(fH, a, b, c, d, m[5], 0xfffa3942u, 4, 32)
    TPI_STEP(fH, d, a, b, c, m[8], 0x8771f681u, 11, 33)
    TPI_STEP(fH, c, d, a, b, m[11], 0x6d9d6122u, 16, 34)
    TPI_STEP(fH, b, c, d, a, m[14], 0xfde5380cu, 23, 35)
    TPI_STEP(fH, a, b, c, d, m[1], 0xa4beea44u, 4, 36)
    TPI_STEP(fH, d, a, b, c, m[4], 0x4bdecfa9u, 11, 37)
    TPI_STEP(fH, c, d, a, b, m[7], 0xf6bb4b60u, 16, 38)
    TPI_STEP(fH, b, c, d, a, m[10], 0xbebfbc70u, 23, 39)
    TPI_STEP(fH, a, b, c, d, m[13], 0x289b7ec6u, 4, 40)
    TPI_STEP(fH, d, a, b, c, m[0], 0xeaa127fau, 11, 41)
    TPI_STEP(fH, c, d, a, b, m[3], 0xd4ef3085u, 16, 42)
    TPI_STEP(fH, b, c, d, a, m[6], 0x04881d05u, 23, 43)
    TPI_STEP(fH, a, b, c, d, m[9], 0xd9d4d039u, 4, 44)
    TPI_STEP(fH, d, a, b, c, m[12], 0xe6db99e5u, 11, 45)
    TPI_STEP(fH, c, d, a, b, m[15], 0x1fa27cf8u, 16, 46)
    TPI_STEP(fH, b, c, d, a, m[2], 0xc4ac5665u, 23, 47)
    // round 4: g = 7i mod 16
    TPI_STEP(fI, a, b, c, d, m[0], 0xf4292244u, 6, 48)
    TPI_STEP(fI, d, a, b, c, m[7], 0x432aff97u, 10, 49)
    TPI_STEP(fI, c, d, a, b, m[14], 0xab9423a7u, 15, 50)
    TPI_STEP(fI, b, c, d, a, m[5], 0xfc93a039u, 21, 51)
    TPI_STEP(fI, a, b, c, d, m[12], 0x655b59c3u, 6, 52)
    TPI_STEP(fI, d, a, b, c, m[3], 0x8f0ccc92u, 10, 53)
    TPI_STEP(fI, c, d, a, b, m[10], 0xffeff47du, 15, 54)
    TPI_STEP(fI, b, c, d, a, m[1], 0x85845dd1u, 21, 55)
    TPI_STEP(fI, a, b, c, d, m[8], 0x6fa87e4fu, 6, 56)
    TPI_STEP(fI, d, a, b, c, m[15], 0xfe2ce6e0u, 10, 57)
    TPI_STEP(fI, c, d, a, b, m[6], 0xa3014314u, 15, 58)
    TPI_STEP(fI, b, c, d, a, m[13], 0x4e0811a1u, 21, 59)
    TPI_STEP(fI, a, b, c, d, m[4], 0xf7537e82u, 6, 60)
    TPI_STEP(fI, d, a, b, c, m[11], 0xbd3af235u, 10, 61)
    TPI_STEP(fI, c, d, a, b, m[2], 0x2ad7d2bbu, 15, 62)
    TPI_STEP(fI, b, c, d, a, m[9], 0xeb86d391u, 21, 63)
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}

// Turns the words of a block that holds only `rem` (0..63) message bytes into the padded block:
// 0x80 after the data, zeros, and -- when the 8 length bytes still fit (rem <= 55) -- the
// message length in bits as a 64-bit little-endian count in words 14/15.  A block that comes
// after the 0x80 block (`after_marker`) is all zeros plus the length.
__device__ __forceinline__ void md5_pad_block(uint32_t (&m)[16], uint32_t rem, bool after_marker, uint64_t total_bytes) {
#pragma unroll
    for (int w = 0; w < 16; ++w) {
        const uint32_t lo = 4u * w;
        uint32_t v = m[w];
        if (after_marker) {
            v = 0;
        } else if (rem <= lo) {
            v = (rem == lo) ? 0x80u : 0u;
        } else if (rem < lo + 4) {
            const uint32_t k = 8u * (rem - lo);                 // 8, 16 or 24 valid bits
            v = (v & ((1u << k) - 1u)) | (0x80u << k);
        }
        m[w] = v;
    }
    if (after_marker || rem <= 55) {
        const uint64_t bits = total_bytes << 3;
        m[14] = (uint32_t)bits;
        m[15] = (uint32_t)(bits >> 32);
    }
}

__device__ __forceinline__ void md5_iv(uint32_t (&st)[4]) {
    st[0] = 0x67452301u; st[1] = 0xefcdab89u; st[2] = 0x98badcfeu; st[3] = 0x10325476u;
}

}  // namespace tpimd5
