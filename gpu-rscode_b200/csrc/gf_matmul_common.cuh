// gf_matmul_common.cuh -- device helpers shared by the split-nibble kernels (register-fed and TMA-staged):
// streaming 128-bit global accesses, shared-memory lookups, the per-word multiply-accumulate on
// split-nibble pages, and the accumulator transpose + store.
#pragma once
#include <stdint.h>

namespace rsgpu {
namespace {

__device__ __forceinline__ uint4 ldg_stream(const uint8_t *p)
{
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

__device__ __forceinline__ void stg_stream(uint8_t *p, uint4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ uint2 lds64(uint32_t addr)
{
    uint2 v;
    asm("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// Accumulators for one 16-byte column: W=4 -> 16 words (byte r of word b = row r at byte position b),
// W=8 -> 16 x 2 words.
template <int W> struct Acc;
template <> struct Acc<4> { uint32_t a[16]; };
template <> struct Acc<8> { uint32_t a[16]; uint32_t b[16]; };

template <int W> __device__ __forceinline__ void acc_clear(Acc<W> &acc);
template <> __device__ __forceinline__ void acc_clear<4>(Acc<4> &acc)
{
#pragma unroll
    for (int i = 0; i < 16; ++i) acc.a[i] = 0;
}
template <> __device__ __forceinline__ void acc_clear<8>(Acc<8> &acc)
{
#pragma unroll
    for (int i = 0; i < 16; ++i) { acc.a[i] = 0; acc.b[i] = 0; }
}

// Shared-memory layout: fragment j owns the 256-byte page at tb + 256*j (tb is 256-byte aligned);
// its low-nibble table (16 entries of W bytes) sits at +0 and its high-nibble table at +16*W.
//
// The 8 lookup addresses of one 32-bit data word.  Two masked copies of the word hold the scaled
// nibbles (nibble * W) of all four bytes; what differs between the variants is how one byte of those
// is turned into an address:
// EX = 0: one PRMT per lookup splices the byte into the page-base register (page bases have
//          a zero low byte), so the address costs a single alu-pipe instruction and the LDS needs no
//          immediate.  The +16*W of the high table is OR-ed into the mask step.
// EX = 2 (default): one DP4A per lookup (dot product with a one-hot byte selector, accumulating onto the base).
// (Two more were measured and dropped: shift+mask per nibble, 1.48 ms vs 1.18 ms on RS(10,14); and a
// PRMT/DP4A mix, 1.23 ms.)
template <int W, int EX>
__device__ __forceinline__ void mac_word(uint32_t w, uint32_t tj, uint32_t hi4, Acc<W> &acc, int pos0)
{
    constexpr int SH = (W == 4) ? 2 : 3;            // log2(W)
    constexpr uint32_t M1 = 0xFu << SH;             // one scaled nibble
    constexpr uint32_t M4 = M1 * 0x01010101u;       // four of them
    constexpr uint32_t HI1 = 16u * W;               // offset of the high-nibble table inside the page
    uint32_t olo[4], ohi[4];
    const uint32_t wl = (w << SH) & M4;
    if constexpr (EX == 0) {
        // hi4 == HI4, but it arrives as a kernel argument: a value the compiler cannot fold keeps
        // (x & M4) | hi4 as ONE three-input LOP3 instead of an AND-immediate plus an OR-immediate
        const uint32_t wh = ((w >> (4 - SH)) & M4) | hi4;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            olo[b] = __byte_perm(wl, tj, 0x7650u + b);
            ohi[b] = __byte_perm(wh, tj, 0x7650u + b);
        }
    } else {
        const uint32_t wh = (w >> (4 - SH)) & M4;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            olo[b] = __dp4a(wl, 1u << (8 * b), tj);
            ohi[b] = __dp4a(wh, 1u << (8 * b), tj + HI1);
        }
    }
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        if constexpr (W == 4) {
            acc.a[pos0 + b] ^= lds32(olo[b]) ^ lds32(ohi[b]);
        } else {
            const uint2 l = lds64(olo[b]), h = lds64(ohi[b]);
            acc.a[pos0 + b] ^= l.x ^ h.x;
            acc.b[pos0 + b] ^= l.y ^ h.y;
        }
    }
}

template <int W, int EX>
__device__ __forceinline__ void mac_vec(const uint4 &d, uint32_t tj, uint32_t hi4, Acc<W> &acc)
{
    mac_word<W, EX>(d.x, tj, hi4, acc, 0);
    mac_word<W, EX>(d.y, tj, hi4, acc, 4);
    mac_word<W, EX>(d.z, tj, hi4, acc, 8);
    mac_word<W, EX>(d.w, tj, hi4, acc, 12);
}

// 16 accumulator words (byte r = row r) -> 4 row vectors of 16 bytes.
__device__ __forceinline__ void transpose_store(const uint32_t *a, uint8_t *out, size_t out_pitch, size_t off,
                                                int row0, int nrows)
{
    uint32_t r[4][4];   // r[row][q]
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint32_t a0 = a[4 * q], a1 = a[4 * q + 1], a2 = a[4 * q + 2], a3 = a[4 * q + 3];
        const uint32_t t0 = __byte_perm(a0, a1, 0x5140), t1 = __byte_perm(a2, a3, 0x5140);
        const uint32_t t2 = __byte_perm(a0, a1, 0x7362), t3 = __byte_perm(a2, a3, 0x7362);
        r[0][q] = __byte_perm(t0, t1, 0x5410);
        r[1][q] = __byte_perm(t0, t1, 0x7632);
        r[2][q] = __byte_perm(t2, t3, 0x5410);
        r[3][q] = __byte_perm(t2, t3, 0x7632);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i < nrows)
            stg_stream(out + (size_t)(row0 + i) * out_pitch + off, make_uint4(r[i][0], r[i][1], r[i][2], r[i][3]));
}

// Load one batch of JB fragment vectors through a running pointer (q += pitch per fragment: two adds
// instead of a 64-bit multiply-add chain per load).  PAD: the last batch may run past fragment k-1;
// padded slots re-read the last valid row (their tables are all zero).
template <int JB, bool PAD>
__device__ __forceinline__ void load_batch(uint4 (&d)[JB], const uint8_t *&q, size_t in_pitch, int &j, int k)
{
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) {
        d[jj] = ldg_stream(q);
        if constexpr (PAD) {
            ++j;
            if (j < k) q += in_pitch;
        } else {
            q += in_pitch;
        }
    }
}

}  // namespace
}  // namespace rsgpu
