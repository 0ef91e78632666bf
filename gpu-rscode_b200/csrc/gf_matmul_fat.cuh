// gf_matmul_fat.cuh -- device code shared by the bank-replicated byte-table kernels
// (gf_matmul_fat_sm100.cu: one stripe per launch; gf_matmul_fat_batch_sm100.cu: a batch of stripes per
// launch): shared-memory plan, table construction, multiply-accumulate, transpose + store, loads.
// The design notes are at the top of gf_matmul_fat_sm100.cu.
#pragma once
#include "rs_internal.h"

namespace rsgpu {
namespace {

constexpr int kSmemBudget = 232448 - 256 - 1024 - 64;   // 227 KiB opt-in limit minus alignment slack and the two bit-7 constant areas (32 B per fragment, k <= 32)

__host__ __device__ constexpr int fat_count(int k, int xb = 0)
{
    // every fragment wants at least a fat7 table (16 KiB); thin pages (256 B) for the rest
    int n = k;
    while (n > 0 && n * 16384 + (k - n) * 256 > kSmemBudget - xb) --n;
    return n;
}
__host__ __device__ constexpr int fat8_count(int k, int xb = 0)
{
    const int n = fat_count(k, xb);
    int left = kSmemBudget - xb - n * 16384 - (k - n) * 256;
    int n8 = left / 16384;
    return n8 > n ? n : n8;
}
__host__ __device__ constexpr int fat_threads(int W, int JB, bool T7)
{
    // register budget: accumulators (16 or 32, +16 with transposed bit-7 terms) + two batches of JB vectors
    if (W == 4 && T7) return JB <= 2 ? 768 : (JB <= 4 ? 640 : (JB <= 8 ? 512 : 384));
    if (W == 4) return JB <= 2 ? 1024 : (JB <= 4 ? 768 : (JB <= 5 ? 640 : (JB <= 8 ? 512 : 384)));
    return JB <= 2 ? 768 : (JB <= 5 ? 512 : (JB <= 8 ? 384 : 256));
}

// block size of the instances that also produce the peeled all-ones row (four more accumulator registers): one
// size down where ptxas -v showed spills at the regular size (padded shapes with transposed bit-7 terms or six and
// more batches, and the 3 x 3 shape)
__host__ __device__ constexpr int fat_threads_x(int W, int JB, int KB, bool T7, bool PAD, bool XR)
{
    int t = fat_threads(W, JB, T7);
    if (XR && ((PAD && (T7 || KB >= 6)) || (JB == 3 && KB == 3))) t -= 128;
    return t;
}

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

template <int OFF> __device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
    uint32_t v;
    asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
    return v;
}
template <int OFF> __device__ __forceinline__ uint2 lds64(uint32_t addr)
{
    uint2 v;
    asm("ld.shared.v2.u32 {%0,%1}, [%2+%3];" : "=r"(v.x), "=r"(v.y) : "r"(addr), "n"(OFF));
    return v;
}
// volatile: keeps the per-fragment constant load next to its use instead of hoisting one register per
// fragment out of the column loop
template <int OFF> __device__ __forceinline__ uint32_t lds32_pinned(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
    return v;
}

template <int W> struct Acc;
// t: bit-7 terms of the fat7 fragments, kept TRANSPOSED (t[4*row + word]: four byte positions of one output
// row per register, i.e. already in store order) -- see mac_vec
// x: the peeled all-ones row (XR instances): plain XOR of the input vectors, already in store order
template <> struct Acc<4> { uint32_t a[16]; uint32_t t[16]; uint4 x; };
template <> struct Acc<8> { uint32_t a[16]; uint32_t b[16]; uint4 x; };

// Shared-memory map for a K-fragment code (all offsets relative to the 256-aligned base tb):
//   fragments [0, N8)        fat8 tables, 32 KiB each
//   fragments [N8, NF)       fat7 tables, 16 KiB each
//   fragments [NF, K)        split-nibble pages, 256 B each
//   then K * 16 bytes: the product of 0x80 with every coefficient (bit-7 term of the fat7 fragments)
// XB: bytes of the shared-memory window kept free for other uses (the cluster kernel's exchange buffers)
template <int K, int XB = 0> struct Map {
    static constexpr int NF = fat_count(K, XB);
    static constexpr int N8 = fat8_count(K, XB);
    static constexpr int THIN0 = N8 * 32768 + (NF - N8) * 16384;
    static constexpr int C7 = THIN0 + (K - NF) * 256;
    static constexpr int C7R = C7 + K * 16;    // the same products, one output row per word, byte-replicated
    static constexpr int BYTES = C7R + K * 16;
    static_assert(BYTES + XB + 256 <= 232448, "shared-memory plan exceeds the 227 KiB opt-in limit");
    __host__ __device__ static constexpr int off(int j)
    {
        return j < N8 ? j * 32768 : (j < NF ? N8 * 32768 + (j - N8) * 16384 : THIN0 + (j - NF) * 256);
    }
};

template <int W, int K, int J, bool T7, int XB = 0>
__device__ __forceinline__ void mac_word(uint32_t w, uint32_t lane_base, uint32_t tb, uint32_t c7a, uint32_t c7b,
                                         Acc<W> &acc, int pos0)
{
    using M = Map<K, XB>;
    constexpr int OFF = M::off(J);
    if constexpr (J < M::NF) {
        constexpr bool F7 = (J >= M::N8);
        const uint32_t wi = F7 ? (w & 0x7F7F7F7Fu) : w;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const uint32_t addr = __dp4a(wi, 0x80u << (8 * b), lane_base);
            uint32_t va, vb = 0;
            if constexpr (W == 4) {
                va = lds32<OFF>(addr);
            } else {
                const uint2 v = lds64<OFF>(addr);
                va = v.x;
                vb = v.y;
            }
            if constexpr (F7 && !T7) {
                // PRMT with the sign-replicate flag (selector nibble 8|b): 0xFFFFFFFF if bit 7 of byte b
                // is set, else 0.  Raw PTX: the __byte_perm intrinsic drops bit 3 of the selector nibbles.
                uint32_t m;
                asm("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(w), "r"(0u), "r"(0x8888u + 0x1111u * b));
                va ^= m & c7a;
                if constexpr (W == 8) vb ^= m & c7b;
            }
            acc.a[pos0 + b] ^= va;
            if constexpr (W == 8) acc.b[pos0 + b] ^= vb;
        }
    } else {
        // split-nibble page (same scheme as gf_matmul_sm100.cu, DP4A extraction)
        constexpr int SH = (W == 4) ? 2 : 3;
        constexpr uint32_t M4 = (0xFu << SH) * 0x01010101u;
        const uint32_t wl = (w << SH) & M4;
        const uint32_t wh = (w >> (4 - SH)) & M4;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const uint32_t alo = __dp4a(wl, 1u << (8 * b), tb);
            const uint32_t ahi = __dp4a(wh, 1u << (8 * b), tb);
            if constexpr (W == 4) {
                acc.a[pos0 + b] ^= lds32<OFF>(alo) ^ lds32<OFF + 16 * W>(ahi);
            } else {
                const uint2 l = lds64<OFF>(alo), h = lds64<OFF + 16 * W>(ahi);
                acc.a[pos0 + b] ^= l.x ^ h.x;
                acc.b[pos0 + b] ^= l.y ^ h.y;
            }
        }
    }
}

template <int W, int K, int J, bool T7, int XB = 0>
__device__ __forceinline__ void mac_vec(const uint4 &d, uint32_t lane_base, uint32_t tb, Acc<W> &acc)
{
    using M = Map<K, XB>;
    uint32_t c7a = 0, c7b = 0;
    if constexpr (J >= M::N8 && J < M::NF) {   // fat7: the bit-7 term
        if constexpr (!T7) {
            c7a = lds32_pinned<M::C7 + J * 16>(tb);
            if constexpr (W == 8) c7b = lds32_pinned<M::C7 + J * 16 + 4>(tb);
        } else {
            // Four output rows: instead of (mask, AND, XOR) per BYTE, work per WORD on transposed
            // accumulators.  One PRMT turns the word into four 0x00/0xFF sign bytes; row r then needs
            // t[r][word] ^= signs & (product of 0x80 with coefficient r, replicated in all four bytes):
            // 5 instructions per 4 input bytes instead of 12, and t is already in store order.
            const uint32_t cr0 = lds32_pinned<M::C7R + J * 16>(tb), cr1 = lds32_pinned<M::C7R + J * 16 + 4>(tb);
            const uint32_t cr2 = lds32_pinned<M::C7R + J * 16 + 8>(tb), cr3 = lds32_pinned<M::C7R + J * 16 + 12>(tb);
            const uint32_t ws[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
            for (int wi = 0; wi < 4; ++wi) {
                uint32_t sg;
                asm("prmt.b32 %0, %1, %2, %3;" : "=r"(sg) : "r"(ws[wi]), "r"(0u), "r"(0xBA98u));
                acc.t[wi] ^= sg & cr0;
                acc.t[4 + wi] ^= sg & cr1;
                acc.t[8 + wi] ^= sg & cr2;
                acc.t[12 + wi] ^= sg & cr3;
            }
        }
    }
    mac_word<W, K, J, T7, XB>(d.x, lane_base, tb, c7a, c7b, acc, 0);
    mac_word<W, K, J, T7, XB>(d.y, lane_base, tb, c7a, c7b, acc, 4);
    mac_word<W, K, J, T7, XB>(d.z, lane_base, tb, c7a, c7b, acc, 8);
    mac_word<W, K, J, T7, XB>(d.w, lane_base, tb, c7a, c7b, acc, 12);
}

// XR: the coding matrix has an all-ones row (row 0 of the reference's Vandermonde generator, (j+1)^0,
// src/matrix.cu:758) that was peeled off the tables: that output row is the XOR of the inputs, one LOP3 per word
// instead of a lookup per byte -- which lets five parity rows run on 32-bit table entries (four looked-up rows +
// the XOR row) instead of 64-bit ones (two shared-memory wavefronts per lookup).
template <bool XR, int W>
__device__ __forceinline__ void xor_row_acc(Acc<W> &acc, const uint4 &d)
{
    if constexpr (XR) {
        acc.x.x ^= d.x;
        acc.x.y ^= d.y;
        acc.x.z ^= d.z;
        acc.x.w ^= d.w;
    }
}

template <bool XR, bool T7, int W>
__device__ __forceinline__ void acc_clear(Acc<W> &acc)
{
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        acc.a[i] = 0;
        if constexpr (W == 8) acc.b[i] = 0;
        else if constexpr (T7) acc.t[i] = 0;
    }
    if constexpr (XR) acc.x = make_uint4(0u, 0u, 0u, 0u);
}

// compile-time loop over the fragments of batch B
template <int W, int K, int JB, int B, int JJ, bool T7, int XB = 0, bool XR = false>
struct BatchMac {
    __device__ __forceinline__ static void run(const uint4 (&d)[JB], uint32_t lane_base, uint32_t tb, Acc<W> &acc)
    {
        mac_vec<W, K, B * JB + JJ, T7, XB>(d[JJ], lane_base, tb, acc);
        xor_row_acc<XR>(acc, d[JJ]);
        if constexpr (JJ + 1 < JB) BatchMac<W, K, JB, B, JJ + 1, T7, XB, XR>::run(d, lane_base, tb, acc);
    }
};

__device__ __forceinline__ void transpose_store(const uint32_t *a, uint8_t *out, size_t out_pitch, size_t off,
                                                int row0, int nrows, const uint32_t *t = nullptr)
{
    uint32_t r[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint32_t a0 = a[4 * q], a1 = a[4 * q + 1], a2 = a[4 * q + 2], a3 = a[4 * q + 3];
        const uint32_t t0 = __byte_perm(a0, a1, 0x5140), t1 = __byte_perm(a2, a3, 0x5140);
        const uint32_t t2 = __byte_perm(a0, a1, 0x7362), t3 = __byte_perm(a2, a3, 0x7362);
        r[0][q] = __byte_perm(t0, t1, 0x5410);
        r[1][q] = __byte_perm(t0, t1, 0x7632);
        r[2][q] = __byte_perm(t2, t3, 0x5410);
        r[3][q] = __byte_perm(t2, t3, 0x7632);
        if (t) {   // compile-time known after inlining: the transposed bit-7 terms
            r[0][q] ^= t[q];
            r[1][q] ^= t[4 + q];
            r[2][q] ^= t[8 + q];
            r[3][q] ^= t[12 + q];
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i < nrows)
            stg_stream(out + (size_t)(row0 + i) * out_pitch + off, make_uint4(r[i][0], r[i][1], r[i][2], r[i][3]));
}

// Besides the register load, ask L2 to fetch the same rows pf_bytes further on: the column this thread
// will load one iteration later (pf_bytes = 0 near the end of the stripe: a harmless prefetch of the line
// being loaded).  One block per SM leaves too few bytes in flight to cover HBM latency from registers
// alone; with the prefetch the later register loads hit L2.  Measured: RS(16,20) 0.82 -> 0.92 of the peak.
// (RS(4,6), already at 0.99 without it, loses 12 % to the extra instructions: PF is per instance.)
// PAD: the instance is compiled for K fragments but the code has only k < K (k = 17 runs on the K = 18
// instance, ...): fragments J0+jj >= k are not loaded -- their tables are all zero, so whatever the
// registers hold contributes nothing -- and q only moves past them, it is never dereferenced there.
template <int JB, bool PF, bool PAD, int J0, typename PFT>
__device__ __forceinline__ void load_batch(uint4 (&d)[JB], const uint8_t *&q, size_t in_pitch, PFT pf_bytes, int k)
{
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) {
        if (!PAD || J0 + jj < k) {
            d[jj] = ldg_stream(q);
            if constexpr (PF) asm volatile("prefetch.global.L2 [%0];" ::"l"(q + pf_bytes));
        } else {
            d[jj] = make_uint4(0u, 0u, 0u, 0u);
        }
        q += in_pitch;
    }
}

// Build the tables of this block from the 32-entry global tables (lo 16 + hi 16 per fragment); the caller
// synchronises the block afterwards.
template <int W, int K, bool PAD, int XB = 0>
__device__ __forceinline__ void build_block_tables(uint32_t tb, const uint32_t *__restrict__ tab, int k)
{
    using M = Map<K, XB>;
    constexpr int WW = W / 4;            // words per entry
    constexpr int REP = 128 / W;         // replicas per entry: 32 (W=4) or 16 (W=8); one 128-byte row per x
    const int tid = threadIdx.x, nt = blockDim.x;
    // ---- build the tables of this block from the 32-entry global tables (lo 16 + hi 16 per fragment) ----
    // One flat loop over 16-byte stores (4 replicas of a 4-byte entry, 2 of an 8-byte entry) so that every
    // thread has several independent table reads in flight; the 32-entry tables (k * 32 * W bytes) stay
    // L1-resident after the first touch.  ~1-2 us per block instead of ~8 us for per-fragment loops.
    {
        constexpr int RV = REP * W / 16;                 // 16-byte stores per table row (= 8)
        constexpr int V8 = 256 * RV, V7 = 128 * RV;      // stores per fat8 / fat7 table
        constexpr int TOTAL = M::N8 * V8 + (M::NF - M::N8) * V7;
#pragma unroll 4
        for (int i = tid; i < TOTAL; i += nt) {
            int j, rem;
            if (i < M::N8 * V8) {
                j = i / V8;
                rem = i - j * V8;
            } else {
                const int i2 = i - M::N8 * V8;
                j = i2 / V7;
                rem = i2 - j * V7;
                j += M::N8;
            }
            const int x = rem / RV, r = rem - x * RV;
            const uint32_t *tj = tab + (size_t)j * 32 * WW;
            uint32_t v0 = 0, v1 = 0;
            if (!PAD || j < k) {      // padded fragments get all-zero tables (tab holds k fragments)
                v0 = __ldg(tj + (x & 15) * WW) ^ __ldg(tj + (16 + (x >> 4)) * WW);
                v1 = v0;
                if constexpr (W == 8) v1 = __ldg(tj + (x & 15) * WW + 1) ^ __ldg(tj + (16 + (x >> 4)) * WW + 1);
            }
            const uint32_t base = tb + (j < M::N8 ? j * 32768 : M::N8 * 32768 + (j - M::N8) * 16384);
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(base + x * 128 + r * 16), "r"(v0), "r"(v1),
                         "r"(v0), "r"(v1) : "memory");
        }
    }
    for (int i = tid; i < (K - M::NF) * 32 * WW; i += nt) {   // thin pages: straight copy, 256-byte stride
        const int j = i / (32 * WW), r = i - j * (32 * WW);
        const uint32_t v = (!PAD || M::NF + j < k) ? __ldg(tab + (size_t)(M::NF + j) * 32 * WW + r) : 0u;
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(tb + M::THIN0 + j * 256 + r * 4), "r"(v) : "memory");
    }
    for (int i = tid; i < K * 4; i += nt) {                   // product of 0x80 = hi-table entry 8
        const int j = i >> 2, q = i & 3;
        const bool real = !PAD || j < k;
        const uint32_t packed = (real && q < WW) ? __ldg(tab + (size_t)j * 32 * WW + (16 + 8) * WW + q) : 0u;
        const uint32_t c = real ? __ldg(tab + (size_t)j * 32 * WW + (16 + 8) * WW) : 0u;
        const uint32_t repl = ((c >> (8 * q)) & 0xFFu) * 0x01010101u;   // rows 0..3, one per word
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(tb + M::C7 + i * 4), "r"(packed) : "memory");
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(tb + M::C7R + i * 4), "r"(repl) : "memory");
    }
}

// Batch shape per k: X(k, JB fragments per batch, KB batches, PAD).  Measured on B200 (tools/sweep.py):
//   * batches of 4 with 768 threads beat batches of 8 with 512 (RS(16,20): 0.888 vs 0.794 of the HBM peak,
//     RS(8,12): 0.898 vs 0.816);
//   * an EVEN batch count matters more than the batch size: with an odd one the two register buffers swap
//     roles per column, the loop is unrolled over two columns and spills (same box, with the L2 prefetch:
//     RS(10,14) 2x5 0.944 vs 5x2 1.005; RS(12,16) 4x3 0.890 vs 3x4 0.989 of the measured copy peak);
//   * hence PAD = true: k without an even shape of its own runs on the next larger one with all-zero tables
//     for the missing fragments (their loads are skipped) -- a few wasted lookups, still one lookup per real
//     input byte.  Single-batch shape -> padded even shape, same box: RS(7,10) 0.822 -> 1.000, RS(11,15)
//     0.937 -> 1.003, RS(13,17) 0.749 -> 0.964, RS(5,8) 0.914 -> 0.972, RS(15,20) 5x3 0.507 -> 0.623;
//     nibble kernel -> padded shape: RS(17,20) 0.654 -> 0.842, RS(19,23) 0.667 -> 0.831, RS(23,27) 0.656 -> 0.762
//     (k = 28 on the 32 shape loses, 0.633 vs 0.674, so 25..29 stay on the nibble kernels).
//     k = 2, 3, 6 were tried on even shapes too: within noise of their single-batch shapes, left alone.
#define RSGPU_FAT_SHAPES(X)                                                                                   \
    X(2, 2, 1, false) X(3, 3, 1, false) X(4, 4, 1, false) X(5, 3, 2, true) X(6, 6, 1, false) X(7, 4, 2, true) \
    X(8, 4, 2, false) X(9, 3, 3, false) X(10, 5, 2, false) X(11, 3, 4, true) X(12, 3, 4, false)               \
    X(13, 7, 2, true) X(14, 7, 2, false) X(15, 4, 4, true) X(16, 4, 4, false) X(17, 3, 6, true)               \
    X(18, 3, 6, false) X(19, 5, 4, true) X(20, 5, 4, false) X(21, 4, 6, true) X(22, 4, 6, true)               \
    X(23, 4, 6, true) X(24, 4, 6, false) X(26, 7, 4, true) X(27, 7, 4, true)      \
    X(29, 4, 8, true) X(30, 4, 8, true) X(31, 4, 8, true) X(32, 4, 8, false)

// which instances use the L2 prefetch / the transposed bit-7 accumulation (measurements: make_fat)
__host__ __device__ constexpr bool fat_use_prefetch(int K) { return K >= 8; }
#ifndef RSGPU_T7_K10
#define RSGPU_T7_K10 0     // experiment switch: transposed bit-7 terms for the k = 10 shape too (sustained rate)
#endif
__host__ __device__ constexpr bool fat_use_t7(int W, int K)
{
    return W == 4 && (K == 12 || K == 14 || K == 16 || K == 18 || K == 20 || (RSGPU_T7_K10 && K == 10));
}

}  // namespace
}  // namespace rsgpu
