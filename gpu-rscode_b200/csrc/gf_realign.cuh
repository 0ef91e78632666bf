// gf_realign.cuh -- in-register realignment shared by the kernels for rows that are not 16-byte aligned
// (gf_matmul_fat_ua_sm100.cu: byte tables; gf_matmul_sm100.cu k_gf_matmul_ua: split-nibble tables).
//
// Geometry (the same in both kernels): the column space behind a 16-byte head (kUaHead) is cut into 16-byte windows;
// window n of a row straddles the aligned vectors n and n+1 of that row's own aligned grid at the row's byte offset s
// (one value per row for the whole launch).  Lane l of a warp loads aligned vector kUaSpan*S + l (S = the warp's span),
// takes vector n+1 from lane l+1 with four SHFL.DOWN, and realign() yields the window; lanes 0..30 hold valid windows.
// Outputs go the same way back: the aligned vector that starts u bytes into window n takes the rest from window n+1, so
// lanes 0..29 store one aligned vector per output row, and a span advances by kUaSpan = 30 windows.
#pragma once
#include <stdint.h>

namespace rsgpu {
namespace {

constexpr int kUaHead = 16;     // logical bytes in front of window 0 (keeps every aligned vector inside its row)
constexpr int kUaSpan = 30;     // windows stored per warp per step

__device__ __forceinline__ uint4 shfl_next(const uint4 &v)
{
    uint4 r;
    r.x = __shfl_down_sync(0xffffffffu, v.x, 1);
    r.y = __shfl_down_sync(0xffffffffu, v.y, 1);
    r.z = __shfl_down_sync(0xffffffffu, v.z, 1);
    r.w = __shfl_down_sync(0xffffffffu, v.w, 1);
    return r;
}

// bytes [s, s + 16) of the 32-byte pair (o, n); 0 <= s <= 15, the same for every lane of the warp -- and of the
// grid: s depends on the row only, so where the row is a compile-time index the three selectors below are loop-invariant
// and ptxas keeps them in uniform registers.  Three PRMT stages (shift by 8, by 4, by 0..3 bytes), no predicates.
__device__ __forceinline__ uint4 realign(const uint4 &o, const uint4 &n, uint32_t s)
{
    const uint32_t sa = (s & 8u) ? 0x7654u : 0x3210u;        // take the word two places up
    const uint32_t sb = (s & 4u) ? 0x7654u : 0x3210u;        // take the word one place up
    const uint32_t sc = 0x3210u + 0x1111u * (s & 3u);        // funnel by 0..3 bytes
    const uint32_t t0 = __byte_perm(o.x, o.z, sa), t1 = __byte_perm(o.y, o.w, sa), t2 = __byte_perm(o.z, n.x, sa),
                   t3 = __byte_perm(o.w, n.y, sa), t4 = __byte_perm(n.x, n.z, sa), t5 = __byte_perm(n.y, n.w, sa);
    const uint32_t u0 = __byte_perm(t0, t1, sb), u1 = __byte_perm(t1, t2, sb), u2 = __byte_perm(t2, t3, sb),
                   u3 = __byte_perm(t3, t4, sb), u4 = __byte_perm(t4, t5, sb);
    return make_uint4(__byte_perm(u0, u1, sc), __byte_perm(u1, u2, sc), __byte_perm(u2, u3, sc), __byte_perm(u3, u4, sc));
}

// 16 accumulator words (byte r of word b = row r at byte position b) -> 4 rows x 4 words; t (optional): terms that
// are already in row order, XOR-ed in
__device__ __forceinline__ void transpose_rows(const uint32_t *a, uint32_t (&r)[4][4], const uint32_t *t)
{
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint32_t a0 = a[4 * q], a1 = a[4 * q + 1], a2 = a[4 * q + 2], a3 = a[4 * q + 3];
        const uint32_t t0 = __byte_perm(a0, a1, 0x5140), t1 = __byte_perm(a2, a3, 0x5140);
        const uint32_t t2 = __byte_perm(a0, a1, 0x7362), t3 = __byte_perm(a2, a3, 0x7362);
        r[0][q] = __byte_perm(t0, t1, 0x5410);
        r[1][q] = __byte_perm(t0, t1, 0x7632);
        r[2][q] = __byte_perm(t2, t3, 0x5410);
        r[3][q] = __byte_perm(t2, t3, 0x7632);
        if (t) {
            r[0][q] ^= t[q];
            r[1][q] ^= t[4 + q];
            r[2][q] ^= t[8 + q];
            r[3][q] ^= t[12 + q];
        }
    }
}

// Rows row0 .. row0+3 (row0 may be -1: the peeled row above the group) of window n: every lane hands its window to the
// lane below, which stores the aligned vector that starts u bytes into its own window.  out_h = out + kUaHead.
__device__ __forceinline__ void store_rows_realigned(const uint32_t (&r)[4][4], uint8_t *out_h, size_t out_pitch, size_t n, int row0,
                                                     int nrows, bool ok)
{
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (i < nrows) {     // uniform
            const uint4 own = make_uint4(r[i][0], r[i][1], r[i][2], r[i][3]);
            const uint4 nx = shfl_next(own);
            uint8_t *b = out_h + (ptrdiff_t)(row0 + i) * (ptrdiff_t)out_pitch + n * 16;
            const uint32_t u = (0u - (uint32_t)reinterpret_cast<uintptr_t>(b)) & 15u;
            const uint4 v = realign(own, nx, u);
            if (ok)
                asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(b + u), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                             : "memory");
        }
    }
}

}  // namespace
}  // namespace rsgpu
