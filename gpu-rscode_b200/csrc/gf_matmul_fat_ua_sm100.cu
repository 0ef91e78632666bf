// gf_matmul_fat_ua_sm100.cu -- the byte-table kernel for rows that are NOT 16-byte aligned.
//
// Why: `rsgpu_encode(ctx, data_d, chunk, code_d, chunk, chunk, s)` with chunk = ceil(size / k) -- what a caller
// holding a file's fragments back to back passes (the reference's own layout, src/encode.cu:325-345) -- gives
// every row its own alignment; chunk is a multiple of 16 one time in sixteen.  The reference drops to its
// byte-per-thread kernel whenever chunk % 8 != 0 (src/matrix.cu:796,811); round 1 of this engine did the same
// (k_gf_matmul_bytes: 0.21 of the HBM peak on RS(10,14), L = 400 MiB + 3).  A 128-bit global access must be
// 16-byte aligned, so the rows are read and written as ALIGNED vectors and shifted into place in registers:
//
//   * logical byte c of input row j sits at A_j + c.  The column space behind a 16-byte head is cut into
//     16-byte windows; window n of row j straddles the aligned vectors n and n+1 of that row's own aligned
//     grid, at the row's byte offset s_j = (A_j + 16) & 15 (the same for every window of the row).
//   * lane l of a warp loads aligned vector 30 S + l of every row (S = span index of the warp); the vector
//     n+1 comes from lane l+1 by four SHFL.DOWN; a branch-free funnel of three PRMT stages (shift by 8, by 4,
//     by 0..3 bytes; s_j is the same for the whole grid, so the selectors live in uniform registers; gf_realign.cuh)
//     yields the window.  Lanes 0..30 of a span hold valid windows; the multiply-accumulate is the byte-table
//     code of gf_matmul_fat.cuh, unchanged.
//   * the outputs are realigned the same way: output row i has its offset u_i to the next 16-byte boundary; the
//     aligned vector that starts there takes bytes [u_i, 16) of window n and [0, u_i) of window n+1 -- lane l+1
//     again -- so lanes 0..29 store one ALIGNED 128-bit vector per row.  Spans advance by 30 windows: 2 of 32
//     lanes redo their neighbours' work (6.7 % more loads), and in return there are no partial stores, no
//     byte loops and no divergence anywhere in the steady state; sector traffic to L2 is that of aligned rows.
//   * the first 32 and the last < 48 bytes of every row (where an aligned vector would reach outside the row)
//     go to the byte kernel, as before.
// Measured: RS(10,14), L = 400 MiB + 3, pitch = L: 0.750 of the HBM peak (byte kernel 0.207, aligned rows 1.0); alu pipe
// 84 %, issue 73 % (profiles/r2_ncu_full_ua_rs1014.csv).  k without a byte-table shape: k_gf_matmul_ua (gf_matmul_sm100.cu).
#include "gf_matmul_fat.cuh"
#include "gf_realign.cuh"

namespace rsgpu {
namespace {

constexpr int kHead = kUaHead;
constexpr int kSpan = kUaSpan;

__host__ __device__ constexpr int fat_threads_ua(int W, int JB, int KB, bool T7)
{
    // the realignment needs ~20 more live registers than the aligned kernel: one block size down, at least 128
    // registers per thread, and one more size down for the long unrolled columns of k >= 18 (ptxas -v: no spills
    // beyond a few bytes with these)
    int t = fat_threads(W, JB, T7) - 128;
    if (W == 4 && t > 512) t = JB <= 2 ? 768 : 512;
    if (JB * KB >= 18) t -= 128;
    return t < 256 ? 256 : t;
}

// q walks the UNALIGNED logical addresses A_j + 16 + 16 n; the load goes to the aligned vector below it
template <int JB, bool PF, bool PAD, int J0, typename PFT>
__device__ __forceinline__ void load_batch_ua(uint4 (&d)[JB], const uint8_t *&q, size_t in_pitch, PFT pf_bytes, int k, bool inrange)
{
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) {
        if ((!PAD || J0 + jj < k) && inrange) {
            const uint8_t *qa = reinterpret_cast<const uint8_t *>(reinterpret_cast<uintptr_t>(q) & ~(uintptr_t)15);
            d[jj] = ldg_stream(qa);
            if constexpr (PF) asm volatile("prefetch.global.L2 [%0];" ::"l"(qa + pf_bytes));
        } else {
            d[jj] = make_uint4(0u, 0u, 0u, 0u);
        }
        q += in_pitch;
    }
}

template <int W, int K, int JB, int B, int JJ, bool T7, bool XR>
struct BatchMacUA {
    __device__ __forceinline__ static void run(const uint4 (&d)[JB], uint32_t s0, uint32_t pm, uint32_t lane_base, uint32_t tb,
                                               Acc<W> &acc)
    {
        constexpr int J = B * JB + JJ;
        const uint32_t s = (s0 + (uint32_t)J * pm) & 15u;          // offset of row J inside its aligned vectors
        const uint4 x = realign(d[JJ], shfl_next(d[JJ]), s);
        mac_vec<W, K, J, T7>(x, lane_base, tb, acc);
        xor_row_acc<XR>(acc, x);
        if constexpr (JJ + 1 < JB) BatchMacUA<W, K, JB, B, JJ + 1, T7, XR>::run(d, s0, pm, lane_base, tb, acc);
    }
};

struct UaGeom {
    const uint8_t *in_h;     // in + kHead
    uint8_t *out_h;          // out + kHead
    size_t in_pitch, out_pitch;
    size_t nv;               // aligned vectors per row that lie inside the row for every row
    size_t nst;              // store vectors per row = nv - 2
    uint32_t s0, pm;         // (in_h & 15), (in_pitch & 15)
    int k, nrows;
};

template <int W, int K, int JB, int KB, int B, int SP, bool PF, bool T7, bool PAD, bool XR>
struct ColumnBatchesUA {
    __device__ __forceinline__ static void run(uint4 (&buf0)[JB], uint4 (&buf1)[JB], const uint8_t *&q, const UaGeom &g, size_t n,
                                               size_t nn, bool have_next, uint32_t lane_base, uint32_t tb, Acc<W> &acc, uint32_t pf0,
                                               uint32_t pf1)
    {
        constexpr int PAR = (SP + B) & 1;
        uint4 (&cur)[JB] = PAR ? buf1 : buf0;
        uint4 (&nxt)[JB] = PAR ? buf0 : buf1;
        if constexpr (B + 1 < KB) {
            load_batch_ua<JB, PF, PAD, (B + 1) * JB>(nxt, q, g.in_pitch, pf0, g.k, n < g.nv);
        } else {
            if (have_next) {     // warp-uniform
                q = g.in_h + nn * 16;
                load_batch_ua<JB, PF, PAD, 0>(nxt, q, g.in_pitch, pf1, g.k, nn < g.nv);
            }
        }
        BatchMacUA<W, K, JB, B, 0, T7, XR>::run(cur, g.s0, g.pm, lane_base, tb, acc);
        if constexpr (B + 1 < KB)
            ColumnBatchesUA<W, K, JB, KB, B + 1, SP, PF, T7, PAD, XR>::run(buf0, buf1, q, g, n, nn, have_next, lane_base, tb, acc, pf0, pf1);
    }
};

template <int W, int K, int JB, int KB, int SP, bool PF, bool T7, bool PAD, bool XR>
__device__ __forceinline__ void do_column_ua(uint4 (&buf0)[JB], uint4 (&buf1)[JB], const uint8_t *&q, const UaGeom &g, size_t n, size_t nn,
                                             bool have_next, bool have_next2, uint32_t lane, uint32_t lane_base, uint32_t tb,
                                             uint32_t pf_stride)
{
    // L2 prefetch one step ahead; per thread, so that no prefetch is aimed past the end of a row
    const uint32_t pf0 = (have_next && nn < g.nv) ? pf_stride : 0u;
    const uint32_t pf1 = (have_next2 && nn + (nn - n) < g.nv) ? pf_stride : 0u;
    Acc<W> acc;
    acc_clear<XR, T7>(acc);
    ColumnBatchesUA<W, K, JB, KB, 0, SP, PF, T7, PAD, XR>::run(buf0, buf1, q, g, n, nn, have_next, lane_base, tb, acc, pf0, pf1);
    const bool ok = lane < (uint32_t)kSpan && n < g.nst;
    uint32_t r[4][4];
    if constexpr (W == 8) {
        transpose_rows(acc.a, r, nullptr);
        store_rows_realigned(r, g.out_h, g.out_pitch, n, 0, g.nrows, ok);
        transpose_rows(acc.b, r, nullptr);
        store_rows_realigned(r, g.out_h, g.out_pitch, n, 4, g.nrows - 4, ok);
    } else {
        transpose_rows(acc.a, r, (T7 && Map<K>::NF > Map<K>::N8) ? acc.t : nullptr);
        store_rows_realigned(r, g.out_h, g.out_pitch, n, 0, g.nrows, ok);
    }
    if constexpr (XR) {      // the peeled all-ones row: the row right above the group's rows
        r[0][0] = acc.x.x;
        r[0][1] = acc.x.y;
        r[0][2] = acc.x.z;
        r[0][3] = acc.x.w;
        store_rows_realigned(r, g.out_h, g.out_pitch, n, -1, 1, ok);
    }
}

template <int W, int JB, int KB, bool PF, bool T7, bool PAD, bool XR>
__global__ void __launch_bounds__(fat_threads_ua(W, JB, KB, T7), 1)
k_gf_matmul_fat_ua(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
                   const uint32_t *__restrict__ tab, int nrows, size_t nv, int k)
{
    constexpr int K = JB * KB;
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;

    build_block_tables<W, K, PAD>(tb, tab, k);
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lane_base = tb + (W == 4 ? lane * 4u : (lane & 15u) * 8u);
    UaGeom g;
    g.in_h = in + kHead;
    g.out_h = out + kHead;
    g.in_pitch = in_pitch;
    g.out_pitch = out_pitch;
    g.nv = nv;
    g.nst = nv - 2;
    g.s0 = (uint32_t)reinterpret_cast<uintptr_t>(g.in_h) & 15u;
    g.pm = (uint32_t)in_pitch & 15u;
    g.k = k;
    g.nrows = nrows;
    const size_t nspan = (g.nst + kSpan - 1) / kSpan;
    const size_t wtotal = (size_t)gridDim.x * (blockDim.x >> 5);
    const size_t nstride = wtotal * kSpan;                 // windows between two steps of a warp
    const uint32_t pf_stride = (uint32_t)(nstride * 16);
    // Loop control depends on blockIdx and kernel arguments only, so every warp of the block runs the same number of
    // steps and the compiler sees converged warps at the shuffles (no WARPSYNC / reconvergence code around them).
    // A warp whose span lies past the end in the last step still runs it: its loads and stores are predicated off.
    const size_t wpb = blockDim.x >> 5;
    size_t S0 = (size_t)blockIdx.x * wpb;                  // span of warp 0 of this block
    size_t n = (S0 + (threadIdx.x >> 5)) * kSpan + lane;
    uint4 buf0[JB], buf1[JB];
    const uint8_t *q = g.in_h + n * 16;
    if (S0 < nspan) load_batch_ua<JB, PF, PAD, 0>(buf0, q, in_pitch, S0 + wtotal < nspan ? pf_stride : 0u, k, n < nv);

    while (S0 < nspan) {
        do_column_ua<W, K, JB, KB, 0, PF, T7, PAD, XR>(buf0, buf1, q, g, n, n + nstride, S0 + wtotal < nspan, S0 + 2 * wtotal < nspan, lane,
                                                       lane_base, tb, pf_stride);
        S0 += wtotal;
        n += nstride;
        if constexpr (KB & 1) {
            if (S0 >= nspan) break;
            do_column_ua<W, K, JB, KB, 1, PF, T7, PAD, XR>(buf0, buf1, q, g, n, n + nstride, S0 + wtotal < nspan, S0 + 2 * wtotal < nspan,
                                                           lane, lane_base, tb, pf_stride);
            S0 += wtotal;
            n += nstride;
        }
    }
}

using FatFn = void (*)(const uint8_t *, size_t, uint8_t *, size_t, const uint32_t *, int, size_t, int);

struct FatKernel {
    FatFn fn;
    int threads;
    int smem;
};

template <int W, int JB, int KB, bool PAD = false, bool XR = false>
FatKernel make_fat_ua()
{
    constexpr int K = JB * KB;
    constexpr bool PF = fat_use_prefetch(K);
    constexpr bool T7 = fat_use_t7(W, K);
    return FatKernel{k_gf_matmul_fat_ua<W, JB, KB, PF, T7, PAD, XR>, fat_threads_ua(W, JB, KB, T7), Map<K>::BYTES + 256};
}

template <int W, bool XR = false>
bool pick_fat_ua(int k, FatKernel *fk)
{
    switch (k) {
#define RSGPU_FAT_CASE(k_, jb_, kb_, pad_) case k_: *fk = make_fat_ua<W, jb_, kb_, pad_, XR>(); return true;
        RSGPU_FAT_SHAPES(RSGPU_FAT_CASE)
#undef RSGPU_FAT_CASE
    default: return false;
    }
}

}  // namespace

// Rows with arbitrary alignment.  Codes the logical columns [lo, hi) it reports back through *vec_lo / *vec_hi
// (the caller sends [0, lo) and [hi, len) to the byte kernel).  cudaErrorNotSupported: no instance for this k, or
// the stripe is too short to hold one span.
cudaError_t launch_matmul_fat_ua(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g, size_t out_pitch,
                                 size_t len, const DeviceInfo &dev, int max_blocks, cudaStream_t s, size_t *vec_lo, size_t *vec_hi)
{
    if (len < (size_t)kHead + 48 + 16 * kSpan) return cudaErrorNotSupported;
    FatKernel fk;
    if (g.xor_row >= 0 && g.width != 4) return cudaErrorNotSupported;
    const bool ok = g.xor_row >= 0 ? pick_fat_ua<4, true>(k, &fk) : (g.width == 4 ? pick_fat_ua<4>(k, &fk) : pick_fat_ua<8>(k, &fk));
    if (!ok || fk.smem > dev.max_smem_optin) return cudaErrorNotSupported;
    int occ = 0;
    cudaError_t e = prepare_launch(dev, (const void *)fk.fn, fk.threads, (size_t)fk.smem, &occ);
    if (e != cudaSuccess) return e;
    const size_t nv = (len - kHead) / 16;          // vectors n = 0 .. nv-1 stay inside every row
    const size_t nst = nv - 2;
    const size_t nspan = (nst + kSpan - 1) / kSpan;
    const size_t wpb = (size_t)fk.threads / 32;
    size_t want = (nspan + wpb - 1) / wpb;
    size_t cap = (size_t)dev.sm_count;
    if (max_blocks > 0 && (size_t)max_blocks < cap) cap = max_blocks;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    fk.fn<<<grid, fk.threads, fk.smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, g.nrows, nv, k);
    // every row is covered from kHead + u_i (u_i <= 15) to kHead + u_i + 16 nst: the caller's byte ranges are the
    // superset [0, kHead + 16) and [kHead + 16 nst, len) -- a few bytes are written twice with the same value
    *vec_lo = kHead + 16;
    *vec_hi = (size_t)kHead + 16 * nst;
    return cudaGetLastError();
}

}  // namespace rsgpu
