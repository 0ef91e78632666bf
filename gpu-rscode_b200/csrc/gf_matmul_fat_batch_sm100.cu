// gf_matmul_fat_batch_sm100.cu -- the byte-table kernel over a BATCH of stripes in one launch.
//
// A storage node codes many stripes of the same (k, p) at a time, and small ones (fragments of 64 KiB - a few
// MiB) are launch-bound when every stripe is its own launch: the ~5 us of launch + table construction + ramp
// dwarf the data time (RS(10,14), 1 MiB fragments: 14.7 MB per launch).  Here one launch builds the tables
// once per SM and walks the VIRTUAL column space of the whole batch: virtual column V = stripe V / nvec,
// column V % nvec of that stripe, so the grid-stride loop stays perfectly balanced whatever the stripe size
// (a stripe of 65536 columns would leave a third of a 94720-thread grid idle if stripes were walked one
// after the other).  A thread keeps (stripe, column) of its current, next and next-but-one columns and
// advances them by the grid stride, pre-split by the host into (stride / nvec, stride % nvec): no division
// in the loop.  Loads, the L2 prefetch one stride ahead, multiply-accumulate, transpose and stores are the
// code of gf_matmul_fat_sm100.cu (gf_matmul_fat.cuh).
//
// Layout: fragment j of stripe s starts at base + s * stripe_pitch + j * frag_pitch (both layouts of a
// batch -- stripe-major and fragment-major -- are instances of this).  No counterpart in the reference,
// which codes one file per process run.
#include "gf_matmul_fat.cuh"

namespace rsgpu {
namespace {

struct Pos {
    uint32_t s, c;   // stripe, 16-byte column within the stripe
};

struct BatchGeom {
    size_t in_stripe, out_stripe;   // bytes between stripes
    uint32_t nvec;                  // 16-byte columns per stripe
    uint32_t n_stripes;
    uint32_t step_s, step_c;        // grid stride split into whole stripes and columns
};

__device__ __forceinline__ void advance(Pos &p, const BatchGeom &g)
{
    p.s += g.step_s;
    p.c += g.step_c;
    if (p.c >= g.nvec) {
        p.c -= g.nvec;
        ++p.s;
    }
}
__device__ __forceinline__ bool valid(const Pos &p, const BatchGeom &g) { return p.s < g.n_stripes; }
__device__ __forceinline__ size_t in_off(const Pos &p, const BatchGeom &g) { return (size_t)p.s * g.in_stripe + (size_t)p.c * 16; }
__device__ __forceinline__ size_t out_off(const Pos &p, const BatchGeom &g) { return (size_t)p.s * g.out_stripe + (size_t)p.c * 16; }

// As ColumnBatches of gf_matmul_fat_sm100.cu, with the next column given as a pointer (nullptr = none) and the
// prefetch distances as 64-bit byte deltas (the next column may sit in another stripe).
template <int W, int K, int JB, int KB, int B, int SP, bool PF, bool T7, bool PAD, bool XR>
struct ColumnBatchesB {
    __device__ __forceinline__ static void run(uint4 (&buf0)[JB], uint4 (&buf1)[JB], const uint8_t *&q, size_t in_pitch,
                                               const uint8_t *q_next, uint32_t lane_base, uint32_t tb, Acc<W> &acc,
                                               ptrdiff_t pf0, ptrdiff_t pf1, int k)
    {
        constexpr int PAR = (SP + B) & 1;
        uint4 (&cur)[JB] = PAR ? buf1 : buf0;
        uint4 (&nxt)[JB] = PAR ? buf0 : buf1;
        if constexpr (B + 1 < KB) {
            load_batch<JB, PF, PAD, (B + 1) * JB>(nxt, q, in_pitch, pf0, k);
        } else {
            if (q_next) {
                q = q_next;
                load_batch<JB, PF, PAD, 0>(nxt, q, in_pitch, pf1, k);
            }
        }
        BatchMac<W, K, JB, B, 0, T7, 0, XR>::run(cur, lane_base, tb, acc);
        if constexpr (B + 1 < KB)
            ColumnBatchesB<W, K, JB, KB, B + 1, SP, PF, T7, PAD, XR>::run(buf0, buf1, q, in_pitch, q_next, lane_base, tb, acc, pf0, pf1, k);
    }
};

template <int W, int K, int JB, int KB, int SP, bool PF, bool T7, bool PAD, bool XR>
__device__ __forceinline__ void do_column_b(uint4 (&buf0)[JB], uint4 (&buf1)[JB], const uint8_t *&q, size_t in_pitch,
                                            const uint8_t *in, uint8_t *out, size_t out_pitch, const Pos &p, const Pos &pn,
                                            const Pos &pnn, const BatchGeom &g, uint32_t lane_base, uint32_t tb, int nrows, int k)
{
    const bool has_n = valid(pn, g), has_nn = valid(pnn, g);
    const size_t o = in_off(p, g), on = in_off(pn, g);
    // L2 prefetch targets: the same rows of the column this thread loads one iteration later (0 = the line
    // being loaded itself, harmless, when there is no such column)
    const ptrdiff_t pf0 = has_n ? (ptrdiff_t)(on - o) : 0;
    const ptrdiff_t pf1 = has_nn ? (ptrdiff_t)(in_off(pnn, g) - on) : 0;
    Acc<W> acc;
    acc_clear<XR, T7>(acc);
    ColumnBatchesB<W, K, JB, KB, 0, SP, PF, T7, PAD, XR>::run(buf0, buf1, q, in_pitch, has_n ? in + on : nullptr, lane_base, tb, acc, pf0, pf1, k);
    const size_t oo = out_off(p, g);
    if constexpr (W == 8) {
        transpose_store(acc.a, out, out_pitch, oo, 0, nrows);
        transpose_store(acc.b, out, out_pitch, oo, 4, nrows - 4);
    } else {
        transpose_store(acc.a, out, out_pitch, oo, 0, nrows, (T7 && Map<K>::NF > Map<K>::N8) ? acc.t : nullptr);
    }
    if constexpr (XR) stg_stream(out - out_pitch + oo, acc.x);   // the peeled all-ones row, right above the group's rows
}

template <int W, int JB, int KB, bool PF, bool T7, bool PAD, bool XR>
__global__ void __launch_bounds__(fat_threads_x(W, JB, KB, T7, PAD, XR), 1)
k_gf_matmul_fat_batch(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
                      const uint32_t *__restrict__ tab, int nrows, BatchGeom g, int k)
{
    constexpr int K = JB * KB;
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;
    build_block_tables<W, K, PAD>(tb, tab, k);
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lane_base = tb + (W == 4 ? lane * 4u : (lane & 15u) * 8u);
    const uint32_t v0 = blockIdx.x * blockDim.x + threadIdx.x;     // < grid stride < 2^32
    Pos p{v0 / g.nvec, v0 % g.nvec};
    Pos pn = p;
    advance(pn, g);
    Pos pnn = pn;
    advance(pnn, g);
    uint4 buf0[JB], buf1[JB];
    const uint8_t *q = in + in_off(p, g);
    if (valid(p, g))
        load_batch<JB, PF, PAD, 0>(buf0, q, in_pitch, valid(pn, g) ? (ptrdiff_t)(in_off(pn, g) - in_off(p, g)) : (ptrdiff_t)0, k);

    while (valid(p, g)) {
        do_column_b<W, K, JB, KB, 0, PF, T7, PAD, XR>(buf0, buf1, q, in_pitch, in, out, out_pitch, p, pn, pnn, g, lane_base, tb, nrows, k);
        p = pn;
        pn = pnn;
        advance(pnn, g);
        if constexpr (KB & 1) {          // odd batch count: the next column starts in the other buffer
            if (!valid(p, g)) break;
            do_column_b<W, K, JB, KB, 1, PF, T7, PAD, XR>(buf0, buf1, q, in_pitch, in, out, out_pitch, p, pn, pnn, g, lane_base, tb, nrows, k);
            p = pn;
            pn = pnn;
            advance(pnn, g);
        }
    }
}

using FatBatchFn = void (*)(const uint8_t *, size_t, uint8_t *, size_t, const uint32_t *, int, BatchGeom, int);

struct FatBatchKernel {
    FatBatchFn fn;
    int threads;
    int smem;
};

template <int W, int JB, int KB, bool PAD, bool XR>
FatBatchKernel make_fat_batch()
{
    constexpr int K = JB * KB;
    constexpr bool PF = fat_use_prefetch(K), T7 = fat_use_t7(W, K);
    return FatBatchKernel{k_gf_matmul_fat_batch<W, JB, KB, PF, T7, PAD, XR>, fat_threads_x(W, JB, KB, T7, PAD, XR), Map<K>::BYTES + 256};
}

// the same shape per k as the single-stripe kernel (RSGPU_FAT_SHAPES, gf_matmul_fat.cuh)
template <int W, bool XR = false>
bool pick_fat_batch(int k, FatBatchKernel *fk)
{
    switch (k) {
#define RSGPU_FAT_CASE(k_, jb_, kb_, pad_) case k_: *fk = make_fat_batch<W, jb_, kb_, pad_, XR>(); return true;
        RSGPU_FAT_SHAPES(RSGPU_FAT_CASE)
#undef RSGPU_FAT_CASE
    default: return false;
    }
}

}  // namespace

// cudaErrorNotSupported when there is no batch instance for this k or the geometry does not fit 32-bit
// column counters (the caller then codes the stripes one launch each).  All pointers, pitches and stripe
// pitches must be 16-byte aligned and len a multiple of 16: nvec = len / 16.
cudaError_t launch_matmul_fat_batch(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, size_t in_stripe,
                                    uint8_t *out_g, size_t out_pitch, size_t out_stripe, size_t nvec, size_t n_stripes,
                                    const DeviceInfo &dev, int max_blocks, cudaStream_t s)
{
    FatBatchKernel fk;
    if (g.xor_row >= 0 && g.width != 4) return cudaErrorNotSupported;
    const bool ok = g.xor_row >= 0 ? pick_fat_batch<4, true>(k, &fk)
                                   : (g.width == 4 ? pick_fat_batch<4>(k, &fk) : pick_fat_batch<8>(k, &fk));
    if (!ok || fk.smem > dev.max_smem_optin) return cudaErrorNotSupported;
    if (nvec == 0 || nvec >= ((size_t)1 << 31) || n_stripes == 0 || n_stripes >= ((size_t)1 << 31)) return cudaErrorNotSupported;
    int occ = 0;
    cudaError_t e = prepare_launch(dev, (const void *)fk.fn, fk.threads, (size_t)fk.smem, &occ);
    if (e != cudaSuccess) return e;
    const size_t total = nvec * n_stripes;
    size_t want = (total + fk.threads - 1) / fk.threads;
    size_t cap = (size_t)dev.sm_count;      // one block per SM: the tables fill its shared memory
    if (max_blocks > 0 && (size_t)max_blocks < cap) cap = max_blocks;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    const size_t stride = (size_t)grid * fk.threads;
    BatchGeom geo;
    geo.in_stripe = in_stripe;
    geo.out_stripe = out_stripe;
    geo.nvec = (uint32_t)nvec;
    geo.n_stripes = (uint32_t)n_stripes;
    geo.step_s = (uint32_t)(stride / nvec);
    geo.step_c = (uint32_t)(stride % nvec);
    fk.fn<<<grid, fk.threads, fk.smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, g.nrows, geo, k);
    return cudaGetLastError();
}

}  // namespace rsgpu
