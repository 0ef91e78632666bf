// gf_matmul_fat_mp_sm100.cu -- the byte-table kernel in PASSES over the fragments, for codes whose tables do not
// fit one SM's shared memory (k >= 25, and every k above 32: the reference's own sweeps go to k = 128,
// doc/result-graph/Total-GPU-encoding-time-2.pdf).  MEASURED, NOT USED BY DEFAULT (tuning variant 9): bit-exact, but
// slower than the nibble kernels on every such code -- RS(28,32) 0.55 vs 0.67, RS(40,44) 0.50 vs 0.66, RS(64,68) 0.48 vs
// 0.64, RS(128,132) 0.42 vs 0.56 of the HBM peak.  ncu (profiles/r2_ncu_full_mp_rs6468.csv): 7.1 instructions per input
// byte against 4.9 for the lookups alone (the table rebuild per pass and its barriers), 12 warps per SM, top stall
// long-scoreboard.  Staging the 32-entry source tables in shared memory and an L2 prefetch of each pass's bytes ahead of
// the rebuild brought 0.43 -> 0.48 on RS(64,68); the rebuild (14336 shared-memory stores per 336 KiB of input) stays.
//
// One lookup per input byte needs a bank-replicated table of 16 KiB per fragment (gf_matmul_fat_sm100.cu), and 227 KiB
// hold 14 of them.  The single-pass kernel therefore leaves the other fragments on split-nibble pages (two lookups,
// the shared-memory pipe saturates: RS(28,32) 0.67, RS(64,68) 0.65, RS(128,132) 0.57 of the HBM peak in round 1).
// Here the fragments are taken 14 at a time: a block builds the byte tables of fragments [14 ps, 14 ps + 14), every
// thread runs those 14 fragments of its C columns through them -- the partial parities of the C columns stay in
// registers across the passes -- then the block rebuilds the tables for the next 14.  Every input byte is still read
// exactly once and every output byte written once: traffic stays (k + rows) * len.  What a pass costs extra is the table
// build (14336 16-byte shared-memory stores per block, shifts-and-masks addressing) and two block barriers, spread over
// 384 threads x C columns x 14 fragments x 16 bytes = 336 KiB of input.
//
// The multiply-accumulate, the loads into two ping-pong register buffers and the transposing stores are the code of
// gf_matmul_fat.cuh on the K = 14 shape (7 fragments per batch, two batches); the first batch of the NEXT column / pass /
// block step is in flight while the current one is consumed, also across the table rebuild.
#include "gf_matmul_fat.cuh"

namespace rsgpu {
namespace {

constexpr int kPassK = 14;       // fragments per pass: all on fat7 tables
constexpr int kPassJB = 7;       // fragments per load batch (two batches per pass)
constexpr int kMpThreads = 384;  // 168 registers per thread: C accumulator sets + two batches of vectors

static_assert(Map<kPassK>::NF == kPassK && Map<kPassK>::N8 == 0, "the pass shape must be all 7-bit byte tables");

// Tables of one pass from the 32-entry global tables of its fragments (lo 16 + hi 16 entries per fragment); fragments
// behind kr (the last pass of a k that is not a multiple of 14) get all-zero tables.  Power-of-two geometry: store i
// is fragment i >> 10, table row (i >> 3) & 127, 16-byte chunk i & 7.
template <int W>
__device__ __forceinline__ void build_pass_tables(uint32_t tb, const uint32_t *__restrict__ tab, int kr)
{
    using M = Map<kPassK>;
    constexpr int WW = W / 4;
    constexpr bool STAGE = (W == 4);              // 14 x 32 x 4 bytes of staging fit behind the tables, 14 x 32 x 8 do not
    constexpr uint32_t SRC = M::BYTES;
    const int tid = threadIdx.x, nt = blockDim.x;
    // the 32-entry tables come from L2 (no L1 to speak of beside 227 KiB of shared memory): fetch them once per pass
    // into shared memory, so that the 14336 table-row stores below read their two source entries at shared-memory latency
    if constexpr (STAGE) {
        for (int i = tid; i < kPassK * 32; i += nt) {
            const uint32_t v = (i >> 5) < kr ? __ldg(tab + i) : 0u;
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(tb + SRC + i * 4), "r"(v) : "memory");
        }
        __syncthreads();
    }
#pragma unroll 4
    for (int i = tid; i < kPassK * 1024; i += nt) {
        const int j = i >> 10, x = (i >> 3) & 127, r = i & 7;
        uint32_t v0 = 0, v1 = 0;
        if constexpr (STAGE) {
            const uint32_t sj = tb + SRC + j * 128;
            uint32_t a, b;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a) : "r"(sj + (x & 15) * 4));
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(b) : "r"(sj + (16 + (x >> 4)) * 4));
            v0 = v1 = a ^ b;
        } else if (j < kr) {
            const uint32_t *tj = tab + (size_t)j * 32 * WW;
            v0 = __ldg(tj + (x & 15) * WW) ^ __ldg(tj + (16 + (x >> 4)) * WW);
            v1 = __ldg(tj + (x & 15) * WW + 1) ^ __ldg(tj + (16 + (x >> 4)) * WW + 1);
        }
        asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(tb + j * 16384 + x * 128 + r * 16), "r"(v0), "r"(v1), "r"(v0),
                     "r"(v1) : "memory");
    }
    for (int i = tid; i < kPassK * 4; i += nt) {          // product of 0x80 (hi-table entry 8): the bit-7 term
        const int j = i >> 2, q = i & 3;
        const bool real = j < kr;
        const uint32_t packed = (real && q < WW) ? __ldg(tab + (size_t)j * 32 * WW + (16 + 8) * WW + q) : 0u;
        const uint32_t c = real ? __ldg(tab + (size_t)j * 32 * WW + (16 + 8) * WW) : 0u;
        const uint32_t repl = ((c >> (8 * q)) & 0xFFu) * 0x01010101u;
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(tb + M::C7 + i * 4), "r"(packed) : "memory");
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(tb + M::C7R + i * 4), "r"(repl) : "memory");
    }
}

// One column of one pass: two batches; the loads that follow the last batch go to q_next (the next column of this
// pass, the first column of the next pass or of the next block step; nullptr = nothing follows).
template <int W, int B, bool XR>
struct ColumnBatchesMP {
    __device__ __forceinline__ static void run(uint4 (&buf0)[kPassJB], uint4 (&buf1)[kPassJB], const uint8_t *&q, size_t in_pitch,
                                               const uint8_t *q_next, uint32_t lane_base, uint32_t tb, Acc<W> &acc, int kr, int kr_next)
    {
        uint4 (&cur)[kPassJB] = (B & 1) ? buf1 : buf0;
        uint4 (&nxt)[kPassJB] = (B & 1) ? buf0 : buf1;
        if constexpr (B == 0) {
            load_batch<kPassJB, false, true, kPassJB>(nxt, q, in_pitch, 0u, kr);
        } else {
            if (q_next) {
                q = q_next;
                load_batch<kPassJB, false, true, 0>(nxt, q, in_pitch, 0u, kr_next);
            }
        }
        BatchMac<W, kPassK, kPassJB, B, 0, false, 0, XR>::run(cur, lane_base, tb, acc);
        if constexpr (B == 0) ColumnBatchesMP<W, 1, XR>::run(buf0, buf1, q, in_pitch, q_next, lane_base, tb, acc, kr, kr_next);
    }
};

template <int W, int C, bool XR>
__global__ void __launch_bounds__(kMpThreads, 1)
k_gf_matmul_fat_mp(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
                   const uint32_t *__restrict__ tab, int nrows, size_t nsteps, int k)
{
    constexpr int WW = W / 4;
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lane_base = tb + (W == 4 ? lane * 4u : (lane & 15u) * 8u);
    const size_t T = blockDim.x;
    const int np = (k + kPassK - 1) / kPassK;
    const int kr_first = k < kPassK ? k : kPassK;

    uint4 buf0[kPassJB], buf1[kPassJB];
    size_t step = blockIdx.x;                 // a block step = T * C consecutive columns
    const uint8_t *q = in + (step * T * C + threadIdx.x) * 16;
    if (step < nsteps) load_batch<kPassJB, false, true, 0>(buf0, q, in_pitch, 0u, kr_first);

    for (; step < nsteps; step += gridDim.x) {
        const size_t v0 = step * T * C + threadIdx.x;
        Acc<W> acc[C];
#pragma unroll
        for (int c = 0; c < C; ++c) acc_clear<XR, false>(acc[c]);
        for (int ps = 0; ps < np; ++ps) {
            const int kr = k - ps * kPassK < kPassK ? k - ps * kPassK : kPassK;
            const uint8_t *in_ps = in + (size_t)ps * kPassK * in_pitch;
            // ask L2 for this thread's bytes of the pass now: the table build below does no global traffic and hides
            // the HBM latency (column 0's first batch is already in registers)
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const uint8_t *pq = in_ps + (v0 + (size_t)c * T) * 16;
                for (int j = 0; j < kr; ++j, pq += in_pitch) asm volatile("prefetch.global.L2 [%0];" ::"l"(pq));
            }
            __syncthreads();                  // every thread has finished with the previous pass's tables
            build_pass_tables<W>(tb, tab + (size_t)ps * kPassK * 32 * WW, kr);
            __syncthreads();
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const uint8_t *qn = nullptr;
                int krn = kr;
                if (c + 1 < C) {
                    qn = in_ps + (v0 + (size_t)(c + 1) * T) * 16;
                } else if (ps + 1 < np) {
                    qn = in_ps + (size_t)kPassK * in_pitch + v0 * 16;
                    krn = k - (ps + 1) * kPassK < kPassK ? k - (ps + 1) * kPassK : kPassK;
                } else if (step + gridDim.x < nsteps) {
                    qn = in + ((step + gridDim.x) * T * C + threadIdx.x) * 16;
                    krn = kr_first;
                }
                ColumnBatchesMP<W, 0, XR>::run(buf0, buf1, q, in_pitch, qn, lane_base, tb, acc[c], kr, krn);
            }
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const size_t off = (v0 + (size_t)c * T) * 16;
            transpose_store(acc[c].a, out, out_pitch, off, 0, nrows);
            if constexpr (W == 8) transpose_store(acc[c].b, out, out_pitch, off, 4, nrows - 4);
            if constexpr (XR) stg_stream(out - out_pitch + off, acc[c].x);
        }
    }
}

}  // namespace

// Codes the first *done_vec 16-byte columns (a whole number of block steps); the caller codes the rest.
// cudaErrorNotSupported: stripe shorter than one block step, or a peeled row with 64-bit entries.
cudaError_t launch_matmul_fat_mp(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g, size_t out_pitch,
                                 size_t nvec, const DeviceInfo &dev, int max_blocks, cudaStream_t s, size_t *done_vec)
{
    *done_vec = 0;
    if (g.xor_row >= 0 && g.width != 4) return cudaErrorNotSupported;
    using Fn = void (*)(const uint8_t *, size_t, uint8_t *, size_t, const uint32_t *, int, size_t, int);
    const int C = g.width == 4 ? 4 : 2;
    const Fn fn = g.width == 4 ? (g.xor_row >= 0 ? (Fn)k_gf_matmul_fat_mp<4, 4, true> : (Fn)k_gf_matmul_fat_mp<4, 4, false>)
                               : (Fn)k_gf_matmul_fat_mp<8, 2, false>;
    const size_t per_step = (size_t)kMpThreads * C;
    const size_t nsteps = nvec / per_step;
    const int smem = Map<kPassK>::BYTES + kPassK * 32 * 4 + 256;   // tables + staging of the 32-entry source tables (32-bit entries)
    if (nsteps == 0 || smem > dev.max_smem_optin) return cudaErrorNotSupported;
    int occ = 0;
    cudaError_t e = prepare_launch(dev, (const void *)fn, kMpThreads, (size_t)smem, &occ);
    if (e != cudaSuccess) return e;
    size_t cap = (size_t)dev.sm_count;
    if (max_blocks > 0 && (size_t)max_blocks < cap) cap = max_blocks;
    const unsigned grid = (unsigned)(nsteps < cap ? nsteps : cap);
    fn<<<grid, kMpThreads, smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, g.nrows, nsteps, k);
    *done_vec = nsteps * per_step;
    return cudaGetLastError();
}

}  // namespace rsgpu
