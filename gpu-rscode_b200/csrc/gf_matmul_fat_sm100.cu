// gf_matmul_fat_sm100.cu -- second generation of the hot kernel: bank-replicated byte tables.
//
// Why: the split-nibble kernel (gf_matmul_sm100.cu) is bound by the shared-memory data pipe, not by
// HBM: every input byte costs two 4-byte lookups = two wavefronts per warp, and at 1 wavefront per clock
// per SM that caps a B200 at ~4.1 TB/s of input, below the ~4.6 TB/s that HBM allows for RS(10,14)
// (profiles/r1_*: lsu pipe 82 %, shared wavefronts 78 % + global 17 %, zero bank conflicts).
// The only way down is ONE lookup per input byte.  A 256-entry table indexed by the whole byte would
// bank-conflict ~3.5-way on random data -- unless every lane owns its bank: the table is stored
// [x][lane], 32 replicas of each 4-byte entry side by side (16 replicas of 8-byte entries for the
// 64-bit width), so lane l always reads bank l whatever x is.  That costs 32 KiB of shared memory per
// input fragment ("fat8"), or 16 KiB if bit 7 of the byte is peeled off and handled in registers
// ("fat7": index = low 7 bits, plus (sign mask & product-of-0x80) from one PRMT and one LOP3).
// With 227 KiB per SM a block holds e.g. 4 fat8 + 6 fat7 tables for k = 10, or 14 fat7 + 2 nibble pages
// for k = 16; fragments that do not fit keep the split-nibble tables.  One block of 384-768 threads per
// SM builds the replicated tables once (from the same 32-entry global tables the other kernel uses) and
// then streams columns like the first kernel: 128-bit loads into two ping-pong register buffers (plus an
// L2 prefetch one grid stride ahead), XOR accumulate, PRMT transpose, 128-bit stores.
//
// Index arithmetic: address = lane_base + x * 128.  __dp4a(w, 0x80 << 8b, lane_base) selects byte b of
// the data word, scales it by 128 and adds the lane base in ONE fmaheavy-pipe instruction, no mask, no
// shift (fat7 masks the word with 0x7F7F7F7F once per four bytes).  The per-fragment table offset is a
// compile-time immediate of the LDS.
#include "gf_matmul_fat.cuh"

namespace rsgpu {
namespace {

// One column = KB batches, all compile time (every LDS offset is an immediate).  The two register
// buffers alternate per batch ("ping-pong"): the loads of batch t+1 go into the buffer batch t-1 has just
// released and are issued BEFORE batch t is consumed, so they have a whole batch of work to land and no
// register-to-register copies are needed.  SP = which buffer holds batch 0 of this column.
template <int W, int K, int JB, int KB, int B, int SP, bool PF, bool T7, bool PAD, bool XR>
struct ColumnBatches {
    __device__ __forceinline__ static void run(uint4 (&buf0)[JB], uint4 (&buf1)[JB], const uint8_t *&q, size_t in_pitch,
                                               const uint8_t *in, size_t vn, size_t nvec, uint32_t lane_base,
                                               uint32_t tb, Acc<W> &acc, uint32_t pf0, uint32_t pf1, int k)
    {
        constexpr int PAR = (SP + B) & 1;
        uint4 (&cur)[JB] = PAR ? buf1 : buf0;
        uint4 (&nxt)[JB] = PAR ? buf0 : buf1;
        if constexpr (B + 1 < KB) {
            load_batch<JB, PF, PAD, (B + 1) * JB>(nxt, q, in_pitch, pf0, k);
        } else {
            if (vn < nvec) {
                q = in + vn * 16;
                load_batch<JB, PF, PAD, 0>(nxt, q, in_pitch, pf1, k);
            }
        }
        BatchMac<W, K, JB, B, 0, T7, 0, XR>::run(cur, lane_base, tb, acc);
        if constexpr (B + 1 < KB)
            ColumnBatches<W, K, JB, KB, B + 1, SP, PF, T7, PAD, XR>::run(buf0, buf1, q, in_pitch, in, vn, nvec, lane_base, tb, acc, pf0, pf1, k);
    }
};

template <int W, int K, int JB, int KB, int SP, bool PF, bool T7, bool PAD, bool XR>
__device__ __forceinline__ void do_column(uint4 (&buf0)[JB], uint4 (&buf1)[JB], const uint8_t *&q, size_t in_pitch,
                                          const uint8_t *in, uint8_t *out, size_t out_pitch, size_t v, size_t vn,
                                          size_t nvec, uint32_t lane_base, uint32_t tb, int nrows, uint32_t pf_stride,
                                          int k)
{
    // L2 prefetch distance in bytes (one grid stride); 0 when that column lies behind the end of the stripe
    const uint32_t pf0 = vn < nvec ? pf_stride : 0u;
    const uint32_t pf1 = vn + (vn - v) < nvec ? pf_stride : 0u;
    Acc<W> acc;
    acc_clear<XR, T7>(acc);
    ColumnBatches<W, K, JB, KB, 0, SP, PF, T7, PAD, XR>::run(buf0, buf1, q, in_pitch, in, vn, nvec, lane_base, tb, acc, pf0, pf1, k);
    if constexpr (W == 8) {
        transpose_store(acc.a, out, out_pitch, v * 16, 0, nrows);
        transpose_store(acc.b, out, out_pitch, v * 16, 4, nrows - 4);
    } else {
        transpose_store(acc.a, out, out_pitch, v * 16, 0, nrows, (T7 && Map<K>::NF > Map<K>::N8) ? acc.t : nullptr);
    }
    if constexpr (XR) stg_stream(out - out_pitch + v * 16, acc.x);   // the peeled row sits right above the group's rows
}

template <int W, int JB, int KB, bool PF, bool T7, bool PAD, bool XR>
__global__ void __launch_bounds__(fat_threads_x(W, JB, KB, T7, PAD, XR), 1)
k_gf_matmul_fat(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
                const uint32_t *__restrict__ tab, int nrows, size_t nvec, int k)
{
    constexpr int K = JB * KB;
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;

    build_block_tables<W, K, PAD>(tb, tab, k);
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lane_base = tb + (W == 4 ? lane * 4u : (lane & 15u) * 8u);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t pf_stride = (uint32_t)(stride * 16);
    uint4 buf0[JB], buf1[JB];
    const uint8_t *q = in + v * 16;
    if (v < nvec) load_batch<JB, PF, PAD, 0>(buf0, q, in_pitch, v + stride < nvec ? pf_stride : 0u, k);

    while (v < nvec) {
        do_column<W, K, JB, KB, 0, PF, T7, PAD, XR>(buf0, buf1, q, in_pitch, in, out, out_pitch, v, v + stride, nvec, lane_base, tb, nrows, pf_stride, k);
        v += stride;
        if constexpr (KB & 1) {          // odd batch count: the next column starts in the other buffer
            if (v >= nvec) break;
            do_column<W, K, JB, KB, 1, PF, T7, PAD, XR>(buf0, buf1, q, in_pitch, in, out, out_pitch, v, v + stride, nvec, lane_base, tb, nrows, pf_stride, k);
            v += stride;
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
FatKernel make_fat()
{
    // L2 prefetch one stride ahead: on from k = 8.  Same box, same build, off -> on (fraction of the measured
    // HBM peak): RS(8,12) 0.894 -> 0.934, RS(10,14) 0.896 -> 0.991, RS(12,16) 0.866 -> 0.913, RS(14,18) 0.796 ->
    // 0.893, RS(16,20) 0.885 -> 0.928, RS(20,24) 0.824 -> 0.858; but RS(6,9) 0.970 -> 0.906, RS(4,6) 0.99 -> 0.87.
    constexpr bool PF = fat_use_prefetch(JB * KB);
    // transposed bit-7 accumulation (mac_vec): fewer alu instructions, 16 more registers.  It pays for the
    // fat7-heavy shapes with an even batch count (same box, off -> on: RS(14,18) 0.888 -> 0.945, RS(16,20)
    // 0.874 -> 0.915, RS(20,24) 0.848 -> 0.863) and loses where the extra registers spill or cost threads
    // (RS(10,14) 0.930 -> 0.754, RS(12,16) 0.878 -> 0.639).
    constexpr int K = JB * KB;
    constexpr bool T7 = fat_use_t7(W, K);
    return FatKernel{k_gf_matmul_fat<W, JB, KB, PF, T7, PAD, XR>, fat_threads_x(W, JB, KB, T7, PAD, XR), Map<K>::BYTES + 256};
}

template <int W, bool XR = false>
bool pick_fat(int k, FatKernel *fk)
{
    switch (k) {
#define RSGPU_FAT_CASE(k_, jb_, kb_, pad_) case k_: *fk = make_fat<W, jb_, kb_, pad_, XR>(); return true;
        RSGPU_FAT_SHAPES(RSGPU_FAT_CASE)
#undef RSGPU_FAT_CASE
    default: return false;
    }
}

}  // namespace

// Returns cudaErrorNotSupported when no bank-replicated instance exists for this k (the caller then
// uses the split-nibble kernel).  Vector part only: nvec 16-byte columns.
cudaError_t launch_matmul_fat(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                              size_t out_pitch, size_t nvec, const DeviceInfo &dev, int max_blocks, int threads,
                              cudaStream_t s)
{
    FatKernel fk;
    if (g.xor_row >= 0 && g.width != 4) return cudaErrorNotSupported;     // the peeled row exists with 32-bit entries only
    const bool ok = g.xor_row >= 0 ? pick_fat<4, true>(k, &fk) : (g.width == 4 ? pick_fat<4>(k, &fk) : pick_fat<8>(k, &fk));
    if (!ok || fk.smem > dev.max_smem_optin) return cudaErrorNotSupported;
    if (threads > 0 && threads < fk.threads) fk.threads = threads;   // tuning: fewer threads than the launch bound
    int occ = 0;   // always 1: the tables fill the SM's shared memory; the call raises the opt-in attribute once
    cudaError_t e = prepare_launch(dev, (const void *)fk.fn, fk.threads, (size_t)fk.smem, &occ);
    if (e != cudaSuccess) return e;
    size_t want = (nvec + fk.threads - 1) / fk.threads;
    size_t cap = (size_t)dev.sm_count;      // one block per SM: the tables fill its shared memory
    if (max_blocks > 0 && (size_t)max_blocks < cap) cap = max_blocks;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    fk.fn<<<grid, fk.threads, fk.smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, g.nrows, nvec, k);
    return cudaGetLastError();
}

}  // namespace rsgpu
