// gf_matmul_tma_sm100.cu -- the split-nibble kernel with TMA staging of the fragment tiles.
//
// north_star asks for "TMA staging of fragment tiles"; this is that design, built and measured so the
// choice between it and the register-fed kernels rests on numbers:
//   * a producer warp streams tiles with cp.async.bulk (the TMA engine's 1-D bulk copy,
//     global -> shared, completion counted on an mbarrier): one 4 KiB row segment per input fragment,
//     JB fragments per stage, NST stages in a ring of full/empty mbarriers;
//   * 256 consumer threads wait on the stage's "full" barrier, read their 16 bytes per fragment from the
//     staged tile (LDS.128, conflict-free), run the same split-nibble multiply-accumulate as
//     gf_matmul_sm100.cu, arrive on the stage's "empty" barrier, and at the end of a column tile do the
//     PRMT transpose and the 128-bit stores.
// No registers are held while data is in flight and the copy engine runs ahead by NST-1 stages; the
// cost is that every input byte is written to and read from shared memory once more -- through the same
// L1/shared data pipe that the table lookups already saturate.
#include "gf_matmul_common.cuh"
#include "rs_internal.h"

namespace rsgpu {
namespace {

constexpr int kTileVec = 256;                  // 16-byte columns per tile = consumer threads
constexpr int kTileBytes = kTileVec * 16;      // bytes per fragment row segment in a stage

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
// TMA 1-D bulk copy global -> shared, completion bytes counted on `bar`
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const uint8_t *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// Requires k == kb * JB (no padded fragments).  Block = 256 consumers + 1 producer warp.
template <int W, int JB, int NST>
__global__ void __launch_bounds__(kTileVec + 32)
k_gf_matmul_tma(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
                const uint32_t *__restrict__ tab, int npages, int kb, int nrows, size_t nvec, uint32_t hi4)
{
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;   // nibble pages
    const uint32_t sb = tb + (uint32_t)npages * 256u;                                  // stage ring
    const uint32_t bar_full = sb + NST * JB * kTileBytes;                              // NST x 8 bytes
    const uint32_t bar_empty = bar_full + NST * 8;
    const int tid = threadIdx.x;
    {
        constexpr int V4 = 32 * W / 16;
        for (int i = tid; i < npages * V4; i += blockDim.x) {
            const int j = i / V4, r = i - j * V4;
            const uint4 t = reinterpret_cast<const uint4 *>(tab)[i];
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(tb + j * 256 + r * 16), "r"(t.x), "r"(t.y),
                         "r"(t.z), "r"(t.w) : "memory");
        }
    }
    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(bar_full + s * 8, 1);            // the producer's arrive.expect_tx
            mbar_init(bar_empty + s * 8, kTileVec);    // every consumer thread releases the stage
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const size_t ntiles = (nvec + kTileVec - 1) / kTileVec;
    int stage = 0;
    uint32_t phase = 0;
    if (tid >= kTileVec) {
        // ---------------- producer warp: one elected lane drives the copy engine ----------------
        if (tid == kTileVec) {
            for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const size_t v0 = tile * kTileVec;
                const uint32_t nv = (uint32_t)(nvec - v0 < (size_t)kTileVec ? nvec - v0 : (size_t)kTileVec);
                const uint32_t bytes = nv * 16u;
                const uint8_t *src = in + v0 * 16;
                for (int b = 0; b < kb; ++b) {
                    mbar_wait(bar_empty + stage * 8, phase ^ 1u);      // stage released by all consumers
                    mbar_expect_tx(bar_full + stage * 8, JB * bytes);
#pragma unroll
                    for (int jj = 0; jj < JB; ++jj) {
                        tma_load_1d(sb + (stage * JB + jj) * kTileBytes, src, bytes, bar_full + stage * 8);
                        src += in_pitch;
                    }
                    if (++stage == NST) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else {
        // ---------------- consumers ----------------
        for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const size_t v = tile * kTileVec + tid;
            const bool active = v < nvec;
            Acc<W> acc;
            acc_clear<W>(acc);
            for (int b = 0; b < kb; ++b) {
                mbar_wait(bar_full + stage * 8, phase);                // the stage's bytes have landed
                if (active) {
                    const uint32_t t0 = tb + (uint32_t)(b * JB) * 256u;
                    const uint32_t d0 = sb + (stage * JB) * kTileBytes + tid * 16;
#pragma unroll
                    for (int jj = 0; jj < JB; ++jj)
                        mac_vec<W, 2>(lds128(d0 + jj * kTileBytes), t0 + jj * 256u, hi4, acc);
                }
                mbar_arrive(bar_empty + stage * 8);
                if (++stage == NST) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
            if (active) {
                transpose_store(acc.a, out, out_pitch, v * 16, 0, nrows);
                if constexpr (W == 8) transpose_store(acc.b, out, out_pitch, v * 16, 4, nrows - 4);
            }
        }
    }
}

using TmaFn = void (*)(const uint8_t *, size_t, uint8_t *, size_t, const uint32_t *, int, int, int, size_t, uint32_t);

template <int W>
TmaFn pick_tma(int jb)
{
    switch (jb) {
    case 4: return k_gf_matmul_tma<W, 4, 4>;
    case 5: return k_gf_matmul_tma<W, 5, 4>;
    case 8: return k_gf_matmul_tma<W, 8, 3>;
    default: return nullptr;
    }
}

}  // namespace

// cudaErrorNotSupported when k has no divisor among the compiled batch sizes (4, 5, 8).
cudaError_t launch_matmul_tma(const RowGroup &g, int k, int kpad, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                              size_t out_pitch, size_t nvec, const DeviceInfo &dev, int max_blocks, cudaStream_t s)
{
    int jb = 0;
    for (int c : {8, 5, 4})
        if (k % c == 0) {
            jb = c;
            break;
        }
    TmaFn fn = jb ? (g.width == 4 ? pick_tma<4>(jb) : pick_tma<8>(jb)) : nullptr;
    if (!fn || g.xor_row >= 0) return cudaErrorNotSupported;    // no peeled-row instances here: the register-fed kernel has them
    const int nst = jb == 8 ? 3 : 4;
    const size_t smem = 256 + (size_t)kpad * 256 + (size_t)nst * jb * kTileBytes + 2 * nst * 8;
    if ((int)smem > dev.max_smem_optin) return cudaErrorNotSupported;
    int occ = 0;
    cudaError_t e = prepare_launch(dev, (const void *)fn, kTileVec + 32, smem, &occ);
    if (e != cudaSuccess) return e;
    const size_t ntiles = (nvec + kTileVec - 1) / kTileVec;
    size_t cap = (size_t)dev.sm_count * occ;
    if (max_blocks > 0 && (size_t)max_blocks < cap) cap = max_blocks;
    const unsigned grid = (unsigned)(ntiles < cap ? ntiles : cap);
    fn<<<grid, kTileVec + 32, smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, kpad, k / jb, g.nrows, nvec,
                                        16u * g.width * 0x01010101u);
    return cudaGetLastError();
}

}  // namespace rsgpu
