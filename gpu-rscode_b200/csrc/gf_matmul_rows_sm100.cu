// gf_matmul_rows_sm100.cu -- encode with one destination pointer PER OUTPUT ROW (rsgpu_encode_rows).
//
// The rows may live on other GPUs of the box: with peer access enabled, a unified-address pointer into
// a peer's memory is written by plain 128-bit stores that travel over NVLink / NVSwitch, so the parity
// fragments land on the devices that will keep them while the kernel is still computing the next
// columns -- no staging copy on the source GPU and no separate transfer step (the storage
// node hand-off; the reference gathers everything on the host, src/encode.cu:410-429).
// Same split-nibble multiply-accumulate as gf_matmul_sm100.cu; the peer stores (NVLink, ~770 GB/s per
// direction measured on this pool) bound it long before the lookups do, so the lean kernel is the
// right base and the byte-table machinery is not needed here.
#include "gf_matmul_common.cuh"
#include "rs_internal.h"

namespace rsgpu {
namespace {

struct OutRows {
    uint8_t *r[8];
    uint8_t *x;      // destination of the peeled all-ones row (plain XOR of the inputs), or nullptr
};

__device__ __forceinline__ void transpose_store_rows(const uint32_t *a, const OutRows &rows, size_t off, int row0, int nrows)
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
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i < nrows) stg_stream(rows.r[row0 + i] + off, make_uint4(r[i][0], r[i][1], r[i][2], r[i][3]));
}

template <int W, int JB, bool PAD>
__global__ void __launch_bounds__(256)
k_gf_matmul_rows(const uint8_t *__restrict__ in, size_t in_pitch, OutRows rows, const uint32_t *__restrict__ tab,
                 int npages, int k, int kb, int nrows, size_t nvec, uint32_t hi4)
{
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;
    {
        constexpr int V4 = 32 * W / 16;
        for (int i = threadIdx.x; i < npages * V4; i += blockDim.x) {
            const int j = i / V4, r = i - j * V4;
            const uint4 t = reinterpret_cast<const uint4 *>(tab)[i];
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(tb + j * 256 + r * 16), "r"(t.x), "r"(t.y),
                         "r"(t.z), "r"(t.w) : "memory");
        }
    }
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint4 nxt[JB];
    const uint8_t *q = in + v * 16;
    int j = 0;
    if (v < nvec) load_batch<JB, PAD>(nxt, q, in_pitch, j, k);
    while (v < nvec) {
        Acc<W> acc;
        acc_clear<W>(acc);
        uint4 x = make_uint4(0u, 0u, 0u, 0u);
        const size_t vn = v + stride;
        for (int b = 0; b < kb; ++b) {
            uint4 cur[JB];
#pragma unroll
            for (int jj = 0; jj < JB; ++jj) cur[jj] = nxt[jj];
            if (b + 1 < kb) {
                load_batch<JB, PAD>(nxt, q, in_pitch, j, k);
            } else if (vn < nvec) {
                q = in + vn * 16;
                j = 0;
                load_batch<JB, PAD>(nxt, q, in_pitch, j, k);
            }
            const uint32_t t0 = tb + (uint32_t)(b * JB) * 256u;
#pragma unroll
            for (int jj = 0; jj < JB; ++jj) {
                mac_vec<W, 2>(cur[jj], t0 + jj * 256u, hi4, acc);
                if (rows.x && (!PAD || b * JB + jj < k)) {   // uniform: this kernel is NVLink-bound, a runtime test is free
                    x.x ^= cur[jj].x;
                    x.y ^= cur[jj].y;
                    x.z ^= cur[jj].z;
                    x.w ^= cur[jj].w;
                }
            }
        }
        transpose_store_rows(acc.a, rows, v * 16, 0, nrows);
        if constexpr (W == 8) transpose_store_rows(acc.b, rows, v * 16, 4, nrows - 4);
        if (rows.x) stg_stream(rows.x + v * 16, x);
        v = vn;
    }
}

// byte tail (len % 16) for row pointers
template <int W>
__global__ void __launch_bounds__(256)
k_gf_matmul_rows_bytes(const uint8_t *__restrict__ in, size_t in_pitch, OutRows rows, const uint32_t *__restrict__ tab,
                       int k, int nrows, size_t c0, size_t len)
{
    for (size_t c = c0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < len; c += (size_t)gridDim.x * blockDim.x) {
        uint32_t a = 0, b = 0;
        unsigned xs = 0;
        for (int j = 0; j < k; ++j) {
            const unsigned x = in[(size_t)j * in_pitch + c];
            xs ^= x;
            const uint32_t *tl = tab + (size_t)j * 32 * (W / 4), *th = tl + 16 * (W / 4);
            if (W == 4) {
                a ^= __ldg(tl + (x & 15u)) ^ __ldg(th + (x >> 4));
            } else {
                a ^= __ldg(tl + 2 * (x & 15u)) ^ __ldg(th + 2 * (x >> 4));
                b ^= __ldg(tl + 2 * (x & 15u) + 1) ^ __ldg(th + 2 * (x >> 4) + 1);
            }
        }
        for (int r = 0; r < nrows; ++r) rows.r[r][c] = (uint8_t)((r < 4 ? a : b) >> (8 * (r & 3)));
        if (rows.x) rows.x[c] = (uint8_t)xs;
    }
}

}  // namespace

// Output rows [g.row0, g.row0 + g.nrows) go to out_rows[g.row0 ...]; data and row pointers 16-byte aligned
// for the vector part (the caller checks), any len.
cudaError_t launch_matmul_rows(const RowGroup &g, int k, int kpad, const uint8_t *in, size_t in_pitch,
                               uint8_t *const *out_rows, size_t len, const DeviceInfo &dev, cudaStream_t s,
                               long long *launches)
{
    OutRows rows;
    bool aligned = (((uintptr_t)in | in_pitch) & 15u) == 0;
    for (int i = 0; i < 8; ++i) {
        rows.r[i] = i < g.nrows ? out_rows[g.row0 + i] : nullptr;
        if (i < g.nrows && ((uintptr_t)rows.r[i] & 15u)) aligned = false;
    }
    rows.x = g.xor_row >= 0 ? out_rows[g.xor_row] : nullptr;
    if ((uintptr_t)rows.x & 15u) aligned = false;
    const size_t nvec = aligned ? len / 16 : 0;
    const uint32_t hi4 = 16u * g.width * 0x01010101u;
    if (nvec) {
        const bool pad = (k % 4) != 0;
        const int kb = (k + 3) / 4;
        using Fn = void (*)(const uint8_t *, size_t, OutRows, const uint32_t *, int, int, int, int, size_t, uint32_t);
        Fn fn = g.width == 4 ? (pad ? k_gf_matmul_rows<4, 4, true> : k_gf_matmul_rows<4, 4, false>)
                             : (pad ? k_gf_matmul_rows<8, 4, true> : k_gf_matmul_rows<8, 4, false>);
        const size_t smem = (size_t)kpad * 256 + 256;
        int occ = 0;
        cudaError_t e = prepare_launch(dev, (const void *)fn, 256, smem, &occ);
        if (e != cudaSuccess) return e;
        const size_t want = (nvec + 255) / 256, cap = (size_t)dev.sm_count * occ;
        fn<<<(unsigned)(want < cap ? want : cap), 256, smem, s>>>(in, in_pitch, rows, g.d_tab, kpad, k, kb, g.nrows, nvec, hi4);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) ++*launches;
    }
    if (nvec * 16 < len) {
        const size_t rem = len - nvec * 16;
        size_t blocks = (rem + 255) / 256;
        if (blocks > (size_t)dev.sm_count * 8) blocks = (size_t)dev.sm_count * 8;
        if (g.width == 4)
            k_gf_matmul_rows_bytes<4><<<(unsigned)blocks, 256, 0, s>>>(in, in_pitch, rows, g.d_tab, k, g.nrows, nvec * 16, len);
        else
            k_gf_matmul_rows_bytes<8><<<(unsigned)blocks, 256, 0, s>>>(in, in_pitch, rows, g.d_tab, k, g.nrows, nvec * 16, len);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
    }
    return cudaSuccess;
}

}  // namespace rsgpu
