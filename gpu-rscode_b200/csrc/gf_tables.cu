// gf_tables.cu -- small set-up kernels: field tables, generator matrix, survivor gather, product tables.
//
// None of these is on the bandwidth path; they run once per context / per erasure pattern.
#include "gf256.cuh"
#include "rs_internal.h"

namespace rsgpu {

// exp/log for GF(2^8).  Replaces setup_tables (reference src/matrix.cu:47-75), which every thread of
// every block re-ran; here one warp builds them once per context.  Lane l walks its own 8-element
// run of the power sequence after jumping to x^(8l) by repeated doubling, so the serial chain is
// 8*31 + 8 steps at worst instead of 255 table-dependent steps -- and it is table-free.
__global__ void k_build_gf_tables(uint8_t *exp, uint8_t *log)
{
    const int lane = threadIdx.x;   // 32 threads
    unsigned v = 1;
    for (int s = 0; s < lane * 8; ++s) {
        v <<= 1;
        if (v & 0x100u) v ^= kPoly;
    }
    for (int t = 0; t < 8; ++t) {
        const int e = lane * 8 + t;
        if (e < 255) {
            exp[e] = (uint8_t)v;
            exp[e + 255] = (uint8_t)v;
            log[v] = (uint8_t)e;
        }
        v <<= 1;
        if (v & 0x100u) v ^= kPoly;
    }
    if (lane == 0) {
        exp[510] = 0;
        exp[511] = 0;
        log[0] = 0;   // reference convention: gf_pow(0, e) reads log[0] = 0 (src/cpu-rs.c:40,160)
    }
}

cudaError_t launch_build_gf_tables(uint8_t *d_exp, uint8_t *d_log, cudaStream_t s)
{
    k_build_gf_tables<<<1, 32, 0, s>>>(d_exp, d_log);
    return cudaGetLastError();
}

// Total encoding matrix [I_k ; V], V[i][j] = (j+1)^i.  Replaces gen_encoding_matrix (reference
// src/matrix.cu:752-759: no bounds check, tables rebuilt by every thread) plus the host-side identity
// rows written by write_metadata (src/encode.cu:78-88).
__global__ void k_gen_total_matrix(uint8_t *total, int k, int p, const uint8_t *exp, const uint8_t *log)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (k + p) * k) return;
    const int r = idx / k, j = idx - r * k;
    total[idx] = (r < k) ? (uint8_t)(r == j) : (uint8_t)gf_pow(exp, log, (unsigned)(j + 1) & 255u, (unsigned)(r - k));
}

cudaError_t launch_gen_total_matrix(uint8_t *d_total, int k, int p, const GfTables &gf, cudaStream_t s)
{
    const int n = (k + p) * k;
    k_gen_total_matrix<<<(n + 255) / 256, 256, 0, s>>>(d_total, k, p, gf.d_exp, gf.d_log);
    return cudaGetLastError();
}

// Cauchy generator (not in the reference): [I_k ; C], C[i][j] = 1 / (x_i + y_j) with
// x_i = k + i and y_j = j -- all distinct for k + p <= 256, so every k x k sub-matrix of [I ; C] is
// invertible (MDS), which [I ; V] is not.
__global__ void k_gen_total_matrix_cauchy(uint8_t *total, int k, int p, const uint8_t *exp, const uint8_t *log)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (k + p) * k) return;
    const int r = idx / k, j = idx - r * k;
    total[idx] = (r < k) ? (uint8_t)(r == j) : (uint8_t)gf_inv(exp, log, (unsigned)(r ^ j));   // r = k + i
}

cudaError_t launch_gen_total_matrix_cauchy(uint8_t *d_total, int k, int p, const GfTables &gf, cudaStream_t s)
{
    const int n = (k + p) * k;
    k_gen_total_matrix_cauchy<<<(n + 255) / 256, 256, 0, s>>>(d_total, k, p, gf.d_exp, gf.d_log);
    return cudaGetLastError();
}

// C (rows x k) = A[sel[0..rows)] (rows of an (n x k) matrix) * B (k x k): the repair matrix
// T[wanted] * inverse(T[survivors]).  Tiny (<= 255 x 255 x 255), one thread per output entry.
__global__ void k_rows_times_matrix(const uint8_t *A, const int *sel, int rows, const uint8_t *B, int k, uint8_t *C,
                                    const uint8_t *exp, const uint8_t *log)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * k) return;
    const int i = idx / k, j = idx - i * k;
    const uint8_t *a = A + (size_t)sel[i] * k;
    unsigned acc = 0;
    for (int t = 0; t < k; ++t) acc ^= gf_mul(exp, log, a[t], B[(size_t)t * k + j]);
    C[idx] = (uint8_t)acc;
}

cudaError_t launch_rows_times_matrix(const uint8_t *d_A, const int *d_sel, int rows, const uint8_t *d_B, int k,
                                     uint8_t *d_C, const GfTables &gf, cudaStream_t s)
{
    const int n = rows * k;
    k_rows_times_matrix<<<(n + 127) / 128, 128, 0, s>>>(d_A, d_sel, rows, d_B, k, d_C, gf.d_exp, gf.d_log);
    return cudaGetLastError();
}

// Row i of the k x k decode input = row rows[i] of the total matrix (reference copy_matrix,
// src/decode.cu:75-81,307).
__global__ void k_gather_rows(const uint8_t *total, int k, const int *rows, int nrows, uint8_t *out)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nrows * k) return;
    const int i = idx / k, j = idx - i * k;
    out[idx] = total[(size_t)rows[i] * k + j];
}

cudaError_t launch_gather_rows(const uint8_t *d_total, int k, const int *d_rows, int nrows,
                               uint8_t *d_out, cudaStream_t s)
{
    const int n = nrows * k;
    k_gather_rows<<<(n + 255) / 256, 256, 0, s>>>(d_total, k, d_rows, nrows, d_out);
    return cudaGetLastError();
}

// Split-nibble product tables for rows [row0, row0+nrows) of a (rows x k) matrix.
// Entry (j, half, n), `width` bytes: byte r = M[row0+r][j] * (half ? n<<4 : n); rows beyond nrows and
// fragments j >= k give 0 (the multiply kernel may run its inner loop up to kpad).
__global__ void k_build_product_tables(const uint8_t *M, int k, int kpad, int row0, int nrows, int width,
                                       uint32_t *tab, const uint8_t *exp, const uint8_t *log)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;   // entry index
    if (e >= kpad * 32) return;
    const int n = e & 15, half = (e >> 4) & 1, j = e >> 5;
    const unsigned x = half ? (unsigned)n << 4 : (unsigned)n;
    const int words = width / 4;
    for (int q = 0; q < words; ++q) {
        uint32_t w = 0;
        for (int b = 0; b < 4; ++b) {
            const int r = q * 4 + b;
            unsigned c = 0;
            if (r < nrows && j < k) c = M[(size_t)(row0 + r) * k + j];
            w |= gf_mul(exp, log, c, x) << (8 * b);
        }
        tab[(size_t)e * words + q] = w;
    }
}

cudaError_t launch_build_product_tables(const uint8_t *d_matrix, int k, int kpad, int row0, int nrows,
                                        int width, uint32_t *d_tab, const GfTables &gf, cudaStream_t s)
{
    const int n = kpad * 32;
    k_build_product_tables<<<(n + 127) / 128, 128, 0, s>>>(d_matrix, k, kpad, row0, nrows, width, d_tab,
                                                           gf.d_exp, gf.d_log);
    return cudaGetLastError();
}

}  // namespace rsgpu
