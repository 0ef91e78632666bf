// gj_inverse.cu -- k x k Gauss-Jordan inversion over GF(2^8) as ONE thread block.
//
// Replaces the reference's CPU_invert_matrix (src/cpu-decode.c:251-298, what its binary really
// calls, src/decode.cu:333) and its dead GPU_invert_matrix (src/matrix.cu:666-744: a host loop with
// one D2H copy, up to three kernel launches and three cudaDeviceSynchronize per row, and a
// cross-block race in eliminate_by_row, src/matrix.cu:525-556).  Here the augmented matrix [A | R]
// lives in shared memory for the whole inversion, the field tables sit beside it, and the only
// global traffic is one read of A and one write of R.
//
// Pivoting is by ROWS (first row >= c with a non-zero in column c), which is a valid elimination
// step; the reference swaps COLUMNS, with a typo that drops the swap on the inverse side
// (src/cpu-decode.c:121-137), so it is wrong whenever a swap is needed.  The inverse of a matrix is
// unique, so wherever the reference is right this kernel returns the same bytes.  A column with no
// pivot sets *flag = 1 (singular) -- the reference would index column -1 there.
#include <limits.h>

#include "gf256.cuh"
#include "rs_internal.h"

namespace rsgpu {

// The elimination itself, shared by both kernels below.  A (n x n, the matrix) and R (n x n, starts as the
// identity) are in shared memory, ex/lg the field tables, fac n bytes of staging, s_piv one shared int.
// Returns false (for every thread) when a column has no pivot; on success R is the inverse and A the identity.
// Five block barriers per column, no global traffic.
__device__ __forceinline__ bool gauss_jordan_smem(uint8_t *A, uint8_t *R, int n, const uint8_t *ex, const uint8_t *lg,
                                                  uint8_t *fac, int *s_piv)
{
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int c = 0; c < n; ++c) {
        if (tid == 0) *s_piv = INT_MAX;
        __syncthreads();
        for (int r = c + tid; r < n; r += nt)
            if (A[r * n + c]) atomicMin(s_piv, r);
        __syncthreads();
        const int piv = *s_piv;
        if (piv == INT_MAX) return false;   // uniform: every thread read the same value
        if (piv != c) {
            for (int col = tid; col < 2 * n; col += nt) {
                uint8_t *M = col < n ? A : R;
                const int cc = col < n ? col : col - n;
                const uint8_t t = M[c * n + cc];
                M[c * n + cc] = M[piv * n + cc];
                M[piv * n + cc] = t;
            }
            __syncthreads();
        }
        const unsigned pinv = gf_inv(ex, lg, A[c * n + c]);
        __syncthreads();               // everyone holds pinv before row c is rewritten
        for (int col = tid; col < 2 * n; col += nt) {
            uint8_t *M = col < n ? A : R;
            const int cc = col < n ? col : col - n;
            M[c * n + cc] = (uint8_t)gf_mul(ex, lg, M[c * n + cc], pinv);
        }
        // column c of the other rows is untouched by the normalisation: stage the factors now
        for (int q = tid; q < n; q += nt) fac[q] = (q == c) ? 0 : A[q * n + c];
        __syncthreads();
        for (int idx = tid; idx < 2 * n * n; idx += nt) {
            const int q = idx / (2 * n), col = idx - q * 2 * n;
            const unsigned f = fac[q];
            if (f) {
                uint8_t *M = col < n ? A : R;
                const int cc = col < n ? col : col - n;
                M[q * n + cc] ^= (uint8_t)gf_mul(ex, lg, f, M[c * n + cc]);
            }
        }
        __syncthreads();
    }
    return true;
}

__global__ void __launch_bounds__(1024, 1)
k_gj_invert(const uint8_t *__restrict__ in, uint8_t *__restrict__ inv, int n, int *flag,
            const uint8_t *__restrict__ gexp, const uint8_t *__restrict__ glog)
{
    extern __shared__ uint8_t sm[];
    uint8_t *A = sm;                  // n*n
    uint8_t *R = A + n * n;           // n*n
    uint8_t *ex = R + n * n;          // 512
    uint8_t *lg = ex + kExpSize;      // 256
    uint8_t *fac = lg + kLogSize;     // n : elimination factor of each row for the current column
    __shared__ int s_piv;

    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < kExpSize; i += nt) ex[i] = gexp[i];
    for (int i = tid; i < kLogSize; i += nt) lg[i] = glog[i];
    for (int i = tid; i < n * n; i += nt) {
        A[i] = in[i];
        R[i] = (uint8_t)((i / n) == (i % n));
    }
    if (tid == 0) *flag = 0;
    __syncthreads();

    if (!gauss_jordan_smem(A, R, n, ex, lg, fac, &s_piv)) {
        if (tid == 0) *flag = 1;
        return;
    }
    for (int i = tid; i < n * n; i += nt) inv[i] = R[i];
}

// Everything rsgpu_set_survivors needs, in ONE launch of one block: gather the survivor rows of the total
// matrix, invert (as k_gj_invert), write the inverse, build the split-nibble product tables of every row group
// of the full k x k plan, pick the rows of the erased natives and build their tables.  Was six launches
// (gather, invert, 1-2 x tables, gather, tables); the index lists travel as kernel parameters instead of two
// host-to-device copies.  Table layout = k_build_product_tables (gf_tables.cu): groups of <= 8 rows back to
// back, group g holding kpad x 32 entries of 4 (<= 4 rows) or 8 bytes.
struct SetupLists {
    uint16_t survivors[256];   // fragment index of conf line i
    uint8_t erased[256];       // natives to rebuild, ascending
};

__device__ __forceinline__ void build_tables_from_smem(const uint8_t *M, int rows, int k, int kpad, uint32_t *tabs,
                                                        const uint8_t *ex, const uint8_t *lg)
{
    // M: rows x k in shared memory.  Entry index e = (j, half, n) as in k_build_product_tables.
    size_t off = 0;
    for (int row0 = 0; row0 < rows; row0 += 8) {
        const int nrows = rows - row0 < 8 ? rows - row0 : 8;
        const int words = nrows > 4 ? 2 : 1;
        for (int e = threadIdx.x; e < kpad * 32; e += blockDim.x) {
            const int n = e & 15, half = (e >> 4) & 1, j = e >> 5;
            const unsigned x = half ? (unsigned)n << 4 : (unsigned)n;
            for (int q = 0; q < words; ++q) {
                uint32_t w = 0;
                for (int b = 0; b < 4; ++b) {
                    const int r = q * 4 + b;
                    unsigned c = 0;
                    if (r < nrows && j < k) c = M[(row0 + r) * k + j];
                    w |= gf_mul(ex, lg, c, x) << (8 * b);
                }
                tabs[off + (size_t)e * words + q] = w;
            }
        }
        off += (size_t)kpad * 32 * words;
    }
}

__global__ void __launch_bounds__(1024, 1)
k_decode_setup(const uint8_t *__restrict__ total, int k, int kpad, SetupLists lists, int n_erased,
               uint8_t *__restrict__ inv, uint32_t *__restrict__ full_tabs, uint8_t *__restrict__ erased_matrix,
               uint32_t *__restrict__ erased_tabs, int *flag, const uint8_t *__restrict__ gexp,
               const uint8_t *__restrict__ glog)
{
    extern __shared__ uint8_t sm[];
    const int n = k;
    uint8_t *A = sm;                  // n*n
    uint8_t *R = A + n * n;           // n*n
    uint8_t *ex = R + n * n;          // 512
    uint8_t *lg = ex + kExpSize;      // 256
    uint8_t *fac = lg + kLogSize;     // n
    __shared__ int s_piv;

    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < kExpSize; i += nt) ex[i] = gexp[i];
    for (int i = tid; i < kLogSize; i += nt) lg[i] = glog[i];
    for (int i = tid; i < n * n; i += nt) {
        const int r = i / n, c = i - r * n;
        A[i] = total[(size_t)lists.survivors[r] * n + c];       // copy_matrix of the reference, src/decode.cu:307
        R[i] = (uint8_t)(r == c);
    }
    if (tid == 0) *flag = 0;
    __syncthreads();

    if (!gauss_jordan_smem(A, R, n, ex, lg, fac, &s_piv)) {
        if (tid == 0) *flag = 1;
        return;
    }
    // R = inverse.  A is free now: reuse it for the rows of the erased natives.
    for (int i = tid; i < n * n; i += nt) inv[i] = R[i];
    for (int i = tid; i < n_erased * n; i += nt) {
        const int r = i / n, c = i - r * n;
        const uint8_t v = R[lists.erased[r] * n + c];
        A[i] = v;
        erased_matrix[i] = v;
    }
    __syncthreads();
    if (full_tabs) build_tables_from_smem(R, n, n, kpad, full_tabs, ex, lg);   // nullptr: built later, on first use
    if (n_erased > 0) build_tables_from_smem(A, n_erased, n, kpad, erased_tabs, ex, lg);
}

cudaError_t launch_decode_setup(const uint8_t *d_total, int k, int kpad, const int *survivors, const int *erased, int n_erased,
                                uint8_t *d_inv, uint32_t *d_full_tabs, uint8_t *d_erased_matrix, uint32_t *d_erased_tabs,
                                int *d_flag, const GfTables &gf, cudaStream_t s)
{
    SetupLists lists;
    for (int i = 0; i < 256; ++i) {
        lists.survivors[i] = (uint16_t)(i < k ? survivors[i] : 0);
        lists.erased[i] = (uint8_t)(i < n_erased ? erased[i] : 0);
    }
    const size_t smem = (size_t)2 * k * k + kExpSize + kLogSize + k;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_decode_setup, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
        if (e != cudaSuccess) return e;
    }
    const int threads = 2 * k * k >= 4096 ? 1024 : 256;
    k_decode_setup<<<1, threads, smem, s>>>(d_total, k, kpad, lists, n_erased, d_inv, d_full_tabs, d_erased_matrix, d_erased_tabs,
                                            d_flag, gf.d_exp, gf.d_log);
    return cudaGetLastError();
}

cudaError_t launch_gj_invert(const uint8_t *d_in, uint8_t *d_inv, int n, int *d_flag,
                             const GfTables &gf, cudaStream_t s)
{
    const size_t smem = (size_t)2 * n * n + kExpSize + kLogSize + n;
    if (smem > 48 * 1024) {   // per-device attribute, so set it on every large launch (cheap, idempotent)
        cudaError_t e = cudaFuncSetAttribute(k_gj_invert, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
        if (e != cudaSuccess) return e;
    }
    int threads = 2 * n * n >= 4096 ? 1024 : (2 * n * n >= 512 ? 256 : 64);
    k_gj_invert<<<1, threads, smem, s>>>(d_in, d_inv, n, d_flag, gf.d_exp, gf.d_log);
    return cudaGetLastError();
}

}  // namespace rsgpu
