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

    for (int c = 0; c < n; ++c) {
        if (tid == 0) s_piv = INT_MAX;
        __syncthreads();
        for (int r = c + tid; r < n; r += nt)
            if (A[r * n + c]) atomicMin(&s_piv, r);
        __syncthreads();
        const int piv = s_piv;
        if (piv == INT_MAX) {          // uniform: every thread read the same value
            if (tid == 0) *flag = 1;
            return;
        }
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
    for (int i = tid; i < n * n; i += nt) inv[i] = R[i];
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
