// ref_gpu_shim.cu -- C entry points onto the REFERENCE's own CUDA launchers (test/bench infrastructure).
//
// oracle/Makefile compiles this file together with the reference's unmodified src/matrix.cu (from where
// it lies under /root/reference; no reference source is copied here) for sm_100a into
// oracle/_ref/libgpurs_ref.so.  It gives two things ("optional same-box GPU
// baseline"):
//   * a second, independent checker on the GPU box: the reference's kernels and ours must write the same
//     bytes (tests/test_gpu_vs_reference_kernels.py);
//   * the recompiled reference kernel's own event-timed rate beside ours (tests/tools/ref_gpu_bench.py).
// Never linked into the product; only tests/ loads it.
//
// The reference declares encode_chunk / decode_chunk with C++ linkage (src/matrix.h:80,97), hence this
// shim.  Their sizes are `int`: k * chunk must stay below 2 GiB.
#include <cuda_runtime.h>
#include <stdint.h>
#include "matrix.h"

extern "C" {

// src/matrix.cu:767 -- returns the reference's own event-timed kernel milliseconds
float refgpu_encode_chunk(unsigned char *data_d, unsigned char *coeff_d, unsigned char *code_d, int k, int p,
                          int chunk, int grid_x)
{
    return encode_chunk(data_d, coeff_d, code_d, k, p, chunk, grid_x, 0);
}

// src/matrix.cu:838 -- data_d receives all k natives, code_d holds the k survivors in conf order
float refgpu_decode_chunk(unsigned char *data_d, unsigned char *inv_d, unsigned char *code_d, int k, int p,
                          int chunk, int grid_x)
{
    return decode_chunk(data_d, inv_d, code_d, k, p, chunk, grid_x, 0);
}

// src/matrix.cu:752 -- launched the way src/encode.cu:134-141 does (block (min(p,16), min(k,16)))
int refgpu_gen_encoding_matrix(uint8_t *matrix_d, int p, int k)
{
    dim3 block(p < 16 ? p : 16, k < 16 ? k : 16);
    dim3 grid((p + block.x - 1) / block.x, (k + block.y - 1) / block.y);
    gen_encoding_matrix<<<grid, block>>>(matrix_d, p, k);
    return (int)cudaDeviceSynchronize();
}

}  // extern "C"
