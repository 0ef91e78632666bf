// membench.cu -- developer micro-benchmark: the memory access pattern of the RS multiply kernel
// (k fragment-major input streams, p output streams, 16 bytes per thread per row, grid-stride) with
// trivial compute (XOR), to find the HBM ceiling of the pattern itself.  Build: nvcc -O3 -arch=sm_100a.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
__device__ __forceinline__ uint4 ldg_stream(const unsigned char *p) {
    uint4 v; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)); return v; }
__device__ __forceinline__ void stg_stream(unsigned char *p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
template <int K, int P, int DEPTH>
__global__ void pat(const unsigned char *in, size_t pitch, unsigned char *out, size_t opitch, size_t nvec, int blockmajor) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t vend = nvec;
    if (blockmajor) { size_t per = (nvec / gridDim.x) / blockDim.x * blockDim.x; v = blockIdx.x * per + threadIdx.x; vend = (blockIdx.x == gridDim.x - 1) ? nvec : (blockIdx.x + 1) * per; stride = blockDim.x; }
    uint4 buf[DEPTH][K];
    #pragma unroll
    for (int d = 0; d < DEPTH; ++d) { size_t vv = v + d * stride; if (vv < vend) { for (int j = 0; j < K; ++j) buf[d][j] = ldg_stream(in + j * pitch + vv * 16); } }
    while (v < vend) {
        uint4 a = make_uint4(0, 0, 0, 0);
        #pragma unroll
        for (int j = 0; j < K; ++j) { a.x ^= buf[0][j].x; a.y ^= buf[0][j].y; a.z ^= buf[0][j].z; a.w ^= buf[0][j].w; }
        #pragma unroll
        for (int d = 0; d + 1 < DEPTH; ++d) for (int j = 0; j < K; ++j) buf[d][j] = buf[d + 1][j];
        size_t vv = v + DEPTH * stride;
        if (vv < vend) { for (int j = 0; j < K; ++j) buf[DEPTH - 1][j] = ldg_stream(in + j * pitch + vv * 16); }
        #pragma unroll
        for (int i = 0; i < P; ++i) { a.x += i; stg_stream(out + i * opitch + v * 16, a); }
        v += stride;
    }
}
template <int K, int P, int DEPTH>
void run(const char *name, size_t L, int threads, int blocks, int blockmajor, int reps = 10) {
    unsigned char *in, *out; cudaMalloc(&in, (size_t)K * L); cudaMalloc(&out, (size_t)P * L);
    cudaMemset(in, 1, (size_t)K * L);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) pat<K, P, DEPTH><<<blocks, threads>>>(in, L, out, L, L / 16, blockmajor);
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) pat<K, P, DEPTH><<<blocks, threads>>>(in, L, out, L, L / 16, blockmajor);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= reps;
    cudaError_t e = cudaGetLastError();
    printf("%s K=%d P=%d depth=%d threads=%d blocks=%d blockmajor=%d: %.3f ms  traffic %.0f GB/s %s\n", name, K, P, DEPTH, threads, blocks, blockmajor, ms,
           (double)(K + P) * L / ms / 1e6, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(in); cudaFree(out);
}
int main(int argc, char **argv) {
    if (argc > 1) {   // sustained mode: ./membench <reps>  -- only the RS(10,14) pattern, for power readings
        run<10, 4, 1>("rs1014", (size_t)400 << 20, 768, 148, 0, atoi(argv[1]));
        return 0;
    }
    size_t L10 = (size_t)400 << 20, L4 = (size_t)512 << 20, L16 = (size_t)256 << 20;
    for (int bm = 0; bm < 2; ++bm) {
        run<4, 2, 1>("rs46", L4, 768, 148, bm);
        run<10, 4, 1>("rs1014", L10, 768, 148, bm);
        run<10, 4, 1>("rs1014", L10, 256, 148 * 3, bm);
        run<10, 4, 1>("rs1014", L10, 512, 148 * 2, bm);
        run<10, 4, 2>("rs1014", L10, 384, 148, bm);
        run<10, 1, 1>("rs1011", L10, 768, 148, bm);
        run<16, 4, 1>("rs1620", L16, 512, 148, bm);
        run<16, 4, 1>("rs1620", L16, 256, 148 * 4, bm);
        run<5, 2, 1>("rs57", L10, 768, 148, bm);
        run<5, 2, 2>("rs57", L10, 768, 148, bm);
    }
    return 0;
}
