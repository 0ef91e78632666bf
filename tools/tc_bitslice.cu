// tc_bitslice.cu -- the tensor-core question, measured (developer tool, GPU box; not part of the product).
//
// north_star: "Tensor cores are used only if a bit-sliced GF(2) formulation is shown by ncu to beat the XOR path".
// This is that formulation for the code where the table kernel is weakest, RS(32,40) (eight parity rows, 64-bit
// table entries, shared-memory bound at 0.46 of the HBM peak): over GF(2) the generator is a 64 x 256 bit matrix
// M (bit b of parity row t = XOR over fragments j and bits c of M[8t+b][8j+c] & bit c of the data byte), so
// parity bits = (M x data bits) mod 2 -- an integer GEMM whose low result bit is kept.  sm_100a has no 1-bit MMA
// (SURVEY H2), so the bits travel as u8 0/1 operands of mma.sync.m16n8k32 (s32 accumulate).
//
//   A (16 x 32, per k-tile)  = data: rows = 16 byte columns, k = 4 fragments x 8 bits   (expanded from bytes in registers)
//   B (32 x 8,  per k/n tile) = M^T: k as above, n = the 8 bits of one parity row       (constant, read from shared memory)
//   D (16 x 8) s32            = how many terms are set; bit 0 is the parity bit        (repacked to bytes, quad shuffle)
//
// Per warp step: 32 columns x 32 fragments = 1024 input bytes -> 16 coalesced 32-bit loads per thread, 64 nibble->4x(0/1)
// expansions (IMAD 0x00204081 + AND), 128 MMAs (2 m-tiles x 8 k-tiles x 8 n-tiles), 64 LDS.64 of B fragments, repack of
// 64 accumulators into 8 words, 16 SHFL, 2 stores.  512 multiply-accumulates per input byte.
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o tools/tc_bitslice tools/tc_bitslice.cu
//   tools/tc_bitslice [frag_mib=64] [reps=10]
// Modes timed: 0 = the whole kernel (checked byte for byte against a plain GF(2^8) kernel on the same buffers),
// 1 = MMAs only (operands constant: the tensor-pipe ceiling of this formulation), 2 = everything but the MMAs.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(1);                                                                           \
        }                                                                                      \
    } while (0)

constexpr int K = 32, P = 8;

static unsigned gf_mul_h(unsigned a, unsigned b)
{
    unsigned r = 0;
    for (int i = 0; i < 8; ++i) {
        if (b & 1) r ^= a;
        b >>= 1;
        a <<= 1;
        if (a & 0x100) a ^= 0x11D;
    }
    return r & 0xFF;
}

__device__ __forceinline__ unsigned gf_mul_d(unsigned a, unsigned b)
{
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (b & 1) r ^= a;
        b >>= 1;
        a <<= 1;
        if (a & 0x100) a ^= 0x11D;
    }
    return r & 0xFF;
}

// plain checker: one thread per byte column
__global__ void k_plain(const uint8_t *in, size_t pitch, uint8_t *out, size_t opitch, const uint8_t *G, size_t len)
{
    for (size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < len; c += (size_t)gridDim.x * blockDim.x) {
        unsigned acc[P] = {0};
        for (int j = 0; j < K; ++j) {
            const unsigned x = in[(size_t)j * pitch + c];
            for (int t = 0; t < P; ++t) acc[t] ^= gf_mul_d(G[t * K + j], x);
        }
        for (int t = 0; t < P; ++t) out[(size_t)t * opitch + c] = (uint8_t)acc[t];
    }
}

__device__ __forceinline__ void mma_u8(int (&d)[4], const uint32_t (&a)[4], const uint2 &b)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

// nibble -> four bytes of 0/1 (bit i in byte i)
__device__ __forceinline__ uint32_t expand4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

// bfrag: [kt 8][t 8][lane 32] uint2 -- the B fragments of M^T in mma layout
template <int MODE>
__global__ void __launch_bounds__(256, 2)
k_bitslice(const uint8_t *__restrict__ in, size_t pitch, uint8_t *__restrict__ out, size_t opitch, const uint2 *__restrict__ bfrag,
           size_t ntiles)
{
    __shared__ uint2 sB[8 * 8 * 32];
    for (int i = threadIdx.x; i < 8 * 8 * 32; i += blockDim.x) sB[i] = bfrag[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 2, tg = lane & 3;
    const size_t wtotal = (size_t)gridDim.x * (blockDim.x >> 5);
    const int f0 = tg >> 1, sh = (tg & 1) * 4;      // fragment within the k-tile (and +2), nibble of the byte
    for (size_t tile = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); tile < ntiles; tile += wtotal) {
        const size_t c0 = tile * 32 + 4 * g;         // this thread's four byte columns
        uint32_t w[8][2];
        if (MODE != 1) {
#pragma unroll
            for (int kt = 0; kt < 8; ++kt) {
                w[kt][0] = *reinterpret_cast<const uint32_t *>(in + (size_t)(4 * kt + f0) * pitch + c0);
                w[kt][1] = *reinterpret_cast<const uint32_t *>(in + (size_t)(4 * kt + f0 + 2) * pitch + c0);
            }
        } else {
#pragma unroll
            for (int kt = 0; kt < 8; ++kt) w[kt][0] = w[kt][1] = 0x01010101u * (unsigned)(lane + kt);
        }
        int d[2][8][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int t = 0; t < 8; ++t)
#pragma unroll
                for (int i = 0; i < 4; ++i) d[m][t][i] = 0;
#pragma unroll
        for (int kt = 0; kt < 8; ++kt) {
            uint32_t a[2][4];
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                // m-tile m: row g -> column 4g + 2m, row g+8 -> column 4g + 2m + 1 (byte 2m / 2m+1 of the loaded words)
                if (MODE != 1) {
                    a[m][0] = expand4((w[kt][0] >> (16 * m + sh)) & 15u);
                    a[m][1] = expand4((w[kt][0] >> (16 * m + 8 + sh)) & 15u);
                    a[m][2] = expand4((w[kt][1] >> (16 * m + sh)) & 15u);
                    a[m][3] = expand4((w[kt][1] >> (16 * m + 8 + sh)) & 15u);
                } else {
                    a[m][0] = a[m][1] = w[kt][0];
                    a[m][2] = a[m][3] = w[kt][1];
                }
            }
            if (MODE != 2) {
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const uint2 b = sB[(kt * 8 + t) * 32 + lane];
                    mma_u8(d[0][t], a[0], b);
                    mma_u8(d[1][t], a[1], b);
                }
            } else {     // no MMAs and no B loads: the expanded operands are only kept alive
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    d[0][kt][i] += (int)a[0][i];
                    d[1][kt][i] += (int)a[1][i];
                }
            }
        }
        // repack: thread holds bits 2tg, 2tg+1 of parity row t for its four columns
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            uint32_t v = 0;
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                v |= (((uint32_t)d[m][t][0] & 1u) | (((uint32_t)d[m][t][1] & 1u) << 1)) << (16 * m);
                v |= (((uint32_t)d[m][t][2] & 1u) | (((uint32_t)d[m][t][3] & 1u) << 1)) << (16 * m + 8);
            }
            v <<= 2 * tg;
            v |= __shfl_xor_sync(0xffffffffu, v, 1);
            v |= __shfl_xor_sync(0xffffffffu, v, 2);
            if ((t >> 1) == tg) *reinterpret_cast<uint32_t *>(out + (size_t)t * opitch + c0) = v;
        }
    }
}

int main(int argc, char **argv)
{
    const size_t L = (size_t)(argc > 1 ? atof(argv[1]) : 64) * (1 << 20) / 32 * 32;
    const int reps = argc > 2 ? atoi(argv[2]) : 10;
    // generator: V[i][j] = (j+1)^i (reference src/matrix.cu:752-759)
    std::vector<uint8_t> G(P * K);
    for (int j = 0; j < K; ++j) {
        unsigned v = 1;
        for (int i = 0; i < P; ++i) {
            G[i * K + j] = (uint8_t)v;
            v = gf_mul_h(v, j + 1);
        }
    }
    // B fragments: b0 bytes i: k = tg*4 + i, n = g; b1: k + 16.  k = fl*8 + bit, fragment 4kt + fl.
    std::vector<uint2> bf(8 * 8 * 32);
    for (int kt = 0; kt < 8; ++kt)
        for (int t = 0; t < 8; ++t)
            for (int lane = 0; lane < 32; ++lane) {
                const int g = lane >> 2, tg = lane & 3;
                uint32_t r[2] = {0, 0};
                for (int h = 0; h < 2; ++h)
                    for (int i = 0; i < 4; ++i) {
                        const int kk = tg * 4 + i + 16 * h, fl = kk >> 3, bit = kk & 7;
                        const unsigned prod = gf_mul_h(G[t * K + 4 * kt + fl], 1u << bit);
                        r[h] |= ((prod >> g) & 1u) << (8 * i);
                    }
                bf[(kt * 8 + t) * 32 + lane] = make_uint2(r[0], r[1]);
            }
    uint8_t *d_in, *d_out, *d_ref, *d_G;
    uint2 *d_bf;
    CK(cudaMalloc(&d_in, (size_t)K * L));
    CK(cudaMalloc(&d_out, (size_t)P * L));
    CK(cudaMalloc(&d_ref, (size_t)P * L));
    CK(cudaMalloc(&d_G, P * K));
    CK(cudaMalloc(&d_bf, bf.size() * sizeof(uint2)));
    CK(cudaMemcpy(d_G, G.data(), P * K, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_bf, bf.data(), bf.size() * sizeof(uint2), cudaMemcpyHostToDevice));
    {
        std::vector<uint8_t> h((size_t)K * L);
        uint64_t s = 0x9E3779B97F4A7C15ull;
        for (size_t i = 0; i < h.size(); i += 8) {
            s += 0x9E3779B97F4A7C15ull;
            uint64_t z = s;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            z ^= z >> 31;
            memcpy(&h[i], &z, 8);
        }
        CK(cudaMemcpy(d_in, h.data(), h.size(), cudaMemcpyHostToDevice));
    }
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const size_t ntiles = L / 32;
    k_plain<<<sms * 8, 256>>>(d_in, L, d_ref, L, d_G, L);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const char *names[3] = {"whole kernel (load, expand, MMA, repack, store)", "MMAs only (constant operands)", "everything but the MMAs"};
    for (int mode = 0; mode < 3; ++mode) {
        CK(cudaMemset(d_out, 0xEE, (size_t)P * L));
        auto launch = [&]() {
            if (mode == 0) k_bitslice<0><<<sms * 2, 256>>>(d_in, L, d_out, L, d_bf, ntiles);
            else if (mode == 1) k_bitslice<1><<<sms * 2, 256>>>(d_in, L, d_out, L, d_bf, ntiles);
            else k_bitslice<2><<<sms * 2, 256>>>(d_in, L, d_out, L, d_bf, ntiles);
        };
        for (int i = 0; i < 3; ++i) launch();
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        for (int i = 0; i < reps; ++i) launch();
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        ms /= reps;
        const double in_gbs = (double)K * L / ms / 1e6, tr = (double)(K + P) * L / ms / 1e6;
        const double tmacs = (double)L / 32 * 128 * 4096 / ms / 1e9;     // 128 MMAs of 16x8x32 per 32 columns
        printf("mode %d  %-50s %8.3f ms  input %7.0f GB/s  traffic %7.0f GB/s  frac of 6448.7 = %.3f  (%.1f T MAC/s)\n", mode,
               names[mode], ms, in_gbs, tr, tr / 6448.7, tmacs);
        if (mode == 0) {
            std::vector<uint8_t> a((size_t)P * L), b((size_t)P * L);
            CK(cudaMemcpy(a.data(), d_out, a.size(), cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(b.data(), d_ref, b.size(), cudaMemcpyDeviceToHost));
            size_t bad = 0, first = 0;
            for (size_t i = 0; i < a.size(); ++i)
                if (a[i] != b[i]) {
                    if (!bad) first = i;
                    ++bad;
                }
            printf("        check against the plain GF(2^8) kernel: %zu mismatching bytes of %zu%s\n", bad, a.size(),
                   bad ? " -- WRONG" : " -- identical");
            if (bad) printf("        first at row %zu byte %zu: got %02x want %02x\n", first / L, first % L, a[first], b[first]);
        }
    }
    return 0;
}
