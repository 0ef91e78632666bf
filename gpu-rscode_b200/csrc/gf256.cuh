// gf256.cuh -- GF(2^8) device primitives (polynomial 0x11D).
//
// Replaces the reference's gf_mul / gf_div / gf_pow (src/matrix.cu:105-110,152-170,204-208).  The
// tables are generated on the device (gf_tables.cu) instead of being pasted literals in constant
// memory (src/matrix.cu:36-37); the hot multiply kernel does not use them at all -- it reads
// per-coefficient split-nibble product tables built from them once per coding matrix.
#pragma once
#include <stdint.h>

namespace rsgpu {

// Carry-less shift-and-reduce multiply: table-free, used only while the tables are being built.
__host__ __device__ __forceinline__ unsigned gf_mul_slow(unsigned a, unsigned b)
{
    unsigned r = 0;
    for (int i = 0; i < 8; ++i) {
        if (b & 1u) r ^= a;
        b >>= 1;
        a <<= 1;
        if (a & 0x100u) a ^= 0x11Du;
    }
    return r & 0xFFu;
}

// exp has 512 entries: exp[e] = exp[e+255] = x^e for e < 255, exp[510] = exp[511] = 0.
__device__ __forceinline__ unsigned gf_mul(const uint8_t *exp, const uint8_t *log, unsigned a, unsigned b)
{
    return (a && b) ? exp[log[a] + log[b]] : 0u;
}

__device__ __forceinline__ unsigned gf_inv(const uint8_t *exp, const uint8_t *log, unsigned a)
{
    return exp[255 - log[a]];   // a != 0
}

// Same convention as the reference (src/cpu-rs.c:158-162): log[0] reads as 0, so pow(0, e) = 1.
__device__ __forceinline__ unsigned gf_pow(const uint8_t *exp, const uint8_t *log, unsigned a, unsigned e)
{
    return exp[(log[a] * e) % 255u];
}

}  // namespace rsgpu
