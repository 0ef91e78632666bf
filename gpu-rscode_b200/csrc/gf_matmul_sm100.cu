// gf_matmul_sm100.cu -- THE hot path: C (rows x len) = M (rows x k) * B (k x len) over GF(2^8).
//
// Replaces the reference's matrix_mul<long long> / matrix_mul<unsigned char> kernels and their
// launchers encode_chunk / decode_chunk (src/matrix.cu:232-323, 335-407, 767-905).
//
// Design (sm_100a, HBM-bound integer/byte work -- no tensor cores):
//   * every thread owns 16-byte columns: one 128-bit coalesced load per input fragment, one
//     128-bit store per output row, grid-stride over the stripe (persistent grid sized from the
//     SM count); input bytes are read from HBM exactly once, outputs written exactly once.
//   * GF multiply = per-coefficient split-nibble product tables in shared memory.  For input
//     fragment j there are two 16-entry tables (low nibble, high nibble); an entry packs the
//     products for ALL output rows of the group, one byte per row (4 rows -> one 32-bit word,
//     8 rows -> 64 bit), so ONE conflict-free shared-memory lookup per nibble yields the
//     contribution to every parity at once: 2 lookups per input byte regardless of the row count.
//     A 16-entry table of 4-byte words spans 16 distinct banks, and equal indices broadcast, so a
//     warp's 32 random nibbles never conflict.
//   * multiply-accumulate is XOR into one accumulator word per byte position (bytes = rows); a
//     4x4 byte transpose with PRMT at the end turns them into the row-major 16-byte stores.
//   * the next batch of fragment loads is issued before the current batch is consumed (register
//     double buffering), so each thread always has JB*16 bytes in flight.
//   The reference instead does, per byte-product, two log lookups, an add and an exp lookup in
//   bank-conflicted shared tables, re-stages A and B into shared memory redundantly in every
//   thread with two __syncthreads per tile, and moves 8 bytes per access.
#include "gf_matmul_common.cuh"
#include "gf_realign.cuh"
#include "rs_internal.h"

namespace rsgpu {
namespace {

// W: table entry bytes (4: up to 4 output rows, 8: up to 8).  JB: fragments per load batch.
// kb = number of batches (kb*JB >= k, kb*JB <= kpad); PAD = (kb*JB > k).
template <int W, int JB, int EX, bool PAD, bool XR = false>
__global__ void __launch_bounds__(256)
k_gf_matmul(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
            const uint32_t *__restrict__ tab, int npages, int k, int kb, int nrows, size_t nvec, uint32_t hi4)
{
    // one 256-byte page per fragment (see mac_word); the launcher adds 256 bytes of slack so the base
    // can be rounded up to a 256-byte boundary whatever offset the dynamic window starts at
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;
    {
        constexpr int V4 = 32 * W / 16;   // uint4 per fragment in the global table (lo + hi tables)
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
        uint4 x = make_uint4(0u, 0u, 0u, 0u);     // XR: the peeled all-ones row (see gf_matmul_fat.cuh)
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
                mac_vec<W, EX>(cur[jj], t0 + jj * 256u, hi4, acc);
                if constexpr (XR) {
                    // padded slots re-read the last valid row: they must not enter the XOR twice
                    if (!PAD || b * JB + jj < k) {
                        x.x ^= cur[jj].x;
                        x.y ^= cur[jj].y;
                        x.z ^= cur[jj].z;
                        x.w ^= cur[jj].w;
                    }
                }
            }
        }
        transpose_store(acc.a, out, out_pitch, v * 16, 0, nrows);
        if constexpr (W == 8) transpose_store(acc.b, out, out_pitch, v * 16, 4, nrows - 4);
        if constexpr (XR) stg_stream(out - out_pitch + v * 16, x);
        v = vn;
    }
}

// Rows of any alignment on split-nibble tables (geometry and realignment: gf_realign.cuh) -- the counterpart of
// gf_matmul_fat_ua_sm100.cu for every k that has no byte-table shape (25, 28, 33 .. 255), so that an odd pitch never
// drops a long stripe onto the byte-per-thread kernel whatever the code.  Runtime loop over the fragments in batches of 4,
// the next batch loaded while the current one is consumed; several blocks per SM.
template <int W, bool XR>
__global__ void __launch_bounds__(256)
k_gf_matmul_ua(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
               const uint32_t *__restrict__ tab, int npages, int k, int nrows, size_t nv, uint32_t hi4)
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

    constexpr int JB = 4;
    const uint32_t lane = threadIdx.x & 31u;
    const uint8_t *in_h = in + kUaHead;
    uint8_t *out_h = out + kUaHead;
    const uint32_t s0 = (uint32_t)reinterpret_cast<uintptr_t>(in_h) & 15u, pm = (uint32_t)in_pitch & 15u;
    const size_t nst = nv - 2, nspan = (nst + kUaSpan - 1) / kUaSpan;
    const size_t wpb = blockDim.x >> 5, wtotal = (size_t)gridDim.x * wpb, nstride = wtotal * kUaSpan;
    // loop control from blockIdx and kernel arguments only: whole warps at every shuffle
    size_t S0 = (size_t)blockIdx.x * wpb;
    size_t n = (S0 + (threadIdx.x >> 5)) * kUaSpan + lane;
    for (; S0 < nspan; S0 += wtotal, n += nstride) {
        Acc<W> acc;
        acc_clear<W>(acc);
        uint4 x = make_uint4(0u, 0u, 0u, 0u);
        const bool inr = n < nv;
        const uint8_t *q = in_h + n * 16;
        uint32_t s = s0;
        // the next batch of four rows is in flight while the current one is consumed (ncu of the first version, which
        // loaded and used batch by batch: long-scoreboard 6.5 stalls per issued instruction)
        auto load4 = [&](uint4 (&d)[JB], int j0) {
#pragma unroll
            for (int jj = 0; jj < JB; ++jj) {
                d[jj] = make_uint4(0u, 0u, 0u, 0u);
                if (j0 + jj < k && inr)
                    d[jj] = ldg_stream(reinterpret_cast<const uint8_t *>(reinterpret_cast<uintptr_t>(q) & ~(uintptr_t)15));
                q += in_pitch;
            }
        };
        uint4 nxt[JB];
        load4(nxt, 0);
        for (int j0 = 0; j0 < k; j0 += JB) {
            uint4 d[JB];
#pragma unroll
            for (int jj = 0; jj < JB; ++jj) d[jj] = nxt[jj];
            if (j0 + JB < k) load4(nxt, j0 + JB);
#pragma unroll
            for (int jj = 0; jj < JB; ++jj) {
                const uint4 w = realign(d[jj], shfl_next(d[jj]), s);     // padded slots: zeros against zero pages
                s = (s + pm) & 15u;
                mac_vec<W, 2>(w, tb + (uint32_t)(j0 + jj) * 256u, hi4, acc);
                if constexpr (XR) {
                    x.x ^= w.x;
                    x.y ^= w.y;
                    x.z ^= w.z;
                    x.w ^= w.w;
                }
            }
        }
        const bool ok = lane < (uint32_t)kUaSpan && n < nst;
        uint32_t r[4][4];
        transpose_rows(acc.a, r, nullptr);
        store_rows_realigned(r, out_h, out_pitch, n, 0, nrows, ok);
        if constexpr (W == 8) {
            transpose_rows(acc.b, r, nullptr);
            store_rows_realigned(r, out_h, out_pitch, n, 4, nrows - 4, ok);
        }
        if constexpr (XR) {
            r[0][0] = x.x;
            r[0][1] = x.y;
            r[0][2] = x.z;
            r[0][3] = x.w;
            store_rows_realigned(r, out_h, out_pitch, n, -1, 1, ok);
        }
    }
}

// One output row whose coefficients are all 1 -- the single parity of RS(k, k+1), or the one lost native of a stripe
// that still has parity row 0 (the commonest repair there is): the row is the XOR of the k inputs, no tables, no
// lookups.  U loads in flight per thread, 8 blocks of 256 threads per SM: plain HBM streaming.
template <int U>
__global__ void __launch_bounds__(256)
k_xor_rows(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, int k, size_t nvec)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
        uint4 acc = make_uint4(0u, 0u, 0u, 0u);
        const uint8_t *q = in + v * 16;
        int j = 0;
        for (; j + U <= k; j += U) {
            uint4 d[U];
#pragma unroll
            for (int u = 0; u < U; ++u) d[u] = ldg_stream(q + (size_t)(j + u) * in_pitch);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                acc.x ^= d[u].x;
                acc.y ^= d[u].y;
                acc.z ^= d[u].z;
                acc.w ^= d[u].w;
            }
        }
        for (; j < k; ++j) {
            const uint4 d = ldg_stream(q + (size_t)j * in_pitch);
            acc.x ^= d.x;
            acc.y ^= d.y;
            acc.z ^= d.z;
            acc.w ^= d.w;
        }
        stg_stream(out + v * 16, acc);
    }
}

// Byte-granular fallback: heads/tails shorter than 16 bytes and buffers that are not 16-byte aligned
// (the reference's matrix_mul<unsigned char>, src/matrix.cu:335-407, used whenever size % 8 != 0).
template <int W>
__global__ void __launch_bounds__(256)
k_gf_matmul_bytes(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
                  const uint32_t *__restrict__ tab, int k, int nrows, size_t c0, size_t len, int xr)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t c = c0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < len; c += stride) {
        uint32_t a = 0, b = 0;
        unsigned xs = 0;       // xr: the peeled all-ones row, written one row above the group's rows
        for (int j = 0; j < k; ++j) {
            const unsigned x = in[(size_t)j * in_pitch + c];
            xs ^= x;
            const uint32_t *tl = tab + (size_t)j * 32 * (W / 4);
            const uint32_t *th = tl + 16 * (W / 4);
            if (W == 4) {
                a ^= __ldg(tl + (x & 15u)) ^ __ldg(th + (x >> 4));
            } else {
                a ^= __ldg(tl + 2 * (x & 15u)) ^ __ldg(th + 2 * (x >> 4));
                b ^= __ldg(tl + 2 * (x & 15u) + 1) ^ __ldg(th + 2 * (x >> 4) + 1);
            }
        }
        for (int r = 0; r < nrows; ++r) {
            const uint32_t wv = r < 4 ? a : b;
            out[(size_t)r * out_pitch + c] = (uint8_t)(wv >> (8 * (r & 3)));
        }
        if (xr) (out - out_pitch)[c] = (uint8_t)xs;
    }
}

// The same for the last (len % 16) bytes of every stripe of a batch: one thread per (stripe, byte).
template <int W>
__global__ void __launch_bounds__(256)
k_gf_matmul_bytes_batch(const uint8_t *__restrict__ in, size_t in_pitch, size_t in_stripe, uint8_t *__restrict__ out,
                        size_t out_pitch, size_t out_stripe, const uint32_t *__restrict__ tab, int k, int nrows, size_t c0,
                        unsigned rem, size_t n_stripes, int xr)
{
    const size_t total = n_stripes * rem, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t st = i / rem, c = c0 + (i - st * rem);
        const uint8_t *src = in + st * in_stripe + c;
        uint32_t a = 0, b = 0;
        unsigned xs = 0;
        for (int j = 0; j < k; ++j) {
            const unsigned x = src[(size_t)j * in_pitch];
            xs ^= x;
            const uint32_t *tl = tab + (size_t)j * 32 * (W / 4);
            const uint32_t *th = tl + 16 * (W / 4);
            if (W == 4) {
                a ^= __ldg(tl + (x & 15u)) ^ __ldg(th + (x >> 4));
            } else {
                a ^= __ldg(tl + 2 * (x & 15u)) ^ __ldg(th + 2 * (x >> 4));
                b ^= __ldg(tl + 2 * (x & 15u) + 1) ^ __ldg(th + 2 * (x >> 4) + 1);
            }
        }
        uint8_t *dst = out + st * out_stripe + c;
        for (int r = 0; r < nrows; ++r) {
            const uint32_t wv = r < 4 ? a : b;
            dst[(size_t)r * out_pitch] = (uint8_t)(wv >> (8 * (r & 3)));
        }
        if (xr) *(dst - out_pitch) = (uint8_t)xs;
    }
}

using MatmulFn = void (*)(const uint8_t *, size_t, uint8_t *, size_t, const uint32_t *, int, int, int, int, size_t, uint32_t);

template <int W, int EX>
MatmulFn pick_jb(int jb, bool pad)
{
    switch (jb) {
    case 1: return k_gf_matmul<W, 1, EX, false>;
    case 2: return k_gf_matmul<W, 2, EX, false>;
    case 3: return k_gf_matmul<W, 3, EX, false>;
    case 4: return pad ? k_gf_matmul<W, 4, EX, true> : k_gf_matmul<W, 4, EX, false>;
    case 5: return pad ? k_gf_matmul<W, 5, EX, true> : k_gf_matmul<W, 5, EX, false>;
    case 6: return pad ? k_gf_matmul<W, 6, EX, true> : k_gf_matmul<W, 6, EX, false>;
    case 7: return pad ? k_gf_matmul<W, 7, EX, true> : k_gf_matmul<W, 7, EX, false>;
    default: return pad ? k_gf_matmul<W, 8, EX, true> : k_gf_matmul<W, 8, EX, false>;
    }
}

// the peeled-row instances: 32-bit entries, DP4A extraction only
MatmulFn pick_kernel_xr(int jb, bool pad)
{
    switch (jb) {
    case 1: return k_gf_matmul<4, 1, 2, false, true>;
    case 2: return k_gf_matmul<4, 2, 2, false, true>;
    case 3: return k_gf_matmul<4, 3, 2, false, true>;
    case 4: return pad ? k_gf_matmul<4, 4, 2, true, true> : k_gf_matmul<4, 4, 2, false, true>;
    case 5: return pad ? k_gf_matmul<4, 5, 2, true, true> : k_gf_matmul<4, 5, 2, false, true>;
    case 6: return pad ? k_gf_matmul<4, 6, 2, true, true> : k_gf_matmul<4, 6, 2, false, true>;
    case 7: return pad ? k_gf_matmul<4, 7, 2, true, true> : k_gf_matmul<4, 7, 2, false, true>;
    default: return pad ? k_gf_matmul<4, 8, 2, true, true> : k_gf_matmul<4, 8, 2, false, true>;
    }
}

MatmulFn pick_kernel(int width, int jb, bool prmt, bool pad)
{
    if (width == 4) return prmt ? pick_jb<4, 0>(jb, pad) : pick_jb<4, 2>(jb, pad);
    return prmt ? pick_jb<8, 0>(jb, pad) : pick_jb<8, 2>(jb, pad);
}

// Fragments per load batch.  A divisor of k in 3..8 if there is one (largest first): no padded slots.
// Otherwise (primes above 8, ...) the batch size in 4..8 that wastes the fewest padded slots; the slots
// behind k re-read the last row against all-zero tables, so waste costs time, not correctness.
int choose_jb(int k)
{
    if (k <= 8) return k;
    for (int jb = 8; jb >= 3; --jb)
        if (k % jb == 0) return jb;
    int best = 8, best_waste = 1 << 30;
    for (int jb = 8; jb >= 4; --jb) {
        const int waste = (k + jb - 1) / jb * jb - k;
        if (waste < best_waste) {
            best = jb;
            best_waste = waste;
        }
    }
    return best;
}

}  // namespace

cudaError_t prepare_launch(const DeviceInfo &dev, const void *fn, int threads, size_t smem, int *occ)
{
    LaunchMemo *m = dev.memo;
    if (m)
        for (int i = 0; i < m->n; ++i)
            if (m->fn[i] == fn && m->threads[i] == threads && m->smem[i] == (int)smem) {
                *occ = m->occ[i];
                return cudaSuccess;
            }
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    int o = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, fn, threads, smem);
    if (e != cudaSuccess) return e;
    if (o < 1) o = 1;
    *occ = o;
    if (m && m->n < LaunchMemo::kSlots) {
        m->fn[m->n] = fn;
        m->threads[m->n] = threads;
        m->smem[m->n] = (int)smem;
        m->occ[m->n] = o;
        ++m->n;
    }
    return cudaSuccess;
}

cudaError_t launch_matmul_group(const RowGroup &g, int k, int kpad,
                                const uint8_t *in, size_t in_pitch, uint8_t *out, size_t out_pitch,
                                size_t len, const Tuning &tune, const DeviceInfo &dev,
                                cudaStream_t s, long long *launches)
{
    if (len == 0) return cudaSuccess;
    uint8_t *out_g = out + (size_t)g.row0 * out_pitch;
    const bool aligned = (((uintptr_t)in | (uintptr_t)out_g | in_pitch | out_pitch) & 15u) == 0;
    const size_t nvec = aligned ? len / 16 : 0;
    const int threads = (tune.threads > 0 && tune.threads <= 256) ? tune.threads : 256;
    const int xr = g.xor_row >= 0;    // this launch also writes the peeled all-ones row, one row above out_g
    if (xr && (g.width != 4 || g.xor_row != g.row0 - 1)) return cudaErrorInvalidValue;

    // byte-granular kernel over the columns [c0, end)
    auto launch_bytes = [&](size_t c0, size_t end) -> cudaError_t {
        if (c0 >= end) return cudaSuccess;
        size_t blocks = (end - c0 + 255) / 256;
        const size_t cap = (size_t)dev.sm_count * 8;
        if (blocks > cap) blocks = cap;
        if (g.width == 4)
            k_gf_matmul_bytes<4><<<(unsigned)blocks, 256, 0, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, k, g.nrows, c0, end, xr);
        else
            k_gf_matmul_bytes<8><<<(unsigned)blocks, 256, 0, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, k, g.nrows, c0, end, xr);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess && launches) ++*launches;
        return e;
    };

    // Rows that are not 16-byte aligned (a base or a pitch that is not a multiple of 16 -- e.g. pitch = ceil(size/k),
    // the layout of a file's fragments held back to back): the byte-table kernel with in-register realignment
    // (gf_matmul_fat_ua_sm100.cu) from 1 MiB per fragment on; its head and tail, shorter stripes and k without an
    // instance stay on the byte kernel.  Variant 6 forbids it (the byte kernel alone, for comparison).
    if (!aligned && tune.variant != 6 && len >= ((size_t)1 << 20)) {
        size_t lo = 0, hi = 0;
        cudaError_t e = launch_matmul_fat_ua(g, k, in, in_pitch, out_g, out_pitch, len, dev, tune.max_blocks, s, &lo, &hi);
        if (e == cudaSuccess) {
            if (launches) ++*launches;
            if ((e = launch_bytes(0, lo)) != cudaSuccess) return e;
            return launch_bytes(hi, len);
        }
        if (e != cudaErrorNotSupported) return e;
        // no byte-table shape for this k: the same realignment on split-nibble tables
        if (len >= (size_t)kUaHead + 48 + 16 * kUaSpan && (size_t)kpad * 256 + 256 <= (size_t)dev.max_smem_optin) {
            using UaFn = void (*)(const uint8_t *, size_t, uint8_t *, size_t, const uint32_t *, int, int, int, size_t, uint32_t);
            const UaFn fn = g.width == 4 ? (xr ? (UaFn)k_gf_matmul_ua<4, true> : (UaFn)k_gf_matmul_ua<4, false>) : (UaFn)k_gf_matmul_ua<8, false>;
            const size_t smem = (size_t)kpad * 256 + 256;
            int occ = 0;
            if ((e = prepare_launch(dev, (const void *)fn, 256, smem, &occ)) != cudaSuccess) return e;
            const size_t nv = (len - kUaHead) / 16, nst = nv - 2, nspan = (nst + kUaSpan - 1) / kUaSpan;
            size_t want = (nspan + 7) / 8, cap = (size_t)dev.sm_count * occ;
            if (tune.max_blocks > 0 && (size_t)tune.max_blocks < cap) cap = tune.max_blocks;
            fn<<<(unsigned)(want < cap ? want : cap), 256, smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, kpad, k, g.nrows, nv,
                                                                    16u * g.width * 0x01010101u);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            if (launches) ++*launches;
            if ((e = launch_bytes(0, kUaHead + 16)) != cudaSuccess) return e;
            return launch_bytes((size_t)kUaHead + 16 * nst, len);
        }
    }

    // a single all-ones row: XOR of the inputs, no lookups (the byte tail below still uses the tables, which exist)
    bool xor_done = false;
    if (nvec && g.all_ones && g.nrows == 1 && !xr && tune.variant == 0) {
        int occ = 0;
        cudaError_t e = prepare_launch(dev, (const void *)k_xor_rows<8>, 256, 0, &occ);
        if (e != cudaSuccess) return e;
        size_t want = (nvec + 255) / 256, cap = (size_t)dev.sm_count * occ;
        if (tune.max_blocks > 0 && (size_t)tune.max_blocks < cap) cap = tune.max_blocks;
        k_xor_rows<8><<<(unsigned)(want < cap ? want : cap), 256, 0, s>>>(in, in_pitch, out_g, k, nvec);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) ++*launches;
        xor_done = true;
    }

    // variant 9: the multi-pass kernel (gf_matmul_fat_mp_sm100.cu) for k > 14 -- the fragments taken 14 at a time through
    // all-byte-table passes, partial parities kept in registers across the table rebuilds.  Measured SLOWER than the
    // nibble kernels it was meant to replace for k >= 25 (RS(28,32) 0.55 vs 0.67, RS(64,68) 0.48 vs 0.64, RS(128,132)
    // 0.42 vs 0.56 of the HBM peak), so it is never chosen automatically; kept as a tested, selectable experiment
    // (DESIGN.md section 4.1 h).  It codes whole block steps; the columns behind them go down the ordinary path.
    if (nvec && tune.variant == 9 && k > 14) {
        size_t done = 0;
        cudaError_t e = launch_matmul_fat_mp(g, k, in, in_pitch, out_g, out_pitch, nvec, dev, tune.max_blocks, s, &done);
        if (e == cudaSuccess) {
            if (launches) ++*launches;
            if (done * 16 >= len) return cudaSuccess;
            Tuning rest = tune;
            rest.variant = 6;
            return launch_matmul_group(g, k, kpad, in + done * 16, in_pitch, out + done * 16, out_pitch, len - done * 16, rest, dev, s,
                                       launches);
        }
        if (e != cudaErrorNotSupported) return e;
    }

    // variant 0 (default): the bank-replicated kernel when an instance exists for this k and the stripe
    // is long enough to pay for building its tables (one 227 KiB fill per SM: measured break-even
    // on B200 at 4-8 MiB per fragment for RS(10,14)); 5 forces it, 6 forbids it
    bool done_vec = xor_done;
    // variant 8: the cluster kernel for k = 17..32 (fragments split over two CTAs, partial parities exchanged
    // through distributed shared memory).  Measured SLOWER than the single-CTA kernel on every such code
    // (RS(24,28) 0.73 vs 0.81, RS(20,24) 0.67 vs 0.87, RS(32,36) 0.66 vs 0.71 of the HBM peak), so it is never
    // chosen automatically; it stays as a tested, selectable experiment (DESIGN.md section 4.1 e)
    if (nvec && tune.variant == 8 && !xr) {
        cudaError_t e = launch_matmul_fat_cluster(g, k, in, in_pitch, out_g, out_pitch, nvec, dev, tune.max_blocks, s);
        if (e == cudaSuccess) {
            done_vec = true;
            if (launches) ++*launches;
        } else if (e != cudaErrorNotSupported) {
            return e;
        }
    }
    if (nvec && !done_vec && (tune.variant == 5 || (tune.variant == 0 && len >= ((size_t)8 << 20)))) {
        cudaError_t e = launch_matmul_fat(g, k, in, in_pitch, out_g, out_pitch, nvec, dev, tune.max_blocks, tune.threads, s);
        if (e == cudaSuccess) {
            done_vec = true;
            if (launches) ++*launches;
        } else if (e != cudaErrorNotSupported) {
            return e;
        }
    }
    // next choice: the nibble kernel fed by TMA bulk copies (2-3 % faster than the register-fed one from
    // ~2 MiB per fragment on: RS(16,20) 0.750 vs 0.731, RS(8,12) 0.860 vs 0.838 of the peak); needs k % 8, 5 or 4 == 0
    if (nvec && !done_vec && !xr && (tune.variant == 7 || (tune.variant == 0 && len >= ((size_t)2 << 20)))) {
        cudaError_t e = launch_matmul_tma(g, k, kpad, in, in_pitch, out_g, out_pitch, nvec, dev, tune.max_blocks, s);
        if (e == cudaSuccess) {
            done_vec = true;
            if (launches) ++*launches;
        } else if (e != cudaErrorNotSupported) {
            return e;
        }
    }
    if (nvec && !done_vec) {
        const int jb = choose_jb(k);
        const int kb = (k + jb - 1) / jb;           // kb*jb <= kpad: choose_jb never wastes more than jb = 8 does
        if (kb * jb > kpad) return cudaErrorInvalidValue;
        // index extraction: DP4A (default; same speed as the PRMT splice on B200 but the alu pipe sits at
        // 45 % instead of 81 %, profiles/r1_pipes_variant*.csv); variant 1 selects the PRMT splice
        MatmulFn fn = xr ? pick_kernel_xr(jb, kb * jb > k) : pick_kernel(g.width, jb, tune.variant == 1, kb * jb > k);
        const size_t smem = (size_t)kpad * 256 + 256;   // one page per fragment + alignment slack
        int occ = 0;
        cudaError_t e = prepare_launch(dev, (const void *)fn, threads, smem, &occ);
        if (e != cudaSuccess) return e;
        size_t want = (nvec + threads - 1) / threads;
        size_t cap = (size_t)dev.sm_count * occ;
        if (tune.max_blocks > 0 && (size_t)tune.max_blocks < cap) cap = tune.max_blocks;
        const unsigned grid = (unsigned)(want < cap ? want : cap);
        fn<<<grid, threads, smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, kpad, k, kb, g.nrows, nvec,
                                       16u * g.width * 0x01010101u);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
    }
    if (nvec * 16 < len) {
        cudaError_t e = launch_bytes(nvec * 16, len);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// Bytes [c0, c0 + rem) of every stripe of a batch (rem = len % 16 <= 15) in one launch.
cudaError_t launch_matmul_bytes_batch(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, size_t in_stripe,
                                      uint8_t *out_g, size_t out_pitch, size_t out_stripe, size_t c0, unsigned rem,
                                      size_t n_stripes, const DeviceInfo &dev, cudaStream_t s)
{
    if (rem == 0 || n_stripes == 0) return cudaSuccess;
    size_t blocks = (n_stripes * rem + 255) / 256;
    const size_t cap = (size_t)dev.sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (g.width == 4)
        k_gf_matmul_bytes_batch<4><<<(unsigned)blocks, 256, 0, s>>>(in, in_pitch, in_stripe, out_g, out_pitch, out_stripe, g.d_tab, k,
                                                                    g.nrows, c0, rem, n_stripes, g.xor_row >= 0);
    else
        k_gf_matmul_bytes_batch<8><<<(unsigned)blocks, 256, 0, s>>>(in, in_pitch, in_stripe, out_g, out_pitch, out_stripe, g.d_tab, k,
                                                                    g.nrows, c0, rem, n_stripes, g.xor_row >= 0);
    return cudaGetLastError();
}

}  // namespace rsgpu
