// gf_matmul_fat_cluster_sm100.cu -- byte-table kernel for k = 17..32: the fragments are split over the two
// CTAs of a thread-block cluster.
//
// Why: one SM's shared memory holds one-lookup byte tables for at most 14 fragments; for larger k the
// single-CTA kernel keeps two-lookup nibble pages for the rest and becomes bound by the shared-memory pipe
// (RS(24,28) 0.81, RS(32,36) 0.71 of the HBM peak).  Here CTA r of a 2-CTA cluster owns fragments
// [r*KH, (r+1)*KH), KH = half the (padded) k: it builds byte tables for those only -- they all fit -- and reads
// only those fragments from HBM, so the traffic stays (k + p) * len.  Both CTAs walk the SAME columns (thread t
// of CTA 0 and thread t of CTA 1 hold partial parities of the same 16-byte column); before the store the pair
// exchanges halves through distributed shared memory: CTA 0 finishes output rows 0-1, CTA 1 rows 2-3, each
// sends its partial of the partner's rows (32 bytes per thread) with st.shared::cluster into the partner's
// exchange slot and signals the partner's per-warp mbarrier (remote arrive, release at cluster scope); the
// partner waits (acquire), XORs and stores.
//
// Ordering without an "empty" signal: the slots are double buffered by column parity.  A thread writes
// column i+2 into buffer i&1 only after it has finished column i+1, which includes waiting for the
// partner's column-i+1 signal -- and the partner sends that signal after it has consumed column i (program
// order).  So a buffer is never overwritten before it was read.  Loops are warp-uniform (a warp runs a
// column as long as its first lane has one) so that both CTAs run the same number of exchanges per warp;
// lanes past the end of the stripe take part in the exchange with junk and store nothing.
//
// Four output rows at most (32-bit table entries); wider groups stay on the single-CTA kernel.
//
// RESULT (B200, tools/sweep.py --variants 5,8): correct on every shape tested, but slower than the single-CTA
// kernel -- RS(17,20) 0.57 vs 0.84, RS(20,24) 0.67 vs 0.87, RS(24,28) 0.73 vs 0.81, RS(32,36) 0.66 vs 0.71 of
// the HBM peak.  The reason is arithmetic, not the exchange: a pair of SMs advances one column at a time, each
// doing k/2 fragments of it, so the device walks 74 column streams instead of 148; the byte-table inner loop
// already runs near its per-SM lookup rate (about 5.1e12 one-byte lookups per second device-wide: RS(12,16),
// RS(16,20)), and halving the work per SM per column cannot raise that rate, it only adds the exchange.  The
// ceiling for RS(24,28) with one lookup per byte is ~0.92 of the HBM peak against 0.81 today, so there is
// little to win.  Kept as variant 8 of rsgpu_set_tuning (never chosen automatically) with its parity tests.
#include "gf_matmul_fat.cuh"

namespace rsgpu {
namespace {

__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank)
{
    uint32_t a;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(smem_addr), "r"(rank));
    return a;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("st.shared::cluster.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}

// as ColumnBatches of gf_matmul_fat_sm100.cu (even batch counts only: batch 0 always sits in buf0)
template <int W, int K, int JB, int KB, int B, bool PF, bool T7, int XB>
struct ColumnBatchesC {
    __device__ __forceinline__ static void run(uint4 (&buf0)[JB], uint4 (&buf1)[JB], const uint8_t *&q, size_t in_pitch,
                                               const uint8_t *in, size_t vn, size_t nvec, uint32_t lane_base, uint32_t tb,
                                               Acc<W> &acc, uint32_t pf0, uint32_t pf1, int k)
    {
        constexpr int PAR = B & 1;
        uint4 (&cur)[JB] = PAR ? buf1 : buf0;
        uint4 (&nxt)[JB] = PAR ? buf0 : buf1;
        if constexpr (B + 1 < KB) {
            load_batch<JB, PF, true, (B + 1) * JB>(nxt, q, in_pitch, pf0, k);
        } else {
            if (vn < nvec) {
                q = in + vn * 16;
                load_batch<JB, PF, true, 0>(nxt, q, in_pitch, pf1, k);
            }
        }
        BatchMac<W, K, JB, B, 0, T7, XB>::run(cur, lane_base, tb, acc);
        if constexpr (B + 1 < KB)
            ColumnBatchesC<W, K, JB, KB, B + 1, PF, T7, XB>::run(buf0, buf1, q, in_pitch, in, vn, nvec, lane_base, tb, acc, pf0, pf1, k);
    }
};

__host__ __device__ constexpr int cluster_threads(int JB) { return fat_threads(4, JB, true); }
__host__ __device__ constexpr int cluster_xb(int JB) { return 2 * cluster_threads(JB) * 32 + 2 * (cluster_threads(JB) / 32) * 8 + 64; }

// JB x KB = KH fragments per CTA (KB even).  k = real fragment count of the whole code (<= 2 * KH).
template <int JB, int KB, bool PF>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(cluster_threads(JB), 1)
k_gf_matmul_fat_cluster(const uint8_t *__restrict__ in, size_t in_pitch, uint8_t *__restrict__ out, size_t out_pitch,
                        const uint32_t *__restrict__ tab, int nrows, size_t nvec, int k)
{
    constexpr int W = 4, KH = JB * KB, NT = cluster_threads(JB), NW = NT / 32, XB = cluster_xb(JB);
    constexpr bool T7 = true;
    static_assert((KB & 1) == 0, "even batch count");
    using M = Map<KH, XB>;
    extern __shared__ __align__(16) uint8_t s_raw[];
    const uint32_t tb = ((uint32_t)__cvta_generic_to_shared(s_raw) + 255u) & ~255u;
    const uint32_t xbuf = tb + ((M::BYTES + 15) & ~15);           // [2][NT] slots of 32 bytes
    const uint32_t bars = xbuf + 2 * NT * 32;                     // [2][NW] mbarriers
    const uint32_t rank = cluster_ctarank(), partner = rank ^ 1u;
    const int tid = threadIdx.x, warp = tid >> 5;

    // this CTA's half of the fragments
    int k_local = k - (int)rank * KH;
    k_local = k_local < 0 ? 0 : (k_local > KH ? KH : k_local);
    in += (size_t)rank * KH * in_pitch;
    tab += (size_t)rank * KH * 32;

    build_block_tables<W, KH, true, XB>(tb, tab, k_local);
    if (tid < 2 * NW) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bars + tid * 8), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    cluster_sync_all();      // both CTAs: tables built, barriers initialised and visible to the partner

    const uint32_t lane = tid & 31u;
    const uint32_t lane_base = tb + lane * 4u;
    const size_t nclusters = gridDim.x >> 1;
    const size_t stride = nclusters * NT;
    size_t v = (size_t)(blockIdx.x >> 1) * NT + tid;
    const uint32_t pf_stride = (uint32_t)(stride * 16);
    const uint32_t my_slot = xbuf + tid * 32, peer_slot = map_to_cta(my_slot, partner);
    const uint32_t my_bar = bars + warp * 8, peer_bar = map_to_cta(my_bar, partner);
    uint8_t *out_mine = out + (size_t)(2 * rank) * out_pitch;      // this CTA finishes rows 2*rank, 2*rank + 1
    const int rows_mine = nrows - 2 * (int)rank;                   // how many of them exist (<= 0: none)

    uint4 buf0[JB], buf1[JB];
    const uint8_t *q = in + v * 16;
    if (v < nvec) load_batch<JB, PF, true, 0>(buf0, q, in_pitch, v + stride < nvec ? pf_stride : 0u, k_local);

    for (uint32_t it = 0; v - lane < nvec; ++it, v += stride) {    // warp-uniform trip count
        const size_t vn = v + stride;
        const uint32_t pf0 = vn < nvec ? pf_stride : 0u;
        const uint32_t pf1 = vn + stride < nvec ? pf_stride : 0u;
        Acc<W> acc;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            acc.a[i] = 0;
            acc.t[i] = 0;
        }
        // lanes of the last, partial warp that lie past the end of the stripe take part in the exchange but must not
        // load: their q was never set for this column (k_eff = 0 makes every load of the column a zero fill)
        const int k_eff = v < nvec ? k_local : 0;
        ColumnBatchesC<W, KH, JB, KB, 0, PF, T7, XB>::run(buf0, buf1, q, in_pitch, in, vn, nvec, lane_base, tb, acc, pf0, pf1, k_eff);

        // 4x4 byte transpose into row order: r[row][word]
        uint32_t r[4][4];
#pragma unroll
        for (int qd = 0; qd < 4; ++qd) {
            const uint32_t a0 = acc.a[4 * qd], a1 = acc.a[4 * qd + 1], a2 = acc.a[4 * qd + 2], a3 = acc.a[4 * qd + 3];
            const uint32_t t0 = __byte_perm(a0, a1, 0x5140), t1 = __byte_perm(a2, a3, 0x5140);
            const uint32_t t2 = __byte_perm(a0, a1, 0x7362), t3 = __byte_perm(a2, a3, 0x7362);
            r[0][qd] = __byte_perm(t0, t1, 0x5410) ^ acc.t[qd];
            r[1][qd] = __byte_perm(t0, t1, 0x7632) ^ acc.t[4 + qd];
            r[2][qd] = __byte_perm(t2, t3, 0x5410) ^ acc.t[8 + qd];
            r[3][qd] = __byte_perm(t2, t3, 0x7632) ^ acc.t[12 + qd];
        }
        // send the partial of the partner's two rows, keep ours
        uint32_t keep[2][4], send[2][4];
#pragma unroll
        for (int qd = 0; qd < 4; ++qd) {
            keep[0][qd] = rank ? r[2][qd] : r[0][qd];
            keep[1][qd] = rank ? r[3][qd] : r[1][qd];
            send[0][qd] = rank ? r[0][qd] : r[2][qd];
            send[1][qd] = rank ? r[1][qd] : r[3][qd];
        }
        const uint32_t b = it & 1u, parity = (it >> 1) & 1u;
        st_cluster_v4(peer_slot + b * (NT * 32), send[0][0], send[0][1], send[0][2], send[0][3]);
        st_cluster_v4(peer_slot + b * (NT * 32) + 16, send[1][0], send[1][1], send[1][2], send[1][3]);
        // one remote arrive per warp, not per lane (a remote arrive is an atomic on the partner's shared-memory
        // port, the resource this kernel is short of): __syncwarp orders the 32 lanes' slot writes before
        // lane 0's release
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(peer_bar + b * (NW * 8));
        mbar_wait_cluster(my_bar + b * (NW * 8), parity);
        uint4 x0, x1;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x0.x), "=r"(x0.y), "=r"(x0.z), "=r"(x0.w) : "r"(my_slot + b * (NT * 32)) : "memory");
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x1.x), "=r"(x1.y), "=r"(x1.z), "=r"(x1.w) : "r"(my_slot + b * (NT * 32) + 16) : "memory");
        if (v < nvec) {
            if (rows_mine > 0)
                stg_stream(out_mine + v * 16, make_uint4(keep[0][0] ^ x0.x, keep[0][1] ^ x0.y, keep[0][2] ^ x0.z, keep[0][3] ^ x0.w));
            if (rows_mine > 1)
                stg_stream(out_mine + out_pitch + v * 16, make_uint4(keep[1][0] ^ x1.x, keep[1][1] ^ x1.y, keep[1][2] ^ x1.z, keep[1][3] ^ x1.w));
        }
    }
    cluster_sync_all();      // no CTA leaves while its partner may still write into its shared memory
}

using ClusterFn = void (*)(const uint8_t *, size_t, uint8_t *, size_t, const uint32_t *, int, size_t, int);

struct ClusterKernel {
    ClusterFn fn;
    int threads;
    int smem;
};

template <int JB, int KB>
ClusterKernel make_cluster()
{
    constexpr int KH = JB * KB;
    return ClusterKernel{k_gf_matmul_fat_cluster<JB, KB, true>, cluster_threads(JB), ((Map<KH, cluster_xb(JB)>::BYTES + 15) & ~15) + cluster_xb(JB) + 256};
}

}  // namespace

// cudaErrorNotSupported unless 17 <= k <= 32 and the group has at most 4 rows (32-bit entries).
cudaError_t launch_matmul_fat_cluster(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                                      size_t out_pitch, size_t nvec, const DeviceInfo &dev, int max_blocks, cudaStream_t s)
{
    if (g.width != 4 || k < 17 || k > 32 || g.xor_row >= 0) return cudaErrorNotSupported;
    // half of k, rounded up to a shape with an even batch count
    ClusterKernel ck = k <= 20 ? make_cluster<5, 2>() : (k <= 24 ? make_cluster<3, 4>() : (k <= 28 ? make_cluster<7, 2>() : make_cluster<4, 4>()));
    if (ck.smem > dev.max_smem_optin) return cudaErrorNotSupported;
    int occ = 0;
    cudaError_t e = prepare_launch(dev, (const void *)ck.fn, ck.threads, (size_t)ck.smem, &occ);
    if (e != cudaSuccess) return e;
    // how many 2-CTA clusters the device runs at once with this much shared memory (one CTA per SM, both
    // SMs in one GPC): size the persistent grid from that
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)dev.sm_count & ~1u);
    cfg.blockDim = dim3((unsigned)ck.threads);
    cfg.dynamicSmemBytes = (size_t)ck.smem;
    int nclusters = 0;
    e = cudaOccupancyMaxActiveClusters(&nclusters, (const void *)ck.fn, &cfg);
    if (e != cudaSuccess || nclusters < 1) {
        cudaGetLastError();
        return cudaErrorNotSupported;
    }
    size_t want = (nvec + ck.threads - 1) / ck.threads;
    size_t cap = (size_t)nclusters;
    if (max_blocks > 1 && (size_t)(max_blocks / 2) < cap) cap = max_blocks / 2;
    const unsigned grid = 2u * (unsigned)(want < cap ? want : cap);
    ck.fn<<<grid, ck.threads, ck.smem, s>>>(in, in_pitch, out_g, out_pitch, g.d_tab, g.nrows, nvec, k);
    return cudaGetLastError();
}

}  // namespace rsgpu
