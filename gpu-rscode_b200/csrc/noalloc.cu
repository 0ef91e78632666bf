// noalloc.cu -- rsgpu_encode_noalloc: RS encode for callers that must not allocate GPU memory.
//
// Who this is for: code that embeds the engine inside a fixed memory budget or a latency-critical
// I/O path -- e.g. a storage daemon whose GPU memory was carved up at start-up and handed out as
// arenas, or a plugin loaded into a process that forbids cudaMalloc after initialisation because
// cudaMalloc/cudaFree serialise the whole device and stall every other stream.  Such callers cannot
// use rsgpu_create(), which allocates the field tables, the generator matrix and the product tables.
//
// Here that scratch lives in ONE file-scope __device__ array of fixed size, part of the library
// image: every GPU that loads librsgpu.so holds these 16 KiB until the process exits, whether or not
// this entry point is ever called.  The price is a hard limit on the code parameters (k <= 32,
// p <= 8) and no concurrency: every call on a device rewrites and then reads the same array.
#include <mutex>

#include "../../include/rsgpu.h"
#include "rs_ctx.h"

using namespace rsgpu;

namespace {

constexpr int kNoallocMaxK = RSGPU_NOALLOC_MAX_K;
constexpr int kNoallocMaxP = RSGPU_NOALLOC_MAX_P;

// Layout of the scratch array (byte offsets, all multiples of 256):
//   [0, 768)        GF(2^8) exp (512) and log (256) tables
//   [1024, 2304)    total matrix [I_k ; V], at most (32 + 8) x 32 bytes
//   [4096, 12288)   product tables, at most kpad(32) x 32 entries x 8 bytes
constexpr size_t kOffGf = 0;
constexpr size_t kOffMatrix = 1024;
constexpr size_t kOffTables = 4096;
constexpr size_t kScratchBytes = 16384;
static_assert(kOffMatrix + (size_t)(kNoallocMaxK + kNoallocMaxP) * kNoallocMaxK <= kOffTables, "matrix region too small");
static_assert(kOffTables + (size_t)kNoallocMaxK * 32 * 8 <= kScratchBytes, "table region too small");

static __device__ __align__(256) uint8_t g_noalloc_scratch[kScratchBytes];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an address a kernel can use.
// The device address differs per GPU, so it is looked up once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_mu;
uint8_t *g_scratch_addr[kMaxDevices] = {};
DeviceInfo g_dev_info[kMaxDevices];
LaunchMemo g_memo[kMaxDevices];   // host memory, one per device; guarded by the caller's one-call-at-a-time rule

int scratch_for_current_device(uint8_t **addr, DeviceInfo *info)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) return fail_msg(RSGPU_ETOOBIG, "rsgpu_encode_noalloc supports device ordinals below 64");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        uint8_t *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        int sms = 0, smem = 0;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess)
            return fail_cuda(e, "cudaDeviceGetAttribute(MultiProcessorCount)");
        if ((e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev)) != cudaSuccess)
            return fail_cuda(e, "cudaDeviceGetAttribute(MaxSharedMemoryPerBlockOptin)");
        g_dev_info[dev].sm_count = sms;
        g_dev_info[dev].max_smem_optin = smem;
        g_dev_info[dev].memo = &g_memo[dev];
        g_scratch_addr[dev] = p;
    }
    *addr = g_scratch_addr[dev];
    *info = g_dev_info[dev];
    return RSGPU_OK;
}

}  // namespace

extern "C" int rsgpu_encode_noalloc(int k, int p, const uint8_t *data_d, size_t data_pitch, uint8_t *parity_d,
                                    size_t parity_pitch, size_t len, void *stream)
{
    if (k < 1 || p < 1) return fail_msg(RSGPU_EINVAL, "need k >= 1 and p >= 1");
    if (k > kNoallocMaxK || p > kNoallocMaxP)
        return fail_msg(RSGPU_ETOOBIG, "rsgpu_encode_noalloc is limited to k <= 32 and p <= 8 (fixed 16 KiB scratch); "
                                       "use rsgpu_create + rsgpu_encode for larger codes");
    if (!data_d || !parity_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if ((k > 1 && data_pitch < len) || (p > 1 && parity_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    if (len == 0) return RSGPU_OK;

    uint8_t *scratch = nullptr;
    DeviceInfo info;
    int rc = scratch_for_current_device(&scratch, &info);
    if (rc != RSGPU_OK) return rc;

    cudaStream_t s = (cudaStream_t)stream;
    GfTables gf;
    gf.d_exp = scratch + kOffGf;
    gf.d_log = scratch + kOffGf + kExpSize;
    uint8_t *d_total = scratch + kOffMatrix;
    RowGroup g;
    g.row0 = 0;
    g.nrows = p;
    g.width = p > 4 ? 8 : 4;
    g.d_tab = reinterpret_cast<uint32_t *>(scratch + kOffTables);
    const int kpad = (k + 7) / 8 * 8;
    g.tab_words = (size_t)kpad * 32 * g.width / 4;

    // Everything the multiply kernel reads from the scratch is rewritten by the three set-up kernels
    // below (every table entry, including the all-zero ones of padded fragments and unused rows), so
    // whatever an earlier call with other (k, p) left behind cannot leak into this one.
    cudaError_t e = launch_build_gf_tables(gf.d_exp, gf.d_log, s);
    if (e == cudaSuccess) e = launch_gen_total_matrix(d_total, k, p, gf, s);
    if (e == cudaSuccess) e = launch_build_product_tables(d_total + (size_t)k * k, k, kpad, 0, p, g.width, g.d_tab, gf, s);
    if (e != cudaSuccess) return fail_cuda(e, "rsgpu_encode_noalloc set-up kernels");
    Tuning tune;
    e = launch_matmul_group(g, k, kpad, data_d, data_pitch, parity_d, parity_pitch, len, tune, info, s, nullptr);
    if (e != cudaSuccess) return fail_cuda(e, "rsgpu_encode_noalloc multiply kernel");
    return RSGPU_OK;
}
