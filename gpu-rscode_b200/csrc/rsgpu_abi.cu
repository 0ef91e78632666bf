// rsgpu_abi.cu -- the extern "C" boundary of librsgpu.so (declared in include/rsgpu.h).
//
// Context management, plan (product table) construction and the device-pointer entry points.
// The host-memory pipeline lives in pipeline.cu.  No torch, no CPU fallback: if CUDA is not usable
// every entry point fails with RSGPU_ECUDA.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>

#include "../../include/rsgpu.h"
#include "rs_ctx.h"

using namespace rsgpu;

namespace rsgpu {

static thread_local char g_err[512] = "";

int fail_cuda(cudaError_t e, const char *what)
{
    snprintf(g_err, sizeof g_err, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
    return RSGPU_ECUDA;
}

int fail_msg(int code, const char *msg)
{
    snprintf(g_err, sizeof g_err, "%s", msg);
    return code;
}

}  // namespace rsgpu

cudaError_t UseTracker::create()
{
    for (auto &e : ev) {
        cudaError_t r = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        if (r != cudaSuccess) return r;
    }
    return cudaSuccess;
}

void UseTracker::destroy()
{
    for (auto &e : ev)
        if (e) {
            cudaEventDestroy(e);
            e = nullptr;
        }
    n = 0;
}

cudaError_t UseTracker::mark(cudaStream_t s)
{
    for (int i = 0; i < n; ++i)
        if (stream[i] == s) return cudaEventRecord(ev[i], s);
    if (n < kSlots) {
        stream[n] = s;
        return cudaEventRecord(ev[n++], s);
    }
    // more streams than slots: s takes over the oldest slot after waiting for what that slot covered, so the
    // slot's new event still stands for both
    const int i = victim;
    victim = (victim + 1) % kSlots;
    cudaError_t r = cudaStreamWaitEvent(s, ev[i], 0);
    if (r != cudaSuccess) return r;
    stream[i] = s;
    return cudaEventRecord(ev[i], s);
}

cudaError_t UseTracker::wait_all(cudaStream_t writer)
{
    for (int i = 0; i < n; ++i) {
        cudaError_t r = cudaStreamWaitEvent(writer, ev[i], 0);
        if (r != cudaSuccess) return r;
    }
    return cudaSuccess;
}

namespace rsgpu {

#define CK(call)                                                   \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return fail_cuda(e__, #call);      \
    } while (0)

void plan_free(Plan &pl)
{
    if (pl.d_matrix) cudaFree(pl.d_matrix);
    if (pl.d_tabs) cudaFree(pl.d_tabs);
    pl = Plan();
}

// Does peeling an all-ones row 0 pay?  Only when it lowers the width class of a launch: rows % 8 == 5 turns
// the 64-bit-entry group of five rows into four looked-up rows on 32-bit entries (one shared-memory wavefront per
// lookup instead of two) plus the XOR row.  For <= 4 rows it would only add instructions.
static bool peel_pays(int rows) { return rows % 8 == 5; }

// Allocate a plan for `rows` x k; rows are cut into groups of 8 (64-bit table entries) with a last
// group of <= 4 rows using 32-bit entries.  peel: row 0 is all ones (the caller knows) and is left out of
// the tables; group 0 then starts at row 1 and its launch also writes row 0 as the XOR of the inputs.
int plan_alloc(Plan &pl, int rows, int k, bool peel = false)
{
    plan_free(pl);
    pl.rows = rows;
    pl.k = k;
    pl.kpad = (k + 7) / 8 * 8;
    pl.ngroups = 0;
    pl.peeled = peel && peel_pays(rows);
    size_t words = 0;
    // with a peeled row the SHORT group comes first (it is the one with 32-bit entries, which the XOR row needs)
    for (int r0 = pl.peeled ? 1 : 0; r0 < rows;) {
        RowGroup &g = pl.grp[pl.ngroups++];
        if (pl.peeled && r0 == 1) {
            g.row0 = 1;
            g.nrows = 4;
            g.width = 4;
            g.xor_row = 0;
            g.tab_words = (size_t)pl.kpad * 32 * g.width / 4;
            words += g.tab_words;
            r0 += g.nrows;
            continue;
        }
        const int left = rows - r0;
        g.row0 = r0;
        g.nrows = left > 8 ? 8 : left;
        g.width = g.nrows > 4 ? 8 : 4;
        g.tab_words = (size_t)pl.kpad * 32 * g.width / 4;
        words += g.tab_words;
        r0 += g.nrows;
    }
    if (peel && rows == 1) pl.grp[0].all_ones = true;   // RS(k, k+1): the one parity row is the XOR of the fragments
    pl.tabs_bytes = words * 4;
    CK(cudaMalloc(&pl.d_matrix, (size_t)rows * k));
    CK(cudaMalloc(&pl.d_tabs, pl.tabs_bytes));
    size_t off = 0;
    for (int g = 0; g < pl.ngroups; ++g) {
        pl.grp[g].d_tab = pl.d_tabs + off;
        off += pl.grp[g].tab_words;
    }
    return RSGPU_OK;
}

// d_matrix must already hold the rows x k matrix (device); builds every group's tables on `s`.
int plan_build_tables(rsgpu_ctx *c, Plan &pl, cudaStream_t s)
{
    for (int g = 0; g < pl.ngroups; ++g) {
        const RowGroup &rg = pl.grp[g];
        CK(launch_build_product_tables(pl.d_matrix, pl.k, pl.kpad, rg.row0, rg.nrows, rg.width, rg.d_tab, c->gf, s));
        ++c->launches;
    }
    pl.valid = true;
    return RSGPU_OK;
}

int plan_apply(rsgpu_ctx *c, const Plan &pl, const uint8_t *in, size_t in_pitch, uint8_t *out, size_t out_pitch,
               size_t len, cudaStream_t s)
{
    if (!pl.valid) return fail_msg(RSGPU_ESTATE, "coding plan not built");
    for (int g = 0; g < pl.ngroups; ++g)
        CK(launch_matmul_group(pl.grp[g], pl.k, pl.kpad, in, in_pitch, out, out_pitch, len, c->tune, c->dev, s,
                               &c->launches));
    return RSGPU_OK;
}

// A batch of stripes through one plan.  One launch per row group when the batch kernel applies (16-byte
// aligned geometry, an instance for k, enough bytes to pay for the table build); otherwise stripe by stripe.
int plan_apply_batch(rsgpu_ctx *c, const Plan &pl, const uint8_t *in, size_t in_pitch, size_t in_stripe, uint8_t *out,
                     size_t out_pitch, size_t out_stripe, size_t len, size_t n_stripes, cudaStream_t s)
{
    if (!pl.valid) return fail_msg(RSGPU_ESTATE, "coding plan not built");
    if (len == 0 || n_stripes == 0) return RSGPU_OK;
    // bases and pitches 16-byte aligned; len itself may be ragged: its last (len % 16) bytes per stripe go to one
    // extra byte-granular launch over the whole batch
    const bool aligned = (((uintptr_t)in | (uintptr_t)out | in_pitch | out_pitch | in_stripe | out_stripe) & 15u) == 0;
    const bool want = n_stripes > 1 && aligned && len >= 16 && (c->tune.variant == 0 || c->tune.variant == 5) &&
                      (c->tune.variant == 5 || len * n_stripes >= ((size_t)2 << 20));
    if (want) {
        bool done = true;
        const size_t nvec = len / 16;
        const unsigned rem = (unsigned)(len - nvec * 16);
        for (int g = 0; g < pl.ngroups && done; ++g) {
            const RowGroup &rg = pl.grp[g];
            uint8_t *out_g = out + (size_t)rg.row0 * out_pitch;
            cudaError_t e = launch_matmul_fat_batch(rg, pl.k, in, in_pitch, in_stripe, out_g, out_pitch, out_stripe, nvec, n_stripes,
                                                    c->dev, c->tune.max_blocks, s);
            if (e == cudaErrorNotSupported) {
                if (g != 0) return fail_msg(RSGPU_ECUDA, "batch kernel support differs between row groups");
                done = false;   // no instance for this k: fall through to the per-stripe launches
                break;
            }
            if (e != cudaSuccess) return fail_cuda(e, "launch_matmul_fat_batch");
            ++c->launches;
            if (rem) {
                e = launch_matmul_bytes_batch(rg, pl.k, in, in_pitch, in_stripe, out_g, out_pitch, out_stripe, nvec * 16, rem,
                                              n_stripes, c->dev, s);
                if (e != cudaSuccess) return fail_cuda(e, "launch_matmul_bytes_batch");
                ++c->launches;
            }
        }
        if (done) return RSGPU_OK;
    }
    for (size_t st = 0; st < n_stripes; ++st) {
        int rc = plan_apply(c, pl, in + st * in_stripe, in_pitch, out + st * out_stripe, out_pitch, len, s);
        if (rc != RSGPU_OK) return rc;
    }
    return RSGPU_OK;
}

// Tables of the full k x k plan of the current decode set, built the first time they are needed -- on the
// context's set-up stream, with an event behind them that EVERY caller's stream waits on: a second
// rsgpu_decode on another stream must not start reading tables whose build is still queued on the first.
int ensure_full_tables(rsgpu_ctx *c, cudaStream_t s)
{
    rsgpu_ctx::DecodeSet &ds = c->dec[c->dec_cur];
    Plan &pl = ds.full;
    if (!pl.valid) {
        int rc = plan_build_tables(c, pl, c->setup_stream);
        if (rc != RSGPU_OK) return rc;
        CK(cudaEventRecord(ds.full_ready, c->setup_stream));
    }
    if (s != c->setup_stream) CK(cudaStreamWaitEvent(s, ds.full_ready, 0));
    return RSGPU_OK;
}

int ctx_bind(const rsgpu_ctx *c)
{
    CK(cudaSetDevice(c->device));
    return RSGPU_OK;
}

static int ctx_create(rsgpu_ctx **out, int device, int k, int p, const uint8_t *total, int generator = RSGPU_GEN_VANDERMONDE)
{
    if (!out) return fail_msg(RSGPU_EINVAL, "out is NULL");
    *out = nullptr;
    if (k < 1 || k > RSGPU_MAX_K || p < 0 || p > 255) return fail_msg(RSGPU_EINVAL, "need 1 <= k <= 255, 0 <= p <= 255");
    if (generator != RSGPU_GEN_VANDERMONDE && generator != RSGPU_GEN_CAUCHY) return fail_msg(RSGPU_EINVAL, "unknown generator kind");
    if (generator == RSGPU_GEN_CAUCHY && k + p > 256) return fail_msg(RSGPU_EINVAL, "a Cauchy generator needs k + p <= 256");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail_msg(RSGPU_EINVAL, "no such CUDA device");
    CK(cudaSetDevice(device));

    rsgpu_ctx *c = new (std::nothrow) rsgpu_ctx();
    if (!c) return fail_msg(RSGPU_ENOMEM, "host allocation failed");
    c->device = device;
    c->k = k;
    c->p = p;
    int rc = RSGPU_OK;
    auto bail = [&](int code) { rsgpu_destroy(c); return code; };

    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return bail(fail_msg(RSGPU_ECUDA, "cudaGetDeviceProperties failed"));
    c->dev.sm_count = prop.multiProcessorCount;
    c->dev.max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    c->dev.memo = &c->memo;

    cudaError_t e;
    if ((e = cudaStreamCreateWithFlags(&c->setup_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(fail_cuda(e, "cudaStreamCreate"));
    uint8_t *d_gf = nullptr;
    if ((e = cudaMalloc(&d_gf, kExpSize + kLogSize)) != cudaSuccess) return bail(fail_cuda(e, "cudaMalloc gf tables"));
    c->gf.d_exp = d_gf;
    c->gf.d_log = d_gf + kExpSize;
    if ((e = cudaMalloc(&c->d_total, (size_t)(k + p) * k)) != cudaSuccess) return bail(fail_cuda(e, "cudaMalloc total matrix"));
    if ((e = cudaMalloc(&c->d_sub, (size_t)k * k)) != cudaSuccess) return bail(fail_cuda(e, "cudaMalloc sub matrix"));
    if ((e = cudaMalloc(&c->d_rows, sizeof(int) * 256)) != cudaSuccess) return bail(fail_cuda(e, "cudaMalloc row list"));
    if ((e = cudaMalloc(&c->d_flag, sizeof(int))) != cudaSuccess) return bail(fail_cuda(e, "cudaMalloc flag"));

    for (cudaEvent_t *ev : {&c->dec[0].full_ready, &c->dec[1].full_ready})
        if ((e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming)) != cudaSuccess) return bail(fail_cuda(e, "cudaEventCreate"));
    for (UseTracker *t : {&c->dec[0].readers, &c->dec[1].readers, &c->repair_readers})
        if ((e = t->create()) != cudaSuccess) return bail(fail_cuda(e, "cudaEventCreate"));

    cudaStream_t s = c->setup_stream;
    if ((e = launch_build_gf_tables(c->gf.d_exp, c->gf.d_log, s)) != cudaSuccess) return bail(fail_cuda(e, "build_gf_tables"));
    ++c->launches;
    if (total) {
        if ((e = cudaMemcpyAsync(c->d_total, total, (size_t)(k + p) * k, cudaMemcpyHostToDevice, s)) != cudaSuccess)
            return bail(fail_cuda(e, "copy total matrix"));
    } else {
        e = generator == RSGPU_GEN_CAUCHY ? launch_gen_total_matrix_cauchy(c->d_total, k, p, c->gf, s)
                                          : launch_gen_total_matrix(c->d_total, k, p, c->gf, s);
        if (e != cudaSuccess) return bail(fail_cuda(e, "gen_total_matrix"));
        ++c->launches;
    }
    if (p > 0) {
        // Row 0 of the generator all ones?  Always for the reference's Vandermonde matrix ((j+1)^0, src/matrix.cu:758);
        // read off a caller-supplied matrix; never for Cauchy.
        bool ones0 = !total && generator == RSGPU_GEN_VANDERMONDE;
        if (total) {
            ones0 = true;
            for (int j = 0; j < k; ++j) ones0 = ones0 && total[(size_t)k * k + j] == 1;
        }
        if ((rc = plan_alloc(c->enc, p, k, ones0)) != RSGPU_OK) return bail(rc);
        if ((e = cudaMemcpyAsync(c->enc.d_matrix, c->d_total + (size_t)k * k, (size_t)p * k, cudaMemcpyDeviceToDevice, s)) != cudaSuccess)
            return bail(fail_cuda(e, "copy generator"));
        if ((rc = plan_build_tables(c, c->enc, s)) != RSGPU_OK) return bail(rc);
    }
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return bail(fail_cuda(e, "context set-up"));
    *out = c;
    return RSGPU_OK;
}

}  // namespace rsgpu

extern "C" {

const char *rsgpu_strerror(int code)
{
    switch (code) {
    case RSGPU_OK: return "ok";
    case RSGPU_EINVAL: return "invalid argument";
    case RSGPU_ECUDA: return "CUDA error";
    case RSGPU_ESINGULAR: return "survivor matrix is singular";
    case RSGPU_ENOMEM: return "out of memory";
    case RSGPU_ESTATE: return "call sequence error";
    case RSGPU_ETOOBIG: return "capacity exceeded";
    default: return "unknown error";
    }
}

const char *rsgpu_last_error(void) { return g_err; }

int rsgpu_device_count(int *count)
{
    if (!count) return fail_msg(RSGPU_EINVAL, "count is NULL");
    *count = 0;
    CK(cudaGetDeviceCount(count));
    return RSGPU_OK;
}

int rsgpu_device_max_grid(int device, int *max_grid)
{
    if (!max_grid) return fail_msg(RSGPU_EINVAL, "max_grid is NULL");
    int gx = 0, gy = 0;
    CK(cudaDeviceGetAttribute(&gx, cudaDevAttrMaxGridDimX, device));
    CK(cudaDeviceGetAttribute(&gy, cudaDevAttrMaxGridDimY, device));
    *max_grid = gx < gy ? gx : gy;
    return RSGPU_OK;
}

int rsgpu_create(rsgpu_ctx **out, int device, int k, int p) { return ctx_create(out, device, k, p, nullptr); }

int rsgpu_create_with_matrix(rsgpu_ctx **out, int device, int k, int p, const uint8_t *total_matrix)
{
    if (!total_matrix) return fail_msg(RSGPU_EINVAL, "total_matrix is NULL");
    return ctx_create(out, device, k, p, total_matrix);
}

int rsgpu_create_generator(rsgpu_ctx **out, int device, int k, int p, int generator)
{
    return ctx_create(out, device, k, p, nullptr, generator);
}

void rsgpu_destroy(rsgpu_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    pipeline_free(c);
    plan_free(c->enc);
    for (auto &ds : c->dec) {
        plan_free(ds.full);
        plan_free(ds.erased);
        ds.readers.destroy();
        if (ds.full_ready) cudaEventDestroy(ds.full_ready);
    }
    c->repair_readers.destroy();
    if (c->d_inv_scratch) cudaFree(c->d_inv_scratch);
    plan_free(c->scratch);
    plan_free(c->repair);
    if (c->gf.d_exp) cudaFree(c->gf.d_exp);
    if (c->d_total) cudaFree(c->d_total);
    if (c->d_sub) cudaFree(c->d_sub);
    if (c->d_rows) cudaFree(c->d_rows);
    if (c->d_flag) cudaFree(c->d_flag);
    if (c->setup_stream) cudaStreamDestroy(c->setup_stream);
    delete c;
}

int rsgpu_k(const rsgpu_ctx *c) { return c ? c->k : RSGPU_EINVAL; }
int rsgpu_p(const rsgpu_ctx *c) { return c ? c->p : RSGPU_EINVAL; }
int rsgpu_device(const rsgpu_ctx *c) { return c ? c->device : RSGPU_EINVAL; }
long long rsgpu_launch_count(const rsgpu_ctx *c) { return c ? c->launches : 0; }

int rsgpu_get_total_matrix(rsgpu_ctx *c, uint8_t *total)
{
    if (!c || !total) return fail_msg(RSGPU_EINVAL, "NULL argument");
    int rc = ctx_bind(c);
    if (rc) return rc;
    CK(cudaMemcpy(total, c->d_total, (size_t)(c->k + c->p) * c->k, cudaMemcpyDeviceToHost));
    return RSGPU_OK;
}

int rsgpu_encode(rsgpu_ctx *c, const uint8_t *data_d, size_t data_pitch, uint8_t *parity_d, size_t parity_pitch,
                 size_t len, void *stream)
{
    if (!c || !data_d || !parity_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (c->p == 0) return RSGPU_OK;
    if ((c->k > 1 && data_pitch < len) || (c->p > 1 && parity_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    return plan_apply(c, c->enc, data_d, data_pitch, parity_d, parity_pitch, len, (cudaStream_t)stream);
}

int rsgpu_encode_batch(rsgpu_ctx *c, const uint8_t *data_d, size_t data_pitch, size_t data_stripe_pitch, uint8_t *parity_d,
                       size_t parity_pitch, size_t parity_stripe_pitch, size_t len, size_t n_stripes, void *stream)
{
    if (!c || !data_d || !parity_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (c->p == 0) return RSGPU_OK;
    if ((c->k > 1 && data_pitch < len) || (c->p > 1 && parity_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    if (n_stripes > 1 && (data_stripe_pitch < len || parity_stripe_pitch < len)) return fail_msg(RSGPU_EINVAL, "stripe pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    return plan_apply_batch(c, c->enc, data_d, data_pitch, data_stripe_pitch, parity_d, parity_pitch, parity_stripe_pitch, len,
                            n_stripes, (cudaStream_t)stream);
}

int rsgpu_enable_peer(rsgpu_ctx *c, int peer_device)
{
    if (!c) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (peer_device == c->device) return RSGPU_OK;
    int rc = ctx_bind(c);
    if (rc) return rc;
    int can = 0;
    CK(cudaDeviceCanAccessPeer(&can, c->device, peer_device));
    if (!can) return fail_msg(RSGPU_EINVAL, "this device cannot access the peer's memory");
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();   // clear the sticky-less error state
        return RSGPU_OK;
    }
    if (e != cudaSuccess) return fail_cuda(e, "cudaDeviceEnablePeerAccess");
    return RSGPU_OK;
}

int rsgpu_encode_rows(rsgpu_ctx *c, const uint8_t *data_d, size_t data_pitch, uint8_t *const *parity_rows, size_t len,
                      void *stream)
{
    if (!c || !data_d || !parity_rows) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (c->p == 0) return RSGPU_OK;
    for (int i = 0; i < c->p; ++i)
        if (!parity_rows[i]) return fail_msg(RSGPU_EINVAL, "NULL row pointer");
    if (c->k > 1 && data_pitch < len) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if (len == 0) return RSGPU_OK;
    for (int g = 0; g < c->enc.ngroups; ++g)
        CK(launch_matmul_rows(c->enc.grp[g], c->k, c->enc.kpad, data_d, data_pitch, parity_rows, len, c->dev,
                              (cudaStream_t)stream, &c->launches));
    return RSGPU_OK;
}

int rsgpu_set_survivors(rsgpu_ctx *c, const int *idx)
{
    if (!c || !idx) return fail_msg(RSGPU_EINVAL, "NULL argument");
    const int k = c->k, n = c->k + c->p;
    bool have[256] = {false};
    for (int i = 0; i < k; ++i) {
        if (idx[i] < 0 || idx[i] >= n) return fail_msg(RSGPU_EINVAL, "fragment index out of range");
        if (idx[i] < k) have[idx[i]] = true;
    }
    int rc = ctx_bind(c);
    if (rc) return rc;
    // Build into the set that is not current; a failure below leaves the context without survivors.
    c->have_survivors = false;
    rsgpu_ctx::DecodeSet &ds = c->dec[c->dec_cur ^ 1];
    ds.full.valid = false;
    ds.erased.valid = false;
    c->n_erased = 0;
    for (int j = 0; j < k; ++j)
        if (!have[j]) c->erased[c->n_erased++] = j;
    memcpy(c->survivors, idx, sizeof(int) * k);

    cudaStream_t s = c->setup_stream;
    CK(ds.readers.wait_all(s));   // every decode that still reads this set's tables, on whatever stream, finishes first
    if (!ds.full.d_matrix && (rc = plan_alloc(ds.full, k, k)) != RSGPU_OK) return rc;
    if (c->n_erased > 0 && (ds.erased.rows != c->n_erased || !ds.erased.d_matrix))
        if ((rc = plan_alloc(ds.erased, c->n_erased, k)) != RSGPU_OK) return rc;
    // one launch: gather the survivor rows, invert, tables of the k x k plan, rows of the inverse that rebuild the
    // erased natives and their tables (the index lists travel as kernel parameters)
    // the tables of the full k x k plan are only needed by the reference-semantic rsgpu_decode[_host]: they are
    // built on its first use (ensure_full_tables), not here
    CK(launch_decode_setup(c->d_total, k, ds.full.kpad, idx, c->erased, c->n_erased, ds.full.d_matrix, nullptr,
                           c->n_erased > 0 ? ds.erased.d_matrix : nullptr, c->n_erased > 0 ? ds.erased.d_tabs : nullptr,
                           c->d_flag, c->gf, s));
    ++c->launches;
    int flag = 0;
    uint8_t one_row[256];
    CK(cudaMemcpyAsync(&flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
    // one native lost: if the row that rebuilds it is all ones (parity row 0 among the survivors, the others natives) the
    // decode is a plain XOR of the survivors -- look at the row while we wait for the flag anyway
    if (c->n_erased == 1) CK(cudaMemcpyAsync(one_row, ds.erased.d_matrix, (size_t)k, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    if (c->n_erased > 0) {
        bool ones = c->n_erased == 1 && !flag;
        for (int j = 0; ones && j < k; ++j) ones = one_row[j] == 1;
        ds.erased.grp[0].all_ones = ones;
    }
    ds.readers.reset();                           // the set-up stream has drained: nobody reads the old tables any more
    if (flag) return fail_msg(RSGPU_ESINGULAR, "the chosen survivors give a singular matrix");
    ds.full.valid = false;                        // matrix there, tables not yet
    if (c->n_erased > 0) ds.erased.valid = true;
    c->dec_cur ^= 1;
    c->have_survivors = true;
    return RSGPU_OK;
}

// After a kernel that reads the current decode set has been queued on `s`: remember it for the ordering
// in rsgpu_set_survivors.
static int mark_decode_use(rsgpu_ctx *c, cudaStream_t s)
{
    CK(c->dec[c->dec_cur].readers.mark(s));
    return RSGPU_OK;
}

int rsgpu_set_repair(rsgpu_ctx *c, const int *survivors, const int *wanted, int n_wanted)
{
    if (!c || !survivors || !wanted) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (n_wanted < 1 || n_wanted > 255) return fail_msg(RSGPU_EINVAL, "need 1 <= n_wanted <= 255");
    for (int i = 0; i < n_wanted; ++i)
        if (wanted[i] < 0 || wanted[i] >= c->k + c->p) return fail_msg(RSGPU_EINVAL, "wanted fragment index out of range");
    c->have_repair = false;
    int rc = rsgpu_set_survivors(c, survivors);   // inverse of the survivor rows (may be RSGPU_ESINGULAR)
    if (rc != RSGPU_OK) return rc;
    const int k = c->k;
    cudaStream_t s = c->setup_stream;
    CK(c->repair_readers.wait_all(s));   // every rsgpu_repair still in flight has finished with the old tables
    if (c->repair.rows != n_wanted || !c->repair.d_matrix)
        if ((rc = plan_alloc(c->repair, n_wanted, k)) != RSGPU_OK) return rc;
    // repair matrix = (rows `wanted` of the total matrix) x inverse: natives come out as rows of the
    // inverse, parities as generator rows applied to the reconstructed natives, in one multiply
    CK(cudaMemcpyAsync(c->d_rows, wanted, sizeof(int) * n_wanted, cudaMemcpyHostToDevice, s));
    CK(launch_rows_times_matrix(c->d_total, c->d_rows, n_wanted, c->dec_full().d_matrix, k, c->repair.d_matrix, c->gf, s));
    ++c->launches;
    if ((rc = plan_build_tables(c, c->repair, s)) != RSGPU_OK) return rc;
    uint8_t one_row[256];
    if (n_wanted == 1) CK(cudaMemcpyAsync(one_row, c->repair.d_matrix, (size_t)k, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    {   // one wanted fragment whose repair row is all ones (parity 0 from the natives, a native from the rest + parity 0): XOR
        bool ones = n_wanted == 1;
        for (int j = 0; ones && j < k; ++j) ones = one_row[j] == 1;
        c->repair.grp[0].all_ones = ones;
    }
    c->repair_readers.reset();
    c->have_repair = true;
    return RSGPU_OK;
}

int rsgpu_repair(rsgpu_ctx *c, const uint8_t *surv_d, size_t surv_pitch, uint8_t *out_d, size_t out_pitch, size_t len,
                 void *stream)
{
    if (!c || !surv_d || !out_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_repair) return fail_msg(RSGPU_ESTATE, "rsgpu_set_repair has not succeeded yet");
    if ((c->k > 1 && surv_pitch < len) || (c->repair.rows > 1 && out_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if ((rc = plan_apply(c, c->repair, surv_d, surv_pitch, out_d, out_pitch, len, (cudaStream_t)stream)) != RSGPU_OK) return rc;
    CK(c->repair_readers.mark((cudaStream_t)stream));
    return RSGPU_OK;
}

int rsgpu_get_decode_matrix(rsgpu_ctx *c, uint8_t *inverse)
{
    if (!c || !inverse) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_survivors) return fail_msg(RSGPU_ESTATE, "rsgpu_set_survivors has not succeeded yet");
    int rc = ctx_bind(c);
    if (rc) return rc;
    CK(cudaMemcpy(inverse, c->dec_full().d_matrix, (size_t)c->k * c->k, cudaMemcpyDeviceToHost));
    return RSGPU_OK;
}

int rsgpu_get_erased(rsgpu_ctx *c, int *idx, int *count)
{
    if (!c || !idx || !count) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_survivors) return fail_msg(RSGPU_ESTATE, "rsgpu_set_survivors has not succeeded yet");
    *count = c->n_erased;
    memcpy(idx, c->erased, sizeof(int) * c->n_erased);
    return RSGPU_OK;
}

int rsgpu_decode(rsgpu_ctx *c, const uint8_t *surv_d, size_t surv_pitch, uint8_t *nat_d, size_t nat_pitch,
                 size_t len, void *stream)
{
    if (!c || !surv_d || !nat_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_survivors) return fail_msg(RSGPU_ESTATE, "rsgpu_set_survivors has not succeeded yet");
    if (c->k > 1 && (surv_pitch < len || nat_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if ((rc = ensure_full_tables(c, (cudaStream_t)stream)) != RSGPU_OK) return rc;
    if ((rc = plan_apply(c, c->dec_full(), surv_d, surv_pitch, nat_d, nat_pitch, len, (cudaStream_t)stream)) != RSGPU_OK) return rc;
    return mark_decode_use(c, (cudaStream_t)stream);
}

int rsgpu_decode_erased(rsgpu_ctx *c, const uint8_t *surv_d, size_t surv_pitch, uint8_t *out_d, size_t out_pitch,
                        size_t len, void *stream)
{
    if (!c || !surv_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_survivors) return fail_msg(RSGPU_ESTATE, "rsgpu_set_survivors has not succeeded yet");
    if (c->n_erased == 0) return RSGPU_OK;
    if (!out_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if ((c->k > 1 && surv_pitch < len) || (c->n_erased > 1 && out_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if ((rc = plan_apply(c, c->dec_erased(), surv_d, surv_pitch, out_d, out_pitch, len, (cudaStream_t)stream)) != RSGPU_OK) return rc;
    return mark_decode_use(c, (cudaStream_t)stream);
}

int rsgpu_decode_erased_batch(rsgpu_ctx *c, const uint8_t *surv_d, size_t surv_pitch, size_t surv_stripe_pitch, uint8_t *out_d,
                              size_t out_pitch, size_t out_stripe_pitch, size_t len, size_t n_stripes, void *stream)
{
    if (!c || !surv_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_survivors) return fail_msg(RSGPU_ESTATE, "rsgpu_set_survivors has not succeeded yet");
    if (c->n_erased == 0) return RSGPU_OK;
    if (!out_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if ((c->k > 1 && surv_pitch < len) || (c->n_erased > 1 && out_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    if (n_stripes > 1 && (surv_stripe_pitch < len || out_stripe_pitch < len)) return fail_msg(RSGPU_EINVAL, "stripe pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if ((rc = plan_apply_batch(c, c->dec_erased(), surv_d, surv_pitch, surv_stripe_pitch, out_d, out_pitch, out_stripe_pitch, len,
                               n_stripes, (cudaStream_t)stream)) != RSGPU_OK)
        return rc;
    return mark_decode_use(c, (cudaStream_t)stream);
}

int rsgpu_invert_matrix(rsgpu_ctx *c, const uint8_t *matrix, uint8_t *inverse, int n)
{
    if (!c || !matrix || !inverse) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (n < 1 || n > 255) return fail_msg(RSGPU_EINVAL, "need 1 <= n <= 255");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if (!c->d_inv_scratch) CK(cudaMalloc(&c->d_inv_scratch, (size_t)2 * 255 * 255));   // once per context, kept
    uint8_t *d = c->d_inv_scratch;
    cudaStream_t s = c->setup_stream;
    int flag = 0;
    cudaError_t e = cudaMemcpyAsync(d, matrix, (size_t)n * n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = launch_gj_invert(d, d + (size_t)n * n, n, c->d_flag, c->gf, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(inverse, d + (size_t)n * n, (size_t)n * n, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return fail_cuda(e, "rsgpu_invert_matrix");
    ++c->launches;
    if (flag) return fail_msg(RSGPU_ESINGULAR, "matrix is singular");
    return RSGPU_OK;
}

int rsgpu_matmul(rsgpu_ctx *c, const uint8_t *matrix, int rows, const uint8_t *in_d, size_t in_pitch,
                 uint8_t *out_d, size_t out_pitch, size_t len, void *stream)
{
    if (!c || !matrix || !in_d || !out_d) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (rows < 1 || rows > 255) return fail_msg(RSGPU_EINVAL, "need 1 <= rows <= 255");
    if ((c->k > 1 && in_pitch < len) || (rows > 1 && out_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if (c->scratch.rows != rows || !c->scratch.d_matrix)
        if ((rc = plan_alloc(c->scratch, rows, c->k)) != RSGPU_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaMemcpyAsync(c->scratch.d_matrix, matrix, (size_t)rows * c->k, cudaMemcpyHostToDevice, s));
    if ((rc = plan_build_tables(c, c->scratch, s)) != RSGPU_OK) return rc;
    if ((rc = plan_apply(c, c->scratch, in_d, in_pitch, out_d, out_pitch, len, s)) != RSGPU_OK) return rc;
    CK(cudaStreamSynchronize(s));
    return RSGPU_OK;
}

int rsgpu_host_alloc(void **ptr, size_t bytes)
{
    if (!ptr) return fail_msg(RSGPU_EINVAL, "NULL argument");
    *ptr = nullptr;
    cudaError_t e = cudaMallocHost(ptr, bytes ? bytes : 1);
    if (e == cudaErrorMemoryAllocation) return fail_msg(RSGPU_ENOMEM, "pinned allocation failed");
    if (e != cudaSuccess) return fail_cuda(e, "cudaMallocHost");
    return RSGPU_OK;
}

int rsgpu_host_free(void *ptr)
{
    if (ptr) CK(cudaFreeHost(ptr));
    return RSGPU_OK;
}

int rsgpu_host_register(void *ptr, size_t bytes)
{
    if (!ptr) return fail_msg(RSGPU_EINVAL, "NULL argument");
    CK(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return RSGPU_OK;
}

int rsgpu_host_unregister(void *ptr)
{
    if (ptr) CK(cudaHostUnregister(ptr));
    return RSGPU_OK;
}

int rsgpu_dev_alloc(rsgpu_ctx *c, void **ptr, size_t bytes)
{
    if (!c || !ptr) return fail_msg(RSGPU_EINVAL, "NULL argument");
    int rc = ctx_bind(c);
    if (rc) return rc;
    *ptr = nullptr;
    cudaError_t e = cudaMalloc(ptr, bytes ? bytes : 1);
    if (e == cudaErrorMemoryAllocation) return fail_msg(RSGPU_ENOMEM, "device allocation failed");
    if (e != cudaSuccess) return fail_cuda(e, "cudaMalloc");
    return RSGPU_OK;
}

int rsgpu_dev_free(rsgpu_ctx *c, void *ptr)
{
    if (!c) return fail_msg(RSGPU_EINVAL, "NULL argument");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if (ptr) CK(cudaFree(ptr));
    return RSGPU_OK;
}

int rsgpu_memcpy_h2d(rsgpu_ctx *c, void *dst_d, const void *src, size_t bytes)
{
    if (!c) return fail_msg(RSGPU_EINVAL, "NULL argument");
    int rc = ctx_bind(c);
    if (rc) return rc;
    CK(cudaMemcpy(dst_d, src, bytes, cudaMemcpyHostToDevice));
    return RSGPU_OK;
}

int rsgpu_memcpy_d2h(rsgpu_ctx *c, void *dst, const void *src_d, size_t bytes)
{
    if (!c) return fail_msg(RSGPU_EINVAL, "NULL argument");
    int rc = ctx_bind(c);
    if (rc) return rc;
    CK(cudaMemcpy(dst, src_d, bytes, cudaMemcpyDeviceToHost));
    return RSGPU_OK;
}

int rsgpu_sync(rsgpu_ctx *c)
{
    if (!c) return fail_msg(RSGPU_EINVAL, "NULL argument");
    int rc = ctx_bind(c);
    if (rc) return rc;
    CK(cudaDeviceSynchronize());
    return RSGPU_OK;
}

int rsgpu_set_tuning(rsgpu_ctx *c, int variant, int threads, int max_blocks)
{
    if (!c) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (threads < 0 || threads > 1024 || (threads % 32) != 0) return fail_msg(RSGPU_EINVAL, "threads must be a multiple of 32, <= 1024");
    if (variant < 0 || max_blocks < 0) return fail_msg(RSGPU_EINVAL, "negative tuning value");
    c->tune.variant = variant;
    c->tune.threads = threads;
    c->tune.max_blocks = max_blocks;
    return RSGPU_OK;
}

}  // extern "C"
