// pipeline.cu -- host-memory coding through a per-device ring of stream slots.
//
// Replaces the per-device encode()/decode() of the reference (src/encode.cu:109-238,
// src/decode.cu:89-196): there every call cudaMallocs per-stream buffers, issues k row copies, launches
// a kernel whose wrapper blocks on cudaEventSynchronize (src/matrix.cu:808-809, which serialises the
// "streams"), copies back and cudaFrees -- all inside the timed region.  Here the slot buffers and
// streams belong to the context (allocated once, reused), a slice is ONE strided 2-D copy in, one
// kernel, one strided 2-D copy out, all asynchronous on the slot's stream, and the only host wait is
// at the very end.  Slices on different slots overlap: H2D of slice i+1, the kernel of slice i and
// D2H of slice i-1 run on the two copy engines and the SMs at the same time.
#include <stdlib.h>

#include "../../include/rsgpu.h"
#include "rs_ctx.h"

namespace rsgpu {

#define CK(call)                                                   \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return fail_cuda(e__, #call);      \
    } while (0)

void pipeline_free(rsgpu_ctx *c)
{
    for (int i = 0; i < kMaxSlots; ++i) {
        if (c->slot_in[i]) cudaFree(c->slot_in[i]);
        if (c->slot_out[i]) cudaFree(c->slot_out[i]);
        if (c->slot_stream[i]) cudaStreamDestroy(c->slot_stream[i]);
        c->slot_in[i] = c->slot_out[i] = nullptr;
        c->slot_stream[i] = nullptr;
    }
    c->n_slots = 0;
    c->slot_pitch = 0;
    c->slot_in_rows = c->slot_out_rows = 0;
}

static int pipeline_ensure(rsgpu_ctx *c, int slots, size_t pitch, int in_rows, int out_rows)
{
    if (c->n_slots >= slots && c->slot_pitch >= pitch && c->slot_in_rows >= in_rows && c->slot_out_rows >= out_rows)
        return RSGPU_OK;
    if (c->n_slots) {
        CK(cudaDeviceSynchronize());
        pipeline_free(c);
    }
    for (int i = 0; i < slots; ++i) {
        cudaError_t e = cudaStreamCreateWithFlags(&c->slot_stream[i], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(&c->slot_in[i], pitch * in_rows);
        if (e == cudaSuccess) e = cudaMalloc(&c->slot_out[i], pitch * out_rows);
        if (e != cudaSuccess) {
            pipeline_free(c);   // a half-built ring is of no use; the next call starts over
            if (e == cudaErrorMemoryAllocation) {
                cudaGetLastError();
                return fail_msg(RSGPU_ENOMEM, "device allocation for pipeline slots failed");
            }
            return fail_cuda(e, "pipeline slot set-up");
        }
    }
    c->n_slots = slots;
    c->slot_pitch = pitch;
    c->slot_in_rows = in_rows;
    c->slot_out_rows = out_rows;
    return RSGPU_OK;
}

// Slice length per fragment row: big enough that a row copy runs at PCIe speed, small enough that the
// ring fills quickly and the tail (last kernel + last D2H, not overlapped) stays short.
static size_t choose_slice(size_t len, int slots)
{
    size_t min_slice = (size_t)1 << 20, max_slice = (size_t)8 << 20;
    if (const char *e = getenv("RSGPU_PIPE_SLICE_MIB"))      // measurement knob: fixed slice length per row
        if (atoi(e) > 0) min_slice = max_slice = (size_t)atoi(e) << 20;
    size_t s = len / ((size_t)slots * 4);
    if (s < min_slice) s = min_slice;
    if (s > max_slice) s = max_slice;
    if (s > len) s = len;
    return (s + 255) / 256 * 256;
}

static int pipeline_run(rsgpu_ctx *c, const Plan &pl, const uint8_t *in, size_t in_pitch, uint8_t *out,
                        size_t out_pitch, size_t len, int slots)
{
    if (len == 0) return RSGPU_OK;
    if (slots <= 0) slots = 4;
    if (slots > kMaxSlots) slots = kMaxSlots;
    const size_t slice = choose_slice(len, slots);
    int rc = pipeline_ensure(c, slots, slice, pl.k, pl.rows);
    if (rc) return rc;
    const size_t pitch = c->slot_pitch;
    // measurement knob: one contiguous copy per fragment row instead of one strided 2-D copy per slice
    static const bool row_copies = getenv("RSGPU_PIPE_ROWCOPY") && atoi(getenv("RSGPU_PIPE_ROWCOPY")) > 0;
    int slot = 0;
    for (size_t off = 0; off < len && rc == RSGPU_OK; off += slice, slot = (slot + 1) % slots) {
        const size_t cur = len - off < slice ? len - off : slice;
        cudaStream_t s = c->slot_stream[slot];
        cudaError_t e = cudaSuccess;
        if (row_copies) {
            for (int j = 0; j < pl.k && e == cudaSuccess; ++j)
                e = cudaMemcpyAsync(c->slot_in[slot] + (size_t)j * pitch, in + (size_t)j * in_pitch + off, cur, cudaMemcpyHostToDevice, s);
        } else {
            e = cudaMemcpy2DAsync(c->slot_in[slot], pitch, in + off, in_pitch, cur, pl.k, cudaMemcpyHostToDevice, s);
        }
        if (e != cudaSuccess) {
            rc = fail_cuda(e, "host-to-device slice copy");
            break;
        }
        if ((rc = plan_apply(c, pl, c->slot_in[slot], pitch, c->slot_out[slot], pitch, cur, s)) != RSGPU_OK) break;
        if (row_copies) {
            for (int i = 0; i < pl.rows && e == cudaSuccess; ++i)
                e = cudaMemcpyAsync(out + (size_t)i * out_pitch + off, c->slot_out[slot] + (size_t)i * pitch, cur, cudaMemcpyDeviceToHost, s);
        } else {
            e = cudaMemcpy2DAsync(out + off, out_pitch, c->slot_out[slot], pitch, cur, pl.rows, cudaMemcpyDeviceToHost, s);
        }
        if (e != cudaSuccess) rc = fail_cuda(e, "device-to-host slice copy");
    }
    // drain every slot even after a failure: the caller's buffers must not be touched after we return
    for (int i = 0; i < slots; ++i) {
        cudaError_t e = cudaStreamSynchronize(c->slot_stream[i]);
        if (e != cudaSuccess && rc == RSGPU_OK) rc = fail_cuda(e, "pipeline drain");
    }
    return rc;
}

}  // namespace rsgpu

using namespace rsgpu;

extern "C" {

int rsgpu_encode_host(rsgpu_ctx *c, const uint8_t *data, size_t data_pitch, uint8_t *parity, size_t parity_pitch,
                      size_t len, int slots)
{
    if (!c || !data || !parity) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (c->p == 0) return RSGPU_OK;
    if ((c->k > 1 && data_pitch < len) || (c->p > 1 && parity_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    return pipeline_run(c, c->enc, data, data_pitch, parity, parity_pitch, len, slots);
}

int rsgpu_decode_host(rsgpu_ctx *c, const uint8_t *surv, size_t surv_pitch, uint8_t *natives, size_t natives_pitch,
                      size_t len, int slots)
{
    if (!c || !surv || !natives) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_survivors) return fail_msg(RSGPU_ESTATE, "rsgpu_set_survivors has not succeeded yet");
    if (c->k > 1 && (surv_pitch < len || natives_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    if ((rc = ensure_full_tables(c, c->setup_stream)) != RSGPU_OK) return rc;
    CK(cudaStreamSynchronize(c->setup_stream));     // the slot streams do not wait on the set-up stream by themselves
    return pipeline_run(c, c->dec_full(), surv, surv_pitch, natives, natives_pitch, len, slots);
}

int rsgpu_decode_erased_host(rsgpu_ctx *c, const uint8_t *surv, size_t surv_pitch, uint8_t *out, size_t out_pitch,
                             size_t len, int slots)
{
    if (!c || !surv) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if (!c->have_survivors) return fail_msg(RSGPU_ESTATE, "rsgpu_set_survivors has not succeeded yet");
    if (c->n_erased == 0) return RSGPU_OK;
    if (!out) return fail_msg(RSGPU_EINVAL, "NULL argument");
    if ((c->k > 1 && surv_pitch < len) || (c->n_erased > 1 && out_pitch < len)) return fail_msg(RSGPU_EINVAL, "pitch < len");
    int rc = ctx_bind(c);
    if (rc) return rc;
    return pipeline_run(c, c->dec_erased(), surv, surv_pitch, out, out_pitch, len, slots);
}

}  // extern "C"
