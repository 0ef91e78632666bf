// rs_ctx.h -- the context object behind the opaque rsgpu_ctx handle (internal).
#pragma once
#include "rs_internal.h"

constexpr int kMaxSlots = 16;

// Which streams have kernels in flight that read a set of tables.  One event per distinct stream (up to
// kSlots; a further stream is chained behind the oldest slot), so that whoever rebuilds the tables can wait
// for ALL of them, not only for the stream that happened to be used last.
struct UseTracker {
    static constexpr int kSlots = 4;
    cudaEvent_t ev[kSlots] = {};
    cudaStream_t stream[kSlots] = {};
    int n = 0;          // slots holding a recorded event
    int victim = 0;     // next slot to take over when all are taken by other streams
    cudaError_t create();
    void destroy();
    cudaError_t mark(cudaStream_t s);            // call after queueing a reader on s
    cudaError_t wait_all(cudaStream_t writer);   // writer waits for every recorded reader
    void reset() { n = 0; }                      // after the writer's stream has been synchronised
};

struct rsgpu_ctx {
    int device = 0;
    int k = 0, p = 0;
    rsgpu::DeviceInfo dev;
    rsgpu::LaunchMemo memo;       // dev.memo points here
    rsgpu::GfTables gf;
    rsgpu::Tuning tune;
    cudaStream_t setup_stream = nullptr;

    uint8_t *d_total = nullptr;   // (k+p) x k : [I ; V] (or caller supplied)
    uint8_t *d_sub = nullptr;     // k x k     : survivor rows, input of the inversion
    int *d_rows = nullptr;        // row-index list for the gather kernel
    int *d_flag = nullptr;        // singular flag written by the inversion kernel

    rsgpu::Plan enc;              // p x k generator
    // Decode plans come in two sets used alternately: rsgpu_set_survivors builds into the set that is NOT
    // current, ordered (cudaStreamWaitEvent) after the last kernel that read that set, and only then makes
    // it current -- a decode still in flight on the caller's stream keeps reading intact tables.
    struct DecodeSet {
        rsgpu::Plan full;         // k x k inverse (reference-semantic decode)
        rsgpu::Plan erased;       // e x k rows of the inverse for the erased natives
        UseTracker readers;            // decodes in flight on this set, per stream
        cudaEvent_t full_ready = nullptr;   // recorded on setup_stream behind the lazily built tables of `full`
    };
    DecodeSet dec[2];
    int dec_cur = 0;
    rsgpu::Plan &dec_full() { return dec[dec_cur].full; }
    rsgpu::Plan &dec_erased() { return dec[dec_cur].erased; }
    rsgpu::Plan scratch;          // rsgpu_matmul's caller-supplied matrix
    rsgpu::Plan repair;           // m x k: T[wanted] * inverse, rebuilds any wanted fragments
    bool have_repair = false;
    UseTracker repair_readers;    // same ordering for rsgpu_set_repair after rsgpu_repair
    uint8_t *d_inv_scratch = nullptr;   // rsgpu_invert_matrix: 2 x 255 x 255 bytes, allocated on its first call, kept

    bool have_survivors = false;
    int survivors[256];
    int erased[256];
    int n_erased = 0;

    long long launches = 0;

    // host-memory pipeline (pipeline.cu): per-slot stream and device staging buffers
    int n_slots = 0;
    cudaStream_t slot_stream[kMaxSlots] = {};
    uint8_t *slot_in[kMaxSlots] = {};
    uint8_t *slot_out[kMaxSlots] = {};
    size_t slot_pitch = 0;        // bytes per fragment row in a slot buffer
    int slot_in_rows = 0, slot_out_rows = 0;
};

namespace rsgpu {
int fail_cuda(cudaError_t e, const char *what);
int fail_msg(int code, const char *msg);
int ctx_bind(const rsgpu_ctx *c);
int plan_apply(rsgpu_ctx *c, const Plan &pl, const uint8_t *in, size_t in_pitch, uint8_t *out, size_t out_pitch,
               size_t len, cudaStream_t s);
void pipeline_free(rsgpu_ctx *c);
int ensure_full_tables(rsgpu_ctx *c, cudaStream_t s);
}  // namespace rsgpu
