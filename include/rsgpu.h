/*
 * rsgpu.h -- thin C ABI of the B200-native Reed-Solomon GF(2^8) coding engine (librsgpu.so).
 *
 * Everything CUDA hides behind these entry points: plain pointers, sizes and int codes, no CUDA or
 * torch types.  The host layer (C: encode_file / decode_file / CLI, see encode.h, decode.h), the
 * Python tests and bench.py all go through exactly this boundary.
 *
 * Each entry point names the reference interface it replaces (paths relative to the reference
 * tree zvonkok/GPU-RSCode).  Conventions:
 *   - every function returns RSGPU_OK (0) or a negative RSGPU_E* code; nothing here calls exit().
 *   - "fragment-major": a stripe is rows of `len` bytes, row j starting at base + j*pitch.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - sizes are 64-bit (the reference's `int` sizes cap it below 2 GiB, src/encode.cu:303-322).
 *   - device-pointer calls are asynchronous on `stream`; host-pointer calls (rsgpu_*_host) block.
 *   - a context is bound to one device and is not thread-safe; use one per host thread / GPU.
 */
#ifndef RSGPU_H_
#define RSGPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RSGPU_OK          0
#define RSGPU_EINVAL     -1   /* bad argument (k, p, pointer, index out of range ...)            */
#define RSGPU_ECUDA      -2   /* a CUDA runtime call failed; see rsgpu_last_error()               */
#define RSGPU_ESINGULAR  -3   /* the k survivors do not determine the data (matrix is singular)   */
#define RSGPU_ENOMEM     -4   /* host or device allocation failed                                  */
#define RSGPU_ESTATE     -5   /* call sequence error, e.g. decode before rsgpu_set_survivors      */
#define RSGPU_ETOOBIG    -6   /* a fixed capacity of this entry point was exceeded                 */

#define RSGPU_MAX_K 255       /* (j+1) must stay a non-zero field element, src/matrix.cu:758      */

typedef struct rsgpu_ctx rsgpu_ctx;

/* Human-readable text for a return code; rsgpu_last_error() adds the CUDA detail (thread-local). */
const char *rsgpu_strerror(int code);
const char *rsgpu_last_error(void);

/* Number of visible CUDA devices.  Replaces cudaGetDeviceCount in src/encode.cu:357, decode.cu:335. */
int rsgpu_device_count(int *count);

/* Largest legal gridDim.x/y of `device` -- the bound the reference prints as "Valid grid size: (0, %d]"
 * (min of maxGridSize[0], maxGridSize[1], src/encode.cu:348-355). */
int rsgpu_device_max_grid(int device, int *max_grid);

/*
 * Create a coding context for RS(k, k+p) on `device`.
 * Builds on the GPU: the GF(2^8) exp/log tables (replaces setup_tables, src/matrix.cu:47-75), the
 * total encoding matrix [I_k ; V], V[i][j] = (j+1)^i (replaces the gen_encoding_matrix kernel,
 * src/matrix.cu:752-759 and its launch in src/encode.cu:134-141) and the per-coefficient
 * split-nibble product tables the multiply kernel reads (replaces the log/exp lookups of gf_mul,
 * src/matrix.cu:105-110).
 */
int rsgpu_create(rsgpu_ctx **out, int device, int k, int p);

/*
 * Same, but with a caller-supplied total matrix ((k+p) x k bytes, row-major, host memory), e.g. the
 * one stored in a METADATA file (src/decode.cu:274-281).  The first k rows need not be identity.
 */
int rsgpu_create_with_matrix(rsgpu_ctx **out, int device, int k, int p, const uint8_t *total_matrix);

/*
 * Same as rsgpu_create with a choice of generator.  RSGPU_GEN_VANDERMONDE is the reference's
 * V[i][j] = (j+1)^i (src/matrix.cu:752-759), which is NOT MDS: some survivor sets are singular (12 of
 * the 1001 four-erasure sets of RS(10,14)).  RSGPU_GEN_CAUCHY builds [I ; C], C[i][j] = 1/((k+i) + j):
 * every set of k fragments decodes (needs k + p <= 256).  No counterpart in the reference (its
 * north_star names "Vandermonde/Cauchy generator construction"); fragments made with it decode through
 * this engine via the matrix stored in METADATA, not through the reference's CPU program, which always
 * regenerates the Vandermonde matrix (src/cpu-rs.c:616-621).
 */
#define RSGPU_GEN_VANDERMONDE 0
#define RSGPU_GEN_CAUCHY      1
int rsgpu_create_generator(rsgpu_ctx **out, int device, int k, int p, int generator);

void rsgpu_destroy(rsgpu_ctx *ctx);

int rsgpu_k(const rsgpu_ctx *ctx);
int rsgpu_p(const rsgpu_ctx *ctx);
int rsgpu_device(const rsgpu_ctx *ctx);

/* Copy the (k+p) x k total matrix to host memory.  Replaces the matrix D2H copy for METADATA,
 * src/encode.cu:151-163.  Rows k.. are the p x k generator the reference calls encodingMatrix. */
int rsgpu_get_total_matrix(rsgpu_ctx *ctx, uint8_t *total_matrix);

/*
 * Encode: parity[i][c] = XOR_j gf_mul(V[i][j], data[j][c]) for c < len.  Device pointers.
 * Replaces encode_chunk (src/matrix.h:80, src/matrix.cu:767-830) and both matrix_mul kernels
 * (src/matrix.cu:232-323, 335-407).  Unlike the reference it overwrites parity (no pre-zeroing
 * needed), takes pitches, 64-bit len, and never synchronises.
 * Alignment: bases and pitches that are multiples of 16 bytes run at the HBM roofline.  Any other alignment -- e.g.
 * pitch = ceil(size / k), a file's fragments held back to back -- is coded by a kernel that reads and writes aligned
 * 128-bit vectors and realigns them in registers (about 0.75 of the roofline from 1 MiB per fragment on; the reference
 * drops to a byte-per-thread kernel whenever the length is not a multiple of 8, src/matrix.cu:796,811).  Same bytes.
 * A code with ONE parity row (p = 1: all coefficients 1) is a plain XOR of the fragments and runs without tables.
 * Byte columns are independent, so a BATCH of stripes stored fragment-major -- row j = fragment j of
 * stripe 0, then of stripe 1, ... (pitch = stripes * stripe_len) -- is one call with len = stripes *
 * stripe_len: one launch, tables built once, which is how small stripes reach the large-stripe rate.
 */
int rsgpu_encode(rsgpu_ctx *ctx, const uint8_t *data_d, size_t data_pitch,
                 uint8_t *parity_d, size_t parity_pitch, size_t len, void *stream);

/*
 * Encode a BATCH of n_stripes stripes of the same code in one call: fragment j of stripe s starts at
 * data_d + s * data_stripe_pitch + j * data_pitch, parity row i of stripe s at parity_d +
 * s * parity_stripe_pitch + i * parity_pitch (stripe-major and fragment-major batches are both instances
 * of this).  When the bases, pitches and stripe pitches are 16-byte aligned the whole batch is ONE kernel
 * launch per group of 8 parity rows (plus one small launch for the last len % 16 bytes of all stripes when
 * len is ragged), with the work spread evenly over the GPU whatever the stripe size -- small stripes
 * (fragments of 64 KiB .. a few MiB) are otherwise bound by the per-launch cost.  Other geometries are coded
 * stripe by stripe, same result.  No counterpart in the reference (one file per process run); SURVEY 8e "device-
 * resident batches".  Asynchronous on `stream`.  rsgpu_decode_erased_batch is the same for the lost natives.
 */
int rsgpu_encode_batch(rsgpu_ctx *ctx, const uint8_t *data_d, size_t data_pitch, size_t data_stripe_pitch,
                       uint8_t *parity_d, size_t parity_pitch, size_t parity_stripe_pitch,
                       size_t len, size_t n_stripes, void *stream);
int rsgpu_decode_erased_batch(rsgpu_ctx *ctx, const uint8_t *survivors_d, size_t survivors_pitch,
                              size_t survivors_stripe_pitch, uint8_t *out_d, size_t out_pitch,
                              size_t out_stripe_pitch, size_t len, size_t n_stripes, void *stream);

/*
 * Encode with one destination pointer per parity row (host array of p device pointers).  The rows may
 * be memory of OTHER GPUs of the box once rsgpu_enable_peer(ctx, that_device) has succeeded: the kernel's
 * 128-bit stores then travel over NVLink / NVSwitch while it computes the following columns, so the
 * parity fragments arrive on the devices (storage nodes) that keep them without a staging buffer or a
 * separate copy.  No counterpart in the reference, which gathers every fragment on the host
 * (src/encode.cu:410-429); SURVEY 8f-3.  Asynchronous on `stream` (a stream of the context's device).
 */
int rsgpu_enable_peer(rsgpu_ctx *ctx, int peer_device);
int rsgpu_encode_rows(rsgpu_ctx *ctx, const uint8_t *data_d, size_t data_pitch,
                      uint8_t *const *parity_rows /* p pointers */, size_t len, void *stream);

/*
 * Choose the k surviving fragments, in conf-file order (index < k: native, >= k: parity
 * index-k), src/decode.cu:302-307.  Gathers those rows of the total matrix, inverts the k x k
 * matrix with a single-block Gauss-Jordan kernel on the GPU (replaces CPU_invert_matrix,
 * src/cpu-decode.c:251-298, and the dead multi-launch GPU_invert_matrix, src/matrix.cu:666-744)
 * and builds the product tables for decoding.  Blocks until the inverse is known.
 * Returns RSGPU_ESINGULAR when the survivors cannot reconstruct the data (the reference cannot
 * detect this).  Row pivoting is used, so the result is the true inverse for every invertible
 * input, and equals the reference's wherever the reference is correct.
 * May be called while an rsgpu_decode / rsgpu_decode_erased of the previous survivor set is still
 * running on the caller's stream: the new tables go to a second buffer (two sets, used alternately,
 * re-use ordered by an event), so that decode finishes with the tables it started with.  After a
 * failure the context has no survivors (decode returns RSGPU_ESTATE) until a call succeeds.
 */
int rsgpu_set_survivors(rsgpu_ctx *ctx, const int *fragment_index /* k entries */);

/* Copy the current k x k decoding matrix (inverse) to host memory. */
int rsgpu_get_decode_matrix(rsgpu_ctx *ctx, uint8_t *inverse);

/* Natives that are NOT among the survivors, ascending; *count may be 0.  idx needs room for k. */
int rsgpu_get_erased(rsgpu_ctx *ctx, int *idx, int *count);

/*
 * Decode, reference semantics: recompute ALL k natives from the k survivors.
 * survivors_d row i is the fragment named on conf line i.  Replaces decode_chunk
 * (src/matrix.h:97, src/matrix.cu:838-905).
 */
int rsgpu_decode(rsgpu_ctx *ctx, const uint8_t *survivors_d, size_t survivors_pitch,
                 uint8_t *natives_d, size_t natives_pitch, size_t len, void *stream);

/*
 * Decode only what was lost: out row t is native rsgpu_get_erased().idx[t].  Same inputs as
 * rsgpu_decode; writes e*len instead of k*len bytes (the reference always rewrites all k,
 * src/matrix.cu:881).  One lost native whose survivors are the other natives plus parity row 0 -- the
 * commonest repair -- is recognised as a plain XOR of the survivors (no lookups, HBM speed for any k).
 */
int rsgpu_decode_erased(rsgpu_ctx *ctx, const uint8_t *survivors_d, size_t survivors_pitch,
                        uint8_t *out_d, size_t out_pitch, size_t len, void *stream);

/*
 * Repair: rebuild ANY fragments, natives or parities, from k survivors in one multiply.
 * survivors: k fragment indices in the order of the rows of survivors_d; wanted: n_wanted indices in
 * [0, k+p).  Output row t of rsgpu_repair is fragment wanted[t].  The e.g. "one storage node lost a
 * parity fragment" case costs (k + 1)*len bytes of traffic instead of a full decode plus a re-encode.
 * The reference can only rebuild all k natives (decode_chunk, src/matrix.cu:838-905) and re-encode.
 * rsgpu_set_repair blocks (it runs the inversion) and may return RSGPU_ESINGULAR.
 */
int rsgpu_set_repair(rsgpu_ctx *ctx, const int *survivors /* k */, const int *wanted, int n_wanted);
int rsgpu_repair(rsgpu_ctx *ctx, const uint8_t *survivors_d, size_t survivors_pitch,
                 uint8_t *out_d, size_t out_pitch, size_t len, void *stream);

/*
 * Stand-alone GF(2^8) matrix inverse on the GPU, host in / host out (n <= 255).
 * Replaces GPU_invert_matrix / CPU_invert_matrix (src/matrix.h:102, src/cpu-decode.h:27);
 * unlike those it does not destroy its input and reports singular matrices.
 */
int rsgpu_invert_matrix(rsgpu_ctx *ctx, const uint8_t *matrix, uint8_t *inverse, int n);

/*
 * Generic C (rows x len) = M (rows x k) * B (k x len) with a caller-supplied host matrix and device
 * buffers; the general form of matrix_mul (src/matrix.cu:232), used by tests and by re-encoding of
 * lost parities.  Builds temporary product tables (one small allocation) and blocks until done.
 */
int rsgpu_matmul(rsgpu_ctx *ctx, const uint8_t *matrix, int rows,
                 const uint8_t *in_d, size_t in_pitch, uint8_t *out_d, size_t out_pitch,
                 size_t len, void *stream);

/*
 * Host-memory coding through the per-device pipeline: the byte range is cut into slices and each
 * slice goes H2D -> kernel -> D2H on one of `slots` streams, so copies of one slice overlap the
 * kernel of another.  Replaces the per-device encode()/decode() of src/encode.cu:109-238 and
 * src/decode.cu:89-196 (the -s stream path).  Host buffers may be pageable or pinned
 * (rsgpu_host_alloc / rsgpu_host_register); pinned memory is what lets the copies overlap.
 * slots <= 0 picks the default (4).  Blocks until the outputs are in host memory.
 */
int rsgpu_encode_host(rsgpu_ctx *ctx, const uint8_t *data, size_t data_pitch,
                      uint8_t *parity, size_t parity_pitch, size_t len, int slots);
int rsgpu_decode_host(rsgpu_ctx *ctx, const uint8_t *survivors, size_t survivors_pitch,
                      uint8_t *natives, size_t natives_pitch, size_t len, int slots);
/* As rsgpu_decode_erased, through the same pipeline: out row t = native rsgpu_get_erased().idx[t]. */
int rsgpu_decode_erased_host(rsgpu_ctx *ctx, const uint8_t *survivors, size_t survivors_pitch,
                             uint8_t *out, size_t out_pitch, size_t len, int slots);

/* Pinned host memory (cudaMallocHost / cudaHostRegister), replaces src/encode.cu:389-390. */
int rsgpu_host_alloc(void **ptr, size_t bytes);
int rsgpu_host_free(void *ptr);
int rsgpu_host_register(void *ptr, size_t bytes);
int rsgpu_host_unregister(void *ptr);

/* Plain device memory helpers for C callers that have no CUDA headers. */
int rsgpu_dev_alloc(rsgpu_ctx *ctx, void **ptr, size_t bytes);
int rsgpu_dev_free(rsgpu_ctx *ctx, void *ptr);
int rsgpu_memcpy_h2d(rsgpu_ctx *ctx, void *dst_d, const void *src, size_t bytes);
int rsgpu_memcpy_d2h(rsgpu_ctx *ctx, void *dst, const void *src_d, size_t bytes);
int rsgpu_sync(rsgpu_ctx *ctx);

/*
 * Encode without allocating anything (no counterpart in the reference, which cudaMallocs on every
 * call, src/encode.cu:168-190).
 *
 * For callers that are not allowed to allocate GPU memory once they are running: engines embedded in
 * a process whose device memory was handed out as fixed arenas at start-up, or sitting in an I/O path
 * where the device-wide serialisation of cudaMalloc/cudaFree would stall every other stream.  Those
 * callers cannot use rsgpu_create.  This call needs no context: the GF tables, the generator matrix
 * [I;V] and the product tables are rebuilt on `stream` into a fixed 16 KiB array that is part of the
 * library image (each GPU that loads librsgpu.so holds it until the process exits), then the same
 * multiply kernel as rsgpu_encode runs.  Same arguments and result bytes as rsgpu_encode; works on the
 * CURRENT device; inputs and outputs are still the caller's buffers.
 *
 * Limits and rules:
 *   - k <= RSGPU_NOALLOC_MAX_K (32) and 1 <= p <= RSGPU_NOALLOC_MAX_P (8); anything larger returns
 *     RSGPU_ETOOBIG and does nothing.
 *   - The first call on each device looks up the array's device address (a one-time host-side query,
 *     made with that device current).  Make that first call before you start capturing a stream.
 *   - All calls on a device share the one array and each call rewrites it: two calls whose GPU work
 *     can overlap (different streams, or different host threads) race on it.  Issue them on one
 *     stream, or otherwise order them.
 *   - It performs no cudaMalloc / cudaFree and calls no allocating library routine.
 */
#define RSGPU_NOALLOC_MAX_K 32
#define RSGPU_NOALLOC_MAX_P 8
int rsgpu_encode_noalloc(int k, int p, const uint8_t *data_d, size_t data_pitch,
                         uint8_t *parity_d, size_t parity_pitch, size_t len, void *stream);

/*
 * Tuning knobs of the multiply kernel (the reference's -p grid cap, src/main.c:75-78, maps to
 * max_blocks).  variant: 0 = default, see DESIGN.md for the list; threads: block size (0 = default);
 * max_blocks: cap on gridDim.x (0 = persistent grid sized from the SM count).
 */
int rsgpu_set_tuning(rsgpu_ctx *ctx, int variant, int threads, int max_blocks);

/* Kernels launched by this context so far (what bench.py reports as gpu_launches). */
long long rsgpu_launch_count(const rsgpu_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* RSGPU_H_ */
