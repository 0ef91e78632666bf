/* sched.c -- stripe scheduling: one worker thread per visible GPU, each owning a contiguous byte
 * range of every fragment and driving its own context (streams, pinned-memory pipeline) with no
 * exchange between GPUs.
 *
 * Replaces the fan-out in the reference's encode_file / decode_file (src/encode.cu:357-429,
 * src/decode.cu:335-405: ThreadDataType, GPU_thread_func, a cudaMallocHost + host-to-host copy of
 * every byte before and after the GPU work, and a file-static pthread barrier that makes the call
 * non re-entrant).  Here the workers read and write the caller's (pinned) buffers in place through
 * pitched 2-D copies, the barrier lives in the job, and errors travel back as codes.
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "rs_host.h"

typedef struct {
    rs_job *job;
    int gpu;
    size_t off, len;
    pthread_barrier_t *barrier;
    double t_start, t_end; /* seconds, CLOCK_MONOTONIC */
    double dev_ms;
    int error;
    char error_text[512];
} worker_t;

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static void worker_fail(worker_t *w, int code, const char *what)
{
    w->error = code;
    snprintf(w->error_text, sizeof w->error_text, "GPU %d: %s: %s (%s)", w->gpu, what, rsgpu_strerror(code),
             rsgpu_last_error());
}

static void *worker_main(void *arg)
{
    worker_t *w = arg;
    rs_job *job = w->job;
    rsgpu_ctx *ctx = NULL;
    int rc;

    /* set-up (context, tables, inverse) is outside the barrier-timed region, as in the reference
     * where cudaSetDevice precedes the barriers (src/encode.cu:243-257) */
    if (job->total_matrix)
        rc = rsgpu_create_with_matrix(&ctx, w->gpu, job->k, job->p, job->total_matrix);
    else
        rc = rsgpu_create(&ctx, w->gpu, job->k, job->p);
    if (rc != RSGPU_OK)
        worker_fail(w, rc, "rsgpu_create");
    if (!w->error && job->grid_cap > 0 && (rc = rsgpu_set_tuning(ctx, 0, 0, job->grid_cap)) != RSGPU_OK)
        worker_fail(w, rc, "rsgpu_set_tuning");
    if (!w->error && w->gpu == 0 && job->matrix_out && (rc = rsgpu_get_total_matrix(ctx, job->matrix_out)) != RSGPU_OK)
        worker_fail(w, rc, "rsgpu_get_total_matrix");

    pthread_barrier_wait(w->barrier);
    w->t_start = now_s();
    pthread_barrier_wait(w->barrier);

    const double t0 = now_s();
    if (!w->error && job->kind == RS_JOB_DECODE_ERASED) {
        /* the inversion is part of decoding time, as in the reference (src/decode.cu:333 precedes
         * the fan-out there; here every GPU inverts redundantly -- microseconds -- instead of
         * receiving a broadcast) */
        if ((rc = rsgpu_set_survivors(ctx, job->survivors)) != RSGPU_OK)
            worker_fail(w, rc, "rsgpu_set_survivors");
    }
    if (!w->error && w->len > 0) {
        if (job->kind == RS_JOB_ENCODE)
            rc = rsgpu_encode_host(ctx, job->in + w->off, job->in_pitch, job->out + w->off, job->out_pitch, w->len,
                                   job->slots);
        else
            rc = rsgpu_decode_erased_host(ctx, job->in + w->off, job->in_pitch, job->out + w->off, job->out_pitch,
                                          w->len, job->slots);
        if (rc != RSGPU_OK)
            worker_fail(w, rc, job->kind == RS_JOB_ENCODE ? "rsgpu_encode_host" : "rsgpu_decode_erased_host");
    }
    w->dev_ms = 1e3 * (now_s() - t0);

    pthread_barrier_wait(w->barrier);
    w->t_end = now_s();
    if (ctx)
        rsgpu_destroy(ctx);
    return NULL;
}

/* How many of the `visible` GPUs a stripe of `len` bytes per fragment is spread over: a GPU is worth
 * starting only for a few MiB, tiny inputs stay on GPU 0. */
int rs_plan_gpus(size_t len, int visible)
{
    const size_t min_range = (size_t)4 << 20;
    int use = visible < 1 ? 1 : visible;
    while (use > 1 && len / (size_t)use < min_range)
        --use;
    return use;
}

/* Byte range [*off, *off + *n) of every fragment that GPU g of `parts` owns: contiguous, boundaries on
 * 256 bytes so every GPU keeps 16-byte aligned rows, the last GPU takes the remainder (the reference
 * splits the same way without the alignment, src/encode.cu:368-381). */
void rs_partition(size_t len, int parts, int g, size_t *off, size_t *n)
{
    const size_t per = (len / (size_t)parts) / 256 * 256;
    *off = (size_t)g * per;
    *n = (g == parts - 1) ? len - *off : per;
}

int rs_run_job(rs_job *job)
{
    job->error = 0;
    job->error_text[0] = 0;
    job->wall_ms = 0;
    int n = 0;
    int rc = rsgpu_device_count(&n);
    if (rc != RSGPU_OK || n < 1) {
        job->error = rc != RSGPU_OK ? rc : RSGPU_ECUDA;
        snprintf(job->error_text, sizeof job->error_text, "no usable CUDA device: %s", rsgpu_last_error());
        return job->error;
    }
    const int use = rs_plan_gpus(job->len, n);
    job->n_gpus = use;

    worker_t *w = calloc((size_t)use, sizeof *w);
    pthread_t *th = calloc((size_t)use, sizeof *th);
    pthread_barrier_t barrier;
    pthread_barrier_init(&barrier, NULL, (unsigned)use);
    for (int g = 0; g < use; ++g) {
        w[g].job = job;
        w[g].gpu = g;
        w[g].barrier = &barrier;
        rs_partition(job->len, use, g, &w[g].off, &w[g].len);
        pthread_create(&th[g], NULL, worker_main, &w[g]);
    }
    for (int g = 0; g < use; ++g)
        pthread_join(th[g], NULL);
    pthread_barrier_destroy(&barrier);

    job->wall_ms = 1e3 * (w[0].t_end - w[0].t_start);
    for (int g = 0; g < use; ++g) {
        printf("Device%d: Total GPU %s time: %fms\n", g, job->kind == RS_JOB_ENCODE ? "encoding" : "decoding",
               w[g].dev_ms);
        if (w[g].error && !job->error) {
            job->error = w[g].error;
            memcpy(job->error_text, w[g].error_text, sizeof job->error_text);
        }
    }
    free(w);
    free(th);
    return job->error;
}
