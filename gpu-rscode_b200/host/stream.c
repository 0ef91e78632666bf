/* stream.c -- streaming slab pipeline of the host layer: file -> pinned ring -> GPU(s) -> files.
 *
 * The reference holds the whole file plus all fragments in RAM twice (pageable + pinned copies,
 * src/encode.cu:325-326,389-390), reads and writes with one thread, and only then / before that runs
 * the GPUs (src/encode.cu:332-345,434-465).  Here the column space [0, chunk) of the stripe is cut into
 * slabs; a slab travels through three overlapped stages on a small ring of pinned buffers:
 *
 *   READ   k positional reads (one per input row), spread over the IO threads
 *   CODE   one GPU worker per visible device takes the next filled slab and runs the pipelined
 *          host-memory entry point on it (rsgpu_encode_host / rsgpu_decode_erased_host)
 *   WRITE  one positional write per output row, spread over the IO threads
 *
 * so disk/page-cache traffic, PCIe copies and kernels of different slabs run at the same time, memory
 * use is O(ring) instead of O(file) (files larger than RAM work), pinned memory is allocated once, and
 * slabs go to GPUs round-robin with no exchange between them.
 */
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include "rs_host.h"

enum { ST_FREE, ST_READING, ST_READY, ST_CODING, ST_WRITING };

typedef struct {
    uint8_t *in, *out;   /* pinned: [in_rows][slab] and [out_rows][slab] */
    long long slab;      /* slab index held, -1 if none */
    size_t len;          /* valid bytes per row in this slab */
    int state;
    int pending;         /* outstanding read / write tasks */
} slot_t;

typedef struct {
    int kind;            /* 0 read, 1 write */
    int slot;
    int row;
} task_t;

struct rs_stream {
    rs_stream_desc d;
    size_t slab_bytes;
    long long n_slabs, next_slab, done_slabs;
    int n_slots, n_gpus, n_io;
    slot_t *slot;
    task_t *queue;       /* ring of IO tasks */
    int q_head, q_tail, q_cap;
    pthread_mutex_t mu;
    pthread_cond_t cv_task, cv_ready, cv_free;
    int stop;
    int error;
    char error_text[512];
    double gpu_ms[64];
    int gpu_slabs[64];   /* slabs each GPU worker coded */
    double t_start;
};

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static void set_error(struct rs_stream *st, int code, const char *text)
{
    /* caller holds st->mu */
    if (!st->error) {
        st->error = code;
        snprintf(st->error_text, sizeof st->error_text, "%s", text);
    }
    st->stop = 1;
    pthread_cond_broadcast(&st->cv_task);
    pthread_cond_broadcast(&st->cv_ready);
    pthread_cond_broadcast(&st->cv_free);
}

static void push_task(struct rs_stream *st, int kind, int slot, int row)
{
    st->queue[st->q_tail] = (task_t){kind, slot, row};
    st->q_tail = (st->q_tail + 1) % st->q_cap;
}

/* full positional transfer; short reads past EOF leave the (pre-zeroed) tail alone */
static int pread_full(int fd, uint8_t *buf, size_t n, long long off)
{
    size_t got = 0;
    while (got < n) {
        ssize_t r = pread(fd, buf + got, n - got, (off_t)(off + (long long)got));
        if (r < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        if (r == 0) break;
        got += (size_t)r;
    }
    if (got < n) memset(buf + got, 0, n - got);
    return 0;
}

static int pwrite_full(int fd, const uint8_t *buf, size_t n, long long off)
{
    size_t put = 0;
    while (put < n) {
        ssize_t r = pwrite(fd, buf + put, n - put, (off_t)(off + (long long)put));
        if (r < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        put += (size_t)r;
    }
    return 0;
}

static void *io_main(void *arg)
{
    struct rs_stream *st = arg;
    const rs_stream_desc *d = &st->d;
    pthread_mutex_lock(&st->mu);
    for (;;) {
        while (!st->stop && st->q_head == st->q_tail)
            pthread_cond_wait(&st->cv_task, &st->mu);
        if (st->q_head == st->q_tail) break; /* stop and drained */
        task_t t = st->queue[st->q_head];
        st->q_head = (st->q_head + 1) % st->q_cap;
        slot_t *s = &st->slot[t.slot];
        const long long col0 = s->slab * (long long)st->slab_bytes;
        const size_t len = s->len;
        pthread_mutex_unlock(&st->mu);

        int rc = 0;
        if (t.kind == 0) {
            const rs_stream_src *src = &d->src[t.row];
            rc = pread_full(src->fd, s->in + (size_t)t.row * st->slab_bytes, len, src->base + col0);
        } else {
            const rs_stream_dst *dst = &d->dst[t.row];
            const uint8_t *buf = dst->from_out ? s->out + (size_t)dst->row * st->slab_bytes
                                               : s->in + (size_t)dst->row * st->slab_bytes;
            /* clip to the row's logical length (the last native of a decode ends at total_size) */
            long long n = (long long)len;
            if (dst->limit >= 0 && col0 + n > dst->limit) n = dst->limit - col0;
            if (n > 0) {
                if (dst->map)
                    memcpy(dst->map + dst->base + col0, buf, (size_t)n);
                else
                    rc = pwrite_full(dst->fd, buf, (size_t)n, dst->base + col0);
            }
        }

        pthread_mutex_lock(&st->mu);
        if (rc != 0) {
            set_error(st, RSGPU_EINVAL, t.kind == 0 ? "fread error!" : "fwrite error!");
            continue;
        }
        if (--s->pending == 0) {
            if (t.kind == 0) {
                s->state = ST_READY;
                pthread_cond_broadcast(&st->cv_ready);
            } else {
                s->state = ST_FREE;
                s->slab = -1;
                ++st->done_slabs;
                pthread_cond_broadcast(&st->cv_free);
            }
        }
    }
    pthread_mutex_unlock(&st->mu);
    return NULL;
}

typedef struct {
    struct rs_stream *st;
    int gpu;
} gpu_arg_t;

static void *gpu_main(void *arg)
{
    gpu_arg_t *ga = arg;
    struct rs_stream *st = ga->st;
    const rs_stream_desc *d = &st->d;
    rsgpu_ctx *ctx = NULL;
    char text[512];
    int rc = d->total_matrix ? rsgpu_create_with_matrix(&ctx, ga->gpu, d->k, d->p, d->total_matrix)
                             : rsgpu_create_generator(&ctx, ga->gpu, d->k, d->p, d->generator);
    if (rc == RSGPU_OK && d->grid_cap > 0) rc = rsgpu_set_tuning(ctx, 0, 0, d->grid_cap);
    if (rc == RSGPU_OK && ga->gpu == 0 && d->matrix_out) rc = rsgpu_get_total_matrix(ctx, d->matrix_out);
    if (rc == RSGPU_OK && d->decode) rc = rsgpu_set_survivors(ctx, d->survivors);
    if (getenv("RS_TIMING"))
        fprintf(stderr, "[rs timing] GPU %d context + tables ready at +%.1f ms\n", ga->gpu, 1e3 * (now_s() - st->t_start));
    pthread_mutex_lock(&st->mu);
    if (rc != RSGPU_OK) {
        snprintf(text, sizeof text, "GPU %d: set-up failed: %s (%s)", ga->gpu, rsgpu_strerror(rc), rsgpu_last_error());
        set_error(st, rc, text);
    }
    for (;;) {
        /* take the READY slab with the lowest index (keeps writes roughly sequential) */
        int pick = -1;
        while (!st->stop) {
            for (int i = 0; i < st->n_slots; ++i)
                if (st->slot[i].state == ST_READY && (pick < 0 || st->slot[i].slab < st->slot[pick].slab)) pick = i;
            if (pick >= 0 || st->done_slabs == st->n_slabs) break;
            pthread_cond_wait(&st->cv_ready, &st->mu);
        }
        if (pick < 0) break;
        slot_t *s = &st->slot[pick];
        s->state = ST_CODING;
        pthread_mutex_unlock(&st->mu);

        const double t0 = now_s();
        if (d->out_rows > 0) {
            if (d->decode)
                rc = rsgpu_decode_erased_host(ctx, s->in, st->slab_bytes, s->out, st->slab_bytes, s->len, d->slots);
            else
                rc = rsgpu_encode_host(ctx, s->in, st->slab_bytes, s->out, st->slab_bytes, s->len, d->slots);
        }
        const double dt = now_s() - t0;

        pthread_mutex_lock(&st->mu);
        st->gpu_ms[ga->gpu] += 1e3 * dt;
        ++st->gpu_slabs[ga->gpu];
        if (rc != RSGPU_OK) {
            snprintf(text, sizeof text, "GPU %d: coding failed: %s (%s)", ga->gpu, rsgpu_strerror(rc), rsgpu_last_error());
            set_error(st, rc, text);
            break;
        }
        s->state = ST_WRITING;
        s->pending = d->n_dst;
        if (d->n_dst == 0) {
            s->state = ST_FREE;
            s->slab = -1;
            ++st->done_slabs;
            pthread_cond_broadcast(&st->cv_free);
        }
        for (int r = 0; r < d->n_dst; ++r) push_task(st, 1, pick, r);
        pthread_cond_broadcast(&st->cv_task);
        if (st->done_slabs == st->n_slabs) pthread_cond_broadcast(&st->cv_ready);
    }
    pthread_mutex_unlock(&st->mu);
    if (ctx) rsgpu_destroy(ctx);
    return NULL;
}

/* Slab length per row: enough slabs to keep every GPU and the ring busy, rows long enough for PCIe. */
size_t rs_stream_slab_bytes(long long chunk, int rows, int n_gpus)
{
    /* pinning costs ~0.6 s per GiB (measured on the B200 box), so the ring is kept small: <= 64 MiB of
     * input per slab, which is still >= 2 MiB per row for k = 32 */
    const long long min_slab = 2LL << 20, max_total = 64LL << 20;
    long long s = chunk / (2LL * n_gpus);
    if (s > max_total / rows) s = max_total / rows;
    if (s < min_slab) s = min_slab;
    if (s > chunk) s = chunk;
    s = (s + 255) / 256 * 256;
    return (size_t)(s > 0 ? s : 256);
}

int rs_stream_run(const rs_stream_desc *desc, double *wall_ms, int *n_gpus_used, char *error_text, size_t error_cap)
{
    struct rs_stream st;
    memset(&st, 0, sizeof st);
    st.d = *desc;
    int visible = 0;
    int rc = rsgpu_device_count(&visible);
    if (rc != RSGPU_OK || visible < 1) {
        snprintf(error_text, error_cap, "no usable CUDA device: %s", rsgpu_last_error());
        return rc != RSGPU_OK ? rc : RSGPU_ECUDA;
    }
    if (visible > 64) visible = 64;
    if (desc->max_gpus > 0 && visible > desc->max_gpus) visible = desc->max_gpus;
    st.n_gpus = rs_plan_gpus((size_t)desc->chunk, visible);
    st.slab_bytes = rs_stream_slab_bytes(desc->chunk, desc->in_rows, st.n_gpus);
    st.n_slabs = desc->chunk > 0 ? (desc->chunk + (long long)st.slab_bytes - 1) / (long long)st.slab_bytes : 0;
    if (st.n_slabs < st.n_gpus) st.n_gpus = st.n_slabs > 0 ? (int)st.n_slabs : 1;
    st.n_slots = st.n_gpus + 2;
    if ((long long)st.n_slots > st.n_slabs) st.n_slots = st.n_slabs > 0 ? (int)st.n_slabs : 1;
    st.n_io = desc->io_threads;
    if (st.n_io <= 0 && getenv("RS_IO_THREADS")) st.n_io = atoi(getenv("RS_IO_THREADS"));
    if (st.n_io <= 0) {   // one per online core, between 4 and 16: the stage is a memcpy through the page cache
        // (16-core B200 host, RS(32,40), 4 GiB in tmpfs: encode pipeline 719 / 320 / 250 / 231 ms with 4 / 8 / 12 / 16)
        long nc = sysconf(_SC_NPROCESSORS_ONLN);
        st.n_io = nc < 4 ? 4 : (nc > 16 ? 16 : (int)nc);
    }
    if (st.n_io > 64) st.n_io = 64;
    if (n_gpus_used) *n_gpus_used = st.n_gpus;

    const double t_alloc0 = now_s();
    st.slot = calloc((size_t)st.n_slots, sizeof(slot_t));
    st.q_cap = st.n_slots * (desc->in_rows + desc->n_dst + 2) + 8;
    st.queue = calloc((size_t)st.q_cap, sizeof(task_t));
    if (!st.slot || !st.queue) {
        free(st.slot);
        free(st.queue);
        snprintf(error_text, error_cap, "out of host memory");
        return RSGPU_ENOMEM;
    }
    for (int i = 0; i < st.n_slots; ++i) {
        st.slot[i].slab = -1;
        st.slot[i].state = ST_FREE;
        rc = rsgpu_host_alloc((void **)&st.slot[i].in, (size_t)desc->in_rows * st.slab_bytes);
        if (rc == RSGPU_OK)
            rc = rsgpu_host_alloc((void **)&st.slot[i].out, (size_t)(desc->out_rows > 0 ? desc->out_rows : 1) * st.slab_bytes);
        if (rc != RSGPU_OK) {
            snprintf(error_text, error_cap, "pinned allocation failed: %s (%s)", rsgpu_strerror(rc), rsgpu_last_error());
            for (int j = 0; j <= i; ++j) {
                rsgpu_host_free(st.slot[j].in);
                rsgpu_host_free(st.slot[j].out);
            }
            free(st.slot);
            free(st.queue);
            return rc;
        }
    }
    pthread_mutex_init(&st.mu, NULL);
    pthread_cond_init(&st.cv_task, NULL);
    pthread_cond_init(&st.cv_ready, NULL);
    pthread_cond_init(&st.cv_free, NULL);

    if (getenv("RS_TIMING"))
        fprintf(stderr, "[rs timing] ring: %d slots x %d rows x %zu B pinned in %.1f ms; %lld slabs, %d GPU(s), %d IO threads\n",
                st.n_slots, desc->in_rows + desc->out_rows, st.slab_bytes, 1e3 * (now_s() - t_alloc0), st.n_slabs, st.n_gpus, st.n_io);
    pthread_t *io = calloc((size_t)st.n_io, sizeof(pthread_t));
    pthread_t *gp = calloc((size_t)st.n_gpus, sizeof(pthread_t));
    gpu_arg_t *ga = calloc((size_t)st.n_gpus, sizeof(gpu_arg_t));
    const double t0 = now_s();
    st.t_start = t0;
    /* a thread that cannot be started is simply not there: fewer IO threads are slower, not wrong; with no
     * IO thread or no GPU worker at all the job is stopped with an error */
    int io_started = 0, gpu_started = 0;
    if (io && gp && ga) {
        for (int i = 0; i < st.n_io; ++i)
            if (pthread_create(&io[io_started], NULL, io_main, &st) == 0) ++io_started;
        for (int g = 0; g < st.n_gpus; ++g) {
            ga[gpu_started].st = &st;
            ga[gpu_started].gpu = g;
            if (pthread_create(&gp[gpu_started], NULL, gpu_main, &ga[gpu_started]) == 0) ++gpu_started;
        }
    }
    st.n_io = io_started;
    st.n_gpus = gpu_started;
    if (io_started == 0 || gpu_started == 0) {
        pthread_mutex_lock(&st.mu);
        set_error(&st, RSGPU_ENOMEM, "could not start the IO / GPU worker threads");
        pthread_mutex_unlock(&st.mu);
    }

    /* feeder: hand slabs to free slots in order */
    pthread_mutex_lock(&st.mu);
    while (!st.stop && st.next_slab < st.n_slabs) {
        int f = -1;
        for (int i = 0; i < st.n_slots; ++i)
            if (st.slot[i].state == ST_FREE) { f = i; break; }
        if (f < 0) {
            pthread_cond_wait(&st.cv_free, &st.mu);
            continue;
        }
        slot_t *s = &st.slot[f];
        s->slab = st.next_slab++;
        const long long col0 = s->slab * (long long)st.slab_bytes;
        s->len = (size_t)(desc->chunk - col0 < (long long)st.slab_bytes ? desc->chunk - col0 : (long long)st.slab_bytes);
        s->state = ST_READING;
        s->pending = desc->in_rows;
        for (int r = 0; r < desc->in_rows; ++r) push_task(&st, 0, f, r);
        pthread_cond_broadcast(&st.cv_task);
    }
    while (!st.stop && st.done_slabs < st.n_slabs) pthread_cond_wait(&st.cv_free, &st.mu);
    st.stop = 1;
    pthread_cond_broadcast(&st.cv_task);
    pthread_cond_broadcast(&st.cv_ready);
    pthread_mutex_unlock(&st.mu);

    const double t_pipe = now_s();
    for (int g = 0; g < st.n_gpus; ++g) pthread_join(gp[g], NULL);
    for (int i = 0; i < st.n_io; ++i) pthread_join(io[i], NULL);
    if (wall_ms) *wall_ms = 1e3 * (t_pipe - t0);
    if (getenv("RS_TIMING"))
        fprintf(stderr, "[rs timing] pipeline %.1f ms, context teardown %.1f ms\n", 1e3 * (t_pipe - t0), 1e3 * (now_s() - t_pipe));
    if (getenv("RS_TIMING"))
        for (int g = 0; g < st.n_gpus; ++g)
            fprintf(stderr, "[rs timing] GPU %d coded %d of %lld slabs\n", g, st.gpu_slabs[g], st.n_slabs);
    for (int g = 0; g < st.n_gpus; ++g)
        printf("Device%d: Total GPU %s time: %fms\n", g, desc->decode ? "decoding" : "encoding", st.gpu_ms[g]);

    for (int i = 0; i < st.n_slots; ++i) {
        rsgpu_host_free(st.slot[i].in);
        rsgpu_host_free(st.slot[i].out);
    }
    free(st.slot);
    free(st.queue);
    free(io);
    free(gp);
    free(ga);
    pthread_mutex_destroy(&st.mu);
    pthread_cond_destroy(&st.cv_task);
    pthread_cond_destroy(&st.cv_ready);
    pthread_cond_destroy(&st.cv_free);
    if (st.error) snprintf(error_text, error_cap, "%s", st.error_text);
    return st.error;
}
