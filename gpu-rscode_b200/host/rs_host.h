/* rs_host.h -- internal declarations of the C host layer (librs_host.so).  Public entry points are
 * include/encode.h and include/decode.h; everything CUDA goes through include/rsgpu.h. */
#ifndef RS_HOST_H_
#define RS_HOST_H_

#include <stddef.h>
#include <stdint.h>

#include "rsgpu.h"

/* ---- metadata.c : "<file>.METADATA" (reference src/encode.cu:61-101, src/decode.cu:257-282) ---- */
typedef struct {
    long long total_size;
    int p, k;
    uint8_t *matrix; /* (k+p) x k when the file carries it (GPU writer), else NULL (CPU writer) */
} rs_metadata;

long long rs_chunk_size(long long total_size, int k);
int rs_write_metadata(const char *path, long long total_size, int p, int k, const uint8_t *total_matrix);
int rs_read_metadata(const char *path, rs_metadata *md); /* 0 ok, -1 cannot open, -2 malformed */
void rs_free_metadata(rs_metadata *md);

/* ---- conf.c : the decode configuration file (reference src/decode.cu:293-319) ---- */
typedef struct {
    int k;
    char **name; /* k fragment file names, conf order */
    int *index;  /* atoi(name + 1) */
} rs_conf;

int rs_read_conf(const char *path, int k, rs_conf *conf); /* 0 ok, -1 cannot open, -2 too few names */
void rs_free_conf(rs_conf *conf);

/* ---- sched.c : how a stripe is spread over the visible GPUs ---- */
int rs_plan_gpus(size_t len, int visible);
void rs_partition(size_t len, int parts, int g, size_t *off, size_t *n);
int rs_file_gpu_budget(long long bytes);
void rs_cli_prepare_gpus(const char *file, int decode); /* RS binary only: narrows CUDA_VISIBLE_DEVICES to the job's GPU budget */

/* ---- stream.c : streaming slab pipeline  file -> pinned ring -> GPU(s) -> files ---- */
typedef struct {
    int fd;          /* input row r of a slab = pread(fd, base + column offset) ; short reads give zeros */
    long long base;
} rs_stream_src;

typedef struct {
    int fd;          /* output row: pwrite(fd, base + column offset) ... */
    uint8_t *map;    /* ... or, when not NULL, memcpy into this shared mapping of the file (many rows of one
                        file: positional writes to a single inode serialise in the kernel, stores do not) */
    long long base;
    int from_out;    /* 1: row `row` of the GPU output, 0: row `row` of the slab's input (pass-through) */
    int row;
    long long limit; /* logical length of this row in bytes (writes are clipped to it); < 0 = whole chunk */
} rs_stream_dst;

typedef struct {
    int k, p;
    int decode;                  /* 0: rsgpu_encode_host, 1: rsgpu_set_survivors + rsgpu_decode_erased_host */
    const uint8_t *total_matrix; /* NULL: generate the total matrix on the GPU */
    int generator;               /* when generating: RSGPU_GEN_VANDERMONDE (reference) or RSGPU_GEN_CAUCHY */
    const int *survivors;        /* decode: k fragment indices in conf order */
    uint8_t *matrix_out;         /* if not NULL: receives GPU 0's (k+p) x k total matrix */
    int in_rows, out_rows;       /* rows read per slab, rows the GPU produces per slab */
    long long chunk;             /* bytes per fragment */
    const rs_stream_src *src;    /* in_rows entries */
    const rs_stream_dst *dst;    /* n_dst entries */
    int n_dst;
    int grid_cap;                /* reference -p, 0 = default */
    int slots;                   /* reference -s, <= 0 = default */
    int io_threads;              /* <= 0 = RS_IO_THREADS, else the online cores clamped to 4..16 */
    int max_gpus;                /* <= 0 = every visible GPU */
} rs_stream_desc;

size_t rs_stream_slab_bytes(long long chunk, int rows, int n_gpus);
/* returns 0 or the first RSGPU_E* code; error_text gets the message */
int rs_stream_run(const rs_stream_desc *desc, double *wall_ms, int *n_gpus_used, char *error_text, size_t error_cap);

#endif
