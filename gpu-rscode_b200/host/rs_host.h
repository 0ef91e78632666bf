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

/* ---- sched.c : stripe scheduling over the visible GPUs ---- */
typedef enum { RS_JOB_ENCODE, RS_JOB_DECODE_ERASED } rs_job_kind;

typedef struct {
    rs_job_kind kind;
    int k, p;
    const uint8_t *total_matrix; /* NULL: generate [I;V] on the GPU */
    const int *survivors;        /* decode: k fragment indices in conf order */
    const uint8_t *in;           /* [in_rows][pitch] host (pinned) */
    uint8_t *out;                /* [out_rows][pitch] host (pinned) */
    size_t in_pitch, out_pitch;
    size_t len;                  /* bytes per fragment row */
    int grid_cap;                /* reference -p, 0 = default */
    int slots;                   /* reference -s, <= 0 = default */
    /* results */
    uint8_t *matrix_out;         /* if not NULL: receives GPU 0's (k+p) x k total matrix */
    double wall_ms;              /* between the two barriers, as the reference measures it */
    int n_gpus;
    int error;                   /* first RSGPU_E* code seen by any worker, 0 if none */
    char error_text[512];
} rs_job;

int rs_run_job(rs_job *job); /* returns job->error */
int rs_plan_gpus(size_t len, int visible);
void rs_partition(size_t len, int parts, int g, size_t *off, size_t *n);

#endif
