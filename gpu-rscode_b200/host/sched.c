/* sched.c -- how work is spread over the visible GPUs (no exchange between GPUs anywhere on the path).
 *
 * Replaces the arithmetic of the fan-out in the reference's encode_file / decode_file
 * (src/encode.cu:357-403, src/decode.cu:335-378).  The pipeline that uses it is stream.c.
 */
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include "rs_host.h"

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


/* How many GPUs a file job of `bytes` should start.  Measured on B200 boxes (tools/run_files.py,
 * RS_TIMING=1): the host side of a file job (page-cache reads and writes through 8 IO threads) moves
 * ~25 GB/s while ONE GPU's pinned H2D+D2H pipeline takes ~50 GB/s, and every additional visible GPU adds
 * 0.6-0.9 s of driver / context start-up per process (4 GiB RS(32,40) encode: 1.78 s with one GPU visible,
 * 3.32 s with two).  So unlike the reference, which starts a thread on every visible GPU
 * (src/encode.cu:357-366), a file job uses one GPU per 32 GiB of input; RS_GPUS=<n> overrides. */
int rs_file_gpu_budget(long long bytes)
{
    const char *e = getenv("RS_GPUS");
    if (e && atoi(e) > 0)
        return atoi(e);
    const long long per_gpu = 32LL << 30;
    long long n = (bytes + per_gpu - 1) / per_gpu;
    return n < 1 ? 1 : (n > 64 ? 64 : (int)n);
}

/* Narrow CUDA_VISIBLE_DEVICES to the first `n` devices (of the user's list if there is one) so that the
 * CUDA start-up of this process only pays for the GPUs it will use.  Only effective before the first
 * CUDA call of the process.  This edits the process environment, so it is called by the RS command line
 * (main.c, a process of our own) and NEVER from encode_file / decode_file: a program that embeds the library
 * keeps its environment, and the library simply creates no context on GPUs beyond the job's budget. */
static void rs_limit_visible_gpus(int n)
{
    const char *cur = getenv("CUDA_VISIBLE_DEVICES");
    char buf[1024];
    size_t len = 0;
    buf[0] = 0;
    if (cur && *cur) {
        int taken = 0;
        const char *p = cur;
        while (*p && taken < n) {
            const char *q = strchr(p, ',');
            size_t w = q ? (size_t)(q - p) : strlen(p);
            if (len + w + 2 >= sizeof buf)
                break;
            if (taken)
                buf[len++] = ',';
            memcpy(buf + len, p, w);
            len += w;
            buf[len] = 0;
            ++taken;
            if (!q)
                break;
            p = q + 1;
        }
    } else {
        for (int i = 0; i < n && len + 8 < sizeof buf; ++i)
            len += (size_t)snprintf(buf + len, sizeof buf - len, i ? ",%d" : "%d", i);
    }
    if (buf[0])
        setenv("CUDA_VISIBLE_DEVICES", buf, 1);
}

/* Command-line helper (main.c): size the job from the file (encode) or from "<file>.METADATA" (decode) and
 * narrow the visible GPUs to its budget before CUDA starts.  Errors are left to encode_file / decode_file. */
void rs_cli_prepare_gpus(const char *file, int decode)
{
    long long bytes = -1;
    if (!file) return;
    if (decode) {
        const size_t nlen = strlen(file) + 32;
        char *name = malloc(nlen);
        if (!name) return;
        snprintf(name, nlen, "%s.METADATA", file);
        FILE *fp = fopen(name, "rb");
        free(name);
        if (fp) {
            if (fscanf(fp, "%lld", &bytes) != 1) bytes = -1;
            fclose(fp);
        }
    } else {
        struct stat sb;
        if (stat(file, &sb) == 0) bytes = (long long)sb.st_size;
    }
    if (bytes >= 0) rs_limit_visible_gpus(rs_file_gpu_budget(bytes));
}
