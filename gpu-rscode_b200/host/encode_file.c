/* encode_file.c -- drop-in for the reference's encode_file (src/encode.cu:300-473, declared in
 * src/encode.h:36): cut the file into k native fragments, compute p parity fragments on the GPU(s),
 * write "_<i>_<file>" and "<file>.METADATA" into the current directory.
 *
 * What differs underneath: nothing is held in RAM whole.  Fragment i is bytes [i*chunk, (i+1)*chunk) of
 * the file, so a slab of the stripe is k positional reads; the streaming pipeline (stream.c) overlaps
 * those reads, the H2D/kernel/D2H pipeline of every GPU and the n positional writes per slab.  The tail
 * of the last native fragment is zero (cpu-rs.c:502; the GPU reference leaves malloc garbage there,
 * src/encode.cu:325).  Sizes are 64-bit.
 */
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include "encode.h"
#include "rs_host.h"

void encode_file(char *fileName, int nativeBlockNum, int parityBlockNum, int gridDimXSize, int streamNum)
{
    const int k = nativeBlockNum, p = parityBlockNum;
    if (k < 1 || k > RSGPU_MAX_K || p < 0 || p > 255) {
        printf("Invalid code parameters: need 1 <= k <= %d and 0 <= n-k <= 255\n", RSGPU_MAX_K);
        exit(0);
    }
    const int fd_in = open(fileName, O_RDONLY);
    if (fd_in < 0) {
        printf("Cannot open source file!\n"); /* src/encode.cu:308-312 */
        exit(0);
    }
    struct stat sb;
    if (fstat(fd_in, &sb) != 0) {
        printf("fseek error!\n");
        exit(0);
    }
    const long long totalSize = (long long)sb.st_size;
    const long long chunk = rs_chunk_size(totalSize, k);

    const int gpu_budget = rs_file_gpu_budget(totalSize);

    int maxGrid = 65535;
    struct timespec ts0, ts1;
    clock_gettime(CLOCK_MONOTONIC, &ts0);
    rsgpu_device_max_grid(0, &maxGrid); /* first CUDA call of the process: pays the driver / context start-up */
    clock_gettime(CLOCK_MONOTONIC, &ts1);
    if (getenv("RS_TIMING"))
        fprintf(stderr, "[rs timing] CUDA start-up: %.1f ms\n",
                1e3 * (double)(ts1.tv_sec - ts0.tv_sec) + 1e-6 * (double)(ts1.tv_nsec - ts0.tv_nsec));
    if (gridDimXSize > maxGrid || gridDimXSize <= 0) {
        printf("Valid grid size: (0, %d]\n", maxGrid); /* src/encode.cu:351-355 */
        gridDimXSize = 0;
    }

    const size_t nlen = strlen(fileName) + 32;
    char *name = malloc(nlen);
    rs_stream_src *src = calloc((size_t)k, sizeof *src);
    rs_stream_dst *dst = calloc((size_t)(k + p), sizeof *dst);
    for (int j = 0; j < k; ++j) {
        src[j].fd = fd_in;
        src[j].base = (long long)j * chunk;
    }
    for (int i = 0; i < k + p; ++i) {
        snprintf(name, nlen, "_%d_%s", i, fileName); /* src/encode.cu:437,453 */
        dst[i].fd = open(name, O_WRONLY | O_CREAT | O_TRUNC, 0644);
        if (dst[i].fd < 0) {
            printf("Cannot open output file!\n");
            exit(0);
        }
        dst[i].base = 0;
        dst[i].from_out = i >= k;
        dst[i].row = i < k ? i : i - k;
        dst[i].limit = -1; /* fragment files carry the padded chunk in full */
    }

    uint8_t *total = malloc((size_t)(k + p) * k);
    rs_stream_desc d;
    memset(&d, 0, sizeof d);
    d.k = k;
    d.p = p;
    d.decode = 0;
    /* RS_GENERATOR=cauchy: MDS generator instead of the reference's Vandermonde rows; the matrix goes
     * into METADATA, which is what this engine's decoder (and the reference's GPU decoder) reads */
    const char *gen = getenv("RS_GENERATOR");
    d.generator = (gen && !strcmp(gen, "cauchy")) ? RSGPU_GEN_CAUCHY : RSGPU_GEN_VANDERMONDE;
    d.matrix_out = total;
    d.in_rows = k;
    d.out_rows = p;
    d.chunk = chunk;
    d.src = src;
    d.dst = dst;
    d.n_dst = k + p;
    d.grid_cap = gridDimXSize;
    d.slots = streamNum;
    d.max_gpus = gpu_budget;
    double wall_ms = 0;
    int used = 0;
    char err[512] = "";
    const int rc = rs_stream_run(&d, &wall_ms, &used, err, sizeof err);
    if (rc != RSGPU_OK) {
        if (!strcmp(err, "fread error!") || !strcmp(err, "fwrite error!")) {
            printf("%s\n", err); /* src/encode.cu:340-344,443-447 */
            exit(0);
        }
        /* CUDA failures: message on stderr, exit(EXIT_FAILURE) (src/helper_cuda.h:964-975) */
        fprintf(stderr, "%s\n", err);
        exit(EXIT_FAILURE);
    }
    printf("Total GPU encoding time using multiple devices: %fms\n", wall_ms); /* src/encode.cu:276 */

    for (int i = 0; i < k + p; ++i) {
        /* a fragment is exactly chunk bytes even when the pipeline had nothing to write (empty file) */
        if (ftruncate(dst[i].fd, (off_t)chunk) != 0 || close(dst[i].fd) != 0) {
            printf("fwrite error!\n");
            exit(0);
        }
    }
    close(fd_in);
    snprintf(name, nlen, "%s.METADATA", fileName);
    if (rs_write_metadata(name, totalSize, p, k, total) != 0) {
        printf("Cannot open META file!\n"); /* src/encode.cu:64-68 */
        exit(0);
    }
    free(name);
    free(total);
    free(src);
    free(dst);
}
