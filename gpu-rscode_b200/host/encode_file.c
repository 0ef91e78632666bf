/* encode_file.c -- drop-in for the reference's encode_file (src/encode.cu:300-473, declared in
 * src/encode.h:36): read the file, cut it into k native fragments, compute p parity fragments on the
 * GPU(s), write "_<i>_<file>" and "<file>.METADATA" into the current directory.
 *
 * What differs underneath: the file is read once, straight into one pinned buffer whose rows ARE the
 * fragments (fragment i = bytes [i*chunk, (i+1)*chunk) of the file, so no per-fragment fseek/fread
 * loop and no pageable->pinned copy); the tail of the last fragment is zero (cpu-rs.c:502; the GPU
 * reference leaves it uninitialised, src/encode.cu:325); every GPU works in place on its byte range of
 * that buffer through the stream pipeline of rsgpu_encode_host; sizes are 64-bit.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "encode.h"
#include "rs_host.h"

static void die_cuda(const char *what, int code)
{
    /* reference convention for CUDA failures: message on stderr, exit(EXIT_FAILURE)
     * (src/helper_cuda.h:964-975) */
    fprintf(stderr, "CUDA error in %s: %s (%s)\n", what, rsgpu_strerror(code), rsgpu_last_error());
    exit(EXIT_FAILURE);
}

void encode_file(char *fileName, int nativeBlockNum, int parityBlockNum, int gridDimXSize, int streamNum)
{
    const int k = nativeBlockNum, p = parityBlockNum;
    if (k < 1 || k > RSGPU_MAX_K || p < 0 || p > 255) {
        printf("Invalid code parameters: need 1 <= k <= %d and 0 <= n-k <= 255\n", RSGPU_MAX_K);
        exit(0);
    }
    FILE *fp_in = fopen(fileName, "rb");
    if (!fp_in) {
        printf("Cannot open source file!\n"); /* src/encode.cu:308-312 */
        exit(0);
    }
    if (fseeko(fp_in, 0, SEEK_END) == -1) {
        printf("fseek error!\n");
        exit(0);
    }
    const long long totalSize = ftello(fp_in);
    const long long chunk = rs_chunk_size(totalSize, k);
    const size_t pitch = (size_t)chunk;

    uint8_t *dataBuf = NULL, *codeBuf = NULL;
    int rc;
    if ((rc = rsgpu_host_alloc((void **)&dataBuf, (size_t)k * pitch)) != RSGPU_OK)
        die_cuda("rsgpu_host_alloc(data)", rc);
    if ((rc = rsgpu_host_alloc((void **)&codeBuf, (size_t)p * pitch)) != RSGPU_OK)
        die_cuda("rsgpu_host_alloc(parity)", rc);

    /* fragments are consecutive ranges of the file: one sequential read fills all k rows */
    rewind(fp_in);
    size_t got = 0;
    while (got < (size_t)totalSize) {
        size_t n = fread(dataBuf + got, 1, (size_t)totalSize - got, fp_in);
        if (n == 0)
            break;
        got += n;
    }
    if (got != (size_t)totalSize) {
        printf("fread error!\n"); /* src/encode.cu:340-344 */
        exit(0);
    }
    fclose(fp_in);
    memset(dataBuf + got, 0, (size_t)k * pitch - got);

    int maxGrid = 65535;
    rsgpu_device_max_grid(0, &maxGrid);
    if (gridDimXSize > maxGrid || gridDimXSize <= 0) {
        printf("Valid grid size: (0, %d]\n", maxGrid); /* src/encode.cu:351-355 */
        gridDimXSize = 0;
    }

    uint8_t *total = malloc((size_t)(k + p) * k);
    rs_job job;
    memset(&job, 0, sizeof job);
    job.kind = RS_JOB_ENCODE;
    job.k = k;
    job.p = p;
    job.in = dataBuf;
    job.out = codeBuf;
    job.in_pitch = job.out_pitch = pitch;
    job.len = pitch;
    job.grid_cap = gridDimXSize;
    job.slots = streamNum;
    job.matrix_out = total;
    if (rs_run_job(&job) != RSGPU_OK) {
        fprintf(stderr, "%s\n", job.error_text);
        exit(EXIT_FAILURE);
    }
    printf("Total GPU encoding time using multiple devices: %fms\n", job.wall_ms); /* src/encode.cu:276 */

    const size_t nlen = strlen(fileName) + 32;
    char *name = malloc(nlen);
    snprintf(name, nlen, "%s.METADATA", fileName);
    if (rs_write_metadata(name, totalSize, p, k, total) != 0) {
        printf("Cannot open META file!\n"); /* src/encode.cu:64-68 */
        exit(0);
    }
    for (int i = 0; i < k + p; ++i) {
        snprintf(name, nlen, "_%d_%s", i, fileName); /* src/encode.cu:437,453 */
        FILE *fp_out = fopen(name, "wb");
        if (!fp_out) {
            printf("Cannot open output file!\n");
            exit(0);
        }
        const uint8_t *src = i < k ? dataBuf + (size_t)i * pitch : codeBuf + (size_t)(i - k) * pitch;
        if (fwrite(src, 1, pitch, fp_out) != pitch) {
            printf("fwrite error!\n");
            exit(0);
        }
        fclose(fp_out);
    }
    free(name);
    free(total);
    rsgpu_host_free(dataBuf);
    rsgpu_host_free(codeBuf);
}
