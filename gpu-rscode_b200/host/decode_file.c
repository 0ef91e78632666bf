/* decode_file.c -- drop-in for the reference's decode_file (src/decode.cu:235-434, declared in
 * src/decode.h:38): read "<in>.METADATA" and the conf file, load the k listed fragments, rebuild the
 * original bytes and write the first totalSize of them to the output.
 *
 * What differs underneath: the k x k inverse is computed on the GPU by one Gauss-Jordan block with
 * row pivoting (the reference calls the CPU routine, src/decode.cu:333, whose column swap is wrong
 * whenever a pivot is missing); only the ERASED natives are recomputed (e x k rows of the inverse) --
 * natives that are among the survivors are written straight from the fragment that was read; a
 * singular survivor set is reported instead of producing garbage; sizes are 64-bit.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "decode.h"
#include "rs_host.h"

void decode_file(char *inFile, char *confFile, char *outFile, int gridDimXSize, int streamNum)
{
    const size_t nlen = strlen(inFile) + 32;
    char *name = malloc(nlen);
    snprintf(name, nlen, "%s.METADATA", inFile);
    rs_metadata md;
    int mrc = rs_read_metadata(name, &md);
    free(name);
    if (mrc == -1) {
        printf("Cannot open metadata file!\n"); /* src/decode.cu:259-263 */
        exit(0);
    }
    if (mrc != 0) {
        printf("Malformed metadata file!\n");
        exit(0);
    }
    const int k = md.k, p = md.p;
    const long long chunk = rs_chunk_size(md.total_size, k);
    const size_t pitch = (size_t)chunk;

    rs_conf conf;
    int crc = rs_read_conf(confFile, k, &conf);
    if (crc == -1) {
        printf("Cannot open configuration file!\n"); /* src/decode.cu:296-300 */
        exit(0);
    }
    if (crc != 0) {
        printf("Configuration file lists fewer than %d fragments!\n", k);
        exit(0);
    }

    int *native_row = malloc(sizeof(int) * (size_t)k); /* conf row that holds native j, or -1 */
    int *erased = malloc(sizeof(int) * (size_t)k);
    int n_erased = 0;
    for (int j = 0; j < k; ++j)
        native_row[j] = -1;
    for (int i = 0; i < k; ++i) {
        if (conf.index[i] < 0 || conf.index[i] >= k + p) {
            printf("Cannot decode: fragment index %d of %s is outside [0, %d)\n", conf.index[i], conf.name[i], k + p);
            exit(1);
        }
        if (conf.index[i] < k && native_row[conf.index[i]] < 0)
            native_row[conf.index[i]] = i;
    }
    for (int j = 0; j < k; ++j)
        if (native_row[j] < 0)
            erased[n_erased++] = j;

    uint8_t *codeBuf = NULL, *recBuf = NULL;
    int rc;
    if ((rc = rsgpu_host_alloc((void **)&codeBuf, (size_t)k * pitch)) != RSGPU_OK ||
        (rc = rsgpu_host_alloc((void **)&recBuf, (size_t)(n_erased ? n_erased : 1) * pitch)) != RSGPU_OK) {
        fprintf(stderr, "CUDA error in rsgpu_host_alloc: %s (%s)\n", rsgpu_strerror(rc), rsgpu_last_error());
        exit(EXIT_FAILURE);
    }
    memset(codeBuf, 0, (size_t)k * pitch); /* short fragment files read as zero-padded, src/decode.cu:290-291 */
    for (int i = 0; i < k; ++i) {
        FILE *fp_in = fopen(conf.name[i], "rb");
        if (!fp_in) {
            printf("Cannot open input file %s!\n", conf.name[i]); /* src/decode.cu:309-313 */
            exit(0);
        }
        size_t got = 0, n;
        while (got < pitch && (n = fread(codeBuf + (size_t)i * pitch + got, 1, pitch - got, fp_in)) > 0)
            got += n;
        fclose(fp_in);
    }

    int maxGrid = 65535;
    rsgpu_device_max_grid(0, &maxGrid);
    if (gridDimXSize > maxGrid || gridDimXSize <= 0) {
        printf("Valid grid size: (0, %d]\n", maxGrid); /* src/decode.cu:324-328 */
        gridDimXSize = 0;
    }

    rs_job job;
    memset(&job, 0, sizeof job);
    job.kind = RS_JOB_DECODE_ERASED;
    job.k = k;
    job.p = p;
    job.total_matrix = md.matrix; /* NULL for the CPU program's two-line METADATA: regenerate [I;V] */
    job.survivors = conf.index;
    job.in = codeBuf;
    job.out = recBuf;
    job.in_pitch = job.out_pitch = pitch;
    job.len = pitch;
    job.grid_cap = gridDimXSize;
    job.slots = streamNum;
    if (rs_run_job(&job) != RSGPU_OK) {
        if (job.error == RSGPU_ESINGULAR) {
            printf("Cannot decode: the listed fragments do not determine the data (singular matrix)\n");
            exit(1);
        }
        fprintf(stderr, "%s\n", job.error_text);
        exit(EXIT_FAILURE);
    }
    printf("Total GPU decoding time using multiple devices: %fms\n", job.wall_ms); /* src/decode.cu:224 */

    const char *out_name = outFile ? outFile : inFile; /* src/decode.cu:410-424 */
    FILE *fp_out = fopen(out_name, "wb");
    if (!fp_out) {
        printf("Cannot open output file %s!\n", out_name);
        exit(0);
    }
    long long left = md.total_size;
    int t = 0;
    for (int j = 0; j < k && left > 0; ++j) {
        const uint8_t *src;
        if (native_row[j] >= 0) {
            src = codeBuf + (size_t)native_row[j] * pitch;
        } else {
            src = recBuf + (size_t)t * pitch;
        }
        if (native_row[j] < 0)
            ++t;
        const size_t n = left < chunk ? (size_t)left : pitch;
        if (fwrite(src, 1, n, fp_out) != n) {
            printf("fwrite error!\n");
            exit(0);
        }
        left -= (long long)n;
    }
    fclose(fp_out);

    rsgpu_host_free(codeBuf);
    rsgpu_host_free(recBuf);
    free(native_row);
    free(erased);
    rs_free_conf(&conf);
    rs_free_metadata(&md);
}
