/* decode_file.c -- drop-in for the reference's decode_file (src/decode.cu:235-434, declared in
 * src/decode.h:38): read "<in>.METADATA" and the conf file, take the k listed fragments, rebuild the
 * original bytes and write the first totalSize of them to the output.
 *
 * What differs underneath: the k x k inverse is computed on the GPU by one Gauss-Jordan block with
 * row pivoting (the reference calls the CPU routine, src/decode.cu:333, whose column swap is wrong
 * whenever a pivot is missing); only the ERASED natives are recomputed (e x k rows of the inverse) --
 * natives that are among the survivors are written straight from the slab that was read; a singular
 * survivor set is reported instead of producing garbage; the work streams slab by slab through
 * stream.c (reads, GPU pipelines and writes overlapped, O(ring) memory); sizes are 64-bit.
 */
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#include "decode.h"
#include "rs_host.h"

void decode_file(char *inFile, char *confFile, char *outFile, int gridDimXSize, int streamNum)
{
    const size_t nlen = strlen(inFile) + 32;
    char *name = malloc(nlen);
    snprintf(name, nlen, "%s.METADATA", inFile);
    rs_metadata md;
    const int mrc = rs_read_metadata(name, &md);
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

    rs_conf conf;
    const int crc = rs_read_conf(confFile, k, &conf);
    if (crc == -1) {
        printf("Cannot open configuration file!\n"); /* src/decode.cu:296-300 */
        exit(0);
    }
    if (crc != 0) {
        printf("Configuration file lists fewer than %d fragments!\n", k);
        exit(0);
    }

    int *native_row = malloc(sizeof(int) * (size_t)k); /* conf row that holds native j, or -1 */
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
        n_erased += native_row[j] < 0;

    rs_stream_src *src = calloc((size_t)k, sizeof *src);
    rs_stream_dst *dst = calloc((size_t)k, sizeof *dst);
    for (int i = 0; i < k; ++i) {
        src[i].fd = open(conf.name[i], O_RDONLY);
        if (src[i].fd < 0) {
            printf("Cannot open input file %s!\n", conf.name[i]); /* src/decode.cu:309-313 */
            exit(0);
        }
        src[i].base = 0; /* short fragment files read as zero-padded, src/decode.cu:290-291 */
    }

    const int gpu_budget = rs_file_gpu_budget(md.total_size);

    int maxGrid = 65535;
    struct timespec ts0, ts1;
    clock_gettime(CLOCK_MONOTONIC, &ts0);
    rsgpu_device_max_grid(0, &maxGrid); /* first CUDA call of the process: pays the driver / context start-up */
    clock_gettime(CLOCK_MONOTONIC, &ts1);
    if (getenv("RS_TIMING"))
        fprintf(stderr, "[rs timing] CUDA start-up: %.1f ms\n",
                1e3 * (double)(ts1.tv_sec - ts0.tv_sec) + 1e-6 * (double)(ts1.tv_nsec - ts0.tv_nsec));
    if (gridDimXSize > maxGrid || gridDimXSize <= 0) {
        printf("Valid grid size: (0, %d]\n", maxGrid); /* src/decode.cu:324-328 */
        gridDimXSize = 0;
    }

    /* the output is opened only after the inputs: without -o it replaces <inFile> (src/decode.cu:410-417) */
    const char *out_name = outFile ? outFile : inFile;
    char *tmp_name = malloc(strlen(out_name) + 16);
    sprintf(tmp_name, "%s.rs-tmp", out_name);
    const int fd_out = open(tmp_name, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (fd_out < 0) {
        printf("Cannot open output file %s!\n", out_name);
        exit(0);
    }
    /* all k natives land in ONE file: map it so the IO threads store into it in parallel (pwrite on a
     * single inode serialises; measured 3 GB/s on tmpfs).  Fall back to pwrite if the mapping fails. */
    /* The blocks are reserved first (posix_fallocate): a store into a sparse mapping of a full or
     * over-quota file system raises SIGBUS, which no error path can catch; if the space cannot be had the
     * pwrite path reports "fwrite error!" the ordinary way. */
    uint8_t *out_map = NULL;
    if (md.total_size > 0 && posix_fallocate(fd_out, 0, (off_t)md.total_size) == 0) {
        void *m = mmap(NULL, (size_t)md.total_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd_out, 0);
        if (m != MAP_FAILED)
            out_map = m;
    }
    int t = 0;
    for (int j = 0; j < k; ++j) {
        dst[j].fd = fd_out;
        dst[j].map = out_map;
        dst[j].base = (long long)j * chunk;
        long long lim = md.total_size - (long long)j * chunk; /* native j holds this many real bytes */
        dst[j].limit = lim < 0 ? 0 : (lim > chunk ? chunk : lim);
        if (native_row[j] >= 0) {
            dst[j].from_out = 0;
            dst[j].row = native_row[j];
        } else {
            dst[j].from_out = 1;
            dst[j].row = t++;
        }
    }

    rs_stream_desc d;
    memset(&d, 0, sizeof d);
    d.k = k;
    d.p = p;
    d.decode = 1;
    d.total_matrix = md.matrix; /* NULL for the CPU program's two-line METADATA: regenerate [I;V] */
    d.survivors = conf.index;
    d.in_rows = k;
    d.out_rows = n_erased;
    d.chunk = chunk;
    d.src = src;
    d.dst = dst;
    d.n_dst = k;
    d.grid_cap = gridDimXSize;
    d.slots = streamNum;
    d.max_gpus = gpu_budget;
    double wall_ms = 0;
    int used = 0;
    char err[512] = "";
    const int rc = rs_stream_run(&d, &wall_ms, &used, err, sizeof err);
    if (rc != RSGPU_OK) {
        close(fd_out);
        unlink(tmp_name);
        if (rc == RSGPU_ESINGULAR) {
            printf("Cannot decode: the listed fragments do not determine the data (singular matrix)\n");
            exit(1);
        }
        if (!strcmp(err, "fread error!") || !strcmp(err, "fwrite error!")) {
            printf("%s\n", err);
            exit(0);
        }
        fprintf(stderr, "%s\n", err);
        exit(EXIT_FAILURE);
    }
    printf("Total GPU decoding time using multiple devices: %fms\n", wall_ms); /* src/decode.cu:224 */

    if (out_map && (msync(out_map, (size_t)md.total_size, MS_SYNC) != 0 || munmap(out_map, (size_t)md.total_size) != 0)) {
        close(fd_out);
        unlink(tmp_name);
        printf("fwrite error!\n"); /* a late write-back failure is a failed write, not a success */
        exit(0);
    }
    if (ftruncate(fd_out, (off_t)md.total_size) != 0 || close(fd_out) != 0 || rename(tmp_name, out_name) != 0) {
        printf("fwrite error!\n");
        exit(0);
    }
    for (int i = 0; i < k; ++i)
        close(src[i].fd);
    free(tmp_name);
    free(native_row);
    free(src);
    free(dst);
    rs_free_conf(&conf);
    rs_free_metadata(&md);
}
