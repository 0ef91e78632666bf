/* metadata.c -- reading and writing "<file>.METADATA".
 *
 * Format (text): "<totalSize>\n" "<p> <k>\n", then -- GPU writer only -- the k identity rows and the
 * p generator rows, every entry printed as "%d " and each row ended by "\n" (reference
 * src/encode.cu:70-98).  The CPU program stops after line 2 (src/cpu-rs.c:473-474) and its reader
 * regenerates the matrix (src/cpu-rs.c:616-621); the GPU reader wants the matrix
 * (src/decode.cu:274-281).  We write the GPU form (a superset both readers accept) and read both.
 * Sizes are written with %lld, which reads back identically through the reference's %d for files
 * below 2 GiB.
 */
#include <stdio.h>
#include <stdlib.h>

#include "rs_host.h"

long long rs_chunk_size(long long total_size, int k)
{
    return (total_size + k - 1) / k; /* src/encode.cu:317 */
}

int rs_write_metadata(const char *path, long long total_size, int p, int k, const uint8_t *total_matrix)
{
    FILE *fp = fopen(path, "wb");
    if (!fp)
        return -1;
    fprintf(fp, "%lld\n", total_size);
    fprintf(fp, "%d %d\n", p, k);
    for (int r = 0; r < k + p; ++r) {
        for (int c = 0; c < k; ++c)
            fprintf(fp, "%d ", (int)total_matrix[(size_t)r * k + c]);
        fprintf(fp, "\n");
    }
    return fclose(fp) == 0 ? 0 : -1;
}

int rs_read_metadata(const char *path, rs_metadata *md)
{
    md->matrix = NULL;
    FILE *fp = fopen(path, "rb");
    if (!fp)
        return -1;
    if (fscanf(fp, "%lld", &md->total_size) != 1 || fscanf(fp, "%d %d", &md->p, &md->k) != 2 ||
        md->k < 1 || md->k > RSGPU_MAX_K || md->p < 0 || md->p > 255 || md->total_size < 0) {
        fclose(fp);
        return -2;
    }
    const size_t n = (size_t)(md->k + md->p) * md->k;
    uint8_t *m = malloc(n ? n : 1);
    size_t got = 0;
    int v;
    int bad = 0;
    while (m && got < n && fscanf(fp, "%d", &v) == 1) {
        if (v < 0 || v > 255) { /* not a field element: refuse rather than truncate silently */
            bad = 1;
            break;
        }
        m[got++] = (uint8_t)v;
    }
    fclose(fp);
    if (bad) {
        free(m);
        return -2;
    }
    if (m && got == n) {
        md->matrix = m;
    } else {
        free(m);
        if (got != 0)
            return -2; /* a truncated matrix is an error; no matrix at all is the CPU program's form */
    }
    return 0;
}

void rs_free_metadata(rs_metadata *md)
{
    free(md->matrix);
    md->matrix = NULL;
}
