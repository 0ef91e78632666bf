/* sched.c -- how work is spread over the visible GPUs (no exchange between GPUs anywhere on the path).
 *
 * Replaces the arithmetic of the fan-out in the reference's encode_file / decode_file
 * (src/encode.cu:357-403, src/decode.cu:335-378).  The pipeline that uses it is stream.c.
 */
#include <stddef.h>

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

