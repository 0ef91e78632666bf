/* abi_example.c -- the C ABI of include/rsgpu.h used from plain C, no CUDA headers, no Python:
 * encode an RS(10,14) stripe on the GPU, lose four fragments, rebuild them, compare.
 *
 *   gcc -O2 -I include examples/abi_example.c -L gpu-rscode_b200 -lrsgpu -Wl,-rpath,$PWD/gpu-rscode_b200 -o abi_example
 *   ./abi_example [fragment_bytes]
 *
 * tests/test_gpu_cli.py builds and runs it on the GPU box.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rsgpu.h"

#define CHECK(call)                                                                          \
    do {                                                                                     \
        int rc__ = (call);                                                                   \
        if (rc__ != RSGPU_OK) {                                                              \
            fprintf(stderr, "%s -> %s (%s)\n", #call, rsgpu_strerror(rc__), rsgpu_last_error()); \
            return 1;                                                                        \
        }                                                                                    \
    } while (0)

int main(int argc, char **argv)
{
    const int k = 10, p = 4;
    const size_t len = argc > 1 ? (size_t)strtoull(argv[1], NULL, 10) : (size_t)1 << 20;
    rsgpu_ctx *ctx = NULL;
    CHECK(rsgpu_create(&ctx, 0, k, p));

    /* host stripe: k fragments of len bytes, any contents */
    uint8_t *data = malloc((size_t)k * len), *back = malloc((size_t)p * len);
    uint64_t x = 0x9E3779B97F4A7C15ull;
    for (size_t i = 0; i < (size_t)k * len; ++i) {
        x ^= x << 13, x ^= x >> 7, x ^= x << 17;
        data[i] = (uint8_t)x;
    }

    /* device buffers through the ABI (a caller with its own allocator passes its own pointers) */
    void *data_d, *parity_d, *surv_d, *lost_d;
    CHECK(rsgpu_dev_alloc(ctx, &data_d, (size_t)k * len));
    CHECK(rsgpu_dev_alloc(ctx, &parity_d, (size_t)p * len));
    CHECK(rsgpu_dev_alloc(ctx, &surv_d, (size_t)k * len));
    CHECK(rsgpu_dev_alloc(ctx, &lost_d, (size_t)p * len));
    CHECK(rsgpu_memcpy_h2d(ctx, data_d, data, (size_t)k * len));

    CHECK(rsgpu_encode(ctx, data_d, len, parity_d, len, len, NULL));
    CHECK(rsgpu_sync(ctx));

    /* natives 0..3 are gone: decode from the four parities and natives 4..9 (conf-file order) */
    int survivors[10] = {10, 11, 12, 13, 4, 5, 6, 7, 8, 9};
    uint8_t *stage = malloc((size_t)k * len);
    CHECK(rsgpu_memcpy_d2h(ctx, stage, parity_d, (size_t)p * len));
    memcpy(stage + (size_t)p * len, data + (size_t)4 * len, (size_t)(k - 4) * len);
    CHECK(rsgpu_memcpy_h2d(ctx, surv_d, stage, (size_t)k * len));
    int rc = rsgpu_set_survivors(ctx, survivors);
    if (rc == RSGPU_ESINGULAR) {
        fprintf(stderr, "these fragments cannot rebuild the data\n");
        return 2;
    }
    CHECK(rc);
    int erased[256], n_erased = 0;
    CHECK(rsgpu_get_erased(ctx, erased, &n_erased));
    CHECK(rsgpu_decode_erased(ctx, surv_d, len, lost_d, len, len, NULL));
    CHECK(rsgpu_memcpy_d2h(ctx, back, lost_d, (size_t)n_erased * len));

    int bad = 0;
    for (int i = 0; i < n_erased; ++i)
        bad |= memcmp(back + (size_t)i * len, data + (size_t)erased[i] * len, len) != 0;
    printf("RS(%d,%d), %zu bytes per fragment: %d lost natives rebuilt, %s; %lld kernel launches\n", k, k + p, len, n_erased,
           bad ? "MISMATCH" : "identical to the originals", rsgpu_launch_count(ctx));

    rsgpu_dev_free(ctx, data_d);
    rsgpu_dev_free(ctx, parity_d);
    rsgpu_dev_free(ctx, surv_d);
    rsgpu_dev_free(ctx, lost_d);
    rsgpu_destroy(ctx);
    free(data);
    free(back);
    free(stage);
    return bad;
}
