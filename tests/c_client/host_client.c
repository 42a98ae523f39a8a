/*
 * A consumer of libgradavg.so written in plain C99, with no CUDA header and no Python: the drop-in
 * boundary is pointers, sizes and ints.  Three ranks of a data-parallel job are hosted on GPU 0;
 * their gradients live in ordinary host memory (the reference's --no-cuda / gloo case,
 * mnist/main.py:181-182) and are averaged by ga_average_host.  The expected values are computed
 * right here the way mnist/main.py:126-127 defines them (rank-order sum, then / (float)world) and
 * compared bit for bit.   Build + run: tests/test_c_client.py.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "gradavg.h"

#define WORLD 3
#define NT 4

static unsigned lcg(unsigned *s) { *s = *s * 1664525u + 1013904223u; return *s; }

int main(void)
{
    const int64_t numel[NT] = {75, 1, 3729, 100003};         /* ragged, not multiples of 4 */
    float *grad[WORLD][NT], *want[NT];
    void *ptrs[WORLD * NT];
    unsigned seed = 12345u;
    ga_ctx *ctx = NULL;

    for (int i = 0; i < NT; ++i) {
        want[i] = (float *)malloc((size_t)numel[i] * sizeof(float));
        for (int r = 0; r < WORLD; ++r) {
            grad[r][i] = (float *)malloc((size_t)numel[i] * sizeof(float));
            for (int64_t k = 0; k < numel[i]; ++k)
                grad[r][i][k] = (float)((int)(lcg(&seed) >> 8) - (1 << 23)) / 65536.0f;
            ptrs[r * NT + i] = grad[r][i];
        }
        for (int64_t k = 0; k < numel[i]; ++k) {
            float acc = grad[0][i][k];
            for (int r = 1; r < WORLD; ++r) acc = acc + grad[r][i][k];
            want[i][k] = acc / (float)WORLD;
        }
    }
    if (ga_create_local(WORLD, 0, (size_t)1 << 20, &ctx) != GA_OK) {
        fprintf(stderr, "ga_create_local: %s\n", ga_last_error());
        return 2;
    }
    for (int variant = GA_VARIANT_AUTO; variant <= GA_VARIANT_TWOSHOT; ++variant) {
        float *work[WORLD][NT];
        void *wptrs[WORLD * NT];
        for (int r = 0; r < WORLD; ++r)
            for (int i = 0; i < NT; ++i) {
                work[r][i] = (float *)malloc((size_t)numel[i] * sizeof(float));
                memcpy(work[r][i], grad[r][i], (size_t)numel[i] * sizeof(float));
                wptrs[r * NT + i] = work[r][i];
            }
        if (ga_average_host(ctx, wptrs, numel, NT, NULL, variant) != GA_OK || ga_synchronize(ctx, NULL) != GA_OK) {
            fprintf(stderr, "variant %d: %s\n", variant, ga_last_error());
            return 3;
        }
        for (int r = 0; r < WORLD; ++r)
            for (int i = 0; i < NT; ++i) {
                if (memcmp(work[r][i], want[i], (size_t)numel[i] * sizeof(float)) != 0) {
                    fprintf(stderr, "variant %d rank %d tensor %d differs\n", variant, r, i);
                    return 4;
                }
                free(work[r][i]);
            }
    }
    /* error path: a set larger than the bucket is refused, nothing is launched */
    {
        const int64_t big[1] = {(int64_t)1 << 22};
        if (ga_average_host(ctx, ptrs, big, 1, NULL, GA_VARIANT_AUTO) != GA_ECAPACITY) return 5;
    }
    printf("C client OK: %d ranks, %d tensors, %llu launches, bit-exact\n", WORLD, NT,
           (unsigned long long)ga_launch_count(ctx));
    ga_destroy(ctx);
    return 0;
}
