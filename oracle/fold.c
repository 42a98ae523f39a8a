/*
 * oracle/fold.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the arithmetic of the reference hot path
 *   /root/reference/mnist/main.py:122-127  (Trainer.average_gradients)
 *     for p in parameters: all_reduce(p.grad, SUM); p.grad /= float(world_size)
 *   /root/reference/toy/main.py:19-20      (int32 all_reduce SUM, no divide)
 * with the cross-rank add order FIXED to ((g0 + g1) + g2) + ... (rank order), which is the
 * order the CUDA kernels are required to reproduce bit for bit.  The reference leaves the
 * order to its backend (gloo/NCCL, inside the torch wheel, not under /root/reference); at
 * world_size 2 every order is this order, at >= 3 gloo's ring differs in the last bits
 *  -- that comparison is done separately against oracle/gloo_path.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * load this file's shared object.  Build: oracle/Makefile (gcc -O2 -ffp-contract=off, no
 * -ffast-math: every add and the divide must be a single IEEE-754 binary32 operation).
 */
#include <stddef.h>
#include <stdint.h>

/* out[i] = (((g[0][i] + g[1][i]) + g[2][i]) + ...) / (float)world      mnist/main.py:126-127 */
void oracle_fold_average_f32(const float *const *g, int world, size_t n, float *out)
{
    const float w = (float)world;               /* float(world_size), mnist/main.py:127 */
    for (size_t i = 0; i < n; ++i) {
        float acc = g[0][i];
        for (int r = 1; r < world; ++r)
            acc = acc + g[r][i];                /* all_reduce(SUM), mnist/main.py:126 */
        out[i] = acc / w;                       /* true division, not a reciprocal multiply */
    }
}

/* out[i] = sum_r g[r][i] in int32 with wrap-around                    toy/main.py:19-20 */
void oracle_fold_sum_i32(const int32_t *const *g, int world, size_t n, int32_t *out)
{
    for (size_t i = 0; i < n; ++i) {
        uint32_t acc = (uint32_t)g[0][i];
        for (int r = 1; r < world; ++r)
            acc += (uint32_t)g[r][i];
        out[i] = (int32_t)acc;
    }
}

/*
 * Oracle of the OPT-IN in-switch variant only (GA_VARIANT_NVLS): the NVSwitch adds the `world`
 * replicas in a wide accumulator and rounds once (measured on this pool: 65 536 of 65 536 results
 * equal the correctly rounded exact sum, profiles/r02_nvlink_probe.md), then the kernel divides by
 * (float)world as mnist/main.py:127 does.  Restated as a binary64 sum rounded to binary32: exact
 * whenever the addends' exponents lie within 29 binades of each other, which holds for gradient
 * data; it is NOT the rank-order fold above and not what the reference's gloo ring computes either
 * (both differ from it in the last bit on about half of the elements).
 */
void oracle_fold_wide_average_f32(const float *const *g, int world, size_t n, float *out, int divide)
{
    const float w = (float)world;
    for (size_t i = 0; i < n; ++i) {
        double acc = (double)g[0][i];
        for (int r = 1; r < world; ++r)
            acc += (double)g[r][i];
        const float s = (float)acc;
        out[i] = divide ? s / w : s;
    }
}
