/*
 * orc_math.h — pinned fp32 arithmetic of the CPU oracle (TEST INFRASTRUCTURE ONLY).
 *
 * The reference kernels are Metal Shading Language built with MTL_FAST_MATH = YES
 * (project.pbxproj:218,273): contraction, reassociation and the accuracy of exp() are
 * unspecified there and cannot be reproduced off an Apple GPU.  The oracle therefore
 * PINS one evaluation order, stated here and in DESIGN.md ("Pinned arithmetic"):
 *
 *   - this file is compiled with -ffp-contract=off -fno-fast-math, so every written
 *     * + - / rounds once (IEEE-754 binary32, round-to-nearest-even);
 *   - a fused multiply-add happens only where orc_fma() is written;
 *   - exp() is orc_expf() below: exp2 of (x * log2(e)) with a degree-5 polynomial, built
 *     from + * fma and one integer add, so any IEEE machine reproduces it bit for bit
 *     (max relative error 1.6e-7 against exp2 on the reduced range; measured in
 *     tests/test_oracle_math.py);
 *   - sqrt and / are the correctly rounded IEEE operations.
 *
 * The CUDA kernels restate the same sequence with __fmaf_rn/__fmul_rn/... so that sort
 * keys, tile lists, tile ranges, lastContribIdx AND the forward image are bit-identical.
 */
#ifndef ORC_MATH_H
#define ORC_MATH_H
#include <stdint.h>
#include <string.h>
#include <math.h>

static inline float orc_fma(float a, float b, float c) { return __builtin_fmaf(a, b, c); }
static inline uint32_t orc_f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float orc_u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* MSL clamp(x, lo, hi) == min(max(x, lo), hi); fmin/fmax return the non-NaN operand. */
static inline float orc_clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

#define ORC_LOG2E   1.44269504088896340736f
#define ORC_EXP2_C1 0x1.62e42ap-1f
#define ORC_EXP2_C2 0x1.ebf9bcp-3f
#define ORC_EXP2_C3 0x1.c6b752p-5f
#define ORC_EXP2_C4 0x1.3cea88p-7f
#define ORC_EXP2_C5 0x1.5bba14p-10f
#define ORC_EXP2_MAGIC 12582912.0f /* 1.5 * 2^23 */

/* 2^t for |t| < 126.  r's low mantissa bits hold round(t); f = t - round(t) in [-0.5, 0.5]. */
static inline float orc_exp2f(float t) {
    float r = t + ORC_EXP2_MAGIC;
    float nf = r - ORC_EXP2_MAGIC;
    float f = t - nf;
    float p = ORC_EXP2_C5;
    p = orc_fma(p, f, ORC_EXP2_C4);
    p = orc_fma(p, f, ORC_EXP2_C3);
    p = orc_fma(p, f, ORC_EXP2_C2);
    p = orc_fma(p, f, ORC_EXP2_C1);
    p = orc_fma(p, f, 1.0f);
    return orc_u2f(orc_f2u(p) + (orc_f2u(r) << 23));
}
/* exp for a blend exponent already known to lie in [-4.5, 0] (no range clamp needed). */
static inline float orc_exp_blend(float x) { return orc_exp2f(x * ORC_LOG2E); }
/* general exp: the exponent is clamped to [-125, 125] first (NaN -> -125), so the integer
 * exponent add can never wrap. */
static inline float orc_expf(float x) {
    return orc_exp2f(fminf(fmaxf(x * ORC_LOG2E, -125.0f), 125.0f));
}

#endif
