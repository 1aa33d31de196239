/*
 * oracle_math.h — TEST INFRASTRUCTURE (CPU oracle).  Not part of the product.
 *
 * "Canonical" float32 transcendental functions for the oracle.
 *
 * Why they exist: the reference shader calls the HLSL intrinsics log, cos, sin,
 * pow and atan2 (resources/shaders/GPU-Utilities.hlsli:59-61,75-76,
 * RaytracerCS.hlsl:80,104).  D3D11 only bounds their error loosely (e.g. sin/cos
 * 0.0008 absolute, log 2^-21), every GPU implements them differently, and no CPU
 * libm reproduces any of them bit for bit.  A bit-level contract therefore has
 * to be *chosen*.  The contract used by this repository is the published Cephes
 * single-precision algorithms (S. Moshier, netlib cephes/single: logf.c, sinf.c,
 * expf.c, atanf.c) evaluated in plain IEEE float32 arithmetic with no FMA
 * contraction.  The CUDA kernels restate the same algorithms; this file is the
 * oracle's independent restatement.  tests/test_oracle_math.py pins these
 * against the host libm (double precision) within a few ulp, so the canonical
 * functions themselves are anchored to an independent implementation.
 *
 * The oracle can also run with the host libm (`math_mode = ORACLE_MATH_LIBM`);
 * GPU-vs-oracle parity is then statistical after the first bounce instead of
 * bit-exact.
 *
 * Compile with -ffp-contract=off (the Makefile does).
 */
#ifndef ORACLE_MATH_H
#define ORACLE_MATH_H

#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

namespace omath {

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

/* x = m * 2^e, m in [0.5, 1) — integer-only frexp for normal/denormal finite x > 0 */
static inline float frexp_pos(float x, int* e)
{
    uint32_t u = f2u(x);
    int bias = 0;
    if ((u & 0x7F800000u) == 0u) {          /* denormal: scale by 2^24 (exact) */
        x = x * 16777216.0f;
        u = f2u(x);
        bias = -24;
    }
    *e = (int)((u >> 23) & 0xFFu) - 126 + bias;
    return u2f((u & 0x807FFFFFu) | 0x3F000000u);
}

/* z * 2^n for results that stay normal or flush gradually; two-step to avoid
 * overflow of the scale factor */
static inline float ldexp_f(float z, int n)
{
    if (n > 127)  { z = z * 1.7014118346046923e38f; /* 2^127 */ n -= 127; if (n > 127) n = 127; }
    if (n < -126) { z = z * 1.1754943508222875e-38f; /* 2^-126 */ n += 126; if (n < -126) n = -126; }
    return z * u2f((uint32_t)(n + 127) << 23);
}

/* Cephes logf */
static inline float c_log(float x)
{
    if (x != x) return x;
    if (x < 0.0f) return std::numeric_limits<float>::quiet_NaN();
    if (x == 0.0f) return -std::numeric_limits<float>::infinity();
    if (x == std::numeric_limits<float>::infinity()) return x;

    int e;
    float m = frexp_pos(x, &e);
    if (m < 0.707106781186547524f) { e -= 1; m = (m + m) - 1.0f; }
    else { m = m - 1.0f; }

    float z = m * m;
    float y = 7.0376836292E-2f;
    y = y * m + -1.1514610310E-1f;
    y = y * m + 1.1676998740E-1f;
    y = y * m + -1.2420140846E-1f;
    y = y * m + 1.4249322787E-1f;
    y = y * m + -1.6668057665E-1f;
    y = y * m + 2.0000714765E-1f;
    y = y * m + -2.4999993993E-1f;
    y = y * m + 3.3333331174E-1f;
    y = y * m;
    y = y * z;

    float fe = (float)e;
    y = y + -2.12194440e-4f * fe;
    y = y + -0.5f * z;
    z = m + y;
    z = z + 0.693359375f * fe;
    return z;
}

/* Cephes expf */
static inline float c_exp(float x)
{
    if (x != x) return x;
    if (x > 88.72283905206835f) return std::numeric_limits<float>::infinity();
    if (x < -103.278929903431851103f) return 0.0f;

    float z = std::floor(1.44269504088896341f * x + 0.5f);
    x = x - z * 0.693359375f;
    x = x - z * -2.12194440e-4f;
    int n = (int)z;

    z = x * x;
    float p = 1.9875691500E-4f;
    p = p * x + 1.3981999507E-3f;
    p = p * x + 8.3334519073E-3f;
    p = p * x + 4.1665795894E-2f;
    p = p * x + 1.6666665459E-1f;
    p = p * x + 5.0000001201E-1f;
    z = p * z + x + 1.0f;
    return ldexp_f(z, n);
}

/* HLSL pow(x, y) == exp2(y * log2(x)): NaN for x < 0, 0 for x == 0 && y > 0 */
static inline float c_pow(float x, float y)
{
    if (x < 0.0f) return std::numeric_limits<float>::quiet_NaN();
    return c_exp(y * c_log(x));
}

static const float C_FOPI = 1.27323954473516f;
static const float C_DP1 = 0.78515625f;
static const float C_DP2 = 2.4187564849853515625e-4f;
static const float C_DP3 = 3.77489497744594108e-8f;

static inline float sin_poly(float x, float z)
{
    float y = -1.9515295891E-4f;
    y = y * z + 8.3321608736E-3f;
    y = y * z + -1.6666654611E-1f;
    y = y * z;
    y = y * x;
    return y + x;
}

static inline float cos_poly(float z)
{
    float y = 2.443315711809948E-005f;
    y = y * z + -1.388731625493765E-003f;
    y = y * z + 4.166664568298827E-002f;
    y = y * z;
    y = y * z;
    y = y - 0.5f * z;
    return y + 1.0f;
}

/* Cephes sinf (valid for |x| < 8192; beyond that still deterministic, less accurate) */
static inline float c_sin(float x)
{
    if (x != x || std::fabs(x) == std::numeric_limits<float>::infinity())
        return std::numeric_limits<float>::quiet_NaN();
    float sign = 1.0f;
    if (x < 0.0f) { sign = -1.0f; x = -x; }
    if (x > 16777215.0f) return 0.0f;
    uint32_t j = (uint32_t)(C_FOPI * x);
    float y = (float)j;
    if (j & 1u) { j += 1u; y = y + 1.0f; }
    j &= 7u;
    if (j > 3u) { sign = -sign; j -= 4u; }
    x = ((x - y * C_DP1) - y * C_DP2) - y * C_DP3;
    float z = x * x;
    float r = (j == 1u || j == 2u) ? cos_poly(z) : sin_poly(x, z);
    return sign < 0.0f ? -r : r;
}

/* Cephes cosf */
static inline float c_cos(float x)
{
    if (x != x || std::fabs(x) == std::numeric_limits<float>::infinity())
        return std::numeric_limits<float>::quiet_NaN();
    float sign = 1.0f;
    if (x < 0.0f) x = -x;
    if (x > 16777215.0f) return 0.0f;
    uint32_t j = (uint32_t)(C_FOPI * x);
    float y = (float)j;
    if (j & 1u) { j += 1u; y = y + 1.0f; }
    j &= 7u;
    if (j > 3u) { j -= 4u; sign = -sign; }
    if (j > 1u) sign = -sign;
    x = ((x - y * C_DP1) - y * C_DP2) - y * C_DP3;
    float z = x * x;
    float r = (j == 1u || j == 2u) ? sin_poly(x, z) : cos_poly(z);
    return sign < 0.0f ? -r : r;
}

/* Cephes atanf */
static inline float c_atan(float xx)
{
    if (xx != xx) return xx;
    float sign = 1.0f, x = xx;
    if (xx < 0.0f) { sign = -1.0f; x = -xx; }
    float y;
    if (x > 2.414213562373095f) { y = 1.5707963267948966192f; x = -(1.0f / x); }
    else if (x > 0.4142135623730950f) { y = 0.7853981633974483096f; x = (x - 1.0f) / (x + 1.0f); }
    else y = 0.0f;
    float z = x * x;
    float p = 8.05374449538e-2f;
    p = p * z - 1.38776856032E-1f;
    p = p * z + 1.99777106478E-1f;
    p = p * z - 3.33329491539E-1f;
    p = p * z;
    p = p * x;
    y = y + (p + x);
    return sign < 0.0f ? -y : y;
}

/* atan2(y, x) with the usual quadrant rules */
static inline float c_atan2(float y, float x)
{
    const float PI_F = 3.14159265358979323846f;
    const float PIO2_F = 1.5707963267948966192f;
    if (x != x || y != y) return std::numeric_limits<float>::quiet_NaN();
    if (x == 0.0f) {
        if (y > 0.0f) return PIO2_F;
        if (y < 0.0f) return -PIO2_F;
        return 0.0f;
    }
    if (y == 0.0f) return x > 0.0f ? 0.0f : PI_F;
    float w = 0.0f;
    if (x < 0.0f) w = (y < 0.0f) ? -PI_F : PI_F;
    return w + c_atan(y / x);
}

} // namespace omath
#endif
