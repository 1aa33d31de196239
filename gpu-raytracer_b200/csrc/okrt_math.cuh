/*
 * okrt_math.cuh — device float32 transcendentals with a bit-level contract.
 *
 * The reference shader uses the HLSL intrinsics log / cos / sin / pow / atan2
 * (resources/shaders/GPU-Utilities.hlsli:59-61,75-76; RaytracerCS.hlsl:80,104).  D3D11 bounds their
 * error only loosely and every implementation differs, so the result bits are not defined by the
 * reference.  This build fixes them: the published Cephes single-precision algorithms (netlib
 * cephes/single logf.c, expf.c, sinf.c, atanf.c), evaluated with individually rounded float32
 * multiplies and adds (the translation unit is compiled with -fmad=false).  The CPU oracle restates
 * the same algorithms, which makes whole paths — not just primary hits — bit-comparable between the
 * GPU and the oracle.  CUDA's own __logf/cosf are not used: they are not reproducible on a CPU.
 */
#pragma once
#include <cstdint>

namespace okm {

__device__ __forceinline__ float nanf_() { return __int_as_float(0x7fc00000); }
__device__ __forceinline__ float inff_() { return __int_as_float(0x7f800000); }

__device__ __forceinline__ float ldexp_(float z, int n)
{
    if (n > 127)  { z = z * 1.7014118346046923e38f; n -= 127; if (n > 127) n = 127; }
    if (n < -126) { z = z * 1.1754943508222875e-38f; n += 126; if (n < -126) n = -126; }
    return z * __int_as_float((n + 127) << 23);
}

__device__ __forceinline__ float log_(float x)
{
    if (x != x) return x;
    if (x < 0.0f) return nanf_();
    if (x == 0.0f) return -inff_();
    if (x == inff_()) return x;

    uint32_t u = (uint32_t)__float_as_int(x);
    int bias = 0;
    if ((u & 0x7F800000u) == 0u) { x = x * 16777216.0f; u = (uint32_t)__float_as_int(x); bias = -24; }
    int e = (int)((u >> 23) & 0xFFu) - 126 + bias;
    float m = __int_as_float((int)((u & 0x807FFFFFu) | 0x3F000000u));

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

    const float fe = (float)e;
    y = y + -2.12194440e-4f * fe;
    y = y + -0.5f * z;
    z = m + y;
    z = z + 0.693359375f * fe;
    return z;
}

__device__ __forceinline__ float exp_(float x)
{
    if (x != x) return x;
    if (x > 88.72283905206835f) return inff_();
    if (x < -103.278929903431851103f) return 0.0f;

    float z = floorf(1.44269504088896341f * x + 0.5f);
    x = x - z * 0.693359375f;
    x = x - z * -2.12194440e-4f;
    const int n = (int)z;

    z = x * x;
    float p = 1.9875691500E-4f;
    p = p * x + 1.3981999507E-3f;
    p = p * x + 8.3334519073E-3f;
    p = p * x + 4.1665795894E-2f;
    p = p * x + 1.6666665459E-1f;
    p = p * x + 5.0000001201E-1f;
    z = p * z + x + 1.0f;
    return ldexp_(z, n);
}

/* HLSL pow(x,y) = exp2(y*log2(x)): NaN for negative x */
__device__ __forceinline__ float pow_(float x, float y)
{
    if (x < 0.0f) return nanf_();
    return exp_(y * log_(x));
}

#define OKM_FOPI 1.27323954473516f
#define OKM_DP1 0.78515625f
#define OKM_DP2 2.4187564849853515625e-4f
#define OKM_DP3 3.77489497744594108e-8f

__device__ __forceinline__ float sinPoly(float x, float z)
{
    float y = -1.9515295891E-4f;
    y = y * z + 8.3321608736E-3f;
    y = y * z + -1.6666654611E-1f;
    y = y * z;
    y = y * x;
    return y + x;
}

__device__ __forceinline__ float cosPoly(float z)
{
    float y = 2.443315711809948E-005f;
    y = y * z + -1.388731625493765E-003f;
    y = y * z + 4.166664568298827E-002f;
    y = y * z;
    y = y * z;
    y = y - 0.5f * z;
    return y + 1.0f;
}

__device__ __forceinline__ float sin_(float x)
{
    if (x != x || fabsf(x) == inff_()) return nanf_();
    float sign = 1.0f;
    if (x < 0.0f) { sign = -1.0f; x = -x; }
    if (x > 16777215.0f) return 0.0f;
    uint32_t j = (uint32_t)(OKM_FOPI * x);
    float y = (float)j;
    if (j & 1u) { j += 1u; y = y + 1.0f; }
    j &= 7u;
    if (j > 3u) { sign = -sign; j -= 4u; }
    x = ((x - y * OKM_DP1) - y * OKM_DP2) - y * OKM_DP3;
    const float z = x * x;
    const float r = (j == 1u || j == 2u) ? cosPoly(z) : sinPoly(x, z);
    return sign < 0.0f ? -r : r;
}

__device__ __forceinline__ float cos_(float x)
{
    if (x != x || fabsf(x) == inff_()) return nanf_();
    float sign = 1.0f;
    if (x < 0.0f) x = -x;
    if (x > 16777215.0f) return 0.0f;
    uint32_t j = (uint32_t)(OKM_FOPI * x);
    float y = (float)j;
    if (j & 1u) { j += 1u; y = y + 1.0f; }
    j &= 7u;
    if (j > 3u) { j -= 4u; sign = -sign; }
    if (j > 1u) sign = -sign;
    x = ((x - y * OKM_DP1) - y * OKM_DP2) - y * OKM_DP3;
    const float z = x * x;
    const float r = (j == 1u || j == 2u) ? sinPoly(x, z) : cosPoly(z);
    return sign < 0.0f ? -r : r;
}

__device__ __forceinline__ float atan_(float xx)
{
    if (xx != xx) return xx;
    float sign = 1.0f, x = xx;
    if (xx < 0.0f) { sign = -1.0f; x = -xx; }
    float y;
    if (x > 2.414213562373095f) { y = 1.5707963267948966192f; x = -(1.0f / x); }
    else if (x > 0.4142135623730950f) { y = 0.7853981633974483096f; x = (x - 1.0f) / (x + 1.0f); }
    else y = 0.0f;
    const float z = x * x;
    float p = 8.05374449538e-2f;
    p = p * z - 1.38776856032E-1f;
    p = p * z + 1.99777106478E-1f;
    p = p * z - 3.33329491539E-1f;
    p = p * z;
    p = p * x;
    y = y + (p + x);
    return sign < 0.0f ? -y : y;
}

__device__ __forceinline__ float atan2_(float y, float x)
{
    const float PI_F = 3.14159265358979323846f;
    const float PIO2_F = 1.5707963267948966192f;
    if (x != x || y != y) return nanf_();
    if (x == 0.0f) {
        if (y > 0.0f) return PIO2_F;
        if (y < 0.0f) return -PIO2_F;
        return 0.0f;
    }
    if (y == 0.0f) return x > 0.0f ? 0.0f : PI_F;
    float w = 0.0f;
    if (x < 0.0f) w = (y < 0.0f) ? -PI_F : PI_F;
    return w + atan_(y / x);
}

} // namespace okm
