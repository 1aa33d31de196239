/*
 * hlsl_shim.h — TEST INFRASTRUCTURE.  Just enough of the HLSL language surface for the reference's own shader text
 * (resources/shaders/RaytracerCS.hlsl, GPU-Utilities.hlsli, GPU-Structs.hlsli) to compile as C++ where it lies
 * (oracle/Makefile target `ref`, oracle/ref_shader_driver.cpp).  Included INSIDE namespace hlslref, after <cstdint>,
 * <cstring>, <cmath> and oracle_math.h.
 *
 * What the shader text fixes when it is compiled this way: every expression, operation order, branch, loop and constant it
 * writes down.  What this header fixes (the same conventions as the restatement in okrt_oracle.cpp, its Q-list; D3D11 leaves
 * them to the driver or the hardware):
 *   - vectors are float32 component-wise; dot = (x*x' + y*y') + z*z'; normalize(v) = v * (1 / sqrt(dot(v, v)));
 *     lerp(a, b, s) = a + s * (b - a); reflect(i, n) = i - n * (2 * dot(i, n)); min / max ignore NaN (fmin / fmax);
 *     saturate / clamp = min(max(x, lo), hi);
 *   - float4x4 in a cbuffer / structured buffer is column-major storage; mul(v, M)_j = ((v.x*M0j + v.y*M1j) + v.z*M2j) + v.w*M3j,
 *     the last term left out when v.w == 0 [Q1];
 *   - log, cos, sin, pow, atan2: the canonical float32 algorithms of oracle_math.h (D3D11 gives error bounds, not bits);
 *   - Texture2DArray.SampleLevel: bilinear over four UNORM8 texels with float32 weights, WRAP, slice index clamped [Q12];
 *     TextureCube.SampleLevel: D3D face selection, bilinear inside the face, edge texels clamp [Q13];
 *   - reading a structured buffer out of range gives zeros (D3D11); `half` is float (cs_5_0 without min-precision).
 */

typedef uint32_t uint;
typedef float half;
#define unorm

struct float2; struct float3; struct float4; struct uint2;

/* ---- swizzle views: trivially copyable overlays of the parent's components */
template <int N, int A, int B> struct SwzF2 {
    float d[N];
    operator float2() const;
    SwzF2& operator=(const float2& v);
    SwzF2& operator/=(const float2& v);
    SwzF2& operator*=(float s);
    SwzF2& operator-=(float s);
};
template <int N, int A, int B, int C> struct SwzF3 {
    float d[N];
    operator float3() const;
};
template <int N, int A, int B> struct SwzU2 {
    uint d[N];
    operator uint2() const;
};

struct uint2 { uint x, y; };
struct uint3 { union { struct { uint x, y, z; }; SwzU2<3, 0, 1> xy; }; };
template <int N, int A, int B> SwzU2<N, A, B>::operator uint2() const { uint2 r; r.x = d[A]; r.y = d[B]; return r; }

struct float2 {
    float x, y;
    float2() = default;
    float2(float x_, float y_) : x(x_), y(y_) {}
    explicit float2(const uint2& u) : x((float)u.x), y((float)u.y) {}
};
struct float3 {
    union {
        struct { float x, y, z; };
        struct { float r, g, b; };
        SwzF2<3, 0, 1> xy;
        SwzF3<3, 0, 1, 0> xyx;
    };
    float3() = default;
    float3(float s) : x(s), y(s), z(s) {}                                  /* HLSL scalar -> vector promotion */
    float3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    float3(const float2& v, float z_) : x(v.x), y(v.y), z(z_) {}
};
struct float4 {
    union {
        struct { float x, y, z, w; };
        struct { float r, g, b, a; };
        SwzF3<4, 0, 1, 2> xyz;
        SwzF3<4, 0, 1, 2> rgb;
    };
    float4() = default;
    float4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    float4(const float3& v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
};
static_assert(sizeof(float2) == 8 && sizeof(float3) == 12 && sizeof(float4) == 16 && sizeof(uint3) == 12, "tight packing");

template <int N, int A, int B> SwzF2<N, A, B>::operator float2() const { return float2(d[A], d[B]); }
template <int N, int A, int B> SwzF2<N, A, B>& SwzF2<N, A, B>::operator=(const float2& v) { d[A] = v.x; d[B] = v.y; return *this; }
template <int N, int A, int B> SwzF2<N, A, B>& SwzF2<N, A, B>::operator/=(const float2& v) { d[A] = d[A] / v.x; d[B] = d[B] / v.y; return *this; }
template <int N, int A, int B> SwzF2<N, A, B>& SwzF2<N, A, B>::operator*=(float s) { d[A] = d[A] * s; d[B] = d[B] * s; return *this; }
template <int N, int A, int B> SwzF2<N, A, B>& SwzF2<N, A, B>::operator-=(float s) { d[A] = d[A] - s; d[B] = d[B] - s; return *this; }
template <int N, int A, int B, int C> SwzF3<N, A, B, C>::operator float3() const { return float3(d[A], d[B], d[C]); }

/* ---- component-wise arithmetic */
inline float2 operator+(float2 a, float2 b) { return float2(a.x + b.x, a.y + b.y); }
inline float2 operator*(float2 a, float s) { return float2(a.x * s, a.y * s); }
inline float3 operator+(float3 a, float3 b) { return float3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline float3 operator-(float3 a, float3 b) { return float3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float3 operator*(float3 a, float3 b) { return float3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float3 operator*(float3 a, float s) { return float3(a.x * s, a.y * s, a.z * s); }
inline float3 operator*(float s, float3 a) { return float3(s * a.x, s * a.y, s * a.z); }
inline float3 operator/(float3 a, float s) { return float3(a.x / s, a.y / s, a.z / s); }
inline float3 operator/(float s, float3 a) { return float3(s / a.x, s / a.y, s / a.z); }
inline float3 operator-(float3 a, float s) { return float3(a.x - s, a.y - s, a.z - s); }
inline float3 operator-(float s, float3 a) { return float3(s - a.x, s - a.y, s - a.z); }
inline float3 operator-(float3 a) { return float3(-a.x, -a.y, -a.z); }
inline float3& operator+=(float3& a, float3 b) { a = a + b; return a; }
inline float3& operator*=(float3& a, float3 b) { a = a * b; return a; }
inline float3& operator*=(float3& a, float s) { a = a * s; return a; }
inline float4 operator/(float4 a, float s) { return float4(a.x / s, a.y / s, a.z / s, a.w / s); }
inline float4& operator+=(float4& a, float4 b) { a = float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); return a; }

/* ---- intrinsics */
inline float min(float a, float b) { return std::fmin(a, b); }
inline float max(float a, float b) { return std::fmax(a, b); }
inline float3 min(float3 a, float3 b) { return float3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline float3 max(float3 a, float3 b) { return float3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float abs(float a) { return std::fabs(a); }
inline float3 abs(float3 a) { return float3(abs(a.x), abs(a.y), abs(a.z)); }
inline float sqrt(float a) { return std::sqrt(a); }
inline float log(float a) { return omath::c_log(a); }
inline float cos(float a) { return omath::c_cos(a); }
inline float sin(float a) { return omath::c_sin(a); }
inline float pow(float a, float b) { return omath::c_pow(a, b); }
inline float atan2(float y, float x) { return omath::c_atan2(y, x); }
inline float clamp(float x, float lo, float hi) { return min(max(x, lo), hi); }
inline float saturate(float x) { return min(max(x, 0.0f), 1.0f); }
inline float3 saturate(float3 v) { return float3(saturate(v.x), saturate(v.y), saturate(v.z)); }
inline float4 saturate(float4 v) { return float4(saturate(v.x), saturate(v.y), saturate(v.z), saturate(v.w)); }
inline float frac(float x) { return x - std::floor(x); }
inline float3 frac(float3 v) { return float3(frac(v.x), frac(v.y), frac(v.z)); }
inline float dot(float3 a, float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float3 cross(float3 a, float3 b) { return float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(float3 a) { return std::sqrt(dot(a, a)); }
inline float3 normalize(float3 a) { const float inv = 1.0f / std::sqrt(dot(a, a)); return a * inv; }
inline float lerp(float a, float b, float s) { return a + s * (b - a); }
inline float3 lerp(float3 a, float3 b, float s) { return float3(lerp(a.x, b.x, s), lerp(a.y, b.y, s), lerp(a.z, b.z, s)); }
inline float3 reflect(float3 i, float3 n) { const float d2 = 2.0f * dot(i, n); return i - n * d2; }

/* ---- matrices */
struct float4x4 { float m[16]; };                     /* column-major storage: element (row k, column j) = m[j * 4 + k] */
struct float3x3 {
    float3 r0, r1, r2;                                /* the constructor takes rows */
    float3x3(float3 a, float3 b, float3 c) : r0(a), r1(b), r2(c) {}
};
inline float4 mul(float4 v, const float4x4& M)
{
    float r[4];
    for (int j = 0; j < 4; j++) {
        const float s = (v.x * M.m[j * 4 + 0] + v.y * M.m[j * 4 + 1]) + v.z * M.m[j * 4 + 2];
        r[j] = v.w == 0.0f ? s : s + v.w * M.m[j * 4 + 3];
    }
    return float4(r[0], r[1], r[2], r[3]);
}
inline float3 mul(float3 v, const float3x3& T)
{
    return float3((v.x * T.r0.x + v.y * T.r1.x) + v.z * T.r2.x,
                  (v.x * T.r0.y + v.y * T.r1.y) + v.z * T.r2.y,
                  (v.x * T.r0.z + v.y * T.r1.z) + v.z * T.r2.z);
}

/* ---- resources */
template <class T> struct StructuredBuffer {
    const T* data = nullptr; uint count = 0;
    T operator[](uint i) const { T z; if (i < count) std::memcpy(&z, data + i, sizeof(T)); else std::memset((void*)&z, 0, sizeof(T)); return z; }
};
template <class T> struct RWTexture2D {
    T* data = nullptr; uint width = 0;
    T& operator[](uint2 p) { return data[(size_t)p.y * width + p.x]; }
};
struct SamplerState {};

inline bool shimFinite(float x) { return x - x == 0.0f; }
inline float3 shimTexel(const uint8_t* p) { return float3((float)p[0] / 255.0f, (float)p[1] / 255.0f, (float)p[2] / 255.0f); }
inline float3 shimBilerp(float3 c00, float3 c10, float3 c01, float3 c11, float fx, float fy)
{
    const float wx0 = 1.0f - fx, wy0 = 1.0f - fy;
    const float3 top = c00 * wx0 + c10 * fx;
    const float3 bot = c01 * wx0 + c11 * fx;
    return top * wy0 + bot * fy;
}
template <class T> struct Texture2DArray {
    const uint8_t* data = nullptr; uint size = 0, layers = 0;          /* layers x size x size RGBA8 */
    float4 SampleLevel(SamplerState, float3 uvw, float) const
    {
        if (!data || layers == 0 || size == 0) return float4(0.f, 0.f, 0.f, 0.f);
        const uint idx = (uint)uvw.z;
        const uint layer = idx < layers ? idx : layers - 1;
        const int N = (int)size;
        const float u = shimFinite(uvw.x) ? uvw.x : 0.0f, v = shimFinite(uvw.y) ? uvw.y : 0.0f;
        float x = u * (float)N - 0.5f, y = v * (float)N - 0.5f;
        if (!(std::fabs(x) < 1.0e9f)) x = 0.0f;
        if (!(std::fabs(y) < 1.0e9f)) y = 0.0f;
        const float xf = std::floor(x), yf = std::floor(y);
        const float fx = x - xf, fy = y - yf;
        int x0 = (int)xf % N; if (x0 < 0) x0 += N;
        int y0 = (int)yf % N; if (y0 < 0) y0 += N;
        const int x1 = x0 + 1 == N ? 0 : x0 + 1;
        const int y1 = y0 + 1 == N ? 0 : y0 + 1;
        const uint8_t* base = data + (size_t)layer * N * N * 4;
        const float3 c = shimBilerp(shimTexel(base + ((size_t)y0 * N + x0) * 4), shimTexel(base + ((size_t)y0 * N + x1) * 4),
                                    shimTexel(base + ((size_t)y1 * N + x0) * 4), shimTexel(base + ((size_t)y1 * N + x1) * 4), fx, fy);
        return float4(c, 1.0f);
    }
};
struct TextureCube {
    const uint8_t* faces = nullptr; uint w = 0, h = 0;                  /* 6 x h x w RGBA8: +X -X +Y -Y +Z -Z */
    float4 SampleLevel(SamplerState, float3 d, float) const
    {
        if (!faces || w == 0 || h == 0) return float4(0.f, 0.f, 0.f, 0.f);
        const float ax = std::fabs(d.x), ay = std::fabs(d.y), az = std::fabs(d.z);
        int face; float sc, tc, ma;
        if (ax >= ay && ax >= az) { ma = ax; if (d.x >= 0) { face = 0; sc = -d.z; tc = -d.y; } else { face = 1; sc = d.z; tc = -d.y; } }
        else if (ay >= az)        { ma = ay; if (d.y >= 0) { face = 2; sc = d.x; tc = d.z; } else { face = 3; sc = d.x; tc = -d.z; } }
        else                      { ma = az; if (d.z >= 0) { face = 4; sc = d.x; tc = -d.y; } else { face = 5; sc = -d.x; tc = -d.y; } }
        if (!(ma > 0.0f) || !shimFinite(ma) || !shimFinite(sc) || !shimFinite(tc)) return float4(0.f, 0.f, 0.f, 0.f);
        const float u = (sc / ma + 1.0f) * 0.5f, v = (tc / ma + 1.0f) * 0.5f;
        const int W = (int)w, H = (int)h;
        const float x = u * (float)W - 0.5f, y = v * (float)H - 0.5f;
        const float xf = std::floor(x), yf = std::floor(y);
        const float fx = x - xf, fy = y - yf;
        int x0 = (int)xf, y0 = (int)yf, x1 = x0 + 1, y1 = y0 + 1;
        x0 = x0 < 0 ? 0 : (x0 > W - 1 ? W - 1 : x0); x1 = x1 < 0 ? 0 : (x1 > W - 1 ? W - 1 : x1);
        y0 = y0 < 0 ? 0 : (y0 > H - 1 ? H - 1 : y0); y1 = y1 < 0 ? 0 : (y1 > H - 1 ? H - 1 : y1);
        const uint8_t* base = faces + (size_t)face * W * H * 4;
        const float3 c = shimBilerp(shimTexel(base + ((size_t)y0 * W + x0) * 4), shimTexel(base + ((size_t)y0 * W + x1) * 4),
                                    shimTexel(base + ((size_t)y1 * W + x0) * 4), shimTexel(base + ((size_t)y1 * W + x1) * 4), fx, fy);
        return float4(c, 1.0f);
    }
};
