// cl_compat.h -- OpenCL-C source compatibility prelude.
//
// Lets Radiance user programs (samples/shader.cl, samples/shader2.cl style: OpenCL C 1.2
// with vector types, swizzles, implicit scalar->vector conversion, `__global`/`__kernel`
// qualifiers) compile as C++ in two places:
//   * NVRTC for sm_100a   (the product path: CreateShaderModule -> rd_shader_module_create)
//   * plain g++ on the host (only used by oracle/ to run the reference's own .cl sources)
//
// The feature list follows SURVEY.md Appendix B (what the reference's six .cl files use),
// reference: radiance/shader/{radiance,data,math,pbr}.cl and samples/shader{,2}.cl.
// Nothing in here is taken from the reference; these are the OpenCL C built-ins restated
// with plain IEEE operations in a fixed left-to-right order so that host and device agree
// bit-for-bit when the device side is built with --fmad=false.
#ifndef RD_CL_COMPAT_H
#define RD_CL_COMPAT_H

#if defined(__CUDACC_RTC__) || defined(__CUDACC__)
#define RD_DEVICE_BUILD 1
#define RD_FDIV(a, b) __fdiv_rn((a), (b))
#define RD_F __device__ __forceinline__
#define RD_DEV __device__
// the traversal core must be seen before the OpenCL spellings float4/uint4 are mapped onto the
// rd_* classes at the end of this file (it uses the CUDA built-in vector types for its loads)
#include "rd_traverse.h"
#else
#define RD_DEVICE_BUILD 0
#define RD_FDIV(a, b) ((a) / (b))
#define RD_F inline
#define RD_DEV
#include <math.h>
#include <float.h>
#include <stdio.h>
#include <stddef.h>
#endif

#ifndef FLT_MAX
#define FLT_MAX 3.402823466e+38f
#endif
#ifndef FLT_MIN
#define FLT_MIN 1.175494351e-38f
#endif
#ifndef FLT_EPSILON
#define FLT_EPSILON 1.192092896e-07f
#endif
#ifndef M_PI_F
#define M_PI_F 3.14159274101257f
#endif

typedef unsigned int uint;
typedef unsigned char uchar;
typedef unsigned short ushort;
typedef unsigned long ulong;

// ---------------------------------------------------------------------------------------
// Vector types. CUDA already owns the names float3/uint3/...; the OpenCL names are mapped
// onto rd_* classes by macros at the end of this block (after the last use of a CUDA name).
// ---------------------------------------------------------------------------------------
struct rd_float2;
struct rd_float3;
struct rd_float4;

// Swizzle proxies: plain arrays that alias the parent vector's storage inside a union.
template <int N, int A, int B>
struct rd_swz2 {
    float d[N];
    RD_F operator rd_float2() const;
    RD_F rd_swz2& operator=(const rd_float2& v);
};
template <int N, int A, int B, int C>
struct rd_swz3 {
    float d[N];
    RD_F operator rd_float3() const;
    RD_F rd_swz3& operator=(const rd_float3& v);
};
template <int OFF>
struct rd_swz4of16 {
    float d[16];
    RD_F operator rd_float4() const;
    RD_F rd_swz4of16& operator=(const rd_float4& v);
};

struct alignas(8) rd_float2 {
    union {
        struct { float x, y; };
        struct { float s0, s1; };
        float s[2];
    };
    RD_F rd_float2() {}
    RD_F rd_float2(float v) : x(v), y(v) {}
    RD_F rd_float2(double v) : x((float)v), y((float)v) {}
    RD_F rd_float2(int v) : x((float)v), y((float)v) {}
    template <class T0, class T1>
    RD_F rd_float2(T0 a, T1 b) : x((float)a), y((float)b) {}
    RD_F float& operator[](int i) { return s[i]; }
    RD_F const float& operator[](int i) const { return s[i]; }
};

struct alignas(16) rd_float3 {
    union {
        struct { float x, y, z; };
        struct { float r, g, b; };
        struct { float s0, s1, s2; };
        float s[4];
        rd_swz2<4, 0, 1> xy;
        rd_swz2<4, 1, 2> yz;
        rd_swz2<4, 0, 2> xz;
        rd_swz3<4, 0, 1, 2> xyz;
        rd_swz3<4, 0, 1, 2> rgb;
        rd_swz3<4, 2, 1, 0> zyx;
    };
    RD_F rd_float3() {}
    RD_F rd_float3(float v) : x(v), y(v), z(v) {}
    RD_F rd_float3(double v) : x((float)v), y((float)v), z((float)v) {}
    RD_F rd_float3(int v) : x((float)v), y((float)v), z((float)v) {}
    template <class T0, class T1, class T2>
    RD_F rd_float3(T0 a, T1 b, T2 c) : x((float)a), y((float)b), z((float)c) {}
    RD_F float& operator[](int i) { return s[i]; }
    RD_F const float& operator[](int i) const { return s[i]; }
};

struct alignas(16) rd_float4 {
    union {
        struct { float x, y, z, w; };
        struct { float r, g, b, a; };
        struct { float s0, s1, s2, s3; };
        float s[4];
        rd_swz2<4, 0, 1> xy;
        rd_swz2<4, 2, 3> zw;
        rd_swz3<4, 0, 1, 2> xyz;
        rd_swz3<4, 0, 1, 2> rgb;
        rd_swz3<4, 1, 2, 3> yzw;
    };
    RD_F rd_float4() {}
    RD_F rd_float4(float v) : x(v), y(v), z(v), w(v) {}
    RD_F rd_float4(double v) : x((float)v), y((float)v), z((float)v), w((float)v) {}
    RD_F rd_float4(int v) : x((float)v), y((float)v), z((float)v), w((float)v) {}
    template <class T0, class T1, class T2, class T3>
    RD_F rd_float4(T0 a, T1 b, T2 c, T3 e) : x((float)a), y((float)b), z((float)c), w((float)e) {}
    RD_F rd_float4(const rd_float3& v, float e) : x(v.x), y(v.y), z(v.z), w(e) {}
    RD_F float& operator[](int i) { return s[i]; }
    RD_F const float& operator[](int i) const { return s[i]; }
};

// float16: only ever used as a row-major 4x4 matrix (reference math.cl:4).
struct alignas(16) rd_float16 {
    union {
        struct { float s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, sa, sb, sc, sd, se, sf; };
        float s[16];
        rd_swz4of16<0> s0123;
        rd_swz4of16<4> s4567;
        rd_swz4of16<8> s89ab;
        rd_swz4of16<12> scdef;
    };
    RD_F rd_float16() {}
    RD_F rd_float16(float v) { for (int i = 0; i < 16; ++i) s[i] = v; }
    RD_F float& operator[](int i) { return s[i]; }
    RD_F const float& operator[](int i) const { return s[i]; }
};

struct alignas(8) rd_uint2 {
    union { struct { uint x, y; }; struct { uint s0, s1; }; uint s[2]; };
    RD_F rd_uint2() {}
    RD_F rd_uint2(uint v) : x(v), y(v) {}
    template <class T0, class T1> RD_F rd_uint2(T0 a, T1 b) : x((uint)a), y((uint)b) {}
    RD_F uint& operator[](int i) { return s[i]; }
};
struct alignas(16) rd_uint3 {
    union { struct { uint x, y, z; }; struct { uint s0, s1, s2; }; uint s[4]; };
    RD_F rd_uint3() {}
    RD_F rd_uint3(uint v) : x(v), y(v), z(v) {}
    RD_F rd_uint3(int v) : x((uint)v), y((uint)v), z((uint)v) {}
    RD_F rd_uint3(float v) : x((uint)v), y((uint)v), z((uint)v) {}
    template <class T0, class T1, class T2>
    RD_F rd_uint3(T0 a, T1 b, T2 c) : x((uint)a), y((uint)b), z((uint)c) {}
    RD_F uint& operator[](int i) { return s[i]; }
};
struct alignas(16) rd_uint4 {
    union { struct { uint x, y, z, w; }; struct { uint s0, s1, s2, s3; }; uint s[4]; };
    RD_F rd_uint4() {}
    RD_F rd_uint4(uint v) : x(v), y(v), z(v), w(v) {}
    RD_F rd_uint4(int v) : x((uint)v), y((uint)v), z((uint)v), w((uint)v) {}
    RD_F rd_uint4(float v) : x((uint)v), y((uint)v), z((uint)v), w((uint)v) {}
    template <class T0, class T1, class T2, class T3>
    RD_F rd_uint4(T0 a, T1 b, T2 c, T3 e) : x((uint)a), y((uint)b), z((uint)c), w((uint)e) {}
    RD_F uint& operator[](int i) { return s[i]; }
};
struct alignas(8) rd_int2 {
    union { struct { int x, y; }; int s[2]; };
    RD_F rd_int2() {}
    RD_F rd_int2(int v) : x(v), y(v) {}
    template <class T0, class T1> RD_F rd_int2(T0 a, T1 b) : x((int)a), y((int)b) {}
};
struct alignas(16) rd_int3 {
    union { struct { int x, y, z; }; int s[4]; };
    RD_F rd_int3() {}
    RD_F rd_int3(int v) : x(v), y(v), z(v) {}
    template <class T0, class T1, class T2>
    RD_F rd_int3(T0 a, T1 b, T2 c) : x((int)a), y((int)b), z((int)c) {}
};
struct alignas(16) rd_int4 {
    union { struct { int x, y, z, w; }; int s[4]; };
    RD_F rd_int4() {}
    RD_F rd_int4(int v) : x(v), y(v), z(v), w(v) {}
    template <class T0, class T1, class T2, class T3>
    RD_F rd_int4(T0 a, T1 b, T2 c, T3 e) : x((int)a), y((int)b), z((int)c), w((int)e) {}
};

// proxy <-> vector conversions
template <int N, int A, int B>
RD_F rd_swz2<N, A, B>::operator rd_float2() const { return rd_float2(d[A], d[B]); }
template <int N, int A, int B>
RD_F rd_swz2<N, A, B>& rd_swz2<N, A, B>::operator=(const rd_float2& v) { d[A] = v.x; d[B] = v.y; return *this; }
template <int N, int A, int B, int C>
RD_F rd_swz3<N, A, B, C>::operator rd_float3() const { return rd_float3(d[A], d[B], d[C]); }
template <int N, int A, int B, int C>
RD_F rd_swz3<N, A, B, C>& rd_swz3<N, A, B, C>::operator=(const rd_float3& v) { d[A] = v.x; d[B] = v.y; d[C] = v.z; return *this; }
template <int OFF>
RD_F rd_swz4of16<OFF>::operator rd_float4() const { return rd_float4(d[OFF], d[OFF + 1], d[OFF + 2], d[OFF + 3]); }
template <int OFF>
RD_F rd_swz4of16<OFF>& rd_swz4of16<OFF>::operator=(const rd_float4& v) { d[OFF] = v.x; d[OFF + 1] = v.y; d[OFF + 2] = v.z; d[OFF + 3] = v.w; return *this; }

// ---------------------------------------------------------------------------------------
// Operators. All are non-template free functions so that swizzle proxies convert implicitly.
// ---------------------------------------------------------------------------------------
#define RD_VEC_OPS(V, S, EXPAND_VV, EXPAND_VS, EXPAND_SV)                                   \
    RD_F V operator+(const V& a, const V& b) { return EXPAND_VV(+); }                       \
    RD_F V operator-(const V& a, const V& b) { return EXPAND_VV(-); }                       \
    RD_F V operator*(const V& a, const V& b) { return EXPAND_VV(*); }                       \
    RD_F V operator/(const V& a, const V& b) { return EXPAND_VV(/); }                       \
    RD_F V operator+(const V& a, S b) { return EXPAND_VS(+); }                              \
    RD_F V operator-(const V& a, S b) { return EXPAND_VS(-); }                              \
    RD_F V operator*(const V& a, S b) { return EXPAND_VS(*); }                              \
    RD_F V operator/(const V& a, S b) { return EXPAND_VS(/); }                              \
    RD_F V operator+(S a, const V& b) { return EXPAND_SV(+); }                              \
    RD_F V operator-(S a, const V& b) { return EXPAND_SV(-); }                              \
    RD_F V operator*(S a, const V& b) { return EXPAND_SV(*); }                              \
    RD_F V operator/(S a, const V& b) { return EXPAND_SV(/); }                              \
    RD_F V& operator+=(V& a, const V& b) { a = a + b; return a; }                           \
    RD_F V& operator-=(V& a, const V& b) { a = a - b; return a; }                           \
    RD_F V& operator*=(V& a, const V& b) { a = a * b; return a; }                           \
    RD_F V& operator/=(V& a, const V& b) { a = a / b; return a; }                           \
    RD_F V& operator+=(V& a, S b) { a = a + b; return a; }                                  \
    RD_F V& operator-=(V& a, S b) { a = a - b; return a; }                                  \
    RD_F V& operator*=(V& a, S b) { a = a * b; return a; }                                  \
    RD_F V& operator/=(V& a, S b) { a = a / b; return a; }

#define RD_VV2(op) rd_float2(a.x op b.x, a.y op b.y)
#define RD_VS2(op) rd_float2(a.x op b, a.y op b)
#define RD_SV2(op) rd_float2(a op b.x, a op b.y)
RD_VEC_OPS(rd_float2, float, RD_VV2, RD_VS2, RD_SV2)
#define RD_VV3(op) rd_float3(a.x op b.x, a.y op b.y, a.z op b.z)
#define RD_VS3(op) rd_float3(a.x op b, a.y op b, a.z op b)
#define RD_SV3(op) rd_float3(a op b.x, a op b.y, a op b.z)
RD_VEC_OPS(rd_float3, float, RD_VV3, RD_VS3, RD_SV3)
#define RD_VV4(op) rd_float4(a.x op b.x, a.y op b.y, a.z op b.z, a.w op b.w)
#define RD_VS4(op) rd_float4(a.x op b, a.y op b, a.z op b, a.w op b)
#define RD_SV4(op) rd_float4(a op b.x, a op b.y, a op b.z, a op b.w)
RD_VEC_OPS(rd_float4, float, RD_VV4, RD_VS4, RD_SV4)
#define RD_UVV3(op) rd_uint3(a.x op b.x, a.y op b.y, a.z op b.z)
#define RD_UVS3(op) rd_uint3(a.x op b, a.y op b, a.z op b)
#define RD_USV3(op) rd_uint3(a op b.x, a op b.y, a op b.z)
RD_VEC_OPS(rd_uint3, uint, RD_UVV3, RD_UVS3, RD_USV3)
#define RD_UVV4(op) rd_uint4(a.x op b.x, a.y op b.y, a.z op b.z, a.w op b.w)
#define RD_UVS4(op) rd_uint4(a.x op b, a.y op b, a.z op b, a.w op b)
#define RD_USV4(op) rd_uint4(a op b.x, a op b.y, a op b.z, a op b.w)
RD_VEC_OPS(rd_uint4, uint, RD_UVV4, RD_UVS4, RD_USV4)

RD_F rd_float2 operator-(const rd_float2& a) { return rd_float2(-a.x, -a.y); }
RD_F rd_float3 operator-(const rd_float3& a) { return rd_float3(-a.x, -a.y, -a.z); }
RD_F rd_float4 operator-(const rd_float4& a) { return rd_float4(-a.x, -a.y, -a.z, -a.w); }

// integer-only operators used by hash functions (reference math.cl:12-15)
RD_F rd_uint3 operator>>(const rd_uint3& a, uint n) { return rd_uint3(a.x >> n, a.y >> n, a.z >> n); }
RD_F rd_uint3 operator<<(const rd_uint3& a, uint n) { return rd_uint3(a.x << n, a.y << n, a.z << n); }
RD_F rd_uint3 operator^(const rd_uint3& a, const rd_uint3& b) { return rd_uint3(a.x ^ b.x, a.y ^ b.y, a.z ^ b.z); }
RD_F rd_uint3 operator&(const rd_uint3& a, const rd_uint3& b) { return rd_uint3(a.x & b.x, a.y & b.y, a.z & b.z); }
RD_F rd_uint3 operator|(const rd_uint3& a, const rd_uint3& b) { return rd_uint3(a.x | b.x, a.y | b.y, a.z | b.z); }
RD_F rd_uint3& operator^=(rd_uint3& a, const rd_uint3& b) { a = a ^ b; return a; }
RD_F rd_uint3& operator>>=(rd_uint3& a, uint n) { a = a >> n; return a; }
RD_F rd_uint4 operator>>(const rd_uint4& a, uint n) { return rd_uint4(a.x >> n, a.y >> n, a.z >> n, a.w >> n); }
RD_F rd_uint4 operator^(const rd_uint4& a, const rd_uint4& b) { return rd_uint4(a.x ^ b.x, a.y ^ b.y, a.z ^ b.z, a.w ^ b.w); }
RD_F rd_uint4& operator^=(rd_uint4& a, const rd_uint4& b) { a = a ^ b; return a; }

// float16 as matrix: only scaling by a scalar is used (reference math.cl:180)
RD_F rd_float16 operator*(const rd_float16& a, float b) { rd_float16 r; for (int i = 0; i < 16; ++i) r.s[i] = a.s[i] * b; return r; }
RD_F rd_float16 operator*(float b, const rd_float16& a) { rd_float16 r; for (int i = 0; i < 16; ++i) r.s[i] = b * a.s[i]; return r; }
RD_F rd_float16 operator+(const rd_float16& a, const rd_float16& b) { rd_float16 r; for (int i = 0; i < 16; ++i) r.s[i] = a.s[i] + b.s[i]; return r; }
RD_F rd_float16 operator-(const rd_float16& a, const rd_float16& b) { rd_float16 r; for (int i = 0; i < 16; ++i) r.s[i] = a.s[i] - b.s[i]; return r; }

// ---------------------------------------------------------------------------------------
// Built-in functions (OpenCL C 1.2 section 6.12) actually used by Radiance programs.
// Every multi-term expression is written left to right with no fused operations.
// ---------------------------------------------------------------------------------------
#if !RD_DEVICE_BUILD
// CUDA already provides scalar min/max overloads in the global namespace.
RD_F float min(float a, float b) { return b < a ? b : a; }
RD_F float max(float a, float b) { return a < b ? b : a; }
RD_F int min(int a, int b) { return b < a ? b : a; }
RD_F int max(int a, int b) { return a < b ? b : a; }
RD_F uint min(uint a, uint b) { return b < a ? b : a; }
RD_F uint max(uint a, uint b) { return a < b ? b : a; }
#endif
RD_F float rd_minf(float a, float b) { return b < a ? b : a; }
RD_F float rd_maxf(float a, float b) { return a < b ? b : a; }

RD_F float clamp(float v, float lo, float hi) { return rd_minf(rd_maxf(v, lo), hi); }
RD_F int clamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
RD_F float mix(float a, float b, float t) { return a + (b - a) * t; }
RD_F float sign(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); }
RD_F float radians(float d) { return d * (M_PI_F / 180.0f); }
RD_F float degrees(float r) { return r * (180.0f / M_PI_F); }
RD_F float step(float edge, float v) { return v < edge ? 0.0f : 1.0f; }

#define RD_MAP2(V, f) V(f(a.x), f(a.y))
#define RD_MAP3(V, f) V(f(a.x), f(a.y), f(a.z))
#define RD_MAP4(V, f) V(f(a.x), f(a.y), f(a.z), f(a.w))
#define RD_UNARY(f, sf)                                                                     \
    RD_F rd_float2 f(const rd_float2& a) { return RD_MAP2(rd_float2, sf); }                 \
    RD_F rd_float3 f(const rd_float3& a) { return RD_MAP3(rd_float3, sf); }                 \
    RD_F rd_float4 f(const rd_float4& a) { return RD_MAP4(rd_float4, sf); }
RD_UNARY(fabs, fabsf)
RD_UNARY(sqrt, sqrtf)
RD_UNARY(sin, sinf)
RD_UNARY(cos, cosf)
RD_UNARY(exp, expf)
RD_UNARY(log, logf)
RD_UNARY(floor, floorf)
RD_UNARY(ceil, ceilf)

#define RD_BINARY(f, sf)                                                                    \
    RD_F rd_float2 f(const rd_float2& a, const rd_float2& b) { return rd_float2(sf(a.x, b.x), sf(a.y, b.y)); } \
    RD_F rd_float3 f(const rd_float3& a, const rd_float3& b) { return rd_float3(sf(a.x, b.x), sf(a.y, b.y), sf(a.z, b.z)); } \
    RD_F rd_float4 f(const rd_float4& a, const rd_float4& b) { return rd_float4(sf(a.x, b.x), sf(a.y, b.y), sf(a.z, b.z), sf(a.w, b.w)); } \
    RD_F rd_float2 f(const rd_float2& a, float b) { return rd_float2(sf(a.x, b), sf(a.y, b)); } \
    RD_F rd_float3 f(const rd_float3& a, float b) { return rd_float3(sf(a.x, b), sf(a.y, b), sf(a.z, b)); } \
    RD_F rd_float4 f(const rd_float4& a, float b) { return rd_float4(sf(a.x, b), sf(a.y, b), sf(a.z, b), sf(a.w, b)); }
RD_BINARY(min, rd_minf)
RD_BINARY(max, rd_maxf)
RD_BINARY(fmin, fminf)
RD_BINARY(fmax, fmaxf)
RD_BINARY(pow, powf)

RD_F rd_float2 clamp(const rd_float2& v, float lo, float hi) { return rd_float2(clamp(v.x, lo, hi), clamp(v.y, lo, hi)); }
RD_F rd_float3 clamp(const rd_float3& v, float lo, float hi) { return rd_float3(clamp(v.x, lo, hi), clamp(v.y, lo, hi), clamp(v.z, lo, hi)); }
RD_F rd_float4 clamp(const rd_float4& v, float lo, float hi) { return rd_float4(clamp(v.x, lo, hi), clamp(v.y, lo, hi), clamp(v.z, lo, hi), clamp(v.w, lo, hi)); }
RD_F rd_float3 clamp(const rd_float3& v, const rd_float3& lo, const rd_float3& hi) { return rd_float3(clamp(v.x, lo.x, hi.x), clamp(v.y, lo.y, hi.y), clamp(v.z, lo.z, hi.z)); }
RD_F rd_float4 clamp(const rd_float4& v, const rd_float4& lo, const rd_float4& hi) { return rd_float4(clamp(v.x, lo.x, hi.x), clamp(v.y, lo.y, hi.y), clamp(v.z, lo.z, hi.z), clamp(v.w, lo.w, hi.w)); }

RD_F rd_float2 mix(const rd_float2& a, const rd_float2& b, float t) { return a + (b - a) * t; }
RD_F rd_float3 mix(const rd_float3& a, const rd_float3& b, float t) { return a + (b - a) * t; }
RD_F rd_float4 mix(const rd_float4& a, const rd_float4& b, float t) { return a + (b - a) * t; }
RD_F rd_float3 mix(const rd_float3& a, const rd_float3& b, const rd_float3& t) { return a + (b - a) * t; }
RD_F rd_float4 mix(const rd_float4& a, const rd_float4& b, const rd_float4& t) { return a + (b - a) * t; }

RD_F float dot(const rd_float2& a, const rd_float2& b) { return a.x * b.x + a.y * b.y; }
RD_F float dot(const rd_float3& a, const rd_float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RD_F float dot(const rd_float4& a, const rd_float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
RD_F rd_float3 cross(const rd_float3& a, const rd_float3& b)
{
    return rd_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
RD_F rd_float4 cross(const rd_float4& a, const rd_float4& b)
{
    return rd_float4(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x, 0.0f);
}
RD_F float length(const rd_float2& a) { return sqrtf(dot(a, a)); }
RD_F float length(const rd_float3& a) { return sqrtf(dot(a, a)); }
RD_F float length(const rd_float4& a) { return sqrtf(dot(a, a)); }
RD_F float distance(const rd_float3& a, const rd_float3& b) { return length(a - b); }
RD_F rd_float2 normalize(const rd_float2& a) { return a / length(a); }
RD_F rd_float3 normalize(const rd_float3& a) { return a / length(a); }
RD_F rd_float4 normalize(const rd_float4& a) { return a / length(a); }

// conversions
RD_F rd_float3 convert_float3(const rd_uint3& a) { return rd_float3((float)a.x, (float)a.y, (float)a.z); }
RD_F rd_float4 convert_float4(const rd_uint4& a) { return rd_float4((float)a.x, (float)a.y, (float)a.z, (float)a.w); }
RD_F rd_uint3 convert_uint3(const rd_float3& a) { return rd_uint3((uint)a.x, (uint)a.y, (uint)a.z); }
RD_F rd_uint4 convert_uint4(const rd_float4& a) { return rd_uint4((uint)a.x, (uint)a.y, (uint)a.z, (uint)a.w); }

// ---------------------------------------------------------------------------------------
// Work-item functions. The host flavour is an NDRange emulator hook (oracle only).
// ---------------------------------------------------------------------------------------
#if RD_DEVICE_BUILD
// Set by the launcher before each launch: first global id of this launch, total work-items,
// and the image width used for the optional 8x4 pixel-tile remap (0 = identity mapping).
extern "C" { __constant__ unsigned int rd_launch_info[8]; }
// Wavefront stages (csrc/rd_wavefront.h) run a path on whatever thread its queue entry lands on; the
// stage stores the path's work-item id here before calling user code and sets rd_launch_info[3].
__shared__ unsigned int rd_wf_gid[1024];   // one slot per thread of the largest block option "block" accepts
RD_F size_t get_global_id(uint dim)
{
    if (dim != 0) return 0;
    if (rd_launch_info[3]) return rd_wf_gid[threadIdx.x];
    unsigned int blk = blockIdx.x;
#ifdef RD_SM_LOCAL
    // ablation: blocks that are resident on one SM together (block b, b + S, b + 2S, ... of a wave of S * RD_SM_LOCAL blocks,
    // S = rd_launch_info[7] SMs) render neighbouring tiles, so the warps of an SM share the nodes in its L1
    {
        const unsigned int S = rd_launch_info[7], G = S * RD_SM_LOCAL;
        if (S && blk < (gridDim.x / G) * G) {
            const unsigned int g = blk / G, r = blk - g * G;
            blk = g * G + (r % S) * RD_SM_LOCAL + r / S;
        }
    }
#endif
    unsigned int t = rd_launch_info[0] + blk * blockDim.x + threadIdx.x;
    unsigned int w = rd_launch_info[2];
    if (w) {
        // t enumerates 8x4 pixel tiles row by row; lanes of a warp cover one tile.
        unsigned int tilesX = w >> 3;
        unsigned int tile = t >> 5, lane = t & 31u;
        unsigned int ty = tile / tilesX, tx = tile - ty * tilesX;
        // multi-GPU interleaved partition (rd_trace_rays_interleaved): this rank's local tile row ty is
        // row (band * world + rank) * bandRows + ty % bandRows of the frame, band = ty / bandRows
        const unsigned int world = rd_launch_info[4];
        if (world > 1u) {
            const unsigned int bandRows = rd_launch_info[6], band = ty / bandRows;
            ty = (band * world + rd_launch_info[5]) * bandRows + (ty - band * bandRows);
        }
        unsigned int px = (tx << 3) + (lane & 7u), py = (ty << 2) + (lane >> 3);
        return (size_t)py * w + px;
    }
    return t;
}
RD_F size_t get_global_size(uint dim) { return dim == 0 ? rd_launch_info[1] : 1; }
RD_F size_t get_local_id(uint) { return 0; }
RD_F size_t get_local_size(uint) { return 1; }
RD_F size_t get_group_id(uint dim) { return get_global_id(dim); }
RD_F size_t get_num_groups(uint dim) { return get_global_size(dim); }
#else
extern thread_local size_t rd_host_global_id;
extern size_t rd_host_global_size;
RD_F size_t get_global_id(uint dim) { return dim == 0 ? rd_host_global_id : 0; }
RD_F size_t get_global_size(uint dim) { return dim == 0 ? rd_host_global_size : 1; }
RD_F size_t get_local_id(uint) { return 0; }
RD_F size_t get_local_size(uint) { return 1; }
RD_F size_t get_group_id(uint dim) { return get_global_id(dim); }
RD_F size_t get_num_groups(uint dim) { return get_global_size(dim); }
#endif

// ---------------------------------------------------------------------------------------
// Images and samplers. An image2d_array_t is a pointer to a small descriptor that lives in
// the same memory space as the texels (HBM on the device, heap in the host oracle); a
// sampler_t is a 64-bit token: [15:0] CL_ADDRESS_*, [31:16] CL_FILTER_*, [32] normalised.
// Reference: RD::CreateImageArray / RD::CreateSampler (radiance/src/radiance.cpp:96-131),
// layered RGBA8 read through read_imageui (samples/shader2.cl:260-264).
// ---------------------------------------------------------------------------------------
struct rd_image_array_desc {
    const uchar* texels;        // RGBA8, layer-major, row-major inside a layer
    unsigned int width, height, layers, pad;
};
typedef const rd_image_array_desc* image2d_array_t;
typedef unsigned long long sampler_t;

#define RD_CL_ADDRESS_NONE            0x1130u
#define RD_CL_ADDRESS_CLAMP_TO_EDGE   0x1131u
#define RD_CL_ADDRESS_CLAMP           0x1132u
#define RD_CL_ADDRESS_REPEAT          0x1133u
#define RD_CL_ADDRESS_MIRRORED_REPEAT 0x1134u
#define RD_CL_FILTER_NEAREST          0x1140u
#define RD_CL_FILTER_LINEAR           0x1141u

RD_F int rd_img_wrap(int i, int n, uint mode, bool* border)
{
    if (mode == RD_CL_ADDRESS_REPEAT) {
        i %= n; if (i < 0) i += n; return i;
    }
    if (mode == RD_CL_ADDRESS_MIRRORED_REPEAT) {
        int p = 2 * n; i %= p; if (i < 0) i += p; return i < n ? i : p - 1 - i;
    }
    if (mode == RD_CL_ADDRESS_CLAMP) {
        if (i < 0 || i >= n) { *border = true; return 0; }
        return i;
    }
    return i < 0 ? 0 : (i >= n ? n - 1 : i);   // CLAMP_TO_EDGE and NONE
}

RD_F rd_uint4 rd_img_fetch(image2d_array_t img, int x, int y, int layer, uint mode)
{
    bool border = false;
    x = rd_img_wrap(x, (int)img->width, mode, &border);
    y = rd_img_wrap(y, (int)img->height, mode, &border);
    if (border) return rd_uint4(0u, 0u, 0u, 0u);
    size_t o = (((size_t)layer * img->height + (size_t)y) * img->width + (size_t)x) * 4;
    const uchar* p = img->texels + o;
    return rd_uint4((uint)p[0], (uint)p[1], (uint)p[2], (uint)p[3]);
}

RD_F rd_uint4 read_imageui(image2d_array_t img, sampler_t smp, const rd_float4& coord)
{
    uint mode = (uint)(smp & 0xffffu), filter = (uint)((smp >> 16) & 0xffffu);
    bool normalised = ((smp >> 32) & 1u) != 0;
    int layer = (int)rintf(coord.z);   // OpenCL 1.2 8.2: l = clamp(rint(w), 0, d - 1), round half to even
    layer = layer < 0 ? 0 : (layer >= (int)img->layers ? (int)img->layers - 1 : layer);
    float u = coord.x, v = coord.y;
    if (normalised) {
        if (mode == RD_CL_ADDRESS_REPEAT) { u = u - floorf(u); v = v - floorf(v); }
        u = u * (float)img->width; v = v * (float)img->height;
    }
    if (filter != RD_CL_FILTER_LINEAR)
        return rd_img_fetch(img, (int)floorf(u), (int)floorf(v), layer, mode);
    float fu = u - 0.5f, fv = v - 0.5f;
    int i0 = (int)floorf(fu), j0 = (int)floorf(fv);
    float a = fu - floorf(fu), b = fv - floorf(fv);
    rd_uint4 t00 = rd_img_fetch(img, i0, j0, layer, mode), t10 = rd_img_fetch(img, i0 + 1, j0, layer, mode);
    rd_uint4 t01 = rd_img_fetch(img, i0, j0 + 1, layer, mode), t11 = rd_img_fetch(img, i0 + 1, j0 + 1, layer, mode);
    rd_uint4 r;
    for (int c = 0; c < 4; ++c) {
        float top = (1.0f - a) * (float)t00.s[c] + a * (float)t10.s[c];
        float bot = (1.0f - a) * (float)t01.s[c] + a * (float)t11.s[c];
        r.s[c] = (uint)((1.0f - b) * top + b * bot + 0.5f);
    }
    return r;
}
RD_F rd_float4 read_imagef(image2d_array_t img, sampler_t smp, const rd_float4& coord)
{
    rd_uint4 t = read_imageui(img, smp, coord);
    return rd_float4((float)t.x / 255.0f, (float)t.y / 255.0f, (float)t.z / 255.0f, (float)t.w / 255.0f);
}

// ---------------------------------------------------------------------------------------
// Qualifiers and the OpenCL spellings of the vector type names. Keep this block last.
// ---------------------------------------------------------------------------------------
#define __global
#define __constant const
#define __private
#define __local
#if !RD_DEVICE_BUILD
// Under CUDA `__global__` / `__constant__` are themselves macros for __attribute__((global)) /
// __attribute__((constant)), so the bare OpenCL keywords cannot be object-like macros there; the
// library rewrites the bare tokens `global`, `constant`, `kernel`, `local`, `private` in the user's
// source before handing it to NVRTC (csrc/rd_program.cu, rewrite_bare_qualifiers).
#define global
#define constant const
#define kernel
#endif
#define restrict __restrict__
#if RD_DEVICE_BUILD
#define __kernel extern "C" __global__
#else
#define __kernel
#endif

#define float2 rd_float2
#define float3 rd_float3
#define float4 rd_float4
#define float16 rd_float16
#define uint2 rd_uint2
#define uint3 rd_uint3
#define uint4 rd_uint4
#define int2 rd_int2
#define int3 rd_int3
#define int4 rd_int4

#endif  // RD_CL_COMPAT_H
