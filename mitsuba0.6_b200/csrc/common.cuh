// common.cuh -- device math for libmtscuda (sm_100a).
//
// fp32 vector helpers, the Mitsuba 0.6 scalar helpers the path-tracing hot path needs, and the
// counter-based RNG.  Each helper cites the reference routine whose results it must reproduce
// (paths relative to the reference root); the parity tests compare them through the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <math.h>

#define MTS_PREFIX mtscuda_
#include "../../include/mtscuda.h"

#define DEVFN __device__ __forceinline__
// Mid-size helpers that one path calls several times (microfacet terms, Fresnel terms, rough-transmittance lookups, texture
// lookups): kept out of line so that a kernel holds ONE copy instead of one per call site -- the shade kernels are bound by
// instruction fetch, not by call overhead.  -DMTS_OUTLINE=0 restores full inlining for A/B runs.
#ifndef MTS_OUTLINE
#define MTS_OUTLINE 1
#endif
#if MTS_OUTLINE
#define DEVOL static __device__ __noinline__
#define DEVOLM __device__ __noinline__
#else
#define DEVOL __device__ __forceinline__
#define DEVOLM __device__ __forceinline__
#endif
#define HDFN __host__ __device__ __forceinline__

namespace mts {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInvTwoPi = 0.15915494309189533577f;
constexpr float kEps = 1e-4f;        // Epsilon, include/mitsuba/core/constants.h:28
constexpr float kShadowEps = 1e-3f;  // ShadowEpsilon, :29
constexpr float kDeltaEps = 1e-3f;   // DeltaEpsilon, :31
constexpr float kInf = __builtin_huge_valf();

// ---- float3 algebra ----------------------------------------------------------------
HDFN float3 f3(float a) { return make_float3(a, a, a); }
HDFN float3 f3(float a, float b, float c) { return make_float3(a, b, c); }
HDFN float3 f3(const float4 &v) { return make_float3(v.x, v.y, v.z); }
HDFN float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
HDFN float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
HDFN float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
HDFN float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
HDFN float3 operator/(float3 a, float3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
HDFN float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
HDFN float3 operator*(float s, float3 a) { return a * s; }
// TVector3::operator/ multiplies by the reciprocal (include/mitsuba/core/vector.h:535-542)
HDFN float3 operator/(float3 a, float s) { float r = 1.0f / s; return a * r; }
HDFN float3 &operator+=(float3 &a, float3 b) { a = a + b; return a; }
HDFN float3 &operator*=(float3 &a, float3 b) { a = a * b; return a; }
HDFN float3 &operator*=(float3 &a, float s) { a = a * s; return a; }
HDFN float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HDFN float absDot(float3 a, float3 b) { return fabsf(dot(a, b)); }
HDFN float3 cross(float3 a, float3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
HDFN float length(float3 a) { return sqrtf(dot(a, a)); }
HDFN float3 normalize(float3 a) { return a / length(a); }
HDFN bool isZero(float3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }
HDFN float maxc(float3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
HDFN float average(float3 a) { return (a.x + a.y + a.z) * (1.0f / 3.0f); }
// RGB luminance, include/mitsuba/core/spectrum.h:726
HDFN float luminance(float3 a) { return a.x * 0.212671f + a.y * 0.715160f + a.z * 0.072169f; }
HDFN float comp(float3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// ---- scalar helpers -----------------------------------------------------------------
// math::fastexp / fastlog evaluate in double on Linux x86-64 (include/mitsuba/core/math.h:176-199).
// fp32 expf/logf of the CUDA math library are within 1-2 ulp of that, well inside the 1e-5 parity bar.
DEVFN float fastexp(float v) { return expf(v); }
DEVFN float fastlog(float v) { return logf(v); }
DEVFN float3 vexp(float3 a) { return f3(fastexp(a.x), fastexp(a.y), fastexp(a.z)); }
DEVFN float safe_sqrt(float v) { return sqrtf(fmaxf(0.0f, v)); }
DEVFN float3 vsafe_sqrt(float3 a) { return f3(safe_sqrt(a.x), safe_sqrt(a.y), safe_sqrt(a.z)); }
DEVFN float safe_acos(float v) { return acosf(fminf(1.0f, fmaxf(-1.0f, v))); }
DEVFN float signum(float v) { return copysignf(1.0f, v); }
DEVFN int floorToInt(float v) { return (int)floorf(v); }
DEVFN int modulo(int a, int b) { int r = a % b; return (r < 0) ? r + b : r; }
// modulo(x, size) for texture wrapping (result in [0, size)).  Power-of-two sizes use a mask; other sizes use a
// float reciprocal with an exact integer correction (texel coordinates are far below 2^23), so no integer division
// is ever executed, whichever side the compiler predicates.
DEVFN int wrapRepeat(int x, int size) {
    if ((size & (size - 1)) == 0) return x & (size - 1);
    int q = __float2int_rd((float)x * (1.0f / (float)size));
    int r = x - q * size;
    if (r < 0) r += size;
    if (r >= size) r -= size;
    return r;
}
DEVFN int clampi(int v, int lo, int hi) { return min(hi, max(lo, v)); }

// src/libcore/math.cpp:74-87
DEVFN float hypot2(float a, float b) {
    float r;
    if (fabsf(a) > fabsf(b)) { r = b / a; r = fabsf(a) * sqrtf(1.0f + r * r); }
    else if (b != 0.0f) { r = a / b; r = fabsf(b) * sqrtf(1.0f + r * r); }
    else r = 0.0f;
    return r;
}

// src/libcore/math.cpp:25-52
DEVFN float erfinv_giles(float x) {
    float w = -fastlog((1.0f - x) * (1.0f + x));
    float p;
    if (w < 5.0f) {
        w = w - 2.5f;
        p = 2.81022636e-08f;
        p = fmaf(p, w, 3.43273939e-07f);
        p = fmaf(p, w, -3.5233877e-06f);
        p = fmaf(p, w, -4.39150654e-06f);
        p = fmaf(p, w, 0.00021858087f);
        p = fmaf(p, w, -0.00125372503f);
        p = fmaf(p, w, -0.00417768164f);
        p = fmaf(p, w, 0.246640727f);
        p = fmaf(p, w, 1.50140941f);
    } else {
        w = sqrtf(w) - 3.0f;
        p = -0.000200214257f;
        p = fmaf(p, w, 0.000100950558f);
        p = fmaf(p, w, 0.00134934322f);
        p = fmaf(p, w, -0.00367342844f);
        p = fmaf(p, w, 0.00573950773f);
        p = fmaf(p, w, -0.0076224613f);
        p = fmaf(p, w, 0.00943887047f);
        p = fmaf(p, w, 1.00167406f);
        p = fmaf(p, w, 2.83297682f);
    }
    return p * x;
}

// src/libcore/math.cpp:54-72 (Abramowitz-Stegun 7.1.26; NOT the libm erf -- the Beckmann
// visible-normal inversion depends on this approximation)
DEVFN float erf_as(float x) {
    const float a1 = 0.254829592f, a2 = -0.284496736f, a3 = 1.421413741f, a4 = -1.453152027f, a5 = 1.061405429f;
    const float p = 0.3275911f;
    float sign = signum(x);
    x = fabsf(x);
    float t = 1.0f / (1.0f + p * x);
    float y = 1.0f - (((((a5 * t + a4) * t) + a3) * t + a2) * t + a1) * t * fastexp(-x * x);
    return sign * y;
}

// src/libcore/util.cpp:447-485
DEVFN bool solveQuadratic(float a, float b, float c, float &x0, float &x1) {
    if (a == 0) {
        if (b != 0) { x0 = x1 = -c / b; return true; }
        return false;
    }
    float discrim = b * b - 4.0f * a * c;
    if (discrim < 0) return false;
    float temp, sq = sqrtf(discrim);
    if (b < 0) temp = -0.5f * (b - sq); else temp = -0.5f * (b + sq);
    x0 = temp / a; x1 = c / temp;
    if (x0 > x1) { float t = x0; x0 = x1; x1 = t; }
    return true;
}
// src/libcore/util.cpp:487-525 (the sphere intersection is solved in fp64, src/shapes/sphere.cpp:163-187)
DEVFN bool solveQuadraticDouble(double a, double b, double c, double &x0, double &x1) {
    if (a == 0) {
        if (b != 0) { x0 = x1 = -c / b; return true; }
        return false;
    }
    double discrim = b * b - 4.0 * a * c;
    if (discrim < 0) return false;
    double temp, sq = sqrt(discrim);
    if (b < 0) temp = -0.5 * (b - sq); else temp = -0.5 * (b + sq);
    x0 = temp / a; x1 = c / temp;
    if (x0 > x1) { double t = x0; x0 = x1; x1 = t; }
    return true;
}

// src/libcore/util.cpp:592-601
DEVFN void coordinateSystem(float3 a, float3 &b, float3 &c) {
    if (fabsf(a.x) > fabsf(a.y)) {
        float invLen = 1.0f / sqrtf(a.x * a.x + a.z * a.z);
        c = f3(a.z * invLen, 0.0f, -a.x * invLen);
    } else {
        float invLen = 1.0f / sqrtf(a.y * a.y + a.z * a.z);
        c = f3(0.0f, a.z * invLen, -a.y * invLen);
    }
    b = cross(c, a);
}

// include/mitsuba/core/frame.h
struct Frame {
    float3 s, t, n;
    DEVFN float3 toLocal(float3 v) const { return f3(dot(v, s), dot(v, t), dot(v, n)); }
    DEVFN float3 toWorld(float3 v) const { return s * v.x + t * v.y + n * v.z; }
};
DEVFN Frame frameFromNormal(float3 n) { Frame f; f.n = n; coordinateSystem(n, f.s, f.t); return f; }
DEVFN float cosTheta(float3 v) { return v.z; }
DEVFN float cosTheta2(float3 v) { return v.z * v.z; }
DEVFN float sinTheta2(float3 v) { return 1.0f - v.z * v.z; }
DEVFN float tanTheta(float3 v) {
    float temp = 1 - v.z * v.z;
    if (temp <= 0.0f) return 0.0f;
    return sqrtf(temp) / v.z;
}
// src/libcore/util.cpp:603-608
DEVFN Frame shadingFrame(float3 n, float3 dpdu) {
    Frame f; f.n = n;
    f.s = normalize(dpdu - n * dot(n, dpdu));
    f.t = cross(f.n, f.s);
    return f;
}

// src/libcore/util.cpp:651-681
DEVOL float fresnelDielectricExt(float cosThetaI_, float &cosThetaT_, float eta) {
    if (eta == 1) { cosThetaT_ = -cosThetaI_; return 0.0f; }
    float scale = (cosThetaI_ > 0) ? 1 / eta : eta,
          cosThetaTSqr = 1 - (1 - cosThetaI_ * cosThetaI_) * (scale * scale);
    if (cosThetaTSqr <= 0.0f) { cosThetaT_ = 0.0f; return 1.0f; }
    float cosThetaI = fabsf(cosThetaI_);
    float cosThetaT = sqrtf(cosThetaTSqr);
    float Rs = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
    float Rp = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
    cosThetaT_ = (cosThetaI_ > 0) ? -cosThetaT : cosThetaT;
    return 0.5f * (Rs * Rs + Rp * Rp);
}
DEVFN float fresnelDielectricExt(float cosThetaI, float eta) { float t; return fresnelDielectricExt(cosThetaI, t, eta); }

// src/libcore/util.cpp:739-761
DEVOL float3 fresnelConductorExact(float cosThetaI, float3 eta, float3 k) {
    float cosThetaI2 = cosThetaI * cosThetaI, sinThetaI2 = 1 - cosThetaI2, sinThetaI4 = sinThetaI2 * sinThetaI2;
    float3 temp1 = eta * eta - k * k - f3(sinThetaI2),
           a2pb2 = vsafe_sqrt(temp1 * temp1 + k * k * eta * eta * 4.0f),
           a = vsafe_sqrt((a2pb2 + temp1) * 0.5f);
    float3 term1 = a2pb2 + f3(cosThetaI2), term2 = a * (2 * cosThetaI);
    float3 Rs2 = (term1 - term2) / (term1 + term2);
    float3 term3 = a2pb2 * cosThetaI2 + f3(sinThetaI4), term4 = term2 * sinThetaI2;
    float3 Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
    return 0.5f * (Rp2 + Rs2);
}

// ---- warps: src/libcore/warp.cpp:25-176, include/mitsuba/core/warp.h:55-76 ----
DEVFN float2 squareToUniformDiskConcentric(float2 s) {
    float r1 = 2.0f * s.x - 1.0f, r2 = 2.0f * s.y - 1.0f;
    float phi, r;
    if (r1 == 0 && r2 == 0) { r = phi = 0; }
    else if (r1 * r1 > r2 * r2) { r = r1; phi = (kPi / 4.0f) * (r2 / r1); }
    else { r = r2; phi = (kPi / 2.0f) - (r1 / r2) * (kPi / 4.0f); }
    float sp, cp; sincosf(phi, &sp, &cp);
    return make_float2(r * cp, r * sp);
}
DEVFN float3 squareToCosineHemisphere(float2 s) {
    float2 p = squareToUniformDiskConcentric(s);
    float z = safe_sqrt(1.0f - p.x * p.x - p.y * p.y);
    if (z == 0) z = 1e-10f;
    return f3(p.x, p.y, z);
}
DEVFN float squareToCosineHemispherePdf(float3 d) { return kInvPi * d.z; }
DEVFN float3 squareToUniformSphere(float2 s) {
    float z = 1.0f - 2.0f * s.y;
    float r = safe_sqrt(1.0f - z * z);
    float sp, cp; sincosf(2.0f * kPi * s.x, &sp, &cp);
    return f3(r * cp, r * sp, z);
}
DEVFN float3 squareToUniformCone(float cosCutoff, float2 s) {
    float ct = (1 - s.x) + s.x * cosCutoff;
    float st = safe_sqrt(1.0f - ct * ct);
    float sp, cp; sincosf(2.0f * kPi * s.y, &sp, &cp);
    return f3(cp * st, sp * st, ct);
}
DEVFN float squareToUniformConePdf(float cosCutoff) { return kInvTwoPi / (1 - cosCutoff); }
DEVFN float2 squareToUniformTriangle(float2 s) {
    float a = safe_sqrt(1.0f - s.x);
    return make_float2(1 - a, a * s.y);
}
DEVFN float intervalToTent(float s) {
    float sign;
    if (s < 0.5f) { sign = 1; s *= 2; } else { sign = -1; s = 2 * (s - 0.5f); }
    return sign * (1 - sqrtf(s));
}

// ---- affine 3x4 transforms (rows of a Mitsuba Transform) ---------------------------------
struct Affine { float4 r0, r1, r2; };
DEVFN float3 xfPoint(const Affine &m, float3 p) {
    return f3(m.r0.x * p.x + m.r0.y * p.y + m.r0.z * p.z + m.r0.w,
              m.r1.x * p.x + m.r1.y * p.y + m.r1.z * p.z + m.r1.w,
              m.r2.x * p.x + m.r2.y * p.y + m.r2.z * p.z + m.r2.w);
}
DEVFN float3 xfVec(const Affine &m, float3 v) {
    return f3(m.r0.x * v.x + m.r0.y * v.y + m.r0.z * v.z,
              m.r1.x * v.x + m.r1.y * v.y + m.r1.z * v.z,
              m.r2.x * v.x + m.r2.y * v.y + m.r2.z * v.z);
}

// ---- Philox4x32-10 counter-based RNG (the reference's `independent` sampler is iid U[0,1)
//      per dimension, src/samplers/independent.cpp:95-103; streams differ by design) ----------
DEVFN uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u; key.y += 0xBB67AE85u;
    }
    return ctr;
}
DEVFN float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }
// 4 uniforms for (pixel, global sample index, dimension block)
DEVFN float4 rand4(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t block) {
    uint4 r = philox4x32(make_uint4(pixel, sample, block, 0x6d747363u), make_uint2(seed, 0x63756461u));
    return make_float4(u01(r.x), u01(r.y), u01(r.z), u01(r.w));
}

// Same draw plus a fifth uniform assembled from the low 8 bits of the four Philox words, which u01() discards
// (24 independent bits): the extra number of EUsesSampler BSDFs (roughdielectric.cpp:405) costs nothing.
DEVFN float4 rand4x(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t block, float &extra) {
    uint4 r = philox4x32(make_uint4(pixel, sample, block, 0x6d747363u), make_uint2(seed, 0x63756461u));
    extra = (float)((r.x & 0xffu) | ((r.y & 0xffu) << 8) | ((r.z & 0xffu) << 16)) * (1.0f / 16777216.0f);
    return make_float4(u01(r.x), u01(r.y), u01(r.z), u01(r.w));
}

// power heuristic, src/integrators/path/path.cpp:296-300
DEVFN float miWeight(float pdfA, float pdfB) { pdfA *= pdfA; pdfB *= pdfB; return pdfA / (pdfA + pdfB); }

} // namespace mts
