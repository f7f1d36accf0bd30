/*
 * oracle_math.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Scalar CPU restatement of the Mitsuba 0.6 math helpers the `path` hot path
 * uses.  Every function names the reference file:line it follows (paths
 * relative to the reference root).  fp32 everywhere, like the reference's
 * SINGLE_PRECISION build; exp/log go through double like the reference's
 * Linux fastexp/fastlog (include/mitsuba/core/math.h:176-199).
 */
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

namespace orc {

static const float kPi = 3.14159265358979323846f;
static const float kInvPi = 0.31830988618379067154f;
static const float kInvTwoPi = 0.15915494309189533577f;
static const float kEps = 1e-4f;        /* constants.h:28 */
static const float kShadowEps = 1e-3f;  /* constants.h:29 */
static const float kDeltaEps = 1e-3f;   /* constants.h:31 */

struct V2 { float x, y; };
struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a) : x(a), y(a), z(a) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    float &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, V3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return a * s; }
/* Mitsuba's TVector3::operator/ multiplies by the reciprocal (include/mitsuba/core/vector.h) */
inline V3 operator/(V3 a, float s) { float r = 1.0f / s; return V3(a.x * r, a.y * r, a.z * r); }
inline V3 &operator+=(V3 &a, V3 b) { a = a + b; return a; }
inline V3 &operator*=(V3 &a, V3 b) { a = a * b; return a; }
inline V3 &operator*=(V3 &a, float s) { a = a * s; return a; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float absDot(V3 a, V3 b) { return std::fabs(dot(a, b)); }
inline V3 cross(V3 a, V3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return a / length(a); }
inline bool isZero(V3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }
inline float maxc(V3 a) { return std::max(a.x, std::max(a.y, a.z)); }
inline float average(V3 a) { return (a.x + a.y + a.z) * (1.0f / 3.0f); } /* spectrum.h average() */
/* include/mitsuba/core/spectrum.h:726 (RGB build) */
inline float luminance(V3 a) { return a.x * 0.212671f + a.y * 0.715160f + a.z * 0.072169f; }
/* spectrum.h:521-526: component-wise math::fastexp */
inline float fastexp(float v);
inline V3 vexp(V3 a);
inline V3 vsafe_sqrt(V3 a) {
    return V3(std::sqrt(std::max(0.0f, a.x)), std::sqrt(std::max(0.0f, a.y)), std::sqrt(std::max(0.0f, a.z)));
}

/* math.h:176-199 (Linux x86-64): exp/log in double */
inline float fastexp(float v) { return (float)::exp((double)v); }
inline float fastlog(float v) { return (float)::log((double)v); }
inline V3 vexp(V3 a) { return V3(fastexp(a.x), fastexp(a.y), fastexp(a.z)); }
inline float safe_sqrt(float v) { return std::sqrt(std::max(0.0f, v)); }
inline float safe_acos(float v) { return std::acos(std::min(1.0f, std::max(-1.0f, v))); }
inline float signum(float v) { return copysignf(1.0f, v); }
inline int floorToInt(float v) { return (int)std::floor(v); }
inline int modulo(int a, int b) { int r = a % b; return (r < 0) ? r + b : r; }
inline int clampi(int v, int lo, int hi) { return std::min(hi, std::max(lo, v)); }

/* src/libcore/math.cpp:74-87 */
inline float hypot2(float a, float b) {
    float r;
    if (std::fabs(a) > std::fabs(b)) { r = b / a; r = std::fabs(a) * std::sqrt(1.0f + r * r); }
    else if (b != 0.0f) { r = a / b; r = std::fabs(b) * std::sqrt(1.0f + r * r); }
    else r = 0.0f;
    return r;
}

/* src/libcore/math.cpp:25-52 (Giles' single-precision erfinv) */
inline float erfinv(float x) {
    float w = -fastlog((1.0f - x) * (1.0f + x));
    float p;
    if (w < 5.0f) {
        w = w - 2.5f;
        p = 2.81022636e-08f;
        p = 3.43273939e-07f + p * w;
        p = -3.5233877e-06f + p * w;
        p = -4.39150654e-06f + p * w;
        p = 0.00021858087f + p * w;
        p = -0.00125372503f + p * w;
        p = -0.00417768164f + p * w;
        p = 0.246640727f + p * w;
        p = 1.50140941f + p * w;
    } else {
        w = std::sqrt(w) - 3.0f;
        p = -0.000200214257f;
        p = 0.000100950558f + p * w;
        p = 0.00134934322f + p * w;
        p = -0.00367342844f + p * w;
        p = 0.00573950773f + p * w;
        p = -0.0076224613f + p * w;
        p = 0.00943887047f + p * w;
        p = 1.00167406f + p * w;
        p = 2.83297682f + p * w;
    }
    return p * x;
}

/* src/libcore/math.cpp:54-72 (A&S 7.1.26) */
inline float erf_as(float x) {
    const float a1 = 0.254829592f, a2 = -0.284496736f, a3 = 1.421413741f, a4 = -1.453152027f, a5 = 1.061405429f;
    const float p = 0.3275911f;
    float sign = signum(x);
    x = std::fabs(x);
    float t = 1.0f / (1.0f + p * x);
    float y = 1.0f - (((((a5 * t + a4) * t) + a3) * t + a2) * t + a1) * t * fastexp(-x * x);
    return sign * y;
}

/* src/libcore/util.cpp:447-485 */
inline bool solveQuadratic(float a, float b, float c, float &x0, float &x1) {
    if (a == 0) {
        if (b != 0) { x0 = x1 = -c / b; return true; }
        return false;
    }
    float discrim = b * b - 4.0f * a * c;
    if (discrim < 0) return false;
    float temp, sq = std::sqrt(discrim);
    if (b < 0) temp = -0.5f * (b - sq); else temp = -0.5f * (b + sq);
    x0 = temp / a; x1 = c / temp;
    if (x0 > x1) std::swap(x0, x1);
    return true;
}
/* src/libcore/util.cpp:487-525 */
inline bool solveQuadraticDouble(double a, double b, double c, double &x0, double &x1) {
    if (a == 0) {
        if (b != 0) { x0 = x1 = -c / b; return true; }
        return false;
    }
    double discrim = b * b - 4.0f * a * c;
    if (discrim < 0) return false;
    double temp, sq = std::sqrt(discrim);
    if (b < 0) temp = -0.5f * (b - sq); else temp = -0.5f * (b + sq);
    x0 = temp / a; x1 = c / temp;
    if (x0 > x1) std::swap(x0, x1);
    return true;
}

/* src/libcore/util.cpp:592-601 */
inline void coordinateSystem(V3 a, V3 &b, V3 &c) {
    if (std::fabs(a.x) > std::fabs(a.y)) {
        float invLen = 1.0f / std::sqrt(a.x * a.x + a.z * a.z);
        c = V3(a.z * invLen, 0.0f, -a.x * invLen);
    } else {
        float invLen = 1.0f / std::sqrt(a.y * a.y + a.z * a.z);
        c = V3(0.0f, a.z * invLen, -a.y * invLen);
    }
    b = cross(c, a);
}

/* include/mitsuba/core/frame.h */
struct Frame {
    V3 s, t, n;
    Frame() {}
    explicit Frame(V3 n_) : n(n_) { coordinateSystem(n, s, t); }
    Frame(V3 s_, V3 t_, V3 n_) : s(s_), t(t_), n(n_) {}
    V3 toLocal(V3 v) const { return V3(dot(v, s), dot(v, t), dot(v, n)); }
    V3 toWorld(V3 v) const { return s * v.x + t * v.y + n * v.z; }
    static float cosTheta(V3 v) { return v.z; }
    static float cosTheta2(V3 v) { return v.z * v.z; }
    static float sinTheta2(V3 v) { return 1.0f - v.z * v.z; }
    static float tanTheta(V3 v) {
        float temp = 1 - v.z * v.z;
        if (temp <= 0.0f) return 0.0f;
        return std::sqrt(temp) / v.z;
    }
};

/* src/libcore/util.cpp:603-608 */
inline void computeShadingFrame(V3 n, V3 dpdu, Frame &frame) {
    frame.n = n;
    frame.s = normalize(dpdu - frame.n * dot(frame.n, dpdu));
    frame.t = cross(frame.n, frame.s);
}

/* src/libcore/util.cpp:651-681 */
inline float fresnelDielectricExt(float cosThetaI_, float &cosThetaT_, float eta) {
    if (eta == 1) { cosThetaT_ = -cosThetaI_; return 0.0f; }
    float scale = (cosThetaI_ > 0) ? 1 / eta : eta,
          cosThetaTSqr = 1 - (1 - cosThetaI_ * cosThetaI_) * (scale * scale);
    if (cosThetaTSqr <= 0.0f) { cosThetaT_ = 0.0f; return 1.0f; }
    float cosThetaI = std::fabs(cosThetaI_);
    float cosThetaT = std::sqrt(cosThetaTSqr);
    float Rs = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
    float Rp = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
    cosThetaT_ = (cosThetaI_ > 0) ? -cosThetaT : cosThetaT;
    return 0.5f * (Rs * Rs + Rp * Rp);
}
inline float fresnelDielectricExt(float cosThetaI, float eta) {
    float t; return fresnelDielectricExt(cosThetaI, t, eta);
}

/* src/libcore/util.cpp:739-761 */
inline V3 fresnelConductorExact(float cosThetaI, V3 eta, V3 k) {
    float cosThetaI2 = cosThetaI * cosThetaI, sinThetaI2 = 1 - cosThetaI2, sinThetaI4 = sinThetaI2 * sinThetaI2;
    V3 temp1 = eta * eta - k * k - V3(sinThetaI2),
       a2pb2 = vsafe_sqrt(temp1 * temp1 + k * k * eta * eta * 4.0f),
       a = vsafe_sqrt((a2pb2 + temp1) * 0.5f);
    V3 term1 = a2pb2 + V3(cosThetaI2), term2 = a * (2 * cosThetaI);
    V3 Rs2 = (term1 - term2) / (term1 + term2);
    V3 term3 = a2pb2 * cosThetaI2 + V3(sinThetaI4), term4 = term2 * sinThetaI2;
    V3 Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
    return 0.5f * (Rp2 + Rs2);
}

/* ---- src/libcore/warp.cpp:25-176, include/mitsuba/core/warp.h:55-76 ---- */
inline V2 squareToUniformDiskConcentric(V2 s) {
    float r1 = 2.0f * s.x - 1.0f, r2 = 2.0f * s.y - 1.0f;
    float phi, r;
    if (r1 == 0 && r2 == 0) { r = phi = 0; }
    else if (r1 * r1 > r2 * r2) { r = r1; phi = (kPi / 4.0f) * (r2 / r1); }
    else { r = r2; phi = (kPi / 2.0f) - (r1 / r2) * (kPi / 4.0f); }
    float sp, cp; sincosf(phi, &sp, &cp);
    return V2{r * cp, r * sp};
}
inline V3 squareToCosineHemisphere(V2 s) {
    V2 p = squareToUniformDiskConcentric(s);
    float z = safe_sqrt(1.0f - p.x * p.x - p.y * p.y);
    if (z == 0) z = 1e-10f;
    return V3(p.x, p.y, z);
}
inline float squareToCosineHemispherePdf(V3 d) { return kInvPi * Frame::cosTheta(d); }
inline V3 squareToUniformSphere(V2 s) {
    float z = 1.0f - 2.0f * s.y;
    float r = safe_sqrt(1.0f - z * z);
    float sp, cp; sincosf(2.0f * kPi * s.x, &sp, &cp);
    return V3(r * cp, r * sp, z);
}
inline V3 squareToUniformCone(float cosCutoff, V2 s) {
    float cosTheta = (1 - s.x) + s.x * cosCutoff;
    float sinTheta = safe_sqrt(1.0f - cosTheta * cosTheta);
    float sp, cp; sincosf(2.0f * kPi * s.y, &sp, &cp);
    return V3(cp * sinTheta, sp * sinTheta, cosTheta);
}
inline float squareToUniformConePdf(float cosCutoff) { return kInvTwoPi / (1 - cosCutoff); }
inline V2 squareToUniformTriangle(V2 s) {
    float a = safe_sqrt(1.0f - s.x);
    return V2{1 - a, a * s.y};
}
inline float intervalToTent(float s) {
    float sign;
    if (s < 0.5f) { sign = 1; s *= 2; } else { sign = -1; s = 2 * (s - 0.5f); }
    return sign * (1 - std::sqrt(s));
}
inline V2 squareToTent(V2 s) { return V2{intervalToTent(s.x), intervalToTent(s.y)}; }

/* 4x4 row-major matrix helpers (Transform, include/mitsuba/core/transform.h) */
struct M44 {
    float m[4][4];
    V3 point(V3 p) const { /* Transform::operator()(Point): with homogeneous divide */
        float x = m[0][0] * p.x + m[0][1] * p.y + m[0][2] * p.z + m[0][3];
        float y = m[1][0] * p.x + m[1][1] * p.y + m[1][2] * p.z + m[1][3];
        float z = m[2][0] * p.x + m[2][1] * p.y + m[2][2] * p.z + m[2][3];
        float w = m[3][0] * p.x + m[3][1] * p.y + m[3][2] * p.z + m[3][3];
        if (w == 1.0f) return V3(x, y, z);
        return V3(x, y, z) / w;
    }
    V3 pointAffine(V3 p) const {
        return V3(m[0][0] * p.x + m[0][1] * p.y + m[0][2] * p.z + m[0][3],
                  m[1][0] * p.x + m[1][1] * p.y + m[1][2] * p.z + m[1][3],
                  m[2][0] * p.x + m[2][1] * p.y + m[2][2] * p.z + m[2][3]);
    }
    V3 vec(V3 v) const {
        return V3(m[0][0] * v.x + m[0][1] * v.y + m[0][2] * v.z,
                  m[1][0] * v.x + m[1][1] * v.y + m[1][2] * v.z,
                  m[2][0] * v.x + m[2][1] * v.y + m[2][2] * v.z);
    }
};
inline M44 m44_from(const float *p) { M44 r; std::memcpy(r.m, p, 64); return r; }
inline M44 m44_inverse(const M44 &a) {
    double w[4][8];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { w[i][j] = a.m[i][j]; w[i][j + 4] = (i == j); }
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        for (int r = c + 1; r < 4; ++r) if (std::fabs(w[r][c]) > std::fabs(w[piv][c])) piv = r;
        for (int j = 0; j < 8; ++j) std::swap(w[c][j], w[piv][j]);
        double d = w[c][c];
        for (int j = 0; j < 8; ++j) w[c][j] /= d;
        for (int r = 0; r < 4; ++r) if (r != c) {
            double f = w[r][c];
            for (int j = 0; j < 8; ++j) w[r][j] -= f * w[c][j];
        }
    }
    M44 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = (float)w[i][j + 4];
    return r;
}

/* round-trip through IEEE half like TMIPMap<Spectrum, half3> storage (mipmap.h) */
inline float roundToHalf(float v) { return (float)(_Float16)v; }

} // namespace orc
