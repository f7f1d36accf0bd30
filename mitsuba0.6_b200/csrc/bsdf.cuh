// bsdf.cuh -- material programs: bitmap/constant textures, microfacet distributions, the rough
// transmittance LUT and the nine in-scope BSDF models as a flattened node tree.
//
// Design: polymorphic BSDF objects of the reference become an array of MtsBsdfNode; the three
// wrappers (coating, blendbsdf, bumpmap) recurse through a depth-bounded template so that the
// call graph is finite (no device-side recursion/stack), the six leaf models are evaluated by
// one switch.  typeMask = EAll and component = -1 throughout, which is how MIPathTracer::Li
// queries BSDFs (src/integrators/path/path.cpp:178-207).
#pragma once
#include "scene.cuh"

namespace mts {

#ifndef BSDF_MAX_DEPTH
#define BSDF_MAX_DEPTH 4
#endif
#ifndef BSDF_NESTED_ATTR
#define BSDF_NESTED_ATTR __noinline__
#endif
constexpr int kMaxBsdfDepth = BSDF_MAX_DEPTH;  // wrapper nesting levels supported above a leaf
#define MTS_MIPMAP_LUT_SIZE 64   // include/mitsuba/render/mipmap.h:37

// ---------------------------------------------------------------------------------------
// Textures: TMIPMap (include/mitsuba/render/mipmap.h:503-834)
// ---------------------------------------------------------------------------------------
// Where a texture is looked up: surface uv and, for camera-ray hits of BSDFs that use ray differentials, the uv partials
// of Intersection::computePartials (dudx, dvdx, dudy, dvdy).  `filtered` = Intersection::hasUVPartials.
struct TexCtx {
    float2 uv;
    float4 duv;
    bool filtered;
};
DEVFN TexCtx texCtx(float2 uv) { TexCtx t; t.uv = uv; t.duv = make_float4(0, 0, 0, 0); t.filtered = false; return t; }

// evalTexel coordinate handling (mipmap.h:503-563): returns the wrapped coordinate, or -1 / -2 when the lookup
// yields the constant 0 / 1 (EZero / EOne boundary conditions)
DEVOL int wrapCoordOutside(int x, int size, int mode) {
    switch (mode) {
        case MTS_WRAP_REPEAT: return wrapRepeat(x, size);
        case MTS_WRAP_CLAMP: return clampi(x, 0, size - 1);
        case MTS_WRAP_MIRROR: { int m = wrapRepeat(x, 2 * size); return m >= size ? 2 * size - m - 1 : m; }
        case MTS_WRAP_ZERO: return -1;
        default: return -2;
    }
}
DEVFN int wrapCoord(int x, int size, int mode) {   // the in-range test stays inline, the boundary conditions are one shared copy
    if (x >= 0 && x < size) return x;
    return wrapCoordOutside(x, size, mode);
}
DEVFN float3 texelAt(const __half *base, int channels, size_t i) {
    if (channels == 1) return f3(__half2float(__ldg(base + i)));
    const uint2 raw = __ldg(reinterpret_cast<const uint2 *>(base) + i);
    const float2 fa = __half22float2(*reinterpret_cast<const __half2 *>(&raw.x)), fb = __half22float2(*reinterpret_cast<const __half2 *>(&raw.y));
    return f3(fa.x, fa.y, fb.x);
}
// The four texels of a bilinear footprint on one pyramid level.  Coordinates are wrapped once per row / column and
// all four loads are issued back to back (one round trip to L2/HBM instead of four dependent ones).
struct Quad { float3 p00, p10, p01, p11; };
// EZero / EOne boundary conditions (rare): a coordinate outside the image yields the constant (mipmap.h:503-563)
DEVOL void quadBoundaryConstants(Quad &q, int xa, int xb, int ya, int yb) {
    const int bx0 = xa < 0 ? xa : 0, bx1 = xb < 0 ? xb : 0, by0 = ya < 0 ? ya : 0, by1 = yb < 0 ? yb : 0;
    auto konst = [](int cx, int cy) { return f3((cx == -2 || (cx == 0 && cy == -2)) ? 1.0f : 0.0f); };  // u is tested first
    if (bx0 | by0) q.p00 = konst(bx0, by0);
    if (bx1 | by0) q.p10 = konst(bx1, by0);
    if (bx0 | by1) q.p01 = konst(bx0, by1);
    if (bx1 | by1) q.p11 = konst(bx1, by1);
}
DEVFN Quad texQuad(const DTexture &t, const __half *base, int w, int h, int x0, int y0) {
    const int xa = wrapCoord(x0, w, t.wrapU), xb = wrapCoord(x0 + 1, w, t.wrapU);
    const int ya = wrapCoord(y0, h, t.wrapV), yb = wrapCoord(y0 + 1, h, t.wrapV);
    const int cxa = max(xa, 0), cxb = max(xb, 0), cya = max(ya, 0), cyb = max(yb, 0);
    const size_t i00 = (size_t)cya * w + cxa, i10 = (size_t)cya * w + cxb, i01 = (size_t)cyb * w + cxa, i11 = (size_t)cyb * w + cxb;
    Quad q;
    if (t.channels == 1) {
        const __half h00 = __ldg(base + i00), h10 = __ldg(base + i10), h01 = __ldg(base + i01), h11 = __ldg(base + i11);
        q.p00 = f3(__half2float(h00)); q.p10 = f3(__half2float(h10)); q.p01 = f3(__half2float(h01)); q.p11 = f3(__half2float(h11));
    } else {
        const uint2 *tx = reinterpret_cast<const uint2 *>(base);
        const uint2 r00 = __ldg(tx + i00), r10 = __ldg(tx + i10), r01 = __ldg(tx + i01), r11 = __ldg(tx + i11);
        auto cvt = [](uint2 raw) {
            float2 fa = __half22float2(*reinterpret_cast<const __half2 *>(&raw.x)), fb = __half22float2(*reinterpret_cast<const __half2 *>(&raw.y));
            return f3(fa.x, fa.y, fb.x);
        };
        q.p00 = cvt(r00); q.p10 = cvt(r10); q.p01 = cvt(r01); q.p11 = cvt(r11);
    }
    if ((xa | xb | ya | yb) < 0) quadBoundaryConstants(q, xa, xb, ya, yb);
    return q;
}
// evalTexel(level, x, y), mipmap.h:503-563
DEVFN float3 texel(const DTexture &t, const __half *base, int w, int h, int x, int y) {
    const int xa = wrapCoord(x, w, t.wrapU);
    if (xa < 0) return f3(xa == -2 ? 1.0f : 0.0f);
    const int ya = wrapCoord(y, h, t.wrapV);
    if (ya < 0) return f3(ya == -2 ? 1.0f : 0.0f);
    return texelAt(base, t.channels, (size_t)ya * w + xa);
}
DEVFN const __half *mipBase(const DTexture &t, int level) { return t.texels + (size_t)t.levelOffset[level] * (t.channels == 1 ? 1 : 4); }
// evalBox(level, uv), mipmap.h:566-569
DEVFN float3 mipBox(const DTexture &t, int level, float2 uv) {
    const int w = t.levelW[level], h = t.levelH[level];
    return texel(t, mipBase(t, level), w, h, floorToInt(uv.x * w), floorToInt(uv.y * h));
}
// evalBilinear(level, uv), mipmap.h:575-596
DEVFN float3 mipBilinear(const DTexture &t, int level, float2 uv) {
    if (!isfinite(uv.x) || !isfinite(uv.y)) return f3(0.0f);
    if (level >= t.levels) return mipBox(t, t.levels - 1, uv);
    const int w = t.levelW[level], h = t.levelH[level];
    float u = uv.x * w - 0.5f, v = uv.y * h - 0.5f;
    int xPos = floorToInt(u), yPos = floorToInt(v);
    float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
    const Quad q = texQuad(t, mipBase(t, level), w, h, xPos, yPos);
    return q.p00 * dx2 * dy2 + q.p01 * dx2 * dy1 + q.p10 * dx1 * dy2 + q.p11 * dx1 * dy1;
}
// evalBilinear(0, uv) / evalBox(0, uv): the unfiltered lookup of BitmapTexture::eval(uv) (bitmap.cpp:431-454)
DEVFN float3 texLookup(const DTexture &t, float2 uv) {
    if (t.filter == MTS_FILTER_NEAREST)
        return texel(t, t.texels, t.width, t.height, floorToInt(uv.x * t.width), floorToInt(uv.y * t.height));
    if (!isfinite(uv.x) || !isfinite(uv.y)) return f3(0.0f);
    float u = uv.x * t.width - 0.5f, v = uv.y * t.height - 0.5f;
    int xPos = floorToInt(u), yPos = floorToInt(v);
    float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
    const Quad q = texQuad(t, t.texels, t.width, t.height, xPos, yPos);
    return q.p00 * dx2 * dy2 + q.p01 * dx2 * dy1 + q.p10 * dx1 * dy2 + q.p11 * dx1 * dy1;
}
// evalGradientBilinear(0, uv), mipmap.h:601-626
DEVFN void texGradient(const DTexture &t, float2 uv, float3 &g0, float3 &g1) {
    if (t.filter == MTS_FILTER_NEAREST || !isfinite(uv.x) || !isfinite(uv.y)) { g0 = g1 = f3(0.0f); return; }
    float u = uv.x * t.width - 0.5f, v = uv.y * t.height - 0.5f;
    int xPos = floorToInt(u), yPos = floorToInt(v);
    float dx = u - xPos, dy = v - yPos;
    const Quad q = texQuad(t, t.texels, t.width, t.height, xPos, yPos);
    float3 tmp = q.p01 + q.p10 - q.p11;
    g0 = (q.p10 + q.p00 * (dy - 1) - tmp * dy) * (float)t.width;
    g1 = (q.p01 + q.p00 * (dx - 1) - tmp * dx) * (float)t.height;
}
// ---- evalEWA(level, uv, A, B, C), mipmap.h:764-834 -------------------------------------------------------------------
// The reference walks the bounding box of the ellipse row by row and weights every texel whose running quadratic q lies
// below the table size.  The walk here visits the same texels in the same order with the same separately rounded
// updates of q (it indexes the weight table, so no FMA contraction), but
//   * with the row's wrapped v coordinate and base pointer hoisted out of the column walk,
//   * and optionally for TWO textures that share their geometry (size, uv transform, wrap modes: the albedo / specular
//     pair of a material), which then share the footprint, the weights and the walk.
DEVFN bool sameTextureGeometry(const DTexture &a, const DTexture &b) {
    return a.width == b.width && a.height == b.height && a.levels == b.levels && a.wrapU == b.wrapU && a.wrapV == b.wrapV &&
           a.filter == b.filter && a.maxAnisotropy == b.maxAnisotropy && a.uvScale.x == b.uvScale.x && a.uvScale.y == b.uvScale.y &&
           a.uvOffset.x == b.uvOffset.x && a.uvOffset.y == b.uvOffset.y;
}
// wrapCoord with the common cases inline: inside the image, or a repeating power-of-two image
DEVFN int wrapCoordFast(int x, int size, int mode) {
    if ((unsigned)x < (unsigned)size) return x;
    if (mode == MTS_WRAP_REPEAT && (size & (size - 1)) == 0) return x & (size - 1);
    return wrapCoordOutside(x, size, mode);
}
// evalEWA on ONE level in three pieces -- the bounding box of the ellipse, one row of the box, the final division -- so
// that the rows of a walk can be shared out among the lanes of a group (k_shade_ewa) or walked by one thread (ewaWalk).
struct EwaLevel {
    int kind;            // EWA_WALK, or the value is known without a walk: EWA_ZERO (non-finite input), EWA_BOX (beyond the pyramid), EWA_BILINEAR (empty box)
    int w, h, u0, u1, v0, v1;
    float u, v, As, Bs, Cs, ddq, uu0;
};
enum { EWA_WALK = 0, EWA_ZERO = 1, EWA_BOX = 2, EWA_BILINEAR = 3 };
DEVFN EwaLevel ewaSetup(const DTexture &ta, int level, float2 uv, float A0, float B0, float C0) {
    EwaLevel L;
    L.kind = EWA_WALK; L.w = L.h = 1; L.u0 = L.v0 = 0; L.u1 = L.v1 = -1; L.u = L.v = L.As = L.Bs = L.Cs = L.ddq = L.uu0 = 0;
    if (!isfinite(A0 + B0 + C0 + uv.x + uv.y)) { L.kind = EWA_ZERO; return L; }
    if (level >= ta.levels) { L.kind = EWA_BOX; return L; }
    L.w = ta.levelW[level]; L.h = ta.levelH[level];
    L.u = uv.x * L.w - 0.5f; L.v = uv.y * L.h - 0.5f;
    const float ratioX = (float)L.w / (float)ta.width, ratioY = (float)L.h / (float)ta.height;   // m_sizeRatio
    const float A = A0 / (ratioX * ratioX), B = B0 / (ratioX * ratioY), C = C0 / (ratioY * ratioY);
    const float invDet = 1.0f / (-B * B + 4.0f * A * C), deltaU = 2.0f * sqrtf(C * invDet), deltaV = 2.0f * sqrtf(A * invDet);
    L.u0 = (int)ceilf(L.u - deltaU); L.u1 = floorToInt(L.u + deltaU); L.v0 = (int)ceilf(L.v - deltaV); L.v1 = floorToInt(L.v + deltaV);
    L.As = A * MTS_MIPMAP_LUT_SIZE; L.Bs = B * MTS_MIPMAP_LUT_SIZE; L.Cs = C * MTS_MIPMAP_LUT_SIZE;
    L.ddq = 2 * L.As; L.uu0 = (float)L.u0 - L.u;
    if (L.u1 < L.u0 || L.v1 < L.v0) L.kind = EWA_BILINEAR;   // no texel: denominator 0 -> evalBilinear (mipmap.h:829-830)
    return L;
}
DEVFN int ewaRows(const EwaLevel &L) { return L.kind == EWA_WALK ? L.v1 - L.v0 + 1 : 0; }
template <bool PAIR>
DEVFN void ewaRow(const DTexture &ta, const DTexture *tb, int level, const EwaLevel &L, int vt, float3 &sumA, float3 &sumB, float &den) {
    const bool monoA = ta.channels == 1, monoB = PAIR ? tb->channels == 1 : false;
    const float vv = (float)vt - L.v;
    float q = __fadd_rn(__fmul_rn(__fmul_rn(L.As, L.uu0), L.uu0), __fmul_rn(__fadd_rn(__fmul_rn(L.Bs, L.uu0), __fmul_rn(L.Cs, vv)), vv));
    float dq = __fadd_rn(__fmul_rn(L.As, __fadd_rn(__fmul_rn(2.0f, L.uu0), 1.0f)), __fmul_rn(L.Bs, vv));
    const int ya = wrapCoordFast(vt, L.h, ta.wrapV);       // hoisted out of the column walk
    const size_t off = (size_t)max(ya, 0) * L.w;
    const __half *rowA = mipBase(ta, level) + off * (monoA ? 1 : 4), *rowB = PAIR ? mipBase(*tb, level) + off * (monoB ? 1 : 4) : nullptr;
    for (int ut = L.u0; ut <= L.u1; ++ut) {
        if (q < (float)MTS_MIPMAP_LUT_SIZE) {
            const uint32_t qi = (uint32_t)q;
            if (qi < MTS_MIPMAP_LUT_SIZE) {
                const float weight = __ldg(ta.weightLut + (int)q);
                const int xa = wrapCoordFast(ut, L.w, ta.wrapU);
                float3 texA, texB = f3(0.0f);
                if (xa < 0) texA = texB = f3(xa == -2 ? 1.0f : 0.0f);          // EZero / EOne: u is tested first (evalTexel, mipmap.h:503-563)
                else if (ya < 0) texA = texB = f3(ya == -2 ? 1.0f : 0.0f);
                else { texA = texelAt(rowA, monoA ? 1 : 3, (size_t)xa); if (PAIR) texB = texelAt(rowB, monoB ? 1 : 3, (size_t)xa); }
                sumA += texA * weight;
                if (PAIR) sumB += texB * weight;
                den += weight;
            }
        }
        q = __fadd_rn(q, dq);
        dq = __fadd_rn(dq, L.ddq);
    }
}
template <bool PAIR>
DEVFN void ewaFinish(const DTexture &ta, const DTexture *tb, int level, float2 uv, const EwaLevel &L, float3 sumA, float3 sumB, float den,
                     float3 &outA, float3 &outB) {
    outB = f3(0.0f);
    if (L.kind == EWA_ZERO) { outA = f3(0.0f); return; }
    if (L.kind == EWA_BOX) { outA = mipBox(ta, ta.levels - 1, uv); if (PAIR) outB = mipBox(*tb, tb->levels - 1, uv); return; }
    if (L.kind == EWA_BILINEAR || den == 0) { outA = mipBilinear(ta, level, uv); if (PAIR) outB = mipBilinear(*tb, level, uv); return; }   // mipmap.h:829-830
    outA = sumA / den;
    if (PAIR) outB = sumB / den;
}
// one thread walks the whole level
template <bool PAIR>
DEVFN void ewaWalk(const DTexture &ta, const DTexture *tb, int level, float2 uv, float A0, float B0, float C0, float3 &outA, float3 &outB) {
    const EwaLevel L = ewaSetup(ta, level, uv, A0, B0, C0);
    float3 sumA = f3(0.0f), sumB = f3(0.0f);
    float den = 0.0f;
    if (L.kind == EWA_WALK) for (int vt = L.v0; vt <= L.v1; ++vt) ewaRow<PAIR>(ta, tb, level, L, vt, sumA, sumB, den);
    ewaFinish<PAIR>(ta, tb, level, uv, L, sumA, sumB, den, outA, outB);
}
// kept out of line: one copy of each walk per kernel.  evalEWA(ilevel) * (1 - a) + evalEWA(ilevel + 1) * a, mipmap.h:712-713.
// (Measured alternatives, both slower on B200 -- profiles/r2h, r2j: the rows of a walk dealt out to groups of 4 / 8 / 16 lanes,
// and one flat loop over the accepted texels of both levels; the plain nested loops below compile to the tightest code.)
static __device__ __noinline__ float3 mipEWA2(const DTexture &t, int ilevel, float2 uv, float A, float B, float C, float a) {
    float3 l0, l1, unused;
    ewaWalk<false>(t, nullptr, ilevel, uv, A, B, C, l0, unused);
    ewaWalk<false>(t, nullptr, ilevel + 1, uv, A, B, C, l1, unused);
    return l0 * (1.0f - a) + l1 * a;
}
static __device__ __noinline__ void mipEWA2Pair(const DTexture &ta, const DTexture &tb, int ilevel, float2 uv, float A, float B, float C, float a,
                                                float3 &ra, float3 &rb) {
    float3 a0, a1, b0, b1;
    ewaWalk<true>(ta, &tb, ilevel, uv, A, B, C, a0, b0);
    ewaWalk<true>(ta, &tb, ilevel + 1, uv, A, B, C, a1, b1);
    ra = a0 * (1.0f - a) + a1 * a;
    rb = b0 * (1.0f - a) + b1 * a;
}
// math::hypot2 (include/mitsuba/core/math.h) with separately rounded operations
DEVFN float hypot2Exact(float a, float b) {
    if (fabsf(a) > fabsf(b)) { const float r = __fdiv_rn(b, a); return __fmul_rn(fabsf(a), sqrtf(__fadd_rn(1.0f, __fmul_rn(r, r)))); }
    if (b != 0.0f) { const float r = __fdiv_rn(a, b); return __fmul_rn(fabsf(b), sqrtf(__fadd_rn(1.0f, __fmul_rn(r, r)))); }
    return 0.0f;
}
// TMIPMap::eval(uv, d0, d1), mipmap.h:629-715: level selection, trilinear interpolation, anisotropy clamp.  What is left
// to do after the footprint analysis: one bilinear lookup, two (trilinear), or the EWA walk of two levels.
struct MipFoot { int kind, ilevel; float a, A, B, C; };
enum { FOOT_BOX0 = 0, FOOT_BILINEAR = 1, FOOT_TRILINEAR = 2, FOOT_EWA = 3 };
DEVFN MipFoot mipFootprint(const DTexture &t, float2 d0, float2 d1) {
    MipFoot f; f.kind = FOOT_BILINEAR; f.ilevel = 0; f.a = 0; f.A = f.B = f.C = 0;
    if (t.filter == MTS_FILTER_NEAREST) { f.kind = FOOT_BOX0; return f; }
    if (t.filter == MTS_FILTER_BILINEAR) return f;
    // Degenerate footprints (one partial zero, F = AC - B^2/4 = 0 up to rounding) decide the branch below by the SIGN of
    // rounding noise: every product and sum is rounded separately, as the host code does (no FMA contraction), so that
    // the same branch is taken as in the reference.
    const float du0 = d0.x * t.width, dv0 = d0.y * t.height, du1 = d1.x * t.width, dv1 = d1.y * t.height;
    float A = __fadd_rn(__fmul_rn(dv0, dv0), __fmul_rn(dv1, dv1)), B = __fmul_rn(-2.0f, __fadd_rn(__fmul_rn(du0, dv0), __fmul_rn(du1, dv1))),
          C = __fadd_rn(__fmul_rn(du0, du0), __fmul_rn(du1, du1)), F = __fsub_rn(__fmul_rn(A, C), __fmul_rn(__fmul_rn(B, B), 0.25f));
    const float root = hypot2Exact(__fsub_rn(A, C), B), Aprime = __fmul_rn(0.5f, __fsub_rn(__fadd_rn(A, C), root)),
                Cprime = __fmul_rn(0.5f, __fadd_rn(__fadd_rn(A, C), root));
    float majorRadius = Aprime != 0 ? sqrtf(__fdiv_rn(F, Aprime)) : 0, minorRadius = Cprime != 0 ? sqrtf(__fdiv_rn(F, Cprime)) : 0;
    const float invLn2 = 1.4426950408889634f;   // math::log2 = fastlog * (1 / log 2), src/libcore/math.cpp:103-106
    if (t.filter == MTS_FILTER_TRILINEAR || !(minorRadius > 0) || !(majorRadius > 0) || F < 0) {
        const float level = fastlog(fmaxf(majorRadius, kEps)) * invLn2;
        const int ilevel = floorToInt(level);
        if (ilevel < 0) return f;
        f.kind = FOOT_TRILINEAR; f.ilevel = ilevel; f.a = level - ilevel;
        return f;
    }
    if (minorRadius * t.maxAnisotropy < majorRadius) {
        minorRadius = majorRadius / t.maxAnisotropy;
        float sinTheta, cosTheta_;
        sincosf(0.5f * atanf(B / (A - C)), &sinTheta, &cosTheta_);
        const float a2 = majorRadius * majorRadius, b2 = minorRadius * minorRadius, sinTheta2 = sinTheta * sinTheta,
                    cosTheta2_ = cosTheta_ * cosTheta_, sin2Theta = 2 * sinTheta * cosTheta_;
        A = a2 * cosTheta2_ + b2 * sinTheta2;
        B = (a2 - b2) * sin2Theta;
        C = a2 * sinTheta2 + b2 * cosTheta2_;
        F = a2 * b2;
    }
    const float scale = 1.0f / F;
    A *= scale; B *= scale; C *= scale;
    const float level = fmaxf(0.0f, fastlog(minorRadius) * invLn2);
    f.ilevel = (int)level;
    f.a = level - f.ilevel;
    if (majorRadius < 1 || !(A > 0 && C > 0)) return f;   // evalBilinear(ilevel, uv)
    f.kind = FOOT_EWA; f.A = A; f.B = B; f.C = C;
    return f;
}
DEVFN float3 mipEvalFoot(const DTexture &t, float2 uv, const MipFoot &f) {
    switch (f.kind) {
    case FOOT_BOX0: return mipBox(t, 0, uv);
    case FOOT_BILINEAR: return mipBilinear(t, f.ilevel, uv);
    case FOOT_TRILINEAR: return mipBilinear(t, f.ilevel, uv) * (1.0f - f.a) + mipBilinear(t, f.ilevel + 1, uv) * f.a;
    default: return mipEWA2(t, f.ilevel, uv, f.A, f.B, f.C, f.a);
    }
}
static __device__ __noinline__ float3 mipEval(const DTexture &t, float2 uv, float2 d0, float2 d1) {
    return mipEvalFoot(t, uv, mipFootprint(t, d0, d1));
}
// Texture2D::eval(its, filter = true) (src/librender/texture.cpp:104-114) + ScalingTexture (src/textures/scale.cpp:86-88)
DEVOL float3 evalTex(const Mats S, const MtsTexRef &r, const TexCtx &tc) {
    float3 s = f3(r.scale[0], r.scale[1], r.scale[2]);
    if (r.tex < 0) return s;
    const DTexture &t = S.textures[r.tex];
    const float2 uv = make_float2(tc.uv.x * t.uvScale.x + t.uvOffset.x, tc.uv.y * t.uvScale.y + t.uvOffset.y);
    if (tc.filtered && t.filter >= MTS_FILTER_TRILINEAR)   // bitmap.cpp:486-499; nearest / bilinear ignore the partials (mipmap.h:630-633)
        return mipEval(t, uv, make_float2(tc.duv.x * t.uvScale.x, tc.duv.y * t.uvScale.y), make_float2(tc.duv.z * t.uvScale.x, tc.duv.w * t.uvScale.y)) * s;
    return texLookup(t, uv) * s;
}
// A filtered lookup whose EWA walk is left to a kernel of its own (k_shade_ewa of the staged shade): only a fraction of
// the lanes of a warp need a walk at all, and those that do wait on dependent texel loads -- as one job per thread pair the
// walks run with full warps at high occupancy.  slotA: which of the two texture slots texA fills (texB always fills slot 1).
struct EwaDefer { int texA, texB, slotA, ilevel; float a, A, B, C; float2 uv; bool pending; };
DEVFN void deferEwa(EwaDefer &d, int texA, int texB, int slotA, const MipFoot &foot, float2 uv) {
    d.pending = true; d.texA = texA; d.texB = texB; d.slotA = slotA; d.ilevel = foot.ilevel;
    d.a = foot.a; d.A = foot.A; d.B = foot.B; d.C = foot.C; d.uv = uv;
}
// Two texture slots of one material at one hit.  Bitmaps that share their geometry (the usual albedo / specular pair) and
// are filtered with EWA share one footprint analysis and one walk over the texels.  With `defer` the walk is not done here.
DEVOL void evalTexPair(const Mats S, const MtsTexRef &ra, const MtsTexRef &rb, const TexCtx &tc, float3 &ca, float3 &cb, EwaDefer *defer) {
    if (tc.filtered) {
        const bool fa = ra.tex >= 0 && S.textures[max(ra.tex, 0)].filter == MTS_FILTER_EWA;
        const bool fb = rb.tex >= 0 && S.textures[max(rb.tex, 0)].filter == MTS_FILTER_EWA;
        if (fa || fb) {
            const DTexture &ta = S.textures[fa ? ra.tex : rb.tex];
            const float2 uv = make_float2(tc.uv.x * ta.uvScale.x + ta.uvOffset.x, tc.uv.y * ta.uvScale.y + ta.uvOffset.y);
            const MipFoot foot = mipFootprint(ta, make_float2(tc.duv.x * ta.uvScale.x, tc.duv.y * ta.uvScale.y),
                                              make_float2(tc.duv.z * ta.uvScale.x, tc.duv.w * ta.uvScale.y));
            if (fa && fb && sameTextureGeometry(ta, S.textures[rb.tex])) {
                const DTexture &tb = S.textures[rb.tex];
                float3 va = f3(0.0f), vb = f3(0.0f);
                if (foot.kind == FOOT_EWA) {
                    if (defer) { deferEwa(*defer, ra.tex, rb.tex, 0, foot, uv); ca = cb = f3(0.0f); return; }
                    mipEWA2Pair(ta, tb, foot.ilevel, uv, foot.A, foot.B, foot.C, foot.a, va, vb);
                } else { va = mipEvalFoot(ta, uv, foot); vb = mipEvalFoot(tb, uv, foot); }
                ca = va * f3(ra.scale[0], ra.scale[1], ra.scale[2]);
                cb = vb * f3(rb.scale[0], rb.scale[1], rb.scale[2]);
                return;
            }
            if (defer && foot.kind == FOOT_EWA) {   // one walk can be deferred: that of the first EWA texture
                deferEwa(*defer, fa ? ra.tex : rb.tex, -1, fa ? 0 : 1, foot, uv);
                if (fa) { ca = f3(0.0f); cb = evalTex(S, rb, tc); } else { cb = f3(0.0f); ca = evalTex(S, ra, tc); }
                return;
            }
        }
    }
    ca = evalTex(S, ra, tc);
    cb = evalTex(S, rb, tc);
}
// Texture2D::evalGradient (texture.cpp:116-124) + ScalingTexture::evalGradient (scale.cpp:90-94)
DEVFN void evalTexGradient(const Mats S, const MtsTexRef &r, float2 uv, float3 &g0, float3 &g1) {
    if (r.tex < 0) { g0 = g1 = f3(0.0f); return; }
    const DTexture &t = S.textures[r.tex];
    texGradient(t, make_float2(uv.x * t.uvScale.x + t.uvOffset.x, uv.y * t.uvScale.y + t.uvOffset.y), g0, g1);
    float3 s = f3(r.scale[0], r.scale[1], r.scale[2]);
    g0 = g0 * t.uvScale.x * s;
    g1 = g1 * t.uvScale.y * s;
}

// ---------------------------------------------------------------------------------------
// MicrofacetDistribution (src/bsdfs/microfacet.h), Beckmann + GGX
// ---------------------------------------------------------------------------------------
struct Microfacet {
    bool ggx, sampleVisible;
    float alphaU, alphaV;
    DEVFN Microfacet(bool g, float au, float av, bool sv) : ggx(g), sampleVisible(sv) {
        alphaU = fmaxf(au, 1e-4f); alphaV = fmaxf(av, 1e-4f);  // :70-71
    }
    DEVFN bool isIsotropic() const { return alphaU == alphaV; }
    DEVFN void scaleAlpha(float v) { alphaU *= v; alphaV *= v; }   // :96-99
    // :191-234
    DEVOLM float eval(float3 m) const {
        if (cosTheta(m) <= 0) return 0.0f;
        float ct2 = cosTheta2(m);
        float be = ((m.x * m.x) / (alphaU * alphaU) + (m.y * m.y) / (alphaV * alphaV)) / ct2;
        float result;
        if (!ggx) result = fastexp(-be) / (kPi * alphaU * alphaV * ct2 * ct2);
        else { float root = (1.0f + be) * ct2; result = 1.0f / (kPi * alphaU * alphaV * root * root); }
        if (result * cosTheta(m) < 1e-20f) result = 0;
        return result;
    }
    // :541-551
    DEVFN float projectRoughness(float3 v) const {
        float invSinTheta2 = 1 / sinTheta2(v);
        if (isIsotropic() || invSinTheta2 <= 0) return alphaU;
        float cosPhi2 = v.x * v.x * invSinTheta2, sinPhi2 = v.y * v.y * invSinTheta2;
        return sqrtf(cosPhi2 * alphaU * alphaU + sinPhi2 * alphaV * alphaV);
    }
    // :477-514
    DEVOLM float smithG1(float3 v, float3 m) const {
        if (dot(v, m) * cosTheta(v) <= 0) return 0.0f;
        float tt = fabsf(tanTheta(v));
        if (tt == 0.0f) return 1.0f;
        float alpha = projectRoughness(v);
        if (!ggx) {
            float a = 1.0f / (alpha * tt);
            if (a >= 1.6f) return 1.0f;
            float aSqr = a * a;
            return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
        }
        float root = alpha * tt;
        return 2.0f / (1.0f + hypot2(1.0f, root));
    }
    DEVFN float G(float3 wi, float3 wo, float3 m) const { return smithG1(wi, m) * smithG1(wo, m); }
    DEVFN float pdfAll(float3 m) const { return eval(m) * cosTheta(m); }  // :404-407
    DEVFN float pdfVisible(float3 wi, float3 m) const {                    // :462-466
        if (cosTheta(wi) == 0) return 0.0f;
        return smithG1(wi, m) * absDot(wi, m) * eval(m) / fabsf(cosTheta(wi));
    }
    DEVFN float pdf(float3 wi, float3 m) const { return sampleVisible ? pdfVisible(wi, m) : pdfAll(m); }
    // :287-395
    DEVOLM float3 sampleAll(float2 sample, float &pdf) const {
        float cosThetaM, sinPhiM, cosPhiM, alphaSqr;
        if (isIsotropic()) {
            sincosf((2.0f * kPi) * sample.y, &sinPhiM, &cosPhiM);
            alphaSqr = alphaU * alphaU;
        } else {
            float phiM = atanf(alphaV / alphaU * tanf(kPi + 2 * kPi * sample.y)) + kPi * floorf(2 * sample.y + 0.5f);
            sincosf(phiM, &sinPhiM, &cosPhiM);
            float cosSc = cosPhiM / alphaU, sinSc = sinPhiM / alphaV;
            alphaSqr = 1.0f / (cosSc * cosSc + sinSc * sinSc);
        }
        if (!ggx) {
            float tanThetaMSqr = alphaSqr * -fastlog(1.0f - sample.x);
            cosThetaM = 1.0f / sqrtf(1.0f + tanThetaMSqr);
            pdf = (1.0f - sample.x) / (kPi * alphaU * alphaV * cosThetaM * cosThetaM * cosThetaM);
        } else {
            float tanThetaMSqr = alphaSqr * sample.x / (1.0f - sample.x);
            cosThetaM = 1.0f / sqrtf(1.0f + tanThetaMSqr);
            float temp = 1 + tanThetaMSqr / alphaSqr;
            pdf = kInvPi / (alphaU * alphaV * cosThetaM * cosThetaM * cosThetaM * temp * temp);
        }
        if (pdf < 1e-20f) pdf = 0;
        float sinThetaM = sqrtf(fmaxf(0.0f, 1 - cosThetaM * cosThetaM));
        return f3(sinThetaM * cosPhiM, sinThetaM * sinPhiM, cosThetaM);
    }
    // :573-697
    DEVFN float2 sampleVisible11(float thetaI, float2 sample) const {
        const float SQRT_PI_INV = 0.56418958354775628f;
        float2 slope;
        if (!ggx) {
            if (thetaI < 1e-4f) {
                float sp, cp;
                float r = sqrtf(-fastlog(1.0f - sample.x));
                sincosf(2 * kPi * sample.y, &sp, &cp);
                return make_float2(r * cp, r * sp);
            }
            float tanThetaI = tanf(thetaI), cotThetaI = 1 / tanThetaI;
            float a = -1, c = erf_as(cotThetaI);
            float sample_x = fmaxf(sample.x, 1e-6f);
            float fit = 1 + thetaI * (-0.876f + thetaI * (0.4265f - 0.0594f * thetaI));
            float b = c - (1 + c) * powf(1 - sample_x, fit);
            float normalization = 1 / (1 + c + SQRT_PI_INV * tanThetaI * expf(-cotThetaI * cotThetaI));
            int it = 0;
            while (++it < 10) {
                if (!(b >= a && b <= c)) b = 0.5f * (a + c);
                float invErf = erfinv_giles(b);
                float value = normalization * (1 + b + SQRT_PI_INV * tanThetaI * expf(-invErf * invErf)) - sample_x;
                float derivative = normalization * (1 - invErf * tanThetaI);
                if (fabsf(value) < 1e-5f) break;
                if (value > 0) c = b; else a = b;
                b -= value / derivative;
            }
            slope.x = erfinv_giles(b);
            slope.y = erfinv_giles(2.0f * fmaxf(sample.y, 1e-6f) - 1.0f);
        } else {
            if (thetaI < 1e-4f) {
                float sp, cp;
                float r = safe_sqrt(sample.x / (1 - sample.x));
                sincosf(2 * kPi * sample.y, &sp, &cp);
                return make_float2(r * cp, r * sp);
            }
            float tanThetaI = tanf(thetaI);
            float a = 1 / tanThetaI;
            float G1 = 2.0f / (1.0f + safe_sqrt(1.0f + 1.0f / (a * a)));
            float A = 2.0f * sample.x / G1 - 1.0f;
            if (fabsf(A) == 1) A -= signum(A) * kEps;
            float tmp = 1.0f / (A * A - 1.0f);
            float B = tanThetaI;
            float D = safe_sqrt(B * B * tmp * tmp - (A * A - B * B) * tmp);
            float slope_x_1 = B * tmp - D, slope_x_2 = B * tmp + D;
            slope.x = (A < 0.0f || slope_x_2 > 1.0f / tanThetaI) ? slope_x_1 : slope_x_2;
            float Sg;
            if (sample.y > 0.5f) { Sg = 1.0f; sample.y = 2.0f * (sample.y - 0.5f); }
            else { Sg = -1.0f; sample.y = 2.0f * (0.5f - sample.y); }
            float z = (sample.y * (sample.y * (sample.y * (-0.365728915865723f) + 0.790235037209296f) - 0.424965825137544f) + 0.000152998850436920f) /
                      (sample.y * (sample.y * (sample.y * (sample.y * 0.169507819808272f - 0.397203533833404f) - 0.232500544458471f) + 1.0f) - 0.539825872510702f);
            slope.y = Sg * z * sqrtf(1.0f + slope.x * slope.x);
        }
        return slope;
    }
    // :421-459
    DEVOLM float3 sampleVisibleNormal(float3 _wi, float2 sample) const {
        float3 wi = normalize(f3(alphaU * _wi.x, alphaV * _wi.y, _wi.z));
        float theta = 0, phi = 0;
        if (wi.z < 0.99999f) { theta = acosf(wi.z); phi = atan2f(wi.y, wi.x); }
        float sp, cp; sincosf(phi, &sp, &cp);
        float2 slope = sampleVisible11(theta, sample);
        slope = make_float2(cp * slope.x - sp * slope.y, sp * slope.x + cp * slope.y);
        slope.x *= alphaU; slope.y *= alphaV;
        float normalization = 1.0f / sqrtf(slope.x * slope.x + slope.y * slope.y + 1.0f);
        return f3(-slope.x * normalization, -slope.y * normalization, normalization);
    }
    // :243-256
    DEVFN float3 sample(float3 wi, float2 s, float &pdf) const {
        if (sampleVisible) { float3 m = sampleVisibleNormal(wi, s); pdf = pdfVisible(wi, m); return m; }
        return sampleAll(s, pdf);
    }
};

// ---------------------------------------------------------------------------------------
// evalCubicInterp1D / 2D (src/libcore/spline.cpp:23-58, 236-297) on [0,1]
// ---------------------------------------------------------------------------------------
DEVFN float cubicInterp1D(float x, const float *values, int size) {
    if (!(x >= 0.0f && x <= 1.0f)) return 0.0f;
    float t = x * (size - 1);
    int k = max(0, min((int)t, size - 2));
    float f0 = __ldg(values + k), f1 = __ldg(values + k + 1), d0, d1;
    if (k > 0) d0 = 0.5f * (f1 - __ldg(values + k - 1)); else d0 = f1 - f0;
    if (k + 2 < size) d1 = 0.5f * (__ldg(values + k + 2) - f0); else d1 = f1 - f0;
    t = t - (float)k;
    float t2 = t * t, t3 = t2 * t;
    return (2 * t3 - 3 * t2 + 1) * f0 + (-2 * t3 + 3 * t2) * f1 + (t3 - 2 * t2 + t) * d0 + (t3 - t2) * d1;
}
DEVFN void cubicWeights(float p, int size, int &knot, float w[4]) {
    float t = p * (size - 1);
    knot = min((int)t, size - 2);
    t = t - (float)knot;
    float t2 = t * t, t3 = t2 * t;
    w[0] = 0.0f; w[1] = 2 * t3 - 3 * t2 + 1; w[2] = -2 * t3 + 3 * t2; w[3] = 0.0f;
    float d0 = t3 - 2 * t2 + t, d1 = t3 - t2;
    if (knot > 0) { w[2] += 0.5f * d0; w[0] -= 0.5f * d0; } else { w[2] += d0; w[1] -= d0; }
    if (knot + 2 < size) { w[3] += 0.5f * d1; w[1] -= 0.5f * d1; } else { w[2] += d1; w[1] -= d1; }
}
DEVFN float cubicInterp2D(float px, float py, const float *values, int sx, int sy) {
    if (!(px >= 0.0f && px <= 1.0f) || !(py >= 0.0f && py <= 1.0f)) return 0.0f;
    float wx[4], wy[4]; int kx, ky;
    cubicWeights(px, sx, kx, wx);
    cubicWeights(py, sy, ky, wy);
    float result = 0.0f;
#pragma unroll
    for (int y = -1; y <= 2; ++y) {
#pragma unroll
        for (int x = -1; x <= 2; ++x) {
            float wxy = wx[x + 1] * wy[y + 1];
            if (wxy == 0) continue;
            result += __ldg(values + (ky + y) * sx + kx + x) * wxy;
        }
    }
    return result;
}
// RoughTransmittance::eval on the host-reduced slices (src/bsdfs/rtrans.h:183-233)
DEVOL float roughTransmittance(const Mats S, const MtsBsdfNode &n, float ct, float alpha) {
    if (!(ct >= 0)) return 0.f;
    float warpedCosTheta = sqrtf(sqrtf(fabsf(ct))), result;  // std::pow(x, 0.25)
    const float *tab = S.pool + n.lut[0];
    if (!(n.flags & MTS_BF_RT_2D)) result = cubicInterp1D(warpedCosTheta, tab, n.lut[1]);
    else {
        float warpedAlpha = sqrtf(sqrtf((alpha - n.f[0]) / (n.f[1] - n.f[0])));
        result = cubicInterp2D(warpedCosTheta, warpedAlpha, tab, n.lut[1], n.lut[2]);
    }
    return fminf(1.0f, fmaxf(0.0f, result));
}
// 1 - m_internalRoughTransmittance->evalDiffuse(alpha)  (roughplastic.cpp:366, rtrans.h:249-285)
DEVFN float roughInternalFdr(const Mats S, const MtsBsdfNode &n, float alpha) {
    if (!(n.flags & MTS_BF_RT_2D)) return n.f[3];
    float warpedAlpha = sqrtf(sqrtf((alpha - n.f[0]) / (n.f[1] - n.f[0])));
    float r = cubicInterp1D(warpedAlpha, S.pool + n.lut[3], n.lut[2]);
    return 1.0f - fminf(1.0f, fmaxf(0.0f, r));
}

DEVFN float3 reflectZ(float3 wi) { return f3(-wi.x, -wi.y, wi.z); }
DEVFN float3 reflectM(float3 wi, float3 m) { return 2 * dot(wi, m) * m - wi; }
// BSDF::getMeasure (include/mitsuba/render/bsdf.h)
DEVFN int measureOf(uint32_t sampledType) { return (sampledType & MTS_ESMOOTH) ? MTS_MEASURE_SOLID_ANGLE : MTS_MEASURE_DISCRETE; }
DEVFN float3 nodeEta(const MtsBsdfNode &n) { return f3(n.eta[0], n.eta[1], n.eta[2]); }
DEVFN float3 nodeK(const MtsBsdfNode &n) { return f3(n.k[0], n.k[1], n.k[2]); }

struct BSample { float3 wo; float pdf; float eta; uint32_t type; float u3; };   // u3: input, the extra number of EUsesSampler models

// The texture parameters of one leaf model at one hit (slot use: include/mtscuda.h), looked up ONCE per hit and leaf:
// eval, pdf and sample of a bounce share them (the reference's lazy per-call Texture::eval returns the same values).
// c3 = roughdielectric's specularTransmittance, kept in the k / lut slots of the node.
struct LeafTex { float3 c0, c1, c2, c3; };
DEVFN MtsTexRef roughDielTransmittance(const MtsBsdfNode &n) { MtsTexRef r; r.scale[0] = n.k[0]; r.scale[1] = n.k[1]; r.scale[2] = n.k[2]; r.tex = n.lut[0]; return r; }
DEVFN LeafTex resolveLeaf(const Mats S, const MtsBsdfNode &n, const TexCtx &tc, EwaDefer *defer = nullptr) {
    LeafTex lt;
    evalTexPair(S, n.t0, n.t1, tc, lt.c0, lt.c1, defer);
    lt.c2 = evalTex(S, n.t2, tc);
    lt.c3 = n.type == MTS_BSDF_ROUGHDIELECTRIC ? evalTex(S, roughDielTransmittance(n), tc) : f3(0.0f);
    return lt;
}

// RoughDielectric (src/bsdfs/roughdielectric.cpp) in functions of its own: the three leaf switches below stay as small
// as they were (inlined into them the model cost every material 6 percent of shade time: registers, instruction cache)
static __device__ __noinline__ float3 roughDielEval(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo) {  // :272-348
    if (cosTheta(wi) == 0) return f3(0.0f);
    const float eta = n.eta[0], invEta = 1 / eta;
    const bool reflect = cosTheta(wi) * cosTheta(wo) > 0;
    float3 H;
    if (reflect) H = normalize(wo + wi);
    else H = normalize(wi + wo * (cosTheta(wi) > 0 ? eta : invEta));
    H = H * signum(cosTheta(H));
    Microfacet distr(n.flags & MTS_BF_GGX, average(lt.c1), average(lt.c2), n.flags & MTS_BF_SAMPLE_VISIBLE);
    const float D = distr.eval(H);
    if (D == 0) return f3(0.0f);
    const float F = fresnelDielectricExt(dot(wi, H), eta);
    const float G = distr.G(wi, wo, H);
    if (reflect) return lt.c0 * (F * D * G / (4.0f * fabsf(cosTheta(wi))));
    const float e = cosTheta(wi) > 0.0f ? eta : invEta;
    const float sqrtDenom = dot(wi, H) + e * dot(wo, H);
    const float value = ((1 - F) * D * G * e * e * dot(wi, H) * dot(wo, H)) / (cosTheta(wi) * sqrtDenom * sqrtDenom);
    const float factor = cosTheta(wi) > 0 ? invEta : eta;   // ERadiance
    return lt.c3 * fabsf(value * factor * factor);
}
static __device__ __noinline__ float roughDielPdf(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo) {  // :350-420
    const float eta = n.eta[0], invEta = 1 / eta;
    const bool reflect = cosTheta(wi) * cosTheta(wo) > 0;
    float3 H; float dwh_dwo;
    if (reflect) {
        H = normalize(wo + wi);
        dwh_dwo = 1.0f / (4.0f * dot(wo, H));
    } else {
        const float e = cosTheta(wi) > 0 ? eta : invEta;
        H = normalize(wi + wo * e);
        const float sqrtDenom = dot(wi, H) + e * dot(wo, H);
        dwh_dwo = (e * e * dot(wo, H)) / (sqrtDenom * sqrtDenom);
    }
    H = H * signum(cosTheta(H));
    Microfacet sampleDistr(n.flags & MTS_BF_GGX, average(lt.c1), average(lt.c2), n.flags & MTS_BF_SAMPLE_VISIBLE);
    if (!sampleDistr.sampleVisible) sampleDistr.scaleAlpha(1.2f - 0.2f * sqrtf(fabsf(cosTheta(wi))));
    float prob = sampleDistr.pdf(wi * signum(cosTheta(wi)), H);
    const float F = fresnelDielectricExt(dot(wi, H), eta);
    prob *= reflect ? F : (1 - F);
    return fabsf(prob * dwh_dwo);
}
static __device__ __noinline__ float3 roughDielSample(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float2 sample, BSample &o) {  // :517-616
    const float eta = n.eta[0], invEta = 1 / eta;
    Microfacet distr(n.flags & MTS_BF_GGX, average(lt.c1), average(lt.c2), n.flags & MTS_BF_SAMPLE_VISIBLE);
    Microfacet sampleDistr = distr;
    if (!distr.sampleVisible) sampleDistr.scaleAlpha(1.2f - 0.2f * sqrtf(fabsf(cosTheta(wi))));
    float microfacetPDF;
    const float3 m = sampleDistr.sample(wi * signum(cosTheta(wi)), sample, microfacetPDF);
    if (microfacetPDF == 0) return f3(0.0f);
    o.pdf = microfacetPDF;
    float cosThetaT;
    const float F = fresnelDielectricExt(dot(wi, m), cosThetaT, eta);
    float3 weight = f3(1.0f);
    bool sampleReflection = true;
    if (o.u3 > F) { sampleReflection = false; o.pdf *= 1 - F; } else o.pdf *= F;   // bRec.sampler->next1D()
    float dwh_dwo;
    if (sampleReflection) {
        o.wo = reflectM(wi, m); o.eta = 1.0f; o.type = MTS_EGLOSSY_REFLECTION;
        if (cosTheta(wi) * cosTheta(o.wo) <= 0) return f3(0.0f);
        weight *= lt.c0;
        dwh_dwo = 1.0f / (4.0f * dot(o.wo, m));
    } else {
        if (cosThetaT == 0) return f3(0.0f);
        const float e = cosThetaT < 0 ? invEta : eta;   // refract(wi, m, eta, cosThetaT), util.cpp:767-772
        o.wo = m * (dot(wi, m) * e + cosThetaT) - wi * e;
        o.eta = cosThetaT < 0 ? eta : invEta; o.type = MTS_EGLOSSY_TRANSMISSION;
        if (cosTheta(wi) * cosTheta(o.wo) >= 0) return f3(0.0f);
        const float factor = cosThetaT < 0 ? invEta : eta;   // ERadiance
        weight *= lt.c3 * (factor * factor);
        const float sqrtDenom = dot(wi, m) + o.eta * dot(o.wo, m);
        dwh_dwo = (o.eta * o.eta * dot(o.wo, m)) / (sqrtDenom * sqrtDenom);
    }
    if (distr.sampleVisible) weight *= distr.smithG1(o.wo, m);
    else weight *= fabsf(distr.eval(m) * distr.G(wi, o.wo, m) * dot(wi, m) / (microfacetPDF * cosTheta(wi)));
    o.pdf *= fabsf(dwh_dwo);
    return weight;
}

// ---------------------------------------------------------------------------------------
// leaf models
// ---------------------------------------------------------------------------------------
// RoughPlastic::eval / pdf in the solid angle measure (src/bsdfs/roughplastic.cpp:326-382, 384-437); inlined into the leaf
// switches AND into roughplastic's sample(), which returns eval / pdf of the sampled direction (:494-500)
DEVOL float3 roughPlasticEval(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo) {
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return f3(0.0f);
    float eta = n.eta[0];
    float a = average(lt.c2);
    Microfacet distr(n.flags & MTS_BF_GGX, a, a, n.flags & MTS_BF_SAMPLE_VISIBLE);
    float3 result = f3(0.0f);
    {
        const float3 H = normalize(wo + wi);
        const float D = distr.eval(H);
        const float F = fresnelDielectricExt(dot(wi, H), eta);
        const float G = distr.G(wi, wo, H);
        float value = F * D * G / (4.0f * cosTheta(wi));
        result += lt.c1 * value;
    }
    {
        float3 diff = lt.c0;
        float T12 = roughTransmittance(S, n, cosTheta(wi), distr.alphaU);
        float T21 = roughTransmittance(S, n, cosTheta(wo), distr.alphaU);
        float Fdr = roughInternalFdr(S, n, distr.alphaU);
        if (n.flags & MTS_BF_NONLINEAR) diff = diff / (f3(1.0f) - diff * Fdr); else diff = diff / (1 - Fdr);
        float invEta2 = 1.0f / (eta * eta);
        result += diff * (kInvPi * cosTheta(wo) * T12 * T21 * invEta2);
    }
    return result;
}
DEVOL float roughPlasticPdf(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo) {
    if (cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return 0.0f;
    float a = average(lt.c2);
    Microfacet distr(n.flags & MTS_BF_GGX, a, a, n.flags & MTS_BF_SAMPLE_VISIBLE);
    const float3 H = normalize(wo + wi);
    float w = n.f[2];
    float probSpecular = 1 - roughTransmittance(S, n, cosTheta(wi), distr.alphaU);
    probSpecular = (probSpecular * w) / (probSpecular * w + (1 - probSpecular) * (1 - w));
    float probDiffuse = 1 - probSpecular;
    const float dwh_dwo = 1.0f / (4.0f * dot(wo, H));
    const float prob = distr.pdf(wi, H);
    float result = prob * dwh_dwo * probSpecular;
    result += probDiffuse * squareToCosineHemispherePdf(wo);
    return result;
}
DEVFN float3 leafEvalImpl(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo, int measure) {
    const bool solid = measure == MTS_MEASURE_SOLID_ANGLE, discrete = measure == MTS_MEASURE_DISCRETE;
    switch (n.type) {
    case MTS_BSDF_DIFFUSE:  // src/bsdfs/diffuse.cpp:110-118
        if (!solid || cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return f3(0.0f);
        return lt.c0 * (kInvPi * cosTheta(wo));
    case MTS_BSDF_CONDUCTOR:  // conductor.cpp:223-237
        if (!discrete || cosTheta(wi) <= 0 || cosTheta(wo) <= 0 || fabsf(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return f3(0.0f);
        return lt.c0 * fresnelConductorExact(cosTheta(wi), nodeEta(n), nodeK(n));
    case MTS_BSDF_ROUGHCONDUCTOR: {  // roughconductor.cpp:257-294
        if (!solid || cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return f3(0.0f);
        float3 H = normalize(wo + wi);
        Microfacet distr(n.flags & MTS_BF_GGX, average(lt.c1), average(lt.c2), n.flags & MTS_BF_SAMPLE_VISIBLE);
        const float D = distr.eval(H);
        if (D == 0) return f3(0.0f);
        const float3 F = fresnelConductorExact(dot(wi, H), nodeEta(n), nodeK(n)) * lt.c0;
        const float G = distr.G(wi, wo, H);
        float model = D * G / (4.0f * cosTheta(wi));
        return F * model;
    }
    case MTS_BSDF_PLASTIC: {  // plastic.cpp:245-278
        if (cosTheta(wo) <= 0 || cosTheta(wi) <= 0) return f3(0.0f);
        float eta = n.eta[0];
        float Fi = fresnelDielectricExt(cosTheta(wi), eta);
        if (discrete) {
            if (fabsf(dot(reflectZ(wi), wo) - 1) < kDeltaEps) return lt.c1 * Fi;
        } else if (solid) {
            float Fo = fresnelDielectricExt(cosTheta(wo), eta);
            float3 diff = lt.c0;
            if (n.flags & MTS_BF_NONLINEAR) diff = diff / (f3(1.0f) - diff * n.f[0]); else diff = diff / (1 - n.f[0]);
            float invEta2 = 1 / (eta * eta);
            return diff * (squareToCosineHemispherePdf(wo) * invEta2 * (1 - Fi) * (1 - Fo));
        }
        return f3(0.0f);
    }
    case MTS_BSDF_ROUGHPLASTIC: return solid ? roughPlasticEval(S, n, lt, wi, wo) : f3(0.0f);
    case MTS_BSDF_DIELECTRIC: {  // dielectric.cpp:228-253
        float eta = n.eta[0], invEta = 1 / eta;
        float cosThetaT;
        float F = fresnelDielectricExt(cosTheta(wi), cosThetaT, eta);
        if (cosTheta(wi) * cosTheta(wo) >= 0) {
            if (!discrete || fabsf(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return f3(0.0f);
            return lt.c0 * F;
        } else {
            float scale = -(cosThetaT < 0 ? invEta : eta);
            float3 refr = f3(scale * wi.x, scale * wi.y, cosThetaT);
            if (!discrete || fabsf(dot(refr, wo) - 1) > kDeltaEps) return f3(0.0f);
            float factor = cosThetaT < 0 ? invEta : eta;  // ERadiance
            return lt.c1 * factor * factor * (1 - F);
        }
    }
    case MTS_BSDF_ROUGHDIELECTRIC: return solid ? roughDielEval(S, n, lt, wi, wo) : f3(0.0f);
    case MTS_BSDF_THINDIELECTRIC: {  // thindielectric.cpp:145-166
        float R = fresnelDielectricExt(fabsf(cosTheta(wi)), n.eta[0]), T = 1 - R;
        if (R < 1) R += T * T * R / (1 - R * R);   // internal reflections: R' = R + TRT + TR^3T + ..
        if (cosTheta(wi) * cosTheta(wo) >= 0) {
            if (!discrete || fabsf(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return f3(0.0f);
            return lt.c0 * R;
        } else {
            if (!discrete || fabsf(dot(-wi, wo) - 1) > kDeltaEps) return f3(0.0f);
            return lt.c1 * (1 - R);
        }
    }
    }
    return f3(0.0f);
}

DEVFN float leafPdfImpl(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo, int measure) {
    const bool solid = measure == MTS_MEASURE_SOLID_ANGLE, discrete = measure == MTS_MEASURE_DISCRETE;
    switch (n.type) {
    case MTS_BSDF_DIFFUSE:  // diffuse.cpp:120-127
        if (!solid || cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return 0.0f;
        return squareToCosineHemispherePdf(wo);
    case MTS_BSDF_CONDUCTOR:  // conductor.cpp:239-252
        if (!discrete || cosTheta(wi) <= 0 || cosTheta(wo) <= 0 || fabsf(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return 0.0f;
        return 1.0f;
    case MTS_BSDF_ROUGHCONDUCTOR: {  // roughconductor.cpp:296-321
        if (!solid || cosTheta(wi) <= 0 || cosTheta(wo) <= 0) return 0.0f;
        float3 H = normalize(wo + wi);
        Microfacet distr(n.flags & MTS_BF_GGX, average(lt.c1), average(lt.c2), n.flags & MTS_BF_SAMPLE_VISIBLE);
        if (distr.sampleVisible) return distr.eval(H) * distr.smithG1(wi, H) / (4.0f * cosTheta(wi));
        return distr.pdf(wi, H) / (4 * absDot(wo, H));
    }
    case MTS_BSDF_PLASTIC: {  // plastic.cpp:280-307
        if (cosTheta(wo) <= 0 || cosTheta(wi) <= 0) return 0.0f;
        float Fi = fresnelDielectricExt(cosTheta(wi), n.eta[0]);
        float w = n.f[2];
        float probSpecular = (Fi * w) / (Fi * w + (1 - Fi) * (1 - w));
        if (discrete) {
            if (fabsf(dot(reflectZ(wi), wo) - 1) < kDeltaEps) return probSpecular;
        } else if (solid) return squareToCosineHemispherePdf(wo) * (1 - probSpecular);
        return 0.0f;
    }
    case MTS_BSDF_ROUGHPLASTIC: return solid ? roughPlasticPdf(S, n, lt, wi, wo) : 0.0f;
    case MTS_BSDF_DIELECTRIC: {  // dielectric.cpp:255-275
        float eta = n.eta[0], invEta = 1 / eta;
        float cosThetaT;
        float F = fresnelDielectricExt(cosTheta(wi), cosThetaT, eta);
        if (cosTheta(wi) * cosTheta(wo) >= 0) {
            if (!discrete || fabsf(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return 0.0f;
            return F;
        } else {
            float scale = -(cosThetaT < 0 ? invEta : eta);
            float3 refr = f3(scale * wi.x, scale * wi.y, cosThetaT);
            if (!discrete || fabsf(dot(refr, wo) - 1) > kDeltaEps) return 0.0f;
            return 1 - F;
        }
    }
    case MTS_BSDF_ROUGHDIELECTRIC: return solid ? roughDielPdf(S, n, lt, wi, wo) : 0.0f;
    case MTS_BSDF_THINDIELECTRIC: {  // thindielectric.cpp:168-189
        float R = fresnelDielectricExt(fabsf(cosTheta(wi)), n.eta[0]), T = 1 - R;
        if (R < 1) R += T * T * R / (1 - R * R);
        if (cosTheta(wi) * cosTheta(wo) >= 0) {
            if (!discrete || fabsf(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return 0.0f;
            return R;
        } else {
            if (!discrete || fabsf(dot(-wi, wo) - 1) > kDeltaEps) return 0.0f;
            return 1 - R;
        }
    }
    }
    return 0.0f;
}

// Entry points.  The bodies above are force-inlined into each of them, so that the fused eval + pdf of the light-sampling
// step (and roughplastic's sample, which returns eval / pdf) computes the shared terms -- half vector, D, G, Fresnel,
// rough-transmittance lookups -- once.
static __device__ __noinline__ float3 leafEval(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo, int measure) {
    return leafEvalImpl(S, n, lt, wi, wo, measure);
}
static __device__ __noinline__ float leafPdf(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo, int measure) {
    return leafPdfImpl(S, n, lt, wi, wo, measure);
}
static __device__ __noinline__ float3 leafEvalPdf(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float3 wo, int measure, float &pdf) {
    pdf = leafPdfImpl(S, n, lt, wi, wo, measure);
    return leafEvalImpl(S, n, lt, wi, wo, measure);
}

static __device__ __noinline__ float3 leafSample(const Mats S, const MtsBsdfNode &n, const LeafTex &lt, float3 wi, float2 sample, BSample &o) {
    switch (n.type) {
    case MTS_BSDF_DIFFUSE:  // diffuse.cpp:141-150
        if (cosTheta(wi) <= 0) return f3(0.0f);
        o.wo = squareToCosineHemisphere(sample);
        o.eta = 1.0f; o.type = MTS_EDIFFUSE_REFLECTION;
        o.pdf = squareToCosineHemispherePdf(o.wo);
        return lt.c0;
    case MTS_BSDF_CONDUCTOR:  // conductor.cpp:270-286
        if (cosTheta(wi) <= 0) return f3(0.0f);
        o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.eta = 1.0f; o.pdf = 1;
        return lt.c0 * fresnelConductorExact(cosTheta(wi), nodeEta(n), nodeK(n));
    case MTS_BSDF_ROUGHCONDUCTOR: {  // roughconductor.cpp:368-410
        if (cosTheta(wi) < 0) return f3(0.0f);
        Microfacet distr(n.flags & MTS_BF_GGX, average(lt.c1), average(lt.c2), n.flags & MTS_BF_SAMPLE_VISIBLE);
        float pdf;
        float3 m = distr.sample(wi, sample, pdf);
        o.pdf = pdf;
        if (pdf == 0) return f3(0.0f);
        o.wo = reflectM(wi, m); o.eta = 1.0f; o.type = MTS_EGLOSSY_REFLECTION;
        if (cosTheta(o.wo) <= 0) return f3(0.0f);
        float3 F = fresnelConductorExact(dot(wi, m), nodeEta(n), nodeK(n)) * lt.c0;
        float weight;
        if (distr.sampleVisible) weight = distr.smithG1(o.wo, m);
        else weight = distr.eval(m) * distr.G(wi, o.wo, m) * dot(wi, m) / (pdf * cosTheta(wi));
        o.pdf = pdf / (4.0f * dot(o.wo, m));
        return F * weight;
    }
    case MTS_BSDF_PLASTIC: {  // plastic.cpp:373-430
        if (cosTheta(wi) <= 0) return f3(0.0f);
        float eta = n.eta[0];
        float Fi = fresnelDielectricExt(cosTheta(wi), eta);
        o.eta = 1.0f;
        float w = n.f[2];
        float probSpecular = (Fi * w) / (Fi * w + (1 - Fi) * (1 - w));
        if (sample.x < probSpecular) {
            o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.pdf = probSpecular;
            return lt.c1 * Fi / probSpecular;
        }
        o.type = MTS_EDIFFUSE_REFLECTION;
        o.wo = squareToCosineHemisphere(make_float2((sample.x - probSpecular) / (1 - probSpecular), sample.y));
        float Fo = fresnelDielectricExt(cosTheta(o.wo), eta);
        float3 diff = lt.c0;
        if (n.flags & MTS_BF_NONLINEAR) diff = diff / (f3(1.0f) - diff * n.f[0]); else diff = diff / (1 - n.f[0]);
        o.pdf = (1 - probSpecular) * squareToCosineHemispherePdf(o.wo);
        float invEta2 = 1 / (eta * eta);
        return diff * (invEta2 * (1 - Fi) * (1 - Fo) / (1 - probSpecular));
    }
    case MTS_BSDF_ROUGHPLASTIC: {  // roughplastic.cpp:439-501
        if (cosTheta(wi) <= 0) return f3(0.0f);
        bool choseSpecular = true;
        float a = average(lt.c2);
        Microfacet distr(n.flags & MTS_BF_GGX, a, a, n.flags & MTS_BF_SAMPLE_VISIBLE);
        float w = n.f[2];
        float probSpecular = 1 - roughTransmittance(S, n, cosTheta(wi), distr.alphaU);
        probSpecular = (probSpecular * w) / (probSpecular * w + (1 - probSpecular) * (1 - w));
        if (sample.y < probSpecular) sample.y /= probSpecular;
        else { sample.y = (sample.y - probSpecular) / (1 - probSpecular); choseSpecular = false; }
        if (choseSpecular) {
            float dummy;
            float3 m = distr.sampleVisible ? distr.sampleVisibleNormal(wi, sample) : distr.sampleAll(sample, dummy);
            o.wo = reflectM(wi, m); o.type = MTS_EGLOSSY_REFLECTION;
            if (cosTheta(o.wo) <= 0) return f3(0.0f);
        } else {
            o.type = MTS_EDIFFUSE_REFLECTION;
            o.wo = squareToCosineHemisphere(sample);
        }
        o.eta = 1.0f;
        o.pdf = roughPlasticPdf(S, n, lt, wi, o.wo);   // roughplastic.cpp:494-500: pdf() and eval() / pdf, solid angle measure
        if (o.pdf == 0) return f3(0.0f);
        return roughPlasticEval(S, n, lt, wi, o.wo) / o.pdf;
    }
    case MTS_BSDF_DIELECTRIC: {  // dielectric.cpp:277-299
        float eta = n.eta[0], invEta = 1 / eta;
        float cosThetaT;
        float F = fresnelDielectricExt(cosTheta(wi), cosThetaT, eta);
        if (sample.x <= F) {
            o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.eta = 1.0f; o.pdf = F;
            return lt.c0;
        }
        o.type = MTS_EDELTA_TRANSMISSION;
        float scale = -(cosThetaT < 0 ? invEta : eta);
        o.wo = f3(scale * wi.x, scale * wi.y, cosThetaT);
        o.eta = cosThetaT < 0 ? eta : invEta;
        o.pdf = 1 - F;
        float factor = cosThetaT < 0 ? invEta : eta;
        return lt.c1 * (factor * factor);
    }
    case MTS_BSDF_ROUGHDIELECTRIC: return roughDielSample(S, n, lt, wi, sample, o);
    case MTS_BSDF_THINDIELECTRIC: {  // thindielectric.cpp:191-221 (the pdf-returning overload path.cpp:209 calls)
        float R = fresnelDielectricExt(fabsf(cosTheta(wi)), n.eta[0]), T = 1 - R;
        if (R < 1) R += T * T * R / (1 - R * R);
        o.eta = 1.0f;
        if (sample.x <= R) {
            o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.pdf = R;
            return lt.c0;
        }
        o.type = MTS_ENULL; o.wo = -wi; o.pdf = 1 - R;
        return lt.c1;
    }
    }
    return f3(0.0f);
}

// ---------------------------------------------------------------------------------------
// wrappers
// ---------------------------------------------------------------------------------------
// Surface context the wrappers need (bumpmap perturbs the frame; textures only need uv).
struct Surf {
    TexCtx tc;
    float3 dpdu, dpdv, geoN;
    Frame sh;
    // one-entry cache of the perturbed frame: eval, pdf and sample of one hit share a single gradient lookup
    mutable int bumpNode;
    mutable Frame bumpCache;
};

// BumpMap::getFrame (src/bsdfs/bumpmap.cpp:135-159)
DEVFN Frame bumpFrame(const Mats S, const MtsBsdfNode &n, const Surf &su) {
    float3 g0, g1;
    evalTexGradient(S, n.t0, su.tc.uv, g0, g1);
    float dDispDu = luminance(g0), dDispDv = luminance(g1);
    float3 dpdu = su.dpdu + su.sh.n * (dDispDu - dot(su.sh.n, su.dpdu));
    float3 dpdv = su.dpdv + su.sh.n * (dDispDv - dot(su.sh.n, su.dpdv));
    Frame r;
    r.n = normalize(cross(dpdu, dpdv));
    r.s = normalize(dpdu - r.n * dot(r.n, dpdu));
    r.t = cross(r.n, r.s);
    if (dot(r.n, su.geoN) < 0) r.n = -r.n;
    return r;
}
// NormalMap::getFrame (src/bsdfs/normalmap.cpp:106-123): the texture is read unfiltered (m_normals->eval(its, false))
DEVFN Frame normalMapFrame(const Mats S, const MtsBsdfNode &n, const Surf &su) {
    const float3 v = evalTex(S, n.t0, texCtx(su.tc.uv));
    const float3 nl = f3(2 * v.x - 1, 2 * v.y - 1, 2 * v.z - 1);
    Frame r;
    r.n = normalize(su.sh.toWorld(nl));
    r.s = normalize(su.dpdu - r.n * dot(r.n, su.dpdu));
    r.t = cross(r.n, r.s);
    return r;
}
// the frame adapters: bumpmap and normalmap differ only in how the perturbed frame is found
DEVFN Frame perturbedFrame(const Mats S, const MtsBsdfNode &n, const Surf &su) {
    return n.type == MTS_BSDF_NORMALMAP ? normalMapFrame(S, n, su) : bumpFrame(S, n, su);
}
DEVFN Frame bumpFrameCached(const Mats S, int node, const MtsBsdfNode &n, const Surf &su) {
    if (su.bumpNode == node) return su.bumpCache;
    Frame f = perturbedFrame(S, n, su);
    su.bumpNode = node; su.bumpCache = f;
    return f;
}
// SmoothCoating::refractIn / refractOut (src/bsdfs/coating.cpp:208-219)
DEVFN float3 coatRefractIn(float eta, float3 wi, float &R) {
    float cosThetaT, invEta = 1 / eta;
    R = fresnelDielectricExt(fabsf(cosTheta(wi)), cosThetaT, eta);
    return f3(invEta * wi.x, invEta * wi.y, -signum(cosTheta(wi)) * cosThetaT);
}
DEVFN float3 coatRefractOut(float eta, float3 wi, float &R) {
    float cosThetaT, invEta = 1 / eta;
    R = fresnelDielectricExt(fabsf(cosTheta(wi)), cosThetaT, invEta);
    return f3(eta * wi.x, eta * wi.y, -signum(cosTheta(wi)) * cosThetaT);
}

// RoughCoating::refractTo (src/bsdfs/roughcoating.cpp:232-253); etaTo = 1/eta towards the interior, eta towards the exterior
DEVFN float3 roughCoatRefract(float etaTo, float3 wi) {
    const float sinThetaTSqr = etaTo * etaTo * sinTheta2(wi);
    if (sinThetaTSqr >= 1.0f) return f3(0.0f);
    const float cosThetaT = sqrtf(1.0f - sinThetaTSqr);
    return f3(etaTo * wi.x, etaTo * wi.y, cosTheta(wi) > 0.0f ? cosThetaT : -cosThetaT);
}
// probability of the specular lobe of roughcoating (roughcoating.cpp:332-343, 388-397)
DEVFN float roughCoatProbSpecular(const Mats S, const MtsBsdfNode &n, float3 wi, float alpha) {
    float probSpecular = 1 - roughTransmittance(S, n, fabsf(cosTheta(wi)), alpha);
    const float w = n.f[2];
    return (probSpecular * w) / (probSpecular * w + (1 - probSpecular) * (1 - w));
}
// MixtureBSDF tables in the float pool (include/mtscuda.h): N child indices, N weights, N + 1 CDF entries
struct MixtureTab {
    const float *p; int n;
    DEVFN MixtureTab(const Mats S, const MtsBsdfNode &nd) : p(S.pool + nd.lut[0]), n(nd.lut[1]) {}
    DEVFN int child(int i) const { return (int)__ldg(p + i); }
    DEVFN float weight(int i) const { return __ldg(p + n + i); }
    DEVFN float cdf(int i) const { return __ldg(p + 2 * n + i); }
    DEVFN float pdf(int i) const { return cdf(i + 1) - cdf(i); }   // DiscreteDistribution::operator[], pmf.h
    // DiscreteDistribution::sampleReuse (pmf.h:124-136, 164-169)
    DEVFN int sampleReuse(float &u) const {
        int lb = 0;                                   // std::lower_bound over the N + 1 CDF entries
        while (lb < n + 1 && cdf(lb) < u) ++lb;
        int index = min(n - 1, max(0, lb - 1));
        while (pdf(index) == 0 && index < n) ++index;
        u = (u - cdf(index)) / (cdf(index + 1) - cdf(index));
        return index;
    }
};

template <int D> struct Bsdf {
    typedef Bsdf<D + 1> Next;

    static __device__ BSDF_NESTED_ATTR float3 eval(const Mats S, int node, const Surf &su, float3 wi, float3 wo, int measure) {
        const MtsBsdfNode &n = S.nodes[node];
        switch (n.type) {
        case MTS_BSDF_COATING: {  // coating.cpp:221-258
            float eta = n.eta[0], invEta = 1 / eta;
            if (measure == MTS_MEASURE_DISCRETE && fabsf(dot(reflectZ(wi), wo) - 1) < kDeltaEps)
                return evalTex(S, n.t1, su.tc) * fresnelDielectricExt(fabsf(cosTheta(wi)), eta);
            float R12, R21;
            float3 wiP = coatRefractIn(eta, wi, R12), woP = coatRefractIn(eta, wo, R21);
            if (R12 == 1 || R21 == 1) return f3(0.0f);
            float3 result = Next::eval(S, n.child[0], su, wiP, woP, measure) * (1 - R12) * (1 - R21);
            float3 sigmaA = evalTex(S, n.t0, su.tc) * n.f[0];
            if (!isZero(sigmaA)) result *= vexp(-sigmaA * (1 / fabsf(cosTheta(wiP)) + 1 / fabsf(cosTheta(woP))));
            if (measure == MTS_MEASURE_SOLID_ANGLE) result *= invEta * invEta * cosTheta(wo) / cosTheta(woP);
            return result;
        }
        case MTS_BSDF_BLEND: {  // blendbsdf.cpp:135-152
            float weight = fminf(1.0f, fmaxf(0.0f, average(evalTex(S, n.t0, su.tc))));
            return Next::eval(S, n.child[0], su, wi, wo, measure) * (1 - weight) + Next::eval(S, n.child[1], su, wi, wo, measure) * weight;
        }
        case MTS_BSDF_MIXTURE: {  // mixturebsdf.cpp:174-189
            const MixtureTab mt(S, n);
            float3 result = f3(0.0f);
            for (int i = 0; i < mt.n; ++i) result += Next::eval(S, mt.child(i), su, wi, wo, measure) * mt.weight(i);
            return result;
        }
        case MTS_BSDF_ROUGHCOATING: {  // roughcoating.cpp:255-314
            const float eta = n.eta[0], invEta = 1 / eta;
            const bool hasNested = (S.nodes[n.child[0]].lobes & MTS_EALL) != 0, hasSpecular = measure == MTS_MEASURE_SOLID_ANGLE;
            const float alpha = average(evalTex(S, n.t2, su.tc));
            const Microfacet distr(n.flags & MTS_BF_GGX, alpha, alpha, n.flags & MTS_BF_SAMPLE_VISIBLE);
            float3 result = f3(0.0f);
            if (hasSpecular && cosTheta(wo) * cosTheta(wi) > 0) {
                const float3 H = normalize(wo + wi) * signum(cosTheta(wo));
                const float Dm = distr.eval(H);
                const float F = fresnelDielectricExt(absDot(wi, H), eta);
                const float G = distr.G(wi, wo, H);
                const float value = F * Dm * G / (4.0f * fabsf(cosTheta(wi)));
                result += evalTex(S, n.t1, su.tc) * value;
            }
            if (hasNested) {
                const float3 wiI = roughCoatRefract(invEta, wi), woI = roughCoatRefract(invEta, wo);
                float3 nested = Next::eval(S, n.child[0], su, wiI, woI, measure) *
                    roughTransmittance(S, n, fabsf(cosTheta(wi)), distr.alphaU) * roughTransmittance(S, n, fabsf(cosTheta(wo)), distr.alphaU);
                const float3 sigmaA = evalTex(S, n.t0, su.tc) * n.f[3];
                if (!isZero(sigmaA)) nested *= vexp(-sigmaA * (1 / fabsf(cosTheta(wiI)) + 1 / fabsf(cosTheta(woI))));
                if (measure == MTS_MEASURE_SOLID_ANGLE) nested *= invEta * invEta * cosTheta(wo) / cosTheta(woI);
                result += nested;
            }
            return result;
        }
        case MTS_BSDF_NORMALMAP:  // normalmap.cpp:162-179 -- same adapter as bumpmap around another frame
        case MTS_BSDF_BUMPMAP: {  // bumpmap.cpp:161-177
            Frame bf = bumpFrameCached(S, node, n, su);
            Surf pert = su;
            pert.sh = bf; pert.bumpNode = -1;
            float3 wiP = pert.sh.toLocal(su.sh.toWorld(wi)), woP = pert.sh.toLocal(su.sh.toWorld(wo));
            if (cosTheta(wo) * cosTheta(woP) <= 0) return f3(0.0f);
            return Next::eval(S, n.child[0], pert, wiP, woP, measure);
        }
        case MTS_BSDF_MASK:        // mask.cpp:106-116
        case MTS_BSDF_TWOSIDED: {  // twosided.cpp:107-119 -- one nested call site for both adapters
            int child = n.child[0];
            float3 scale = f3(1.0f);
            if (n.type == MTS_BSDF_MASK) {
                scale = evalTex(S, n.t0, su.tc);
                if (measure != MTS_MEASURE_SOLID_ANGLE)
                    return (measure == MTS_MEASURE_DISCRETE && fabsf(1 - dot(wi, -wo)) < kDeltaEps) ? f3(1.0f) - scale : f3(0.0f);
            } else if (!(cosTheta(wi) > 0)) { wi.z = -wi.z; wo.z = -wo.z; child = n.child[1]; }
            return Next::eval(S, child, su, wi, wo, measure) * scale;
        }
        default: return leafEval(S, n, resolveLeaf(S, n, su.tc), wi, wo, measure);
        }
    }

    static __device__ BSDF_NESTED_ATTR float pdf(const Mats S, int node, const Surf &su, float3 wi, float3 wo, int measure) {
        const MtsBsdfNode &n = S.nodes[node];
        switch (n.type) {
        case MTS_BSDF_COATING: {  // coating.cpp:260-296
            float eta = n.eta[0], invEta = 1 / eta;
            float R12;
            float3 wiP = coatRefractIn(eta, wi, R12);
            float w = n.f[2];
            float probSpecular = (R12 * w) / (R12 * w + (1 - R12) * (1 - w));
            if (measure == MTS_MEASURE_DISCRETE && fabsf(dot(reflectZ(wi), wo) - 1) < kDeltaEps) return probSpecular;
            float R21;
            float3 woP = coatRefractIn(eta, wo, R21);
            if (R12 == 1 || R21 == 1) return 0.0f;
            float p = Next::pdf(S, n.child[0], su, wiP, woP, measure);
            if (measure == MTS_MEASURE_SOLID_ANGLE) p *= invEta * invEta * cosTheta(wo) / cosTheta(woP);
            return p * (1 - probSpecular);
        }
        case MTS_BSDF_BLEND: {  // blendbsdf.cpp:154-173
            float weight = fminf(1.0f, fmaxf(0.0f, average(evalTex(S, n.t0, su.tc))));
            return Next::pdf(S, n.child[0], su, wi, wo, measure) * (1 - weight) + Next::pdf(S, n.child[1], su, wi, wo, measure) * weight;
        }
        case MTS_BSDF_MIXTURE: {  // mixturebsdf.cpp:191-206
            const MixtureTab mt(S, n);
            float result = 0.0f;
            for (int i = 0; i < mt.n; ++i) result += Next::pdf(S, mt.child(i), su, wi, wo, measure) * mt.pdf(i);
            return result;
        }
        case MTS_BSDF_ROUGHCOATING: {  // roughcoating.cpp:316-376
            const float eta = n.eta[0], invEta = 1 / eta;
            const bool hasNested = (S.nodes[n.child[0]].lobes & MTS_EALL) != 0, hasSpecular = measure == MTS_MEASURE_SOLID_ANGLE;
            const float3 H = normalize(wo + wi) * signum(cosTheta(wo));
            const float alpha = average(evalTex(S, n.t2, su.tc));
            const Microfacet distr(n.flags & MTS_BF_GGX, alpha, alpha, n.flags & MTS_BF_SAMPLE_VISIBLE);
            float probNested = 1.0f, probSpecular = 1.0f;
            if (hasSpecular && hasNested) { probSpecular = roughCoatProbSpecular(S, n, wi, distr.alphaU); probNested = 1 - probSpecular; }
            float result = 0.0f;
            if (hasSpecular && cosTheta(wo) * cosTheta(wi) > 0) {
                const float dwh_dwo = 1.0f / (4.0f * absDot(wo, H));
                result = distr.pdf(wi, H) * dwh_dwo * probSpecular;
            }
            if (hasNested) {
                const float3 wiI = roughCoatRefract(invEta, wi), woI = roughCoatRefract(invEta, wo);
                float prob = Next::pdf(S, n.child[0], su, wiI, woI, measure);
                if (measure == MTS_MEASURE_SOLID_ANGLE) prob *= invEta * invEta * cosTheta(wo) / cosTheta(woI);
                result += prob * probNested;
            }
            return result;
        }
        case MTS_BSDF_NORMALMAP:  // normalmap.cpp:181-197
        case MTS_BSDF_BUMPMAP: {  // bumpmap.cpp:179-197
            Frame bf = bumpFrameCached(S, node, n, su);
            Surf pert = su;
            pert.sh = bf; pert.bumpNode = -1;
            float3 wiP = pert.sh.toLocal(su.sh.toWorld(wi)), woP = pert.sh.toLocal(su.sh.toWorld(wo));
            if (cosTheta(wo) * cosTheta(woP) <= 0) return 0;
            return Next::pdf(S, n.child[0], pert, wiP, woP, measure);
        }
        case MTS_BSDF_MASK:        // mask.cpp:118-143
        case MTS_BSDF_TWOSIDED: {  // twosided.cpp:121-133
            int child = n.child[0];
            float prob = 1.0f;
            if (n.type == MTS_BSDF_MASK) {
                prob = luminance(evalTex(S, n.t0, su.tc));
                if (measure != MTS_MEASURE_SOLID_ANGLE)
                    return (measure == MTS_MEASURE_DISCRETE && fabsf(1 - dot(wi, -wo)) < kDeltaEps) ? 1 - prob : 0.0f;
            } else if (!(wi.z > 0)) { wi.z = -wi.z; wo.z = -wo.z; child = n.child[1]; }
            return Next::pdf(S, child, su, wi, wo, measure) * prob;
        }
        default: return leafPdf(S, n, resolveLeaf(S, n, su.tc), wi, wo, measure);
        }
    }

    static __device__ BSDF_NESTED_ATTR float3 sample(const Mats S, int node, const Surf &su, float3 wi, float2 sample, BSample &o) {
        const MtsBsdfNode &n = S.nodes[node];
        switch (n.type) {
        case MTS_BSDF_COATING: {  // coating.cpp:298-371
            float eta = n.eta[0], invEta = 1 / eta;
            float R12;
            float3 wiPrime = coatRefractIn(eta, wi, R12);
            float w = n.f[2];
            float probSpecular = (R12 * w) / (R12 * w + (1 - R12) * (1 - w));
            bool choseSpecular = true;
            if (sample.x < probSpecular) sample.x /= probSpecular;
            else { sample.x = (sample.x - probSpecular) / (1 - probSpecular); choseSpecular = false; }
            if (choseSpecular) {
                o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.eta = 1.0f; o.pdf = probSpecular;
                return evalTex(S, n.t1, su.tc) * (R12 / o.pdf);
            }
            if (R12 == 1.0f) return f3(0.0f);
            float3 result = Next::sample(S, n.child[0], su, wiPrime, sample, o);
            if (isZero(result)) return f3(0.0f);
            float3 woPrime = o.wo;
            float3 sigmaA = evalTex(S, n.t0, su.tc) * n.f[0];
            if (!isZero(sigmaA)) result *= vexp(-sigmaA * (1 / fabsf(cosTheta(wiPrime)) + 1 / fabsf(cosTheta(woPrime))));
            float R21;
            o.wo = coatRefractOut(eta, woPrime, R21);
            if (R21 == 1.0f) return f3(0.0f);
            o.pdf *= 1.0f - probSpecular;
            result = result / (1.0f - probSpecular);
            result *= (1 - R12) * (1 - R21);
            if (measureOf(o.type) == MTS_MEASURE_SOLID_ANGLE) o.pdf *= invEta * invEta * cosTheta(o.wo) / cosTheta(woPrime);
            return result;
        }
        case MTS_BSDF_BLEND: {  // blendbsdf.cpp:223-262
            float w1 = fminf(1.0f, fmaxf(0.0f, average(evalTex(S, n.t0, su.tc))));
            float w0 = 1 - w1;
            int entry;
            if (sample.x < w0) { entry = 0; sample.x /= w0; } else { entry = 1; sample.x = (sample.x - w0) / w1; }
            float wEntry = entry ? w1 : w0, wOther = entry ? w0 : w1;
            float3 result = Next::sample(S, n.child[entry], su, wi, sample, o);
            if (isZero(result)) return result;
            result *= wEntry * o.pdf;
            o.pdf *= wEntry;
            int measure = measureOf(o.type);
            o.pdf += Next::pdf(S, n.child[1 - entry], su, wi, o.wo, measure) * wOther;
            result += Next::eval(S, n.child[1 - entry], su, wi, o.wo, measure) * wOther;
            return result / o.pdf;
        }
        case MTS_BSDF_MIXTURE: {  // mixturebsdf.cpp:244-268 (the pdf-returning overload path.cpp:209 calls)
            const MixtureTab mt(S, n);
            const int entry = mt.sampleReuse(sample.x);
            float3 result = Next::sample(S, mt.child(entry), su, wi, sample, o);
            if (isZero(result)) return result;
            result *= mt.weight(entry) * o.pdf;
            o.pdf *= mt.pdf(entry);
            const int measure = measureOf(o.type);
            for (int i = 0; i < mt.n; ++i) {
                if (i == entry) continue;
                o.pdf += Next::pdf(S, mt.child(i), su, wi, o.wo, measure) * mt.pdf(i);
                result += Next::eval(S, mt.child(i), su, wi, o.wo, measure) * mt.weight(i);
            }
            return result / o.pdf;
        }
        case MTS_BSDF_ROUGHCOATING: {  // roughcoating.cpp:378-448
            const float eta = n.eta[0], invEta = 1 / eta;
            const bool hasNested = (S.nodes[n.child[0]].lobes & MTS_EALL) != 0;
            bool choseSpecular = true;
            const float alpha = average(evalTex(S, n.t2, su.tc));
            const Microfacet distr(n.flags & MTS_BF_GGX, alpha, alpha, n.flags & MTS_BF_SAMPLE_VISIBLE);
            if (hasNested) {
                const float probSpecular = roughCoatProbSpecular(S, n, wi, distr.alphaU);
                if (sample.y < probSpecular) sample.y /= probSpecular;
                else { sample.y = (sample.y - probSpecular) / (1 - probSpecular); choseSpecular = false; }
            }
            if (choseSpecular) {
                float dummy;
                const float3 m = distr.sampleVisible ? distr.sampleVisibleNormal(wi, sample) : distr.sampleAll(sample, dummy);
                o.wo = reflectM(wi, m); o.type = MTS_EGLOSSY_REFLECTION; o.eta = 1.0f;
                if (cosTheta(o.wo) * cosTheta(wi) <= 0) return f3(0.0f);
            } else {
                const float3 result = Next::sample(S, n.child[0], su, roughCoatRefract(invEta, wi), sample, o);
                if (isZero(result)) return f3(0.0f);
                o.wo = roughCoatRefract(eta, o.wo);
                if (isZero(o.wo)) return f3(0.0f);
            }
            const int measure = measureOf(o.type);   // "guard against numerical imprecisions": pdf() and eval() / pdf
            o.pdf = pdf(S, node, su, wi, o.wo, measure);
            if (o.pdf == 0) return f3(0.0f);
            return eval(S, node, su, wi, o.wo, measure) / o.pdf;
        }
        case MTS_BSDF_NORMALMAP:  // normalmap.cpp:218-239
        case MTS_BSDF_BUMPMAP: {  // bumpmap.cpp:221-240
            Frame bf = bumpFrameCached(S, node, n, su);
            Surf pert = su;
            pert.sh = bf; pert.bumpNode = -1;
            float3 wiP = pert.sh.toLocal(su.sh.toWorld(wi));
            float3 result = Next::sample(S, n.child[0], pert, wiP, sample, o);
            if (!isZero(result)) {
                float3 woP = o.wo;
                o.wo = su.sh.toLocal(pert.sh.toWorld(woP));
                if (cosTheta(o.wo) * cosTheta(woP) <= 0) return f3(0.0f);
            }
            return result;
        }
        case MTS_BSDF_MASK:        // mask.cpp:187-226 (pdf-returning overload)
        case MTS_BSDF_TWOSIDED: {  // twosided.cpp:158-180
            const bool mask = n.type == MTS_BSDF_MASK;
            int child = n.child[0];
            float3 opacity = f3(1.0f);
            float prob = 1.0f;
            bool flipped = false;
            if (mask) {
                opacity = evalTex(S, n.t0, su.tc);
                prob = luminance(opacity);
                if (!(sample.x < prob)) {
                    o.wo = -wi; o.eta = 1.0f; o.type = MTS_ENULL; o.pdf = 1 - prob;
                    return (f3(1.0f) - opacity) / o.pdf;
                }
                sample.x /= prob;
            } else if (cosTheta(wi) < 0) { flipped = true; wi.z = -wi.z; child = n.child[1]; }
            float3 result = Next::sample(S, child, su, wi, sample, o);
            if (mask) { result = result * opacity / prob; o.pdf *= prob; }
            else if (flipped && !isZero(result) && o.pdf != 0) o.wo.z = -o.wo.z;
            return result;
        }
        default: return leafSample(S, n, resolveLeaf(S, n, su.tc), wi, sample, o);
        }
    }
};

// recursion floor: below kMaxBsdfDepth wrappers only leaves are legal (set_bsdfs enforces it)
template <> struct Bsdf<kMaxBsdfDepth> {
    static __device__ BSDF_NESTED_ATTR float3 eval(const Mats S, int node, const Surf &su, float3 wi, float3 wo, int measure) {
        return leafEval(S, S.nodes[node], resolveLeaf(S, S.nodes[node], su.tc), wi, wo, measure);
    }
    static __device__ BSDF_NESTED_ATTR float pdf(const Mats S, int node, const Surf &su, float3 wi, float3 wo, int measure) {
        return leafPdf(S, S.nodes[node], resolveLeaf(S, S.nodes[node], su.tc), wi, wo, measure);
    }
    static __device__ BSDF_NESTED_ATTR float3 sample(const Mats S, int node, const Surf &su, float3 wi, float2 sample, BSample &o) {
        return leafSample(S, S.nodes[node], resolveLeaf(S, S.nodes[node], su.tc), wi, sample, o);
    }
};

typedef Bsdf<0> BsdfRoot;

// ---------------------------------------------------------------------------------------
// Shade classes: what MIPathTracer::Li needs from the BSDF of one hit (eval + pdf of the light sample, sample of the
// next direction), specialised by the shape of the BSDF tree so that the common cases run without the generic
// wrapper recursion -- no per-level Surf copies on the stack, leaf textures fetched once per hit, far less code
// per kernel.  mtscuda_build() classifies every root BSDF (shadeClassOf, mtscuda.cu) and orders the material sort
// keys by class, so each class is one contiguous range of the sorted queue and gets its own k_shade<> instance.
//   SHADE_LAMBERT  root is `diffuse`
//   SHADE_LEAF     root is any leaf model, optionally below the frame / side adapters twosided -> bumpmap
//                  (leaf | bumpmap(leaf) | twosided(x [, y]) with x, y = leaf | bumpmap(leaf))
//   SHADE_GENERIC  everything else (coating, blendbsdf, mask, deeper nestings): Bsdf<0> recursion
// ---------------------------------------------------------------------------------------
enum { SHADE_LAMBERT = 0, SHADE_LEAF = 1, SHADE_GENERIC = 2, SHADE_CLASSES = 3 };
// evalPdf(): when the pdf is wanted.  The fused classes always produce it (it shares its terms with the value); the
// generic class evaluates it separately and skips it when nothing will use it.
enum { PDF_NONE = 0, PDF_IF_LIT = 1, PDF_ALWAYS = 2 };

template <int CLASS> struct Material;

template <> struct Material<SHADE_GENERIC> {
    const Mats S; int node; const Surf &su;
    DEVFN Material(const Mats S_, int node_, const Surf &su_, float3) : S(S_), node(node_), su(su_) {}
    DEVFN float3 evalPdf(float3 wi, float3 wo, int measure, int pdfMode, float &pdf) const {
        pdf = 0.0f;
        const float3 v = BsdfRoot::eval(S, node, su, wi, wo, measure);
        if (pdfMode == PDF_ALWAYS || (pdfMode == PDF_IF_LIT && !isZero(v))) pdf = BsdfRoot::pdf(S, node, su, wi, wo, measure);
        return v;
    }
    DEVFN float3 sample(float3 wi, float2 u, BSample &bs) const { return BsdfRoot::sample(S, node, su, wi, u, bs); }
};

template <> struct Material<SHADE_LEAF> {
    const Mats S;
    const MtsBsdfNode *n;   // the leaf
    LeafTex lt;
    Frame sh, pf;           // shading frame of the hit, perturbed frame of the bumpmap / normalmap adapter
    float3 wiLeaf;          // incident direction as the leaf sees it
    bool bump, flipped;
    DEVFN Material(const Mats S_, int root, const Surf &su, float3 wi, EwaDefer *defer = nullptr) : S(S_), sh(su.sh), bump(false), flipped(false) {
        const MtsBsdfNode *r = &S.nodes[root];
        if (r->type == MTS_BSDF_TWOSIDED) {   // twosided.cpp:107-180: the back side sees the mirrored configuration
            int child = r->child[0];
            if (!(cosTheta(wi) > 0)) { wi.z = -wi.z; flipped = true; child = r->child[1]; }
            r = &S.nodes[child];
        }
        if (r->type == MTS_BSDF_BUMPMAP || r->type == MTS_BSDF_NORMALMAP) {   // bumpmap.cpp:161-240, normalmap.cpp:162-239: same directions expressed in the perturbed frame
            bump = true;
            pf = perturbedFrame(S, *r, su);
            wi = pf.toLocal(sh.toWorld(wi));
            r = &S.nodes[r->child[0]];
        }
        n = r; wiLeaf = wi;
        lt = resolveLeaf(S, *n, su.tc, defer);
    }
    // the staged shade kernels resolve a hit once (constructor above, k_surface) and rebuild the material from the stored
    // surface record in the scatter kernel
    DEVFN Material(const Mats S_, int leaf, const LeafTex &lt_, const Frame &sh_, const Frame &pf_, float3 wiLeaf_, bool bump_, bool flipped_)
        : S(S_), n(&S_.nodes[leaf]), lt(lt_), sh(sh_), pf(pf_), wiLeaf(wiLeaf_), bump(bump_), flipped(flipped_) {}
    DEVFN int leafIndex() const { return (int)(n - S.nodes); }
    DEVFN float3 evalPdf(float3, float3 wo, int measure, int, float &pdf) const {
        if (flipped) wo.z = -wo.z;
        float3 woL = wo;
        if (bump) {
            woL = pf.toLocal(sh.toWorld(wo));
            if (cosTheta(wo) * cosTheta(woL) <= 0) { pdf = 0.0f; return f3(0.0f); }
        }
        return leafEvalPdf(S, *n, lt, wiLeaf, woL, measure, pdf);
    }
    DEVFN float3 sample(float3, float2 u, BSample &bs) const {
        const float3 r = leafSample(S, *n, lt, wiLeaf, u, bs);
        if (isZero(r)) return r;
        if (bump) {
            const float3 woP = bs.wo;
            bs.wo = sh.toLocal(pf.toWorld(woP));
            if (cosTheta(bs.wo) * cosTheta(woP) <= 0) return f3(0.0f);
        }
        if (flipped && bs.pdf != 0) bs.wo.z = -bs.wo.z;
        return r;
    }
};

template <> struct Material<SHADE_LAMBERT> {   // src/bsdfs/diffuse.cpp:110-150, everything inline
    float3 reflectance;
    DEVFN Material(const Mats S, int root, const Surf &su, float3) { reflectance = evalTex(S, S.nodes[root].t0, su.tc); }
    DEVFN float3 evalPdf(float3 wi, float3 wo, int measure, int, float &pdf) const {
        if (measure != MTS_MEASURE_SOLID_ANGLE || cosTheta(wi) <= 0 || cosTheta(wo) <= 0) { pdf = 0.0f; return f3(0.0f); }
        pdf = squareToCosineHemispherePdf(wo);
        return reflectance * (kInvPi * cosTheta(wo));
    }
    DEVFN float3 sample(float3 wi, float2 u, BSample &bs) const {
        if (cosTheta(wi) <= 0) return f3(0.0f);
        bs.wo = squareToCosineHemisphere(u);
        bs.eta = 1.0f; bs.type = MTS_EDIFFUSE_REFLECTION;
        bs.pdf = squareToCosineHemispherePdf(bs.wo);
        return reflectance;
    }
};

inline bool isWrapperBsdf(int t) {
    return t == MTS_BSDF_COATING || t == MTS_BSDF_BLEND || t == MTS_BSDF_BUMPMAP || t == MTS_BSDF_TWOSIDED || t == MTS_BSDF_MASK ||
           t == MTS_BSDF_MIXTURE || t == MTS_BSDF_ROUGHCOATING || t == MTS_BSDF_NORMALMAP;
}
// Classification of a root BSDF (host side: mtscuda_build; also used by the per-function test entry points)
inline int shadeClassOf(const MtsBsdfNode *nodes, int root) {
    auto isLeaf = [&](int i) { return !isWrapperBsdf(nodes[i].type); };
    auto leafOrBump = [&](int i) { return isLeaf(i) || ((nodes[i].type == MTS_BSDF_BUMPMAP || nodes[i].type == MTS_BSDF_NORMALMAP) && isLeaf(nodes[i].child[0])); };
    if (nodes[root].type == MTS_BSDF_DIFFUSE) return SHADE_LAMBERT;
    if (leafOrBump(root)) return SHADE_LEAF;
    if (nodes[root].type == MTS_BSDF_TWOSIDED && leafOrBump(nodes[root].child[0]) && leafOrBump(nodes[root].child[1])) return SHADE_LEAF;
    return SHADE_GENERIC;
}


} // namespace mts
