// emitter.cuh -- direct illumination sampling for area, point and environment-map emitters.
#pragma once
#include "geometry.cuh"
#include "bsdf.cuh"

namespace mts {

struct DRec { float3 ref, refN, p, n, d; float dist, pdf; int measure; int emitter; };

// DiscreteDistribution::sample / sampleReuse (include/mitsuba/core/pmf.h:125-188); cdf has n+1 entries
DEVFN int cdfSampleReuse(const float *cdf, int n, float &v, float *pdf) {
    // std::lower_bound: first entry >= v
    int lo = 0, hi = n + 1;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (__ldg(cdf + mid) < v) lo = mid + 1; else hi = mid; }
    int index = min(n - 1, max(0, lo - 1));
    while (index < n && __ldg(cdf + index + 1) - __ldg(cdf + index) == 0) ++index;
    float c0 = __ldg(cdf + index), c1 = __ldg(cdf + index + 1);
    if (pdf) *pdf = c1 - c0;
    v = (v - c0) / (c1 - c0);
    return index;
}

// Shape::samplePosition: TriMesh (src/librender/trimesh.cpp:412-424 + src/libcore/triangle.cpp:23-56),
// Rectangle (src/shapes/rectangle.cpp:210-216), Sphere (src/shapes/sphere.cpp:257-266)
DEVFN void samplePosition(const DScene &S, const DShape &s, float2 sample, float3 &p, float3 &n, float &pdf) {
    if (s.type == SHAPE_MESH) {
        int tri = cdfSampleReuse(S.areaCdfs + s.cdfOffset, (int)s.ntris, sample.y, nullptr);
        size_t gt = (size_t)s.triOffset + tri;
        uint32_t i0 = __ldg(S.indices + 3 * gt), i1 = __ldg(S.indices + 3 * gt + 1), i2 = __ldg(S.indices + 3 * gt + 2);
        float3 p0 = loadF3(S.positions, i0), p1 = loadF3(S.positions, i1), p2 = loadF3(S.positions, i2);
        float2 bary = squareToUniformTriangle(sample);
        float3 sideA = p1 - p0, sideB = p2 - p0;
        p = p0 + (sideA * bary.x) + (sideB * bary.y);
        if (s.flags & 1)
            n = normalize(loadF3(S.normals, i0) * (1.0f - bary.x - bary.y) + loadF3(S.normals, i1) * bary.x + loadF3(S.normals, i2) * bary.y);
        else n = normalize(cross(sideA, sideB));
    } else if (s.type == SHAPE_RECT) {
        p = xfPoint(s.o2w, f3(sample.x * 2 - 1, sample.y * 2 - 1, 0));
        n = f3(s.normal);
    } else {
        float3 v = squareToUniformSphere(sample);
        p = v * s.radius + f3(s.center); n = v;
        if (s.flip) n = -n;
    }
    pdf = s.invArea;
}

// Shape::sampleDirect (src/librender/shape.cpp:99-112), Sphere::sampleDirect (sphere.cpp:286-346)
DEVFN void shapeSampleDirect(const DScene &S, const DShape &s, DRec &r, float2 sample) {
    if (s.type != SHAPE_SPHERE) {
        samplePosition(S, s, sample, r.p, r.n, r.pdf);
        r.d = r.p - r.ref;
        float distSquared = dot(r.d, r.d);
        r.dist = sqrtf(distSquared);
        r.d = r.d / r.dist;
        float dp = absDot(r.d, r.n);
        r.pdf *= dp != 0 ? (distSquared / dp) : 0.0f;
        r.measure = MTS_MEASURE_SOLID_ANGLE;
        return;
    }
    const float3 center = f3(s.center);
    const float3 refToCenter = center - r.ref;
    const float refDist2 = dot(refToCenter, refToCenter);
    const float invRefDist = 1.0f / sqrtf(refDist2);
    const float sinAlpha = s.radius * invRefDist;
    if (sinAlpha < 1 - kEps) {
        float cosAlpha = safe_sqrt(1.0f - sinAlpha * sinAlpha);
        r.d = frameFromNormal(refToCenter * invRefDist).toWorld(squareToUniformCone(cosAlpha, sample));
        r.pdf = squareToUniformConePdf(cosAlpha);
        const float projDist = dot(refToCenter, r.d);
        const float baseT = refDist2 / projDist;
        const float3 query = r.ref + r.d * baseT;
        const float3 queryToCenter = center - query;
        const float queryDist2 = dot(queryToCenter, queryToCenter);
        const float queryProjDist = dot(queryToCenter, r.d);
        float A = 1.0f, B = -2 * queryProjDist, C = queryDist2 - s.radius * s.radius;
        float nearT, farT;
        if (!solveQuadratic(A, B, C, nearT, farT)) nearT = queryProjDist;
        r.dist = baseT + nearT;
        r.n = normalize(r.d * nearT - queryToCenter);
        r.p = center + r.n * s.radius;
    } else {
        float3 d = squareToUniformSphere(sample);
        r.p = center + d * s.radius;
        r.n = d;
        r.d = r.p - r.ref;
        float dist2 = dot(r.d, r.d);
        r.dist = sqrtf(dist2);
        r.d = r.d / r.dist;
        r.pdf = s.invArea * dist2 / absDot(r.d, r.n);
    }
    if (s.flip) r.n = -r.n;
    r.measure = MTS_MEASURE_SOLID_ANGLE;
}

// Shape::pdfDirect (shape.cpp:114-124), Sphere::pdfDirect (sphere.cpp:348-377); solid-angle measure
DEVFN float shapePdfDirect(const DShape &s, float3 ref, float3 d, float3 n, float dist) {
    if (s.type != SHAPE_SPHERE) return s.invArea * (dist * dist) / absDot(d, n);
    const float3 refToCenter = f3(s.center) - ref;
    const float invRefDist = 1.0f / length(refToCenter);
    const float sinAlpha = s.radius * invRefDist;
    if (sinAlpha < 1 - kEps) return squareToUniformConePdf(safe_sqrt(1 - sinAlpha * sinAlpha));
    return s.invArea * dist * dist / absDot(d, n);
}

// ---- EnvironmentMap (src/emitters/envmap.cpp) -------------------------------------------------
DEVFN float3 envTexel(const DEnvmap &e, int x, int y) {  // repeat-u / clamp-v (:145-146,179)
    x = wrapRepeat(x, e.width); y = clampi(y, 0, e.height - 1);
    uint2 raw = __ldg(reinterpret_cast<const uint2 *>(e.texels) + (size_t)y * e.width + x);
    float2 fa = __half22float2(*reinterpret_cast<__half2 *>(&raw.x)), fb = __half22float2(*reinterpret_cast<__half2 *>(&raw.y));
    return f3(fa.x, fa.y, fb.x);
}
DEVFN uint32_t envSampleReuse(const float *cdf, uint32_t size, float &sample, const uint16_t *guide, int G) {  // :657-662
    // std::lower_bound over the size+1 CDF entries, started from the guide cell of the sample
    const int k = min((int)(sample * (float)G), G - 1);
    uint32_t lo = __ldg(guide + k), hi = __ldg(guide + k + 1);
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (__ldg(cdf + mid) < sample) lo = mid + 1; else hi = mid; }
    uint32_t index = min((uint32_t)max((int)lo - 1, 0), size - 1);
    float c0 = __ldg(cdf + index), c1 = __ldg(cdf + index + 1);
    sample = (sample - c0) / (c1 - c0);
    return index;
}
DEVFN void envBilinear(const DEnvmap &e, float px, float py, float3 &value1, float3 &value2, int &yPos) {
    int xPos = floorToInt(px); yPos = floorToInt(py);
    float dx1 = px - xPos, dx2 = 1.0f - dx1, dy1 = py - yPos, dy2 = 1.0f - dy1;
    value1 = envTexel(e, xPos, yPos) * dx2 * dy2 + envTexel(e, xPos + 1, yPos) * dx1 * dy2;
    value2 = envTexel(e, xPos, yPos + 1) * dx2 * dy1 + envTexel(e, xPos + 1, yPos + 1) * dx1 * dy1;
}
// internalSampleDirection (:567-600)
DEVFN void envSampleDirection(const DEnvmap &e, float2 sample, float3 &d, float3 &value, float &pdf) {
    uint32_t row = envSampleReuse(e.cdfRows, e.height, sample.y, e.guideRows, kGuideRows);
    uint32_t col = envSampleReuse(e.cdfCols + (size_t)row * (e.width + 1), e.width, sample.x, e.guideCols + (size_t)row * (kGuideCols + 1), kGuideCols);
    float posx = (float)col + intervalToTent(sample.x), posy = (float)row + intervalToTent(sample.y);
    float3 value1, value2; int yPos;
    envBilinear(e, posx, posy, value1, value2, yPos);
    value = (value1 + value2) * e.scale;
    pdf = (luminance(value1) * __ldg(e.rowWeights + clampi(yPos, 0, e.height - 1)) +
           luminance(value2) * __ldg(e.rowWeights + clampi(yPos + 1, 0, e.height - 1))) * e.normalization;
    float sp, cp, st, ct;
    sincosf((2 * kPi / e.width) * (posx + 0.5f), &sp, &cp);
    sincosf((kPi / e.height) * (posy + 0.5f), &st, &ct);
    d = f3(sp * st, ct, -cp * st);
    pdf /= fmaxf(fabsf(st), kEps);
}
// internalPdfDirection (:603-633); d in emitter-local space
DEVFN float envPdfDirection(const DEnvmap &e, float3 d) {
    float uvx = atan2f(d.x, -d.z) * kInvTwoPi, uvy = safe_acos(d.y) * kInvPi;
    if (!isfinite(uvx) || !isfinite(uvy)) return 0.0f;
    float3 value1, value2; int yPos;
    envBilinear(e, uvx * e.width - 0.5f, uvy * e.height - 0.5f, value1, value2, yPos);
    float st = safe_sqrt(1 - d.y * d.y);
    return (luminance(value1) * __ldg(e.rowWeights + clampi(yPos, 0, e.height - 1)) +
            luminance(value2) * __ldg(e.rowWeights + clampi(yPos + 1, 0, e.height - 1))) * e.normalization / fmaxf(fabsf(st), kEps);
}
// evalEnvironment without ray differentials (:380-410): bilinear lookup at level 0
DEVFN float3 envEval(const DEnvmap &e, float3 dWorld) {
    float3 v = xfVec(e.toLocal, dWorld);
    float uvx = atan2f(v.x, -v.z) * kInvTwoPi, uvy = safe_acos(v.y) * kInvPi;
    if (!isfinite(uvx) || !isfinite(uvy)) return f3(0.0f);
    float3 value1, value2; int yPos;
    envBilinear(e, uvx * e.width - 0.5f, uvy * e.height - 0.5f, value1, value2, yPos);
    return (value1 + value2) * e.scale;
}
// evalEnvironment of a ray with differentials (:391-407): texture-space partials of the lat-long map, then TMIPMap::eval
DEVFN float3 envEvalDiff(const DScene &S, const DEnvmap &e, float3 dWorld, float3 rxD, float3 ryD) {
    const float3 v = xfVec(e.toLocal, dWorld);
    const float2 uv = make_float2(atan2f(v.x, -v.z) * kInvTwoPi, safe_acos(v.y) * kInvPi);
    const float3 dvdx = xfVec(e.toLocal, rxD) - v, dvdy = xfVec(e.toLocal, ryD) - v;
    const float t1 = kInvTwoPi / (v.x * v.x + v.z * v.z), t2 = -kInvPi / fmaxf(safe_sqrt(1.0f - v.y * v.y), kEps);
    const float2 d0 = make_float2(t1 * (dvdx.z * v.x - dvdx.x * v.z), t2 * dvdx.y);
    const float2 d1 = make_float2(t1 * (dvdy.z * v.x - dvdy.x * v.z), t2 * dvdy.y);
    return mipEval(*e.mip, uv, d0, d1) * e.scale;
}
// BSphere::rayIntersect (include/mitsuba/core/bsphere.h:88-95)
DEVFN bool bsphereIntersect(const DEnvmap &e, float3 o, float3 d, float &nearT, float &farT) {
    float3 oc = o - e.bsCenter;
    return solveQuadratic(dot(d, d), 2 * dot(oc, d), dot(oc, oc) - e.bsRadius * e.bsRadius, nearT, farT);
}

// Scene::sampleEmitterDirect without the shadow ray (src/librender/scene.cpp:828-852) on top of
// Emitter::sampleDirect (area.cpp:158-172, point.cpp:131-147, envmap.cpp:516-543).
// Returns value / (pdf * emitter probability); r.pdf = 0 signals failure.
DEVFN float3 sampleEmitterDirect(const DScene &S, DRec &r, float2 sample) {
    float emPdf;
    int index = cdfSampleReuse(S.emitterCdf, S.numEmitters, sample.x, &emPdf);
    const DEmitter &e = S.emitters[index];
    float3 value;
    if (e.type == MTS_EMITTER_AREA) {
        shapeSampleDirect(S, S.shapes[e.shape], r, sample);
        if (dot(r.d, r.refN) >= 0 && dot(r.d, r.n) < 0 && r.pdf != 0) value = e.radiance / r.pdf;
        else { r.pdf = 0.0f; value = f3(0.0f); }
    } else if (e.type == MTS_EMITTER_POINT) {
        r.p = e.position; r.pdf = 1.0f; r.measure = MTS_MEASURE_DISCRETE;
        r.d = r.p - r.ref;
        r.dist = length(r.d);
        float invDist = 1.0f / r.dist;
        r.d *= invDist; r.n = f3(0.0f);
        value = e.radiance * (invDist * invDist);
    } else {
        float3 d; float pdf;
        envSampleDirection(S.env, sample, d, value, pdf);
        float3 dw = xfVec(S.env.toWorld, d);
        float nearT, farT;
        if (isZero(value) || pdf == 0 || !bsphereIntersect(S.env, r.ref, dw, nearT, farT) || nearT >= 0 || farT <= 0) {
            r.pdf = 0.0f; value = f3(0.0f);
        } else {
            r.pdf = pdf; r.p = r.ref + dw * farT;
            r.n = normalize(S.env.bsCenter - r.p);
            r.dist = farT; r.d = dw; r.measure = MTS_MEASURE_SOLID_ANGLE;
            value = value / pdf;
        }
    }
    if (r.pdf != 0) {
        r.emitter = index;
        r.pdf *= emPdf;
        return value / emPdf;
    }
    return f3(0.0f);
}

// AreaLight::eval (src/emitters/area.cpp:104-109)
DEVFN float3 areaLe(const DEmitter &e, float3 shN, float3 d) { return dot(shN, d) <= 0 ? f3(0.0f) : e.radiance; }

} // namespace mts
