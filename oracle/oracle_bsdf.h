/*
 * oracle_bsdf.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement of the nine in-scope Mitsuba 0.6 BSDF models, the microfacet
 * distribution helpers, the rough-transmittance LUT lookup and the bitmap
 * texture lookup.  Each block cites the reference file:line it follows.
 * typeMask is always EAll and component is always -1, which is how
 * MIPathTracer::Li queries BSDFs (src/integrators/path/path.cpp:178-207).
 */
#pragma once
#include <vector>
#include "oracle_math.h"
#include "oracle_mipmap.h"
#define MTS_PREFIX mtsoracle_
#include "../include/mtscuda.h"

namespace orc {

struct Tex {
    MtsTextureDesc d;
    MipMap mip;   /* level 0 (+ the Lanczos pyramid for trilinear / ewa), values rounded through half */
};

/* Subset of `Intersection` (include/mitsuba/render/shape.h:36-172) */
struct Its {
    V3 p; float t;
    Frame geo, sh;
    V2 uv;
    V3 dpdu, dpdv;
    V3 wi;
    int shape, prim;
    bool hasUVPartials = false;                      /* shape.h:154-166: set by computePartials for camera-ray hits */
    float dudx = 0, dvdx = 0, dudy = 0, dvdy = 0;
    V3 toLocal(V3 v) const { return sh.toLocal(v); }
    V3 toWorld(V3 v) const { return sh.toWorld(v); }
};

struct Materials {
    std::vector<MtsBsdfNode> nodes;
    std::vector<Tex> textures;
    std::vector<float> pool;
};

/* BitmapTexture::eval(uv) (src/textures/bitmap.cpp:431-454): unfiltered = bilinear at level 0, box for `nearest` */
inline V3 texLookup(const Tex &t, V2 uv) {
    if (t.d.filter == MTS_FILTER_NEAREST) return t.mip.evalBox(0, uv);
    return t.mip.evalBilinear(0, uv);
}
/* BitmapTexture::evalGradient (bitmap.cpp:456-480) -> mipmap.h:601-626 evalGradientBilinear(0, uv) */
inline void texGradient(const Tex &t, V2 uv, V3 grad[2]) {
    if (t.d.filter == MTS_FILTER_NEAREST) { grad[0] = grad[1] = V3(0.0f); return; }
    t.mip.evalGradientBilinear(uv, grad);
}

/* Texture2D::eval(its, filter = true) (src/librender/texture.cpp:104-114) + ScalingTexture (scale.cpp:86-88);
 * constants: basicshader.h:40,97.  With uv partials the lookup goes through TMIPMap::eval (bitmap.cpp:486-499). */
inline V3 evalTex(const Materials &m, const MtsTexRef &r, const Its &its) {
    V3 s(r.scale[0], r.scale[1], r.scale[2]);
    if (r.tex < 0) return s;
    const Tex &t = m.textures[r.tex];
    V2 uv{its.uv.x * t.d.uv_scale[0] + t.d.uv_offset[0], its.uv.y * t.d.uv_scale[1] + t.d.uv_offset[1]};
    if (its.hasUVPartials)
        return t.mip.eval(uv, V2{its.dudx * t.d.uv_scale[0], its.dvdx * t.d.uv_scale[1]},
                          V2{its.dudy * t.d.uv_scale[0], its.dvdy * t.d.uv_scale[1]}) * s;
    return texLookup(t, uv) * s;
}
/* Texture2D::evalGradient (texture.cpp:116-124) + ScalingTexture::evalGradient (scale.cpp:90-94);
 * a constant texture has zero gradient (Texture::evalGradient finite differences). */
inline void evalTexGradient(const Materials &m, const MtsTexRef &r, const Its &its, V3 grad[2]) {
    if (r.tex < 0) { grad[0] = grad[1] = V3(0.0f); return; }
    const Tex &t = m.textures[r.tex];
    V2 uv{its.uv.x * t.d.uv_scale[0] + t.d.uv_offset[0], its.uv.y * t.d.uv_scale[1] + t.d.uv_offset[1]};
    texGradient(t, uv, grad);
    V3 s(r.scale[0], r.scale[1], r.scale[2]);
    grad[0] = grad[0] * t.d.uv_scale[0] * s;
    grad[1] = grad[1] * t.d.uv_scale[1] * s;
}

/* ---- MicrofacetDistribution (src/bsdfs/microfacet.h) ---- */
struct Microfacet {
    bool ggx, sampleVisible;
    float alphaU, alphaV;
    Microfacet(bool ggx_, float au, float av, bool sv) : ggx(ggx_), sampleVisible(sv) {
        alphaU = std::max(au, 1e-4f); alphaV = std::max(av, 1e-4f); /* :70-71 */
    }
    bool isIsotropic() const { return alphaU == alphaV; }
    void scaleAlpha(float v) { alphaU *= v; alphaV *= v; } /* :178-180 */
    /* :191-234 */
    float eval(V3 m) const {
        if (Frame::cosTheta(m) <= 0) return 0.0f;
        float cosTheta2 = Frame::cosTheta2(m);
        float beckmannExponent = ((m.x * m.x) / (alphaU * alphaU) + (m.y * m.y) / (alphaV * alphaV)) / cosTheta2;
        float result;
        if (!ggx) {
            result = fastexp(-beckmannExponent) / (kPi * alphaU * alphaV * cosTheta2 * cosTheta2);
        } else {
            float root = (1.0f + beckmannExponent) * cosTheta2;
            result = 1.0f / (kPi * alphaU * alphaV * root * root);
        }
        if (result * Frame::cosTheta(m) < 1e-20f) result = 0;
        return result;
    }
    /* :541-551 */
    float projectRoughness(V3 v) const {
        float invSinTheta2 = 1 / Frame::sinTheta2(v);
        if (isIsotropic() || invSinTheta2 <= 0) return alphaU;
        float cosPhi2 = v.x * v.x * invSinTheta2, sinPhi2 = v.y * v.y * invSinTheta2;
        return std::sqrt(cosPhi2 * alphaU * alphaU + sinPhi2 * alphaV * alphaV);
    }
    /* :477-514 */
    float smithG1(V3 v, V3 m) const {
        if (dot(v, m) * Frame::cosTheta(v) <= 0) return 0.0f;
        float tanTheta = std::fabs(Frame::tanTheta(v));
        if (tanTheta == 0.0f) return 1.0f;
        float alpha = projectRoughness(v);
        if (!ggx) {
            float a = 1.0f / (alpha * tanTheta);
            if (a >= 1.6f) return 1.0f;
            float aSqr = a * a;
            return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
        } else {
            float root = alpha * tanTheta;
            return 2.0f / (1.0f + hypot2(1.0f, root));
        }
    }
    float G(V3 wi, V3 wo, V3 m) const { return smithG1(wi, m) * smithG1(wo, m); }
    /* :404-407 */
    float pdfAll(V3 m) const { return eval(m) * Frame::cosTheta(m); }
    /* :462-466 */
    float pdfVisible(V3 wi, V3 m) const {
        if (Frame::cosTheta(wi) == 0) return 0.0f;
        return smithG1(wi, m) * absDot(wi, m) * eval(m) / std::fabs(Frame::cosTheta(wi));
    }
    float pdf(V3 wi, V3 m) const { return sampleVisible ? pdfVisible(wi, m) : pdfAll(m); }
    /* :287-395 (Beckmann and GGX branches) */
    V3 sampleAll(V2 sample, float &pdf) const {
        float cosThetaM, sinPhiM, cosPhiM, alphaSqr;
        if (isIsotropic()) {
            sincosf((2.0f * kPi) * sample.y, &sinPhiM, &cosPhiM);
            alphaSqr = alphaU * alphaU;
        } else {
            float phiM = std::atan(alphaV / alphaU * std::tan(kPi + 2 * kPi * sample.y)) + kPi * std::floor(2 * sample.y + 0.5f);
            sincosf(phiM, &sinPhiM, &cosPhiM);
            float cosSc = cosPhiM / alphaU, sinSc = sinPhiM / alphaV;
            alphaSqr = 1.0f / (cosSc * cosSc + sinSc * sinSc);
        }
        if (!ggx) {
            float tanThetaMSqr = alphaSqr * -fastlog(1.0f - sample.x);
            cosThetaM = 1.0f / std::sqrt(1.0f + tanThetaMSqr);
            pdf = (1.0f - sample.x) / (kPi * alphaU * alphaV * cosThetaM * cosThetaM * cosThetaM);
        } else {
            float tanThetaMSqr = alphaSqr * sample.x / (1.0f - sample.x);
            cosThetaM = 1.0f / std::sqrt(1.0f + tanThetaMSqr);
            float temp = 1 + tanThetaMSqr / alphaSqr;
            pdf = kInvPi / (alphaU * alphaV * cosThetaM * cosThetaM * cosThetaM * temp * temp);
        }
        if (pdf < 1e-20f) pdf = 0;
        float sinThetaM = std::sqrt(std::max(0.0f, 1 - cosThetaM * cosThetaM));
        return V3(sinThetaM * cosPhiM, sinThetaM * sinPhiM, cosThetaM);
    }
    /* :573-697 */
    V2 sampleVisible11(float thetaI, V2 sample) const {
        const float SQRT_PI_INV = 1 / std::sqrt(kPi);
        V2 slope;
        if (!ggx) {
            if (thetaI < 1e-4f) {
                float sinPhi, cosPhi;
                float r = std::sqrt(-fastlog(1.0f - sample.x));
                sincosf(2 * kPi * sample.y, &sinPhi, &cosPhi);
                return V2{r * cosPhi, r * sinPhi};
            }
            float tanThetaI = std::tan(thetaI), cotThetaI = 1 / tanThetaI;
            float a = -1, c = erf_as(cotThetaI);
            float sample_x = std::max(sample.x, 1e-6f);
            float fit = 1 + thetaI * (-0.876f + thetaI * (0.4265f - 0.0594f * thetaI));
            float b = c - (1 + c) * std::pow(1 - sample_x, fit);
            float normalization = 1 / (1 + c + SQRT_PI_INV * tanThetaI * std::exp(-cotThetaI * cotThetaI));
            int it = 0;
            while (++it < 10) {
                if (!(b >= a && b <= c)) b = 0.5f * (a + c);
                float invErf = erfinv(b);
                float value = normalization * (1 + b + SQRT_PI_INV * tanThetaI * std::exp(-invErf * invErf)) - sample_x;
                float derivative = normalization * (1 - invErf * tanThetaI);
                if (std::fabs(value) < 1e-5f) break;
                if (value > 0) c = b; else a = b;
                b -= value / derivative;
            }
            slope.x = erfinv(b);
            slope.y = erfinv(2.0f * std::max(sample.y, 1e-6f) - 1.0f);
        } else {
            if (thetaI < 1e-4f) {
                float sinPhi, cosPhi;
                float r = safe_sqrt(sample.x / (1 - sample.x));
                sincosf(2 * kPi * sample.y, &sinPhi, &cosPhi);
                return V2{r * cosPhi, r * sinPhi};
            }
            float tanThetaI = std::tan(thetaI);
            float a = 1 / tanThetaI;
            float G1 = 2.0f / (1.0f + safe_sqrt(1.0f + 1.0f / (a * a)));
            float A = 2.0f * sample.x / G1 - 1.0f;
            if (std::fabs(A) == 1) A -= signum(A) * kEps;
            float tmp = 1.0f / (A * A - 1.0f);
            float B = tanThetaI;
            float D = safe_sqrt(B * B * tmp * tmp - (A * A - B * B) * tmp);
            float slope_x_1 = B * tmp - D, slope_x_2 = B * tmp + D;
            slope.x = (A < 0.0f || slope_x_2 > 1.0f / tanThetaI) ? slope_x_1 : slope_x_2;
            float S;
            if (sample.y > 0.5f) { S = 1.0f; sample.y = 2.0f * (sample.y - 0.5f); }
            else { S = -1.0f; sample.y = 2.0f * (0.5f - sample.y); }
            float z = (sample.y * (sample.y * (sample.y * (-0.365728915865723f) + 0.790235037209296f) - 0.424965825137544f) + 0.000152998850436920f) /
                      (sample.y * (sample.y * (sample.y * (sample.y * 0.169507819808272f - 0.397203533833404f) - 0.232500544458471f) + 1.0f) - 0.539825872510702f);
            slope.y = S * z * std::sqrt(1.0f + slope.x * slope.x);
        }
        return slope;
    }
    /* :421-459 */
    V3 sampleVisibleNormal(V3 _wi, V2 sample) const {
        V3 wi = normalize(V3(alphaU * _wi.x, alphaV * _wi.y, _wi.z));
        float theta = 0, phi = 0;
        if (wi.z < 0.99999f) { theta = std::acos(wi.z); phi = std::atan2(wi.y, wi.x); }
        float sinPhi, cosPhi; sincosf(phi, &sinPhi, &cosPhi);
        V2 slope = sampleVisible11(theta, sample);
        slope = V2{cosPhi * slope.x - sinPhi * slope.y, sinPhi * slope.x + cosPhi * slope.y};
        slope.x *= alphaU; slope.y *= alphaV;
        float normalization = 1.0f / std::sqrt(slope.x * slope.x + slope.y * slope.y + 1.0f);
        return V3(-slope.x * normalization, -slope.y * normalization, normalization);
    }
    /* :243-256 */
    V3 sample(V3 wi, V2 s, float &pdf) const {
        if (sampleVisible) { V3 m = sampleVisibleNormal(wi, s); pdf = pdfVisible(wi, m); return m; }
        return sampleAll(s, pdf);
    }
};

/* ---- evalCubicInterp1D/2D (src/libcore/spline.cpp:23-58, 236-297), min=0,max=1 ---- */
inline float cubicInterp1D(float x, const float *values, int size) {
    if (!(x >= 0.0f && x <= 1.0f)) return 0.0f;
    float t = (x * (size - 1));
    int k = std::max(0, std::min((int)t, size - 2));
    float f0 = values[k], f1 = values[k + 1], d0, d1;
    if (k > 0) d0 = 0.5f * (values[k + 1] - values[k - 1]); else d0 = values[k + 1] - values[k];
    if (k + 2 < size) d1 = 0.5f * (values[k + 2] - values[k]); else d1 = values[k + 1] - values[k];
    t = t - (float)k;
    float t2 = t * t, t3 = t2 * t;
    return (2 * t3 - 3 * t2 + 1) * f0 + (-2 * t3 + 3 * t2) * f1 + (t3 - 2 * t2 + t) * d0 + (t3 - t2) * d1;
}
inline float cubicInterp2D(float px, float py, const float *values, int sx, int sy) {
    float w[2][4]; int knot[2];
    const float p[2] = {px, py}; const int size[2] = {sx, sy};
    for (int dim = 0; dim < 2; ++dim) {
        float *weights = w[dim];
        if (!(p[dim] >= 0.0f && p[dim] <= 1.0f)) return 0.0f;
        float t = p[dim] * (size[dim] - 1);
        knot[dim] = std::min((int)t, size[dim] - 2);
        t = t - (float)knot[dim];
        float t2 = t * t, t3 = t2 * t;
        weights[0] = 0.0f; weights[1] = 2 * t3 - 3 * t2 + 1; weights[2] = -2 * t3 + 3 * t2; weights[3] = 0.0f;
        float d0 = t3 - 2 * t2 + t, d1 = t3 - t2;
        if (knot[dim] > 0) { weights[2] += 0.5f * d0; weights[0] -= 0.5f * d0; }
        else { weights[2] += d0; weights[1] -= d0; }
        if (knot[dim] + 2 < size[dim]) { weights[3] += 0.5f * d1; weights[1] -= 0.5f * d1; }
        else { weights[2] += d1; weights[1] -= d1; }
    }
    float result = 0.0f;
    for (int y = -1; y <= 2; ++y) {
        float wy = w[1][y + 1];
        for (int x = -1; x <= 2; ++x) {
            float wxy = w[0][x + 1] * wy;
            if (wxy == 0) continue;
            result += values[(knot[1] + y) * sx + knot[0] + x] * wxy;
        }
    }
    return result;
}

/* RoughTransmittance::eval for the reduced tables (src/bsdfs/rtrans.h:183-233):
 * 1-D slice (eta and alpha fixed) or 2-D slice (eta fixed). */
inline float roughTransmittance(const Materials &m, const MtsBsdfNode &n, float cosTheta, float alpha) {
    float warpedCosTheta = std::pow(std::fabs(cosTheta), 0.25f), result;
    if (!(cosTheta >= 0)) return 0.f;
    const float *tab = &m.pool[n.lut[0]];
    if (!(n.flags & MTS_BF_RT_2D)) {
        result = cubicInterp1D(warpedCosTheta, tab, n.lut[1]);
    } else {
        float warpedAlpha = std::pow((alpha - n.f[0]) / (n.f[1] - n.f[0]), 0.25f);
        result = cubicInterp2D(warpedCosTheta, warpedAlpha, tab, n.lut[1], n.lut[2]);
    }
    return std::min(1.0f, std::max(0.0f, result));
}
/* 1 - m_internalRoughTransmittance->evalDiffuse(alpha) (roughplastic.cpp:366, rtrans.h:249-285) */
inline float roughInternalFdr(const Materials &m, const MtsBsdfNode &n, float alpha) {
    if (!(n.flags & MTS_BF_RT_2D)) return n.f[3];
    float warpedAlpha = std::pow((alpha - n.f[0]) / (n.f[1] - n.f[0]), 0.25f);
    float r = cubicInterp1D(warpedAlpha, &m.pool[n.lut[3]], n.lut[2]);
    return 1.0f - std::min(1.0f, std::max(0.0f, r));
}

inline V3 reflectZ(V3 wi) { return V3(-wi.x, -wi.y, wi.z); }
inline V3 reflectM(V3 wi, V3 m) { return 2 * dot(wi, m) * m - wi; }
/* BSDF::getMeasure (include/mitsuba/render/bsdf.h) */
inline int measureOf(uint32_t sampledType) {
    return (sampledType & MTS_ESMOOTH) ? MTS_MEASURE_SOLID_ANGLE : MTS_MEASURE_DISCRETE;
}

struct BSample { V3 wo; float pdf; float eta; uint32_t type; float u3 = 0; }; /* u3: input, bRec.sampler->next1D() of EUsesSampler models */
inline MtsTexRef roughDielTransmittance(const MtsBsdfNode &n) { MtsTexRef r; r.scale[0] = n.k[0]; r.scale[1] = n.k[1]; r.scale[2] = n.k[2]; r.tex = n.lut[0]; return r; }

V3 bsdfEval(const Materials &M, int node, const Its &its, V3 wi, V3 wo, int measure);
float bsdfPdf(const Materials &M, int node, const Its &its, V3 wi, V3 wo, int measure);
V3 bsdfSample(const Materials &M, int node, const Its &its, V3 wi, V2 u, BSample &out);

/* BumpMap::getFrame (src/bsdfs/bumpmap.cpp:135-159) */
inline Frame bumpFrame(const Materials &M, const MtsBsdfNode &n, const Its &its) {
    V3 grad[2];
    evalTexGradient(M, n.t0, its, grad);
    float dDispDu = luminance(grad[0]), dDispDv = luminance(grad[1]);
    V3 dpdu = its.dpdu + its.sh.n * (dDispDu - dot(its.sh.n, its.dpdu));
    V3 dpdv = its.dpdv + its.sh.n * (dDispDv - dot(its.sh.n, its.dpdv));
    Frame r;
    r.n = normalize(cross(dpdu, dpdv));
    r.s = normalize(dpdu - r.n * dot(r.n, dpdu));
    r.t = cross(r.n, r.s);
    if (dot(r.n, its.geo.n) < 0) r.n *= -1;
    return r;
}

/* NormalMap::getFrame (src/bsdfs/normalmap.cpp:106-123); the texture is read unfiltered (`eval(its, false)`) */
inline Frame normalMapFrame(const Materials &M, const MtsBsdfNode &n, const Its &its) {
    Its plain(its);
    plain.hasUVPartials = false;
    V3 v = evalTex(M, n.t0, plain);
    V3 nl(2 * v.x - 1, 2 * v.y - 1, 2 * v.z - 1);
    Frame r;
    r.n = normalize(its.sh.toWorld(nl));
    r.s = normalize(its.dpdu - r.n * dot(r.n, its.dpdu));
    r.t = cross(r.n, r.s);
    return r;
}
inline Frame perturbedFrame(const Materials &M, const MtsBsdfNode &n, const Its &its) {
    return n.type == MTS_BSDF_NORMALMAP ? normalMapFrame(M, n, its) : bumpFrame(M, n, its);
}

/* RoughCoating::refractTo (src/bsdfs/roughcoating.cpp:232-253); etaTo = 1/eta towards the interior, eta towards the exterior */
inline V3 roughCoatRefract(float etaTo, V3 wi) {
    float sinThetaTSqr = etaTo * etaTo * Frame::sinTheta2(wi);
    if (sinThetaTSqr >= 1.0f) return V3(0.0f);
    float cosThetaT = std::sqrt(1.0f - sinThetaTSqr);
    return V3(etaTo * wi.x, etaTo * wi.y, Frame::cosTheta(wi) > 0.0f ? cosThetaT : -cosThetaT);
}
/* roughcoating.cpp:332-343, 388-397 */
inline float roughCoatProbSpecular(const Materials &M, const MtsBsdfNode &n, V3 wi, float alpha) {
    float probSpecular = 1 - roughTransmittance(M, n, std::fabs(Frame::cosTheta(wi)), alpha);
    float w = n.f[2];
    return (probSpecular * w) / (probSpecular * w + (1 - probSpecular) * (1 - w));
}
/* MixtureBSDF tables in the float pool (include/mtscuda.h): N child indices, N weights, N + 1 CDF entries */
struct MixtureTab {
    const float *p; int n;
    MixtureTab(const Materials &M, const MtsBsdfNode &nd) : p(M.pool.data() + nd.lut[0]), n(nd.lut[1]) {}
    int child(int i) const { return (int)p[i]; }
    float weight(int i) const { return p[n + i]; }
    float cdf(int i) const { return p[2 * n + i]; }
    float pdf(int i) const { return cdf(i + 1) - cdf(i); }
    /* DiscreteDistribution::sampleReuse (include/mitsuba/core/pmf.h:124-136, 164-169) */
    int sampleReuse(float &u) const {
        const float *b = p + 2 * n, *e = b + n + 1;
        const float *entry = std::lower_bound(b, e, u);
        int index = (int)std::min((ptrdiff_t)(n - 1), std::max((ptrdiff_t)0, entry - b - 1));
        while (pdf(index) == 0 && index < n) ++index;
        u = (u - cdf(index)) / (cdf(index + 1) - cdf(index));
        return index;
    }
};

/* SmoothCoating::refractIn/refractOut (src/bsdfs/coating.cpp:208-219) */
inline V3 coatRefractIn(float eta, V3 wi, float &R) {
    float cosThetaT; float invEta = 1 / eta;
    R = fresnelDielectricExt(std::fabs(Frame::cosTheta(wi)), cosThetaT, eta);
    return V3(invEta * wi.x, invEta * wi.y, -signum(Frame::cosTheta(wi)) * cosThetaT);
}
inline V3 coatRefractOut(float eta, V3 wi, float &R) {
    float cosThetaT; float invEta = 1 / eta;
    R = fresnelDielectricExt(std::fabs(Frame::cosTheta(wi)), cosThetaT, invEta);
    return V3(eta * wi.x, eta * wi.y, -signum(Frame::cosTheta(wi)) * cosThetaT);
}

inline V3 bsdfEval(const Materials &M, int node, const Its &its, V3 wi, V3 wo, int measure) {
    const MtsBsdfNode &n = M.nodes[node];
    const bool solid = measure == MTS_MEASURE_SOLID_ANGLE, discrete = measure == MTS_MEASURE_DISCRETE;
    switch (n.type) {
    case MTS_BSDF_DIFFUSE: { /* diffuse.cpp:110-118 */
        if (!solid || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0) return V3(0.0f);
        return evalTex(M, n.t0, its) * (kInvPi * Frame::cosTheta(wo));
    }
    case MTS_BSDF_CONDUCTOR: { /* conductor.cpp:223-237 */
        if (!discrete || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0 ||
            std::fabs(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return V3(0.0f);
        return evalTex(M, n.t0, its) * fresnelConductorExact(Frame::cosTheta(wi), V3(n.eta[0], n.eta[1], n.eta[2]), V3(n.k[0], n.k[1], n.k[2]));
    }
    case MTS_BSDF_ROUGHCONDUCTOR: { /* roughconductor.cpp:257-294 */
        if (!solid || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0) return V3(0.0f);
        V3 H = normalize(wo + wi);
        Microfacet distr(n.flags & MTS_BF_GGX, average(evalTex(M, n.t1, its)), average(evalTex(M, n.t2, its)), n.flags & MTS_BF_SAMPLE_VISIBLE);
        const float D = distr.eval(H);
        if (D == 0) return V3(0.0f);
        const V3 F = fresnelConductorExact(dot(wi, H), V3(n.eta[0], n.eta[1], n.eta[2]), V3(n.k[0], n.k[1], n.k[2])) * evalTex(M, n.t0, its);
        const float G = distr.G(wi, wo, H);
        float model = D * G / (4.0f * Frame::cosTheta(wi));
        return F * model;
    }
    case MTS_BSDF_PLASTIC: { /* plastic.cpp:245-278 */
        if (Frame::cosTheta(wo) <= 0 || Frame::cosTheta(wi) <= 0) return V3(0.0f);
        float eta = n.eta[0];
        float Fi = fresnelDielectricExt(Frame::cosTheta(wi), eta);
        if (discrete) {
            if (std::fabs(dot(reflectZ(wi), wo) - 1) < kDeltaEps) return evalTex(M, n.t1, its) * Fi;
        } else if (solid) {
            float Fo = fresnelDielectricExt(Frame::cosTheta(wo), eta);
            V3 diff = evalTex(M, n.t0, its);
            if (n.flags & MTS_BF_NONLINEAR) diff = diff / (V3(1.0f) - diff * n.f[0]);
            else diff = diff / (1 - n.f[0]);
            float invEta2 = 1 / (eta * eta);
            return diff * (squareToCosineHemispherePdf(wo) * invEta2 * (1 - Fi) * (1 - Fo));
        }
        return V3(0.0f);
    }
    case MTS_BSDF_ROUGHPLASTIC: { /* roughplastic.cpp:326-382 */
        if (!solid || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0) return V3(0.0f);
        float eta = n.eta[0];
        float a = average(evalTex(M, n.t2, its));
        Microfacet distr(n.flags & MTS_BF_GGX, a, a, n.flags & MTS_BF_SAMPLE_VISIBLE);
        V3 result(0.0f);
        {
            const V3 H = normalize(wo + wi);
            const float D = distr.eval(H);
            const float F = fresnelDielectricExt(dot(wi, H), eta);
            const float G = distr.G(wi, wo, H);
            float value = F * D * G / (4.0f * Frame::cosTheta(wi));
            result += evalTex(M, n.t1, its) * value;
        }
        {
            V3 diff = evalTex(M, n.t0, its);
            float T12 = roughTransmittance(M, n, Frame::cosTheta(wi), distr.alphaU);
            float T21 = roughTransmittance(M, n, Frame::cosTheta(wo), distr.alphaU);
            float Fdr = roughInternalFdr(M, n, distr.alphaU);
            if (n.flags & MTS_BF_NONLINEAR) diff = diff / (V3(1.0f) - diff * Fdr);
            else diff = diff / (1 - Fdr);
            float invEta2 = 1.0f / (eta * eta);
            result += diff * (kInvPi * Frame::cosTheta(wo) * T12 * T21 * invEta2);
        }
        return result;
    }
    case MTS_BSDF_DIELECTRIC: { /* dielectric.cpp:228-253 */
        float eta = n.eta[0], invEta = 1 / eta;
        float cosThetaT;
        float F = fresnelDielectricExt(Frame::cosTheta(wi), cosThetaT, eta);
        if (Frame::cosTheta(wi) * Frame::cosTheta(wo) >= 0) {
            if (!discrete || std::fabs(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return V3(0.0f);
            return evalTex(M, n.t0, its) * F;
        } else {
            float scale = -(cosThetaT < 0 ? invEta : eta);
            V3 refr(scale * wi.x, scale * wi.y, cosThetaT);
            if (!discrete || std::fabs(dot(refr, wo) - 1) > kDeltaEps) return V3(0.0f);
            float factor = cosThetaT < 0 ? invEta : eta; /* ERadiance */
            return evalTex(M, n.t1, its) * factor * factor * (1 - F);
        }
    }
    case MTS_BSDF_ROUGHDIELECTRIC: { /* roughdielectric.cpp:272-348 */
        if (!solid || Frame::cosTheta(wi) == 0) return V3(0.0f);
        const float eta = n.eta[0], invEta = 1 / eta;
        const bool reflect = Frame::cosTheta(wi) * Frame::cosTheta(wo) > 0;
        V3 H;
        if (reflect) H = normalize(wo + wi);
        else H = normalize(wi + wo * (Frame::cosTheta(wi) > 0 ? eta : invEta));
        H = H * signum(Frame::cosTheta(H));
        Microfacet distr(n.flags & MTS_BF_GGX, average(evalTex(M, n.t1, its)), average(evalTex(M, n.t2, its)), n.flags & MTS_BF_SAMPLE_VISIBLE);
        const float D = distr.eval(H);
        if (D == 0) return V3(0.0f);
        const float F = fresnelDielectricExt(dot(wi, H), eta);
        const float G = distr.G(wi, wo, H);
        if (reflect) return evalTex(M, n.t0, its) * (F * D * G / (4.0f * std::fabs(Frame::cosTheta(wi))));
        const float e = Frame::cosTheta(wi) > 0.0f ? eta : invEta;
        const float sqrtDenom = dot(wi, H) + e * dot(wo, H);
        const float value = ((1 - F) * D * G * e * e * dot(wi, H) * dot(wo, H)) / (Frame::cosTheta(wi) * sqrtDenom * sqrtDenom);
        const float factor = Frame::cosTheta(wi) > 0 ? invEta : eta; /* ERadiance */
        return evalTex(M, roughDielTransmittance(n), its) * std::fabs(value * factor * factor);
    }
    case MTS_BSDF_THINDIELECTRIC: { /* thindielectric.cpp:145-166 */
        float R = fresnelDielectricExt(std::fabs(Frame::cosTheta(wi)), n.eta[0]), T = 1 - R;
        if (R < 1) R += T * T * R / (1 - R * R);
        if (Frame::cosTheta(wi) * Frame::cosTheta(wo) >= 0) {
            if (!discrete || std::fabs(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return V3(0.0f);
            return evalTex(M, n.t0, its) * R;
        } else {
            if (!discrete || std::fabs(dot(V3(-wi.x, -wi.y, -wi.z), wo) - 1) > kDeltaEps) return V3(0.0f);
            return evalTex(M, n.t1, its) * (1 - R);
        }
    }
    case MTS_BSDF_COATING: { /* coating.cpp:221-258 */
        float eta = n.eta[0], invEta = 1 / eta;
        if (discrete && std::fabs(dot(reflectZ(wi), wo) - 1) < kDeltaEps) {
            return evalTex(M, n.t1, its) * fresnelDielectricExt(std::fabs(Frame::cosTheta(wi)), eta);
        } else {
            float R12, R21;
            V3 wiP = coatRefractIn(eta, wi, R12), woP = coatRefractIn(eta, wo, R21);
            if (R12 == 1 || R21 == 1) return V3(0.0f);
            V3 result = bsdfEval(M, n.child[0], its, wiP, woP, measure) * (1 - R12) * (1 - R21);
            V3 sigmaA = evalTex(M, n.t0, its) * n.f[0];
            if (!isZero(sigmaA))
                result *= vexp(-sigmaA * (1 / std::fabs(Frame::cosTheta(wiP)) + 1 / std::fabs(Frame::cosTheta(woP))));
            if (solid) result *= invEta * invEta * Frame::cosTheta(wo) / Frame::cosTheta(woP);
            return result;
        }
    }
    case MTS_BSDF_BLEND: { /* blendbsdf.cpp:135-152 */
        float weight = std::min(1.0f, std::max(0.0f, average(evalTex(M, n.t0, its))));
        return bsdfEval(M, n.child[0], its, wi, wo, measure) * (1 - weight) + bsdfEval(M, n.child[1], its, wi, wo, measure) * weight;
    }
    case MTS_BSDF_MIXTURE: { /* mixturebsdf.cpp:174-189 */
        MixtureTab mt(M, n);
        V3 result(0.0f);
        for (int i = 0; i < mt.n; ++i) result += bsdfEval(M, mt.child(i), its, wi, wo, measure) * mt.weight(i);
        return result;
    }
    case MTS_BSDF_ROUGHCOATING: { /* roughcoating.cpp:255-314 */
        float eta = n.eta[0], invEta = 1 / eta;
        bool hasNested = (M.nodes[n.child[0]].lobes & MTS_EALL) != 0, hasSpecular = solid;
        float alpha = average(evalTex(M, n.t2, its));
        Microfacet distr(n.flags & MTS_BF_GGX, alpha, alpha, n.flags & MTS_BF_SAMPLE_VISIBLE);
        V3 result(0.0f);
        if (hasSpecular && Frame::cosTheta(wo) * Frame::cosTheta(wi) > 0) {
            V3 H = normalize(wo + wi) * signum(Frame::cosTheta(wo));
            float D = distr.eval(H);
            float F = fresnelDielectricExt(absDot(wi, H), eta);
            float G = distr.G(wi, wo, H);
            float value = F * D * G / (4.0f * std::fabs(Frame::cosTheta(wi)));
            result += evalTex(M, n.t1, its) * value;
        }
        if (hasNested) {
            V3 wiI = roughCoatRefract(invEta, wi), woI = roughCoatRefract(invEta, wo);
            V3 nested = bsdfEval(M, n.child[0], its, wiI, woI, measure) *
                roughTransmittance(M, n, std::fabs(Frame::cosTheta(wi)), distr.alphaU) *
                roughTransmittance(M, n, std::fabs(Frame::cosTheta(wo)), distr.alphaU);
            V3 sigmaA = evalTex(M, n.t0, its) * n.f[3];
            if (!isZero(sigmaA))
                nested *= vexp(-sigmaA * (1 / std::fabs(Frame::cosTheta(wiI)) + 1 / std::fabs(Frame::cosTheta(woI))));
            if (solid) nested *= invEta * invEta * Frame::cosTheta(wo) / Frame::cosTheta(woI);
            result += nested;
        }
        return result;
    }
    case MTS_BSDF_NORMALMAP: /* normalmap.cpp:162-179 */
    case MTS_BSDF_BUMPMAP: { /* bumpmap.cpp:161-177 */
        Its pert(its);
        pert.sh = perturbedFrame(M, n, its);
        V3 wiP = pert.toLocal(its.toWorld(wi)), woP = pert.toLocal(its.toWorld(wo));
        if (Frame::cosTheta(wo) * Frame::cosTheta(woP) <= 0) return V3(0.0f);
        return bsdfEval(M, n.child[0], pert, wiP, woP, measure);
    }
    case MTS_BSDF_MASK: { /* mask.cpp:106-116 */
        V3 opacity = evalTex(M, n.t0, its);
        if (measure == MTS_MEASURE_SOLID_ANGLE) return bsdfEval(M, n.child[0], its, wi, wo, measure) * opacity;
        if (measure == MTS_MEASURE_DISCRETE && std::fabs(1 - dot(wi, V3(-wo.x, -wo.y, -wo.z))) < kDeltaEps) return V3(1.0f) - opacity;
        return V3(0.0f);
    }
    case MTS_BSDF_TWOSIDED: { /* twosided.cpp:107-119 */
        if (Frame::cosTheta(wi) > 0) return bsdfEval(M, n.child[0], its, wi, wo, measure);
        return bsdfEval(M, n.child[1], its, V3(wi.x, wi.y, -wi.z), V3(wo.x, wo.y, -wo.z), measure);
    }
    }
    return V3(0.0f);
}

inline float bsdfPdf(const Materials &M, int node, const Its &its, V3 wi, V3 wo, int measure) {
    const MtsBsdfNode &n = M.nodes[node];
    const bool solid = measure == MTS_MEASURE_SOLID_ANGLE, discrete = measure == MTS_MEASURE_DISCRETE;
    switch (n.type) {
    case MTS_BSDF_DIFFUSE: /* diffuse.cpp:120-127 */
        if (!solid || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0) return 0.0f;
        return squareToCosineHemispherePdf(wo);
    case MTS_BSDF_CONDUCTOR: /* conductor.cpp:239-252 */
        if (!discrete || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0 ||
            std::fabs(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return 0.0f;
        return 1.0f;
    case MTS_BSDF_ROUGHCONDUCTOR: { /* roughconductor.cpp:296-321 */
        if (!solid || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0) return 0.0f;
        V3 H = normalize(wo + wi);
        Microfacet distr(n.flags & MTS_BF_GGX, average(evalTex(M, n.t1, its)), average(evalTex(M, n.t2, its)), n.flags & MTS_BF_SAMPLE_VISIBLE);
        if (distr.sampleVisible) return distr.eval(H) * distr.smithG1(wi, H) / (4.0f * Frame::cosTheta(wi));
        return distr.pdf(wi, H) / (4 * absDot(wo, H));
    }
    case MTS_BSDF_PLASTIC: { /* plastic.cpp:280-307 */
        if (Frame::cosTheta(wo) <= 0 || Frame::cosTheta(wi) <= 0) return 0.0f;
        float Fi = fresnelDielectricExt(Frame::cosTheta(wi), n.eta[0]);
        float w = n.f[2];
        float probSpecular = (Fi * w) / (Fi * w + (1 - Fi) * (1 - w));
        if (discrete) {
            if (std::fabs(dot(reflectZ(wi), wo) - 1) < kDeltaEps) return probSpecular;
        } else if (solid) {
            return squareToCosineHemispherePdf(wo) * (1 - probSpecular);
        }
        return 0.0f;
    }
    case MTS_BSDF_ROUGHPLASTIC: { /* roughplastic.cpp:384-437 */
        if (!solid || Frame::cosTheta(wi) <= 0 || Frame::cosTheta(wo) <= 0) return 0.0f;
        float a = average(evalTex(M, n.t2, its));
        Microfacet distr(n.flags & MTS_BF_GGX, a, a, n.flags & MTS_BF_SAMPLE_VISIBLE);
        const V3 H = normalize(wo + wi);
        float w = n.f[2];
        float probSpecular = 1 - roughTransmittance(M, n, Frame::cosTheta(wi), distr.alphaU);
        probSpecular = (probSpecular * w) / (probSpecular * w + (1 - probSpecular) * (1 - w));
        float probDiffuse = 1 - probSpecular;
        const float dwh_dwo = 1.0f / (4.0f * dot(wo, H));
        const float prob = distr.pdf(wi, H);
        float result = prob * dwh_dwo * probSpecular;
        result += probDiffuse * squareToCosineHemispherePdf(wo);
        return result;
    }
    case MTS_BSDF_DIELECTRIC: { /* dielectric.cpp:255-275 */
        float eta = n.eta[0], invEta = 1 / eta;
        float cosThetaT;
        float F = fresnelDielectricExt(Frame::cosTheta(wi), cosThetaT, eta);
        if (Frame::cosTheta(wi) * Frame::cosTheta(wo) >= 0) {
            if (!discrete || std::fabs(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return 0.0f;
            return F;
        } else {
            float scale = -(cosThetaT < 0 ? invEta : eta);
            V3 refr(scale * wi.x, scale * wi.y, cosThetaT);
            if (!discrete || std::fabs(dot(refr, wo) - 1) > kDeltaEps) return 0.0f;
            return 1 - F;
        }
    }
    case MTS_BSDF_ROUGHDIELECTRIC: { /* roughdielectric.cpp:350-420 */
        if (!solid) return 0.0f;
        const float eta = n.eta[0], invEta = 1 / eta;
        const bool reflect = Frame::cosTheta(wi) * Frame::cosTheta(wo) > 0;
        V3 H; float dwh_dwo;
        if (reflect) {
            H = normalize(wo + wi);
            dwh_dwo = 1.0f / (4.0f * dot(wo, H));
        } else {
            const float e = Frame::cosTheta(wi) > 0 ? eta : invEta;
            H = normalize(wi + wo * e);
            const float sqrtDenom = dot(wi, H) + e * dot(wo, H);
            dwh_dwo = (e * e * dot(wo, H)) / (sqrtDenom * sqrtDenom);
        }
        H = H * signum(Frame::cosTheta(H));
        Microfacet sampleDistr(n.flags & MTS_BF_GGX, average(evalTex(M, n.t1, its)), average(evalTex(M, n.t2, its)), n.flags & MTS_BF_SAMPLE_VISIBLE);
        if (!sampleDistr.sampleVisible) sampleDistr.scaleAlpha(1.2f - 0.2f * std::sqrt(std::fabs(Frame::cosTheta(wi))));
        float prob = sampleDistr.pdf(wi * signum(Frame::cosTheta(wi)), H);
        const float F = fresnelDielectricExt(dot(wi, H), eta);
        prob *= reflect ? F : (1 - F);
        return std::fabs(prob * dwh_dwo);
    }
    case MTS_BSDF_THINDIELECTRIC: { /* thindielectric.cpp:168-189 */
        float R = fresnelDielectricExt(std::fabs(Frame::cosTheta(wi)), n.eta[0]), T = 1 - R;
        if (R < 1) R += T * T * R / (1 - R * R);
        if (Frame::cosTheta(wi) * Frame::cosTheta(wo) >= 0) {
            if (!discrete || std::fabs(dot(reflectZ(wi), wo) - 1) > kDeltaEps) return 0.0f;
            return R;
        } else {
            if (!discrete || std::fabs(dot(V3(-wi.x, -wi.y, -wi.z), wo) - 1) > kDeltaEps) return 0.0f;
            return 1 - R;
        }
    }
    case MTS_BSDF_COATING: { /* coating.cpp:260-296 */
        float eta = n.eta[0], invEta = 1 / eta;
        float R12;
        V3 wiP = coatRefractIn(eta, wi, R12);
        float w = n.f[2];
        float probSpecular = (R12 * w) / (R12 * w + (1 - R12) * (1 - w));
        if (discrete && std::fabs(dot(reflectZ(wi), wo) - 1) < kDeltaEps) {
            return probSpecular;
        } else {
            float R21;
            V3 woP = coatRefractIn(eta, wo, R21);
            if (R12 == 1 || R21 == 1) return 0.0f;
            float pdf = bsdfPdf(M, n.child[0], its, wiP, woP, measure);
            if (solid) pdf *= invEta * invEta * Frame::cosTheta(wo) / Frame::cosTheta(woP);
            return pdf * (1 - probSpecular);
        }
    }
    case MTS_BSDF_BLEND: { /* blendbsdf.cpp:154-173 */
        float weight = std::min(1.0f, std::max(0.0f, average(evalTex(M, n.t0, its))));
        return bsdfPdf(M, n.child[0], its, wi, wo, measure) * (1 - weight) + bsdfPdf(M, n.child[1], its, wi, wo, measure) * weight;
    }
    case MTS_BSDF_MIXTURE: { /* mixturebsdf.cpp:191-206 */
        MixtureTab mt(M, n);
        float result = 0.0f;
        for (int i = 0; i < mt.n; ++i) result += bsdfPdf(M, mt.child(i), its, wi, wo, measure) * mt.pdf(i);
        return result;
    }
    case MTS_BSDF_ROUGHCOATING: { /* roughcoating.cpp:316-376 */
        float eta = n.eta[0], invEta = 1 / eta;
        bool hasNested = (M.nodes[n.child[0]].lobes & MTS_EALL) != 0, hasSpecular = measure == MTS_MEASURE_SOLID_ANGLE;
        V3 H = normalize(wo + wi) * signum(Frame::cosTheta(wo));
        float alpha = average(evalTex(M, n.t2, its));
        Microfacet distr(n.flags & MTS_BF_GGX, alpha, alpha, n.flags & MTS_BF_SAMPLE_VISIBLE);
        float probNested = 1.0f, probSpecular = 1.0f;
        if (hasSpecular && hasNested) { probSpecular = roughCoatProbSpecular(M, n, wi, distr.alphaU); probNested = 1 - probSpecular; }
        float result = 0.0f;
        if (hasSpecular && Frame::cosTheta(wo) * Frame::cosTheta(wi) > 0) {
            float dwh_dwo = 1.0f / (4.0f * absDot(wo, H));
            result = distr.pdf(wi, H) * dwh_dwo * probSpecular;
        }
        if (hasNested) {
            V3 wiI = roughCoatRefract(invEta, wi), woI = roughCoatRefract(invEta, wo);
            float prob = bsdfPdf(M, n.child[0], its, wiI, woI, measure);
            if (measure == MTS_MEASURE_SOLID_ANGLE) prob *= invEta * invEta * Frame::cosTheta(wo) / Frame::cosTheta(woI);
            result += prob * probNested;
        }
        return result;
    }
    case MTS_BSDF_NORMALMAP: /* normalmap.cpp:181-197 */
    case MTS_BSDF_BUMPMAP: { /* bumpmap.cpp:179-197 */
        Its pert(its);
        pert.sh = perturbedFrame(M, n, its);
        V3 wiP = pert.toLocal(its.toWorld(wi)), woP = pert.toLocal(its.toWorld(wo));
        if (Frame::cosTheta(wo) * Frame::cosTheta(woP) <= 0) return 0;
        return bsdfPdf(M, n.child[0], pert, wiP, woP, measure);
    }
    case MTS_BSDF_MASK: { /* mask.cpp:118-143 */
        float prob = luminance(evalTex(M, n.t0, its));
        if (measure == MTS_MEASURE_SOLID_ANGLE) return bsdfPdf(M, n.child[0], its, wi, wo, measure) * prob;
        if (measure == MTS_MEASURE_DISCRETE && std::fabs(1 - dot(wi, V3(-wo.x, -wo.y, -wo.z))) < kDeltaEps) return 1 - prob;
        return 0.0f;
    }
    case MTS_BSDF_TWOSIDED: { /* twosided.cpp:121-133 */
        if (wi.z > 0) return bsdfPdf(M, n.child[0], its, wi, wo, measure);
        return bsdfPdf(M, n.child[1], its, V3(wi.x, wi.y, -wi.z), V3(wo.x, wo.y, -wo.z), measure);
    }
    }
    return 0.0f;
}

/* BSDF::sample(bRec, pdf, sample): returns the weight (f*cos/pdf), fills wo/pdf/eta/sampledType */
inline V3 bsdfSample(const Materials &M, int node, const Its &its, V3 wi, V2 sample, BSample &o) {
    const MtsBsdfNode &n = M.nodes[node];
    switch (n.type) {
    case MTS_BSDF_DIFFUSE: { /* diffuse.cpp:141-150 */
        if (Frame::cosTheta(wi) <= 0) return V3(0.0f);
        o.wo = squareToCosineHemisphere(sample);
        o.eta = 1.0f; o.type = MTS_EDIFFUSE_REFLECTION;
        o.pdf = squareToCosineHemispherePdf(o.wo);
        return evalTex(M, n.t0, its);
    }
    case MTS_BSDF_CONDUCTOR: { /* conductor.cpp:270-286 */
        if (Frame::cosTheta(wi) <= 0) return V3(0.0f);
        o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.eta = 1.0f; o.pdf = 1;
        return evalTex(M, n.t0, its) * fresnelConductorExact(Frame::cosTheta(wi), V3(n.eta[0], n.eta[1], n.eta[2]), V3(n.k[0], n.k[1], n.k[2]));
    }
    case MTS_BSDF_ROUGHCONDUCTOR: { /* roughconductor.cpp:368-410 */
        if (Frame::cosTheta(wi) < 0) return V3(0.0f);
        Microfacet distr(n.flags & MTS_BF_GGX, average(evalTex(M, n.t1, its)), average(evalTex(M, n.t2, its)), n.flags & MTS_BF_SAMPLE_VISIBLE);
        float pdf;
        V3 m = distr.sample(wi, sample, pdf);
        o.pdf = pdf;
        if (pdf == 0) return V3(0.0f);
        o.wo = reflectM(wi, m); o.eta = 1.0f; o.type = MTS_EGLOSSY_REFLECTION;
        if (Frame::cosTheta(o.wo) <= 0) return V3(0.0f);
        V3 F = fresnelConductorExact(dot(wi, m), V3(n.eta[0], n.eta[1], n.eta[2]), V3(n.k[0], n.k[1], n.k[2])) * evalTex(M, n.t0, its);
        float weight;
        if (distr.sampleVisible) weight = distr.smithG1(o.wo, m);
        else weight = distr.eval(m) * distr.G(wi, o.wo, m) * dot(wi, m) / (pdf * Frame::cosTheta(wi));
        o.pdf = pdf / (4.0f * dot(o.wo, m));
        return F * weight;
    }
    case MTS_BSDF_PLASTIC: { /* plastic.cpp:373-430 */
        if (Frame::cosTheta(wi) <= 0) return V3(0.0f);
        float eta = n.eta[0];
        float Fi = fresnelDielectricExt(Frame::cosTheta(wi), eta);
        o.eta = 1.0f;
        float w = n.f[2];
        float probSpecular = (Fi * w) / (Fi * w + (1 - Fi) * (1 - w));
        if (sample.x < probSpecular) {
            o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.pdf = probSpecular;
            return evalTex(M, n.t1, its) * Fi / probSpecular;
        } else {
            o.type = MTS_EDIFFUSE_REFLECTION;
            o.wo = squareToCosineHemisphere(V2{(sample.x - probSpecular) / (1 - probSpecular), sample.y});
            float Fo = fresnelDielectricExt(Frame::cosTheta(o.wo), eta);
            V3 diff = evalTex(M, n.t0, its);
            if (n.flags & MTS_BF_NONLINEAR) diff = diff / (V3(1.0f) - diff * n.f[0]);
            else diff = diff / (1 - n.f[0]);
            o.pdf = (1 - probSpecular) * squareToCosineHemispherePdf(o.wo);
            float invEta2 = 1 / (eta * eta);
            return diff * (invEta2 * (1 - Fi) * (1 - Fo) / (1 - probSpecular));
        }
    }
    case MTS_BSDF_ROUGHPLASTIC: { /* roughplastic.cpp:439-501 */
        if (Frame::cosTheta(wi) <= 0) return V3(0.0f);
        bool choseSpecular = true;
        float a = average(evalTex(M, n.t2, its));
        Microfacet distr(n.flags & MTS_BF_GGX, a, a, n.flags & MTS_BF_SAMPLE_VISIBLE);
        float w = n.f[2];
        float probSpecular = 1 - roughTransmittance(M, n, Frame::cosTheta(wi), distr.alphaU);
        probSpecular = (probSpecular * w) / (probSpecular * w + (1 - probSpecular) * (1 - w));
        if (sample.y < probSpecular) sample.y /= probSpecular;
        else { sample.y = (sample.y - probSpecular) / (1 - probSpecular); choseSpecular = false; }
        if (choseSpecular) {
            float dummy;
            V3 m = distr.sampleVisible ? distr.sampleVisibleNormal(wi, sample) : distr.sampleAll(sample, dummy);
            o.wo = reflectM(wi, m); o.type = MTS_EGLOSSY_REFLECTION;
            if (Frame::cosTheta(o.wo) <= 0) return V3(0.0f);
        } else {
            o.type = MTS_EDIFFUSE_REFLECTION;
            o.wo = squareToCosineHemisphere(sample);
        }
        o.eta = 1.0f;
        o.pdf = bsdfPdf(M, node, its, wi, o.wo, MTS_MEASURE_SOLID_ANGLE);
        if (o.pdf == 0) return V3(0.0f);
        return bsdfEval(M, node, its, wi, o.wo, MTS_MEASURE_SOLID_ANGLE) / o.pdf;
    }
    case MTS_BSDF_DIELECTRIC: { /* dielectric.cpp:277-299 */
        float eta = n.eta[0], invEta = 1 / eta;
        float cosThetaT;
        float F = fresnelDielectricExt(Frame::cosTheta(wi), cosThetaT, eta);
        if (sample.x <= F) {
            o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.eta = 1.0f; o.pdf = F;
            return evalTex(M, n.t0, its);
        } else {
            o.type = MTS_EDELTA_TRANSMISSION;
            float scale = -(cosThetaT < 0 ? invEta : eta);
            o.wo = V3(scale * wi.x, scale * wi.y, cosThetaT);
            o.eta = cosThetaT < 0 ? eta : invEta;
            o.pdf = 1 - F;
            float factor = cosThetaT < 0 ? invEta : eta;
            return evalTex(M, n.t1, its) * (factor * factor);
        }
    }
    case MTS_BSDF_ROUGHDIELECTRIC: { /* roughdielectric.cpp:517-616 (pdf-returning overload) */
        const float eta = n.eta[0], invEta = 1 / eta;
        Microfacet distr(n.flags & MTS_BF_GGX, average(evalTex(M, n.t1, its)), average(evalTex(M, n.t2, its)), n.flags & MTS_BF_SAMPLE_VISIBLE);
        Microfacet sampleDistr = distr;
        if (!distr.sampleVisible) sampleDistr.scaleAlpha(1.2f - 0.2f * std::sqrt(std::fabs(Frame::cosTheta(wi))));
        float microfacetPDF;
        const V3 m = sampleDistr.sample(wi * signum(Frame::cosTheta(wi)), sample, microfacetPDF);
        if (microfacetPDF == 0) return V3(0.0f);
        o.pdf = microfacetPDF;
        float cosThetaT;
        const float F = fresnelDielectricExt(dot(wi, m), cosThetaT, eta);
        V3 weight(1.0f);
        bool sampleReflection = true;
        if (o.u3 > F) { sampleReflection = false; o.pdf *= 1 - F; } else o.pdf *= F; /* bRec.sampler->next1D() */
        float dwh_dwo;
        if (sampleReflection) {
            o.wo = reflectM(wi, m); o.eta = 1.0f; o.type = MTS_EGLOSSY_REFLECTION;
            if (Frame::cosTheta(wi) * Frame::cosTheta(o.wo) <= 0) return V3(0.0f);
            weight *= evalTex(M, n.t0, its);
            dwh_dwo = 1.0f / (4.0f * dot(o.wo, m));
        } else {
            if (cosThetaT == 0) return V3(0.0f);
            const float e = cosThetaT < 0 ? invEta : eta; /* refract(wi, m, eta, cosThetaT), util.cpp:767-772 */
            o.wo = m * (dot(wi, m) * e + cosThetaT) - wi * e;
            o.eta = cosThetaT < 0 ? eta : invEta; o.type = MTS_EGLOSSY_TRANSMISSION;
            if (Frame::cosTheta(wi) * Frame::cosTheta(o.wo) >= 0) return V3(0.0f);
            const float factor = cosThetaT < 0 ? invEta : eta; /* ERadiance */
            weight *= evalTex(M, roughDielTransmittance(n), its) * (factor * factor);
            const float sqrtDenom = dot(wi, m) + o.eta * dot(o.wo, m);
            dwh_dwo = (o.eta * o.eta * dot(o.wo, m)) / (sqrtDenom * sqrtDenom);
        }
        if (distr.sampleVisible) weight *= distr.smithG1(o.wo, m);
        else weight *= std::fabs(distr.eval(m) * distr.G(wi, o.wo, m) * dot(wi, m) / (microfacetPDF * Frame::cosTheta(wi)));
        o.pdf *= std::fabs(dwh_dwo);
        return weight;
    }
    case MTS_BSDF_THINDIELECTRIC: { /* thindielectric.cpp:191-221 */
        float R = fresnelDielectricExt(std::fabs(Frame::cosTheta(wi)), n.eta[0]), T = 1 - R;
        if (R < 1) R += T * T * R / (1 - R * R);
        o.eta = 1.0f;
        if (sample.x <= R) {
            o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.pdf = R;
            return evalTex(M, n.t0, its);
        }
        o.type = MTS_ENULL; o.wo = V3(-wi.x, -wi.y, -wi.z); o.pdf = 1 - R;
        return evalTex(M, n.t1, its);
    }
    case MTS_BSDF_COATING: { /* coating.cpp:298-371 */
        float eta = n.eta[0], invEta = 1 / eta;
        float R12;
        V3 wiPrime = coatRefractIn(eta, wi, R12);
        float w = n.f[2];
        float probSpecular = (R12 * w) / (R12 * w + (1 - R12) * (1 - w));
        bool choseSpecular = true;
        if (sample.x < probSpecular) sample.x /= probSpecular;
        else { sample.x = (sample.x - probSpecular) / (1 - probSpecular); choseSpecular = false; }
        if (choseSpecular) {
            o.type = MTS_EDELTA_REFLECTION; o.wo = reflectZ(wi); o.eta = 1.0f; o.pdf = probSpecular;
            return evalTex(M, n.t1, its) * (R12 / o.pdf);
        } else {
            if (R12 == 1.0f) return V3(0.0f);
            V3 result = bsdfSample(M, n.child[0], its, wiPrime, sample, o);
            if (isZero(result)) return V3(0.0f);
            V3 woPrime = o.wo;
            V3 sigmaA = evalTex(M, n.t0, its) * n.f[0];
            if (!isZero(sigmaA))
                result *= vexp(-sigmaA * (1 / std::fabs(Frame::cosTheta(wiPrime)) + 1 / std::fabs(Frame::cosTheta(woPrime))));
            float R21;
            o.wo = coatRefractOut(eta, woPrime, R21);
            if (R21 == 1.0f) return V3(0.0f);
            o.pdf *= 1.0f - probSpecular;
            result = result / (1.0f - probSpecular);
            result *= (1 - R12) * (1 - R21);
            if (measureOf(o.type) == MTS_MEASURE_SOLID_ANGLE)
                o.pdf *= invEta * invEta * Frame::cosTheta(o.wo) / Frame::cosTheta(woPrime);
            return result;
        }
    }
    case MTS_BSDF_BLEND: { /* blendbsdf.cpp:223-262 */
        float weights[2];
        weights[1] = std::min(1.0f, std::max(0.0f, average(evalTex(M, n.t0, its))));
        weights[0] = 1 - weights[1];
        int entry;
        if (sample.x < weights[0]) { entry = 0; sample.x /= weights[0]; }
        else { entry = 1; sample.x = (sample.x - weights[0]) / weights[1]; }
        V3 result = bsdfSample(M, n.child[entry], its, wi, sample, o);
        if (isZero(result)) return result;
        result *= weights[entry] * o.pdf;
        o.pdf *= weights[entry];
        int measure = measureOf(o.type);
        int other = 1 - entry;
        o.pdf += bsdfPdf(M, n.child[other], its, wi, o.wo, measure) * weights[other];
        result += bsdfEval(M, n.child[other], its, wi, o.wo, measure) * weights[other];
        return result / o.pdf;
    }
    case MTS_BSDF_MIXTURE: { /* mixturebsdf.cpp:244-268 (the pdf-returning overload path.cpp:209 calls) */
        MixtureTab mt(M, n);
        int entry = mt.sampleReuse(sample.x);
        V3 result = bsdfSample(M, mt.child(entry), its, wi, sample, o);
        if (isZero(result)) return result;
        result *= mt.weight(entry) * o.pdf;
        o.pdf *= mt.pdf(entry);
        int measure = measureOf(o.type);
        for (int i = 0; i < mt.n; ++i) {
            if (i == entry) continue;
            o.pdf += bsdfPdf(M, mt.child(i), its, wi, o.wo, measure) * mt.pdf(i);
            result += bsdfEval(M, mt.child(i), its, wi, o.wo, measure) * mt.weight(i);
        }
        return result / o.pdf;
    }
    case MTS_BSDF_ROUGHCOATING: { /* roughcoating.cpp:378-448 */
        float eta = n.eta[0], invEta = 1 / eta;
        bool hasNested = (M.nodes[n.child[0]].lobes & MTS_EALL) != 0;
        bool choseSpecular = true;
        float alpha = average(evalTex(M, n.t2, its));
        Microfacet distr(n.flags & MTS_BF_GGX, alpha, alpha, n.flags & MTS_BF_SAMPLE_VISIBLE);
        if (hasNested) {
            float probSpecular = roughCoatProbSpecular(M, n, wi, distr.alphaU);
            if (sample.y < probSpecular) sample.y /= probSpecular;
            else { sample.y = (sample.y - probSpecular) / (1 - probSpecular); choseSpecular = false; }
        }
        if (choseSpecular) {
            float dummy;
            V3 m = distr.sampleVisible ? distr.sampleVisibleNormal(wi, sample) : distr.sampleAll(sample, dummy);
            o.wo = reflectM(wi, m); o.type = MTS_EGLOSSY_REFLECTION; o.eta = 1.0f;
            if (Frame::cosTheta(o.wo) * Frame::cosTheta(wi) <= 0) return V3(0.0f);
        } else {
            V3 result = bsdfSample(M, n.child[0], its, roughCoatRefract(invEta, wi), sample, o);
            if (isZero(result)) return V3(0.0f);
            o.wo = roughCoatRefract(eta, o.wo);
            if (isZero(o.wo)) return V3(0.0f);
        }
        int measure = measureOf(o.type);
        o.pdf = bsdfPdf(M, node, its, wi, o.wo, measure);
        if (o.pdf == 0) return V3(0.0f);
        return bsdfEval(M, node, its, wi, o.wo, measure) / o.pdf;
    }
    case MTS_BSDF_NORMALMAP: /* normalmap.cpp:218-239 */
    case MTS_BSDF_BUMPMAP: { /* bumpmap.cpp:221-240 */
        Its pert(its);
        pert.sh = perturbedFrame(M, n, its);
        V3 wiP = pert.toLocal(its.toWorld(wi));
        V3 result = bsdfSample(M, n.child[0], pert, wiP, sample, o);
        if (!isZero(result)) {
            V3 woP = o.wo;
            o.wo = its.toLocal(pert.toWorld(woP));
            if (Frame::cosTheta(o.wo) * Frame::cosTheta(woP) <= 0) return V3(0.0f);
        }
        return result;
    }
    case MTS_BSDF_MASK: { /* mask.cpp:187-226 */
        V3 opacity = evalTex(M, n.t0, its);
        float prob = luminance(opacity);
        if (sample.x < prob) {
            sample.x /= prob;
            V3 result = bsdfSample(M, n.child[0], its, wi, sample, o) * opacity / prob;
            o.pdf *= prob;
            return result;
        }
        o.wo = V3(-wi.x, -wi.y, -wi.z); o.eta = 1.0f; o.type = MTS_ENULL; o.pdf = 1 - prob;
        return (V3(1.0f) - opacity) / o.pdf;
    }
    case MTS_BSDF_TWOSIDED: { /* twosided.cpp:158-180 (the pdf-returning overload path.cpp:236 calls) */
        bool flipped = Frame::cosTheta(wi) < 0;
        V3 wiF = flipped ? V3(wi.x, wi.y, -wi.z) : wi;
        V3 result = bsdfSample(M, n.child[flipped ? 1 : 0], its, wiF, sample, o);
        if (flipped && !isZero(result) && o.pdf != 0) o.wo.z *= -1;
        return result;
    }
    }
    return V3(0.0f);
}

} // namespace orc
