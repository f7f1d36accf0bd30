/*
 * oracle_mipmap.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement of the reference's MIP map: pyramid construction with the 2-lobe Lanczos resampler
 * (include/mitsuba/render/mipmap.h:143-303, include/mitsuba/core/rfilter.h:107-460, src/rfilters/lanczos.cpp:38-50,
 * src/libcore/bitmap.cpp:2230-2329) and the filtered lookups (mipmap.h:503-715 evalTexel / evalBox / evalBilinear /
 * eval, :764-834 evalEWA).  Used by bitmap textures (src/textures/bitmap.cpp:431-499) and by the environment map
 * (src/emitters/envmap.cpp:138-182, 380-410).
 */
#pragma once
#include <cmath>
#include <vector>
#include "oracle_math.h"
#define MTS_PREFIX mtsoracle_
#include "../include/mtscuda.h"

namespace orc {

constexpr int kMipLutSize = 64;   /* MTS_MIPMAP_LUT_SIZE, mipmap.h:37 */

/* LanczosSincFilter::eval with m_radius = lobes = 2 (src/rfilters/lanczos.cpp:38-50) */
inline float lanczosEval(float x, float radius) {
    x = std::fabs(x);
    if (x < kEps) return 1.0f;
    else if (x > radius) return 0.0f;
    float x1 = kPi * x, x2 = x1 / radius;
    return (std::sin(x1) * std::sin(x2)) / (x1 * x2);
}

/* Resampler<float> in resampling mode (rfilter.h:123-176): start index and normalised weights of every target sample */
struct Resampler {
    int sourceRes, targetRes, taps, bc;
    std::vector<int> start;
    std::vector<float> weights;
    Resampler(int bc_, int sourceRes_, int targetRes_) : sourceRes(sourceRes_), targetRes(targetRes_), bc(bc_) {
        const float radius = 2.0f;
        float filterRadius = radius, scale = 1.0f, invScale = 1.0f;
        if (targetRes < sourceRes) {
            scale = (float)sourceRes / (float)targetRes;
            invScale = 1 / scale;
            filterRadius *= scale;
        }
        taps = (int)std::ceil(filterRadius * 2);
        start.resize(targetRes); weights.resize((size_t)taps * targetRes);
        for (int i = 0; i < targetRes; i++) {
            float center = (i + 0.5f) / targetRes * sourceRes;
            start[i] = floorToInt(center - filterRadius + 0.5f);
            float sum = 0;
            for (int j = 0; j < taps; j++) {
                float pos = start[i] + j + 0.5f - center;
                float weight = lanczosEval(pos * invScale, radius);
                weights[(size_t)i * taps + j] = weight;
                sum += weight;
            }
            float normalization = 1.0f / sum;
            for (int j = 0; j < taps; j++) weights[(size_t)i * taps + j] *= normalization;
        }
    }
    /* Resampler::lookup (rfilter.h:432-458) */
    float lookup(const float *source, int pos, size_t stride, int ch) const {
        if (pos < 0 || pos >= sourceRes) {
            switch (bc) {
                case MTS_WRAP_CLAMP: pos = clampi(pos, 0, sourceRes - 1); break;
                case MTS_WRAP_REPEAT: pos = modulo(pos, sourceRes); break;
                case MTS_WRAP_MIRROR: pos = modulo(pos, 2 * sourceRes); if (pos >= sourceRes) pos = 2 * sourceRes - pos - 1; break;
                case MTS_WRAP_ZERO: return 0.0f;
                case MTS_WRAP_ONE: return 1.0f;
            }
        }
        return source[stride * pos + ch];
    }
    /* resampleAndClamp / resample (rfilter.h:229-340): sequential fp32 accumulation over the taps */
    void run(const float *source, size_t sourceStride, float *target, size_t targetStride, int channels, bool clamp, float lo, float hi) const {
        sourceStride *= channels; targetStride *= channels;
        for (int i = 0; i < targetRes; ++i) {
            for (int ch = 0; ch < channels; ++ch) {
                float result = 0;
                for (int j = 0; j < taps; ++j) result += lookup(source, start[i] + j, sourceStride, ch) * weights[(size_t)i * taps + j];
                target[targetStride * i + ch] = clamp ? std::min(hi, std::max(lo, result)) : result;
            }
        }
    }
};

struct MipLevel { int w = 0, h = 0; std::vector<float> tx; /* channels*w*h, rounded through half */ };

struct MipMap {
    int channels = 3, bcu = MTS_WRAP_REPEAT, bcv = MTS_WRAP_REPEAT, filter = MTS_FILTER_BILINEAR;
    float maxAnisotropy = 1.0f;
    std::vector<MipLevel> levels;
    float weightLut[kMipLutSize];

    /* TMIPMap ctor (mipmap.h:162-303): level 0 = the input rounded to half; every further level is resampled from the
       previous *float* bitmap (Bitmap::resample, bitmap.cpp:2230-2329: rows first, then columns, clamped to
       [0, maxValue]) and then rounded to half */
    void build(const float *texels, int w, int h, int channels_, int bcu_, int bcv_, int filter_, float maxAniso, float maxValue) {
        channels = channels_; bcu = bcu_; bcv = bcv_; filter = filter_;
        maxAnisotropy = filter == MTS_FILTER_EWA ? maxAniso : 1.0f;   /* bitmap.cpp:232-233 */
        levels.clear();
        std::vector<float> cur(texels, texels + (size_t)w * h * channels);
        for (float &v : cur) if (v < 0) v = 0;   /* clampNegative, mipmap.h:228-236 */
        auto push = [&](const std::vector<float> &f, int lw, int lh) {
            MipLevel l; l.w = lw; l.h = lh; l.tx.resize(f.size());
            for (size_t i = 0; i < f.size(); ++i) l.tx[i] = roundToHalf(f[i]);
            levels.push_back(std::move(l));
        };
        push(cur, w, h);
        if (filter == MTS_FILTER_TRILINEAR || filter == MTS_FILTER_EWA) {
            const bool clamp = true;   /* minValue = 0 is finite, so resampleAndClamp always runs */
            int sw = w, sh = h;
            while (sw > 1 || sh > 1) {
                int tw = std::max(1, (sw + 1) / 2), th = std::max(1, (sh + 1) / 2);
                std::vector<float> tmp, next;
                const std::vector<float> *src = &cur;
                if (sw != tw) {   /* along x: each source row -> target row (bitmap.cpp:2259-2303) */
                    Resampler r(bcu, sw, tw);
                    tmp.resize((size_t)tw * sh * channels);
                    for (int y = 0; y < sh; ++y) r.run(cur.data() + (size_t)y * sw * channels, 1, tmp.data() + (size_t)y * tw * channels, 1, channels, clamp, 0.0f, maxValue);
                    src = &tmp;
                }
                if (sh != th) {   /* along y: each column (bitmap.cpp:2305-2328) */
                    Resampler r(bcv, sh, th);
                    next.resize((size_t)tw * th * channels);
                    for (int x = 0; x < tw; ++x) r.run(src->data() + (size_t)x * channels, tw, next.data() + (size_t)x * channels, tw, channels, clamp, 0.0f, maxValue);
                } else next = *src;
                cur.swap(next); sw = tw; sh = th;
                push(cur, sw, sh);
            }
        }
        if (filter == MTS_FILTER_EWA)   /* mipmap.h:296-302; math::fastexp evaluates in double (math.h:176-199) */
            for (int i = 0; i < kMipLutSize; ++i) {
                float r2 = (float)i / (float)(kMipLutSize - 1);
                weightLut[i] = (float)std::exp((double)(-2.0f * r2)) - (float)std::exp((double)-2.0f);
            }
    }

    /* mipmap.h:503-563 */
    V3 evalTexel(int level, int x, int y) const {
        const MipLevel &l = levels[level];
        if (x < 0 || x >= l.w) {
            switch (bcu) {
                case MTS_WRAP_REPEAT: x = modulo(x, l.w); break;
                case MTS_WRAP_CLAMP: x = clampi(x, 0, l.w - 1); break;
                case MTS_WRAP_MIRROR: x = modulo(x, 2 * l.w); if (x >= l.w) x = 2 * l.w - x - 1; break;
                case MTS_WRAP_ZERO: return V3(0.0f);
                case MTS_WRAP_ONE: return V3(1.0f);
            }
        }
        if (y < 0 || y >= l.h) {
            switch (bcv) {
                case MTS_WRAP_REPEAT: y = modulo(y, l.h); break;
                case MTS_WRAP_CLAMP: y = clampi(y, 0, l.h - 1); break;
                case MTS_WRAP_MIRROR: y = modulo(y, 2 * l.h); if (y >= l.h) y = 2 * l.h - y - 1; break;
                case MTS_WRAP_ZERO: return V3(0.0f);
                case MTS_WRAP_ONE: return V3(1.0f);
            }
        }
        const float *p = &l.tx[((size_t)y * l.w + x) * channels];
        return channels == 1 ? V3(p[0]) : V3(p[0], p[1], p[2]);
    }
    /* :566-569 */
    V3 evalBox(int level, V2 uv) const {
        const MipLevel &l = levels[level];
        return evalTexel(level, floorToInt(uv.x * l.w), floorToInt(uv.y * l.h));
    }
    /* :575-596 */
    V3 evalBilinear(int level, V2 uv) const {
        if (!std::isfinite(uv.x) || !std::isfinite(uv.y)) return V3(0.0f);
        if (level >= (int)levels.size()) return evalBox((int)levels.size() - 1, uv);
        const MipLevel &l = levels[level];
        float u = uv.x * l.w - 0.5f, v = uv.y * l.h - 0.5f;
        int xPos = floorToInt(u), yPos = floorToInt(v);
        float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
        return evalTexel(level, xPos, yPos) * dx2 * dy2 + evalTexel(level, xPos, yPos + 1) * dx2 * dy1 +
               evalTexel(level, xPos + 1, yPos) * dx1 * dy2 + evalTexel(level, xPos + 1, yPos + 1) * dx1 * dy1;
    }
    /* :601-626 at level 0 */
    void evalGradientBilinear(V2 uv, V3 grad[2]) const {
        if (!std::isfinite(uv.x) || !std::isfinite(uv.y)) { grad[0] = grad[1] = V3(0.0f); return; }
        const MipLevel &l = levels[0];
        float u = uv.x * l.w - 0.5f, v = uv.y * l.h - 0.5f;
        int xPos = floorToInt(u), yPos = floorToInt(v);
        float dx = u - xPos, dy = v - yPos;
        V3 p00 = evalTexel(0, xPos, yPos), p10 = evalTexel(0, xPos + 1, yPos);
        V3 p01 = evalTexel(0, xPos, yPos + 1), p11 = evalTexel(0, xPos + 1, yPos + 1);
        V3 tmp = p01 + p10 - p11;
        grad[0] = (p10 + p00 * (dy - 1) - tmp * dy) * (float)l.w;
        grad[1] = (p01 + p00 * (dx - 1) - tmp * dx) * (float)l.h;
    }
    /* :764-834 */
    V3 evalEWA(int level, V2 uv, float A, float B, float C) const {
        if (!std::isfinite(A + B + C + uv.x + uv.y)) return V3(0.0f);
        if (level >= (int)levels.size()) return evalBox((int)levels.size() - 1, uv);
        const MipLevel &l = levels[level];
        float u = uv.x * l.w - 0.5f, v = uv.y * l.h - 0.5f;
        const float ratioX = (float)l.w / (float)levels[0].w, ratioY = (float)l.h / (float)levels[0].h;   /* m_sizeRatio, :258-260 */
        A /= ratioX * ratioX; B /= ratioX * ratioY; C /= ratioY * ratioY;
        float invDet = 1.0f / (-B * B + 4.0f * A * C), deltaU = 2.0f * std::sqrt(C * invDet), deltaV = 2.0f * std::sqrt(A * invDet);
        int u0 = (int)std::ceil(u - deltaU), u1 = floorToInt(u + deltaU);
        int v0 = (int)std::ceil(v - deltaV), v1 = floorToInt(v + deltaV);
        float As = A * kMipLutSize, Bs = B * kMipLutSize, Cs = C * kMipLutSize;
        V3 result(0.0f);
        float denominator = 0.0f;
        float ddq = 2 * As, uu0 = (float)u0 - u;
        for (int vt = v0; vt <= v1; ++vt) {
            const float vv = (float)vt - v;
            float q = As * uu0 * uu0 + (Bs * uu0 + Cs * vv) * vv;
            float dq = As * (2 * uu0 + 1) + Bs * vv;
            for (int ut = u0; ut <= u1; ++ut) {
                if (q < (float)kMipLutSize) {
                    uint32_t qi = (uint32_t)q;
                    if (qi < (uint32_t)kMipLutSize) {
                        const float weight = weightLut[(int)q];
                        result += evalTexel(level, ut, vt) * weight;
                        denominator += weight;
                    }
                }
                q += dq;
                dq += ddq;
            }
        }
        if (denominator == 0) return evalBilinear(level, uv);
        return result / denominator;
    }
    /* :629-715 */
    V3 eval(V2 uv, V2 d0, V2 d1) const {
        if (filter == MTS_FILTER_NEAREST) return evalBox(0, uv);
        else if (filter == MTS_FILTER_BILINEAR) return evalBilinear(0, uv);
        const MipLevel &l0 = levels[0];
        float du0 = d0.x * l0.w, dv0 = d0.y * l0.h, du1 = d1.x * l0.w, dv1 = d1.y * l0.h;
        float A = dv0 * dv0 + dv1 * dv1, B = -2.0f * (du0 * dv0 + du1 * dv1), C = du0 * du0 + du1 * du1, F = A * C - B * B * 0.25f;
        float root = hypot2(A - C, B), Aprime = 0.5f * (A + C - root), Cprime = 0.5f * (A + C + root),
              majorRadius = Aprime != 0 ? std::sqrt(F / Aprime) : 0, minorRadius = Cprime != 0 ? std::sqrt(F / Cprime) : 0;
        const float invLn2 = 1.0f / std::log(2.0f);   /* math::log2 = fastlog * invLn2 (src/libcore/math.cpp:103-106) */
        if (filter == MTS_FILTER_TRILINEAR || !(minorRadius > 0) || !(majorRadius > 0) || F < 0) {
            float level = (float)std::log((double)std::max(majorRadius, kEps)) * invLn2;
            int ilevel = floorToInt(level);
            if (ilevel < 0) return evalBilinear(0, uv);
            float a = level - ilevel;
            return evalBilinear(ilevel, uv) * (1.0f - a) + evalBilinear(ilevel + 1, uv) * a;
        }
        if (minorRadius * maxAnisotropy < majorRadius) {
            minorRadius = majorRadius / maxAnisotropy;
            float theta = 0.5f * std::atan(B / (A - C)), sinTheta = std::sin(theta), cosTheta = std::cos(theta);
            float a2 = majorRadius * majorRadius, b2 = minorRadius * minorRadius, sinTheta2 = sinTheta * sinTheta,
                  cosTheta2 = cosTheta * cosTheta, sin2Theta = 2 * sinTheta * cosTheta;
            A = a2 * cosTheta2 + b2 * sinTheta2;
            B = (a2 - b2) * sin2Theta;
            C = a2 * sinTheta2 + b2 * cosTheta2;
            F = a2 * b2;
        }
        float scale = 1.0f / F;
        A *= scale; B *= scale; C *= scale;
        float level = std::max(0.0f, (float)std::log((double)minorRadius) * invLn2);
        int ilevel = (int)level;
        float a = level - ilevel;
        if (majorRadius < 1 || !(A > 0 && C > 0)) return evalBilinear(ilevel, uv);
        return evalEWA(ilevel, uv, A, B, C) * (1.0f - a) + evalEWA(ilevel + 1, uv, A, B, C) * a;
    }
};

} // namespace orc
