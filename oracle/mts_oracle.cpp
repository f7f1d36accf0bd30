/*
 * mts_oracle.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement ("port") of the Mitsuba 0.6 unidirectional path-tracing hot
 * path: SamplingIntegrator::renderBlock -> MIPathTracer::Li -> ShapeKDTree
 * ray queries -> BSDF / emitter sampling -> ImageBlock::put.  It exports the
 * C ABI of include/mtscuda.h with the prefix `mtsoracle_` so the parity tests
 * can feed identical inputs to this code and to the CUDA library.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (libmtscuda) never does.
 *
 * Parity pinning: see oracle/README.md -- this restatement is checked against
 * the reference's own known-answer tests (test_dgeom, chi-square / sample==eval/pdf
 * identities of test_chisquare and test_microfacet, the rough-transmittance LUT)
 * and, where oracle/_ref could be built from the reference sources, against the
 * reference itself.
 *
 * All citations are paths relative to the reference root.
 */
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "oracle_bsdf.h"

using namespace orc;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string &msg) { g_err = msg; return code; }

/* ------------------------------------------------------------------------ */
struct Shape {
    int type = 0; /* 0 mesh, 1 rectangle, 2 sphere */
    int bsdf = 0;
    int emitter = -1;
    /* mesh */
    std::vector<V3> pos, nrm;
    std::vector<V2> uv;
    std::vector<uint32_t> idx;
    std::vector<V3> tang; /* 2 per triangle */
    std::vector<float> areaCdf;
    float invArea = 0;
    /* rectangle / sphere */
    M44 o2w, w2o;
    V3 dpdu, dpdv; Frame frame;
    V3 center; float radius = 0; bool flip = false;
};

struct Emitter {
    int type = 0, shape = -1;
    V3 radiance, position;
    float weight = 1, scale = 1;
    M44 toWorld, toLocal;
    int w = 0, h = 0;
    std::vector<float> tex;                       /* w*h*3, half-rounded (= mip level 0) */
    MipMap mip;                                   /* m_mipmap: EWA pyramid for camera rays with differentials */
    std::vector<float> cdfRows, cdfCols, rowWeights;
    float normalization = 0;
    V3 bsCenter; float bsRadius = 0;              /* m_sceneBSphere */
};

/* TriAccel (include/mitsuba/render/triaccel.h:37-51) */
struct TriAccel {
    uint32_t k;
    float n_u, n_v, n_d, a_u, a_v, b_nu, b_nv, c_nu, c_nv;
    uint32_t shapeIndex, primIndex;
};
const uint32_t kNoTri = 0xFFFFFFFFu;

struct BNode { float lo[3], hi[3]; int left, right, first, count; };

} // namespace

struct MtsContext {
    MtsCamera cam; bool hasCam = false;
    MtsFilm film; bool hasFilm = false;
    Materials mat;
    std::vector<Shape> shapes;
    std::vector<Emitter> emitters;
    std::vector<float> emitterCdf;
    int envIndex = -1;
    std::vector<TriAccel> prims;
    std::vector<BNode> nodes;
    V3 sceneLo, sceneHi;
    bool built = false;
    std::vector<float> filmBuf; /* w*h*4 */
    std::vector<float> alphaBuf; /* w*h, films with an alpha channel only */
    MtsStats stats;
    std::atomic<uint64_t> normalRays{0}, shadowRays{0};
};

namespace {

/* ---- TriAccel::load / rayIntersect (triaccel.h:61-158) ------------------- */
int triLoad(TriAccel &ta, V3 A, V3 B, V3 C) {
    static const int waldModulo[4] = {1, 2, 0, 1};
    V3 b = C - A, c = B - A, N = cross(c, b);
    int k = 0;
    for (int j = 0; j < 3; j++) if (std::fabs(N[j]) > std::fabs(N[k])) k = j;
    uint32_t u = waldModulo[k], v = waldModulo[k + 1];
    const float n_k = N[k], denom = b[u] * c[v] - b[v] * c[u];
    if (denom == 0) { ta.k = 3; return 1; }
    ta.k = k;
    ta.n_u = N[u] / n_k; ta.n_v = N[v] / n_k; ta.n_d = dot(A, N) / n_k;
    ta.b_nu = b[u] / denom; ta.b_nv = -b[v] / denom;
    ta.a_u = A[u]; ta.a_v = A[v];
    ta.c_nu = c[v] / denom; ta.c_nv = -c[u] / denom;
    return 0;
}
bool triIntersect(const TriAccel &ta, V3 o, V3 d, float mint, float maxt, float &u, float &v, float &t) {
    float o_u, o_v, o_k, d_u, d_v, d_k;
    switch (ta.k) {
        case 0: o_u = o.y; o_v = o.z; o_k = o.x; d_u = d.y; d_v = d.z; d_k = d.x; break;
        case 1: o_u = o.z; o_v = o.x; o_k = o.y; d_u = d.z; d_v = d.x; d_k = d.y; break;
        case 2: o_u = o.x; o_v = o.y; o_k = o.z; d_u = d.x; d_v = d.y; d_k = d.z; break;
        default: return false;
    }
    t = (ta.n_d - o_u * ta.n_u - o_v * ta.n_v - o_k) / (d_u * ta.n_u + d_v * ta.n_v + d_k);
    if (t < mint || t > maxt) return false;
    const float hu = o_u + t * d_u - ta.a_u, hv = o_v + t * d_v - ta.a_v;
    u = hv * ta.b_nu + hu * ta.b_nv;
    v = hu * ta.c_nu + hv * ta.c_nv;
    return u >= 0 && v >= 0 && u + v <= 1.0f;
}

/* ---- Rectangle::rayIntersect (src/shapes/rectangle.cpp:125-148) ---------- */
bool rectIntersect(const Shape &s, V3 o, V3 d, float mint, float maxt, float &t, float *tmp) {
    V3 lo = s.w2o.pointAffine(o), ld = s.w2o.vec(d);
    float hit = -lo.z / ld.z;
    if (!(hit >= mint && hit <= maxt)) return false;
    V3 local = lo + ld * hit;
    if (std::fabs(local.x) <= 1 && std::fabs(local.y) <= 1) {
        t = hit;
        if (tmp) { tmp[0] = local.x; tmp[1] = local.y; }
        return true;
    }
    return false;
}
/* ---- Sphere::rayIntersect (src/shapes/sphere.cpp:163-207), fp64 quadratic -- */
bool sphereIntersect(const Shape &s, V3 o, V3 d, float mint, float maxt, float &t) {
    double ox = (double)o.x - (double)s.center.x, oy = (double)o.y - (double)s.center.y, oz = (double)o.z - (double)s.center.z;
    double dx = d.x, dy = d.y, dz = d.z;
    double A = dx * dx + dy * dy + dz * dz;
    double B = 2 * (ox * dx + oy * dy + oz * dz);
    double C = ox * ox + oy * oy + oz * oz - s.radius * s.radius;
    double nearT, farT;
    if (!solveQuadraticDouble(A, B, C, nearT, farT)) return false;
    if (!(nearT <= maxt && farT >= mint)) return false;
    if (nearT < mint) {
        if (farT > maxt) return false;
        t = (float)farT;
    } else t = (float)nearT;
    return true;
}

struct Hit { float t, u, v; int prim; float tmp[2]; };

bool primIntersect(const MtsContext &c, int pi, V3 o, V3 d, float mint, float maxt, Hit &h) {
    const TriAccel &ta = c.prims[pi];
    if (ta.k != kNoTri) {
        float u, v, t;
        if (triIntersect(ta, o, d, mint, maxt, u, v, t)) { h.t = t; h.u = u; h.v = v; h.prim = pi; return true; }
        return false;
    }
    const Shape &s = c.shapes[ta.shapeIndex];
    float t, tmp[2] = {0, 0};
    bool ok = s.type == 1 ? rectIntersect(s, o, d, mint, maxt, t, tmp) : sphereIntersect(s, o, d, mint, maxt, t);
    if (ok) { h.t = t; h.u = h.v = 0; h.prim = pi; h.tmp[0] = tmp[0]; h.tmp[1] = tmp[1]; }
    return ok;
}

/* ---- acceleration structure: a plain median-split BVH.  The reference uses a SAH
 * kd-tree (ShapeKDTree); closest-hit results do not depend on the structure. ---- */
void primBounds(const MtsContext &c, int pi, float lo[3], float hi[3]) {
    const TriAccel &ta = c.prims[pi];
    const Shape &s = c.shapes[ta.shapeIndex];
    V3 pts[4]; int np;
    if (s.type == 0) {
        for (int j = 0; j < 3; ++j) pts[j] = s.pos[s.idx[3 * ta.primIndex + j]];
        np = 3;
    } else if (s.type == 1) {
        pts[0] = s.o2w.point(V3(-1, -1, 0)); pts[1] = s.o2w.point(V3(1, -1, 0));
        pts[2] = s.o2w.point(V3(1, 1, 0)); pts[3] = s.o2w.point(V3(-1, 1, 0));
        np = 4;
    } else {
        pts[0] = s.center - V3(s.radius); pts[1] = s.center + V3(s.radius); np = 2;
    }
    for (int a = 0; a < 3; ++a) { lo[a] = INFINITY; hi[a] = -INFINITY; }
    for (int j = 0; j < np; ++j) for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], pts[j][a]); hi[a] = std::max(hi[a], pts[j][a]); }
}

struct PB { float lo[3], hi[3], c[3]; int prim; };

int buildRec(MtsContext &c, std::vector<PB> &pb, int b, int e) {
    BNode n;
    for (int a = 0; a < 3; ++a) { n.lo[a] = INFINITY; n.hi[a] = -INFINITY; }
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = b; i < e; ++i) for (int a = 0; a < 3; ++a) {
        n.lo[a] = std::min(n.lo[a], pb[i].lo[a]); n.hi[a] = std::max(n.hi[a], pb[i].hi[a]);
        clo[a] = std::min(clo[a], pb[i].c[a]); chi[a] = std::max(chi[a], pb[i].c[a]);
    }
    n.left = n.right = -1; n.first = b; n.count = e - b;
    int id = (int)c.nodes.size();
    c.nodes.push_back(n);
    if (e - b <= 4) return id;
    int axis = 0;
    for (int a = 1; a < 3; ++a) if (chi[a] - clo[a] > chi[axis] - clo[axis]) axis = a;
    int mid = (b + e) / 2;
    std::nth_element(pb.begin() + b, pb.begin() + mid, pb.begin() + e, [axis](const PB &x, const PB &y) { return x.c[axis] < y.c[axis]; });
    int l = buildRec(c, pb, b, mid);
    int r = buildRec(c, pb, mid, e);
    c.nodes[id].left = l; c.nodes[id].right = r; c.nodes[id].count = 0;
    return id;
}

inline bool boxHit(const BNode &n, V3 o, V3 inv, float mint, float maxt) {
    float t0 = mint, t1 = maxt;
    for (int a = 0; a < 3; ++a) {
        float tn = (n.lo[a] - o[a]) * inv[a], tf = (n.hi[a] - o[a]) * inv[a];
        if (tn > tf) std::swap(tn, tf);
        /* NaN (0*inf) compares false and leaves the interval untouched */
        if (tn > t0) t0 = tn;
        if (tf < t1) t1 = tf;
    }
    /* small conservative slack so that boundary hits are never culled */
    return t0 <= t1 * 1.0000004f + 1e-30f;
}

/* Closest hit over [mint, maxt]; shadow: any hit. */
bool traverse(const MtsContext &c, V3 o, V3 d, float mint, float maxt, bool shadow, Hit &best) {
    if (c.nodes.empty()) return false;
    V3 inv(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[128]; int sp = 0; stack[sp++] = 0;
    bool found = false; best.t = maxt;
    while (sp) {
        const BNode &n = c.nodes[stack[--sp]];
        if (!boxHit(n, o, inv, mint, best.t)) continue;
        if (n.left < 0) {
            for (int i = n.first; i < n.first + n.count; ++i) {
                Hit h;
                if (primIntersect(c, i, o, d, mint, best.t, h)) {
                    if (shadow) return true;
                    /* the kd-tree keeps the first hit among exact ties; prefer the strictly closer one */
                    if (!found || h.t < best.t || (h.t == best.t && h.prim < best.prim)) best = h;
                    found = true;
                }
            }
        } else { stack[sp++] = n.left; stack[sp++] = n.right; }
    }
    return found;
}

thread_local uint64_t t_normalRays = 0, t_shadowRays = 0;

/* ShapeKDTree::rayIntersect epsilon logic (src/librender/skdtree.cpp:112-142, 207-226).
 * The AABB clip only bounds [mint,maxt]; primitives all lie inside, so it is omitted. */
bool sceneIntersect(MtsContext &c, V3 o, V3 d, float rmint, float rmaxt, bool shadow, Hit &h) {
    float rayMinT = rmint;
    if (rayMinT == kEps) {
        float m = std::max(std::max(std::fabs(o.x), std::fabs(o.y)), std::fabs(o.z));
        if (!shadow) m = std::max(m, kEps);
        rayMinT *= m;
    }
    if (shadow) ++t_shadowRays; else ++t_normalRays; /* per-thread; summed after the parallel loop */
    if (!(rmaxt > rayMinT)) return false;
    return traverse(c, o, d, rayMinT, rmaxt, shadow, h);
}

/* fillIntersectionRecord<true> (include/mitsuba/render/skdtree.h:343-428) */
void fillIts(const MtsContext &c, V3 o, V3 d, const Hit &h, Its &its) {
    const TriAccel &ta = c.prims[h.prim];
    const Shape &s = c.shapes[ta.shapeIndex];
    its.t = h.t; its.shape = ta.shapeIndex;
    if (s.type == 0) {
        const V3 b(1 - h.u - h.v, h.u, h.v);
        uint32_t i0 = s.idx[3 * ta.primIndex], i1 = s.idx[3 * ta.primIndex + 1], i2 = s.idx[3 * ta.primIndex + 2];
        V3 p0 = s.pos[i0], p1 = s.pos[i1], p2 = s.pos[i2];
        its.p = p0 * b.x + p1 * b.y + p2 * b.z;
        V3 side1 = p1 - p0, side2 = p2 - p0;
        V3 faceNormal = cross(side1, side2);
        float len = length(faceNormal);
        if (!isZero(faceNormal)) faceNormal = faceNormal / len;
        if (!s.tang.empty()) { its.dpdu = s.tang[2 * ta.primIndex]; its.dpdv = s.tang[2 * ta.primIndex + 1]; }
        else { its.dpdu = side1; its.dpdv = side2; }
        if (!s.nrm.empty()) {
            its.sh.n = normalize(s.nrm[i0] * b.x + s.nrm[i1] * b.y + s.nrm[i2] * b.z);
            if (dot(faceNormal, its.sh.n) < 0) faceNormal = -faceNormal;
        } else its.sh.n = faceNormal;
        its.geo = Frame(faceNormal);
        if (!s.uv.empty()) {
            V2 t0 = s.uv[i0], t1 = s.uv[i1], t2 = s.uv[i2];
            its.uv = V2{t0.x * b.x + t1.x * b.y + t2.x * b.z, t0.y * b.x + t1.y * b.y + t2.y * b.z};
        } else its.uv = V2{b.y, b.z};
        its.prim = ta.primIndex;
    } else if (s.type == 1) { /* rectangle.cpp:155-168 */
        its.geo = s.frame; its.sh.n = its.geo.n;
        its.dpdu = s.dpdu; its.dpdv = s.dpdv;
        its.uv = V2{0.5f * (h.tmp[0] + 1), 0.5f * (h.tmp[1] + 1)};
        its.p = o + d * its.t;
        its.prim = -1;
    } else { /* sphere.cpp:209-255 */
        its.p = o + d * its.t;
        its.p = s.center + normalize(its.p - s.center) * s.radius;
        V3 local = s.w2o.vec(its.p - s.center);
        float theta = safe_acos(local.z / s.radius);
        float phi = std::atan2(local.y, local.x);
        if (phi < 0) phi += 2 * kPi;
        its.uv.x = phi * (0.5f * kInvPi); its.uv.y = theta * kInvPi;
        its.dpdu = s.o2w.vec(V3(-local.y, local.x, 0) * (2 * kPi));
        its.geo.n = normalize(its.p - s.center);
        float zrad = std::sqrt(local.x * local.x + local.y * local.y);
        if (zrad > 0) {
            float invZRad = 1.0f / zrad, cosPhi = local.x * invZRad, sinPhi = local.y * invZRad;
            its.dpdv = s.o2w.vec(V3(local.z * cosPhi, local.z * sinPhi, -std::sin(theta) * s.radius) * kPi);
            its.geo.s = normalize(its.dpdu); its.geo.t = normalize(its.dpdv);
        } else {
            const float cosPhi = 0, sinPhi = 1;
            its.dpdv = s.o2w.vec(V3(local.z * cosPhi, local.z * sinPhi, -std::sin(theta) * s.radius) * kPi);
            coordinateSystem(its.geo.n, its.geo.s, its.geo.t);
        }
        if (s.flip) its.geo.n *= -1;
        its.sh.n = its.geo.n;
        its.prim = -1;
    }
    computeShadingFrame(its.sh.n, its.dpdu, its.sh);
    its.wi = its.toLocal(-d);
}

/* ---- emitters ---------------------------------------------------------------- */
struct DRec { V3 ref, refN, p, n, d; float dist, pdf; int measure; int emitter; };

/* DiscreteDistribution::sample / sampleReuse (include/mitsuba/core/pmf.h) */
size_t cdfSample(const std::vector<float> &cdf, float v) {
    auto entry = std::lower_bound(cdf.begin(), cdf.end(), v);
    size_t index = std::min(cdf.size() - 2, (size_t)std::max((ptrdiff_t)0, entry - cdf.begin() - 1));
    while (cdf[index + 1] - cdf[index] == 0 && index < cdf.size() - 1) ++index;
    return index;
}
size_t cdfSampleReuse(const std::vector<float> &cdf, float &v, float *pdf = nullptr) {
    size_t index = cdfSample(cdf, v);
    if (pdf) *pdf = cdf[index + 1] - cdf[index];
    v = (v - cdf[index]) / (cdf[index + 1] - cdf[index]);
    return index;
}

/* Shape::samplePosition for the three shape kinds */
void samplePosition(const Shape &s, V2 sample, V3 &p, V3 &n, float &pdf) {
    if (s.type == 0) { /* TriMesh::samplePosition (trimesh.cpp:412-424), Triangle::sample (triangle.cpp:23-56) */
        size_t index = cdfSampleReuse(s.areaCdf, sample.y);
        uint32_t i0 = s.idx[3 * index], i1 = s.idx[3 * index + 1], i2 = s.idx[3 * index + 2];
        V3 p0 = s.pos[i0], p1 = s.pos[i1], p2 = s.pos[i2];
        V2 bary = squareToUniformTriangle(sample);
        V3 sideA = p1 - p0, sideB = p2 - p0;
        p = p0 + (sideA * bary.x) + (sideB * bary.y);
        if (!s.nrm.empty()) n = normalize(s.nrm[i0] * (1.0f - bary.x - bary.y) + s.nrm[i1] * bary.x + s.nrm[i2] * bary.y);
        else n = normalize(cross(sideA, sideB));
        pdf = s.invArea;
    } else if (s.type == 1) { /* rectangle.cpp:210-216 */
        p = s.o2w.point(V3(sample.x * 2 - 1, sample.y * 2 - 1, 0));
        n = s.frame.n; pdf = s.invArea;
    } else { /* sphere.cpp:257-266 */
        V3 v = squareToUniformSphere(sample);
        p = v * s.radius + s.center; n = v;
        if (s.flip) n *= -1;
        pdf = s.invArea;
    }
}

/* Shape::sampleDirect (shape.cpp:99-112) / Sphere::sampleDirect (sphere.cpp:286-346) */
void shapeSampleDirect(const Shape &s, DRec &dRec, V2 sample) {
    if (s.type != 2) {
        samplePosition(s, sample, dRec.p, dRec.n, dRec.pdf);
        dRec.d = dRec.p - dRec.ref;
        float distSquared = dot(dRec.d, dRec.d);
        dRec.dist = std::sqrt(distSquared);
        dRec.d = dRec.d / dRec.dist;
        float dp = absDot(dRec.d, dRec.n);
        dRec.pdf *= dp != 0 ? (distSquared / dp) : 0.0f;
        dRec.measure = MTS_MEASURE_SOLID_ANGLE;
        return;
    }
    const V3 refToCenter = s.center - dRec.ref;
    const float refDist2 = dot(refToCenter, refToCenter);
    const float invRefDist = 1.0f / std::sqrt(refDist2);
    const float sinAlpha = s.radius * invRefDist;
    if (sinAlpha < 1 - kEps) {
        float cosAlpha = safe_sqrt(1.0f - sinAlpha * sinAlpha);
        dRec.d = Frame(refToCenter * invRefDist).toWorld(squareToUniformCone(cosAlpha, sample));
        dRec.pdf = squareToUniformConePdf(cosAlpha);
        const float projDist = dot(refToCenter, dRec.d);
        const float baseT = refDist2 / projDist;
        const V3 query = dRec.ref + dRec.d * baseT;
        const V3 queryToCenter = s.center - query;
        const float queryDist2 = dot(queryToCenter, queryToCenter);
        const float queryProjDist = dot(queryToCenter, dRec.d);
        float A = 1.0f, B = -2 * queryProjDist, C = queryDist2 - s.radius * s.radius;
        float nearT, farT;
        if (!solveQuadratic(A, B, C, nearT, farT)) nearT = queryProjDist;
        dRec.dist = baseT + nearT;
        dRec.n = normalize(dRec.d * nearT - queryToCenter);
        dRec.p = s.center + dRec.n * s.radius;
    } else {
        V3 d = squareToUniformSphere(sample);
        dRec.p = s.center + d * s.radius;
        dRec.n = d;
        dRec.d = dRec.p - dRec.ref;
        float dist2 = dot(dRec.d, dRec.d);
        dRec.dist = std::sqrt(dist2);
        dRec.d = dRec.d / dRec.dist;
        dRec.pdf = s.invArea * dist2 / absDot(dRec.d, dRec.n);
    }
    if (s.flip) dRec.n *= -1;
    dRec.measure = MTS_MEASURE_SOLID_ANGLE;
}

/* Shape::pdfDirect (shape.cpp:114-124) / Sphere::pdfDirect (sphere.cpp:348-377), solid angle measure */
float shapePdfDirect(const Shape &s, const DRec &dRec) {
    if (s.type != 2) return s.invArea * (dRec.dist * dRec.dist) / absDot(dRec.d, dRec.n);
    const V3 refToCenter = s.center - dRec.ref;
    const float invRefDist = 1.0f / length(refToCenter);
    const float sinAlpha = s.radius * invRefDist;
    if (sinAlpha < 1 - kEps) {
        float cosAlpha = safe_sqrt(1 - sinAlpha * sinAlpha);
        return squareToUniformConePdf(cosAlpha);
    }
    return s.invArea * dRec.dist * dRec.dist / absDot(dRec.d, dRec.n);
}

/* EnvironmentMap helpers (src/emitters/envmap.cpp) */
V3 envTexel(const Emitter &e, int x, int y) { /* evalTexel with repeat-u / clamp-v (:145-146,179) */
    x = modulo(x, e.w); y = clampi(y, 0, e.h - 1);
    const float *p = &e.tex[((size_t)y * e.w + x) * 3];
    return V3(p[0], p[1], p[2]);
}
uint32_t envSampleReuse(const float *cdf, uint32_t size, float &sample) { /* :657-662 */
    const float *entry = std::lower_bound(cdf, cdf + size + 1, sample);
    uint32_t index = std::min((uint32_t)std::max((ptrdiff_t)0, entry - cdf - 1), size - 1);
    sample = (sample - cdf[index]) / (cdf[index + 1] - cdf[index]);
    return index;
}
/* :567-600 */
void envSampleDirection(const Emitter &e, V2 sample, V3 &d, V3 &value, float &pdf) {
    uint32_t row = envSampleReuse(e.cdfRows.data(), e.h, sample.y);
    uint32_t col = envSampleReuse(e.cdfCols.data() + (size_t)row * (e.w + 1), e.w, sample.x);
    V2 tent = squareToTent(sample);
    float posx = (float)col + tent.x, posy = (float)row + tent.y;
    int xPos = floorToInt(posx), yPos = floorToInt(posy);
    float dx1 = posx - xPos, dx2 = 1.0f - dx1, dy1 = posy - yPos, dy2 = 1.0f - dy1;
    V3 value1 = envTexel(e, xPos, yPos) * dx2 * dy2 + envTexel(e, xPos + 1, yPos) * dx1 * dy2;
    V3 value2 = envTexel(e, xPos, yPos + 1) * dx2 * dy1 + envTexel(e, xPos + 1, yPos + 1) * dx1 * dy1;
    value = (value1 + value2) * e.scale;
    pdf = (luminance(value1) * e.rowWeights[clampi(yPos, 0, e.h - 1)] +
           luminance(value2) * e.rowWeights[clampi(yPos + 1, 0, e.h - 1)]) * e.normalization;
    float sinPhi, cosPhi, sinTheta, cosTheta;
    sincosf((2 * kPi / e.w) * (posx + 0.5f), &sinPhi, &cosPhi);
    sincosf((kPi / e.h) * (posy + 0.5f), &sinTheta, &cosTheta);
    d = V3(sinPhi * sinTheta, cosTheta, -cosPhi * sinTheta);
    pdf /= std::max(std::fabs(sinTheta), kEps);
}
/* :603-633 */
float envPdfDirection(const Emitter &e, V3 d) {
    float uvx = std::atan2(d.x, -d.z) * kInvTwoPi, uvy = safe_acos(d.y) * kInvPi;
    if (!std::isfinite(uvx) || !std::isfinite(uvy)) return 0.0f;
    float u = uvx * e.w - 0.5f, v = uvy * e.h - 0.5f;
    int xPos = floorToInt(u), yPos = floorToInt(v);
    float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
    V3 value1 = envTexel(e, xPos, yPos) * dx2 * dy2 + envTexel(e, xPos + 1, yPos) * dx1 * dy2;
    V3 value2 = envTexel(e, xPos, yPos + 1) * dx2 * dy1 + envTexel(e, xPos + 1, yPos + 1) * dx1 * dy1;
    float sinTheta = safe_sqrt(1 - d.y * d.y);
    return (luminance(value1) * e.rowWeights[clampi(yPos, 0, e.h - 1)] +
            luminance(value2) * e.rowWeights[clampi(yPos + 1, 0, e.h - 1)]) * e.normalization / std::max(std::fabs(sinTheta), kEps);
}
/* evalEnvironment without differentials (:380-410): bilinear level-0 lookup */
V3 envEval(const Emitter &e, V3 dWorld) {
    V3 v = e.toLocal.vec(dWorld);
    float uvx = std::atan2(v.x, -v.z) * kInvTwoPi, uvy = safe_acos(v.y) * kInvPi;
    if (!std::isfinite(uvx) || !std::isfinite(uvy)) return V3(0.0f);
    float u = uvx * e.w - 0.5f, w = uvy * e.h - 0.5f;
    int xPos = floorToInt(u), yPos = floorToInt(w);
    float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = w - yPos, dy2 = 1.0f - dy1;
    V3 r = envTexel(e, xPos, yPos) * dx2 * dy2 + envTexel(e, xPos, yPos + 1) * dx2 * dy1 +
           envTexel(e, xPos + 1, yPos) * dx1 * dy2 + envTexel(e, xPos + 1, yPos + 1) * dx1 * dy1;
    return r * e.scale;
}
/* evalEnvironment of a ray with differentials (:391-407) */
V3 envEvalDiff(const Emitter &e, V3 dWorld, V3 rxD, V3 ryD) {
    V3 v = e.toLocal.vec(dWorld);
    V2 uv{std::atan2(v.x, -v.z) * kInvTwoPi, safe_acos(v.y) * kInvPi};
    V3 dvdx = e.toLocal.vec(rxD) - v, dvdy = e.toLocal.vec(ryD) - v;
    float t1 = kInvTwoPi / (v.x * v.x + v.z * v.z), t2 = -kInvPi / std::max(safe_sqrt(1.0f - v.y * v.y), kEps);
    V2 dudx{t1 * (dvdx.z * v.x - dvdx.x * v.z), t2 * dvdx.y}, dudy{t1 * (dvdy.z * v.x - dvdy.x * v.z), t2 * dvdy.y};
    return e.mip.eval(uv, dudx, dudy) * e.scale;
}
/* BSphere::rayIntersect (include/mitsuba/core/bsphere.h:88-95) */
bool bsphereIntersect(const Emitter &e, V3 o, V3 d, float &nearT, float &farT) {
    V3 oc = o - e.bsCenter;
    return solveQuadratic(dot(d, d), 2 * dot(oc, d), dot(oc, oc) - e.bsRadius * e.bsRadius, nearT, farT);
}

/* Emitter::sampleDirect (area.cpp:158-172, point.cpp:131-147, envmap.cpp:516-543) */
V3 emitterSampleDirect(const MtsContext &c, const Emitter &e, DRec &dRec, V2 sample) {
    if (e.type == MTS_EMITTER_AREA) {
        shapeSampleDirect(c.shapes[e.shape], dRec, sample);
        if (dot(dRec.d, dRec.refN) >= 0 && dot(dRec.d, dRec.n) < 0 && dRec.pdf != 0) return e.radiance / dRec.pdf;
        dRec.pdf = 0.0f; return V3(0.0f);
    } else if (e.type == MTS_EMITTER_POINT) {
        dRec.p = e.position; dRec.pdf = 1.0f; dRec.measure = MTS_MEASURE_DISCRETE;
        dRec.d = dRec.p - dRec.ref;
        dRec.dist = length(dRec.d);
        float invDist = 1.0f / dRec.dist;
        dRec.d *= invDist; dRec.n = V3(0.0f);
        return e.radiance * (invDist * invDist);
    } else {
        V3 value, d; float pdf;
        envSampleDirection(e, sample, d, value, pdf);
        V3 dw = e.toWorld.vec(d);
        float nearT, farT;
        if (isZero(value) || pdf == 0 || !bsphereIntersect(e, dRec.ref, dw, nearT, farT) || nearT >= 0 || farT <= 0) {
            dRec.pdf = 0.0f; return V3(0.0f);
        }
        dRec.pdf = pdf; dRec.p = dRec.ref + dw * farT;
        dRec.n = normalize(e.bsCenter - dRec.p);
        dRec.dist = farT; dRec.d = dw; dRec.measure = MTS_MEASURE_SOLID_ANGLE;
        return value / pdf;
    }
}
/* Emitter::pdfDirect (area.cpp:174-183, point.cpp:149-151, envmap.cpp:545-556) */
float emitterPdfDirect(const MtsContext &c, const Emitter &e, const DRec &dRec) {
    if (e.type == MTS_EMITTER_AREA) {
        if (dot(dRec.d, dRec.refN) >= 0 && dot(dRec.d, dRec.n) < 0) return shapePdfDirect(c.shapes[e.shape], dRec);
        return 0.0f;
    } else if (e.type == MTS_EMITTER_POINT) {
        return dRec.measure == MTS_MEASURE_DISCRETE ? 1.0f : 0.0f;
    }
    return envPdfDirection(e, e.toLocal.vec(dRec.d));
}

/* Scene::sampleEmitterDirect (src/librender/scene.cpp:828-852) */
V3 sampleEmitterDirect(MtsContext &c, DRec &dRec, V2 sample, bool testVisibility) {
    float emPdf;
    size_t index = cdfSampleReuse(c.emitterCdf, sample.x, &emPdf);
    const Emitter &e = c.emitters[index];
    V3 value = emitterSampleDirect(c, e, dRec, sample);
    if (dRec.pdf != 0) {
        if (testVisibility) {
            Hit h;
            if (sceneIntersect(c, dRec.ref, dRec.d, kEps, dRec.dist * (1 - kShadowEps), true, h)) return V3(0.0f);
        }
        dRec.emitter = (int)index;
        dRec.pdf *= emPdf;
        value = value / emPdf;
        return value;
    }
    return V3(0.0f);
}

struct Rng { /* PCG32; the reference uses SFMT (streams differ by design) */
    uint64_t state, inc;
    Rng(uint64_t seed, uint64_t seq) { state = 0; inc = (seq << 1) | 1; next(); state += seed; next(); }
    uint32_t next() {
        uint64_t old = state; state = old * 6364136223846793005ULL + inc;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((-rot) & 31));
    }
    float next1D() { return (next() >> 8) * (1.0f / 16777216.0f); }
    V2 next2D() { float a = next1D(); float b = next1D(); return V2{a, b}; }
};

/* PerspectiveCameraImpl::sampleRayDifferential (src/sensors/perspective.cpp:271-297) */
void sampleRay(const MtsCamera &cam, V2 samplePos, V3 &o, V3 &d, float &mint, float &maxt, V3 *rx = nullptr, V3 *ry = nullptr) {
    M44 s2c = m44_from(cam.sample_to_camera), c2w = m44_from(cam.camera_to_world);
    V3 nearP = s2c.point(V3(samplePos.x * cam.inv_resolution[0], samplePos.y * cam.inv_resolution[1], 0.0f));
    V3 dir = normalize(nearP);
    float invZ = 1.0f / dir.z;
    mint = cam.near_clip * invZ; maxt = cam.far_clip * invZ;
    o = c2w.pointAffine(V3(0.0f));
    d = c2w.vec(dir);
    if (rx) *rx = c2w.vec(normalize(nearP + V3(cam.dx[0], cam.dx[1], cam.dx[2])));
    if (ry) *ry = c2w.vec(normalize(nearP + V3(cam.dy[0], cam.dy[1], cam.dy[2])));
}

/* The offset rays of a camera ray after RayDifferential::scaleDifferential (ray.h:163-168; integrator.cpp:144-145,181) */
struct RayDiff { bool has = false; V3 o, rxD, ryD; };
RayDiff sampleRayDiff(const MtsCamera &cam, V2 samplePos, float scale) {
    RayDiff r; V3 d, rx, ry; float mint, maxt;
    sampleRay(cam, samplePos, r.o, d, mint, maxt, &rx, &ry);
    r.rxD = d + (rx - d) * scale; r.ryD = d + (ry - d) * scale; r.has = true;
    return r;
}
inline float diffScaleOf(const MtsRenderParams &P) {
    int n = P.sample_count > 0 ? P.sample_count : P.spp * (P.sample_stride > 0 ? P.sample_stride : 1);
    return 1.0f / std::sqrt((float)n);
}
inline float comp(V3 v, int a) { return a == 0 ? v.x : a == 1 ? v.y : v.z; }
/* Intersection::computePartials (src/librender/intersection.cpp:5-76) */
void computePartials(Its &its, const RayDiff &ray) {
    if (its.hasUVPartials || !ray.has) return;
    its.hasUVPartials = true;
    if (isZero(its.dpdu) && isZero(its.dpdv)) { its.dudx = its.dvdx = its.dudy = its.dvdy = 0.0f; return; }
    const V3 n = its.geo.n;
    const float pp = dot(n, its.p), pox = dot(n, ray.o), poy = dot(n, ray.o), prx = dot(n, ray.rxD), pry = dot(n, ray.ryD);
    if (prx == 0 || pry == 0) { its.dudx = its.dvdx = its.dudy = its.dvdy = 0.0f; return; }
    const float tx = (pp - pox) / prx, ty = (pp - poy) / pry;
    float absX = std::fabs(n.x), absY = std::fabs(n.y), absZ = std::fabs(n.z);
    int a0, a1;
    if (absX > absY && absX > absZ) { a0 = 1; a1 = 2; } else if (absY > absZ) { a0 = 0; a1 = 2; } else { a0 = 0; a1 = 1; }
    const float A00 = comp(its.dpdu, a0), A01 = comp(its.dpdv, a0), A10 = comp(its.dpdu, a1), A11 = comp(its.dpdv, a1);
    const V3 px = ray.o + ray.rxD * tx, py = ray.o + ray.ryD * ty;
    const float Bx0 = comp(px, a0) - comp(its.p, a0), Bx1 = comp(px, a1) - comp(its.p, a1);
    const float By0 = comp(py, a0) - comp(its.p, a0), By1 = comp(py, a1) - comp(its.p, a1);
    /* solveLinearSystem2x2 (src/libcore/util.cpp:527-539); RCPOVERFLOW = 2^-128 in single precision */
    const float det = A00 * A11 - A01 * A10;
    if (std::fabs(det) <= 2.93873587705571876e-39f) {
        its.dudx = 1; its.dvdx = 0;
        its.dudy = 1; /* intersection.cpp:72-74 assigns dudy twice (0, then 1); dvdy keeps its initial value */
        return;
    }
    const float inverse = 1.0f / det;
    its.dudx = (A11 * Bx0 - A01 * Bx1) * inverse; its.dvdx = (A00 * Bx1 - A10 * Bx0) * inverse;
    its.dudy = (A11 * By0 - A01 * By1) * inverse; its.dvdy = (A00 * By1 - A10 * By0) * inverse;
}

inline float miWeight(float pdfA, float pdfB) { pdfA *= pdfA; pdfB *= pdfB; return pdfA / (pdfA + pdfB); }

V3 emitterLe(const MtsContext &c, const Its &its, V3 d) { /* AreaLight::eval (area.cpp:104-109) */
    const Emitter &e = c.emitters[c.shapes[its.shape].emitter];
    if (dot(its.sh.n, d) <= 0) return V3(0.0f);
    return e.radiance;
}

/* MIPathTracer::Li (src/integrators/path/path.cpp:119-294) */
V3 Li(MtsContext &c, const MtsRenderParams &P, V3 o, V3 d, float mint, float maxt, Rng &rng, int &depthOut, float *alphaOut = nullptr,
      RayDiff ray = RayDiff()) {
    Its its; V3 L(0.0f); bool scattered = false;
    Hit h;
    bool valid = sceneIntersect(c, o, d, mint, maxt, false, h);
    if (alphaOut) *alphaOut = valid ? 1.0f : 0.0f;   /* RadianceQueryRecord::rayIntersect with EOpacity (records.inl:121-137) */
    if (valid) fillIts(c, o, d, h, its);
    V3 throughput(1.0f); float eta = 1.0f;
    int depth = 1; bool emittedOk = true;
    const Emitter *env = c.envIndex >= 0 ? &c.emitters[c.envIndex] : nullptr;
    while (depth <= P.max_depth || P.max_depth < 0) {
        if (!valid) {
            if (emittedOk && (!P.hide_emitters || scattered) && env)
                L += throughput * (ray.has ? envEvalDiff(*env, d, ray.rxD, ray.ryD) : envEval(*env, d));   /* path.cpp:136-143 */
            break;
        }
        const Shape &shape = c.shapes[its.shape];
        const int bsdf = shape.bsdf;
        /* Intersection::getBSDF(ray) (records.inl:69-75): partials only when the BSDF asks for ray differentials */
        if (c.mat.nodes[bsdf].flags & MTS_BF_RAY_DIFF) computePartials(its, ray);
        const uint32_t lobes = c.mat.nodes[bsdf].lobes;
        if (shape.emitter >= 0 && emittedOk && (!P.hide_emitters || scattered)) L += throughput * emitterLe(c, its, -d);
        if ((depth >= P.max_depth && P.max_depth > 0) ||
            (P.strict_normals && dot(d, its.geo.n) * Frame::cosTheta(its.wi) >= 0)) break;

        /* direct illumination sampling */
        DRec dRec; dRec.ref = its.p; dRec.refN = V3(0.0f); /* records.inl:160-164 */
        if ((lobes & (MTS_ETRANSMISSION | MTS_EBACKSIDE)) == 0) dRec.refN = its.sh.n;
        if (lobes & MTS_ESMOOTH) {
            V3 value = sampleEmitterDirect(c, dRec, rng.next2D(), true);
            if (!isZero(value)) {
                const Emitter &em = c.emitters[dRec.emitter];
                V3 wo = its.toLocal(dRec.d);
                const V3 bsdfVal = bsdfEval(c.mat, bsdf, its, its.wi, wo, MTS_MEASURE_SOLID_ANGLE);
                if (!isZero(bsdfVal) && (!P.strict_normals || dot(its.geo.n, dRec.d) * Frame::cosTheta(wo) > 0)) {
                    /* area and envmap emitters are "on surface"; the point emitter is not (emitter.h) */
                    float bsdfPdfV = (em.type != MTS_EMITTER_POINT && dRec.measure == MTS_MEASURE_SOLID_ANGLE)
                        ? bsdfPdf(c.mat, bsdf, its, its.wi, wo, MTS_MEASURE_SOLID_ANGLE) : 0;
                    float weight = miWeight(dRec.pdf, bsdfPdfV);
                    L += throughput * value * bsdfVal * weight;
                }
            }
        }

        /* BSDF sampling */
        BSample bs; bs.pdf = 0; bs.eta = 1; bs.type = 0;
        const V2 uBsdf = rng.next2D();
        if (lobes & MTS_EGLOSSY_TRANSMISSION) bs.u3 = rng.next1D(); /* roughdielectric.cpp:405 */
        V3 bsdfWeight = bsdfSample(c.mat, bsdf, its, its.wi, uBsdf, bs);
        if (isZero(bsdfWeight)) break;
        scattered |= bs.type != MTS_ENULL;
        const V3 wo = its.toWorld(bs.wo);
        float woDotGeoN = dot(its.geo.n, wo);
        if (P.strict_normals && woDotGeoN * Frame::cosTheta(bs.wo) <= 0) break;

        bool hitEmitter = false; V3 value;
        o = its.p; d = wo;
        ray.has = false;   /* `ray = Ray(its.p, wo, ray.time)` drops the differentials (path.cpp:225, ray.h:196-209) */
        valid = sceneIntersect(c, o, d, kEps, INFINITY, false, h);
        if (valid) {
            its = Its();
            fillIts(c, o, d, h, its);
            if (c.shapes[its.shape].emitter >= 0) {
                value = emitterLe(c, its, -d);
                /* dRec.setQuery (records.inl:170-178) */
                dRec.p = its.p; dRec.n = its.sh.n; dRec.measure = MTS_MEASURE_SOLID_ANGLE;
                dRec.emitter = c.shapes[its.shape].emitter; dRec.d = d; dRec.dist = its.t;
                hitEmitter = true;
            }
        } else {
            if (env) {
                if (P.hide_emitters && !scattered) break;
                value = envEval(*env, d);
                float nearT, farT; /* fillDirectSamplingRecord (envmap.cpp:358-374) */
                if (!bsphereIntersect(*env, o, d, nearT, farT) || nearT > 0 || farT < 0) break;
                dRec.p = o + d * farT; dRec.n = normalize(env->bsCenter - dRec.p);
                dRec.measure = MTS_MEASURE_SOLID_ANGLE; dRec.emitter = c.envIndex; dRec.d = d; dRec.dist = farT;
                hitEmitter = true;
            } else break;
        }
        throughput *= bsdfWeight;
        eta *= bs.eta;
        if (hitEmitter) {
            /* Scene::pdfEmitterDirect (scene.cpp:949-952) */
            float lumPdf = 0;
            if (!(bs.type & MTS_EDELTA)) {
                const Emitter &em = c.emitters[dRec.emitter];
                float emProb = c.emitterCdf[dRec.emitter + 1] - c.emitterCdf[dRec.emitter];
                lumPdf = emitterPdfDirect(c, em, dRec) * emProb;
            }
            L += throughput * value * miWeight(bs.pdf, lumPdf);
        }
        if (!valid) break;
        emittedOk = false; /* rRec.type = ERadianceNoEmission */
        if (depth++ >= P.rr_depth) {
            float q = std::min(maxc(throughput) * eta * eta, 0.95f);
            if (rng.next1D() >= q) break;
            throughput = throughput / q;
        }
    }
    depthOut = depth;
    return L;
}

/* ImageBlock::put (include/mitsuba/render/imageblock.h:144-204) on a border-less full-frame
 * accumulator; equivalent to per-block put + Film::put(block). */
void splat(const MtsFilm &f, float *buf, V2 samplePos, V3 value, float *alphaBuf = nullptr, float alpha = 1.0f) {
    float vals[4] = {value.x, value.y, value.z, 1.0f};
    for (int i = 0; i < 4; ++i) if (!std::isfinite(vals[i]) || vals[i] < 0) return;
    const float r = f.filter_radius;
    const float scaleFactor = MTS_FILTER_RESOLUTION / r;
    const float px = samplePos.x - 0.5f, py = samplePos.y - 0.5f;
    int minx = std::max((int)std::ceil(px - r), 0), miny = std::max((int)std::ceil(py - r), 0);
    int maxx = std::min((int)std::floor(px + r), f.width - 1), maxy = std::min((int)std::floor(py + r), f.height - 1);
    for (int y = miny; y <= maxy; ++y) {
        float wy = f.filter_table[std::min((int)std::fabs((y - py) * scaleFactor), MTS_FILTER_RESOLUTION)];
        for (int x = minx; x <= maxx; ++x) {
            float wx = f.filter_table[std::min((int)std::fabs((x - px) * scaleFactor), MTS_FILTER_RESOLUTION)];
            float w = wx * wy;
            float *dst = buf + ((size_t)y * f.width + x) * 4;
            for (int k = 0; k < 4; ++k) dst[k] += w * vals[k];
            if (alphaBuf) alphaBuf[(size_t)y * f.width + x] += w * alpha;
        }
    }
}

float triArea(V3 p0, V3 p1, V3 p2) { return 0.5f * length(cross(p1 - p0, p2 - p0)); } /* triangle.cpp surfaceArea */

} // namespace

/* ======================================================================== */
extern "C" {

MTS_API int32_t mtsoracle_abi_version(void) { return MTS_ABI_VERSION; }
MTS_API int mtsoracle_create_multi(const int32_t *, int32_t, MtsContext **) { return fail(MTS_EINVAL, "create_multi: devices are a CUDA-library concept"); }
MTS_API int mtsoracle_device_count(MtsContext *) { return 1; }
MTS_API int mtsoracle_trace_noalloc(MtsContext *, int64_t, const float *, int32_t, float *, int32_t *) { return fail(MTS_EINVAL, "trace_noalloc: device scratch memory is a CUDA-library concept"); }
MTS_API const char *mtsoracle_last_error(void) { return g_err.c_str(); }

MTS_API int mtsoracle_create(int, MtsContext **out) {
    if (!out) return fail(MTS_EINVAL, "null out");
    *out = new MtsContext();
    std::memset(&(*out)->stats, 0, sizeof(MtsStats));
    return MTS_OK;
}
MTS_API void mtsoracle_destroy(MtsContext *c) { delete c; }

MTS_API int mtsoracle_set_camera(MtsContext *c, const MtsCamera *cam) { c->cam = *cam; c->hasCam = true; return MTS_OK; }
MTS_API int mtsoracle_set_film(MtsContext *c, const MtsFilm *f) {
    if (f->width <= 0 || f->height <= 0 || !(f->filter_radius > 0)) return fail(MTS_EINVAL, "bad film");
    c->film = *f; c->hasFilm = true;
    c->filmBuf.assign((size_t)f->width * f->height * 4, 0.0f);
    c->alphaBuf.clear();
    if (f->has_alpha) c->alphaBuf.assign((size_t)f->width * f->height, 0.0f);
    return MTS_OK;
}
MTS_API int mtsoracle_add_texture(MtsContext *c, const MtsTextureDesc *d, const float *texels) {
    if (d->channels != 1 && d->channels != 3) return fail(MTS_EINVAL, "texture channels must be 1 or 3");
    if (d->filter < MTS_FILTER_NEAREST || d->filter > MTS_FILTER_EWA) return fail(MTS_EINVAL, "unknown texture filter type");
    Tex t; t.d = *d;
    std::vector<float> fromHalf;
    if (d->texel_format == MTS_TEXEL_F16) {   /* IEEE half input: widen (exact) */
        const uint16_t *h = reinterpret_cast<const uint16_t *>(texels);
        fromHalf.resize((size_t)d->width * d->height * d->channels);
        for (size_t i = 0; i < fromHalf.size(); ++i) {
            const uint32_t sign = (uint32_t)(h[i] & 0x8000u) << 16, e = (h[i] >> 10) & 0x1fu, m = h[i] & 0x3ffu;
            float v;
            if (e == 0) v = std::ldexp((float)m, -24);
            else if (e == 31) v = m ? NAN : INFINITY;
            else v = std::ldexp((float)(m | 0x400u), (int)e - 25);
            fromHalf[i] = sign ? -v : v;
        }
        texels = fromHalf.data();
    } else if (d->texel_format != MTS_TEXEL_F32) return fail(MTS_EINVAL, "unknown texel format");
    t.mip.build(texels, d->width, d->height, d->channels, d->wrap_u, d->wrap_v, d->filter, d->max_anisotropy, 1.0f /* TMIPMap maxValue default */);
    c->mat.textures.push_back(std::move(t));
    return (int)c->mat.textures.size() - 1;
}
MTS_API int mtsoracle_set_float_pool(MtsContext *c, const float *data, int64_t n) { c->mat.pool.assign(data, data + n); return MTS_OK; }
MTS_API int mtsoracle_set_bsdfs(MtsContext *c, const MtsBsdfNode *nodes, int32_t n) { c->mat.nodes.assign(nodes, nodes + n); return MTS_OK; }

MTS_API int mtsoracle_add_mesh(MtsContext *c, const MtsMeshDesc *m) {
    Shape s; s.type = 0; s.bsdf = m->bsdf;
    s.pos.resize(m->nverts);
    for (uint32_t i = 0; i < m->nverts; ++i) s.pos[i] = V3(m->positions[3 * i], m->positions[3 * i + 1], m->positions[3 * i + 2]);
    if (m->normals) { s.nrm.resize(m->nverts); for (uint32_t i = 0; i < m->nverts; ++i) s.nrm[i] = V3(m->normals[3 * i], m->normals[3 * i + 1], m->normals[3 * i + 2]); }
    if (m->texcoords) { s.uv.resize(m->nverts); for (uint32_t i = 0; i < m->nverts; ++i) s.uv[i] = V2{m->texcoords[2 * i], m->texcoords[2 * i + 1]}; }
    s.idx.assign(m->indices, m->indices + 3 * (size_t)m->ntris);
    if (m->tangents) { s.tang.resize(2 * (size_t)m->ntris); for (size_t i = 0; i < 2 * (size_t)m->ntris; ++i) s.tang[i] = V3(m->tangents[3 * i], m->tangents[3 * i + 1], m->tangents[3 * i + 2]); }
    c->shapes.push_back(std::move(s)); c->built = false;
    return (int)c->shapes.size() - 1;
}
MTS_API int mtsoracle_add_rectangle(MtsContext *c, const float tw[16], int32_t bsdf) {
    Shape s; s.type = 1; s.bsdf = bsdf;
    s.o2w = m44_from(tw); s.w2o = m44_inverse(s.o2w);
    /* Rectangle::configure (rectangle.cpp:102-113) */
    s.dpdu = s.o2w.vec(V3(2, 0, 0)); s.dpdv = s.o2w.vec(V3(0, 2, 0));
    const M44 &it = s.w2o; /* Transform::operator()(Normal): inverse transpose */
    V3 nn(it.m[2][0], it.m[2][1], it.m[2][2]);
    s.frame = Frame(normalize(s.dpdu), normalize(s.dpdv), normalize(nn));
    s.invArea = 1.0f / (length(s.dpdu) * length(s.dpdv));
    c->shapes.push_back(std::move(s)); c->built = false;
    return (int)c->shapes.size() - 1;
}
MTS_API int mtsoracle_add_sphere(MtsContext *c, const float o2w[16], float radius, int32_t flip, int32_t bsdf) {
    if (!(radius > 0)) return fail(MTS_EINVAL, "Cannot create spheres of radius <= 0");
    Shape s; s.type = 2; s.bsdf = bsdf; s.radius = radius; s.flip = flip != 0;
    s.o2w = m44_from(o2w); s.w2o = m44_inverse(s.o2w);
    s.center = s.o2w.point(V3(0, 0, 0));
    s.invArea = 1 / (4 * kPi * radius * radius);
    c->shapes.push_back(std::move(s)); c->built = false;
    return (int)c->shapes.size() - 1;
}
MTS_API int mtsoracle_add_emitter(MtsContext *c, const MtsEmitterDesc *d) {
    Emitter e; e.type = d->type; e.shape = d->shape; e.weight = d->sampling_weight; e.scale = d->scale;
    e.radiance = V3(d->radiance[0], d->radiance[1], d->radiance[2]);
    e.position = V3(d->position[0], d->position[1], d->position[2]);
    if (d->type == MTS_EMITTER_AREA) {
        if (d->shape < 0 || d->shape >= (int)c->shapes.size()) return fail(MTS_EINVAL, "area emitter: bad shape id");
        c->shapes[d->shape].emitter = (int)c->emitters.size();
    } else if (d->type == MTS_EMITTER_ENVMAP) {
        if (c->envIndex >= 0) return fail(MTS_EINVAL, "only one environment emitter is allowed");
        e.toWorld = m44_from(d->to_world); e.toLocal = m44_inverse(e.toWorld);
        e.w = d->env_width; e.h = d->env_height;
        /* envmap.cpp:138-182: EWA MIP map, repeat-u / clamp-v, maxAnisotropy 10, unbounded values */
        e.mip.build(d->env_texels, e.w, e.h, 3, MTS_WRAP_REPEAT, MTS_WRAP_CLAMP, MTS_FILTER_EWA, 10.0f, INFINITY);
        e.tex = e.mip.levels[0].tx;
        c->envIndex = (int)c->emitters.size();
    }
    c->emitters.push_back(std::move(e)); c->built = false;
    return (int)c->emitters.size() - 1;
}

MTS_API int mtsoracle_build(MtsContext *c) {
    auto t0 = std::chrono::steady_clock::now();
    if (c->emitters.empty()) return fail(MTS_EINVAL, "scene has no emitters (the reference would add a sunsky emitter, which is out of scope)");
    c->prims.clear(); c->nodes.clear();
    /* ShapeKDTree::build TriAccel table (src/librender/skdtree.cpp:80-106) */
    for (size_t si = 0; si < c->shapes.size(); ++si) {
        Shape &s = c->shapes[si];
        if (s.type == 0) {
            size_t nt = s.idx.size() / 3;
            for (size_t j = 0; j < nt; ++j) {
                TriAccel ta; std::memset(&ta, 0, sizeof(ta));
                triLoad(ta, s.pos[s.idx[3 * j]], s.pos[s.idx[3 * j + 1]], s.pos[s.idx[3 * j + 2]]);
                ta.shapeIndex = (uint32_t)si; ta.primIndex = (uint32_t)j;
                c->prims.push_back(ta);
            }
            if (s.emitter >= 0) { /* TriMesh::prepareSamplingTable (trimesh.cpp:388-403) */
                s.areaCdf.assign(1, 0.0f);
                for (size_t j = 0; j < nt; ++j)
                    s.areaCdf.push_back(s.areaCdf.back() + triArea(s.pos[s.idx[3 * j]], s.pos[s.idx[3 * j + 1]], s.pos[s.idx[3 * j + 2]]));
                float sum = s.areaCdf.back();
                float norm = 1.0f / sum;
                for (size_t j = 1; j < s.areaCdf.size(); ++j) s.areaCdf[j] *= norm;
                s.areaCdf.back() = 1.0f;
                s.invArea = 1.0f / sum;
            }
        } else {
            TriAccel ta; std::memset(&ta, 0, sizeof(ta));
            ta.shapeIndex = (uint32_t)si; ta.k = kNoTri;
            c->prims.push_back(ta);
        }
    }
    std::vector<PB> pb(c->prims.size());
    c->sceneLo = V3(INFINITY); c->sceneHi = V3(-INFINITY);
    for (size_t i = 0; i < pb.size(); ++i) {
        primBounds(*c, (int)i, pb[i].lo, pb[i].hi); pb[i].prim = (int)i;
        for (int a = 0; a < 3; ++a) {
            pb[i].c[a] = 0.5f * (pb[i].lo[a] + pb[i].hi[a]);
            c->sceneLo[a] = std::min(c->sceneLo[a], pb[i].lo[a]); c->sceneHi[a] = std::max(c->sceneHi[a], pb[i].hi[a]);
        }
    }
    if (!pb.empty()) {
        buildRec(*c, pb, 0, (int)pb.size());
        std::vector<TriAccel> ordered(pb.size());
        for (size_t i = 0; i < pb.size(); ++i) ordered[i] = c->prims[pb[i].prim];
        c->prims.swap(ordered);
    }
    /* emitter pmf (src/librender/scene.cpp:375-380) */
    c->emitterCdf.assign(1, 0.0f);
    for (auto &e : c->emitters) c->emitterCdf.push_back(c->emitterCdf.back() + e.weight);
    { float sum = c->emitterCdf.back(), norm = 1.0f / sum;
      for (size_t j = 1; j < c->emitterCdf.size(); ++j) c->emitterCdf[j] *= norm;
      c->emitterCdf.back() = 1.0f; }
    /* EnvironmentMap::configure CDFs (src/emitters/envmap.cpp:260-329) + bounding sphere (:331-341) */
    if (c->envIndex >= 0) {
        Emitter &e = c->emitters[c->envIndex];
        e.cdfCols.assign((size_t)(e.w + 1) * e.h, 0.0f); e.cdfRows.assign(e.h + 1, 0.0f); e.rowWeights.assign(e.h, 0.0f);
        size_t colPos = 0, rowPos = 0; float rowSum = 0.0f;
        e.cdfRows[rowPos++] = 0;
        for (int y = 0; y < e.h; ++y) {
            float colSum = 0;
            e.cdfCols[colPos++] = 0;
            for (int x = 0; x < e.w; ++x) {
                const float *p = &e.tex[((size_t)y * e.w + x) * 3];
                colSum += luminance(V3(p[0], p[1], p[2]));
                e.cdfCols[colPos++] = colSum;
            }
            float normalization = 1.0f / colSum;
            for (int x = 1; x < e.w; ++x) e.cdfCols[colPos - x - 1] *= normalization;
            e.cdfCols[colPos - 1] = 1.0f;
            float weight = std::sin((y + 0.5f) * kPi / e.h);
            e.rowWeights[y] = weight;
            rowSum += colSum * weight;
            e.cdfRows[rowPos++] = rowSum;
        }
        float normalization = 1.0f / rowSum;
        for (int y = 1; y < e.h; ++y) e.cdfRows[rowPos - y - 1] *= normalization;
        e.cdfRows[rowPos - 1] = 1.0f;
        if (rowSum == 0) return fail(MTS_EINVAL, "The environment map is completely black -- this is not allowed.");
        e.normalization = 1.0f / (rowSum * (2 * kPi / e.w) * (kPi / e.h));
        /* scene AABB incl. the sensor position (scene.cpp:386-398), AABB::getBSphere, radius*1.5 */
        V3 lo = c->sceneLo, hi = c->sceneHi;
        if (!pb.empty()) { /* the kd-tree enlarges its AABB (include/mitsuba/render/gkdtree.h:1218-1221, eps = 1e-3) */
            const float eps = 1e-3f;
            lo = lo - ((hi - lo) * eps + V3(eps));
            hi = hi + ((hi - lo) * eps + V3(eps));   /* uses the already enlarged minimum, as the reference does */
        }
        if (c->hasCam) {
            V3 camP = m44_from(c->cam.camera_to_world).pointAffine(V3(0.0f));
            for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], camP[a]); hi[a] = std::max(hi[a], camP[a]); }
        }
        if (pb.empty() && !c->hasCam) { lo = V3(0.0f); hi = V3(0.0f); }
        e.bsCenter = (lo + hi) * 0.5f;
        e.bsRadius = std::max(kEps, length(hi - e.bsCenter) * 1.5f);
    }
    c->built = true;
    c->stats.build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return MTS_OK;
}

MTS_API int mtsoracle_render(MtsContext *c, const MtsRenderParams *P, volatile const int *cancel, MtsStats *stats) {
    if (!c->built || !c->hasCam || !c->hasFilm) return fail(MTS_ESTATE, "render before build/set_camera/set_film");
    if (P->rr_depth <= 0) return fail(MTS_EINVAL, "'rrDepth' must be set to a value greater than zero!");
    if (P->max_depth <= 0 && P->max_depth != -1) return fail(MTS_EINVAL, "'maxDepth' must be set to -1 (infinite) or a value greater than zero!");
    const int W = c->film.width, H = c->film.height;
    if (P->clear_film) { std::fill(c->filmBuf.begin(), c->filmBuf.end(), 0.0f); std::fill(c->alphaBuf.begin(), c->alphaBuf.end(), 0.0f); }
    const bool hasAlpha = !c->alphaBuf.empty();
    c->normalRays = 0; c->shadowRays = 0;
    auto t0 = std::chrono::steady_clock::now();
    uint64_t pathLen = 0;
    int nthreads = 1;
#ifdef _OPENMP
    nthreads = omp_get_max_threads();
#endif
    std::vector<std::vector<float>> tfilm(nthreads), talpha(hasAlpha ? nthreads : 0);
    bool cancelled = false;
    const float diffScale = diffScaleOf(*P);
#pragma omp parallel reduction(+ : pathLen)
    {
        int tid = 0;
#ifdef _OPENMP
        tid = omp_get_thread_num();
#endif
        std::vector<float> &buf = tfilm[tid];
        buf.assign((size_t)W * H * 4, 0.0f);
        float *abuf = nullptr;
        if (hasAlpha) { talpha[tid].assign((size_t)W * H, 0.0f); abuf = talpha[tid].data(); }
        t_normalRays = t_shadowRays = 0;
#pragma omp for schedule(dynamic, 1)
        for (int y = 0; y < H; ++y) {
            if (cancel && *cancel) { cancelled = true; continue; }
            for (int x = 0; x < W; ++x) {
                Rng rng(P->seed * 0x9E3779B97F4A7C15ULL + (uint64_t)P->sample_offset, (uint64_t)y * W + x);
                for (int s = 0; s < P->spp; ++s) {
                    V2 u = rng.next2D();
                    V2 samplePos{x + u.x, y + u.y};
                    V3 o, d; float mint, maxt;
                    sampleRay(c->cam, samplePos, o, d, mint, maxt);
                    int depth; float alpha = 1.0f;
                    V3 L = Li(*c, *P, o, d, mint, maxt, rng, depth, hasAlpha ? &alpha : nullptr, sampleRayDiff(c->cam, samplePos, diffScale));
                    pathLen += depth;
                    splat(c->film, buf.data(), samplePos, L, abuf, alpha);
                }
            }
        }
        c->normalRays.fetch_add(t_normalRays); c->shadowRays.fetch_add(t_shadowRays);
        t_normalRays = t_shadowRays = 0;
    }
    /* ImageBlock merge (Film::put(block)): per-thread accumulators summed in thread order */
    const long long nf = (long long)c->filmBuf.size();
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < nf; ++i) { float acc = c->filmBuf[i]; for (auto &b : tfilm) if (!b.empty()) acc += b[i]; c->filmBuf[i] = acc; }
    for (auto &b : talpha) if (!b.empty()) for (size_t i = 0; i < b.size(); ++i) c->alphaBuf[i] += b[i];
    MtsStats &st = c->stats;
    st.camera_samples = (uint64_t)W * H * P->spp;
    st.normal_rays = c->normalRays; st.shadow_rays = c->shadowRays; st.path_length_sum = pathLen;
    st.render_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (stats) *stats = st;
    return cancelled ? fail(MTS_ECANCELLED, "cancelled") : MTS_OK;
}

MTS_API int mtsoracle_read_film(MtsContext *c, float *out) { std::memcpy(out, c->filmBuf.data(), c->filmBuf.size() * sizeof(float)); return MTS_OK; }
MTS_API int mtsoracle_read_film_alpha(MtsContext *c, float *out) {
    if (c->alphaBuf.empty()) return fail(MTS_ESTATE, "the film has no alpha plane (MtsFilm.has_alpha = 0)");
    std::memcpy(out, c->alphaBuf.data(), c->alphaBuf.size() * sizeof(float)); return MTS_OK;
}
MTS_API int mtsoracle_bind_film(MtsContext *, void *) { return fail(MTS_EINVAL, "bind_film: device memory is a CUDA-library concept"); }
MTS_API int mtsoracle_set_stream(MtsContext *, void *) { return fail(MTS_EINVAL, "set_stream: CUDA streams are a CUDA-library concept"); }
MTS_API int mtsoracle_get_stats(MtsContext *c, MtsStats *s) { *s = c->stats; return MTS_OK; }

MTS_API int mtsoracle_sample_rays(MtsContext *c, int64_t n, const float *sp, float *out) {
    if (!c->hasCam) return fail(MTS_ESTATE, "no camera");
    for (int64_t i = 0; i < n; ++i) {
        V3 o, d, rx, ry; float mint, maxt;
        sampleRay(c->cam, V2{sp[2 * i], sp[2 * i + 1]}, o, d, mint, maxt, &rx, &ry);
        float *q = out + 16 * i;
        q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = mint; q[4] = d.x; q[5] = d.y; q[6] = d.z; q[7] = maxt;
        q[8] = rx.x; q[9] = rx.y; q[10] = rx.z; q[11] = 0; q[12] = ry.x; q[13] = ry.y; q[14] = ry.z; q[15] = 0;
    }
    return MTS_OK;
}

MTS_API int mtsoracle_trace(MtsContext *c, int64_t n, const float *rays, int32_t shadow, float *hits, int32_t *ids) {
    if (!c->built) return fail(MTS_ESTATE, "trace before build");
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        Hit h; h.prim = -1;
        bool ok = sceneIntersect(*c, V3(r[0], r[1], r[2]), V3(r[4], r[5], r[6]), r[3], r[7], shadow != 0, h);
        float *q = hits + 4 * i;
        if (shadow) { q[0] = ok ? 0.0f : INFINITY; q[1] = q[2] = q[3] = 0; if (ids) { ids[2 * i] = ids[2 * i + 1] = -1; } continue; }
        if (ok) {
            q[0] = h.t; q[1] = h.u; q[2] = h.v; q[3] = 0;
            if (ids) { const TriAccel &ta = c->prims[h.prim]; ids[2 * i] = ta.shapeIndex; ids[2 * i + 1] = ta.k == kNoTri ? 0 : (int)ta.primIndex; }
        } else { q[0] = INFINITY; q[1] = q[2] = q[3] = 0; if (ids) { ids[2 * i] = ids[2 * i + 1] = -1; } }
    }
    return MTS_OK;
}

MTS_API int mtsoracle_hit_records(MtsContext *c, int64_t n, const float *rays, float *out, int32_t *ids) {
    if (!c->built) return fail(MTS_ESTATE, "hit_records before build");
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        V3 o(r[0], r[1], r[2]), d(r[4], r[5], r[6]);
        Hit h; float *q = out + 24 * i;
        if (!sceneIntersect(*c, o, d, r[3], r[7], false, h)) {
            for (int k = 0; k < 24; ++k) q[k] = 0;
            q[3] = INFINITY;
            if (ids) { ids[2 * i] = ids[2 * i + 1] = -1; }
            continue;
        }
        Its its; fillIts(*c, o, d, h, its);
        q[0] = its.p.x; q[1] = its.p.y; q[2] = its.p.z; q[3] = its.t;
        q[4] = its.geo.n.x; q[5] = its.geo.n.y; q[6] = its.geo.n.z; q[7] = its.uv.x;
        q[8] = its.sh.n.x; q[9] = its.sh.n.y; q[10] = its.sh.n.z; q[11] = its.uv.y;
        q[12] = its.dpdu.x; q[13] = its.dpdu.y; q[14] = its.dpdu.z; q[15] = 0;
        q[16] = its.dpdv.x; q[17] = its.dpdv.y; q[18] = its.dpdv.z; q[19] = 0;
        q[20] = its.wi.x; q[21] = its.wi.y; q[22] = its.wi.z; q[23] = 0;
        if (ids) { ids[2 * i] = its.shape; ids[2 * i + 1] = its.prim < 0 ? 0 : its.prim; }
    }
    return MTS_OK;
}

static void surfToIts(const float *s, Its &its) {
    its.uv = V2{s[0], s[1]};
    if (s[15] != 0.0f) { its.hasUVPartials = true; its.dudx = s[2]; its.dvdx = s[3]; its.dudy = s[7]; its.dvdy = s[11]; }
    its.dpdu = V3(s[4], s[5], s[6]); its.dpdv = V3(s[8], s[9], s[10]);
    V3 n(s[12], s[13], s[14]);
    its.geo = Frame(n);
    computeShadingFrame(n, its.dpdu, its.sh);
    its.p = V3(0.0f); its.t = 1; its.shape = -1; its.prim = -1;
}

MTS_API int mtsoracle_bsdf_eval_pdf(MtsContext *c, int32_t bsdf, int64_t n, const float *surf, const float *wi, const float *wo,
                                    int32_t measure, float *out_value, float *out_pdf) {
    if (bsdf < 0 || bsdf >= (int)c->mat.nodes.size()) return fail(MTS_EINVAL, "bad bsdf index");
    for (int64_t i = 0; i < n; ++i) {
        Its its; surfToIts(surf + 16 * i, its);
        V3 a(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), b(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
        its.wi = a;
        V3 v = bsdfEval(c->mat, bsdf, its, a, b, measure);
        out_value[3 * i] = v.x; out_value[3 * i + 1] = v.y; out_value[3 * i + 2] = v.z;
        out_pdf[i] = bsdfPdf(c->mat, bsdf, its, a, b, measure);
    }
    return MTS_OK;
}
MTS_API int mtsoracle_bsdf_sample(MtsContext *c, int32_t bsdf, int64_t n, const float *surf, const float *wi, const float *u, float *out) {
    if (bsdf < 0 || bsdf >= (int)c->mat.nodes.size()) return fail(MTS_EINVAL, "bad bsdf index");
    for (int64_t i = 0; i < n; ++i) {
        Its its; surfToIts(surf + 16 * i, its);
        V3 a(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
        its.wi = a;
        BSample bs; bs.wo = V3(0.0f); bs.pdf = 0; bs.eta = 1; bs.type = 0;
        { uint32_t xb, yb; std::memcpy(&xb, u + 2 * i, 4); std::memcpy(&yb, u + 2 * i + 1, 4); MTS_PROBE_U3(xb, yb, bs.u3); }
        V3 w = bsdfSample(c->mat, bsdf, its, a, V2{u[2 * i], u[2 * i + 1]}, bs);
        float *q = out + 12 * i;
        if (isZero(w)) { for (int k = 0; k < 12; ++k) q[k] = 0; continue; }
        q[0] = w.x; q[1] = w.y; q[2] = w.z; q[3] = bs.pdf; q[4] = bs.wo.x; q[5] = bs.wo.y; q[6] = bs.wo.z;
        q[7] = bs.eta; q[8] = (float)bs.type; q[9] = q[10] = q[11] = 0;
    }
    return MTS_OK;
}
MTS_API int mtsoracle_sample_emitter_direct(MtsContext *c, int64_t n, const float *ref, const float *refn, const float *u, float *out) {
    if (!c->built) return fail(MTS_ESTATE, "sample_emitter_direct before build");
    for (int64_t i = 0; i < n; ++i) {
        DRec dRec; dRec.ref = V3(ref[3 * i], ref[3 * i + 1], ref[3 * i + 2]); dRec.refN = V3(refn[3 * i], refn[3 * i + 1], refn[3 * i + 2]);
        dRec.pdf = 0; dRec.emitter = -1; dRec.measure = 0; dRec.dist = 0;
        V3 v = sampleEmitterDirect(*c, dRec, V2{u[2 * i], u[2 * i + 1]}, false);
        float *q = out + 16 * i;
        if (isZero(v) && dRec.pdf == 0) { for (int k = 0; k < 16; ++k) q[k] = 0; q[15] = -1; continue; }
        q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = dRec.pdf; q[4] = dRec.d.x; q[5] = dRec.d.y; q[6] = dRec.d.z; q[7] = dRec.dist;
        q[8] = dRec.p.x; q[9] = dRec.p.y; q[10] = dRec.p.z; q[11] = (float)dRec.measure; q[12] = dRec.n.x; q[13] = dRec.n.y; q[14] = dRec.n.z;
        q[15] = (float)dRec.emitter;
    }
    return MTS_OK;
}
MTS_API int mtsoracle_eval_environment(MtsContext *c, int64_t n, const float *d, float *out) {
    if (!c->built || c->envIndex < 0) return fail(MTS_ESTATE, "no environment emitter");
    const Emitter &e = c->emitters[c->envIndex];
    float emProb = c->emitterCdf[c->envIndex + 1] - c->emitterCdf[c->envIndex];
    for (int64_t i = 0; i < n; ++i) {
        V3 dir(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        V3 v = envEval(e, dir);
        out[4 * i] = v.x; out[4 * i + 1] = v.y; out[4 * i + 2] = v.z;
        out[4 * i + 3] = envPdfDirection(e, e.toLocal.vec(dir)) * emProb;
    }
    return MTS_OK;
}
MTS_API int mtsoracle_texture_eval(MtsContext *c, int32_t tex, int64_t n, const float *uv, const float *duv, float *out) {
    if (tex < 0 || tex >= (int)c->mat.textures.size()) return fail(MTS_EINVAL, "bad texture index");
    MtsTexRef r; r.scale[0] = r.scale[1] = r.scale[2] = 1.0f; r.tex = tex;
    for (int64_t i = 0; i < n; ++i) {
        Its its; its.uv = V2{uv[2 * i], uv[2 * i + 1]};
        if (duv) { its.hasUVPartials = true; its.dudx = duv[4 * i]; its.dvdx = duv[4 * i + 1]; its.dudy = duv[4 * i + 2]; its.dvdy = duv[4 * i + 3]; }
        V3 v = evalTex(c->mat, r, its);
        out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
    }
    return MTS_OK;
}
MTS_API int mtsoracle_hit_partials(MtsContext *c, int64_t n, const float *sp, int32_t sample_count, float *out) {
    if (!c->built) return fail(MTS_ESTATE, "hit_partials before build");
    if (!c->hasCam) return fail(MTS_ESTATE, "no camera set");
    if (sample_count <= 0) return fail(MTS_EINVAL, "sample_count must be positive");
    const float scale = 1.0f / std::sqrt((float)sample_count);
    for (int64_t i = 0; i < n; ++i) {
        V2 pos{sp[2 * i], sp[2 * i + 1]};
        V3 o, d; float mint, maxt;
        sampleRay(c->cam, pos, o, d, mint, maxt);
        float *q = out + 8 * i;
        Hit h;
        if (!sceneIntersect(*c, o, d, mint, maxt, false, h)) { for (int k = 0; k < 8; ++k) q[k] = 0; continue; }
        Its its; fillIts(*c, o, d, h, its);
        computePartials(its, sampleRayDiff(c->cam, pos, scale));
        q[0] = its.dudx; q[1] = its.dvdx; q[2] = its.dudy; q[3] = its.dvdy; q[4] = its.t; q[5] = 1.0f; q[6] = its.uv.x; q[7] = its.uv.y;
    }
    return MTS_OK;
}
MTS_API int mtsoracle_eval_environment_camera(MtsContext *c, int64_t n, const float *sp, int32_t sample_count, float *out) {
    if (!c->built || c->envIndex < 0) return fail(MTS_ESTATE, "the scene has no environment emitter");
    if (!c->hasCam) return fail(MTS_ESTATE, "no camera set");
    if (sample_count <= 0) return fail(MTS_EINVAL, "sample_count must be positive");
    const float scale = 1.0f / std::sqrt((float)sample_count);
    const Emitter &e = c->emitters[c->envIndex];
    for (int64_t i = 0; i < n; ++i) {
        V2 pos{sp[2 * i], sp[2 * i + 1]};
        V3 o, d; float mint, maxt;
        sampleRay(c->cam, pos, o, d, mint, maxt);
        RayDiff rd = sampleRayDiff(c->cam, pos, scale);
        V3 v = envEvalDiff(e, d, rd.rxD, rd.ryD);
        out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
    }
    return MTS_OK;
}
MTS_API int mtsoracle_splat(MtsContext *c, int64_t n, const float *pos, const float *value, int32_t clear_first) {
    if (!c->hasFilm) return fail(MTS_ESTATE, "no film");
    if (clear_first) std::fill(c->filmBuf.begin(), c->filmBuf.end(), 0.0f);
    for (int64_t i = 0; i < n; ++i)
        splat(c->film, c->filmBuf.data(), V2{pos[2 * i], pos[2 * i + 1]}, V3(value[3 * i], value[3 * i + 1], value[3 * i + 2]));
    return MTS_OK;
}

} // extern "C"
