// geometry.cuh -- ray/primitive tests, software BVH traversal (B200 has no RT cores) and the
// reconstruction of the full hit record from (primitive slot, barycentrics).
#pragma once
#include "scene.cuh"

namespace mts {

struct RayHit { float t, u, v; int prim; };  // prim = slot in the leaf-ordered PrimRecord array, -1 = miss

// Rectangle::rayIntersect (src/shapes/rectangle.cpp:125-148); u,v receive the local xy hit position
DEVFN bool rectIntersect(const DShape &s, float3 o, float3 d, float mint, float maxt, float &t, float &lx, float &ly) {
    float3 lo = xfPoint(s.w2o, o), ld = xfVec(s.w2o, d);
    float hit = -lo.z / ld.z;
    if (!(hit >= mint && hit <= maxt)) return false;
    float x = lo.x + ld.x * hit, y = lo.y + ld.y * hit;
    if (fabsf(x) <= 1 && fabsf(y) <= 1) { t = hit; lx = x; ly = y; return true; }
    return false;
}
// Sphere::rayIntersect (src/shapes/sphere.cpp:163-207): the quadratic is solved in fp64 like the reference
DEVFN bool sphereIntersect(const DShape &s, float3 o, float3 d, float mint, float maxt, float &t) {
    double ox = (double)o.x - (double)s.center.x, oy = (double)o.y - (double)s.center.y, oz = (double)o.z - (double)s.center.z;
    double dx = d.x, dy = d.y, dz = d.z;
    double A = dx * dx + dy * dy + dz * dz;
    double B = 2 * (ox * dx + oy * dy + oz * dz);
    double C = ox * ox + oy * oy + oz * oz - (double)(s.radius * s.radius);
    double nearT, farT;
    if (!solveQuadraticDouble(A, B, C, nearT, farT)) return false;
    if (!(nearT <= (double)maxt && farT >= (double)mint)) return false;
    if (nearT < (double)mint) {
        if (farT > (double)maxt) return false;
        t = (float)farT;
    } else t = (float)nearT;
    return true;
}

// One primitive of a leaf.  Triangles use the Moeller-Trumbore test on the 48-byte (v0,e1,e2)
// record (the reference's TriAccel is the same 48 bytes, include/mitsuba/render/triaccel.h:37-51);
// u,v are the barycentrics of v1,v2 exactly as TriAccel::rayIntersect returns them.
DEVFN bool primIntersect(const DScene &S, int slot, float3 o, float3 d, float mint, float maxt, RayHit &h) {
    const float4 *rec = reinterpret_cast<const float4 *>(S.prims + slot);
    float4 p0 = __ldg(rec), p1 = __ldg(rec + 1), p2 = __ldg(rec + 2);
    int type = __float_as_int(p2.w);
    if (type == 0) {
        float3 e1 = f3(p1), e2 = f3(p2);
        float3 pv = cross(d, e2);
        float det = dot(e1, pv);
        if (det == 0.0f) return false;
        float inv = 1.0f / det;
        float3 tv = o - f3(p0);
        float u = dot(tv, pv) * inv;
        float3 qv = cross(tv, e1);
        float v = dot(d, qv) * inv;
        float t = dot(e2, qv) * inv;
        if (!(u >= 0.0f && v >= 0.0f && u + v <= 1.0f)) return false;
        if (t < mint || t > maxt) return false;
        h.t = t; h.u = u; h.v = v; h.prim = slot;
        return true;
    }
    const DShape &s = S.shapes[__float_as_int(p0.w)];
    float t, lx = 0, ly = 0;
    bool ok = (type == 1) ? rectIntersect(s, o, d, mint, maxt, t, lx, ly) : sphereIntersect(s, o, d, mint, maxt, t);
    if (ok) { h.t = t; h.u = lx; h.v = ly; h.prim = slot; }
    return ok;
}

DEVFN float2 slab(float lox, float loy, float loz, float hix, float hiy, float hiz, float3 o, float3 inv, float tmin, float tmax) {
    float tx0 = (lox - o.x) * inv.x, tx1 = (hix - o.x) * inv.x;
    float ty0 = (loy - o.y) * inv.y, ty1 = (hiy - o.y) * inv.y;
    float tz0 = (loz - o.z) * inv.z, tz1 = (hiz - o.z) * inv.z;
    float tn = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), tmin));
    float tf = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), tmax));
    return make_float2(tn, tf);
}

struct TraceCounters { unsigned nodes, prims; };

// Stack-based while-while traversal of the two-wide BVH.  ANY = shadow ray (first hit terminates).
// The 1.0000004 slack on the far distance keeps boundary hits from being culled by the box test.
template <bool ANY, bool COUNT>
DEVFN bool traverse(const DScene &S, float3 o, float3 d, float mint, float maxt, RayHit &best, TraceCounters *ctr) {
    best.prim = -1; best.t = maxt;
    float3 inv = f3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[64];
    int sp = 0;
    int node = 0;
    while (true) {
        const float4 *np = reinterpret_cast<const float4 *>(S.bvh + node);
        float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2);
        int4 link = __ldg(reinterpret_cast<const int4 *>(np + 3));
        if (COUNT) ctr->nodes++;
        float2 t0 = slab(a.x, a.y, a.z, a.w, b.x, b.y, o, inv, mint, best.t);
        float2 t1 = slab(b.z, b.w, c.x, c.y, c.z, c.w, o, inv, mint, best.t);
        bool h0 = t0.x <= t0.y * 1.0000004f, h1 = t1.x <= t1.y * 1.0000004f;
        int next = -1;
        // leaves are intersected immediately, inner children are ordered near-first
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            bool hk = k ? h1 : h0;
            int cnt = k ? link.w : link.z;
            int first = k ? link.y : link.x;
            if (hk && cnt > 0) {
                for (int i = first; i < first + cnt; ++i) {
                    if (COUNT) ctr->prims++;
                    RayHit h;
                    if (primIntersect(S, i, o, d, mint, best.t, h)) {
                        if (ANY) { best = h; return true; }
                        if (h.t < best.t || best.prim < 0) best = h;
                    }
                }
            }
        }
        bool in0 = h0 && link.z == 0 && link.x >= 0, in1 = h1 && link.w == 0 && link.y >= 0;  // link < 0: empty child
        if (in0 && in1) {
            bool firstIs0 = t0.x <= t1.x;
            next = firstIs0 ? link.x : link.y;
            if (sp < 64) stack[sp++] = firstIs0 ? link.y : link.x;
        } else if (in0) next = link.x;
        else if (in1) next = link.y;
        if (next < 0) {
            if (sp == 0) break;
            next = stack[--sp];
        }
        node = next;
    }
    return best.prim >= 0;
}

// ShapeKDTree::rayIntersect epsilon handling (src/librender/skdtree.cpp:112-142 closest hit,
// :207-226 shadow rays): a ray whose mint is exactly Epsilon gets the adaptive scale
// max(|o|_inf, Epsilon) (closest hit) or |o|_inf without the clamp (shadow).
DEVFN float adaptiveMinT(float3 o, float mint, bool shadow) {
    if (mint == kEps) {
        float m = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
        if (!shadow) m = fmaxf(m, kEps);
        mint *= m;
    }
    return mint;
}
template <bool ANY, bool COUNT>
DEVFN bool sceneIntersect(const DScene &S, float3 o, float3 d, float mint, float maxt, RayHit &h, TraceCounters *ctr = nullptr) {
    float rayMinT = adaptiveMinT(o, mint, ANY);
    h.prim = -1; h.t = kInf; h.u = h.v = 0;
    if (S.numPrims == 0 || !(maxt > rayMinT)) return false;
    return traverse<ANY, COUNT>(S, o, d, rayMinT, maxt, h, ctr);
}

DEVFN float3 loadF3(const float *base, uint32_t i) {
    return f3(__ldg(base + 3 * (size_t)i), __ldg(base + 3 * (size_t)i + 1), __ldg(base + 3 * (size_t)i + 2));
}

// fillIntersectionRecord<true> (include/mitsuba/render/skdtree.h:343-428) + the per-shape
// fillIntersectionRecord of rectangle.cpp:155-168 and sphere.cpp:209-255.
DEVFN void fillHit(const DScene &S, float3 o, float3 d, const RayHit &rh, Hit &its) {
    const float4 *rec = reinterpret_cast<const float4 *>(S.prims + rh.prim);
    float4 p0 = __ldg(rec), p1 = __ldg(rec + 1), p2 = __ldg(rec + 2);
    int type = __float_as_int(p2.w);
    its.shape = __float_as_int(p0.w);
    its.t = rh.t;
    const DShape &s = S.shapes[its.shape];
    if (type == 0) {
        uint32_t tri = (uint32_t)__float_as_int(p1.w);
        its.prim = (int)tri;
        size_t gt = (size_t)s.triOffset + tri;
        uint32_t i0 = __ldg(S.indices + 3 * gt), i1 = __ldg(S.indices + 3 * gt + 1), i2 = __ldg(S.indices + 3 * gt + 2);
        float bx = 1 - rh.u - rh.v, by = rh.u, bz = rh.v;
        float3 q0 = loadF3(S.positions, i0), q1 = loadF3(S.positions, i1), q2 = loadF3(S.positions, i2);
        its.p = q0 * bx + q1 * by + q2 * bz;
        float3 side1 = q1 - q0, side2 = q2 - q0;
        float3 faceNormal = cross(side1, side2);
        float len = length(faceNormal);
        if (!isZero(faceNormal)) faceNormal = faceNormal / len;
        if (s.flags & 4) {
            its.dpdu = loadF3(S.tangents, 2 * (uint32_t)gt);
            its.dpdv = loadF3(S.tangents, 2 * (uint32_t)gt + 1);
        } else { its.dpdu = side1; its.dpdv = side2; }
        float3 shN;
        if (s.flags & 1) {
            shN = normalize(loadF3(S.normals, i0) * bx + loadF3(S.normals, i1) * by + loadF3(S.normals, i2) * bz);
            if (dot(faceNormal, shN) < 0) faceNormal = -faceNormal;
        } else shN = faceNormal;
        its.geoN = faceNormal;
        if (s.flags & 2) {
            float2 t0 = make_float2(__ldg(S.texcoords + 2 * (size_t)i0), __ldg(S.texcoords + 2 * (size_t)i0 + 1));
            float2 t1 = make_float2(__ldg(S.texcoords + 2 * (size_t)i1), __ldg(S.texcoords + 2 * (size_t)i1 + 1));
            float2 t2 = make_float2(__ldg(S.texcoords + 2 * (size_t)i2), __ldg(S.texcoords + 2 * (size_t)i2 + 1));
            its.uv = make_float2(t0.x * bx + t1.x * by + t2.x * bz, t0.y * bx + t1.y * by + t2.y * bz);
        } else its.uv = make_float2(by, bz);
        its.sh = shadingFrame(shN, its.dpdu);
    } else if (type == 1) {
        its.prim = 0;
        its.geoN = f3(s.normal);
        its.dpdu = f3(s.dpdu); its.dpdv = f3(s.dpdv);
        its.uv = make_float2(0.5f * (rh.u + 1), 0.5f * (rh.v + 1));
        its.p = o + d * rh.t;
        its.sh = shadingFrame(its.geoN, its.dpdu);
    } else {
        its.prim = 0;
        float3 center = f3(s.center);
        float3 p = o + d * rh.t;
        p = center + normalize(p - center) * s.radius;  // re-projection, sphere.cpp:213-216
        its.p = p;
        float3 local = xfVec(s.w2o, p - center);
        float theta = safe_acos(local.z / s.radius);
        float phi = atan2f(local.y, local.x);
        if (phi < 0) phi += 2 * kPi;
        its.uv = make_float2(phi * (0.5f * kInvPi), theta * kInvPi);
        its.dpdu = xfVec(s.o2w, f3(-local.y, local.x, 0) * (2 * kPi));
        float3 gn = normalize(p - center);
        float zrad = sqrtf(local.x * local.x + local.y * local.y);
        float cosPhi = 0, sinPhi = 1;
        if (zrad > 0) { float invZRad = 1.0f / zrad; cosPhi = local.x * invZRad; sinPhi = local.y * invZRad; }
        its.dpdv = xfVec(s.o2w, f3(local.z * cosPhi, local.z * sinPhi, -sinf(theta) * s.radius) * kPi);
        if (s.flip) gn = -gn;
        its.geoN = gn;
        its.sh = shadingFrame(gn, its.dpdu);
    }
    its.wi = its.sh.toLocal(-d);
}

// PerspectiveCameraImpl::sampleRayDifferential (src/sensors/perspective.cpp:271-297)
DEVFN void cameraRay(const MtsCamera &cam, float2 samplePos, float3 &o, float3 &d, float &mint, float &maxt,
                     float3 *rx = nullptr, float3 *ry = nullptr) {
    const float *m = cam.sample_to_camera;
    float px = samplePos.x * cam.inv_resolution[0], py = samplePos.y * cam.inv_resolution[1];
    float x = m[0] * px + m[1] * py + m[3], y = m[4] * px + m[5] * py + m[7];
    float z = m[8] * px + m[9] * py + m[11], w = m[12] * px + m[13] * py + m[15];
    float3 nearP = f3(x, y, z);
    if (w != 1.0f) nearP = nearP / w;
    float3 dir = normalize(nearP);
    float invZ = 1.0f / dir.z;
    mint = cam.near_clip * invZ; maxt = cam.far_clip * invZ;
    const float *c = cam.camera_to_world;
    o = f3(c[3], c[7], c[11]);
    auto toWorld = [&](float3 v) {
        return f3(c[0] * v.x + c[1] * v.y + c[2] * v.z, c[4] * v.x + c[5] * v.y + c[6] * v.z, c[8] * v.x + c[9] * v.y + c[10] * v.z);
    };
    d = toWorld(dir);
    if (rx) *rx = toWorld(normalize(nearP + f3(cam.dx[0], cam.dx[1], cam.dx[2])));
    if (ry) *ry = toWorld(normalize(nearP + f3(cam.dy[0], cam.dy[1], cam.dy[2])));
}

} // namespace mts
