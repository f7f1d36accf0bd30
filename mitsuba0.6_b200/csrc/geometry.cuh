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

// Per-ray constants of the watertight triangle test (Woop, Benthin, Wald 2013): the permutation that makes z the dominant
// axis of the direction, and the shear that maps the direction onto +z.
struct RayShear { int kx, ky, kz; float Sx, Sy, Sz; };
DEVFN float comp3(float3 v, int a) { return a == 0 ? v.x : a == 1 ? v.y : v.z; }
DEVFN RayShear rayShear(float3 d) {
    RayShear r;
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    r.kz = (ax > ay && ax > az) ? 0 : (ay > az ? 1 : 2);
    r.kx = r.kz == 2 ? 0 : r.kz + 1; r.ky = r.kx == 2 ? 0 : r.kx + 1;
    const float dz = comp3(d, r.kz);
    if (dz < 0.0f) { const int t = r.kx; r.kx = r.ky; r.ky = t; }   // preserve the winding
    r.Sz = 1.0f / dz; r.Sx = comp3(d, r.kx) * r.Sz; r.Sy = comp3(d, r.ky) * r.Sz;
    return r;
}
// One primitive of a leaf.  Triangles: the 48-byte record holds the three vertices (the reference's TriAccel is the same
// 48 bytes, include/mitsuba/render/triaccel.h:37-51); the test is the watertight one -- the scaled barycentrics U, V, W
// are 2-D edge functions of the sheared vertices, each computed from two vertices only and with separately rounded
// products, so the two triangles that share an edge get the exact negative of each other's value and a ray can never
// slip between them.  (The Moeller-Trumbore test used in round 1 leaked 6e-5 of the camera rays through the
// 9,999,392-triangle terrain: found by tests/test_gpu_fullsize.py against the reference's kd-tree.)  u, v are the
// barycentrics of v1, v2 exactly as TriAccel::rayIntersect returns them.
DEVFN bool primIntersect(const DScene &S, int slot, float3 o, float3 d, const RayShear &rs, float mint, float maxt, RayHit &h) {
    const float4 *rec = reinterpret_cast<const float4 *>(S.prims + slot);
    float4 p0 = __ldg(rec), p1 = __ldg(rec + 1), p2 = __ldg(rec + 2);
    int type = __float_as_int(p2.w);
#if !TRI_TEST_WOOP
    if (type == 0) {
        // Moeller-Trumbore on (v0, e1, e2), made conservative: u, v and 1 - u - v may be negative by `tol`, an upper bound
        // of the fp32 error of their own evaluation (products of |tv| ~ distance and |e| ~ triangle size divided by det:
        // 1e-4 ... 1e-3 for a 0.009-unit triangle seen from 15 units).  The two triangles of an edge then overlap by at
        // least the error either of them makes, and a ray can not pass between them.
        const float3 e1 = f3(p1), e2 = f3(p2);
        const float3 pv = cross(d, e2);
        const float det = dot(e1, pv);
        if (det == 0.0f) return false;
        const float inv = 1.0f / det;
        const float3 tv = o - f3(p0);
        const float u = dot(tv, pv) * inv;
        const float3 qv = cross(tv, e1);
        const float v = dot(d, qv) * inv;
        const float m1 = fabsf(e1.x) + fabsf(e1.y) + fabsf(e1.z), m2 = fabsf(e2.x) + fabsf(e2.y) + fabsf(e2.z);
        const float mt = fabsf(tv.x) + fabsf(tv.y) + fabsf(tv.z), me = fmaxf(m1, m2);
        const float tol = fminf(1.9e-6f * me * (mt + 2.0f * me) * fabsf(inv), 1e-2f);     // 32 eps * |e| * (|tv| + 2 |e|) / |det|
        if (!(u >= -tol && v >= -tol && u + v <= 1.0f + 2.0f * tol)) return false;
        const float t = dot(e2, qv) * inv;
        if (!(t >= mint && t <= maxt)) return false;
        h.t = t; h.u = u; h.v = v; h.prim = slot;
        return true;
    }
#else
    if (type == 0) {
        const float3 A = f3(p0) - o, B = f3(p1) - o, C = f3(p2) - o;
        const float Akz = comp3(A, rs.kz), Bkz = comp3(B, rs.kz), Ckz = comp3(C, rs.kz);
        const float Ax = comp3(A, rs.kx) - rs.Sx * Akz, Ay = comp3(A, rs.ky) - rs.Sy * Akz;
        const float Bx = comp3(B, rs.kx) - rs.Sx * Bkz, By = comp3(B, rs.ky) - rs.Sy * Bkz;
        const float Cx = comp3(C, rs.kx) - rs.Sx * Ckz, Cy = comp3(C, rs.ky) - rs.Sy * Ckz;
        float U = __fsub_rn(__fmul_rn(Cx, By), __fmul_rn(Cy, Bx));
        float V = __fsub_rn(__fmul_rn(Ax, Cy), __fmul_rn(Ay, Cx));
        float W = __fsub_rn(__fmul_rn(Bx, Ay), __fmul_rn(By, Ax));
        // an exact zero (the ray passes through an edge or a vertex in fp32) counts as inside for BOTH triangles of the edge:
        // a double hit at the same distance instead of Woop's double-precision tie-break -- never a leak
        if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
        const float det = U + V + W;
        if (det == 0.0f) return false;
        const float inv = 1.0f / det;
        const float t = (U * (rs.Sz * Akz) + V * (rs.Sz * Bkz) + W * (rs.Sz * Ckz)) * inv;
        if (!(t >= mint && t <= maxt)) return false;
        h.t = t; h.u = V * inv; h.v = W * inv; h.prim = slot;
        return true;
    }
#endif
    const DShape &s = S.shapes[__float_as_int(p0.w)];
    float t, lx = 0, ly = 0;
    bool ok = (type == 1) ? rectIntersect(s, o, d, mint, maxt, t, lx, ly) : sphereIntersect(s, o, d, mint, maxt, t);
    if (ok) { h.t = t; h.u = lx; h.v = ly; h.prim = slot; }
    return ok;
}

struct TraceCounters { unsigned nodes, prims; };

DEVFN uint32_t byteOf(uint32_t v, int j) { return __byte_perm(v, 0u, 0x4440u | j); }   // one PRMT
// 1 + (byte j of v) * 2^-15, exact, in one PRMT: the byte is spliced into mantissa bits 8..15 of 1.0f.  The constant
// offset is folded into the plane equation below, so decoding a quantised plane costs no floating-point instruction.
// `one` holds the bits of 1.0f in a register the compiler cannot fold (see wideNodeHits), so that the selector is
// the PRMT immediate; with a literal 1.0f ptxas spends a MOV per PRMT on the selector instead.
DEVFN float byteToUnit(uint32_t v, int j, uint32_t one) { return __uint_as_float(__byte_perm(v, one, 0x7604u | (j << 4))); }
// Blackwell packed fp32 FMA (FFMA2): d = a * (b, b) + c on a register pair -- entry and exit distance of one child
// slab in one instruction.
DEVFN float2 ffma2(float qn, float qf, float b, float2 c) {
    unsigned long long ua, ub, uc, ud;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ua) : "f"(qn), "f"(qf));
    asm("mov.b64 %0, {%1, %1};" : "=l"(ub) : "f"(b));
    asm("mov.b64 %0, {%1, %2};" : "=l"(uc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(ud) : "l"(ua), "l"(ub), "l"(uc));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(ud));
    return d;
}

// Intersects the ray with the eight quantised child boxes of one wide node and returns the hit mask:
// bits 24..31 = inner children ordered front to back (bit 24 + (slot ^ oct)), bits 0..23 = primitives of the
// leaf children that were hit (offsets from the node's first primitive).
//
// A child plane sits at origin + q * 2^e (q = 0..255), i.e. at ray distance t = q * a + b with a = 2^e / d and
// b = (origin - o) / d.  With u = 1 + q * 2^-15 (byteToUnit) this is t = u * A + (b - A), A = a * 2^15.  The
// subtraction b - A rounds at the magnitude of A (128 node extents), an error of at most 2^-9 of one
// quantisation step; the entry planes are therefore moved out by delta = |A| * 2^-22 towards the ray origin and the
// exit planes away from it, which keeps the test conservative (watertight) at no per-child cost.
DEVFN uint32_t wideNodeHits(const float4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4,
                            float3 o, float3 idir, uint32_t oct, float tmin, float tmax) {
    const uint32_t ew = (uint32_t)__float_as_int(n0.w);
    // 2^(e + 15); the builder keeps e + 15 below the exponent range (scene extents are nowhere near 2^112)
    const float sx = __int_as_float((int)(((ew & 0xffu) + 15u) << 23)), sy = __int_as_float((int)((((ew >> 8) & 0xffu) + 15u) << 23)),
                sz = __int_as_float((int)((((ew >> 16) & 0xffu) + 15u) << 23));
    const float Ax = sx * idir.x, Ay = sy * idir.y, Az = sz * idir.z;
    const float bx = fmaf(n0.x - o.x, idir.x, -Ax), by = fmaf(n0.y - o.y, idir.y, -Ay), bz = fmaf(n0.z - o.z, idir.z, -Az);
    const float kSlack = 2.3841858e-7f;  // 2^-22
    const float dx = fabsf(Ax) * kSlack, dy = fabsf(Ay) * kSlack, dz = fabsf(Az) * kSlack;
    const float2 Bx = make_float2(bx - dx, bx + dx), By = make_float2(by - dy, by + dy), Bz = make_float2(bz - dz, bz + dz);
    const uint32_t oct4 = oct * 0x01010101u;
    const float tmaxS = tmax * 1.0000004f;   // slack: boundary hits are never culled by the box test
    const uint32_t one = 0x3F800000u | (n1.x >> 31);   // child indices stay below 2^31: always the bits of 1.0f
    uint32_t hitmask = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const uint32_t meta4 = half ? n1.w : n1.z;
        const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t innerMask4 = (isInner4 >> 4) * 0xffu;
        const uint32_t bitIndex4 = (meta4 ^ (oct4 & innerMask4)) & 0x1f1f1f1fu;
        const uint32_t childBits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
        const uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
        // the plane the ray enters through depends on the sign of the direction
        const uint32_t nx = idir.x < 0 ? qhix : qlox, fx = idir.x < 0 ? qlox : qhix;
        const uint32_t ny = idir.y < 0 ? qhiy : qloy, fy = idir.y < 0 ? qloy : qhiy;
        const uint32_t nz = idir.z < 0 ? qhiz : qloz, fz = idir.z < 0 ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 tx = ffma2(byteToUnit(nx, j, one), byteToUnit(fx, j, one), Ax, Bx);
            const float2 ty = ffma2(byteToUnit(ny, j, one), byteToUnit(fy, j, one), Ay, By);
            const float2 tz = ffma2(byteToUnit(nz, j, one), byteToUnit(fz, j, one), Az, Bz);
            float cmin = fmaxf(fmaxf(tx.x, ty.x), fmaxf(tz.x, tmin));
            float cmax = fminf(fminf(tx.y, ty.y), fminf(tz.y, tmaxS));
            if (cmin <= cmax) hitmask |= byteOf(childBits4, j) << byteOf(bitIndex4, j);
        }
    }
    return hitmask;
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

// Stack-based traversal of the eight-wide compressed BVH, one ray per thread, as a resumable state machine so
// that the queue kernels can refill idle lanes of a warp between steps.  The state is a node group (first
// child index + pending front-to-back hit bits + inner mask); groups that still hold work are pushed when a
// nearer child is entered.  ANY = shadow ray (the first hit terminates).
#ifndef NODE_ITERS
#define NODE_ITERS 4
#endif
constexpr int kNodeIters = NODE_ITERS;

// Traversal stack: the first SH entries of every thread live in shared memory (column `threadIdx.x` of a
// [SH][kTraceBlock] uint2 array: conflict-free whatever the lanes' stack depths are), deeper entries spill to a
// per-thread local array.  A purely local stack costs one 32-byte sector per LANE per push/pop (the lanes of a warp
// sit at different depths), which made the stack more than half of the trace kernels' L1 traffic
// (profiles/r1e_ncu_full_raw.txt: 1.6 useful bytes per local sector).
#ifndef TRACE_SHEAR_IN_STATE
#define TRACE_SHEAR_IN_STATE 0
#endif
#ifndef TRACE_SMEM_NODES
#define TRACE_SMEM_NODES 0
#endif
#ifndef TRACE_SMEM_STACK
#define TRACE_SMEM_STACK 10
#endif
constexpr int kSmemStack = TRACE_SMEM_STACK;
constexpr int kTraceBlock = 128;                   // block size of every kernel that uses the shared stack
constexpr int kStackCapacity = 32;
DEVFN void stsStack(uint32_t addr, uint2 v) { asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(addr), "r"(v.x), "r"(v.y) : "memory"); }
DEVFN uint2 ldsStack(uint32_t addr) { uint2 v; asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory"); return v; }

template <bool ANY, bool COUNT, int SH = 0>
struct Traversal {
    float3 o, d, idir;
    float mint;
    uint32_t oct;
    RayHit best;
    uint2 ngroup;
    int sp;
#if TRACE_SHEAR_IN_STATE
    RayShear shear;
#endif
    uint32_t smBase;                               // shared-memory address of this thread's stack column (SH > 0)
    const float4 *topNodes = nullptr;              // per-CTA shared-memory copy of the top of the tree (TRACE_SMEM_NODES > 0)
    uint2 stack[kStackCapacity - SH];

    DEVFN void push(uint2 v) {
        if (SH > 0 && sp < SH) stsStack(smBase + (uint32_t)sp * (kTraceBlock * 8u), v);
        else stack[sp - SH] = v;
        ++sp;
    }
    DEVFN uint2 pop() {
        --sp;
        if (SH > 0 && sp < SH) return ldsStack(smBase + (uint32_t)sp * (kTraceBlock * 8u));
        return stack[sp - SH];
    }

    // returns false when the ray is finished before it starts (empty scene / empty interval)
    DEVFN bool init(const DScene &S, float3 o_, float3 d_, float rayMinT, float rayMaxT) {
        o = o_; d = d_;
        mint = adaptiveMinT(o, rayMinT, ANY);
        best.prim = -1; best.t = rayMaxT; best.u = best.v = 0;
        if (S.numPrims == 0 || !(rayMaxT > mint)) { best.t = kInf; return false; }
        const float tiny = 1e-24f;   // keeps 2^15 * extent / d finite (wideNodeHits)
        idir = f3(1.0f / (fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x)),
                  1.0f / (fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y)),
                  1.0f / (fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z)));
        oct = (idir.x >= 0 ? 1u : 0u) | (idir.y >= 0 ? 2u : 0u) | (idir.z >= 0 ? 4u : 0u);
        sp = 0;
        ngroup = make_uint2(0u, 0x80000000u);
#if TRACE_SHEAR_IN_STATE
        shear = rayShear(d);
#endif
        return true;
    }

    // "while-while" step: visit nodes until one yields primitive hits (at most kNodeIters visits, so that the lanes of
    // a warp reach the primitive tests together instead of trickling in one by one), then test those primitives.
    // Returns true when the traversal has finished.
    DEVFN bool step(const DScene &S, TraceCounters *ctr) {
        uint32_t pending = 0, primBase = 0;
#pragma unroll 1
        for (int it = 0; it < kNodeIters; ++it) {
            // enter the nearest pending inner child of the current node group
            const int bit = 31 - __clz((int)ngroup.y);
            ngroup.y &= ~(1u << bit);
            const uint32_t slot = (uint32_t)(bit - 24) ^ oct;
            const uint32_t rel = __popc(ngroup.y & 0xffu & ~(0xffffffffu << slot));
            const uint32_t nodeIdx = ngroup.x + rel;
            if ((ngroup.y & 0xff000000u) && sp < kStackCapacity) push(ngroup);
            float4 n0; uint4 n1, n2, n3, n4;
#if TRACE_SMEM_NODES > 0
            // experiment (north_star: "BVH nodes staged through shared memory"): the first TRACE_SMEM_NODES nodes -- the top
            // levels, stored breadth first -- are read from a per-CTA copy in shared memory.  Measured: no gain, see DESIGN 4.
            if (nodeIdx < (uint32_t)TRACE_SMEM_NODES && topNodes) {
                const float4 *sp4 = topNodes + 5 * nodeIdx;
                n0 = sp4[0]; n1 = *reinterpret_cast<const uint4 *>(sp4 + 1); n2 = *reinterpret_cast<const uint4 *>(sp4 + 2);
                n3 = *reinterpret_cast<const uint4 *>(sp4 + 3); n4 = *reinterpret_cast<const uint4 *>(sp4 + 4);
            } else
#endif
            {
                const float4 *np = reinterpret_cast<const float4 *>(S.bvh + nodeIdx);
                n0 = __ldg(np);
                n1 = __ldg(reinterpret_cast<const uint4 *>(np + 1)); n2 = __ldg(reinterpret_cast<const uint4 *>(np + 2));
                n3 = __ldg(reinterpret_cast<const uint4 *>(np + 3)); n4 = __ldg(reinterpret_cast<const uint4 *>(np + 4));
            }
            if (COUNT) ctr->nodes++;
            const uint32_t hits = wideNodeHits(n0, n1, n2, n3, n4, o, idir, oct, mint, best.t);
            ngroup = make_uint2(n1.x, (hits & 0xff000000u) | ((uint32_t)__float_as_int(n0.w) >> 24));
            pending = hits & 0x00ffffffu; primBase = n1.y;
            if (pending) break;
            if ((ngroup.y & 0xff000000u) == 0) {
                if (sp == 0) return true;
                ngroup = pop();
            }
        }
#if TRACE_SHEAR_IN_STATE
        const RayShear rs = shear;
#elif TRI_TEST_WOOP
        const RayShear rs = rayShear(d);   // recomputed per leaf batch (a few instructions) rather than held in registers across node visits
#else
        RayShear rs; rs.kx = rs.ky = rs.kz = 0; rs.Sx = rs.Sy = rs.Sz = 0;   // unused by the Moeller-Trumbore form
#endif
        while (pending) {
            const int j = __ffs((int)pending) - 1;
            pending &= pending - 1;
            if (COUNT) ctr->prims++;
            RayHit h;
            if (primIntersect(S, (int)(primBase + j), o, d, rs, mint, best.t, h)) {
                if (ANY) { best = h; return true; }
                if (h.t < best.t || best.prim < 0) best = h;
            }
        }
        if ((ngroup.y & 0xff000000u) == 0) {
            if (sp == 0) return true;
            ngroup = pop();
        }
        return false;
    }
    DEVFN bool hit() const { return best.prim >= 0; }
    DEVFN RayHit result() const { RayHit r = best; if (r.prim < 0) { r.t = kInf; r.u = r.v = 0; } return r; }
};

template <bool ANY, bool COUNT>
DEVFN bool sceneIntersect(const DScene &S, float3 o, float3 d, float mint, float maxt, RayHit &h, TraceCounters *ctr = nullptr) {
    Traversal<ANY, COUNT> tr;
    if (tr.init(S, o, d, mint, maxt)) while (!tr.step(S, ctr)) { }
    h = tr.result();
    return tr.hit();
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
        // All per-vertex / per-triangle attribute arrays are index-aligned with the positions (zeros where a mesh has
        // none), so every gather is issued unconditionally and back to back -- one memory round trip -- and the
        // shape flags only select what is used.
        // Sectors per hit: the vertex positions come with the primitive record already loaded above (bit-identical copies),
        // normals + texcoords from the interleaved array (one 32-byte sector per vertex instead of three 12 / 12 / 8-byte
        // gathers from separate arrays).
#if TRI_TEST_WOOP
        float3 q0 = f3(p0), q1 = f3(p1), q2 = f3(p2);
#else
        float3 q0 = loadF3(S.positions, i0), q1 = loadF3(S.positions, i1), q2 = loadF3(S.positions, i2);
#endif
        const float4 a0 = __ldg(S.vtxAttr + 2 * (size_t)i0), a1 = __ldg(S.vtxAttr + 2 * (size_t)i1), a2 = __ldg(S.vtxAttr + 2 * (size_t)i2);
        const float v0 = __ldg(reinterpret_cast<const float *>(S.vtxAttr + 2 * (size_t)i0 + 1)), v1 = __ldg(reinterpret_cast<const float *>(S.vtxAttr + 2 * (size_t)i1 + 1)),
                    v2 = __ldg(reinterpret_cast<const float *>(S.vtxAttr + 2 * (size_t)i2 + 1));
        float3 n0 = f3(a0), n1 = f3(a1), n2 = f3(a2);
        float2 t0 = make_float2(a0.w, v0), t1 = make_float2(a1.w, v1), t2 = make_float2(a2.w, v2);
        float3 tgu = loadF3(S.tangents, 2 * (uint32_t)gt), tgv = loadF3(S.tangents, 2 * (uint32_t)gt + 1);
        its.p = q0 * bx + q1 * by + q2 * bz;
        float3 side1 = q1 - q0, side2 = q2 - q0;
        float3 faceNormal = cross(side1, side2);
        float len = length(faceNormal);
        if (!isZero(faceNormal)) faceNormal = faceNormal / len;
        if (s.flags & 4) { its.dpdu = tgu; its.dpdv = tgv; }
        else { its.dpdu = side1; its.dpdv = side2; }
        float3 shN;
        if (s.flags & 1) {
            shN = normalize(n0 * bx + n1 * by + n2 * bz);
            if (dot(faceNormal, shN) < 0) faceNormal = -faceNormal;
        } else shN = faceNormal;
        its.geoN = faceNormal;
        if (s.flags & 2) its.uv = make_float2(t0.x * bx + t1.x * by + t2.x * bz, t0.y * bx + t1.y * by + t2.y * bz);
        else its.uv = make_float2(by, bz);
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
