// scene.cuh -- device-side scene layout (the data layout in HBM; see DESIGN.md "Data layout").
#pragma once
#include "common.cuh"

namespace mts {

// Bitmap texture = TMIPMap (include/mitsuba/render/mipmap.h): half-precision texels like the reference stores them;
// 3-channel texels are padded to half4 so one bilinear tap is one aligned 8-byte load.  Level 0 starts at `texels`;
// ewa / trilinear textures carry the whole pyramid (level l at texel offset levelOffset[l], row-major).
struct DTexture {
    const __half *texels;
    int width, height, channels;  // channels: 1 or 3 (3 -> stored as 4 halves)
    int wrapU, wrapV, filter;
    float2 uvScale, uvOffset;
    int levels;                   // 1 for nearest / bilinear
    float maxAnisotropy;
    const float *weightLut;       // EWA Gaussian weight table, MTS_MIPMAP_LUT_SIZE entries (mipmap.h:296-302)
    uint32_t levelOffset[MTS_MAX_MIP_LEVELS];
    uint16_t levelW[MTS_MAX_MIP_LEVELS], levelH[MTS_MAX_MIP_LEVELS];
};

constexpr int kGuideRows = 256, kGuideCols = 64;
enum { SHAPE_MESH = 0, SHAPE_RECT = 1, SHAPE_SPHERE = 2 };

struct DShape {
    int type, bsdf, emitter, flip;
    uint32_t sortKey;                             // material sort key of the root BSDF (1 + rank in shade-class order)
    uint32_t pad0[3];
    // mesh: offsets into the concatenated vertex / triangle arrays
    uint32_t vtxOffset, triOffset, ntris, flags;  // flags: 1 normals, 2 uvs, 4 tangents
    uint32_t cdfOffset;                           // area CDF (ntris+1 floats) when the mesh is an emitter
    float invArea;
    float radius;
    float pad;
    Affine o2w, w2o;                              // rectangle / sphere
    float4 dpdu, dpdv, normal, center;            // rectangle frame / sphere centre
};

struct DEmitter {
    int type, shape;
    float3 radiance;   // area radiance / point intensity
    float3 position;
    float prob;        // discrete selection probability (Scene::m_emitterPDF)
};

struct DEnvmap {
    const __half *texels;    // half4 per texel
    const float *cdfRows, *cdfCols, *rowWeights;
    // guide tables: guide[k] = lower_bound(cdf, k / G); a lookup narrows the binary search of a sample in
    // [k/G, (k+1)/G) to [guide[k], guide[k+1]] (same result, ~4x fewer dependent loads)
    const uint16_t *guideRows, *guideCols;   // (kGuideRows + 1), height * (kGuideCols + 1)
    int width, height;
    float normalization, scale;
    Affine toWorld, toLocal;
    float3 bsCenter; float bsRadius;
    int emitterIndex;
    float prob;
    const DTexture *mip;     // (device memory) the same texels as a MIP map (repeat-u / clamp-v, EWA, maxAnisotropy 10; envmap.cpp:138-182):
                             // camera rays carry differentials and are looked up filtered (:391-407)
};

// Eight-wide BVH node with quantised child boxes, 80 bytes = five 16-byte loads per visit (the layout of
// Ylitie, Karras, Laine: "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs", HPG 2017).
//   n0   origin p.xyz of the node box; w = ex | ey << 8 | ez << 16 | imask << 24
//        (e* = biased exponents of the per-axis grid scale 2^e, imask bit s = slot s holds an inner node)
//   n1   x = index of the first child node, y = slot of the first primitive of the leaf children,
//        z,w = meta[0..3], meta[4..7]: inner child in slot s -> 001 | (24 + s); leaf -> unary prim count (1..3) in
//        the top three bits | offset from n1.y in the low five; empty slot -> 0
//   n2   qlo.x[0..7], qlo.y[0..7]   n3   qlo.z[0..7], qhi.x[0..7]   n4   qhi.y[0..7], qhi.z[0..7]   (one byte per slot)
// Slot s lies towards (s&1 ? +x : -x, s&2 ? +y : -y, s&4 ? +z : -z) of the node centre, so that slot ^ rayOctant
// orders the children front to back without sorting.
struct WideNode {
    float4 n0;
    uint4 n1, n2, n3, n4;
};

// Triangle test of the trace kernels (geometry.cuh primIntersect), which also decides what a triangle's record holds:
//   0  conservative Moeller-Trumbore: record = (v0, e1, e2); a hit is accepted when the barycentrics lie inside the
//      triangle enlarged by the fp32 error bound of their own evaluation -- neighbouring triangles overlap by that
//      margin, so no ray slips between them (a ray in the margin hits both, at the same distance)
//   1  Woop / Benthin / Wald 2013: record = (v0, v1, v2); sign-consistent edge functions on the shared vertices (default)
// Measured on cfg3 (profiles/r2z_tritest.txt): form 0 is 3.4 % faster on the frame, but 1.7 % of the rays then report the
// NEIGHBOUR of the reference's triangle (same distance, inside the overlap margin) against 1.3e-4 with form 1.
#ifndef TRI_TEST_WOOP
#define TRI_TEST_WOOP 1
#endif
// 48-byte primitive record in leaf order.
//   triangle: p0 = (v0, shape), p1 = (e1 = v1-v0 | v1, prim index), p2 = (e2 = v2-v0 | v2, 0)   (see TRI_TEST_WOOP)
//   analytic: p0.w = shape, p2.w = 1 (rectangle) / 2 (sphere)
struct PrimRecord { float4 p0, p1, p2; };

// What the BSDF code reads of the scene.  Passed BY VALUE (three pointers in registers) into the BSDF functions: a
// reference to the kernel-parameter DScene handed to a non-inlined function would make the compiler copy the whole
// scene descriptor into every thread's local memory.
struct Mats {
    const MtsBsdfNode *nodes;
    const DTexture *textures;
    const float *pool;
};

struct DScene : Mats {
    int numNodes, numTextures;
    // geometry
    const DShape *shapes; int numShapes;
    const float *positions;   // 3 floats / vertex
    const float *normals;     // 3 floats / vertex (valid where the shape has normals)
    const float *texcoords;   // 2 floats / vertex
    const float4 *vtxAttr;    // the same normals and texcoords interleaved, 32 bytes = one sector per vertex: (nx, ny, nz, u), (v, 0, 0, 0);
                              // what the shade kernels gather (built by mtscuda_build; the separate arrays stay for the emitter code)
    const uint32_t *indices;  // 3 / triangle, already offset into the global vertex arrays
    const float *tangents;    // 6 floats / triangle
    const float *areaCdfs;
    const WideNode *bvh; const PrimRecord *prims; int numPrims, numNodesBvh;
    // emitters
    const DEmitter *emitters; const float *emitterCdf; int numEmitters;
    DEnvmap env; int hasEnv;
    // sensor + film
    MtsCamera cam;
    float4 *film; int filmW, filmH;
    float *filmAlpha;   // sum w * alpha, only for films with an alpha channel (else nullptr)
    float filterRadius; float filterTable[MTS_FILTER_RESOLUTION + 1];
};

// Full hit record (subset of `Intersection`, include/mitsuba/render/shape.h:36-172)
struct Hit {
    float3 p; float t;
    float3 geoN;
    Frame sh;
    float2 uv;
    float3 dpdu, dpdv;
    float3 wi;
    int shape, prim;
};

} // namespace mts
