// mtscuda.cu -- libmtscuda: context management, scene upload, the wavefront render loop and the
// C ABI of include/mtscuda.h.  sm_100a only; there is no CPU path in this library.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <mutex>
#include <thread>
#include <vector>
#include "bvh_build.h"
#include "devmem.h"
#include "wavefront.cuh"
#include "shade_launch.h"

using namespace mts;

namespace {

thread_local std::string g_err;
constexpr int kPollRing = 8;   // bounces in flight between a queue-size copy and the host looking at it
int fail(int code, const std::string &msg) { g_err = msg; return code; }

#define CUDA_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) \
    return fail(MTS_ECUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); } while (0)

struct HostShape {
    DShape d;
    bool isEmitter = false;
    uint32_t nverts = 0;  // mesh
    float lo[3], hi[3];  // analytic bounds
};

struct HostTex { MtsTextureDesc d; DTexture dt; size_t bytes = 0; };  // dt.texels: the MIP pyramid, built and kept on the device

// Growable device array: scene buffers are appended straight from the caller's host pointers (no host staging
// copy), so the plugin-facing upload runs at PCIe speed.
template <class T> struct DevVec {
    T *data = nullptr; size_t size = 0, cap = 0;
    cudaError_t reserve(size_t n, cudaStream_t st) {
        if (n <= cap) return cudaSuccess;
        size_t ncap = std::max(n, cap + cap / 2);
        T *nd = nullptr;
        cudaError_t e = devAlloc(&nd, sizeof(T) * ncap, st);
        if (e != cudaSuccess) return e;
        if (size) e = cudaMemcpyAsync(nd, data, sizeof(T) * size, cudaMemcpyDeviceToDevice, st);
        devFree(data, st);   // stream-ordered: released after the copy above
        data = nd; cap = ncap;
        return e;
    }
    // append n elements from host memory (or zeros when host == nullptr)
    cudaError_t append(const T *host, size_t n, cudaStream_t st) {
        cudaError_t e = reserve(size + n, st);
        if (e != cudaSuccess || n == 0) return e;
        if (host) e = cudaMemcpyAsync(data + size, host, sizeof(T) * n, cudaMemcpyDefault, st);   // host or (UVA) device memory
        else e = cudaMemsetAsync(data + size, 0, sizeof(T) * n, st);
        size += n;
        return e;
    }
    void release(cudaStream_t st) { devFree(data, st); data = nullptr; size = cap = 0; }
};

struct HostEmitter {
    MtsEmitterDesc d;
    float *envRaw = nullptr;   // device copy of the caller's rgb float texels (freed after build)
};

// double-precision 4x4 helpers for the one-off host set-up
struct M4 { double m[4][4]; };
M4 m4From(const float *p) { M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = p[4 * i + j]; return r; }
M4 m4Inverse(const M4 &a) {
    double w[4][8];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { w[i][j] = a.m[i][j]; w[i][j + 4] = (i == j); }
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        for (int r = c + 1; r < 4; ++r) if (std::fabs(w[r][c]) > std::fabs(w[piv][c])) piv = r;
        for (int j = 0; j < 8; ++j) std::swap(w[c][j], w[piv][j]);
        double d = w[c][c];
        for (int j = 0; j < 8; ++j) w[c][j] /= d;
        for (int r = 0; r < 4; ++r) if (r != c) { double f = w[r][c]; for (int j = 0; j < 8; ++j) w[r][j] -= f * w[c][j]; }
    }
    M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = w[i][j + 4];
    return r;
}
Affine toAffine(const M4 &a) {
    Affine r;
    r.r0 = make_float4((float)a.m[0][0], (float)a.m[0][1], (float)a.m[0][2], (float)a.m[0][3]);
    r.r1 = make_float4((float)a.m[1][0], (float)a.m[1][1], (float)a.m[1][2], (float)a.m[1][3]);
    r.r2 = make_float4((float)a.m[2][0], (float)a.m[2][1], (float)a.m[2][2], (float)a.m[2][3]);
    return r;
}
void hPoint(const Affine &m, const float p[3], float out[3]) {
    out[0] = m.r0.x * p[0] + m.r0.y * p[1] + m.r0.z * p[2] + m.r0.w;
    out[1] = m.r1.x * p[0] + m.r1.y * p[1] + m.r1.z * p[2] + m.r1.w;
    out[2] = m.r2.x * p[0] + m.r2.y * p[1] + m.r2.z * p[2] + m.r2.w;
}
void hVec(const Affine &m, const float p[3], float out[3]) {
    out[0] = m.r0.x * p[0] + m.r0.y * p[1] + m.r0.z * p[2];
    out[1] = m.r1.x * p[0] + m.r1.y * p[1] + m.r1.z * p[2];
    out[2] = m.r2.x * p[0] + m.r2.y * p[1] + m.r2.z * p[2];
}
float hLen(const float v[3]) { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }
float hLum(const float *p) { return p[0] * 0.212671f + p[1] * 0.715160f + p[2] * 0.072169f; }

} // namespace

struct MtsContext {
    int device = 0;
    cudaStream_t stream = nullptr, ownStream = nullptr;
    // Texture uploads (H2D copy, half conversion, MIP pyramid) run on a stream of their own and are joined by build():
    // the copy engine then works on the 0.9 GB of texels of cfg3 while the SMs build the BVH (e2e -30 ms per frame).
    cudaStream_t upStream = nullptr;
    cudaEvent_t upDone = nullptr;
    bool upPending = false;
    int numSMs = 148;
    bool deferEwa = true;      // staged shade: EWA texture walks as a kernel of their own (MTS_EWA_INLINE=1: inside k_shade_surface)
    bool raySort = false;      // MTS_RAY_SORT=1: bin the rays of bounce >= 1 by octant and origin cell before k_intersect
    float sceneLo[3] = {0, 0, 0}, sceneHi[3] = {1, 1, 1};   // bounds of the geometry (BVH root)
    bool leafTextured = false; // a leaf-class material of the built scene reads a bitmap texture
    bool stagedShade = true;   // leaf class as k_shade_surface (+ k_shade_ewa) + k_shade_scatter; default: where its materials are textured
                               // (MTS_SHADE_FUSED=1 / MTS_SHADE_STAGED=1 in the environment force one form)
    size_t l2Bytes = 0, l2WindowBytes = 0, maxWindow = 0; float l2HitRatio = 0;   // persisting-L2 set-up (0 = off)
    // staged scene
    MtsCamera cam; bool hasCam = false;
    MtsFilm film; bool hasFilm = false;
    std::vector<MtsBsdfNode> nodes;
    std::vector<float> pool;
    std::vector<HostTex> textures;
    std::vector<HostShape> shapes;
    DevVec<float> positions, normals, texcoords, tangents;   // concatenated over all meshes, on the device
    DevVec<uint32_t> indices;                                 // already offset into the global vertex arrays
    int *errFlag = nullptr;                                   // device flag written by validation kernels
    std::vector<HostEmitter> emitters;
    int envIndex = -1;
    // device scene
    DScene ds;
    std::vector<void *> sceneAllocs;
    bool built = false;
    bool peerFilm = false;     // peer of a multi-device group: film / alpha plane from cudaMalloc, readable by the primary over NVLink
    float4 *filmInternal = nullptr, *filmBound = nullptr;
    float *filmAlpha = nullptr;                               // sum w * alpha (films with an alpha channel only)
    uint32_t *hostPoll = nullptr;                             // pinned: next-queue sizes polled without blocking
    std::vector<cudaEvent_t> eventPool; size_t eventsUsed = 0;
    int bvhDepth = 0;
    float *ewaLut = nullptr;                                  // EWA weight table shared by all MIP maps
    uint32_t numKeys = 1, classKey[SHADE_CLASSES + 1] = {1, 1, 1, 1};   // material sort keys (set by build)
    // multi-device group (mtscuda_create_multi): this context is the primary; every peer is a full single-device context on
    // another GPU of the box that receives a copy of the scene over NVLink and renders its share of the samples
    std::vector<MtsContext *> peers;
    // wavefront
    Wavefront wf; std::vector<void *> wfAllocs; uint32_t wfCapacity = 0;
    MtsStats stats;

    MtsContext() { std::memset(&ds, 0, sizeof(ds)); std::memset(&wf, 0, sizeof(wf)); std::memset(&stats, 0, sizeof(stats)); }
    void freeScene() { for (void *p : sceneAllocs) devFree(p, stream); sceneAllocs.clear(); built = false; }
    void freeGeometry() {
        if (upStream && upPending) { cudaStreamSynchronize(upStream); upPending = false; }   // texels may still be in flight
        positions.release(stream); normals.release(stream); texcoords.release(stream); tangents.release(stream); indices.release(stream);
        for (auto &t : textures) devFree((void *)t.dt.texels, stream);
        textures.clear();
        devFree(errFlag, stream);
        errFlag = nullptr;
        for (auto &e : emitters) { devFree(e.envRaw, stream); e.envRaw = nullptr; }
    }
    void freeWavefront() { for (void *p : wfAllocs) devFree(p, stream); wfAllocs.clear(); wfCapacity = 0; }
    float4 *filmPtr() const { return filmBound ? filmBound : filmInternal; }
    void freeFilm() {
        if (peerFilm) { if (filmInternal || filmAlpha) cudaStreamSynchronize(stream); cudaFree(filmInternal); cudaFree(filmAlpha); }
        else { devFree(filmInternal, stream); devFree(filmAlpha, stream); }
        filmInternal = nullptr; filmAlpha = nullptr;
    }
};

namespace {

template <class T> int devUpload(MtsContext *c, const T *host, size_t n, const T **out) {
    *out = nullptr;
    if (n == 0) return MTS_OK;
    T *d = nullptr;
    CUDA_TRY(devAlloc(&d, sizeof(T) * n, c->stream));
    c->sceneAllocs.push_back(d);
    CUDA_TRY(cudaMemcpyAsync(d, host, sizeof(T) * n, cudaMemcpyHostToDevice, c->stream));
    *out = d;
    return MTS_OK;
}

// Wrapper levels above the deepest leaf below node i; >= 1000 for malformed trees (bad child index, cycle, bad mixture table).
int bsdfDepth(const std::vector<MtsBsdfNode> &nodes, const std::vector<float> &pool, int i, int guard = 0) {
    if (i < 0 || i >= (int)nodes.size() || guard > 64) return 1000;
    const MtsBsdfNode &n = nodes[i];
    if (n.type == MTS_BSDF_COATING || n.type == MTS_BSDF_BUMPMAP || n.type == MTS_BSDF_MASK || n.type == MTS_BSDF_ROUGHCOATING || n.type == MTS_BSDF_NORMALMAP)
        return 1 + bsdfDepth(nodes, pool, n.child[0], guard + 1);
    if (n.type == MTS_BSDF_BLEND || n.type == MTS_BSDF_TWOSIDED)
        return 1 + std::max(bsdfDepth(nodes, pool, n.child[0], guard + 1), bsdfDepth(nodes, pool, n.child[1], guard + 1));
    if (n.type == MTS_BSDF_MIXTURE) {
        const int cnt = n.lut[1], off = n.lut[0];
        if (cnt < 1 || cnt > MTS_MIXTURE_MAX || off < 0 || (size_t)off + 3 * (size_t)cnt + 1 > pool.size()) return 1000;
        int d = 0;
        for (int k = 0; k < cnt; ++k) d = std::max(d, bsdfDepth(nodes, pool, (int)pool[off + k], guard + 1));
        return 1 + d;
    }
    return 0;
}

const char *kDepthError = "BSDF nesting deeper than 4 wrapper levels (coating/blendbsdf/bumpmap/twosided/mask/mixturebsdf/roughcoating/normalmap) is not supported";

int ensureWavefront(MtsContext *c, uint32_t cap) {
    const uint32_t numKeys = c->numKeys;
    for (int k = 0; k <= SHADE_CLASSES; ++k) c->wf.classKey[k] = c->classKey[k];
    const bool needRec = c->stagedShade && c->classKey[SHADE_LEAF] < c->classKey[SHADE_LEAF + 1];
    if (c->wfCapacity >= cap && c->wf.numKeys == numKeys && needRec == (c->wf.rec != nullptr) && (needRec && c->deferEwa) == (c->wf.ewaJobs != nullptr) &&
        c->raySort == (c->wf.raySortBin != nullptr)) return MTS_OK;
    c->freeWavefront();
    Wavefront &w = c->wf;
    auto alloc = [&](void **p, size_t bytes) -> cudaError_t {
        cudaError_t e = devAlloc(p, bytes, c->stream);
        if (e == cudaSuccess) c->wfAllocs.push_back(*p);
        return e;
    };
    CUDA_TRY(alloc((void **)&w.rayO, sizeof(float4) * cap)); CUDA_TRY(alloc((void **)&w.rayD, sizeof(float4) * cap));
    CUDA_TRY(alloc((void **)&w.hit, sizeof(float4) * cap)); CUDA_TRY(alloc((void **)&w.thr, sizeof(float4) * cap));
    CUDA_TRY(alloc((void **)&w.L, sizeof(float4) * cap)); CUDA_TRY(alloc((void **)&w.meta, sizeof(uint4) * cap));
    CUDA_TRY(alloc((void **)&w.shO, sizeof(float4) * cap)); CUDA_TRY(alloc((void **)&w.shD, sizeof(float4) * cap));
    CUDA_TRY(alloc((void **)&w.shC, sizeof(float4) * cap));
    w.rec = nullptr;
    if (needRec)   // surface records of k_shade_surface -> k_shade_scatter
        CUDA_TRY(alloc((void **)&w.rec, sizeof(float4) * (size_t)kSurfRows * cap));
    w.raySortBin = nullptr;
    if (c->raySort) {
        CUDA_TRY(alloc((void **)&w.raySortBin, sizeof(uint32_t) * 3 * kRaySortKeys));
        CUDA_TRY(cudaMemsetAsync(w.raySortBin, 0, sizeof(uint32_t) * 3 * kRaySortKeys, c->stream));
    }
    w.ewaJobs = nullptr; w.ewaC = nullptr; w.ewaOrder = nullptr;
    if (needRec && c->deferEwa) {
        CUDA_TRY(alloc((void **)&w.ewaJobs, sizeof(float4) * 2 * (size_t)cap));
        CUDA_TRY(alloc((void **)&w.ewaC, sizeof(float) * (size_t)cap));
        CUDA_TRY(alloc((void **)&w.ewaOrder, sizeof(uint32_t) * (size_t)cap));
    }
    CUDA_TRY(alloc((void **)&w.queue[0], 4 * (size_t)cap)); CUDA_TRY(alloc((void **)&w.queue[1], 4 * (size_t)cap));
    CUDA_TRY(alloc((void **)&w.doneQ, 4 * (size_t)cap));
    CUDA_TRY(alloc((void **)&w.keys, 4 * (size_t)cap)); CUDA_TRY(alloc((void **)&w.sorted, 4 * (size_t)cap));
    CUDA_TRY(alloc((void **)&w.keyCount, 4 * 3 * (size_t)numKeys));
    CUDA_TRY(cudaMemsetAsync(w.keyCount, 0, 4 * 3 * (size_t)numKeys, c->stream));
    w.numKeys = numKeys;
    for (int k = 0; k <= SHADE_CLASSES; ++k) w.classKey[k] = c->classKey[k];
    CUDA_TRY(alloc((void **)&w.ctr, sizeof(Counters)));
    w.capacity = cap;
    c->wfCapacity = cap;
    return MTS_OK;
}

// ---- upload helpers --------------------------------------------------------------------------
// float texels -> half storage (TMIPMap keeps half precision); RGB is padded to half4
__global__ void k_texels_to_half(const float *__restrict__ src, __half *__restrict__ dst, size_t npix, int channels) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    if (channels == 1) dst[i] = __float2half_rn(src[i]);
    else {
        dst[4 * i] = __float2half_rn(src[3 * i]); dst[4 * i + 1] = __float2half_rn(src[3 * i + 1]);
        dst[4 * i + 2] = __float2half_rn(src[3 * i + 2]); dst[4 * i + 3] = __float2half_rn(0.0f);
    }
}
__global__ void k_half_to_float(const __half *__restrict__ src, float *__restrict__ dst, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = __half2float(src[i]);
}
// normals + texcoords of every vertex interleaved into one 32-byte record (DScene::vtxAttr)
__global__ void k_pack_vertex_attrs(const float *__restrict__ nrm, const float *__restrict__ uv, size_t nverts, float4 *__restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nverts) return;
    out[2 * i] = make_float4(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2], uv[2 * i]);
    out[2 * i + 1] = make_float4(uv[2 * i + 1], 0.0f, 0.0f, 0.0f);
}
// mesh-local vertex indices -> global ones, with the range check of the mesh loaders
__global__ void k_offset_indices(uint32_t *idx, size_t n, uint32_t nverts, uint32_t vOff, int *err) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t v = idx[i];
    if (v >= nverts) { *err = 1; v = 0; }
    idx[i] = v + vOff;
}

// TriMesh::computeUVTangents (src/librender/trimesh.cpp:683-743), one triangle per thread; the products and
// differences are rounded separately (no FMA contraction) so that the result equals the host computation bit for bit.
__global__ void k_uv_tangents(const uint32_t *__restrict__ idx, size_t ntris, const float *__restrict__ pos,
                              const float *__restrict__ uv, float *__restrict__ tang) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ntris) return;
    const uint32_t i0 = idx[3 * i], i1 = idx[3 * i + 1], i2 = idx[3 * i + 2];
    const float3 v0 = f3(pos[3 * (size_t)i0], pos[3 * (size_t)i0 + 1], pos[3 * (size_t)i0 + 2]);
    const float3 v1 = f3(pos[3 * (size_t)i1], pos[3 * (size_t)i1 + 1], pos[3 * (size_t)i1 + 2]);
    const float3 v2 = f3(pos[3 * (size_t)i2], pos[3 * (size_t)i2 + 1], pos[3 * (size_t)i2 + 2]);
    const float2 t0 = reinterpret_cast<const float2 *>(uv)[i0], t1 = reinterpret_cast<const float2 *>(uv)[i1], t2 = reinterpret_cast<const float2 *>(uv)[i2];
    const float3 dP1 = v1 - v0, dP2 = v2 - v0;
    const float du1 = t1.x - t0.x, dv1 = t1.y - t0.y, du2 = t2.x - t0.x, dv2 = t2.y - t0.y;
    auto mul = [](float a, float b) { return __fmul_rn(a, b); };
    auto sub = [](float a, float b) { return __fsub_rn(a, b); };
    const float3 n = f3(sub(mul(dP1.y, dP2.z), mul(dP1.z, dP2.y)), sub(mul(dP1.z, dP2.x), mul(dP1.x, dP2.z)), sub(mul(dP1.x, dP2.y), mul(dP1.y, dP2.x)));
    const float len = sqrtf(__fadd_rn(__fadd_rn(mul(n.x, n.x), mul(n.y, n.y)), mul(n.z, n.z)));
    float3 dpdu = f3(0.0f), dpdv = f3(0.0f);
    if (len != 0) {
        const float det = sub(mul(du1, dv2), mul(dv1, du2));
        if (det == 0) {
            coordinateSystem(n / len, dpdu, dpdv);
        } else {
            const float inv = 1.0f / det;
            dpdu = f3(mul(sub(mul(dv2, dP1.x), mul(dv1, dP2.x)), inv), mul(sub(mul(dv2, dP1.y), mul(dv1, dP2.y)), inv), mul(sub(mul(dv2, dP1.z), mul(dv1, dP2.z)), inv));
            dpdv = f3(mul(__fadd_rn(mul(-du2, dP1.x), mul(du1, dP2.x)), inv), mul(__fadd_rn(mul(-du2, dP1.y), mul(du1, dP2.y)), inv), mul(__fadd_rn(mul(-du2, dP1.z), mul(du1, dP2.z)), inv));
        }
    }
    float *o = tang + 6 * i;
    o[0] = dpdu.x; o[1] = dpdu.y; o[2] = dpdu.z; o[3] = dpdv.x; o[4] = dpdv.y; o[5] = dpdv.z;
}

// EnvironmentMap::configure (src/emitters/envmap.cpp:260-329) on the device.  The reference accumulates every CDF
// sequentially in fp32; one thread per image row reproduces exactly that order (products and sums rounded
// separately, no FMA contraction), so the tables equal the host computation bit for bit.  The same thread rounds
// the row's texels to half (TMIPMap storage) and fills the row's guide table.
__global__ void k_env_rows(const float *__restrict__ raw, int w, int h, const float *__restrict__ rowWeights, __half *__restrict__ tex,
                           float *__restrict__ cdfCols, float *__restrict__ rowTerm, uint16_t *__restrict__ guideCols, int G) {
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= h) return;
    float *cdf = cdfCols + (size_t)y * (w + 1);
    float colSum = 0.0f;
    cdf[0] = 0.0f;
    for (int x = 0; x < w; ++x) {
        const size_t i = (size_t)y * w + x;
        const __half r = __float2half_rn(raw[3 * i]), g = __float2half_rn(raw[3 * i + 1]), b = __float2half_rn(raw[3 * i + 2]);
        tex[4 * i] = r; tex[4 * i + 1] = g; tex[4 * i + 2] = b; tex[4 * i + 3] = __float2half_rn(0.0f);
        const float lum = __fadd_rn(__fadd_rn(__fmul_rn(__half2float(r), 0.212671f), __fmul_rn(__half2float(g), 0.715160f)), __fmul_rn(__half2float(b), 0.072169f));
        colSum = __fadd_rn(colSum, lum);
        cdf[x + 1] = colSum;
    }
    const float normalization = __fdiv_rn(1.0f, colSum);
    for (int x = 1; x < w; ++x) cdf[x] = __fmul_rn(cdf[x], normalization);
    cdf[w] = 1.0f;
    rowTerm[y] = __fmul_rn(colSum, rowWeights[y]);
    // guide[k] = std::lower_bound(cdf, cdf + w + 1, k / G) - cdf
    uint16_t *guide = guideCols + (size_t)y * (G + 1);
    int pos = 0;
    for (int k = 0; k <= G; ++k) {
        const float v = __fdiv_rn((float)k, (float)G);
        while (pos <= w && cdf[pos] < v) ++pos;
        guide[k] = (uint16_t)min(pos, w);
    }
}
__global__ void k_env_finish(const float *__restrict__ rowTerm, int h, float *__restrict__ cdfRows, uint16_t *__restrict__ guideRows, int G, float *rowSumOut) {
    if (blockIdx.x || threadIdx.x) return;
    float rowSum = 0.0f;
    cdfRows[0] = 0.0f;
    for (int y = 0; y < h; ++y) { rowSum = __fadd_rn(rowSum, rowTerm[y]); cdfRows[y + 1] = rowSum; }
    const float normalization = __fdiv_rn(1.0f, rowSum);
    for (int y = 1; y < h; ++y) cdfRows[y] = __fmul_rn(cdfRows[y], normalization);
    cdfRows[h] = 1.0f;
    int pos = 0;
    for (int k = 0; k <= G; ++k) {
        const float v = __fdiv_rn((float)k, (float)G);
        while (pos <= h && cdfRows[pos] < v) ++pos;
        guideRows[k] = (uint16_t)min(pos, h);
    }
    *rowSumOut = rowSum;
}

// ---- MIP pyramid (TMIPMap ctor, include/mitsuba/render/mipmap.h:162-303) ------------------------------------------
__global__ void k_clamp_negative(float *p, size_t n) {   // Spectrum::clampNegative of the base image (:228-236)
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && p[i] < 0) p[i] = 0.0f;
}
// One axis of Bitmap::resample (src/libcore/bitmap.cpp:2259-2328) = Resampler::resampleAndClamp
// (include/mitsuba/core/rfilter.h:229-340, boundary handling :432-458).  The taps are accumulated in the host code's
// order with separately rounded products and sums, so every level equals the host computation bit for bit.
__global__ void k_resample_axis(const float *__restrict__ src, float *__restrict__ dst, int sw, int sh, int tw, int th, int ch,
                                int axis, int bc, const int *__restrict__ start, const float *__restrict__ weights, int taps, float hi) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)tw * th) return;
    const int x = (int)(idx % tw), y = (int)(idx / tw);
    const int t = axis == 0 ? x : y, res = axis == 0 ? sw : sh;
    const int s0 = start[t];
    const float *wts = weights + (size_t)t * taps;
    for (int c = 0; c < ch; ++c) {
        float result = 0.0f;
        for (int j = 0; j < taps; ++j) {
            int pos = s0 + j;
            float v;
            bool constant = false;
            if (pos < 0 || pos >= res) {
                switch (bc) {
                    case MTS_WRAP_CLAMP: pos = min(max(pos, 0), res - 1); break;
                    case MTS_WRAP_REPEAT: pos = wrapRepeat(pos, res); break;
                    case MTS_WRAP_MIRROR: pos = wrapRepeat(pos, 2 * res); if (pos >= res) pos = 2 * res - pos - 1; break;
                    case MTS_WRAP_ZERO: constant = true; v = 0.0f; break;
                    default: constant = true; v = 1.0f; break;
                }
            }
            if (!constant) v = axis == 0 ? src[((size_t)y * sw + pos) * ch + c] : src[((size_t)pos * sw + x) * ch + c];
            result = __fadd_rn(result, __fmul_rn(v, wts[j]));
        }
        dst[idx * ch + c] = fminf(hi, fmaxf(0.0f, result));
    }
}

// LanczosSincFilter::eval, 2 lobes (src/rfilters/lanczos.cpp:38-50)
static float lanczos2(float x) {
    const float radius = 2.0f;
    x = std::fabs(x);
    if (x < kEps) return 1.0f;
    else if (x > radius) return 0.0f;
    float x1 = kPi * x, x2 = x1 / radius;
    return (std::sin(x1) * std::sin(x2)) / (x1 * x2);
}
// Resampler ctor in resampling mode (rfilter.h:123-176)
static void resamplerWeights(int sourceRes, int targetRes, int &taps, std::vector<int> &start, std::vector<float> &weights) {
    float filterRadius = 2.0f, scale = 1.0f, invScale = 1.0f;
    if (targetRes < sourceRes) { scale = (float)sourceRes / (float)targetRes; invScale = 1 / scale; filterRadius *= scale; }
    taps = (int)std::ceil(filterRadius * 2);
    start.resize(targetRes); weights.resize((size_t)taps * targetRes);
    for (int i = 0; i < targetRes; i++) {
        float center = (i + 0.5f) / targetRes * sourceRes;
        start[i] = (int)std::floor(center - filterRadius + 0.5f);
        float sum = 0;
        for (int j = 0; j < taps; j++) {
            float pos = start[i] + j + 0.5f - center;
            float weight = lanczos2(pos * invScale);
            weights[(size_t)i * taps + j] = weight;
            sum += weight;
        }
        float normalization = 1.0f / sum;
        for (int j = 0; j < taps; j++) weights[(size_t)i * taps + j] *= normalization;
    }
}

// Builds the pyramid of `raw` (w*h*ch floats on the device; negative values are clamped in place) into one half buffer
// and describes it in `out`.  trilinear / ewa: every further level is resampled from the previous FLOAT level (x pass,
// then y pass, clamped to [0, maxValue]) and rounded to half.
static int buildPyramid(MtsContext *c, float *raw, int w, int h, int ch, int bcu, int bcv, int filter, float maxAniso, float maxValue, DTexture &out,
                        size_t *outBytes = nullptr) {
    cudaStream_t st = c->stream;
    const int stride = ch == 1 ? 1 : 4;
    std::vector<int> lw(1, w), lh(1, h);
    if (filter == MTS_FILTER_TRILINEAR || filter == MTS_FILTER_EWA)
        while (lw.back() > 1 || lh.back() > 1) { int nw = std::max(1, (lw.back() + 1) / 2), nh = std::max(1, (lh.back() + 1) / 2); lw.push_back(nw); lh.push_back(nh); }
    if ((int)lw.size() > MTS_MAX_MIP_LEVELS || w > 65535 || h > 65535) return fail(MTS_EINVAL, "texture too large for the MIP pyramid");
    size_t total = 0;
    for (size_t l = 0; l < lw.size(); ++l) { out.levelOffset[l] = (uint32_t)total; out.levelW[l] = (uint16_t)lw[l]; out.levelH[l] = (uint16_t)lh[l]; total += (size_t)lw[l] * lh[l]; }
    if (total > 0xffffffffull) return fail(MTS_EINVAL, "texture too large for the MIP pyramid");
    for (size_t l = lw.size(); l < MTS_MAX_MIP_LEVELS; ++l) { out.levelOffset[l] = 0; out.levelW[l] = out.levelH[l] = 0; }
    __half *pyr = nullptr;
    CUDA_TRY(devAlloc(&pyr, sizeof(__half) * total * stride, st));
    auto grid = [](size_t n) { return (unsigned)((n + 255) / 256); };
    const size_t n0 = (size_t)w * h;
    k_clamp_negative<<<grid(n0 * ch), 256, 0, st>>>(raw, n0 * ch);
    k_texels_to_half<<<grid(n0), 256, 0, st>>>(raw, pyr, n0, ch);
    float *cur = raw; bool ownCur = false;
    cudaError_t e = cudaSuccess;
    for (size_t l = 1; l < lw.size() && e == cudaSuccess; ++l) {
        const int sw = lw[l - 1], sh = lh[l - 1], tw = lw[l], th = lh[l];
        float *tmp = nullptr, *next = nullptr; int *dStart = nullptr; float *dW = nullptr;
        const float *src = cur;
        auto pass = [&](int axis, int sW, int sH, int tW, int tH, int bc, float **dst) {
            int taps; std::vector<int> start; std::vector<float> weights;
            resamplerWeights(axis == 0 ? sW : sH, axis == 0 ? tW : tH, taps, start, weights);
            if (e == cudaSuccess) e = devAlloc(dst, sizeof(float) * (size_t)tW * tH * ch, st);
            if (e == cudaSuccess) e = devAlloc(&dStart, sizeof(int) * start.size(), st);
            if (e == cudaSuccess) e = devAlloc(&dW, sizeof(float) * weights.size(), st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(dStart, start.data(), sizeof(int) * start.size(), cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(dW, weights.data(), sizeof(float) * weights.size(), cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) k_resample_axis<<<grid((size_t)tW * tH), 256, 0, st>>>(src, *dst, sW, sH, tW, tH, ch, axis, bc, dStart, dW, taps, maxValue);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);   // the host weight vectors go out of scope
            devFree(dStart, st); devFree(dW, st); dStart = nullptr; dW = nullptr;
        };
        if (sw != tw) { pass(0, sw, sh, tw, sh, bcu, &tmp); src = tmp; }
        if (sh != th) pass(1, tw, sh, tw, th, bcv, &next);
        else { next = tmp; tmp = nullptr; }
        if (tmp) devFree(tmp, st);
        if (ownCur) devFree(cur, st);
        cur = next; ownCur = true;
        if (e == cudaSuccess && cur) k_texels_to_half<<<grid((size_t)tw * th), 256, 0, st>>>(cur, pyr + (size_t)out.levelOffset[l] * stride, (size_t)tw * th, ch);
    }
    if (ownCur && cur) devFree(cur, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { devFree(pyr, st); return fail(MTS_ECUDA, std::string("MIP pyramid: ") + cudaGetErrorString(e)); }
    out.texels = pyr; out.width = w; out.height = h; out.channels = ch; out.wrapU = bcu; out.wrapV = bcv; out.filter = filter;
    out.uvScale = make_float2(1, 1); out.uvOffset = make_float2(0, 0);
    out.levels = (int)lw.size();
    out.maxAnisotropy = filter == MTS_FILTER_EWA ? maxAniso : 1.0f;   // bitmap.cpp:230-233
    out.weightLut = c->ewaLut;
    if (outBytes) *outBytes = sizeof(__half) * total * stride;
    return MTS_OK;
}

// ---- kernels behind the per-function entry points -------------------------------------------
__global__ void k_hook_sample_rays(const DScene S, int64_t n, const float *sp, float *out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float3 o, d, rx, ry; float mint, maxt;
    cameraRay(S.cam, make_float2(sp[2 * i], sp[2 * i + 1]), o, d, mint, maxt, &rx, &ry);
    float *q = out + 16 * i;
    q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = mint; q[4] = d.x; q[5] = d.y; q[6] = d.z; q[7] = maxt;
    q[8] = rx.x; q[9] = rx.y; q[10] = rx.z; q[11] = 0; q[12] = ry.x; q[13] = ry.y; q[14] = ry.z; q[15] = 0;
}

// mtscuda_trace drives the PRODUCTION queue kernels (k_intersect / k_shadow: persistent warps, lane refill, shared
// stack), so the bit-exact intersection tests cover exactly what the renderer runs.  k_trace_load fills the wavefront
// arrays the way raygen / shade do, k_trace_export turns the hit records back into (t, u, v) + (shape, primitive).
__global__ void k_trace_load(const Wavefront W, uint32_t n, const float *rays, int shadow) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        Counters *c = W.ctr;
        c->qCount[0] = shadow ? 0u : n; c->qCount[1] = 0; c->shadowCount = shadow ? n : 0u; c->doneCount = 0; c->cursor[0] = c->cursor[1] = 0;
    }
    if (i >= n) return;
    const float *r = rays + 8 * (size_t)i;
    const float4 o = make_float4(r[0], r[1], r[2], r[3]), d = make_float4(r[4], r[5], r[6], r[7]);
    if (shadow) {
        W.shO[i] = o; W.shD[i] = d; W.shC[i] = make_float4(1.0f, 0.0f, 0.0f, __int_as_float((int)i));
        W.L[i] = make_float4(0, 0, 0, 0);
    } else {
        W.rayO[i] = o; W.rayD[i] = d; W.queue[0][i] = i;
    }
}
__global__ void k_trace_export(const DScene S, const Wavefront W, uint32_t n, int shadow, float *hits, int32_t *ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float *q = hits + 4 * (size_t)i;
    q[1] = q[2] = q[3] = 0;
    ids[2 * (size_t)i] = ids[2 * (size_t)i + 1] = -1;
    if (shadow) { q[0] = W.L[i].x != 0.0f ? kInf : 0.0f; return; }   // k_shadow adds the contribution of unoccluded rays
    const float4 h = W.hit[i];
    const int prim = __float_as_int(h.w);
    if (prim < 0) { q[0] = kInf; return; }
    const float4 *rec = reinterpret_cast<const float4 *>(S.prims + prim);
    const float4 p0 = rec[0], p1 = rec[1], p2 = rec[2];
    const bool tri = __float_as_int(p2.w) == 0;
    q[0] = h.x; q[1] = tri ? h.y : 0.0f; q[2] = tri ? h.z : 0.0f;
    ids[2 * (size_t)i] = __float_as_int(p0.w); ids[2 * (size_t)i + 1] = tri ? __float_as_int(p1.w) : 0;
}

__global__ void k_hook_hit_records(const DScene S, int64_t n, const float *rays, float *out, int32_t *ids) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *r = rays + 8 * i;
    float3 o = f3(r[0], r[1], r[2]), d = f3(r[4], r[5], r[6]);
    RayHit h; float *q = out + 24 * i;
    if (!sceneIntersect<false, false>(S, o, d, r[3], r[7], h)) {
        for (int k = 0; k < 24; ++k) q[k] = 0;
        q[3] = kInf;
        if (ids) { ids[2 * i] = ids[2 * i + 1] = -1; }
        return;
    }
    Hit its; fillHit(S, o, d, h, its);
    q[0] = its.p.x; q[1] = its.p.y; q[2] = its.p.z; q[3] = its.t;
    q[4] = its.geoN.x; q[5] = its.geoN.y; q[6] = its.geoN.z; q[7] = its.uv.x;
    q[8] = its.sh.n.x; q[9] = its.sh.n.y; q[10] = its.sh.n.z; q[11] = its.uv.y;
    q[12] = its.dpdu.x; q[13] = its.dpdu.y; q[14] = its.dpdu.z; q[15] = 0;
    q[16] = its.dpdv.x; q[17] = its.dpdv.y; q[18] = its.dpdv.z; q[19] = 0;
    q[20] = its.wi.x; q[21] = its.wi.y; q[22] = its.wi.z; q[23] = 0;
    if (ids) { ids[2 * i] = its.shape; ids[2 * i + 1] = its.prim; }
}

__global__ void k_hook_texture_eval(const DScene S, int tex, int64_t n, const float *uv, const float *duv, float *out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    TexCtx tc = texCtx(make_float2(uv[2 * i], uv[2 * i + 1]));
    if (duv) { tc.duv = make_float4(duv[4 * i], duv[4 * i + 1], duv[4 * i + 2], duv[4 * i + 3]); tc.filtered = true; }
    MtsTexRef r; r.scale[0] = r.scale[1] = r.scale[2] = 1.0f; r.tex = tex;
    const float3 v = evalTex(S, r, tc);
    out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
}
__global__ void k_hook_hit_partials(const DScene S, int64_t n, const float *sp, float diffScale, float *out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float3 o, d, rxD, ryD; float mint, maxt;
    cameraRay(S.cam, make_float2(sp[2 * i], sp[2 * i + 1]), o, d, mint, maxt);
    float3 o2, d2;
    cameraRayDiff(S.cam, make_float2(sp[2 * i], sp[2 * i + 1]), diffScale, o2, d2, rxD, ryD);
    float *q = out + 8 * i;
    RayHit h;
    if (!sceneIntersect<false, false>(S, o, d, mint, maxt, h)) { for (int k = 0; k < 8; ++k) q[k] = 0; return; }
    Hit its; fillHit(S, o, d, h, its);
    const float4 p = computePartials(its.p, its.geoN, its.dpdu, its.dpdv, o, rxD, ryD);
    q[0] = p.x; q[1] = p.y; q[2] = p.z; q[3] = p.w; q[4] = its.t; q[5] = 1.0f; q[6] = its.uv.x; q[7] = its.uv.y;
}
__global__ void k_hook_eval_env_camera(const DScene S, int64_t n, const float *sp, float diffScale, float *out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float3 o, d, rxD, ryD;
    cameraRayDiff(S.cam, make_float2(sp[2 * i], sp[2 * i + 1]), diffScale, o, d, rxD, ryD);
    const float3 v = envEvalDiff(S, S.env, d, rxD, ryD);
    out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
}
__global__ void k_hook_sample_emitter(const DScene S, int64_t n, const float *ref, const float *refn, const float *u, float *out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    DRec r; r.ref = f3(ref[3 * i], ref[3 * i + 1], ref[3 * i + 2]); r.refN = f3(refn[3 * i], refn[3 * i + 1], refn[3 * i + 2]);
    r.pdf = 0; r.emitter = -1; r.measure = 0; r.dist = 0; r.d = r.n = r.p = f3(0.0f);
    float3 v = sampleEmitterDirect(S, r, make_float2(u[2 * i], u[2 * i + 1]));
    float *q = out + 16 * i;
    if (isZero(v) && r.pdf == 0) { for (int k = 0; k < 16; ++k) q[k] = 0; q[15] = -1; return; }
    q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = r.pdf; q[4] = r.d.x; q[5] = r.d.y; q[6] = r.d.z; q[7] = r.dist;
    q[8] = r.p.x; q[9] = r.p.y; q[10] = r.p.z; q[11] = (float)r.measure; q[12] = r.n.x; q[13] = r.n.y; q[14] = r.n.z;
    q[15] = (float)r.emitter;
}
__global__ void k_hook_eval_env(const DScene S, int64_t n, const float *d, float *out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float3 dir = f3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    float3 v = envEval(S.env, dir);
    out[4 * i] = v.x; out[4 * i + 1] = v.y; out[4 * i + 2] = v.z;
    out[4 * i + 3] = envPdfDirection(S.env, xfVec(S.env.toLocal, dir)) * S.env.prob;
}
__global__ void k_hook_splat(const DScene S, int64_t n, const float *pos, const float *val) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    splatSample(S, make_float2(pos[2 * i], pos[2 * i + 1]), f3(val[3 * i], val[3 * i + 1], val[3 * i + 2]));
}

// device staging of host arrays for the hooks
struct Staged {
    cudaStream_t st;
    std::vector<void *> ptrs;
    explicit Staged(cudaStream_t s) : st(s) {}
    ~Staged() { for (void *p : ptrs) devFree(p, st); }
    template <class T> T *in(const T *h, size_t n, cudaStream_t) {
        T *d = nullptr;
        if (devAlloc(&d, sizeof(T) * std::max<size_t>(n, 1), st) != cudaSuccess) return nullptr;
        ptrs.push_back(d);
        if (n) cudaMemcpyAsync(d, h, sizeof(T) * n, cudaMemcpyHostToDevice, st);
        return d;
    }
    template <class T> T *out(size_t n) {
        T *d = nullptr;
        if (devAlloc(&d, sizeof(T) * std::max<size_t>(n, 1), st) != cudaSuccess) return nullptr;
        ptrs.push_back(d);
        return d;
    }
};
inline int gridFor(int64_t n, int bs = 128) { return (int)std::max<int64_t>(1, (n + bs - 1) / bs); }

int syncCheck(MtsContext *c, const char *what) {
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MTS_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
    return MTS_OK;
}

} // namespace

// =============================================================================================
extern "C" {

MTS_API int32_t mtscuda_abi_version(void) { return MTS_ABI_VERSION; }
MTS_API const char *mtscuda_last_error(void) { return g_err.c_str(); }

MTS_API int mtscuda_create(int device, MtsContext **out) {
    if (!out) return fail(MTS_EINVAL, "null out pointer");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(MTS_ECUDA, std::string("no CUDA device available (libmtscuda has no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(MTS_EINVAL, "bad device index");
    CUDA_TRY(cudaSetDevice(device));
    MtsContext *c = new MtsContext();
    c->device = device;
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    c->numSMs = prop.multiProcessorCount;
    // Opt-in (MTS_L2_WINDOW=1): measured on cfg3 it LOSES 4.5 % -- the traversal gains nothing (its hot nodes already
    // stay in the 126 MB L2) while the carve-out halves the L2 left to the shade stage.
    if (getenv("MTS_L2_WINDOW") && prop.persistingL2CacheMaxSize > 0) {
        c->l2Bytes = std::min<size_t>((size_t)prop.persistingL2CacheMaxSize, (size_t)prop.l2CacheSize / 2);
        c->maxWindow = (size_t)prop.accessPolicyMaxWindowSize;
        if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, c->l2Bytes) != cudaSuccess) { c->l2Bytes = 0; cudaGetLastError(); }
    }
    devPoolInit(device);
    CUDA_TRY(cudaStreamCreateWithFlags(&c->ownStream, cudaStreamNonBlocking));
    c->stream = c->ownStream;
    CUDA_TRY(cudaStreamCreateWithFlags(&c->upStream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&c->upDone, cudaEventDisableTiming));
    CUDA_TRY(cudaHostAlloc((void **)&c->hostPoll, sizeof(uint32_t) * kPollRing, cudaHostAllocDefault));
    {   // EWA Gaussian weight table (mipmap.h:296-302; math::fastexp evaluates in double, include/mitsuba/core/math.h:185-187)
        float lut[MTS_MIPMAP_LUT_SIZE];
        for (int i = 0; i < MTS_MIPMAP_LUT_SIZE; ++i) {
            const float r2 = (float)i / (float)(MTS_MIPMAP_LUT_SIZE - 1);
            lut[i] = (float)std::exp((double)(-2.0f * r2)) - (float)std::exp((double)-2.0f);
        }
        CUDA_TRY(devAlloc(&c->ewaLut, sizeof(lut), c->stream));
        CUDA_TRY(cudaMemcpyAsync(c->ewaLut, lut, sizeof(lut), cudaMemcpyHostToDevice, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    *out = c;
    return MTS_OK;
}

MTS_API void mtscuda_destroy(MtsContext *c) {
    if (!c) return;
    for (MtsContext *p : c->peers) mtscuda_destroy(p);
    c->peers.clear();
    cudaSetDevice(c->device);
    c->freeScene(); c->freeWavefront(); c->freeGeometry();
    c->freeFilm(); devFree(c->ewaLut, c->stream);
    cudaStreamSynchronize(c->stream);
    for (cudaEvent_t e : c->eventPool) cudaEventDestroy(e);
    if (c->hostPoll) cudaFreeHost(c->hostPoll);
    if (c->upDone) cudaEventDestroy(c->upDone);
    if (c->upStream) cudaStreamDestroy(c->upStream);
    if (c->ownStream) cudaStreamDestroy(c->ownStream);
    if (getenv("MTS_POOL_TRIM")) devPoolTrim(c->device);
    delete c;
}

MTS_API int mtscuda_set_camera(MtsContext *c, const MtsCamera *cam) {
    c->cam = *cam; c->hasCam = true; c->ds.cam = *cam;
    return MTS_OK;
}

MTS_API int mtscuda_set_film(MtsContext *c, const MtsFilm *f) {
    if (f->width <= 0 || f->height <= 0 || !(f->filter_radius > 0)) return fail(MTS_EINVAL, "bad film description");
    CUDA_TRY(cudaSetDevice(c->device));
    c->freeFilm();
    c->film = *f; c->hasFilm = true; c->filmBound = nullptr;
    size_t bytes = sizeof(float4) * (size_t)f->width * f->height;
    if (c->peerFilm) CUDA_TRY(cudaMalloc((void **)&c->filmInternal, bytes));
    else CUDA_TRY(devAlloc(&c->filmInternal, bytes, c->stream));
    CUDA_TRY(cudaMemsetAsync(c->filmInternal, 0, bytes, c->stream));
    if (f->has_alpha) {
        if (c->peerFilm) CUDA_TRY(cudaMalloc((void **)&c->filmAlpha, bytes / 4));
        else CUDA_TRY(devAlloc(&c->filmAlpha, bytes / 4, c->stream));
        CUDA_TRY(cudaMemsetAsync(c->filmAlpha, 0, bytes / 4, c->stream));
    }
    c->ds.filmAlpha = c->filmAlpha;
    c->ds.film = c->filmInternal; c->ds.filmW = f->width; c->ds.filmH = f->height;
    c->ds.filterRadius = f->filter_radius;
    std::memcpy(c->ds.filterTable, f->filter_table, sizeof(f->filter_table));
    return MTS_OK;
}

MTS_API int mtscuda_add_texture(MtsContext *c, const MtsTextureDesc *d, const float *texels) {
    if (d->channels != 1 && d->channels != 3) return fail(MTS_EINVAL, "texture channels must be 1 or 3");
    if (d->width <= 0 || d->height <= 0) return fail(MTS_EINVAL, "bad texture size");
    if (d->filter < MTS_FILTER_NEAREST || d->filter > MTS_FILTER_EWA) return fail(MTS_EINVAL, "unknown texture filter type");
    if (d->texel_format != MTS_TEXEL_F32 && d->texel_format != MTS_TEXEL_F16) return fail(MTS_EINVAL, "unknown texel format");
    CUDA_TRY(cudaSetDevice(c->device));
    HostTex t; t.d = *d;
    const size_t npix = (size_t)d->width * d->height;
    float *raw = nullptr;
    // everything below is issued on the upload stream (ordered after what the main stream has done so far)
    struct StreamSwap {
        MtsContext *c; cudaStream_t saved;
        StreamSwap(MtsContext *c_) : c(c_), saved(c_->stream) {
            if (!c->upPending) { cudaEventRecord(c->upDone, c->stream); cudaStreamWaitEvent(c->upStream, c->upDone, 0); }
            c->stream = c->upStream; c->upPending = true;
        }
        ~StreamSwap() { c->stream = saved; }
    } swap(c);
    CUDA_TRY(devAlloc(&raw, sizeof(float) * npix * d->channels, c->stream));
    cudaError_t e;
    if (d->texel_format == MTS_TEXEL_F16) {   // half texels: half the bytes over PCIe, widened on the device (exact)
        __half *rawHalf = nullptr;
        CUDA_TRY(devAlloc(&rawHalf, sizeof(__half) * npix * d->channels, c->stream));
        e = cudaMemcpyAsync(rawHalf, texels, sizeof(__half) * npix * d->channels, cudaMemcpyDefault, c->stream);
        const size_t nval = npix * d->channels;
        if (e == cudaSuccess) k_half_to_float<<<(unsigned)((nval + 255) / 256), 256, 0, c->stream>>>(rawHalf, raw, nval);
        devFree(rawHalf, c->stream);
    } else
        e = cudaMemcpyAsync(raw, texels, sizeof(float) * npix * d->channels, cudaMemcpyDefault, c->stream);
    int rc = MTS_OK;
    if (e == cudaSuccess) rc = buildPyramid(c, raw, d->width, d->height, d->channels, d->wrap_u, d->wrap_v, d->filter, d->max_anisotropy, 1.0f /* TMIPMap maxValue default */, t.dt, &t.bytes);
    devFree(raw, c->stream);
    if (e != cudaSuccess) return fail(MTS_ECUDA, std::string("add_texture: ") + cudaGetErrorString(e));
    if (rc != MTS_OK) return rc;
    t.dt.uvScale = make_float2(d->uv_scale[0], d->uv_scale[1]); t.dt.uvOffset = make_float2(d->uv_offset[0], d->uv_offset[1]);
    c->textures.push_back(t); c->built = false;
    return (int)c->textures.size() - 1;
}

MTS_API int mtscuda_set_float_pool(MtsContext *c, const float *data, int64_t n) { c->pool.assign(data, data + n); c->built = false; return MTS_OK; }

MTS_API int mtscuda_set_bsdfs(MtsContext *c, const MtsBsdfNode *nodes, int32_t n) {
    c->nodes.assign(nodes, nodes + n); c->built = false;
    bool hasMixture = false;
    for (int i = 0; i < n; ++i) {
        if (nodes[i].type < MTS_BSDF_DIFFUSE || nodes[i].type > MTS_BSDF_NORMALMAP) return fail(MTS_EINVAL, "unknown BSDF type");
        hasMixture |= nodes[i].type == MTS_BSDF_MIXTURE;
    }
    // mixturebsdf keeps its child list in the float pool, which may arrive later: such node sets are checked by build()
    for (int i = 0; i < n && !hasMixture; ++i)
        if (bsdfDepth(c->nodes, c->pool, i) > kMaxBsdfDepth) return fail(MTS_EINVAL, kDepthError);
    return MTS_OK;
}

MTS_API int mtscuda_add_mesh(MtsContext *c, const MtsMeshDesc *m) {
    if (!m->positions || !m->indices || m->ntris == 0) return fail(MTS_EINVAL, "Encountered an empty triangle mesh!");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    HostShape s; std::memset(&s.d, 0, sizeof(DShape));
    s.d.type = SHAPE_MESH; s.d.bsdf = m->bsdf; s.d.emitter = -1;
    const uint32_t vOff = (uint32_t)(c->positions.size / 3), tOff = (uint32_t)(c->indices.size / 3);
    s.d.vtxOffset = vOff; s.d.triOffset = tOff; s.d.ntris = m->ntris; s.nverts = m->nverts;
    const bool deviceTangents = m->uv_tangents == 1 && m->texcoords;
    s.d.flags = (m->normals ? 1u : 0u) | (m->texcoords ? 2u : 0u) | ((m->tangents || deviceTangents) ? 4u : 0u);
    if (!c->errFlag) { CUDA_TRY(devAlloc(&c->errFlag, sizeof(int), st)); CUDA_TRY(cudaMemsetAsync(c->errFlag, 0, sizeof(int), st)); }
    // the optional per-vertex / per-triangle arrays are kept index-aligned with positions / indices (zeros where absent)
    CUDA_TRY(c->positions.append(m->positions, 3 * (size_t)m->nverts, st));
    CUDA_TRY(c->normals.append(m->normals, 3 * (size_t)m->nverts, st));
    CUDA_TRY(c->texcoords.append(m->texcoords, 2 * (size_t)m->nverts, st));
    CUDA_TRY(c->indices.append(m->indices, 3 * (size_t)m->ntris, st));
    CUDA_TRY(c->tangents.append(deviceTangents ? nullptr : m->tangents, 6 * (size_t)m->ntris, st));
    const size_t ni = 3 * (size_t)m->ntris;
    k_offset_indices<<<(unsigned)((ni + 255) / 256), 256, 0, st>>>(c->indices.data + 3 * (size_t)tOff, ni, m->nverts, vOff, c->errFlag);
    if (deviceTangents)
        k_uv_tangents<<<(unsigned)((m->ntris + 255) / 256), 256, 0, st>>>(c->indices.data + 3 * (size_t)tOff, m->ntris, c->positions.data,
                                                                          c->texcoords.data, c->tangents.data + 6 * (size_t)tOff);
    int err = 0;
    CUDA_TRY(cudaMemcpyAsync(&err, c->errFlag, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));   // also guarantees the caller's buffers have been consumed
    if (err) { CUDA_TRY(cudaMemsetAsync(c->errFlag, 0, sizeof(int), st)); return fail(MTS_EINVAL, "mesh index out of range"); }
    c->shapes.push_back(s); c->built = false;
    return (int)c->shapes.size() - 1;
}

MTS_API int mtscuda_add_rectangle(MtsContext *c, const float tw[16], int32_t bsdf) {
    HostShape s; std::memset(&s.d, 0, sizeof(DShape));
    s.d.type = SHAPE_RECT; s.d.bsdf = bsdf; s.d.emitter = -1;
    M4 o2w = m4From(tw), w2o = m4Inverse(o2w);
    s.d.o2w = toAffine(o2w); s.d.w2o = toAffine(w2o);
    // Rectangle::configure (src/shapes/rectangle.cpp:102-113)
    const float ex[3] = {2, 0, 0}, ey[3] = {0, 2, 0};
    float du[3], dv[3]; hVec(s.d.o2w, ex, du); hVec(s.d.o2w, ey, dv);
    float nn[3] = {(float)w2o.m[2][0], (float)w2o.m[2][1], (float)w2o.m[2][2]};  // inverse-transpose * (0,0,1)
    float nl = hLen(nn);
    s.d.dpdu = make_float4(du[0], du[1], du[2], 0); s.d.dpdv = make_float4(dv[0], dv[1], dv[2], 0);
    s.d.normal = make_float4(nn[0] / nl, nn[1] / nl, nn[2] / nl, 0);
    s.d.invArea = 1.0f / (hLen(du) * hLen(dv));
    const float corners[4][3] = {{-1, -1, 0}, {1, -1, 0}, {1, 1, 0}, {-1, 1, 0}};
    for (int a = 0; a < 3; ++a) { s.lo[a] = INFINITY; s.hi[a] = -INFINITY; }
    for (auto &cr : corners) { float p[3]; hPoint(s.d.o2w, cr, p); for (int a = 0; a < 3; ++a) { s.lo[a] = std::min(s.lo[a], p[a]); s.hi[a] = std::max(s.hi[a], p[a]); } }
    c->shapes.push_back(s); c->built = false;
    return (int)c->shapes.size() - 1;
}

MTS_API int mtscuda_add_sphere(MtsContext *c, const float o2wf[16], float radius, int32_t flip, int32_t bsdf) {
    if (!(radius > 0)) return fail(MTS_EINVAL, "Cannot create spheres of radius <= 0");
    HostShape s; std::memset(&s.d, 0, sizeof(DShape));
    s.d.type = SHAPE_SPHERE; s.d.bsdf = bsdf; s.d.emitter = -1; s.d.flip = flip != 0; s.d.radius = radius;
    M4 o2w = m4From(o2wf), w2o = m4Inverse(o2w);
    s.d.o2w = toAffine(o2w); s.d.w2o = toAffine(w2o);
    const float zero[3] = {0, 0, 0}; float ctr[3]; hPoint(s.d.o2w, zero, ctr);
    s.d.center = make_float4(ctr[0], ctr[1], ctr[2], 0);
    s.d.invArea = 1.0f / (4 * kPi * radius * radius);
    for (int a = 0; a < 3; ++a) { s.lo[a] = ctr[a] - radius; s.hi[a] = ctr[a] + radius; }
    c->shapes.push_back(s); c->built = false;
    return (int)c->shapes.size() - 1;
}

MTS_API int mtscuda_add_emitter(MtsContext *c, const MtsEmitterDesc *d) {
    HostEmitter e; e.d = *d; e.d.env_texels = nullptr;
    if (d->type == MTS_EMITTER_AREA) {
        if (d->shape < 0 || d->shape >= (int)c->shapes.size()) return fail(MTS_EINVAL, "area emitter: bad shape id");
        c->shapes[d->shape].d.emitter = (int)c->emitters.size();
        c->shapes[d->shape].isEmitter = true;
    } else if (d->type == MTS_EMITTER_ENVMAP) {
        if (c->envIndex >= 0) return fail(MTS_EINVAL, "only one environment emitter is allowed");
        if (!d->env_texels || d->env_width <= 0 || d->env_height <= 0) return fail(MTS_EINVAL, "envmap without texels");
        size_t n = (size_t)d->env_width * d->env_height * 3;
        if (d->env_width > 65534 || d->env_height > 65534) return fail(MTS_EINVAL, "Environment maps images must be smaller than 65536 pixels in width and height");
        CUDA_TRY(cudaSetDevice(c->device));
        CUDA_TRY(devAlloc(&e.envRaw, sizeof(float) * n, c->stream));   // straight from the caller's buffer; rounded to half on the device
        CUDA_TRY(cudaMemcpyAsync(e.envRaw, d->env_texels, sizeof(float) * n, cudaMemcpyDefault, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        c->envIndex = (int)c->emitters.size();
    } else if (d->type != MTS_EMITTER_POINT) return fail(MTS_EINVAL, "unsupported emitter type");
    c->emitters.push_back(std::move(e)); c->built = false;
    return (int)c->emitters.size() - 1;
}

static int buildOne(MtsContext *c) {
    CUDA_TRY(cudaSetDevice(c->device));
    if (c->emitters.empty())
        return fail(MTS_EINVAL, "scene has no emitters (the reference would add a sunsky emitter, which is out of scope)");
    if (c->nodes.empty()) return fail(MTS_EINVAL, "no BSDFs set");
    for (auto &s : c->shapes) if (s.d.bsdf < 0 || s.d.bsdf >= (int)c->nodes.size()) return fail(MTS_EINVAL, "shape references a bad BSDF index");
    for (int i = 0; i < (int)c->nodes.size(); ++i) {
        const int d = bsdfDepth(c->nodes, c->pool, i);
        if (d >= 1000) return fail(MTS_EINVAL, "malformed BSDF tree (bad child index, cycle or mixturebsdf table)");
        if (d > kMaxBsdfDepth) return fail(MTS_EINVAL, kDepthError);
    }
    for (auto &n : c->nodes) {
        const MtsTexRef *tr[3] = {&n.t0, &n.t1, &n.t2};
        for (auto t : tr) if (t->tex >= (int)c->textures.size()) return fail(MTS_EINVAL, "BSDF references a bad texture index");
    }
    c->freeScene();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, c->stream);
    struct JoinUploads {   // every way out of build() leaves the main stream ordered after the texture uploads
        MtsContext *c;
        ~JoinUploads() { if (c->upPending) { cudaEventRecord(c->upDone, c->upStream); cudaStreamWaitEvent(c->stream, c->upDone, 0); c->upPending = false; } }
    } join{c};
    DScene &S = c->ds;
    int rc;
    // geometry already lives on the device (appended by add_mesh)
    S.positions = c->positions.data; S.normals = c->normals.data; S.texcoords = c->texcoords.data;
    S.indices = c->indices.data; S.tangents = c->tangents.data;
    {
        const size_t nverts = c->positions.size / 3;
        float4 *va = nullptr;
        if (nverts) {
            CUDA_TRY(devAlloc(&va, sizeof(float4) * 2 * nverts, c->stream));
            c->sceneAllocs.push_back(va);
            k_pack_vertex_attrs<<<(unsigned)((nverts + 255) / 256), 256, 0, c->stream>>>(c->normals.data, c->texcoords.data, nverts, va);
        }
        S.vtxAttr = va;
    }
    // area CDFs of emitter meshes: TriMesh::prepareSamplingTable (src/librender/trimesh.cpp:388-403),
    // sequential fp32 accumulation like DiscreteDistribution::append/normalize (host, from a read-back of that mesh)
    std::vector<float> cdfs;
    BvhBuildInput bin;
    uint32_t primOffset = 0;
    for (size_t si = 0; si < c->shapes.size(); ++si) {
        HostShape &s = c->shapes[si];
        if (s.d.type == SHAPE_MESH) {
            bin.meshes.push_back({s.d.triOffset, s.d.ntris, primOffset, (int)si});
            primOffset += s.d.ntris;
            if (s.isEmitter) {
                std::vector<float> pos(3 * (size_t)s.nverts); std::vector<uint32_t> idx(3 * (size_t)s.d.ntris);
                CUDA_TRY(cudaMemcpyAsync(pos.data(), c->positions.data + 3 * (size_t)s.d.vtxOffset, sizeof(float) * pos.size(), cudaMemcpyDeviceToHost, c->stream));
                CUDA_TRY(cudaMemcpyAsync(idx.data(), c->indices.data + 3 * (size_t)s.d.triOffset, sizeof(uint32_t) * idx.size(), cudaMemcpyDeviceToHost, c->stream));
                CUDA_TRY(cudaStreamSynchronize(c->stream));
                s.d.cdfOffset = (uint32_t)cdfs.size();
                cdfs.push_back(0.0f);
                float acc = 0.0f;
                for (uint32_t t = 0; t < s.d.ntris; ++t) {
                    const uint32_t *ix = &idx[3 * (size_t)t];
                    const float *p0 = &pos[3 * (size_t)(ix[0] - s.d.vtxOffset)], *p1 = &pos[3 * (size_t)(ix[1] - s.d.vtxOffset)], *p2 = &pos[3 * (size_t)(ix[2] - s.d.vtxOffset)];
                    float a[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]}, b[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]};
                    float cr[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
                    acc += 0.5f * hLen(cr);
                    cdfs.push_back(acc);
                }
                float norm = 1.0f / acc;
                for (uint32_t t = 1; t <= s.d.ntris; ++t) cdfs[s.d.cdfOffset + t] *= norm;
                cdfs[s.d.cdfOffset + s.d.ntris] = 1.0f;
                s.d.invArea = 1.0f / acc;
            }
        } else {
            BvhAnalytic a; a.primOffset = primOffset++; a.shape = (int)si; a.type = s.d.type;
            for (int k = 0; k < 3; ++k) { a.lo[k] = s.lo[k]; a.hi[k] = s.hi[k]; }
            bin.analytic.push_back(a);
        }
    }
    if ((rc = devUpload(c, cdfs.data(), cdfs.size(), &S.areaCdfs))) return rc;
    // material sort keys: the distinct root BSDFs ordered by shade class, so that every k_shade<CLASS> instance owns one
    // contiguous range of the sorted queue (key 0 = escaped rays)
    {
        std::vector<int> roots;
        for (auto &s : c->shapes) roots.push_back(s.d.bsdf);
        std::sort(roots.begin(), roots.end());
        roots.erase(std::unique(roots.begin(), roots.end()), roots.end());
        std::stable_sort(roots.begin(), roots.end(), [&](int a, int b) { return shadeClassOf(c->nodes.data(), a) < shadeClassOf(c->nodes.data(), b); });
        std::vector<uint32_t> keyOf(c->nodes.size(), 0u);
        for (int k = 0; k <= SHADE_CLASSES; ++k) c->classKey[k] = (uint32_t)roots.size() + 1;
        for (size_t r = 0; r < roots.size(); ++r) {
            keyOf[roots[r]] = (uint32_t)r + 1;
            const int cls = shadeClassOf(c->nodes.data(), roots[r]);
            for (int k = 0; k <= cls; ++k) c->classKey[k] = std::min(c->classKey[k], (uint32_t)r + 1);
        }
        // classes without any BSDF start where the next populated class starts
        for (int k = SHADE_CLASSES - 1; k >= 0; --k) c->classKey[k] = std::min(c->classKey[k], c->classKey[k + 1]);
        c->numKeys = (uint32_t)roots.size() + 1;
        for (auto &s : c->shapes) s.d.sortKey = keyOf[s.d.bsdf];
        // Staged or fused shade of the leaf class?  Measured (profiles/r2e, r2m): the staged form wins by 2x where hits
        // look up bitmap textures (the texture / MIP / EWA code is what outgrows the instruction caches), the fused,
        // fully inlined kernel by 10-25 % where the materials are constants (no handover of surface records).
        c->leafTextured = false;
        for (int root : roots) {
            if (shadeClassOf(c->nodes.data(), root) != SHADE_LEAF) continue;
            std::vector<int> todo(1, root);
            while (!todo.empty()) {
                const MtsBsdfNode &n = c->nodes[todo.back()]; todo.pop_back();
                if (n.t0.tex >= 0 || n.t1.tex >= 0 || n.t2.tex >= 0 || (n.type == MTS_BSDF_ROUGHDIELECTRIC && n.lut[0] >= 0)) c->leafTextured = true;
                if (isWrapperBsdf(n.type) && n.type != MTS_BSDF_MIXTURE) for (int ch : n.child) if (ch >= 0) todo.push_back(ch);
            }
        }
    }
    std::vector<DShape> dshapes; for (auto &s : c->shapes) dshapes.push_back(s.d);
    if ((rc = devUpload(c, dshapes.data(), dshapes.size(), &S.shapes))) return rc;
    S.numShapes = (int)dshapes.size();
    // acceleration structure
    bin.numPrims = (int)primOffset; bin.positions = S.positions; bin.indices = S.indices;
    BvhBuildOutput bout; std::string berr;
    if (!buildBvh(bin, bout, c->stream, berr)) return fail(MTS_ECUDA, "BVH build failed: " + berr);
    if (bout.nodes) c->sceneAllocs.push_back(bout.nodes);
    if (bout.prims) c->sceneAllocs.push_back(bout.prims);
    c->bvhDepth = bout.depth;
    if (bout.depth > kStackCapacity)   // the traversal stack holds one node group per level; deeper trees would drop siblings
        return fail(MTS_EINVAL, "BVH is " + std::to_string(bout.depth) + " wide levels deep, more than the " + std::to_string(kStackCapacity) +
                    "-entry traversal stack holds (degenerate geometry, e.g. a long chain of nested slivers)");
    S.bvh = bout.nodes; S.prims = bout.prims; S.numPrims = bin.numPrims; S.numNodesBvh = (int)bout.numNodes;
    for (int a = 0; a < 3; ++a) { c->sceneLo[a] = bin.numPrims ? bout.lo[a] : 0.0f; c->sceneHi[a] = bin.numPrims ? bout.hi[a] : 1.0f; }
    if (c->l2Bytes && bout.numNodes > 0) {
        // nodes are stored level by level (breadth first), so a prefix window covers the top of the tree
        c->l2WindowBytes = std::min<size_t>(sizeof(WideNode) * (size_t)bout.numNodes, c->maxWindow);
        c->l2HitRatio = (float)std::min(1.0, (double)c->l2Bytes / (double)c->l2WindowBytes);
    }
    // materials
    if ((rc = devUpload(c, c->nodes.data(), c->nodes.size(), &S.nodes))) return rc;
    S.numNodes = (int)c->nodes.size();
    if ((rc = devUpload(c, c->pool.data(), c->pool.size(), &S.pool))) return rc;
    std::vector<DTexture> dtex;
    for (auto &t : c->textures) dtex.push_back(t.dt);
    if ((rc = devUpload(c, dtex.data(), dtex.size(), &S.textures))) return rc;
    S.numTextures = (int)dtex.size();
    // emitter pmf (src/librender/scene.cpp:375-380)
    std::vector<float> ecdf(1, 0.0f);
    for (auto &e : c->emitters) ecdf.push_back(ecdf.back() + e.d.sampling_weight);
    { float sum = ecdf.back(); if (!(sum > 0)) return fail(MTS_EINVAL, "emitter sampling weights sum to zero");
      float norm = 1.0f / sum; for (size_t j = 1; j < ecdf.size(); ++j) ecdf[j] *= norm; ecdf.back() = 1.0f; }
    std::vector<DEmitter> dem;
    for (size_t i = 0; i < c->emitters.size(); ++i) {
        const MtsEmitterDesc &d = c->emitters[i].d;
        DEmitter e; e.type = d.type; e.shape = d.shape;
        e.radiance = make_float3(d.radiance[0], d.radiance[1], d.radiance[2]);
        e.position = make_float3(d.position[0], d.position[1], d.position[2]);
        e.prob = ecdf[i + 1] - ecdf[i];
        dem.push_back(e);
    }
    if ((rc = devUpload(c, dem.data(), dem.size(), &S.emitters))) return rc;
    if ((rc = devUpload(c, ecdf.data(), ecdf.size(), &S.emitterCdf))) return rc;
    S.numEmitters = (int)dem.size();
    // environment map: EnvironmentMap::configure CDFs (src/emitters/envmap.cpp:260-329), fp32 sequential sums
    S.hasEnv = 0;
    if (c->envIndex >= 0) {
        HostEmitter &he = c->emitters[c->envIndex];
        const int w = he.d.env_width, h = he.d.env_height;
        std::vector<float> rowWeights(h);
        for (int y = 0; y < h; ++y) rowWeights[y] = std::sin((y + 0.5f) * kPi / h);
        DEnvmap &E = S.env;
        if ((rc = devUpload(c, rowWeights.data(), rowWeights.size(), &E.rowWeights))) return rc;
        auto devAlloc = [&](size_t bytes, void **out) -> int {
            CUDA_TRY(mts::devAlloc(out, bytes, c->stream)); c->sceneAllocs.push_back(*out); return MTS_OK;
        };
        __half *tex = nullptr; float *cdfCols = nullptr, *cdfRows = nullptr, *rowTerm = nullptr; uint16_t *gRows = nullptr, *gCols = nullptr;
        // the texels as a TMIPMap: repeat-u / clamp-v, EWA, maxAnisotropy 10, unbounded values (envmap.cpp:138-182)
        DTexture envMip;
        if ((rc = buildPyramid(c, he.envRaw, w, h, 3, MTS_WRAP_REPEAT, MTS_WRAP_CLAMP, MTS_FILTER_EWA, 10.0f, INFINITY, envMip))) return rc;
        c->sceneAllocs.push_back((void *)envMip.texels);
        if ((rc = devUpload(c, &envMip, 1, &E.mip))) return rc;
        tex = const_cast<__half *>(envMip.texels);   // k_env_rows rewrites level 0 with the same values
        if ((rc = devAlloc(sizeof(float) * (size_t)(w + 1) * h, (void **)&cdfCols))) return rc;
        if ((rc = devAlloc(sizeof(float) * (size_t)(h + 1), (void **)&cdfRows))) return rc;
        if ((rc = devAlloc(sizeof(float) * (size_t)(h + 1), (void **)&rowTerm))) return rc;   // [h] receives the total
        if ((rc = devAlloc(sizeof(uint16_t) * (size_t)(kGuideRows + 1), (void **)&gRows))) return rc;
        if ((rc = devAlloc(sizeof(uint16_t) * (size_t)h * (kGuideCols + 1), (void **)&gCols))) return rc;
        k_env_rows<<<(h + 63) / 64, 64, 0, c->stream>>>(he.envRaw, w, h, E.rowWeights, tex, cdfCols, rowTerm, gCols, kGuideCols);
        k_env_finish<<<1, 32, 0, c->stream>>>(rowTerm, h, cdfRows, gRows, kGuideRows, rowTerm + h);
        float rowSum = 0.0f;
        CUDA_TRY(cudaMemcpyAsync(&rowSum, rowTerm + h, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        if (rowSum == 0) return fail(MTS_EINVAL, "The environment map is completely black -- this is not allowed.");
        if (!std::isfinite(rowSum)) return fail(MTS_EINVAL, "The environment map contains an invalid floating point value (nan/inf) -- giving up.");
        E.texels = tex; E.cdfRows = cdfRows; E.cdfCols = cdfCols; E.guideRows = gRows; E.guideCols = gCols;
        E.width = w; E.height = h; E.scale = he.d.scale;
        E.normalization = 1.0f / (rowSum * (2 * kPi / w) * (kPi / h));
        M4 tw = m4From(he.d.to_world), tl = m4Inverse(tw);
        E.toWorld = toAffine(tw); E.toLocal = toAffine(tl);
        // m_sceneBSphere: scene AABB incl. the sensor position, radius * 1.5 (envmap.cpp:331-341, scene.cpp:386-398)
        float lo[3], hi[3];
        for (int a = 0; a < 3; ++a) { lo[a] = bin.numPrims ? bout.lo[a] : 0.0f; hi[a] = bin.numPrims ? bout.hi[a] : 0.0f; }
        if (bin.numPrims) {   // the kd-tree enlarges its AABB (include/mitsuba/render/gkdtree.h:1218-1221, eps = 1e-3)
            const float eps = 1e-3f;
            for (int a = 0; a < 3; ++a) {
                lo[a] = lo[a] - ((hi[a] - lo[a]) * eps + eps);
                hi[a] = hi[a] + ((hi[a] - lo[a]) * eps + eps);   // uses the already enlarged minimum, as the reference does
            }
        }
        if (c->hasCam) {
            const float camP[3] = {c->cam.camera_to_world[3], c->cam.camera_to_world[7], c->cam.camera_to_world[11]};
            for (int a = 0; a < 3; ++a) {
                if (!bin.numPrims) lo[a] = hi[a] = camP[a];
                lo[a] = std::min(lo[a], camP[a]); hi[a] = std::max(hi[a], camP[a]);
            }
        }
        float ctr[3], ext[3];
        for (int a = 0; a < 3; ++a) { ctr[a] = (lo[a] + hi[a]) * 0.5f; ext[a] = hi[a] - ctr[a]; }
        E.bsCenter = make_float3(ctr[0], ctr[1], ctr[2]);
        E.bsRadius = std::max(kEps, hLen(ext) * 1.5f);
        E.emitterIndex = c->envIndex; E.prob = ecdf[c->envIndex + 1] - ecdf[c->envIndex];
        S.hasEnv = 1;
    }
    if (c->upPending) { cudaEventRecord(c->upDone, c->upStream); cudaStreamWaitEvent(c->stream, c->upDone, 0); c->upPending = false; }
    cudaEventRecord(e1, c->stream);
    if ((rc = syncCheck(c, "build"))) return rc;
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    c->stats.build_ms = ms;
    c->stats.kernel_launches = (uint64_t)bout.launches;
    c->built = true;
    return MTS_OK;
}

MTS_API int mtscuda_bind_film(MtsContext *c, void *device_ptr) {
    if (!c->hasFilm) return fail(MTS_ESTATE, "bind_film before set_film");
    c->filmBound = (float4 *)device_ptr;
    c->ds.film = c->filmPtr();
    return MTS_OK;
}

// L2-persisting access window over the wide BVH nodes for the traversal kernels: the top levels of the tree are
// touched by every ray and should not be evicted by the streaming gathers of the shade stage.
static void setBvhWindow(MtsContext *c, cudaStream_t st, bool on) {
    if (!c->l2Bytes) return;
    cudaStreamAttrValue v;
    std::memset(&v, 0, sizeof(v));
    if (on) {
        v.accessPolicyWindow.base_ptr = (void *)c->ds.bvh;
        v.accessPolicyWindow.num_bytes = c->l2WindowBytes;
        v.accessPolicyWindow.hitRatio = c->l2HitRatio;
        v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    }
    cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v);
}

MTS_API int mtscuda_set_stream(MtsContext *c, void *cuda_stream) {
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->ownStream;
    return MTS_OK;
}

static int renderOne(MtsContext *c, const MtsRenderParams *P, volatile const int *cancel, MtsStats *stats) {
    if (!c->built || !c->hasCam || !c->hasFilm) return fail(MTS_ESTATE, "render before build/set_camera/set_film");
    // MonteCarloIntegrator parameter validation (src/librender/integrator.cpp:190-225)
    if (P->rr_depth <= 0) return fail(MTS_EINVAL, "'rrDepth' must be set to a value greater than zero!");
    if (P->max_depth <= 0 && P->max_depth != -1) return fail(MTS_EINVAL, "'maxDepth' must be set to -1 (infinite) or a value greater than zero!");
    if (P->spp < 0) return fail(MTS_EINVAL, "negative sample count");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    DScene &S = c->ds;
    S.film = c->filmPtr(); S.cam = c->cam;
    const int W = c->film.width, H = c->film.height;
    RenderConst R;
    R.maxDepth = P->max_depth; R.rrDepth = P->rr_depth; R.strictNormals = P->strict_normals; R.hideEmitters = P->hide_emitters;
    R.seed = P->seed; R.sampleOffset = P->sample_offset; R.sampleStride = P->sample_stride > 0 ? P->sample_stride : 1;
    R.tilesX = (W + 7) / 8; R.tilesY = (H + 3) / 4; R.paddedPixels = (uint32_t)R.tilesX * R.tilesY * 32u;
    R.diffScale = 1.0f / std::sqrt((float)(P->sample_count > 0 ? P->sample_count : P->spp * R.sampleStride));   // integrator.cpp:144-145
    const unsigned long long totalItems = (unsigned long long)R.paddedPixels * (unsigned long long)P->spp;
    uint32_t cap = P->wavefront_size > 0 ? (uint32_t)P->wavefront_size : (1u << 25);   // 32 Mi path slots (5.5 GB of state): large wavefronts amortise the latency-bound tail bounces
    cap = (uint32_t)std::min<unsigned long long>(std::max<unsigned long long>(totalItems, 32), cap);
    cap = (cap + 31u) & ~31u;
    int rc;
    c->stagedShade = getenv("MTS_SHADE_FUSED") ? false : getenv("MTS_SHADE_STAGED") ? true : c->leafTextured;
    c->deferEwa = getenv("MTS_EWA_INLINE") == nullptr;
    c->raySort = getenv("MTS_RAY_SORT") != nullptr;
    if ((rc = ensureWavefront(c, cap))) return rc;
    const Wavefront &Wf = c->wf;
    if (P->clear_film) CUDA_TRY(cudaMemsetAsync(S.film, 0, sizeof(float4) * (size_t)W * H, st));
    S.filmAlpha = c->filmAlpha;
    if (P->clear_film && c->filmAlpha) CUDA_TRY(cudaMemsetAsync(c->filmAlpha, 0, sizeof(float) * (size_t)W * H, st));
    CUDA_TRY(cudaMemsetAsync(Wf.ctr, 0, sizeof(Counters), st));
    CUDA_TRY(cudaMemsetAsync(Wf.keyCount, 0, 4 * 3 * (size_t)Wf.numKeys, st));

    // CUDA events around every stage (pooled in the context: creating ~50 events per frame showed up in short renders)
    struct Span { cudaEvent_t a, b; int kind; };
    std::vector<Span> spans;
    c->eventsUsed = 0;
    auto newEvent = [&]() {
        if (c->eventsUsed == c->eventPool.size()) { cudaEvent_t e; cudaEventCreate(&e); c->eventPool.push_back(e); }
        return c->eventPool[c->eventsUsed++];
    };
    auto begin = [&](int kind) { Span s; s.a = newEvent(); s.b = newEvent(); s.kind = kind; cudaEventRecord(s.a, st); spans.push_back(s); };
    auto end = [&]() { cudaEventRecord(spans.back().b, st); };
    cudaEvent_t ev0 = newEvent(), ev1 = newEvent();
    cudaEvent_t pollEv[kPollRing];
    for (int i = 0; i < kPollRing; ++i) pollEv[i] = newEvent();
    cudaEventRecord(ev0, st);

    // 4 CTAs/SM (128 registers) measured best on cfg3; 3 (168 registers, fewer spills) kept as the measured alternative
    const bool traceBounces = getenv("MTS_TRACE_BOUNCES") != nullptr;   // per-bounce queue sizes on stderr (matches ncu launches to ray counts)
    const int gridTrace = c->numSMs * TRACE_MINB, gridWide = c->numSMs * 8;
    uint64_t launches = 0;
    bool cancelled = false;
    // The bounce loop never waits for the device.  Every kernel reads its trip count from device memory, so a bounce
    // whose queue is empty costs a few empty launches; the loop runs maxDepth bounces (a path of depth maxDepth spawns
    // no further ray, path.cpp:156-165) and leaves early as soon as a *non-blocking* look at a queue size copied back
    // a few bounces ago shows zero.  Only unbounded renders (maxDepth = -1) fall back to waiting, after kPatience bounces.
    const int kPatience = 48, kGiveUp = 4096;
    for (unsigned long long first = 0; first < totalItems && !cancelled; first += cap) {
        uint32_t count = (uint32_t)std::min<unsigned long long>(cap, totalItems - first);
        int cur = 0;
        k_begin_wavefront<<<1, 1, 0, st>>>(Wf.ctr);
        begin(0);
        k_raygen<<<gridWide, 256, 0, st>>>(S, Wf, R, first, count, 0);
        end();
        launches += 2;
        for (int i = 0; i < kPollRing; ++i) c->hostPoll[i] = 0xffffffffu;
        for (int bounce = 0; P->max_depth < 0 || bounce < P->max_depth; ++bounce) {
            if (cancel && *cancel) { cancelled = true; break; }
            if (bounce >= 1) {
                // slot (j % ring) = number of rays bounce j left for bounce j + 1; a queue that was empty once stays empty
                bool empty = false;
                for (int j = bounce - 1; j >= 0 && j > bounce - 4 && !empty; --j)
                    empty = cudaEventQuery(pollEv[j % kPollRing]) == cudaSuccess && c->hostPoll[j % kPollRing] == 0;
                if (!empty && P->max_depth < 0 && bounce >= kPatience) {
                    CUDA_TRY(cudaEventSynchronize(pollEv[(bounce - 1) % kPollRing]));
                    empty = c->hostPoll[(bounce - 1) % kPollRing] == 0;
                }
                cudaGetLastError();   // cudaErrorNotReady of the queries is not an error
                if (empty) break;
                if (bounce > kGiveUp) {
                    k_abandon<<<gridWide, 256, 0, st>>>(Wf, cur);
                    k_splat<<<gridWide, 256, 0, st>>>(S, Wf, R);
                    launches += 2;
                    break;
                }
            }
            k_begin_bounce<<<1, 1, 0, st>>>(Wf.ctr, cur);
            Wavefront Wb = Wf;   // this bounce's view of the wavefront
            if (Wf.raySortBin && bounce >= 1) {
                // ray sort: queue[cur] -> `sorted` by (octant, origin cell); for the rest of the bounce the two buffers swap roles
                // (k_intersect reads the reordered queue, the material sort writes back into the old queue buffer)
                begin(5);
                Wavefront Ws = Wf; Ws.keyCount = Wf.raySortBin; Ws.numKeys = kRaySortKeys;
                RaySortGrid G;
                G.lo = make_float3(c->sceneLo[0], c->sceneLo[1], c->sceneLo[2]);
                G.invCell = make_float3(4.0f / std::max(c->sceneHi[0] - c->sceneLo[0], 1e-20f), 4.0f / std::max(c->sceneHi[1] - c->sceneLo[1], 1e-20f),
                                        4.0f / std::max(c->sceneHi[2] - c->sceneLo[2], 1e-20f));
                k_raysort_keys<<<gridWide, 256, 0, st>>>(Ws, cur, G);
                k_raysort_offsets<<<1, 32, 0, st>>>(Ws);
                k_sort_scatter<<<gridWide, 256, 0, st>>>(Ws, cur);
                end();
                launches += 3;
                Wb.queue[cur] = Wf.sorted; Wb.sorted = Wf.queue[cur];
            }
            begin(1);
            setBvhWindow(c, st, true);
            if (P->instrument) k_intersect<true><<<gridTrace, 128, 0, st>>>(S, Wb, cur);
            else k_intersect<false><<<gridTrace, 128, 0, st>>>(S, Wb, cur);
            setBvhWindow(c, st, false);
            end();
            begin(5);
            k_sort_offsets<<<1, 32, 0, st>>>(Wb);
            k_sort_scatter<<<gridWide, 256, 0, st>>>(Wb, cur);
            end();
            begin(6); k_shade_miss<<<gridWide, 256, 0, st>>>(S, Wb, R, cur); end();
            begin(2);
            // one instance per shade class that has BSDFs in this scene (each walks its own range of the sorted queue)
            if (Wb.classKey[SHADE_LAMBERT] < Wb.classKey[SHADE_LAMBERT + 1]) { launchShade<SHADE_LAMBERT>(S, Wb, R, cur, c->numSMs, st); ++launches; }
            if (Wb.classKey[SHADE_LEAF] < Wb.classKey[SHADE_LEAF + 1]) {
                if (Wb.rec) {
                    launchSurface(S, Wb, R, cur, c->numSMs, st);
                    if (Wb.ewaJobs && bounce == 0) { launchEwa(S, Wb, c->numSMs, st); launches += 4; }   // only camera rays carry differentials (path.cpp:225)
                    launchScatter(S, Wb, R, cur, c->numSMs, st);
                    launches += 2;
                }
                else { launchShade<SHADE_LEAF>(S, Wb, R, cur, c->numSMs, st); ++launches; }
            }
            if (Wb.classKey[SHADE_GENERIC] < Wb.classKey[SHADE_GENERIC + 1]) { launchShade<SHADE_GENERIC>(S, Wb, R, cur, c->numSMs, st); ++launches; }
            end();
            begin(3);
            setBvhWindow(c, st, true);
            if (P->instrument) k_shadow<true><<<gridTrace, 128, 0, st>>>(S, Wb);
            else k_shadow<false><<<gridTrace, 128, 0, st>>>(S, Wb);
            setBvhWindow(c, st, false);
            end();
            begin(4); k_splat<<<gridWide, 256, 0, st>>>(S, Wb, R); end();
            launches += 7;
            CUDA_TRY(cudaMemcpyAsync(&c->hostPoll[bounce % kPollRing], &Wf.ctr->qCount[cur ^ 1], sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaEventRecord(pollEv[bounce % kPollRing], st));
            if (traceBounces) {
                Counters hc;
                CUDA_TRY(cudaMemcpyAsync(&hc, Wf.ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
                CUDA_TRY(cudaStreamSynchronize(st));
                std::vector<uint32_t> off(Wf.numKeys + 1);
                CUDA_TRY(cudaMemcpyAsync(off.data(), Wf.keyCount + Wf.numKeys, sizeof(uint32_t) * Wf.numKeys, cudaMemcpyDeviceToHost, st));
                CUDA_TRY(cudaStreamSynchronize(st));
                off[Wf.numKeys] = hc.qCount[cur];
                uint32_t perClass[SHADE_CLASSES];
                for (int k = 0; k < SHADE_CLASSES; ++k) perClass[k] = off[std::min(Wf.classKey[k + 1], Wf.numKeys)] - off[std::min(Wf.classKey[k], Wf.numKeys)];
                fprintf(stderr, "[mtscuda] wavefront %llu bounce %d: closest-hit rays %u, shadow rays %u, continuing %u, escaped so far %llu, "
                        "shaded lambert/leaf/generic %u %u %u\n",
                        first / cap, bounce, hc.qCount[cur], hc.shadowCount, hc.qCount[cur ^ 1], hc.missRays, perClass[0], perClass[1], perClass[2]);
            }
            cur ^= 1;
        }
    }
    k_begin_wavefront<<<1, 1, 0, st>>>(Wf.ctr);   // folds the last bounce's shadow rays into the totals
    ++launches;
    cudaEventRecord(ev1, st);
    if ((rc = syncCheck(c, "render"))) return rc;
    MtsStats &sx = c->stats;
    float build_ms = sx.build_ms;
    std::memset(&sx, 0, sizeof(sx));
    sx.build_ms = build_ms;
    cudaEventElapsedTime(&sx.render_ms, ev0, ev1);
    for (auto &s : spans) {
        float ms = 0; cudaEventElapsedTime(&ms, s.a, s.b);
        float *dst = s.kind == 0 ? &sx.raygen_ms : s.kind == 1 ? &sx.intersect_ms : s.kind == 2 ? &sx.shade_ms : s.kind == 3 ? &sx.shadow_ms :
                     s.kind == 4 ? &sx.splat_ms : s.kind == 5 ? &sx.sort_ms : &sx.miss_ms;
        *dst += ms;
        if (s.kind == 1) sx.intersect_launches++; else if (s.kind == 2) sx.shade_launches++;
        else if (s.kind == 3) sx.shadow_launches++; else if (s.kind == 4) sx.splat_launches++;
    }
    Counters hc;
    CUDA_TRY(cudaMemcpy(&hc, Wf.ctr, sizeof(Counters), cudaMemcpyDeviceToHost));
    sx.camera_samples = (uint64_t)W * H * (uint64_t)P->spp;
    sx.normal_rays = hc.normalRays; sx.shadow_rays = hc.shadowRays;
    sx.path_length_sum = hc.pathLenSum;
    sx.miss_paths = hc.missRays; sx.shaded_paths = hc.normalRays - hc.missRays;
    sx.kernel_launches = launches;
    sx.bvh_nodes_visited = hc.nodesVisited; sx.bvh_prims_tested = hc.primsTested;
    sx.bvh_nodes_visited_shadow = hc.nodesVisitedShadow; sx.bvh_prims_tested_shadow = hc.primsTestedShadow;
    if (stats) *stats = sx;
    return cancelled ? fail(MTS_ECANCELLED, "render cancelled") : MTS_OK;
}

// ---- multi-device groups --------------------------------------------------------------------------------------------
// The reference scales a render over the cores of a box with `-p N` (src/mitsuba/mitsuba.cpp:277-327: one worker per
// core, image blocks handed out by RenderProcess).  cudapath scales it over the GPUs of a box: every device renders the
// FULL frame with its own subset of the sample indices (sample s goes to device s mod N, so the union is exactly the
// single-device sample set), and the films are summed on the primary over NVLink.  One host thread per device.
} // extern "C"
namespace {

// Runs fn(device index) for every device of the group concurrently; returns the first error (code + message).
template <class Fn> int forEachDevice(MtsContext *c, Fn fn) {
    const int n = 1 + (int)c->peers.size();
    std::vector<int> rc(n, MTS_OK);
    std::vector<std::string> msg(n);
    std::vector<std::thread> th;
    for (int d = 1; d < n; ++d)
        th.emplace_back([&, d]() { rc[d] = fn(d); if (rc[d] != MTS_OK) msg[d] = g_err; });   // g_err is thread-local
    rc[0] = fn(0);
    if (rc[0] != MTS_OK) msg[0] = g_err;
    for (auto &t : th) t.join();
    for (int d = 0; d < n; ++d) if (rc[d] != MTS_OK) return fail(rc[d], "device " + std::to_string(d ? c->peers[d - 1]->device : c->device) + ": " + msg[d]);
    return MTS_OK;
}

template <class T> cudaError_t peerCopy(DevVec<T> &dst, int dstDev, cudaStream_t dstStream, const DevVec<T> &src, int srcDev) {
    dst.release(dstStream);
    if (!src.size) return cudaSuccess;
    cudaError_t e = dst.reserve(src.size, dstStream);
    if (e != cudaSuccess) return e;
    dst.size = src.size;
    return cudaMemcpyPeerAsync(dst.data, dstDev, src.data, srcDev, sizeof(T) * src.size, dstStream);
}

// Gives peer `p` the staged scene of the primary: host-side descriptions are copied, device-resident inputs (vertex /
// index / tangent arrays, texture pyramids, raw envmap texels) travel GPU to GPU.  The peer then runs its own build().
int replicateScene(const MtsContext *c, MtsContext *p) {
    CUDA_TRY(cudaSetDevice(p->device));
    cudaStream_t st = p->stream;
    p->freeScene(); p->freeGeometry();
    p->cam = c->cam; p->hasCam = c->hasCam;
    if (c->hasFilm && (!p->hasFilm || std::memcmp(&p->film, &c->film, sizeof(MtsFilm)) != 0)) {
        int rc = mtscuda_set_film(p, &c->film);
        if (rc != MTS_OK) return rc;
    }
    p->nodes = c->nodes; p->pool = c->pool; p->shapes = c->shapes; p->envIndex = c->envIndex;
    CUDA_TRY(peerCopy(p->positions, p->device, st, c->positions, c->device));
    CUDA_TRY(peerCopy(p->normals, p->device, st, c->normals, c->device));
    CUDA_TRY(peerCopy(p->texcoords, p->device, st, c->texcoords, c->device));
    CUDA_TRY(peerCopy(p->tangents, p->device, st, c->tangents, c->device));
    CUDA_TRY(peerCopy(p->indices, p->device, st, c->indices, c->device));
    for (const HostTex &t : c->textures) {
        HostTex q = t;
        __half *px = nullptr;
        CUDA_TRY(devAlloc(&px, t.bytes, st));
        q.dt.texels = px; q.dt.weightLut = p->ewaLut;
        p->textures.push_back(q);
        CUDA_TRY(cudaMemcpyPeerAsync(px, p->device, t.dt.texels, c->device, t.bytes, st));
    }
    for (const HostEmitter &e : c->emitters) {
        HostEmitter q; q.d = e.d;
        if (e.envRaw) {
            const size_t bytes = sizeof(float) * 3 * (size_t)e.d.env_width * e.d.env_height;
            CUDA_TRY(devAlloc(&q.envRaw, bytes, st));
            CUDA_TRY(cudaMemcpyPeerAsync(q.envRaw, p->device, e.envRaw, c->device, bytes, st));
        }
        p->emitters.push_back(q);
    }
    CUDA_TRY(cudaStreamSynchronize(st));
    return MTS_OK;
}

// film[primary] += sum of the peers' films, read over NVLink (peer access is enabled by create_multi); fixed order
struct PeerFilms { const float4 *film[15]; const float *alpha[15]; int n; };
__global__ void k_film_reduce(float4 *__restrict__ dst, float *__restrict__ dstAlpha, const PeerFilms P, size_t npix) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
        float4 a = dst[i];
        for (int k = 0; k < P.n; ++k) { const float4 b = P.film[k][i]; a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
        dst[i] = a;
        if (dstAlpha) { float v = dstAlpha[i]; for (int k = 0; k < P.n; ++k) v += P.alpha[k][i]; dstAlpha[i] = v; }
    }
}

} // namespace
extern "C" {

MTS_API int mtscuda_create_multi(const int32_t *devices, int32_t n, MtsContext **out) {
    if (!out || !devices || n < 1) return fail(MTS_EINVAL, "create_multi: bad arguments");
    if (n > 16) return fail(MTS_EINVAL, "create_multi: at most 16 devices");
    for (int i = 0; i < n; ++i) for (int j = 0; j < i; ++j) if (devices[i] == devices[j]) return fail(MTS_EINVAL, "create_multi: device listed twice");
    MtsContext *c = nullptr;
    int rc = mtscuda_create(devices[0], &c);
    if (rc != MTS_OK) return rc;
    for (int i = 1; i < n; ++i) {
        MtsContext *p = nullptr;
        rc = mtscuda_create(devices[i], &p);
        if (rc == MTS_OK) {
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[0], devices[i]);
            if (!can) { rc = fail(MTS_EINVAL, "create_multi: no peer access between devices " + std::to_string(devices[0]) + " and " + std::to_string(devices[i])); mtscuda_destroy(p); }
        }
        if (rc != MTS_OK) { mtscuda_destroy(c); return rc; }
        c->peers.push_back(p);
        // the primary reads the peers' films; the peers pull the scene from the primary
        cudaSetDevice(devices[0]); if (cudaDeviceEnablePeerAccess(devices[i], 0) != cudaSuccess) cudaGetLastError();   // already enabled is fine
        cudaSetDevice(devices[i]); if (cudaDeviceEnablePeerAccess(devices[0], 0) != cudaSuccess) cudaGetLastError();
        // The peers' films are read by the primary's reduce kernel.  Memory of the stream-ordered pools (devmem.h) is private to
        // its device -- cudaDeviceEnablePeerAccess covers cudaMalloc memory only, and opening the pools with
        // cudaMemPoolSetAccess made the large allocations of a 10 M-triangle scene fail with "out of memory" -- so the film of
        // a peer context is a plain cudaMalloc allocation (set_film).
        p->peerFilm = true;
    }
    cudaSetDevice(devices[0]);
    *out = c;
    return MTS_OK;
}

MTS_API int mtscuda_device_count(MtsContext *c) { return 1 + (int)c->peers.size(); }

MTS_API int mtscuda_build(MtsContext *c) {
    if (c->peers.empty()) return buildOne(c);
    // peers copy the staged inputs and build their own acceleration structure while the primary builds its own
    return forEachDevice(c, [&](int d) -> int {
        if (d == 0) return buildOne(c);
        MtsContext *p = c->peers[d - 1];
        int rc = replicateScene(c, p);
        return rc != MTS_OK ? rc : buildOne(p);
    });
}

MTS_API int mtscuda_render(MtsContext *c, const MtsRenderParams *P, volatile const int *cancel, MtsStats *stats) {
    if (c->peers.empty()) return renderOne(c, P, cancel, stats);
    if (!c->built) return fail(MTS_ESTATE, "call build() first");
    const int n = 1 + (int)c->peers.size();
    const int stride = P->sample_stride > 0 ? P->sample_stride : 1;
    std::vector<MtsStats> st(n);
    std::vector<MtsRenderParams> pp(n, *P);
    for (int d = 0; d < n; ++d) {
        // device d renders samples d, d + n, d + 2n, ... of this call's sample set
        pp[d].spp = P->spp / n + (d < P->spp % n ? 1 : 0);
        pp[d].sample_offset = P->sample_offset + d * stride;
        pp[d].sample_stride = stride * n;
        pp[d].sample_count = P->sample_count > 0 ? P->sample_count : P->spp * stride;
        if (d > 0) pp[d].clear_film = 1;
        std::memset(&st[d], 0, sizeof(MtsStats));
    }
    int rc = forEachDevice(c, [&](int d) -> int {
        MtsContext *x = d ? c->peers[d - 1] : c;
        if (pp[d].spp == 0) {
            if (pp[d].clear_film) {   // nothing to render here, but the film that gets reduced must not hold an earlier frame
                cudaSetDevice(x->device);
                cudaMemsetAsync(x->filmPtr(), 0, sizeof(float4) * (size_t)x->film.width * x->film.height, x->stream);
                if (x->filmAlpha) cudaMemsetAsync(x->filmAlpha, 0, sizeof(float) * (size_t)x->film.width * x->film.height, x->stream);
                cudaStreamSynchronize(x->stream);
            }
            return MTS_OK;
        }
        return renderOne(x, &pp[d], cancel, &st[d]);
    });
    if (rc != MTS_OK && rc != MTS_ECANCELLED) return rc;
    // film reduce on the primary over NVLink
    CUDA_TRY(cudaSetDevice(c->device));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    PeerFilms pf; pf.n = n - 1;
    for (int k = 0; k < n - 1; ++k) { pf.film[k] = c->peers[k]->filmPtr(); pf.alpha[k] = c->peers[k]->filmAlpha; }
    const size_t npix = (size_t)c->film.width * c->film.height;
    cudaEventRecord(e0, c->stream);
    k_film_reduce<<<c->numSMs * 4, 256, 0, c->stream>>>(c->filmPtr(), c->filmAlpha, pf, npix);
    cudaEventRecord(e1, c->stream);
    int rc2 = syncCheck(c, "film reduce");
    float reduceMs = 0; cudaEventElapsedTime(&reduceMs, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rc2 != MTS_OK) return rc2;
    // merged statistics: counts add up, times are those of the slowest device
    MtsStats m = st[0];
    for (int d = 1; d < n; ++d) {
        const MtsStats &s = st[d];
        m.camera_samples += s.camera_samples; m.normal_rays += s.normal_rays; m.shadow_rays += s.shadow_rays;
        m.path_length_sum += s.path_length_sum; m.kernel_launches += s.kernel_launches; m.bvh_nodes_visited += s.bvh_nodes_visited;
        m.bvh_prims_tested += s.bvh_prims_tested; m.bvh_nodes_visited_shadow += s.bvh_nodes_visited_shadow; m.bvh_prims_tested_shadow += s.bvh_prims_tested_shadow;
        m.shaded_paths += s.shaded_paths; m.miss_paths += s.miss_paths;
        m.render_ms = std::max(m.render_ms, s.render_ms);
    }
    m.render_ms += reduceMs;
    m.kernel_launches += 1;
    m.build_ms = c->stats.build_ms;
    c->stats = m;
    if (stats) *stats = m;
    return rc;
}

MTS_API int mtscuda_read_film(MtsContext *c, float *out) {
    if (!c->hasFilm) return fail(MTS_ESTATE, "no film");
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    CUDA_TRY(cudaMemcpy(out, c->filmPtr(), sizeof(float4) * (size_t)c->film.width * c->film.height, cudaMemcpyDeviceToHost));
    return MTS_OK;
}

MTS_API int mtscuda_read_film_alpha(MtsContext *c, float *out) {
    if (!c->hasFilm || !c->filmAlpha) return fail(MTS_ESTATE, "the film has no alpha plane (MtsFilm.has_alpha = 0)");
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    CUDA_TRY(cudaMemcpy(out, c->filmAlpha, sizeof(float) * (size_t)c->film.width * c->film.height, cudaMemcpyDeviceToHost));
    return MTS_OK;
}

MTS_API int mtscuda_get_stats(MtsContext *c, MtsStats *s) { *s = c->stats; return MTS_OK; }

// ---- per-function entry points --------------------------------------------------------------
#define HOOK_PREAMBLE(needBuilt) \
    if ((needBuilt) && !c->built) return fail(MTS_ESTATE, "call build() first"); \
    CUDA_TRY(cudaSetDevice(c->device)); \
    cudaStream_t st = c->stream; Staged sg(st); c->ds.film = c->filmPtr(); c->ds.cam = c->cam;
#define NEED(p) if (!(p)) return fail(MTS_ECUDA, "device allocation failed")

MTS_API int mtscuda_sample_rays(MtsContext *c, int64_t n, const float *sp, float *out) {
    if (!c->hasCam) return fail(MTS_ESTATE, "no camera");
    HOOK_PREAMBLE(false)
    float *dsp = sg.in(sp, 2 * n, st), *dout = sg.out<float>(16 * n); NEED(dsp && dout);
    k_hook_sample_rays<<<gridFor(n), 128, 0, st>>>(c->ds, n, dsp, dout);
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 16 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "sample_rays");
}

MTS_API int mtscuda_trace(MtsContext *c, int64_t n, const float *rays, int32_t shadow, float *hits, int32_t *ids) {
    HOOK_PREAMBLE(true)
    if (n <= 0) return MTS_OK;
    const uint32_t chunk = (uint32_t)std::min<int64_t>(n, 1 << 22);
    int rc;
    if ((rc = ensureWavefront(c, (chunk + 31u) & ~31u))) return rc;
    const Wavefront &Wf = c->wf;
    float *dr = sg.out<float>(8 * (size_t)chunk), *dh = sg.out<float>(4 * (size_t)chunk); int32_t *di = sg.out<int32_t>(2 * (size_t)chunk);
    NEED(dr && dh && di);
    CUDA_TRY(cudaMemsetAsync(Wf.ctr, 0, sizeof(Counters), st));
    CUDA_TRY(cudaMemsetAsync(Wf.keyCount, 0, 4 * 3 * (size_t)Wf.numKeys, st));
    const int gridTrace = c->numSMs * TRACE_MINB;
    std::vector<int32_t> idTmp;
    for (int64_t first = 0; first < n; first += chunk) {
        const uint32_t m = (uint32_t)std::min<int64_t>(chunk, n - first);
        CUDA_TRY(cudaMemcpyAsync(dr, rays + 8 * first, sizeof(float) * 8 * m, cudaMemcpyHostToDevice, st));
        k_trace_load<<<gridFor(m), 128, 0, st>>>(Wf, m, dr, shadow);
        if (shadow) k_shadow<false><<<gridTrace, 128, 0, st>>>(c->ds, Wf);
        else k_intersect<false><<<gridTrace, 128, 0, st>>>(c->ds, Wf, 0);
        k_trace_export<<<gridFor(m), 128, 0, st>>>(c->ds, Wf, m, shadow, dh, di);
        CUDA_TRY(cudaMemcpyAsync(hits + 4 * first, dh, sizeof(float) * 4 * m, cudaMemcpyDeviceToHost, st));
        if (ids) CUDA_TRY(cudaMemcpyAsync(ids + 2 * first, di, sizeof(int32_t) * 2 * m, cudaMemcpyDeviceToHost, st));
        if ((rc = syncCheck(c, "trace"))) return rc;
    }
    CUDA_TRY(cudaMemsetAsync(Wf.keyCount, 0, 4 * 3 * (size_t)Wf.numKeys, st));   // k_intersect left its key histogram behind
    return syncCheck(c, "trace");
}

// ---- mtscuda_trace_noalloc: the trace entry point for callers that must not allocate -------------------------------
// No counterpart in the reference; added (CUDA-graph capture, tight serving loops
// and harnesses that forbid cudaMalloc).  Same kernels and results as mtscuda_trace, but the wavefront slice the queue
// kernels work in (ray state, hit records, queues, sort keys, counters) and the staging of the rays / results live in ONE
// file-scope device array of fixed size instead of pool allocations: at most MTS_NOALLOC_MAX_RAYS rays per call, scenes
// with at most kNoAllocMaxKeys material sort keys.  Every GPU that loads the library holds the array (2.9 MB) until exit.
// All calls on a device share it -- calls that overlap on the GPU (different streams / contexts) race -- and it keeps
// the previous call's contents: the counters and the key histogram are cleared at the start of every call, everything
// else is written before it is read.
constexpr uint32_t kNoAllocMaxRays = MTS_NOALLOC_MAX_RAYS;
constexpr uint32_t kNoAllocMaxKeys = 512;
struct NoAllocScratch {
    float4 rayO[kNoAllocMaxRays], rayD[kNoAllocMaxRays], hit[kNoAllocMaxRays];                           // closest-hit ray state
    float4 shO[kNoAllocMaxRays], shD[kNoAllocMaxRays], shC[kNoAllocMaxRays], L[kNoAllocMaxRays];         // shadow-ray state
    float raysIn[8 * kNoAllocMaxRays], hitsOut[4 * kNoAllocMaxRays];                                     // staging of the caller's buffers
    int32_t idsOut[2 * kNoAllocMaxRays];
    uint32_t queue[kNoAllocMaxRays], keys[kNoAllocMaxRays];
    uint32_t keyCount[3 * kNoAllocMaxKeys];
    Counters ctr;
};
static __device__ __align__(256) unsigned char g_noAllocScratch[sizeof(NoAllocScratch)];
namespace {
std::mutex g_noAllocMutex;
void *g_noAllocAddr[64] = {};   // device address of g_noAllocScratch per device ordinal (the name itself is a host-side stand-in)
// once per device, with that device current
int noAllocScratch(MtsContext *c, NoAllocScratch **out) {
    if (c->device < 0 || c->device >= 64) return fail(MTS_EINVAL, "trace_noalloc: device ordinal out of range");
    CUDA_TRY(cudaSetDevice(c->device));
    std::lock_guard<std::mutex> lock(g_noAllocMutex);
    if (!g_noAllocAddr[c->device]) {
        void *p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_noAllocScratch));
        if (!p) return fail(MTS_ECUDA, "trace_noalloc: cudaGetSymbolAddress returned a null address");
        g_noAllocAddr[c->device] = p;
    }
    *out = reinterpret_cast<NoAllocScratch *>(g_noAllocAddr[c->device]);
    return MTS_OK;
}
} // namespace

MTS_API int mtscuda_trace_noalloc(MtsContext *c, int64_t n, const float *rays, int32_t shadow, float *hits, int32_t *ids) {
    if (!c->built) return fail(MTS_ESTATE, "call build() first");
    if (n < 0 || n > (int64_t)kNoAllocMaxRays)
        return fail(MTS_EINVAL, "trace_noalloc: " + std::to_string(n) + " rays exceed the fixed scratch of " + std::to_string(kNoAllocMaxRays) +
                                " rays per call (MTS_NOALLOC_MAX_RAYS); split the batch or use mtscuda_trace");
    if (c->numKeys > kNoAllocMaxKeys)
        return fail(MTS_EINVAL, "trace_noalloc: the scene has " + std::to_string(c->numKeys) + " material sort keys, the fixed scratch holds " +
                                std::to_string(kNoAllocMaxKeys) + "; use mtscuda_trace");
    if (n == 0) return MTS_OK;
    if (!rays || !hits) return fail(MTS_EINVAL, "trace_noalloc: rays and hits must not be NULL");
    NoAllocScratch *s = nullptr;   // a DEVICE pointer: only its members' addresses are taken on the host
    int rc;
    if ((rc = noAllocScratch(c, &s))) return rc;
    cudaStream_t st = c->stream;
    Wavefront W;
    std::memset(&W, 0, sizeof(W));
    W.rayO = s->rayO; W.rayD = s->rayD; W.hit = s->hit; W.shO = s->shO; W.shD = s->shD; W.shC = s->shC; W.L = s->L;
    W.queue[0] = W.queue[1] = s->queue; W.keys = s->keys; W.keyCount = s->keyCount; W.numKeys = c->numKeys;
    for (int k = 0; k <= SHADE_CLASSES; ++k) W.classKey[k] = c->classKey[k];
    W.ctr = &s->ctr; W.capacity = kNoAllocMaxRays;
    const uint32_t m = (uint32_t)n;
    // what must start at zero: the counters / cursors and the key histogram k_intersect adds to
    CUDA_TRY(cudaMemsetAsync(&s->ctr, 0, sizeof(Counters), st));
    CUDA_TRY(cudaMemsetAsync(s->keyCount, 0, sizeof(uint32_t) * 3 * (size_t)c->numKeys, st));
    CUDA_TRY(cudaMemcpyAsync(s->raysIn, rays, sizeof(float) * 8 * m, cudaMemcpyDefault, st));    // host or device memory (UVA)
    k_trace_load<<<gridFor(m), 128, 0, st>>>(W, m, s->raysIn, shadow);
    const int gridTrace = c->numSMs * TRACE_MINB;
    if (shadow) k_shadow<false><<<gridTrace, 128, 0, st>>>(c->ds, W);
    else k_intersect<false><<<gridTrace, 128, 0, st>>>(c->ds, W, 0);
    k_trace_export<<<gridFor(m), 128, 0, st>>>(c->ds, W, m, shadow, s->hitsOut, s->idsOut);
    CUDA_TRY(cudaMemcpyAsync(hits, s->hitsOut, sizeof(float) * 4 * m, cudaMemcpyDefault, st));
    if (ids) CUDA_TRY(cudaMemcpyAsync(ids, s->idsOut, sizeof(int32_t) * 2 * m, cudaMemcpyDefault, st));
    return syncCheck(c, "trace_noalloc");
}

MTS_API int mtscuda_hit_records(MtsContext *c, int64_t n, const float *rays, float *out, int32_t *ids) {
    HOOK_PREAMBLE(true)
    float *dr = sg.in(rays, 8 * n, st), *dout = sg.out<float>(24 * n); int32_t *di = sg.out<int32_t>(2 * n); NEED(dr && dout && di);
    k_hook_hit_records<<<gridFor(n), 128, 0, st>>>(c->ds, n, dr, dout, di);
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 24 * n, cudaMemcpyDeviceToHost, st));
    if (ids) CUDA_TRY(cudaMemcpyAsync(ids, di, sizeof(int32_t) * 2 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "hit_records");
}

MTS_API int mtscuda_bsdf_eval_pdf(MtsContext *c, int32_t bsdf, int64_t n, const float *surf, const float *wi, const float *wo,
                                  int32_t measure, float *out_value, float *out_pdf) {
    if (bsdf < 0 || bsdf >= (int)c->nodes.size()) return fail(MTS_EINVAL, "bad bsdf index");
    HOOK_PREAMBLE(true)
    float *ds_ = sg.in(surf, 16 * n, st), *dwi = sg.in(wi, 3 * n, st), *dwo = sg.in(wo, 3 * n, st);
    float *dv = sg.out<float>(3 * n), *dp = sg.out<float>(n); NEED(ds_ && dwi && dwo && dv && dp);
    switch (shadeClassOf(c->nodes.data(), bsdf)) {   // the same code path k_shade<CLASS> runs for this BSDF
    case SHADE_LAMBERT: launchBsdfEvalPdf<SHADE_LAMBERT>(c->ds, bsdf, n, ds_, dwi, dwo, measure, dv, dp, st); break;
    case SHADE_LEAF: launchBsdfEvalPdf<SHADE_LEAF>(c->ds, bsdf, n, ds_, dwi, dwo, measure, dv, dp, st); break;
    default: launchBsdfEvalPdf<SHADE_GENERIC>(c->ds, bsdf, n, ds_, dwi, dwo, measure, dv, dp, st); break;
    }
    CUDA_TRY(cudaMemcpyAsync(out_value, dv, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(out_pdf, dp, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "bsdf_eval_pdf");
}

MTS_API int mtscuda_bsdf_sample(MtsContext *c, int32_t bsdf, int64_t n, const float *surf, const float *wi, const float *u, float *out) {
    if (bsdf < 0 || bsdf >= (int)c->nodes.size()) return fail(MTS_EINVAL, "bad bsdf index");
    HOOK_PREAMBLE(true)
    float *ds_ = sg.in(surf, 16 * n, st), *dwi = sg.in(wi, 3 * n, st), *du = sg.in(u, 2 * n, st), *dout = sg.out<float>(12 * n);
    NEED(ds_ && dwi && du && dout);
    switch (shadeClassOf(c->nodes.data(), bsdf)) {
    case SHADE_LAMBERT: launchBsdfSample<SHADE_LAMBERT>(c->ds, bsdf, n, ds_, dwi, du, dout, st); break;
    case SHADE_LEAF: launchBsdfSample<SHADE_LEAF>(c->ds, bsdf, n, ds_, dwi, du, dout, st); break;
    default: launchBsdfSample<SHADE_GENERIC>(c->ds, bsdf, n, ds_, dwi, du, dout, st); break;
    }
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 12 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "bsdf_sample");
}

MTS_API int mtscuda_sample_emitter_direct(MtsContext *c, int64_t n, const float *ref, const float *refn, const float *u, float *out) {
    HOOK_PREAMBLE(true)
    float *dr = sg.in(ref, 3 * n, st), *dn = sg.in(refn, 3 * n, st), *du = sg.in(u, 2 * n, st), *dout = sg.out<float>(16 * n);
    NEED(dr && dn && du && dout);
    k_hook_sample_emitter<<<gridFor(n), 128, 0, st>>>(c->ds, n, dr, dn, du, dout);
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 16 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "sample_emitter_direct");
}

MTS_API int mtscuda_eval_environment(MtsContext *c, int64_t n, const float *d, float *out) {
    if (!c->built || !c->ds.hasEnv) return fail(MTS_ESTATE, "no environment emitter");
    HOOK_PREAMBLE(true)
    float *dd = sg.in(d, 3 * n, st), *dout = sg.out<float>(4 * n); NEED(dd && dout);
    k_hook_eval_env<<<gridFor(n), 128, 0, st>>>(c->ds, n, dd, dout);
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 4 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "eval_environment");
}

MTS_API int mtscuda_texture_eval(MtsContext *c, int32_t tex, int64_t n, const float *uv, const float *duv, float *out) {
    if (tex < 0 || tex >= (int)c->textures.size()) return fail(MTS_EINVAL, "bad texture index");
    HOOK_PREAMBLE(true)
    float *duv_ = sg.in(uv, 2 * n, st), *dd = duv ? sg.in(duv, 4 * n, st) : nullptr, *dout = sg.out<float>(3 * n);
    NEED(duv_ && dout && (!duv || dd));
    k_hook_texture_eval<<<gridFor(n), 128, 0, st>>>(c->ds, tex, n, duv_, dd, dout);
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "texture_eval");
}

MTS_API int mtscuda_hit_partials(MtsContext *c, int64_t n, const float *sample_pos, int32_t sample_count, float *out) {
    if (sample_count <= 0) return fail(MTS_EINVAL, "sample_count must be positive");
    HOOK_PREAMBLE(true)
    if (!c->hasCam) return fail(MTS_ESTATE, "no camera set");
    c->ds.cam = c->cam;
    float *dsp = sg.in(sample_pos, 2 * n, st), *dout = sg.out<float>(8 * n); NEED(dsp && dout);
    k_hook_hit_partials<<<gridFor(n), 128, 0, st>>>(c->ds, n, dsp, 1.0f / std::sqrt((float)sample_count), dout);
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 8 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "hit_partials");
}

MTS_API int mtscuda_eval_environment_camera(MtsContext *c, int64_t n, const float *sample_pos, int32_t sample_count, float *out) {
    if (sample_count <= 0) return fail(MTS_EINVAL, "sample_count must be positive");
    HOOK_PREAMBLE(true)
    if (!c->hasCam) return fail(MTS_ESTATE, "no camera set");
    if (!c->ds.hasEnv) return fail(MTS_ESTATE, "the scene has no environment emitter");
    c->ds.cam = c->cam;
    float *dsp = sg.in(sample_pos, 2 * n, st), *dout = sg.out<float>(3 * n); NEED(dsp && dout);
    k_hook_eval_env_camera<<<gridFor(n), 128, 0, st>>>(c->ds, n, dsp, 1.0f / std::sqrt((float)sample_count), dout);
    CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, st));
    return syncCheck(c, "eval_environment_camera");
}

MTS_API int mtscuda_splat(MtsContext *c, int64_t n, const float *pos, const float *value, int32_t clear_first) {
    if (!c->hasFilm) return fail(MTS_ESTATE, "no film");
    HOOK_PREAMBLE(false)
    if (clear_first) CUDA_TRY(cudaMemsetAsync(c->ds.film, 0, sizeof(float4) * (size_t)c->film.width * c->film.height, st));
    float *dp = sg.in(pos, 2 * n, st), *dv = sg.in(value, 3 * n, st); NEED(dp && dv);
    k_hook_splat<<<gridFor(n), 128, 0, st>>>(c->ds, n, dp, dv);
    return syncCheck(c, "splat");
}

} // extern "C"
