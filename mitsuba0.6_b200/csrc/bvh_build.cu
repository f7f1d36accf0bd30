// bvh_build.cu -- GPU construction of the acceleration structure (replaces ShapeKDTree::build,
// include/mitsuba/render/gkdtree.h:959-1264, for the cudapath plugin).
//
// Level-synchronous top-down binned-SAH build, entirely on the device:
//   per level   bin      : every primitive adds its box to 3 axes x 16 centroid bins of its node
//                          (integer-ordered float atomics),
//               split    : one thread per open node sweeps the bins, picks the cheapest SAH plane
//                          (or makes a leaf / falls back to an object-median split), allocates the
//                          two children and their exact boxes,
//               partition: stable in-segment partition of the primitive index array through one
//                          global exclusive scan (cub::DeviceScan) of the "goes left" flags.
// The result is emitted as 64-byte two-wide nodes (both child boxes per node) plus the 48-byte
// primitive records in leaf order.
#include <cub/device/device_scan.cuh>
#include <vector>
#include "bvh_build.h"

namespace mts {

constexpr int NB = 16;            // SAH bins per axis
constexpr int kMaxLeaf = 4;       // primitives per leaf upper bound when the SAH prefers a leaf
constexpr float kCostTrav = 1.0f, kCostIsect = 1.2f;

struct BuildNode {
    float lo[3], hi[3];     // box
    float clo[3], chi[3];   // centroid bounds used for binning
    int begin, end;         // range in the index array
    int left, right;        // children (build-node ids), -1 for leaves
    int axis, splitBin;     // chosen plane; splitBin < 0 -> object median
    int nLeft;
    int pad;
};

struct Bin { int count; int lo[3], hi[3]; };

__device__ __forceinline__ int ordF(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float unordF(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__device__ __forceinline__ int binOf(float c, float clo, float chi) {
    float ext = chi - clo;
    if (!(ext > 0.0f)) return 0;
    int b = (int)((c - clo) * ((float)NB / ext));
    return min(NB - 1, max(0, b));
}

// ---- primitive bounds -------------------------------------------------------------------
__global__ void k_tri_bounds(const float *__restrict__ pos, const uint32_t *__restrict__ idx, uint32_t triOffset,
                             uint32_t ntris, uint32_t primOffset, float4 *__restrict__ pLo, float4 *__restrict__ pHi) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ntris) return;
    size_t gt = (size_t)triOffset + i;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int k = 0; k < 3; ++k) {
        uint32_t v = idx[3 * gt + k];
        for (int a = 0; a < 3; ++a) { float c = pos[3 * (size_t)v + a]; lo[a] = fminf(lo[a], c); hi[a] = fmaxf(hi[a], c); }
    }
    pLo[primOffset + i] = make_float4(lo[0], lo[1], lo[2], 0);
    pHi[primOffset + i] = make_float4(hi[0], hi[1], hi[2], 0);
}

__global__ void k_init(int n, uint32_t *idx, int *nodeOf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { idx[i] = i; nodeOf[i] = 0; }
}

// root bounds: block reduce + atomics into an ordered-int box
__global__ void k_root_bounds(int n, const float4 *__restrict__ pLo, const float4 *__restrict__ pHi, int *box /*12*/) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < n) {
        float4 l = pLo[i], h = pHi[i];
        lo[0] = l.x; lo[1] = l.y; lo[2] = l.z; hi[0] = h.x; hi[1] = h.y; hi[2] = h.z;
        for (int a = 0; a < 3; ++a) clo[a] = chi[a] = 0.5f * (lo[a] + hi[a]);
    }
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
            clo[a] = fminf(clo[a], __shfl_xor_sync(0xffffffffu, clo[a], o));
            chi[a] = fmaxf(chi[a], __shfl_xor_sync(0xffffffffu, chi[a], o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(box + a, ordF(lo[a])); atomicMax(box + 3 + a, ordF(hi[a]));
            atomicMin(box + 6 + a, ordF(clo[a])); atomicMax(box + 9 + a, ordF(chi[a]));
        }
    }
}

__global__ void k_set_root(BuildNode *nodes, const int *box, int n) {
    BuildNode &r = nodes[0];
    for (int a = 0; a < 3; ++a) {
        r.lo[a] = unordF(box[a]); r.hi[a] = unordF(box[3 + a]);
        r.clo[a] = unordF(box[6 + a]); r.chi[a] = unordF(box[9 + a]);
    }
    r.begin = 0; r.end = n; r.left = r.right = -1; r.axis = 0; r.splitBin = -1; r.nLeft = 0;
}

__global__ void k_clear_bins(Bin *bins, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        Bin b; b.count = 0;
        for (int a = 0; a < 3; ++a) { b.lo[a] = 0x7fffffff; b.hi[a] = (int)0x80000000; }
        bins[i] = b;
    }
}

// ---- per level ------------------------------------------------------------------------------
__global__ void k_bin(int n, const uint32_t *__restrict__ idx, const int *__restrict__ nodeOf,
                      const float4 *__restrict__ pLo, const float4 *__restrict__ pHi,
                      const BuildNode *__restrict__ nodes, int levelStart, Bin *bins) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int node = nodeOf[i];
    if (node < levelStart) return;  // finished (leaf) or not part of this level
    const BuildNode &nd = nodes[node];
    uint32_t p = idx[i];
    float4 l = pLo[p], h = pHi[p];
    float c[3] = {0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)};
    int ol[3] = {ordF(l.x), ordF(l.y), ordF(l.z)}, oh[3] = {ordF(h.x), ordF(h.y), ordF(h.z)};
    Bin *nb = bins + (size_t)(node - levelStart) * 3 * NB;
    for (int a = 0; a < 3; ++a) {
        Bin *b = nb + a * NB + binOf(c[a], nd.clo[a], nd.chi[a]);
        atomicAdd(&b->count, 1);
        for (int k = 0; k < 3; ++k) { atomicMin(&b->lo[k], ol[k]); atomicMax(&b->hi[k], oh[k]); }
    }
}

__device__ __forceinline__ float halfArea(const float lo[3], const float hi[3]) {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_split(int levelStart, int levelEnd, BuildNode *nodes, const Bin *__restrict__ bins, int *nodeCount) {
    int node = levelStart + blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= levelEnd) return;
    BuildNode nd = nodes[node];
    const int count = nd.end - nd.begin;
    const Bin *nb = bins + (size_t)(node - levelStart) * 3 * NB;
    float bestCost = INFINITY; int bestAxis = -1, bestBin = -1, bestNL = 0;
    float bLlo[3], bLhi[3], bRlo[3], bRhi[3];
    const float parentArea = fmaxf(halfArea(nd.lo, nd.hi), 1e-30f);
    for (int a = 0; a < 3; ++a) {
        // suffix boxes
        float rlo[NB][3], rhi[NB][3]; int rcnt[NB];
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}; int cnt = 0;
        for (int b = NB - 1; b >= 1; --b) {
            const Bin &bn = nb[a * NB + b];
            if (bn.count) for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], unordF(bn.lo[k])); hi[k] = fmaxf(hi[k], unordF(bn.hi[k])); }
            cnt += bn.count;
            for (int k = 0; k < 3; ++k) { rlo[b][k] = lo[k]; rhi[b][k] = hi[k]; }
            rcnt[b] = cnt;
        }
        float llo[3] = {INFINITY, INFINITY, INFINITY}, lhi[3] = {-INFINITY, -INFINITY, -INFINITY}; int lcnt = 0;
        for (int s = 1; s < NB; ++s) {  // plane between bin s-1 and s
            const Bin &bn = nb[a * NB + s - 1];
            if (bn.count) for (int k = 0; k < 3; ++k) { llo[k] = fminf(llo[k], unordF(bn.lo[k])); lhi[k] = fmaxf(lhi[k], unordF(bn.hi[k])); }
            lcnt += bn.count;
            if (lcnt == 0 || rcnt[s] == 0) continue;
            float cost = halfArea(llo, lhi) * lcnt + halfArea(rlo[s], rhi[s]) * rcnt[s];
            if (cost < bestCost) {
                bestCost = cost; bestAxis = a; bestBin = s; bestNL = lcnt;
                for (int k = 0; k < 3; ++k) { bLlo[k] = llo[k]; bLhi[k] = lhi[k]; bRlo[k] = rlo[s][k]; bRhi[k] = rhi[s][k]; }
            }
        }
    }
    bool leaf = false, median = false;
    if (bestAxis < 0) {
        // all centroids fall into one bin on every axis
        if (count <= kMaxLeaf) leaf = true; else median = true;
    } else {
        float splitCost = kCostTrav + kCostIsect * bestCost / parentArea;
        float leafCost = kCostIsect * count;
        if (count <= kMaxLeaf && leafCost <= splitCost) leaf = true;
    }
    if (count <= 1) { leaf = true; median = false; }
    if (leaf) { nd.left = nd.right = -1; nodes[node] = nd; return; }
    int c0 = atomicAdd(nodeCount, 2);
    BuildNode L, R;
    L.left = L.right = R.left = R.right = -1; L.axis = R.axis = 0; L.splitBin = R.splitBin = -1; L.nLeft = R.nLeft = 0; L.pad = R.pad = 0;
    if (median) {
        // object median on the index order: children inherit the parent box (conservative) --
        // happens only for coincident centroids
        nd.axis = 0; nd.splitBin = -1; nd.nLeft = count / 2;
        for (int k = 0; k < 3; ++k) {
            L.lo[k] = R.lo[k] = nd.lo[k]; L.hi[k] = R.hi[k] = nd.hi[k];
            L.clo[k] = R.clo[k] = nd.clo[k]; L.chi[k] = R.chi[k] = nd.chi[k];
        }
    } else {
        nd.axis = bestAxis; nd.splitBin = bestBin; nd.nLeft = bestNL;
        float plane = nd.clo[bestAxis] + (nd.chi[bestAxis] - nd.clo[bestAxis]) * ((float)bestBin / NB);
        for (int k = 0; k < 3; ++k) {
            L.lo[k] = bLlo[k]; L.hi[k] = bLhi[k]; R.lo[k] = bRlo[k]; R.hi[k] = bRhi[k];
            // centroids lie inside the child's box and inside the parent's centroid bounds
            L.clo[k] = fmaxf(nd.clo[k], bLlo[k]); L.chi[k] = fminf(nd.chi[k], bLhi[k]);
            R.clo[k] = fmaxf(nd.clo[k], bRlo[k]); R.chi[k] = fminf(nd.chi[k], bRhi[k]);
        }
        L.chi[bestAxis] = fminf(L.chi[bestAxis], plane);
        R.clo[bestAxis] = fmaxf(R.clo[bestAxis], plane);
    }
    L.begin = nd.begin; L.end = nd.begin + nd.nLeft; R.begin = L.end; R.end = nd.end;
    nd.left = c0; nd.right = c0 + 1;
    nodes[c0] = L; nodes[c0 + 1] = R;
    nodes[node] = nd;
}

__global__ void k_flags(int n, const uint32_t *__restrict__ idx, const int *__restrict__ nodeOf,
                        const float4 *__restrict__ pLo, const float4 *__restrict__ pHi,
                        const BuildNode *__restrict__ nodes, int levelStart, int *flags) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int node = nodeOf[i];
    int f = 0;
    if (node >= levelStart) {
        const BuildNode &nd = nodes[node];
        if (nd.left >= 0) {
            if (nd.splitBin < 0) f = (i - nd.begin) < nd.nLeft;
            else {
                uint32_t p = idx[i];
                float4 l = pLo[p], h = pHi[p];
                float c = nd.axis == 0 ? 0.5f * (l.x + h.x) : (nd.axis == 1 ? 0.5f * (l.y + h.y) : 0.5f * (l.z + h.z));
                f = binOf(c, nd.clo[nd.axis], nd.chi[nd.axis]) < nd.splitBin;
            }
        }
    }
    flags[i] = f;
}

__global__ void k_scatter(int n, const uint32_t *__restrict__ idxIn, const int *__restrict__ nodeOfIn,
                          const int *__restrict__ flags, const int *__restrict__ scan,
                          const BuildNode *__restrict__ nodes, int levelStart, uint32_t *idxOut, int *nodeOfOut) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int node = nodeOfIn[i];
    if (node < levelStart) { idxOut[i] = idxIn[i]; nodeOfOut[i] = -1; return; }
    const BuildNode &nd = nodes[node];
    if (nd.left < 0) { idxOut[i] = idxIn[i]; nodeOfOut[i] = -1; return; }  // became a leaf
    int leftBefore = scan[i] - scan[nd.begin];
    int dst;
    if (flags[i]) dst = nd.begin + leftBefore;
    else dst = nd.begin + nd.nLeft + ((i - nd.begin) - leftBefore);
    idxOut[dst] = idxIn[i];
    nodeOfOut[dst] = flags[i] ? nd.left : nd.right;
}

// ---- emission -----------------------------------------------------------------------------
__global__ void k_inner_flags(int nNodes, const BuildNode *__restrict__ nodes, int *flags) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nNodes) flags[i] = nodes[i].left >= 0;
}

__global__ void k_emit_nodes(int nNodes, const BuildNode *__restrict__ nodes, const int *__restrict__ innerIndex, BvhNode *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nNodes) return;
    const BuildNode &nd = nodes[i];
    if (nd.left < 0) return;
    const BuildNode &L = nodes[nd.left], &R = nodes[nd.right];
    BvhNode o;
    o.a = make_float4(L.lo[0], L.lo[1], L.lo[2], L.hi[0]);
    o.b = make_float4(L.hi[1], L.hi[2], R.lo[0], R.lo[1]);
    o.c = make_float4(R.lo[2], R.hi[0], R.hi[1], R.hi[2]);
    o.link.x = L.left >= 0 ? innerIndex[nd.left] : L.begin;
    o.link.z = L.left >= 0 ? 0 : (L.end - L.begin);
    o.link.y = R.left >= 0 ? innerIndex[nd.right] : R.begin;
    o.link.w = R.left >= 0 ? 0 : (R.end - R.begin);
    out[innerIndex[i]] = o;
}

// root is a leaf (tiny scenes): one node whose first child holds everything
__global__ void k_emit_single(const BuildNode *nodes, BvhNode *out) {
    const BuildNode &r = nodes[0];
    BvhNode o;
    o.a = make_float4(r.lo[0], r.lo[1], r.lo[2], r.hi[0]);
    o.b = make_float4(r.hi[1], r.hi[2], 0, 0);
    o.c = make_float4(0, 0, 0, 0);
    o.link = make_int4(r.begin, -1, r.end - r.begin, 0);
    out[0] = o;
}

__global__ void k_emit_prims(int n, const uint32_t *__restrict__ order, const PrimSource *__restrict__ src,
                             const float *__restrict__ pos, const uint32_t *__restrict__ idx, PrimRecord *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    PrimSource s = src[order[i]];
    PrimRecord r;
    if (s.type == 0) {
        size_t gt = (size_t)s.triOffset + s.prim;
        uint32_t i0 = idx[3 * gt], i1 = idx[3 * gt + 1], i2 = idx[3 * gt + 2];
        float3 v0 = make_float3(pos[3 * (size_t)i0], pos[3 * (size_t)i0 + 1], pos[3 * (size_t)i0 + 2]);
        float3 v1 = make_float3(pos[3 * (size_t)i1], pos[3 * (size_t)i1 + 1], pos[3 * (size_t)i1 + 2]);
        float3 v2 = make_float3(pos[3 * (size_t)i2], pos[3 * (size_t)i2 + 1], pos[3 * (size_t)i2 + 2]);
        r.p0 = make_float4(v0.x, v0.y, v0.z, __int_as_float(s.shape));
        r.p1 = make_float4(v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, __int_as_float((int)s.prim));
        r.p2 = make_float4(v2.x - v0.x, v2.y - v0.y, v2.z - v0.z, __int_as_float(0));
    } else {
        r.p0 = make_float4(0, 0, 0, __int_as_float(s.shape));
        r.p1 = make_float4(0, 0, 0, __int_as_float(0));
        r.p2 = make_float4(0, 0, 0, __int_as_float(s.type));
    }
    out[i] = r;
}

__global__ void k_fill_tri_sources(uint32_t ntris, uint32_t primOffset, int shape, uint32_t triOffset, PrimSource *src) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ntris) return;
    PrimSource s; s.type = 0; s.shape = shape; s.prim = i; s.triOffset = triOffset;
    src[primOffset + i] = s;
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = std::string(#x) + ": " + cudaGetErrorString(e_); goto fail; } } while (0)

static inline int blocks(size_t n, int bs = 256) { return (int)((n + bs - 1) / bs); }

bool buildBvh(const BvhBuildInput &in, BvhBuildOutput &out, cudaStream_t st, std::string &err) {
    const int n = in.numPrims;
    out = BvhBuildOutput();
    float4 *pLo = nullptr, *pHi = nullptr; uint32_t *idxA = nullptr, *idxB = nullptr; int *nodeA = nullptr, *nodeB = nullptr;
    BuildNode *nodes = nullptr; int *flags = nullptr, *scan = nullptr, *box = nullptr, *nodeCount = nullptr;
    Bin *bins = nullptr; size_t binCap = 0; void *cubTmp = nullptr; size_t cubBytes = 0;
    PrimSource *src = nullptr; int *innerIdx = nullptr;
    int launches = 0;
    if (n == 0) { out.nodes = nullptr; out.prims = nullptr; out.numNodes = 0; return true; }
    {
    const size_t maxNodes = 2 * (size_t)n + 2;
    CK(cudaMalloc(&pLo, sizeof(float4) * n)); CK(cudaMalloc(&pHi, sizeof(float4) * n));
    CK(cudaMalloc(&idxA, 4 * (size_t)n)); CK(cudaMalloc(&idxB, 4 * (size_t)n));
    CK(cudaMalloc(&nodeA, 4 * (size_t)n)); CK(cudaMalloc(&nodeB, 4 * (size_t)n));
    CK(cudaMalloc(&nodes, sizeof(BuildNode) * maxNodes));
    CK(cudaMalloc(&flags, 4 * maxNodes)); CK(cudaMalloc(&scan, 4 * maxNodes));
    CK(cudaMalloc(&box, 4 * 12)); CK(cudaMalloc(&nodeCount, 4));
    CK(cudaMalloc(&src, sizeof(PrimSource) * n));
    CK(cub::DeviceScan::ExclusiveSum(nullptr, cubBytes, flags, scan, (int)maxNodes, st));
    CK(cudaMalloc(&cubTmp, cubBytes));

    // primitive sources + bounds
    for (const auto &m : in.meshes) {
        if (!m.ntris) continue;
        k_tri_bounds<<<blocks(m.ntris), 256, 0, st>>>(in.positions, in.indices, m.triOffset, m.ntris, m.primOffset, pLo, pHi);
        k_fill_tri_sources<<<blocks(m.ntris), 256, 0, st>>>(m.ntris, m.primOffset, m.shape, m.triOffset, src);
        launches += 2;
    }
    for (const auto &a : in.analytic) {
        float4 lo = make_float4(a.lo[0], a.lo[1], a.lo[2], 0), hi = make_float4(a.hi[0], a.hi[1], a.hi[2], 0);
        PrimSource s; s.type = a.type; s.shape = a.shape; s.prim = 0; s.triOffset = 0;
        CK(cudaMemcpyAsync(pLo + a.primOffset, &lo, sizeof(lo), cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(pHi + a.primOffset, &hi, sizeof(hi), cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(src + a.primOffset, &s, sizeof(s), cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
    }
    {
        int hbox[12];
        for (int a = 0; a < 3; ++a) { hbox[a] = hbox[6 + a] = 0x7fffffff; hbox[3 + a] = hbox[9 + a] = (int)0x80000000; }
        CK(cudaMemcpyAsync(box, hbox, sizeof(hbox), cudaMemcpyHostToDevice, st));
        int one = 1;
        CK(cudaMemcpyAsync(nodeCount, &one, 4, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
    }
    k_init<<<blocks(n), 256, 0, st>>>(n, idxA, nodeA);
    k_root_bounds<<<blocks(n), 256, 0, st>>>(n, pLo, pHi, box);
    k_set_root<<<1, 1, 0, st>>>(nodes, box, n);
    launches += 3;

    int levelStart = 0, levelEnd = 1;
    while (levelStart < levelEnd) {
        size_t nb = (size_t)(levelEnd - levelStart) * 3 * NB;
        if (nb > binCap) {
            if (bins) CK(cudaFree(bins));
            binCap = nb + nb / 2;
            CK(cudaMalloc(&bins, sizeof(Bin) * binCap));
        }
        k_clear_bins<<<blocks(nb), 256, 0, st>>>(bins, nb);
        k_bin<<<blocks(n), 256, 0, st>>>(n, idxA, nodeA, pLo, pHi, nodes, levelStart, bins);
        k_split<<<blocks(levelEnd - levelStart, 64), 64, 0, st>>>(levelStart, levelEnd, nodes, bins, nodeCount);
        k_flags<<<blocks(n), 256, 0, st>>>(n, idxA, nodeA, pLo, pHi, nodes, levelStart, flags);
        CK(cub::DeviceScan::ExclusiveSum(cubTmp, cubBytes, flags, scan, n, st));
        k_scatter<<<blocks(n), 256, 0, st>>>(n, idxA, nodeA, flags, scan, nodes, levelStart, idxB, nodeB);
        launches += 6;
        std::swap(idxA, idxB); std::swap(nodeA, nodeB);
        int cnt = 0;
        CK(cudaMemcpyAsync(&cnt, nodeCount, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        levelStart = levelEnd; levelEnd = cnt;
        if ((size_t)cnt > maxNodes) { err = "BVH build: node overflow"; goto fail; }
    }
    const int nBuild = levelEnd;
    // compact inner nodes -> traversal nodes
    k_inner_flags<<<blocks(nBuild), 256, 0, st>>>(nBuild, nodes, flags);
    CK(cub::DeviceScan::ExclusiveSum(cubTmp, cubBytes, flags, scan, nBuild, st));
    int lastFlag = 0, lastScan = 0;
    CK(cudaMemcpyAsync(&lastFlag, flags + nBuild - 1, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&lastScan, scan + nBuild - 1, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    int nInner = lastFlag + lastScan;
    out.numNodes = nInner > 0 ? nInner : 1;
    CK(cudaMalloc(&out.nodes, sizeof(BvhNode) * out.numNodes));
    CK(cudaMalloc(&out.prims, sizeof(PrimRecord) * n));
    if (nInner > 0) k_emit_nodes<<<blocks(nBuild), 256, 0, st>>>(nBuild, nodes, scan, out.nodes);
    else k_emit_single<<<1, 1, 0, st>>>(nodes, out.nodes);
    k_emit_prims<<<blocks(n), 256, 0, st>>>(n, idxA, src, in.positions, in.indices, out.prims);
    launches += 4;
    BuildNode root;
    CK(cudaMemcpyAsync(&root, nodes, sizeof(root), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    for (int a = 0; a < 3; ++a) { out.lo[a] = root.lo[a]; out.hi[a] = root.hi[a]; }
    out.launches = launches;
    out.buildNodes = nBuild;
    }
    cudaFree(pLo); cudaFree(pHi); cudaFree(idxA); cudaFree(idxB); cudaFree(nodeA); cudaFree(nodeB); cudaFree(nodes);
    cudaFree(flags); cudaFree(scan); cudaFree(box); cudaFree(nodeCount); cudaFree(bins); cudaFree(cubTmp); cudaFree(src); cudaFree(innerIdx);
    return true;
fail:
    cudaFree(pLo); cudaFree(pHi); cudaFree(idxA); cudaFree(idxB); cudaFree(nodeA); cudaFree(nodeB); cudaFree(nodes);
    cudaFree(flags); cudaFree(scan); cudaFree(box); cudaFree(nodeCount); cudaFree(bins); cudaFree(cubTmp); cudaFree(src); cudaFree(innerIdx);
    if (out.nodes) cudaFree(out.nodes);
    if (out.prims) cudaFree(out.prims);
    out = BvhBuildOutput();
    return false;
}

} // namespace mts
