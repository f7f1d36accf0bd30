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
#include "devmem.h"
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>

namespace mts {

constexpr int NB = 16;            // SAH bins per axis
constexpr int kMaxLeaf = 3;       // primitives per leaf (a wide-node leaf child encodes 1..3 in unary)
#ifndef BVH_COST_ISECT
#define BVH_COST_ISECT 1.2f
#endif
constexpr float kCostTrav = 1.0f, kCostIsect = BVH_COST_ISECT;
constexpr int kSmall = 32;        // nodes of at most one primitive per lane are finished by k_small_subtrees

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

// root bounds: grid-stride accumulation, warp shuffle + shared-memory block reduce, 12 atomics per block into an
// ordered-int box
__global__ void __launch_bounds__(256) k_root_bounds(int n, const float4 *__restrict__ pLo, const float4 *__restrict__ pHi, int *box /*12*/) {
    __shared__ float sred[8][12];
    float v[12];   // lo, hi, clo, chi
    for (int a = 0; a < 3; ++a) { v[a] = INFINITY; v[3 + a] = -INFINITY; v[6 + a] = INFINITY; v[9 + a] = -INFINITY; }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 l = pLo[i], h = pHi[i];
        const float lo[3] = {l.x, l.y, l.z}, hi[3] = {h.x, h.y, h.z};
        for (int a = 0; a < 3; ++a) {
            const float c = 0.5f * (lo[a] + hi[a]);
            v[a] = fminf(v[a], lo[a]); v[3 + a] = fmaxf(v[3 + a], hi[a]);
            v[6 + a] = fminf(v[6 + a], c); v[9 + a] = fmaxf(v[9 + a], c);
        }
    }
    for (int k = 0; k < 12; ++k) {
        const bool isMin = (k / 3) % 2 == 0;
        for (int o = 16; o > 0; o >>= 1) {
            const float t = __shfl_xor_sync(0xffffffffu, v[k], o);
            v[k] = isMin ? fminf(v[k], t) : fmaxf(v[k], t);
        }
        if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5][k] = v[k];
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        const int k = threadIdx.x; const bool isMin = (k / 3) % 2 == 0;
        float r = sred[0][k];
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = isMin ? fminf(r, sred[w][k]) : fmaxf(r, sred[w][k]);
        if (isMin) atomicMin(box + k, ordF(r)); else atomicMax(box + k, ordF(r));
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
// Every block bins a contiguous chunk of the (node-partitioned) primitive order.  A node's primitives are
// contiguous, so when the first and last primitive of the chunk belong to the same open node the whole chunk does:
// it is then binned in shared memory and flushed with one global atomic per non-empty bin entry, which removes the
// same-address contention of the top levels (all 10^7 primitives in a handful of nodes).  Mixed chunks (deep
// levels, many small nodes) use global atomics directly.
constexpr int kBinChunk = 2048;
__global__ void __launch_bounds__(256) k_bin(int n, const uint32_t *__restrict__ idx, const int *__restrict__ nodeOf,
                      const float4 *__restrict__ pLo, const float4 *__restrict__ pHi,
                      const BuildNode *__restrict__ nodes, int levelStart, Bin *bins) {
    __shared__ Bin sbin[3 * NB];
    const int c0 = blockIdx.x * kBinChunk, c1 = min(n, c0 + kBinChunk);
    if (c0 >= n) return;
    const int nodeFirst = nodeOf[c0], nodeLast = nodeOf[c1 - 1];
    const bool uniform = nodeFirst == nodeLast && nodeFirst >= levelStart;   // finished leaves all read -1
    if (uniform) {
        for (int i = threadIdx.x; i < 3 * NB; i += blockDim.x) {
            Bin b; b.count = 0;
            for (int a = 0; a < 3; ++a) { b.lo[a] = 0x7fffffff; b.hi[a] = (int)0x80000000; }
            sbin[i] = b;
        }
        __syncthreads();
    }
    // The primitive order is node-partitioned and spatially coherent, so the 32 primitives of a warp fall into a
    // handful of (node, bin) pairs: each distinct pair is reduced across the warp with redux.sync (identity values for
    // the lanes outside the group) and its leader issues the seven atomics -- instead of seven contended atomics per lane.
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    for (int base = c0; base < c1; base += blockDim.x) {
        const int i = base + threadIdx.x;
        bool valid = i < c1;
        int node = -1;
        if (valid) { node = uniform ? nodeFirst : nodeOf[i]; valid = node >= levelStart; }  // else finished (leaf) / not part of this level
        float c[3] = {0, 0, 0}; int ol[3] = {0, 0, 0}, oh[3] = {0, 0, 0};
        float clo[3] = {0, 0, 0}, chi[3] = {0, 0, 0};
        if (valid) {
            const BuildNode &nd = nodes[node];
            const uint32_t p = idx[i];
            const float4 l = pLo[p], h = pHi[p];
            c[0] = 0.5f * (l.x + h.x); c[1] = 0.5f * (l.y + h.y); c[2] = 0.5f * (l.z + h.z);
            ol[0] = ordF(l.x); ol[1] = ordF(l.y); ol[2] = ordF(l.z); oh[0] = ordF(h.x); oh[1] = ordF(h.y); oh[2] = ordF(h.z);
            for (int a = 0; a < 3; ++a) { clo[a] = nd.clo[a]; chi[a] = nd.chi[a]; }
        }
        const long long nodeBase = uniform ? 0 : (long long)(node - levelStart) * 3 * NB;
        for (int a = 0; a < 3; ++a) {
            const long long key = valid ? nodeBase + a * NB + binOf(c[a], clo[a], chi[a]) : -1;
            unsigned todo = __ballot_sync(full, valid);
            while (todo) {
                const int leader = __ffs(todo) - 1;
                const long long k = __shfl_sync(full, key, leader);
                const bool mine = valid && key == k;
                const unsigned grp = __ballot_sync(full, mine);
                const int l0 = __reduce_min_sync(full, mine ? ol[0] : 0x7fffffff), l1 = __reduce_min_sync(full, mine ? ol[1] : 0x7fffffff),
                          l2 = __reduce_min_sync(full, mine ? ol[2] : 0x7fffffff);
                const int h0 = __reduce_max_sync(full, mine ? oh[0] : (int)0x80000000), h1 = __reduce_max_sync(full, mine ? oh[1] : (int)0x80000000),
                          h2 = __reduce_max_sync(full, mine ? oh[2] : (int)0x80000000);
                if (lane == leader) {
                    Bin *b = (uniform ? sbin : bins) + k;
                    atomicAdd(&b->count, __popc(grp));
                    atomicMin(&b->lo[0], l0); atomicMin(&b->lo[1], l1); atomicMin(&b->lo[2], l2);
                    atomicMax(&b->hi[0], h0); atomicMax(&b->hi[1], h1); atomicMax(&b->hi[2], h2);
                }
                todo &= ~grp;
            }
        }
    }
    if (uniform) {
        __syncthreads();
        Bin *nb = bins + (size_t)(nodeFirst - levelStart) * 3 * NB;
        for (int i = threadIdx.x; i < 3 * NB; i += blockDim.x) {
            const Bin &b = sbin[i];
            if (!b.count) continue;
            atomicAdd(&nb[i].count, b.count);
            for (int k = 0; k < 3; ++k) { atomicMin(&nb[i].lo[k], b.lo[k]); atomicMax(&nb[i].hi[k], b.hi[k]); }
        }
    }
}

__device__ __forceinline__ float halfArea(const float lo[3], const float hi[3]) {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_split(int levelStart, int levelEnd, BuildNode *nodes, const Bin *__restrict__ bins, int *nodeCount,
                        int *smallList, int *smallCount) {
    int node = levelStart + blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= levelEnd) return;
    BuildNode nd = nodes[node];
    const int count = nd.end - nd.begin;
    if (count <= kSmall) {
        // finished for the level-synchronous phase (left = right = -1 since creation); k_small_subtrees builds the
        // whole subtree of this node with one warp once the large nodes are done
        if (count > 1) smallList[atomicAdd(smallCount, 1)] = node;
        return;
    }
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

// ---- small subtrees: one warp finishes a node of <= 32 primitives --------------------------------------
// The level-synchronous passes above touch all n primitives per level, which is wasteful once the open nodes hold
// a few dozen primitives each (two thirds of the levels of a 10^7-triangle tree).  Here every such node is handed
// to one warp: the primitives live in shared memory (one per lane), splits are chosen by an exact SAH sweep over the
// three axes (every primitive's centroid is a candidate plane; ties broken by position, so a split always exists),
// the partition is a ballot/popc prefix, and the warp recurses over an explicit stack until only leaves remain.
__global__ void __launch_bounds__(128) k_small_subtrees(const int *__restrict__ smallList, int nSmall, BuildNode *nodes, uint32_t *idx,
                                                        const float4 *__restrict__ pLo, const float4 *__restrict__ pHi, int *nodeCount) {
    __shared__ float4 sLo[4][32], sHi[4][32];   // .w of sLo = primitive id bits
    __shared__ int3 sStack[4][32];
    const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int w = blockIdx.x * 4 + wib;
    if (w >= nSmall) return;
    const unsigned full = 0xffffffffu;
    const int root = smallList[w];
    const int b0 = nodes[root].begin, m = nodes[root].end - b0;
    if (lane < m) {
        uint32_t p = idx[b0 + lane];
        float4 l = pLo[p], h = pHi[p];
        l.w = __uint_as_float(p);
        sLo[wib][lane] = l; sHi[wib][lane] = h;
    }
    int sp = 0;
    if (lane == 0) sStack[wib][0] = make_int3(root, 0, m);
    sp = 1;
    __syncwarp();
    while (sp > 0) {
        const int3 top = sStack[wib][--sp];
        const int nid = top.x, b = top.y, e = top.z, cnt = e - b;
        if (cnt <= 1) continue;
        const bool in = lane >= b && lane < e;
        const float4 myLo = sLo[wib][lane], myHi = sHi[wib][lane];
        const float myC[3] = {0.5f * (myLo.x + myHi.x), 0.5f * (myLo.y + myHi.y), 0.5f * (myLo.z + myHi.z)};
        float bestCost = INFINITY; int bestAxis = 0;
        float bL[6], bR[6]; int bNL = 0;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            float llo[3] = {INFINITY, INFINITY, INFINITY}, lhi[3] = {-INFINITY, -INFINITY, -INFINITY};
            float rlo[3] = {INFINITY, INFINITY, INFINITY}, rhi[3] = {-INFINITY, -INFINITY, -INFINITY};
            int nL = 0;
            for (int j = b; j < e; ++j) {
                const float4 l = sLo[wib][j], h = sHi[wib][j];   // broadcast reads
                const float cj = a == 0 ? 0.5f * (l.x + h.x) : (a == 1 ? 0.5f * (l.y + h.y) : 0.5f * (l.z + h.z));
                const bool left = cj < myC[a] || (cj == myC[a] && j <= lane);
                if (left) {
                    llo[0] = fminf(llo[0], l.x); llo[1] = fminf(llo[1], l.y); llo[2] = fminf(llo[2], l.z);
                    lhi[0] = fmaxf(lhi[0], h.x); lhi[1] = fmaxf(lhi[1], h.y); lhi[2] = fmaxf(lhi[2], h.z);
                    ++nL;
                } else {
                    rlo[0] = fminf(rlo[0], l.x); rlo[1] = fminf(rlo[1], l.y); rlo[2] = fminf(rlo[2], l.z);
                    rhi[0] = fmaxf(rhi[0], h.x); rhi[1] = fmaxf(rhi[1], h.y); rhi[2] = fmaxf(rhi[2], h.z);
                }
            }
            if (in && nL < cnt) {
                const float cost = halfArea(llo, lhi) * nL + halfArea(rlo, rhi) * (cnt - nL);
                if (cost < bestCost) {
                    bestCost = cost; bestAxis = a; bNL = nL;
                    for (int k = 0; k < 3; ++k) { bL[k] = llo[k]; bL[3 + k] = lhi[k]; bR[k] = rlo[k]; bR[3 + k] = rhi[k]; }
                }
            }
        }
        // warp argmin (ties -> lowest lane)
        float c = bestCost; int who = lane;
        for (int o = 16; o > 0; o >>= 1) {
            const float oc = __shfl_xor_sync(full, c, o); const int ow = __shfl_xor_sync(full, who, o);
            if (oc < c || (oc == c && ow < who)) { c = oc; who = ow; }
        }
        // parent box for the leaf criterion of k_split
        float plo[3], phi[3];
        for (int k = 0; k < 3; ++k) { plo[k] = __ldcg(&nodes[nid].lo[k]); phi[k] = __ldcg(&nodes[nid].hi[k]); }   // written by lane 0 one level up
        const float parentArea = fmaxf(halfArea(plo, phi), 1e-30f);
        const float splitCost = kCostTrav + kCostIsect * c / parentArea, leafCost = kCostIsect * cnt;
        if (!(c < INFINITY) || (cnt <= kMaxLeaf && leafCost <= splitCost)) {
            if (cnt <= kMaxLeaf) continue;   // leaf
            // no finite-cost split (NaN/inf boxes) on a node that is too big for a leaf: halve by position
            who = b + cnt / 2 - 1; bestAxis = -1;
        }
        const int axis = __shfl_sync(full, bestAxis, who & 31);
        int nL; bool goLeft;
        if (axis < 0) { nL = cnt / 2; goLeft = lane < b + nL; for (int k = 0; k < 3; ++k) { bL[k] = bR[k] = plo[k]; bL[3 + k] = bR[3 + k] = phi[k]; } }
        else {
            const float cs = __shfl_sync(full, axis == 0 ? myC[0] : (axis == 1 ? myC[1] : myC[2]), who);
            const float mc = axis == 0 ? myC[0] : (axis == 1 ? myC[1] : myC[2]);
            goLeft = mc < cs || (mc == cs && lane <= who);
            nL = __shfl_sync(full, bNL, who);
            for (int k = 0; k < 6; ++k) { bL[k] = __shfl_sync(full, bL[k], who); bR[k] = __shfl_sync(full, bR[k], who); }
        }
        const unsigned inMask = __ballot_sync(full, in);
        const unsigned leftMask = __ballot_sync(full, in && goLeft);
        const unsigned rightMask = inMask & ~leftMask;
        const unsigned below = (1u << lane) - 1u;
        const int dst = goLeft ? b + __popc(leftMask & below) : b + nL + __popc(rightMask & below);
        __syncwarp();
        if (in) { sLo[wib][dst] = myLo; sHi[wib][dst] = myHi; }
        int c0 = 0;
        if (lane == 0) {
            c0 = atomicAdd(nodeCount, 2);
            BuildNode L, R;
            L.left = L.right = R.left = R.right = -1; L.axis = R.axis = 0; L.splitBin = R.splitBin = -1; L.nLeft = R.nLeft = 0; L.pad = R.pad = 0;
            for (int k = 0; k < 3; ++k) {
                L.lo[k] = L.clo[k] = bL[k]; L.hi[k] = L.chi[k] = bL[3 + k];
                R.lo[k] = R.clo[k] = bR[k]; R.hi[k] = R.chi[k] = bR[3 + k];
            }
            L.begin = b0 + b; L.end = b0 + b + nL; R.begin = L.end; R.end = b0 + e;
            nodes[c0] = L; nodes[c0 + 1] = R;
            nodes[nid].left = c0; nodes[nid].right = c0 + 1; nodes[nid].axis = max(axis, 0); nodes[nid].nLeft = nL;
            sStack[wib][sp] = make_int3(c0 + 1, b + nL, e);
            sStack[wib][sp + 1] = make_int3(c0, b, b + nL);
        }
        sp += 2;
        __syncwarp();
    }
    if (lane < m) idx[b0 + lane] = __float_as_uint(sLo[wib][lane].w);
}

// ---- collapse: binary tree -> eight-wide compressed nodes ----------------------------------------
// One thread per wide node, one launch per level of the wide tree.  Wide node k is built from the binary
// subtree rooted at srcNode[k]: its two children are expanded greedily (largest surface area first) until
// eight are open or only leaves remain, the children are assigned to octant slots, their boxes are quantised
// to the 8-bit grid of the node, inner children are appended to the node array (contiguous, slot order) and the
// primitives of leaf children are appended to the primitive array (contiguous, slot order).
__device__ __forceinline__ PrimRecord makePrimRecord(const PrimSource &s, const float *__restrict__ pos, const uint32_t *__restrict__ idx) {
    PrimRecord r;
    if (s.type == 0) {
        size_t gt = (size_t)s.triOffset + s.prim;
        uint32_t i0 = idx[3 * gt], i1 = idx[3 * gt + 1], i2 = idx[3 * gt + 2];
        float3 v0 = make_float3(pos[3 * (size_t)i0], pos[3 * (size_t)i0 + 1], pos[3 * (size_t)i0 + 2]);
        float3 v1 = make_float3(pos[3 * (size_t)i1], pos[3 * (size_t)i1 + 1], pos[3 * (size_t)i1 + 2]);
        float3 v2 = make_float3(pos[3 * (size_t)i2], pos[3 * (size_t)i2 + 1], pos[3 * (size_t)i2 + 2]);
        r.p0 = make_float4(v0.x, v0.y, v0.z, __int_as_float(s.shape));
        // the three vertices themselves (not v0 + edges): neighbouring triangles then hold bit-identical copies of their
        // shared vertices, which is what makes the edge tests of primIntersect watertight
#if TRI_TEST_WOOP
        r.p1 = make_float4(v1.x, v1.y, v1.z, __int_as_float((int)s.prim));
        r.p2 = make_float4(v2.x, v2.y, v2.z, __int_as_float(0));
#else
        r.p1 = make_float4(v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, __int_as_float((int)s.prim));
        r.p2 = make_float4(v2.x - v0.x, v2.y - v0.y, v2.z - v0.z, __int_as_float(0));
#endif
    } else {
        r.p0 = make_float4(0, 0, 0, __int_as_float(s.shape));
        r.p1 = make_float4(0, 0, 0, __int_as_float(0));
        r.p2 = make_float4(0, 0, 0, __int_as_float(s.type));
    }
    return r;
}

// smallest biased exponent e such that the 8-bit grid origin + q * 2^(e-127), q in 0..255, covers [lo, hi]
__device__ __forceinline__ int gridExponent(float lo, float hi) {
    float ext = fmaxf(hi - lo, 1e-30f);
    int e;
    frexpf(ext / 255.0f, &e);            // ext/255 = m * 2^e, m in [0.5, 1)  ->  2^e >= ext/255
    e = max(-126, min(127, e));
    while (e < 127 && lo + 255.0f * __int_as_float((e + 127) << 23) < hi) ++e;   // guard fp rounding
    return e + 127;
}

__global__ void k_collapse(int levelStart, int levelEnd, const BuildNode *__restrict__ bn, const uint32_t *__restrict__ order,
                           int *srcNode, WideNode *out, int *nodeCounter, int *primCounter,
                           const PrimSource *__restrict__ src, const float *__restrict__ pos, const uint32_t *__restrict__ idx,
                           PrimRecord *prims) {
    int k = levelStart + blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= levelEnd) return;
    const int root = srcNode[k];
    const BuildNode N = bn[root];
    int child[8]; int n = 0;
    if (N.left < 0) { child[0] = root; n = 1; }   // the whole scene is one leaf
    else {
        child[0] = N.left; child[1] = N.right; n = 2;
        while (n < 8) {
            int best = -1; float bestArea = -1.0f;
            for (int i = 0; i < n; ++i) {
                const BuildNode &c = bn[child[i]];
                if (c.left >= 0) { float a = halfArea(c.lo, c.hi); if (a > bestArea) { bestArea = a; best = i; } }
            }
            if (best < 0) break;
            int c = child[best];
            child[best] = bn[c].left; child[n++] = bn[c].right;
        }
    }
    // octant slot assignment: greedily take the (child, slot) pair whose centroid offset projects furthest onto the
    // slot's diagonal direction
    float cx[8], cy[8], cz[8];
    const float ncx = 0.5f * (N.lo[0] + N.hi[0]), ncy = 0.5f * (N.lo[1] + N.hi[1]), ncz = 0.5f * (N.lo[2] + N.hi[2]);
    for (int i = 0; i < n; ++i) {
        const BuildNode &c = bn[child[i]];
        cx[i] = 0.5f * (c.lo[0] + c.hi[0]) - ncx; cy[i] = 0.5f * (c.lo[1] + c.hi[1]) - ncy; cz[i] = 0.5f * (c.lo[2] + c.hi[2]) - ncz;
    }
    int slotOf[8]; int childIn[8];
    for (int i = 0; i < 8; ++i) { slotOf[i] = -1; childIn[i] = -1; }
    for (int it = 0; it < n; ++it) {
        float bestC = -INFINITY; int bi = -1, bs = -1;
        for (int i = 0; i < n; ++i) {
            if (slotOf[i] >= 0) continue;
            for (int sl = 0; sl < 8; ++sl) {
                if (childIn[sl] >= 0) continue;
                float c = ((sl & 1) ? cx[i] : -cx[i]) + ((sl & 2) ? cy[i] : -cy[i]) + ((sl & 4) ? cz[i] : -cz[i]);
                if (c > bestC) { bestC = c; bi = i; bs = sl; }
            }
        }
        slotOf[bi] = bs; childIn[bs] = bi;
    }
    // grid
    const int ex = gridExponent(N.lo[0], N.hi[0]), ey = gridExponent(N.lo[1], N.hi[1]), ez = gridExponent(N.lo[2], N.hi[2]);
    const float sx = __int_as_float(ex << 23), sy = __int_as_float(ey << 23), sz = __int_as_float(ez << 23);
    uint32_t imask = 0; int nInner = 0, nPrims = 0;
    for (int sl = 0; sl < 8; ++sl) {
        int i = childIn[sl];
        if (i < 0) continue;
        const BuildNode &c = bn[child[i]];
        if (c.left >= 0) { imask |= 1u << sl; ++nInner; } else nPrims += c.end - c.begin;
    }
    const int childBase = nInner ? atomicAdd(nodeCounter, nInner) : 0;
    const int primBase = nPrims ? atomicAdd(primCounter, nPrims) : 0;
    uint32_t meta[2] = {0, 0}, q[6][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}};
    int innerRank = 0, primOff = 0;
    for (int sl = 0; sl < 8; ++sl) {
        int i = childIn[sl];
        const int w = sl >> 2, sh = (sl & 3) * 8;
        if (i < 0) {   // empty slot: inverted box, never hit
            for (int a = 0; a < 3; ++a) { q[a][w] |= 255u << sh; }
            continue;
        }
        const BuildNode &c = bn[child[i]];
        // conservative quantisation: lo rounds down, hi rounds up, then verified in the arithmetic of the traversal
        const float lo[3] = {N.lo[0], N.lo[1], N.lo[2]}, sc[3] = {sx, sy, sz};
        for (int a = 0; a < 3; ++a) {
            int ql = (int)floorf((c.lo[a] - lo[a]) / sc[a]), qh = (int)ceilf((c.hi[a] - lo[a]) / sc[a]);
            ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
            while (ql > 0 && lo[a] + (float)ql * sc[a] > c.lo[a]) --ql;
            while (qh < 255 && lo[a] + (float)qh * sc[a] < c.hi[a]) ++qh;
            q[a][w] |= (uint32_t)ql << sh; q[3 + a][w] |= (uint32_t)qh << sh;
        }
        if (c.left >= 0) {
            meta[w] |= (uint32_t)((1u << 5) | (24u + sl)) << sh;
            srcNode[childBase + innerRank] = child[i];
            ++innerRank;
        } else {
            const int cnt = c.end - c.begin;   // 1..3
            const uint32_t unary = cnt == 1 ? 1u : (cnt == 2 ? 3u : 7u);
            meta[w] |= (uint32_t)((unary << 5) | (uint32_t)primOff) << sh;
            for (int j = 0; j < cnt; ++j) prims[primBase + primOff + j] = makePrimRecord(src[order[c.begin + j]], pos, idx);
            primOff += cnt;
        }
    }
    WideNode o;
    o.n0 = make_float4(N.lo[0], N.lo[1], N.lo[2], __int_as_float((int)((uint32_t)ex | ((uint32_t)ey << 8) | ((uint32_t)ez << 16) | (imask << 24))));
    o.n1 = make_uint4((uint32_t)childBase, (uint32_t)primBase, meta[0], meta[1]);
    o.n2 = make_uint4(q[0][0], q[0][1], q[1][0], q[1][1]);
    o.n3 = make_uint4(q[2][0], q[2][1], q[3][0], q[3][1]);
    o.n4 = make_uint4(q[4][0], q[4][1], q[5][0], q[5][1]);
    out[k] = o;
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
    PrimSource *src = nullptr; WideNode *wide = nullptr; int *srcNode = nullptr, *primCounter = nullptr;
    int *smallList = nullptr, *smallCount = nullptr;
    int launches = 0;
    if (n == 0) { out.nodes = nullptr; out.prims = nullptr; out.numNodes = 0; return true; }
    const bool trace = getenv("MTS_BUILD_TRACE") != nullptr;
    auto now = [&]() { if (trace) cudaStreamSynchronize(st); return std::chrono::steady_clock::now(); };
    auto t0 = now();
    auto lap = [&](const char *what) { if (!trace) return; auto t1 = now(); fprintf(stderr, "[bvh] %-10s %.2f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count()); t0 = t1; };
    {
    const size_t maxNodes = 2 * (size_t)n + 2;
    CK(devAlloc(&pLo, sizeof(float4) * n, st)); CK(devAlloc(&pHi, sizeof(float4) * n, st));
    CK(devAlloc(&idxA, 4 * (size_t)n, st)); CK(devAlloc(&idxB, 4 * (size_t)n, st));
    CK(devAlloc(&nodeA, 4 * (size_t)n, st)); CK(devAlloc(&nodeB, 4 * (size_t)n, st));
    CK(devAlloc(&nodes, sizeof(BuildNode) * maxNodes, st));
    CK(devAlloc(&flags, 4 * maxNodes, st)); CK(devAlloc(&scan, 4 * maxNodes, st));
    CK(devAlloc(&box, 4 * 12, st)); CK(devAlloc(&nodeCount, 4, st));
    CK(devAlloc(&src, sizeof(PrimSource) * n, st));
    CK(devAlloc(&smallList, 4 * ((size_t)n / 2 + 1), st)); CK(devAlloc(&smallCount, 4, st));
    CK(cudaMemsetAsync(smallCount, 0, 4, st));
    CK(cub::DeviceScan::ExclusiveSum(nullptr, cubBytes, flags, scan, (int)maxNodes, st));
    CK(devAlloc(&cubTmp, cubBytes, st));

    lap("alloc");
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
    k_root_bounds<<<std::min(blocks(n), 148 * 8), 256, 0, st>>>(n, pLo, pHi, box);
    k_set_root<<<1, 1, 0, st>>>(nodes, box, n);
    launches += 3;

    lap("bounds");
    int levelStart = 0, levelEnd = 1;
    while (levelStart < levelEnd) {
        size_t nb = (size_t)(levelEnd - levelStart) * 3 * NB;
        if (nb > binCap) {
            devFree(bins, st);
            binCap = nb + nb / 2;
            CK(devAlloc(&bins, sizeof(Bin) * binCap, st));
        }
        k_clear_bins<<<blocks(nb), 256, 0, st>>>(bins, nb);
        k_bin<<<blocks(n, kBinChunk), 256, 0, st>>>(n, idxA, nodeA, pLo, pHi, nodes, levelStart, bins);
        k_split<<<blocks(levelEnd - levelStart, 64), 64, 0, st>>>(levelStart, levelEnd, nodes, bins, nodeCount, smallList, smallCount);
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
    lap("levels");
    int nBuild = levelEnd;
    {
        int nSmall = 0;
        CK(cudaMemcpyAsync(&nSmall, smallCount, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (nSmall > 0) {
            k_small_subtrees<<<blocks(nSmall, 4), 128, 0, st>>>(smallList, nSmall, nodes, idxA, pLo, pHi, nodeCount);
            ++launches;
            CK(cudaMemcpyAsync(&nBuild, nodeCount, 4, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if ((size_t)nBuild > maxNodes) { err = "BVH build: node overflow"; goto fail; }
        }
    }
    lap("small");
    // collapse into eight-wide nodes, level by level (the wide tree of 10M triangles is ~10 levels deep)
    {
        const size_t maxWide = (size_t)nBuild / 2 + 2;   // every wide node consumes at least one binary inner node
        CK(devAlloc(&wide, sizeof(WideNode) * maxWide, st));
        CK(devAlloc(&srcNode, 4 * maxWide, st));
        CK(devAlloc(&primCounter, 4, st));
        CK(devAlloc(&out.prims, sizeof(PrimRecord) * n, st));
        int zero = 0, one = 1;
        CK(cudaMemcpyAsync(srcNode, &zero, 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(primCounter, &zero, 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(nodeCount, &one, 4, cudaMemcpyHostToDevice, st));
        int ws = 0, we = 1;
        while (ws < we) {
            k_collapse<<<blocks(we - ws, 64), 64, 0, st>>>(ws, we, nodes, idxA, srcNode, wide, nodeCount, primCounter, src,
                                                         in.positions, in.indices, out.prims);
            ++launches;
            int cnt = 0;
            CK(cudaMemcpyAsync(&cnt, nodeCount, 4, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if ((size_t)cnt > maxWide) { err = "BVH collapse: node overflow"; goto fail; }
            ws = we; we = cnt;
            ++out.depth;
        }
        out.numNodes = we;
        CK(devAlloc(&out.nodes, sizeof(WideNode) * (size_t)we, st));
        CK(cudaMemcpyAsync(out.nodes, wide, sizeof(WideNode) * (size_t)we, cudaMemcpyDeviceToDevice, st));
        CK(cudaStreamSynchronize(st));
    }
    BuildNode root;
    CK(cudaMemcpyAsync(&root, nodes, sizeof(root), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    for (int a = 0; a < 3; ++a) { out.lo[a] = root.lo[a]; out.hi[a] = root.hi[a]; }
    out.launches = launches;
    out.buildNodes = nBuild;
    lap("collapse");
    }
    devFree(pLo, st); devFree(pHi, st); devFree(idxA, st); devFree(idxB, st); devFree(nodeA, st); devFree(nodeB, st); devFree(nodes, st);
    devFree(flags, st); devFree(scan, st); devFree(box, st); devFree(nodeCount, st); devFree(bins, st); devFree(cubTmp, st); devFree(src, st); devFree(wide, st); devFree(srcNode, st); devFree(primCounter, st); devFree(smallList, st); devFree(smallCount, st);
    lap("free");
    return true;
fail:
    devFree(pLo, st); devFree(pHi, st); devFree(idxA, st); devFree(idxB, st); devFree(nodeA, st); devFree(nodeB, st); devFree(nodes, st);
    devFree(flags, st); devFree(scan, st); devFree(box, st); devFree(nodeCount, st); devFree(bins, st); devFree(cubTmp, st); devFree(src, st); devFree(wide, st); devFree(srcNode, st); devFree(primCounter, st); devFree(smallList, st); devFree(smallCount, st);
    if (out.nodes) devFree(out.nodes, st);
    if (out.prims) devFree(out.prims, st);
    out = BvhBuildOutput();
    return false;
}

} // namespace mts
