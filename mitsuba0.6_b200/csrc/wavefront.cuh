// wavefront.cuh -- the path-tracing hot path as a wavefront of sm_100a kernels.
//
// MIPathTracer::Li (src/integrators/path/path.cpp:119-294) is a per-sample recursive loop of
// virtual calls.  Here one wavefront holds P path slots in SoA float4 arrays; every bounce runs
//   intersect (closest hit) -> shade (emission + MIS, NEE sample, BSDF sample, Russian roulette)
//   -> shadow (any-hit for the NEE rays) -> splat (finished paths into the film)
// over compacted index queues built with warp-aggregated atomics (__ballot_sync/__popc, one
// atomicAdd per warp).  All kernels are persistent grid-stride kernels whose trip count is read
// from device memory, so a bounce needs no host round trip to size its launches.
#pragma once
#include "bsdf.cuh"
#include "emitter.cuh"

namespace mts {

// Every counter that a kernel's warps bump concurrently sits in a 128-byte line of its own (COUNTER_LINE): the queue
// appends of one shade kernel (next / shadow / done) are three streams of one-per-warp atomics, and atomics on one line
// are served one after the other by the same L2 slice.
#ifndef COUNTER_LINE
#define COUNTER_LINE 128
#endif
struct Counters {
    alignas(COUNTER_LINE) uint32_t qCount[2];     // active queue sizes (ping-pong: one is read, the other appended to)
    alignas(COUNTER_LINE) uint32_t shadowCount;
    alignas(COUNTER_LINE) uint32_t doneCount;
    alignas(COUNTER_LINE) uint32_t ewaCount;      // deferred EWA texture walks of the current bounce (k_shade_surface -> k_shade_ewa)
    alignas(COUNTER_LINE) uint32_t cursor[2];     // next unclaimed ray of the intersect / shadow queue (persistent warps pull from these)
    alignas(COUNTER_LINE) uint32_t ewaBin[3 * 16]; // size-class histogram, offsets and cursors of the walks (kEwaBins each; all zero between uses)
    alignas(COUNTER_LINE) unsigned long long missRays;                 // rays that escaped (sort key 0), total over all bounces
    unsigned long long normalRays, shadowRays;   // totals over all bounces ("Normal/Shadow rays traced", skdtree.cpp:46-47), kept on the device
    unsigned long long pathLenSum;
    unsigned long long nodesVisited, primsTested;  // instrumented traces only (closest hit)
    unsigned long long nodesVisitedShadow, primsTestedShadow;
};

struct Wavefront {
    float4 *rayO, *rayD;     // o.xyz + mint, d.xyz + maxt
    float4 *hit;             // t, u, v, prim slot (int bits)
    float4 *thr;             // throughput rgb, eta
    float4 *L;               // radiance rgb, pdf of the BSDF sample that produced the current ray
    uint4 *meta;             // pixel index, global sample index, depth | flags << 16, unused
    float4 *shO, *shD, *shC; // shadow ray origin/dir (+mint/maxt), contribution rgb + path slot
    float4 *rec;             // surface records of the staged shade kernels: kSurfRows rows of `capacity` float4, indexed by sorted-queue position
    float4 *ewaJobs;         // deferred EWA walks: 2 rows of `capacity` float4 (target / textures / level / lerp weight; uv, A, B) ...
    float *ewaC;             // ... and the ellipse coefficient C
    uint32_t *ewaOrder;      // the walks ordered by size class
    uint32_t *queue[2];
    uint32_t *doneQ;
    // material sort of the active queue: key 0 = miss, DShape::sortKey otherwise (root BSDFs ordered by shade class)
    uint32_t *keys;          // per queue position, written by intersect
    uint32_t *sorted;        // queue[cur] reordered by key; what shade consumes
    uint32_t *keyCount;      // numKeys histogram, then numKeys offsets, then numKeys cursors
    uint32_t *raySortBin;    // the same three arrays for the ray sort (kRaySortKeys each; MTS_RAY_SORT=1), else null
    uint32_t numKeys;
    uint32_t classKey[SHADE_CLASSES + 1];   // sort keys [classKey[c], classKey[c+1]) belong to shade class c (key 0 = miss)
    Counters *ctr;
    uint32_t capacity;
};

struct RenderConst {
    int maxDepth, rrDepth, strictNormals, hideEmitters;
    uint32_t seed;
    int sampleOffset, sampleStride;
    int tilesX, tilesY;      // 8x4 pixel tiles
    uint32_t paddedPixels;   // tilesX*tilesY*32
    float diffScale;         // 1/sqrt(sampler sample count): RayDifferential::scaleDifferential, integrator.cpp:144-145,181
};

enum { PF_SCATTERED = 1, PF_PREV_DELTA = 2, PF_PREV_REFN_OK = 4,
       PF_CAMERA_HIT = 8 };   // the camera ray hit a surface: alpha = 1 (RadianceQueryRecord::rayIntersect, records.inl:121-137)

// One atomicAdd per warp: returns this thread's slot (valid only where pred is true).
DEVFN uint32_t warpAppend(uint32_t *counter, bool pred) {
    unsigned active = __activemask();
    unsigned mask = __ballot_sync(active, pred);
    if (!pred) return 0xffffffffu;
    int lane = threadIdx.x & 31;
    int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(mask, base, leader);
    return base + __popc(mask & ((1u << lane) - 1));
}

// One atomicAdd per CTA: every thread of the block must call it (same trip count for all warps).  The light kernels
// (raygen, escaped rays) bump one global counter once per warp otherwise -- half a million atomics on one address in
// half a millisecond.
DEVFN uint32_t blockAppend(uint32_t *counter, bool pred) {
    __shared__ uint32_t sWarpCount[32], sBlockBase;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (lane == 0) sWarpCount[warp] = __popc(mask);
    __syncthreads();
    if (warp == 0) {
        uint32_t v = lane < nwarps ? sWarpCount[lane] : 0, incl = v;
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        uint32_t base = 0;
        if (lane == 0 && total) base = atomicAdd(counter, total);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (lane < nwarps) sWarpCount[lane] = base + incl - v;
    }
    __syncthreads();
    const uint32_t r = sWarpCount[warp] + __popc(mask & ((1u << lane) - 1));
    __syncthreads();   // the next call reuses the shared array
    return pred ? r : 0xffffffffu;
}

DEVFN Surf surfOf(const Hit &h) {
    Surf s; s.tc = texCtx(h.uv); s.dpdu = h.dpdu; s.dpdv = h.dpdv; s.geoN = h.geoN; s.sh = h.sh; s.bumpNode = -1; return s;
}

// The camera ray through `samplePos` with its differentials scaled like SamplingIntegrator::renderBlock does
// (sampleRayDifferential, perspective.cpp:271-297; scaleDifferential, ray.h:163-168).  Origins coincide with o.
DEVFN void cameraRayDiff(const MtsCamera &cam, float2 samplePos, float diffScale, float3 &o, float3 &d, float3 &rxD, float3 &ryD) {
    float mint, maxt;
    cameraRay(cam, samplePos, o, d, mint, maxt, &rxD, &ryD);
    rxD = d + (rxD - d) * diffScale;
    ryD = d + (ryD - d) * diffScale;
}

// Intersection::computePartials (src/librender/intersection.cpp:5-76): uv partials wrt. the screen position from the
// offset rays' intersections with the tangent plane.  Returns dudx dvdx dudy dvdy.
DEVFN float4 computePartials(float3 p, float3 n, float3 dpdu, float3 dpdv, float3 o, float3 rxD, float3 ryD) {
    if (isZero(dpdu) && isZero(dpdv)) return make_float4(0, 0, 0, 0);
    const float pp = dot(n, p), po = dot(n, o), prx = dot(n, rxD), pry = dot(n, ryD);
    if (prx == 0 || pry == 0) return make_float4(0, 0, 0, 0);
    const float tx = (pp - po) / prx, ty = (pp - po) / pry;
    const float ax = fabsf(n.x), ay = fabsf(n.y), az = fabsf(n.z);
    int a0, a1;
    if (ax > ay && ax > az) { a0 = 1; a1 = 2; } else if (ay > az) { a0 = 0; a1 = 2; } else { a0 = 0; a1 = 1; }
    auto comp = [](float3 v, int a) { return a == 0 ? v.x : a == 1 ? v.y : v.z; };
    const float A00 = comp(dpdu, a0), A01 = comp(dpdv, a0), A10 = comp(dpdu, a1), A11 = comp(dpdv, a1);
    const float3 px = o + rxD * tx, py = o + ryD * ty;
    const float bx0 = comp(px, a0) - comp(p, a0), bx1 = comp(px, a1) - comp(p, a1);
    const float by0 = comp(py, a0) - comp(p, a0), by1 = comp(py, a1) - comp(p, a1);
    const float det = A00 * A11 - A01 * A10;
    // solveLinearSystem2x2 (src/libcore/util.cpp:527-539); the failure values are the reference's (:64-74, including
    // its `dudy = 0; dudy = 1` slip, which leaves dvdy at its initial 0)
    if (fabsf(det) <= 2.93873587705571876e-39f) return make_float4(1, 0, 1, 0);
    const float inv = 1.0f / det;
    return make_float4((A11 * bx0 - A01 * bx1) * inv, (A00 * bx1 - A10 * bx0) * inv,
                       (A11 * by0 - A01 * by1) * inv, (A00 * by1 - A10 * by0) * inv);
}

// ---- raygen: SamplingIntegrator::renderBlock sample loop + sampleRayDifferential -------------
static __global__ void __launch_bounds__(256) k_raygen(const DScene S, const Wavefront W, const RenderConst R,
                                                unsigned long long firstItem, uint32_t count, int qsel) {
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < ((count + 255u) & ~255u); j += gridDim.x * blockDim.x) {   // 256-thread CTAs, whole CTAs per trip
        bool valid = j < count;
        uint32_t pixel = 0, sample = 0;
        int x = 0, y = 0;
        if (valid) {
            unsigned long long g = firstItem + j;
            uint32_t sLocal = (uint32_t)(g / R.paddedPixels), pp = (uint32_t)(g % R.paddedPixels);
            uint32_t tile = pp >> 5, lane = pp & 31;
            x = (int)(tile % R.tilesX) * 8 + (int)(lane & 7);
            y = (int)(tile / R.tilesX) * 4 + (int)(lane >> 3);
            valid = x < S.filmW && y < S.filmH;
            pixel = (uint32_t)y * S.filmW + x;
            sample = (uint32_t)(R.sampleOffset + (int)sLocal * R.sampleStride);
        }
        if (valid) {
            float4 u = rand4(R.seed, pixel, sample, 0);
            float3 o, d; float mint, maxt;
            cameraRay(S.cam, make_float2(x + u.x, y + u.y), o, d, mint, maxt);
            W.rayO[j] = make_float4(o.x, o.y, o.z, mint);
            W.rayD[j] = make_float4(d.x, d.y, d.z, maxt);
            W.thr[j] = make_float4(1, 1, 1, 1);
            W.L[j] = make_float4(0, 0, 0, 0);
            W.meta[j] = make_uint4(pixel, sample, 1u, 0u);
        }
        uint32_t slot = blockAppend(&W.ctr->qCount[qsel], valid);
        if (valid) W.queue[qsel][slot] = j;
    }
}

// Lanes of a warp that hold the same key elect a leader that reserves room for all of them with one atomicAdd.
// Returns this lane's position (valid only where key != kNoKey).
constexpr uint32_t kNoKey = 0xffffffffu;
constexpr uint32_t kSortSmemKeys = 512;   // sort keys (distinct root BSDFs + 1) up to which the sort counts per CTA in shared memory
DEVFN uint32_t keyedAppend(uint32_t *counters, uint32_t key) {
    unsigned active = __activemask();
    unsigned peers = __match_any_sync(active, key);
    if (key == kNoKey) return 0;
    int lane = threadIdx.x & 31;
    int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counters + key, __popc(peers));
    base = __shfl_sync(peers, base, leader);
    return base + __popc(peers & ((1u << lane) - 1));
}

// ---- persistent-warp ray queue driver -----------------------------------------------------------
// Every warp keeps 32 traversals in flight: lanes whose ray has finished pull the next ray index from a global
// cursor (one atomicAdd per warp) as soon as fewer than kRefillBelow lanes are still busy, instead of idling
// until the slowest ray of the warp is done.  LOAD(i, traversal) starts ray i, DONE(i, traversal) consumes it.
#ifndef TRACE_MINB
#define TRACE_MINB 8
#endif
#ifndef REFILL_BELOW
#define REFILL_BELOW 24
#endif
constexpr int kRefillBelow = REFILL_BELOW;

template <bool ANY, bool COUNT, class Load, class Done>
DEVFN void traceQueue(const DScene &S, uint32_t count, uint32_t *cursor, TraceCounters *tc, Load load, Done done) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    __shared__ uint2 sstack[kSmemStack > 0 ? kSmemStack * kTraceBlock : 1];
    Traversal<ANY, COUNT, kSmemStack> tr;
    tr.smBase = (uint32_t)__cvta_generic_to_shared(&sstack[threadIdx.x]);
#if TRACE_SMEM_NODES > 0
    __shared__ float4 stop[5 * TRACE_SMEM_NODES];
    {
        const uint32_t nTop = min((uint32_t)TRACE_SMEM_NODES, (uint32_t)S.numNodesBvh);
        const float4 *src = reinterpret_cast<const float4 *>(S.bvh);
        for (uint32_t k = threadIdx.x; k < 5 * nTop; k += blockDim.x) stop[k] = __ldg(src + k);
        __syncthreads();
        tr.topNodes = nTop == (uint32_t)TRACE_SMEM_NODES ? stop : nullptr;
    }
#endif
    uint32_t item = 0;
    bool busy = false, exhausted = false;
    while (true) {
        if (!exhausted) {
            // refill idle lanes
            const unsigned idle = __ballot_sync(full, !busy);
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(cursor, (uint32_t)__popc(idle));
            base = __shfl_sync(full, base, 0);
            if (!busy) {
                item = base + __popc(idle & ((1u << lane) - 1));
                if (item < count) {
                    busy = load(item, tr);
                    if (!busy) done(item, tr);   // finished before it started (empty interval)
                }
            }
            exhausted = base + __popc(idle) >= count;
        }
        unsigned active = __ballot_sync(full, busy);
        if (!active) { if (exhausted) break; else continue; }
        // traverse until too few lanes are busy (and there is more work to fetch)
        do {
            if (busy && tr.step(S, tc)) { done(item, tr); busy = false; }
            active = __ballot_sync(full, busy);
        } while (active && (exhausted || __popc(active) >= kRefillBelow));
    }
}

// ---- intersect: Scene::rayIntersect -> ShapeKDTree::rayIntersect(ray, its) --------------------
// Also emits the shade sort key of every ray (0 = escaped, 1 + BSDF root of the hit shape) and its histogram.
template <bool COUNT>
static __global__ void __launch_bounds__(128, TRACE_MINB) k_intersect(const DScene S, const Wavefront W, int qsel) {
    const uint32_t count = W.ctr->qCount[qsel];
    TraceCounters tc = {0, 0};
    const uint32_t *queue = W.queue[qsel];
    // (the key histogram stays in global memory: counting it per CTA in shared memory cost 0.8 ms per wavefront in spills,
    //  profiles/r3a_sort.txt -- unlike k_sort_scatter, where the per-CTA form is 8x faster)
    traceQueue<false, COUNT>(S, count, &W.ctr->cursor[0], &tc,
        [&](uint32_t i, Traversal<false, COUNT, kSmemStack> &tr) {
            uint32_t slot = queue[i];
            float4 ro = W.rayO[slot], rd = W.rayD[slot];
            return tr.init(S, f3(ro), f3(rd), ro.w, rd.w);
        },
        [&](uint32_t i, const Traversal<false, COUNT, kSmemStack> &tr) {
            uint32_t slot = queue[i];
            RayHit h = tr.result();
            W.hit[slot] = make_float4(h.t, h.u, h.v, __int_as_float(h.prim));
            uint32_t key = 0;
            if (h.prim >= 0) {
                int shape = __float_as_int(__ldg(reinterpret_cast<const float4 *>(S.prims + h.prim)).w);
                key = S.shapes[shape].sortKey;
            }
            W.keys[i] = key;
            keyedAppend(W.keyCount, key);   // lanes finishing in the same step share one atomicAdd per key
        });
    if (COUNT) { atomicAdd(&W.ctr->nodesVisited, (unsigned long long)tc.nodes); atomicAdd(&W.ctr->primsTested, (unsigned long long)tc.prims); }
}

// ---- shadow: ShapeKDTree::rayIntersect(ray) for Scene::sampleEmitterDirect --------------------
template <bool COUNT>
static __global__ void __launch_bounds__(128, TRACE_MINB) k_shadow(const DScene S, const Wavefront W) {
    const uint32_t count = W.ctr->shadowCount;
    TraceCounters tc = {0, 0};
    traceQueue<true, COUNT>(S, count, &W.ctr->cursor[1], &tc,
        [&](uint32_t i, Traversal<true, COUNT, kSmemStack> &tr) {
            float4 ro = W.shO[i], rd = W.shD[i];
            return tr.init(S, f3(ro), f3(rd), ro.w, rd.w);
        },
        [&](uint32_t i, const Traversal<true, COUNT, kSmemStack> &tr) {
            if (!tr.hit()) {
                float4 c = W.shC[i];
                uint32_t slot = (uint32_t)__float_as_int(c.w);
                float4 l = W.L[slot];
                l.x += c.x; l.y += c.y; l.z += c.z;
                W.L[slot] = l;
            }
        });
    if (COUNT) { atomicAdd(&W.ctr->nodesVisitedShadow, (unsigned long long)tc.nodes); atomicAdd(&W.ctr->primsTestedShadow, (unsigned long long)tc.prims); }
}

// ---- material sort: counting sort of the active queue by key (histogram from intersect) --------
static __global__ void k_sort_offsets(const Wavefront W) {
    // one warp; numKeys is small (BSDF nodes + 1)
    uint32_t *cnt = W.keyCount, *off = W.keyCount + W.numKeys, *cur = W.keyCount + 2 * W.numKeys;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < W.numKeys; base += 32) {
        uint32_t k = base + threadIdx.x;
        uint32_t v = k < W.numKeys ? cnt[k] : 0, incl = v;
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)threadIdx.x >= o) incl += t; }
        if (k < W.numKeys) { off[k] = carry + incl - v; cur[k] = 0; cnt[k] = 0; }
        if (k == 0) W.ctr->missRays += v;
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
}
// Positions are handed out per CTA: the 256 entries of a trip count their keys in shared memory (warp-aggregated), one
// global atomicAdd per key and CTA reserves the range, and every entry lands at range start + its rank within the CTA.
// (One global atomic per key and WARP on a handful of hot counters was atomic-bound: 2.1 ms per wavefront for 35 M
// entries, 200 GB/s.)  Scenes with more sort keys than kSortSmemKeys fall back to the per-warp form.
static __global__ void __launch_bounds__(256) k_sort_scatter(const Wavefront W, int qsel) {
    const uint32_t count = W.ctr->qCount[qsel];
    const uint32_t *off = W.keyCount + W.numKeys;
    uint32_t *cur = W.keyCount + 2 * W.numKeys;
    if (W.numKeys > kSortSmemKeys) {
        const uint32_t rounded = (count + 31u) & ~31u;
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < rounded; i += gridDim.x * blockDim.x) {
            uint32_t key = kNoKey, slot = 0;
            if (i < count) { key = W.keys[i]; slot = W.queue[qsel][i]; }
            uint32_t pos = keyedAppend(cur, key);
            if (key != kNoKey) W.sorted[off[key] + pos] = slot;
        }
        return;
    }
    __shared__ uint32_t sCnt[kSortSmemKeys], sBase[kSortSmemKeys];
    const uint32_t rounded = (count + 255u) & ~255u;
    for (uint32_t base = blockIdx.x * blockDim.x; base < rounded; base += gridDim.x * blockDim.x) {   // whole CTAs take part in every trip
        const uint32_t i = base + threadIdx.x;
        for (uint32_t k = threadIdx.x; k < W.numKeys; k += blockDim.x) sCnt[k] = 0;
        __syncthreads();
        uint32_t key = kNoKey, slot = 0;
        if (i < count) { key = W.keys[i]; slot = W.queue[qsel][i]; }
        const uint32_t pos = keyedAppend(sCnt, key);          // rank within the CTA (shared-memory atomics)
        __syncthreads();
        for (uint32_t k = threadIdx.x; k < W.numKeys; k += blockDim.x) { const uint32_t c = sCnt[k]; if (c) sBase[k] = atomicAdd(cur + k, c); }
        __syncthreads();
        if (key != kNoKey) W.sorted[off[key] + sBase[key] + pos] = slot;
        __syncthreads();
    }
}

// ---- ray sort (MTS_RAY_SORT=1) -----------------------------------------------------------
// Bins the rays of a bounce by direction octant and origin cell (8 x 4x4x4 = 512 keys, counting sort like the material
// sort) before k_intersect, so that the 32 rays a warp pulls from the queue start in the same region of the tree and
// visit the children of a node in the same order.  Camera rays are coherent already: used from the first bounce on.
constexpr uint32_t kRaySortKeys = 512;
struct RaySortGrid { float3 lo, invCell; };
static __global__ void __launch_bounds__(256) k_raysort_keys(const Wavefront W, int qsel, const RaySortGrid G) {
    __shared__ uint32_t sCnt[kRaySortKeys];
    const uint32_t count = W.ctr->qCount[qsel];
    for (uint32_t k = threadIdx.x; k < kRaySortKeys; k += blockDim.x) sCnt[k] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const uint32_t slot = W.queue[qsel][i];
        const float4 o = W.rayO[slot], d = W.rayD[slot];
        const int cx = min(3, max(0, (int)((o.x - G.lo.x) * G.invCell.x))), cy = min(3, max(0, (int)((o.y - G.lo.y) * G.invCell.y))),
                  cz = min(3, max(0, (int)((o.z - G.lo.z) * G.invCell.z)));
        const uint32_t oct = (d.x >= 0 ? 1u : 0u) | (d.y >= 0 ? 2u : 0u) | (d.z >= 0 ? 4u : 0u);
        const uint32_t key = (oct << 6) | (uint32_t)(cz << 4 | cy << 2 | cx);
        W.keys[i] = key;
        atomicAdd(&sCnt[key], 1u);
    }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < kRaySortKeys; k += blockDim.x) { const uint32_t c = sCnt[k]; if (c) atomicAdd(W.keyCount + k, c); }
}
static __global__ void k_raysort_offsets(const Wavefront W) {   // one warp; W.keyCount / numKeys = the ray-sort bins
    uint32_t *cnt = W.keyCount, *off = W.keyCount + W.numKeys, *cur = W.keyCount + 2 * W.numKeys;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < W.numKeys; base += 32) {
        const uint32_t k = base + threadIdx.x;
        uint32_t v = k < W.numKeys ? cnt[k] : 0, incl = v;
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)threadIdx.x >= o) incl += t; }
        if (k < W.numKeys) { off[k] = carry + incl - v; cur[k] = 0; cnt[k] = 0; }
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
}

// ---- shade: one iteration of the Li loop ----------------------------------------------------
struct ShadeOut { bool next, shadow, done; float4 shO, shD, shC; };

// What the second half of an iteration (light sampling, BSDF sampling) needs to know about the hit and the path.
struct PathVertex {
    float3 p, geoN, wi;
    Frame sh;
    float3 thr, L;
    float eta;
    uint32_t depth, flags, lobes;
    uint4 meta;
    uint32_t touched;   // what shadeVertex changed of the stored path state: 1 = L, 2 = throughput, 4 = flags (staged shade writes only these back)
};

// First half of an iteration: hit record, emission + MIS of the previous BSDF sample, Russian roulette, depth / normal
// tests (path.cpp:145-165, 257-264, 276-286).  Returns false when the path ends here (its state has been finished).
// On success `su` is the surface context of the hit (with the uv partials of a camera-ray hit) and `bsdf` its root BSDF.
DEVFN bool shadeVertex(const DScene &S, const Wavefront &W, const RenderConst &R, uint32_t slot, PathVertex &v, Surf &su, int &bsdf) {
    float4 ro = W.rayO[slot], rd = W.rayD[slot], hv = W.hit[slot];
    float4 tv = W.thr[slot], lv = W.L[slot];
    uint4 meta = W.meta[slot];
    float3 o = f3(ro), d = f3(rd);
    float3 thr = f3(tv), L = f3(lv);
    float eta = tv.w, prevPdf = lv.w;
    uint32_t depth = meta.z & 0xffffu, flags = meta.z >> 16;
    const int prim = __float_as_int(hv.w);
    const bool first = depth == 1;
    uint32_t touched = 0;
    if (first) { flags |= PF_CAMERA_HIT; touched |= 4u; }

    auto finish = [&]() { W.L[slot] = make_float4(L.x, L.y, L.z, 0); W.meta[slot] = make_uint4(meta.x, meta.y, depth | (flags << 16), 0); };

    RayHit rh; rh.t = hv.x; rh.u = hv.y; rh.v = hv.z; rh.prim = prim;
    Hit its;
    fillHit(S, o, d, rh, its);
    const DShape &shape = S.shapes[its.shape];

    if (shape.emitter >= 0) {
        const DEmitter &em = S.emitters[shape.emitter];
        float3 Le = areaLe(em, its.sh.n, -d);
        touched |= 1u;
        if (first) {
            if (!R.hideEmitters) L += thr * Le;  // path.cpp:148-150
        } else {
            // path.cpp:257-264 with Scene::pdfEmitterDirect (scene.cpp:949-952) and AreaLight::pdfDirect (area.cpp:174-183)
            float lumPdf = 0.0f;
            if (!(flags & PF_PREV_DELTA) && (flags & PF_PREV_REFN_OK) && dot(d, its.sh.n) < 0)
                lumPdf = shapePdfDirect(shape, o, d, its.sh.n, its.t) * em.prob;
            L += thr * Le * miWeight(prevPdf, lumPdf);
        }
    }

    if (!first) {
        // Russian roulette of the previous iteration (path.cpp:276-286): `depth` was already incremented
        if ((int)(depth - 1) >= R.rrDepth) {
            float q = fminf(maxc(thr) * eta * eta, 0.95f);
            float4 u = rand4(R.seed, meta.x, meta.y, 2 * depth - 2);
            if (u.x >= q) { finish(); return false; }
            thr = thr / q;
            touched |= 2u;
        }
    }

    // path.cpp:156-165
    if ((R.maxDepth > 0 && (int)depth >= R.maxDepth) ||
        (R.strictNormals && dot(d, its.geoN) * cosTheta(its.wi) >= 0)) { finish(); return false; }

    bsdf = shape.bsdf;
    su = surfOf(its);
    if (first && (S.nodes[bsdf].flags & MTS_BF_RAY_DIFF)) {
        // Intersection::getBSDF(ray) (records.inl:69-75): only the camera ray carries differentials (path.cpp:225 drops them)
        const float4 us = rand4(R.seed, meta.x, meta.y, 0);
        const int x = (int)(meta.x % (uint32_t)S.filmW), y = (int)(meta.x / (uint32_t)S.filmW);
        float3 co, cd, rxD, ryD;
        cameraRayDiff(S.cam, make_float2(x + us.x, y + us.y), R.diffScale, co, cd, rxD, ryD);
        su.tc.duv = computePartials(its.p, its.geoN, its.dpdu, its.dpdv, co, rxD, ryD);
        su.tc.filtered = true;
    }
    v.p = its.p; v.geoN = its.geoN; v.wi = its.wi; v.sh = its.sh;
    v.thr = thr; v.L = L; v.eta = eta; v.depth = depth; v.flags = flags; v.meta = meta;
    v.lobes = S.nodes[bsdf].lobes;
    v.touched = touched;
    return true;
}

// Second half: direct illumination sampling (path.cpp:172-200) and BSDF sampling (path.cpp:203-253).
template <class Mat>
DEVFN void shadeScatter(const DScene &S, const Wavefront &W, const RenderConst &R, uint32_t slot, const PathVertex &v, const Mat &mat, ShadeOut &out) {
    float3 thr = v.thr;
    const float3 L = v.L;
    float eta = v.eta;
    uint32_t depth = v.depth, flags = v.flags;
    const uint4 meta = v.meta;
    const uint32_t lobes = v.lobes;
    auto finish = [&]() { W.L[slot] = make_float4(L.x, L.y, L.z, 0); W.meta[slot] = make_uint4(meta.x, meta.y, depth | (flags << 16), 0); };
    float uExtra;
    float4 u = rand4x(R.seed, meta.x, meta.y, 2 * depth - 1, uExtra);

    // ---- direct illumination sampling (path.cpp:172-200) ----
    float3 refN = ((lobes & (MTS_ETRANSMISSION | MTS_EBACKSIDE)) == 0) ? v.sh.n : f3(0.0f);  // records.inl:160-164
    if (lobes & MTS_ESMOOTH) {
        DRec dr; dr.ref = v.p; dr.refN = refN; dr.pdf = 0; dr.emitter = -1; dr.measure = 0; dr.dist = 0;
        dr.d = f3(0.0f); dr.n = f3(0.0f); dr.p = f3(0.0f);
        float3 value = sampleEmitterDirect(S, dr, make_float2(u.x, u.y));
        if (!isZero(value)) {
            float3 wo = v.sh.toLocal(dr.d);
            // area and environment emitters are "on surface"; the point emitter is not
            const bool wantPdf = S.emitters[dr.emitter].type != MTS_EMITTER_POINT && dr.measure == MTS_MEASURE_SOLID_ANGLE;
            float bsdfPdf;
            float3 bsdfVal = mat.evalPdf(v.wi, wo, MTS_MEASURE_SOLID_ANGLE, wantPdf ? PDF_IF_LIT : PDF_NONE, bsdfPdf);
            if (!isZero(bsdfVal) && (!R.strictNormals || dot(v.geoN, dr.d) * cosTheta(wo) > 0)) {
                if (!wantPdf) bsdfPdf = 0.0f;
                float3 c = thr * value * bsdfVal * miWeight(dr.pdf, bsdfPdf);
                out.shadow = true;
                out.shO = make_float4(dr.ref.x, dr.ref.y, dr.ref.z, kEps);
                out.shD = make_float4(dr.d.x, dr.d.y, dr.d.z, dr.dist * (1 - kShadowEps));
                out.shC = make_float4(c.x, c.y, c.z, __int_as_float((int)slot));
            }
        }
    }

    // ---- BSDF sampling (path.cpp:203-253) ----
    BSample bs; bs.pdf = 0; bs.eta = 1; bs.type = 0; bs.wo = f3(0.0f);
    bs.u3 = uExtra;   // EUsesSampler models (roughdielectric.cpp:405 bRec.sampler->next1D()): spare bits of the same Philox draw
    float3 bsdfWeight = mat.sample(v.wi, make_float2(u.z, u.w), bs);
    if (isZero(bsdfWeight)) { finish(); return; }
    if (bs.type != MTS_ENULL) flags |= PF_SCATTERED;   // path.cpp:213
    const float3 wo = v.sh.toWorld(bs.wo);
    if (R.strictNormals && dot(v.geoN, wo) * cosTheta(bs.wo) <= 0) { finish(); return; }

    thr *= bsdfWeight;
    eta *= bs.eta;
    flags &= ~(uint32_t)(PF_PREV_DELTA | PF_PREV_REFN_OK);
    if (bs.type & MTS_EDELTA) flags |= PF_PREV_DELTA;
    if (dot(wo, refN) >= 0) flags |= PF_PREV_REFN_OK;
    depth += 1;

    W.rayO[slot] = make_float4(v.p.x, v.p.y, v.p.z, kEps);
    W.rayD[slot] = make_float4(wo.x, wo.y, wo.z, kInf);
    W.thr[slot] = make_float4(thr.x, thr.y, thr.z, eta);
    W.L[slot] = make_float4(L.x, L.y, L.z, bs.pdf);
    W.meta[slot] = make_uint4(meta.x, meta.y, depth | (flags << 16), 0);
    out.next = true; out.done = false;
}

template <int CLASS>
DEVFN void shadePath(const DScene &S, const Wavefront &W, const RenderConst &R, uint32_t slot, ShadeOut &out) {
    out.next = false; out.shadow = false; out.done = true;
    PathVertex v; Surf su; int bsdf;
    if (!shadeVertex(S, W, R, slot, v, su, bsdf)) return;
    const Material<CLASS> mat(S, bsdf, su, v.wi);
    shadeScatter(S, W, R, slot, v, mat, out);
}

// Escaped rays (sort key 0) are shaded by their own small kernel: no hit record, no BSDF code, 40 registers, full
// occupancy -- after the first bounce three of four rays of an outdoor scene end here.
// path.cpp:136-143 (camera ray) / :238-248,257-264 (BSDF-sampled ray escaping to the envmap)
static __global__ void __launch_bounds__(256) k_shade_miss(const DScene S, const Wavefront W, const RenderConst R, int qsel) {
    const uint32_t total = W.ctr->qCount[qsel];
    const uint32_t count = W.numKeys > 1 ? min(W.keyCount[W.numKeys + 1], total) : total;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ((count + 255u) & ~255u); i += gridDim.x * blockDim.x) {   // whole CTAs per trip
        const bool valid = i < count;
        uint32_t slot = 0;
        if (valid) {
            slot = W.sorted[i];
            const uint4 meta = W.meta[slot];
            uint32_t depth = meta.z & 0xffffu;
            const uint32_t flags = meta.z >> 16;
            const bool first = depth == 1;
            // the reference bumps rRec.depth only after a valid hit (path.cpp:272-276); `depth` was already
            // incremented when this ray was queued, so undo that for the path-length statistic
            if (!first) depth -= 1;
            float4 lv = W.L[slot];
            if (S.hasEnv && !(R.hideEmitters && !(flags & PF_SCATTERED))) {   // path.cpp:140, 238-239
                const float3 d = f3(W.rayD[slot]);
                const float3 thr = f3(W.thr[slot]);
                float3 L = f3(lv);
                if (first) {
                    // Emitter::evalEnvironment(RayDifferential) of the camera ray: EWA lookup (envmap.cpp:380-410)
                    const float4 us = rand4(R.seed, meta.x, meta.y, 0);
                    const int x = (int)(meta.x % (uint32_t)S.filmW), y = (int)(meta.x / (uint32_t)S.filmW);
                    float3 co, cd, rxD, ryD;
                    cameraRayDiff(S.cam, make_float2(x + us.x, y + us.y), R.diffScale, co, cd, rxD, ryD);
                    L += thr * envEvalDiff(S, S.env, d, rxD, ryD);
                } else {
                    const float3 o = f3(W.rayO[slot]);
                    float nearT, farT;  // fillDirectSamplingRecord, envmap.cpp:358-374
                    if (bsphereIntersect(S.env, o, d, nearT, farT) && !(nearT > 0) && !(farT < 0)) {
                        float lumPdf = (flags & PF_PREV_DELTA) ? 0.0f : envPdfDirection(S.env, xfVec(S.env.toLocal, d)) * S.env.prob;
                        L += thr * envEval(S.env, d) * miWeight(lv.w, lumPdf);
                    }
                }
                lv = make_float4(L.x, L.y, L.z, 0);
            }
            W.L[slot] = make_float4(lv.x, lv.y, lv.z, 0);
            W.meta[slot] = make_uint4(meta.x, meta.y, depth | (flags << 16), 0);
        }
        uint32_t c = blockAppend(&W.ctr->doneCount, valid);
        if (valid) W.doneQ[c] = slot;
    }
}

// One instance per shade class; the class's paths are the contiguous range of the sorted queue whose keys lie in
// [classKey[CLASS], classKey[CLASS + 1]).
template <int CLASS> struct ShadeLaunch;
template <> struct ShadeLaunch<SHADE_LAMBERT> { static constexpr int kThreads = 128, kMinBlocks = 5; };
template <> struct ShadeLaunch<SHADE_LEAF>    { static constexpr int kThreads = 128, kMinBlocks = 4; };
template <> struct ShadeLaunch<SHADE_GENERIC> { static constexpr int kThreads = 128, kMinBlocks = 4; };

DEVFN uint32_t sortedBegin(const Wavefront &W, uint32_t key, uint32_t count) {   // first queue position with sort key >= key
    return key < W.numKeys ? min(W.keyCount[W.numKeys + key], count) : count;
}

template <int CLASS>
static __global__ void __launch_bounds__(ShadeLaunch<CLASS>::kThreads, ShadeLaunch<CLASS>::kMinBlocks)
k_shade(const DScene S, const Wavefront W, const RenderConst R, int qsel) {
    const uint32_t count = W.ctr->qCount[qsel];
    const uint32_t first = sortedBegin(W, W.classKey[CLASS], count), last = sortedBegin(W, W.classKey[CLASS + 1], count);
    const uint32_t rounded = first + ((last - first + 31u) & ~31u);
    for (uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x; i < rounded; i += gridDim.x * blockDim.x) {
        ShadeOut out; out.next = out.shadow = out.done = false;
        uint32_t slot = 0;
        if (i < last) {
            slot = W.sorted[i];
            shadePath<CLASS>(S, W, R, slot, out);
        }
        uint32_t a = warpAppend(&W.ctr->qCount[qsel ^ 1], out.next);
        if (out.next) W.queue[qsel ^ 1][a] = slot;
        uint32_t b = warpAppend(&W.ctr->shadowCount, out.shadow);
        if (out.shadow) { W.shO[b] = out.shO; W.shD[b] = out.shD; W.shC[b] = out.shC; }
        uint32_t c = warpAppend(&W.ctr->doneCount, out.done);
        if (out.done) W.doneQ[c] = slot;
    }
}

// ---- staged shade of the leaf class -------------------------------------------------------------------------------
// k_shade<SHADE_LEAF> as one kernel is half a megabyte of machine code (hit record, ray differentials, MIP / EWA texture
// filtering, bump frames, emitter sampling, nine BSDF models): the warps of an SM sit at different places of a program far
// larger than the instruction caches and stall on instruction fetch (profiles/r2b: stall_no_instruction 14 cycles per
// issue at the camera-ray bounce).  The staged form runs the iteration as two kernels over the same range of the sorted
// queue, handing over one "surface record" per queue position (coalesced float4 rows, kSurfRows x capacity):
//   k_shade_surface  hit record, emission + MIS, Russian roulette, depth tests, uv partials, textures and perturbed frame;
//                    EWA texture walks of camera-ray hits are queued, not done
//   k_shade_ewa      the queued EWA walks, one thread pair per lookup (camera-ray bounce only)
//   k_shade_scatter  emitter sampling + BSDF eval/pdf (shadow ray), BSDF sampling (next ray)
// Same arithmetic on the same fp32 values as the fused kernel.
constexpr int kSurfRows = 10;
#ifndef SURFACE_MINB
#define SURFACE_MINB 4
#endif
#ifndef SCATTER_MINB
#define SCATTER_MINB 6
#endif
enum { SR_BUMP = 1u << 24, SR_FLIPPED = 1u << 25, SR_ALIVE = 1u << 26, SR_NFLIP = 1u << 27,
       SR_SMOOTH = 1u << 28,    // the root BSDF has smooth lobes (light sampling applies, path.cpp:172)
       SR_ONESIDED = 1u << 29,  // no transmission / back-side lobes: refN = shading normal (records.inl:160-164)
       SR_NODE_MASK = 0xffffffu };

template <int CLASS>
static __global__ void __launch_bounds__(128, SURFACE_MINB) k_shade_surface(const DScene S, const Wavefront W, const RenderConst R, int qsel) {
    const uint32_t count = W.ctr->qCount[qsel];
    const uint32_t first = sortedBegin(W, W.classKey[SHADE_LEAF], count), last = sortedBegin(W, W.classKey[SHADE_LEAF + 1], count);
    const uint32_t rounded = first + ((last - first + 31u) & ~31u);
    const size_t cap = W.capacity;
    for (uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x; i < rounded; i += gridDim.x * blockDim.x) {
        bool done = false;
        uint32_t slot = 0;
        if (i < last) {
            slot = W.sorted[i];
            PathVertex v; Surf su; int bsdf;
            if (!shadeVertex(S, W, R, slot, v, su, bsdf)) {
                done = true;
                W.rec[i] = make_float4(0, 0, 0, __uint_as_float(0u));
            } else {
                EwaDefer ewa; ewa.pending = false;
                const Material<SHADE_LEAF> mat(S, bsdf, su, v.wi, W.ewaJobs ? &ewa : nullptr);
                if (ewa.pending) {   // the walk runs in k_shade_ewa, which fills c0 / c1 of this record
                    const uint32_t j = warpAppend(&W.ctr->ewaCount, true);   // one atomic per warp (the lanes that queue a walk)
                    W.ewaJobs[j] = make_float4(__uint_as_float(i), __uint_as_float((uint32_t)ewa.texA | ((uint32_t)(ewa.texB & 0xffff) << 16)),
                                               __uint_as_float((uint32_t)ewa.ilevel | ((uint32_t)ewa.slotA << 8)), ewa.a);
                    W.ewaJobs[cap + j] = make_float4(ewa.uv.x, ewa.uv.y, ewa.A, ewa.B);
                    W.ewaC[j] = ewa.C;
                }
                uint32_t info = (uint32_t)mat.leafIndex() | SR_ALIVE | (mat.bump ? SR_BUMP : 0u) | (mat.flipped ? SR_FLIPPED : 0u);
                if (v.lobes & MTS_ESMOOTH) info |= SR_SMOOTH;
                if ((v.lobes & (MTS_ETRANSMISSION | MTS_EBACKSIDE)) == 0) info |= SR_ONESIDED;
                if (mat.bump && dot(cross(mat.pf.n, mat.pf.s), mat.pf.t) < 0) info |= SR_NFLIP;   // bumpmap.cpp:156-157 flips n after t was built
                float4 *r = W.rec + i;
                r[0] = make_float4(v.p.x, v.p.y, v.p.z, __uint_as_float(info));
                r[cap] = make_float4(v.sh.n.x, v.sh.n.y, v.sh.n.z, v.geoN.x);
                r[2 * cap] = make_float4(v.sh.s.x, v.sh.s.y, v.sh.s.z, v.geoN.y);
                r[3 * cap] = make_float4(v.wi.x, v.wi.y, v.wi.z, v.geoN.z);
                r[4 * cap] = make_float4(mat.lt.c0.x, mat.lt.c0.y, mat.lt.c0.z, mat.wiLeaf.x);
                r[5 * cap] = make_float4(mat.lt.c1.x, mat.lt.c1.y, mat.lt.c1.z, mat.wiLeaf.y);
                r[6 * cap] = make_float4(mat.lt.c2.x, mat.lt.c2.y, mat.lt.c2.z, mat.wiLeaf.z);
                if (mat.bump) {
                    r[7 * cap] = make_float4(mat.pf.n.x, mat.pf.n.y, mat.pf.n.z, 0);
                    r[8 * cap] = make_float4(mat.pf.s.x, mat.pf.s.y, mat.pf.s.z, 0);
                }
                if (mat.n->type == MTS_BSDF_ROUGHDIELECTRIC) r[9 * cap] = make_float4(mat.lt.c3.x, mat.lt.c3.y, mat.lt.c3.z, 0);
                // path state the first half changed (the scatter kernel reads it back by slot)
                if (v.touched & 1u) W.L[slot] = make_float4(v.L.x, v.L.y, v.L.z, 0);
                if (v.touched & 2u) W.thr[slot] = make_float4(v.thr.x, v.thr.y, v.thr.z, v.eta);
                if (v.touched & 4u) W.meta[slot] = make_uint4(v.meta.x, v.meta.y, v.depth | (v.flags << 16), 0);
            }
        }
        uint32_t c = warpAppend(&W.ctr->doneCount, done);
        if (done) W.doneQ[c] = slot;
    }
}

// The deferred EWA walks: one thread per lookup, both pyramid levels (evalEWA(ilevel) * (1 - a) + evalEWA(ilevel + 1) * a,
// mipmap.h:712-713), scaled (ScalingTexture) and stored into the c0 / c1 rows of the surface record.  Footprints vary by
// more than an order of magnitude between neighbouring paths (unsorted, a warp ran with 10 of 32 lanes busy), so the
// lookups are first binned by the number of texels their two bounding boxes hold (k_ewa_bin: log2 bins, counting sort
// like the material sort) and a warp walks 32 lookups of one size class.
constexpr int kEwaBins = 16;
#ifndef EWA_MINB
#define EWA_MINB 4
#endif
DEVFN uint32_t ewaJobSize(const DScene &S, const float4 j0, const float4 j1, float C) {
    const uint32_t tex = __float_as_uint(j0.y), li = __float_as_uint(j0.z);
    const DTexture &ta = S.textures[tex & 0xffffu];
    const int ilevel = (int)(li & 0xffu);
    const EwaLevel L0 = ewaSetup(ta, ilevel, make_float2(j1.x, j1.y), j1.z, j1.w, C), L1 = ewaSetup(ta, ilevel + 1, make_float2(j1.x, j1.y), j1.z, j1.w, C);
    return (uint32_t)(ewaRows(L0) * max(L0.u1 - L0.u0 + 1, 0) + ewaRows(L1) * max(L1.u1 - L1.u0 + 1, 0));
}
static __global__ void __launch_bounds__(256) k_ewa_bin(const DScene S, const Wavefront W) {
    const uint32_t count = W.ctr->ewaCount;
    const size_t cap = W.capacity;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < ((count + 31u) & ~31u); j += gridDim.x * blockDim.x) {
        uint32_t key = kNoKey;
        if (j < count) {
            const uint32_t taps = ewaJobSize(S, W.ewaJobs[j], W.ewaJobs[cap + j], W.ewaC[j]);
            key = min((uint32_t)kEwaBins - 1u, 32u - (uint32_t)__clz(taps));   // 0 taps -> bin 0, else 1 + floor(log2)
            W.keys[j] = key;    // (the material sort is done with `keys` by the time the shade kernels run)
        }
        keyedAppend(W.ctr->ewaBin, key);
    }
}
static __global__ void k_ewa_offsets(const Wavefront W) {   // one warp: exclusive scan of the kEwaBins counts
    uint32_t *cnt = W.ctr->ewaBin, *off = cnt + kEwaBins, *cur = cnt + 2 * kEwaBins;
    const uint32_t k = threadIdx.x;
    uint32_t v = k < kEwaBins ? cnt[k] : 0, incl = v;
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)threadIdx.x >= o) incl += t; }
    if (k < kEwaBins) { off[k] = incl - v; cur[k] = 0; cnt[k] = 0; }
}
static __global__ void __launch_bounds__(256) k_ewa_scatter(const Wavefront W) {
    const uint32_t count = W.ctr->ewaCount;
    const uint32_t *off = W.ctr->ewaBin + kEwaBins;
    uint32_t *cur = W.ctr->ewaBin + 2 * kEwaBins;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < ((count + 31u) & ~31u); j += gridDim.x * blockDim.x) {
        const uint32_t key = j < count ? W.keys[j] : kNoKey;
        const uint32_t pos = keyedAppend(cur, key);
        // largest size class first: the long walks start while the machine is full
        if (key != kNoKey) W.ewaOrder[count - 1u - (off[key] + pos)] = j;
    }
}
template <int CLASS>
static __global__ void __launch_bounds__(256, EWA_MINB) k_shade_ewa(const DScene S, const Wavefront W) {
    const uint32_t count = W.ctr->ewaCount;
    const size_t cap = W.capacity;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < count; t += gridDim.x * blockDim.x) {
        const uint32_t j = W.ewaOrder[t];
        const float4 j0 = W.ewaJobs[j], j1 = W.ewaJobs[cap + j];
        const float C = W.ewaC[j], a = j0.w;
        const uint32_t tex = __float_as_uint(j0.y), li = __float_as_uint(j0.z);
        const int texA = (int)(tex & 0xffffu), texB = (int)(tex >> 16), ilevel = (int)(li & 0xffu);
        const bool pair = texB != 0xffff;
        const float2 uv = make_float2(j1.x, j1.y);
        float3 va, vb = f3(0.0f);
        if (pair) mipEWA2Pair(S.textures[texA], S.textures[texB], ilevel, uv, j1.z, j1.w, C, a, va, vb);
        else va = mipEWA2(S.textures[texA], ilevel, uv, j1.z, j1.w, C, a);
        const uint32_t pos = __float_as_uint(j0.x), slotA = (li >> 8) & 1u;
        float4 *r = W.rec + pos;
        const MtsBsdfNode &n = S.nodes[__float_as_uint(r[0].w) & SR_NODE_MASK];
        const MtsTexRef &ra = slotA ? n.t1 : n.t0;
        const float3 ca = va * f3(ra.scale[0], ra.scale[1], ra.scale[2]);
        float4 *rowA = r + (4 + slotA) * cap;
        float4 cur = *rowA;
        *rowA = make_float4(ca.x, ca.y, ca.z, cur.w);
        if (pair) {
            const float3 cb = vb * f3(n.t1.scale[0], n.t1.scale[1], n.t1.scale[2]);
            float4 *rowB = r + 5 * cap;
            cur = *rowB;
            *rowB = make_float4(cb.x, cb.y, cb.z, cur.w);
        }
    }
}

template <int CLASS>
static __global__ void __launch_bounds__(128, SCATTER_MINB) k_shade_scatter(const DScene S, const Wavefront W, const RenderConst R, int qsel) {
    const uint32_t count = W.ctr->qCount[qsel];
    const uint32_t first = sortedBegin(W, W.classKey[SHADE_LEAF], count), last = sortedBegin(W, W.classKey[SHADE_LEAF + 1], count);
    const uint32_t rounded = first + ((last - first + 127u) & ~127u);   // whole 128-thread CTAs per trip (blockAppend)
    const size_t cap = W.capacity;
    for (uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x; i < rounded; i += gridDim.x * blockDim.x) {
        ShadeOut out; out.next = out.shadow = out.done = false;
        uint32_t slot = 0;
        if (i < last) {
            const float4 *r = W.rec + i;
            const float4 r0 = r[0];
            const uint32_t info = __float_as_uint(r0.w);
            if (info & SR_ALIVE) {
                slot = W.sorted[i];
                const float4 r1 = r[cap], r2 = r[2 * cap], r3 = r[3 * cap], r4 = r[4 * cap], r5 = r[5 * cap], r6 = r[6 * cap];
                const float4 tv = W.thr[slot], lv = W.L[slot];
                const uint4 meta = W.meta[slot];
                PathVertex v;
                v.p = f3(r0); v.geoN = f3(r1.w, r2.w, r3.w); v.wi = f3(r3);
                v.sh.n = f3(r1); v.sh.s = f3(r2); v.sh.t = cross(v.sh.n, v.sh.s);
                v.thr = f3(tv); v.eta = tv.w; v.L = f3(lv);
                v.depth = meta.z & 0xffffu; v.flags = meta.z >> 16; v.meta = meta;
                const int leaf = (int)(info & SR_NODE_MASK);
                LeafTex lt; lt.c0 = f3(r4); lt.c1 = f3(r5); lt.c2 = f3(r6); lt.c3 = f3(0.0f);
                if (S.nodes[leaf].type == MTS_BSDF_ROUGHDIELECTRIC) lt.c3 = f3(r[9 * cap]);
                Frame pf = v.sh;
                if (info & SR_BUMP) {
                    pf.n = f3(r[7 * cap]); pf.s = f3(r[8 * cap]);
                    pf.t = cross((info & SR_NFLIP) ? -pf.n : pf.n, pf.s);
                }
                // shadeScatter only asks the two questions the flags answer
                v.lobes = ((info & SR_SMOOTH) ? MTS_EDIFFUSE_REFLECTION : 0u) | ((info & SR_ONESIDED) ? 0u : MTS_EBACKSIDE);
                const Material<SHADE_LEAF> mat(S, leaf, lt, v.sh, pf, f3(r4.w, r5.w, r6.w), (info & SR_BUMP) != 0, (info & SR_FLIPPED) != 0);
                out.done = true;
                shadeScatter(S, W, R, slot, v, mat, out);
            }
        }
        // one atomicAdd per CTA and queue (measured against one per warp: k_shade_scatter -0.5 ms per 16.6M-path wavefront of cfg3)
        uint32_t a = blockAppend(&W.ctr->qCount[qsel ^ 1], out.next);
        uint32_t b = blockAppend(&W.ctr->shadowCount, out.shadow);
        uint32_t c = blockAppend(&W.ctr->doneCount, out.done);
        if (out.next) W.queue[qsel ^ 1][a] = slot;
        if (out.shadow) { W.shO[b] = out.shO; W.shD[b] = out.shD; W.shC[b] = out.shC; }
        if (out.done) W.doneQ[c] = slot;
    }
}

// ---- splat: ImageBlock::put(pos, spec, alpha) (include/mitsuba/render/imageblock.h:124-204) ----
DEVFN void filmAdd(float4 *dst, float4 v) {
#if __CUDA_ARCH__ >= 900
    atomicAdd(dst, v);  // one 16-byte vector reduction per tap
#else
    atomicAdd(&dst->x, v.x); atomicAdd(&dst->y, v.y); atomicAdd(&dst->z, v.z); atomicAdd(&dst->w, v.w);
#endif
}
DEVFN void splatSample(const DScene &S, float2 samplePos, float3 value, float alpha = 1.0f) {
    // reject invalid samples like put() does (:147-151)
    if (!(isfinite(value.x) && isfinite(value.y) && isfinite(value.z)) || value.x < 0 || value.y < 0 || value.z < 0) return;
    const float r = S.filterRadius;
    const float scaleFactor = MTS_FILTER_RESOLUTION / r;
    const float px = samplePos.x - 0.5f, py = samplePos.y - 0.5f;
    int minx = max((int)ceilf(px - r), 0), miny = max((int)ceilf(py - r), 0);
    int maxx = min((int)floorf(px + r), S.filmW - 1), maxy = min((int)floorf(py + r), S.filmH - 1);
    for (int y = miny; y <= maxy; ++y) {
        float wy = S.filterTable[min((int)fabsf((y - py) * scaleFactor), MTS_FILTER_RESOLUTION)];
        for (int x = minx; x <= maxx; ++x) {
            float w = S.filterTable[min((int)fabsf((x - px) * scaleFactor), MTS_FILTER_RESOLUTION)] * wy;
            if (w == 0.0f) continue;
            filmAdd(S.film + (size_t)y * S.filmW + x, make_float4(w * value.x, w * value.y, w * value.z, w));
            if (S.filmAlpha && alpha != 0.0f) atomicAdd(S.filmAlpha + (size_t)y * S.filmW + x, w * alpha);   // imageblock.h:176-186, alpha channel
        }
    }
}
static __global__ void __launch_bounds__(256) k_splat(const DScene S, const Wavefront W, const RenderConst R) {
    const uint32_t count = W.ctr->doneCount;
    unsigned long long lenSum = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        uint32_t slot = W.doneQ[i];
        float4 l = W.L[slot];
        uint4 meta = W.meta[slot];
        float4 u = rand4(R.seed, meta.x, meta.y, 0);
        int x = (int)(meta.x % (uint32_t)S.filmW), y = (int)(meta.x / (uint32_t)S.filmW);
        splatSample(S, make_float2(x + u.x, y + u.y), f3(l), ((meta.z >> 16) & PF_CAMERA_HIT) ? 1.0f : 0.0f);
        lenSum += meta.z & 0xffffu;
    }
    for (int o = 16; o > 0; o >>= 1) lenSum += __shfl_xor_sync(0xffffffffu, lenSum, o);
    if ((threadIdx.x & 31) == 0 && lenSum) atomicAdd(&W.ctr->pathLenSum, lenSum);
}

// Bounce bookkeeping on the device (the host never reads a counter inside the bounce loop): add the rays of the
// bounce that is about to run (closest hit) and of the one that just ran (shadow) to the totals, reset the queues.
static __global__ void k_begin_bounce(Counters *c, int cur) {
    c->normalRays += c->qCount[cur]; c->shadowRays += c->shadowCount;
    c->qCount[cur ^ 1] = 0; c->shadowCount = 0; c->doneCount = 0; c->ewaCount = 0; c->cursor[0] = c->cursor[1] = 0;
}
static __global__ void k_begin_wavefront(Counters *c) {
    c->shadowRays += c->shadowCount;
    c->qCount[0] = c->qCount[1] = 0; c->shadowCount = 0; c->doneCount = 0; c->ewaCount = 0; c->cursor[0] = c->cursor[1] = 0;
}
// Safety net of unbounded renders (maxDepth = -1 with Russian roulette pushed out of reach): the paths still alive
// when the host gives up are finished with what they have gathered so far and splatted, never dropped.
static __global__ void __launch_bounds__(256) k_abandon(const Wavefront W, int qsel) {
    const uint32_t count = W.ctr->qCount[qsel];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ((count + 31u) & ~31u); i += gridDim.x * blockDim.x) {
        const bool valid = i < count;
        const uint32_t slot = valid ? W.queue[qsel][i] : 0u;
        uint32_t c = warpAppend(&W.ctr->doneCount, valid);
        if (valid) W.doneQ[c] = slot;
    }
}

} // namespace mts
