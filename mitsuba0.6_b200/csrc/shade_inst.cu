// shade_inst.cu -- one shade class per translation unit (see shade_launch.h):
//   nvcc -DSHADE_CLASS={0,1,2} -DSHADE_PART={0,1}   part 0 = k_shade<CLASS>, part 1 = the BSDF test entry points;
//   class 1 (leaf) also has part 2 = k_shade_surface and part 3 = k_shade_scatter, the staged form of its k_shade
#include <algorithm>
#include "shade_launch.h"

#ifndef SHADE_CLASS
#error "compile with -DSHADE_CLASS=0|1|2 -DSHADE_PART=0|1"
#endif

namespace mts {

#if SHADE_PART == 0

template <> void launchShade<SHADE_CLASS>(const DScene &S, const Wavefront &W, const RenderConst &R, int qsel, int numSMs, cudaStream_t st) {
    k_shade<SHADE_CLASS><<<numSMs * ShadeLaunch<SHADE_CLASS>::kMinBlocks, ShadeLaunch<SHADE_CLASS>::kThreads, 0, st>>>(S, W, R, qsel);
}

#elif SHADE_PART == 2

void launchSurface(const DScene &S, const Wavefront &W, const RenderConst &R, int qsel, int numSMs, cudaStream_t st) {
    k_shade_surface<SHADE_CLASS><<<numSMs * SURFACE_MINB, 128, 0, st>>>(S, W, R, qsel);
}

#elif SHADE_PART == 4

void launchEwa(const DScene &S, const Wavefront &W, int numSMs, cudaStream_t st) {
    k_ewa_bin<<<numSMs * 8, 256, 0, st>>>(S, W);
    k_ewa_offsets<<<1, 32, 0, st>>>(W);
    k_ewa_scatter<<<numSMs * 8, 256, 0, st>>>(W);
    k_shade_ewa<SHADE_CLASS><<<numSMs * EWA_MINB, 256, 0, st>>>(S, W);
}

#elif SHADE_PART == 3

void launchScatter(const DScene &S, const Wavefront &W, const RenderConst &R, int qsel, int numSMs, cudaStream_t st) {
    k_shade_scatter<SHADE_CLASS><<<numSMs * SCATTER_MINB, 128, 0, st>>>(S, W, R, qsel);
}

#else

DEVFN Surf surfFromArray(const float *s) {
    Surf su;
    su.tc = texCtx(make_float2(s[0], s[1]));
    if (s[15] != 0.0f) { su.tc.duv = make_float4(s[2], s[3], s[7], s[11]); su.tc.filtered = true; }   // Intersection::hasUVPartials
    su.dpdu = f3(s[4], s[5], s[6]); su.dpdv = f3(s[8], s[9], s[10]);
    float3 n = f3(s[12], s[13], s[14]);
    su.geoN = n;
    su.sh = shadingFrame(n, su.dpdu);
    su.bumpNode = -1;
    return su;
}
// The per-function BSDF entry points run the same Material<CLASS> code as k_shade<CLASS> (class of the root BSDF).
template <int CLASS>
static __global__ void k_hook_bsdf_eval_pdf(const DScene S, int bsdf, int64_t n, const float *surf, const float *wi, const float *wo,
                                     int measure, float *outV, float *outP) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Surf su = surfFromArray(surf + 16 * i);
    float3 a = f3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), b = f3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    const Material<CLASS> mat(S, bsdf, su, a);
    float pdf;
    float3 v = mat.evalPdf(a, b, measure, PDF_ALWAYS, pdf);
    outV[3 * i] = v.x; outV[3 * i + 1] = v.y; outV[3 * i + 2] = v.z;
    outP[i] = pdf;
}
template <int CLASS>
static __global__ void k_hook_bsdf_sample(const DScene S, int bsdf, int64_t n, const float *surf, const float *wi, const float *u, float *out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Surf su = surfFromArray(surf + 16 * i);
    float3 a = f3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
    BSample bs; bs.wo = f3(0.0f); bs.pdf = 0; bs.eta = 1; bs.type = 0;
    MTS_PROBE_U3(__float_as_uint(u[2 * i]), __float_as_uint(u[2 * i + 1]), bs.u3);
    const Material<CLASS> mat(S, bsdf, su, a);
    float3 w = mat.sample(a, make_float2(u[2 * i], u[2 * i + 1]), bs);
    float *q = out + 12 * i;
    if (isZero(w)) { for (int k = 0; k < 12; ++k) q[k] = 0; return; }
    q[0] = w.x; q[1] = w.y; q[2] = w.z; q[3] = bs.pdf; q[4] = bs.wo.x; q[5] = bs.wo.y; q[6] = bs.wo.z;
    q[7] = bs.eta; q[8] = (float)bs.type; q[9] = q[10] = q[11] = 0;
}

static int gridFor(int64_t n, int bs = 128) { return (int)std::max<int64_t>(1, (n + bs - 1) / bs); }

template <> void launchBsdfEvalPdf<SHADE_CLASS>(const DScene &S, int bsdf, int64_t n, const float *surf, const float *wi, const float *wo,
                                                int measure, float *outV, float *outP, cudaStream_t st) {
    k_hook_bsdf_eval_pdf<SHADE_CLASS><<<gridFor(n), 128, 0, st>>>(S, bsdf, n, surf, wi, wo, measure, outV, outP);
}
template <> void launchBsdfSample<SHADE_CLASS>(const DScene &S, int bsdf, int64_t n, const float *surf, const float *wi, const float *u,
                                               float *out, cudaStream_t st) {
    k_hook_bsdf_sample<SHADE_CLASS><<<gridFor(n), 128, 0, st>>>(S, bsdf, n, surf, wi, u, out);
}

#endif

} // namespace mts
