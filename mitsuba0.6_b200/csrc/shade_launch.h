// shade_launch.h -- host-side launchers of the per-shade-class kernels.  Every class (and, within a class, the wavefront
// kernel and the per-function test entry points) is its own translation unit (shade_inst.cu compiled with
// -DSHADE_CLASS=c -DSHADE_PART=p), so the expensive BSDF code is compiled in parallel and only once per kernel.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "wavefront.cuh"

namespace mts {

// k_shade<CLASS> over the class's range of the sorted queue (persistent grid sized from the SM count)
template <int CLASS> void launchShade(const DScene &S, const Wavefront &W, const RenderConst &R, int qsel, int numSMs, cudaStream_t st);
// the staged form of the leaf class: k_shade_surface (part 2) then k_shade_scatter (part 3), each its own translation unit
void launchSurface(const DScene &S, const Wavefront &W, const RenderConst &R, int qsel, int numSMs, cudaStream_t st);
void launchEwa(const DScene &S, const Wavefront &W, int numSMs, cudaStream_t st);
void launchScatter(const DScene &S, const Wavefront &W, const RenderConst &R, int qsel, int numSMs, cudaStream_t st);
// mtscuda_bsdf_eval_pdf / mtscuda_bsdf_sample through Material<CLASS> (all pointers are device pointers)
template <int CLASS> void launchBsdfEvalPdf(const DScene &S, int bsdf, int64_t n, const float *surf, const float *wi, const float *wo,
                                            int measure, float *outV, float *outP, cudaStream_t st);
template <int CLASS> void launchBsdfSample(const DScene &S, int bsdf, int64_t n, const float *surf, const float *wi, const float *u,
                                           float *out, cudaStream_t st);

} // namespace mts
