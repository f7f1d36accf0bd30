// devmem.h -- stream-ordered device allocations for libmtscuda.
//
// The plugin creates and destroys one context per render (like every Mitsuba integrator allocates its state in
// render()), so the library allocates and frees several GB of scene and wavefront state per frame.  cudaMalloc /
// cudaFree synchronise the device and go to the OS every time (the 32 Mi-slot wavefront alone cost ~25 ms per render
// at round 1); cudaMallocAsync takes the blocks from the device's default memory pool instead, in stream order, and
// with an unlimited release threshold the pool keeps them across contexts of the same process.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mts {

inline void devPoolInit(int device) {
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t keep = UINT64_MAX;   // never hand memory back to the OS between renders
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    } else cudaGetLastError();
}
// MTS_POOL_TRIM=1 (memory-constrained hosts): give the cached blocks back when a context is destroyed
inline void devPoolTrim(int device) {
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
}
template <class T> inline cudaError_t devAlloc(T **p, size_t bytes, cudaStream_t st) {
    return cudaMallocAsync((void **)p, bytes ? bytes : 16, st);
}
inline void devFree(void *p, cudaStream_t st) { if (p) cudaFreeAsync(p, st); }

} // namespace mts
