// bvh_build.h -- interface of the GPU BVH builder (bvh_build.cu).
#pragma once
#include <string>
#include <vector>
#include "scene.cuh"

namespace mts {

struct PrimSource { int type; int shape; uint32_t prim; uint32_t triOffset; };  // type: 0 tri, 1 rect, 2 sphere

struct BvhMeshRange { uint32_t triOffset, ntris, primOffset; int shape; };
struct BvhAnalytic { uint32_t primOffset; int shape, type; float lo[3], hi[3]; };

struct BvhBuildInput {
    int numPrims = 0;
    const float *positions = nullptr;   // device, global vertex array
    const uint32_t *indices = nullptr;  // device, global (already offset) triangle indices
    std::vector<BvhMeshRange> meshes;
    std::vector<BvhAnalytic> analytic;
};

struct BvhBuildOutput {
    WideNode *nodes = nullptr;    // device (cudaMallocAsync on `stream`); caller frees with cudaFreeAsync
    PrimRecord *prims = nullptr;  // device, in wide-node leaf order
    int numNodes = 0;
    int buildNodes = 0;
    int launches = 0;
    int depth = 0;                // levels of the wide tree = upper bound of the traversal stack depth
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
};

bool buildBvh(const BvhBuildInput &in, BvhBuildOutput &out, cudaStream_t stream, std::string &err);

} // namespace mts
