// bvh.h — device BVH builder interface (bvh.cu). Stands in for Accel::build (include/nori/accel.h:45).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/nori_gpu.h"

namespace ng {

struct BvhParams {
    int plocRadius = 16;       // PLOC search radius in Morton order
    int maxLeaf = 4;           // triangles per leaf (<= 8: 3 bits in the leaf link)
    float costInner = 1.0f;    // SAH: one node visit (two child boxes)
    float costTri = 1.0f;      // SAH: one Moeller-Trumbore test
    float inflate = 4e-6f;     // child boxes grow by inflate * max(scene diagonal, max |coordinate|): ~60 float ulps
};

struct BvhDevice {
    float4 *nodes = nullptr;       // 4 x float4 per node (traverse.cuh)
    float4 *leafTris = nullptr;    // 3 x float4 per triangle, leaf order
    uint32_t nNodes = 0, nTris = 0;
    float sceneLo[3] = {0, 0, 0}, sceneHi[3] = {0, 0, 0};
};

/// Builds the tree for triangles tris[0..nTris) over vpos. Frees nothing it did not allocate;
/// the caller owns out.nodes / out.leafTris (cudaFree).
bool buildBvh(const float4 *d_vpos, const uint4 *d_tris, uint32_t nTris, const BvhParams &p, cudaStream_t stream,
              BvhDevice &out, NoriGpuAccelStats &stats, std::string &err);

} // namespace ng
