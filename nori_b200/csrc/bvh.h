// bvh.h — device BVH builder interface (bvh.cu). Stands in for Accel::build (include/nori/accel.h:45).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/nori_gpu.h"

namespace ng {

// float4 slots per leaf triangle: 3 = (p0|gid, p1, p2), 48 B, three 128-bit loads; 4 = padded to 64 B, two 256-bit loads
// (experiment, -DNORI_TRI_STRIDE=4: profiles/r01_trace.md)
#ifndef NORI_TRI_STRIDE
#define NORI_TRI_STRIDE 3
#endif
constexpr int kTriStride = NORI_TRI_STRIDE;

struct BvhParams {
    int plocRadius = 16;       // PLOC search radius in Morton order
    int maxLeaf = 4;           // triangles per leaf (<= 8: 3 bits in the leaf link)
    float costInner = 1.0f;    // SAH: one node visit (two child boxes)
    float costTri = 1.0f;      // SAH: one Moeller-Trumbore test
    int rotationPasses = 0;    // bottom-up tree-rotation passes over the PLOC tree before the leaf collapse (k_rotate)
    float inflate = 4e-6f;     // child boxes grow by inflate * max(scene diagonal, max |coordinate|): ~60 float ulps
};

struct BvhDevice {
    float4 *nodes = nullptr;       // 4 x float4 per node (traverse.cuh)
    float4 *leafTris = nullptr;    // 3 x float4 per triangle, leaf order
    uint32_t nNodes = 0, nTris = 0;
    float sceneLo[3] = {0, 0, 0}, sceneHi[3] = {0, 0, 0};
    // optional second copy of the nodes, 32 B each: child boxes on a 15-bit grid over the scene box (quantiseNodes)
    uint4 *qnodes = nullptr;
    float qBase[3] = {0, 0, 0}, qExt[3] = {1, 1, 1};   // coordinate = qBase + m * qExt, m in [1, 2) (traverse_persistent.cuh)
    // optional third copy: the tree collapsed to 4-wide nodes, 128 B each (wideNodes)
    float4 *wnodes = nullptr;
    uint32_t nWide = 0, wideDepth = 0;
};

/// Builds the tree for triangles tris[0..nTris) over vpos. Frees nothing it did not allocate;
/// the caller owns out.nodes / out.leafTris (cudaFree).
bool buildBvh(const float4 *d_vpos, const uint4 *d_tris, uint32_t nTris, const BvhParams &p, cudaStream_t stream,
              BvhDevice &out, NoriGpuAccelStats &stats, std::string &err);

/// NORI_ACCEL_QUANTISED: writes bvh.qnodes from bvh.nodes (same topology, same links). Per node
///   q0 = (L.x | L.y | L.z | linkL), q1 = (R.x | R.y | R.z | linkR), each axis word = lo16 | hi16 << 16,
///   a 16-bit field = 0x8000 | q with q the 15-bit grid index of the bound, rounded OUTWARDS (the float boxes already
///   carry the builder's inflation, ~0.1 cell, which covers the 0.004-cell error of the decode),
/// so that the two bytes drop into a float mantissa (0x3F.. .. .. 00 = 1 + q / 32768) with one PRMT.
bool quantiseNodes(BvhDevice &bvh, float inflate, cudaStream_t stream, std::string &err);

/// NORI_ACCEL_WIDE4: writes bvh.wnodes, the binary tree collapsed top-down into 4-wide nodes (bvh.cu). Leaves are unchanged.
constexpr int kWideEmpty = 0x7ffffffe;       // link of an unused child slot
constexpr int kWideStackSize = 128;          // traversal stack entries of the wide layout (three pushes per visit)
bool wideNodes(BvhDevice &bvh, cudaStream_t stream, std::string &err);

} // namespace ng
