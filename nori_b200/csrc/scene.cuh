// scene.cuh — device-resident scene: the flattened form of Scene/Mesh/BSDF/Emitter/Camera
// (reference include/nori/scene.h:75-127, mesh.h:172-175) as the kernels see it.
#pragma once
#include "nmath.cuh"

namespace ng {

/// Per-mesh record (Mesh::getBSDF / getEmitter / m_N.size() / m_UV.size()).
struct MeshInfo {
    int bsdfType;
    float albedo[3];
    float alpha, intIOR, extIOR, ks;
    int emitter;                 // index into DeviceScene::emitters, -1: none (Mesh::isEmitter)
    uint32_t hasN, hasUV;
    uint32_t triOffset, nTris;   // global triangle id range of this mesh (concatenation order)
    uint32_t vertOffset, nVerts;
    uint32_t pad;
};

/// `area` emitter on a mesh: radiance + triangle-area CDF (DiscretePDF, dpdf.h).
struct EmitterInfo {
    int mesh;
    float radiance[3];
    float area;                  // DiscretePDF::normalize() return value = sum of triangle areas
    uint32_t cdfOffset;          // into DeviceScene::emitterCdf: nTris + 1 entries
    uint32_t nTris, triOffset;
};

/// PerspectiveCamera (src/perspective.cpp:53-109)
struct Camera {
    float s2c[16], c2w[16];
    float invW, invH, nearClip, farClip;
    int width, height;
    V3 origin;                   // cameraToWorld * (0,0,0), constant over samples (perspective.cpp:100)
};

/// ImageBlock filter state (src/block.cpp:31-39)
struct FilterTab {
    float radius, lookupFactor;
    int border;
    float table[33];
};

struct DeviceScene {
    // geometry, all meshes concatenated; vertex arrays padded to 16 B for vector loads
    const float4 *vpos;          // nVertsTotal: (x, y, z, 0)
    const float4 *vnrm;          // nVertsTotal: (nx, ny, nz, 0) (zero where the mesh has no normals)
    const float2 *vuv;           // nVertsTotal
    const uint4 *tris;           // nTris: (i0, i1, i2, mesh) with GLOBAL vertex indices
    const MeshInfo *meshes;
    const EmitterInfo *emitters;
    const float *emitterCdf;
    int nMeshes, nEmitters;
    uint32_t nTris, nVerts;
    // acceleration structure (bvh.cu): 64-byte nodes, leaf-ordered pre-gathered triangles
    const float4 *nodes;         // 4 x float4 per node
    const float4 *leafTris;      // 3 x float4 per triangle slot: (p0, gid) (p1, -) (p2, -)
    uint32_t nNodes;
    const uint4 *qnodes;         // quantised copy of the nodes (NORI_ACCEL_QUANTISED) or nullptr; grid: coordinate = qBase + m qExt
    float qBase[3], qExt[3];
    const float4 *wnodes;        // 4-wide copy of the tree (NORI_ACCEL_WIDE4) or nullptr: 8 x float4 per node
    int isectMode;               // NORI_ISECT_*: ray-triangle test used by every traversal (0 = the reference's Moeller-Trumbore)
    int plainBsdfs;              // 1: no mesh has a microfacet BSDF (host-side census of the uploaded meshes): the shade kernel without the Beckmann code is launched
    Camera cam;
    FilterTab filter;
};

/// Transform * Point3f (include/nori/transform.h:76-80): Eigen evaluates the fixed 4x4 * 4x1
/// product column-wise, ((c0*x + c1*y) + c2*z) + c3*1, then divides by w.
NG_HD V3 mulPoint(const float *m, V3 p) {
    float r[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        r[i] = ((m[4 * i + 0] * p.x + m[4 * i + 1] * p.y) + m[4 * i + 2] * p.z) + m[4 * i + 3] * 1.0f;
    return mk(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
}
/// Transform * Vector3f (transform.h:66-68): coefficient-wise 3x3 product, Eigen's unrolled
/// reduction x0 + (x1 + x2) (Redux.h:94-103) — pinned by tests/golden/ref_vectors.npz (bunny, table).
NG_HD V3 mulVector(const float *m, V3 v) {
    float r[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        r[i] = m[4 * i + 0] * v.x + (m[4 * i + 1] * v.y + m[4 * i + 2] * v.z);
    return mk(r[0], r[1], r[2]);
}

/// PerspectiveCamera::sampleRay (src/perspective.cpp:88-109)
NG_HD Ray sampleRay(const Camera &c, V2 samplePosition) {
    V3 nearP = mulPoint(c.s2c, mk(samplePosition.x * c.invW, samplePosition.y * c.invH, 0.0f));
    V3 d = normalized(nearP);
    float invZ = 1.0f / d.z;
    Ray ray;
    ray.o = c.origin;
    ray.d = mulVector(c.c2w, d);
    ray.mint = c.nearClip * invZ;
    ray.maxt = c.farClip * invZ;
    return ray;
}

/// Intersection (include/nori/mesh.h:35-64) as produced by src/accel.cpp:57-108
struct Hit {
    V3 p; float t; V2 uv; Frame sh, geo; int mesh; uint32_t prim;
};

NG_D V3 ld3(const float4 *p, uint32_t i) { float4 v = __ldg(p + i); return mk(v.x, v.y, v.z); }

/// src/accel.cpp:57-108 for global triangle `gid` hit at (t, u, v).
NG_D void fillHit(const DeviceScene &s, uint32_t gid, float t, float u, float v, Hit &its) {
    uint4 tri = __ldg(s.tris + gid);
    const MeshInfo &m = s.meshes[tri.w];
    its.t = t; its.mesh = (int) tri.w; its.prim = gid - m.triOffset;
    float b0 = 1 - (u + v), b1 = u, b2 = v;
    V3 p0 = ld3(s.vpos, tri.x), p1 = ld3(s.vpos, tri.y), p2 = ld3(s.vpos, tri.z);
    its.p = (b0 * p0 + b1 * p1) + b2 * p2;
    if (m.hasUV) {
        float2 t0 = __ldg(s.vuv + tri.x), t1 = __ldg(s.vuv + tri.y), t2 = __ldg(s.vuv + tri.z);
        its.uv.x = (b0 * t0.x + b1 * t1.x) + b2 * t2.x;
        its.uv.y = (b0 * t0.y + b1 * t1.y) + b2 * t2.y;
    } else { its.uv.x = u; its.uv.y = v; }
    its.geo = Frame(normalized(cross(p1 - p0, p2 - p0)));
    if (m.hasN) {
        V3 n0 = ld3(s.vnrm, tri.x), n1 = ld3(s.vnrm, tri.y), n2 = ld3(s.vnrm, tri.z);
        its.sh = Frame(normalizedDynamic((b0 * n0 + b1 * n1) + b2 * n2));
    } else its.sh = its.geo;
}

} // namespace ng
