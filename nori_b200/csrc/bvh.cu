// bvh.cu — Accel::build on the device (the reference's is a no-op, src/accel.cpp:31-33, and its
// rayIntersect is a brute-force loop): a PLOC (parallel locally-ordered clustering) bottom-up
// builder over Morton-sorted triangle boxes, SAH leaf collapse, and emission of the 64-byte
// traversal nodes + leaf-ordered pre-gathered triangles described in traverse.cuh.
//
// Inputs follow Mesh::getBoundingBox(i) / getCentroid(i) (src/mesh.cpp:90-102). Every step is
// deterministic (no order-dependent atomics decide the topology): the same scene always yields
// the same tree.
#include "bvh.h"
#include <cub/cub.cuh>
#include <vector>
#include <cstdio>
#include <cstdlib>

namespace ng {

#define BVH_CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = std::string(#x) + ": " + cudaGetErrorString(e_); return false; } } while (0)

// monotone float <-> uint mapping for atomicMin/Max on floats
__device__ __forceinline__ uint32_t fkey(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
static inline float fkeyInv(uint32_t k) { uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; float f; memcpy(&f, &u, 4); return f; }

__device__ __forceinline__ float boxArea(float3 lo, float3 hi) {
    float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return 2.0f * (dx * dy + dy * dz + dz * dx);              // bbox.h:86-99 (3D)
}

/// Per-triangle AABB (Mesh::getBoundingBox(i), src/mesh.cpp:90-95) + scene bounds.
__global__ void k_triBounds(const float4 *__restrict__ vpos, const uint4 *__restrict__ tris, uint32_t n,
                            float4 *__restrict__ tlo, float4 *__restrict__ thi, uint32_t *__restrict__ bounds /*6 keys*/) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float3 lo = make_float3(INFINITY, INFINITY, INFINITY), hi = make_float3(-INFINITY, -INFINITY, -INFINITY);
    if (i < n) {
        uint4 t = tris[i];
        float4 a = vpos[t.x], b = vpos[t.y], c = vpos[t.z];
        lo = make_float3(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)));
        hi = make_float3(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)));
        tlo[i] = make_float4(lo.x, lo.y, lo.z, 0.f);
        thi[i] = make_float4(hi.x, hi.y, hi.z, 0.f);
    }
    // warp reduce, then one atomic per warp
    for (int o = 16; o > 0; o >>= 1) {
        lo.x = fminf(lo.x, __shfl_xor_sync(~0u, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(~0u, lo.y, o)); lo.z = fminf(lo.z, __shfl_xor_sync(~0u, lo.z, o));
        hi.x = fmaxf(hi.x, __shfl_xor_sync(~0u, hi.x, o)); hi.y = fmaxf(hi.y, __shfl_xor_sync(~0u, hi.y, o)); hi.z = fmaxf(hi.z, __shfl_xor_sync(~0u, hi.z, o));
    }
    if ((threadIdx.x & 31) == 0 && lo.x <= hi.x) {
        atomicMin(bounds + 0, fkey(lo.x)); atomicMin(bounds + 1, fkey(lo.y)); atomicMin(bounds + 2, fkey(lo.z));
        atomicMax(bounds + 3, fkey(hi.x)); atomicMax(bounds + 4, fkey(hi.y)); atomicMax(bounds + 5, fkey(hi.z));
    }
}

__device__ __forceinline__ uint64_t expand21(uint32_t v) {
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffULL;
    x = (x | x << 16) & 0x1f0000ff0000ffULL;
    x = (x | x << 8) & 0x100f00f00f00f00fULL;
    x = (x | x << 4) & 0x10c30c30c30c30c3ULL;
    x = (x | x << 2) & 0x1249249249249249ULL;
    return x;
}

/// 63-bit Morton code of the box centre inside the scene bounds.
__global__ void k_morton(const float4 *__restrict__ tlo, const float4 *__restrict__ thi, uint32_t n,
                         float3 slo, float3 sinv, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 lo = tlo[i], hi = thi[i];
    float cx = ((lo.x + hi.x) * 0.5f - slo.x) * sinv.x, cy = ((lo.y + hi.y) * 0.5f - slo.y) * sinv.y, cz = ((lo.z + hi.z) * 0.5f - slo.z) * sinv.z;
    uint32_t qx = (uint32_t) fminf(fmaxf(cx * 2097152.0f, 0.0f), 2097151.0f);
    uint32_t qy = (uint32_t) fminf(fmaxf(cy * 2097152.0f, 0.0f), 2097151.0f);
    uint32_t qz = (uint32_t) fminf(fmaxf(cz * 2097152.0f, 0.0f), 2097151.0f);
    keys[i] = (expand21(qx) << 2) | (expand21(qy) << 1) | expand21(qz);
    vals[i] = i;
}

/// Leaves of the build tree: node i (< N) = triangle vals[i] in Morton order.
__global__ void k_initLeaves(const float4 *__restrict__ tlo, const float4 *__restrict__ thi, const uint32_t *__restrict__ vals, uint32_t n,
                             float4 *__restrict__ blo, float4 *__restrict__ bhi, int *__restrict__ parent, uint32_t *__restrict__ cluster) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t g = vals[i];
    float4 lo = tlo[g], hi = thi[g];
    lo.w = __int_as_float(-1); hi.w = __int_as_float(-1);
    blo[i] = lo; bhi[i] = hi; parent[i] = -1; cluster[i] = i;
}

constexpr int kPlocBlock = 256;
constexpr int kPlocMaxRadius = 32;

/// PLOC step 1: nearest neighbour (smallest merged surface area) within +-radius in cluster order.
__global__ void k_plocNearest(const uint32_t *__restrict__ cluster, uint32_t n, int radius,
                              const float4 *__restrict__ blo, const float4 *__restrict__ bhi, uint32_t *__restrict__ nn) {
    __shared__ float3 slo[kPlocBlock + 2 * kPlocMaxRadius], shi[kPlocBlock + 2 * kPlocMaxRadius];
    const int base = (int) (blockIdx.x * kPlocBlock) - radius;
    for (int k = threadIdx.x; k < kPlocBlock + 2 * radius; k += kPlocBlock) {
        int j = base + k;
        if (j >= 0 && j < (int) n) {
            uint32_t c = cluster[j];
            float4 lo = blo[c], hi = bhi[c];
            slo[k] = make_float3(lo.x, lo.y, lo.z); shi[k] = make_float3(hi.x, hi.y, hi.z);
        }
    }
    __syncthreads();
    const int i = (int) (blockIdx.x * kPlocBlock + threadIdx.x);
    if (i >= (int) n) return;
    const float3 lo = slo[threadIdx.x + radius], hi = shi[threadIdx.x + radius];
    float bestA = INFINITY; int bestJ = -1;
    for (int dj = -radius; dj <= radius; ++dj) {
        int j = i + dj;
        if (dj == 0 || j < 0 || j >= (int) n) continue;
        const float3 l2 = slo[threadIdx.x + radius + dj], h2 = shi[threadIdx.x + radius + dj];
        float a = boxArea(make_float3(fminf(lo.x, l2.x), fminf(lo.y, l2.y), fminf(lo.z, l2.z)),
                          make_float3(fmaxf(hi.x, h2.x), fmaxf(hi.y, h2.y), fmaxf(hi.z, h2.z)));
        if (a < bestA) { bestA = a; bestJ = j; }       // ties: lowest j (loop order)
    }
    nn[i] = (uint32_t) bestJ;
}

/// PLOC step 2: mutual nearest neighbours merge; lower index keeps the slot. val = valid | merge<<32.
__global__ void k_plocFlags(const uint32_t *__restrict__ nn, uint32_t n, uint64_t *__restrict__ val) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t m = nn[i];
    bool mutual = m < n && nn[m] == i;
    uint64_t valid = !(mutual && i > m), merge = mutual && i < m;
    val[i] = valid | (merge << 32);
}

/// PLOC step 3: create the merged nodes (ids allocated by scan rank => deterministic) and compact.
__global__ void k_plocMerge(const uint32_t *__restrict__ cluster, const uint32_t *__restrict__ nn, const uint64_t *__restrict__ val,
                            const uint64_t *__restrict__ scan, uint32_t n, uint32_t nextNode,
                            float4 *__restrict__ blo, float4 *__restrict__ bhi, int *__restrict__ parent, uint32_t *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t v = val[i], s = scan[i];
    if (!(v & 1)) return;
    uint32_t pos = (uint32_t) (s & 0xffffffffu);
    if (v >> 32) {
        uint32_t id = nextNode + (uint32_t) (s >> 32);
        uint32_t a = cluster[i], b = cluster[nn[i]];
        float4 la = blo[a], ha = bhi[a], lb = blo[b], hb = bhi[b];
        blo[id] = make_float4(fminf(la.x, lb.x), fminf(la.y, lb.y), fminf(la.z, lb.z), __int_as_float((int) a));
        bhi[id] = make_float4(fmaxf(ha.x, hb.x), fmaxf(ha.y, hb.y), fmaxf(ha.z, hb.z), __int_as_float((int) b));
        parent[a] = (int) id; parent[b] = (int) id; parent[id] = -1;
        out[pos] = id;
    } else out[pos] = cluster[i];
}

/// Optional refinement of the PLOC tree (BvhParams::rotationPasses): tree rotations after Kensler 2008, bottom-up. At an inner
/// node X = (A, B) with B = (B0, B1) the subtree A may change places with B0 or B1 (and symmetrically for A's children) when that
/// shrinks the surface area of the child that is rebuilt — X's own box, hence everything above it, stays as it is. The second
/// thread to arrive at a node processes it (as k_collapse): both subtrees are complete and nobody else touches them.
__global__ void k_rotate(uint32_t n, float4 *__restrict__ blo, float4 *__restrict__ bhi, int *__restrict__ parent, uint32_t *__restrict__ visit, uint32_t *__restrict__ nRotations) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int cur = parent[i];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(visit + cur, 1u) == 0) return;
        __threadfence();
        const int c[2] = {__float_as_int(blo[cur].w), __float_as_int(bhi[cur].w)};
        float bestDelta = 0.f; int bestSide = -1, bestKeep = 0;
        for (int s = 0; s < 2; ++s) {
            const int inner = c[s], other = c[1 - s];
            if (inner < (int) n) continue;                                   // a leaf has no children to trade
            const float4 ilo = blo[inner], ihi = bhi[inner], olo = blo[other], ohi = bhi[other];
            const float aOld = boxArea(make_float3(ilo.x, ilo.y, ilo.z), make_float3(ihi.x, ihi.y, ihi.z));
            const int g[2] = {__float_as_int(ilo.w), __float_as_int(ihi.w)};
            for (int keep = 0; keep < 2; ++keep) {                            // `keep` stays below `inner`, the other grandchild moves up
                const float4 klo = blo[g[keep]], khi = bhi[g[keep]];
                const float a = boxArea(make_float3(fminf(klo.x, olo.x), fminf(klo.y, olo.y), fminf(klo.z, olo.z)),
                                        make_float3(fmaxf(khi.x, ohi.x), fmaxf(khi.y, ohi.y), fmaxf(khi.z, ohi.z)));
                const float d = a - aOld;
                if (d < bestDelta) { bestDelta = d; bestSide = s; bestKeep = keep; }
            }
        }
        if (bestSide >= 0) {
            const int inner = c[bestSide], other = c[1 - bestSide];
            const float4 ilo = blo[inner], ihi = bhi[inner];
            const int g[2] = {__float_as_int(ilo.w), __float_as_int(ihi.w)};
            const int keep = g[bestKeep], up = g[1 - bestKeep];
            const float4 klo = blo[keep], khi = bhi[keep], olo = blo[other], ohi = bhi[other];
            // inner = (keep, other); cur = (inner, up) on the sides they had
            blo[inner] = make_float4(fminf(klo.x, olo.x), fminf(klo.y, olo.y), fminf(klo.z, olo.z), __int_as_float(keep));
            bhi[inner] = make_float4(fmaxf(khi.x, ohi.x), fmaxf(khi.y, ohi.y), fmaxf(khi.z, ohi.z), __int_as_float(other));
            parent[other] = inner; parent[up] = cur;
            float4 plo = blo[cur], phi = bhi[cur];
            if (bestSide == 0) phi.w = __int_as_float(up); else plo.w = __int_as_float(up);
            blo[cur] = plo; bhi[cur] = phi;
            atomicAdd(nRotations, 1u);
        }
        cur = parent[cur];
    }
}

/// Bottom-up SAH pass: subtree triangle counts, SAH cost, and the leaf-collapse decision.
/// The second thread to arrive at a node processes it (both children are then complete).
__global__ void k_collapse(uint32_t n, const float4 *__restrict__ blo, const float4 *__restrict__ bhi, const int *__restrict__ parent,
                           uint32_t *__restrict__ visit, volatile uint32_t *count, volatile float *cost, uint8_t *__restrict__ collapsed,
                           int maxLeaf, float ci, float ct) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    {
        float4 lo = blo[i], hi = bhi[i];
        count[i] = 1; cost[i] = ct * boxArea(make_float3(lo.x, lo.y, lo.z), make_float3(hi.x, hi.y, hi.z));
    }
    int cur = parent[i];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(visit + cur, 1u) == 0) return;
        __threadfence();
        float4 lo = blo[cur], hi = bhi[cur];
        int l = __float_as_int(lo.w), r = __float_as_int(hi.w);
        uint32_t cnt = count[l] + count[r];
        float a = boxArea(make_float3(lo.x, lo.y, lo.z), make_float3(hi.x, hi.y, hi.z));
        float cInner = ci * a + (cost[l] + cost[r]), cLeaf = ct * a * (float) cnt;
        bool leaf = cnt <= (uint32_t) maxLeaf && cLeaf <= cInner;
        collapsed[cur] = leaf ? 1 : 0;
        count[cur] = cnt; cost[cur] = leaf ? cLeaf : cInner;
        cur = parent[cur];
    }
}

/// Walk from every node to the root: DFS position of its first triangle, whether an ancestor was
/// collapsed into a leaf, depth. Leaves also write their triangle into its DFS slot.
__global__ void k_walkUp(uint32_t n /*leaves*/, uint32_t total, const float4 *__restrict__ blo, const float4 *__restrict__ bhi,
                         const int *__restrict__ parent, const uint32_t *__restrict__ count, const uint8_t *__restrict__ collapsed,
                         const uint32_t *__restrict__ vals, const float4 *__restrict__ vpos, const uint4 *__restrict__ tris,
                         uint32_t *__restrict__ start, uint8_t *__restrict__ effective, float4 *__restrict__ leafTris, uint32_t *__restrict__ maxDepth) {
    uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= total) return;
    uint32_t offset = 0, depth = 0; bool under = false;
    int cur = (int) x;
    for (int p = parent[cur]; p >= 0; p = parent[cur]) {
        if (__float_as_int(bhi[p].w) == cur) offset += count[__float_as_int(blo[p].w)];
        if (collapsed[p]) under = true;
        cur = p; ++depth;
    }
    start[x] = offset;
    if (x < n) {
        uint32_t g = vals[x];
        uint4 t = tris[g];
        float4 a = vpos[t.x], b = vpos[t.y], c = vpos[t.z];
        a.w = __uint_as_float(g); b.w = 0.f; c.w = 0.f;
        leafTris[kTriStride * (size_t) offset] = a; leafTris[kTriStride * (size_t) offset + 1] = b; leafTris[kTriStride * (size_t) offset + 2] = c;
        if (kTriStride == 4) leafTris[kTriStride * (size_t) offset + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
        effective[x] = 0;
    } else {
        bool eff = !collapsed[x] && !under;
        effective[x] = eff ? 1 : 0;
        if (eff) atomicMax(maxDepth, depth + 1);
    }
}

/// flagsRev[k] = effective[total-1-k] for the inner nodes, so that the root (highest id) gets id 0
/// and parents precede children.
__global__ void k_revFlags(uint32_t n, uint32_t total, const uint8_t *__restrict__ effective, uint32_t *__restrict__ flagsRev) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= total - n) return;
    flagsRev[k] = effective[total - 1 - k];
}

__global__ void k_emit(uint32_t n, uint32_t total, const float4 *__restrict__ blo, const float4 *__restrict__ bhi,
                       const uint32_t *__restrict__ count, const uint8_t *__restrict__ collapsed, const uint8_t *__restrict__ effective,
                       const uint32_t *__restrict__ start, const uint32_t *__restrict__ newIdRev, float delta,
                       float4 *__restrict__ nodes, uint32_t *__restrict__ leafStats /*[0]=leaves [1]=max leaf size*/) {
    uint32_t x = n + blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= total || !effective[x]) return;
    uint32_t id = newIdRev[total - 1 - x];
    int c[2] = {__float_as_int(blo[x].w), __float_as_int(bhi[x].w)};
    int link[2]; float4 clo[2], chi[2];
    for (int k = 0; k < 2; ++k) {
        clo[k] = blo[c[k]]; chi[k] = bhi[c[k]];
        if (c[k] < (int) n || collapsed[c[k]]) {
            link[k] = ~(int) ((start[c[k]] << 3) | (count[c[k]] - 1));
            atomicAdd(leafStats, 1u); atomicMax(leafStats + 1, count[c[k]]);
        } else link[k] = (int) newIdRev[total - 1 - (uint32_t) c[k]];
    }
    float4 *o = nodes + 4 * (size_t) id;
    o[0] = make_float4(clo[0].x - delta, chi[0].x + delta, clo[0].y - delta, chi[0].y + delta);
    o[1] = make_float4(clo[1].x - delta, chi[1].x + delta, clo[1].y - delta, chi[1].y + delta);
    o[2] = make_float4(clo[0].z - delta, chi[0].z + delta, clo[1].z - delta, chi[1].z + delta);
    o[3] = make_float4(__int_as_float(link[0]), __int_as_float(link[1]), 0.f, 0.f);
}

/// Degenerate scene of one triangle: a root whose right child is an empty (inverted) box.
__global__ void k_emitSingle(const float4 *__restrict__ blo, const float4 *__restrict__ bhi, float delta, float4 *__restrict__ nodes) {
    float4 lo = blo[0], hi = bhi[0];
    nodes[0] = make_float4(lo.x - delta, hi.x + delta, lo.y - delta, hi.y + delta);
    nodes[1] = make_float4(INFINITY, -INFINITY, INFINITY, -INFINITY);
    nodes[2] = make_float4(lo.z - delta, hi.z + delta, INFINITY, -INFINITY);
    nodes[3] = make_float4(__int_as_float(~0), __int_as_float(~0), 0.f, 0.f);
}

namespace {
struct Scratch {
    std::vector<void *> ptrs;
    ~Scratch() { for (void *p : ptrs) cudaFree(p); }
    template <class T> bool alloc(T *&p, size_t n) {
        void *q = nullptr;
        if (cudaMalloc(&q, (n ? n : 1) * sizeof(T)) != cudaSuccess) return false;
        ptrs.push_back(q); p = (T *) q; return true;
    }
};
inline uint32_t cdiv(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
}

bool buildBvh(const float4 *d_vpos, const uint4 *d_tris, uint32_t N, const BvhParams &prm, cudaStream_t st,
              BvhDevice &out, NoriGpuAccelStats &stats, std::string &err) {
    memset(&stats, 0, sizeof(stats));
    if (N == 0) { err = "Accel::build: the scene has no triangles"; return false; }
    if (prm.maxLeaf < 1 || prm.maxLeaf > 8 || prm.plocRadius < 1 || prm.plocRadius > kPlocMaxRadius) { err = "bad BVH parameters"; return false; }
    cudaEvent_t e0, e1;
    BVH_CHECK(cudaEventCreate(&e0)); BVH_CHECK(cudaEventCreate(&e1));
    BVH_CHECK(cudaEventRecord(e0, st));

    const uint32_t total = 2 * N - 1, T = 256;
    Scratch S;
    float4 *tlo, *thi, *blo, *bhi; uint32_t *bounds, *vals, *valsSorted, *clusterA, *clusterB, *nn, *count, *visit, *start, *flagsRev, *newIdRev, *misc;
    uint64_t *keys, *keysSorted, *val, *scan; int *parent; float *cost; uint8_t *collapsed, *effective;
    bool ok = S.alloc(tlo, N) && S.alloc(thi, N) && S.alloc(blo, total) && S.alloc(bhi, total) && S.alloc(bounds, 6) && S.alloc(vals, N) &&
              S.alloc(valsSorted, N) && S.alloc(clusterA, N) && S.alloc(clusterB, N) && S.alloc(nn, N) && S.alloc(count, total) &&
              S.alloc(visit, total) && S.alloc(start, total) && S.alloc(flagsRev, N) && S.alloc(newIdRev, N) && S.alloc(misc, 4) &&
              S.alloc(keys, N) && S.alloc(keysSorted, N) && S.alloc(val, N + 1) && S.alloc(scan, N + 1) && S.alloc(parent, total) &&
              S.alloc(cost, total) && S.alloc(collapsed, total) && S.alloc(effective, total);
    if (!ok) { err = "cudaMalloc failed in buildBvh"; return false; }

    // 1. triangle boxes + scene bounds
    uint32_t hb[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    BVH_CHECK(cudaMemcpyAsync(bounds, hb, sizeof(hb), cudaMemcpyHostToDevice, st));
    k_triBounds<<<cdiv(N, T), T, 0, st>>>(d_vpos, d_tris, N, tlo, thi, bounds);
    BVH_CHECK(cudaMemcpyAsync(hb, bounds, sizeof(hb), cudaMemcpyDeviceToHost, st));
    BVH_CHECK(cudaStreamSynchronize(st));
    float lo[3], hi[3], diag2 = 0, maxAbs = 0;
    for (int k = 0; k < 3; ++k) {
        lo[k] = fkeyInv(hb[k]); hi[k] = fkeyInv(hb[3 + k]);
        if (!(lo[k] <= hi[k]) || !std::isfinite(lo[k]) || !std::isfinite(hi[k])) { err = "Accel::build: non-finite vertex positions"; return false; }
        diag2 += (hi[k] - lo[k]) * (hi[k] - lo[k]);
        maxAbs = std::max(maxAbs, std::max(std::fabs(lo[k]), std::fabs(hi[k])));
        out.sceneLo[k] = lo[k]; out.sceneHi[k] = hi[k];
    }
    const float delta = prm.inflate * std::max(std::sqrt(diag2), maxAbs);
    float3 slo = make_float3(lo[0], lo[1], lo[2]);
    float3 sinv = make_float3(hi[0] > lo[0] ? 1.0f / (hi[0] - lo[0]) : 0.f, hi[1] > lo[1] ? 1.0f / (hi[1] - lo[1]) : 0.f, hi[2] > lo[2] ? 1.0f / (hi[2] - lo[2]) : 0.f);

    // 2. Morton order
    k_morton<<<cdiv(N, T), T, 0, st>>>(tlo, thi, N, slo, sinv, keys, vals);
    size_t tmpBytes = 0, tmpScan = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, keys, keysSorted, vals, valsSorted, (int) N, 0, 63, st);
    cub::DeviceScan::ExclusiveSum(nullptr, tmpScan, val, scan, (int) N + 1, st);
    tmpBytes = std::max(tmpBytes, tmpScan);
    { size_t t2 = 0; cub::DeviceScan::ExclusiveSum(nullptr, t2, flagsRev, newIdRev, (int) N, st); tmpBytes = std::max(tmpBytes, t2); }
    uint8_t *tmp; if (!S.alloc(tmp, tmpBytes)) { err = "cudaMalloc failed in buildBvh"; return false; }
    BVH_CHECK(cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, keys, keysSorted, vals, valsSorted, (int) N, 0, 63, st));
    k_initLeaves<<<cdiv(N, T), T, 0, st>>>(tlo, thi, valsSorted, N, blo, bhi, parent, clusterA);

    // 3. PLOC: merge mutual nearest neighbours until one cluster is left
    uint32_t n = N, nextNode = N, iters = 0;
    uint32_t *cin = clusterA, *cout = clusterB;
    while (n > 1) {
        k_plocNearest<<<cdiv(n, kPlocBlock), kPlocBlock, 0, st>>>(cin, n, prm.plocRadius, blo, bhi, nn);
        k_plocFlags<<<cdiv(n, T), T, 0, st>>>(nn, n, val);
        BVH_CHECK(cudaMemsetAsync(val + n, 0, sizeof(uint64_t), st));
        BVH_CHECK(cub::DeviceScan::ExclusiveSum(tmp, tmpBytes, val, scan, (int) n + 1, st));
        k_plocMerge<<<cdiv(n, T), T, 0, st>>>(cin, nn, val, scan, n, nextNode, blo, bhi, parent, cout);
        uint64_t tot = 0;
        BVH_CHECK(cudaMemcpyAsync(&tot, scan + n, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
        BVH_CHECK(cudaStreamSynchronize(st));
        uint32_t nValid = (uint32_t) (tot & 0xffffffffu), nMerged = (uint32_t) (tot >> 32);
        if (nMerged == 0 || nValid != n - nMerged) { err = "PLOC made no progress (internal error)"; return false; }
        nextNode += nMerged; n = nValid; std::swap(cin, cout); ++iters;
    }
    if (nextNode != total) { err = "PLOC node count mismatch (internal error)"; return false; }
    const uint32_t root = total - 1;

    // 3b. optional rotations (off by default: measured, profiles/r02_notes.md)
    for (int pass = 0; pass < prm.rotationPasses; ++pass) {
        BVH_CHECK(cudaMemsetAsync(visit, 0, total * sizeof(uint32_t), st));
        BVH_CHECK(cudaMemsetAsync(misc, 0, 4 * sizeof(uint32_t), st));
        k_rotate<<<cdiv(N, T), T, 0, st>>>(N, blo, bhi, parent, visit, misc + 3);
    }

    // 4. SAH collapse, DFS layout, emission
    BVH_CHECK(cudaMemsetAsync(visit, 0, total * sizeof(uint32_t), st));
    BVH_CHECK(cudaMemsetAsync(collapsed, 0, total, st));
    BVH_CHECK(cudaMemsetAsync(misc, 0, 4 * sizeof(uint32_t), st));
    k_collapse<<<cdiv(N, T), T, 0, st>>>(N, blo, bhi, parent, visit, count, cost, collapsed, prm.maxLeaf, prm.costInner, prm.costTri);
    BVH_CHECK(cudaMemsetAsync(collapsed + root, 0, 1, st));       // the root is always an inner node
    float4 *leafTris = nullptr, *nodes = nullptr;
    BVH_CHECK(cudaMalloc((void **) &leafTris, kTriStride * (size_t) N * sizeof(float4)));
    k_walkUp<<<cdiv(total, T), T, 0, st>>>(N, total, blo, bhi, parent, count, collapsed, valsSorted, d_vpos, d_tris, start, effective, leafTris, misc + 2);
    uint32_t nNodes = 1;
    if (N == 1) {
        BVH_CHECK(cudaMalloc((void **) &nodes, 4 * sizeof(float4)));
        k_emitSingle<<<1, 1, 0, st>>>(blo, bhi, delta, nodes);
        stats.n_leaves = 1; stats.max_leaf_size = 1; stats.depth = 1;
    } else {
        k_revFlags<<<cdiv(N - 1, T), T, 0, st>>>(N, total, effective, flagsRev);
        BVH_CHECK(cudaMemsetAsync(flagsRev + (N - 1), 0, sizeof(uint32_t), st));
        BVH_CHECK(cub::DeviceScan::ExclusiveSum(tmp, tmpBytes, flagsRev, newIdRev, (int) N, st));
        BVH_CHECK(cudaMemcpyAsync(&nNodes, newIdRev + (N - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        BVH_CHECK(cudaStreamSynchronize(st));
        BVH_CHECK(cudaMalloc((void **) &nodes, 4 * (size_t) nNodes * sizeof(float4)));
        k_emit<<<cdiv(N - 1, T), T, 0, st>>>(N, total, blo, bhi, count, collapsed, effective, start, newIdRev, delta, nodes, misc);
        uint32_t hm[4]; float rootCost = 0; float4 rlo, rhi;
        BVH_CHECK(cudaMemcpyAsync(hm, misc, sizeof(hm), cudaMemcpyDeviceToHost, st));
        BVH_CHECK(cudaMemcpyAsync(&rootCost, cost + root, 4, cudaMemcpyDeviceToHost, st));
        BVH_CHECK(cudaMemcpyAsync(&rlo, blo + root, 16, cudaMemcpyDeviceToHost, st));
        BVH_CHECK(cudaMemcpyAsync(&rhi, bhi + root, 16, cudaMemcpyDeviceToHost, st));
        BVH_CHECK(cudaStreamSynchronize(st));
        stats.n_leaves = hm[0]; stats.max_leaf_size = hm[1]; stats.depth = hm[2];
        float dx = rhi.x - rlo.x, dy = rhi.y - rlo.y, dz = rhi.z - rlo.z, ra = 2.0f * (dx * dy + dy * dz + dz * dx);
        stats.sah_cost = ra > 0 ? rootCost / ra : 0.0f;
    }
    BVH_CHECK(cudaEventRecord(e1, st));
    BVH_CHECK(cudaStreamSynchronize(st));
    BVH_CHECK(cudaGetLastError());
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1); cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (stats.depth > 62) { cudaFree(nodes); cudaFree(leafTris); err = "BVH deeper than the traversal stack (62)"; return false; }
    out.nodes = nodes; out.leafTris = leafTris; out.nNodes = nNodes; out.nTris = N;
    stats.n_triangles = N; stats.n_nodes = nNodes; stats.build_ms = ms;
    (void) iters;
    return true;
}

// ---------------------------------------------------------------------------------------------
// Quantised node layout (include/nori_gpu.h: NORI_ACCEL_QUANTISED)
__global__ void k_quantiseNodes(const float4 *__restrict__ nodes, uint32_t nNodes, double3 lo, double3 invCell, uint4 *__restrict__ q) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nNodes) return;
    const float4 n0 = nodes[4 * (size_t) i], n1 = nodes[4 * (size_t) i + 1], n2 = nodes[4 * (size_t) i + 2], n3 = nodes[4 * (size_t) i + 3];
    auto qmin = [](float v, double l, double ic) { const double x = floor(((double) v - l) * ic); return (uint32_t) fmin(fmax(x, 0.0), 32767.0) | 0x8000u; };
    auto qmax = [](float v, double l, double ic) { const double x = ceil(((double) v - l) * ic); return (uint32_t) fmin(fmax(x, 0.0), 32767.0) | 0x8000u; };
    auto axis = [&](float mn, float mx, double l, double ic) { return qmin(mn, l, ic) | (qmax(mx, l, ic) << 16); };
    uint4 a, b;
    a.x = axis(n0.x, n0.y, lo.x, invCell.x); a.y = axis(n0.z, n0.w, lo.y, invCell.y); a.z = axis(n2.x, n2.y, lo.z, invCell.z); a.w = __float_as_uint(n3.x);
    b.x = axis(n1.x, n1.y, lo.x, invCell.x); b.y = axis(n1.z, n1.w, lo.y, invCell.y); b.z = axis(n2.z, n2.w, lo.z, invCell.z); b.w = __float_as_uint(n3.y);
    q[2 * (size_t) i] = a; q[2 * (size_t) i + 1] = b;
}

bool quantiseNodes(BvhDevice &bvh, float inflate, cudaStream_t st, std::string &err) {
    if (!bvh.nodes || !bvh.nNodes) { err = "quantiseNodes: no BVH"; return false; }
    if (bvh.qnodes) return true;
    double diag2 = 0, maxAbs = 0;
    for (int k = 0; k < 3; ++k) {
        diag2 += ((double) bvh.sceneHi[k] - bvh.sceneLo[k]) * ((double) bvh.sceneHi[k] - bvh.sceneLo[k]);
        maxAbs = std::max(maxAbs, (double) std::max(std::fabs(bvh.sceneLo[k]), std::fabs(bvh.sceneHi[k])));
    }
    // the stored child boxes are the exact ones grown by inflate * scale (buildBvh): the grid must contain them
    const double margin = 4.0 * inflate * std::max(std::sqrt(diag2), maxAbs) + 1e-30;
    double lo[3], cell[3];
    for (int k = 0; k < 3; ++k) {
        const double ext = std::max(((double) bvh.sceneHi[k] - bvh.sceneLo[k]) + 2.0 * margin, 1e-20);
        lo[k] = (double) bvh.sceneLo[k] - margin - 2.0 * ext / 32760.0;            // two spare cells on either side
        const float E = (float) (ext * (32768.0 / 32760.0) * (1.0 + 4.0 / 32760.0));
        // what the traversal reconstructs: qBase + m * qExt with m = 1 + q / 32768, i.e. lo + q * E / 32768
        bvh.qExt[k] = E; bvh.qBase[k] = (float) lo[k] - E;
        lo[k] = (double) bvh.qBase[k] + (double) E;                                     // the grid as the FLOATS define it
        cell[k] = (double) E / 32768.0;
    }
    uint4 *q = nullptr;
    if (cudaMalloc((void **) &q, 2 * (size_t) bvh.nNodes * sizeof(uint4)) != cudaSuccess) { err = "quantiseNodes: out of device memory"; return false; }
    k_quantiseNodes<<<(bvh.nNodes + 255u) / 256u, 256, 0, st>>>(bvh.nodes, bvh.nNodes, make_double3(lo[0], lo[1], lo[2]),
                                                                make_double3(1.0 / cell[0], 1.0 / cell[1], 1.0 / cell[2]), q);
    if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) { cudaFree(q); err = "quantiseNodes: kernel failed"; return false; }
    bvh.qnodes = q;
    return true;
}

// ---------------------------------------------------------------------------------------------
// 4-wide node layout (include/nori_gpu.h: NORI_ACCEL_WIDE4): the binary tree collapsed top-down. A wide node is rooted at a
// binary inner node; its children start as that node's two and, while fewer than four, the inner child with the largest
// surface area is replaced by its own two (the greedy collapse of Wald et al. 2008 / Dammertz et al. 2008). One launch per
// level of the wide tree: a job = (binary node, wide node index); every inner child found gets the next wide index and becomes
// a job of the next level.
//   wide node = 8 x float4 (128 B, four 256-bit loads): lo.x[4] hi.x[4] lo.y[4] hi.y[4] lo.z[4] hi.z[4] link[4] -,
//   link as in the binary tree (>= 0: wide node index, < 0: leaf code), kWideEmpty for an unused child.
__device__ __forceinline__ float wideArea(const float *b) {          // b = lo.xyz hi.xyz
    const float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
    return dx * dy + dy * dz + dz * dx;
}
__device__ __forceinline__ void wideChildren(const float4 *__restrict__ nodes, uint32_t bin, float (*box)[6], int *link, int a, int b) {
    const float4 n0 = nodes[4 * (size_t) bin], n1 = nodes[4 * (size_t) bin + 1], n2 = nodes[4 * (size_t) bin + 2], n3 = nodes[4 * (size_t) bin + 3];
    box[a][0] = n0.x; box[a][3] = n0.y; box[a][1] = n0.z; box[a][4] = n0.w; box[a][2] = n2.x; box[a][5] = n2.y; link[a] = __float_as_int(n3.x);
    box[b][0] = n1.x; box[b][3] = n1.y; box[b][1] = n1.z; box[b][4] = n1.w; box[b][2] = n2.z; box[b][5] = n2.w; link[b] = __float_as_int(n3.y);
}
__global__ void k_wideLevel(const float4 *__restrict__ nodes, const uint2 *__restrict__ jobsIn, uint32_t nIn, uint2 *__restrict__ jobsOut,
                            uint32_t *__restrict__ counters /* [0] next wide index, [1] jobs written */, float4 *__restrict__ wnodes) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nIn) return;
    const uint2 job = jobsIn[t];
    float box[4][6]; int link[4]; int cnt = 2;
    wideChildren(nodes, job.x, box, link, 0, 1);
    while (cnt < 4) {
        int pick = -1; float best = -1.0f;
        for (int c = 0; c < cnt; ++c) if (link[c] >= 0) { const float a = wideArea(box[c]); if (a > best) { best = a; pick = c; } }
        if (pick < 0) break;
        wideChildren(nodes, (uint32_t) link[pick], box, link, pick, cnt);
        ++cnt;
    }
    for (int c = 0; c < cnt; ++c) if (link[c] >= 0) {
        const uint32_t w = atomicAdd(&counters[0], 1u), pos = atomicAdd(&counters[1], 1u);
        jobsOut[pos] = make_uint2((uint32_t) link[c], w);
        link[c] = (int) w;
    }
    for (int c = cnt; c < 4; ++c) { for (int k = 0; k < 6; ++k) box[c][k] = 0.0f; link[c] = kWideEmpty; }
    float4 *o = wnodes + 8 * (size_t) job.y;
    for (int k = 0; k < 3; ++k) {
        o[2 * k] = make_float4(box[0][k], box[1][k], box[2][k], box[3][k]);
        o[2 * k + 1] = make_float4(box[0][3 + k], box[1][3 + k], box[2][3 + k], box[3][3 + k]);
    }
    o[6] = make_float4(__int_as_float(link[0]), __int_as_float(link[1]), __int_as_float(link[2]), __int_as_float(link[3]));
    o[7] = make_float4(0.f, 0.f, 0.f, 0.f);
}

bool wideNodes(BvhDevice &bvh, cudaStream_t st, std::string &err) {
    if (!bvh.nodes || !bvh.nNodes) { err = "wideNodes: no BVH"; return false; }
    if (bvh.wnodes) return true;
    const uint32_t n = bvh.nNodes;
    float4 *w = nullptr; uint2 *jobs[2] = {nullptr, nullptr}; uint32_t *counters = nullptr;
    auto fail = [&](const char *what) { cudaFree(w); cudaFree(jobs[0]); cudaFree(jobs[1]); cudaFree(counters); err = what; return false; };
    if (cudaMalloc((void **) &w, 8 * (size_t) n * sizeof(float4)) != cudaSuccess || cudaMalloc((void **) &jobs[0], (size_t) n * sizeof(uint2)) != cudaSuccess ||
        cudaMalloc((void **) &jobs[1], (size_t) n * sizeof(uint2)) != cudaSuccess || cudaMalloc((void **) &counters, 2 * sizeof(uint32_t)) != cudaSuccess)
        return fail("wideNodes: out of device memory");
    const uint2 root = make_uint2(0u, 0u);
    uint32_t hc[2] = {1u, 0u}, nIn = 1, depth = 0;
    if (cudaMemcpyAsync(jobs[0], &root, sizeof(root), cudaMemcpyHostToDevice, st) != cudaSuccess) return fail("wideNodes: copy failed");
    while (nIn) {
        hc[1] = 0;
        if (cudaMemcpyAsync(counters, hc, sizeof(hc), cudaMemcpyHostToDevice, st) != cudaSuccess) return fail("wideNodes: copy failed");
        k_wideLevel<<<(nIn + 127u) / 128u, 128, 0, st>>>(bvh.nodes, jobs[depth & 1], nIn, jobs[(depth + 1) & 1], counters, w);
        if (cudaMemcpyAsync(hc, counters, sizeof(hc), cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess)
            return fail("wideNodes: kernel failed");
        nIn = hc[1]; ++depth;
        if (hc[0] > n || depth > 4096) return fail("wideNodes: the binary tree is not a tree (internal error)");
    }
    // a visit pushes at most three entries: the traversal's stack (kWideStackSize) must hold 3 per level
    if (3u * depth + 1u > (uint32_t) kWideStackSize) return fail("wideNodes: the collapsed tree is too deep for the traversal stack");
    cudaFree(jobs[0]); cudaFree(jobs[1]); cudaFree(counters);
    bvh.wnodes = w; bvh.nWide = hc[0]; bvh.wideDepth = depth;
    return true;
}

} // namespace ng
