// render.cu — the wavefront path tracer: everything the reference runs inside
// `tbb::parallel_for` (src/main.cpp:93-121: renderBlock -> Camera::sampleRay -> Integrator::Li ->
// ImageBlock::put) as a fixed pool of path slots advanced by one bounce per iteration:
//
//   k_prepare      (1 thread)  statistics, ring bookkeeping, sample issue limit
//   k_traceClosest (slot)      Accel::rayIntersect(ray, its, false) for every live path
//   k_shadeRegen   (slot)      one Li loop body per path (emitter, next-event estimation -> shadow
//                              queue, BSDF sample, Russian roulette), material-sorted per CTA; paths
//                              that end write their sample record and their slot draws the next
//                              camera sample in place (Sampler stream + PerspectiveCamera::sampleRay)
//   k_traceShadow  (queue)     Accel::rayIntersect(ray, its, true); add unoccluded contributions
//   k_splatPlan / k_splatPixel ImageBlock::put for all samples of a finished 32x32 tile, accumulated
//                              in registers and shared memory, then added to the full-frame block
//
// Integrator semantics follow oracle/noracle_integrators.h statement for statement (the reference
// ships no Integrator subclass: SURVEY.md F3), including the order in which a path consumes its
// random numbers, so that a camera sample yields the same radiance on both sides.
#include "engine.h"
#include <cstdlib>
#include "traverse_persistent.cuh"
#include <cuda_pipeline.h>

namespace ng {

__device__ __forceinline__ unsigned laneId() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemaskLt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

/// Warp-aggregated `atomicAdd(counter, 1)` for the lanes with pred set; returns this lane's index.
__device__ __forceinline__ uint32_t warpAppend(unsigned int *counter, bool pred) {
    unsigned m = __ballot_sync(0xffffffffu, pred);
    if (!m) return 0;
    int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if ((int) laneId() == leader) base = atomicAdd(counter, (unsigned) __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(m & lanemaskLt());
}

// ------------------------------------------------------------------------------------------
__global__ void k_prepare(Counters *c, RenderParams rp) {
    unsigned rays = 0;
    for (int k = 0; k < kNumKeys; ++k) { rays += c->sortCount[k]; c->sortCount[k] = 0; }
    c->closestRays += rays; c->segments += rays; c->liveLast = rays;
    c->shadowRays += c->shadowCount; c->shadowCount = 0;
    c->iterations++;
    c->traceCursor[0] = 0; c->traceCursor[1] = 0;      // [0]: for the NEXT iteration's closest-hit pass (it runs before k_prepare)
    c->readyCount = 0;
    // lanes that drew an index at or past the previous limit discarded it: hand those indices out again
    if (c->issue > c->issueLimit) c->issue = c->issueLimit;
    c->passHint = rp.listPixel ? 0u : (unsigned) (c->issue / rp.passSlots);   // pass of the next index to hand out
    unsigned cp = c->completedPasses;
    while (cp < rp.numPasses && c->tilesSplatted[cp % rp.ring] >= rp.numTiles * (cp / rp.ring + 1)) ++cp;
    c->completedPasses = cp;
    for (int s = 0; s < rp.ring; ++s) c->slotPass[s] = s + rp.ring * (c->tilesSplatted[s] / rp.numTiles);
    unsigned long long limitPass = min((unsigned long long) rp.numPasses, (unsigned long long) cp + rp.ring);
    unsigned long long limit = limitPass == rp.numPasses ? rp.totalIssue : limitPass * rp.passSlots;
    if (rp.listPixel) limit = rp.totalIssue;
    c->issueLimit = limit;
    // Ring about to run dry while the oldest pass still waits for a few long paths: close that pass
    // now (k_splat takes the samples it has); its stragglers are splatted one by one when they finish.
    c->forcePass = (!rp.listPixel && cp < rp.numPasses && limit < rp.totalIssue && c->issue + rp.poolCapacity > limit) ? cp : 0xffffffffu;
}

/// Pool::state = the sampler stream of a path and its flags in one 16-byte word. pcg32's increment is
/// (initseq << 1) | 1 with initseq a sample index or a tile offset: 32 bits are enough (checked on the host).
__device__ __forceinline__ uint4 packState(const Pcg32 &rng, uint32_t flags) {
    return make_uint4((uint32_t) rng.state, (uint32_t) (rng.state >> 32), (uint32_t) rng.inc, flags);
}
__device__ __forceinline__ Pcg32 unpackRng(const uint4 &st) {
    Pcg32 r; r.state = (uint64_t) st.x | ((uint64_t) st.y << 32); r.inc = (uint64_t) st.z;
    return r;
}

/// Sort word of a bounce ray: bin (2 bits of origin cell per axis of the scene box, two direction sign bits) << 24 | slot.
#ifndef NORI_BIN_DIRS
#define NORI_BIN_DIRS 0      // which two direction signs: 0 = (x, y), 1 = (x, z), 2 = (y, z)
#endif
__device__ __forceinline__ uint32_t binWord(const Pool &pool, const Ray &ray, uint32_t slot) {
    const int cx = min(max((int) ((ray.o.x - pool.binLo[0]) * pool.binInv[0]), 0), 3);
    const int cy = min(max((int) ((ray.o.y - pool.binLo[1]) * pool.binInv[1]), 0), 3);
    const int cz = min(max((int) ((ray.o.z - pool.binLo[2]) * pool.binInv[2]), 0), 3);
#if NORI_BIN_DIRS >= 3      // 3 / 4: the full direction octant, paid for with one bit of the z / y cell
    const int cell = NORI_BIN_DIRS == 3 ? (cx | (cy << 2) | ((cz >> 1) << 4)) : (cx | (cz << 2) | ((cy >> 1) << 4));
    const uint32_t bin = min((uint32_t) ((cell << 3) | (ray.d.x < 0.f ? 1 : 0) | (ray.d.y < 0.f ? 2 : 0) | (ray.d.z < 0.f ? 4 : 0)), 0xfdu);
#else
    const float da = NORI_BIN_DIRS == 2 ? ray.d.y : ray.d.x, db = NORI_BIN_DIRS == 0 ? ray.d.y : ray.d.z;
    const uint32_t bin = min((uint32_t) (((cx | (cy << 2) | (cz << 4)) << 2) | (da < 0.f ? 1 : 0) | (db < 0.f ? 2 : 0)), 0xfdu);
#endif
    return (bin << 24) | slot;
}

/// Pixel of slot r (< 1024) inside a tile: warps cover 8x4 pixel patches (coherent primary rays).
__device__ __forceinline__ void tilePixel(uint32_t pix, int &lx, int &ly) {
    uint32_t w = pix >> 5, l = pix & 31u;
    lx = (int) ((w & 3u) * 8u + (l & 7u));
    ly = (int) ((w >> 2) * 4u + (l >> 3));
}

/// ImageBlock::put(pos, value) (src/block.cpp:74-103) straight into the full-frame block with float
/// atomics: only for the few samples that finish after their pass was closed.
__device__ void splatDirect(const DeviceScene &sc, const RenderParams &rp, Counters *c, float2 ps, V3 val) {
    if (!isValidColor(val)) { atomicAdd(&c->dropped, 1ull); return; }
    const int b = sc.filter.border;
    const float radius = sc.filter.radius, lookup = sc.filter.lookupFactor;
    const float px = ps.x - 0.5f - (float) (0 - b), py = ps.y - 0.5f - (float) (0 - b);
    const int x0 = max((int) ceilf(px - radius), 0), y0 = max((int) ceilf(py - radius), 0);
    const int x1 = min((int) floorf(px + radius), rp.blockCols - 1), y1 = min((int) floorf(py + radius), rp.blockRows - 1);
    for (int y = y0; y <= y1; ++y) {
        const float wy = sc.filter.table[(int) (fabsf((float) y - py) * lookup)];
        for (int x = x0; x <= x1; ++x) {
            const float wx = sc.filter.table[(int) (fabsf((float) x - px) * lookup)];
            float4 *dst = reinterpret_cast<float4 *>(rp.block) + (size_t) y * rp.blockCols + x;
            atomicAdd(dst, make_float4(val.x * wx * wy, val.y * wx * wy, val.z * wx * wy, 1.0f * wx * wy));
        }
    }
}

// ------------------------------------------------------------------------------------------
struct ClosestIO {
    Pool pool;
    bool sorted;        // the order list was written by the bin sort of this iteration: entries below Counters::traceLimit are live slots
    /// slot: in = k, out = the path slot the k-th entry of the order list names (the handle store() gets back)
    __device__ __forceinline__ bool load(uint32_t k, uint32_t &slot, V3 &o, V3 &d, float &mint, float &maxt) const {
        NG_CHECK(k < pool.capacity);
        const uint32_t i = pool.order ? (pool.order[k] & kBinSlotMask) : k;
        NG_CHECK(i < pool.capacity);
        // a sorted list holds the live slots first (fresh and bounce bins), the empty ones behind traceLimit: no need to ask
        if (!sorted) { if (!(pool.state[i].w & kFlagAlive)) return false; }
        else NG_CHECK(pool.state[i].w & kFlagAlive);
        const float4 ro = pool.rayO[i], rd = pool.rayD[i];
        o = mk(ro.x, ro.y, ro.z); d = mk(rd.x, rd.y, rd.z); mint = ro.w; maxt = rd.w;
        slot = i;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t slot, uint32_t gid, float t, float u, float v) const {
        NG_CHECK(slot < pool.capacity);
        pool.hit[slot] = make_float4(t, u, v, __uint_as_float(gid));
    }
};

template <> struct IoIndirect<ClosestIO> { static constexpr bool value = true; };

/// Accel::rayIntersect(ray, its, false) for every live path (persistent warps, dynamic fetch).
__device__ __forceinline__ QGrid sceneGrid(const DeviceScene &sc) { QGrid g = {sc.qBase[0], sc.qBase[1], sc.qBase[2], sc.qExt[0], sc.qExt[1], sc.qExt[2]}; return g; }

template <int MINB, bool WT = false, int QN = 0>
__global__ void __launch_bounds__(128, MINB) k_traceClosest(DeviceScene sc, Pool pool, Counters *c, int fetchThreshold) {
    ClosestIO io; io.pool = pool; io.sorted = pool.order && c->orderSorted;
    const uint32_t nRays = pool.order ? min(c->traceLimit, pool.capacity) : pool.capacity;
    tracePersistent<false, WT, QN>(QN == 2 ? sc.wnodes : (QN == 1 ? reinterpret_cast<const float4 *>(sc.qnodes) : sc.nodes), sc.leafTris, nRays, &c->traceCursor[0], io, fetchThreshold, sceneGrid(sc));
    // This grid is a programmatic dependent of the splat kernels (they execute launch_dependents on entry). It must
    // not COMPLETE before they have: k_prepare, a normal launch behind it, is ordered after this grid only, and it
    // resets counters the splat still reads. Waiting here, after the last hit is stored, keeps the overlap and makes
    // stream order transitive again (a no-op when the launch carried no programmatic attribute).
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ------------------------------------------------------------------------------------------
// `area` emitter (oracle/noracle_integrators.h:55-88, noracle_scene.h:135-147)
struct LightSample { V3 p, n, wi; float dist, pdf; V3 Le; };

/// DiscretePDF::sample (include/nori/dpdf.h:118-123): lower_bound on the CDF.
__device__ __forceinline__ uint32_t cdfSample(const float *__restrict__ cdf, uint32_t nEntries /* size()+1 */, float v) {
    uint32_t lo = 0, len = nEntries;
    while (len > 0) {                              // std::lower_bound
        uint32_t half = len >> 1;
        if (__ldg(cdf + lo + half) < v) { lo += half + 1; len -= half + 1; } else len = half;
    }
    int idx = max(0, (int) lo - 1);
    return min((uint32_t) idx, nEntries - 2);
}

__device__ __forceinline__ void emitterSamplePoint(const DeviceScene &s, const EmitterInfo &em, V2 xi, V3 &p, V3 &n) {
    const float *cdf = s.emitterCdf + em.cdfOffset;
    float x = xi.x;
    uint32_t f = cdfSample(cdf, em.nTris + 1, x);
    float c0 = __ldg(cdf + f), c1 = __ldg(cdf + f + 1);
    x = (x - c0) / (c1 - c0);                      // DiscretePDF::sampleReuse (dpdf.h:151-156)
    float su = sqrtf(1.0f - x);
    float b0 = 1.0f - su, b1 = xi.y * su, b2 = 1.0f - b0 - b1;
    uint4 tri = __ldg(s.tris + em.triOffset + f);
    V3 p0 = ld3(s.vpos, tri.x), p1 = ld3(s.vpos, tri.y), p2 = ld3(s.vpos, tri.z);
    p = (b0 * p0 + b1 * p1) + b2 * p2;
    if (s.meshes[em.mesh].hasN) n = normalized((b0 * ld3(s.vnrm, tri.x) + b1 * ld3(s.vnrm, tri.y)) + b2 * ld3(s.vnrm, tri.z));
    else n = normalized(cross(p1 - p0, p2 - p0));
}

/// One emitter picked uniformly from u, then a uniform point by area (noracle_integrators.h:73-88).
__device__ __forceinline__ bool sampleLights(const DeviceScene &s, V3 ref, float u, V2 xi, LightSample &ls) {
    int n = s.nEmitters;
    if (n == 0) return false;
    int e = min((int) (u * (float) n), n - 1);
    const EmitterInfo &em = s.emitters[e];
    emitterSamplePoint(s, em, xi, ls.p, ls.n);
    V3 d = ls.p - ref;
    float dist2 = squaredNorm(d);
    ls.dist = sqrtf(dist2);
    ls.wi = d / ls.dist;
    float cosL = dot(ls.n, -ls.wi);
    if (!(cosL > 0.0f) || !(ls.dist > 0.0f)) { ls.pdf = 0.0f; return false; }
    ls.pdf = dist2 / (cosL * em.area) / (float) n;
    ls.Le = mk(em.radiance[0], em.radiance[1], em.radiance[2]);
    return true;
}

__device__ __forceinline__ float emitterPdfSolidAngle(const EmitterInfo &em, V3 ref, V3 p, V3 n) {
    V3 d = p - ref;
    float dist2 = squaredNorm(d);
    float dist = sqrtf(dist2);
    float cosL = dot(n, -(d / dist));
    if (!(cosL > 0.0f)) return 0.0f;
    return dist2 / (cosL * em.area);
}

__device__ __forceinline__ Bsdf meshBsdf(const MeshInfo &m) {
    Bsdf b;
    b.type = m.bsdfType; b.albedo = mk(m.albedo[0], m.albedo[1], m.albedo[2]);
    b.alpha = m.alpha; b.intIOR = m.intIOR; b.extIOR = m.extIOR; b.ks = m.ks;
    return b;
}

struct ShadowOut { bool cast; V3 o, d; float maxt; V3 contrib; };

/// Body of the Li loops of oracle/noracle_integrators.h for ONE path vertex. Returns true while
/// the path continues (a new ray was written to `ray`).
/// What the path-graph tool's LPoint records want of a vertex (k_pathVertices only; nullptr everywhere else: the code folds away).
struct LightOut { V3 Lem, Ldirect; float lightPdf; };

/// PLAIN: the caller knows that no mesh of the scene has a microfacet BSDF (DeviceScene::plainBsdfs); the Beckmann branches of
/// Bsdf::eval / pdf / sample — by far the largest — are then dead code and leave the kernel (same results: they would never have run).
template <int INTEGRATOR, bool PLAIN = false>
__device__ __forceinline__ bool shadeVertex(const DeviceScene &sc, const RenderParams &rp, Ray &ray, uint32_t gid, float t, float u, float v,
                                            Pcg32 &rng, V3 &T, V3 &L, float &eta, float &bsdfPdf, uint32_t &flags, ShadowOut &sh,
                                            const float4 *lazyOrigin = nullptr, float *rrProb = nullptr, LightOut *lo = nullptr) {
    // lazyOrigin: where the incoming ray's origin can be read when `ray.o` was not loaded. Only the MIS weight of an
    // emitter hit needs it, so the wavefront shade kernel leaves those 16 bytes per path in memory.
    sh.cast = false;
    if (gid == kNoHit) return false;
    Hit its;
    fillHit(sc, gid, t, u, v, its);
    const MeshInfo &mesh = sc.meshes[its.mesh];
    const int nLights = sc.nEmitters;

    if (INTEGRATOR == NORI_INT_NORMALS) {                        // LiNormals
        L = mk(fabsf(its.sh.n.x), fabsf(its.sh.n.y), fabsf(its.sh.n.z));
        return false;
    }
    if (INTEGRATOR == NORI_INT_SIMPLE) {                         // LiSimple
        V3 d = rp.lightPos - its.p;
        float dist2 = squaredNorm(d), dist = sqrtf(dist2);
        V3 wi = d / dist;
        float c = dot(its.sh.n, wi);
        if (!(c > 0.0f)) return false;
        sh.cast = true; sh.o = its.p; sh.d = wi; sh.maxt = dist - kEpsilon;
        sh.contrib = rp.lightEnergy * (c / (4.0f * kPi * kPi * dist2));
        return false;
    }
    if (INTEGRATOR == NORI_INT_AO) {                             // LiAo
        V3 dl = squareToCosineHemisphere(rng.next2D());
        sh.cast = true; sh.o = its.p; sh.d = its.sh.toWorld(dl); sh.maxt = INFINITY; sh.contrib = mk(1, 1, 1);
        return false;
    }
    if (INTEGRATOR == NORI_INT_PRIMARY_SHADOW) {                 // LiPrimaryShadow
        if (nLights == 0) return false;
        float uu = rng.next1D();
        V2 xi = rng.next2D();
        LightSample ls;
        if (!sampleLights(sc, its.p, uu, xi, ls)) return false;
        float c = fabsf(dot(its.sh.n, ls.wi));
        sh.cast = true; sh.o = its.p; sh.d = ls.wi; sh.maxt = ls.dist - kEpsilon; sh.contrib = mk(c, c, c);
        return false;
    }
    Bsdf bsdf = meshBsdf(mesh);
    if (PLAIN && bsdf.type == kMicrofacet) bsdf.type = kDiffuse; // a no-op under the caller's promise; tells the compiler the type's range
    const V3 wiLocal = its.sh.toLocal(-ray.d);

    if (INTEGRATOR == NORI_INT_WHITTED) {                        // LiWhitted
        if (bsdf.isDiffuse()) {
            flags |= kFlagMulT;                                  // value = T * (Le + direct)
            if (mesh.emitter >= 0) {
                const EmitterInfo &em = sc.emitters[mesh.emitter];
                if (dot(its.sh.n, -ray.d) > 0.0f) L = L + mk(em.radiance[0], em.radiance[1], em.radiance[2]);
            }
            if (nLights > 0) {
                float uu = rng.next1D();
                V2 xi = rng.next2D();
                LightSample ls;
                if (sampleLights(sc, its.p, uu, xi, ls)) {
                    BsdfQuery q; q.wi = wiLocal; q.wo = its.sh.toLocal(ls.wi); q.eta = 1.0f; q.measure = kSolidAngle;
                    V3 f = bsdf.eval(q);
                    if (!isZero(f)) {
                        sh.cast = true; sh.o = its.p; sh.d = ls.wi; sh.maxt = ls.dist - kEpsilon;
                        sh.contrib = f * ls.Le * (fabsf(q.wo.z) / ls.pdf);
                    }
                }
            }
            return false;
        }
        if (!(rng.next1D() < 0.95f)) { L = mk(0, 0, 0); T = mk(0, 0, 0); flags |= kFlagMulT; return false; }
        BsdfQuery q; q.wi = wiLocal; q.eta = 1.0f; q.measure = kUnknownMeasure;
        V3 wgt = bsdf.sample(q, rng.next2D());
        if (isZero(wgt)) { L = mk(0, 0, 0); T = mk(0, 0, 0); flags |= kFlagMulT; return false; }
        T = T * wgt / 0.95f;
        ray = mkRay(its.p, its.sh.toWorld(q.wo));
        return true;
    }

    // ---- LiPath: path_mats / path_ems / path_mis ------------------------------------------
    constexpr bool useEms = INTEGRATOR == NORI_INT_PATH_EMS || INTEGRATOR == NORI_INT_PATH_MIS;
    constexpr bool useMats = INTEGRATOR == NORI_INT_PATH_MATS || INTEGRATOR == NORI_INT_PATH_MIS;
    constexpr bool mis = INTEGRATOR == NORI_INT_PATH_MIS;
    const bool discrete = flags & kFlagDiscrete;
    uint32_t depth = flags & kDepthMask;

    if (mesh.emitter >= 0) {
        const EmitterInfo &em = sc.emitters[mesh.emitter];
        if (dot(its.sh.n, -ray.d) > 0.0f) {
            V3 Le = mk(em.radiance[0], em.radiance[1], em.radiance[2]);
            if (lo) lo->Lem = Le;
            if (!isZero(Le)) {
                if (discrete || !useEms) L = L + T * Le;
                else if (useMats) {
                    V3 rayOrigin = ray.o;
                    if (lazyOrigin) { const float4 o4 = *lazyOrigin; rayOrigin = mk(o4.x, o4.y, o4.z); }
                    float pl = emitterPdfSolidAngle(em, rayOrigin, its.p, its.sh.n) / (float) nLights;
                    float wgt = bsdfPdf / (bsdfPdf + pl);
                    L = L + T * Le * wgt;
                }
            }
        }
    }
    if (rp.maxDepth > 0 && (int) depth + 1 >= rp.maxDepth) return false;

    if (useEms && bsdf.isDiffuse() && nLights > 0) {
        float uu = rng.next1D();
        V2 xi = rng.next2D();
        LightSample ls;
        if (sampleLights(sc, its.p, uu, xi, ls)) {
            if (lo) { lo->Ldirect = ls.Le; lo->lightPdf = ls.pdf; }
            BsdfQuery q; q.wi = wiLocal; q.wo = its.sh.toLocal(ls.wi); q.eta = 1.0f; q.measure = kSolidAngle;
            V3 f = bsdf.eval(q);
            if (!isZero(f)) {
                float wgt = 1.0f;
                if (mis) { float pb = bsdf.pdf(q); wgt = ls.pdf / (ls.pdf + pb); }
                float c = fabsf(q.wo.z);
                sh.cast = true; sh.o = its.p; sh.d = ls.wi; sh.maxt = ls.dist - kEpsilon;
                sh.contrib = T * f * ls.Le * (c * wgt / ls.pdf);
            }
        }
    }

    BsdfQuery q; q.wi = wiLocal; q.eta = 1.0f; q.measure = kUnknownMeasure;
    V3 wgt = bsdf.sample(q, rng.next2D());
    if (isZero(wgt)) return false;
    T = T * wgt;
    eta *= q.eta;
    const bool nowDiscrete = q.measure == kDiscrete;
    bsdfPdf = nowDiscrete ? 0.0f : bsdf.pdf(q);
    ray = mkRay(its.p, its.sh.toWorld(q.wo));
    ++depth;
    flags = (flags & ~(kDepthMask | kFlagDiscrete)) | (depth & kDepthMask) | (nowDiscrete ? kFlagDiscrete : 0u);
    if ((int) depth >= rp.rrStart) {
        float qc = fminf(maxCoeff(T) * eta * eta, rp.rrMax);
        if (rrProb) *rrProb = qc;
        if (!(rng.next1D() < qc)) return false;
        T = T / qc;
    }
    return true;
}

#ifndef NORI_SHADOW_PREFETCH
#define NORI_SHADOW_PREFETCH 8192      // any-hit stage 29.06 -> 28.55 ms at 512 spp (65 536: the same); 0 = off
#endif
constexpr uint32_t kShadowToRecord = 0x80000000u;          // shadow-queue target is a sample record, not a path slot
struct ShadowIO {
    Pool pool; float4 *recA; float *recB; size_t recCount;
    __device__ __forceinline__ bool load(uint32_t i, V3 &o, V3 &d, float &mint, float &maxt) const {
        NG_CHECK(i < pool.capacity);
#if NORI_SHADOW_PREFETCH > 0      // the queue is read front to back: the entry a few thousand places on goes to L2 now
        if (i + NORI_SHADOW_PREFETCH < pool.capacity) asm volatile("prefetch.global.L2 [%0];" :: "l"(pool.shadow + 3 * (size_t) (i + NORI_SHADOW_PREFETCH)));
#endif
        const float4 a = pool.shadow[3 * (size_t) i], b = pool.shadow[3 * (size_t) i + 1];
        o = mk(a.x, a.y, a.z); d = mk(b.x, b.y, b.z); mint = kEpsilon; maxt = a.w;
        return true;
    }
    /// Unoccluded: the light sample's contribution joins the path's radiance (one shadow ray per
    /// path and iteration, so no two lanes touch the same slot) or, when the path ended at this
    /// vertex and its slot was handed to a new camera sample, the sample record it left behind.
    __device__ __forceinline__ void store(uint32_t i, uint32_t gid, float, float, float) const {
        if (gid != kNoHit) return;
        const float4 b = pool.shadow[3 * (size_t) i + 1], k = pool.shadow[3 * (size_t) i + 2];
        const uint32_t target = __float_as_uint(b.w);
        if (target & kShadowToRecord) {
            const size_t r = target & ~kShadowToRecord;
            NG_CHECK(r < recCount);
            float4 a = recA[r];
            a.z = a.z + k.x; a.w = a.w + k.y;
            recA[r] = a;
            recB[r] = recB[r] + k.z;
        } else {
            NG_CHECK(target < pool.capacity);
            float4 r = pool.rad[target];
            r.x = r.x + k.x; r.y = r.y + k.y; r.z = r.z + k.z;
            pool.rad[target] = r;
        }
    }
};

/// Accel::rayIntersect(ray, its, true) for the queued shadow rays.
template <int MINB, bool WT = false, int QN = 0>
__global__ void __launch_bounds__(128, MINB) k_traceShadow(DeviceScene sc, Pool pool, float4 *recA, float *recB, size_t recCount, Counters *c, int fetchThreshold) {
    ShadowIO io; io.pool = pool; io.recA = recA; io.recB = recB; io.recCount = recCount;
    tracePersistent<true, WT, QN>(QN == 2 ? sc.wnodes : (QN == 1 ? reinterpret_cast<const float4 *>(sc.qnodes) : sc.nodes), sc.leafTris, c->shadowCount, &c->traceCursor[1], io, fetchThreshold, sceneGrid(sc));
}

/// Hands path slot `dst` to camera sample index g of the issue order (pass, tile, sample, pixel):
/// Sampler stream + PerspectiveCamera::sampleRay (src/main.cpp:49-56). Indices past `limit` or in
/// the padding of a border tile leave the slot empty; it draws again next iteration.
__device__ __forceinline__ void regenSlot(const DeviceScene &sc, const Pool &pool, const RenderParams &rp, uint32_t dst,
                                          unsigned long long g, unsigned long long limit, uint32_t passHint, bool wasEmpty) {
    const int W = sc.cam.width, H = sc.cam.height;
    uint32_t x = 0, y = 0, sampleK = 0, rec = 0, ring = 0, tileSlot = 0;
    int lx = 0, ly = 0, ox = 0, oy = 0;
    bool ok = false;
    if (g < limit) {
        if (rp.listPixel) {
            const uint32_t p = rp.listPixel[g];
            x = p % (uint32_t) W; y = p / (uint32_t) W; sampleK = rp.listSample[g]; rec = (uint32_t) g;
            ok = true;
        } else {
            // g lies in the hinted pass or just after it (a pass holds several pools): no 64-bit division
            uint32_t pass = passHint;
            unsigned long long j64 = g - (unsigned long long) pass * rp.passSlots;
            while (j64 >= rp.passSlots) { j64 -= rp.passSlots; ++pass; }
            const uint32_t j = (uint32_t) j64;
            const uint32_t Kp = pass == rp.numPasses - 1 ? rp.Klast : rp.K;
            const uint32_t tile = j / (1024u * Kp), r = j - tile * (1024u * Kp);
            tilePixel(r & 1023u, lx, ly);
            const uint32_t tyi = tile / (uint32_t) rp.tilesX, txi = tile - tyi * (uint32_t) rp.tilesX;
            ox = (int) txi * kTile; oy = (int) tyi * kTile;
            x = (uint32_t) (ox + lx); y = (uint32_t) (oy + ly);
            sampleK = pass * rp.K + (r >> 10); rec = j; ring = pass % (uint32_t) rp.ring;
            tileSlot = ring * rp.numTiles + tile;
            ok = tile < rp.numTiles && x < (uint32_t) W && y < (uint32_t) H;
        }
    }
    if (!ok) {
        if (!wasEmpty) pool.state[dst] = make_uint4(0, 0, 0, 0);
        if (pool.order) pool.key[dst] = (kBinEmpty << 24) | dst;      // whatever the slot held before (a path moved away by the tail compaction keeps its word): empty slots sort last
        return;
    }
    NG_CHECK(dst < pool.capacity && (rp.listPixel || (tileSlot < (uint32_t) rp.ring * rp.numTiles && rec < rp.passSlots)));
    Pcg32 rng;
    const uint32_t sampleIndex = rp.sampleFirst + sampleK;
    if (rp.streamMode == NORI_STREAM_PER_SAMPLE || rp.listPixel) {
        rng.seed((uint64_t) y * (uint64_t) W + x, (uint64_t) sampleIndex);
    } else {
        // src/independent.cpp:48-53: seed(offset.x, offset.y); renderBlock (src/main.cpp:49-66)
        // consumes 4 values per camera sample in (y, x, sample) order when Li draws none.
        const int bw = min(W - ox, kTile);
        rng.seed((uint64_t) (int64_t) ox, (uint64_t) (int64_t) oy);
        rng.advance(4ll * (((long long) ly * bw + lx) * (long long) rp.sampleCount + sampleK));
    }
    const V2 r2 = rng.next2D();
    V2 ps; ps.x = (float) x + r2.x; ps.y = (float) y + r2.y;      // src/main.cpp:52
    rng.next2D();                                                // apertureSample (main.cpp:53)
    const Ray ray = sampleRay(sc.cam, ps);
    pool.rayO[dst] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.mint);
    pool.rayD[dst] = make_float4(ray.d.x, ray.d.y, ray.d.z, ray.maxt);
    pool.thr[dst] = make_float4(1.f, 1.f, 1.f, 0.f);
    pool.rad[dst] = make_float4(0.f, 0.f, 0.f, 1.f);
    pool.pos[dst] = make_float2(ps.x, ps.y);
    pool.state[dst] = packState(rng, kFlagAlive | kFlagDiscrete | (ring << 24));
    pool.info[dst] = make_uint4(tileSlot, sampleIndex, rec, 0u);
    if (pool.order) pool.key[dst] = (kBinFresh << 24) | dst;
}

// ------------------------------------------------------------------------------------------
/// Shade + record + regenerate in one pass over the pool. A CTA owns 256 consecutive path slots
/// and sorts them in shared memory by what has to happen to them: BSDF of the surface hit (keys
/// 0-3), ray escaped (4), no live path (5), so that warps run one code path.
///   * Slots of keys 4 and 5 are free as soon as their sample is recorded: their warps draw new
///     camera samples (consecutive indices, full warps) while the other warps still shade.
///   * Slots whose path ends at a surface are recorded and regenerated in place after the shade
///     step. If the last vertex still has a shadow ray in flight, the record is left behind and
///     the shadow kernel adds the light sample there (kShadowToRecord); only what the record
///     cannot express (value = T * L, late samples) parks the slot until the next iteration.
/// One atomic per CTA each for: sample indices (twice), shadow queue space, finished count.
#ifndef NORI_SHADE_PREFETCH
#define NORI_SHADE_PREFETCH 4
#endif
#ifndef NORI_SHADE_PREFETCH_AHEAD
#define NORI_SHADE_PREFETCH_AHEAD 592       // CTAs: one wave of 4 x 148 (296 and 1184 measure the same)
#endif
#ifndef NORI_SHADE_THREADS_PER_SM
#define NORI_SHADE_THREADS_PER_SM 1024      // 64 registers per thread
#endif
template <int INTEGRATOR, int NT, bool PLAIN = false>
__global__ void __launch_bounds__(NT, NORI_SHADE_THREADS_PER_SM / NT) k_shadeRegen(DeviceScene sc, Pool pool, RenderParams rp, Counters *c) {
    constexpr int NK = kNumKeys + 1, NW = NT / 32;
    static_assert(NK * NW <= 64 && (NW == 8 || NW == 4), "the warp scan below keeps the NK x NW sort counts in two registers per lane");
    __shared__ unsigned sCnt[NW][NK], sWarpShadow[NW], sWarpRegen[NW], sShadowBase;
    __shared__ unsigned long long sIssueA, sIssueB;
    __shared__ unsigned short sPerm[NT], sRegen[NT];
    const unsigned tid = threadIdx.x, warp = tid >> 5, lane = laneId();
    // Experiment (-DNORI_SHADE_DESCENDING): CTAs walk the pool from its END: the closest-hit pass before this kernel fetched slots in ascending order, so
    // the rays and hit records it touched last are the ones still in L2; and this kernel then finishes at the low
    // slots, where the next closest-hit pass starts.
#ifdef NORI_SHADE_DESCENDING
    const uint32_t blockBase = (gridDim.x - 1u - blockIdx.x) * blockDim.x;
#else
    const uint32_t blockBase = blockIdx.x * blockDim.x;
#endif
    const unsigned long long limit = c->issueLimit;
    const uint32_t passHint = c->passHint, lateBelow = c->completedPasses * rp.K;   // samples of closed passes are late
    // ---- CTA-local counting sort by what the slot needs -----------------------------------
    int key0 = kNumKeys;
    {
#if NORI_SHADE_PREFETCH >= 4      // the first loads of a CTA that starts a few waves later: into L2
        {
            const size_t ahead = (size_t) blockBase + tid + (size_t) NORI_SHADE_PREFETCH_AHEAD * NT;
            if (ahead < pool.capacity) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(pool.state + ahead));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(pool.hit + ahead));
#if NORI_SHADE_PREFETCH >= 5      // and what its shade step reads
                asm volatile("prefetch.global.L2 [%0];" :: "l"(pool.rayD + ahead));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(pool.thr + ahead));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(pool.rad + ahead));
#endif
            }
        }
#endif
        const uint32_t flags0 = pool.state[blockBase + tid].w;
        const uint32_t gid = __float_as_uint(pool.hit[blockBase + tid].w);
#ifndef NORI_SHADE_NO_PREFETCH
        // What the shade step reads of this slot after the sort (by ANOTHER thread of this CTA, two barriers from here): asked for
        // now, so that the DRAM round trip runs under the key's dependent loads and the sort instead of after them.
        if (flags0 & kFlagAlive) {
            asm volatile("prefetch.global.L1 [%0];" :: "l"(pool.rayD + blockBase + tid));
            asm volatile("prefetch.global.L1 [%0];" :: "l"(pool.thr + blockBase + tid));
            asm volatile("prefetch.global.L1 [%0];" :: "l"(pool.rad + blockBase + tid));
#if NORI_SHADE_PREFETCH >= 2      // + what a camera sample reads when it ends (a third of the live slots per pass)
            asm volatile("prefetch.global.L1 [%0];" :: "l"(pool.info + blockBase + tid));
            asm volatile("prefetch.global.L1 [%0];" :: "l"(pool.pos + blockBase + tid));
#endif
        }
#endif
        if (flags0 & kFlagAlive) {
            if (gid == kNoHit) key0 = kKeyMiss;
            else {
                const uint4 tri = __ldg(sc.tris + gid);
#if NORI_SHADE_PREFETCH == 3      // + the triangle's vertices (fillHit): measured, much slower (profiles/r02_notes.md)
                asm volatile("prefetch.global.L1 [%0];" :: "l"(sc.vpos + tri.x));
                asm volatile("prefetch.global.L1 [%0];" :: "l"(sc.vpos + tri.y));
                asm volatile("prefetch.global.L1 [%0];" :: "l"(sc.vpos + tri.z));
#endif
                key0 = sc.meshes[tri.w].bsdfType;
            }
        }
    }
    unsigned rank = 0;
#pragma unroll
    for (int k = 0; k < NK; ++k) {
        const unsigned m = __ballot_sync(0xffffffffu, key0 == k);
        if (key0 == k) rank = __popc(m & lanemaskLt());
        if (lane == 0) sCnt[warp][k] = __popc(m);
    }
    __syncthreads();
    // Every warp scans the NK x NW (key-major, warp-minor) counts itself: entry i = key * NW + warp sits in
    // lane i & 31 of e0 (i < 32) or e1. No serial section and no second barrier.
    unsigned aBegin, alive;
    {
        constexpr unsigned NE = NK * NW;
        const unsigned i1 = lane + 32u;
        const unsigned v0 = lane < NE ? sCnt[lane % NW][lane / NW] : 0u, v1 = i1 < NE ? sCnt[i1 % NW][i1 / NW] : 0u;
        unsigned s0 = v0, s1 = v1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned t0 = __shfl_up_sync(0xffffffffu, s0, d), t1 = __shfl_up_sync(0xffffffffu, s1, d);
            if ((int) lane >= d) { s0 += t0; s1 += t1; }
        }
        s1 += __shfl_sync(0xffffffffu, s0, 31);
        const unsigned e0 = s0 - v0, e1 = s1 - v1;                 // exclusive prefixes
        auto entry = [&](unsigned i) {                               // exclusive prefix of entry i (i may differ per lane)
            const unsigned x0 = __shfl_sync(0xffffffffu, e0, i & 31u), x1 = __shfl_sync(0xffffffffu, e1, i & 31u);
            return i < 32u ? x0 : x1;
        };
        const unsigned dst = entry((unsigned) key0 * NW + warp) + rank;
        aBegin = entry(kKeyMiss * NW);                               // first slot of key kKeyMiss
        alive = entry(kNumKeys * NW);                                // first slot of key kNumKeys = live paths of this CTA
        NG_CHECK(dst < (unsigned) NT);
        sPerm[dst] = (unsigned short) tid;
    }
    // Sample indices for the slots known free up front: requested now, consumed after the shade step,
    // so that the round trip of the atomic hides behind it. (Indices past the limit are discarded by
    // regenSlot and handed out again: k_prepare clamps the counter.)
    // The last thread of the CTA is one of those free slots whenever there is any (they sort to the end).
    const unsigned nA = (unsigned) NT - aBegin;
    if (tid == 0 && alive) atomicAdd(&c->sortCount[0], alive);   // closest rays traced = live slots
    if (tid == (unsigned) NT - 1u && nA) sIssueA = atomicAdd(&c->issue, (unsigned long long) nA);
    __syncthreads();
    const unsigned local = sPerm[tid];
    const uint32_t slot = blockBase + local;

    ShadowOut sh; sh.cast = false;
    uint32_t shadowTarget = slot;
    bool fin = false, needB = false;
    V3 finL = mk(0, 0, 0);
    const uint4 st = pool.state[slot];
    uint4 info = make_uint4(0, 0, 0, 0);                     // read only when a camera sample ends
    if (tid < aBegin) {
        // ---- a surface was hit: one Li loop body ----------------------------------------------
        const float4 rd = pool.rayD[slot], h = pool.hit[slot], th = pool.thr[slot], ra = pool.rad[slot];
        Ray ray; ray.o = mk(0.f, 0.f, 0.f); ray.d = mk(rd.x, rd.y, rd.z); ray.mint = 0.f; ray.maxt = rd.w;   // origin: see lazyOrigin
        Pcg32 rng = unpackRng(st);
        V3 T = mk(th.x, th.y, th.z), L = mk(ra.x, ra.y, ra.z);
        float bsdfPdf = th.w, eta = ra.w;
        uint32_t flags = st.w;
        const bool cont = shadeVertex<INTEGRATOR, PLAIN>(sc, rp, ray, __float_as_uint(h.w), h.x, h.y, h.z, rng, T, L, eta, bsdfPdf, flags, sh, pool.rayO + slot);
        if (cont) {
            pool.rayO[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.mint);
            pool.rayD[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, ray.maxt);
            pool.thr[slot] = make_float4(T.x, T.y, T.z, bsdfPdf);
            pool.rad[slot] = make_float4(L.x, L.y, L.z, eta);
            pool.state[slot] = packState(rng, flags);
            if (pool.order) pool.key[slot] = binWord(pool, ray, slot);
        } else {
            if (pool.order) pool.key[slot] = (kBinEmpty << 24) | slot;   // ended or parked; a regeneration later in this kernel overwrites it
            info = pool.info[slot];
            const bool mulT = flags & kFlagMulT;
            const bool late = !rp.listPixel && info.y - rp.sampleFirst < lateBelow;
            const size_t r = (size_t) (st.w >> 24) * rp.passSlots + info.z;
            if (!sh.cast) { fin = true; needB = true; finL = mulT ? T * L : L; }
            else if (!mulT && !late && r < (size_t) kShadowToRecord) {
                fin = true; needB = true; finL = L;
                shadowTarget = kShadowToRecord | (uint32_t) r;
            } else {                                          // park: recorded by the next iteration's pass
                pool.thr[slot] = make_float4(T.x, T.y, T.z, bsdfPdf);
                pool.rad[slot] = make_float4(L.x, L.y, L.z, eta);
                pool.state[slot] = packState(rng, (flags & ~kFlagAlive) | kFlagHasSample);
            }
        }
    } else {
        sRegen[tid - aBegin] = (unsigned short) (local | (st.w == 0 ? 0x8000u : 0u));   // free slot: regenerated below
        if (st.w & (kFlagAlive | kFlagHasSample)) {
            info = pool.info[slot];
            // ---- ray escaped (Li returns what the path gathered) or a parked sample -----------------
            const float4 ra = pool.rad[slot];
            finL = mk(ra.x, ra.y, ra.z);
            if (st.w & kFlagMulT) { const float4 t = pool.thr[slot]; finL = mk(t.x, t.y, t.z) * finL; }
            fin = true;
        }
    }
    // ---- record the camera samples that ended (src/main.cpp:59-62 -> block.put) ---------------
    bool counted = false;
    if (fin) {
        const float2 ps = pool.pos[slot];
        const bool late = !rp.listPixel && info.y - rp.sampleFirst < lateBelow;
        if (late) {
            splatDirect(sc, rp, c, ps, finL);
            atomicAdd(&c->lateSamples, 1u);
        } else {
            const size_t r = (size_t) (st.w >> 24) * rp.passSlots + info.z;
            NG_CHECK(r < rp.recCount && (rp.listPixel || info.x < (uint32_t) rp.ring * rp.numTiles));
            rp.recA[r] = make_float4(ps.x, ps.y, finL.x, finL.y);
            rp.recB[r] = finL.z;
            counted = !rp.listPixel;
        }
    }
    if (__any_sync(0xffffffffu, counted)) {       // one atomic per (ring slot, tile) present in the warp
        const uint32_t tile = counted ? info.x : 0xffffffffu;
        const unsigned peers = __match_any_sync(0xffffffffu, tile);
        if (counted && (int) lane == __ffs(peers) - 1) atomicAdd(rp.tileDone + tile, (unsigned) __popc(peers));
    }

    // ---- compaction: shadow queue entries, slots freed by the shade step ----------------------
    const unsigned mShadow = __ballot_sync(0xffffffffu, sh.cast), mB = __ballot_sync(0xffffffffu, needB);
    if (lane == 0) { sWarpShadow[warp] = __popc(mShadow); sWarpRegen[warp] = __popc(mB); }
    const int nfin = __syncthreads_count(fin);
    unsigned preS = 0, preR = 0, totS = 0, totR = 0;
#pragma unroll
    for (unsigned w = 0; w < (unsigned) NW; ++w) {
        const unsigned a = sWarpShadow[w], b = sWarpRegen[w];
        if (w < warp) { preS += a; preR += b; }
        totS += a; totR += b;
    }
    if (tid == 0) {
        sShadowBase = totS ? atomicAdd(&c->shadowCount, totS) : 0u;
        sIssueB = totR ? atomicAdd(&c->issue, (unsigned long long) totR) : ~0ull;
        if (nfin) atomicAdd(&c->finished, (unsigned long long) nfin);
    }
    if (needB) sRegen[nA + preR + __popc(mB & lanemaskLt())] = (unsigned short) local;
    __syncthreads();
    if (sh.cast) {
        const uint32_t q = sShadowBase + preS + __popc(mShadow & lanemaskLt());
        NG_CHECK(q < pool.capacity);
        float4 *e = pool.shadow + 3 * (size_t) q;
        e[0] = make_float4(sh.o.x, sh.o.y, sh.o.z, sh.maxt);
        e[1] = make_float4(sh.d.x, sh.d.y, sh.d.z, __uint_as_float(shadowTarget));
        e[2] = make_float4(sh.contrib.x, sh.contrib.y, sh.contrib.z, 0.f);
    }
    // ---- every free slot of the CTA draws its next camera sample: one dense list, full warps ----
    if (tid < nA + totR) {
        const unsigned e = sRegen[tid];
        const unsigned long long g = tid < nA ? sIssueA + tid : (sIssueB == ~0ull ? ~0ull : sIssueB + (tid - nA));
        regenSlot(sc, pool, rp, blockBase + (e & 0x7fffu), g, limit, passHint, (e & 0x8000u) != 0u);
    }
}

// ------------------------------------------------------------------------------------------
/// ImageBlock::put(pos, value) (src/block.cpp:74-103) for every sample of a finished tile, into a
/// shared-memory ImageBlock of (tile + 2*border)^2 cells kept as four planes (R, G, B, W), then
/// `result.put(block)` (src/block.cpp:105-114) into the full-frame accumulator.
constexpr int kSplatSide = kTile + 2 * kMaxBorder;
constexpr int kSplatSplit = 8;      // CTAs sharing one tile (each takes the samples k = z, z + Z, ...)
/// Which (ring slot, tile) pairs are splatted in this iteration: complete tiles, and every pending
/// tile of a pass that is being closed early. One thread per pair; the splat kernels then run
/// only over this list with a grid sized to the machine.
__global__ void k_splatPlan(DeviceScene sc, RenderParams rp, Counters *c) {
    const uint32_t ti = blockIdx.x * blockDim.x + threadIdx.x;
    bool ready = false;
    if (ti < (uint32_t) rp.ring * rp.numTiles) {
        const uint32_t slot = ti / rp.numTiles, tile = ti - slot * rp.numTiles;
        const uint32_t pass = c->slotPass[slot];
        if (pass < rp.numPasses && rp.tileGen[ti] == pass / (uint32_t) rp.ring) {
            const int ox = (int) (tile % (uint32_t) rp.tilesX) * kTile, oy = (int) (tile / (uint32_t) rp.tilesX) * kTile;
            const int tw = min(sc.cam.width - ox, kTile), th = min(sc.cam.height - oy, kTile);
            const uint32_t Kp = pass == rp.numPasses - 1 ? rp.Klast : rp.K;
            ready = pass == c->forcePass || rp.tileDone[ti] == (uint32_t) (tw * th) * Kp;
        }
    }
    const uint32_t q = warpAppend(&c->readyCount, ready);
    NG_CHECK(!ready || q < (uint32_t) rp.ring * rp.numTiles);
    if (ready) rp.readyTiles[q] = ti;
}

__global__ void __launch_bounds__(256) k_splat(DeviceScene sc, RenderParams rp, Counters *c) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");     // see k_splatPixel
    __shared__ float plane[4][kSplatSide * kSplatSide];
    __shared__ float sTable[33];
    __shared__ unsigned sDropped;
    const uint32_t Z = min((uint32_t) kSplatSplit, rp.K);
    for (uint32_t work = blockIdx.x; work < c->readyCount * Z; work += gridDim.x) {
    const uint32_t tiq = rp.readyTiles[work / Z], z = work % Z;
    NG_CHECK(tiq < (uint32_t) rp.ring * rp.numTiles);
    const uint32_t slot = tiq / rp.numTiles, tile = tiq - slot * rp.numTiles;
    const uint32_t pass = c->slotPass[slot];
    const int W = sc.cam.width, H = sc.cam.height, b = sc.filter.border;
    const int ox = (int) (tile % (uint32_t) rp.tilesX) * kTile, oy = (int) (tile / (uint32_t) rp.tilesX) * kTile;
    const int tw = min(W - ox, kTile), th = min(H - oy, kTile);
    const uint32_t Kp = pass == rp.numPasses - 1 ? rp.Klast : rp.K;
    const uint32_t ti = slot * rp.numTiles + tile;
    const int cols = tw + 2 * b, rows = th + 2 * b;
    __syncthreads();                                                            // shared tile is reused across work items
    if (z < Kp) {
        for (int k = threadIdx.x; k < cols * rows; k += blockDim.x) { plane[0][k] = 0.f; plane[1][k] = 0.f; plane[2][k] = 0.f; plane[3][k] = 0.f; }
        if (threadIdx.x < 33) sTable[threadIdx.x] = sc.filter.table[threadIdx.x];
        if (threadIdx.x == 0) sDropped = 0;
        __syncthreads();
        const float radius = sc.filter.radius, lookup = sc.filter.lookupFactor;
        const size_t base = (size_t) slot * rp.passSlots + (size_t) tile * 1024u * Kp;
        for (uint32_t k = z; k < Kp; k += Z) {
            for (uint32_t pix = threadIdx.x; pix < 1024u; pix += blockDim.x) {
                int lx, ly; tilePixel(pix, lx, ly);
                if (lx >= tw || ly >= th) continue;
                const size_t r = base + k * 1024u + pix;
                NG_CHECK(r < rp.recCount);
                const float4 a = rp.recA[r];
                if (isnan(a.x)) continue;                                        // sample still in flight (forced close)
                const V3 val = mk(a.z, a.w, rp.recB[r]);
                rp.recA[r].x = __int_as_float(0x7fc00000);                       // consumed: free for pass + ring
                if (!isValidColor(val)) { atomicAdd(&sDropped, 1u); continue; }  // block.cpp:75-79
                const float px = a.x - 0.5f - (float) (ox - b), py = a.y - 0.5f - (float) (oy - b);   // block.cpp:82-85
                const int x0 = max((int) ceilf(px - radius), 0), y0 = max((int) ceilf(py - radius), 0);
                const int x1 = min((int) floorf(px + radius), cols - 1), y1 = min((int) floorf(py + radius), rows - 1);
                for (int y = y0; y <= y1; ++y) {
                    const float wy = sTable[(int) (fabsf((float) y - py) * lookup)];
                    for (int x = x0; x <= x1; ++x) {
                        const float wx = sTable[(int) (fabsf((float) x - px) * lookup)];
                        const int cell = y * cols + x;
                        NG_CHECK(cell >= 0 && cell < kSplatSide * kSplatSide && (int) (fabsf((float) x - px) * lookup) <= 32 && (int) (fabsf((float) y - py) * lookup) <= 32);
                        atomicAdd(&plane[0][cell], val.x * wx * wy);             // block.cpp:100-102
                        atomicAdd(&plane[1][cell], val.y * wx * wy);
                        atomicAdd(&plane[2][cell], val.z * wx * wy);
                        atomicAdd(&plane[3][cell], 1.0f * wx * wy);
                    }
                }
            }
        }
        __syncthreads();
        // result.put(block): tile cell (x, y) is full-frame cell (ox + x, oy + y); borders of
        // neighbouring tiles overlap and Z CTAs share the tile, hence atomics.
        for (int k = threadIdx.x; k < cols * rows; k += blockDim.x) {
            const int x = k % cols, y = k / cols;
            const float4 v = make_float4(plane[0][k], plane[1][k], plane[2][k], plane[3][k]);
            if (v.w == 0.f && v.x == 0.f && v.y == 0.f && v.z == 0.f) continue;
            NG_CHECK(oy + y < rp.blockRows && ox + x < rp.blockCols);
            float4 *dst = reinterpret_cast<float4 *>(rp.block) + (size_t) (oy + y) * rp.blockCols + (ox + x);
            atomicAdd(dst, v);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (z < Kp && sDropped) atomicAdd(&c->dropped, (unsigned long long) sDropped);
        __threadfence();
        if (atomicAdd(rp.tileArrive + ti, 1u) == Z - 1) {                        // last CTA of this tile
            rp.tileArrive[ti] = 0;
            rp.tileDone[ti] = 0;
            rp.tileGen[ti] = pass / (uint32_t) rp.ring + 1;
            __threadfence();
            atomicAdd(&c->tilesSplatted[slot], 1u);
        }
    }
    }
}

/// The same stage for filters whose border is at most 2 pixels (all of the reference's: gaussian
/// and mitchell radius 2, tent 1, box 0.5): one thread per PIXEL of a 16x16 quadrant of the tile
/// walks the K samples of its pixel and accumulates their (2B+1)^2 footprint in registers; only
/// then does it touch the shared-memory ImageBlock (one atomic per cell and channel per pass
/// instead of one per sample), which the CTA finally adds to the full-frame block.
/// A tile is shared by 4 quadrants x kSplatSlices sample slices (slice z takes the samples k = z, z + S, ... of every pixel).
/// Measured (profiles/r02_notes.md): more slices make the step SLOWER (closest + splat stage 202.0 / 204.3 / 207.3 / 213.4 ms for
/// S = 1 / 2 / 4 / 8). A splat CTA holds 248 registers x 256 threads = the whole register file of its SM, so no traversal CTA can
/// share the SM with it: what the splat costs the overlapped traversal is the SM-time it occupies, and spreading the same work
/// over more SMs for a shorter time only adds the per-CTA set-up and flush. S = 1 stays.
#ifndef NORI_SPLAT_SLICES
#define NORI_SPLAT_SLICES 1
#endif
constexpr uint32_t kSplatSlices = NORI_SPLAT_SLICES;
#ifndef NORI_SPLAT_THREADS
#define NORI_SPLAT_THREADS 256
#endif
constexpr int kSplatThreads = NORI_SPLAT_THREADS;
template <int B, int NT>
__global__ void __launch_bounds__(NT) k_splatPixel(DeviceScene sc, RenderParams rp, Counters *c) {
    // Programmatic dependent launch: the closest-hit pass queued behind this kernel does not read
    // anything the splat writes, so it may start filling the SMs this grid leaves free right away.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    // NT threads = a 16 x QH block of the tile (NT = 256: a quadrant; NT = 128: half a quadrant, 8 CTAs per tile, half the
    // registers per CTA so that traversal CTAs can share the SM with it)
    constexpr int WN = 2 * B + 1, Q = 16, QH = NT / 16, SIDE = Q + 2 * B, SIDEH = QH + 2 * B;
    constexpr uint32_t NSUB = 1024u / NT;
    static_assert(NT == 256 || NT == 128, "one thread per pixel of a 16 x 16 or 16 x 8 block");
    __shared__ float plane[4][SIDE * SIDEH];
    __shared__ float sTable[33];
    __shared__ unsigned sDropped;
    // Per-thread ring of sample records, filled with cp.async kSplatAhead - 1 samples ahead of the one being splatted: the kernel
    // runs 8 warps per SM (248 registers) and used to wait for every record load in turn (20 bytes per sample, latency-bound).
    // Each thread reads back only what it copied itself: no barrier, just its own async-copy groups. [stage][thread]: conflict-free.
#ifndef NORI_SPLAT_AHEAD
#define NORI_SPLAT_AHEAD 4
#endif
    constexpr uint32_t kSplatAhead = NORI_SPLAT_AHEAD;
    static_assert((kSplatAhead & (kSplatAhead - 1)) == 0, "ring size must be a power of two");
    __shared__ float4 sRecA[kSplatAhead > 1 ? kSplatAhead : 1][NT];
    __shared__ float sRecB[kSplatAhead > 1 ? kSplatAhead : 1][NT];
    for (uint32_t work = blockIdx.x; work < c->readyCount * NSUB * kSplatSlices; work += gridDim.x) {
    const uint32_t tiq = rp.readyTiles[work / (NSUB * kSplatSlices)], z = (work / kSplatSlices) % NSUB, slice = work % kSplatSlices;
    NG_CHECK(tiq < (uint32_t) rp.ring * rp.numTiles);
    const uint32_t slot = tiq / rp.numTiles, tile = tiq - slot * rp.numTiles;
    const uint32_t pass = c->slotPass[slot];
    const int W = sc.cam.width, H = sc.cam.height;
    const int ox = (int) (tile % (uint32_t) rp.tilesX) * kTile, oy = (int) (tile / (uint32_t) rp.tilesX) * kTile;
    const int tw = min(W - ox, kTile), th = min(H - oy, kTile);
    const uint32_t Kp = pass == rp.numPasses - 1 ? rp.Klast : rp.K;
    const uint32_t ti = slot * rp.numTiles + tile;
    __syncthreads();                                                            // shared tile is reused across work items
    const int qx = (int) (z & 1u), qy = (int) (z >> 1);
    if (qx * Q < tw && qy * QH < th) {
        for (int k = threadIdx.x; k < SIDE * SIDEH; k += blockDim.x) { plane[0][k] = 0.f; plane[1][k] = 0.f; plane[2][k] = 0.f; plane[3][k] = 0.f; }
        if (threadIdx.x < 33) sTable[threadIdx.x] = sc.filter.table[threadIdx.x];
        if (threadIdx.x == 0) sDropped = 0;
        __syncthreads();
        // thread -> record slot of its pixel (inverse of tilePixel: warps are 8x4 pixel patches)
        const uint32_t wi = threadIdx.x >> 5;
        const uint32_t wq = (wi & 1u) + 2u * (uint32_t) qx + 4u * ((wi >> 1) + (uint32_t) (QH / 4) * (uint32_t) qy);   // 8 x 4 pixel patches: 4 across, 8 down the tile
        const uint32_t pix = wq * 32u + (threadIdx.x & 31u);
        int lx, ly; tilePixel(pix, lx, ly);
        const int tx = lx - qx * Q, ty = ly - qy * QH;              // position inside the quadrant
        float acc[WN][WN][4];
#pragma unroll
        for (int i = 0; i < WN; ++i)
#pragma unroll
            for (int j = 0; j < WN; ++j) { acc[i][j][0] = 0.f; acc[i][j][1] = 0.f; acc[i][j][2] = 0.f; acc[i][j][3] = 0.f; }
        const bool inside = lx < tw && ly < th;
        if (inside) {
            const float radius = sc.filter.radius, lookup = sc.filter.lookupFactor;
            const int cols = tw + 2 * B, rows = th + 2 * B;
            const size_t base = (size_t) slot * rp.passSlots + (size_t) tile * 1024u * Kp + pix;
            unsigned dropped = 0;
#ifdef NORI_SPLAT_SKIP          // timing experiment: what the stage costs without the sample loop (images are wrong)
            const uint32_t nK = 0u;
#else
            const uint32_t nK = Kp > slice ? (Kp - slice + kSplatSlices - 1u) / kSplatSlices : 0u;      // samples of this slice: k = slice + j * S
#endif
            auto fetch = [&](uint32_t j) {
                const size_t r = base + (size_t) (slice + j * kSplatSlices) * 1024u;
                __pipeline_memcpy_async(&sRecA[j & (kSplatAhead - 1u)][threadIdx.x], rp.recA + r, 16);
                __pipeline_memcpy_async(&sRecB[j & (kSplatAhead - 1u)][threadIdx.x], rp.recB + r, 4);
            };
            if (kSplatAhead > 1) {
                for (uint32_t j = 0; j + 1u < kSplatAhead; ++j) { if (j < nK) fetch(j); __pipeline_commit(); }
            }
#pragma unroll 2
            for (uint32_t j = 0; j < nK; ++j) {
                const uint32_t k = slice + j * kSplatSlices;
                const size_t r = base + (size_t) k * 1024u;
                NG_CHECK(r < rp.recCount);
                float4 a; float recb;
                if (kSplatAhead > 1) {
                    if (j + kSplatAhead - 1u < nK) fetch(j + kSplatAhead - 1u);
                    __pipeline_commit();
                    __pipeline_wait_prior(kSplatAhead - 1u);                     // the group that carried sample j has landed
                    a = sRecA[j & (kSplatAhead - 1u)][threadIdx.x]; recb = sRecB[j & (kSplatAhead - 1u)][threadIdx.x];
                } else { a = rp.recA[r]; recb = rp.recB[r]; }
                if (isnan(a.x)) continue;                                        // still in flight (forced close)
                const V3 val = mk(a.z, a.w, recb);
                rp.recA[r].x = __int_as_float(0x7fc00000);                       // consumed
                if (!isValidColor(val)) { ++dropped; continue; }                 // block.cpp:75-79
                const float px = a.x - 0.5f - (float) (ox - B), py = a.y - 0.5f - (float) (oy - B);   // block.cpp:82-85
                const int x0 = max((int) ceilf(px - radius), 0), y0 = max((int) ceilf(py - radius), 0);
                const int x1 = min((int) floorf(px + radius), cols - 1), y1 = min((int) floorf(py + radius), rows - 1);
                float wx[WN], wy[WN];
#pragma unroll
                for (int j = 0; j < WN; ++j) {
                    const int x = lx + j, y = ly + j;                            // window cell = pixel + B - B + j
                    wx[j] = (x >= x0 && x <= x1) ? sTable[min((int) (fabsf((float) x - px) * lookup), 32)] : 0.f;
                    wy[j] = (y >= y0 && y <= y1) ? sTable[min((int) (fabsf((float) y - py) * lookup), 32)] : 0.f;
                }
#pragma unroll
                for (int i = 0; i < WN; ++i)
#pragma unroll
                    for (int j = 0; j < WN; ++j) {
                        acc[i][j][0] += val.x * wx[j] * wy[i];                   // block.cpp:100-102
                        acc[i][j][1] += val.y * wx[j] * wy[i];
                        acc[i][j][2] += val.z * wx[j] * wy[i];
                        acc[i][j][3] += 1.0f * wx[j] * wy[i];
                    }
            }
            if (dropped) atomicAdd(&sDropped, dropped);
#pragma unroll
            for (int i = 0; i < WN; ++i)
#pragma unroll
                for (int j = 0; j < WN; ++j) {
                    const int cell = (ty + i) * SIDE + tx + j;
                    NG_CHECK(cell >= 0 && cell < SIDE * SIDEH);
                    if (acc[i][j][3] != 0.f) {
                        atomicAdd(&plane[0][cell], acc[i][j][0]); atomicAdd(&plane[1][cell], acc[i][j][1]);
                        atomicAdd(&plane[2][cell], acc[i][j][2]); atomicAdd(&plane[3][cell], acc[i][j][3]);
                    }
                }
        }
        __syncthreads();
        // block cell (x, y) is full-frame cell (ox + 16 qx + x, oy + QH qy + y)
        for (int k = threadIdx.x; k < SIDE * SIDEH; k += blockDim.x) {
            const float4 v = make_float4(plane[0][k], plane[1][k], plane[2][k], plane[3][k]);
            if (v.w == 0.f) continue;
            const int x = k % SIDE, y = k / SIDE;
            NG_CHECK(oy + qy * QH + y < rp.blockRows && ox + qx * Q + x < rp.blockCols);
            float4 *dst = reinterpret_cast<float4 *>(rp.block) + (size_t) (oy + qy * QH + y) * rp.blockCols + (ox + qx * Q + x);
            atomicAdd(dst, v);
        }
        __syncthreads();
        if (threadIdx.x == 0 && sDropped) atomicAdd(&c->dropped, (unsigned long long) sDropped);
    }
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(rp.tileArrive + ti, 1u) == NSUB * kSplatSlices - 1u) {     // last block and slice of this tile
            rp.tileArrive[ti] = 0;
            rp.tileDone[ti] = 0;
            rp.tileGen[ti] = pass / (uint32_t) rp.ring + 1;
            __threadfence();
            atomicAdd(&c->tilesSplatted[slot], 1u);
        }
    }
    }
}

// ------------------------------------------------------------------------------------------
template <int I> static void launchShade(const DeviceScene &sc, const Pool &pool, const RenderParams &rp, Counters *c, cudaStream_t st) {
#ifdef NORI_GPU_SWEEP
    static const bool small = getenv("NORI_GPU_SHADE_NT") && atoi(getenv("NORI_GPU_SHADE_NT")) == 128;   // sweep: profiles/r01_pipeline.md
    if (small) { k_shadeRegen<I, 128><<<pool.capacity / 128, 128, 0, st>>>(sc, pool, rp, c); return; }
#endif
    // scenes without microfacet surfaces (the Cornell box of the headline config: diffuse walls, a mirror and a glass sphere): the path
    // integrators' kernel without the Beckmann code. Same results bit for bit (tests/test_gpu_render.py: test_plain_bsdf_kernel_...).
    constexpr bool kPath = I == NORI_INT_PATH_MATS || I == NORI_INT_PATH_EMS || I == NORI_INT_PATH_MIS;
    if constexpr (kPath) {
        if (sc.plainBsdfs) { k_shadeRegen<I, 256, true><<<pool.capacity / 256, 256, 0, st>>>(sc, pool, rp, c); return; }
    }
    k_shadeRegen<I, 256><<<pool.capacity / 256, 256, 0, st>>>(sc, pool, rp, c);
}

// ------------------------------------------------------------------------------------------
/// The last paths of a frame: with Russian roulette capped near 1 a handful of paths live for hundreds
/// of bounces, and a wavefront iteration (six launches) per bounce is launch-bound long before that.
/// Once the pool has drained to a few thousand paths, ONE launch follows every remaining path to its
/// end, one thread per path: the same traverse() / shadeVertex() steps in the same order, so each
/// sample gets the value the wavefront would have produced.
template <int INTEGRATOR>
__global__ void __launch_bounds__(128) k_finishTail(DeviceScene sc, Pool pool, RenderParams rp, Counters *c) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= pool.capacity) return;
    const uint4 st = pool.state[slot];
    if (!(st.w & (kFlagAlive | kFlagHasSample))) return;
    const uint4 info = pool.info[slot];
    const float4 th = pool.thr[slot], ra = pool.rad[slot];
    V3 T = mk(th.x, th.y, th.z), L = mk(ra.x, ra.y, ra.z);
    uint32_t flags = st.w;
    unsigned closest = 0, shadows = 0;
    if (flags & kFlagAlive) {
        const float4 ro = pool.rayO[slot], rd = pool.rayD[slot];
        Ray ray; ray.o = mk(ro.x, ro.y, ro.z); ray.d = mk(rd.x, rd.y, rd.z); ray.mint = ro.w; ray.maxt = rd.w;
        Pcg32 rng = unpackRng(st);
        float bsdfPdf = th.w, eta = ra.w;
        for (;;) {
            float t, u, v;
            const uint32_t gid = traverse<false>(sc.nodes, sc.leafTris, ray.o, ray.d, ray.mint, ray.maxt, t, u, v, nullptr, sc.isectMode == NORI_ISECT_WATERTIGHT);
            ++closest;
            ShadowOut sh;
            const bool cont = shadeVertex<INTEGRATOR>(sc, rp, ray, gid, t, u, v, rng, T, L, eta, bsdfPdf, flags, sh);
            if (sh.cast) {
                float st, su, sv;
                ++shadows;
                if (traverse<true>(sc.nodes, sc.leafTris, sh.o, sh.d, kEpsilon, sh.maxt, st, su, sv, nullptr, sc.isectMode == NORI_ISECT_WATERTIGHT) == kNoHit) L = L + sh.contrib;
            }
            if (!cont) break;
        }
    }
    const V3 finL = (flags & kFlagMulT) ? T * L : L;
    const float2 ps = pool.pos[slot];
    const bool late = !rp.listPixel && info.y - rp.sampleFirst < c->completedPasses * rp.K;
    if (late) {
        splatDirect(sc, rp, c, ps, finL);
        atomicAdd(&c->lateSamples, 1u);
    } else {
        const size_t r = (size_t) (st.w >> 24) * rp.passSlots + info.z;
        NG_CHECK(r < rp.recCount && (rp.listPixel || info.x < (uint32_t) rp.ring * rp.numTiles));
        rp.recA[r] = make_float4(ps.x, ps.y, finL.x, finL.y);
        rp.recB[r] = finL.z;
        if (!rp.listPixel) atomicAdd(rp.tileDone + info.x, 1u);
    }
    pool.state[slot] = make_uint4(0, 0, 0, 0);
    atomicAdd(&c->finished, 1ull);
    if (closest) { atomicAdd(&c->closestRays, (unsigned long long) closest); atomicAdd(&c->segments, (unsigned long long) closest); }
    if (shadows) atomicAdd(&c->shadowRays, (unsigned long long) shadows);
}

int enqueueFinishTail(const DeviceScene &sc, const Pool &pool, const RenderParams &rp, Counters *c, cudaStream_t st) {
    const unsigned g = (pool.capacity + 127u) / 128u;
    switch (rp.integrator) {
    case NORI_INT_NORMALS: k_finishTail<NORI_INT_NORMALS><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    case NORI_INT_PATH_MATS: k_finishTail<NORI_INT_PATH_MATS><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    case NORI_INT_PATH_EMS: k_finishTail<NORI_INT_PATH_EMS><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    case NORI_INT_PATH_MIS: k_finishTail<NORI_INT_PATH_MIS><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    case NORI_INT_WHITTED: k_finishTail<NORI_INT_WHITTED><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    case NORI_INT_PRIMARY_SHADOW: k_finishTail<NORI_INT_PRIMARY_SHADOW><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    case NORI_INT_SIMPLE: k_finishTail<NORI_INT_SIMPLE><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    default: k_finishTail<NORI_INT_AO><<<g, 128, 0, st>>>(sc, pool, rp, c); break;
    }
    return 1;
}

// ------------------------------------------------------------------------------------------
/// Integrator::Li(scene, sampler, ray) (include/nori/integrator.h:54) for n given rays, each with the state of the Sampler
/// (pcg32: state, inc) it draws from; the advanced state is written back, so the caller's sampler continues exactly where
/// a CPU Li would have left it (src/ttest.cpp:168-175 draws the next ray from the same sampler).
template <int INTEGRATOR>
__global__ void __launch_bounds__(128) k_li(DeviceScene sc, RenderParams rp, int n, const float4 *__restrict__ rays, unsigned long long *__restrict__ rngState, float *__restrict__ rgb) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 ro = rays[2 * (size_t) i], rd = rays[2 * (size_t) i + 1];
    Ray ray; ray.o = mk(ro.x, ro.y, ro.z); ray.d = mk(rd.x, rd.y, rd.z); ray.mint = ro.w; ray.maxt = rd.w;
    Pcg32 rng; rng.state = rngState[2 * (size_t) i]; rng.inc = rngState[2 * (size_t) i + 1];
    V3 T = mk(1.f, 1.f, 1.f), L = mk(0.f, 0.f, 0.f);
    float bsdfPdf = 0.f, eta = 1.f;
    uint32_t flags = kFlagAlive | kFlagDiscrete;
    for (;;) {
        float t, u, v;
        const uint32_t gid = traverse<false>(sc.nodes, sc.leafTris, ray.o, ray.d, ray.mint, ray.maxt, t, u, v, nullptr, sc.isectMode == NORI_ISECT_WATERTIGHT);
        ShadowOut sh;
        const bool cont = shadeVertex<INTEGRATOR>(sc, rp, ray, gid, t, u, v, rng, T, L, eta, bsdfPdf, flags, sh);
        if (sh.cast) {
            float st, su, sv;
            if (traverse<true>(sc.nodes, sc.leafTris, sh.o, sh.d, kEpsilon, sh.maxt, st, su, sv, nullptr, sc.isectMode == NORI_ISECT_WATERTIGHT) == kNoHit) L = L + sh.contrib;
        }
        if (!cont) break;
    }
    const V3 finL = (flags & kFlagMulT) ? T * L : L;
    rgb[3 * (size_t) i] = finL.x; rgb[3 * (size_t) i + 1] = finL.y; rgb[3 * (size_t) i + 2] = finL.z;
    rngState[2 * (size_t) i] = rng.state; rngState[2 * (size_t) i + 1] = rng.inc;
}

void enqueueLi(const DeviceScene &sc, const RenderParams &rp, int n, const float4 *rays, unsigned long long *rngState, float *rgb, cudaStream_t st) {
    const unsigned g = (unsigned) ((n + 127) / 128);
    switch (rp.integrator) {
    case NORI_INT_NORMALS: k_li<NORI_INT_NORMALS><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    case NORI_INT_PATH_MATS: k_li<NORI_INT_PATH_MATS><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    case NORI_INT_PATH_EMS: k_li<NORI_INT_PATH_EMS><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    case NORI_INT_PATH_MIS: k_li<NORI_INT_PATH_MIS><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    case NORI_INT_WHITTED: k_li<NORI_INT_WHITTED><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    case NORI_INT_PRIMARY_SHADOW: k_li<NORI_INT_PRIMARY_SHADOW><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    case NORI_INT_SIMPLE: k_li<NORI_INT_SIMPLE><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    default: k_li<NORI_INT_AO><<<g, 128, 0, st>>>(sc, rp, n, rays, rngState, rgb); break;
    }
}

// ------------------------------------------------------------------------------------------
/// Path vertices for the path-graph post-process (include/nori_gpu.h: NoriGpuSPoint = the reference's SPoint,
/// include/nori/shadingPoint.h:125-147). One thread follows one camera sample's path with the same traverse() /
/// shadeVertex() steps and random numbers as the wavefront, so these are the vertices nori_gpu_render visits.
/// WRITE = false: counts[g] = surface vertices of path g. WRITE = true: records at offsets[g] + v (path-contiguous).
template <int INTEGRATOR, bool WRITE>
__global__ void __launch_bounds__(128) k_pathVertices(DeviceScene sc, RenderParams rp, uint32_t spp, unsigned long long nPaths,
                                                      unsigned int *__restrict__ counts, const unsigned long long *__restrict__ offsets, NoriGpuSPoint *__restrict__ out,
                                                      NoriGpuLPoint *__restrict__ lights) {
    const unsigned long long g = (unsigned long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nPaths) return;
    const uint32_t pixel = (uint32_t) (g / spp), k = (uint32_t) (g % spp);
    const uint32_t x = pixel % (uint32_t) sc.cam.width, y = pixel / (uint32_t) sc.cam.width;
    Pcg32 rng;
    rng.seed((uint64_t) y * (uint64_t) sc.cam.width + x, (uint64_t) (rp.sampleFirst + k));
    const V2 r2 = rng.next2D();
    V2 ps; ps.x = (float) x + r2.x; ps.y = (float) y + r2.y;
    rng.next2D();
    Ray ray = sampleRay(sc.cam, ps);
    V3 T = mk(1.f, 1.f, 1.f), L = mk(0.f, 0.f, 0.f);
    float bsdfPdf = 0.f, eta = 1.f;
    uint32_t flags = kFlagAlive | kFlagDiscrete;
    unsigned int nv = 0;
    const unsigned long long base = WRITE ? offsets[g] : 0ull;
    for (;;) {
        float t, u, v;
        const uint32_t gid = traverse<false>(sc.nodes, sc.leafTris, ray.o, ray.d, ray.mint, ray.maxt, t, u, v, nullptr, sc.isectMode == NORI_ISECT_WATERTIGHT);
        if (gid == kNoHit) break;
        NoriGpuSPoint *rec = nullptr;
        if (WRITE) {
            if (nv > 0) out[base + nv - 1].nidx = (int32_t) (base + nv);   // the previous vertex has a successor: its record index (pbsdf.cu:1749 compares nidx with an index)
            rec = out + base + nv;
            Hit its; fillHit(sc, gid, t, u, v, its);
            const MeshInfo &m = sc.meshes[its.mesh];
            NoriGpuSPoint r;
            r.pos[0] = its.p.x; r.pos[1] = its.p.y; r.pos[2] = its.p.z;
            r.wi[0] = r.wi[1] = r.wi[2] = 0.f; r.wi_d[0] = r.wi_d[1] = r.wi_d[2] = 0.f;
            r.wo[0] = -ray.d.x; r.wo[1] = -ray.d.y; r.wo[2] = -ray.d.z;
            r.shN[0] = its.sh.n.x; r.shN[1] = its.sh.n.y; r.shN[2] = its.sh.n.z;
            r.geoN[0] = its.geo.n.x; r.geoN[1] = its.geo.n.y; r.geoN[2] = its.geo.n.z;
            const bool dif = m.bsdfType == NORI_BSDF_DIFFUSE, mic = m.bsdfType == NORI_BSDF_MICROFACET;
            const float spec = mic ? m.ks : (dif ? 0.f : 1.f), ratio = (m.bsdfType == NORI_BSDF_DIELECTRIC || mic) ? m.intIOR / m.extIOR : 0.f;
            for (int c = 0; c < 3; ++c) {
                r.diffuse[c] = (dif || mic) ? m.albedo[c] : 0.f; r.specular[c] = spec;
                r.eLi[c] = 0.f; r.eLd[c] = 0.f; r.eta[c] = ratio; r.k[c] = 0.f;
            }
            r.roughness = mic ? m.alpha : 0.f; r.pdf = 0.f; r.rrpdf = 0.f; r.nidx = 0; r.groupIdx = (int32_t) g;
            r.bsdf_type = dif ? 'd' : (m.bsdfType == NORI_BSDF_MIRROR ? 'c' : (mic ? 'o' : 't'));
            *rec = r;
        }
        ++nv;
        ShadowOut sh;
        float rr = 1.0f;
        const V3 Tb = T, Lb = L;                                     // throughput up to this vertex, radiance credited before it
        LightOut lo; lo.Lem = mk(0.f, 0.f, 0.f); lo.Ldirect = mk(0.f, 0.f, 0.f); lo.lightPdf = 0.f;
        const bool cont = shadeVertex<INTEGRATOR>(sc, rp, ray, gid, t, u, v, rng, T, L, eta, bsdfPdf, flags, sh, nullptr, &rr, WRITE ? &lo : nullptr);
        if (WRITE) {
            bool lit = false;
            if (sh.cast) {
                rec->wi_d[0] = sh.d.x; rec->wi_d[1] = sh.d.y; rec->wi_d[2] = sh.d.z;
                float st, su, sv;                                    // the light sample counts if its shadow ray is free (as k_li)
                lit = traverse<true>(sc.nodes, sc.leafTris, sh.o, sh.d, kEpsilon, sh.maxt, st, su, sv, nullptr, sc.isectMode == NORI_ISECT_WATERTIGHT) == kNoHit;
                if (lit) L = L + sh.contrib;
            }
            if (lights) {                                            // LPoint (include/nori/shadingPoint.h:149-156) of this vertex
                NoriGpuLPoint lp;
                const V3 Ld = lit ? lo.Ldirect : mk(0.f, 0.f, 0.f);
                lp.L_directsample[0] = Ld.x; lp.L_directsample[1] = Ld.y; lp.L_directsample[2] = Ld.z;
                lp.L_bsdfsample[0] = lp.L_bsdfsample[1] = lp.L_bsdfsample[2] = 0.f;      // known at the next vertex (below)
                lp.L_em[0] = lo.Lem.x; lp.L_em[1] = lo.Lem.y; lp.L_em[2] = lo.Lem.z;
                lp.lightpdf = sh.cast ? lo.lightPdf : 0.f; lp.bsdfpdf = 0.f;
                lights[base + nv - 1] = lp;
                if (nv > 1) { NoriGpuLPoint *prev = lights + base + nv - 2; prev->L_bsdfsample[0] = lo.Lem.x; prev->L_bsdfsample[1] = lo.Lem.y; prev->L_bsdfsample[2] = lo.Lem.z; }
            }
            if (cont) { rec->wi[0] = ray.d.x; rec->wi[1] = ray.d.y; rec->wi[2] = ray.d.z; rec->pdf = bsdfPdf; rec->rrpdf = rr; if (lights) lights[base + nv - 1].bsdfpdf = bsdfPdf; }
            // eLd: what the estimator credits to this vertex itself (emission seen from the previous vertex with its MIS weight +
            // the light sample), as radiance leaving towards wo = the credited amount / the throughput up to here. eLi holds that
            // throughput until the path has ended (the pass below turns it into the radiance of the whole path suffix).
            const V3 dL = L - Lb;
            rec->eLd[0] = Tb.x > 0.f ? dL.x / Tb.x : 0.f; rec->eLd[1] = Tb.y > 0.f ? dL.y / Tb.y : 0.f; rec->eLd[2] = Tb.z > 0.f ? dL.z / Tb.z : 0.f;
            rec->eLi[0] = Tb.x; rec->eLi[1] = Tb.y; rec->eLi[2] = Tb.z;
        }
        if (!cont) break;
    }
    if (WRITE) {
        // eLi[j] = sum over the vertices m >= j of what the path credits to m, over the throughput up to j: the path's own
        // estimate of the radiance leaving vertex j towards wo (eLi of a path's first vertex = the sample's radiance).
        V3 acc = mk(0.f, 0.f, 0.f);
        for (unsigned int j = nv; j-- > 0;) {
            NoriGpuSPoint *rec = out + base + j;
            const V3 Tj = mk(rec->eLi[0], rec->eLi[1], rec->eLi[2]);
            acc = acc + mk(rec->eLd[0], rec->eLd[1], rec->eLd[2]) * Tj;
            rec->eLi[0] = Tj.x > 0.f ? acc.x / Tj.x : 0.f; rec->eLi[1] = Tj.y > 0.f ? acc.y / Tj.y : 0.f; rec->eLi[2] = Tj.z > 0.f ? acc.z / Tj.z : 0.f;
        }
    } else counts[g] = nv;
}

bool enqueuePathVertices(const DeviceScene &sc, const RenderParams &rp, uint32_t spp, unsigned long long nPaths, unsigned int *counts,
                         const unsigned long long *offsets, NoriGpuSPoint *out, cudaStream_t st, NoriGpuLPoint *lights) {
    const unsigned g = (unsigned) ((nPaths + 127ull) / 128ull);
#define PV(I) do { if (out) k_pathVertices<I, true><<<g, 128, 0, st>>>(sc, rp, spp, nPaths, counts, offsets, out, lights); \
                   else k_pathVertices<I, false><<<g, 128, 0, st>>>(sc, rp, spp, nPaths, counts, offsets, out, nullptr); } while (0)
    switch (rp.integrator) {
    case NORI_INT_PATH_MATS: PV(NORI_INT_PATH_MATS); break;
    case NORI_INT_PATH_EMS: PV(NORI_INT_PATH_EMS); break;
    case NORI_INT_PATH_MIS: PV(NORI_INT_PATH_MIS); break;
    default: return false;                                   // the path graph is built from path-traced vertices only
    }
#undef PV
    return true;
}

/// Exclusive prefix sums of the per-path vertex counts (64-bit): one CTA, sequential over tiles (a few million paths).
__global__ void __launch_bounds__(1024) k_scanCounts(const unsigned int *__restrict__ counts, unsigned long long nPaths, unsigned long long *__restrict__ offsets) {
    __shared__ unsigned long long sWarp[32];
    __shared__ unsigned long long sCarry;
    if (threadIdx.x == 0) sCarry = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (unsigned long long start = 0; start < nPaths; start += 1024ull) {
        const unsigned long long i = start + threadIdx.x;
        const unsigned long long v = i < nPaths ? counts[i] : 0ull;
        unsigned long long s = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const unsigned long long t = __shfl_up_sync(0xffffffffu, s, d); if ((int) lane >= d) s += t; }
        if (lane == 31) sWarp[warp] = s;
        __syncthreads();
        if (warp == 0) {
            unsigned long long w = sWarp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const unsigned long long t = __shfl_up_sync(0xffffffffu, w, d); if ((int) lane >= d) w += t; }
            sWarp[lane] = w;
        }
        __syncthreads();
        const unsigned long long carry = sCarry, base = warp ? sWarp[warp - 1] : 0ull;
        if (i < nPaths) offsets[i] = carry + base + s - v;
        __syncthreads();
        if (threadIdx.x == 1023) sCarry = carry + sWarp[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) offsets[nPaths] = sCarry;
}
void enqueueScanCounts(const unsigned int *counts, unsigned long long nPaths, unsigned long long *offsets, cudaStream_t st) {
    k_scanCounts<<<1, 1024, 0, st>>>(counts, nPaths, offsets);
}

/// Tail compaction, step 1: list the free slots below `newCap` and the occupied ones at or above it
/// (lists live in the idle shadow queue: [0, cap) and [cap, 2 cap) as uint32).
__global__ void k_compactScan(Pool pool, uint32_t newCap, Counters *c) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t *freeList = reinterpret_cast<uint32_t *>(pool.shadow), *liveList = freeList + pool.capacity;
    const bool occupied = i < pool.capacity && (pool.state[i].w & (kFlagAlive | kFlagHasSample));
    const bool isFree = i < newCap && !occupied, isLive = i >= newCap && occupied;
    const uint32_t qf = warpAppend(&c->compactCount[0], isFree), ql = warpAppend(&c->compactCount[1], isLive);
    NG_CHECK((!isFree || qf < pool.capacity) && (!isLive || ql < pool.capacity));
    if (isFree) freeList[qf] = i;
    if (isLive) liveList[ql] = i;
}
/// Step 2: move the k-th listed path into the k-th listed free slot.
__global__ void k_compactMove(Pool pool, Counters *c) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= c->compactCount[1] || k >= c->compactCount[0]) return;   // never read past either list (the host sizes newCap so that free >= live)
    const uint32_t *freeList = reinterpret_cast<const uint32_t *>(pool.shadow), *liveList = freeList + pool.capacity;
    NG_CHECK(k < c->compactCount[0]);                 // a free slot below the new size for every path above it
    const uint32_t src = liveList[k], dst = freeList[k];
    NG_CHECK(src < pool.capacity && dst < pool.capacity && (pool.state[src].w & (kFlagAlive | kFlagHasSample)) && !(pool.state[dst].w & (kFlagAlive | kFlagHasSample)));
    pool.rayO[dst] = pool.rayO[src]; pool.rayD[dst] = pool.rayD[src]; pool.thr[dst] = pool.thr[src]; pool.rad[dst] = pool.rad[src];
    pool.info[dst] = pool.info[src]; pool.pos[dst] = pool.pos[src]; pool.state[dst] = pool.state[src];
    pool.state[src] = make_uint4(0, 0, 0, 0);
}
int enqueueCompaction(const Pool &pool, uint32_t newCapacity, Counters *c, cudaStream_t st) {
    cudaMemsetAsync(&c->compactCount[0], 0, 2 * sizeof(unsigned int), st);
    k_compactScan<<<pool.capacity / 256, 256, 0, st>>>(pool, newCapacity, c);
    k_compactMove<<<(pool.capacity - newCapacity + 255u) / 256u, 256, 0, st>>>(pool, c);
    return 2;
}

__global__ void k_poolInit(Pool pool, Counters *c) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    pool.state[i] = make_uint4(0, 0, 0, 0);
    if (pool.order) { pool.key[i] = (kBinEmpty << 24) | i; pool.order[i] = i; }
    if (i == 0) {                                   // what k_prepare would have left: the first splat pass runs before it
        c->traceLimit = pool.capacity; c->orderSorted = 0;
        c->forcePass = 0xffffffffu;
        for (int s = 0; s < kMaxRing; ++s) c->slotPass[s] = (unsigned) s;
    }
}
__global__ void k_identityOrder(Pool pool, uint32_t n, Counters *c) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pool.order[i] = i;
    if (i == 0) { c->traceLimit = n; c->orderSorted = 0; }     // slot order: the closest-hit pass asks every slot whether it is alive
}
void enqueueIdentityOrder(const Pool &pool, uint32_t n, Counters *c, cudaStream_t st) { if (pool.order && n) k_identityOrder<<<(n + 255u) / 256u, 256, 0, st>>>(pool, n, c); }
// ---- ray binning: a counting sort of the (bin << 24 | slot) words by bin, specialised for 256 bins -----------------------------
// Three small kernels instead of a general radix pass (cub::DeviceRadixSort::SortKeys over bits 24..31: profiles/r02_notes.md).
// One CTA per chunk of kBinChunk consecutive words:
//   k_binCount    shared-memory histogram of the chunk -> counts[bin][chunk]            (no global atomics)
//   k_binScan     one warp per bin: exclusive offsets over the chunks (in place), totals[bin] behind the matrix
//   k_binScatter  the chunk again, sorted by bin inside shared memory first, so that consecutive threads write consecutive words
//                 of a (bin, chunk) run (a word stored straight to its place costs a 32-byte sector per lane: 3x slower).
//                 The lanes of a warp that hold fresh camera rays take one range of the cursor and keep their order, so the
//                 primary rays of one pixel patch stay neighbours; inside the other bins the order is whatever the
//                 shared-memory atomics make it, which only decides which lanes trace which rays.
constexpr int kBinChunk = 8192, kBinThreads = 512, kBins = 256;

__global__ void __launch_bounds__(kBinThreads) k_binCount(const uint32_t *__restrict__ key, uint32_t n, uint32_t nChunks, uint32_t *__restrict__ counts, const uint4 *__restrict__ state) {
    __shared__ unsigned int hist[kBins];
    for (int b = threadIdx.x; b < kBins; b += kBinThreads) hist[b] = 0;
    __syncthreads();
    const uint32_t lo = blockIdx.x * kBinChunk, hi = min(n, lo + kBinChunk);
    uint32_t k[kBinChunk / kBinThreads];
#pragma unroll
    for (int j = 0; j < kBinChunk / kBinThreads; ++j) { const uint32_t i = lo + j * kBinThreads + threadIdx.x; k[j] = i < hi ? key[i] : 0u; }
#pragma unroll
    for (int j = 0; j < kBinChunk / kBinThreads; ++j) if (lo + j * kBinThreads + threadIdx.x < hi) {
        atomicAdd(&hist[k[j] >> 24], 1u);
        // the closest-hit pass trusts the sorted list: a word is in the empty bin exactly if its slot holds no live path
        NG_CHECK((k[j] & kBinSlotMask) == lo + j * kBinThreads + threadIdx.x);
        NG_CHECK(((k[j] >> 24) == kBinEmpty) == !(state[k[j] & kBinSlotMask].w & kFlagAlive));
    }
    __syncthreads();
    for (int b = threadIdx.x; b < kBins; b += kBinThreads) counts[b * nChunks + blockIdx.x] = hist[b];
}
__device__ __forceinline__ unsigned int warpInclusive(unsigned int v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += t; }
    return v;
}
__global__ void __launch_bounds__(256) k_binScan(uint32_t *__restrict__ counts /* in: counts, out: offsets; then totals[kBins] */, uint32_t nChunks, uint32_t n, Counters *c) {
    const int bin = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    uint32_t *row = counts + (size_t) bin * nChunks;
    unsigned int carry = 0;
#pragma unroll 4
    for (uint32_t c0 = 0; c0 < nChunks; c0 += 32) {
        const uint32_t c = c0 + lane;
        const unsigned int v = c < nChunks ? row[c] : 0u, incl = warpInclusive(v, lane);
        if (c < nChunks) row[c] = carry + incl - v;
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) {
        counts[(size_t) kBins * nChunks + bin] = carry;
        // the empty slots sort last: the next closest-hit pass stops in front of them and trusts the list (ClosestIO::sorted)
        if (bin == (int) kBinEmpty) { c->traceLimit = n - carry; c->orderSorted = 1; }
    }
}
__global__ void __launch_bounds__(kBinThreads) k_binScatter(const uint32_t *__restrict__ key, uint32_t n, uint32_t nChunks, const uint32_t *__restrict__ offsets, uint32_t *__restrict__ order) {
    __shared__ uint32_t stage[kBinChunk];
    __shared__ unsigned int cursor[kBins], shift[kBins], wsumA[kBins / 32], wsumB[kBins / 32];
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t chunk = blockIdx.x, lo = chunk * kBinChunk, hi = min(n, lo + kBinChunk);
    uint32_t k[kBinChunk / kBinThreads];
#pragma unroll
    for (int j = 0; j < kBinChunk / kBinThreads; ++j) { const uint32_t i = lo + j * kBinThreads + tid; k[j] = i < hi ? key[i] : 0u; }
    // two exclusive scans over the bins: the totals (where a bin starts in the output) and this chunk's counts (where it starts in stage[])
    unsigned int tot = 0, cnt = 0, goff = 0, inclA = 0, inclB = 0;
    if (tid < kBins) {
        tot = offsets[(size_t) kBins * nChunks + tid];
        goff = offsets[(size_t) tid * nChunks + chunk];
        cnt = (chunk + 1 < nChunks ? offsets[(size_t) tid * nChunks + chunk + 1] : tot) - goff;
        inclA = warpInclusive(tot, lane); inclB = warpInclusive(cnt, lane);
        if (lane == 31) { wsumA[tid >> 5] = inclA; wsumB[tid >> 5] = inclB; }
    }
    __syncthreads();
    if (tid < kBins) {
        unsigned int preA = 0, preB = 0;
        for (int w = 0; w < (tid >> 5); ++w) { preA += wsumA[w]; preB += wsumB[w]; }
        const unsigned int lstart = preB + inclB - cnt;                    // where the bin starts in stage[]
        cursor[tid] = lstart;
        shift[tid] = preA + inclA - tot + goff - lstart;                    // stage[] index -> output index (mod 2^32)
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kBinChunk / kBinThreads; ++j) {
        const bool valid = lo + j * kBinThreads + tid < hi;
        const uint32_t bin = k[j] >> 24;
        const unsigned int fresh = __ballot_sync(0xffffffffu, valid && bin == kBinFresh);
        unsigned int baseF = 0;
        if (lane == 0 && fresh) baseF = atomicAdd(&cursor[kBinFresh], (unsigned int) __popc(fresh));
        baseF = __shfl_sync(0xffffffffu, baseF, 0);
        if (valid) stage[bin == kBinFresh ? baseF + __popc(fresh & ((1u << lane) - 1u)) : atomicAdd(&cursor[bin], 1u)] = k[j];
    }
    __syncthreads();
    for (uint32_t t = tid; t < hi - lo; t += kBinThreads) {
        const uint32_t w = stage[t], bin = w >> 24;
        order[shift[bin] + t] = w;
    }
}
size_t binSortTmpBytes(uint32_t n) { return (size_t) ((n + kBinChunk - 1) / kBinChunk + 1) * kBins * sizeof(uint32_t); }
static void enqueueBinSort(const Pool &pool, Counters *c, cudaStream_t st) {
    uint32_t *counts = static_cast<uint32_t *>(pool.sortTmp);
    const uint32_t n = pool.capacity, nChunks = (n + kBinChunk - 1) / kBinChunk;
    k_binCount<<<nChunks, kBinThreads, 0, st>>>(pool.key, n, nChunks, counts, pool.state);
    k_binScan<<<kBins / 8, 256, 0, st>>>(counts, nChunks, n, c);
    k_binScatter<<<nChunks, kBinThreads, 0, st>>>(pool.key, n, nChunks, counts, pool.order);
}
void enqueuePoolInit(const Pool &pool, Counters *c, cudaStream_t st) { k_poolInit<<<pool.capacity / 256, 256, 0, st>>>(pool, c); }

/// One CTA slot per resident block: 8 CTAs of 128 threads per SM (launch bounds of the trace kernels).
static unsigned persistentGrid() {
    static unsigned g = 0;
    if (!g) { int dev = 0, sms = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); g = (unsigned) sms * 8u; }
    return g;
}

/// Tunables of the traversal kernels (environment overrides exist only in -DNORI_GPU_SWEEP builds: build_variant).
struct Tune { int fetchClosest, fetchShadow; unsigned ctasPerSm, ctasClosest; bool overlapSplat; };
static const Tune &tune() {
    static Tune t = [] {
        // any-hit: 10 CTAs of 128 threads per SM (48 registers); closest-hit: 9 (56 registers) since the rays are binned (75.4 vs 77.9 ms at 512 spp)
        Tune v = {kTraceTuning, kTraceTuning, 10u, 9u, true};
#ifdef NORI_GPU_SWEEP
        if (getenv("NORI_GPU_NO_OVERLAP")) v.overlapSplat = false;
        if (const char *s = getenv("NORI_GPU_FETCH")) { v.fetchClosest = (v.fetchClosest & ~0xff) | atoi(s); v.fetchShadow = (v.fetchShadow & ~0xff) | atoi(s); }
        if (const char *s = getenv("NORI_GPU_FETCH_SHADOW")) v.fetchShadow = (v.fetchShadow & ~0xff) | atoi(s);
        if (const char *s = getenv("NORI_GPU_NODE_EXIT")) { v.fetchClosest = (v.fetchClosest & 0xff) | atoi(s) << 8; v.fetchShadow = (v.fetchShadow & 0xff) | atoi(s) << 8; }
        if (const char *s = getenv("NORI_GPU_NODE_EXIT_CLOSEST")) v.fetchClosest = (v.fetchClosest & 0xff) | atoi(s) << 8;
        if (const char *s = getenv("NORI_GPU_NODE_EXIT_SHADOW")) v.fetchShadow = (v.fetchShadow & 0xff) | atoi(s) << 8;
        if (const char *s = getenv("NORI_GPU_TRACE_CTAS")) v.ctasPerSm = v.ctasClosest = (unsigned) atoi(s);
        if (const char *s = getenv("NORI_GPU_TRACE_CTAS_CLOSEST")) v.ctasClosest = (unsigned) atoi(s);
#endif
        return v;
    }();
    return t;
}

/// Launches `kernel` so that it may begin while the kernel queued before it on `st` is still running
/// (programmatic stream serialisation): only for kernels that do not consume that kernel's output.
template <class... Params, class... Args>
static void launchOverlapped(void (*kernel)(Params...), unsigned grid, unsigned block, cudaStream_t st, bool overlap, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = overlap ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, args...);
}

// One iteration. The splat of the PREVIOUS iteration's finished tiles is queued first and the
// closest-hit traversal directly behind it as a programmatic dependent launch: the traversal touches
// only the path pool, the splat only sample records, tile counters and the frame, so the traversal's
// CTAs start on the SMs the (few, register-heavy) splat CTAs leave free instead of waiting for them.
// k_prepare is a normal launch: it starts when both are complete.
// ev (profiling): [0] begin [1] splat(i-1) + trace end [2] shade end [3] shadow end.
int enqueueIteration(const DeviceScene &sc, const Pool &pool, const RenderParams &rp, Counters *c, cudaStream_t st, cudaEvent_t *ev) {
    int launches = 0;
    const bool splat = !rp.listPixel;
    if (ev) cudaEventRecord(ev[0], st);
    if (splat) {
        k_splatPlan<<<((uint32_t) rp.ring * rp.numTiles + 255u) / 256u, 256, 0, st>>>(sc, rp, c); ++launches;
        const unsigned g = persistentGrid() / 8u;          // one resident CTA per SM (248 registers x 256 threads): a single wave, so
                                                           // that every CTA reaches launch_dependents at once
        switch (sc.filter.border) {
        case 0: k_splatPixel<0, kSplatThreads><<<g, kSplatThreads, 0, st>>>(sc, rp, c); break;
        case 1: k_splatPixel<1, kSplatThreads><<<g, kSplatThreads, 0, st>>>(sc, rp, c); break;
        case 2: k_splatPixel<2, kSplatThreads><<<g, kSplatThreads, 0, st>>>(sc, rp, c); break;
        default: k_splat<<<g, 256, 0, st>>>(sc, rp, c); break;
        }
        ++launches;
    }
    {
        const unsigned b = (sc.isectMode == NORI_ISECT_WATERTIGHT || sc.qnodes || sc.wnodes) ? 8u : tune().ctasClosest, g = persistentGrid() * b / 8u;
        const bool overlap = splat && tune().overlapSplat;
        if (sc.isectMode == NORI_ISECT_WATERTIGHT) launchOverlapped(k_traceClosest<8, true>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        else if (sc.qnodes) launchOverlapped(k_traceClosest<8, false, 1>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        else if (sc.wnodes) launchOverlapped(k_traceClosest<8, false, 2>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        else if (b == 10) launchOverlapped(k_traceClosest<10>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        else if (b == 12) launchOverlapped(k_traceClosest<12>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        else if (b == 9) launchOverlapped(k_traceClosest<9>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        else if (b == 11) launchOverlapped(k_traceClosest<11>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        else launchOverlapped(k_traceClosest<8>, g, 128, st, overlap, sc, pool, c, tune().fetchClosest);
        ++launches;
    }
    if (ev) cudaEventRecord(ev[1], st);
    k_prepare<<<1, 1, 0, st>>>(c, rp); ++launches;
    switch (rp.integrator) {
    case NORI_INT_NORMALS: launchShade<NORI_INT_NORMALS>(sc, pool, rp, c, st); break;
    case NORI_INT_PATH_MATS: launchShade<NORI_INT_PATH_MATS>(sc, pool, rp, c, st); break;
    case NORI_INT_PATH_EMS: launchShade<NORI_INT_PATH_EMS>(sc, pool, rp, c, st); break;
    case NORI_INT_PATH_MIS: launchShade<NORI_INT_PATH_MIS>(sc, pool, rp, c, st); break;
    case NORI_INT_WHITTED: launchShade<NORI_INT_WHITTED>(sc, pool, rp, c, st); break;
    case NORI_INT_PRIMARY_SHADOW: launchShade<NORI_INT_PRIMARY_SHADOW>(sc, pool, rp, c, st); break;
    case NORI_INT_SIMPLE: launchShade<NORI_INT_SIMPLE>(sc, pool, rp, c, st); break;
    default: launchShade<NORI_INT_AO>(sc, pool, rp, c, st); break;
    }
    ++launches;
    if (pool.order) {                               // one 8-bit radix pass over (bin << 24 | slot): slot numbers ordered by bin for the next closest-hit pass
        enqueueBinSort(pool, c, st);
        launches += 3;
    }
    if (ev) cudaEventRecord(ev[2], st);
    if (rp.integrator != NORI_INT_NORMALS) {
        const unsigned b = (sc.isectMode == NORI_ISECT_WATERTIGHT || sc.qnodes || sc.wnodes) ? 8u : tune().ctasPerSm, g = persistentGrid() * b / 8u;
        if (sc.isectMode == NORI_ISECT_WATERTIGHT) k_traceShadow<8, true><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        else if (sc.qnodes) k_traceShadow<8, false, 1><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        else if (sc.wnodes) k_traceShadow<8, false, 2><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        else if (b == 10) k_traceShadow<10><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        else if (b == 12) k_traceShadow<12><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        else if (b == 9) k_traceShadow<9><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        else if (b == 11) k_traceShadow<11><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        else k_traceShadow<8><<<g, 128, 0, st>>>(sc, pool, rp.recA, rp.recB, rp.recCount, c, tune().fetchShadow);
        ++launches;
    }
    if (ev) cudaEventRecord(ev[3], st);
    return launches;
}

unsigned renderChecks(unsigned *line) { return fetchChecks(line); }

} // namespace ng
