// noracle_integrators.h — CPU restatement of the Integrator::Li implementations the reference
// declares but does not ship (TEST INFRASTRUCTURE; see noracle_math.h for the usage rule).
//
// Contract followed: include/nori/integrator.h:40,54 (`preprocess`, `Li(scene, sampler, ray)`),
// call pattern src/main.cpp:57-63 and src/ttest.cpp:168-175. SURVEY.md §8c-S4/S5 lists the
// definitions; the choices no reference code pins (marked UNPINNED) are fixed here once and
// used identically by the CUDA kernels (DESIGN.md "Unpinned semantics").
//
// Written once against a `World` concept so that the same code runs over
//   (a) the flattened scene of the standalone port (noracle_scene.h), and
//   (b) the compiled reference's Scene/Mesh/BSDF objects (oracle/ref/plugins.cpp),
// which is how the reference's own ttest/chi2 scene fixtures pin it.
//
// Smp concept: float next1D(); V2 next2D();   (Sampler::next1D/next2D, include/nori/sampler.h:88-91)
// World concept:
//   bool  intersect(const Ray&, Hit&) const;          // Accel::rayIntersect(ray, its, false)
//   bool  occluded(const Ray&) const;                 // Accel::rayIntersect(ray, its, true)
//   C3    bsdfSample(int mesh, BsdfQuery&, V2) const; // BSDF::sample
//   C3    bsdfEval(int mesh, const BsdfQuery&) const; // BSDF::eval
//   float bsdfPdf(int mesh, const BsdfQuery&) const;  // BSDF::pdf
//   bool  bsdfIsDiffuse(int mesh) const;              // BSDF::isDiffuse
//   int   emitterCount() const;
//   int   meshEmitter(int mesh) const;                // -1 when the mesh carries no emitter
//   C3    emitterRadiance(int e) const;
//   float emitterArea(int e) const;                   // sum of triangle areas (Mesh::surfaceArea)
//   void  emitterSamplePoint(int e, V2 xi, V3 &p, V3 &n) const;  // uniform by area
#pragma once
#include "noracle_math.h"

namespace noracle {

struct Hit {
    V3 p; float t; V2 uv; Frame sh, geo; int mesh; uint32_t prim;
    V3 toLocal(V3 d) const { return sh.toLocal(d); }
    V3 toWorld(V3 d) const { return sh.toWorld(d); }
};

enum IntegratorType { kNormals = 0, kPathMats = 1, kPathEms = 2, kPathMis = 3, kWhitted = 4,
                      kPrimaryShadow = 5, kSimple = 6, kAo = 7 };

struct IntegratorParams {
    int type;
    int rrStart;          // UNPINNED: Russian roulette starts once depth >= rrStart (3)
    float rrMax;          // UNPINNED: continuation probability cap (0.99)
    int maxDepth;         // UNPINNED: hard cap on path vertices; <=0 means none
    V3 lightPos; C3 lightEnergy;   // `simple` (scenes/pa3/ajax-simple.xml:9-10)
    IntegratorParams() : type(kPathMis), rrStart(3), rrMax(0.99f), maxDepth(0) {
        lightPos = mk(0, 0, 0); lightEnergy = mkc(0, 0, 0);
    }
};

/// Per-call counters (rays, segments) used for Mrays/s and Gsegments/s and the roofline bytes.
struct PathStats { uint64_t closest, shadow, segments; PathStats() : closest(0), shadow(0), segments(0) {} };

// --- S4: area emitter ------------------------------------------------------------------
/// Radiance leaving a point with shading normal n towards w (one-sided).
template <class World> inline C3 emitterEval(const World &w, int e, V3 n, V3 wOut) {
    return dot(n, wOut) > 0.0f ? w.emitterRadiance(e) : mkc(0, 0, 0);
}
/// Solid-angle density of sampling point p (normal n) on emitter e from ref, for ONE emitter.
template <class World> inline float emitterPdfSolidAngle(const World &w, int e, V3 ref, V3 p, V3 n) {
    V3 d = p - ref;
    float dist2 = squaredNorm(d);
    float dist = std::sqrt(dist2);
    float cosL = dot(n, -(d / dist));
    if (!(cosL > 0.0f)) return 0.0f;
    return dist2 / (cosL * w.emitterArea(e));
}

struct LightSample { V3 p, n, wi; float dist, pdf; C3 Le; int emitter; };

/// UNPINNED: one emitter picked uniformly (pdf 1/nLights) from u; then a uniform point by area.
template <class World> inline bool sampleLights(const World &w, V3 ref, float u, V2 xi, LightSample &ls) {
    int n = w.emitterCount();
    if (n == 0) return false;
    int e = std::min((int) (u * (float) n), n - 1);
    ls.emitter = e;
    w.emitterSamplePoint(e, xi, ls.p, ls.n);
    V3 d = ls.p - ref;
    float dist2 = squaredNorm(d);
    ls.dist = std::sqrt(dist2);
    ls.wi = d / ls.dist;
    float cosL = dot(ls.n, -ls.wi);
    if (!(cosL > 0.0f) || !(ls.dist > 0.0f)) { ls.pdf = 0.0f; return false; }
    ls.pdf = dist2 / (cosL * w.emitterArea(e)) / (float) n;
    ls.Le = w.emitterRadiance(e);
    return true;
}

// --- S5: integrators --------------------------------------------------------------------
template <class World> inline C3 LiNormals(const World &w, const Ray &ray, PathStats &st) {
    Hit its; st.closest++; st.segments++;
    if (!w.intersect(ray, its)) return mkc(0, 0, 0);
    return mkc(std::abs(its.sh.n.x), std::abs(its.sh.n.y), std::abs(its.sh.n.z));
}

/// Unidirectional path tracer; `type` selects mats / ems / mis. Draw order per vertex:
/// [next1D light select, next2D light point] (only at non-discrete BSDFs, ems/mis),
/// [next2D BSDF], [next1D roulette once depth >= rrStart].
template <class World, class Smp>
inline C3 LiPath(const World &w, Smp &smp, Ray ray, const IntegratorParams &ip, PathStats &st) {
    const bool useEms = ip.type == kPathEms || ip.type == kPathMis;
    const bool useMats = ip.type == kPathMats || ip.type == kPathMis;
    const bool mis = ip.type == kPathMis;
    C3 L = mkc(0, 0, 0), T = mkc(1, 1, 1);
    float eta = 1.0f, bsdfPdf = 0.0f;
    bool discrete = true;      // camera vertex counts as discrete: emitted light is not double counted
    int nLights = w.emitterCount();
    for (int depth = 0; ; ) {
        Hit its; st.closest++; st.segments++;
        if (!w.intersect(ray, its)) break;

        int e = w.meshEmitter(its.mesh);
        if (e >= 0) {
            C3 Le = emitterEval(w, e, its.sh.n, -ray.d);
            if (!isZero(Le)) {
                if (discrete || !useEms) L = L + T * Le;
                else if (useMats) {   // MIS: BSDF-sampled hit
                    float pl = emitterPdfSolidAngle(w, e, ray.o, its.p, its.sh.n) / (float) nLights;
                    float wgt = bsdfPdf / (bsdfPdf + pl);
                    L = L + T * Le * wgt;
                }
            }
        }
        if (ip.maxDepth > 0 && depth + 1 >= ip.maxDepth) break;

        V3 wiLocal = its.toLocal(-ray.d);
        if (useEms && w.bsdfIsDiffuse(its.mesh) && nLights > 0) {
            float u = smp.next1D();
            V2 xi = smp.next2D();
            LightSample ls;
            if (sampleLights(w, its.p, u, xi, ls)) {
                BsdfQuery q; q.wi = wiLocal; q.wo = its.toLocal(ls.wi); q.eta = 1.0f; q.measure = kSolidAngle;
                C3 f = w.bsdfEval(its.mesh, q);
                if (!isZero(f)) {
                    Ray sr(its.p, ls.wi, kEpsilon, ls.dist - kEpsilon);
                    st.shadow++;
                    if (!w.occluded(sr)) {
                        float wgt = 1.0f;
                        if (mis) { float pb = w.bsdfPdf(its.mesh, q); wgt = ls.pdf / (ls.pdf + pb); }
                        float c = std::abs(q.wo.z);
                        L = L + T * f * ls.Le * (c * wgt / ls.pdf);
                    }
                }
            }
        }

        BsdfQuery q; q.wi = wiLocal; q.eta = 1.0f; q.measure = kUnknownMeasure;
        C3 wgt = w.bsdfSample(its.mesh, q, smp.next2D());
        if (isZero(wgt)) break;
        T = T * wgt;
        eta *= q.eta;
        discrete = q.measure == kDiscrete;
        bsdfPdf = discrete ? 0.0f : w.bsdfPdf(its.mesh, q);
        ray = Ray(its.p, its.toWorld(q.wo));
        ++depth;
        if (depth >= ip.rrStart) {
            float qc = std::min(maxCoeff(T) * eta * eta, ip.rrMax);
            if (!(smp.next1D() < qc)) break;
            T = T / qc;
        }
    }
    return L;
}

/// `whitted` (scenes/pa4): emitted + direct light at diffuse surfaces; recursion through discrete
/// BSDFs with probability 0.95 (pinned by pa4/tests/test-mesh-furnace.xml:16 = 1.5, 1.8).
template <class World, class Smp>
inline C3 LiWhitted(const World &w, Smp &smp, Ray ray, PathStats &st) {
    C3 T = mkc(1, 1, 1);
    int nLights = w.emitterCount();
    for (;;) {
        Hit its; st.closest++; st.segments++;
        if (!w.intersect(ray, its)) return mkc(0, 0, 0);
        V3 wiLocal = its.toLocal(-ray.d);
        if (w.bsdfIsDiffuse(its.mesh)) {
            C3 L = mkc(0, 0, 0);
            int e = w.meshEmitter(its.mesh);
            if (e >= 0) L = L + emitterEval(w, e, its.sh.n, -ray.d);
            if (nLights > 0) {
                float u = smp.next1D();
                V2 xi = smp.next2D();
                LightSample ls;
                if (sampleLights(w, its.p, u, xi, ls)) {
                    BsdfQuery q; q.wi = wiLocal; q.wo = its.toLocal(ls.wi); q.eta = 1.0f; q.measure = kSolidAngle;
                    C3 f = w.bsdfEval(its.mesh, q);
                    if (!isZero(f)) {
                        Ray sr(its.p, ls.wi, kEpsilon, ls.dist - kEpsilon);
                        st.shadow++;
                        if (!w.occluded(sr)) L = L + f * ls.Le * (std::abs(q.wo.z) / ls.pdf);
                    }
                }
            }
            return T * L;
        }
        if (!(smp.next1D() < 0.95f)) return mkc(0, 0, 0);
        BsdfQuery q; q.wi = wiLocal; q.eta = 1.0f; q.measure = kUnknownMeasure;
        C3 wgt = w.bsdfSample(its.mesh, q, smp.next2D());
        if (isZero(wgt)) return mkc(0, 0, 0);
        T = T * wgt / 0.95f;
        ray = Ray(its.p, its.toWorld(q.wo));
    }
}

/// C4 workload: one primary ray plus one shadow ray towards a sampled light point;
/// value = unshadowed ? |cos| : 0 as grey — defined here, no reference counterpart.
template <class World, class Smp>
inline C3 LiPrimaryShadow(const World &w, Smp &smp, const Ray &ray, PathStats &st) {
    Hit its; st.closest++; st.segments++;
    if (!w.intersect(ray, its)) return mkc(0, 0, 0);
    if (w.emitterCount() == 0) return mkc(0, 0, 0);
    float u = smp.next1D();
    V2 xi = smp.next2D();
    LightSample ls;
    if (!sampleLights(w, its.p, u, xi, ls)) return mkc(0, 0, 0);
    Ray sr(its.p, ls.wi, kEpsilon, ls.dist - kEpsilon);
    st.shadow++;
    if (w.occluded(sr)) return mkc(0, 0, 0);
    float c = std::abs(dot(its.sh.n, ls.wi));
    return mkc(c, c, c);
}

/// `simple` (scenes/pa3/ajax-simple.xml): point light, Phi/(4 pi^2) * max(0,cos)/d^2 * visibility.
template <class World>
inline C3 LiSimple(const World &w, const Ray &ray, const IntegratorParams &ip, PathStats &st) {
    Hit its; st.closest++; st.segments++;
    if (!w.intersect(ray, its)) return mkc(0, 0, 0);
    V3 d = ip.lightPos - its.p;
    float dist2 = squaredNorm(d), dist = std::sqrt(dist2);
    V3 wi = d / dist;
    float c = dot(its.sh.n, wi);
    if (!(c > 0.0f)) return mkc(0, 0, 0);
    Ray sr(its.p, wi, kEpsilon, dist - kEpsilon);
    st.shadow++;
    if (w.occluded(sr)) return mkc(0, 0, 0);
    return ip.lightEnergy * (c / (4.0f * kPi * kPi * dist2));
}

/// `ao` (scenes/pa3/ajax-ao.xml): cosine-weighted hemisphere visibility.
template <class World, class Smp>
inline C3 LiAo(const World &w, Smp &smp, const Ray &ray, PathStats &st) {
    Hit its; st.closest++; st.segments++;
    if (!w.intersect(ray, its)) return mkc(0, 0, 0);
    V3 dl = squareToCosineHemisphere(smp.next2D());
    Ray sr(its.p, its.toWorld(dl));
    st.shadow++;
    return w.occluded(sr) ? mkc(0, 0, 0) : mkc(1, 1, 1);
}

template <class World, class Smp>
inline C3 Li(const World &w, Smp &smp, const Ray &ray, const IntegratorParams &ip, PathStats &st) {
    switch (ip.type) {
    case kNormals:       return LiNormals(w, ray, st);
    case kWhitted:       return LiWhitted(w, smp, ray, st);
    case kPrimaryShadow: return LiPrimaryShadow(w, smp, ray, st);
    case kSimple:        return LiSimple(w, ray, ip, st);
    case kAo:            return LiAo(w, smp, ray, st);
    default:             return LiPath(w, smp, ray, ip, st);
    }
}

} // namespace noracle
