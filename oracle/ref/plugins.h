// plugins.h — declarations shared by the Nori-side oracle plugins and the C glue
// (TEST INFRASTRUCTURE: compiled only into oracle/_ref/libnori_ref.so, together with the
// unmodified reference sources taken from /root/reference where they lie).
#pragma once
#include <nori/bsdf.h>
#include <nori/emitter.h>
#include <nori/integrator.h>
#include <nori/scene.h>
#include <nori/mesh.h>
#include <nori/dpdf.h>
#include <nori/sampler.h>
#include <nori/warp.h>
#include "../noracle_integrators.h"

NORI_NAMESPACE_BEGIN

inline noracle::V3 toO(const Vector3f &v) { return noracle::mk(v.x(), v.y(), v.z()); }
inline noracle::V3 toO(const Point3f &v) { return noracle::mk(v.x(), v.y(), v.z()); }
inline noracle::V3 toO(const Normal3f &v) { return noracle::mk(v.x(), v.y(), v.z()); }
inline Vector3f toN(noracle::V3 v) { return Vector3f(v.x, v.y, v.z); }
inline Color3f toN(noracle::C3 c) { return Color3f(c.r, c.g, c.b); }
inline noracle::C3 toO(const Color3f &c) { return noracle::mkc(c.r(), c.g(), c.b()); }

/// BSDFs restated by the oracle expose their parameters so the glue can flatten a scene.
class OracleBSDF : public BSDF {
public:
    noracle::Bsdf m_model;
    Color3f sample(BSDFQueryRecord &bRec, const Point2f &sample) const;
    Color3f eval(const BSDFQueryRecord &bRec) const;
    float pdf(const BSDFQueryRecord &bRec) const;
    bool isDiffuse() const { return m_model.isDiffuse(); }
};

/// S4: `area` emitter. The reference's Emitter base class is empty (include/nori/emitter.h:28-36);
/// the interface below is the oracle's ("the emitter interface is the builder's to define").
class AreaLight : public Emitter {
public:
    AreaLight(const PropertyList &props);
    void setParent(NoriObject *parent);
    std::string toString() const;
    Color3f m_radiance;
    const Mesh *m_mesh = nullptr;
};

/// Integrators restated by the oracle: one class, type chosen by the registered name.
class OracleIntegrator : public Integrator {
public:
    noracle::IntegratorParams m_params;
    OracleIntegrator(int type, const PropertyList &props);
    void preprocess(const Scene *scene);
    Color3f Li(const Scene *scene, Sampler *sampler, const Ray3f &ray) const;
    std::string toString() const;
};

/// World concept (noracle_integrators.h) over the reference's Scene graph.
struct NoriWorld {
    struct Em { int mesh; const AreaLight *light; DiscretePDF pdf; float area; };
    const Scene *scene;
    std::vector<const Mesh *> meshes;
    std::vector<int> meshEm;
    std::vector<Em> emitters;
    explicit NoriWorld(const Scene *s);
    static const NoriWorld &get(const Scene *s);     // lazily built, cached per Scene

    int meshIndex(const Mesh *m) const { for (size_t i = 0; i < meshes.size(); ++i) if (meshes[i] == m) return (int) i; return -1; }
    bool intersect(const noracle::Ray &r, noracle::Hit &h) const;
    bool occluded(const noracle::Ray &r) const;
    noracle::C3 bsdfSample(int mesh, noracle::BsdfQuery &q, noracle::V2 s) const;
    noracle::C3 bsdfEval(int mesh, const noracle::BsdfQuery &q) const;
    float bsdfPdf(int mesh, const noracle::BsdfQuery &q) const;
    bool bsdfIsDiffuse(int mesh) const { return meshes[mesh]->getBSDF()->isDiffuse(); }
    int emitterCount() const { return (int) emitters.size(); }
    int meshEmitter(int mesh) const { return meshEm[mesh]; }
    noracle::C3 emitterRadiance(int e) const { return toO(emitters[e].light->m_radiance); }
    float emitterArea(int e) const { return emitters[e].area; }
    void emitterSamplePoint(int e, noracle::V2 xi, noracle::V3 &p, noracle::V3 &n) const;
};


/// Multi-mesh Accel side state (oracle/ref/accel_multi.cpp).
void accelSetMode(int useBvh);                         // process-wide: 0 brute force, 1 SAH BVH
void accelCounters(uint64_t *out3, int reset);         // {rays, boxTests, triTests} (when enabled)
void accelEnableCounters(int on);
/// Raw hit of the multi-mesh Accel: mesh index in Scene::getMeshes() order + local triangle id.
bool accelTraceRaw(const Scene *scene, const Ray3f &ray, bool shadow, uint32_t &mesh, uint32_t &tri, float &t, float &u, float &v);

NORI_NAMESPACE_END
