// accel_multi.cpp — S6: multi-mesh nori::Accel for the _ref build (TEST INFRASTRUCTURE).
//
// The reference's Accel (src/accel.cpp:24-55) accepts one mesh and loops over all triangles.
// Every scene on the hot path has several meshes, and brute force at C1 size never finishes,
// so the _ref build links THIS implementation of the class declared in include/nori/accel.h
// (same signatures; the header is used unmodified, extra state lives in a side table):
//   mode 0  brute force over the concatenation of all meshes, calling the reference's own
//           Mesh::rayIntersect (src/mesh.cpp:51-88) per triangle — the hit-ID ground truth;
//   mode 1  binary SAH BVH of the standalone port (../noracle_scene.h) for CPU-baseline timing.
// The unmodified single-mesh class is also in the library as nori::StockAccel
// (stock_accel.cpp) so tests can pin mode 0 against it on single-mesh scenes.
#include "plugins.h"
#include "../noracle_scene.h"
#include <nori/accel.h>
#include <Eigen/Geometry>
#include <map>
#include <mutex>
#include <atomic>

NORI_NAMESPACE_BEGIN

namespace {
struct AccelState {
    std::vector<Mesh *> meshes;
    noracle::Scene shadow;          // geometry copy feeding the port's BVH
    bool built = false;
};
std::mutex g_mtx;
std::map<const Accel *, AccelState *> g_state;
std::atomic<int> g_useBvh(1), g_count(0);
std::atomic<uint64_t> g_rays(0), g_box(0), g_tri(0);

AccelState *stateOf(const Accel *a, bool create) {
    thread_local const Accel *last = nullptr; thread_local AccelState *lastState = nullptr;
    if (last == a && lastState) return lastState;
    std::lock_guard<std::mutex> lk(g_mtx);
    AccelState *&s = g_state[a];
    if (!s && create) s = new AccelState();
    last = a; lastState = s;
    return s;
}
} // namespace

void accelSetMode(int useBvh) { g_useBvh = useBvh; }
void accelEnableCounters(int on) { g_count = on; }
void accelCounters(uint64_t *out3, int reset) {
    out3[0] = g_rays; out3[1] = g_box; out3[2] = g_tri;
    if (reset) { g_rays = 0; g_box = 0; g_tri = 0; }
}

void Accel::addMesh(Mesh *mesh) {
    AccelState *s = stateOf(this, true);
    s->meshes.push_back(mesh);
    if (!m_mesh) m_mesh = mesh;
    m_bbox.expandBy(mesh->getBoundingBox());
}

void Accel::build() {
    AccelState *s = stateOf(this, true);
    s->shadow.meshes.clear();
    for (Mesh *m : s->meshes) {
        noracle::MeshData d;
        const MatrixXf &V = m->getVertexPositions(); const MatrixXu &F = m->getIndices();
        d.V.assign(V.data(), V.data() + V.size());
        d.F.assign(F.data(), F.data() + F.size());
        d.bsdf = -1; d.emitter = -1;
        s->shadow.meshes.push_back(d);
    }
    s->shadow.accel.useBvh = true;
    s->shadow.accel.build(&s->shadow);
    s->built = true;
}

static bool traceRaw(AccelState *s, const Ray3f &ray_, bool shadowRay, uint32_t &mi, uint32_t &f, float &tO, float &uO, float &vO) {
    if (g_useBvh && s->built) {
        noracle::Ray r(toO(ray_.o), toO(ray_.d), ray_.mint, ray_.maxt);
        uint32_t gid;
        if (g_count) {      // counters are test-only and single-threaded
            noracle::TraversalCounters tc;
            s->shadow.accel.counters = &tc;
            gid = s->shadow.accel.trace(r, shadowRay, tO, uO, vO);
            s->shadow.accel.counters = nullptr;
            g_rays += tc.rays; g_box += tc.boxTests; g_tri += tc.triTests;
        } else gid = s->shadow.accel.trace(r, shadowRay, tO, uO, vO);
        if (gid == 0xffffffffu) return false;
        mi = s->shadow.accel.triMesh[gid]; f = s->shadow.accel.triLocal[gid];
        return true;
    }
    Ray3f ray(ray_);
    bool found = false;
    for (uint32_t m = 0; m < (uint32_t) s->meshes.size(); ++m) {
        const Mesh *mesh = s->meshes[m];
        for (uint32_t idx = 0; idx < mesh->getTriangleCount(); ++idx) {
            float u, v, t;
            if (mesh->rayIntersect(idx, ray, u, v, t)) {
                tO = t; uO = u; vO = v; mi = m; f = idx;
                if (shadowRay) return true;
                ray.maxt = t;
                found = true;
            }
        }
    }
    return found;
}

bool accelTraceRaw(const Scene *scene, const Ray3f &ray, bool shadow, uint32_t &mesh, uint32_t &tri, float &t, float &u, float &v) {
    // Scene::m_accel is private; reach the state through the one Accel whose first mesh matches.
    AccelState *st = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_mtx);
        for (auto &kv : g_state)
            if (kv.second && !kv.second->meshes.empty() && !scene->getMeshes().empty() &&
                kv.second->meshes.size() == scene->getMeshes().size() && kv.second->meshes[0] == scene->getMeshes()[0]) { st = kv.second; break; }
    }
    if (!st) return false;
    return traceRaw(st, ray, shadow, mesh, tri, t, u, v);
}

bool Accel::rayIntersect(const Ray3f &ray_, Intersection &its, bool shadowRay) const {
    AccelState *s = stateOf(this, false);
    if (!s) return false;
    uint32_t mi = 0, f = 0; float t = 0, u = 0, v = 0;
    if (!traceRaw(s, ray_, shadowRay, mi, f, t, u, v)) return false;
    if (shadowRay) return true;

    // Intersection record, as src/accel.cpp:57-108 derives it from (mesh, f, u, v, t).
    const Mesh *mesh = s->meshes[mi];
    its.t = t; its.mesh = mesh;
    Vector3f bary(1 - Point2f(u, v).sum(), u, v);
    const MatrixXf &V = mesh->getVertexPositions(), &N = mesh->getVertexNormals(), &UV = mesh->getVertexTexCoords();
    const MatrixXu &F = mesh->getIndices();
    uint32_t i0 = F(0, f), i1 = F(1, f), i2 = F(2, f);
    Point3f p0 = V.col(i0), p1 = V.col(i1), p2 = V.col(i2);
    its.p = bary.x() * p0 + bary.y() * p1 + bary.z() * p2;
    if (UV.size() > 0) its.uv = bary.x() * UV.col(i0) + bary.y() * UV.col(i1) + bary.z() * UV.col(i2);
    else its.uv = Point2f(u, v);
    its.geoFrame = Frame((p1 - p0).cross(p2 - p0).normalized());
    if (N.size() > 0)
        its.shFrame = Frame((bary.x() * N.col(i0) + bary.y() * N.col(i1) + bary.z() * N.col(i2)).normalized());
    else
        its.shFrame = its.geoFrame;
    return true;
}

NORI_NAMESPACE_END
