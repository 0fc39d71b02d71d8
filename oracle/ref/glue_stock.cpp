// glue_stock.cpp — C access to the reference's UNMODIFIED single-mesh brute-force Accel
// (TEST INFRASTRUCTURE). The Makefile compiles this file and $(REF)/src/accel.cpp with
// -DAccel=StockAccel, so the stock class lives beside the multi-mesh replacement
// (accel_multi.cpp) as nori::StockAccel. Never include <nori/scene.h> here: its inline
// Scene::rayIntersect would be emitted against the renamed class.
#include <nori/accel.h>
#include <nori/mesh.h>
#include <cstring>
#include <algorithm>

using namespace nori;

extern "C" {

/// Accel::rayIntersect (src/accel.cpp:35-111) on ONE mesh: out n x 24
/// [p t uv sh.s sh.t sh.n geo.s geo.t geo.n], valid n.
void ref_stock_intersect(void *mesh_, const float *rays, int n, int shadow, float *out, int *valid) {
    Mesh *mesh = (Mesh *) mesh_;
    Accel accel;
    accel.addMesh(mesh);
    accel.build();
    for (int i = 0; i < n; ++i) {
        const float *p = rays + 8 * (size_t) i;
        Ray3f ray(Point3f(p[0], p[1], p[2]), Vector3f(p[4], p[5], p[6]), p[3], p[7]);
        Intersection its;
        valid[i] = accel.rayIntersect(ray, its, shadow != 0) ? 1 : 0;
        float *o = out + 24 * (size_t) i;
        std::fill(o, o + 24, 0.0f);
        if (!valid[i] || shadow) continue;
        const float v[24] = {its.p.x(), its.p.y(), its.p.z(), its.t, its.uv.x(), its.uv.y(),
            its.shFrame.s.x(), its.shFrame.s.y(), its.shFrame.s.z(), its.shFrame.t.x(), its.shFrame.t.y(), its.shFrame.t.z(),
            its.shFrame.n.x(), its.shFrame.n.y(), its.shFrame.n.z(),
            its.geoFrame.s.x(), its.geoFrame.s.y(), its.geoFrame.s.z(), its.geoFrame.t.x(), its.geoFrame.t.y(), its.geoFrame.t.z(),
            its.geoFrame.n.x(), its.geoFrame.n.y(), its.geoFrame.n.z()};
        std::memcpy(o, v, sizeof(v));
    }
}

} // extern "C"
