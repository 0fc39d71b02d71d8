// glue.cpp — C API over the COMPILED REFERENCE (TEST INFRASTRUCTURE; oracle/_ref/libnori_ref.so).
// Everything called here is the reference's own object code (parser, OBJ loader, PCG32 sampler,
// camera, Mesh::rayIntersect, ImageBlock, filters, Diffuse/Mirror, fresnel, ...) except the
// plugins the reference leaves unimplemented, which come from plugins.cpp / accel_multi.cpp.
// Used to (1) pin the standalone port and the CUDA path bit-exactly on the REF functions,
// (2) run the reference's chi2/t-test scene fixtures over the restated plugins,
// (3) flatten reference scenes into committed fixtures, (4) time the CPU baseline ("reference").
#include "plugins.h"
#include <nori/parser.h>
#include <nori/camera.h>
#include <nori/block.h>
#include <nori/rfilter.h>
#include <nori/timer.h>
#include <filesystem/resolver.h>
#include <pcg32.h>
#include <tbb/parallel_for.h>
#include <tbb/blocked_range.h>
#include <tbb/task_scheduler_init.h>
#include <thread>
#include <cstring>

namespace nori {
bool refCameraState(const Camera *cam, float *s2c, float *c2w, float *nearClip, float *farClip, float *fov);
bool refDiffuseAlbedo(const BSDF *bsdf, float *rgb);
}
using namespace nori;

static thread_local std::string g_err;
#define GUARD(expr, fail) try { expr; } catch (const std::exception &e) { g_err = e.what(); return fail; }

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

/// loadFromXML (src/parser.cpp:28). Tests inside the XML run during activate(), as in `nori file.xml`.
void *ref_load_xml(const char *path) {
    GUARD({
        filesystem::path p(path);
        getFileResolver()->prepend(p.parent_path());      // src/main.cpp:181
        return loadFromXML(path);
    }, nullptr)
}
void ref_object_destroy(void *o) { delete (NoriObject *) o; }
int ref_object_class_type(void *o) { return (int) ((NoriObject *) o)->getClassType(); }

static const Scene *S(void *o) { return static_cast<const Scene *>((NoriObject *) o); }

int ref_scene_mesh_count(void *o) { return (int) S(o)->getMeshes().size(); }
void *ref_scene_mesh_ptr(void *o, int i) { return S(o)->getMeshes()[i]; }

/// bsdfType: noracle::BsdfType; params = {r,g,b (albedo|kd), alpha, intIOR, extIOR}
int ref_scene_mesh_info(void *o, int i, uint32_t *nV, uint32_t *nF, int *hasN, int *hasUV,
                        int *bsdfType, float *params, int *isEmitter, float *radiance) {
    const Mesh *m = S(o)->getMeshes()[i];
    *nV = m->getVertexCount(); *nF = m->getTriangleCount();
    *hasN = m->getVertexNormals().size() > 0; *hasUV = m->getVertexTexCoords().size() > 0;
    for (int k = 0; k < 6; ++k) params[k] = 0;
    const BSDF *b = m->getBSDF();
    if (const OracleBSDF *ob = dynamic_cast<const OracleBSDF *>(b)) {
        *bsdfType = ob->m_model.type;
        params[0] = ob->m_model.albedo.r; params[1] = ob->m_model.albedo.g; params[2] = ob->m_model.albedo.b;
        params[3] = ob->m_model.alpha; params[4] = ob->m_model.intIOR; params[5] = ob->m_model.extIOR;
    } else if (refDiffuseAlbedo(b, params)) *bsdfType = noracle::kDiffuse;
    else if (b->toString() == "Mirror[]") *bsdfType = noracle::kMirror;
    else { g_err = "unknown BSDF: " + b->toString(); return -1; }
    *isEmitter = m->isEmitter();
    radiance[0] = radiance[1] = radiance[2] = 0;
    if (m->isEmitter()) {
        const AreaLight *a = static_cast<const AreaLight *>(m->getEmitter());
        radiance[0] = a->m_radiance.r(); radiance[1] = a->m_radiance.g(); radiance[2] = a->m_radiance.b();
    }
    return 0;
}

/// Copies Mesh::m_V/m_N/m_UV/m_F verbatim (column-major; include/nori/mesh.h:172-175).
void ref_scene_mesh_copy(void *o, int i, float *V, float *N, float *UV, uint32_t *F) {
    const Mesh *m = S(o)->getMeshes()[i];
    std::memcpy(V, m->getVertexPositions().data(), m->getVertexPositions().size() * 4);
    if (N && m->getVertexNormals().size()) std::memcpy(N, m->getVertexNormals().data(), m->getVertexNormals().size() * 4);
    if (UV && m->getVertexTexCoords().size()) std::memcpy(UV, m->getVertexTexCoords().data(), m->getVertexTexCoords().size() * 4);
    std::memcpy(F, m->getIndices().data(), m->getIndices().size() * 4);
}

int ref_scene_camera(void *o, float *s2c, float *c2w, int *W, int *H, float *nearClip, float *farClip, float *fov) {
    const Camera *c = S(o)->getCamera();
    *W = c->getOutputSize().x(); *H = c->getOutputSize().y();
    return refCameraState(c, s2c, c2w, nearClip, farClip, fov) ? 0 : -1;
}

/// ImageBlock's tabulation (src/block.cpp:31-39) of the camera's reconstruction filter.
void ref_scene_filter(void *o, float *radius, float *table33, int *border) {
    const ReconstructionFilter *f = S(o)->getCamera()->getReconstructionFilter();
    *radius = f->getRadius();
    *border = (int) std::ceil(*radius - 0.5f);
    for (int i = 0; i < NORI_FILTER_RESOLUTION; ++i) table33[i] = f->eval((*radius * i) / NORI_FILTER_RESOLUTION);
    table33[NORI_FILTER_RESOLUTION] = 0.0f;
}

int ref_scene_sample_count(void *o) { return (int) S(o)->getSampler()->getSampleCount(); }

/// integrator: {type, rrStart, maxDepth}, fparams: {rrMax, lightPos xyz, lightEnergy rgb}
int ref_scene_integrator(void *o, int *iparams, float *fparams) {
    const OracleIntegrator *it = dynamic_cast<const OracleIntegrator *>(S(o)->getIntegrator());
    if (!it) return -1;
    const noracle::IntegratorParams &p = it->m_params;
    iparams[0] = p.type; iparams[1] = p.rrStart; iparams[2] = p.maxDepth;
    fparams[0] = p.rrMax; fparams[1] = p.lightPos.x; fparams[2] = p.lightPos.y; fparams[3] = p.lightPos.z;
    fparams[4] = p.lightEnergy.r; fparams[5] = p.lightEnergy.g; fparams[6] = p.lightEnergy.b;
    return 0;
}

// ---- REF operators -------------------------------------------------------------------------
void ref_pcg32_uints(uint64_t initstate, uint64_t initseq, int64_t advance, int n, uint32_t *out) {
    pcg32 r(initstate, initseq);
    if (advance) r.advance(advance);
    for (int i = 0; i < n; ++i) out[i] = r.nextUInt();
}
void ref_pcg32_floats(uint64_t initstate, uint64_t initseq, int64_t advance, int n, float *out) {
    pcg32 r(initstate, initseq);
    if (advance) r.advance(advance);
    for (int i = 0; i < n; ++i) out[i] = r.nextFloat();
}

/// Independent::prepare(block at offset) then n x next2D (src/independent.cpp:48-67): out n x 2.
void ref_sampler_next2d(int ox, int oy, int n, float *out) {
    std::unique_ptr<Sampler> s(static_cast<Sampler *>(NoriObjectFactory::createInstance("independent", PropertyList())));
    ImageBlock b(Vector2i(NORI_BLOCK_SIZE), nullptr);
    b.setOffset(Point2i(ox, oy));
    s->prepare(b);
    for (int i = 0; i < n; ++i) { Point2f p = s->next2D(); out[2 * i] = p.x(); out[2 * i + 1] = p.y(); }
}

/// Camera::sampleRay (src/perspective.cpp:88-109): rays n x 8 [o mint d maxt]
void ref_camera_rays(void *o, int n, const float *pixelSamples, float *rays) {
    const Camera *c = S(o)->getCamera();
    for (int i = 0; i < n; ++i) {
        Ray3f r;
        c->sampleRay(r, Point2f(pixelSamples[2 * i], pixelSamples[2 * i + 1]), Point2f(0.5f, 0.5f));
        float *p = rays + 8 * (size_t) i;
        p[0] = r.o.x(); p[1] = r.o.y(); p[2] = r.o.z(); p[3] = r.mint;
        p[4] = r.d.x(); p[5] = r.d.y(); p[6] = r.d.z(); p[7] = r.maxt;
    }
}

static Ray3f ray8(const float *p) { return Ray3f(Point3f(p[0], p[1], p[2]), Vector3f(p[4], p[5], p[6]), p[3], p[7]); }

void ref_accel_mode(int useBvh) { accelSetMode(useBvh); }
void ref_accel_counters(int enable, uint64_t *out3, int reset) { accelEnableCounters(enable); if (out3) accelCounters(out3, reset); }

/// Hit ids of the multi-mesh Accel (mode: 0 brute force over Mesh::rayIntersect, 1 BVH).
void ref_trace(void *o, const float *rays, int n, int shadow, int mode, uint32_t *tri, uint32_t *mesh, float *tuv) {
    accelSetMode(mode);
    for (int i = 0; i < n; ++i) {
        uint32_t m = 0, f = 0; float t = 0, u = 0, v = 0;
        bool hit = accelTraceRaw(S(o), ray8(rays + 8 * (size_t) i), shadow != 0, m, f, t, u, v);
        tri[i] = hit ? f : 0xffffffffu; mesh[i] = hit ? m : 0xffffffffu;
        if (tuv) { tuv[3 * i] = hit ? t : 0; tuv[3 * i + 1] = hit ? u : 0; tuv[3 * i + 2] = hit ? v : 0; }
    }
}

static void packIts(const Intersection &its, float *o) {
    const float v[24] = {its.p.x(), its.p.y(), its.p.z(), its.t, its.uv.x(), its.uv.y(),
        its.shFrame.s.x(), its.shFrame.s.y(), its.shFrame.s.z(), its.shFrame.t.x(), its.shFrame.t.y(), its.shFrame.t.z(),
        its.shFrame.n.x(), its.shFrame.n.y(), its.shFrame.n.z(),
        its.geoFrame.s.x(), its.geoFrame.s.y(), its.geoFrame.s.z(), its.geoFrame.t.x(), its.geoFrame.t.y(), its.geoFrame.t.z(),
        its.geoFrame.n.x(), its.geoFrame.n.y(), its.geoFrame.n.z()};
    std::memcpy(o, v, sizeof(v));
}

/// Scene::rayIntersect(ray, its) records: out n x 24 (layout as noracle_intersect), valid n.
void ref_intersect(void *o, const float *rays, int n, int mode, float *out, int *valid) {
    accelSetMode(mode);
    for (int i = 0; i < n; ++i) {
        Intersection its;
        valid[i] = S(o)->rayIntersect(ray8(rays + 8 * (size_t) i), its) ? 1 : 0;
        if (valid[i]) packIts(its, out + 24 * (size_t) i); else std::fill(out + 24 * (size_t) i, out + 24 * (size_t) (i + 1), 0.0f);
    }
}

/// BSDF of mesh i through the BSDF virtuals (include/nori/bsdf.h:71-112).
void ref_bsdf_sample(void *o, int mi, int n, const float *wi, const float *xi, float *wo, float *weight, float *pdf, float *eta, int *measure) {
    const BSDF *b = S(o)->getMeshes()[mi]->getBSDF();
    for (int i = 0; i < n; ++i) {
        BSDFQueryRecord q(Vector3f(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]));
        Color3f w = b->sample(q, Point2f(xi[2 * i], xi[2 * i + 1]));
        bool ok = !(w.r() == 0 && w.g() == 0 && w.b() == 0);
        for (int k = 0; k < 3; ++k) { wo[3 * i + k] = ok ? q.wo[k] : 0; weight[3 * i + k] = w[k]; }
        pdf[i] = (ok && q.measure == ESolidAngle) ? b->pdf(q) : 0.0f;
        eta[i] = ok ? q.eta : 1.0f; measure[i] = ok ? (int) q.measure : 0;
    }
}
void ref_bsdf_eval_pdf(void *o, int mi, int n, const float *wi, const float *wo, float *f, float *pdf) {
    const BSDF *b = S(o)->getMeshes()[mi]->getBSDF();
    for (int i = 0; i < n; ++i) {
        BSDFQueryRecord q(Vector3f(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), Vector3f(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), ESolidAngle);
        Color3f v = b->eval(q);
        for (int k = 0; k < 3; ++k) f[3 * i + k] = v[k];
        pdf[i] = b->pdf(q);
    }
}

void ref_fresnel(int n, const float *c, float extIOR, float intIOR, float *out) { for (int i = 0; i < n; ++i) out[i] = fresnel(c[i], extIOR, intIOR); }
/// Frame(n): out n x 6 = s, t (include/nori/frame.h:48-50, src/common.cpp:260-269)
void ref_frames(int n, const float *nrm, float *out) {
    for (int i = 0; i < n; ++i) {
        Frame f(Vector3f(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]));
        for (int k = 0; k < 3; ++k) { out[6 * i + k] = f.s[k]; out[6 * i + 3 + k] = f.t[k]; }
    }
}

/// ReconstructionFilter::eval through the factory: name in {gaussian, mitchell, tent, box}.
int ref_filter_table(const char *name, float radius, float p0, float p1, float *outRadius, float *table33) {
    GUARD({
        PropertyList pl;
        if (std::string(name) == "gaussian") { pl.setFloat("radius", radius); pl.setFloat("stddev", p0); }
        if (std::string(name) == "mitchell") { pl.setFloat("radius", radius); pl.setFloat("B", p0); pl.setFloat("C", p1); }
        std::unique_ptr<ReconstructionFilter> f(static_cast<ReconstructionFilter *>(NoriObjectFactory::createInstance(name, pl)));
        *outRadius = f->getRadius();
        for (int i = 0; i < NORI_FILTER_RESOLUTION; ++i) table33[i] = f->eval((f->getRadius() * i) / NORI_FILTER_RESOLUTION);
        table33[NORI_FILTER_RESOLUTION] = 0;
        return 0;
    }, -1)
}

/// BlockGenerator::next order (src/block.cpp:133-164): 4 ints per block (ox, oy, w, h).
int ref_block_order(int W, int H, int bs, int *out, int cap) {
    BlockGenerator gen(Vector2i(W, H), bs);
    ImageBlock b(Vector2i(bs), nullptr);
    int n = gen.getBlockCount();
    for (int i = 0; i < n && i < cap; ++i) {
        gen.next(b);
        out[4 * i] = b.getOffset().x(); out[4 * i + 1] = b.getOffset().y(); out[4 * i + 2] = b.getSize().x(); out[4 * i + 3] = b.getSize().y();
    }
    return n;
}

/// ImageBlock::put(pos, value) n times into a W x H block with filter `name` (src/block.cpp:74-103).
int ref_splat(int W, int H, const char *name, int n, const float *pos, const float *rgb, float *rgbw) {
    GUARD({
        std::unique_ptr<ReconstructionFilter> f(static_cast<ReconstructionFilter *>(NoriObjectFactory::createInstance(name, PropertyList())));
        ImageBlock b(Vector2i(W, H), f.get());
        b.clear();
        for (int i = 0; i < n; ++i) b.put(Point2f(pos[2 * i], pos[2 * i + 1]), Color3f(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]));
        std::memcpy(rgbw, b.data(), (size_t) b.rows() * b.cols() * 16);
        return 0;
    }, -1)
}

} // extern "C"
