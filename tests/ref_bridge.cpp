// ref_bridge.cpp — INTEGRATION.md sections 2-3 as compiled code (TEST INFRASTRUCTURE): the binding a maintainer of the
// reference adds to run its render path on libnori_b200.so. Built by `make -C oracle ref` against the REFERENCE's own headers
// and linked into oracle/_ref/libnori_ref.so next to the reference's unmodified parser, OBJ loader, camera, filters:
//
//     reference loadFromXML (src/parser.cpp:28)  ->  Scene / Mesh / Camera / BSDF objects of the reference
//         -> the PODs of include/nori_gpu.h (this file)  ->  nori_gpu_upload_meshes / set_camera / set_filter / build_accel
//         -> nori_gpu_render  ->  ImageBlock::data() of the reference's layout
//
// It replaces render()'s body between the BlockGenerator and toBitmap (src/main.cpp:75-127). The product library is loaded
// with dlopen at run time (the product never links oracle/; this test code loads the product, not the other way round).
// Private members the reference offers no getter for (PerspectiveCamera's matrices, Diffuse::m_albedo) are read through the
// accessor shims of oracle/ref/glue_private.cpp (reference sources unmodified).
#include "../oracle/ref/plugins.h"
#include <nori/camera.h>
#include <nori/rfilter.h>
#include <nori/block.h>
#include "../include/nori_gpu.h"
#include <dlfcn.h>
#include <cstring>
#include <string>
#include <vector>

namespace nori {
bool refCameraState(const Camera *cam, float *s2c, float *c2w, float *nearClip, float *farClip, float *fov);   // glue_private.cpp
bool refDiffuseAlbedo(const BSDF *bsdf, float *rgb);
}
using namespace nori;

namespace {
std::string g_bridgeError;

struct Api {
    void *lib = nullptr;
    decltype(&nori_gpu_create) create; decltype(&nori_gpu_destroy) destroy; decltype(&nori_gpu_last_error) last_error;
    decltype(&nori_gpu_upload_meshes) upload_meshes; decltype(&nori_gpu_set_camera) set_camera; decltype(&nori_gpu_set_filter) set_filter;
    decltype(&nori_gpu_build_accel) build_accel; decltype(&nori_gpu_render) render; decltype(&nori_gpu_download_block) download_block;
    bool load(const char *path) {
        lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
        if (!lib) { g_bridgeError = dlerror(); return false; }
#define SYM(n) n = (decltype(n)) dlsym(lib, "nori_gpu_" #n); if (!n) { g_bridgeError = "libnori_b200 lacks nori_gpu_" #n; return false; }
        SYM(create) SYM(destroy) SYM(last_error) SYM(upload_meshes) SYM(set_camera) SYM(set_filter) SYM(build_accel) SYM(render) SYM(download_block)
#undef SYM
        return true;
    }
};

/// Mesh::getBSDF() -> NORI_BSDF_* + parameters (albedo|kd rgb, alpha, intIOR, extIOR).
bool describeBsdf(const BSDF *b, int32_t &type, float params[6]) {
    for (int k = 0; k < 6; ++k) params[k] = 0;
    if (const OracleBSDF *ob = dynamic_cast<const OracleBSDF *>(b)) {        // dielectric, microfacet: the reference's own throw
        type = ob->m_model.type;
        params[0] = ob->m_model.albedo.r; params[1] = ob->m_model.albedo.g; params[2] = ob->m_model.albedo.b;
        params[3] = ob->m_model.alpha; params[4] = ob->m_model.intIOR; params[5] = ob->m_model.extIOR;
        return true;
    }
    if (refDiffuseAlbedo(b, params)) { type = NORI_BSDF_DIFFUSE; return true; }              // src/diffuse.cpp, unmodified
    if (b->toString() == "Mirror[]") { type = NORI_BSDF_MIRROR; return true; }                 // src/mirror.cpp, unmodified
    g_bridgeError = "unknown BSDF: " + b->toString();
    return false;
}
} // namespace

extern "C" const char *ref_bridge_error() { return g_bridgeError.c_str(); }

/// Renders samples [s0, s1) of every pixel of a scene loaded by the REFERENCE's parser on the device, per-sample streams.
/// rgbw: ImageBlock::data() layout, (H + 2b) x (W + 2b) x 4. Returns 0 on success.
extern "C" int ref_bridge_render(void *sceneObject, const char *libnori_b200, int device, uint32_t s0, uint32_t s1, float *rgbw, NoriGpuRenderStats *stats) {
    const Scene *scene = static_cast<const Scene *>((NoriObject *) sceneObject);
    Api api;
    if (!api.load(libnori_b200)) return -1;
    nori_gpu_ctx *ctx = nullptr;
    if (api.create(&ctx, device) != 0) { g_bridgeError = api.last_error(nullptr); return -2; }
    struct Guard { Api &a; nori_gpu_ctx *c; ~Guard() { a.destroy(c); } } guard{api, ctx};
#define GPU(call) do { if ((call) != 0) { g_bridgeError = api.last_error(ctx); return -3; } } while (0)
    // ---- the scene as PODs -----------------------------------------------------------
    const std::vector<Mesh *> &meshes = scene->getMeshes();
    std::vector<NoriGpuMeshDesc> descs(meshes.size());
    for (size_t i = 0; i < meshes.size(); ++i) {
        const Mesh *m = meshes[i];
        NoriGpuMeshDesc &d = descs[i];
        std::memset(&d, 0, sizeof(d));
        d.V = m->getVertexPositions().data();                                           // MatrixXf 3 x nV, column-major = xyz triples
        d.N = m->getVertexNormals().size() ? m->getVertexNormals().data() : nullptr;
        d.UV = m->getVertexTexCoords().size() ? m->getVertexTexCoords().data() : nullptr;
        d.F = m->getIndices().data();                                                   // MatrixXu 3 x nF
        d.nV = m->getVertexCount(); d.nF = m->getTriangleCount();
        if (!describeBsdf(m->getBSDF(), d.bsdf_type, d.bsdf_params)) return -4;
        d.is_emitter = m->isEmitter();
        if (m->isEmitter()) { const AreaLight *a = static_cast<const AreaLight *>(m->getEmitter()); d.radiance[0] = a->m_radiance.r(); d.radiance[1] = a->m_radiance.g(); d.radiance[2] = a->m_radiance.b(); }
    }
    GPU(api.upload_meshes(ctx, descs.data(), (int) descs.size()));
    const Camera *camera = scene->getCamera();
    NoriGpuCameraDesc cam; float fov;
    if (!refCameraState(camera, cam.sample_to_camera, cam.camera_to_world, &cam.near_clip, &cam.far_clip, &fov)) { g_bridgeError = "camera is not a PerspectiveCamera"; return -5; }
    cam.width = camera->getOutputSize().x(); cam.height = camera->getOutputSize().y();
    GPU(api.set_camera(ctx, &cam));
    const ReconstructionFilter *f = camera->getReconstructionFilter();
    float table[NORI_FILTER_RESOLUTION + 1];                                            // ImageBlock's tabulation, src/block.cpp:33-39
    for (int i = 0; i < NORI_FILTER_RESOLUTION; ++i) table[i] = f->eval((f->getRadius() * i) / NORI_FILTER_RESOLUTION);
    table[NORI_FILTER_RESOLUTION] = 0.0f;
    GPU(api.set_filter(ctx, f->getRadius(), table));
    GPU(api.build_accel(ctx, nullptr));                                                 // Accel::build (src/scene.cpp:43)
    // ---- section 3: the block loop of render() replaced by one call ------------------------------------------
    const OracleIntegrator *it = dynamic_cast<const OracleIntegrator *>(scene->getIntegrator());
    if (!it) { g_bridgeError = "integrator has no device counterpart"; return -6; }
    NoriGpuRenderDesc rd; std::memset(&rd, 0, sizeof(rd));
    rd.integrator = it->m_params.type; rd.stream_mode = NORI_STREAM_PER_SAMPLE; rd.sample_first = s0; rd.sample_last = s1;
    rd.rr_start = it->m_params.rrStart; rd.rr_max = it->m_params.rrMax; rd.max_depth = it->m_params.maxDepth; rd.intersect_mode = NORI_ISECT_REFERENCE_MT;
    rd.light_pos[0] = it->m_params.lightPos.x; rd.light_pos[1] = it->m_params.lightPos.y; rd.light_pos[2] = it->m_params.lightPos.z;
    rd.light_energy[0] = it->m_params.lightEnergy.r; rd.light_energy[1] = it->m_params.lightEnergy.g; rd.light_energy[2] = it->m_params.lightEnergy.b;
    rd.clear = 1;
    NoriGpuRenderStats st;
    GPU(api.render(ctx, &rd, &st));
    if (stats) *stats = st;
    ImageBlock result(camera->getOutputSize(), f);                                      // same (H + 2b) x (W + 2b) RGBW layout (src/block.cpp:48)
    GPU(api.download_block(ctx, (float *) result.data()));
    std::memcpy(rgbw, result.data(), (size_t) result.rows() * result.cols() * 16);
#undef GPU
    return 0;
}
