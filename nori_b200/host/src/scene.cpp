// scene.cpp — Scene, ImageBlock and the bridge from the object graph to the C ABI.
// Reference: src/scene.cpp:26-100 (child bookkeeping, defaults), src/block.cpp:27-63 (filter table,
// toBitmap), src/main.cpp:69-151 (the render() this replaces).
#include <nori/scene.h>
#include <chrono>
#include <nori_gpu.h>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <thread>
#include <fstream>
#include <iostream>

NORI_NAMESPACE_BEGIN

// ---- file resolver ---------------------------------------------------------------------------------
std::string FileResolver::resolve(const std::string &name) const {
    if (!name.empty() && name[0] == '/') return name;
    for (const std::string &dir : m_paths) {
        const std::string candidate = dir.empty() ? name : dir + "/" + name;
        if (std::ifstream(candidate).good()) return candidate;
    }
    return name;
}
FileResolver *getFileResolver() { static FileResolver *r = new FileResolver(); return r; }

// ---- errors: C codes -> NoriException ----------------------------------------------------------------
void gpuCheck(nori_gpu_ctx *ctx, int rc, const char *what) {
    if (rc != NORI_GPU_OK) throw NoriException("%s: %s", what, nori_gpu_last_error(ctx));
}
static void check(nori_gpu_ctx *ctx, int rc, const char *what) { gpuCheck(ctx, rc, what); }

// ---- Scene -------------------------------------------------------------------------------------------
Scene::Scene(const PropertyList &) {}

Scene::~Scene() {
    for (auto &kv : m_ctx) nori_gpu_destroy(kv.second);
    delete m_sampler; delete m_camera; delete m_integrator;
    for (Mesh *m : m_meshes) delete m;                       // (the reference leaks its meshes)
}

void Scene::addChild(NoriObject *obj) {
    switch (obj->getClassType()) {
        case EMesh: m_meshes.push_back(static_cast<Mesh *>(obj)); break;
        case EEmitter: throw NoriException("Scene::addChild(): emitters must be attached to a mesh");
        case ESampler:
            if (m_sampler) throw NoriException("There can only be one sampler per scene!");
            m_sampler = static_cast<Sampler *>(obj); break;
        case ECamera:
            if (m_camera) throw NoriException("There can only be one camera per scene!");
            m_camera = static_cast<Camera *>(obj); break;
        case EIntegrator:
            if (m_integrator) throw NoriException("There can only be one integrator per scene!");
            m_integrator = static_cast<Integrator *>(obj); break;
        default: throw NoriException("Scene::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
    }
}

void Scene::activate() {
    if (!m_integrator) throw NoriException("No integrator was specified!");
    if (!m_camera) throw NoriException("No camera was specified!");
    if (!m_sampler) m_sampler = static_cast<Sampler *>(NoriObjectFactory::createInstance("independent", PropertyList()));
    std::cout << std::endl << "Configuration: " << toString() << std::endl << std::endl;
}

std::string Scene::toString() const {
    std::string meshes;
    for (size_t i = 0; i < m_meshes.size(); ++i) {
        meshes += std::string("  ") + indent(m_meshes[i]->toString(), 2);
        if (i + 1 < m_meshes.size()) meshes += ",";
        meshes += "\n";
    }
    return "Scene[\n  integrator = " + indent(m_integrator->toString()) + ",\n  sampler = " + indent(m_sampler->toString())
         + "\n  camera = " + indent(m_camera->toString()) + ",\n  meshes = {\n  " + indent(meshes) + "  }\n]";
}

nori_gpu_ctx *Scene::gpu(int device) const {
    static std::mutex lock;                                   // renderGpuMulti: one host thread per device
    {
        std::lock_guard<std::mutex> g(lock);
        auto it = m_ctx.find(device);
        if (it != m_ctx.end()) return it->second;
    }
    nori_gpu_ctx *ctx = nullptr;
    if (nori_gpu_create(&ctx, device) != NORI_GPU_OK)
        throw NoriException("nori_gpu_create: %s (nori_b200 has no CPU path)", nori_gpu_last_error(nullptr));
    try {
        std::vector<NoriGpuMeshDesc> descs(m_meshes.size());
        for (size_t i = 0; i < m_meshes.size(); ++i) {
            const Mesh *m = m_meshes[i];
            NoriGpuMeshDesc &d = descs[i];
            memset(&d, 0, sizeof(d));
            d.V = m->getVertexPositions().data();
            d.N = m->getVertexNormals().empty() ? nullptr : m->getVertexNormals().data();
            d.UV = m->getVertexTexCoords().empty() ? nullptr : m->getVertexTexCoords().data();
            d.F = m->getIndices().data();
            d.nV = m->getVertexCount(); d.nF = m->getTriangleCount();
            m->getBSDF()->describe(d.bsdf_type, d.bsdf_params);
            d.is_emitter = m->isEmitter() ? 1 : 0;
            if (m->isEmitter()) { const Color3f r = m->getEmitter()->getRadiance(); d.radiance[0] = r.r; d.radiance[1] = r.g; d.radiance[2] = r.b; }
        }
        check(ctx, nori_gpu_upload_meshes(ctx, descs.data(), (int) descs.size()), "nori_gpu_upload_meshes");   // Scene::addChild + Accel::addMesh
        NoriGpuCameraDesc cam;
        memset(&cam, 0, sizeof(cam));
        float fov;
        m_camera->describe(cam.sample_to_camera, cam.camera_to_world, cam.near_clip, cam.far_clip, fov);
        cam.width = m_camera->getOutputSize().x; cam.height = m_camera->getOutputSize().y;
        check(ctx, nori_gpu_set_camera(ctx, &cam), "nori_gpu_set_camera");
        const ImageBlock table(m_camera->getOutputSize(), m_camera->getReconstructionFilter());
        check(ctx, nori_gpu_set_filter(ctx, table.filterRadius(), table.filterTable()), "nori_gpu_set_filter");
        NoriGpuAccelStats st;
        check(ctx, nori_gpu_build_accel(ctx, &st), "nori_gpu_build_accel");                                       // Accel::build (src/scene.cpp:41)
        std::cout << "Built the device BVH: " << st.n_triangles << " triangles, " << st.n_nodes << " nodes, depth " << st.depth
                  << ", SAH cost " << st.sah_cost << " (took " << timeString(st.build_ms, true) << ")" << std::endl;
    } catch (...) {
        nori_gpu_destroy(ctx);
        throw;
    }
    std::lock_guard<std::mutex> g(lock);
    m_ctx[device] = ctx;
    return ctx;
}

bool Scene::rayIntersect(const Ray3f &ray, Intersection &its) const {
    nori_gpu_ctx *ctx = gpu();
    const float r[8] = {ray.o.x, ray.o.y, ray.o.z, ray.mint, ray.d.x, ray.d.y, ray.d.z, ray.maxt};
    float out[24]; int valid = 0; uint32_t tri = 0, mesh = 0; float tuv[3];
    check(ctx, nori_gpu_intersect(ctx, r, 1, out, &valid), "nori_gpu_intersect");
    if (!valid) return false;
    check(ctx, nori_gpu_trace(ctx, r, 1, 0, &tri, &mesh, tuv), "nori_gpu_trace");
    // out: p(3) t(1) uv(2) shading frame s,t,n (9) geometric frame s,t,n (9)
    its.p = Point3f(out[0], out[1], out[2]); its.t = out[3]; its.uv = Point2f(out[4], out[5]);
    its.shS = Vector3f(out[6], out[7], out[8]); its.shT = Vector3f(out[9], out[10], out[11]); its.shN = Vector3f(out[12], out[13], out[14]);
    its.geoS = Vector3f(out[15], out[16], out[17]); its.geoT = Vector3f(out[18], out[19], out[20]); its.geoN = Vector3f(out[21], out[22], out[23]);
    its.mesh = mesh < m_meshes.size() ? m_meshes[mesh] : nullptr;
    return true;
}

bool Scene::rayIntersect(const Ray3f &ray) const {
    nori_gpu_ctx *ctx = gpu();
    const float r[8] = {ray.o.x, ray.o.y, ray.o.z, ray.mint, ray.d.x, ray.d.y, ray.d.z, ray.maxt};
    uint32_t tri = 0, mesh = 0; float tuv[3];
    check(ctx, nori_gpu_trace(ctx, r, 1, 1, &tri, &mesh, tuv), "nori_gpu_trace");
    return tri != 0xffffffffu;
}

// ---- BSDF operators: evaluated by the device code the renderer uses ------------------------------------
static bool g_deviceObjDedup = false;
bool deviceObjDedup() { return g_deviceObjDedup; }
void setDeviceObjDedup(bool on) { g_deviceObjDedup = on; }

nori_gpu_ctx *operatorContext() {
    static nori_gpu_ctx *ctx = nullptr;
    if (!ctx && nori_gpu_create(&ctx, 0) != NORI_GPU_OK)
        throw NoriException("nori_gpu_create: %s (nori_b200 has no CPU path)", nori_gpu_last_error(nullptr));
    return ctx;
}

Color3f BSDF::sample(BSDFQueryRecord &bRec, const Point2f &sample) const {
    int type; float params[6];
    describe(type, params);
    const float wi[3] = {bRec.wi.x, bRec.wi.y, bRec.wi.z}, xi[2] = {sample.x, sample.y};
    float wo[3], weight[3], pdf, eta; int measure;
    nori_gpu_ctx *ctx = operatorContext();
    check(ctx, nori_gpu_bsdf_sample(ctx, type, params, 1, wi, xi, wo, weight, &pdf, &eta, &measure), "nori_gpu_bsdf_sample");
    bRec.wo = Vector3f(wo[0], wo[1], wo[2]); bRec.eta = eta; bRec.measure = (EMeasure) measure;
    return Color3f(weight[0], weight[1], weight[2]);
}

Color3f BSDF::eval(const BSDFQueryRecord &bRec) const {
    if (bRec.measure != ESolidAngle) return Color3f(0.0f);
    int type; float params[6];
    describe(type, params);
    const float wi[3] = {bRec.wi.x, bRec.wi.y, bRec.wi.z}, wo[3] = {bRec.wo.x, bRec.wo.y, bRec.wo.z};
    float f[3], pdf;
    nori_gpu_ctx *ctx = operatorContext();
    check(ctx, nori_gpu_bsdf_eval_pdf(ctx, type, params, 1, wi, wo, f, &pdf), "nori_gpu_bsdf_eval_pdf");
    return Color3f(f[0], f[1], f[2]);
}

float BSDF::pdf(const BSDFQueryRecord &bRec) const {
    if (bRec.measure != ESolidAngle) return 0.0f;
    int type; float params[6];
    describe(type, params);
    const float wi[3] = {bRec.wi.x, bRec.wi.y, bRec.wi.z}, wo[3] = {bRec.wo.x, bRec.wo.y, bRec.wo.z};
    float f[3], pdf;
    nori_gpu_ctx *ctx = operatorContext();
    check(ctx, nori_gpu_bsdf_eval_pdf(ctx, type, params, 1, wi, wo, f, &pdf), "nori_gpu_bsdf_eval_pdf");
    return pdf;
}

NORI_REGISTER_CLASS(Scene, "scene")

// ---- ImageBlock ------------------------------------------------------------------------------------------
ImageBlock::ImageBlock(const Vector2i &size, const ReconstructionFilter *filter) : m_size(size) {
    for (float &t : m_table) t = 0.0f;
    if (filter) {
        m_radius = filter->getRadius();
        m_border = (int) std::ceil(m_radius - 0.5f);
        for (int i = 0; i < NORI_FILTER_RESOLUTION; ++i) m_table[i] = filter->eval((m_radius * i) / NORI_FILTER_RESOLUTION);
        m_table[NORI_FILTER_RESOLUTION] = 0.0f;
    }
    m_data.assign((size_t) rows() * cols() * 4, 0.0f);
}

std::unique_ptr<Bitmap> ImageBlock::toBitmap() const {
    std::unique_ptr<Bitmap> bmp(new Bitmap(m_size.x, m_size.y));
    for (int y = 0; y < m_size.y; ++y)
        for (int x = 0; x < m_size.x; ++x) {
            const float *c = &m_data[4 * ((size_t) (y + m_border) * cols() + (x + m_border))];
            float *o = bmp->data() + 3 * ((size_t) y * m_size.x + x);
            if (c[3] != 0) { o[0] = c[0] / c[3]; o[1] = c[1] / c[3]; o[2] = c[2] / c[3]; }   // Color4f::divideByFilterWeight
            else o[0] = o[1] = o[2] = 0.0f;
        }
    return bmp;
}

// ---- render ------------------------------------------------------------------------------------------------
static NoriGpuRenderDesc renderDesc(const Scene *scene, uint32_t sampleFirst, uint32_t sampleLast) {
    const Integrator *integrator = scene->getIntegrator();
    if (!integrator) throw NoriException("No integrator was specified!");
    NoriGpuRenderDesc d;
    memset(&d, 0, sizeof(d));
    d.integrator = integrator->gpuType();
    d.stream_mode = NORI_STREAM_PER_SAMPLE;
    d.sample_first = sampleFirst; d.sample_last = sampleLast;
    d.rr_start = integrator->rrStart(); d.rr_max = integrator->rrMax(); d.max_depth = integrator->maxDepth();
    d.intersect_mode = integrator->intersectMode();
    d.clear = 1;
    integrator->describe(d.light_pos, d.light_energy);
    return d;
}

Color3f Integrator::Li(const Scene *scene, Sampler *sampler, const Ray3f &ray) const {
    if (!scene || !sampler) throw NoriException("Integrator::Li: scene and sampler are required");
    if (scene->getIntegrator() != this) throw NoriException("Integrator::Li: the device evaluates the scene's own integrator");
    uint64_t st[2];
    if (!sampler->getStreamState(st[0], st[1])) throw NoriException("Integrator::Li: this sampler has no pcg32 stream the device could continue");
    nori_gpu_ctx *ctx = scene->gpu(0);
    const NoriGpuRenderDesc d = renderDesc(scene, 0, 1);
    const float r[8] = {ray.o.x, ray.o.y, ray.o.z, ray.mint, ray.d.x, ray.d.y, ray.d.z, ray.maxt};
    float rgb[3];
    check(ctx, nori_gpu_li(ctx, &d, 1, r, st, rgb), "nori_gpu_li");
    sampler->setStreamState(st[0], st[1]);
    return Color3f(rgb[0], rgb[1], rgb[2]);
}

void sampleRadianceGpu(const Scene *scene, int n, const uint32_t *pixel, const uint32_t *sample, float *rgb, int device) {
    nori_gpu_ctx *ctx = scene->gpu(device);
    const NoriGpuRenderDesc d = renderDesc(scene, 0, 1);
    // a path pool sized to the request (the default pool is sized for full-frame renders)
    const uint32_t slots = (uint32_t) std::min<uint64_t>(1u << 20, std::max<uint64_t>(4096, ((uint64_t) n + 255) / 256 * 256));
    check(ctx, nori_gpu_set_pool_size(ctx, slots), "nori_gpu_set_pool_size");
    check(ctx, nori_gpu_sample_radiance(ctx, &d, n, pixel, sample, rgb, nullptr), "nori_gpu_sample_radiance");
}

void renderGpuMulti(const Scene *scene, ImageBlock &result, uint32_t sampleFirst, uint32_t sampleLast, const std::vector<int> &devices,
                    RenderReport *report) {
    const int n = (int) devices.size();
    if (n < 1) throw NoriException("renderGpuMulti: no devices");
    const uint32_t total = sampleLast - sampleFirst;
    if ((uint32_t) n > total) throw NoriException("renderGpuMulti: %d devices for %d samples per pixel", n, (int) total);
    std::vector<nori_gpu_ctx *> ctxs((size_t) n);
    std::vector<NoriGpuRenderStats> stats((size_t) n);
    std::vector<std::string> errors((size_t) n);
    std::vector<std::thread> workers;
    for (int k = 0; k < n; ++k)
        workers.emplace_back([&, k]() {
            try {
                nori_gpu_ctx *ctx = scene->gpu(devices[(size_t) k]);                      // upload + accel build per device (replicated scene)
                ctxs[(size_t) k] = ctx;
                const uint32_t s0 = sampleFirst + (uint32_t) ((uint64_t) total * k / n), s1 = sampleFirst + (uint32_t) ((uint64_t) total * (k + 1) / n);
                const NoriGpuRenderDesc d = renderDesc(scene, s0, s1);
                check(ctx, nori_gpu_render(ctx, &d, &stats[(size_t) k]), "nori_gpu_render");
            } catch (const std::exception &e) { errors[(size_t) k] = e.what(); }
        });
    for (std::thread &w : workers) w.join();
    for (const std::string &e : errors) if (!e.empty()) throw NoriException("%s", e);
    check(ctxs[0], nori_gpu_reduce_blocks(ctxs.data(), n, 0), "nori_gpu_reduce_blocks");   // result.put(block) across devices
    check(ctxs[0], nori_gpu_download_block(ctxs[0], result.data()), "nori_gpu_download_block");
    if (report) {
        *report = RenderReport();
        uint64_t samples = 0, rays = 0, segs = 0;
        for (const NoriGpuRenderStats &st : stats) {
            report->ms = std::max(report->ms, (double) st.ms_total);                        // devices run side by side
            samples += st.camera_samples; rays += st.closest_rays + st.shadow_rays; segs += st.segments;
            report->dropped += st.dropped; report->launches += st.kernel_launches;
        }
        if (report->ms > 0) { report->msamples = samples / report->ms / 1e3; report->mrays = rays / report->ms / 1e3; report->gsegments = segs / report->ms / 1e6; }
    }
}

void saveCheckpoint(const std::string &path, const ImageBlock &block, uint32_t sampleFirst, uint32_t sampleLast) {
    std::ofstream os(path, std::ios::binary);
    if (os.fail()) throw NoriException("Unable to open \"%s\" for writing", path);
    const int32_t dims[3] = {block.rows(), block.cols(), block.getBorderSize()};
    const uint32_t range[2] = {sampleFirst, sampleLast};
    os.write("NB2BLK1", 8); os.write((const char *) dims, sizeof(dims)); os.write((const char *) range, sizeof(range));
    os.write((const char *) block.data(), (std::streamsize) ((size_t) block.rows() * block.cols() * 4 * sizeof(float)));
}

void loadCheckpoint(const std::string &path, ImageBlock &block, uint32_t &sampleFirst, uint32_t &sampleLast) {
    std::ifstream is(path, std::ios::binary);
    char magic[8]; int32_t dims[3]; uint32_t range[2];
    if (is.fail() || !is.read(magic, 8) || memcmp(magic, "NB2BLK1", 8) != 0 || !is.read((char *) dims, sizeof(dims)) || !is.read((char *) range, sizeof(range)))
        throw NoriException("\"%s\" is not an ImageBlock checkpoint", path);
    if (dims[0] != block.rows() || dims[1] != block.cols() || dims[2] != block.getBorderSize())
        throw NoriException("checkpoint \"%s\" holds a %dx%d block (border %d); the camera needs %dx%d (border %d)", path, dims[1], dims[0], dims[2],
                            block.cols(), block.rows(), block.getBorderSize());
    if (!is.read((char *) block.data(), (std::streamsize) ((size_t) block.rows() * block.cols() * 4 * sizeof(float))))
        throw NoriException("checkpoint \"%s\" is truncated", path);
    sampleFirst = range[0]; sampleLast = range[1];
}

void renderGpu(const Scene *scene, ImageBlock &result, uint32_t sampleFirst, uint32_t sampleLast, int device, RenderReport *report,
               const ImageBlock *resumeFrom) {
    nori_gpu_ctx *ctx = scene->gpu(device);
    NoriGpuRenderDesc d = renderDesc(scene, sampleFirst, sampleLast);
    if (resumeFrom) {
        if (resumeFrom->rows() != result.rows() || resumeFrom->cols() != result.cols()) throw NoriException("renderGpu: the checkpoint does not match the camera");
        check(ctx, nori_gpu_upload_block(ctx, resumeFrom->data(), 0), "nori_gpu_upload_block");
        d.clear = 0;
    }
    NoriGpuRenderStats st;
    check(ctx, nori_gpu_render(ctx, &d, &st), "nori_gpu_render");           // replaces tbb::parallel_for(range, map)
    int rows = 0, cols = 0, border = 0;
    check(ctx, nori_gpu_block_dims(ctx, &rows, &cols, &border), "nori_gpu_block_dims");
    if (rows != result.rows() || cols != result.cols()) throw NoriException("renderGpu: the image block does not match the camera");
    check(ctx, nori_gpu_download_block(ctx, result.data()), "nori_gpu_download_block");
    if (report) {
        report->ms = st.ms_total;
        report->msamples = st.ms_total > 0 ? st.camera_samples / st.ms_total / 1e3 : 0;
        report->mrays = st.ms_total > 0 ? (st.closest_rays + st.shadow_rays) / st.ms_total / 1e3 : 0;
        report->gsegments = st.ms_total > 0 ? st.segments / st.ms_total / 1e6 : 0;
        report->dropped = st.dropped; report->launches = st.kernel_launches;
    }
}

PathGraphImages pathGraphGpu(const Scene *scene, uint32_t sampleFirst, uint32_t sampleLast, int k, int iterations, int mode, int device) {
    if (sampleLast <= sampleFirst) throw NoriException("pathGraphGpu: empty sample range");
    if (k < 1 || iterations < 1) throw NoriException("pathGraphGpu: k and the number of iterations must be positive");
    nori_gpu_ctx *ctx = scene->gpu(device);
    const NoriGpuRenderDesc d = renderDesc(scene, sampleFirst, sampleLast);
    const Vector2i size = scene->getCamera()->getOutputSize();
    const size_t plane = (size_t) size.x * size.y * 3;
    if (mode != NORI_PG_SCATTER && mode != NORI_PG_MIS && mode != NORI_PG_CLUSTER) throw NoriException("pathGraphGpu: unknown mode %d", mode);
    const int nImages = mode == NORI_PG_CLUSTER ? 6 : 5;
    std::vector<float> images(nImages * plane);
    NoriGpuPathGraphStats st;
    // records, grid, neighbours or clusters, scatter / gather and the images in one call: nothing but the images leaves the device
    check(ctx, nori_gpu_pathgraph_render(ctx, &d, mode, k, iterations, images.data(), &st), "nori_gpu_pathgraph_render");
    PathGraphImages out;
    out.vertices = st.vertices; out.msVertices = st.ms_vertices; out.msKnn = st.ms_knn; out.msScatter = st.ms_stage;
    std::unique_ptr<Bitmap> *slots[6] = {&out.full, &out.indirect, &out.pt, &out.blur, &out.init, &out.direct};
    for (int i = 0; i < nImages; ++i) {
        slots[i]->reset(new Bitmap(size.x, size.y));
        std::copy(images.begin() + (std::ptrdiff_t) (i * plane), images.begin() + (std::ptrdiff_t) ((i + 1) * plane), (*slots[i])->data());
    }
    return out;
}

NORI_NAMESPACE_END
