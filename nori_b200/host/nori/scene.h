// nori/scene.h — the scene-graph classes of the host layer: same class names, XML names,
// properties and defaults as the reference; the numerical work behind their operators is done by
// libnori_b200.so (include/nori_gpu.h).
//
// Reference: include/nori/{scene,mesh,bsdf,emitter,camera,rfilter,sampler,integrator,block,bitmap}.h
#pragma once
#include "object.h"
#include <map>
#include <memory>

struct nori_gpu_ctx;

NORI_NAMESPACE_BEGIN

class Mesh; class BSDF; class Emitter; class Scene; class Sampler;

// ---- reconstruction filters (include/nori/rfilter.h:34-52, src/rfilter.cpp) ---------------------
#define NORI_FILTER_RESOLUTION 32
#define NORI_BLOCK_SIZE 32          /* include/nori/block.h:29 */
class ReconstructionFilter : public NoriObject {
public:
    float getRadius() const { return m_radius; }
    virtual float eval(float x) const = 0;
    EClassType getClassType() const override { return EReconstructionFilter; }
protected:
    float m_radius = 0;
};

// ---- BSDFs (include/nori/bsdf.h:29-112) ---------------------------------------------------------
enum EMeasure { EUnknownMeasure = 0, ESolidAngle, EDiscrete };

struct BSDFQueryRecord {
    Vector3f wi, wo;
    float eta = 1.0f;
    EMeasure measure = EUnknownMeasure;
    explicit BSDFQueryRecord(const Vector3f &wi_) : wi(wi_) {}
    BSDFQueryRecord(const Vector3f &wi_, const Vector3f &wo_, EMeasure m) : wi(wi_), wo(wo_), measure(m) {}
};

class BSDF : public NoriObject {
public:
    /// Device-evaluated (nori_gpu_bsdf_sample / nori_gpu_bsdf_eval_pdf); arguments as in the reference.
    Color3f sample(BSDFQueryRecord &bRec, const Point2f &sample) const;
    Color3f eval(const BSDFQueryRecord &bRec) const;
    float pdf(const BSDFQueryRecord &bRec) const;
    virtual bool isDiffuse() const { return false; }
    EClassType getClassType() const override { return EBSDF; }
    /// Tagged POD view for NoriGpuMeshDesc: NORI_BSDF_* and {albedo|kd rgb, alpha, intIOR, extIOR}
    virtual void describe(int &type, float params[6]) const = 0;
};

// ---- emitters (include/nori/emitter.h:28-36 is an empty base; `area` is defined by the scenes) ---
class Emitter : public NoriObject {
public:
    EClassType getClassType() const override { return EEmitter; }
    virtual Color3f getRadiance() const = 0;
};

// ---- meshes (include/nori/mesh.h:35-178) ----------------------------------------------------------
struct Intersection {
    Point3f p; float t = 0; Point2f uv;
    Vector3f shS, shT, shN, geoS, geoT, geoN;      // shading and geometric frames
    const Mesh *mesh = nullptr;
};

class Mesh : public NoriObject {
public:
    ~Mesh() override;
    uint32_t getTriangleCount() const { return (uint32_t) (m_F.size() / 3); }
    uint32_t getVertexCount() const { return (uint32_t) (m_V.size() / 3); }
    /// 3 x nV positions, column-major (== nV consecutive xyz triples), as Mesh::m_V
    const std::vector<float> &getVertexPositions() const { return m_V; }
    const std::vector<float> &getVertexNormals() const { return m_N; }
    const std::vector<float> &getVertexTexCoords() const { return m_UV; }
    const std::vector<uint32_t> &getIndices() const { return m_F; }
    bool isEmitter() const { return m_emitter != nullptr; }
    const Emitter *getEmitter() const { return m_emitter; }
    const BSDF *getBSDF() const { return m_bsdf; }
    const std::string &getName() const { return m_name; }
    void getBoundingBox(Point3f &lo, Point3f &hi) const { lo = m_lo; hi = m_hi; }
    void addChild(NoriObject *child) override;
    void activate() override;                        // default diffuse BSDF (src/mesh.cpp:35-41)
    std::string toString() const override;
    EClassType getClassType() const override { return EMesh; }
protected:
    std::string m_name;
    std::vector<float> m_V, m_N, m_UV;
    std::vector<uint32_t> m_F;
    BSDF *m_bsdf = nullptr;
    Emitter *m_emitter = nullptr;
    Point3f m_lo = Point3f(INFINITY), m_hi = Point3f(-INFINITY);
};

// ---- camera (include/nori/camera.h:34-72, src/perspective.cpp) -------------------------------------
class Camera : public NoriObject {
public:
    const Vector2i &getOutputSize() const { return m_outputSize; }
    const ReconstructionFilter *getReconstructionFilter() const { return m_rfilter; }
    EClassType getClassType() const override { return ECamera; }
    /// include/nori/camera.h:56-58 — importance-sample a ray through the film position `samplePosition` (pixels)
    virtual Color3f sampleRay(Ray3f &ray, const Point2f &samplePosition, const Point2f &apertureSample) const = 0;
    /// State after activate(): the two matrices and clip planes NoriGpuCameraDesc carries
    virtual void describe(float sampleToCamera[16], float cameraToWorld[16], float &nearClip, float &farClip, float &fov) const = 0;
protected:
    Vector2i m_outputSize;
    ReconstructionFilter *m_rfilter = nullptr;
};

// ---- sampler (include/nori/sampler.h:45-104, src/independent.cpp) ------------------------------------
class ImageBlock;
class Sampler : public NoriObject {
public:
    /// Exact copy, random number state included (one per worker in the reference: src/main.cpp:102)
    virtual std::unique_ptr<Sampler> clone() const = 0;
    /// Deterministic re-seed for a new image block (src/independent.cpp:48-53: pcg32::seed(offset.x, offset.y))
    virtual void prepare(const ImageBlock &block) = 0;
    virtual void generate() = 0;
    virtual void advance() = 0;
    virtual float next1D() = 0;
    virtual Point2f next2D() = 0;
    virtual size_t getSampleCount() const { return m_sampleCount; }
    EClassType getClassType() const override { return ESampler; }
    /// Device bridge: the pcg32 (state, inc) the sampler draws from, so that Integrator::Li can continue the SAME stream on the
    /// GPU and hand it back advanced. A sampler without such a state returns false (Li then throws).
    virtual bool getStreamState(uint64_t &state, uint64_t &inc) const { (void) state; (void) inc; return false; }
    virtual void setStreamState(uint64_t state, uint64_t inc) { (void) state; (void) inc; }
protected:
    size_t m_sampleCount = 1;
};

// ---- integrators (include/nori/integrator.h:30-66; the reference ships no subclass) ----------------
class Integrator : public NoriObject {
public:
    EClassType getClassType() const override { return EIntegrator; }
    virtual void preprocess(const Scene *) {}
    /// include/nori/integrator.h:54 — radiance arriving along `ray`, drawing from `sampler`. One ray through the device
    /// operator nori_gpu_li: the sampler's pcg32 state goes in and comes back advanced by exactly the numbers the path drew,
    /// so a loop written against the reference (src/main.cpp:49-62, src/ttest.cpp:168-175) produces the same sequence of
    /// values as the CPU integrator would. (Batches: sampleRadianceGpu / renderGpu.)
    virtual Color3f Li(const Scene *scene, Sampler *sampler, const Ray3f &ray) const;
    /// NORI_INT_* of include/nori_gpu.h
    virtual int gpuType() const = 0;
    /// Extra parameters some integrators carry (`simple`: position, energy)
    virtual void describe(float lightPos[3], float lightEnergy[3]) const { for (int i = 0; i < 3; ++i) lightPos[i] = lightEnergy[i] = 0; }
    /// Path termination: Russian roulette from bounce `rrStart` with survival min(max(T) eta^2, rrMax);
    /// maxDepth 0 = unbounded (properties rrStart / rrMax / maxDepth, as oracle/ref/plugins.cpp)
    int rrStart() const { return m_rrStart; }
    float rrMax() const { return m_rrMax; }
    int maxDepth() const { return m_maxDepth; }
    /// NORI_ISECT_* (property `watertight`, default false = the reference's Moeller-Trumbore, src/mesh.cpp:51-88)
    int intersectMode() const { return m_watertight ? 1 : 0; }
protected:
    void readTermination(const PropertyList &p) {
        m_rrStart = p.getInteger("rrStart", 3); m_rrMax = p.getFloat("rrMax", 0.99f); m_maxDepth = p.getInteger("maxDepth", 0);
        m_watertight = p.getBoolean("watertight", false);
    }
    int m_rrStart = 3; float m_rrMax = 0.99f; int m_maxDepth = 0; bool m_watertight = false;
};

// ---- image block and bitmap (include/nori/block.h:44-113, include/nori/bitmap.h:30-58) ---------------
/// Interleaved RGB float image; saveEXR / savePNG as src/bitmap.cpp:81-134.
class Bitmap {
public:
    Bitmap(int width = 0, int height = 0) : m_width(width), m_height(height), m_data((size_t) width * height * 3, 0.0f) {}
    int cols() const { return m_width; }
    int rows() const { return m_height; }
    float *data() { return m_data.data(); }
    const float *data() const { return m_data.data(); }
    Color3f coeff(int y, int x) const { const float *p = &m_data[3 * ((size_t) y * m_width + x)]; return Color3f(p[0], p[1], p[2]); }
    /// 3-channel FLOAT scanline OpenEXR, uncompressed; appends ".exr" like the reference
    void saveEXR(const std::string &filename) const;
    /// 8-bit sRGB PNG (Color3f::toSRGB, clamp); appends ".png"
    void savePNG(const std::string &filename) const;
private:
    int m_width, m_height;
    std::vector<float> m_data;
};

/// The full-frame accumulator as the device keeps it: (H + 2b) x (W + 2b) x RGBW, row-major.
class ImageBlock {
public:
    ImageBlock(const Vector2i &size, const ReconstructionFilter *filter);
    const Vector2i &getSize() const { return m_size; }
    /// Offset of the block inside the frame (include/nori/block.h:62-66); Sampler::prepare seeds from it
    const Point2i &getOffset() const { return m_offset; }
    void setOffset(const Point2i &offset) { m_offset = offset; }
    int getBorderSize() const { return m_border; }
    int rows() const { return m_size.y + 2 * m_border; }
    int cols() const { return m_size.x + 2 * m_border; }
    float *data() { return m_data.data(); }
    const float *data() const { return m_data.data(); }
    /// 33-entry filter table and radius exactly as src/block.cpp:31-39 builds them
    const float *filterTable() const { return m_table; }
    float filterRadius() const { return m_radius; }
    /// rgb / w per pixel, border cropped (src/block.cpp:57-63, include/nori/color.h:112-117)
    std::unique_ptr<Bitmap> toBitmap() const;
private:
    Vector2i m_size; Point2i m_offset; int m_border = 0; float m_radius = 0; float m_table[NORI_FILTER_RESOLUTION + 1];
    std::vector<float> m_data;
};

// ---- scene (include/nori/scene.h:34-127, src/scene.cpp) -----------------------------------------------
class Scene : public NoriObject {
public:
    explicit Scene(const PropertyList &);
    ~Scene() override;
    const Integrator *getIntegrator() const { return m_integrator; }
    const Camera *getCamera() const { return m_camera; }
    const Sampler *getSampler() const { return m_sampler; }
    const std::vector<Mesh *> &getMeshes() const { return m_meshes; }
    /// Accel::rayIntersect through the device BVH (one ray; batches go through nori_gpu_trace)
    bool rayIntersect(const Ray3f &ray, Intersection &its) const;
    bool rayIntersect(const Ray3f &ray) const;
    void addChild(NoriObject *obj) override;
    void activate() override;
    std::string toString() const override;
    EClassType getClassType() const override { return EScene; }
    /// Device context holding this scene (created and filled on first use). Throws NoriException
    /// when no CUDA device or libnori_b200.so is available: there is no CPU path.
    nori_gpu_ctx *gpu(int device = 0) const;
private:
    std::vector<Mesh *> m_meshes;
    Integrator *m_integrator = nullptr;
    Sampler *m_sampler = nullptr;
    Camera *m_camera = nullptr;
    mutable std::map<int, nori_gpu_ctx *> m_ctx;        // device -> context holding this scene
};

/// `--device-dedup`: the OBJ loader hands corner de-duplication and the vertex gather (src/obj.cpp:96-161) to the device
/// (nori_gpu_obj_dedup) instead of its host hash map; same m_F / m_V / m_N / m_UV. An explicit choice of the caller, never automatic.
bool deviceObjDedup();
void setDeviceObjDedup(bool on);

/// Device context for operators that need no scene (BSDF::sample/eval/pdf, Warp::*): created on first use.
nori_gpu_ctx *operatorContext();
/// Turns a non-zero return of the C ABI into a NoriException carrying nori_gpu_last_error().
void gpuCheck(nori_gpu_ctx *ctx, int rc, const char *what);

/// Statistics nori_gpu_render reports (subset printed by the command line tool)
struct RenderReport { double ms = 0, msamples = 0, mrays = 0, gsegments = 0; uint64_t dropped = 0, launches = 0; };

/// Replacement for the reference's render() block loop (src/main.cpp:86-127): the whole frame,
/// samples [sampleFirst, sampleLast) of every pixel, into `result`.
/// `resumeFrom`: an ImageBlock holding the samples rendered so far (a checkpoint); the new samples are added to it
/// on the device (nori_gpu_upload_block + NoriGpuRenderDesc::clear = 0).
void renderGpu(const Scene *scene, ImageBlock &result, uint32_t sampleFirst, uint32_t sampleLast, int device, RenderReport *report,
               const ImageBlock *resumeFrom = nullptr);
/// The same over several GPUs of the box: device k renders the k-th contiguous part of [sampleFirst, sampleLast)
/// (per-sample streams: the union of the parts is the single-GPU sample set) on its own host thread, then ONE
/// nori_gpu_reduce_blocks (ncclReduce over NVLink) sums the accumulators into the first device's, which is downloaded.
void renderGpuMulti(const Scene *scene, ImageBlock &result, uint32_t sampleFirst, uint32_t sampleLast, const std::vector<int> &devices,
                    RenderReport *report);
/// Checkpoint file of an ImageBlock (RGBW accumulator + the sample range it holds): raw little-endian.
void saveCheckpoint(const std::string &path, const ImageBlock &block, uint32_t sampleFirst, uint32_t sampleLast);
void loadCheckpoint(const std::string &path, ImageBlock &block, uint32_t &sampleFirst, uint32_t &sampleLast);

/// Integrator::Li for n individual camera samples (pixel index y * width + x, sample index) with the
/// per-sample streams of the render path: rgb n x 3. Used by the scene t-tests (src/ttest.cpp:136-176).
void sampleRadianceGpu(const Scene *scene, int n, const uint32_t *pixel, const uint32_t *sample, float *rgb, int device = 0);

/// The fork's path-graph tool (`pg`, src/cluster.cpp:156-253, mode "knn") on the tracer's own paths instead of another renderer's
/// dumps, in ONE device call (nori_gpu_pathgraph_render): path vertices of the camera samples [sampleFirst, sampleLast) of every
/// pixel, grid as PathGraph::computeDimensions sizes it, kNN + `iterations` passes (mis = true: the MIS gather the tool's "knn" mode runs, computeMISAllOnGPURecord; false: the
/// scatter driver computeScatterAllOnGPURecord), then the images writeFullLight / writeIndirectLight / writeFullinit compose from each path's
/// FIRST vertex (src/cluster.cpp:23-107): full = eLd + mc, indirect = mc, pt = eLi - eLd, blur, init = eLd, all of the last
/// pass. With several samples per pixel the reference's loop keeps the last path of a pixel; here they are averaged.
/// mode NORI_PG_CLUSTER (2) = the tool's DEFAULT "opt" mode (src/cluster.cpp:204-218): SPoints::BuildClusters with k members per cluster +
/// ClusterScatterWithDirectOptNR, with `direct` = the filtered direct light writeDirectLight writes (:102-126) beside the five.
struct PathGraphImages {
    std::unique_ptr<Bitmap> full, indirect, pt, blur, init, direct;
    uint64_t vertices = 0; double msVertices = 0, msKnn = 0, msScatter = 0;
};
PathGraphImages pathGraphGpu(const Scene *scene, uint32_t sampleFirst, uint32_t sampleLast, int k, int iterations, int mode = 1 /* NORI_PG_MIS; 0 scatter, 2 cluster */, int device = 0);

/// Search path for files named in scene descriptions (the reference prepends the scene's directory)
class FileResolver {
public:
    void prepend(const std::string &dir) { m_paths.insert(m_paths.begin(), dir); }
    std::string resolve(const std::string &name) const;
private:
    std::vector<std::string> m_paths;
};
FileResolver *getFileResolver();

NORI_NAMESPACE_END
