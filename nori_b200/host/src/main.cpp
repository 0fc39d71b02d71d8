// main.cpp — `nori_b200 scene.xml`: the headless counterpart of the reference's `nori scene.xml`
// (src/main.cpp:153-213): parse the scene description, render it on the GPU through the C ABI, write
// <scene>.exr and <scene>.png next to the scene file.
//
//   nori_b200 [--device N] [--spp N] [--first A --last B] [--out NAME] [--no-png] scene.xml
//   nori_b200 --devices 0,1,2,3 scene.xml                              sample ranges over several GPUs, one NCCL reduce
//   nori_b200 --first 0 --last 64 --checkpoint part.blk scene.xml      render a sample range, keep the accumulator
//   nori_b200 --resume part.blk --last 256 scene.xml                   continue from it (samples [64, 256) are added)
//   nori_b200 --device-dedup ...        OBJ corner de-duplication + vertex gather on the GPU (nori_gpu_obj_dedup) instead of the loader's hash map
//   nori_b200 --dump FILE scene.xml     parse only; write the flattened descriptors (tests; no GPU)
//   nori_b200 test.xml                  a <test> root (chi2test, ttest, warptest) runs in activate(), as in the reference
//   nori_b200 --li-stream FILE [--spp N] scene.xml   the reference's renderBlock loop (src/main.cpp:38-67) written against the
//                                       plugin contracts: Sampler::prepare / next2D, Camera::sampleRay, Integrator::Li (one ray per
//                                       call through nori_gpu_li); writes (pixelSample.xy, rgb) float32 x 5 per camera sample
//   nori_b200 --warptest NAME [P [P2]]  the reference's `warptest NAME P P2` command line mode (src/warptest.cpp:979-1006)
#include <nori/scene.h>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>

using namespace nori;

namespace {

template <class T> void put(std::ofstream &os, const T &v) { os.write((const char *) &v, sizeof(T)); }
template <class T> void putArray(std::ofstream &os, const T *p, size_t n) { os.write((const char *) p, (std::streamsize) (n * sizeof(T))); }

/// Everything that crosses the C ABI for this scene, as one little-endian binary file
/// (layout read by tests/test_host_cpp.py).
void dumpScene(const Scene *scene, const std::string &path) {
    std::ofstream os(path, std::ios::binary);
    if (os.fail()) throw NoriException("Unable to open \"%s\" for writing", path);
    os.write("NB2DUMP1", 8);
    const Camera *cam = scene->getCamera();
    float s2c[16], c2w[16], nearClip, farClip, fov;
    cam->describe(s2c, c2w, nearClip, farClip, fov);
    put<int32_t>(os, (int32_t) scene->getMeshes().size());
    put<int32_t>(os, cam->getOutputSize().x); put<int32_t>(os, cam->getOutputSize().y);
    put(os, nearClip); put(os, farClip); put(os, fov);
    putArray(os, s2c, 16); putArray(os, c2w, 16);
    const ImageBlock block(cam->getOutputSize(), cam->getReconstructionFilter());
    put(os, block.filterRadius()); putArray(os, block.filterTable(), NORI_FILTER_RESOLUTION + 1);
    const Integrator *in = scene->getIntegrator();
    float lp[3], le[3];
    in->describe(lp, le);
    put<int32_t>(os, in->gpuType()); put<int32_t>(os, in->rrStart()); put<int32_t>(os, in->maxDepth());
    put<int32_t>(os, (int32_t) scene->getSampler()->getSampleCount());
    put(os, in->rrMax()); putArray(os, lp, 3); putArray(os, le, 3);
    for (const Mesh *m : scene->getMeshes()) {
        put<uint32_t>(os, m->getVertexCount()); put<uint32_t>(os, m->getTriangleCount());
        put<uint32_t>(os, m->getVertexNormals().empty() ? 0u : 1u); put<uint32_t>(os, m->getVertexTexCoords().empty() ? 0u : 1u);
        int type; float params[6];
        m->getBSDF()->describe(type, params);
        put<int32_t>(os, type); putArray(os, params, 6);
        put<int32_t>(os, m->isEmitter() ? 1 : 0);
        const Color3f r = m->isEmitter() ? m->getEmitter()->getRadiance() : Color3f(0.0f);
        put(os, r.r); put(os, r.g); put(os, r.b);
        putArray(os, m->getVertexPositions().data(), m->getVertexPositions().size());
        putArray(os, m->getVertexNormals().data(), m->getVertexNormals().size());
        putArray(os, m->getVertexTexCoords().data(), m->getVertexTexCoords().size());
        putArray(os, m->getIndices().data(), m->getIndices().size());
    }
}

void render(const Scene *scene, const std::string &outName, int device, int sppOverride, long first, long last, bool png,
            const std::string &resumeFile, const std::string &checkpointFile, const std::vector<int> &devices) {
    const Camera *camera = scene->getCamera();
    const uint32_t spp = sppOverride > 0 ? (uint32_t) sppOverride : (uint32_t) scene->getSampler()->getSampleCount();
    uint32_t s0 = first >= 0 ? (uint32_t) first : 0u, s1 = last >= 0 ? (uint32_t) last : spp, held0 = 0;
    ImageBlock result(camera->getOutputSize(), camera->getReconstructionFilter());
    std::unique_ptr<ImageBlock> held;
    if (!resumeFile.empty()) {                                // continue after the samples the checkpoint holds
        held.reset(new ImageBlock(camera->getOutputSize(), camera->getReconstructionFilter()));
        uint32_t heldLast = 0;
        loadCheckpoint(resumeFile, *held, held0, heldLast);
        if (first >= 0 && (uint32_t) first != heldLast) throw NoriException("the checkpoint ends at sample %d; cannot resume at %d", (int) heldLast, (int) first);
        s0 = heldLast;
    }
    if (s1 <= s0) throw NoriException("empty sample range [%d, %d)", (int) s0, (int) s1);
    RenderReport rep;
    if (devices.size() > 1) {
        if (held) throw NoriException("--resume and --devices cannot be combined");
        std::cout << "Rendering on " << devices.size() << " GPUs .. " << std::flush;
        renderGpuMulti(scene, result, s0, s1, devices, &rep);
    } else {
        scene->gpu(device);                                   // upload + Accel::build (Scene::activate does this in the reference)
        std::cout << "Rendering .. " << std::flush;
        renderGpu(scene, result, s0, s1, device, &rep, held.get());
    }
    if (!checkpointFile.empty()) saveCheckpoint(checkpointFile, result, held ? held0 : s0, s1);
    std::cout << "done. (took " << timeString(rep.ms) << ": " << rep.msamples << " Msamples/s, " << rep.mrays << " Mrays/s, "
              << rep.gsegments << " Gsegments/s";
    if (rep.dropped) std::cout << ", " << rep.dropped << " invalid samples dropped";
    std::cout << ")" << std::endl;
    std::unique_ptr<Bitmap> bitmap = result.toBitmap();
    bitmap->saveEXR(outName);
    if (png) bitmap->savePNG(outName);
}

/// renderBlock (src/main.cpp:38-67) for every 32 x 32 block of the frame, exactly as a reference plugin user writes it: one
/// `independent` sampler re-seeded per block from the block offset, four numbers per camera sample, then whatever Li draws.
/// This is the reference's own sample stream for ANY integrator (the wavefront path offers it for normals / simple only): the
/// serial dependence between consecutive samples is carried by the Sampler, whose pcg32 state travels through nori_gpu_li.
void liStream(const Scene *scene, const std::string &path, int sppOverride) {
    const Camera *camera = scene->getCamera();
    const Integrator *integrator = scene->getIntegrator();
    const Vector2i size = camera->getOutputSize();
    const uint32_t spp = sppOverride > 0 ? (uint32_t) sppOverride : (uint32_t) scene->getSampler()->getSampleCount();
    std::unique_ptr<Sampler> sampler(scene->getSampler()->clone());
    std::ofstream os(path, std::ios::binary);
    if (os.fail()) throw NoriException("Unable to open \"%s\" for writing", path);
    scene->gpu(0);
    for (int oy = 0; oy < size.y; oy += NORI_BLOCK_SIZE)
        for (int ox = 0; ox < size.x; ox += NORI_BLOCK_SIZE) {
            ImageBlock block(Vector2i(std::min(NORI_BLOCK_SIZE, size.x - ox), std::min(NORI_BLOCK_SIZE, size.y - oy)), camera->getReconstructionFilter());
            block.setOffset(Point2i(ox, oy));
            sampler->prepare(block);
            for (int y = 0; y < block.getSize().y; ++y)
                for (int x = 0; x < block.getSize().x; ++x)
                    for (uint32_t i = 0; i < spp; ++i) {
                        const Point2f r = sampler->next2D();
                        const Point2f pixelSample((float) (x + ox) + r.x, (float) (y + oy) + r.y);
                        const Point2f apertureSample = sampler->next2D();
                        Ray3f ray;
                        Color3f value = camera->sampleRay(ray, pixelSample, apertureSample);
                        const Color3f li = integrator->Li(scene, sampler.get(), ray);
                        value = Color3f(value.r * li.r, value.g * li.g, value.b * li.b);
                        const float rec[5] = {pixelSample.x, pixelSample.y, value.r, value.g, value.b};
                        putArray(os, rec, 5);
                    }
        }
}

/// The `pg` tool's "knn" mode (src/cluster.cpp:156-253) on this scene's own paths: <out>_k-K_full / _indirect / _indirect_pt /
/// _indirect_blur and <out>_Le_init, named as the reference names them.
void pathGraph(const Scene *scene, const std::string &outName, int device, int sppOverride, int k, int iterations, int mode) {
    const uint32_t spp = sppOverride > 0 ? (uint32_t) sppOverride : 1u;
    const char *what = mode == 2 ? "clusters" : (mode == 1 ? "MIS gather" : "scatter");
    std::cout << "Path graph: " << spp << " spp, k = " << k << ", " << iterations << " iterations (" << what << ") .. " << std::flush;
    PathGraphImages r = pathGraphGpu(scene, 0, spp, k, iterations, mode, device);
    std::cout << "done. (" << r.vertices << " vertices in " << timeString(r.msVertices) << ", " << (mode == 2 ? "clusters " : "kNN ") << r.msKnn << " ms, scatter " << r.msScatter << " ms)" << std::endl;
    const std::string stem = outName + "_k-" + std::to_string(k);
    r.full->saveEXR(stem + "_full"); r.indirect->saveEXR(stem + "_indirect"); r.pt->saveEXR(stem + "_indirect_pt");
    r.blur->saveEXR(stem + "_indirect_blur"); r.init->saveEXR(outName + "_Le_init");
    if (r.direct) { r.direct->saveEXR(stem + "_direct"); r.init->saveEXR(stem + "_direct_o"); }      // writeDirectLight (src/cluster.cpp:102-126)
}

} // namespace

int main(int argc, char **argv) {
    std::string sceneFile, dumpFile, outName, resumeFile, checkpointFile, liFile;
    int pgMode = 1;                                      // 1: the tool's "knn" mode (MIS gather), 0: scatter driver, 2: its default "opt" mode (clusters)
    int device = 0, spp = 0, pgK = 0, pgIterations = 1; long first = -1, last = -1; bool png = true;
    std::vector<int> devices;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto value = [&](const char *what) -> std::string {
            if (i + 1 >= argc) { std::cerr << "Missing value for " << what << std::endl; exit(-1); }
            return argv[++i];
        };
        if (a == "--device") device = atoi(value("--device").c_str());
        else if (a == "--devices") { for (const std::string &t : tokenize(value("--devices"), ",")) devices.push_back(toInt(t)); if (!devices.empty()) device = devices[0]; }
        else if (a == "--spp") spp = atoi(value("--spp").c_str());
        else if (a == "--first") first = atol(value("--first").c_str());
        else if (a == "--last") last = atol(value("--last").c_str());
        else if (a == "--out") outName = value("--out");
        else if (a == "--dump") dumpFile = value("--dump");
        else if (a == "--li-stream") liFile = value("--li-stream");
        else if (a == "--resume") resumeFile = value("--resume");
        else if (a == "--checkpoint") checkpointFile = value("--checkpoint");
        else if (a == "--no-png") png = false;
        else if (a == "--device-dedup") setDeviceObjDedup(true);
        else if (a == "--pathgraph-scatter") pgMode = 0;
        else if (a == "--pathgraph-cluster") pgMode = 2;
        else if (a == "--pathgraph") { pgK = atoi(value("--pathgraph").c_str()); if (i + 1 < argc && argv[i + 1][0] != '-' && isdigit((unsigned char) argv[i + 1][0]) && i + 2 < argc) pgIterations = atoi(argv[++i]); }
        else if (a == "--warptest") {
            try {
                PropertyList list;
                list.setString("warp", value("--warptest"));
                if (i + 1 < argc) list.setFloat("param", toFloat(argv[++i]));
                if (i + 1 < argc) list.setFloat("param2", toFloat(argv[++i]));
                std::unique_ptr<NoriObject> test(NoriObjectFactory::createInstance("warptest", list));
                test->activate();
                return 0;
            } catch (const std::exception &e) {
                std::cout << e.what() << std::endl;
                return 1;
            }
        }
        else if (!a.empty() && a[0] == '-') { std::cerr << "Unknown option " << a << std::endl; return -1; }
        else sceneFile = a;
    }
    if (sceneFile.empty()) {
        std::cerr << "Syntax: " << argv[0] << " [--device N | --devices A,B,..] [--spp N] [--first A --last B] [--out NAME] [--no-png] [--resume FILE] [--checkpoint FILE] [--dump FILE] [--pathgraph K [ITERATIONS] [--pathgraph-scatter | --pathgraph-cluster]] <scene.xml>" << std::endl;
        return -1;
    }
    try {
        const size_t slash = sceneFile.find_last_of('/');
        getFileResolver()->prepend(slash == std::string::npos ? std::string(".") : sceneFile.substr(0, slash));
        std::unique_ptr<NoriObject> root(loadFromXML(sceneFile));
        if (root->getClassType() != NoriObject::EScene) return 0;      // e.g. a <test>: nothing to render
        const Scene *scene = static_cast<const Scene *>(root.get());
        if (!dumpFile.empty()) { dumpScene(scene, dumpFile); return 0; }
        if (!liFile.empty()) { liStream(scene, liFile, spp); return 0; }
        if (outName.empty()) {
            const size_t dot = sceneFile.find_last_of('.');
            outName = dot == std::string::npos || (slash != std::string::npos && dot < slash) ? sceneFile : sceneFile.substr(0, dot);
        }
        if (pgK > 0) { pathGraph(scene, outName, device, spp, pgK, pgIterations, pgMode); return 0; }
        render(scene, outName, device, spp, first, last, png, resumeFile, checkpointFile, devices);
    } catch (const std::exception &e) {
        std::cerr << "Fatal error: " << e.what() << std::endl;
        return -1;
    }
    return 0;
}
