// glue_render.cpp — headless restatement of the reference's render driver (TEST INFRASTRUCTURE).
// src/main.cpp cannot be compiled here (nanogui/OpenGL); this follows its loop —
// BlockGenerator -> per-worker ImageBlock + cloned sampler -> prepare(block) -> per pixel, per
// sample: 2x next2D, Camera::sampleRay, Integrator::Li, ImageBlock::put -> result.put(block)
// (src/main.cpp:38-67, 93-121) — over the reference's own compiled classes and TBB.
#include "plugins.h"
#include <nori/camera.h>
#include <nori/block.h>
#include <nori/timer.h>
#include <pcg32.h>
#include <tbb/parallel_for.h>
#include <tbb/blocked_range.h>
#include <tbb/task_scheduler_init.h>
#include <cstring>
#include <chrono>
#include <thread>
#include <mutex>
#include <set>

using namespace nori;

namespace {
/// `independent` with an explicit sample count and, optionally, one PCG32 stream per camera
/// sample: seed(initstate = y*W + x, initseq = sample index) (SURVEY §8d PER_SAMPLE_STREAM).
class StreamSampler : public Sampler {
public:
    explicit StreamSampler(size_t spp) { m_sampleCount = spp; }
    std::unique_ptr<Sampler> clone() const { return std::unique_ptr<Sampler>(new StreamSampler(*this)); }
    void prepare(const ImageBlock &block) { m_random.seed(block.getOffset().x(), block.getOffset().y()); }   // independent.cpp:48-53
    void seedSample(uint64_t pixel, uint64_t sample) { m_random.seed(pixel, sample); }
    void generate() {}
    void advance() {}
    float next1D() { return m_random.nextFloat(); }
    /// same draw order as the compiled `independent` plugin: first draw -> y, second -> x
    Point2f next2D() { float y = m_random.nextFloat(); float x = m_random.nextFloat(); return Point2f(x, y); }
    std::string toString() const { return "StreamSampler[]"; }
private:
    pcg32 m_random;
};
} // namespace

namespace {
std::mutex g_workerLock;
std::set<std::thread::id> g_workers;        // TBB workers that rendered at least one block in the last ref_render
} // namespace

extern "C" {

/// mode 0: the reference's per-block stream; mode 1: per-sample streams, samples [s0, s1).
/// rgbw = ImageBlock::data() of the merged result, (H+2b) x (W+2b) x 4. Returns render seconds.
/// With threads == 1 blocks are merged in generator order (deterministic); otherwise in
/// completion order under the ImageBlock mutex exactly as the reference does.
double ref_render(void *o, int mode, uint32_t s0, uint32_t s1, int threads, int useBvh, float *rgbw) {
    const Scene *scene = static_cast<const Scene *>((NoriObject *) o);
    const Camera *camera = scene->getCamera();
    const Integrator *integrator = scene->getIntegrator();
    accelSetMode(useBvh);
    const_cast<Integrator *>(integrator)->preprocess(scene);
    Vector2i outputSize = camera->getOutputSize();
    BlockGenerator blockGenerator(outputSize, NORI_BLOCK_SIZE);
    ImageBlock result(outputSize, camera->getReconstructionFilter());
    result.clear();
    const int W = outputSize.x();

    { std::lock_guard<std::mutex> g(g_workerLock); g_workers.clear(); }
    tbb::task_scheduler_init init(threads > 0 ? threads : tbb::task_scheduler_init::automatic);
    auto t0 = std::chrono::steady_clock::now();
    tbb::blocked_range<int> range(0, blockGenerator.getBlockCount());
    auto map = [&](const tbb::blocked_range<int> &r) {
        ImageBlock block(Vector2i(NORI_BLOCK_SIZE), camera->getReconstructionFilter());
        // mode 0 draws through the reference's own `independent` plugin (src/independent.cpp)
        StreamSampler streams(s1 - s0);
        std::unique_ptr<Sampler> indep(static_cast<Sampler *>(NoriObjectFactory::createInstance("independent", PropertyList())));
        Sampler &sampler = mode == 1 ? static_cast<Sampler &>(streams) : *indep;
        { std::lock_guard<std::mutex> g(g_workerLock); g_workers.insert(std::this_thread::get_id()); }
        for (int i = r.begin(); i < r.end(); ++i) {
            blockGenerator.next(block);
            sampler.prepare(block);
            Point2i offset = block.getOffset(); Vector2i size = block.getSize();
            block.clear();
            for (int y = 0; y < size.y(); ++y) for (int x = 0; x < size.x(); ++x)
                for (uint32_t k = s0; k < s1; ++k) {
                    int gx = x + offset.x(), gy = y + offset.y();
                    if (mode == 1) streams.seedSample((uint64_t) gy * (uint64_t) W + (uint64_t) gx, k);
                    Point2f pixelSample = Point2f((float) gx, (float) gy) + sampler.next2D();
                    Point2f apertureSample = sampler.next2D();
                    Ray3f ray;
                    Color3f value = camera->sampleRay(ray, pixelSample, apertureSample);
                    value *= integrator->Li(scene, &sampler, ray);
                    block.put(pixelSample, value);
                }
            result.put(block);
        }
    };
    if (threads == 1) map(range); else tbb::parallel_for(range, map);
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rgbw) std::memcpy(rgbw, result.data(), (size_t) result.rows() * result.cols() * 16);
    return sec;
}

/// Radiance of individual camera samples under per-sample streams: rgb n x 3, pos n x 2.
void ref_sample_radiance(void *o, int n, const uint32_t *pixel, const uint32_t *sample, int useBvh, float *rgb, float *pos) {
    const Scene *scene = static_cast<const Scene *>((NoriObject *) o);
    const Camera *camera = scene->getCamera();
    accelSetMode(useBvh);
    StreamSampler sampler(1);
    const uint32_t W = (uint32_t) camera->getOutputSize().x();
    for (int i = 0; i < n; ++i) {
        sampler.seedSample(pixel[i], sample[i]);
        Point2f ps = Point2f((float) (pixel[i] % W), (float) (pixel[i] / W)) + sampler.next2D();
        Point2f ap = sampler.next2D();
        Ray3f ray;
        Color3f v = camera->sampleRay(ray, ps, ap);
        v *= scene->getIntegrator()->Li(scene, &sampler, ray);
        for (int k = 0; k < 3; ++k) rgb[3 * i + k] = v[k];
        if (pos) { pos[2 * i] = ps.x(); pos[2 * i + 1] = ps.y(); }
    }
}

int ref_hardware_threads() { return (int) std::thread::hardware_concurrency(); }
/// Distinct threads that took part in the last ref_render (what `cores` of a CPU baseline line should say).
int ref_render_threads_used() { std::lock_guard<std::mutex> g(g_workerLock); return (int) g_workers.size(); }

} // extern "C"
