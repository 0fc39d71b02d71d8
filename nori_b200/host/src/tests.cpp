// tests.cpp — the reference's statistical <test> plugins on the device operators:
//
//   chi2test  (src/chi2test.cpp:42-228)   does BSDF::sample() draw from the density BSDF::pdf() claims?
//   ttest     (src/ttest.cpp:44-220)      does a BSDF's / a scene's estimator converge to a reference value?
//   warptest  (src/warptest.cpp:96-262, the CLI mode :979-1006)   the same chi^2 test for Warp::squareTo*
//
// Same XML names, properties, defaults, decision rules and console output as the reference, so its test
// scenes (scenes/pa4/tests, scenes/pa5/tests) run unchanged: `nori_b200 chi2test-microfacet.xml`.
// What changes is where the numbers come from. The reference draws and evaluates one sample at a time
// on the CPU; here every batch is ONE call of the C ABI (nori_gpu_bsdf_sample, nori_gpu_bsdf_eval_pdf,
// nori_gpu_warp, nori_gpu_warp_pdf, nori_gpu_sample_radiance), i.e. the device code the renderer uses:
//   * samples: the host PCG32 stream of the reference (default-seeded pcg32, two floats per sample) is
//     generated up front and handed over as one array, so a BSDF test sees the reference's inputs;
//   * expected frequencies: the reference integrates the density over each cell with a nested adaptive
//     Simpson rule, one pdf() call at a time; here every cell gets a composite Simpson grid (33 x 33 nodes for
//     the 10 x 20 BSDF tables, 17 x 17 for the 51 x 102 warp tables), the nodes of ALL cells are evaluated
//     in one batch, and the cells where that rule disagrees with the next coarser one are integrated again
//     on a 257 x 257 grid (adaptivity at cell granularity, integrateCells below);
//   * scene t-tests: the reference threads one serial sampler through all paths; paths here are
//     camera samples k = 0 .. n-1 of a uniformly drawn pixel with the per-sample streams of the render path.
#include <nori/scene.h>
#include <nori/hypothesis.h>
#include <nori_gpu.h>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>

NORI_NAMESPACE_BEGIN

namespace {

const double kPi = 3.14159265358979323846;

/// chi2_dump (hypothesis.h:246-283): both tables as an Octave / MATLAB script.
void dumpTables(int rows, int cols, const std::vector<double> &obs, const std::vector<double> &expd, const std::string &filename) {
    std::ofstream f(filename);
    auto table = [&](const char *name, const std::vector<double> &t) {
        f << name << " = [ ";
        for (int i = 0; i < rows; ++i) {
            for (int j = 0; j < cols; ++j) f << t[(size_t) i * cols + j] << (j + 1 < cols ? ", " : "");
            if (i + 1 < rows) f << "; ";
        }
        f << " ];" << std::endl;
    };
    table("obsFrequencies", obs); table("expFrequencies", expd);
    f << "colormap(jet);\nclf; subplot(2,1,1);\nimagesc(obsFrequencies);\ntitle('Observed frequencies');\naxis equal;\n"
         "subplot(2,1,2);\nimagesc(expFrequencies);\naxis equal;\ntitle('Expected frequencies');" << std::endl;
}

/// Integrates a density over every cell of a rows x cols table on the device.
/// node(u, v, out3): maps table coordinates u in [0, 1) (columns), v in [0, 1) (rows) to the point the
/// density is evaluated at; eval(points, n, pdf): the batched density; returns integral * scale per cell.
/// Pass 1 puts a composite Simpson grid of q x q intervals on every cell (one batch) and compares it with
/// the q/2 x q/2 rule on the same nodes; cells where the two disagree by more than 1e-4 (a density edge
/// runs through them: the horizon of a hemisphere, the rim of the disk) are integrated again with
/// 256 x 256 intervals — the role the recursion plays in the reference's adaptive rule (hypothesis.h:65-109).
template <class Node, class Eval>
std::vector<double> integrateCells(int rows, int cols, int q, double scale, const Node &node, const Eval &eval) {
    std::vector<double> out((size_t) rows * cols), coarse(out.size());
    // integrates the listed cells with q x q intervals; `half` (optional) receives the q/2 x q/2 estimate
    auto run = [&](const std::vector<int> &cells, int q, std::vector<double> &full, std::vector<double> *half) {
        const int qn = q + 1;
        const size_t perCell = (size_t) qn * qn, n = perCell * cells.size();
        std::vector<float> pts(3 * n), pdf(n);
        size_t k = 0;
        for (int cell : cells) {
            const int r = cell / cols, c = cell % cols;
            for (int j = 0; j < qn; ++j)
                for (int i = 0; i < qn; ++i, ++k)
                    node((c + i / (double) q) / cols, (r + j / (double) q) / rows, &pts[3 * k]);
        }
        eval(pts.data(), (int) n, pdf.data());
        const std::vector<double> wu = hypothesis::simpsonWeights(q, 0.0, 1.0 / cols), wv = hypothesis::simpsonWeights(q, 0.0, 1.0 / rows);
        const std::vector<double> hu = hypothesis::simpsonWeights(q / 2, 0.0, 1.0 / cols), hv = hypothesis::simpsonWeights(q / 2, 0.0, 1.0 / rows);
        k = 0;
        for (int cell : cells) {
            double acc = 0, accHalf = 0;
            for (int j = 0; j < qn; ++j) {
                double row = 0, rowHalf = 0;
                for (int i = 0; i < qn; ++i, ++k) {
                    row += wu[(size_t) i] * (double) pdf[k];
                    if (!(i & 1)) rowHalf += hu[(size_t) (i / 2)] * (double) pdf[k];
                }
                acc += wv[(size_t) j] * row;
                if (!(j & 1)) accHalf += hv[(size_t) (j / 2)] * rowHalf;
            }
            full[(size_t) cell] = acc * scale;
            if (half) (*half)[(size_t) cell] = accHalf * scale;
        }
    };
    std::vector<int> all(out.size());
    for (size_t i = 0; i < all.size(); ++i) all[i] = (int) i;
    run(all, q, out, &coarse);
    std::vector<int> refine;
    for (size_t i = 0; i < out.size(); ++i)
        if (std::fabs(out[i] - coarse[i]) > 1e-4 * std::max(std::fabs(out[i]), std::fabs(coarse[i]))) refine.push_back((int) i);
    const int fine = 256;
    const size_t perBatch = std::max<size_t>(1, (size_t) 1 << 23) / ((size_t) (fine + 1) * (fine + 1));   // about 8 M nodes per call
    for (size_t at = 0; at < refine.size(); at += perBatch) {
        const std::vector<int> part(refine.begin() + (long) at, refine.begin() + (long) std::min(refine.size(), at + perBatch));
        run(part, fine, out, nullptr);
    }
    return out;
}

/// (cos theta in [-1, 1], phi in [0, 2 pi)) -> unit vector, computed in double as the reference's integrands do.
void sphereNode(double cosTheta, double phi, float *out3) {
    const double sinTheta = std::sqrt(std::max(0.0, 1.0 - cosTheta * cosTheta));
    out3[0] = (float) (sinTheta * std::cos(phi)); out3[1] = (float) (sinTheta * std::sin(phi)); out3[2] = (float) cosTheta;
}

struct BsdfBatch {
    int type; float params[6];
    explicit BsdfBatch(const BSDF *b) { b->describe(type, params); }
    /// n x BSDF::sample for one wi; xi: n x 2. Outputs n x 3 each (+ n).
    void sample(const Vector3f &wi, const std::vector<float> &xi, std::vector<float> &wo, std::vector<float> &weight) const {
        const int n = (int) (xi.size() / 2);
        std::vector<float> wis(3 * (size_t) n), pdf((size_t) n), eta((size_t) n); std::vector<int32_t> measure((size_t) n);
        for (int i = 0; i < n; ++i) { wis[3 * (size_t) i] = wi.x; wis[3 * (size_t) i + 1] = wi.y; wis[3 * (size_t) i + 2] = wi.z; }
        wo.resize(3 * (size_t) n); weight.resize(3 * (size_t) n);
        nori_gpu_ctx *ctx = operatorContext();
        gpuCheck(ctx, nori_gpu_bsdf_sample(ctx, type, params, n, wis.data(), xi.data(), wo.data(), weight.data(), pdf.data(), eta.data(), measure.data()),
                 "nori_gpu_bsdf_sample");
    }
    /// n x BSDF::pdf (solid angle measure) for one wi
    void pdf(const Vector3f &wi, const float *wo, int n, float *out) const {
        std::vector<float> wis(3 * (size_t) n), f(3 * (size_t) n);
        for (int i = 0; i < n; ++i) { wis[3 * (size_t) i] = wi.x; wis[3 * (size_t) i + 1] = wi.y; wis[3 * (size_t) i + 2] = wi.z; }
        nori_gpu_ctx *ctx = operatorContext();
        gpuCheck(ctx, nori_gpu_bsdf_eval_pdf(ctx, type, params, n, wis.data(), wo, f.data(), out), "nori_gpu_bsdf_eval_pdf");
    }
};

/// Knuth's running mean / variance (TAOCP vol. 2, 3rd ed., p. 232), in the sample order given.
void meanVariance(const std::vector<double> &x, double &mean, double &variance) {
    mean = 0; variance = 0;
    for (size_t k = 0; k < x.size(); ++k) {
        const double delta = x[k] - mean;
        mean += delta / (double) (k + 1);
        variance += delta * (x[k] - mean);
    }
    variance /= (double) x.size() - 1.0;
}

} // namespace

// ------------------------------------------------------------------------------------------------
class ChiSquareTest : public NoriObject {
public:
    explicit ChiSquareTest(const PropertyList &p) {
        m_significanceLevel = p.getFloat("significanceLevel", 0.01f);
        m_cosThetaResolution = p.getInteger("resolution", 10);
        m_minExpFrequency = p.getInteger("minExpFrequency", 5);
        m_sampleCount = p.getInteger("sampleCount", -1);
        m_testCount = p.getInteger("testCount", 5);
        m_phiResolution = 2 * m_cosThetaResolution;
        if (m_sampleCount < 0) m_sampleCount = m_cosThetaResolution * m_phiResolution * 5000;   // ~5K samples per bin
    }
    ~ChiSquareTest() override { for (BSDF *b : m_bsdfs) delete b; }

    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != EBSDF)
            throw NoriException("ChiSquareTest::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        m_bsdfs.push_back(static_cast<BSDF *>(obj));
    }

    void activate() override {
        int passed = 0, total = 0;
        const int rows = m_cosThetaResolution, cols = m_phiResolution;
        Pcg32 random;
        for (const BSDF *bsdf : m_bsdfs) {
            const BsdfBatch batch(bsdf);
            for (int l = 0; l < m_testCount; ++l) {
                std::cout << "------------------------------------------------------" << std::endl;
                std::cout << "Testing: " << bsdf->toString() << std::endl;
                ++total;
                const float cosTheta = random.nextFloat();
                const float sinTheta = std::sqrt(std::max(0.0f, 1 - cosTheta * cosTheta));
                float sinPhi, cosPhi;
                sincosf(2.0f * (float) M_PI * random.nextFloat(), &sinPhi, &cosPhi);
                const Vector3f wi(cosPhi * sinTheta, sinPhi * sinTheta, cosTheta);

                std::cout << "Accumulating " << m_sampleCount << " samples into a " << rows << "x" << cols << " contingency table .. " << std::flush;
                std::vector<float> xi(2 * (size_t) m_sampleCount), wo, weight;
                for (float &v : xi) v = random.nextFloat();
                batch.sample(wi, xi, wo, weight);
                std::vector<double> obs((size_t) rows * cols, 0.0);
                for (int i = 0; i < m_sampleCount; ++i) {
                    const float *w = &weight[3 * (size_t) i], *o = &wo[3 * (size_t) i];
                    if (w[0] == 0 && w[1] == 0 && w[2] == 0) continue;
                    const int rbin = std::min(std::max(0, (int) std::floor((o[2] * 0.5f + 0.5f) * rows)), rows - 1);
                    float scaledPhi = std::atan2(o[1], o[0]) * (0.5f / (float) M_PI);
                    if (scaledPhi < 0) scaledPhi += 1;
                    const int cbin = std::min(std::max(0, (int) std::floor(scaledPhi * cols)), cols - 1);
                    obs[(size_t) rbin * cols + cbin] += 1;
                }
                std::cout << "done." << std::endl;

                std::cout << "Integrating expected frequencies .. " << std::flush;
                // cell (r, c) = cos theta in [-1 + 2r/rows, ..) x phi in [2 pi c/cols, ..): area element d(cos theta) d(phi)
                const std::vector<double> expd = integrateCells(rows, cols, 32, 2.0 * 2.0 * kPi * m_sampleCount,
                    [](double u, double v, float *out3) { sphereNode(-1.0 + 2.0 * v, 2.0 * kPi * u, out3); },
                    [&](const float *pts, int n, float *pdf) { batch.pdf(wi, pts, n, pdf); });
                std::cout << "done." << std::endl;

                dumpTables(rows, cols, obs, expd, "chi2test_" + std::to_string(total) + ".m");
                const std::pair<bool, std::string> result = hypothesis::chi2_test(rows * cols, obs.data(), expd.data(), m_sampleCount,
                    m_minExpFrequency, m_significanceLevel, m_testCount * (int) m_bsdfs.size());
                if (result.first) ++passed;
                std::cout << result.second << std::endl;
            }
        }
        std::cout << "Passed " << passed << "/" << total << " tests." << std::endl;
        if (passed < total) throw std::runtime_error("Some tests failed :(");
    }

    std::string toString() const override {
        char buf[512];
        snprintf(buf, sizeof(buf), "ChiSquareTest[\n  thetaResolution = %i,\n  phiResolution = %i,\n  minExpFrequency = %i,\n  sampleCount = %i,\n"
                 "  testCount = %i,\n  significanceLevel = %f\n]", m_cosThetaResolution, m_phiResolution, m_minExpFrequency, m_sampleCount, m_testCount,
                 m_significanceLevel);
        return buf;
    }
    EClassType getClassType() const override { return ETest; }

private:
    int m_cosThetaResolution, m_phiResolution, m_minExpFrequency, m_sampleCount, m_testCount;
    float m_significanceLevel;
    std::vector<BSDF *> m_bsdfs;
};

// ------------------------------------------------------------------------------------------------
class StudentsTTest : public NoriObject {
public:
    explicit StudentsTTest(const PropertyList &p) {
        m_significanceLevel = p.getFloat("significanceLevel", 0.01f);
        for (const std::string &a : tokenize(p.getString("angles", ""))) m_angles.push_back(toFloat(a));
        for (const std::string &r : tokenize(p.getString("references", ""))) m_references.push_back(toFloat(r));
        m_sampleCount = p.getInteger("sampleCount", 100000);
    }
    ~StudentsTTest() override {
        for (BSDF *b : m_bsdfs) delete b;
        for (Scene *s : m_scenes) delete s;
    }

    void addChild(NoriObject *obj) override {
        switch (obj->getClassType()) {
        case EBSDF: m_bsdfs.push_back(static_cast<BSDF *>(obj)); break;
        case EScene: m_scenes.push_back(static_cast<Scene *>(obj)); break;
        default: throw NoriException("StudentsTTest::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        }
    }

    void activate() override {
        int total = 0, passed = 0;
        Pcg32 random;
        auto judge = [&](const std::vector<double> &values, float reference) {
            double mean, variance;
            meanVariance(values, mean, variance);
            const std::pair<bool, std::string> result = hypothesis::students_t_test(mean, variance, reference, m_sampleCount, m_significanceLevel,
                                                                                   (int) m_references.size());
            if (result.first) ++passed;
            std::cout << result.second << std::endl;
        };
        if (!m_bsdfs.empty()) {
            if (m_references.size() * m_bsdfs.size() != m_angles.size())
                throw NoriException("Specified a different number of angles and reference values!");
            if (!m_scenes.empty()) throw NoriException("Cannot test BSDFs and scenes at the same time!");
            size_t ctr = 0;
            for (const BSDF *bsdf : m_bsdfs) {
                const BsdfBatch batch(bsdf);
                for (size_t i = 0; i < m_references.size(); ++i) {
                    const float angle = m_angles[i], reference = m_references[ctr++];
                    std::cout << "------------------------------------------------------" << std::endl;
                    std::cout << "Testing (angle=" << angle << "): " << bsdf->toString() << std::endl;
                    ++total;
                    const Vector3f wi = sphericalDirection(degToRad(angle), 0);
                    std::cout << "Drawing " << m_sampleCount << " samples .. " << std::endl;
                    std::vector<float> xi(2 * (size_t) m_sampleCount), wo, weight;
                    for (float &v : xi) v = random.nextFloat();
                    batch.sample(wi, xi, wo, weight);
                    std::vector<double> values((size_t) m_sampleCount);
                    for (int k = 0; k < m_sampleCount; ++k)
                        values[(size_t) k] = (double) Color3f(weight[3 * (size_t) k], weight[3 * (size_t) k + 1], weight[3 * (size_t) k + 2]).getLuminance();
                    judge(values, reference);
                }
            }
        } else {
            if (m_references.size() != m_scenes.size())
                throw NoriException("Specified a different number of scenes and reference values!");
            size_t ctr = 0;
            for (const Scene *scene : m_scenes) {
                const float reference = m_references[ctr++];
                std::cout << "------------------------------------------------------" << std::endl;
                std::cout << "Testing scene: " << scene->toString() << std::endl;
                ++total;
                std::cout << "Generating " << m_sampleCount << " paths.. " << std::endl;
                const Vector2i size = scene->getCamera()->getOutputSize();
                std::vector<uint32_t> pixel((size_t) m_sampleCount), sample((size_t) m_sampleCount);
                for (int k = 0; k < m_sampleCount; ++k) {
                    const int x = std::min((int) (random.nextFloat() * size.x), size.x - 1), y = std::min((int) (random.nextFloat() * size.y), size.y - 1);
                    pixel[(size_t) k] = (uint32_t) (y * size.x + x); sample[(size_t) k] = (uint32_t) k;
                }
                std::vector<float> rgb(3 * (size_t) m_sampleCount);
                sampleRadianceGpu(scene, m_sampleCount, pixel.data(), sample.data(), rgb.data());
                std::vector<double> values((size_t) m_sampleCount);
                for (int k = 0; k < m_sampleCount; ++k)
                    values[(size_t) k] = (double) Color3f(rgb[3 * (size_t) k], rgb[3 * (size_t) k + 1], rgb[3 * (size_t) k + 2]).getLuminance();
                judge(values, reference);
            }
        }
        std::cout << "Passed " << passed << "/" << total << " tests." << std::endl;
        if (passed < total) throw std::runtime_error("Some tests failed :(");
    }

    std::string toString() const override {
        char buf[256];
        snprintf(buf, sizeof(buf), "StudentsTTest[\n  significanceLevel = %f,\n  sampleCount= %i\n]", m_significanceLevel, m_sampleCount);
        return buf;
    }
    EClassType getClassType() const override { return ETest; }

private:
    std::vector<BSDF *> m_bsdfs;
    std::vector<Scene *> m_scenes;
    std::vector<float> m_angles, m_references;
    float m_significanceLevel;
    int m_sampleCount;
};

// ------------------------------------------------------------------------------------------------
/// The chi^2 test of src/warptest.cpp's CLI mode as a <test type="warptest"> (and `nori_b200 --warptest`):
/// string `warp` (one of the names below), float `param` (Beckmann alpha / microfacet alpha), float
/// `param2` (microfacet kd), float `angle` (microfacet incident angle, degrees).
class WarpTest : public NoriObject {
public:
    explicit WarpTest(const PropertyList &p) {
        static const char *names[] = {"square", "tent", "disk", "uniform_sphere", "uniform_hemisphere", "cosine_hemisphere", "beckmann", "microfacet_brdf"};
        const std::string warp = p.getString("warp");
        m_type = -1;
        for (int i = 0; i < 8; ++i) if (warp == names[i]) m_type = i;
        if (m_type < 0) throw NoriException("Invalid warp type!");
        m_name = warp;
        m_param = p.getFloat("param", 0.0f); m_param2 = p.getFloat("param2", 0.0f); m_angle = p.getFloat("angle", 0.0f);
        m_xres = m_yres = 51;                                           // warptest.cpp:97-98
        if (m_type > 2) m_xres *= 2;
    }

    std::pair<bool, std::string> run() const {
        const int res = m_xres * m_yres, sampleCount = 1000 * res;
        Pcg32 rng;
        std::vector<float> xi(2 * (size_t) sampleCount);
        for (float &v : xi) v = rng.nextFloat();
        std::vector<float> pts(3 * (size_t) sampleCount), value((size_t) sampleCount, 1.0f);
        nori_gpu_ctx *ctx = operatorContext();
        std::unique_ptr<BSDF> bsdf;
        Vector3f wi;
        if (m_type == 7) {                                              // warptest.cpp:301-313
            PropertyList list;
            list.setFloat("alpha", m_param); list.setColor("kd", Color3f(m_param2));
            bsdf.reset(static_cast<BSDF *>(NoriObjectFactory::createInstance("microfacet", list)));
            const float a = degToRad(m_angle);
            wi = Vector3f(std::sin(a), 0.f, std::max(std::cos(a), 1e-4f)).normalized();
            std::vector<float> weight;
            BsdfBatch(bsdf.get()).sample(wi, xi, pts, weight);
            for (int i = 0; i < sampleCount; ++i)
                value[(size_t) i] = Color3f(weight[3 * (size_t) i], weight[3 * (size_t) i + 1], weight[3 * (size_t) i + 2]).getLuminance();
        } else {
            std::vector<float> pdf((size_t) sampleCount);
            gpuCheck(ctx, nori_gpu_warp(ctx, m_type, m_param, sampleCount, xi.data(), pts.data(), pdf.data()), "nori_gpu_warp");
        }
        std::vector<double> obs((size_t) res, 0.0);
        for (int i = 0; i < sampleCount; ++i) {
            if (value[(size_t) i] == 0) continue;
            const float *s = &pts[3 * (size_t) i];
            float x, y;
            if (m_type == 0) { x = s[0]; y = s[1]; }
            else if (m_type <= 2) { x = s[0] * 0.5f + 0.5f; y = s[1] * 0.5f + 0.5f; }
            else { x = std::atan2(s[1], s[0]) * (0.5f / (float) M_PI); if (x < 0) x += 1; y = s[2] * 0.5f + 0.5f; }
            const int xbin = std::min(m_xres - 1, std::max(0, (int) std::floor(x * m_xres))), ybin = std::min(m_yres - 1, std::max(0, (int) std::floor(y * m_yres)));
            obs[(size_t) ybin * m_xres + xbin] += 1;
        }
        const double scale = (double) sampleCount * (m_type == 0 ? 1.0 : (m_type <= 2 ? 4.0 : 4.0 * kPi));
        const int type = m_type; const float param = m_param;
        const BSDF *b = bsdf.get();
        const std::vector<double> expd = integrateCells(m_yres, m_xres, 16, scale,     // cells are 1/51 wide: 17 x 17 nodes each

            [type](double u, double v, float *out3) {
                if (type == 0) { out3[0] = (float) u; out3[1] = (float) v; out3[2] = 0; }
                else if (type <= 2) { out3[0] = (float) (u * 2 - 1); out3[1] = (float) (v * 2 - 1); out3[2] = 0; }
                else sphereNode(v * 2 - 1, u * 2 * kPi, out3);
            },
            [&](const float *p3, int n, float *pdf) {
                if (type == 7) BsdfBatch(b).pdf(wi, p3, n, pdf);
                else gpuCheck(ctx, nori_gpu_warp_pdf(ctx, type, param, n, p3, pdf), "nori_gpu_warp_pdf");
            });
        for (double e : expd) if (e < 0) throw NoriException("The Pdf() function returned negative values!");
        dumpTables(m_yres, m_xres, obs, expd, "chitest.m");
        return hypothesis::chi2_test(res, obs.data(), expd.data(), sampleCount, 5, 0.01f, 1);
    }

    void activate() override {
        char buf[256];
        snprintf(buf, sizeof(buf), "Testing warp %s, parameter value = %f", m_name.c_str(), m_param);
        std::cout << buf;
        if (m_param2 > 0) { snprintf(buf, sizeof(buf), ", second parameter value = %f", m_param2); std::cout << buf; }
        std::cout << std::endl;
        const std::pair<bool, std::string> res = run();
        std::cout << res.second << std::endl;
        if (!res.first) throw std::runtime_error("warptest failed: " + res.second);
    }
    std::string toString() const override { return "WarpTest[" + m_name + "]"; }
    EClassType getClassType() const override { return ETest; }

private:
    int m_type, m_xres, m_yres; float m_param, m_param2, m_angle; std::string m_name;
};

NORI_REGISTER_CLASS(ChiSquareTest, "chi2test")
NORI_REGISTER_CLASS(StudentsTTest, "ttest")
NORI_REGISTER_CLASS(WarpTest, "warptest")

NORI_NAMESPACE_END
