// plugins.cpp — the classes scene files name by `type`: reconstruction filters, BSDFs, the area
// emitter, the independent sampler and the integrators. They hold parameters (same property names
// and defaults as the reference) and describe themselves to the device layer; none of them computes
// on the host.
//
// Reference: src/rfilter.cpp:25-120, src/diffuse.cpp:28-33, src/mirror.cpp, src/dielectric.cpp:27-33,
// src/microfacet.cpp:27-48, src/independent.cpp:35-38; the `area` emitter and all integrators are
// absent from the reference: their names and properties are those of its scene files
// (scenes/pa3, pa4, pa5), their semantics those of oracle/noracle_integrators.h.
#include <nori/scene.h>
#include <nori_gpu.h>
#include <algorithm>

NORI_NAMESPACE_BEGIN

namespace {

std::string fmt(const char *f, ...) {
    char buf[512];
    va_list ap; va_start(ap, f); vsnprintf(buf, sizeof(buf), f, ap); va_end(ap);
    return buf;
}

// ---- reconstruction filters ------------------------------------------------------------------------
class GaussianFilter : public ReconstructionFilter {
public:
    explicit GaussianFilter(const PropertyList &p) { m_radius = p.getFloat("radius", 2.0f); m_stddev = p.getFloat("stddev", 0.5f); }
    float eval(float x) const override {
        const float alpha = -1.0f / (2.0f * m_stddev * m_stddev);
        return std::max(0.0f, std::exp(alpha * x * x) - std::exp(alpha * m_radius * m_radius));
    }
    std::string toString() const override { return fmt("GaussianFilter[radius=%f, stddev=%f]", m_radius, m_stddev); }
private:
    float m_stddev;
};

class MitchellNetravaliFilter : public ReconstructionFilter {
public:
    explicit MitchellNetravaliFilter(const PropertyList &p) {
        m_radius = p.getFloat("radius", 2.0f); m_B = p.getFloat("B", 1.0f / 3.0f); m_C = p.getFloat("C", 1.0f / 3.0f);
    }
    float eval(float x) const override {
        x = std::abs(2.0f * x / m_radius);
        const float x2 = x * x, x3 = x2 * x;
        if (x < 1) return 1.0f / 6.0f * ((12 - 9 * m_B - 6 * m_C) * x3 + (-18 + 12 * m_B + 6 * m_C) * x2 + (6 - 2 * m_B));
        if (x < 2) return 1.0f / 6.0f * ((-m_B - 6 * m_C) * x3 + (6 * m_B + 30 * m_C) * x2 + (-12 * m_B - 48 * m_C) * x + (8 * m_B + 24 * m_C));
        return 0.0f;
    }
    std::string toString() const override { return fmt("MitchellNetravaliFilter[radius=%f, B=%f, C=%f]", m_radius, m_B, m_C); }
private:
    float m_B, m_C;
};

class TentFilter : public ReconstructionFilter {
public:
    explicit TentFilter(const PropertyList &) { m_radius = 1.0f; }
    float eval(float x) const override { return std::max(0.0f, 1.0f - std::abs(x)); }
    std::string toString() const override { return "TentFilter[]"; }
};

class BoxFilter : public ReconstructionFilter {
public:
    explicit BoxFilter(const PropertyList &) { m_radius = 0.5f; }
    float eval(float) const override { return 1.0f; }
    std::string toString() const override { return "BoxFilter[]"; }
};

// ---- BSDFs ---------------------------------------------------------------------------------------------
class Diffuse : public BSDF {
public:
    explicit Diffuse(const PropertyList &p) { m_albedo = p.getColor("albedo", Color3f(0.5f)); }
    bool isDiffuse() const override { return true; }
    void describe(int &type, float q[6]) const override {
        type = NORI_BSDF_DIFFUSE; q[0] = m_albedo.r; q[1] = m_albedo.g; q[2] = m_albedo.b; q[3] = q[4] = q[5] = 0;
    }
    std::string toString() const override { return fmt("Diffuse[\n  albedo = [%f, %f, %f]\n]", m_albedo.r, m_albedo.g, m_albedo.b); }
private:
    Color3f m_albedo;
};

class Mirror : public BSDF {
public:
    explicit Mirror(const PropertyList &) {}
    void describe(int &type, float q[6]) const override { type = NORI_BSDF_MIRROR; for (int i = 0; i < 6; ++i) q[i] = 0; }
    std::string toString() const override { return "Mirror[]"; }
};

class Dielectric : public BSDF {
public:
    explicit Dielectric(const PropertyList &p) {
        m_intIOR = p.getFloat("intIOR", 1.5046f);     // BK7 glass
        m_extIOR = p.getFloat("extIOR", 1.000277f);   // air
    }
    void describe(int &type, float q[6]) const override {
        type = NORI_BSDF_DIELECTRIC; q[0] = q[1] = q[2] = q[3] = 0; q[4] = m_intIOR; q[5] = m_extIOR;
    }
    std::string toString() const override { return fmt("Dielectric[\n  intIOR = %f,\n  extIOR = %f\n]", m_intIOR, m_extIOR); }
private:
    float m_intIOR, m_extIOR;
};

class Microfacet : public BSDF {
public:
    explicit Microfacet(const PropertyList &p) {
        m_alpha = p.getFloat("alpha", 0.1f);
        m_intIOR = p.getFloat("intIOR", 1.5046f);
        m_extIOR = p.getFloat("extIOR", 1.000277f);
        m_kd = p.getColor("kd", Color3f(0.5f));
        // ks = 1 - max(kd) is derived on the device side from kd (src/microfacet.cpp:47)
    }
    bool isDiffuse() const override { return true; }   // src/microfacet.cpp:84-90
    void describe(int &type, float q[6]) const override {
        type = NORI_BSDF_MICROFACET; q[0] = m_kd.r; q[1] = m_kd.g; q[2] = m_kd.b; q[3] = m_alpha; q[4] = m_intIOR; q[5] = m_extIOR;
    }
    std::string toString() const override {
        return fmt("Microfacet[\n  alpha = %f,\n  intIOR = %f,\n  extIOR = %f,\n  kd = [%f, %f, %f],\n  ks = %f\n]",
                   m_alpha, m_intIOR, m_extIOR, m_kd.r, m_kd.g, m_kd.b, 1.0f - m_kd.maxCoeff());
    }
private:
    float m_alpha, m_intIOR, m_extIOR; Color3f m_kd;
};

// ---- emitter ---------------------------------------------------------------------------------------------
class AreaLight : public Emitter {
public:
    explicit AreaLight(const PropertyList &p) { m_radiance = p.getColor("radiance"); }
    Color3f getRadiance() const override { return m_radiance; }
    std::string toString() const override { return fmt("AreaLight[\n  radiance = [%f, %f, %f]\n]", m_radiance.r, m_radiance.g, m_radiance.b); }
private:
    Color3f m_radiance;
};

// ---- sampler -----------------------------------------------------------------------------------------------
/// src/independent.cpp:33-77: independent samples from one pcg32 per sampler instance.
class Independent : public Sampler {
public:
    explicit Independent(const PropertyList &p) { m_sampleCount = (size_t) p.getInteger("sampleCount", 1); }
    std::unique_ptr<Sampler> clone() const override {
        std::unique_ptr<Independent> cloned(new Independent());
        cloned->m_sampleCount = m_sampleCount;
        cloned->m_random = m_random;
        return std::unique_ptr<Sampler>(cloned.release());
    }
    void prepare(const ImageBlock &block) override { m_random.seed((uint64_t) (int64_t) block.getOffset().x, (uint64_t) (int64_t) block.getOffset().y); }
    void generate() override {}
    void advance() override {}
    float next1D() override { return m_random.nextFloat(); }
    /// The reference writes Point2f(m_random.nextFloat(), m_random.nextFloat()): x86-64 g++ evaluates the arguments right to left,
    /// so y is drawn first (pinned against the compiled reference: tests/golden/ref_vectors.npz, sampler_next2d).
    Point2f next2D() override { const float y = m_random.nextFloat(); const float x = m_random.nextFloat(); return Point2f(x, y); }
    bool getStreamState(uint64_t &state, uint64_t &inc) const override { state = m_random.state; inc = m_random.inc; return true; }
    void setStreamState(uint64_t state, uint64_t inc) override { m_random.state = state; m_random.inc = inc; }
    std::string toString() const override { return fmt("Independent[sampleCount=%d]", (int) m_sampleCount); }
protected:
    Independent() {}
private:
    Pcg32 m_random;
};

// ---- integrators ---------------------------------------------------------------------------------------------
template <int TYPE> class TaggedIntegrator : public Integrator {
public:
    explicit TaggedIntegrator(const PropertyList &p) { readTermination(p); }
    int gpuType() const override { return TYPE; }
    std::string toString() const override {
        static const char *names[] = {"NormalIntegrator", "PathMats", "PathEms", "PathMis", "Whitted", "PrimaryShadow", "SimpleIntegrator", "AmbientOcclusion"};
        return std::string(names[TYPE]) + "[]";
    }
};
typedef TaggedIntegrator<NORI_INT_NORMALS> NormalIntegrator;
typedef TaggedIntegrator<NORI_INT_PATH_MATS> PathMatsIntegrator;
typedef TaggedIntegrator<NORI_INT_PATH_EMS> PathEmsIntegrator;
typedef TaggedIntegrator<NORI_INT_PATH_MIS> PathMisIntegrator;
typedef TaggedIntegrator<NORI_INT_WHITTED> WhittedIntegrator;
typedef TaggedIntegrator<NORI_INT_PRIMARY_SHADOW> PrimaryShadowIntegrator;
typedef TaggedIntegrator<NORI_INT_AO> AmbientOcclusionIntegrator;

/// scenes/pa3/ajax-simple.xml: point light with `position` and `energy`
class SimpleIntegrator : public Integrator {
public:
    explicit SimpleIntegrator(const PropertyList &p) { readTermination(p); m_position = p.getPoint("position"); m_energy = p.getColor("energy"); }
    int gpuType() const override { return NORI_INT_SIMPLE; }
    void describe(float pos[3], float energy[3]) const override {
        pos[0] = m_position.x; pos[1] = m_position.y; pos[2] = m_position.z;
        energy[0] = m_energy.r; energy[1] = m_energy.g; energy[2] = m_energy.b;
    }
    std::string toString() const override { return "SimpleIntegrator[]"; }
private:
    Point3f m_position; Color3f m_energy;
};

} // namespace

NORI_REGISTER_CLASS(GaussianFilter, "gaussian")
NORI_REGISTER_CLASS(MitchellNetravaliFilter, "mitchell")
NORI_REGISTER_CLASS(TentFilter, "tent")
NORI_REGISTER_CLASS(BoxFilter, "box")
NORI_REGISTER_CLASS(Diffuse, "diffuse")
NORI_REGISTER_CLASS(Mirror, "mirror")
NORI_REGISTER_CLASS(Dielectric, "dielectric")
NORI_REGISTER_CLASS(Microfacet, "microfacet")
NORI_REGISTER_CLASS(AreaLight, "area")
NORI_REGISTER_CLASS(Independent, "independent")
NORI_REGISTER_CLASS(NormalIntegrator, "normals")
NORI_REGISTER_CLASS(PathMatsIntegrator, "path_mats")
NORI_REGISTER_CLASS(PathEmsIntegrator, "path_ems")
NORI_REGISTER_CLASS(PathMisIntegrator, "path_mis")
NORI_REGISTER_CLASS(WhittedIntegrator, "whitted")
NORI_REGISTER_CLASS(PrimaryShadowIntegrator, "primary_shadow")
NORI_REGISTER_CLASS(AmbientOcclusionIntegrator, "ao")
NORI_REGISTER_CLASS(SimpleIntegrator, "simple")

NORI_NAMESPACE_END
