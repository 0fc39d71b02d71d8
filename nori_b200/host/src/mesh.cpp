// mesh.cpp — Mesh base class, the Wavefront OBJ loader ("obj") and the perspective camera.
// Reference: src/mesh.cpp:28-41,104-146; src/obj.cpp:30-173; src/perspective.cpp:33-86.
#include <nori/scene.h>
#include <nori_gpu.h>
#include <chrono>
#include <fstream>
#include <iostream>
#include <sstream>
#include <unordered_map>

NORI_NAMESPACE_BEGIN

Mesh::~Mesh() { delete m_bsdf; delete m_emitter; }

void Mesh::activate() {
    if (!m_bsdf) m_bsdf = static_cast<BSDF *>(NoriObjectFactory::createInstance("diffuse", PropertyList()));
}

void Mesh::addChild(NoriObject *obj) {
    switch (obj->getClassType()) {
        case EBSDF:
            if (m_bsdf) throw NoriException("Mesh: tried to register multiple BSDF instances!");
            m_bsdf = static_cast<BSDF *>(obj);
            break;
        case EEmitter:
            if (m_emitter) throw NoriException("Mesh: tried to register multiple Emitter instances!");
            m_emitter = static_cast<Emitter *>(obj);
            break;
        default:
            throw NoriException("Mesh::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
    }
}

std::string Mesh::toString() const {
    char head[256];
    snprintf(head, sizeof(head), "Mesh[\n  name = \"%s\",\n  vertexCount = %u,\n  triangleCount = %u,\n", m_name.c_str(), getVertexCount(), getTriangleCount());
    return std::string(head) + "  bsdf = " + (m_bsdf ? indent(m_bsdf->toString()) : std::string("null")) + ",\n  emitter = "
         + (m_emitter ? indent(m_emitter->toString()) : std::string("null")) + "\n]";
}

namespace {

/// One corner of an OBJ face: 1-based position / texcoord / normal indices, ~0 when absent.
struct ObjCorner {
    uint32_t p = ~0u, n = ~0u, uv = ~0u;
    ObjCorner() {}
    explicit ObjCorner(const std::string &s) {
        const std::vector<std::string> tok = tokenize(s, "/", true);
        if (tok.empty() || tok.size() > 3) throw NoriException("Invalid vertex data: \"%s\"", s);
        p = parseIndex(tok[0]);
        if (tok.size() >= 2 && !tok[1].empty()) uv = parseIndex(tok[1]);
        if (tok.size() >= 3 && !tok[2].empty()) n = parseIndex(tok[2]);
    }
    static uint32_t parseIndex(const std::string &t) {
        char *end = nullptr;
        const unsigned long v = strtoul(t.c_str(), &end, 10);
        if (*end != '\0') throw NoriException("Could not parse integer value \"%s\"", t);
        return (uint32_t) v;
    }
    bool operator==(const ObjCorner &o) const { return p == o.p && n == o.n && uv == o.uv; }
};
struct ObjCornerHash {
    size_t operator()(const ObjCorner &c) const { return ((size_t) c.p * 37u + c.uv) * 37u + c.n; }
};

/// `<mesh type="obj">`: properties `filename`, `toWorld`.
class WavefrontOBJ : public Mesh {
public:
    explicit WavefrontOBJ(const PropertyList &props) {
        const std::string path = getFileResolver()->resolve(props.getString("filename"));
        std::ifstream is(path);
        if (is.fail()) throw NoriException("Unable to open OBJ file \"%s\"!", path);
        const Transform trafo = props.getTransform("toWorld", Transform());
        std::cout << "Loading \"" << path << "\" .. " << std::flush;
        const auto t0 = std::chrono::steady_clock::now();

        std::vector<Point3f> positions; std::vector<Point2f> texcoords; std::vector<Normal3f> normals;
        std::vector<ObjCorner> corners;                      // unique corners in order of first use
        std::unordered_map<ObjCorner, uint32_t, ObjCornerHash> cornerIndex;
        const bool onDevice = deviceObjDedup();              // --device-dedup: the hash map below becomes one nori_gpu_obj_dedup call
        std::vector<uint32_t> cornerList;                    // (p, uv, n) of every face corner in file order
        std::string lineStr;
        while (std::getline(is, lineStr)) {
            std::istringstream line(lineStr);
            std::string prefix;
            line >> prefix;
            if (prefix == "v") {
                Point3f p; line >> p.x >> p.y >> p.z;
                p = trafo.applyPoint(p);
                m_lo = Point3f(std::fmin(m_lo.x, p.x), std::fmin(m_lo.y, p.y), std::fmin(m_lo.z, p.z));
                m_hi = Point3f(std::fmax(m_hi.x, p.x), std::fmax(m_hi.y, p.y), std::fmax(m_hi.z, p.z));
                positions.push_back(p);
            } else if (prefix == "vt") {
                Point2f tc; line >> tc.x >> tc.y;
                texcoords.push_back(tc);
            } else if (prefix == "vn") {
                Normal3f n; line >> n.x >> n.y >> n.z;
                normals.push_back(trafo.applyNormal(n).normalized());
            } else if (prefix == "f") {
                std::string v[4];
                line >> v[0] >> v[1] >> v[2] >> v[3];
                ObjCorner quad[6] = {ObjCorner(v[0]), ObjCorner(v[1]), ObjCorner(v[2])};
                int count = 3;
                if (!v[3].empty()) {                         // quad -> (0 1 2) (3 0 2)
                    quad[3] = ObjCorner(v[3]); quad[4] = quad[0]; quad[5] = quad[2];
                    count = 6;
                }
                if (onDevice) {
                    for (int i = 0; i < count; ++i) { cornerList.push_back(quad[i].p); cornerList.push_back(quad[i].uv); cornerList.push_back(quad[i].n); }
                    continue;
                }
                for (int i = 0; i < count; ++i) {
                    auto it = cornerIndex.find(quad[i]);
                    if (it == cornerIndex.end()) {
                        const uint32_t id = (uint32_t) corners.size();
                        cornerIndex.emplace(quad[i], id);
                        corners.push_back(quad[i]);
                        m_F.push_back(id);
                    } else m_F.push_back(it->second);
                }
            }
        }
        if (onDevice) {
            if (cornerList.empty()) throw NoriException("OBJ file \"%s\" has no faces", path);
            const size_t n = cornerList.size() / 3;
            std::vector<float> P(3 * positions.size()), T(2 * texcoords.size()), Nr(3 * normals.size());
            for (size_t i = 0; i < positions.size(); ++i) { P[3 * i] = positions[i].x; P[3 * i + 1] = positions[i].y; P[3 * i + 2] = positions[i].z; }
            for (size_t i = 0; i < texcoords.size(); ++i) { T[2 * i] = texcoords[i].x; T[2 * i + 1] = texcoords[i].y; }
            for (size_t i = 0; i < normals.size(); ++i) { Nr[3 * i] = normals[i].x; Nr[3 * i + 1] = normals[i].y; Nr[3 * i + 2] = normals[i].z; }
            m_F.resize(n); m_V.resize(3 * n);
            if (!normals.empty()) m_N.resize(3 * n);
            if (!texcoords.empty()) m_UV.resize(2 * n);
            uint32_t nV = 0;
            nori_gpu_ctx *ctx = operatorContext();
            gpuCheck(ctx, nori_gpu_obj_dedup(ctx, cornerList.data(), (uint32_t) n, P.data(), (uint32_t) positions.size(), T.empty() ? nullptr : T.data(), (uint32_t) texcoords.size(),
                                             Nr.empty() ? nullptr : Nr.data(), (uint32_t) normals.size(), m_F.data(), m_V.data(), m_N.empty() ? nullptr : m_N.data(),
                                             m_UV.empty() ? nullptr : m_UV.data(), &nV, nullptr), "nori_gpu_obj_dedup");
            m_V.resize(3 * (size_t) nV);
            if (!m_N.empty()) m_N.resize(3 * (size_t) nV);
            if (!m_UV.empty()) m_UV.resize(2 * (size_t) nV);
        }
        if (!onDevice) m_V.resize(3 * corners.size());
        for (size_t i = 0; i < corners.size(); ++i) {
            const Point3f &p = positions.at(corners[i].p - 1);
            m_V[3 * i] = p.x; m_V[3 * i + 1] = p.y; m_V[3 * i + 2] = p.z;
        }
        if (!onDevice && !normals.empty()) {
            m_N.resize(3 * corners.size());
            for (size_t i = 0; i < corners.size(); ++i) {
                const Normal3f &n = normals.at(corners[i].n - 1);
                m_N[3 * i] = n.x; m_N[3 * i + 1] = n.y; m_N[3 * i + 2] = n.z;
            }
        }
        if (!onDevice && !texcoords.empty()) {
            m_UV.resize(2 * corners.size());
            for (size_t i = 0; i < corners.size(); ++i) {
                const Point2f &t = texcoords.at(corners[i].uv - 1);
                m_UV[2 * i] = t.x; m_UV[2 * i + 1] = t.y;
            }
        }
        m_name = path;
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::cout << "done. (V=" << getVertexCount() << ", F=" << getTriangleCount() << ", took " << timeString(ms) << " and "
                  << memString(m_F.size() * sizeof(uint32_t) + sizeof(float) * (m_V.size() + m_N.size() + m_UV.size())) << ")" << std::endl;
    }
};

/// `<camera type="perspective">`: width, height, toWorld, fov (horizontal, degrees), nearClip, farClip.
class PerspectiveCamera : public Camera {
public:
    explicit PerspectiveCamera(const PropertyList &p) {
        m_outputSize = Vector2i(p.getInteger("width", 1280), p.getInteger("height", 720));
        m_cameraToWorld = p.getTransform("toWorld", Transform());
        m_fov = p.getFloat("fov", 30.0f);
        m_nearClip = p.getFloat("nearClip", 1e-4f);
        m_farClip = p.getFloat("farClip", 1e4f);
    }

    void activate() override {
        const float aspect = m_outputSize.x / (float) m_outputSize.y;
        const float recip = 1.0f / (m_farClip - m_nearClip), cot = 1.0f / std::tan(degToRad(m_fov / 2.0f));
        // camera space -> clip space on the plane z = 1, then -> [0,1]^2 sample space:
        // scale(-0.5, -0.5 aspect, 1) * translate(-1, -1/aspect, 0) * perspective
        Matrix4f persp = Matrix4f::identity();
        persp(0, 0) = cot; persp(1, 1) = cot;
        persp(2, 2) = m_farClip * recip; persp(2, 3) = -m_nearClip * m_farClip * recip;
        persp(3, 2) = 1.0f; persp(3, 3) = 0.0f;
        const Vector3f s(-0.5f, -0.5f * aspect, 1.0f), t(-1.0f, -1.0f / aspect, 0.0f);
        Matrix4f scaleTranslate = Matrix4f::identity();      // scale * translate: linear part s, translation s * t
        for (int i = 0; i < 3; ++i) { scaleTranslate(i, i) = s[i]; scaleTranslate(i, 3) = s[i] * t[i]; }
        m_sampleToCamera = Transform(scaleTranslate * persp).inverse();
        if (!m_rfilter) m_rfilter = static_cast<ReconstructionFilter *>(NoriObjectFactory::createInstance("gaussian", PropertyList()));
    }

    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != EReconstructionFilter)
            throw NoriException("Camera::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        if (m_rfilter) throw NoriException("Camera: tried to register multiple reconstruction filters!");
        m_rfilter = static_cast<ReconstructionFilter *>(obj);
    }

    /// src/perspective.cpp:88-109, in the evaluation orders the compiled reference uses (the device's sampleRay, csrc/scene.cuh)
    Color3f sampleRay(Ray3f &ray, const Point2f &samplePosition, const Point2f &) const override {
        const float invW = 1.0f / (float) m_outputSize.x, invH = 1.0f / (float) m_outputSize.y;
        const Point3f nearP = m_sampleToCamera.applyPoint(Point3f(samplePosition.x * invW, samplePosition.y * invH, 0.0f));
        const Vector3f d = Vector3f(nearP.x, nearP.y, nearP.z).normalized();
        const float invZ = 1.0f / d.z;
        ray.o = m_cameraToWorld.applyPoint(Point3f(0, 0, 0));
        ray.d = m_cameraToWorld.applyVector(d);
        ray.mint = m_nearClip * invZ;
        ray.maxt = m_farClip * invZ;
        return Color3f(1.0f);
    }

    void describe(float s2c[16], float c2w[16], float &nearClip, float &farClip, float &fov) const override {
        m_sampleToCamera.getMatrix().toRowMajor(s2c);
        m_cameraToWorld.getMatrix().toRowMajor(c2w);
        nearClip = m_nearClip; farClip = m_farClip; fov = m_fov;
    }

    std::string toString() const override {
        char buf[512];
        snprintf(buf, sizeof(buf), "PerspectiveCamera[\n  cameraToWorld = %s,\n  outputSize = [%d, %d],\n  fov = %f,\n  clip = [%f, %f],\n  rfilter = %s\n]",
                 indent(m_cameraToWorld.toString(), 18).c_str(), m_outputSize.x, m_outputSize.y, m_fov, m_nearClip, m_farClip,
                 m_rfilter ? indent(m_rfilter->toString()).c_str() : "null");
        return buf;
    }
private:
    Transform m_sampleToCamera, m_cameraToWorld;
    float m_fov, m_nearClip, m_farClip;
};

} // namespace

NORI_REGISTER_CLASS(WavefrontOBJ, "obj")
NORI_REGISTER_CLASS(PerspectiveCamera, "perspective")

NORI_NAMESPACE_END
