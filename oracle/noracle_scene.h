// noracle_scene.h — flattened scene, Accel (brute force + SAH BVH with counters), camera,
// ImageBlock and the render loop of the standalone CPU port (TEST INFRASTRUCTURE; see
// noracle_math.h for the usage rule). Citations relative to /root/reference.
#pragma once
#include "noracle_integrators.h"
#include <string>

namespace noracle {

// ---------------------------------------------------------------------------------------
// Mesh buffers exactly as Mesh stores them (include/nori/mesh.h:172-175): column-major
// V 3xnV, N 3xnV (may be empty), UV 2xnV (may be empty), F 3xnF u32.
// ---------------------------------------------------------------------------------------
struct MeshData {
    std::vector<float> V, N, UV;
    std::vector<uint32_t> F;
    int bsdf, emitter;          // indices into Scene::bsdfs / Scene::emitters (-1: none)
    uint32_t nTris() const { return (uint32_t) (F.size() / 3); }
    uint32_t nVerts() const { return (uint32_t) (V.size() / 3); }
    V3 pos(uint32_t i) const { return mk(V[3 * i], V[3 * i + 1], V[3 * i + 2]); }
    V3 nrm(uint32_t i) const { return mk(N[3 * i], N[3 * i + 1], N[3 * i + 2]); }
    /// Mesh::surfaceArea (src/mesh.cpp:43-49)
    float surfaceArea(uint32_t f) const {
        V3 p0 = pos(F[3 * f]), p1 = pos(F[3 * f + 1]), p2 = pos(F[3 * f + 2]);
        return 0.5f * norm(cross(p1 - p0, p2 - p0));
    }
};

struct EmitterData { int mesh; C3 radiance; DiscretePdf triPdf; float area; };

/// PerspectiveCamera state after activate() (src/perspective.cpp:53-86); matrices row-major.
struct CameraData {
    float sampleToCamera[16], cameraToWorld[16];
    float invW, invH, nearClip, farClip;
    int width, height;
};

/// Transform * Point3f (include/nori/transform.h:76-80): homogeneous 4x4 product, then the
/// divide. Eigen evaluates the fixed-size 4x4 * 4x1 product column-wise:
/// ((c0*x + c1*y) + c2*z) + c3*w — pinned against the compiled reference.
inline V3 mulPoint(const float *m, V3 p) {
    float r[4];
    for (int i = 0; i < 4; ++i)
        r[i] = ((m[4 * i + 0] * p.x + m[4 * i + 1] * p.y) + m[4 * i + 2] * p.z) + m[4 * i + 3] * 1.0f;
    return mk(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
}
/// Transform * Vector3f (transform.h:66-68): topLeftCorner<3,3>() * v. Eigen evaluates this small
/// product coefficient-wise with the unrolled fixed-size reduction x0 + (x1 + x2) (Redux.h:94-103),
/// like dot() — pinned against the compiled reference on a rotated camera (bunny, table).
inline V3 mulVector(const float *m, V3 v) {
    float r[3];
    for (int i = 0; i < 3; ++i)
        r[i] = m[4 * i + 0] * v.x + (m[4 * i + 1] * v.y + m[4 * i + 2] * v.z);
    return mk(r[0], r[1], r[2]);
}

/// PerspectiveCamera::sampleRay (src/perspective.cpp:88-109)
inline Ray sampleRay(const CameraData &c, V2 samplePosition) {
    V3 nearP = mulPoint(c.sampleToCamera, mk(samplePosition.x * c.invW, samplePosition.y * c.invH, 0.0f));
    V3 d = normalized(nearP);
    float invZ = 1.0f / d.z;
    Ray ray;
    ray.o = mulPoint(c.cameraToWorld, mk(0, 0, 0));
    ray.d = mulVector(c.cameraToWorld, d);
    ray.mint = c.nearClip * invZ;
    ray.maxt = c.farClip * invZ;
    ray.update();
    return ray;
}

// ---------------------------------------------------------------------------------------
// Accel: S6 multi-mesh extension of src/accel.cpp:35-111 (brute force) + binary SAH BVH
// ---------------------------------------------------------------------------------------
struct BvhNode {                 // 32 B of bounds+link per node: the "one child AABB + link" unit of §8d
    BBox box;
    uint32_t left;               // inner: index of left child (right = left+1); leaf: first prim ref
    uint32_t count;              // 0: inner node; >0: number of prim refs in the leaf
};

struct TraversalCounters { uint64_t rays, boxTests, triTests; TraversalCounters() : rays(0), boxTests(0), triTests(0) {} };

struct Scene;
struct Accel {
    const Scene *scene;
    std::vector<uint32_t> triMesh, triLocal;   // global triangle id -> (mesh, local index)
    std::vector<uint32_t> refs;                // BVH leaf order -> global triangle id
    std::vector<BvhNode> nodes;
    bool useBvh;
    int isectMode;                             // 0: Moeller-Trumbore of src/mesh.cpp:51-88; 1: watertight (noracle_math.h)
    mutable TraversalCounters *counters;       // optional (single-threaded callers only)
    mutable TraversalCounters *shadowCounters; // optional: shadow rays counted apart from closest-hit rays
    Accel() : scene(0), useBvh(true), isectMode(0), counters(0), shadowCounters(0), active(0) {}
    void build(const Scene *s, int leafSize = 4);
    mutable TraversalCounters *active;         // counter set of the ray being traced
    bool testTri(uint32_t gid, const Ray &ray, float &u, float &v, float &t) const;
    /// closest hit: returns global triangle id or 0xffffffff; any hit when shadowRay.
    uint32_t trace(const Ray &ray_, bool shadowRay, float &tOut, float &uOut, float &vOut) const;
};

struct Scene {
    std::vector<MeshData> meshes;
    std::vector<Bsdf> bsdfs;
    std::vector<EmitterData> emitters;
    CameraData camera;
    float filterRadius, filterTable[33];      // ImageBlock's tabulated filter (src/block.cpp:31-39)
    IntegratorParams integrator;
    int sampleCount;
    Accel accel;
    Scene() : sampleCount(1) { setFilter(gaussianFilter()); }
    void setFilter(const Filter &f) {
        filterRadius = f.radius;
        for (int i = 0; i < 32; ++i) filterTable[i] = f.eval((f.radius * i) / 32);
        filterTable[32] = 0.0f;
    }

    /// Build emitter CDFs (S4) and the acceleration structure; call after filling the arrays.
    void activate(bool useBvh = true) {
        for (size_t e = 0; e < emitters.size(); ++e) {
            EmitterData &em = emitters[e];
            const MeshData &m = meshes[em.mesh];
            em.triPdf.clear();
            for (uint32_t f = 0; f < m.nTris(); ++f) em.triPdf.append(m.surfaceArea(f));
            em.area = em.triPdf.normalize();
        }
        accel.useBvh = useBvh;
        accel.build(this);
    }

    // ---- World concept (noracle_integrators.h) ----
    bool intersect(const Ray &ray, Hit &its) const;
    bool occluded(const Ray &ray) const { float t, u, v; return accel.trace(ray, true, t, u, v) != 0xffffffffu; }
    C3 bsdfSample(int mesh, BsdfQuery &q, V2 s) const { return bsdfs[meshes[mesh].bsdf].sample(q, s); }
    C3 bsdfEval(int mesh, const BsdfQuery &q) const { return bsdfs[meshes[mesh].bsdf].eval(q); }
    float bsdfPdf(int mesh, const BsdfQuery &q) const { return bsdfs[meshes[mesh].bsdf].pdf(q); }
    bool bsdfIsDiffuse(int mesh) const { return bsdfs[meshes[mesh].bsdf].isDiffuse(); }
    int emitterCount() const { return (int) emitters.size(); }
    int meshEmitter(int mesh) const { return meshes[mesh].emitter; }
    C3 emitterRadiance(int e) const { return emitters[e].radiance; }
    float emitterArea(int e) const { return emitters[e].area; }
    void emitterSamplePoint(int e, V2 xi, V3 &p, V3 &n) const {
        const EmitterData &em = emitters[e];
        const MeshData &m = meshes[em.mesh];
        float x = xi.x;
        uint32_t f = (uint32_t) em.triPdf.sampleReuse(x);
        float su = std::sqrt(1.0f - x);
        float b0 = 1.0f - su, b1 = xi.y * su, b2 = 1.0f - b0 - b1;
        uint32_t i0 = m.F[3 * f], i1 = m.F[3 * f + 1], i2 = m.F[3 * f + 2];
        V3 p0 = m.pos(i0), p1 = m.pos(i1), p2 = m.pos(i2);
        p = (b0 * p0 + b1 * p1) + b2 * p2;
        if (!m.N.empty()) n = normalized((b0 * m.nrm(i0) + b1 * m.nrm(i1)) + b2 * m.nrm(i2));
        else n = normalized(cross(p1 - p0, p2 - p0));
    }
};

inline bool Accel::testTri(uint32_t gid, const Ray &ray, float &u, float &v, float &t) const {
    const MeshData &m = scene->meshes[triMesh[gid]];
    uint32_t f = triLocal[gid];
    if (active) active->triTests++;
    if (isectMode == 1) return rayTriangleWatertight(m.pos(m.F[3 * f]), m.pos(m.F[3 * f + 1]), m.pos(m.F[3 * f + 2]), ray, u, v, t);
    return rayTriangle(m.pos(m.F[3 * f]), m.pos(m.F[3 * f + 1]), m.pos(m.F[3 * f + 2]), ray, u, v, t);
}

inline uint32_t Accel::trace(const Ray &ray_, bool shadowRay, float &tOut, float &uOut, float &vOut) const {
    Ray ray(ray_);
    uint32_t best = 0xffffffffu;
    active = (shadowRay && shadowCounters) ? shadowCounters : counters;
    if (active) active->rays++;
    if (!useBvh) {
        // src/accel.cpp:42-55 over the concatenation of all meshes (mesh order, then index):
        // `t <= maxt` with maxt tightened per hit => among equal t the highest id wins.
        for (uint32_t gid = 0; gid < (uint32_t) triMesh.size(); ++gid) {
            float u, v, t;
            if (testTri(gid, ray, u, v, t)) {
                if (shadowRay) { tOut = t; uOut = u; vOut = v; return gid; }
                ray.maxt = t; tOut = t; uOut = u; vOut = v; best = gid;
            }
        }
        return best;
    }
    if (nodes.empty()) return best;
    uint32_t stack[64]; int sp = 0; stack[sp++] = 0;
    while (sp > 0) {
        const BvhNode &n = nodes[stack[--sp]];
        if (active) active->boxTests++;
        if (!n.box.rayIntersect(ray)) continue;
        if (n.count > 0) {
            for (uint32_t i = 0; i < n.count; ++i) {
                uint32_t gid = refs[n.left + i];
                float u, v, t;
                if (testTri(gid, ray, u, v, t)) {
                    if (shadowRay) { tOut = t; uOut = u; vOut = v; return gid; }
                    // same tie rule as the brute-force loop: equal t -> highest id wins
                    if (t < ray.maxt || best == 0xffffffffu || gid > best) {
                        ray.maxt = t; tOut = t; uOut = u; vOut = v; best = gid;
                    }
                }
            }
        } else {
            // near child first along the split of the children's centres
            const BvhNode &a = nodes[n.left], &b = nodes[n.left + 1];
            int ax = n.box.largestAxis();
            float ca = a.box.min[ax] + a.box.max[ax], cb = b.box.min[ax] + b.box.max[ax];
            bool aFirst = (ray.d[ax] >= 0) == (ca <= cb);
            stack[sp++] = aFirst ? n.left + 1 : n.left;
            stack[sp++] = aFirst ? n.left : n.left + 1;
        }
    }
    return best;
}

/// Binned SAH build (16 bins), leaves of <= leafSize triangles; boxes from
/// Mesh::getBoundingBox(i)/getCentroid(i) (src/mesh.cpp:90-102).
inline void Accel::build(const Scene *s, int leafSize) {
    scene = s;
    triMesh.clear(); triLocal.clear(); refs.clear(); nodes.clear();
    for (size_t m = 0; m < s->meshes.size(); ++m)
        for (uint32_t f = 0; f < s->meshes[m].nTris(); ++f) { triMesh.push_back((uint32_t) m); triLocal.push_back(f); }
    uint32_t n = (uint32_t) triMesh.size();
    if (n == 0) return;
    std::vector<BBox> tb(n); std::vector<V3> tc(n);
    for (uint32_t g = 0; g < n; ++g) {
        const MeshData &m = s->meshes[triMesh[g]]; uint32_t f = triLocal[g];
        V3 p0 = m.pos(m.F[3 * f]), p1 = m.pos(m.F[3 * f + 1]), p2 = m.pos(m.F[3 * f + 2]);
        tb[g].expandBy(p0); tb[g].expandBy(p1); tb[g].expandBy(p2);
        tc[g] = (1.0f / 3.0f) * ((p0 + p1) + p2);
    }
    refs.resize(n);
    for (uint32_t g = 0; g < n; ++g) refs[g] = g;
    nodes.reserve(2 * n);
    nodes.push_back(BvhNode());
    struct Job { uint32_t node, lo, hi; };
    std::vector<Job> todo; Job j0 = {0, 0, n}; todo.push_back(j0);
    const int NB = 16;
    // Node boxes are stored grown by a scene-relative delta so that the slab test can never cull a
    // triangle the Moeller-Trumbore test would report (its t may round an ulp outside the exact
    // box): the BVH must return what the reference's brute-force loop returns.
    float delta = 0.0f;
    {
        BBox all; for (uint32_t g = 0; g < n; ++g) all.expandBy(tb[g]);
        V3 e = all.max - all.min;
        float diag = std::sqrt(dot(e, e)), maxAbs = 0.0f;
        for (int k = 0; k < 3; ++k) maxAbs = std::max(maxAbs, std::max(std::abs(all.min[k]), std::abs(all.max[k])));
        delta = 4e-6f * std::max(diag, maxAbs);
    }
    while (!todo.empty()) {
        Job j = todo.back(); todo.pop_back();
        BBox box, cbox;
        for (uint32_t i = j.lo; i < j.hi; ++i) { box.expandBy(tb[refs[i]]); cbox.expandBy(tc[refs[i]]); }
        nodes[j.node].box = box;
        nodes[j.node].box.min = box.min - mk(delta, delta, delta); nodes[j.node].box.max = box.max + mk(delta, delta, delta);
        uint32_t cnt = j.hi - j.lo;
        bool leaf = cnt <= (uint32_t) leafSize;
        int bestAxis = -1, bestBin = -1; float bestCost = kInf;
        if (!leaf) {
            for (int ax = 0; ax < 3; ++ax) {
                float lo = cbox.min[ax], ext = cbox.max[ax] - lo;
                if (!(ext > 0.0f)) continue;
                BBox bb[NB]; uint32_t bc[NB] = {0};
                for (uint32_t i = j.lo; i < j.hi; ++i) {
                    int b = std::min(NB - 1, (int) ((tc[refs[i]][ax] - lo) / ext * NB));
                    bb[b].expandBy(tb[refs[i]]); bc[b]++;
                }
                float rightArea[NB]; uint32_t rightCnt[NB]; BBox acc; uint32_t c = 0;
                for (int b = NB - 1; b > 0; --b) { acc.expandBy(bb[b]); c += bc[b]; rightArea[b] = c ? acc.surfaceArea() : 0; rightCnt[b] = c; }
                acc.reset(); c = 0;
                for (int b = 0; b < NB - 1; ++b) {
                    if (bc[b]) acc.expandBy(bb[b]);
                    c += bc[b];
                    if (c == 0 || rightCnt[b + 1] == 0) continue;
                    float cost = acc.surfaceArea() * c + rightArea[b + 1] * rightCnt[b + 1];
                    if (cost < bestCost) { bestCost = cost; bestAxis = ax; bestBin = b; }
                }
            }
        }
        uint32_t mid = j.lo;
        if (!leaf && bestAxis >= 0) {
            float lo = cbox.min[bestAxis], ext = cbox.max[bestAxis] - lo;
            uint32_t *first = &refs[0] + j.lo, *last = &refs[0] + j.hi;
            const int axis = bestAxis, bin = bestBin;
            uint32_t *m = std::partition(first, last, [&](uint32_t g) {
                return std::min(NB - 1, (int) ((tc[g][axis] - lo) / ext * NB)) <= bin; });
            mid = (uint32_t) (m - &refs[0]);
        }
        if (!leaf && (mid == j.lo || mid == j.hi)) {
            if (cnt <= 16) leaf = true;         // degenerate: identical centroids
            else mid = j.lo + cnt / 2;
        }
        if (leaf) { nodes[j.node].left = j.lo; nodes[j.node].count = cnt; continue; }
        uint32_t l = (uint32_t) nodes.size();
        nodes.push_back(BvhNode()); nodes.push_back(BvhNode());
        nodes[j.node].left = l; nodes[j.node].count = 0;
        Job a = {l, j.lo, mid}, b = {l + 1, mid, j.hi};
        todo.push_back(a); todo.push_back(b);
    }
}

/// Intersection post-processing (src/accel.cpp:57-108).
inline void fillHit(const Scene &s, uint32_t gid, float t, float u, float v, Hit &its) {
    uint32_t mi = s.accel.triMesh[gid], f = s.accel.triLocal[gid];
    const MeshData &m = s.meshes[mi];
    its.t = t; its.mesh = (int) mi; its.prim = f;
    V3 bary = mk(1 - (u + v), u, v);             // `1-its.uv.sum()`: 2-vector sum = u+v
    uint32_t i0 = m.F[3 * f], i1 = m.F[3 * f + 1], i2 = m.F[3 * f + 2];
    V3 p0 = m.pos(i0), p1 = m.pos(i1), p2 = m.pos(i2);
    its.p = (bary.x * p0 + bary.y * p1) + bary.z * p2;
    if (!m.UV.empty()) {
        its.uv.x = (bary.x * m.UV[2 * i0] + bary.y * m.UV[2 * i1]) + bary.z * m.UV[2 * i2];
        its.uv.y = (bary.x * m.UV[2 * i0 + 1] + bary.y * m.UV[2 * i1 + 1]) + bary.z * m.UV[2 * i2 + 1];
    } else { its.uv.x = u; its.uv.y = v; }
    its.geo = Frame(normalized(cross(p1 - p0, p2 - p0)));
    if (!m.N.empty())
        its.sh = Frame(normalizedDynamic((bary.x * m.nrm(i0) + bary.y * m.nrm(i1)) + bary.z * m.nrm(i2)));
    else its.sh = its.geo;
}

inline bool Scene::intersect(const Ray &ray, Hit &its) const {
    float t, u, v;
    uint32_t gid = accel.trace(ray, false, t, u, v);
    if (gid == 0xffffffffu) return false;
    fillHit(*this, gid, t, u, v, its);
    return true;
}

// ---------------------------------------------------------------------------------------
// REF: ImageBlock (src/block.cpp:27-114) and BlockGenerator (src/block.cpp:121-164)
// ---------------------------------------------------------------------------------------
static const int kBlockSize = 32;          // NORI_BLOCK_SIZE (include/nori/block.h:29)
static const int kFilterRes = 32;          // NORI_FILTER_RESOLUTION (include/nori/rfilter.h:24)

struct FilterTable {
    float radius, lookupFactor, table[kFilterRes + 1];
    int border;
    void init(const Filter &f) {                              // block.cpp:31-39
        float t[kFilterRes + 1];
        for (int i = 0; i < kFilterRes; ++i) t[i] = f.eval((f.radius * i) / kFilterRes);
        t[kFilterRes] = 0.0f;
        init(f.radius, t);
    }
    void init(float r, const float *t) {
        radius = r;
        border = (int) std::ceil(radius - 0.5f);
        std::memcpy(table, t, sizeof(table));
        lookupFactor = kFilterRes / radius;
    }
};

struct ImageBlock {
    int ox, oy, w, h, border;         // offset, size (without border)
    FilterTable ft;
    std::vector<float> data;          // row-major (h+2b) x (w+2b) x RGBW
    uint64_t dropped;
    ImageBlock() : ox(0), oy(0), w(0), h(0), border(0), dropped(0) {}
    void init(int w_, int h_, const FilterTable &f) { ft = f; border = f.border; w = w_; h = h_; data.assign((size_t) rows() * cols() * 4, 0.0f); }
    int rows() const { return h + 2 * border; }
    int cols() const { return w + 2 * border; }
    void clear() { std::fill(data.begin(), data.end(), 0.0f); }

    void put(V2 _pos, C3 value) {                              // block.cpp:74-103
        if (!isValid(value)) { dropped++; return; }
        float px = _pos.x - 0.5f - (ox - border), py = _pos.y - 0.5f - (oy - border);
        int x0 = (int) std::ceil(px - ft.radius), y0 = (int) std::ceil(py - ft.radius);
        int x1 = (int) std::floor(px + ft.radius), y1 = (int) std::floor(py + ft.radius);
        x0 = std::max(x0, 0); y0 = std::max(y0, 0);
        x1 = std::min(x1, cols() - 1); y1 = std::min(y1, rows() - 1);
        float wx[16], wy[16];
        for (int x = x0, i = 0; x <= x1; ++x) wx[i++] = ft.table[(int) (std::abs(x - px) * ft.lookupFactor)];
        for (int y = y0, i = 0; y <= y1; ++y) wy[i++] = ft.table[(int) (std::abs(y - py) * ft.lookupFactor)];
        const float val[4] = {value.r, value.g, value.b, 1.0f};
        for (int y = y0, yr = 0; y <= y1; ++y, ++yr)
            for (int x = x0, xr = 0; x <= x1; ++x, ++xr) {
                float *p = &data[((size_t) y * cols() + x) * 4];
                for (int c = 0; c < 4; ++c) p[c] += val[c] * wx[xr] * wy[yr];
            }
    }
    void put(const ImageBlock &b) {                            // block.cpp:105-114
        int offx = b.ox - ox + border - b.border, offy = b.oy - oy + border - b.border;
        for (int y = 0; y < b.rows(); ++y)
            for (int x = 0; x < b.cols(); ++x)
                for (int c = 0; c < 4; ++c)
                    data[((size_t) (offy + y) * cols() + offx + x) * 4 + c] += b.data[((size_t) y * b.cols() + x) * 4 + c];
    }
    /// toBitmap (block.cpp:57-63, color.h:112-117): rgb / w, 0 where w == 0; row-major h x w x 3
    void toBitmap(std::vector<float> &rgb) const {
        rgb.assign((size_t) w * h * 3, 0.0f);
        for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
            const float *p = &data[((size_t) (y + border) * cols() + x + border) * 4];
            if (p[3] != 0) for (int c = 0; c < 3; ++c) rgb[((size_t) y * w + x) * 3 + c] = p[c] / p[3];
        }
    }
};

/// Spiral block order from the image centre (block.cpp:121-164). Returns (ox, oy, w, h) per block.
inline void blockOrder(int W, int H, int bs, std::vector<int> &out) {
    out.clear();
    int nbx = (int) std::ceil(W / (float) bs), nby = (int) std::ceil(H / (float) bs);
    int left = nbx * nby, dir = 0 /*right*/, bx = nbx / 2, by = nby / 2, stepsLeft = 1, numSteps = 1;
    while (left > 0) {
        int px = bx * bs, py = by * bs;
        out.push_back(px); out.push_back(py);
        out.push_back(std::min(W - px, bs)); out.push_back(std::min(H - py, bs));
        if (--left == 0) break;
        do {
            switch (dir) { case 0: ++bx; break; case 1: ++by; break; case 2: --bx; break; case 3: --by; break; }
            if (--stepsLeft == 0) {
                dir = (dir + 1) % 4;
                if (dir == 2 || dir == 0) ++numSteps;
                stepsLeft = numSteps;
            }
        } while (bx < 0 || by < 0 || bx >= nbx || by >= nby);
    }
}

// ---------------------------------------------------------------------------------------
// Render loop (src/main.cpp:38-67 renderBlock, :93-121 block loop). Two RNG stream modes
// (SURVEY §8d "Value distributions / seeds"):
//   kBlockStream     the reference's: one PCG32 per block, seed(offset.x, offset.y), consumed
//                    serially (src/independent.cpp:48-53) — bit-exact with the reference.
//   kPerSampleStream seed(initstate = y*W + x, initseq = sample index) per camera sample —
//                    what the GPU throughput path uses; independent of any partitioning.
// ---------------------------------------------------------------------------------------
enum StreamMode { kBlockStream = 0, kPerSampleStream = 1 };

struct RenderJob { int mode; uint32_t sampleFirst, sampleLast; };   // samples [first, last) of every pixel

inline void renderBlock(const Scene &s, const RenderJob &job, ImageBlock &block, PathStats &st) {
    Sampler smp;
    if (job.mode == kBlockStream) smp.rng.seed((uint64_t) (int64_t) block.ox, (uint64_t) (int64_t) block.oy);
    block.clear();
    for (int y = 0; y < block.h; ++y) for (int x = 0; x < block.w; ++x)
        for (uint32_t i = job.sampleFirst; i < job.sampleLast; ++i) {
            int gx = x + block.ox, gy = y + block.oy;
            if (job.mode == kPerSampleStream)
                smp.rng.seed((uint64_t) gy * (uint64_t) s.camera.width + (uint64_t) gx, (uint64_t) i);
            V2 r = smp.next2D();
            V2 pixelSample = {(float) gx + r.x, (float) gy + r.y};
            smp.next2D();                                       // apertureSample (unused by the pinhole)
            Ray ray = sampleRay(s.camera, pixelSample);
            C3 value = Li(s, smp, ray, s.integrator, st);       // camera weight is 1 (perspective.cpp:108)
            block.put(pixelSample, value);
        }
}

} // namespace noracle
