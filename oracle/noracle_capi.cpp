// noracle_capi.cpp — C API over the standalone CPU port (TEST INFRASTRUCTURE).
// Builds into oracle/liboracle.so (oracle/Makefile). Loaded with ctypes by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs ONLY.
// The product library (libnori_b200.so) never links or loads this file.
#include "noracle_scene.h"
#include "noracle_pathgraph.h"
#include "noracle_pathgraph_scatter.h"
#include <thread>
#include <mutex>
#include <atomic>

using namespace noracle;

extern "C" {

void *noracle_scene_create() { return new Scene(); }
void noracle_scene_destroy(void *s) { delete (Scene *) s; }

int noracle_scene_add_bsdf(void *s_, int type, const float *p /* albedo/kd rgb, alpha, intIOR, extIOR */) {
    Scene *s = (Scene *) s_;
    Bsdf b;
    switch (type) {
    case kDiffuse:    b = makeDiffuse(mkc(p[0], p[1], p[2])); break;
    case kMirror:     b = makeMirror(); break;
    case kDielectric: b = makeDielectric(p[4], p[5]); break;
    default:          b = makeMicrofacet(p[3], p[4], p[5], mkc(p[0], p[1], p[2])); break;
    }
    s->bsdfs.push_back(b);
    return (int) s->bsdfs.size() - 1;
}

int noracle_scene_add_mesh(void *s_, const float *V, uint32_t nV, const float *N, const float *UV,
                           const uint32_t *F, uint32_t nF, int bsdf) {
    Scene *s = (Scene *) s_;
    s->meshes.push_back(MeshData());
    MeshData &m = s->meshes.back();
    m.V.assign(V, V + 3 * (size_t) nV);
    if (N) m.N.assign(N, N + 3 * (size_t) nV);
    if (UV) m.UV.assign(UV, UV + 2 * (size_t) nV);
    m.F.assign(F, F + 3 * (size_t) nF);
    m.bsdf = bsdf; m.emitter = -1;
    return (int) s->meshes.size() - 1;
}

int noracle_scene_add_emitter(void *s_, int mesh, const float *radiance) {
    Scene *s = (Scene *) s_;
    EmitterData e; e.mesh = mesh; e.radiance = mkc(radiance[0], radiance[1], radiance[2]); e.area = 0;
    s->emitters.push_back(e);
    s->meshes[mesh].emitter = (int) s->emitters.size() - 1;
    return s->meshes[mesh].emitter;
}

void noracle_scene_set_camera(void *s_, const float *s2c, const float *c2w, int W, int H, float nearClip, float farClip) {
    CameraData &c = ((Scene *) s_)->camera;
    std::memcpy(c.sampleToCamera, s2c, 64); std::memcpy(c.cameraToWorld, c2w, 64);
    c.width = W; c.height = H; c.nearClip = nearClip; c.farClip = farClip;
    c.invW = 1.0f / (float) W; c.invH = 1.0f / (float) H;     // perspective.cpp:38 cwiseInverse
}

void noracle_scene_set_filter(void *s_, int type, float radius, float p0, float p1) {
    Filter f = {type, radius, p0, p1};
    ((Scene *) s_)->setFilter(f);
}
void noracle_scene_set_filter_table(void *s_, float radius, const float *table33) {
    Scene *s = (Scene *) s_;
    s->filterRadius = radius; std::memcpy(s->filterTable, table33, sizeof(s->filterTable));
}

void noracle_scene_set_integrator(void *s_, int type, int rrStart, float rrMax, int maxDepth,
                                  const float *lightPos, const float *lightEnergy) {
    IntegratorParams &ip = ((Scene *) s_)->integrator;
    ip.type = type; ip.rrStart = rrStart; ip.rrMax = rrMax; ip.maxDepth = maxDepth;
    if (lightPos) ip.lightPos = mk(lightPos[0], lightPos[1], lightPos[2]);
    if (lightEnergy) ip.lightEnergy = mkc(lightEnergy[0], lightEnergy[1], lightEnergy[2]);
}

void noracle_scene_activate(void *s_, int useBvh) { ((Scene *) s_)->activate(useBvh != 0); }
void noracle_scene_set_use_bvh(void *s_, int useBvh) { ((Scene *) s_)->accel.useBvh = useBvh != 0; }
void noracle_scene_set_intersect_mode(void *s_, int mode) { ((Scene *) s_)->accel.isectMode = mode; }
uint32_t noracle_scene_triangle_count(void *s_) { return (uint32_t) ((Scene *) s_)->accel.triMesh.size(); }
uint32_t noracle_scene_bvh_node_count(void *s_) { return (uint32_t) ((Scene *) s_)->accel.nodes.size(); }

// ---- operators ---------------------------------------------------------------------------
void noracle_pcg32_uints(uint64_t initstate, uint64_t initseq, int64_t advance, int n, uint32_t *out) {
    Pcg32 r(initstate, initseq);
    if (advance) r.advance(advance);
    for (int i = 0; i < n; ++i) out[i] = r.nextUInt();
}
void noracle_pcg32_floats(uint64_t initstate, uint64_t initseq, int64_t advance, int n, float *out) {
    Pcg32 r(initstate, initseq);
    if (advance) r.advance(advance);
    for (int i = 0; i < n; ++i) out[i] = r.nextFloat();
}

/// rays: n x 8 floats [ox oy oz mint dx dy dz maxt]
void noracle_camera_rays(void *s_, int n, const float *pixelSamples, float *rays) {
    const Scene *s = (const Scene *) s_;
    for (int i = 0; i < n; ++i) {
        V2 ps = {pixelSamples[2 * i], pixelSamples[2 * i + 1]};
        Ray r = sampleRay(s->camera, ps);
        float *o = rays + 8 * (size_t) i;
        o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.mint;
        o[4] = r.d.x; o[5] = r.d.y; o[6] = r.d.z; o[7] = r.maxt;
    }
}

static Ray rayFrom8(const float *p) { return Ray(mk(p[0], p[1], p[2]), mk(p[4], p[5], p[6]), p[3], p[7]); }

/// Accel::rayIntersect operator. tri = mesh-local triangle index, 0xffffffff on miss.
/// counters (nullable) receives {rays, boxTests, triTests} (adds; single-threaded).
void noracle_trace(void *s_, const float *rays, int n, int shadow, uint32_t *tri, uint32_t *mesh, float *tuv, uint64_t *counters) {
    Scene *s = (Scene *) s_;
    TraversalCounters tc;
    s->accel.counters = counters ? &tc : 0;
    for (int i = 0; i < n; ++i) {
        float t = 0, u = 0, v = 0;
        uint32_t gid = s->accel.trace(rayFrom8(rays + 8 * (size_t) i), shadow != 0, t, u, v);
        if (gid == 0xffffffffu) { tri[i] = mesh[i] = 0xffffffffu; if (tuv) tuv[3 * i] = tuv[3 * i + 1] = tuv[3 * i + 2] = 0; }
        else { tri[i] = s->accel.triLocal[gid]; mesh[i] = s->accel.triMesh[gid]; if (tuv) { tuv[3 * i] = t; tuv[3 * i + 1] = u; tuv[3 * i + 2] = v; } }
    }
    s->accel.counters = 0;
    if (counters) { counters[0] += tc.rays; counters[1] += tc.boxTests; counters[2] += tc.triTests; }
}

/// Full Intersection record for closest hits: out n x 24 floats
/// [p(3) t uv(2) sh.s(3) sh.t(3) sh.n(3) geo.s(3) geo.t(3) geo.n(3)], valid[i] = hit.
void noracle_intersect(void *s_, const float *rays, int n, float *out, int *valid) {
    const Scene *s = (const Scene *) s_;
    for (int i = 0; i < n; ++i) {
        Hit h; float *o = out + 24 * (size_t) i;
        valid[i] = s->intersect(rayFrom8(rays + 8 * (size_t) i), h) ? 1 : 0;
        if (!valid[i]) { std::fill(o, o + 24, 0.0f); continue; }
        const float v[24] = {h.p.x, h.p.y, h.p.z, h.t, h.uv.x, h.uv.y, h.sh.s.x, h.sh.s.y, h.sh.s.z, h.sh.t.x, h.sh.t.y, h.sh.t.z,
                             h.sh.n.x, h.sh.n.y, h.sh.n.z, h.geo.s.x, h.geo.s.y, h.geo.s.z, h.geo.t.x, h.geo.t.y, h.geo.t.z,
                             h.geo.n.x, h.geo.n.y, h.geo.n.z};
        std::memcpy(o, v, sizeof(v));
    }
}

static Bsdf bsdfFrom(int type, const float *p) {
    switch (type) {
    case kDiffuse:    return makeDiffuse(mkc(p[0], p[1], p[2]));
    case kMirror:     return makeMirror();
    case kDielectric: return makeDielectric(p[4], p[5]);
    default:          return makeMicrofacet(p[3], p[4], p[5], mkc(p[0], p[1], p[2]));
    }
}

/// BSDF::sample on n (wi, xi) pairs: wo n x 3, weight n x 3, pdf n (BSDF::pdf of the sampled
/// direction; 0 for discrete), eta n, measure n.
void noracle_bsdf_sample(int type, const float *params, int n, const float *wi, const float *xi,
                         float *wo, float *weight, float *pdf, float *eta, int *measure) {
    Bsdf b = bsdfFrom(type, params);
    for (int i = 0; i < n; ++i) {
        BsdfQuery q; q.wi = mk(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]); q.wo = mk(0, 0, 0); q.eta = 1.0f; q.measure = kUnknownMeasure;
        V2 s = {xi[2 * i], xi[2 * i + 1]};
        C3 w = b.sample(q, s);
        bool ok = !isZero(w);
        wo[3 * i] = ok ? q.wo.x : 0; wo[3 * i + 1] = ok ? q.wo.y : 0; wo[3 * i + 2] = ok ? q.wo.z : 0;
        weight[3 * i] = w.r; weight[3 * i + 1] = w.g; weight[3 * i + 2] = w.b;
        pdf[i] = (ok && q.measure == kSolidAngle) ? b.pdf(q) : 0.0f;
        eta[i] = ok ? q.eta : 1.0f; measure[i] = ok ? q.measure : kUnknownMeasure;
    }
}

/// BSDF::eval and BSDF::pdf (solid-angle measure) on n (wi, wo) pairs.
void noracle_bsdf_eval_pdf(int type, const float *params, int n, const float *wi, const float *wo, float *f, float *pdf) {
    Bsdf b = bsdfFrom(type, params);
    for (int i = 0; i < n; ++i) {
        BsdfQuery q; q.wi = mk(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]); q.wo = mk(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
        q.eta = 1.0f; q.measure = kSolidAngle;
        C3 v = b.eval(q);
        f[3 * i] = v.r; f[3 * i + 1] = v.g; f[3 * i + 2] = v.b; pdf[i] = b.pdf(q);
    }
}

/// Warp::squareTo* and its pdf. ids follow src/warptest.cpp:91-94 order:
/// 0 square 1 tent 2 disk 3 uniform_sphere 4 uniform_hemisphere 5 cosine_hemisphere 6 beckmann
void noracle_warp(int id, float param, int n, const float *xi, float *out3, float *pdf) {
    for (int i = 0; i < n; ++i) {
        V2 s = {xi[2 * i], xi[2 * i + 1]}; V3 r = mk(0, 0, 0); float p = 0;
        switch (id) {
        case 0: r = mk(s.x, s.y, 0); p = (s.x >= 0 && s.x <= 1 && s.y >= 0 && s.y <= 1) ? 1.0f : 0.0f; break;
        case 1: { V2 t = squareToTent(s); r = mk(t.x, t.y, 0); p = squareToTentPdf(t); break; }
        case 2: { V2 t = squareToUniformDisk(s); r = mk(t.x, t.y, 0); p = squareToUniformDiskPdf(t); break; }
        case 3: r = squareToUniformSphere(s); p = squareToUniformSpherePdf(r); break;
        case 4: r = squareToUniformHemisphere(s); p = squareToUniformHemispherePdf(r); break;
        case 5: r = squareToCosineHemisphere(s); p = squareToCosineHemispherePdf(r); break;
        default: r = squareToBeckmann(s, param); p = squareToBeckmannPdf(r, param); break;
        }
        out3[3 * i] = r.x; out3[3 * i + 1] = r.y; out3[3 * i + 2] = r.z; pdf[i] = p;
    }
}

void noracle_fresnel(int n, const float *cosThetaI, float extIOR, float intIOR, float *out) {
    for (int i = 0; i < n; ++i) out[i] = fresnel(cosThetaI[i], extIOR, intIOR);
}

void noracle_filter_table(int type, float radius, float p0, float p1, float *table33, int *border) {
    Filter f = {type, radius, p0, p1}; FilterTable ft; ft.init(f);
    std::memcpy(table33, ft.table, sizeof(ft.table)); *border = ft.border;
}

/// BlockGenerator order: out receives 4 ints (ox, oy, w, h) per block; returns the block count.
int noracle_block_order(int W, int H, int bs, int *out, int cap) {
    std::vector<int> v; blockOrder(W, H, bs, v);
    int n = (int) v.size() / 4;
    for (int i = 0; i < std::min(n, cap) * 4; ++i) out[i] = v[i];
    return n;
}

/// ImageBlock::put over n samples in order into one full-frame block; rgbw is (H+2b)x(W+2b)x4.
uint64_t noracle_splat(int W, int H, float radius, const float *table33, int n, const float *pos, const float *rgb, float *rgbw) {
    FilterTable ft; ft.init(radius, table33);
    ImageBlock b; b.init(W, H, ft);
    for (int i = 0; i < n; ++i) { V2 p = {pos[2 * i], pos[2 * i + 1]}; b.put(p, mkc(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2])); }
    std::memcpy(rgbw, b.data.data(), b.data.size() * 4);
    return b.dropped;
}

/// Radiance of individual camera samples in kPerSampleStream mode: pixel index (y*W+x) and
/// sample index per entry -> rgb n x 3, pos n x 2 (the pixelSample passed to ImageBlock::put).
void noracle_sample_radiance(void *s_, int n, const uint32_t *pixel, const uint32_t *sample, float *rgb, float *pos, uint64_t *stats) {
    const Scene *s = (const Scene *) s_;
    PathStats st;
    for (int i = 0; i < n; ++i) {
        Sampler smp; smp.rng.seed((uint64_t) pixel[i], (uint64_t) sample[i]);
        int gx = (int) (pixel[i] % (uint32_t) s->camera.width), gy = (int) (pixel[i] / (uint32_t) s->camera.width);
        V2 r = smp.next2D(); V2 ps = {(float) gx + r.x, (float) gy + r.y};
        smp.next2D();
        C3 v = Li(*s, smp, sampleRay(s->camera, ps), s->integrator, st);
        rgb[3 * i] = v.r; rgb[3 * i + 1] = v.g; rgb[3 * i + 2] = v.b;
        if (pos) { pos[2 * i] = ps.x; pos[2 * i + 1] = ps.y; }
    }
    if (stats) { stats[0] += st.closest; stats[1] += st.shadow; stats[2] += st.segments; }
}

/// The reference's block loop (src/main.cpp:93-121) with `threads` workers pulling blocks in
/// BlockGenerator order. Per-block results are merged in generator order afterwards so the
/// image does not depend on the thread count (the reference merges in completion order).
/// rgbw: (H+2b) x (W+2b) x 4 == ImageBlock::data(). stats: {closest, shadow, segments, dropped}.
void noracle_render(void *s_, int mode, uint32_t sampleFirst, uint32_t sampleLast, int threads, float *rgbw, uint64_t *stats) {
    const Scene *s = (const Scene *) s_;
    FilterTable ft; ft.init(s->filterRadius, s->filterTable);
    int W = s->camera.width, H = s->camera.height;
    std::vector<int> order; blockOrder(W, H, kBlockSize, order);
    int nb = (int) order.size() / 4;
    std::vector<ImageBlock> blocks(nb);
    std::atomic<int> next(0);
    std::mutex mtx; PathStats total;
    RenderJob job = {mode, sampleFirst, sampleLast};
    auto worker = [&]() {
        PathStats st;
        for (;;) {
            int i = next.fetch_add(1);
            if (i >= nb) break;
            ImageBlock &b = blocks[i];
            b.init(order[4 * i + 2], order[4 * i + 3], ft);
            b.ox = order[4 * i]; b.oy = order[4 * i + 1];
            renderBlock(*s, job, b, st);
        }
        std::lock_guard<std::mutex> lk(mtx);
        total.closest += st.closest; total.shadow += st.shadow; total.segments += st.segments;
    };
    if (threads <= 0) threads = (int) std::thread::hardware_concurrency();
    if (threads <= 1) worker();
    else { std::vector<std::thread> pool; for (int t = 0; t < threads; ++t) pool.emplace_back(worker); for (auto &t : pool) t.join(); }
    ImageBlock result; result.init(W, H, ft);
    uint64_t dropped = 0;
    for (int i = 0; i < nb; ++i) { result.put(blocks[i]); dropped += blocks[i].dropped; }
    std::memcpy(rgbw, result.data.data(), result.data.size() * 4);
    if (stats) { stats[0] = total.closest; stats[1] = total.shadow; stats[2] = total.segments; stats[3] = dropped; }
}

/// Single-threaded render with traversal work counters, closest-hit and shadow rays apart:
/// counters = {rays, boxTests, triTests} x {closest, shadow}. Source of the algorithmic bytes per
/// ray of SURVEY.md 8(d) (box tests and triangle tests of the oracle's binary SAH BVH, <= 4 tris/leaf).
void noracle_render_counted(void *s_, int mode, uint32_t sampleFirst, uint32_t sampleLast, int stride, uint64_t *counters6, uint64_t *stats3) {
    Scene *s = (Scene *) s_;
    FilterTable ft; ft.init(s->filterRadius, s->filterTable);
    int W = s->camera.width, H = s->camera.height;
    std::vector<int> order; blockOrder(W, H, kBlockSize, order);
    int nb = (int) order.size() / 4;
    TraversalCounters c0, c1; PathStats st;
    s->accel.counters = &c0; s->accel.shadowCounters = &c1;
    RenderJob job = {mode, sampleFirst, sampleLast};
    for (int i = 0; i < nb; i += (stride > 0 ? stride : 1)) {
        ImageBlock b; b.init(order[4 * i + 2], order[4 * i + 3], ft);
        b.ox = order[4 * i]; b.oy = order[4 * i + 1];
        renderBlock(*s, job, b, st);
    }
    s->accel.counters = 0; s->accel.shadowCounters = 0;
    counters6[0] = c0.rays; counters6[1] = c0.boxTests; counters6[2] = c0.triTests;
    counters6[3] = c1.rays; counters6[4] = c1.boxTests; counters6[5] = c1.triTests;
    if (stats3) { stats3[0] = st.closest; stats3[1] = st.shadow; stats3[2] = st.segments; }
}

/// Path-graph hash grid + kNN (noracle_pathgraph.h). hp = {dim[3]} ints, {cellsize[3], minb[3]} floats, k, dk.
void noracle_pathgraph_knn(const float *pts, int n, const int *dim, const float *cellsize, const float *minb, int k, int dk,
                           int32_t *neighbors, int32_t *knn, int32_t *dknn, float *dists) {
    HashParams h;
    for (int c = 0; c < 3; ++c) { h.dim[c] = dim[c]; h.cellsize[c] = cellsize[c]; h.minb[c] = minb[c]; }
    h.k = k; h.dk = dk;
    pathgraphKnn(pts, n, h, neighbors, knn, dknn, dists);
}

/// BuildClusters restated (noracle_pathgraph.h). mode 0: the whole builder (index-ordered lists) -> returns the number of clusters, fills
/// group_idx (n), lists (n), offsets (capacity n + n_seeds), *n_listed. mode 1: only the nearest-seed assignment (group_idx).
/// mode 2: only one subdivision of the clustering given in group_idx / lists / offsets (n_seeds = its number of clusters, *n_listed = its
/// listed members): group_idx rewritten, returns the new number of clusters.
int noracle_pathgraph_clusters(int mode, const float *pts, int n, const int *dim, const float *cellsize, const float *minb, const int32_t *seeds, int n_seeds,
                               int K, int32_t *group_idx, int32_t *lists, int32_t *offsets, int *n_listed) {
    HashParams h;
    for (int c = 0; c < 3; ++c) { h.dim[c] = dim[c]; h.cellsize[c] = cellsize[c]; h.minb[c] = minb[c]; }
    h.k = h.dk = K;
    if (mode == 1) { pgClusterAssign(pts, n, h, seeds, n_seeds, group_idx); return n_seeds; }
    if (mode == 2) return pgClusterSubdivide(pts, n, group_idx, lists, *n_listed, offsets, n_seeds, K);
    return pathgraphBuildClusters(pts, n, h, seeds, n_seeds, K, group_idx, lists, offsets, n_listed);
}

/// Path-graph marginal densities + radiance scatter (noracle_pathgraph_scatter.h). sps: n SPoint records of 168 bytes.
void noracle_pathgraph_scatter(const void *sps, int n, const int32_t *neighbors, int row_ints, const int32_t *knn, int iterations,
                               int sequential, float *pdfmarginal, float *blur, float *mc) {
    pathgraphScatter(static_cast<const SPointRec *>(sps), n, neighbors, row_ints, knn, iterations, sequential != 0, pdfmarginal, blur, mc);
}
/// The MIS gather variant (computeMISAllOnGPURecord).
void noracle_pathgraph_mis(const void *sps, int n, const int32_t *neighbors, int row_ints, const int32_t *knn, int iterations,
                           int sequential, float *blur, float *mc) {
    pathgraphMis(static_cast<const SPointRec *>(sps), n, neighbors, row_ints, knn, iterations, sequential != 0, blur, mc);
}
/// The direct-light scatter (computeDirectScatterAllOnGPU) and, iterations >= 0, the indirect scatter that uses its result
/// (computeScatterAllOnGPURecordWithDirectOpt). lps: n LPoint records (44 bytes). marginal: n x 2 or NULL.
void noracle_pathgraph_direct(const void *sps, const void *lps, int n, const int32_t *neighbors, int row_ints, const int32_t *knn, const int32_t *dknn,
                              int iterations, int sequential, float *marginal, float *direct, float *blur, float *mc) {
    pathgraphDirect(static_cast<const SPointRec *>(sps), static_cast<const LPointRec *>(lps), n, neighbors, row_ints, dknn, sequential != 0,
                    marginal, direct);
    if (iterations >= 0)
        pathgraphScatter(static_cast<const SPointRec *>(sps), n, neighbors, row_ints, knn, iterations, sequential != 0, nullptr, blur, mc, direct);
}
/// The clustered variant (SPoints::ClusterScatterWithDirectOptNR on given clusters). direct / blur / mc: n x 3.
void noracle_pathgraph_cluster(const void *sps, const void *lps, int n, const int32_t *clusters, int n_listed, const int32_t *offsets, int n_clusters,
                               int iterations, int sequential, float *direct, float *blur, float *mc) {
    pathgraphCluster(static_cast<const SPointRec *>(sps), static_cast<const LPointRec *>(lps), n, clusters, n_listed, offsets, n_clusters, iterations,
                     sequential != 0, direct, blur, mc);
}
/// The cluster modes without direct light (order 0: ClusterScatter2, 1: ClusterScatter / loadClusterScatter). blur / mc: iterations x n x 3.
void noracle_pathgraph_cluster_scatter(int order, const void *sps, int n, const int32_t *clusters, int n_listed, const int32_t *offsets, int n_clusters,
                                       int iterations, int sequential, float *blur, float *mc) {
    pathgraphClusterScatter(static_cast<const SPointRec *>(sps), n, clusters, n_listed, offsets, n_clusters, iterations, order != 0, sequential != 0, blur, mc);
}
/// The tool's evaluations at one vertex each: pdf[i] = pdf_device(sps[i], wi[i]), bsdf[i] = bsdfeval_device(sps[i], wi[i]).
void noracle_pathgraph_eval(const void *sps, int n, const float *wi, float *pdf, float *bsdf) {
    const SPointRec *p = static_cast<const SPointRec *>(sps);
    for (int i = 0; i < n; ++i) { pdf[i] = pgs::pdfAt(p[i], wi + 3 * (size_t) i); pgs::bsdfAt(p[i], wi + 3 * (size_t) i, bsdf + 3 * (size_t) i); }
}

} // extern "C"
