// operators.cu — the reference's per-call operators as array kernels: Accel::rayIntersect,
// PerspectiveCamera::sampleRay, BSDF::sample/eval/pdf, Warp::squareTo*, pcg32, ImageBlock::put.
// They run the same device functions as the wavefront renderer (traverse.cuh, nmath.cuh,
// scene.cuh) and exist for parity tests, the chi2/t-test harness and Mrays/s measurements.
#include "engine.h"
#include "traverse_persistent.cuh"

namespace ng {

static inline unsigned gridFor(int n, int t = 256) { return (unsigned) ((n + t - 1) / t); }

/// Accel::rayIntersect (src/accel.cpp:35-55); rays: (o, mint) (d, maxt); hits: (t, u, v, gid).
template <bool SHADOW, bool COUNT>
__global__ void __launch_bounds__(256) k_trace(DeviceScene sc, const float4 *__restrict__ rays, int n, float4 *__restrict__ hits,
                                               unsigned long long *counts) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    TraceStats ts; ts.nodes = 0; ts.tris = 0;
    if (i < n) {
        const float4 ro = __ldg(rays + 2 * (size_t) i), rd = __ldg(rays + 2 * (size_t) i + 1);
        float t = 0.f, u = 0.f, v = 0.f;
        uint32_t gid = traverse<SHADOW, COUNT>(sc.nodes, sc.leafTris, mk(ro.x, ro.y, ro.z), mk(rd.x, rd.y, rd.z), ro.w, rd.w, t, u, v, &ts, sc.isectMode == NORI_ISECT_WATERTIGHT);
        if (gid == kNoHit) { t = 0.f; u = 0.f; v = 0.f; }
        hits[i] = make_float4(t, u, v, __uint_as_float(gid));
    }
    if (COUNT) {
        unsigned a = ts.nodes, b = ts.tris;
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(~0u, a, o); b += __shfl_xor_sync(~0u, b, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(counts, (unsigned long long) a); atomicAdd(counts + 1, (unsigned long long) b); }
    }
}

/// The same operator on the renderer's traversal kernel (persistent warps, dynamic ray fetch).
struct RayArrayIO {
    const float4 *rays; float4 *hits; uint32_t n;
    __device__ __forceinline__ bool load(uint32_t i, V3 &o, V3 &d, float &mint, float &maxt) const {
        NG_CHECK(i < n);
        const float4 ro = __ldg(rays + 2 * (size_t) i), rd = __ldg(rays + 2 * (size_t) i + 1);
        o = mk(ro.x, ro.y, ro.z); d = mk(rd.x, rd.y, rd.z); mint = ro.w; maxt = rd.w;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, uint32_t gid, float t, float u, float v) const {
        NG_CHECK(i < n);
        hits[i] = gid == kNoHit ? make_float4(0.f, 0.f, 0.f, __uint_as_float(gid)) : make_float4(t, u, v, __uint_as_float(gid));
    }
};
template <bool SHADOW, bool WT, int QN = 0>
__global__ void __launch_bounds__(128, QN == 2 ? 8 : (SHADOW ? 10 : 9)) k_tracePersistent(DeviceScene sc, const float4 *__restrict__ rays, int n, float4 *__restrict__ hits, unsigned int *cursor) {
    RayArrayIO io; io.rays = rays; io.hits = hits; io.n = (uint32_t) n;
    const QGrid g = {sc.qBase[0], sc.qBase[1], sc.qBase[2], sc.qExt[0], sc.qExt[1], sc.qExt[2]};
    tracePersistent<SHADOW, WT, QN>(QN == 2 ? sc.wnodes : (QN == 1 ? reinterpret_cast<const float4 *>(sc.qnodes) : sc.nodes), sc.leafTris, (uint32_t) n, cursor, io, kTraceTuning, g);
}

void launchTrace(const DeviceScene &sc, const float4 *rays, int n, bool shadow, float4 *hits, unsigned int *cursor, cudaStream_t st) {
    if (n <= 0) return;
    static unsigned grid = 0;
    if (!grid) { int dev = 0, sms = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); grid = (unsigned) sms * 10u; }   // 10 resident CTAs per SM (48 registers), as the renderer's trace kernels
    cudaMemsetAsync(cursor, 0, sizeof(unsigned int), st);
    const bool wt = sc.isectMode == NORI_ISECT_WATERTIGHT;
    if (sc.qnodes && !wt) {                                   // quantised node layout
        if (shadow) k_tracePersistent<true, false, 1><<<grid, 128, 0, st>>>(sc, rays, n, hits, cursor);
        else k_tracePersistent<false, false, 1><<<grid, 128, 0, st>>>(sc, rays, n, hits, cursor);
        return;
    }
    if (sc.wnodes && !wt) {                                   // 4-wide node layout (64 registers: 8 CTAs per SM)
        if (shadow) k_tracePersistent<true, false, 2><<<grid * 8u / 10u, 128, 0, st>>>(sc, rays, n, hits, cursor);
        else k_tracePersistent<false, false, 2><<<grid * 8u / 10u, 128, 0, st>>>(sc, rays, n, hits, cursor);
        return;
    }
    // closest hit: 9 resident CTAs per SM (56 registers), any hit: 10 (48), as the renderer's kernels
    if (shadow) { if (wt) k_tracePersistent<true, true><<<grid, 128, 0, st>>>(sc, rays, n, hits, cursor); else k_tracePersistent<true, false><<<grid, 128, 0, st>>>(sc, rays, n, hits, cursor); }
    else { if (wt) k_tracePersistent<false, true><<<grid * 9u / 10u, 128, 0, st>>>(sc, rays, n, hits, cursor); else k_tracePersistent<false, false><<<grid * 9u / 10u, 128, 0, st>>>(sc, rays, n, hits, cursor); }
}
void launchTraceCount(const DeviceScene &sc, const float4 *rays, int n, bool shadow, float4 *hits, unsigned long long *counts2, cudaStream_t st) {
    if (n <= 0) return;
    if (shadow) k_trace<true, true><<<gridFor(n), 256, 0, st>>>(sc, rays, n, hits, counts2);
    else k_trace<false, true><<<gridFor(n), 256, 0, st>>>(sc, rays, n, hits, counts2);
}

/// Closest hit + Intersection post-processing (src/accel.cpp:57-108).
__global__ void __launch_bounds__(256) k_intersect(DeviceScene sc, const float4 *__restrict__ rays, int n, float *__restrict__ out, int *__restrict__ valid) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 ro = __ldg(rays + 2 * (size_t) i), rd = __ldg(rays + 2 * (size_t) i + 1);
    float t, u, v;
    uint32_t gid = traverse<false>(sc.nodes, sc.leafTris, mk(ro.x, ro.y, ro.z), mk(rd.x, rd.y, rd.z), ro.w, rd.w, t, u, v, nullptr, sc.isectMode == NORI_ISECT_WATERTIGHT);
    float *o = out + 24 * (size_t) i;
    valid[i] = gid != kNoHit;
    if (gid == kNoHit) { for (int k = 0; k < 24; ++k) o[k] = 0.f; return; }
    Hit h; fillHit(sc, gid, t, u, v, h);
    const float vals[24] = {h.p.x, h.p.y, h.p.z, h.t, h.uv.x, h.uv.y, h.sh.s.x, h.sh.s.y, h.sh.s.z, h.sh.t.x, h.sh.t.y, h.sh.t.z,
                            h.sh.n.x, h.sh.n.y, h.sh.n.z, h.geo.s.x, h.geo.s.y, h.geo.s.z, h.geo.t.x, h.geo.t.y, h.geo.t.z,
                            h.geo.n.x, h.geo.n.y, h.geo.n.z};
    for (int k = 0; k < 24; ++k) o[k] = vals[k];
}
void launchIntersect(const DeviceScene &sc, const float4 *rays, int n, float *out24, int *valid, cudaStream_t st) {
    if (n > 0) k_intersect<<<gridFor(n), 256, 0, st>>>(sc, rays, n, out24, valid);
}

__global__ void k_cameraRays(DeviceScene sc, const float2 *__restrict__ ps, int n, float4 *__restrict__ rays) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V2 p; p.x = ps[i].x; p.y = ps[i].y;
    Ray r = sampleRay(sc.cam, p);
    rays[2 * (size_t) i] = make_float4(r.o.x, r.o.y, r.o.z, r.mint);
    rays[2 * (size_t) i + 1] = make_float4(r.d.x, r.d.y, r.d.z, r.maxt);
}
void launchCameraRays(const DeviceScene &sc, const float2 *ps, int n, float4 *rays, cudaStream_t st) {
    if (n > 0) k_cameraRays<<<gridFor(n), 256, 0, st>>>(sc, ps, n, rays);
}

__global__ void k_bsdfSample(Bsdf b, int n, const float *__restrict__ wi, const float *__restrict__ xi, float *wo, float *weight, float *pdf, float *eta, int *measure) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BsdfQuery q; q.wi = mk(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]); q.wo = mk(0, 0, 0); q.eta = 1.0f; q.measure = kUnknownMeasure;
    V2 s; s.x = xi[2 * i]; s.y = xi[2 * i + 1];
    V3 w = b.sample(q, s);
    bool ok = !isZero(w);
    wo[3 * i] = ok ? q.wo.x : 0; wo[3 * i + 1] = ok ? q.wo.y : 0; wo[3 * i + 2] = ok ? q.wo.z : 0;
    weight[3 * i] = w.x; weight[3 * i + 1] = w.y; weight[3 * i + 2] = w.z;
    pdf[i] = (ok && q.measure == kSolidAngle) ? b.pdf(q) : 0.0f;
    eta[i] = ok ? q.eta : 1.0f; measure[i] = ok ? q.measure : kUnknownMeasure;
}
void launchBsdfSample(Bsdf b, int n, const float *wi, const float *xi, float *wo, float *weight, float *pdf, float *eta, int *measure, cudaStream_t st) {
    if (n > 0) k_bsdfSample<<<gridFor(n), 256, 0, st>>>(b, n, wi, xi, wo, weight, pdf, eta, measure);
}

__global__ void k_bsdfEvalPdf(Bsdf b, int n, const float *__restrict__ wi, const float *__restrict__ wo, float *f, float *pdf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BsdfQuery q; q.wi = mk(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]); q.wo = mk(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    q.eta = 1.0f; q.measure = kSolidAngle;
    V3 v = b.eval(q);
    f[3 * i] = v.x; f[3 * i + 1] = v.y; f[3 * i + 2] = v.z; pdf[i] = b.pdf(q);
}
void launchBsdfEvalPdf(Bsdf b, int n, const float *wi, const float *wo, float *f, float *pdf, cudaStream_t st) {
    if (n > 0) k_bsdfEvalPdf<<<gridFor(n), 256, 0, st>>>(b, n, wi, wo, f, pdf);
}

__global__ void k_warp(int id, float param, int n, const float *__restrict__ xi, float *out3, float *pdf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V2 s; s.x = xi[2 * i]; s.y = xi[2 * i + 1];
    V3 r = mk(0, 0, 0); float p = 0;
    switch (id) {
    case 0: r = mk(s.x, s.y, 0); p = (s.x >= 0 && s.x <= 1 && s.y >= 0 && s.y <= 1) ? 1.0f : 0.0f; break;   // warp.cpp:25-31
    case 1: { V2 t = squareToTent(s); r = mk(t.x, t.y, 0); p = squareToTentPdf(t); break; }
    case 2: { V2 t = squareToUniformDisk(s); r = mk(t.x, t.y, 0); p = squareToUniformDiskPdf(t); break; }
    case 3: r = squareToUniformSphere(s); p = squareToUniformSpherePdf(r); break;
    case 4: r = squareToUniformHemisphere(s); p = squareToUniformHemispherePdf(r); break;
    case 5: r = squareToCosineHemisphere(s); p = squareToCosineHemispherePdf(r); break;
    default: r = squareToBeckmann(s, param); p = squareToBeckmannPdf(r, param); break;
    }
    out3[3 * i] = r.x; out3[3 * i + 1] = r.y; out3[3 * i + 2] = r.z; pdf[i] = p;
}
void launchWarp(int id, float param, int n, const float *xi, float *out3, float *pdf, cudaStream_t st) {
    if (n > 0) k_warp<<<gridFor(n), 256, 0, st>>>(id, param, n, xi, out3, pdf);
}

/// Warp::squareTo*Pdf (include/nori/warp.h:33-69) at given points (2D warps read x, y only).
__global__ void k_warpPdf(int id, float param, int n, const float *__restrict__ pts3, float *pdf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 v = mk(pts3[3 * i], pts3[3 * i + 1], pts3[3 * i + 2]);
    V2 q; q.x = v.x; q.y = v.y;
    float p;
    switch (id) {
    case 0: p = (q.x >= 0 && q.x <= 1 && q.y >= 0 && q.y <= 1) ? 1.0f : 0.0f; break;
    case 1: p = squareToTentPdf(q); break;
    case 2: p = squareToUniformDiskPdf(q); break;
    case 3: p = squareToUniformSpherePdf(v); break;
    case 4: p = squareToUniformHemispherePdf(v); break;
    case 5: p = squareToCosineHemispherePdf(v); break;
    default: p = squareToBeckmannPdf(v, param); break;
    }
    pdf[i] = p;
}
void launchWarpPdf(int id, float param, int n, const float *pts3, float *pdf, cudaStream_t st) {
    if (n > 0) k_warpPdf<<<gridFor(n), 256, 0, st>>>(id, param, n, pts3, pdf);
}

/// Thread i produces the (advance + i)-th output of the stream through pcg32::advance.
__global__ void k_pcg32(uint64_t initstate, uint64_t initseq, int64_t advance, int n, uint32_t *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Pcg32 r; r.seed(initstate, initseq);
    r.advance(advance + i);
    out[i] = r.nextUInt();
}
void launchPcg32(uint64_t initstate, uint64_t initseq, int64_t advance, int n, uint32_t *out, cudaStream_t st) {
    if (n > 0) k_pcg32<<<gridFor(n), 256, 0, st>>>(initstate, initseq, advance, n, out);
}

/// ImageBlock::put(pos, value) (src/block.cpp:74-103) into a full-frame block, one thread per
/// sample, global float atomics (order not reproducible).
__global__ void k_splatList(DeviceScene sc, int n, const float2 *__restrict__ pos, const float *__restrict__ rgb, float *block, unsigned long long *dropped) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = sc.filter.border, cols = sc.cam.width + 2 * b, rows = sc.cam.height + 2 * b;
    const float radius = sc.filter.radius, lookup = sc.filter.lookupFactor;
    V3 val = mk(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
    if (!isValidColor(val)) { atomicAdd(dropped, 1ull); return; }
    const float px = pos[i].x - 0.5f - (float) (0 - b), py = pos[i].y - 0.5f - (float) (0 - b);
    const int x0 = max((int) ceilf(px - radius), 0), y0 = max((int) ceilf(py - radius), 0);
    const int x1 = min((int) floorf(px + radius), cols - 1), y1 = min((int) floorf(py + radius), rows - 1);
    for (int y = y0; y <= y1; ++y) {
        float wy = sc.filter.table[(int) (fabsf((float) y - py) * lookup)];
        for (int x = x0; x <= x1; ++x) {
            float wx = sc.filter.table[(int) (fabsf((float) x - px) * lookup)];
            NG_CHECK((int) (fabsf((float) y - py) * lookup) <= 32 && (int) (fabsf((float) x - px) * lookup) <= 32 && y < rows && x < cols);
            float *p = block + ((size_t) y * cols + x) * 4;
            atomicAdd(p, val.x * wx * wy); atomicAdd(p + 1, val.y * wx * wy); atomicAdd(p + 2, val.z * wx * wy); atomicAdd(p + 3, 1.0f * wx * wy);
        }
    }
}
void launchSplatList(const DeviceScene &sc, int n, const float2 *pos, const float *rgb, float *block, unsigned long long *dropped, cudaStream_t st) {
    if (n > 0) k_splatList<<<gridFor(n), 256, 0, st>>>(sc, n, pos, rgb, block, dropped);
}

/// The same operator with the reference's summation order: one thread per accumulator cell
/// walks the samples in list order, so every cell sees `+=` in exactly the order
/// ImageBlock::put would apply them (bit-reproducible; O(cells x n), for parity checks).
__global__ void k_splatOrdered(DeviceScene sc, int n, const float2 *__restrict__ pos, const float *__restrict__ rgb, float *block, unsigned long long *dropped) {
    const int b = sc.filter.border, cols = sc.cam.width + 2 * b, rows = sc.cam.height + 2 * b;
    int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= cols * rows) return;
    const int cx = cell % cols, cy = cell / cols;
    const float radius = sc.filter.radius, lookup = sc.filter.lookupFactor;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    unsigned long long drop = 0;
    for (int i = 0; i < n; ++i) {
        V3 val = mk(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
        if (!isValidColor(val)) { ++drop; continue; }
        const float px = pos[i].x - 0.5f - (float) (0 - b), py = pos[i].y - 0.5f - (float) (0 - b);
        const int x0 = max((int) ceilf(px - radius), 0), y0 = max((int) ceilf(py - radius), 0);
        const int x1 = min((int) floorf(px + radius), cols - 1), y1 = min((int) floorf(py + radius), rows - 1);
        if (cx < x0 || cx > x1 || cy < y0 || cy > y1) continue;
        float wx = sc.filter.table[(int) (fabsf((float) cx - px) * lookup)], wy = sc.filter.table[(int) (fabsf((float) cy - py) * lookup)];
        acc[0] += val.x * wx * wy; acc[1] += val.y * wx * wy; acc[2] += val.z * wx * wy; acc[3] += 1.0f * wx * wy;
    }
    float *p = block + (size_t) cell * 4;
    p[0] += acc[0]; p[1] += acc[1]; p[2] += acc[2]; p[3] += acc[3];
    if (cell == 0 && drop) atomicAdd(dropped, drop);
}
void launchSplatOrdered(const DeviceScene &sc, int n, const float2 *pos, const float *rgb, float *block, unsigned long long *dropped, cudaStream_t st) {
    const int b = sc.filter.border, cells = (sc.cam.width + 2 * b) * (sc.cam.height + 2 * b);
    if (n > 0) k_splatOrdered<<<gridFor(cells), 256, 0, st>>>(sc, n, pos, rgb, block, dropped);
}

/// Scene flattening on the device: the Mesh buffers as the reference keeps them (m_V / m_N 3 x nV, m_UV 2 x nV,
/// m_F 3 x nF, column-major, include/nori/mesh.h:172-175) arrive as raw copies; these two kernels write the
/// padded, concatenated arrays the traversal and shading kernels read (16-byte vertex records, global indices).
__global__ void k_packVerts(const float *__restrict__ V, const float *__restrict__ N, const float *__restrict__ UV, uint32_t nV,
                            float4 *__restrict__ vpos, float4 *__restrict__ vnrm, float2 *__restrict__ vuv) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nV) return;
    vpos[v] = make_float4(V[3 * (size_t) v], V[3 * (size_t) v + 1], V[3 * (size_t) v + 2], 0.f);
    vnrm[v] = N ? make_float4(N[3 * (size_t) v], N[3 * (size_t) v + 1], N[3 * (size_t) v + 2], 0.f) : make_float4(0.f, 0.f, 0.f, 0.f);
    vuv[v] = UV ? make_float2(UV[2 * (size_t) v], UV[2 * (size_t) v + 1]) : make_float2(0.f, 0.f);
}
__global__ void k_packTris(const uint32_t *__restrict__ F, uint32_t nF, uint32_t nV, uint32_t vertOffset, uint32_t mesh,
                           uint4 *__restrict__ tris, unsigned int *badIndex) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nF) return;
    const uint32_t a = F[3 * (size_t) f], b = F[3 * (size_t) f + 1], c = F[3 * (size_t) f + 2];
    if (a >= nV || b >= nV || c >= nV) { atomicAdd(badIndex, 1u); tris[f] = make_uint4(vertOffset, vertOffset, vertOffset, mesh); return; }
    tris[f] = make_uint4(vertOffset + a, vertOffset + b, vertOffset + c, mesh);
}
void launchPackMesh(const float *V, const float *N, const float *UV, const uint32_t *F, uint32_t nV, uint32_t nF, uint32_t vertOffset, uint32_t mesh,
                    float4 *vpos, float4 *vnrm, float2 *vuv, uint4 *tris, unsigned int *badIndex, cudaStream_t st) {
    if (nV) k_packVerts<<<gridFor((int) nV), 256, 0, st>>>(V, N, UV, nV, vpos, vnrm, vuv);
    if (nF) k_packTris<<<gridFor((int) nF), 256, 0, st>>>(F, nF, nV, vertOffset, mesh, tris, badIndex);
}

/// ImageBlock::toBitmap (src/block.cpp:57-63; Color4f::divideByFilterWeight, color.h:112-117)
__global__ void k_toBitmap(const float4 *__restrict__ block, int W, int H, int border, float *__restrict__ rgb) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    int x = i % W, y = i / W;
    float4 c = block[(size_t) (y + border) * (W + 2 * border) + x + border];
    float r = 0.f, g = 0.f, b = 0.f;
    if (c.w != 0.f) { r = c.x / c.w; g = c.y / c.w; b = c.z / c.w; }
    rgb[3 * (size_t) i] = r; rgb[3 * (size_t) i + 1] = g; rgb[3 * (size_t) i + 2] = b;
}
void launchToBitmap(const float *block, int W, int H, int border, float *rgb, cudaStream_t st) {
    k_toBitmap<<<gridFor(W * H), 256, 0, st>>>(reinterpret_cast<const float4 *>(block), W, H, border, rgb);
}

/// toBitmap followed by the PNG path of Bitmap::savePNG (src/bitmap.cpp:117-127): Color3f::toSRGB
/// (src/common.cpp:178-192), scale by 255, clamp to [0, 255], truncate.
__device__ __forceinline__ float toSRGB1(float v) { return v <= 0.0031308f ? 12.92f * v : (1.0f + 0.055f) * powf(v, 1.0f / 2.4f) - 0.055f; }
__global__ void k_toSRGB8(const float4 *__restrict__ block, int W, int H, int border, uint8_t *__restrict__ rgb8) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    int x = i % W, y = i / W;
    float4 c = block[(size_t) (y + border) * (W + 2 * border) + x + border];
    float v[3] = {0.f, 0.f, 0.f};
    if (c.w != 0.f) { v[0] = c.x / c.w; v[1] = c.y / c.w; v[2] = c.z / c.w; }
#pragma unroll
    for (int k = 0; k < 3; ++k) rgb8[3 * (size_t) i + k] = (uint8_t) fminf(fmaxf(255.f * toSRGB1(v[k]), 0.f), 255.f);
}
void launchToSRGB8(const float *block, int W, int H, int border, uint8_t *rgb8, cudaStream_t st) {
    k_toSRGB8<<<gridFor(W * H), 256, 0, st>>>(reinterpret_cast<const float4 *>(block), W, H, border, rgb8);
}

/// ImageBlock::put(ImageBlock &) (src/block.cpp:105-114) for two full-frame blocks: dst += src.
__global__ void k_addBlock(float4 *__restrict__ dst, const float4 *__restrict__ src, size_t n) {
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = dst[i], b = src[i];
    dst[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
void launchAddBlock(float *dst, const float *src, size_t nFloats, cudaStream_t st) {
    const size_t n = nFloats / 4;
    if (n) k_addBlock<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(reinterpret_cast<float4 *>(dst), reinterpret_cast<const float4 *>(src), n);
}

unsigned operatorChecks(unsigned *line) { return fetchChecks(line); }

} // namespace ng
