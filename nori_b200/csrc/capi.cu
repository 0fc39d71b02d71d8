// capi.cu — the C ABI of include/nori_gpu.h: context, scene upload, render loop, operators.
// No CPU path exists here: every entry point either runs CUDA kernels or fails.
#include "engine.h"
#include "traverse_persistent.cuh"
#include "pathgraph.h"
#include <vector>
#include <string>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <algorithm>
#include <map>
#include <mutex>
#include <thread>
#include <dlfcn.h>

using namespace ng;

static thread_local std::string g_createError;

struct nori_gpu_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    cudaDeviceProp prop;
    // scene
    std::vector<MeshInfo> meshes;
    std::vector<EmitterInfo> emitters;
    float4 *d_vpos = nullptr, *d_vnrm = nullptr; float2 *d_vuv = nullptr; uint4 *d_tris = nullptr;
    MeshInfo *d_meshes = nullptr; EmitterInfo *d_emitters = nullptr; float *d_cdf = nullptr;
    uint32_t nTris = 0, nVerts = 0;
    BvhDevice bvh; BvhParams bvhParams; bool haveBvh = false, haveCamera = false, haveFilter = false;
    Camera cam; FilterTab filter;
    // accumulator
    float *d_block = nullptr; size_t blockFloats = 0;
    // wavefront pool
    Pool pool; bool havePool = false; uint32_t poolSize = 1u << 23;   // 8 Mi path slots (sweep: profiles/r01_pipeline.md)
    bool poolAuto = true;                       // no nori_gpu_set_pool_size yet: renders of >= 3 * 2^26 (201 M) camera samples take 16 Mi slots, short ones a third of their samples (2 Mi .. 8 Mi): profiles/r02_notes.md
    Counters *d_counters = nullptr; Counters *h_counters = nullptr;   // pinned mirrors: two (one per batch in flight)
    // host -> device staging for large pageable arrays (stagedCopy): kUploadWorkers host threads x 2 pinned buffers each (sweep on the GPU box, C4 240 MB: 4 x 8 MB 19.3 ms, 8 x 4 MB 13.4 ms, 8 x 8 15.8, 12 x 4 14.4, 8 x 2 15.1, 16 x 2 18.3; plain pageable copies 26 ms)
#ifndef NORI_UPLOAD_WORKERS
#define NORI_UPLOAD_WORKERS 8
#endif
#ifndef NORI_UPLOAD_CHUNK_MB
#define NORI_UPLOAD_CHUNK_MB 4
#endif
    static constexpr int kUploadWorkers = NORI_UPLOAD_WORKERS; static constexpr size_t kUploadChunk = (size_t) NORI_UPLOAD_CHUNK_MB << 20;
    void *upBuf[kUploadWorkers][2] = {}; cudaEvent_t upEv[kUploadWorkers][2] = {}; cudaStream_t upStream[kUploadWorkers] = {};
    cudaEvent_t upDone[kUploadWorkers] = {}, upGate = nullptr; bool haveUpload = false;
    float4 *recA = nullptr; float *recB = nullptr; size_t recCap = 0;
    uint32_t *tileDone = nullptr, *tileGen = nullptr, *tileArrive = nullptr, *readyTiles = nullptr; size_t tileDoneCap = 0;
    bool profiling = false;
    bool plainKernels = true;                   // scenes without microfacet surfaces render with the shade kernel that carries no Beckmann code (nori_gpu_set_plain_bsdf_kernels)
    bool rayBins = true;                        // bounce rays binned by origin cell + direction signs between shade and the next closest-hit pass (nori_gpu_set_ray_binning)
    int accelLayout = NORI_ACCEL_FLOAT;         // node layout the persistent traversal kernels read
    int isectMode = NORI_ISECT_REFERENCE_MT;   // used by the raw operators; nori_gpu_render takes it from its descriptor
};

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { ctx->err = std::string(#call) + " failed: " + cudaGetErrorString(e_); return NORI_GPU_ECUDA; } } while (0)
#define REQUIRE(cond, msg) do { if (!(cond)) { ctx->err = (msg); return NORI_GPU_EINVAL; } } while (0)

/// Checked build only (csrc/check.cuh): turns range-check violations counted by the kernels into an error.
static int reportChecks(nori_gpu_ctx *ctx, const char *what) {
    unsigned lineR = 0, lineO = 0;
    const unsigned nR = renderChecks(&lineR), nO = operatorChecks(&lineO);
    if (!nR && !nO) return NORI_GPU_OK;
    ctx->err = std::string(what) + ": device range check failed (" + std::to_string(nR) + " at render.cu/traverse:" + std::to_string(lineR) +
               ", " + std::to_string(nO) + " at operators.cu/traverse:" + std::to_string(lineO) + ")";
    return NORI_GPU_ECUDA;
}

namespace {

template <class T> struct DevBuf {          // scoped device scratch
    T *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc((void **) &p, std::max<size_t>(n, 1) * sizeof(T)); }
};

DeviceScene deviceScene(const nori_gpu_ctx *c) {
    DeviceScene s;
    memset(&s, 0, sizeof(s));
    s.vpos = c->d_vpos; s.vnrm = c->d_vnrm; s.vuv = c->d_vuv; s.tris = c->d_tris;
    s.meshes = c->d_meshes; s.emitters = c->d_emitters; s.emitterCdf = c->d_cdf;
    s.nMeshes = (int) c->meshes.size(); s.nEmitters = (int) c->emitters.size();
    s.nTris = c->nTris; s.nVerts = c->nVerts;
    s.nodes = c->bvh.nodes; s.leafTris = c->bvh.leafTris; s.nNodes = c->bvh.nNodes;
    s.cam = c->cam; s.filter = c->filter; s.isectMode = c->isectMode;
    s.plainBsdfs = c->plainKernels && !c->meshes.empty();
    for (const MeshInfo &m : c->meshes) if (m.bsdfType == NORI_BSDF_MICROFACET) s.plainBsdfs = 0;
    s.qnodes = c->accelLayout == NORI_ACCEL_QUANTISED ? c->bvh.qnodes : nullptr;
    s.wnodes = c->accelLayout == NORI_ACCEL_WIDE4 ? c->bvh.wnodes : nullptr;
    for (int k = 0; k < 3; ++k) { s.qBase[k] = c->bvh.qBase[k]; s.qExt[k] = c->bvh.qExt[k]; }
    return s;
}

void freeScene(nori_gpu_ctx *c) {
    cudaFree(c->d_vpos); cudaFree(c->d_vnrm); cudaFree(c->d_vuv); cudaFree(c->d_tris);
    cudaFree(c->d_meshes); cudaFree(c->d_emitters); cudaFree(c->d_cdf);
    cudaFree(c->bvh.nodes); cudaFree(c->bvh.leafTris); cudaFree(c->bvh.qnodes); cudaFree(c->bvh.wnodes);
    c->d_vpos = c->d_vnrm = nullptr; c->d_vuv = nullptr; c->d_tris = nullptr; c->d_meshes = nullptr; c->d_emitters = nullptr; c->d_cdf = nullptr;
    c->bvh = BvhDevice(); c->haveBvh = false; c->meshes.clear(); c->emitters.clear(); c->nTris = c->nVerts = 0;
}

void freePool(nori_gpu_ctx *c) {
    if (!c->havePool) return;
    Pool &p = c->pool;
    cudaFree(p.rayO); cudaFree(p.rayD); cudaFree(p.hit); cudaFree(p.thr); cudaFree(p.rad); cudaFree(p.state); cudaFree(p.info);
    cudaFree(p.pos); cudaFree(p.shadow);
    cudaFree(p.key); cudaFree(p.order); cudaFree(p.sortTmp);
    c->havePool = false;
}

int ensurePool(nori_gpu_ctx *ctx) {
    // the sort word keeps the slot number in 24 bits: pools beyond 16 Mi slots render without binning
    const bool bins = ctx->rayBins && ctx->poolSize <= kBinSlotMask + 1u;
    if (ctx->havePool && ctx->pool.capacity == ctx->poolSize && (ctx->pool.order != nullptr) == bins) return 0;
    freePool(ctx);
    Pool &p = ctx->pool; const size_t n = ctx->poolSize;
    memset(&p, 0, sizeof(p));
    p.capacity = ctx->poolSize;
    CK(cudaMalloc((void **) &p.rayO, n * 16)); CK(cudaMalloc((void **) &p.rayD, n * 16)); CK(cudaMalloc((void **) &p.hit, n * 16));
    CK(cudaMalloc((void **) &p.thr, n * 16)); CK(cudaMalloc((void **) &p.rad, n * 16)); CK(cudaMalloc((void **) &p.state, n * 16));
    CK(cudaMalloc((void **) &p.info, n * 16)); CK(cudaMalloc((void **) &p.pos, n * 8));
    CK(cudaMalloc((void **) &p.shadow, n * 48));
    if (bins) {
        CK(cudaMalloc((void **) &p.key, n * 4)); CK(cudaMalloc((void **) &p.order, n * 4));
        p.sortTmpBytes = binSortTmpBytes(ctx->poolSize);
        CK(cudaMalloc(&p.sortTmp, p.sortTmpBytes));
    }
    ctx->havePool = true;
    return 0;
}

/// Host -> device copy of a large PAGEABLE array at pinned-memory speed. A plain cudaMemcpyAsync from pageable memory goes
/// through one driver staging buffer at ~9 GB/s (C4's 240 MB took 26 ms). Here kUploadWorkers host threads each copy
/// their 8 MB chunks into two pinned buffers of their own and queue the DMA on their own stream, so that the CPU copies of
/// several chunks and the DMA of the previous ones overlap; ctx->stream then waits for all of them. Small arrays take the plain path.
int stagedCopy(nori_gpu_ctx *ctx, void *dst, const void *src, size_t bytes) {
    constexpr int W = nori_gpu_ctx::kUploadWorkers; constexpr size_t C = nori_gpu_ctx::kUploadChunk;
    if (bytes < 4 * C) { CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream)); return 0; }
    if (!ctx->haveUpload) {
        for (int w = 0; w < W; ++w) {
            CK(cudaStreamCreateWithFlags(&ctx->upStream[w], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&ctx->upDone[w], cudaEventDisableTiming));
            for (int b = 0; b < 2; ++b) { CK(cudaHostAlloc(&ctx->upBuf[w][b], C, cudaHostAllocDefault)); CK(cudaEventCreateWithFlags(&ctx->upEv[w][b], cudaEventDisableTiming)); }
        }
        CK(cudaEventCreateWithFlags(&ctx->upGate, cudaEventDisableTiming));
        ctx->haveUpload = true;
    }
    // the staging streams must not run ahead of work already queued on ctx->stream that still reads or writes dst
    CK(cudaEventRecord(ctx->upGate, ctx->stream));
    for (int w = 0; w < W; ++w) CK(cudaStreamWaitEvent(ctx->upStream[w], ctx->upGate, 0));
    const size_t nChunks = (bytes + C - 1) / C;
    cudaError_t errs[W];
    std::thread workers[W];
    for (int w = 0; w < W; ++w)
        workers[w] = std::thread([=, &errs]() {
            cudaError_t e = cudaSetDevice(ctx->device);
            int turn = 0;
            for (size_t c = (size_t) w; c < nChunks && e == cudaSuccess; c += W, ++turn) {
                const int b = turn & 1;
                const size_t off = c * C, n = std::min(C, bytes - off);
                e = cudaEventSynchronize(ctx->upEv[w][b]);        // the DMA that last read this buffer (this call or an earlier one) is done
                if (e != cudaSuccess) break;
                memcpy(ctx->upBuf[w][b], static_cast<const unsigned char *>(src) + off, n);
                e = cudaMemcpyAsync(static_cast<unsigned char *>(dst) + off, ctx->upBuf[w][b], n, cudaMemcpyHostToDevice, ctx->upStream[w]);
                if (e == cudaSuccess) e = cudaEventRecord(ctx->upEv[w][b], ctx->upStream[w]);
            }
            errs[w] = e;
        });
    for (int w = 0; w < W; ++w) workers[w].join();
    for (int w = 0; w < W; ++w) CK(errs[w]);
    for (int w = 0; w < W; ++w) {                 // everything queued behind this call on ctx->stream sees the data
        CK(cudaEventRecord(ctx->upDone[w], ctx->upStream[w]));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->upDone[w], 0));
    }
    return 0;
}

int ensureBlock(nori_gpu_ctx *ctx) {
    size_t need = (size_t) (ctx->cam.width + 2 * ctx->filter.border) * (ctx->cam.height + 2 * ctx->filter.border) * 4;
    if (ctx->d_block && ctx->blockFloats == need) return 0;
    cudaFree(ctx->d_block); ctx->d_block = nullptr;
    CK(cudaMalloc((void **) &ctx->d_block, need * sizeof(float)));
    CK(cudaMemsetAsync(ctx->d_block, 0, need * sizeof(float), ctx->stream));
    ctx->blockFloats = need;
    return 0;
}

} // namespace

extern "C" {

int nori_gpu_create(nori_gpu_ctx **out, int device_id) {
    if (!out) return NORI_GPU_EINVAL;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_createError = std::string("no CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") + "): libnori_b200 has no CPU path";
        return NORI_GPU_ENODEVICE;
    }
    if (device_id < 0 || device_id >= n) { g_createError = "device id out of range"; return NORI_GPU_EINVAL; }
    if ((e = cudaSetDevice(device_id)) != cudaSuccess) { g_createError = cudaGetErrorString(e); return NORI_GPU_ECUDA; }
    nori_gpu_ctx *ctx = new nori_gpu_ctx();
    ctx->device = device_id;
    cudaGetDeviceProperties(&ctx->prop, device_id);
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaMalloc((void **) &ctx->d_counters, sizeof(Counters))) != cudaSuccess ||
        (e = cudaMallocHost((void **) &ctx->h_counters, 2 * sizeof(Counters))) != cudaSuccess) {
        g_createError = cudaGetErrorString(e); delete ctx; return NORI_GPU_ECUDA;
    }
    memset(&ctx->cam, 0, sizeof(ctx->cam)); memset(&ctx->filter, 0, sizeof(ctx->filter)); memset(&ctx->pool, 0, sizeof(ctx->pool));
#ifdef NORI_GPU_SWEEP   // parameter sweeps (build_variant): the default build reads no tuning from the environment
    if (const char *s = getenv("NORI_GPU_POOL")) { ctx->poolSize = std::max(256u, (uint32_t) strtoul(s, nullptr, 10) / 256u * 256u); ctx->poolAuto = false; }
    if (const char *s = getenv("NORI_GPU_PLOC_RADIUS")) ctx->bvhParams.plocRadius = atoi(s);
    if (const char *s = getenv("NORI_GPU_MAX_LEAF")) ctx->bvhParams.maxLeaf = atoi(s);
    if (const char *s = getenv("NORI_GPU_ROTATIONS")) ctx->bvhParams.rotationPasses = atoi(s);
    if (const char *s = getenv("NORI_GPU_BINS")) ctx->rayBins = atoi(s) != 0;
    if (const char *s = getenv("NORI_GPU_PLAIN")) ctx->plainKernels = atoi(s) != 0;
    if (const char *s = getenv("NORI_GPU_SAH_CI")) ctx->bvhParams.costInner = (float) atof(s);
    if (const char *s = getenv("NORI_GPU_SAH_CT")) ctx->bvhParams.costTri = (float) atof(s);
    if (const char *e = getenv("NORI_GPU_ACCEL_LAYOUT")) ctx->accelLayout = atoi(e) == 1 ? NORI_ACCEL_QUANTISED : (atoi(e) == 2 ? NORI_ACCEL_WIDE4 : NORI_ACCEL_FLOAT);
#endif
    *out = ctx;
    return NORI_GPU_OK;
}

void nori_gpu_destroy(nori_gpu_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    freeScene(ctx); freePool(ctx);
    cudaFree(ctx->d_block); cudaFree(ctx->d_counters); cudaFreeHost(ctx->h_counters);
    if (ctx->haveUpload)
        for (int w = 0; w < nori_gpu_ctx::kUploadWorkers; ++w) {
            for (int b = 0; b < 2; ++b) { cudaFreeHost(ctx->upBuf[w][b]); cudaEventDestroy(ctx->upEv[w][b]); }
            cudaStreamDestroy(ctx->upStream[w]); cudaEventDestroy(ctx->upDone[w]);
        }
    if (ctx->haveUpload) cudaEventDestroy(ctx->upGate);
    cudaFree(ctx->recA); cudaFree(ctx->recB); cudaFree(ctx->tileDone); cudaFree(ctx->tileGen); cudaFree(ctx->tileArrive); cudaFree(ctx->readyTiles);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *nori_gpu_last_error(const nori_gpu_ctx *ctx) { return ctx ? ctx->err.c_str() : g_createError.c_str(); }

int nori_gpu_device_name(const nori_gpu_ctx *ctx, char *buf, int cap) {
    if (!ctx || !buf || cap <= 0) return NORI_GPU_EINVAL;
    snprintf(buf, (size_t) cap, "%s (sm_%d%d, %d SMs)", ctx->prop.name, ctx->prop.major, ctx->prop.minor, ctx->prop.multiProcessorCount);
    return NORI_GPU_OK;
}

int nori_gpu_set_profiling(nori_gpu_ctx *ctx, int on) { if (!ctx) return NORI_GPU_EINVAL; ctx->profiling = on != 0; return NORI_GPU_OK; }
int nori_gpu_set_pool_size(nori_gpu_ctx *ctx, uint32_t n) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(n >= 256, "pool size must be >= 256");
    ctx->poolSize = n / 256u * 256u; ctx->poolAuto = false;
    return NORI_GPU_OK;
}

/// Scene::addChild(EMesh) + Accel::addMesh for every mesh (src/scene.cpp:58-66, accel.cpp:24-29,
/// without the reference's single-mesh limit), Mesh::addChild(BSDF/Emitter) (mesh.cpp:104-126).
static int uploadMeshes(nori_gpu_ctx *ctx, const NoriGpuMeshDesc *meshes, int n) {
    REQUIRE(meshes && n > 0, "upload_meshes: no meshes");
    CK(cudaSetDevice(ctx->device));
    freeScene(ctx);
    // Every descriptor is validated before anything is allocated (the face indices of non-emitter meshes are
    // range-checked on the device while they are packed; that path frees the scene as well).
    size_t nV = 0, nF = 0;
    for (int i = 0; i < n; ++i) {
        REQUIRE(meshes[i].V && meshes[i].F && meshes[i].nV > 0, "upload_meshes: mesh without vertices or faces");
        REQUIRE(meshes[i].bsdf_type >= 0 && meshes[i].bsdf_type <= 3, "upload_meshes: unknown BSDF type");
        if (meshes[i].is_emitter) {
            REQUIRE(meshes[i].nF > 0, "upload_meshes: emitter mesh without triangles");
            for (size_t k = 0; k < 3 * (size_t) meshes[i].nF; ++k) REQUIRE(meshes[i].F[k] < meshes[i].nV, "upload_meshes: face index out of range");
        }
        nV += meshes[i].nV; nF += meshes[i].nF;
    }
    REQUIRE(nF > 0 && nF < (1u << 28), "upload_meshes: triangle count out of range");
    // The Mesh buffers cross the bus as they are (12 + 12 + 8 B per vertex, 12 B per face) and are padded,
    // concatenated and range-checked on the device (k_packVerts / k_packTris); only the per-mesh records and
    // the emitters' area tables are put together here.
    ctx->nTris = (uint32_t) nF; ctx->nVerts = (uint32_t) nV;
    CK(cudaMalloc((void **) &ctx->d_vpos, nV * sizeof(float4))); CK(cudaMalloc((void **) &ctx->d_vnrm, nV * sizeof(float4)));
    CK(cudaMalloc((void **) &ctx->d_vuv, nV * sizeof(float2))); CK(cudaMalloc((void **) &ctx->d_tris, nF * sizeof(uint4)));
    size_t stageFloats = 0, stageIdx = 0;
    for (int i = 0; i < n; ++i) { stageFloats = std::max(stageFloats, (size_t) meshes[i].nV * 8); stageIdx = std::max(stageIdx, (size_t) meshes[i].nF * 3); }
    DevBuf<float> stageV[2]; DevBuf<uint32_t> stageF[2]; DevBuf<unsigned int> bad;
    for (int k = 0; k < (n > 1 ? 2 : 1); ++k) { CK(stageV[k].alloc(std::max<size_t>(1, stageFloats))); CK(stageF[k].alloc(std::max<size_t>(1, stageIdx))); }
    CK(bad.alloc(1)); CK(cudaMemsetAsync(bad.p, 0, sizeof(unsigned int), ctx->stream));
    struct StageEvents { cudaEvent_t e[2] = {nullptr, nullptr}; ~StageEvents() { for (cudaEvent_t x : e) if (x) cudaEventDestroy(x); } } stageEvents;   // freed on every return path
    cudaEvent_t *stageFree = stageEvents.e;
    std::vector<float> cdf;
    size_t vo = 0, fo = 0;
    for (int i = 0; i < n; ++i) {
        const NoriGpuMeshDesc &m = meshes[i];
        MeshInfo mi; memset(&mi, 0, sizeof(mi));
        mi.bsdfType = m.bsdf_type;
        Bsdf b = makeBsdf(m.bsdf_type, m.bsdf_params);
        mi.albedo[0] = b.albedo.x; mi.albedo[1] = b.albedo.y; mi.albedo[2] = b.albedo.z;
        mi.alpha = b.alpha; mi.intIOR = b.intIOR; mi.extIOR = b.extIOR; mi.ks = b.ks;
        mi.emitter = -1; mi.hasN = m.N != nullptr; mi.hasUV = m.UV != nullptr;
        mi.triOffset = (uint32_t) fo; mi.nTris = m.nF; mi.vertOffset = (uint32_t) vo; mi.nVerts = m.nV;
        {   // two staging buffers: the copy of mesh i + 1 overlaps the packing of mesh i
            const int k = i & 1;
            if (stageFree[k]) CK(cudaEventSynchronize(stageFree[k])); else CK(cudaEventCreateWithFlags(&stageFree[k], cudaEventDisableTiming));
            float *dV = stageV[k].p, *dN = m.N ? dV + 3 * (size_t) m.nV : nullptr, *dUV = m.UV ? dV + 6 * (size_t) m.nV : nullptr;
            if (int r = stagedCopy(ctx, dV, m.V, (size_t) m.nV * 12)) return r;
            if (m.N) if (int r = stagedCopy(ctx, dN, m.N, (size_t) m.nV * 12)) return r;
            if (m.UV) if (int r = stagedCopy(ctx, dUV, m.UV, (size_t) m.nV * 8)) return r;
            if (m.nF) if (int r = stagedCopy(ctx, stageF[k].p, m.F, (size_t) m.nF * 12)) return r;
            launchPackMesh(dV, dN, dUV, stageF[k].p, m.nV, m.nF, (uint32_t) vo, (uint32_t) i, ctx->d_vpos + vo, ctx->d_vnrm + vo, ctx->d_vuv + vo,
                           ctx->d_tris + fo, bad.p, ctx->stream);
            CK(cudaEventRecord(stageFree[k], ctx->stream));
        }
        if (m.is_emitter) {
            // `area` emitter: DiscretePDF over triangle areas (Mesh::surfaceArea src/mesh.cpp:43-49;
            // append/normalize include/nori/dpdf.h:57-59,94-108). Sequential float sums on the host
            // so the table equals the reference's bit for bit.
            EmitterInfo e; memset(&e, 0, sizeof(e));
            e.mesh = i; e.radiance[0] = m.radiance[0]; e.radiance[1] = m.radiance[1]; e.radiance[2] = m.radiance[2];
            e.cdfOffset = (uint32_t) cdf.size(); e.nTris = m.nF; e.triOffset = (uint32_t) fo;
            cdf.push_back(0.0f);
            for (uint32_t f = 0; f < m.nF; ++f) {
                const uint32_t ia = m.F[3 * (size_t) f], ib = m.F[3 * (size_t) f + 1], ic = m.F[3 * (size_t) f + 2];
                const float *A = m.V + 3 * (size_t) ia, *B = m.V + 3 * (size_t) ib, *C = m.V + 3 * (size_t) ic;
                V3 p0 = mk(A[0], A[1], A[2]), p1 = mk(B[0], B[1], B[2]), p2 = mk(C[0], C[1], C[2]);
                volatile float area = 0.5f * sqrtf(squaredNorm(cross(p1 - p0, p2 - p0)));
                cdf.push_back(cdf.back() + area);
            }
            float sum = cdf.back();
            if (sum > 0) {
                float norm = 1.0f / sum;
                for (size_t k = e.cdfOffset + 1; k < cdf.size(); ++k) cdf[k] *= norm;
                cdf.back() = 1.0f;
            }
            e.area = sum;
            mi.emitter = (int) ctx->emitters.size();
            ctx->emitters.push_back(e);
        }
        ctx->meshes.push_back(mi);
        vo += m.nV; fo += m.nF;
    }
    CK(cudaMalloc((void **) &ctx->d_meshes, ctx->meshes.size() * sizeof(MeshInfo)));
    CK(cudaMalloc((void **) &ctx->d_emitters, std::max<size_t>(1, ctx->emitters.size()) * sizeof(EmitterInfo)));
    CK(cudaMalloc((void **) &ctx->d_cdf, std::max<size_t>(1, cdf.size()) * sizeof(float)));
    CK(cudaMemcpyAsync(ctx->d_meshes, ctx->meshes.data(), ctx->meshes.size() * sizeof(MeshInfo), cudaMemcpyHostToDevice, ctx->stream));
    if (!ctx->emitters.empty()) CK(cudaMemcpyAsync(ctx->d_emitters, ctx->emitters.data(), ctx->emitters.size() * sizeof(EmitterInfo), cudaMemcpyHostToDevice, ctx->stream));
    if (!cdf.empty()) CK(cudaMemcpyAsync(ctx->d_cdf, cdf.data(), cdf.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    unsigned int nBad = 0;
    CK(cudaMemcpyAsync(&nBad, bad.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    if (nBad) { ctx->err = "upload_meshes: face index out of range"; return NORI_GPU_EINVAL; }
    return NORI_GPU_OK;
}

int nori_gpu_upload_meshes(nori_gpu_ctx *ctx, const NoriGpuMeshDesc *meshes, int n) {
    if (!ctx) return NORI_GPU_EINVAL;
    const int r = uploadMeshes(ctx, meshes, n);
    if (r != NORI_GPU_OK) freeScene(ctx);          // no error return leaves a half-uploaded scene behind
    return r;
}

int nori_gpu_set_camera(nori_gpu_ctx *ctx, const NoriGpuCameraDesc *cam) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(cam && cam->width > 0 && cam->height > 0, "set_camera: bad output size");
    memcpy(ctx->cam.s2c, cam->sample_to_camera, 64); memcpy(ctx->cam.c2w, cam->camera_to_world, 64);
    ctx->cam.width = cam->width; ctx->cam.height = cam->height;
    ctx->cam.nearClip = cam->near_clip; ctx->cam.farClip = cam->far_clip;
    ctx->cam.invW = 1.0f / (float) cam->width; ctx->cam.invH = 1.0f / (float) cam->height;   // perspective.cpp:38
    ctx->cam.origin = mulPoint(ctx->cam.c2w, mk(0, 0, 0));
    ctx->haveCamera = true;
    return NORI_GPU_OK;
}

int nori_gpu_set_filter(nori_gpu_ctx *ctx, float radius, const float table33[33]) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(table33 && radius > 0, "set_filter: bad filter");
    ctx->filter.radius = radius;
    ctx->filter.border = (int) std::ceil(radius - 0.5f);                   // block.cpp:32
    ctx->filter.lookupFactor = 32 / radius;                                // block.cpp:38 (NORI_FILTER_RESOLUTION 32)
    memcpy(ctx->filter.table, table33, 33 * sizeof(float));
    REQUIRE(ctx->filter.border <= kMaxBorder, "set_filter: filter radius above 4.5 is not supported");
    ctx->haveFilter = true;
    return NORI_GPU_OK;
}

int nori_gpu_build_accel(nori_gpu_ctx *ctx, NoriGpuAccelStats *out) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(ctx->d_tris, "build_accel: upload meshes first");
    CK(cudaSetDevice(ctx->device));
    cudaFree(ctx->bvh.nodes); cudaFree(ctx->bvh.leafTris); cudaFree(ctx->bvh.qnodes); cudaFree(ctx->bvh.wnodes); ctx->bvh = BvhDevice(); ctx->haveBvh = false;
    NoriGpuAccelStats st;
    if (!buildBvh(ctx->d_vpos, ctx->d_tris, ctx->nTris, ctx->bvhParams, ctx->stream, ctx->bvh, st, ctx->err)) return NORI_GPU_ECUDA;
    ctx->haveBvh = true;
    if (ctx->accelLayout == NORI_ACCEL_QUANTISED && !quantiseNodes(ctx->bvh, ctx->bvhParams.inflate, ctx->stream, ctx->err)) return NORI_GPU_ECUDA;
    if (ctx->accelLayout == NORI_ACCEL_WIDE4 && !wideNodes(ctx->bvh, ctx->stream, ctx->err)) return NORI_GPU_ECUDA;
    if (out) *out = st;
    return NORI_GPU_OK;
}

int nori_gpu_set_accel_layout(nori_gpu_ctx *ctx, int layout) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(layout == NORI_ACCEL_FLOAT || layout == NORI_ACCEL_QUANTISED || layout == NORI_ACCEL_WIDE4, "set_accel_layout: unknown layout");
    CK(cudaSetDevice(ctx->device));
    ctx->accelLayout = layout;
    if (layout == NORI_ACCEL_QUANTISED && ctx->haveBvh && !quantiseNodes(ctx->bvh, ctx->bvhParams.inflate, ctx->stream, ctx->err)) return NORI_GPU_ECUDA;
    if (layout == NORI_ACCEL_WIDE4 && ctx->haveBvh && !wideNodes(ctx->bvh, ctx->stream, ctx->err)) return NORI_GPU_ECUDA;
    return NORI_GPU_OK;
}

// ---------------------------------------------------------------------------------------------
// The render call: replaces the body of render() between BlockGenerator construction and
// toBitmap (src/main.cpp:75-127).
static int runWavefront(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *d, const uint32_t *d_listPixel, const uint32_t *d_listSample, uint32_t listCount,
                        NoriGpuRenderStats *out) {
    const bool list = d_listPixel != nullptr;
    REQUIRE(ctx->haveBvh && ctx->haveCamera && ctx->haveFilter, "render: scene, camera, filter and accel must be set first");
    REQUIRE(d->integrator >= NORI_INT_NORMALS && d->integrator <= NORI_INT_AO, "render: unknown integrator");
    REQUIRE(list || d->sample_last > d->sample_first, "render: empty sample range");
    REQUIRE(list || d->sample_last <= 0x7fffffffu, "render: sample indices must stay below 2^31 (pcg32 stream ids are kept in 32 bits)");
    REQUIRE(d->intersect_mode == NORI_ISECT_REFERENCE_MT || d->intersect_mode == NORI_ISECT_WATERTIGHT, "render: unknown intersect_mode");
    if (!list && d->stream_mode == NORI_STREAM_BLOCK && d->integrator != NORI_INT_NORMALS && d->integrator != NORI_INT_SIMPLE) {
        ctx->err = "render: NORI_STREAM_BLOCK (one serial PCG32 stream per block) is only reproducible for integrators that draw no random numbers (normals, simple)";
        return NORI_GPU_EUNSUPPORTED;
    }
    if (!list && d->stream_mode == NORI_STREAM_BLOCK && d->sample_first != 0) {
        // the block stream is advanced by 4 * (pixel * spp + k): a sub-range would need the total spp, which the descriptor does not carry
        ctx->err = "render: NORI_STREAM_BLOCK renders the sample range [0, spp) only (sample_first must be 0; shard with NORI_STREAM_PER_SAMPLE)";
        return NORI_GPU_EUNSUPPORTED;
    }
    CK(cudaSetDevice(ctx->device));
    const int W = ctx->cam.width, H = ctx->cam.height;
    // Path slots this render works on. Unless the caller fixed the pool size: a third of the render's samples within [2 Mi, 8 Mi] (ramp-up, tail, the
    // per-iteration sort and the pool initialisation grow with the pool: 7.7 M samples take 11.7 ms on 2 Mi slots, 14.0 on 8 Mi), 16 Mi from 201 M
    // samples on (1024 spp: -2 %). The allocation only ever grows; a shorter render uses the front of it.
    uint32_t renderSlots = ctx->poolSize;
    if (ctx->poolAuto) {
        const uint64_t total = list ? (uint64_t) listCount : (uint64_t) W * H * (d->sample_last - d->sample_first);
        renderSlots = (!list && total >= (3ull << 26)) ? (1u << 24) : (uint32_t) std::min<uint64_t>(1u << 23, std::max<uint64_t>(1u << 21, total / 3 / 256 * 256));
        ctx->poolSize = std::max(ctx->poolSize, renderSlots);
    }
    if (int r = ensurePool(ctx)) return r;
    if (int r = ensureBlock(ctx)) return r;
    cudaStream_t st = ctx->stream;

    RenderParams rp; memset(&rp, 0, sizeof(rp));
    rp.integrator = d->integrator; rp.streamMode = d->stream_mode;
    rp.sampleFirst = d->sample_first; rp.sampleCount = list ? 1 : d->sample_last - d->sample_first;
    rp.rrStart = d->rr_start; rp.rrMax = d->rr_max; rp.maxDepth = d->max_depth;
    rp.lightPos = mk(d->light_pos[0], d->light_pos[1], d->light_pos[2]);
    rp.lightEnergy = mk(d->light_energy[0], d->light_energy[1], d->light_energy[2]);
    rp.tilesX = (W + kTile - 1) / kTile; rp.tilesY = (H + kTile - 1) / kTile; rp.numTiles = (uint32_t) (rp.tilesX * rp.tilesY);
    rp.block = ctx->d_block; rp.blockCols = W + 2 * ctx->filter.border; rp.blockRows = H + 2 * ctx->filter.border;
    uint64_t totalSamples;
    // Slots used by this render: never more than there are camera samples (small frames, sample lists)
    const uint64_t wanted = list ? (uint64_t) listCount : (uint64_t) W * H * (d->sample_last - d->sample_first);
    Pool pool = ctx->pool;                       // .capacity also shrinks in the tail of the render (enqueueCompaction)
    for (int k = 0; k < 3; ++k) {
        const float ext = ctx->bvh.sceneHi[k] - ctx->bvh.sceneLo[k];
        pool.binLo[k] = ctx->bvh.sceneLo[k]; pool.binInv[k] = ext > 0.f ? 4.0f / ext : 0.f;
    }
    pool.capacity = (uint32_t) std::min<uint64_t>(std::min(ctx->pool.capacity, renderSlots), std::max<uint64_t>(4096, (wanted + 255) / 256 * 256));
    if (list) {
        rp.K = rp.Klast = 1; rp.numPasses = 1; rp.passSlots = listCount; rp.totalIssue = listCount; rp.ring = 1;
        rp.listPixel = d_listPixel; rp.listSample = d_listSample; totalSamples = listCount;
    } else {
        // pass = K samples of every pixel; sized to a few pool refills so stragglers never stall the ring
        const uint64_t slotsPerSpp = (uint64_t) rp.numTiles * 1024u;
        uint32_t K = (uint32_t) std::max<uint64_t>(1, (4ull * pool.capacity + slotsPerSpp - 1) / slotsPerSpp);
        K = std::min(K, rp.sampleCount);
        rp.K = K; rp.numPasses = (rp.sampleCount + K - 1) / K; rp.Klast = rp.sampleCount - (rp.numPasses - 1) * K;
        rp.passSlots = slotsPerSpp * K; rp.ring = kMaxRing;
        rp.totalIssue = (uint64_t) (rp.numPasses - 1) * rp.passSlots + slotsPerSpp * rp.Klast;
        totalSamples = (uint64_t) W * H * rp.sampleCount;
    }
    const size_t recNeed = (size_t) rp.ring * rp.passSlots;
    if (recNeed > ctx->recCap) {
        cudaFree(ctx->recA); cudaFree(ctx->recB); ctx->recA = nullptr; ctx->recB = nullptr; ctx->recCap = 0;
        CK(cudaMalloc((void **) &ctx->recA, recNeed * sizeof(float4))); CK(cudaMalloc((void **) &ctx->recB, recNeed * sizeof(float)));
        ctx->recCap = recNeed;
    }
    const size_t tdNeed = (size_t) rp.ring * rp.numTiles;
    if (tdNeed > ctx->tileDoneCap) {
        cudaFree(ctx->tileDone); cudaFree(ctx->tileGen); cudaFree(ctx->tileArrive); cudaFree(ctx->readyTiles); ctx->tileDone = ctx->tileGen = ctx->tileArrive = ctx->readyTiles = nullptr;
        CK(cudaMalloc((void **) &ctx->readyTiles, tdNeed * sizeof(uint32_t)));
        CK(cudaMalloc((void **) &ctx->tileArrive, tdNeed * sizeof(uint32_t)));
        CK(cudaMalloc((void **) &ctx->tileDone, tdNeed * sizeof(uint32_t))); CK(cudaMalloc((void **) &ctx->tileGen, tdNeed * sizeof(uint32_t)));
        ctx->tileDoneCap = tdNeed;
    }
    rp.recA = ctx->recA; rp.recB = ctx->recB; rp.tileDone = ctx->tileDone; rp.tileGen = ctx->tileGen; rp.tileArrive = ctx->tileArrive; rp.readyTiles = ctx->readyTiles; rp.poolCapacity = pool.capacity; rp.recCount = recNeed;
    CK(cudaMemsetAsync(ctx->tileDone, 0, tdNeed * sizeof(uint32_t), st));
    CK(cudaMemsetAsync(ctx->tileGen, 0, tdNeed * sizeof(uint32_t), st));
    CK(cudaMemsetAsync(ctx->tileArrive, 0, tdNeed * sizeof(uint32_t), st));
    if (!list) CK(cudaMemsetAsync(ctx->recA, 0xff, recNeed * sizeof(float4), st));        // NaN position = no sample recorded
    CK(cudaMemsetAsync(ctx->d_counters, 0, sizeof(Counters), st));
    enqueuePoolInit(pool, ctx->d_counters, st);
    if (d->clear && !list) CK(cudaMemsetAsync(ctx->d_block, 0, ctx->blockFloats * sizeof(float), st));

    DeviceScene sc = deviceScene(ctx);
    sc.isectMode = d->intersect_mode;
    // Stage events (profiling): 4 per iteration of a batch, read back after the batch's one
    // synchronisation, so that profiling does not change how the kernels are queued.
    const int batch = 8;
    // Two batches are kept in flight while camera samples are still being handed out: the host reads the counters of batch k
    // while batch k + 1 already runs, so the device never idles for the round trip (26 of them per 1024-spp frame, ~0.27 ms each).
    // Nothing the host decides in that phase depends on the counters. Once every sample has been issued (tail: compaction,
    // k_finishTail, termination) the loop falls back to one batch of 2 iterations at a time.
    struct Events {                              // destroyed on every return path
        cudaEvent_t e0 = nullptr, e1 = nullptr, evCopy[2] = {nullptr, nullptr}; std::vector<cudaEvent_t> stage;
        ~Events() { for (cudaEvent_t e : {e0, e1, evCopy[0], evCopy[1]}) if (e) cudaEventDestroy(e); for (cudaEvent_t e : stage) if (e) cudaEventDestroy(e); }
    } evs;
    evs.stage.assign(ctx->profiling ? 2 * 4 * batch : 0, nullptr);
    cudaEvent_t &e0 = evs.e0, &e1 = evs.e1;
    std::vector<cudaEvent_t> &stage = evs.stage;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (cudaEvent_t &e : evs.evCopy) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : stage) CK(cudaEventCreate(&e));
    StageTimes tm = {0, 0, 0, 0, 0};
    CK(cudaEventRecord(e0, st));
    uint64_t launches = 0;
    bool done = false;
#ifdef NORI_GPU_SWEEP
    const bool debug = getenv("NORI_GPU_DEBUG") != nullptr;
    const uint32_t kTailPaths = getenv("NORI_GPU_TAIL") ? (uint32_t) atoi(getenv("NORI_GPU_TAIL")) : 262144u;   // sweep: profiles/r01_pipeline.md
    const int depthSteady = getenv("NORI_GPU_NO_PIPELINE") ? 1 : 2;
#else
    const bool debug = false;
    const uint32_t kTailPaths = 262144u;
    const int depthSteady = 2;
#endif
    uint64_t guard = 0;
    bool drainedBefore = false, tailDone = false, drainedSeen = false;
    int batchNow = batch;                        // iterations per batch
    int inFlight = 0, head = 0, sizes[2] = {0, 0};   // ring of batches whose counters have not been looked at yet: [head, head + inFlight)
    const Counters *last = ctx->h_counters;
    while (!done) {
        const int depth = drainedSeen ? 1 : depthSteady;
        while (inFlight < depth) {
            const int b = (head + inFlight) & 1;
            for (int k = 0; k < batchNow; ++k)
                launches += (uint64_t) enqueueIteration(sc, pool, rp, ctx->d_counters, st, ctx->profiling ? stage.data() + 4 * (b * batch + k) : nullptr);
            CK(cudaMemcpyAsync(ctx->h_counters + b, ctx->d_counters, sizeof(Counters), cudaMemcpyDeviceToHost, st));
            CK(cudaEventRecord(evs.evCopy[b], st));
            sizes[b] = batchNow; ++inFlight;
        }
        const int b = head;
        CK(cudaEventSynchronize(evs.evCopy[b]));
        CK(cudaGetLastError());
        head ^= 1; --inFlight;
        if (ctx->profiling) {
            for (int k = 0; k < sizes[b]; ++k) {
                cudaEvent_t *ev = stage.data() + 4 * (b * batch + k); float ms;
                cudaEventElapsedTime(&ms, ev[0], ev[1]); tm.traceClosest += ms;    // splat of the previous iteration runs under it
                cudaEventElapsedTime(&ms, ev[1], ev[2]); tm.shade += ms;           // k_prepare + k_shadeRegen
                cudaEventElapsedTime(&ms, ev[2], ev[3]); tm.traceShadow += ms;
            }
        }
        const Counters &hc = ctx->h_counters[b];
        last = &hc;
        if (debug) { float tms = 0; cudaEventRecord(e1, st); cudaEventSynchronize(e1); cudaEventElapsedTime(&tms, e0, e1); fprintf(stderr, "[nori_gpu] t=%.3fms cap=%u live=%u ", tms, pool.capacity, hc.liveLast); }
        if (debug) fprintf(stderr, "[nori_gpu] it=%llu issue=%llu/%llu limit=%llu finished=%llu/%llu passes=%u/%u splat=[%u %u %u %u]\n", (unsigned long long) hc.iterations, (unsigned long long) hc.issue, (unsigned long long) rp.totalIssue, (unsigned long long) hc.issueLimit, (unsigned long long) hc.finished, (unsigned long long) totalSamples, hc.completedPasses, rp.numPasses, hc.tilesSplatted[0], hc.tilesSplatted[1], hc.tilesSplatted[2], hc.tilesSplatted[3]);
        done = hc.finished >= totalSamples && (list || hc.completedPasses >= rp.numPasses);
        // Tail: every camera sample has been handed out and the pool drains. Pack the surviving
        // paths into the front of the pool so that the kernels stop scanning empty slots.
        // (liveLast bounds the occupied slots only once a whole batch has run without regeneration.)
        // (issue overshoots its limit transiently: every free slot adds to it before k_prepare clamps it)
        const bool drained = std::min<unsigned long long>(hc.issue, hc.issueLimit) >= rp.totalIssue;
        if (drained && !drainedSeen) {           // first sight of the tail: from here on the counters steer the loop, no look-ahead
            drainedSeen = true; batchNow = 2;
            if (inFlight) { drainedBefore = true; continue; }   // the batch already queued behind this one ran without regeneration: examine it before anything is decided
        }
        if (!done && drained && drainedBefore && inFlight == 0 && pool.capacity > 4096u && 2ull * hc.liveLast <= pool.capacity) {
            const uint32_t newCap = std::max(4096u, (hc.liveLast + 255u) / 256u * 256u);
            launches += (uint64_t) enqueueCompaction(pool, newCap, ctx->d_counters, st);
            pool.capacity = newCap;
            enqueueIdentityOrder(pool, newCap, ctx->d_counters, st);         // paths changed slots: the next closest-hit pass walks the pool in slot order
            // A few thousand paths left: one launch follows each of them to its end (k_finishTail)
            // instead of one launch-bound wavefront iteration per bounce of the longest path.
            if (newCap <= kTailPaths && !tailDone) { launches += (uint64_t) enqueueFinishTail(sc, pool, rp, ctx->d_counters, st); tailDone = true; }
        }
        drainedBefore = drained;
        if (++guard > 4000000ull) { ctx->err = "render: wavefront did not terminate (internal error)"; return NORI_GPU_ECUDA; }
    }
    while (inFlight) {                           // a look-ahead batch still queued when the end was seen: its counters are the final ones
        const int b2 = head;
        CK(cudaEventSynchronize(evs.evCopy[b2]));
        last = ctx->h_counters + b2; head ^= 1; --inFlight;
    }
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    if (int r = reportChecks(ctx, "render")) return r;
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    if (out) {
        const Counters &hc = *last;
        memset(out, 0, sizeof(*out));
        uint64_t rays = hc.closestRays, shadows = hc.shadowRays + hc.shadowCount;
        for (int k = 0; k < kNumKeys; ++k) rays += hc.sortCount[k];
        out->camera_samples = hc.finished; out->closest_rays = rays; out->shadow_rays = shadows; out->segments = rays;
        out->dropped = hc.dropped; out->iterations = hc.iterations; out->kernel_launches = launches; out->ms_total = ms;
        out->ms_generate = tm.generate; out->ms_trace_closest = tm.traceClosest; out->ms_shade = tm.shade;
        out->ms_trace_shadow = tm.traceShadow; out->ms_splat = tm.splat; out->late_samples = hc.lateSamples;
    }
    return NORI_GPU_OK;
}

int nori_gpu_render(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, NoriGpuRenderStats *out) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(desc, "render: null descriptor");
    if (desc->deterministic) { ctx->err = "render: deterministic splat order is only offered by nori_gpu_splat"; return NORI_GPU_EUNSUPPORTED; }
    return runWavefront(ctx, desc, nullptr, nullptr, 0, out);
}

int nori_gpu_sample_radiance(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, int n, const uint32_t *pixel, const uint32_t *sample, float *rgb, float *pos) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(desc && n > 0 && pixel && sample && rgb, "sample_radiance: bad arguments");
    for (int i = 0; i < n; ++i) REQUIRE(sample[i] <= 0x7fffffffu, "sample_radiance: sample indices must stay below 2^31");
    CK(cudaSetDevice(ctx->device));
    DevBuf<uint32_t> dp, ds;
    CK(dp.alloc(n)); CK(ds.alloc(n));
    CK(cudaMemcpy(dp.p, pixel, (size_t) n * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(ds.p, sample, (size_t) n * 4, cudaMemcpyHostToDevice));
    NoriGpuRenderStats st;
    if (int r = runWavefront(ctx, desc, dp.p, ds.p, (uint32_t) n, &st)) return r;
    std::vector<float4> a(n); std::vector<float> b(n);
    CK(cudaMemcpy(a.data(), ctx->recA, (size_t) n * 16, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(b.data(), ctx->recB, (size_t) n * 4, cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; ++i) {
        rgb[3 * i] = a[i].z; rgb[3 * i + 1] = a[i].w; rgb[3 * i + 2] = b[i];
        if (pos) { pos[2 * i] = a[i].x; pos[2 * i + 1] = a[i].y; }
    }
    return NORI_GPU_OK;
}

int nori_gpu_li(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *d, int n, const float *rays, uint64_t *rng_state, float *rgb) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(d && n > 0 && rays && rng_state && rgb, "li: bad arguments");
    REQUIRE(ctx->haveBvh, "li: scene and accel must be set first");
    REQUIRE(d->integrator >= NORI_INT_NORMALS && d->integrator <= NORI_INT_AO, "li: unknown integrator");
    REQUIRE(d->intersect_mode == NORI_ISECT_REFERENCE_MT || d->intersect_mode == NORI_ISECT_WATERTIGHT, "li: unknown intersect_mode");
    CK(cudaSetDevice(ctx->device));
    RenderParams rp; memset(&rp, 0, sizeof(rp));
    rp.integrator = d->integrator; rp.rrStart = d->rr_start; rp.rrMax = d->rr_max; rp.maxDepth = d->max_depth;
    rp.lightPos = mk(d->light_pos[0], d->light_pos[1], d->light_pos[2]);
    rp.lightEnergy = mk(d->light_energy[0], d->light_energy[1], d->light_energy[2]);
    DeviceScene sc = deviceScene(ctx); sc.isectMode = d->intersect_mode;
    DevBuf<float4> dr; DevBuf<unsigned long long> ds; DevBuf<float> dc;
    CK(dr.alloc(2 * (size_t) n)); CK(ds.alloc(2 * (size_t) n)); CK(dc.alloc(3 * (size_t) n));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(dr.p, rays, (size_t) n * 32, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ds.p, rng_state, (size_t) n * 16, cudaMemcpyHostToDevice, st));
    enqueueLi(sc, rp, n, dr.p, ds.p, dc.p, st);
    CK(cudaMemcpyAsync(rgb, dc.p, (size_t) n * 12, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(rng_state, ds.p, (size_t) n * 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (int r = reportChecks(ctx, "li")) return r;
    return NORI_GPU_OK;
}

int nori_gpu_set_ray_binning(nori_gpu_ctx *ctx, int on) {
    if (!ctx) return NORI_GPU_EINVAL;
    ctx->rayBins = on != 0;          // takes effect at the next render (ensurePool reallocates the two index arrays)
    return NORI_GPU_OK;
}
int nori_gpu_set_plain_bsdf_kernels(nori_gpu_ctx *ctx, int on) {
    if (!ctx) return NORI_GPU_EINVAL;
    ctx->plainKernels = on != 0;
    return NORI_GPU_OK;
}
int nori_gpu_set_intersect_mode(nori_gpu_ctx *ctx, int mode) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(mode == NORI_ISECT_REFERENCE_MT || mode == NORI_ISECT_WATERTIGHT, "set_intersect_mode: unknown mode");
    ctx->isectMode = mode;
    return NORI_GPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Cross-device sum of ImageBlocks inside one process: `result.put(block)` (src/block.cpp:105-114) between
// the contexts of several GPUs as ONE ncclReduce over NVLink / NVSwitch. (One process per GPU — bench.py,
// torchrun — does the same reduce through torch.distributed on nori_gpu_block_device_ptr.)
// NCCL is resolved at run time (libnccl.so.2): a single-device build or box never needs it.
namespace {
typedef struct ncclComm *ncclComm_t;                  // declarations after nccl.h (NCCL 2.x, stable C ABI)
struct NcclApi {
    void *lib = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int /*ncclDataType_t*/, int /*ncclRedOp_t*/, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string error;
    std::map<std::vector<int>, std::vector<ncclComm_t>> comms;   // one communicator set per device list
    std::mutex lock;
};
constexpr int kNcclFloat32 = 7, kNcclSum = 0;         // nccl.h: ncclFloat32, ncclSum

NcclApi &nccl() {
    static NcclApi api;
    if (api.lib || !api.error.empty()) return api;
    const char *names[] = {getenv("NORI_GPU_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) { if (nm && (api.lib = dlopen(nm, RTLD_NOW | RTLD_LOCAL))) break; }
    if (!api.lib) { api.error = std::string("NCCL is not available (") + dlerror() + ")"; return api; }
    auto sym = [&](const char *n) { void *f = dlsym(api.lib, n); if (!f) api.error = std::string("libnccl lacks ") + n; return f; };
    api.CommInitAll = (decltype(api.CommInitAll)) sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy)) sym("ncclCommDestroy");
    api.Reduce = (decltype(api.Reduce)) sym("ncclReduce");
    api.GroupStart = (decltype(api.GroupStart)) sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd)) sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString)) sym("ncclGetErrorString");
    return api;
}
} // namespace

int nori_gpu_reduce_blocks(nori_gpu_ctx *const *ctxs, int n, int root) {
    if (!ctxs || n < 1 || root < 0 || root >= n || !ctxs[root]) return NORI_GPU_EINVAL;
    nori_gpu_ctx *ctx = ctxs[root];                   // error messages land on the root context
    std::vector<int> devs((size_t) n);
    for (int i = 0; i < n; ++i) {
        REQUIRE(ctxs[i] && ctxs[i]->d_block, "reduce_blocks: a context has no ImageBlock yet");
        REQUIRE(ctxs[i]->blockFloats == ctx->blockFloats, "reduce_blocks: the contexts hold ImageBlocks of different sizes");
        devs[(size_t) i] = ctxs[i]->device;
        for (int j = 0; j < i; ++j) REQUIRE(ctxs[j]->device != ctxs[i]->device, "reduce_blocks: one context per device");
    }
    if (n == 1) return NORI_GPU_OK;
    NcclApi &api = nccl();
    if (!api.error.empty()) { ctx->err = "reduce_blocks: " + api.error; return NORI_GPU_EUNSUPPORTED; }
    std::lock_guard<std::mutex> guard(api.lock);
#define NCK(call) do { int r_ = (call); if (r_ != 0) { ctx->err = std::string(#call) + " failed: " + api.GetErrorString(r_); return NORI_GPU_ECUDA; } } while (0)
    auto it = api.comms.find(devs);
    if (it == api.comms.end()) {
        std::vector<ncclComm_t> comms((size_t) n);
        NCK(api.CommInitAll(comms.data(), n, devs.data()));
        it = api.comms.emplace(devs, comms).first;
    }
    for (int i = 0; i < n; ++i) { CK(cudaSetDevice(devs[(size_t) i])); CK(cudaStreamSynchronize(ctxs[i]->stream)); }
    NCK(api.GroupStart());
    for (int i = 0; i < n; ++i) {
        CK(cudaSetDevice(devs[(size_t) i]));
        NCK(api.Reduce(ctxs[i]->d_block, ctxs[i]->d_block, ctx->blockFloats, kNcclFloat32, kNcclSum, root, it->second[(size_t) i], ctxs[i]->stream));
    }
    NCK(api.GroupEnd());
    for (int i = 0; i < n; ++i) { CK(cudaSetDevice(devs[(size_t) i])); CK(cudaStreamSynchronize(ctxs[i]->stream)); }
    CK(cudaSetDevice(ctx->device));
#undef NCK
    return NORI_GPU_OK;
}

int nori_gpu_device_count(void) {
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

int nori_gpu_block_dims(const nori_gpu_ctx *ctx, int *rows, int *cols, int *border) {
    if (!ctx || !ctx->haveCamera || !ctx->haveFilter) return NORI_GPU_EINVAL;
    if (rows) *rows = ctx->cam.height + 2 * ctx->filter.border;
    if (cols) *cols = ctx->cam.width + 2 * ctx->filter.border;
    if (border) *border = ctx->filter.border;
    return NORI_GPU_OK;
}

int nori_gpu_block_device_ptr(nori_gpu_ctx *ctx, void **dptr, uint64_t *n_floats) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(ctx->haveCamera && ctx->haveFilter, "block_device_ptr: camera and filter must be set first");
    CK(cudaSetDevice(ctx->device));
    if (int r = ensureBlock(ctx)) return r;
    CK(cudaStreamSynchronize(ctx->stream));
    if (dptr) *dptr = ctx->d_block;
    if (n_floats) *n_floats = ctx->blockFloats;
    return NORI_GPU_OK;
}

int nori_gpu_download_block(nori_gpu_ctx *ctx, float *rgbw) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(rgbw && ctx->d_block, "download_block: nothing rendered");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(rgbw, ctx->d_block, ctx->blockFloats * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return NORI_GPU_OK;
}

int nori_gpu_download_bitmap(nori_gpu_ctx *ctx, float *rgb) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(rgb && ctx->d_block, "download_bitmap: nothing rendered");
    CK(cudaSetDevice(ctx->device));
    const size_t n = (size_t) ctx->cam.width * ctx->cam.height * 3;
    DevBuf<float> d; CK(d.alloc(n));
    launchToBitmap(ctx->d_block, ctx->cam.width, ctx->cam.height, ctx->filter.border, d.p, ctx->stream);
    CK(cudaMemcpyAsync(rgb, d.p, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return NORI_GPU_OK;
}

int nori_gpu_download_srgb8(nori_gpu_ctx *ctx, uint8_t *rgb8) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(rgb8 && ctx->d_block, "download_srgb8: nothing rendered");
    CK(cudaSetDevice(ctx->device));
    const size_t n = (size_t) ctx->cam.width * ctx->cam.height * 3;
    DevBuf<uint8_t> d; CK(d.alloc(n));
    launchToSRGB8(ctx->d_block, ctx->cam.width, ctx->cam.height, ctx->filter.border, d.p, ctx->stream);
    CK(cudaMemcpyAsync(rgb8, d.p, n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return NORI_GPU_OK;
}

int nori_gpu_upload_block(nori_gpu_ctx *ctx, const float *rgbw, int accumulate) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(rgbw, "upload_block: null block");
    REQUIRE(ctx->haveCamera && ctx->haveFilter, "upload_block: camera and filter must be set first");
    CK(cudaSetDevice(ctx->device));
    if (int r = ensureBlock(ctx)) return r;
    if (!accumulate) {
        CK(cudaMemcpyAsync(ctx->d_block, rgbw, ctx->blockFloats * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    } else {
        DevBuf<float> d; CK(d.alloc(ctx->blockFloats));
        CK(cudaMemcpyAsync(d.p, rgbw, ctx->blockFloats * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
        launchAddBlock(ctx->d_block, d.p, ctx->blockFloats, ctx->stream);
        CK(cudaStreamSynchronize(ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Raw operators. Host arrays in, host arrays out; the work itself runs on the device.
#define NEED_SCENE() REQUIRE(ctx->haveBvh, "operator needs an uploaded scene with a built accel")

static void splitHits(const std::vector<float4> &h, const nori_gpu_ctx *ctx, int n, uint32_t *tri_id, uint32_t *mesh_id, float *tuv) {
    for (int i = 0; i < n; ++i) {
        uint32_t gid; memcpy(&gid, &h[i].w, 4);
        if (gid == 0xffffffffu) { tri_id[i] = 0xffffffffu; if (mesh_id) mesh_id[i] = 0xffffffffu; }
        else {
            // global id -> (mesh, local index): meshes are concatenated in upload order
            size_t lo = 0, hi = ctx->meshes.size();
            while (hi - lo > 1) { size_t mid = (lo + hi) / 2; if (ctx->meshes[mid].triOffset <= gid) lo = mid; else hi = mid; }
            tri_id[i] = gid - ctx->meshes[lo].triOffset; if (mesh_id) mesh_id[i] = (uint32_t) lo;
        }
        if (tuv) { tuv[3 * i] = h[i].x; tuv[3 * i + 1] = h[i].y; tuv[3 * i + 2] = h[i].z; }
    }
}

int nori_gpu_trace(nori_gpu_ctx *ctx, const float *rays, int n, int shadow, uint32_t *tri_id, uint32_t *mesh_id, float *tuv) {
    if (!ctx) return NORI_GPU_EINVAL;
    NEED_SCENE(); REQUIRE(rays && tri_id && n >= 0, "trace: bad arguments");
    if (n == 0) return NORI_GPU_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf<float4> dr, dh; CK(dr.alloc(2 * (size_t) n)); CK(dh.alloc(n));
    CK(cudaMemcpyAsync(dr.p, rays, (size_t) n * 32, cudaMemcpyHostToDevice, ctx->stream));
    launchTrace(deviceScene(ctx), dr.p, n, shadow != 0, dh.p, &ctx->d_counters->traceCursor[0], ctx->stream);
    std::vector<float4> h(n);
    CK(cudaMemcpyAsync(h.data(), dh.p, (size_t) n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    if (int r = reportChecks(ctx, "trace")) return r;
    splitHits(h, ctx, n, tri_id, mesh_id, tuv);
    return NORI_GPU_OK;
}

int nori_gpu_trace_device(nori_gpu_ctx *ctx, const void *d_rays, int n, int shadow, void *d_hits, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    NEED_SCENE(); REQUIRE(d_rays && d_hits && n > 0, "trace_device: bad arguments");
    CK(cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, ctx->stream));
    launchTrace(deviceScene(ctx), (const float4 *) d_rays, n, shadow != 0, (float4 *) d_hits, &ctx->d_counters->traceCursor[0], ctx->stream);
    CK(cudaEventRecord(e1, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    float t = 0; cudaEventElapsedTime(&t, e0, e1); cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (ms) *ms = t;
    return NORI_GPU_OK;
}

/// Traversal work counters of the device BVH for n rays: counts[0] = node visits (two child
/// boxes each), counts[1] = triangle tests. Feeds the algorithmic-bytes figure of the roofline.
int nori_gpu_trace_count(nori_gpu_ctx *ctx, const float *rays, int n, int shadow, uint64_t counts[2]) {
    if (!ctx) return NORI_GPU_EINVAL;
    NEED_SCENE(); REQUIRE(rays && counts && n > 0, "trace_count: bad arguments");
    CK(cudaSetDevice(ctx->device));
    DevBuf<float4> dr, dh; DevBuf<unsigned long long> dc; CK(dr.alloc(2 * (size_t) n)); CK(dh.alloc(n)); CK(dc.alloc(2));
    CK(cudaMemcpyAsync(dr.p, rays, (size_t) n * 32, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemsetAsync(dc.p, 0, 16, ctx->stream));
    launchTraceCount(deviceScene(ctx), dr.p, n, shadow != 0, dh.p, dc.p, ctx->stream);
    unsigned long long h[2];
    CK(cudaMemcpyAsync(h, dc.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    if (int r = reportChecks(ctx, "trace_count")) return r;
    counts[0] = h[0]; counts[1] = h[1];
    return NORI_GPU_OK;
}

int nori_gpu_intersect(nori_gpu_ctx *ctx, const float *rays, int n, float *out24, int32_t *valid) {
    if (!ctx) return NORI_GPU_EINVAL;
    NEED_SCENE(); REQUIRE(rays && out24 && valid && n >= 0, "intersect: bad arguments");
    if (n == 0) return NORI_GPU_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf<float4> dr; DevBuf<float> dout; DevBuf<int> dv; CK(dr.alloc(2 * (size_t) n)); CK(dout.alloc(24 * (size_t) n)); CK(dv.alloc(n));
    CK(cudaMemcpyAsync(dr.p, rays, (size_t) n * 32, cudaMemcpyHostToDevice, ctx->stream));
    launchIntersect(deviceScene(ctx), dr.p, n, dout.p, dv.p, ctx->stream);
    CK(cudaMemcpyAsync(out24, dout.p, (size_t) n * 96, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(valid, dv.p, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

int nori_gpu_camera_rays(nori_gpu_ctx *ctx, const float *pixel_samples, int n, float *rays) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(ctx->haveCamera, "camera_rays: set the camera first"); REQUIRE(pixel_samples && rays && n >= 0, "camera_rays: bad arguments");
    if (n == 0) return NORI_GPU_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf<float2> dp; DevBuf<float4> dr; CK(dp.alloc(n)); CK(dr.alloc(2 * (size_t) n));
    CK(cudaMemcpyAsync(dp.p, pixel_samples, (size_t) n * 8, cudaMemcpyHostToDevice, ctx->stream));
    launchCameraRays(deviceScene(ctx), dp.p, n, dr.p, ctx->stream);
    CK(cudaMemcpyAsync(rays, dr.p, (size_t) n * 32, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

int nori_gpu_bsdf_sample(nori_gpu_ctx *ctx, int bsdf_type, const float params[6], int n, const float *wi, const float *xi,
                         float *wo, float *weight, float *pdf, float *eta, int32_t *measure) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(bsdf_type >= 0 && bsdf_type <= 3 && params && n > 0 && wi && xi && wo && weight && pdf && eta && measure, "bsdf_sample: bad arguments");
    CK(cudaSetDevice(ctx->device));
    DevBuf<float> dwi, dxi, dwo, dw, dpdf, deta; DevBuf<int> dm;
    CK(dwi.alloc(3 * (size_t) n)); CK(dxi.alloc(2 * (size_t) n)); CK(dwo.alloc(3 * (size_t) n)); CK(dw.alloc(3 * (size_t) n));
    CK(dpdf.alloc(n)); CK(deta.alloc(n)); CK(dm.alloc(n));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(dwi.p, wi, (size_t) n * 12, cudaMemcpyHostToDevice, st)); CK(cudaMemcpyAsync(dxi.p, xi, (size_t) n * 8, cudaMemcpyHostToDevice, st));
    launchBsdfSample(makeBsdf(bsdf_type, params), n, dwi.p, dxi.p, dwo.p, dw.p, dpdf.p, deta.p, dm.p, st);
    CK(cudaMemcpyAsync(wo, dwo.p, (size_t) n * 12, cudaMemcpyDeviceToHost, st)); CK(cudaMemcpyAsync(weight, dw.p, (size_t) n * 12, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(pdf, dpdf.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st)); CK(cudaMemcpyAsync(eta, deta.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(measure, dm.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

int nori_gpu_bsdf_eval_pdf(nori_gpu_ctx *ctx, int bsdf_type, const float params[6], int n, const float *wi, const float *wo, float *f, float *pdf) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(bsdf_type >= 0 && bsdf_type <= 3 && params && n > 0 && wi && wo && f && pdf, "bsdf_eval_pdf: bad arguments");
    CK(cudaSetDevice(ctx->device));
    DevBuf<float> dwi, dwo, df, dpdf;
    CK(dwi.alloc(3 * (size_t) n)); CK(dwo.alloc(3 * (size_t) n)); CK(df.alloc(3 * (size_t) n)); CK(dpdf.alloc(n));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(dwi.p, wi, (size_t) n * 12, cudaMemcpyHostToDevice, st)); CK(cudaMemcpyAsync(dwo.p, wo, (size_t) n * 12, cudaMemcpyHostToDevice, st));
    launchBsdfEvalPdf(makeBsdf(bsdf_type, params), n, dwi.p, dwo.p, df.p, dpdf.p, st);
    CK(cudaMemcpyAsync(f, df.p, (size_t) n * 12, cudaMemcpyDeviceToHost, st)); CK(cudaMemcpyAsync(pdf, dpdf.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

int nori_gpu_warp(nori_gpu_ctx *ctx, int warp_id, float param, int n, const float *xi, float *out3, float *pdf) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(warp_id >= 0 && warp_id <= 6 && n > 0 && xi && out3 && pdf, "warp: bad arguments");
    CK(cudaSetDevice(ctx->device));
    DevBuf<float> dxi, dout, dpdf; CK(dxi.alloc(2 * (size_t) n)); CK(dout.alloc(3 * (size_t) n)); CK(dpdf.alloc(n));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(dxi.p, xi, (size_t) n * 8, cudaMemcpyHostToDevice, st));
    launchWarp(warp_id, param, n, dxi.p, dout.p, dpdf.p, st);
    CK(cudaMemcpyAsync(out3, dout.p, (size_t) n * 12, cudaMemcpyDeviceToHost, st)); CK(cudaMemcpyAsync(pdf, dpdf.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

int nori_gpu_warp_pdf(nori_gpu_ctx *ctx, int warp_id, float param, int n, const float *points3, float *pdf) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(warp_id >= 0 && warp_id <= 6 && n > 0 && points3 && pdf, "warp_pdf: bad arguments");
    CK(cudaSetDevice(ctx->device));
    DevBuf<float> dp, dpdf; CK(dp.alloc(3 * (size_t) n)); CK(dpdf.alloc(n));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(dp.p, points3, (size_t) n * 12, cudaMemcpyHostToDevice, st));
    launchWarpPdf(warp_id, param, n, dp.p, dpdf.p, st);
    CK(cudaMemcpyAsync(pdf, dpdf.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

int nori_gpu_pcg32(nori_gpu_ctx *ctx, uint64_t initstate, uint64_t initseq, int64_t advance, int n, uint32_t *out) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(n > 0 && out, "pcg32: bad arguments");
    CK(cudaSetDevice(ctx->device));
    DevBuf<uint32_t> d; CK(d.alloc(n));
    launchPcg32(initstate, initseq, advance, n, d.p, ctx->stream);
    CK(cudaMemcpyAsync(out, d.p, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}

int nori_gpu_splat(nori_gpu_ctx *ctx, int n, const float *pos, const float *rgb, int deterministic, float *rgbw, uint64_t *dropped) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(ctx->haveCamera && ctx->haveFilter, "splat: set camera and filter first"); REQUIRE(n > 0 && pos && rgb && rgbw, "splat: bad arguments");
    CK(cudaSetDevice(ctx->device));
    const size_t cells = (size_t) (ctx->cam.width + 2 * ctx->filter.border) * (ctx->cam.height + 2 * ctx->filter.border);
    DevBuf<float2> dp; DevBuf<float> dc, db; DevBuf<unsigned long long> dd;
    CK(dp.alloc(n)); CK(dc.alloc(3 * (size_t) n)); CK(db.alloc(cells * 4)); CK(dd.alloc(1));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(dp.p, pos, (size_t) n * 8, cudaMemcpyHostToDevice, st)); CK(cudaMemcpyAsync(dc.p, rgb, (size_t) n * 12, cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(db.p, 0, cells * 16, st)); CK(cudaMemsetAsync(dd.p, 0, 8, st));
    if (deterministic) launchSplatOrdered(deviceScene(ctx), n, dp.p, dc.p, db.p, dd.p, st);
    else launchSplatList(deviceScene(ctx), n, dp.p, dc.p, db.p, dd.p, st);
    unsigned long long hd = 0;
    CK(cudaMemcpyAsync(rgbw, db.p, cells * 16, cudaMemcpyDeviceToHost, st)); CK(cudaMemcpyAsync(&hd, dd.p, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (int r = reportChecks(ctx, "splat")) return r;
    if (dropped) *dropped = hd;
    return NORI_GPU_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// The step before the path: OBJ corner de-duplication on the device (objdedup.cu).
extern "C" int nori_gpu_obj_dedup(nori_gpu_ctx *ctx, const uint32_t *corners, uint32_t n_corners, const float *positions, uint32_t n_positions,
                                  const float *texcoords, uint32_t n_texcoords, const float *normals, uint32_t n_normals,
                                  uint32_t *F, float *V, float *N, float *UV, uint32_t *n_vertices, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(corners && n_corners > 0 && n_corners % 3 == 0 && positions && n_positions > 0 && F && V && n_vertices, "obj_dedup: bad arguments");
    REQUIRE((n_texcoords == 0) == (UV == nullptr) && (n_normals == 0) == (N == nullptr), "obj_dedup: UV / N outputs must be given exactly when the file has vt / vn lines");
    REQUIRE((n_texcoords == 0 || texcoords) && (n_normals == 0 || normals), "obj_dedup: missing texcoord / normal array");
    REQUIRE(n_corners < (1u << 31), "obj_dedup: too many corners");
    CK(cudaSetDevice(ctx->device));
    const size_t n = n_corners;
    DevBuf<uint32_t> dC, dF; DevBuf<float> dP, dT, dNr, dV, dN, dUV;
    CK(dC.alloc(3 * n)); CK(dF.alloc(n)); CK(dP.alloc(3 * (size_t) n_positions)); CK(dV.alloc(3 * n));
    if (n_texcoords) { CK(dT.alloc(2 * (size_t) n_texcoords)); CK(dUV.alloc(2 * n)); }
    if (n_normals) { CK(dNr.alloc(3 * (size_t) n_normals)); CK(dN.alloc(3 * n)); }
    cudaStream_t st = ctx->stream;
    if (int r = stagedCopy(ctx, dC.p, corners, 12 * n)) return r;
    if (int r = stagedCopy(ctx, dP.p, positions, 12 * (size_t) n_positions)) return r;
    if (n_texcoords) if (int r = stagedCopy(ctx, dT.p, texcoords, 8 * (size_t) n_texcoords)) return r;
    if (n_normals) if (int r = stagedCopy(ctx, dNr.p, normals, 12 * (size_t) n_normals)) return r;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    uint32_t nV = 0;
    if (!objDedup(dC.p, n_corners, dP.p, n_positions, dT.p, n_texcoords, dNr.p, n_normals, dF.p, dV.p, dN.p, dUV.p, &nV, st, ctx->err)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    CK(cudaMemcpyAsync(F, dF.p, 4 * n, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(V, dV.p, 12 * (size_t) nV, cudaMemcpyDeviceToHost, st));
    if (n_normals) CK(cudaMemcpyAsync(N, dN.p, 12 * (size_t) nV, cudaMemcpyDeviceToHost, st));
    if (n_texcoords) CK(cudaMemcpyAsync(UV, dUV.p, 8 * (size_t) nV, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    *n_vertices = nV;
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    return NORI_GPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Path-graph post-process, first slice (src/pbsdf.cu): vertex records out of the tracer, hash grid + kNN.
extern "C" int nori_gpu_path_vertices_lights(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *d, NoriGpuSPoint *out, NoriGpuLPoint *lights, uint64_t capacity,
                                             uint64_t *n_vertices) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(d && n_vertices, "path_vertices: bad arguments");
    REQUIRE(ctx->haveBvh && ctx->haveCamera, "path_vertices: scene, camera and accel must be set first");
    REQUIRE(d->sample_last > d->sample_first && d->sample_last <= 0x7fffffffu, "path_vertices: bad sample range");
    REQUIRE(d->integrator == NORI_INT_PATH_MATS || d->integrator == NORI_INT_PATH_EMS || d->integrator == NORI_INT_PATH_MIS,
            "path_vertices: the path graph is built from path_mats / path_ems / path_mis paths");
    CK(cudaSetDevice(ctx->device));
    const uint32_t spp = d->sample_last - d->sample_first;
    const unsigned long long nPaths = (unsigned long long) ctx->cam.width * ctx->cam.height * spp;
    REQUIRE(nPaths < (1ull << 31), "path_vertices: more than 2^31 camera samples in one call");
    RenderParams rp; memset(&rp, 0, sizeof(rp));
    rp.integrator = d->integrator; rp.streamMode = NORI_STREAM_PER_SAMPLE; rp.sampleFirst = d->sample_first; rp.sampleCount = spp;
    rp.rrStart = d->rr_start; rp.rrMax = d->rr_max; rp.maxDepth = d->max_depth;
    DeviceScene sc = deviceScene(ctx); sc.isectMode = d->intersect_mode;
    DevBuf<unsigned int> counts; DevBuf<unsigned long long> offsets;
    CK(counts.alloc(nPaths)); CK(offsets.alloc(nPaths + 1));
    cudaStream_t st = ctx->stream;
    enqueuePathVertices(sc, rp, spp, nPaths, counts.p, nullptr, nullptr, st);
    enqueueScanCounts(counts.p, nPaths, offsets.p, st);
    unsigned long long total = 0;
    CK(cudaMemcpyAsync(&total, offsets.p + nPaths, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    *n_vertices = total;
    if (!out) return NORI_GPU_OK;
    if (total > capacity) { ctx->err = "path_vertices: " + std::to_string(total) + " vertices do not fit the given capacity of " + std::to_string(capacity); return NORI_GPU_EINVAL; }
    if (total == 0) return NORI_GPU_OK;
    DevBuf<NoriGpuSPoint> recs; DevBuf<NoriGpuLPoint> lrecs; CK(recs.alloc(total));
    if (lights) CK(lrecs.alloc(total));
    enqueuePathVertices(sc, rp, spp, nPaths, counts.p, offsets.p, recs.p, st, lights ? lrecs.p : nullptr);
    CK(cudaMemcpyAsync(out, recs.p, total * sizeof(NoriGpuSPoint), cudaMemcpyDeviceToHost, st));
    if (lights) CK(cudaMemcpyAsync(lights, lrecs.p, total * sizeof(NoriGpuLPoint), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    return NORI_GPU_OK;
}
extern "C" int nori_gpu_path_vertices(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *d, NoriGpuSPoint *out, uint64_t capacity, uint64_t *n_vertices) {
    return nori_gpu_path_vertices_lights(ctx, d, out, nullptr, capacity, n_vertices);
}

namespace {
/// SPoints::BuildKNN's grid (shadingPoint.h:794-801) over the box [lo, hi] of n points; dim == NULL: PathGraph::computeDimensions.
int pathgraphGridFromBox(const float lo[3], const float hi[3], int n, const int32_t dim[3], int k, int dk, NoriGpuHashParams *out) {
    int dm[3];
    if (dim) { dm[0] = dim[0]; dm[1] = dim[1]; dm[2] = dim[2]; }
    else {   // PathGraph::computeDimensions (src/pathgraph.cpp:40-50): about n cells for a cubic box, proportional to the extents
        const float ext[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
        const float longest = std::max(ext[0], std::max(ext[1], ext[2]));
        const float dd = (float) (std::pow((double) n, 1.0 / 3.0) + 1.0);
        for (int c = 0; c < 3; ++c) dm[c] = (int) ((longest > 0 ? ext[c] / longest : 0.f) * dd + 1.0f);
    }
    for (int c = 0; c < 3; ++c) {
        if (dm[c] < 1) return NORI_GPU_EINVAL;
        out->dim[c] = dm[c]; out->minb[c] = lo[c];
        out->cellsize[c] = (hi[c] - lo[c]) / (float) dm[c];              // SPoints::BuildKNN, include/nori/shadingPoint.h:794
        if (!(out->cellsize[c] > 0.f)) out->cellsize[c] = 1.0f;          // flat point set: one layer of unit cells
    }
    out->k = k; out->dk = dk;
    return NORI_GPU_OK;
}
}

extern "C" int nori_gpu_pathgraph_grid(const void *points, uint64_t stride, int n, const int32_t dim[3], int k, int dk, NoriGpuHashParams *out) {
    if (!points || n <= 0 || stride < 12 || !out) return NORI_GPU_EINVAL;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < n; ++i) {
        float p[3]; memcpy(p, static_cast<const unsigned char *>(points) + (size_t) i * stride, 12);
        for (int c = 0; c < 3; ++c) { lo[c] = std::min(lo[c], p[c]); hi[c] = std::max(hi[c], p[c]); }
    }
    return pathgraphGridFromBox(lo, hi, n, dim, k, dk, out);
}

extern "C" int nori_gpu_pathgraph_knn(nori_gpu_ctx *ctx, const void *points, uint64_t stride, int n, const NoriGpuHashParams *hp,
                                      int32_t *neighbors, int32_t *knn, int32_t *dknn, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(points && hp && neighbors && knn && dknn && n > 0 && stride >= 12 && stride % 4 == 0, "pathgraph_knn: bad arguments");
    REQUIRE(hp->k >= 1 && hp->dk >= 1 && hp->dim[0] >= 1 && hp->dim[1] >= 1 && hp->dim[2] >= 1, "pathgraph_knn: bad grid parameters");
    const unsigned long long cells = (unsigned long long) hp->dim[0] * hp->dim[1] * hp->dim[2];
    REQUIRE(cells < (1ull << 30), "pathgraph_knn: more than 2^30 grid cells");
    REQUIRE(hp->cellsize[0] > 0 && hp->cellsize[1] > 0 && hp->cellsize[2] > 0, "pathgraph_knn: cell sizes must be positive");
    CK(cudaSetDevice(ctx->device));
    PgParams P;
    for (int c = 0; c < 3; ++c) { P.minb[c] = hp->minb[c]; P.cell[c] = hp->cellsize[c]; P.dim[c] = hp->dim[c]; }
    P.k = hp->k; P.dk = hp->dk; P.maxK = std::max(hp->k, hp->dk); P.nCells = (int) cells;
    REQUIRE(P.maxK <= kPgMaxK, "pathgraph_knn: max(k, dk) above 200 is not supported");
    DevBuf<unsigned char> pts, scratch; DevBuf<int> nb, dk1, dk2;
    CK(pts.alloc((size_t) n * stride)); CK(scratch.alloc(pgScratchBytes(n, P.nCells)));
    CK(nb.alloc((size_t) n * P.maxK)); CK(dk1.alloc(n)); CK(dk2.alloc(n));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(pts.p, points, (size_t) n * stride, cudaMemcpyHostToDevice, st));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    if (!launchPathgraphKnn(pts.p, (size_t) stride, n, P, scratch.p, nb.p, P.maxK, dk1.p, dk2.p, st, ctx->err)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    CK(cudaMemcpyAsync(neighbors, nb.p, (size_t) n * P.maxK * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(knn, dk1.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(dknn, dk2.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    unsigned line = 0;
    if (const unsigned bad = pathgraphChecks(&line)) { ctx->err = "pathgraph_knn: device range check failed (" + std::to_string(bad) + " at pathgraph.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}

extern "C" int nori_gpu_pathgraph_stride_seeds(int n, int n_seeds, int32_t *seeds) {
    if (n <= 0 || n_seeds <= 0 || n_seeds > n || !seeds) return NORI_GPU_EINVAL;
    const long long step = pathgraphSeedStep(n);
    for (int s = 0; s < n_seeds; ++s) seeds[s] = (int32_t) (((long long) s * step) % n);
    return NORI_GPU_OK;
}

extern "C" int nori_gpu_pathgraph_build_clusters(nori_gpu_ctx *ctx, const void *points, uint64_t stride, int n, const NoriGpuHashParams *hp, const int32_t *seeds,
                                                 int n_seeds, int32_t *cluster_of, int32_t *clusters, int32_t *offsets, int offsets_capacity, int *n_clusters,
                                                 int *n_listed, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(points && hp && seeds && cluster_of && clusters && offsets && n_clusters && n_listed && n > 0 && stride >= 12 && stride % 4 == 0, "pathgraph_build_clusters: bad arguments");
    REQUIRE(hp->k >= 2 && hp->dim[0] >= 1 && hp->dim[1] >= 1 && hp->dim[2] >= 1, "pathgraph_build_clusters: bad grid parameters (k = members per cluster >= 2)");
    REQUIRE(n_seeds >= 1 && n_seeds <= n && offsets_capacity >= n_seeds, "pathgraph_build_clusters: need 1 <= n_seeds <= n and offsets_capacity >= n_seeds");
    const unsigned long long cells = (unsigned long long) hp->dim[0] * hp->dim[1] * hp->dim[2];
    REQUIRE(cells < (1ull << 30), "pathgraph_build_clusters: more than 2^30 grid cells");
    REQUIRE(hp->cellsize[0] > 0 && hp->cellsize[1] > 0 && hp->cellsize[2] > 0, "pathgraph_build_clusters: cell sizes must be positive");
    for (int s = 0; s < n_seeds; ++s) REQUIRE(seeds[s] >= 0 && seeds[s] < n, "pathgraph_build_clusters: seed outside [0, n)");
    CK(cudaSetDevice(ctx->device));
    PgParams P;
    for (int c = 0; c < 3; ++c) { P.minb[c] = hp->minb[c]; P.cell[c] = hp->cellsize[c]; P.dim[c] = hp->dim[c]; }
    P.k = hp->k; P.dk = hp->k; P.maxK = hp->k; P.nCells = (int) cells;
    DevBuf<unsigned char> pts, scratch; DevBuf<int> sd, g, lists, off;
    CK(pts.alloc((size_t) n * stride)); CK(scratch.alloc(pgBuildClustersScratchBytes(n, n_seeds, offsets_capacity)));
    CK(sd.alloc(n_seeds)); CK(g.alloc(n)); CK(lists.alloc(n)); CK(off.alloc((size_t) offsets_capacity + 1));
    cudaStream_t st = ctx->stream;
    if (int rc = stagedCopy(ctx, pts.p, points, (size_t) n * stride)) return rc;
    CK(cudaMemcpyAsync(sd.p, seeds, (size_t) n_seeds * 4, cudaMemcpyHostToDevice, st));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    int nc = 0, nl = 0;
    if (!launchPathgraphBuildClusters(pts.p, (size_t) stride, n, P, sd.p, n_seeds, hp->k, scratch.p, g.p, lists.p, off.p, offsets_capacity, &nc, &nl, st, ctx->err)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    CK(cudaMemcpyAsync(cluster_of, g.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    if (nl > 0) CK(cudaMemcpyAsync(clusters, lists.p, (size_t) nl * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(offsets, off.p, (size_t) nc * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    *n_clusters = nc; *n_listed = nl;
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    unsigned line = 0;
    if (const unsigned bad = pathgraphChecks(&line)) { ctx->err = "pathgraph_build_clusters: device range check failed (" + std::to_string(bad) + " at pathgraph.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}

namespace {
/// Device side of the scatter stage shared by the two entry points: rows already on the device and validated.
int pathgraphScatterOnDevice(nori_gpu_ctx *ctx, int mode, const NoriGpuSPoint *d_sps, int n, const int *d_nb, int rowInts, const int *d_knn, int iterations,
                             float *pdfmarginal, float *blur, float *mc, float *ms) {
    DevBuf<unsigned char> scratch; DevBuf<float> pm, db, dm;
    CK(scratch.alloc(mode == NORI_PG_MIS ? pgMisScratchBytes(n, rowInts) : pgScatterScratchBytes(n))); CK(pm.alloc(n));
    const size_t resultFloats = (size_t) std::max(iterations, 1) * 3 * (size_t) n;
    CK(db.alloc(resultFloats)); CK(dm.alloc(resultFloats));
    cudaStream_t st = ctx->stream;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    if (mode == NORI_PG_MIS) { if (!launchPathgraphMis(d_sps, n, d_nb, rowInts, d_knn, iterations, scratch.p, db.p, dm.p, st, ctx->err)) return NORI_GPU_EINVAL; }
    else if (!launchPathgraphScatter(d_sps, n, d_nb, rowInts, d_knn, iterations, scratch.p, pm.p, db.p, dm.p, st, ctx->err)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    if (pdfmarginal && mode != NORI_PG_MIS) CK(cudaMemcpyAsync(pdfmarginal, pm.p, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    if (iterations > 0) {
        CK(cudaMemcpyAsync(blur, db.p, (size_t) iterations * 3 * n * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(mc, dm.p, (size_t) iterations * 3 * n * 4, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    unsigned line = 0;
    if (const unsigned bad = pathgraphScatterChecks(&line)) { ctx->err = "pathgraph_scatter: device range check failed (" + std::to_string(bad) + " at pathgraph_scatter.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}
}

namespace {
int pathgraphFromHostRows(nori_gpu_ctx *ctx, int mode, const NoriGpuSPoint *points, int n, const int32_t *neighbors, int row_ints,
                          const int32_t *knn, int iterations, float *pdfmarginal, float *blur, float *mc, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(points && neighbors && knn && n > 0 && row_ints >= 1 && iterations >= 0, "pathgraph_scatter: bad arguments");
    REQUIRE(iterations == 0 || (blur && mc), "pathgraph_scatter: blur and mc must hold iterations x n x 3 floats");
    REQUIRE(iterations <= 64, "pathgraph_scatter: more than 64 iterations");
    CK(cudaSetDevice(ctx->device));
    DevBuf<NoriGpuSPoint> sps; DevBuf<int> nb, kn, flag;
    CK(sps.alloc(n)); CK(nb.alloc((size_t) n * row_ints)); CK(kn.alloc(n)); CK(flag.alloc(1));
    cudaStream_t st = ctx->stream;
    if (int rc = stagedCopy(ctx, sps.p, points, (size_t) n * sizeof(NoriGpuSPoint))) return rc;
    if (int rc = stagedCopy(ctx, nb.p, neighbors, (size_t) n * row_ints * 4)) return rc;
    CK(cudaMemcpyAsync(kn.p, knn, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    // the kernels trust the rows: checked once, on the device copy
    launchPathgraphValidate(nb.p, row_ints, row_ints, kn.p, n, flag.p, st);
    int bad = 0;
    CK(cudaMemcpyAsync(&bad, flag.p, sizeof(int), cudaMemcpyDeviceToHost, st)); CK(cudaStreamSynchronize(st));
    REQUIRE(!(bad & 1), "pathgraph_scatter: knn[i] outside [0, row_ints]");
    REQUIRE(!(bad & 2), "pathgraph_scatter: neighbour index outside [0, n)");
    return pathgraphScatterOnDevice(ctx, mode, sps.p, n, nb.p, row_ints, kn.p, iterations, pdfmarginal, blur, mc, ms);
}
}

extern "C" int nori_gpu_pathgraph_scatter(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, int n, const int32_t *neighbors, int row_ints,
                                          const int32_t *knn, int iterations, float *pdfmarginal, float *blur, float *mc, float *ms) {
    return pathgraphFromHostRows(ctx, NORI_PG_SCATTER, points, n, neighbors, row_ints, knn, iterations, pdfmarginal, blur, mc, ms);
}
extern "C" int nori_gpu_pathgraph_mis(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, int n, const int32_t *neighbors, int row_ints,
                                      const int32_t *knn, int iterations, float *blur, float *mc, float *ms) {
    return pathgraphFromHostRows(ctx, NORI_PG_MIS, points, n, neighbors, row_ints, knn, iterations, nullptr, blur, mc, ms);
}

namespace {
/// The direct-light filter alone (iterations < 0) or followed by the indirect scatter that uses it (SPoints::
/// computeScatterRadianceAOGWithProcessRecordingAndDirectLightOPT, include/nori/shadingPoint.h:462-476).
int pathgraphDirectFromHost(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, const NoriGpuLPoint *lights, int n, const int32_t *neighbors, int row_ints,
                            const int32_t *knn, const int32_t *dknn, int iterations, float *direct, float *blur, float *mc, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    const bool indirect = iterations >= 0;
    REQUIRE(points && lights && neighbors && dknn && direct && n > 0 && row_ints >= 1, "pathgraph_direct: bad arguments");
    REQUIRE(!indirect || (knn && iterations <= 64 && (iterations == 0 || (blur && mc))), "pathgraph_direct: knn, blur and mc (iterations x n x 3 floats, <= 64 iterations)");
    CK(cudaSetDevice(ctx->device));
    DevBuf<NoriGpuSPoint> sps; DevBuf<NoriGpuLPoint> lps; DevBuf<int> nb, kn, dkn, flag; DevBuf<unsigned char> scratch; DevBuf<float> out, pm, db, dm;
    CK(sps.alloc(n)); CK(lps.alloc(n)); CK(nb.alloc((size_t) n * row_ints)); CK(dkn.alloc(n)); CK(flag.alloc(1));
    CK(scratch.alloc(std::max(pgDirectScratchBytes(n), pgScatterScratchBytes(n)))); CK(out.alloc((size_t) 3 * n));
    cudaStream_t st = ctx->stream;
    if (int rc = stagedCopy(ctx, sps.p, points, (size_t) n * sizeof(NoriGpuSPoint))) return rc;
    if (int rc = stagedCopy(ctx, lps.p, lights, (size_t) n * sizeof(NoriGpuLPoint))) return rc;
    if (int rc = stagedCopy(ctx, nb.p, neighbors, (size_t) n * row_ints * 4)) return rc;
    CK(cudaMemcpyAsync(dkn.p, dknn, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    launchPathgraphValidate(nb.p, row_ints, row_ints, dkn.p, n, flag.p, st);
    int bad = 0;
    CK(cudaMemcpyAsync(&bad, flag.p, sizeof(int), cudaMemcpyDeviceToHost, st)); CK(cudaStreamSynchronize(st));
    REQUIRE(!(bad & 1), "pathgraph_direct: dknn[i] outside [0, row_ints]");
    REQUIRE(!(bad & 2), "pathgraph_direct: neighbour index outside [0, n)");
    if (indirect) {
        CK(kn.alloc(n)); CK(pm.alloc(n));
        const size_t resultFloats = (size_t) std::max(iterations, 1) * 3 * (size_t) n;
        CK(db.alloc(resultFloats)); CK(dm.alloc(resultFloats));
        CK(cudaMemcpyAsync(kn.p, knn, (size_t) n * 4, cudaMemcpyHostToDevice, st));
        launchPathgraphValidate(nb.p, row_ints, row_ints, kn.p, n, flag.p, st);
        CK(cudaMemcpyAsync(&bad, flag.p, sizeof(int), cudaMemcpyDeviceToHost, st)); CK(cudaStreamSynchronize(st));
        REQUIRE(!(bad & 1), "pathgraph_direct: knn[i] outside [0, row_ints]");
        REQUIRE(!(bad & 2), "pathgraph_direct: neighbour index outside [0, n)");
    }
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    if (!launchPathgraphDirect(sps.p, lps.p, n, nb.p, row_ints, dkn.p, scratch.p, out.p, st, ctx->err)) return NORI_GPU_EINVAL;
    if (indirect && !launchPathgraphScatter(sps.p, n, nb.p, row_ints, kn.p, iterations, scratch.p, pm.p, db.p, dm.p, st, ctx->err, out.p)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    CK(cudaMemcpyAsync(direct, out.p, (size_t) 3 * n * 4, cudaMemcpyDeviceToHost, st));
    if (indirect && iterations > 0) {
        CK(cudaMemcpyAsync(blur, db.p, (size_t) iterations * 3 * n * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(mc, dm.p, (size_t) iterations * 3 * n * 4, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    unsigned line = 0;
    if (const unsigned b = pathgraphScatterChecks(&line)) { ctx->err = "pathgraph_direct: device range check failed (" + std::to_string(b) + " at pathgraph_scatter.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}
}

extern "C" int nori_gpu_pathgraph_direct(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, const NoriGpuLPoint *lights, int n, const int32_t *neighbors,
                                         int row_ints, const int32_t *dknn, float *direct, float *ms) {
    return pathgraphDirectFromHost(ctx, points, lights, n, neighbors, row_ints, nullptr, dknn, -1, direct, nullptr, nullptr, ms);
}
extern "C" int nori_gpu_pathgraph_scatter_direct(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, const NoriGpuLPoint *lights, int n, const int32_t *neighbors,
                                                 int row_ints, const int32_t *knn, const int32_t *dknn, int iterations, float *direct, float *blur, float *mc,
                                                 float *ms) {
    if (ctx) REQUIRE(iterations >= 0, "pathgraph_scatter_direct: iterations < 0");
    return pathgraphDirectFromHost(ctx, points, lights, n, neighbors, row_ints, knn, dknn, iterations, direct, blur, mc, ms);
}

extern "C" int nori_gpu_pathgraph_cluster(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, const NoriGpuLPoint *lights, int n, const int32_t *clusters,
                                          int n_listed, const int32_t *offsets, int n_clusters, int iterations, float *direct, float *blur, float *mc, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(points && lights && direct && blur && mc && n > 0, "pathgraph_cluster: bad arguments");
    REQUIRE(n_clusters >= 0 && n_listed >= 0 && n_listed <= n && (n_clusters == 0 ? n_listed == 0 : (clusters && offsets)), "pathgraph_cluster: bad cluster lists");
    REQUIRE(iterations >= 1 && iterations <= 64, "pathgraph_cluster: iterations outside [1, 64]");
    CK(cudaSetDevice(ctx->device));
    DevBuf<NoriGpuSPoint> sps; DevBuf<NoriGpuLPoint> lps; DevBuf<int> cl, off, listed, flag; DevBuf<unsigned char> scratch; DevBuf<float> out;
    CK(sps.alloc(n)); CK(lps.alloc(n)); CK(cl.alloc(std::max(n_listed, 1))); CK(off.alloc(std::max(n_clusters, 1))); CK(listed.alloc(n)); CK(flag.alloc(2));
    CK(scratch.alloc(pgClusterScratchBytes(n, n_clusters))); CK(out.alloc((size_t) 9 * n));
    cudaStream_t st = ctx->stream;
    if (int rc = stagedCopy(ctx, sps.p, points, (size_t) n * sizeof(NoriGpuSPoint))) return rc;
    if (int rc = stagedCopy(ctx, lps.p, lights, (size_t) n * sizeof(NoriGpuLPoint))) return rc;
    if (n_listed > 0) if (int rc = stagedCopy(ctx, cl.p, clusters, (size_t) n_listed * 4)) return rc;
    if (n_clusters > 0) CK(cudaMemcpyAsync(off.p, offsets, (size_t) n_clusters * 4, cudaMemcpyHostToDevice, st));
    // the kernels trust the lists (a partition of the vertices that name a cluster): checked once, on the device copy
    launchPathgraphClusterValidate(sps.p, n, cl.p, n_listed, off.p, n_clusters, listed.p, flag.p, st);
    int chk[2] = {0, 0};
    CK(cudaMemcpyAsync(chk, flag.p, sizeof chk, cudaMemcpyDeviceToHost, st)); CK(cudaStreamSynchronize(st));
    REQUIRE(!(chk[0] & 1), "pathgraph_cluster: offsets must start at 0 and ascend inside [0, n_listed]");
    REQUIRE(!(chk[0] & 32), "pathgraph_cluster: a cluster of more than 8192 members");
    REQUIRE(!(chk[0] & 2), "pathgraph_cluster: cluster member outside [0, n)");
    REQUIRE(!(chk[0] & 16), "pathgraph_cluster: groupIdx outside [-1, n_clusters)");
    REQUIRE(!(chk[0] & 4), "pathgraph_cluster: a member's groupIdx is not the cluster that lists it");
    REQUIRE(!(chk[0] & 8), "pathgraph_cluster: a vertex is listed twice");
    REQUIRE(chk[1] == n_listed, "pathgraph_cluster: a vertex names a cluster that does not list it");
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    if (!launchPathgraphCluster(sps.p, lps.p, n, cl.p, n_listed, off.p, n_clusters, iterations, scratch.p, out.p, out.p + (size_t) 3 * n, out.p + (size_t) 6 * n, st, ctx->err))
        return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    CK(cudaMemcpyAsync(direct, out.p, (size_t) 3 * n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(blur, out.p + (size_t) 3 * n, (size_t) 3 * n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(mc, out.p + (size_t) 6 * n, (size_t) 3 * n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    unsigned line = 0;
    if (const unsigned b = pathgraphScatterChecks(&line)) { ctx->err = "pathgraph_cluster: device range check failed (" + std::to_string(b) + " at pathgraph_scatter.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}

extern "C" int nori_gpu_pathgraph_cluster_scatter(nori_gpu_ctx *ctx, int order, const NoriGpuSPoint *points, int n, const int32_t *clusters, int n_listed,
                                                  const int32_t *offsets, int n_clusters, int iterations, float *blur, float *mc, float *ms) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(points && blur && mc && n > 0, "pathgraph_cluster_scatter: bad arguments");
    REQUIRE(order == NORI_PG_CLUSTER_ELEMENTS || order == NORI_PG_CLUSTER_SCATTER, "pathgraph_cluster_scatter: unknown order");
    REQUIRE(n_clusters >= 0 && n_listed >= 0 && n_listed <= n && (n_clusters == 0 ? n_listed == 0 : (clusters && offsets)), "pathgraph_cluster_scatter: bad cluster lists");
    REQUIRE(iterations >= 1 && iterations <= 64, "pathgraph_cluster_scatter: iterations outside [1, 64]");
    CK(cudaSetDevice(ctx->device));
    DevBuf<NoriGpuSPoint> sps; DevBuf<int> cl, off, listed, flag; DevBuf<unsigned char> scratch; DevBuf<float> out;
    CK(sps.alloc(n)); CK(cl.alloc(std::max(n_listed, 1))); CK(off.alloc(std::max(n_clusters, 1))); CK(listed.alloc(n)); CK(flag.alloc(2));
    CK(scratch.alloc(pgClusterScratchBytes(n, n_clusters))); CK(out.alloc((size_t) 6 * n * iterations));
    cudaStream_t st = ctx->stream;
    if (int rc = stagedCopy(ctx, sps.p, points, (size_t) n * sizeof(NoriGpuSPoint))) return rc;
    if (n_listed > 0) if (int rc = stagedCopy(ctx, cl.p, clusters, (size_t) n_listed * 4)) return rc;
    if (n_clusters > 0) CK(cudaMemcpyAsync(off.p, offsets, (size_t) n_clusters * 4, cudaMemcpyHostToDevice, st));
    launchPathgraphClusterValidate(sps.p, n, cl.p, n_listed, off.p, n_clusters, listed.p, flag.p, st);
    int chk[2] = {0, 0};
    CK(cudaMemcpyAsync(chk, flag.p, sizeof chk, cudaMemcpyDeviceToHost, st)); CK(cudaStreamSynchronize(st));
    REQUIRE(!(chk[0] & 1), "pathgraph_cluster_scatter: offsets must start at 0 and ascend inside [0, n_listed]");
    REQUIRE(!(chk[0] & 32), "pathgraph_cluster_scatter: a cluster of more than 8192 members");
    REQUIRE(!(chk[0] & 2), "pathgraph_cluster_scatter: cluster member outside [0, n)");
    REQUIRE(!(chk[0] & 16), "pathgraph_cluster_scatter: groupIdx outside [-1, n_clusters)");
    REQUIRE(!(chk[0] & 4), "pathgraph_cluster_scatter: a member's groupIdx is not the cluster that lists it");
    REQUIRE(!(chk[0] & 8), "pathgraph_cluster_scatter: a vertex is listed twice");
    REQUIRE(chk[1] == n_listed, "pathgraph_cluster_scatter: a vertex names a cluster that does not list it");
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    const size_t rows = (size_t) 3 * n * iterations;
    if (!launchPathgraphClusterScatter(sps.p, n, cl.p, n_listed, off.p, n_clusters, iterations, order, scratch.p, out.p, out.p + rows, st, ctx->err)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    CK(cudaMemcpyAsync(blur, out.p, rows * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(mc, out.p + rows, rows * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    unsigned line = 0;
    if (const unsigned b = pathgraphScatterChecks(&line)) { ctx->err = "pathgraph_cluster_scatter: device range check failed (" + std::to_string(b) + " at pathgraph_scatter.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}

extern "C" int nori_gpu_pathgraph_filter(nori_gpu_ctx *ctx, int mode, const NoriGpuSPoint *points, int n, const NoriGpuHashParams *hp, int iterations,
                                         float *pdfmarginal, float *blur, float *mc, float ms[2]) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(mode == NORI_PG_SCATTER || mode == NORI_PG_MIS, "pathgraph_filter: unknown mode");
    REQUIRE(points && hp && n > 0 && iterations >= 0 && iterations <= 64, "pathgraph_filter: bad arguments");
    REQUIRE(iterations == 0 || (blur && mc), "pathgraph_filter: blur and mc must hold iterations x n x 3 floats");
    REQUIRE(hp->k >= 1 && hp->dk >= 1 && hp->dim[0] >= 1 && hp->dim[1] >= 1 && hp->dim[2] >= 1, "pathgraph_filter: bad grid parameters");
    const unsigned long long cells = (unsigned long long) hp->dim[0] * hp->dim[1] * hp->dim[2];
    REQUIRE(cells < (1ull << 30), "pathgraph_filter: more than 2^30 grid cells");
    REQUIRE(hp->cellsize[0] > 0 && hp->cellsize[1] > 0 && hp->cellsize[2] > 0, "pathgraph_filter: cell sizes must be positive");
    CK(cudaSetDevice(ctx->device));
    PgParams P;
    for (int c = 0; c < 3; ++c) { P.minb[c] = hp->minb[c]; P.cell[c] = hp->cellsize[c]; P.dim[c] = hp->dim[c]; }
    P.k = hp->k; P.dk = hp->dk; P.maxK = std::max(hp->k, hp->dk); P.nCells = (int) cells;
    REQUIRE(P.maxK <= kPgMaxK, "pathgraph_filter: max(k, dk) above 200 is not supported");
    DevBuf<NoriGpuSPoint> sps; DevBuf<unsigned char> scratch; DevBuf<int> nb, dk1, dk2;
    CK(sps.alloc(n)); CK(scratch.alloc(pgScratchBytes(n, P.nCells))); CK(nb.alloc((size_t) n * P.maxK)); CK(dk1.alloc(n)); CK(dk2.alloc(n));
    cudaStream_t st = ctx->stream;
    if (int rc = stagedCopy(ctx, sps.p, points, (size_t) n * sizeof(NoriGpuSPoint))) return rc;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Ev { cudaEvent_t a, b; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); } } evGuard{e0, e1};
    CK(cudaEventRecord(e0, st));
    if (!launchPathgraphKnn(sps.p, sizeof(NoriGpuSPoint), n, P, scratch.p, nb.p, P.maxK, dk1.p, dk2.p, st, ctx->err)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(e1, st));
    float msScatter = 0.f;
    if (int rc = pathgraphScatterOnDevice(ctx, mode, sps.p, n, nb.p, P.maxK, dk1.p, iterations, pdfmarginal, blur, mc, &msScatter)) return rc;
    if (ms) { cudaEventElapsedTime(&ms[0], e0, e1); ms[1] = msScatter; }
    unsigned line = 0;
    if (const unsigned bad = pathgraphChecks(&line)) { ctx->err = "pathgraph_filter: device range check failed (" + std::to_string(bad) + " at pathgraph.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}

extern "C" int nori_gpu_pathgraph_render(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *d, int mode, int k, int iterations, float *images,
                                         NoriGpuPathGraphStats *stats) {
    if (!ctx) return NORI_GPU_EINVAL;
    REQUIRE(d && images, "pathgraph_render: bad arguments");
    REQUIRE(mode == NORI_PG_SCATTER || mode == NORI_PG_MIS || mode == NORI_PG_CLUSTER, "pathgraph_render: unknown mode");
    REQUIRE(mode != NORI_PG_CLUSTER || k >= 2, "pathgraph_render: clusters need k >= 2 members");
    REQUIRE(k >= 1 && k <= kPgMaxK && iterations >= 1 && iterations <= 64, "pathgraph_render: k in [1, 200], iterations in [1, 64]");
    REQUIRE(ctx->haveBvh && ctx->haveCamera, "pathgraph_render: scene, camera and accel must be set first");
    REQUIRE(d->sample_last > d->sample_first && d->sample_last <= 0x7fffffffu, "pathgraph_render: bad sample range");
    REQUIRE(d->integrator == NORI_INT_PATH_MATS || d->integrator == NORI_INT_PATH_EMS || d->integrator == NORI_INT_PATH_MIS,
            "pathgraph_render: the path graph is built from path_mats / path_ems / path_mis paths");
    CK(cudaSetDevice(ctx->device));
    const uint32_t spp = d->sample_last - d->sample_first;
    const int nPixels = ctx->cam.width * ctx->cam.height;
    const unsigned long long nPaths = (unsigned long long) nPixels * spp;
    REQUIRE(nPaths < (1ull << 31), "pathgraph_render: more than 2^31 camera samples in one call");
    RenderParams rp; memset(&rp, 0, sizeof(rp));
    rp.integrator = d->integrator; rp.streamMode = NORI_STREAM_PER_SAMPLE; rp.sampleFirst = d->sample_first; rp.sampleCount = spp;
    rp.rrStart = d->rr_start; rp.rrMax = d->rr_max; rp.maxDepth = d->max_depth;
    DeviceScene sc = deviceScene(ctx); sc.isectMode = d->intersect_mode;
    cudaStream_t st = ctx->stream;
    cudaEvent_t ev[4];
    for (auto &e : ev) CK(cudaEventCreate(&e));
    struct Ev { cudaEvent_t *e; ~Ev() { for (int i = 0; i < 4; ++i) cudaEventDestroy(e[i]); } } evGuard{ev};
    // 1. the paths' vertices, never leaving the device: count, scan, write
    DevBuf<unsigned int> counts, bounds; DevBuf<unsigned long long> offsets;
    CK(counts.alloc(nPaths)); CK(offsets.alloc(nPaths + 1)); CK(bounds.alloc(6));
    CK(cudaEventRecord(ev[0], st));
    enqueuePathVertices(sc, rp, spp, nPaths, counts.p, nullptr, nullptr, st);
    enqueueScanCounts(counts.p, nPaths, offsets.p, st);
    unsigned long long total = 0;
    CK(cudaMemcpyAsync(&total, offsets.p + nPaths, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    REQUIRE(total > 0, "pathgraph_render: no camera sample hit a surface");
    REQUIRE(total <= 0x7fffffffull, "pathgraph_render: more than 2^31 path vertices; render fewer samples per call");
    const int n = (int) total;
    DevBuf<NoriGpuSPoint> recs; CK(recs.alloc(total));
    DevBuf<NoriGpuLPoint> lrecs;
    if (mode == NORI_PG_CLUSTER) CK(lrecs.alloc(total));             // the clustered mode filters direct light too: light records beside the vertices
    enqueuePathVertices(sc, rp, spp, nPaths, counts.p, offsets.p, recs.p, st, mode == NORI_PG_CLUSTER ? lrecs.p : nullptr);
    // 2. grid as SPoints::BuildKNN / PathGraph::computeDimensions size it, from the records' bounding box
    launchPathgraphBounds(recs.p, n, bounds.p, st);
    unsigned int hb[6];
    CK(cudaMemcpyAsync(hb, bounds.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(ev[1], st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    float lo[3], hi[3];
    for (int c = 0; c < 3; ++c) { lo[c] = pathgraphOrderedToFloat(hb[c]); hi[c] = pathgraphOrderedToFloat(hb[3 + c]); }
    NoriGpuHashParams hp;
    REQUIRE(pathgraphGridFromBox(lo, hi, n, nullptr, k, k, &hp) == NORI_GPU_OK, "pathgraph_render: degenerate vertex set");
    PgParams P;
    for (int c = 0; c < 3; ++c) { P.minb[c] = hp.minb[c]; P.cell[c] = hp.cellsize[c]; P.dim[c] = hp.dim[c]; }
    const unsigned long long cells = (unsigned long long) hp.dim[0] * hp.dim[1] * hp.dim[2];
    REQUIRE(cells < (1ull << 30), "pathgraph_render: more than 2^30 grid cells");
    P.k = k; P.dk = k; P.maxK = k; P.nCells = (int) cells;
    // 3. neighbours, 4. scatter or MIS gather, 5. images: all on the records where they lie
    DevBuf<unsigned char> scratch, stage; DevBuf<int> nb, dk1, dk2; DevBuf<float> pm, db, dm, img;
    if (mode == NORI_PG_CLUSTER) {
        // 3'. BuildClusters from n / k + 1 scattered seeds, 4'. ClusterScatterWithDirectOptNR, 5'. the five images + the filtered direct light
        const int nSeeds = n / k + 1, cap = n / k + 2 * nSeeds + 2;
        DevBuf<int> seeds, of, lists, off;
        CK(seeds.alloc(nSeeds)); CK(of.alloc(n)); CK(lists.alloc(n)); CK(off.alloc((size_t) cap + 1));
        CK(scratch.alloc(pgBuildClustersScratchBytes(n, nSeeds, cap)));
        launchPathgraphStrideSeeds(seeds.p, nSeeds, n, st);
        int nc = 0, nl = 0, largest = 0;
        if (!launchPathgraphBuildClusters(recs.p, sizeof(NoriGpuSPoint), n, P, seeds.p, nSeeds, k, scratch.p, of.p, lists.p, off.p, cap, &nc, &nl, st, ctx->err, &largest)) return NORI_GPU_EINVAL;
        REQUIRE(largest <= kPgMaxClusterSize, "pathgraph_render: the builder left a cluster of more than 8192 members (coincident vertices?); use another k");
        CK(cudaEventRecord(ev[2], st));
        CK(stage.alloc(pgClusterScratchBytes(n, nc))); CK(db.alloc((size_t) 9 * n)); CK(img.alloc((size_t) 6 * nPixels * 3));
        if (!launchPathgraphCluster(recs.p, lrecs.p, n, lists.p, nl, off.p, nc, iterations, stage.p, db.p, db.p + (size_t) 3 * n, db.p + (size_t) 6 * n, st, ctx->err, of.p))
            return NORI_GPU_EINVAL;
        launchPathgraphCompose(recs.p, offsets.p, counts.p, db.p + (size_t) 3 * n, db.p + (size_t) 6 * n, nPixels, spp, img.p, st, db.p);
        CK(cudaEventRecord(ev[3], st));
        CK(cudaMemcpyAsync(images, img.p, (size_t) 6 * nPixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
        if (stats) {
            stats->vertices = total;
            cudaEventElapsedTime(&stats->ms_vertices, ev[0], ev[1]); cudaEventElapsedTime(&stats->ms_knn, ev[1], ev[2]); cudaEventElapsedTime(&stats->ms_stage, ev[2], ev[3]);
            for (int c = 0; c < 3; ++c) stats->dim[c] = hp.dim[c];
        }
        unsigned line = 0;
        if (const unsigned bad = pathgraphChecks(&line)) { ctx->err = "pathgraph_render: device range check failed (" + std::to_string(bad) + " at pathgraph.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
        if (const unsigned bad = pathgraphScatterChecks(&line)) { ctx->err = "pathgraph_render: device range check failed (" + std::to_string(bad) + " at pathgraph_scatter.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
        return NORI_GPU_OK;
    }
    CK(scratch.alloc(pgScratchBytes(n, P.nCells))); CK(nb.alloc((size_t) n * k)); CK(dk1.alloc(n)); CK(dk2.alloc(n));
    if (!launchPathgraphKnn(recs.p, sizeof(NoriGpuSPoint), n, P, scratch.p, nb.p, k, dk1.p, dk2.p, st, ctx->err)) return NORI_GPU_EINVAL;
    CK(cudaEventRecord(ev[2], st));
    CK(stage.alloc(mode == NORI_PG_MIS ? pgMisScratchBytes(n, k) : pgScatterScratchBytes(n))); CK(pm.alloc(n));
    CK(db.alloc((size_t) iterations * 3 * n)); CK(dm.alloc((size_t) iterations * 3 * n)); CK(img.alloc((size_t) 5 * nPixels * 3));
    if (mode == NORI_PG_MIS) { if (!launchPathgraphMis(recs.p, n, nb.p, k, dk1.p, iterations, stage.p, db.p, dm.p, st, ctx->err)) return NORI_GPU_EINVAL; }
    else if (!launchPathgraphScatter(recs.p, n, nb.p, k, dk1.p, iterations, stage.p, pm.p, db.p, dm.p, st, ctx->err)) return NORI_GPU_EINVAL;
    launchPathgraphCompose(recs.p, offsets.p, counts.p, db.p + (size_t) (iterations - 1) * 3 * n, dm.p + (size_t) (iterations - 1) * 3 * n, nPixels, spp, img.p, st);
    CK(cudaEventRecord(ev[3], st));
    CK(cudaMemcpyAsync(images, img.p, (size_t) 5 * nPixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    if (stats) {
        stats->vertices = total;
        cudaEventElapsedTime(&stats->ms_vertices, ev[0], ev[1]); cudaEventElapsedTime(&stats->ms_knn, ev[1], ev[2]); cudaEventElapsedTime(&stats->ms_stage, ev[2], ev[3]);
        for (int c = 0; c < 3; ++c) stats->dim[c] = hp.dim[c];
    }
    unsigned line = 0;
    if (const unsigned bad = pathgraphChecks(&line)) { ctx->err = "pathgraph_render: device range check failed (" + std::to_string(bad) + " at pathgraph.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    if (const unsigned bad = pathgraphScatterChecks(&line)) { ctx->err = "pathgraph_render: device range check failed (" + std::to_string(bad) + " at pathgraph_scatter.cu:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}

// ---------------------------------------------------------------------------------------------
// nori_gpu_trace_static: Accel::rayIntersect for callers that must not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc). Everything nori_gpu_trace keeps in cudaMalloc'ed scratch — the device
// copy of the rays, the hit records, the work cursor of the persistent traversal, the split output arrays — lives in
// ONE file-scope device array of fixed size. Every GPU that loads the library holds it (4.4 MB) until exit.
// The array is shared by every context and every call on a device and keeps the previous call's contents: the
// work cursor is cleared at the start of each call, everything else is overwritten before it is read. Two calls
// that overlap on one GPU (different streams / host threads) race on it: callers serialise them.
namespace {

struct TraceStaticScratch {
    unsigned int cursor; unsigned int pad[63];                  // work counter of tracePersistent; must start at zero
    float4 rays[2 * NORI_GPU_TRACE_STATIC_MAX_RAYS];            // (o, mint) (d, maxt)
    float4 hits[NORI_GPU_TRACE_STATIC_MAX_RAYS];                // (t, u, v, bits(gid))
    uint32_t tri[NORI_GPU_TRACE_STATIC_MAX_RAYS], mesh[NORI_GPU_TRACE_STATIC_MAX_RAYS];
    float tuv[3 * NORI_GPU_TRACE_STATIC_MAX_RAYS];
};
static __device__ __align__(256) unsigned char g_traceStaticScratch[sizeof(TraceStaticScratch)];

/// Device address of the array on the CURRENT device. In host code `g_traceStaticScratch` is only a stand-in for the
/// symbol, not a device address: it is resolved once per device, with that device current, and cached.
constexpr int kMaxDevices = 64;
TraceStaticScratch *g_traceStaticPtr[kMaxDevices];
std::mutex g_traceStaticLock;

cudaError_t traceStaticScratch(int device, TraceStaticScratch **out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> guard(g_traceStaticLock);
    if (!g_traceStaticPtr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_traceStaticScratch);     // the caller made `device` current
        if (e != cudaSuccess) return e;
        g_traceStaticPtr[device] = static_cast<TraceStaticScratch *>(p);
    }
    *out = g_traceStaticPtr[device];
    return cudaSuccess;
}

struct StaticRayIO {
    TraceStaticScratch *s; uint32_t n;
    __device__ __forceinline__ bool load(uint32_t i, V3 &o, V3 &d, float &mint, float &maxt) const {
        NG_CHECK(i < n);
        const float4 ro = s->rays[2 * (size_t) i], rd = s->rays[2 * (size_t) i + 1];
        o = mk(ro.x, ro.y, ro.z); d = mk(rd.x, rd.y, rd.z); mint = ro.w; maxt = rd.w;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, uint32_t gid, float t, float u, float v) const {
        NG_CHECK(i < n);
        s->hits[i] = gid == kNoHit ? make_float4(0.f, 0.f, 0.f, __uint_as_float(gid)) : make_float4(t, u, v, __uint_as_float(gid));
    }
};

__global__ void k_traceStaticReset(TraceStaticScratch *s) { s->cursor = 0u; }

/// The renderer's traversal kernel (persistent warps, dynamic ray fetch) on the rays staged in the static array.
template <bool SHADOW, bool WT>
__global__ void __launch_bounds__(128, 10) k_traceStatic(DeviceScene sc, TraceStaticScratch *s, int n) {
    StaticRayIO io; io.s = s; io.n = (uint32_t) n;
    tracePersistent<SHADOW, WT, false>(sc.nodes, sc.leafTris, (uint32_t) n, &s->cursor, io, kTraceTuning, QGrid());
}

/// Global triangle id -> (mesh, mesh-local index) and (t, u, v) as the separate arrays the caller receives.
__global__ void k_traceStaticSplit(DeviceScene sc, TraceStaticScratch *s, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 h = s->hits[i];
    const uint32_t gid = __float_as_uint(h.w);
    uint32_t tri = 0xffffffffu, mesh = 0xffffffffu;
    if (gid != kNoHit) { mesh = sc.tris[gid].w; tri = gid - sc.meshes[mesh].triOffset; }
    s->tri[i] = tri; s->mesh[i] = mesh;
    s->tuv[3 * (size_t) i] = h.x; s->tuv[3 * (size_t) i + 1] = h.y; s->tuv[3 * (size_t) i + 2] = h.z;
}

} // namespace

extern "C" int nori_gpu_trace_static(nori_gpu_ctx *ctx, const float *rays, int n, int shadow, uint32_t *tri_id, uint32_t *mesh_id, float *tuv) {
    if (!ctx) return NORI_GPU_EINVAL;
    NEED_SCENE(); REQUIRE(rays && tri_id && n >= 0, "trace_static: bad arguments");
    if (n > NORI_GPU_TRACE_STATIC_MAX_RAYS) {
        ctx->err = "trace_static: " + std::to_string(n) + " rays exceed the static scratch array (NORI_GPU_TRACE_STATIC_MAX_RAYS = " +
                   std::to_string(NORI_GPU_TRACE_STATIC_MAX_RAYS) + " per call): split the batch or use nori_gpu_trace";
        return NORI_GPU_EINVAL;
    }
    if (n == 0) return NORI_GPU_OK;
    CK(cudaSetDevice(ctx->device));
    TraceStaticScratch *s = nullptr;
    CK(traceStaticScratch(ctx->device, &s));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(s->rays, rays, (size_t) n * 32, cudaMemcpyHostToDevice, st));
    k_traceStaticReset<<<1, 1, 0, st>>>(s);
    const DeviceScene sc = deviceScene(ctx);
    const unsigned grid = (unsigned) ctx->prop.multiProcessorCount * 10u;      // 10 resident CTAs of 128 threads per SM, as launchTrace
    const bool wt = sc.isectMode == NORI_ISECT_WATERTIGHT;
    if (shadow) { if (wt) k_traceStatic<true, true><<<grid, 128, 0, st>>>(sc, s, n); else k_traceStatic<true, false><<<grid, 128, 0, st>>>(sc, s, n); }
    else { if (wt) k_traceStatic<false, true><<<grid, 128, 0, st>>>(sc, s, n); else k_traceStatic<false, false><<<grid, 128, 0, st>>>(sc, s, n); }
    k_traceStaticSplit<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(sc, s, n);
    CK(cudaMemcpyAsync(tri_id, s->tri, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    if (mesh_id) CK(cudaMemcpyAsync(mesh_id, s->mesh, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    if (tuv) CK(cudaMemcpyAsync(tuv, s->tuv, (size_t) n * 12, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st)); CK(cudaGetLastError());
    unsigned line = 0;
    if (const unsigned bad = fetchChecks(&line)) { ctx->err = "trace_static: device range check failed (" + std::to_string(bad) + " at capi.cu/traverse:" + std::to_string(line) + ")"; return NORI_GPU_ECUDA; }
    return NORI_GPU_OK;
}
