// engine.h — host-side declarations shared by capi.cu, render.cu and operators.cu.
#pragma once
#include "scene.cuh"
#include "bvh.h"
#include <string>

namespace ng {

// Pool::state.w bits of a path slot
constexpr uint32_t kFlagAlive = 1u << 16;       // has a ray in flight
constexpr uint32_t kFlagHasSample = 1u << 17;   // finished camera sample waiting to be recorded
constexpr uint32_t kFlagDiscrete = 1u << 18;    // last bounce was a discrete BSDF (or the camera)
constexpr uint32_t kFlagMulT = 1u << 19;        // final value = throughput * radiance (whitted)
constexpr uint32_t kDepthMask = 0xffffu;
constexpr int kNumKeys = 5;                     // diffuse, mirror, dielectric, microfacet, miss
constexpr int kKeyMiss = 4;
constexpr int kMaxRing = 4;
constexpr int kTile = 32;                       // NORI_BLOCK_SIZE (include/nori/block.h:29)
constexpr int kMaxBorder = 4;

/// Wavefront path pool: structure of arrays, one entry per path slot.
struct Pool {
    uint32_t capacity;
    float4 *rayO;        // o.xyz, mint
    float4 *rayD;        // d.xyz, maxt
    float4 *hit;         // t, u, v, bits(global triangle id | kNoHit)
    float4 *thr;         // throughput rgb, BSDF pdf of the last sampled direction
    float4 *rad;         // radiance rgb, eta product
    uint4 *state;        // pcg32 state lo, hi, pcg32 inc (32 bits), flags | depth | ring slot << 24   (read and written every bounce)
    uint4 *info;         // ring slot * numTiles + tile, sample index, record index, -   (written when the path starts, read when it ends)
    float2 *pos;         // pixelSample handed to ImageBlock::put
    float4 *shadow;      // 3 x capacity: (o, maxt) (d, bits(slot)) (contribution, -)
    // Ray binning (off unless `order` is set): the shade kernel writes (bin << 24 | slot) per slot — bin = origin cell, 2 bits per
    // axis of the scene box, + two direction sign bits; fresh camera rays and empty slots in bins of their own —, ONE 8-bit radix
    // pass over those words orders the slot numbers by bin, and the closest-hit pass walks `order` instead of the pool.
    uint32_t *key, *order; void *sortTmp; size_t sortTmpBytes;
    float binLo[3], binInv[3];
};
constexpr uint32_t kBinFresh = 0xfeu, kBinEmpty = 0xffu, kBinSlotMask = 0x00ffffffu;

struct Counters {
    unsigned long long issue, issueLimit, finished;
    unsigned long long closestRays, shadowRays, segments, dropped, iterations;
    unsigned int sortCount[8];
    unsigned int shadowCount;
    unsigned int completedPasses;
    unsigned int tilesSplatted[kMaxRing];   // cumulative per ring slot
    unsigned int slotPass[kMaxRing];        // pass currently held by a ring slot
    unsigned int forcePass;                 // pass whose tiles are splatted now although samples are still in flight
    unsigned int lateSamples;               // samples that finished after their pass was closed (direct splat)
    unsigned int readyCount;                // entries of RenderParams::readyTiles (tiles splatted this iteration)
    unsigned int passHint;                  // pass containing the next sample index to be issued
    unsigned int traceCursor[2];            // next ray of the persistent traversal kernels (closest, shadow)
    unsigned int liveLast;                  // live paths shaded by the previous iteration
    unsigned int compactCount[2];           // k_compactScan: free slots below / occupied slots above the new pool size
    unsigned int traceLimit, orderSorted;   // ray binning: entries of Pool::order the closest-hit pass walks; 1 if the bin sort wrote them (live slots first)
};

struct RenderParams {
    int integrator, streamMode;
    uint32_t sampleFirst, sampleCount;
    int rrStart; float rrMax; int maxDepth;
    V3 lightPos, lightEnergy;
    // issue order: pass-major, tile-major, sample, pixel-in-tile
    int tilesX, tilesY; uint32_t numTiles;
    uint32_t K, Klast, numPasses;
    unsigned long long passSlots, totalIssue;
    int ring;
    // explicit sample list (nori_gpu_sample_radiance)
    const uint32_t *listPixel, *listSample;
    // sample records of the passes in flight
    float4 *recA; float *recB; size_t recCount;   // ring * passSlots records
    uint32_t *tileDone;                      // ring x numTiles: finished samples of the pass held by the slot
    uint32_t *tileGen;                       // ring x numTiles: passes already splatted for this (slot, tile)
    uint32_t *readyTiles;                    // ring x numTiles: (slot * numTiles + tile) entries ready to splat
    uint32_t *tileArrive;                    // ring x numTiles: k_splat CTAs of the tile that finished (reset by the last)
    uint32_t poolCapacity;
    float *block;                            // (H+2b) x (W+2b) x RGBW accumulator
    int blockCols, blockRows;
};

struct StageTimes { float generate, traceClosest, shade, traceShadow, splat; };

/// order[i] = i for i < n (ray binning: after the tail compaction moved paths between slots).
void enqueueIdentityOrder(const Pool &pool, uint32_t n, Counters *c, cudaStream_t st);
/// Bytes of scratch the per-iteration radix pass needs for n slots.
size_t binSortTmpBytes(uint32_t n);

/// Marks every slot of the pool empty (start of a render).
void enqueuePoolInit(const Pool &pool, Counters *d_counters, cudaStream_t st);

/// Tail of a render (no camera sample left to hand out): moves the occupied slots at or above
/// `newCapacity` into free slots below it, so that the following iterations run over
/// `newCapacity` slots only. `pool.capacity` is the size the pool was run with so far.
int enqueueCompaction(const Pool &pool, uint32_t newCapacity, Counters *d_counters, cudaStream_t st);

/// Follows every path still in `pool` (capacity = the compacted size) to its end in one launch.
int enqueueFinishTail(const DeviceScene &sc, const Pool &pool, const RenderParams &rp, Counters *d_counters, cudaStream_t st);

/// One wavefront iteration, enqueued on `st`. Returns the number of kernel launches.
int enqueueIteration(const DeviceScene &sc, const Pool &pool, const RenderParams &rp, Counters *d_counters, cudaStream_t st,
                     cudaEvent_t *stageEvents /* 4 or nullptr */);

/// Integrator::Li for n rays (2 float4 each) with per-ray pcg32 (state, inc) pairs read and written back; rgb n x 3.
void enqueueLi(const DeviceScene &sc, const RenderParams &rp, int n, const float4 *rays, unsigned long long *rngState, float *rgb, cudaStream_t st);

/// Path vertices (NoriGpuSPoint records) of nPaths = W * H * spp camera samples: counts (out == nullptr) or records at offsets.
bool enqueuePathVertices(const DeviceScene &sc, const RenderParams &rp, uint32_t spp, unsigned long long nPaths, unsigned int *counts,
                         const unsigned long long *offsets, NoriGpuSPoint *out, cudaStream_t st, NoriGpuLPoint *lights = nullptr);
/// offsets[0..nPaths] = exclusive prefix sums of counts (offsets[nPaths] = total).
void enqueueScanCounts(const unsigned int *counts, unsigned long long nPaths, unsigned long long *offsets, cudaStream_t st);

/// OBJ corner de-duplication in first-use order + vertex gather (objdedup.cu; src/obj.cpp:96-161). Device pointers.
bool objDedup(const uint32_t *corners, uint32_t n, const float *pos, uint32_t nP, const float *tex, uint32_t nT, const float *nrm, uint32_t nN,
              uint32_t *F, float *V, float *N, float *UV, uint32_t *nVertsOut, cudaStream_t st, std::string &err);

/// Checked build (check.cuh): violations counted by the kernels of render.cu / operators.cu since the last call.
unsigned renderChecks(unsigned *line);
unsigned operatorChecks(unsigned *line);

// raw operators (operators.cu); all pointers are device pointers
void launchTrace(const DeviceScene &sc, const float4 *rays /* 2 per ray */, int n, bool shadow, float4 *hits, unsigned int *cursor /* device scratch */, cudaStream_t st);
void launchTraceCount(const DeviceScene &sc, const float4 *rays, int n, bool shadow, float4 *hits, unsigned long long *counts2, cudaStream_t st);
void launchIntersect(const DeviceScene &sc, const float4 *rays, int n, float *out24, int *valid, cudaStream_t st);
void launchCameraRays(const DeviceScene &sc, const float2 *ps, int n, float4 *rays, cudaStream_t st);
void launchBsdfSample(Bsdf b, int n, const float *wi, const float *xi, float *wo, float *weight, float *pdf, float *eta, int *measure, cudaStream_t st);
void launchBsdfEvalPdf(Bsdf b, int n, const float *wi, const float *wo, float *f, float *pdf, cudaStream_t st);
void launchWarp(int id, float param, int n, const float *xi, float *out3, float *pdf, cudaStream_t st);
void launchWarpPdf(int id, float param, int n, const float *pts3, float *pdf, cudaStream_t st);
void launchPcg32(uint64_t initstate, uint64_t initseq, int64_t advance, int n, uint32_t *out, cudaStream_t st);
void launchSplatList(const DeviceScene &sc, int n, const float2 *pos, const float *rgb, float *block, unsigned long long *dropped, cudaStream_t st);
void launchSplatOrdered(const DeviceScene &sc, int n, const float2 *pos, const float *rgb, float *block, unsigned long long *dropped, cudaStream_t st);
void launchPackMesh(const float *V, const float *N, const float *UV, const uint32_t *F, uint32_t nV, uint32_t nF, uint32_t vertOffset, uint32_t mesh,
                    float4 *vpos, float4 *vnrm, float2 *vuv, uint4 *tris, unsigned int *badIndex, cudaStream_t st);
void launchToBitmap(const float *block, int W, int H, int border, float *rgb, cudaStream_t st);
void launchToSRGB8(const float *block, int W, int H, int border, uint8_t *rgb8, cudaStream_t st);
void launchAddBlock(float *dst, const float *src, size_t nFloats, cudaStream_t st);

} // namespace ng
