// pathgraph.cu — first slice of the fork's path-graph post-process as sm_100a kernels: the uniform hash
// grid over the path vertices and the batched k-nearest-neighbour search, i.e. what the reference runs in
//   buildHashGridAndFindKNN      src/pbsdf.cu:3173-3235   (host driver)
//   HashGridSize / getOffset / buildHash   src/pbsdf.cu:844-887   (cell histogram, per-point offsets, scatter)
//   batchHashKNN / batchNearestNeighbor    src/pbsdf.cu:2834-2896, 1089-1178
// Differences in HOW (not in WHAT is computed):
//   * the cell offsets are scanned on the device (the reference copies the histogram to the host, scans it with
//     thrust::exclusive_scan on the CPU and copies it back: pbsdf.cu:3207-3216);
//   * queries are processed in CELL order (one thread per grid entry), so the lanes of a warp walk the same few cells and
//     their candidate loads are broadcasts out of L1; the reference walks queries in input order;
//   * the K best candidates of a query live in a max-heap held in SHARED memory as [entry][thread] columns (bank = thread:
//     conflict-free whatever the lanes' heap positions are) — no global scratch: the reference gives every thread
//     8 x (largest cell) floats + ints of global memory and runs quickselect on them (pbsdf.cu:2855-2886, 743-760), in
//     batches sized by cudaMemGetInfo;
//   * selection is by the total order (distance, index): the result is deterministic and independent of the order in which
//     atomics filled a cell. The reference's quickselect breaks distance ties arbitrarily.
// The cell arithmetic (getKeys / getKeysAndDirs / makesureGrid, pbsdf.cu:630-661, 762-777) and the distance
// (dev_distance, pbsdf.cu:667-675, as nvcc contracts it: fused multiply-adds) are restated operation for operation.
#include <cub/cub.cuh>
#include "engine.h"
#include "check.cuh"
#include "pathgraph.h"

namespace ng {

namespace {

struct I3 { int x, y, z; };

/// getKeys (src/pbsdf.cu:630-634): cell of a position, clamped into the grid.
__device__ __forceinline__ I3 pgKeys(const PgParams &P, float x, float y, float z) {
    I3 k;
    k.x = max(0, min((int) floorf((x - P.minb[0]) / P.cell[0]), P.dim[0] - 1));
    k.y = max(0, min((int) floorf((y - P.minb[1]) / P.cell[1]), P.dim[1] - 1));
    k.z = max(0, min((int) floorf((z - P.minb[2]) / P.cell[2]), P.dim[2] - 1));
    return k;
}
/// getKey (src/pbsdf.cu:663-665)
__device__ __forceinline__ int pgKey(const PgParams &P, I3 k) { return P.dim[2] * P.dim[1] * k.x + P.dim[2] * k.y + k.z; }

/// The half of its cell a coordinate lies in (getKeysAndDirs, src/pbsdf.cu:643-660): t = x - cellsize * key - minb, with the
/// product contracted into the subtraction as nvcc's default -fmad=true does for the reference; +1 if t > cellsize / 2.
__device__ __forceinline__ int pgDir(float x, float cell, float minb, int key) {
    const float t = __fsub_rn(__fmaf_rn(-cell, (float) key, x), minb);
    return (double) t > (double) cell / 2.0 ? 1 : -1;
}
/// checkOneDim (src/pbsdf.cu:762-771): a clamped one-cell range at the border of the grid grows inwards.
__device__ __forceinline__ void pgCheckOneDim(int &l, int &h, int dim) {
    if (l == 0 && (h - l < 1) && h < dim - 1) { h += 1; return; }
    if (h == dim - 1 && (h - l < 1) && l > 0) { l -= 1; return; }
}
/// dev_distance (src/pbsdf.cu:667-675): a = candidate, b = query; dist += (a-b)*(a-b) contracted to FMAs, then sqrtf.
__device__ __forceinline__ float pgDistance(float ax, float ay, float az, float bx, float by, float bz) {
    const float dx = __fsub_rn(ax, bx), dy = __fsub_rn(ay, by), dz = __fsub_rn(az, bz);
    float d = __fmaf_rn(dx, dx, 0.0f);
    d = __fmaf_rn(dy, dy, d);
    d = __fmaf_rn(dz, dz, d);
    return __fsqrt_rn(d);
}

/// HashGridSize (src/pbsdf.cu:844-851): points per cell. `pos` = the position field of n records `stride` bytes apart
/// (SPoint::pos at offset 0 of a 168-byte record, include/nori/shadingPoint.h:125-147, or packed float3).
__global__ void k_pgCount(const unsigned char *__restrict__ pos, size_t stride, int n, PgParams P, int *__restrict__ cellCount) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *p = reinterpret_cast<const float *>(pos + (size_t) i * stride);
    atomicAdd(&cellCount[pgKey(P, pgKeys(P, p[0], p[1], p[2]))], 1);
}

// ---- exclusive scan of the cell histogram on the device (replaces the D2H + thrust::exclusive_scan + H2D of pbsdf.cu:3207-3216)
constexpr int kScanThreads = 256, kScanItems = 8, kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int blockExclusiveScan(int v, int *sWarp, int &total) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    int s = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, d); if ((int) lane >= d) s += t; }
    if (lane == 31) sWarp[warp] = s;
    __syncthreads();
    if (warp == 0) {
        int w = lane < (unsigned) (blockDim.x >> 5) ? sWarp[lane] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, d); if ((int) lane >= d) w += t; }
        sWarp[lane] = w;                                             // inclusive prefix of the warp sums
    }
    __syncthreads();
    total = sWarp[(blockDim.x >> 5) - 1];
    const int base = warp ? sWarp[warp - 1] : 0;
    __syncthreads();
    return base + s - v;
}
/// Pass 1: every CTA scans its tile of 2048 counts in place (exclusive) and writes the tile's sum.
__global__ void __launch_bounds__(kScanThreads) k_pgScanTiles(int *__restrict__ data, int n, int *__restrict__ tileSums) {
    __shared__ int sWarp[32];
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int v[kScanItems], sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) { v[k] = base + k < n ? data[base + k] : 0; sum += v[k]; }
    int total;
    int run = blockExclusiveScan(sum, sWarp, total);
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) { if (base + k < n) data[base + k] = run; run += v[k]; }
    if (threadIdx.x == 0) tileSums[blockIdx.x] = total;
}
/// Pass 2: one CTA turns the tile sums into exclusive tile offsets (loops when there are more than 256 x 8 tiles).
__global__ void __launch_bounds__(kScanThreads) k_pgScanSums(int *__restrict__ tileSums, int nTiles) {
    __shared__ int sWarp[32];
    int carry = 0;
    for (int start = 0; start < nTiles; start += kScanTile) {
        const int base = start + threadIdx.x * kScanItems;
        int v[kScanItems], sum = 0;
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) { v[k] = base + k < nTiles ? tileSums[base + k] : 0; sum += v[k]; }
        int total;
        int run = carry + blockExclusiveScan(sum, sWarp, total);
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) { if (base + k < nTiles) tileSums[base + k] = run; run += v[k]; }
        carry += total;
    }
}
/// Pass 3: add the tile offsets; entry n (one past the last cell) receives the grand total = number of points.
__global__ void __launch_bounds__(kScanThreads) k_pgScanAdd(int *__restrict__ data, int n, const int *__restrict__ tileSums, int nPoints) {
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    const int off = tileSums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) if (base + k < n) data[base + k] += off;
    if (blockIdx.x == 0 && threadIdx.x == 0) data[n] = nPoints;
}

/// getOffset + buildHash (src/pbsdf.cu:863-887): every point takes the next free slot of its cell. entries[slot] =
/// (position, bits(index)) = the reference's GridEntry (include/nori/shadingPoint.h:26-29), 16 bytes, one vector load.
__global__ void k_pgScatter(const unsigned char *__restrict__ pos, size_t stride, int n, PgParams P, const int *__restrict__ cellStart,
                            int *__restrict__ cellFill, float4 *__restrict__ entries) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *p = reinterpret_cast<const float *>(pos + (size_t) i * stride);
    const float x = p[0], y = p[1], z = p[2];
    const int key = pgKey(P, pgKeys(P, x, y, z));
    const int slot = cellStart[key] + atomicAdd(&cellFill[key], 1);
    NG_CHECK(slot >= 0 && slot < n);
    entries[slot] = make_float4(x, y, z, __int_as_float(i));
}

// ---- k nearest neighbours ------------------------------------------------------------------------------------------
/// Candidate order: by distance, ties by point index (a total order: the selection does not depend on how a cell was filled).
__device__ __forceinline__ bool pgLess(float da, int ia, float db, int ib) { return da < db || (da == db && ia < ib); }

/// batchNearestNeighbor (src/pbsdf.cu:1089-1178). One thread per grid entry (cell order). The 2 x 2 x 2 block of cells
/// around the query (the neighbours on the side of the cell the point lies in) is scanned; the maxK = max(k, dk) nearest
/// candidates are kept in a max-heap in shared memory, heap-sorted at the end and written nearest first, so that the first
/// min(k, dk) entries are the min(k, dk) nearest as the reference's second selection guarantees (pbsdf.cu:1138-1139).
/// neighbors: n rows of `pitchInts` ints (row = ORIGINAL point index, like the reference's pitched array); unused entries -1.
template <int NT>
__global__ void __launch_bounds__(NT) k_pgKnn(const float4 *__restrict__ entries, const int *__restrict__ cellStart, int n, PgParams P,
                                              int *__restrict__ neighbors, int pitchInts, int *__restrict__ knn, int *__restrict__ dknn) {
    extern __shared__ unsigned char pgSmem[];
    const int K = P.maxK;
    float *hd = reinterpret_cast<float *>(pgSmem);                 // [K][NT]
    int *hi = reinterpret_cast<int *>(pgSmem) + (size_t) K * NT;   // [K][NT]
    const int tid = threadIdx.x, slot = blockIdx.x * NT + tid;
    if (slot >= n) return;
    const float4 q = __ldg(entries + slot);
    const int Index = __float_as_int(q.w);
    const I3 key = pgKeys(P, q.x, q.y, q.z);
    const int dx = pgDir(q.x, P.cell[0], P.minb[0], key.x), dy = pgDir(q.y, P.cell[1], P.minb[1], key.y), dz = pgDir(q.z, P.cell[2], P.minb[2], key.z);
    int lx = max(0, min(key.x, key.x + dx)), ly = max(0, min(key.y, key.y + dy)), lz = max(0, min(key.z, key.z + dz));
    int hx = min(P.dim[0] - 1, max(key.x, key.x + dx)), hy = min(P.dim[1] - 1, max(key.y, key.y + dy)), hz = min(P.dim[2] - 1, max(key.z, key.z + dz));
    pgCheckOneDim(lx, hx, P.dim[0]); pgCheckOneDim(ly, hy, P.dim[1]); pgCheckOneDim(lz, hz, P.dim[2]);      // makesureGrid

#define HD(j) hd[(size_t) (j) * NT + tid]
#define HI(j) hi[(size_t) (j) * NT + tid]
    int count = 0, total = 0;
    for (int ix = lx; ix <= hx; ++ix)
        for (int iy = ly; iy <= hy; ++iy) {
            // cells (ix, iy, lz..hz) are consecutive keys: one contiguous range of entries
            I3 c0 = {ix, iy, lz}, c1 = {ix, iy, hz};
            const int begin = __ldg(cellStart + pgKey(P, c0)), end = __ldg(cellStart + pgKey(P, c1) + 1);
            total += end - begin;
            for (int e = begin; e < end; ++e) {
                const float4 g = __ldg(entries + e);
                const float d = pgDistance(g.x, g.y, g.z, q.x, q.y, q.z);
                const int id = __float_as_int(g.w);
                if (count < K) {                                   // push: sift up
                    int j = count++;
                    while (j > 0) {
                        const int p = (j - 1) >> 1;
                        const float pd = HD(p); const int pi = HI(p);
                        if (!pgLess(pd, pi, d, id)) break;         // parent >= new: stop
                        HD(j) = pd; HI(j) = pi; j = p;
                    }
                    HD(j) = d; HI(j) = id;
                } else if (pgLess(d, id, HD(0), HI(0))) {          // better than the worst kept: replace the root, sift down
                    int j = 0;
                    for (;;) {
                        int c = 2 * j + 1;
                        if (c >= K) break;
                        float cd = HD(c); int ci = HI(c);
                        if (c + 1 < K) { const float rd = HD(c + 1); const int ri = HI(c + 1); if (pgLess(cd, ci, rd, ri)) { ++c; cd = rd; ci = ri; } }
                        if (!pgLess(d, id, cd, ci)) break;         // new >= larger child: stop
                        HD(j) = cd; HI(j) = ci; j = c;
                    }
                    HD(j) = d; HI(j) = id;
                }
            }
        }
    // heap sort in place: ascending (distance, index)
    for (int m = count - 1; m > 0; --m) {
        const float d = HD(m); const int id = HI(m);
        HD(m) = HD(0); HI(m) = HI(0);
        int j = 0;
        for (;;) {
            int c = 2 * j + 1;
            if (c >= m) break;
            float cd = HD(c); int ci = HI(c);
            if (c + 1 < m) { const float rd = HD(c + 1); const int ri = HI(c + 1); if (pgLess(cd, ci, rd, ri)) { ++c; cd = rd; ci = ri; } }
            if (!pgLess(d, id, cd, ci)) break;
            HD(j) = cd; HI(j) = ci; j = c;
        }
        HD(j) = d; HI(j) = id;
    }
    // "ensure the first point is myself" (pbsdf.cu:1168-1176): swap the query's own entry to the front
    for (int j = 0; j < count; ++j)
        if (HI(j) == Index) { HI(j) = HI(0); HI(0) = Index; const float t = HD(j); HD(j) = HD(0); HD(0) = t; break; }
    int *row = neighbors + (size_t) Index * pitchInts;
    for (int j = 0; j < K; ++j) row[j] = j < count ? HI(j) : -1;
    // neighbour counts for the two consumers (k: indirect scatter, dk: direct light; pbsdf.cu:1141-1164)
    const int maxK = K, minK = min(P.k, P.dk);
    int nk, ndk;
    if (maxK < total) { nk = P.k; ndk = P.dk; }
    else if (minK < total) { if (P.dk > P.k) { nk = P.k; ndk = total; } else { nk = total; ndk = P.dk; } }
    else { nk = total; ndk = total; }
    knn[Index] = nk; dknn[Index] = ndk;
#undef HD
#undef HI
}

} // namespace

size_t pgScratchBytes(int n, int nCells) {
    const size_t tiles = ((size_t) nCells + kScanTile - 1) / kScanTile;
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    return al(((size_t) nCells + 1) * 4) + al((size_t) nCells * 4) + al(tiles * 4) + al((size_t) n * 16);
}

namespace {
/// Sum over the cells of (points in the cell)^2: an eighth of the candidate distances the search will compute, roughly.
__global__ void k_pgSearchCost(const int *__restrict__ cellStart, int nCells, unsigned long long *__restrict__ out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long v = 0;
    if (c < nCells) { const unsigned long long m = (unsigned long long) (cellStart[c + 1] - cellStart[c]); v = m * m; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(out, v);
}
}

/// Grid build + kNN on device-resident positions. `scratch`: pgScratchBytes(n, nCells) bytes of device memory.
/// Returns false (with a message) for parameters the kernels do not cover.
bool launchPathgraphKnn(const void *d_pos, size_t stride, int n, const PgParams &P, void *scratch, int *d_neighbors, int pitchInts, int *d_knn, int *d_dknn,
                        cudaStream_t st, std::string &err) {
    if (P.maxK < 1 || P.maxK > kPgMaxK) { err = "pathgraph_knn: max(k, dk) must be in [1, " + std::to_string(kPgMaxK) + "]"; return false; }
    const int nCells = P.nCells;
    const int tiles = (nCells + kScanTile - 1) / kScanTile;
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    unsigned char *s = static_cast<unsigned char *>(scratch);
    int *cellStart = reinterpret_cast<int *>(s); s += al(((size_t) nCells + 1) * 4);
    int *cellFill = reinterpret_cast<int *>(s); s += al((size_t) nCells * 4);
    int *tileSums = reinterpret_cast<int *>(s); s += al((size_t) tiles * 4);
    float4 *entries = reinterpret_cast<float4 *>(s);
    cudaMemsetAsync(cellStart, 0, ((size_t) nCells + 1) * 4, st);
    cudaMemsetAsync(cellFill, 0, (size_t) nCells * 4, st);
    const unsigned g = (unsigned) ((n + 255) / 256);
    k_pgCount<<<g, 256, 0, st>>>(static_cast<const unsigned char *>(d_pos), stride, n, P, cellStart);
    k_pgScanTiles<<<tiles, kScanThreads, 0, st>>>(cellStart, nCells, tileSums);
    k_pgScanSums<<<1, kScanThreads, 0, st>>>(tileSums, tiles);
    k_pgScanAdd<<<tiles, kScanThreads, 0, st>>>(cellStart, nCells, tileSums, n);
    k_pgScatter<<<g, 256, 0, st>>>(static_cast<const unsigned char *>(d_pos), stride, n, P, cellStart, cellFill, entries);
    {   // every query scans the cells around it: with nearly all points in a few cells that is quadratic. Refuse what would run for many seconds
        // (the grid PathGraph::computeDimensions derives from the points' bounding box keeps this sum near n x a few hundred).
        unsigned long long *cost = reinterpret_cast<unsigned long long *>(tileSums), h = 0;          // tileSums is free again
        static_assert(sizeof(unsigned long long) <= 256, "fits the smallest tileSums allocation");
        cudaMemsetAsync(cost, 0, sizeof(unsigned long long), st);
        k_pgSearchCost<<<(unsigned) ((nCells + 255) / 256), 256, 0, st>>>(cellStart, nCells, cost);
        if (cudaMemcpyAsync(&h, cost, sizeof h, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) { err = "pathgraph_knn: cost read failed"; return false; }
        if ((double) h > 5e11) { err = "pathgraph_knn: the grid puts too many points into too few cells (sum of squared cell counts " + std::to_string(h) + "): choose a finer grid"; return false; }
    }
    constexpr int NT = 128;
    const size_t smem = (size_t) P.maxK * NT * 8;
    cudaFuncSetAttribute(k_pgKnn<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPgMaxK * NT * 8);      // per device; cheap
    k_pgKnn<NT><<<(unsigned) ((n + NT - 1) / NT), NT, smem, st>>>(entries, cellStart, n, P, d_neighbors, pitchInts, d_knn, d_dknn);
    return true;
}

// ---- SPoints::BuildClusters (include/nori/shadingPoint.h:637-760) ---------------------------------------------------------------
// seeds -> nearest seed through a hashed table of the seeds' cells (buildBatchClusters, src/pbsdf.cu:3555-3622) -> two rounds of splitting
// every cluster around its first size / K + 1 members (countMyClusters + subClusters, :3236-3310) -> member lists (FinalizeCluster).
// HOW it differs: the reference fills its table buckets and member lists with atomics (buildHashSub :888-902, SaveClusters :987-1006), so
// the order inside a bucket or a cluster — and with it which members seed a split — depends on scheduling, and it scans sizes into
// offsets through thrust with host round trips. Here both are STABLE radix sorts by bucket / cluster (members in index order: one of the
// orders the reference can produce) with the offsets read off the sorted keys by binary search: deterministic, no atomics.
namespace {

__device__ __forceinline__ unsigned int pgHashBucket(unsigned int key, unsigned int table) { return (key * 5039u + 39916801u) % table; }   // hashbmap, :798-800
__device__ __forceinline__ const float *pgPosOf(const unsigned char *pos, size_t stride, int i) { return reinterpret_cast<const float *>(pos + (size_t) i * stride); }

/// HashTableSize / buildHashSub (src/pbsdf.cu:853-861, 888-902): bucket of every seed; sorted by bucket afterwards.
__global__ void k_pgSeedKeys(const unsigned char *__restrict__ pos, size_t stride, const int *__restrict__ seeds, int nSeeds, int n, PgParams P,
                             unsigned int *__restrict__ keys, int *__restrict__ vals) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nSeeds) return;
    const int i = seeds[s];
    NG_CHECK(i >= 0 && i < n);
    const float *p = pgPosOf(pos, stride, i);
    keys[s] = pgHashBucket((unsigned int) pgKey(P, pgKeys(P, p[0], p[1], p[2])), (unsigned int) nSeeds);
    vals[s] = s;
}
/// start[b] = first position in the ascending keys with key >= b, for b in [0, nKeys]: offsets of buckets / clusters without a scan.
__global__ void k_pgLowerBounds(const unsigned int *__restrict__ sorted, int nSorted, int nKeys, int *__restrict__ start) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > nKeys) return;
    int lo = 0, hi = nSorted;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (sorted[mid] < (unsigned int) b) lo = mid + 1; else hi = mid; }
    start[b] = lo;
}
__global__ void k_pgSeedEntries(const unsigned char *__restrict__ pos, size_t stride, const int *__restrict__ seeds, const int *__restrict__ sortedSeed, int nSeeds,
                                float4 *__restrict__ entries) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nSeeds) return;
    const int s = sortedSeed[e];
    const float *p = pgPosOf(pos, stride, seeds[s]);
    entries[e] = make_float4(p[0], p[1], p[2], __int_as_float(s));
}
/// Cluster (src/pbsdf.cu:1008-1044): nearest seed among the table buckets of the 2 x 2 x 2 cell block around the vertex; `<=`: the last
/// of equals wins; -1 if none within minDist (twice the cell diagonal).
__global__ void k_pgAssign(const unsigned char *__restrict__ pos, size_t stride, int n, PgParams P, const float4 *__restrict__ entries,
                           const int *__restrict__ bucketStart, int nSeeds, float minDist, int *__restrict__ groupIdx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *p = pgPosOf(pos, stride, i);
    const float qx = p[0], qy = p[1], qz = p[2];
    const I3 key = pgKeys(P, qx, qy, qz);
    const int dx = pgDir(qx, P.cell[0], P.minb[0], key.x), dy = pgDir(qy, P.cell[1], P.minb[1], key.y), dz = pgDir(qz, P.cell[2], P.minb[2], key.z);
    int lx = max(0, min(key.x, key.x + dx)), ly = max(0, min(key.y, key.y + dy)), lz = max(0, min(key.z, key.z + dz));
    int hx = min(P.dim[0] - 1, max(key.x, key.x + dx)), hy = min(P.dim[1] - 1, max(key.y, key.y + dy)), hz = min(P.dim[2] - 1, max(key.z, key.z + dz));
    pgCheckOneDim(lx, hx, P.dim[0]); pgCheckOneDim(ly, hy, P.dim[1]); pgCheckOneDim(lz, hz, P.dim[2]);
    float best = minDist; int who = -1;
    for (int ix = lx; ix <= hx; ++ix)
        for (int iy = ly; iy <= hy; ++iy)
            for (int iz = lz; iz <= hz; ++iz) {
                I3 c = {ix, iy, iz};
                const unsigned int b = pgHashBucket((unsigned int) pgKey(P, c), (unsigned int) nSeeds);
                const int begin = __ldg(bucketStart + b), end = __ldg(bucketStart + b + 1);
                for (int e = begin; e < end; ++e) {
                    const float4 g = __ldg(entries + e);
                    const float d = pgDistance(g.x, g.y, g.z, qx, qy, qz);
                    if (d <= best) { best = d; who = __float_as_int(g.w); }
                }
            }
    groupIdx[i] = who;
}
__global__ void k_pgGroupKeys(const int *__restrict__ groupIdx, int n, unsigned int *__restrict__ keys, int *__restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = (unsigned int) groupIdx[i];          // -1 -> 0xffffffff: behind every cluster
    vals[i] = i;
}
/// countClusters (src/pbsdf.cu:910-920): a cluster of `size` members becomes size / K + 1 clusters. small[nClusters] = 0 (so that the
/// exclusive sum over nClusters + 1 entries ends with the total).
__global__ void k_pgSmall(const int *__restrict__ offsets /* nClusters + 1 */, int nClusters, int K, int *__restrict__ small) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c > nClusters) return;
    small[c] = c == nClusters ? 0 : max((offsets[c + 1] - offsets[c]) / K + 1, 1);
}
/// SubdivideClusters (src/pbsdf.cu:942-974): a member joins the nearest of its cluster's first `pnum` listed members (strict <: the first
/// of equals; nothing nearer than 262144 -> the first).
__global__ void k_pgSubdivide(const unsigned char *__restrict__ pos, size_t stride, int n, const int *__restrict__ lists, const int *__restrict__ offsets,
                              const int *__restrict__ first /* nClusters + 1 */, int *__restrict__ groupIdx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c = groupIdx[i];
    if (c < 0) return;
    const int f = first[c], pnum = first[c + 1] - f;
    int g = f;
    if (pnum > 1) {
        const float *p = pgPosOf(pos, stride, i);
        const int *mem = lists + offsets[c];
        float best = 262144.0f; int who = 0;
        for (int t = 0; t < pnum; ++t) {
            const float *q = pgPosOf(pos, stride, mem[t]);
            const float d = pgDistance(q[0], q[1], q[2], p[0], p[1], p[2]);
            if (d < best) { best = d; who = t; }
        }
        g += who;
    }
    groupIdx[i] = g;
}

/// Largest cluster of a listing (offsets[nClusters] = members listed).
__global__ void k_pgMaxCluster(const int *__restrict__ offsets, int nClusters, int *__restrict__ out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    int m = c < nClusters ? offsets[c + 1] - offsets[c] : 0;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

inline size_t pgAl(size_t b) { return (b + 255) / 256 * 256; }
size_t pgSortTempBytes(int n) {
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, (const unsigned int *) nullptr, (unsigned int *) nullptr, (const int *) nullptr, (int *) nullptr, n);
    cub::DeviceScan::ExclusiveSum(nullptr, b, (const int *) nullptr, (int *) nullptr, n + 1);
    return std::max(a, b);
}

} // namespace

size_t pgBuildClustersScratchBytes(int n, int nSeeds, int cap) {
    const int m = std::max(n, nSeeds), c = std::max(cap, nSeeds) + 2;
    return 4 * pgAl((size_t) m * 4) + pgAl((size_t) nSeeds * 16) + 3 * pgAl((size_t) c * 4) + pgAl(pgSortTempBytes(std::max(m, c))) + 1024;
}

/// BuildClusters on device-resident positions. d_groupIdx: n ints (cluster of every vertex or -1), d_lists: n ints, d_offsets: `cap` + 1 ints
/// (offsets[nClusters] = members listed). Synchronises the stream three times to read a cluster count (the reference: after every kernel).
bool launchPathgraphBuildClusters(const void *d_pos, size_t stride, int n, const PgParams &P, const int *d_seeds, int nSeeds, int K, void *scratch,
                                  int *d_groupIdx, int *d_lists, int *d_offsets, int cap, int *nClustersOut, int *nListedOut, cudaStream_t st, std::string &err,
                                  int *maxClusterOut) {
    if (n <= 0 || nSeeds <= 0 || nSeeds > n || K < 2) { err = "pathgraph_build_clusters: need 0 < n_seeds <= n and K >= 2"; return false; }
    const int m = std::max(n, nSeeds), c = std::max(cap, nSeeds) + 2;
    unsigned char *s = static_cast<unsigned char *>(scratch);
    unsigned int *keysA = reinterpret_cast<unsigned int *>(s); s += pgAl((size_t) m * 4);
    unsigned int *keysB = reinterpret_cast<unsigned int *>(s); s += pgAl((size_t) m * 4);
    int *valsA = reinterpret_cast<int *>(s); s += pgAl((size_t) m * 4);
    int *valsB = reinterpret_cast<int *>(s); s += pgAl((size_t) m * 4);
    float4 *entries = reinterpret_cast<float4 *>(s); s += pgAl((size_t) nSeeds * 16);
    int *offs = reinterpret_cast<int *>(s); s += pgAl((size_t) c * 4);
    int *small = reinterpret_cast<int *>(s); s += pgAl((size_t) c * 4);
    int *first = reinterpret_cast<int *>(s); s += pgAl((size_t) c * 4);
    void *tmp = s;
    size_t tmpBytes = pgSortTempBytes(std::max(m, c));
    const unsigned char *pos = static_cast<const unsigned char *>(d_pos);
    auto grid = [](int work) { return (unsigned) ((work + 255) / 256); };
    auto fail = [&](const char *what, cudaError_t e) { err = std::string("pathgraph_build_clusters: ") + what + ": " + cudaGetErrorString(e); return false; };
    // table of the seeds (size nSeeds: shadingPoint.h:667), buckets in ascending seed order
    k_pgSeedKeys<<<grid(nSeeds), 256, 0, st>>>(pos, stride, d_seeds, nSeeds, n, P, keysA, valsA);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, keysA, keysB, valsA, valsB, nSeeds, 0, 32, st);
    if (e != cudaSuccess) return fail("sort", e);
    k_pgLowerBounds<<<grid(nSeeds + 1), 256, 0, st>>>(keysB, nSeeds, nSeeds, offs);
    k_pgSeedEntries<<<grid(nSeeds), 256, 0, st>>>(pos, stride, d_seeds, valsB, nSeeds, entries);
    const float diag = std::sqrt(P.cell[0] * P.cell[0] + P.cell[1] * P.cell[1] + P.cell[2] * P.cell[2]);
    const float minDist = (float) ((double) diag * 2.0);
    k_pgAssign<<<grid(n), 256, 0, st>>>(pos, stride, n, P, entries, offs, nSeeds, minDist, d_groupIdx);
    int nClusters = nSeeds;
    for (int round = 0; round < 3; ++round) {
        // member lists in index order + offsets (countMyClusters' scan + SaveClusters; FinalizeCluster in the last round)
        int *offOut = round == 2 ? d_offsets : offs;
        k_pgGroupKeys<<<grid(n), 256, 0, st>>>(d_groupIdx, n, keysA, valsA);
        tmpBytes = pgSortTempBytes(std::max(m, c));
        e = cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, keysA, keysB, valsA, d_lists, n, 0, 32, st);
        if (e != cudaSuccess) return fail("sort", e);
        k_pgLowerBounds<<<grid(nClusters + 1), 256, 0, st>>>(keysB, n, nClusters, offOut);
        if (round == 2) break;
        k_pgSmall<<<grid(nClusters + 1), 256, 0, st>>>(offs, nClusters, K, small);
        tmpBytes = pgSortTempBytes(std::max(m, c));
        e = cub::DeviceScan::ExclusiveSum(tmp, tmpBytes, small, first, nClusters + 1, st);
        if (e != cudaSuccess) return fail("scan", e);
        int next = 0, biggest = 0;
        cudaMemsetAsync(small, 0, sizeof(int), st);              // (k_pgSmall's values are consumed: the scan above read them)
        k_pgMaxCluster<<<grid(nClusters), 256, 0, st>>>(offs, nClusters, small);
        cudaMemcpyAsync(&biggest, small, sizeof(int), cudaMemcpyDeviceToHost, st);
        if ((e = cudaMemcpyAsync(&next, first + nClusters, sizeof(int), cudaMemcpyDeviceToHost, st)) != cudaSuccess || (e = cudaStreamSynchronize(st)) != cudaSuccess) return fail("count", e);
        // a split compares every member with the cluster's size / K + 1 first members: refuse what would run for seconds (coincident vertices, seeds far from the data)
        if ((double) biggest * ((double) biggest / K + 1.0) > 1.4e11) { err = "pathgraph_build_clusters: a cluster of " + std::to_string(biggest) + " members is too large to split (are the seeds spread over the vertices?)"; return false; }
        if (next > cap) { err = "pathgraph_build_clusters: " + std::to_string(next) + " clusters exceed the capacity " + std::to_string(cap); return false; }
        k_pgSubdivide<<<grid(n), 256, 0, st>>>(pos, stride, n, d_lists, offs, first, d_groupIdx);
        nClusters = next;
    }
    int listed = 0, largest = 0;
    cudaMemsetAsync(small, 0, sizeof(int), st);                  // `small` is free again: the largest cluster goes there
    k_pgMaxCluster<<<grid(nClusters), 256, 0, st>>>(d_offsets, nClusters, small);
    cudaMemcpyAsync(&largest, small, sizeof(int), cudaMemcpyDeviceToHost, st);
    if ((e = cudaMemcpyAsync(&listed, d_offsets + nClusters, sizeof(int), cudaMemcpyDeviceToHost, st)) != cudaSuccess || (e = cudaStreamSynchronize(st)) != cudaSuccess) return fail("count", e);
    if (maxClusterOut) *maxClusterOut = largest;
    if ((e = cudaGetLastError()) != cudaSuccess) return fail("launch", e);
    *nClustersOut = nClusters; *nListedOut = listed;
    return true;
}

namespace {
__global__ void k_pgStrideSeeds(int *__restrict__ seeds, int nSeeds, int n, long long step) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nSeeds) seeds[s] = (int) (((long long) s * step) % n);
}
}
/// The smallest step >= 0.618 n that shares no factor with n: s -> (s * step) mod n then visits n distinct vertices, consecutive seeds far apart
/// in the record order (= pixel order), so the seeds spread over the image the way a shuffled list's head does.
long long pathgraphSeedStep(int n) {
    long long step = std::max(1LL, (long long) ((double) n * 0.6180339887498949));
    auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
    while (gcd(step, n) != 1) ++step;
    return step;
}
void launchPathgraphStrideSeeds(int *d_seeds, int nSeeds, int n, cudaStream_t st) {
    k_pgStrideSeeds<<<(unsigned) ((nSeeds + 255) / 256), 256, 0, st>>>(d_seeds, nSeeds, n, pathgraphSeedStep(n));
}

unsigned pathgraphChecks(unsigned *line) { return fetchChecks(line); }

} // namespace ng
