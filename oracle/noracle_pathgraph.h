// noracle_pathgraph.h — CPU restatement of the path-graph tool's hash grid + k-nearest-neighbour search (TEST
// INFRASTRUCTURE; never part of the product). Follows the reference's CUDA code statement for statement:
//   getKeys / getKeysAndDirs / getKey     src/pbsdf.cu:630-665
//   checkOneDim / makesureGrid            src/pbsdf.cu:762-777
//   dev_distance                          src/pbsdf.cu:667-675
//   batchNearestNeighbor                  src/pbsdf.cu:1089-1178  (candidate block, k / dk counts, "myself first")
//   pointInCell + the cell offsets        src/pbsdf.cu:820-826, 3195-3216
// The reference is device code compiled by nvcc with its default -fmad=true: `a - b * c` and `s += d * d` are single-rounding
// fused multiply-adds there. They are written as fmaf() here so that an x86-64 build computes the same bits.
// Pinned by: oracle/_ref/libpbsdf_ref.so = the reference's own src/pbsdf.cu compiled for sm_100a (oracle/Makefile refcuda),
// run on the GPU box (tests/test_pathgraph.py::test_knn_vs_compiled_reference_cuda) and the fixture it produced there
// (tests/golden/pathgraph_ref.npz, oracle/make_pathgraph_golden.py).
// The one place where the reference is not a function of its input: its quickselect (findKSmallest, pbsdf.cu:743-760)
// resolves EQUAL distances at the k-th place arbitrarily, can stop early when equal distances occur among a row's candidates
// (`if(lb==p && dists[lb] == dists[k] && flag == 0) break`: it then keeps a farther point instead of a nearer one; < 1 % of the
// rows of a point set with 200 exact duplicates, none without duplicates), and the order of its result rows is an artefact of
// the partitioning. Grids whose fullest cell holds fewer than 16 points crash it (integer division by zero, pbsdf.cu:2855-2861). This restatement selects by (distance, index) and returns rows nearest first; comparisons against the
// reference are therefore set comparisons that allow members at exactly the boundary distance to differ.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>
#include <algorithm>

namespace noracle {

struct HashParams { int dim[3]; float cellsize[3]; float minb[3]; int k, dk; };   // hparam, include/nori/shadingPoint.h:18-24

inline int pgCell(const HashParams &h, int c, float x) {
    return std::max(0, std::min((int) std::floor((x - h.minb[c]) / h.cellsize[c]), h.dim[c] - 1));
}
inline int pgKey(const HashParams &h, int kx, int ky, int kz) { return h.dim[2] * h.dim[1] * kx + h.dim[2] * ky + kz; }
inline int pgDir(const HashParams &h, int c, float x, int key) {
    const float t = fmaf(-h.cellsize[c], (float) key, x) - h.minb[c];
    return (double) t > (double) h.cellsize[c] / 2.0 ? 1 : -1;
}
inline void pgCheckOneDim(int &l, int &hgh, int dim) {
    if (l == 0 && (hgh - l < 1) && hgh < dim - 1) { hgh += 1; return; }
    if (hgh == dim - 1 && (hgh - l < 1) && l > 0) { l -= 1; return; }
}
inline float pgDistance(const float *a, const float *b) {
    float dist = 0.0f;
    for (int i = 0; i < 3; ++i) { const float d = a[i] - b[i]; dist = fmaf(d, d, dist); }
    return std::sqrt(dist);
}

/// points: n x 3. neighbors: n x max(k, dk) (unused entries -1), knn / dknn: n. dists (optional): the distances of the rows.
inline void pathgraphKnn(const float *pts, int n, const HashParams &h, int32_t *neighbors, int32_t *knn, int32_t *dknn, float *dists = nullptr) {
    const int nCells = h.dim[0] * h.dim[1] * h.dim[2];
    const int maxK = std::max(h.k, h.dk), minK = std::min(h.k, h.dk);
    std::vector<int> start(nCells + 1, 0), key(n);
    for (int i = 0; i < n; ++i) {
        key[i] = pgKey(h, pgCell(h, 0, pts[3 * i]), pgCell(h, 1, pts[3 * i + 1]), pgCell(h, 2, pts[3 * i + 2]));
        start[key[i] + 1]++;
    }
    for (int c = 0; c < nCells; ++c) start[c + 1] += start[c];
    std::vector<int> fill(start.begin(), start.end() - 1), entry(n);
    for (int i = 0; i < n; ++i) entry[fill[key[i]]++] = i;
    std::vector<std::pair<float, int> > cand;
    for (int i = 0; i < n; ++i) {
        const float *p = pts + 3 * (size_t) i;
        int lo[3], hi[3];
        for (int c = 0; c < 3; ++c) {
            const int k = pgCell(h, c, p[c]), d = pgDir(h, c, p[c], k);
            lo[c] = std::max(0, std::min(k, k + d)); hi[c] = std::min(h.dim[c] - 1, std::max(k, k + d));
            pgCheckOneDim(lo[c], hi[c], h.dim[c]);
        }
        cand.clear();
        for (int ix = lo[0]; ix <= hi[0]; ++ix) for (int iy = lo[1]; iy <= hi[1]; ++iy) for (int iz = lo[2]; iz <= hi[2]; ++iz) {
            const int c = pgKey(h, ix, iy, iz);
            for (int e = start[c]; e < start[c + 1]; ++e) cand.push_back(std::make_pair(pgDistance(pts + 3 * (size_t) entry[e], p), entry[e]));
        }
        const int total = (int) cand.size(), m = std::min(maxK, total);
        std::partial_sort(cand.begin(), cand.begin() + m, cand.end());          // (distance, index) lexicographic
        int32_t *row = neighbors + (size_t) i * maxK;
        for (int j = 0; j < maxK; ++j) { row[j] = j < m ? cand[j].second : -1; if (dists) dists[(size_t) i * maxK + j] = j < m ? cand[j].first : -1.0f; }
        for (int j = 0; j < m; ++j) if (row[j] == i) {                           // "ensure the first point is myself"
            row[j] = row[0]; row[0] = i;
            if (dists) std::swap(dists[(size_t) i * maxK + j], dists[(size_t) i * maxK]);
            break;
        }
        if (maxK < total) { knn[i] = h.k; dknn[i] = h.dk; }
        else if (minK < total) { if (h.dk > h.k) { knn[i] = h.k; dknn[i] = total; } else { knn[i] = total; dknn[i] = h.dk; } }
        else { knn[i] = total; dknn[i] = total; }
    }
}


// ---- SPoints::BuildClusters (include/nori/shadingPoint.h:637-760) -------------------------------------------------------------------
// Seeds (the first n / K + 1 entries of a shuffled index list: the reference draws it with libstdc++'s random_shuffle under srand(1994);
// here the seeds are an INPUT) -> every vertex joins the nearest seed found through a hashed table of the seeds' grid cells
// (buildBatchClusters, src/pbsdf.cu:3555-3622) -> two rounds that split every cluster of more than K members around its first
// size / K + 1 members (countMyClusters + subClusters, :3236-3310) -> member lists (FinalizeCluster, :3312-3332).
// The reference fills its lists with atomics (SaveClusters, :987-1006), so the order of a cluster's members — and with it WHICH members
// seed the split — depends on scheduling. The step functions below take the lists as they are given (to be checked against the reference's
// own intermediate state); the driver at the end builds them in index order, which is one of the orders the reference can produce.

/// hashbmap (src/pbsdf.cu:798-800)
inline uint32_t pgHashBucket(uint32_t key, uint32_t tableSize) { return (key * 5039u + 39916801u) % tableSize; }

/// buildBatchClusters: groupIdx[i] = index (into seeds) of the nearest seed among the table buckets of the 2 x 2 x 2 cell block around
/// vertex i (Cluster, src/pbsdf.cu:1008-1044: `dist <= mindistance`, so among equals the last one visited wins; buckets hold the seeds
/// in ascending order here), or -1 if none lies within twice the cell diagonal. Table size = nSeeds (shadingPoint.h:667).
inline void pgClusterAssign(const float *pts, int n, const HashParams &h, const int32_t *seeds, int nSeeds, int32_t *groupIdx) {
    const uint32_t table = (uint32_t) nSeeds;
    std::vector<std::vector<int> > bucket(nSeeds);
    for (int s = 0; s < nSeeds; ++s) {
        const float *p = pts + 3 * (size_t) seeds[s];
        bucket[pgHashBucket((uint32_t) pgKey(h, pgCell(h, 0, p[0]), pgCell(h, 1, p[1]), pgCell(h, 2, p[2])), table)].push_back(s);
    }
    const float diag = std::sqrt(h.cellsize[0] * h.cellsize[0] + h.cellsize[1] * h.cellsize[1] + h.cellsize[2] * h.cellsize[2]);
    const float minDist = (float) ((double) diag * 2.0);
    for (int i = 0; i < n; ++i) {
        const float *p = pts + 3 * (size_t) i;
        int lo[3], hi[3];
        for (int c = 0; c < 3; ++c) {
            const int k = pgCell(h, c, p[c]), d = pgDir(h, c, p[c], k);
            lo[c] = std::max(0, std::min(k, k + d)); hi[c] = std::min(h.dim[c] - 1, std::max(k, k + d));
            pgCheckOneDim(lo[c], hi[c], h.dim[c]);
        }
        float best = minDist; int who = -1;
        for (int ix = lo[0]; ix <= hi[0]; ++ix) for (int iy = lo[1]; iy <= hi[1]; ++iy) for (int iz = lo[2]; iz <= hi[2]; ++iz) {
            const std::vector<int> &b = bucket[pgHashBucket((uint32_t) pgKey(h, ix, iy, iz), table)];
            for (size_t e = 0; e < b.size(); ++e) {
                const float d = pgDistance(pts + 3 * (size_t) seeds[b[e]], p);
                if (d <= best) { best = d; who = b[e]; }
            }
        }
        groupIdx[i] = who;
    }
}

/// Member lists in index order + offsets (what SaveClusters leaves when its atomics arrive in index order). Returns the number listed.
inline int pgClusterLists(const int32_t *groupIdx, int n, int nClusters, int32_t *lists, int32_t *offsets) {
    std::vector<int> count(nClusters + 1, 0);
    for (int i = 0; i < n; ++i) if (groupIdx[i] >= 0) count[groupIdx[i] + 1]++;
    for (int c = 0; c < nClusters; ++c) count[c + 1] += count[c];
    for (int c = 0; c < nClusters; ++c) offsets[c] = count[c];
    std::vector<int> fill(count.begin(), count.end() - 1);
    for (int i = 0; i < n; ++i) if (groupIdx[i] >= 0) lists[fill[groupIdx[i]]++] = i;
    return count[nClusters];
}

/// countClusters + SubdivideClusters (src/pbsdf.cu:910-920, 942-974): cluster c of `size` members becomes size / K + 1 clusters around its
/// first size / K + 1 listed members; a member joins the nearest of them (strict <: the first of equals; nothing nearer than 262144 -> the
/// first). groupIdx is rewritten to the new numbering (cluster c's children are consecutive). Returns the new number of clusters.
inline int pgClusterSubdivide(const float *pts, int n, int32_t *groupIdx, const int32_t *lists, int nListed, const int32_t *offsets, int nClusters, int K) {
    std::vector<int> first(nClusters + 1, 0);
    for (int c = 0; c < nClusters; ++c) {
        const int size = (c == nClusters - 1 ? nListed : offsets[c + 1]) - offsets[c];
        first[c + 1] = first[c] + std::max(size / K + 1, 1);
    }
    for (int i = 0; i < n; ++i) {
        const int c = groupIdx[i];
        if (c < 0) continue;
        const int pnum = first[c + 1] - first[c];
        int g = first[c];
        if (pnum > 1) {
            float best = 262144.0f; int who = 0;
            for (int t = 0; t < pnum; ++t) {
                const float d = pgDistance(pts + 3 * (size_t) lists[offsets[c] + t], pts + 3 * (size_t) i);
                if (d < best) { best = d; who = t; }
            }
            g += who;
        }
        groupIdx[i] = g;
    }
    return first[nClusters];
}

/// BuildClusters with index-ordered lists. lists: n ints, offsets: room for the final number of clusters (<= 2 n / K + nSeeds: every split adds
/// at most n / K; n + nSeeds is always enough for K >= 2). Returns the number of clusters; *nListed = members listed (vertices with groupIdx != -1).
inline int pathgraphBuildClusters(const float *pts, int n, const HashParams &h, const int32_t *seeds, int nSeeds, int K, int32_t *groupIdx,
                                  int32_t *lists, int32_t *offsets, int *nListed) {
    pgClusterAssign(pts, n, h, seeds, nSeeds, groupIdx);
    int nClusters = nSeeds;
    for (int round = 0; round < 2; ++round) {
        std::vector<int32_t> off(nClusters);
        const int listed = pgClusterLists(groupIdx, n, nClusters, lists, off.data());
        nClusters = pgClusterSubdivide(pts, n, groupIdx, lists, listed, off.data(), nClusters, K);
    }
    *nListed = pgClusterLists(groupIdx, n, nClusters, lists, offsets);
    return nClusters;
}

} // namespace noracle
