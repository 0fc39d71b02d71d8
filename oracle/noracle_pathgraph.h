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

} // namespace noracle
