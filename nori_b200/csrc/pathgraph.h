// pathgraph.h — host-side interface of pathgraph.cu (hash grid + kNN over path vertices; reference src/pbsdf.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/nori_gpu.h"

namespace ng {

constexpr int kPgMaxClusterSize = 8192;   // clustered drivers: one warp takes a cluster's size^2 pairs (67 M at the limit); the tool's builder cuts clusters down to ~K members
constexpr int kPgMaxK = 200;          // max(k, dk): the per-thread heaps of a 128-thread CTA must fit shared memory (8 * K * 128 bytes)

/// KPara (src/pbsdf.cu:16-24) = hparam (include/nori/shadingPoint.h:18-24) + derived sizes.
struct PgParams {
    float minb[3], cell[3];
    int dim[3];
    int k, dk, maxK;
    int nCells;
};

size_t pgScratchBytes(int n, int nCells);
bool launchPathgraphKnn(const void *d_pos, size_t strideBytes, int n, const PgParams &P, void *d_scratch, int *d_neighbors, int pitchInts,
                        int *d_knn, int *d_dknn, cudaStream_t st, std::string &err);
/// SPoints::BuildClusters (include/nori/shadingPoint.h:637-760) with given seeds; deterministic (members in index order).
size_t pgBuildClustersScratchBytes(int n, int nSeeds, int cap);
bool launchPathgraphBuildClusters(const void *d_pos, size_t strideBytes, int n, const PgParams &P, const int *d_seeds, int nSeeds, int K, void *d_scratch,
                                  int *d_groupIdx, int *d_lists, int *d_offsets, int cap, int *nClusters, int *nListed, cudaStream_t st, std::string &err,
                                  int *maxCluster = nullptr);
/// seeds[s] = (s * step) mod n, step = pathgraphSeedStep(n): a fixed scattered choice of distinct vertices (the reference: the head of a shuffled list).
long long pathgraphSeedStep(int n);
void launchPathgraphStrideSeeds(int *d_seeds, int nSeeds, int n, cudaStream_t st);
unsigned pathgraphChecks(unsigned *line);

/// pathgraph_scatter.cu: marginal densities + iterated radiance scatter over the kNN graph (src/pbsdf.cu:2317-2368).
size_t pgScatterScratchBytes(int n);
bool launchPathgraphScatter(const NoriGpuSPoint *d_sps, int n, const int *d_neighbors, int pitchInts, const int *d_knn, int iterations,
                            void *d_scratch, float *d_pdfmarginal, float *d_blur, float *d_mc, cudaStream_t st, std::string &err,
                            const float *d_direct = nullptr);
size_t pgMisScratchBytes(int n, int rowInts);
bool launchPathgraphMis(const NoriGpuSPoint *d_sps, int n, const int *d_neighbors, int pitchInts, const int *d_knn, int iterations,
                        void *d_scratch, float *d_blur, float *d_mc, cudaStream_t st, std::string &err);
size_t pgDirectScratchBytes(int n);
bool launchPathgraphDirect(const NoriGpuSPoint *d_sps, const NoriGpuLPoint *d_lps, int n, const int *d_neighbors, int pitchInts, const int *d_knn,
                           void *d_scratch, float *d_direct, cudaStream_t st, std::string &err);
/// clustered variant (SPoints::ClusterScatterWithDirectOptNR): the lists are validated once, then trusted.
void launchPathgraphClusterValidate(const NoriGpuSPoint *d_sps, int n, const int *d_clusters, int nListed, const int *d_offsets, int nClusters,
                                    int *d_listed, int *d_flag2, cudaStream_t st);
size_t pgClusterScratchBytes(int n, int nClusters);
bool launchPathgraphCluster(const NoriGpuSPoint *d_sps, const NoriGpuLPoint *d_lps, int n, const int *d_clusters, int nListed, const int *d_offsets,
                            int nClusters, int iterations, void *d_scratch, float *d_direct, float *d_blur, float *d_mc, cudaStream_t st, std::string &err,
                            const int *d_clusterOf = nullptr);
bool launchPathgraphClusterScatter(const NoriGpuSPoint *d_sps, int n, const int *d_clusters, int nListed, const int *d_offsets, int nClusters, int iterations,
                                   int scatterOrder, void *d_scratch, float *d_blur, float *d_mc, cudaStream_t st, std::string &err);
void launchPathgraphValidate(const int *d_neighbors, int pitchInts, int rowInts, const int *d_knn, int n, int *d_flag, cudaStream_t st);
/// Whole pipeline on the device (nori_gpu_pathgraph_render): bounding box of the records, image composition from first vertices.
void launchPathgraphBounds(const NoriGpuSPoint *d_sps, int n, unsigned int *d_bounds6, cudaStream_t st);
float pathgraphOrderedToFloat(unsigned int ordered);
void launchPathgraphCompose(const NoriGpuSPoint *d_sps, const unsigned long long *d_offsets, const unsigned int *d_counts, const float *d_blur, const float *d_mc,
                            int nPixels, unsigned int spp, float *d_images, cudaStream_t st, const float *d_direct = nullptr);
unsigned pathgraphScatterChecks(unsigned *line);

} // namespace ng
