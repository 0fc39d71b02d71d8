// pbsdf_glue.cu — C entry point over the REFERENCE's own CUDA tool (TEST INFRASTRUCTURE). oracle/Makefile `refcuda` compiles
// /root/reference/src/pbsdf.cu where it lies (unmodified, nvcc, sm_100a) together with this file into
// oracle/_ref/libpbsdf_ref.so: the oracle and the same-box baseline of the path-graph slice (hash grid + kNN).
// It needs a GPU, so it runs on the GPU box only (tests -m gpu, scripts/bench_pathgraph.py); nothing in nori_b200/ loads it.
#include "nori/shadingPoint.h"
#include <cuda_runtime.h>
#include <chrono>
#include <cstring>

extern "C" int refcuda_sizeof_spoint() { return (int) sizeof(SPoint); }

/// SPoints::BuildKNN (include/nori/shadingPoint.h:776-806) on n positions: device SPoint array, pitched neighbour rows,
/// buildHashGridAndFindKNN (src/pbsdf.cu:3173-3235). neighbors: n x max(k, dk) ints (rows beyond knn/dknn are whatever the
/// reference left in its uninitialised allocation; this glue clears the array first so that they read -1).
/// seconds: wall time of buildHashGridAndFindKNN alone (it synchronises internally), the reference's own pipeline incl. its host scan.
extern "C" int refcuda_knn(const float *pos, int n, const int *dim, const float *cellsize, const float *minb, int k, int dk,
                           int *neighbors, int *knn, int *dknn, double *seconds) {
    SPoint *h = new SPoint[n];
    for (int i = 0; i < n; ++i) { std::memset((void *) &h[i], 0, sizeof(SPoint)); h[i].pos[0] = pos[3 * i]; h[i].pos[1] = pos[3 * i + 1]; h[i].pos[2] = pos[3 * i + 2]; h[i].bsdf_type = 'd'; }
    SPoint *cu_sps = nullptr; int *cu_neighbors = nullptr, *cu_knn = nullptr, *cu_dknn = nullptr; size_t pitch_n = 0;
    const int maxK = k > dk ? k : dk;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) { delete[] h; return -1; }
    cudaMemcpy(cu_sps, h, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    delete[] h;
    if (cudaMallocPitch((void **) &cu_neighbors, &pitch_n, maxK * sizeof(int), n) != cudaSuccess) return -2;
    cudaMemset2D(cu_neighbors, pitch_n, 0xff, maxK * sizeof(int), n);
    cudaMalloc((void **) &cu_knn, sizeof(int) * (size_t) n); cudaMalloc((void **) &cu_dknn, sizeof(int) * (size_t) n);
    hparam hp{{dim[0], dim[1], dim[2]}, {cellsize[0], cellsize[1], cellsize[2]}, {minb[0], minb[1], minb[2]}, k, dk};
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    buildHashGridAndFindKNN(cu_sps, hp, cu_neighbors, n, cu_knn, cu_dknn, pitch_n);
    cudaDeviceSynchronize();
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    cudaError_t e = cudaGetLastError();
    cudaMemcpy2D(neighbors, maxK * sizeof(int), cu_neighbors, pitch_n, maxK * sizeof(int), n, cudaMemcpyDeviceToHost);
    cudaMemcpy(knn, cu_knn, sizeof(int) * (size_t) n, cudaMemcpyDeviceToHost);
    cudaMemcpy(dknn, cu_dknn, sizeof(int) * (size_t) n, cudaMemcpyDeviceToHost);
    cudaFree(cu_sps); cudaFree(cu_neighbors); cudaFree(cu_knn); cudaFree(cu_dknn);
    return e == cudaSuccess ? 0 : -3;
}
