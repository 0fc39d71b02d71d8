// pbsdf_glue.cu — C entry point over the REFERENCE's own CUDA tool (TEST INFRASTRUCTURE). oracle/Makefile `refcuda` compiles
// /root/reference/src/pbsdf.cu where it lies (unmodified, nvcc, sm_100a) together with this file into
// oracle/_ref/libpbsdf_ref.so: the oracle and the same-box baseline of the path-graph slice (hash grid + kNN).
// It needs a GPU, so it runs on the GPU box only (tests -m gpu, scripts/bench_pathgraph.py); nothing in nori_b200/ loads it.
#include "nori/shadingPoint.h"
#include <cuda_runtime.h>
#include <chrono>
#include <cstring>

extern "C" int refcuda_sizeof_spoint() { return (int) sizeof(SPoint); }

// The boundary records of include/nori_gpu.h ARE the reference's structs: checked here, where both headers are in sight, when this file is compiled.
#include "../../include/nori_gpu.h"
#include <cstddef>
#define SAME_FIELD(A, B, f) static_assert(offsetof(A, f) == offsetof(B, f) && sizeof(((A *) 0)->f) == sizeof(((B *) 0)->f), "layout of " #A "::" #f)
static_assert(sizeof(SPoint) == sizeof(NoriGpuSPoint) && sizeof(LPoint) == sizeof(NoriGpuLPoint) && sizeof(hparam) == sizeof(NoriGpuHashParams), "record sizes");
SAME_FIELD(SPoint, NoriGpuSPoint, pos); SAME_FIELD(SPoint, NoriGpuSPoint, wi); SAME_FIELD(SPoint, NoriGpuSPoint, wi_d); SAME_FIELD(SPoint, NoriGpuSPoint, wo);
SAME_FIELD(SPoint, NoriGpuSPoint, shN); SAME_FIELD(SPoint, NoriGpuSPoint, geoN); SAME_FIELD(SPoint, NoriGpuSPoint, diffuse); SAME_FIELD(SPoint, NoriGpuSPoint, specular);
SAME_FIELD(SPoint, NoriGpuSPoint, eLi); SAME_FIELD(SPoint, NoriGpuSPoint, eLd); SAME_FIELD(SPoint, NoriGpuSPoint, eta); SAME_FIELD(SPoint, NoriGpuSPoint, k);
SAME_FIELD(SPoint, NoriGpuSPoint, roughness); SAME_FIELD(SPoint, NoriGpuSPoint, pdf); SAME_FIELD(SPoint, NoriGpuSPoint, rrpdf); SAME_FIELD(SPoint, NoriGpuSPoint, nidx);
SAME_FIELD(SPoint, NoriGpuSPoint, groupIdx); SAME_FIELD(SPoint, NoriGpuSPoint, bsdf_type);
SAME_FIELD(LPoint, NoriGpuLPoint, L_directsample); SAME_FIELD(LPoint, NoriGpuLPoint, L_bsdfsample); SAME_FIELD(LPoint, NoriGpuLPoint, L_em);
SAME_FIELD(LPoint, NoriGpuLPoint, lightpdf); SAME_FIELD(LPoint, NoriGpuLPoint, bsdfpdf);
SAME_FIELD(hparam, NoriGpuHashParams, dim); SAME_FIELD(hparam, NoriGpuHashParams, cellsize); SAME_FIELD(hparam, NoriGpuHashParams, minb);
SAME_FIELD(hparam, NoriGpuHashParams, k); SAME_FIELD(hparam, NoriGpuHashParams, dk);
#undef SAME_FIELD

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

// ---- second slice: marginal densities + radiance scatter ---------------------------------------------------------------------
// The reference's kernels, device functions and __constant__ parameters are reached across translation units (the Makefile
// compiles with -rdc=true); nothing of them is restated here.
extern __global__ void allGPUPdfMarginal(const SPoint *sps, const int *neighbors, const int *knn, float *pdfmarginal);   // src/pbsdf.cu:1738
extern __device__ void pdf_device(const SPoint sp, const float *wi, float *pdf);                                         // src/pbsdf.cu:562
extern __device__ void bsdfeval_device(const SPoint sp, const float *wi, float *bsdf, bool t);                          // src/pbsdf.cu:464

namespace {
__global__ void glueEval(const SPoint *sps, int n, const float *wi, float *pdf, float *bsdf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float p = 0.0f, f[3] = {0.0f, 0.0f, 0.0f};           // what every caller in pbsdf.cu passes in
    pdf_device(sps[i], wi + 3 * (size_t) i, &p);
    bsdfeval_device(sps[i], wi + 3 * (size_t) i, f, false);
    pdf[i] = p; bsdf[3 * (size_t) i] = f[0]; bsdf[3 * (size_t) i + 1] = f[1]; bsdf[3 * (size_t) i + 2] = f[2];
}
}

/// pdf_device / bsdfeval_device (src/pbsdf.cu:464-628) at n records, one direction each.
extern "C" int refcuda_eval(const void *sps, int n, const float *wi, float *pdf, float *bsdf) {
    SPoint *cu_sps = nullptr; float *cu_wi = nullptr, *cu_pdf = nullptr, *cu_bsdf = nullptr;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) return -1;
    cudaMalloc((void **) &cu_wi, 12 * (size_t) n); cudaMalloc((void **) &cu_pdf, 4 * (size_t) n); cudaMalloc((void **) &cu_bsdf, 12 * (size_t) n);
    cudaMemcpy(cu_sps, sps, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMemcpy(cu_wi, wi, 12 * (size_t) n, cudaMemcpyHostToDevice);
    glueEval<<<(n + 127) / 128, 128>>>(cu_sps, n, cu_wi, cu_pdf, cu_bsdf);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(pdf, cu_pdf, 4 * (size_t) n, cudaMemcpyDeviceToHost); cudaMemcpy(bsdf, cu_bsdf, 12 * (size_t) n, cudaMemcpyDeviceToHost);
    cudaFree(cu_sps); cudaFree(cu_wi); cudaFree(cu_pdf); cudaFree(cu_bsdf);
    return e == cudaSuccess ? 0 : -3;
}

/// computeScatterAllOnGPURecord (src/pbsdf.cu:2317-2368) on n records and their neighbour rows (row_ints ints each, knn[i] used),
/// then allGPUPdfMarginal once more into a buffer of this glue (the driver frees its own before returning; the __constant__
/// sizes it set are still in place). blur / mc: iterations x n x 3 floats. seconds: wall time of the reference's driver alone
/// (it synchronises and copies its results to the host inside, as the tool does).
extern "C" int refcuda_scatter(const void *sps, int n, const int *neighbors, int row_ints, const int *knn, int iterations,
                               float *pdfmarginal, float *blur, float *mc, double *seconds) {
    SPoint *cu_sps = nullptr; int *cu_neighbors = nullptr, *cu_knn = nullptr; float *cu_pm = nullptr; size_t pitch_n = 0;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) return -1;
    cudaMemcpy(cu_sps, sps, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    if (cudaMallocPitch((void **) &cu_neighbors, &pitch_n, row_ints * sizeof(int), n) != cudaSuccess) return -2;
    cudaMemcpy2D(cu_neighbors, pitch_n, neighbors, row_ints * sizeof(int), row_ints * sizeof(int), n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_knn, sizeof(int) * (size_t) n); cudaMemcpy(cu_knn, knn, sizeof(int) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_pm, sizeof(float) * (size_t) n);
    ResultSpace ret;
    for (int i = 0; i < iterations; ++i) {
        ret.blur_results.push_back((val3f *) malloc(sizeof(val3f) * (size_t) n));
        ret.mc_results.push_back((val3f *) malloc(sizeof(val3f) * (size_t) n));
    }
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    computeScatterAllOnGPURecord(cu_sps, cu_neighbors, cu_knn, n, iterations, row_ints, ret, pitch_n);
    cudaDeviceSynchronize();
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    allGPUPdfMarginal<<<n / 512 + 1, 512>>>(cu_sps, cu_neighbors, cu_knn, cu_pm);
    cudaError_t e = cudaDeviceSynchronize();
    if (pdfmarginal) cudaMemcpy(pdfmarginal, cu_pm, sizeof(float) * (size_t) n, cudaMemcpyDeviceToHost);
    for (int i = 0; i < iterations; ++i) {
        static_assert(sizeof(val3f) == 12, "val3f is three floats");
        std::memcpy(blur + (size_t) i * 3 * n, ret.blur_results[i], sizeof(val3f) * (size_t) n);
        std::memcpy(mc + (size_t) i * 3 * n, ret.mc_results[i], sizeof(val3f) * (size_t) n);
        free(ret.blur_results[i]); free(ret.mc_results[i]);
    }
    cudaFree(cu_sps); cudaFree(cu_neighbors); cudaFree(cu_knn); cudaFree(cu_pm);
    return e == cudaSuccess ? 0 : -3;
}

/// computeMISAllOnGPURecord (src/pbsdf.cu:2922-2965; what the `pg` tool's "knn" mode runs) on n records and their neighbour rows.
/// blur / mc: iterations x n x 3 floats. seconds: wall time of the reference's driver alone.
extern "C" int refcuda_mis(const void *sps, int n, const int *neighbors, int row_ints, const int *knn, int iterations,
                           float *blur, float *mc, double *seconds) {
    SPoint *cu_sps = nullptr; int *cu_neighbors = nullptr, *cu_knn = nullptr; size_t pitch_n = 0;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) return -1;
    cudaMemcpy(cu_sps, sps, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    if (cudaMallocPitch((void **) &cu_neighbors, &pitch_n, row_ints * sizeof(int), n) != cudaSuccess) return -2;
    cudaMemcpy2D(cu_neighbors, pitch_n, neighbors, row_ints * sizeof(int), row_ints * sizeof(int), n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_knn, sizeof(int) * (size_t) n); cudaMemcpy(cu_knn, knn, sizeof(int) * (size_t) n, cudaMemcpyHostToDevice);
    ResultSpace ret;
    for (int i = 0; i < iterations; ++i) {
        ret.blur_results.push_back((val3f *) malloc(sizeof(val3f) * (size_t) n));
        ret.mc_results.push_back((val3f *) malloc(sizeof(val3f) * (size_t) n));
    }
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    computeMISAllOnGPURecord(cu_sps, cu_neighbors, cu_knn, n, iterations, row_ints, ret, pitch_n);
    cudaError_t e = cudaDeviceSynchronize();
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int i = 0; i < iterations; ++i) {
        std::memcpy(blur + (size_t) i * 3 * n, ret.blur_results[i], sizeof(val3f) * (size_t) n);
        std::memcpy(mc + (size_t) i * 3 * n, ret.mc_results[i], sizeof(val3f) * (size_t) n);
        free(ret.blur_results[i]); free(ret.mc_results[i]);
    }
    cudaFree(cu_sps); cudaFree(cu_neighbors); cudaFree(cu_knn);
    return e == cudaSuccess ? 0 : -3;
}

/// computeDirectScatterAllOnGPU on host arrays (the reference's driver, src/pbsdf.cu:2370-2395; direct: n x 3 floats) and, iterations >= 0,
/// computeScatterAllOnGPURecordWithDirectOpt (:2488-2540) on its result, as SPoints::computeScatterRadianceAOGWithProcessRecordingAndDirectLightOPT
/// chains them (include/nori/shadingPoint.h:462-476).
extern "C" int refcuda_direct(const void *sps, const void *lps, int n, const int *neighbors, int row_ints, const int *knn, const int *dknn,
                              int iterations, float *direct, float *blur, float *mc, double *seconds) {
    SPoint *cu_sps = nullptr; LPoint *cu_lps = nullptr; int *cu_neighbors = nullptr, *cu_knn = nullptr, *cu_dknn = nullptr; val3f *cu_direct = nullptr; size_t pitch_n = 0;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) return -1;
    cudaMemcpy(cu_sps, sps, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_lps, sizeof(LPoint) * (size_t) n); cudaMemcpy(cu_lps, lps, sizeof(LPoint) * (size_t) n, cudaMemcpyHostToDevice);
    if (cudaMallocPitch((void **) &cu_neighbors, &pitch_n, row_ints * sizeof(int), n) != cudaSuccess) return -2;
    cudaMemcpy2D(cu_neighbors, pitch_n, neighbors, row_ints * sizeof(int), row_ints * sizeof(int), n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_dknn, sizeof(int) * (size_t) n); cudaMemcpy(cu_dknn, dknn, sizeof(int) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_direct, sizeof(val3f) * (size_t) n);
    ResultSpace ret;
    if (iterations >= 0) {
        cudaMalloc((void **) &cu_knn, sizeof(int) * (size_t) n); cudaMemcpy(cu_knn, knn, sizeof(int) * (size_t) n, cudaMemcpyHostToDevice);
        for (int i = 0; i < iterations; ++i) {
            ret.blur_results.push_back((val3f *) malloc(sizeof(val3f) * (size_t) n));
            ret.mc_results.push_back((val3f *) malloc(sizeof(val3f) * (size_t) n));
        }
    }
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    computeDirectScatterAllOnGPU(cu_sps, cu_lps, cu_neighbors, cu_dknn, n, row_ints, cu_direct, pitch_n);
    if (iterations >= 0) computeScatterAllOnGPURecordWithDirectOpt(cu_sps, cu_neighbors, cu_knn, n, iterations, row_ints, cu_direct, ret, pitch_n);
    cudaError_t e = cudaDeviceSynchronize();
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    cudaMemcpy(direct, cu_direct, sizeof(val3f) * (size_t) n, cudaMemcpyDeviceToHost);
    for (int i = 0; i < iterations; ++i) {
        std::memcpy(blur + (size_t) i * 3 * n, ret.blur_results[i], sizeof(val3f) * (size_t) n);
        std::memcpy(mc + (size_t) i * 3 * n, ret.mc_results[i], sizeof(val3f) * (size_t) n);
        free(ret.blur_results[i]); free(ret.mc_results[i]);
    }
    cudaFree(cu_sps); cudaFree(cu_lps); cudaFree(cu_neighbors); cudaFree(cu_knn); cudaFree(cu_dknn); cudaFree(cu_direct);
    return e == cudaSuccess ? 0 : -3;
}

// ---- clustered variant ------------------------------------------------------------------------------------------------------------
// The drivers below read three __constant__ sizes that only the reference's cluster BUILDERS set (src/pbsdf.cu:3047-3049, 3259, 3331);
// with given clusters (SPoints::loadClusters, shadingPoint.h:626-635 — which leaves them unset too) this glue sets them itself.
extern __constant__ int d_sizeofsp;              // src/pbsdf.cu:36-42
extern __constant__ int d_num_clusters;
extern __constant__ int d_num_clusters_point;

/// SPoints::ClusterScatterWithDirectOptNR (include/nori/shadingPoint.h:600-624) on host arrays: ClusterDirect, MatrixElementsNumber,
/// precomputedMatrixElemtns, ClusterIterations3. sps[i].groupIdx = cluster of i or -1. direct / blur / mc: n x 3 floats.
extern "C" int refcuda_cluster(const void *sps, const void *lps, int n, const int *clusters, int n_listed, const int *offsets, int n_clusters,
                               int iterations, float *direct, float *blur, float *mc, double *seconds) {
    SPoint *cu_sps = nullptr; LPoint *cu_lps = nullptr; int *cu_clusters = nullptr, *cu_offsets = nullptr; val3f *cu_direct = nullptr;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) return -1;
    cudaMemcpy(cu_sps, sps, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_lps, sizeof(LPoint) * (size_t) n); cudaMemcpy(cu_lps, lps, sizeof(LPoint) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_clusters, sizeof(int) * (size_t) n_listed); cudaMemcpy(cu_clusters, clusters, sizeof(int) * (size_t) n_listed, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_offsets, sizeof(int) * (size_t) n_clusters); cudaMemcpy(cu_offsets, offsets, sizeof(int) * (size_t) n_clusters, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_direct, sizeof(val3f) * (size_t) n);
    cudaMemcpyToSymbol(d_sizeofsp, &n, sizeof(int));
    cudaMemcpyToSymbol(d_num_clusters, &n_clusters, sizeof(int));
    cudaMemcpyToSymbol(d_num_clusters_point, &n_listed, sizeof(int));
    ResultSpace ret;
    for (int i = 0; i < iterations; ++i) {
        ret.blur_results.push_back((val3f *) calloc((size_t) n, sizeof(val3f)));
        ret.mc_results.push_back((val3f *) calloc((size_t) n, sizeof(val3f)));
    }
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    ClusterDirect(cu_sps, cu_lps, cu_clusters, cu_offsets, n, n_clusters, cu_direct);
    size_t *cu_element_offset = nullptr;
    cudaMalloc((void **) &cu_element_offset, sizeof(size_t) * (size_t) n);
    const size_t elements = MatrixElementsNumber(cu_sps, cu_offsets, n_clusters, cu_element_offset, n);
    val3f *cu_elements = nullptr;
    if (cudaMalloc((void **) &cu_elements, sizeof(val3f) * elements) != cudaSuccess) return -2;
    precomputedMatrixElemtns(cu_sps, cu_clusters, cu_offsets, cu_element_offset, n_clusters, n, (int) elements, cu_elements);
    ClusterIterations3(cu_sps, cu_clusters, cu_offsets, n_clusters, cu_element_offset, n, cu_elements, cu_direct, iterations, ret);
    cudaError_t e = cudaDeviceSynchronize();
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    cudaMemcpy(direct, cu_direct, sizeof(val3f) * (size_t) n, cudaMemcpyDeviceToHost);
    std::memcpy(blur, ret.blur_results[iterations - 1], sizeof(val3f) * (size_t) n);
    std::memcpy(mc, ret.mc_results[iterations - 1], sizeof(val3f) * (size_t) n);
    for (int i = 0; i < iterations; ++i) { free(ret.blur_results[i]); free(ret.mc_results[i]); }
    cudaFree(cu_sps); cudaFree(cu_lps); cudaFree(cu_clusters); cudaFree(cu_offsets); cudaFree(cu_direct); cudaFree(cu_element_offset); cudaFree(cu_elements);
    return e == cudaSuccess ? 0 : -3;
}

namespace {
void glueGroupIdx(const SPoint *cu_sps, int n, int *out) {           // SPoint::groupIdx of every record (what the builders rewrite)
    SPoint *h = new SPoint[n];
    cudaMemcpy((void *) h, cu_sps, sizeof(SPoint) * (size_t) n, cudaMemcpyDeviceToHost);
    for (int i = 0; i < n; ++i) out[i] = h[i].groupIdx;
    delete[] h;
}
}

/// SPoints::BuildClusters (include/nori/shadingPoint.h:637-760) step by step on n positions with GIVEN seeds (the reference shuffles an
/// index list with std::random_shuffle; `seeds` takes the place of its first n_seeds entries), with the intermediate state the steps
/// leave — which depends on the order its atomics arrive in — handed out so that each step can be checked on its own inputs:
///   g1 (n): groupIdx after buildBatchClusters;  lists1 (n) / off1 (n_seeds): the member lists countMyClusters built from it;
///   g2 (n): groupIdx after the first subClusters, counts[0] clusters;  lists2 (n) / off2 (capacity cap): the lists of those;
///   g3 (n), lists3 (n), off3 (cap): the final clustering (second subClusters + FinalizeCluster), counts[1] clusters, counts[2] members listed.
extern "C" int refcuda_build_clusters(const float *pos, int n, const int *dim, const float *cellsize, const float *minb, const int *seeds, int n_seeds, int K,
                                      int *g1, int *lists1, int *off1, int *g2, int *lists2, int *off2, int *g3, int *lists3, int *off3, int cap, int *counts,
                                      double *seconds) {
    SPoint *h = new SPoint[n];
    for (int i = 0; i < n; ++i) { std::memset((void *) &h[i], 0, sizeof(SPoint)); h[i].pos[0] = pos[3 * i]; h[i].pos[1] = pos[3 * i + 1]; h[i].pos[2] = pos[3 * i + 2]; h[i].bsdf_type = 'd'; }
    SPoint *cu_sps = nullptr;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) { delete[] h; return -1; }
    cudaMemcpy(cu_sps, h, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    delete[] h;
    hparam hp{{dim[0], dim[1], dim[2]}, {cellsize[0], cellsize[1], cellsize[2]}, {minb[0], minb[1], minb[2]}, K, K};
    int *cu_clusters = nullptr, *cu_np = nullptr, *cu_small = nullptr, *cu_new1 = nullptr, *cu_small2 = nullptr, *cu_new2 = nullptr;
    cudaMalloc((void **) &cu_clusters, sizeof(int) * (size_t) n); cudaMemset(cu_clusters, 0xff, sizeof(int) * (size_t) n);
    cudaMalloc((void **) &cu_np, sizeof(int) * (size_t) n_seeds);
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    buildBatchClusters(cu_sps, n, n_seeds, hp, n_seeds, const_cast<int *>(seeds), cu_clusters, cu_np);
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    glueGroupIdx(cu_sps, n, g1);
    t0 = std::chrono::steady_clock::now();
    cudaMalloc((void **) &cu_small, sizeof(int) * (size_t) n_seeds);
    const int new1 = countMyClusters(cu_sps, n, hp, n_seeds, cu_clusters, cu_np, cu_small);
    sec += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (new1 > cap) return -4;
    cudaMemcpy(lists1, cu_clusters, sizeof(int) * (size_t) n, cudaMemcpyDeviceToHost);
    cudaMemcpy(off1, cu_np, sizeof(int) * (size_t) n_seeds, cudaMemcpyDeviceToHost);
    t0 = std::chrono::steady_clock::now();
    cudaMalloc((void **) &cu_new1, sizeof(int) * (size_t) new1);
    subClusters(cu_sps, n, hp, n_seeds, new1, cu_clusters, cu_np, cu_new1, cu_small);
    sec += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    glueGroupIdx(cu_sps, n, g2);
    t0 = std::chrono::steady_clock::now();
    cudaMalloc((void **) &cu_small2, sizeof(int) * (size_t) new1);
    const int new2 = countMyClusters(cu_sps, n, hp, new1, cu_clusters, cu_new1, cu_small2);
    sec += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (new2 > cap) return -4;
    cudaMemcpy(lists2, cu_clusters, sizeof(int) * (size_t) n, cudaMemcpyDeviceToHost);
    cudaMemcpy(off2, cu_new1, sizeof(int) * (size_t) new1, cudaMemcpyDeviceToHost);
    t0 = std::chrono::steady_clock::now();
    cudaMalloc((void **) &cu_new2, sizeof(int) * (size_t) new2);
    subClusters(cu_sps, n, hp, new1, new2, cu_clusters, cu_new1, cu_new2, cu_small2);
    int lastSize = 0;
    cudaMemcpy(&lastSize, cu_new2 + new2 - 1, sizeof(int), cudaMemcpyDeviceToHost);
    FinalizeCluster(cu_sps, n, hp, new2, cu_clusters, cu_new2);
    cudaError_t e = cudaDeviceSynchronize();
    sec += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (seconds) *seconds = sec;
    glueGroupIdx(cu_sps, n, g3);
    cudaMemcpy(lists3, cu_clusters, sizeof(int) * (size_t) n, cudaMemcpyDeviceToHost);
    cudaMemcpy(off3, cu_new2, sizeof(int) * (size_t) new2, cudaMemcpyDeviceToHost);
    counts[0] = new1; counts[1] = new2; counts[2] = off3[new2 - 1] + lastSize;
    cudaFree(cu_sps); cudaFree(cu_clusters); cudaFree(cu_np); cudaFree(cu_small); cudaFree(cu_new1); cudaFree(cu_small2); cudaFree(cu_new2);
    return e == cudaSuccess ? 0 : -3;
}

/// The tool's cluster modes without direct light on host arrays. order 0: SPoints::ClusterScatter2 (MatrixElementsNumber, precomputedMatrixElemtns,
/// ClusterIterations); order 1: SPoints::ClusterScatter (computeClusterScatterAllOnGPURecord). blur / mc: iterations x n x 3 floats.
extern "C" int refcuda_cluster_scatter(int order, const void *sps, int n, const int *clusters, int n_listed, const int *offsets, int n_clusters, int iterations,
                                       float *blur, float *mc, double *seconds) {
    SPoint *cu_sps = nullptr; int *cu_clusters = nullptr, *cu_offsets = nullptr;
    if (cudaMalloc((void **) &cu_sps, sizeof(SPoint) * (size_t) n) != cudaSuccess) return -1;
    cudaMemcpy(cu_sps, sps, sizeof(SPoint) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_clusters, sizeof(int) * (size_t) n_listed); cudaMemcpy(cu_clusters, clusters, sizeof(int) * (size_t) n_listed, cudaMemcpyHostToDevice);
    cudaMalloc((void **) &cu_offsets, sizeof(int) * (size_t) n_clusters); cudaMemcpy(cu_offsets, offsets, sizeof(int) * (size_t) n_clusters, cudaMemcpyHostToDevice);
    cudaMemcpyToSymbol(d_sizeofsp, &n, sizeof(int));
    cudaMemcpyToSymbol(d_num_clusters, &n_clusters, sizeof(int));
    cudaMemcpyToSymbol(d_num_clusters_point, &n_listed, sizeof(int));
    ResultSpace ret;
    for (int i = 0; i < iterations; ++i) {
        ret.blur_results.push_back((val3f *) calloc((size_t) n, sizeof(val3f)));
        ret.mc_results.push_back((val3f *) calloc((size_t) n, sizeof(val3f)));
    }
    cudaDeviceSynchronize();
    auto t0 = std::chrono::steady_clock::now();
    if (order == 1) computeClusterScatterAllOnGPURecord(cu_sps, n_clusters, cu_clusters, cu_offsets, n, iterations, ret);
    else {
        size_t *cu_element_offset = nullptr;
        cudaMalloc((void **) &cu_element_offset, sizeof(size_t) * (size_t) n);
        const size_t elements = MatrixElementsNumber(cu_sps, cu_offsets, n_clusters, cu_element_offset, n);
        val3f *cu_elements = nullptr;
        if (cudaMalloc((void **) &cu_elements, sizeof(val3f) * elements) != cudaSuccess) return -2;
        precomputedMatrixElemtns(cu_sps, cu_clusters, cu_offsets, cu_element_offset, n_clusters, n, (int) elements, cu_elements);
        ClusterIterations(cu_sps, cu_clusters, cu_offsets, n_clusters, cu_element_offset, n, cu_elements, iterations, ret);
        cudaFree(cu_element_offset); cudaFree(cu_elements);
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int i = 0; i < iterations; ++i) {
        std::memcpy(blur + (size_t) i * 3 * n, ret.blur_results[i], sizeof(val3f) * (size_t) n);
        std::memcpy(mc + (size_t) i * 3 * n, ret.mc_results[i], sizeof(val3f) * (size_t) n);
        free(ret.blur_results[i]); free(ret.mc_results[i]);
    }
    cudaFree(cu_sps); cudaFree(cu_clusters); cudaFree(cu_offsets);
    return e == cudaSuccess ? 0 : -3;
}
