// objdedup.cu — the step before the path: OBJ corner de-duplication and vertex gather on the device.
//
// The reference's loader keeps a hash map from (p, uv, n) index triples to vertex ids and hands out ids in order of FIRST USE
// while it reads the `f` lines (src/obj.cpp:32-48 OBJVertex + hash, :96-117 the face loop), then gathers m_V / m_N / m_UV through
// the unique corners (:136-161). For a 10 M-triangle mesh that is 30 M hash probes on one host thread. Here:
//   1. the corners are sorted by (p, uv, n) with three stable 32-bit radix passes (least significant field first), carrying the
//      corner number; inside a run of equal triples the corner numbers stay ascending, so the head of a run is its first use;
//   2. run heads are flagged and scanned into run numbers;
//   3. the runs are sorted by the corner number of their head: the rank of a run in that order IS the id the reference's map
//      would have handed out;
//   4. every corner receives the id of its run (the index buffer m_F), every run gathers its position / normal / texcoord.
// Result: F, V, N, UV bit-equal to the host loader's (tests/test_objdedup.py, and through it to the compiled reference's:
// tests/test_host_cpp.py::test_descriptors_bit_equal_to_reference_parser).
#include "engine.h"
#include <cub/cub.cuh>
#include <algorithm>

namespace ng {

namespace {

__global__ void k_ddInit(const uint32_t *__restrict__ corners, uint32_t n, int field, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = corners[3 * (size_t) i + field];
    vals[i] = i;
}
__global__ void k_ddGatherKey(const uint32_t *__restrict__ corners, const uint32_t *__restrict__ order, uint32_t n, int field, uint32_t *__restrict__ keys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = corners[3 * (size_t) order[i] + field];
}
/// head[i] = 1 where the sorted corner i starts a run of equal (p, uv, n) triples; also range-checks the indices
/// (the reference's positions.at(p - 1) etc. throw: src/obj.cpp:139-158).
__global__ void k_ddHeads(const uint32_t *__restrict__ corners, const uint32_t *__restrict__ order, uint32_t n, uint32_t nP, uint32_t nT, uint32_t nN,
                          uint32_t *__restrict__ head, unsigned int *bad) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t *c = corners + 3 * (size_t) order[i];
    uint32_t h = 1;
    if (i > 0) { const uint32_t *d = corners + 3 * (size_t) order[i - 1]; h = c[0] != d[0] || c[1] != d[1] || c[2] != d[2]; }
    head[i] = h;
    if (h) {
        const bool ok = c[0] >= 1 && c[0] <= nP && (nT == 0 || (c[1] >= 1 && c[1] <= nT)) && (nN == 0 || (c[2] >= 1 && c[2] <= nN));
        if (!ok) atomicAdd(bad, 1u);
    }
}
/// firstCorner[run] = corner number of the run's head; runIds[run] = run.
__global__ void k_ddRuns(const uint32_t *__restrict__ order, const uint32_t *__restrict__ head, const uint32_t *__restrict__ runOf /* exclusive scan of head */, uint32_t n,
                         uint32_t *__restrict__ firstCorner, uint32_t *__restrict__ runIds) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !head[i]) return;
    const uint32_t r = runOf[i];
    firstCorner[r] = order[i]; runIds[r] = r;
}
__global__ void k_ddRank(const uint32_t *__restrict__ runSorted, uint32_t nRuns, uint32_t *__restrict__ newId) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < nRuns) newId[runSorted[r]] = r;
}
__global__ void k_ddIndex(const uint32_t *__restrict__ order, const uint32_t *__restrict__ head, const uint32_t *__restrict__ runOf, const uint32_t *__restrict__ newId, uint32_t n,
                          uint32_t *__restrict__ F) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F[order[i]] = newId[runOf[i] - (head[i] ? 0u : 1u)];            // runOf is exclusive: non-heads belong to the run before their count
}
/// m_V / m_N / m_UV (src/obj.cpp:136-161): vertex r = the corner that used it first.
__global__ void k_ddGather(const uint32_t *__restrict__ corners, const uint32_t *__restrict__ firstSorted, uint32_t nRuns,
                           const float *__restrict__ pos, const float *__restrict__ tex, const float *__restrict__ nrm,
                           float *__restrict__ V, float *__restrict__ UV, float *__restrict__ N) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nRuns) return;
    const uint32_t *c = corners + 3 * (size_t) firstSorted[r];
    const size_t p = (size_t) c[0] - 1;
    V[3 * (size_t) r] = pos[3 * p]; V[3 * (size_t) r + 1] = pos[3 * p + 1]; V[3 * (size_t) r + 2] = pos[3 * p + 2];
    if (nrm) { const size_t k = (size_t) c[2] - 1; N[3 * (size_t) r] = nrm[3 * k]; N[3 * (size_t) r + 1] = nrm[3 * k + 1]; N[3 * (size_t) r + 2] = nrm[3 * k + 2]; }
    if (tex) { const size_t k = (size_t) c[1] - 1; UV[2 * (size_t) r] = tex[2 * k]; UV[2 * (size_t) r + 1] = tex[2 * k + 1]; }
}

} // namespace

#define DD_CHECK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = std::string(#call) + " failed: " + cudaGetErrorString(e_); cleanup(); return false; } } while (0)

/// All pointers are device pointers. corners: n x (p, uv, n), 1-based (absent fields may hold anything when nT / nN == 0).
/// F: n ids; V / N / UV: room for n vertices (N, UV may be null). *nVertsOut = number of unique corners.
bool objDedup(const uint32_t *corners, uint32_t n, const float *pos, uint32_t nP, const float *tex, uint32_t nT, const float *nrm, uint32_t nN,
              uint32_t *F, float *V, float *N, float *UV, uint32_t *nVertsOut, cudaStream_t st, std::string &err) {
    uint32_t *keys[2] = {nullptr, nullptr}, *vals[2] = {nullptr, nullptr}, *head = nullptr, *runOf = nullptr, *aux[3] = {nullptr, nullptr, nullptr};
    void *tmp = nullptr; unsigned int *bad = nullptr;
    auto cleanup = [&]() { for (auto p : {keys[0], keys[1], vals[0], vals[1], head, runOf, aux[0], aux[1], aux[2]}) cudaFree(p); cudaFree(tmp); cudaFree(bad); };
    for (auto pp : {&keys[0], &keys[1], &vals[0], &vals[1], &head, &runOf, &aux[0], &aux[1], &aux[2]}) DD_CHECK(cudaMalloc((void **) pp, ((size_t) n + 1) * 4));
    DD_CHECK(cudaMalloc((void **) &bad, 4)); DD_CHECK(cudaMemsetAsync(bad, 0, 4, st));
    size_t tmpBytes = 0, t2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, keys[0], keys[1], vals[0], vals[1], (int) n, 0, 32, st);
    cub::DeviceScan::ExclusiveSum(nullptr, t2, head, runOf, (int) n + 1, st);
    tmpBytes = std::max(tmpBytes, t2);
    DD_CHECK(cudaMalloc(&tmp, tmpBytes));
    const unsigned g = (n + 255u) / 256u;
    // 1. stable LSD sort by (p, uv, n): fields in the order n (2), uv (1), p (0)
    const int fields[3] = {2, 1, 0};
    uint32_t *order = vals[0];
    for (int pass = 0; pass < 3; ++pass) {
        if (pass == 0) k_ddInit<<<g, 256, 0, st>>>(corners, n, fields[0], keys[0], vals[0]);
        else k_ddGatherKey<<<g, 256, 0, st>>>(corners, vals[0], n, fields[pass], keys[0]);
        DD_CHECK(cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, keys[0], keys[1], vals[0], vals[1], (int) n, 0, 32, st));
        std::swap(vals[0], vals[1]);
        order = vals[0];
    }
    // 2. run heads and run numbers
    k_ddHeads<<<g, 256, 0, st>>>(corners, order, n, nP, nT, nN, head, bad);
    DD_CHECK(cudaMemsetAsync(head + n, 0, 4, st));
    DD_CHECK(cub::DeviceScan::ExclusiveSum(tmp, tmpBytes, head, runOf, (int) n + 1, st));
    uint32_t nRuns = 0; unsigned int nBad = 0;
    DD_CHECK(cudaMemcpyAsync(&nRuns, runOf + n, 4, cudaMemcpyDeviceToHost, st));
    DD_CHECK(cudaMemcpyAsync(&nBad, bad, 4, cudaMemcpyDeviceToHost, st));
    DD_CHECK(cudaStreamSynchronize(st));
    if (nBad) { err = "obj_dedup: " + std::to_string(nBad) + " face corners name a position / texcoord / normal that does not exist"; cleanup(); return false; }
    // 3. first-use order of the runs
    uint32_t *firstCorner = aux[0], *runIds = aux[1], *newId = aux[2];
    k_ddRuns<<<g, 256, 0, st>>>(order, head, runOf, n, firstCorner, runIds);
    DD_CHECK(cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, firstCorner, keys[0], runIds, keys[1], (int) nRuns, 0, 32, st));   // keys[0] = first corners ascending, keys[1] = their runs
    const unsigned gr = (nRuns + 255u) / 256u;
    k_ddRank<<<gr, 256, 0, st>>>(keys[1], nRuns, newId);
    // 4. index buffer and vertex arrays
    k_ddIndex<<<g, 256, 0, st>>>(order, head, runOf, newId, n, F);
    k_ddGather<<<gr, 256, 0, st>>>(corners, keys[0], nRuns, pos, nT ? tex : nullptr, nN ? nrm : nullptr, V, UV, N);
    DD_CHECK(cudaStreamSynchronize(st)); DD_CHECK(cudaGetLastError());
    *nVertsOut = nRuns;
    cleanup();
    return true;
}

} // namespace ng
