// pathgraph_scatter.cu — second slice of the fork's path-graph post-process as sm_100a kernels: the
// marginal densities and the iterated radiance scatter over the k-nearest-neighbour graph, i.e. what the reference runs in
//   computeScatterAllOnGPURecord   src/pbsdf.cu:2317-2368   (host driver: marginal, then per iteration scatter + "last run")
//   allGPUPdfMarginal              src/pbsdf.cu:1738-1756   pdfmarginal[i] = sum over i's neighbours j of pdf_j(wi_i) * rrpdf_j
//   allGPUScatterRadiance          src/pbsdf.cu:1398-1431   radiance[j] += temp[i + 1] * bsdf_j(wi_i) / pdfmarginal[i]  for j in N(i)
//   updateRadiance / lastRun       src/pbsdf.cu:2185-2194, 1497-1526
// Differences in HOW (not in WHAT is computed):
//   * the reference runs one thread per vertex and copies whole 168-byte SPoint records by value for every neighbour it
//     visits (`SPoint spo = sps[index]`); here a WARP owns a vertex, its lanes take the neighbours, and what an evaluation reads
//     of a neighbour comes from two packed arrays: 64 bytes that every evaluation needs (normals, wo, albedo, roughness,
//     rrpdf, nidx, type tag) and 48 bytes only the non-diffuse types read (specular colour, eta, k);
//   * radiance is kept as float4 and scattered with ONE vector reduction per neighbour (float4 atomicAdd: REDG.E.ADD.F32x4) instead of three
//     scalar atomics;
//   * results stay on the device between the stages (the reference synchronises and copies after every kernel);
//   * the marginal sums are taken in a fixed order (lane l adds neighbours l, l + 32, ... in turn, then a butterfly over the
//     lanes), so they do not depend on scheduling; the reference adds them in list order in one thread. The scatter sums are
//     atomic in both: their order is not defined by either implementation.
#include <algorithm>
#include <cstring>
#include "engine.h"
#include "check.cuh"
#include "pathgraph.h"
#include "pathgraph_bsdf.cuh"

namespace ng {

namespace {

/// Packed vertex: A = 4 x float4 (shN | roughness, geoN | rrpdf, wo | type, diffuse | nidx), B = 3 x float4 (specular | k.x, eta | k.y, k.z).
__global__ void k_pgPack(const NoriGpuSPoint *__restrict__ sps, int n, float4 *__restrict__ A, float4 *__restrict__ B, float4 *__restrict__ wi,
                         float4 *__restrict__ eLi, float4 *__restrict__ eLd) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const NoriGpuSPoint &s = sps[i];
    A[4 * (size_t) i + 0] = make_float4(s.shN[0], s.shN[1], s.shN[2], s.roughness);
    A[4 * (size_t) i + 1] = make_float4(s.geoN[0], s.geoN[1], s.geoN[2], s.rrpdf);
    A[4 * (size_t) i + 2] = make_float4(s.wo[0], s.wo[1], s.wo[2], __int_as_float((int) (unsigned char) s.bsdf_type));
    A[4 * (size_t) i + 3] = make_float4(s.diffuse[0], s.diffuse[1], s.diffuse[2], __int_as_float(s.nidx));
    B[3 * (size_t) i + 0] = make_float4(s.specular[0], s.specular[1], s.specular[2], s.k[0]);
    B[3 * (size_t) i + 1] = make_float4(s.eta[0], s.eta[1], s.eta[2], s.k[1]);
    B[3 * (size_t) i + 2] = make_float4(s.k[2], 0.f, 0.f, 0.f);
    wi[i] = make_float4(s.wi[0], s.wi[1], s.wi[2], 0.f);
    eLi[i] = make_float4(s.eLi[0], s.eLi[1], s.eLi[2], 0.f);
    eLd[i] = make_float4(s.eLd[0], s.eLd[1], s.eLd[2], 0.f);
}

__device__ __forceinline__ void pgLoadA(const float4 *__restrict__ A, int j, PgVertex &v) {
    const float4 a0 = __ldg(A + 4 * (size_t) j), a1 = __ldg(A + 4 * (size_t) j + 1), a2 = __ldg(A + 4 * (size_t) j + 2), a3 = __ldg(A + 4 * (size_t) j + 3);
    v.shN[0] = a0.x; v.shN[1] = a0.y; v.shN[2] = a0.z; v.roughness = a0.w;
    v.geoN[0] = a1.x; v.geoN[1] = a1.y; v.geoN[2] = a1.z; v.rrpdf = a1.w;
    v.wo[0] = a2.x; v.wo[1] = a2.y; v.wo[2] = a2.z; v.type = __float_as_int(a2.w);
    v.diffuse[0] = a3.x; v.diffuse[1] = a3.y; v.diffuse[2] = a3.z; v.nidx = __float_as_int(a3.w);
    v.specular[0] = v.specular[1] = v.specular[2] = 0.f; v.eta[0] = v.eta[1] = v.eta[2] = 0.f; v.k[0] = v.k[1] = v.k[2] = 0.f;
}
__device__ __forceinline__ void pgLoadB(const float4 *__restrict__ B, int j, PgVertex &v) {
    if (v.type == 'd') return;                              // the diffuse evaluations read nothing of B
    const float4 b0 = __ldg(B + 3 * (size_t) j), b1 = __ldg(B + 3 * (size_t) j + 1), b2 = __ldg(B + 3 * (size_t) j + 2);
    v.specular[0] = b0.x; v.specular[1] = b0.y; v.specular[2] = b0.z;
    v.eta[0] = b1.x; v.eta[1] = b1.y; v.eta[2] = b1.z;
    v.k[0] = b0.w; v.k[1] = b1.w; v.k[2] = b2.x;
}

constexpr int kPgWarps = 8;             // warps (= vertices) per CTA

/// allGPUPdfMarginal (src/pbsdf.cu:1738-1756). Lane l sums its neighbours l, l + 32, ... in order; lanes are then combined with a
/// butterfly (xor 16, 8, 4, 2, 1): a fixed order, restated by the oracle.
__global__ void __launch_bounds__(kPgWarps * 32) k_pgPdfMarginal(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                                 const int *__restrict__ neighbors, int pitchInts, const int *__restrict__ knn, int n,
                                                                 float *__restrict__ pdfmarginal) {
    const int i = blockIdx.x * kPgWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const int k = knn[i];
    const float4 w4 = wiAll[i];
    const float wi[3] = {w4.x, w4.y, w4.z};
    const int *row = neighbors + (size_t) i * pitchInts;
    float sum = 0.0f;
    for (int t = lane; t < k; t += 32) {
        const int j = row[t];
        NG_CHECK(j >= 0 && j < n);
        PgVertex v; pgLoadA(A, j, v);
        if (v.nidx == i) continue;                          // j's own continuation leads to i: skipped (pbsdf.cu:1749)
        pgLoadB(B, j, v);
        sum += pgPdf(v, wi) * v.rrpdf;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
    if (lane == 0) pdfmarginal[i] = sum;
}

/// float4 atomicAdd (sm_90+): one vector reduction for the three colour channels (the fourth adds 0).
__device__ __forceinline__ void pgRedAdd(float4 *dst, float x, float y, float z) {
    atomicAdd(dst, make_float4(x, y, z, 0.0f));            // result unused: compiles to one vector RED
}

/// allGPUScatterRadiance (src/pbsdf.cu:1398-1431): vertex i hands the radiance arriving along its continuation (temp of the
/// next record) to all its neighbours, each evaluating its own BSDF for i's direction, divided by i's marginal density.
__global__ void __launch_bounds__(kPgWarps * 32) k_pgScatter(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                             const float *__restrict__ pdfmarginal, const int *__restrict__ neighbors, int pitchInts,
                                                             const int *__restrict__ knn, int n, const float4 *__restrict__ temp, float4 *__restrict__ radiance) {
    const int i = blockIdx.x * kPgWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n || i + 1 >= n) return;                       // temp[i + 1] must exist (a well-formed last record has nidx == 0 anyway)
    const float pm = pdfmarginal[i];
    const float4 a1 = __ldg(A + 4 * (size_t) i + 1), a3 = __ldg(A + 4 * (size_t) i + 3);
    if (!(pm > 0.0f && __float_as_int(a3.w) > 0 && a1.w > 1e-7f)) return;
    const float inv = (float) (1.0 / (double) pm);          // `1.0 / (pdfmarginal[Index])`: double, rounded on assignment
    const int k = knn[i];
    const float4 w4 = wiAll[i], src = temp[i + 1];
    const float wi[3] = {w4.x, w4.y, w4.z};
    const int *row = neighbors + (size_t) i * pitchInts;
    for (int t = lane; t < k; t += 32) {
        const int j = row[t];
        NG_CHECK(j >= 0 && j < n);
        if (j == i + 1) continue;
        PgVertex v; pgLoadA(A, j, v); pgLoadB(B, j, v);
        float f[3]; pgBsdf(v, wi, f);
        pgRedAdd(radiance + j, (src.x * f[0]) * inv, (src.y * f[1]) * inv, (src.z * f[2]) * inv);
    }
}

/// updateRadiance (src/pbsdf.cu:2185-2194): temp = eLi in pass 0, radiance + eLd afterwards.
__global__ void k_pgUpdate(const float4 *__restrict__ eLi, const float4 *__restrict__ eLd, const float4 *__restrict__ radiance, float4 *__restrict__ temp, int n, int pass) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (pass == 0) temp[i] = eLi[i];
    else { const float4 r = radiance[i], d = eLd[i]; temp[i] = make_float4(r.x + d.x, r.y + d.y, r.z + d.z, 0.f); }
}

/// updateWithOptDirectRadiance (src/pbsdf.cu:2196-2199): temp = filtered direct light (n x 3 floats) + radiance, in every pass.
__global__ void k_pgUpdateDirect(const float *__restrict__ direct, const float4 *__restrict__ radiance, float4 *__restrict__ temp, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 r = radiance[i];
    temp[i] = make_float4(direct[3 * (size_t) i] + r.x, direct[3 * (size_t) i + 1] + r.y, direct[3 * (size_t) i + 2] + r.z, 0.f);
}

/// lastRun (src/pbsdf.cu:1497-1526): the plain Monte-Carlo estimate of the same quantity from the vertex's own continuation:
/// temp[i + 1] * bsdf_i(wi_i) / (pdf_i(wi_i) * rrpdf_i). Writes radiance (the reference adds into a cleared array).
__global__ void k_pgLastRun(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll, const float4 *__restrict__ temp,
                            float4 *__restrict__ radiance, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
    PgVertex v; pgLoadA(A, i, v); pgLoadB(B, i, v);
    const float4 w4 = wiAll[i];
    const float wi[3] = {w4.x, w4.y, w4.z};
    const float spdf = pgPdf(v, wi);
    if (spdf > 0.0f && v.nidx > 0 && i + 1 < n) {
        const float inv = (float) (1.0 / (double) (spdf * v.rrpdf));
        float f[3]; pgBsdf(v, wi, f);
        const float4 src = temp[i + 1];
        out = make_float4((src.x * f[0]) * inv, (src.y * f[1]) * inv, (src.z * f[2]) * inv, 0.f);
    }
    radiance[i] = out;
}

/// float4 rows -> the val3f rows (include/nori/shadingPoint.h:35-37) the reference's ResultSpace holds
__global__ void k_pgUnpack(const float4 *__restrict__ src, float *__restrict__ dst, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = src[i];
    dst[3 * (size_t) i] = v.x; dst[3 * (size_t) i + 1] = v.y; dst[3 * (size_t) i + 2] = v.z;
}

/// allGPUPdfSum (src/pbsdf.cu:1600-1617): pdfsum[i][a] = sum over i's neighbours b of pdf_b(wi_a) * rrpdf_b, for every neighbour a of i
/// whose path continues (nidx > 0, rrpdf > 1e-7): the density with which ANY vertex of the neighbourhood would have produced a's
/// direction — the denominator of the tool's MIS-style gather. k^2 evaluations per vertex: a warp per vertex, the outer loop
/// over a is warp-uniform, the lanes take the b (for k <= 32 each lane keeps "its" neighbour's record in registers for the whole
/// loop); sums in the same fixed order as the marginal densities (lane-strided, then a butterfly). Entries of neighbours that
/// do not continue stay 0 (the array is cleared first, as the reference does).
__global__ void __launch_bounds__(kPgWarps * 32) k_pgPdfSum(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                            const int *__restrict__ neighbors, int pitchInts, const int *__restrict__ knn, int n,
                                                            int rowInts, float *__restrict__ pdfsum) {
    const int i = blockIdx.x * kPgWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const int k = knn[i];
    const int *row = neighbors + (size_t) i * pitchInts;
    PgVertex mine; mine.type = 0; mine.rrpdf = 0.f;
    if (k <= 32) {
        // One neighbour record per lane, kept in registers for the whole loop. Rows of at most 16 (8) entries split the warp into
        // 2 (4) groups that work on different a at the same time; a group's butterfly adds the same terms in the same order as the
        // whole warp's would (the other lanes hold zeros), so the grouping does not change a bit of the result.
        const int G = k <= 8 ? 8 : (k <= 16 ? 16 : 32), group = lane / G, l = lane % G, groups = 32 / G;
        if (l < k) { const int jb = row[l]; NG_CHECK(jb >= 0 && jb < n); pgLoadA(A, jb, mine); pgLoadB(B, jb, mine); }
        for (int a0 = 0; a0 < k; a0 += groups) {
            const int a = a0 + group;
            float sum = 0.0f;
            bool want = false;
            if (a < k) {
                const int ja = row[a];
                NG_CHECK(ja >= 0 && ja < n);
                const float4 a1 = __ldg(A + 4 * (size_t) ja + 1), a3 = __ldg(A + 4 * (size_t) ja + 3);
                want = __float_as_int(a3.w) > 0 && a1.w > 1e-7f;
                if (want && l < k) {
                    const float4 w4 = __ldg(wiAll + ja);
                    const float wi[3] = {w4.x, w4.y, w4.z};
                    sum += pgPdf(mine, wi) * mine.rrpdf;
                }
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) { const float t = __shfl_xor_sync(0xffffffffu, sum, d); if (d < G) sum += t; }
            if (want && l == 0) pdfsum[(size_t) i * rowInts + a] = sum;
        }
        return;
    }
    for (int a = 0; a < k; ++a) {
        const int ja = row[a];
        NG_CHECK(ja >= 0 && ja < n);
        const float4 a1 = __ldg(A + 4 * (size_t) ja + 1), a3 = __ldg(A + 4 * (size_t) ja + 3);
        if (!(__float_as_int(a3.w) > 0 && a1.w > 1e-7f)) continue;
        const float4 w4 = __ldg(wiAll + ja);
        const float wi[3] = {w4.x, w4.y, w4.z};
        float sum = 0.0f;
        for (int b = lane; b < k; b += 32) {
            const int jb = row[b];
            NG_CHECK(jb >= 0 && jb < n);
            PgVertex v; pgLoadA(A, jb, v); pgLoadB(B, jb, v);
            sum += pgPdf(v, wi) * v.rrpdf;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
        if (lane == 0) pdfsum[(size_t) i * rowInts + a] = sum;
    }
}

/// allGPUMISRadiance (src/pbsdf.cu:1346-1370): vertex i GATHERS from its neighbours a: bsdf_i(wi_a) * temp[a + 1] / pdfsum[i][a].
/// No atomics: lane l adds its neighbours l, l + 32, ... in order, then a butterfly per channel.
__global__ void __launch_bounds__(kPgWarps * 32) k_pgMisGather(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                               const float *__restrict__ pdfsum, int rowInts, const int *__restrict__ neighbors, int pitchInts,
                                                               const int *__restrict__ knn, int n, const float4 *__restrict__ temp, float4 *__restrict__ radiance) {
    const int i = blockIdx.x * kPgWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const int k = knn[i];
    const int *row = neighbors + (size_t) i * pitchInts;
    PgVertex self; pgLoadA(A, i, self); pgLoadB(B, i, self);
    float r0 = 0.f, r1 = 0.f, r2 = 0.f;
    for (int a = lane; a < k; a += 32) {
        const int ja = row[a];
        NG_CHECK(ja >= 0 && ja < n);
        const float4 a1 = __ldg(A + 4 * (size_t) ja + 1), a3 = __ldg(A + 4 * (size_t) ja + 3);
        const float ps = pdfsum[(size_t) i * rowInts + a];
        if (!(__float_as_int(a3.w) > 0 && ps > 0.0f && a1.w > 1e-7f) || ja + 1 >= n) continue;
        const float4 w4 = __ldg(wiAll + ja), src = temp[ja + 1];
        const float wi[3] = {w4.x, w4.y, w4.z};
        float f[3]; pgBsdf(self, wi, f);
        const float inv = (float) (1.0 / (double) ps);
        r0 += (f[0] * src.x) * inv; r1 += (f[1] * src.y) * inv; r2 += (f[2] * src.z) * inv;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        r0 += __shfl_xor_sync(0xffffffffu, r0, d); r1 += __shfl_xor_sync(0xffffffffu, r1, d); r2 += __shfl_xor_sync(0xffffffffu, r2, d);
    }
    if (lane == 0) radiance[i] = make_float4(r0, r1, r2, 0.f);
}

/// Direct-light scatter (computeDirectScatterAllOnGPU, src/pbsdf.cu:2370-2395). allGPUDirectPdfMarginal (:1881-1897):
/// m.x = the vertex's light-sample density added up once per neighbour (the reference adds the same float k times: so does lane 0),
/// m.y = sum over the neighbours j of pdf_j(wi_i) — no rrpdf, no exclusion here — in the fixed lane order of the other sums.
__global__ void __launch_bounds__(kPgWarps * 32) k_pgDirectMarginal(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                                    const NoriGpuLPoint *__restrict__ lps, const int *__restrict__ neighbors, int pitchInts,
                                                                    const int *__restrict__ knn, int n, float2 *__restrict__ marginal) {
    const int i = blockIdx.x * kPgWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const int k = knn[i];
    const float4 w4 = wiAll[i];
    const float wi[3] = {w4.x, w4.y, w4.z};
    const int *row = neighbors + (size_t) i * pitchInts;
    float sum = 0.0f;
    for (int t = lane; t < k; t += 32) {
        const int j = row[t];
        NG_CHECK(j >= 0 && j < n);
        PgVertex v; pgLoadA(A, j, v); pgLoadB(B, j, v);
        sum += pgPdf(v, wi);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
    if (lane == 0) {
        const float lp = lps[i].lightpdf;
        float x = 0.0f;
        for (int t = 0; t < k; ++t) x += lp;
        marginal[i] = make_float2(x, sum);
    }
}

/// allGPUDirectScatterRadiance (:1651-1687): vertex i hands its light sample and what its continuation found on an emitter to all its
/// neighbours, each evaluating its own BSDF for i's two directions; float4 reduction per neighbour as in the indirect scatter.
__global__ void __launch_bounds__(kPgWarps * 32) k_pgDirectScatter(const float4 *__restrict__ A, const float4 *__restrict__ B, const NoriGpuSPoint *__restrict__ sps,
                                                                   const NoriGpuLPoint *__restrict__ lps, const float2 *__restrict__ marginal,
                                                                   const int *__restrict__ neighbors, int pitchInts, const int *__restrict__ knn, int n,
                                                                   float4 *__restrict__ radiance) {
    const int i = blockIdx.x * kPgWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const float2 m = marginal[i];
    const float invL = (float) (1.0 / (double) m.x), invB = (float) (1.0 / (double) m.y);
    const NoriGpuLPoint lp = lps[i];
    const float wid[3] = {sps[i].wi_d[0], sps[i].wi_d[1], sps[i].wi_d[2]}, wi[3] = {sps[i].wi[0], sps[i].wi[1], sps[i].wi[2]};
    const int k = knn[i];
    const int *row = neighbors + (size_t) i * pitchInts;
    for (int t = lane; t < k; t += 32) {
        const int j = row[t];
        NG_CHECK(j >= 0 && j < n);
        PgVertex v; pgLoadA(A, j, v); pgLoadB(B, j, v);
        float out[3] = {0.f, 0.f, 0.f};
        if (m.x > 0.0f) {
            float f[3]; pgBsdf(v, wid, f);
            for (int c = 0; c < 3; ++c) out[c] = out[c] + (lp.L_directsample[c] * f[c]) * invL;
        }
        if (m.y > 0.0f) {
            float f[3]; pgBsdf(v, wi, f);
            for (int c = 0; c < 3; ++c) out[c] = out[c] + (lp.L_bsdfsample[c] * f[c]) * invB;
        }
        pgRedAdd(radiance + j, out[0], out[1], out[2]);
    }
}
/// addEmitterToDirectLight (:813-818) + the val3f rows of the result
__global__ void k_pgDirectFinish(const float4 *__restrict__ radiance, const NoriGpuLPoint *__restrict__ lps, int n, float *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 r = radiance[i];
    out[3 * (size_t) i] = r.x + lps[i].L_em[0]; out[3 * (size_t) i + 1] = r.y + lps[i].L_em[1]; out[3 * (size_t) i + 2] = r.z + lps[i].L_em[2];
}

/// The neighbour rows are trusted by the kernels above: every knn[i] in [0, rowInts] and every used entry in [0, n), else flag = 1.
__global__ void k_pgValidate(const int *__restrict__ neighbors, int pitchInts, int rowInts, const int *__restrict__ knn, int n, int *__restrict__ flag) {
    const int i = blockIdx.x * kPgWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const int k = knn[i];
    if (k < 0 || k > rowInts) { if (lane == 0) atomicOr(flag, 1); return; }
    for (int t = lane; t < k; t += 32) { const int j = neighbors[(size_t) i * pitchInts + t]; if (j < 0 || j >= n) atomicOr(flag, 2); }
}

} // namespace

/// flag (device int, cleared here): bit 0 = a knn[i] outside [0, rowInts], bit 1 = a neighbour index outside [0, n).
void launchPathgraphValidate(const int *d_neighbors, int pitchInts, int rowInts, const int *d_knn, int n, int *d_flag, cudaStream_t st) {
    cudaMemsetAsync(d_flag, 0, sizeof(int), st);
    k_pgValidate<<<(unsigned) ((n + kPgWarps - 1) / kPgWarps), kPgWarps * 32, 0, st>>>(d_neighbors, pitchInts, rowInts, d_knn, n, d_flag);
}

namespace {

/// Bounding box of the record positions (SPoints::BuildKNN reads it off the loaded set, shadingPoint.h:786-793): block-wise
/// min / max, then one ordered-integer atomic per block and coordinate (floats compare like their sign-fixed bit patterns).
__device__ __forceinline__ unsigned int pgOrdered(float f) { const unsigned int u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__global__ void __launch_bounds__(256) k_pgBounds(const NoriGpuSPoint *__restrict__ sps, int n, unsigned int *__restrict__ bounds /* lo.xyz (min), hi.xyz (max), ordered */) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int c = 0; c < 3; ++c) { const float p = sps[i].pos[c]; lo[c] = fminf(lo[c], p); hi[c] = fmaxf(hi[c], p); }
    for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], d)); hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], d)); }
        if ((threadIdx.x & 31) == 0) { atomicMin(bounds + c, pgOrdered(lo[c])); atomicMax(bounds + 3 + c, pgOrdered(hi[c])); }
    }
}

/// The images the tool composes from each path's FIRST vertex (src/cluster.cpp:23-107): full = eLd + mc, indirect = mc, pt = eLi - eLd,
/// blur, init = eLd, averaged over the pixel's samples in sample order (one thread per pixel: no atomics, a fixed order).
/// offsets[g] / counts[g]: first record and number of records of camera sample g = pixel * spp + k (the vertex pass's own scan).
__global__ void k_pgCompose(const NoriGpuSPoint *__restrict__ sps, const unsigned long long *__restrict__ offsets, const unsigned int *__restrict__ counts,
                            const float *__restrict__ blur, const float *__restrict__ mc, int nPixels, unsigned int spp, float *__restrict__ images) {
    const int pixel = blockIdx.x * blockDim.x + threadIdx.x;
    if (pixel >= nPixels) return;
    float acc[5][3] = {};
    const float inv = 1.0f / (float) spp;
    for (unsigned int k = 0; k < spp; ++k) {
        const unsigned long long g = (unsigned long long) pixel * spp + k;
        if (!counts[g]) continue;
        const unsigned long long i = offsets[g];
        for (int c = 0; c < 3; ++c) {
            const float eLd = sps[i].eLd[c], eLi = sps[i].eLi[c], m = mc[3 * i + c], b = blur[3 * i + c];
            acc[0][c] += (eLd + m) * inv; acc[1][c] += m * inv; acc[2][c] += (eLi - eLd) * inv; acc[3][c] += b * inv; acc[4][c] += eLd * inv;
        }
    }
    for (int im = 0; im < 5; ++im) for (int c = 0; c < 3; ++c) images[((size_t) im * nPixels + pixel) * 3 + c] = acc[im][c];
}

} // namespace

void launchPathgraphBounds(const NoriGpuSPoint *d_sps, int n, unsigned int *d_bounds6, cudaStream_t st) {
    const unsigned int init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    cudaMemcpyAsync(d_bounds6, init, sizeof(init), cudaMemcpyHostToDevice, st);
    k_pgBounds<<<(unsigned) std::min((n + 255) / 256, 148 * 8), 256, 0, st>>>(d_sps, n, d_bounds6);
}
float pathgraphOrderedToFloat(unsigned int u) { const unsigned int b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u; float f; memcpy(&f, &b, 4); return f; }
void launchPathgraphCompose(const NoriGpuSPoint *d_sps, const unsigned long long *d_offsets, const unsigned int *d_counts, const float *d_blur, const float *d_mc,
                            int nPixels, unsigned int spp, float *d_images, cudaStream_t st) {
    k_pgCompose<<<(unsigned) ((nPixels + 127) / 128), 128, 0, st>>>(d_sps, d_offsets, d_counts, d_blur, d_mc, nPixels, spp, d_images);
}

size_t pgScatterScratchBytes(int n) {
    // A (64) + B (48) + wi, eLi, eLd, temp, radiance (16 each) per vertex
    return (size_t) n * (64 + 48 + 5 * 16);
}

/// computeScatterAllOnGPURecord (src/pbsdf.cu:2317-2368). d_blur / d_mc: iterations x n x 3 floats (device); d_pdfmarginal: n floats.
/// Sequence kept as the reference has it, including that its update step after the first scatter passes the LOOP index (0) again,
/// so pass 1 scatters the same temp = eLi as pass 0 (pbsdf.cu:2344-2356).
/// d_direct != nullptr: computeScatterAllOnGPURecordWithDirectOpt (:2488-2540) — the same passes with the filtered direct light (n x 3 floats,
/// launchPathgraphDirect's result) in the place of eLd: temp = direct + radiance after EVERY scatter (updateWithOptDirectRadiance).
bool launchPathgraphScatter(const NoriGpuSPoint *d_sps, int n, const int *d_neighbors, int pitchInts, const int *d_knn, int iterations,
                            void *d_scratch, float *d_pdfmarginal, float *d_blur, float *d_mc, cudaStream_t st, std::string &err, const float *d_direct) {
    if (n <= 0 || iterations < 0) { err = "pathgraph_scatter: bad sizes"; return false; }
    float4 *A = static_cast<float4 *>(d_scratch), *B = A + 4 * (size_t) n, *wi = B + 3 * (size_t) n, *eLi = wi + n, *eLd = eLi + n, *temp = eLd + n, *radiance = temp + n;
    const unsigned gt = (unsigned) ((n + 255) / 256), gw = (unsigned) ((n + kPgWarps - 1) / kPgWarps);
    k_pgPack<<<gt, 256, 0, st>>>(d_sps, n, A, B, wi, eLi, eLd);
    k_pgPdfMarginal<<<gw, kPgWarps * 32, 0, st>>>(A, B, wi, d_neighbors, pitchInts, d_knn, n, d_pdfmarginal);
    cudaMemsetAsync(radiance, 0, (size_t) n * sizeof(float4), st);
    k_pgUpdate<<<gt, 256, 0, st>>>(eLi, eLd, radiance, temp, n, 0);
    for (int it = 0; it < iterations; ++it) {
        cudaMemsetAsync(radiance, 0, (size_t) n * sizeof(float4), st);
        k_pgScatter<<<gw, kPgWarps * 32, 0, st>>>(A, B, wi, d_pdfmarginal, d_neighbors, pitchInts, d_knn, n, temp, radiance);
        k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_blur + (size_t) it * 3 * n, n);
        if (d_direct) k_pgUpdateDirect<<<gt, 256, 0, st>>>(d_direct, radiance, temp, n);
        else k_pgUpdate<<<gt, 256, 0, st>>>(eLi, eLd, radiance, temp, n, it);
        k_pgLastRun<<<gt, 256, 0, st>>>(A, B, wi, temp, radiance, n);
        k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_mc + (size_t) it * 3 * n, n);
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("pathgraph_scatter: ") + cudaGetErrorString(e); return false; }
    return true;
}

size_t pgMisScratchBytes(int n, int rowInts) { return pgScatterScratchBytes(n) + (size_t) n * rowInts * sizeof(float); }

/// computeMISAllOnGPURecord (src/pbsdf.cu:2922-2965), what the `pg` tool runs in its "knn" mode (src/cluster.cpp:223,
/// SPoints::computeMISRadianceAOGWithProcessRecording): pdfsum once, then per pass gather -> blur[it], temp = radiance + eLd
/// (the update gets it + 1 here, so only the initial temp is eLi), lastRun -> mc[it].
bool launchPathgraphMis(const NoriGpuSPoint *d_sps, int n, const int *d_neighbors, int pitchInts, const int *d_knn, int iterations,
                        void *d_scratch, float *d_blur, float *d_mc, cudaStream_t st, std::string &err) {
    if (n <= 0 || iterations < 0 || pitchInts < 1) { err = "pathgraph_mis: bad sizes"; return false; }
    float4 *A = static_cast<float4 *>(d_scratch), *B = A + 4 * (size_t) n, *wi = B + 3 * (size_t) n, *eLi = wi + n, *eLd = eLi + n, *temp = eLd + n, *radiance = temp + n;
    float *pdfsum = reinterpret_cast<float *>(radiance + n);
    const unsigned gt = (unsigned) ((n + 255) / 256), gw = (unsigned) ((n + kPgWarps - 1) / kPgWarps);
    k_pgPack<<<gt, 256, 0, st>>>(d_sps, n, A, B, wi, eLi, eLd);
    cudaMemsetAsync(pdfsum, 0, (size_t) n * pitchInts * sizeof(float), st);
    k_pgPdfSum<<<gw, kPgWarps * 32, 0, st>>>(A, B, wi, d_neighbors, pitchInts, d_knn, n, pitchInts, pdfsum);
    cudaMemsetAsync(radiance, 0, (size_t) n * sizeof(float4), st);
    k_pgUpdate<<<gt, 256, 0, st>>>(eLi, eLd, radiance, temp, n, 0);
    for (int it = 0; it < iterations; ++it) {
        k_pgMisGather<<<gw, kPgWarps * 32, 0, st>>>(A, B, wi, pdfsum, pitchInts, d_neighbors, pitchInts, d_knn, n, temp, radiance);
        k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_blur + (size_t) it * 3 * n, n);
        k_pgUpdate<<<gt, 256, 0, st>>>(eLi, eLd, radiance, temp, n, it + 1);
        k_pgLastRun<<<gt, 256, 0, st>>>(A, B, wi, temp, radiance, n);
        k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_mc + (size_t) it * 3 * n, n);
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("pathgraph_mis: ") + cudaGetErrorString(e); return false; }
    return true;
}

size_t pgDirectScratchBytes(int n) { return (size_t) n * (64 + 48 + 16 + 16 + 16 + 16 + 8); }

/// computeDirectScatterAllOnGPU (src/pbsdf.cu:2370-2395): d_direct = n x 3 floats (device), d_marginal (optional) = n float2.
bool launchPathgraphDirect(const NoriGpuSPoint *d_sps, const NoriGpuLPoint *d_lps, int n, const int *d_neighbors, int pitchInts, const int *d_knn,
                           void *d_scratch, float *d_direct, cudaStream_t st, std::string &err) {
    if (n <= 0 || pitchInts < 1) { err = "pathgraph_direct: bad sizes"; return false; }
    float4 *A = static_cast<float4 *>(d_scratch), *B = A + 4 * (size_t) n, *wi = B + 3 * (size_t) n, *eLi = wi + n, *eLd = eLi + n, *radiance = eLd + n;
    float2 *marginal = reinterpret_cast<float2 *>(radiance + n);
    const unsigned gt = (unsigned) ((n + 255) / 256), gw = (unsigned) ((n + kPgWarps - 1) / kPgWarps);
    k_pgPack<<<gt, 256, 0, st>>>(d_sps, n, A, B, wi, eLi, eLd);
    k_pgDirectMarginal<<<gw, kPgWarps * 32, 0, st>>>(A, B, wi, d_lps, d_neighbors, pitchInts, d_knn, n, marginal);
    cudaMemsetAsync(radiance, 0, (size_t) n * sizeof(float4), st);
    k_pgDirectScatter<<<gw, kPgWarps * 32, 0, st>>>(A, B, d_sps, d_lps, marginal, d_neighbors, pitchInts, d_knn, n, radiance);
    k_pgDirectFinish<<<gt, 256, 0, st>>>(radiance, d_lps, n, d_direct);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("pathgraph_direct: ") + cudaGetErrorString(e); return false; }
    return true;
}

unsigned pathgraphScatterChecks(unsigned *line) { return fetchChecks(line); }

} // namespace ng
