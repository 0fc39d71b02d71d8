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
        // What the loop needs of neighbour a — does its path continue, and in which direction — is fetched ONCE, by lane a, all k chains of
        // dependent loads (row -> record -> direction) in flight together, and handed to the group that works on a by shuffle.
        float4 wMine = make_float4(0.f, 0.f, 0.f, 0.f);           // wi of neighbour `lane` | 1 if it continues
        if (lane < k) {
            const int ja = row[lane];
            NG_CHECK(ja >= 0 && ja < n);
            const float4 a1 = __ldg(A + 4 * (size_t) ja + 1), a3 = __ldg(A + 4 * (size_t) ja + 3);
            if (__float_as_int(a3.w) > 0 && a1.w > 1e-7f) { wMine = __ldg(wiAll + ja); wMine.w = 1.0f; }
        }
        for (int a0 = 0; a0 < k; a0 += groups) {
            const int a = a0 + group, src = min(a, 31);
            float sum = 0.0f;
            const float wx = __shfl_sync(0xffffffffu, wMine.x, src), wy = __shfl_sync(0xffffffffu, wMine.y, src), wz = __shfl_sync(0xffffffffu, wMine.z, src);
            const float goes = __shfl_sync(0xffffffffu, wMine.w, src);          // (every lane takes part in the shuffles: no short circuit)
            const bool want = a < k && goes != 0.0f;
            if (want && l < k) {
                const float wi[3] = {wx, wy, wz};
                sum += pgPdf(mine, wi) * mine.rrpdf;
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
                            const float *__restrict__ blur, const float *__restrict__ mc, const float *__restrict__ direct, int nPixels, unsigned int spp,
                            float *__restrict__ images) {
    const int pixel = blockIdx.x * blockDim.x + threadIdx.x;
    if (pixel >= nPixels) return;
    float acc[6][3] = {};
    const float inv = 1.0f / (float) spp;
    for (unsigned int k = 0; k < spp; ++k) {
        const unsigned long long g = (unsigned long long) pixel * spp + k;
        if (!counts[g]) continue;
        const unsigned long long i = offsets[g];
        for (int c = 0; c < 3; ++c) {
            const float eLd = sps[i].eLd[c], eLi = sps[i].eLi[c], m = mc[3 * i + c], b = blur[3 * i + c];
            acc[0][c] += (eLd + m) * inv; acc[1][c] += m * inv; acc[2][c] += (eLi - eLd) * inv; acc[3][c] += b * inv; acc[4][c] += eLd * inv;
            if (direct) acc[5][c] += direct[3 * i + c] * inv;                 // writeDirectLight (src/cluster.cpp:102-126): the filtered direct light
        }
    }
    for (int im = 0; im < (direct ? 6 : 5); ++im) for (int c = 0; c < 3; ++c) images[((size_t) im * nPixels + pixel) * 3 + c] = acc[im][c];
}

} // namespace

void launchPathgraphBounds(const NoriGpuSPoint *d_sps, int n, unsigned int *d_bounds6, cudaStream_t st) {
    const unsigned int init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    cudaMemcpyAsync(d_bounds6, init, sizeof(init), cudaMemcpyHostToDevice, st);
    k_pgBounds<<<(unsigned) std::min((n + 255) / 256, 148 * 8), 256, 0, st>>>(d_sps, n, d_bounds6);
}
float pathgraphOrderedToFloat(unsigned int u) { const unsigned int b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u; float f; memcpy(&f, &b, 4); return f; }
void launchPathgraphCompose(const NoriGpuSPoint *d_sps, const unsigned long long *d_offsets, const unsigned int *d_counts, const float *d_blur, const float *d_mc,
                            int nPixels, unsigned int spp, float *d_images, cudaStream_t st, const float *d_direct) {
    k_pgCompose<<<(unsigned) ((nPixels + 127) / 128), 128, 0, st>>>(d_sps, d_offsets, d_counts, d_blur, d_mc, d_direct, nPixels, spp, d_images);
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

// ---- clustered variant (the `pg` tool's default "opt" mode after BuildClusters: SPoints::ClusterScatterWithDirectOptNR,
// include/nori/shadingPoint.h:600-624) --------------------------------------------------------------------------------------
// The vertices are partitioned into clusters (clusters[offsets[c] ...] = the members of cluster c; SPoint::groupIdx = the vertex's
// cluster or -1) and every vertex exchanges light with ALL members of its cluster: a dense block per cluster. The reference runs a
// thread per vertex, stores the block's elements (MatrixElementsNumber / precomputedMatrixElemtns, src/pbsdf.cu:3344-3357, 3535-3553),
// multiplies with float atomics (MX, :1265-1307) and adds the clusters' in / out sums with more atomics (clampCluster, :2127-2144).
// Here a WARP owns a cluster: its lanes own the receiving members, the sending members are walked in list order, so every sum of the
// block — what a member receives, the cluster's in / out totals — is taken inside the warp in a fixed order: no atomics, no stored
// elements (the K evaluations per member are redone each pass: cheaper on this machine than streaming 12 bytes per element), and
// results that do not depend on scheduling.
namespace {

__device__ __forceinline__ int pgClusterSize(const int *__restrict__ offsets, int c, int nClusters, int nListed) {   // getCellSize, :828-834
    return (c == nClusters - 1 ? nListed : offsets[c + 1]) - offsets[c];
}

/// What the clustered kernels read of the light records, 16-byte aligned: wi_d, L_directsample | lightpdf, L_bsdfsample | bsdfpdf, L_em | cluster.
__global__ void k_pgPackLights(const NoriGpuSPoint *__restrict__ sps, const NoriGpuLPoint *__restrict__ lps, const int *__restrict__ clusterOf, int n,
                               float4 *__restrict__ wid, float4 *__restrict__ Ld, float4 *__restrict__ Lb, float4 *__restrict__ Lem) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    NoriGpuLPoint l = {};                                       // no light records (the drivers without direct light): zeros
    if (lps) l = lps[i];
    wid[i] = make_float4(sps[i].wi_d[0], sps[i].wi_d[1], sps[i].wi_d[2], 0.f);
    Ld[i] = make_float4(l.L_directsample[0], l.L_directsample[1], l.L_directsample[2], l.lightpdf);
    Lb[i] = make_float4(l.L_bsdfsample[0], l.L_bsdfsample[1], l.L_bsdfsample[2], l.bsdfpdf);
    Lem[i] = make_float4(l.L_em[0], l.L_em[1], l.L_em[2], __int_as_float(clusterOf ? clusterOf[i] : sps[i].groupIdx));
}

/// The lists are trusted by the kernels below. flag bits: 1 = offsets not ascending inside [0, nListed], 2 = a member outside [0, n),
/// 4 = a member whose groupIdx is not the cluster that lists it, 8 = a vertex listed twice, 16 = a groupIdx outside [-1, nClusters),
/// 32 = a cluster of more than kPgMaxClusterSize members.
/// listed: cleared array of n ints (members are counted in it); the host compares the number of members with the vertices that name a cluster.
__global__ void k_pgClusterValidate(const NoriGpuSPoint *__restrict__ sps, int n, const int *__restrict__ clusters, int nListed, const int *__restrict__ offsets,
                                    int nClusters, int *__restrict__ listed, int *__restrict__ flag, int *__restrict__ nNamed) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid < nClusters) {
        const int lo = offsets[tid], hi = tid == nClusters - 1 ? nListed : offsets[tid + 1];
        if (lo < 0 || hi < lo || hi > nListed || (tid == 0 && lo != 0)) { atomicOr(flag, 1); return; }
        if (hi - lo > kPgMaxClusterSize) { atomicOr(flag, 32); return; }          // a warp works through a cluster's size^2 pairs
        for (int t = lo; t < hi; ++t) {
            const int j = clusters[t];
            if (j < 0 || j >= n) { atomicOr(flag, 2); continue; }
            if (sps[j].groupIdx != tid) atomicOr(flag, 4);
            if (atomicAdd(listed + j, 1) != 0) atomicOr(flag, 8);
        }
    }
    if (tid < n) {
        const int g = sps[tid].groupIdx;
        if (g < -1 || g >= nClusters) atomicOr(flag, 16);
        else if (g >= 0) atomicAdd(nNamed, 1);
    }
}

/// Sub-warp groups. A cluster of at most 8 (16) members is worked on by a quarter (half) of a warp, so that a warp takes up to four
/// clusters at once; the sums are those of a full warp whose other lanes hold +0: `+ 0.0f` stands for the butterfly steps that would
/// have added those lanes (it also turns -0 into +0 exactly as they would), then xor W/2 .. 1 inside the group.
template <int W> __device__ __forceinline__ float pgGroupSum(unsigned int mask, float v) {
    if (W < 32) v += 0.0f;
#pragma unroll
    for (int d = W / 2; d > 0; d >>= 1) v += __shfl_xor_sync(mask, v, d);
    return v;
}
template <int W> __device__ __forceinline__ void pgGroupSum3(unsigned int mask, float *v) {
    v[0] = pgGroupSum<W>(mask, v[0]); v[1] = pgGroupSum<W>(mask, v[1]); v[2] = pgGroupSum<W>(mask, v[2]);
}
/// A warp takes four consecutive clusters: all of at most 8 members -> one per quarter warp; else pair by pair: both of at most 16
/// members -> one per half warp, else one after the other on the whole warp. F::run<W>(c, lane in group, mask of the group).
/// The loops are NOT unrolled: one copy of the body per width. The bodies are thousands of instructions long and these kernels wait
/// on instruction fetch as it is: with the seven copies full unrolling makes they ran 45 % longer (r02_notes.md); a single copy with
/// W as a run-time value was slower again (4.6 vs 4.2 ms).
template <class F> __device__ __forceinline__ void pgClusterDispatch(const F &f, const int *__restrict__ offsets, int nClusters, int nListed) {
    const int c4 = (blockIdx.x * kPgWarps + (threadIdx.x >> 5)) * 4, lane = threadIdx.x & 31;
    if (c4 >= nClusters) return;
    int m[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) m[q] = c4 + q < nClusters ? pgClusterSize(offsets, c4 + q, nClusters, nListed) : 0;
    if (max(max(m[0], m[1]), max(m[2], m[3])) <= 8) { f.template run<8>(c4 + (lane >> 3), lane & 7, 0xffu << (lane & 24)); return; }
    const int pairMax[2] = {max(m[0], m[1]), max(m[2], m[3])};
#pragma unroll 1
    for (int pr = 0; pr < 2; ++pr) {
        if ((pr ? pairMax[1] : pairMax[0]) <= 16) f.template run<16>(c4 + 2 * pr + (lane >> 4), lane & 15, 0xffffu << (lane & 16));
        else {
#pragma unroll 1
            for (int h = 0; h < 2; ++h) f.template run<32>(c4 + 2 * pr + h, lane, 0xffffffffu);
        }
    }
}

/// The senders of a cluster are staged W at a time: lane q of the group loads sender s0 + q — W independent loads in flight instead of
/// a chain of dependent ones per receiver — and hands it to the group's lanes by shuffle when its turn in the list order comes.
template <int W> __device__ __forceinline__ float4 pgBcast4(unsigned int mask, const float4 &v, int q) {
    return make_float4(__shfl_sync(mask, v.x, q, W), __shfl_sync(mask, v.y, q, W), __shfl_sync(mask, v.z, q, W), __shfl_sync(mask, v.w, q, W));
}
template <int W> __device__ __forceinline__ float2 pgBcast2(unsigned int mask, const float2 &v, int q) {
    return make_float2(__shfl_sync(mask, v.x, q, W), __shfl_sync(mask, v.y, q, W));
}

/// ClusterDirectPdfMarginal (src/pbsdf.cu:1899-1937), clustered branch: for the sending member i, over all members j in list order,
/// x += lightpdf_i (if wi_d_i lies above j's shading normal) + pdf_j(wi_d_i), y += bsdfpdf_i (if wi_i does) + pdf_j(wi_i).
/// Lane l takes j = l, l + 32, ...; butterfly over the lanes.
struct PgClusterDirectMarginal {
    const float4 *A;
    const float4 *B;
    const float4 *wiAll;
    const float4 *widAll;
    const float4 *Ld;
    const float4 *Lb;
    const int *clusters;
    int nListed;
    const int *offsets;
    int nClusters;
    float2 *dm;
    template <int W> __device__ __forceinline__ void run(int c, int lane, unsigned int mask) const {
        if (c >= nClusters) return;
        const int m = pgClusterSize(offsets, c, nClusters, nListed);
        const int *mem = clusters + offsets[c];
        PgVertex own;
        if (m <= W && lane < m) { const int j = mem[lane]; pgLoadA(A, j, own); pgLoadB(B, j, own); }     // the usual case: one member per lane, loaded once
        for (int s0 = 0; s0 < m; s0 += W) {
            int iMine = 0; float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f), d4 = w4;            // sender s0 + lane: wi | lightpdf, wi_d | bsdfpdf
            if (s0 + lane < m) { iMine = mem[s0 + lane]; w4 = wiAll[iMine]; d4 = widAll[iMine]; w4.w = Ld[iMine].w; d4.w = Lb[iMine].w; }
            const int cnt = min(W, m - s0);
            for (int q = 0; q < cnt; ++q) {
                const int i = __shfl_sync(mask, iMine, q, W);
                const float4 sw = pgBcast4<W>(mask, w4, q), sd = pgBcast4<W>(mask, d4, q);
                const float wi[3] = {sw.x, sw.y, sw.z}, wid[3] = {sd.x, sd.y, sd.z};
                const float lp = sw.w, bp = sd.w;
                float x = 0.0f, y = 0.0f;
                auto term = [&](const PgVertex &v) {
                    const float pb = pgPdf(v, wi), pl = pgPdf(v, wid);
                    const float cosh1 = pgDot(wi, v.shN), cosh2 = pgDot(wid, v.shN);
                    if (cosh2 > 0.0f) x += lp;
                    if (cosh1 > 0.0f) y += bp;
                    x += pl; y += pb;
                };
                if (m <= W) { if (lane < m) term(own); }
                else for (int t = lane; t < m; t += W) { const int j = mem[t]; PgVertex v; pgLoadA(A, j, v); pgLoadB(B, j, v); term(v); }
                x = pgGroupSum<W>(mask, x); y = pgGroupSum<W>(mask, y);
                // what ClusterDirectRadiance divides by, once per sender instead of once per pair: 1 / x, 1 / y in double, rounded (-1: term skipped)
                if (lane == 0) dm[i] = make_float2(x > 0.0f ? (float) (1.0 / (double) x) : -1.0f, y > 0.0f ? (float) (1.0 / (double) y) : -1.0f);
            }
        }
    }
};
__global__ void __launch_bounds__(kPgWarps * 32) k_pgClusterDirectMarginal(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                                           const float4 *__restrict__ widAll, const float4 *__restrict__ Ld, const float4 *__restrict__ Lb,
                                                                           const int *__restrict__ clusters, int nListed, const int *__restrict__ offsets, int nClusters,
                                                                           float2 *__restrict__ dm) {
    pgClusterDispatch(PgClusterDirectMarginal{A, B, wiAll, widAll, Ld, Lb, clusters, nListed, offsets, nClusters, dm}, offsets, nClusters, nListed);
}

/// computeRatio (src/pbsdf.cu:2146-2172): a cluster that hands out more than it took in is scaled back, channel by channel.
__device__ __forceinline__ void pgRatio(const float *in, const float *out, float *ratio) {
    for (int c = 0; c < 3; ++c) ratio[c] = in[c] < out[c] ? in[c] / out[c] : 1.0f;
}

/// ClusterDirect for one cluster (src/pbsdf.cu:2400-2446): ClusterDirectRadiance (:1939-2010) — member j receives, from every member i in
/// list order, L_direct_i bsdf_j(wi_d_i) / x_i + L_bsdf_i bsdf_j(wi_i) / y_i —, clampDirectCluster (:2100-2125: in = sum of the members'
/// L_directsample and L_bsdfsample, out = sum of what they received), computeRatio, updateComputeCluster, addEmitterToDirectLight.
/// raw: float4 scratch per vertex (a lane re-reads only what it wrote itself).
struct PgClusterDirect {
    const float4 *A;
    const float4 *B;
    const float4 *wiAll;
    const float4 *widAll;
    const float4 *Ld;
    const float4 *Lb;
    const float4 *Lem;
    const float2 *dm;
    const int *clusters;
    int nListed;
    const int *offsets;
    int nClusters;
    float4 *raw;
    float4 *direct;
    template <int W> __device__ __forceinline__ void run(int c, int lane, unsigned int mask) const {
        if (c >= nClusters) return;
        const int m = pgClusterSize(offsets, c, nClusters, nListed);
        const int *mem = clusters + offsets[c];
        float in[3] = {0.f, 0.f, 0.f}, out[3] = {0.f, 0.f, 0.f};
        for (int t0 = 0; t0 < m; t0 += W) {
            const int t = t0 + lane;
            const bool valid = t < m;
            const int j = mem[valid ? t : 0];
            PgVertex v; pgLoadA(A, j, v); pgLoadB(B, j, v);
            float acc[3] = {0.f, 0.f, 0.f};
            for (int s0 = 0; s0 < m; s0 += W) {
                float2 mg = make_float2(-1.f, -1.f); float4 d4 = make_float4(0.f, 0.f, 0.f, 0.f), w4 = d4, La = d4, Lc = d4;      // sender s0 + lane
                if (s0 + lane < m) { const int i = mem[s0 + lane]; mg = dm[i]; if (mg.x >= 0.0f) { d4 = widAll[i]; La = Ld[i]; } if (mg.y >= 0.0f) { w4 = wiAll[i]; Lc = Lb[i]; } }
                const int cnt = min(W, m - s0);
                for (int q = 0; q < cnt; ++q) {
                    const float2 g = pgBcast2<W>(mask, mg, q);
                    float fin[3] = {0.f, 0.f, 0.f};
                    if (g.x >= 0.0f) {
                        const float4 sd = pgBcast4<W>(mask, d4, q), L = pgBcast4<W>(mask, La, q);
                        const float wid[3] = {sd.x, sd.y, sd.z};
                        float f[3]; pgBsdf(v, wid, f);
                        fin[0] = fin[0] + (L.x * f[0]) * g.x; fin[1] = fin[1] + (L.y * f[1]) * g.x; fin[2] = fin[2] + (L.z * f[2]) * g.x;
                    }
                    if (g.y >= 0.0f) {
                        const float4 sw = pgBcast4<W>(mask, w4, q), L = pgBcast4<W>(mask, Lc, q);
                        const float wi[3] = {sw.x, sw.y, sw.z};
                        float f[3]; pgBsdf(v, wi, f);
                        fin[0] = fin[0] + (L.x * f[0]) * g.y; fin[1] = fin[1] + (L.y * f[1]) * g.y; fin[2] = fin[2] + (L.z * f[2]) * g.y;
                    }
                    acc[0] += fin[0]; acc[1] += fin[1]; acc[2] += fin[2];
                }
            }
            if (valid) {
                raw[j] = make_float4(acc[0], acc[1], acc[2], 0.f);
                out[0] += acc[0]; out[1] += acc[1]; out[2] += acc[2];
                const float4 a = Ld[j], b = Lb[j];
                in[0] += a.x; in[1] += a.y; in[2] += a.z;
                in[0] += b.x; in[1] += b.y; in[2] += b.z;
            }
        }
        pgGroupSum3<W>(mask, in); pgGroupSum3<W>(mask, out);
        float ratio[3]; pgRatio(in, out, ratio);
        for (int t = lane; t < m; t += W) {
            const int j = mem[t];
            const float4 r = raw[j], e = Lem[j];
            direct[j] = make_float4(r.x * ratio[0] + e.x, r.y * ratio[1] + e.y, r.z * ratio[2] + e.z, 0.f);
        }
    }
};
__global__ void __launch_bounds__(kPgWarps * 32) k_pgClusterDirect(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                                   const float4 *__restrict__ widAll, const float4 *__restrict__ Ld, const float4 *__restrict__ Lb,
                                                                   const float4 *__restrict__ Lem, const float2 *__restrict__ dm, const int *__restrict__ clusters,
                                                                   int nListed, const int *__restrict__ offsets, int nClusters, float4 *__restrict__ raw,
                                                                   float4 *__restrict__ direct) {
    pgClusterDispatch(PgClusterDirect{A, B, wiAll, widAll, Ld, Lb, Lem, dm, clusters, nListed, offsets, nClusters, raw, direct}, offsets, nClusters, nListed);
}
/// Vertices outside every cluster keep their own estimate (ClusterDirectRadiance's else branch: eLd) + their emission.
__global__ void k_pgLooseDirect(const float4 *__restrict__ eLd, const float4 *__restrict__ Lem, int n, float4 *__restrict__ direct) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || __float_as_int(Lem[i].w) != -1) return;
    const float4 d = eLd[i], e = Lem[i];
    direct[i] = make_float4(d.x + e.x, d.y + e.y, d.z + e.z, 0.f);
}

/// What a pass needs of a sender, once per vertex instead of once per pair: x = the factor computeNoneZeroElements multiplies the
/// evaluations by, 1 / pm_i in double, rounded (-1 where it leaves its cleared elements untouched: pm_i <= 0 or rrpdf_i <= 1e-7),
/// y = 1 if the vertex has a successor whose radiance it hands on (MX's `nidx > 0`), else 0.
__device__ __forceinline__ float2 pgSenderTerms(const float4 *__restrict__ A, int i, int n, float pmI) {
    const float4 a1 = __ldg(A + 4 * (size_t) i + 1), a3 = __ldg(A + 4 * (size_t) i + 3);
    const bool live = pmI > 0.0f && a1.w > 1e-7f;
    return make_float2(live ? (float) (1.0 / (double) pmI) : -1.0f, (__float_as_int(a3.w) > 0 && i + 1 < n) ? 1.0f : 0.0f);
}

/// allGPUClusterPdfMarginal (src/pbsdf.cu:1239-1263): pm[i] = sum over the members j of i's cluster, except i's predecessor on its
/// path, of pdf_j(wi_i) rrpdf_j; for a vertex outside every cluster its own pdf_i(wi_i) rrpdf_i.
struct PgClusterMarginal {
    const float4 *A;
    const float4 *B;
    const float4 *wiAll;
    const int *clusters;
    int nListed;
    const int *offsets;
    int nClusters;
    int n;
    float2 *pm;
    template <int W> __device__ __forceinline__ void run(int c, int lane, unsigned int mask) const {
        if (c >= nClusters) return;
        const int m = pgClusterSize(offsets, c, nClusters, nListed);
        const int *mem = clusters + offsets[c];
        PgVertex own;
        if (m <= W && lane < m) { const int j = mem[lane]; pgLoadA(A, j, own); pgLoadB(B, j, own); }
        for (int s0 = 0; s0 < m; s0 += W) {
            float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f);                                     // sender s0 + lane: wi | index
            if (s0 + lane < m) { const int i = mem[s0 + lane]; w4 = wiAll[i]; w4.w = __int_as_float(i); }
            const int cnt = min(W, m - s0);
            for (int q = 0; q < cnt; ++q) {
                const float4 sw = pgBcast4<W>(mask, w4, q);
                const int i = __float_as_int(sw.w);
                const float wi[3] = {sw.x, sw.y, sw.z};
                float sum = 0.0f;
                if (m <= W) { if (lane < m && own.nidx != i) sum += pgPdf(own, wi) * own.rrpdf; }
                else for (int t = lane; t < m; t += W) {
                    const int j = mem[t];
                    PgVertex v; pgLoadA(A, j, v);
                    if (v.nidx == i) continue;
                    pgLoadB(B, j, v);
                    sum += pgPdf(v, wi) * v.rrpdf;
                }
                sum = pgGroupSum<W>(mask, sum);
                if (lane == 0) pm[i] = pgSenderTerms(A, i, n, sum);
            }
        }
    }
};
__global__ void __launch_bounds__(kPgWarps * 32) k_pgClusterMarginal(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                                     const int *__restrict__ clusters, int nListed, const int *__restrict__ offsets, int nClusters,
                                                                     int n, float2 *__restrict__ pm) {
    pgClusterDispatch(PgClusterMarginal{A, B, wiAll, clusters, nListed, offsets, nClusters, n, pm}, offsets, nClusters, nListed);
}
__global__ void k_pgLooseMarginal(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll, const float4 *__restrict__ Lem,
                                  int n, float2 *__restrict__ pm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || __float_as_int(Lem[i].w) != -1) return;
    PgVertex v; pgLoadA(A, i, v); pgLoadB(B, i, v);
    const float4 w4 = wiAll[i];
    const float wi[3] = {w4.x, w4.y, w4.z};
    pm[i] = pgSenderTerms(A, i, n, pgPdf(v, wi) * v.rrpdf);
}

/// One pass over one cluster: MX with the elements of computeNoneZeroElements (src/pbsdf.cu:1265-1307, 1059-1085) — member j receives
/// (bsdf_j(wi_i) / pm_i) * temp[i + 1] from every member i that has a successor, in list order; the element is 0 where the reference
/// left its cleared array untouched (pm_i <= 0 or rrpdf_i <= 1e-7) —, clampCluster (:2127-2144: in += temp[i + 1] of those members,
/// out += what the members received; both totals RUN ON over the passes, as the reference never clears them inside its loop, :3492-3516),
/// computeRatio, updateComputeCluster, and, when another pass follows, updateWithOptDirectRadiance: tempOut = direct + radiance.
struct PgClusterPass {
    const float4 *A;
    const float4 *B;
    const float4 *wiAll;
    const float2 *pm;
    const int *clusters;
    int nListed;
    const int *offsets;
    int nClusters;
    int n;
    const float4 *tempIn;
    const float4 *direct;
    float4 *clusterIn;
    float4 *clusterOut;
    float4 *radiance;
    float4 *tempOut;
    int scatterOrder;            // 0: MX on computeNoneZeroElements' elements, (bsdf * inv) * temp; 1: allGPUClusterScatterRadiance, (bsdf * temp) * inv, not to the sender's own successor
    template <int W> __device__ __forceinline__ void run(int c, int lane, unsigned int mask) const {
        if (c >= nClusters) return;
        const int m = pgClusterSize(offsets, c, nClusters, nListed);
        const int *mem = clusters + offsets[c];
        float in[3] = {0.f, 0.f, 0.f}, out[3] = {0.f, 0.f, 0.f};
        for (int t0 = 0; t0 < m; t0 += W) {
            const int t = t0 + lane;
            const bool valid = t < m;
            const int j = mem[valid ? t : 0];
            PgVertex v; pgLoadA(A, j, v); pgLoadB(B, j, v);
            float acc[3] = {0.f, 0.f, 0.f};
            for (int s0 = 0; s0 < m; s0 += W) {
                float2 ps = make_float2(-1.f, 0.f); float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f), src = w4;      // sender s0 + lane (pgSenderTerms, wi | index, temp[i + 1])
                if (s0 + lane < m) { const int i = mem[s0 + lane]; ps = pm[i]; if (ps.y != 0.0f) { src = tempIn[i + 1]; if (ps.x >= 0.0f) { w4 = wiAll[i]; w4.w = __int_as_float(i); } } }
                const int cnt = min(W, m - s0);
                for (int q = 0; q < cnt; ++q) {
                    const float2 g = pgBcast2<W>(mask, ps, q);
                    if (g.y == 0.0f) continue;                   // no successor: hands nothing on (the same for every lane of the group)
                    if (scatterOrder && g.x < 0.0f) continue;    // allGPUClusterScatterRadiance: a sender without a usable density does nothing at all
                    float e[3] = {0.f, 0.f, 0.f};
                    const float4 sr = pgBcast4<W>(mask, src, q);
                    if (g.x >= 0.0f) {
                        const float4 sw = pgBcast4<W>(mask, w4, q);
                        const float wi[3] = {sw.x, sw.y, sw.z};
                        float f[3]; pgBsdf(v, wi, f);
                        if (scatterOrder) {
                            if (j != __float_as_int(sw.w) + 1) { acc[0] += (f[0] * sr.x) * g.x; acc[1] += (f[1] * sr.y) * g.x; acc[2] += (f[2] * sr.z) * g.x; }
                            continue;
                        }
                        e[0] = f[0] * g.x; e[1] = f[1] * g.x; e[2] = f[2] * g.x;
                    }
                    acc[0] += e[0] * sr.x; acc[1] += e[1] * sr.y; acc[2] += e[2] * sr.z;
                }
            }
            if (valid) {
                radiance[j] = make_float4(acc[0], acc[1], acc[2], 0.f);
                out[0] += acc[0]; out[1] += acc[1]; out[2] += acc[2];
                if (v.nidx > 0 && j + 1 < n) { const float4 sr = tempIn[j + 1]; in[0] += sr.x; in[1] += sr.y; in[2] += sr.z; }
            }
        }
        pgGroupSum3<W>(mask, in); pgGroupSum3<W>(mask, out);
        const float4 pin = clusterIn[c], pout = clusterOut[c];
        in[0] = pin.x + in[0]; in[1] = pin.y + in[1]; in[2] = pin.z + in[2];
        out[0] = pout.x + out[0]; out[1] = pout.y + out[1]; out[2] = pout.z + out[2];
        float ratio[3]; pgRatio(in, out, ratio);
        __syncwarp(mask);
        if (lane == 0) { clusterIn[c] = make_float4(in[0], in[1], in[2], 0.f); clusterOut[c] = make_float4(out[0], out[1], out[2], 0.f); }
        for (int t = lane; t < m; t += W) {
            const int j = mem[t];
            float4 r = radiance[j];
            r = make_float4(r.x * ratio[0], r.y * ratio[1], r.z * ratio[2], 0.f);
            radiance[j] = r;
            if (tempOut) { const float4 d = direct[j]; tempOut[j] = make_float4(d.x + r.x, d.y + r.y, d.z + r.z, 0.f); }
        }
    }
};
__global__ void __launch_bounds__(kPgWarps * 32) k_pgClusterPass(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll,
                                                                 const float2 *__restrict__ pm, const int *__restrict__ clusters, int nListed,
                                                                 const int *__restrict__ offsets, int nClusters, int n, const float4 *__restrict__ tempIn,
                                                                 const float4 *__restrict__ direct, float4 *__restrict__ clusterIn, float4 *__restrict__ clusterOut,
                                                                 float4 *__restrict__ radiance, float4 *__restrict__ tempOut, int scatterOrder) {
    pgClusterDispatch(PgClusterPass{A, B, wiAll, pm, clusters, nListed, offsets, nClusters, n, tempIn, direct, clusterIn, clusterOut, radiance, tempOut, scatterOrder}, offsets, nClusters, nListed);
}
/// The same pass for the vertices outside every cluster (MX's else branch: the vertex's own element, no ratio).
__global__ void k_pgLoosePass(const float4 *__restrict__ A, const float4 *__restrict__ B, const float4 *__restrict__ wiAll, const float4 *__restrict__ Lem,
                              const float2 *__restrict__ pm, int n, const float4 *__restrict__ tempIn, const float4 *__restrict__ direct,
                              float4 *__restrict__ radiance, float4 *__restrict__ tempOut, int scatterOrder) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || __float_as_int(Lem[i].w) != -1) return;
    PgVertex v; pgLoadA(A, i, v); pgLoadB(B, i, v);
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    const float2 ps = pm[i];
    if (ps.y != 0.0f) {
        float e[3] = {0.f, 0.f, 0.f};
        if (ps.x >= 0.0f) {
            const float inv = ps.x;
            const float4 w4 = wiAll[i];
            const float wi[3] = {w4.x, w4.y, w4.z};
            float f[3]; pgBsdf(v, wi, f);
            e[0] = f[0] * inv; e[1] = f[1] * inv; e[2] = f[2] * inv;
            if (scatterOrder) { const float4 src = tempIn[i + 1]; e[0] = (f[0] * src.x) * inv; e[1] = (f[1] * src.y) * inv; e[2] = (f[2] * src.z) * inv; }
        }
        const float4 src = tempIn[i + 1];
        r = scatterOrder ? make_float4(e[0], e[1], e[2], 0.f) : make_float4(e[0] * src.x, e[1] * src.y, e[2] * src.z, 0.f);
    }
    radiance[i] = r;
    if (tempOut) { const float4 d = direct[i]; tempOut[i] = make_float4(d.x + r.x, d.y + r.y, d.z + r.z, 0.f); }
}

} // namespace

void launchPathgraphClusterValidate(const NoriGpuSPoint *d_sps, int n, const int *d_clusters, int nListed, const int *d_offsets, int nClusters,
                                    int *d_listed, int *d_flag2 /* flag, number of vertices that name a cluster */, cudaStream_t st) {
    cudaMemsetAsync(d_listed, 0, (size_t) n * sizeof(int), st);
    cudaMemsetAsync(d_flag2, 0, 2 * sizeof(int), st);
    const int work = std::max(n, nClusters);
    k_pgClusterValidate<<<(unsigned) ((work + 255) / 256), 256, 0, st>>>(d_sps, n, d_clusters, nListed, d_offsets, nClusters, d_listed, d_flag2, d_flag2 + 1);
}

size_t pgClusterScratchBytes(int n, int nClusters) {
    // A, B, wi, eLi, eLd (pack) + wid, Ld, Lb, Lem + direct, tempA, tempB, radiance (16 each) + dm (8) + pm (8) per vertex; in / out per cluster
    return (size_t) n * (64 + 48 + 3 * 16 + 4 * 16 + 4 * 16 + 8 + 8) + (size_t) nClusters * 32 + 256;
}

/// SPoints::ClusterScatterWithDirectOptNR (include/nori/shadingPoint.h:600-624): ClusterDirect, then ClusterIterations3 (src/pbsdf.cu:3475-3532)
/// on the same clusters. d_direct / d_blur / d_mc: n x 3 floats (device): the filtered direct light, the last pass's radiance
/// (ret.blur_results[iterations - 1]) and lastRun on the temp that pass read (ret.mc_results[iterations - 1]). iterations >= 1.
bool launchPathgraphCluster(const NoriGpuSPoint *d_sps, const NoriGpuLPoint *d_lps, int n, const int *d_clusters, int nListed, const int *d_offsets,
                            int nClusters, int iterations, void *d_scratch, float *d_direct, float *d_blur, float *d_mc, cudaStream_t st, std::string &err,
                            const int *d_clusterOf) {
    if (n <= 0 || nClusters < 0 || iterations < 1) { err = "pathgraph_cluster: bad sizes"; return false; }
    float4 *A = static_cast<float4 *>(d_scratch), *B = A + 4 * (size_t) n, *wi = B + 3 * (size_t) n, *eLi = wi + n, *eLd = eLi + n;
    float4 *wid = eLd + n, *Ld = wid + n, *Lb = Ld + n, *Lem = Lb + n, *direct = Lem + n, *tempA = direct + n, *tempB = tempA + n, *radiance = tempB + n;
    float4 *cin = radiance + n, *cout = cin + nClusters;
    float2 *dm = reinterpret_cast<float2 *>(cout + nClusters);
    float2 *pm = dm + n;
    const unsigned gt = (unsigned) ((n + 255) / 256), gc = (unsigned) (((nClusters + 3) / 4 + kPgWarps - 1) / kPgWarps);    // a warp takes four clusters
    k_pgPack<<<gt, 256, 0, st>>>(d_sps, n, A, B, wi, eLi, eLd);
    k_pgPackLights<<<gt, 256, 0, st>>>(d_sps, d_lps, d_clusterOf, n, wid, Ld, Lb, Lem);
    if (nClusters > 0) {
        k_pgClusterDirectMarginal<<<gc, kPgWarps * 32, 0, st>>>(A, B, wi, wid, Ld, Lb, d_clusters, nListed, d_offsets, nClusters, dm);
        k_pgClusterDirect<<<gc, kPgWarps * 32, 0, st>>>(A, B, wi, wid, Ld, Lb, Lem, dm, d_clusters, nListed, d_offsets, nClusters, radiance, direct);
        k_pgClusterMarginal<<<gc, kPgWarps * 32, 0, st>>>(A, B, wi, d_clusters, nListed, d_offsets, nClusters, n, pm);
        cudaMemsetAsync(cin, 0, (size_t) nClusters * 2 * sizeof(float4), st);
    }
    k_pgLooseDirect<<<gt, 256, 0, st>>>(eLd, Lem, n, direct);
    k_pgLooseMarginal<<<gt, 256, 0, st>>>(A, B, wi, Lem, n, pm);
    k_pgUnpack<<<gt, 256, 0, st>>>(direct, d_direct, n);
    const float4 *tin = eLi;                                 // updateRadiance(..., 0): temp = eLi
    for (int it = 0; it < iterations; ++it) {
        float4 *tout = it == iterations - 1 ? nullptr : ((it & 1) ? tempB : tempA);
        if (nClusters > 0)
            k_pgClusterPass<<<gc, kPgWarps * 32, 0, st>>>(A, B, wi, pm, d_clusters, nListed, d_offsets, nClusters, n, tin, direct, cin, cout, radiance, tout, 0);
        k_pgLoosePass<<<gt, 256, 0, st>>>(A, B, wi, Lem, pm, n, tin, direct, radiance, tout, 0);
        if (tout) tin = tout;
    }
    k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_blur, n);
    k_pgLastRun<<<gt, 256, 0, st>>>(A, B, wi, tin, radiance, n);
    k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_mc, n);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("pathgraph_cluster: ") + cudaGetErrorString(e); return false; }
    return true;
}

/// The tool's cluster modes without direct light: scatterOrder 0 = SPoints::ClusterScatter2 (mode "n": precomputedMatrixElemtns + ClusterIterations,
/// src/pbsdf.cu:3359-3418), 1 = SPoints::ClusterScatter / loadClusterScatter (modes "t", "l": computeClusterScatterAllOnGPURecord / ClusterScatterRecord,
/// :3108-3172, 3030-3106, with allGPUClusterScatterRadiance :1309-1345). Per pass: the cluster block, the clusters' running totals and ratio -> blur[it];
/// temp = radiance + eLd (updateRadiance with it + 1); lastRun -> mc[it]. d_blur / d_mc: iterations x n x 3 floats. Scratch: pgClusterScratchBytes.
bool launchPathgraphClusterScatter(const NoriGpuSPoint *d_sps, int n, const int *d_clusters, int nListed, const int *d_offsets, int nClusters, int iterations,
                                   int scatterOrder, void *d_scratch, float *d_blur, float *d_mc, cudaStream_t st, std::string &err) {
    if (n <= 0 || nClusters < 0 || iterations < 1) { err = "pathgraph_cluster_scatter: bad sizes"; return false; }
    float4 *A = static_cast<float4 *>(d_scratch), *B = A + 4 * (size_t) n, *wi = B + 3 * (size_t) n, *eLi = wi + n, *eLd = eLi + n;
    float4 *wid = eLd + n, *Ld = wid + n, *Lb = Ld + n, *Lem = Lb + n, *direct = Lem + n, *temp = direct + n, *tempB = temp + n, *radiance = tempB + n;
    float4 *cin = radiance + n, *cout = cin + nClusters;
    float2 *dm = reinterpret_cast<float2 *>(cout + nClusters);
    float2 *pm = dm + n;
    const unsigned gt = (unsigned) ((n + 255) / 256), gc = (unsigned) (((nClusters + 3) / 4 + kPgWarps - 1) / kPgWarps);
    k_pgPack<<<gt, 256, 0, st>>>(d_sps, n, A, B, wi, eLi, eLd);
    k_pgPackLights<<<gt, 256, 0, st>>>(d_sps, nullptr, nullptr, n, wid, Ld, Lb, Lem);          // only Lem.w = the vertex's cluster is read
    if (nClusters > 0) {
        k_pgClusterMarginal<<<gc, kPgWarps * 32, 0, st>>>(A, B, wi, d_clusters, nListed, d_offsets, nClusters, n, pm);
        cudaMemsetAsync(cin, 0, (size_t) nClusters * 2 * sizeof(float4), st);
    }
    k_pgLooseMarginal<<<gt, 256, 0, st>>>(A, B, wi, Lem, n, pm);
    k_pgUpdate<<<gt, 256, 0, st>>>(eLi, eLd, radiance, temp, n, 0);                            // temp = eLi
    for (int it = 0; it < iterations; ++it) {
        if (nClusters > 0)
            k_pgClusterPass<<<gc, kPgWarps * 32, 0, st>>>(A, B, wi, pm, d_clusters, nListed, d_offsets, nClusters, n, temp, nullptr, cin, cout, radiance, nullptr, scatterOrder);
        k_pgLoosePass<<<gt, 256, 0, st>>>(A, B, wi, Lem, pm, n, temp, nullptr, radiance, nullptr, scatterOrder);
        k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_blur + (size_t) it * 3 * n, n);
        k_pgUpdate<<<gt, 256, 0, st>>>(eLi, eLd, radiance, temp, n, it + 1);                   // temp = radiance + eLd
        k_pgLastRun<<<gt, 256, 0, st>>>(A, B, wi, temp, radiance, n);
        k_pgUnpack<<<gt, 256, 0, st>>>(radiance, d_mc + (size_t) it * 3 * n, n);
    }
    (void) direct; (void) tempB;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("pathgraph_cluster_scatter: ") + cudaGetErrorString(e); return false; }
    return true;
}

unsigned pathgraphScatterChecks(unsigned *line) { return fetchChecks(line); }

} // namespace ng
