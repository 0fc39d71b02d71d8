// noracle_pathgraph_scatter.h — CPU restatement of the path-graph tool's marginal densities and radiance scatter (TEST
// INFRASTRUCTURE; never part of the product). Follows the reference's CUDA code:
//   dot ... refract (clamped dot product, Beckmann-shaped lobe, Fresnel terms, Smith G1)   src/pbsdf.cu:257-461
//   bsdfeval_device / pdf_device                                                           src/pbsdf.cu:464-628
//   allGPUPdfMarginal / allGPUScatterRadiance / lastRun / updateRadiance                   src/pbsdf.cu:1738-1756, 1398-1431, 1497-1526, 2185-2194
//   computeScatterAllOnGPURecord (the sequence of the passes)                              src/pbsdf.cu:2317-2368
//   allGPUPdfSum / allGPUMISRadiance / computeMISAllOnGPURecord (the MIS gather variant)   src/pbsdf.cu:1600-1617, 1346-1370, 2922-2965
// Float / double mixing is written out where the reference's literals (1.0, 4.0, M_PI, M_1_PI) promote an expression to double.
// The reference is compiled by nvcc with -fmad=true, this file with -ffp-contract=off: agreement with the reference is to
// rounding (tests: rtol 2e-4 on the radiance rows, which also covers the undefined order of its float atomics), agreement with
// the product's kernels (compiled -fmad=false, same operation order) is to the bit for the marginal densities and the mc rows.
// Summation orders: the marginal density adds neighbours t = l, l + 32, ... per "lane" l and combines the 32 partial sums with
// a butterfly (xor 16, 8, 4, 2, 1) as the product's warp does; `sequential = true` adds them in list order as the reference's
// single thread does (used against the reference). The scatter adds in vertex order (the reference: atomics, any order).
// Pinned by: oracle/_ref/libpbsdf_ref.so (the reference's own src/pbsdf.cu, compiled for sm_100a) run on the GPU box
// (tests/test_pathgraph_scatter.py) and the fixture it produced there (tests/golden/pathgraph_scatter_ref.npz).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <vector>

namespace noracle {

#pragma pack(push, 1)
struct SPointRec {           // SPoint, include/nori/shadingPoint.h:125-147 (168 bytes with its trailing padding)
    float pos[3], wi[3], wi_d[3], wo[3], shN[3], geoN[3], diffuse[3], specular[3], eLi[3], eLd[3], eta[3], k[3];
    float roughness, pdf, rrpdf;
    int32_t nidx, groupIdx;
    char bsdf_type; char pad[3];
};
#pragma pack(pop)
static_assert(sizeof(SPointRec) == 168, "SPoint is 168 bytes");

namespace pgs {

// `ret += a[i] * b[i]` is a chain of fused multiply-adds in the reference's build (nvcc contracts it): written as fmaf here,
// because the half-vector cosines feed 1 - c * c, which turns an ulp of c into 1e-4 of a near-specular lobe.
inline float cdot(const float *a, const float *b) {
    float r = a[0] * b[0];
    r = std::fmaf(a[1], b[1], r); r = std::fmaf(a[2], b[2], r);
    return std::fmin(std::fmax(0.0f, r), 1.0f);
}
inline void unit(float *a) {
    float t = a[0] * a[0];
    t = std::fmaf(a[1], a[1], t); t = std::fmaf(a[2], a[2], t);
    t = std::sqrt(t);
    a[0] /= t; a[1] /= t; a[2] /= t;
}
inline float hyp(float a, float b) {
    if (std::fabs(a) > std::fabs(b)) { float r = b / a; return std::fabs(a) * std::sqrt(1.0f + r * r); }
    if (b != 0.0f) { float r = a / b; return std::fabs(b) * std::sqrt(1.0f + r * r); }
    return 0.0f;
}
inline float lobe(float cosh, float alpha) {
    if (cosh <= 0.0f) return 0.0f;
    float c2 = cosh * cosh;
    float ex = (float) ((1.0 - (double) c2) / (double) (c2 * alpha * alpha));
    float root = (1.0f + ex) * c2;
    return (float) ((double) 1.0f / (M_PI * (double) alpha * (double) alpha * (double) root * (double) root));
}
inline void conductor(float *out, float ci, const float *eta, const float *k) {
    float c2 = ci * ci, s2 = 1 - c2, s4 = s2 * s2;
    for (int i = 0; i < 3; ++i) {
        float t1 = eta[i] * eta[i] - k[i] * k[i] - s2;
        float a2pb2 = (float) std::sqrt((double) (t1 * t1) + (double) (k[i] * k[i] * eta[i] * eta[i]) * 4.0);
        float a = std::sqrt((a2pb2 + t1) * 0.5f);
        float term1 = a2pb2 + c2, term2 = a * (2 * ci);
        float Rs2 = (term1 - term2) / (term1 + term2);
        float term3 = a2pb2 * c2 + s4, term4 = term2 * s2;
        float Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
        out[i] = 0.5f * (Rp2 + Rs2);
    }
}
inline float dielectric(float ci_, float ETA, float *ct_ = nullptr, bool identityShortcut = false) {
    if (identityShortcut && ETA == 1.0f) { *ct_ = -ci_; return 0.0f; }
    float scale = (ci_ > 0.0f) ? (float) (1.0 / (double) ETA) : ETA;
    float ct2 = (float) (1.0 - (1.0 - (double) (ci_ * ci_)) * (double) (scale * scale));
    if (ct2 <= 0.0f) { if (ct_) *ct_ = 0.0f; return 1.0f; }
    float ci = std::fabs(ci_), ct = std::sqrt(ct2);
    float Rs = (ci - ETA * ct) / (ci + ETA * ct), Rp = (ETA * ci - ct) / (ETA * ci + ct);
    if (ct_) *ct_ = (ci_ > 0) ? -ct : ct;
    return 0.5f * (Rs * Rs + Rp * Rp);
}
inline float g1(float dotWH, float dotWN, float alpha) {
    if (dotWH * dotWN <= 0.0f) return 0.0f;
    float c2 = dotWN * dotWN;
    float tan2 = (float) ((1.0 - (double) c2) / (double) c2);
    float tanTheta = std::fabs(std::sqrt(tan2));
    if (tanTheta == 0.0f) return 1.0f;
    return 2.0f / (1.0f + hyp(1.0f, alpha * tanTheta));
}
inline float lobePdf(float dotWoN, float dotWoH, float dotWhN, float alpha) {
    if (dotWoN == 0) return 0.0f;
    return g1(dotWoH, dotWoN, alpha) * std::fabs(dotWoH) * lobe(dotWhN, alpha) / std::fabs(dotWoN);
}

inline float pdfAt(const SPointRec &sp, const float *wi) {
    float dotWiN = cdot(wi, sp.shN), dotWiG = cdot(wi, sp.geoN), dotWoN = cdot(sp.wo, sp.shN);
    if (sp.bsdf_type == 't') {
        float ct = 0.0f;
        float F = dielectric(dotWoN, sp.eta[0], &ct, true);
        if (dotWiN * dotWoN >= 0.0f) {
            float dir[3] = {-sp.wo[0], -sp.wo[1], sp.wo[2]};
            return std::fabs(cdot(dir, wi) - 1.0f) > 0.00001f ? 0.0f : F;
        }
        float scale = (-ct < 0) ? (float) (1.0 / (double) sp.eta[0]) : sp.eta[0];
        float dir[3] = {scale * sp.wo[0], scale * sp.wo[1], ct};
        return std::fabs(cdot(dir, wi) - 1.0f) > 0.00001f ? 0.0f : (float) (1.0 - (double) F);
    }
    if (dotWiG * dotWiN <= 0.0f) return 0.0f;
    if (dotWiN <= 0.0f || dotWoN <= 0.0f) return 0.0f;
    float diffuse = (float) ((double) dotWiN * M_1_PI);
    if (sp.bsdf_type == 'd') return diffuse;
    if (sp.bsdf_type != 'o' && sp.bsdf_type != 'c') return 0.0f;
    float wh[3] = {wi[0] + sp.wo[0], wi[1] + sp.wo[1], wi[2] + sp.wo[2]};
    unit(wh);
    float dotWhN = cdot(wh, sp.shN), dotWiH = cdot(wi, wh), dotWoH = cdot(sp.wo, wh);
    float inv = (float) (1.0 / (4.0 * (double) dotWiH));
    float prob = lobePdf(dotWoN, dotWoH, dotWhN, sp.roughness);
    if (sp.bsdf_type == 'c') return prob * inv;
    float ps = dielectric(dotWoN, 1.5f);
    float pd = std::fmax(std::fmax(sp.diffuse[0], sp.diffuse[1]), sp.diffuse[2]);
    ps = ps / (ps + pd);
    pd = (float) (1.0 - (double) ps);
    float specular = prob * inv * ps;
    return specular + diffuse * pd;
}

inline void bsdfAt(const SPointRec &sp, const float *wi, float *out) {
    out[0] = out[1] = out[2] = 0.0f;
    float dotWiN = cdot(wi, sp.shN), dotWiG = cdot(wi, sp.geoN), dotWoN = cdot(sp.wo, sp.shN);
    if (sp.bsdf_type == 't') {
        float ct = 0.0f;
        float F = dielectric(dotWoN, sp.eta[0], &ct, true);
        if (dotWiN * dotWoN >= 0.0f) {
            float dir[3] = {-sp.wo[0], -sp.wo[1], sp.wo[2]};
            if (std::fabs(cdot(wi, dir) - 1.0f) > 0.00001f) return;
            for (int i = 0; i < 3; ++i) out[i] = sp.specular[i] * F;
        } else {
            float scale = (-ct < 0) ? (float) (1.0 / (double) sp.eta[0]) : sp.eta[0];
            float dir[3] = {scale * sp.wo[0], scale * sp.wo[1], ct};
            if (std::fabs(cdot(dir, wi) - 1.0f) > 0.00001f) return;
            float factor = (ct < 0.0f) ? (float) (1.0 / (double) sp.eta[0]) : sp.eta[0];
            float f = (float) ((double) (factor * factor) * (1.0 - (double) F));
            for (int i = 0; i < 3; ++i) out[i] = sp.diffuse[i] * f;
        }
        return;
    }
    if (dotWiG * dotWiN <= 0.0f) return;
    if (dotWiN <= 0.0f || dotWoN <= 0.0f) return;
    float dc = (float) (M_1_PI * (double) dotWiN);
    float diffuse[3] = {sp.diffuse[0] * dc, sp.diffuse[1] * dc, sp.diffuse[2] * dc};
    if (sp.bsdf_type == 'd') { for (int i = 0; i < 3; ++i) out[i] = diffuse[i]; return; }
    if (sp.bsdf_type != 'o' && sp.bsdf_type != 'c') return;
    float wh[3] = {wi[0] + sp.wo[0], wi[1] + sp.wo[1], wi[2] + sp.wo[2]};
    unit(wh);
    float dotWhN = cdot(wh, sp.shN);
    float D = lobe(dotWhN, sp.roughness);
    if (sp.bsdf_type == 'c' && D == 0.0f) return;
    float dotWoH = cdot(sp.wo, wh), dotWiH = cdot(wi, wh);
    float G = g1(dotWoH, dotWoN, sp.roughness) * g1(dotWiH, dotWiN, sp.roughness);
    if (sp.bsdf_type == 'o') {
        float F = dielectric(dotWoH, 1.5f);
        float sc = F * G * D / (4.0f * dotWoN);
        float T1221 = (float) ((1.0 - (double) dielectric(dotWoN, 1.5f)) * (1.0 - (double) dielectric(dotWiN, 1.5f)));
        for (int i = 0; i < 3; ++i) out[i] = diffuse[i] * T1221 + sp.specular[i] * sc;
    } else {
        float model = D * G / (4.0f * dotWoN);
        float F[3];
        conductor(F, dotWoH, sp.eta, sp.k);
        for (int i = 0; i < 3; ++i) out[i] = (F[i] * sp.specular[i]) * model;
    }
}

} // namespace pgs

/// computeScatterAllOnGPURecord. neighbors: n rows of rowInts ints; blur / mc: iterations x n x 3.
inline void pathgraphScatter(const SPointRec *sps, int n, const int32_t *neighbors, int rowInts, const int32_t *knn, int iterations,
                             bool sequential, float *pdfmarginal, float *blur, float *mc, const float *direct = nullptr) {
    using namespace pgs;
    std::vector<float> pm(n), temp(3 * (size_t) n), rad(3 * (size_t) n);
    for (int i = 0; i < n; ++i) {
        const int32_t *row = neighbors + (size_t) i * rowInts;
        if (sequential) {
            float s = 0.0f;
            for (int t = 0; t < knn[i]; ++t) {
                const SPointRec &spo = sps[row[t]];
                if (spo.nidx == i) continue;
                s += pdfAt(spo, sps[i].wi) * spo.rrpdf;
            }
            pm[i] = s;
        } else {
            float lane[32];
            for (int l = 0; l < 32; ++l) {
                float s = 0.0f;
                for (int t = l; t < knn[i]; t += 32) {
                    const SPointRec &spo = sps[row[t]];
                    if (spo.nidx == i) continue;
                    s += pdfAt(spo, sps[i].wi) * spo.rrpdf;
                }
                lane[l] = s;
            }
            for (int d = 16; d > 0; d >>= 1) {
                float nxt[32];
                for (int l = 0; l < 32; ++l) nxt[l] = lane[l] + lane[l ^ d];
                std::memcpy(lane, nxt, sizeof lane);
            }
            pm[i] = lane[0];
        }
        if (pdfmarginal) pdfmarginal[i] = pm[i];
    }
    for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) temp[3 * (size_t) i + c] = sps[i].eLi[c];
    for (int it = 0; it < iterations; ++it) {
        std::fill(rad.begin(), rad.end(), 0.0f);
        for (int i = 0; i + 1 < n; ++i) {
            const SPointRec &sp = sps[i];
            if (!(pm[i] > 0.0f && sp.nidx > 0 && sp.rrpdf > 1e-7f)) continue;
            const float inv = (float) (1.0 / (double) pm[i]);
            const int32_t *row = neighbors + (size_t) i * rowInts;
            for (int t = 0; t < knn[i]; ++t) {
                const int j = row[t];
                if (j == i + 1) continue;
                float f[3]; bsdfAt(sps[j], sp.wi, f);
                for (int c = 0; c < 3; ++c) rad[3 * (size_t) j + c] += (temp[3 * (size_t) (i + 1) + c] * f[c]) * inv;
            }
        }
        std::memcpy(blur + (size_t) it * 3 * n, rad.data(), sizeof(float) * 3 * (size_t) n);
        for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c)
            temp[3 * (size_t) i + c] = direct ? direct[3 * (size_t) i + c] + rad[3 * (size_t) i + c]      // updateWithOptDirectRadiance (:2196-2199)
                                              : (it == 0 ? sps[i].eLi[c] : rad[3 * (size_t) i + c] + sps[i].eLd[c]);
        float *out = mc + (size_t) it * 3 * n;
        for (int i = 0; i < n; ++i) {
            out[3 * (size_t) i] = out[3 * (size_t) i + 1] = out[3 * (size_t) i + 2] = 0.0f;
            const SPointRec &sp = sps[i];
            const float spdf = pdfAt(sp, sp.wi);
            if (spdf > 0.0f && sp.nidx > 0 && i + 1 < n) {
                const float inv = (float) (1.0 / (double) (spdf * sp.rrpdf));
                float f[3]; bsdfAt(sp, sp.wi, f);
                for (int c = 0; c < 3; ++c) out[3 * (size_t) i + c] = (temp[3 * (size_t) (i + 1) + c] * f[c]) * inv;
            }
        }
    }
}

/// computeMISAllOnGPURecord (src/pbsdf.cu:2922-2965): allGPUPdfSum (:1600-1617), then per pass allGPUMISRadiance (:1346-1370),
/// updateRadiance with it + 1, lastRun. Sums in lane order + butterfly (the product's warp) or, `sequential`, in list order (the
/// reference's single thread).
inline void pathgraphMis(const SPointRec *sps, int n, const int32_t *neighbors, int rowInts, const int32_t *knn, int iterations,
                         bool sequential, float *blur, float *mc) {
    using namespace pgs;
    std::vector<float> pdfsum((size_t) n * rowInts, 0.0f), temp(3 * (size_t) n), rad(3 * (size_t) n);
    auto butterfly = [](float *lane) {
        for (int d = 16; d > 0; d >>= 1) { float nxt[32]; for (int l = 0; l < 32; ++l) nxt[l] = lane[l] + lane[l ^ d]; std::memcpy(lane, nxt, sizeof nxt); }
        return lane[0];
    };
    for (int i = 0; i < n; ++i) {
        const int32_t *row = neighbors + (size_t) i * rowInts; const int k = knn[i];
        for (int a = 0; a < k; ++a) {
            const SPointRec &spa = sps[row[a]];
            if (!(spa.nidx > 0 && spa.rrpdf > 1e-7f)) continue;
            if (sequential) {
                float s = 0.0f;
                for (int b = 0; b < k; ++b) s += pdfAt(sps[row[b]], spa.wi) * sps[row[b]].rrpdf;
                pdfsum[(size_t) i * rowInts + a] = s;
            } else {
                float lane[32];
                for (int l = 0; l < 32; ++l) { float s = 0.0f; for (int b = l; b < k; b += 32) s += pdfAt(sps[row[b]], spa.wi) * sps[row[b]].rrpdf; lane[l] = s; }
                pdfsum[(size_t) i * rowInts + a] = butterfly(lane);
            }
        }
    }
    for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) temp[3 * (size_t) i + c] = sps[i].eLi[c];
    for (int it = 0; it < iterations; ++it) {
        for (int i = 0; i < n; ++i) {
            const int32_t *row = neighbors + (size_t) i * rowInts; const int k = knn[i];
            float lanes[3][32];
            for (int c = 0; c < 3; ++c) for (int l = 0; l < 32; ++l) lanes[c][l] = 0.0f;
            float seq[3] = {0.0f, 0.0f, 0.0f};
            for (int a = 0; a < k; ++a) {
                const int ja = row[a];
                const SPointRec &spa = sps[ja];
                const float ps = pdfsum[(size_t) i * rowInts + a];
                if (!(spa.nidx > 0 && ps > 0.0f && spa.rrpdf > 1e-7f) || ja + 1 >= n) continue;
                float f[3]; bsdfAt(sps[i], spa.wi, f);
                const float inv = (float) (1.0 / (double) ps);
                for (int c = 0; c < 3; ++c) {
                    const float term = (f[c] * temp[3 * (size_t) (ja + 1) + c]) * inv;
                    seq[c] += term; lanes[c][a & 31] += term;          // lane a % 32 adds its neighbours in order
                }
            }
            for (int c = 0; c < 3; ++c) rad[3 * (size_t) i + c] = sequential ? seq[c] : butterfly(lanes[c]);
        }
        std::memcpy(blur + (size_t) it * 3 * n, rad.data(), sizeof(float) * 3 * (size_t) n);
        for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) temp[3 * (size_t) i + c] = rad[3 * (size_t) i + c] + sps[i].eLd[c];
        float *out = mc + (size_t) it * 3 * n;
        for (int i = 0; i < n; ++i) {
            out[3 * (size_t) i] = out[3 * (size_t) i + 1] = out[3 * (size_t) i + 2] = 0.0f;
            const SPointRec &sp = sps[i];
            const float spdf = pdfAt(sp, sp.wi);
            if (spdf > 0.0f && sp.nidx > 0 && i + 1 < n) {
                const float inv = (float) (1.0 / (double) (spdf * sp.rrpdf));
                float f[3]; bsdfAt(sp, sp.wi, f);
                for (int c = 0; c < 3; ++c) out[3 * (size_t) i + c] = (temp[3 * (size_t) (i + 1) + c] * f[c]) * inv;
            }
        }
    }
}


/// LPoint, include/nori/shadingPoint.h:148-155 (44 bytes)
struct LPointRec { float L_directsample[3], L_bsdfsample[3], L_em[3], lightpdf, bsdfpdf; };
static_assert(sizeof(LPointRec) == 44, "LPoint is 44 bytes");

/// computeDirectScatterAllOnGPU (src/pbsdf.cu:2370-2395): allGPUDirectPdfMarginal (:1881-1897), allGPUDirectScatterRadiance
/// (:1651-1687), addEmitterToDirectLight (:813-818). marginal (optional): n x 2 = {k additions of the vertex's light density, the
/// neighbours' densities of wi summed in lane order + butterfly or, `sequential`, in list order}. The scatter adds in vertex order i,
/// neighbour order t (the reference and the product: float atomics, any order).
inline void pathgraphDirect(const SPointRec *sps, const LPointRec *lps, int n, const int32_t *neighbors, int rowInts, const int32_t *knn,
                            bool sequential, float *marginal, float *direct) {
    using namespace pgs;
    std::vector<float> m(2 * (size_t) n, 0.0f);
    for (int i = 0; i < n; ++i) {
        const int32_t *row = neighbors + (size_t) i * rowInts; const int k = knn[i];
        float x = 0.0f, y = 0.0f;
        for (int t = 0; t < k; ++t) x += lps[i].lightpdf;
        if (sequential) {
            for (int t = 0; t < k; ++t) y += pdfAt(sps[row[t]], sps[i].wi);
        } else {
            float lane[32];
            for (int l = 0; l < 32; ++l) { float s = 0.0f; for (int t = l; t < k; t += 32) s += pdfAt(sps[row[t]], sps[i].wi); lane[l] = s; }
            for (int d = 16; d > 0; d >>= 1) { float nxt[32]; for (int l = 0; l < 32; ++l) nxt[l] = lane[l] + lane[l ^ d]; std::memcpy(lane, nxt, sizeof nxt); }
            y = lane[0];
        }
        m[2 * (size_t) i] = x; m[2 * (size_t) i + 1] = y;
    }
    if (marginal) std::memcpy(marginal, m.data(), sizeof(float) * 2 * (size_t) n);
    for (size_t q = 0; q < 3 * (size_t) n; ++q) direct[q] = 0.0f;
    for (int i = 0; i < n; ++i) {
        const int32_t *row = neighbors + (size_t) i * rowInts; const int k = knn[i];
        const float mx = m[2 * (size_t) i], my = m[2 * (size_t) i + 1];
        const float invL = (float) (1.0 / (double) mx), invB = (float) (1.0 / (double) my);
        for (int t = 0; t < k; ++t) {
            const int j = row[t];
            float out[3] = {0.0f, 0.0f, 0.0f};
            if (mx > 0.0f) { float f[3]; bsdfAt(sps[j], sps[i].wi_d, f); for (int c = 0; c < 3; ++c) out[c] = out[c] + (lps[i].L_directsample[c] * f[c]) * invL; }
            if (my > 0.0f) { float f[3]; bsdfAt(sps[j], sps[i].wi, f); for (int c = 0; c < 3; ++c) out[c] = out[c] + (lps[i].L_bsdfsample[c] * f[c]) * invB; }
            for (int c = 0; c < 3; ++c) direct[3 * (size_t) j + c] += out[c];
        }
    }
    for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) direct[3 * (size_t) i + c] += lps[i].L_em[c];
}


/// SPoints::ClusterScatterWithDirectOptNR (include/nori/shadingPoint.h:600-624) on given clusters: ClusterDirect (src/pbsdf.cu:2400-2446:
/// ClusterDirectPdfMarginal :1899-1937, ClusterDirectRadiance :1939-2010, clampDirectCluster :2100-2125, computeRatio :2146-2172,
/// updateComputeCluster :2174-2183, addEmitterToDirectLight :813-818), then the elements of computeNoneZeroElements (:1059-1085) with
/// allGPUClusterPdfMarginal (:1239-1263) applied by ClusterIterations3 (:3475-3532: MX :1265-1307, clampCluster :2127-2144 with totals that
/// run on over the passes, updateWithOptDirectRadiance, lastRun). clusters[offsets[c] ..] = members of cluster c (the last up to nListed),
/// sps[i].groupIdx = i's cluster or -1. What a member receives is added in list order of the senders. `sequential`: every other sum in
/// list order too (one thread's order in the reference; its atomics have none); otherwise as the product's warp takes them: lane l adds the
/// members l, l + 32, ..., butterfly over the lanes, the clusters' running totals + the pass's sum.
inline void pathgraphCluster(const SPointRec *sps, const LPointRec *lps, int n, const int32_t *clusters, int nListed, const int32_t *offsets, int nClusters,
                             int iterations, bool sequential, float *direct, float *blur, float *mc) {
    using namespace pgs;
    auto size = [&](int c) { return (c == nClusters - 1 ? nListed : offsets[c + 1]) - offsets[c]; };
    auto butterfly = [](float *lane) {
        for (int d = 16; d > 0; d >>= 1) { float nxt[32]; for (int l = 0; l < 32; ++l) nxt[l] = lane[l] + lane[l ^ d]; std::memcpy(lane, nxt, sizeof nxt); }
        return lane[0];
    };
    // sum of term(t) over t < m: list order, or lane-strided partial sums + butterfly
    auto total = [&](int m, const std::function<void(int, float &)> &term) {
        if (sequential) { float s = 0.0f; for (int t = 0; t < m; ++t) term(t, s); return s; }
        float lane[32];
        for (int l = 0; l < 32; ++l) { float s = 0.0f; for (int t = l; t < m; t += 32) term(t, s); lane[l] = s; }
        return butterfly(lane);
    };
    auto ratioOf = [](float in, float out) { return in < out ? in / out : 1.0f; };
    std::vector<float> dm(2 * (size_t) n, 0.0f), pm(n, 0.0f), raw(3 * (size_t) n, 0.0f), rad(3 * (size_t) n, 0.0f), temp(3 * (size_t) n), next(3 * (size_t) n);
    std::vector<float> cin(3 * (size_t) std::max(nClusters, 1), 0.0f), cout(3 * (size_t) std::max(nClusters, 1), 0.0f);
    // ---- direct light
    for (int c = 0; c < nClusters; ++c) {
        const int32_t *mem = clusters + offsets[c]; const int m = size(c);
        for (int s = 0; s < m; ++s) {
            const int i = mem[s];
            dm[2 * (size_t) i] = total(m, [&](int t, float &x) {
                const SPointRec &o = sps[mem[t]];
                if (cdot(sps[i].wi_d, o.shN) > 0.0f) x += lps[i].lightpdf;
                x += pdfAt(o, sps[i].wi_d); });
            dm[2 * (size_t) i + 1] = total(m, [&](int t, float &y) {
                const SPointRec &o = sps[mem[t]];
                if (cdot(sps[i].wi, o.shN) > 0.0f) y += lps[i].bsdfpdf;
                y += pdfAt(o, sps[i].wi); });
        }
        for (int t = 0; t < m; ++t) {
            const int j = mem[t];
            float acc[3] = {0.0f, 0.0f, 0.0f};
            for (int s = 0; s < m; ++s) {
                const int i = mem[s];
                const float mx = dm[2 * (size_t) i], my = dm[2 * (size_t) i + 1];
                float fin[3] = {0.0f, 0.0f, 0.0f};
                if (mx > 0.0f) { const float inv = (float) (1.0 / (double) mx); float f[3]; bsdfAt(sps[j], sps[i].wi_d, f); for (int k = 0; k < 3; ++k) fin[k] = fin[k] + (lps[i].L_directsample[k] * f[k]) * inv; }
                if (my > 0.0f) { const float inv = (float) (1.0 / (double) my); float f[3]; bsdfAt(sps[j], sps[i].wi, f); for (int k = 0; k < 3; ++k) fin[k] = fin[k] + (lps[i].L_bsdfsample[k] * f[k]) * inv; }
                for (int k = 0; k < 3; ++k) acc[k] += fin[k];
            }
            for (int k = 0; k < 3; ++k) raw[3 * (size_t) j + k] = acc[k];
        }
        float ratio[3];
        for (int k = 0; k < 3; ++k) {
            const float in = total(m, [&](int t, float &a) { a += lps[mem[t]].L_directsample[k]; a += lps[mem[t]].L_bsdfsample[k]; });
            const float out = total(m, [&](int t, float &a) { a += raw[3 * (size_t) mem[t] + k]; });
            ratio[k] = ratioOf(in, out);
        }
        for (int t = 0; t < m; ++t) for (int k = 0; k < 3; ++k) direct[3 * (size_t) mem[t] + k] = raw[3 * (size_t) mem[t] + k] * ratio[k] + lps[mem[t]].L_em[k];
    }
    for (int i = 0; i < n; ++i) if (sps[i].groupIdx == -1) for (int k = 0; k < 3; ++k) direct[3 * (size_t) i + k] = sps[i].eLd[k] + lps[i].L_em[k];
    // ---- densities of the continuation directions inside the clusters
    for (int c = 0; c < nClusters; ++c) {
        const int32_t *mem = clusters + offsets[c]; const int m = size(c);
        for (int s = 0; s < m; ++s) {
            const int i = mem[s];
            pm[i] = total(m, [&](int t, float &a) { const SPointRec &o = sps[mem[t]]; if (o.nidx == i) return; a += pdfAt(o, sps[i].wi) * o.rrpdf; });
        }
    }
    for (int i = 0; i < n; ++i) if (sps[i].groupIdx == -1) pm[i] = pdfAt(sps[i], sps[i].wi) * sps[i].rrpdf;
    // ---- passes
    auto element = [&](int i, int j, float *e) {          // what computeNoneZeroElements stores for sender i, receiver j (0 where it writes nothing)
        e[0] = e[1] = e[2] = 0.0f;
        if (pm[i] > 0.0f && sps[i].rrpdf > 1e-7f) { const float inv = (float) (1.0 / (double) pm[i]); float f[3]; bsdfAt(sps[j], sps[i].wi, f); for (int k = 0; k < 3; ++k) e[k] = f[k] * inv; }
    };
    for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) temp[3 * (size_t) i + k] = sps[i].eLi[k];
    for (int it = 0; it < iterations; ++it) {
        for (int c = 0; c < nClusters; ++c) {
            const int32_t *mem = clusters + offsets[c]; const int m = size(c);
            for (int t = 0; t < m; ++t) {
                const int j = mem[t];
                float acc[3] = {0.0f, 0.0f, 0.0f};
                for (int s = 0; s < m; ++s) {
                    const int i = mem[s];
                    if (sps[i].nidx <= 0 || i + 1 >= n) continue;
                    float e[3]; element(i, j, e);
                    for (int k = 0; k < 3; ++k) acc[k] += e[k] * temp[3 * (size_t) (i + 1) + k];
                }
                for (int k = 0; k < 3; ++k) rad[3 * (size_t) j + k] = acc[k];
            }
            for (int k = 0; k < 3; ++k) {
                const float in = total(m, [&](int t, float &a) { const int i = mem[t]; if (sps[i].nidx > 0 && i + 1 < n) a += temp[3 * (size_t) (i + 1) + k]; });
                const float out = total(m, [&](int t, float &a) { a += rad[3 * (size_t) mem[t] + k]; });
                cin[3 * (size_t) c + k] = cin[3 * (size_t) c + k] + in; cout[3 * (size_t) c + k] = cout[3 * (size_t) c + k] + out;
                const float ratio = ratioOf(cin[3 * (size_t) c + k], cout[3 * (size_t) c + k]);
                for (int t = 0; t < m; ++t) rad[3 * (size_t) mem[t] + k] *= ratio;
            }
        }
        for (int i = 0; i < n; ++i) if (sps[i].groupIdx == -1) {
            float e[3] = {0.0f, 0.0f, 0.0f};
            const bool has = sps[i].nidx > 0 && i + 1 < n;
            if (has) element(i, i, e);
            for (int k = 0; k < 3; ++k) rad[3 * (size_t) i + k] = has ? e[k] * temp[3 * (size_t) (i + 1) + k] : 0.0f;
        }
        if (it != iterations - 1) { for (size_t q = 0; q < 3 * (size_t) n; ++q) next[q] = direct[q] + rad[q]; temp.swap(next); }
    }
    std::memcpy(blur, rad.data(), sizeof(float) * 3 * (size_t) n);
    for (int i = 0; i < n; ++i) {
        mc[3 * (size_t) i] = mc[3 * (size_t) i + 1] = mc[3 * (size_t) i + 2] = 0.0f;
        const SPointRec &sp = sps[i];
        const float spdf = pdfAt(sp, sp.wi);
        if (spdf > 0.0f && sp.nidx > 0 && i + 1 < n) {
            const float inv = (float) (1.0 / (double) (spdf * sp.rrpdf));
            float f[3]; bsdfAt(sp, sp.wi, f);
            for (int k = 0; k < 3; ++k) mc[3 * (size_t) i + k] = (temp[3 * (size_t) (i + 1) + k] * f[k]) * inv;
        }
    }
}


/// The tool's cluster modes without direct light. scatterOrder false: SPoints::ClusterScatter2 (mode "n": precomputedMatrixElemtns + ClusterIterations,
/// src/pbsdf.cu:3359-3418, the elements applied by MX); true: SPoints::ClusterScatter / loadClusterScatter (modes "t" / "l":
/// computeClusterScatterAllOnGPURecord :3108-3172, allGPUClusterScatterRadiance :1309-1345: (bsdf * temp) * inv, nothing to the sender's own successor,
/// nothing at all from a sender whose density is unusable). Per pass: blocks, running cluster totals + ratio -> blur[it]; temp = radiance + eLd; lastRun -> mc[it].
inline void pathgraphClusterScatter(const SPointRec *sps, int n, const int32_t *clusters, int nListed, const int32_t *offsets, int nClusters, int iterations,
                                    bool scatterOrder, bool sequential, float *blur, float *mc) {
    using namespace pgs;
    auto size = [&](int c) { return (c == nClusters - 1 ? nListed : offsets[c + 1]) - offsets[c]; };
    auto butterfly = [](float *lane) {
        for (int d = 16; d > 0; d >>= 1) { float nxt[32]; for (int l = 0; l < 32; ++l) nxt[l] = lane[l] + lane[l ^ d]; std::memcpy(lane, nxt, sizeof nxt); }
        return lane[0];
    };
    auto total = [&](int m, const std::function<void(int, float &)> &term) {
        if (sequential) { float s = 0.0f; for (int t = 0; t < m; ++t) term(t, s); return s; }
        float lane[32];
        for (int l = 0; l < 32; ++l) { float s = 0.0f; for (int t = l; t < m; t += 32) term(t, s); lane[l] = s; }
        return butterfly(lane);
    };
    std::vector<float> pm(n, 0.0f), rad(3 * (size_t) n, 0.0f), temp(3 * (size_t) n);
    std::vector<float> cin(3 * (size_t) std::max(nClusters, 1), 0.0f), cout(3 * (size_t) std::max(nClusters, 1), 0.0f);
    for (int c = 0; c < nClusters; ++c) {
        const int32_t *mem = clusters + offsets[c]; const int m = size(c);
        for (int s = 0; s < m; ++s) {
            const int i = mem[s];
            pm[i] = total(m, [&](int t, float &a) { const SPointRec &o = sps[mem[t]]; if (o.nidx == i) return; a += pdfAt(o, sps[i].wi) * o.rrpdf; });
        }
    }
    for (int i = 0; i < n; ++i) if (sps[i].groupIdx == -1) pm[i] = pdfAt(sps[i], sps[i].wi) * sps[i].rrpdf;
    // what sender i hands to receiver j from temp; returns false if nothing is added at all
    auto handed = [&](int i, int j, const float *src, float *out) {
        const bool live = pm[i] > 0.0f && sps[i].rrpdf > 1e-7f;
        out[0] = out[1] = out[2] = 0.0f;
        if (scatterOrder) {
            if (!live || j == i + 1) return false;
            const float inv = (float) (1.0 / (double) pm[i]); float f[3]; bsdfAt(sps[j], sps[i].wi, f);
            for (int k = 0; k < 3; ++k) out[k] = (f[k] * src[k]) * inv;
            return true;
        }
        float e[3] = {0.0f, 0.0f, 0.0f};
        if (live) { const float inv = (float) (1.0 / (double) pm[i]); float f[3]; bsdfAt(sps[j], sps[i].wi, f); for (int k = 0; k < 3; ++k) e[k] = f[k] * inv; }
        for (int k = 0; k < 3; ++k) out[k] = e[k] * src[k];
        return true;
    };
    for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) temp[3 * (size_t) i + k] = sps[i].eLi[k];
    for (int it = 0; it < iterations; ++it) {
        for (int c = 0; c < nClusters; ++c) {
            const int32_t *mem = clusters + offsets[c]; const int m = size(c);
            for (int t = 0; t < m; ++t) {
                const int j = mem[t];
                float acc[3] = {0.0f, 0.0f, 0.0f};
                for (int s = 0; s < m; ++s) {
                    const int i = mem[s];
                    if (sps[i].nidx <= 0 || i + 1 >= n) continue;
                    float h[3];
                    if (handed(i, j, &temp[3 * (size_t) (i + 1)], h)) for (int k = 0; k < 3; ++k) acc[k] += h[k];
                }
                for (int k = 0; k < 3; ++k) rad[3 * (size_t) j + k] = acc[k];
            }
            for (int k = 0; k < 3; ++k) {
                const float in = total(m, [&](int t, float &a) { const int i = mem[t]; if (sps[i].nidx > 0 && i + 1 < n) a += temp[3 * (size_t) (i + 1) + k]; });
                const float out = total(m, [&](int t, float &a) { a += rad[3 * (size_t) mem[t] + k]; });
                cin[3 * (size_t) c + k] = cin[3 * (size_t) c + k] + in; cout[3 * (size_t) c + k] = cout[3 * (size_t) c + k] + out;
                const float ratio = cin[3 * (size_t) c + k] < cout[3 * (size_t) c + k] ? cin[3 * (size_t) c + k] / cout[3 * (size_t) c + k] : 1.0f;
                for (int t = 0; t < m; ++t) rad[3 * (size_t) mem[t] + k] *= ratio;
            }
        }
        for (int i = 0; i < n; ++i) if (sps[i].groupIdx == -1) {
            float h[3] = {0.0f, 0.0f, 0.0f};
            if (sps[i].nidx > 0 && i + 1 < n) {
                // the vertex's own element; in scatter order there is no successor exception here (allGPUClusterScatterRadiance's else branch)
                const bool live = pm[i] > 0.0f && sps[i].rrpdf > 1e-7f;
                if (live) {
                    const float inv = (float) (1.0 / (double) pm[i]); float f[3]; bsdfAt(sps[i], sps[i].wi, f);
                    for (int k = 0; k < 3; ++k) h[k] = scatterOrder ? (f[k] * temp[3 * (size_t) (i + 1) + k]) * inv : (f[k] * inv) * temp[3 * (size_t) (i + 1) + k];
                } else if (!scatterOrder) for (int k = 0; k < 3; ++k) h[k] = 0.0f * temp[3 * (size_t) (i + 1) + k];
            }
            for (int k = 0; k < 3; ++k) rad[3 * (size_t) i + k] = h[k];
        }
        std::memcpy(blur + (size_t) it * 3 * n, rad.data(), sizeof(float) * 3 * (size_t) n);
        for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) temp[3 * (size_t) i + k] = rad[3 * (size_t) i + k] + sps[i].eLd[k];
        float *out = mc + (size_t) it * 3 * n;
        for (int i = 0; i < n; ++i) {
            out[3 * (size_t) i] = out[3 * (size_t) i + 1] = out[3 * (size_t) i + 2] = 0.0f;
            const SPointRec &sp = sps[i];
            const float spdf = pdfAt(sp, sp.wi);
            if (spdf > 0.0f && sp.nidx > 0 && i + 1 < n) {
                const float inv = (float) (1.0 / (double) (spdf * sp.rrpdf));
                float f[3]; bsdfAt(sp, sp.wi, f);
                for (int k = 0; k < 3; ++k) out[3 * (size_t) i + k] = (temp[3 * (size_t) (i + 1) + k] * f[k]) * inv;
            }
        }
    }
}

} // namespace noracle
