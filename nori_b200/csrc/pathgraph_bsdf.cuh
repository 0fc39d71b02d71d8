// pathgraph_bsdf.cuh — the path-graph tool's own BSDF / density evaluation at a recorded vertex (reference src/pbsdf.cu:257-628),
// restated for the scatter stage (pathgraph_scatter.cu). These are NOT Nori's BSDF plugins: the tool re-evaluates the materials
// from the fields of the vertex record (type tag, albedo, specular colour, Beckmann roughness, eta, k), in WORLD space, with its
// own conventions, all of which are kept because the stage's output is defined by them:
//   * every dot product is clamped into [0, 1] (pbsdf.cu:257-264), so a direction below the surface reads as grazing;
//   * float / double mixing as written there: literals such as 1.0, 4.0, M_PI, M_1_PI promote the surrounding expression to
//     double, the result is rounded to float on assignment. Each such place is marked `(double)` below;
//   * the dielectric branch compares wi with a "mirror direction" built from wo's components as if wo were in the local frame
//     (pbsdf.cu:450-461, 475-480): for world-space records it almost never matches, the branch returns 0. Kept as it is.
// The file is compiled with -fmad=false: every operation below is one IEEE operation, fused only where written (the oracle restates them the same way,
// so the two agree to the bit; the reference itself is compiled with nvcc's default contraction: tolerance in the tests).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace ng {

/// What the evaluations read of a vertex (a subset of NoriGpuSPoint = SPoint, include/nori/shadingPoint.h:125-147).
struct PgVertex {
    float shN[3], geoN[3], wo[3];
    float diffuse[3], specular[3], eta[3], k[3];
    float roughness, rrpdf;
    int nidx;
    int type;                   // 'd', 'o', 'c', 't'
};

#define PG_FN __device__ __forceinline__

/// dot (pbsdf.cu:257-264): clamped into [0, 1]. `ret += a[i] * b[i]` is a chain of fused multiply-adds in the reference's
/// build; kept explicitly, because the half-vector cosines feed 1 - c * c (an ulp of c = 1e-4 of a near-specular lobe).
PG_FN float pgDot(const float *a, const float *b) {
    float r = a[0] * b[0];
    r = __fmaf_rn(a[1], b[1], r); r = __fmaf_rn(a[2], b[2], r);
    return fminf(fmaxf(0.0f, r), 1.0f);
}
/// normalize (pbsdf.cu:266-275)
PG_FN void pgNormalize(float *a) {
    float t = a[0] * a[0];
    t = __fmaf_rn(a[1], a[1], t); t = __fmaf_rn(a[2], a[2], t);
    t = sqrtf(t);
    a[0] /= t; a[1] /= t; a[2] /= t;
}
/// hypot2 (pbsdf.cu:326-338)
PG_FN float pgHypot2(float a, float b) {
    float r;
    if (fabsf(a) > fabsf(b)) { r = b / a; r = fabsf(a) * sqrtf(1.0f + r * r); }
    else if (b != 0.0f) { r = a / b; r = fabsf(b) * sqrtf(1.0f + r * r); }
    else r = 0.0f;
    return r;
}
/// distreval (pbsdf.cu:340-350): Beckmann-shaped lobe of the tool (GGX-like normalisation), cosh = cos of the half vector
PG_FN float pgDistrEval(float cosh, float roughness) {
    if (cosh <= 0.0f) return 0.0f;
    const float c2 = cosh * cosh;
    const float ex = (float) ((1.0 - (double) c2) / (double) (c2 * roughness * roughness));                 // (double)
    const float root = (1.0f + ex) * c2;
    return (float) ((double) 1.0f / (M_PI * (double) roughness * (double) roughness * (double) root * (double) root));   // (double)
}
/// fresnelConductorExact (pbsdf.cu:354-374)
PG_FN void pgFresnelConductor(float *result, float cosThetaI, const float *eta, const float *k) {
    const float c2 = cosThetaI * cosThetaI, s2 = 1 - c2, s4 = s2 * s2;
    for (int i = 0; i < 3; ++i) {
        const float t1 = eta[i] * eta[i] - k[i] * k[i] - s2;
        const float a2pb2 = (float) sqrt((double) (t1 * t1) + (double) (k[i] * k[i] * eta[i] * eta[i]) * 4.0);     // (double)
        const float a = sqrtf((a2pb2 + t1) * 0.5f);
        const float term1 = a2pb2 + c2, term2 = a * (2 * cosThetaI);
        const float Rs2 = (term1 - term2) / (term1 + term2);
        const float term3 = a2pb2 * c2 + s4, term4 = term2 * s2;
        const float Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
        result[i] = 0.5f * (Rp2 + Rs2);
    }
}
/// fresnelDielectricExt2 (pbsdf.cu:376-407): reflectance and the signed cosine of the transmitted direction
PG_FN float pgFresnelDielectric2(float cosThetaI_, float *cosThetaT_, float ETA) {
    if (ETA == 1.0f) { *cosThetaT_ = -cosThetaI_; return 0.0f; }
    const float scale = (cosThetaI_ > 0.0f) ? (float) (1.0 / (double) ETA) : ETA;                           // (double)
    const float cosThetaTSqr = (float) (1.0 - (1.0 - (double) (cosThetaI_ * cosThetaI_)) * (double) (scale * scale));   // (double)
    if (cosThetaTSqr <= 0.0f) { *cosThetaT_ = 0.0f; return 1.0f; }
    const float cosThetaI = fabsf(cosThetaI_), cosThetaT = sqrtf(cosThetaTSqr);
    const float Rs = (cosThetaI - ETA * cosThetaT) / (cosThetaI + ETA * cosThetaT);
    const float Rp = (ETA * cosThetaI - cosThetaT) / (ETA * cosThetaI + cosThetaT);
    *cosThetaT_ = (cosThetaI_ > 0) ? -cosThetaT : cosThetaT;
    return 0.5f * (Rs * Rs + Rp * Rp);
}
/// fresnelDielectricExt (pbsdf.cu:409-430)
PG_FN float pgFresnelDielectric(float cosThetaI_, float ETA) {
    const float scale = (cosThetaI_ > 0.0f) ? (float) (1.0 / (double) ETA) : ETA;                           // (double)
    const float cosThetaTSqr = (float) (1.0 - (1.0 - (double) (cosThetaI_ * cosThetaI_)) * (double) (scale * scale));   // (double)
    if (cosThetaTSqr <= 0.0f) return 1.0f;
    const float cosThetaI = fabsf(cosThetaI_), cosThetaT = sqrtf(cosThetaTSqr);
    const float Rs = (cosThetaI - ETA * cosThetaT) / (cosThetaI + ETA * cosThetaT);
    const float Rp = (ETA * cosThetaI - cosThetaT) / (ETA * cosThetaI + cosThetaT);
    return 0.5f * (Rs * Rs + Rp * Rp);
}
/// smithG1 (pbsdf.cu:432-443)
PG_FN float pgSmithG1(float dotWH, float dotWShN, float roughness) {
    if (dotWH * dotWShN <= 0.0f) return 0.0f;
    const float c2 = dotWShN * dotWShN;
    const float tan2 = (float) ((1.0 - (double) c2) / (double) c2);                                         // (double)
    const float tanTheta = fabsf(sqrtf(tan2));
    if (tanTheta == 0.0f) return 1.0f;
    const float root = roughness * tanTheta;
    return 2.0f / (1.0f + pgHypot2(1.0f, root));
}
/// distrpdf (pbsdf.cu:445-448): density of the visible-normal style sampling the tool assumes
PG_FN float pgDistrPdf(float dotWoShN, float dotWoH, float dotWhShN, float roughness) {
    if (dotWoShN == 0) return 0.0f;
    return pgSmithG1(dotWoH, dotWoShN, roughness) * fabsf(dotWoH) * pgDistrEval(dotWhShN, roughness) / fabsf(dotWoShN);
}

/// pdf_device (pbsdf.cu:562-628): density of direction wi at vertex sp. The reference leaves *pdf untouched on its early
/// returns and every caller passes 0: returned as 0 here.
PG_FN float pgPdf(const PgVertex &sp, const float *wi) {
    const float dotWiShN = pgDot(wi, sp.shN), dotWiGeoN = pgDot(wi, sp.geoN), dotWoShN = pgDot(sp.wo, sp.shN);
    if (sp.type == 't') {
        float cosT;
        const float F = pgFresnelDielectric2(dotWoShN, &cosT, sp.eta[0]);
        // dotWiShN * dotWoShN >= 0 always holds for clamped dots: the "reflection" side is the only one ever taken (pbsdf.cu:572)
        if (dotWiShN * dotWoShN >= 0.0f) {
            const float dir[3] = {-sp.wo[0], -sp.wo[1], sp.wo[2]};
            if (fabsf(pgDot(dir, wi) - 1.0f) > 0.00001f) return 0.0f;
            return F;
        }
        const float scale = (-cosT < 0) ? (float) (1.0 / (double) sp.eta[0]) : sp.eta[0];
        const float dir[3] = {scale * sp.wo[0], scale * sp.wo[1], cosT};
        if (fabsf(pgDot(dir, wi) - 1.0f) > 0.00001f) return 0.0f;
        return (float) (1.0 - (double) F);
    }
    if (dotWiGeoN * dotWiShN <= 0.0f) return 0.0f;
    if (dotWiShN <= 0.0f || dotWoShN <= 0.0f) return 0.0f;
    const float diffuse = (float) ((double) dotWiShN * M_1_PI);                                             // (double)
    if (sp.type == 'd') return diffuse;
    if (sp.type == 'o') {
        float pspecular = pgFresnelDielectric(dotWoShN, 1.5f);
        float pdiffuse = fmaxf(fmaxf(sp.diffuse[0], sp.diffuse[1]), sp.diffuse[2]);
        pspecular = pspecular / (pspecular + pdiffuse);
        pdiffuse = (float) (1.0 - (double) pspecular);                                                      // (double)
        float wh[3] = {wi[0] + sp.wo[0], wi[1] + sp.wo[1], wi[2] + sp.wo[2]};
        pgNormalize(wh);
        const float dotWhShN = pgDot(wh, sp.shN), dotWiH = pgDot(wi, wh), dotWoH = pgDot(sp.wo, wh);
        const float inv = (float) (1.0 / (4.0 * (double) dotWiH));                                          // (double)
        const float prob = pgDistrPdf(dotWoShN, dotWoH, dotWhShN, sp.roughness);
        const float specular = prob * inv * pspecular;
        return specular + diffuse * pdiffuse;
    }
    if (sp.type == 'c') {
        float wh[3] = {wi[0] + sp.wo[0], wi[1] + sp.wo[1], wi[2] + sp.wo[2]};
        pgNormalize(wh);
        const float dotWhShN = pgDot(wh, sp.shN), dotWiH = pgDot(wi, wh), dotWoH = pgDot(sp.wo, wh);
        const float inv = (float) (1.0 / (4.0 * (double) dotWiH));                                          // (double)
        return pgDistrPdf(dotWoShN, dotWoH, dotWhShN, sp.roughness) * inv;
    }
    return 0.0f;
}

/// bsdfeval_device (pbsdf.cu:464-560): BSDF x cosine at vertex sp for incident direction wi (out starts at 0 as every caller's
/// array does).
PG_FN void pgBsdf(const PgVertex &sp, const float *wi, float *out) {
    out[0] = out[1] = out[2] = 0.0f;
    const float dotWiShN = pgDot(wi, sp.shN), dotWiGeoN = pgDot(wi, sp.geoN), dotWoShN = pgDot(sp.wo, sp.shN);
    if (sp.type == 't') {
        float cosT = 0.0f;
        const float F = pgFresnelDielectric2(dotWoShN, &cosT, sp.eta[0]);
        if (dotWiShN * dotWoShN >= 0.0f) {
            const float dir[3] = {-sp.wo[0], -sp.wo[1], sp.wo[2]};
            if (fabsf(pgDot(wi, dir) - 1.0f) > 0.00001f) return;
            out[0] = sp.specular[0] * F; out[1] = sp.specular[1] * F; out[2] = sp.specular[2] * F;
        } else {
            const float scale = (-cosT < 0) ? (float) (1.0 / (double) sp.eta[0]) : sp.eta[0];
            const float dir[3] = {scale * sp.wo[0], scale * sp.wo[1], cosT};
            if (fabsf(pgDot(dir, wi) - 1.0f) > 0.00001f) return;
            const float factor = (cosT < 0.0f) ? (float) (1.0 / (double) sp.eta[0]) : sp.eta[0];
            const float f = (float) ((double) (factor * factor) * (1.0 - (double) F));
            out[0] = sp.diffuse[0] * f; out[1] = sp.diffuse[1] * f; out[2] = sp.diffuse[2] * f;
        }
        return;
    }
    if (dotWiGeoN * dotWiShN <= 0.0f) return;
    if (dotWiShN <= 0.0f || dotWoShN <= 0.0f) return;
    const float diffuseconst = (float) (M_1_PI * (double) dotWiShN);                                        // (double)
    float diffuse[3] = {sp.diffuse[0] * diffuseconst, sp.diffuse[1] * diffuseconst, sp.diffuse[2] * diffuseconst};
    if (sp.type == 'd') { out[0] = diffuse[0]; out[1] = diffuse[1]; out[2] = diffuse[2]; return; }
    if (sp.type == 'o') {
        float wh[3] = {wi[0] + sp.wo[0], wi[1] + sp.wo[1], wi[2] + sp.wo[2]};
        pgNormalize(wh);
        const float dotWhShN = pgDot(wh, sp.shN);
        const float D = pgDistrEval(dotWhShN, sp.roughness);
        const float dotWoH = pgDot(sp.wo, wh), dotWiH = pgDot(wi, wh);
        const float F = pgFresnelDielectric(dotWoH, 1.5f);
        const float G = pgSmithG1(dotWoH, dotWoShN, sp.roughness) * pgSmithG1(dotWiH, dotWiShN, sp.roughness);
        const float sc = F * G * D / (4.0f * dotWoShN);
        const float T1221 = (float) ((1.0 - (double) pgFresnelDielectric(dotWoShN, 1.5f)) * (1.0 - (double) pgFresnelDielectric(dotWiShN, 1.5f)));   // (double)
        for (int i = 0; i < 3; ++i) out[i] = diffuse[i] * T1221 + sp.specular[i] * sc;
        return;
    }
    if (sp.type == 'c') {
        float wh[3] = {wi[0] + sp.wo[0], wi[1] + sp.wo[1], wi[2] + sp.wo[2]};
        pgNormalize(wh);
        const float dotWhShN = pgDot(wh, sp.shN);
        const float D = pgDistrEval(dotWhShN, sp.roughness);
        if (D == 0.0f) return;
        const float dotWoH = pgDot(sp.wo, wh), dotWiH = pgDot(wi, wh);
        const float G = pgSmithG1(dotWoH, dotWoShN, sp.roughness) * pgSmithG1(dotWiH, dotWiShN, sp.roughness);
        const float model = D * G / (4.0f * dotWoShN);
        float F[3];
        pgFresnelConductor(F, dotWoH, sp.eta, sp.k);
        for (int i = 0; i < 3; ++i) out[i] = (F[i] * sp.specular[i]) * model;
    }
}

} // namespace ng
