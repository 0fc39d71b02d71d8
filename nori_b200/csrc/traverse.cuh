// traverse.cuh — Accel::rayIntersect on the device: stack traversal of the flattened BVH with
// the reference's Moeller-Trumbore test fused into the leaf loop.
//
// Replaces the brute-force loop of src/accel.cpp:42-55 (closest hit and the shadow-ray early out
// at :47-48). The result is the same triangle the loop would report: minimum t and, among equal
// t, the highest global triangle id (the loop keeps the LAST triangle satisfying `t <= maxt`,
// src/mesh.cpp:87 with maxt tightened at accel.cpp:49).
//
// Node layout (64 B, four 16-byte loads; bvh.cu writes it):
//   n0 = (L.min.x, L.max.x, L.min.y, L.max.y)      n1 = (R.min.x, R.max.x, R.min.y, R.max.y)
//   n2 = (L.min.z, L.max.z, R.min.z, R.max.z)      n3 = (bits(linkL), bits(linkR), -, -)
//   link >= 0: inner node index;  link < 0: leaf, ~link = (firstTriSlot << 3) | (count - 1)
// Child boxes are stored inflated by a scene-relative delta (bvh.cu), which makes the slab test
// below conservative with respect to the rounding of BOTH the slab arithmetic and the
// Moeller-Trumbore arithmetic: a box is never culled while one of its triangles would be
// reported by the reference test. Slabs use fused multiply-adds on purpose (conservative side
// only); the triangle test does not.
#pragma once
#include "scene.cuh"
#include "bvh.h"
#include "check.cuh"

namespace ng {

constexpr uint32_t kNoHit = 0xffffffffu;
constexpr int kStackSize = 64;

struct TraceStats { uint32_t nodes, tris; };

/// One 64-byte node as TWO 256-bit loads (sm_100 `ld.global.nc.v8.f32`, SASS LDG.E.256) instead of four 128-bit ones:
/// the traversal is bound by the L1/TEX pipe (ncu: 82-85 % on every ray type, DRAM 4-6 %), where every lane's load
/// of its own node costs a tag lookup per instruction; nodes are 64-byte aligned (cudaMalloc + 64 B stride).
#ifndef NORI_NODE_LOAD_128
NG_D void loadNode(const float4 *n, float4 &n0, float4 &n1, float4 &n2, float4 &n3) {
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(n0.x), "=f"(n0.y), "=f"(n0.z), "=f"(n0.w), "=f"(n1.x), "=f"(n1.y), "=f"(n1.z), "=f"(n1.w) : "l"(n));
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(n2.x), "=f"(n2.y), "=f"(n2.z), "=f"(n2.w), "=f"(n3.x), "=f"(n3.y), "=f"(n3.z), "=f"(n3.w) : "l"(n + 2));
}
#else
NG_D void loadNode(const float4 *n, float4 &n0, float4 &n1, float4 &n2, float4 &n3) { n0 = __ldg(n); n1 = __ldg(n + 1); n2 = __ldg(n + 2); n3 = __ldg(n + 3); }
#endif

NG_D void loadTri(const float4 *tp, float4 &a, float4 &b, float4 &c) {
    if (kTriStride == 4) { float4 pad; loadNode(tp, a, b, c, pad); }
    else { a = __ldg(tp); b = __ldg(tp + 1); c = __ldg(tp + 2); }
}

template <bool ANY_HIT, bool COUNT = false>
NG_D uint32_t traverse(const float4 *__restrict__ nodes, const float4 *__restrict__ leafTris,
                       V3 o, V3 d, float mint, float maxt, float &tOut, float &uOut, float &vOut,
                       TraceStats *stats = nullptr, bool watertight = false) {
    WtRay wt = {0, 1, 2, 0.f, 0.f, 0.f};
    if (watertight) wt = wtSetup(d);
    const float big = 1e30f;
    const float idx = fabsf(d.x) > 1e-30f ? 1.0f / d.x : copysignf(big, d.x);
    const float idy = fabsf(d.y) > 1e-30f ? 1.0f / d.y : copysignf(big, d.y);
    const float idz = fabsf(d.z) > 1e-30f ? 1.0f / d.z : copysignf(big, d.z);
    const float oodx = o.x * idx, oody = o.y * idy, oodz = o.z * idz;

    int stack[kStackSize];
    int sp = 0;
    int cur = 0;
    uint32_t best = kNoHit;
    float tbest = maxt;

    for (;;) {
        if (cur >= 0) {
            const float4 *n = nodes + 4 * (size_t) cur;
            float4 n0, n1, n2, n3; loadNode(n, n0, n1, n2, n3);
            if (COUNT) stats->nodes++;
            const float lx0 = __fmaf_rn(n0.x, idx, -oodx), lx1 = __fmaf_rn(n0.y, idx, -oodx);
            const float ly0 = __fmaf_rn(n0.z, idy, -oody), ly1 = __fmaf_rn(n0.w, idy, -oody);
            const float lz0 = __fmaf_rn(n2.x, idz, -oodz), lz1 = __fmaf_rn(n2.y, idz, -oodz);
            const float rx0 = __fmaf_rn(n1.x, idx, -oodx), rx1 = __fmaf_rn(n1.y, idx, -oodx);
            const float ry0 = __fmaf_rn(n1.z, idy, -oody), ry1 = __fmaf_rn(n1.w, idy, -oody);
            const float rz0 = __fmaf_rn(n2.z, idz, -oodz), rz1 = __fmaf_rn(n2.w, idz, -oodz);
            const float lmin = fmaxf(fmaxf(fminf(lx0, lx1), fminf(ly0, ly1)), fmaxf(fminf(lz0, lz1), mint));
            const float lmax = fminf(fminf(fmaxf(lx0, lx1), fmaxf(ly0, ly1)), fminf(fmaxf(lz0, lz1), tbest));
            const float rmin = fmaxf(fmaxf(fminf(rx0, rx1), fminf(ry0, ry1)), fmaxf(fminf(rz0, rz1), mint));
            const float rmax = fminf(fminf(fmaxf(rx0, rx1), fmaxf(ry0, ry1)), fminf(fmaxf(rz0, rz1), tbest));
            const bool hl = lmin <= lmax, hr = rmin <= rmax;
            int cl = __float_as_int(n3.x), cr = __float_as_int(n3.y);
            if (hl && hr) {
                if (rmin < lmin) { int tmp = cl; cl = cr; cr = tmp; }
                NG_CHECK(sp < kStackSize);
                stack[sp++] = cr;
                cur = cl;
            } else if (hl) cur = cl;
            else if (hr) cur = cr;
            else {
                if (sp == 0) break;
                cur = stack[--sp];
            }
        } else {
            const uint32_t code = ~(uint32_t) cur;
            const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
            for (uint32_t i = 0; i < cnt; ++i) {
                const float4 *tp = leafTris + kTriStride * (size_t) (first + i);
                float4 a, b, c; loadTri(tp, a, b, c);
                if (COUNT) stats->tris++;
                float u, v, t;
                const bool hit = watertight ? rayTriangleWatertight(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), o, wt, mint, tbest, u, v, t)
                                            : rayTriangle(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), o, d, mint, tbest, u, v, t);
                if (hit) {
                    const uint32_t gid = __float_as_uint(a.w);
                    if (ANY_HIT) { tOut = t; uOut = u; vOut = v; return gid; }
                    if (t < tbest || best == kNoHit || gid > best) {
                        tbest = t; best = gid; tOut = t; uOut = u; vOut = v;
                    }
                }
            }
            if (sp == 0) break;
            cur = stack[--sp];
        }
    }
    return best;
}

} // namespace ng
