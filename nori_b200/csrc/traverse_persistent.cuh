// traverse_persistent.cuh — persistent-warp traversal with dynamic ray fetch (the "while-while"
// scheme of Aila & Laine): a warp keeps its 32 lanes busy by refilling lanes whose ray has
// terminated from a global work counter as soon as enough of them are idle, instead of idling
// until the slowest ray of the warp is done (ncu on the one-ray-per-thread kernel: 6.4 of 32
// lanes active per issued instruction, profiles/r01_trace.md).
//
// Same node layout, slab and Moeller-Trumbore arithmetic and tie rule as traverse() in
// traverse.cuh, i.e. the result is the triangle the reference's brute-force loop reports
// (src/accel.cpp:42-55, src/mesh.cpp:51-88).
//
// IO concept:  bool load(uint32_t i, V3 &o, V3 &d, float &mint, float &maxt)   false: no ray at i
//              void store(uint32_t i, uint32_t gid, float t, float u, float v)
#pragma once
#include "traverse.cuh"

namespace ng {

#ifndef NORI_TRACE_SHARED_STACK
#define NORI_TRACE_SHARED_STACK 0
#endif
constexpr int kFetchThreshold = 20;   // default: refill when fewer than this many lanes still traverse
// Leave the inner-node loop for the postponed leaves as soon as at most this many lanes still descend (0: only when none does).
// Measured with binned rays (profiles/r02_notes.md): cbox closest-hit stage 80.4 -> 77.8 ms, any-hit 34.9 -> 30.9 at 512 spp.
constexpr int kNodeExit = 3;
constexpr int kTraceTuning = kFetchThreshold | (kNodeExit << 8);     // the `tuning` argument of tracePersistent

/// Grid of the quantised node layout (bvh.h: quantiseNodes): coordinate = base + m * ext, m in [1, 2).
struct QGrid { float bx, by, bz, ex, ey, ez; };

/// A 16-bit field of a quantised node (0x8000 | q) dropped into a float mantissa: 1 + q / 32768. One PRMT.
template <bool HIGH> NG_D float qfield(uint32_t w) { return __uint_as_float(__byte_perm(w, 0x3F000000u, HIGH ? 0x7324u : 0x7104u)); }

/// IO types whose load() turns the ray number into another handle for store() (the renderer's ClosestIO: the path slot the
/// sorted order list names) specialise this to true and take the handle as load()'s second argument.
template <class IO> struct IoIndirect { static constexpr bool value = false; };

/// QN = 2: `nodes` holds the 4-wide nodes of bvh.cu: wideNodes (128 bytes, four 256-bit loads per visit, up to three pushes).
/// QN = 1: `nodes` holds the 32-byte quantised nodes (two 16-byte loads per visit instead of four); the slab test is the
/// same fused multiply-add per plane with per-ray constants that fold the grid in: t = m * (ext / d) + (base - o) / d.
template <bool ANY_HIT, bool WT = false, int QN = 0, class IO>
NG_D void tracePersistent(const float4 *__restrict__ nodes, const float4 *__restrict__ leafTris, uint32_t nRays,
                          unsigned int *workCounter, IO &io, int tuning = kTraceTuning, QGrid grid = QGrid()) {
    const int fetchThreshold = tuning & 0xff;                  // refill when fewer lanes than this still hold a ray
    const int nodeExit = (tuning >> 8) & 0xff;                 // leave the node loop when fewer lanes than this still descend
    const unsigned lane = threadIdx.x & 31u;
    unsigned ltMask; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(ltMask));
    // Traversal stack: the first kSharedStack entries of every lane live in shared memory as [entry][thread], so a
    // push or pop is one conflict-free wavefront whatever the lanes' stack depths are (bank = thread % 32); in
    // local memory lanes at different depths touch different lines. Deeper entries (rare) overflow to local memory.
    int sp = 0, cur = 0, leaf = 0;                             // leaf: postponed leaf link (< 0) or 0
    constexpr int kSharedStack = NORI_TRACE_SHARED_STACK;
    __shared__ int sStack[kSharedStack > 0 ? kSharedStack : 1][128];
    constexpr int kStack = QN == 2 ? kWideStackSize : kStackSize;
    int stack[kStack - kSharedStack];
    const unsigned tix = threadIdx.x;
    auto push = [&](int v) {
        NG_CHECK(sp < kStack);
        if (kSharedStack > 0 && sp < kSharedStack) sStack[sp][tix] = v; else stack[sp - kSharedStack] = v;
        ++sp;
    };
    auto pop = [&]() -> int {
        --sp;
        return (kSharedStack > 0 && sp < kSharedStack) ? sStack[sp][tix] : stack[sp - kSharedStack];
    };
    constexpr int kDone = 0x7fffffff;                          // no node left for this lane
    cur = kDone;
    bool active = false;
    uint32_t rayIdx = 0xffffffffu, best = kNoHit;              // rayIdx: the handle io.load returned for io.store (path slot or ray number)
    V3 o = mk(0, 0, 0), d = mk(0, 0, 1);
    float idx = 0, idy = 0, idz = 0, oodx = 0, oody = 0, oodz = 0, mint = 0, tbest = 0, bu = 0, bv = 0;
    bool exhausted = false;
    WtRay wt = {0, 1, 2, 0.f, 0.f, 0.f};                      // NORI_ISECT_WATERTIGHT only (WT): per-ray shear

    for (;;) {
        // ---- refill idle lanes ------------------------------------------------------------
        if (!exhausted) {
            const unsigned idleMask = __ballot_sync(0xffffffffu, !active);
            if (idleMask) {
                unsigned base = 0;
                const int leader = __ffs(idleMask) - 1;
                if ((int) lane == leader) base = atomicAdd(workCounter, (unsigned) __popc(idleMask));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (base >= nRays) exhausted = true;
                else if (!active) {
                    const uint32_t i = base + __popc(idleMask & ltMask);
                    float maxt;
                    // (two copies of the same few lines: the register allocation of the 48-register kernels is at its limit, and a shared
                    // `handle` variable alone costs the any-hit kernel a 16-byte spill)
#define NG_START_RAY(HANDLE) do {                                                                                        \
                        const float big = 1e30f;                                                                         \
                        idx = fabsf(d.x) > 1e-30f ? 1.0f / d.x : copysignf(big, d.x);                                    \
                        idy = fabsf(d.y) > 1e-30f ? 1.0f / d.y : copysignf(big, d.y);                                    \
                        idz = fabsf(d.z) > 1e-30f ? 1.0f / d.z : copysignf(big, d.z);                                    \
                        if (QN == 1) {   /* plane at m: ((base + m ext) - o) / d = m * (ext / d) - (o - base) / d */          \
                            oodx = (o.x - grid.bx) * idx; oody = (o.y - grid.by) * idy; oodz = (o.z - grid.bz) * idz;    \
                            idx *= grid.ex; idy *= grid.ey; idz *= grid.ez;                                              \
                        } else { oodx = o.x * idx; oody = o.y * idy; oodz = o.z * idz; }                                 \
                        if (WT) wt = wtSetup(d);                                                                         \
                        rayIdx = (HANDLE); tbest = maxt; best = kNoHit; sp = 0; cur = 0; leaf = 0; active = true;        \
                    } while (0)
                    if constexpr (IoIndirect<IO>::value) {      // load() names the handle store() gets back (the path slot)
                        uint32_t handle;
                        if (i < nRays && io.load(i, handle, o, d, mint, maxt)) NG_START_RAY(handle);
                    } else {
                        if (i < nRays && io.load(i, o, d, mint, maxt)) NG_START_RAY(i);
                    }
#undef NG_START_RAY
                }
            }
        }
        if (!__any_sync(0xffffffffu, active)) {
            if (exhausted) break;
            continue;
        }
        // ---- traverse until too few lanes are left ------------------------------------------
        // Speculative while-while: a lane that reaches a leaf postpones it (`leaf`) and keeps
        // descending, so the node loop runs with (nearly) all lanes until every lane holds a leaf
        // or has emptied its stack; only then are the postponed leaves intersected. The slab test
        // of a speculating lane uses its not-yet-tightened tbest: conservative, never wrong.
        for (;;) {
            for (;;) {                                         // inner nodes (warp-uniform loop)
                if (QN == 2) {
                    if (cur >= 0 && cur != kDone) {
                        const float4 *n = nodes + 8 * (size_t) cur;
                        float4 lox, hix, loy, hiy, loz, hiz, lk, pad;
                        loadNode(n, lox, hix, loy, hiy);
                        loadNode(n + 4, loz, hiz, lk, pad);
                        // slab test of the four children (fused multiply-adds on inflated boxes, as the binary visit)
                        float key[4]; int link[4];
#define NG_WIDE_CHILD(C, M)                                                                                              \
                        {                                                                                                \
                            const float x0 = __fmaf_rn(lox.M, idx, -oodx), x1 = __fmaf_rn(hix.M, idx, -oodx);            \
                            const float y0 = __fmaf_rn(loy.M, idy, -oody), y1 = __fmaf_rn(hiy.M, idy, -oody);            \
                            const float z0 = __fmaf_rn(loz.M, idz, -oodz), z1 = __fmaf_rn(hiz.M, idz, -oodz);            \
                            const float tmn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), mint));    \
                            const float tmx = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tbest));   \
                            link[C] = __float_as_int(lk.M);                                                              \
                            key[C] = (tmn <= tmx && link[C] != kWideEmpty) ? tmn : INFINITY;                             \
                        }
                        NG_WIDE_CHILD(0, x) NG_WIDE_CHILD(1, y) NG_WIDE_CHILD(2, z) NG_WIDE_CHILD(3, w)
#undef NG_WIDE_CHILD
                        // nearest first: a five-comparator network on (entry distance, link); misses carry +inf and end up last
#define NG_WIDE_SWAP(A, B) if (key[B] < key[A]) { const float tk = key[A]; key[A] = key[B]; key[B] = tk; const int tl = link[A]; link[A] = link[B]; link[B] = tl; }
                        NG_WIDE_SWAP(0, 1) NG_WIDE_SWAP(2, 3) NG_WIDE_SWAP(0, 2) NG_WIDE_SWAP(1, 3) NG_WIDE_SWAP(1, 2)
#undef NG_WIDE_SWAP
                        if (key[3] < INFINITY) push(link[3]);
                        if (key[2] < INFINITY) push(link[2]);
                        if (key[1] < INFINITY) push(link[1]);
                        cur = key[0] < INFINITY ? link[0] : (sp > 0 ? pop() : kDone);
                        if (cur < 0 && leaf == 0) {            // first leaf: postpone, keep descending
                            leaf = cur;
                            cur = sp > 0 ? pop() : kDone;
                        }
                    }
                } else
                if (cur >= 0 && cur != kDone) {
                    float4 n0, n1, n2; int cl, cr;
                    if (QN == 1) {
                        const uint4 *qn = reinterpret_cast<const uint4 *>(nodes) + 2 * (size_t) cur;
                        const uint4 a = __ldg(qn), b = __ldg(qn + 1);
                        n0 = make_float4(qfield<false>(a.x), qfield<true>(a.x), qfield<false>(a.y), qfield<true>(a.y));
                        n1 = make_float4(qfield<false>(b.x), qfield<true>(b.x), qfield<false>(b.y), qfield<true>(b.y));
                        n2 = make_float4(qfield<false>(a.z), qfield<true>(a.z), qfield<false>(b.z), qfield<true>(b.z));
                        cl = (int) a.w; cr = (int) b.w;
                    } else {
                        const float4 *n = nodes + 4 * (size_t) cur;
                        float4 n3; loadNode(n, n0, n1, n2, n3);
                        cl = __float_as_int(n3.x); cr = __float_as_int(n3.y);
                    }
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
                    if (hl && hr) {
                        if (rmin < lmin) { int tmp = cl; cl = cr; cr = tmp; }
                        push(cr);
                        cur = cl;
                    } else if (hl) cur = cl;
                    else if (hr) cur = cr;
                    else cur = sp > 0 ? pop() : kDone;
                    if (cur < 0 && leaf == 0) {                // first leaf: postpone, keep descending
                        leaf = cur;
                        cur = sp > 0 ? pop() : kDone;
                    }
                }
                const bool searching = leaf == 0 && cur >= 0 && cur != kDone;
                const unsigned sm = __ballot_sync(0xffffffffu, searching);
                if (__popc(sm) <= nodeExit) break;
            }
            while (leaf < 0) {                                 // postponed leaves
                const uint32_t code = ~(uint32_t) leaf;
                const uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
                for (uint32_t k = 0; k < cnt; ++k) {
                    const float4 *tp = leafTris + kTriStride * (size_t) (first + k);
                    float4 a, b, c; loadTri(tp, a, b, c);
                    float u, v, t;
                    const bool hit = WT ? rayTriangleWatertight(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), o, wt, mint, tbest, u, v, t)
                                        : rayTriangle(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), o, d, mint, tbest, u, v, t);
                    if (hit) {
                        const uint32_t gid = __float_as_uint(a.w);
                        if (ANY_HIT) { best = gid; tbest = t; bu = u; bv = v; break; }
                        if (t < tbest || best == kNoHit || gid > best) { tbest = t; best = gid; bu = u; bv = v; }
                    }
                }
                leaf = 0;
                if (ANY_HIT && best != kNoHit) { cur = kDone; sp = 0; }
                else if (cur < 0) { leaf = cur; cur = sp > 0 ? pop() : kDone; }   // a second leaf was reached meanwhile
            }
            if (active && cur == kDone) {
                active = false;
                io.store(rayIdx, best, best == kNoHit ? 0.f : tbest, bu, bv);
            }
            const unsigned am = __ballot_sync(0xffffffffu, active);
            if (!am || (!exhausted && __popc(am) < fetchThreshold)) break;
        }
    }
}

} // namespace ng

