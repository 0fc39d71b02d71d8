// check.cuh — device-side range checks of the CHECKED build (`NORI_GPU_CHECKED=1 python -m
// nori_b200.build` -> libnori_b200_checked.so; select it with NORI_GPU_LIB). Every index into the path
// pool, the shadow queue, the sample records, the tile counters, the traversal stack and the frame is
// tested before use; violations are counted per translation unit and nori_gpu_render / the raw
// operators report them as NORI_GPU_ECUDA ("device range check failed at <file id>:<line>").
// In the normal build NG_CHECK expands to nothing.
#pragma once
#include <cuda_runtime.h>

namespace ng {
#ifdef NORI_GPU_CHECKED
static __device__ unsigned int g_check[2];     // [0] violations, [1] line of the first one
#define NG_CHECK(cond) do { if (!(cond)) { if (atomicAdd(&ng::g_check[0], 1u) == 0u) ng::g_check[1] = (unsigned) __LINE__; } } while (0)
/// Reads and clears this translation unit's counters (host).
static inline unsigned fetchChecks(unsigned *line) {
    unsigned h[2] = {0, 0}, z[2] = {0, 0};
    cudaMemcpyFromSymbol(h, g_check, sizeof h);
    cudaMemcpyToSymbol(g_check, z, sizeof z);
    *line = h[1];
    return h[0];
}
#else
#define NG_CHECK(cond) ((void) 0)
static inline unsigned fetchChecks(unsigned *line) { *line = 0; return 0; }
#endif
} // namespace ng
