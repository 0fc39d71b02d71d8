/* nori_gpu.h — C ABI of libnori_b200.so, the B200 (sm_100a) execution engine behind Nori's
 * render path.
 *
 * The reference (joyDeng/nori) has no FFI: its plugins are C++ translation units that register
 * with NoriObjectFactory (include/nori/object.h:142-161) and are driven by render()
 * (src/main.cpp:69-151). This header is the boundary a maintainer binds instead: the host side
 * (XML parser, object factory, OBJ loader — unchanged) flattens its Scene into the PODs below and
 * ONE call, nori_gpu_render, replaces everything inside `tbb::parallel_for` (src/main.cpp:93-121).
 * Each entry point cites the reference interface it stands in for; INTEGRATION.md shows the
 * reference-side stub.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success and a negative
 * NORI_GPU_E* code otherwise; nori_gpu_last_error() returns the message (the C++ wrapper rethrows it
 * as NoriException, include/nori/common.h:147-152). Host arrays are borrowed for the duration of a
 * call. One ctx = one CUDA device; calls on one ctx are serialised by the caller. There is no CPU
 * fallback: without a usable CUDA device nori_gpu_create fails.
 */
#ifndef NORI_GPU_H
#define NORI_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nori_gpu_ctx nori_gpu_ctx;

enum {
    NORI_GPU_OK = 0,
    NORI_GPU_EINVAL = -1,     /* bad argument / call order */
    NORI_GPU_ECUDA = -2,      /* CUDA runtime error (message has the call site) */
    NORI_GPU_ENODEVICE = -3,  /* no CUDA device: the product has no CPU path */
    NORI_GPU_EUNSUPPORTED = -4
};

/* BSDF models: src/diffuse.cpp, src/mirror.cpp, src/dielectric.cpp, src/microfacet.cpp */
enum { NORI_BSDF_DIFFUSE = 0, NORI_BSDF_MIRROR = 1, NORI_BSDF_DIELECTRIC = 2, NORI_BSDF_MICROFACET = 3 };

/* Integrator::Li implementations (include/nori/integrator.h:54), selected by the XML type name */
enum {
    NORI_INT_NORMALS = 0, NORI_INT_PATH_MATS = 1, NORI_INT_PATH_EMS = 2, NORI_INT_PATH_MIS = 3,
    NORI_INT_WHITTED = 4, NORI_INT_PRIMARY_SHADOW = 5, NORI_INT_SIMPLE = 6, NORI_INT_AO = 7
};

/* Sampler stream modes (src/independent.cpp:48-53).
 * BLOCK: the reference's stream — one PCG32 per 32x32 block seeded (offset.x, offset.y), consumed
 *        serially; reproducible on the device only for integrators that draw a fixed number of
 *        values per camera sample (normals, simple), via pcg32::advance (ext/pcg32/pcg32.h:140-161).
 * PER_SAMPLE: seed(initstate = y*W + x, initseq = sample index) per camera sample; independent of
 *        any partitioning of the work over GPUs. */
enum { NORI_STREAM_BLOCK = 0, NORI_STREAM_PER_SAMPLE = 1 };

/* Ray-triangle test: REFERENCE_MT reproduces src/mesh.cpp:51-88 operation for operation (the parity mode).
 * WATERTIGHT: Woop/Benthin/Wald 2013 — a ray never slips between two triangles sharing an edge or a
 * vertex (Moeller-Trumbore can); not in the reference, specified by oracle/noracle_math.h. */
enum { NORI_ISECT_REFERENCE_MT = 0, NORI_ISECT_WATERTIGHT = 1 };

/* Node layout read by the traversal kernels of nori_gpu_render / nori_gpu_trace*:
 * FLOAT: 64-byte nodes, two child boxes in single precision (four 16-byte loads per visit).
 * QUANTISED: 32-byte nodes, child boxes on a 15-bit grid over the scene box, rounded outwards (two loads per visit;
 * half the node traffic for scenes whose tree does not fit the caches). Same tree, same triangle tests: same hits.
 * WIDE4: the same tree collapsed on the device into 4-wide nodes of 128 bytes (four child boxes in single precision, half as many
 * dependent node fetches per ray). Same leaves, same triangle tests: same closest hits; any-hit rays agree on occlusion.
 * Measured no faster than FLOAT on this hardware (profiles/r02_notes.md); FLOAT is the default. */
enum { NORI_ACCEL_FLOAT = 0, NORI_ACCEL_QUANTISED = 1, NORI_ACCEL_WIDE4 = 2 };

/* One Mesh (include/nori/mesh.h:172-175): column-major Eigen buffers == C arrays below. */
typedef struct {
    const float *V;        /* 3 x nV positions (world space; obj.cpp applies toWorld at load) */
    const float *N;        /* 3 x nV vertex normals, or NULL */
    const float *UV;       /* 2 x nV texture coordinates, or NULL */
    const uint32_t *F;     /* 3 x nF vertex indices */
    uint32_t nV, nF;
    int32_t bsdf_type;     /* NORI_BSDF_* (Mesh::getBSDF; default diffuse, mesh.cpp:35-41) */
    float bsdf_params[6];  /* albedo|kd rgb, alpha, intIOR, extIOR */
    int32_t is_emitter;    /* Mesh::isEmitter (mesh.h:140) : `area` emitter attached */
    float radiance[3];
} NoriGpuMeshDesc;

/* PerspectiveCamera after activate() (src/perspective.cpp:53-86); matrices row-major 4x4. */
typedef struct {
    float sample_to_camera[16];
    float camera_to_world[16];
    int32_t width, height;
    float near_clip, far_clip;
} NoriGpuCameraDesc;

typedef struct {
    int32_t integrator;      /* NORI_INT_* */
    int32_t stream_mode;     /* NORI_STREAM_* */
    uint32_t sample_first;   /* samples [first, last) of every pixel (Sampler::getSampleCount range) */
    uint32_t sample_last;
    int32_t rr_start;        /* Russian roulette from this depth (oracle: 3) */
    float rr_max;            /* continuation probability cap (oracle: 0.99) */
    int32_t max_depth;       /* <= 0: unbounded */
    int32_t intersect_mode;  /* NORI_ISECT_* */
    float light_pos[3];      /* `simple` integrator (scenes/pa3/ajax-simple.xml:9-10) */
    float light_energy[3];
    int32_t clear;           /* 1: zero the accumulated ImageBlock first (ImageBlock::clear) */
    int32_t deterministic;   /* 1: splat in the reference's sample order (bit-reproducible image) */
} NoriGpuRenderDesc;

typedef struct {
    uint64_t camera_samples;
    uint64_t closest_rays;
    uint64_t shadow_rays;
    uint64_t segments;        /* path vertices processed */
    uint64_t dropped;         /* invalid radiance samples (src/block.cpp:75-79) */
    uint64_t iterations;      /* wavefront iterations */
    uint64_t kernel_launches;
    float ms_total;           /* device time of the whole call (CUDA events) */
    float ms_trace_closest, ms_trace_shadow, ms_shade, ms_generate, ms_splat;  /* filled when profiling=1 */
    uint64_t late_samples;    /* samples splatted one by one after their pass had been closed */
} NoriGpuRenderStats;

typedef struct {
    uint32_t n_triangles, n_nodes, n_leaves, max_leaf_size, depth;
    float sah_cost;
    float build_ms;
} NoriGpuAccelStats;

/* ---- lifetime ------------------------------------------------------------------------------- */
int nori_gpu_create(nori_gpu_ctx **out, int device_id);
void nori_gpu_destroy(nori_gpu_ctx *ctx);
const char *nori_gpu_last_error(const nori_gpu_ctx *ctx);   /* ctx may be NULL: last create() error */
int nori_gpu_device_name(const nori_gpu_ctx *ctx, char *buf, int cap);

/* ---- scene upload: replaces Scene::addChild/activate (src/scene.cpp:39-96) ------------------- */
int nori_gpu_upload_meshes(nori_gpu_ctx *ctx, const NoriGpuMeshDesc *meshes, int n);   /* Accel::addMesh (accel.h:39) for all meshes, emitters incl. */
int nori_gpu_set_camera(nori_gpu_ctx *ctx, const NoriGpuCameraDesc *cam);              /* camera.h:56-64 */
int nori_gpu_set_filter(nori_gpu_ctx *ctx, float radius, const float table33[33]);     /* ImageBlock ctor table, src/block.cpp:31-39 */
int nori_gpu_build_accel(nori_gpu_ctx *ctx, NoriGpuAccelStats *out);                   /* Accel::build (accel.h:45) — device BVH */
int nori_gpu_set_accel_layout(nori_gpu_ctx *ctx, int layout);                          /* NORI_ACCEL_*; before or after build_accel */

/* ---- the hot path: replaces render()'s parallel_for (src/main.cpp:93-121) -------------------- */
int nori_gpu_render(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, NoriGpuRenderStats *out);
/* ImageBlock::data() of the full-frame result, RowMajor (H+2b) x (W+2b) x RGBW (src/block.cpp:48) */
int nori_gpu_download_block(nori_gpu_ctx *ctx, float *rgbw);
/* ImageBlock::toBitmap (src/block.cpp:57-63): H x W x RGB, rgb/w, border cropped */
int nori_gpu_download_bitmap(nori_gpu_ctx *ctx, float *rgb);
/* toBitmap + the 8-bit path of Bitmap::savePNG (src/bitmap.cpp:117-127): Color3f::toSRGB (src/common.cpp:178-192),
 * x 255, clamp, truncate — H x W x RGB8, tonemapped on the device (what the reference's preview shows, src/gui.cpp:64-78). */
int nori_gpu_download_srgb8(nori_gpu_ctx *ctx, uint8_t *rgb8);
/* Checkpoint / resume and merging of partial results: sets the accumulated ImageBlock (layout of
 * nori_gpu_download_block) or, with accumulate = 1, adds to it — ImageBlock::put(ImageBlock &), src/block.cpp:105-114.
 * A render continues on top of it with NoriGpuRenderDesc::clear = 0 and the next sample range. */
int nori_gpu_upload_block(nori_gpu_ctx *ctx, const float *rgbw, int accumulate);
/* Device-side access for the cross-GPU sum of ImageBlocks (`result.put(block)`, block.cpp:105-114):
 * the accumulator is a device array of block_floats() floats that NCCL reduces in place. */
int nori_gpu_block_device_ptr(nori_gpu_ctx *ctx, void **dptr, uint64_t *n_floats);
int nori_gpu_block_dims(const nori_gpu_ctx *ctx, int *rows, int *cols, int *border);
/* The same sum inside ONE process holding a context per GPU: the accumulators of ctxs[0..n) (same camera and
 * filter, one device each) are summed into ctxs[root] with one ncclReduce over NVLink (libnccl.so.2, resolved at
 * run time; n == 1 is a no-op). The other contexts keep their own partial block. */
int nori_gpu_reduce_blocks(nori_gpu_ctx *const *ctxs, int n, int root);
/* Number of CUDA devices visible to the library (0 without a driver). */
int nori_gpu_device_count(void);
int nori_gpu_set_profiling(nori_gpu_ctx *ctx, int on);   /* per-stage CUDA-event timing (serialises stages) */
/* Path slots of the wavefront (a multiple of 256). Without this call the library sizes the pool per render: a third of the render's
 * camera samples, at least 2 Mi and at most 8 Mi slots, and 16 Mi for renders of at least 3 * 2^26 (201 M) camera samples (long renders
 * gain from the larger wavefront, short ones lose to its ramp-up and tail). */
int nori_gpu_set_pool_size(nori_gpu_ctx *ctx, uint32_t n_paths);
/* Ray binning of the wavefront renderer (default on): after every shade pass the path slots are counting-sorted by
 * (origin cell of the scene box, two direction signs) and the next closest-hit pass walks them in that order, so that the
 * lanes of a warp traverse neighbouring rays. Changes which lane traces which ray, never a result: images are identical with
 * it on or off. No counterpart in the reference (its render loop, src/main.cpp:57-100, traces one ray at a time per thread).
 * Pools beyond 16 Mi slots render without it. */
int nori_gpu_set_ray_binning(nori_gpu_ctx *ctx, int on);
/* Scenes without a microfacet BSDF (src/microfacet.cpp) — the Cornell box among them — are shaded by an instance of the path integrators'
 * kernel that carries no Beckmann code (default on; the census is taken from the uploaded meshes). The branches left out would never have
 * run: the radiance of every camera sample is identical bit for bit with it on or off. */
int nori_gpu_set_plain_bsdf_kernels(nori_gpu_ctx *ctx, int on);
/* Mesh::rayIntersect variant (src/mesh.cpp:51-88) used by the raw operators below (trace, trace_device,
 * trace_count, intersect): NORI_ISECT_*. nori_gpu_render takes it from NoriGpuRenderDesc::intersect_mode. */
int nori_gpu_set_intersect_mode(nori_gpu_ctx *ctx, int mode);

/* ---- raw operators (parity and Mrays/s) ----------------------------------------------------- */
/* Accel::rayIntersect (src/accel.cpp:35-55). rays: n x 8 [o.xyz, mint, d.xyz, maxt] (HOST memory).
 * tri_id: mesh-local triangle index or 0xffffffff; mesh_id likewise; tuv: n x 3 (t,u,v) or NULL. */
int nori_gpu_trace(nori_gpu_ctx *ctx, const float *rays, int n, int shadow, uint32_t *tri_id, uint32_t *mesh_id, float *tuv);
/* NO COUNTERPART IN THE REFERENCE. nori_gpu_trace for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): same
 * arguments, same results bit for bit, but the call allocates nothing — no cudaMalloc / cudaFree, no allocating
 * library routine (the error paths alone format a message). All device scratch (staged rays, hit records, the traversal's work cursor, the
 * split output arrays) lives in one file-scope `static __device__` array of the library, sized for at most
 * NORI_GPU_TRACE_STATIC_MAX_RAYS rays per call (4.4 MB, held by every GPU that loads the library until exit);
 * more rays are rejected with NORI_GPU_EINVAL. The array is shared by every context on a device and keeps the last
 * call's contents (the work cursor is cleared by each call): calls that overlap on one GPU race and must be
 * serialised by the caller. Always reads the float node layout (same tree, same hits as the quantised one). */
#define NORI_GPU_TRACE_STATIC_MAX_RAYS 65536
int nori_gpu_trace_static(nori_gpu_ctx *ctx, const float *rays, int n, int shadow, uint32_t *tri_id, uint32_t *mesh_id, float *tuv);
/* Same on DEVICE buffers already resident (no copies): for kernel-only timing. ms = device time. */
int nori_gpu_trace_device(nori_gpu_ctx *ctx, const void *d_rays, int n, int shadow, void *d_hits /* n x 16 B: t,u,v,gid */, float *ms);
/* Work counters of the device BVH for n rays: counts[0] = node visits (two child boxes per visit),
 * counts[1] = ray-triangle tests (Mesh::rayIntersect calls). Source of the roofline's bytes per ray. */
int nori_gpu_trace_count(nori_gpu_ctx *ctx, const float *rays, int n, int shadow, uint64_t counts[2]);
/* Accel::rayIntersect post-processing (src/accel.cpp:57-108): n x 24 floats
 * [p(3) t uv(2) sh.s sh.t sh.n geo.s geo.t geo.n], valid[i] = hit. */
int nori_gpu_intersect(nori_gpu_ctx *ctx, const float *rays, int n, float *out24, int32_t *valid);
/* PerspectiveCamera::sampleRay (src/perspective.cpp:88-109): pixel samples n x 2 -> rays n x 8 */
int nori_gpu_camera_rays(nori_gpu_ctx *ctx, const float *pixel_samples, int n, float *rays);
/* BSDF::sample (include/nori/bsdf.h:71) for n (wi, xi) pairs; pdf = BSDF::pdf of the sampled
 * direction (0 for discrete); used by the chi2/t-test harness (src/chi2test.cpp, src/ttest.cpp). */
int nori_gpu_bsdf_sample(nori_gpu_ctx *ctx, int bsdf_type, const float params[6], int n, const float *wi, const float *xi,
                         float *wo, float *weight, float *pdf, float *eta, int32_t *measure);
/* BSDF::eval / BSDF::pdf (bsdf.h:82,99) in solid-angle measure */
int nori_gpu_bsdf_eval_pdf(nori_gpu_ctx *ctx, int bsdf_type, const float params[6], int n, const float *wi, const float *wo,
                           float *f, float *pdf);
/* Warp::squareTo* and *Pdf (include/nori/warp.h:30-69); ids in src/warptest.cpp:91-94 order:
 * 0 square 1 tent 2 disk 3 uniform_sphere 4 uniform_hemisphere 5 cosine_hemisphere 6 beckmann */
int nori_gpu_warp(nori_gpu_ctx *ctx, int warp_id, float param, int n, const float *xi, float *out3, float *pdf);
/* Warp::squareTo*Pdf (include/nori/warp.h:33-69) evaluated at n given points (n x 3; the 2D warps read x, y):
 * the density the chi^2 tests integrate over their cells (src/warptest.cpp:121-155). */
int nori_gpu_warp_pdf(nori_gpu_ctx *ctx, int warp_id, float param, int n, const float *points3, float *pdf);
/* pcg32 (ext/pcg32/pcg32.h:55-112,140-161): seed, advance, n x nextUInt */
int nori_gpu_pcg32(nori_gpu_ctx *ctx, uint64_t initstate, uint64_t initseq, int64_t advance, int n, uint32_t *out);
/* ImageBlock::put(pos, value) (src/block.cpp:74-103) for n samples into a zeroed full-frame block
 * of the current camera/filter; rgbw as nori_gpu_download_block; dropped = invalid samples. */
int nori_gpu_splat(nori_gpu_ctx *ctx, int n, const float *pos, const float *rgb, int deterministic, float *rgbw, uint64_t *dropped);
/* Radiance of individual camera samples (PER_SAMPLE stream): pixel index y*W+x and sample index
 * per entry -> rgb n x 3 and the pixelSample position n x 2 handed to ImageBlock::put. */
int nori_gpu_sample_radiance(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, int n, const uint32_t *pixel, const uint32_t *sample,
                             float *rgb, float *pos);

/* ---- the step before the path: OBJ corner de-duplication on the device (src/obj.cpp:32-48, 96-117, 136-161) ------------------- */
/* The loader's OBJVertex hash map and vertex gather for one mesh. corners: n_corners x 3 = the (p, uv, n) index triples of the `f`
 * lines in file order, 1-based as written (quads already split as obj.cpp:101-107 does; uv / n entries are ignored when the file has
 * no vt / vn lines); positions / texcoords / normals: the file's v / vt / vn arrays (toWorld already applied, obj.cpp:66-84).
 * Outputs, bit-equal to what the reference's loader builds: F (n_corners vertex ids, handed out in order of FIRST USE), V (3 floats
 * per unique corner), N, UV (NULL exactly when the file has no vn / vt); the caller provides room for n_corners vertices.
 * Corners naming a non-existent v / vt / vn are an error (the reference's .at() throws). ms (optional): device time of the sort,
 * scan and gather. All arrays in HOST memory. */
int nori_gpu_obj_dedup(nori_gpu_ctx *ctx, const uint32_t *corners, uint32_t n_corners, const float *positions, uint32_t n_positions,
                       const float *texcoords, uint32_t n_texcoords, const float *normals, uint32_t n_normals,
                       uint32_t *F, float *V, float *N, float *UV, uint32_t *n_vertices, float *ms);

/* Integrator::Li(const Scene *, Sampler *, const Ray3f &) (include/nori/integrator.h:54) for n rays [o.xyz, mint, d.xyz, maxt], each
 * drawing from its own Sampler: rng_state holds n (pcg32 state, pcg32 inc) pairs (ext/pcg32/pcg32.h:55-57), read before and written
 * back after the call, so that a host Sampler continues exactly where Li left it (src/ttest.cpp:168-175, src/main.cpp:57-59).
 * rgb: n x 3 radiance. Uses integrator, rr_*, max_depth, intersect_mode, light_* of the descriptor. */
int nori_gpu_li(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, int n, const float *rays, uint64_t *rng_state, float *rgb);

/* ---- path-graph post-process, first slice (the fork's own CUDA tool, src/pbsdf.cu) -------------------- */
/* One path vertex = the reference's SPoint (include/nori/shadingPoint.h:125-147), same field order, 168 bytes. The reference
 * only READS such records from .bin dumps written by another renderer (src/pathgraph.cpp:22); here the wavefront tracer's own
 * paths produce them (nori_gpu_path_vertices). Direction conventions as bsdfeval_device uses them (src/pbsdf.cu:464-468):
 * wo = unit vector towards the previous vertex (the camera side), wi = sampled continuation towards the next vertex. */
typedef struct {
    float pos[3];
    float wi[3];        /* sampled continuation direction (world); 0 when the path ends here */
    float wi_d[3];      /* direction of the light sample of next-event estimation (world); 0 if none was drawn */
    float wo[3];        /* towards the previous vertex (world) */
    float shN[3], geoN[3];
    float diffuse[3];   /* albedo (diffuse), kd (microfacet), 0 otherwise */
    float specular[3];  /* ks (microfacet), 1 (mirror, dielectric), 0 (diffuse) */
    float eLi[3], eLd[3];   /* eLd: what the path estimator credits to this vertex itself (emission seen from the previous vertex with its MIS
                             * weight + the light sample if unoccluded) as radiance leaving towards wo; eLi: the same for the whole path
                             * suffix from this vertex on (eLi of a path's first vertex = the camera sample's radiance) */
    float eta[3];       /* intIOR / extIOR (dielectric, microfacet), else 0 */
    float k[3];         /* conductor absorption: 0 (Nori has no conductor) */
    float roughness;    /* Beckmann alpha (microfacet), else 0 */
    float pdf;          /* solid-angle density of wi (0: discrete) */
    float rrpdf;        /* Russian-roulette survival probability applied after this vertex (1 before roulette starts, 0 if the path ends) */
    int32_t nidx;       /* record index of the next vertex of the same path (= own index + 1: pbsdf.cu:1408 reads Index + 1, :1749 compares nidx with an index), 0: the path ends here */
    int32_t groupIdx;   /* camera sample the vertex belongs to: pixel * spp + k */
    char bsdf_type;     /* 'd' diffuse, 'c' mirror (conductor, roughness 0), 't' dielectric, 'o' microfacet (opaque rough) */
} NoriGpuSPoint;

/* One light record per path vertex = the reference's LPoint (include/nori/shadingPoint.h:149-156), same field order, 44 bytes: what the
 * direct-light scatter of the tool reads beside the SPoint (src/pbsdf.cu:1651-1687, 1881-1897). As the tracer fills them:
 * L_directsample = radiance of the light sample drawn at this vertex (direction SPoint::wi_d) if its shadow ray is free, else 0;
 * lightpdf = that sample's solid-angle density; L_bsdfsample = radiance the continuation (SPoint::wi) finds on an emitter at the
 * next vertex, else 0; bsdfpdf = SPoint::pdf; L_em = radiance this vertex emits towards wo. */
typedef struct {
    float L_directsample[3];
    float L_bsdfsample[3];
    float L_em[3];
    float lightpdf;
    float bsdfpdf;
} NoriGpuLPoint;

/* hparam (include/nori/shadingPoint.h:18-24): uniform grid of dim cells of size cellsize from minb; k / dk neighbours for the
 * indirect / direct-light scatter. SPoints::BuildKNN (shadingPoint.h:776-806) derives it from the points' bounding box. */
typedef struct {
    int32_t dim[3];
    float cellsize[3];
    float minb[3];
    int32_t k, dk;
} NoriGpuHashParams;

/* Path vertices of the camera samples [sample_first, sample_last) of every pixel (PER_SAMPLE streams; the same paths
 * nori_gpu_render traces), in path order, the vertices of one path contiguous. out == NULL: only counts. Returns the number of
 * vertices in *n_vertices; if it exceeds `capacity` nothing is written and NORI_GPU_EINVAL is returned (call again with room). */
int nori_gpu_path_vertices(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, NoriGpuSPoint *out, uint64_t capacity, uint64_t *n_vertices);
/* The same, with the light record of every vertex beside it (lights: `capacity` records or NULL). */
int nori_gpu_path_vertices_lights(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, NoriGpuSPoint *out, NoriGpuLPoint *lights, uint64_t capacity,
                                  uint64_t *n_vertices);
/* buildHashGridAndFindKNN (src/pbsdf.cu:3173-3235): uniform hash grid over n points + max(k, dk) nearest neighbours of every
 * point inside the 2x2x2 cell block around it (batchNearestNeighbor, pbsdf.cu:1089-1178). points: n records `stride` bytes
 * apart whose first 12 bytes are the position (NoriGpuSPoint array: stride 168; packed float3: stride 12), HOST memory.
 * neighbors: n x max(k, dk) ints (row i = point i; nearest first by (distance, index), the point itself swapped to the front as
 * the reference does; unused entries -1). knn / dknn: neighbours available for the k / dk consumers (pbsdf.cu:1141-1164).
 * ms (optional): device time of grid build + search. max(k, dk) <= 200. A grid whose squared cell counts sum to more than 5e11 (nearly all
 * points in a few cells: every query would scan them all) is refused with NORI_GPU_EINVAL before the search starts. */
int nori_gpu_pathgraph_knn(nori_gpu_ctx *ctx, const void *points, uint64_t stride, int n, const NoriGpuHashParams *hp,
                           int32_t *neighbors, int32_t *knn, int32_t *dknn, float *ms);
/* SPoints::BuildKNN's grid (shadingPoint.h:794-801) for a cell count per axis: bounding box of the points, cellsize = extent / dim. */
int nori_gpu_pathgraph_grid(const void *points, uint64_t stride, int n, const int32_t dim[3], int k, int dk, NoriGpuHashParams *out);
/* computeScatterAllOnGPURecord (src/pbsdf.cu:2317-2368), the second slice of the tool: the marginal density of every vertex's
 * continuation direction over its neighbourhood (allGPUPdfMarginal, :1738-1756), then `iterations` passes of
 *   blur[it] = radiance scattered over the k-nearest-neighbour graph (allGPUScatterRadiance, :1398-1431: vertex i hands
 *              temp[i + 1] * bsdf_j(wi_i) / pdfmarginal[i] to each neighbour j);
 *   temp     = eLi, or blur[it] + eLd once it >= 1 (updateRadiance, :2185-2194, called with the loop index: passes 0 and 1
 *              therefore scatter the same temp = eLi, :2344-2356);
 *   mc[it]   = the vertex's own one-sample estimate temp[i + 1] * bsdf_i(wi_i) / (pdf_i(wi_i) * rrpdf_i) with that updated temp
 *              (lastRun, :1497-1526).
 * bsdf / pdf are the TOOL's evaluations from the record fields (bsdfeval_device / pdf_device, :464-628: clamped dot products,
 * its own microfacet and conductor models), not Nori's BSDF plugins. points: n records (HOST); neighbors: n rows of row_ints
 * ints, the first knn[i] of row i are read (what nori_gpu_pathgraph_knn returns); pdfmarginal: n floats or NULL; blur, mc:
 * iterations x n x 3 floats. The scatter sums are atomic, here as in the reference: equal up to the order of float additions.
 * ms: device time without the copies. */
int nori_gpu_pathgraph_scatter(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, int n, const int32_t *neighbors, int row_ints,
                               const int32_t *knn, int iterations, float *pdfmarginal, float *blur, float *mc, float *ms);
/* computeMISAllOnGPURecord (src/pbsdf.cu:2922-2965), the variant the `pg` tool runs in its "knn" mode (src/cluster.cpp:223 ->
 * SPoints::computeMISRadianceAOGWithProcessRecording, shadingPoint.h:504-510): a GATHER instead of a scatter. Once:
 *   pdfsum[i][a] = sum over i's neighbours b of pdf_b(wi_a) * rrpdf_b   for every neighbour a of i whose path continues
 *                  (allGPUPdfSum, :1600-1617; k^2 evaluations per vertex);
 * per pass: blur[it][i] = sum over a of bsdf_i(wi_a) * temp[a + 1] / pdfsum[i][a] (allGPUMISRadiance, :1346-1370), temp = eLi
 * before the first pass and blur[it] + eLd after each, mc[it] as above (lastRun). No atomics: results are deterministic. */
int nori_gpu_pathgraph_mis(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, int n, const int32_t *neighbors, int row_ints,
                           const int32_t *knn, int iterations, float *blur, float *mc, float *ms);
/* computeDirectScatterAllOnGPU (src/pbsdf.cu:2370-2395): the tool's filter for DIRECT light. Per vertex i the marginal densities
 * (k x lightpdf_i, sum_j pdf_j(wi_i): allGPUDirectPdfMarginal, :1881-1897), then every neighbour j of i receives
 * L_directsample_i * bsdf_j(wi_d_i) / (k lightpdf_i) + L_bsdfsample_i * bsdf_j(wi_i) / sum_j pdf_j(wi_i) (allGPUDirectScatterRadiance,
 * :1651-1687), and each vertex its own emission L_em (:813-818). dknn: the dk consumer's counts of nori_gpu_pathgraph_knn.
 * direct: n x 3 floats. Float atomics, here as in the reference: equal up to the order of the additions. */
int nori_gpu_pathgraph_direct(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, const NoriGpuLPoint *lights, int n, const int32_t *neighbors,
                              int row_ints, const int32_t *dknn, float *direct, float *ms);
/* SPoints::computeScatterRadianceAOGWithProcessRecordingAndDirectLightOPT (include/nori/shadingPoint.h:462-476): the direct-light filter
 * above (result in `direct`, the reference's m_result.blur_direct), then computeScatterAllOnGPURecordWithDirectOpt (src/pbsdf.cu:2488-2540):
 * the iterated scatter of nori_gpu_pathgraph_scatter with the FILTERED direct light in the place of the vertices' own eLd — after every
 * pass temp = direct + scattered radiance (updateWithOptDirectRadiance, :2196-2199). knn / dknn: the two consumers' counts of the rows. */
int nori_gpu_pathgraph_scatter_direct(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, const NoriGpuLPoint *lights, int n, const int32_t *neighbors,
                                      int row_ints, const int32_t *knn, const int32_t *dknn, int iterations, float *direct, float *blur, float *mc,
                                      float *ms);
/* SPoints::BuildClusters (include/nori/shadingPoint.h:637-760): the clustering the `pg` tool's default mode works on. `seeds` = the vertices
 * the clusters grow from (the reference takes the first n / K + 1 entries of an index list shuffled by std::random_shuffle under srand(1994),
 * :640-647; here they are an input), hp->k = K = the members a cluster should have, hp->dim / cellsize / minb = the grid (nori_gpu_pathgraph_grid).
 *   1. every vertex joins the nearest seed found through a hashed table (n_seeds buckets) of the seeds' grid cells, searched over the 2x2x2 cell
 *      block around the vertex, or stays outside (-1) if none lies within twice the cell diagonal (buildBatchClusters, src/pbsdf.cu:3555-3622);
 *   2. twice: every cluster of `size` members is split into size / K + 1 clusters around its first size / K + 1 listed members (countMyClusters +
 *      subClusters, :3236-3310);   3. member lists and offsets (FinalizeCluster, :3312-3332).
 * The reference lists a cluster's members in the order its atomics arrive; here they are listed in index order (stable sorts), one of the
 * orders it can produce, so the result is deterministic. cluster_of: n ints (SPoint::groupIdx as SaveClusters leaves it), clusters: n ints
 * (*n_listed used), offsets: offsets_capacity ints (*n_clusters used; NORI_GPU_EINVAL if more clusters arise — 2 n / K + n_seeds + 2 always fits:
 * a split of clusters with n members in all adds at most n / K of them).
 * points / stride as nori_gpu_pathgraph_knn. */
int nori_gpu_pathgraph_build_clusters(nori_gpu_ctx *ctx, const void *points, uint64_t stride, int n, const NoriGpuHashParams *hp, const int32_t *seeds,
                                      int n_seeds, int32_t *cluster_of, int32_t *clusters, int32_t *offsets, int offsets_capacity, int *n_clusters,
                                      int *n_listed, float *ms);
/* Seeds for nori_gpu_pathgraph_build_clusters when the caller has no shuffled list of its own: seeds[s] = (s * step) mod n with step the
 * smallest integer >= 0.618 n coprime to n — n_seeds distinct vertices, consecutive ones far apart in record order. Host-only, no context. */
int nori_gpu_pathgraph_stride_seeds(int n, int n_seeds, int32_t *seeds);
/* SPoints::ClusterScatterWithDirectOptNR (include/nori/shadingPoint.h:600-624), what the `pg` tool runs in its default "opt" mode once the
 * vertices are clustered (src/cluster.cpp:217): every vertex exchanges light with ALL members of its cluster instead of its k nearest
 * neighbours. clusters[offsets[c] .. offsets[c + 1]) (the last one up to n_listed) = the members of cluster c, a partition of the vertices
 * whose SPoint::groupIdx names a cluster (groupIdx = c for a member of c, -1 for a vertex outside every cluster: SaveClusters,
 * src/pbsdf.cu:987-1006) — the arrays SPoints::loadClusters uploads (shadingPoint.h:626-635). Anything else is refused (NORI_GPU_EINVAL), and so is
 * a cluster of more than 8192 members (the work is quadratic in a cluster's size; the tool's builder cuts clusters down to about K members).
 *   direct = ClusterDirect (src/pbsdf.cu:2400-2446): direct light filtered inside the clusters, each cluster scaled back to the light its
 *            members sampled (clampDirectCluster / computeRatio / updateComputeCluster) + emission       [m_result.blur_direct]
 *   blur   = the radiance after `iterations` passes of ClusterIterations3 (:3475-3532): per pass the cluster's dense block
 *            bsdf_j(wi_i) / pm_i applied to temp[i + 1] (MX), the same scaling, temp = direct + radiance   [blur_results[iterations - 1]]
 *   mc     = lastRun on the temp the last pass read                                                        [mc_results[iterations - 1]]
 * each n x 3 floats. The sums inside a cluster are taken in list order here (the reference: float atomics, any order). 1 <= iterations <= 64. */
int nori_gpu_pathgraph_cluster(nori_gpu_ctx *ctx, const NoriGpuSPoint *points, const NoriGpuLPoint *lights, int n, const int32_t *clusters,
                               int n_listed, const int32_t *offsets, int n_clusters, int iterations, float *direct, float *blur, float *mc, float *ms);
/* The tool's other cluster modes, without direct light (src/cluster.cpp:219-224), on the same cluster arrays:
 *   NORI_PG_CLUSTER_ELEMENTS  mode "n", SPoints::ClusterScatter2 (include/nori/shadingPoint.h:556-573): the blocks' elements bsdf_j(wi_i) / pm_i of
 *                             precomputedMatrixElemtns applied by ClusterIterations (src/pbsdf.cu:3359-3418);
 *   NORI_PG_CLUSTER_SCATTER   modes "t" / "l", SPoints::ClusterScatter / loadClusterScatter (:536-554): computeClusterScatterAllOnGPURecord /
 *                             ClusterScatterRecord (src/pbsdf.cu:3108-3172, 3030-3106) with allGPUClusterScatterRadiance (:1309-1345): (bsdf_j(wi_i) * temp) / pm_i,
 *                             nothing to the sender's own successor, nothing at all from a sender without a usable density.
 * Per pass: the cluster blocks, each cluster scaled back by its running in / out totals -> blur[it]; temp = radiance + eLd; lastRun -> mc[it].
 * blur / mc: iterations x n x 3 floats. */
#define NORI_PG_CLUSTER_ELEMENTS 0
#define NORI_PG_CLUSTER_SCATTER 1
int nori_gpu_pathgraph_cluster_scatter(nori_gpu_ctx *ctx, int order, const NoriGpuSPoint *points, int n, const int32_t *clusters, int n_listed,
                                       const int32_t *offsets, int n_clusters, int iterations, float *blur, float *mc, float *ms);
/* The tool's whole pipeline in one call, as SPoints::BuildKNN + one of the two drivers run it (include/nori/shadingPoint.h:776-806,
 * 496-510): hash grid + kNN over the records' positions, then the scatter (mode NORI_PG_SCATTER) or the MIS gather (NORI_PG_MIS)
 * on the rows the search left ON THE DEVICE (the k consumer's count knn[i] of each row, as the reference passes cu_knn). Same
 * results as nori_gpu_pathgraph_knn followed by nori_gpu_pathgraph_scatter / _mis without the round trip of the n x max(k, dk)
 * rows through host memory. pdfmarginal (optional) is written in scatter mode only. ms[0]: search, ms[1]: scatter / gather. */
#define NORI_PG_SCATTER 0
#define NORI_PG_MIS 1
#define NORI_PG_CLUSTER 2       /* nori_gpu_pathgraph_render only: BuildClusters + ClusterScatterWithDirectOptNR, the tool's default "opt" mode */
int nori_gpu_pathgraph_filter(nori_gpu_ctx *ctx, int mode, const NoriGpuSPoint *points, int n, const NoriGpuHashParams *hp, int iterations,
                              float *pdfmarginal, float *blur, float *mc, float ms[2]);
/* The `pg` tool end to end on the scene's own paths (src/cluster.cpp:156-253 in its "knn" mode, minus the file I/O), everything
 * on the device: the path vertices of the camera samples [sample_first, sample_last) of every pixel (nori_gpu_path_vertices; the
 * 168-byte records never leave the device), their bounding box and the grid PathGraph::computeDimensions derives from it, k
 * nearest neighbours, `iterations` passes of the scatter (NORI_PG_SCATTER) or the MIS gather (NORI_PG_MIS), and the images
 * writeFullLight / writeIndirectLight / writeFullinit compose from each path's first vertex (src/cluster.cpp:23-107).
 * images: 5 x height x width x 3 floats (HOST): full = eLd + mc, indirect = mc, pt = eLi - eLd, blur, init = eLd, all of the last
 * pass, averaged over a pixel's samples in sample order (the reference keeps the last path of a pixel).
 * NORI_PG_CLUSTER = the tool's default "opt" mode (src/cluster.cpp:204-218): light records beside the vertices (nori_gpu_path_vertices_lights),
 * BuildClusters with k members per cluster from n / k + 1 seeds (nori_gpu_pathgraph_stride_seeds), ClusterScatterWithDirectOptNR; `images` then
 * holds SIX images: the five above + direct = the filtered direct light writeDirectLight writes (:102-126). ms_knn = the cluster build. */
typedef struct {
    uint64_t vertices;
    float ms_vertices, ms_knn, ms_stage;    /* device time: records + bounding box, grid + search, scatter / gather + images */
    int32_t dim[3];                         /* grid the search used */
} NoriGpuPathGraphStats;
int nori_gpu_pathgraph_render(nori_gpu_ctx *ctx, const NoriGpuRenderDesc *desc, int mode, int k, int iterations, float *images,
                              NoriGpuPathGraphStats *stats);

#ifdef __cplusplus
}
#endif
#endif /* NORI_GPU_H */
