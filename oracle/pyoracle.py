"""ctypes bindings of the parity oracle (TEST INFRASTRUCTURE).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module;
the product package (nori_b200) never does.

* ``Port``  — oracle/liboracle.so, the standalone CPU restatement (builds anywhere).
* ``Ref``   — oracle/_ref/libnori_ref.so, the reference's own sources compiled by
  oracle/Makefile (present in the authoring container; travels prebuilt to GPU boxes).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from nori_b200.scene import FlatMesh, FlatScene

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libnori_ref.so")
REFERENCE_ROOT = "/root/reference"

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, np.float32)
    return a if shape is None else a.reshape(shape)


def build_port() -> str:
    subprocess.check_call(["make", "-s", "-C", HERE, "port"])
    return PORT_SO


def build_ref() -> str:
    if not os.path.isdir(REFERENCE_ROOT):
        raise RuntimeError("reference sources absent: oracle/_ref can only be built where /root/reference exists")
    subprocess.check_call(["make", "-s", "-j8", "-C", HERE, "ref"])
    return REF_SO


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def _optional(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class PortScene:
    """A FlatScene loaded into the standalone port."""

    def __init__(self, lib: "Port", scene: FlatScene, use_bvh: bool = True):
        self.lib, self.L, self.scene = lib, lib.L, scene
        L = self.L
        self.h = C.c_void_p(L.noracle_scene_create())
        for m in scene.meshes:
            b = L.noracle_scene_add_bsdf(self.h, int(m.bsdf_type), _f32(m.bsdf_params))
            mi = L.noracle_scene_add_mesh(self.h, m.V, m.V.shape[0], _optional(m.N), _optional(m.UV), m.F, m.F.shape[0], b)
            if m.is_emitter:
                L.noracle_scene_add_emitter(self.h, mi, _f32(m.radiance))
        L.noracle_scene_set_camera(self.h, _f32(scene.sample_to_camera).ravel(), _f32(scene.camera_to_world).ravel(),
                                   scene.width, scene.height, C.c_float(scene.near_clip), C.c_float(scene.far_clip))
        L.noracle_scene_set_filter_table(self.h, C.c_float(scene.filter_radius), _f32(scene.filter_table))
        L.noracle_scene_set_integrator(self.h, scene.integrator, scene.rr_start, C.c_float(scene.rr_max), scene.max_depth,
                                       _f32(scene.light_pos), _f32(scene.light_energy))
        L.noracle_scene_activate(self.h, int(use_bvh))

    def __del__(self):
        try:
            self.L.noracle_scene_destroy(self.h)
        except Exception:
            pass

    def set_integrator(self, integrator: int):
        s = self.scene
        self.L.noracle_scene_set_integrator(self.h, integrator, s.rr_start, C.c_float(s.rr_max), s.max_depth,
                                            _f32(s.light_pos), _f32(s.light_energy))

    def set_intersect_mode(self, mode: int):
        """0: Moeller-Trumbore (src/mesh.cpp:51-88), 1: watertight (noracle_math.h)."""
        self.L.noracle_scene_set_intersect_mode(self.h, int(mode))

    def camera_rays(self, pixel_samples):
        ps = _f32(pixel_samples, (-1, 2)); out = np.empty((ps.shape[0], 8), np.float32)
        self.L.noracle_camera_rays(self.h, ps.shape[0], ps, out)
        return out

    def trace(self, rays, shadow=False, use_bvh=True, count=False):
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        tri = np.empty(n, np.uint32); mesh = np.empty(n, np.uint32); tuv = np.empty((n, 3), np.float32)
        counters = np.zeros(3, np.uint64)
        self.L.noracle_scene_set_use_bvh(self.h, int(use_bvh))
        self.L.noracle_trace(self.h, rays, n, int(shadow), tri, mesh, tuv, counters.ctypes.data_as(C.c_void_p) if count else None)
        return (tri, mesh, tuv, counters) if count else (tri, mesh, tuv)

    def intersect(self, rays, use_bvh=True):
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        out = np.empty((n, 24), np.float32); valid = np.empty(n, np.int32)
        self.L.noracle_scene_set_use_bvh(self.h, int(use_bvh))
        self.L.noracle_intersect(self.h, rays, n, out, valid)
        return out, valid.astype(bool)

    def sample_radiance(self, pixel, sample):
        pixel = np.ascontiguousarray(pixel, np.uint32); sample = np.ascontiguousarray(sample, np.uint32)
        n = pixel.shape[0]
        rgb = np.empty((n, 3), np.float32); pos = np.empty((n, 2), np.float32); stats = np.zeros(3, np.uint64)
        self.L.noracle_sample_radiance(self.h, n, pixel, sample, rgb, pos, stats)
        return rgb, pos, stats

    def render_counted(self, mode=1, sample_first=0, sample_last=1, stride=1):
        """Traversal work of the oracle BVH over a render: ({rays, box, tri} closest, same for shadow)."""
        c = np.zeros(6, np.uint64); st = np.zeros(3, np.uint64)
        self.L.noracle_render_counted(self.h, mode, sample_first, sample_last, stride, c, st)
        return c[:3].astype(np.int64), c[3:].astype(np.int64)

    def render(self, mode=0, sample_first=0, sample_last=None, threads=0):
        if sample_last is None:
            sample_last = self.scene.sample_count
        out = np.empty(self.scene.block_shape(), np.float32); stats = np.zeros(4, np.uint64)
        self.L.noracle_render(self.h, mode, sample_first, sample_last, threads, out, stats)
        return out, stats


class Port:
    def __init__(self, build: bool = True):
        if build or not os.path.exists(PORT_SO):
            build_port()
        L = self.L = C.CDLL(PORT_SO)
        L.noracle_scene_create.restype = C.c_void_p
        L.noracle_scene_destroy.argtypes = [C.c_void_p]
        L.noracle_scene_add_bsdf.argtypes = [C.c_void_p, C.c_int, _f32p]
        L.noracle_scene_add_mesh.argtypes = [C.c_void_p, _f32p, C.c_uint32, C.c_void_p, C.c_void_p, _u32p, C.c_uint32, C.c_int]
        L.noracle_scene_add_emitter.argtypes = [C.c_void_p, C.c_int, _f32p]
        L.noracle_scene_set_camera.argtypes = [C.c_void_p, _f32p, _f32p, C.c_int, C.c_int, C.c_float, C.c_float]
        L.noracle_scene_set_filter.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float]
        L.noracle_scene_set_filter_table.argtypes = [C.c_void_p, C.c_float, _f32p]
        L.noracle_scene_set_integrator.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int, _f32p, _f32p]
        L.noracle_scene_activate.argtypes = [C.c_void_p, C.c_int]
        L.noracle_scene_set_use_bvh.argtypes = [C.c_void_p, C.c_int]
        L.noracle_scene_set_intersect_mode.argtypes = [C.c_void_p, C.c_int]
        L.noracle_pcg32_uints.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int, _u32p]
        L.noracle_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int, _f32p]
        L.noracle_camera_rays.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p]
        L.noracle_trace.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _u32p, _u32p, _f32p, C.c_void_p]
        L.noracle_intersect.argtypes = [C.c_void_p, _f32p, C.c_int, _f32p, _i32p]
        L.noracle_bsdf_sample.argtypes = [C.c_int, _f32p, C.c_int, _f32p, _f32p, _f32p, _f32p, _f32p, _f32p, _i32p]
        L.noracle_bsdf_eval_pdf.argtypes = [C.c_int, _f32p, C.c_int, _f32p, _f32p, _f32p, _f32p]
        L.noracle_warp.argtypes = [C.c_int, C.c_float, C.c_int, _f32p, _f32p, _f32p]
        L.noracle_fresnel.argtypes = [C.c_int, _f32p, C.c_float, C.c_float, _f32p]
        L.noracle_filter_table.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float, _f32p, _i32p]
        L.noracle_block_order.argtypes = [C.c_int, C.c_int, C.c_int, _i32p, C.c_int]
        L.noracle_splat.argtypes = [C.c_int, C.c_int, C.c_float, _f32p, C.c_int, _f32p, _f32p, _f32p]
        L.noracle_splat.restype = C.c_uint64
        L.noracle_sample_radiance.argtypes = [C.c_void_p, C.c_int, _u32p, _u32p, _f32p, _f32p, _u64p]
        L.noracle_render.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_int, _f32p, _u64p]
        L.noracle_render_counted.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_int, _u64p, _u64p]
        L.noracle_pathgraph_knn.argtypes = [_f32p, C.c_int, _i32p, _f32p, _f32p, C.c_int, C.c_int, _i32p, _i32p, _i32p, _f32p]
        L.noracle_pathgraph_scatter.argtypes = [C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, C.c_int, _f32p, _f32p, _f32p]
        L.noracle_pathgraph_eval.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p]
        L.noracle_pathgraph_mis.argtypes = [C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, C.c_int, _f32p, _f32p]
        L.noracle_pathgraph_clusters.argtypes = [C.c_int, _f32p, C.c_int, _i32p, _f32p, _f32p, _i32p, C.c_int, C.c_int, _i32p, _i32p, _i32p, C.POINTER(C.c_int)]
        L.noracle_pathgraph_clusters.restype = C.c_int
        L.noracle_pathgraph_cluster_scatter.argtypes = [C.c_int, C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, C.c_int, C.c_int, _f32p, _f32p]
        L.noracle_pathgraph_cluster.argtypes = [C.c_void_p, C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, C.c_int, C.c_int, _f32p, _f32p, _f32p]
        L.noracle_pathgraph_direct.argtypes = [C.c_void_p, C.c_void_p, C.c_int, _i32p, C.c_int, C.c_void_p, _i32p, C.c_int, C.c_int, _f32p, _f32p, C.c_void_p, C.c_void_p]

    def pathgraph_knn(self, pts, dim, cellsize, minb, k, dk):
        """noracle_pathgraph.h: hash grid + kNN. Returns (neighbors, knn, dknn, dists) rows nearest first, self swapped to the front."""
        pts = _f32(pts, (-1, 3)); n = pts.shape[0]; mk = max(k, dk)
        nb = np.empty((n, mk), np.int32); a = np.empty(n, np.int32); b = np.empty(n, np.int32); d = np.empty((n, mk), np.float32)
        self.L.noracle_pathgraph_knn(pts, n, np.asarray(dim, np.int32), _f32(cellsize), _f32(minb), k, dk, nb, a, b, d)
        return nb, a, b, d

    def pathgraph_scatter(self, sps, neighbors, knn, iterations, sequential=False):
        """noracle_pathgraph_scatter.h: (pdfmarginal, blur (it, n, 3), mc (it, n, 3)). sequential: the marginal sums in list order
        (the reference's single thread) instead of the product's lane order."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168
        nb = np.ascontiguousarray(neighbors, np.int32); kn = np.ascontiguousarray(knn, np.int32)
        pm = np.empty(n, np.float32); blur = np.empty((max(iterations, 1), n, 3), np.float32); mc = np.empty((max(iterations, 1), n, 3), np.float32)
        self.L.noracle_pathgraph_scatter(sps.ctypes.data_as(C.c_void_p), n, nb, nb.shape[1], kn, iterations, int(sequential), pm, blur.reshape(-1), mc.reshape(-1))
        return pm, blur[:iterations], mc[:iterations]

    def pathgraph_mis(self, sps, neighbors, knn, iterations, sequential=False):
        """computeMISAllOnGPURecord restated: (blur (it, n, 3), mc (it, n, 3))."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168
        nb = np.ascontiguousarray(neighbors, np.int32); kn = np.ascontiguousarray(knn, np.int32)
        blur = np.empty((max(iterations, 1), n, 3), np.float32); mc = np.empty((max(iterations, 1), n, 3), np.float32)
        self.L.noracle_pathgraph_mis(sps.ctypes.data_as(C.c_void_p), n, nb, nb.shape[1], kn, iterations, int(sequential), blur.reshape(-1), mc.reshape(-1))
        return blur[:iterations], mc[:iterations]

    def pathgraph_direct(self, sps, lps, neighbors, dknn, sequential=False, knn=None, iterations=None):
        """computeDirectScatterAllOnGPU restated: (marginal (n, 2), direct (n, 3)). lps: n 44-byte LPoint records. With knn and
        iterations also computeScatterAllOnGPURecordWithDirectOpt on the result: (marginal, direct, blur (it, n, 3), mc (it, n, 3))."""
        sps = np.ascontiguousarray(sps); lps = np.ascontiguousarray(lps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168 and lps.dtype.itemsize == 44 and lps.shape[0] == n
        nb = np.ascontiguousarray(neighbors, np.int32); dk = np.ascontiguousarray(dknn, np.int32)
        m = np.empty((n, 2), np.float32); out = np.empty((n, 3), np.float32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        if iterations is None:
            self.L.noracle_pathgraph_direct(vp(sps), vp(lps), n, nb, nb.shape[1], None, dk, -1, int(sequential), m.reshape(-1), out.reshape(-1), None, None)
            return m, out
        kn = np.ascontiguousarray(knn, np.int32)
        blur = np.empty((max(iterations, 1), n, 3), np.float32); mc = np.empty((max(iterations, 1), n, 3), np.float32)
        self.L.noracle_pathgraph_direct(vp(sps), vp(lps), n, nb, nb.shape[1], vp(kn), dk, iterations, int(sequential), m.reshape(-1), out.reshape(-1), vp(blur), vp(mc))
        return m, out, blur[:iterations], mc[:iterations]

    def pathgraph_build_clusters(self, pts, dim, cellsize, minb, seeds, K):
        """SPoints::BuildClusters restated with index-ordered lists: (cluster_of (n,), members, offsets)."""
        pts = _f32(pts, (-1, 3)); n = pts.shape[0]; seeds = np.ascontiguousarray(seeds, np.int32)
        g = np.empty(n, np.int32); lists = np.empty(n, np.int32); off = np.empty(n + len(seeds), np.int32); nl = C.c_int()
        nc = self.L.noracle_pathgraph_clusters(0, pts, n, np.asarray(dim, np.int32), _f32(cellsize), _f32(minb), seeds, len(seeds), K, g, lists, off, C.byref(nl))
        return g, lists[:nl.value].copy(), off[:nc].copy()

    def pathgraph_cluster_assign(self, pts, dim, cellsize, minb, seeds):
        """buildBatchClusters restated: index of the nearest seed (or -1) per point."""
        pts = _f32(pts, (-1, 3)); n = pts.shape[0]; seeds = np.ascontiguousarray(seeds, np.int32)
        g = np.empty(n, np.int32); d = np.zeros(1, np.int32); nl = C.c_int()
        self.L.noracle_pathgraph_clusters(1, pts, n, np.asarray(dim, np.int32), _f32(cellsize), _f32(minb), seeds, len(seeds), 2, g, d, d, C.byref(nl))
        return g

    def pathgraph_cluster_subdivide(self, pts, cluster_of, members, offsets, K):
        """countClusters + SubdivideClusters restated on a given clustering (lists in the given order): (new cluster_of, new count)."""
        pts = _f32(pts, (-1, 3)); n = pts.shape[0]
        g = np.ascontiguousarray(cluster_of, np.int32).copy(); lists = np.ascontiguousarray(members, np.int32); off = np.ascontiguousarray(offsets, np.int32)
        nl = C.c_int(len(lists)); z3 = np.zeros(3, np.float32)
        nc = self.L.noracle_pathgraph_clusters(2, pts, n, np.ones(3, np.int32), z3, z3, np.zeros(1, np.int32), len(off), K, g, lists, off, C.byref(nl))
        return g, nc

    def pathgraph_cluster(self, sps, lps, clusters, offsets, iterations, sequential=False):
        """SPoints::ClusterScatterWithDirectOptNR restated: (direct, blur, mc), each (n, 3). sps["groupIdx"] = cluster or -1."""
        sps = np.ascontiguousarray(sps); lps = np.ascontiguousarray(lps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168 and lps.dtype.itemsize == 44 and lps.shape[0] == n and iterations >= 1
        cl = np.ascontiguousarray(clusters, np.int32).reshape(-1); off = np.ascontiguousarray(offsets, np.int32).reshape(-1)
        cl_ = cl if len(cl) else np.zeros(1, np.int32); off_ = off if len(off) else np.zeros(1, np.int32)
        d = np.empty((n, 3), np.float32); b = np.empty((n, 3), np.float32); m = np.empty((n, 3), np.float32)
        self.L.noracle_pathgraph_cluster(sps.ctypes.data_as(C.c_void_p), lps.ctypes.data_as(C.c_void_p), n, cl_, len(cl), off_, len(off), iterations,
                                         int(sequential), d.reshape(-1), b.reshape(-1), m.reshape(-1))
        return d, b, m

    def pathgraph_cluster_scatter(self, sps, clusters, offsets, iterations, order="elements", sequential=False):
        """The cluster modes without direct light restated ("elements": ClusterScatter2, "scatter": ClusterScatter): (blur, mc), each (it, n, 3)."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168 and iterations >= 1
        cl = np.ascontiguousarray(clusters, np.int32).reshape(-1); off = np.ascontiguousarray(offsets, np.int32).reshape(-1)
        cl_ = cl if len(cl) else np.zeros(1, np.int32); off_ = off if len(off) else np.zeros(1, np.int32)
        b = np.empty((iterations, n, 3), np.float32); m = np.empty((iterations, n, 3), np.float32)
        self.L.noracle_pathgraph_cluster_scatter({"elements": 0, "scatter": 1}[order], sps.ctypes.data_as(C.c_void_p), n, cl_, len(cl), off_, len(off), iterations,
                                                 int(sequential), b.reshape(-1), m.reshape(-1))
        return b, m

    def pathgraph_eval(self, sps, wi):
        """pdf_device / bsdfeval_device restated: (pdf (n,), bsdf (n, 3)) of record i for direction wi[i]."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]; wi = _f32(wi, (-1, 3))
        pdf = np.empty(n, np.float32); f = np.empty((n, 3), np.float32)
        self.L.noracle_pathgraph_eval(sps.ctypes.data_as(C.c_void_p), n, wi, pdf, f.reshape(-1))
        return pdf, f

    def scene(self, scene: FlatScene, use_bvh=True) -> PortScene:
        return PortScene(self, scene, use_bvh)

    def pcg32_uints(self, initstate, initseq, n, advance=0):
        out = np.empty(n, np.uint32); self.L.noracle_pcg32_uints(initstate, initseq, advance, n, out); return out

    def pcg32_floats(self, initstate, initseq, n, advance=0):
        out = np.empty(n, np.float32); self.L.noracle_pcg32_floats(initstate, initseq, advance, n, out); return out

    def bsdf_sample(self, btype, params, wi, xi):
        wi = _f32(wi, (-1, 3)); xi = _f32(xi, (-1, 2)); n = wi.shape[0]
        wo = np.empty((n, 3), np.float32); w = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        eta = np.empty(n, np.float32); meas = np.empty(n, np.int32)
        self.L.noracle_bsdf_sample(btype, _f32(params), n, wi, xi, wo, w, pdf, eta, meas)
        return wo, w, pdf, eta, meas

    def bsdf_eval_pdf(self, btype, params, wi, wo):
        wi = _f32(wi, (-1, 3)); wo = _f32(wo, (-1, 3)); n = wi.shape[0]
        f = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        self.L.noracle_bsdf_eval_pdf(btype, _f32(params), n, wi, wo, f, pdf)
        return f, pdf

    def warp(self, wid, xi, param=0.0):
        xi = _f32(xi, (-1, 2)); n = xi.shape[0]
        out = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        self.L.noracle_warp(wid, param, n, xi, out, pdf)
        return out, pdf

    def fresnel(self, cos_i, ext_ior, int_ior):
        c = _f32(cos_i).ravel(); out = np.empty_like(c); self.L.noracle_fresnel(c.shape[0], c, ext_ior, int_ior, out); return out

    def filter_table(self, ftype, radius, p0=0.0, p1=0.0):
        t = np.empty(33, np.float32); b = np.zeros(1, np.int32)
        self.L.noracle_filter_table(ftype, radius, p0, p1, t, b)
        return t, int(b[0])

    def block_order(self, W, H, bs=32):
        cap = ((W + bs - 1) // bs) * ((H + bs - 1) // bs)
        out = np.empty((cap, 4), np.int32)
        n = self.L.noracle_block_order(W, H, bs, out, cap)
        return out[:n]

    def splat(self, W, H, radius, table, pos, rgb):
        pos = _f32(pos, (-1, 2)); rgb = _f32(rgb, (-1, 3))
        b = int(np.ceil(np.float32(radius) - np.float32(0.5)))
        out = np.empty((H + 2 * b, W + 2 * b, 4), np.float32)
        dropped = self.L.noracle_splat(W, H, radius, _f32(table), pos.shape[0], pos, rgb, out)
        return out, int(dropped)


REFCUDA_SO = os.path.join(HERE, "_ref", "libpbsdf_ref.so")


class RefCuda:
    """The reference's own CUDA tool (src/pbsdf.cu compiled unmodified, oracle/Makefile refcuda). Needs a GPU."""
    def __init__(self):
        if not os.path.exists(REFCUDA_SO):
            raise RuntimeError("oracle/_ref/libpbsdf_ref.so missing (run `make -C oracle refcuda` where /root/reference exists)")
        self.L = C.CDLL(REFCUDA_SO)
        self.L.refcuda_knn.argtypes = [_f32p, C.c_int, _i32p, _f32p, _f32p, C.c_int, C.c_int, _i32p, _i32p, _i32p, C.POINTER(C.c_double)]
        self.L.refcuda_scatter.argtypes = [C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, _f32p, _f32p, _f32p, C.POINTER(C.c_double)]
        self.L.refcuda_eval.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, _f32p]
        self.L.refcuda_mis.argtypes = [C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, _f32p, _f32p, C.POINTER(C.c_double)]
        if hasattr(self.L, 'refcuda_build_clusters'):
            self.L.refcuda_build_clusters.argtypes = [_f32p, C.c_int, _i32p, _f32p, _f32p, _i32p, C.c_int, C.c_int] + [_i32p] * 9 + [C.c_int, _i32p, C.POINTER(C.c_double)]
        if hasattr(self.L, 'refcuda_cluster_scatter'):
            self.L.refcuda_cluster_scatter.argtypes = [C.c_int, C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, C.c_int, _f32p, _f32p, C.POINTER(C.c_double)]
        if hasattr(self.L, 'refcuda_cluster'):
            self.L.refcuda_cluster.argtypes = [C.c_void_p, C.c_void_p, C.c_int, _i32p, C.c_int, _i32p, C.c_int, C.c_int, _f32p, _f32p, _f32p, C.POINTER(C.c_double)]
        if hasattr(self.L, 'refcuda_direct'):
            self.L.refcuda_direct.argtypes = [C.c_void_p, C.c_void_p, C.c_int, _i32p, C.c_int, C.c_void_p, _i32p, C.c_int, _f32p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]

    def knn(self, pts, dim, cellsize, minb, k, dk):
        """buildHashGridAndFindKNN. Returns (neighbors n x max(k,dk), knn, dknn, seconds)."""
        pts = _f32(pts, (-1, 3)); n = pts.shape[0]; mk = max(k, dk)
        nb = np.empty((n, mk), np.int32); a = np.empty(n, np.int32); b = np.empty(n, np.int32); sec = C.c_double()
        rc = self.L.refcuda_knn(pts, n, np.asarray(dim, np.int32), _f32(cellsize), _f32(minb), k, dk, nb, a, b, C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"refcuda_knn failed ({rc})")
        return nb, a, b, sec.value


    def scatter(self, sps, neighbors, knn, iterations):
        """computeScatterAllOnGPURecord + allGPUPdfMarginal: (pdfmarginal, blur (it, n, 3), mc (it, n, 3), seconds)."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168
        nb = np.ascontiguousarray(neighbors, np.int32); kn = np.ascontiguousarray(knn, np.int32)
        pm = np.empty(n, np.float32); blur = np.empty((max(iterations, 1), n, 3), np.float32); mc = np.empty((max(iterations, 1), n, 3), np.float32); sec = C.c_double()
        rc = self.L.refcuda_scatter(sps.ctypes.data_as(C.c_void_p), n, nb, nb.shape[1], kn, iterations, pm, blur.reshape(-1), mc.reshape(-1), C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"refcuda_scatter failed ({rc})")
        return pm, blur[:iterations], mc[:iterations], sec.value

    def mis(self, sps, neighbors, knn, iterations):
        """computeMISAllOnGPURecord of the compiled reference: (blur (it, n, 3), mc (it, n, 3), seconds)."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168
        nb = np.ascontiguousarray(neighbors, np.int32); kn = np.ascontiguousarray(knn, np.int32)
        blur = np.empty((max(iterations, 1), n, 3), np.float32); mc = np.empty((max(iterations, 1), n, 3), np.float32); sec = C.c_double()
        rc = self.L.refcuda_mis(sps.ctypes.data_as(C.c_void_p), n, nb, nb.shape[1], kn, iterations, blur.reshape(-1), mc.reshape(-1), C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"refcuda_mis failed ({rc})")
        return blur[:iterations], mc[:iterations], sec.value

    def direct(self, sps, lps, neighbors, dknn, knn=None, iterations=None):
        """computeDirectScatterAllOnGPU: (direct (n, 3), seconds); with knn and iterations followed by
        computeScatterAllOnGPURecordWithDirectOpt: (direct, blur (it, n, 3), mc (it, n, 3), seconds)."""
        sps = np.ascontiguousarray(sps); lps = np.ascontiguousarray(lps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168 and lps.dtype.itemsize == 44
        nb = np.ascontiguousarray(neighbors, np.int32); dk = np.ascontiguousarray(dknn, np.int32)
        out = np.empty((n, 3), np.float32); sec = C.c_double()
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        if iterations is None:
            rc = self.L.refcuda_direct(vp(sps), vp(lps), n, nb, nb.shape[1], None, dk, -1, out.reshape(-1), None, None, C.byref(sec))
            if rc != 0:
                raise RuntimeError(f"refcuda_direct failed ({rc})")
            return out, sec.value
        kn = np.ascontiguousarray(knn, np.int32)
        blur = np.empty((max(iterations, 1), n, 3), np.float32); mc = np.empty((max(iterations, 1), n, 3), np.float32)
        rc = self.L.refcuda_direct(vp(sps), vp(lps), n, nb, nb.shape[1], vp(kn), dk, iterations, out.reshape(-1), vp(blur), vp(mc), C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"refcuda_direct failed ({rc})")
        return out, blur[:iterations], mc[:iterations], sec.value

    def build_clusters(self, pts, dim, cellsize, minb, seeds, K):
        """SPoints::BuildClusters step by step with given seeds. Returns a dict: g1, lists1, off1 (after buildBatchClusters + countMyClusters),
        g2, lists2, off2 (after the first subdivision), g3, lists3, off3 (final), seconds."""
        pts = _f32(pts, (-1, 3)); n = pts.shape[0]; seeds = np.ascontiguousarray(seeds, np.int32); cap = n + len(seeds)
        a = {k: np.full(n, -1, np.int32) for k in ("g1", "lists1", "g2", "lists2", "g3", "lists3")}
        off1 = np.zeros(len(seeds), np.int32); off2 = np.zeros(cap, np.int32); off3 = np.zeros(cap, np.int32); counts = np.zeros(3, np.int32); sec = C.c_double()
        rc = self.L.refcuda_build_clusters(pts, n, np.asarray(dim, np.int32), _f32(cellsize), _f32(minb), seeds, len(seeds), K, a["g1"], a["lists1"], off1,
                                           a["g2"], a["lists2"], off2, a["g3"], a["lists3"], off3, cap, counts, C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"refcuda_build_clusters failed ({rc})")
        a["off1"] = off1; a["off2"] = off2[:counts[0]].copy(); a["off3"] = off3[:counts[1]].copy(); a["n_listed"] = int(counts[2]); a["seconds"] = sec.value
        a["lists3"] = a["lists3"][:counts[2]].copy()
        return a

    def cluster_scatter(self, sps, clusters, offsets, iterations, order="elements"):
        """SPoints::ClusterScatter2 ("elements") / ClusterScatter ("scatter") on given clusters: (blur, mc (it, n, 3), seconds)."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168
        cl = np.ascontiguousarray(clusters, np.int32).reshape(-1); off = np.ascontiguousarray(offsets, np.int32).reshape(-1)
        b = np.empty((iterations, n, 3), np.float32); m = np.empty((iterations, n, 3), np.float32); sec = C.c_double()
        rc = self.L.refcuda_cluster_scatter({"elements": 0, "scatter": 1}[order], sps.ctypes.data_as(C.c_void_p), n, cl, len(cl), off, len(off), iterations,
                                            b.reshape(-1), m.reshape(-1), C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"refcuda_cluster_scatter failed ({rc})")
        return b, m, sec.value

    def cluster(self, sps, lps, clusters, offsets, iterations):
        """SPoints::ClusterScatterWithDirectOptNR's drivers on given clusters: (direct, blur, mc, seconds)."""
        sps = np.ascontiguousarray(sps); lps = np.ascontiguousarray(lps); n = sps.shape[0]
        assert sps.dtype.itemsize == 168 and lps.dtype.itemsize == 44
        cl = np.ascontiguousarray(clusters, np.int32).reshape(-1); off = np.ascontiguousarray(offsets, np.int32).reshape(-1)
        d = np.empty((n, 3), np.float32); b = np.empty((n, 3), np.float32); m = np.empty((n, 3), np.float32); sec = C.c_double()
        rc = self.L.refcuda_cluster(sps.ctypes.data_as(C.c_void_p), lps.ctypes.data_as(C.c_void_p), n, cl, len(cl), off, len(off), iterations,
                                    d.reshape(-1), b.reshape(-1), m.reshape(-1), C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"refcuda_cluster failed ({rc})")
        return d, b, m, sec.value

    def eval(self, sps, wi):
        """pdf_device / bsdfeval_device of the compiled reference: (pdf (n,), bsdf (n, 3))."""
        sps = np.ascontiguousarray(sps); n = sps.shape[0]; wi = _f32(wi, (-1, 3))
        pdf = np.empty(n, np.float32); f = np.empty((n, 3), np.float32)
        rc = self.L.refcuda_eval(sps.ctypes.data_as(C.c_void_p), n, wi, pdf, f.reshape(-1))
        if rc != 0:
            raise RuntimeError(f"refcuda_eval failed ({rc})")
        return pdf, f


def have_refcuda() -> bool:
    return os.path.exists(REFCUDA_SO)


class RefScene:
    """A scene loaded by the reference's own parser (loadFromXML)."""

    def __init__(self, lib: "Ref", handle):
        self.lib, self.L, self.h = lib, lib.L, C.c_void_p(handle)

    def flatten(self) -> FlatScene:
        L = self.L
        meshes = []
        for i in range(L.ref_scene_mesh_count(self.h)):
            nV, nF = C.c_uint32(), C.c_uint32()
            hasN, hasUV, btype, isEm = C.c_int(), C.c_int(), C.c_int(), C.c_int()
            params = np.zeros(6, np.float32); rad = np.zeros(3, np.float32)
            if L.ref_scene_mesh_info(self.h, i, C.byref(nV), C.byref(nF), C.byref(hasN), C.byref(hasUV),
                                     C.byref(btype), params, C.byref(isEm), rad) != 0:
                raise RuntimeError(L.ref_last_error().decode())
            V = np.empty((nV.value, 3), np.float32); F = np.empty((nF.value, 3), np.uint32)
            N = np.empty((nV.value, 3), np.float32) if hasN.value else None
            UV = np.empty((nV.value, 2), np.float32) if hasUV.value else None
            L.ref_scene_mesh_copy(self.h, i, V.ctypes.data_as(C.c_void_p), _optional(N), _optional(UV), F.ctypes.data_as(C.c_void_p))
            meshes.append(FlatMesh(V=V, F=F, N=N, UV=UV, bsdf_type=btype.value, bsdf_params=params,
                                   is_emitter=bool(isEm.value), radiance=rad))
        s2c = np.zeros(16, np.float32); c2w = np.zeros(16, np.float32)
        W, H = C.c_int(), C.c_int(); near, far, fov = C.c_float(), C.c_float(), C.c_float()
        if L.ref_scene_camera(self.h, s2c, c2w, C.byref(W), C.byref(H), C.byref(near), C.byref(far), C.byref(fov)) != 0:
            raise RuntimeError("camera is not a PerspectiveCamera")
        radius = C.c_float(); table = np.zeros(33, np.float32); border = C.c_int()
        L.ref_scene_filter(self.h, C.byref(radius), table, C.byref(border))
        ip = np.zeros(3, np.int32); fp = np.zeros(7, np.float32)
        if L.ref_scene_integrator(self.h, ip, fp) != 0:
            raise RuntimeError("integrator is not an oracle integrator")
        return FlatScene(meshes=meshes, sample_to_camera=s2c.reshape(4, 4), camera_to_world=c2w.reshape(4, 4),
                         width=W.value, height=H.value, near_clip=near.value, far_clip=far.value, fov=fov.value,
                         filter_radius=radius.value, filter_table=table, integrator=int(ip[0]), rr_start=int(ip[1]),
                         max_depth=int(ip[2]), rr_max=float(fp[0]), light_pos=fp[1:4].copy(), light_energy=fp[4:7].copy(),
                         sample_count=L.ref_scene_sample_count(self.h))

    def camera_rays(self, pixel_samples):
        ps = _f32(pixel_samples, (-1, 2)); out = np.empty((ps.shape[0], 8), np.float32)
        self.L.ref_camera_rays(self.h, ps.shape[0], ps, out)
        return out

    def trace(self, rays, shadow=False, use_bvh=False):
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        tri = np.empty(n, np.uint32); mesh = np.empty(n, np.uint32); tuv = np.empty((n, 3), np.float32)
        self.L.ref_trace(self.h, rays, n, int(shadow), int(use_bvh), tri, mesh, tuv)
        return tri, mesh, tuv

    def intersect(self, rays, use_bvh=False):
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        out = np.empty((n, 24), np.float32); valid = np.empty(n, np.int32)
        self.L.ref_intersect(self.h, rays, n, int(use_bvh), out, valid)
        return out, valid.astype(bool)

    def stock_intersect(self, mesh_index, rays, shadow=False):
        """The UNMODIFIED reference Accel (single mesh, brute force) on mesh `mesh_index`."""
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        out = np.empty((n, 24), np.float32); valid = np.empty(n, np.int32)
        self.L.ref_stock_intersect(C.c_void_p(self.L.ref_scene_mesh_ptr(self.h, mesh_index)), rays, n, int(shadow), out, valid)
        return out, valid.astype(bool)

    def bsdf_sample(self, mesh_index, wi, xi):
        wi = _f32(wi, (-1, 3)); xi = _f32(xi, (-1, 2)); n = wi.shape[0]
        wo = np.empty((n, 3), np.float32); w = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        eta = np.empty(n, np.float32); meas = np.empty(n, np.int32)
        self.L.ref_bsdf_sample(self.h, mesh_index, n, wi, xi, wo, w, pdf, eta, meas)
        return wo, w, pdf, eta, meas

    def bsdf_eval_pdf(self, mesh_index, wi, wo):
        wi = _f32(wi, (-1, 3)); wo = _f32(wo, (-1, 3)); n = wi.shape[0]
        f = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        self.L.ref_bsdf_eval_pdf(self.h, mesh_index, n, wi, wo, f, pdf)
        return f, pdf

    def sample_radiance(self, pixel, sample, use_bvh=True):
        pixel = np.ascontiguousarray(pixel, np.uint32); sample = np.ascontiguousarray(sample, np.uint32)
        n = pixel.shape[0]
        rgb = np.empty((n, 3), np.float32); pos = np.empty((n, 2), np.float32)
        self.L.ref_sample_radiance(self.h, n, pixel, sample, int(use_bvh), rgb, pos)
        return rgb, pos

    def render(self, shape, mode=0, sample_first=0, sample_last=1, threads=0, use_bvh=True):
        out = np.empty(shape, np.float32)
        sec = self.L.ref_render(self.h, mode, sample_first, sample_last, threads, int(use_bvh), out)
        return out, sec

    def bridge_render(self, shape, libnori_b200: str, sample_first=0, sample_last=1, device=0):
        """tests/ref_bridge.cpp: this scene (objects of the reference's own parser) rendered by libnori_b200.so through the C ABI."""
        out = np.empty(shape, np.float32)
        self.L.ref_bridge_render.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_uint32, C.c_uint32, _f32p, C.c_void_p]
        self.L.ref_bridge_error.restype = C.c_char_p
        rc = self.L.ref_bridge_render(self.h, libnori_b200.encode(), device, sample_first, sample_last, out, None)
        if rc != 0:
            raise RuntimeError(f"ref_bridge_render failed ({rc}): {self.L.ref_bridge_error().decode()}")
        return out

    def threads_used(self) -> int:
        """Distinct TBB workers that rendered blocks in the last render() (0 with a library built before this existed)."""
        return int(self.L.ref_render_threads_used()) if hasattr(self.L, "ref_render_threads_used") else 0


class Ref:
    def __init__(self, build: bool = False):
        if build:
            build_ref()
        if not os.path.exists(REF_SO):
            raise RuntimeError("oracle/_ref/libnori_ref.so missing (run `make -C oracle ref` where /root/reference exists)")
        L = self.L = C.CDLL(REF_SO)
        L.ref_last_error.restype = C.c_char_p
        L.ref_load_xml.restype = C.c_void_p; L.ref_load_xml.argtypes = [C.c_char_p]
        L.ref_scene_mesh_ptr.restype = C.c_void_p; L.ref_scene_mesh_ptr.argtypes = [C.c_void_p, C.c_int]
        L.ref_scene_mesh_count.argtypes = [C.c_void_p]
        L.ref_scene_mesh_info.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [_f32p, C.c_void_p, _f32p]
        L.ref_scene_mesh_copy.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_scene_camera.argtypes = [C.c_void_p, _f32p, _f32p] + [C.c_void_p] * 5
        L.ref_scene_filter.argtypes = [C.c_void_p, C.c_void_p, _f32p, C.c_void_p]
        L.ref_scene_sample_count.argtypes = [C.c_void_p]
        L.ref_scene_integrator.argtypes = [C.c_void_p, _i32p, _f32p]
        L.ref_pcg32_uints.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int, _u32p]
        L.ref_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int, _f32p]
        L.ref_sampler_next2d.argtypes = [C.c_int, C.c_int, C.c_int, _f32p]
        L.ref_camera_rays.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p]
        L.ref_trace.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_int, _u32p, _u32p, _f32p]
        L.ref_intersect.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _f32p, _i32p]
        L.ref_stock_intersect.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _f32p, _i32p]
        L.ref_bsdf_sample.argtypes = [C.c_void_p, C.c_int, C.c_int, _f32p, _f32p, _f32p, _f32p, _f32p, _f32p, _i32p]
        L.ref_bsdf_eval_pdf.argtypes = [C.c_void_p, C.c_int, C.c_int, _f32p, _f32p, _f32p, _f32p]
        L.ref_fresnel.argtypes = [C.c_int, _f32p, C.c_float, C.c_float, _f32p]
        L.ref_frames.argtypes = [C.c_int, _f32p, _f32p]
        L.ref_filter_table.argtypes = [C.c_char_p, C.c_float, C.c_float, C.c_float, C.c_void_p, _f32p]
        L.ref_block_order.argtypes = [C.c_int, C.c_int, C.c_int, _i32p, C.c_int]
        L.ref_splat.argtypes = [C.c_int, C.c_int, C.c_char_p, C.c_int, _f32p, _f32p, _f32p]
        L.ref_render.restype = C.c_double
        L.ref_render.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_int, _f32p]
        L.ref_sample_radiance.argtypes = [C.c_void_p, C.c_int, _u32p, _u32p, C.c_int, _f32p, _f32p]
        L.ref_accel_counters.argtypes = [C.c_int, C.c_void_p, C.c_int]

    def load_xml(self, path: str) -> RefScene:
        """loadFromXML; statistical tests embedded in the file run during activate()."""
        h = self.L.ref_load_xml(path.encode())
        if not h:
            raise RuntimeError(self.L.ref_last_error().decode())
        return RefScene(self, h)

    def pcg32_uints(self, initstate, initseq, n, advance=0):
        out = np.empty(n, np.uint32); self.L.ref_pcg32_uints(initstate, initseq, advance, n, out); return out

    def pcg32_floats(self, initstate, initseq, n, advance=0):
        out = np.empty(n, np.float32); self.L.ref_pcg32_floats(initstate, initseq, advance, n, out); return out

    def sampler_next2d(self, ox, oy, n):
        out = np.empty((n, 2), np.float32); self.L.ref_sampler_next2d(ox, oy, n, out); return out

    def fresnel(self, cos_i, ext_ior, int_ior):
        c = _f32(cos_i).ravel(); out = np.empty_like(c); self.L.ref_fresnel(c.shape[0], c, ext_ior, int_ior, out); return out

    def frames(self, normals):
        nrm = _f32(normals, (-1, 3)); out = np.empty((nrm.shape[0], 6), np.float32)
        self.L.ref_frames(nrm.shape[0], nrm, out); return out

    def filter_table(self, name, radius=2.0, p0=0.5, p1=0.0):
        r = C.c_float(); t = np.empty(33, np.float32)
        if self.L.ref_filter_table(name.encode(), radius, p0, p1, C.byref(r), t) != 0:
            raise RuntimeError(self.L.ref_last_error().decode())
        return r.value, t

    def block_order(self, W, H, bs=32):
        cap = ((W + bs - 1) // bs) * ((H + bs - 1) // bs)
        out = np.empty((cap, 4), np.int32)
        n = self.L.ref_block_order(W, H, bs, out, cap)
        return out[:n]

    def splat(self, W, H, name, pos, rgb, border):
        pos = _f32(pos, (-1, 2)); rgb = _f32(rgb, (-1, 3))
        out = np.empty((H + 2 * border, W + 2 * border, 4), np.float32)
        if self.L.ref_splat(W, H, name.encode(), pos.shape[0], pos, rgb, out) != 0:
            raise RuntimeError(self.L.ref_last_error().decode())
        return out

    def traversal_counters(self, enable=True, reset=False):
        out = np.zeros(3, np.uint64)
        self.L.ref_accel_counters(int(enable), out.ctypes.data_as(C.c_void_p), int(reset))
        return out
