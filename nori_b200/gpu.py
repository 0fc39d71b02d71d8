"""ctypes binding of libnori_b200.so (include/nori_gpu.h) — the Python view of the C ABI.

The product path: there is no CPU fallback. Loading fails loudly when the library has not been
built (``python -m nori_b200.build``) and ``Engine()`` fails loudly when no CUDA device is present.

The method names mirror the reference operators they stand in for (Accel::rayIntersect ->
``trace``/``intersect``, PerspectiveCamera::sampleRay -> ``camera_rays``, BSDF::sample ->
``bsdf_sample``, ImageBlock::put -> ``splat``, render() -> ``render``), so the parity tests read
like calls into the reference.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .scene import FlatScene

HERE = os.path.dirname(os.path.abspath(__file__))
# NORI_GPU_LIB: another build of the SAME engine (the range-checked one, libnori_b200_checked.so); never a different implementation
LIB_PATH = os.path.join(HERE, os.path.basename(os.environ.get("NORI_GPU_LIB", "libnori_b200.so")))

EXPORTS = [
    "nori_gpu_create", "nori_gpu_destroy", "nori_gpu_last_error", "nori_gpu_device_name",
    "nori_gpu_upload_meshes", "nori_gpu_set_camera", "nori_gpu_set_filter", "nori_gpu_build_accel", "nori_gpu_set_accel_layout",
    "nori_gpu_render", "nori_gpu_download_block", "nori_gpu_download_bitmap", "nori_gpu_download_srgb8", "nori_gpu_upload_block", "nori_gpu_block_device_ptr",
    "nori_gpu_block_dims", "nori_gpu_reduce_blocks", "nori_gpu_device_count", "nori_gpu_set_profiling", "nori_gpu_set_pool_size", "nori_gpu_set_ray_binning", "nori_gpu_set_plain_bsdf_kernels", "nori_gpu_set_intersect_mode",
    "nori_gpu_trace", "nori_gpu_trace_device", "nori_gpu_trace_count", "nori_gpu_intersect", "nori_gpu_camera_rays",
    "nori_gpu_bsdf_sample", "nori_gpu_bsdf_eval_pdf", "nori_gpu_warp", "nori_gpu_warp_pdf", "nori_gpu_pcg32", "nori_gpu_splat",
    "nori_gpu_sample_radiance", "nori_gpu_trace_static",
    "nori_gpu_li", "nori_gpu_obj_dedup", "nori_gpu_path_vertices", "nori_gpu_path_vertices_lights", "nori_gpu_pathgraph_direct", "nori_gpu_pathgraph_scatter_direct", "nori_gpu_pathgraph_cluster", "nori_gpu_pathgraph_cluster_scatter", "nori_gpu_pathgraph_build_clusters", "nori_gpu_pathgraph_stride_seeds", "nori_gpu_pathgraph_knn", "nori_gpu_pathgraph_grid", "nori_gpu_pathgraph_scatter", "nori_gpu_pathgraph_mis", "nori_gpu_pathgraph_filter", "nori_gpu_pathgraph_render",
]


class NoriGpuError(RuntimeError):
    """Raised for every non-zero return of the C ABI (the C++ wrapper's NoriException)."""


class PathGraphStats(C.Structure):
    _fields_ = [("vertices", C.c_uint64), ("ms_vertices", C.c_float), ("ms_knn", C.c_float), ("ms_stage", C.c_float), ("dim", C.c_int32 * 3)]


class MeshDesc(C.Structure):
    _fields_ = [("V", C.c_void_p), ("N", C.c_void_p), ("UV", C.c_void_p), ("F", C.c_void_p),
                ("nV", C.c_uint32), ("nF", C.c_uint32), ("bsdf_type", C.c_int32), ("bsdf_params", C.c_float * 6),
                ("is_emitter", C.c_int32), ("radiance", C.c_float * 3)]


class CameraDesc(C.Structure):
    _fields_ = [("sample_to_camera", C.c_float * 16), ("camera_to_world", C.c_float * 16),
                ("width", C.c_int32), ("height", C.c_int32), ("near_clip", C.c_float), ("far_clip", C.c_float)]


class RenderDesc(C.Structure):
    _fields_ = [("integrator", C.c_int32), ("stream_mode", C.c_int32), ("sample_first", C.c_uint32), ("sample_last", C.c_uint32),
                ("rr_start", C.c_int32), ("rr_max", C.c_float), ("max_depth", C.c_int32), ("intersect_mode", C.c_int32),
                ("light_pos", C.c_float * 3), ("light_energy", C.c_float * 3), ("clear", C.c_int32), ("deterministic", C.c_int32)]


class RenderStats(C.Structure):
    _fields_ = [("camera_samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("segments", C.c_uint64),
                ("dropped", C.c_uint64), ("iterations", C.c_uint64), ("kernel_launches", C.c_uint64), ("ms_total", C.c_float),
                ("ms_trace_closest", C.c_float), ("ms_trace_shadow", C.c_float), ("ms_shade", C.c_float),
                ("ms_generate", C.c_float), ("ms_splat", C.c_float), ("late_samples", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class AccelStats(C.Structure):
    _fields_ = [("n_triangles", C.c_uint32), ("n_nodes", C.c_uint32), ("n_leaves", C.c_uint32), ("max_leaf_size", C.c_uint32),
                ("depth", C.c_uint32), ("sah_cost", C.c_float), ("build_ms", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class HashParams(C.Structure):
    """hparam (include/nori/shadingPoint.h:18-24)."""
    _fields_ = [("dim", C.c_int32 * 3), ("cellsize", C.c_float * 3), ("minb", C.c_float * 3), ("k", C.c_int32), ("dk", C.c_int32)]


# NoriGpuSPoint = the reference's SPoint (include/nori/shadingPoint.h:125-147), 168 bytes
SPOINT_DTYPE = np.dtype([("pos", "<f4", 3), ("wi", "<f4", 3), ("wi_d", "<f4", 3), ("wo", "<f4", 3), ("shN", "<f4", 3), ("geoN", "<f4", 3),
                         ("diffuse", "<f4", 3), ("specular", "<f4", 3), ("eLi", "<f4", 3), ("eLd", "<f4", 3), ("eta", "<f4", 3), ("k", "<f4", 3),
                         ("roughness", "<f4"), ("pdf", "<f4"), ("rrpdf", "<f4"), ("nidx", "<i4"), ("groupIdx", "<i4"), ("bsdf_type", "S1"), ("_pad", "V3")])
assert SPOINT_DTYPE.itemsize == 168
# NoriGpuLPoint = the reference's LPoint (include/nori/shadingPoint.h:148-155), 44 bytes
LPOINT_DTYPE = np.dtype([("L_directsample", "<f4", 3), ("L_bsdfsample", "<f4", 3), ("L_em", "<f4", 3), ("lightpdf", "<f4"), ("bsdfpdf", "<f4")])
assert LPOINT_DTYPE.itemsize == 44

_lib = None


def load_library() -> C.CDLL:
    """dlopen libnori_b200.so; never falls back to anything else."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NoriGpuError(f"{LIB_PATH} is missing: build it with `python -m nori_b200.build` (nvcc, sm_100a). "
                           "nori_b200 has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, ip = C.c_void_p, C.c_int
    L.nori_gpu_create.argtypes = [C.POINTER(vp), ip]
    L.nori_gpu_destroy.argtypes = [vp]; L.nori_gpu_destroy.restype = None
    L.nori_gpu_last_error.argtypes = [vp]; L.nori_gpu_last_error.restype = C.c_char_p
    L.nori_gpu_device_name.argtypes = [vp, C.c_char_p, ip]
    L.nori_gpu_upload_meshes.argtypes = [vp, C.POINTER(MeshDesc), ip]
    L.nori_gpu_set_camera.argtypes = [vp, C.POINTER(CameraDesc)]
    L.nori_gpu_set_filter.argtypes = [vp, C.c_float, vp]
    L.nori_gpu_build_accel.argtypes = [vp, C.POINTER(AccelStats)]
    L.nori_gpu_set_accel_layout.argtypes = [vp, ip]
    L.nori_gpu_render.argtypes = [vp, C.POINTER(RenderDesc), C.POINTER(RenderStats)]
    L.nori_gpu_download_block.argtypes = [vp, vp]
    L.nori_gpu_download_bitmap.argtypes = [vp, vp]
    L.nori_gpu_download_srgb8.argtypes = [vp, vp]
    L.nori_gpu_upload_block.argtypes = [vp, vp, ip]
    L.nori_gpu_block_device_ptr.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_uint64)]
    L.nori_gpu_block_dims.argtypes = [vp, C.POINTER(ip), C.POINTER(ip), C.POINTER(ip)]
    L.nori_gpu_set_profiling.argtypes = [vp, ip]
    L.nori_gpu_reduce_blocks.argtypes = [C.POINTER(vp), ip, ip]
    L.nori_gpu_device_count.argtypes = []
    L.nori_gpu_set_pool_size.argtypes = [vp, C.c_uint32]
    L.nori_gpu_set_ray_binning.argtypes = [vp, C.c_int]
    L.nori_gpu_set_plain_bsdf_kernels.argtypes = [vp, C.c_int]
    L.nori_gpu_set_intersect_mode.argtypes = [vp, ip]
    L.nori_gpu_trace.argtypes = [vp, vp, ip, ip, vp, vp, vp]
    L.nori_gpu_trace_static.argtypes = [vp, vp, ip, ip, vp, vp, vp]
    L.nori_gpu_trace_device.argtypes = [vp, vp, ip, ip, vp, C.POINTER(C.c_float)]
    L.nori_gpu_trace_count.argtypes = [vp, vp, ip, ip, vp]
    L.nori_gpu_intersect.argtypes = [vp, vp, ip, vp, vp]
    L.nori_gpu_camera_rays.argtypes = [vp, vp, ip, vp]
    L.nori_gpu_bsdf_sample.argtypes = [vp, ip, vp, ip, vp, vp, vp, vp, vp, vp, vp]
    L.nori_gpu_bsdf_eval_pdf.argtypes = [vp, ip, vp, ip, vp, vp, vp, vp]
    L.nori_gpu_warp.argtypes = [vp, ip, C.c_float, ip, vp, vp, vp]
    L.nori_gpu_warp_pdf.argtypes = [vp, ip, C.c_float, ip, vp, vp]
    L.nori_gpu_pcg32.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_int64, ip, vp]
    L.nori_gpu_splat.argtypes = [vp, ip, vp, vp, ip, vp, C.POINTER(C.c_uint64)]
    L.nori_gpu_sample_radiance.argtypes = [vp, C.POINTER(RenderDesc), ip, vp, vp, vp, vp]
    L.nori_gpu_obj_dedup.argtypes = [vp, vp, C.c_uint32, vp, C.c_uint32, vp, C.c_uint32, vp, C.c_uint32, vp, vp, vp, vp, C.POINTER(C.c_uint32), C.POINTER(C.c_float)]
    L.nori_gpu_li.argtypes = [vp, C.POINTER(RenderDesc), ip, vp, vp, vp]
    L.nori_gpu_path_vertices.argtypes = [vp, C.POINTER(RenderDesc), vp, C.c_uint64, C.POINTER(C.c_uint64)]
    L.nori_gpu_path_vertices_lights.argtypes = [vp, C.POINTER(RenderDesc), vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
    L.nori_gpu_pathgraph_direct.argtypes = [vp, vp, vp, ip, vp, ip, vp, vp, C.POINTER(C.c_float)]
    L.nori_gpu_pathgraph_build_clusters.argtypes = [vp, vp, C.c_uint64, ip, C.POINTER(HashParams), vp, ip, vp, vp, vp, ip, C.POINTER(ip), C.POINTER(ip), C.POINTER(C.c_float)]
    L.nori_gpu_pathgraph_cluster_scatter.argtypes = [vp, ip, vp, ip, vp, ip, vp, ip, ip, vp, vp, C.POINTER(C.c_float)]
    L.nori_gpu_pathgraph_stride_seeds.argtypes = [ip, ip, vp]
    L.nori_gpu_pathgraph_cluster.argtypes = [vp, vp, vp, ip, vp, ip, vp, ip, ip, vp, vp, vp, C.POINTER(C.c_float)]
    L.nori_gpu_pathgraph_scatter_direct.argtypes = [vp, vp, vp, ip, vp, ip, vp, vp, ip, vp, vp, vp, C.POINTER(C.c_float)]
    L.nori_gpu_pathgraph_knn.argtypes = [vp, vp, C.c_uint64, ip, C.POINTER(HashParams), vp, vp, vp, C.POINTER(C.c_float)]
    L.nori_gpu_pathgraph_grid.argtypes = [vp, C.c_uint64, ip, vp, ip, ip, C.POINTER(HashParams)]
    L.nori_gpu_pathgraph_scatter.argtypes = [vp, vp, ip, vp, ip, vp, ip, vp, vp, vp, C.POINTER(C.c_float)]
    L.nori_gpu_pathgraph_filter.argtypes = [vp, ip, vp, ip, C.POINTER(HashParams), ip, vp, vp, vp, vp]
    L.nori_gpu_pathgraph_render.argtypes = [vp, C.POINTER(RenderDesc), ip, ip, ip, vp, C.POINTER(PathGraphStats)]
    L.nori_gpu_pathgraph_mis.argtypes = [vp, vp, ip, vp, ip, vp, ip, vp, vp, C.POINTER(C.c_float)]
    _lib = L
    return L


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, np.float32)
    return a if shape is None else a.reshape(shape)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pathgraph_stride_seeds(n: int, n_seeds: int):
    """nori_gpu_pathgraph_stride_seeds: n_seeds distinct scattered vertex indices (what the cluster mode of pathgraph_render seeds with)."""
    out = np.empty(n_seeds, np.int32)
    if load_library().nori_gpu_pathgraph_stride_seeds(int(n), int(n_seeds), _p(out)) != 0:
        raise NoriGpuError("pathgraph_stride_seeds: need 0 < n_seeds <= n")
    return out


def pathgraph_grid(points, k: int, dk: int, dim=None) -> HashParams:
    """SPoints::BuildKNN's grid over the bounding box of the points; dim=None: PathGraph::computeDimensions (about n cells)."""
    pts = points if points.dtype == SPOINT_DTYPE else _f32(points, (-1, 3))
    pts = np.ascontiguousarray(pts)
    hp = HashParams()
    d = None if dim is None else (C.c_int32 * 3)(*[int(x) for x in dim])
    rc = load_library().nori_gpu_pathgraph_grid(_p(pts), pts.strides[0], pts.shape[0], d, int(k), int(dk), C.byref(hp))
    if rc != 0:
        raise NoriGpuError(f"[{rc}] pathgraph_grid: bad arguments")
    return hp


def device_count() -> int:
    return int(load_library().nori_gpu_device_count())


def reduce_blocks(engines, root: int = 0):
    """Sum the ImageBlocks of several engines (one per GPU, same process) into engines[root]: one ncclReduce."""
    L = load_library()
    hs = (C.c_void_p * len(engines))(*[e.h for e in engines])
    rc = L.nori_gpu_reduce_blocks(hs, len(engines), int(root))
    if rc != 0:
        raise NoriGpuError(f"[{rc}] {L.nori_gpu_last_error(engines[root].h).decode()}")


class Engine:
    """One nori_gpu_ctx on one CUDA device."""

    def __init__(self, device: int = 0):
        self.L = load_library()
        h = C.c_void_p()
        rc = self.L.nori_gpu_create(C.byref(h), device)
        if rc != 0:
            raise NoriGpuError(f"nori_gpu_create failed ({rc}): {self.L.nori_gpu_last_error(None).decode()}")
        self.h = h
        self.scene: FlatScene | None = None
        self.accel_stats: dict | None = None

    def close(self):
        if getattr(self, "h", None):
            self.L.nori_gpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise NoriGpuError(f"[{rc}] {self.L.nori_gpu_last_error(self.h).decode()}")

    @property
    def device_name(self) -> str:
        buf = C.create_string_buffer(256)
        self._ck(self.L.nori_gpu_device_name(self.h, buf, 256))
        return buf.value.decode()

    # ---- scene ---------------------------------------------------------------------------
    def load_scene(self, scene: FlatScene) -> dict:
        """Scene::activate: upload meshes/materials/emitters, camera, filter; build the accel."""
        descs = (MeshDesc * len(scene.meshes))()
        keep = []
        for d, m in zip(descs, scene.meshes):
            keep.append(m)
            d.V, d.F = _p(m.V), _p(m.F)
            d.N, d.UV = _p(m.N), _p(m.UV)
            d.nV, d.nF = m.V.shape[0], m.F.shape[0]
            d.bsdf_type = int(m.bsdf_type)
            d.bsdf_params = (C.c_float * 6)(*[float(x) for x in m.bsdf_params])
            d.is_emitter = int(m.is_emitter)
            d.radiance = (C.c_float * 3)(*[float(x) for x in m.radiance])
        self._ck(self.L.nori_gpu_upload_meshes(self.h, descs, len(scene.meshes)))
        self.scene = scene
        self.set_camera(scene.sample_to_camera, scene.camera_to_world, scene.width, scene.height, scene.near_clip, scene.far_clip)
        self.set_filter(scene.filter_radius, scene.filter_table)
        st = AccelStats()
        self._ck(self.L.nori_gpu_build_accel(self.h, C.byref(st)))
        self.accel_stats = st.as_dict()
        return self.accel_stats

    def set_camera(self, s2c, c2w, width, height, near_clip, far_clip):
        cam = CameraDesc()
        cam.sample_to_camera = (C.c_float * 16)(*_f32(s2c).ravel().tolist())
        cam.camera_to_world = (C.c_float * 16)(*_f32(c2w).ravel().tolist())
        cam.width, cam.height, cam.near_clip, cam.far_clip = int(width), int(height), float(near_clip), float(far_clip)
        self._ck(self.L.nori_gpu_set_camera(self.h, C.byref(cam)))
        self._wh = (int(width), int(height))

    def set_filter(self, radius, table33):
        t = _f32(table33, (33,))
        self._ck(self.L.nori_gpu_set_filter(self.h, C.c_float(float(radius)), _p(t)))

    def set_profiling(self, on: bool):
        self._ck(self.L.nori_gpu_set_profiling(self.h, int(on)))

    def set_pool_size(self, n: int):
        self._ck(self.L.nori_gpu_set_pool_size(self.h, int(n)))

    def set_ray_binning(self, on: bool):
        """Counting sort of the path slots by origin cell and direction signs between bounces (default on; never changes an image)."""
        self._ck(self.L.nori_gpu_set_ray_binning(self.h, int(bool(on))))

    def set_plain_bsdf_kernels(self, on: bool):
        """Scenes without microfacet surfaces use the shade kernel without the Beckmann code (default on; never changes a sample's radiance)."""
        self._ck(self.L.nori_gpu_set_plain_bsdf_kernels(self.h, int(bool(on))))

    def block_dims(self):
        r, c, b = C.c_int(), C.c_int(), C.c_int()
        self._ck(self.L.nori_gpu_block_dims(self.h, C.byref(r), C.byref(c), C.byref(b)))
        return r.value, c.value, b.value

    def block_device_ptr(self):
        p, n = C.c_void_p(), C.c_uint64()
        self._ck(self.L.nori_gpu_block_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    # ---- render --------------------------------------------------------------------------
    def render_desc(self, integrator=None, stream_mode=1, sample_first=0, sample_last=None, clear=True, intersect_mode=0) -> RenderDesc:
        s = self.scene
        d = RenderDesc()
        d.integrator = int(s.integrator if integrator is None else integrator)
        d.stream_mode = int(stream_mode)
        d.sample_first = int(sample_first)
        d.sample_last = int(s.sample_count if sample_last is None else sample_last)
        d.rr_start, d.rr_max, d.max_depth, d.intersect_mode = int(s.rr_start), float(s.rr_max), int(s.max_depth), int(intersect_mode)
        d.light_pos = (C.c_float * 3)(*[float(x) for x in s.light_pos])
        d.light_energy = (C.c_float * 3)(*[float(x) for x in s.light_energy])
        d.clear, d.deterministic = int(clear), 0
        return d

    def render(self, integrator=None, stream_mode=1, sample_first=0, sample_last=None, clear=True, download=True, intersect_mode=0):
        """render(): all samples [first, last) of every pixel. Returns (ImageBlock data, stats)."""
        d = self.render_desc(integrator, stream_mode, sample_first, sample_last, clear, intersect_mode)
        st = RenderStats()
        self._ck(self.L.nori_gpu_render(self.h, C.byref(d), C.byref(st)))
        return (self.download_block() if download else None), st.as_dict()

    def download_block(self):
        r, c, _ = self.block_dims()
        out = np.empty((r, c, 4), np.float32)
        self._ck(self.L.nori_gpu_download_block(self.h, _p(out)))
        return out

    def download_bitmap(self):
        w, h = self._wh
        out = np.empty((h, w, 3), np.float32)
        self._ck(self.L.nori_gpu_download_bitmap(self.h, _p(out)))
        return out

    def set_accel_layout(self, layout: int):
        """0: 64-byte float nodes, 1: 32-byte quantised nodes (NORI_ACCEL_*)."""
        self._ck(self.L.nori_gpu_set_accel_layout(self.h, int(layout)))

    def set_intersect_mode(self, mode: int):
        """Ray-triangle test of the raw operators: 0 = the reference's Moeller-Trumbore, 1 = watertight."""
        self._ck(self.L.nori_gpu_set_intersect_mode(self.h, int(mode)))

    def download_srgb8(self):
        """ImageBlock::toBitmap + Color3f::toSRGB + 8-bit quantisation on the device (Bitmap::savePNG's pixels)."""
        w, h = self._wh
        out = np.empty((h, w, 3), np.uint8)
        self._ck(self.L.nori_gpu_download_srgb8(self.h, _p(out)))
        return out

    def upload_block(self, rgbw, accumulate=False):
        """Restore (or add to) the accumulated ImageBlock: checkpoint / resume, merging partial renders."""
        a = _f32(rgbw)
        r, c, _ = self.block_dims()
        if a.size != r * c * 4:
            raise NoriGpuError(f"upload_block: expected {r}x{c}x4 floats")
        self._ck(self.L.nori_gpu_upload_block(self.h, _p(a), int(accumulate)))

    def sample_radiance(self, pixel, sample, integrator=None, intersect_mode=0):
        pixel = np.ascontiguousarray(pixel, np.uint32); sample = np.ascontiguousarray(sample, np.uint32)
        n = pixel.shape[0]
        rgb = np.empty((n, 3), np.float32); pos = np.empty((n, 2), np.float32)
        d = self.render_desc(integrator, intersect_mode=intersect_mode)
        self._ck(self.L.nori_gpu_sample_radiance(self.h, C.byref(d), n, _p(pixel), _p(sample), _p(rgb), _p(pos)))
        return rgb, pos

    def obj_dedup(self, corners, positions, texcoords=None, normals=None):
        """The OBJ loader's corner de-duplication (first-use order) + vertex gather. corners: (n, 3) uint32 (p, uv, n), 1-based.
        Returns (F (n/3, 3), V, N or None, UV or None, device ms)."""
        c = np.ascontiguousarray(corners, np.uint32).reshape(-1, 3); n = c.shape[0]
        P = _f32(positions, (-1, 3)); T = None if texcoords is None else _f32(texcoords, (-1, 2)); Nr = None if normals is None else _f32(normals, (-1, 3))
        F = np.empty(n, np.uint32); V = np.empty((n, 3), np.float32)
        N = None if Nr is None else np.empty((n, 3), np.float32); UV = None if T is None else np.empty((n, 2), np.float32)
        nv, ms = C.c_uint32(), C.c_float()
        self._ck(self.L.nori_gpu_obj_dedup(self.h, _p(c), n, _p(P), P.shape[0], _p(T), 0 if T is None else T.shape[0], _p(Nr), 0 if Nr is None else Nr.shape[0],
                                           _p(F), _p(V), _p(N), _p(UV), C.byref(nv), C.byref(ms)))
        k = nv.value
        return F.reshape(-1, 3), V[:k].copy(), None if N is None else N[:k].copy(), None if UV is None else UV[:k].copy(), ms.value

    def li(self, rays, rng_state, integrator=None, intersect_mode=0):
        """Integrator::Li for n rays, each with its sampler's (pcg32 state, inc). Returns (rgb n x 3, advanced states n x 2 uint64)."""
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        st = np.array(rng_state, np.uint64).reshape(n, 2).copy()
        rgb = np.empty((n, 3), np.float32)
        d = self.render_desc(integrator, intersect_mode=intersect_mode)
        self._ck(self.L.nori_gpu_li(self.h, C.byref(d), n, _p(rays), _p(st), _p(rgb)))
        return rgb, st

    # ---- path graph (src/pbsdf.cu) -------------------------------------------------------------
    def path_vertices(self, integrator=None, sample_first=0, sample_last=1, lights=False):
        """SPoint records of every surface vertex of the camera samples [first, last) of every pixel, path-contiguous.
        lights=True: (SPoint records, LPoint records) — the light record of every vertex beside it."""
        d = self.render_desc(integrator, 1, sample_first, sample_last)
        n = C.c_uint64()
        self._ck(self.L.nori_gpu_path_vertices(self.h, C.byref(d), None, 0, C.byref(n)))
        out = np.zeros(int(n.value), SPOINT_DTYPE)
        lps = np.zeros(int(n.value), LPOINT_DTYPE) if lights else None
        if n.value:
            if lights:
                self._ck(self.L.nori_gpu_path_vertices_lights(self.h, C.byref(d), _p(out), _p(lps), n.value, C.byref(n)))
            else:
                self._ck(self.L.nori_gpu_path_vertices(self.h, C.byref(d), _p(out), n.value, C.byref(n)))
        return (out, lps) if lights else out

    def pathgraph_direct(self, points, lights, neighbors, dknn, knn=None, iterations=None):
        """computeDirectScatterAllOnGPU: SPoint + LPoint records and the kNN rows (dk consumer's counts) -> (direct (n, 3), ms).
        With knn and iterations the chained driver (computeScatterRadianceAOGWithProcessRecordingAndDirectLightOPT): the indirect
        scatter over the k consumer's counts using the filtered direct light -> (direct, blur (it, n, 3), mc (it, n, 3), ms)."""
        assert points.dtype == SPOINT_DTYPE and lights.dtype == LPOINT_DTYPE
        pts = np.ascontiguousarray(points); lps = np.ascontiguousarray(lights); n = pts.shape[0]
        nb = np.ascontiguousarray(neighbors, np.int32); dk = np.ascontiguousarray(dknn, np.int32)
        assert lps.shape == (n,) and nb.ndim == 2 and nb.shape[0] == n and dk.shape == (n,)
        out = np.empty((n, 3), np.float32); ms = C.c_float()
        if iterations is None:
            self._ck(self.L.nori_gpu_pathgraph_direct(self.h, _p(pts), _p(lps), n, _p(nb), nb.shape[1], _p(dk), _p(out), C.byref(ms)))
            return out, ms.value
        kn = np.ascontiguousarray(knn, np.int32)
        assert kn.shape == (n,)
        blur = np.empty((iterations, n, 3), np.float32); mc = np.empty((iterations, n, 3), np.float32)
        self._ck(self.L.nori_gpu_pathgraph_scatter_direct(self.h, _p(pts), _p(lps), n, _p(nb), nb.shape[1], _p(kn), _p(dk), int(iterations),
                                                          _p(out), _p(blur), _p(mc), C.byref(ms)))
        return out, blur, mc, ms.value

    def pathgraph_build_clusters(self, points, hp: "HashParams", seeds):
        """SPoints::BuildClusters with given seeds (hp.k = members per cluster): (cluster_of (n,), members, offsets, ms)."""
        pts = points if points.dtype == SPOINT_DTYPE else _f32(points, (-1, 3))
        pts = np.ascontiguousarray(pts); n, stride = pts.shape[0], pts.strides[0]
        sd = np.ascontiguousarray(seeds, np.int32).reshape(-1); cap = 2 * (n // max(hp.k, 1)) + len(sd) + 4              # two splits add at most n / K clusters each
        of = np.empty(n, np.int32); lists = np.empty(n, np.int32); off = np.empty(cap, np.int32); nc = C.c_int(); nl = C.c_int(); ms = C.c_float()
        self._ck(self.L.nori_gpu_pathgraph_build_clusters(self.h, _p(pts), stride, n, C.byref(hp), _p(sd), len(sd), _p(of), _p(lists), _p(off), cap,
                                                          C.byref(nc), C.byref(nl), C.byref(ms)))
        return of, lists[:nl.value].copy(), off[:nc.value].copy(), ms.value

    def pathgraph_cluster(self, points, lights, clusters, offsets, iterations: int, cluster_of=None):
        """SPoints::ClusterScatterWithDirectOptNR (the pg tool's default mode on given clusters): members of cluster c =
        clusters[offsets[c]:offsets[c + 1]]; points["groupIdx"] = the vertex's cluster or -1 (cluster_of, if given, replaces it in a
        copy). Returns (direct (n, 3), blur (n, 3), mc (n, 3), ms)."""
        assert points.dtype == SPOINT_DTYPE and lights.dtype == LPOINT_DTYPE
        pts = np.ascontiguousarray(points); lps = np.ascontiguousarray(lights); n = pts.shape[0]
        if cluster_of is not None:
            pts = pts.copy(); pts["groupIdx"] = np.asarray(cluster_of, np.int32)
        cl = np.ascontiguousarray(clusters, np.int32).reshape(-1); off = np.ascontiguousarray(offsets, np.int32).reshape(-1)
        direct = np.empty((n, 3), np.float32); blur = np.empty((n, 3), np.float32); mc = np.empty((n, 3), np.float32); ms = C.c_float()
        self._ck(self.L.nori_gpu_pathgraph_cluster(self.h, _p(pts), _p(lps), n, _p(cl) if len(cl) else None, len(cl), _p(off) if len(off) else None, len(off),
                                                   int(iterations), _p(direct), _p(blur), _p(mc), C.byref(ms)))
        return direct, blur, mc, ms.value

    def pathgraph_cluster_scatter(self, points, clusters, offsets, iterations: int, order: str = "elements", cluster_of=None):
        """The pg tool's cluster modes without direct light: order "elements" = ClusterScatter2 (mode n), "scatter" = ClusterScatter /
        loadClusterScatter (modes t, l). Returns (blur (it, n, 3), mc (it, n, 3), ms)."""
        assert points.dtype == SPOINT_DTYPE
        pts = np.ascontiguousarray(points); n = pts.shape[0]
        if cluster_of is not None:
            pts = pts.copy(); pts["groupIdx"] = np.asarray(cluster_of, np.int32)
        cl = np.ascontiguousarray(clusters, np.int32).reshape(-1); off = np.ascontiguousarray(offsets, np.int32).reshape(-1)
        blur = np.empty((iterations, n, 3), np.float32); mc = np.empty((iterations, n, 3), np.float32); ms = C.c_float()
        self._ck(self.L.nori_gpu_pathgraph_cluster_scatter(self.h, {"elements": 0, "scatter": 1}[order], _p(pts), n, _p(cl) if len(cl) else None, len(cl),
                                                           _p(off) if len(off) else None, len(off), int(iterations), _p(blur), _p(mc), C.byref(ms)))
        return blur, mc, ms.value

    def pathgraph_knn(self, points, hp: "HashParams"):
        """buildHashGridAndFindKNN: points = (n, 3) float32 or an SPoint record array. Returns (neighbors, knn, dknn, ms)."""
        pts = points if points.dtype == SPOINT_DTYPE else _f32(points, (-1, 3))
        pts = np.ascontiguousarray(pts)
        n, stride = pts.shape[0], pts.strides[0]
        mk = max(hp.k, hp.dk)
        nb = np.empty((n, mk), np.int32); knn = np.empty(n, np.int32); dknn = np.empty(n, np.int32); ms = C.c_float()
        self._ck(self.L.nori_gpu_pathgraph_knn(self.h, _p(pts), stride, n, C.byref(hp), _p(nb), _p(knn), _p(dknn), C.byref(ms)))
        return nb, knn, dknn, ms.value

    def pathgraph_scatter(self, points, neighbors, knn, iterations: int):
        """computeScatterAllOnGPURecord: SPoint records + the kNN rows -> (pdfmarginal (n,), blur (it, n, 3), mc (it, n, 3), ms)."""
        assert points.dtype == SPOINT_DTYPE
        pts = np.ascontiguousarray(points); n = pts.shape[0]
        nb = np.ascontiguousarray(neighbors, np.int32); kn = np.ascontiguousarray(knn, np.int32)
        assert nb.ndim == 2 and nb.shape[0] == n and kn.shape == (n,)
        pm = np.empty(n, np.float32); blur = np.empty((iterations, n, 3), np.float32); mc = np.empty((iterations, n, 3), np.float32); ms = C.c_float()
        self._ck(self.L.nori_gpu_pathgraph_scatter(self.h, _p(pts), n, _p(nb), nb.shape[1], _p(kn), int(iterations), _p(pm), _p(blur), _p(mc), C.byref(ms)))
        return pm, blur, mc, ms.value

    def pathgraph_render(self, k: int, iterations: int, mode: str = "mis", integrator=None, sample_first=0, sample_last=1):
        """The pg tool end to end on the device: {"full", "indirect", "pt", "blur", "init"} images (H, W, 3) + stats; mode "cluster"
        (the tool's default: BuildClusters + ClusterScatterWithDirectOptNR) adds "direct"."""
        d = self.render_desc(integrator, 1, sample_first, sample_last)
        W, H = self._wh
        names = ("full", "indirect", "pt", "blur", "init") + (("direct",) if mode == "cluster" else ())
        img = np.empty((len(names), H, W, 3), np.float32); st = PathGraphStats()
        self._ck(self.L.nori_gpu_pathgraph_render(self.h, C.byref(d), {"scatter": 0, "mis": 1, "cluster": 2}[mode], int(k), int(iterations), _p(img), C.byref(st)))
        out = dict(zip(names, img))
        return out, {"vertices": int(st.vertices), "ms_vertices": st.ms_vertices, "ms_knn": st.ms_knn, "ms_stage": st.ms_stage, "dim": list(st.dim)}

    def pathgraph_mis(self, points, neighbors, knn, iterations: int):
        """computeMISAllOnGPURecord (the pg tool's knn mode): SPoint records + the kNN rows -> (blur (it, n, 3), mc (it, n, 3), ms)."""
        assert points.dtype == SPOINT_DTYPE
        pts = np.ascontiguousarray(points); n = pts.shape[0]
        nb = np.ascontiguousarray(neighbors, np.int32); kn = np.ascontiguousarray(knn, np.int32)
        assert nb.ndim == 2 and nb.shape[0] == n and kn.shape == (n,)
        blur = np.empty((iterations, n, 3), np.float32); mc = np.empty((iterations, n, 3), np.float32); ms = C.c_float()
        self._ck(self.L.nori_gpu_pathgraph_mis(self.h, _p(pts), n, _p(nb), nb.shape[1], _p(kn), int(iterations), _p(blur), _p(mc), C.byref(ms)))
        return blur, mc, ms.value

    def pathgraph_filter(self, points, hp: "HashParams", iterations: int, mode: str = "scatter"):
        """kNN + scatter ("scatter") or MIS gather ("mis") in one call, the rows staying on the device:
        (pdfmarginal (scatter mode; zeros otherwise), blur, mc, (ms_knn, ms_stage))."""
        assert points.dtype == SPOINT_DTYPE
        pts = np.ascontiguousarray(points); n = pts.shape[0]
        pm = np.zeros(n, np.float32); blur = np.empty((iterations, n, 3), np.float32); mc = np.empty((iterations, n, 3), np.float32)
        ms = (C.c_float * 2)()
        self._ck(self.L.nori_gpu_pathgraph_filter(self.h, {"scatter": 0, "mis": 1}[mode], _p(pts), n, C.byref(hp), int(iterations), _p(pm), _p(blur), _p(mc), ms))
        return pm, blur, mc, (ms[0], ms[1])

    # ---- operators -----------------------------------------------------------------------
    def camera_rays(self, pixel_samples):
        ps = _f32(pixel_samples, (-1, 2)); out = np.empty((ps.shape[0], 8), np.float32)
        self._ck(self.L.nori_gpu_camera_rays(self.h, _p(ps), ps.shape[0], _p(out)))
        return out

    def trace(self, rays, shadow=False):
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        tri = np.empty(n, np.uint32); mesh = np.empty(n, np.uint32); tuv = np.empty((n, 3), np.float32)
        self._ck(self.L.nori_gpu_trace(self.h, _p(rays), n, int(shadow), _p(tri), _p(mesh), _p(tuv)))
        return tri, mesh, tuv

    TRACE_STATIC_MAX_RAYS = 65536

    def trace_static(self, rays, shadow=False):
        """nori_gpu_trace without any allocation on the library side (static device scratch, <= 65 536 rays per call)."""
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        tri = np.empty(n, np.uint32); mesh = np.empty(n, np.uint32); tuv = np.empty((n, 3), np.float32)
        self._ck(self.L.nori_gpu_trace_static(self.h, _p(rays), n, int(shadow), _p(tri), _p(mesh), _p(tuv)))
        return tri, mesh, tuv

    def trace_count(self, rays, shadow=False):
        rays = _f32(rays, (-1, 8)); counts = np.zeros(2, np.uint64)
        self._ck(self.L.nori_gpu_trace_count(self.h, _p(rays), rays.shape[0], int(shadow), _p(counts)))
        return int(counts[0]), int(counts[1])

    def trace_device(self, d_rays_ptr: int, n: int, d_hits_ptr: int, shadow=False) -> float:
        ms = C.c_float()
        self._ck(self.L.nori_gpu_trace_device(self.h, C.c_void_p(d_rays_ptr), n, int(shadow), C.c_void_p(d_hits_ptr), C.byref(ms)))
        return ms.value

    def intersect(self, rays):
        rays = _f32(rays, (-1, 8)); n = rays.shape[0]
        out = np.empty((n, 24), np.float32); valid = np.empty(n, np.int32)
        self._ck(self.L.nori_gpu_intersect(self.h, _p(rays), n, _p(out), _p(valid)))
        return out, valid.astype(bool)

    def bsdf_sample(self, btype, params, wi, xi):
        wi = _f32(wi, (-1, 3)); xi = _f32(xi, (-1, 2)); n = wi.shape[0]; prm = _f32(params, (6,))
        wo = np.empty((n, 3), np.float32); w = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        eta = np.empty(n, np.float32); meas = np.empty(n, np.int32)
        self._ck(self.L.nori_gpu_bsdf_sample(self.h, int(btype), _p(prm), n, _p(wi), _p(xi), _p(wo), _p(w), _p(pdf), _p(eta), _p(meas)))
        return wo, w, pdf, eta, meas

    def bsdf_eval_pdf(self, btype, params, wi, wo):
        wi = _f32(wi, (-1, 3)); wo = _f32(wo, (-1, 3)); n = wi.shape[0]; prm = _f32(params, (6,))
        f = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        self._ck(self.L.nori_gpu_bsdf_eval_pdf(self.h, int(btype), _p(prm), n, _p(wi), _p(wo), _p(f), _p(pdf)))
        return f, pdf

    def warp(self, wid, xi, param=0.0):
        xi = _f32(xi, (-1, 2)); n = xi.shape[0]
        out = np.empty((n, 3), np.float32); pdf = np.empty(n, np.float32)
        self._ck(self.L.nori_gpu_warp(self.h, int(wid), C.c_float(param), n, _p(xi), _p(out), _p(pdf)))
        return out, pdf

    def warp_pdf(self, wid, points3, param=0.0):
        pts = _f32(points3, (-1, 3)); n = pts.shape[0]
        pdf = np.empty(n, np.float32)
        self._ck(self.L.nori_gpu_warp_pdf(self.h, int(wid), C.c_float(param), n, _p(pts), _p(pdf)))
        return pdf

    def pcg32_uints(self, initstate, initseq, n, advance=0):
        out = np.empty(n, np.uint32)
        self._ck(self.L.nori_gpu_pcg32(self.h, initstate, initseq, advance, n, _p(out)))
        return out

    def splat(self, pos, rgb, deterministic=False):
        pos = _f32(pos, (-1, 2)); rgb = _f32(rgb, (-1, 3))
        r, c, _ = self.block_dims()
        out = np.empty((r, c, 4), np.float32); dropped = C.c_uint64()
        self._ck(self.L.nori_gpu_splat(self.h, pos.shape[0], _p(pos), _p(rgb), int(deterministic), _p(out), C.byref(dropped)))
        return out, int(dropped.value)
