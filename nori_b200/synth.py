"""Procedural stand-in scenes for the configurations whose meshes the reference tree lacks
(SURVEY.md F6, 8d): C1 an Ajax-class closed mesh (`scenes/pa2/ajax-normals.xml` references an
`ajax.obj` that is absent), C4 a ~10 M-triangle tessellated height field. All randomness comes from
a seeded numpy generator, so the meshes are reproducible everywhere."""
from __future__ import annotations

import numpy as np

from .scene import FlatMesh, FlatScene, BSDF_DIFFUSE


def _value_noise(u, v, rng, octaves=3, base=8):
    out = np.zeros_like(u, dtype=np.float64)
    amp = 1.0
    for o in range(octaves):
        n = base << o
        g = rng.random((n + 1, n + 1))
        g[:, -1] = g[:, 0]                       # periodic in u (closed sphere seam)
        x, y = u * n, v * n
        x0, y0 = np.minimum(x.astype(int), n - 1), np.minimum(y.astype(int), n - 1)
        fx, fy = x - x0, y - y0
        fx, fy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
        val = (g[y0, x0] * (1 - fx) + g[y0, x0 + 1] * fx) * (1 - fy) + (g[y0 + 1, x0] * (1 - fx) + g[y0 + 1, x0 + 1] * fx) * fy
        out += amp * (val - 0.5)
        amp *= 0.5
    return out


def grid_faces(nu, nv, wrap_u=False):
    """Two triangles per quad of an (nv+1) x (nu+1 | nu) vertex grid."""
    cols = nu if wrap_u else nu + 1
    j, i = np.meshgrid(np.arange(nv), np.arange(nu), indexing="ij")
    i1 = (i + 1) % cols if wrap_u else i + 1
    a, b, c, d = j * cols + i, j * cols + i1, (j + 1) * cols + i1, (j + 1) * cols + i
    F = np.stack([np.stack([a, b, c], -1), np.stack([a, c, d], -1)], 2).reshape(-1, 3)
    return F.astype(np.uint32)


def displaced_sphere(n=522, radius=1.0, amplitude=0.08, seed=11, center=(0, 0, 0)):
    """Closed smooth-normal UV sphere with n x n quads (2 n^2 triangles; n=522 -> 544 968)."""
    rng = np.random.default_rng(seed)
    v, u = np.meshgrid(np.linspace(0, 1, n + 1), np.arange(n) / n, indexing="ij")
    r = radius * (1 + amplitude * _value_noise(u, v, rng))
    r[0, :] = r[0, 0]; r[-1, :] = r[-1, 0]
    th, ph = v * np.pi, u * 2 * np.pi
    P = np.stack([r * np.sin(th) * np.cos(ph), r * np.cos(th), r * np.sin(th) * np.sin(ph)], -1).reshape(-1, 3) + np.asarray(center)
    F = grid_faces(n, n, wrap_u=True)
    V = P.astype(np.float32)
    N = vertex_normals(V, F)
    return V, N, F


def height_field(n=2236, size=2.0, amplitude=0.15, seed=4):
    """n x n quads over [-size/2, size/2]^2 (2 n^2 triangles; n=2236 -> 9 998 592)."""
    rng = np.random.default_rng(seed)
    v, u = np.meshgrid(np.linspace(0, 1, n + 1), np.linspace(0, 1, n + 1), indexing="ij")
    h = amplitude * _value_noise(u * 0.999, v, rng, octaves=4)
    V = np.stack([(u - 0.5) * size, h, (v - 0.5) * size], -1).reshape(-1, 3).astype(np.float32)
    F = grid_faces(n, n)
    return V, vertex_normals(V, F), F


def vertex_normals(V, F):
    fn = np.cross(V[F[:, 1]] - V[F[:, 0]], V[F[:, 2]] - V[F[:, 0]]).astype(np.float64)
    N = np.zeros((V.shape[0], 3))
    for k in range(3):
        np.add.at(N, F[:, k], fn)
    N /= np.maximum(np.linalg.norm(N, axis=1, keepdims=True), 1e-30)
    return N.astype(np.float32)


def look_at_camera(origin, target, up, fov, width, height, near=1e-4, far=1e4):
    """PerspectiveCamera::activate (src/perspective.cpp:53-86) + the parser's lookat
    (src/parser.cpp:286-301) in float32; returns (sampleToCamera, cameraToWorld) row-major."""
    f32 = np.float32
    o, t, upv = [np.asarray(x, np.float64) for x in (origin, target, up)]
    d = t - o; d /= np.linalg.norm(d)
    left = np.cross(upv / np.linalg.norm(upv), d); left /= np.linalg.norm(left)
    newup = np.cross(d, left)
    c2w = np.eye(4); c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = left, newup, d, o
    aspect = width / float(height)
    recip = 1.0 / (far - near); cot = 1.0 / np.tan(np.deg2rad(fov / 2.0))
    persp = np.array([[cot, 0, 0, 0], [0, cot, 0, 0], [0, 0, far * recip, -near * far * recip], [0, 0, 1, 0]])
    scale = np.diag([-0.5, -0.5 * aspect, 1.0, 1.0])
    trans = np.eye(4); trans[:3, 3] = [-1.0, -1.0 / aspect, 0.0]
    s2c = np.linalg.inv(scale @ trans @ persp)
    return s2c.astype(f32), c2w.astype(f32)


def gaussian_table(radius=2.0, stddev=0.5):
    """ImageBlock's tabulated Gaussian (src/rfilter.cpp:37-42, src/block.cpp:33-37), float32."""
    f = np.float32
    alpha = f(-1.0) / (f(2.0) * f(stddev) * f(stddev))
    x = (f(radius) * np.arange(32, dtype=np.float32)) / f(32)
    t = np.maximum(f(0), np.exp(alpha * x * x, dtype=np.float32) - np.exp(alpha * f(radius) * f(radius), dtype=np.float32))
    return np.concatenate([t, [f(0)]]).astype(np.float32)


def ajax_class_scene(n=522, width=768, height=768, spp=32) -> FlatScene:
    """C1: the camera of scenes/pa2/ajax-normals.xml:21-32 looking at a 544 968-triangle stand-in."""
    origin, target = np.array([-65.6055, 47.5762, 24.3583]), np.array([-64.8161, 47.2211, 23.8576])
    dirv = (target - origin) / np.linalg.norm(target - origin)
    center = origin + dirv * 80.0
    V, N, F = displaced_sphere(n, radius=18.0, amplitude=0.25, center=center)
    s2c, c2w = look_at_camera(origin, target, (0, 1, 0), 30.0, width, height)
    mesh = FlatMesh(V=V, F=F, N=N, bsdf_type=BSDF_DIFFUSE)
    return FlatScene(meshes=[mesh], sample_to_camera=s2c, camera_to_world=c2w, width=width, height=height, fov=30.0,
                     filter_radius=2.0, filter_table=gaussian_table(), integrator=0, sample_count=spp)


def terrain_scene(n=2236, width=3840, height=2160, spp=1) -> FlatScene:
    """C4: ~10 M-triangle height field, one diffuse material, one small area light above it."""
    V, N, F = height_field(n)
    lv = np.array([[-0.1, 1.2, -0.1], [0.1, 1.2, -0.1], [0.1, 1.2, 0.1], [-0.1, 1.2, 0.1]], np.float32)
    lf = np.array([[0, 1, 2], [0, 2, 3]], np.uint32)            # normal pointing down
    light = FlatMesh(V=lv, F=lf, bsdf_type=BSDF_DIFFUSE, is_emitter=True, radiance=np.array([40, 40, 40], np.float32))
    s2c, c2w = look_at_camera((0.0, 0.9, -1.6), (0.0, 0.0, 0.0), (0, 1, 0), 40.0, width, height)
    return FlatScene(meshes=[FlatMesh(V=V, F=F, N=N), light], sample_to_camera=s2c, camera_to_world=c2w, width=width, height=height,
                     fov=40.0, filter_radius=2.0, filter_table=gaussian_table(), integrator=5, sample_count=spp)
