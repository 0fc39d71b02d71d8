"""NORI_ISECT_WATERTIGHT — the second ray-triangle mode of the C ABI (include/nori_gpu.h).

The reference has one test, Moeller-Trumbore (src/mesh.cpp:51-88): rays aimed at an edge or a vertex
shared by two triangles can be rejected by both (`u < 0` on one, `v < 0` on the other) and leak
through a closed surface. The watertight mode (Woop, Benthin, Wald 2013) is specified by the oracle
(oracle/noracle_math.h: rayTriangleWatertight; parity unpinned by the reference, which has no such
code). Checked here:
  * CPU: the oracle's watertight test agrees with Moeller-Trumbore away from edges, its BVH returns
    what its brute-force loop returns, and no ray aimed at the edges / vertices of a closed mesh
    leaks (Moeller-Trumbore does leak on the same rays: the count is asserted non-zero so the test
    keeps meaning something);
  * GPU (through the C ABI): hit ids and (t,u,v) bit-exact against the oracle's brute-force loop in
    watertight mode, the same no-leak property, and per-sample radiance of a render in that mode.
"""
import copy

import numpy as np
import pytest

from conftest import random_rays

NOHIT = 0xffffffff


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def sphere_scene(scenes):
    """The Cornell box's mirror sphere (628 vertices, 1088 triangles, closed) on its own."""
    s = scenes("cbox_mis")
    one = copy.copy(s)
    one.meshes = [s.meshes[3]]
    return one


def edge_and_vertex_rays(scene, n_origins=24, seed=3):
    """Rays from outside towards every vertex and three points on every edge of the closed mesh,
    kept when the target faces the origin (so the target IS the first intersection).
    Returns rays (n, 8) and the distance to the target."""
    m = scene.meshes[0]
    V, F = m.V.astype(np.float32), m.F
    c = V.mean(0)
    tg = [V]
    for a, b in ((0, 1), (1, 2), (2, 0)):
        for w in (np.float32(0.5), np.float32(0.25), np.float32(0.8125)):
            tg.append(V[F[:, a]] * (np.float32(1) - w) + V[F[:, b]] * w)
    tg = np.concatenate(tg).astype(np.float32)
    nrm = tg - c
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    radius = float(np.linalg.norm(V - c, axis=1).max())
    rng = np.random.default_rng(seed)
    rays, dist = [], []
    for _ in range(n_origins):
        dirn = rng.normal(size=3); dirn /= np.linalg.norm(dirn)
        o = (c + dirn * rng.uniform(3 * radius, 9 * radius)).astype(np.float32)
        d = tg - o
        L = np.linalg.norm(d, axis=1).astype(np.float32)
        d = (d / L[:, None]).astype(np.float32)
        keep = (nrm * d).sum(1) < -0.3
        r = np.zeros((int(keep.sum()), 8), np.float32)
        r[:, :3] = o; r[:, 3] = 1e-4; r[:, 4:7] = d[keep]; r[:, 7] = np.inf
        rays.append(r); dist.append(L[keep])
    return np.concatenate(rays), np.concatenate(dist)


def leaks(tri, tuv, dist):
    return int(((tri == NOHIT) | (np.abs(tuv[:, 0] - dist) > 1e-3 * dist)).sum())


# ---- CPU: the oracle's statement of the mode ---------------------------------------------------
def test_oracle_watertight_agrees_with_moeller_trumbore_off_edges(port, scenes):
    s = scenes("cbox_mis"); ps = port.scene(s)
    rng = np.random.default_rng(1)
    rays = ps.camera_rays((rng.random((20000, 2)) * [s.width, s.height]).astype(np.float32))
    ps.set_intersect_mode(0); mt = ps.trace(rays, use_bvh=False)
    ps.set_intersect_mode(1); wt = ps.trace(rays, use_bvh=False); wb = ps.trace(rays, use_bvh=True)
    ps.set_intersect_mode(0)
    assert (mt[0] != wt[0]).sum() <= 2 and (mt[1] != wt[1]).sum() <= 2       # random film positions: no edge-on rays
    same = mt[0] == wt[0]
    np.testing.assert_allclose(wt[2][same], mt[2][same], rtol=1e-4, atol=2e-4)     # two float evaluations of the same hit
    np.testing.assert_array_equal(wt[0], wb[0])                              # BVH == brute force in this mode too
    np.testing.assert_array_equal(bits(wt[2]), bits(wb[2]))


def test_oracle_watertight_has_no_leaks_where_moeller_trumbore_does(port, scenes):
    one = sphere_scene(scenes); ps = port.scene(one)
    rays, dist = edge_and_vertex_rays(one)
    ps.set_intersect_mode(0); tri, _, tuv = ps.trace(rays, use_bvh=False); n_mt = leaks(tri, tuv, dist)
    ps.set_intersect_mode(1); tri, _, tuv = ps.trace(rays, use_bvh=False); n_wt = leaks(tri, tuv, dist)
    assert rays.shape[0] > 50000
    assert n_wt == 0, f"watertight mode leaked {n_wt} of {rays.shape[0]} edge / vertex rays"
    assert n_mt > 0, "Moeller-Trumbore no longer leaks on these rays: the property test lost its teeth"


# ---- GPU: the C ABI in watertight mode against the oracle ---------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cbox_mis", "bunny_normals"])
def test_gpu_watertight_hits_bit_exact_vs_oracle_brute_force(gpu_scene, port, scenes, name):
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    rng = np.random.default_rng(2)
    prim = ps.camera_rays((rng.random((15000, 2)) * [s.width, s.height]).astype(np.float32))
    rays = np.concatenate([prim, random_rays(s, 15000, 3)])
    ps.set_intersect_mode(1); e.set_intersect_mode(1)
    try:
        tri, mesh, tuv = e.trace(rays)
        rtri, rmesh, rtuv = ps.trace(rays, use_bvh=False)
        np.testing.assert_array_equal(mesh, rmesh)
        np.testing.assert_array_equal(tri, rtri)
        np.testing.assert_array_equal(bits(tuv), bits(rtuv))
        assert (tri != NOHIT).mean() > 0.2
        sh = rays.copy(); sh[:, 7] = np.random.default_rng(4).uniform(0.05, 1.5, size=sh.shape[0]).astype(np.float32)
        occ = e.trace(sh, shadow=True)[0] != NOHIT
        rocc = ps.trace(sh, shadow=True, use_bvh=False)[0] != NOHIT
        np.testing.assert_array_equal(occ, rocc)
        out, valid = e.intersect(prim)                        # Intersection records built from watertight (t,u,v)
        rout, rvalid = ps.intersect(prim, use_bvh=False)
        np.testing.assert_array_equal(valid, rvalid)
        np.testing.assert_array_equal(bits(out[valid]), bits(rout[rvalid]))
    finally:
        e.set_intersect_mode(0); ps.set_intersect_mode(0)


@pytest.mark.gpu
def test_gpu_watertight_no_leaks_on_shared_edges_and_vertices(engine, gpu_scene, port, scenes):
    one = sphere_scene(scenes)
    gpu_scene.invalidate()
    engine.load_scene(one)
    try:
        rays, dist = edge_and_vertex_rays(one)
        engine.set_intersect_mode(0); tri, _, tuv = engine.trace(rays); n_mt = leaks(tri, tuv, dist)
        engine.set_intersect_mode(1); tri, _, tuv = engine.trace(rays); n_wt = leaks(tri, tuv, dist)
        ps = port.scene(one); ps.set_intersect_mode(1)
        rtri, _, rtuv = ps.trace(rays, use_bvh=False)
        np.testing.assert_array_equal(tri, rtri)
        np.testing.assert_array_equal(bits(tuv), bits(rtuv))
        assert n_wt == 0, f"watertight mode leaked {n_wt} of {rays.shape[0]} edge / vertex rays"
        assert n_mt > 0
    finally:
        engine.set_intersect_mode(0)


@pytest.mark.gpu
def test_gpu_watertight_render_matches_oracle(engine, gpu_scene, port, scenes):
    """path_mis in watertight mode: per-sample radiance against the oracle run in the same mode (same
    tolerance as the Moeller-Trumbore render tests: device sin/cos/log vs glibc)."""
    s = copy.copy(scenes("cbox_mis")); s.width, s.height = 96, 72
    gpu_scene.invalidate()
    engine.load_scene(s)
    if True:
        ps = port.scene(s); ps.set_intersect_mode(1)
        rng = np.random.default_rng(7)
        pixel = rng.integers(0, s.width * s.height, 6000).astype(np.uint32)
        sample = rng.integers(0, 64, 6000).astype(np.uint32)
        rgb, pos = engine.sample_radiance(pixel, sample, intersect_mode=1)
        want, wpos, _ = ps.sample_radiance(pixel, sample)
        np.testing.assert_array_equal(bits(pos), bits(wpos))
        rel = np.abs(rgb - want).max(1) / (np.abs(want).max(1) + 1e-3)
        assert (rel < 1e-4).mean() > 0.995
        assert abs(rgb.mean() - want.mean()) < 0.02 * want.mean()
        block, st = engine.render(sample_first=0, sample_last=4, intersect_mode=1)
        ref, _ = ps.render(mode=1, sample_first=0, sample_last=4)
        np.testing.assert_allclose(block[..., 3], ref[..., 3], rtol=1e-5, atol=1e-6)
        err = float(np.mean((block[..., :3] - ref[..., :3]) ** 2) / np.mean(ref[..., :3] ** 2))
        assert err < 2e-3, err
