"""GPU parity: every operator of the hot path through the C ABI (nori_b200.gpu.Engine) against the
CPU oracle (oracle/liboracle.so, pinned to the compiled reference by tests/test_oracle_pins.py) on
the same seeded inputs. Integer/index results bit-exact; float results bit-exact where only
+,-,*,/,sqrt are involved, otherwise within the stated tolerance (libm sin/cos/log/exp)."""
import numpy as np
import pytest

from conftest import random_rays

pytestmark = pytest.mark.gpu

PCG32_DEMO_ROUND1 = [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]   # ext/pcg32/pcg32-demo.out:8


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_pcg32_known_answer(engine, port):
    assert engine.pcg32_uints(42, 54, 6).tolist() == PCG32_DEMO_ROUND1
    for st, seq, adv in [(0, 0, 0), (17, 5, 1000), (800 * 600 - 1, 1023, 7), (2**40 + 3, 2**33 + 1, 123456789)]:
        np.testing.assert_array_equal(engine.pcg32_uints(st, seq, 4096, advance=adv), port.pcg32_uints(st, seq, 4096, advance=adv))


@pytest.mark.parametrize("name", ["cbox_mis", "bunny_normals", "table_mis"])
def test_camera_rays_bit_exact(gpu_scene, port, scenes, name):
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    rng = np.random.default_rng(1)
    px = (rng.random((50000, 2)) * [s.width, s.height]).astype(np.float32)
    np.testing.assert_array_equal(bits(e.camera_rays(px)), bits(ps.camera_rays(px)))


@pytest.mark.parametrize("name", ["cbox_mis", "bunny_normals", "veach_mis"])
def test_primary_hits_bit_exact_vs_brute_force(gpu_scene, port, scenes, name):
    """Hit (mesh, triangle) ids and (t,u,v) of camera rays equal the reference's brute-force loop
    (src/accel.cpp:42-55) bit for bit."""
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    rng = np.random.default_rng(2)
    px = (rng.random((20000, 2)) * [s.width, s.height]).astype(np.float32)
    rays = ps.camera_rays(px)
    tri, mesh, tuv = e.trace(rays)
    rtri, rmesh, rtuv = ps.trace(rays, use_bvh=False)
    np.testing.assert_array_equal(mesh, rmesh)
    np.testing.assert_array_equal(tri, rtri)
    np.testing.assert_array_equal(bits(tuv), bits(rtuv))
    assert (tri != 0xffffffff).mean() > 0.2


@pytest.mark.parametrize("name", ["cbox_mis", "bunny_normals"])
def test_incoherent_and_shadow_rays_vs_brute_force(gpu_scene, port, scenes, name):
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    rays = random_rays(s, 20000, 3)
    tri, mesh, tuv = e.trace(rays)
    rtri, rmesh, rtuv = ps.trace(rays, use_bvh=False)
    np.testing.assert_array_equal(mesh, rmesh)
    np.testing.assert_array_equal(tri, rtri)
    np.testing.assert_array_equal(bits(tuv), bits(rtuv))
    # shadow rays: bounded segments, any hit — occlusion flag must agree
    rays[:, 7] = np.random.default_rng(4).uniform(0.05, 1.5, size=rays.shape[0]).astype(np.float32)
    occ = e.trace(rays, shadow=True)[0] != 0xffffffff
    rocc = ps.trace(rays, shadow=True, use_bvh=False)[0] != 0xffffffff
    np.testing.assert_array_equal(occ, rocc)
    assert 0.05 < occ.mean() < 0.999


def test_table_hits_vs_oracle_bvh(gpu_scene, port, scenes):
    """22 766 triangles: brute force on a subset, the oracle's SAH BVH on the rest."""
    e = gpu_scene("table_mis"); s = scenes("table_mis"); ps = port.scene(s)
    rng = np.random.default_rng(5)
    px = (rng.random((60000, 2)) * [s.width, s.height]).astype(np.float32)
    rays = np.concatenate([ps.camera_rays(px), random_rays(s, 40000, 6)])
    tri, mesh, tuv = e.trace(rays)
    rtri, rmesh, rtuv = ps.trace(rays, use_bvh=True)
    np.testing.assert_array_equal(mesh, rmesh); np.testing.assert_array_equal(tri, rtri)
    np.testing.assert_array_equal(bits(tuv), bits(rtuv))
    sub = slice(0, 100000, 25)
    btri, bmesh, btuv = ps.trace(rays[sub], use_bvh=False)
    np.testing.assert_array_equal(tri[sub], btri); np.testing.assert_array_equal(mesh[sub], bmesh)


@pytest.mark.parametrize("name", ["cbox_mis", "bunny_normals", "table_mis"])
def test_intersection_records_bit_exact(gpu_scene, port, scenes, name):
    """Intersection post-processing (src/accel.cpp:57-108): p, uv, shading and geometric frames."""
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    rng = np.random.default_rng(7)
    px = (rng.random((20000, 2)) * [s.width, s.height]).astype(np.float32)
    rays = np.concatenate([ps.camera_rays(px), random_rays(s, 5000, 8)])
    out, valid = e.intersect(rays)
    rout, rvalid = ps.intersect(rays, use_bvh=True)
    np.testing.assert_array_equal(valid, rvalid)
    np.testing.assert_array_equal(bits(out), bits(rout))


BSDFS = {
    "diffuse": (0, [0.3, 0.6, 0.2, 0, 0, 0]),
    "mirror": (1, [0, 0, 0, 0, 0, 0]),
    "dielectric": (2, [0, 0, 0, 0, 1.5046, 1.000277]),
    "microfacet": (3, [0.1, 0.2, 0.15, 0.1, 1.5, 1.000277]),
    "microfacet_rough": (3, [0.0, 0.0, 0.0, 0.6, 1.5046, 1.000277]),
}


@pytest.mark.parametrize("name", list(BSDFS))
def test_bsdf_sample_eval_pdf(engine, port, name):
    """BSDF::sample/eval/pdf. Tolerance 2e-5 relative (device sinf/cosf/logf/expf vs glibc);
    diffuse/mirror/dielectric decisions (measure, eta) must agree exactly."""
    btype, prm = BSDFS[name]
    rng = np.random.default_rng(9)
    n = 100000
    wi = rng.normal(size=(n, 3)).astype(np.float32); wi[:, 2] = np.abs(wi[:, 2]) if btype != 2 else wi[:, 2]
    wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    xi = rng.random((n, 2)).astype(np.float32)
    wo, w, pdf, eta, meas = engine.bsdf_sample(btype, prm, wi, xi)
    rwo, rw, rpdf, reta, rmeas = port.bsdf_sample(btype, prm, wi, xi)
    np.testing.assert_array_equal(meas, rmeas)
    np.testing.assert_array_equal(eta, reta)
    np.testing.assert_allclose(wo, rwo, rtol=0, atol=5e-5)   # z = sqrt(1 - x^2 - y^2) amplifies ulps near grazing
    np.testing.assert_allclose(w, rw, rtol=5e-4, atol=1e-6)
    np.testing.assert_allclose(pdf, rpdf, rtol=5e-4, atol=1e-5)
    wo2 = rng.normal(size=(n, 3)).astype(np.float32); wo2[:, 2] = np.abs(wo2[:, 2]); wo2 /= np.linalg.norm(wo2, axis=1, keepdims=True)
    f, p = engine.bsdf_eval_pdf(btype, prm, wi, wo2)
    rf, rp = port.bsdf_eval_pdf(btype, prm, wi, wo2)
    np.testing.assert_allclose(f, rf, rtol=2e-5, atol=1e-7)
    np.testing.assert_allclose(p, rp, rtol=2e-5, atol=1e-7)


@pytest.mark.parametrize("wid,param", [(0, 0), (1, 0), (2, 0), (3, 0), (4, 0), (5, 0), (6, 0.1), (6, 0.5)])
def test_warps(engine, port, wid, param):
    xi = np.random.default_rng(10).random((100000, 2)).astype(np.float32)
    out, pdf = engine.warp(wid, xi, param)
    rout, rpdf = port.warp(wid, xi, param)
    np.testing.assert_allclose(out, rout, rtol=0, atol=3e-5)    # sqrt(1 - r^2) amplifies ulps near grazing
    np.testing.assert_allclose(pdf, rpdf, rtol=3e-5, atol=1e-5)


def test_splat_reference_order_bit_exact(gpu_scene, port, scenes):
    """ImageBlock::put (src/block.cpp:74-103) incl. invalid-sample rejection; the ordered kernel
    reproduces the reference's accumulation order bit for bit, the atomic one within 1e-5."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    e.set_camera(s.sample_to_camera, s.camera_to_world, 96, 64, s.near_clip, s.far_clip)
    try:
        rng = np.random.default_rng(11)
        n = 20000
        pos = (rng.random((n, 2)) * [96, 64]).astype(np.float32)
        rgb = rng.random((n, 3)).astype(np.float32) * 3
        rgb[5] = [np.nan, 1, 1]; rgb[6] = [-1, 0, 0]; rgb[7] = [np.inf, 0, 0]
        pos[8] = [0.0, 0.0]; pos[9] = [95.99999, 63.99999]
        want, wdrop = port.splat(96, 64, s.filter_radius, s.filter_table, pos, rgb)
        got, drop = e.splat(pos, rgb, deterministic=True)
        assert drop == wdrop == 3
        np.testing.assert_array_equal(bits(got), bits(want))
        got2, drop2 = e.splat(pos, rgb, deterministic=False)
        assert drop2 == 3
        np.testing.assert_allclose(got2, want, rtol=1e-5, atol=1e-6)
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)


@pytest.mark.parametrize("name", ["cbox_mis", "bunny_normals", "table_mis"])
def test_quantised_node_layout_returns_the_same_hits(engine, gpu_scene, port, scenes, name):
    """NORI_ACCEL_QUANTISED: 32-byte nodes with child boxes on a 15-bit grid, rounded outwards. The tree and the
    triangle tests are the same, so hit ids and (t,u,v) must equal the float layout's bit for bit — and the
    reference's brute-force loop."""
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    rng = np.random.default_rng(11)
    prim = ps.camera_rays((rng.random((20000, 2)) * [s.width, s.height]).astype(np.float32))
    rays = np.concatenate([prim, random_rays(s, 20000, 12)])
    want = e.trace(rays)
    sh = rays.copy(); sh[:, 7] = rng.uniform(0.05, 1.5, size=sh.shape[0]).astype(np.float32)
    want_occ = e.trace(sh, shadow=True)[0] != 0xffffffff
    e.set_accel_layout(1)
    try:
        got = e.trace(rays)
        for a, b in zip(got, want):
            np.testing.assert_array_equal(bits(a) if a.dtype == np.float32 else a, bits(b) if b.dtype == np.float32 else b)
        np.testing.assert_array_equal(e.trace(sh, shadow=True)[0] != 0xffffffff, want_occ)
        if name != "table_mis":
            rtri, rmesh, rtuv = ps.trace(rays[:8000], use_bvh=False)
            np.testing.assert_array_equal(got[0][:8000], rtri); np.testing.assert_array_equal(bits(got[2][:8000]), bits(rtuv))
        # a render through the quantised layout gives every camera sample the same radiance
        pixel = rng.integers(0, s.width * s.height, 3000).astype(np.uint32); sample = rng.integers(0, 32, 3000).astype(np.uint32)
        rgb_q, _ = e.sample_radiance(pixel, sample)
        e.set_accel_layout(0)
        rgb_f, _ = e.sample_radiance(pixel, sample)
        np.testing.assert_array_equal(bits(rgb_q), bits(rgb_f))
    finally:
        e.set_accel_layout(0)


@pytest.mark.parametrize("name", ["cbox_mis", "bunny_normals", "table_mis"])
def test_wide_node_layout_returns_the_same_hits(engine, gpu_scene, port, scenes, name):
    """NORI_ACCEL_WIDE4: the same tree collapsed into 4-wide nodes (128 bytes, child boxes unchanged), the same leaves and
    triangle tests in another visiting order: closest hits (id and t, u, v) equal the binary layout's bit for bit — the
    reference's rule (minimum t, among equal t the highest triangle id) does not depend on the order —, any-hit rays agree on
    occlusion, and a render gives every camera sample the same radiance."""
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    rng = np.random.default_rng(13)
    prim = ps.camera_rays((rng.random((20000, 2)) * [s.width, s.height]).astype(np.float32))
    rays = np.concatenate([prim, random_rays(s, 20000, 14)])
    want = e.trace(rays)
    sh = rays.copy(); sh[:, 7] = rng.uniform(0.05, 1.5, size=sh.shape[0]).astype(np.float32)
    want_occ = e.trace(sh, shadow=True)[0] != 0xffffffff
    e.set_accel_layout(2)
    try:
        got = e.trace(rays)
        for a, b in zip(got, want):
            np.testing.assert_array_equal(bits(a) if a.dtype == np.float32 else a, bits(b) if b.dtype == np.float32 else b)
        np.testing.assert_array_equal(e.trace(sh, shadow=True)[0] != 0xffffffff, want_occ)
        if name != "table_mis":
            rtri, rmesh, rtuv = ps.trace(rays[:8000], use_bvh=False)
            np.testing.assert_array_equal(got[0][:8000], rtri); np.testing.assert_array_equal(bits(got[2][:8000]), bits(rtuv))
        pixel = rng.integers(0, s.width * s.height, 3000).astype(np.uint32); sample = rng.integers(0, 32, 3000).astype(np.uint32)
        rgb_w, _ = e.sample_radiance(pixel, sample)
        e.set_accel_layout(0)
        rgb_f, _ = e.sample_radiance(pixel, sample)
        np.testing.assert_array_equal(bits(rgb_w), bits(rgb_f))
    finally:
        e.set_accel_layout(0)
