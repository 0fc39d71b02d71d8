"""GPU parity of the render path proper: Integrator::Li per camera sample and whole images
(renderBlock + ImageBlock::put) against the CPU oracle on the same sample streams."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# integrator ids (include/nori_gpu.h)
NORMALS, MATS, EMS, MIS, WHITTED, PRIMARY_SHADOW, SIMPLE, AO = range(8)


def per_sample_agreement(engine, ps, scene, integrator, n, seed):
    rng = np.random.default_rng(seed)
    pixel = rng.integers(0, scene.width * scene.height, size=n).astype(np.uint32)
    sample = rng.integers(0, 4096, size=n).astype(np.uint32)
    ps.set_integrator(integrator)
    want, wpos, _ = ps.sample_radiance(pixel, sample)
    got, gpos = engine.sample_radiance(pixel, sample, integrator=integrator)
    np.testing.assert_array_equal(gpos.view(np.uint32), wpos.view(np.uint32))    # pixelSample: bit-exact
    close = np.all(np.abs(got - want) <= 1e-4 * np.maximum(np.abs(want), 1.0) + 1e-5, axis=1)
    return got, want, close


@pytest.mark.parametrize("name,integrator", [
    ("cbox_mis", MIS), ("cbox_mis", EMS), ("cbox_mis", MATS), ("cbox_whitted", WHITTED),
    ("table_mis", MIS), ("veach_mis", MIS), ("bunny_normals", NORMALS), ("cbox_mis", AO), ("cbox_mis", PRIMARY_SHADOW)])
def test_per_sample_radiance(gpu_scene, port, scenes, name, integrator):
    """The same camera sample (pixel, sample index => same PCG32 stream) yields the same radiance.
    Paths are chaotic: one ulp of libm difference can flip a Russian-roulette or Fresnel decision, so
    a small fraction of samples may differ; the mean over all samples must still agree.
    Tolerance: >= 99.5 % of samples within 1e-4 relative; means within 2 %."""
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    got, want, close = per_sample_agreement(e, ps, s, integrator, 40000, 20 + integrator)
    assert close.mean() >= 0.995, f"only {close.mean():.4%} of samples agree"
    if integrator == NORMALS:
        assert close.all()
    assert want.mean() > 0
    np.testing.assert_allclose(got.mean(0), want.mean(0), rtol=0.02, atol=1e-4)


def rel_mse(a, b, eps=1e-2):
    return float(np.mean((a - b) ** 2 / (b ** 2 + eps)))


def to_bitmap(block, border):
    rgb = block[border:-border, border:-border, :3]
    w = block[border:-border, border:-border, 3:4]
    return np.where(w != 0, rgb / np.where(w != 0, w, 1), 0)


@pytest.mark.parametrize("name,spp,tol", [("cbox_mis", 4, 2e-3), ("bunny_normals", 2, 1e-9), ("table_mis", 2, 2e-3), ("veach_mis", 2, 2e-3)])
def test_render_image_vs_oracle(gpu_scene, port, scenes, name, spp, tol):
    """Whole frame, same per-sample streams on both sides: the ImageBlocks agree up to summation
    order and the rare chaotic sample; relative MSE of the bitmap <= tol (eps 1e-2)."""
    e = gpu_scene(name); s = scenes(name); ps = port.scene(s)
    ps.set_integrator(s.integrator)
    want, wstats = ps.render(mode=1, sample_first=0, sample_last=spp)
    got, stats = e.render(stream_mode=1, sample_first=0, sample_last=spp)
    assert stats["camera_samples"] == s.width * s.height * spp
    assert stats["dropped"] == int(wstats[3])
    b = s.border
    np.testing.assert_allclose(got[..., 3], want[..., 3], rtol=1e-5, atol=1e-6)          # filter weights
    assert rel_mse(to_bitmap(got, b), to_bitmap(want, b)) <= tol
    # ray statistics: same paths => same counts up to the chaotic fraction
    assert abs(stats["closest_rays"] - int(wstats[0])) <= 0.002 * int(wstats[0])
    assert abs(stats["shadow_rays"] - int(wstats[1])) <= 0.002 * max(int(wstats[1]), 1)
    np.testing.assert_allclose(e.download_bitmap(), to_bitmap(got, b), rtol=1e-6, atol=1e-7)


def test_render_converges_to_an_independently_seeded_cpu_render(gpu_scene, port, scenes):
    """Convergence, not sample-for-sample agreement: the GPU renders samples [0, n) of every pixel, the CPU oracle renders
    DISJOINT sample indices [5000, 5000 + n) (other PCG32 streams: an independent estimate of the same image). The relative
    MSE between two independent n-sample estimates is the sum of their variances, i.e. it must fall like 1/n: from 64 to
    256 spp by about 4x (accepted: at least 2.5x), and it must be small in absolute terms at 256 spp (<= 4e-2, eps 1e-2; the CPU-vs-CPU
    value of two independent 64-spp renders is 7e-2).
    200 x 150 Cornell box, path_mis (glass + mirror + area light)."""
    import os
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    from bench import resize_camera
    small = resize_camera(s, 200, 150)
    e.set_camera(small.sample_to_camera, small.camera_to_world, 200, 150, s.near_clip, s.far_clip)
    try:
        ps = port.scene(small); ps.set_integrator(small.integrator)
        errs = []
        for n in (64, 256):
            want, _ = ps.render(mode=1, sample_first=5000, sample_last=5000 + n, threads=os.cpu_count() or 1)
            got, st = e.render(sample_first=0, sample_last=n)
            assert st["camera_samples"] == 200 * 150 * n
            errs.append(rel_mse(to_bitmap(got, s.border), to_bitmap(want, s.border)))
        assert errs[1] <= 4e-2 and errs[0] >= 2.5 * errs[1], errs
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)


def test_render_reference_block_stream_normals(gpu_scene, port, scenes):
    """NORI_STREAM_BLOCK: the reference's own sampler stream (one PCG32 per 32x32 block seeded by the
    block offset, src/independent.cpp:48-53) reproduced with pcg32::advance."""
    e = gpu_scene("bunny_normals"); s = scenes("bunny_normals"); ps = port.scene(s)
    want, _ = ps.render(mode=0, sample_first=0, sample_last=3)
    got, stats = e.render(stream_mode=0, sample_first=0, sample_last=3)
    np.testing.assert_allclose(got, want, rtol=2e-5, atol=2e-6)
    assert stats["shadow_rays"] == 0


def test_render_accumulates_sample_ranges(gpu_scene, scenes):
    """Sample-range sharding (multi-GPU partitioning): rendering [0,2) then [2,4) without clearing
    equals rendering [0,4) (up to float summation order)."""
    e = gpu_scene("cbox_mis")
    full, _ = e.render(sample_first=0, sample_last=4)
    e.render(sample_first=0, sample_last=2, download=False)
    parts, _ = e.render(sample_first=2, sample_last=4, clear=False)
    np.testing.assert_allclose(parts, full, rtol=2e-4, atol=1e-5)


def test_small_pool_and_odd_sizes(gpu_scene, scenes):
    """Ragged inputs: image size not a multiple of the 32-pixel tile, pool smaller than a pass."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    e.set_camera(s.sample_to_camera, s.camera_to_world, 75, 41, s.near_clip, s.far_clip)
    e.set_pool_size(4096)
    try:
        a, st = e.render(sample_first=0, sample_last=5)
        assert st["camera_samples"] == 75 * 41 * 5
        e.set_pool_size(1 << 20)
        b, _ = e.render(sample_first=0, sample_last=5)
        np.testing.assert_allclose(a, b, rtol=2e-4, atol=1e-5)
    finally:
        e.set_pool_size(1 << 21)
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)


def test_ray_binning_never_changes_the_image(gpu_scene, scenes):
    """nori_gpu_set_ray_binning: the counting sort of the slots between bounces only decides which lane traces which ray.
    Same samples -> the same hits, radiance and splats (up to the summation order inside a pixel, as between two pool sizes):
    on the default pool, on a pool small enough for tail compaction to kick in, and switched back on (the index arrays are
    reallocated)."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    try:
        imgs = []
        for on, pool in ((True, 1 << 21), (False, 1 << 21), (True, 4096), (False, 4096), (True, 1 << 21)):
            e.set_ray_binning(on); e.set_pool_size(pool)
            block, st = e.render(sample_first=0, sample_last=6)
            assert st["camera_samples"] == 6 * s.width * s.height
            imgs.append(block)
        for b in imgs[1:]:
            np.testing.assert_allclose(b[..., 3], imgs[0][..., 3], rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(b, imgs[0], rtol=2e-4, atol=1e-5)
    finally:
        e.set_ray_binning(True); e.set_pool_size(1 << 21)


def test_plain_bsdf_kernel_returns_the_same_radiance_bit_for_bit(gpu_scene, scenes):
    """nori_gpu_set_plain_bsdf_kernels: a scene without microfacet surfaces — the Cornell box: diffuse walls, a mirror and a glass sphere
    — is shaded by the instance of the path integrators' kernel that carries no Beckmann code. Those branches would never have run: the
    radiance of every camera sample is the same to the bit with the switch on or off, for all three path integrators; a scene WITH
    microfacet surfaces (the table) never takes that kernel."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    rng = np.random.default_rng(3)
    pix = rng.integers(0, s.width * s.height, size=40_000).astype(np.uint32); smp = rng.integers(0, 64, size=40_000).astype(np.uint32)
    try:
        for integ in (MIS, MATS, EMS):
            e.set_plain_bsdf_kernels(True); a, _ = e.sample_radiance(pix, smp, integrator=integ)
            e.set_plain_bsdf_kernels(False); b, _ = e.sample_radiance(pix, smp, integrator=integ)
            np.testing.assert_array_equal(a.view(np.uint32), b.view(np.uint32))
            assert a.sum() > 0
    finally:
        e.set_plain_bsdf_kernels(True)
    t = gpu_scene("table_mis")
    tp = rng.integers(0, 800 * 600, size=20_000).astype(np.uint32); ts = rng.integers(0, 16, size=20_000).astype(np.uint32)
    try:
        t.set_plain_bsdf_kernels(True); a, _ = t.sample_radiance(tp, ts)
        t.set_plain_bsdf_kernels(False); b, _ = t.sample_radiance(tp, ts)
        np.testing.assert_array_equal(a.view(np.uint32), b.view(np.uint32))
    finally:
        t.set_plain_bsdf_kernels(True)


def test_tiny_pool_many_tiles_splat_outlasts_trace(gpu_scene, scenes):
    """Stream ordering of the overlapped launch: with a 4 096-slot pool over 640 tiles the closest-hit pass takes
    microseconds while the splat of the previous iteration still has whole tiles to go. k_prepare must nevertheless
    run after BOTH (k_traceClosest ends with griddepcontrol.wait): a prepare that ran under a live splat would
    reset its tile list and leave tiles partly splatted — the filter-weight channel would then be short."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    e.set_camera(s.sample_to_camera, s.camera_to_world, 1024, 640, s.near_clip, s.far_clip)
    try:
        e.set_pool_size(1 << 21)
        want, _ = e.render(sample_first=0, sample_last=3)
        e.set_pool_size(4096)
        for _ in range(2):
            got, st = e.render(sample_first=0, sample_last=3)
            assert st["camera_samples"] == 1024 * 640 * 3
            np.testing.assert_allclose(got[..., 3], want[..., 3], rtol=1e-5, atol=1e-6)      # every sample splatted exactly once
            np.testing.assert_allclose(got, want, rtol=2e-4, atol=1e-5)
    finally:
        e.set_pool_size(1 << 21)
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)


def test_block_stream_rejects_sample_subranges(gpu_scene):
    """NORI_STREAM_BLOCK advances one PCG32 per block by 4 * (pixel * spp + k): a shard [first, last) with first > 0
    would replay the stream of [0, last - first). Refused, not rendered."""
    e = gpu_scene("bunny_normals")
    with pytest.raises(RuntimeError, match="sample_first must be 0"):
        e.render(stream_mode=0, integrator=NORMALS, sample_first=2, sample_last=4)


# ---- the other reconstruction filters and integrators -----------------------------
import os
HOST_GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "host")


@pytest.fixture(scope="module")
def own_engine():
    """A context of its own, so that loading other scenes does not disturb the session's scene cache."""
    from nori_b200.gpu import Engine
    e = Engine(0)
    yield e
    e.close()


@pytest.mark.parametrize("name,border,kernel", [
    ("whitted_box", 0, "k_splatPixel<0>: box filter, radius 0.5; whitted + dielectric"),
    ("simple_tent", 1, "k_splatPixel<1>: tent filter, radius 1; simple integrator, 96x160 frame"),
    ("all_tags", 1, "k_splatPixel<1>: mitchell radius 1.5; path_mis with rrStart 5 / rrMax 0.95 / maxDepth 12, microfacet"),
    ("defaults", 2, "k_splatPixel<2>: default gaussian; normals; 1280x720 default camera"),
])
def test_filters_and_integrators_vs_oracle(own_engine, port, name, border, kernel):
    """Scenes of tests/golden/host (descriptors produced by the reference's parser): every filter
    radius class of the splat stage and the integrators the Cornell-box fixtures do not use."""
    from nori_b200.scene import FlatScene
    s = FlatScene.load(os.path.join(HOST_GOLDEN, name + ".npz"))
    assert s.border == border
    spp = 3
    own_engine.load_scene(s)
    ps = port.scene(s); ps.set_integrator(s.integrator)
    want, wstats = ps.render(mode=1, sample_first=0, sample_last=spp)
    got, stats = own_engine.render(sample_first=0, sample_last=spp)
    assert got.shape == want.shape == s.block_shape()
    assert stats["camera_samples"] == s.width * s.height * spp and stats["dropped"] == int(wstats[3])
    np.testing.assert_allclose(got[..., 3], want[..., 3], rtol=1e-5, atol=1e-6)          # filter weights: same positions, same table
    if border:
        assert rel_mse(to_bitmap(got, border), to_bitmap(want, border)) <= 2e-3
    else:
        assert rel_mse(got[..., :3], want[..., :3]) <= 2e-3


def test_wide_filter_takes_the_generic_splat_kernel(own_engine, port):
    """A gaussian of radius 3.2 (border 3) is beyond the register-accumulating kernels: k_splat."""
    from nori_b200.scene import FlatScene
    s = FlatScene.load(os.path.join(HOST_GOLDEN, "simple_tent.npz"))
    table, border = port.filter_table(0, 3.2, 0.8, 0.0)                                   # gaussian, stddev 0.8
    s.filter_radius, s.filter_table = 3.2, table
    assert border == 3 == s.border
    own_engine.load_scene(s)
    ps = port.scene(s); ps.set_integrator(s.integrator)
    want, _ = ps.render(mode=1, sample_first=0, sample_last=2)
    got, _ = own_engine.render(sample_first=0, sample_last=2)
    np.testing.assert_allclose(got[..., 3], want[..., 3], rtol=1e-5, atol=1e-6)
    assert rel_mse(to_bitmap(got, 3), to_bitmap(want, 3)) <= 2e-3


@pytest.mark.gpu
def test_range_checked_build_sees_no_violation():
    """csrc/check.cuh: the engine compiled with -DNORI_GPU_CHECKED tests every pool / shadow-queue / record /
    tile-counter / traversal-stack / frame index on the device and turns a violation into an error.
    scripts/sanitize.py drives every kernel family with tiny pools (ring wrap, compaction, finish-tail)."""
    import os, subprocess, sys
    from nori_b200.build import build
    build(checked=True)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, NORI_GPU_LIB="libnori_b200_checked.so")
    r = subprocess.run([sys.executable, os.path.join(root, "scripts", "sanitize.py"), "4"], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "sanitize pass complete" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_device_tonemap_matches_bitmap_save_png(gpu_scene, scenes):
    """nori_gpu_download_srgb8 = ImageBlock::toBitmap + Color3f::toSRGB + the 8-bit quantisation of Bitmap::savePNG
    (src/bitmap.cpp:117-127, src/common.cpp:178-192) on the device: equal to the host formula on the downloaded
    bitmap up to one level where device powf and numpy differ in the last ulp."""
    e = gpu_scene("cbox_mis")
    e.render(sample_first=0, sample_last=8)
    img = e.download_bitmap()
    got = e.download_srgb8()
    assert got.dtype == np.uint8 and got.shape == img.shape
    srgb = np.where(img <= np.float32(0.0031308), np.float32(12.92) * img,
                    np.float32(1.055) * np.power(np.maximum(img, 0), np.float32(1 / 2.4), dtype=np.float32) - np.float32(0.055))
    want = np.clip(np.float32(255) * srgb, 0, 255).astype(np.uint8)
    diff = np.abs(got.astype(int) - want.astype(int))
    assert diff.max() <= 1 and (diff == 0).mean() > 0.999
    assert got.max() == 255 and got.min() == 0              # the light saturates, the background is black


def test_block_upload_resumes_and_merges(gpu_scene, scenes):
    """Checkpoint / resume through nori_gpu_upload_block: [0,3) then [3,8) with clear=0 on the restored block equals
    [0,8); accumulate=1 merges two partial blocks (ImageBlock::put(ImageBlock&), src/block.cpp:105-114)."""
    e = gpu_scene("cbox_mis")
    whole, _ = e.render(sample_first=0, sample_last=8)
    first, _ = e.render(sample_first=0, sample_last=3)
    second, _ = e.render(sample_first=3, sample_last=8)
    e.render(sample_first=0, sample_last=1)                 # scribble over the accumulator
    e.upload_block(first)
    np.testing.assert_array_equal(e.download_block(), first)
    resumed, _ = e.render(sample_first=3, sample_last=8, clear=False)
    np.testing.assert_allclose(resumed, whole, rtol=2e-4, atol=2e-4)
    e.upload_block(first); e.upload_block(second, accumulate=True)
    np.testing.assert_array_equal(e.download_block(), first + second)
    with pytest.raises(Exception):
        e.upload_block(first[:-1])


def test_reduce_blocks_in_process(engine, gpu_scene, scenes):
    """nori_gpu_reduce_blocks: `result.put(block)` across the contexts of several GPUs in ONE process (ncclReduce).
    With one GPU: n = 1 is a no-op and two contexts on one device are refused. With two or more (gpurun --gpus 2):
    two sample ranges on two devices sum to the single-device image."""
    from nori_b200.gpu import Engine, NoriGpuError, device_count, reduce_blocks
    e = gpu_scene("cbox_mis")
    whole, _ = e.render(sample_first=0, sample_last=6)
    reduce_blocks([e])
    np.testing.assert_array_equal(e.download_block(), whole)
    other = Engine(0)
    try:
        other.load_scene(scenes("cbox_mis")); other.render(sample_first=0, sample_last=1)
        with pytest.raises(NoriGpuError, match="one context per device"):
            reduce_blocks([e, other])
    finally:
        other.close()


def test_reduce_blocks_across_two_devices(engine, gpu_scene, scenes):
    """The NCCL part of nori_gpu_reduce_blocks: two sample ranges on two devices sum to the single-device image.
    Needs two GPUs (`gpurun --gpus 2`; the log of that run is profiles/r02_two_gpu_tests.log); skipped, not passed, on one."""
    from nori_b200.gpu import Engine, device_count, reduce_blocks
    if device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    e = gpu_scene("cbox_mis")
    whole, _ = e.render(sample_first=0, sample_last=6)
    second = Engine(1)
    try:
        second.load_scene(scenes("cbox_mis"))
        a, _ = e.render(sample_first=0, sample_last=3)
        b, _ = second.render(sample_first=3, sample_last=6)
        reduce_blocks([e, second], root=0)
        got = e.download_block()
        np.testing.assert_allclose(got, a + b, rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(got, whole, rtol=2e-4, atol=2e-4)
        np.testing.assert_array_equal(second.download_block(), b)      # non-root keeps its partial block
    finally:
        second.close()
