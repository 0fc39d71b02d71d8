"""The two operator contracts the north star keeps on the host: `Color3f Integrator::Li(const Scene *, Sampler *, const Ray3f &)`
(include/nori/integrator.h:54) and `Sampler::{clone, prepare, next1D, next2D}` (include/nori/sampler.h:68-98), restored in
nori_b200/host/nori/scene.h on top of the device operator nori_gpu_li (the sampler's pcg32 state travels with every ray)."""
import os
import subprocess

import numpy as np
import pytest

from nori_b200.hostdump import HOST_BIN, parse_scene

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
HOST = os.path.join(HERE, "golden", "host")
MIS = 3


class Pcg32:
    """ext/pcg32/pcg32.h:55-112 (host side of the test: seeds and the four numbers renderBlock draws before Li)."""
    M = (1 << 64) - 1

    def __init__(self, initstate, initseq):
        self.state, self.inc = 0, ((initseq << 1) | 1) & self.M
        self.next_uint(); self.state = (self.state + initstate) & self.M; self.next_uint()

    def next_uint(self):
        old = self.state
        self.state = (old * 0x5851f42d4c957f2d + self.inc) & self.M
        x = (((old >> 18) ^ old) >> 27) & 0xffffffff
        rot = old >> 59
        return ((x >> rot) | (x << ((-rot) & 31))) & 0xffffffff

    def next_float(self):
        return np.frombuffer(np.uint32((self.next_uint() >> 9) | 0x3f800000).tobytes(), np.float32)[0] - np.float32(1.0)


def test_li_operator_continues_the_sampler_stream(gpu_scene, scenes):
    """nori_gpu_li on (ray, sampler state) pairs = the Li part of a camera sample: same radiance, bit for bit, as
    nori_gpu_sample_radiance (which seeds the same stream itself), and a returned state that has moved."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    rng = np.random.default_rng(11)
    n = 3000
    pixel = rng.integers(0, s.width * s.height, n).astype(np.uint32); sample = rng.integers(0, 512, n).astype(np.uint32)
    pos = np.empty((n, 2), np.float32); states = np.empty((n, 2), np.uint64)
    for i in range(n):
        r = Pcg32(int(pixel[i]), int(sample[i]))
        y, x = r.next_float(), r.next_float()                        # next2D draws y first (x86-64 argument order, pinned in r1)
        r.next_float(); r.next_float()                               # apertureSample
        pos[i] = [pixel[i] % s.width + x, pixel[i] // s.width + y]
        states[i] = [r.state, r.inc]
    want, wpos = e.sample_radiance(pixel, sample, integrator=MIS)
    np.testing.assert_array_equal(pos.view(np.uint32), wpos.view(np.uint32))
    got, after = e.li(e.camera_rays(pos), states, integrator=MIS)
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
    assert (after[:, 1] == states[:, 1]).all() and (after[:, 0] != states[:, 0]).mean() > 0.8     # paths that reach a diffuse surface drew numbers (the rest left the box or hit the light)
    # one sampler serving two consecutive rays (what a host loop does): the second call starts where the first ended
    a, st1 = e.li(e.camera_rays(pos[:1]), states[:1], integrator=MIS)
    b, st2 = e.li(e.camera_rays(pos[1:2]), st1, integrator=MIS)
    c, _ = e.li(e.camera_rays(pos[1:2]), st1, integrator=MIS)
    np.testing.assert_array_equal(b, c)
    assert (st2 != st1).any()


def test_host_loop_over_li_is_the_reference_block_stream(port, tmp_path):
    """`nori_b200 --li-stream`: renderBlock (src/main.cpp:38-67) written against the plugin contracts — one `independent` sampler
    per block seeded from the block offset (src/independent.cpp:48-53), Camera::sampleRay, Integrator::Li per ray. That is the
    reference's own sample stream for a PATH integrator (serial: sample k + 1 starts where sample k's path stopped drawing).
    The oracle renders the same stream on the CPU (mode 0); the splatted images agree up to libm ulps on chaotic paths."""
    xml = os.path.join(HOST, "all_tags.xml")
    flat = parse_scene(xml, str(tmp_path / "d.bin"))
    out = str(tmp_path / "li.bin")
    r = subprocess.run([HOST_BIN, "--li-stream", out, "--spp", "1", xml], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr
    rec = np.fromfile(out, np.float32).reshape(-1, 5)
    assert len(rec) == flat.width * flat.height
    ps = port.scene(flat)
    ps.set_integrator(flat.integrator)
    want, _ = ps.render(mode=0, sample_first=0, sample_last=1)
    got = port.splat(flat.width, flat.height, flat.filter_radius, flat.filter_table, rec[:, 0:2].copy(), rec[:, 2:5].copy())
    got = got[0] if isinstance(got, tuple) else got
    np.testing.assert_allclose(got[..., 3], want[..., 3], rtol=1e-5, atol=1e-6)          # same pixelSample positions
    err = np.abs(got[..., :3] - want[..., :3])
    assert (err <= 1e-3 * np.abs(want[..., :3]) + 1e-4).mean() > 0.995                  # a chaotic sample touches (2b+1)^2 cells
    assert abs(got[..., :3].mean() - want[..., :3].mean()) <= 0.01 * want[..., :3].mean()


def test_reference_parser_through_the_compiled_bridge(ref, tmp_path):
    """tests/ref_bridge.cpp (INTEGRATION.md sections 2-3 as compiled code, built against the reference's headers into oracle/_ref):
    the REFERENCE's loadFromXML builds the scene, the bridge flattens its objects into the PODs of include/nori_gpu.h and calls
    nori_gpu_upload_meshes / set_camera / set_filter / build_accel / render / download_block. The image must be the one this
    repository's own C++ host route produces for the same file (same descriptors bit for bit, hence the same samples; the
    accumulators differ by float-atomic summation order only)."""
    from nori_b200.gpu import Engine, LIB_PATH
    xml = os.path.join(HOST, "all_tags.xml")
    rs = ref.load_xml(xml)
    flat = parse_scene(xml, str(tmp_path / "d.bin"))                 # the host route's descriptors
    got = rs.bridge_render(flat.block_shape(), LIB_PATH, 0, 8)
    e = Engine(0)
    try:
        e.load_scene(flat)
        want, st = e.render(sample_first=0, sample_last=8)
    finally:
        e.close()
    assert st["camera_samples"] == flat.width * flat.height * 8
    np.testing.assert_allclose(got[..., 3], want[..., 3], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(got, want, rtol=2e-4, atol=1e-5)
    assert got[..., :3].mean() > 1e-3
