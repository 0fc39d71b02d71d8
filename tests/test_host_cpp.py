"""The C++ host front end (nori_b200/host -> nori_b200/bin/nori_b200): NoriObject factory, XML scene
parser, OBJ loader, camera / filter setup, EXR + PNG writers above the C ABI.

CPU part: `nori_b200 --dump` (parse only, no GPU) must hand the C ABI exactly the descriptors the
COMPILED REFERENCE's own parser produces for the same scene files (tests/golden/host/*.npz, made by
oracle/make_host_golden.py), and must reject malformed input the way src/parser.cpp does.
GPU part: the command line tool renders through libnori_b200.so and writes the image files."""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from nori_b200.hostdump import HOST_BIN, parse_scene, read_dump
from nori_b200.scene import FlatScene

HERE = os.path.dirname(os.path.abspath(__file__))
HOST = os.path.join(HERE, "golden", "host")
SCENES = ["all_tags", "defaults", "simple_tent", "whitted_box"]


@pytest.fixture(scope="module", autouse=True)
def _built():
    from nori_b200.build import build, build_host
    build(); build_host()
    assert os.path.exists(HOST_BIN)


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("name", SCENES)
def test_descriptors_bit_equal_to_reference_parser(name, tmp_path):
    got = parse_scene(os.path.join(HOST, name + ".xml"), str(tmp_path / "d.bin"))
    want = FlatScene.load(os.path.join(HOST, name + ".npz"))
    assert (got.width, got.height, got.integrator, got.sample_count) == (want.width, want.height, want.integrator, want.sample_count)
    assert (got.rr_start, got.max_depth) == (want.rr_start, want.max_depth)
    for field in ("near_clip", "far_clip", "fov", "filter_radius", "rr_max"):
        assert np.float32(getattr(got, field)) == np.float32(getattr(want, field)), field
    # camera matrices (4x4 products, lookat, the blockwise float inverse) and the 33-entry filter table
    np.testing.assert_array_equal(_bits(got.sample_to_camera), _bits(want.sample_to_camera))
    np.testing.assert_array_equal(_bits(got.camera_to_world), _bits(want.camera_to_world))
    np.testing.assert_array_equal(_bits(got.filter_table), _bits(want.filter_table))
    np.testing.assert_array_equal(_bits(got.light_pos), _bits(want.light_pos))
    np.testing.assert_array_equal(_bits(got.light_energy), _bits(want.light_energy))
    assert len(got.meshes) == len(want.meshes)
    for g, w in zip(got.meshes, want.meshes):
        np.testing.assert_array_equal(g.F, w.F)                      # corner de-duplication in first-use order
        np.testing.assert_array_equal(_bits(g.V), _bits(w.V))        # toWorld applied to points
        assert (g.N is None) == (w.N is None) and (g.UV is None) == (w.UV is None)
        if w.N is not None:
            np.testing.assert_array_equal(_bits(g.N), _bits(w.N))    # inverse-transpose, renormalised
        if w.UV is not None:
            np.testing.assert_array_equal(_bits(g.UV), _bits(w.UV))
        assert g.bsdf_type == w.bsdf_type and g.is_emitter == w.is_emitter
        np.testing.assert_array_equal(_bits(g.bsdf_params), _bits(w.bsdf_params))
        np.testing.assert_array_equal(_bits(g.radiance), _bits(w.radiance))


BAD = {
    "unknown tag": ('<scene><integrator type="normals"/><camera type="perspective"/><sphere/></scene>', 'unexpected tag "sphere"'),
    "unknown class": ('<scene><integrator type="photonmapper"/><camera type="perspective"/></scene>', 'A constructor for class "photonmapper" could not be found!'),
    "missing attribute": ('<scene><integrator type="normals"/><camera type="perspective"><integer name="width"/></camera></scene>', 'missing attribute "value" in "integer"'),
    "unexpected attribute": ('<scene><integrator type="normals" id="x"/><camera type="perspective"/></scene>', 'unexpected attribute "id" in "integrator"'),
    "bad float": ('<scene><integrator type="normals"/><camera type="perspective"><float name="fov" value="30deg"/></camera></scene>', 'Could not parse floating point value "30deg"'),
    "wrong property type": ('<scene><integrator type="normals"/><camera type="perspective"><float name="width" value="30"/></camera></scene>', "Property 'width' has the wrong type! (expected <integer>)!"),
    "missing property": ('<scene><integrator type="simple"/><camera type="perspective"/></scene>', "Property 'position' is missing!"),
    "transform op outside transform": ('<scene><integrator type="normals"/><camera type="perspective"><translate value="1,2,3"/></camera></scene>', "transform nodes can only contain transform operations"),
    "wrong class for tag": ('<scene><integrator type="diffuse"/><camera type="perspective"/></scene>', "Unexpectedly constructed an object of type <bsdf> (expected type <integrator>)"),
    "no integrator": ('<scene><camera type="perspective"/></scene>', "No integrator was specified!"),
    "two cameras": ('<scene><integrator type="normals"/><camera type="perspective"/><camera type="perspective"/></scene>', "There can only be one camera per scene!"),
    "three values expected": ('<scene><integrator type="simple"><point name="position" value="1,2"/><color name="energy" value="1,1,1"/></integrator><camera type="perspective"/></scene>', "Expected 3 values"),
    "missing obj": ('<scene><integrator type="normals"/><camera type="perspective"/><mesh type="obj"><string name="filename" value="nope.obj"/></mesh></scene>', 'Unable to open OBJ file'),
    "not xml": ('<scene><integrator type="normals"></scene>', "closing tag"),
    "root is a property": ('<float name="x" value="1"/>', 'root element "float" must be a Nori object'),
}


@pytest.mark.parametrize("case", sorted(BAD))
def test_malformed_scenes_are_rejected_like_the_reference(case, tmp_path):
    text, expect = BAD[case]
    path = tmp_path / "bad.xml"
    path.write_text(text)
    r = subprocess.run([HOST_BIN, "--dump", str(tmp_path / "d.bin"), str(path)], capture_output=True, text=True)
    assert r.returncode != 0
    assert "Fatal error:" in r.stderr and expect in r.stderr, r.stderr       # src/main.cpp:196-199
    assert 'Error while parsing "' in r.stderr and "(at line " in r.stderr  # src/parser.cpp:307-310
    assert not (tmp_path / "d.bin").exists()


def test_missing_file_and_usage(tmp_path):
    r = subprocess.run([HOST_BIN, str(tmp_path / "absent.xml")], capture_output=True, text=True)
    assert r.returncode != 0 and "Fatal error:" in r.stderr
    r = subprocess.run([HOST_BIN], capture_output=True, text=True)
    assert r.returncode != 0 and "Syntax:" in r.stderr


# ---- readers for the two image formats the tool writes ----------------------------------------
def read_exr(path):
    """Uncompressed scanline OpenEXR with FLOAT channels (what Bitmap::saveEXR writes)."""
    b = open(path, "rb").read()
    assert struct.unpack_from("<i", b, 0)[0] == 20000630 and struct.unpack_from("<i", b, 4)[0] == 2
    off, attrs = 8, {}
    while b[off] != 0:
        e = b.index(b"\0", off); name = b[off:e].decode(); off = e + 1
        e = b.index(b"\0", off); typ = b[off:e].decode(); off = e + 1
        size = struct.unpack_from("<i", b, off)[0]; off += 4
        attrs[name] = (typ, b[off:off + size]); off += size
    off += 1
    assert attrs["compression"][1] == b"\0" and attrs["comments"][1] == b"Generated by Nori"
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    W, H = x1 - x0 + 1, y1 - y0 + 1
    names, p, ch = [], 0, attrs["channels"][1]
    while ch[p] != 0:
        e = ch.index(b"\0", p); names.append(ch[p:e].decode()); p = e + 1
        assert struct.unpack_from("<i", ch, p)[0] == 2                      # FLOAT
        p += 16
    assert names == ["B", "G", "R"]
    offsets = struct.unpack_from("<%dQ" % H, b, off)
    img = np.zeros((H, W, 3), np.float32)
    for y in range(H):
        yy, size = struct.unpack_from("<2i", b, offsets[y])
        assert yy == y and size == W * 12
        line = np.frombuffer(b, "<f4", 3 * W, offsets[y] + 8).reshape(3, W)
        img[y, :, 2], img[y, :, 1], img[y, :, 0] = line[0], line[1], line[2]
    return img


def read_png(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    off, idat, W, H = 8, b"", 0, 0
    while off < len(b):
        n, typ = struct.unpack_from(">I4s", b, off)
        data = b[off + 8:off + 8 + n]
        assert struct.unpack_from(">I", b, off + 8 + n)[0] == zlib.crc32(typ + data) & 0xffffffff
        if typ == b"IHDR":
            W, H, depth, ctype = struct.unpack_from(">IIBB", data)
            assert (depth, ctype) == (8, 2)
        elif typ == b"IDAT":
            idat += data
        off += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(H, 1 + 3 * W)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(H, W, 3)


@pytest.mark.gpu
def test_cli_renders_scene_and_writes_exr_png(tmp_path):
    xml = os.path.join(HOST, "all_tags.xml")
    out = str(tmp_path / "img")
    r = subprocess.run([HOST_BIN, "--out", out, "--spp", "16", xml], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Rendering .. done." in r.stdout and "Msamples/s" in r.stdout
    img = read_exr(out + ".exr")
    flat = FlatScene.load(os.path.join(HOST, "all_tags.npz"))
    assert img.shape == (flat.height, flat.width, 3)
    # the same render through the Python binding of the same C ABI: same samples, same kernels
    from nori_b200.gpu import Engine
    engine = Engine(0)
    engine.load_scene(flat)
    block, _ = engine.render(sample_first=0, sample_last=16)       # integrator and termination come from the scene
    b = flat.border
    rgbw = block[b:b + flat.height, b:b + flat.width]
    want = np.where(rgbw[..., 3:] != 0, rgbw[..., :3] / np.where(rgbw[..., 3:] != 0, rgbw[..., 3:], 1), 0).astype(np.float32)
    engine.close()
    np.testing.assert_allclose(img, want, rtol=2e-4, atol=1e-5)              # float atomics: summation order only
    assert np.isfinite(img).all() and img.mean() > 1e-3
    png = read_png(out + ".png")
    srgb = np.where(img <= 0.0031308, 12.92 * img, 1.055 * np.power(np.maximum(img, 0), 1 / 2.4) - 0.055)
    ref8 = np.clip(255.0 * srgb, 0, 255).astype(np.uint8)
    assert np.abs(png.astype(int) - ref8.astype(int)).max() <= 1


@pytest.mark.gpu
def test_cli_sample_range_and_no_gpu_message(tmp_path):
    xml = os.path.join(HOST, "defaults.xml")
    out = str(tmp_path / "n")
    r = subprocess.run([HOST_BIN, "--out", out, "--first", "2", "--last", "6", "--no-png", xml], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert os.path.exists(out + ".exr") and not os.path.exists(out + ".png")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([HOST_BIN, "--out", out, xml], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode != 0 and "Fatal error:" in r.stderr and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_cli_checkpoint_and_resume_equal_one_render(tmp_path):
    """--checkpoint keeps the RGBW accumulator of a sample range, --resume continues on top of it on the device:
    [0, 6) + [6, 16) must be the image of [0, 16) (same samples; float atomics: summation order only)."""
    xml = os.path.join(HOST, "all_tags.xml")
    whole, part, rest, blk = (str(tmp_path / n) for n in ("whole", "part", "rest", "part.blk"))
    def run(*args):
        r = subprocess.run([HOST_BIN, "--no-png"] + list(args) + [xml], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
    run("--out", whole, "--spp", "16")
    run("--out", part, "--first", "0", "--last", "6", "--checkpoint", blk)
    run("--out", rest, "--resume", blk, "--last", "16")
    a, b, c = read_exr(whole + ".exr"), read_exr(rest + ".exr"), read_exr(part + ".exr")
    np.testing.assert_allclose(b, a, rtol=2e-4, atol=1e-5)
    assert np.abs(c - a).mean() > 1e-4                       # the 6-sample image alone is a different (noisier) image
    r = subprocess.run([HOST_BIN, "--resume", blk, "--first", "3", "--last", "16", xml], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "cannot resume at 3" in r.stderr


@pytest.mark.gpu
def test_cli_renders_over_several_devices(tmp_path):
    """--devices A,B: sample ranges on one host thread per GPU + nori_gpu_reduce_blocks (NCCL) = the single-GPU image."""
    from nori_b200.gpu import device_count
    if device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    xml = os.path.join(HOST, "all_tags.xml")
    one, two = str(tmp_path / "one"), str(tmp_path / "two")
    for out, extra in ((one, []), (two, ["--devices", "0,1"])):
        r = subprocess.run([HOST_BIN, "--no-png", "--out", out, "--spp", "16"] + extra + [xml], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
    assert "Rendering on 2 GPUs" in r.stdout
    np.testing.assert_allclose(read_exr(two + ".exr"), read_exr(one + ".exr"), rtol=2e-4, atol=1e-5)


def test_cli_rejects_a_bad_checkpoint_before_touching_the_device(tmp_path):
    """--resume reads and validates the checkpoint (magic, block size of this camera) before any GPU work."""
    xml = os.path.join(HOST, "defaults.xml")
    bad = tmp_path / "bad.blk"; bad.write_bytes(b"not a block")
    r = subprocess.run([HOST_BIN, "--resume", str(bad), xml], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "is not an ImageBlock checkpoint" in r.stderr
    import struct
    wrong = tmp_path / "wrong.blk"
    wrong.write_bytes(b"NB2BLK1\0" + struct.pack("<3i2I", 7, 9, 2, 0, 4) + b"\0" * (7 * 9 * 16))
    r = subprocess.run([HOST_BIN, "--resume", str(wrong), xml], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "holds a 9x7 block" in r.stderr


@pytest.mark.gpu
def test_cli_pathgraph_tool_writes_the_reference_tools_images(tmp_path):
    """`nori_b200 --pathgraph K ITER`: the fork's `pg` tool in its "knn" mode (src/cluster.cpp:156-253) on the scene's own paths.
    Images as the tool composes them from each path's first vertex (src/cluster.cpp:23-107): with one pass, mc = the vertex's own
    estimate, so on diffuse first hits full = eLd + mc equals eLi = init + pt (the path tracer's value of that sample); the
    blurred estimate has the same overall brightness."""
    xml = os.path.join(HOST, "all_tags.xml")
    out = str(tmp_path / "pg")
    r = subprocess.run([HOST_BIN, "--out", out, "--spp", "1", "--pathgraph", "12", "1", xml], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Path graph: 1 spp, k = 12, 1 iterations (MIS gather)" in r.stdout and "vertices in" in r.stdout and "scatter" in r.stdout, r.stdout
    full, ind, pt, blur, init = (read_exr(out + s + ".exr") for s in ("_k-12_full", "_k-12_indirect", "_k-12_indirect_pt", "_k-12_indirect_blur", "_Le_init"))
    flat = FlatScene.load(os.path.join(HOST, "all_tags.npz"))
    assert full.shape == (flat.height, flat.width, 3)
    for img in (full, ind, pt, blur, init):
        assert np.isfinite(img).all() and (img >= 0).all()
    np.testing.assert_allclose(full, init + ind, rtol=1e-5, atol=1e-6)
    # the same pipeline through the Python binding of the same C ABI, composed as src/cluster.cpp:23-107 does
    from nori_b200.gpu import Engine, pathgraph_grid
    engine = Engine(0)
    engine.load_scene(flat)
    v = engine.path_vertices(sample_first=0, sample_last=1)
    pm, wblur, wmc, _ = engine.pathgraph_filter(v, pathgraph_grid(v, 12, 12), 1, mode="mis")      # the tool's knn mode: the MIS gather
    engine.close()
    first = np.concatenate([[True], v["groupIdx"][1:] != v["groupIdx"][:-1]])
    pix = v["groupIdx"][first]
    def image(rows):
        img = np.zeros((flat.height * flat.width, 3), np.float32); img[pix] = rows[first]
        return img.reshape(flat.height, flat.width, 3)
    np.testing.assert_array_equal(init, image(v["eLd"]))
    np.testing.assert_array_equal(pt, image(v["eLi"] - v["eLd"]))
    np.testing.assert_array_equal(ind, image(wmc[0]))                 # a gather without atomics: deterministic
    np.testing.assert_array_equal(blur, image(wblur[0]))
    assert init.mean() > 1e-3 and blur.mean() > 0 and first.sum() > 0.2 * flat.height * flat.width


@pytest.mark.gpu
def test_cli_pathgraph_cluster_mode_is_the_tools_default_mode(tmp_path):
    """`nori_b200 --pathgraph K ITER --pathgraph-cluster`: the `pg` tool's DEFAULT "opt" mode (src/cluster.cpp:204-218: BuildClusters +
    ClusterScatterWithDirectOptNR) on the scene's own paths; beside the five images it writes the filtered direct light as
    writeDirectLight does (<out>_k-K_direct, and the unfiltered one as _direct_o). Equal, to the bit, to the same stages called one by
    one through the Python binding of the same C ABI."""
    xml = os.path.join(HOST, "all_tags.xml")
    out = str(tmp_path / "pgc")
    r = subprocess.run([HOST_BIN, "--out", out, "--spp", "1", "--pathgraph", "10", "2", "--pathgraph-cluster", xml], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Path graph: 1 spp, k = 10, 2 iterations (clusters)" in r.stdout and "clusters " in r.stdout, r.stdout
    names = ("_k-10_full", "_k-10_indirect", "_k-10_indirect_pt", "_k-10_indirect_blur", "_Le_init", "_k-10_direct", "_k-10_direct_o")
    full, ind, pt, blur, init, direct, direct_o = (read_exr(out + s + ".exr") for s in names)
    flat = FlatScene.load(os.path.join(HOST, "all_tags.npz"))
    from nori_b200.gpu import Engine, pathgraph_grid, pathgraph_stride_seeds
    engine = Engine(0)
    engine.load_scene(flat)
    v, lp = engine.path_vertices(sample_first=0, sample_last=1, lights=True)
    hp = pathgraph_grid(v, 10, 10)
    of, members, offsets, _ = engine.pathgraph_build_clusters(v, hp, pathgraph_stride_seeds(len(v), len(v) // 10 + 1))
    wdirect, wblur, wmc, _ = engine.pathgraph_cluster(v, lp, members, offsets, 2, cluster_of=of)
    engine.close()
    first = np.concatenate([[True], v["groupIdx"][1:] != v["groupIdx"][:-1]])
    pix = v["groupIdx"][first]
    def image(rows):
        img = np.zeros((flat.height * flat.width, 3), np.float32); img[pix] = rows[first]
        return img.reshape(flat.height, flat.width, 3)
    np.testing.assert_array_equal(direct, image(wdirect))
    np.testing.assert_array_equal(direct_o, image(v["eLd"]))
    np.testing.assert_array_equal(init, direct_o)
    np.testing.assert_array_equal(blur, image(wblur))
    np.testing.assert_array_equal(ind, image(wmc))
    np.testing.assert_array_equal(full, image(v["eLd"] + wmc))
    assert np.isfinite(full).all(), "non-finite pixels"
    assert direct.mean() > 1e-4 and blur.mean() > 0, (float(direct.mean()), float(blur.mean()), float(init.mean()))
