"""The step before the path: OBJ corner de-duplication + vertex gather on the device (nori_gpu_obj_dedup,
csrc/objdedup.cu) against the loader's hash map (src/obj.cpp:32-48, 96-117, 136-161: ids in order of first use)."""
import os
import subprocess

import numpy as np
import pytest

from nori_b200.hostdump import HOST_BIN, parse_scene, read_dump

pytestmark = pytest.mark.gpu
HOST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "host")


def first_use_dedup(corners):
    """The reference's loop in plain Python: a dict from (p, uv, n) to the next free id."""
    ids, F, uniq = {}, np.empty(len(corners), np.uint32), []
    for i, c in enumerate(map(tuple, corners.tolist())):
        j = ids.get(c)
        if j is None:
            j = ids[c] = len(uniq); uniq.append(c)
        F[i] = j
    return F.reshape(-1, 3), np.array(uniq, np.uint32)


def test_dedup_equals_the_first_use_hash_map(engine):
    rng = np.random.default_rng(3)
    nP, nT, nN, nF = 5000, 3000, 4000, 40000
    P = rng.normal(size=(nP, 3)).astype(np.float32); T = rng.random((nT, 2)).astype(np.float32); Nr = rng.normal(size=(nN, 3)).astype(np.float32)
    p = rng.integers(1, nP + 1, 3 * nF)
    corners = np.stack([p, (p * 7 + rng.integers(0, 2, 3 * nF)) % nT + 1, (p * 3 + rng.integers(0, 3, 3 * nF)) % nN + 1], 1).astype(np.uint32)   # few distinct triples per position
    F, V, N, UV, ms = engine.obj_dedup(corners, P, T, Nr)
    wF, uniq = first_use_dedup(corners)
    np.testing.assert_array_equal(F, wF)
    np.testing.assert_array_equal(V.view(np.uint32), P[uniq[:, 0] - 1].view(np.uint32))
    np.testing.assert_array_equal(UV.view(np.uint32), T[uniq[:, 1] - 1].view(np.uint32))
    np.testing.assert_array_equal(N.view(np.uint32), Nr[uniq[:, 2] - 1].view(np.uint32))
    assert len(V) < len(corners) and ms > 0
    # positions only (a file without vt / vn lines: the other two fields are ignored markers)
    c2 = corners.copy(); c2[:, 1:] = 0xffffffff
    F2, V2, N2, UV2, _ = engine.obj_dedup(c2, P)
    wF2, u2 = first_use_dedup(c2)
    np.testing.assert_array_equal(F2, wF2); assert N2 is None and UV2 is None
    np.testing.assert_array_equal(V2.view(np.uint32), P[u2[:, 0] - 1].view(np.uint32))


def test_dedup_rejects_corners_that_name_nothing(engine):
    P = np.zeros((4, 3), np.float32)
    bad = np.array([[1, 0, 0], [2, 0, 0], [5, 0, 0]], np.uint32)            # position 5 of 4
    with pytest.raises(RuntimeError, match="does not exist"):
        engine.obj_dedup(bad, P)
    with pytest.raises(RuntimeError, match="does not exist"):
        engine.obj_dedup(np.array([[1, 0, 0], [2, 0, 0], [0, 0, 0]], np.uint32), P)     # OBJ indices are 1-based


@pytest.mark.parametrize("name", ["all_tags", "defaults", "simple_tent", "whitted_box"])
def test_loader_with_device_dedup_hands_over_the_same_descriptors(name, tmp_path):
    """`nori_b200 --device-dedup --dump`: the C++ loader with the device operator in place of its hash map produces byte-identical
    descriptors (and those are bit-equal to the compiled reference's parser: tests/test_host_cpp.py)."""
    xml = os.path.join(HOST, name + ".xml")
    a, b = str(tmp_path / "host.bin"), str(tmp_path / "dev.bin")
    parse_scene(xml, a)
    r = subprocess.run([HOST_BIN, "--device-dedup", "--dump", b, xml], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert open(a, "rb").read() == open(b, "rb").read()


def test_dedup_ten_million_triangles(engine):
    """C4-class input: every triangle of the 9 999 392-triangle terrain written OBJ-style with its own three corners (30 M corners naming
    5 M positions). Ids must come out in first-use order: for this grid that is a closed form (checked on a prefix and by properties)."""
    from nori_b200 import synth
    V, N, F = synth.height_field(2236)
    corners = np.empty((F.size, 3), np.uint32)
    corners[:, 0] = F.ravel() + 1; corners[:, 1] = 0xffffffff; corners[:, 2] = F.ravel() + 1
    gF, gV, gN, gUV, ms = engine.obj_dedup(corners, V, None, N)
    assert len(gV) == len(V) and gUV is None
    np.testing.assert_array_equal(gV[gF.ravel()].view(np.uint32), V[F.ravel()].view(np.uint32))       # same geometry through the new ids
    np.testing.assert_array_equal(gN[gF.ravel()].view(np.uint32), N[F.ravel()].view(np.uint32))
    firsts = np.full(len(V), -1, np.int64)
    order = np.unique(gF.ravel(), return_index=True)[1]                                              # first corner using each new id
    assert (np.diff(order) > 0).all()                                                                # ids ascend with first use
    wF, _ = first_use_dedup(corners[:300000])
    np.testing.assert_array_equal(gF[:100000], wF)
    print(f"obj_dedup: {len(corners)} corners -> {len(gV)} vertices in {ms:.1f} ms on the device")
