"""Shared fixtures. `-m "not gpu"` runs on CPU (oracle, host logic, ABI symbols); `-m gpu` are the
parity tests proper: the CUDA path through the C ABI against the oracle (oracle/ is test
infrastructure and is only ever used as the checker)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLDEN, "scenes")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref (the compiled reference; authoring container only)")


@pytest.fixture(scope="session")
def port():
    from oracle.pyoracle import Port
    return Port(build=True)


@pytest.fixture(scope="session")
def ref():
    from oracle.pyoracle import Ref, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref/libnori_ref.so not built (needs /root/reference)")
    return Ref()


def load_scene(name):
    from nori_b200.scene import FlatScene
    return FlatScene.load(os.path.join(SCENES, name + ".npz"))


@pytest.fixture(scope="session")
def scenes():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_scene(name)
        return cache[name]
    return get


@pytest.fixture(scope="session")
def engine():
    """One GPU context for the whole session. No fallback: a missing library or device is an error."""
    from nori_b200.gpu import Engine
    e = Engine(0)
    yield e
    e.close()


@pytest.fixture(scope="session")
def gpu_scene(engine):
    """Loads a fixture scene into the engine (one resident scene at a time)."""
    state = {"name": None}

    def load(name):
        if state["name"] != name:
            engine.load_scene(load_scene(name))
            state["name"] = name
        return engine
    load.invalidate = lambda: state.update(name=None)    # for tests that load a scene of their own into `engine`
    return load


def random_rays(scene, n, seed, port_scene=None):
    """Primary rays through random film positions plus incoherent rays from points inside the
    scene bounds — [o, mint, d, maxt] rows."""
    rng = np.random.default_rng(seed)
    V = np.concatenate([m.V for m in scene.meshes])
    lo, hi = V.min(0), V.max(0)
    o = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.empty((n, 8), np.float32)
    rays[:, 0:3] = o; rays[:, 3] = 1e-4; rays[:, 4:7] = d.astype(np.float32); rays[:, 7] = np.inf
    return rays


def assert_hits_equal_reference(got, ps, rays, max_bvh_misses=5):
    """got = (tri, mesh, tuv) from the GPU. Compared with the oracle's SAH BVH; wherever the two
    differ the reference's brute-force loop (src/accel.cpp:42-55) is the judge: the GPU must equal
    it bit for bit. (The oracle BVH uses exact boxes with the reference slab test and can, very
    rarely, cull a triangle whose Moeller-Trumbore t lies an ulp outside its box; the device boxes
    are inflated, so the GPU may be right where the oracle BVH is not.)"""
    tri, mesh, tuv = got
    rtri, rmesh, rtuv = ps.trace(rays, use_bvh=True)
    bad = np.nonzero((tri != rtri) | (mesh != rmesh) | (tuv.view(np.uint32) != rtuv.view(np.uint32)).any(axis=1))[0]
    assert bad.size <= max_bvh_misses, f"{bad.size} rays differ from the oracle BVH"
    if bad.size:
        btri, bmesh, btuv = ps.trace(rays[bad], use_bvh=False)
        np.testing.assert_array_equal(tri[bad], btri)
        np.testing.assert_array_equal(mesh[bad], bmesh)
        np.testing.assert_array_equal(tuv[bad].view(np.uint32), btuv.view(np.uint32))
    return bad.size
