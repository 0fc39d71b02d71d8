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
