"""Experiment: traversal work per ray of the device's PLOC tree against the oracle's top-down SAH tree on the same rays
(box tests and triangle tests per ray: what a better builder could still buy).  python scripts/exp_tree_quality.py [scene ...]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
from oracle.pyoracle import Port
port = Port(build=False)
for name in (sys.argv[1:] or ["cbox_mis", "table_mis"]):
    s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", name + ".npz"))
    e = Engine(0); acc = e.load_scene(s); ps = port.scene(s)
    rng = np.random.default_rng(3); n = 200_000
    film = np.stack([rng.uniform(0, s.width, n), rng.uniform(0, s.height, n)], -1).astype(np.float32)
    prim = e.camera_rays(film)
    out, valid = e.intersect(prim)
    P = out[valid, 0:3]; N = out[valid, 12:15]; m = len(P)
    u1, u2 = rng.uniform(size=m), rng.uniform(size=m)
    r, phi = np.sqrt(u1), 2 * np.pi * u2
    loc = np.stack([r * np.cos(phi), r * np.sin(phi), np.sqrt(np.maximum(0, 1 - u1))], -1)
    a = np.where(np.abs(N[:, :1]) > 0.9, np.array([[0, 1, 0]]), np.array([[1, 0, 0]]))
    T = np.cross(a, N); T /= np.linalg.norm(T, axis=1, keepdims=True); B = np.cross(N, T)
    D = (loc[:, :1] * T + loc[:, 1:2] * B + loc[:, 2:3] * N).astype(np.float32)
    bounce = np.concatenate([P, np.full((m, 1), 1e-4, np.float32), D, np.full((m, 1), np.inf, np.float32)], 1).astype(np.float32)
    print(name, "device tree: sah", round(acc["sah_cost"], 3), "nodes", acc["n_nodes"], "depth", acc["depth"])
    for label, rays in (("primary", prim), ("bounce", bounce)):
        for shadow in (False, True):
            db, dt = e.trace_count(rays, shadow=shadow)
            c = ps.trace(rays, shadow=shadow, count=True)[3]
            k = len(rays)
            print(f"  {label:8s} {'any-hit' if shadow else 'closest'}: device box {db / k:7.2f} tri {dt / k:6.2f} | oracle box {c[1] / c[0]:7.2f} tri {c[2] / c[0]:6.2f}", flush=True)
    e.close()
