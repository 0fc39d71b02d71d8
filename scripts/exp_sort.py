"""Experiment: how much faster does the closest-hit kernel trace incoherent bounce rays when they are
ordered by (origin cell, direction octant)?  python scripts/exp_sort.py [scene]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
name = sys.argv[1] if len(sys.argv) > 1 else "cbox_mis"
s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", name + ".npz"))
e = Engine(0); e.load_scene(s)
rng = np.random.default_rng(3)
n = 4_000_000
ps = np.stack([rng.uniform(0, s.width, n), rng.uniform(0, s.height, n)], -1).astype(np.float32)
prim = e.camera_rays(ps)
out, valid = e.intersect(prim)
P = out[valid, 0:3]; N = out[valid, 12:15]
m = len(P)
# cosine-distributed direction about the shading normal
u1, u2 = rng.uniform(size=m), rng.uniform(size=m)
r, phi = np.sqrt(u1), 2 * np.pi * u2
loc = np.stack([r * np.cos(phi), r * np.sin(phi), np.sqrt(np.maximum(0, 1 - u1))], -1)
a = np.where(np.abs(N[:, :1]) > 0.9, np.array([[0, 1, 0]]), np.array([[1, 0, 0]]))
T = np.cross(a, N); T /= np.linalg.norm(T, axis=1, keepdims=True); B = np.cross(N, T)
D = (loc[:, :1] * T + loc[:, 1:2] * B + loc[:, 2:3] * N).astype(np.float32)
rays = np.concatenate([P, np.full((m, 1), 1e-4, np.float32), D, np.full((m, 1), np.inf, np.float32)], 1).astype(np.float32)
lo, hi = P.min(0), P.max(0)
def morton(q, bits):
    k = np.zeros(len(q), np.uint64)
    for b in range(bits):
        for ax in range(3):
            k |= ((q[:, ax] >> b) & 1).astype(np.uint64) << np.uint64(3 * b + ax)
    return k
def order(cell_bits):
    q = np.clip(((P - lo) / (hi - lo + 1e-9) * (1 << cell_bits)).astype(np.int64), 0, (1 << cell_bits) - 1)
    octant = ((D[:, 0] < 0).astype(np.uint64) | ((D[:, 1] < 0).astype(np.uint64) << np.uint64(1)) | ((D[:, 2] < 0).astype(np.uint64) << np.uint64(2)))
    key = (morton(q, cell_bits) << np.uint64(3)) | octant
    return np.argsort(key, kind="stable")
def time_it(arr, label):
    dr = torch.from_numpy(np.ascontiguousarray(arr)).cuda(); dh = torch.empty((len(arr), 4), dtype=torch.float32, device="cuda")
    ms = [e.trace_device(dr.data_ptr(), len(arr), dh.data_ptr()) for _ in range(6)][2:]
    print(f"{label:40s} {np.mean(ms):7.3f} ms  {len(arr) / np.mean(ms) / 1e3:8.1f} Mrays/s", flush=True)
time_it(prim[valid], "primary rays (random film positions)")
time_it(rays, "bounce rays, primary-hit order")
time_it(rays[rng.permutation(m)], "bounce rays, random order")
for bits in (2, 3, 4, 6):
    time_it(rays[order(bits)], f"bounce rays, sorted: {bits}-bit cell + octant")
dirkey = np.argsort((np.arctan2(D[:, 1], D[:, 0]) * 8 // np.pi).astype(np.int64) * 64 + (np.arccos(np.clip(D[:, 2], -1, 1)) * 8 // np.pi).astype(np.int64), kind="stable")
time_it(rays[dirkey], "bounce rays, sorted by direction bins only")

# What a CTA of the shade kernel could do on its own: order the rays of each block of 256 consecutive slots
octant = ((D[:, 0] < 0).astype(np.int64) | ((D[:, 1] < 0).astype(np.int64) << 1) | ((D[:, 2] < 0).astype(np.int64) << 2))
base = rays[rng.permutation(m)]; oc = ((base[:, 4] < 0).astype(np.int64) | ((base[:, 5] < 0).astype(np.int64) << 1) | ((base[:, 6] < 0).astype(np.int64) << 2))
for blk in (256, 1024, 4096):
    key = (np.arange(m) // blk) * 8 + oc
    time_it(base[np.argsort(key, kind="stable")], f"random order, octant-sorted within blocks of {blk}")
q2 = np.clip(((base[:, 0:3] - lo) / (hi - lo + 1e-9) * 4).astype(np.int64), 0, 3)
cell = q2[:, 0] | (q2[:, 1] << 2) | (q2[:, 2] << 4)
for blk in (256, 4096):
    key = (np.arange(m) // blk) * 512 + cell * 8 + oc
    time_it(base[np.argsort(key, kind="stable")], f"random order, cell+octant-sorted within blocks of {blk}")

# Global binning by origin cell alone (what per-bin queues filled by the shade kernel would give: no order inside a bin)
for bits in (1, 2, 3):
    q = np.clip(((base[:, 0:3] - lo) / (hi - lo + 1e-9) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    key = q[:, 0] | (q[:, 1] << bits) | (q[:, 2] << (2 * bits))
    time_it(base[np.argsort(key, kind="stable")], f"random order, globally binned by {bits}-bit origin cell ({len(np.unique(key))} bins used)")
    key = key * 8 + oc
    time_it(base[np.argsort(key, kind="stable")], f"random order, globally binned by {bits}-bit cell + octant ({len(np.unique(key))} bins used)")
