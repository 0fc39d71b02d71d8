"""Golden vectors of the path-graph kNN from the REFERENCE's own CUDA code (oracle/_ref/libpbsdf_ref.so = src/pbsdf.cu compiled
unmodified). Needs a GPU: run on the GPU box, the result comes back through gpurun_out/ and is committed as
tests/golden/pathgraph_ref.npz.   python oracle/make_pathgraph_golden.py [out.npz]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.pyoracle import RefCuda


def point_sets():
    rng = np.random.default_rng(2024)
    a = rng.random((6000, 3)).astype(np.float32) * np.array([2.0, 1.0, 0.5], np.float32) + np.array([-1.0, 0.25, 3.0], np.float32)
    # points on a few planes and a sphere (surface-like, as path vertices are), with exact duplicates
    u = rng.random((5000, 2)).astype(np.float32)
    plane = np.stack([u[:, 0], np.zeros(5000, np.float32), u[:, 1]], 1)
    d = rng.normal(size=(3000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    sph = (0.5 + 0.3 * d).astype(np.float32)
    b = np.concatenate([plane, sph, plane[:200]]).astype(np.float32)
    return {"box": a, "surf": b}


def grid(pts, dim):
    lo, hi = pts.min(0), pts.max(0)
    dim = np.asarray(dim, np.int32)
    cs = ((hi - lo) / dim.astype(np.float32)).astype(np.float32)
    cs[cs <= 0] = 1.0
    return dim, cs, lo.astype(np.float32)


# NB: the reference divides by (largest cell population * 64 / 1024) in integer arithmetic (batchHashKNN, src/pbsdf.cu:2855-2861):
# grids whose fullest cell holds fewer than 16 points make it divide by zero (SIGFPE). The grids below are coarse enough.
def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "pathgraph_ref.npz")
    ref = RefCuda()
    res = {}
    for name, pts in point_sets().items():
        for tag, dim, k, dk in (("a", (12, 9, 7), 8, 8), ("b", (8, 6, 8), 6, 16), ("c", (3, 3, 3), 32, 4)):
            dm, cs, lo = grid(pts, dim)
            nb, knn, dknn, sec = ref.knn(pts, dm, cs, lo, k, dk)
            key = f"{name}_{tag}"
            res[key + "_pts"] = pts; res[key + "_dim"] = dm; res[key + "_cs"] = cs; res[key + "_minb"] = lo
            res[key + "_kdk"] = np.array([k, dk], np.int32); res[key + "_nb"] = nb; res[key + "_knn"] = knn; res[key + "_dknn"] = dknn
    np.savez_compressed(out, **res)
    print("wrote", out, sorted(res)[:4], "...")


if __name__ == "__main__":
    main()
