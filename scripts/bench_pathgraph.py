"""Path-graph slice (hash grid + kNN) on the GPU box: this library's kernels beside the reference's own src/pbsdf.cu
(oracle/_ref/libpbsdf_ref.so, compiled unmodified for sm_100a) on the same B200 and the same points: the path vertices of the
Cornell box (path_mis, 800x600, --spp samples per pixel) produced by nori_gpu_path_vertices.
    python scripts/bench_pathgraph.py [--spp 2] [--k 16] [--out profiles/pathgraph_r02.json]
Times: this library = CUDA events around grid build + search (device-resident points; the H2D copy of the call is outside);
reference = wall clock around buildHashGridAndFindKNN (it synchronises internally; includes its D2H / host scan / H2D)."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nori_b200.gpu import Engine, pathgraph_grid
from nori_b200.scene import FlatScene
from oracle.pyoracle import RefCuda, have_refcuda      # baseline leg only


def arg(name, default):
    return type(default)(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default


def main():
    spp, k, out = arg("--spp", 2), arg("--k", 16), arg("--out", "")
    s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "cbox_mis.npz"))
    e = Engine(0); e.load_scene(s)
    t0 = time.perf_counter(); v = e.path_vertices(sample_first=0, sample_last=spp); t_vert = time.perf_counter() - t0
    pts = np.ascontiguousarray(v["pos"])
    hp = pathgraph_grid(pts, k, k)
    ms = []
    for _ in range(5):
        nb, knn, dknn, t = e.pathgraph_knn(pts, hp)
        ms.append(t)
    mine = float(np.median(ms[1:]))
    res = {"workload": f"path vertices of cbox path_mis 800x600, {spp} spp: {len(pts)} points, k = dk = {k}, grid {list(hp.dim)} (PathGraph::computeDimensions)",
           "points": int(len(pts)), "k": k, "vertices_wall_s_incl_download": t_vert,
           "nori_b200": {"ms": mine, "Mpoints_per_s": len(pts) / mine / 1e3, "timing": "CUDA events: histogram + device scan + scatter + kNN"},
           "mean_neighbours": float(knn.mean())}
    if have_refcuda():
        ref = RefCuda()
        secs = []
        for _ in range(3):
            rnb, rknn, rdknn, sec = ref.knn(pts, list(hp.dim), list(hp.cellsize), list(hp.minb), k, k)
            secs.append(sec)
        r = float(np.median(secs[1:]))
        same = float(np.mean([set(a.tolist()) == set(b.tolist()) for a, b in zip(nb[::97], rnb[::97])]))
        res["reference_pbsdf_cu"] = {"ms": r * 1e3, "Mpoints_per_s": len(pts) / r / 1e6, "timing": "wall clock around buildHashGridAndFindKNN (src/pbsdf.cu:3173-3235), same GPU"}
        res["speedup_vs_reference_kernels"] = r * 1e3 / mine
        res["rows_identical_as_sets_sampled"] = same
        res["counts_equal"] = bool((knn == rknn).all() and (dknn == rdknn).all())
    print(json.dumps(res))
    if out:
        json.dump(res, open(out, "w"), indent=1)


main()
