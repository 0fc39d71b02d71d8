"""Golden vectors of the path-graph scatter stage from the REFERENCE's own CUDA code (oracle/_ref/libpbsdf_ref.so = src/pbsdf.cu
compiled unmodified). Needs a GPU: run on the GPU box, the result comes back through gpurun_out/ and is committed as
tests/golden/pathgraph_scatter_ref.npz.   python oracle/make_pathgraph_scatter_golden.py [out.npz]

Two parts: (eval_*) pdf_device / bsdfeval_device at synthetic records of all four material tags, one direction each; (sc_*)
computeScatterAllOnGPURecord + allGPUPdfMarginal and (mis_*) computeMISAllOnGPURecord on path-vertex records. The records of the second part are INPUT only: they
are whole paths of the table scene (diffuse + microfacet vertices) taken from nori_gpu_path_vertices, their neighbour rows come
from the reference's own buildHashGridAndFindKNN."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.pyoracle import RefCuda
from nori_b200.gpu import SPOINT_DTYPE, LPOINT_DTYPE


def unit(v):
    return (v / np.linalg.norm(v, axis=-1, keepdims=True)).astype(np.float32)


def synthetic_records(n, seed):
    """Records that reach every branch of the tool's evaluations: four tags, grazing and below-surface directions, shading
    normals that differ from the geometric ones, roughness from near-specular to rough, eta / k of metals and glass."""
    rng = np.random.default_rng(seed)
    s = np.zeros(n, SPOINT_DTYPE)
    s["pos"] = rng.random((n, 3))
    s["geoN"] = unit(rng.normal(size=(n, 3)))
    s["shN"] = unit(s["geoN"] + 0.2 * rng.normal(size=(n, 3)))
    s["wo"] = unit(s["shN"] + rng.normal(size=(n, 3)) * rng.choice([0.3, 1.0, 3.0], size=(n, 1)))
    s["wi"] = unit(s["shN"] + rng.normal(size=(n, 3)) * rng.choice([0.3, 1.0, 3.0], size=(n, 1)))
    s["diffuse"] = rng.random((n, 3)); s["specular"] = rng.random((n, 3))
    s["roughness"] = rng.choice([0.02, 0.1, 0.3, 0.7], size=n) * (0.5 + rng.random(n))
    s["eta"] = np.where(rng.random((n, 1)) < 0.5, 1.5, 0.2 + 2.0 * rng.random((n, 3)))
    s["k"] = 1.0 + 3.0 * rng.random((n, 3))
    s["rrpdf"] = rng.choice([1.0, 0.95, 0.5], size=n); s["pdf"] = rng.random(n)
    s["bsdf_type"] = rng.choice(np.array([b"d", b"o", b"c", b"t"]), size=n, p=[0.35, 0.3, 0.3, 0.05])
    # a few exact mirror configurations for the 't' branch as the tool builds them (components of wo negated: src/pbsdf.cu:450-454)
    t = np.nonzero(s["bsdf_type"] == b"t")[0][::2]
    s["wi"][t] = s["wo"][t] * np.array([-1, -1, 1], np.float32)
    return s


def synthetic_lights(n, seed):
    """LPoint records for the direct-light scatter: a third of the vertices without a light sample (density 0: the reference skips
    that term), a few emitters."""
    rng = np.random.default_rng(seed)
    lp = np.zeros(n, LPOINT_DTYPE)
    has = rng.random(n) < 0.66
    lp["L_directsample"] = rng.random((n, 3)) * 20 * has[:, None]
    lp["lightpdf"] = (0.05 + rng.random(n) * 3) * has
    lp["L_bsdfsample"] = rng.random((n, 3)) * 15 * (rng.random((n, 1)) < 0.2)
    lp["L_em"] = rng.random((n, 3)) * 15 * (rng.random((n, 1)) < 0.05)
    lp["bsdfpdf"] = rng.random(n)
    return lp


def main_direct(out):
    """(dir_*) computeDirectScatterAllOnGPU (+ computeScatterAllOnGPURecordWithDirectOpt on its result) on the committed fixture's records and rows (tests/golden/pathgraph_scatter_ref.npz) with
    seeded light records: python oracle/make_pathgraph_scatter_golden.py --direct [out.npz]."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "pathgraph_scatter_ref.npz"))
    v = np.ascontiguousarray(z["sc_sps"]).view(SPOINT_DTYPE).reshape(-1)
    nb, knn = z["sc_nb"], z["sc_knn"]
    rng = np.random.default_rng(23)
    dknn = np.minimum(knn, rng.integers(1, nb.shape[1] + 1, size=len(v))).astype(np.int32)
    lp = synthetic_lights(len(v), 29)
    ref = RefCuda()
    direct, sec = ref.direct(v, lp, nb, dknn)
    direct2, blur, mc, _ = ref.direct(v, lp, nb, dknn, knn=knn, iterations=3)     # ... + computeScatterAllOnGPURecordWithDirectOpt
    np.savez_compressed(out, dir_lps=lp.view(np.uint8).reshape(len(lp), 44), dir_dknn=dknn, dir_out=direct, dir_blur=blur, dir_mc=mc)
    print("wrote", out, len(v), "records; rows with light", int((np.abs(direct).sum(1) > 0).sum()), "non-finite", int((~np.isfinite(direct)).any(1).sum()))


def grid_clusters(pos, cells, max_size, loose, seed):
    """A clustering for the fixtures and benches (NOT the reference's BuildClusters, whose seeds come from libstdc++'s random_shuffle):
    vertices sorted by grid cell, every cell's run cut into clusters of at most max_size members; a fraction `loose` stays outside
    every cluster. Returns (cluster_of (n,), clusters, offsets) in the layout SPoints::loadClusters uploads."""
    rng = np.random.default_rng(seed)
    n = len(pos)
    lo, hi = pos.min(0), pos.max(0)
    cell = np.minimum(((pos - lo) / np.maximum(hi - lo, 1e-20) * cells).astype(np.int64), cells - 1)
    key = (cell[:, 0] * cells + cell[:, 1]) * cells + cell[:, 2]
    inside = np.nonzero(rng.random(n) >= loose)[0]
    order = inside[np.argsort(key[inside], kind="stable")]
    k = key[order]
    cluster_of = np.full(n, -1, np.int32); offsets = []; c = 0; a = 0
    while a < len(order):
        b = a + 1
        size = int(rng.integers(1, max_size + 1))
        while b < len(order) and k[b] == k[a] and b - a < size:
            b += 1
        offsets.append(a); cluster_of[order[a:b]] = c; c += 1; a = b
    return cluster_of, order.astype(np.int32), np.asarray(offsets, np.int32)


def main_cluster(out):
    """(cl_*) SPoints::ClusterScatterWithDirectOptNR's drivers (ClusterDirect, precomputedMatrixElemtns, ClusterIterations3) on the committed
    fixture's records with seeded light records and a seeded grid clustering (members up to 45 per cluster: more than a warp; 6 % of the
    vertices outside every cluster): python oracle/make_pathgraph_scatter_golden.py --cluster [out.npz]."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "pathgraph_scatter_ref.npz"))
    v = np.ascontiguousarray(z["sc_sps"]).view(SPOINT_DTYPE).reshape(-1).copy()
    lp = synthetic_lights(len(v), 31)
    cluster_of, clusters, offsets = grid_clusters(v["pos"], 6, 45, 0.06, 37)
    # consecutive path vertices rarely share a grid cell: move 300 successors into their predecessor's cluster, so that the drivers'
    # different treatment of a sender's own successor (allGPUClusterScatterRadiance skips it, MX does not) shows in the vectors
    rng = np.random.default_rng(43)
    inner = np.nonzero((v["nidx"] > 0) & (cluster_of >= 0))[0]
    for i in rng.choice(inner, 300, replace=False):
        cluster_of[i + 1] = cluster_of[i]
    order = np.argsort(np.where(cluster_of < 0, np.iinfo(np.int32).max, cluster_of), kind="stable")
    clusters = order[:(cluster_of >= 0).sum()].astype(np.int32)
    offsets = np.searchsorted(cluster_of[clusters], np.arange(len(offsets))).astype(np.int32)
    v["groupIdx"] = cluster_of
    ref = RefCuda()
    res = {"cl_of": cluster_of, "cl_members": clusters, "cl_offsets": offsets, "cl_lps": lp.view(np.uint8).reshape(len(lp), 44)}
    for it in (1, 3):
        d, b, m, _ = ref.cluster(v, lp, clusters, offsets, it)
        res[f"cl_direct_{it}"] = d; res[f"cl_blur_{it}"] = b; res[f"cl_mc_{it}"] = m
    for order in ("elements", "scatter"):                     # the tool's modes "n" (ClusterScatter2) and "t" / "l" (ClusterScatter), 3 passes
        b, m, _ = ref.cluster_scatter(v, clusters, offsets, 3, order=order)
        res[f"cl_{order}_blur"] = b; res[f"cl_{order}_mc"] = m
    np.savez_compressed(out, **res)
    sizes = np.diff(np.append(offsets, len(clusters)))
    print("wrote", out, len(v), "records,", len(offsets), "clusters, sizes", int(sizes.min()), "..", int(sizes.max()), "loose", int((cluster_of < 0).sum()),
          "non-finite", int((~np.isfinite(res["cl_blur_3"])).any(1).sum()))


def main_build_clusters(out):
    """(bc_*) SPoints::BuildClusters of the compiled reference step by step (buildBatchClusters, countMyClusters, subClusters x 2, FinalizeCluster)
    with seeded seeds, and the intermediate state each step left: python oracle/make_pathgraph_scatter_golden.py --build-clusters [out.npz].
    Positions: the fixture's 7 000 path vertices + 9 000 uniform points in their bounding box (cells with many and with few points)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "pathgraph_scatter_ref.npz"))
    v = np.ascontiguousarray(z["sc_sps"]).view(SPOINT_DTYPE).reshape(-1)
    rng = np.random.default_rng(41)
    lo, hi = v["pos"].min(0), v["pos"].max(0)
    pts = np.concatenate([v["pos"], (lo + rng.random((9000, 3)) * (hi - lo)).astype(np.float32)]).astype(np.float32)
    K = 12
    dim = np.array([7, 5, 6], np.int32)
    lo, hi = pts.min(0), pts.max(0)
    cs = ((hi - lo) / dim).astype(np.float32)
    seeds = rng.permutation(len(pts))[:len(pts) // K + 1].astype(np.int32)
    r = RefCuda().build_clusters(pts, dim, cs, lo.astype(np.float32), seeds, K)
    res = {"bc_pts": pts, "bc_dim": dim, "bc_cellsize": cs, "bc_minb": lo.astype(np.float32), "bc_seeds": seeds, "bc_K": np.int32(K)}
    for k in ("g1", "lists1", "off1", "g2", "lists2", "off2", "g3", "lists3", "off3"):
        res["bc_" + k] = r[k]
    np.savez_compressed(out, **res)
    sizes = np.diff(np.append(r["off3"], r["n_listed"]))
    print("wrote", out, len(pts), "points,", len(seeds), "seeds ->", len(r["off2"]), "->", len(r["off3"]), "clusters, sizes", int(sizes.min()), "..", int(sizes.max()),
          "outside", int((r["g3"] < 0).sum()))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--build-clusters":
        return main_build_clusters(sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "pathgraph_clusters_ref.npz"))
    if len(sys.argv) > 1 and sys.argv[1] == "--cluster":
        return main_cluster(sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "pathgraph_cluster_ref.npz"))
    if len(sys.argv) > 1 and sys.argv[1] == "--direct":
        return main_direct(sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "pathgraph_direct_ref.npz"))
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "pathgraph_scatter_ref.npz")
    ref = RefCuda()
    res = {}
    ev = synthetic_records(6000, 11)
    pdf, f = ref.eval(ev, ev["wi"])
    res["eval_sps"] = ev.view(np.uint8).reshape(len(ev), 168); res["eval_pdf"] = pdf; res["eval_bsdf"] = f

    from nori_b200.gpu import Engine
    from nori_b200.scene import FlatScene
    s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "table_mis.npz"))
    e = Engine(0); e.load_scene(s)
    e.set_camera(s.sample_to_camera, s.camera_to_world, 96, 72, s.near_clip, s.far_clip)
    v = e.path_vertices(sample_first=0, sample_last=1)
    keep = v["bsdf_type"] != b"t"                       # the tool prints a line per dielectric evaluation: whole paths without glass only
    bad_groups = np.unique(v["groupIdx"][~keep])
    v = v[~np.isin(v["groupIdx"], bad_groups)]
    # re-link: nidx = index of the successor inside the filtered array
    nxt = np.nonzero(v["nidx"] > 0)[0]
    v["nidx"][nxt] = nxt + 1
    v = v[:7000]
    last = np.nonzero(v["nidx"] == 0)[0][-1]
    v = np.ascontiguousarray(v[:last + 1])             # whole paths only
    pts = np.ascontiguousarray(v["pos"])
    lo, hi = pts.min(0), pts.max(0); dim = np.array([4, 4, 4], np.int32)
    cs = ((hi - lo) / dim).astype(np.float32)
    k = 12
    nb, knn, dknn, _ = ref.knn(pts, dim, cs, lo.astype(np.float32), k, k)
    it = 3
    pm, blur, mc, sec = ref.scatter(v, nb, knn, it)
    res["sc_sps"] = v.view(np.uint8).reshape(len(v), 168); res["sc_nb"] = nb; res["sc_knn"] = knn
    res["sc_pm"] = pm; res["sc_blur"] = blur; res["sc_mc"] = mc
    mblur, mmc, _ = ref.mis(v, nb, knn, it)                 # computeMISAllOnGPURecord on the same records and rows
    res["mis_blur"] = mblur; res["mis_mc"] = mmc
    np.savez_compressed(out, **res)
    print("wrote", out, len(v), "records", {t: int((v["bsdf_type"] == t).sum()) for t in np.unique(v["bsdf_type"])}, "nan rows", int(np.isnan(blur).any(-1).sum()))


if __name__ == "__main__":
    main()
