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


def scatter_leg(spp, k, iterations):
    """Second slice (marginal densities + iterated radiance scatter) on the veach scene's path vertices (diffuse + microfacet:
    the reference prints a line per dielectric evaluation, so scenes with glass would time its printf)."""
    s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "veach_mis.npz"))
    e = Engine(0); e.load_scene(s)
    v = e.path_vertices(sample_first=0, sample_last=spp)
    hp = pathgraph_grid(v, k, k)
    nb, knn, dknn, _ = e.pathgraph_knn(v, hp)
    ms, wall = [], []
    for _ in range(4):
        t0 = time.perf_counter(); pm, blur, mc, t = e.pathgraph_scatter(v, nb, knn, iterations); wall.append(time.perf_counter() - t0); ms.append(t)
    mine, mine_wall = float(np.median(ms[1:])), float(np.median(wall[1:]))
    fwall = []
    for _ in range(4):
        t0 = time.perf_counter(); fpm, fblur, fmc, fms = e.pathgraph_filter(v, hp, iterations); fwall.append(time.perf_counter() - t0)
    filter_wall = float(np.median(fwall[1:]))
    mms = []
    for _ in range(3):
        mblur, mmc, t = e.pathgraph_mis(v, nb, knn, iterations); mms.append(t)
    mis_ms = float(np.median(mms[1:]))
    evals = float(knn.sum()) * (1 + iterations) + len(v) * 2.0 * iterations          # pdf per edge once, bsdf per edge per pass, lastRun
    out = {"workload": f"path vertices of veach path_mis {s.width}x{s.height}, {spp} spp: {len(v)} vertices, k = {k} ({float(knn.mean()):.1f} used), {iterations} iterations",
           "vertices": int(len(v)), "edges": int(knn.sum()), "iterations": iterations,
           "nori_b200": {"ms_device": mine, "ms_call_wall": mine_wall * 1e3, "Gevals_per_s_device": evals / mine / 1e6,
                         "filter_call_wall_ms": filter_wall * 1e3, "filter_device_ms": [float(fms[0]), float(fms[1])],
                         "filter": "nori_gpu_pathgraph_filter: kNN + scatter in one call from the host records (H2D of 168 B per vertex, rows never leave the device, D2H of the result rows)",
                         "timing": "ms_device: CUDA events around pack + marginal + iterations x (scatter, update, lastRun), records resident; ms_call_wall: the whole nori_gpu_pathgraph_scatter call from host buffers (H2D of records and rows, D2H of all result rows)"}}
    t0 = time.perf_counter(); img, rst = e.pathgraph_render(k, iterations, mode="mis", sample_first=0, sample_last=spp); t_first = time.perf_counter() - t0
    t0 = time.perf_counter(); img, rst = e.pathgraph_render(k, iterations, mode="mis", sample_first=0, sample_last=spp); t_render = time.perf_counter() - t0
    t0 = time.perf_counter()
    v2 = e.path_vertices(sample_first=0, sample_last=spp); hp2 = pathgraph_grid(v2, k, k); e.pathgraph_filter(v2, hp2, iterations, mode="mis")
    t_staged = time.perf_counter() - t0
    out["end_to_end"] = {"what": "the pg tool's knn mode from the loaded scene to the five images on the host (MIS gather)",
                         "pathgraph_render_wall_ms": t_render * 1e3, "device_ms": [rst["ms_vertices"], rst["ms_knn"], rst["ms_stage"]],
                         "staged_calls_wall_ms": t_staged * 1e3, "staged": "nori_gpu_path_vertices (records to the host) + nori_gpu_pathgraph_grid + nori_gpu_pathgraph_filter (records back to the device, result rows to the host)"}
    out["mis"] = {"what": "computeMISAllOnGPURecord (the pg tool's knn mode): pdfsum (k^2 evaluations per vertex) + iterations x (gather, update, lastRun)",
                  "nori_b200_ms_device": mis_ms, "Gevals_per_s_device": (float((knn.astype(np.float64) ** 2).sum()) + float(knn.sum()) * iterations) / mis_ms / 1e6}
    v3, lp = e.path_vertices(sample_first=0, sample_last=spp, lights=True)
    dms = []
    for _ in range(3):
        direct, dblur, dmc, t = e.pathgraph_direct(v3, lp, nb, dknn, knn=knn, iterations=iterations); dms.append(t)
    direct_ms = float(np.median(dms[1:]))
    out["direct"] = {"what": "computeDirectScatterAllOnGPU + computeScatterAllOnGPURecordWithDirectOpt (SPoints::computeScatterRadianceAOGWithProcessRecordingAndDirectLightOPT) "
                             "on the tracer's own light records", "nori_b200_ms_device": direct_ms,
                     "vertices_with_light_sample": int((lp["lightpdf"] > 0).sum()), "emitter_hits_by_continuation": int((lp["L_bsdfsample"].sum(1) > 0).sum())}
    # the tool's default mode: cluster builder + clustered scatter (seeds: every vertex index s * step mod n, K = k members)
    from nori_b200.gpu import pathgraph_stride_seeds
    seeds = pathgraph_stride_seeds(len(v3), len(v3) // k + 1)
    bms, cms = [], []
    for _ in range(3):
        of, members, offsets, t = e.pathgraph_build_clusters(v3, hp, seeds); bms.append(t)
        cdirect, cblur, cmc, t = e.pathgraph_cluster(v3, lp, members, offsets, iterations, cluster_of=of); cms.append(t)
    t0 = time.perf_counter(); cimg, cst = e.pathgraph_render(k, iterations, mode="cluster", sample_first=0, sample_last=spp); t_cr = time.perf_counter() - t0
    sizes = np.diff(np.append(offsets, len(members)))
    out["cluster"] = {"what": "SPoints::BuildClusters + ClusterScatterWithDirectOptNR (the pg tool's default mode)", "clusters": int(len(offsets)),
                      "members_mean_max": [float(sizes.mean()), int(sizes.max())], "pairs": int((sizes.astype(np.int64) ** 2).sum()),
                      "build_ms_device": float(np.median(bms[1:])), "scatter_ms_device": float(np.median(cms[1:])),
                      "render_wall_ms": t_cr * 1e3, "render_device_ms": [cst["ms_vertices"], cst["ms_knn"], cst["ms_stage"]]}
    if have_refcuda():
        ref = RefCuda()
        vc = v3.copy(); vc["groupIdx"] = of
        csec = []
        for _ in range(2):
            rd, rb, rm, sec = ref.cluster(vc, lp, members, offsets, iterations); csec.append(sec)
        out["cluster"]["reference_scatter_ms"] = float(csec[-1]) * 1e3
        out["cluster"]["scatter_speedup_vs_reference"] = float(csec[-1]) * 1e3 / out["cluster"]["scatter_ms_device"]
        out["cluster"]["max_rel_row_error_blur"] = float((np.abs(cblur - rb) / (np.abs(rb).max(-1, keepdims=True) + 1e-3)).max())
        rb_ = ref.build_clusters(np.ascontiguousarray(v3["pos"]), list(hp.dim), list(hp.cellsize), list(hp.minb), seeds, k)
        out["cluster"]["reference_build_ms"] = rb_["seconds"] * 1e3
        out["cluster"]["reference_clusters"] = int(len(rb_["off3"]))
        out["cluster"]["first_level_equal"] = bool(np.array_equal(rb_["g3"] >= 0, of >= 0))
        dsec = []
        for _ in range(2):
            rdirect, rdblur, rdmc, sec = ref.direct(v3, lp, nb, dknn, knn=knn, iterations=iterations); dsec.append(sec)
        out["direct"]["reference_ms"] = float(dsec[-1]) * 1e3
        out["direct"]["speedup_vs_reference_device"] = float(dsec[-1]) * 1e3 / direct_ms
        out["direct"]["max_rel_row_error_direct"] = float((np.abs(direct - rdirect) / (np.abs(rdirect).max(-1, keepdims=True) + 1e-3)).max())
        out["direct"]["max_rel_row_error_last_blur"] = float((np.abs(dblur[-1] - rdblur[-1]) / (np.abs(rdblur[-1]).max(-1, keepdims=True) + 1e-3)).max())
        msec = []
        for _ in range(2):
            rmblur, rmmc, sec = ref.mis(v, nb, knn, iterations); msec.append(sec)
        out["mis"]["reference_ms"] = float(msec[-1]) * 1e3
        out["mis"]["speedup_vs_reference_device"] = float(msec[-1]) * 1e3 / mis_ms
        out["mis"]["max_rel_row_error_last_blur"] = float((np.abs(mblur[-1] - rmblur[-1]) / (np.abs(rmblur[-1]).max(-1, keepdims=True) + 1e-3)).max())
        secs = []
        for _ in range(3):
            rpm, rblur, rmc, sec = ref.scatter(v, nb, knn, iterations)
            secs.append(sec)
        r = float(np.median(secs[1:]))
        scale = np.abs(rblur[-1]).max(-1, keepdims=True) + 1e-3
        out["reference_pbsdf_cu"] = {"ms": r * 1e3, "timing": "wall clock around computeScatterAllOnGPURecord (src/pbsdf.cu:2317-2368), same GPU, records and rows resident; includes its per-iteration synchronisation and D2H copies of the result rows"}
        out["speedup_vs_reference_device"] = r * 1e3 / mine
        out["speedup_vs_reference_call"] = r / mine_wall
        out["max_rel_row_error_last_blur"] = float((np.abs(blur[-1] - rblur[-1]) / scale).max())
        out["max_rel_error_pdfmarginal"] = float((np.abs(pm - rpm) / np.maximum(np.abs(rpm), 1e-20)).max())
    return out


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
    res["scatter"] = scatter_leg(arg("--scatter-spp", 2), k, arg("--iterations", 3))
    print(json.dumps(res))
    if out:
        json.dump(res, open(out, "w"), indent=1)


main()
