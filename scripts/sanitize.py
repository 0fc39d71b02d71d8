"""Small pass over every kernel family, meant for the checked build (NORI_GPU_CHECKED=1 python -m
nori_b200.build: every pool / record / queue / frame index is range-checked on the device and violations
are counted and reported by nori_gpu_render):

    python scripts/sanitize.py

Tiny frames and a tiny path pool so the ring, forced pass close, compaction and the finish-tail launch
all run in a few thousand kernel launches. No oracle involved: the only question is whether the kernels
touch memory they should not."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from nori_b200.gpu import Engine          # noqa: E402
from nori_b200.scene import FlatScene     # noqa: E402

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")


def main():
    spp = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    eng = Engine(0)
    rng = np.random.default_rng(5)
    for name, w, h, pool in (("cbox_mis", 80, 60, 4096), ("table_mis", 64, 48, 2048), ("veach_mis", 48, 32, 0),
                             ("cbox_whitted", 40, 40, 1024), ("cbox_ems", 33, 47, 0), ("cbox_mats", 64, 36, 8192),
                             ("bunny_normals", 50, 50, 0)):
        sc = FlatScene.load(os.path.join(SCENES, name + ".npz"))
        sc.width, sc.height = w, h
        if pool:
            eng.set_pool_size(pool)
        print(name, eng.load_scene(sc), flush=True)
        block, st = eng.render(sample_first=0, sample_last=spp)
        block2, st2 = eng.render(sample_first=spp, sample_last=spp + 2, clear=False)
        assert np.isfinite(block2).all() and st2["camera_samples"] == w * h * 2
        print("  render", st["iterations"], "iterations,", st["kernel_launches"], "launches,", st["late_samples"], "late", flush=True)
        rays = eng.camera_rays((rng.random((3000, 2)) * [w, h]).astype(np.float32))
        eng.trace(rays); eng.trace(rays, shadow=True); eng.trace_count(rays); eng.intersect(rays)
        pix = rng.integers(0, w * h, 500).astype(np.uint32)
        eng.sample_radiance(pix, rng.integers(0, 50, 500).astype(np.uint32))
        eng.download_bitmap(); eng.download_srgb8()
        eng.upload_block(block); eng.upload_block(block2, accumulate=True)
        if name in ("cbox_mis", "table_mis"):                                  # watertight ray-triangle mode
            eng.render(sample_first=0, sample_last=2, intersect_mode=1)
            eng.set_intersect_mode(1); eng.trace(rays); eng.trace(rays, shadow=True); eng.intersect(rays); eng.set_intersect_mode(0)
        if name in ("cbox_mis", "table_mis"):                                  # 4-wide node layout; renders without the ray binning
            eng.set_accel_layout(2); eng.render(sample_first=0, sample_last=2); eng.trace(rays); eng.trace(rays, shadow=True); eng.set_accel_layout(0)
            eng.set_ray_binning(False); eng.render(sample_first=0, sample_last=2); eng.set_ray_binning(True)
            eng.trace_static(rays); eng.trace_static(rays, shadow=True)
        if name in ("cbox_mis", "table_mis", "veach_mis"):                     # path-graph tool: records, grid + kNN, scatter, MIS gather
            from nori_b200.gpu import pathgraph_grid
            v = eng.path_vertices(sample_first=0, sample_last=2)
            hp = pathgraph_grid(v, 9, 35)
            nb, knn, dknn, _ = eng.pathgraph_knn(v, hp)
            eng.pathgraph_scatter(v, nb, knn, 2); eng.pathgraph_mis(v, nb, dknn, 2)         # dknn: rows longer than a warp
            eng.pathgraph_filter(v, hp, 1); eng.pathgraph_filter(v, hp, 1, mode="mis")
            v2, lp = eng.path_vertices(sample_first=0, sample_last=2, lights=True)  # light records + direct-light filter + chained scatter
            assert len(v2) == len(v)
            eng.pathgraph_direct(v2, lp, nb, dknn); eng.pathgraph_direct(v2, lp, nb, dknn, knn=knn, iterations=2)
            order = np.argsort(np.floor(v2["pos"] * 3).astype(np.int64) @ np.array([1, 64, 4096]), kind="stable").astype(np.int32)   # clusters: runs of 40 in cell order
            offs = np.arange(0, len(order) - 100, 40, dtype=np.int32)                                                              # ... the last one long, 0 loose
            of = np.empty(len(order), np.int32); of[order] = np.minimum(np.arange(len(order)) // 40, len(offs) - 1)
            eng.pathgraph_cluster(v2, lp, order, offs, 2, cluster_of=of)
            from nori_b200.gpu import pathgraph_stride_seeds
            hpc = pathgraph_grid(v2, 12, 12)
            of2, mem2, off2, _ = eng.pathgraph_build_clusters(v2, hpc, pathgraph_stride_seeds(len(v2), len(v2) // 12 + 1))
            eng.pathgraph_cluster(v2, lp, mem2, off2, 1, cluster_of=of2)
            eng.pathgraph_cluster_scatter(v2, mem2, off2, 2, order="elements", cluster_of=of2); eng.pathgraph_cluster_scatter(v2, order, offs, 2, order="scatter", cluster_of=of)
            eng.pathgraph_render(12, 2, mode="cluster", sample_first=0, sample_last=2)
            print("  pathgraph", len(v), "vertices", flush=True)
        if name == "bunny_normals":
            eng.render(stream_mode=0, sample_first=0, sample_last=3)          # block-seeded stream
        pos = (rng.random((4000, 2)) * [w, h]).astype(np.float32)
        eng.splat(pos, rng.random((4000, 3)).astype(np.float32))
        eng.splat(pos, rng.random((4000, 3)).astype(np.float32), deterministic=True)
    for integ in (5, 6, 7):                                                    # primary_shadow, simple, ao
        eng.render(integrator=integ, sample_first=0, sample_last=2)
    wi = rng.normal(size=(2000, 3)).astype(np.float32); wi[:, 2] = np.abs(wi[:, 2]); wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    xi = rng.random((2000, 2)).astype(np.float32)
    for bt in range(4):
        p = np.array([0.3, 0.4, 0.5, 0.2, 1.5, 1.0], np.float32)
        out = eng.bsdf_sample(bt, p, wi, xi)
        eng.bsdf_eval_pdf(bt, p, wi, out[0])
    for wid in range(7):
        pts, _ = eng.warp(wid, xi, 0.3)
        eng.warp_pdf(wid, pts, 0.3)
    eng.pcg32_uints(42, 54, 100, advance=1000)
    eng.close()
    print("sanitize pass complete")


if __name__ == "__main__":
    main()
