"""Path-graph post-process: the cluster BUILDER (SPoints::BuildClusters, include/nori/shadingPoint.h:637-760; reference kernels
src/pbsdf.cu:853-861, 888-1044, drivers :3236-3332, 3555-3622) — what the `pg` tool's default mode runs before its clustered scatter.
The reference lists the members of a cluster in the order its atomics arrive, and the members listed first seed the next split: its result
depends on scheduling. So the pin is step by step. tests/golden/pathgraph_clusters_ref.npz (oracle/make_pathgraph_scatter_golden.py
--build-clusters, made on the GPU box by the compiled reference) holds the state after EVERY step of one run; the oracle's step functions
must reproduce each step exactly from the state the reference itself left before it. The product (index-ordered lists: one of the orders
the reference can produce) is then compared with the oracle's whole builder, integer for integer."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

NPZ = os.path.join(GOLDEN, "pathgraph_clusters_ref.npz")


def case():
    z = np.load(NPZ)
    return z, z["bc_pts"], z["bc_dim"], z["bc_cellsize"], z["bc_minb"], z["bc_seeds"], int(z["bc_K"])


def members_of(lists, offsets, n_listed):
    ends = np.append(offsets[1:], n_listed)
    return [lists[a:b] for a, b in zip(offsets, ends)]


def check_lists(cluster_of, lists, offsets):
    """lists / offsets are a listing of cluster_of: every vertex that names a cluster appears once, in its cluster's range."""
    named = np.nonzero(cluster_of >= 0)[0]
    assert len(lists) == len(named) and np.array_equal(np.sort(lists), named)
    ends = np.append(offsets[1:], len(lists))
    assert offsets[0] == 0 and (ends >= offsets).all()
    owner = np.repeat(np.arange(len(offsets)), ends - offsets)
    np.testing.assert_array_equal(cluster_of[lists], owner)


def test_oracle_steps_reproduce_the_references_own_run(port):
    z, pts, dim, cs, lo, seeds, K = case()
    n = len(pts)
    # step 1: nearest seed through the hashed table (buildBatchClusters)
    g1 = port.pathgraph_cluster_assign(pts, dim, cs, lo, seeds)
    np.testing.assert_array_equal(g1, z["bc_g1"])
    assert (g1[seeds] == np.arange(len(seeds))).mean() > 0.99 and (g1 >= 0).mean() > 0.95
    lists1 = z["bc_lists1"][:(g1 >= 0).sum()]
    check_lists(z["bc_g1"], lists1, z["bc_off1"])
    assert not np.array_equal(lists1, np.sort(lists1))                       # the reference's order is NOT the index order ...
    # step 2: first split, from the lists in the order the reference left them
    g2, n2 = port.pathgraph_cluster_subdivide(pts, z["bc_g1"], lists1, z["bc_off1"], K)
    np.testing.assert_array_equal(g2, z["bc_g2"])
    assert n2 == len(z["bc_off2"]) > len(seeds)
    lists2 = z["bc_lists2"][:(g2 >= 0).sum()]
    check_lists(z["bc_g2"], lists2, z["bc_off2"])
    # step 3: second split + final lists
    g3, n3 = port.pathgraph_cluster_subdivide(pts, z["bc_g2"], lists2, z["bc_off2"], K)
    np.testing.assert_array_equal(g3, z["bc_g3"])
    assert n3 == len(z["bc_off3"])
    check_lists(z["bc_g3"], z["bc_lists3"], z["bc_off3"])
    sizes3 = np.diff(np.append(z["bc_off3"], len(z["bc_lists3"])))
    sizes1 = np.diff(np.append(z["bc_off1"], len(lists1)))
    assert sizes3.max() < sizes1.max() and sizes3.max() <= 3 * K


def test_oracle_builder_is_a_partition_with_small_clusters(port):
    """The whole builder with index-ordered lists: same first level as the reference's run, a valid listing, clusters cut down to about K
    members; a vertex further than twice the cell diagonal from every seed of its cell block stays outside (-1)."""
    z, pts, dim, cs, lo, seeds, K = case()
    g, lists, off = port.pathgraph_build_clusters(pts, dim, cs, lo, seeds, K)
    check_lists(g, lists, off)
    np.testing.assert_array_equal(g >= 0, z["bc_g3"] >= 0)
    sizes = np.diff(np.append(off, len(lists)))
    assert sizes.max() <= 3 * K and abs(len(off) - len(z["bc_off3"])) < 0.05 * len(off)
    far = np.concatenate([pts, (pts.max(0) + 10 * cs * dim)[None, :]]).astype(np.float32)     # one point far outside the grid's seeds
    g_far, _, _ = port.pathgraph_build_clusters(far, dim, cs, lo, seeds, K)
    assert g_far[-1] == -1 and np.array_equal(g_far[:-1] >= 0, g >= 0)


def test_stride_seeds_are_distinct_and_spread():
    """nori_gpu_pathgraph_stride_seeds (host-only, no GPU): n_seeds DISTINCT vertex indices s * step mod n with step coprime to n near 0.618 n —
    what the cluster mode of nori_gpu_pathgraph_render seeds BuildClusters with in place of the reference's shuffled list."""
    import math
    from nori_b200.gpu import NoriGpuError, pathgraph_stride_seeds
    for n, m in ((1, 1), (2, 2), (16, 16), (97, 40), (1000, 1000), (1_555_934, 97_246), (2**20, 2**16)):
        s = pathgraph_stride_seeds(n, m)
        assert s.shape == (m,) and s.min() >= 0 and s.max() < n and len(np.unique(s)) == m
        if m > 1:
            step = int(s[1])
            assert math.gcd(step, n) == 1 and (n <= 2 or abs(step - 0.618 * n) < 0.05 * n + 2)
            np.testing.assert_array_equal(s, (np.arange(m, dtype=np.int64) * step) % n)
    spread = np.sort(pathgraph_stride_seeds(100_000, 1000))
    assert np.diff(spread).max() < 400                                        # no large stretch of the record order without a seed
    for bad in ((0, 1), (10, 0), (10, 11)):
        with pytest.raises(NoriGpuError):
            pathgraph_stride_seeds(*bad)


@pytest.mark.gpu
def test_gpu_builder_equals_the_oracle(engine, port):
    """nori_gpu_pathgraph_build_clusters == the oracle's builder, integer for integer (cluster of every vertex, member lists, offsets), on
    the fixture's points and on SPoint records (stride 168); bad arguments are refused."""
    from nori_b200.gpu import HashParams, NoriGpuError, SPOINT_DTYPE
    z, pts, dim, cs, lo, seeds, K = case()
    hp = HashParams(); hp.dim[:] = [int(d) for d in dim]; hp.cellsize[:] = [float(c) for c in cs]; hp.minb[:] = [float(m) for m in lo]; hp.k = K; hp.dk = K
    g, lists, off, ms = engine.pathgraph_build_clusters(pts, hp, seeds)
    og, olists, ooff = port.pathgraph_build_clusters(pts, dim, cs, lo, seeds, K)
    np.testing.assert_array_equal(g, og); np.testing.assert_array_equal(lists, olists); np.testing.assert_array_equal(off, ooff)
    check_lists(g, lists, off)
    np.testing.assert_array_equal(g >= 0, z["bc_g3"] >= 0)
    rec = np.zeros(len(pts), SPOINT_DTYPE); rec["pos"] = pts
    g2, lists2, off2, _ = engine.pathgraph_build_clusters(rec, hp, seeds)
    np.testing.assert_array_equal(g2, g); np.testing.assert_array_equal(lists2, lists); np.testing.assert_array_equal(off2, off)
    for K2 in (2, 40):                                                   # many tiny clusters / hardly any split
        hp.k = hp.dk = K2
        sd = seeds[:len(pts) // K2 + 1] if K2 > K else np.random.default_rng(3).permutation(len(pts))[:len(pts) // K2 + 1].astype(np.int32)
        a = engine.pathgraph_build_clusters(pts, hp, sd); b = port.pathgraph_build_clusters(pts, dim, cs, lo, sd, K2)
        for x, y in zip(a[:3], b):
            np.testing.assert_array_equal(x, y)
    hp.k = hp.dk = K                                                     # a point nowhere near a seed stays outside every cluster
    far = np.concatenate([pts, (pts.max(0) + 10 * cs * dim)[None, :]]).astype(np.float32)
    a = engine.pathgraph_build_clusters(far, hp, seeds); b = port.pathgraph_build_clusters(far, dim, cs, lo, seeds, K)
    for x, y in zip(a[:3], b):
        np.testing.assert_array_equal(x, y)
    assert a[0][-1] == -1 and len(a[1]) == len(pts)
    with pytest.raises(NoriGpuError, match="seed outside"):
        engine.pathgraph_build_clusters(pts, hp, np.array([0, len(pts)], np.int32))
    hp.k = 1
    with pytest.raises(NoriGpuError, match="k = members per cluster"):
        engine.pathgraph_build_clusters(pts, hp, seeds)


@pytest.mark.gpu
def test_gpu_builder_vs_compiled_reference_live(engine, port):
    """The compiled reference's BuildClusters steps run on this GPU beside the product: the first level (nearest seed) is the same
    assignment; the reference's later levels, fed to the oracle's step function in the order this very run left them, come out as the
    reference computed them; both end with every vertex of the first level still in a cluster."""
    from oracle.pyoracle import RefCuda, have_refcuda
    if not have_refcuda():
        pytest.skip("oracle/_ref/libpbsdf_ref.so not built")
    from nori_b200.gpu import HashParams
    rng = np.random.default_rng(77)
    n, K = 60_000, 16
    pts = (rng.random((n, 3)) ** 2).astype(np.float32)                       # denser towards one corner
    dim = np.array([12, 12, 12], np.int32); lo = pts.min(0); cs = ((pts.max(0) - lo) / dim).astype(np.float32)
    seeds = rng.permutation(n)[:n // K + 1].astype(np.int32)
    r = RefCuda().build_clusters(pts, dim, cs, lo, seeds, K)
    hp = HashParams(); hp.dim[:] = [int(d) for d in dim]; hp.cellsize[:] = [float(c) for c in cs]; hp.minb[:] = [float(m) for m in lo]; hp.k = K; hp.dk = K
    g, lists, off, ms = engine.pathgraph_build_clusters(pts, hp, seeds)
    check_lists(g, lists, off)
    mine = port.pathgraph_cluster_assign(pts, dim, cs, lo, seeds)
    differ = np.nonzero(mine != r["g1"])[0]                              # only where two seeds are EXACTLY as far: the reference takes whichever its
    assert len(differ) <= 5                                              # atomics listed last in the bucket (`<=`), the oracle the higher seed index
    for i in differ:
        da, db = (np.linalg.norm(pts[i].astype(np.float64) - pts[seeds[q]]) for q in (mine[i], r["g1"][i]))
        assert abs(da - db) <= 1e-6 * max(da, 1e-12), (i, da, db)
    np.testing.assert_array_equal(g >= 0, r["g1"] >= 0)
    lists1 = r["lists1"][:(r["g1"] >= 0).sum()]
    g2, n2 = port.pathgraph_cluster_subdivide(pts, r["g1"], lists1, r["off1"], K)
    np.testing.assert_array_equal(g2, r["g2"])
    lists2 = r["lists2"][:(r["g2"] >= 0).sum()]
    g3, n3 = port.pathgraph_cluster_subdivide(pts, r["g2"], lists2, r["off2"], K)
    np.testing.assert_array_equal(g3, r["g3"])
    assert n3 == len(r["off3"]) and abs(len(off) - n3) < 0.05 * n3


@pytest.mark.gpu
def test_gpu_cluster_mode_end_to_end_equals_the_stages(gpu_scene, scenes, port):
    """nori_gpu_pathgraph_render in NORI_PG_CLUSTER mode = the `pg` tool's default mode on the scene's own paths, everything on the device:
    the six images equal — to the bit, there is no atomic anywhere in this mode — what the host composes (src/cluster.cpp:23-126) from the
    stages called one by one: vertex + light records, grid, stride seeds, cluster builder, clustered scatter. The oracle's pipeline on
    the same records agrees to the bit as well; the filter moves light between the vertices of a cluster but (the ratio clamps) never
    creates any: every cluster's filtered direct light stays within what its members sampled."""
    from nori_b200.gpu import pathgraph_grid, pathgraph_stride_seeds
    e = gpu_scene("table_mis"); s = scenes("table_mis")
    W, H, spp, K, it = 160, 120, 2, 12, 2
    e.set_camera(s.sample_to_camera, s.camera_to_world, W, H, s.near_clip, s.far_clip)
    try:
        v, lp = e.path_vertices(sample_first=0, sample_last=spp, lights=True)
        n = len(v)
        hp = pathgraph_grid(v, K, K)
        seeds = pathgraph_stride_seeds(n, n // K + 1)
        of, members, offsets, _ = e.pathgraph_build_clusters(v, hp, seeds)
        check_lists(of, members, offsets)
        direct, blur, mc, _ = e.pathgraph_cluster(v, lp, members, offsets, it, cluster_of=of)
        img, st = e.pathgraph_render(K, it, mode="cluster", sample_first=0, sample_last=spp)
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)
    assert st["vertices"] == n and st["dim"] == list(hp.dim) and set(img) == {"full", "indirect", "pt", "blur", "init", "direct"}
    first = np.nonzero(np.concatenate([[True], v["groupIdx"][1:] != v["groupIdx"][:-1]]))[0]
    g = v["groupIdx"][first]

    def image(rows):
        out = np.zeros((H * W, 3), np.float32)
        for kk in range(spp):                               # sample order, float32 accumulation as the kernel's
            sel = g % spp == kk
            out[g[sel] // spp] += rows[first[sel]] * np.float32(1.0 / spp)
        return out.reshape(H, W, 3)
    np.testing.assert_array_equal(img["direct"], image(direct))
    np.testing.assert_array_equal(img["blur"], image(blur))
    np.testing.assert_array_equal(img["indirect"], image(mc))
    np.testing.assert_array_equal(img["full"], image(v["eLd"] + mc))
    np.testing.assert_array_equal(img["init"], image(v["eLd"]))
    assert np.isfinite(img["full"]).all() and img["direct"].mean() > 0 and img["blur"].mean() > 0
    vv = v.copy(); vv["groupIdx"] = of
    od, ob, om = port.pathgraph_cluster(vv, lp, members, offsets, it, sequential=False)
    for got, want in ((direct, od), (blur, ob), (mc, om)):
        np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
    og, olists, ooff = port.pathgraph_build_clusters(np.ascontiguousarray(v["pos"]), list(hp.dim), list(hp.cellsize), list(hp.minb), seeds, K)
    np.testing.assert_array_equal(of, og); np.testing.assert_array_equal(members, olists); np.testing.assert_array_equal(offsets, ooff)
    inside = of >= 0
    took = np.zeros((len(offsets), 3)); np.add.at(took, of[inside], lp["L_directsample"][inside].astype(np.float64) + lp["L_bsdfsample"][inside])
    gave = np.zeros((len(offsets), 3)); np.add.at(gave, of[inside], (direct[inside] - lp["L_em"][inside]).astype(np.float64))
    assert (gave <= took * (1 + 1e-4) + 1e-4).all()
