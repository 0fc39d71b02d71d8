"""Path-graph post-process, first slice (reference src/pbsdf.cu:844-909, 1089-1178, 3173-3235): uniform hash
grid + k nearest neighbours over path vertices, and the vertex records produced by the tracer.
CPU: the oracle restatement (oracle/noracle_pathgraph.h) against brute force and against vectors produced by the reference's own
CUDA code. GPU: the sm_100a kernels through the C ABI against the oracle (exact) and the compiled reference (sets)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

REF_NPZ = os.path.join(GOLDEN, "pathgraph_ref.npz")


def grid(pts, dim):
    lo, hi = pts.min(0), pts.max(0)
    dim = np.asarray(dim, np.int32)
    cs = ((hi - lo) / dim.astype(np.float32)).astype(np.float32)
    cs[cs <= 0] = 1.0
    return dim, cs, lo.astype(np.float32)


def dist32(pts, i, j):
    """dev_distance (src/pbsdf.cu:667-675) up to the fused multiply-adds: float32 differences, exact squares summed in float64."""
    d = (pts[j].astype(np.float32) - pts[i].astype(np.float32)).astype(np.float64)
    return np.sqrt((d * d).sum(-1))


def assert_same_sets(pts, got, want, counts_got, counts_want, upto):
    """Rows as SETS of the first `upto[i]` entries. Members may differ only where equal distances are involved: the reference's
    quickselect (findKSmallest, src/pbsdf.cu:743-760) resolves equal distances at the k-th place arbitrarily and, with equal
    distances INSIDE a row's candidates, can stop early (`if(lb==p && dists[lb] == dists[k] && flag == 0) break`) and keep a
    farther point instead of a nearer one. Returns (rows differing only by boundary ties, rows hit by the early exit)."""
    np.testing.assert_array_equal(counts_got, counts_want)
    ties = early = 0
    for i in range(len(pts)):
        m = int(upto[i])
        a, b = set(got[i, :m].tolist()), set(want[i, :m].tolist())
        assert -1 not in a and -1 not in b
        if a == b:
            continue
        only_a, only_b = np.array(sorted(a - b)), np.array(sorted(b - a))
        assert len(only_a) == len(only_b)
        da, db = dist32(pts, i, only_a), dist32(pts, i, only_b)
        dmax = max(dist32(pts, i, np.array(sorted(a))).max(), dist32(pts, i, np.array(sorted(b))).max())
        if np.allclose(da, dmax, rtol=1e-6, atol=0) and np.allclose(db, dmax, rtol=1e-6, atol=0):
            ties += 1
            continue
        dall = np.sort(dist32(pts, i, np.array(sorted(a | b))).astype(np.float32))
        assert (np.diff(dall) == 0).any(), f"row {i}: {only_a} ({da}) vs {only_b} ({db}) without any equal distances among the candidates"
        assert da.max() <= db.max() + 1e-7                            # this side holds the nearer points
        early += 1
    return ties, early


# ---- CPU: the oracle ---------------------------------------------------------------------------------------------------
def test_oracle_knn_one_cell_equals_brute_force(port):
    """With a 1 x 1 x 1 grid every point is a candidate: the rows must be the k nearest by (distance, index)."""
    rng = np.random.default_rng(5)
    pts = rng.random((700, 3)).astype(np.float32)
    pts[100:104] = pts[0]                                            # exact duplicates (fewer than k: all of them are in every row)
    dim, cs, lo = grid(pts, (1, 1, 1))
    nb, knn, dknn, d = port.pathgraph_knn(pts, dim, cs, lo, 9, 5)
    assert (knn == 9).all() and (dknn == 5).all() and (nb[:, 0] == np.arange(700)).all()
    for i in range(0, 700, 7):
        dd = dist32(pts, i, np.arange(700))
        order = np.lexsort((np.arange(700), dd.astype(np.float32)))[:9]
        assert set(order.tolist()) == set(nb[i].tolist())
        np.testing.assert_allclose(np.sort(d[i]), np.sort(dd[order]), rtol=1e-6)


def test_oracle_knn_rows_are_consistent(port):
    """Finer grid: neighbours come from the 2x2x2 block around the query, rows are sorted by distance behind the query itself,
    and the k / dk counts follow src/pbsdf.cu:1141-1164."""
    rng = np.random.default_rng(6)
    pts = rng.random((4000, 3)).astype(np.float32)
    dim, cs, lo = grid(pts, (9, 7, 5))
    nb, knn, dknn, d = port.pathgraph_knn(pts, dim, cs, lo, 4, 40)
    cell = np.clip(np.floor((pts - lo) / cs).astype(int), 0, dim - 1)
    for i in range(4000):
        m = int(max(knn[i], dknn[i]))
        row = nb[i, :m]
        assert row[0] == i and (row >= 0).all() and (nb[i, m:] == -1).all() and len(set(row.tolist())) == m
        assert (np.abs(cell[row] - cell[i]) <= 1).all()
        assert (np.diff(d[i, 1:m]) >= 0).all()
        assert knn[i] == min(4, m) and (dknn[i] == 40 or dknn[i] == m)


@pytest.mark.skipif(not os.path.exists(REF_NPZ), reason="tests/golden/pathgraph_ref.npz not generated yet (oracle/make_pathgraph_golden.py on a GPU box)")
def test_oracle_knn_vs_reference_cuda_vectors(port):
    """Pin: the reference's own src/pbsdf.cu (compiled unmodified, run on a B200) produced these neighbour rows."""
    z = np.load(REF_NPZ)
    keys = sorted(k[:-4] for k in z.files if k.endswith("_pts"))
    assert keys
    for key in keys:
        pts, (k, dk) = z[key + "_pts"], z[key + "_kdk"]
        nb, knn, dknn, _ = port.pathgraph_knn(pts, z[key + "_dim"], z[key + "_cs"], z[key + "_minb"], int(k), int(dk))
        np.testing.assert_array_equal(dknn, z[key + "_dknn"])
        for m_of in (np.maximum, np.minimum):                        # all max(k, dk) entries; and the first min(k, dk) are the nearest
            ties, early = assert_same_sets(pts, nb, z[key + "_nb"], knn, z[key + "_knn"], m_of(knn, dknn))
            if key.startswith("box"):
                assert ties == 0 and early == 0                      # distinct points: the reference is a function of its input, rows are EQUAL as sets
            assert early <= 0.01 * len(pts)                          # 'surf' holds 200 exact duplicates: the early exit shows in < 1 % of the rows


# ---- GPU ------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("n,dim,k,dk", [(20000, (17, 13, 9), 8, 8), (20000, (30, 30, 30), 5, 24), (3000, (2, 2, 2), 64, 3), (50, (4, 4, 4), 16, 16), (1, (1, 1, 1), 4, 4)])
def test_gpu_knn_equals_oracle(engine, port, n, dim, k, dk):
    """Both sides select by the same total order (distance, index): rows, counts and row order are EQUAL."""
    from nori_b200.gpu import pathgraph_grid
    rng = np.random.default_rng(n + k)
    pts = (rng.random((n, 3)) * [3.0, 1.0, 2.0] - [1.0, 0.0, 5.0]).astype(np.float32)
    if n > 100:
        pts[50:60] = pts[3]                                          # duplicates
    hp = pathgraph_grid(pts, k, dk, dim)
    nb, knn, dknn, ms = engine.pathgraph_knn(pts, hp)
    wnb, wknn, wdknn, _ = port.pathgraph_knn(pts, list(hp.dim), list(hp.cellsize), list(hp.minb), k, dk)
    np.testing.assert_array_equal(knn, wknn); np.testing.assert_array_equal(dknn, wdknn)
    np.testing.assert_array_equal(nb, wnb)
    assert ms >= 0


@pytest.mark.gpu
def test_gpu_path_vertices_and_their_graph(gpu_scene, scenes, port):
    """Vertex records out of the tracer (NoriGpuSPoint = SPoint, 168 bytes): first vertices are the primary hits, consecutive
    vertices of a path are joined by the sampled direction, and the kNN over > 1 M of them equals the oracle's."""
    from nori_b200.gpu import pathgraph_grid
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    v = e.path_vertices(sample_first=0, sample_last=1)
    assert len(v) > 1_000_000 and v.dtype.itemsize == 168
    first = np.concatenate([[True], v["groupIdx"][1:] != v["groupIdx"][:-1]])
    assert (np.diff(v["groupIdx"]) >= 0).all()                        # path-contiguous, in camera-sample order
    # every non-last vertex of a path points at its successor
    nxt = np.nonzero(v["nidx"][:-1] > 0)[0]
    assert (v["nidx"][nxt] == nxt + 1).all()                          # nidx = record index of the successor
    assert (v["groupIdx"][nxt + 1] == v["groupIdx"][nxt]).all() and (~first[nxt + 1]).all()
    seg = v["pos"][nxt + 1] - v["pos"][nxt]
    length = np.linalg.norm(seg, axis=1)
    far = length > 1e-2                                               # positions are barycentric interpolations: short segments are noisy
    cosang = (seg[far] / length[far, None] * v["wi"][nxt][far]).sum(1)
    assert far.mean() > 0.9 and np.abs(cosang - 1).max() < 1e-3 and np.abs((v["wo"][nxt + 1] + v["wi"][nxt])).max() < 1e-6
    last = np.nonzero(v["nidx"] == 0)[0]
    assert ((last + 1 == len(v)) | first[np.minimum(last + 1, len(v) - 1)]).all()
    assert set(np.unique(v["bsdf_type"]).tolist()) <= {b"d", b"c", b"t"} and (np.abs(np.linalg.norm(v["shN"], axis=1) - 1) < 1e-4).all()
    # first vertices = primary hits of the same camera samples (Intersection::p of the raw operator, bit-exact)
    g = v["groupIdx"][first][:5000]
    ps_rng = port.pcg32_floats
    pos = np.empty((len(g), 2), np.float32)
    for i, gi in enumerate(g):
        r = ps_rng(int(gi), 0, 2)                                     # seed(pixel, sample 0): next2D draws y first (pinned in r1)
        pos[i] = [gi % s.width + r[1], gi // s.width + r[0]]
    its, valid = e.intersect(e.camera_rays(pos))
    assert valid.all()
    np.testing.assert_array_equal(its[:, 0:3].view(np.uint32), v["pos"][first][:5000].view(np.uint32))
    # the graph over all of them
    hp = pathgraph_grid(v, 12, 12)
    nb, knn, dknn, ms = e.pathgraph_knn(v, hp)
    sub = np.ascontiguousarray(v["pos"])
    wnb, wknn, wdknn, _ = port.pathgraph_knn(sub, list(hp.dim), list(hp.cellsize), list(hp.minb), 12, 12)
    np.testing.assert_array_equal(knn, wknn); np.testing.assert_array_equal(nb, wnb)


@pytest.mark.gpu
def test_knn_rejects_what_it_does_not_cover(engine):
    from nori_b200.gpu import pathgraph_grid
    pts = np.random.default_rng(1).random((100, 3)).astype(np.float32)
    with pytest.raises(RuntimeError, match="above 200"):
        engine.pathgraph_knn(pts, pathgraph_grid(pts, 201, 4, (2, 2, 2)))
    hp = pathgraph_grid(pts, 4, 4, (2, 2, 2)); hp.cellsize[1] = 0.0
    with pytest.raises(RuntimeError, match="cell sizes must be positive"):
        engine.pathgraph_knn(pts, hp)
    # a grid that puts 800 000 points into one cell would make every query scan all of them (6.4e11 distances): refused before it starts
    many = np.random.default_rng(2).random((800_000, 3)).astype(np.float32)
    with pytest.raises(RuntimeError, match="too many points into too few cells"):
        engine.pathgraph_knn(many, pathgraph_grid(many, 4, 4, (1, 1, 1)))
    nb, knn, dknn, _ = engine.pathgraph_knn(many, pathgraph_grid(many, 4, 4))              # the tool's own grid for them is fine
    assert (knn == 4).mean() > 0.9


@pytest.mark.gpu
def test_knn_vs_compiled_reference_cuda(engine):
    """The reference's own kernels (src/pbsdf.cu compiled unmodified for sm_100a, oracle/_ref/libpbsdf_ref.so) on the same GPU:
    same neighbour SETS (equal-distance members at the k-th place may differ: its quickselect is not a function of the input),
    same knn / dknn counts."""
    from oracle.pyoracle import RefCuda, have_refcuda
    from nori_b200.gpu import pathgraph_grid
    if not have_refcuda():
        pytest.skip("oracle/_ref/libpbsdf_ref.so not built (needs /root/reference at build time)")
    rng = np.random.default_rng(77)
    u = rng.random((150000, 2)).astype(np.float32)
    pts = np.concatenate([np.stack([u[:, 0], 0 * u[:, 0], u[:, 1]], 1), rng.random((150000, 3)).astype(np.float32)]).astype(np.float32)
    ref = RefCuda()
    for k, dk in ((8, 8), (4, 20)):
        hp = pathgraph_grid(pts, k, dk)
        nb, knn, dknn, _ = engine.pathgraph_knn(pts, hp)
        rnb, rknn, rdknn, _ = ref.knn(pts, list(hp.dim), list(hp.cellsize), list(hp.minb), k, dk)
        np.testing.assert_array_equal(dknn, rdknn)
        sel = rng.choice(len(pts), 20000, replace=False)
        idx = np.sort(sel)
        sub_got, sub_ref = nb[idx], rnb[idx]
        for m_of in (np.maximum, np.minimum):
            upto = m_of(knn, dknn)[idx]
            for r, i in enumerate(idx):
                a, b = set(sub_got[r, :upto[r]].tolist()), set(sub_ref[r, :upto[r]].tolist())
                if a != b:
                    oa, ob = np.array(sorted(a - b)), np.array(sorted(b - a))
                    dmax = dist32(pts, i, np.array(sorted(a))).max()
                    assert len(oa) == len(ob) and np.allclose(dist32(pts, i, oa), dmax, rtol=1e-6) and np.allclose(dist32(pts, i, ob), dmax, rtol=1e-6), (i, oa, ob)
        np.testing.assert_array_equal(knn, rknn)
