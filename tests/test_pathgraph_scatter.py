"""Path-graph post-process, second slice (reference src/pbsdf.cu:464-628, 1398-1431, 1497-1526, 1738-1756,
2185-2194, 2317-2368): the tool's own BSDF / density evaluations at a recorded vertex, the marginal densities over the
k-nearest-neighbour graph and the iterated radiance scatter.
CPU: the oracle restatement (oracle/noracle_pathgraph_scatter.h) against vectors produced by the reference's own CUDA code on
the GPU box (tests/golden/pathgraph_scatter_ref.npz, oracle/make_pathgraph_scatter_golden.py) and against closed forms.
GPU: the sm_100a kernels through the C ABI against the oracle (marginal densities and first mc rows to the bit) and against the
compiled reference running on the same GPU; the eLi / eLd fields of nori_gpu_path_vertices against the renderer's own radiance.

Tolerances. The reference is compiled with nvcc's default -fmad=true, the oracle and the product without contraction, and the
reference's scatter adds with float atomics in no fixed order: rows agree to RTOL = 2e-4 of the row's largest component
(+ 1e-6 absolute); the deterministic quantities (evaluations, marginal densities, mc rows) to 2e-5."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

REF_NPZ = os.path.join(GOLDEN, "pathgraph_scatter_ref.npz")
CLUSTER_NPZ = os.path.join(GOLDEN, "pathgraph_cluster_ref.npz")    # ClusterScatterWithDirectOptNR's drivers on the same records (--cluster)
DIRECT_NPZ = os.path.join(GOLDEN, "pathgraph_direct_ref.npz")      # computeDirectScatterAllOnGPU on the same records (--direct)
RTOL_SUM, RTOL_DET = 2e-4, 2e-5


def spoint_dtype():
    from nori_b200.gpu import SPOINT_DTYPE
    return SPOINT_DTYPE


def records(raw):
    return np.ascontiguousarray(raw).view(spoint_dtype()).reshape(-1)


def close_rows(got, want, rtol, what):
    """Rows (n, 3) equal up to rtol of the row's largest component; NaN / inf patterns must coincide."""
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    np.testing.assert_array_equal(np.isfinite(got), np.isfinite(want), err_msg=what + ": non-finite pattern")
    fin = np.isfinite(want).all(-1)
    scale = np.abs(want[fin]).max(-1, keepdims=True)
    err = np.abs(got[fin] - want[fin]) / (scale + 1e-6 / rtol)
    assert err.max(initial=0.0) <= rtol, f"{what}: max relative row error {err.max():.3g} at row {np.nonzero(fin)[0][err.max(-1).argmax()]}"


def unit(v):
    return (v / np.linalg.norm(v, axis=-1, keepdims=True)).astype(np.float32)


def light_records(raw):
    from nori_b200.gpu import LPOINT_DTYPE
    return np.ascontiguousarray(raw).view(LPOINT_DTYPE).reshape(-1)


def direct_case():
    """Records, rows and light records of the direct-light fixture."""
    z, d = np.load(REF_NPZ), np.load(DIRECT_NPZ)
    return records(z["sc_sps"]), light_records(d["dir_lps"]), z["sc_nb"], d["dir_dknn"], d["dir_out"]


# ---- CPU: the oracle ---------------------------------------------------------------------------------------------------
def test_oracle_eval_closed_forms(port):
    """Diffuse vertices: pdf = cos / pi, bsdf = albedo * cos / pi with the cosine CLAMPED into [0, 1] (src/pbsdf.cu:257-264);
    directions below the shading or the geometric side evaluate to 0; an unknown tag evaluates to 0."""
    rng = np.random.default_rng(3)
    n = 2000
    s = np.zeros(n, spoint_dtype())
    s["shN"] = unit(rng.normal(size=(n, 3))); s["geoN"] = s["shN"]
    s["wo"] = unit(s["shN"] + 0.7 * rng.normal(size=(n, 3)))
    wi = unit(rng.normal(size=(n, 3)))
    s["diffuse"] = rng.random((n, 3)); s["rrpdf"] = 1; s["bsdf_type"] = b"d"
    pdf, f = port.pathgraph_eval(s, wi)
    cos_i = (wi * s["shN"]).sum(1); cos_o = (s["wo"] * s["shN"]).sum(1)
    lit = (cos_i > 0) & (cos_o > 0)
    assert lit.sum() > 300 and (~lit).sum() > 300
    np.testing.assert_allclose(pdf[lit], np.minimum(cos_i[lit], 1) / np.pi, rtol=1e-5)
    np.testing.assert_allclose(f[lit], s["diffuse"][lit] * (np.minimum(cos_i[lit], 1) / np.pi)[:, None], rtol=1e-5)
    assert (pdf[~lit] == 0).all() and (f[~lit] == 0).all()
    s["bsdf_type"] = b"x"
    pdf, f = port.pathgraph_eval(s, wi)
    assert (pdf == 0).all() and (f == 0).all()


def test_oracle_eval_vs_reference_vectors(port):
    """pdf_device / bsdfeval_device of the compiled reference (run on the GPU box) at 6 000 synthetic records of all four tags."""
    z = np.load(REF_NPZ)
    s = records(z["eval_sps"])
    pdf, f = port.pathgraph_eval(s, s["wi"])
    tags = {t: int((s["bsdf_type"] == t).sum()) for t in (b"d", b"o", b"c", b"t")}
    assert min(tags.values()) > 100
    want_pdf, want_f = z["eval_pdf"], z["eval_bsdf"]
    np.testing.assert_array_equal(want_pdf == 0, pdf == 0)             # the same branches taken
    np.testing.assert_allclose(pdf, want_pdf, rtol=RTOL_DET, atol=1e-30)
    close_rows(f, want_f, RTOL_DET, "bsdf")
    for t in tags:                                                      # every tag contributes non-zero values
        assert (want_pdf[s["bsdf_type"] == t] > 0).sum() > 20, t


def test_oracle_scatter_vs_reference_vectors(port):
    """computeScatterAllOnGPURecord + allGPUPdfMarginal of the compiled reference on ~7 000 path vertices (whole paths of the
    table scene: diffuse and microfacet) with the reference's own neighbour rows, three iterations."""
    z = np.load(REF_NPZ)
    s = records(z["sc_sps"]); nb, knn = z["sc_nb"], z["sc_knn"]
    assert {b"d", b"o"} <= set(np.unique(s["bsdf_type"]).tolist())
    it = z["sc_blur"].shape[0]
    pm, blur, mc = port.pathgraph_scatter(s, nb, knn, it, sequential=True)
    np.testing.assert_allclose(pm, z["sc_pm"], rtol=RTOL_DET, atol=1e-30)
    assert (z["sc_pm"] > 0).mean() > 0.5
    for i in range(it):
        close_rows(blur[i], z["sc_blur"][i], RTOL_SUM, f"blur[{i}]")
        close_rows(mc[i], z["sc_mc"][i], RTOL_SUM if i else RTOL_DET, f"mc[{i}]")
    assert np.abs(z["sc_blur"][0]).sum() > 0 and np.abs(z["sc_mc"][2]).sum() > 0
    # passes 0 and 1 scatter the same temp = eLi (the reference's update passes its loop index, src/pbsdf.cu:2344-2356)
    close_rows(z["sc_blur"][1], z["sc_blur"][0], RTOL_SUM, "reference blur[1] vs blur[0]")
    np.testing.assert_array_equal(blur[1], blur[0])
    # the lane order of the product's warps changes the marginal sums by rounding only
    pm_lane, _, _ = port.pathgraph_scatter(s, nb, knn, 0, sequential=False)
    np.testing.assert_allclose(pm_lane, pm, rtol=2e-6, atol=1e-30)


def test_oracle_mis_gather_vs_reference_vectors(port):
    """computeMISAllOnGPURecord of the compiled reference (what the `pg` tool's "knn" mode runs: allGPUPdfSum + allGPUMISRadiance)
    on the fixture's records and rows. A gather without atomics: the reference is deterministic here, so everything agrees to the
    rounding of its fused multiply-adds."""
    z = np.load(REF_NPZ)
    s = records(z["sc_sps"]); nb, knn = z["sc_nb"], z["sc_knn"]
    it = z["mis_blur"].shape[0]
    blur, mc = port.pathgraph_mis(s, nb, knn, it, sequential=True)
    for i in range(it):
        close_rows(blur[i], z["mis_blur"][i], RTOL_DET * 2, f"mis blur[{i}]")
        close_rows(mc[i], z["mis_mc"][i], RTOL_DET * 2, f"mis mc[{i}]")
    assert np.abs(z["mis_blur"][0]).sum() > 0 and not np.allclose(z["mis_blur"][0], z["sc_blur"][0])
    # the lane order of the product's warps changes the sums by rounding only
    lblur, lmc = port.pathgraph_mis(s, nb, knn, it, sequential=False)
    close_rows(lblur[it - 1], blur[it - 1], 1e-5, "lane order vs list order")


def test_oracle_direct_scatter_vs_reference_vectors(port):
    """computeDirectScatterAllOnGPU of the compiled reference (src/pbsdf.cu:2370-2395) on the fixture's records and rows with seeded
    light records (a third of them without a light sample, dk counts below the k counts): the restatement agrees to the rounding
    of the reference's atomics; and on self-only rows it is the closed form L_d f(wi_d) / p_light + L_b f(wi) / p(wi) + L_em."""
    s, lp, nb, dknn, want = direct_case()
    assert (dknn <= z_knn()).all() and (dknn < z_knn()).any() and (lp["lightpdf"] == 0).mean() > 0.2
    m, got = port.pathgraph_direct(s, lp, nb, dknn, sequential=True)
    close_rows(got, want, RTOL_SUM, "direct vs reference")
    assert np.isfinite(want).all() and (np.abs(want).sum(1) > 0).mean() > 0.5
    np.testing.assert_allclose(m[:, 0], dknn * lp["lightpdf"], rtol=2e-6)
    m2, got2 = port.pathgraph_direct(s, lp, nb, dknn, sequential=False)
    np.testing.assert_allclose(m2, m, rtol=2e-6, atol=1e-30)
    close_rows(got2, got, 1e-5, "lane order vs list order")
    n = len(s); idx = np.arange(n, dtype=np.int32)[:, None].copy()
    m1, self_only = port.pathgraph_direct(s, lp, idx, np.ones(n, np.int32), sequential=True)
    pdf, f = port.pathgraph_eval(s, s["wi"]); _, fd = port.pathgraph_eval(s, s["wi_d"])
    np.testing.assert_array_equal(m1[:, 1], pdf)
    closed = lp["L_em"].astype(np.float64)
    a = lp["lightpdf"] > 0; closed[a] += lp["L_directsample"][a] * fd[a] / lp["lightpdf"][a, None]
    b = pdf > 0; closed[b] += lp["L_bsdfsample"][b] * f[b] / pdf[b, None]
    close_rows(self_only, closed, 1e-5, "self-only rows vs closed form")
    # ... and the indirect scatter chained to it (computeScatterAllOnGPURecordWithDirectOpt, :2488-2540): the filtered direct light
    # takes the place of eLd after EVERY pass, so pass 0 equals the plain driver's and the later ones do not
    d = np.load(DIRECT_NPZ); z = np.load(REF_NPZ)
    it = d["dir_blur"].shape[0]
    _, got3, blur, mc = port.pathgraph_direct(s, lp, nb, dknn, sequential=True, knn=z["sc_knn"], iterations=it)
    np.testing.assert_array_equal(got3, got)
    for i in range(it):
        close_rows(blur[i], d["dir_blur"][i], RTOL_SUM, f"direct-opt blur[{i}]")
        close_rows(mc[i], d["dir_mc"][i], RTOL_SUM, f"direct-opt mc[{i}]")
    close_rows(d["dir_blur"][0], z["sc_blur"][0], RTOL_SUM, "pass 0 scatters eLi in both drivers")
    assert not np.allclose(d["dir_blur"][1], z["sc_blur"][1], rtol=1e-3) and not np.allclose(d["dir_mc"][0], z["sc_mc"][0], rtol=1e-3)


def cluster_case():
    """Records (groupIdx = cluster or -1), light records, member lists and offsets of the clustered fixture."""
    z, c = np.load(REF_NPZ), np.load(CLUSTER_NPZ)
    s = records(z["sc_sps"]).copy(); s["groupIdx"] = c["cl_of"]
    return s, light_records(c["cl_lps"]), c["cl_members"], c["cl_offsets"], c


def test_oracle_cluster_scatter_vs_reference_vectors(port):
    """SPoints::ClusterScatterWithDirectOptNR's drivers of the compiled reference (ClusterDirect, precomputedMatrixElemtns, ClusterIterations3:
    the `pg` tool's default mode once the clusters exist) on the fixture's records with a seeded clustering — 321 clusters of 1 .. 45
    members, 6 % of the vertices outside every cluster — for 1 and 3 passes. All its sums are float atomics: rows agree to rounding. And
    with every vertex a cluster of its own the direct part is the closed form min(1, in / out) (L_d f(wi_d) / x + L_b f(wi) / y) + L_em."""
    s, lp, members, offsets, c = cluster_case()
    sizes = np.diff(np.append(offsets, len(members)))
    assert sizes.max() > 32 and sizes.min() <= 1 and (s["groupIdx"] == -1).sum() > 100
    for it in (1, 3):
        d, b, m = port.pathgraph_cluster(s, lp, members, offsets, it, sequential=True)
        close_rows(d, c[f"cl_direct_{it}"], RTOL_SUM, f"cluster direct ({it})")
        close_rows(b, c[f"cl_blur_{it}"], RTOL_SUM, f"cluster blur ({it})")
        close_rows(m, c[f"cl_mc_{it}"], RTOL_SUM, f"cluster mc ({it})")
        d2, b2, m2 = port.pathgraph_cluster(s, lp, members, offsets, it, sequential=False)
        close_rows(b2, b, 1e-5, "lane order vs list order")
    assert np.abs(c["cl_blur_3"]).sum() > 0 and not np.allclose(c["cl_blur_3"], c["cl_blur_1"], rtol=1e-3)
    close_rows(c["cl_direct_3"], c["cl_direct_1"], 1e-5, "two runs of the reference's ClusterDirect (atomics: equal to rounding only)")
    loose = s["groupIdx"] == -1
    np.testing.assert_allclose(c["cl_direct_1"][loose], s["eLd"][loose] + lp["L_em"][loose], rtol=1e-6)
    n = len(s); one = s.copy(); one["groupIdx"] = np.arange(n); idx = np.arange(n, dtype=np.int32)
    d1, b1, m1 = port.pathgraph_cluster(one, lp, idx, idx, 1, sequential=True)
    pdf, f = port.pathgraph_eval(s, s["wi"]); pdfd, fd = port.pathgraph_eval(s, s["wi_d"])
    x = np.where((s["wi_d"] * s["shN"]).sum(1) > 0, lp["lightpdf"], 0) + pdfd
    y = np.where((s["wi"] * s["shN"]).sum(1) > 0, lp["bsdfpdf"], 0) + pdf
    raw = np.zeros((n, 3))
    a = x > 0; raw[a] += lp["L_directsample"][a] * fd[a] / x[a, None]
    b = y > 0; raw[b] += lp["L_bsdfsample"][b] * f[b] / y[b, None]
    took = lp["L_directsample"].astype(np.float64) + lp["L_bsdfsample"]
    want = raw * np.where(took < raw, took / np.maximum(raw, 1e-300), 1.0) + lp["L_em"]
    close_rows(d1, want, 1e-5, "singleton clusters vs closed form")


def test_oracle_other_cluster_modes_vs_reference_vectors(port):
    """The `pg` tool's cluster modes without direct light on the same clustered fixture: "n" = SPoints::ClusterScatter2 (stored elements + MX,
    ClusterIterations) and "t" / "l" = SPoints::ClusterScatter (allGPUClusterScatterRadiance). They differ in the order of the products, in
    whether a sender reaches its own successor, and in what a sender without a usable density does; both record every pass."""
    s, lp, members, offsets, c = cluster_case()
    for order in ("elements", "scatter"):
        b, m = port.pathgraph_cluster_scatter(s, members, offsets, 3, order=order, sequential=True)
        for i in range(3):
            close_rows(b[i], c[f"cl_{order}_blur"][i], RTOL_SUM, f"{order} blur[{i}]")
            close_rows(m[i], c[f"cl_{order}_mc"][i], RTOL_SUM, f"{order} mc[{i}]")
        b2, m2 = port.pathgraph_cluster_scatter(s, members, offsets, 3, order=order, sequential=False)
        close_rows(b2[2], b[2], 1e-5, "lane order vs list order")
    same = (s["nidx"] > 0) & (s["groupIdx"] >= 0) & (np.roll(s["groupIdx"], -1) == s["groupIdx"])
    assert same.sum() >= 250                                                    # senders whose own successor sits in their cluster: the fixture forces 300
    # ... yet the two modes agree: on an opaque surface the successor sees its predecessor's direction arrive from BELOW (it points into the
    # surface it hits), so its own BSDF is zero for it, and skipping it or not changes nothing. The exception only matters behind glass.
    close_rows(c["cl_scatter_blur"][0], c["cl_elements_blur"][0], RTOL_SUM, "mode t vs mode n, pass 0")
    close_rows(c["cl_elements_blur"][0], c["cl_blur_1"], RTOL_SUM, "pass 0 of mode n = the one pass of the default mode (temp = eLi in both)")


def z_knn():
    return np.load(REF_NPZ)["sc_knn"]


def test_oracle_mc_row_is_the_paths_own_estimate(port):
    """lastRun on diffuse vertices: bsdf / pdf = albedo, so mc[0][i] = eLi[i + 1] * albedo / rrpdf — with records whose eLi obeys
    the path recursion eLi[i] = eLd[i] + albedo * eLi[i + 1] / rrpdf this reads mc[0] + eLd = eLi."""
    rng = np.random.default_rng(9)
    paths, length = 300, 5
    n = paths * length
    s = np.zeros(n, spoint_dtype())
    s["shN"] = unit(rng.normal(size=(n, 3))); s["geoN"] = s["shN"]
    s["wo"] = unit(s["shN"] + 0.5 * rng.normal(size=(n, 3))); s["wi"] = unit(s["shN"] + 0.5 * rng.normal(size=(n, 3)))
    s["diffuse"] = 0.2 + 0.6 * rng.random((n, 3)); s["rrpdf"] = rng.choice([1.0, 0.9], size=n); s["bsdf_type"] = b"d"
    s["eLd"] = rng.random((n, 3)); s["pos"] = rng.random((n, 3))
    idx = np.arange(n)
    inner = idx % length != length - 1
    s["nidx"][inner] = idx[inner] + 1
    s["wi"][~inner] = 0; s["rrpdf"][~inner] = 0
    eli = s["eLd"].astype(np.float64).copy()
    for j in range(length - 2, -1, -1):
        rows = idx[idx % length == j]
        eli[rows] += s["diffuse"][rows] * eli[rows + 1] / s["rrpdf"][rows, None]
    s["eLi"] = eli
    nb = np.tile(idx[:, None], (1, 4)).astype(np.int32); nb[:, 1:] = (idx[:, None] + np.array([7, 13, 29])) % n
    knn = np.full(n, 4, np.int32)
    pm, blur, mc = port.pathgraph_scatter(s, nb, knn, 1)
    lit = inner & ((s["wi"] * s["shN"]).sum(1) > 0) & ((s["wo"] * s["shN"]).sum(1) > 0)
    assert lit.sum() > n // 2
    np.testing.assert_allclose(mc[0][lit] + s["eLd"][lit], s["eLi"][lit], rtol=1e-5)
    assert (mc[0][~inner] == 0).all()


# ---- GPU: the kernels ---------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_scatter_vs_oracle_on_reference_vectors(engine, port):
    """The fixture's records through nori_gpu_pathgraph_scatter: marginal densities and the first mc rows equal the oracle's
    (same operations in the same order) to the bit; everything that went through the scatter's atomics to rounding; and all of
    it agrees with what the compiled reference produced on the GPU box."""
    z = np.load(REF_NPZ)
    s = records(z["sc_sps"]); nb, knn = z["sc_nb"], z["sc_knn"]
    it = z["sc_blur"].shape[0]
    pm, blur, mc, ms = engine.pathgraph_scatter(s, nb, knn, it)
    wpm, wblur, wmc = port.pathgraph_scatter(s, nb, knn, it, sequential=False)
    np.testing.assert_array_equal(pm.view(np.uint32), wpm.view(np.uint32))
    np.testing.assert_array_equal(mc[0].view(np.uint32), wmc[0].view(np.uint32))
    for i in range(it):
        close_rows(blur[i], wblur[i], RTOL_SUM, f"blur[{i}] vs oracle")
        close_rows(mc[i], wmc[i], RTOL_SUM, f"mc[{i}] vs oracle")
        close_rows(blur[i], z["sc_blur"][i], RTOL_SUM, f"blur[{i}] vs reference")
        close_rows(mc[i], z["sc_mc"][i], RTOL_SUM, f"mc[{i}] vs reference")
    np.testing.assert_allclose(pm, z["sc_pm"], rtol=RTOL_DET, atol=1e-30)


@pytest.mark.gpu
def test_gpu_mis_gather_vs_oracle_and_reference(engine, port):
    """nori_gpu_pathgraph_mis on the fixture's records: no atomics anywhere, so EVERY row of every pass equals the oracle's
    (same operations, same summation order) to the bit — also on synthetic records of all four tags with rows longer than a
    warp —, and agrees with the compiled reference's vectors to rounding."""
    z = np.load(REF_NPZ)
    s = records(z["sc_sps"]); nb, knn = z["sc_nb"], z["sc_knn"]
    it = z["mis_blur"].shape[0]
    blur, mc, ms = engine.pathgraph_mis(s, nb, knn, it)
    wblur, wmc = port.pathgraph_mis(s, nb, knn, it, sequential=False)
    np.testing.assert_array_equal(blur.view(np.uint32), wblur.view(np.uint32))
    np.testing.assert_array_equal(mc.view(np.uint32), wmc.view(np.uint32))
    for i in range(it):
        close_rows(blur[i], z["mis_blur"][i], RTOL_DET * 2, f"mis blur[{i}] vs reference")
        close_rows(mc[i], z["mis_mc"][i], RTOL_DET * 2, f"mis mc[{i}] vs reference")
    t = records(z["eval_sps"])[:3000].copy()
    n = len(t); idx = np.arange(n)
    inner = idx % 4 != 3
    t["nidx"] = 0; t["nidx"][inner] = idx[inner] + 1
    rng = np.random.default_rng(5)
    t["eLi"] = rng.random((n, 3)); t["eLd"] = rng.random((n, 3))
    k = 45
    tnb = rng.integers(0, n, size=(n, k)).astype(np.int32); tnb[:, 0] = idx
    tknn = rng.integers(1, k + 1, size=n).astype(np.int32)
    blur, mc, _ = engine.pathgraph_mis(t, tnb, tknn, 2)
    wblur, wmc = port.pathgraph_mis(t, tnb, tknn, 2, sequential=False)
    np.testing.assert_array_equal(blur.view(np.uint32), wblur.view(np.uint32))
    np.testing.assert_array_equal(mc.view(np.uint32), wmc.view(np.uint32))
    assert np.isfinite(blur).all() and np.abs(blur[1]).sum() > 0


@pytest.mark.gpu
def test_gpu_scatter_all_tags_vs_oracle(engine, port):
    """Synthetic records of all four tags (the fixture's eval set, given paths of four vertices and random neighbour rows): the
    kernels' evaluations agree with the oracle's to the bit wherever no atomic is involved."""
    z = np.load(REF_NPZ)
    s = records(z["eval_sps"]).copy()
    n = len(s); idx = np.arange(n)
    inner = idx % 4 != 3
    s["nidx"] = 0; s["nidx"][inner] = idx[inner] + 1
    rng = np.random.default_rng(21)
    s["eLi"] = rng.random((n, 3)); s["eLd"] = rng.random((n, 3))
    k = 40                                                              # more than one lane round
    nb = rng.integers(0, n, size=(n, k)).astype(np.int32); nb[:, 0] = idx
    knn = rng.integers(1, k + 1, size=n).astype(np.int32)
    pm, blur, mc, _ = engine.pathgraph_scatter(s, nb, knn, 2)
    wpm, wblur, wmc = port.pathgraph_scatter(s, nb, knn, 2, sequential=False)
    np.testing.assert_array_equal(pm.view(np.uint32), wpm.view(np.uint32))
    np.testing.assert_array_equal(mc[0].view(np.uint32), wmc[0].view(np.uint32))
    for i in range(2):
        close_rows(blur[i], wblur[i], RTOL_SUM, f"blur[{i}]")
    assert (pm > 0).mean() > 0.3 and np.abs(blur[0]).sum() > 0


@pytest.mark.gpu
def test_gpu_scatter_vs_compiled_reference_cuda(gpu_scene, scenes, port):
    """The reference's own kernels (src/pbsdf.cu compiled unmodified, oracle/_ref/libpbsdf_ref.so) on the same GPU, same records
    (path vertices of the veach scene: diffuse and microfacet plates under four lights), same neighbour rows."""
    from oracle.pyoracle import RefCuda, have_refcuda
    from nori_b200.gpu import pathgraph_grid
    if not have_refcuda():
        pytest.skip("oracle/_ref/libpbsdf_ref.so not built (needs /root/reference at build time)")
    e = gpu_scene("veach_mis"); s = scenes("veach_mis")
    e.set_camera(s.sample_to_camera, s.camera_to_world, 256, 160, s.near_clip, s.far_clip)
    try:
        v = e.path_vertices(sample_first=0, sample_last=2)
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)
    assert len(v) > 100_000 and set(np.unique(v["bsdf_type"]).tolist()) <= {b"d", b"o"}
    hp = pathgraph_grid(v, 16, 16)
    nb, knn, dknn, _ = e.pathgraph_knn(v, hp)
    pm, blur, mc, ms = e.pathgraph_scatter(v, nb, knn, 3)
    rpm, rblur, rmc, sec = RefCuda().scatter(v, nb, knn, 3)
    np.testing.assert_allclose(pm, rpm, rtol=RTOL_DET, atol=1e-30)
    for i in range(3):
        close_rows(blur[i], rblur[i], RTOL_SUM, f"blur[{i}]")
        close_rows(mc[i], rmc[i], RTOL_SUM, f"mc[{i}]")
    assert np.abs(rblur[2]).sum() > 0 and (rpm > 0).mean() > 0.3
    mblur, mmc, _ = e.pathgraph_mis(v, nb, knn, 2)
    rmblur, rmmc, _ = RefCuda().mis(v, nb, knn, 2)
    for i in range(2):
        close_rows(mblur[i], rmblur[i], RTOL_DET * 2, f"mis blur[{i}]")
        close_rows(mmc[i], rmmc[i], RTOL_DET * 2, f"mis mc[{i}]")
    fpm, fblur, fmc, _ = e.pathgraph_filter(v, hp, 2, mode="mis")
    np.testing.assert_array_equal(fblur.view(np.uint32), mblur.view(np.uint32))       # same rows, same kernels: deterministic


@pytest.mark.gpu
def test_gpu_path_vertices_carry_the_paths_radiance(gpu_scene, scenes, port):
    """eLi / eLd of nori_gpu_path_vertices: eLi of a path's first vertex is the radiance nori_gpu_sample_radiance returns for
    that camera sample; along a path eLi[i] = eLd[i] + w_i * eLi[i + 1] with w_i = the sampling weight of the continuation, which
    at diffuse vertices is albedo / rrpdf — the same statement lastRun makes: mc[0] + eLd = eLi there."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    e.set_camera(s.sample_to_camera, s.camera_to_world, 160, 120, s.near_clip, s.far_clip)
    try:
        v = e.path_vertices(sample_first=0, sample_last=2)
        first = np.concatenate([[True], v["groupIdx"][1:] != v["groupIdx"][:-1]])
        g = v["groupIdx"][first]
        rad, _ = e.sample_radiance(g // 2, g % 2)
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)
    np.testing.assert_allclose(v["eLi"][first], rad, rtol=2e-4, atol=1e-5)
    assert (v["eLd"] >= 0).all() and (np.abs(v["eLi"]).sum(1) > 0).mean() > 0.5
    idx = np.arange(len(v))
    nb = idx[:, None].astype(np.int32).copy(); knn = np.ones(len(v), np.int32)
    pm, blur, mc, _ = e.pathgraph_scatter(v, nb, knn, 1)
    dif = (v["bsdf_type"] == b"d") & (v["nidx"] > 0) & (np.abs(mc[0]).sum(1) > 0)
    assert dif.sum() > 10_000
    np.testing.assert_allclose(mc[0][dif] + v["eLd"][dif], v["eLi"][dif], rtol=5e-4, atol=1e-5)


@pytest.mark.gpu
def test_gpu_direct_scatter_vs_oracle_and_reference(engine, port):
    """nori_gpu_pathgraph_direct on the direct-light fixture: equal to the oracle (same evaluations, same marginal sums; the scatter's
    atomics add in another order) and to the compiled reference's vectors to rounding; rows of one vertex — no two additions to the
    same address — to the bit, also for synthetic records of all four tags with rows longer than a warp; bad rows are refused."""
    s, lp, nb, dknn, want = direct_case()
    got, ms = engine.pathgraph_direct(s, lp, nb, dknn)
    _, ora = port.pathgraph_direct(s, lp, nb, dknn, sequential=False)
    close_rows(got, ora, RTOL_SUM, "direct vs oracle")
    close_rows(got, want, RTOL_SUM, "direct vs reference")
    n = len(s); idx = np.arange(n, dtype=np.int32)[:, None].copy(); one = np.ones(n, np.int32)
    got1, _ = engine.pathgraph_direct(s, lp, idx, one)
    _, ora1 = port.pathgraph_direct(s, lp, idx, one, sequential=False)
    np.testing.assert_array_equal(got1.view(np.uint32), ora1.view(np.uint32))
    d = np.load(DIRECT_NPZ); knn = np.load(REF_NPZ)["sc_knn"]; it = d["dir_blur"].shape[0]
    got3, blur, mc, _ = engine.pathgraph_direct(s, lp, nb, dknn, knn=knn, iterations=it)     # nori_gpu_pathgraph_scatter_direct
    _, ora3, oblur, omc = port.pathgraph_direct(s, lp, nb, dknn, sequential=False, knn=knn, iterations=it)
    close_rows(got3, ora, RTOL_SUM, "chained: direct vs oracle")
    for i in range(it):
        close_rows(blur[i], oblur[i], RTOL_SUM, f"direct-opt blur[{i}] vs oracle")
        close_rows(mc[i], omc[i], RTOL_SUM, f"direct-opt mc[{i}] vs oracle")
        close_rows(blur[i], d["dir_blur"][i], RTOL_SUM, f"direct-opt blur[{i}] vs reference")
        close_rows(mc[i], d["dir_mc"][i], RTOL_SUM, f"direct-opt mc[{i}] vs reference")
    # all four tags, long rows, every vertex scattering to distinct targets (a permutation per column: one addition per column and address)
    t = records(np.load(REF_NPZ)["eval_sps"])[:3000].copy(); n = len(t)
    rng = np.random.default_rng(9)
    t["wi_d"] = unit(t["shN"] + rng.normal(size=(n, 3)))
    from nori_b200.gpu import LPOINT_DTYPE
    tl = np.zeros(n, LPOINT_DTYPE)
    tl["L_directsample"] = rng.random((n, 3)); tl["L_bsdfsample"] = rng.random((n, 3)); tl["L_em"] = rng.random((n, 3))
    tl["lightpdf"] = rng.random(n) * (rng.random(n) < 0.7)
    k = 40
    tnb = np.stack([rng.permutation(n) for _ in range(k)], 1).astype(np.int32)
    tk = rng.integers(0, k + 1, size=n).astype(np.int32)
    gt, _ = engine.pathgraph_direct(t, tl, tnb, tk)
    _, ot = port.pathgraph_direct(t, tl, tnb, tk, sequential=False)
    close_rows(gt, ot, 2e-5, "all tags vs oracle")
    assert np.isfinite(gt).all() and np.abs(gt).sum() > 0
    from nori_b200.gpu import NoriGpuError
    bad = tnb.copy(); bad[7, 0] = n
    with pytest.raises(NoriGpuError, match="neighbour index"):
        engine.pathgraph_direct(t, tl, bad, np.maximum(tk, 1))
    with pytest.raises(NoriGpuError, match="dknn"):
        engine.pathgraph_direct(t, tl, tnb, tk + k)


@pytest.mark.gpu
def test_gpu_cluster_scatter_vs_oracle_and_reference(engine, port):
    """nori_gpu_pathgraph_cluster on the clustered fixture. A warp owns a cluster and takes every sum in a fixed order, so the device
    equals the oracle (same evaluations, same order) TO THE BIT on all three results, for clusters below and above a warp's width and
    for the vertices outside every cluster; and it agrees with the compiled reference's vectors to the rounding of that side's atomics.
    Lists that are not a partition of the vertices naming a cluster are refused."""
    s, lp, members, offsets, c = cluster_case()
    for it in (1, 3):
        d, b, m, ms = engine.pathgraph_cluster(s, lp, members, offsets, it)
        od, ob, om = port.pathgraph_cluster(s, lp, members, offsets, it, sequential=False)
        np.testing.assert_array_equal(d.view(np.uint32), od.view(np.uint32))
        np.testing.assert_array_equal(b.view(np.uint32), ob.view(np.uint32))
        np.testing.assert_array_equal(m.view(np.uint32), om.view(np.uint32))
        close_rows(d, c[f"cl_direct_{it}"], RTOL_SUM, f"cluster direct ({it}) vs reference")
        close_rows(b, c[f"cl_blur_{it}"], RTOL_SUM, f"cluster blur ({it}) vs reference")
        close_rows(m, c[f"cl_mc_{it}"], RTOL_SUM, f"cluster mc ({it}) vs reference")
    for order in ("elements", "scatter"):                     # the tool's modes "n" and "t" / "l": every pass recorded, equal to the oracle to the bit
        b, m, _ = engine.pathgraph_cluster_scatter(s, members, offsets, 3, order=order)
        ob, om = port.pathgraph_cluster_scatter(s, members, offsets, 3, order=order, sequential=False)
        np.testing.assert_array_equal(b.view(np.uint32), ob.view(np.uint32))
        np.testing.assert_array_equal(m.view(np.uint32), om.view(np.uint32))
        for i in range(3):
            close_rows(b[i], c[f"cl_{order}_blur"][i], RTOL_SUM, f"{order} blur[{i}] vs reference")
            close_rows(m[i], c[f"cl_{order}_mc"][i], RTOL_SUM, f"{order} mc[{i}] vs reference")
    # all four tags, cluster_of passed beside the records; no clusters at all = every vertex keeps its own estimate
    t = records(np.load(REF_NPZ)["eval_sps"])[:3000].copy(); n = len(t); idx = np.arange(n)
    rng = np.random.default_rng(13)
    inner = idx % 4 != 3
    t["nidx"] = 0; t["nidx"][inner] = idx[inner] + 1
    t["wi_d"] = unit(t["shN"] + rng.normal(size=(n, 3))); t["eLi"] = rng.random((n, 3)); t["eLd"] = rng.random((n, 3))
    from nori_b200.gpu import LPOINT_DTYPE, NoriGpuError
    tl = np.zeros(n, LPOINT_DTYPE)
    tl["L_directsample"] = rng.random((n, 3)); tl["L_bsdfsample"] = rng.random((n, 3)); tl["L_em"] = rng.random((n, 3))
    tl["lightpdf"] = rng.random(n) * (rng.random(n) < 0.7); tl["bsdfpdf"] = rng.random(n)
    perm = np.concatenate([idx[:600], 600 + rng.permutation(n - 600)]).astype(np.int32)[:2800]   # the first clusters are runs of consecutive vertices:
                                                                                                   # a sender's own successor sits in its cluster
    offs = np.concatenate([[0], np.cumsum(rng.integers(1, 70, size=200))]); offs = offs[offs < len(perm)].astype(np.int32)
    of = np.full(n, -1, np.int32)
    for ci, (a, b_) in enumerate(zip(offs, np.append(offs[1:], len(perm)))):
        of[perm[a:b_]] = ci
    d, b, m, _ = engine.pathgraph_cluster(t, tl, perm, offs, 2, cluster_of=of)
    tt = t.copy(); tt["groupIdx"] = of
    od, ob, om = port.pathgraph_cluster(tt, tl, perm, offs, 2, sequential=False)
    for got, want in ((d, od), (b, ob), (m, om)):
        np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.isfinite(b).all() and np.abs(b).sum() > 0
    modes = {}
    for order in ("elements", "scatter"):
        b, m, _ = engine.pathgraph_cluster_scatter(t, perm, offs, 2, order=order, cluster_of=of)
        ob, om = port.pathgraph_cluster_scatter(tt, perm, offs, 2, order=order, sequential=False)
        np.testing.assert_array_equal(b.view(np.uint32), ob.view(np.uint32))
        np.testing.assert_array_equal(m.view(np.uint32), om.view(np.uint32))
        modes[order] = b
    # with random directions a successor does receive from its predecessor: here allGPUClusterScatterRadiance's exception shows ...
    assert not np.allclose(modes["elements"][0], modes["scatter"][0], rtol=1e-3)
    # ... and is the compiled reference's, run live on the same records (without the glass ones: it prints a line per dielectric evaluation)
    from oracle.pyoracle import RefCuda, have_refcuda
    if have_refcuda():
        keep = np.nonzero(tt["bsdf_type"] != b"t")[0]
        renum = np.full(n, -1, np.int64); renum[keep] = np.arange(len(keep))
        u = tt[keep].copy()
        nxt = (u["nidx"] > 0) & (np.append(keep[1:], -1) == keep + 1)                 # successor kept too
        u["nidx"] = np.where(nxt, np.arange(len(keep)) + 1, 0)
        lists_u = renum[perm[np.isin(perm, keep)]].astype(np.int32)
        offs_u = np.searchsorted(u["groupIdx"][lists_u], np.arange(len(offs))).astype(np.int32)
        for order in ("elements", "scatter"):
            rb, rm, _ = RefCuda().cluster_scatter(u, lists_u, offs_u, 2, order=order)
            gb, gm, _ = engine.pathgraph_cluster_scatter(u, lists_u, offs_u, 2, order=order)
            close_rows(gb[1], rb[1], RTOL_SUM, f"{order} blur vs live reference"); close_rows(gm[1], rm[1], RTOL_SUM, f"{order} mc vs live reference")
    d0, b0, m0, _ = engine.pathgraph_cluster(t, tl, [], [], 1, cluster_of=np.full(n, -1, np.int32))
    np.testing.assert_array_equal(d0, t["eLd"] + tl["L_em"])
    bad = of.copy(); bad[perm[5]] = (of[perm[5]] + 1) % len(offs)
    with pytest.raises(NoriGpuError, match="not the cluster that lists it"):
        engine.pathgraph_cluster(t, tl, perm, offs, 1, cluster_of=bad)
    twice = perm.copy(); twice[1] = twice[0]
    with pytest.raises(NoriGpuError, match="listed twice|not the cluster"):
        engine.pathgraph_cluster(t, tl, twice, offs, 1, cluster_of=of)
    names = of.copy(); names[np.nonzero(of == -1)[0][0]] = 0
    with pytest.raises(NoriGpuError, match="does not list it"):
        engine.pathgraph_cluster(t, tl, perm, offs, 1, cluster_of=names)
    with pytest.raises(NoriGpuError, match="offsets"):
        engine.pathgraph_cluster(t, tl, perm, offs[::-1].copy(), 1, cluster_of=of)
    with pytest.raises(NoriGpuError, match="iterations"):
        engine.pathgraph_cluster(t, tl, perm, offs, 0, cluster_of=of)
    big = records(np.load(REF_NPZ)["eval_sps"])                           # one cluster of 6 000 is fine, one of more than 8 192 is refused
    big = np.concatenate([big, big])[:9000].copy(); big["nidx"] = 0
    bl = np.zeros(len(big), LPOINT_DTYPE)
    with pytest.raises(NoriGpuError, match="more than 8192 members"):
        engine.pathgraph_cluster(big, bl, np.arange(9000, dtype=np.int32), [0], 1, cluster_of=np.zeros(9000, np.int32))


@pytest.mark.gpu
def test_gpu_path_vertices_light_records(gpu_scene, scenes, port):
    """nori_gpu_path_vertices_lights: the LPoint of every vertex beside its SPoint. bsdfpdf = SPoint::pdf; L_bsdfsample = L_em of
    the path's next vertex; at a path's first vertex off the emitters the estimator's own credit eLd is the light sample under the
    balance heuristic, L_directsample f(wi_d) / (lightpdf + p(wi_d)), with the tool's evaluations f and p; on an emitter it is L_em;
    and the records without lights are the ones nori_gpu_path_vertices returns."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    e.set_camera(s.sample_to_camera, s.camera_to_world, 160, 120, s.near_clip, s.far_clip)
    try:
        v, lp = e.path_vertices(sample_first=0, sample_last=2, lights=True)
        v0 = e.path_vertices(sample_first=0, sample_last=2)
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)
    np.testing.assert_array_equal(v.view(np.uint8), v0.view(np.uint8))
    assert len(lp) == len(v) > 30_000
    np.testing.assert_array_equal(lp["bsdfpdf"], v["pdf"])
    inner = np.nonzero(v["nidx"] > 0)[0]
    np.testing.assert_array_equal(v["nidx"][inner], inner + 1)
    np.testing.assert_array_equal(lp["L_bsdfsample"][inner], lp["L_em"][inner + 1])
    assert (lp["L_bsdfsample"][v["nidx"] == 0] == 0).all()
    first = np.concatenate([[True], v["groupIdx"][1:] != v["groupIdx"][:-1]])
    em = first & (lp["L_em"].sum(1) > 0)
    assert em.sum() > 50
    np.testing.assert_allclose(v["eLd"][em], lp["L_em"][em], rtol=1e-6)
    lit = first & ~em & (v["bsdf_type"] == b"d") & (lp["L_directsample"].sum(1) > 0)
    assert lit.sum() > 5_000 and (lp["lightpdf"][lit] > 0).all()
    p, f = port.pathgraph_eval(v[lit], v["wi_d"][lit])
    np.testing.assert_allclose(v["eLd"][lit], lp["L_directsample"][lit] * f / (lp["lightpdf"][lit] + p)[:, None], rtol=3e-4, atol=1e-6)
    shadowed = first & ~em & (lp["L_directsample"].sum(1) == 0)
    assert shadowed.sum() > 100 and np.abs(v["eLd"][shadowed]).max() == 0


@pytest.mark.gpu
def test_gpu_filter_equals_knn_then_scatter(gpu_scene, scenes):
    """nori_gpu_pathgraph_filter = the two stages back to back with the rows kept on the device: same marginal densities and mc[0]
    rows to the bit (no atomics involved), the scattered rows to rounding."""
    from nori_b200.gpu import pathgraph_grid
    e = gpu_scene("table_mis"); s = scenes("table_mis")
    e.set_camera(s.sample_to_camera, s.camera_to_world, 200, 150, s.near_clip, s.far_clip)
    try:
        v = e.path_vertices(sample_first=0, sample_last=1)
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)
    hp = pathgraph_grid(v, 10, 14)                                      # k < dk: the scatter must read the k consumer's counts
    nb, knn, dknn, _ = e.pathgraph_knn(v, hp)
    pm, blur, mc, _ = e.pathgraph_scatter(v, nb, knn, 2)
    fpm, fblur, fmc, ms = e.pathgraph_filter(v, hp, 2)
    np.testing.assert_array_equal(fpm.view(np.uint32), pm.view(np.uint32))
    np.testing.assert_array_equal(fmc[0].view(np.uint32), mc[0].view(np.uint32))
    for i in range(2):
        close_rows(fblur[i], blur[i], RTOL_SUM, f"blur[{i}]")
    assert ms[0] > 0 and ms[1] > 0


@pytest.mark.gpu
def test_gpu_pathgraph_render_equals_the_stages_one_by_one(gpu_scene, scenes):
    """nori_gpu_pathgraph_render keeps records, rows and results on the device and returns the five images only: the same images
    as composing them on the host (src/cluster.cpp:23-107) from nori_gpu_path_vertices + nori_gpu_pathgraph_filter, to the bit in
    MIS mode (no atomics anywhere), to rounding in the blurred image of scatter mode; two samples per pixel are averaged in
    sample order."""
    from nori_b200.gpu import pathgraph_grid
    e = gpu_scene("veach_mis"); s = scenes("veach_mis")
    W, H, spp, k = 192, 128, 2, 10
    e.set_camera(s.sample_to_camera, s.camera_to_world, W, H, s.near_clip, s.far_clip)
    try:
        v = e.path_vertices(sample_first=0, sample_last=spp)
        hp = pathgraph_grid(v, k, k)
        for mode in ("mis", "scatter"):
            _, blur, mc, _ = e.pathgraph_filter(v, hp, 2, mode=mode)
            img, st = e.pathgraph_render(k, 2, mode=mode, sample_first=0, sample_last=spp)
            assert st["vertices"] == len(v) and st["dim"] == list(hp.dim) and st["ms_knn"] > 0
            first = np.nonzero(np.concatenate([[True], v["groupIdx"][1:] != v["groupIdx"][:-1]]))[0]
            g = v["groupIdx"][first]
            def image(rows):
                out = np.zeros((H * W, 3), np.float32)
                for kk in range(spp):                       # sample order, float32 accumulation as the kernel's
                    sel = g % spp == kk
                    out[g[sel] // spp] += rows[first[sel]] * np.float32(1.0 / spp)
                return out.reshape(H, W, 3)
            np.testing.assert_array_equal(img["init"], image(v["eLd"]))
            np.testing.assert_array_equal(img["pt"], image(v["eLi"] - v["eLd"]))
            if mode == "mis":
                np.testing.assert_array_equal(img["indirect"], image(mc[1]))
                np.testing.assert_array_equal(img["blur"], image(blur[1]))
                np.testing.assert_array_equal(img["full"], image(v["eLd"] + mc[1]))
            else:
                close_rows(img["blur"].reshape(-1, 3), image(blur[1]).reshape(-1, 3), RTOL_SUM, "blur image")
                close_rows(img["indirect"].reshape(-1, 3), image(mc[1]).reshape(-1, 3), RTOL_SUM, "indirect image")
            assert img["full"].mean() > 0 and np.isfinite(img["full"]).all()
    finally:
        e.set_camera(s.sample_to_camera, s.camera_to_world, s.width, s.height, s.near_clip, s.far_clip)


@pytest.mark.gpu
def test_gpu_scatter_rejects_bad_rows(engine):
    z = np.load(REF_NPZ)
    s = records(z["sc_sps"])[:64].copy(); s["nidx"][63] = 0
    nb = np.zeros((64, 4), np.int32); knn = np.full(64, 4, np.int32)
    nb[5, 2] = 64
    with pytest.raises(RuntimeError, match="neighbour index outside"):
        engine.pathgraph_scatter(s, nb, knn, 1)
    nb[5, 2] = 0; knn[7] = 5
    with pytest.raises(RuntimeError, match="knn"):
        engine.pathgraph_scatter(s, nb, knn, 1)
