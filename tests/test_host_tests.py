"""The reference's statistical <test> plugins (src/chi2test.cpp, src/ttest.cpp, src/warptest.cpp CLI mode)
in the C++ host layer (nori_b200/host/src/tests.cpp, nori/hypothesis.h).

CPU: the statistics header against scipy.stats and hand-computed decisions (a probe program compiled
here with g++). GPU: the `nori_b200` binary runs the reference's own test descriptions — the three
microfacet configurations of scenes/pa5/tests/chi2test-microfacet.xml, the angles / golden reference
values of scenes/pa5/tests/ttest-microfacet.xml, a furnace t-test after scenes/pa5/tests/test-furnace.xml —
on device-generated samples, and must reach the reference's verdict ("Passed n/n tests."); a scene
whose reference value is wrong must be rejected."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_BIN = os.path.join(ROOT, "nori_b200", "bin", "nori_b200")
CASES = os.path.join(ROOT, "tests", "golden", "host", "tests")


@pytest.fixture(scope="module")
def probe(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("probe") / "hypothesis_probe")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "nori_b200", "host"),
                           os.path.join(ROOT, "tests", "hypothesis_probe.cpp"), "-o", exe])

    def run(*args):
        return subprocess.run([exe] + [str(a) for a in args], capture_output=True, text=True, check=True).stdout
    return run


def test_distribution_functions_match_scipy(probe):
    from scipy import stats
    for x, dof in [(0.5, 1), (3.0, 2), (10.0, 5), (150.0, 199), (250.0, 199), (190.0, 199), (5300.0, 5201), (1e-3, 7)]:
        assert float(probe("chi2", x, dof)) == pytest.approx(stats.chi2.cdf(x, dof), rel=1e-10, abs=1e-13)
    for x, dof in [(0.0, 5), (0.3, 3), (2.5, 10), (-1.7, 30), (3.2, 99999), (0.01, 99999), (4.5, 99999), (12.0, 4)]:
        assert float(probe("t", x, dof)) == pytest.approx(stats.t.cdf(x, dof), rel=2e-9, abs=1e-13)   # lgamma differences at a = 5e4
    # the t-test's p-value is 2 (1 - cdf): the upper tail must keep its relative accuracy where decisions are made
    for x in (2.0, 3.0, 3.5, 4.0):
        tail = 1.0 - float(probe("t", x, 99999))
        assert tail == pytest.approx(stats.t.sf(x, 99999), rel=1e-6)


def test_quadrature(probe):
    from scipy import integrate, special
    a, b, c = [float(v) for v in probe("simpson").split()]
    assert a == pytest.approx(0.5 * np.sqrt(np.pi) * (special.erf(2.0) + special.erf(1.0)), abs=1e-6)
    want, _ = integrate.dblquad(lambda y, x: np.sin(x) * np.cos(y) + x * y, 0.0, 1.5, 0.0, 0.75)
    assert b == pytest.approx(want, abs=1e-6)
    assert c == pytest.approx(4.0, abs=1e-12)               # Simpson is exact for cubics


def test_chi2_decision_rule_pools_low_expectation_cells(probe):
    """hypothesis.h:143-243: cells sorted by expected frequency; 1.0 and 1.5 are below the minimum (5), 3.5 joins
    because the pool (2.5) is still below it, 18 joins because the pool (6.0 >= 5) is not -> wait: 6.0 >= 5, so 18 is a term."""
    from scipy import stats
    out = probe("chi2test").splitlines()
    assert out[0] == "1"
    assert "Pooled 3 " in out[1]
    chsq = (6 - 6.0) ** 2 / 6.0 + (15 - 18) ** 2 / 18 + (35 - 30) ** 2 / 30 + (48 - 50) ** 2 / 50
    assert f"d.o.f. = 3" in out[2]
    stat = float(out[2].split("=")[1].split("(")[0])
    assert stat == pytest.approx(chsq, rel=1e-5)
    pval = float(out[3].split("p-value = ")[1].split(",")[0])
    assert pval == pytest.approx(stats.chi2.sf(chsq, 3), rel=1e-4)
    alpha = float(out[3].split("significance level = ")[1].rstrip(")"))
    assert alpha == pytest.approx(1 - (1 - 0.01) ** (1 / 3), rel=1e-4)


def test_t_decision_rule(probe):
    ok = probe("ttest", 0.2071, 0.05, 0.207067).splitlines()
    assert ok[0] == "1" and "Accepted" in ok[-1]
    bad = probe("ttest", 0.2171, 0.05, 0.207067).splitlines()          # 14 standard errors off
    assert bad[0] == "0" and "Rejected" in bad[-1]


def test_statistics_header_pinned_to_the_reference(probe, tmp_path):
    """nori/hypothesis.h against tests/golden/hypothesis_ref.json: verdicts, pooled-cell counts, statistics, degrees of
    freedom, p-values and Sidak levels of the REFERENCE's ext/hypothesis/hypothesis.h (compiled into oracle/_ref,
    oracle/make_hypothesis_golden.py) on 24 seeded contingency tables and 16 t-tests around the decision boundary."""
    import json, re
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "hypothesis_ref.json")))

    def numbers(msg):
        out = {"pooled": None, "statistic": None, "dof": None, "pvalue": None, "alpha": None}
        m = re.search(r"Pooled (\d+)", msg); out["pooled"] = int(m.group(1)) if m else None
        m = re.search(r"statistic = ([-\d.e+naif]+) \(d\.o\.f\. = (-?\d+)\)", msg)
        if m: out["statistic"], out["dof"] = float(m.group(1)), int(m.group(2))
        m = re.search(r"p-value = ([-\d.e+naif]+), significance level = ([-\d.e+]+)", msg)
        if m: out["pvalue"], out["alpha"] = float(m.group(1)), float(m.group(2))
        return out

    def same(got, want, case):
        assert got["pooled"] == want["pooled"] and got["dof"] == want["dof"], case
        for k in ("statistic", "pvalue", "alpha"):
            if want[k] is None:
                assert got[k] is None, case
            else:
                assert got[k] == pytest.approx(want[k], rel=2e-5, abs=1e-12), (k, case)   # both print 6 significant digits

    for i, c in enumerate(gold["chi2"]):
        f = tmp_path / f"t{i}.txt"
        f.write_text("%d %d %r %r %d\n%s\n%s\n" % (c["n"], c["sampleCount"], c["minExp"], c["significance"], c["numTests"],
                                                  " ".join(repr(x) for x in c["obs"]), " ".join(repr(x) for x in c["exp"])))
        out = probe("chi2file", f)
        assert int(out.splitlines()[0]) == c["accept"], (i, out, c["message"])
        same(numbers(out), c, i)
        if c["statistic"] is None:                             # rejection shortcuts: same reason given
            assert out.split("\n", 1)[1].split(".")[0][:30] == c["message"].split(".")[0][:30]
    assert any(c["pooled"] for c in gold["chi2"]) and any(not c["accept"] for c in gold["chi2"])
    for i, c in enumerate(gold["ttest"]):
        out = probe("ttestn", repr(c["mean"]), repr(c["variance"]), repr(c["reference"]), c["sampleCount"], c["significance"], c["numTests"])
        assert int(out.splitlines()[0]) == c["accept"], (i, out, c["message"])
        same(numbers(out), c, i)
    for x, dof, want in gold["chi2_cdf"]:
        assert float(probe("chi2", x, dof)) == pytest.approx(want, rel=1e-9, abs=1e-13)
    for x, dof, want in gold["t_cdf"]:
        assert float(probe("t", x, dof)) == pytest.approx(want, rel=2e-9, abs=1e-13)


def test_plugins_are_registered_and_fail_loudly_without_a_device():
    """Parsing works anywhere; activate() needs the device (there is no CPU path)."""
    import ctypes
    try:
        have_gpu = ctypes.CDLL("libcuda.so.1").cuInit(0) == 0
    except OSError:
        have_gpu = False
    if have_gpu:
        pytest.skip("a CUDA device is present: covered by the gpu tests")
    r = subprocess.run([HOST_BIN, os.path.join(CASES, "ttest-microfacet.xml")], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "no CPU path" in r.stderr and "Testing (angle=0): Microfacet[" in r.stdout
    r = subprocess.run([HOST_BIN, "--warptest", "no_such_warp"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "Invalid warp type!" in r.stdout


# ---- GPU: the reference's test descriptions through the binary ---------------------------------------
def run_case(name, tmp_path, timeout=600):
    return subprocess.run([HOST_BIN, os.path.join(CASES, name)], capture_output=True, text=True, timeout=timeout, cwd=str(tmp_path))


@pytest.mark.gpu
def test_chi2test_microfacet(tmp_path):
    r = run_case("chi2test-microfacet.xml", tmp_path)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr
    assert "Passed 15/15 tests." in r.stdout
    assert r.stdout.count("Accepted the null hypothesis") == 15
    assert os.path.exists(tmp_path / "chi2test_15.m")               # hypothesis::chi2_dump


@pytest.mark.gpu
def test_chi2test_diffuse(tmp_path):
    r = run_case("chi2test-diffuse.xml", tmp_path)
    assert r.returncode == 0 and "Passed 3/3 tests." in r.stdout, r.stdout[-2000:] + r.stderr


@pytest.mark.gpu
def test_ttest_microfacet_against_the_reference_values(tmp_path):
    r = run_case("ttest-microfacet.xml", tmp_path)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr
    assert "Passed 5/5 tests." in r.stdout


@pytest.mark.gpu
def test_ttest_furnace_scenes(tmp_path):
    r = run_case("ttest-furnace.xml", tmp_path)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr
    assert "Passed 6/6 tests." in r.stdout


@pytest.mark.gpu
def test_ttest_glass_furnace_on_device_samples(tmp_path):
    """The dielectric pin (closed furnace seen through a glass sphere: 1 / (1 - a) whatever the eta^2 bookkeeping, provided it is
    consistent) on DEVICE-generated samples; tests/test_oracle_pins.py runs the same file through the reference's own t-test on the CPU."""
    r = run_case("ttest-furnace-glass.xml", tmp_path)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr
    assert "Passed 6/6 tests." in r.stdout


@pytest.mark.gpu
def test_ttest_rejects_a_wrong_reference_value(tmp_path):
    r = run_case("ttest-furnace-wrong.xml", tmp_path)
    assert r.returncode != 0 and "***** Rejected *****" in r.stdout and "Passed 0/1 tests." in r.stdout
    assert "Some tests failed" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("args", [["square"], ["tent"], ["disk"], ["uniform_sphere"], ["uniform_hemisphere"], ["cosine_hemisphere"],
                                  ["beckmann", "0.05"], ["beckmann", "0.5"], ["microfacet_brdf", "0.3", "0.4"]])
def test_warptest_command_line_mode(args, tmp_path):
    r = subprocess.run([HOST_BIN, "--warptest"] + args, capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr
    assert "Accepted the null hypothesis" in r.stdout and f"Testing warp {args[0]}" in r.stdout
