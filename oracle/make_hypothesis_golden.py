"""Golden vectors for the host layer's statistics header (nori_b200/host/nori/hypothesis.h), produced by the REFERENCE's
own ext/hypothesis/hypothesis.h + cephes.h compiled into oracle/_ref (oracle/ref/glue_hypothesis.cpp):
    python oracle/make_hypothesis_golden.py        ->  tests/golden/hypothesis_ref.json
Seeded random contingency tables (some with empty, tiny and zero-expectation cells so that the pooling rule and the
rejection shortcuts are exercised), t-tests around the decision boundary, and CDF values. Authoring container only."""
import ctypes as C, json, os, re
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libnori_ref.so"))
L.ref_chi2_cdf.restype = C.c_double; L.ref_chi2_cdf.argtypes = [C.c_double, C.c_int]
L.ref_students_t_cdf.restype = C.c_double; L.ref_students_t_cdf.argtypes = [C.c_double, C.c_int]
dp = np.ctypeslib.ndpointer(np.float64, flags="C")
L.ref_chi2_test.argtypes = [C.c_int, dp, dp, C.c_int, C.c_double, C.c_double, C.c_int, C.c_char_p, C.c_int]
L.ref_students_t_test.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int, C.c_double, C.c_int, C.c_char_p, C.c_int]

def numbers(msg):
    out = {"pooled": None, "statistic": None, "dof": None, "pvalue": None, "alpha": None}
    m = re.search(r"Pooled (\d+)", msg); out["pooled"] = int(m.group(1)) if m else None
    m = re.search(r"statistic = ([-\d.e+naif]+) \(d\.o\.f\. = (-?\d+)\)", msg)
    if m: out["statistic"], out["dof"] = float(m.group(1)), int(m.group(2))
    m = re.search(r"p-value = ([-\d.e+naif]+), significance level = ([-\d.e+]+)", msg)
    if m: out["pvalue"], out["alpha"] = float(m.group(1)), float(m.group(2))
    return out

rng = np.random.default_rng(2024)
gold = {"source": "reference ext/hypothesis/hypothesis.h via oracle/_ref (oracle/make_hypothesis_golden.py)", "chi2": [], "ttest": [], "chi2_cdf": [], "t_cdf": []}
for case in range(24):
    n = int(rng.integers(6, 220)); sc = int(rng.integers(2000, 400000))
    p = rng.random(n) ** rng.choice([1, 3, 6]); p /= p.sum()
    if case % 4 == 1: p[rng.integers(0, n, 3)] = 0; p /= p.sum()                 # zero-expectation cells
    expd = p * sc
    q = p.copy()
    if case % 3 == 2: q = q * (1 + 0.08 * rng.normal(size=n)).clip(0.1); q /= q.sum()   # a wrong model: should be rejected
    obs = rng.multinomial(sc, q).astype(np.float64)
    if case == 5: obs[np.argmin(expd + (expd == 0) * 1e9)] += 0; obs[np.nonzero(expd == 0)[0][:1]] += 50   # samples where none may fall
    minexp, sig, nt = float(rng.choice([5, 5, 10])), float(rng.choice([0.01, 0.05])), int(rng.choice([1, 5, 15]))
    buf = C.create_string_buffer(2048)
    ok = L.ref_chi2_test(n, np.ascontiguousarray(obs), np.ascontiguousarray(expd), sc, minexp, sig, nt, buf, 2048)
    gold["chi2"].append({"n": n, "sampleCount": sc, "minExp": minexp, "significance": sig, "numTests": nt, "obs": obs.tolist(),
                         "exp": [float(x) for x in expd], "accept": int(ok), "message": buf.value.decode(), **numbers(buf.value.decode())})
for case in range(16):
    ref = float(rng.uniform(0.05, 2.0)); var = float(rng.uniform(0.01, 4.0)); nsamp = int(rng.choice([1000, 100000, 1000000]))
    mean = ref + float(rng.choice([0, 1, 2, 3, 3.5, 4, 6])) * (var / nsamp) ** 0.5 * float(rng.choice([-1, 1]))
    sig, nt = 0.01, int(rng.choice([1, 5, 6, 15]))
    buf = C.create_string_buffer(2048)
    ok = L.ref_students_t_test(mean, var, ref, nsamp, sig, nt, buf, 2048)
    gold["ttest"].append({"mean": mean, "variance": var, "reference": ref, "sampleCount": nsamp, "significance": sig, "numTests": nt,
                          "accept": int(ok), "message": buf.value.decode(), **numbers(buf.value.decode())})
for x, dof in [(0.5, 1), (3.0, 2), (10.0, 5), (150.0, 199), (250.0, 199), (190.0, 199), (5300.0, 5201), (1e-3, 7), (40.0, 12)]:
    gold["chi2_cdf"].append([x, dof, L.ref_chi2_cdf(x, dof)])
for x, dof in [(0.0, 5), (0.3, 3), (2.5, 10), (-1.7, 30), (3.2, 99999), (0.01, 99999), (4.5, 99999), (12.0, 4), (3.0, 999)]:
    gold["t_cdf"].append([x, dof, L.ref_students_t_cdf(x, dof)])
out = os.path.join(ROOT, "tests", "golden", "hypothesis_ref.json")
json.dump(gold, open(out, "w"), indent=0)
print(out, len(gold["chi2"]), "chi2 tables,", sum(c["accept"] for c in gold["chi2"]), "accepted;", len(gold["ttest"]), "t-tests,", sum(c["accept"] for c in gold["ttest"]), "accepted")
