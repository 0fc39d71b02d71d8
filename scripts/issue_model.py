"""Per-ray instruction and DRAM figures of the wavefront kernels, from ncu (feeds bench.py's roofline: profiles/issue_model.json).

On the GPU box, the whole render under ncu with four cheap counters (one replay pass per kernel):
    ncu --metrics smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
        --clock-control none -k regex:'k_trace|k_shade|k_bin' --csv --log-file gpurun_out/<tag>_issue.csv python scripts/issue_model.py run <tag>
`run` renders the Cornell box (path_mis, 800x600, 64 spp, 8 Mi path slots) once and writes the render's own counters (closest rays, shadow
rays) to gpurun_out/<tag>_issue_stats.json. Here, afterwards:
    python scripts/issue_model.py parse <tag>        -> profiles/issue_model.json
Sums over ALL launches of the render divided by the render's ray counts: exact per-ray figures, ramp-up and tail included."""
import collections, csv, io, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mode, tag = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "r02")
GO = os.path.join(ROOT, "gpurun_out")
SPP = 64

if mode == "run":
    from nori_b200.gpu import Engine
    from nori_b200.scene import FlatScene
    s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "cbox_mis.npz"))
    e = Engine(0); e.load_scene(s)
    _, st = e.render(sample_first=0, sample_last=SPP, download=False)
    os.makedirs(GO, exist_ok=True)
    json.dump(st, open(os.path.join(GO, tag + "_issue_stats.json"), "w"))
    print(st)
else:
    st = json.load(open(os.path.join(GO, tag + "_issue_stats.json")))
    txt = open(os.path.join(GO, tag + "_issue.csv")).read()
    rows = csv.DictReader(io.StringIO(txt[txt.index('"ID"'):]))
    agg = collections.defaultdict(lambda: collections.defaultdict(float)); launches = collections.Counter(); seen = set()
    for r in rows:
        k = r["Kernel Name"].split("(")[0].replace("void ", "").replace("ng::", "").split("<")[0]
        v = float(r["Metric Value"].replace(",", ""))
        m, u = r["Metric Name"], r["Metric Unit"]
        if m.startswith("dram__bytes"): v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        if m == "gpu__time_duration.sum": v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1)      # -> ms
        agg[k][m] += v
        if (r["ID"], k) not in seen: seen.add((r["ID"], k)); launches[k] += 1
    units = {"k_traceClosest": ("closest ray", st["closest_rays"]), "k_shadeRegen": ("path segment shaded", st["closest_rays"]), "k_traceShadow": ("shadow ray", st["shadow_rays"]),
             "k_binCount": ("closest ray (sort of the slots between bounces)", st["closest_rays"]), "k_binScan": ("closest ray (sort of the slots between bounces)", st["closest_rays"]),
             "k_binScatter": ("closest ray (sort of the slots between bounces)", st["closest_rays"])}
    out = {"source": f"ncu ({tag}): smsp__inst_executed.sum, smsp__thread_inst_executed.sum, dram__bytes_read.sum + dram__bytes_write.sum summed over every launch of one "
                     f"Cornell-box path_mis render (800x600, {SPP} spp, 8 Mi path slots) and divided by that render's own ray counts; raw list profiles/{tag}_issue.csv",
           "sm_count": 148, "render": {k: st[k] for k in ("camera_samples", "closest_rays", "shadow_rays", "iterations")}, "kernels": {}}
    for k, (uname, n) in units.items():
        a = agg[k]
        if not a: continue
        d = {"unit": uname, "launches": launches[k], "warp_inst_per_unit": a["smsp__inst_executed.sum"] / n, "lanes_per_inst": a["smsp__thread_inst_executed.sum"] / a["smsp__inst_executed.sum"],
             "dram_bytes_per_unit": (a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]) / n, "ncu_ms_total": a["gpu__time_duration.sum"]}
        if k == "k_traceClosest": d["warp_inst_per_ray"], d["dram_bytes_per_ray"] = d["warp_inst_per_unit"], d["dram_bytes_per_unit"]
        out["kernels"][k] = d
    json.dump(out, open(os.path.join(ROOT, "profiles", "issue_model.json"), "w"), indent=1)
    import shutil; shutil.copy(os.path.join(GO, tag + "_issue.csv"), os.path.join(ROOT, "profiles", tag + "_issue.csv"))
    print(json.dumps(out, indent=1))
