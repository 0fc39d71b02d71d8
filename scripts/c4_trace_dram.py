"""DRAM bytes per ray of the C4 traversal launches, from ncu (feeds bench.py: profiles/c4_trace_dram.json).
On the GPU box:
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum \
        --clock-control none -k regex:k_tracePersistent --csv --log-file gpurun_out/<tag>_c4dram.csv python scripts/c4_trace_dram.py run <tag>
`run` launches the Accel::rayIntersect operator once per ray set (primary, towards the light, bounce) on the 9 999 392-triangle
terrain AFTER all set-up launches; `parse` (here) takes the LAST three k_tracePersistent launches of the list."""
import csv, io, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mode, tag = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "r02")
GO = os.path.join(ROOT, "gpurun_out")
NAMES = ["primary_closest", "shadow_anyhit", "bounce_closest"]
if mode == "run":
    import torch
    from bench import c4_rays
    from nori_b200 import synth
    from nori_b200.gpu import Engine
    scene = synth.terrain_scene()
    e = Engine(0); e.load_scene(scene)
    sets = c4_rays(e, scene)
    bufs = [(torch.from_numpy(r).cuda(), torch.empty((len(r), 4), dtype=torch.float32, device="cuda")) for r in sets]
    torch.cuda.synchronize()
    ms = [e.trace_device(dr.data_ptr(), dr.shape[0], dh.data_ptr(), shadow=(i == 1)) for i, (dr, dh) in enumerate(bufs)]
    json.dump({"rays": [int(len(r)) for r in sets], "ms_under_ncu": ms}, open(os.path.join(GO, tag + "_c4dram_stats.json"), "w"))
else:
    st = json.load(open(os.path.join(GO, tag + "_c4dram_stats.json")))
    txt = open(os.path.join(GO, tag + "_c4dram.csv")).read()
    per = {}
    for r in csv.DictReader(io.StringIO(txt[txt.index('"ID"'):])):
        v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
        if r["Metric Name"].startswith("dram__bytes"): v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        if r["Metric Name"] == "gpu__time_duration.sum": v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1)
        per.setdefault(int(r["ID"]), {})[r["Metric Name"]] = v
    last = [per[k] for k in sorted(per)][-3:]
    out = {"source": f"ncu ({tag}): dram__bytes_read.sum + dram__bytes_write.sum of one k_tracePersistent launch per ray set on the C4 terrain (raw list profiles/{tag}_c4dram.csv)", "rays": {}}
    for name, n, m in zip(NAMES, st["rays"], last):
        out["rays"][name] = {"rays": n, "dram_bytes_per_ray": (m["dram__bytes_read.sum"] + m["dram__bytes_write.sum"]) / n, "ncu_ms": m["gpu__time_duration.sum"],
                             "warp_inst_per_ray": m["smsp__inst_executed.sum"] / n, "lanes_per_inst": m["smsp__thread_inst_executed.sum"] / m["smsp__inst_executed.sum"]}
    json.dump(out, open(os.path.join(ROOT, "profiles", "c4_trace_dram.json"), "w"), indent=1)
    import shutil; shutil.copy(os.path.join(GO, tag + "_c4dram.csv"), os.path.join(ROOT, "profiles", tag + "_c4dram.csv"))
    print(json.dumps(out, indent=1))
