"""Turn the ncu captures a gpurun call left in gpurun_out/ into the tracked summaries under profiles/.
Runs in the build container (ncu reads .ncu-rep files without a GPU):
    python scripts/make_profile_summaries.py [round-tag]
Inputs : gpurun_out/<tag>_launches.csv (ncu --metrics gpu__time_duration.sum --csv)
         gpurun_out/<tag>_full.ncu-rep (ncu --set full --import-source on)
Outputs: profiles/<tag>_launches.csv (copy), profiles/<tag>_launches.md (per-kernel share of the step),
         profiles/<tag>_kernels.md/.csv (per-kernel counters), profiles/traffic.json (DRAM bytes per launch)."""
import collections, csv, io, json, os, shutil, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
GO, PR = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")

def short(name):
    n = name.split("(")[0].replace("void ", "").replace("ng::", "")
    return n

def launches():
    src = os.path.join(GO, tag + "_launches.csv")
    txt = open(src).read()
    shutil.copy(src, os.path.join(PR, tag + "_launches.csv"))
    rows = csv.DictReader(io.StringIO(txt[txt.index('"ID"'):]))
    agg = collections.OrderedDict()
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum": continue
        v = float(r["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r["Metric Unit"], 1.0)
        if "at::" in r["Kernel Name"]: continue            # bench.py's L2 flush (torch), not a kernel of this library
        a = agg.setdefault(short(r["Kernel Name"]), [0, 0.0]); a[0] += 1; a[1] += v
    tot = sum(a[1] for a in agg.values())
    out = ["# %s: launch list of `python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-configs`" % tag, "",
           "`ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 400` (steady state of the wavefront",
           "loop, the tail of one frame and the start of the next; 8 Mi path slots; times are cold-cache and serialised,",
           "so read the SHARES). Raw list: `%s_launches.csv`." % tag, "",
           "| kernel | launches | total us | avg us | share |", "|---|---:|---:|---:|---:|"]
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        out.append("| `%s` | %d | %.1f | %.1f | %.1f %% |" % (k, n, t, t / n, 100 * t / tot))
    open(os.path.join(PR, tag + "_launches.md"), "w").write("\n".join(out) + "\n")
    return {k: t / tot for k, (n, t) in agg.items()}

WANT = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
        ("lts__t_bytes.sum", "L2 bytes"), ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue slots busy %"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes / instruction"),
        ("l1tex__t_sector_hit_rate.pct", "L1 hit %"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1 throughput %"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
        ("smsp__inst_executed.sum", "warp instructions")]

def kernels():
    rep = os.path.join(GO, tag + "_full.ncu-rep")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u, data = rows[0], rows[1], rows[2:]
    ix = {k: i for i, k in enumerate(h)}
    per = collections.OrderedDict()
    for d in data: per.setdefault(short(d[ix["Kernel Name"]]), []).append(d)
    md = ["# %s: `ncu --set full` counters (per launch, averaged over the captured launches)" % tag, "",
          "Command: `python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-configs`, `-k regex:k_trace|k_shade|k_splatPixel|k_bin -s 70 -c 14` (r01 / r02: without k_bin, `-s 60 -c 8`), cbox path_mis, 8 Mi path slots (128 spp: below the 16 Mi rule)", "(about 5.4 M live rays per launch).", ""]
    names = list(per)
    md += ["| counter | " + " | ".join("`%s`" % n for n in names) + " |", "|---|" + "---:|" * len(names)]
    table = {}
    for key, label in WANT:
        if key not in ix: continue
        vals = []
        for n in names:
            xs = [float(d[ix[key]].replace(",", "")) for d in per[n] if d[ix[key]] not in ("", "n/a")]
            vals.append(sum(xs) / len(xs) if xs else float("nan"))
        table[key] = dict(zip(names, vals))
        md.append("| %s [%s] | " % (label, u[ix[key]]) + " | ".join("%.4g" % v for v in vals) + " |")
    open(os.path.join(PR, tag + "_kernels.md"), "w").write("\n".join(md) + "\n")
    with open(os.path.join(PR, tag + "_kernels.csv"), "w") as f:
        w = csv.writer(f); w.writerow(["counter", "unit"] + names)
        for key, _ in WANT:
            if key in table: w.writerow([key, u[ix[key]]] + ["%.6g" % table[key][n] for n in names])
    def to_bytes(key, n):
        v = table[key][n]; unit = u[ix[key]].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(unit, 1)
    traffic = {"source": "profiles/%s_kernels.csv (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch)" % tag,
               "workload": "cbox path_mis 800x600, 8 Mi path slots, steady state (about 5.4 M live rays per launch)"}
    for n in names:
        traffic[n + "_dram_bytes_per_launch"] = to_bytes("dram__bytes_read.sum", n) + to_bytes("dram__bytes_write.sum", n)
    json.dump(traffic, open(os.path.join(PR, "traffic.json"), "w"), indent=1)
    return table

if __name__ == "__main__":
    sh = launches(); kernels()
    print({k: round(v, 3) for k, v in sh.items()})
