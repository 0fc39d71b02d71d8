"""Per-CUDA-source-line totals from `ncu --page source --csv --print-source cuda,sass`:
python scripts/ncu_lines.py dump.csv [N] -> top-N lines by warp instructions executed and by stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
cur_file = None; lines = []
hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = {k: i for i, k in enumerate(r)}; continue
    if hdr and r[0] not in ("", "-"):
        try:
            lines.append((cur_file, int(r[0]), r[1].strip()[:90], float(r[hdr["Instructions Executed"]] or 0), float(r[hdr["# Samples"]] or 0),
                          float(r[hdr["Thread Instructions Executed"]] or 0)))
        except ValueError:
            pass
ti = sum(l[3] for l in lines); ts = sum(l[4] for l in lines)
print(f"total warp inst {ti:.0f}, samples {ts:.0f}")
print("-- by instructions")
for f, ln, src, ins, smp, thr in sorted(lines, key=lambda l: -l[3])[:n]:
    print(f"{ins / ti * 100:5.1f}% inst {smp / ts * 100:5.1f}% smp lanes={thr / max(ins, 1):4.1f} {f}:{ln}  {src}")
print("-- by stall samples")
for f, ln, src, ins, smp, thr in sorted(lines, key=lambda l: -l[4])[:n]:
    print(f"{smp / ts * 100:5.1f}% smp {ins / ti * 100:5.1f}% inst lanes={thr / max(ins, 1):4.1f} {f}:{ln}  {src}")
