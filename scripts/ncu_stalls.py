"""Summarise an `ncu --page source --csv` dump: stall-reason totals and the top-N SASS instructions by
stall samples. Usage: python scripts/ncu_stalls.py dump.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
h = rows[1]; ix = {k: i for i, k in enumerate(h)}
data = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name": break            # first launch in the dump only
    if len(r) == len(h): data.append(r)
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
tot = {k: sum(float(r[ix[k]] or 0) for r in data) for k in stalls}
S = sum(tot.values())
print("kernel:", rows[0][1][:90])
print("samples", int(S), " ".join(f"{k[6:]}={v/S*100:.1f}%" for k, v in sorted(tot.items(), key=lambda x: -x[1]) if v / S > 0.01))
inst = sum(float(r[ix["Instructions Executed"]] or 0) for r in data)
thr = sum(float(r[ix["Thread Instructions Executed"]] or 0) for r in data)
print(f"warp inst {inst:.0f} thread inst {thr:.0f} lanes/inst {thr / max(inst, 1):.2f}")
def samples(r): return float(r[ix["# Samples"]] or 0)
for r in sorted(data, key=samples, reverse=True)[:n]:
    top = sorted(((float(r[ix[k]] or 0), k[6:]) for k in stalls), reverse=True)[:2]
    print(f"{samples(r) / S * 100:5.1f}% {r[ix['Address']][-5:]} {r[ix['Source']][:70]:70s} thr/inst={r[ix['Avg. Threads Executed']]:>5s} " + " ".join(f"{k}:{v:.0f}" for v, k in top))
