"""SASS summary of the wavefront kernels of the shipped build -> profiles/<tag>_sass.md
(python scripts/sass_extract.py [tag]; needs cuobjdump, no GPU). Per kernel: instruction count, mix by mnemonic, every
global / local / shared memory, prefetch, barrier and atomic instruction form, registers and spills as ptxas reports them."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02b"
obj = os.path.join(ROOT, "nori_b200", "build", "render.o")
sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", obj], capture_output=True, text=True).stdout
usage = {}
for m in re.finditer(r"Function (\S+):\s*\n\s*(REG:\d+[^\n]*)", res):
    usage[m.group(1)] = m.group(2)
WANT = [("k_traceClosestILi9ELb0ELi0", "k_traceClosest<9, false, 0>"), ("k_traceShadowILi10ELb0ELi0", "k_traceShadow<10, false, 0>"),
        ("k_shadeRegenILi3ELi256ELb1", "k_shadeRegen<3, 256, true> (scenes without microfacet surfaces: the Cornell box)"), ("k_shadeRegenILi3ELi256ELb0", "k_shadeRegen<3, 256, false>"), ("k_binScatter", "k_binScatter"), ("k_splatPixelILi2ELi256", "k_splatPixel<2, 256>")]
funcs = collections.OrderedDict(); cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m: cur = m.group(1); funcs[cur] = []; continue
    m = re.search(r"/\*[0-9a-f]{4,6}\*/\s+(.*?);", line)
    if cur and m: funcs[cur].append(m.group(1).strip())
out = [f"# {tag}: SASS of the wavefront kernels of the shipped build (`cuobjdump -sass nori_b200/build/render.o`, sm_100a; `scripts/sass_extract.py`)", "",
       "What to read off: nodes are fetched with `LDG.E.ENL2.256` (two per visit), leaf triangles with `LDG.E.128`; the traversal kernels' only local-memory traffic (`STL` / `LDL`)",
       "is the per-lane stack (no spills: SPILL columns of `cuobjdump -res-usage` below); the shade kernel's software prefetches are the `CCTL` / `PREFETCH` forms;",
       "its barriers are `BAR.SYNC` / `BAR.RED`.", ""]
for key, title in WANT:
    name = next((f for f in funcs if key in f), None)
    if not name: continue
    ins = funcs[name]
    mn = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", i).split()[0].split(".")[0] for i in ins)
    forms = collections.Counter()
    for i in ins:
        op = re.sub(r"^@!?U?P\d+\s+", "", i).split()[0]
        if re.match(r"(LDG|STG|LDL|STL|LDS|STS|ATOM|RED|BAR|CCTL|PREFETCH|LDGSTS|LDC|ATOMS|MATCH)", op): forms[op] += 1
    out += [f"## `{title}` — {len(ins)} SASS instructions", "", f"resources: {usage.get(name, 'n/a')}", "",
            "mix: " + ", ".join(f"{k} {v}" for k, v in mn.most_common(24)), "",
            "memory / barrier forms: " + ", ".join(f"`{k}` x{v}" for k, v in sorted(forms.items())), ""]
open(os.path.join(ROOT, "profiles", tag + "_sass.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out)[:3000])
