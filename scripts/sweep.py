"""Parameter sweep helper (GPU box): one subprocess per environment setting, short cbox render,
prints stage times. Usage: python scripts/sweep.py KEY=v1,v2 [KEY2=...] [--spp N] [--scene name]"""
import itertools, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, json; sys.path.insert(0, %r)
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
s = FlatScene.load(%r)
e = Engine(0); acc = e.load_scene(s); e.set_profiling(True)
e.render(sample_first=0, sample_last=%d, download=False)
_, st = e.render(sample_first=0, sample_last=%d, download=False)
print(json.dumps({"ms": st["ms_total"], "closest": st["ms_trace_closest"], "shadow": st["ms_trace_shadow"], "shade": st["ms_shade"],
                  "regen": st["ms_generate"], "splat": st["ms_splat"], "sah": acc["sah_cost"], "nodes": acc["n_nodes"], "depth": acc["depth"], "build_ms": acc["build_ms"]}))
'''
def main():
    spp, scene, axes = 128, "cbox_mis", []
    args = sys.argv[1:]
    while args:
        a = args.pop(0)
        if a == "--spp": spp = int(args.pop(0))
        elif a == "--scene": scene = args.pop(0)
        else:
            k, v = a.split("="); axes.append([(k, x) for x in v.split(",")])
    path = os.path.join(ROOT, "tests", "golden", "scenes", scene + ".npz")
    for combo in itertools.product(*axes):
        env = dict(os.environ); env.update({k: v for k, v in combo})
        out = subprocess.run([sys.executable, "-c", CHILD % (ROOT, path, spp, spp)], env=env, capture_output=True, text=True, timeout=300)
        line = out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-300:]
        print(" ".join(f"{k}={v}" for k, v in combo), "->", line, flush=True)
main()
