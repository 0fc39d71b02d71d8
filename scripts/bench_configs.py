"""Throughput of the other BASELINE configs through the full render path (GPU box), one JSON line each:
C1 ajax-class normals 768^2 32 spp, C3 table_mis 800x600, C4 terrain primary+shadow 3840x2160, veach_mis.
    python scripts/bench_configs.py [c1 c3 c4 veach cbox] [--spp N]
Device time of nori_gpu_render (CUDA events inside the call), second of two renders."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
from nori_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = [a for a in sys.argv[1:] if not a.startswith("--")]
spp_override = int(sys.argv[sys.argv.index("--spp") + 1]) if "--spp" in sys.argv else 0
which = args or ["c1", "c3", "c4", "veach"]
e = Engine(0)
for w in which:
    if w == "c1": s, spp, name = synth.ajax_class_scene(), 32, "C1 ajax-class 544 968 tris, normals, 768x768"
    elif w == "c4": s, spp, name = synth.terrain_scene(), 4, "C4 terrain 9 999 394 tris, primary+shadow, 3840x2160"
    elif w == "c3": s, spp, name = FlatScene.load(os.path.join(ROOT, "tests/golden/scenes/table_mis.npz")), 256, "C3 table_mis 800x600"
    elif w == "veach": s, spp, name = FlatScene.load(os.path.join(ROOT, "tests/golden/scenes/veach_mis.npz")), 256, "veach_mis"
    else: s, spp, name = FlatScene.load(os.path.join(ROOT, "tests/golden/scenes/cbox_mis.npz")), 256, "C2 cbox_mis 800x600"
    spp = spp_override or spp
    acc = e.load_scene(s); e.set_profiling(True)
    e.render(sample_first=0, sample_last=spp, download=False)
    _, st = e.render(sample_first=0, sample_last=spp, download=False)
    ms = st["ms_total"]
    print(json.dumps({"config": name, "spp": spp, "ms": round(ms, 2), "Msamples_per_s": round(st["camera_samples"] / ms / 1e3, 1),
                      "Mrays_per_s": round((st["closest_rays"] + st["shadow_rays"]) / ms / 1e3, 1), "segments_per_sample": round(st["segments"] / st["camera_samples"], 3),
                      "stages_ms": {"closest": round(st["ms_trace_closest"], 1), "shade": round(st["ms_shade"], 1), "shadow": round(st["ms_trace_shadow"], 1)},
                      "iterations": st["iterations"], "launches": st["kernel_launches"], "late": st["late_samples"],
                      "build_ms": round(acc["build_ms"], 1), "nodes": acc["n_nodes"]}), flush=True)
