"""Kernel-only traversal throughput in the HBM-bound regime: Mrays/s of the
Accel::rayIntersect operator (nori_gpu_trace_device, rays resident in HBM, CUDA events) for coherent
primary rays, shadow rays towards the light and incoherent bounce rays, with the algorithmic bytes per
ray (48 + 32 N_box + 36 N_tri, counted by the device BVH: nori_gpu_trace_count) against the measured
HBM peak.   python scripts/bench_trace.py [c4|c1] [--layout 0|1] [--out profiles/trace_<cfg>.json]
--layout 1: NORI_ACCEL_QUANTISED (32-byte nodes), 2: NORI_ACCEL_WIDE4 (the tree collapsed to 4-wide nodes; N_box counts stay the binary tree's); the algorithmic bytes stay those of SURVEY 8d (32 B per box test)."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from nori_b200.gpu import Engine
from nori_b200 import synth

def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        for k in ("hbm_gbs", "hbm_gbps", "hbm_copy_gbs"):
            if k in d: return float(d[k]), "measured (MEASURED_PEAKS.json)"
    return 6500.0, "fallback (B200_PROFILING.md)"

def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "c4"
    out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
    scene = synth.terrain_scene() if cfg == "c4" else synth.ajax_class_scene()
    layout = int(sys.argv[sys.argv.index("--layout") + 1]) if "--layout" in sys.argv else 0
    eng = Engine(0)
    eng.set_accel_layout(layout)
    t0 = time.perf_counter(); accel = eng.load_scene(scene); load_s = time.perf_counter() - t0
    W, H = scene.width, scene.height
    rng = np.random.default_rng(7)
    ys, xs = np.mgrid[0:H, 0:W]
    # 8x4 pixel patches per warp, like the renderer's issue order
    ps = np.stack([xs + 0.5, ys + 0.5], -1).astype(np.float32)
    ps = ps.reshape(H // 4, 4, W // 8, 8, 2).transpose(0, 2, 1, 3, 4).reshape(-1, 2)
    primary = eng.camera_rays(ps)
    tri, mesh, tuv = eng.trace(primary)
    hit = tri != 0xffffffff
    P = primary[hit, 0:3] + tuv[hit, 0:1] * primary[hit, 4:7]
    light = np.array([0.0, 1.2, 0.0], np.float32) if cfg == "c4" else primary[0, 0:3]
    d = light - P; dist = np.linalg.norm(d, axis=1, keepdims=True)
    shadow = np.concatenate([P, np.full((len(P), 1), 1e-4, np.float32), d / dist, dist - 1e-4], 1).astype(np.float32)
    v = rng.normal(size=P.shape).astype(np.float32); v /= np.linalg.norm(v, axis=1, keepdims=True)
    if cfg == "c4": v[:, 1] = np.abs(v[:, 1]) * 0.3          # grazing bounce rays over the terrain
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    bounce = np.concatenate([P, np.full((len(P), 1), 1e-4, np.float32), v, np.full((len(P), 1), np.inf, np.float32)], 1).astype(np.float32)
    hbm, src = peak()
    res = {"config": cfg, "node_layout": ["float 64 B", "quantised 32 B", "4-wide float 128 B"][layout], "triangles": accel["n_triangles"], "accel": accel, "load_and_build_s": load_s, "hbm_peak_gbs": hbm, "peak_source": src, "rays": {}}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for name, rays, sh in (("primary_closest", primary, False), ("shadow_anyhit", shadow, True), ("bounce_closest", bounce, False)):
        n = len(rays)
        sub = rays[rng.choice(n, min(n, 1 << 20), replace=False)] if n > (1 << 20) else rays
        nb, nt = eng.trace_count(sub, shadow=sh)
        bpr = 48 + 32 * (2 * nb / len(sub)) + 36 * (nt / len(sub))      # two child boxes per node visit
        dr = torch.from_numpy(rays).cuda(); dh = torch.empty((n, 4), dtype=torch.float32, device="cuda")
        ms = []
        for it in range(6):
            flush.zero_(); torch.cuda.synchronize()
            t = eng.trace_device(dr.data_ptr(), n, dh.data_ptr(), shadow=sh)
            if it >= 3: ms.append(t)
        t = float(np.mean(ms))
        res["rays"][name] = {"n": n, "ms": t, "Mrays_per_s": n / t / 1e3, "n_box_per_ray": 2 * nb / len(sub), "n_tri_per_ray": nt / len(sub),
                             "bytes_per_ray": bpr, "algorithmic_GBs": n * bpr / t / 1e6, "frac_of_hbm_peak": n * bpr / t / 1e6 / hbm}
        del dr, dh
    print(json.dumps(res))
    if out: json.dump(res, open(out, "w"), indent=1)

main()
