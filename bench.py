#!/usr/bin/env python
"""bench.py — headline benchmark of the Nori hot path on B200 (contract: task brief + SURVEY.md 8d).

Workload (BASELINE.json configs[1], "C2"): Cornell box `scenes/pa5/cbox/cbox_mis.xml` (fixture
tests/golden/scenes/cbox_mis.npz: 6 meshes, 2 188 triangles), `path_mis`, 800x600, 1024 samples per
pixel per step, PER_SAMPLE sampler streams. One step = one full render of that workload on each GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--spp S] [--impl reference]

N > 1 (torchrun, one process per GPU): rank r renders sample indices [r*S, (r+1)*S) of every pixel
(weak scaling: per-GPU work fixed) and the ImageBlocks are summed with one NCCL reduce per step —
the multi-GPU form of `result.put(block)` (src/block.cpp:105-114) — inside the timed region.

value      Msamples/s (camera samples) device-timed with the scene resident in HBM.
e2e        the same metric through the C ABI from HOST buffers: scene upload + accel build +
           render + ImageBlock download inside the timed region.
roofline   k_traceClosest (dominant kernel): algorithmic bytes per ray (profiles/algorithmic_bytes.json,
           counted by the oracle BVH) x rays per launch / live CUDA-event duration of that stage.
cpu_baseline / --impl reference: the reference's TBB render loop (oracle/_ref, real reference sources
           + restated plugins) or the oracle port on the host cores, bounded spp.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from nori_b200.scene import FlatScene  # noqa: E402

SCENE = "cbox_mis"
METRIC = "Msamples/s"


def load_scene():
    return FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", SCENE + ".npz"))


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu: int):
        super().__init__(daemon=True)
        self.gpu, self.rows, self.stop_flag = gpu, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
class _QuietStdout:
    """The compiled reference prints its scene report to fd 1; bench.py's stdout carries one JSON line only."""
    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(null, 1); os.close(null)
    def __exit__(self, *exc):
        os.dup2(self.saved, 1); os.close(self.saved)


_REF_CACHE = {}


def cpu_reference(scene: FlatScene, spp: int, repeat: int = 1):
    """The reference's CPU render loop on all host cores. kind "reference": oracle/_ref/libnori_ref.so, i.e.
    the reference's own sources compiled as they are (parser, OBJ loader, camera, sampler, ImageBlock,
    BlockGenerator, TBB parallel_for over blocks: src/main.cpp:86-127) plus the restated plugins the
    reference lacks (integrators, area emitter, S6 BVH); the scene reaches it as XML + OBJ files written
    from the same FlatScene (oracle/export_xml.py). kind "port": the oracle port (std::thread pool over
    the same block order) where _ref was not built. Returns (Msamples/s, kind, cores, seconds)."""
    cores = os.cpu_count() or 1
    samples = scene.width * scene.height * spp
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libnori_ref.so")
    if "render" not in _REF_CACHE:
        if os.path.exists(ref_so):
            import tempfile
            from oracle.export_xml import write_scene_files      # oracle: CPU baseline leg only
            from oracle.pyoracle import Ref
            tmp = tempfile.mkdtemp(prefix="nori_ref_scene_")
            with _QuietStdout():
                rs = Ref(build=False).load_xml(write_scene_files(scene, tmp))
            _REF_CACHE["kind"] = "reference"
            _REF_CACHE["render"] = lambda: rs.render(scene.block_shape(), mode=1, sample_first=0, sample_last=spp, threads=0)[1]
        else:
            from oracle.pyoracle import Port                       # oracle: CPU baseline leg only
            ps = Port(build=not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so"))).scene(scene)
            def run():
                t0 = time.perf_counter()
                ps.render(mode=1, sample_first=0, sample_last=spp, threads=cores)
                return time.perf_counter() - t0
            _REF_CACHE["kind"] = "port"
            _REF_CACHE["render"] = run
    best = None
    for _ in range(repeat):
        dt = _REF_CACHE["render"]()
        best = dt if best is None else min(best, dt)
    return samples / best / 1e6, _REF_CACHE["kind"], cores, best


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scene = load_scene()
    spp = max(1, args.ref_spp)
    times = []
    kind, cores = "port", os.cpu_count() or 1
    for i in range(args.warmup + args.steps):
        v, kind, cores, dt = cpu_reference(scene, spp)
        if i >= args.warmup:
            times.append(dt)
    samples = scene.width * scene.height * spp
    ms = 1e3 * sum(times) / len(times)
    value = samples / (ms / 1e3) / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": f"cbox path_mis 800x600 (C2), CPU sample of {spp} spp per step", "scene": SCENE, "spp": spp},
            "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": kind,
                             "sample": f"{spp} spp of the 800x600 frame per step ({samples} camera samples), {cores} threads"},
            "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
class _CudaArray:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from nori_b200.gpu import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — nori_b200 has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    scene = load_scene()
    spp = args.spp
    eng = Engine(local)
    if args.pool:
        eng.set_pool_size(args.pool)
    accel = eng.load_scene(scene)
    eng.set_profiling(True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2
    ptr, nfl = eng.block_device_ptr()
    block_t = torch.as_tensor(_CudaArray(ptr, nfl), device="cuda") if world > 1 else None
    first, last = rank * spp, (rank + 1) * spp

    def step():
        flush.zero_()
        torch.cuda.synchronize()
        _, st = eng.render(sample_first=first, sample_last=last, clear=True, download=False)
        red_ms = 0.0
        if world > 1:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); dist.reduce(block_t, dst=0, op=dist.ReduceOp.SUM); e1.record(); e1.synchronize()
            red_ms = e0.elapsed_time(e1)
        return st, red_ms

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local); sampler.start()
    t0 = time.perf_counter()
    dev_ms, stats, red_total = 0.0, [], 0.0
    for _ in range(args.steps):
        st, red = step()
        stats.append(st); dev_ms += st["ms_total"] + red; red_total += red
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.summary()

    # ---- e2e: HOST buffers through the C ABI (upload + build + render + download per step) -----
    e2e_steps = max(1, min(args.steps, 2))
    barrier()
    t1 = time.perf_counter()
    for _ in range(e2e_steps):
        eng.load_scene(scene)
        eng.render(sample_first=first, sample_last=last, clear=True, download=False)
        if world > 1:
            dist.reduce(block_t, dst=0, op=dist.ReduceOp.SUM)
        blk = eng.download_block()
    barrier()
    e2e_ms = (time.perf_counter() - t1) * 1e3 / e2e_steps
    h2d = sum(m.V.nbytes + m.F.nbytes + (m.N.nbytes if m.N is not None else 0) + (m.UV.nbytes if m.UV is not None else 0) for m in scene.meshes) + 33 * 4 + 136
    d2h = blk.nbytes

    tmax = torch.tensor([dev_ms, wall_ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_ms = [float(x) for x in tmax.tolist()]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    samples_per_step = scene.width * scene.height * spp * world
    ms_per_step = dev_ms / args.steps
    value = samples_per_step / (ms_per_step / 1e3) / 1e6
    tot = {k: sum(s[k] for s in stats) for k in stats[0]}
    rays = tot["closest_rays"] + tot["shadow_rays"]
    # roofline of the dominant kernel
    alg = json.load(open(os.path.join(ROOT, "profiles", "algorithmic_bytes.json")))["scenes"][SCENE]
    peak, peak_src = measured_peak()
    # the splat of the previous iteration runs under the closest-hit pass (programmatic dependent launch)
    stage_ms = {"traceClosest+splat": tot["ms_trace_closest"], "prepare+shadeRegen": tot["ms_shade"], "traceShadow": tot["ms_trace_shadow"]}
    launches_closest = tot["iterations"]
    bytes_per_launch = alg["closest"]["bytes_per_ray"] * tot["closest_rays"] / launches_closest
    avg_launch_ms = tot["ms_trace_closest"] / launches_closest
    achieved = bytes_per_launch / (avg_launch_ms / 1e3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        traffic = next((v for k, v in tj.items() if k.startswith("k_traceClosest") and k.endswith("_dram_bytes_per_launch")), None)
    cpu_v, cpu_kind, cores, cpu_s = cpu_reference(scene, args.ref_spp) if world == 1 and not args.no_cpu else (None, None, None, None)
    line = {
        "metric": METRIC, "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"cbox path_mis 800x600 {spp}spp per GPU (BASELINE configs[1], C2)", "scene": SCENE, "triangles": scene.n_triangles,
                   "spp_per_gpu": spp, "sampler": "pcg32 per-sample streams", "partition": "sample-index ranges, one NCCL reduce of the ImageBlock per step",
                   "l2": "256 MiB flush buffer written between steps; path pool + sample records (> 400 MB) exceed L2, BVH+geometry (0.25 MB) are cache-resident by nature of the scene",
                   "pool_paths": args.pool or 1 << 23, "accel": accel},
        "e2e": {"value": samples_per_step / (e2e_ms / 1e3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms, "includes": "scene upload, accel build, render, NCCL reduce, ImageBlock download"},
        "gpu_launches": int(tot["kernel_launches"]),
        "clocks": clocks,
        "roofline": {"kernel": "k_traceClosest", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "bytes_per_ray": alg["closest"]["bytes_per_ray"],
                     "rays_per_launch": tot["closest_rays"] / launches_closest, "avg_launch_ms": avg_launch_ms,
                     "timing": "CUDA events around k_traceClosest on its stream; the interval also contains the previous iteration's splat kernels, which run concurrently with it (programmatic dependent launch), so `achieved` is a lower bound",
                     "traffic_capture": "ncu --set full of one steady-state launch at 8 Mi path slots, about 5.4 M live rays (profiles/r01b_kernels.md): ray records in/out only",
                     "note": "algorithmic bytes are served by L1 (BVH + triangles of this scene are 0.17 MB): the kernel is bound by instruction issue (70 %) and the L1 LSU data pipe (80 %), not by HBM (ncu: 6 % DRAM); see DESIGN.md and profiles/r01_trace.md"},
        "stages_ms_per_step": {k: v / args.steps for k, v in stage_ms.items()},
        "Mrays_per_s": rays * world / (dev_ms / 1e3) / 1e6, "Gsegments_per_s": tot["segments"] * world / (dev_ms / 1e3) / 1e9,
        "segments_per_sample": tot["segments"] / max(tot["camera_samples"], 1), "reduce_ms_per_step": red_total / args.steps,
        "wall_ms_per_step": wall_ms / args.steps, "dropped_samples": int(tot["dropped"]),
    }
    if cpu_v is not None:
        line["cpu_baseline"] = {"value": cpu_v, "unit": "Msamples/s", "cores": cores, "kind": cpu_kind,
                                "sample": f"{args.ref_spp} spp of the same 800x600 frame ({scene.width * scene.height * args.ref_spp} camera samples, {cpu_s:.1f} s)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--ref-spp", type=int, default=96, help="spp of the bounded CPU sample")
    ap.add_argument("--pool", type=int, default=0)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
