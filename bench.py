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
roofline   k_traceClosest (dominant kernel). The Cornell box's tree (0.17 MB) lives in L1, so HBM does not bound it
           (ncu: 6 % DRAM): the binding limit is instruction issue. `achieved` = warp instructions per second
           (instructions per live ray from the committed ncu capture, profiles/issue_model.json, x rays per launch /
           live CUDA-event duration of the stage) against SMs x 4 schedulers x SM clock; the algorithmic-byte rate
           of SURVEY 8d (`algorithmic_gbs`) and the measured DRAM fraction (`dram_frac`) stand beside it.
rel_mse    GPU image vs the CPU reference's image of the same samples (same per-sample streams, --ref-spp).
configs    N = 1 only, after the timed steps, each with its own clock record: C1 (544 968 triangles, normals, 768^2,
           32 spp), C3 (table_mis 800x600, 4096 spp), C4 (9 999 392 triangles, 3840x2160: upload + BVH build, the
           primary+shadow render, and kernel-only traversal of primary / shadow / incoherent rays with N_box, N_tri).
strong     N > 1 only: strong scaling of C3 (4096 spp split into sample ranges; rank 0 alone renders the same frame in
           the same run as the 1-GPU leg) and, at N = 8, C5 (Cornell box 3840x2160, 16 384 spp).
cpu_baseline / --impl reference: the reference's TBB render loop (oracle/_ref, real reference sources
           + restated plugins) or the oracle port on the host cores, bounded spp; `cores` = the TBB workers that
           actually rendered blocks.
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
    """SM clock + throttle reasons DURING a timed region. NVML in-process (nvidia_ml_py: a query takes ~1 ms, sampled every 50 ms,
    so that even a 0.3 s region gets several samples); falls back to polling nvidia-smi (one query can take a second on an 8-GPU box)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}   # nvml.h nvmlClocksEventReason*
    _nvml = None

    def __init__(self, gpu: int):
        super().__init__(daemon=True)
        self.gpu, self.sm, self.mx, self.reasons, self.stop_flag, self.source = gpu, [], [], set(), threading.Event(), "nvidia-smi"
        self.handle = None
        try:
            import pynvml
            if ClockSampler._nvml is None:
                pynvml.nvmlInit(); ClockSampler._nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(gpu)
            self.source = "nvml"
        except Exception:
            self.handle = None

    def _sample_nvml(self):
        n = ClockSampler._nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        self.mx.append(float(n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)))
        get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(n, "nvmlDeviceGetCurrentClocksThrottleReasons")
        mask = int(get(self.handle))
        self.reasons |= {k for k, bit in self.BITS.items() if mask & bit}

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        r = [x.strip() for x in out.split(",")] if out else []
        if r and r[0].replace(".", "").isdigit():
            self.sm.append(float(r[0]))
        if len(r) > 1 and r[1].replace(".", "").isdigit():
            self.mx.append(float(r[1]))
        self.reasons |= {n for i, n in enumerate(self.NAMES) if len(r) > 2 + i and r[2 + i].lower().startswith("active")}

    def run(self):
        while not self.stop_flag.is_set():
            try:
                self._sample_nvml() if self.handle is not None else self._sample_smi()
            except Exception:
                if self.handle is not None:
                    self.handle, self.source = None, "nvidia-smi"         # NVML refused: poll nvidia-smi instead
            self.stop_flag.wait(0.05 if self.handle is not None else 0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": [n for n in self.NAMES if n in self.reasons], "samples": len(self.sm), "source": self.source}


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
    the same block order) where _ref was not built. Returns (Msamples/s, kind, cores, seconds); the image of the
    last render stays in _REF_CACHE["image"], `cores` = threads that rendered blocks (not os.cpu_count())."""
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
            def run_ref():
                img, sec = rs.render(scene.block_shape(), mode=1, sample_first=0, sample_last=spp, threads=0)
                _REF_CACHE["image"], _REF_CACHE["threads"] = img, rs.threads_used() or cores
                return sec
            _REF_CACHE["render"] = run_ref
        else:
            from oracle.pyoracle import Port                       # oracle: CPU baseline leg only
            ps = Port(build=not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so"))).scene(scene)
            def run():
                t0 = time.perf_counter()
                _REF_CACHE["image"] = ps.render(mode=1, sample_first=0, sample_last=spp, threads=cores)[0]
                _REF_CACHE["threads"] = cores
                return time.perf_counter() - t0
            _REF_CACHE["kind"] = "port"
            _REF_CACHE["render"] = run
    best = None
    for _ in range(repeat):
        dt = _REF_CACHE["render"]()
        best = dt if best is None else min(best, dt)
    return samples / best / 1e6, _REF_CACHE["kind"], _REF_CACHE.get("threads", cores), best


def host_cpu_info():
    """What the box offers next to what the CPU arm used: logical CPUs, affinity mask, cgroup CPU quota."""
    info = {"logical_cpus": os.cpu_count()}
    try:
        info["affinity"] = len(os.sched_getaffinity(0))
    except Exception:
        pass
    try:
        q, p = open("/sys/fs/cgroup/cpu.max").read().split()[:2]
        info["cgroup_cpu_quota"] = None if q == "max" else round(float(q) / float(p), 2)
    except Exception:
        pass
    return info


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
            "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": kind, "host": host_cpu_info(),
                             "sample": f"{spp} spp of the 800x600 frame per step ({samples} camera samples), {cores} TBB workers rendered blocks"},
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
        with _QuietStdout():                       # NCCL announces its version on stdout when the communicator comes up: stdout carries one JSON line only
            dist.barrier()
            torch.cuda.synchronize()

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

    strong = None
    if world > 1 and not args.no_configs:
        strong = measure_strong(eng, torch, dist, rank, world, local, args)

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
    # ---- roofline of the dominant kernel (k_traceClosest): instruction issue, with the byte rates beside it ----
    alg = json.load(open(os.path.join(ROOT, "profiles", "algorithmic_bytes.json")))["scenes"][SCENE]
    peak, peak_src = measured_peak()
    # the splat of the previous iteration runs under the closest-hit pass (programmatic dependent launch)
    stage_ms = {"traceClosest+splat": tot["ms_trace_closest"], "prepare+shadeRegen": tot["ms_shade"], "traceShadow": tot["ms_trace_shadow"]}
    launches_closest = tot["iterations"]
    rays_per_launch = tot["closest_rays"] / launches_closest
    avg_launch_ms = tot["ms_trace_closest"] / launches_closest
    algorithmic_gbs = alg["closest"]["bytes_per_ray"] * rays_per_launch / (avg_launch_ms / 1e3) / 1e9
    im = json.load(open(os.path.join(ROOT, "profiles", "issue_model.json")))
    imk = im["kernels"]["k_traceClosest"]
    sm_count = int(im.get("sm_count", 148))
    sm_mhz = clocks["sm_mhz"] or clocks["sm_max_mhz"] or 1965.0
    issue_peak = sm_count * 4 * sm_mhz * 1e6 / 1e9                                  # G warp-instructions / s the chip can issue
    issue_achieved = imk["warp_inst_per_ray"] * rays_per_launch / (avg_launch_ms / 1e3) / 1e9
    traffic = imk["dram_bytes_per_ray"] * rays_per_launch                            # ncu dram__bytes_read + write, per launch
    roofline = {"kernel": "k_traceClosest", "bound": "issue", "achieved": issue_achieved, "peak": issue_peak, "unit": "Gwarp-inst/s",
                "frac": issue_achieved / issue_peak, "lanes_per_inst": imk["lanes_per_inst"], "warp_inst_per_ray": imk["warp_inst_per_ray"],
                "useful_lane_frac": issue_achieved / issue_peak * imk["lanes_per_inst"] / 32.0,
                "peak_definition": f"{sm_count} SMs x 4 schedulers x {sm_mhz:.0f} MHz (median SM clock sampled during the timed region)",
                "traffic": traffic, "dram_frac": traffic / (avg_launch_ms / 1e3) / 1e9 / peak, "hbm_peak_gbs": peak, "peak_source": peak_src,
                "algorithmic_gbs": algorithmic_gbs, "algorithmic_bytes_per_ray": alg["closest"]["bytes_per_ray"],
                "rays_per_launch": rays_per_launch, "avg_launch_ms": avg_launch_ms, "counter_source": im["source"],
                "timing": "CUDA events around k_traceClosest on its stream; the interval also contains the previous iteration's splat kernels, which run concurrently with it (programmatic dependent launch), so `achieved` is a lower bound",
                "note": "the Cornell box's nodes + triangles (0.17 MB) are L1-resident: the SURVEY 8d algorithmic bytes (algorithmic_gbs) are served on chip and exceed the HBM peak by nature of the scene; what binds the kernel is instruction issue at partially filled warps (frac), DRAM carries the ray records only (dram_frac)"}
    # the other two stages against the same two limits (shade: DRAM stream of the path pool)
    other = {}
    for kname, stage, units in (("k_shadeRegen", "ms_shade", float(tot["closest_rays"])), ("k_traceShadow", "ms_trace_shadow", float(tot["shadow_rays"]))):
        k = im["kernels"].get(kname)
        if not k or not tot[stage]:
            continue
        sec = tot[stage] / 1e3
        other[kname] = {"issue_frac": k["warp_inst_per_unit"] * units / sec / 1e9 / issue_peak, "dram_frac": k["dram_bytes_per_unit"] * units / sec / 1e9 / peak,
                        "lanes_per_inst": k["lanes_per_inst"], "unit": k["unit"], "ms_per_step": tot[stage] / args.steps}
    roofline["other_kernels"] = other
    # Since the ray binning of round 2 the shade stage is as long as the closest-hit stage. If it is the longer one, it is the
    # dominant kernel and heads the roofline object: its byte stream through HBM (ncu dram bytes per shaded segment x segments /
    # live stage time) against the measured HBM peak, the traversal's issue roofline moving to `other_kernels`.
    ks = im["kernels"].get("k_shadeRegen")
    if ks and tot["ms_shade"] > tot["ms_trace_closest"]:
        sec = tot["ms_shade"] / 1e3
        seg = float(tot["closest_rays"])
        shade_gbs = ks["dram_bytes_per_unit"] * seg / sec / 1e9
        trav = dict(roofline); trav.pop("other_kernels", None)
        other["k_traceClosest"] = trav
        roofline = {"kernel": "k_shadeRegen", "bound": "hbm", "achieved": shade_gbs, "peak": peak, "unit": "GB/s", "frac": shade_gbs / peak,
                    "traffic": ks["dram_bytes_per_unit"] * seg / launches_closest, "peak_source": peak_src,
                    "issue_frac": ks["warp_inst_per_unit"] * seg / sec / 1e9 / issue_peak, "lanes_per_inst": ks["lanes_per_inst"],
                    "algorithmic_bytes_per_segment": alg["bytes_per_segment"] - alg["closest"]["bytes_per_ray"] - alg["shadow_per_closest"] * alg["shadow"]["bytes_per_ray"],
                    "dram_bytes_per_segment_ncu": ks["dram_bytes_per_unit"], "segments_per_launch": seg / launches_closest, "avg_launch_ms": tot["ms_shade"] / launches_closest,
                    "counter_source": im["source"],
                    "timing": "CUDA events around k_prepare + k_shadeRegen + the three bin-sort kernels on their stream (the stage), so `achieved` is a lower bound for the kernel",
                    "note": "neither DRAM (frac) nor instruction issue (issue_frac) binds this kernel: it waits on dependent loads and on its CTA barriers (ncu: long scoreboard 34 %, barrier 24 % of the stall samples); both fractions are reported so that the distance to either roofline can be read off",
                    "other_kernels": other}

    cpu_v, cpu_kind, cores, cpu_s = cpu_reference(scene, args.ref_spp) if world == 1 and not args.no_cpu else (None, None, None, None)
    line = {
        "metric": METRIC, "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"cbox path_mis 800x600 {spp}spp per GPU (BASELINE configs[1], C2)", "scene": SCENE, "triangles": scene.n_triangles,
                   "spp_per_gpu": spp, "sampler": "pcg32 per-sample streams", "partition": "sample-index ranges, one NCCL reduce of the ImageBlock per step",
                   "l2": "256 MiB flush buffer written between steps; path pool + sample records (> 400 MB) exceed L2, BVH+geometry (0.25 MB) are cache-resident by nature of the scene",
                   "pool_paths": args.pool or (1 << 24 if scene.width * scene.height * spp >= 3 << 26 else min(1 << 23, max(1 << 21, scene.width * scene.height * spp // 3 // 256 * 256))),
                   "pool_rule": "library default: 16 Mi path slots for renders of >= 3 * 2^26 camera samples, below that a third of the samples within [2 Mi, 8 Mi] (nori_gpu_set_pool_size overrides)", "accel": accel},
        "e2e": {"value": samples_per_step / (e2e_ms / 1e3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms, "includes": "scene upload, accel build, render, NCCL reduce, ImageBlock download"},
        "gpu_launches": int(tot["kernel_launches"]),
        "clocks": clocks,
        "roofline": roofline,
        "stages_ms_per_step": {k: v / args.steps for k, v in stage_ms.items()},
        "Mrays_per_s": rays * world / (dev_ms / 1e3) / 1e6, "Gsegments_per_s": tot["segments"] * world / (dev_ms / 1e3) / 1e9,
        "segments_per_sample": tot["segments"] / max(tot["camera_samples"], 1), "reduce_ms_per_step": red_total / args.steps,
        "wall_ms_per_step": wall_ms / args.steps, "dropped_samples": int(tot["dropped"]),
    }
    if strong is not None:
        line["strong"] = strong
    if cpu_v is not None:
        line["cpu_baseline"] = {"value": cpu_v, "unit": "Msamples/s", "cores": cores, "kind": cpu_kind, "host": host_cpu_info(),
                                "sample": f"{args.ref_spp} spp of the same 800x600 frame ({scene.width * scene.height * args.ref_spp} camera samples, {cpu_s:.1f} s), {cores} TBB workers rendered blocks"}
        # rel-MSE (BASELINE.json metric): the GPU's image of the SAME camera samples (same per-sample streams) against the CPU reference's
        eng.load_scene(scene)
        gblk, _ = eng.render(sample_first=0, sample_last=args.ref_spp, clear=True)
        line["rel_mse"] = {"value": rel_mse(to_bitmap(gblk, scene.border), to_bitmap(_REF_CACHE["image"], scene.border)), "spp": args.ref_spp,
                           "definition": "mean((gpu - ref)^2 / (ref^2 + 1e-2)) over the 800x600 RGB bitmap, both sides rendering sample indices [0, spp) of every pixel with the same PCG32 streams",
                           "against": cpu_kind}
    if world == 1 and not args.no_configs:
        line["configs"] = measure_configs(eng, torch, local, peak)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def resize_camera(scene, width, height):
    """PerspectiveCamera::activate (src/perspective.cpp:53-86) for another output size: sampleToCamera depends on the aspect."""
    import copy
    s = copy.copy(scene)
    aspect = width / float(height)
    recip = 1.0 / (s.far_clip - s.near_clip)
    cot = 1.0 / np.tan(np.deg2rad(s.fov / 2.0))
    persp = np.array([[cot, 0, 0, 0], [0, cot, 0, 0], [0, 0, s.far_clip * recip, -s.near_clip * s.far_clip * recip], [0, 0, 1, 0]], np.float64)
    scale = np.diag([-0.5, -0.5 * aspect, 1.0, 1.0])
    trans = np.eye(4); trans[0, 3] = -1.0; trans[1, 3] = -1.0 / aspect
    s.sample_to_camera = np.linalg.inv(scale @ trans @ persp).astype(np.float32)
    s.width, s.height = width, height
    return s


def strong_leg(eng, torch, dist, scene, total_spp, one_gpu_spp, rank, world, gpu, name):
    """A FIXED total of `total_spp` samples per pixel split into contiguous sample-index ranges over the ranks
    (nori_b200.shard.sample_range), ONE NCCL sum of the ImageBlocks; then rank 0 ALONE renders `one_gpu_spp` of the same
    frame in the same run (the other ranks idle) as the 1-GPU leg. Device times: CUDA events inside nori_gpu_render plus
    events around the reduce, max over ranks."""
    from nori_b200.shard import sample_range
    eng.load_scene(scene)
    ptr, nfl = eng.block_device_ptr()
    block_t = torch.as_tensor(_CudaArray(ptr, nfl), device="cuda")
    first, last = sample_range(rank, world, total_spp)
    eng.render(sample_first=first, sample_last=first + max(1, (last - first) // 8), clear=True, download=False)      # warm-up
    dist.reduce(block_t, dst=0, op=dist.ReduceOp.SUM)
    dist.barrier(); torch.cuda.synchronize()
    cs = ClockSampler(gpu); cs.start()
    _, st = eng.render(sample_first=first, sample_last=last, clear=True, download=False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); dist.reduce(block_t, dst=0, op=dist.ReduceOp.SUM); e1.record(); e1.synchronize()
    red_ms = e0.elapsed_time(e1)
    clocks = cs.summary()
    mean_n = float(to_bitmap(eng.download_block(), scene.border).mean()) if rank == 0 else 0.0
    t = torch.tensor([st["ms_total"] + red_ms, st["ms_total"], red_ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n = torch.tensor([st["camera_samples"], st["closest_rays"] + st["shadow_rays"]], dtype=torch.float64, device="cuda")
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    one = None
    if rank == 0:
        cs1 = ClockSampler(gpu); cs1.start()
        _, s1 = eng.render(sample_first=0, sample_last=one_gpu_spp, clear=True, download=False)
        one = {"spp": one_gpu_spp, "ms": s1["ms_total"], "Msamples_per_s": s1["camera_samples"] / s1["ms_total"] / 1e3, "clocks": cs1.summary(),
               "image_mean": float(to_bitmap(eng.download_block(), scene.border).mean())}
    dist.barrier(); torch.cuda.synchronize()
    if rank != 0:
        return None
    ms, render_ms, red_ms = [float(x) for x in t.tolist()]
    samples, rays = [float(x) for x in n.tolist()]
    ms_one_full = one["ms"] * total_spp / one_gpu_spp
    return {"workload": name, "scaling": "strong", "n_gpus": world, "spp_total": total_spp, "spp_per_gpu": last - first, "ms": ms,
            "render_ms_max_over_ranks": render_ms, "reduce_ms": red_ms, "reduce_bytes_per_rank": int(nfl) * 4,
            "Msamples_per_s": samples / ms / 1e3, "Mrays_per_s": rays / ms / 1e3, "image_mean": mean_n, "clocks": clocks,
            "one_gpu": one, "one_gpu_ms_for_spp_total": ms_one_full,
            "one_gpu_note": ("rank 0 alone rendered the whole frame at spp_total" if one_gpu_spp == total_spp else
                             f"rank 0 alone rendered {one_gpu_spp} of the {total_spp} spp (what the time budget of a default run allows); its time is scaled by {total_spp // one_gpu_spp}: the single-GPU throughput does not depend on spp at this size"),
            "speedup": ms_one_full / ms, "efficiency": ms_one_full / ms / world}


def measure_strong(eng, torch, dist, rank, world, gpu, args):
    """Strong scaling of the two sharded BASELINE configs: C3 at every N > 1, C5 at N = 8 (or --c5)."""
    out = {}
    c3 = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "table_mis.npz"))
    r = strong_leg(eng, torch, dist, c3, 4096, 4096, rank, world, gpu, "C3: table_mis 800x600, path_mis, 4096 spp in total")
    if r:
        out["C3"] = r
    if world == 8 or args.c5:
        c5 = resize_camera(load_scene(), 3840, 2160)
        r = strong_leg(eng, torch, dist, c5, 16384, 16384 // 8, rank, world, gpu, "C5: Cornell box path_mis 3840x2160, 16 384 spp in total")
        if r:
            out["C5"] = r
    return out


def rel_mse(a, b, eps=1e-2):
    return float(np.mean((a - b) ** 2 / (b ** 2 + eps)))


def to_bitmap(block, border):
    """ImageBlock::toBitmap (src/block.cpp:57-63): rgb / w, border cropped."""
    rgb = block[border:-border, border:-border, :3] if border else block[..., :3]
    w = block[border:-border, border:-border, 3:4] if border else block[..., 3:4]
    return np.where(w != 0, rgb / np.where(w != 0, w, 1), 0)


# ------------------------------------------------------------------------------------------------
def timed_render(eng, scene, spp, warm_spp, gpu):
    """Loads `scene`, one warm-up render, one timed render under a clock sampler. Device time from the call's own CUDA events."""
    t0 = time.perf_counter(); eng.load_scene(scene); first_ms = (time.perf_counter() - t0) * 1e3     # incl. first-use allocations (pinned staging ring)
    t0 = time.perf_counter(); accel = eng.load_scene(scene); load_ms = (time.perf_counter() - t0) * 1e3
    eng.set_profiling(True)
    eng.render(sample_first=0, sample_last=warm_spp, download=False)
    cs = ClockSampler(gpu); cs.start()
    t0, reps, ms_sum = time.perf_counter(), 0, 0.0
    while reps == 0 or (time.perf_counter() - t0 < 1.5 and reps < 200):      # short renders repeat for ~1.5 s so that the clock sampler sees them
        _, st = eng.render(sample_first=0, sample_last=spp, download=False)
        ms_sum += st["ms_total"]; reps += 1
    clocks = cs.summary()
    ms = ms_sum / reps
    return {"spp": spp, "ms": ms, "renders_averaged": reps, "Msamples_per_s": st["camera_samples"] / ms / 1e3, "Mrays_per_s": (st["closest_rays"] + st["shadow_rays"]) / ms / 1e3,
            "Gsegments_per_s": st["segments"] / ms / 1e6, "segments_per_sample": st["segments"] / max(st["camera_samples"], 1),
            "stages_ms": {"closest+splat": st["ms_trace_closest"], "shade": st["ms_shade"], "shadow": st["ms_trace_shadow"]},
            "iterations": st["iterations"], "gpu_launches": st["kernel_launches"], "triangles": accel["n_triangles"], "nodes": accel["n_nodes"],
            "build_ms_device": accel["build_ms"], "upload_and_build_ms_wall": load_ms, "upload_and_build_ms_wall_first_call": first_ms, "sah_cost": accel["sah_cost"], "clocks": clocks}


def measure_pathgraph(eng, gpu, spp=2, k=16, iterations=3):
    """SURVEY 8f row 4 — the fork's path-graph tool (src/pbsdf.cu, src/cluster.cpp) on the veach scene's own path vertices, every stage
    through the C ABI under the clock sampler. Device times are the calls' own CUDA events (records and rows resident); *_wall_ms are whole
    calls from / to host buffers. The reference's drivers on the same GPU are builder-run only (scripts/bench_pathgraph.py ->
    profiles/pathgraph_r02g.json): bench.py does not execute oracle/_ref outside its cpu_baseline / --impl reference legs."""
    from nori_b200.gpu import pathgraph_grid, pathgraph_stride_seeds
    scene = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "veach_mis.npz"))
    eng.load_scene(scene)
    med = lambda xs: float(np.median(xs[1:])) if len(xs) > 1 else float(xs[0])
    v, lp = eng.path_vertices(sample_first=0, sample_last=spp, lights=True)
    hp = pathgraph_grid(v, k, k)
    seeds = pathgraph_stride_seeds(len(v), len(v) // k + 1)
    cs = ClockSampler(gpu); cs.start()
    t = {"knn": [], "scatter": [], "mis": [], "direct": [], "build": [], "cluster": [], "render_mis": [], "render_cluster": []}
    for _ in range(3):
        nb, knn, dknn, ms = eng.pathgraph_knn(v, hp); t["knn"].append(ms)
        t["scatter"].append(eng.pathgraph_scatter(v, nb, knn, iterations)[3])
        t["mis"].append(eng.pathgraph_mis(v, nb, knn, iterations)[2])
        t["direct"].append(eng.pathgraph_direct(v, lp, nb, dknn, knn=knn, iterations=iterations)[3])
        of, members, offsets, ms = eng.pathgraph_build_clusters(v, hp, seeds); t["build"].append(ms)
        t["cluster"].append(eng.pathgraph_cluster(v, lp, members, offsets, iterations, cluster_of=of)[3])
        for mode in ("mis", "cluster"):
            t0 = time.perf_counter(); img, st = eng.pathgraph_render(k, iterations, mode=mode, sample_first=0, sample_last=spp)
            t["render_" + mode].append((time.perf_counter() - t0) * 1e3)
    clocks = cs.summary()
    sizes = np.diff(np.append(offsets, len(members)))
    return {"workload": f"F4: path-graph tool on veach path_mis {scene.width}x{scene.height}, {spp} spp: {len(v)} path vertices, k = {k}, {iterations} passes",
            "vertices": int(len(v)), "edges": int(knn.sum()), "clusters": int(len(offsets)), "cluster_pairs": int((sizes.astype(np.int64) ** 2).sum()),
            "device_ms": {"hash_grid_knn": med(t["knn"]), "scatter": med(t["scatter"]), "mis_gather": med(t["mis"]), "direct_filter_and_scatter": med(t["direct"]),
                          "cluster_build": med(t["build"]), "cluster_direct_and_scatter": med(t["cluster"])},
            "Mpoints_per_s_knn": len(v) / med(t["knn"]) / 1e3,
            "tool_end_to_end_wall_ms": {"knn_mode": med(t["render_mis"]), "default_cluster_mode": med(t["render_cluster"])},
            "tool_end_to_end_device_ms_cluster_mode": [st["ms_vertices"], st["ms_knn"], st["ms_stage"]],
            "image_mean_full": float(img["full"].mean()), "clocks": clocks}


def measure_configs(eng, torch, gpu, hbm):
    """The other single-GPU BASELINE configs through the same C ABI (outside the timed region of the headline)."""
    from nori_b200 import synth
    out = {}
    c1 = synth.ajax_class_scene()
    r = timed_render(eng, c1, 32, 8, gpu)
    r["workload"] = "C1: 544 968-triangle stand-in for ajax.obj (absent from the reference tree), normals, 768x768, 32 spp: primary-ray traversal only"
    out["C1"] = r
    c3 = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "table_mis.npz"))
    r = timed_render(eng, c3, 4096, 128, gpu)
    r["workload"] = "C3: scenes/pa5/table/table_mis.xml (22 766 triangles, microfacet + dielectric), path_mis, 800x600, 4096 spp on ONE GPU"
    out["C3"] = r
    c4 = synth.terrain_scene()
    r = timed_render(eng, c4, 4, 1, gpu)
    r["workload"] = "C4: 9 999 392-triangle height field + area light, primary + shadow rays (PRIMARY_SHADOW), 3840x2160, 4 spp"
    r["h2d_bytes"] = int(sum(m.V.nbytes + m.F.nbytes + (m.N.nbytes if m.N is not None else 0) for m in c4.meshes))
    r["traversal_kernel_only"] = trace_only(eng, torch, c4, gpu, hbm)
    out["C4"] = r
    out["F4"] = measure_pathgraph(eng, gpu)
    return out


def c4_rays(eng, scene):
    """The three ray sets of the C4 traversal stress: primary rays in the renderer's issue order, any-hit rays from their hit points
    towards the light, grazing bounce rays off the terrain."""
    W, H = scene.width, scene.height
    rng = np.random.default_rng(8)
    ys, xs = np.mgrid[0:H, 0:W]
    ps = np.stack([xs + 0.5, ys + 0.5], -1).astype(np.float32)
    ps = ps.reshape(H // 4, 4, W // 8, 8, 2).transpose(0, 2, 1, 3, 4).reshape(-1, 2)          # 8x4 pixel patches per warp, the renderer's issue order
    primary = eng.camera_rays(ps)
    tri, _, tuv = eng.trace(primary)
    hit = tri != 0xffffffff
    P = primary[hit, 0:3] + tuv[hit, 0:1] * primary[hit, 4:7]
    light = np.array([0.0, 1.2, 0.0], np.float32)
    d = light - P; dist = np.linalg.norm(d, axis=1, keepdims=True)
    shadow = np.concatenate([P, np.full((len(P), 1), 1e-4, np.float32), d / dist, dist - 1e-4], 1).astype(np.float32)
    v = rng.normal(size=P.shape).astype(np.float32); v /= np.linalg.norm(v, axis=1, keepdims=True)
    v[:, 1] = np.abs(v[:, 1]) * 0.3; v /= np.linalg.norm(v, axis=1, keepdims=True)           # grazing bounce rays over the terrain
    bounce = np.concatenate([P, np.full((len(P), 1), 1e-4, np.float32), v, np.full((len(P), 1), np.inf, np.float32)], 1).astype(np.float32)
    return primary, shadow, bounce


def trace_only(eng, torch, scene, gpu, hbm):
    """C4 traversal stress: the Accel::rayIntersect operator on rays resident in HBM (nori_gpu_trace_device, CUDA
    events, L2 flushed between launches): coherent primary rays, any-hit rays towards the light, incoherent bounce rays.
    N_box / N_tri are counted by the device BVH on a 1 Mi-ray subset; algorithmic bytes = 48 + 32 N_box + 36 N_tri."""
    rng = np.random.default_rng(7)
    primary, shadow, bounce = c4_rays(eng, scene)
    ncu = {}
    pth = os.path.join(ROOT, "profiles", "c4_trace_dram.json")
    if os.path.exists(pth):
        ncu = json.load(open(pth))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    res = {"hbm_peak_gbs": hbm}
    cs = ClockSampler(gpu); cs.start()
    for name, rays, sh in (("primary_closest", primary, False), ("shadow_anyhit", shadow, True), ("bounce_closest", bounce, False)):
        n = len(rays)
        sub = rays[rng.choice(n, 1 << 20, replace=False)] if n > (1 << 20) else rays
        nb, nt = eng.trace_count(sub, shadow=sh)
        bpr = 48 + 32 * (2 * nb / len(sub)) + 36 * (nt / len(sub))                              # two child boxes per node visit
        dr = torch.from_numpy(rays).cuda(); dh = torch.empty((n, 4), dtype=torch.float32, device="cuda")
        ms = []
        for it in range(40):
            flush.zero_(); torch.cuda.synchronize()
            t = eng.trace_device(dr.data_ptr(), n, dh.data_ptr(), shadow=sh)
            if it >= 3:
                ms.append(t)
        t = float(np.mean(ms))
        e = {"rays": n, "launches_averaged": len(ms), "ms": t, "Mrays_per_s": n / t / 1e3, "n_box_per_ray": 2 * nb / len(sub), "n_tri_per_ray": nt / len(sub),
             "algorithmic_bytes_per_ray": bpr, "algorithmic_gbs": n * bpr / t / 1e6, "algorithmic_frac_of_hbm": n * bpr / t / 1e6 / hbm}
        k = ncu.get("rays", {}).get(name)
        if k:   # DRAM bytes per ray of the same launch under ncu (dram__bytes_read.sum + dram__bytes_write.sum), applied to the live time
            e["dram_bytes_per_ray_ncu"] = k["dram_bytes_per_ray"]
            e["dram_frac"] = k["dram_bytes_per_ray"] * n / t / 1e6 / hbm
        res[name] = e
        del dr, dh
    res["clocks"] = cs.summary()
    if ncu:
        res["dram_source"] = ncu.get("source")
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--ref-spp", type=int, default=96, help="spp of the bounded CPU sample")
    ap.add_argument("--pool", type=int, default=0)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (and rel_mse)")
    ap.add_argument("--c5", action="store_true", help="run the C5 strong-scaling leg at any N > 1 (default: N = 8 only)")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C3 / C4 legs (N = 1) and the strong-scaling legs (N > 1)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
