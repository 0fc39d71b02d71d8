"""Strong-scaling lines of BASELINE configs C3 and C5: a FIXED total sample count per pixel is split
into contiguous sample-index ranges over the ranks (nori_b200.shard.sample_range), each rank renders its range with the
per-sample streams and ONE NCCL reduce sums the ImageBlocks on rank 0.

    C3  scenes/pa5/table/table_mis.xml, 800x600, 4096 spp
    C5  Cornell box path_mis at 3840x2160, 16 384 spp (133 MB ImageBlock per rank)

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_strong.py --config c5 [--spp S]          (or plain `python scripts/bench_strong.py` for N = 1)

Timing: CUDA events inside nori_gpu_render (device time of the call) + CUDA events around the reduce, max over ranks;
one untimed warm-up render of 1/8 of the range first. Prints one JSON line on rank 0."""
import copy, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
from nori_b200.shard import sample_range


class _CudaArray:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}


def main():
    cfg = sys.argv[sys.argv.index("--config") + 1] if "--config" in sys.argv else "c3"
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if cfg == "c5":
        scene = copy.copy(FlatScene.load(os.path.join(ROOT, "tests/golden/scenes/cbox_mis.npz")))
        # same camera at 3840x2160: fov is horizontal (src/perspective.cpp:43), the aspect changes sampleToCamera (resize below)
        spp, name = 16384, "C5 cbox path_mis 3840x2160"
    else:
        scene = FlatScene.load(os.path.join(ROOT, "tests/golden/scenes/table_mis.npz"))
        spp, name = 4096, "C3 table_mis 800x600"
    if "--spp" in sys.argv:
        spp = int(sys.argv[sys.argv.index("--spp") + 1])
    if cfg == "c5":
        scene = resize(scene, 3840, 2160)
    eng = Engine(local)
    accel = eng.load_scene(scene)
    ptr, nfl = eng.block_device_ptr()
    block_t = torch.as_tensor(_CudaArray(ptr, nfl), device="cuda")
    first, last = sample_range(rank, world, spp)
    eng.render(sample_first=first, sample_last=first + max(1, (last - first) // 8), download=False)       # warm-up
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _, st = eng.render(sample_first=first, sample_last=last, clear=True, download=False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if world > 1:
        dist.reduce(block_t, dst=0, op=dist.ReduceOp.SUM)
    e1.record(); e1.synchronize()
    red_ms = e0.elapsed_time(e1)
    img = eng.download_bitmap() if rank == 0 else None
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([st["ms_total"], red_ms, wall_ms], dtype=torch.float64, device="cuda")
    tot = torch.tensor([st["camera_samples"], st["closest_rays"] + st["shadow_rays"], st["segments"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    if rank == 0:
        render_ms, red_ms, wall_ms = [float(x) for x in t.tolist()]
        samples, rays, segs = [float(x) for x in tot.tolist()]
        ms = render_ms + red_ms
        print(json.dumps({"config": name, "scaling": "strong", "n_gpus": world, "spp_total": spp, "spp_per_gpu": last - first,
                          "triangles": scene.n_triangles, "ms": round(ms, 2), "render_ms_max_over_ranks": round(render_ms, 2), "reduce_ms": round(red_ms, 3),
                          "wall_ms_incl_download": round(wall_ms, 1), "reduce_bytes_per_rank": int(nfl) * 4,
                          "Msamples_per_s": round(samples / ms / 1e3, 1), "Mrays_per_s": round(rays / ms / 1e3, 1), "Gsegments_per_s": round(segs / ms / 1e6, 3),
                          "image_mean": float(img.mean()), "image_finite": bool((img == img).all()), "accel": accel}), flush=True)
    if world > 1:
        dist.destroy_process_group()


def resize(scene, width, height):
    """PerspectiveCamera::activate (src/perspective.cpp:53-86) for another output size: sampleToCamera depends on the aspect."""
    import numpy as np
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


main()
