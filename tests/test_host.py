"""Host-side logic: flattened scenes, sample-range sharding, the gloo stand-in for the NCCL reduce."""
import os

import numpy as np
import pytest

from conftest import SCENES, load_scene


def test_flat_scene_round_trip(tmp_path):
    from nori_b200.scene import FlatScene
    s = load_scene("cbox_mis")
    assert (s.width, s.height, s.n_triangles, len(s.meshes)) == (800, 600, 2188, 6)
    assert s.border == 2 and s.block_shape() == (604, 804, 4)
    p = str(tmp_path / "s.npz"); s.save(p); t = FlatScene.load(p)
    assert t.integrator == s.integrator and t.sample_count == s.sample_count
    for a, b in zip(s.meshes, t.meshes):
        np.testing.assert_array_equal(a.V, b.V); np.testing.assert_array_equal(a.F, b.F)
        assert (a.N is None) == (b.N is None) and a.bsdf_type == b.bsdf_type and a.is_emitter == b.is_emitter


def test_fixture_scenes_present():
    for n in ["bunny_normals", "cbox_mis", "cbox_ems", "cbox_mats", "cbox_whitted", "table_mis", "veach_mis"]:
        assert os.path.exists(os.path.join(SCENES, n + ".npz"))
    assert load_scene("table_mis").n_triangles == 22766


@pytest.mark.parametrize("spp,world", [(1024, 8), (10, 4), (3, 8), (7, 2)])
def test_sample_ranges_partition(spp, world):
    from nori_b200.shard import sample_range
    got = [sample_range(r, world, spp) for r in range(world)]
    assert got[0][0] == 0 and got[-1][1] == spp
    for (a, b), (c, d) in zip(got, got[1:]):
        assert b == c and a <= b
    assert max(b - a for a, b in got) - min(b - a for a, b in got) <= 1


def _worker(rank, world, port_file, spp, out):
    import torch
    import torch.distributed as dist
    from nori_b200.shard import reduce_block, sample_range
    from oracle.pyoracle import Port
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = port_file
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = load_scene("cbox_mis"); s.width, s.height = 48, 36
    ps = Port(build=False).scene(s)
    a, b = sample_range(rank, world, spp)
    blk, _ = ps.render(mode=1, sample_first=a, sample_last=b, threads=1)       # oracle stands in for the GPU render
    t = torch.from_numpy(blk.copy())
    reduce_block(t, dst=0)
    if rank == 0:
        np.save(out, t.numpy())
    dist.destroy_process_group()


def test_two_rank_reduce_equals_single_render(port, tmp_path):
    """N>1 host path with gloo, world_size 2: sample-index ranges per rank + one sum-reduce of the
    ImageBlock == the single-process render of the union (per-sample streams make them identical)."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "blk.npy")
    mp.spawn(_worker, args=(2, "29613", 6, out), nprocs=2, join=True)
    s = load_scene("cbox_mis"); s.width, s.height = 48, 36
    want, _ = port.scene(s).render(mode=1, sample_first=0, sample_last=6, threads=1)
    np.testing.assert_allclose(np.load(out), want, rtol=1e-5, atol=1e-6)


# ---- bench.py host logic (round 2) -------------------------------------------------------------------------------------
def test_bench_helpers():
    """bench.py pieces that need no GPU: the camera resize reproduces PerspectiveCamera::activate for the fixture's own size
    (so the C5 frame is the same camera at 3840x2160), rel-MSE / toBitmap, the host CPU record, and the issue model the roofline reads."""
    import json, os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import bench
    s = bench.load_scene()
    same = bench.resize_camera(s, s.width, s.height)
    np.testing.assert_allclose(same.sample_to_camera, s.sample_to_camera, rtol=2e-6, atol=1e-7)
    big = bench.resize_camera(s, 3840, 2160)
    assert (big.width, big.height) == (3840, 2160) and big.block_shape() == (2164, 3844, 4)
    blk = np.zeros((8, 8, 4), np.float32); blk[..., :3] = 2.0; blk[..., 3] = 4.0; blk[3, 3, 3] = 0.0
    bmp = bench.to_bitmap(blk, 2)
    assert bmp.shape == (4, 4, 3) and bmp[0, 0, 0] == 0.5 and bmp[1, 1, 0] == 0.0          # rgb / w, zero weight -> 0
    assert bench.rel_mse(bmp, bmp) == 0.0 and bench.rel_mse(bmp + 0.1, bmp) > 0
    info = bench.host_cpu_info()
    assert info["logical_cpus"] >= 1
    im = json.load(open(os.path.join(root, "profiles", "issue_model.json")))
    k = im["kernels"]["k_traceClosest"]
    assert 50 < k["warp_inst_per_ray"] < 400 and 1 <= k["lanes_per_inst"] <= 32 and k["dram_bytes_per_ray"] > 0
    assert {"k_shadeRegen", "k_traceShadow"} <= set(im["kernels"])
