"""Golden vectors produced by the COMPILED REFERENCE (oracle/_ref: the reference's own sources
where they exist — parser, OBJ loader, camera, pcg32/independent sampler, Mesh::rayIntersect and the
brute-force Accel loop, diffuse/mirror BSDFs, ImageBlock, BlockGenerator, rfilter — plus the plugin
restatements the reference's chi2/t-test scene files validate). TEST INFRASTRUCTURE.

Runs only where /root/reference and oracle/_ref exist (the authoring container); the vectors are
committed under tests/golden/ref_vectors.npz so that the oracle port (CPU tests, anywhere) and the
CUDA path (GPU box, where /root/reference does not exist) are both checked against the reference.

    python -m oracle.make_golden
"""
import os

import numpy as np

from oracle.make_fixtures import SCENES
from oracle.pyoracle import Ref, REFERENCE_ROOT

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "ref_vectors.npz")
N_RAYS, N_SAMPLES, N_BSDF = 2100, 1500, 1500


def scene_rays(rs, fs, rng):
    px = (rng.random((N_RAYS, 2)) * [fs.width, fs.height]).astype(np.float32)
    cam = rs.camera_rays(px)
    V = np.concatenate([m.V for m in fs.meshes]); lo, hi = V.min(0), V.max(0)
    o = rng.uniform(lo, hi, size=(N_RAYS // 3, 3)).astype(np.float32)
    d = rng.normal(size=(N_RAYS // 3, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
    inc = np.empty((N_RAYS // 3, 8), np.float32)
    inc[:, :3] = o; inc[:, 3] = 1e-4; inc[:, 4:7] = d; inc[:, 7] = np.inf
    return px, cam, np.concatenate([cam, inc])


def main():
    ref = Ref()
    g = {}
    # pcg32 (ext/pcg32/pcg32.h) and Independent::next2D order (src/independent.cpp:62-67)
    seeds = np.array([[42, 54, 0], [0, 0, 0], [17, 5, 1000], [479999, 1023, 7], [2**40 + 3, 2**33 + 1, 123456789]], np.int64)
    g["pcg32_seeds"] = seeds
    g["pcg32_uints"] = np.stack([ref.pcg32_uints(int(a), int(b), 64, advance=int(c)) for a, b, c in seeds])
    g["next2d_block_32_64"] = ref.sampler_next2d(32, 64, 16)
    g["block_order_800x600"] = ref.block_order(800, 600)
    g["block_order_75x41"] = ref.block_order(75, 41)
    for name, prm in [("gaussian", (2.0, 0.5, 0.0)), ("mitchell", (2.0, 1 / 3, 1 / 3)), ("tent", (1.0, 0, 0)), ("box", (0.5, 0, 0))]:
        r, t = ref.filter_table(name, *prm)
        g[f"filter_{name}"] = np.concatenate([[r], t]).astype(np.float32)
    rng = np.random.default_rng(2026)
    c = rng.uniform(-1, 1, 4000).astype(np.float32)
    g["fresnel_cos"] = c
    g["fresnel_glass"] = ref.fresnel(c, 1.000277, 1.5046)
    nrm = rng.normal(size=(2000, 3)).astype(np.float32); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    g["frame_n"] = nrm; g["frame_st"] = ref.frames(nrm)

    for name in ["cbox_mis", "bunny_normals", "table_mis", "veach_mis", "cbox_whitted"]:
        rs = ref.load_xml(os.path.join(REFERENCE_ROOT, SCENES[name]))
        fs = rs.flatten()
        px, cam, rays = scene_rays(rs, fs, rng)
        g[f"{name}/pixel_samples"] = px
        g[f"{name}/camera_rays"] = cam
        g[f"{name}/rays"] = rays
        brute = fs.n_triangles <= 3000           # the reference's O(N) loop; SAH BVH of the glue beyond that
        tri, mesh, tuv = rs.trace(rays, shadow=False, use_bvh=not brute)
        g[f"{name}/hit_tri"], g[f"{name}/hit_mesh"], g[f"{name}/hit_tuv"] = tri, mesh, tuv
        g[f"{name}/hit_brute_force"] = np.array(brute)
        srays = rays.copy(); srays[:, 7] = rng.uniform(0.05, 1.5, size=rays.shape[0]).astype(np.float32)
        g[f"{name}/shadow_rays"] = srays
        g[f"{name}/shadow_occluded"] = rs.trace(srays, shadow=True, use_bvh=not brute)[0] != 0xffffffff
        its, valid = rs.intersect(rays[:1200], use_bvh=not brute)
        g[f"{name}/its"], g[f"{name}/its_valid"] = its, valid
        pixel = rng.integers(0, fs.width * fs.height, size=N_SAMPLES).astype(np.uint32)
        sample = rng.integers(0, 4096, size=N_SAMPLES).astype(np.uint32)
        rgb, pos = rs.sample_radiance(pixel, sample, use_bvh=True)
        g[f"{name}/li_pixel"], g[f"{name}/li_sample"], g[f"{name}/li_rgb"], g[f"{name}/li_pos"] = pixel, sample, rgb, pos
        # BSDF::sample / eval / pdf of every distinct BSDF in the scene
        seen = set()
        for mi, m in enumerate(fs.meshes):
            key = (m.bsdf_type, tuple(np.round(m.bsdf_params, 6)))
            if key in seen:
                continue
            seen.add(key)
            wi = rng.normal(size=(N_BSDF, 3)).astype(np.float32)
            if m.bsdf_type != 2:
                wi[:, 2] = np.abs(wi[:, 2])
            wi /= np.linalg.norm(wi, axis=1, keepdims=True)
            xi = rng.random((N_BSDF, 2)).astype(np.float32)
            wo2 = rng.normal(size=(N_BSDF, 3)).astype(np.float32); wo2[:, 2] = np.abs(wo2[:, 2]); wo2 /= np.linalg.norm(wo2, axis=1, keepdims=True)
            wo, w, pdf, eta, meas = rs.bsdf_sample(mi, wi, xi)
            f, p = rs.bsdf_eval_pdf(mi, wi, wo2)
            pre = f"{name}/bsdf{mi}"
            g[pre + "_type"] = np.array(m.bsdf_type); g[pre + "_params"] = m.bsdf_params
            g[pre + "_wi"], g[pre + "_xi"], g[pre + "_wo2"] = wi, xi, wo2
            g[pre + "_wo"], g[pre + "_w"], g[pre + "_pdf"], g[pre + "_eta"], g[pre + "_meas"] = wo, w, pdf, eta, meas
            g[pre + "_f"], g[pre + "_p"] = f, p
    # ImageBlock::put on a 96x64 frame incl. invalid samples (src/block.cpp:74-103)
    n = 6000
    pos = (rng.random((n, 2)) * [96, 64]).astype(np.float32)
    rgb = (rng.random((n, 3)) * 3).astype(np.float32)
    rgb[5] = [np.nan, 1, 1]; rgb[6] = [-1, 0, 0]; rgb[7] = [np.inf, 0, 0]
    pos[8] = [0.0, 0.0]; pos[9] = [95.99999, 63.99999]
    g["splat_pos"], g["splat_rgb"] = pos, rgb
    g["splat_gaussian_96x64"] = ref.splat(96, 64, "gaussian", pos, rgb, 2)
    # a whole frame through the reference's own renderBlock/ImageBlock loop, block stream (normals)
    rs = ref.load_xml(os.path.join(REFERENCE_ROOT, SCENES["bunny_normals"]))
    fs = rs.flatten()
    blk, _ = rs.render(fs.block_shape(), mode=0, sample_first=0, sample_last=1, threads=0, use_bvh=True)
    g["bunny_normals/block_stream_1spp_bitmap_u8"] = np.clip(np.where(blk[2:-2, 2:-2, 3:] > 0, blk[2:-2, 2:-2, :3] / np.maximum(blk[2:-2, 2:-2, 3:], 1e-20), 0) * 255 + 0.5, 0, 255).astype(np.uint8)[::4, ::4]
    g["bunny_normals/block_stream_1spp_sum"] = blk.sum(axis=(0, 1)).astype(np.float64)
    np.savez_compressed(OUT, **g)
    print(f"{OUT}: {len(g)} arrays, {os.path.getsize(OUT) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
