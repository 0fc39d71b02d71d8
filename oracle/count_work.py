"""Per-ray traversal work of the ORACLE's BVH (binary SAH, <= 4 triangles per leaf) on the rays a
render actually traces (TEST INFRASTRUCTURE; SURVEY.md 8(d) "algorithmic bytes per unit of work").

    python -m oracle.count_work          -> profiles/algorithmic_bytes.json

B_ray = 32 (ray in) + 16 (hit out) + 32 * N_box + 36 * N_tri, N_box / N_tri = ray-box and
ray-triangle tests per ray counted by oracle/noracle_scene.h Accel::trace. bench.py multiplies these
per-ray figures by the rays one launch of the traversal kernel processes.
"""
import json
import os

from nori_b200.scene import FlatScene
from oracle.pyoracle import Port

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def main():
    port = Port(build=True)
    out = {"definition": "B_ray = 32 + 16 + 32*N_box + 36*N_tri; B_segment = B_ray(closest) + shadow_per_closest*B_ray(shadow) + 128; "
                         "B_sample = segments_per_sample*B_segment + 16",
           "counted_by": "oracle/noracle_scene.h Accel::trace (binary SAH BVH, <= 4 tris/leaf), 1 spp of the full frame, PER_SAMPLE streams",
           "scenes": {}}
    for name in ["cbox_mis", "table_mis", "veach_mis", "bunny_normals"]:
        s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", name + ".npz"))
        ps = port.scene(s)
        c, sh = ps.render_counted(1, 0, 1, stride=1)
        samples = s.width * s.height
        nb, nt = c[1] / c[0], c[2] / c[0]
        sb, stt = (sh[1] / sh[0], sh[2] / sh[0]) if sh[0] else (0.0, 0.0)
        b_closest = 48 + 32 * nb + 36 * nt
        b_shadow = 48 + 32 * sb + 36 * stt
        spc = sh[0] / c[0]
        b_seg = b_closest + spc * b_shadow + 128
        seg_per_sample = c[0] / samples
        out["scenes"][name] = {
            "integrator": int(s.integrator), "closest": {"n_box": nb, "n_tri": nt, "bytes_per_ray": b_closest},
            "shadow": {"n_box": sb, "n_tri": stt, "bytes_per_ray": b_shadow},
            "shadow_per_closest": spc, "segments_per_sample": seg_per_sample,
            "bytes_per_segment": b_seg, "bytes_per_sample": seg_per_sample * b_seg + 16}
        print(name, json.dumps(out["scenes"][name]))
    with open(os.path.join(ROOT, "profiles", "algorithmic_bytes.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
