"""Flatten reference scenes into committed fixtures (TEST INFRASTRUCTURE).

Runs only where /root/reference and oracle/_ref exist (the authoring container). The scenes
are parsed by the reference's OWN parser / OBJ loader / camera (oracle/_ref) and written as
tests/golden/scenes/<name>.npz (nori_b200.scene.FlatScene). The GPU box has no
/root/reference, so tests and bench.py read these files instead.

    python -m oracle.make_fixtures
"""
import os
import numpy as np
from oracle.pyoracle import Ref, REFERENCE_ROOT

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "scenes")
SCENES = {
    "bunny_normals": "scenes/pa1/bunny.xml",
    "cbox_mis": "scenes/pa5/cbox/cbox_mis.xml",
    "cbox_ems": "scenes/pa5/cbox/cbox_ems.xml",
    "cbox_mats": "scenes/pa5/cbox/cbox_mats.xml",
    "cbox_whitted": "scenes/pa4/cbox/cbox-whitted.xml",
    "table_mis": "scenes/pa5/table/table_mis.xml",
    "veach_mis": "scenes/pa5/veach_mi/veach_mis.xml",
}


def main():
    ref = Ref()
    os.makedirs(OUT, exist_ok=True)
    for name, rel in SCENES.items():
        fs = ref.load_xml(os.path.join(REFERENCE_ROOT, rel)).flatten()
        path = os.path.join(OUT, name + ".npz")
        fs.save(path)
        print(f"{name}: {len(fs.meshes)} meshes, {fs.n_triangles} tris, {fs.width}x{fs.height}, "
              f"spp {fs.sample_count}, integrator {fs.integrator}, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
