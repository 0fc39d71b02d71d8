"""One cbox path_mis render for profiling (GPU box): python scripts/prof_render.py [spp] [scene]."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
scene = sys.argv[2] if len(sys.argv) > 2 else "cbox_mis"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
s = FlatScene.load(os.path.join(root, "tests", "golden", "scenes", scene + ".npz"))
e = Engine(0); e.load_scene(s); e.set_profiling(True)
_, st = e.render(sample_first=0, sample_last=spp, download=False)
print({k: round(v, 2) if isinstance(v, float) else v for k, v in st.items()})
