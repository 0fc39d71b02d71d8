"""Iteration timeline of one render (sweep build only: NORI_GPU_LIB=libnori_b200_sweep.so NORI_GPU_DEBUG=1): where ramp-up and
the tail of a short render go.  python scripts/timeline.py [scene] [spp]   (the library prints one line per batch of 8 iterations)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("NORI_GPU_LIB", "libnori_b200_sweep.so"); os.environ.setdefault("NORI_GPU_DEBUG", "1")
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
name = sys.argv[1] if len(sys.argv) > 1 else "table_mis"; spp = int(sys.argv[2]) if len(sys.argv) > 2 else 512
s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", name + ".npz"))
e = Engine(0); e.load_scene(s); e.set_profiling(True)
pass
_, st = e.render(sample_first=0, sample_last=spp, download=False)
print("TOTAL", st)
