"""Scratch driver for profiling the clustered path-graph mode: one build + one clustered scatter on veach path vertices."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nori_b200.gpu import Engine, pathgraph_grid, pathgraph_stride_seeds
from nori_b200.scene import FlatScene
s = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "veach_mis.npz"))
e = Engine(0); e.load_scene(s)
k = 16
v, lp = e.path_vertices(sample_first=0, sample_last=2, lights=True)
hp = pathgraph_grid(v, k, k)
of, members, offsets, t = e.pathgraph_build_clusters(v, hp, pathgraph_stride_seeds(len(v), len(v) // k + 1))
for _ in range(2):
    d, b, m, ms = e.pathgraph_cluster(v, lp, members, offsets, 3, cluster_of=of)
sizes = np.diff(np.append(offsets, len(members)))
print(len(v), len(offsets), ms, np.bincount(np.minimum((sizes + 7) // 8, 5)))
