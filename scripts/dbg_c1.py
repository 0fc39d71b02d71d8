import sys; sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from nori_b200 import synth
from nori_b200.gpu import Engine
from oracle.pyoracle import Port
s = synth.ajax_class_scene(); e = Engine(0); print(e.load_scene(s)); ps = Port(build=False).scene(s)
want, _ = ps.render(mode=0, sample_first=0, sample_last=1)
got, st = e.render(stream_mode=0, sample_first=0, sample_last=1)
bad = np.argwhere(~np.isclose(got, want, rtol=2e-5, atol=2e-6))
print(len(bad), bad[:40].tolist())
ys, xs = bad[:, 0], bad[:, 1]
print('bbox', ys.min(), ys.max(), xs.min(), xs.max())
# all pixel samples of the block stream in that neighbourhood: compare per-sample through trace
cy, cx = int(ys.mean()) - 2, int(xs.mean()) - 2
pix = [(y, x) for y in range(cy - 3, cy + 4) for x in range(cx - 3, cx + 4)]
from oracle.pyoracle import Port
P = Port(build=False)
pos = []
for y, x in pix:
    ox, oy = (x // 32) * 32, (y // 32) * 32
    bw = min(768 - ox, 32)
    n = ((y - oy) * bw + (x - ox)) * 1
    f = P.pcg32_floats(ox, oy, 2, advance=4 * n)
    pos.append([x + f[1], y + f[0]])
pos = np.array(pos, np.float32)
rays = ps.camera_rays(pos)
g = e.trace(rays); o = ps.trace(rays, use_bvh=True); b = ps.trace(rays, use_bvh=False)
for k in range(len(pix)):
    if g[0][k] != o[0][k] or g[0][k] != b[0][k]:
        print(pix[k], 'gpu', g[0][k], g[2][k], 'obvh', o[0][k], o[2][k], 'brute', b[0][k], b[2][k])
gi, gv = e.intersect(rays); oi, ov = ps.intersect(rays)
d = np.argwhere(gi.view(np.uint32) != oi.view(np.uint32))
print('its diffs', d[:20].tolist())
for k in sorted(set(d[:, 0].tolist()))[:3]:
    print(pix[k], gi[k], oi[k])
