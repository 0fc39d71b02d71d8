"""Time of the step before the path: nori_gpu_upload_meshes + nori_gpu_build_accel from host
buffers, for C4 (10 M triangles), C1 (545 k) and the Cornell box. GPU box: python scripts/bench_upload.py"""
import ctypes as C, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nori_b200.gpu import Engine, MeshDesc, AccelStats, _p
from nori_b200.scene import FlatScene
from nori_b200 import synth
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
e = Engine(0)
for name, s in (("cbox", FlatScene.load(os.path.join(ROOT, "tests/golden/scenes/cbox_mis.npz"))), ("C1", synth.ajax_class_scene()), ("C4", synth.terrain_scene())):
    descs = (MeshDesc * len(s.meshes))()
    for d, m in zip(descs, s.meshes):
        d.V, d.F, d.N, d.UV = _p(m.V), _p(m.F), _p(m.N), _p(m.UV)
        d.nV, d.nF = m.V.shape[0], m.F.shape[0]
        d.bsdf_type = int(m.bsdf_type); d.bsdf_params = (C.c_float * 6)(*[float(x) for x in m.bsdf_params])
        d.is_emitter = int(m.is_emitter); d.radiance = (C.c_float * 3)(*[float(x) for x in m.radiance])
    best = None
    for rep in range(3):
        t0 = time.perf_counter(); e._ck(e.L.nori_gpu_upload_meshes(e.h, descs, len(s.meshes))); t1 = time.perf_counter()
        st = AccelStats(); e._ck(e.L.nori_gpu_build_accel(e.h, C.byref(st))); t2 = time.perf_counter()
        cur = ((t1 - t0) * 1e3, (t2 - t1) * 1e3, st.build_ms)
        best = cur if best is None or cur[0] + cur[1] < best[0] + best[1] else best
    raw = sum(m.V.nbytes + m.F.nbytes + (m.N.nbytes if m.N is not None else 0) + (m.UV.nbytes if m.UV is not None else 0) for m in s.meshes)
    print(json.dumps({"scene": name, "triangles": s.n_triangles, "raw_MB": round(raw / 1e6, 1), "upload_ms": round(best[0], 2), "build_wall_ms": round(best[1], 2),
                      "build_device_ms": round(best[2], 2), "upload_GBps": round(raw / best[0] / 1e6, 2)}), flush=True)
