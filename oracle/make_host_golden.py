"""Generates tests/golden/host/: scene descriptions of our own making that exercise every tag, transform
operation, BSDF, filter and OBJ feature the parser knows, plus what the COMPILED REFERENCE (oracle/_ref,
its own parser.cpp / obj.cpp / perspective.cpp / rfilter.cpp, built in this container) makes of them.
The C++ host front end must reproduce those descriptors bit for bit (tests/test_host_cpp.py).

Run here (needs /root/reference for oracle/_ref):  python oracle/make_host_golden.py
TEST INFRASTRUCTURE ONLY."""
import math, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.pyoracle import Ref

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "host")

def write_plane_obj(path):          # quads, texture coordinates and normals, shared corners
    n = 5
    with open(path, "w") as f:
        f.write("# 5x5 grid of quads with vt and vn\n")
        for j in range(n + 1):
            for i in range(n + 1):
                f.write(f"v {i / n - 0.5:.6f} {0.02 * math.sin(3.0 * i + j):.6f} {j / n - 0.5:.6f}\n")
        for j in range(n + 1):
            for i in range(n + 1):
                f.write(f"vt {i / n:.6f} {j / n:.6f}\n")
        for j in range(n + 1):
            for i in range(n + 1):
                f.write(f"vn {0.1 * math.cos(i):.6f} 1.0 {0.1 * math.sin(j):.6f}\n")      # not unit length on purpose
        for j in range(n):
            for i in range(n):
                a = j * (n + 1) + i + 1; b = a + 1; c = a + n + 2; d = a + n + 1
                f.write(f"f {a}/{a}/{a} {b}/{b}/{b} {c}/{c}/{c} {d}/{d}/{d}\n")

def write_tetra_obj(path):          # triangles only, no vt / vn, positions reused by several faces
    with open(path, "w") as f:
        f.write("v 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\n\nf 1 3 2\nf 1 2 4\nf 2 3 4\nf 3 1 4\n")

def write_ball_obj(path):           # lat-long sphere with normals (v//vn form), triangles at the poles, quads elsewhere
    nu, nv = 12, 8
    with open(path, "w") as f:
        pts = [(0.0, 1.0, 0.0)]
        for j in range(1, nv):
            th = math.pi * j / nv
            for i in range(nu):
                ph = 2 * math.pi * i / nu
                pts.append((math.sin(th) * math.cos(ph), math.cos(th), math.sin(th) * math.sin(ph)))
        pts.append((0.0, -1.0, 0.0))
        for p in pts: f.write("v %.7f %.7f %.7f\n" % p)
        for p in pts: f.write("vn %.7f %.7f %.7f\n" % p)
        def v(k): return f"{k}//{k}"
        ring = lambda j, i: 2 + (j - 1) * nu + (i % nu)
        for i in range(nu): f.write(f"f {v(1)} {v(ring(1, i + 1))} {v(ring(1, i))}\n")
        for j in range(1, nv - 1):
            for i in range(nu): f.write(f"f {v(ring(j, i))} {v(ring(j, i + 1))} {v(ring(j + 1, i + 1))} {v(ring(j + 1, i))}\n")
        last = len(pts)
        for i in range(nu): f.write(f"f {v(last)} {v(ring(nv - 1, i))} {v(ring(nv - 1, i + 1))}\n")

SCENES = {
"all_tags.xml": """<?xml version="1.0" encoding="utf-8"?>
<!-- every tag and transform operation; mitchell filter; path_mis with explicit termination -->
<scene>
    <integrator type="path_mis">
        <integer name="rrStart" value="5"/>
        <float name="rrMax" value="0.95"/>
        <integer name="maxDepth" value="12"/>
    </integrator>
    <sampler type="independent"><integer name="sampleCount" value="48"/></sampler>
    <camera type="perspective">
        <transform name="toWorld">
            <scale value="-1,1,1"/>
            <lookat target="0.1, 0.3, 0.2" origin="2.5, 1.75, -3.25" up="0.1, 1, 0"/>
        </transform>
        <float name="fov" value="37.5"/>
        <float name="nearClip" value="0.01"/>
        <float name="farClip" value="250"/>
        <integer name="width" value="200"/>
        <integer name="height" value="120"/>
        <rfilter type="mitchell"><float name="radius" value="1.5"/><float name="B" value="0.4"/><float name="C" value="0.3"/></rfilter>
    </camera>
    <mesh type="obj">
        <string name="filename" value="plane.obj"/>
        <transform name="toWorld">
            <scale value="4, 1.5, 3"/>
            <rotate angle="33" axis="0.2, 1, 0.1"/>
            <translate value="0.25, -0.5, 0.125"/>
        </transform>
        <bsdf type="diffuse"><color name="albedo" value="0.725 0.71 0.68"/></bsdf>
    </mesh>
    <mesh type="obj">
        <string name="filename" value="ball.obj"/>
        <transform name="toWorld">
            <scale value="0.3, 0.45, 0.3"/>
            <matrix value="0.8 0 0.6 0.1  0 1 0 0.2  -0.6 0 0.8 0.3  0 0 0 1"/>
            <rotate angle="-75.5" axis="1, 0, 0"/>
        </transform>
        <bsdf type="microfacet"><float name="alpha" value="0.23"/><color name="kd" value="0.2, 0.3, 0.1"/><float name="intIOR" value="1.7"/></bsdf>
    </mesh>
    <mesh type="obj">
        <string name="filename" value="ball.obj"/>
        <transform name="toWorld"><scale value="0.2, 0.2, 0.2"/><translate value="-0.75, 0.1, 0.5"/></transform>
        <bsdf type="dielectric"><float name="extIOR" value="1.33"/></bsdf>
    </mesh>
    <mesh type="obj">
        <string name="filename" value="tetra.obj"/>
        <transform name="toWorld"><translate value="0.5, 0, -0.5"/></transform>
        <bsdf type="mirror"/>
    </mesh>
    <mesh type="obj">
        <string name="filename" value="tetra.obj"/>
        <transform name="toWorld"><scale value="0.5,0.1,0.5"/><translate value="0, 2, 0"/></transform>
        <emitter type="area"><color name="radiance" value="17 12 8.5"/></emitter>
    </mesh>
</scene>
""",
"defaults.xml": """<scene>
    <!-- defaults everywhere: gaussian filter, independent sampler with one sample, diffuse 0.5 -->
    <integrator type="normals"/>
    <camera type="perspective"/>
    <mesh type="obj"><string name="filename" value="tetra.obj"/></mesh>
</scene>
""",
"simple_tent.xml": """<scene>
    <integrator type="simple"><point name="position" value="-20, 40, 20"/><color name="energy" value="3.76e4, 3.76e4, 3.76e4"/></integrator>
    <sampler type="independent"><integer name="sampleCount" value="7"/></sampler>
    <camera type="perspective">
        <transform name="toWorld"><lookat origin="0 1 -4" target="0 0 0" up="0 1 0"/></transform>
        <integer name="width" value="96"/><integer name="height" value="160"/>
        <rfilter type="tent"/>
    </camera>
    <mesh type="obj"><string name="filename" value="ball.obj"/></mesh>
    <mesh type="obj"><string name="filename" value="plane.obj"/><transform name="toWorld"><scale value="6 1 6"/><translate value="0 -1 0"/></transform></mesh>
</scene>
""",
"whitted_box.xml": """<scene>
    <integrator type="whitted"/>
    <camera type="perspective">
        <transform name="toWorld"><lookat origin="1 2 3" target="0 0 0" up="0 0 1"/></transform>
        <integer name="width" value="64"/><integer name="height" value="64"/><float name="fov" value="60"/>
        <rfilter type="box"/>
    </camera>
    <mesh type="obj"><string name="filename" value="ball.obj"/><bsdf type="dielectric"/></mesh>
    <mesh type="obj"><string name="filename" value="tetra.obj"/><transform name="toWorld"><scale value="3 3 3"/><translate value="-1 -1 -2"/></transform>
        <emitter type="area"><color name="radiance" value="5 5 5"/></emitter></mesh>
</scene>
""",
}

def main():
    os.makedirs(OUT, exist_ok=True)
    write_plane_obj(os.path.join(OUT, "plane.obj")); write_tetra_obj(os.path.join(OUT, "tetra.obj")); write_ball_obj(os.path.join(OUT, "ball.obj"))
    ref = Ref(build=True)
    for name, text in SCENES.items():
        path = os.path.join(OUT, name)
        open(path, "w").write(text)
        flat = ref.load_xml(path).flatten()
        flat.save(os.path.join(OUT, name.replace(".xml", ".npz")))
        print(name, len(flat.meshes), "meshes", flat.n_triangles, "triangles")

if __name__ == "__main__":
    main()
