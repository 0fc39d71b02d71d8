"""Writes a FlatScene back out as a Nori scene description (scene.xml + one OBJ per mesh) so that the
COMPILED REFERENCE (oracle/_ref/libnori_ref.so: its own parser, OBJ loader, camera, TBB render loop)
can load it on machines where /root/reference/scenes does not exist (the GPU box).
TEST / BASELINE INFRASTRUCTURE ONLY (bench.py --impl reference, cpu_baseline kind "reference").

Geometry is written in world space with %.9g (float32 round-trips exactly), faces as i/i/i so that the
loader's corner de-duplication reproduces the vertex order; the camera's toWorld goes in as <matrix>.
Only the default reconstruction filter (gaussian, radius 2, stddev 0.5) is expressible from a FlatScene."""
from __future__ import annotations

import os

import numpy as np

from nori_b200.scene import FlatScene, INTEGRATOR_NAMES

_BSDF = {0: "diffuse", 1: "mirror", 2: "dielectric", 3: "microfacet"}


def _f(x) -> str:
    return "%.9g" % float(np.float32(x))


def write_scene_files(scene: FlatScene, out_dir: str) -> str:
    os.makedirs(out_dir, exist_ok=True)
    name = {v: k for k, v in INTEGRATOR_NAMES.items()}[scene.integrator]
    xml = ["<scene>", f'  <integrator type="{name}">',
           f'    <integer name="rrStart" value="{scene.rr_start}"/><float name="rrMax" value="{_f(scene.rr_max)}"/><integer name="maxDepth" value="{scene.max_depth}"/>']
    if name == "simple":
        xml.append('    <point name="position" value="%s"/><color name="energy" value="%s"/>' % (
            ", ".join(_f(v) for v in scene.light_pos), ", ".join(_f(v) for v in scene.light_energy)))
    xml += ["  </integrator>", f'  <sampler type="independent"><integer name="sampleCount" value="{scene.sample_count}"/></sampler>',
            '  <camera type="perspective">', '    <transform name="toWorld"><matrix value="%s"/></transform>' % " ".join(_f(v) for v in np.asarray(scene.camera_to_world).ravel()),
            f'    <float name="fov" value="{_f(scene.fov)}"/><float name="nearClip" value="{_f(scene.near_clip)}"/><float name="farClip" value="{_f(scene.far_clip)}"/>',
            f'    <integer name="width" value="{scene.width}"/><integer name="height" value="{scene.height}"/>', "  </camera>"]
    for i, m in enumerate(scene.meshes):
        obj = f"mesh{i}.obj"
        with open(os.path.join(out_dir, obj), "w") as f:
            for v in m.V: f.write("v %s %s %s\n" % tuple(_f(c) for c in v))
            if m.UV is not None:
                for t in m.UV: f.write("vt %s %s\n" % tuple(_f(c) for c in t))
            if m.N is not None:
                for n in m.N: f.write("vn %s %s %s\n" % tuple(_f(c) for c in n))
            for a, b, c in m.F + 1:
                if m.UV is not None and m.N is not None: f.write(f"f {a}/{a}/{a} {b}/{b}/{b} {c}/{c}/{c}\n")
                elif m.N is not None: f.write(f"f {a}//{a} {b}//{b} {c}//{c}\n")
                elif m.UV is not None: f.write(f"f {a}/{a} {b}/{b} {c}/{c}\n")
                else: f.write(f"f {a} {b} {c}\n")
        p = m.bsdf_params
        xml.append(f'  <mesh type="obj"><string name="filename" value="{obj}"/>')
        kind = _BSDF[int(m.bsdf_type)]
        if kind == "diffuse": xml.append('    <bsdf type="diffuse"><color name="albedo" value="%s, %s, %s"/></bsdf>' % (_f(p[0]), _f(p[1]), _f(p[2])))
        elif kind == "mirror": xml.append('    <bsdf type="mirror"/>')
        elif kind == "dielectric": xml.append('    <bsdf type="dielectric"><float name="intIOR" value="%s"/><float name="extIOR" value="%s"/></bsdf>' % (_f(p[4]), _f(p[5])))
        else: xml.append('    <bsdf type="microfacet"><color name="kd" value="%s, %s, %s"/><float name="alpha" value="%s"/><float name="intIOR" value="%s"/><float name="extIOR" value="%s"/></bsdf>'
                         % (_f(p[0]), _f(p[1]), _f(p[2]), _f(p[3]), _f(p[4]), _f(p[5])))
        if m.is_emitter: xml.append('    <emitter type="area"><color name="radiance" value="%s, %s, %s"/></emitter>' % tuple(_f(c) for c in m.radiance))
        xml.append("  </mesh>")
    xml.append("</scene>")
    path = os.path.join(out_dir, "scene.xml")
    open(path, "w").write("\n".join(xml) + "\n")
    return path
