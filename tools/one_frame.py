"""Render the C4 frame a few times (for ncu captures: short, no benchmarking around it).  usage: python tools/one_frame.py dense|culled|half"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "python-ray-tracer_b200")); sys.path.insert(0, ROOT)
from b200rt import render_host, _lib
from b200rt.workloads import workload

mode = sys.argv[1] if len(sys.argv) > 1 else "culled"
w = workload("c4"); cam = w["cameras"][0]; W, H = cam.resolution
sp, li, pl = w["scene"].generate_scene()
args = (None, cam.position, cam.rotation, sp, li, pl, w["amb"], w["lamb"], w["refl"], w["refl_depth"], w["aliasing"])
flags = _lib.F_DENSE if mode == "dense" else 0
strips = (16, 0, 8) if mode == "eighth" else None
for _ in range(2 if mode == "dense" else 3):
    out = render_host(*args, grid=cam.grid(), shape=(W, H), flags=flags | _lib.F_NO_GRAPH, strips=strips)
print(mode, int(out.sum()))
