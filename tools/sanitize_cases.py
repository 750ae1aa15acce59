"""Small renders that touch every kernel family, for compute-sanitizer (one tool per gpurun call).  usage: python tools/sanitize_cases.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "python-ray-tracer_b200")); sys.path.insert(0, ROOT)
from b200rt import render_host, render_frame, _lib
from b200rt.scene import Scene, Camera


def synthetic(S, L, W, H, seed):
    rng = np.random.default_rng(seed)
    sp = np.zeros((7, S), np.float32)
    sp[0] = rng.uniform(0.5, 9.0, S); sp[1] = rng.uniform(-3.5, 3.5, S); sp[3] = rng.uniform(0.03, 0.25, S)
    sp[2] = sp[3] + rng.uniform(0.0, 1.5, S); sp[4:7] = rng.integers(40, 256, (3, S))
    li = np.zeros((3, L), np.float32); li[0] = rng.uniform(-2, 8, L); li[1] = rng.uniform(-4, 4, L); li[2] = rng.uniform(3, 7, L)
    pl = np.array([[5, 0, 0, 0, 0, 1, 128, 128, 128]], np.float32).T.copy()
    cam = Camera(resolution=(W, H), position=[-2, 0, 2.0], euler=[0, -30, 0])
    return cam, (cam.position, cam.rotation, sp, li, pl, 0.05, 0.6, 0.3)


tot = 0
# C1-like: default scene, AA (ordered compaction of the primary queue), depth 2
scene = Scene.default_scene(); cam = Camera(resolution=(96, 48), position=[-2, 0, 2.0], euler=[0, -30, 0])
for _ in range(3):                                      # third call replays the captured graph
    tot += int(render_frame(scene, cam, 0.0, 0.6, 0.3, 2, True).sum())
# C4-like: 1024 spheres: beam culling + look-back compaction + light grids + regrouping (big enough for the sort) / grid walk (small frames)
cam, a = synthetic(1024, 3, 256, 128, 1)
for flags in (0, _lib.F_NO_CGRID, _lib.F_NO_SGRID, _lib.F_DENSE):
    tot += int(render_host(None, *a, 3, False, grid=cam.grid(), shape=(256, 128), flags=flags).sum())
cam, a = synthetic(300, 9, 128, 64, 2)                   # > 7 lights: two compaction rounds (scan path), AA
tot += int(render_host(None, *a, 2, True, grid=cam.grid(), shape=(128, 64), flags=_lib.F_NO_SGRID).sum())
tot += int(render_host(None, *a, 2, True, grid=cam.grid(), shape=(128, 64)).sum())
cam, a = synthetic(2100, 2, 96, 64, 3)                   # S > 1024: grid walk at every depth
tot += int(render_host(None, *a, 2, False, grid=cam.grid(), shape=(96, 64)).sum())
f = _lib.lib().b200rt_debug_check_failures()
print("sanitize cases done, checksum", tot, "| index-assertion failures:", "not a checked build" if f == 2**64 - 1 else f)
sys.exit(0 if f in (0, 2**64 - 1) else 3)
