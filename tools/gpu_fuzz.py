"""Fuzz the exact-culling paths: random scenes (lights among the spheres, inside them, far away; clustered and huge spheres),
default path vs brute force (B200RT_F_DENSE), the scan path (B200RT_F_NO_SGRID), the path without the closest-hit grid / the fused
small-scene kernels (B200RT_F_NO_CGRID | B200RT_F_NO_FUSED), a repeated (cached, graph-replayed) call and - for S <= 700 - the exact
test on every sphere without any filter (B200RT_F_NO_FILTER) must be byte-identical.  usage: python tools/gpu_fuzz.py [n_scenes] [first_seed]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "python-ray-tracer_b200")); sys.path.insert(0, ROOT)
from b200rt import render_host, _lib
from b200rt.scene import Camera

def scene(seed):
    """Random scene `seed` -> (args of render_host after pixel_loc, grid, (W, H), description)."""
    rng = np.random.default_rng(1000 + seed)
    S = int(rng.choice([1, 3, 17, 63, 64, 200, 700, 1500, 2600])); L = int(rng.integers(1, 6)); Pl = int(rng.integers(0, 3))
    W, H = 16 * int(rng.integers(6, 14)), 16 * int(rng.integers(4, 9))
    sp = np.zeros((7, S), np.float32)
    spread = float(rng.choice([1.0, 4.0, 12.0]))
    sp[0] = rng.uniform(0.5, 0.5 + 2 * spread, S); sp[1] = rng.uniform(-spread, spread, S)
    sp[3] = rng.uniform(0.02, float(rng.choice([0.1, 0.4, 1.5])), S); sp[2] = sp[3] * rng.uniform(0.2, 1.0, S) + rng.uniform(0, 2.0, S)
    sp[4:7] = rng.integers(40, 256, (3, S))
    li = np.zeros((3, L), np.float32)
    for j in range(L):
        mode = rng.integers(0, 5)
        if mode == 0:   li[:, j] = sp[0:3, rng.integers(0, S)] + rng.normal(0, 0.05, 3)          # inside / at a sphere
        elif mode == 1: li[:, j] = [rng.uniform(-2, 2 * spread), rng.uniform(-spread, spread), rng.uniform(0.05, 1.0)]   # among the spheres
        elif mode == 2: li[:, j] = [rng.uniform(-300, 300), rng.uniform(-300, 300), rng.uniform(1, 400)]              # far (tier 2 / out of range)
        else:           li[:, j] = [rng.uniform(-2, 8), rng.uniform(-4, 4), rng.uniform(2, 7)]
    pl = np.zeros((9, Pl), np.float32)
    for p in range(Pl):
        nrm = np.array([0, 0, 1.0]) if p == 0 else rng.normal(0, 1, 3); nrm /= np.linalg.norm(nrm)
        pl[:, p] = [5, 0, -0.1 * p - (0 if p == 0 else 3), *nrm, 128, 160, 128]
    cam = Camera(resolution=(W, H), position=[-2, 0, 2.0], euler=[0, -30, 0])
    depth = int(rng.integers(0, 4)); aa = bool(rng.integers(0, 4) == 0)
    args = (cam.position, cam.rotation, sp, li, pl, 0.05, 0.6, 0.3, depth, aa)
    return args, np.array(cam.grid()), (W, H), f"S={S} L={L} Pl={Pl} {W}x{H} depth={depth} aa={aa}"


def check(seed):
    args, grid, shape, desc = scene(seed)
    a = render_host(None, *args, grid=grid, shape=shape)
    b = render_host(None, *args, grid=grid, shape=shape, flags=_lib.F_DENSE)
    c = render_host(None, *args, grid=grid, shape=shape, flags=_lib.F_NO_SGRID)
    # a, b and c all run the conservative FFMA2 filter; B200RT_F_NO_FILTER runs sphere_exact on every sphere (scalar debug
    # kernel, slow: S <= 700 only) and is the check that the filter's margins never drop a hit
    d = render_host(None, *args, grid=grid, shape=shape, flags=_lib.F_NO_FILTER) if args[2].shape[1] <= 700 else a
    # the structures of round 2 one by one: the per-ray grid walk off (beam-culled scan instead), the fused small-scene kernels off (wavefront instead),
    # and the cached / graph-replayed second call of the same arguments
    e = render_host(None, *args, grid=grid, shape=shape, flags=_lib.F_NO_CGRID | _lib.F_NO_FUSED)
    f = render_host(None, *args, grid=grid, shape=shape)
    return (np.array_equal(a, b) and np.array_equal(a, c) and np.array_equal(a, d) and np.array_equal(a, e) and np.array_equal(a, f), desc, int(a.sum()), int((a != b).sum()),
            int((a != c).sum()) + int((a != d).sum()) + int((a != e).sum()) + int((a != f).sum()))


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    bad = 0
    for seed in range(first, first + n):
        ok, desc, chk, d1, d2 = check(seed)
        bad += 0 if ok else 1
        print(f"seed {seed}: {desc} sum={chk} -> {'ok' if ok else f'MISMATCH {d1} / {d2}'}", flush=True)
    print("mismatching scenes:", bad)
    sys.exit(1 if bad else 0)
