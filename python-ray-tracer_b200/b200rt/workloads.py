"""The synthetic workloads of BASELINE.json `configs` as frozen in SURVEY §8(d) (deterministic, no files)."""
import numpy as np
from .scene import Scene, Camera, Sphere, Light, Plane
from .scene.colors import GREY

SHADER = dict(amb=0.0, lamb=0.6, refl=0.3)               # reference src/main.py:11


def main_camera(res):
    return Camera(resolution=res, position=[-2, 0, 2.0], euler=[0, -30, 0])      # src/main.py:24


def c4_scene(n=1024, seed=1024):
    rng = np.random.default_rng(seed)
    r = rng.uniform(0.10, 0.45, n)
    cx = rng.uniform(-2, 22, n)
    cy = rng.uniform(-9, 9, n)
    cz = r + rng.uniform(0, 2.5, n)
    col = rng.integers(40, 256, (n, 3))
    d = Scene.default_scene()
    return Scene(d.lights, [Sphere([cx[i], cy[i], cz[i]], r[i], col[i]) for i in range(n)], [Plane([5, 0, 0], [0, 0, 1], GREY)])


def c3_scene():
    d = Scene.default_scene()
    lights = [Light([4 * np.cos(2 * np.pi * k / 8), 4 * np.sin(2 * np.pi * k / 8), 5.0]) for k in range(8)]
    return Scene(lights, d.spheres, d.planes)


def c5_camera(k, res=(3840, 2160)):
    th = 2 * np.pi * k / 64
    return Camera(res, [0.2 + 6 * np.cos(th), 6 * np.sin(th), 3.0],
                  [0, -np.degrees(np.arctan2(3.0 - 0.5, 6)), np.degrees(th) + 180])


def workload(name):
    """-> dict(scene, cameras [list], amb, lamb, refl, refl_depth, aliasing, label)"""
    name = name.lower()
    if name == "c1":
        return dict(scene=Scene.default_scene(), cameras=[main_camera((320, 180))], **SHADER, refl_depth=2, aliasing=True,
                    label="C1 default scene 320x180 depth 2 AA")
    if name == "c2":
        return dict(scene=Scene.default_scene(), cameras=[main_camera((1920, 1080))], **SHADER, refl_depth=1, aliasing=False,
                    label="C2 default scene 1920x1080 depth 1")
    if name == "c3":
        return dict(scene=c3_scene(), cameras=[main_camera((3840, 2160))], amb=0.0, lamb=0.225, refl=0.3, refl_depth=3,
                    aliasing=False, label="C3 default spheres + 8 lights 3840x2160 depth 3")
    if name == "c4":
        return dict(scene=c4_scene(), cameras=[Camera((7680, 4320), [-6, 0, 5.0], [0, -25, 0])], **SHADER, refl_depth=3,
                    aliasing=False, label="C4 1024 random spheres 7680x4320 depth 3")
    if name == "c5":
        return dict(scene=Scene.default_scene(), cameras=[c5_camera(k) for k in range(64)], **SHADER, refl_depth=2,
                    aliasing=False, label="C5 64-frame orbit of the default scene 3840x2160 depth 2")
    if name in ("c4_s2048", "c4_s4096"):          # not a BASELINE config: the C4 generator with more spheres, for the S > 1024 path (grid walk at every depth)
        n = int(name.split("_s")[1])
        return dict(scene=c4_scene(n=n, seed=n), cameras=[Camera((3840, 2160), [-6, 0, 5.0], [0, -25, 0])], **SHADER, refl_depth=3,
                    aliasing=False, label=f"C4-like: {n} random spheres 3840x2160 depth 3")
    raise ValueError(name)


def flops_per_ray(S, Pl):
    """Algorithmic FLOPs of one get_intersection query, canonical brute-force form: 16*S + 14*Pl."""
    return 16 * S + 14 * Pl
