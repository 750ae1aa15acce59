"""tests/golden/make_golden.py — generate the committed golden vectors by RUNNING THE REFERENCE.

Run in the build container only (needs /root/reference and numba):

    python tests/golden/make_golden.py [--only NAME ...] [--procs 8]

Every .npz written here holds the exact inputs (scene arrays, camera, scalars) next to the
reference's outputs, so the tests on the GPU box need neither numba nor the reference tree.
The reference is imported in place (oracle/sim_harness.py) and executed unmodified under
NUMBA_ENABLE_CUDASIM=1; scene/camera inputs come from the reference's own `scene` package.
"""
import argparse
import os
import sys
import time
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import sim_harness as SH  # noqa: E402  (sets NUMBA_ENABLE_CUDASIM=1)

RS = SH.reference_scene_module()
Scene, Camera, Sphere, Light, Plane = RS.Scene, RS.Camera, RS.Sphere, RS.Light, RS.Plane
GREY = [125, 125, 125]


def grid_of(cam):
    """(f, ar, sy, sz) closed form of Camera.generate_pixel_locations (reference camera.py:18-26)."""
    W, H = cam.resolution
    ar = int(W / H)
    f = 1 / np.tan(np.radians(cam.field_of_view) / 2)
    return np.array([f, float(ar), (-ar - ar) / float(W - 1), (-1 - 1) / float(H - 1)], dtype=np.float64)


def frame_case(name, scene, cam, amb, lamb, refl, depth, aa, regions=None, procs=None, cols_per_job=4):
    """Render whole frame (regions=None) or a list of (x0,x1,y0,y1) crops through the simulator."""
    sp, li, pl = scene.generate_scene()
    P = cam.generate_pixel_locations()
    W, H = cam.resolution
    t0 = time.time()
    if regions is None:
        regions = [(0, W, 0, H)]
    outs = []
    for (x0, x1, y0, y1) in regions:
        outs.append(SH.render_reference_sim(P, cam.position, cam.rotation, sp, li, pl, amb, lamb, refl, depth, aa,
                                            x_range=(x0, x1), y_range=(y0, y1), procs=procs, cols_per_job=cols_per_job))
    d = dict(W=W, H=H, spheres=sp, lights=li, planes=pl,
             cam_origin=np.asarray(cam.position, dtype=np.float64), cam_rotation=np.asarray(cam.rotation, dtype=np.float64),
             grid=grid_of(cam), amb=float(amb), lamb=float(lamb), refl=float(refl), refl_depth=int(depth), aliasing=bool(aa),
             regions=np.asarray(regions, dtype=np.int64), seconds=time.time() - t0)
    # check that the closed-form grid reproduces the reference's pixel_loc bit for bit
    g = d["grid"]
    assert np.array_equal(P[0], np.full((W, H), g[0]))
    assert np.array_equal(P[1], (np.arange(W) * g[2] + g[1])[:, None] * np.ones((1, H)))
    assert np.array_equal(P[2], np.ones((W, 1)) * (np.arange(H) * g[3] + 1.0)[None, :])
    if len(outs) == 1 and regions[0] == (0, W, 0, H):
        d["result"] = outs[0]
    else:
        for k, o in enumerate(outs):
            d[f"result_{k}"] = o
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
    print(f"[golden] {name}: {time.time() - t0:.1f} s", flush=True)


def random_scene(rng, S, L, Pl, spread=3.0):
    spheres = [Sphere(list(rng.uniform(-spread, spread, 2)) + [float(rng.uniform(0.2, 1.5))],
                      float(rng.uniform(0.2, 0.9)), [int(v) for v in rng.integers(30, 256, 3)]) for _ in range(S)]
    lights = [Light([float(rng.uniform(-4, 4)), float(rng.uniform(-4, 4)), float(rng.uniform(3, 7))]) for _ in range(L)]
    planes = [Plane([0, 0, 0], [0, 0, 1], GREY)]
    if Pl >= 2:
        planes.append(Plane([6.0, 0, 0], [-1.0, 0.2, 0.3], [200, 180, 90]))
    if Pl >= 3:
        planes.append(Plane([0, 5.0, 0], [0.1, -1.0, 0.2], [90, 180, 200]))
    return Scene(lights, spheres, planes[:Pl])


def c4_scene():
    """1024 random spheres, default lights, grey floor."""
    rng = np.random.default_rng(1024)
    r = rng.uniform(0.10, 0.45, 1024)
    cx = rng.uniform(-2, 22, 1024)
    cy = rng.uniform(-9, 9, 1024)
    cz = r + rng.uniform(0, 2.5, 1024)
    col = rng.integers(40, 256, (1024, 3))
    d = Scene.default_scene()
    spheres = [Sphere([cx[i], cy[i], cz[i]], r[i], col[i]) for i in range(1024)]
    return Scene(d.lights, spheres, [Plane([5, 0, 0], [0, 0, 1], GREY)])


def c3_scene():
    d = Scene.default_scene()
    lights = [Light([4 * np.cos(2 * np.pi * k / 8), 4 * np.sin(2 * np.pi * k / 8), 5.0]) for k in range(8)]
    return Scene(lights, d.spheres, d.planes)


def c5_camera(k, res=(3840, 2160)):
    th = 2 * np.pi * k / 64
    return Camera(res, [0.2 + 6 * np.cos(th), 6 * np.sin(th), 3.0],
                  [0, -np.degrees(np.arctan2(3.0 - 0.5, 6)), np.degrees(th) + 180])


def random_regions(rng, W, H, n, size, include_edges=True):
    regs = []
    if include_edges:
        regs.append((0, size, 0, size))                     # touches x=0 / y=0 (un-aliased border)
        regs.append((W - size, W, H - size, H))             # touches x=W-1 / y=H-1 (flat-wrap reads)
    while len(regs) < n:
        x0 = int(rng.integers(0, W - size + 1)); y0 = int(rng.integers(0, H - size + 1))
        regs.append((x0, x0 + size, y0, y0 + size))
    return regs


def function_vectors():
    """Known-answer vectors from the reference's device functions called directly (simulator mode).

    Ray data is handed over as np.float64 scalars (what the kernel path sees: camera arrays are float64),
    scene data as float32 arrays; amb/lamb/refl as python floats, as src/main.py:11 does.
    """
    F = SH.reference_device_functions()
    rng = np.random.default_rng(7)
    out = {}

    def unit(v):
        v = np.asarray(v, dtype=np.float64)
        return v / np.sqrt((v * v).sum())

    def tup(v):
        return tuple(np.float64(x) for x in v)

    # --- intersect_ray_sphere: random + inside-sphere + behind-origin + tangent-ish + non-unit direction
    O, D, Cc, R, T = [], [], [], [], []
    for i in range(256):
        o = rng.uniform(-5, 5, 3); c = rng.uniform(-5, 5, 3).astype(np.float32); r = np.float32(rng.uniform(0.1, 2.0))
        kind = i % 6
        if kind == 0:
            d = unit(rng.normal(size=3))
        elif kind == 1:                                   # aimed at the sphere
            d = unit(c.astype(np.float64) - o + rng.normal(size=3) * 0.3 * r)
        elif kind == 2:                                   # origin inside the sphere
            o = c.astype(np.float64) + unit(rng.normal(size=3)) * float(r) * rng.uniform(0, 0.95); d = unit(rng.normal(size=3))
        elif kind == 3:                                   # sphere behind the origin
            d = -unit(c.astype(np.float64) - o)
        elif kind == 4:                                   # grazing: offset by ~r perpendicular
            to_c = c.astype(np.float64) - o; perp = unit(np.cross(to_c, rng.normal(size=3)))
            d = unit(to_c + perp * float(r) * (1 + rng.normal() * 1e-3))
        else:                                             # direction not normalised (function re-normalises)
            d = rng.normal(size=3) * 3.0
        t = F["intersect_ray_sphere"](tup(o), tup(d), c, r)
        O.append(o); D.append(d); Cc.append(c); R.append(r); T.append(float(t))
    out.update(sph_O=np.array(O), sph_D=np.array(D), sph_C=np.array(Cc, dtype=np.float32), sph_r=np.array(R, dtype=np.float32), sph_t=np.array(T))

    # --- intersect_ray_plane: random + near-parallel around the 1e-3 threshold + behind
    O, D, Po, Pn, T = [], [], [], [], []
    for i in range(256):
        o = rng.uniform(-5, 5, 3); po = rng.uniform(-5, 5, 3).astype(np.float32)
        n = unit(rng.normal(size=3)).astype(np.float32)
        if i % 4 == 1:                                    # |denom| close to EPS
            a = unit(np.cross(n.astype(np.float64), rng.normal(size=3)))
            d = unit(a + n.astype(np.float64) * rng.uniform(-2e-3, 2e-3))
        else:
            d = unit(rng.normal(size=3))
        t = F["intersect_ray_plane"](tup(o), tup(d), po, n)
        O.append(o); D.append(d); Po.append(po); Pn.append(n); T.append(float(t))
    out.update(pln_O=np.array(O), pln_D=np.array(D), pln_o=np.array(Po, dtype=np.float32), pln_n=np.array(Pn, dtype=np.float32), pln_t=np.array(T))

    # --- get_intersection / trace / sample on two scenes
    for tag, scene in (("def", Scene.default_scene()), ("rnd", random_scene(np.random.default_rng(11), 9, 3, 2))):
        sp, li, pl = scene.generate_scene()
        O, D, GI, TR, SA, SAW = [], [], [], [], [], []
        for i in range(96):
            o = np.array([-2, 0, 2.0]) + rng.normal(size=3) * 0.3
            d = unit(np.array([0.8, 0, -0.5]) + rng.normal(size=3) * 0.5)
            t, idx, typ = F["get_intersection"](tup(o), tup(d), sp, pl)
            rgb, p, r = F["trace"](tup(o), tup(d), sp, li, pl, 0.0, 0.6)
            s0 = F["sample"](tup(o), tup(d), sp, li, pl, 0.0, 0.6, 0.3, 3)
            s1 = F["sample"](tup(o), tup(d), sp, li, pl, 0.1, 0.6, 0.3, 2)      # amb != 0: simulator weak-f32 path
            O.append(o); D.append(d); GI.append([float(t), idx, typ])
            TR.append([float(v) for v in rgb] + [float(v) for v in p] + [float(v) for v in r])
            SA.append([float(v) for v in s0]); SAW.append([float(v) for v in s1])
        out.update({f"{tag}_spheres": sp, f"{tag}_lights": li, f"{tag}_planes": pl, f"{tag}_O": np.array(O), f"{tag}_D": np.array(D),
                    f"{tag}_gi": np.array(GI), f"{tag}_trace": np.array(TR), f"{tag}_sample_d3": np.array(SA),
                    f"{tag}_sample_amb01_d2": np.array(SAW)})

    # --- clip_color: ties (half-even), negatives, saturation
    xs = np.array([-3.2, -0.5, -0.0, 0.0, 0.49999, 0.5, 0.50001, 1.5, 2.5, 3.5, 126.5, 127.5, 253.5, 254.5, 254.51, 255.0,
                   255.49, 255.5, 256.0, 1000.7, 64.259224, 99.5, 100.5], dtype=np.float64)
    out.update(clip_in=xs, clip_out=np.array([int(F["clip_color"](np.float64(x))) for x in xs], dtype=np.int64))
    np.savez_compressed(os.path.join(HERE, "functions.npz"), **out)
    print("[golden] functions.npz", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*", default=None)
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    a = ap.parse_args()
    p = a.procs
    default = Scene.default_scene()
    main_cam = lambda res: Camera(resolution=res, position=[-2, 0, 2.0], euler=[0, -30, 0])   # src/main.py:24
    rng = np.random.default_rng(2024)
    cases = {
        "functions": function_vectors,
        # C1: BASELINE.json configs[0] — default scene 320x180 under the simulator, AA off and on (main.py:11-12)
        "c1_aa0": lambda: frame_case("c1_aa0", default, main_cam((320, 180)), 0.0, 0.6, 0.3, 2, False, procs=p),
        "c1_aa1": lambda: frame_case("c1_aa1", default, main_cam((320, 180)), 0.0, 0.6, 0.3, 2, True, procs=p),
        # small random scenes, whole frames
        "rand_a": lambda: frame_case("rand_a", random_scene(np.random.default_rng(1), 12, 2, 2), Camera((96, 54), [-4, 0.5, 2.5], [0, -25, 5], fov=60.),
                                     0.1, 0.5, 0.4, 3, False, procs=p),
        "rand_b": lambda: frame_case("rand_b", random_scene(np.random.default_rng(2), 5, 4, 1), Camera((64, 48), [-3, -1, 3.0], [5, -35, 10]),
                                     0.05, 0.3, 0.5, 1, True, procs=p),
        "rand_c": lambda: frame_case("rand_c", random_scene(np.random.default_rng(3), 20, 3, 3), Camera((128, 32), [-5, 0, 2.0], [0, -15, 0], fov=70.),
                                     0.0, 0.6, 0.3, 4, False, procs=p),
        # degenerate object counts: loops are `range(shape[1])` (trace.py:22,32,86)
        "edge_s0": lambda: frame_case("edge_s0", random_scene(np.random.default_rng(4), 0, 2, 1), Camera((32, 16), [-3, 0, 2.0], [0, -30, 0]),
                                      0.0, 0.6, 0.3, 2, True, procs=p),
        "edge_l0": lambda: frame_case("edge_l0", random_scene(np.random.default_rng(5), 4, 0, 1), Camera((32, 16), [-3, 0, 2.0], [0, -30, 0]),
                                      0.2, 0.6, 0.3, 2, False, procs=p),
        "edge_p0": lambda: frame_case("edge_p0", random_scene(np.random.default_rng(6), 6, 2, 0), Camera((32, 16), [-3, 0, 2.0], [0, -20, 0]),
                                      0.0, 0.6, 0.3, 2, False, procs=p),
        "edge_d0": lambda: frame_case("edge_d0", default, main_cam((48, 27)), 0.0, 0.6, 0.3, 0, True, procs=p),
        # crops of the large configs; whole frames are infeasible under the simulator
        "c2_aa0_crops": lambda: frame_case("c2_aa0_crops", default, main_cam((1920, 1080)), 0.0, 0.6, 0.3, 1, False,
                                           regions=random_regions(rng, 1920, 1080, 8, 32), procs=p),
        "c2_aa1_crops": lambda: frame_case("c2_aa1_crops", default, main_cam((1920, 1080)), 0.0, 0.6, 0.3, 1, True,
                                           regions=random_regions(rng, 1920, 1080, 4, 24), procs=p),
        "c3_crops": lambda: frame_case("c3_crops", c3_scene(), main_cam((3840, 2160)), 0.0, 0.225, 0.3, 3, False,
                                       regions=random_regions(rng, 3840, 2160, 8, 32), procs=p),
        "c4_crops": lambda: frame_case("c4_crops", c4_scene(), Camera((7680, 4320), [-6, 0, 5.0], [0, -25, 0]), 0.0, 0.6, 0.3, 3, False,
                                       regions=random_regions(rng, 7680, 4320, 16, 12), procs=p, cols_per_job=1),
        # round 2: the simulator pin of the full-size C4 frame grown 4x (28 more crops of 16x16 = 7168 px, ~2.4 core-s per pixel
        # at S=1024: ~5 core-hours), and C4 WITH anti-aliasing (9 samples per pixel: ~22 core-s per pixel), incl. the flat-wrap corner
        "c4_crops_b": lambda: frame_case("c4_crops_b", c4_scene(), Camera((7680, 4320), [-6, 0, 5.0], [0, -25, 0]), 0.0, 0.6, 0.3, 3, False,
                                         regions=random_regions(np.random.default_rng(4102), 7680, 4320, 28, 16, include_edges=False), procs=p, cols_per_job=1),
        "c4_aa1_crops": lambda: frame_case("c4_aa1_crops", c4_scene(), Camera((7680, 4320), [-6, 0, 5.0], [0, -25, 0]), 0.0, 0.6, 0.3, 3, True,
                                           regions=random_regions(np.random.default_rng(4103), 7680, 4320, 5, 8), procs=p, cols_per_job=1),
        "c4_small": lambda: frame_case("c4_small", c4_scene(), Camera((64, 36), [-6, 0, 5.0], [0, -25, 0]), 0.0, 0.6, 0.3, 3, False,
                                       procs=p, cols_per_job=1),
        "c5_crops_k0": lambda: frame_case("c5_crops_k0", default, c5_camera(0), 0.0, 0.6, 0.3, 2, False,
                                          regions=random_regions(rng, 3840, 2160, 4, 32), procs=p),
        "c5_crops_k23": lambda: frame_case("c5_crops_k23", default, c5_camera(23), 0.0, 0.6, 0.3, 2, False,
                                           regions=random_regions(rng, 3840, 2160, 4, 32), procs=p),
    }
    names = a.only or list(cases)
    for n in names:
        cases[n]()


if __name__ == "__main__":
    main()


def viewer_golden():
    """tests/golden/viewer_square.npz: the reference's convert_array_to_image (src/viewer/image.py:7-19) on random
    square uint8 arrays.  Run with:  python -c "import make_golden as m; m.viewer_golden()"  (needs /root/reference)."""
    sys.path.insert(0, "/root/reference/src")
    from viewer import convert_array_to_image
    rng = np.random.default_rng(7)
    out = {}
    for n in (96, 130):
        x = rng.integers(0, 256, (3, n, n), dtype=np.uint8)
        out[f"x_{n}"] = x
        out[f"img_{n}"] = np.asarray(convert_array_to_image(x))
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "viewer_square.npz"), **out)
