"""Host-side API mirror (b200rt.scene) against the reference's values captured in the goldens and, when the
reference tree is present (build container only), against the reference package itself."""
import importlib.util
import os
import sys
import numpy as np
import pytest
from conftest import load_golden, pixel_loc_of

from b200rt.scene import Scene, Camera, Sphere, Light, Plane, euler_rotation
from b200rt.scene import colors


def test_default_scene_matches_reference_packing():
    g = load_golden("c1_aa0")
    sp, li, pl = Scene.default_scene().generate_scene()
    for got, name in ((sp, "spheres"), (li, "lights"), (pl, "planes")):
        assert got.dtype == np.float32 and got.flags.c_contiguous
        assert np.array_equal(got, g[name]), name
    assert sp.shape == (7, 6) and li.shape == (3, 3) and pl.shape == (9, 1)


def test_camera_matches_reference():
    g = load_golden("c1_aa0")
    cam = Camera(resolution=(320, 180), position=[-2, 0, 2.0], euler=[0, -30, 0])
    assert cam.position.dtype == np.float64 and np.array_equal(cam.position, g["cam_origin"])
    assert np.array_equal(cam.rotation, g["cam_rotation"])
    assert cam.field_of_view == 45. and cam.resolution == (320, 180)
    P = cam.generate_pixel_locations()
    assert P.shape == (3, 320, 180) and P.dtype == np.float64
    assert np.array_equal(np.asarray(P), pixel_loc_of(g))
    assert np.array_equal(np.array(cam.grid()), g["grid"])
    # the grid-tagged array is read-only; any derived array is a plain writable ndarray without the tag
    assert P.grid is not None and not P.flags.writeable
    Q = P * 1.0
    assert getattr(Q, "grid", None) is None
    with pytest.raises(ValueError):
        P[0, 0, 0] = 1.0


@pytest.mark.parametrize("res,fov", [((1920, 1080), 45.), ((128, 32), 70.), ((64, 48), 45.), ((96, 54), 60.)])
def test_pixel_grid_aspect_ratio_is_integer_division(res, fov):
    """camera.py:22: AR = int(W/H) (16:9 -> 1, 4:1 -> 4)."""
    cam = Camera(res, [0, 0, 1.0], [0, 0, 0], fov=fov)
    P = np.asarray(cam.generate_pixel_locations())
    ar = int(res[0] / res[1])
    assert P[1, 0, 0] == ar and P[1, -1, 0] == pytest.approx(-ar, abs=1e-12)
    assert P[2, 0, 0] == 1 and P[2, 0, -1] == pytest.approx(-1, abs=1e-12)
    assert np.all(P[0] == 1 / np.tan(np.radians(fov) / 2))


def test_dataclasses_and_packing_rules():
    s = Sphere([1, 2, 3], 0.5, [10, 20, 30])
    assert Sphere.data_length == 7 and Light.data_length == 3 and Plane.data_length == 9
    assert np.array_equal(s.to_array(), np.array([1, 2, 3, 0.5, 10, 20, 30], dtype=np.float32))
    p = Plane([0, 0, 0], [0, 3, 4], [1, 2, 3]).to_array()
    assert np.allclose(p[3:6], [0, 0.6, 0.8]) and p.dtype == np.float32     # normalised in float64, stored float32
    assert np.array_equal(Light(np.array([1.5, 2, 3])).to_array(), np.array([1.5, 2, 3], dtype=np.float32))
    sc = Scene([], [], [])
    sp, li, pl = sc.generate_scene()
    assert sp.shape == (7, 0) and li.shape == (3, 0) and pl.shape == (9, 0)
    assert colors.RED == [255, 70, 70] and colors.GREY == [125, 125, 125] and colors.MAGENTA == [139, 0, 139]


def test_euler_rotation_conventions():
    R = euler_rotation(0, -30, 0)
    c, s = np.cos(np.deg2rad(-30)), np.sin(np.deg2rad(-30))
    assert np.allclose(R, [[c, 0, -s], [0, 1, 0], [s, 0, c]])             # rotation.py:18-20 (transposed Ry)
    assert np.allclose(euler_rotation(0.1, 0.2, 0.3, is_radians=True), euler_rotation(*np.rad2deg([0.1, 0.2, 0.3])))


REF = "/root/reference/src"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_against_reference_package_in_place():
    spec = importlib.util.spec_from_file_location("ref_scene", os.path.join(REF, "scene", "__init__.py"),
                                                  submodule_search_locations=[os.path.join(REF, "scene")])
    ref = importlib.util.module_from_spec(spec); sys.modules["ref_scene"] = ref; spec.loader.exec_module(ref)
    rng = np.random.default_rng(0)
    for _ in range(5):
        e = rng.uniform(-90, 90, 3); pos = rng.uniform(-5, 5, 3); res = (int(rng.integers(8, 300)), int(rng.integers(8, 200)))
        fov = float(rng.uniform(20, 100))
        a, b = ref.Camera(res, list(pos), list(e), fov), Camera(res, list(pos), list(e), fov)
        assert np.array_equal(a.rotation, b.rotation) and np.array_equal(a.position, b.position)
        assert np.array_equal(a.generate_pixel_locations(), np.asarray(b.generate_pixel_locations()))
    for res in ((1, 1), (5, 1), (1, 5), (2, 1)):                    # one-point axes: mgrid returns `start`, no division by W-1 = 0
        a, b = ref.Camera(res, [0, 0, 1.0], [0, 0, 0]), Camera(res, [0, 0, 1.0], [0, 0, 0])
        Pa, Pb = a.generate_pixel_locations(), np.asarray(b.generate_pixel_locations())
        assert np.all(np.isfinite(Pb)) and np.array_equal(Pa, Pb), res
        assert np.all(np.isfinite(b.grid()))
    sa = ref.Scene([ref.Light([1, 2, 3])], [ref.Sphere([0.1, 0.2, 0.3], 0.7, [1, 2, 3])], [ref.Plane([0, 0, 1], [1, 2, 3], [9, 8, 7])])
    sb = Scene([Light([1, 2, 3])], [Sphere([0.1, 0.2, 0.3], 0.7, [1, 2, 3])], [Plane([0, 0, 1], [1, 2, 3], [9, 8, 7])])
    for x, y in zip(sa.generate_scene(), sb.generate_scene()):
        assert np.array_equal(x, y) and x.dtype == y.dtype


def test_from_arrays_packs_the_same_bytes_as_the_object_path():
    """vectorised packing == per-object to_array() packing, byte for byte."""
    from b200rt.scene import Scene, Sphere, Light, Plane
    rng = np.random.default_rng(11)
    n = 500
    c = rng.uniform(-20, 20, (n, 3)); r = rng.uniform(0.05, 2, n); col = rng.integers(0, 256, (n, 3))
    li = rng.uniform(-5, 5, (4, 3))
    po = rng.uniform(-5, 5, (3, 3)); pn = rng.normal(size=(3, 3)); pc = rng.integers(0, 256, (3, 3))
    a = Scene([Light(list(v)) for v in li], [Sphere(list(c[i]), r[i], list(col[i])) for i in range(n)],
              [Plane(list(po[i]), list(pn[i]), list(pc[i])) for i in range(3)])
    b = Scene.from_arrays(c, r, col, li, po, pn, pc)
    for x, y in zip(a.generate_scene(), b.generate_scene()):
        assert x.dtype == y.dtype == np.float32 and x.shape == y.shape
        assert x.tobytes() == y.tobytes()
    e = Scene.from_arrays()
    assert [v.shape for v in e.generate_scene()] == [(7, 0), (3, 0), (9, 0)]
