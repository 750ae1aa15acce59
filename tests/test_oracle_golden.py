"""The float64 oracle (oracle/render_f64.c) against every golden produced by running the reference itself
(tests/golden/make_golden.py, numba CUDA simulator).  Bit-exact at the uint8 level; function vectors to 1e-12."""
import numpy as np
import pytest
from conftest import load_golden, golden_regions, pixel_loc_of, have_golden, FRAME_GOLDENS, CROP_GOLDENS


def _render(oracle, g, x0, x1, y0, y1, use_grid):
    W, H = int(g["W"]), int(g["H"])
    weak = float(g["amb"]) != 0.0     # simulator-only NumPy weak-promotion artefact, see render_f64.c header
    kw = dict(x_range=(x0, x1), y_range=(y0, y1), numpy_weak=weak)
    args = (g["cam_origin"], g["cam_rotation"], g["spheres"], g["lights"], g["planes"],
            float(g["amb"]), float(g["lamb"]), float(g["refl"]), int(g["refl_depth"]), bool(g["aliasing"]))
    if use_grid:
        out = oracle.render(None, *args, grid=g["grid"], shape=(W, H), **kw)
    else:
        out = oracle.render(pixel_loc_of(g), *args, **kw)
    return out[:, x0:x1, y0:y1]


@pytest.mark.parametrize("name", FRAME_GOLDENS)
def test_oracle_matches_reference_frames(oracle, name):
    if not have_golden(name):
        pytest.skip("golden not generated")
    g = load_golden(name)
    for (x0, x1, y0, y1, ref) in golden_regions(g):
        got = _render(oracle, g, x0, x1, y0, y1, use_grid=False)
        assert np.array_equal(got, ref), f"{name}: {np.count_nonzero(got != ref)} bytes differ"
        got2 = _render(oracle, g, x0, x1, y0, y1, use_grid=True)
        assert np.array_equal(got2, ref)


@pytest.mark.parametrize("name", CROP_GOLDENS)
def test_oracle_matches_reference_crops(oracle, name):
    if not have_golden(name):
        pytest.skip("golden not generated")
    g = load_golden(name)
    for (x0, x1, y0, y1, ref) in golden_regions(g):
        got = _render(oracle, g, x0, x1, y0, y1, use_grid=True)   # closed-form grid: no 796 MB array at 8K
        assert np.array_equal(got, ref), f"{name} crop {(x0, x1, y0, y1)}: {np.count_nonzero(got != ref)} bytes differ"


def test_function_vectors(oracle):
    f = load_golden("functions")
    for i in range(len(f["sph_t"])):
        t = oracle.intersect_ray_sphere(f["sph_O"][i], f["sph_D"][i], f["sph_C"][i], f["sph_r"][i])
        assert t == f["sph_t"][i], (i, t, f["sph_t"][i])
    for i in range(len(f["pln_t"])):
        t = oracle.intersect_ray_plane(f["pln_O"][i], f["pln_D"][i], f["pln_o"][i], f["pln_n"][i])
        assert t == f["pln_t"][i], (i, t, f["pln_t"][i])
    for tag in ("def", "rnd"):
        sp, li, pl = f[f"{tag}_spheres"], f[f"{tag}_lights"], f[f"{tag}_planes"]
        for i in range(len(f[f"{tag}_O"])):
            O, D = f[f"{tag}_O"][i], f[f"{tag}_D"][i]
            t, idx, typ = oracle.get_intersection(O, D, sp, pl)
            assert (t, idx, typ) == tuple(f[f"{tag}_gi"][i]), (tag, i)
            rgb, P, R = oracle.trace(O, D, sp, li, pl, 0.0, 0.6, numpy_weak=True)
            assert np.array_equal(np.concatenate([rgb, P, R]), f[f"{tag}_trace"][i]), (tag, i)
            s, _ = oracle.sample(O, D, sp, li, pl, 0.0, 0.6, 0.3, 3, numpy_weak=True)
            assert np.array_equal(s, f[f"{tag}_sample_d3"][i]), (tag, i)
            s, _ = oracle.sample(O, D, sp, li, pl, 0.1, 0.6, 0.3, 2, numpy_weak=True)
            assert np.array_equal(s, f[f"{tag}_sample_amb01_d2"][i]), (tag, i)
            # compiled-kernel semantics (float64 amb): differs from the simulator by float32 rounding of amb*colour only
            s64, _ = oracle.sample(O, D, sp, li, pl, 0.1, 0.6, 0.3, 2, numpy_weak=False)
            assert np.allclose(s64, s, rtol=0, atol=1e-4)
    for x, y in zip(f["clip_in"], f["clip_out"]):
        assert oracle.clip_color(x) == y, (x, y)


def test_ray_counts_match_survey(oracle):
    """SURVEY §3.3 probe: default scene/camera 64x36 no AA -> 6.751 rays/px at depth 1, 8.908 at depth 3."""
    from b200rt.scene import Scene, Camera
    sp, li, pl = Scene.default_scene().generate_scene()
    cam = Camera((64, 36), [-2, 0, 2.0], [0, -30, 0])
    for depth, expect in ((1, 6.751), (3, 8.908)):
        _, cnt = oracle.render(cam.generate_pixel_locations(), cam.position, cam.rotation, sp, li, pl,
                               0.0, 0.6, 0.3, depth, False, want_counts=True)
        assert abs(int(cnt.sum()) / (64 * 36) - expect) < 5e-4
