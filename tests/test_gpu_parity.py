"""Parity tests proper: the CUDA path (through the C ABI) against the reference's goldens and the float64 oracle.

Tolerance (BASELINE.json north_star): max per-channel error <= 1/255 after uint8 quantisation on >= 99.9 % of
pixels; every other pixel is listed (parity_report()['bad']).  The kernel computes in FP32, the reference in
float64, so bit-equality is not expected: a .5 rounding tie costs one level, a hit/miss flip at a silhouette or
shadow edge costs more and is what the 0.1 % budget is for.
"""
import math
import numpy as np
import pytest
from conftest import (load_golden, golden_regions, pixel_loc_of, have_golden, parity_report,
                      FRAME_GOLDENS, CROP_GOLDENS)

pytestmark = pytest.mark.gpu
TOL_FRAC = 1e-3


def _budget(n):
    # 0.1 % of the pixels; frames under 2000 px get an absolute allowance of 2 flipped pixels
    return max(2, math.floor(TOL_FRAC * n))


def _args(g):
    return (g["cam_origin"], g["cam_rotation"], g["spheres"], g["lights"], g["planes"],
            float(g["amb"]), float(g["lamb"]), float(g["refl"]), int(g["refl_depth"]), bool(g["aliasing"]))


def _cuda(g, tile=None, use_grid=True, flags=0, stats=False):
    from b200rt import render_host
    W, H = int(g["W"]), int(g["H"])
    if use_grid:
        return render_host(None, *_args(g), grid=g["grid"], shape=(W, H), tile=tile, flags=flags, stats=stats)
    return render_host(pixel_loc_of(g), *_args(g), tile=tile, flags=flags, stats=stats)


@pytest.mark.parametrize("name", FRAME_GOLDENS)
def test_frames_vs_reference_golden(name):
    if not have_golden(name):
        pytest.skip("golden not generated")
    g = load_golden(name)
    ref = golden_regions(g)[0][4]
    got = _cuda(g)
    rep = parity_report(got, ref)
    print(name, rep)
    assert rep["n_bad"] <= _budget(rep["n"]), rep


@pytest.mark.parametrize("name", CROP_GOLDENS)
def test_crops_vs_reference_golden(name):
    if not have_golden(name):
        pytest.skip("golden not generated")
    g = load_golden(name)
    tot = dict(n=0, n_bad=0, max=0)
    for (x0, x1, y0, y1, ref) in golden_regions(g):
        got = _cuda(g, tile=(x0, x1))[:, x0:x1, y0:y1]
        rep = parity_report(got, ref)
        tot["n"] += rep["n"]; tot["n_bad"] += rep["n_bad"]; tot["max"] = max(tot["max"], rep["max"])
    print(name, tot)
    assert tot["n_bad"] <= _budget(tot["n"]), tot


@pytest.mark.parametrize("name", ["c1_aa0", "rand_a", "rand_b", "c4_small", "edge_s0"])
def test_pixel_loc_array_path_equals_grid_path(name):
    g = load_golden(name)
    assert np.array_equal(_cuda(g, use_grid=True), _cuda(g, use_grid=False))


@pytest.mark.parametrize("name", ["c1_aa1", "rand_c", "c4_small"])
def test_filter_is_exact(name):
    """The 7-FFMA conservative filter only SELECTS spheres for the exact test: disabling it changes no byte."""
    if not have_golden(name):
        pytest.skip("golden not generated")
    from b200rt import _lib
    g = load_golden(name)
    assert np.array_equal(_cuda(g), _cuda(g, flags=_lib.F_NO_FILTER))


@pytest.mark.parametrize("name", ["c1_aa1", "rand_a", "rand_c", "c4_small", "edge_l0", "edge_p0", "edge_d0"])
def test_scan_path_equals_grid_path(name):
    """Shadow queries by k_isect3<shadow> + k_shade (B200RT_F_NO_SGRID, the path of rounds before the light grids) and
    brute force (B200RT_F_DENSE) write the same bytes as the default path on the reference's golden scenes."""
    from b200rt import _lib
    if not have_golden(name):
        pytest.skip("golden missing")
    g = load_golden(name)
    got = _cuda(g)
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_SGRID))
    assert np.array_equal(got, _cuda(g, flags=_lib.F_DENSE))


def test_deterministic_and_tiled_equals_whole():
    g = load_golden("c4_small")
    a, b = _cuda(g), _cuda(g)
    assert np.array_equal(a, b)
    W = int(g["W"])
    out = np.zeros_like(a)
    from b200rt import render_host
    for x0 in range(0, W, 7):                                  # ragged column tiles, AA-free frame
        render_host(None, *_args(g), grid=g["grid"], shape=(W, int(g["H"])), tile=(x0, min(W, x0 + 7)), out=out)
    assert np.array_equal(out, a)
    g = load_golden("rand_b")                                  # anti-aliased frame: tile-edge neighbours are index arithmetic
    a = _cuda(g); W = int(g["W"]); out = np.zeros_like(a)
    for x0 in range(0, W, 5):
        render_host(None, *_args(g), grid=g["grid"], shape=(W, int(g["H"])), tile=(x0, min(W, x0 + 5)), out=out)
    assert np.array_equal(out, a)


def test_ray_counts_match_oracle(oracle):
    for name in ("c1_aa0", "c4_small", "rand_b"):
        g = load_golden(name)
        _, st = _cuda(g, stats=True)
        _, cnt = oracle.render(pixel_loc_of(g), *_args(g), want_counts=True)
        assert st["primary"] == int(cnt[0])
        assert abs(st["shadow"] - int(cnt[1])) <= max(3, 1e-3 * int(cnt[1])), (st, cnt)
        assert abs(st["bounce"] - int(cnt[2])) <= max(3, 1e-3 * int(cnt[2])), (st, cnt)


def test_numba_style_entry_point_matches_host_call():
    """render[grid, block](12 args) with cuda.to_device arrays == the one-call host path (src/main.py:19-51 sequence)."""
    import torch
    from b200rt import cuda
    from b200rt.ray_tracing import render
    from b200rt.scene import Scene, Camera
    scene = Scene.default_scene()
    sp_h, li_h, pl_h = scene.generate_scene()
    spheres, lights, planes = cuda.to_device(sp_h), cuda.to_device(li_h), cuda.to_device(pl_h)
    w, h = 200, 120
    camera = Camera(resolution=(w, h), position=[-2, 0, 2.0], euler=[0, -30, 0])
    camera_origin, camera_rotation = cuda.to_device(camera.position), cuda.to_device(camera.rotation)
    pixel_loc = cuda.to_device(camera.generate_pixel_locations())
    result = cuda.to_device(np.zeros((3, w, h), dtype=np.uint8))
    tpb = (16, 16)
    bpg = (int(np.ceil(w / 16)), int(np.ceil(h / 16)))
    render[bpg, tpb](pixel_loc, result, camera_origin, camera_rotation, spheres, lights, planes, 0.0, 0.6, 0.3, 2, True)
    out = result.copy_to_host()
    from b200rt import render_frame
    assert np.array_equal(out, render_frame(scene, camera, 0.0, 0.6, 0.3, 2, True))
    # plain (writable, untagged) ndarray pixel grid staged explicitly, torch tensors and host result array
    P = np.array(camera.generate_pixel_locations())
    res2 = np.zeros((3, w, h), dtype=np.uint8)
    render[bpg, tpb](torch.from_numpy(P).cuda(), res2, camera.position, camera.rotation, sp_h, li_h, pl_h, 0.0, 0.6, 0.3, 2, True)
    assert np.array_equal(res2, out)
    with pytest.raises(ValueError):
        render[bpg, tpb](pixel_loc, cuda.to_device(np.zeros((3, w, h + 1), dtype=np.uint8)), camera_origin, camera_rotation,
                         spheres, lights, planes, 0.0, 0.6, 0.3, 2, True)


def test_full_size_c4_crops_vs_oracle(oracle):
    """BASELINE config 4 at its full 7680x4320 size: column tiles rendered on the GPU, 64x64 windows compared
    with the float64 oracle (itself pinned to the simulator on crops of this very frame: c4_crops golden)."""
    g = load_golden("c4_crops")
    rng = np.random.default_rng(5)
    W, H = int(g["W"]), int(g["H"])
    tot = dict(n=0, n_bad=0, max=0, bad=[])
    for _ in range(6):
        x0 = int(rng.integers(0, W - 64)); y0 = int(rng.integers(0, H - 64))
        got = _cuda(g, tile=(x0, x0 + 64))[:, x0:x0 + 64, y0:y0 + 64]
        ref = oracle.render(None, *_args(g), grid=g["grid"], shape=(W, H), x_range=(x0, x0 + 64), y_range=(y0, y0 + 64))[:, x0:x0 + 64, y0:y0 + 64]
        rep = parity_report(got, ref)
        tot["n"] += rep["n"]; tot["n_bad"] += rep["n_bad"]; tot["max"] = max(tot["max"], rep["max"]); tot["bad"] += rep["bad"][:5]
    print("c4 full-size crops", tot)
    assert tot["n_bad"] <= _budget(tot["n"]), tot


def _synthetic(S, L, Pl, W, H, depth, aa, seed):
    """Random scene in front of the main.py camera, packed (7,S)/(3,L)/(9,Pl) float32 like Scene.generate_scene()."""
    from b200rt.scene import Camera
    rng = np.random.default_rng(seed)
    sp = np.zeros((7, S), np.float32)
    if S:
        sp[0] = rng.uniform(0.5, 9.0, S); sp[1] = rng.uniform(-3.5, 3.5, S); sp[3] = rng.uniform(0.03, 0.25, S)
        sp[2] = sp[3] + rng.uniform(0.0, 1.5, S); sp[4:7] = rng.integers(40, 256, (3, S))
    li = np.zeros((3, L), np.float32)
    if L:
        li[0] = rng.uniform(-2, 8, L); li[1] = rng.uniform(-4, 4, L); li[2] = rng.uniform(3, 7, L)
    pl = np.zeros((9, Pl), np.float32)
    for p in range(Pl):
        pl[:, p] = [5, 0, -0.1 * p, 0, 0, 1, 128, 128, 128]
    cam = Camera(resolution=(W, H), position=[-2, 0, 2.0], euler=[0, -30, 0])
    return dict(W=W, H=H, grid=np.array(cam.grid()), cam_origin=cam.position, cam_rotation=cam.rotation, spheres=sp, lights=li,
                planes=pl, amb=0.05, lamb=0.6, refl=0.3, refl_depth=depth, aliasing=aa)


@pytest.mark.parametrize("S,L,Pl", [(4001, 2, 1),      # more spheres than the constant-memory table: two sphere chunks per query
                                    (2000, 5, 1),      # light tables do not fit side by side: one upload per light
                                    (13, 17, 2),       # many lights, odd sphere count (padding pair half empty)
                                    (1, 1, 0), (0, 2, 1), (7, 0, 1)])
def test_launch_plans_vs_oracle_and_unfiltered(oracle, S, L, Pl):
    from b200rt import _lib
    g = _synthetic(S, L, Pl, 96, 64, 2, False, seed=S * 31 + L)
    got = _cuda(g)
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_FILTER)), "filter/pivot path differs from exact-test-on-everything"
    ref = oracle.render(None, *_args(g), grid=g["grid"], shape=(96, 64))
    rep = parity_report(got, ref)
    print((S, L, Pl), rep)
    assert rep["n_bad"] <= _budget(rep["n"]), rep
    _, st = _cuda(g, stats=True)
    _, cnt = oracle.render(None, *_args(g), grid=g["grid"], shape=(96, 64), want_counts=True)
    assert st["primary"] == int(cnt[0])
    assert abs(st["shadow"] - int(cnt[1])) <= max(3, 2e-3 * int(cnt[1])) and abs(st["bounce"] - int(cnt[2])) <= max(3, 2e-3 * int(cnt[2]))


def test_small_queue_variants_bit_identical():
    """Rays-per-thread variants of k_isect3 (chosen by queue size, forced here through the tile width) agree."""
    g = load_golden("c4_small")
    whole = _cuda(g)
    W = int(g["W"])
    out = np.zeros_like(whole)
    from b200rt import render_host
    render_host(None, *_args(g), grid=g["grid"], shape=(W, int(g["H"])), tile=(0, 3), out=out)     # tiny queue -> NR=2
    assert np.array_equal(out[:, :3], whole[:, :3])


def test_batch_equals_single_frames():
    """b200rt_render_batch_host (scene staged once, double-buffered D2H) == frame-by-frame renders."""
    from b200rt import render_batch, render_frame
    from b200rt.scene import Scene
    from b200rt.workloads import c5_camera
    scene = Scene.default_scene()
    cams = [c5_camera(k, res=(192, 108)) for k in (0, 7, 23, 40, 63)]
    got = render_batch(scene, cams, 0.0, 0.6, 0.3, 2, False)
    assert got.shape == (5, 3, 192, 108)
    for k, cam in enumerate(cams):
        assert np.array_equal(got[k], render_frame(scene, cam, 0.0, 0.6, 0.3, 2, False)), k
    two = render_batch(scene, cams, 0.0, 0.6, 0.3, 2, False, devices=[0, 0])      # two host threads, one device
    assert np.array_equal(two, got)


@pytest.mark.parametrize("S,L,W,H,depth,aa", [(1024, 3, 512, 256, 3, False),    # C4-like: shadow passes, beam culling, look-back over 512 chunks
                                              (700, 2, 272, 144, 2, True),      # anti-aliased: ordered compaction of the primary queue (border samples skipped)
                                              (600, 9, 208, 112, 1, False)])    # more than 7 lights: two compaction rounds per chunk
def test_culling_and_passes_do_not_change_the_image(oracle, S, L, W, H, depth, aa):
    """Work the kernel proves unnecessary and skips (shadow passes over the sorted light tables, beam culling, the
    regions of the survivor lists) must leave every byte as the exact-test-on-every-sphere path writes it, at a size
    where queues span hundreds of 256-ray chunks (ordered compaction with look-back, several regions per list)."""
    from b200rt import _lib
    g = _synthetic(S, L, 1, W, H, depth, aa, seed=S + 7 * L)
    got, st = _cuda(g, stats=True)
    nocull, st_nc = _cuda(g, flags=_lib.F_NO_CULL, stats=True)
    assert np.array_equal(got, nocull), "beam culling changed the image"
    nosort, st_ns = _cuda(g, flags=_lib.F_NO_SORT, stats=True)
    assert np.array_equal(got, nosort), "regrouping the bounce-depth hits changed the image"
    assert st["rays"] == st_ns["rays"] and st["shadow_traced"] == st_ns["shadow_traced"]
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_FILTER)), "filtered path differs from exact-test-on-everything"
    assert st["rays"] == st_nc["rays"]
    nogrid, st_ng = _cuda(g, flags=_lib.F_NO_SGRID, stats=True)       # shadow queries by k_isect3 (passes + beam culling) instead of the light grids
    assert np.array_equal(got, nogrid), "the lights' direction grids changed the image"
    assert st_ng["rays"] == st["rays"] and st_ng["shadow_traced"] == st["shadow_traced"]
    dense, st_d = _cuda(g, flags=_lib.F_DENSE, stats=True)            # brute force: the mode bench.py's FP32-peak figure is measured in
    assert np.array_equal(got, dense), "dense (brute-force) mode differs from the production path"
    assert st_d["rays"] == st["rays"] and st_d["shadow_traced"] == st_d["shadow"]
    assert st_d["tests_pivot"] + st_d["tests_general"] >= st["rays"] * S       # every specified test is executed
    canonical = st["rays"] * S
    executed, executed_nc = st["tests_pivot"] + st["tests_general"], st_nc["tests_pivot"] + st_nc["tests_general"]
    print(dict(S=S, L=L, canonical=canonical, executed=executed, executed_no_cull=executed_nc))
    assert executed < executed_nc <= 1.1 * canonical + 8 * 256 * S        # (padding and partially filled groups count as executed)
    x0, y0 = W // 3, H // 4
    ref = oracle.render(None, *_args(g), grid=g["grid"], shape=(W, H), x_range=(x0, x0 + 64), y_range=(y0, y0 + 64))[:, x0:x0 + 64, y0:y0 + 64]
    rep = parity_report(got[:, x0:x0 + 64, y0:y0 + 64], ref)
    assert rep["n_bad"] <= _budget(rep["n"]), rep


@pytest.mark.parametrize("case", ["light_inside_sphere", "light_touching_sphere", "cluster_in_one_cell", "far_light", "light_below_plane", "grazing"])
def test_shadow_grid_adversarial_scenes(case):
    """The per-light direction grids (csrc/shadow_grid.cuh) must never drop a sphere a shadow ray can hit: placements that
    stress the cone bounds (light inside / on a sphere, hundreds of spheres behind one cell -> list overflow, light outside
    the grid's coordinate range -> grid disabled, shadow rays leaving from far-away plane points, grazing directions)
    rendered with the grids and brute force (B200RT_F_DENSE) must give the same bytes."""
    from b200rt import _lib
    S, L, W, H = 300, 3, 256, 144
    g = _synthetic(S, L, 1, W, H, 2, False, seed=11)
    sp, li = g["spheres"], g["lights"]
    rng = np.random.default_rng(3)
    if case == "light_inside_sphere":
        sp[0:3, 0] = li[:, 0]; sp[3, 0] = 0.8                          # light 0 at the centre of sphere 0
        sp[0:3, 1] = li[:, 1] + np.float32(0.2); sp[3, 1] = 1.0        # light 1 off-centre inside sphere 1
    elif case == "light_touching_sphere":
        sp[0:3, 0] = li[:, 0] + np.array([0.5, 0, 0], np.float32); sp[3, 0] = 0.5          # light exactly on the surface
        sp[0:3, 1] = li[:, 1] + np.array([0, 0.5004, 0], np.float32); sp[3, 1] = 0.5       # just outside
        sp[0:3, 2] = li[:, 2] + np.array([0, 0, -0.4996], np.float32); sp[3, 2] = 0.5      # just inside
    elif case == "cluster_in_one_cell":
        c = np.array([4.0, 0.5, 0.6], np.float32)                       # 250 small spheres in a 0.4-wide ball: one grid cell of every light
        sp[0:3, :250] = c[:, None] + rng.uniform(-0.2, 0.2, (3, 250)).astype(np.float32); sp[3, :250] = rng.uniform(0.01, 0.03, 250)
    elif case == "far_light":
        li[:, 0] = [300.0, 40.0, 200.0]                                 # beyond the grid's coordinate range: every cell scans all spheres
        li[:, 1] = [-2.0, 0.0, 260.0]                                   # |V - P| > 256 for every hit: per-ray fallback
    elif case == "light_below_plane":
        li[:, 0] = [3.0, 0.0, -2.0]; li[:, 1] = [3.0, 1.0, 0.05]        # light under / just above the ground plane
    elif case == "grazing":
        li[:, 0] = [-30.0, 0.0, 0.3]; li[:, 1] = [40.0, 25.0, 0.2]; li[:, 2] = [4.0, 0.0, 0.26]     # lights at sphere height: rays graze the whole field
    got, st = _cuda(g, stats=True)
    dense = _cuda(g, flags=_lib.F_DENSE)
    assert np.array_equal(got, dense), case
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_SGRID)), case
    assert st["shadow_traced"] > 0


def test_fuzz_default_equals_brute_force_and_scan_path():
    """Random scenes (tools/gpu_fuzz.py: lights inside / among / far from the spheres, clustered and large spheres, 0-2 planes,
    depth 0-3, anti-aliasing on and off): the default path, B200RT_F_DENSE and B200RT_F_NO_SGRID write the same bytes."""
    import importlib.util, os
    spec = importlib.util.spec_from_file_location("gpu_fuzz", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "gpu_fuzz.py"))
    fz = importlib.util.module_from_spec(spec); spec.loader.exec_module(fz)
    total = 0
    for seed in range(500, 532):
        ok, desc, chk, d1, d2 = fz.check(seed)
        assert ok, (seed, desc, d1, d2)
        total += chk                       # (a single scene may be all black: spheres off camera, no plane)
    assert total > 0


@pytest.fixture
def max_paths():
    """Sets B200RT_MAX_PATHS (read by the library at every call) and restores it."""
    import os
    old = os.environ.get("B200RT_MAX_PATHS")

    def set_(v):
        if v is None:
            os.environ.pop("B200RT_MAX_PATHS", None)
        else:
            os.environ["B200RT_MAX_PATHS"] = str(v)
    yield set_
    set_(old)


@pytest.mark.parametrize("name", ["c1_aa1", "rand_b", "c4_small", "synthetic_512x256", "synthetic_aa_272x144"])
def test_multi_chunk_wavefront_equals_one_chunk(max_paths, name):
    """A frame whose paths exceed the chunk capacity is rendered as several wavefronts over column ranges (C4 with
    anti-aliasing: 7 chunks of <= 48 Mi paths).  Forced here with B200RT_MAX_PATHS: every byte equals the one-chunk render,
    with tiled (16-column multiples) and column-major slot orders, anti-aliased (ordered compaction per chunk) or not."""
    if name.startswith("synthetic_aa"):
        g = _synthetic(700, 2, 1, 272, 144, 2, True, seed=91)
    elif name.startswith("synthetic"):
        g = _synthetic(1024, 3, 1, 512, 256, 3, False, seed=90)
    else:
        g = load_golden(name)
    max_paths(None)
    one, st1 = _cuda(g, stats=True)
    H, ns = int(g["H"]), 9 if bool(g["aliasing"]) else 1
    for paths in (4096, 16 * H * ns + 5, 40 * H * ns):           # 1-2 columns per chunk (untiled), exactly 16, 32 (tiled: rounded down to 16s)
        max_paths(max(1024, paths))
        many, stn = _cuda(g, stats=True)
        assert np.array_equal(one, many), (name, paths)
        assert stn["rays"] == st1["rays"] and stn["primary"] == st1["primary"], (name, paths)
        assert np.array_equal(one, _cuda(g)), (name, paths, "production kernels")
    max_paths(None)


def test_c4_antialiased_crops_vs_simulator_golden():
    """C4 with anti-aliasing (SURVEY §8(d) 'secondary'): crops of the full 7680x4320 frame, 9 samples per pixel, against the
    reference itself under the simulator — including the flat-wrap corner x = W-1 / y = H-1."""
    if not have_golden("c4_aa1_crops"):
        pytest.skip("golden not generated")
    g = load_golden("c4_aa1_crops")
    assert bool(g["aliasing"])
    tot = dict(n=0, n_bad=0, max=0)
    for (x0, x1, y0, y1, ref) in golden_regions(g):
        xa = (x0 // 16) * 16
        got = _cuda(g, tile=(xa, min(int(g["W"]), xa + 32)))[:, x0:x1, y0:y1]
        rep = parity_report(got, ref)
        tot["n"] += rep["n"]; tot["n_bad"] += rep["n_bad"]; tot["max"] = max(tot["max"], rep["max"])
    print("c4 AA crops", tot)
    assert tot["n_bad"] <= _budget(tot["n"]), tot


def test_c4_second_simulator_pin():
    """28 more 16x16 crops of the full-size C4 frame rendered by the reference under the simulator (round 2: the pin grown 4x)."""
    if not have_golden("c4_crops_b"):
        pytest.skip("golden not generated")
    g = load_golden("c4_crops_b")
    tot = dict(n=0, n_bad=0, max=0)
    for (x0, x1, y0, y1, ref) in golden_regions(g):
        xa = (x0 // 16) * 16
        got = _cuda(g, tile=(xa, min(int(g["W"]), xa + 32)))[:, x0:x1, y0:y1]
        rep = parity_report(got, ref)
        tot["n"] += rep["n"]; tot["n_bad"] += rep["n_bad"]; tot["max"] = max(tot["max"], rep["max"])
    print("c4 crops b", tot)
    assert tot["n_bad"] <= _budget(tot["n"]), tot


def test_bit_stable_on_multi_stream_frame_sizes():
    """Determinism where the launch sequence runs on several streams (S >= 64, <= 12 Mi paths: chain / shade streams) and
    on a strip-sharded share of the full-size C4 frame: repeated renders are byte-identical."""
    g = _synthetic(1024, 3, 1, 1024, 512, 3, False, seed=5)
    a = _cuda(g)
    for _ in range(3):
        assert np.array_equal(a, _cuda(g))
    from b200rt import render_host
    from b200rt.workloads import workload
    w = workload("c4"); cam = w["cameras"][0]; W, H = cam.resolution
    sp, li, pl = w["scene"].generate_scene()
    args = (cam.position, cam.rotation, sp, li, pl, w["amb"], w["lamb"], w["refl"], w["refl_depth"], w["aliasing"])
    outs = [render_host(None, *args, grid=cam.grid(), shape=(W, H), strips=(16, 5, 8)) for _ in range(3)]
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
    assert outs[0][:, 80:96].any() and not outs[0][:, 0:80].any()           # only rank 5's strips were written


@pytest.mark.parametrize("case", ["far_plane_origins", "tiny_radii", "coords_near_250", "camera_inside_sphere", "huge_and_tiny"])
def test_filter_margin_stress_scenes(case):
    """Scenes that stress the hand-derived margins of the conservative FFMA2 filter (isect3.cuh): the default path, the scan
    path and brute force all run that filter, so they are compared here with B200RT_F_NO_FILTER — sphere_exact on EVERY
    sphere, no filter at all.  A margin that is too tight would drop a hit and change a byte."""
    from b200rt import _lib
    from b200rt.scene import Camera
    S, L, W, H = 400, 3, 256, 144
    g = _synthetic(S, L, 1, W, H, 3, False, seed=23)
    sp, li = g["spheres"], g["lights"]
    rng = np.random.default_rng(17)
    if case == "far_plane_origins":             # camera almost parallel to the plane: bounce / shadow origins up to ~900 units out
        cam = Camera(resolution=(W, H), position=[-2, 0, 0.6], euler=[0, -0.05, 0])
        g.update(grid=np.array(cam.grid()), cam_origin=cam.position, cam_rotation=cam.rotation)
        sp[0] = rng.uniform(5, 900, S); sp[1] = rng.uniform(-60, 60, S); sp[3] = rng.uniform(0.3, 6.0, S); sp[2] = sp[3] + rng.uniform(0, 3, S)
    elif case == "tiny_radii":
        sp[3] = rng.uniform(1e-3, 4e-3, S); sp[2] = sp[3] + rng.uniform(0, 0.02, S)
        sp[0] = rng.uniform(0.5, 3.0, S); sp[1] = rng.uniform(-1, 1, S)
    elif case == "coords_near_250":             # |C| close to the light grids' coordinate range, camera and lights moved along
        off = np.array([240.0, -230.0, 0.0], np.float32)
        sp[0] += off[0]; sp[1] += off[1]; li[0] += off[0]; li[1] += off[1]
        cam = Camera(resolution=(W, H), position=[-2 + 240.0, -230.0, 2.0], euler=[0, -30, 0])
        g.update(grid=np.array(cam.grid()), cam_origin=cam.position, cam_rotation=cam.rotation)
        for p in range(g["planes"].shape[1]):
            g["planes"][0, p] += 240.0; g["planes"][1, p] -= 230.0
    elif case == "camera_inside_sphere":        # the far root is taken (intersections.py:33-36) for every primary ray
        sp[0:3, 0] = [-2.0, 0.0, 2.0]; sp[3, 0] = 6.0
        sp[0:3, 1] = [-1.5, 0.2, 1.8]; sp[3, 1] = 0.45      # and a small sphere just in front of the camera, inside the big one
    elif case == "huge_and_tiny":
        sp[3, :20] = rng.uniform(20, 120, 20); sp[0, :20] = rng.uniform(100, 400, 20); sp[2, :20] = sp[3, :20] * rng.uniform(0.2, 1.0, 20)
        sp[3, 20:200] = rng.uniform(2e-3, 1e-2, 180)
    got = _cuda(g)
    assert got.any()
    exact = _cuda(g, flags=_lib.F_NO_FILTER)
    assert np.array_equal(got, exact), (case, int((got != exact).sum()))
    assert np.array_equal(got, _cuda(g, flags=_lib.F_DENSE)), case
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_SGRID)), case


@pytest.mark.parametrize("S,W,H,depth,aa,seed", [(1024, 512, 256, 3, False, 31), (300, 256, 144, 2, True, 32), (2500, 192, 128, 2, False, 33), (64, 128, 96, 4, False, 34)])
def test_closest_hit_grid_equals_beam_scan_and_brute_force(S, W, H, depth, aa, seed):
    """Bounce rays of small frames find their closest hit by a per-ray walk through a uniform grid of the spheres
    (csrc/closest_grid.cuh).  Same bytes as the beam-culled scan (B200RT_F_NO_CGRID), brute force and the unfiltered exact test."""
    from b200rt import _lib
    g = _synthetic(S, 3, 1, W, H, depth, aa, seed=seed)
    got, st = _cuda(g, stats=True)
    scan, st_scan = _cuda(g, flags=_lib.F_NO_CGRID, stats=True)
    assert np.array_equal(got, scan), "grid walk differs from the beam-culled scan"
    assert st["rays"] == st_scan["rays"]
    assert st["tests_general"] < st_scan["tests_general"]            # the bounce queries did not run through k_isect3
    assert np.array_equal(got, _cuda(g, flags=_lib.F_DENSE))
    if S <= 1024:
        assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_FILTER))


@pytest.mark.parametrize("case", ["huge_sphere_on_big_list", "out_of_range_falls_back", "flat_scene", "camera_inside_sphere", "lists_overflow_capacity", "nan_sphere"])
def test_closest_hit_grid_edge_cases(case):
    from b200rt import _lib
    from b200rt.scene import Camera
    g = _synthetic(200, 2, 1, 256, 144, 3, False, seed=41)
    sp = g["spheres"]
    if case == "huge_sphere_on_big_list":
        sp[0:3, 0] = [6.0, 0.0, -40.0]; sp[3, 0] = 41.0             # a dome under the scene: covers most cells
        sp[0:3, 1] = [4.0, 1.0, 3.0]; sp[3, 1] = 2.5
    elif case == "out_of_range_falls_back":
        sp[0:3, 5] = [300.0, 10.0, 2.0]; sp[3, 5] = 1.0              # max |C| > 256: the grid is unusable, k_isect3 takes the queue
    elif case == "flat_scene":
        sp[2] = sp[3]                                                # all spheres resting on the plane: one layer of cells
    elif case == "camera_inside_sphere":
        sp[0:3, 0] = [-2.0, 0.0, 2.0]; sp[3, 0] = 5.0
    elif case == "lists_overflow_capacity":                          # every sphere covers hundreds of cells: the id lists exceed the grid's capacity ->
        sp[3] = np.random.default_rng(4).uniform(1.0, 1.3, sp.shape[1])  # the grid marks itself unusable on the device and every ray tests every sphere
        sp[2] = sp[3] + 0.1
    elif case == "nan_sphere":
        sp[0, 3] = np.nan; sp[3, 7] = np.nan                          # never hit, by any path
    got = _cuda(g)
    assert got.any()
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_CGRID)), case
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_FILTER)), case


@pytest.mark.parametrize("res", [(1, 1), (5, 1), (1, 5), (2, 3)])
def test_one_point_axes_render_like_the_oracle(oracle, res):
    """W == 1 or H == 1: the closed-form grid has no step to divide by (Camera.grid: mgrid returns `start`)."""
    from b200rt import render_frame
    from b200rt.scene import Scene, Camera
    scene = Scene.default_scene()
    for aa in (False, True):
        cam = Camera(resolution=res, position=[-2, 0, 2.0], euler=[0, -30, 0])
        got = render_frame(scene, cam, 0.1, 0.6, 0.3, 2, aa)
        sp, li, pl = scene.generate_scene()
        ref = oracle.render(np.asarray(cam.generate_pixel_locations()), cam.position, cam.rotation, sp, li, pl, 0.1, 0.6, 0.3, 2, aa)
        assert got.shape == (3, *res)
        assert np.abs(got.astype(int) - ref.astype(int)).max() <= 1, (res, aa)


@pytest.mark.parametrize("name", ["c1_aa0", "c1_aa1", "rand_a", "rand_b", "rand_c", "edge_s0", "edge_l0", "edge_p0", "edge_d0", "synthetic_13_17_2", "synthetic_63_3_1_aa"])
def test_fused_small_scene_path_equals_wavefront(name):
    """Scenes with fewer than 64 spheres run one fused kernel per depth (csrc/fused_small.cuh): same bytes and the same ray counts as the
    wavefront kernels (B200RT_F_NO_FUSED), brute force and the unfiltered exact test."""
    from b200rt import _lib
    if name == "synthetic_13_17_2":
        g = _synthetic(13, 17, 2, 192, 112, 3, False, seed=77)
    elif name == "synthetic_63_3_1_aa":
        g = _synthetic(63, 3, 1, 160, 96, 2, True, seed=78)
    else:
        g = load_golden(name)
    got, st = _cuda(g, stats=True)
    wave, st_w = _cuda(g, flags=_lib.F_NO_FUSED, stats=True)
    assert np.array_equal(got, wave), "fused small-scene path differs from the wavefront kernels"
    for k in ("primary", "shadow", "bounce", "rays", "shadow_traced"):
        assert st[k] == st_w[k], (k, st[k], st_w[k])
    assert np.array_equal(got, _cuda(g, flags=_lib.F_DENSE))
    assert np.array_equal(got, _cuda(g, flags=_lib.F_NO_FILTER))
    assert np.array_equal(got, _cuda(g))                                    # production launch (no stats), replayed graph on the repeat
    assert np.array_equal(got, _cuda(g))
