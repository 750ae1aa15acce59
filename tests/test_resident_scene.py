"""Resident scene tables (rebuilt only when k_scene_cmp finds the arrays changed) and CUDA-graph replay of the launch sequence.

Every frame rendered through the cached / replayed path must equal, byte for byte, the same frame rendered with
B200RT_F_NO_CACHE (all tables rebuilt, plain launches) — for a moving camera, for scene arrays changed in place behind the same
device pointers, for host scenes, and when launch sequences with different flags interleave."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scene(S, seed, L=3):
    rng = np.random.default_rng(seed)
    sp = np.zeros((7, S), np.float32)
    sp[0] = rng.uniform(0.5, 9.0, S); sp[1] = rng.uniform(-3.5, 3.5, S); sp[3] = rng.uniform(0.05, 0.3, S)
    sp[2] = sp[3] + rng.uniform(0.0, 1.5, S); sp[4:7] = rng.integers(40, 256, (3, S))
    li = np.zeros((3, L), np.float32)
    li[0] = rng.uniform(-2, 8, L); li[1] = rng.uniform(-4, 4, L); li[2] = rng.uniform(3, 7, L)
    pl = np.array([[5, 0, 0, 0, 0, 1, 128, 128, 128]], np.float32).T.copy()
    return sp, li, pl


def _cam(k, res):
    from b200rt.scene import Camera
    return Camera(resolution=res, position=[-2 - 0.1 * k, 0.05 * k, 2.0 + 0.02 * k], euler=[0, -30 + k, 2 * k])


@pytest.mark.parametrize("S,res,depth,aa", [(1024, (512, 256), 3, False), (6, (320, 180), 2, True), (200, (256, 144), 2, False)])
def test_moving_camera_replays_graph_and_equals_uncached(S, res, depth, aa):
    from b200rt import render_host, _lib
    L = _lib.lib()
    sp, li, pl = _scene(S, 3)
    r0 = L.b200rt_graph_replays(0)
    for k in range(6):
        cam = _cam(k, res)
        args = (None, cam.position, cam.rotation, sp, li, pl, 0.05, 0.6, 0.3, depth, aa)
        got = render_host(*args, grid=cam.grid(), shape=res)
        ref = render_host(*args, grid=cam.grid(), shape=res, flags=_lib.F_NO_CACHE)
        assert got.any()
        assert np.array_equal(got, ref), k
    if S < 64:                                            # (frames on the multi-stream schedule keep plain launches)
        assert L.b200rt_graph_replays(0) - r0 >= 4        # frame 0 plain, frame 1 captured + replayed, 2.. replayed


def test_scene_changed_in_place_behind_the_same_device_pointers():
    """The numba-style entry point with device-resident arrays: the arrays' CONTENTS change between launches (same pointers,
    same shapes).  The device-side comparison must notice and rebuild — inside a replayed graph too."""
    import torch
    from b200rt import render_host, _lib
    from b200rt.ray_tracing import render
    from b200rt.cuda import pixel_grid
    res = (384, 192)
    cam = _cam(0, res)
    sp, li, pl = _scene(300, 5)
    d_sp, d_li, d_pl = [torch.from_numpy(a.copy()).cuda() for a in (sp, li, pl)]
    result = torch.zeros((3, *res), dtype=torch.uint8, device="cuda")
    grid = pixel_grid(cam, 0)
    rng = np.random.default_rng(1)
    for k in range(7):
        if k in (2, 3, 5):                                # move spheres / a light / recolour, in place on the device
            sp[0:3, rng.integers(0, 300, 20)] += rng.normal(0, 0.2, (3, 20)).astype(np.float32)
            li[:, k % 3] += np.float32(0.3)
            sp[4:7, 7] = [10 * k, 255 - 10 * k, 99]
            d_sp.copy_(torch.from_numpy(sp)); d_li.copy_(torch.from_numpy(li))
        render[(1, 1), (16, 16)](grid, result, cam.position, cam.rotation, d_sp, d_li, d_pl, 0.0, 0.6, 0.3, 2, False)
        got = result.cpu().numpy()
        ref = render_host(None, cam.position, cam.rotation, sp, li, pl, 0.0, 0.6, 0.3, 2, False, grid=cam.grid(), shape=res, flags=_lib.F_NO_CACHE)
        assert np.array_equal(got, ref), k


def test_interleaved_launch_sequences_and_host_scene_changes():
    """Sequences with different flags / shapes share the resident tables: a scene change seen by one of them (which does not
    build the light grids) must still reach the others."""
    from b200rt import render_host, _lib
    res = (256, 128)
    cam = _cam(1, res)
    spA, liA, plA = _scene(400, 11)
    spB, liB, plB = _scene(400, 12)
    spC, liC, plC = _scene(90, 13, L=2)

    def r(scene, flags=0):
        sp, li, pl = scene
        return render_host(None, cam.position, cam.rotation, sp, li, pl, 0.0, 0.6, 0.3, 2, False, grid=cam.grid(), shape=res, flags=flags)

    A, B, Cc = (spA, liA, plA), (spB, liB, plB), (spC, liC, plC)
    refs = {id(s): r(s, _lib.F_NO_CACHE) for s in (A, B, Cc)}
    seq = [(A, 0), (A, 0), (A, 0), (B, _lib.F_DENSE), (A, 0), (B, _lib.F_NO_SGRID), (B, 0), (B, 0), (Cc, 0), (A, _lib.F_DENSE), (A, 0), (Cc, 0), (Cc, 0), (Cc, 0)]
    for i, (s, fl) in enumerate(seq):
        assert np.array_equal(r(s, fl), refs[id(s)]), (i, fl)
    # same host array object, contents edited
    spA[0, :50] += 0.25
    ref = r(A, _lib.F_NO_CACHE)
    assert not np.array_equal(ref, refs[id(A)])
    for _ in range(3):
        assert np.array_equal(r(A), ref)


def test_strips_and_batch_through_the_cached_path():
    from b200rt import render_host, render_batch, render_frame, _lib
    from b200rt.scene import Scene
    from b200rt.workloads import c5_camera
    res = (512, 256)
    cam = _cam(2, res)
    sp, li, pl = _scene(1024, 21)
    whole = render_host(None, cam.position, cam.rotation, sp, li, pl, 0.0, 0.6, 0.3, 3, False, grid=cam.grid(), shape=res, flags=_lib.F_NO_CACHE)
    for rep in range(3):
        out = np.zeros_like(whole)
        for rank in range(4):
            render_host(None, cam.position, cam.rotation, sp, li, pl, 0.0, 0.6, 0.3, 3, False, grid=cam.grid(), shape=res, out=out, strips=(16, rank, 4))
        assert np.array_equal(out, whole), rep
    scene = Scene.default_scene()
    cams = [c5_camera(k, res=(192, 108)) for k in range(9)]
    got = render_batch(scene, cams, 0.0, 0.6, 0.3, 2, False)
    for k, c in enumerate(cams):
        assert np.array_equal(got[k], render_frame(scene, c, 0.0, 0.6, 0.3, 2, False, flags=_lib.F_NO_CACHE)), k
