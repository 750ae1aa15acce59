"""oracle/sim_harness.py — run the UNMODIFIED reference `render` under the numba CUDA simulator.

TEST INFRASTRUCTURE ONLY (see oracle/render_f64.c header).  This module is usable only where
/root/reference exists (the build container); it is what generates the committed goldens under
tests/golden/ (tests/golden/make_golden.py).  Nothing on the GPU box imports it.

What it does, without editing the reference:
  * sets NUMBA_ENABLE_CUDASIM=1 before numba is imported and puts /root/reference/src on sys.path;
  * re-creates main.py's call sequence (src/main.py:19-51) with an EXACT launch grid, so that no
    thread has x >= W or y >= H (the kernel's `<=` bounds tests, kernels.py:13,29, otherwise raise
    IndexError in the simulator);
  * feeds the kernel a padded pixel_loc, pad[c,x,y] = flat[c*W*H + x*H + y] for x <= W, y <= H with
    elements past the end of the buffer defined as 0.0 — exactly what a GPU's unchecked flat
    addressing reads for the anti-alias neighbours at x+1 == W / y+1 == H;
  * shards column tiles over a process pool (the simulator runs one core per process); each
    worker gets a one-pixel halo so that only the global x == 0 column / y == 0 row are un-aliased.
"""
import os
os.environ["NUMBA_ENABLE_CUDASIM"] = "1"
import sys
import time
import multiprocessing as mp
import numpy as np

REFERENCE_SRC = os.environ.get("B200RT_REFERENCE_SRC", "/root/reference/src")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "ray_tracing"))


def _import_reference():
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    from numba import cuda  # noqa: F401  (simulator)
    import ray_tracing  # the reference package, imported in place
    return cuda, ray_tracing


def padded_pixel_loc(pixel_loc: np.ndarray, x_lo=0, x_hi=None, y_lo=0, y_hi=None) -> np.ndarray:
    """pad[c,x,y] = flat[c*W*H + x*H + y] for x in [x_lo, x_hi], y in [y_lo, y_hi] (inclusive; defaults
    0..W and 0..H, i.e. shape (3,W+1,H+1)).  Reproduces GPU flat-wrap reads; past-the-end elements are 0.0."""
    _, W, H = pixel_loc.shape
    x_hi = W if x_hi is None else x_hi
    y_hi = H if y_hi is None else y_hi
    flat = np.ascontiguousarray(pixel_loc, dtype=np.float64).reshape(-1)
    c = np.arange(3, dtype=np.int64)[:, None, None]
    x = np.arange(x_lo, x_hi + 1, dtype=np.int64)[None, :, None]
    y = np.arange(y_lo, y_hi + 1, dtype=np.int64)[None, None, :]
    idx = c * W * H + x * H + y
    ok = idx < flat.size
    return np.where(ok, flat[np.minimum(idx, flat.size - 1)], 0.0)


def _largest_divisor(n: int, cap: int) -> int:
    for d in range(min(n, cap), 0, -1):
        if n % d == 0:
            return d
    return 1


def _run_tile(job):
    (pad_tile, hx, hy, cam_o, cam_r, spheres, lights, planes, amb, lamb, refl, depth, aliasing) = job
    cuda, ray_tracing = _import_reference()
    render = ray_tracing.render
    # pad_tile covers local x in [0, nx] and y in [0, ny] (the last index is the +1 neighbour only).
    nx, ny = pad_tile.shape[1] - 1, pad_tile.shape[2] - 1
    res = cuda.to_device(np.zeros((3, nx, ny), dtype=np.uint8))
    by = _largest_divisor(ny, 128)
    if not aliasing:
        # without anti-aliasing the kernel never reads neighbours: hand it the exact-size array
        pad_tile = np.ascontiguousarray(pad_tile[:, :nx, :ny])
    render[(nx, ny // by), (1, by)](cuda.to_device(pad_tile), res, cuda.to_device(cam_o), cuda.to_device(cam_r),
                                    cuda.to_device(spheres), cuda.to_device(lights), cuda.to_device(planes),
                                    amb, lamb, refl, depth, aliasing)
    out = res.copy_to_host()
    return out[:, hx:, hy:]


def render_reference_sim(pixel_loc, cam_origin, cam_rotation, spheres, lights, planes,
                         amb, lamb, refl, refl_depth, aliasing,
                         x_range=None, y_range=None, procs=None, cols_per_job=4, verbose=False):
    """Render region [x0,x1) x [y0,y1) of the frame with the reference kernel under the simulator.

    Scalars are passed as python float/int/bool exactly as src/main.py:11-12,41-42 does.
    Returns uint8 (3, x1-x0, y1-y0) indexed [channel, x, y] (channel wiring as stored by the
    reference: R,B,G — SURVEY F5/F6).
    """
    assert reference_available(), "reference tree not present"
    pixel_loc = np.ascontiguousarray(pixel_loc, dtype=np.float64)
    _, W, H = pixel_loc.shape
    x0, x1 = x_range if x_range is not None else (0, W)
    y0, y1 = y_range if y_range is not None else (0, H)
    hy = 1 if y0 > 0 else 0
    hx0 = 1 if x0 > 0 else 0
    pad = padded_pixel_loc(pixel_loc, x0 - hx0, x1, y0 - hy, y1)   # region + halo only
    px0 = x0 - hx0                                                 # pad's x origin in frame coordinates
    amb, lamb, refl = float(amb), float(lamb), float(refl)
    refl_depth, aliasing = int(refl_depth), bool(aliasing)
    jobs, spans = [], []
    for xa in range(x0, x1, cols_per_job):
        xb = min(xa + cols_per_job, x1)
        hx = 1 if xa > 0 else 0
        tile = np.ascontiguousarray(pad[:, xa - hx - px0: xb + 1 - px0, :])
        jobs.append((tile, hx, hy, np.asarray(cam_origin), np.asarray(cam_rotation),
                     np.asarray(spheres), np.asarray(lights), np.asarray(planes),
                     amb, lamb, refl, refl_depth, aliasing))
        spans.append((xa, xb))
    out = np.zeros((3, x1 - x0, y1 - y0), dtype=np.uint8)
    procs = procs or os.cpu_count() or 1
    t0 = time.time()
    if procs == 1 or len(jobs) == 1:
        results = [_run_tile(j) for j in jobs]
    else:
        ctx = mp.get_context("spawn")
        with ctx.Pool(min(procs, len(jobs))) as pool:
            results = pool.map(_run_tile, jobs, chunksize=1)
    for (xa, xb), r in zip(spans, results):
        out[:, xa - x0: xb - x0, :] = r
    if verbose:
        print(f"[sim] {x1 - x0}x{y1 - y0} px region in {time.time() - t0:.1f} s on {procs} procs", flush=True)
    return out


def reference_device_functions():
    """The reference's @cuda.jit(device=True) functions as plain callables (simulator mode)."""
    _import_reference()
    from ray_tracing import trace as T, intersections as I, common as C
    return {
        "intersect_ray_sphere": I.intersect_ray_sphere,
        "intersect_ray_plane": I.intersect_ray_plane,
        "get_intersection": T.get_intersection,
        "trace": T.trace,
        "sample": T.sample,
        "clip_color": C.clip_color,
        "clip_color_vector": C.clip_color_vector,
    }


def reference_scene_module():
    """The reference's `scene` package (Scene, Camera, ...) imported in place."""
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import scene
    return scene
