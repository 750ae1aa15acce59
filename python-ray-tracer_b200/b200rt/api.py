"""Convenience front ends over the C ABI: whole-frame and tiled renders from host data.

`render_host` is the to_device / launch / copy_to_host sequence of reference src/main.py:19-51 as ONE call
into libb200rt (b200rt_render_host): host buffers in, host uint8 (3,W,H) out.  Pure ctypes + numpy.
"""
import ctypes as C
import numpy as np
from . import _lib


def _fill_common(a, W, H, cam_origin, cam_rotation, spheres, lights, planes, amb, lamb, refl, refl_depth, aliasing,
                 device, flags, tile, strips=None):
    a.struct_size = C.sizeof(_lib.Args)
    a.flags = flags
    a.device = int(device)
    a.W, a.H = int(W), int(H)
    a.x_begin, a.x_end = (0, W) if tile is None else (int(tile[0]), int(tile[1]))
    if strips is not None:                       # (strip_width, rank, world): interleaved column strips of this tile
        a.strip_width, a.strip_rank, a.strip_world = [int(v) for v in strips]
    co = np.ascontiguousarray(cam_origin, dtype=np.float64).ravel()
    cr = np.ascontiguousarray(cam_rotation, dtype=np.float64).ravel()
    if co.size != 3 or cr.size != 9:
        raise ValueError("camera_origin must have 3 elements and camera_rotation 9")
    for i in range(3):
        a.cam_origin[i] = co[i]
    for i in range(9):
        a.cam_rot[i] = cr[i]
    for name, arr, rows in (("spheres", spheres, 7), ("lights", lights, 3), ("planes", planes, 9)):
        if arr.ndim != 2 or arr.shape[0] != rows:
            raise ValueError(f"{name} must have shape ({rows}, N), got {arr.shape}")
    a.spheres, a.S = (spheres.ctypes.data if spheres.size else None), spheres.shape[1]
    a.lights, a.L = (lights.ctypes.data if lights.size else None), lights.shape[1]
    a.planes, a.Pl = (planes.ctypes.data if planes.size else None), planes.shape[1]
    a.amb, a.lamb, a.refl = float(amb), float(lamb), float(refl)
    a.refl_depth, a.aliasing = int(refl_depth), int(bool(aliasing))


def render_host(pixel_loc, cam_origin, cam_rotation, spheres, lights, planes, amb, lamb, refl, refl_depth, aliasing,
                *, grid=None, shape=None, out=None, device=0, tile=None, strips=None, flags=0, stats=False):
    """Host arrays in, host uint8 (3,W,H) out, through b200rt_render_host.

    pixel_loc: float64 (3,W,H) host array, or None with grid=(f, ar, sy, sz) and shape=(W, H).
    Returns `out` (or (out, stats_dict) when stats=True).  Only the tile's columns are written.
    """
    L = _lib.lib()
    a = _lib.Args()
    if pixel_loc is not None and getattr(pixel_loc, "grid", None) is not None and grid is None:
        grid, shape, pixel_loc = pixel_loc.grid, pixel_loc.shape[1:], None
    keep = []
    if pixel_loc is not None:
        pl = np.ascontiguousarray(pixel_loc, dtype=np.float64)
        if pl.ndim != 3 or pl.shape[0] != 3:
            raise ValueError(f"pixel_loc must have shape (3, W, H), got {pl.shape}")
        W, H = pl.shape[1:]
        a.pixel_loc = pl.ctypes.data
        keep.append(pl)
    else:
        if grid is None or shape is None:
            raise ValueError("either pixel_loc or grid+shape is required")
        W, H = shape
        a.pixel_loc = None
        a.grid = _lib.Grid(*[float(v) for v in grid])
    sp = np.ascontiguousarray(spheres, dtype=np.float32)
    li = np.ascontiguousarray(lights, dtype=np.float32)
    pn = np.ascontiguousarray(planes, dtype=np.float32)
    keep += [sp, li, pn]
    _fill_common(a, W, H, cam_origin, cam_rotation, sp, li, pn, amb, lamb, refl, refl_depth, aliasing, device, flags, tile, strips)
    if out is None:
        out = np.zeros((3, W, H), dtype=np.uint8)
    if out.shape != (3, W, H) or out.dtype != np.uint8 or not out.flags.c_contiguous:
        raise ValueError("out must be a C-contiguous uint8 array of shape (3, W, H)")
    a.result = out.ctypes.data
    st = _lib.Stats() if stats else None
    a.stats = C.pointer(st) if st is not None else None
    _lib.check(L.b200rt_render_host(C.byref(a)))
    return (out, st.as_dict()) if stats else out


def render_frame(scene, camera, amb=0.0, lamb=0.6, refl=0.3, refl_depth=2, aliasing=True, devices=None, strip=None, **kw):
    """Scene + Camera -> uint8 (3,W,H) host array (channel wiring as the reference stores it: R,B,G).

    devices=[d0, d1, ...]: the frame is cut into interleaved column strips (sharding.strip_ranges), device i renders
    strips s % N == i, and every device copies its strips straight into the one host array (direct D2H gather,
    SURVEY §8(e)); one host thread per device (ctypes releases the GIL).  Output is bit-identical to one device."""
    spheres, lights, planes = scene.generate_scene()
    W, H = camera.resolution
    common = (None, camera.position, camera.rotation, spheres, lights, planes, amb, lamb, refl, refl_depth, aliasing)
    if not devices or len(devices) == 1:
        if devices:
            kw["device"] = devices[0]
        return render_host(*common, grid=camera.grid(), shape=(W, H), **kw)
    from concurrent.futures import ThreadPoolExecutor
    from .sharding import DEFAULT_STRIP
    out = kw.pop("out", None)
    if out is None:
        out = np.zeros((3, W, H), dtype=np.uint8)
    n = len(devices)
    sw = strip or DEFAULT_STRIP

    def one(i):
        return render_host(*common, grid=camera.grid(), shape=(W, H), out=out, device=devices[i], strips=(sw, i, n), **kw)

    with ThreadPoolExecutor(n) as ex:
        res = list(ex.map(one, range(n)))
    if kw.get("stats"):
        tot = {k: sum(r[1][k] for r in res) for k in ("primary", "shadow", "bounce", "rays", "slow_path", "exact_tests", "shadow_traced")}
        tot["ms_total"] = max(r[1]["ms_total"] for r in res)
        return out, tot
    return out


def render_batch(scene, cameras, amb=0.0, lamb=0.6, refl=0.3, refl_depth=2, aliasing=True, devices=(0,), out=None, flags=0):
    """Camera-pose batch: frame k renders on devices[k % N] (sharding.frames_of_rank); per device ONE
    call into b200rt_render_batch_host (scene staged once, D2H of frame k overlapped with the render of frame k+1).
    All cameras must share the resolution.  Returns uint8 (n, 3, W, H)."""
    from concurrent.futures import ThreadPoolExecutor
    from .sharding import frames_of_rank
    devices = list(devices)
    n = len(cameras)
    W, H = cameras[0].resolution
    if any(tuple(c.resolution) != (W, H) for c in cameras):
        raise ValueError("all cameras of a batch must have the same resolution")
    sp, li, pn = [np.ascontiguousarray(a, dtype=np.float32) for a in scene.generate_scene()]
    if out is None:
        out = np.zeros((n, 3, W, H), dtype=np.uint8)
    if out.shape != (n, 3, W, H) or out.dtype != np.uint8 or not out.flags.c_contiguous:
        raise ValueError("out must be a C-contiguous uint8 array of shape (n, 3, W, H)")
    L = _lib.lib()

    def one(i):
        ks = frames_of_rank(n, i, len(devices))
        if not ks:
            return
        a = _lib.Args()
        a.pixel_loc = None
        a.grid = _lib.Grid(*[float(v) for v in cameras[0].grid()])
        _fill_common(a, W, H, cameras[0].position, cameras[0].rotation, sp, li, pn, amb, lamb, refl, refl_depth, aliasing,
                     devices[i], flags, None)
        org = np.ascontiguousarray([np.asarray(cameras[k].position, dtype=np.float64).ravel() for k in ks])
        rot = np.ascontiguousarray([np.asarray(cameras[k].rotation, dtype=np.float64).ravel() for k in ks])
        contiguous = len(devices) == 1
        res = out if contiguous else np.zeros((len(ks), 3, W, H), dtype=np.uint8)
        a.result = res.ctypes.data                      # (validated as non-NULL; the batch writes through `results`)
        _lib.check(L.b200rt_render_batch_host(C.byref(a), len(ks), org.ctypes.data, rot.ctypes.data, res.ctypes.data))
        if not contiguous:
            out[ks] = res

    if len(devices) == 1:
        one(0)
    else:
        with ThreadPoolExecutor(len(devices)) as ex:           # ctypes releases the GIL; one batch in flight per device
            list(ex.map(one, range(len(devices))))
    return out


def fp32_probe(device=0):
    """Measured FP32 CUDA-core peak of the device: dict(ffma_tflops, ffma2_tflops, sm_mhz)."""
    f1, f2, mhz = C.c_double(), C.c_double(), C.c_double()
    _lib.check(_lib.lib().b200rt_fp32_probe(int(device), C.byref(f1), C.byref(f2), C.byref(mhz)))
    return dict(ffma_tflops=f1.value, ffma2_tflops=f2.value, sm_mhz=mhz.value)
