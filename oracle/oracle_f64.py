"""oracle/oracle_f64.py — ctypes front end of oracle/render_f64.c (float64 CPU restatement).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs.  The product package (python-ray-tracer_b200/b200rt) never imports this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "render_f64.c")
_LIB = os.path.join(_HERE, "liboracle_f64.so")


def build(force: bool = False) -> str:
    """gcc -O2 -fopenmp -ffp-contract=off  (no FMA contraction: the simulator runs plain Python floats)."""
    if force or not os.path.exists(_LIB) or (os.path.exists(_SRC) and os.path.getmtime(_SRC) > os.path.getmtime(_LIB)):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
                               "-fvisibility=hidden", "-o", _LIB, _SRC, "-lm"])
    return _LIB


class OracleArgs(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("W", C.c_int), ("H", C.c_int),
        ("x_begin", C.c_int), ("x_end", C.c_int), ("y_begin", C.c_int), ("y_end", C.c_int),
        ("pixel_loc", C.c_void_p),
        ("grid_f", C.c_double), ("grid_ar", C.c_double), ("grid_sy", C.c_double), ("grid_sz", C.c_double),
        ("cam_origin", C.c_double * 3), ("cam_rot", C.c_double * 9),
        ("spheres", C.c_void_p), ("S", C.c_int),
        ("lights", C.c_void_p), ("L", C.c_int),
        ("planes", C.c_void_p), ("Pl", C.c_int),
        ("amb", C.c_double), ("lamb", C.c_double), ("refl", C.c_double),
        ("refl_depth", C.c_int64), ("aliasing", C.c_int), ("numpy_weak", C.c_int), ("threads", C.c_int),
        ("result", C.c_void_p), ("result_f64", C.c_void_p), ("counts", C.c_void_p),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.oracle_render.argtypes = [C.POINTER(OracleArgs)]
        _lib.oracle_render.restype = C.c_int
        dp, fp = C.POINTER(C.c_double), C.POINTER(C.c_float)
        _lib.oracle_intersect_ray_sphere.argtypes = [dp, dp, fp, C.c_float]
        _lib.oracle_intersect_ray_sphere.restype = C.c_double
        _lib.oracle_intersect_ray_plane.argtypes = [dp, dp, fp, fp]
        _lib.oracle_intersect_ray_plane.restype = C.c_double
        _lib.oracle_get_intersection.argtypes = [dp, dp, fp, C.c_int, fp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        _lib.oracle_get_intersection.restype = C.c_double
        _lib.oracle_trace.argtypes = [dp, dp, fp, C.c_int, fp, C.c_int, fp, C.c_int, C.c_double, C.c_double, C.c_int, dp]
        _lib.oracle_trace.restype = None
        _lib.oracle_sample.argtypes = [dp, dp, fp, C.c_int, fp, C.c_int, fp, C.c_int, C.c_double, C.c_double, C.c_double,
                                       C.c_int64, C.c_int, dp, C.POINTER(C.c_uint64)]
        _lib.oracle_sample.restype = None
        _lib.oracle_clip_color.argtypes = [C.c_double]
        _lib.oracle_clip_color.restype = C.c_int
        _lib.oracle_max_threads.restype = C.c_int
        _lib.oracle_args_size.restype = C.c_uint32
        assert _lib.oracle_args_size() == C.sizeof(OracleArgs), "oracle_args layout mismatch"
    return _lib


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def max_threads() -> int:
    return int(lib().oracle_max_threads())


def render(pixel_loc, cam_origin, cam_rotation, spheres, lights, planes, amb, lamb, refl, refl_depth, aliasing,
           *, grid=None, shape=None, x_range=None, y_range=None, numpy_weak=False, threads=0,
           want_f64=False, want_counts=False, out=None):
    """Render with the float64 oracle.

    pixel_loc: (3,W,H) float64 array, or None with grid=(f, ar, sy, sz) and shape=(W,H)
    (Camera.generate_pixel_locations closed form, reference src/scene/camera.py:18-26).
    Returns uint8 (3,W,H) [+ float64 (3,W,H) pre-clip values] [+ uint64[3] ray counts].
    Only the requested region is computed; the rest of the frame stays 0.
    """
    L_ = lib()
    a = OracleArgs()
    a.struct_size = C.sizeof(OracleArgs)
    keep = []
    if pixel_loc is not None:
        pl = _f64(pixel_loc); keep.append(pl)
        _, W, H = pl.shape
        a.pixel_loc = pl.ctypes.data
    else:
        W, H = shape
        a.pixel_loc = None
        a.grid_f, a.grid_ar, a.grid_sy, a.grid_sz = [float(v) for v in grid]
    a.W, a.H = int(W), int(H)
    a.x_begin, a.x_end = x_range if x_range is not None else (0, W)
    a.y_begin, a.y_end = y_range if y_range is not None else (0, H)
    co, cr = _f64(cam_origin).ravel(), _f64(cam_rotation).ravel()
    for i in range(3):
        a.cam_origin[i] = co[i]
    for i in range(9):
        a.cam_rot[i] = cr[i]
    sp, li, pl_ = _f32(spheres), _f32(lights), _f32(planes)
    keep += [sp, li, pl_]
    a.spheres, a.S = sp.ctypes.data, sp.shape[1]
    a.lights, a.L = li.ctypes.data, li.shape[1]
    a.planes, a.Pl = pl_.ctypes.data, pl_.shape[1]
    a.amb, a.lamb, a.refl = float(amb), float(lamb), float(refl)
    a.refl_depth, a.aliasing = int(refl_depth), int(bool(aliasing))
    a.numpy_weak, a.threads = int(bool(numpy_weak)), int(threads)
    res = out if out is not None else np.zeros((3, W, H), dtype=np.uint8)
    a.result = res.ctypes.data
    f64 = np.zeros((3, W, H), dtype=np.float64) if want_f64 else None
    a.result_f64 = f64.ctypes.data if want_f64 else None
    counts = np.zeros(3, dtype=np.uint64) if want_counts else None
    a.counts = counts.ctypes.data if want_counts else None
    rc = L_.oracle_render(C.byref(a))
    if rc != 0:
        raise RuntimeError(f"oracle_render failed: {rc}")
    ret = [res]
    if want_f64:
        ret.append(f64)
    if want_counts:
        ret.append(counts)
    return ret[0] if len(ret) == 1 else tuple(ret)


def intersect_ray_sphere(O, D, Cc, r):
    O, D, Cc = _f64(O), _f64(D), _f32(Cc)
    return float(lib().oracle_intersect_ray_sphere(_dp(O), _dp(D), _fp(Cc), C.c_float(float(np.float32(r)))))


def intersect_ray_plane(O, D, po, pn):
    O, D, po, pn = _f64(O), _f64(D), _f32(po), _f32(pn)
    return float(lib().oracle_intersect_ray_plane(_dp(O), _dp(D), _fp(po), _fp(pn)))


def get_intersection(O, D, spheres, planes):
    O, D, sp, pl = _f64(O), _f64(D), _f32(spheres), _f32(planes)
    idx, typ = C.c_int(), C.c_int()
    d = lib().oracle_get_intersection(_dp(O), _dp(D), _fp(sp), sp.shape[1], _fp(pl), pl.shape[1], C.byref(idx), C.byref(typ))
    return float(d), idx.value, typ.value


def trace(O, D, spheres, lights, planes, amb, lamb, numpy_weak=False):
    O, D, sp, li, pl = _f64(O), _f64(D), _f32(spheres), _f32(lights), _f32(planes)
    out = np.zeros(9)
    lib().oracle_trace(_dp(O), _dp(D), _fp(sp), sp.shape[1], _fp(li), li.shape[1], _fp(pl), pl.shape[1],
                       float(amb), float(lamb), int(numpy_weak), _dp(out))
    return out[0:3].copy(), out[3:6].copy(), out[6:9].copy()


def sample(O, D, spheres, lights, planes, amb, lamb, refl, refl_depth, numpy_weak=False):
    O, D, sp, li, pl = _f64(O), _f64(D), _f32(spheres), _f32(lights), _f32(planes)
    out = np.zeros(3)
    cnt = np.zeros(3, dtype=np.uint64)
    lib().oracle_sample(_dp(O), _dp(D), _fp(sp), sp.shape[1], _fp(li), li.shape[1], _fp(pl), pl.shape[1],
                        float(amb), float(lamb), float(refl), int(refl_depth), int(numpy_weak), _dp(out),
                        cnt.ctypes.data_as(C.POINTER(C.c_uint64)))
    return out, cnt


def clip_color(c):
    return int(lib().oracle_clip_color(float(c)))


def convert_array_to_image_array(x):
    """numpy restatement of reference src/viewer/image.py:7-19 for SQUARE input: stack (3,W,H) -> (W,H,3), cast to
    uint8, then mirror(rotate(270)) == transpose of the first two axes.  Returns the (H, W, 3) uint8 array the PIL image
    holds.  (Pinned by tests/golden/viewer_square.npz, generated with the reference's own function.)"""
    import numpy as np
    y = np.zeros((x.shape[1], x.shape[2], 3))
    for i in range(3):
        y[:, :, i] = x[i, :, :]
    return np.ascontiguousarray(y.astype(np.uint8).transpose(1, 0, 2))
