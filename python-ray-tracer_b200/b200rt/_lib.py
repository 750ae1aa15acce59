"""ctypes binding of libb200rt.so (include/b200rt.h).  No fallback: a missing library is a hard error."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200rt.so")

F_SCENE_ON_HOST, F_NO_FILTER, F_UNSWAP_GB, F_NO_CULL, F_NO_SORT, F_TIME_ONLY, F_DENSE, F_NO_SGRID = 0x1, 0x2, 0x4, 0x8, 0x10, 0x20, 0x40, 0x80
F_NO_GRAPH, F_NO_CACHE, F_NO_CGRID, F_NO_FUSED = 0x100, 0x200, 0x400, 0x800


class Grid(C.Structure):
    _fields_ = [("f", C.c_double), ("ar", C.c_double), ("sy", C.c_double), ("sz", C.c_double)]


class Stats(C.Structure):
    _fields_ = [("primary", C.c_uint64), ("shadow", C.c_uint64), ("bounce", C.c_uint64), ("slow_path", C.c_uint64), ("exact_tests", C.c_uint64), ("shadow_traced", C.c_uint64), ("tests_pivot", C.c_uint64), ("tests_general", C.c_uint64),
                ("kernel_ms", C.c_float * 4)]

    def as_dict(self):
        return dict(primary=int(self.primary), shadow=int(self.shadow), bounce=int(self.bounce),
                    rays=int(self.primary + self.shadow + self.bounce), slow_path=int(self.slow_path), exact_tests=int(self.exact_tests), shadow_traced=int(self.shadow_traced), tests_pivot=int(self.tests_pivot), tests_general=int(self.tests_general),
                    ms_total=float(self.kernel_ms[0]), ms_closest=float(self.kernel_ms[1]),
                    ms_shadow=float(self.kernel_ms[2]), ms_other=float(self.kernel_ms[3]))


class Args(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("flags", C.c_uint32), ("device", C.c_int32), ("reserved0", C.c_int32),
        ("stream", C.c_void_p),
        ("pixel_loc", C.c_void_p), ("grid", Grid), ("result", C.c_void_p),
        ("W", C.c_int32), ("H", C.c_int32), ("x_begin", C.c_int32), ("x_end", C.c_int32),
        ("cam_origin", C.c_double * 3), ("cam_rot", C.c_double * 9),
        ("spheres", C.c_void_p), ("S", C.c_int32),
        ("lights", C.c_void_p), ("L", C.c_int32),
        ("planes", C.c_void_p), ("Pl", C.c_int32),
        ("amb", C.c_double), ("lamb", C.c_double), ("refl", C.c_double),
        ("refl_depth", C.c_int64), ("aliasing", C.c_int32), ("reserved1", C.c_int32),
        ("strip_width", C.c_int32), ("strip_rank", C.c_int32), ("strip_world", C.c_int32), ("reserved2", C.c_int32),
        ("stats", C.POINTER(Stats)),
    ]


_lib = None


def lib():
    """Load libb200rt.so.  Raises ImportError with build instructions when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: the CUDA library has not been built. "
                              "Run `python -c 'import __graft_entry__ as g; g.build()'` at the repo root "
                              "(there is no CPU fallback).")
        L = C.CDLL(LIB_PATH)
        L.b200rt_version.restype = C.c_int
        L.b200rt_last_error.restype = C.c_char_p
        L.b200rt_args_size.restype = C.c_uint32
        L.b200rt_render.argtypes = [C.POINTER(Args)]; L.b200rt_render.restype = C.c_int
        L.b200rt_render_host.argtypes = [C.POINTER(Args)]; L.b200rt_render_host.restype = C.c_int
        L.b200rt_render_batch_host.argtypes = [C.POINTER(Args), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]; L.b200rt_render_batch_host.restype = C.c_int
        L.b200rt_fp32_probe.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.b200rt_fp32_probe.restype = C.c_int
        L.b200rt_to_image.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint32]; L.b200rt_to_image.restype = C.c_int
        L.b200rt_to_image_host.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint32]; L.b200rt_to_image_host.restype = C.c_int
        L.b200rt_to_image_host_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint32]; L.b200rt_to_image_host_noalloc.restype = C.c_int
        L.b200rt_alloc_host.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]; L.b200rt_alloc_host.restype = C.c_int
        L.b200rt_free_host.argtypes = [C.c_void_p]; L.b200rt_free_host.restype = C.c_int
        L.b200rt_host_register.argtypes = [C.c_void_p, C.c_size_t]; L.b200rt_host_register.restype = C.c_int
        L.b200rt_host_unregister.argtypes = [C.c_void_p]; L.b200rt_host_unregister.restype = C.c_int
        L.b200rt_device_count.restype = C.c_int
        L.b200rt_sm_count.argtypes = [C.c_int]; L.b200rt_sm_count.restype = C.c_int
        L.b200rt_launch_count.restype = C.c_ulonglong
        L.b200rt_debug_check_failures.restype = C.c_ulonglong
        L.b200rt_graph_replays.argtypes = [C.c_int]; L.b200rt_graph_replays.restype = C.c_ulonglong
        L.b200rt_shutdown.restype = C.c_int
        if L.b200rt_args_size() != C.sizeof(Args):
            raise ImportError("b200rt_args layout mismatch between _lib.py and libb200rt.so")
        _lib = L
    return _lib


def check(rc: int):
    """0 -> ok; <0 -> ValueError (argument errors); >0 -> RuntimeError carrying the CUDA error string."""
    if rc == 0:
        return
    msg = lib().b200rt_last_error().decode("utf-8", "replace")
    if rc < 0:
        raise ValueError(f"b200rt: {msg} (code {rc})")
    raise RuntimeError(f"b200rt: CUDA error {rc}: {msg}")
