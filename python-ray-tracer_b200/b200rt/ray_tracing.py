"""The render entry point with the reference's call shape (src/ray_tracing/__init__.py:1, kernels.py:6-7):

    render[blockspergrid, threadsperblock](pixel_loc, result, camera_origin, camera_rotation,
                                           spheres, lights, planes, amb, lamb, refl, refl_depth, aliasing)

12 positional arguments, returns None, writes `result` in place, asynchronous on the stream (numba semantics,
src/main.py:41-42).  Grid/block are accepted and ignored: libb200rt owns its launch configuration (which also
removes the out-of-range threads of the reference's `<=` bounds tests, kernels.py:13).

Everything computes in libb200rt.so (hand-written CUDA, sm_100a) through ctypes; there is no fallback.
"""
import ctypes as C
import numpy as np
from . import _lib
from .cuda import DeviceArray, PixelGrid, to_device


def _is_torch_tensor(x):
    return type(x).__module__.startswith("torch") and hasattr(x, "data_ptr")


def _shape(x):
    return tuple(x.shape)


def _check_shapes(pixel_loc, result, cam_o, cam_r, spheres, lights, planes):
    ps = _shape(pixel_loc)
    if len(ps) != 3 or ps[0] != 3:
        raise ValueError(f"pixel_loc must have shape (3, W, H), got {ps}")
    if _shape(result) != ps:
        raise ValueError(f"result must have shape {ps} (3, W, H), got {_shape(result)}")
    if int(np.prod(_shape(cam_o))) != 3:
        raise ValueError(f"camera_origin must have 3 elements, got shape {_shape(cam_o)}")
    if _shape(cam_r) != (3, 3):
        raise ValueError(f"camera_rotation must have shape (3, 3), got {_shape(cam_r)}")
    for name, arr, rows in (("spheres", spheres, 7), ("lights", lights, 3), ("planes", planes, 9)):
        s = _shape(arr)
        if len(s) != 2 or s[0] != rows:
            raise ValueError(f"{name} must have shape ({rows}, N), got {s}")
    return ps[1], ps[2]


def _small_host(x, dtype):
    """Host copy of a tiny array argument (camera origin / rotation)."""
    if isinstance(x, DeviceArray):
        h = x._host_small if x._host_small is not None else x.copy_to_host()
    elif _is_torch_tensor(x):
        h = x.detach().cpu().numpy()
    else:
        h = np.asarray(x)
    return np.ascontiguousarray(h, dtype=dtype)


def _device_index(x, default):
    if isinstance(x, DeviceArray):
        return x.device_index
    if _is_torch_tensor(x) and x.is_cuda:
        return x.device.index
    return default


class _Launch:
    def __init__(self, stream=None, flags=0, stats=False, tile=None, strips=None):
        self.stream, self.flags, self.want_stats, self.tile, self.strips = stream, flags, stats, tile, strips
        self.stats = None

    def __call__(self, pixel_loc, result, camera_origin, camera_rotation, spheres, lights, planes,
                 amb, lamb, refl, refl_depth, aliasing):
        import torch
        W, H = _check_shapes(pixel_loc, result, camera_origin, camera_rotation, spheres, lights, planes)
        dev = _device_index(result, _device_index(spheres, torch.cuda.current_device()))
        keep = []

        def dev_ptr(x, np_dtype, torch_dtype):
            """Device pointer of an array argument; host arrays are staged, other dtypes are cast (the reference
            JIT-specialises on whatever dtype arrives; this path is float64 grid / float32 scene)."""
            if isinstance(x, DeviceArray):
                x = x.tensor
            if not _is_torch_tensor(x):
                x = torch.from_numpy(np.ascontiguousarray(np.asarray(x), dtype=np_dtype)).to(f"cuda:{dev}")
            if x.dtype != torch_dtype or not x.is_contiguous() or not x.is_cuda:
                x = x.to(device=f"cuda:{dev}", dtype=torch_dtype).contiguous()
            keep.append(x)
            return x.data_ptr() if x.numel() else None

        a = _lib.Args()
        a.struct_size = C.sizeof(_lib.Args)
        a.flags = self.flags
        a.device = dev
        st = self.stream
        if st is None or st == 0:
            st = torch.cuda.current_stream(dev)
        a.stream = int(st.cuda_stream) if hasattr(st, "cuda_stream") else int(st)

        grid = getattr(pixel_loc, "grid", None) if isinstance(pixel_loc, (DeviceArray, PixelGrid)) else None
        if grid is not None:
            a.pixel_loc = None
            a.grid = _lib.Grid(*grid)
        else:
            a.pixel_loc = dev_ptr(pixel_loc, np.float64, torch.float64)

        host_result = None
        if isinstance(result, DeviceArray) or _is_torch_tensor(result):
            rt = result.tensor if isinstance(result, DeviceArray) else result
            if rt.dtype != torch.uint8 or not rt.is_cuda or not rt.is_contiguous():
                raise ValueError("result must be a contiguous uint8 CUDA array of shape (3, W, H)")
            keep.append(rt)
        else:                                        # host ndarray: numba would copy in and back; so do we
            host_result = result
            if host_result.dtype != np.uint8:
                raise ValueError("result must be uint8")
            rt = torch.from_numpy(np.ascontiguousarray(host_result)).to(f"cuda:{dev}")
            keep.append(rt)
        a.result = rt.data_ptr()
        a.W, a.H = W, H
        a.x_begin, a.x_end = (0, W) if self.tile is None else (int(self.tile[0]), int(self.tile[1]))
        if self.strips is not None:
            a.strip_width, a.strip_rank, a.strip_world = [int(v) for v in self.strips]
        co = _small_host(camera_origin, np.float64).ravel()
        cr = _small_host(camera_rotation, np.float64).ravel()
        for i in range(3):
            a.cam_origin[i] = co[i]
        for i in range(9):
            a.cam_rot[i] = cr[i]
        a.spheres, a.S = dev_ptr(spheres, np.float32, torch.float32), _shape(spheres)[1]
        a.lights, a.L = dev_ptr(lights, np.float32, torch.float32), _shape(lights)[1]
        a.planes, a.Pl = dev_ptr(planes, np.float32, torch.float32), _shape(planes)[1]
        a.amb, a.lamb, a.refl = float(amb), float(lamb), float(refl)
        a.refl_depth, a.aliasing = int(refl_depth), int(bool(aliasing))
        stats = _lib.Stats() if self.want_stats else None
        a.stats = C.pointer(stats) if stats is not None else None
        _lib.check(_lib.lib().b200rt_render(C.byref(a)))
        # staged temporaries must outlive the asynchronous launch: tie them to the launch stream
        if a.stream != int(torch.cuda.current_stream(dev).cuda_stream):
            ext = torch.cuda.ExternalStream(a.stream, device=dev) if a.stream else torch.cuda.default_stream(dev)
            for t in keep:
                t.record_stream(ext)
        if stats is not None:
            self.stats = stats.as_dict()
        if host_result is not None:
            host_result[...] = rt.cpu().numpy()
        return None


class _Render:
    """`render[grid, block]`, `render[grid, block, stream]` -> launcher (numba dispatcher syntax)."""

    def __getitem__(self, cfg):
        stream = None
        if isinstance(cfg, tuple) and len(cfg) >= 3:
            stream = cfg[2]
        return _Launch(stream=stream)

    def configure(self, stream=None, flags=0, stats=False, tile=None, strips=None):
        """Library-specific launcher: explicit stream, B200RT_F_* flags, ray statistics, column tile,
        interleaved strips (strip_width, rank, world) for multi-GPU sharding."""
        return _Launch(stream=stream, flags=flags, stats=stats, tile=tile, strips=strips)

    def __call__(self, *args, **kw):
        raise TypeError("render is a kernel: launch it as render[blockspergrid, threadsperblock](...)")


render = _Render()
