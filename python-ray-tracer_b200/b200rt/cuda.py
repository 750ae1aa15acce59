"""Staging shim with the three numba.cuda names the reference's driver uses (src/main.py:19-21,27-29,32,51):
`to_device(ndarray)` -> object with `.shape`, `.dtype`, `.copy_to_host()`; plus `synchronize()`.

Device memory is a torch CUDA tensor (torch is plumbing here: allocator + streams).  Nothing in this module
computes; there is no CPU fallback — without a CUDA device `to_device` raises.
"""
import numpy as np


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("b200rt.cuda: no CUDA device available (this path has no CPU fallback)")
    return torch


class PixelGrid(np.ndarray):
    """float64 (3,W,H) array produced by Camera.generate_pixel_locations() that remembers its closed form
    (f, ar, sy, sz).  Read-only, so the descriptor cannot go stale; any copy / arithmetic result is a plain
    ndarray without it.  Lets render() skip the 24*W*H-byte H2D copy (796 MB at 8K)."""
    grid = None

    def __array_finalize__(self, obj):
        self.grid = None


def as_pixel_grid(arr: np.ndarray, grid) -> "PixelGrid":
    out = np.ascontiguousarray(arr, dtype=np.float64).view(PixelGrid)
    out.grid = tuple(float(v) for v in grid)
    out.flags.writeable = False
    return out


class DeviceArray:
    """Device-resident array: the object `to_device` returns."""

    def __init__(self, tensor=None, host_small=None, grid=None, shape=None, dtype=None, device=0, lazy_host=None):
        self._t = tensor
        self._host_small = host_small        # host copy of tiny arrays (camera origin / rotation): no D2H sync in render()
        self.grid = grid                     # closed-form pixel grid, if known
        self._lazy_host = lazy_host          # grid-backed pixel_loc: materialised on the device only on demand
        self.shape = tuple(shape if shape is not None else tensor.shape)
        self.dtype = np.dtype(dtype) if dtype is not None else None
        self.device_index = device

    @property
    def tensor(self):
        if self._t is None and self._lazy_host is not None:
            torch = _torch()
            self._t = torch.from_numpy(np.array(self._lazy_host)).to(f"cuda:{self.device_index}")
        return self._t

    @property
    def ndim(self):
        return len(self.shape)

    def data_ptr(self) -> int:
        return int(self.tensor.data_ptr())

    @property
    def __cuda_array_interface__(self):
        return self.tensor.__cuda_array_interface__

    def copy_to_host(self, ary=None):
        """Device -> host copy; synchronises like numba's copy_to_host (reference src/main.py:51)."""
        out = np.array(self._lazy_host) if (self._t is None and self._lazy_host is not None) else self.tensor.cpu().numpy()
        if ary is not None:
            ary[...] = out
            return ary
        return out


def to_device(ary, device=None, stream=None):
    """numba.cuda.to_device equivalent (reference src/main.py:19-21,27-29,32)."""
    torch = _torch()
    dev = torch.cuda.current_device() if device is None else int(device)
    if isinstance(ary, DeviceArray):
        return ary
    grid = getattr(ary, "grid", None) if isinstance(ary, PixelGrid) else None
    host = np.ascontiguousarray(np.asarray(ary))
    if grid is not None:
        return DeviceArray(None, grid=grid, shape=host.shape, dtype=host.dtype, device=dev, lazy_host=ary)
    t = torch.from_numpy(host.copy() if not host.flags.writeable else host).to(f"cuda:{dev}", non_blocking=False)
    small = host.copy() if host.size <= 16 else None
    return DeviceArray(t, host_small=small, shape=host.shape, dtype=host.dtype, device=dev)


def pixel_grid(camera, device=None):
    """Device-side stand-in for to_device(camera.generate_pixel_locations()) that never materialises the
    24*W*H-byte array: carries only the closed form (f, ar, sy, sz).  render() accepts it as `pixel_loc`."""
    W, H = camera.resolution
    dev = 0 if device is None else int(device)
    return DeviceArray(None, grid=tuple(camera.grid()), shape=(3, W, H), dtype=np.float64, device=dev,
                       lazy_host=_LazyGrid(camera))


class _LazyGrid:
    """np.array(lazy) materialises the grid only if somebody really asks for the data."""

    def __init__(self, camera):
        self._camera = camera

    def __array__(self, dtype=None, copy=None):
        return np.array(self._camera.generate_pixel_locations(), dtype=dtype or np.float64)


def device_array(shape, dtype=np.float64, device=None):
    torch = _torch()
    dev = torch.cuda.current_device() if device is None else int(device)
    tdt = {np.dtype(np.uint8): torch.uint8, np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64,
           np.dtype(np.int32): torch.int32, np.dtype(np.int64): torch.int64}[np.dtype(dtype)]
    return DeviceArray(torch.zeros(tuple(shape), dtype=tdt, device=f"cuda:{dev}"), shape=shape, dtype=dtype, device=dev)


class _PinnedOwner:
    """Frees a b200rt_alloc_host block when the last numpy view of it goes away."""

    def __init__(self, ptr, nbytes):
        self.ptr, self.nbytes = ptr, nbytes

    def __del__(self):
        try:
            from . import _lib
            _lib.lib().b200rt_free_host(self.ptr)
        except Exception:
            pass


def pinned_array(shape, dtype=np.float64):
    """numba.cuda.pinned_array equivalent: a page-locked host ndarray (b200rt_alloc_host).  Passing one as `out=` of
    render_host / render_frame / render_batch makes the framebuffer's device-to-host copy a single DMA (reference
    src/main.py:51 `result.copy_to_host()`); allocate it once and reuse it across frames (pinning 100 MB costs tens of ms)."""
    import ctypes as C
    from . import _lib
    shape = (int(shape),) if np.isscalar(shape) else tuple(int(v) for v in shape)
    dt = np.dtype(dtype)
    nbytes = max(1, int(np.prod(shape)) * dt.itemsize)
    p = C.c_void_p()
    _lib.check(_lib.lib().b200rt_alloc_host(nbytes, C.byref(p)))
    owner = _PinnedOwner(p.value, nbytes)
    buf = (C.c_uint8 * nbytes).from_address(p.value)
    buf._owner = owner                                   # the ctypes buffer keeps the block alive; numpy keeps the buffer
    arr = np.frombuffer(buf, dtype=dt, count=int(np.prod(shape))).reshape(shape)
    arr[...] = 0
    return arr


def synchronize():
    _torch().cuda.synchronize()
