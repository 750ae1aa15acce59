"""Camera of the scene API (reference src/scene/camera.py:7-26): same constructor, attributes and arrays."""
from typing import Tuple
import numpy as np
from .common import Vector3D
from .rotation import euler_rotation


class Camera:
    def __init__(self, resolution: Tuple[int, int], position: Vector3D, euler: Vector3D, fov: float = 45.):
        self.resolution = resolution
        self._position = position
        self.rotation = euler_rotation(euler[0], euler[1], euler[2])
        self.field_of_view = fov

    @property
    def position(self):
        return np.array(self._position)

    def grid(self):
        """(f, ar, sy, sz): closed form of generate_pixel_locations(), the b200rt_grid of include/b200rt.h.

        pixel_loc[0] = f, pixel_loc[1][x, :] = x*sy + ar, pixel_loc[2][:, y] = y*sz + 1 — numpy.mgrid with a
        complex step evaluates index*step + start with step = (stop-start)/(n-1) (reference camera.py:21-24).
        """
        width, height = self.resolution
        ar = int(width / height)                                   # camera.py:22 (sic: integer aspect ratio)
        f = 1 / np.tan(np.radians(self.field_of_view) / 2)
        # a one-point axis: mgrid leaves its step at 1 (`if step != 1: step = (stop - start) / (step - 1)`) and the only
        # index is 0, so the value is `start`; any finite step reproduces that (a division by W-1 = 0 would give NaN)
        sy = np.float64(-ar - ar) / np.float64(width - 1) if width > 1 else np.float64(1.0)
        sz = np.float64(-1 - 1) / np.float64(height - 1) if height > 1 else np.float64(1.0)
        return float(f), float(ar), float(sy), float(sz)

    def generate_pixel_locations(self):
        """float64 (3, W, H) pixel positions on the camera plane (reference camera.py:18-26)."""
        width, height = self.resolution
        f, ar, sy, sz = self.grid()
        out = np.empty((3, width, height), dtype=np.float64)
        out[0] = f
        out[1] = (np.arange(width, dtype=np.float64) * sy + ar)[:, None]
        out[2] = (np.arange(height, dtype=np.float64) * sz + 1.0)[None, :]
        from ..cuda import as_pixel_grid          # ndarray subclass that remembers (f, ar, sy, sz); read-only
        return as_pixel_grid(out, (f, ar, sy, sz))
