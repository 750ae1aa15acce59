"""b200rt — B200-native render path behind the python-ray-tracer API.

    from b200rt import cuda                      # to_device / copy_to_host shim (was: from numba import cuda)
    from b200rt.ray_tracing import render        # render[grid, block](12 args)
    from b200rt.scene import Scene, Camera       # same API as the reference's `scene` package
"""
from . import scene
from .api import render_host, render_frame, render_batch, fp32_probe

__all__ = ["scene", "render_host", "render_frame", "render_batch", "fp32_probe"]
