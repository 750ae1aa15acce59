"""Presentation step with the reference's name and signature (src/viewer/image.py:7-19, src/viewer/__init__.py:1).

    convert_array_to_image(x) -> PIL.Image        x: uint8 (3, W, H) as render() leaves it

The array work (planar [c][x][y] -> interleaved RGB rows) runs in libb200rt (k_to_image); PIL only wraps the
finished buffer.  For the square frames the reference renders, `ImageOps.mirror(im.rotate(270))` is a transpose, and
that is what is produced; a non-square frame comes out whole, (H rows x W pixels), where the reference's `rotate`
(no expand=True) would crop it.  No CPU fallback: without the CUDA library this raises.
"""
import ctypes as C
import numpy as np
from .. import _lib


def to_image_array(x, unswap_gb=False, device=0):
    """uint8 (3,W,H) host array or b200rt DeviceArray / torch CUDA tensor -> uint8 (H, W, 3) host array."""
    L = _lib.lib()
    flags = _lib.F_UNSWAP_GB if unswap_gb else 0
    t = getattr(x, "tensor", x)
    if type(t).__module__.startswith("torch") and getattr(t, "is_cuda", False):
        import torch
        if t.dtype != torch.uint8 or t.dim() != 3 or t.shape[0] != 3:
            raise ValueError("expected a uint8 (3, W, H) array")
        t = t.contiguous()
        _, W, H = t.shape
        out = torch.empty((H, W, 3), dtype=torch.uint8, device=t.device)
        st = torch.cuda.current_stream(t.device)
        _lib.check(L.b200rt_to_image(t.device.index, C.c_void_p(st.cuda_stream), C.c_void_p(t.data_ptr()),
                                     C.c_void_p(out.data_ptr()), int(W), int(H), flags))
        return out.cpu().numpy()
    a = np.ascontiguousarray(np.asarray(x))
    if a.dtype != np.uint8:
        a = a.astype(np.uint8)                              # reference: y.astype(np.uint8), src/viewer/image.py:16
    if a.ndim != 3 or a.shape[0] != 3:
        raise ValueError(f"expected a (3, W, H) array, got {a.shape}")
    _, W, H = a.shape
    out = np.empty((H, W, 3), dtype=np.uint8)
    _lib.check(L.b200rt_to_image_host(int(device), a.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), int(W), int(H), flags))
    return out


def convert_array_to_image(x, unswap_gb=False):
    """Takes the (3,W,H) uint8 array and converts it to an RGB PIL image (reference signature)."""
    from PIL import Image
    return Image.fromarray(to_image_array(x, unswap_gb=unswap_gb), mode="RGB")
