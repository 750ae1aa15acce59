"""b200rt_to_image_host_noalloc: the allocation-free variant of b200rt_to_image_host (static __device__ scratch).

Collected last on purpose (a failure here must not hide other tests).  Called through the
C ABI; the output must equal the existing entry point's byte for byte (the tolerance of the viewer tests: exact),
on two calls in a row that share the static scratch."""
import ctypes as C
import numpy as np
import pytest

from b200rt import _lib


def _call(fn, x, flags=0):
    x = np.ascontiguousarray(x, np.uint8)
    _, W, H = x.shape
    out = np.full((H, W, 3), 0xA5, np.uint8)
    rc = fn(0, x.ctypes.data, out.ctypes.data, W, H, flags)
    _lib.check(rc)
    return out


@pytest.mark.gpu
def test_noalloc_to_image_matches_existing_entry_point_twice():
    L = _lib.lib()
    rng = np.random.default_rng(77)
    # second call smaller than the first: stale bytes of call 1 stay in the shared array and must not leak into call 2
    frames = [rng.integers(0, 256, (3, 1000, 1000), dtype=np.uint8), rng.integers(0, 256, (3, 257, 65), dtype=np.uint8)]
    for k, x in enumerate(frames):
        for flags in (0, _lib.F_UNSWAP_GB):
            exp = _call(L.b200rt_to_image_host, x, flags)
            got = _call(L.b200rt_to_image_host_noalloc, x, flags)
            assert np.array_equal(got, exp), (k, flags)
            ref = np.ascontiguousarray(x.transpose(2, 1, 0))
            assert np.array_equal(got, ref[:, :, [0, 2, 1]] if flags else ref), (k, flags)
    # the maximum size is accepted ...
    x = rng.integers(0, 256, (3, 1024, 1024), dtype=np.uint8)
    assert np.array_equal(_call(L.b200rt_to_image_host_noalloc, x), _call(L.b200rt_to_image_host, x))


@pytest.mark.gpu
def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    L = _lib.lib()
    x = np.zeros((3, 1025, 1024), np.uint8)
    out = np.zeros((1024, 1025, 3), np.uint8)
    rc = L.b200rt_to_image_host_noalloc(0, x.ctypes.data, out.ctypes.data, 1025, 1024, 0)
    assert rc == -3
    msg = L.b200rt_last_error().decode()
    assert "B200RT_NOALLOC_MAX_PIXELS" in msg and "1049600" in msg
    with pytest.raises(ValueError):
        _lib.check(rc)


def test_noalloc_argument_errors_without_gpu():
    """Size / pointer validation happens before any CUDA call."""
    L = _lib.lib()
    buf = np.zeros(16, np.uint8)
    assert L.b200rt_to_image_host_noalloc(0, None, buf.ctypes.data, 2, 2, 0) == -1
    assert L.b200rt_to_image_host_noalloc(0, buf.ctypes.data, buf.ctypes.data, 0, 2, 0) == -3
    assert L.b200rt_to_image_host_noalloc(0, buf.ctypes.data, buf.ctypes.data, 4096, 4096, 0) == -3
