"""The C-ABI shared library loads and exports every symbol include/b200rt.h declares (no compute without a GPU)."""
import ctypes as C
import os
import re
import numpy as np
import pytest
from conftest import ROOT

from b200rt import _lib, _build


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200rt.h")).read()
    return sorted(set(re.findall(r"B200RT_API\s+[\w\s\*]+?\b(b200rt_\w+)\s*\(", src)))


def test_header_symbols_exported():
    _build.build()
    L = C.CDLL(_build.LIB_PATH)
    syms = _declared_symbols()
    assert len(syms) >= 9
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/b200rt.h but not exported"


def test_struct_layout_and_version():
    L = _lib.lib()
    assert L.b200rt_version() == 1
    assert L.b200rt_args_size() == C.sizeof(_lib.Args)


def test_argument_errors_without_gpu():
    """Argument validation happens before any CUDA call: same ValueError on a CPU box."""
    L = _lib.lib()
    assert L.b200rt_render(None) == -1
    a = _lib.Args()
    a.struct_size = 12
    assert L.b200rt_render(C.byref(a)) == -2
    a.struct_size = C.sizeof(_lib.Args)
    assert L.b200rt_render(C.byref(a)) == -1                       # result NULL
    buf = np.zeros(16, dtype=np.uint8)
    a.result = buf.ctypes.data
    assert L.b200rt_render(C.byref(a)) == -3                       # W,H = 0
    a.W, a.H, a.x_begin, a.x_end = 4, 4, 0, 5
    assert L.b200rt_render(C.byref(a)) == -3                       # tile outside the frame
    a.x_end = 4
    a.refl_depth = -1
    assert L.b200rt_render(C.byref(a)) == -4
    a.refl_depth = 1
    a.S = 1                                                        # spheres pointer missing
    assert L.b200rt_render(C.byref(a)) == -1
    with pytest.raises(ValueError):
        _lib.check(-3)
    assert b"" != L.b200rt_last_error()


def test_python_entry_validates_shapes():
    from b200rt import render_host
    z = np.zeros
    with pytest.raises(ValueError):
        render_host(z((2, 4, 4)), z(3), z((3, 3)), z((7, 1), np.float32), z((3, 1), np.float32), z((9, 1), np.float32), 0, 0.6, 0.3, 1, False)
    with pytest.raises(ValueError):
        render_host(z((3, 4, 4)), z(3), z((3, 3)), z((6, 1), np.float32), z((3, 1), np.float32), z((9, 1), np.float32), 0, 0.6, 0.3, 1, False)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libb200rt.so")
    with pytest.raises(ImportError):
        _lib.lib()


def test_flag_constants_match_header():
    """The B200RT_F_* values of include/b200rt.h and the constants the Python binding passes are the same numbers."""
    import re
    from b200rt import _lib
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "b200rt.h")).read()
    flags = {m.group(1): int(m.group(2), 16) for m in re.finditer(r"#define\s+B200RT_F_(\w+)\s+0x([0-9a-fA-F]+)u", hdr)}
    assert len(flags) >= 8 and len(set(flags.values())) == len(flags)          # distinct bits
    for name, val in flags.items():
        assert getattr(_lib, "F_" + name) == val, name
