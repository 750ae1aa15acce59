"""Build the checked variant of the library (-DB200RT_CHECK: index assertions, see csrc/rt_math.cuh) as tools/ab/libb200rt_check.so."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "python-ray-tracer_b200"))
from b200rt import _build
out = os.path.join(ROOT, "tools", "ab", "libb200rt_check.so")
os.makedirs(os.path.dirname(out), exist_ok=True)
subprocess.check_call(["nvcc"] + _build.NVCC_FLAGS + ["-DB200RT_CHECK", "--threads", "2", "-o", out] + [os.path.join(_build.SRC_DIR, f) for f in _build.SOURCES])
print(out)
