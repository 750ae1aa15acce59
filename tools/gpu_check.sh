#!/bin/bash
# The library's own index assertions (checked build -DB200RT_CHECK, built in the container as tools/ab/libb200rt_check.so by
# `python tools/build_checked.py`) over tools/sanitize_cases.py and the GPU parity tests; compute-sanitizer is closed on this pool.
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
cp python-ray-tracer_b200/b200rt/libb200rt.so /tmp/orig.so
cp tools/ab/libb200rt_check.so python-ray-tracer_b200/b200rt/libb200rt.so; touch python-ray-tracer_b200/b200rt/libb200rt.so
{ python tools/sanitize_cases.py; echo "sanitize_cases rc=$?"
  B200RT_REPORT_CHECKS=1 python -m pytest tests/test_gpu_parity.py tests/test_resident_scene.py tests/test_full_frame_parity.py -m gpu -x -q 2>&1 | tail -4; } > gpurun_out/r02_check.log 2>&1
cp /tmp/orig.so python-ray-tracer_b200/b200rt/libb200rt.so
cat gpurun_out/r02_check.log
