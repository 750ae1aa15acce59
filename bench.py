#!/usr/bin/env python
"""bench.py — the headline benchmark of the render path (BASELINE.json metric: Mrays/s and ms/frame).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4] [--impl ours|reference]

A "step" is one full frame of the workload (default C4: 1024 random spheres, 7680x4320, 3 bounces, no AA —
the configuration BASELINE.json's target is quoted on; it fits one GPU).  Rank r of N renders the interleaved
column strips s with s % N == r; timing is CUDA events on the launch stream, barrier +
synchronize on both sides, max over ranks.  One JSON line on rank 0.

    value     whole-job Mrays/s with scene and camera resident in HBM (ray = one get_intersection query)
    e2e       the same metric through the host-buffer C-ABI call (b200rt_render_host): H2D of the scene arrays
              and D2H of the uint8 framebuffer inside the timed region
    roofline  FP32 CUDA-core roofline of the dominant kernel (k_intersect): algorithmic FLOPs = rays x (16 S + 14 Pl)
    cpu_baseline  the float64 oracle port (OpenMP, all host threads) on a bounded sample of the same frame

`--impl reference` times that CPU port alone (the reference itself is a numba-CUDA Python kernel whose only
CPU-executable form is the ~160 px/s simulator; it cannot travel to the GPU box — DESIGN.md).
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "python-ray-tracer_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

STRIP = 32           # columns per interleaved strip


def strips_for_rank(W, rank, world):
    return [(x0, min(W, x0 + STRIP)) for i, x0 in enumerate(range(0, W, STRIP)) if i % world == rank]


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons during the timed region (pynvml; falls back to nvidia-smi)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        self.power = []

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8: "hw_slowdown",
                     0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                        else nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for bit, nm in names.items():
                        if r & bit:
                            self.reasons.add(nm)
                except Exception:
                    pass
                time.sleep(0.05)
        except Exception as e:  # pragma: no cover
            self.reasons.add(f"sampler_error:{type(e).__name__}")

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "power_w_max": (max(self.power) if self.power else None), "samples": len(s)}


def oracle_sample(w, seconds_target=15.0, threads=0):
    """Time the float64 oracle (all host threads) on a bounded window of the workload's first frame.
    Returns dict(value Mrays/s, cores, sample, seconds)."""
    from oracle import oracle_f64 as O
    sp, li, pl = w["scene"].generate_scene()
    cam = w["cameras"][0]
    W, H = cam.resolution
    args = (cam.position, cam.rotation, sp, li, pl, w["amb"], w["lamb"], w["refl"], w["refl_depth"], w["aliasing"])
    x0, y0 = W // 2, H // 2

    def run(nx, ny):
        t = time.perf_counter()
        _, cnt = O.render(None, *args, grid=cam.grid(), shape=(W, H), x_range=(x0, x0 + nx), y_range=(y0, y0 + ny),
                          want_counts=True, threads=threads)
        return time.perf_counter() - t, int(cnt.sum())

    cores = O.max_threads() if threads == 0 else threads
    nx, ny = min(W - x0, 2 * cores), min(H - y0, 64)
    t, rays = run(nx, ny)                                         # calibration window
    scale = max(1.0, seconds_target / max(t, 1e-3))
    nx2 = int(min(W - x0, max(nx, nx * scale ** 0.5)))
    ny2 = int(min(H - y0, max(ny, ny * scale / max(1.0, nx2 / nx))))
    t, rays = run(nx2, ny2)
    return dict(value=rays / t / 1e6, unit="Mrays/s", cores=cores, kind="port",
                sample=f"{nx2}x{ny2} px window at ({x0},{y0}) of the {W}x{H} frame, {rays} rays in {t:.1f} s", seconds=t)
