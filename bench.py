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

STRIP = int(os.environ.get("B200RT_STRIP", "16"))    # columns per interleaved strip (multi-GPU load balance); multiple of 16


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


def workload_config(w, world, rays_total=None):
    """The `config` object shared by both arms."""
    cam = w["cameras"][0]
    W, H = cam.resolution
    sp, li, pl = w["scene"].generate_scene()
    cfg = {"workload": w["label"], "resolution": [W, H], "spheres": sp.shape[1], "planes": pl.shape[1], "lights": li.shape[1],
           "refl_depth": w["refl_depth"], "aliasing": bool(w["aliasing"]),
           "parallelism": f"column strips of {STRIP} interleaved over {world} GPU(s)" if world > 1 else "1 GPU, whole frame",
           "l2": "working set (ray queues, ~100 B/path x %.1f M paths) exceeds the 126 MB L2; no flush needed" % (W * H * (9 if w["aliasing"] else 1) / world / 1e6)}
    if rays_total is not None:
        cfg["rays_per_frame"] = rays_total
        cfg["rays_per_pixel"] = rays_total / (W * H)
    return cfg


def run_reference(args, w):
    """--impl reference: the CPU port of the reference path on all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count())          # torchrun pins it to 1; this arm is the all-host-threads CPU run
    per_step = max(2.0, min(20.0, 60.0 / max(1, args.steps + args.warmup)))
    res = None
    vals = []
    for i in range(args.warmup + args.steps):
        res = oracle_sample(w, seconds_target=per_step)
        if i >= args.warmup:
            vals.append(res)
    rays = sum(float(v["value"]) * v["seconds"] for v in vals)
    secs = sum(v["seconds"] for v in vals)
    value = rays / secs
    cam = w["cameras"][0]
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(1, len(vals)),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(w, args.gpus),
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": res["cores"], "kind": "port", "sample": res["sample"]},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_ours(args, w):
    import torch
    import torch.distributed as dist
    from b200rt import _lib, render_host, fp32_probe
    from b200rt.ray_tracing import render
    from b200rt.workloads import flops_per_ray

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    cams = w["cameras"]
    cam = cams[0]
    W, H = cam.resolution
    sp_h, li_h, pl_h = w["scene"].generate_scene()
    S, L, Pl = sp_h.shape[1], li_h.shape[1], pl_h.shape[1]
    spheres, lights, planes = [torch.from_numpy(a).to(dev) for a in (sp_h, li_h, pl_h)]
    result = torch.zeros((3, W, H), dtype=torch.uint8, device=dev)
    strips = (STRIP, rank, world) if world > 1 else None
    emu = int(os.environ.get("B200RT_EMULATE_WORLD", "0"))      # tuning aid: render rank 0's share of an N-GPU frame on one GPU
    if world == 1 and emu > 1:
        strips = (STRIP, int(os.environ.get("B200RT_EMULATE_RANK", "0")), emu)
    shade = (w["amb"], w["lamb"], w["refl"], w["refl_depth"], w["aliasing"])
    stream = torch.cuda.current_stream(dev)

    from b200rt.cuda import pixel_grid
    grids = [pixel_grid(c, local) for c in cams]               # closed-form pixel grid: no 24*W*H-byte array

    def frame(k=0, stats=False, flags=0):
        c = cams[k % len(cams)]
        ln = render.configure(stream=stream, strips=strips, stats=stats, flags=flags)
        ln(grids[k % len(cams)], result, c.position, c.rotation, spheres, lights, planes, *shade)
        return ln.stats

    # one instrumented frame: ray counts (deterministic) and per-kernel-class device time
    st = frame(0, stats=True)
    rays_rank = torch.tensor([st["primary"], st["shadow"], st["bounce"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(rays_rank)
    rays_total = float(rays_rank.sum().item())
    for i in range(args.warmup):
        frame(i)
    sampler = ClockSampler(local); sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib.lib().b200rt_launch_count()                # counted by the library at every <<< >>> (this rank)
    e0.record(stream)
    t_host = time.perf_counter()
    for i in range(args.steps):
        frame(i)
    host_enqueue_ms = (time.perf_counter() - t_host) * 1e3 / args.steps      # host time to enqueue one frame (no sync inside)
    e1.record(stream)
    launches = int(_lib.lib().b200rt_launch_count() - launches0)
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = float(ms.item()) / args.steps
    sampler.stop_flag = True; sampler.join(timeout=2)

    # kernel-class times of an instrumented frame after warm-up (dominant kernel = k_intersect, both variants)
    # (a) executed-work counters from the diagnostic kernel build; (b) times from the production kernels, launches serialised
    st2 = frame(0, stats=True)
    st3 = frame(0, stats=True, flags=_lib.F_TIME_ONLY)
    for k in ("ms_total", "ms_closest", "ms_shadow", "ms_other"):
        st2[k] = st3[k]
    isect_ms = st2["ms_closest"] + st2["ms_shadow"]
    rays_this_rank = st2["rays"]
    # (c) the brute-force mode (B200RT_F_DENSE: every specified query scans every sphere, nothing culled): the figure the
    # FP32-peak target of north_star is about.  Counters from one diagnostic frame, times from production-kernel frames.
    dense = None
    if rank == 0 and S >= 64 and not args.no_dense:
        sd = frame(0, stats=True, flags=_lib.F_DENSE)
        frame(0, flags=_lib.F_DENSE)
        torch.cuda.synchronize()
        d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_dense = max(1, min(args.steps, 3))
        d0.record(stream)
        for i in range(n_dense):
            frame(i, flags=_lib.F_DENSE)
        d1.record(stream)
        torch.cuda.synchronize()
        sdt = frame(0, stats=True, flags=_lib.F_DENSE | _lib.F_TIME_ONLY)
        dense = {"ms_per_frame": d0.elapsed_time(d1) / n_dense, "kernel_ms_per_frame": sdt["ms_closest"] + sdt["ms_shadow"],
                 "tests_pivot": sd["tests_pivot"], "tests_general": sd["tests_general"], "rays": sd["rays"]}

    # end to end: host buffers through b200rt_render_host, copies inside the timed region; N>1: + NCCL gather of strips
    host_out = torch.zeros((3, W, H), dtype=torch.uint8).pin_memory().numpy()      # pinned: the D2H of the framebuffer is a DMA
    h2d = sp_h.nbytes + li_h.nbytes + pl_h.nbytes
    from b200rt.sharding import columns_of_rank
    ncols_rank = len(columns_of_rank(W, rank, world, STRIP))
    d2h = 3 * ncols_rank * H

    def frame_e2e(k=0):
        c = cams[k % len(cams)]
        render_host(None, c.position, c.rotation, sp_h, li_h, pl_h, *shade, grid=c.grid(), shape=(W, H), out=host_out,
                    device=local, strips=strips)

    if len(cams) > 1:
        # camera-pose batch (C5): frames sharded by rank, one b200rt_render_batch_host call per rank (scene staged once,
        # D2H of frame k overlapped with frame k+1); per-frame time = batch time / frames in the batch
        from b200rt import render_batch
        from b200rt.sharding import frames_of_rank
        mine = [cams[k] for k in frames_of_rank(len(cams), rank, world)]
        batch_out = torch.zeros((len(mine), 3, W, H), dtype=torch.uint8).pin_memory().numpy()
        render_batch(w["scene"], mine[:2], *shade, devices=[local], out=batch_out[:2])
        barrier()
        t0 = time.perf_counter()
        render_batch(w["scene"], mine, *shade, devices=[local], out=batch_out)
        barrier()
        e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / len(cams)], dtype=torch.float64, device=dev)
        d2h = 3 * W * H * len(mine) // max(1, len(cams)) * 1
    else:
        frame_e2e(0)
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(1, min(args.steps, 5))
        for i in range(n_e2e):
            frame_e2e(i)
        barrier()
        e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / n_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_ms.item())

    if rank == 0:
        probe = fp32_probe(local)
        sms = _lib.lib().b200rt_sm_count(local)
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        nominal_tf = 2.0 * 128 * sms * sm_max * 1e6 / 1e12
        fpr = flops_per_ray(S, Pl)
        traffic, traffic_src = None, None          # ncu dram bytes of the k_isect3 launches of one frame (committed capture, 1 GPU, C4)
        tp = os.path.join(ROOT, "profiles", "r01_v12_traffic.json")
        if os.path.exists(tp) and args.workload == "c4" and world == 1:
            tj = json.load(open(tp))
            traffic, traffic_src = tj["traffic_bytes_per_frame"], "profiles/r01_v12_traffic.json (ncu dram__bytes_read+write, the 4 k_isect3 + 4 k_shadow_shade launches of one frame)"
        ach_tf = rays_this_rank * fpr / (isect_ms * 1e-3) / 1e12
        clocks = sampler.summary()
        line = {
            "metric": "Mrays/s", "value": rays_total / (ms_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(w, world, rays_total),
            "e2e": {"value": rays_total / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": launches,                         # kernels launched inside the timed region (rank 0; b200rt_launch_count)
            "clocks": clocks, "host_enqueue_ms_per_step": host_enqueue_ms,
            "roofline": {"bound": "fp32", "kernel": "k_isect3 (closest-hit queries) + k_shadow_shade (shadow queries through the light grids, with the Lambert sum)", "achieved": ach_tf, "peak": nominal_tf,
                         "unit": "TFLOP/s", "frac": ach_tf / nominal_tf, "traffic": traffic, "traffic_source": traffic_src,
                         "algorithmic_bytes_per_frame": int(rays_this_rank * 40),
                         "peak_def": f"2 FLOP x 128 lanes x {sms} SMs x {sm_max:.0f} MHz (nominal max clock; MEASURED_PEAKS has no FP32 entry)",
                         "peak_measured_ffma": probe["ffma_tflops"], "peak_measured_ffma2": probe["ffma2_tflops"],
                         "probe_sm_mhz": probe["sm_mhz"], "frac_of_measured_ffma": ach_tf / max(probe["ffma_tflops"], 1e-9),
                         "algorithmic_flops_per_ray": fpr, "kernel_ms_per_frame": isect_ms, "kernel_ms_closest": st2["ms_closest"], "kernel_ms_shadow": st2["ms_shadow"], "other_ms": st2["ms_other"], "kernel_share_of_frame": isect_ms / st2["ms_total"],
                         "slow_path_entries_per_ray": st2["slow_path"] / max(1, rays_this_rank), "exact_tests_per_ray": st2["exact_tests"] / max(1, rays_this_rank), "shadow_queries_traced_fraction": st2["shadow_traced"] / max(1, st2["shadow"]),
                         # work actually executed (the early-out shadow passes and the culled shadow queries skip tests):
                         "executed_tests_pivot_form": st2["tests_pivot"], "executed_tests_general_form": st2["tests_general"],
                         "executed_sphere_tests": st2["tests_pivot"] + st2["tests_general"], "canonical_sphere_tests": rays_this_rank * S,
                         "executed_tflops": (8.0 * st2["tests_pivot"] + 16.0 * st2["tests_general"]) / (isect_ms * 1e-3) / 1e12,
                         "executed_frac": (8.0 * st2["tests_pivot"] + 16.0 * st2["tests_general"]) / (isect_ms * 1e-3) / 1e12 / nominal_tf,
                         "frac_note": "frac = canonical brute-force FLOPs (16 per ray-sphere test, SURVEY 8d) / kernel time / peak; it exceeds the FMA-pipe utilisation because the kernel executes 8 FLOP per test in the pivot form and skips tests that cannot change the image (executed_frac = FLOPs actually issued / peak)",
                         "dense_mode": None if dense is None else {
                             "what": "B200RT_F_DENSE: brute force as the reference specifies it - every query scans every sphere (no beam culling, regrouping, shadow passes or lit-pair lists); same image",
                             "ms_per_frame": dense["ms_per_frame"], "mrays_per_s": dense["rays"] / (dense["ms_per_frame"] * 1e-3) / 1e6,
                             "kernel_ms_per_frame": dense["kernel_ms_per_frame"],
                             "algorithmic_tflops": dense["rays"] * fpr / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12,
                             "frac": dense["rays"] * fpr / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12 / nominal_tf,
                             "executed_tflops": (8.0 * dense["tests_pivot"] + 16.0 * dense["tests_general"]) / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12,
                             "executed_frac": (8.0 * dense["tests_pivot"] + 16.0 * dense["tests_general"]) / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12 / nominal_tf},
                         "framebuffer_write_gbs": 3.0 * W * H / world / (st2["ms_total"] * 1e-3) / 1e9,
                         "hbm_peak_gbs": peaks.get("hbm_gbs")},
        }
        if world == 1 and not args.no_cpu:
            os.environ.setdefault("OMP_NUM_THREADS", str(os.cpu_count()))
            cb = oracle_sample(w, seconds_target=12.0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-dense", action="store_true", help="skip the brute-force (B200RT_F_DENSE) measurement")
    args = ap.parse_args()
    from b200rt.workloads import workload
    w = workload(args.workload)
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
