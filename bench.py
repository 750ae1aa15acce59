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
    roofline  two labelled objects: `brute_force` (B200RT_F_DENSE: every query scans every sphere; roofline.frac = its executed FP32 FLOPs / kernel time /
              FP32 peak) and `exact_culled` (the shipped default: per-class times, executed fraction, DRAM bytes, Mrays/s as specified and as traced)
    cpu_baseline  the float64 oracle port (OpenMP, all host threads) on a fixed, representative sample of the same frame (every n-th 16-column strip)

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


def sample_stride(w):
    """Every `stride`-th interleaved 16-column strip of the frame is the CPU arm's sample: the same ray mix as the whole frame
    (strips are spread evenly over it), fixed for a workload, sized to ~2 s of the oracle on 16 threads per pass."""
    cam = w["cameras"][0]
    W, H = cam.resolution
    sp, _, pl = w["scene"].generate_scene()
    cost = float(W) * H * (sp.shape[1] + pl.shape[1] + 1) * (9 if w["aliasing"] else 1)
    stride = 1
    while cost / stride > 6e8 and stride < 256:
        stride *= 2
    return stride


def oracle_sample(w, stride=None, rank=0, passes=1, threads=0):
    """Time the float64 oracle (all host threads) on strips s % stride == rank of the workload's first frame, `passes` times.
    Returns dict(value Mrays/s, cores, sample, seconds, rays)."""
    from oracle import oracle_f64 as O
    sp, li, pl = w["scene"].generate_scene()
    cam = w["cameras"][0]
    W, H = cam.resolution
    stride = sample_stride(w) if stride is None else stride
    args = (cam.position, cam.rotation, sp, li, pl, w["amb"], w["lamb"], w["refl"], w["refl_depth"], w["aliasing"])
    cores = O.max_threads() if threads == 0 else threads
    rays, ncols = 0, 0
    t0 = time.perf_counter()
    for _ in range(passes):
        for s_, x0 in enumerate(range(0, W, STRIP)):
            if s_ % stride != rank % stride:
                continue
            _, cnt = O.render(None, *args, grid=cam.grid(), shape=(W, H), x_range=(x0, min(W, x0 + STRIP)), want_counts=True, threads=threads)
            rays += int(cnt.sum()); ncols += min(W, x0 + STRIP) - x0
    t = time.perf_counter() - t0
    return dict(value=rays / t / 1e6, unit="Mrays/s", cores=cores, kind="port", rays=rays, seconds=t,
                sample=f"every {stride}th interleaved {STRIP}-column strip of the {W}x{H} frame ({ncols // passes} columns = 1/{stride} of the frame, "
                       f"same ray mix as the frame) x {passes} pass(es): {rays} rays in {t:.1f} s")


def workload_config(w, world):
    """The `config` object: identical in both arms (`--impl ours` / `--impl reference`)."""
    cam = w["cameras"][0]
    W, H = cam.resolution
    sp, li, pl = w["scene"].generate_scene()
    cfg = {"workload": w["label"], "resolution": [W, H], "spheres": sp.shape[1], "planes": pl.shape[1], "lights": li.shape[1],
           "refl_depth": w["refl_depth"], "aliasing": bool(w["aliasing"]),
           "parallelism": ("1 GPU, whole frame" if world == 1 else f"frames of the batch sharded over {world} GPU(s), frame k on rank k % {world}" if len(w["cameras"]) > 1
                           else f"column strips of {STRIP} interleaved over {world} GPU(s)"),
           "l2": "working set (ray queues, ~100 B/path x %.1f M paths) exceeds the 126 MB L2; no flush needed" % (W * H * (9 if w["aliasing"] else 1) / world / 1e6)}
    return cfg


def run_reference(args, w):
    """--impl reference: the CPU port of the reference path on all host threads.  One step = one pass over the fixed sample
    of the frame (sample_stride: every 64th 16-column strip at C4 = 1/64 of the frame, same ray mix), same `config`."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count())          # torchrun pins it to 1; this arm is the all-host-threads CPU run
    res = None
    vals = []
    for i in range(args.warmup + args.steps):
        res = oracle_sample(w)
        if i >= args.warmup:
            vals.append(res)
    rays = sum(v["rays"] for v in vals)
    secs = sum(v["seconds"] for v in vals)
    value = rays / secs / 1e6
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
    batch = len(cams) > 1                                      # camera-pose batch (C5): frames are sharded, frame k on rank k % N, each whole
    strips = (STRIP, rank, world) if (world > 1 and not batch) else None
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

    # instrumented frames: ray counts (deterministic) and per-kernel-class device time.  Single-frame workloads: `value` is rays of
    # THE frame / time per frame (all ranks render their strips of every step's frame).  Batch: step k = frame k of the orbit on
    # rank k % N; rays are the mean per frame over the K timed frames.
    my_steps = [k for k in range(args.steps) if (k % world == rank or not batch)]
    if batch:
        rr = np.zeros(3)
        for k in my_steps:
            st = frame(k, stats=True)
            rr += [st["primary"], st["shadow"], st["bounce"]]
        rays_rank = torch.tensor(rr / max(1, args.steps), dtype=torch.float64, device=dev)
    else:
        st = frame(0, stats=True)
        rays_rank = torch.tensor([st["primary"], st["shadow"], st["bounce"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(rays_rank)
    rays_total = float(rays_rank.sum().item())                  # per step
    for i in range(args.warmup):
        frame(i)
    sampler = ClockSampler(local); sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib.lib().b200rt_launch_count()                # counted by the library at every <<< >>> (this rank)
    e0.record(stream)
    t_host = time.perf_counter()
    for i in my_steps:
        frame(i)
    host_enqueue_ms = (time.perf_counter() - t_host) * 1e3 / max(1, len(my_steps))      # host time to enqueue one frame (no sync inside)
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
    # FP32-peak target of north_star is about, and the bench line's roofline.frac.  >= 10 timed frames; executed-work counters from
    # one diagnostic frame, kernel-class times from a production-kernel frame (launches serialised, event pairs per class).
    dense = None
    if rank == 0 and S >= 64 and not args.no_dense:
        sd = frame(0, stats=True, flags=_lib.F_DENSE)
        for i in range(2):
            frame(i, flags=_lib.F_DENSE)
        torch.cuda.synchronize()
        d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_dense = max(10, args.steps)
        d0.record(stream)
        for i in range(n_dense):
            frame(i, flags=_lib.F_DENSE)
        d1.record(stream)
        torch.cuda.synchronize()
        kms = []
        for i in range(3):
            sdt = frame(0, stats=True, flags=_lib.F_DENSE | _lib.F_TIME_ONLY)
            kms.append(sdt["ms_closest"] + sdt["ms_shadow"])
        dense = {"ms_per_frame": d0.elapsed_time(d1) / n_dense, "frames_timed": n_dense, "kernel_ms_per_frame": sorted(kms)[1],
                 "tests_pivot": sd["tests_pivot"], "tests_general": sd["tests_general"], "rays": sd["rays"]}

    # ---- end to end: host buffers through b200rt_render_host, H2D of the scene and D2H of the framebuffer inside the timed region.
    # N > 1: every rank copies its strips straight into ONE shared, pinned host frame (sharding.SharedFrame: the direct
    # device-to-host gather of the tiles, north_star / SURVEY §8(e)); after the closing barrier rank 0 holds the assembled
    # (3,W,H) frame — the `result.copy_to_host()` of reference src/main.py:51 — and its checksum is compared with a 1-GPU render.
    from b200rt.cuda import pinned_array
    from b200rt.sharding import columns_of_rank, SharedFrame
    import hashlib
    shared = None
    if world > 1 and not batch:
        shm_name = "b200rt_bench_%s" % os.environ.get("MASTER_PORT", "0")
        if rank == 0:
            shared = SharedFrame(shm_name, W, H, create=True)
        barrier()                                                  # the file exists before anyone else maps it
        if rank != 0:
            shared = SharedFrame(shm_name, W, H, create=False)
        host_out = shared.array
    elif batch:
        host_out = None
    else:
        host_out = pinned_array((3, W, H), np.uint8)               # pinned: the D2H of the framebuffer is one DMA
    h2d = sp_h.nbytes + li_h.nbytes + pl_h.nbytes
    ncols_rank = len(columns_of_rank(W, rank, world, STRIP)) if strips is None or world > 1 else len(columns_of_rank(W, strips[1], strips[2], STRIP))
    d2h = 3 * ncols_rank * H

    def frame_e2e(k=0):
        c = cams[k % len(cams)]
        render_host(None, c.position, c.rotation, sp_h, li_h, pl_h, *shade, grid=c.grid(), shape=(W, H), out=host_out,
                    device=local, strips=strips)

    assembled = None
    if batch:
        # camera-pose batch (C5): frames sharded by rank (frame k on rank k % N), one b200rt_render_batch_host call per rank (scene
        # staged once, D2H of frame k overlapped with frame k+1); per-frame time = batch time / frames of the whole batch
        from b200rt import render_batch
        from b200rt.sharding import frames_of_rank
        mine = [cams[k] for k in frames_of_rank(len(cams), rank, world)]
        batch_out = pinned_array((len(mine), 3, W, H), np.uint8)
        render_batch(w["scene"], mine[:2], *shade, devices=[local], out=batch_out[:2])
        barrier()
        t0 = time.perf_counter()
        render_batch(w["scene"], mine, *shade, devices=[local], out=batch_out)
        barrier()
        e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / len(cams)], dtype=torch.float64, device=dev)
        d2h = 3 * W * H * len(mine) // max(1, len(cams))
        # device-timed batch through the same sharding: the C5 `value` (frames / s over all ranks)
    else:
        host_out[...] = 0
        frame_e2e(0)
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(1, min(args.steps, 10))
        for i in range(n_e2e):
            frame_e2e(i)
        barrier()
        e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / n_e2e], dtype=torch.float64, device=dev)
        if (rank == 0 and strips is None) or world > 1:
            # the assembled frame (all ranks' strips, in rank 0's mapping) against a whole-frame render of this process's GPU
            if rank == 0:
                sha_e2e = hashlib.sha256(np.ascontiguousarray(host_out).tobytes()).hexdigest()
                whole = render_host(None, cam.position, cam.rotation, sp_h, li_h, pl_h, *shade, grid=cam.grid(), shape=(W, H), device=local)
                sha_one = hashlib.sha256(whole.tobytes()).hexdigest()
                assembled = {"assembled": True, "how": ("every rank's b200rt_render_host copies its strips into one shared pinned host mapping (/dev/shm + b200rt_host_register); rank 0 holds the (3,W,H) frame after the barrier"
                                                        if world > 1 else "one b200rt_render_host call into a pinned (3,W,H) host array"),
                             "frame_sha256": sha_e2e, "one_gpu_frame_sha256": sha_one, "checksum_matches_one_gpu_frame": sha_e2e == sha_one}
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_ms.item())
    if shared is not None:
        barrier()
        shared.close()

    if rank == 0:
        probe = fp32_probe(local)
        sms = _lib.lib().b200rt_sm_count(local)
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        hbm_peak = float(peaks.get("hbm_gbs", 6551.7))
        nominal_tf = 2.0 * 128 * sms * sm_max * 1e6 / 1e12
        fpr = flops_per_ray(S, Pl)
        # ncu DRAM bytes per kernel class of one frame (committed capture of the same command, 1 GPU, C4): profiles/r02_traffic.json
        traffic = None
        for tname in ("r02_traffic.json", "r01_v12_traffic.json"):
            tp = os.path.join(ROOT, "profiles", tname)
            if os.path.exists(tp) and args.workload == "c4" and world == 1 and strips is None:
                traffic = dict(json.load(open(tp)), source="profiles/" + tname)
                break
        exec_flops = 8.0 * st2["tests_pivot"] + 16.0 * st2["tests_general"]
        rays_traced = st2["primary"] + st2["bounce"] + st2["shadow_traced"]
        clocks = sampler.summary()
        culled = {
            "what": "the shipped default path: exact culling (beam / cluster culling of the closest-hit scans, per-light direction grids for the shadow queries, shadow queries of unlit pairs skipped); image byte-identical to brute force",
            "bound": "issue / latency + HBM (ncu: k_isect3 bounce launches 70-77 % issue-active, k_shadow_shade long-scoreboard bound, k_setup at 74-80 % of HBM peak; profiles/r02_summary.md) - NOT the FP32 pipe",
            "ms_per_frame": ms_step, "mrays_per_s_as_specified": rays_total / (ms_step * 1e-3) / 1e6,
            "mrays_per_s_as_traced": (rays_traced * (rays_total / max(1, rays_this_rank))) / (ms_step * 1e-3) / 1e6,
            "rays_traced_fraction": rays_traced / max(1, rays_this_rank),
            "shadow_queries_traced_fraction": st2["shadow_traced"] / max(1, st2["shadow"]),
            "kernel_ms": {"closest_hit": st2["ms_closest"], "shadow_shade": st2["ms_shadow"], "other": st2["ms_other"], "total_serialised": st2["ms_total"]},
            "executed_sphere_tests": st2["tests_pivot"] + st2["tests_general"], "canonical_sphere_tests": rays_this_rank * S,
            "executed_tests_pivot_form": st2["tests_pivot"], "executed_tests_general_form": st2["tests_general"],
            "executed_tflops": exec_flops / (isect_ms * 1e-3) / 1e12, "executed_frac": exec_flops / (isect_ms * 1e-3) / 1e12 / nominal_tf,
            "canonical_tflops_not_a_utilisation": rays_this_rank * fpr / (isect_ms * 1e-3) / 1e12,
            "slow_path_entries_per_ray": st2["slow_path"] / max(1, rays_this_rank), "exact_tests_per_ray": st2["exact_tests"] / max(1, rays_this_rank),
        }
        if traffic is not None and "dram_bytes_per_class" in traffic:
            tb = traffic["dram_bytes_per_class"]
            culled["dram_bytes_per_frame"] = tb
            culled["hbm_frac"] = {"closest_hit": tb.get("closest_hit", 0) / (st2["ms_closest"] * 1e-3) / 1e9 / hbm_peak,
                                  "shadow_shade": tb.get("shadow_shade", 0) / (st2["ms_shadow"] * 1e-3) / 1e9 / hbm_peak,
                                  "other": tb.get("other", 0) / (st2["ms_other"] * 1e-3) / 1e9 / hbm_peak,
                                  "frame": sum(tb.values()) / (st2["ms_total"] * 1e-3) / 1e9 / hbm_peak}
        brute = None
        if dense is not None:
            dflops = 8.0 * dense["tests_pivot"] + 16.0 * dense["tests_general"]
            brute = {"what": "B200RT_F_DENSE: brute force as the reference specifies it - every query (shadow queries of unlit pairs included) scans every sphere; no beam culling, regrouping, passes, lists or grids; same image (tested)",
                     "bound": "fp32", "frames_timed": dense["frames_timed"], "ms_per_frame": dense["ms_per_frame"],
                     "mrays_per_s": dense["rays"] / (dense["ms_per_frame"] * 1e-3) / 1e6, "kernel_ms_per_frame": dense["kernel_ms_per_frame"],
                     "executed_sphere_tests": dense["tests_pivot"] + dense["tests_general"], "canonical_sphere_tests": dense["rays"] * S,
                     "executed_tflops": dflops / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12,
                     "executed_frac": dflops / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12 / nominal_tf,
                     "executed_frac_of_measured_ffma2": dflops / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12 / max(probe["ffma2_tflops"], 1e-9),
                     "canonical_tflops": dense["rays"] * fpr / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12,
                     "canonical_frac": dense["rays"] * fpr / (dense["kernel_ms_per_frame"] * 1e-3) / 1e12 / nominal_tf}
        # roofline (headline object): the FP32 roofline of the dominant kernel k_isect3, measured in the brute-force mode — the only
        # mode in which "fraction of FP32 peak" is a utilisation: achieved = FLOPs the kernel executes (its own counters: 8 per test in
        # the pivot form, 16 in the origin-free form) / its device time, measured live here.  The shipped path's numbers are `exact_culled`.
        if brute is not None:
            roof = {"bound": "fp32", "mode": "brute_force", "kernel": "k_isect3 (closest-hit + shadow queries, dense scan of every sphere)",
                    "achieved": brute["executed_tflops"], "peak": nominal_tf, "unit": "TFLOP/s", "frac": brute["executed_frac"]}
        else:
            roof = {"bound": "issue/latency", "mode": "exact_culled", "kernel": "k_isect3 + k_shadow_shade", "achieved": culled["executed_tflops"],
                    "peak": nominal_tf, "unit": "TFLOP/s", "frac": culled["executed_frac"]}
        roof.update({"traffic": None if traffic is None else traffic.get("traffic_bytes_per_frame"), "traffic_source": None if traffic is None else traffic.get("source"),
                     "algorithmic_flops_per_ray": fpr, "algorithmic_bytes_per_frame": int(3 * W * H / world),
                     "peak_def": f"2 FLOP x 128 lanes x {sms} SMs x {sm_max:.0f} MHz (nominal max clock; MEASURED_PEAKS has no FP32 entry)",
                     "peak_measured_ffma": probe["ffma_tflops"], "peak_measured_ffma2": probe["ffma2_tflops"], "probe_sm_mhz": probe["sm_mhz"],
                     "brute_force": brute, "exact_culled": culled,
                     "framebuffer_write_gbs": 3.0 * W * H / world / (ms_step * 1e-3) / 1e9, "hbm_peak_gbs": hbm_peak})
        line = {
            "metric": "Mrays/s", "value": rays_total / (ms_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(w, world),
            "rays_per_frame": rays_total, "rays_per_pixel": rays_total / (W * H),
            "rays_note": "rays as the reference specifies them (one per get_intersection query, trace.py:53/:92); %.1f %% of them are shadow queries whose answer the reference discards (lamb*(L.N) <= 0) and which the shipped path does not trace: see roofline.exact_culled.mrays_per_s_as_traced" % (100.0 * (1.0 - rays_traced / max(1, rays_this_rank))),
            "e2e": dict({"value": rays_total / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms,
                         "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}, **(assembled or {})),
            "gpu_launches": launches,                         # kernels launched inside the timed region (rank 0; b200rt_launch_count)
            "clocks": clocks, "host_enqueue_ms_per_step": host_enqueue_ms,
            "roofline": roof,
        }
        if world == 1 and not args.no_cpu:
            os.environ.setdefault("OMP_NUM_THREADS", str(os.cpu_count()))
            cb = oracle_sample(w, stride=max(1, sample_stride(w) // 8))
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
