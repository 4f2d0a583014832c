#!/usr/bin/env python
"""Benchmark of the splat rendering hot path (BASELINE.json metric: frames/s and Mgaussians/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4|c5]

A step = one frame of the hot path (K1 preprocess .. K5 blend) over one synthetic scene.
N=1 default workload = BASELINE.json configs[1]: 1M gaussians, 1920x1080, TILE mode.
N>1 default: every rank renders the same scene from its own orbit cameras (views sharded,
no data-path collective, weak scaling); `--workload c3` is the 4K/6M single frame split into
tile-row bands with an NCCL gather of the framebuffer (strong scaling).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    #       n_gaussians  w     h     mode  seed_cfg  name
    "c1": (100_000, 1280, 720, 0, 1, "100K-gaussian synthetic scene, 1280x720, TILE"),
    "c2": (1_000_000, 1920, 1080, 0, 2, "1M-gaussian synthetic scene, 1920x1080, single frame, TILE"),
    "c3": (6_000_000, 3840, 2160, 0, 3, "6M-gaussian synthetic scene, 3840x2160, tile-row bands + NCCL gather"),
    "c4": (3_000_000, 1920, 1080, 0, 4, "3M-gaussian scene, orbit camera views at 1080p, views sharded"),
    "c5": (2_000_000, 1920, 1080, 1, 5, "2M-gaussian scene, 50-layer variant (K=50), 1080p"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback", 1965.0


def scene_cache(n, seed):
    """Synthetic scene, cached on disk under gpurun_out/ for repeated runs."""
    from gaussian_splat_b200 import scene
    inst, _ = scene.make_instances(n, seed)
    return inst


def orbit_camera(k, n_views, w, h):
    """Config-4 cameras: setPosition(5 sin t, 2, 5 cos t), target origin."""
    from gaussian_splat_b200 import camera
    t = 2.0 * math.pi * k / n_views
    return camera.app_camera(w, h, eye=(5.0 * math.sin(t), 2.0, 5.0 * math.cos(t)))


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake_slowdown",
        }
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def algorithmic_bytes(n, visible, D, n_pass, P):
    """SURVEY 8(d) per-stage algorithmic bytes of one frame."""
    return {
        "preprocess": n * (56 + 4) + visible * 48,
        "scan_duplicate": n * 8 + n * 4 + visible * 12 + D * 12,
        "sort": D * (8 + n_pass * 24),
        "ranges": D * 8,
        "blend": D * (4 + 48) + P * 8,
    }


def run_reference(args):
    """--impl reference: the CPU implementation of the path (Oracle-B, the C++ restatement of
    the reference shaders; the Metal reference itself cannot run on this box) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    from gaussian_splat_b200 import camera
    n, w, h, mode, cfg, name = WORKLOADS[args.workload or "c2"]
    cores = os.cpu_count() or 1
    O.set_threads(cores)
    # bounded sample: a scene of n/10 gaussians with the same generator at the same resolution
    # costs ~1/10 of the per-splat work; frames/s is reported for the sample and labelled.
    from gaussian_splat_b200 import scene
    frac = args.ref_fraction
    ns = max(1, int(n * frac))
    inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[cfg])
    inst = inst[:ns] if frac < 1.0 else inst
    V, P = camera.app_camera(w, h)
    times = []
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        F = O.Frame(inst, V, P, w, h, mode=mode)
        F.blend(n_threads=cores)
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    fps = 1e3 / ms
    sample = f"full frame of the first {ns} of {n} gaussians ({name}), {cores} threads"
    line = {
        "impl": "reference", "metric": "frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32+f16", "data": "synthetic",
        "mgaussians_per_s": ns * fps / 1e6,
        "config": {"workload": name, "n_gaussians": ns, "width": w, "height": h, "mode": "LAYER50" if mode else "TILE",
                   "reference_kind": "CPU restatement (oracle/oracle_b.cpp); Metal shaders unavailable on this box"},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def cpu_baseline_leg(inst, V, P, w, h, mode, name):
    """Oracle-B, single thread (the named baseline), one full frame of the same workload."""
    from oracle import oracle as O
    O.set_threads(1)
    t0 = time.perf_counter()
    F = O.Frame(inst, V, P, w, h, mode=mode)
    F.blend(n_threads=1)
    dt = time.perf_counter() - t0
    return {"value": 1.0 / dt, "unit": "frames/s", "cores": 1, "kind": "port",
            "sample": f"1 full frame, single thread, {name} ({dt:.2f} s)"}


def sort_bench(ctx, torch, n=1 << 26, reps=5):
    """Stand-alone 'sort HBM GB/s': full-entropy 64-bit keys, 8 passes, D*(8+8*24) bytes."""
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    hi = torch.randint(-2 ** 31, 2 ** 31, (n,), generator=g, dtype=torch.int64, device="cuda")
    lo = torch.randint(0, 2 ** 32, (n,), generator=g, dtype=torch.int64, device="cuda")
    src = (hi << 32) | lo
    del hi, lo
    k = torch.empty_like(src); ka = torch.empty_like(src)
    v = torch.arange(n, dtype=torch.int32, device="cuda"); va = torch.empty_like(v)
    s = torch.cuda.current_stream().cuda_stream
    ts = []
    for it in range(reps + 2):
        k.copy_(src)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.sort_pairs(s, k.data_ptr(), v.data_ptr(), ka.data_ptr(), va.data_ptr(), n, 0, 64)
        e1.record(); torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    by = n * (8 + 8 * 24)
    return {"n_pairs": n, "n_pass": 8, "ms": ms, "gb_per_s": by / ms / 1e6, "mkeys_per_s": n / ms / 1e3,
            "bytes": by, "note": "includes the histogram pass (the +8 B/pair)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--ref-fraction", type=float, default=1.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sort-bench", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from gaussian_splat_b200 import camera, capi, scene

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    wl = args.workload or "c2"
    n, w, h, mode, cfg, name = WORKLOADS[wl]
    bands = wl == "c3" and world > 1
    inst = scene_cache(n, scene.CONFIG_SEEDS[cfg])
    n = inst.shape[0]
    ctx = capi.Context(local)
    ctx.upload(inst)
    stream = torch.cuda.current_stream().cuda_stream
    tiles_y = (h + 15) // 16
    row0, row1 = 0, tiles_y
    if bands:
        row0, row1 = tiles_y * rank // world, tiles_y * (rank + 1) // world
    fb = torch.zeros((h, w, 4), dtype=torch.float16, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    n_views = 64
    # views: N=1 uses the app camera (configs 2/3/5); view-sharded runs walk the orbit
    def cam(step):
        if world > 1 and not bands or wl == "c4":
            return orbit_camera((step * world + rank) % n_views, n_views, w, h)
        return camera.app_camera(w, h)
    opts = capi.make_opts(mode=mode, tile_row_begin=row0 if bands else 0, tile_row_end=row1 if bands else 0,
                          flags=capi.FLAG_TIMING)
    gather_list = None
    if bands:
        rows_px = [(min(tiles_y * r // world * 16, h), min(tiles_y * (r + 1) // world * 16, h)) for r in range(world)]

    def step(i):
        V, P = cam(i)
        ctx.render(stream, fb.data_ptr(), V, P, w, h, opts)
        if bands:      # NCCL gather of the framebuffer bands to rank 0 (the path's one exchange step)
            a, b = rows_px[rank]
            if rank == 0:
                ops = [dist.P2POp(dist.irecv, fb[rows_px[r][0]:rows_px[r][1]], r) for r in range(1, world)
                       if rows_px[r][1] > rows_px[r][0]]
            else:
                ops = [dist.P2POp(dist.isend, fb[a:b], 0)] if b > a else []
            if ops:
                for wk in dist.batch_isend_irecv(ops):
                    wk.wait()

    # warm-up (also grows the pair buffers if the first frame overflowed)
    for i in range(args.warmup):
        step(i)
        torch.cuda.synchronize()
        st = ctx.status()
    assert not st.overflow, "pair buffers still overflow after warm-up"

    sampler = ClockSampler(local)
    stage_acc = {}
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        flush.fill_(i & 0xFF)                    # L2 flush between timed iterations (not timed)
        ev[i][0].record()
        step(args.warmup + i)
        ev[i][1].record()
        ev[i][1].synchronize()
        for k, v in ctx.stage_times().items():
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall0
    sampler.stop_flag = True
    sampler.join()
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    frames_per_step = 1 if bands else world
    fps = frames_per_step * 1e3 / ms_per_step
    st = ctx.status()
    D = int(st.n_pairs); n_pass = int(st.sort_passes)

    # e2e: the public host-buffer call (splat_render_host): uniforms from host memory, framebuffer
    # copied back to pinned host memory inside the timed region; scene resident as in the reference
    # (uploaded once by initialize(), src/instanced_splat_renderer.mm:399-422).
    host_fb = torch.empty((h, w, 4), dtype=torch.float16).pin_memory()
    eopts = capi.make_opts(mode=mode, tile_row_begin=row0 if bands else 0, tile_row_end=row1 if bands else 0)
    for i in range(3):
        ctx.render_host(host_fb.data_ptr(), *cam(i), w, h, eopts)
    if world > 1:
        dist.barrier()
    e2e_steps = max(5, min(args.steps, 50))
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        ctx.render_host(host_fb.data_ptr(), *cam(i), w, h, eopts)
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_fps = frames_per_step * e2e_steps / e2e_s
    band_px = (min(row1 * 16, h) - min(row0 * 16, h)) * w

    if rank == 0:
        hbm_peak, peak_kind, sm_max = peaks()
        visible = int((ctx.fetch(capi.FETCH_TILES_TOUCHED) > 0).sum())
        P = band_px if bands else w * h
        ab = algorithmic_bytes(n, visible, D, n_pass, P)
        stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
        stages = {}
        for k in ("preprocess", "scan_duplicate", "sort", "ranges", "blend"):
            ms = stage_ms.get(k, 0.0)
            gbs = ab[k] / ms / 1e6 if ms > 0 else 0.0
            stages[k] = {"ms": ms, "alg_bytes": ab[k], "gb_per_s": gbs, "frac_of_hbm_peak": gbs / hbm_peak}
        dom = max(("preprocess", "scan_duplicate", "sort", "ranges", "blend"), key=lambda k: stage_ms.get(k, 0.0))
        roof = {"kernel": dom, "bound": "hbm", "achieved": stages[dom]["gb_per_s"], "peak": hbm_peak, "unit": "GB/s",
                "frac": stages[dom]["frac_of_hbm_peak"], "traffic": None, "peak_kind": peak_kind,
                "stage_ms_total": stage_ms.get("total")}
        line = {
            "metric": "frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if bands else "weak", "vs_baseline": None, "dtype": "f32+f16", "data": "synthetic",
            "mgaussians_per_s": n * fps / 1e6,
            "config": {"workload": name, "n_gaussians": n, "width": w, "height": h,
                       "mode": "LAYER50" if mode else "TILE", "order": "near_first", "pairs_D": D, "visible": visible,
                       "sort_passes": n_pass, "parallelism": ("bands%d+nccl_gather" % world) if bands else ("views%d" % world),
                       "l2": "flushed between timed steps (256 MiB write, untimed)"},
            "roofline": roof, "stages": stages,
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": 128,
                    "d2h_bytes_per_step": band_px * 8 + 64, "steps": e2e_steps,
                    "note": "splat_render_host: matrices from host, framebuffer to pinned host; scene resident"},
            "gpu_launches": args.steps * (4 + n_pass),
            "clocks": sampler.result(), "wall_s": t_wall,
        }
        if world == 1 and not args.no_sort_bench:
            line["sort_bench"] = sort_bench(ctx, torch)
            line["sort_bench"]["frac_of_hbm_peak"] = line["sort_bench"]["gb_per_s"] / hbm_peak
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_leg(inst, *camera.app_camera(w, h), w, h, mode, name)
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
