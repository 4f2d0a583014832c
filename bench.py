#!/usr/bin/env python
"""Benchmark of the splat rendering hot path (BASELINE.json metric: frames/s and Mgaussians/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4|c5]

A step = one frame of the hot path (K1 preprocess .. K5 blend) over one synthetic scene.
N=1 default workload = BASELINE.json configs[1]: 1M gaussians, 1920x1080, TILE mode, single frame.
N>1 default: every rank renders the same scene from its own orbit cameras (views sharded, no
data-path collective, weak scaling: value = frames of all ranks / max-over-ranks time);
`--workload c3` is the 4K/6M single frame split into tile-row bands with an NCCL gather of the
framebuffer (strong scaling). Prints ONE JSON line on rank 0.
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
STAGES = ("preprocess", "scan_duplicate", "sort", "ranges", "blend")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def measured_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full`
    capture of this workload (profiles/traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(workload, {}).get(kernel)
    except Exception:
        return None


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
        names = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
                 0x80: "hw_power_brake_slowdown"}
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
            time.sleep(0.004)

    def result(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def algorithmic_bytes(n, visible, D, C, P, sh_degree=0, layer=False):
    """Per-stage algorithmic bytes of one frame (SURVEY 8(d) adapted to the per-tile binning).
    n splats, V visible, C candidate (splat, tile-of-rectangle) slots, D pairs that pass the tile test, P pixels."""
    sh = {0: 0, 1: 36, 2: 96, 3: 180}[sh_degree]
    return {
        "preprocess": n * (56 + 4) + visible * (48 + sh),
        # bin words in, the visible splats' record rows for the tile test, candidates parked (key + tile|slot),
        # parked pairs read back, pairs scattered
        "scan_duplicate": n * 8 + visible * 64 + C * 16 + C * 16 + D * 8,
        "sort": D * 16 + (D * 4 if layer else 0),                   # every tile list read and written once (+ LAYER50's arrival list)
        "ranges": 0,
        "blend": D * (8 + 48) + (D * (4 + 48) if layer else 0) + P * 8,
    }


_REAL_STDOUT = None


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def _cpu_frame(O, inst, V, P, w, h, mode, threads, rows=None):
    """One frame of the CPU restatement (all stages + blend); rows = (begin, end) tile-row band or None."""
    kw = {} if rows is None else {"tile_row_begin": rows[0], "tile_row_end": rows[1]}
    t0 = time.perf_counter()
    F = O.Frame(inst, V, P, w, h, mode=mode, **kw)
    F.blend(n_threads=threads)
    return time.perf_counter() - t0


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on all host cores. The reference's
    own GPU path (Metal shaders) is unavailable on this box, so this is Oracle-B, the C++ restatement
    of those shaders that tests pin against the shader text (kind "port"). Each step is one full frame of
    the same workload when the whole run fits ~2 minutes; otherwise a centred band of tile rows covering
    a stated fraction of the frame (all splats are still preprocessed), reported as measured on the
    sample: frames/s = fraction / time."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    from gaussian_splat_b200 import camera, scene
    n, w, h, mode, cfg, name = WORKLOADS[args.workload or "c2"]
    cores = os.cpu_count() or 1
    O.set_threads(cores)
    inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[cfg])
    V, P = camera.app_camera(w, h)
    tiles_y = (h + 15) // 16
    t_full = _cpu_frame(O, inst, V, P, w, h, mode, cores)            # also the first warm-up
    budget = 120.0 / max(1, args.steps + args.warmup)
    frac, rows = 1.0, None
    if t_full > budget:
        nrows = max(1, min(tiles_y, int(tiles_y * budget / t_full)))
        b = (tiles_y - nrows) // 2
        rows, frac = (b, b + nrows), nrows / tiles_y
    times = []
    for it in range(args.warmup + args.steps):
        dt = _cpu_frame(O, inst, V, P, w, h, mode, cores, rows)
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    fps = frac * 1e3 / ms
    sample = (f"full frames, {cores} threads" if rows is None else
              f"tile rows [{rows[0]},{rows[1]}) of {tiles_y} ({frac:.3f} of the frame; all {n} splats preprocessed), "
              f"{cores} threads; frames/s = fraction / time") + f"; one full frame took {t_full:.2f} s"
    line = {
        "impl": "reference", "metric": "frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32+f16", "data": "synthetic",
        "mgaussians_per_s": n * fps / 1e6,
        "config": {"workload": name, "n_gaussians": n, "width": w, "height": h, "mode": "LAYER50" if mode else "TILE",
                   "order": "near_first",
                   "reference_kind": "CPU restatement of the reference shaders (oracle/oracle_b.cpp); the Metal path is unavailable on this box"},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def cpu_baseline_leg(inst, V, P, w, h, mode, name):
    """Oracle-B, single thread (the named baseline): one full frame of the same workload if it takes
    under ~25 s, else a centred band of tile rows sized for ~15 s."""
    from oracle import oracle as O
    O.set_threads(1)
    tiles_y = (h + 15) // 16
    probe_rows = (tiles_y // 2, tiles_y // 2 + 1)
    t_row = _cpu_frame(O, inst, V, P, w, h, mode, 1, probe_rows)       # K1 over all splats + one tile row
    t_pre = _cpu_frame(O, inst, V, P, w, h, mode, 1, (tiles_y - 1, tiles_y)) if tiles_y > 1 else 0.0
    est_full = t_pre + max(0.0, t_row - t_pre) * tiles_y * 0.6
    if est_full <= 25.0:
        dt = _cpu_frame(O, inst, V, P, w, h, mode, 1)
        return {"value": 1.0 / dt, "unit": "frames/s", "cores": 1, "kind": "port",
                "sample": f"1 full frame, single thread, {name} ({dt:.2f} s)"}
    nrows = max(1, min(tiles_y, int(tiles_y * 15.0 / est_full)))
    b = (tiles_y - nrows) // 2
    dt = _cpu_frame(O, inst, V, P, w, h, mode, 1, (b, b + nrows))
    frac = nrows / tiles_y
    return {"value": frac / dt, "unit": "frames/s", "cores": 1, "kind": "port",
            "sample": f"tile rows [{b},{b + nrows}) of {tiles_y} ({frac:.3f} of the frame, all splats preprocessed), "
                      f"single thread, {dt:.2f} s; frames/s = fraction / time; {name}"}


def sort_bench(ctx, torch, n=1 << 26, reps=5):
    """Stand-alone 'sort HBM GB/s' of the 64-bit entry point: full-entropy keys, 8 passes, D*(8+8*24) bytes."""
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
            "bytes": by, "note": "8 onesweep passes + the histogram pass (the +8 B/pair); inputs larger than L2"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--sh-degree", type=int, default=0, choices=[0, 1, 2, 3])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sort-bench", action="store_true")
    ap.add_argument("--even-bands", action="store_true", help="c3: equal tile-row bands instead of pair-balanced ones")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    # stdout carries exactly ONE line (the JSON): libraries that print there (NCCL's version banner,
    # the reference loader's progress) are sent to stderr for the duration of the run.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from gaussian_splat_b200 import camera, capi, multigpu, scene

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = multigpu.bind_to_gpu_numa_node(local) if world > 1 else None      # host buffers land next to this rank's GPU
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    wl = args.workload or "c2"
    n, w, h, mode, cfg, name = WORKLOADS[wl]
    bands = wl == "c3" and world > 1
    inst, sh = scene.make_instances(n, scene.CONFIG_SEEDS[cfg], sh_rest=args.sh_degree > 0)
    n = inst.shape[0]
    ctx = capi.Context(local)
    ctx.upload(inst)
    if args.sh_degree:
        ctx.upload_sh(sh, 3)
    stream = torch.cuda.current_stream().cuda_stream
    tiles_x, tiles_y = (w + 15) // 16, (h + 15) // 16
    fb = torch.zeros((h, w, 4), dtype=torch.float16, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    n_views = 64
    view_sharded = (world > 1 and not bands) or wl == "c4"

    def cam(step):      # N=1 uses the app camera (configs 2/3/5); view-sharded runs walk the orbit
        if view_sharded:
            return orbit_camera((step * world + rank) % n_views, n_views, w, h)
        return camera.app_camera(w, h)

    # ---- tile-row bands (c3, N>1): boundaries balanced on the pairs per tile row of a first full frame
    row0, row1, band_list = 0, tiles_y, [(0, tiles_y)]
    if bands:
        ctx.render(stream, fb.data_ptr(), *cam(0), w, h, capi.make_opts(mode=mode))
        torch.cuda.synchronize()
        if ctx.status().overflow:
            ctx.render(stream, fb.data_ptr(), *cam(0), w, h, capi.make_opts(mode=mode)); torch.cuda.synchronize(); ctx.status()
        r = ctx.fetch(capi.FETCH_TILE_RANGES).astype(np.int64)
        row_pairs = (r[:, 1] - r[:, 0]).reshape(tiles_y, tiles_x).sum(axis=1)
        band_list = multigpu.band_rows(tiles_y, world, None if args.even_bands else row_pairs)
        if not args.even_bands:
            # Measured-time balancing (untimed set-up, like the pair count above): pairs per row are not the cost
            # (rows of large near splats blend slower per pair), so three rounds of "render the band, exchange
            # the band-proportional stage times, move the boundaries" follow.
            wts = row_pairs.astype(np.float64)
            for _ in range(3):
                a, b = band_list[rank]
                o = capi.make_opts(mode=mode, sh_degree=args.sh_degree, tile_row_begin=a, tile_row_end=b, flags=capi.FLAG_TIMING)
                acc = []
                for it in range(6):
                    var = 0.0
                    if b > a:
                        ctx.render(stream, fb.data_ptr(), *cam(0), w, h, o); torch.cuda.synchronize()
                        st_ = ctx.stage_times()
                        var = st_["sort"] + st_["ranges"] + st_["blend"] + 0.5 * st_["scan_duplicate"]
                    if it >= 2:
                        acc.append(var)
                tt = torch.tensor([float(np.median(acc))], dtype=torch.float64, device=dev)
                allt = [torch.zeros_like(tt) for _ in range(world)]
                dist.all_gather(allt, tt)
                wts = multigpu.reweight_rows(wts, band_list, [float(x.item()) for x in allt])
                band_list = multigpu.band_rows(tiles_y, world, wts)
        row0, row1 = band_list[rank]
    rows_px = multigpu.bands_to_pixel_rows(band_list, h) if bands else [(0, h)]
    have_rows = row1 > row0

    def opts_for(flags):
        return capi.make_opts(mode=mode, sh_degree=args.sh_degree, tile_row_begin=row0 if bands else 0,
                              tile_row_end=row1 if bands else 0, flags=flags)
    opts = opts_for(capi.FLAG_TIMING)

    def step(i):
        if have_rows:
            ctx.render(stream, fb.data_ptr(), *cam(i), w, h, opts)
        if bands:      # NCCL gather of the framebuffer bands to rank 0 (the path's one exchange step)
            multigpu.gather_bands(fb, rows_px, rank, world, dst=0)

    # warm-up (also grows the pair buffers if the first frame overflowed)
    for i in range(args.warmup):
        step(i)
        torch.cuda.synchronize()
        st = ctx.status()
    assert not st.overflow, "pair buffers still overflow after warm-up"

    sampler = ClockSampler(local)
    stage_acc = {}
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if bands:
        # Single frame over tile-row bands: frame i's NCCL gather (side stream, framebuffer i&1) overlaps
        # frame i+1's kernels. Timed as one region of K frames on the device (events on the render
        # stream, which finally waits for the last gather). No L2 flush: scene planes + records
        # (6M x 98 B) are several times the 126 MB L2.
        fbs = [fb, torch.zeros_like(fb)]
        comm = torch.cuda.Stream(device=dev)
        ev_r = [torch.cuda.Event() for _ in range(2)]; ev_g = [torch.cuda.Event() for _ in range(2)]
        main = torch.cuda.current_stream()
        plain = capi.make_opts(mode=mode, sh_degree=args.sh_degree, tile_row_begin=row0, tile_row_end=row1)

        def frame(i):
            b = i & 1
            if i >= 2:
                main.wait_event(ev_g[b])
            if have_rows:
                ctx.render(stream, fbs[b].data_ptr(), *cam(i), w, h, plain)
            ev_r[b].record(main)
            with torch.cuda.stream(comm):
                comm.wait_event(ev_r[b])
                multigpu.gather_bands(fbs[b], rows_px, rank, world, dst=0)
                ev_g[b].record(comm)
        for i in range(4):
            frame(i)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.start()
        t_wall0 = time.perf_counter()
        e0.record(main)
        for i in range(args.steps):
            frame(i)
        main.wait_stream(comm)
        e1.record(main)
        torch.cuda.synchronize()
        dist.barrier()
        t_wall = time.perf_counter() - t_wall0
        sampler.stop_flag = True; sampler.join()
        total_ms = e0.elapsed_time(e1)
        for i in range(10):                        # per-stage times from a short separate, synchronising loop
            step(i)
            torch.cuda.synchronize()
            if have_rows:
                for k, v in ctx.stage_times().items():
                    stage_acc[k] = stage_acc.get(k, 0.0) + v * args.steps / 10.0
    else:
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
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
        per_step = [a.elapsed_time(b) for a, b in ev]
        total_ms = sum(per_step)
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    step_pct = None
    if not bands:      # this rank's per-step distribution (median and p10 / p90)
        q = np.percentile(np.asarray(per_step), [10, 50, 90])
        step_pct = {"p10": float(q[0]), "p50": float(q[1]), "p90": float(q[2])}
    frames_per_step = 1 if bands else world
    fps = frames_per_step * 1e3 / ms_per_step
    st = ctx.status()
    D = int(st.n_pairs); n_pass = int(st.sort_passes); C = int(getattr(st, 'n_candidates', 0)) or D

    # ---- e2e: through the public host-buffer API. Scene resident (uploaded once, as the reference's
    # initialize() does, src/instanced_splat_renderer.mm:399-422); every step copies that step's
    # matrices from host memory and brings the finished framebuffer back to pinned host memory.
    e2e_steps = max(8, min(args.steps, 64))
    eopts = opts_for(0)
    band_px = (rows_px[rank][1] - rows_px[rank][0]) * w if bands else w * h
    if bands:
        # Host delivery needs no collective: every rank copies its own band over its own PCIe link into ONE
        # page-locked host frame shared by the ranks (multigpu.SharedHostFrame); frame i's copy (side stream,
        # device framebuffer i & 1) overlaps frame i+1's kernels. Two shared frames alternate so a consumer
        # could read frame i while i+1 lands.
        frames = [multigpu.SharedHostFrame("splat_bench_%d_%d" % (os.getppid(), k), h, w, rank, world) for k in range(2)]
        efbs = [fb, torch.zeros_like(fb)]
        copy_s = torch.cuda.Stream(device=dev)
        ev_r2 = [torch.cuda.Event() for _ in range(2)]; ev_c2 = [torch.cuda.Event() for _ in range(2)]
        main2 = torch.cuda.current_stream()
        eplain = capi.make_opts(mode=mode, sh_degree=args.sh_degree, tile_row_begin=row0, tile_row_end=row1)

        def e2e_run(k):
            for i in range(k):
                b = i & 1
                if i >= 2:
                    main2.wait_event(ev_c2[b])
                if have_rows:
                    ctx.render(stream, efbs[b].data_ptr(), *cam(i), w, h, eplain)      # matrices come from host memory
                ev_r2[b].record(main2)
                with torch.cuda.stream(copy_s):
                    copy_s.wait_event(ev_r2[b])
                    frames[b].deliver(efbs[b], rows_px[rank])
                    ev_c2[b].record(copy_s)
            torch.cuda.synchronize()
        e2e_run(4)
        dist.barrier()
        t0 = time.perf_counter(); e2e_run(e2e_steps); dist.barrier(); e2e_s = time.perf_counter() - t0
        if rank == 0:       # the delivered frame is the whole frame: compare its last copy with a gathered one
            multigpu_check = frames[(e2e_steps - 1) & 1].tensor.clone()
        step(e2e_steps - 1)
        torch.cuda.synchronize(); dist.barrier()
        if rank == 0:
            assert torch.equal(multigpu_check, fb.cpu()), "shared host frame differs from the gathered frame"
        for fr in frames:
            fr.close()
        d2h = band_px * 8
        e2e_note = ("per frame: matrices from host, band rendered, band copied device -> shared page-locked host frame by every "
                    "rank in parallel (no collective; copies overlap the next frame); d2h bytes are this rank's band")
        lat_ms = None
    else:
        host_fbs = torch.empty((e2e_steps, h, w, 4), dtype=torch.float16).pin_memory()
        ptrs = [host_fbs[k].data_ptr() for k in range(e2e_steps)]
        cams = [cam(i) for i in range(e2e_steps)]
        views = np.stack([c[0] for c in cams]); projs = np.stack([c[1] for c in cams])
        ctx.render_batch_host(ptrs[:4], views[:4], projs[:4], w, h, eopts)
        t0 = time.perf_counter()
        for i in range(8):
            ctx.render_host(ptrs[i], views[i], projs[i], w, h, eopts)
        lat_ms = 1e3 * (time.perf_counter() - t0) / 8
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        ctx.render_batch_host(ptrs, views, projs, w, h, eopts)
        e2e_s = time.perf_counter() - t0
        # the same, delivered in the reference drawable's format (BGRA8Unorm, 4 B / pixel, packed on the device)
        bgra_fbs = torch.empty((e2e_steps, h, w, 4), dtype=torch.uint8).pin_memory()
        bptrs = [bgra_fbs[k].data_ptr() for k in range(e2e_steps)]
        bopts = opts_for(capi.FLAG_HOST_BGRA8)
        ctx.render_batch_host(bptrs[:4], views[:4], projs[:4], w, h, bopts)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        ctx.render_batch_host(bptrs, views, projs, w, h, bopts)
        e2e_bgra_s = time.perf_counter() - t0
        del bgra_fbs
        d2h = w * h * 8 + 16
        e2e_note = ("splat_render_batch_host: %d frames, each frame's matrices from host and its framebuffer to pinned "
                    "host memory, copies pipelined behind the next frame's kernels" % e2e_steps)
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_fps = frames_per_step * e2e_steps / e2e_s
    e2e_bgra_fps = None
    if not bands:
        if world > 1:
            t = torch.tensor([e2e_bgra_s], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_bgra_s = float(t.item())
        e2e_bgra_fps = frames_per_step * e2e_steps / e2e_bgra_s

    # ---- device-resident camera batch (splat_render_batch: two frames in flight on two internal streams)
    inflight = None
    if not bands and world == 1:
        nb = 8
        dfb = torch.zeros((nb, h, w, 4), dtype=torch.float16, device=dev)
        dptrs = [dfb[k % nb].data_ptr() for k in range(e2e_steps)]
        ctx.render_batch(stream, dptrs[:4], views[:4], projs[:4], w, h, eopts)
        torch.cuda.synchronize()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record()
        ctx.render_batch(stream, dptrs, views, projs, w, h, eopts)
        b1.record(); torch.cuda.synchronize()
        two = (not mode) and n <= 2_500_000 and w * h <= 2_400_000      # the library's own rule (splat_api.cu: two_frames_in_flight)
        inflight = {"frames_per_s": e2e_steps * 1e3 / b0.elapsed_time(b1), "frames": e2e_steps, "frames_in_flight": 2 if two else 1,
                    "note": "splat_render_batch, framebuffers stay on the device; no L2 flush (a frame touches ~2x the L2)"}
        del dfb

    if rank == 0:
        hbm_peak, peak_kind, sm_max = peaks()
        visible = int((ctx.fetch(capi.FETCH_TILES_TOUCHED) > 0).sum())
        P = band_px if bands else w * h
        ab = algorithmic_bytes(n, visible, D, C, P, args.sh_degree, layer=bool(mode))
        stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
        stages = {}
        for k in STAGES:
            ms = stage_ms.get(k, 0.0)
            gbs = ab[k] / ms / 1e6 if ms > 0 else 0.0
            stages[k] = {"ms": ms, "alg_bytes": ab[k], "gb_per_s": gbs, "frac_of_hbm_peak": gbs / hbm_peak}
        dom = max(STAGES, key=lambda k: stage_ms.get(k, 0.0))
        clocks = sampler.result()
        roof = {"kernel": {"blend": "blend_kernel", "sort": "tile_sort_kernel", "preprocess": "preprocess_kernel",
                           "scan_duplicate": "bin_pairs_kernel", "ranges": "-"}[dom],
                "bound": "hbm", "achieved": stages[dom]["gb_per_s"], "peak": hbm_peak, "unit": "GB/s",
                "frac": stages[dom]["frac_of_hbm_peak"], "frac_of_8TBs_nominal": stages[dom]["gb_per_s"] / 8000.0,
                "traffic": measured_traffic(wl, dom), "peak_kind": peak_kind,
                "alg_bytes_per_launch": ab[dom], "ms_per_launch": stage_ms.get(dom),
                "share_of_step": stage_ms.get(dom, 0.0) / max(1e-9, stage_ms.get("total", 0.0)),
                "note": ("the blend is bound by SM instruction issue, not HBM (ncu: issue slots ~89% busy, DRAM ~2%); "
                         "its HBM fraction is reported as the contract asks, sm_issue below is the number that bounds it")
                if dom == "blend" else ""}
        if dom == "blend":
            # unit = one (pixel, splat) test at ~16 fp32 ops; lane peak = SMs * 128 lanes * clock / 16
            sm_mhz = clocks["sm_mhz"] or sm_max
            lane_peak = 148 * 128 * sm_mhz * 1e6 / 16.0
            tests = float(D) * 256.0                                  # upper bound: every pair against the whole tile
            roof["sm_issue"] = {"pair_tile_tests_per_s": tests / (stage_ms["blend"] * 1e-3), "lane_peak_tests_per_s": lane_peak,
                                "frac": tests / (stage_ms["blend"] * 1e-3) / lane_peak,
                                "note": "D*256 tests counted although 8x4-block culling skips most of them"}
        line = {
            "metric": "frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if bands else "weak", "vs_baseline": None, "dtype": "f32+f16", "data": "synthetic",
            "mgaussians_per_s": n * fps / 1e6, "ms_per_step_percentiles": step_pct,
            "config": {"workload": name, "n_gaussians": n, "width": w, "height": h,
                       "mode": "LAYER50" if mode else "TILE", "order": "near_first", "sh_degree": args.sh_degree,
                       "pairs_D": D, "visible": visible, "sort_passes": n_pass,
                       "parallelism": ("bands%d+nccl_gather" % world) if bands else ("views%d" % world),
                       "bands": band_list if bands else None,
                       "l2": ("inputs larger than L2 (scene planes + records ~590 MB); frames back to back, gather of frame i "
                              "overlapped with frame i+1") if bands else "flushed between timed steps (256 MiB write, untimed)"},
            "roofline": roof, "stages": stages,
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": 128, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "single_frame_ms": lat_ms, "note": e2e_note,
                    "bgra8": None if bands else {
                        "value": e2e_bgra_fps, "d2h_bytes_per_step": w * h * 4 + 16,
                        "note": "same call with SPLAT_FLAG_HOST_BGRA8: frames delivered as the reference drawable's BGRA8Unorm"}},
            "camera_batch_device": inflight,
            "gpu_launches": args.steps * 7,      # preprocess, tile scan, bin pairs, tile offsets, scatter, tile sort, blend (+1 memset node, not counted)
            "clocks": clocks, "wall_s": t_wall, "numa": numa,
        }
        if world == 1 and not args.no_sort_bench:
            sb = sort_bench(ctx, torch)
            sb["frac_of_hbm_peak"] = sb["gb_per_s"] / hbm_peak
            sb["frac_of_8TBs_nominal"] = sb["gb_per_s"] / 8000.0
            line["sort_bench"] = sb
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_leg(inst, *camera.app_camera(w, h), w, h, mode, name)
        emit(line)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
