#!/usr/bin/env python
"""Benchmark of the splat rendering hot path (BASELINE.json metric: frames/s and Mgaussians/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4|c5]

A step = one frame of the hot path (K1 preprocess .. K5 blend) over one synthetic scene.
N=1 default workload = BASELINE.json configs[1]: 1M gaussians, 1920x1080, TILE mode, single frame.
N>1 default: every rank renders the same scene from its own orbit cameras (views sharded, no
data-path collective, weak scaling: value = frames of all ranks / max-over-ranks time), and in the
same invocation the config-3 leg - the 4K/6M single frame split into tile-row bands with an NCCL
gather of the framebuffer (strong scaling), its one-GPU rate measured by rank 0 in the same run -
reported under config.c3; `--workload c3` makes that leg the headline. Prints ONE JSON line on rank 0.
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
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample, "cpu_model": cpu_model(),
                         "note": "K1 and the blend use every core; the restatement's scan / duplication / sort / tile ranges are single-threaded"},
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


def sort_bench(ctx, torch, n=1 << 26, reps=7):
    """Stand-alone 'sort HBM GB/s' of the 64-bit entry point: full-entropy keys, 8 passes, D*(8+8*24) bytes."""
    torch.cuda.empty_cache()       # fresh allocations for the 1.6 GB of buffers (recycled, fragmented blocks of the caching allocator cost 2-3%)
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


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for ln in f:
                if ln.lower().startswith("model name"):
                    return ln.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def c3_bands_leg(ctx, torch, dist, args, dev, rank, world, steps, sh_degree=0, even_bands=False):
    """Config 3: the 6M-gaussian / 3840x2160 single frame split into tile-row bands over the ranks with the NCCL gather of
    the framebuffer (the one multi-GPU path north_star defines; strong scaling). In the same run: the one-GPU frame rate
    (rank 0 alone, the others idle at a barrier) and a check that the gathered frame equals the one-GPU frame bit for bit.
    Returns the result dict (meaningful on rank 0)."""
    from gaussian_splat_b200 import camera, capi, multigpu, scene
    n, w, h, mode, cfg, name = WORKLOADS["c3"]
    inst, sh = scene.make_instances(n, scene.CONFIG_SEEDS[cfg], sh_rest=sh_degree > 0)
    n = inst.shape[0]
    ctx.upload(inst)
    if sh_degree:
        ctx.upload_sh(sh, 3)
    del inst
    stream = torch.cuda.current_stream().cuda_stream
    tiles_x, tiles_y = (w + 15) // 16, (h + 15) // 16
    V, P = camera.app_camera(w, h)
    fb = torch.zeros((h, w, 4), dtype=torch.float16, device=dev)
    full = capi.make_opts(mode=mode, sh_degree=sh_degree)

    # ---- the whole frame on one GPU (every rank renders it once: warm-up, buffer growth, pairs per tile row)
    for _ in range(3):
        ctx.render(stream, fb.data_ptr(), V, P, w, h, full); torch.cuda.synchronize()
        st_full = ctx.status()
        if not st_full.overflow:
            break
    assert not st_full.overflow
    r = ctx.fetch(capi.FETCH_TILE_RANGES).astype(np.int64)
    row_pairs = (r[:, 1] - r[:, 0]).reshape(tiles_y, tiles_x).sum(axis=1)
    fb_single = fb.clone() if rank == 0 else None
    one_gpu = None
    dist.barrier()
    if rank == 0:       # same run, same box: rank 0 alone, frames back to back like the banded loop below
        tim = capi.make_opts(mode=mode, sh_degree=sh_degree, flags=capi.FLAG_TIMING)
        acc = {}
        for it in range(8):
            ctx.render(stream, fb.data_ptr(), V, P, w, h, tim); torch.cuda.synchronize()
            if it >= 3:
                for k, v in ctx.stage_times().items():
                    acc[k] = acc.get(k, 0.0) + v / 5
        # frames back to back through splat_render_batch, the same call the banded loop below makes (the library keeps two
        # frames in flight on two frame sets)
        fb1 = [fb] + [torch.zeros_like(fb) for _ in range(3)]
        vs1, ps1 = np.stack([V] * steps), np.stack([P] * steps)
        ptr1 = [fb1[i & 3].data_ptr() for i in range(steps)]
        ctx.render_batch(stream, ptr1[:4], vs1[:4], ps1[:4], w, h, full); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.render_batch(stream, ptr1, vs1, ps1, w, h, full)
        e1.record(); torch.cuda.synchronize()
        assert torch.equal(fb1[1], fb_single) and torch.equal(fb1[2], fb_single)
        del fb1
        one_gpu = {"frames_per_s": steps * 1e3 / e0.elapsed_time(e1), "ms_per_frame": e0.elapsed_time(e1) / steps,
                   "stage_us": {k: round(v * 1e3, 1) for k, v in acc.items()}, "pairs_D": int(st_full.n_pairs)}
    dist.barrier()

    # ---- bands: boundaries balanced on the pairs per tile row, then six rounds of measured-time balancing (untimed
    # set-up: pairs are not the cost, rows of large near splats blend slower per pair)
    band_list = multigpu.band_rows(tiles_y, world, None if even_bands else row_pairs)
    if not even_bands:
        wts = row_pairs.astype(np.float64)
        for _ in range(6):
            a, b = band_list[rank]
            o = capi.make_opts(mode=mode, sh_degree=sh_degree, tile_row_begin=a, tile_row_end=b, flags=capi.FLAG_TIMING)
            acc = []
            for it in range(6):
                var = 0.0
                if b > a:
                    ctx.render(stream, fb.data_ptr(), V, P, w, h, o); torch.cuda.synchronize()
                    st_ = ctx.stage_times()
                    var = st_["total"] - 0.6 * st_["preprocess"]      # the part of the band's time that moves with its rows
                if it >= 2:
                    acc.append(var)
            tt = torch.tensor([float(np.median(acc))], dtype=torch.float64, device=dev)
            allt = [torch.zeros_like(tt) for _ in range(world)]
            dist.all_gather(allt, tt)
            wts = multigpu.reweight_rows(wts, band_list, [float(x.item()) for x in allt])
            band_list = multigpu.band_rows(tiles_y, world, wts)
    GROUP = 8                                   # frames per splat_render_batch call; two groups of framebuffers alternate
    fbs = [fb] + [torch.zeros_like(fb) for _ in range(2 * GROUP - 1)]
    comm = torch.cuda.Stream(device=dev)
    ev_r = [torch.cuda.Event() for _ in range(2)]; ev_g = [torch.cuda.Event() for _ in range(2)]
    main_s = torch.cuda.current_stream()
    vsg, psg = np.stack([V] * GROUP), np.stack([P] * GROUP)
    st8 = {}                                    # the current partition: set by use_bands()

    def use_bands(bl):
        a, b = bl[rank]
        st8.update(row0=a, row1=b, rows_px=multigpu.bands_to_pixel_rows(bl, h), have_rows=b > a,
                   plain=capi.make_opts(mode=mode, sh_degree=sh_degree, tile_row_begin=a, tile_row_end=b))

    peer = multigpu.PeerFrames(2 * GROUP, h, w, rank, world, dev, dst=0)      # rank 0's framebuffers, mapped by every rank

    def group(g, ev_pair=None, exchange="peer"):
        """GROUP frames of this rank's band (the library keeps two in flight), then the exchange on the side stream:
        "peer"  the blend wrote the band straight into rank 0's framebuffer over NVLink; a 4-byte all-reduce signals completion
        "nccl"  the bands are gathered on rank 0 by grouped ncclSend / ncclRecv (one NCCL group for the GROUP frames)
        ev_pair: timing events recorded around the rendering alone (after the wait for the framebuffers)."""
        b = g & 1
        mine_fbs = fbs[b * GROUP:(b + 1) * GROUP]
        if g >= 2:
            main_s.wait_event(ev_g[b])          # the exchange that last touched these framebuffers
        if ev_pair:
            ev_pair[0].record(main_s)
        if st8["have_rows"]:
            ptrs = [peer.ptr(b * GROUP + k) for k in range(GROUP)] if exchange == "peer" else [f.data_ptr() for f in mine_fbs]
            ctx.render_batch(stream, ptrs, vsg, psg, w, h, st8["plain"])
        if ev_pair:
            ev_pair[1].record(main_s)
        ev_r[b].record(main_s)
        with torch.cuda.stream(comm):
            comm.wait_event(ev_r[b])
            if exchange == "peer":
                peer.signal()
            elif exchange == "nccl":
                multigpu.gather_bands_batch(mine_fbs, st8["rows_px"], rank, world, dst=0)
            ev_g[b].record(comm)

    use_bands(band_list)
    if not even_bands:
        # three more rounds on what each rank's rendering takes INSIDE the real loop: two frames in flight, and on rank 0
        # the gather's receive kernels sharing the SMs (the serial stage times above see neither)
        for _ in range(3):
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(6)]
            for g in range(6):
                group(g, evs[g])
            torch.cuda.synchronize(); dist.barrier()
            mine_t = float(np.median([a_.elapsed_time(b_) for a_, b_ in evs[2:]])) / GROUP
            tt = torch.tensor([mine_t], dtype=torch.float64, device=dev)
            allt = [torch.zeros_like(tt) for _ in range(world)]
            dist.all_gather(allt, tt)
            wts = multigpu.reweight_rows(wts, band_list, [float(x.item()) for x in allt], damping=0.5)
            band_list = multigpu.band_rows(tiles_y, world, wts)
            use_bands(band_list)
    row0, row1, rows_px, have_rows, plain = st8["row0"], st8["row1"], st8["rows_px"], st8["have_rows"], st8["plain"]
    timed = capi.make_opts(mode=mode, sh_degree=sh_degree, tile_row_begin=row0, tile_row_end=row1, flags=capi.FLAG_TIMING)

    # ---- parity of the real NCCL path: bands rendered by the ranks, gathered on rank 0, against rank 0's own whole frame
    fb.zero_()
    if have_rows:
        ctx.render(stream, fb.data_ptr(), V, P, w, h, plain)
    multigpu.gather_bands(fb, rows_px, rank, world, dst=0)
    torch.cuda.synchronize(); dist.barrier()
    equal = bool(torch.equal(fb, fb_single)) if rank == 0 else True
    assert equal, "the NCCL-gathered band frame differs from the single-GPU frame"

    # ---- timed: groups of eight frames through splat_render_batch (two frames in flight inside the library); a group's
    # gathers (side stream) overlap the next group's kernels; one region of K frames, device-timed, the render stream finally
    # waits for the last gather. No L2 flush: scene planes + records are several times the 126 MB L2.
    n_groups = max(1, steps // GROUP)
    steps = n_groups * GROUP
    sampler = ClockSampler(dev.index)

    def timed_loop(exchange, sample=False):
        for g in range(4):
            group(g, exchange=exchange)
        torch.cuda.synchronize(); dist.barrier()
        if rank == 0:       # every framebuffer of both groups holds a complete frame, equal to the one-GPU frame
            frames_ = [peer.tensor(k) for k in range(2 * GROUP)] if exchange == "peer" else fbs
            assert all(torch.equal(f, fb_single) for f in frames_), "a pipelined band frame (%s) differs from the single-GPU frame" % exchange
        t0_, t1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if sample:
            sampler.start()
        t0_.record(main_s)
        for g in range(n_groups):
            group(g, exchange=exchange)
        main_s.wait_stream(comm)
        t1_.record(main_s)
        torch.cuda.synchronize()
        dist.barrier()
        if sample:
            sampler.stop_flag = True; sampler.join()
        t_ = torch.tensor([t0_.elapsed_time(t1_)], dtype=torch.float64, device=dev)
        dist.all_reduce(t_, op=dist.ReduceOp.MAX)
        return float(t_.item()) / steps

    ms_nccl = timed_loop("nccl")
    ms_per_frame = timed_loop("peer", sample=True)
    fbs = fbs[:2]

    # ---- per-stage times of every rank's band (a short separate, synchronising loop) and the gather on its own
    acc = {}
    for it in range(10):
        if have_rows:
            ctx.render(stream, fb.data_ptr(), V, P, w, h, timed)
        torch.cuda.synchronize()
        if have_rows and it >= 3:
            for k, v in ctx.stage_times().items():
                acc[k] = acc.get(k, 0.0) + v / 7
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gts = []
    for it in range(6):
        dist.barrier(); torch.cuda.synchronize()
        g0.record(); multigpu.gather_bands(fb, rows_px, rank, world, dst=0); g1.record(); torch.cuda.synchronize()
        gts.append(g0.elapsed_time(g1))
    st_band = ctx.status()
    mine = {"rank": rank, "rows": [row0, row1], "pairs": int(st_band.n_pairs), "live_blocks": int(st_band.n_live_blocks),
            "stage_us": {k: round(v * 1e3, 1) for k, v in acc.items()}, "gather_us": round(float(np.median(gts[1:])) * 1e3, 1)}
    everyone = [None] * world
    dist.all_gather_object(everyone, mine)

    # ---- end to end: host delivery needs no collective. Every rank copies its own band over its own PCIe link into ONE
    # page-locked host frame shared by the ranks; frame i's copy (side stream) overlaps frame i+1's kernels.
    e2e_steps = max(8, min(steps, 64))
    frames = [multigpu.SharedHostFrame("splat_bench_%d_%d" % (os.getppid(), k), h, w, rank, world) for k in range(2)]
    copy_s = torch.cuda.Stream(device=dev)
    ev_r2 = [torch.cuda.Event() for _ in range(2)]; ev_c2 = [torch.cuda.Event() for _ in range(2)]

    def e2e_run(k):
        for i in range(k):
            b = i & 1
            if i >= 2:
                main_s.wait_event(ev_c2[b])
            if have_rows:
                ctx.render(stream, fbs[b].data_ptr(), V, P, w, h, plain)      # matrices come from host memory
            ev_r2[b].record(main_s)
            with torch.cuda.stream(copy_s):
                copy_s.wait_event(ev_r2[b])
                frames[b].deliver(fbs[b], rows_px[rank])
                ev_c2[b].record(copy_s)
        torch.cuda.synchronize()
    e2e_run(4)
    dist.barrier()
    t0 = time.perf_counter(); e2e_run(e2e_steps); dist.barrier(); e2e_s = time.perf_counter() - t0
    host_equal = True
    if rank == 0:
        host_equal = bool(torch.equal(frames[(e2e_steps - 1) & 1].tensor, fb_single.cpu()))
        assert host_equal, "the shared host frame differs from the single-GPU frame"
    for fr in frames:
        fr.close()
    peer.close()
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_fps = e2e_steps / float(t.item())
    fps = 1e3 / ms_per_frame
    slow = max(everyone, key=lambda d: d["stage_us"].get("total", 0.0))
    out = {
        "workload": name, "n_gaussians": n, "width": w, "height": h, "mode": "TILE", "parallelism": "bands%d+nccl_gather" % world,
        "frames_per_s": fps, "ms_per_frame": ms_per_frame, "steps": steps,
        "exchange": "fused: every rank's blend kernel writes its band straight into rank 0's framebuffer over NVLink (CUDA IPC peer "
                    "memory, splat_peer_alloc / splat_peer_open); one 4-byte NCCL all-reduce per group of %d frames is the completion signal" % GROUP,
        "frames_per_s_nccl_gather": 1e3 / ms_nccl, "speedup_nccl_gather": None,
        "nccl_gather_note": "the same loop with the bands gathered on rank 0 by grouped ncclSend / ncclRecv (splat_gather_bands_batch) instead",
        "frames_per_s_1gpu": None, "speedup": None, "one_gpu": None,
        "bands": band_list, "band_balance": "even" if even_bands else "pairs per tile row, 6 rounds on serial stage times, 3 rounds on each rank's render time inside the pipelined loop",
        "per_rank": everyone, "slowest_rank_stage_us": slow["stage_us"], "gather_us": max(d["gather_us"] for d in everyone),
        "gather_bytes_per_frame": w * h * 8,
        "e2e_frames_per_s": e2e_fps, "e2e_steps": e2e_steps, "e2e_d2h_bytes_per_step_per_rank": (rows_px[rank][1] - rows_px[rank][0]) * w * 8,
        "band_frame_equals_single_gpu": equal, "shared_host_frame_equals_single_gpu": host_equal,
        "clocks": sampler.result(),
        "l2": "inputs larger than L2 (scene planes + records ~700 MB); frames back to back, the gather of frame i overlapped with frame i+1",
    }
    if rank == 0:
        out["frames_per_s_1gpu"] = one_gpu["frames_per_s"]; out["speedup"] = fps / one_gpu["frames_per_s"]; out["one_gpu"] = one_gpu
        out["speedup_nccl_gather"] = out["frames_per_s_nccl_gather"] / one_gpu["frames_per_s"]
    return out


def c4_leg(ctx, torch, dist, dev, rank, world, reps=3):
    """Config 4: 3M gaussians, 64 orbit views at 1080p, views block-sharded over the ranks (8 per GPU at N = 8), no data-path
    collective. Every rank renders its views through splat_render_batch into device framebuffers; frames/s = 64 / the slowest
    rank's device time (median of `reps` passes)."""
    from gaussian_splat_b200 import capi, multigpu, scene
    n, w, h, mode, cfg, name = WORKLOADS["c4"]
    inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[cfg])
    n = inst.shape[0]
    ctx.upload(inst)
    del inst
    n_views = 64
    mine = multigpu.shard_views(n_views, rank, world)
    cams = [orbit_camera(k, n_views, w, h) for k in mine]
    views = np.stack([c[0] for c in cams]); projs = np.stack([c[1] for c in cams])
    fbs = torch.zeros((min(8, len(mine)), h, w, 4), dtype=torch.float16, device=dev)
    ptrs = [fbs[k % fbs.shape[0]].data_ptr() for k in range(len(mine))]
    stream = torch.cuda.current_stream().cuda_stream
    opts = capi.make_opts(mode=mode)
    for _ in range(2):                      # warm-up; grows the pair buffers if a view overflowed
        ctx.render_batch(stream, ptrs, views, projs, w, h, opts); torch.cuda.synchronize()
        st = ctx.status()
    assert not st.overflow
    ts = []
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.render_batch(stream, ptrs, views, projs, w, h, opts)
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t.item()))
    ms = float(np.median(ts))
    return {"workload": name, "n_gaussians": n, "width": w, "height": h, "mode": "TILE", "views": n_views, "views_per_gpu": len(mine),
            "parallelism": "views%d" % world, "frames_per_s": n_views * 1e3 / ms, "mgaussians_per_s": n * n_views * 1e3 / ms / 1e6,
            "ms_for_all_views": ms, "pairs_D_last_view": int(st.n_pairs),
            "note": "splat_render_batch (two frames in flight), framebuffers stay on the device; max over ranks, median of %d passes" % reps}


def c5_leg(ctx, torch, dev, steps=60):
    """Config 5: 2M gaussians, the 50-layer variant (K = 50; the mode the reference app executes), 1080p, one GPU: single frames
    with an L2 flush between them, as the headline loop."""
    from gaussian_splat_b200 import camera, capi, scene
    n, w, h, mode, cfg, name = WORKLOADS["c5"]
    inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[cfg])
    n = inst.shape[0]
    ctx.upload(inst)
    del inst
    V, P = camera.app_camera(w, h)
    fb = torch.zeros((h, w, 4), dtype=torch.float16, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    opts = capi.make_opts(mode=mode, flags=capi.FLAG_TIMING)
    for _ in range(4):
        ctx.render(stream, fb.data_ptr(), V, P, w, h, opts); torch.cuda.synchronize()
        st = ctx.status()
    assert not st.overflow
    acc, total = {}, 0.0
    for i in range(steps):
        flush.fill_(i & 0xFF)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.render(stream, fb.data_ptr(), V, P, w, h, opts); e1.record(); e1.synchronize()
        total += e0.elapsed_time(e1)
        for k, v in ctx.stage_times().items():
            acc[k] = acc.get(k, 0.0) + v / steps
    ms = total / steps
    return {"workload": name, "n_gaussians": n, "width": w, "height": h, "mode": "LAYER50", "K": 50, "order": "near_first",
            "frames_per_s": 1e3 / ms, "mgaussians_per_s": n / ms / 1e3, "ms_per_frame": ms, "pairs_D": int(st.n_pairs),
            "stage_us": {k: round(v * 1e3, 1) for k, v in acc.items()}, "steps": steps,
            "l2": "flushed between timed frames (256 MiB write, untimed)"}


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
    ap.add_argument("--no-c3-leg", action="store_true", help="N > 1: skip the config-3 band + gather leg reported in config.c3")
    ap.add_argument("--no-extra-legs", action="store_true", help="skip the config-4 (and, at N = 1, config-5) legs reported in config.c4 / config.c5")
    ap.add_argument("--even-bands", action="store_true", help="c3: equal tile-row bands instead of balanced ones")
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
    ctx = capi.Context(local)
    hbm_peak, peak_kind, sm_max = peaks()

    if wl == "c3" and world > 1:
        # --workload c3 at N > 1: the band + gather leg IS the headline (strong scaling)
        leg = c3_bands_leg(ctx, torch, dist, args, dev, rank, world, args.steps, args.sh_degree, args.even_bands)
        if rank == 0:
            n, w, h = leg["n_gaussians"], leg["width"], leg["height"]
            line = {
                "metric": "frames_per_s", "value": leg["frames_per_s"], "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": leg["ms_per_frame"], "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32+f16", "data": "synthetic", "mgaussians_per_s": n * leg["frames_per_s"] / 1e6,
                "config": {"workload": leg["workload"], "n_gaussians": n, "width": w, "height": h, "mode": "TILE", "order": "near_first",
                           "parallelism": leg["parallelism"], "c3": leg, "l2": leg["l2"]},
                "e2e": {"value": leg["e2e_frames_per_s"], "unit": "frames/s", "h2d_bytes_per_step": 128,
                        "d2h_bytes_per_step": leg["e2e_d2h_bytes_per_step_per_rank"], "steps": leg["e2e_steps"],
                        "note": "per frame: matrices from host, band rendered, band copied device -> shared page-locked host frame by "
                                "every rank in parallel (no collective; copies overlap the next frame); d2h bytes are rank 0's band"},
                "gpu_launches": args.steps * 7, "clocks": leg["clocks"], "numa": numa,
            }
            emit(line)
        ctx.close()
        dist.barrier(); dist.destroy_process_group()
        return

    n, w, h, mode, cfg, name = WORKLOADS[wl]
    inst, sh = scene.make_instances(n, scene.CONFIG_SEEDS[cfg], sh_rest=args.sh_degree > 0)
    n = inst.shape[0]
    ctx.upload(inst)
    if args.sh_degree:
        ctx.upload_sh(sh, 3)
    stream = torch.cuda.current_stream().cuda_stream
    fb = torch.zeros((h, w, 4), dtype=torch.float16, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    n_views = 64
    view_sharded = world > 1 or wl == "c4"

    def cam(step):      # N=1 uses the app camera (configs 2/3/5); view-sharded runs walk the 64-view orbit
        if view_sharded:
            return orbit_camera((step * world + rank) % n_views, n_views, w, h)
        return camera.app_camera(w, h)

    opts = capi.make_opts(mode=mode, sh_degree=args.sh_degree, flags=capi.FLAG_TIMING)

    def step(i):
        ctx.render(stream, fb.data_ptr(), *cam(i), w, h, opts)

    # warm-up (also grows the pair buffers if a frame overflowed: status() reports it and the frame is rendered again)
    for i in range(max(args.warmup, n_views // world if view_sharded else 0)):
        step(i)
        torch.cuda.synchronize()
        if ctx.status().overflow:
            step(i); torch.cuda.synchronize(); ctx.status()
    st = ctx.status()
    assert not st.overflow, "pair buffers still overflow after warm-up"

    sampler = ClockSampler(local)
    stage_acc = {}
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
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
    q = np.percentile(np.asarray(per_step), [10, 50, 90])      # this rank's per-step distribution
    step_pct = {"p10": float(q[0]), "p50": float(q[1]), "p90": float(q[2])}
    fps = world * 1e3 / ms_per_step
    # the app-camera frame's counts (view-sharded runs end on some orbit view: render the app camera once more for them)
    ctx.render(stream, fb.data_ptr(), *camera.app_camera(w, h), w, h, capi.make_opts(mode=mode, sh_degree=args.sh_degree,
                                                                                       flags=capi.FLAG_COUNT_WORK))
    torch.cuda.synchronize()
    st = ctx.status()
    D = int(st.n_pairs); slots = int(st.n_candidates); warp_cand = int(st.blend_warp_candidates)
    visible = int((ctx.fetch(capi.FETCH_TILES_TOUCHED) > 0).sum()) if rank == 0 else 0

    # ---- e2e: through the public host-buffer API. Scene resident (uploaded once, as the reference's
    # initialize() does, src/instanced_splat_renderer.mm:399-422); every step copies that step's
    # matrices from host memory and brings the finished framebuffer back to pinned host memory.
    e2e_steps = max(8, min(args.steps, 64))
    eopts = capi.make_opts(mode=mode, sh_degree=args.sh_degree)
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
    del host_fbs
    # the same, delivered in the reference drawable's format (BGRA8Unorm, 4 B / pixel, packed on the device)
    bgra_fbs = torch.empty((e2e_steps, h, w, 4), dtype=torch.uint8).pin_memory()
    bptrs = [bgra_fbs[k].data_ptr() for k in range(e2e_steps)]
    bopts = capi.make_opts(mode=mode, sh_degree=args.sh_degree, flags=capi.FLAG_HOST_BGRA8)
    ctx.render_batch_host(bptrs[:4], views[:4], projs[:4], w, h, bopts)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    ctx.render_batch_host(bptrs, views, projs, w, h, bopts)
    e2e_bgra_s = time.perf_counter() - t0
    # the box's host-write ceiling: the same number of plain device-to-pinned-host copies of one half4 frame, all ranks at once
    if world > 1:
        dist.barrier()
    hb = torch.empty((4, h, w, 4), dtype=torch.uint8).pin_memory()
    src = torch.zeros((h, w, 4), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        hb[i & 3].copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    d2h_s = time.perf_counter() - t0
    del bgra_fbs, hb, src
    if world > 1:
        t = torch.tensor([e2e_s, e2e_bgra_s, d2h_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s, e2e_bgra_s, d2h_s = (float(x) for x in t.tolist())
    e2e_fps = world * e2e_steps / e2e_s
    e2e_bgra_fps = world * e2e_steps / e2e_bgra_s
    bgra_ceiling_fps = world * e2e_steps / d2h_s

    # ---- device-resident camera batch (splat_render_batch: two frames in flight on two internal streams)
    inflight = None
    if world == 1:
        nb = 8
        dfb = torch.zeros((nb, h, w, 4), dtype=torch.float16, device=dev)
        dptrs = [dfb[k % nb].data_ptr() for k in range(e2e_steps)]
        ctx.render_batch(stream, dptrs[:4], views[:4], projs[:4], w, h, eopts)
        torch.cuda.synchronize()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record()
        ctx.render_batch(stream, dptrs, views, projs, w, h, eopts)
        b1.record(); torch.cuda.synchronize()
        two = os.environ.get("SPLAT_TWO_FRAMES", "1") != "0"      # the library's own rule (splat_api.cu: two_frames_in_flight)
        inflight = {"frames_per_s": e2e_steps * 1e3 / b0.elapsed_time(b1), "frames": e2e_steps, "frames_in_flight": 2 if two else 1,
                    "note": "splat_render_batch, framebuffers stay on the device; no L2 flush (a frame touches ~2x the L2)"}
        del dfb

    sb = None
    if world == 1 and not args.no_sort_bench:
        sb = sort_bench(ctx, torch)
        sb["frac_of_hbm_peak"] = sb["gb_per_s"] / hbm_peak
        sb["frac_of_8TBs_nominal"] = sb["gb_per_s"] / 8000.0
    cpu_leg = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_leg = cpu_baseline_leg(inst, *camera.app_camera(w, h), w, h, mode, name)
        cpu_leg["cpu_model"] = cpu_model()
        cpu_leg["note"] = "single thread = the named baseline; the --impl reference arm uses all cores (its binning stages stay single-threaded)"
    del inst
    c4 = c5 = None
    if wl == "c2" and not args.no_extra_legs:
        del fb, flush
        torch.cuda.empty_cache()
        fb = flush = None
        c4 = c4_leg(ctx, torch, dist, dev, rank, world)
        if world == 1:
            c5 = c5_leg(ctx, torch, dev)
    c3 = None
    if world > 1 and wl == "c2" and not args.no_c3_leg:
        fb = flush = None
        torch.cuda.empty_cache()
        c3 = c3_bands_leg(ctx, torch, dist, args, dev, rank, world, min(args.steps, 100), 0, args.even_bands)

    if rank == 0:
        P = w * h
        ab = algorithmic_bytes(n, visible, D, slots, P, args.sh_degree, layer=bool(mode))
        stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
        stages = {}
        for k in STAGES:
            ms = stage_ms.get(k, 0.0)
            gbs = ab[k] / ms / 1e6 if ms > 0 else 0.0
            stages[k] = {"ms": ms, "alg_bytes": ab[k], "gb_per_s": gbs, "frac_of_hbm_peak": gbs / hbm_peak}
        dom = max(STAGES, key=lambda k: stage_ms.get(k, 0.0))
        clocks = sampler.result()
        kernel_of = {"blend": "blend_kernel", "sort": "tile_sort_kernel", "preprocess": "preprocess_kernel",
                     "scan_duplicate": "scatter_pairs_kernel", "ranges": "-"}
        hbm = {"achieved": stages[dom]["gb_per_s"], "peak": hbm_peak, "unit": "GB/s", "frac": stages[dom]["frac_of_hbm_peak"],
               "frac_of_8TBs_nominal": stages[dom]["gb_per_s"] / 8000.0, "peak_kind": peak_kind, "alg_bytes_per_launch": ab[dom]}
        roof = {"kernel": kernel_of[dom], "traffic": measured_traffic(wl, dom), "ms_per_launch": stage_ms.get(dom),
                "share_of_step": stage_ms.get(dom, 0.0) / max(1e-9, stage_ms.get("total", 0.0))}
        if dom == "blend":
            # The blend is bound by SM instruction issue, not HBM (ncu: issue slots ~86% busy, DRAM ~4%). Its work unit
            #  is one (pixel, splat) test at ~16 fp32 ops; EXECUTED tests = the (8x4-pixel warp, splat)
            # candidates the kernel actually ran (counted by the kernel itself in an untimed frame) x 32 lanes.
            # Peak = SMs x 128 lanes x clock / 16 ops.
            sm_mhz = clocks["sm_mhz"] or sm_max
            lane_peak = 148 * 128 * sm_mhz * 1e6 / 16.0
            tests_per_s = warp_cand * 32.0 / (stage_ms["blend"] * 1e-3)
            roof.update({"bound": "sm_issue", "achieved": tests_per_s, "peak": lane_peak, "unit": "lane-tests/s", "frac": tests_per_s / lane_peak,
                         "warp_candidates_per_frame": warp_cand, "hbm": hbm,
                         "note": "bound = SM instruction issue; achieved counts executed lane-tests only (warp candidates x 32), not D x 256; "
                                 "the HBM view of the same kernel is under 'hbm' (small by construction: records are shared through L2)"})
        else:
            roof.update({"bound": "hbm", **hbm})
        config = {"workload": name, "n_gaussians": n, "width": w, "height": h,
                  "mode": "LAYER50" if mode else "TILE", "order": "near_first", "sh_degree": args.sh_degree,
                  "pairs_D": D, "parking_slots": slots, "visible": visible, "sort_passes": 1,
                  "live_blocks": int(st.n_live_blocks), "blocks": (n + 63) // 64, "longest_tile_list": int(st.max_tile_pairs),
                  "parallelism": "views%d" % world,
                  "l2": "flushed between timed steps (256 MiB write, untimed)",
                  "e2e_bgra8_frames_per_s": e2e_bgra_fps, "e2e_half4_frames_per_s": e2e_fps,
                  "host_write_ceiling_frames_per_s": bgra_ceiling_fps,
                  "host_write_ceiling_note": "plain cudaMemcpyAsync of one BGRA8-sized frame (w*h*4 B) per step to pinned host memory on every rank at once: what the box's host memory takes"}
        if sb is not None:
            config.update({"sort_bench_gb_per_s": sb["gb_per_s"], "sort_bench_frac_measured": sb["frac_of_hbm_peak"],
                           "sort_bench_frac_8TBs": sb["frac_of_8TBs_nominal"], "sort_bench_n_pairs": sb["n_pairs"], "sort_bench_ms": sb["ms"]})
        if c4 is not None:
            config["c4"] = c4
        if c5 is not None:
            config["c5"] = c5
        if c3 is not None:
            config["c3"] = c3
            config["band_frame_equals_single_gpu"] = c3["band_frame_equals_single_gpu"]
        line = {
            "metric": "frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32+f16", "data": "synthetic",
            "mgaussians_per_s": n * fps / 1e6, "ms_per_step_percentiles": step_pct,
            "config": config, "roofline": roof, "stages": stages,
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": 128, "d2h_bytes_per_step": w * h * 8 + 24,
                    "steps": e2e_steps, "single_frame_ms": lat_ms,
                    "note": "splat_render_batch_host: %d frames, each frame's matrices from host and its framebuffer to pinned "
                            "host memory, copies pipelined behind the next frame's kernels" % e2e_steps,
                    "bgra8": {"value": e2e_bgra_fps, "d2h_bytes_per_step": w * h * 4 + 24,
                              "note": "same call with SPLAT_FLAG_HOST_BGRA8: frames delivered as the reference drawable's BGRA8Unorm"}},
            "camera_batch_device": inflight,
            "gpu_launches": args.steps * 7,      # block cull, preprocess, its tail launch (present when the hinted grid is smaller than the scene), tile offsets, scatter, tile sort, blend (+1 memset node, not counted)
            "clocks": clocks, "wall_s": t_wall, "numa": numa,
        }
        if sb is not None:
            line["sort_bench"] = sb
        if cpu_leg is not None:
            line["cpu_baseline"] = cpu_leg
        emit(line)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
