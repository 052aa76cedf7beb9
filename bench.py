#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native pathtrace sample loop.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched under torch.distributed.run)
  python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): the reference's default scene (src/world.rs:65-241) at 1920x1080,
1024 spp, MIS, Russian roulette from depth 4, soft max depth 50.  A "step" renders that whole frame once.
metric = Msamples/s = W*H*spp / time.  N GPUs split the samples per pixel (disjoint Philox sample indices),
one NCCL reduce(sum) of the W*H float4 buffer to rank 0, resolve on rank 0 — total work fixed => "strong".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
# stdout carries exactly one JSON line: file descriptor 1 is pointed at stderr for the whole run (NCCL prints its version
# banner with the C runtime, past sys.stdout and past NCCL_DEBUG_FILE) and the line is written to the saved descriptor.
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(line: dict):
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


import numpy as np  # noqa: E402

WIDTH, HEIGHT, SPP = 1920, 1080, 1024
WORKLOAD = "default Cornell scene (13 objects, world.rs:65-241) 1920x1080 1024spp MIS max_depth=50"


# ------------------------------------------------------------------------------------------------ helpers
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index=0):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2])); power.append(float(r[3]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "MEASURED_PEAKS.json"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` on this workload, from the committed
    `ncu --set full` capture (profiles/roofline_traffic.json; null when no capture of this kernel exists)."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(kernel, {}).get("bytes_per_launch")
    except (OSError, ValueError):
        return None


def default_scene_arrays():
    from pathtrace_b200 import scenes
    return scenes.default_scene(WIDTH, HEIGHT)


def oracle_cam(O, cam):
    c = O.CameraDesc()
    for name in ("origin", "lower_left_corner", "horizontal", "vertical"):
        v = getattr(cam, name)
        getattr(c, name)[:] = [v.x, v.y, v.z]
    c.width, c.height = cam.width, cam.height
    return c


# ------------------------------------------------------------------------------------------ other configurations
def _frame(api, dist, torch, ctx, scene, cam, spp, rank, world, ext, dev, reps=2, **kw):
    """One frame of `spp` samples per pixel split over the ranks (spp split + one reduce + resolve on rank 0), device
    resident, CUDA events on the library stream, max over ranks; best of `reps`.  Returns (ms, stats of this rank)."""
    from pathtrace_b200.distributed import split_spp
    W, H = cam.width, cam.height
    spp_g, off_g = split_spp(spp, world, rank)
    with torch.cuda.stream(ext):
        accum = torch.empty(W * H, 4, device=dev, dtype=torch.float32)
        rgba = torch.empty(W * H, 4, device=dev, dtype=torch.uint8)
    best, st_best = None, None
    for _ in range(reps + 1):                      # first pass = warm-up
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = None
        with torch.cuda.stream(ext):
            e0.record(ext)
            accum.zero_()
            if spp_g:
                st = api.render_accumulate_device(ctx, scene, cam, api.render_params(spp=spp_g, sample_offset=off_g, **kw), accum.data_ptr())
            if world > 1:
                dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                api.resolve_device(ctx, accum.data_ptr(), W, H, spp, 0, rgba.data_ptr())
            e1.record(ext)
        ext.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if best is None or ms.item() < best:
            best, st_best = float(ms.item()), st
    return best, st_best


def _sum_over_ranks(torch, dist, dev, world, st, keys):
    v = torch.tensor([float(st[k]) if st else 0.0 for k in keys], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(v, op=dist.ReduceOp.SUM)
    return dict(zip(keys, v.tolist()))


def other_configs(args, rank, world, ctx, ext, dev):
    """BASELINE.json configs[2..4] next to the headline (configs[1]): each at its own resolution and object count with spp
    reduced so that the whole block stays under a minute.  Not bench lines of their own — sub-records of the one JSON line."""
    import torch
    import torch.distributed as dist
    from pathtrace_b200 import api, scenes
    peaks, peaks_src = measured_peaks()
    keys = ("samples", "rays_primary", "rays_shadow", "rays_scattered", "path_vertices", "kernel_launches", "nodes_visited", "prim_tests")
    out = {}

    # ---- C1: the reference's own default workload, world.rs:16-18: 400x400 at 3000 spp (full size: 4.8e8 samples)
    sh, mt, cam = scenes.default_scene(400, 400)
    sc = api.Scene(ctx, sh, mt)
    ms, st = _frame(api, dist, torch, ctx, sc, cam, 3000, rank, world, ext, dev, reps=1)
    tot = _sum_over_ranks(torch, dist, dev, world, st, keys)
    out["C1"] = {"workload": "default Cornell scene 400x400 3000spp MIS (WIDTH/HEIGHT/SAMPLE_NUM of world.rs:16-18, full size)", "n_gpus": world,
                 "value": 400 * 400 * 3000 / ms / 1e3, "unit": "Msamples/s", "ms": ms, "kernel_launches": int(tot["kernel_launches"]),
                 "mrays_per_s": (tot["rays_primary"] + tot["rays_shadow"] + tot["rays_scattered"]) / ms / 1e3}
    sc.close()

    # ---- C3: 495-object random spheres, 3840x2160 (formulation B, LBVH in shared memory)
    sh, mt, cam = scenes.random_spheres_scene(3840, 2160)
    sc = api.Scene(ctx, sh, mt)
    spp = args.c3_spp
    ms, st = _frame(api, dist, torch, ctx, sc, cam, spp, rank, world, ext, dev)
    tot = _sum_over_ranks(torch, dist, dev, world, st, keys)
    _, stc = _frame(api, dist, torch, ctx, sc, cam, max(world, 4), rank, world, ext, dev, reps=0, counters=True)   # node / primitive counters (slower pass)
    cnt = _sum_over_ranks(torch, dist, dev, world, stc, keys)
    rays_c = max(cnt["rays_primary"] + cnt["rays_shadow"] + cnt["rays_scattered"], 1.0)
    rays = tot["rays_primary"] + tot["rays_shadow"] + tot["rays_scattered"]
    # algorithmic flops of a BVH scene: nodes visited x 2 box tests x 18 + primitive tests x the oracle's per-test
    # cost + shading per vertex; per-ray averages from the counter pass
    fl_node, fl_prim, fl_shade = 36.0, 40.0, 320.0
    flops = rays * (cnt["nodes_visited"] / rays_c * fl_node + cnt["prim_tests"] / rays_c * fl_prim) + tot["path_vertices"] * fl_shade
    fp32_nominal = 148 * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12
    out["C3"] = {"workload": f"random spheres ({len(sh)} objects) 3840x2160 {spp}spp MIS (of the config's 4096 spp)", "n_gpus": world,
                 "value": 3840 * 2160 * spp / ms / 1e3, "unit": "Msamples/s", "mrays_per_s": rays / ms / 1e3, "ms": ms,
                 "kernel_launches": int(tot["kernel_launches"]), "lbvh_build_ms": sc.build_ms,
                 "nodes_per_ray": cnt["nodes_visited"] / rays_c, "prim_tests_per_ray": cnt["prim_tests"] / rays_c,
                 "roofline": {"bound": "fp32", "achieved": flops / ms / 1e9, "peak": fp32_nominal, "unit": "TFLOP/s", "frac": flops / ms / 1e9 / fp32_nominal,
                              "peak_source": "nominal 148 SM x 128 lanes x 2 x max SM clock", "flops_per_unit": {"node_visit": fl_node, "prim_test": fl_prim, "vertex": fl_shade}}}
    sc.close()

    # ---- C4: 1,002,530-triangle heightfield at 1920x1080 (formulation B, LBVH walked through L1/L2): HBM roofline
    sh, mt, cam = scenes.heightfield_scene(1920, 1080)
    sc = api.Scene(ctx, sh, mt)
    spp = args.c4_spp
    ms, st = _frame(api, dist, torch, ctx, sc, cam, spp, rank, world, ext, dev)
    tot = _sum_over_ranks(torch, dist, dev, world, st, keys)
    msc, stc = _frame(api, dist, torch, ctx, sc, cam, max(world, 4), rank, world, ext, dev, reps=0, counters=True)
    cnt = _sum_over_ranks(torch, dist, dev, world, stc, keys)
    rays_c = max(cnt["rays_primary"] + cnt["rays_shadow"] + cnt["rays_scattered"], 1.0)
    rays = tot["rays_primary"] + tot["rays_shadow"] + tot["rays_scattered"]
    node_b, prim_b = 64.0, 48.0                      # bytes per node visit / per triangle test
    alg_bytes = rays * (cnt["nodes_visited"] / rays_c * node_b + cnt["prim_tests"] / rays_c * prim_b)
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            import hashlib
            from oracle import pyoracle as O
            O.build()
            got = sc.bvh()
            ref = O.lbvh_build(sh)
            h = lambda b: hashlib.sha256(b["nodes"].tobytes() + b["prims"].tobytes() + b["codes"].tobytes()).hexdigest()[:16]
            _, _, cam64 = scenes.heightfield_scene(64, 64)
            ids, t, _ = api.trace_primary(ctx, sc, cam64, api.render_params(traversal="bvh"))
            rid, rt, _ = O.trace_primary(oracle_cam(O, cam64), sh, mt, precision="f32")
            hit = rid >= 0
            parity = {"lbvh_hash_gpu": h(got), "lbvh_hash_cpu_twin": h(ref), "lbvh_bit_exact": h(got) == h(ref),
                      "sampled_primary_rays": int(ids.size), "id_mismatches_vs_fp32_oracle_brute_force": int((ids != rid).sum()),
                      "t_bit_exact": bool(np.array_equal(t[hit].view(np.uint32), rt[hit].astype(np.float32).view(np.uint32)))}
        except Exception as e:
            parity = {"error": str(e)}
    out["C4"] = {"workload": f"heightfield ({len(sh)} triangles) 1920x1080 {spp}spp MIS", "n_gpus": world,
                 "value": 1920 * 1080 * spp / ms / 1e3, "unit": "Msamples/s", "mrays_per_s": rays / ms / 1e3, "ms": ms,
                 "kernel_launches": int(tot["kernel_launches"]), "lbvh_build_ms": sc.build_ms, "lbvh_build_launches": sc.build_launches,
                 "nodes_per_ray": cnt["nodes_visited"] / rays_c, "prim_tests_per_ray": cnt["prim_tests"] / rays_c,
                 "roofline": {"bound": "hbm", "achieved": alg_bytes / ms / 1e6, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                              "frac": alg_bytes / ms / 1e6 / peaks["hbm_gbs"], "peak_source": peaks_src,
                              "bytes_per_unit": {"node_visit": node_b, "prim_test": prim_b},
                              "note": "algorithmic bytes = nodes visited x 64 + primitive tests x 48 from device counters (PT_FLAG_COUNTERS pass); "
                                      "dram__bytes of the walk kernels: profiles/"},
                 "parity": parity}
    sc.close()

    # ---- C5: convergence sweep points on the default scene at 1920x1080 (spp split over the ranks + one reduce each)
    sh, mt, cam = scenes.default_scene(1920, 1080)
    sc = api.Scene(ctx, sh, mt)
    pts = []
    for spp in (1, 16, 256, 4096):
        ms, st = _frame(api, dist, torch, ctx, sc, cam, spp, rank, world, ext, dev, reps=2 if spp < 4096 else 1)
        pts.append({"spp": spp, "ms": ms, "value": 1920 * 1080 * spp / ms / 1e3, "unit": "Msamples/s"})
    out["C5"] = {"workload": "default scene 1920x1080, spp sweep (render + reduce + resolve per frame)", "n_gpus": world, "points": pts}
    sc.close()
    return out


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args, rank):
    """The reference's own CPU implementation of the path, restated (oracle/: f64, brute-force hit_scene, literal
    recursive MIS incl. the repeated look-ahead trace, one task per pixel over all host threads).  The Rust
    reference itself cannot be built on this image (no rustc/cargo; needs a window) — DESIGN.md."""
    if rank != 0:
        return
    from oracle import pyoracle as O
    O.build()
    sh, mt, cam = default_scene_arrays()
    oc = oracle_cam(O, cam)
    spp_step = args.ref_spp
    threads = os.cpu_count() or 1          # torchrun exports OMP_NUM_THREADS=1: ask for every host core explicitly
    samples = WIDTH * HEIGHT * spp_step
    for w in range(args.warmup):
        O.render(oc, sh, mt, O.RenderParams(spp=1, sample_offset=w), "f64", recursive=True, nthreads=threads)
    t0 = time.perf_counter()
    rays = 0
    for k in range(args.steps):
        r = O.render(oc, sh, mt, O.RenderParams(spp=spp_step, sample_offset=1000 + k * spp_step), "f64", recursive=True, nthreads=threads)
        c = r["counters"]
        rays += c["primary"] + c["shadow"] + c["scattered"]
    dt = time.perf_counter() - t0
    val = samples * args.steps / dt / 1e6
    sample = f"{WIDTH}x{HEIGHT} at {spp_step} spp per step (of the workload's {SPP}); samples/s is spp-independent"
    emit(({
        "impl": "reference", "metric": "Msamples/s", "value": val, "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": {"workload": WORKLOAD, "step_sample": sample},
        "mrays_per_s": rays / dt / 1e6,
        "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args, rank, world, local_rank):
    import gc
    import torch
    import torch.distributed as dist
    from pathtrace_b200 import api

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = api.Context(local_rank)
    sh, mt, cam = default_scene_arrays()
    scene = api.Scene(ctx, sh, mt)
    try:
        _bench_ours(args, rank, world, local_rank, ctx, scene, sh, mt, cam)
    finally:
        # every tensor that touched the library's stream must be gone before that stream is destroyed
        gc.collect()
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        scene.close()
        ctx.close()


def _bench_ours(args, rank, world, local_rank, ctx, scene, sh, mt, cam):
    import torch
    import torch.distributed as dist
    from pathtrace_b200 import api

    dev = torch.device("cuda", local_rank)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    npix = WIDTH * HEIGHT
    spp_total = args.spp
    from pathtrace_b200.distributed import split_spp
    spp_g, off_g = split_spp(spp_total, world, rank)
    accum = torch.zeros(npix, 4, device=dev, dtype=torch.float32)
    radiance = torch.empty(npix, 3, device=dev, dtype=torch.float32)
    rgba = torch.empty(npix, 4, device=dev, dtype=torch.uint8)
    h_radiance = torch.empty(npix, 3, dtype=torch.float32).pin_memory()
    h_rgba = torch.empty(npix, 4, dtype=torch.uint8).pin_memory()
    flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)   # > 126 MB L2
    params = api.render_params(spp=spp_g, sample_offset=off_g, path_pool=args.pool, profile=args.profile_in_timed, hbm_queues=args.hbm_queues)
    totals = {"launches": 0, "rays": 0, "samples": 0, "stats": []}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step(sc):
        """inputs resident in HBM: zero sum buffer, render this rank's spp share, reduce to rank 0, resolve."""
        with torch.cuda.stream(ext):
            flush.zero_()            # L2 flush between timed iterations
            accum.zero_()
            st = api.render_accumulate_device(ctx, sc, cam, params, accum.data_ptr())
            if world > 1:
                dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                api.resolve_device(ctx, accum.data_ptr(), WIDTH, HEIGHT, spp_total, radiance.data_ptr(), rgba.data_ptr())
        totals["launches"] += st["kernel_launches"] + 3 + (1 if rank == 0 else 0)
        totals["rays"] += st["rays_primary"] + st["rays_shadow"] + st["rays_scattered"]
        totals["samples"] += st["samples"]
        totals["stats"].append(st)
        return st

    # ---- value: device-resident timing
    for _ in range(args.warmup):
        device_step(scene)
    totals.update(launches=0, rays=0, samples=0, stats=[])
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(args.steps):
        device_step(scene)
    e1.record(ext)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    agg = torch.tensor([totals["launches"], totals["rays"], totals["samples"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
    ms_total = float(ms.item())
    launches, rays, samples = (float(x) for x in agg.tolist())
    value = npix * spp_total * args.steps / (ms_total * 1e-3) / 1e6
    dev_stats = list(totals["stats"])

    # ---- e2e: host buffers in, host buffers out, through the public API
    h2d = sh.nbytes + mt.nbytes + 56 + 48
    d2h = h_radiance.numel() * 4 + h_rgba.numel()

    def e2e_step():
        sc = api.Scene(ctx, sh, mt)               # host arrays -> HBM + LBVH build
        if world == 1:
            api.render(ctx, sc, cam, params)      # pt_render: host radiance + RGBA8 buffers (D2H inside)
        else:
            device_step(sc)
            if rank == 0:
                with torch.cuda.stream(ext):
                    h_radiance.copy_(radiance, non_blocking=True)
                    h_rgba.copy_(rgba, non_blocking=True)
                ext.synchronize()
        sc.close()

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = npix * spp_total * args.steps / float(dt.item()) / 1e6

    # ---- roofline of the dominant kernel (rank 0): per-kernel CUDA-event times + algorithmic flops from the oracle
    roofline = roofline_hbm = cpu_baseline = None
    kernel_ms = None
    if rank == 0:
        peaks, peaks_src = measured_peaks()
        if args.profile_in_timed:
            prof_stats = dev_stats
        else:   # same workload, PT_FLAG_PROFILE on: events around every launch (outside the timed region)
            pp = api.render_params(spp=spp_g, sample_offset=off_g, path_pool=args.pool, profile=True, hbm_queues=args.hbm_queues)
            with torch.cuda.stream(ext):
                accum.zero_()
                prof_stats = [api.render_accumulate_device(ctx, scene, cam, pp, accum.data_ptr())]
                # per-phase cycle shares come from the instrumented kernel variant in a pass of its own
                pc = api.render_params(spp=max(spp_g // 8, 1), sample_offset=off_g, path_pool=args.pool, phase_clocks=True, hbm_queues=args.hbm_queues)
                prof_stats[0]["phase_share"] = api.render_accumulate_device(ctx, scene, cam, pc, accum.data_ptr())["phase_share"]
        kernel_ms = {k: sum(s["ms_kernel"][k] for s in prof_stats) for k in api.KERNEL_CLASSES}
        n_kernel = {k: sum(s["n_kernel"][k] for s in prof_stats) for k in api.KERNEL_CLASSES}
        fp32_tf, _ = api.measure_fp32_peak(ctx)
        fp32_nominal = 148 * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12
        try:
            from oracle import pyoracle as O
            O.build()
            fl = O.count_flops(oracle_cam(O, cam), sh, mt, O.RenderParams(spp=1), stride=6)
            r_closest = sum(s["rays_primary"] + s["rays_scattered"] for s in prof_stats)
            r_shadow = sum(s["rays_shadow"] for s in prof_stats)
            verts = sum(s["path_vertices"] for s in prof_stats)
            flops = {"trace_closest": fl["flops_closest_per_ray"] * r_closest, "trace_shadow": fl["flops_shadow_per_ray"] * r_shadow,
                     "shade": fl["flops_shade_per_vertex"] * verts, "generate": 0.0}
            flops["smem_wavefront"] = flops["trace_closest"] + flops["trace_shadow"] + flops["shade"]   # all stages in one launch
            dom = max(kernel_ms, key=kernel_ms.get)
            ach = flops[dom] / (kernel_ms[dom] * 1e-3) / 1e12
            whole = fl["flops_per_sample"] * value * 1e6 / 1e12
            roofline = {"bound": "fp32", "kernel": "k_" + dom, "achieved": ach, "peak": fp32_tf, "unit": "TFLOP/s", "frac": ach / fp32_tf,
                        "traffic": ncu_traffic("k_" + dom), "peak_source": "in-run FFMA microbenchmark (pt_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 entry",
                        "peak_nominal": fp32_nominal, "frac_of_nominal": ach / fp32_nominal,
                        "launches": n_kernel[dom], "avg_launch_ms": kernel_ms[dom] / max(n_kernel[dom], 1),
                        "algorithmic_flops_per_unit": {"closest_ray": fl["flops_closest_per_ray"], "shadow_ray": fl["flops_shadow_per_ray"],
                                                       "shade_vertex": fl["flops_shade_per_vertex"], "sample": fl["flops_per_sample"]},
                        "units_per_launch": {"closest_rays": r_closest / max(n_kernel[dom], 1), "shadow_rays": r_shadow / max(n_kernel[dom], 1),
                                             "vertices": verts / max(n_kernel[dom], 1)},
                        "phase_share": prof_stats[0].get("phase_share"),
                        "kernel_share_of_step": kernel_ms[dom] / max(sum(kernel_ms.values()), 1e-9),
                        "whole_step": {"flops_per_sample": fl["flops_per_sample"], "achieved": whole, "frac": whole / fp32_tf},
                        "note": "algorithmic flops counted from the reference's arithmetic by the oracle's counting scalar (FMA=2, no re-trace)"}
            # queue traffic (implementation overhead, reported separately; never added to roofline.achieved)
            t_all = sum(kernel_ms.values()) * 1e-3
            if dom == "smem_wavefront":
                # queues live in shared memory: HBM sees only the float4 accumulate atomics (16 B each, L2-resident frame)
                qb = 16.0 * (r_shadow + sum(s["samples"] for s in prof_stats))
                roofline_hbm = {"bound": "hbm", "what": "upper bound on accumulate traffic: one 16-byte float4 atomic per finished path and per shadow ray; "
                                "ray / hit / path-state / shadow queues are in shared memory and never reach HBM", "achieved": qb / t_all / 1e9,
                                "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": qb / t_all / 1e9 / peaks["hbm_gbs"], "peak_source": peaks_src}
            else:
                qb = 56.0 * r_closest + 200.0 * verts + 64.0 * r_shadow
                roofline_hbm = {"bound": "hbm", "what": "wavefront queue traffic (path state, hits, shadow queue)", "achieved": qb / t_all / 1e9,
                                "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": qb / t_all / 1e9 / peaks["hbm_gbs"], "peak_source": peaks_src,
                                "bytes_per_ray_closest": 56, "bytes_per_vertex_shade": 200, "bytes_per_shadow_ray": 64}
            # ---- CPU baseline: the oracle (port of the reference's CPU path) on this box's host cores, bounded sample
            if world == 1 and not args.no_cpu_baseline:
                oc = oracle_cam(O, cam)
                ncpu = os.cpu_count() or 1
                O.render(oc, sh, mt, O.RenderParams(spp=1), "f64", recursive=True, nthreads=ncpu)
                r = O.render(oc, sh, mt, O.RenderParams(spp=args.cpu_spp, sample_offset=7), "f64", recursive=True, nthreads=ncpu)
                cpu_baseline = {"value": npix * args.cpu_spp / r["seconds"] / 1e6, "unit": "Msamples/s", "cores": ncpu,
                                "kind": "port", "sample": f"{WIDTH}x{HEIGHT} at {args.cpu_spp} spp (of {SPP}); f64, brute-force scan, recursive MIS with re-trace, OpenMP dynamic per pixel; g++ -O3 -march=x86-64-v3 -ffp-contract=off",
                                "seconds": r["seconds"]}
        except Exception as e:   # the oracle is a checker; the bench line survives without it
            roofline = {"bound": "fp32", "error": f"oracle unavailable: {e}", "peak": fp32_tf, "unit": "TFLOP/s"}

    configs = None
    if not args.no_configs:
        try:
            configs = other_configs(args, rank, world, ctx, ext, dev)
        except Exception as e:           # the headline line survives a failing side configuration
            configs = {"error": repr(e)}
    if rank == 0:
        st0 = dev_stats[0]
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "width": WIDTH, "height": HEIGHT, "spp": spp_total, "spp_per_gpu": spp_g, "strategy": "mis",
                       "parallelism": f"spp-split x{world} + 1 NCCL reduce(sum)", "path_pool": st0["path_pool"],
                       "formulation": "hbm-queue wavefront" if args.hbm_queues else "shared-memory wavefront (one persistent launch)",
                       "scan_mode": st0.get("scan_mode"),
                       "l2": "explicit 256 MiB write between steps"},
            "mrays_per_s": rays / (ms_total * 1e-3) / 1e6, "rays_per_sample": rays / max(samples, 1),
            "gpu_launches": int(launches), "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "what": "pt_scene_create(host arrays)+pt_render(host radiance+RGBA8 buffers)" if world == 1 else
                            "scene upload per rank + render + NCCL reduce + resolve + D2H on rank 0"},
            "roofline": roofline, "roofline_hbm": roofline_hbm, "kernel_ms": kernel_ms, "cpu_baseline": cpu_baseline,
            "configs": configs,
        }
        lib_override = os.environ.get("PT_LIB")
        if lib_override:
            line["config"]["lib_override"] = lib_override
        emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--spp", type=int, default=SPP)
    ap.add_argument("--pool", type=int, default=0, help="wavefront path pool (0 = library default)")
    ap.add_argument("--cpu-spp", type=int, default=16, help="spp of the bounded CPU-baseline sample")
    ap.add_argument("--ref-spp", type=int, default=2, help="spp per step of the --impl reference arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C3 / C4 / C5 sub-records")
    ap.add_argument("--c3-spp", type=int, default=256, help="C3 sub-record: samples per pixel of the 3840x2160 frame (the reduce of its 133 MB sum buffer is a fixed ~6 ms per frame at N > 1)")
    ap.add_argument("--c4-spp", type=int, default=16)
    ap.add_argument("--hbm-queues", action="store_true", help="force the HBM-queue wavefront (one launch per stage per bounce)")
    ap.add_argument("--profile-in-timed", action="store_true", help="record per-kernel events inside the timed steps")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
