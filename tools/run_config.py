"""Runs one of the configurations of SURVEY §8(d) that are not the bench line (C3, C4, C5) and prints JSON lines.

  python tools/run_config.py --config C3 [--spp 4096]                 one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
         tools/run_config.py --config C5                               spp split over the ranks, one NCCL reduce per frame

  C3  random-spheres scene (495 objects), 3840x2160, 4096 spp
  C4  1 002 530-triangle heightfield, 1920x1080, 64 spp (LBVH build timed separately)
  C5  default scene, spp = 1, 2, 4, ..., 16384: time per frame at 1920x1080 (render, reduce, resolve) and the RMSE of a
      160x160 render against the committed f64 reference image (tests/golden/cornell_160x160_f64.npz)

Times are CUDA events on the library's stream (torch.cuda.ExternalStream), maximum over ranks.  Nothing here reads the CPU
oracle; the RMSE reference is a committed fixture.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["C3", "C4", "C5"])
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--max-spp", type=int, default=16384)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    from pathtrace_b200 import api, scenes
    from pathtrace_b200.distributed import split_spp

    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ctx = api.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def frame(scene, cam, spp, want_image=False):
        """One distributed frame; returns (ms_render, ms_reduce, ms_resolve, stats, radiance or None), ms = max over ranks."""
        W, H = cam.width, cam.height
        spp_g, off_g = split_spp(spp, world, rank)
        accum = torch.zeros(W * H, 4, device=dev, dtype=torch.float32)
        radiance = torch.empty(W * H, 3, device=dev, dtype=torch.float32)
        rgba = torch.empty(W * H, 4, device=dev, dtype=torch.uint8)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        stats = None
        with torch.cuda.stream(stream):
            ev[0].record()
            if spp_g > 0:
                stats = api.render_accumulate_device(ctx, scene, cam, api.render_params(spp=spp_g, sample_offset=off_g), accum.data_ptr())
            ev[1].record()
            if world > 1:
                dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            ev[2].record()
            if rank == 0:
                api.resolve_device(ctx, accum.data_ptr(), W, H, spp, radiance.data_ptr(), rgba.data_ptr())
            ev[3].record()
        stream.synchronize()
        torch.cuda.synchronize()
        ms = torch.tensor([ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])], device=dev)
        cnt = torch.tensor([float(stats[k]) if stats else 0.0 for k in ("rays_primary", "rays_shadow", "rays_scattered")], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        img = radiance.view(H, W, 3).cpu().numpy() if (want_image and rank == 0) else None
        return ms.tolist(), float(cnt.sum().item()), stats, img

    def emit(d):
        if rank == 0:
            print(json.dumps(d), flush=True)

    if a.config in ("C3", "C4"):
        W, H = (a.width or (3840 if a.config == "C3" else 1920)), (a.height or (2160 if a.config == "C3" else 1080))
        spp = a.spp or (4096 if a.config == "C3" else 64)
        sh, mt, cam = (scenes.random_spheres_scene if a.config == "C3" else scenes.heightfield_scene)(W, H)
        t0 = time.time(); scene = api.Scene(ctx, sh, mt); build_s = time.time() - t0
        frame(scene, cam, max(world, min(spp, 4 * world)))                 # warm-up
        best = None
        for _ in range(a.reps):
            ms, rays, st, _ = frame(scene, cam, spp)
            if best is None or sum(ms) < sum(best[0]):
                best = (ms, rays, st)
        ms, rays, st = best
        samples = W * H * spp
        emit({"config": a.config, "n_gpus": world, "objects": len(sh), "width": W, "height": H, "spp": spp,
              "msamples_per_s": samples / sum(ms) / 1e3, "mrays_per_s": rays / sum(ms) / 1e3, "ms_render": ms[0], "ms_reduce": ms[1],
              "ms_resolve": ms[2], "scene_create_s_rank0": build_s, "lbvh_build_ms": scene.build_ms if hasattr(scene, "build_ms") else None,
              "formulation": "smem" if (st and st["n_kernel"].get("smem_wavefront")) else "hbm-queues",
              "kernel_launches": st["kernel_launches"] if st else 0})
        scene.close()
    else:
        g = np.load(os.path.join(ROOT, "tests", "golden", "cornell_160x160_f64.npz"))
        ref = g["mean"].astype(np.float64)
        ref_se = float(np.sqrt(g["var_of_mean"].astype(np.float64).mean()))
        sh, mt, cam_big = scenes.default_scene(a.width or 1920, a.height or 1080)
        _, _, cam_small = scenes.default_scene(160, 160)
        scene = api.Scene(ctx, sh, mt)
        frame(scene, cam_big, 8 * world)                                 # warm-up
        spp = 1
        while spp <= a.max_spp:
            best = None
            for _ in range(a.reps if spp >= 64 else a.reps + 2):
                ms, rays, st, _ = frame(scene, cam_big, spp)
                if best is None or sum(ms) < sum(best[0]):
                    best = (ms, rays)
            ms, rays = best
            _, _, _, img = frame(scene, cam_small, spp, want_image=True)
            rec = {"config": "C5", "n_gpus": world, "spp": spp, "width": cam_big.width, "height": cam_big.height,
                   "ms_render": ms[0], "ms_reduce": ms[1], "ms_resolve": ms[2],
                   "msamples_per_s": cam_big.width * cam_big.height * spp / sum(ms) / 1e3, "mrays_per_s": rays / sum(ms) / 1e3}
            if rank == 0:
                d = img.astype(np.float64) - ref
                rec["rmse_160x160_rgb"] = [float(x) for x in np.sqrt((d ** 2).mean(axis=(0, 1)))]
                rec["reference"] = {"spp": int(g["spp"]), "mean_standard_error": ref_se}
            emit(rec)
            spp *= 2
        scene.close()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
