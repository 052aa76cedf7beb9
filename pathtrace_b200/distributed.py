"""Multi-GPU render: split the samples per pixel across ranks, ONE reduce of the sum buffer.

One process per GPU (torchrun).  Rank g renders the full frame for the global sample indices
[offset_g, offset_g + spp_g); the Philox counters are keyed by the global sample index, so the union over ranks is
exactly the single-GPU sample set (samples are independent: src/world.rs:296-313 just sums them).  The per-pixel
float4 sums are combined with one ``torch.distributed.reduce(SUM)`` to rank 0 (NCCL over NVLink on GPUs; gloo in the
CPU tests), and rank 0 resolves: mean, sqrt gamma, clamp, RGBA8 (src/world.rs:315-332).

There is no CPU render path here: ``render_distributed`` needs the CUDA library.  ``split_spp`` and ``reduce_sum`` are
backend-agnostic host logic and are what tests/test_distributed_cpu.py exercises under gloo.
"""
from __future__ import annotations

from typing import Optional, Tuple


def split_spp(spp: int, world: int, rank: int) -> Tuple[int, int]:
    """(spp_g, offset_g): rank g gets floor(spp/G) samples, the first spp % G ranks one more; offsets are contiguous."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError(f"bad split: spp={spp} world={world} rank={rank}")
    base, rem = divmod(spp, world)
    return base + (1 if rank < rem else 0), rank * base + min(rank, rem)


def reduce_sum(accum, dst: int = 0, group=None):
    """In-place SUM-reduce of the per-pixel sum buffer to rank ``dst`` (a no-op without an initialised process group)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def render_distributed(ctx, scene, cam, spp: int, *, rank: int, world: int, params_kw: Optional[dict] = None, group=None,
                       want_host: bool = True):
    """Render ``spp`` samples per pixel over ``world`` ranks.  Returns (radiance [H,W,3] f32, rgba [H,W,4] u8, stats) on
    rank 0 and (None, None, stats) elsewhere.  ``ctx`` / ``scene`` are this rank's api.Context / api.Scene."""
    import torch
    from . import api

    W, H = cam.width, cam.height
    dev = torch.device("cuda", ctx.device)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    spp_g, off_g = split_spp(spp, world, rank)
    stats = None
    # the library's stream is non-blocking (no implicit ordering against torch's default stream): allocate and clear the sum
    # buffer on the stream that renders into it
    with torch.cuda.stream(stream):
        accum = torch.empty(W * H, 4, device=dev, dtype=torch.float32)
        accum.zero_()
        if spp_g > 0:
            p = api.render_params(spp=spp_g, sample_offset=off_g, **(params_kw or {}))
            stats = api.render_accumulate_device(ctx, scene, cam, p, accum.data_ptr())
        reduce_sum(accum, 0, group)
        if rank != 0:
            stream.synchronize()          # accum is released on return: the reduce must have consumed it
            return None, None, stats
        radiance = torch.empty(W * H, 3, device=dev, dtype=torch.float32)
        rgba = torch.empty(W * H, 4, device=dev, dtype=torch.uint8)
        api.resolve_device(ctx, accum.data_ptr(), W, H, spp, radiance.data_ptr(), rgba.data_ptr())
    stream.synchronize()
    if not want_host:
        return radiance.view(H, W, 3), rgba.view(H, W, 4), stats
    return radiance.view(H, W, 3).cpu().numpy(), rgba.view(H, W, 4).cpu().numpy(), stats
