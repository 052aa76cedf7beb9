"""pt_render_device_noalloc against pt_render (run on a B200 with -m gpu; collected last on purpose).

The entry point has no counterpart in the reference: it is pt_render with the scratch memory (frame sum buffer, work counter,
per-CTA statistics) in a static device array of the library, for callers that cannot call cudaMalloc.  Same samples, same
kernels, so its frame equals pt_render's within the tolerance the suite uses for two renders of the same samples
(test_gpu_outputs.py: rtol 2e-5 / atol 1e-6 — float atomics commute, they do not associate), on two calls in a row that share
the array: the second call uses another frame size, scene and strategy, so anything left over from the first would show.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL, ATOL = 2e-5, 1e-6


def _device_outputs(gpu_ctx, H, W):
    import torch
    dev = torch.device("cuda", gpu_ctx.device)
    # filled with garbage: every pixel must be written by the call
    d_rad = torch.full((H, W, 3), float("nan"), device=dev, dtype=torch.float32)
    d_rgba = torch.full((H, W, 4), 7, device=dev, dtype=torch.uint8)
    torch.cuda.synchronize(dev)
    return d_rad, d_rgba


def _check_against_render(gpu_ctx, sc, cam, params):
    from pathtrace_b200 import api
    H, W = cam.height, cam.width
    d_rad, d_rgba = _device_outputs(gpu_ctx, H, W)
    api.render_device_noalloc(gpu_ctx, sc, cam, params, d_rad.data_ptr(), d_rgba.data_ptr())
    gpu_ctx.synchronize()
    rad, rgba = d_rad.cpu().numpy(), d_rgba.cpu().numpy()
    want_rad, want_rgba, st = api.render(gpu_ctx, sc, cam, params)
    assert st["n_kernel"]["smem_wavefront"] == 1, st      # the shared-memory wavefront served pt_render too
    assert np.isfinite(rad).all()
    assert np.allclose(rad, want_rad, rtol=RTOL, atol=ATOL), float(np.abs(rad - want_rad).max())
    assert (rgba[..., 3] == 255).all()
    assert np.abs(rgba.astype(np.int16) - want_rgba.astype(np.int16)).max() <= 1    # a sum that differs in the last bit may cross a truncation boundary
    assert (rgba == want_rgba).mean() > 0.999
    assert len(np.unique(rgba[..., :3])) > 50                                       # a real image
    return rad


def test_noalloc_equals_render_on_two_calls_in_a_row(gpu_ctx):
    import torch
    from pathtrace_b200 import api, scenes
    dev = torch.device("cuda", gpu_ctx.device)
    # call 1: the default scene at the largest frame the scratch holds (256 x 256 == PT_NOALLOC_MAX_PIXELS), MIS
    sh, mt, cam = scenes.default_scene(256, 256)
    assert cam.width * cam.height == api.NOALLOC_MAX_PIXELS
    sc = api.Scene(gpu_ctx, sh, mt)
    first = _check_against_render(gpu_ctx, sc, cam, api.render_params(spp=32))
    # call 2, straight after: smaller frame, other scene, BRDF-only sampling, other sample range
    sh2, mt2, cam2 = scenes.material_zoo_scene(96, 64)
    sc2 = api.Scene(gpu_ctx, sh2, mt2)
    _check_against_render(gpu_ctx, sc2, cam2, api.render_params(spp=24, sample_offset=5, strategy="brdf_only"))
    # and the first frame again: equal to its first rendering, and the call took no device memory
    d_rad, d_rgba = _device_outputs(gpu_ctx, 256, 256)
    free_before, _ = torch.cuda.mem_get_info(dev)
    api.render_device_noalloc(gpu_ctx, sc, cam, api.render_params(spp=32), d_rad.data_ptr(), d_rgba.data_ptr())
    gpu_ctx.synchronize()
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert free_after >= free_before, (free_before, free_after)
    assert np.allclose(d_rad.cpu().numpy(), first, rtol=RTOL, atol=ATOL)
    sc.close(); sc2.close()


def test_noalloc_rejects_what_the_static_scratch_cannot_hold(gpu_ctx):
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(512, 256)                # 131072 pixels > PT_NOALLOC_MAX_PIXELS
    sc = api.Scene(gpu_ctx, sh, mt)
    d_rad, d_rgba = _device_outputs(gpu_ctx, 256, 512)
    with pytest.raises(api.PathtraceError, match="PT_NOALLOC_MAX_PIXELS"):
        api.render_device_noalloc(gpu_ctx, sc, cam, api.render_params(spp=1), d_rad.data_ptr(), d_rgba.data_ptr())
    _, _, small = scenes.default_scene(64, 64)
    with pytest.raises(api.PathtraceError, match="flags must be 0"):
        api.render_device_noalloc(gpu_ctx, sc, small, api.render_params(spp=1, profile=True), d_rad.data_ptr(), d_rgba.data_ptr())
    with pytest.raises(api.PathtraceError, match="output pointers"):
        api.render_device_noalloc(gpu_ctx, sc, small, api.render_params(spp=1), 0, 0)
    with pytest.raises(api.PathtraceError, match="exceed the static scratch"):
        api.render_device_noalloc(gpu_ctx, sc, small, api.render_params(spp=1024, path_pool=3072 * 2000), d_rad.data_ptr(), d_rgba.data_ptr())
    sc.close()
    # a scene too large for shared memory needs the HBM path queues, which are cudaMalloc'd
    from conftest import random_scene
    shapes, mats = random_scene(3, 4000)
    big = api.Scene(gpu_ctx, shapes, mats)
    with pytest.raises(api.PathtraceError, match="does not fit the shared-memory wavefront"):
        api.render_device_noalloc(gpu_ctx, big, small, api.render_params(spp=1), d_rad.data_ptr(), d_rgba.data_ptr())
    big.close()
