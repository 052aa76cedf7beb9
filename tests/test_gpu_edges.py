"""Edge cases of the render path on the GPU (run with -m gpu), through the C ABI: progressive rendering in sample chunks
(`sample_offset` replaces the window's progressive display), ragged and tiny frames, empty and lightless
scenes, the tie rule of world.rs:281-287 on coincident triangles, and error codes instead of panics."""
import ctypes as C

import numpy as np
import pytest

from conftest import oracle_camera, random_scene

pytestmark = pytest.mark.gpu


def _default(w, h):
    from pathtrace_b200 import scenes
    return scenes.default_scene(w, h)


@pytest.mark.parametrize("hbm_queues", [False, True])
def test_progressive_chunks_equal_one_shot(hbm_queues, gpu_ctx):
    """Samples [0, 12) rendered as 5 + 4 + 3 with increasing sample_offset are the one-shot sample set: the counter-based RNG is
    keyed by the global sample index, so the weighted mean of the chunk means equals the one-shot mean up to the
    order of the float additions."""
    from pathtrace_b200 import api
    sh, mt, cam = _default(96, 64)
    sc = api.Scene(gpu_ctx, sh, mt)
    one, _, st1 = api.render(gpu_ctx, sc, cam, api.render_params(spp=12, hbm_queues=hbm_queues))
    acc = np.zeros_like(one, dtype=np.float64)
    rays = 0
    off = 0
    for n in (5, 4, 3):
        rad, _, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=n, sample_offset=off, hbm_queues=hbm_queues))
        acc += rad.astype(np.float64) * n
        rays += st["rays_primary"] + st["rays_shadow"] + st["rays_scattered"]
        off += n
    sc.close()
    assert rays == st1["rays_primary"] + st1["rays_shadow"] + st1["rays_scattered"]     # the same paths, exactly
    assert np.allclose(acc / 12, one, rtol=2e-5, atol=1e-6)


@pytest.mark.parametrize("size", [(2, 2), (33, 17), (5, 129)])
@pytest.mark.parametrize("hbm_queues", [False, True])
def test_ragged_frames(size, hbm_queues, oracle, gpu_ctx):
    """Frames that are not a multiple of a warp, a tile or a CTA: every pixel is written once and equals the oracle's."""
    from pathtrace_b200 import api
    w, h = size
    sh, mt, cam = _default(w, h)
    sc = api.Scene(gpu_ctx, sh, mt)
    rad, rgba, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=4, hbm_queues=hbm_queues))
    sc.close()
    assert rad.shape == (h, w, 3) and rgba.shape == (h, w, 4)
    assert st["samples"] == 4 * w * h and st["rays_primary"] == 4 * w * h
    assert np.all(rgba[..., 3] == 255)                                                    # world.rs:330
    assert np.isfinite(rad).all() and rad.min() >= 0.0
    # same sample set, same FP32 paths: the frame matches the oracle path by path up to rare flipped decisions
    ref = oracle.render(oracle_camera(oracle, cam), sh, mt, oracle.RenderParams(spp=4), "f32", recursive=False)["sum"] / 4
    close = np.all(np.abs(rad - ref) <= 2e-3 * np.maximum(np.abs(ref), 1e-3), axis=2)
    assert close.mean() > 0.97


@pytest.mark.parametrize("size", [(1, 7), (31, 1), (0, 4)])
def test_one_pixel_wide_frames_are_refused(size, gpu_ctx):
    """camera.rs:143-144 divides by (W-1) and (H-1): a frame one pixel wide has no defined rays in the reference (NaN
    directions); the library refuses it with a status instead of rendering NaNs."""
    from pathtrace_b200 import api
    import dataclasses
    sh, mt, cam = _default(8, 8)
    cam = dataclasses.replace(cam, width=size[0], height=size[1])
    sc = api.Scene(gpu_ctx, sh, mt)
    with pytest.raises(api.PathtraceError, match="camera"):
        api.render(gpu_ctx, sc, cam, api.render_params(spp=1))
    sc.close()


@pytest.mark.parametrize("hbm_queues", [False, True])
def test_empty_scene_is_black(hbm_queues, gpu_ctx):
    from pathtrace_b200 import api
    sh, mt = random_scene(1, 0)
    _, _, cam = _default(40, 24)
    sc = api.Scene(gpu_ctx, sh, mt)
    assert sc.n_lights == 0
    rad, rgba, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=3, hbm_queues=hbm_queues))
    sc.close()
    assert not rad.any() and not rgba[..., :3].any() and np.all(rgba[..., 3] == 255)     # rendering.rs:140-141
    assert st["rays_primary"] == 3 * 40 * 24 and st["rays_shadow"] == 0 and st["rays_scattered"] == 0


@pytest.mark.parametrize("strategy", ["mis", "brdf_only"])
def test_scene_without_lights_is_black_and_terminates(strategy, gpu_ctx):
    """No emitter: world.rs:251-253 returns no light sample, every path ends by roulette or a miss with zero radiance."""
    from pathtrace_b200 import api
    sh, mt = random_scene(11, 40, with_light=False)
    was_light = mt["kind"] == 1
    mt["kind"][was_light] = 0                                                             # emissive -> Lambertian
    mt["p"][was_light, :3] = 0.5
    _, _, cam = _default(64, 48)
    sc = api.Scene(gpu_ctx, sh, mt)
    assert sc.n_lights == 0
    rad, _, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=8, strategy=strategy))
    sc.close()
    assert not rad.any() and st["rays_shadow"] == 0 and st["rays_scattered"] > 0


@pytest.mark.parametrize("traversal", ["brute", "bvh"])
def test_coincident_triangles_higher_index_wins(traversal, oracle, gpu_ctx):
    """world.rs:281-287 scans in index order with an inclusive t_max, so of two identical triangles the LATER one is the hit
     — whichever order the scan table or the LBVH visits them in."""
    from pathtrace_b200 import api
    from pathtrace_b200.objects import MATERIAL_DTYPE, SHAPE_DTYPE
    tri = np.array([-1, -1, -3, 1, -1, -3, 0, 1, -3], np.float32)
    far = np.array([-2, -2, -4, 2, -2, -4, 0, 2, -4], np.float32)
    order = [tri, far, tri, far, tri]                                                     # ids 0, 2, 4 coincide; 1, 3 coincide behind them
    sh = np.zeros(len(order), SHAPE_DTYPE); mt = np.zeros(len(order), MATERIAL_DTYPE)
    for i, v in enumerate(order):
        sh[i]["kind"] = 1; sh[i]["v"] = v
        mt[i]["kind"] = 0; mt[i]["p"][:3] = 0.5
    _, _, cam = _default(64, 64)
    sc = api.Scene(gpu_ctx, sh, mt)
    ids, t, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal=traversal))
    sc.close()
    rid, rt, _ = oracle.trace_primary(oracle_camera(oracle, cam), sh, mt, precision="f32")
    assert np.array_equal(ids, rid)
    assert set(np.unique(ids)) <= {-1, 3, 4} and (ids == 4).any() and (ids == 3).any()


def test_invalid_arguments_return_error_codes(gpu_ctx):
    """no exception or abort crosses the ABI; a bad call returns a status and leaves a message."""
    from pathtrace_b200 import api
    lib = gpu_ctx.lib
    sh, mt, cam = _default(16, 16)
    sc = api.Scene(gpu_ctx, sh, mt)
    cs = api.camera_struct(cam)
    prm = api.render_params(spp=1)
    rad = np.zeros((16, 16, 3), np.float32)
    # null scene / null camera
    assert lib.pt_render(gpu_ctx.handle, None, C.byref(cs), C.byref(prm), rad.ctypes.data_as(C.c_void_p), None, None) != 0
    assert lib.pt_last_error(gpu_ctx.handle)
    assert lib.pt_render(gpu_ctx.handle, sc.handle, None, C.byref(prm), rad.ctypes.data_as(C.c_void_p), None, None) != 0
    # zero-sized frame
    bad = api.camera_struct(cam); bad.width = 0
    assert lib.pt_render(gpu_ctx.handle, sc.handle, C.byref(bad), C.byref(prm), rad.ctypes.data_as(C.c_void_p), None, None) != 0
    # the context is still usable afterwards
    out, _, st = api.render(gpu_ctx, sc, cam, prm)
    sc.close()
    assert st["samples"] == 256 and np.isfinite(out).all()
