"""Multi-GPU through the C ABI (pt_group_*), run on a box with >= 2 B200 (-m gpu; skipped below two GPUs).

The path shards by samples per pixel: every device renders the whole frame for its share of the global sample indices, ONE
ncclReduce(sum) inside the library combines the float4 sum buffers, the root resolves.  The Philox counters are keyed by the
global sample index, so the N-GPU sample set IS the 1-GPU sample set: the frames agree up to the order of the float adds.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("scene_name,spp", [("default", 64), ("default", 3), ("spheres", 16)])
def test_group_frame_equals_single_gpu_frame(scene_name, spp, gpu_ctx):
    from pathtrace_b200 import api, scenes
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    sh, mt, cam = scenes.default_scene(160, 120) if scene_name == "default" else scenes.random_spheres_scene(160, 90)
    sc = api.Scene(gpu_ctx, sh, mt)
    one, rgba1, st1 = api.render(gpu_ctx, sc, cam, api.render_params(spp=spp, sample_offset=2))
    sc.close()
    for devs in ([0, 1], list(range(min(n, 8)))):
        g = api.Group(devs)
        assert g.size == len(devs)
        gs = g.scene_create(sh, mt)
        rad, rgba, st = g.render(gs, cam, api.render_params(spp=spp, sample_offset=2))
        g.scene_destroy(gs)
        g.close()
        for k in ("samples", "rays_primary", "rays_shadow", "rays_scattered", "path_vertices"):
            assert st[k] == st1[k], (devs, k, st[k], st1[k])          # the same paths, exactly, whatever the split
        assert np.allclose(rad, one, rtol=5e-5, atol=1e-6), np.abs(rad - one).max()
        assert (np.abs(rgba.astype(int) - rgba1.astype(int)) <= 1).all() and (rgba[..., 3] == 255).all()


def test_group_of_one_is_pt_render(gpu_ctx):
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(96, 64)
    sc = api.Scene(gpu_ctx, sh, mt)
    one, rgba1, st1 = api.render(gpu_ctx, sc, cam, api.render_params(spp=8))
    sc.close()
    g = api.Group([0])
    gs = g.scene_create(sh, mt)
    rad, rgba, st = g.render(gs, cam, api.render_params(spp=8))
    g.scene_destroy(gs)
    g.close()
    assert np.array_equal(rgba, rgba1) and np.allclose(rad, one, rtol=1e-6, atol=1e-7)
    assert st["samples"] == st1["samples"]


def test_group_rejects_bad_arguments():
    from pathtrace_b200 import api
    with pytest.raises(api.PathtraceError, match="duplicate"):
        api.Group([0, 0])
    with pytest.raises(api.PathtraceError):
        api.Group([])
    with pytest.raises(api.PathtraceError, match="out of range|device"):
        api.Group([0, 99])
