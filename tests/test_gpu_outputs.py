"""The two outputs of the render path, byte for byte (run on a B200 with -m gpu):

  World.data           RGBA8 = mean -> sqrt gamma -> clamp -> *255 truncated, alpha 255     world.rs:315-332
  luminance.csv        "x,y,r,g,b,luminance" with {:.6}                                      world.rs:344-369

pt_resolve_device / pt_render's RGBA8 must equal the oracle's tone map of the SAME linear radiance for every byte, including
values that sit exactly on a truncation boundary (sqrt(m) * 255 an integer), and the CSV the library writes must equal the
oracle's file for the same buffer.
"""
import os

import numpy as np
import pytest

from conftest import oracle_camera

pytestmark = pytest.mark.gpu


def _boundary_means():
    """Radiance values straddling every truncation boundary of (u8)(sqrt(m) * 255): m = (k / 255)^2 and its FP32 neighbours."""
    k = np.arange(0, 256, dtype=np.float64)
    exact = ((k / 255.0) ** 2).astype(np.float32)
    vals = [exact, np.nextafter(exact, np.float32(0)), np.nextafter(exact, np.float32(2)), np.nextafter(np.nextafter(exact, np.float32(2)), np.float32(2))]
    extra = np.array([0.0, 1.0, 1.0000001, 4.0, 1e-9, 1e-30, 0.25, 0.5, -0.0, -1.0, np.nan, np.inf, 3.0e38], np.float32)
    return np.concatenate(vals + [extra]).astype(np.float32)


@pytest.mark.parametrize("spp", [1, 3, 1024])
def test_resolve_equals_oracle_tonemap_on_identical_radiance(spp, oracle, gpu_ctx):
    import torch
    from pathtrace_b200 import api
    m = _boundary_means()
    n = (len(m) + 2) // 3 * 3
    m = np.resize(m, n)
    W, H = n // 3 // 2 * 2, 2                           # any frame shape: resolve is per pixel
    npix = W * H
    vals = np.resize(m, npix * 3).reshape(npix, 3)
    accum = np.zeros((npix, 4), np.float32)
    accum[:, :3] = vals * np.float32(spp)               # the sum buffer the kernels would have produced
    dev = torch.device("cuda", gpu_ctx.device)
    stream = torch.cuda.ExternalStream(gpu_ctx.stream, device=dev)
    with torch.cuda.stream(stream):
        d_acc = torch.from_numpy(accum).to(dev)
        d_rad = torch.empty(npix, 3, device=dev, dtype=torch.float32)
        d_rgba = torch.empty(npix, 4, device=dev, dtype=torch.uint8)
        api.resolve_device(gpu_ctx, d_acc.data_ptr(), W, H, spp, d_rad.data_ptr(), d_rgba.data_ptr())
    stream.synchronize()
    rad = d_rad.cpu().numpy(); rgba = d_rgba.cpu().numpy()
    # the mean is the correctly rounded FP32 quotient (world.rs:315 divides in f64; FP32 is this path's arithmetic)
    with np.errstate(all="ignore"):
        want_mean = accum[:, :3] / np.float32(spp)
    fin = ~np.isnan(want_mean)                         # (CUDA returns the canonical NaN, x86 keeps the payload)
    assert np.array_equal(np.isnan(rad), ~fin)
    assert np.array_equal(rad[fin].view(np.uint32), want_mean[fin].view(np.uint32))
    # gamma / clamp / truncation: byte-equal to world.rs:321-332 applied to that radiance
    want = oracle.tonemap(rad.astype(np.float64))
    assert np.array_equal(rgba, want), np.argwhere(rgba != want)[:8]
    assert (rgba[:, 3] == 255).all()


@pytest.mark.parametrize("hbm_queues", [False, True])
def test_render_rgba_is_the_tonemap_of_its_radiance(hbm_queues, oracle, gpu_ctx):
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(160, 120)
    sc = api.Scene(gpu_ctx, sh, mt)
    rad, rgba, _ = api.render(gpu_ctx, sc, cam, api.render_params(spp=16, hbm_queues=hbm_queues))
    sc.close()
    want = oracle.tonemap(rad.astype(np.float64))
    assert np.array_equal(rgba, want)
    assert len(np.unique(rgba[..., :3])) > 100          # a real image, not a constant


def test_luminance_csv_bytes_equal_oracle(tmp_path, oracle, gpu_ctx):
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(48, 32)
    sc = api.Scene(gpu_ctx, sh, mt)
    rad, _, _ = api.render(gpu_ctx, sc, cam, api.render_params(spp=8))
    sc.close()
    rad = rad.copy()
    rad[0, 0] = [np.nan, np.inf, -np.inf]               # Rust prints NaN / inf / -inf (world.rs:361-365 uses {:.6})
    rad[0, 1] = [-0.0, 1e-7, 123456.789]
    a, b = os.path.join(tmp_path, "gpu.csv"), os.path.join(tmp_path, "oracle.csv")
    api.export_luminance(a, rad)
    oracle.export_luminance(b, rad.astype(np.float64))
    ba, bb = open(a, "rb").read(), open(b, "rb").read()
    assert ba == bb
    lines = ba.decode().splitlines()
    assert lines[0] == "x,y,r,g,b,luminance" and len(lines) == 1 + 48 * 32
    assert lines[1] == "0,0,NaN,inf,-inf,NaN"
    assert lines[2].startswith("1,0,-0.000000,0.000000,123456.789062,")


def test_moments_equal_plain_render_and_give_a_variance(gpu_ctx):
    """pt_render_moments renders the same samples as pt_render (the RNG is keyed by the global sample index) and its second
    moment with batches == spp is the literal per-sample sum of squares."""
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(64, 48)
    sc = api.Scene(gpu_ctx, sh, mt)
    spp = 64
    one, _, st1 = api.render(gpu_ctx, sc, cam, api.render_params(spp=spp, sample_offset=9))
    rad, m2, st = api.render_moments(gpu_ctx, sc, cam, api.render_params(spp=spp, sample_offset=9), 8)
    assert st["samples"] == st1["samples"] and st["rays_scattered"] == st1["rays_scattered"] and st["rays_shadow"] == st1["rays_shadow"]
    assert np.allclose(rad, one, rtol=2e-5, atol=1e-6)
    # batches == spp: moment2 is sum of squared per-sample radiance; check against single-sample renders of a few samples
    rad1, m21, _ = api.render_moments(gpu_ctx, sc, cam, api.render_params(spp=4, sample_offset=0), 4)
    acc = np.zeros_like(rad1, dtype=np.float64); sq = np.zeros_like(acc)
    for s in range(4):
        r, _, _ = api.render(gpu_ctx, sc, cam, api.render_params(spp=1, sample_offset=s))
        acc += r; sq += r.astype(np.float64) ** 2
    sc.close()
    assert np.allclose(rad1, acc / 4, rtol=1e-5, atol=1e-6)
    assert np.allclose(m21, sq, rtol=1e-4, atol=1e-6)
    v = api.variance_of_mean(rad, m2, spp, 8)
    assert v.shape == rad.shape and (v >= 0).all() and v.mean() > 0
