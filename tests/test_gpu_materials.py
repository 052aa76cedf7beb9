"""GPU-vs-oracle parity of every material model, both light shapes and both integrators (run on a B200 with -m gpu).

The default scene only exercises Lambertian walls, one rough-glass sphere and triangle lights under MIS.  Here the CUDA path
meets the oracle on: Oren-Nayar (material.rs:166-296, the device evaluates the azimuth term algebraically), metallic GGX with
coloured F0 (mirror.rs:126-132, :187, :227), dielectric GGX, sphere lights (shape.rs:91-145: cone sampling, re-intersection
with the un-normalised direction at :132), triangle lights, and BrdfOnlyStrategy (rendering.rs:213-265) next to MisStrategy —
two ways each: path by path against the FP32 oracle (same counter-based RNG => the same path), and as a converged image
against committed f64 oracle renders (tests/golden/make_golden_images.py) with the Monte Carlo noise of both sides removed.
"""
import os

import numpy as np
import pytest

from conftest import oracle_camera

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _scene(name, w=None, h=None):
    from pathtrace_b200 import scenes
    if name == "zoo":
        return scenes.material_zoo_scene(w or 128, h or 128)
    return scenes.random_spheres_scene(w or 200, h or 112)


def _object_classes(name, mt):
    """object id -> label of what the primary ray sees there (so that agreement is asserted per material, not on average)."""
    lab = {}
    for i, m in enumerate(mt):
        k = int(m["kind"])
        if k == 0: lab[i] = "lambert"
        elif k == 1: lab[i] = "light"
        elif k == 2: lab[i] = "ggx_metal" if m["p"][4] > 0.99 else "ggx_glass"
        else: lab[i] = "oren_nayar"
    return lab


@pytest.mark.parametrize("strategy", ["mis", "brdf_only"])
@pytest.mark.parametrize("scene_name,mode", [("zoo", "smem_scan"), ("zoo", "smem_bvh"), ("zoo", "hbm"), ("spheres", "hbm")])
def test_single_samples_match_fp32_oracle_all_materials(scene_name, mode, strategy, oracle, gpu_ctx):
    """Sample s of pixel p is the same path on the GPU and in the FP32 oracle.  One-sample frames agree pixel by pixel except
    where a rounding difference flips a discrete decision; asserted per class of the primary hit, so a wrong Oren-Nayar,
    metal, glass or sphere-light formula cannot hide behind the Lambertian majority."""
    from pathtrace_b200 import api
    sh, mt, cam = _scene(scene_name, 96, 96) if scene_name == "zoo" else _scene(scene_name, 200, 112)
    ocam = oracle_camera(oracle, cam)
    sc = api.Scene(gpu_ctx, sh, mt)
    kw = dict(hbm_queues=(mode == "hbm"), traversal="bvh" if mode == "smem_bvh" else "auto")
    ids, _, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(), jitter=False)
    labels = _object_classes(scene_name, mt)
    lab_img = np.array([labels.get(int(i), "miss") for i in ids.ravel()], dtype=object).reshape(ids.shape)
    agree = {}
    sum_g = sum_c = 0.0
    s_strategy = {"mis": 0, "brdf_only": 1}[strategy]
    samples = (0, 3, 200, 4095) if scene_name == "zoo" else (1, 77)
    for s in samples:
        rad, _, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=1, sample_offset=s, strategy=strategy, **kw))
        if mode == "smem_scan":
            assert st["n_kernel"]["smem_wavefront"] == 1 and st["scan_mode"] == 4
        elif mode == "smem_bvh":
            assert st["n_kernel"]["smem_wavefront"] == 1 and st["scan_mode"] == 0
        else:
            assert st["n_kernel"]["smem_wavefront"] == 0
        r = oracle.render(ocam, sh, mt, oracle.RenderParams(spp=1, sample_offset=s, strategy=s_strategy), "f32", recursive=False)
        c = r["sum"]
        close = np.all(np.abs(rad - c) <= 3e-3 * np.maximum(np.abs(c), 1e-3), axis=2)
        for name in set(labels.values()):
            m = lab_img == name
            if m.any():
                a, n = agree.get(name, (0, 0))
                agree[name] = (a + int(close[m].sum()), n + int(m.sum()))
        sum_g += float(rad.sum()); sum_c += float(c.sum())
        npx = cam.width * cam.height
        assert st["rays_primary"] == npx
        # The oracle walks a path whose GGX sample failed (the reference's (normal, 0, 1, 0) tuple, mirror.rs:258-265) on with zero
        # throughput until roulette kills it; the device ends it at once (same radiance): it shades a few vertices less.
        dv = r["counters"]["vertices"] - st["path_vertices"]
        assert -0.01 * npx <= dv <= 0.04 * npx, dv
        if strategy == "brdf_only":
            assert st["rays_shadow"] == 0                                  # no NEE in BrdfOnlyStrategy
    sc.close()
    frac = {k: a / n for k, (a, n) in agree.items()}
    print(scene_name, mode, strategy, "agreement by primary-hit class:", {k: round(v, 4) for k, v in frac.items()}, "energy", sum_g / sum_c)
    need = {"lambert", "oren_nayar", "ggx_metal", "ggx_glass", "light"} if scene_name == "zoo" else {"lambert", "ggx_metal", "ggx_glass"}
    assert need <= set(frac), set(frac)
    for k, v in frac.items():
        assert v > (0.97 if k.startswith("ggx") else 0.985), (k, v)      # GGX paths are the longest: most chances for a flipped decision
    # a flipped decision moves one path's energy, so the frame energy agrees to a few percent (brdf_only: few bright paths)
    assert abs(sum_g / sum_c - 1.0) < (0.02 if strategy == "mis" else 0.08)


def _converged(gpu_ctx, name, fixture, strategy, spp, batches, tile):
    from pathtrace_b200 import api
    g = np.load(os.path.join(G, fixture))
    ref, ref_var = g["mean"].astype(np.float64), g["var_of_mean"].astype(np.float64)
    H, W = ref.shape[:2]
    sh, mt, cam = _scene(name, W, H)
    sc = api.Scene(gpu_ctx, sh, mt)
    rad, m2, st = api.render_moments(gpu_ctx, sc, cam, api.render_params(spp=spp, strategy=strategy, seed=0xB200), batches)
    sc.close()
    assert st["samples"] == W * H * spp
    gpu = rad.astype(np.float64)
    gpu_var = api.variance_of_mean(rad, m2, spp, batches)
    diff, tot = gpu - ref, gpu_var + ref_var
    # (1) RMSE with the noise of both sides removed; pixels whose variance is far above the rest (light edges, caustic
    # fireflies) would dominate the estimator's own error and are judged by their z-scores instead
    lum_var = tot.sum(axis=2)
    body = lum_var < 30 * np.median(lum_var)
    debiased = (diff ** 2 - tot)[body].mean(axis=0)
    se = np.sqrt(2.0 * (tot[body] ** 2).sum(axis=0)) / body.sum()
    print(name, strategy, "spp", spp, "body", body.mean(), "debiased MSE", debiased, "SE", se, "RMSE", np.sqrt(np.maximum(debiased, 0)))
    assert body.mean() > 0.8
    assert np.all(debiased <= (1e-3) ** 2 + 3.0 * se), (debiased, se)
    # (2) tile-aggregated z-scores ~ N(0,1): no excess error, no spatial structure
    th, tw = H // tile, W // tile
    num = diff[:th * tile, :tw * tile].reshape(th, tile, tw, tile, 3).sum(axis=(1, 3))
    den = np.sqrt(tot[:th * tile, :tw * tile].reshape(th, tile, tw, tile, 3).sum(axis=(1, 3)))
    Z = num / np.maximum(den, 1e-30)
    print("tile Z rms", np.sqrt((Z ** 2).mean()), "mean", Z.mean(), "max", np.abs(Z).max())
    assert 0.6 < np.sqrt((Z ** 2).mean()) < 1.4
    assert np.abs(Z).max() < 5.0 and abs(Z.mean()) < 0.6
    # (3) per-pixel z-scores with the batch-moment variance (batches - 1 degrees of freedom): the bulk is
    # standard normal; the tail is heavier than a normal's by the t-distribution and by the skew of the estimator
    # (pixels that look straight into a light have zero variance on both sides and differ by FP32 rounding only: the floor
    # keeps them from dividing by zero)
    z = diff / np.sqrt(tot + (1e-5 * np.maximum(np.abs(ref), 1.0)) ** 2)
    zb = z[body]
    print("pixel z: rms", np.sqrt((zb ** 2).mean()), "|z|>3", (np.abs(zb) > 3).mean(), "|z|>5", (np.abs(zb) > 5).mean())
    assert 0.8 < np.sqrt((zb ** 2).mean()) < 1.25
    assert (np.abs(zb) > 3).mean() < 0.02 and (np.abs(zb) > 6).mean() < 1e-3
    # (4) energy
    assert abs(gpu.mean() - ref.mean()) / ref.mean() < 5e-3
    return st


@pytest.mark.parametrize("strategy,spp", [("mis", 16384), ("brdf_only", 32768)])
def test_converged_material_zoo_matches_f64_oracle(strategy, spp, gpu_ctx):
    fixture = f"zoo_128x128_{'mis' if strategy == 'mis' else 'brdf'}_f64.npz"
    st = _converged(gpu_ctx, "zoo", fixture, strategy, spp, 32, 16)
    assert st["scan_mode"] == 4                    # the scan the benchmark runs: exact pair test for primary rays, planar entries for the rest


@pytest.mark.parametrize("strategy,spp", [("mis", 4096), ("brdf_only", 8192)])
def test_converged_random_spheres_matches_f64_oracle(strategy, spp, gpu_ctx):
    """Config C3's scene (4 emissive spheres, 15 % metallic spheres, glass): formulation B, LBVH in shared memory."""
    fixture = f"spheres_200x112_{'mis' if strategy == 'mis' else 'brdf'}_f64.npz"
    st = _converged(gpu_ctx, "spheres", fixture, strategy, spp, 32, 8)
    assert st["n_kernel"]["smem_wavefront"] == 0


def test_hit_distance_bar_on_sphere_scenes(oracle, gpu_ctx):
    """North-star bar |dt|/t <= 1e-5 vs the reference's f64 numerics on a sphere-heavy scene (C3 at 400x224).  The bar holds for
    every ray that meets its surface at |n.d| >= 0.1; a ray grazing a silhouette is ill-conditioned in any FP32 scheme (the
    rounding of the ray direction itself moves the root by eps / |n.d|), so there the bound scales with 1 / |n.d|."""
    from pathtrace_b200 import api
    sh, mt, cam = _scene("spheres", 400, 224)
    ocam = oracle_camera(oracle, cam)
    sc = api.Scene(gpu_ctx, sh, mt)
    ids, t, dirs = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="bvh"), want_dirs=True)
    sc.close()
    did, dt, _ = oracle.trace_primary(ocam, sh, mt, precision="f64")
    same = (did == ids) & (ids >= 0)
    assert same.sum() > 0.99 * (ids >= 0).sum() and (ids >= 0).mean() > 0.5 and (did != ids).mean() < 5e-3
    rel = np.abs(t.astype(np.float64) - dt) / np.where(same, dt, 1.0)
    o = np.array([cam.origin.x, cam.origin.y, cam.origin.z])
    p = o + dirs.astype(np.float64) * dt[..., None]
    kind = sh["kind"][np.maximum(ids, 0)]
    c = sh["v"][np.maximum(ids, 0)][..., :3].astype(np.float64); r = sh["v"][np.maximum(ids, 0)][..., 3].astype(np.float64)
    v0 = sh["v"][np.maximum(ids, 0)].astype(np.float64)
    n_tri = np.cross(v0[..., 3:6] - v0[..., 0:3], v0[..., 6:9] - v0[..., 0:3])
    n_tri /= np.maximum(np.linalg.norm(n_tri, axis=-1, keepdims=True), 1e-300)
    n = np.where((kind == 0)[..., None], (p - c) / np.where(r == 0, 1, r)[..., None], n_tri)
    cosi = np.abs((n * dirs.astype(np.float64)).sum(-1))
    bound = 1e-5 * np.maximum(1.0, 0.1 / np.maximum(cosi, 1e-12))
    bad = same & (rel > bound)
    print("pixels", same.sum(), "beyond 1e-5:", (same & (rel > 1e-5)).sum(), "of them grazing (|n.d| < 0.1):", (same & (rel > 1e-5) & (cosi < 0.1)).sum(),
          "max rel", rel[same].max())
    assert not bad.any(), (rel[bad], cosi[bad])
    assert (same & (rel > 1e-5)).sum() <= 1e-4 * same.sum()


@pytest.mark.parametrize("scene_name", ["default", "zoo"])
@pytest.mark.parametrize("jitter", [False, True])
def test_secondary_ray_scan_against_f64_oracle(scene_name, jitter, oracle, gpu_ctx):
    """The planar-entry test that shadow and scattered rays use in scan mode 4 (pt_wave.cuh: fast_pair; one entry per triangle or
    per parallelogram of two triangles), put in front of the reference's f64 numerics through PT_TRAVERSAL_BRUTE_FAST: the same
    object everywhere except on shared edges / silhouettes, hit distance within 1e-5 relative, no ray lost on a diagonal."""
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(320, 240) if scene_name == "default" else scenes.material_zoo_scene(320, 240)
    ocam = oracle_camera(oracle, cam)
    sc = api.Scene(gpu_ctx, sh, mt)
    ids, t, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="brute_fast"), jitter=jitter, sample=5)
    eid, et, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="brute"), jitter=jitter, sample=5)
    sc.close()
    did, dt, _ = oracle.trace_primary(ocam, sh, mt, precision="f64", jitter=jitter, sample=5)
    same = (did == ids) & (ids >= 0)
    print("id mismatches vs f64:", (did != ids).sum(), "exact scan vs f64:", (did != eid).sum(), "fast vs exact:", (ids != eid).sum())
    assert (did != ids).mean() < 2e-3                       # as few as the exact FP32 scan has against f64 (shared edges only)
    assert (did != ids).sum() <= (did != eid).sum() + 0.0005 * ids.size
    assert ((ids < 0) & (did >= 0)).sum() <= ((eid < 0) & (did >= 0)).sum() + 2      # no cracks along the diagonals of the quads
    rel = np.abs(t[same].astype(np.float64) - dt[same]) / dt[same]
    assert rel.max() <= 1e-5, rel.max()


def test_fast_and_exact_secondary_scans_render_the_same_image(gpu_ctx):
    """scan mode 4 (default) against scan mode 3 (bit-exact Moller-Trumbore for every ray): the same estimator, so the same
    image up to the rounding of secondary hits — per-pixel agreement on almost every single-sample path, equal ray statistics
    to 1e-3, equal converged means."""
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(160, 120)
    sc = api.Scene(gpu_ctx, sh, mt)
    a, _, sa = api.render(gpu_ctx, sc, cam, api.render_params(spp=1, sample_offset=11, traversal="brute"))
    b, _, sb = api.render(gpu_ctx, sc, cam, api.render_params(spp=1, sample_offset=11, traversal="brute_exact"))
    assert sa["scan_mode"] == 4 and sb["scan_mode"] == 3
    close = np.all(np.abs(a - b) <= 2e-3 * np.maximum(np.abs(b), 1e-3), axis=2)
    assert close.mean() > 0.997, close.mean()
    for k in ("rays_shadow", "rays_scattered", "path_vertices"):
        assert abs(sa[k] - sb[k]) <= 1e-3 * sb[k] + 4, (k, sa[k], sb[k])
    a, _, _ = api.render(gpu_ctx, sc, cam, api.render_params(spp=4096, traversal="brute"))
    b, _, _ = api.render(gpu_ctx, sc, cam, api.render_params(spp=4096, traversal="brute_exact", seed=77))
    sc.close()
    assert abs(a.mean() - b.mean()) / b.mean() < 2e-3
    num = (a - b).reshape(15, 8, 20, 8, 3).mean(axis=(1, 3))
    assert np.abs(num).max() < 0.02 * max(b.mean(), 1e-6) + 0.01
