"""Committed fixtures (tests/golden/, produced by tests/golden/make_golden.py from the CPU oracle).

CPU part: the oracle still reproduces them (guards the checker against drift).
GPU part (-m gpu): the CUDA path against the converged f64 image — the north-star image bar: per-channel RMSE <= 1e-3
once the Monte Carlo noise of both sides is subtracted, plus the per-pixel z-score test.
"""
import os

import numpy as np
import pytest

from conftest import oracle_camera

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(G, name))


def test_oracle_reproduces_primary_hits(oracle):
    from pathtrace_b200 import scenes
    g = _load("primary_default_96x64.npz")
    sh, mt, cam = scenes.default_scene(96, 64)
    oc = oracle_camera(oracle, cam)
    ids, t, d = oracle.trace_primary(oc, sh, mt, precision="f32", want_dirs=True)
    assert np.array_equal(ids, g["id_f32"])
    assert np.array_equal(t.astype(np.float32).view(np.uint32), g["t_f32"].view(np.uint32))
    assert np.array_equal(d.view(np.uint32), g["dirs_f32"].view(np.uint32))
    ids, t, _ = oracle.trace_primary(oc, sh, mt, precision="f64")
    assert np.array_equal(ids, g["id_f64"]) and np.array_equal(t, g["t_f64"])
    # the frame shows every object class: walls, light, glass sphere, background
    assert {-1, 12}.issubset(set(np.unique(g["id_f32"]).tolist()))


def test_oracle_reproduces_lbvh(oracle):
    from pathtrace_b200 import scenes
    g = _load("lbvh_default.npz")
    b = oracle.lbvh_build(scenes.default_scene(96, 64)[0])
    for k in ("codes", "prims", "nodes"):
        assert np.array_equal(b[k], g[k]), k
    assert np.array_equal(b["bounds"].view(np.uint32), g["bounds"].view(np.uint32))


def test_oracle_reproduces_individual_paths(oracle):
    from pathtrace_b200 import scenes
    g = _load("samples_default_f64.npz")
    sh, mt, cam = scenes.default_scene(64, 64)
    oc = oracle_camera(oracle, cam)
    prm = oracle.RenderParams(spp=1, seed=0x5EED)
    for form, key in ((True, "recursive"), (False, "iterative")):
        got = np.array([oracle.sample(oc, sh, mt, prm, int(x), int(y), int(s), "f64", form) for x, y, s in zip(g["x"], g["y"], g["s"])])
        assert np.allclose(got, g[key], rtol=1e-12, atol=1e-14), key
    # the literal recursion (rendering.rs:34-142) and the iterative form the GPU executes agree path by path
    assert np.allclose(g["recursive"], g["iterative"], rtol=1e-10, atol=1e-12)
    assert (g["recursive"].sum(axis=1) > 0).sum() >= 10


def test_reference_image_fixture_is_sane():
    g = _load("cornell_160x160_f64.npz")
    m, v = g["mean"], g["var_of_mean"]
    assert m.shape == (160, 160, 3) and v.shape == m.shape and int(g["spp"]) == 8192
    assert np.isfinite(m).all() and (m >= 0).all() and (v >= 0).all()
    assert 0.1 < m.mean() < 1.0
    assert np.sqrt(v).mean() < 1e-2          # mean standard error of the reference itself


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("hbm_queues", [False, True])
def test_converged_image_matches_reference(hbm_queues, gpu_ctx):
    """Image bar of the north star at 160x160: RMSE (noise removed) <= 1e-3 per channel; z-scores look normal."""
    from pathtrace_b200 import api, scenes
    g = _load("cornell_160x160_f64.npz")
    ref, ref_var = g["mean"].astype(np.float64), g["var_of_mean"].astype(np.float64)
    sh, mt, cam = scenes.default_scene(160, 160)
    sc = api.Scene(gpu_ctx, sh, mt)
    spp_g = 32768 if not hbm_queues else 8192
    # two independent GPU renders (disjoint sample ranges) give the GPU-side variance of the mean without a sumsq buffer
    a, _, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=spp_g // 2, sample_offset=0, hbm_queues=hbm_queues))
    b, _, _ = api.render(gpu_ctx, sc, cam, api.render_params(spp=spp_g // 2, sample_offset=spp_g // 2, hbm_queues=hbm_queues))
    sc.close()
    a = a.astype(np.float64); b = b.astype(np.float64)
    gpu = 0.5 * (a + b)
    gpu_var = 0.25 * (a - b) ** 2                  # unbiased estimate of Var[(a+b)/2] per pixel (1 dof: used in aggregates only)
    diff, tot = gpu - ref, gpu_var + ref_var
    # (1) RMSE with the Monte Carlo noise of both sides removed.  The ~180 pixels straddling the edge of the light carry a
    # binomial variance 1000x the rest (15 or ~0.3 per sample) and would dominate the estimator's own error, so they are
    # judged by their z-scores instead.
    lum_var = ref_var.sum(axis=2)
    body = lum_var < 30 * np.median(lum_var)
    assert 50 < (~body).sum() < 400
    debiased = (diff ** 2 - tot)[body].mean(axis=0)
    se = np.sqrt(2.0 * (tot[body] ** 2).sum(axis=0)) / body.sum()       # standard error of `debiased` itself
    print("spp", spp_g, "debiased MSE", debiased, "its SE", se, "RMSE", np.sqrt(np.maximum(debiased, 0)))
    assert np.all(debiased <= (1e-3) ** 2 + 3.0 * se), (debiased, se)
    z_edge = diff[~body] / np.sqrt(tot[~body] + 1e-30)
    assert np.sqrt((z_edge ** 2).mean()) < 1.6 and np.abs(z_edge).max() < 6.0
    # (2) z-scores aggregated over 16x16 tiles: Z = sum(diff) / sqrt(sum(var)).  (A per-pixel studentised z is biased for
    # this right-skewed estimator because a pixel's variance estimate is correlated with its mean; sums over a tile are not.)
    num = diff.reshape(10, 16, 10, 16, 3).sum(axis=(1, 3))
    den = np.sqrt(tot.reshape(10, 16, 10, 16, 3).sum(axis=(1, 3)))
    Z = num / den
    print("tile Z: rms", np.sqrt((Z ** 2).mean()), "mean", Z.mean(), "max", np.abs(Z).max())
    assert 0.7 < np.sqrt((Z ** 2).mean()) < 1.35          # ~N(0,1): no excess error, no spatial structure
    assert np.abs(Z).max() < 4.5 and abs(Z.mean()) < 0.6
    frac3 = (np.abs(Z) > 3).mean()
    assert frac3 < 0.02                                     # normal tail: 0.27 %
    # (3) energy: global mean within 0.2 %
    assert abs(gpu.mean() - ref.mean()) / ref.mean() < 2e-3


@pytest.mark.gpu
def test_primary_hits_match_committed_fixture(gpu_ctx):
    from pathtrace_b200 import api, scenes
    g = _load("primary_default_96x64.npz")
    sh, mt, cam = scenes.default_scene(96, 64)
    sc = api.Scene(gpu_ctx, sh, mt)
    for trav in ("brute", "bvh"):
        ids, t, d = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal=trav), want_dirs=True)
        assert np.array_equal(ids, g["id_f32"]), trav
        hit = ids >= 0
        assert np.array_equal(t[hit].view(np.uint32), g["t_f32"][hit].view(np.uint32)), trav
        assert np.array_equal(d.view(np.uint32), g["dirs_f32"].view(np.uint32))
        same = hit & (g["id_f64"] == ids)
        rel = np.abs(t[same].astype(np.float64) - g["t_f64"][same]) / g["t_f64"][same]
        assert rel.max() <= 1e-5, rel.max()       # hit distance bar vs the reference's f64 numerics
        assert (g["id_f64"] != ids).mean() < 5e-3  # ids differ from f64 only on shared edges
    got = sc.bvh()
    l = _load("lbvh_default.npz")
    assert np.array_equal(got["nodes"], l["nodes"]) and np.array_equal(got["prims"], l["prims"]) and np.array_equal(got["codes"], l["codes"])
    sc.close()
