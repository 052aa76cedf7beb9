"""GPU parity tests (run on a B200 with -m gpu): CUDA path through the C ABI vs the CPU oracle.

Bars (BASELINE.json north_star): LBVH build and primary closest-hit object ids bit-exact (vs the FP32
oracle), hit distances within 1e-5 relative of the f64 oracle.
"""
import numpy as np
import pytest

from conftest import oracle_camera, random_scene

pytestmark = pytest.mark.gpu


def _scenes():
    from pathtrace_b200 import scenes
    from pathtrace_b200.objects import Camera, Vector3
    out = {}
    sh, mt, cam = scenes.default_scene(160, 120)
    out["default"] = (sh, mt, cam)
    sh, mt, cam = scenes.random_spheres_scene(200, 112)
    out["spheres"] = (sh, mt, cam)
    out["zoo"] = scenes.material_zoo_scene(160, 120)
    for seed, n in ((1, 2), (2, 3), (3, 17), (4, 300), (5, 5000)):
        sh, mt = random_scene(seed, n)
        out[f"random{n}"] = (sh, mt, Camera.new(Vector3.new(0, 0, 2), 96, 64, 1.0, 35.0))
    return out


@pytest.mark.parametrize("name", ["default", "spheres", "random2", "random3", "random17", "random300", "random5000"])
def test_lbvh_bit_exact(name, oracle, gpu_ctx):
    from pathtrace_b200 import api
    sh, mt, cam = _scenes()[name]
    ref = oracle.lbvh_build(sh)
    sc = api.Scene(gpu_ctx, sh, mt)
    got = sc.bvh()
    sc.close()
    assert np.array_equal(got["bounds"].view(np.uint32), ref["bounds"].view(np.uint32))
    assert np.array_equal(got["codes"], ref["codes"])
    assert np.array_equal(got["prims"], ref["prims"])
    assert np.array_equal(got["nodes"], ref["nodes"])


@pytest.mark.parametrize("n", [0, 1])
def test_lbvh_trivial(n, oracle, gpu_ctx):
    from pathtrace_b200 import api
    sh, mt = random_scene(7, n)
    ref = oracle.lbvh_build(sh)
    sc = api.Scene(gpu_ctx, sh, mt)
    got = sc.bvh()
    sc.close()
    assert np.array_equal(got["nodes"], ref["nodes"])


@pytest.mark.parametrize("name", ["default", "zoo", "spheres", "random2", "random17", "random300", "random5000"])
@pytest.mark.parametrize("traversal", ["bvh", "brute", "brute_exact", "brute_table", "brute_param_checked"])
@pytest.mark.parametrize("jitter", [False, True])
def test_primary_hits(name, traversal, jitter, oracle, gpu_ctx):
    """`brute` is the scan the render (and the benchmark) runs on that scene: for the default scene and the zoo the triangle
    pairs come from the constant bank with the in-domain pair test (pt_stats.scan_mode 3, closest_hit_param<false>);
    `brute_param_checked` / `brute_table` are its conservative forms; all must give the FP32 oracle's ids and t bit for bit."""
    from pathtrace_b200 import api
    sh, mt, cam = _scenes()[name]
    if traversal.startswith("brute") and len(sh) > 1000:
        pytest.skip("brute force over 5000 objects is covered by the bvh run")
    if traversal in ("brute_exact", "brute_table", "brute_param_checked") and name not in ("default", "zoo", "random17"):
        pytest.skip("the forced scan forms are exercised on the scenes whose pairs fit the constant bank")
    ocam = oracle_camera(oracle, cam)
    sc = api.Scene(gpu_ctx, sh, mt)
    ids, t, dirs = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal=traversal), jitter=jitter, sample=3, want_dirs=True)
    sc.close()
    rid, rt, rdirs = oracle.trace_primary(ocam, sh, mt, precision="f32", jitter=jitter, sample=3, want_dirs=True)
    assert np.array_equal(dirs.view(np.uint32), rdirs.view(np.uint32)), "camera ray directions must be bit-exact"
    assert np.array_equal(ids, rid), f"{(ids != rid).sum()} object-id mismatches vs the FP32 oracle"
    hit = rid >= 0
    assert np.array_equal(t[hit].view(np.uint32), rt[hit].astype(np.float32).view(np.uint32)), "t must be bit-exact vs FP32 oracle"
    assert np.all(np.isinf(t[~hit]))
    # vs the reference's numerics (f64): t within 1e-5 relative wherever both agree on the object
    did, dt, _ = oracle.trace_primary(ocam, sh, mt, precision="f64", jitter=jitter, sample=3)
    same = (did == ids) & hit
    rel = np.abs(t[same].astype(np.float64) - dt[same]) / dt[same]
    # north-star bar: 1e-5 relative.  Held everywhere on the default scene; on sphere-heavy scenes a ray that
    # grazes a sphere's silhouette is ill-conditioned (dt ~ sqrt of the FP32 rounding of the ray direction
    # itself), so there the bar is: >= 99.9 % of pixels within 1e-5 and none beyond 1e-3.
    if name in ("default", "zoo"):
        assert rel.max() <= 1e-5, rel.max()
    else:   # the conditioning-aware form of the bar is asserted in test_gpu_materials.py::test_hit_distance_bar_on_sphere_scenes
        assert (rel > 1e-5).mean() < 1e-3 and rel.max() < 1e-3, ((rel > 1e-5).mean(), rel.max())
    # id disagreements with f64 only on shared edges: a handful
    assert (did != ids).mean() < 2e-2


@pytest.mark.parametrize("scene_name", ["default", "spheres", "random17"])
@pytest.mark.parametrize("strategy", ["mis", "brdf_only"])
def test_smem_wavefront_equals_hbm_wavefront(scene_name, strategy, gpu_ctx):
    """The two wavefront formulations (queues in shared memory, one persistent launch / queues in HBM, one launch per stage
    per bounce) run the same per-(pixel, sample, bounce) paths: identical ray counts, images equal up to float add order."""
    from pathtrace_b200 import api
    sh, mt, cam = _scenes()[scene_name]
    sc = api.Scene(gpu_ctx, sh, mt)
    out = {}
    trav = "brute_exact" if scene_name == "default" else "auto"     # the bit-exact scan on both sides (scan mode 3 / LBVH)
    for hbm in (False, True):
        out[hbm] = api.render(gpu_ctx, sc, cam, api.render_params(spp=48, strategy=strategy, hbm_queues=hbm, sample_offset=5, traversal=trav))
    sc.close()
    (ra, ga, sa), (rb, gb, sb) = out[False], out[True]
    if scene_name != "spheres":     # 495 objects + LBVH do not fit the shared-memory formulation: both runs take the HBM path
        assert sa["n_kernel"]["smem_wavefront"] == 1 and sa["kernel_launches"] == 2    # + k_resolve
    assert sb["n_kernel"]["smem_wavefront"] == 0 and sb["kernel_launches"] > 3
    for k in ("samples", "rays_primary", "rays_shadow", "rays_scattered", "path_vertices"):
        assert sa[k] == sb[k], (k, sa[k], sb[k])
    assert np.isfinite(ra).all() and np.isfinite(rb).all()
    scale = max(float(np.abs(rb).max()), 1.0)
    assert np.abs(ra - rb).max() <= 2e-4 * scale, np.abs(ra - rb).max()
    assert (np.abs(ga.astype(int) - gb.astype(int)) <= 1).all()


@pytest.mark.parametrize("scene_name", ["spheres", "random300", "random5000"])
def test_bvh_stream_render_equals_brute_force_render(scene_name, gpu_ctx):
    """Formulation B walks the LBVH with the warp-persistent stream kernels (nodes in shared memory for the small scenes,
    through L1/L2 for 5000 objects); the brute-force scan of the same scene takes the one-thread-per-slot kernels.  Same
    closest hits and occlusion answers => same paths: identical ray counts, images equal up to float add order."""
    from pathtrace_b200 import api
    sh, mt, cam = _scenes()[scene_name]
    sc = api.Scene(gpu_ctx, sh, mt)
    out = {}
    for trav in ("bvh", "brute"):
        out[trav] = api.render(gpu_ctx, sc, cam, api.render_params(spp=24, hbm_queues=True, traversal=trav, sample_offset=3))
    sc.close()
    (ra, ga, sa), (rb, gb, sb) = out["bvh"], out["brute"]
    assert sa["n_kernel"]["smem_wavefront"] == 0 and sb["n_kernel"]["smem_wavefront"] == 0
    for k in ("samples", "rays_primary", "rays_shadow", "rays_scattered", "path_vertices"):
        assert sa[k] == sb[k], (k, sa[k], sb[k])
    assert np.isfinite(ra).all() and np.isfinite(rb).all()
    scale = max(float(np.abs(rb).max()), 1.0)
    assert np.abs(ra - rb).max() <= 2e-4 * scale, np.abs(ra - rb).max()
    assert (np.abs(ga.astype(int) - gb.astype(int)) <= 1).all()


@pytest.mark.parametrize("hbm_queues", [False, True])
def test_render_smoke_matches_oracle_mean(hbm_queues, oracle, gpu_ctx):
    """Small frame, moderate spp: GPU mean radiance vs f64 oracle (different sample counts, same estimator)."""
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(64, 64)
    ocam = oracle_camera(oracle, cam)
    sc = api.Scene(gpu_ctx, sh, mt)
    spp_g, spp_c = 2048, 256
    rad, rgba, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=spp_g, hbm_queues=hbm_queues))
    sc.close()
    assert np.isfinite(rad).all()
    assert st["samples"] == 64 * 64 * spp_g
    r = oracle.render(ocam, sh, mt, oracle.RenderParams(spp=spp_c, seed=123), "f64", recursive=False)   # iterative form counts vertices
    mean_c = r["sum"] / spp_c
    var_c = np.maximum(r["sumsq"] / spp_c - mean_c ** 2, 0)
    se2 = var_c / spp_c + var_c / spp_g
    mse = ((rad - mean_c) ** 2).mean(axis=(0, 1))
    debiased = mse - se2.mean(axis=(0, 1))
    print("mse", mse, "noise", se2.mean(axis=(0, 1)), "debiased", debiased, "stats", st)
    assert np.all(debiased < (3e-3) ** 2 + 0.15 * se2.mean(axis=(0, 1)))
    # global mean agrees to ~1%
    assert abs(rad.mean() - mean_c.mean()) / mean_c.mean() < 0.01
    # same rays/sample as the oracle's iterative form, statistically
    # same path statistics as the oracle: scattered rays and shaded vertices per sample (the GPU emits a shadow ray
    # only when the vertex survives roulette with a non-zero contribution, so its shadow count is lower by design)
    c = r["counters"]
    n_c = 64 * 64 * spp_c
    assert abs(st["rays_scattered"] / st["samples"] - c["scattered"] / n_c) / (c["scattered"] / n_c) < 0.01
    assert abs(st["path_vertices"] / st["samples"] - c["vertices"] / n_c) / (c["vertices"] / n_c) < 0.01
    assert st["rays_shadow"] <= st["path_vertices"]


@pytest.mark.parametrize("hbm_queues", [False, True])
def test_single_samples_match_fp32_oracle(hbm_queues, oracle, gpu_ctx):
    """Same counter-based RNG on both sides: sample s of pixel p is the SAME path on the GPU and in the FP32 oracle, so
    one-sample images agree pixel by pixel except where a rounding difference flips a discrete decision (roulette,
    lobe choice, an edge hit) — a small fraction, and unbiased."""
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(96, 64)
    ocam = oracle_camera(oracle, cam)
    sc = api.Scene(gpu_ctx, sh, mt)
    agree, total, sum_g, sum_c = 0, 0, 0.0, 0.0
    for s in (0, 1, 17, 4095):
        rad, _, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=1, sample_offset=s, hbm_queues=hbm_queues))
        r = oracle.render(ocam, sh, mt, oracle.RenderParams(spp=1, sample_offset=s), "f32", recursive=False)
        c = r["sum"]
        close = np.all(np.abs(rad - c) <= 2e-3 * np.maximum(np.abs(c), 1e-3), axis=2)
        agree += int(close.sum()); total += close.size
        sum_g += float(rad.sum()); sum_c += float(c.sum())
        n = 96 * 64
        assert st["rays_primary"] == n
        assert abs(st["rays_scattered"] - r["counters"]["scattered"]) <= 0.01 * n
        assert abs(st["path_vertices"] - r["counters"]["vertices"]) <= 0.01 * n
    sc.close()
    print("pixels agreeing:", agree / total, "energy ratio", sum_g / sum_c)
    assert agree / total > 0.995
    assert abs(sum_g / sum_c - 1.0) < 0.01


def test_benchmark_scan_mode_is_the_one_trace_primary_checks(gpu_ctx):
    """The default scene renders through scan mode 4 (primary rays: constant-bank pairs, in-domain pair test; shadow and
    scattered rays: planar entries); `brute_exact`, `brute_param_checked` and `brute_table` select the bit-exact modes 3, 2
    and 1, which trace the same paths as the LBVH walk."""
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(96, 64)
    sc = api.Scene(gpu_ctx, sh, mt)
    out = {}
    for trav, mode in (("auto", 4), ("brute", 4), ("brute_exact", 3), ("brute_param_checked", 2), ("brute_table", 1), ("bvh", 0)):
        rad, _, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=8, traversal=trav))
        assert st["scan_mode"] == mode and st["n_kernel"]["smem_wavefront"] == 1, (trav, st["scan_mode"])
        out[trav] = (rad, st)
    sc.close()
    for trav in ("brute_param_checked", "brute_table", "bvh"):           # the bit-exact forms trace the very same paths
        for k in ("rays_primary", "rays_shadow", "rays_scattered", "path_vertices"):
            assert out[trav][1][k] == out["brute_exact"][1][k], (trav, k)
        assert np.allclose(out[trav][0], out["brute_exact"][0], rtol=1e-4, atol=1e-5)
    assert out["brute"][1]["rays_primary"] == out["brute_exact"][1]["rays_primary"]
    assert abs(out["brute"][1]["rays_scattered"] - out["brute_exact"][1]["rays_scattered"]) <= 1e-3 * out["brute_exact"][1]["rays_scattered"]


@pytest.mark.parametrize("n", [0, 1])
@pytest.mark.parametrize("hbm_queues", [False, True])
def test_forced_bvh_on_trivial_scenes(n, hbm_queues, oracle, gpu_ctx):
    """A tree over fewer than two objects has no inner node: PT_TRAVERSAL_BVH must fall back to the scan instead of walking
    the placeholder node (whose missing child is an out-of-range primitive)."""
    from pathtrace_b200 import api
    from pathtrace_b200.objects import Camera, Vector3
    sh, mt = random_scene(7, n)
    cam = Camera.new(Vector3.new(0, 0, 2), 48, 32, 1.0, 35.0)
    sc = api.Scene(gpu_ctx, sh, mt)
    ids, t, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="bvh"))
    rid, rt, _ = oracle.trace_primary(oracle_camera(oracle, cam), sh, mt, precision="f32")
    assert np.array_equal(ids, rid)
    hit = rid >= 0
    assert np.array_equal(t[hit].view(np.uint32), rt[hit].astype(np.float32).view(np.uint32))
    rad, _, st = api.render(gpu_ctx, sc, cam, api.render_params(spp=4, traversal="bvh", hbm_queues=hbm_queues))
    sc.close()
    assert np.isfinite(rad).all() and st["rays_primary"] == 4 * 48 * 32
    if n == 0:
        assert not rad.any()


def _odd_cameras():
    from pathtrace_b200.objects import Camera, Vector3 as V
    W, H = 192, 108
    return {
        "default": Camera.new(V.new(0, 0, 2), W, H, 1.0, 35.0),
        "inside_room": Camera.new(V.new(0.1, -0.2, -1.6), W, H, 1.0, 80.0),                       # walls extend behind the camera plane
        "oblique": Camera.look_at(V.new(2.5, 1.5, 1.0), V.new(0, -0.3, -2), V.new(0, 1, 0), W, H, 40.0),
        "rolled": Camera.look_at(V.new(-1.5, 0.2, 1.5), V.new(0.4, -0.6, -2), V.new(0.3, 1, 0.1), W, H, 25.0),
        "edge_on_floor": Camera.new(V.new(0, -1.0 + 1e-4, 2), W, H, 1.0, 35.0),                   # the floor's plane passes (almost) through the eye
        "in_left_wall_plane": Camera.new(V.new(-1.0, 0, 2), W, H, 1.0, 50.0),                     # exactly in a wall's plane
        "from_behind": Camera.look_at(V.new(0, 0, -6), V.new(0, 0, -2), V.new(0, 1, 0), W, H, 35.0),
        "looking_away": Camera.look_at(V.new(0, 0, 2), V.new(0, 0, 6), V.new(0, 1, 0), W, H, 35.0),   # everything behind
        "inside_sphere": Camera.look_at(V.new(0.4, -0.6, -2.0), V.new(0, 0.5, -3), V.new(0, 1, 0), W, H, 70.0),
        "telephoto_on_light": Camera.look_at(V.new(0, -0.9, 0.5), V.new(0, 0.99, -2), V.new(0, 1, 0), W, H, 6.0),
    }


@pytest.mark.parametrize("cam_name", list(_odd_cameras()))
@pytest.mark.parametrize("scene_name", ["default", "zoo"])
def test_primary_cull_is_exact_for_any_camera(cam_name, scene_name, oracle, gpu_ctx):
    """The constant-bank scan of camera rays skips entries whose projected pixel rectangle no lane of the warp touches
    (pt_wave.cuh: closest_hit_param_culled).  For every camera pose — including those where the host cannot bound an entry and
    must fall back to the full screen — ids and t equal the scan table's (no cull) and the FP32 oracle's bit for bit."""
    from pathtrace_b200 import api, scenes
    sh, mt, _ = scenes.default_scene(8, 8) if scene_name == "default" else scenes.material_zoo_scene(8, 8)
    cam = _odd_cameras()[cam_name]
    sc = api.Scene(gpu_ctx, sh, mt)
    runs = [dict(jitter=True, sample=s) for s in (0, 1, 7)] + [dict(jitter=False, ox=ox, oy=oy) for ox, oy in ((0.0, 0.0), (1.0, 1.0), (0.5, 0.5), (0.0, 1.0))]
    n_hit = 0
    for kw in runs:
        ids, t, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="brute_exact"), **kw)
        for other in ("brute_table", "brute_param_checked"):
            ids2, t2, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal=other), **kw)
            assert np.array_equal(ids, ids2), (kw, other, int((ids != ids2).sum()))
            assert np.array_equal(t.view(np.uint32), t2.view(np.uint32)), (kw, other)
        n_hit += int((ids >= 0).sum())
    rid, rt, _ = oracle.trace_primary(oracle_camera(oracle, cam), sh, mt, precision="f32", jitter=True, sample=7)
    ids, t, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="brute_exact"), jitter=True, sample=7)
    sc.close()
    assert np.array_equal(ids, rid)
    assert np.array_equal(t[rid >= 0].view(np.uint32), rt[rid >= 0].astype(np.float32).view(np.uint32))
    if cam_name != "looking_away":
        assert n_hit > 0
    else:
        assert n_hit == 0


@pytest.mark.parametrize("pose", ["far", "axis_aligned"])
def test_bvh_stream_walk_is_exact_for_far_and_axis_parallel_rays(pose, gpu_ctx):
    """The shared-memory walk tests boxes with one FMA per plane (pt_wave.cuh: slab_fma): exact for ray origins within four
    scene radii thanks to the padded boxes, and through a slack on the comparison beyond.  A camera 300 scene radii away
    (telephoto) takes the second branch for every primary ray; a camera looking straight down an axis produces rays with
    zero direction components (infinite reciprocals).  Same paths as the brute-force scan in both cases."""
    from pathtrace_b200 import api
    from pathtrace_b200.objects import Camera, Vector3 as V
    sh, mt, _ = _scenes()["random300"]
    if pose == "far":
        cam = Camera.look_at(V.new(300.0, 200.0, 900.0), V.new(0, 0, -2.5), V.new(0, 1, 0), 96, 64, 0.2)
    else:
        cam = Camera.new(V.new(0, 0, 2), 97, 65, 1.0, 35.0)          # odd frame: the centre column / row look exactly along -z
    sc = api.Scene(gpu_ctx, sh, mt)
    out = {}
    for trav in ("bvh", "brute"):
        out[trav] = api.render(gpu_ctx, sc, cam, api.render_params(spp=16, hbm_queues=True, traversal=trav))
    ids, t, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="bvh"))
    ids2, t2, _ = api.trace_primary(gpu_ctx, sc, cam, api.render_params(traversal="brute"))
    sc.close()
    assert np.array_equal(ids, ids2) and np.array_equal(t.view(np.uint32), t2.view(np.uint32))
    assert (ids >= 0).sum() > 50
    (ra, _, sa), (rb, _, sb) = out["bvh"], out["brute"]
    for k in ("samples", "rays_primary", "rays_shadow", "rays_scattered", "path_vertices"):
        assert sa[k] == sb[k], (k, sa[k], sb[k])
    assert np.abs(ra - rb).max() <= 2e-4 * max(float(np.abs(rb).max()), 1.0)
