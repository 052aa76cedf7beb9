"""The camera rays' screen-space cull, checked on the CPU (no device): pt_primary_cull_rects returns the pixel rectangle the
library computes for every entry of the constant-bank scan (csrc/pt_wave.cuh: make_cull_param); a warp skips an entry when none
of its pixels lies inside.  That is only exact if EVERY ray the FP32 oracle sees hitting an object starts in a pixel inside
the rectangle of that object's entry — for any camera, any offset inside the pixel cell, and for the hit the oracle would report
if the object were alone in the scene (the cull must not depend on occlusion).  The GPU twin of this test is
tests/test_gpu_parity.py::test_primary_cull_is_exact_for_any_camera.
"""
import numpy as np
import pytest

from conftest import oracle_camera


def _cameras(rng, W, H, n):
    from pathtrace_b200.objects import Camera, Vector3 as V
    cams = [Camera.new(V.new(0, 0, 2), W, H, 1.0, 35.0),
            Camera.new(V.new(0.1, -0.2, -1.6), W, H, 1.0, 80.0),                      # inside the room
            Camera.new(V.new(0, -1.0 + 1e-4, 2), W, H, 1.0, 35.0),                     # the floor seen edge-on
            Camera.new(V.new(-1.0, 0, 2), W, H, 1.0, 50.0),                            # in a wall's plane
            Camera.look_at(V.new(0.4, -0.6, -2.0), V.new(0, 0.5, -3), V.new(0, 1, 0), W, H, 70.0)]   # inside the sphere
    for _ in range(n):
        o = rng.uniform(-3.5, 3.5, 3) + np.array([0, 0, -2.0])
        tgt = rng.uniform(-1, 1, 3) + np.array([0, 0, -2.0])
        up = rng.normal(size=3) * 0.3 + np.array([0, 1, 0])
        if np.linalg.norm(np.cross(tgt - o, up)) < 1e-3:
            continue
        cams.append(Camera.look_at(V.new(*o), V.new(*tgt), V.new(*up), W, H, float(rng.uniform(8, 110))))
    return cams


def _scenes():
    from pathtrace_b200 import scenes
    sh, mt, _ = scenes.default_scene(8, 8)
    out = {"default": (sh, mt)}
    sh, mt, _ = scenes.material_zoo_scene(8, 8)
    out["zoo"] = (sh, mt)
    return out


@pytest.mark.parametrize("scene_name", ["default", "zoo"])
def test_every_oracle_hit_lies_inside_its_entrys_rectangle(scene_name, oracle):
    from pathtrace_b200 import api
    sh, mt = _scenes()[scene_name]
    W, H = 96, 54
    rng = np.random.default_rng(7 if scene_name == "default" else 8)
    n_cam = n_hits = n_culled_entries = n_entries_total = 0
    for cam in _cameras(rng, W, H, 36):
        rects, ent = api.primary_cull_rects(sh, cam)
        assert rects.shape[1] == 4 and len(ent) == len(sh)
        assert ((rects[:, 0] >= 0) & (rects[:, 1] >= 0) & (rects[:, 2] <= W - 1) & (rects[:, 3] <= H - 1)).all()
        ocam = oracle_camera(oracle, cam)
        runs = [dict(jitter=True, sample=5)] + [dict(jitter=False, ox=a, oy=b) for a, b in ((0.0, 0.0), (1.0, 1.0), (0.0, 1.0), (1.0, 0.0))]
        ys, xs = np.mgrid[0:H, 0:W]
        yflip = H - 1 - ys
        for i in range(len(sh)):
            # the object alone: occlusion must not matter to a cull
            one_s, one_m = sh[i:i + 1], mt[i:i + 1]
            x0, y0, x1, y1 = rects[ent[i]]
            inside = (xs >= x0) & (xs <= x1) & (yflip >= y0) & (yflip <= y1)
            for kw in runs:
                ids, _, _ = oracle.trace_primary(ocam, one_s, one_m, precision="f32", **kw)
                hit = ids >= 0
                n_hits += int(hit.sum())
                assert not (hit & ~inside).any(), (scene_name, n_cam, i, kw, rects[ent[i]].tolist(), np.argwhere(hit & ~inside)[:4].tolist())
        full = (rects[:, 0] == 0) & (rects[:, 1] == 0) & (rects[:, 2] == W - 1) & (rects[:, 3] == H - 1)
        n_culled_entries += int((~full).sum()); n_entries_total += len(rects)
        n_cam += 1
    assert n_cam > 30 and n_hits > 100000
    assert n_culled_entries > 0.3 * n_entries_total          # the rectangles do cull: they are not all the full screen


def test_cull_rects_of_the_default_view_are_tight(oracle):
    """Default camera: every rectangle contains the pixels that see its entry, and for the entries wholly on screen it is their
    bounding box up to the 2-pixel pad (+ rounding)."""
    from pathtrace_b200 import api, scenes
    sh, mt, cam = scenes.default_scene(192, 108)
    rects, ent = api.primary_cull_rects(sh, cam)
    ocam = oracle_camera(oracle, cam)
    ys, xs = np.mgrid[0:108, 0:192]
    yflip = 107 - ys
    n_tight = 0
    for e in range(len(rects)):
        objs = np.nonzero(ent == e)[0]
        hit = np.zeros((108, 192), bool)
        for i in objs:
            ids, _, _ = oracle.trace_primary(ocam, sh[i:i + 1], mt[i:i + 1], precision="f32", jitter=False, ox=0.5, oy=0.5)
            hit |= ids >= 0
        assert hit.any()
        bx0, bx1, by0, by1 = xs[hit].min(), xs[hit].max(), yflip[hit].min(), yflip[hit].max()
        x0, y0, x1, y1 = rects[e]
        assert x0 <= bx0 and x1 >= bx1 and y0 <= by0 and y1 >= by1
        if x0 > 0 and y0 > 0 and x1 < 191 and y1 < 107:      # wholly on screen (light, sphere, back wall): nothing of it is clipped away
            n_tight += 1
            # a sphere is bounded through the corners of its box, which project ~20 % wider than its silhouette
            slack = 4 if sh[objs[0]]["kind"] == 1 else 4 + 0.3 * max(bx1 - bx0, by1 - by0)
            assert max(bx0 - x0, x1 - bx1, by0 - y0, y1 - by1) <= slack, (e, rects[e].tolist(), (bx0, by0, bx1, by1))
    assert n_tight >= 3


def test_cull_rects_reject_scenes_without_a_constant_bank_scan():
    from pathtrace_b200 import api, scenes
    from conftest import random_scene
    _, _, cam = scenes.default_scene(32, 32)
    shapes, _ = random_scene(1, 40, tri_frac=1.0)             # 20 pairs > 8
    with pytest.raises(api.PathtraceError, match="no constant-bank scan"):
        api.primary_cull_rects(shapes, cam)
