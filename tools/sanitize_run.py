"""Small workload touching every kernel of libpathtrace_b200.so, for tools/sanitize.sh (compute-sanitizer)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from conftest import random_scene
from pathtrace_b200 import api, scenes
from pathtrace_b200.objects import Camera, Vector3

ctx = api.Context(0)
cam64 = scenes.default_camera(64, 64)
jobs = [("default", scenes.default_scene(64, 64)[:2]), ("zoo", scenes.material_zoo_scene(64, 64)[:2]), ("random300", random_scene(4, 300)),
        ("random5000", random_scene(5, 5000)), ("one", random_scene(7, 1)), ("empty", random_scene(7, 0))]
for name, (sh, mt) in jobs:
    sc = api.Scene(ctx, sh, mt)                     # pack + LBVH build (k_prim_aabb .. k_refit, k_brute_table)
    sc.bvh()
    for trav in ("auto", "bvh", "brute", "brute_exact", "brute_table", "brute_param_checked"):
        if trav.startswith("brute") and len(sh) > 1024:
            continue
        for hbm in (False, True):
            for strat in ("mis", "brdf_only"):
                rad, rgba, st = api.render(ctx, sc, cam64, api.render_params(spp=4, traversal=trav, hbm_queues=hbm, strategy=strat, counters=hbm))
                assert np.isfinite(rad).all()
        api.trace_primary(ctx, sc, cam64, api.render_params(traversal=trav), jitter=True, sample=1, want_dirs=True)
    if name in ("default", "zoo"):
        api.trace_primary(ctx, sc, cam64, api.render_params(traversal="brute_fast"))
        api.render_moments(ctx, sc, cam64, api.render_params(spp=8), 4)
        api.render(ctx, sc, cam64, api.render_params(spp=2, profile=True))
    sc.close()
    print("ok", name, flush=True)
ctx.close()
