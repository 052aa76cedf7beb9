"""Scene upload + LBVH build timing of config C4's 1 M-triangle heightfield (development helper)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtrace_b200 import api, scenes
ctx = api.Context(0)
sh, mt, cam = scenes.heightfield_scene(1920, 1080) if len(sys.argv) < 2 else scenes.random_spheres_scene(1920, 1080)
for rep in range(3):
    t0 = time.time(); sc = api.Scene(ctx, sh, mt); t1 = time.time()
    print(f"{len(sh)} objects: pt_scene_create {1e3 * (t1 - t0):.1f} ms wall, pack + LBVH build {sc.build_ms:.2f} ms on the device, {sc.build_launches} launches")
    sc.close()
