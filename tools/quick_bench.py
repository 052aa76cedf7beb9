"""Quick device timing of pt_render on the default scene (development helper, not the bench contract)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtrace_b200 import api, scenes

W, H, spp = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
trav = sys.argv[4] if len(sys.argv) > 4 else "auto"
pool = int(sys.argv[5]) if len(sys.argv) > 5 else 0
scene_name = sys.argv[6] if len(sys.argv) > 6 else "default"
ctx = api.Context(0)
if scene_name == "default":
    sh, mt, cam = scenes.default_scene(W, H)
elif scene_name == "spheres":
    sh, mt, cam = scenes.random_spheres_scene(W, H)
else:
    sh, mt, cam = scenes.heightfield_scene(W, H)
t0 = time.time(); sc = api.Scene(ctx, sh, mt); t1 = time.time()
print(f"scene {scene_name}: {len(sh)} objects, {sc.n_lights} lights, build {1e3*(t1-t0):.1f} ms")
for rep in range(3):
    rad, rgba, st = api.render(ctx, sc, cam, api.render_params(spp=spp, traversal=trav, path_pool=pool))
    rays = st["rays_primary"] + st["rays_shadow"] + st["rays_scattered"]
    print(f"rep{rep} {W}x{H}x{spp} trav={trav} pool={pool}: render {st['ms_render']:.1f} ms total {st['ms_total']:.1f} ms  "
          f"{st['samples']/st['ms_render']/1e3:.1f} Msamples/s {rays/st['ms_render']/1e3:.1f} Mrays/s  iters {st['wavefront_iters']} launches {st['kernel_launches']} "
          f"rays/sample {rays/st['samples']:.3f} mean {rad.mean():.5f} nan {np.isnan(rad).sum()}")
os.makedirs("gpurun_out", exist_ok=True)
api.write_ppm(f"gpurun_out/{scene_name}_{W}x{H}.ppm", rgba)
