"""Per-kernel CUDA-event times of one render (PT_FLAG_PROFILE) — development helper."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtrace_b200 import api, scenes
W, H, spp, pool = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
scene = sys.argv[5] if len(sys.argv) > 5 else "default"
trav = sys.argv[6] if len(sys.argv) > 6 else "auto"
hbm = len(sys.argv) > 7 and sys.argv[7] == "hbm"
ctx = api.Context(0)
sh, mt, cam = {"default": scenes.default_scene, "spheres": scenes.random_spheres_scene, "height": scenes.heightfield_scene}[scene](W, H)
sc = api.Scene(ctx, sh, mt)
for prof in (False, True, False, True):
    rad, rgba, st = api.render(ctx, sc, cam, api.render_params(spp=spp, path_pool=pool, profile=prof, traversal=trav, hbm_queues=hbm))
    k = st["ms_kernel"]; n = st["n_kernel"]
    print(f"profile={prof} render {st['ms_render']:.2f} ms {st['samples']/st['ms_render']/1e3:.0f} Msamples/s iters {st['wavefront_iters']} | " +
          " ".join(f"{c}={k[c]:.2f}ms/{n[c]}" for c in k) + f" | sum {sum(k.values()):.2f} | phases {st['phase_share']}")
