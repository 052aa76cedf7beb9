"""Per-kernel-class device time of one frame (PT_FLAG_PROFILE), development helper: python tools/kernel_split.py C3|C4 [spp]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pathtrace_b200 import api, scenes
cfg = sys.argv[1]; spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
ctx = api.Context(0)
sh, mt, cam = scenes.random_spheres_scene(3840, 2160) if cfg == "C3" else scenes.heightfield_scene(1920, 1080)
sc = api.Scene(ctx, sh, mt)
api.render(ctx, sc, cam, api.render_params(spp=4), want_radiance=False)
for prof in (False, True):
    _, _, st = api.render(ctx, sc, cam, api.render_params(spp=spp, profile=prof), want_radiance=False)
    rays = st["rays_primary"] + st["rays_shadow"] + st["rays_scattered"]
    print(f"{cfg} {spp}spp profile={prof}: render {st['ms_render']:.1f} ms {st['samples']/st['ms_render']/1e3:.0f} Msamples/s {rays/st['ms_render']/1e3:.0f} Mrays/s launches {st['kernel_launches']} iters {st['wavefront_iters']} pool {st['path_pool']}")
    if prof:
        print("  ms by kernel class:", {k: round(v, 1) for k, v in st["ms_kernel"].items()}, "launches", st["n_kernel"])
        print(f"  rays: primary {st['rays_primary']/1e6:.1f}M shadow {st['rays_shadow']/1e6:.1f}M scattered {st['rays_scattered']/1e6:.1f}M vertices {st['path_vertices']/1e6:.1f}M")
