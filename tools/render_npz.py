"""Render the default scene and save linear radiance to an .npz (development helper for image-parity analysis)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pathtrace_b200 import api, scenes
W, H, spp, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
strategy = sys.argv[5] if len(sys.argv) > 5 else "mis"
ctx = api.Context(0)
sh, mt, cam = scenes.default_scene(W, H)
sc = api.Scene(ctx, sh, mt)
a, _, st = api.render(ctx, sc, cam, api.render_params(spp=spp // 2, sample_offset=0, strategy=strategy))
b, _, _ = api.render(ctx, sc, cam, api.render_params(spp=spp // 2, sample_offset=spp // 2, strategy=strategy))
np.savez_compressed(out, a=a, b=b, spp=spp)
print(st["ms_render"], "ms", a.mean(), b.mean())
