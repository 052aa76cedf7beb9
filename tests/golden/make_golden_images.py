"""Converged f64 oracle images of the parity scenes beyond the default one (run from the repo root:
python tests/golden/make_golden_images.py [name ...]).  About 15 minutes on 8 cores.

Same caveat as make_golden.py: these are outputs of the ORACLE (the reference cannot run here); they give the GPU tests a
converged image of every material / light / strategy combination the object API can build.  The oracle is fed the scene in
f64 exactly as world.rs would hold it (oracle.describe_f64: no f32 narrowing); the GPU gets the f32 rows.

  zoo_128x128_{mis,brdf}_f64.npz       pathtrace_b200.scenes.material_zoo_scene: Oren-Nayar, metallic + dielectric GGX, sphere + triangle lights
  spheres_200x112_{mis,brdf}_f64.npz   config C3's 495-object random-spheres scene
each: mean, var_of_mean (float32 [H,W,3]), spp, seed
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from conftest import oracle_camera  # noqa: E402
from oracle import pyoracle as O  # noqa: E402
from pathtrace_b200 import scenes  # noqa: E402

JOBS = {
    "zoo_128x128_mis": (scenes.material_zoo_objects, lambda: scenes.material_zoo_scene(128, 128)[2], 0, 8192),
    "zoo_128x128_brdf": (scenes.material_zoo_objects, lambda: scenes.material_zoo_scene(128, 128)[2], 1, 16384),
    "spheres_200x112_mis": (scenes.random_spheres_objects, lambda: scenes.random_spheres_scene(200, 112)[2], 0, 2048),
    "spheres_200x112_brdf": (scenes.random_spheres_objects, lambda: scenes.random_spheres_scene(200, 112)[2], 1, 4096),
}
SEED = 0x601D


def main():
    O.build()
    for name in (sys.argv[1:] or JOBS):
        objects, camera, strategy, spp = JOBS[name]
        sh, mt = O.describe_f64(objects())
        oc = oracle_camera(O, camera())
        t0 = time.time()
        r = O.render(oc, sh, mt, O.RenderParams(spp=spp, seed=SEED, strategy=strategy), "f64", recursive=False)
        mean = r["sum"] / spp
        var_mean = np.maximum(r["sumsq"] / spp - mean ** 2, 0.0) / (spp - 1)
        np.savez_compressed(os.path.join(HERE, name + "_f64.npz"), mean=mean.astype(np.float32), var_of_mean=var_mean.astype(np.float32),
                            spp=spp, seed=SEED, strategy=strategy)
        print(f"{name}: {spp} spp in {time.time() - t0:.1f} s, mean {mean.mean():.5f}, mean standard error {np.sqrt(var_mean).mean():.2e}", flush=True)


if __name__ == "__main__":
    main()
