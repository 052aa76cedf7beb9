"""Generates the committed golden fixtures from the CPU oracle (run from the repo root: python tests/golden/make_golden.py).

The reference (Rust, needs a window) cannot run anywhere in this pipeline and its own tests pin nothing on this path
except Vector3, so these fixtures are outputs of the ORACLE, not of the reference: they pin the oracle
against drift and give the GPU tests a converged image without rendering it on the GPU box's CPU.  "Parity unpinned"
 still applies to what the oracle itself restates.

  primary_default_96x64.npz   closest-hit ids / t of one primary ray per pixel, FP32 and f64 oracle, pixel centres
  lbvh_default.npz            CPU-twin LBVH of the default scene (sorted codes, prim order, nodes, bounds)
  samples_default_f64.npz     radiance of 48 individual (x, y, sample) paths, f64, recursive and iterative forms
  cornell_160x160_f64.npz     converged image: mean and variance-of-the-mean per pixel/channel, f64 oracle, 8192 spp
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

REF_W, REF_H, REF_SPP, REF_SEED = 160, 160, 8192, 0x601D


def main():
    O.build()
    only = sys.argv[1] if len(sys.argv) > 1 else None      # e.g. "samples": regenerate just that fixture
    if only not in (None, "samples"):
        raise SystemExit("usage: make_golden.py [samples]")
    # 1. primary hits
    sh, mt, cam = scenes.default_scene(96, 64)
    oc = oracle_camera(O, cam)
    id32, t32, d32 = O.trace_primary(oc, sh, mt, precision="f32", want_dirs=True)
    id64, t64, _ = O.trace_primary(oc, sh, mt, precision="f64")
    np.savez_compressed(os.path.join(HERE, "primary_default_96x64.npz"), id_f32=id32, t_f32=t32.astype(np.float32), dirs_f32=d32,
                        id_f64=id64, t_f64=t64)
    # 2. LBVH twin
    b = O.lbvh_build(sh)
    np.savez_compressed(os.path.join(HERE, "lbvh_default.npz"), **b)
    # 3. individual paths
    sh, mt, cam = scenes.default_scene(64, 64)
    oc = oracle_camera(O, cam)
    rng = np.random.default_rng(12345)
    xs = rng.integers(8, 56, 48); ys = rng.integers(8, 56, 48); ss = rng.integers(0, 4096, 48)
    prm = O.RenderParams(spp=1, seed=0x5EED)
    rec = np.array([O.sample(oc, sh, mt, prm, int(x), int(y), int(s), "f64", True) for x, y, s in zip(xs, ys, ss)])
    itr = np.array([O.sample(oc, sh, mt, prm, int(x), int(y), int(s), "f64", False) for x, y, s in zip(xs, ys, ss)])
    np.savez_compressed(os.path.join(HERE, "samples_default_f64.npz"), x=xs, y=ys, s=ss, recursive=rec, iterative=itr)
    if only == "samples":
        return
    # 4. converged image
    sh, mt, cam = scenes.default_scene(REF_W, REF_H)
    oc = oracle_camera(O, cam)
    t0 = time.time()
    r = O.render(oc, sh, mt, O.RenderParams(spp=REF_SPP, seed=REF_SEED), "f64", recursive=False)
    mean = r["sum"] / REF_SPP
    var_mean = np.maximum(r["sumsq"] / REF_SPP - mean ** 2, 0.0) / REF_SPP
    np.savez_compressed(os.path.join(HERE, "cornell_160x160_f64.npz"), mean=mean.astype(np.float32), var_of_mean=var_mean.astype(np.float32),
                        spp=REF_SPP, seed=REF_SEED)
    print(f"reference image: {REF_W}x{REF_H}x{REF_SPP} in {time.time() - t0:.1f} s, mean {mean.mean():.5f}, "
          f"mean standard error {np.sqrt(var_mean).mean():.2e}")


if __name__ == "__main__":
    main()
