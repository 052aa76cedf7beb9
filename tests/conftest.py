import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (oracle/): checker only."""
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


@pytest.fixture(scope="session")
def gpu_ctx():
    from pathtrace_b200 import api
    ctx = api.Context(0)
    yield ctx
    ctx.close()


def oracle_camera(oracle, cam):
    """pathtrace_b200.objects.Camera (f64 fields) -> oracle CameraDesc."""
    c = oracle.CameraDesc()
    for name in ("origin", "lower_left_corner", "horizontal", "vertical"):
        v = getattr(cam, name)
        getattr(c, name)[:] = [v.x, v.y, v.z]
    c.width, c.height = cam.width, cam.height
    return c


def random_scene(seed, n, tri_frac=0.5, with_light=True):
    """n random spheres/triangles in [-1,1]^3 in front of the default camera; mixed materials."""
    from pathtrace_b200.objects import MATERIAL_DTYPE, SHAPE_DTYPE
    rng = np.random.default_rng(seed)
    shapes = np.zeros(n, SHAPE_DTYPE); mats = np.zeros(n, MATERIAL_DTYPE)
    for i in range(n):
        c = rng.uniform(-1, 1, 3) + np.array([0, 0, -2.5])
        if rng.uniform() < tri_frac:
            shapes[i]["kind"] = 1
            shapes[i]["v"] = np.concatenate([c + rng.uniform(-0.3, 0.3, 3) for _ in range(3)]).astype(np.float32)
        else:
            shapes[i]["kind"] = 0
            shapes[i]["v"][:3] = c.astype(np.float32); shapes[i]["v"][3] = np.float32(rng.uniform(0.05, 0.25))
        k = rng.integers(0, 4)
        if i == 0 and with_light:
            k = 1
        mats[i]["kind"] = k
        if k == 0: mats[i]["p"][:3] = rng.uniform(0.1, 0.9, 3)
        elif k == 1: mats[i]["p"][:3] = rng.uniform(2, 10, 3)
        elif k == 2: mats[i]["p"][:6] = [rng.uniform(0.05, 0.6), *rng.uniform(0.5, 1, 3), float(rng.integers(0, 2)), 1.5]
        else:
            r = rng.uniform(0, 1); s2 = r * r
            mats[i]["p"][:6] = [*rng.uniform(0.1, 0.9, 3), r, 1.0 - 0.5 * s2 / (s2 + 0.33), 0.45 * s2 / (s2 + 0.09)]
    return shapes, mats
