"""CPU tests: host-side object API mirror, scene builders, and that the C-ABI library loads and exports every
symbol include/pathtrace_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from pathtrace_b200 import api, build
    build.build_library()
    lib = api.load_library()
    header = open(os.path.join(ROOT, "include", "pathtrace_b200.h")).read()
    declared = set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", header))
    assert declared == set(api.EXPORTS), declared ^ set(api.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.pt_version()


def test_no_device_fails_loudly():
    """No CPU fallback: without a GPU pt_context_create must return an error, not render on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from pathtrace_b200 import api
    with pytest.raises(api.PathtraceError, match="no CUDA device|no CPU fallback"):
        api.Context(0)


def test_struct_layouts_match_header():
    from pathtrace_b200 import api
    from pathtrace_b200.objects import MATERIAL_DTYPE, SHAPE_DTYPE
    assert SHAPE_DTYPE.itemsize == 40 and MATERIAL_DTYPE.itemsize == 36
    assert ctypes.sizeof(api.pt_camera) == 56
    assert ctypes.sizeof(api.pt_render_params) == 48
    p = api.render_params()
    assert (p.spp, p.min_depth, p.max_depth, p.strategy) == (3000, 4, 50, 0)     # world.rs:18, rendering.rs:6-7, Cargo.toml:7
    assert abs(p.t_min - 1e-3) < 1e-9


def test_default_scene_is_world_new():
    """src/world.rs:75-225: 13 objects, lights = [10, 11], glass sphere last."""
    from pathtrace_b200 import scenes
    sh, mt, cam = scenes.default_scene()
    assert len(sh) == 13 and (cam.width, cam.height) == (400, 400)
    assert sh["kind"].tolist() == [1] * 12 + [0]
    assert mt["kind"].tolist() == [0] * 10 + [1, 1, 2]
    assert np.allclose(sh[12]["v"][:4], [0.4, -0.6, -2.0, 0.4])
    assert np.allclose(mt[12]["p"][:6], [0.3, 1, 1, 1, 0.0, 1.5])
    assert np.allclose(sh[10]["v"], [-0.3, 0.99, -2.3, 0.3, 0.99, -2.3, 0.3, 0.99, -1.7])
    assert np.allclose(mt[0]["p"][:3], [0.8, 0.1, 0.1]) and np.allclose(mt[4]["p"][:3], [0.2, 0.2, 0.8])
    assert np.allclose(mt[10]["p"][:3], [15, 15, 15])


def test_oren_nayar_coefficients():
    from pathtrace_b200.objects import OrenNayar
    m = OrenNayar.new((0.5, 0.5, 0.5), 0.4)   # material.rs:182-196
    s2 = 0.16
    assert m.a == pytest.approx(1 - 0.5 * s2 / (s2 + 0.33)) and m.b == pytest.approx(0.45 * s2 / (s2 + 0.09))
    assert m.describe()[1][3:6] == [0.4, m.a, m.b]


def test_random_spheres_scene_shape():
    from pathtrace_b200 import scenes
    sh, mt, cam = scenes.random_spheres_scene(192, 108)
    assert len(sh) == 2 + 484 + 3 + 2 + 4
    assert (mt["kind"] == 1).sum() == 6
    # deterministic
    sh2, mt2, _ = scenes.random_spheres_scene(192, 108)
    assert sh.tobytes() == sh2.tobytes() and mt.tobytes() == mt2.tobytes()


def test_heightfield_fast_path_equals_object_api():
    """The vectorised C4 builder writes the rows Object.new(TriangleShape.new(..), Lambertian..) would."""
    from pathtrace_b200 import scenes
    from pathtrace_b200.objects import LambertianCosineWeighted, Object, TriangleShape, describe_objects
    sh, mt, cam = scenes.heightfield_scene(64, 36, cells=5)
    assert len(sh) == 5 * 5 * 2 + 2
    objs = [Object.new(TriangleShape.new(tuple(map(float, r["v"][0:3])), tuple(map(float, r["v"][3:6])), tuple(map(float, r["v"][6:9]))),
                       LambertianCosineWeighted.new((0.6, 0.6, 0.6))) for r in sh[:50]]
    sh2, mt2 = describe_objects(objs)
    assert sh2.tobytes() == sh[:50].tobytes() and mt2.tobytes() == mt[:50].tobytes()
    full = scenes.heightfield_scene(64, 36)
    assert len(full[0]) == 708 * 708 * 2 + 2 == 1002530


def test_splitmix_vector_twin():
    from pathtrace_b200.scenes import SplitMix64, splitmix64_array
    r = SplitMix64(0x5EED0004)
    a = np.array([r.uniform() for _ in range(100)])
    assert np.array_equal(a, splitmix64_array(0x5EED0004, 100))


def test_luminance_csv_writer_equals_oracle_bytes(tmp_path, oracle):
    """world.rs:344-369.  pt_export_luminance is host code (no device needed): its file equals the oracle's byte for byte,
    Rust's NaN / inf / -inf spellings included."""
    from pathtrace_b200 import api
    rng = np.random.default_rng(5)
    rad = rng.uniform(0, 3, (7, 5, 3)).astype(np.float32)
    rad[0, 0] = [np.nan, np.inf, -np.inf]
    rad[1, 2] = [-0.0, 1e-7, 123456.789]
    rad[2, 2] = [0.0000005, 0.0000015, 2.5]             # exact-decimal rounding of the f32 values promoted to f64
    a, b = os.path.join(tmp_path, "lib.csv"), os.path.join(tmp_path, "oracle.csv")
    api.export_luminance(a, rad)
    oracle.export_luminance(b, rad.astype(np.float64))
    assert open(a, "rb").read() == open(b, "rb").read()
    lines = open(a).read().splitlines()
    assert lines[0] == "x,y,r,g,b,luminance" and lines[1] == "0,0,NaN,inf,-inf,NaN" and len(lines) == 36


def test_header_structs_match_ctypes_mirror():
    """Field order, offsets and sizes of every struct in include/pathtrace_b200.h against the ctypes mirror in api.py, computed
    from the header text with the C layout rules (natural alignment)."""
    from pathtrace_b200 import api
    header = open(os.path.join(ROOT, "include", "pathtrace_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    sizes = {"uint32_t": 4, "int32_t": 4, "uint64_t": 8, "float": 4, "uint8_t": 1}

    def layout(name):
        body = re.search(r"typedef struct \{([^{}]*)\}\s*" + name + r"\s*;", header).group(1)
        fields, off, align = [], 0, 1
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            ty, rest = decl.split(None, 1)
            for item in rest.split(","):
                m = re.match(r"\s*(\w+)(?:\[(\d+)\])?\s*$", item)
                n = int(m.group(2) or 1)
                sz = sizes[ty]
                off = (off + sz - 1) // sz * sz
                fields.append((m.group(1), off, sz * n))
                off += sz * n
                align = max(align, sz)
        return fields, (off + align - 1) // align * align

    for name, mirror in (("pt_camera", api.pt_camera), ("pt_render_params", api.pt_render_params), ("pt_stats", api.pt_stats)):
        fields, size = layout(name)
        assert size == ctypes.sizeof(mirror), (name, size, ctypes.sizeof(mirror))
        assert [f[0] for f in fields] == [f[0] for f in mirror._fields_], name
        for fname, off, sz in fields:
            d = getattr(mirror, fname)
            assert (d.offset, d.size) == (off, sz), (name, fname, off, sz, d.offset, d.size)
