"""The compiled host side above the C ABI (host/pathtrace_host.hpp + host/main.cpp, C++17): the reference's World / Object /
Camera API (src/world.rs, src/objects, src/camera.rs) restated in the compiled language this image has (the reference is
Rust; no rustc here).  CPU: it builds against the in-tree library and flattens World::new() to exactly the rows the Python
mirror produces.  GPU: the headless program renders through the C ABI and writes the same bytes as the Python path."""
import os
import subprocess

import numpy as np
import pytest


def _host():
    from pathtrace_b200 import build
    return build.build_host()


def test_cpp_world_new_flattens_like_the_python_mirror():
    from pathtrace_b200 import api, scenes
    out = subprocess.run([_host(), "--describe"], capture_output=True, text=True, check=True).stdout.splitlines()
    sh, mt, cam = scenes.default_scene()
    rows = {(r.split()[0], int(r.split()[1])): r.split()[2:] for r in out if r.startswith(("shape", "material"))}
    assert len(rows) == 2 * len(sh) == 26
    for i in range(len(sh)):
        k, *words = rows[("shape", i)]
        assert int(k) == sh[i]["kind"] and [int(w, 16) for w in words] == sh[i]["v"].view(np.uint32).tolist(), i
        k, *words = rows[("material", i)]
        assert int(k) == mt[i]["kind"] and [int(w, 16) for w in words] == mt[i]["p"].view(np.uint32).tolist(), i
    c = next(r for r in out if r.startswith("camera")).split()[1:]
    cs = api.camera_struct(cam)
    assert (int(c[0]), int(c[1])) == (cs.width, cs.height) == (400, 400)
    want = np.array(list(cs.origin) + list(cs.lower_left_corner) + list(cs.horizontal) + list(cs.vertical), np.float32).view(np.uint32).tolist()
    assert [int(w, 16) for w in c[2:]] == want
    assert next(r for r in out if r.startswith("lights")).split()[1:] == ["10", "11"]       # world.rs:214-225


def test_cpp_host_fails_loudly_without_a_device(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([_host(), "--width", "32", "--height", "32", "--spp", "1", "--out", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 1 and "pt_context_create" in r.stderr        # no CPU fallback
    assert not (tmp_path / "frame.ppm").exists()


@pytest.mark.gpu
@pytest.mark.parametrize("strategy", ["mis", "brdf_only"])
def test_cpp_host_renders_the_same_frame_as_the_python_path(tmp_path, strategy):
    from pathtrace_b200 import api
    w, h, spp = 96, 64, 8
    r = subprocess.run([_host(), "--width", str(w), "--height", str(h), "--spp", str(spp), "--strategy", strategy, "--out", str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Rendering completed" in r.stdout and "Luminance data exported" in r.stdout      # main.rs:63, :67
    world = api.World.new(w, h, spp).render(strategy=strategy)
    # Two renders of the same sample set differ only in the order of the float atomic adds into the pixel sums (last-ulp noise):
    # layout, header and coordinates are compared exactly, the values to one step of what each file format can show.
    ppm = (tmp_path / "frame.ppm").read_bytes()
    head = f"P6\n{w} {h}\n255\n".encode()
    assert ppm.startswith(head) and len(ppm) == len(head) + 3 * w * h
    got = np.frombuffer(ppm[len(head):], np.uint8).astype(np.int16).reshape(h, w, 3)
    assert np.abs(got - world.data[..., :3].astype(np.int16)).max() <= 1
    ref_csv = tmp_path / "ref.csv"
    world.export_luminance(str(ref_csv))
    a = (tmp_path / "luminance.csv").read_text().splitlines()
    b = ref_csv.read_text().splitlines()
    assert a[0] == b[0] and len(a) == len(b) == 1 + w * h                                   # world.rs:352
    va = np.array([ln.split(",") for ln in a[1:]], dtype=np.float64)
    vb = np.array([ln.split(",") for ln in b[1:]], dtype=np.float64)
    assert np.array_equal(va[:, :2], vb[:, :2])
    assert np.abs(va[:, 2:] - vb[:, 2:]).max() <= 1e-5 * max(1.0, np.abs(vb[:, 2:]).max())
    assert all(len(f.split(".")[1]) == 6 for f in a[1].split(",")[2:])                      # {:.6}, world.rs:359-363
