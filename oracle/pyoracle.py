"""ctypes loader for the CPU oracle (oracle/_build/libpt_oracle.so).

ORACLE / TEST INFRASTRUCTURE.  Import this only from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; the product package pathtrace_b200 never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libpt_oracle.so")

# Layout-identical to pt_shape / pt_material in include/pathtrace_b200.h
SHAPE_DTYPE = np.dtype([("kind", "<u4"), ("v", "<f4", (9,))])
MATERIAL_DTYPE = np.dtype([("kind", "<u4"), ("p", "<f4", (8,))])
# The reference's own f64 scene data (oracle-only: ShapeDescD / MaterialDescD in pt_oracle.hpp; 4 bytes of padding after kind)
SHAPE_DTYPE_F64 = np.dtype({"names": ["kind", "v"], "formats": ["<u4", ("<f8", (9,))], "offsets": [0, 8], "itemsize": 80})
MATERIAL_DTYPE_F64 = np.dtype({"names": ["kind", "p"], "formats": ["<u4", ("<f8", (8,))], "offsets": [0, 8], "itemsize": 72})


class CameraDesc(C.Structure):
    _fields_ = [("origin", C.c_double * 3), ("lower_left_corner", C.c_double * 3), ("horizontal", C.c_double * 3),
                ("vertical", C.c_double * 3), ("width", C.c_uint32), ("height", C.c_uint32)]


class RenderParams(C.Structure):
    _fields_ = [("spp", C.c_uint32), ("sample_offset", C.c_uint32), ("min_depth", C.c_uint32), ("max_depth", C.c_uint32),
                ("strategy", C.c_uint32), ("seed", C.c_uint64), ("t_min", C.c_float)]

    def __init__(self, spp=3000, sample_offset=0, min_depth=4, max_depth=50, strategy=0, seed=0x5EED, t_min=1e-3):
        super().__init__(spp, sample_offset, min_depth, max_depth, strategy, seed, t_min)


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (g++ -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp)."""
    srcs = [os.path.join(_HERE, f) for f in ("pt_oracle_capi.cpp", "pt_oracle.hpp", "pt_lbvh_ref.hpp", "Makefile")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "CXX=g++"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.pto_render.restype = C.c_double
        _lib.pto_render_d.restype = C.c_double
        _lib.pto_shape_hit.restype = C.c_int
        _lib.pto_export_luminance.restype = C.c_int
        _lib.pto_num_threads.restype = C.c_int
    return _lib


def _p(a, t=C.c_void_p):
    return a.ctypes.data_as(t) if a is not None else None


def philox(ctr, key, rounds=10):
    """Philox4x32 with 10 rounds (the published known answers) or 7 (the variant the render draws from)."""
    assert rounds in (7, 10)
    c = np.asarray(ctr, dtype=np.uint32); k = np.asarray(key, dtype=np.uint32); o = np.zeros(4, np.uint32)
    lib().pto_philox4x32(C.c_int(rounds), _p(c), _p(k), _p(o))
    return o


def vec3_op(op, a, b=(0, 0, 0), s=0.0):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64); o = np.zeros(3, np.float64)
    lib().pto_vec3_op(C.c_int(op), _p(a), _p(b), C.c_double(s), _p(o))
    return o


def camera_new(origin, w, h, screen_distance=1.0, fov_deg=35.0) -> CameraDesc:
    o = np.asarray(origin, np.float64); cam = CameraDesc()
    lib().pto_camera_new(_p(o), C.c_uint32(w), C.c_uint32(h), C.c_double(screen_distance), C.c_double(fov_deg), C.byref(cam))
    return cam


def camera_look_at(origin, target, up, w, h, fov_deg) -> CameraDesc:
    o = np.asarray(origin, np.float64); t = np.asarray(target, np.float64); u = np.asarray(up, np.float64); cam = CameraDesc()
    lib().pto_camera_look_at(_p(o), _p(t), _p(u), C.c_uint32(w), C.c_uint32(h), C.c_double(fov_deg), C.byref(cam))
    return cam


def describe_f64(objects):
    """Flatten a list of pathtrace_b200.objects.Object into the f64 rows (no f32 narrowing): the scene exactly as
    world.rs:80-211 holds it."""
    n = len(objects)
    shapes = np.zeros(n, SHAPE_DTYPE_F64); mats = np.zeros(n, MATERIAL_DTYPE_F64)
    for i, o in enumerate(objects):
        k, v = o.shape.describe(); shapes[i]["kind"] = k; shapes[i]["v"] = np.asarray(v, np.float64)
        k, p = o.material.describe(); mats[i]["kind"] = k; mats[i]["p"] = np.asarray(p, np.float64)
    return shapes, mats


def _scene(shapes, mats):
    """-> (shapes, mats, n, suffix): f32 rows (pt_shape / pt_material layout) or the f64 rows of describe_f64()."""
    if shapes.dtype == SHAPE_DTYPE_F64:
        shapes = np.ascontiguousarray(shapes); mats = np.ascontiguousarray(mats, dtype=MATERIAL_DTYPE_F64)
        sfx = "_d"
    else:
        shapes = np.ascontiguousarray(shapes, dtype=SHAPE_DTYPE); mats = np.ascontiguousarray(mats, dtype=MATERIAL_DTYPE)
        sfx = ""
    assert shapes.shape == mats.shape
    return shapes, mats, C.c_uint32(shapes.shape[0]), sfx


def trace_primary(cam, shapes, mats, params=None, precision="f32", jitter=False, ox=0.5, oy=0.5, sample=0, want_dirs=False):
    """Closest hit of one primary ray per pixel: (obj_id int32 [H,W] (-1 = miss), t float64 [H,W], dirs or None)."""
    params = params or RenderParams()
    shapes, mats, n, sfx = _scene(shapes, mats)
    H, W = cam.height, cam.width
    ids = np.empty((H, W), np.int32); t = np.empty((H, W), np.float64)
    dirs = np.empty((H, W, 3), np.float32) if want_dirs else None
    getattr(lib(), "pto_trace_primary" + sfx)(C.c_int(0 if precision == "f64" else 1), C.byref(cam), _p(shapes), _p(mats), n, C.byref(params),
                            C.c_int(int(jitter)), C.c_float(ox), C.c_float(oy), C.c_uint32(sample), _p(ids), _p(t), _p(dirs))
    return ids, t, dirs


def render(cam, shapes, mats, params, precision="f64", recursive=True, nthreads=0, window=None):
    """Returns dict(sum [H,W,3] f64, sumsq [H,W,3] f64, counters dict, seconds).  sum = Σ over spp."""
    shapes, mats, n, sfx = _scene(shapes, mats)
    H, W = cam.height, cam.width
    x0, y0, x1, y1 = window or (0, 0, W, H)
    s = np.zeros((H, W, 3), np.float64); q = np.zeros((H, W, 3), np.float64); cnt = np.zeros(5, np.uint64)
    sec = getattr(lib(), "pto_render" + sfx)(C.c_int(0 if precision == "f64" else 1), C.c_int(int(recursive)), C.byref(cam), _p(shapes), _p(mats), n,
                           C.byref(params), C.c_int(nthreads), C.c_uint32(x0), C.c_uint32(y0), C.c_uint32(x1), C.c_uint32(y1),
                           _p(s), _p(q), _p(cnt))
    names = ("primary", "shadow", "scattered", "retrace", "vertices")
    return dict(sum=s, sumsq=q, counters={k: int(v) for k, v in zip(names, cnt)}, seconds=float(sec))


def sample(cam, shapes, mats, params, x, y, s, precision="f64", recursive=True):
    shapes, mats, n, sfx = _scene(shapes, mats)
    rgb = np.zeros(3, np.float64)
    getattr(lib(), "pto_sample" + sfx)(C.c_int(0 if precision == "f64" else 1), C.c_int(int(recursive)), C.byref(cam), _p(shapes), _p(mats), n,
                     C.byref(params), C.c_uint32(x), C.c_uint32(y), C.c_uint32(s), _p(rgb))
    return rgb


def count_flops(cam, shapes, mats, params, stride=1):
    """Counting-scalar run: dict(flops_per_sample, rays_per_sample, vertices_per_sample, samples)."""
    shapes, mats, n, sfx = _scene(shapes, mats)
    assert sfx == "", "count_flops takes the f32 rows"
    out = np.zeros(10, np.float64)
    lib().pto_count_flops(C.byref(cam), _p(shapes), _p(mats), n, C.byref(params), C.c_uint32(stride), _p(out))
    return dict(flops_per_sample=out[0] / out[1], rays_per_sample=out[2] / out[1], vertices_per_sample=out[3] / out[1],
                samples=int(out[1]), flops_shade_per_vertex=out[4] / max(out[3], 1), flops_closest_per_ray=out[5] / max(out[7] + out[9], 1),
                flops_shadow_per_ray=out[6] / max(out[8], 1), flops_shade=out[4], flops_closest=out[5], flops_shadow=out[6],
                rays_primary=out[7], rays_shadow=out[8], rays_scattered=out[9], vertices=out[3])


def tonemap(mean):
    mean = np.ascontiguousarray(mean, np.float64)
    rgba = np.empty(mean.shape[:-1] + (4,), np.uint8)
    lib().pto_tonemap(_p(mean), C.c_uint64(mean.size // 3), _p(rgba))
    return rgba


def export_luminance(path, mean):
    mean = np.ascontiguousarray(mean, np.float64)
    H, W = mean.shape[:2]
    rc = lib().pto_export_luminance(path.encode(), _p(mean), C.c_uint32(W), C.c_uint32(H))
    if rc:
        raise OSError(f"cannot write {path}")


def lbvh_build(shapes):
    shapes = np.ascontiguousarray(shapes, dtype=SHAPE_DTYPE)
    n = shapes.shape[0]
    codes = np.zeros(n, np.uint32); prims = np.zeros(n, np.uint32)
    nodes = np.zeros((max(n - 1, 1), 16), np.uint32); bounds = np.zeros(6, np.float32)
    lib().pto_lbvh_build(_p(shapes), C.c_uint32(n), _p(codes), _p(prims), _p(nodes), _p(bounds))
    return dict(codes=codes, prims=prims, nodes=nodes, bounds=bounds)


# ---- unit hooks (tests/test_oracle_pinning.py): one routine of shape.rs / material.rs / mirror.rs per call --------------------
def _v3(a):
    return np.ascontiguousarray(a, np.float64)


def _mat_row(kind, p):
    m = np.zeros(1, MATERIAL_DTYPE_F64); m[0]["kind"] = kind; m[0]["p"][:len(p)] = p
    return m


def _shape_row(kind, v):
    s = np.zeros(1, SHAPE_DTYPE_F64); s[0]["kind"] = kind; s[0]["v"][:len(v)] = v
    return s


def bsdf_pdf(kind, p, rd, eta, o, n, precision="f64"):
    """Material::bsdf_pdf -> (f [3], pdf).  rd = incoming ray direction (normalised by Ray::new)."""
    m = _mat_row(kind, p); f = np.zeros(3); pdf = C.c_double()
    lib().pto_bsdf_pdf(C.c_int(0 if precision == "f64" else 1), _p(m), _p(_v3(rd)), C.c_double(eta), _p(_v3(o)), _p(_v3(n)), _p(f), C.byref(pdf))
    return f, pdf.value


def bsdf_pdf_sample(kind, p, rd, eta, n, r1, r2, r3, precision="f64"):
    """Material::bsdf_pdf_sample -> (dir [3], f [3], pdf, cos_theta)."""
    m = _mat_row(kind, p); d = np.zeros(3); f = np.zeros(3); pdf = C.c_double(); ct = C.c_double()
    lib().pto_bsdf_pdf_sample(C.c_int(0 if precision == "f64" else 1), _p(m), _p(_v3(rd)), C.c_double(eta), _p(_v3(n)), C.c_double(r1),
                              C.c_double(r2), C.c_double(r3), _p(d), _p(f), C.byref(pdf), C.byref(ct))
    return d, f, pdf.value, ct.value


def shape_hit(kind, v, ro, rd, t_min, t_max, precision="f64"):
    """Shape::hit + HitRecord::new -> None or dict(t, point, normal, front_face)."""
    s = _shape_row(kind, v); t = C.c_double(); pt = np.zeros(3); nr = np.zeros(3); fr = C.c_int()
    ok = lib().pto_shape_hit(C.c_int(0 if precision == "f64" else 1), _p(s), _p(_v3(ro)), _p(_v3(rd)), C.c_double(t_min), C.c_double(t_max),
                             C.byref(t), _p(pt), _p(nr), C.byref(fr))
    return dict(t=t.value, point=pt, normal=nr, front_face=bool(fr.value)) if ok else None


def shape_sample(kind, v, from_point, r1=0.0, r2=0.0, target=None, precision="f64"):
    """Shape::sample_surface_from_point -> dict(point, normal, pdf_omega, light_dir, d)."""
    s = _shape_row(kind, v); pt = np.zeros(3); nr = np.zeros(3); ld = np.zeros(3); pdf = C.c_double(); d = C.c_double()
    tg = _v3(target) if target is not None else None
    lib().pto_shape_sample(C.c_int(0 if precision == "f64" else 1), _p(s), _p(_v3(from_point)), _p(tg), C.c_double(r1), C.c_double(r2), _p(pt),
                           _p(nr), C.byref(pdf), _p(ld), C.byref(d))
    return dict(point=pt, normal=nr, pdf_omega=pdf.value, light_dir=ld, d=d.value)
