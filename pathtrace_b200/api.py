"""ctypes binding of libpathtrace_b200.so and the host-side ``World`` that mirrors the reference's.

Reference: src/world.rs:54-62 (World{data, luminance_data, camera, objects, light_indices}), :65 new,
:293 render_pixel, :335 draw, :344 export_luminance; src/main.rs:43-60 (the pixel loop this replaces).
Everything numerical happens in the CUDA library; there is NO CPU fallback — a missing library or
device raises PathtraceError.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

from . import build as _build
from .objects import MATERIAL_DTYPE, SHAPE_DTYPE, Camera, describe_objects
from . import scenes as _scenes


class PathtraceError(RuntimeError):
    pass


class pt_camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("lower_left_corner", C.c_float * 3), ("horizontal", C.c_float * 3),
                ("vertical", C.c_float * 3), ("width", C.c_uint32), ("height", C.c_uint32)]


class pt_render_params(C.Structure):
    _fields_ = [("spp", C.c_uint32), ("sample_offset", C.c_uint32), ("min_depth", C.c_uint32), ("max_depth", C.c_uint32),
                ("n_light_samples", C.c_uint32), ("strategy", C.c_uint32), ("traversal", C.c_uint32), ("flags", C.c_uint32),
                ("seed", C.c_uint64), ("t_min", C.c_float), ("path_pool", C.c_uint32)]


class pt_stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_scattered", C.c_uint64),
                ("path_vertices", C.c_uint64), ("wavefront_iters", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("prim_tests", C.c_uint64),
                ("ms_total", C.c_float), ("ms_upload", C.c_float), ("ms_render", C.c_float), ("ms_download", C.c_float),
                ("ms_kernel", C.c_float * 5), ("n_kernel", C.c_uint32 * 5), ("path_pool", C.c_uint32),
                ("phase_share", C.c_float * 3), ("scan_mode", C.c_uint32)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["ms_kernel"] = dict(zip(KERNEL_CLASSES, list(self.ms_kernel)))
        d["n_kernel"] = dict(zip(KERNEL_CLASSES, list(self.n_kernel)))
        d["phase_share"] = dict(zip(("closest_hit", "shade", "wait"), list(self.phase_share)))
        return d


KERNEL_CLASSES = ("generate", "trace_closest", "shade", "trace_shadow", "smem_wavefront")


STRATEGY = {"mis": 0, "brdf_only": 1}
TRAVERSAL = {"auto": 0, "bvh": 1, "brute": 2, "brute_table": 3, "brute_param_checked": 4, "brute_exact": 5, "brute_fast": 6}

# every symbol include/pathtrace_b200.h declares
EXPORTS = ["pt_context_create", "pt_context_destroy", "pt_last_error", "pt_scene_create", "pt_scene_destroy", "pt_render",
           "pt_render_moments", "pt_render_accumulate_device", "pt_resolve_device", "pt_context_stream", "pt_context_synchronize",
           "pt_trace_primary", "pt_bvh_download", "pt_scene_num_objects", "pt_scene_num_lights", "pt_scene_build_ms",
           "pt_scene_build_launches", "pt_export_luminance", "pt_write_ppm", "pt_version", "pt_measure_fp32_peak",
           "pt_group_create", "pt_group_destroy", "pt_group_size", "pt_group_context", "pt_group_last_error",
           "pt_group_scene_create", "pt_group_scene_destroy", "pt_group_render", "pt_render_device_noalloc", "pt_primary_cull_rects"]

_lib = None


def load_library() -> C.CDLL:
    """dlopen the in-tree CUDA library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(_build.LIB):
            raise PathtraceError(f"{_build.LIB} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                 "(there is no CPU fallback)")
        lib = C.CDLL(_build.LIB)
        lib.pt_last_error.restype = C.c_char_p
        lib.pt_last_error.argtypes = [C.c_void_p]
        lib.pt_version.restype = C.c_char_p
        lib.pt_context_stream.restype = C.c_void_p
        lib.pt_context_stream.argtypes = [C.c_void_p]
        for name in ("pt_context_destroy", "pt_scene_destroy"):
            getattr(lib, name).restype = None
            getattr(lib, name).argtypes = [C.c_void_p]
        lib.pt_scene_num_objects.argtypes = [C.c_void_p]
        lib.pt_scene_num_lights.argtypes = [C.c_void_p]
        lib.pt_scene_num_objects.restype = C.c_uint32
        lib.pt_scene_num_lights.restype = C.c_uint32
        lib.pt_scene_build_ms.argtypes = [C.c_void_p]
        lib.pt_scene_build_ms.restype = C.c_float
        lib.pt_scene_build_launches.argtypes = [C.c_void_p]
        lib.pt_scene_build_launches.restype = C.c_uint32
        lib.pt_group_last_error.restype = C.c_char_p
        lib.pt_group_last_error.argtypes = [C.c_void_p]
        lib.pt_group_context.restype = C.c_void_p
        lib.pt_group_context.argtypes = [C.c_void_p, C.c_int32]
        for name in ("pt_group_destroy", "pt_group_scene_destroy"):
            getattr(lib, name).restype = None
            getattr(lib, name).argtypes = [C.c_void_p]
        lib.pt_group_size.argtypes = [C.c_void_p]
        _lib = lib
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def camera_struct(cam: Camera) -> pt_camera:
    """f64 -> f32 narrowing happens once, here."""
    c = pt_camera()
    for name in ("origin", "lower_left_corner", "horizontal", "vertical"):
        v = getattr(cam, name)
        getattr(c, name)[:] = [np.float32(v.x), np.float32(v.y), np.float32(v.z)]
    c.width, c.height = cam.width, cam.height
    return c


def render_params(spp=3000, sample_offset=0, min_depth=4, max_depth=50, strategy="mis", traversal="auto", seed=0x5EED,
                  t_min=1e-3, path_pool=0, profile=False, hbm_queues=False, counters=False, n_light_samples=1, phase_clocks=False) -> pt_render_params:
    s = STRATEGY[strategy] if isinstance(strategy, str) else int(strategy)
    t = TRAVERSAL[traversal] if isinstance(traversal, str) else int(traversal)
    flags = (1 if profile else 0) | (2 if hbm_queues else 0) | (4 if counters else 0) | (8 if phase_clocks else 0)
    return pt_render_params(spp, sample_offset, min_depth, max_depth, n_light_samples, s, t, flags, seed, t_min, path_pool)


class Context:
    """One CUDA device (one process per GPU)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.pt_context_create(C.c_int32(device), C.byref(h))
        if rc:
            raise PathtraceError(f"pt_context_create failed ({rc}): {self.lib.pt_last_error(None).decode()}")
        self.handle = h
        self.device = device

    def check(self, rc, what):
        if rc:
            raise PathtraceError(f"{what} failed ({rc}): {self.lib.pt_last_error(self.handle).decode()}")

    @property
    def stream(self) -> int:
        return self.lib.pt_context_stream(self.handle)

    def synchronize(self):
        self.check(self.lib.pt_context_synchronize(self.handle), "pt_context_synchronize")

    def close(self):
        if getattr(self, "handle", None):
            self.lib.pt_context_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Scene:
    """Device-resident scene + LBVH (replaces World.objects / light_indices, world.rs:58-61)."""

    def __init__(self, ctx: Context, shapes: np.ndarray, mats: np.ndarray):
        self.ctx = ctx
        shapes = np.ascontiguousarray(shapes, dtype=SHAPE_DTYPE)
        mats = np.ascontiguousarray(mats, dtype=MATERIAL_DTYPE)
        if shapes.shape != mats.shape or shapes.ndim != 1:
            raise PathtraceError("shapes and materials must be 1-D arrays of equal length")
        h = C.c_void_p()
        ctx.check(ctx.lib.pt_scene_create(ctx.handle, _ptr(shapes), _ptr(mats), C.c_uint32(shapes.shape[0]), C.byref(h)), "pt_scene_create")
        self.handle = h
        self.n_objects = shapes.shape[0]

    @property
    def n_lights(self):
        return self.ctx.lib.pt_scene_num_lights(self.handle)

    @property
    def build_ms(self) -> float:
        """Device time of scene packing + LBVH build inside pt_scene_create (CUDA events)."""
        return float(self.ctx.lib.pt_scene_build_ms(self.handle))

    @property
    def build_launches(self) -> int:
        return int(self.ctx.lib.pt_scene_build_launches(self.handle))

    def bvh(self):
        n = self.n_objects
        codes = np.zeros(n, np.uint32); prims = np.zeros(n, np.uint32)
        nodes = np.zeros((max(n - 1, 1), 16), np.uint32); bounds = np.zeros(6, np.float32)
        self.ctx.check(self.ctx.lib.pt_bvh_download(self.ctx.handle, self.handle, _ptr(codes), _ptr(prims), _ptr(nodes), _ptr(bounds)), "pt_bvh_download")
        return dict(codes=codes, prims=prims, nodes=nodes, bounds=bounds)

    def close(self):
        if getattr(self, "handle", None) and getattr(self.ctx, "handle", None):
            self.ctx.lib.pt_scene_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render(ctx: Context, scene: Scene, cam: Camera, params: pt_render_params, want_radiance=True, want_rgba=True):
    """One pt_render call with HOST buffers: returns (radiance [H,W,3] f32 | None, rgba [H,W,4] u8 | None, stats dict)."""
    H, W = cam.height, cam.width
    rad = np.empty((H, W, 3), np.float32) if want_radiance else None
    rgba = np.empty((H, W, 4), np.uint8) if want_rgba else None
    st = pt_stats()
    c = camera_struct(cam)
    ctx.check(ctx.lib.pt_render(ctx.handle, scene.handle, C.byref(c), C.byref(params), _ptr(rad), _ptr(rgba), C.byref(st)), "pt_render")
    return rad, rgba, st.as_dict()


def render_moments(ctx: Context, scene: Scene, cam: Camera, params: pt_render_params, batches: int):
    """pt_render_moments: (radiance [H,W,3], moment2 [H,W,3] = sum over batches of (batch sum)^2, stats)."""
    H, W = cam.height, cam.width
    rad = np.empty((H, W, 3), np.float32); m2 = np.empty((H, W, 3), np.float32)
    st = pt_stats()
    c = camera_struct(cam)
    ctx.check(ctx.lib.pt_render_moments(ctx.handle, scene.handle, C.byref(c), C.byref(params), C.c_uint32(batches), _ptr(rad), _ptr(m2), C.byref(st)),
              "pt_render_moments")
    return rad, m2, st.as_dict()


def variance_of_mean(rad: np.ndarray, m2: np.ndarray, spp: int, batches: int) -> np.ndarray:
    """Per-pixel variance of the mean radiance from pt_render_moments' outputs (batches - 1 degrees of freedom)."""
    B, n = batches, spp // batches
    mean_batch_sum = rad.astype(np.float64) * spp / B
    var_batch_sum = np.maximum(m2.astype(np.float64) / B - mean_batch_sum ** 2, 0.0) * B / (B - 1)
    return var_batch_sum / B / n ** 2


def render_accumulate_device(ctx: Context, scene: Scene, cam: Camera, params: pt_render_params, d_accum_ptr: int):
    """Adds this call's samples into a device float4 sum buffer (multi-GPU spp split)."""
    st = pt_stats()
    c = camera_struct(cam)
    ctx.check(ctx.lib.pt_render_accumulate_device(ctx.handle, scene.handle, C.byref(c), C.byref(params), C.c_void_p(d_accum_ptr), C.byref(st)),
              "pt_render_accumulate_device")
    return st.as_dict()


def resolve_device(ctx: Context, d_accum_ptr: int, width: int, height: int, total_spp: int, d_radiance_ptr: int = 0, d_rgba_ptr: int = 0):
    ctx.check(ctx.lib.pt_resolve_device(ctx.handle, C.c_void_p(d_accum_ptr), C.c_uint32(width), C.c_uint32(height), C.c_uint32(total_spp),
                                        C.c_void_p(d_radiance_ptr or None), C.c_void_p(d_rgba_ptr or None)), "pt_resolve_device")


NOALLOC_MAX_PIXELS = 65536       # PT_NOALLOC_MAX_PIXELS


def render_device_noalloc(ctx: Context, scene: Scene, cam: Camera, params: pt_render_params, d_radiance_ptr: int = 0, d_rgba_ptr: int = 0):
    """pt_render_device_noalloc: renders into the caller's device buffers with the library's static scratch array (no
    cudaMalloc on the path); enqueued on ctx.stream, not waited for."""
    c = camera_struct(cam)
    ctx.check(ctx.lib.pt_render_device_noalloc(ctx.handle, scene.handle, C.byref(c), C.byref(params), C.c_void_p(d_radiance_ptr or None),
                                               C.c_void_p(d_rgba_ptr or None)), "pt_render_device_noalloc")


def primary_cull_rects(shapes: np.ndarray, cam: Camera):
    """pt_primary_cull_rects (host only, no device): (rects[n_entries, 4] = x0, y0, x1, y1 inclusive with y from the bottom row,
    entry_of_object[n])."""
    lib = load_library()
    shapes = np.ascontiguousarray(shapes, dtype=SHAPE_DTYPE)
    n = shapes.shape[0]
    rects = np.zeros((max(n, 1), 4), np.int32); ent = np.zeros(max(n, 1), np.int32); ne = C.c_uint32(0)
    c = camera_struct(cam)
    rc = lib.pt_primary_cull_rects(_ptr(shapes), C.c_uint32(n), C.byref(c), _ptr(rects), _ptr(ent), C.byref(ne))
    if rc:
        raise PathtraceError(f"pt_primary_cull_rects failed ({rc}): {lib.pt_last_error(None).decode()}")
    return rects[:ne.value].copy(), ent[:n].copy()


def trace_primary(ctx: Context, scene: Scene, cam: Camera, params: pt_render_params = None, jitter=False, ox=0.5, oy=0.5, sample=0,
                  want_dirs=False):
    params = params or render_params()
    H, W = cam.height, cam.width
    ids = np.empty((H, W), np.int32); t = np.empty((H, W), np.float32)
    dirs = np.empty((H, W, 3), np.float32) if want_dirs else None
    c = camera_struct(cam)
    ctx.check(ctx.lib.pt_trace_primary(ctx.handle, scene.handle, C.byref(c), C.byref(params), C.c_int32(int(jitter)), C.c_float(ox),
                                       C.c_float(oy), C.c_uint32(sample), _ptr(ids), _ptr(t), _ptr(dirs)), "pt_trace_primary")
    return ids, t, dirs


class Group:
    """pt_group_*: one host process driving several GPUs; spp split + one ncclReduce inside the library."""

    def __init__(self, device_ids):
        self.lib = load_library()
        ids = (C.c_int32 * len(device_ids))(*device_ids)
        h = C.c_void_p()
        rc = self.lib.pt_group_create(ids, C.c_int32(len(device_ids)), C.byref(h))
        if rc:
            raise PathtraceError(f"pt_group_create failed ({rc}): {self.lib.pt_group_last_error(None).decode()}")
        self.handle = h
        self.scene = None

    def check(self, rc, what):
        if rc:
            raise PathtraceError(f"{what} failed ({rc}): {self.lib.pt_group_last_error(self.handle).decode()}")

    @property
    def size(self):
        return self.lib.pt_group_size(self.handle)

    def scene_create(self, shapes, mats):
        shapes = np.ascontiguousarray(shapes, dtype=SHAPE_DTYPE); mats = np.ascontiguousarray(mats, dtype=MATERIAL_DTYPE)
        h = C.c_void_p()
        self.check(self.lib.pt_group_scene_create(self.handle, _ptr(shapes), _ptr(mats), C.c_uint32(shapes.shape[0]), C.byref(h)), "pt_group_scene_create")
        return h

    def scene_destroy(self, h):
        self.lib.pt_group_scene_destroy(h)

    def render(self, scene_handle, cam: Camera, params: pt_render_params, want_radiance=True, want_rgba=True):
        H, W = cam.height, cam.width
        rad = np.empty((H, W, 3), np.float32) if want_radiance else None
        rgba = np.empty((H, W, 4), np.uint8) if want_rgba else None
        st = pt_stats()
        c = camera_struct(cam)
        self.check(self.lib.pt_group_render(self.handle, scene_handle, C.byref(c), C.byref(params), _ptr(rad), _ptr(rgba), C.byref(st)), "pt_group_render")
        return rad, rgba, st.as_dict()

    def close(self):
        if getattr(self, "handle", None):
            self.lib.pt_group_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def measure_fp32_peak(ctx: Context):
    """(TFLOP/s, ms) of the FFMA microbenchmark on this device."""
    tf, ms = C.c_float(), C.c_float()
    ctx.check(ctx.lib.pt_measure_fp32_peak(ctx.handle, C.byref(tf), C.byref(ms)), "pt_measure_fp32_peak")
    return tf.value, ms.value


def export_luminance(path: str, radiance: np.ndarray):
    radiance = np.ascontiguousarray(radiance, np.float32)
    H, W = radiance.shape[:2]
    rc = load_library().pt_export_luminance(path.encode(), _ptr(radiance), C.c_uint32(W), C.c_uint32(H))
    if rc:
        raise PathtraceError(f"pt_export_luminance failed ({rc})")


def write_ppm(path: str, rgba: np.ndarray):
    rgba = np.ascontiguousarray(rgba, np.uint8)
    H, W = rgba.shape[:2]
    rc = load_library().pt_write_ppm(path.encode(), _ptr(rgba), C.c_uint32(W), C.c_uint32(H))
    if rc:
        raise PathtraceError(f"pt_write_ppm failed ({rc})")


@dataclass
class World:
    """Mirror of the reference's World (src/world.rs:54-62).  ``World.new()`` builds the default Cornell box."""
    camera: Camera
    objects: list
    width: int = _scenes.WIDTH
    height: int = _scenes.HEIGHT
    sample_num: int = _scenes.SAMPLE_NUM
    data: np.ndarray = None             # RGBA8 [H,W,4]  == World.data
    luminance_data: np.ndarray = None   # linear mean radiance [H,W,3] == World.luminance_data
    stats: dict = None

    @staticmethod
    def new(width=_scenes.WIDTH, height=_scenes.HEIGHT, sample_num=_scenes.SAMPLE_NUM):
        return World(_scenes.default_camera(width, height), _scenes.default_objects(), width, height, sample_num)

    def render(self, device=0, strategy="mis", **kw):
        """Replaces the render thread body, src/main.rs:43-60: one pt_render call for the whole frame."""
        ctx = Context(device)
        try:
            shapes, mats = describe_objects(self.objects)
            scene = Scene(ctx, shapes, mats)
            self.luminance_data, self.data, self.stats = render(ctx, scene, self.camera, render_params(spp=self.sample_num, strategy=strategy, **kw))
            scene.close()
        finally:
            ctx.close()
        return self

    def draw(self, frame: np.ndarray):
        """src/world.rs:335-341: copy RGBA8 into a caller frame."""
        frame.reshape(-1)[:] = self.data.reshape(-1)

    def export_luminance(self, path: str):
        """src/world.rs:344-369."""
        export_luminance(path, self.luminance_data)
