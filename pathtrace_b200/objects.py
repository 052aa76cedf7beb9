"""Host-side mirror of the reference's scene / object API (constructors and field names unchanged).

Reference (roxas1533/pathtrace, Rust): src/math.rs (Vector3), src/camera.rs (Camera), src/objects/shape.rs
(SphereShape, TriangleShape), src/objects/material.rs (LambertianCosineWeighted, Emissive, OrenNayar),
src/objects/mirror.rs (Mirror), src/objects/object.rs (Object).  The one additive method the reference
needs to cross an FFI boundary is ``describe()``: it flattens a shape / material into the
plain-data ``pt_shape`` / ``pt_material`` rows of include/pathtrace_b200.h.  Scene construction stays on
the host in f64; values are narrowed to f32 once, here.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Sequence

import numpy as np

# Layout-identical to include/pathtrace_b200.h
SHAPE_DTYPE = np.dtype([("kind", "<u4"), ("v", "<f4", (9,))])          # pt_shape, 40 B
MATERIAL_DTYPE = np.dtype([("kind", "<u4"), ("p", "<f4", (8,))])       # pt_material, 36 B
PT_SHAPE_SPHERE, PT_SHAPE_TRIANGLE = 0, 1
PT_MAT_LAMBERTIAN, PT_MAT_EMISSIVE, PT_MAT_MIRROR, PT_MAT_OREN_NAYAR = 0, 1, 2, 3


@dataclass(frozen=True)
class Vector3:
    """src/math.rs:4-8 (three f64)."""
    x: float
    y: float
    z: float

    @staticmethod
    def new(x, y, z): return Vector3(float(x), float(y), float(z))
    @staticmethod
    def zero(): return Vector3(0.0, 0.0, 0.0)
    @staticmethod
    def one(): return Vector3(1.0, 1.0, 1.0)
    def __add__(self, o): return Vector3(self.x + o.x, self.y + o.y, self.z + o.z)
    def __sub__(self, o): return Vector3(self.x - o.x, self.y - o.y, self.z - o.z)
    def __neg__(self): return Vector3(-self.x, -self.y, -self.z)
    def __mul__(self, s):
        if isinstance(s, Vector3): return Vector3(self.x * s.x, self.y * s.y, self.z * s.z)
        return Vector3(self.x * s, self.y * s, self.z * s)
    __rmul__ = __mul__
    def __truediv__(self, s):
        if isinstance(s, Vector3): return Vector3(self.x / s.x, self.y / s.y, self.z / s.z)
        return Vector3(self.x / s, self.y / s, self.z / s)
    def dot(self, o): return self.x * o.x + self.y * o.y + self.z * o.z
    def cross(self, o): return Vector3(self.y * o.z - self.z * o.y, self.z * o.x - self.x * o.z, self.x * o.y - self.y * o.x)
    def length(self): return math.sqrt(self.x * self.x + self.y * self.y + self.z * self.z)
    def normalize(self):
        n = self.length()
        return self / n if n > 0.0 else self
    def luminance(self): return 0.2126 * self.x + 0.7152 * self.y + 0.0722 * self.z
    def tuple(self): return (self.x, self.y, self.z)


def _v(a) -> Vector3:
    return a if isinstance(a, Vector3) else Vector3.new(*a)


# --------------------------------------------------------------------------------------------- shapes
@dataclass(frozen=True)
class SphereShape:
    """src/objects/shape.rs:38-51."""
    center: Vector3
    radius: float

    @staticmethod
    def new(center, radius): return SphereShape(_v(center), float(radius))
    def describe(self):
        return (PT_SHAPE_SPHERE, [*self.center.tuple(), self.radius, 0, 0, 0, 0, 0])


@dataclass(frozen=True)
class TriangleShape:
    """src/objects/shape.rs:148-158."""
    v0: Vector3
    v1: Vector3
    v2: Vector3

    @staticmethod
    def new(v0, v1, v2): return TriangleShape(_v(v0), _v(v1), _v(v2))
    def describe(self):
        return (PT_SHAPE_TRIANGLE, [*self.v0.tuple(), *self.v1.tuple(), *self.v2.tuple()])


# ------------------------------------------------------------------------------------------ materials
@dataclass(frozen=True)
class LambertianCosineWeighted:
    """src/objects/material.rs:67-77."""
    albedo: Vector3

    @staticmethod
    def new(albedo): return LambertianCosineWeighted(_v(albedo))
    def describe(self): return (PT_MAT_LAMBERTIAN, [*self.albedo.tuple(), 0, 0, 0, 0, 0])


@dataclass(frozen=True)
class Emissive:
    """src/objects/material.rs:126-136."""
    emission: Vector3

    @staticmethod
    def new(emission): return Emissive(_v(emission))
    def describe(self): return (PT_MAT_EMISSIVE, [*self.emission.tuple(), 0, 0, 0, 0, 0])


@dataclass(frozen=True)
class Mirror:
    """src/objects/mirror.rs:5-14 — GGX metal (metallic=1) / rough dielectric (metallic=0); struct literal, all fields public."""
    roughness: float
    color: Vector3
    metallic: float
    ior: float

    def __post_init__(self): object.__setattr__(self, "color", _v(self.color))
    def describe(self): return (PT_MAT_MIRROR, [self.roughness, *self.color.tuple(), self.metallic, self.ior, 0, 0])


@dataclass(frozen=True)
class OrenNayar:
    """src/objects/material.rs:166-196 — a, b precomputed in new() exactly as the reference does."""
    albedo: Vector3
    roughness: float
    a: float
    b: float

    @staticmethod
    def new(albedo, roughness):
        s2 = float(roughness) * float(roughness)
        return OrenNayar(_v(albedo), float(roughness), 1.0 - 0.5 * s2 / (s2 + 0.33), 0.45 * s2 / (s2 + 0.09))
    def describe(self): return (PT_MAT_OREN_NAYAR, [*self.albedo.tuple(), self.roughness, self.a, self.b, 0, 0])


@dataclass(frozen=True)
class Object:
    """src/objects/object.rs:9-25 — shape + material."""
    shape: object
    material: object

    @staticmethod
    def new(shape, material): return Object(shape, material)


def describe_objects(objects: Sequence[Object]):
    """Flatten Vec<Object> into (pt_shape[n], pt_material[n]); object id = index, as in World.objects."""
    n = len(objects)
    shapes = np.zeros(n, SHAPE_DTYPE); mats = np.zeros(n, MATERIAL_DTYPE)
    for i, o in enumerate(objects):
        k, v = o.shape.describe(); shapes[i]["kind"] = k; shapes[i]["v"] = np.asarray(v, np.float64).astype(np.float32)
        k, p = o.material.describe(); mats[i]["kind"] = k; mats[i]["p"] = np.asarray(p, np.float64).astype(np.float32)
    return shapes, mats


# --------------------------------------------------------------------------------------------- camera
@dataclass(frozen=True)
class Camera:
    """src/camera.rs:27-148 — pinhole; only the four cached vectors + width/height cross the boundary."""
    origin: Vector3
    width: int
    height: int
    lower_left_corner: Vector3
    horizontal: Vector3
    vertical: Vector3

    @staticmethod
    def new(origin, width, height, screen_distance, fov_degrees):
        """camera.rs:50-82.  The 'horizontal' fov sets the VERTICAL extent (camera.rs:61-62)."""
        origin = _v(origin)
        fov = math.radians(fov_degrees)
        aspect = width / height
        vh = 2.0 * math.tan(fov / 2.0) * screen_distance
        vw = vh * aspect
        hz, vt = Vector3(vw, 0.0, 0.0), Vector3(0.0, vh, 0.0)
        llc = origin - hz / 2.0 - vt / 2.0 - Vector3(0.0, 0.0, screen_distance)
        return Camera(origin, int(width), int(height), llc, hz, vt)

    @staticmethod
    def look_at(origin, target, up, width, height, fov_degrees):
        """camera.rs:94-130."""
        origin, target, up = _v(origin), _v(target), _v(up)
        fov = math.radians(fov_degrees)
        aspect = width / height
        w = (origin - target).normalize(); u = up.cross(w).normalize(); v = w.cross(u)
        vh = 2.0 * math.tan(fov / 2.0) * 1.0
        vw = vh * aspect
        hz, vt = u * vw, v * vh
        llc = origin - hz / 2.0 - vt / 2.0 - w * 1.0
        return Camera(origin, int(width), int(height), llc, hz, vt)
