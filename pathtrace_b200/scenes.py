"""Scene builders: the reference's default scene and the synthetic benchmark scenes of BASELINE.json.

default_scene()      — World::new's 13-object Cornell box, src/world.rs:65-241
material_zoo()       — the same box with every material model and both light shapes the object API can build (parity scene)
random_spheres()     — config C3 (~500 spheres, mixed diffuse/metal/dielectric), SURVEY §8(d)
heightfield()        — config C4 (1,002,528 triangles + emissive quad), SURVEY §8(d)
Each returns (objects-or-arrays, Camera factory); all go through the object API of objects.py.
"""
from __future__ import annotations

import numpy as np

from .objects import (MATERIAL_DTYPE, PT_MAT_EMISSIVE, PT_MAT_LAMBERTIAN, PT_SHAPE_TRIANGLE, SHAPE_DTYPE, Camera, Emissive,
                      LambertianCosineWeighted, Mirror, Object, OrenNayar, SphereShape, TriangleShape, Vector3, describe_objects)

V = Vector3.new

# world.rs:16-18
WIDTH, HEIGHT, SAMPLE_NUM = 400, 400, 3000


def default_objects():
    """src/world.rs:75-211, object order preserved (object id = index)."""
    bs, bd, ls = 1.0, -2.0, 0.3
    tri = lambda a, b, c, m: Object.new(TriangleShape.new(V(*a), V(*b), V(*c)), m)
    red = LambertianCosineWeighted.new(V(0.8, 0.1, 0.1)); green = LambertianCosineWeighted.new(V(0.1, 0.8, 0.1))
    back = LambertianCosineWeighted.new(V(0.2, 0.2, 0.8)); floor = LambertianCosineWeighted.new(V(0.2, 0.8, 0.8))
    ceil = LambertianCosineWeighted.new(V(0.8, 0.8, 0.8)); light = Emissive.new(V(15.0, 15.0, 15.0))
    zb, zf, ly = bd - bs, bd + bs, bs - 0.01
    return [
        tri((-bs, -bs, zb), (-bs, bs, zb), (-bs, bs, zf), red),        # 0 left wall
        tri((-bs, -bs, zb), (-bs, bs, zf), (-bs, -bs, zf), red),       # 1
        tri((bs, -bs, zb), (bs, bs, zf), (bs, bs, zb), green),         # 2 right wall
        tri((bs, -bs, zb), (bs, -bs, zf), (bs, bs, zf), green),        # 3
        tri((-bs, -bs, zb), (bs, -bs, zb), (bs, bs, zb), back),        # 4 back wall
        tri((-bs, -bs, zb), (bs, bs, zb), (-bs, bs, zb), back),        # 5
        tri((-bs, -bs, zb), (bs, -bs, zf), (bs, -bs, zb), floor),      # 6 floor
        tri((-bs, -bs, zb), (-bs, -bs, zf), (bs, -bs, zf), floor),     # 7
        tri((-bs, bs, zb), (bs, bs, zb), (bs, bs, zf), ceil),          # 8 ceiling
        tri((-bs, bs, zb), (bs, bs, zf), (-bs, bs, zf), ceil),         # 9
        tri((-ls, ly, bd - ls), (ls, ly, bd - ls), (ls, ly, bd + ls), light),    # 10 light
        tri((-ls, ly, bd - ls), (ls, ly, bd + ls), (-ls, ly, bd + ls), light),   # 11 light
        Object.new(SphereShape.new(V(0.4, -0.6, bd), 0.4), Mirror(roughness=0.3, color=V(1.0, 1.0, 1.0), metallic=0.0, ior=1.5)),  # 12 glass
    ]


def default_camera(width=WIDTH, height=HEIGHT):
    """world.rs:67-73."""
    return Camera.new(V(0.0, 0.0, 2.0), width, height, 1.0, 35.0)


def default_scene(width=WIDTH, height=HEIGHT):
    shapes, mats = describe_objects(default_objects())
    return shapes, mats, default_camera(width, height)


def material_zoo_objects():
    """A 16-object parity scene that puts every constructible material and both light shapes in front of the oracle:
    Lambertian walls (material.rs:67-123), Oren-Nayar floor (sigma 0.8) and back wall (sigma 0.2) (material.rs:166-296),
    a metallic GGX sphere with coloured F0 (mirror.rs:126-132, metallic > 0.99 branch :187/:227), a rough-glass sphere
    (metallic 0), a rough metallic triangle, a two-triangle light and a SPHERE light (shape.rs:91-145) — the commented-out
    variants of world.rs:183-200 made live.  Object order = object id."""
    bs, bd, ls = 1.0, -2.0, 0.3
    tri = lambda a, b, c, m: Object.new(TriangleShape.new(V(*a), V(*b), V(*c)), m)
    red = LambertianCosineWeighted.new(V(0.8, 0.1, 0.1)); green = LambertianCosineWeighted.new(V(0.1, 0.8, 0.1))
    back = OrenNayar.new(V(0.3, 0.3, 0.8), 0.2); floor = OrenNayar.new(V(0.7, 0.7, 0.6), 0.8)
    ceil = LambertianCosineWeighted.new(V(0.8, 0.8, 0.8)); light = Emissive.new(V(12.0, 12.0, 12.0))
    zb, zf, ly = bd - bs, bd + bs, bs - 0.01
    return [
        tri((-bs, -bs, zb), (-bs, bs, zb), (-bs, bs, zf), red),        # 0 left wall
        tri((-bs, -bs, zb), (-bs, bs, zf), (-bs, -bs, zf), red),       # 1
        tri((bs, -bs, zb), (bs, bs, zf), (bs, bs, zb), green),         # 2 right wall
        tri((bs, -bs, zb), (bs, -bs, zf), (bs, bs, zf), green),        # 3
        tri((-bs, -bs, zb), (bs, -bs, zb), (bs, bs, zb), back),        # 4 back wall, Oren-Nayar 0.2
        tri((-bs, -bs, zb), (bs, bs, zb), (-bs, bs, zb), back),        # 5
        tri((-bs, -bs, zb), (bs, -bs, zf), (bs, -bs, zb), floor),      # 6 floor, Oren-Nayar 0.8
        tri((-bs, -bs, zb), (-bs, -bs, zf), (bs, -bs, zf), floor),     # 7
        tri((-bs, bs, zb), (bs, bs, zb), (bs, bs, zf), ceil),          # 8 ceiling
        tri((-bs, bs, zb), (bs, bs, zf), (-bs, bs, zf), ceil),         # 9
        tri((0.1, ly, bd - ls), (0.7, ly, bd - ls), (0.7, ly, bd + ls), light),     # 10 triangle light
        tri((0.1, ly, bd - ls), (0.7, ly, bd + ls), (0.1, ly, bd + ls), light),     # 11
        Object.new(SphereShape.new(V(-0.55, 0.35, -2.4), 0.15), Emissive.new(V(20.0, 16.0, 10.0))),                                  # 12 sphere light
        Object.new(SphereShape.new(V(0.45, -0.6, -1.9), 0.4), Mirror(roughness=0.25, color=V(0.95, 0.64, 0.54), metallic=1.0, ior=1.5)),   # 13 copper
        Object.new(SphereShape.new(V(-0.45, -0.65, -1.7), 0.35), Mirror(roughness=0.15, color=V(1.0, 1.0, 1.0), metallic=0.0, ior=1.5)),   # 14 glass
        tri((-0.95, -0.2, -2.9), (-0.2, -0.2, -2.95), (-0.6, 0.7, -2.6), Mirror(roughness=0.5, color=V(0.9, 0.9, 0.3), metallic=1.0, ior=1.5)),  # 15 brass panel
    ]


def material_zoo_scene(width=128, height=128):
    shapes, mats = describe_objects(material_zoo_objects())
    return shapes, mats, default_camera(width, height)


# ----------------------------------------------------------------------------------------- SplitMix64
class SplitMix64:
    def __init__(self, seed): self.s = seed & 0xFFFFFFFFFFFFFFFF
    def next_u64(self):
        self.s = (self.s + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return z ^ (z >> 31)
    def uniform(self): return (self.next_u64() >> 11) * (1.0 / 9007199254740992.0)


def splitmix64_array(seed, n):
    """n uniforms in [0,1) — vectorised twin of SplitMix64.uniform()."""
    with np.errstate(over="ignore"):
        s = np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * np.arange(1, n + 1, dtype=np.uint64)
        z = (s ^ (s >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def random_spheres_objects(seed=0x5EED0003):
    """Config C3: 484 small spheres + 3 big + ground quad + emissive quad + 4 emissive spheres."""
    r = SplitMix64(seed)
    objs = []
    g = LambertianCosineWeighted.new(V(0.5, 0.5, 0.5))
    objs.append(Object.new(TriangleShape.new(V(-12, 0, -12), V(-12, 0, 12), V(12, 0, 12)), g))
    objs.append(Object.new(TriangleShape.new(V(-12, 0, -12), V(12, 0, 12), V(12, 0, -12)), g))
    for i in range(-11, 11):
        for j in range(-11, 11):
            u1, u2, u3 = r.uniform(), r.uniform(), r.uniform()
            c = V(i + 0.9 * u1, 0.2, j + 0.9 * u2)
            if u3 < 0.80:
                m = LambertianCosineWeighted.new(V(r.uniform() * r.uniform(), r.uniform() * r.uniform(), r.uniform() * r.uniform()))
            elif u3 < 0.95:
                m = Mirror(roughness=0.05 + 0.45 * r.uniform(), color=V(0.5 + 0.5 * r.uniform(), 0.5 + 0.5 * r.uniform(), 0.5 + 0.5 * r.uniform()),
                           metallic=1.0, ior=1.5)
            else:
                m = Mirror(roughness=0.05 + 0.25 * r.uniform(), color=V(1, 1, 1), metallic=0.0, ior=1.5)
            objs.append(Object.new(SphereShape.new(c, 0.2), m))
    objs.append(Object.new(SphereShape.new(V(0, 1, 0), 1.0), Mirror(roughness=0.05, color=V(1, 1, 1), metallic=0.0, ior=1.5)))
    objs.append(Object.new(SphereShape.new(V(-4, 1, 0), 1.0), LambertianCosineWeighted.new(V(0.4, 0.2, 0.1))))
    objs.append(Object.new(SphereShape.new(V(4, 1, 0), 1.0), Mirror(roughness=0.1, color=V(0.7, 0.6, 0.5), metallic=1.0, ior=1.5)))
    e = Emissive.new(V(15, 15, 15))
    objs.append(Object.new(TriangleShape.new(V(-3, 8, -3), V(3, 8, -3), V(3, 8, 3)), e))
    objs.append(Object.new(TriangleShape.new(V(-3, 8, -3), V(3, 8, 3), V(-3, 8, 3)), e))
    es = Emissive.new(V(8, 8, 8))
    for c in ((-6, 2.5, -6), (6, 2.5, -6), (-6, 2.5, 6), (6, 2.5, 6)):
        objs.append(Object.new(SphereShape.new(V(*c), 0.3), es))
    return objs


def random_spheres_scene(width=3840, height=2160, seed=0x5EED0003):
    shapes, mats = describe_objects(random_spheres_objects(seed))
    return shapes, mats, Camera.look_at(V(13, 2, 3), V(0, 0, 0), V(0, 1, 0), width, height, 20.0)


def heightfield_scene(width=1920, height=1080, cells=708, seed=0x5EED0004):
    """Config C4: cells x cells x 2 Lambertian triangles over +-10 in x,z + an emissive 6x6 quad at y=8.

    Vectorised: writes the same pt_shape / pt_material rows that Object.new(TriangleShape.new(..),
    LambertianCosineWeighted.new(..)) + describe_objects() would (tests/test_scenes.py checks a small grid).
    """
    n = cells + 1
    xs = np.linspace(-10.0, 10.0, n); zs = np.linspace(-10.0, 10.0, n)
    X, Z = np.meshgrid(xs, zs, indexing="ij")
    noise = splitmix64_array(seed, n * n).reshape(n, n)
    Y = 0.5 * np.sin(1.3 * X) * np.cos(0.9 * Z) + 0.1 * noise
    P = np.stack([X, Y, Z], axis=-1)                      # [n, n, 3] f64
    p00, p10, p01, p11 = P[:-1, :-1], P[1:, :-1], P[:-1, 1:], P[1:, 1:]
    t1 = np.concatenate([p00, p01, p11], axis=-1).reshape(-1, 9)   # winding gives +y normals
    t2 = np.concatenate([p00, p11, p10], axis=-1).reshape(-1, 9)
    tris = np.empty((2 * cells * cells, 9), np.float64); tris[0::2] = t1; tris[1::2] = t2
    ntri = tris.shape[0]
    shapes = np.zeros(ntri + 2, SHAPE_DTYPE); mats = np.zeros(ntri + 2, MATERIAL_DTYPE)
    shapes["kind"][:] = PT_SHAPE_TRIANGLE
    shapes["v"][:ntri] = tris.astype(np.float32)
    mats["kind"][:ntri] = PT_MAT_LAMBERTIAN
    mats["p"][:ntri, :3] = np.float32(0.6)
    quad = np.array([[-3, 8, -3, 3, 8, -3, 3, 8, 3], [-3, 8, -3, 3, 8, 3, -3, 8, 3]], np.float32)
    shapes["v"][ntri:] = quad
    mats["kind"][ntri:] = PT_MAT_EMISSIVE
    mats["p"][ntri:, :3] = np.float32(15.0)
    cam = Camera.look_at(V(0, 9, 16), V(0, 0, 0), V(0, 1, 0), width, height, 40.0)
    return shapes, mats, cam
