"""Checks that pin the oracle WITHOUT sharing its code (CPU suite).

The reference's tests pin nothing on this path except Vector3 and the reference cannot run here, so the oracle
(oracle/pt_oracle.hpp) is one reading of the Rust.  This file is a second, independent reading plus properties no reading can
fake:

  1. NumPy/f64 transcriptions written HERE, line by line from /root/reference/src/objects/{shape,material,mirror}.rs (cited at
     each function), compared with the oracle's routines on seeded random inputs to 1e-12;
  2. every sampler's returned pdf IS the density of the directions it generates (the Jacobian of (r1, r2) -> direction, by
     central differences), and integrates to one over its support (quadrature);
  3. single-scatter energy: E[f cos / pdf] equals the quadrature of f cos over the hemisphere (the white-furnace identity),
     is exactly the albedo for Lambertian, and never exceeds one for the GGX lobes; an integrator-level furnace (uniform
     emissive enclosure, BrdfOnlyStrategy) returns albedo x emission on a Lambertian sphere to rounding;
  4. Helmholtz reciprocity of the reflection lobes (GGX brdf, Oren-Nayar, Lambertian).
"""
import math

import numpy as np
import pytest

PI = math.pi
LAMBERT, EMISSIVE, MIRROR, ORENNAYAR = 0, 1, 2, 3
SPHERE, TRIANGLE = 0, 1


# ====================================================================================== transcription: src/math.rs
def dot(a, b): return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]                                            # math.rs:24-26
def cross(a, b): return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])   # :29-35
def length(a): return math.sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2])                                  # :38-40
def normalize(a):                                                                                          # :48-51
    n = length(a)
    return a / n if n > 0.0 else a
def V(x, y, z): return np.array([x, y, z], np.float64)


# ====================================================================================== transcription: src/objects/shape.rs + base.rs
def hit_record(point, outward, t, rd):                      # base.rs:19-34
    front = dot(rd, outward) < 0.0
    return dict(t=t, point=point, normal=outward if front else -outward, front_face=front)


def sphere_hit(center, radius, ro, rd, t_min, t_max):       # shape.rs:53-89 (rd already normalised by Ray::new)
    oc = ro - center
    a = dot(rd, rd)
    half_b = dot(oc, rd)
    c = dot(oc, oc) - radius * radius
    disc = half_b * half_b - a * c
    if disc < 0.0:
        return None
    sqrtd = math.sqrt(disc)
    root = (-half_b - sqrtd) / a
    if root < t_min or t_max < root:
        root = (-half_b + sqrtd) / a
        if root < t_min or t_max < root:
            return None
    p = ro + rd * root
    return hit_record(p, (p - center) / radius, root, rd)


def triangle_hit(v0, v1, v2, ro, rd, t_min, t_max):         # shape.rs:161-198
    e1, e2 = v1 - v0, v2 - v0
    h = cross(rd, e2)
    a = dot(e1, h)
    if abs(a) < 1e-8:
        return None
    f = 1.0 / a
    s = ro - v0
    u = f * dot(s, h)
    if not (0.0 <= u <= 1.0):
        return None
    q = cross(s, e1)
    v = f * dot(rd, q)
    if v < 0.0 or u + v > 1.0:
        return None
    t = f * dot(e2, q)
    if t < t_min or t > t_max:
        return None
    return hit_record(ro + rd * t, normalize(cross(e1, e2)), t, rd)


def sphere_sample(center, radius, x, r1, r2, target=None):   # shape.rs:91-145
    to_center = center - x
    dist_sq = dot(to_center, to_center)
    sin2 = (radius * radius) / dist_sq
    cos_max = math.sqrt(max(1.0 - sin2, 0.0))
    pdf_omega = 1.0 / (2.0 * PI * (1.0 - cos_max))
    if target is not None:
        sp = target
    else:
        cos_t = 1.0 - r1 + r1 * cos_max
        sin_t = math.sqrt(max(1.0 - cos_t * cos_t, 0.0))
        phi = 2.0 * PI * r2
        w = normalize(to_center)
        up = V(1, 0, 0) if abs(w[1]) > 0.999 else V(0, 1, 0)
        u = normalize(cross(up, w))
        v = cross(w, u)
        direction = u * (sin_t * math.cos(phi)) + v * (sin_t * math.sin(phi)) + w * cos_t
        dn = normalize(direction)                           # Ray::new normalises (camera.rs:10-16)
        oc = x - center
        a = dot(direction, direction)                       # :132 — the UN-normalised direction
        half_b = dot(oc, direction)
        c = dot(oc, oc) - radius * radius
        disc = half_b * half_b - a * c
        t = (-half_b - math.sqrt(disc)) / a
        sp = x + dn * t
    ld = sp - x
    return dict(point=sp, normal=normalize(sp - center), pdf_omega=pdf_omega, light_dir=normalize(ld), d=length(ld))


def triangle_sample(v0, v1, v2, x, r1, r2, target=None):     # shape.rs:200-242
    if target is not None:
        sp = target
    else:
        sq = math.sqrt(r1)
        u, v = 1.0 - sq, r2 * sq
        sp = v0 + (v1 - v0) * u + (v2 - v0) * v
    e1, e2 = v1 - v0, v2 - v0
    n = normalize(cross(e1, e2))
    area = length(cross(e1, e2)) * 0.5
    tl = sp - x
    d = length(tl)
    ld = tl / d
    cos_light = abs(dot(n, -ld))
    pdf_omega = (1.0 / area) * (d * d) / cos_light if cos_light > 1e-8 else 1e-8
    return dict(point=sp, normal=n, pdf_omega=pdf_omega, light_dir=ld, d=d)


# ====================================================================================== transcription: src/objects/material.rs
def frame(n):                                               # material.rs:105-119 (== mirror.rs:21-27, material.rs:215-222)
    up = V(1, 0, 0) if abs(n[1]) > 0.999 else V(0, 1, 0)
    tg = normalize(cross(up, n))
    return tg, cross(n, tg)


def cosine_direction(n, r1, r2):                            # material.rs:93-122 / :268-295
    phi = 2.0 * PI * r1
    ct = math.sqrt(r2)
    st = math.sqrt(1.0 - ct * ct)
    tg, bt = frame(n)
    return normalize(tg * (st * math.cos(phi)) + bt * (st * math.sin(phi)) + n * ct)


def lambert_bsdf_pdf(albedo, o, n):                         # material.rs:80-91
    return albedo / PI, max(dot(o, n), 0.0) / PI


def orennayar_ab(sigma):                                    # material.rs:182-196
    s2 = sigma * sigma
    return 1.0 - 0.5 * s2 / (s2 + 0.33), 0.45 * s2 / (s2 + 0.09)


def orennayar_bsdf_pdf(albedo, a, b, rd, o, n):             # material.rs:226-266
    i = -rd
    ci, co = max(dot(i, n), 0.0), max(dot(o, n), 0.0)
    si, so = math.sqrt(max(1.0 - ci * ci, 0.0)), math.sqrt(max(1.0 - co * co, 0.0))
    tg, _ = frame(n)
    bt = cross(n, tg)
    phi_i = math.atan2(dot(i, bt), dot(i, tg))
    phi_o = math.atan2(dot(o, bt), dot(o, tg))
    cpd = max(math.cos(phi_i - phi_o), 0.0)
    if ci > co:
        tan_beta = si / ci if ci > 1e-6 else 0.0
        sin_alpha = so
    else:
        tan_beta = so / co if co > 1e-6 else 0.0
        sin_alpha = si
    term = a + b * cpd * sin_alpha * tan_beta
    return albedo * (term / PI), co / PI


# ====================================================================================== transcription: src/objects/mirror.rs
class Mirror:
    def __init__(self, roughness, color, metallic, ior):
        self.roughness, self.color, self.metallic, self.ior = roughness, np.asarray(color, np.float64), metallic, ior

    def params(self): return [self.roughness, *self.color, self.metallic, self.ior]

    def get_f(self, c):                                     # :126-132
        f0d = ((1.0 - self.ior) / (1.0 + self.ior)) ** 2
        f0 = V(f0d, f0d, f0d) * (1.0 - self.metallic) + self.color * self.metallic
        return f0 + (V(1, 1, 1) - f0) * (1.0 - c) ** 5

    def get_g1(self, c):                                    # :136-149
        if c <= 0.0:
            return 0.0
        a2 = (self.roughness * self.roughness) ** 2
        return 2.0 * c / (c + math.sqrt(a2 + (1.0 - a2) * c * c))

    def get_g(self, ci, co):                                # :153-175
        if ci <= 0.0 or co <= 0.0:
            return 0.0
        a2 = (self.roughness * self.roughness) ** 2
        lam = lambda c: (math.sqrt(a2 + (1.0 - a2) * c * c) - c) / (2.0 * c)
        return 1.0 / (1.0 + lam(ci) + lam(co))

    def d(self, n_h):
        a2 = (self.roughness * self.roughness) ** 2
        den = (n_h * n_h) * (a2 - 1.0) + 1.0
        return a2 / (PI * den * den)

    def sample_vndf(self, view, n, r1, r2):                 # :17-60
        alpha = self.roughness * self.roughness
        tg, bt = frame(n)
        vl = V(dot(view, tg), dot(view, bt), dot(view, n))
        vh = normalize(V(alpha * vl[0], alpha * vl[1], vl[2]))
        lensq = vh[0] * vh[0] + vh[1] * vh[1]
        t1 = V(-vh[1], vh[0], 0.0) * (1.0 / math.sqrt(lensq)) if lensq > 0.0 else V(1, 0, 0)
        t2 = cross(vh, t1)
        r = math.sqrt(r1)
        phi = 2.0 * PI * r2
        c1, c2 = r * math.cos(phi), r * math.sin(phi)
        s = 0.5 * (1.0 + vh[2])
        c2 = (1.0 - s) * math.sqrt(1.0 - c1 * c1) + s * c2
        nh = t1 * c1 + t2 * c2 + vh * math.sqrt(max(1.0 - c1 * c1 - c2 * c2, 0.0))
        ne = normalize(V(alpha * nh[0], alpha * nh[1], max(nh[2], 0.0)))
        return normalize(tg * ne[0] + bt * ne[1] + n * ne[2])

    def brdf(self, rd, o, n):                               # :62-88
        i = -rd
        h = normalize(i + o)
        d = self.d(dot(n, h))
        i_n, o_n = max(dot(n, i), 0.0), max(dot(n, o), 0.0)
        g = self.get_g(i_n, o_n)
        f = self.get_f(max(dot(i, h), 0.0))
        with np.errstate(all="ignore"):
            brdf = (d * g) * f / np.float64(4.0 * i_n * o_n)
        return brdf, d * abs(dot(n, h)) / (4.0 * abs(dot(i, h)))

    def btdf(self, rd, eta, o, n):                          # :90-124
        i = -rd
        h = -normalize(i * eta + o)
        n_h = dot(n, h)
        d = self.d(n_h)
        i_n, o_n = abs(dot(n, i)), abs(dot(n, o))
        g = self.get_g(i_n, o_n)
        i_h, o_h = dot(i, h), dot(o, h)
        dt = eta * i_h + o_h
        f = self.get_f(abs(i_h))
        btdf = (V(1, 1, 1) - f) * d * g * abs(i_h) * abs(o_h) / (i_n * o_n * dt * dt)
        return btdf, d * abs(n_h) * (abs(o_h) / (dt * dt))

    def bsdf_pdf(self, rd, eta, o, n):                      # :179-198
        i = -rd
        refl = dot(i, n) * dot(o, n) > 0.0
        if self.metallic > 0.99 and not refl:
            return V(0, 0, 0), 1.0
        return self.brdf(rd, o, n) if refl else self.btdf(rd, eta, o, n)

    def bsdf_pdf_sample(self, rd, eta, n, r1, r2, r3):      # :200-305
        fail = (n, V(0, 0, 0), 1.0, 0.0)
        i = -rd
        i_dot_n = dot(i, n)
        h = self.sample_vndf(i, n, r1, r2)
        i_h = dot(i, h)
        if i_h <= 0.0:
            return fail
        fr = self.get_f(i_h)
        cos2_t = 1.0 - (eta * eta) * (1.0 - i_h * i_h)
        rr_f = fr[0]
        if cos2_t < 0.0 or self.metallic > 0.99:
            rr_f, fr = 1.0, V(1, 1, 1)
        d = self.d(dot(n, h))
        with np.errstate(all="ignore"):
            if r3 < rr_f:
                on = normalize(2.0 * i_h * h - i)
                o_n, i_n = max(dot(n, on), 0.0), max(i_dot_n, 0.0)
                g = self.get_g(i_n, o_n)
                f = fr * d * g / np.float64(4.0 * i_n * o_n * rr_f)
                pdf = np.float64(self.get_g1(i_n) * d * max(i_h, 0.0)) / np.float64(i_n) / (4.0 * abs(i_h))
            else:
                on = normalize(h * (eta * i_h - math.sqrt(cos2_t)) - i * eta)
                o_h = dot(on, h)
                o_n, i_n = abs(dot(n, on)), abs(i_dot_n)
                dt = eta * i_h + o_h
                g = self.get_g(i_n, o_n)
                f = (V(1, 1, 1) - fr) * d * g * abs(i_h) * abs(o_h) / np.float64(i_n * o_n * dt * dt * (1.0 - rr_f))
                pdf = np.float64(self.get_g1(i_n) * d * max(i_h, 0.0)) / np.float64(i_n) * (abs(o_h) / (dt * dt))
        if not (np.isfinite(f).all() and np.isfinite(pdf) and pdf > 0.0):
            return fail
        return on, f, float(pdf), o_n


# ====================================================================================== helpers
def rand_unit(rng):
    v = rng.normal(size=3)
    return v / np.linalg.norm(v)


def rand_hemi(rng, n, min_cos=0.05):
    while True:
        v = rand_unit(rng)
        if dot(v, n) < 0:
            v = -v
        if dot(v, n) > min_cos:
            return v


def close(a, b, tol=1e-12):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.all(np.abs(a - b) <= tol * np.maximum(1.0, np.maximum(np.abs(a), np.abs(b))))


# ====================================================================================== 1. known answers from the second reading
def test_shape_hit_matches_independent_transcription(oracle):
    rng = np.random.default_rng(101)
    hits = 0
    for k in range(400):
        ro = rng.uniform(-2, 2, 3)
        rd = rand_unit(rng)
        if k % 2 == 0:
            c, r = rng.uniform(-1, 1, 3) + ro + rd * rng.uniform(0.5, 3), rng.uniform(0.1, 1.5)
            want = sphere_hit(c, r, ro, rd, 0.001, math.inf)
            got = oracle.shape_hit(SPHERE, [*c, r], ro, rd, 0.001, math.inf)
        else:
            ctr = ro + rd * rng.uniform(0.5, 3)
            v = [ctr + rng.uniform(-1, 1, 3) for _ in range(3)]
            want = triangle_hit(v[0], v[1], v[2], ro, rd, 0.001, math.inf)
            got = oracle.shape_hit(TRIANGLE, np.concatenate(v), ro, rd, 0.001, math.inf)
        assert (want is None) == (got is None), k
        if want:
            hits += 1
            assert close(got["t"], want["t"]) and close(got["point"], want["point"]) and close(got["normal"], want["normal"])
            assert got["front_face"] == want["front_face"]
    assert hits > 150


def test_shape_hit_range_and_tie_semantics(oracle):
    """t == t_max is accepted (shape.rs:77 / :190 reject only strictly outside): the root of world.rs:281-287's tie rule (Q9);
    the sphere falls back to its far root when the near one is out of range; inside hits face-forward the normal (base.rs)."""
    tri = np.array([-1, -1, -3, 1, -1, -3, 0, 1, -3], np.float64)
    ro, rd = V(0, 0, 0), V(0, 0, -1)
    h = oracle.shape_hit(TRIANGLE, tri, ro, rd, 0.001, math.inf)
    assert h["t"] == 3.0
    assert oracle.shape_hit(TRIANGLE, tri, ro, rd, 0.001, 3.0) is not None            # inclusive upper bound
    assert oracle.shape_hit(TRIANGLE, tri, ro, rd, 0.001, np.nextafter(3.0, 0)) is None
    assert oracle.shape_hit(TRIANGLE, tri, ro, rd, 3.0, math.inf) is not None            # inclusive lower bound
    sph = [0, 0, 0, 1.0]
    inside = oracle.shape_hit(SPHERE, sph, V(0, 0, 0), V(0, 0, -1), 0.001, math.inf)      # near root negative -> far root
    assert inside["t"] == 1.0 and not inside["front_face"] and close(inside["normal"], [0, 0, 1])
    assert oracle.shape_hit(SPHERE, sph, V(0, 0, 3), V(0, 0, -1), 0.001, 2.0)["t"] == 2.0  # near root == t_max accepted
    assert oracle.shape_hit(SPHERE, sph, V(0, 0, 3), V(0, 0, -1), 0.001, 1.999) is None
    # |a| < 1e-8: ray parallel to the triangle
    assert oracle.shape_hit(TRIANGLE, tri, V(0, 0, 0), V(1, 0, 0), 0.001, math.inf) is None


def test_light_sampling_matches_independent_transcription(oracle):
    rng = np.random.default_rng(202)
    for k in range(200):
        x = rng.uniform(-1, 1, 3)
        r1, r2 = rng.uniform(0, 1, 2)
        if k % 2 == 0:
            c, r = x + rand_unit(rng) * rng.uniform(1.0, 4.0), rng.uniform(0.05, 0.8)
            want = sphere_sample(c, r, x, r1, r2)
            got = oracle.shape_sample(SPHERE, [*c, r], x, r1, r2)
            tgt = c + rand_unit(rng) * r
            want_t = sphere_sample(c, r, x, 0, 0, target=tgt)
            got_t = oracle.shape_sample(SPHERE, [*c, r], x, target=tgt)
        else:
            v = [x + rng.uniform(-2, 2, 3) + V(0, 2, 0) for _ in range(3)]
            want = triangle_sample(v[0], v[1], v[2], x, r1, r2)
            got = oracle.shape_sample(TRIANGLE, np.concatenate(v), x, r1, r2)
            tgt = v[0] + (v[1] - v[0]) * 0.3 + (v[2] - v[0]) * 0.2
            want_t = triangle_sample(v[0], v[1], v[2], x, 0, 0, target=tgt)
            got_t = oracle.shape_sample(TRIANGLE, np.concatenate(v), x, target=tgt)
        for w, g in ((want, got), (want_t, got_t)):
            for key in ("point", "normal", "pdf_omega", "light_dir", "d"):
                assert close(g[key], w[key], 1e-11), (k, key, g[key], w[key])


MIRRORS = [Mirror(0.3, (1, 1, 1), 0.0, 1.5), Mirror(0.25, (0.95, 0.64, 0.54), 1.0, 1.5), Mirror(0.05, (1, 1, 1), 0.0, 1.33),
           Mirror(0.6, (0.9, 0.9, 0.3), 1.0, 1.5), Mirror(0.15, (0.7, 0.8, 0.9), 0.5, 2.0)]


def test_material_eval_matches_independent_transcription(oracle):
    rng = np.random.default_rng(303)
    for k in range(300):
        n = rand_unit(rng)
        i = rand_hemi(rng, n)
        rd = -i
        o = rand_unit(rng)
        albedo = rng.uniform(0.1, 0.9, 3)
        f, pdf = oracle.bsdf_pdf(LAMBERT, albedo, rd, 1.0, o, n)
        wf, wp = lambert_bsdf_pdf(albedo, o, n)
        assert close(f, wf) and close(pdf, wp)
        sigma = rng.uniform(0, 1)
        a, b = orennayar_ab(sigma)
        f, pdf = oracle.bsdf_pdf(ORENNAYAR, [*albedo, sigma, a, b], rd, 1.0, o, n)
        wf, wp = orennayar_bsdf_pdf(albedo, a, b, rd, o, n)
        assert close(f, wf, 1e-11) and close(pdf, wp)
        m = MIRRORS[k % len(MIRRORS)]
        eta = [1.0, 1.0 / m.ior, m.ior][k % 3]
        f, pdf = oracle.bsdf_pdf(MIRROR, m.params(), rd, eta, o, n)
        wf, wp = m.bsdf_pdf(rd, eta, o, n)
        if np.isfinite(wf).all() and np.isfinite(wp):
            assert close(f, wf, 1e-10) and close(pdf, wp, 1e-10), (k, f, wf, pdf, wp)
        else:
            assert not (np.isfinite(f).all() and np.isfinite(pdf))
    f, pdf = oracle.bsdf_pdf(EMISSIVE, [15, 15, 15], V(0, 0, -1), 1.0, V(0, 0, 1), V(0, 0, 1))   # material.rs:139-148
    assert f.tolist() == [0, 0, 0] and pdf == 1.0


def test_material_sampling_matches_independent_transcription(oracle):
    rng = np.random.default_rng(404)
    n_refl = n_refr = n_fail = 0
    for k in range(600):
        n = rand_unit(rng)
        i = rand_hemi(rng, n, 0.02)
        rd = -i
        r1, r2, r3 = rng.uniform(0, 1, 3)
        albedo = rng.uniform(0.1, 0.9, 3)
        d, f, pdf, ct = oracle.bsdf_pdf_sample(LAMBERT, albedo, rd, 1.0, n, r1, r2, r3)
        wd = cosine_direction(n, r1, r2)
        assert close(d, wd) and close(ct, max(dot(wd, n), 0.0)) and close(pdf, max(dot(wd, n), 0.0) / PI) and close(f, albedo / PI)
        m = MIRRORS[k % len(MIRRORS)]
        front = k % 2 == 0
        eta = 1.0 / m.ior if front else m.ior
        d, f, pdf, ct = oracle.bsdf_pdf_sample(MIRROR, m.params(), rd, eta, n, r1, r2, r3)
        wd, wf, wp, wc = m.bsdf_pdf_sample(rd, eta, n, r1, r2, r3)
        assert close(d, wd, 1e-10) and close(f, wf, 1e-9) and close(pdf, wp, 1e-9) and close(ct, wc, 1e-10), (k, d, wd, f, wf)
        if wp == 1.0 and wc == 0.0:
            n_fail += 1
        elif dot(wd, n) > 0:
            n_refl += 1
        else:
            n_refr += 1
    assert n_refl > 100 and n_refr > 100      # both lobes and (usually) the failure tuple were exercised
    # the failure tuple (normal, 0, 1, 0) of mirror.rs:258-265: the reflection of a rough metal falls below the surface
    # (o_n = 0 -> g = 0 and denom_brdf = 0 -> brdf = 0/0, not finite)
    m = Mirror(0.9, (1, 1, 1), 1.0, 1.5)
    n = V(0, 0, 1)
    i = normalize(V(math.sqrt(1 - 0.25), 0.0, 0.5))
    found = 0
    for r1 in np.linspace(0.01, 0.99, 20):
        for r2 in np.linspace(0.01, 0.99, 20):
            wd, wf, wp, wc = m.bsdf_pdf_sample(-i, 1.0, n, r1, r2, 0.5)
            d, f, pdf, ct = oracle.bsdf_pdf_sample(MIRROR, m.params(), -i, 1.0, n, r1, r2, 0.5)
            assert close(d, wd, 1e-9) and close(pdf, wp, 1e-8)
            if wp == 1.0 and wc == 0.0 and not wf.any():
                found += 1
                assert close(d, n) and not f.any() and pdf == 1.0 and ct == 0.0
    assert found > 40


# ====================================================================================== 2. pdfs are densities and integrate to one
def _jacobian_density(fn, r1, r2, eps=2e-6):
    """1 / |d(direction)/d(r1, r2)| on the unit sphere by central differences: the density of the sampler at (r1, r2)."""
    d1 = (fn(r1 + eps, r2) - fn(r1 - eps, r2)) / (2 * eps)
    d2 = (fn(r1, r2 + eps) - fn(r1, r2 - eps)) / (2 * eps)
    return 1.0 / np.linalg.norm(np.cross(d1, d2))


@pytest.mark.parametrize("case", ["cosine", "oren_nayar", "ggx_reflect_metal", "ggx_reflect_glass", "ggx_refract_enter", "ggx_refract_exit",
                                  "sphere_cone", "triangle"])
def test_returned_pdf_is_the_density_of_the_sampler(case, oracle):
    """Independent of any formula: perturb (r1, r2), watch the generated direction move, and compare the inverse area
    element with the pdf the routine returns (cosine material.rs:93-122; VNDF reflect / refract mirror.rs:200-305; cone
    shape.rs:97-104; triangle area -> solid angle shape.rs:215-240)."""
    import zlib
    rng = np.random.default_rng(zlib.crc32(case.encode()))          # (hash() of a str changes per process)
    n = normalize(V(0.2, 0.9, -0.3))
    checked = 0
    for _ in range(80):
        i = rand_hemi(rng, n, 0.25)
        r1, r2 = rng.uniform(0.08, 0.92, 2)
        if case in ("cosine", "oren_nayar"):
            kind, p = (LAMBERT, [0.5, 0.5, 0.5]) if case == "cosine" else (ORENNAYAR, [0.5, 0.5, 0.5, 0.6, *orennayar_ab(0.6)])
            fn = lambda a, b: oracle.bsdf_pdf_sample(kind, p, -i, 1.0, n, a, b, 0.5)[0]
            pdf = oracle.bsdf_pdf_sample(kind, p, -i, 1.0, n, r1, r2, 0.5)[2]
        elif case.startswith("ggx"):
            m = Mirror(0.45, (1, 1, 1), 1.0 if case.endswith("metal") else 0.0, 1.5)
            eta = m.ior if case.endswith("exit") else 1.0 / m.ior
            r3 = 0.0 if "reflect" in case else 0.999999       # r3 < F -> reflect; r3 ~ 1 -> refract (unless TIR)
            out = oracle.bsdf_pdf_sample(MIRROR, m.params(), -i, eta, n, r1, r2, r3)
            if out[2] == 1.0 and out[3] == 0.0:
                continue                                       # failure tuple
            want_side = 1 if "reflect" in case else -1
            if np.sign(dot(out[0], n)) != want_side:
                continue                                       # total internal reflection took the other lobe
            fn = lambda a, b: oracle.bsdf_pdf_sample(MIRROR, m.params(), -i, eta, n, a, b, r3)[0]
            pdf = out[2]
        elif case == "sphere_cone":
            x, c, r = V(0, 0, 0), rand_unit(rng) * rng.uniform(1.5, 3.0), rng.uniform(0.2, 0.9)
            fn = lambda a, b: oracle.shape_sample(SPHERE, [*c, r], x, a, b)["light_dir"]
            pdf = oracle.shape_sample(SPHERE, [*c, r], x, r1, r2)["pdf_omega"]
        else:
            x = V(0, 0, 0)
            v = np.concatenate([rand_unit(rng) * 2 + V(0, 3, 0) for _ in range(3)])
            fn = lambda a, b: oracle.shape_sample(TRIANGLE, v, x, a, b)["light_dir"]
            pdf = oracle.shape_sample(TRIANGLE, v, x, r1, r2)["pdf_omega"]
        dens = _jacobian_density(fn, r1, r2)
        assert abs(dens / pdf - 1.0) < 2e-3, (case, dens, pdf)
        checked += 1
    assert checked >= 20


def _sphere_grid(nt=400, npn=800):
    th = (np.arange(nt) + 0.5) * PI / nt
    ph = (np.arange(npn) + 0.5) * 2 * PI / npn
    T, P = np.meshgrid(th, ph, indexing="ij")
    dirs = np.stack([np.sin(T) * np.cos(P), np.sin(T) * np.sin(P), np.cos(T)], -1)
    return dirs, np.sin(T) * (PI / nt) * (2 * PI / npn)


def test_pdfs_integrate_to_one(oracle):
    n = V(0, 0, 1)
    dirs, dw = _sphere_grid(120, 240)
    # cosine-weighted pdf over the sphere (zero below the horizon), material.rs:80-83
    pdf = np.array([[oracle.bsdf_pdf(LAMBERT, [0.5, 0.5, 0.5], V(0, 0, -1), 1.0, o, n)[1] for o in row] for row in dirs[::4, ::4]])
    assert abs((pdf * dw[::4, ::4] * 16).sum() - 1.0) < 2e-3
    # cone pdf x cone solid angle (shape.rs:97-104), for several distances
    for dist, r in ((2.0, 0.5), (1.2, 1.0), (10.0, 0.3)):
        s = oracle.shape_sample(SPHERE, [0, 0, dist, r], V(0, 0, 0), 0.3, 0.6)
        cos_max = math.sqrt(1 - (r / dist) ** 2)
        assert s["pdf_omega"] * 2 * PI * (1 - cos_max) == pytest.approx(1.0, abs=1e-12)
    # triangle: integral of pdf_omega over the solid angle it subtends == integral of pdf_area over the area == 1
    v0, v1, v2 = V(-1, 2, -1), V(1.5, 2.2, -0.5), V(0, 2.5, 1.2)
    x = V(0.1, 0, 0.2)
    total, N = 0.0, 200
    area = length(cross(v1 - v0, v2 - v0)) * 0.5
    nrm = normalize(cross(v1 - v0, v2 - v0))
    for a in range(N):
        for b in range(N - a):
            u, v = (a + 1 / 3) / N, (b + 1 / 3) / N        # centroid of the lower sub-triangle; upper ones mirrored below
            for uu, vv in ((u, v), ((a + 2 / 3) / N, (b + 2 / 3) / N)):
                if uu + vv >= 1:
                    continue
                p = v0 + (v1 - v0) * uu + (v2 - v0) * vv
                s = oracle.shape_sample(TRIANGLE, np.concatenate([v0, v1, v2]), x, target=p)
                d_omega = (area / N ** 2) * abs(dot(nrm, s["light_dir"])) / s["d"] ** 2
                total += s["pdf_omega"] * d_omega
    assert total == pytest.approx(1.0, abs=2e-3)
    # VNDF reflection pdf over ALL outgoing directions (mirror.rs:249-255 evaluated through bsdf_pdf's brdf branch gives the
    # half-vector pdf D|n.h|/(4|i.h|), not the VNDF one; the sampler's own pdf is checked by quadrature of the transcription)
    m = Mirror(0.5, (1, 1, 1), 1.0, 1.5)
    i = normalize(V(0.6, 0.0, 0.8))
    D2, dw2 = _sphere_grid(300, 600)
    H = D2 + i
    H /= np.linalg.norm(H, axis=-1, keepdims=True)
    i_h = H @ i
    n_h = H[..., 2]
    a2 = (m.roughness ** 2) ** 2
    Dv = a2 / (PI * ((n_h * n_h) * (a2 - 1) + 1) ** 2)
    pdf_o = np.where((i_h > 0) & (n_h > 0), m.get_g1(i[2]) * Dv * np.maximum(i_h, 0) / i[2] / (4 * np.abs(i_h)), 0.0)
    assert abs((pdf_o * dw2).sum() - 1.0) < 5e-3


# ====================================================================================== 3. energy (white furnace)
def _strat(nn):
    g = (np.arange(nn) + 0.5) / nn
    return [(a, b) for a in g for b in g]


def test_single_scatter_energy_identity(oracle):
    """E[f cos / pdf] over the sampler == integral of f cos over the hemisphere (quadrature of bsdf_pdf): the estimator the
    integrator multiplies into the throughput (rendering.rs:89) is unbiased for each reflection model."""
    n = V(0, 0, 1)
    dirs, dw = _sphere_grid(90, 180)
    for kind, p, name in ((LAMBERT, [0.8, 0.5, 0.2], "lambert"), (ORENNAYAR, [0.8, 0.5, 0.2, 0.8, *orennayar_ab(0.8)], "oren_nayar"),
                          (MIRROR, Mirror(0.5, (1.0, 1.0, 1.0), 1.0, 1.5).params(), "ggx_metal")):
        for i in (normalize(V(0.3, 0.1, 0.95)), normalize(V(0.8, -0.2, 0.4))):
            est = np.zeros(3)
            pts = _strat(48)
            for r1, r2 in pts:
                d, f, pdf, ct = oracle.bsdf_pdf_sample(kind, p, -i, 1.0, n, r1, r2, 0.0)
                est += f * ct / pdf
            est /= len(pts)
            quad = np.zeros(3)
            for row, wrow in zip(dirs, dw):
                for o, w in zip(row, wrow):
                    if o[2] <= 0:
                        continue
                    f, _ = oracle.bsdf_pdf(kind, p, -i, 1.0, o, n)
                    quad += f * o[2] * w
            assert np.allclose(est, quad, rtol=2e-2), (name, est, quad)
            if name == "lambert":
                assert np.allclose(est, p[:3], rtol=1e-12)                 # exactly the albedo: zero-variance estimator
            assert (est <= 1.0 + 1e-9).all()                                # no lobe creates energy


def test_metal_sampling_is_colour_blind_quirk(oracle):
    """Reference quirk (Q10, mirror.rs:226-230): with metallic > 0.99 bsdf_pdf_sample replaces the Fresnel VECTOR by (1, 1, 1),
    so the BSDF-sampled weight of a coloured metal is white, while bsdf_pdf (the NEE side, mirror.rs:62-88) keeps the coloured
    F0.  Both readings agree on it; the GPU reproduces it (pt_shade.cuh: mirror_sample)."""
    n = V(0, 0, 1)
    m = Mirror(0.5, (0.9, 0.6, 0.3), 1.0, 1.5)
    i = normalize(V(0.3, 0.1, 0.95))
    d, f, pdf, ct = oracle.bsdf_pdf_sample(MIRROR, m.params(), -i, 1.0, n, 0.37, 0.61, 0.5)
    assert pdf != 1.0 and f[0] == f[1] == f[2]                                # sampled weight: colour-blind
    fe, _ = oracle.bsdf_pdf(MIRROR, m.params(), -i, 1.0, d, n)
    assert fe[0] > fe[1] > fe[2] and fe[0] < f[0]                             # evaluated BRDF: coloured, and below the white one
    wd, wf, wp, wc = m.bsdf_pdf_sample(-i, 1.0, n, 0.37, 0.61, 0.5)
    assert close(f, wf, 1e-11)


def test_dielectric_lobes_conserve_energy(oracle):
    """Rough glass with white colour absorbs nothing: reflected + transmitted single-scatter energy <= 1 and close to 1 at
    moderate roughness (the deficit is the energy of multiple scattering GGX drops).  f is pre-divided by the lobe
    probability (mirror.rs:249, :281), so the plain mean over a stratified r3 IS the sum over lobes."""
    n = V(0, 0, 1)
    m = Mirror(0.3, (1, 1, 1), 0.0, 1.5)
    for i, eta in ((normalize(V(0.2, 0.1, 0.97)), 1 / 1.5), (normalize(V(0.6, 0.0, 0.8)), 1 / 1.5), (normalize(V(0.2, 0.1, 0.97)), 1.5)):
        tot, cnt = 0.0, 0
        for r1, r2 in _strat(40):
            for r3 in (0.05, 0.15, 0.25, 0.35, 0.45, 0.55, 0.65, 0.75, 0.85, 0.95):
                d, f, pdf, ct = oracle.bsdf_pdf_sample(MIRROR, m.params(), -i, eta, n, r1, r2, r3)
                tot += f[0] * ct / pdf
                cnt += 1
        e = tot / cnt
        assert 0.85 < e <= 1.0 + 5e-3, (i, eta, e)


def test_integrator_white_furnace(oracle):
    """Uniform emissive enclosure (a sphere light around everything), BrdfOnlyStrategy (rendering.rs:213-265): a convex
    Lambertian sphere of albedo rho shows exactly rho x Le on every pixel and every sample — the throughput bookkeeping of
    rendering.rs:89 + :131-133 conserves energy; an Oren-Nayar sphere shows at most Le."""
    from pathtrace_b200.objects import Camera, Emissive, LambertianCosineWeighted, Object, OrenNayar, SphereShape, Vector3
    from conftest import oracle_camera
    Le, rho = 2.0, np.array([0.75, 0.5, 0.25])
    cam = Camera.new(Vector3.new(0, 0, 2), 32, 32, 1.0, 35.0)
    oc = oracle_camera(oracle, cam)
    for mat, exact in ((LambertianCosineWeighted.new(tuple(rho)), True), (OrenNayar.new(tuple(rho), 0.7), False)):
        objs = [Object.new(SphereShape.new((0, 0, 0), 50.0), Emissive.new((Le, Le, Le))), Object.new(SphereShape.new((0, 0, -2.5), 0.7), mat)]
        sh, mt = oracle.describe_f64(objs)
        ids, _, _ = oracle.trace_primary(oc, sh, mt, precision="f64")
        r = oracle.render(oc, sh, mt, oracle.RenderParams(spp=16, strategy=1), "f64", recursive=True)
        mean = r["sum"] / 16
        def interior(mask):                                                   # jitter moves silhouette pixels across the edge
            m = np.pad(mask, 1, constant_values=False)
            return np.logical_and.reduce([m[1 + dy:m.shape[0] - 1 + dy, 1 + dx:m.shape[1] - 1 + dx] for dy in (-1, 0, 1) for dx in (-1, 0, 1)])
        on_sphere, off_sphere = interior(ids == 1), interior(ids == 0)
        assert on_sphere.sum() > 100 and off_sphere.sum() > 300
        assert np.allclose(mean[off_sphere], Le, rtol=1e-14)                   # primary ray sees the enclosure
        if exact:
            assert np.allclose(mean[on_sphere], rho * Le, rtol=1e-12)          # every sample: f cos / pdf = rho, then Le
            var = r["sumsq"] / 16 - mean ** 2
            assert np.abs(var[on_sphere]).max() < 1e-12
        else:
            assert (mean[on_sphere] <= Le * rho.max() * 1.05).all() and mean[on_sphere].mean() > 0.5 * Le * rho.mean()


# ====================================================================================== 4. reciprocity
def test_reflection_lobes_are_reciprocal(oracle):
    """f(i -> o) == f(o -> i) for the reflection models (Lambertian; Oren-Nayar material.rs:226-266; GGX brdf mirror.rs:62-88 with
    its symmetric D, G2 and F(i.h))."""
    rng = np.random.default_rng(505)
    n = normalize(V(-0.3, 0.8, 0.5))
    for k in range(40):
        i, o = rand_hemi(rng, n, 0.1), rand_hemi(rng, n, 0.1)
        for kind, p in ((LAMBERT, [0.6, 0.5, 0.4]), (ORENNAYAR, [0.6, 0.5, 0.4, 0.9, *orennayar_ab(0.9)]),
                        (MIRROR, Mirror(0.35, (0.9, 0.6, 0.3), 1.0, 1.5).params()), (MIRROR, Mirror(0.2, (1, 1, 1), 0.0, 1.5).params())):
            f_io, _ = oracle.bsdf_pdf(kind, p, -i, 1.0, o, n)
            f_oi, _ = oracle.bsdf_pdf(kind, p, -o, 1.0, i, n)
            assert close(f_io, f_oi, 1e-11), (kind, f_io, f_oi)


# ====================================================================================== 5. the oracle's two scene inputs
def test_f64_scene_rows_vs_f32_rows(oracle):
    """The oracle can be fed the scene as world.rs holds it (f64 literals such as 0.99, 0.3) or the f32 rows that cross the C
    ABI; the closest hits differ only by the narrowing (<= 1e-6 relative in t, ids equal away from edges)."""
    from conftest import oracle_camera
    from pathtrace_b200 import scenes
    from pathtrace_b200.objects import describe_objects
    objs = scenes.default_objects()
    s64, m64 = oracle.describe_f64(objs)
    s32, m32 = describe_objects(objs)
    assert s64["v"][10][1] == 0.99 and np.float64(s32["v"][10][1]) != 0.99         # 0.99 is not an f32
    cam = scenes.default_camera(96, 64)
    oc = oracle_camera(oracle, cam)
    ia, ta, _ = oracle.trace_primary(oc, s64, m64, precision="f64")
    ib, tb, _ = oracle.trace_primary(oc, s32, m32, precision="f64")
    same = (ia == ib) & (ia >= 0)
    assert (ia != ib).mean() < 5e-3
    assert (np.abs(ta[same] - tb[same]) / ta[same]).max() < 1e-6
    a = oracle.sample(oc, s64, m64, oracle.RenderParams(spp=1), 40, 30, 3, "f64", True)
    b = oracle.sample(oc, s64, m64, oracle.RenderParams(spp=1), 40, 30, 3, "f64", False)
    assert np.allclose(a, b, rtol=1e-10, atol=1e-12)
