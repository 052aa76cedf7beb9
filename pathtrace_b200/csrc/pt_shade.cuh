// pt_shade.cuh — device restatement of the reference's light sampling and material models (FP32).
// Citations are into /root/reference/src.  Statistical parity only (image RMSE / z-test vs the f64
// oracle); see pt_math.cuh for the bit-exact part.
#pragma once
#include "pt_math.cuh"

namespace ptd {

// Material record, 32 B: m0 = (color.rgb, kind bits), m1 = (roughness, metallic | on_a, ior | on_b, -)
#define PT_MATK_LAMBERT 0u
#define PT_MATK_EMISSIVE 1u
#define PT_MATK_MIRROR 2u
#define PT_MATK_ORENNAYAR 3u

struct Mat {
    float3 color;
    uint32_t kind;
    float roughness, metallic, ior;   // Mirror; OrenNayar reuses metallic = a, ior = b
};
PT_DEV Mat load_mat(const float4* __restrict__ mats, int obj) {
    float4 m0 = __ldg(mats + 2 * obj), m1 = __ldg(mats + 2 * obj + 1);
    Mat m;
    m.color = v3(m0); m.kind = __float_as_uint(m0.w); m.roughness = m1.x; m.metallic = m1.y; m.ior = m1.z;
    return m;
}
// material.rs:51-53, mirror.rs:317-319
PT_DEV float mat_eta(const Mat& m) { return m.kind == PT_MATK_MIRROR ? m.ior : 1.0f; }

// Shared tangent frame: material.rs:105-119, mirror.rs:21-27, material.rs:215-222.  up x n written out for the two axes
// `up` can be (the products with its 0 and 1 components are exact, so this is the reference's cross product).
PT_DEV void tangent_frame(float3 n, float3& tg, float3& bt) {
    const bool up_x = fabsf(n.y) > 0.999f;
    tg = normalize(up_x ? v3(0.f, -n.z, n.y) : v3(n.z, 0.f, -n.x));
    bt = cross(n, tg);
}
// sin and cos of 2 pi r for r in (0, 1): the argument is folded to (-pi, pi), where MUFU.SIN / MUFU.COS are accurate to
// 2^-21 absolute — far inside what a sampled direction needs (the library's sin-cos-of-pi-x costs ~30 instructions, this 6).
PT_DEV void sincos_2pi(float r, float& s, float& c) {
    const float phi = (r - 0.5f) * (2.0f * PT_PI);
    s = -__sinf(phi);
    c = -__cosf(phi);
}
// material.rs:93-122 / :268-295 — phi from r1, cos(theta) = sqrt(r2)
PT_DEV float3 cosine_sample(float3 n, float r1, float r2) {
    float sp, cp;
    sincos_2pi(r1, sp, cp);
    float ct = sqrtf(r2);
    float st = sqrtf(fmaxf(1.0f - ct * ct, 0.0f));
    float3 tg, bt;
    tangent_frame(n, tg, bt);
    return normalize(tg * (st * cp) + bt * (st * sp) + n * ct);
}

// ------------------------------------------------------------------------------------ Mirror (GGX), mirror.rs
PT_DEV float3 mirror_f(const Mat& m, float cos_theta) {   // :126-132
    float r = (1.0f - m.ior) / (1.0f + m.ior);
    float f0d = r * r;
    float3 f0 = v3(f0d, f0d, f0d) * (1.0f - m.metallic) + m.color * m.metallic;
    float c = 1.0f - cos_theta;
    float c2 = c * c;
    float c5 = c2 * c2 * c;
    return f0 + (v3(1.f, 1.f, 1.f) - f0) * c5;
}
PT_DEV float mirror_g1(float alpha2, float c) {   // :136-149
    if (c <= 0.0f) return 0.0f;
    return 2.0f * c / (c + sqrtf(alpha2 + (1.0f - alpha2) * c * c));
}
PT_DEV float mirror_g(float alpha2, float ci, float co) {   // :153-175
    if (ci <= 0.0f || co <= 0.0f) return 0.0f;
    float li = (sqrtf(alpha2 + (1.0f - alpha2) * ci * ci) - ci) / (2.0f * ci);
    float lo = (sqrtf(alpha2 + (1.0f - alpha2) * co * co) - co) / (2.0f * co);
    return 1.0f / (1.0f + li + lo);
}
PT_DEV float ggx_d(float alpha2, float n_h) {
    float denom = n_h * n_h * (alpha2 - 1.0f) + 1.0f;
    return alpha2 / (PT_PI * denom * denom);
}
PT_DEV float3 sample_ggx_vndf(const Mat& m, float3 view, float3 n, float r1, float r2) {   // :17-60
    float alpha = m.roughness * m.roughness;
    float3 tg, bt;
    tangent_frame(n, tg, bt);
    float3 vl = v3(dot(view, tg), dot(view, bt), dot(view, n));
    float3 vh = normalize(v3(alpha * vl.x, alpha * vl.y, vl.z));
    float lensq = vh.x * vh.x + vh.y * vh.y;
    float3 t1 = lensq > 0.0f ? v3(-vh.y, vh.x, 0.f) * rsqrtf(lensq) : v3(1.f, 0.f, 0.f);
    float3 t2 = cross(vh, t1);
    float r = sqrtf(r1);
    float sp, cp;
    sincos_2pi(r2, sp, cp);
    float c1 = r * cp, c2 = r * sp;
    float s = 0.5f * (1.0f + vh.z);
    c2 = (1.0f - s) * sqrtf(fmaxf(1.0f - c1 * c1, 0.0f)) + s * c2;
    float3 nh = t1 * c1 + t2 * c2 + vh * sqrtf(fmaxf(1.0f - c1 * c1 - c2 * c2, 0.0f));
    float3 ne = normalize(v3(alpha * nh.x, alpha * nh.y, fmaxf(nh.z, 0.0f)));
    return normalize(tg * ne.x + bt * ne.y + n * ne.z);
}
// Material::bsdf_pdf for Mirror: :179-198 dispatching to brdf :62-88 / btdf :90-124.  rd = incoming ray direction.
PT_DEV void mirror_eval(const Mat& m, float3 rd, float eta, float3 o, float3 n, float3& f, float& pdf) {
    float3 i = -rd;
    float alpha = m.roughness * m.roughness, alpha2 = alpha * alpha;
    bool refl = dot(i, n) * dot(o, n) > 0.0f;
    if (m.metallic > 0.99f && !refl) { f = v3(0.f, 0.f, 0.f); pdf = 1.0f; return; }
    if (refl) {
        float3 h = normalize(i + o);
        float n_h = dot(n, h);
        float d = ggx_d(alpha2, n_h);
        float i_n = fmaxf(dot(n, i), 0.0f), o_n = fmaxf(dot(n, o), 0.0f);
        float g = mirror_g(alpha2, i_n, o_n);
        float i_h = dot(i, h);
        float3 fr = mirror_f(m, fmaxf(i_h, 0.0f));
        f = fr * (d * g / (4.0f * i_n * o_n));
        pdf = d * fabsf(n_h) / (4.0f * fabsf(i_h));
    } else {
        float3 h = -normalize(i * eta + o);
        float n_h = dot(n, h);
        float d = ggx_d(alpha2, n_h);
        float i_n = fabsf(dot(n, i)), o_n = fabsf(dot(n, o));
        float g = mirror_g(alpha2, i_n, o_n);
        float i_h = dot(i, h), o_h = dot(o, h);
        float dt = eta * i_h + o_h;
        float3 fr = mirror_f(m, fabsf(i_h));
        f = (v3(1.f, 1.f, 1.f) - fr) * (d * g * fabsf(i_h) * fabsf(o_h) / (i_n * o_n * dt * dt));
        pdf = d * fabsf(n_h) * (fabsf(o_h) / (dt * dt));
    }
}
// Mirror::bsdf_pdf_sample, :200-305.  Returns false for the reference's (normal, 0, 1, 0) failure tuple.
PT_DEV bool mirror_sample(const Mat& m, float3 rd, float eta, float3 n, float r1, float r2, float r3, float3& dir, float3& f,
                          float& pdf, float& cos_theta) {
    float3 i = -rd;
    float i_dot_n = dot(i, n);
    float3 h = sample_ggx_vndf(m, i, n, r1, r2);
    float i_h = dot(i, h);
    if (i_h <= 0.0f) return false;                       // :214-217, before the lobe draw
    float3 fr = mirror_f(m, i_h);
    float sin2_i = 1.0f - i_h * i_h;
    float cos2_t = 1.0f - eta * eta * sin2_i;
    bool tir = cos2_t < 0.0f;
    float rr_f = fr.x;                                   // lobe choice by the red channel, :226
    if (tir || m.metallic > 0.99f) { rr_f = 1.0f; fr = v3(1.f, 1.f, 1.f); }
    bool is_reflect = r3 < rr_f;
    float alpha = m.roughness * m.roughness, alpha2 = alpha * alpha;
    float d = ggx_d(alpha2, dot(n, h));
    if (is_reflect) {
        dir = normalize(h * (2.0f * i_h) - i);
        float o_n = fmaxf(dot(n, dir), 0.0f), i_n = fmaxf(i_dot_n, 0.0f);
        float g = mirror_g(alpha2, i_n, o_n);
        f = fr * (d * g / (4.0f * i_n * o_n * rr_f));    // pre-divided by the lobe probability, :249
        pdf = mirror_g1(alpha2, i_n) * d * i_h / i_n / (4.0f * i_h);
        cos_theta = o_n;
    } else {
        float cos_t = sqrtf(cos2_t);
        dir = normalize(h * (eta * i_h - cos_t) - i * eta);
        float o_h = dot(dir, h);
        float o_n = fabsf(dot(n, dir)), i_n = fabsf(i_dot_n);
        float dt = eta * i_h + o_h;
        float g = mirror_g(alpha2, i_n, o_n);
        f = (v3(1.f, 1.f, 1.f) - fr) * (d * g * i_h * fabsf(o_h) / (i_n * o_n * dt * dt * (1.0f - rr_f)));   // :281-282
        pdf = mirror_g1(alpha2, i_n) * d * i_h / i_n * (fabsf(o_h) / (dt * dt));
        cos_theta = o_n;
    }
    return finite3(f) && isfinite(pdf) && pdf > 0.0f;    // :258-265, :293-300
}

// ------------------------------------------------------------------------------------ OrenNayar, material.rs:226-266
PT_DEV float3 orennayar_f(const Mat& m, float3 rd, float3 o, float3 n) {
    float3 i = -rd;
    float ci = fmaxf(dot(i, n), 0.0f), co = fmaxf(dot(o, n), 0.0f);
    float si = sqrtf(fmaxf(1.0f - ci * ci, 0.0f)), so = sqrtf(fmaxf(1.0f - co * co, 0.0f));
    float3 tg, bt;
    tangent_frame(n, tg, bt);
    // cos(phi_i - phi_o) of material.rs:244-249 without the two atan2 and the cosine: with (x, y) = the tangent-plane
    // coordinates, cos(phi_i - phi_o) = (xi xo + yi yo) / (|pi| |po|); a zero projection has phi = atan2(0, 0) = 0.
    float xi = dot(i, tg), yi = dot(i, bt), xo = dot(o, tg), yo = dot(o, bt);
    float ri2 = xi * xi + yi * yi, ro2 = xo * xo + yo * yo;
    if (ri2 == 0.0f) { xi = 1.0f; yi = 0.0f; ri2 = 1.0f; }
    if (ro2 == 0.0f) { xo = 1.0f; yo = 0.0f; ro2 = 1.0f; }
    float cpd = fmaxf((xi * xo + yi * yo) * rsqrtf(ri2 * ro2), 0.0f);
    float sin_alpha, tan_beta;
    if (ci > co) { tan_beta = ci > 1e-6f ? si / ci : 0.0f; sin_alpha = so; }
    else         { tan_beta = co > 1e-6f ? so / co : 0.0f; sin_alpha = si; }
    float term = m.metallic + m.ior * cpd * sin_alpha * tan_beta;   // a + b * ...
    return m.color * (term * PT_INV_PI);
}

// ------------------------------------------------------------------------------------ Material dispatch
// Material::bsdf_pdf (material.rs:86-91, :139-148, :226-266; mirror.rs:179-198)
PT_DEV void bsdf_eval(const Mat& m, float3 rd, float eta, float3 o, float3 n, float3& f, float& pdf) {
    switch (m.kind) {
        case PT_MATK_LAMBERT: f = m.color * PT_INV_PI; pdf = fmaxf(dot(o, n), 0.0f) * PT_INV_PI; break;
        case PT_MATK_MIRROR: mirror_eval(m, rd, eta, o, n, f, pdf); break;
        case PT_MATK_ORENNAYAR: f = orennayar_f(m, rd, o, n); pdf = fmaxf(dot(o, n), 0.0f) * PT_INV_PI; break;
        default: f = v3(0.f, 0.f, 0.f); pdf = 1.0f; break;
    }
}
// Material::bsdf_pdf_sample (default material.rs:29-40; Mirror override mirror.rs:200-305)
PT_DEV bool bsdf_sample(const Mat& m, float3 rd, float eta, float3 n, float r1, float r2, float r3, float3& dir, float3& f,
                        float& pdf, float& cos_theta) {
    if (m.kind == PT_MATK_MIRROR) return mirror_sample(m, rd, eta, n, r1, r2, r3, dir, f, pdf, cos_theta);
    if (m.kind == PT_MATK_EMISSIVE) return false;   // f = 0: zero throughput (never reached in MIS, rendering.rs:42-49)
    dir = cosine_sample(n, r1, r2);
    cos_theta = fmaxf(dot(dir, n), 0.0f);
    pdf = cos_theta * PT_INV_PI;
    f = m.kind == PT_MATK_LAMBERT ? m.color * PT_INV_PI : orennayar_f(m, rd, dir, n);
    return pdf > 0.0f;
}

// ------------------------------------------------------------------------------------ light sampling, shape.rs
// TriangleShape::sample_surface_from_point pdf part, :215-240: solid-angle pdf of point p seen from x.
// `area` = |e1 x e2| / 2, stored with the primitive (q3.w) when the scene is packed.
PT_DEV float triangle_pdf_omega(float area, float3 nrm, float3 x, float3 p, float3& ldir, float& d) {
    float3 tl = p - x;
    d = length(tl);
    ldir = tl / d;
    float cos_light = fabsf(dot(nrm, ldir));   // two-sided emitters, :231
    return cos_light > 1e-8f ? (d * d) / (area * cos_light) : 1e-8f;
}
// SphereShape::sample_surface_from_point pdf part, :97-104
PT_DEV float sphere_pdf_omega(float3 c, float radius, float3 x) {
    float3 tc = c - x;
    float sin2 = radius * radius / dot(tc, tc);
    float cos_max = sqrtf(fmaxf(1.0f - sin2, 0.0f));
    return 1.0f / (2.0f * PT_PI * (1.0f - cos_max));
}
// Samples a point on light primitive (q0,q1,q2,q3) as seen from x.  shape.rs:91-145 / :200-242.
PT_DEV void sample_light_shape(float4 q0, float4 q1, float4 q2, float4 q3, float3 x, float r1, float r2, float3& ldir, float& d,
                               float& pdf_omega) {
    if (__float_as_uint(q0.w) == PT_KIND_TRIANGLE) {
        float sq = sqrtf(r1);
        float u = 1.0f - sq, v = r2 * sq;
        float3 p = v3(q0) + v3(q1) * u + v3(q2) * v;
        pdf_omega = triangle_pdf_omega(q3.w, v3(q3), x, p, ldir, d);
    } else {
        float3 c = v3(q0);
        float radius = q1.x;
        float3 tc = c - x;
        float sin2 = radius * radius / dot(tc, tc);
        float cos_max = sqrtf(fmaxf(1.0f - sin2, 0.0f));
        pdf_omega = 1.0f / (2.0f * PT_PI * (1.0f - cos_max));
        float ct = 1.0f - r1 + r1 * cos_max;
        float st = sqrtf(fmaxf(1.0f - ct * ct, 0.0f));
        float sp, cp;
        sincos_2pi(r2, sp, cp);
        float3 w = normalize(tc);
        float3 up = fabsf(w.y) > 0.999f ? v3(1.f, 0.f, 0.f) : v3(0.f, 1.f, 0.f);
        float3 uu = normalize(cross(up, w));
        float3 vv = cross(w, uu);
        float3 dir = uu * (st * cp) + vv * (st * sp) + w * ct;
        float3 dn = normalize(dir);
        float3 oc = x - c;
        float a = dot(dir, dir), hb = dot(oc, dir), cc = dot(oc, oc) - radius * radius;   // un-normalised dir, :132
        float t = (-hb - sqrtf(hb * hb - a * cc)) / a;    // NaN when the cone sample grazes past: contribution dropped below
        float3 p = x + dn * t;
        float3 tl = p - x;
        d = length(tl);
        ldir = normalize(tl);
    }
}

}  // namespace ptd
