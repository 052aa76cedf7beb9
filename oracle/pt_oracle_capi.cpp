// pt_oracle_capi.cpp — extern "C" face of the CPU oracle for ctypes (tests/, smoke(), bench.py's
// cpu_baseline and --impl reference legs ONLY).  ORACLE / test infrastructure, not product.
// Build: make -C oracle   ->  oracle/_build/libpt_oracle.so   (g++ -O2 -ffp-contract=off -fopenmp)
#include <omp.h>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>

#include "pt_lbvh_ref.hpp"
#include "pt_oracle.hpp"

using namespace pto;

namespace {
template <class T, class SD, class MD>
void trace_primary_t(const CameraDesc& cam, const SD* sh, const MD* mt, uint32_t n, const RenderParams& p,
                     int jitter, float ox, float oy, uint32_t sample, int32_t* obj_id, double* t_out, float* dir_out) {
    World<T> w(cam, sh, mt, n, p);
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t y = 0; y < (int64_t)cam.height; ++y)
        for (uint32_t x = 0; x < cam.width; ++x) {
            size_t i = (size_t)y * cam.width + x;
            Ray<T> ray;
            if (jitter) {
                ray = w.primary_ray(x, (uint32_t)y, w.key(x, (uint32_t)y, sample));
            } else {
                ray = w.camera.get_ray_with_offset(x, cam.height - 1 - (uint32_t)y, T((double)ox), T((double)oy));
            }
            if (dir_out) { dir_out[3 * i] = (float)to_double(ray.direction.x); dir_out[3 * i + 1] = (float)to_double(ray.direction.y); dir_out[3 * i + 2] = (float)to_double(ray.direction.z); }
            HitRecord<T> h; uint32_t o = 0;
            if (w.hit_scene(ray, w.t_min, World<T>::inf(), h, o)) { obj_id[i] = (int32_t)o; t_out[i] = to_double(h.t); }
            else { obj_id[i] = -1; t_out[i] = std::numeric_limits<double>::infinity(); }
        }
}

template <class T, class SD, class MD>
void render_t(const CameraDesc& cam, const SD* sh, const MD* mt, uint32_t n, const RenderParams& p,
              int recursive, int nthreads, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, double* sum, double* sumsq,
              uint64_t* counters) {
    World<T> w(cam, sh, mt, n, p);
    uint32_t W = cam.width;
    uint64_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, plen = 0;
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    int64_t npix = (int64_t)(x1 - x0) * (y1 - y0);
    // one task per pixel, dynamic schedule: the analogue of rayon's par_iter at main.rs:48
#pragma omp parallel for schedule(dynamic, 8) num_threads(nthreads) reduction(+ : c0, c1, c2, c3, plen)
    for (int64_t k = 0; k < npix; ++k) {
        uint32_t x = x0 + (uint32_t)(k % (x1 - x0)), y = y0 + (uint32_t)(k / (x1 - x0));
        RayCounters rc;
        double s[3] = {0, 0, 0}, q[3] = {0, 0, 0};
        for (uint32_t sp = 0; sp < p.spp; ++sp) {   // world.rs:296
            uint32_t pl = 0;
            Vec3<T> c = w.sample_radiance(x, y, p.sample_offset + sp, recursive != 0, &rc, &pl);
            double r = to_double(c.x), g = to_double(c.y), b = to_double(c.z);
            s[0] += r; s[1] += g; s[2] += b;
            q[0] += r * r; q[1] += g * g; q[2] += b * b;
            plen += pl;
        }
        size_t i = (size_t)y * W + x;
        for (int ch = 0; ch < 3; ++ch) { sum[3 * i + ch] = s[ch]; if (sumsq) sumsq[3 * i + ch] = q[ch]; }
        c0 += rc.primary; c1 += rc.shadow; c2 += rc.scattered; c3 += rc.retrace;
    }
    if (counters) { counters[0] = c0; counters[1] = c1; counters[2] = c2; counters[3] = c3; counters[4] = plen; }
}
}  // namespace

extern "C" {

// rounds: 10 (the published known answers) or 7 (what the render draws from)
void pto_philox4x32(int rounds, const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    U4 r = rounds == 10 ? philox4x32<10>(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]) : philox4x32<7>(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    std::memcpy(out, r.w, 16);
}
void pto_u01(uint32_t x, float* f, double* d) { *f = u01<float>(x); *d = u01<double>(x); }

// Vector3 known-answer runner (math.rs:250-345).  op: 0 add 1 sub 2 mul_scalar 3 scalar_mul 4 mul_vec
// 5 div_scalar 6 neg 7 dot(->out[0]) 8 cross 9 length(->out[0]) 10 normalize 11 luminance(->out[0]) 12 div_vec
void pto_vec3_op(int op, const double a[3], const double b[3], double s, double out[3]) {
    Vec3<double> A(a[0], a[1], a[2]), B(b[0], b[1], b[2]), R;
    switch (op) {
        case 0: R = A + B; break;
        case 1: R = A - B; break;
        case 2: R = A * s; break;
        case 3: R = s * A; break;
        case 4: R = A * B; break;
        case 5: R = A / s; break;
        case 6: R = -A; break;
        case 7: R = Vec3<double>(A.dot(B), 0, 0); break;
        case 8: R = A.cross(B); break;
        case 9: R = Vec3<double>(A.length(), 0, 0); break;
        case 10: R = A.normalize(); break;
        case 11: R = Vec3<double>(A.luminance(), 0, 0); break;
        case 12: R = A / B; break;
    }
    out[0] = R.x; out[1] = R.y; out[2] = R.z;
}

void pto_camera_new(const double origin[3], uint32_t w, uint32_t h, double screen_distance, double fov_deg, CameraDesc* out) {
    *out = camera_new(origin, w, h, screen_distance, fov_deg);
}
void pto_camera_look_at(const double origin[3], const double target[3], const double up[3], uint32_t w, uint32_t h,
                        double fov_deg, CameraDesc* out) {
    *out = camera_look_at(origin, target, up, w, h, fov_deg);
}

// precision: 0 = f64 (reference numerics), 1 = f32 (device-matching).  jitter != 0: offsets from the
// RNG camera block of `sample`; else the fixed (ox, oy).  dir_out may be null.
void pto_trace_primary(int precision, const CameraDesc* cam, const ShapeDesc* sh, const MaterialDesc* mt, uint32_t n,
                       const RenderParams* p, int jitter, float ox, float oy, uint32_t sample, int32_t* obj_id, double* t,
                       float* dir_out) {
    if (precision == 0) trace_primary_t<double>(*cam, sh, mt, n, *p, jitter, ox, oy, sample, obj_id, t, dir_out);
    else trace_primary_t<float>(*cam, sh, mt, n, *p, jitter, ox, oy, sample, obj_id, t, dir_out);
}

// Renders the window [x0,x1) x [y0,y1) of the frame; sum / sumsq are full-frame W*H*3 doubles holding
// the per-pixel SUM over spp of sample radiance and of its square (untouched outside the window).
// counters[5] = {primary, shadow, scattered, retrace (recursive only), total path vertices}.
// Returns wall seconds.
double pto_render(int precision, int recursive, const CameraDesc* cam, const ShapeDesc* sh, const MaterialDesc* mt, uint32_t n,
                  const RenderParams* p, int nthreads, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, double* sum,
                  double* sumsq, uint64_t* counters) {
    auto t0 = std::chrono::steady_clock::now();
    if (precision == 0) render_t<double>(*cam, sh, mt, n, *p, recursive, nthreads, x0, y0, x1, y1, sum, sumsq, counters);
    else render_t<float>(*cam, sh, mt, n, *p, recursive, nthreads, x0, y0, x1, y1, sum, sumsq, counters);
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// One sample of one pixel in both control-flow forms (consistency test).
void pto_sample(int precision, int recursive, const CameraDesc* cam, const ShapeDesc* sh, const MaterialDesc* mt, uint32_t n,
                const RenderParams* p, uint32_t x, uint32_t y, uint32_t sample, double rgb[3]) {
    if (precision == 0) {
        World<double> w(*cam, sh, mt, n, *p);
        Vec3<double> c = w.sample_radiance(x, y, sample, recursive != 0, nullptr);
        rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
    } else {
        World<float> w(*cam, sh, mt, n, *p);
        Vec3<float> c = w.sample_radiance(x, y, sample, recursive != 0, nullptr);
        rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
    }
}

// The same three entry points on the reference's own f64 scene data (ShapeDescD / MaterialDescD): what "the reference's
// numerics" means literally — world.rs:80-211 builds the scene from f64 literals such as 0.99 and 0.3 that f32 cannot hold.
void pto_trace_primary_d(int precision, const CameraDesc* cam, const ShapeDescD* sh, const MaterialDescD* mt, uint32_t n,
                         const RenderParams* p, int jitter, float ox, float oy, uint32_t sample, int32_t* obj_id, double* t, float* dir_out) {
    if (precision == 0) trace_primary_t<double>(*cam, sh, mt, n, *p, jitter, ox, oy, sample, obj_id, t, dir_out);
    else trace_primary_t<float>(*cam, sh, mt, n, *p, jitter, ox, oy, sample, obj_id, t, dir_out);
}
double pto_render_d(int precision, int recursive, const CameraDesc* cam, const ShapeDescD* sh, const MaterialDescD* mt, uint32_t n,
                    const RenderParams* p, int nthreads, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, double* sum, double* sumsq,
                    uint64_t* counters) {
    auto t0 = std::chrono::steady_clock::now();
    if (precision == 0) render_t<double>(*cam, sh, mt, n, *p, recursive, nthreads, x0, y0, x1, y1, sum, sumsq, counters);
    else render_t<float>(*cam, sh, mt, n, *p, recursive, nthreads, x0, y0, x1, y1, sum, sumsq, counters);
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
void pto_sample_d(int precision, int recursive, const CameraDesc* cam, const ShapeDescD* sh, const MaterialDescD* mt, uint32_t n,
                  const RenderParams* p, uint32_t x, uint32_t y, uint32_t sample, double rgb[3]) {
    if (precision == 0) {
        World<double> w(*cam, sh, mt, n, *p);
        Vec3<double> c = w.sample_radiance(x, y, sample, recursive != 0, nullptr);
        rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
    } else {
        World<float> w(*cam, sh, mt, n, *p);
        Vec3<float> c = w.sample_radiance(x, y, sample, recursive != 0, nullptr);
        rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
    }
}

}  // extern "C"

// ---- unit hooks for tests/test_oracle_pinning.py: one shape / material routine per call, f64 in and out ----------------------
namespace {
template <class T> Vec3<T> vin(const double v[3]) { return Vec3<T>(T(v[0]), T(v[1]), T(v[2])); }
template <class T> void vout(const Vec3<T>& v, double o[3]) { o[0] = to_double(v.x); o[1] = to_double(v.y); o[2] = to_double(v.z); }
template <class T> Ray<T> ray_in(const double ro[3], const double rd[3], double eta) {
    Ray<T> r(vin<T>(ro), vin<T>(rd));   // Ray::new normalises, camera.rs:10-16
    r.eta_ratio = T(eta);
    return r;
}
template <class T>
void bsdf_pdf_t(const MaterialDescD& m, const double rd[3], double eta, const double o[3], const double n[3], double f[3], double* pdf) {
    const double zero[3] = {0, 0, 0};
    BsdfEval<T> e = Material<T>(m).bsdf_pdf(ray_in<T>(zero, rd, eta), vin<T>(o), vin<T>(n));
    vout(e.f, f); *pdf = to_double(e.pdf);
}
template <class T>
void bsdf_sample_t(const MaterialDescD& m, const double rd[3], double eta, const double n[3], double r1, double r2, double r3, double dir[3],
                   double f[3], double* pdf, double* cos_theta) {
    const double zero[3] = {0, 0, 0};
    BsdfSample<T> s = Material<T>(m).bsdf_pdf_sample(ray_in<T>(zero, rd, eta), vin<T>(n), T(r1), T(r2), T(r3));
    vout(s.dir, dir); vout(s.f, f); *pdf = to_double(s.pdf); *cos_theta = to_double(s.cos_theta);
}
template <class T>
int shape_hit_t(const ShapeDescD& sd, const double ro[3], const double rd[3], double t_min, double t_max, double* t, double point[3], double normal[3],
                int* front) {
    HitRecord<T> h;
    if (!Shape<T>(sd).hit(ray_in<T>(ro, rd, 1.0), T(t_min), T(t_max), h)) return 0;
    *t = to_double(h.t); vout(h.point, point); vout(h.normal, normal); *front = h.front_face ? 1 : 0;
    return 1;
}
template <class T>
void shape_sample_t(const ShapeDescD& sd, const double from[3], const double* target, double r1, double r2, double point[3], double normal[3],
                    double* pdf_omega, double ldir[3], double* d) {
    HitRecord<T> hf, ht;
    hf.point = vin<T>(from);
    if (target) ht.point = vin<T>(target);
    SurfaceSample<T> s = Shape<T>(sd).sample_surface_from_point(hf, target ? &ht : nullptr, T(r1), T(r2));
    vout(s.point, point); vout(s.normal, normal); *pdf_omega = to_double(s.pdf_omega); vout(s.light_dir, ldir); *d = to_double(s.d);
}
}  // namespace
extern "C" {
// Material::bsdf_pdf (material.rs:86-91, :139-148, :226-266; mirror.rs:179-198).  rd = incoming ray direction.
void pto_bsdf_pdf(int precision, const MaterialDescD* m, const double rd[3], double eta, const double o[3], const double n[3], double f[3],
                  double* pdf) {
    if (precision == 0) bsdf_pdf_t<double>(*m, rd, eta, o, n, f, pdf); else bsdf_pdf_t<float>(*m, rd, eta, o, n, f, pdf);
}
// Material::bsdf_pdf_sample (material.rs:29-40; mirror.rs:200-305); r1, r2, r3 = the uniforms in the reference's draw order.
void pto_bsdf_pdf_sample(int precision, const MaterialDescD* m, const double rd[3], double eta, const double n[3], double r1, double r2,
                         double r3, double dir[3], double f[3], double* pdf, double* cos_theta) {
    if (precision == 0) bsdf_sample_t<double>(*m, rd, eta, n, r1, r2, r3, dir, f, pdf, cos_theta);
    else bsdf_sample_t<float>(*m, rd, eta, n, r1, r2, r3, dir, f, pdf, cos_theta);
}
// Shape::hit (shape.rs:53-89, :161-198) + HitRecord::new (base.rs:19-34).  Returns 1 on a hit.
int pto_shape_hit(int precision, const ShapeDescD* s, const double ro[3], const double rd[3], double t_min, double t_max, double* t,
                  double point[3], double normal[3], int* front) {
    return precision == 0 ? shape_hit_t<double>(*s, ro, rd, t_min, t_max, t, point, normal, front)
                          : shape_hit_t<float>(*s, ro, rd, t_min, t_max, t, point, normal, front);
}
// Shape::sample_surface_from_point (shape.rs:91-145, :200-242); target == null: sample with (r1, r2); else pdf at that point.
void pto_shape_sample(int precision, const ShapeDescD* s, const double from[3], const double* target_or_null, double r1, double r2,
                      double point[3], double normal[3], double* pdf_omega, double ldir[3], double* d) {
    if (precision == 0) shape_sample_t<double>(*s, from, target_or_null, r1, r2, point, normal, pdf_omega, ldir, d);
    else shape_sample_t<float>(*s, from, target_or_null, r1, r2, point, normal, pdf_omega, ldir, d);
}

// Counting-scalar run: iterative form (no redundant re-trace, Q8) over a pixel window.
// out[0] = flops, out[1] = samples, out[2] = unique rays, out[3] = path vertices,
// out[4..6] = flops in shading / closest-hit scans / shadow scans, out[7..9] = primary / shadow / scattered rays.
void pto_count_flops(const CameraDesc* cam, const ShapeDesc* sh, const MaterialDesc* mt, uint32_t n, const RenderParams* p,
                     uint32_t stride, double out[10]) {
    World<Counted> w(*cam, sh, mt, n, *p);
    RayCounters rc;
    uint64_t samples = 0, plen = 0;
    for (int c = 0; c < 3; ++c) Counted::counters()[c] = 0;
    Counted::cat() = 0;
    for (uint32_t y = 0; y < cam->height; y += stride)
        for (uint32_t x = 0; x < cam->width; x += stride)
            for (uint32_t sp = 0; sp < p->spp; ++sp) {
                uint32_t pl = 0;
                w.sample_radiance(x, y, p->sample_offset + sp, false, &rc, &pl);
                samples++; plen += pl;
            }
    uint64_t* fc = Counted::counters();
    out[0] = (double)(fc[0] + fc[1] + fc[2]); out[1] = (double)samples;
    out[2] = (double)(rc.primary + rc.shadow + rc.scattered); out[3] = (double)plen;
    out[4] = (double)fc[0]; out[5] = (double)fc[1]; out[6] = (double)fc[2];
    out[7] = (double)rc.primary; out[8] = (double)rc.shadow; out[9] = (double)rc.scattered;
}

// world.rs:315-332 on a buffer of per-pixel means (3 doubles per pixel) -> RGBA8
void pto_tonemap(const double* mean, uint64_t npix, uint8_t* rgba) {
    for (uint64_t i = 0; i < npix; ++i) {
        for (int c = 0; c < 3; ++c) rgba[4 * i + c] = to_u8_gamma2(mean[3 * i + c]);
        rgba[4 * i + 3] = 255;
    }
}
}  // extern "C"
// world.rs:344-369, byte-compatible with Rust's "{},{},{:.6},{:.6},{:.6},{:.6}": fixed notation, 6 decimals, exact decimal
// expansion rounded half to even (glibc's %.6f does the same); Rust prints non-finite values as NaN, inf, -inf.
static std::string rust6(double v) {
    if (std::isnan(v)) return "NaN";
    if (std::isinf(v)) return v > 0 ? "inf" : "-inf";
    char buf[400];
    std::snprintf(buf, sizeof buf, "%.6f", v);
    return buf;
}
extern "C" {
int pto_export_luminance(const char* path, const double* mean, uint32_t W, uint32_t H) {
    FILE* f = std::fopen(path, "w");
    if (!f) return 1;
    std::fprintf(f, "x,y,r,g,b,luminance\n");
    for (uint32_t y = 0; y < H; ++y)
        for (uint32_t x = 0; x < W; ++x) {
            const double* v = mean + 3 * ((size_t)y * W + x);
            double lum = 0.2126 * v[0] + 0.7152 * v[1] + 0.0722 * v[2];
            std::fprintf(f, "%u,%u,%s,%s,%s,%s\n", x, y, rust6(v[0]).c_str(), rust6(v[1]).c_str(), rust6(v[2]).c_str(), rust6(lum).c_str());
        }
    std::fclose(f);
    return 0;
}

// CPU twin of the GPU LBVH build.  nodes must hold max(n-1,1)*16 words.
void pto_lbvh_build(const ShapeDesc* sh, uint32_t n, uint32_t* codes_sorted, uint32_t* prim_sorted, uint32_t* nodes,
                    float bounds[6]) {
    LbvhRef r = lbvh_build_ref(sh, n);
    if (n) { std::memcpy(codes_sorted, r.codes_sorted.data(), 4ull * n); std::memcpy(prim_sorted, r.prim_sorted.data(), 4ull * n); }
    std::memcpy(nodes, r.nodes.data(), 4ull * r.nodes.size());
    for (int k = 0; k < 3; ++k) { bounds[k] = r.smin[k]; bounds[3 + k] = r.smax[k]; }
}

int pto_num_threads(void) { return omp_get_max_threads(); }

}  // extern "C"
