// pt_oracle_capi.cpp — extern "C" face of the CPU oracle for ctypes (tests/, smoke(), bench.py's
// cpu_baseline and --impl reference legs ONLY).  ORACLE / test infrastructure, not product.
// Build: make -C oracle   ->  oracle/_build/libpt_oracle.so   (g++ -O2 -ffp-contract=off -fopenmp)
#include <omp.h>

#include <chrono>
#include <cstdio>
#include <cstring>

#include "pt_lbvh_ref.hpp"
#include "pt_oracle.hpp"

using namespace pto;

namespace {
template <class T>
void trace_primary_t(const CameraDesc& cam, const ShapeDesc* sh, const MaterialDesc* mt, uint32_t n, const RenderParams& p,
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

template <class T>
void render_t(const CameraDesc& cam, const ShapeDesc* sh, const MaterialDesc* mt, uint32_t n, const RenderParams& p,
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

void pto_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    U4 r = philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
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
// world.rs:344-369, byte-compatible with Rust's "{},{},{:.6},{:.6},{:.6},{:.6}"
int pto_export_luminance(const char* path, const double* mean, uint32_t W, uint32_t H) {
    FILE* f = std::fopen(path, "w");
    if (!f) return 1;
    std::fprintf(f, "x,y,r,g,b,luminance\n");
    for (uint32_t y = 0; y < H; ++y)
        for (uint32_t x = 0; x < W; ++x) {
            const double* v = mean + 3 * ((size_t)y * W + x);
            double lum = 0.2126 * v[0] + 0.7152 * v[1] + 0.0722 * v[2];
            std::fprintf(f, "%u,%u,%.6f,%.6f,%.6f,%.6f\n", x, y, v[0], v[1], v[2], lum);
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
