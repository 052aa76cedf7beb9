// pt_oracle.hpp — CPU ORACLE (test infrastructure, NOT product code).
//
// A line-by-line C++ restatement of the hot path of roxas1533/pathtrace (a Rust f64 CPU path
// tracer).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, link or call anything under oracle/.  The CUDA product under
// pathtrace_b200/csrc/ never includes this file.
//
// PARITY STATUS: "parity unpinned".  The reference's own tests (src/math.rs:246-418) pin only Vector3
// arithmetic; those known answers are replayed in tests/test_oracle_kat.py.  Nothing in the
// reference pins intersection, sampling, materials, integrator or image (its world.rs tests have
// no assertions and do not compile), and the reference cannot be built here (no rustc/cargo,
// needs a window).  What guards this restatement instead (all in the CPU suite):
//   tests/test_oracle_pinning.py  independent NumPy f64 transcriptions of shape.rs / material.rs /
//                                 mirror.rs written in the test file from the Rust lines and compared
//                                 with this code to 1e-12; every sampler's pdf integrates to one and
//                                 its sample histogram follows its pdf; white-furnace energy
//                                 conservation; reciprocity of the GGX reflection lobe
//   tests/test_oracle_kat.py      Vector3 + Philox known answers; recursive == iterative
//   tests/test_golden.py          committed outputs (drift guard)
// Third-party arithmetic absent from /root/reference: rand 0.9.2 / rand_chacha 0.9.0 /
// rand_core 0.9.3 (Cargo.lock:1092-1115; StdRng = ChaCha12).  Not restated: the north star
// mandates a counter-based RNG, so image parity is distributional.  The RNG here is
// Philox4x32-7 (Salmon et al., SC'11); its round function is pinned by the published Random123 10-round known answers.
//
// Templated on the scalar type T:
//   double          — the reference's numerics (image oracle, timed CPU baseline)
//   float           — same operation order; dot/cross/at use fused multiply-add exactly as the
//                     device code does, making it the bit-exact id oracle for the GPU
//   Counted         — double that counts one flop per arithmetic op (roofline numerator)
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include <type_traits>
#include <vector>

namespace pto {

// ----------------------------------------------------------------------------------------------
// scalar helpers
// ----------------------------------------------------------------------------------------------
struct Counted {
    double v;
    Counted() : v(0) {}
    Counted(double x) : v(x) {}
    // flop counters by category: 0 = shading / everything else, 1 = closest-hit scans, 2 = shadow scans
    static int& cat() { static thread_local int c = 0; return c; }
    static uint64_t* counters() { static thread_local uint64_t n[3] = {0, 0, 0}; return n; }
    static uint64_t& flops() { return counters()[cat()]; }
};
struct CatScope { int prev; explicit CatScope(int c) : prev(Counted::cat()) { Counted::cat() = c; } ~CatScope() { Counted::cat() = prev; } };
inline Counted operator+(Counted a, Counted b) { Counted::flops()++; return Counted(a.v + b.v); }
inline Counted operator-(Counted a, Counted b) { Counted::flops()++; return Counted(a.v - b.v); }
inline Counted operator*(Counted a, Counted b) { Counted::flops()++; return Counted(a.v * b.v); }
inline Counted operator/(Counted a, Counted b) { Counted::flops()++; return Counted(a.v / b.v); }
inline Counted operator-(Counted a) { return Counted(-a.v); }
inline bool operator<(Counted a, Counted b) { Counted::flops()++; return a.v < b.v; }
inline bool operator>(Counted a, Counted b) { Counted::flops()++; return a.v > b.v; }
inline bool operator<=(Counted a, Counted b) { Counted::flops()++; return a.v <= b.v; }
inline bool operator>=(Counted a, Counted b) { Counted::flops()++; return a.v >= b.v; }
inline Counted& operator+=(Counted& a, Counted b) { a = a + b; return a; }

template <class T> inline double to_double(T x) { return (double)x; }
template <> inline double to_double<Counted>(Counted x) { return x.v; }

#define PTO_UNARY(name, expr)                                                       \
    inline float name(float x) { return expr; }                                     \
    inline double name(double x) { return expr; }                                   \
    inline Counted name(Counted c) { Counted::flops()++; double x = c.v; return Counted(expr); }
PTO_UNARY(m_sqrt, std::sqrt(x))
PTO_UNARY(m_sin, std::sin(x))
PTO_UNARY(m_cos, std::cos(x))
PTO_UNARY(m_abs, std::fabs(x))
#undef PTO_UNARY
inline float m_atan2(float y, float x) { return std::atan2(y, x); }
inline double m_atan2(double y, double x) { return std::atan2(y, x); }
inline Counted m_atan2(Counted y, Counted x) { Counted::flops()++; return Counted(std::atan2(y.v, x.v)); }
// Rust f64::max / f64::min return the non-NaN operand; std::fmax/fmin have the same rule.
inline float m_max(float a, float b) { return std::fmax(a, b); }
inline double m_max(double a, double b) { return std::fmax(a, b); }
inline Counted m_max(Counted a, Counted b) { Counted::flops()++; return Counted(std::fmax(a.v, b.v)); }
inline float m_min(float a, float b) { return std::fmin(a, b); }
inline double m_min(double a, double b) { return std::fmin(a, b); }
inline Counted m_min(Counted a, Counted b) { Counted::flops()++; return Counted(std::fmin(a.v, b.v)); }
inline bool m_finite(float a) { return std::isfinite(a); }
inline bool m_finite(double a) { return std::isfinite(a); }
inline bool m_finite(Counted a) { return std::isfinite(a.v); }

// a*b + c.  The float instantiation fuses (one rounding), matching the device's __fmaf_rn in
// dot / cross / ray.at; double and Counted evaluate exactly as the reference writes it.
template <class T> inline T madd(T a, T b, T c) { return a * b + c; }
template <> inline float madd<float>(float a, float b, float c) { return std::fmaf(a, b, c); }

template <class T> inline T pi() { return T(3.14159265358979323846); }   // std::f64::consts::PI

// ----------------------------------------------------------------------------------------------
// Vector3 — src/math.rs:4-51, :133-135, :138-244 (the subset the render path uses)
// ----------------------------------------------------------------------------------------------
template <class T> struct Vec3 {
    T x, y, z;
    Vec3() : x(T(0)), y(T(0)), z(T(0)) {}
    Vec3(T a, T b, T c) : x(a), y(b), z(c) {}
    static Vec3 zero() { return Vec3(T(0), T(0), T(0)); }   // math.rs:15
    static Vec3 one() { return Vec3(T(1), T(1), T(1)); }    // math.rs:19
    // math.rs:24-26: x*ox + y*oy + z*oz, left to right
    T dot(const Vec3& o) const { return madd(z, o.z, madd(y, o.y, x * o.x)); }
    // math.rs:29-35
    Vec3 cross(const Vec3& o) const {
        return Vec3(madd(y, o.z, -(z * o.y)), madd(z, o.x, -(x * o.z)), madd(x, o.y, -(y * o.x)));
    }
    T length() const { return m_sqrt(madd(z, z, madd(y, y, x * x))); }   // math.rs:38-40
    // math.rs:48-51: returns self unchanged when the length is zero
    Vec3 normalize() const {
        T len = length();
        if (len > T(0)) return *this / len;
        return *this;
    }
    T luminance() const {   // math.rs:133-135 (Rec.709)
        return T(0.2126) * x + T(0.7152) * y + T(0.0722) * z;
    }
    Vec3 operator+(const Vec3& o) const { return Vec3(x + o.x, y + o.y, z + o.z); }   // :139-149
    Vec3 operator-(const Vec3& o) const { return Vec3(x - o.x, y - o.y, z - o.z); }   // :160-170
    Vec3 operator*(T s) const { return Vec3(x * s, y * s, z * s); }                   // :173-183
    Vec3 operator*(const Vec3& o) const { return Vec3(x * o.x, y * o.y, z * o.z); }   // :195-205
    Vec3 operator/(T s) const { return Vec3(x / s, y / s, z / s); }                   // :208-218
    Vec3 operator/(const Vec3& o) const { return Vec3(x / o.x, y / o.y, z / o.z); }   // :221-231
    Vec3 operator-() const { return Vec3(-x, -y, -z); }                               // :234-244
    Vec3& operator+=(const Vec3& o) { x = x + o.x; y = y + o.y; z = z + o.z; return *this; }   // :151-157
};
template <class T> inline Vec3<T> operator*(T s, const Vec3<T>& v) { return v * s; }   // math.rs:186-192

// ----------------------------------------------------------------------------------------------
// Counter-based RNG: Philox4x32 (7 rounds in the render).  Replaces rand::StdRng (main.rs:51-52) by design.
// key = (seed_lo, seed_hi); counter = (pixel, sample, depth, block).
// Slot map (SURVEY Appendix A.5 draw order):
//   camera : depth=0xFFFFFFFF block 0 -> [ox, oy, -, -]                      world.rs:299
//   vertex d block 0 -> [light index, light r1, light r2, bsdf r1]           world.rs:255, shape.rs:111-112/:211-212
//   vertex d block 1 -> [bsdf r2, lobe u, rr u, -]                           mirror.rs:232, rendering.rs:100
// ----------------------------------------------------------------------------------------------
struct U4 { uint32_t w[4]; };
template <int R>
inline U4 philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < R; ++r) {
        uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return U4{{c0, c1, c2, c3}};
}
// The render draws from the 7-round variant (the fastest Crush-resistant one, SC'11 Table 2), like the device; the 10-round
// instantiation exists for the published known answers, which pin the round function both share.
constexpr int kRngRounds = 7;
inline U4 rng_block(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) { return philox4x32<kRngRounds>(c0, c1, c2, c3, k0, k1); }
// 23-bit uniform in the open interval (0,1): ((x>>9)+0.5)/2^23 — exact in float and double.
// (rand's random::<f64>() is [0,1) with 53 bits; the open interval avoids r=0 singularities
//  that are 2^-53-rare in the reference but would be 2^-23-common in FP32)
template <class T> inline T u01(uint32_t x) { return T(((double)(x >> 9) + 0.5) * (1.0 / 8388608.0)); }
// rng.random_range(0..n) (world.rs:255): widening multiply
inline uint32_t urange(uint32_t x, uint32_t n) { return (uint32_t)(((uint64_t)x * n) >> 32); }
// Random word for the light pick: the low byte of each word of block 0 (u01 uses bits 9..31 of the same words, so the
// light index is independent of the four uniforms).  Block 0 = [light r1, light r2, bsdf r1, bsdf r2]; block 1 =
// [lobe u, roulette u, -, -] and is only needed by GGX surfaces and from the roulette depth on (the device skips it otherwise).
inline uint32_t light_word(const U4& b) { return (b.w[0] & 0xffu) | ((b.w[1] & 0xffu) << 8) | ((b.w[2] & 0xffu) << 16) | ((b.w[3] & 0xffu) << 24); }

struct RngKey {
    uint32_t k0, k1, pixel, sample;
    U4 block(uint32_t depth, uint32_t b) const { return rng_block(pixel, sample, depth, b, k0, k1); }
};

// ----------------------------------------------------------------------------------------------
// Ray — src/camera.rs:3-25
// ----------------------------------------------------------------------------------------------
template <class T> struct Ray {
    Vec3<T> origin, direction;
    T eta_ratio;
    Ray() : eta_ratio(T(1)) {}
    // camera.rs:10-16: direction is normalised, eta_ratio starts at 1.0
    Ray(const Vec3<T>& o, const Vec3<T>& d) : origin(o), direction(d.normalize()), eta_ratio(T(1)) {}
    // camera.rs:18-20: origin + direction * t
    Vec3<T> at(T t) const {
        return Vec3<T>(madd(direction.x, t, origin.x), madd(direction.y, t, origin.y), madd(direction.z, t, origin.z));
    }
};

// ----------------------------------------------------------------------------------------------
// Camera — src/camera.rs:27-148.  Construction is host-side f64 (camera.rs:50-82, :94-130);
// only the four cached vectors + width/height reach the sample loop.
// ----------------------------------------------------------------------------------------------
struct CameraDesc {   // same field order as pt_camera in include/pathtrace_b200.h
    double origin[3], lower_left_corner[3], horizontal[3], vertical[3];
    uint32_t width, height;
};
inline CameraDesc camera_new(const double origin[3], uint32_t width, uint32_t height, double screen_distance,
                             double fov_degrees) {   // camera.rs:50-82
    double fov = fov_degrees * (3.14159265358979323846 / 180.0);   // to_radians
    double aspect = (double)width / (double)height;
    double vh = 2.0 * std::tan(fov / 2.0) * screen_distance;   // documented horizontal, sets VERTICAL extent
    double vw = vh * aspect;
    CameraDesc c{};
    double hz[3] = {vw, 0, 0}, vt[3] = {0, vh, 0}, fw[3] = {0, 0, screen_distance};
    for (int i = 0; i < 3; ++i) {
        c.origin[i] = origin[i]; c.horizontal[i] = hz[i]; c.vertical[i] = vt[i];
        c.lower_left_corner[i] = origin[i] - hz[i] / 2.0 - vt[i] / 2.0 - fw[i];
    }
    c.width = width; c.height = height;
    return c;
}
inline CameraDesc camera_look_at(const double origin[3], const double target[3], const double up[3], uint32_t width,
                                 uint32_t height, double fov_degrees) {   // camera.rs:94-130
    using V = Vec3<double>;
    V o(origin[0], origin[1], origin[2]), tg(target[0], target[1], target[2]), upv(up[0], up[1], up[2]);
    double fov = fov_degrees * (3.14159265358979323846 / 180.0);
    double aspect = (double)width / (double)height;
    V w = (o - tg).normalize(), u = upv.cross(w).normalize(), v = w.cross(u);
    double vh = 2.0 * std::tan(fov / 2.0) * 1.0, vw = vh * aspect;
    V hz = u * vw, vt = v * vh;
    V llc = o - hz / 2.0 - vt / 2.0 - w * 1.0;
    CameraDesc c{};
    double* dst[4] = {c.origin, c.lower_left_corner, c.horizontal, c.vertical};
    V src[4] = {o, llc, hz, vt};
    for (int k = 0; k < 4; ++k) { dst[k][0] = src[k].x; dst[k][1] = src[k].y; dst[k][2] = src[k].z; }
    c.width = width; c.height = height;
    return c;
}
template <class T> struct Camera {
    Vec3<T> origin, llc, horizontal, vertical;
    uint32_t width, height;
    explicit Camera(const CameraDesc& d)
        : origin(T(d.origin[0]), T(d.origin[1]), T(d.origin[2])),
          llc(T(d.lower_left_corner[0]), T(d.lower_left_corner[1]), T(d.lower_left_corner[2])),
          horizontal(T(d.horizontal[0]), T(d.horizontal[1]), T(d.horizontal[2])),
          vertical(T(d.vertical[0]), T(d.vertical[1]), T(d.vertical[2])), width(d.width), height(d.height) {}
    // camera.rs:139-147.  Called with y = H-1-y_img (world.rs:299).  Note the W-1 / H-1 divisors.
    Ray<T> get_ray_with_offset(uint32_t x, uint32_t y, T ox, T oy) const {
        T u = (T((double)x) + ox) / T((double)(width - 1));
        T v = (T((double)y) + oy) / T((double)(height - 1));
        Vec3<T> dir = llc + horizontal * u + vertical * v - origin;
        return Ray<T>(origin, dir);
    }
};

// ----------------------------------------------------------------------------------------------
// HitRecord — src/objects/base.rs:6-34
// ----------------------------------------------------------------------------------------------
template <class T> struct HitRecord {
    Vec3<T> point, normal;
    T t;
    bool front_face;
    HitRecord() : t(T(0)), front_face(false) {}
    HitRecord(const Vec3<T>& p, const Vec3<T>& outward, T tt, const Ray<T>& ray) : point(p), t(tt) {
        front_face = ray.direction.dot(outward) < T(0);   // base.rs:20
        normal = front_face ? outward : -outward;         // base.rs:21-25
    }
};

// ----------------------------------------------------------------------------------------------
// Shapes — src/objects/shape.rs.  kind 0 = SphereShape{center,radius} (:38-43),
// kind 1 = TriangleShape{v0,v1,v2} (:148-152).
// ----------------------------------------------------------------------------------------------
enum : uint32_t { SHAPE_SPHERE = 0, SHAPE_TRIANGLE = 1 };
struct ShapeDesc { uint32_t kind; float v[9]; };    // mirrors pt_shape (what crosses the C ABI: f32)
struct ShapeDescD { uint32_t kind; double v[9]; };  // the reference's own f64 scene data (world.rs:80-211 builds from f64 literals)

template <class T> struct SurfaceSample { Vec3<T> point, normal; T pdf_omega; Vec3<T> light_dir; T d; };

template <class T> struct Shape {
    uint32_t kind;
    Vec3<T> a, b, c;   // sphere: a = center, b.x = radius; triangle: v0, v1, v2
    template <class SD> explicit Shape(const SD& s) : kind(s.kind) {
        a = Vec3<T>(T((double)s.v[0]), T((double)s.v[1]), T((double)s.v[2]));
        b = Vec3<T>(T((double)s.v[3]), T((double)s.v[4]), T((double)s.v[5]));
        c = Vec3<T>(T((double)s.v[6]), T((double)s.v[7]), T((double)s.v[8]));
    }
    bool hit(const Ray<T>& ray, T t_min, T t_max, HitRecord<T>& out) const {
        return kind == SHAPE_SPHERE ? hit_sphere(ray, t_min, t_max, out) : hit_triangle(ray, t_min, t_max, out);
    }
    // shape.rs:53-89.  double / Counted: the reference's half-b quadratic, literally.
    // float: the cancellation-free form of the same quadratic (Haines et al., Ray Tracing Gems ch. 7) that the
    // device uses — the naive FP32 discriminant loses ~|oc|^2/r^2 ulps and cannot hold 1e-5 relative on t for
    // small distant spheres.  Same roots, same near/far/range selection.
    bool hit_sphere(const Ray<T>& ray, T t_min, T t_max, HitRecord<T>& out) const {
        const Vec3<T>& center = a; T radius = b.x;
        Vec3<T> oc = ray.origin - center;
        T aa = ray.direction.dot(ray.direction);
        T root;
        if constexpr (std::is_same<T, float>::value) {
            T bp = -oc.dot(ray.direction);
            T k = bp / aa;
            Vec3<T> l(madd(k, ray.direction.x, oc.x), madd(k, ray.direction.y, oc.y), madd(k, ray.direction.z, oc.z));
            T disc = radius * radius - l.dot(l);
            if (disc < T(0)) return false;
            T cc = oc.dot(oc) - radius * radius;
            T q = bp + std::copysign(m_sqrt(aa * disc), bp);
            T t_near = bp >= T(0) ? cc / q : q / aa;
            T t_far = bp >= T(0) ? q / aa : cc / q;
            root = t_near;
            if (root < t_min || t_max < root) {
                root = t_far;
                if (root < t_min || t_max < root) return false;
            }
        } else {
            T half_b = oc.dot(ray.direction);
            T cc = oc.dot(oc) - radius * radius;
            T disc = half_b * half_b - aa * cc;
            if (disc < T(0)) return false;
            T sqrtd = m_sqrt(disc);
            root = (-half_b - sqrtd) / aa;
            if (root < t_min || t_max < root) {
                root = (-half_b + sqrtd) / aa;
                if (root < t_min || t_max < root) return false;
            }
        }
        Vec3<T> p = ray.at(root);
        Vec3<T> outward = (p - center) / radius;
        out = HitRecord<T>(p, outward, root, ray);
        return true;
    }
    // shape.rs:161-198 (Möller–Trumbore, two-sided)
    bool hit_triangle(const Ray<T>& ray, T t_min, T t_max, HitRecord<T>& out) const {
        Vec3<T> e1 = b - a, e2 = c - a;
        Vec3<T> h = ray.direction.cross(e2);
        T det = e1.dot(h);
        if (m_abs(det) < T(1e-8)) return false;
        T f = T(1) / det;
        Vec3<T> s = ray.origin - a;
        T u = f * s.dot(h);
        if (!(u >= T(0) && u <= T(1))) return false;   // !(0.0..=1.0).contains(&u): NaN misses
        Vec3<T> q = s.cross(e1);
        T v = f * ray.direction.dot(q);
        if (v < T(0) || u + v > T(1)) return false;
        T t = f * e2.dot(q);
        if (t < t_min || t > t_max) return false;
        Vec3<T> p = ray.at(t);
        Vec3<T> outward = e1.cross(e2).normalize();
        out = HitRecord<T>(p, outward, t, ray);
        return true;
    }
    // shape.rs:91-145 (sphere) and :200-242 (triangle).  target != nullptr: pdf evaluation at that
    // point, no random numbers used.  r1, r2 are the two uniforms the reference draws in order.
    SurfaceSample<T> sample_surface_from_point(const HitRecord<T>& from, const HitRecord<T>* target, T r1, T r2) const {
        return kind == SHAPE_SPHERE ? sample_sphere(from, target, r1, r2) : sample_triangle(from, target, r1, r2);
    }
    SurfaceSample<T> sample_sphere(const HitRecord<T>& from, const HitRecord<T>* target, T r1, T r2) const {
        const Vec3<T>& center = a; T radius = b.x;
        Vec3<T> to_center = center - from.point;
        T dist_sq = to_center.dot(to_center);
        T sin2max = (radius * radius) / dist_sq;
        T cos_max = m_sqrt(m_max(T(1) - sin2max, T(0)));
        T solid = T(2) * pi<T>() * (T(1) - cos_max);
        T pdf_omega = T(1) / solid;
        Vec3<T> sp;
        if (target) {
            sp = target->point;
        } else {
            T cos_t = T(1) - r1 + r1 * cos_max;
            T sin_t = m_sqrt(m_max(T(1) - cos_t * cos_t, T(0)));
            T phi = T(2) * pi<T>() * r2;
            Vec3<T> w = to_center.normalize();
            Vec3<T> up = m_abs(w.y) > T(0.999) ? Vec3<T>(T(1), T(0), T(0)) : Vec3<T>(T(0), T(1), T(0));
            Vec3<T> u = up.cross(w).normalize();
            Vec3<T> v = w.cross(u);
            Vec3<T> dir = u * (sin_t * m_cos(phi)) + v * (sin_t * m_sin(phi)) + w * cos_t;
            Ray<T> sray(from.point, dir);
            Vec3<T> oc = sray.origin - center;
            T aa = dir.dot(dir);            // shape.rs:132 uses the un-normalised direction here
            T hb = oc.dot(dir);
            T cc = oc.dot(oc) - radius * radius;
            T disc = hb * hb - aa * cc;
            T t = (-hb - m_sqrt(disc)) / aa;
            sp = sray.at(t);
        }
        SurfaceSample<T> s;
        s.point = sp;
        s.normal = (sp - center).normalize();
        Vec3<T> ld = sp - from.point;
        s.d = ld.length();
        s.pdf_omega = pdf_omega;
        s.light_dir = ld.normalize();
        return s;
    }
    SurfaceSample<T> sample_triangle(const HitRecord<T>& from, const HitRecord<T>* target, T r1, T r2) const {
        Vec3<T> sp;
        if (target) {
            sp = target->point;
        } else {
            T sq = m_sqrt(r1);
            T u = T(1) - sq;
            T v = r2 * sq;
            sp = a + (b - a) * u + (c - a) * v;
        }
        Vec3<T> e1 = b - a, e2 = c - a;
        Vec3<T> n = e1.cross(e2).normalize();
        T area = e1.cross(e2).length() * T(0.5);
        Vec3<T> to_light = sp - from.point;
        T d = to_light.length();
        Vec3<T> ld = to_light / d;
        T cos_light = m_abs(n.dot(-ld));   // two-sided emitters
        T pdf_area = T(1) / area;
        T pdf_omega = cos_light > T(1e-8) ? pdf_area * (d * d) / cos_light : T(1e-8);
        return SurfaceSample<T>{sp, n, pdf_omega, ld, d};
    }
};

// ----------------------------------------------------------------------------------------------
// Materials — src/objects/material.rs, src/objects/mirror.rs
// kind 0 LambertianCosineWeighted{albedo}            p = [albedo.xyz]
// kind 1 Emissive{emission}                          p = [emission.xyz]
// kind 2 Mirror{roughness,color,metallic,ior}        p = [roughness, color.xyz, metallic, ior]
// kind 3 OrenNayar{albedo,roughness,a,b}             p = [albedo.xyz, roughness, a, b]
// ----------------------------------------------------------------------------------------------
enum : uint32_t { MAT_LAMBERT = 0, MAT_EMISSIVE = 1, MAT_MIRROR = 2, MAT_ORENNAYAR = 3 };
struct MaterialDesc { uint32_t kind; float p[8]; };    // mirrors pt_material
struct MaterialDescD { uint32_t kind; double p[8]; };  // f64 parameters as the reference holds them

template <class T> struct BsdfSample { Vec3<T> dir, f; T pdf, cos_theta; };
template <class T> struct BsdfEval { Vec3<T> f; T pdf; };

// Shared tangent frame (material.rs:105-119, mirror.rs:21-27, material.rs:215-222)
template <class T> inline void tangent_frame(const Vec3<T>& n, Vec3<T>& tangent, Vec3<T>& bitangent) {
    Vec3<T> up = m_abs(n.y) > T(0.999) ? Vec3<T>(T(1), T(0), T(0)) : Vec3<T>(T(0), T(1), T(0));
    tangent = up.cross(n).normalize();
    bitangent = n.cross(tangent);
}
// material.rs:93-122 / :268-295 — cosine-weighted hemisphere sample (phi from r1, cos from r2)
template <class T> inline Vec3<T> cosine_sample(const Vec3<T>& n, T r1, T r2) {
    T phi = T(2) * pi<T>() * r1;
    T cos_t = m_sqrt(r2);
    T sin_t = m_sqrt(T(1) - cos_t * cos_t);
    T x = sin_t * m_cos(phi), y = sin_t * m_sin(phi), z = cos_t;
    Vec3<T> tg, bt;
    tangent_frame(n, tg, bt);
    return (tg * x + bt * y + n * z).normalize();
}
template <class T> inline T powi2(T x) { return x * x; }
template <class T> inline T powi5(T x) { T x2 = x * x; return x2 * x2 * x; }   // powi(5)

template <class T> struct Material {
    uint32_t kind;
    Vec3<T> color;          // albedo / emission / Mirror.color
    T roughness, metallic, ior, on_a, on_b;
    template <class MD> explicit Material(const MD& m) : kind(m.kind), roughness(T(0)), metallic(T(0)), ior(T(1)), on_a(T(1)), on_b(T(0)) {
        auto P = [&](int i) { return T((double)m.p[i]); };
        switch (kind) {
            case MAT_LAMBERT: case MAT_EMISSIVE: color = Vec3<T>(P(0), P(1), P(2)); break;
            case MAT_MIRROR: roughness = P(0); color = Vec3<T>(P(1), P(2), P(3)); metallic = P(4); ior = P(5); break;
            case MAT_ORENNAYAR: color = Vec3<T>(P(0), P(1), P(2)); roughness = P(3); on_a = P(4); on_b = P(5); break;
        }
    }
    // material.rs:51-53, mirror.rs:317-319
    T get_eta() const { return kind == MAT_MIRROR ? ior : T(1); }
    // material.rs:62-64, :160-162
    Vec3<T> emit() const { return kind == MAT_EMISSIVE ? color : Vec3<T>::zero(); }

    // ---- Mirror internals --------------------------------------------------------------------
    Vec3<T> get_f(T cos_theta) const {   // mirror.rs:126-132
        T f0d = powi2((T(1) - ior) / (T(1) + ior));
        Vec3<T> f0 = Vec3<T>(f0d, f0d, f0d) * (T(1) - metallic) + color * metallic;
        return f0 + (Vec3<T>::one() - f0) * powi5(T(1) - cos_theta);
    }
    T get_g1(T cos_theta) const {   // mirror.rs:136-149
        if (cos_theta <= T(0)) return T(0);
        T alpha = roughness * roughness, alpha2 = alpha * alpha;
        T term = alpha2 + (T(1) - alpha2) * (cos_theta * cos_theta);
        return T(2) * cos_theta / (cos_theta + m_sqrt(term));
    }
    T get_g(T ci, T co) const {   // mirror.rs:153-175
        if (ci <= T(0) || co <= T(0)) return T(0);
        T alpha = roughness * roughness, alpha2 = alpha * alpha;
        auto lambda = [&](T c) {
            T num = m_sqrt(alpha2 + (T(1) - alpha2) * (c * c));
            return (num - c) / (T(2) * c);
        };
        return T(1) / (T(1) + lambda(ci) + lambda(co));
    }
    Vec3<T> sample_ggx_vndf(const Vec3<T>& view, const Vec3<T>& n, T r1, T r2) const {   // mirror.rs:17-60
        T alpha = roughness * roughness;
        Vec3<T> tg, bt;
        tangent_frame(n, tg, bt);
        Vec3<T> vl(view.dot(tg), view.dot(bt), view.dot(n));
        Vec3<T> vh = Vec3<T>(alpha * vl.x, alpha * vl.y, vl.z).normalize();
        T lensq = vh.x * vh.x + vh.y * vh.y;
        Vec3<T> t1 = lensq > T(0) ? Vec3<T>(-vh.y, vh.x, T(0)) * (T(1) / m_sqrt(lensq)) : Vec3<T>(T(1), T(0), T(0));
        Vec3<T> t2 = vh.cross(t1);
        T r = m_sqrt(r1);
        T phi = T(2) * pi<T>() * r2;
        T c1 = r * m_cos(phi);
        T c2 = r * m_sin(phi);
        T s = T(0.5) * (T(1) + vh.z);
        c2 = (T(1) - s) * m_sqrt(T(1) - c1 * c1) + s * c2;
        Vec3<T> nh = t1 * c1 + t2 * c2 + vh * m_sqrt(m_max(T(1) - c1 * c1 - c2 * c2, T(0)));
        Vec3<T> ne = Vec3<T>(alpha * nh.x, alpha * nh.y, m_max(nh.z, T(0))).normalize();
        return (tg * ne.x + bt * ne.y + n * ne.z).normalize();
    }
    BsdfEval<T> mirror_brdf(const Ray<T>& ray, const Vec3<T>& o, const Vec3<T>& n) const {   // mirror.rs:62-88
        Vec3<T> i = -ray.direction;
        T alpha = roughness * roughness, alpha2 = alpha * alpha;
        Vec3<T> h = (i + o).normalize();
        T denom = (n.dot(h) * n.dot(h)) * (alpha2 - T(1)) + T(1);
        T d = alpha2 / (pi<T>() * denom * denom);
        T i_n = m_max(n.dot(i), T(0)), o_n = m_max(n.dot(o), T(0));
        T g = get_g(i_n, o_n);
        T cos_theta = m_max(i.dot(h), T(0));
        Vec3<T> f = get_f(cos_theta);
        T denom_brdf = T(4) * i_n * o_n;
        Vec3<T> brdf = (d * g) * f / denom_brdf;   // d * g * f / denom_brdf, left to right
        T i_h = m_abs(i.dot(h));
        T pdf = d * m_abs(n.dot(h)) / (T(4) * i_h);
        return BsdfEval<T>{brdf, pdf};
    }
    BsdfEval<T> mirror_btdf(const Ray<T>& ray, const Vec3<T>& o, const Vec3<T>& n) const {   // mirror.rs:90-124
        Vec3<T> i = -ray.direction;
        T eta = ray.eta_ratio;
        Vec3<T> h = -((i * eta + o).normalize());
        T alpha = roughness * roughness, alpha2 = alpha * alpha;
        T n_h = n.dot(h);
        T denom = (n_h * n_h) * (alpha2 - T(1)) + T(1);
        T d = alpha2 / (pi<T>() * denom * denom);
        T i_n = m_abs(n.dot(i)), o_n = m_abs(n.dot(o));
        T g = get_g(i_n, o_n);
        T i_h = i.dot(h), o_h = o.dot(h);
        T cos_theta = m_abs(i_h);
        T dt = eta * i_h + o_h;
        Vec3<T> f = get_f(cos_theta);
        // (1 - f) * d * g * |i.h| * |o.h| / (i_n * o_n * dt * dt)
        Vec3<T> btdf = (Vec3<T>::one() - f) * d * g * m_abs(i_h) * m_abs(o_h) / (i_n * o_n * dt * dt);
        T jac = m_abs(o_h) / (dt * dt);
        T pdf = d * m_abs(n_h) * jac;
        return BsdfEval<T>{btdf, pdf};
    }

    // ---- Material::bsdf_pdf ------------------------------------------------------------------
    BsdfEval<T> bsdf_pdf(const Ray<T>& ray, const Vec3<T>& o, const Vec3<T>& n) const {
        switch (kind) {
            case MAT_LAMBERT: {   // material.rs:86-91, :80-83
                return BsdfEval<T>{color / pi<T>(), m_max(o.dot(n), T(0)) / pi<T>()};
            }
            case MAT_EMISSIVE: return BsdfEval<T>{Vec3<T>::zero(), T(1)};   // material.rs:139-148
            case MAT_MIRROR: {   // mirror.rs:179-198
                Vec3<T> i = -ray.direction;
                bool refl = i.dot(n) * o.dot(n) > T(0);
                if (metallic > T(0.99) && !refl) return BsdfEval<T>{Vec3<T>::zero(), T(1)};
                return refl ? mirror_brdf(ray, o, n) : mirror_btdf(ray, o, n);
            }
            default: {   // OrenNayar, material.rs:226-266
                Vec3<T> i = -ray.direction;
                T ci = m_max(i.dot(n), T(0)), co = m_max(o.dot(n), T(0));
                T si = m_sqrt(m_max(T(1) - ci * ci, T(0))), so = m_sqrt(m_max(T(1) - co * co, T(0)));
                Vec3<T> tg, bt;
                tangent_frame(n, tg, bt);   // compute_tangent :215-222; compute_azimuth :207-212
                T phi_i = m_atan2(i.dot(bt), i.dot(tg));
                T phi_o = m_atan2(o.dot(bt), o.dot(tg));
                T cpd = m_max(m_cos(phi_i - phi_o), T(0));
                T sin_alpha, tan_beta;
                if (ci > co) { tan_beta = ci > T(1e-6) ? si / ci : T(0); sin_alpha = so; }
                else         { tan_beta = co > T(1e-6) ? so / co : T(0); sin_alpha = si; }
                T term = on_a + on_b * cpd * sin_alpha * tan_beta;
                return BsdfEval<T>{color * (term / pi<T>()), co / pi<T>()};
            }
        }
    }
    // ---- Material::bsdf_pdf_sample -----------------------------------------------------------
    // r1, r2: the sampler's two uniforms in draw order; r3: Mirror's lobe uniform (mirror.rs:232).
    BsdfSample<T> bsdf_pdf_sample(const Ray<T>& ray, const Vec3<T>& n, T r1, T r2, T r3) const {
        if (kind == MAT_MIRROR) return mirror_sample(ray, n, r1, r2, r3);
        // default impl, material.rs:29-40
        Vec3<T> dir = (kind == MAT_EMISSIVE) ? n : cosine_sample(n, r1, r2);   // material.rs:150-158
        BsdfEval<T> e = bsdf_pdf(ray, dir, n);
        return BsdfSample<T>{dir, e.f, e.pdf, m_max(dir.dot(n), T(0))};
    }
    BsdfSample<T> mirror_sample(const Ray<T>& ray, const Vec3<T>& n, T r1, T r2, T r3) const {   // mirror.rs:200-305
        const BsdfSample<T> fail{n, Vec3<T>::zero(), T(1), T(0)};
        Vec3<T> i = -ray.direction;
        T i_dot_n = i.dot(n);
        T eta = ray.eta_ratio;
        Vec3<T> h = sample_ggx_vndf(i, n, r1, r2);
        T i_h = i.dot(h);
        if (i_h <= T(0)) return fail;
        Vec3<T> fr = get_f(i_h);
        T sin2_i = T(1) - i_h * i_h;
        T cos2_t = T(1) - (eta * eta) * sin2_i;
        bool tir = cos2_t < T(0);
        T rr_f = fr.x;
        if (tir || metallic > T(0.99)) { rr_f = T(1); fr = Vec3<T>::one(); }
        bool is_reflect = r3 < rr_f;
        T alpha = roughness * roughness, alpha2 = alpha * alpha;
        T n_h = n.dot(h);
        T denom = (n_h * n_h) * (alpha2 - T(1)) + T(1);
        T d = alpha2 / (pi<T>() * denom * denom);
        if (is_reflect) {
            Vec3<T> o = (T(2) * i_h) * h - i;
            Vec3<T> on = o.normalize();
            T o_n = m_max(n.dot(on), T(0));
            T i_n = m_max(i_dot_n, T(0));
            T g = get_g(i_n, o_n);
            T denom_brdf = T(4) * i_n * o_n;
            Vec3<T> brdf = fr * d * g / (denom_brdf * rr_f);
            T g1 = get_g1(i_n);
            T pdf_vndf = g1 * d * m_max(i_h, T(0)) / i_n;
            T pdf = pdf_vndf / (T(4) * m_abs(i_h));
            if (!m_finite(brdf.x) || !m_finite(brdf.y) || !m_finite(brdf.z) || !m_finite(pdf) || pdf <= T(0)) return fail;
            return BsdfSample<T>{on, brdf, pdf, o_n};
        } else {
            T cos_t = m_sqrt(cos2_t);
            Vec3<T> o = h * (eta * i_h - cos_t) - i * eta;
            Vec3<T> on = o.normalize();
            T o_h = on.dot(h);
            T o_n = m_abs(n.dot(on));
            T i_n = m_abs(i_dot_n);
            T dt = eta * i_h + o_h;
            T g = get_g(i_n, o_n);
            Vec3<T> one_f = Vec3<T>::one() - fr;
            Vec3<T> btdf = one_f * d * g * m_abs(i_h) * m_abs(o_h) / (i_n * o_n * dt * dt * (T(1) - rr_f));
            T jac = m_abs(o_h) / (dt * dt);
            T g1 = get_g1(i_n);
            T pdf_vndf = g1 * d * m_max(i_h, T(0)) / i_n;
            T pdf = pdf_vndf * jac;
            if (!m_finite(btdf.x) || !m_finite(btdf.y) || !m_finite(btdf.z) || !m_finite(pdf) || pdf <= T(0)) return fail;
            return BsdfSample<T>{on, btdf, pdf, o_n};
        }
    }
};

// ----------------------------------------------------------------------------------------------
// World — src/world.rs (scene container, hit_scene, sample_light_point, render_pixel)
// ----------------------------------------------------------------------------------------------
struct RenderParams {   // mirrors pt_render_params; defaults = the reference's constants
    uint32_t spp = 3000;            // SAMPLE_NUM world.rs:18
    uint32_t sample_offset = 0;
    uint32_t min_depth = 4;         // MIN_DEPTH rendering.rs:6
    uint32_t max_depth = 50;        // MAX_DEPTH rendering.rs:7
    uint32_t strategy = 0;          // 0 = mis (default feature, Cargo.toml:7), 1 = brdf_only
    uint64_t seed = 0x5EED;
    float t_min = 0.001f;           // rendering.rs:41, :64, :105
};
enum : uint32_t { STRATEGY_MIS = 0, STRATEGY_BRDF_ONLY = 1 };

struct RayCounters { uint64_t primary = 0, shadow = 0, scattered = 0, retrace = 0; };

template <class T> struct LightSample { Vec3<T> point, emission; T pdf; };

template <class T> struct World {
    Camera<T> camera;
    std::vector<Shape<T>> shapes;
    std::vector<Material<T>> materials;
    std::vector<uint32_t> light_indices;
    RenderParams prm;
    T t_min;

    template <class SD, class MD> World(const CameraDesc& cam, const SD* sh, const MD* mt, uint32_t n, const RenderParams& p)
        : camera(cam), prm(p), t_min(T((double)p.t_min)) {
        for (uint32_t i = 0; i < n; ++i) { shapes.emplace_back(sh[i]); materials.emplace_back(mt[i]); }
        // world.rs:214-225: a light is any object whose emit() has non-zero length
        for (uint32_t i = 0; i < n; ++i)
            if (to_double(materials[i].emit().length()) > 0.0) light_indices.push_back(i);
    }
    static T inf() { return T(std::numeric_limits<double>::infinity()); }

    // world.rs:270-290 — brute force, shrinking t_max; ties go to the later object (Q9)
    bool hit_scene(const Ray<T>& ray, T tmin, T tmax, HitRecord<T>& rec, uint32_t& obj) const {
        bool any = false;
        T closest = tmax;
        HitRecord<T> tmp;
        for (uint32_t i = 0; i < shapes.size(); ++i) {
            if (shapes[i].hit(ray, tmin, closest, tmp)) { closest = tmp.t; rec = tmp; obj = i; any = true; }
        }
        return any;
    }
    // world.rs:251-267
    bool sample_light_point(const HitRecord<T>& hit, uint32_t w_idx, T r1, T r2, LightSample<T>& out) const {
        if (light_indices.empty()) return false;
        uint32_t n = (uint32_t)light_indices.size();
        uint32_t li = light_indices[urange(w_idx, n)];
        SurfaceSample<T> s = shapes[li].sample_surface_from_point(hit, nullptr, r1, r2);
        out.point = s.point;
        out.emission = materials[li].emit();
        out.pdf = s.pdf_omega / T((double)n);
        return true;
    }
    // rendering.rs:20-25
    T get_eta_from_object(uint32_t obj, const HitRecord<T>& hit) const {
        return hit.front_face ? T(1) / materials[obj].get_eta() : materials[obj].get_eta();
    }
    // rendering.rs:91-98
    T rr_prob(const Vec3<T>& next_throughput, uint32_t depth) const {
        if (depth < prm.min_depth) return T(1);
        T l = m_min(next_throughput.luminance(), T(1));
        if (depth >= prm.max_depth) {
            T p = T(1);
            for (uint32_t k = 0; k < depth - prm.min_depth; ++k) p = p * T(0.5);   // 0.5.powi(depth-4)
            return l * p;
        }
        return l;
    }

    // ---- MisStrategy::ray_color, literal recursive form — rendering.rs:34-142 -----------------
    // Deviation (documented): a non-positive or non-finite pdf / throughput terminates
    // the path with zero contribution instead of propagating NaN.
    Vec3<T> ray_color_mis(Ray<T>& ray, uint32_t depth, const RngKey& rng, Vec3<T> throughput, RayCounters* rc) const {
        HitRecord<T> hit; uint32_t obj = 0;
        if (rc) { if (depth == 0) rc->primary++; else rc->retrace++; }
        if (!hit_scene(ray, t_min, inf(), hit, obj)) return Vec3<T>::zero();   // :140-141
        Vec3<T> emitted = materials[obj].emit();
        if (to_double(emitted.length()) > 0.0) return depth == 0 ? emitted : Vec3<T>::zero();   // :42-49
        U4 b0 = rng.block(depth, 0), b1 = rng.block(depth, 1);
        Vec3<T> total = Vec3<T>::zero(), direct = Vec3<T>::zero();
        LightSample<T> ls;
        if (sample_light_point(hit, light_word(b0), u01<T>(b0.w[0]), u01<T>(b0.w[1]), ls)) {   // :56
            Vec3<T> to_light = ls.point - hit.point;
            T distance = to_light.length();
            Vec3<T> light_dir = to_light.normalize();
            Ray<T> shadow(hit.point, light_dir);
            HitRecord<T> sh; uint32_t so = 0;
            if (rc) rc->shadow++;
            bool visible = !hit_scene(shadow, t_min, distance - T(0.001), sh, so);   // :62-65
            if (visible) {
                T cos_theta = m_abs(hit.normal.dot(light_dir));
                BsdfEval<T> e = materials[obj].bsdf_pdf(ray, light_dir, hit.normal);   // inherited eta (Q3)
                T w_nee = ls.pdf / (ls.pdf + e.pdf);
                direct += w_nee * e.f * ls.emission * cos_theta / ls.pdf;   // :75-76
            }
        }
        total += direct / T(1);   // NUM_LIGHT_SAMPLES = 1, :81
        ray.eta_ratio = get_eta_from_object(obj, hit);   // :83
        BsdfSample<T> bs = materials[obj].bsdf_pdf_sample(ray, hit.normal, u01<T>(b0.w[2]), u01<T>(b0.w[3]), u01<T>(b1.w[0]));
        Ray<T> scattered(hit.point, bs.dir);
        scattered.eta_ratio = get_eta_from_object(obj, hit);   // :87
        Vec3<T> next_tp = throughput * bs.f * bs.cos_theta / bs.pdf;   // :89
        if (!(bs.pdf > T(0)) || !m_finite(next_tp.x) || !m_finite(next_tp.y) || !m_finite(next_tp.z)) return Vec3<T>::zero();
        T q = rr_prob(next_tp, depth);
        if (u01<T>(b1.w[1]) > q) return Vec3<T>::zero();   // :100-102 (drops `direct` too, Q1)
        HitRecord<T> h2; uint32_t o2 = 0;
        if (rc) rc->scattered++;
        if (hit_scene(scattered, t_min, inf(), h2, o2)) {   // :104-105
            Vec3<T> le = materials[o2].emit();
            if (to_double(le.length()) > 0.0) {
                SurfaceSample<T> s = shapes[o2].sample_surface_from_point(hit, &h2, T(0), T(0));   // :114-116
                T w_bsdf = bs.pdf / (bs.pdf + s.pdf_omega);   // not divided by n_lights (Q2)
                total += w_bsdf * bs.f * le * bs.cos_theta / (bs.pdf * q);
            } else {
                Vec3<T> incoming = ray_color_mis(scattered, depth + 1, rng, next_tp / q, rc);   // re-traces (Q8)
                total += bs.f * incoming * bs.cos_theta / (bs.pdf * q);
            }
        }
        return total;
    }
    // ---- BrdfOnlyStrategy::ray_color, literal — rendering.rs:213-265 --------------------------
    Vec3<T> ray_color_brdf_only(Ray<T>& ray, uint32_t depth, const RngKey& rng, Vec3<T> throughput, RayCounters* rc) const {
        HitRecord<T> hit; uint32_t obj = 0;
        if (rc) { if (depth == 0) rc->primary++; else rc->scattered++; }
        if (!hit_scene(ray, t_min, inf(), hit, obj)) return Vec3<T>::zero();
        Vec3<T> emitted = materials[obj].emit();
        if (to_double(emitted.length()) > 0.0) return emitted;
        U4 b0 = rng.block(depth, 0), b1 = rng.block(depth, 1);
        ray.eta_ratio = get_eta_from_object(obj, hit);
        BsdfSample<T> bs = materials[obj].bsdf_pdf_sample(ray, hit.normal, u01<T>(b0.w[2]), u01<T>(b0.w[3]), u01<T>(b1.w[0]));
        Ray<T> scattered(hit.point, bs.dir);
        scattered.eta_ratio = get_eta_from_object(obj, hit);
        Vec3<T> next_tp = throughput * bs.f * bs.cos_theta / bs.pdf;
        if (!(bs.pdf > T(0)) || !m_finite(next_tp.x) || !m_finite(next_tp.y) || !m_finite(next_tp.z)) return Vec3<T>::zero();
        T q = rr_prob(next_tp, depth);
        if (u01<T>(b1.w[1]) > q) return Vec3<T>::zero();
        Vec3<T> incoming = ray_color_brdf_only(scattered, depth + 1, rng, next_tp / q, rc);
        return bs.f * incoming * bs.cos_theta / (bs.pdf * q);
    }

    // ---- Iterative form of both strategies: the shape the GPU executes ---
    // No re-trace (Q8); radiance accumulated as beta-weighted sums.  Must agree with the recursive
    // form sample-for-sample up to floating-point re-association (tests/test_oracle_kat.py: test_recursive_equals_iterative).
    Vec3<T> radiance_iterative(Ray<T> ray, const RngKey& rng, RayCounters* rc, uint32_t* path_len = nullptr) const {
        const bool mis = prm.strategy == STRATEGY_MIS;
        Vec3<T> beta = Vec3<T>::one(), L = Vec3<T>::zero();
        HitRecord<T> hit; uint32_t obj = 0;
        if (rc) rc->primary++;
        if (path_len) *path_len = 0;
        bool hit0;
        { CatScope cs(1); hit0 = hit_scene(ray, t_min, inf(), hit, obj); }
        if (!hit0) return L;
        Vec3<T> le0 = materials[obj].emit();
        if (to_double(le0.length()) > 0.0) return le0;
        for (uint32_t depth = 0;; ++depth) {
            if (path_len) *path_len = depth + 1;
            U4 b0 = rng.block(depth, 0), b1 = rng.block(depth, 1);
            Vec3<T> D = Vec3<T>::zero();
            LightSample<T> ls;
            if (mis && sample_light_point(hit, light_word(b0), u01<T>(b0.w[0]), u01<T>(b0.w[1]), ls)) {
                Vec3<T> to_light = ls.point - hit.point;
                T distance = to_light.length();
                Vec3<T> light_dir = to_light.normalize();
                Ray<T> shadow(hit.point, light_dir);
                HitRecord<T> sh; uint32_t so = 0;
                if (rc) rc->shadow++;
                bool occluded;
                { CatScope cs(2); occluded = hit_scene(shadow, t_min, distance - T(0.001), sh, so); }
                if (!occluded) {
                    T cos_theta = m_abs(hit.normal.dot(light_dir));
                    BsdfEval<T> e = materials[obj].bsdf_pdf(ray, light_dir, hit.normal);
                    T w_nee = ls.pdf / (ls.pdf + e.pdf);
                    D = w_nee * e.f * ls.emission * cos_theta / ls.pdf;
                }
            }
            ray.eta_ratio = get_eta_from_object(obj, hit);
            BsdfSample<T> bs = materials[obj].bsdf_pdf_sample(ray, hit.normal, u01<T>(b0.w[2]), u01<T>(b0.w[3]), u01<T>(b1.w[0]));
            Vec3<T> tn = beta * bs.f * bs.cos_theta / bs.pdf;
            if (!(bs.pdf > T(0)) || !m_finite(tn.x) || !m_finite(tn.y) || !m_finite(tn.z)) return L;
            T q = rr_prob(tn, depth);
            if (u01<T>(b1.w[1]) > q) return L;
            L += beta * D;
            beta = tn / q;
            Ray<T> scattered(hit.point, bs.dir);
            scattered.eta_ratio = ray.eta_ratio;
            HitRecord<T> h2; uint32_t o2 = 0;
            if (rc) rc->scattered++;
            bool hit2;
            { CatScope cs(1); hit2 = hit_scene(scattered, t_min, inf(), h2, o2); }
            if (!hit2) return L;
            Vec3<T> le = materials[o2].emit();
            if (to_double(le.length()) > 0.0) {
                if (!mis) return L + beta * le;
                SurfaceSample<T> s = shapes[o2].sample_surface_from_point(hit, &h2, T(0), T(0));
                return L + beta * (bs.pdf / (bs.pdf + s.pdf_omega)) * le;
            }
            ray = scattered; hit = h2; obj = o2;
        }
    }

    RngKey key(uint32_t x, uint32_t y_img, uint32_t sample) const {
        return RngKey{(uint32_t)prm.seed, (uint32_t)(prm.seed >> 32), y_img * camera.width + x, sample};
    }
    Ray<T> primary_ray(uint32_t x, uint32_t y_img, const RngKey& k) const {   // world.rs:297-299
        U4 c = k.block(0xFFFFFFFFu, 0);
        return camera.get_ray_with_offset(x, camera.height - 1 - y_img, u01<T>(c.w[0]), u01<T>(c.w[1]));
    }
    // One sample of pixel (x, y_img); recursive = literal reference control flow.
    Vec3<T> sample_radiance(uint32_t x, uint32_t y_img, uint32_t sample, bool recursive, RayCounters* rc, uint32_t* plen = nullptr) const {
        RngKey k = key(x, y_img, sample);
        Ray<T> ray = primary_ray(x, y_img, k);
        if (!recursive) return radiance_iterative(ray, k, rc, plen);
        return prm.strategy == STRATEGY_MIS ? ray_color_mis(ray, 0, k, Vec3<T>::one(), rc)
                                            : ray_color_brdf_only(ray, 0, k, Vec3<T>::one(), rc);
    }
};

// world.rs:315-332: mean -> sqrt gamma -> clamp -> *255 -> truncating cast (NaN casts to 0 in Rust)
inline uint8_t to_u8_gamma2(double mean) {
    double g = std::sqrt(mean);
    if (g != g) return 0;
    g = g < 0.0 ? 0.0 : (g > 1.0 ? 1.0 : g);
    return (uint8_t)(g * 255.0);
}

}  // namespace pto
