// pathtrace_host.hpp — C++17 host side above the C ABI (include/pathtrace_b200.h), mirroring the reference's scene / object
// API type for type so that a program written against roxas1533/pathtrace's `World` reads the same here.  The reference is
// Rust (no toolchain in this image; INTEGRATION.md carries the Rust binding); this is the compiled-language host the
// parity tests and the headless `pathtrace_host` program (host/main.cpp, the stand-in for src/main.rs) drive.
//
//   reference                                            here (namespace pathtrace)
//   Vector3 (src/math.rs:4-244, used subset)             Vector3: new_/zero/one, + - * / neg, dot, cross, length, normalize, luminance
//   Camera::new / look_at (src/camera.rs:50-130)         Camera::new_ / look_at; cached vectors cross the ABI as pt_camera
//   SphereShape::new, TriangleShape::new (shape.rs)      SphereShape::new_, TriangleShape::new_            : Shape
//   LambertianCosineWeighted::new, Emissive::new,        LambertianCosineWeighted::new_, Emissive::new_,
//   OrenNayar::new, Mirror{..} (material.rs, mirror.rs)  OrenNayar::new_, Mirror{roughness,color,metallic,ior} : Material
//   Object::new(Box<dyn Shape>, Box<dyn Material>)       Object::new_(unique_ptr<Shape>, unique_ptr<Material>)
//   World::new / render_pixel loop / draw /              World::new_ / render (ONE pt_render call replaces main.rs:43-60 +
//   export_luminance (src/world.rs)                      world.rs:293-333) / draw / export_luminance
//
// The one additive method the reference's traits need to cross an FFI boundary is describe(): it flattens a
// shape / material into the plain-data pt_shape / pt_material rows.  Scene construction stays in f64 on the host; values
// are narrowed to f32 once, in describe().  Errors: the reference panics (unwrap); here a failed pt_* call throws
// pathtrace::Error carrying pt_last_error().  There is no CPU fallback.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../include/pathtrace_b200.h"

namespace pathtrace {

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };

// ---------------------------------------------------------------------------------------------- src/math.rs
struct Vector3 {
    double x = 0, y = 0, z = 0;
    static Vector3 new_(double x, double y, double z) { return {x, y, z}; }
    static Vector3 zero() { return {0, 0, 0}; }
    static Vector3 one() { return {1, 1, 1}; }
    double dot(const Vector3& o) const { return x * o.x + y * o.y + z * o.z; }                                   // :24-26
    Vector3 cross(const Vector3& o) const { return {y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x}; }  // :29-35
    double length() const { return std::sqrt(x * x + y * y + z * z); }                                            // :38-40
    Vector3 normalize() const { double n = length(); return n > 0.0 ? Vector3{x / n, y / n, z / n} : *this; }     // :48-51
    double luminance() const { return 0.2126 * x + 0.7152 * y + 0.0722 * z; }                                     // :133-135
};
inline Vector3 operator+(Vector3 a, Vector3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vector3 operator-(Vector3 a, Vector3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vector3 operator-(Vector3 a) { return {-a.x, -a.y, -a.z}; }
inline Vector3 operator*(Vector3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vector3 operator*(double s, Vector3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline Vector3 operator*(Vector3 a, Vector3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline Vector3 operator/(Vector3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline Vector3 operator/(Vector3 a, Vector3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline bool operator==(Vector3 a, Vector3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

// -------------------------------------------------------------------------------------------- src/camera.rs
struct Ray {                                          // :3-25
    Vector3 origin, direction; double eta_ratio = 1.0;
    static Ray new_(Vector3 o, Vector3 d) { return {o, d.normalize(), 1.0}; }
    Vector3 at(double t) const { return origin + direction * t; }
    void set_eta_ratio(double e) { eta_ratio = e; }
};
struct Camera {                                       // :27-39
    Vector3 origin, lower_left_corner, horizontal, vertical;
    uint32_t width = 0, height = 0;
    // :50-82 — axis-aligned pinhole looking down -Z; the fov sets the VERTICAL extent (:61-62)
    static Camera new_(Vector3 origin, uint32_t width, uint32_t height, double screen_distance, double fov_degrees) {
        const double kPi = 3.14159265358979323846;
        double fov = fov_degrees * kPi / 180.0, aspect = (double)width / (double)height;
        double vh = 2.0 * std::tan(fov / 2.0) * screen_distance, vw = vh * aspect;
        Vector3 hz{vw, 0, 0}, vt{0, vh, 0};
        return {origin, origin - hz / 2.0 - vt / 2.0 - Vector3{0, 0, screen_distance}, hz, vt, width, height};
    }
    // :94-130
    static Camera look_at(Vector3 origin, Vector3 target, Vector3 up, uint32_t width, uint32_t height, double fov_degrees) {
        const double kPi = 3.14159265358979323846;
        double fov = fov_degrees * kPi / 180.0, aspect = (double)width / (double)height;
        Vector3 w = (origin - target).normalize(), u = up.cross(w).normalize(), v = w.cross(u);
        double vh = 2.0 * std::tan(fov / 2.0), vw = vh * aspect;
        Vector3 hz = u * vw, vt = v * vh;
        return {origin, origin - hz / 2.0 - vt / 2.0 - w, hz, vt, width, height};
    }
    // :139-147 (host-side twin of the device's camera_dir; used by tests and tools, never by the render path)
    Ray get_ray_with_offset(uint32_t x, uint32_t y, double ox, double oy) const {
        double u = ((double)x + ox) / (double)(width - 1), v = ((double)y + oy) / (double)(height - 1);
        return Ray::new_(origin, lower_left_corner + horizontal * u + vertical * v - origin);
    }
    pt_camera describe() const {
        pt_camera c;
        auto put = [](float* d, Vector3 s) { d[0] = (float)s.x; d[1] = (float)s.y; d[2] = (float)s.z; };
        put(c.origin, origin); put(c.lower_left_corner, lower_left_corner); put(c.horizontal, horizontal); put(c.vertical, vertical);
        c.width = width; c.height = height;
        return c;
    }
};

// ------------------------------------------------------------------------------------- src/objects/shape.rs
struct Shape {                                        // trait Shape :8-35 (+ describe)
    virtual ~Shape() = default;
    virtual pt_shape describe() const = 0;
};
struct SphereShape : Shape {                          // :38-51
    Vector3 center; double radius;
    SphereShape(Vector3 c, double r) : center(c), radius(r) {}
    static std::unique_ptr<Shape> new_(Vector3 center, double radius) { return std::make_unique<SphereShape>(center, radius); }
    pt_shape describe() const override {
        pt_shape s; std::memset(&s, 0, sizeof s);
        s.kind = PT_SHAPE_SPHERE; s.v[0] = (float)center.x; s.v[1] = (float)center.y; s.v[2] = (float)center.z; s.v[3] = (float)radius;
        return s;
    }
};
struct TriangleShape : Shape {                        // :148-158
    Vector3 v0, v1, v2;
    TriangleShape(Vector3 a, Vector3 b, Vector3 c) : v0(a), v1(b), v2(c) {}
    static std::unique_ptr<Shape> new_(Vector3 v0, Vector3 v1, Vector3 v2) { return std::make_unique<TriangleShape>(v0, v1, v2); }
    pt_shape describe() const override {
        pt_shape s; std::memset(&s, 0, sizeof s);
        s.kind = PT_SHAPE_TRIANGLE;
        const Vector3* v[3] = {&v0, &v1, &v2};
        for (int k = 0; k < 3; ++k) { s.v[3 * k] = (float)v[k]->x; s.v[3 * k + 1] = (float)v[k]->y; s.v[3 * k + 2] = (float)v[k]->z; }
        return s;
    }
};

// ------------------------------------------------------------ src/objects/material.rs, src/objects/mirror.rs
struct Material {                                     // trait Material material.rs:5-65 (+ describe)
    virtual ~Material() = default;
    virtual pt_material describe() const = 0;
    virtual Vector3 emit() const { return Vector3::zero(); }          // material.rs:62-64
    virtual double get_eta() const { return 1.0; }                    // material.rs:51-53
};
namespace detail {
inline pt_material mat(uint32_t kind, std::initializer_list<double> p) {
    pt_material m; std::memset(&m, 0, sizeof m);
    m.kind = kind; int i = 0;
    for (double v : p) m.p[i++] = (float)v;
    return m;
}
}  // namespace detail
struct LambertianCosineWeighted : Material {          // material.rs:67-77
    Vector3 albedo;
    explicit LambertianCosineWeighted(Vector3 a) : albedo(a) {}
    static std::unique_ptr<Material> new_(Vector3 albedo) { return std::make_unique<LambertianCosineWeighted>(albedo); }
    pt_material describe() const override { return detail::mat(PT_MAT_LAMBERTIAN, {albedo.x, albedo.y, albedo.z}); }
};
struct Emissive : Material {                          // material.rs:126-136, emit :160-162
    Vector3 emission;
    explicit Emissive(Vector3 e) : emission(e) {}
    static std::unique_ptr<Material> new_(Vector3 emission) { return std::make_unique<Emissive>(emission); }
    pt_material describe() const override { return detail::mat(PT_MAT_EMISSIVE, {emission.x, emission.y, emission.z}); }
    Vector3 emit() const override { return emission; }
};
struct OrenNayar : Material {                         // material.rs:166-196 — a, b precomputed in new() as the reference does
    Vector3 albedo; double roughness, a, b;
    OrenNayar(Vector3 al, double r) : albedo(al), roughness(r) {
        double s2 = r * r;
        a = 1.0 - 0.5 * s2 / (s2 + 0.33); b = 0.45 * s2 / (s2 + 0.09);
    }
    static std::unique_ptr<Material> new_(Vector3 albedo, double roughness) { return std::make_unique<OrenNayar>(albedo, roughness); }
    pt_material describe() const override { return detail::mat(PT_MAT_OREN_NAYAR, {albedo.x, albedo.y, albedo.z, roughness, a, b}); }
};
struct Mirror : Material {                            // mirror.rs:5-14 — struct literal with public fields; get_eta :317-319
    double roughness; Vector3 color; double metallic, ior;
    Mirror(double roughness_, Vector3 color_, double metallic_, double ior_) : roughness(roughness_), color(color_), metallic(metallic_), ior(ior_) {}
    pt_material describe() const override { return detail::mat(PT_MAT_MIRROR, {roughness, color.x, color.y, color.z, metallic, ior}); }
    double get_eta() const override { return ior; }
};

// ------------------------------------------------------------------------------------ src/objects/object.rs
struct Object {                                       // :9-25
    std::unique_ptr<Shape> shape; std::unique_ptr<Material> material;
    static Object new_(std::unique_ptr<Shape> s, std::unique_ptr<Material> m) { return Object{std::move(s), std::move(m)}; }
};
// Vec<Object> -> (pt_shape[n], pt_material[n]); object id = index, as in World.objects.
inline void describe_objects(const std::vector<Object>& objects, std::vector<pt_shape>& shapes, std::vector<pt_material>& mats) {
    shapes.clear(); mats.clear();
    for (const Object& o : objects) { shapes.push_back(o.shape->describe()); mats.push_back(o.material->describe()); }
}

// --------------------------------------------------------------------------------------------- src/world.rs
struct Color { uint8_t r = 0, g = 0, b = 0, a = 0; };                 // :20-26 (RGBA8)

enum class Strategy : uint32_t { Mis = PT_STRATEGY_MIS, BrdfOnly = PT_STRATEGY_BRDF_ONLY };   // cargo features mis | brdf_only, Cargo.toml:6-10

struct RenderOptions {                                 // the reference's compile-time constants as run-time fields
    uint32_t sample_num = 3000;                        // SAMPLE_NUM world.rs:18
    uint32_t sample_offset = 0;                        // first global sample index (progressive chunks / spp split)
    uint32_t min_depth = 4, max_depth = 50;            // rendering.rs:6-7
    Strategy strategy = Strategy::Mis;
    uint64_t seed = 0x5EED;
    int device = 0;
};

struct World {                                         // :54-62
    static constexpr uint32_t WIDTH = 400, HEIGHT = 400, SAMPLE_NUM = 3000;   // :16-18
    Camera camera;
    std::vector<Object> objects;
    std::vector<size_t> light_indices;
    std::vector<Color> data;                           // RGBA8, row-major, y down
    std::vector<Vector3> luminance_data;               // linear mean radiance
    pt_stats last_stats{};

    // World::new (:65-241): camera :67-73, 10 wall triangles, 2 light triangles, the glass sphere, light discovery :214-225
    static World new_(uint32_t width = WIDTH, uint32_t height = HEIGHT) {
        World w;
        w.camera = Camera::new_(Vector3::new_(0.0, 0.0, 2.0), width, height, 1.0, 35.0);
        const double bs = 1.0, bd = -2.0, ls = 0.3, zb = bd - bs, zf = bd + bs, ly = bs - 0.01;
        auto V = [](double x, double y, double z) { return Vector3::new_(x, y, z); };
        auto tri = [&](Vector3 a, Vector3 b, Vector3 c, std::unique_ptr<Material> m) { w.objects.push_back(Object::new_(TriangleShape::new_(a, b, c), std::move(m))); };
        auto lam = [&](double r, double g, double b) { return LambertianCosineWeighted::new_(V(r, g, b)); };
        tri(V(-bs, -bs, zb), V(-bs, bs, zb), V(-bs, bs, zf), lam(0.8, 0.1, 0.1));     // 0 left wall (red)
        tri(V(-bs, -bs, zb), V(-bs, bs, zf), V(-bs, -bs, zf), lam(0.8, 0.1, 0.1));    // 1
        tri(V(bs, -bs, zb), V(bs, bs, zf), V(bs, bs, zb), lam(0.1, 0.8, 0.1));        // 2 right wall (green)
        tri(V(bs, -bs, zb), V(bs, -bs, zf), V(bs, bs, zf), lam(0.1, 0.8, 0.1));       // 3
        tri(V(-bs, -bs, zb), V(bs, -bs, zb), V(bs, bs, zb), lam(0.2, 0.2, 0.8));      // 4 back wall
        tri(V(-bs, -bs, zb), V(bs, bs, zb), V(-bs, bs, zb), lam(0.2, 0.2, 0.8));      // 5
        tri(V(-bs, -bs, zb), V(bs, -bs, zf), V(bs, -bs, zb), lam(0.2, 0.8, 0.8));     // 6 floor
        tri(V(-bs, -bs, zb), V(-bs, -bs, zf), V(bs, -bs, zf), lam(0.2, 0.8, 0.8));    // 7
        tri(V(-bs, bs, zb), V(bs, bs, zb), V(bs, bs, zf), lam(0.8, 0.8, 0.8));        // 8 ceiling
        tri(V(-bs, bs, zb), V(bs, bs, zf), V(-bs, bs, zf), lam(0.8, 0.8, 0.8));       // 9
        tri(V(-ls, ly, bd - ls), V(ls, ly, bd - ls), V(ls, ly, bd + ls), Emissive::new_(V(15.0, 15.0, 15.0)));    // 10 light
        tri(V(-ls, ly, bd - ls), V(ls, ly, bd + ls), V(-ls, ly, bd + ls), Emissive::new_(V(15.0, 15.0, 15.0)));   // 11 light
        w.objects.push_back(Object::new_(SphereShape::new_(V(0.4, -0.6, bd), 0.4), std::make_unique<Mirror>(0.3, V(1.0, 1.0, 1.0), 0.0, 1.5)));   // 12 glass
        w.find_lights();
        w.data.assign((size_t)width * height, Color{});
        w.luminance_data.assign((size_t)width * height, Vector3::zero());
        return w;
    }
    // :214-225 — lights = objects whose emission has non-zero length (call again after editing `objects`)
    void find_lights() {
        light_indices.clear();
        for (size_t i = 0; i < objects.size(); ++i) if (objects[i].material->emit().length() > 0.0) light_indices.push_back(i);
    }

    // Replaces the render thread's par_iter().for_each(render_pixel) (main.rs:43-60, world.rs:293-333): pt_scene_create once,
    // ONE pt_render for the frame, results copied into `data` / `luminance_data` exactly where render_pixel stores them.
    void render(const RenderOptions& opt = RenderOptions{}) {
        pt_context* ctx = nullptr;
        if (int rc = pt_context_create(opt.device, &ctx)) throw Error(std::string("pt_context_create: ") + pt_last_error(nullptr) + " (status " + std::to_string(rc) + ")");
        pt_scene* scene = nullptr;
        std::string err;
        try {
            std::vector<pt_shape> shapes; std::vector<pt_material> mats;
            describe_objects(objects, shapes, mats);
            if (pt_scene_create(ctx, shapes.data(), mats.data(), (uint32_t)shapes.size(), &scene)) throw Error(std::string("pt_scene_create: ") + pt_last_error(ctx));
            if (pt_scene_num_lights(scene) != light_indices.size()) throw Error("light discovery differs between host and library");
            pt_camera cam = camera.describe();
            pt_render_params p; std::memset(&p, 0, sizeof p);
            p.spp = opt.sample_num; p.sample_offset = opt.sample_offset; p.min_depth = opt.min_depth; p.max_depth = opt.max_depth;
            p.n_light_samples = 1;                                                                                  // NUM_LIGHT_SAMPLES, rendering.rs:10
            p.strategy = (uint32_t)opt.strategy; p.traversal = PT_TRAVERSAL_AUTO; p.seed = opt.seed; p.t_min = 0.001f;
            const size_t npix = (size_t)cam.width * cam.height;
            std::vector<float> rad(npix * 3); std::vector<uint8_t> rgba(npix * 4);
            if (pt_render(ctx, scene, &cam, &p, rad.data(), rgba.data(), &last_stats)) throw Error(std::string("pt_render: ") + pt_last_error(ctx));
            data.resize(npix); luminance_data.resize(npix);
            for (size_t i = 0; i < npix; ++i) {
                data[i] = Color{rgba[4 * i], rgba[4 * i + 1], rgba[4 * i + 2], rgba[4 * i + 3]};                 // main.rs:58-59
                luminance_data[i] = Vector3{rad[3 * i], rad[3 * i + 1], rad[3 * i + 2]};                           // world.rs:318-319
            }
        } catch (...) {
            if (scene) pt_scene_destroy(scene);
            pt_context_destroy(ctx);
            throw;
        }
        pt_scene_destroy(scene);
        pt_context_destroy(ctx);
    }

    // world.rs:335-341 — copies the RGBA8 framebuffer into `frame` (4 bytes per pixel)
    void draw(uint8_t* frame) const { std::memcpy(frame, data.data(), data.size() * 4); }

    // world.rs:344-369 — "x,y,r,g,b,luminance", one row per pixel, row-major, {:.6}
    void export_luminance(const std::string& path) const {
        std::vector<float> rad(luminance_data.size() * 3);
        for (size_t i = 0; i < luminance_data.size(); ++i) { rad[3 * i] = (float)luminance_data[i].x; rad[3 * i + 1] = (float)luminance_data[i].y; rad[3 * i + 2] = (float)luminance_data[i].z; }
        if (pt_export_luminance(path.c_str(), rad.data(), camera.width, camera.height)) throw Error("export_luminance: cannot write " + path);
    }
    // headless stand-in for the window: binary PPM of `data`
    void write_ppm(const std::string& path) const {
        if (pt_write_ppm(path.c_str(), reinterpret_cast<const uint8_t*>(data.data()), camera.width, camera.height)) throw Error("write_ppm: cannot write " + path);
    }
};

}  // namespace pathtrace
