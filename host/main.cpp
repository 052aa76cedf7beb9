// pathtrace_host — headless stand-in for the reference's src/main.rs: builds World::new(), renders it through the C ABI
// (one pt_render call where main.rs:43-60 runs its rayon pixel loop), exports luminance.csv (main.rs:63-67) and writes the
// RGBA8 framebuffer as a PPM where main.rs blits it to a window (graphics stack: unavailable on this box).
//
//   pathtrace_host [--width W] [--height H] [--spp N] [--strategy mis|brdf_only] [--seed S] [--device D] [--out DIR]
//   pathtrace_host --describe          prints the flattened pt_shape / pt_material / pt_camera rows of World::new() (no GPU needed)
#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "pathtrace_host.hpp"

using namespace pathtrace;

int main(int argc, char** argv) {
    uint32_t width = World::WIDTH, height = World::HEIGHT;
    RenderOptions opt;
    std::string out = ".";
    bool describe = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "missing value after %s\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--width") width = (uint32_t)std::strtoul(next(), nullptr, 0);
        else if (a == "--height") height = (uint32_t)std::strtoul(next(), nullptr, 0);
        else if (a == "--spp") opt.sample_num = (uint32_t)std::strtoul(next(), nullptr, 0);
        else if (a == "--seed") opt.seed = std::strtoull(next(), nullptr, 0);
        else if (a == "--device") opt.device = std::atoi(next());
        else if (a == "--out") out = next();
        else if (a == "--strategy") { std::string s = next(); opt.strategy = s == "brdf_only" ? Strategy::BrdfOnly : Strategy::Mis; }
        else if (a == "--describe") describe = true;
        else { std::fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
    }
    World world = World::new_(width, height);
    if (describe) {
        std::vector<pt_shape> shapes; std::vector<pt_material> mats;
        describe_objects(world.objects, shapes, mats);
        for (size_t i = 0; i < shapes.size(); ++i) {
            std::printf("shape %zu %u", i, shapes[i].kind);
            for (float v : shapes[i].v) { uint32_t b; std::memcpy(&b, &v, 4); std::printf(" %08x", b); }
            std::printf("\nmaterial %zu %u", i, mats[i].kind);
            for (float v : mats[i].p) { uint32_t b; std::memcpy(&b, &v, 4); std::printf(" %08x", b); }
            std::printf("\n");
        }
        pt_camera c = world.camera.describe();
        std::printf("camera %u %u", c.width, c.height);
        for (const float* p : {c.origin, c.lower_left_corner, c.horizontal, c.vertical})
            for (int k = 0; k < 3; ++k) { uint32_t b; std::memcpy(&b, &p[k], 4); std::printf(" %08x", b); }
        std::printf("\nlights");
        for (size_t l : world.light_indices) std::printf(" %zu", l);
        std::printf("\n");
        return 0;
    }
    try {
        world.render(opt);
        world.export_luminance(out + "/luminance.csv");                      // main.rs:63-67
        world.write_ppm(out + "/frame.ppm");
    } catch (const Error& e) {
        std::fprintf(stderr, "pathtrace_host: %s\n", e.what());
        return 1;
    }
    const pt_stats& st = world.last_stats;
    double rays = (double)(st.rays_primary + st.rays_shadow + st.rays_scattered);
    std::printf("Rendering completed! %ux%u x %u spp: %.1f ms on the device, %.1f Msamples/s, %.1f Mrays/s\n", width, height, opt.sample_num,
                st.ms_render, (double)st.samples / st.ms_render / 1e3, rays / st.ms_render / 1e3);
    std::printf("Luminance data exported to %s/luminance.csv\n", out.c_str());
    return 0;
}
