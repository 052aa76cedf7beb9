// capi.cu — extern "C" entry points of libpathtrace_b200.so (see include/pathtrace_b200.h).
// Host-side glue only: argument checks, uploads, launches, downloads.  No CPU fallback of any kind.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "pt_internal.h"

namespace {
thread_local std::string g_last_error;

int fail(pt_context* ctx, int code, const std::string& msg) {
    if (ctx) ctx->err = msg;
    g_last_error = msg;
    return code;
}
int keep(pt_context* ctx, int rc) {
    if (rc && ctx) g_last_error = ctx->err;
    return rc;
}
// NUM_LIGHT_SAMPLES is the compile-time constant 1 in the reference (rendering.rs:10); min_depth must fit the 8-bit depth of
// the path state (pt_wave.cuh: pack_meta); the global sample index is packed into 24 bits of it.
bool valid_params(const pt_render_params* p) {
    return p && p->strategy <= PT_STRATEGY_BRDF_ONLY && p->traversal < PT_TRAVERSAL_COUNT && std::isfinite(p->t_min) && p->t_min >= 0.0f &&
           p->n_light_samples <= 1 && p->min_depth <= 254 && (uint64_t)p->sample_offset + p->spp <= (1ull << 24);
}
const char* kBadParams = "bad render params (strategy / traversal out of range, t_min not finite or negative, n_light_samples > 1, "
                         "min_depth > 254, or sample_offset + spp > 2^24)";
bool valid_camera(const pt_camera* c) { return c && c->width >= 2 && c->height >= 2 && (uint64_t)c->width * c->height < (1ull << 31); }

// Rust's `{:.6}` (world.rs:361-365): fixed notation with 6 decimals; non-finite values print as NaN / inf / -inf.
void fmt6(char* out, size_t n, double v) {
    if (v != v) std::snprintf(out, n, "NaN");
    else if (std::isinf(v)) std::snprintf(out, n, v > 0 ? "inf" : "-inf");
    else std::snprintf(out, n, "%.6f", v);
}
}  // namespace

int ensure_frame(pt_context* ctx, size_t npix) {
    if (npix > ctx->accum_px) {
        if (ctx->d_accum) cudaFree(ctx->d_accum);
        ctx->d_accum = nullptr; ctx->accum_px = 0;
        PT_CUDA_CHECK(ctx, cudaMalloc(&ctx->d_accum, sizeof(float4) * npix));
        ctx->accum_px = npix;
    }
    if (npix > ctx->out_px) {
        if (ctx->d_radiance) cudaFree(ctx->d_radiance);
        if (ctx->d_rgba) cudaFree(ctx->d_rgba);
        ctx->d_radiance = nullptr; ctx->d_rgba = nullptr; ctx->out_px = 0;
        PT_CUDA_CHECK(ctx, cudaMalloc(&ctx->d_radiance, sizeof(float) * 3 * npix));
        PT_CUDA_CHECK(ctx, cudaMalloc(&ctx->d_rgba, 4 * npix));
        ctx->out_px = npix;
    }
    return PT_OK;
}
// Planar entries for the secondary-ray scan.  An entry is a parallelogram (two consecutive triangles A = (a, s0, s1) and
// B = (b, s0, s1) with a + b == s0 + s1) or a single triangle, stored as the rows of the affine map world -> (u, v, w) with
// P = a + u e_u + v e_v + w n (a precomputed-transform test in the manner of Baldwin & Weber 2016): the ray meets the plane
// at t = -o'_w / d'_w and hits when (u, v) lies in the unit square (parallelogram) or under u + v <= 1 (triangle).
void build_fast_table(const pt_shape* shapes, uint32_t n, SceneDev& out) {
    struct Entry { double m[12]; int id_a, id_b; float lim; };
    std::vector<Entry> entries;
    auto make = [](const double a[3], const double eu[3], const double ev[3], Entry& e) {
        double nn[3] = {eu[1] * ev[2] - eu[2] * ev[1], eu[2] * ev[0] - eu[0] * ev[2], eu[0] * ev[1] - eu[1] * ev[0]};
        const double len = std::sqrt(nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
        if (!(len > 0.0) || !std::isfinite(len)) return false;
        for (double& c : nn) c /= len;
        // inverse of the matrix with columns (eu, ev, nn): rows are cross products over the determinant
        const double det = eu[0] * (ev[1] * nn[2] - ev[2] * nn[1]) - eu[1] * (ev[0] * nn[2] - ev[2] * nn[0]) + eu[2] * (ev[0] * nn[1] - ev[1] * nn[0]);
        if (!(std::fabs(det) > 0.0)) return false;
        const double r0[3] = {(ev[1] * nn[2] - ev[2] * nn[1]) / det, (ev[2] * nn[0] - ev[0] * nn[2]) / det, (ev[0] * nn[1] - ev[1] * nn[0]) / det};
        const double r1[3] = {(nn[1] * eu[2] - nn[2] * eu[1]) / det, (nn[2] * eu[0] - nn[0] * eu[2]) / det, (nn[0] * eu[1] - nn[1] * eu[0]) / det};
        const double r2[3] = {(eu[1] * ev[2] - eu[2] * ev[1]) / det, (eu[2] * ev[0] - eu[0] * ev[2]) / det, (eu[0] * ev[1] - eu[1] * ev[0]) / det};
        const double* rows[3] = {r0, r1, r2};
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) e.m[4 * r + c] = rows[r][c];
            e.m[4 * r + 3] = -(rows[r][0] * a[0] + rows[r][1] * a[1] + rows[r][2] * a[2]);
        }
        for (double v : e.m) if (!std::isfinite(v) || std::fabs(v) > 1e18) return false;
        return true;
    };
    out.fast_host.clear(); out.n_fast_pairs = 0; out.n_fast_spheres = 0;
    bool ok = true;
    for (uint32_t i = 0; i < n && ok; ++i) {
        if (shapes[i].kind != PT_SHAPE_TRIANGLE) continue;
        double A[3][3];
        for (int v = 0; v < 3; ++v) for (int c = 0; c < 3; ++c) A[v][c] = shapes[i].v[3 * v + c];
        Entry e;
        bool quad = false;
        if (i + 1 < n && shapes[i + 1].kind == PT_SHAPE_TRIANGLE) {
            double B[3][3];
            for (int v = 0; v < 3; ++v) for (int c = 0; c < 3; ++c) B[v][c] = shapes[i + 1].v[3 * v + c];
            int sa[2] = {0, 0}, sb[2] = {0, 0}, ns = 0;
            for (int va = 0; va < 3; ++va)
                for (int vb = 0; vb < 3; ++vb)
                    if (A[va][0] == B[vb][0] && A[va][1] == B[vb][1] && A[va][2] == B[vb][2]) { if (ns < 2) { sa[ns] = va; sb[ns] = vb; } ++ns; }
            if (ns == 2 && sa[0] != sa[1] && sb[0] != sb[1]) {   // exactly one shared edge: the diagonal
                const int ia = 3 - sa[0] - sa[1], ib = 3 - sb[0] - sb[1];
                const double *a = A[ia], *b = B[ib], *s0 = A[sa[0]], *s1 = A[sa[1]];
                double scale = 0, err = 0;
                for (int c = 0; c < 3; ++c) { err = std::fmax(err, std::fabs(a[c] + b[c] - s0[c] - s1[c])); scale = std::fmax(scale, std::fmax(std::fabs(s0[c] - a[c]), std::fabs(s1[c] - a[c]))); }
                if (err <= 1e-6 * scale) {
                    double eu[3], ev[3];
                    for (int c = 0; c < 3; ++c) { eu[c] = s0[c] - a[c]; ev[c] = s1[c] - a[c]; }
                    if (make(a, eu, ev, e)) { e.id_a = (int)i; e.id_b = (int)i + 1; e.lim = 2.0f; quad = true; }
                }
            }
        }
        if (quad) { entries.push_back(e); ++i; continue; }
        double eu[3], ev[3];
        for (int c = 0; c < 3; ++c) { eu[c] = A[1][c] - A[0][c]; ev[c] = A[2][c] - A[0][c]; }
        if (!make(A[0], eu, ev, e)) {           // degenerate triangle: the exact test rejects it too (|a| < 1e-8); keep an entry that never hits
            for (double& v : e.m) v = 0.0;
            e.m[11] = 1.0;
        }
        e.id_a = e.id_b = (int)i; e.lim = 1.0f;
        entries.push_back(e);
    }
    if (entries.empty() || entries.size() > 2u * kFastParamPairs) return;
    uint32_t n_sph = 0;
    for (uint32_t i = 0; i < n; ++i) n_sph += shapes[i].kind == PT_SHAPE_SPHERE;
    if (n_sph > (uint32_t)kFastParamSpheres) return;
    for (uint32_t i = 0; i < n; ++i)
        if (shapes[i].kind == PT_SHAPE_SPHERE) {
            out.fast_spheres[out.n_fast_spheres] = make_float4(shapes[i].v[0], shapes[i].v[1], shapes[i].v[2], shapes[i].v[3]);
            out.fast_sphere_ids[out.n_fast_spheres++] = (int)i;
        }
    // (u, v) are stored centred: the unit square becomes [-1/2, 1/2]^2 (one |.| <= 1/2 test), u + v <= lim becomes u' + v' <= lim - 1
    for (Entry& e : entries) { e.m[3] -= 0.5; e.m[7] -= 0.5; e.lim -= 1.0f; }
    if (entries.size() & 1u) { Entry d; for (double& v : d.m) v = 0.0; d.m[11] = 1.0; d.id_a = d.id_b = -1; d.lim = 0.0f; entries.push_back(d); }   // w == 1 everywhere: never hit
    out.n_fast_pairs = (uint32_t)entries.size() / 2;
    out.fast_host.resize(8 * (size_t)out.n_fast_pairs);
    for (uint32_t p = 0; p < out.n_fast_pairs; ++p) {
        const Entry &x = entries[2 * p], &y = entries[2 * p + 1];
        float4* E = out.fast_host.data() + 8 * (size_t)p;
        for (int r = 0; r < 3; ++r) {
            E[2 * r] = make_float4((float)x.m[4 * r], (float)y.m[4 * r], (float)x.m[4 * r + 1], (float)y.m[4 * r + 1]);
            E[2 * r + 1] = make_float4((float)x.m[4 * r + 2], (float)y.m[4 * r + 2], (float)x.m[4 * r + 3], (float)y.m[4 * r + 3]);
        }
        int ia[4] = {x.id_a, y.id_a, x.id_b, y.id_b};
        std::memcpy(&E[6], ia, 16);
        E[7] = make_float4(x.lim, y.lim, 0.f, 0.f);
    }
}

extern "C" {

const char* pt_version(void) { return "pathtrace_b200 0.2 (sm_100a)"; }

const char* pt_last_error(const pt_context* ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

int32_t pt_context_create(int32_t device_id, pt_context** out) {
    if (!out) return fail(nullptr, PT_ERR_INVALID, "pt_context_create: out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return fail(nullptr, PT_ERR_NO_DEVICE, std::string("no CUDA device: ") + cudaGetErrorString(e) + " (there is no CPU fallback)");
    if (device_id < 0 || device_id >= count) return fail(nullptr, PT_ERR_INVALID, "pt_context_create: device_id out of range");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_id) != cudaSuccess) return fail(nullptr, PT_ERR_NO_DEVICE, "cudaGetDeviceProperties failed");
    if (prop.major != 10) return fail(nullptr, PT_ERR_NO_DEVICE, std::string("device is sm_") + std::to_string(prop.major) + std::to_string(prop.minor) + "; this library is built for sm_100a only");
    pt_context* ctx = new (std::nothrow) pt_context();
    if (!ctx) return fail(nullptr, PT_ERR_OOM, "out of host memory");
    ctx->device = device_id;
    ctx->sm_count = prop.multiProcessorCount;
    if (cudaSetDevice(device_id) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return fail(nullptr, PT_ERR_CUDA, "cannot create stream");
    }
    for (auto& ev : ctx->ev) cudaEventCreate(&ev);
    *out = ctx;
    return PT_OK;
}

void pt_context_destroy(pt_context* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->pool_mem) cudaFree(ctx->pool_mem);
    if (ctx->d_accum) cudaFree(ctx->d_accum);
    if (ctx->d_radiance) cudaFree(ctx->d_radiance);
    if (ctx->d_rgba) cudaFree(ctx->d_rgba);
    if (ctx->d_counters) cudaFree(ctx->d_counters);
    if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
    for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->prof_events) cudaEventDestroy(ev);
    if (ctx->ev_shaded) cudaEventDestroy(ctx->ev_shaded);
    if (ctx->ev_shadowed) cudaEventDestroy(ctx->ev_shadowed);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

void* pt_context_stream(pt_context* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int32_t pt_context_synchronize(pt_context* ctx) {
    if (!ctx) return PT_ERR_INVALID;
    PT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return PT_OK;
}

int32_t pt_scene_create(pt_context* ctx, const pt_shape* shapes, const pt_material* materials, uint32_t n, pt_scene** out) {
    if (!ctx || !out || (n && (!shapes || !materials))) return fail(ctx, PT_ERR_INVALID, "pt_scene_create: null argument");
    *out = nullptr;
    if (n >= (1u << 30)) return fail(ctx, PT_ERR_INVALID, "pt_scene_create: too many objects");
    for (uint32_t i = 0; i < n; ++i) {
        if (shapes[i].kind > PT_SHAPE_TRIANGLE) return fail(ctx, PT_ERR_INVALID, "pt_scene_create: unknown shape kind at object " + std::to_string(i));
        if (materials[i].kind > PT_MAT_OREN_NAYAR) return fail(ctx, PT_ERR_INVALID, "pt_scene_create: unknown material kind at object " + std::to_string(i));
    }
    cudaSetDevice(ctx->device);
    pt_scene* sc = new (std::nothrow) pt_scene();
    if (!sc) return fail(ctx, PT_ERR_OOM, "out of host memory");
    sc->ctx = ctx;
    sc->dev.n_objects = n;
    // world.rs:214-225: lights = objects whose emission has non-zero length
    std::vector<uint32_t> lights;
    for (uint32_t i = 0; i < n; ++i)
        if (materials[i].kind == PT_MAT_EMISSIVE) {
            const float* e = materials[i].p;
            if (std::sqrt((double)e[0] * e[0] + (double)e[1] * e[1] + (double)e[2] * e[2]) > 0.0) lights.push_back(i);
        }
    sc->dev.n_lights = (uint32_t)lights.size();
    sc->dev.fp32_domain = true;
    for (uint32_t i = 0; i < n && sc->dev.fp32_domain; ++i)
        for (int k = 0; k < (shapes[i].kind == PT_SHAPE_TRIANGLE ? 9 : 4); ++k)
            if (!(std::fabs(shapes[i].v[k]) <= kFp32DomainMax)) { sc->dev.fp32_domain = false; break; }
    if (n && n <= 1024) {   // brute-force scan table: consecutive triangles share an entry (lbvh.cu: k_brute_table walks the same way)
        uint32_t e = 0, pairs = 0;
        for (uint32_t i = 0; i < n; ++e) {
            bool pair = i + 1 < n && shapes[i].kind == PT_SHAPE_TRIANGLE && shapes[i + 1].kind == PT_SHAPE_TRIANGLE;
            pairs += pair; i += pair ? 2 : 1;
        }
        sc->dev.n_brute = e; sc->dev.n_pairs = pairs;
        sc->dev.shapes_host.assign(shapes, shapes + n);
    }
    if (sc->dev.n_brute && sc->dev.fp32_domain) build_fast_table(shapes, n, sc->dev);
    pt_shape* d_shapes = nullptr;
    pt_material* d_mats = nullptr;
    auto cleanup = [&](int rc) { if (d_shapes) cudaFree(d_shapes); if (d_mats) cudaFree(d_mats); if (rc) { pt_scene_destroy(sc); } return keep(ctx, rc); };
    size_t n1 = n ? n : 1;
    cudaError_t e;
    if ((e = cudaMalloc(&d_shapes, sizeof(pt_shape) * n1)) != cudaSuccess || (e = cudaMalloc(&d_mats, sizeof(pt_material) * n1)) != cudaSuccess ||
        (e = cudaMalloc(&sc->dev.light_ids, 4 * (lights.size() ? lights.size() : 1))) != cudaSuccess) {
        ctx->err = std::string("pt_scene_create: ") + cudaGetErrorString(e);
        return cleanup(e == cudaErrorMemoryAllocation ? PT_ERR_OOM : PT_ERR_CUDA);
    }
    if (n) {
        cudaMemcpyAsync(d_shapes, shapes, sizeof(pt_shape) * n, cudaMemcpyHostToDevice, ctx->stream);
        cudaMemcpyAsync(d_mats, materials, sizeof(pt_material) * n, cudaMemcpyHostToDevice, ctx->stream);
    }
    if (!lights.empty()) cudaMemcpyAsync(sc->dev.light_ids, lights.data(), 4 * lights.size(), cudaMemcpyHostToDevice, ctx->stream);
    int rc = scene_pack_launch(ctx, d_shapes, d_mats, n, sc->dev);
    if (rc) return cleanup(rc);
    uint64_t launches = 0;
    rc = lbvh_build_device(ctx, d_shapes, n, sc->dev, &launches);   // records ev[2] / ev[3] around its kernels and synchronises
    if (rc) return cleanup(rc);
    cudaEventElapsedTime(&sc->build_ms, ctx->ev[2], ctx->ev[3]);
    sc->build_launches = (uint32_t)launches;
    cleanup(0);
    *out = sc;
    return PT_OK;
}

void pt_scene_destroy(pt_scene* sc) {
    if (!sc) return;
    if (sc->ctx) cudaSetDevice(sc->ctx->device);
    void* ptrs[] = {sc->dev.prims, sc->dev.mats, sc->dev.light_ids, sc->dev.nodes, sc->dev.codes_sorted, sc->dev.prim_sorted, sc->dev.brute};
    for (void* p : ptrs) if (p) cudaFree(p);
    delete sc;
}

uint32_t pt_scene_num_objects(const pt_scene* s) { return s ? s->dev.n_objects : 0; }
uint32_t pt_scene_num_lights(const pt_scene* s) { return s ? s->dev.n_lights : 0; }
float pt_scene_build_ms(const pt_scene* s) { return s ? s->build_ms : 0.f; }
uint32_t pt_scene_build_launches(const pt_scene* s) { return s ? s->build_launches : 0; }

int32_t pt_render_accumulate_device(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, void* d_accum,
                                    pt_stats* stats) {
    if (!ctx || !scene || !d_accum) return fail(ctx, PT_ERR_INVALID, "pt_render_accumulate_device: null argument");
    if (!valid_camera(cam)) return fail(ctx, PT_ERR_INVALID, "pt_render: bad camera (width/height must be >= 2)");
    if (!valid_params(p)) return fail(ctx, PT_ERR_INVALID, std::string("pt_render: ") + kBadParams);
    if (scene->ctx != ctx) return fail(ctx, PT_ERR_INVALID, "pt_render: scene belongs to another context");
    cudaSetDevice(ctx->device);
    if (stats) std::memset(stats, 0, sizeof(*stats));
    RenderLaunch rl{&scene->dev, *cam, *p, static_cast<float4*>(d_accum)};
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
    int rc = wavefront_render(ctx, rl, stats);
    if (rc) return keep(ctx, rc);
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
    PT_CUDA_CHECK(ctx, cudaEventSynchronize(ctx->ev[1]));
    if (stats) { cudaEventElapsedTime(&stats->ms_render, ctx->ev[0], ctx->ev[1]); stats->ms_total = stats->ms_render; }
    return PT_OK;
}

int32_t pt_resolve_device(pt_context* ctx, const void* d_accum, uint32_t w, uint32_t h, uint32_t total_spp, void* d_rad, void* d_rgba) {
    if (!ctx || !d_accum || !total_spp || !w || !h) return fail(ctx, PT_ERR_INVALID, "pt_resolve_device: bad argument");
    cudaSetDevice(ctx->device);
    return keep(ctx, resolve_launch(ctx, static_cast<const float4*>(d_accum), w, h, total_spp, static_cast<float*>(d_rad), static_cast<uint8_t*>(d_rgba)));
}

int32_t pt_render(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, float* radiance, uint8_t* rgba,
                  pt_stats* stats) {
    if (!ctx || !scene) return fail(ctx, PT_ERR_INVALID, "pt_render: null argument");
    if (!valid_camera(cam)) return fail(ctx, PT_ERR_INVALID, "pt_render: bad camera (width/height must be >= 2)");
    if (!valid_params(p) || p->spp == 0) return fail(ctx, PT_ERR_INVALID, std::string("pt_render: spp must be > 0; ") + kBadParams);
    cudaSetDevice(ctx->device);
    size_t npix = (size_t)cam->width * cam->height;
    int rc = ensure_frame(ctx, npix);
    if (rc) return keep(ctx, rc);
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
    PT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_accum, 0, sizeof(float4) * npix, ctx->stream));
    pt_stats local;
    rc = pt_render_accumulate_device(ctx, scene, cam, p, ctx->d_accum, &local);
    if (rc) return rc;
    rc = resolve_launch(ctx, ctx->d_accum, cam->width, cam->height, p->spp, radiance ? ctx->d_radiance : nullptr, rgba ? ctx->d_rgba : nullptr);
    if (rc) return keep(ctx, rc);
    local.kernel_launches += 1;
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
    if (radiance) PT_CUDA_CHECK(ctx, cudaMemcpyAsync(radiance, ctx->d_radiance, sizeof(float) * 3 * npix, cudaMemcpyDeviceToHost, ctx->stream));
    if (rgba) PT_CUDA_CHECK(ctx, cudaMemcpyAsync(rgba, ctx->d_rgba, 4 * npix, cudaMemcpyDeviceToHost, ctx->stream));
    PT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
    PT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    if (stats) {
        *stats = local;
        cudaEventElapsedTime(&stats->ms_total, ctx->ev[2], ctx->ev[1]);
        cudaEventElapsedTime(&stats->ms_download, ctx->ev[3], ctx->ev[1]);
    }
    return PT_OK;
}

// No counterpart in the reference: see include/pathtrace_b200.h and the
// scratch array in smem_wavefront.cu.
int32_t pt_render_device_noalloc(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, void* d_radiance,
                                 void* d_rgba) {
    if (!ctx || !scene) return fail(ctx, PT_ERR_INVALID, "pt_render_device_noalloc: null argument");
    if (!d_radiance && !d_rgba) return fail(ctx, PT_ERR_INVALID, "pt_render_device_noalloc: both output pointers are null");
    if (!valid_camera(cam)) return fail(ctx, PT_ERR_INVALID, "pt_render_device_noalloc: bad camera (width/height must be >= 2)");
    if (!valid_params(p) || p->spp == 0) return fail(ctx, PT_ERR_INVALID, std::string("pt_render_device_noalloc: spp must be > 0; ") + kBadParams);
    if (scene->ctx != ctx) return fail(ctx, PT_ERR_INVALID, "pt_render_device_noalloc: scene belongs to another context");
    cudaSetDevice(ctx->device);
    return keep(ctx, smem_wavefront_render_noalloc(ctx, scene->dev, *cam, *p, static_cast<float*>(d_radiance), static_cast<uint8_t*>(d_rgba)));
}

int32_t pt_render_moments(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, uint32_t batches,
                          float* radiance, float* moment2, pt_stats* stats) {
    if (!ctx || !scene) return fail(ctx, PT_ERR_INVALID, "pt_render_moments: null argument");
    if (!valid_camera(cam)) return fail(ctx, PT_ERR_INVALID, "pt_render_moments: bad camera (width/height must be >= 2)");
    if (!valid_params(p) || p->spp == 0) return fail(ctx, PT_ERR_INVALID, std::string("pt_render_moments: spp must be > 0; ") + kBadParams);
    if (batches == 0 || p->spp % batches != 0) return fail(ctx, PT_ERR_INVALID, "pt_render_moments: batches must divide spp");
    cudaSetDevice(ctx->device);
    const size_t npix = (size_t)cam->width * cam->height;
    int rc = ensure_frame(ctx, npix);
    if (rc) return keep(ctx, rc);
    DevTemps tmp;
    float4 *d_batch = nullptr, *d_m2 = nullptr;
    float* d_m2rgb = nullptr;
    PT_CUDA_CHECK(ctx, tmp.alloc(&d_batch, sizeof(float4) * npix));
    PT_CUDA_CHECK(ctx, tmp.alloc(&d_m2, sizeof(float4) * npix));
    PT_CUDA_CHECK(ctx, tmp.alloc(&d_m2rgb, sizeof(float) * 3 * npix));
    PT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_accum, 0, sizeof(float4) * npix, ctx->stream));
    PT_CUDA_CHECK(ctx, cudaMemsetAsync(d_batch, 0, sizeof(float4) * npix, ctx->stream));
    PT_CUDA_CHECK(ctx, cudaMemsetAsync(d_m2, 0, sizeof(float4) * npix, ctx->stream));
    pt_stats total;
    std::memset(&total, 0, sizeof total);
    pt_render_params bp = *p;
    bp.spp = p->spp / batches;
    for (uint32_t b = 0; b < batches; ++b) {
        bp.sample_offset = p->sample_offset + b * bp.spp;
        pt_stats st;
        rc = pt_render_accumulate_device(ctx, scene, cam, &bp, d_batch, &st);
        if (rc) return rc;
        rc = fold_moments_launch(ctx, d_batch, ctx->d_accum, d_m2, (uint32_t)npix);
        if (rc) return keep(ctx, rc);
        total.samples += st.samples; total.rays_primary += st.rays_primary; total.rays_shadow += st.rays_shadow;
        total.rays_scattered += st.rays_scattered; total.path_vertices += st.path_vertices; total.wavefront_iters += st.wavefront_iters;
        total.kernel_launches += st.kernel_launches + 1; total.ms_render += st.ms_render; total.path_pool = st.path_pool; total.scan_mode = st.scan_mode;
    }
    rc = resolve_launch(ctx, ctx->d_accum, cam->width, cam->height, p->spp, ctx->d_radiance, nullptr);
    if (rc) return keep(ctx, rc);
    rc = moments_out_launch(ctx, d_m2, (uint32_t)npix, d_m2rgb);
    if (rc) return keep(ctx, rc);
    total.kernel_launches += 2;
    if (radiance) PT_CUDA_CHECK(ctx, cudaMemcpyAsync(radiance, ctx->d_radiance, sizeof(float) * 3 * npix, cudaMemcpyDeviceToHost, ctx->stream));
    if (moment2) PT_CUDA_CHECK(ctx, cudaMemcpyAsync(moment2, d_m2rgb, sizeof(float) * 3 * npix, cudaMemcpyDeviceToHost, ctx->stream));
    PT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    total.ms_total = total.ms_render;
    if (stats) *stats = total;
    return PT_OK;
}

int32_t pt_trace_primary(pt_context* ctx, const pt_scene* scene, const pt_camera* cam, const pt_render_params* p, int32_t jitter, float ox,
                         float oy, uint32_t sample, int32_t* obj_id, float* t, float* dir) {
    if (!ctx || !scene || !obj_id || !t) return fail(ctx, PT_ERR_INVALID, "pt_trace_primary: null argument");
    if (!valid_camera(cam) || !valid_params(p)) return fail(ctx, PT_ERR_INVALID, "pt_trace_primary: bad camera or params");
    cudaSetDevice(ctx->device);
    size_t npix = (size_t)cam->width * cam->height;
    DevTemps tmp;
    int32_t* d_obj = nullptr; float *d_t = nullptr, *d_dir = nullptr;
    PT_CUDA_CHECK(ctx, tmp.alloc(&d_obj, 4 * npix));
    PT_CUDA_CHECK(ctx, tmp.alloc(&d_t, 4 * npix));
    if (dir) PT_CUDA_CHECK(ctx, tmp.alloc(&d_dir, 12 * npix));
    int rc = wavefront_trace_primary(ctx, scene->dev, *cam, *p, jitter, ox, oy, sample, d_obj, d_t, d_dir, nullptr);
    if (!rc) {
        cudaMemcpyAsync(obj_id, d_obj, 4 * npix, cudaMemcpyDeviceToHost, ctx->stream);
        cudaMemcpyAsync(t, d_t, 4 * npix, cudaMemcpyDeviceToHost, ctx->stream);
        if (dir) cudaMemcpyAsync(dir, d_dir, 12 * npix, cudaMemcpyDeviceToHost, ctx->stream);
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { ctx->err = std::string("pt_trace_primary: ") + cudaGetErrorString(e); rc = PT_ERR_CUDA; }
    }
    return keep(ctx, rc);
}

int32_t pt_primary_cull_rects(const pt_shape* shapes, uint32_t n, const pt_camera* cam, int32_t* rects, int32_t* entry_of_object, uint32_t* n_entries) {
    if (!shapes || !rects || !entry_of_object || !n_entries) return fail(nullptr, PT_ERR_INVALID, "pt_primary_cull_rects: null argument");
    if (!valid_camera(cam)) return fail(nullptr, PT_ERR_INVALID, "pt_primary_cull_rects: bad camera (width/height must be >= 2)");
    for (uint32_t i = 0; i < n; ++i) if (shapes[i].kind > PT_SHAPE_TRIANGLE) return fail(nullptr, PT_ERR_INVALID, "pt_primary_cull_rects: unknown shape kind");
    if (!primary_cull_rects_host(shapes, n, *cam, rects, entry_of_object, n_entries))
        return fail(nullptr, PT_ERR_INVALID, "pt_primary_cull_rects: the scene has no constant-bank scan (empty, more than 1024 objects or more than 8 triangle pairs)");
    return PT_OK;
}

int32_t pt_bvh_download(pt_context* ctx, const pt_scene* scene, uint32_t* codes, uint32_t* prims, uint32_t* nodes, float* bounds) {
    if (!ctx || !scene) return fail(ctx, PT_ERR_INVALID, "pt_bvh_download: null argument");
    cudaSetDevice(ctx->device);
    const SceneDev& d = scene->dev;
    if (codes && d.n_objects) PT_CUDA_CHECK(ctx, cudaMemcpy(codes, d.codes_sorted, 4 * (size_t)d.n_objects, cudaMemcpyDeviceToHost));
    if (prims && d.n_objects) PT_CUDA_CHECK(ctx, cudaMemcpy(prims, d.prim_sorted, 4 * (size_t)d.n_objects, cudaMemcpyDeviceToHost));
    if (nodes) PT_CUDA_CHECK(ctx, cudaMemcpy(nodes, d.nodes, 64 * (size_t)d.n_nodes, cudaMemcpyDeviceToHost));
    if (bounds) std::memcpy(bounds, d.bounds, 24);
    return PT_OK;
}

int32_t pt_measure_fp32_peak(pt_context* ctx, float* tflops, float* ms) {
    if (!ctx) return fail(nullptr, PT_ERR_INVALID, "pt_measure_fp32_peak: null context");
    cudaSetDevice(ctx->device);
    return keep(ctx, fp32_peak_launch(ctx, tflops, ms));
}

// world.rs:344-369
int32_t pt_export_luminance(const char* path, const float* rad, uint32_t W, uint32_t H) {
    if (!path || !rad) return fail(nullptr, PT_ERR_INVALID, "pt_export_luminance: null argument");
    FILE* f = std::fopen(path, "w");
    if (!f) return fail(nullptr, PT_ERR_INVALID, std::string("pt_export_luminance: cannot open ") + path);
    std::fprintf(f, "x,y,r,g,b,luminance\n");
    for (uint32_t y = 0; y < H; ++y)
        for (uint32_t x = 0; x < W; ++x) {
            const float* v = rad + 3 * ((size_t)y * W + x);
            double r = v[0], g = v[1], b = v[2];
            double lum = 0.2126 * r + 0.7152 * g + 0.0722 * b;
            char sr[64], sg[64], sb[64], sl[64];
            fmt6(sr, sizeof sr, r); fmt6(sg, sizeof sg, g); fmt6(sb, sizeof sb, b); fmt6(sl, sizeof sl, lum);
            std::fprintf(f, "%u,%u,%s,%s,%s,%s\n", x, y, sr, sg, sb, sl);
        }
    std::fclose(f);
    return PT_OK;
}

int32_t pt_write_ppm(const char* path, const uint8_t* rgba, uint32_t W, uint32_t H) {
    if (!path || !rgba) return fail(nullptr, PT_ERR_INVALID, "pt_write_ppm: null argument");
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(nullptr, PT_ERR_INVALID, std::string("pt_write_ppm: cannot open ") + path);
    std::fprintf(f, "P6\n%u %u\n255\n", W, H);
    std::vector<uint8_t> row(3 * (size_t)W);
    for (uint32_t y = 0; y < H; ++y) {
        for (uint32_t x = 0; x < W; ++x) for (int c = 0; c < 3; ++c) row[3 * x + c] = rgba[4 * ((size_t)y * W + x) + c];
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    return PT_OK;
}

}  // extern "C"
