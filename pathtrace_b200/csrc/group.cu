// group.cu — multi-GPU through the C ABI (include/pathtrace_b200.h: pt_group_*).
//
// The path shards by samples per pixel (world.rs:296-313 just sums independent samples): device g renders the whole frame
// for the global sample indices [off_g, off_g + spp_g), into its own W*H float4 sum buffer; ONE ncclReduce(sum) to the root
// combines the buffers over NVLink; the root resolves (world.rs:315-332).  No other exchange exists on the path.  One host
// thread per device drives its context (a context is single-threaded by contract), the reduce is issued for all devices
// inside one ncclGroupStart / ncclGroupEnd on the contexts' own streams, so it is stream-ordered behind each render.
//
// NCCL is bound at run time with dlopen("libnccl.so.2"): the library keeps no link-time dependency on it (a host that never
// creates a group of more than one device never needs it, and a process that already loaded an NCCL — e.g. PyTorch's —
// shares that copy instead of loading a second one).
#include <dlfcn.h>
#include <nccl.h>   // types and enums only; every entry point is resolved with dlsym

#include <chrono>
#include <cstring>
#include <new>
#include <thread>

#include "pt_internal.h"

namespace {
thread_local std::string g_group_error;

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string& err) {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (handle) break;
        }
        if (!handle) { err = std::string("cannot load NCCL (libnccl.so.2): ") + dlerror(); return false; }
        auto sym = [&](const char* n) { void* p = dlsym(handle, n); if (!p && err.empty()) err = std::string("NCCL symbol missing: ") + n; return p; };
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(sym("ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
        Reduce = reinterpret_cast<decltype(Reduce)>(sym("ncclReduce"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
        return err.empty();
    }
};
}  // namespace

struct pt_group {
    std::vector<pt_context*> ctx;
    std::vector<ncclComm_t> comm;
    NcclApi nccl;
    std::string err;
};
struct pt_group_scene {
    pt_group* group = nullptr;
    std::vector<pt_scene*> scene;
};

namespace {
int gfail(pt_group* g, int code, const std::string& msg) {
    if (g) g->err = msg;
    g_group_error = msg;
    return code;
}
}  // namespace

extern "C" {

const char* pt_group_last_error(const pt_group* g) { return g ? g->err.c_str() : g_group_error.c_str(); }
int32_t pt_group_size(const pt_group* g) { return g ? (int32_t)g->ctx.size() : 0; }
pt_context* pt_group_context(pt_group* g, int32_t i) { return (g && i >= 0 && i < (int32_t)g->ctx.size()) ? g->ctx[(size_t)i] : nullptr; }

void pt_group_destroy(pt_group* g) {
    if (!g) return;
    for (size_t i = 0; i < g->comm.size(); ++i)
        if (g->comm[i]) { cudaSetDevice(g->ctx[i]->device); g->nccl.CommDestroy(g->comm[i]); }
    for (pt_context* c : g->ctx) pt_context_destroy(c);
    delete g;
}

int32_t pt_group_create(const int32_t* device_ids, int32_t n, pt_group** out) {
    if (!out) return gfail(nullptr, PT_ERR_INVALID, "pt_group_create: out is null");
    *out = nullptr;
    if (!device_ids || n < 1) return gfail(nullptr, PT_ERR_INVALID, "pt_group_create: need at least one device id");
    for (int32_t i = 0; i < n; ++i)
        for (int32_t j = 0; j < i; ++j)
            if (device_ids[i] == device_ids[j]) return gfail(nullptr, PT_ERR_INVALID, "pt_group_create: duplicate device id");
    pt_group* g = new (std::nothrow) pt_group();
    if (!g) return gfail(nullptr, PT_ERR_OOM, "out of host memory");
    for (int32_t i = 0; i < n; ++i) {
        pt_context* c = nullptr;
        int rc = pt_context_create(device_ids[i], &c);
        if (rc) { std::string m = pt_last_error(nullptr); pt_group_destroy(g); return gfail(nullptr, rc, "pt_group_create: device " + std::to_string(device_ids[i]) + ": " + m); }
        g->ctx.push_back(c);
    }
    if (n > 1) {
        std::string err;
        if (!g->nccl.load(err)) { pt_group_destroy(g); return gfail(nullptr, PT_ERR_CUDA, "pt_group_create: " + err + " (a multi-device group needs NCCL; there is no fallback)"); }
        g->comm.assign((size_t)n, nullptr);
        std::vector<int> devs(device_ids, device_ids + n);
        ncclResult_t r = g->nccl.CommInitAll(g->comm.data(), n, devs.data());
        if (r != ncclSuccess) {
            std::string m = std::string("ncclCommInitAll: ") + g->nccl.GetErrorString(r);
            g->comm.clear();
            pt_group_destroy(g);
            return gfail(nullptr, PT_ERR_CUDA, m);
        }
    }
    *out = g;
    return PT_OK;
}

void pt_group_scene_destroy(pt_group_scene* s) {
    if (!s) return;
    for (pt_scene* sc : s->scene) pt_scene_destroy(sc);
    delete s;
}

int32_t pt_group_scene_create(pt_group* g, const pt_shape* shapes, const pt_material* materials, uint32_t n_objects, pt_group_scene** out) {
    if (!g || !out) return gfail(g, PT_ERR_INVALID, "pt_group_scene_create: null argument");
    *out = nullptr;
    pt_group_scene* s = new (std::nothrow) pt_group_scene();
    if (!s) return gfail(g, PT_ERR_OOM, "out of host memory");
    s->group = g;
    const size_t n = g->ctx.size();
    s->scene.assign(n, nullptr);
    // scene and LBVH are replicated: every device builds its own copy (deterministic, so identical)
    std::vector<int> rc(n, 0);
    std::vector<std::thread> th;
    for (size_t i = 0; i < n; ++i) th.emplace_back([&, i] { rc[i] = pt_scene_create(g->ctx[i], shapes, materials, n_objects, &s->scene[i]); });
    for (auto& t : th) t.join();
    for (size_t i = 0; i < n; ++i)
        if (rc[i]) { std::string m = pt_last_error(g->ctx[i]); pt_group_scene_destroy(s); return gfail(g, rc[i], "pt_group_scene_create: device " + std::to_string(g->ctx[i]->device) + ": " + m); }
    *out = s;
    return PT_OK;
}

int32_t pt_group_render(pt_group* g, const pt_group_scene* scene, const pt_camera* cam, const pt_render_params* p, float* radiance, uint8_t* rgba,
                        pt_stats* stats) {
    if (!g || !scene || !cam || !p) return gfail(g, PT_ERR_INVALID, "pt_group_render: null argument");
    if (scene->group != g) return gfail(g, PT_ERR_INVALID, "pt_group_render: scene belongs to another group");
    if (p->spp == 0) return gfail(g, PT_ERR_INVALID, "pt_group_render: spp must be > 0");
    const size_t n = g->ctx.size();
    if (n == 1) {
        int rc = pt_render(g->ctx[0], scene->scene[0], cam, p, radiance, rgba, stats);
        if (rc) g->err = pt_last_error(g->ctx[0]);
        return rc;
    }
    const size_t npix = (size_t)cam->width * cam->height;
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<int> rc(n, 0);
    std::vector<pt_stats> st(n);
    std::vector<std::thread> th;
    const uint32_t base = p->spp / (uint32_t)n, rem = p->spp % (uint32_t)n;
    for (size_t i = 0; i < n; ++i)
        th.emplace_back([&, i] {
            pt_context* c = g->ctx[i];
            std::memset(&st[i], 0, sizeof(pt_stats));
            cudaSetDevice(c->device);
            if ((rc[i] = ensure_frame(c, npix)) != 0) return;
            if (cudaMemsetAsync(c->d_accum, 0, sizeof(float4) * npix, c->stream) != cudaSuccess) { c->err = "cudaMemsetAsync failed"; rc[i] = PT_ERR_CUDA; return; }
            pt_render_params q = *p;
            q.spp = base + (i < rem ? 1u : 0u);
            q.sample_offset = p->sample_offset + (uint32_t)i * base + (uint32_t)(i < rem ? i : rem);
            if (q.spp) rc[i] = pt_render_accumulate_device(c, scene->scene[i], cam, &q, c->d_accum, &st[i]);
        });
    for (auto& t : th) t.join();
    for (size_t i = 0; i < n; ++i)
        if (rc[i]) return gfail(g, rc[i], "pt_group_render: device " + std::to_string(g->ctx[i]->device) + ": " + pt_last_error(g->ctx[i]));
    // the one collective of the path
    ncclResult_t r = g->nccl.GroupStart();
    for (size_t i = 0; i < n && r == ncclSuccess; ++i)
        r = g->nccl.Reduce(g->ctx[i]->d_accum, g->ctx[i]->d_accum, npix * 4, ncclFloat32, ncclSum, 0, g->comm[i], g->ctx[i]->stream);
    ncclResult_t r2 = g->nccl.GroupEnd();
    if (r == ncclSuccess) r = r2;
    if (r != ncclSuccess) return gfail(g, PT_ERR_CUDA, std::string("ncclReduce: ") + g->nccl.GetErrorString(r));
    // root: mean, gamma, clamp, RGBA8 (world.rs:315-332), then D2H
    pt_context* root = g->ctx[0];
    cudaSetDevice(root->device);
    int rrc = resolve_launch(root, root->d_accum, cam->width, cam->height, p->spp, radiance ? root->d_radiance : nullptr, rgba ? root->d_rgba : nullptr);
    if (rrc) return gfail(g, rrc, root->err);
    if (radiance) cudaMemcpyAsync(radiance, root->d_radiance, sizeof(float) * 3 * npix, cudaMemcpyDeviceToHost, root->stream);
    if (rgba) cudaMemcpyAsync(rgba, root->d_rgba, 4 * npix, cudaMemcpyDeviceToHost, root->stream);
    for (size_t i = 0; i < n; ++i) {
        cudaSetDevice(g->ctx[i]->device);
        cudaError_t e = cudaStreamSynchronize(g->ctx[i]->stream);
        if (e != cudaSuccess) return gfail(g, PT_ERR_CUDA, std::string("pt_group_render: device ") + std::to_string(g->ctx[i]->device) + ": " + cudaGetErrorString(e));
    }
    if (stats) {
        pt_stats t;
        std::memset(&t, 0, sizeof t);
        for (size_t i = 0; i < n; ++i) {
            t.samples += st[i].samples; t.rays_primary += st[i].rays_primary; t.rays_shadow += st[i].rays_shadow; t.rays_scattered += st[i].rays_scattered;
            t.path_vertices += st[i].path_vertices; t.kernel_launches += st[i].kernel_launches; t.nodes_visited += st[i].nodes_visited; t.prim_tests += st[i].prim_tests;
            if (st[i].wavefront_iters > t.wavefront_iters) t.wavefront_iters = st[i].wavefront_iters;
            if (st[i].ms_render > t.ms_render) t.ms_render = st[i].ms_render;
            t.path_pool += st[i].path_pool;
        }
        t.kernel_launches += 1;
        t.scan_mode = st[0].scan_mode;
        t.ms_total = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        *stats = t;
    }
    return PT_OK;
}

}  // extern "C"
