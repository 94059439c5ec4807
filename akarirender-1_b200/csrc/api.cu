// api.cu — the C ABI of libakari_b200.so (include/akari_b200.h): device/scene lifecycle, HBM residency, the render
// driver (wavefront bounce loop or megakernel) and the batch ray-query entry points. No CPU fallback: every compute
// entry fails with AKB_NO_DEVICE / AKB_CUDA_ERROR when the GPU path is unavailable.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "bvh_build.cuh"
#include "kernels.cuh"

using namespace akb;

namespace {

thread_local std::string g_err;
int g_device = -1;
int g_n_sms = 0;
constexpr size_t kMaxBatch = (size_t)1 << 30;
bool g_profile_stages = false;
size_t g_trace_chunk = (size_t)1 << 21; // AKB_TRACE_CHUNK: rays per staging slot of akb_trace (host batches)
int g_check_every = 4;                 // AKB_CHECK_EVERY: bounces per graph launch / queue-length poll (even)
int g_finish_div = 4;                  // AKB_FINISH_DIV: a pipeline whose extension queue is below slots / this is finished by wf_finish_kernel (0: never)
int g_finish_max = 1 << 16;            // AKB_FINISH_MAX: ... and below this many paths (about one wave of wf_finish_kernel threads)
int g_timeline = 0;                    // AKB_TIMELINE=1: print every poll's device time and queue lengths (tuning aid)
int g_graphs = 1;                      // AKB_GRAPHS=0: launch every bounce's kernels one by one (tuning aid)
int g_poll_lag = 1;                    // AKB_POLL_LAG=0: wait for every queue-length poll (tuning aid)
int g_pipelines = 2;                   // AKB_PIPELINES: half-frame pipelines of the wavefront (1 or 2)
int g_pipeline_min_slots = 128 * 1024; // AKB_PIPELINE_MIN_SLOTS: below this one pipeline keeps the launches large

akb_status set_err(akb_status st, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return st;
}

#define CU(call)                                                                                                       \
    do {                                                                                                               \
        cudaError_t e_ = (call);                                                                                       \
        if (e_ != cudaSuccess)                                                                                         \
            return set_err(e_ == cudaErrorMemoryAllocation ? AKB_OOM : AKB_CUDA_ERROR, "%s: %s (%s:%d)", #call,        \
                           cudaGetErrorString(e_), __FILE__, __LINE__);                                                \
    } while (0)

} // namespace

namespace akb {
akb_status set_error(akb_status st, const std::string &msg) { // used by host/frontend_api.cpp
    g_err = msg;
    return st;
}
} // namespace akb

struct akb_scene {
    SceneDev dev;
    std::vector<void *> allocs;
    uint64_t n_tris = 0, n_nodes = 0, bvh_bytes = 0, n_lights = 0, tri_bytes = 0, bvh_depth = 0, build_us = 0, built_by = 0;
    int width = 0, height = 0;
    // render workspace (lazily sized)
    Workspace ws;
    std::vector<void *> ws_allocs;
    int ws_slots = 0;
    Counters *d_ctr = nullptr, *h_ctr = nullptr; // [2]: one per half-frame pipeline
    cudaStream_t aux_stream = nullptr;           // second pipeline's stream
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_poll[2] = {nullptr, nullptr};
    float4 *d_film = nullptr;
    size_t d_film_px = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // akb_trace (host ray batches): two staging slots of pinned host + device buffers, each on its own stream, so that
    // the host-side copies, both PCIe directions and the traversal kernel of consecutive chunks overlap
    struct TraceSlot {
        char *h_in = nullptr, *h_out = nullptr; // pinned: org | dir, results
        float4 *d_o = nullptr, *d_d = nullptr;
        void *d_out = nullptr;
        uint32_t *cursor = nullptr;
        cudaStream_t st = nullptr;
    } trace_slot[2];
    size_t trace_cap = 0; // rays per slot
    // one CUDA graph per half-frame pipeline: check_every bounces (trace + shade each) of the wavefront loop, captured
    // once and relaunched for the whole frame (and later frames with the same parameters)
    struct BounceGraph {
        cudaGraphExec_t exec = nullptr;
        RenderConfig rc;
        Workspace ws;
        int n_bounces = 0;
    } bounce_graph[2];
    cudaStream_t cap_stream = nullptr;
    void free_graphs() {
        for (BounceGraph &g : bounce_graph) {
            if (g.exec) cudaGraphExecDestroy(g.exec);
            g.exec = nullptr;
        }
    }

    void free_trace() {
        for (TraceSlot &t : trace_slot) {
            if (t.h_in) cudaFreeHost(t.h_in);
            if (t.h_out) cudaFreeHost(t.h_out);
            if (t.d_o) cudaFree(t.d_o);
            if (t.d_d) cudaFree(t.d_d);
            if (t.d_out) cudaFree(t.d_out);
            if (t.cursor) cudaFree(t.cursor);
            if (t.st) cudaStreamDestroy(t.st);
            t = TraceSlot();
        }
        trace_cap = 0;
    }
    void free_ws() {
        for (void *p : ws_allocs) cudaFree(p);
        ws_allocs.clear();
        ws_slots = 0;
    }
    ~akb_scene() {
        free_graphs();
        if (cap_stream) cudaStreamDestroy(cap_stream);
        free_ws();
        free_trace();
        for (void *p : allocs) cudaFree(p);
        if (d_ctr) cudaFree(d_ctr);
        if (h_ctr) cudaFreeHost(h_ctr);
        if (d_film) cudaFree(d_film);
        if (aux_stream) cudaStreamDestroy(aux_stream);
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (ev_join) cudaEventDestroy(ev_join);
        for (cudaEvent_t e : ev_poll)
            if (e) cudaEventDestroy(e);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
    }
};

namespace {

// memcpy over a few host threads: one core moves ~10 GB/s, less than one direction of the PCIe link
void host_copy(void *dst, const void *src, size_t bytes) {
    const size_t piece = (size_t)4 << 20;
    const int n_threads = (int)std::min<size_t>(8, bytes / piece);
    if (n_threads <= 1) {
        memcpy(dst, src, bytes);
        return;
    }
    std::vector<std::thread> th;
    const size_t per = ((bytes / n_threads) + 4095) & ~(size_t)4095;
    for (int i = 0; i < n_threads; i++) {
        const size_t b = (size_t)i * per, e = std::min(bytes, b + per);
        if (b >= e) break;
        th.emplace_back([=]() { memcpy((char *)dst + b, (const char *)src + b, e - b); });
    }
    for (auto &t : th) t.join();
}


template <class T> akb_status upload(akb_scene *sc, const T *src, size_t n, const T **dst) {
    void *p = nullptr;
    size_t bytes = std::max<size_t>(n * sizeof(T), 16);
    CU(cudaMalloc(&p, bytes));
    sc->allocs.push_back(p);
    if (n) CU(cudaMemcpy(p, src, n * sizeof(T), cudaMemcpyHostToDevice));
    *dst = (const T *)p;
    return AKB_OK;
}

template <class T> akb_status ws_alloc(akb_scene *sc, size_t n, T **dst) {
    void *p = nullptr;
    CU(cudaMalloc(&p, std::max<size_t>(n * sizeof(T), 16)));
    sc->ws_allocs.push_back(p);
    *dst = (T *)p;
    return AKB_OK;
}

akb_status ensure_workspace(akb_scene *sc, int n_slots) {
    if (sc->ws_slots >= n_slots && sc->ws_slots > 0) {
        sc->ws.n_slots = n_slots;
        return AKB_OK;
    }
    sc->free_ws();
    Workspace &w = sc->ws;
    const size_t n = (size_t)std::max(n_slots, 32);
    akb_status st;
#define WS(field, type)                                                                                                \
    if ((st = ws_alloc<type>(sc, n, &w.field)) != AKB_OK) return st;
    WS(ray_o, float4) WS(ray_d, float4) WS(hit, float4) WS(beta, float4) WS(L, float4) WS(rng, unsigned long long)
    WS(sh_o, float4) WS(sh_d, float4) WS(sh_c, float4) WS(sh_b, float4) WS(acc, float4)
    for (int b = 0; b < 2; b++) {
        WS(q_ext[b], uint32_t) WS(q_sh[b], uint32_t)
        for (int k = 0; k < kNumKeys; k++) { WS(q_shade[b][k], uint32_t) }
    }
#undef WS
    w.ctr = sc->d_ctr;
    w.slot_base = 0;
    w.n_slots = n_slots;
    sc->ws_slots = n_slots;
    return AKB_OK;
}

// The slots [base, base + n) of the frame's workspace as a workspace of their own (own queues, own counters).
Workspace workspace_view(const Workspace &w, int base, int n, Counters *ctr) {
    Workspace v = w;
    v.n_slots = n, v.slot_base = w.slot_base + base, v.ctr = ctr;
    v.ray_o += base, v.ray_d += base, v.hit += base, v.beta += base, v.L += base, v.rng += base;
    v.sh_o += base, v.sh_d += base, v.sh_c += base, v.sh_b += base, v.acc += base;
    for (int b = 0; b < 2; b++) {
        v.q_ext[b] += base, v.q_sh[b] += base;
        for (int k = 0; k < kNumKeys; k++) v.q_shade[b][k] += base;
    }
    return v;
}

akb_status check_ready(const char *who) {
    if (g_device < 0) return set_err(AKB_NO_DEVICE, "%s: akb_init() has not succeeded (no CUDA device bound; there is no CPU fallback)", who);
    return AKB_OK;
}

} // namespace

extern "C" {

const char *akb_last_error(void) { return g_err.c_str(); }

int akb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

akb_status akb_init(int device_index) {
    int n = akb_device_count();
    if (n <= 0) return set_err(AKB_NO_DEVICE, "akb_init: no CUDA device visible (this backend has no CPU fallback)");
    if (device_index < 0 || device_index >= n) return set_err(AKB_INVALID_ARG, "akb_init: device %d out of range [0,%d)", device_index, n);
    CU(cudaSetDevice(device_index));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device_index));
    if (prop.major < 10)
        return set_err(AKB_NO_DEVICE, "akb_init: device %d is sm_%d%d; this library carries sm_100a code only", device_index, prop.major, prop.minor);
    g_n_sms = prop.multiProcessorCount;
    // arithmetic self-test: the parity contract needs un-contracted fp32 (see vecmath.h)
    int *d_res = nullptr, h_res = 0;
    CU(cudaMalloc(&d_res, sizeof(int)));
    cudaError_t e = launch_selftest(d_res, 0);
    if (e == cudaSuccess) e = cudaMemcpy(&h_res, d_res, sizeof(int), cudaMemcpyDeviceToHost);
    cudaFree(d_res);
    if (e != cudaSuccess) return set_err(AKB_CUDA_ERROR, "akb_init: self-test launch failed: %s", cudaGetErrorString(e));
    if (h_res != 1) return set_err(AKB_CUDA_ERROR, "akb_init: kernels were built with FMA contraction; rebuild with -fmad=false");
    const char *pf = getenv("AKB_PROFILE_STAGES");
    g_profile_stages = pf && pf[0] == '1';
    if (const char *e = getenv("AKB_PIPELINES")) g_pipelines = atoi(e);
    if (const char *e = getenv("AKB_POLL_LAG")) g_poll_lag = atoi(e) != 0;
    if (const char *e = getenv("AKB_GRAPHS")) g_graphs = atoi(e) != 0;
    if (const char *e = getenv("AKB_TIMELINE")) g_timeline = atoi(e) != 0;
    if (const char *e = getenv("AKB_FINISH_DIV")) g_finish_div = std::max(0, atoi(e));
    if (const char *e = getenv("AKB_FINISH_MAX")) g_finish_max = std::max(1, atoi(e));
    if (const char *e = getenv("AKB_CHECK_EVERY")) g_check_every = std::max(2, atoi(e) & ~1);
    if (const char *e = getenv("AKB_TRACE_CHUNK")) g_trace_chunk = std::max<size_t>(1024, (size_t)atoll(e));
    if (const char *e = getenv("AKB_PIPELINE_MIN_SLOTS")) g_pipeline_min_slots = atoi(e);
    if (const char *rb = getenv("AKB_REFILL_BELOW")) {
        int v = atoi(rb);
        if (v >= 1 && v <= 32) CU(set_refill_below(v));
    }
    if (const char *e = getenv("AKB_NO_PARK")) CU(set_no_park(atoi(e) != 0));
    if (const char *fc = getenv("AKB_FETCH_CHUNK")) {
        int v = atoi(fc);
        if (v >= 0 && v <= 4096) CU(set_fetch_chunk(v));
    }
    g_device = device_index;
    return AKB_OK;
}

akb_status akb_shutdown(void) {
    if (g_device >= 0) cudaDeviceSynchronize();
    g_device = -1;
    return AKB_OK;
}

namespace {
// AKB_BVH_DEVICE: csrc/bvh_build.cu, with the host builder as the fallback for trees deeper than the stack
bool device_bvh_builder(const float *lo, const float *hi, size_t n, float pad, int n_threads, Bvh8 *out, std::string *err) {
    const auto t0 = std::chrono::steady_clock::now();
    float node_cost, prim_cost;
    int max_leaf;
    bvh8_collapse_params(&node_cost, &prim_cost, &max_leaf);
    bool fallback = false;
    DeviceBuildTimes times;
    const cudaError_t e = build_bvh8_device(lo, hi, n, pad, node_cost, prim_cost, max_leaf, out, &fallback, &times);
    if (e != cudaSuccess) {
        *err = std::string("device BVH build: ") + cudaGetErrorString(e);
        cudaGetLastError();
        return false;
    }
    if (fallback) {
        build_bvh8(lo, hi, n, pad, n_threads, out);
        return true;
    }
    out->built_by = AKB_BVH_DEVICE;
    out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (getenv("AKB_BVH_VERBOSE"))
        fprintf(stderr, "[akb] device BVH build: %zu triangles -> %zu nodes, depth %d (%s, %d rounds); upload %.2f sort %.2f tree %.2f fit %.2f emit %.2f download %.2f ms\n",
                n, out->nodes.size(), out->depth, times.ploc_rounds ? "PLOC" : "radix tree", times.ploc_rounds, times.ms_upload, times.ms_sort, times.ms_tree, times.ms_fit, times.ms_emit, times.ms_download);
    return true;
}
} // namespace

akb_status akb_scene_create(const akb_scene_desc *desc, akb_scene **out_scene) {
    if (!out_scene) return set_err(AKB_INVALID_ARG, "akb_scene_create: out_scene == NULL");
    *out_scene = nullptr;
    akb_status st = check_ready("akb_scene_create");
    if (st != AKB_OK) return st;
    HostScene hs;
    std::string err;
    int build = desc ? (int)desc->bvh_build : 0;
    if (build == AKB_BVH_DEFAULT) {
        // host SAH for scenes it builds in about a second, the device build above that (measured on B200: C3's 1 M
        // triangles 0.43 s vs 0.03 s to build but 1 % slower to render (C5: 2 %); C4's 10 M: 4.4 s vs 0.08 s, 2.4 %
        // slower; a 1 M-triangle clustered soup: 0.9 s vs 0.03 s, 1-8 % FASTER to traverse)
        uint64_t nt = 0;
        if (desc && desc->meshes)
            for (uint32_t g = 0; g < desc->n_meshes; g++) nt += desc->meshes[g].n_triangles;
        const char *e = getenv("AKB_BVH_BUILD");
        if (e && !strcmp(e, "device"))
            build = AKB_BVH_DEVICE;
        else if (e && !strcmp(e, "host"))
            build = AKB_BVH_HOST_SAH;
        else
            build = nt >= (4u << 20) ? AKB_BVH_DEVICE : AKB_BVH_HOST_SAH;
    }
    if (build != AKB_BVH_HOST_SAH && build != AKB_BVH_DEVICE) return set_err(AKB_INVALID_ARG, "akb_scene_create: bad bvh_build %d", build);
    const auto t_create = std::chrono::steady_clock::now();
    st = flatten_scene(desc, 0, &hs, &err, build == AKB_BVH_DEVICE ? device_bvh_builder : nullptr);
    if (st != AKB_OK) return set_err(st, "%s", err.c_str());
    const auto t_flat = std::chrono::steady_clock::now();
    akb_scene *sc = new akb_scene();
    SceneDev &S = sc->dev;
    memset(&S, 0, sizeof S);
    { // BVH nodes and leaf triangles share one allocation (the traversal working set: 56 MB for 1 M triangles)
        const size_t node_f4 = hs.bvh.nodes.size() * 5, node_pad = (node_f4 + 15) & ~(size_t)15, tri_f4 = hs.tris.size();
        float4 *buf = nullptr;
        const size_t bytes = std::max<size_t>((node_pad + tri_f4) * sizeof(float4), 16);
        cudaError_t e = cudaMalloc((void **)&buf, bytes);
        if (e == cudaSuccess) sc->allocs.push_back(buf);
        if (e == cudaSuccess && node_f4) e = cudaMemcpy(buf, hs.bvh.nodes.data(), node_f4 * sizeof(float4), cudaMemcpyHostToDevice);
        if (e == cudaSuccess && tri_f4) e = cudaMemcpy(buf + node_pad, hs.tris.data(), tri_f4 * sizeof(float4), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) st = set_err(e == cudaErrorMemoryAllocation ? AKB_OOM : AKB_CUDA_ERROR, "akb_scene_create: %s", cudaGetErrorString(e));
        S.nodes = buf, S.tris = buf + node_pad;
    }
    if (st == AKB_OK) st = upload(sc, (const float4 *)hs.shading.data(), hs.shading.size(), &S.shading);
    if (st == AKB_OK) st = upload(sc, hs.mats.data(), hs.mats.size(), &S.mats);
    if (st == AKB_OK) st = upload(sc, (const float4 *)hs.lights.data(), hs.lights.size(), &S.lights);
    if (st == AKB_OK) st = upload(sc, hs.light_cdf.data(), hs.light_cdf.size(), &S.light_cdf);
    if (st == AKB_OK) st = upload(sc, hs.light_func.data(), hs.light_func.size(), &S.light_func);
    if (st == AKB_OK && !hs.tex_info.empty()) { // image textures: texels, per-texture info, per-triangle / per-light texcoords
        st = upload(sc, (const float4 *)hs.texels.data(), hs.texels.size(), &S.texels);
        if (st == AKB_OK) st = upload(sc, (const int4 *)hs.tex_info.data(), hs.tex_info.size() / 4, &S.tex_info);
        if (st == AKB_OK) st = upload(sc, (const float4 *)hs.tri_tc.data(), hs.tri_tc.size(), &S.tri_tc);
        if (st == AKB_OK) st = upload(sc, (const float4 *)hs.light_tc.data(), hs.light_tc.size(), &S.light_tc);
    }
    if (st == AKB_OK) {
        cudaError_t e = cudaMalloc((void **)&sc->d_ctr, 2 * sizeof(Counters));
        if (e == cudaSuccess) e = cudaMallocHost((void **)&sc->h_ctr, 2 * sizeof(Counters));
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&sc->aux_stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sc->ev_fork, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sc->ev_join, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&sc->cap_stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sc->ev_poll[0], g_timeline ? cudaEventDefault : cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sc->ev_poll[1], g_timeline ? cudaEventDefault : cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreate(&sc->ev0);
        if (e == cudaSuccess) e = cudaEventCreate(&sc->ev1);
        if (e != cudaSuccess) st = set_err(AKB_CUDA_ERROR, "akb_scene_create: %s", cudaGetErrorString(e));
    }
    if (st != AKB_OK) {
        delete sc;
        return st;
    }
    if (getenv("AKB_BVH_VERBOSE")) {
        const auto t_up = std::chrono::steady_clock::now();
        fprintf(stderr, "[akb] scene create: flatten + BVH %.3f s (BVH %.3f s), upload %.3f s\n", std::chrono::duration<double>(t_flat - t_create).count(),
                hs.bvh.build_seconds, std::chrono::duration<double>(t_up - t_flat).count());
    }
    S.light_func_int = hs.light_func_int;
    S.n_lights = (int)hs.light_func.size();
    S.n_mats = (int)hs.mats.size();
    // if-if traversal for real trees, fused step for tiny ones (AKB_TRAV_IFIF=0/1 forces either; tuning aid)
    {
        const char *e = getenv("AKB_TRAV_IFIF");
        const bool ifif = e ? e[0] == '1' : hs.bvh.nodes.size() >= 64;
        S.trav_post_limit = ifif ? std::max(0, kStackSize - hs.bvh.depth - 1) : -1;
    }
    S.n_textures = (int)(hs.tex_info.size() / 4);
    S.cam = hs.cam;
    sc->n_tris = hs.n_tris, sc->n_nodes = hs.bvh.nodes.size(), sc->bvh_bytes = hs.bvh.nodes.size() * sizeof(Node8);
    sc->n_lights = hs.light_func.size(), sc->tri_bytes = hs.tris.size() * sizeof(f4);
    sc->bvh_depth = (uint64_t)hs.bvh.depth, sc->build_us = (uint64_t)(hs.bvh.build_seconds * 1e6);
    sc->built_by = (uint64_t)hs.bvh.built_by;
    sc->width = hs.cam.width, sc->height = hs.cam.height;
    *out_scene = sc;
    return AKB_OK;
}

akb_status akb_scene_destroy(akb_scene *scene) {
    if (!scene) return AKB_OK;
    if (g_device >= 0) cudaDeviceSynchronize();
    delete scene;
    return AKB_OK;
}

uint64_t akb_scene_info(const akb_scene *sc, int what) {
    if (!sc) return 0;
    switch (what) {
    case 0: return sc->n_tris;
    case 1: return sc->n_nodes;
    case 2: return sc->bvh_bytes;
    case 3: return sc->n_lights;
    case 4: return sc->tri_bytes;
    case 5: return sc->bvh_depth;
    case 6: return sc->build_us;
    case 7: return sc->built_by;
    }
    return 0;
}

akb_status akb_render_device(akb_scene *sc, const akb_render_params *p, float *d_out_rgbw, void *cuda_stream,
                             akb_render_stats *stats) {
    akb_status st = check_ready("akb_render_device");
    if (st != AKB_OK) return st;
    if (!sc || !p || !d_out_rgbw) return set_err(AKB_INVALID_ARG, "akb_render_device: null argument");
    if (p->integrator != AKB_INTEGRATOR_PATH && p->integrator != AKB_INTEGRATOR_AO) return set_err(AKB_INVALID_ARG, "akb_render_device: unknown integrator %d", p->integrator);
    if (p->integrator == AKB_INTEGRATOR_AO && !(p->ao_occlude > 0.0f)) return set_err(AKB_INVALID_ARG, "akb_render_device: AO.occlude must be positive");
    if (p->spp < 0 || p->max_depth < 0) return set_err(AKB_INVALID_ARG, "akb_render_device: negative spp/max_depth");
    if (p->sampler != AKB_SAMPLER_LCG && p->sampler != AKB_SAMPLER_PCG) return set_err(AKB_INVALID_ARG, "akb_render_device: bad sampler");
    const int world = p->world_size > 0 ? p->world_size : 1;
    if (p->rank < 0 || p->rank >= world) return set_err(AKB_INVALID_ARG, "akb_render_device: rank %d outside world_size %d", p->rank, world);
    const int tile = p->tile_size > 0 ? p->tile_size : 32;
    if ((tile & (tile - 1)) != 0 || tile < 8 || tile > 1024) return set_err(AKB_INVALID_ARG, "akb_render_device: tile_size must be a power of two in [8,1024]");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    RenderConfig rc;
    memset(&rc, 0, sizeof rc); // compared bytewise with the cached graphs' copy
    if (p->split != AKB_SPLIT_TILE && p->split != AKB_SPLIT_SAMPLE) return set_err(AKB_INVALID_ARG, "akb_render_device: bad split %d", p->split);
    const Shard shard = make_shard(sc->width, sc->height, p);
    rc.spp = shard.spp, rc.max_depth = p->max_depth, rc.rr_start = p->rr_start_depth, rc.sampler = p->sampler;
    rc.sort_materials = p->sort_materials, rc.ray_clamp = p->ray_clamp;
    rc.seed_offset = shard.seed_offset;
    rc.map = shard.map;
    rc.ao = p->integrator == AKB_INTEGRATOR_AO, rc.ao_occlude = p->ao_occlude;
    if (rc.ao) rc.sort_materials = 0, rc.ray_clamp = 0.0f; // AO: no materials, samples are exact 0/1
    const size_t n_px = (size_t)sc->width * sc->height;
    uint64_t launches = 0;
    std::vector<cudaEvent_t> evs;
    CU(cudaMemsetAsync(d_out_rgbw, 0, n_px * sizeof(float4), stream));
    CU(cudaMemsetAsync(sc->d_ctr, 0, 2 * sizeof(Counters), stream));
    CU(cudaEventRecord(sc->ev0, stream));
    if (p->integrator == AKB_INTEGRATOR_AO && p->mode == AKB_MODE_MEGAKERNEL) {
        CU(launch_ao(sc->dev, rc, p->ao_occlude, (float4 *)d_out_rgbw, sc->d_ctr, stream));
        launches = 1;
    } else if (p->mode == AKB_MODE_MEGAKERNEL) {
        CU(launch_megakernel(sc->dev, rc, (float4 *)d_out_rgbw, sc->d_ctr, stream));
        launches = 1;
    } else {
        const int n_total = rc.map.n_slots();
        st = ensure_workspace(sc, n_total);
        if (st != AKB_OK) return st;
        // Two half-frame pipelines on two streams: while one half's traversal launch drains (persistent warps running
        // their last, longest rays at a few lanes each) the other half's kernels fill the machine. Halves are whole
        // tiles, so every slot keeps its pixel, seed and accumulation order: the film is unchanged.
        const int tt = rc.map.tile * rc.map.tile;
        const int n_pipes = (g_pipelines >= 2 && !g_profile_stages && rc.map.n_owned_tiles >= 2 && n_total >= g_pipeline_min_slots) ? 2 : 1;
        const int base1 = n_pipes == 2 ? (rc.map.n_owned_tiles / 2) * tt : n_total;
        struct Pipe {
            Workspace ws;
            cudaStream_t st;
            bool live;
        } pipe[2];
        pipe[0] = Pipe{workspace_view(sc->ws, 0, base1, sc->d_ctr), stream, n_total > 0 && rc.spp > 0};
        pipe[1] = Pipe{workspace_view(sc->ws, base1, n_total - base1, sc->d_ctr + 1), sc->aux_stream, n_pipes == 2 && rc.spp > 0};
        if (n_pipes == 2) {
            CU(cudaEventRecord(sc->ev_fork, stream));
            CU(cudaStreamWaitEvent(sc->aux_stream, sc->ev_fork, 0));
        }
        for (int q = 0; q < n_pipes; q++) {
            CU(launch_wf_init(sc->dev, rc, pipe[q].ws, g_n_sms, pipe[q].st));
            launches++;
        }
        const long max_it = (long)rc.spp * (rc.max_depth + 2) + 2;
        const int check_every = g_check_every; // even: a snapshot taken after an odd bounce is read after an odd bounce
        bool poll_pending[2] = {false, false};
        auto mark = [&]() {
            if (!g_profile_stages) return cudaSuccess;
            cudaEvent_t e;
            cudaError_t r = cudaEventCreate(&e);
            if (r != cudaSuccess) return r;
            evs.push_back(e);
            return cudaEventRecord(e, stream);
        };
        CU(mark());
        const bool use_graphs = g_graphs && !g_profile_stages;
        const int finish_div = (rc.ao || g_profile_stages) ? 0 : g_finish_div; // the AO integrator keeps to its queues
        if (use_graphs)
            for (int q = 0; q < n_pipes; q++) {
                akb_scene::BounceGraph &g = sc->bounce_graph[q];
                if (g.exec && g.n_bounces == check_every && !memcmp(&g.rc, &rc, sizeof rc) && !memcmp(&g.ws, &pipe[q].ws, sizeof(Workspace))) continue;
                if (g.exec) cudaGraphExecDestroy(g.exec);
                g.exec = nullptr;
                CU(cudaStreamBeginCapture(sc->cap_stream, cudaStreamCaptureModeThreadLocal));
                cudaError_t ce = cudaSuccess;
                for (int k = 0; k < check_every && ce == cudaSuccess; k++) {
                    ce = launch_wf_trace(sc->dev, rc, pipe[q].ws, k & 1, g_n_sms, sc->cap_stream);
                    if (ce == cudaSuccess) ce = launch_wf_shade(sc->dev, rc, pipe[q].ws, k & 1, g_n_sms, sc->cap_stream);
                }
                cudaGraph_t graph = nullptr;
                const cudaError_t ee = cudaStreamEndCapture(sc->cap_stream, &graph);
                if (ce == cudaSuccess) ce = ee;
                if (ce == cudaSuccess) ce = cudaGraphInstantiate(&g.exec, graph, 0);
                if (graph) cudaGraphDestroy(graph);
                CU(ce);
                g.rc = rc, g.ws = pipe[q].ws, g.n_bounces = check_every;
            }
        for (long it = 0; it < max_it && (pipe[0].live || pipe[1].live); it++) {
            const int b = (int)(it & 1);
            if (use_graphs) {
                if ((it % check_every) == 0)
                    for (int q = 0; q < n_pipes; q++)
                        if (pipe[q].live) {
                            CU(cudaGraphLaunch(sc->bounce_graph[q].exec, pipe[q].st));
                            launches += 2 * check_every;
                        }
            } else {
                for (int q = 0; q < n_pipes; q++)
                    if (pipe[q].live) {
                        CU(launch_wf_trace(sc->dev, rc, pipe[q].ws, b, g_n_sms, pipe[q].st));
                        launches++;
                    }
                CU(mark());
                for (int q = 0; q < n_pipes; q++)
                    if (pipe[q].live) {
                        CU(launch_wf_shade(sc->dev, rc, pipe[q].ws, b, g_n_sms, pipe[q].st));
                        launches++;
                    }
                CU(mark());
            }
            if ((it % check_every) == check_every - 1) {
                // The queue lengths are read one poll late: the copy issued here is looked at only after the next
                // check_every bounces have been enqueued, so the GPU never idles while the host waits for it. A
                // finished pipeline runs at most check_every more bounces on empty queues (microseconds).
                for (int q = 0; q < n_pipes; q++) {
                    if (!pipe[q].live) continue;
                    if (poll_pending[q] || !g_poll_lag) {
                        if (!g_poll_lag) {
                            CU(cudaMemcpyAsync(sc->h_ctr + q, sc->d_ctr + q, sizeof(Counters), cudaMemcpyDeviceToHost, pipe[q].st));
                            CU(cudaStreamSynchronize(pipe[q].st));
                        } else {
                            CU(cudaEventSynchronize(sc->ev_poll[q]));
                        }
                        poll_pending[q] = false;
                        if (g_timeline) {
                            float ms = 0.f;
                            cudaEventElapsedTime(&ms, sc->ev0, sc->ev_poll[q]);
                            fprintf(stderr, "akb timeline: pipe %d after bounce %ld: %.3f ms, n_ext %u n_sh %u\n", q, it - check_every,
                                    ms, sc->h_ctr[q].n_ext[b ^ 1], sc->h_ctr[q].n_sh[b ^ 1]);
                        }
                        const uint32_t left = sc->h_ctr[q].n_ext[b ^ 1];
                        if (left == 0 && sc->h_ctr[q].n_sh[b ^ 1] == 0) {
                            pipe[q].live = false;
                        } else if (finish_div > 0 && left <= (uint32_t)std::min(pipe[q].ws.n_slots / finish_div, g_finish_max)) {
                            // few paths left (the count only falls from here on): one thread per path runs them out.
                            // Everything enqueued so far ends on a shade step of parity b, so the live queues are b ^ 1.
                            CU(launch_wf_finish(sc->dev, rc, pipe[q].ws, b ^ 1, left, pipe[q].st));
                            launches++;
                            pipe[q].live = false;
                        }
                    }
                    if (pipe[q].live && g_poll_lag) {
                        CU(cudaMemcpyAsync(sc->h_ctr + q, sc->d_ctr + q, sizeof(Counters), cudaMemcpyDeviceToHost, pipe[q].st));
                        CU(cudaEventRecord(sc->ev_poll[q], pipe[q].st));
                        poll_pending[q] = true;
                    }
                }
            }
        }
        if (n_pipes == 2) {
            CU(cudaEventRecord(sc->ev_join, sc->aux_stream));
            CU(cudaStreamWaitEvent(stream, sc->ev_join, 0));
        }
        for (int q = 0; q < n_pipes; q++) {
            CU(launch_wf_resolve(rc, pipe[q].ws, (float4 *)d_out_rgbw, stream));
            launches++;
        }
    }
    CU(cudaEventRecord(sc->ev1, stream));
    CU(cudaMemcpyAsync(sc->h_ctr, sc->d_ctr, 2 * sizeof(Counters), cudaMemcpyDeviceToHost, stream));
    CU(cudaStreamSynchronize(stream));
    if (stats) {
        memset(stats, 0, sizeof *stats);
        for (int q = 0; q < 2; q++) {
            stats->n_samples += sc->h_ctr[q].stat_samples;
            stats->n_closest_rays += sc->h_ctr[q].stat_closest - sc->h_ctr[q].stat_skipped;
            stats->n_shadow_rays += sc->h_ctr[q].stat_shadow;
        }
        stats->n_kernel_launches = launches;
        CU(cudaEventElapsedTime(&stats->ms_total, sc->ev0, sc->ev1));
        if (evs.size() >= 3) {
            for (size_t i = 1; i + 1 < evs.size(); i += 2) {
                float a = 0, b2 = 0;
                cudaEventElapsedTime(&a, evs[i - 1], evs[i]);
                cudaEventElapsedTime(&b2, evs[i], evs[i + 1]);
                stats->ms_extend += a, stats->ms_shade += b2;
            }
        }
    }
    for (cudaEvent_t e : evs) cudaEventDestroy(e);
    return AKB_OK;
}

akb_status akb_render(akb_scene *sc, const akb_render_params *p, float *out_rgbw, akb_render_stats *stats) {
    akb_status st = check_ready("akb_render");
    if (st != AKB_OK) return st;
    if (!sc || !p || !out_rgbw) return set_err(AKB_INVALID_ARG, "akb_render: null argument");
    const size_t n_px = (size_t)sc->width * sc->height;
    if (sc->d_film_px < n_px) {
        if (sc->d_film) cudaFree(sc->d_film);
        sc->d_film = nullptr, sc->d_film_px = 0;
        CU(cudaMalloc((void **)&sc->d_film, n_px * sizeof(float4)));
        sc->d_film_px = n_px;
    }
    st = akb_render_device(sc, p, (float *)sc->d_film, nullptr, stats);
    if (st != AKB_OK) return st;
    CU(cudaMemcpy(out_rgbw, sc->d_film, n_px * sizeof(float4), cudaMemcpyDeviceToHost));
    return AKB_OK;
}

akb_status akb_trace_device(akb_scene *sc, const float *d_org, const float *d_dir, size_t n, int any_hit, akb_hit *d_hits,
                            int32_t *d_occ, void *cuda_stream, float *ms_kernel) {
    akb_status st = check_ready("akb_trace_device");
    if (st != AKB_OK) return st;
    if (!sc || (n && (!d_org || !d_dir)) || (n && !any_hit && !d_hits) || (n && any_hit && !d_occ))
        return set_err(AKB_INVALID_ARG, "akb_trace_device: null argument");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    if (ms_kernel) CU(cudaEventRecord(sc->ev0, stream));
    for (size_t off = 0; off < n; off += kMaxBatch) { // the kernel indexes rays with 32 bits
        const size_t m = std::min(kMaxBatch, n - off);
        CU(launch_trace_batch(sc->dev, (const float4 *)d_org + off, (const float4 *)d_dir + off, m, any_hit,
                              d_hits ? d_hits + off : nullptr, d_occ ? d_occ + off : nullptr, &sc->d_ctr->fetch_trace, g_n_sms, stream));
    }
    if (ms_kernel) CU(cudaEventRecord(sc->ev1, stream));
    CU(cudaStreamSynchronize(stream));
    if (ms_kernel) CU(cudaEventElapsedTime(ms_kernel, sc->ev0, sc->ev1));
    return AKB_OK;
}

akb_status akb_trace(akb_scene *sc, const float *org, const float *dir, size_t n, int any_hit, akb_hit *out_hits,
                     int32_t *out_occ) {
    akb_status st = check_ready("akb_trace");
    if (st != AKB_OK) return st;
    if (!sc || (n && (!org || !dir)) || (n && !any_hit && !out_hits) || (n && any_hit && !out_occ))
        return set_err(AKB_INVALID_ARG, "akb_trace: null argument");
    if (n == 0) return AKB_OK;
    const size_t out_elt = any_hit ? sizeof(int32_t) : sizeof(akb_hit);
    const size_t chunk = std::min<size_t>(n, g_trace_chunk);
    cudaError_t e = cudaSuccess;
    if (sc->trace_cap < chunk) { // (re)size the staging slots; kept with the scene: pinning memory is slow
        sc->free_trace();
        for (akb_scene::TraceSlot &t : sc->trace_slot) {
            if (e == cudaSuccess) e = cudaMallocHost((void **)&t.h_in, chunk * 2 * sizeof(float4));
            if (e == cudaSuccess) e = cudaMallocHost((void **)&t.h_out, chunk * sizeof(akb_hit));
            if (e == cudaSuccess) e = cudaMalloc((void **)&t.d_o, chunk * sizeof(float4));
            if (e == cudaSuccess) e = cudaMalloc((void **)&t.d_d, chunk * sizeof(float4));
            if (e == cudaSuccess) e = cudaMalloc(&t.d_out, chunk * sizeof(akb_hit));
            if (e == cudaSuccess) e = cudaMalloc((void **)&t.cursor, sizeof(uint32_t));
            if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&t.st, cudaStreamNonBlocking);
        }
        if (e == cudaSuccess) sc->trace_cap = chunk;
    }
    char *out = any_hit ? (char *)out_occ : (char *)out_hits;
    const size_t n_chunks = (n + chunk - 1) / chunk;
    auto drain = [&](size_t k) { // chunk k's results: wait for its slot, then pinned -> caller
        akb_scene::TraceSlot &t = sc->trace_slot[k & 1];
        cudaError_t r = cudaStreamSynchronize(t.st);
        if (r != cudaSuccess) return r;
        const size_t off = k * chunk, m = std::min(chunk, n - off);
        host_copy(out + off * out_elt, t.h_out, m * out_elt);
        return cudaSuccess;
    };
    for (size_t k = 0; k < n_chunks && e == cudaSuccess; k++) {
        akb_scene::TraceSlot &t = sc->trace_slot[k & 1];
        if (k >= 2) e = drain(k - 2);
        if (e != cudaSuccess) break;
        const size_t off = k * chunk, m = std::min(chunk, n - off);
        host_copy(t.h_in, org + 4 * off, m * sizeof(float4));
        host_copy(t.h_in + chunk * sizeof(float4), dir + 4 * off, m * sizeof(float4));
        e = cudaMemcpyAsync(t.d_o, t.h_in, m * sizeof(float4), cudaMemcpyHostToDevice, t.st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(t.d_d, t.h_in + chunk * sizeof(float4), m * sizeof(float4), cudaMemcpyHostToDevice, t.st);
        if (e == cudaSuccess) e = launch_trace_batch(sc->dev, t.d_o, t.d_d, m, any_hit, (akb_hit *)t.d_out, (int32_t *)t.d_out, t.cursor, g_n_sms, t.st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(t.h_out, t.d_out, m * out_elt, cudaMemcpyDeviceToHost, t.st);
    }
    for (size_t k = n_chunks >= 2 ? n_chunks - 2 : 0; k < n_chunks && e == cudaSuccess; k++) e = drain(k);
    if (e != cudaSuccess) {
        cudaStreamSynchronize(sc->trace_slot[0].st), cudaStreamSynchronize(sc->trace_slot[1].st);
        if (e == cudaErrorMemoryAllocation) sc->free_trace();
        return set_err(e == cudaErrorMemoryAllocation ? AKB_OOM : AKB_CUDA_ERROR, "akb_trace: %s", cudaGetErrorString(e));
    }
    return AKB_OK;
}

akb_status akb_generate_rays(akb_scene *sc, const int32_t *xy, const float *u2, size_t n, float *org, float *dir) {
    akb_status st = check_ready("akb_generate_rays");
    if (st != AKB_OK) return st;
    if (!sc || (n && (!xy || !u2 || !org || !dir))) return set_err(AKB_INVALID_ARG, "akb_generate_rays: null argument");
    if (n == 0) return AKB_OK;
    int32_t *d_xy = nullptr;
    float *d_u = nullptr;
    float4 *d_o = nullptr, *d_d = nullptr;
    cudaError_t e = cudaMalloc((void **)&d_xy, n * 8);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_u, n * 8);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_o, n * 16);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_d, n * 16);
    if (e == cudaSuccess) e = cudaMemcpy(d_xy, xy, n * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_u, u2, n * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = launch_generate_rays(sc->dev, d_xy, d_u, n, d_o, d_d, 0);
    if (e == cudaSuccess) e = cudaMemcpy(org, d_o, n * 16, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(dir, d_d, n * 16, cudaMemcpyDeviceToHost);
    cudaFree(d_xy), cudaFree(d_u), cudaFree(d_o), cudaFree(d_d);
    if (e != cudaSuccess) return set_err(AKB_CUDA_ERROR, "akb_generate_rays: %s", cudaGetErrorString(e));
    return AKB_OK;
}

akb_status akb_film_to_srgb8(const float *rgbw, int32_t width, int32_t height, uint8_t *out_rgb8) {
    akb_status st = check_ready("akb_film_to_srgb8");
    if (st != AKB_OK) return st;
    if (!rgbw || !out_rgb8 || width <= 0 || height <= 0) return set_err(AKB_INVALID_ARG, "akb_film_to_srgb8: bad argument");
    const size_t n = (size_t)width * height;
    float4 *d_in = nullptr;
    uint8_t *d_out = nullptr;
    cudaError_t e = cudaMalloc((void **)&d_in, n * 16);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_out, n * 3);
    if (e == cudaSuccess) e = cudaMemcpy(d_in, rgbw, n * 16, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = launch_film_to_srgb8(d_in, n, d_out, 0);
    if (e == cudaSuccess) e = cudaMemcpy(out_rgb8, d_out, n * 3, cudaMemcpyDeviceToHost);
    cudaFree(d_in), cudaFree(d_out);
    if (e != cudaSuccess) return set_err(AKB_CUDA_ERROR, "akb_film_to_srgb8: %s", cudaGetErrorString(e));
    return AKB_OK;
}

} // extern "C"
