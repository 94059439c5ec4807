// api.cu — the C ABI of libakari_b200.so (include/akari_b200.h): device/scene lifecycle, HBM residency, the render
// driver (wavefront bounce loop or megakernel) on one GPU or split over several GPUs of this process with one film
// collective, and the batch ray-query entry points. No CPU fallback: every compute entry fails with AKB_NO_DEVICE /
// AKB_CUDA_ERROR when the GPU path is unavailable. The library never reads the environment (akb_set_option).
#include <dlfcn.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "bvh_build.cuh"
#include "kernels.cuh"
#include "options.h"

using namespace akb;

namespace {

thread_local std::string g_err;
constexpr size_t kMaxBatch = (size_t)1 << 30;
constexpr int kMaxDevices = 16;
constexpr int kMaxPipes = 4; // wavefront pipelines per GPU (option pipelines)

// A CUDA device this process was bound to by akb_init / akb_init_device.
struct Device {
    int index = -1;
    int n_sms = 0;
    bool peers_ok = false; // every other bound device maps this one's memory (NVLink / PCIe peer access)
};
std::vector<Device> g_devs;

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

// NVTX ranges around the entry points and the per-GPU work (the reference wraps every launch in one,
// kernel/cuda/launch.h:74-79). Header-only NVTX3: a no-op unless a profiler injects itself.
struct NvtxRange {
    bool on;
    explicit NvtxRange(const char *name) : on(options().nvtx != 0) {
        if (on) nvtxRangePushA(name);
    }
    ~NvtxRange() {
        if (on) nvtxRangePop();
    }
};

// ---- NCCL, loaded at run time (libnccl.so.2: the one already in the process when the host is a torch program, the
// system's otherwise), so that a single-GPU host needs no NCCL at all. Only the five entry points the film combine uses.
struct NcclApi {
    void *lib = nullptr;
    typedef struct ncclComm *comm_t;
    int (*CommInitAll)(comm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(comm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, comm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
    std::string why;
    bool ok() const { return lib != nullptr; }
};
constexpr int kNcclFloat32 = 7, kNcclSum = 0; // ncclDataType_t / ncclRedOp_t values of nccl.h (stable since 2.0)

NcclApi &nccl() {
    static NcclApi api = [] {
        NcclApi a;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            a.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (a.lib) break;
        }
        if (!a.lib) {
            a.why = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "not found");
            return a;
        }
        bool all = true;
        auto sym = [&](const char *n) {
            void *p = dlsym(a.lib, n);
            if (!p) all = false, a.why = std::string("libnccl lacks ") + n;
            return p;
        };
        a.CommInitAll = (decltype(a.CommInitAll))sym("ncclCommInitAll");
        a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
        a.GroupStart = (decltype(a.GroupStart))sym("ncclGroupStart");
        a.GroupEnd = (decltype(a.GroupEnd))sym("ncclGroupEnd");
        a.Reduce = (decltype(a.Reduce))sym("ncclReduce");
        a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
        a.GetVersion = (decltype(a.GetVersion))sym("ncclGetVersion");
        if (!all) a.lib = nullptr;
        return a;
    }();
    return api;
}

// One communicator per bound device, created on the first multi-GPU frame (ncclCommInitAll: single process).
std::vector<NcclApi::comm_t> g_comms;

void destroy_comms() {
    if (!g_comms.empty() && nccl().ok())
        for (auto c : g_comms)
            if (c) nccl().CommDestroy(c);
    g_comms.clear();
}

} // namespace

namespace akb {
akb_status set_error(akb_status st, const std::string &msg) { // used by host/frontend_api.cpp
    g_err = msg;
    return st;
}
} // namespace akb

// Everything a scene keeps on ONE GPU: the flattened scene, the render workspace, queues' counters, streams, graphs.
struct Replica {
    Device dev;
    SceneDev S;
    std::vector<void *> allocs;
    // render workspace (lazily sized)
    Workspace ws;
    std::vector<void *> ws_allocs;
    int ws_slots = 0;
    Counters *d_ctr = nullptr, *h_ctr = nullptr; // [kMaxPipes]: one per pipeline
    cudaStream_t main_stream = nullptr;          // multi-GPU frames and host-buffer entry points run here
    cudaStream_t aux_stream[kMaxPipes - 1] = {}; // the further pipelines' streams
    cudaEvent_t ev_fork = nullptr, ev_join[kMaxPipes - 1] = {}, ev_poll[kMaxPipes] = {};
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
    } bounce_graph[kMaxPipes];
    cudaStream_t cap_stream = nullptr;
    // streaming wavefront (option stream): the two rings and their control block, sized with the workspace
    uint32_t *r_ring = nullptr;
    uint4 *q_ring = nullptr;
    int ring_shift = 0;
    StreamCtl *d_stream = nullptr, *h_stream = nullptr;
    // pinned staging of the film for akb_render's device-to-host copy
    char *h_film = nullptr;
    size_t h_film_bytes = 0;
    std::vector<cudaEvent_t> ev_chunk;

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
        free_rings();
    }
    void free_rings() {
        if (r_ring) cudaFree(r_ring);
        if (q_ring) cudaFree(q_ring);
        r_ring = nullptr, q_ring = nullptr, ring_shift = 0;
    }
    ~Replica() {
        if (dev.index >= 0) cudaSetDevice(dev.index);
        free_graphs();
        if (cap_stream) cudaStreamDestroy(cap_stream);
        free_ws();
        free_trace();
        for (void *p : allocs) cudaFree(p);
        if (d_ctr) cudaFree(d_ctr);
        if (h_ctr) cudaFreeHost(h_ctr);
        if (d_stream) cudaFree(d_stream);
        if (h_stream) cudaFreeHost(h_stream);
        if (d_film) cudaFree(d_film);
        if (h_film) cudaFreeHost(h_film);
        for (cudaEvent_t e : ev_chunk) cudaEventDestroy(e);
        if (main_stream) cudaStreamDestroy(main_stream);
        for (cudaStream_t a : aux_stream)
            if (a) cudaStreamDestroy(a);
        if (ev_fork) cudaEventDestroy(ev_fork);
        for (cudaEvent_t e : ev_join)
            if (e) cudaEventDestroy(e);
        for (cudaEvent_t e : ev_poll)
            if (e) cudaEventDestroy(e);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
    }
};

struct akb_scene {
    std::vector<std::unique_ptr<Replica>> rep; // rep[i] lives on g_devs[i]
    uint64_t n_tris = 0, n_nodes = 0, bvh_bytes = 0, n_lights = 0, tri_bytes = 0, bvh_depth = 0, build_us = 0, built_by = 0,
             upload_us = 0;
    int width = 0, height = 0;
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

template <class T> akb_status upload(Replica *r, const T *src, size_t n, const T **dst) {
    void *p = nullptr;
    size_t bytes = std::max<size_t>(n * sizeof(T), 16);
    CU(cudaMalloc(&p, bytes));
    r->allocs.push_back(p);
    if (n) CU(cudaMemcpy(p, src, n * sizeof(T), cudaMemcpyHostToDevice));
    *dst = (const T *)p;
    return AKB_OK;
}

template <class T> akb_status ws_alloc(Replica *r, size_t n, T **dst) {
    void *p = nullptr;
    CU(cudaMalloc(&p, std::max<size_t>(n * sizeof(T), 16)));
    r->ws_allocs.push_back(p);
    *dst = (T *)p;
    return AKB_OK;
}

akb_status ensure_workspace(Replica *r, int n_slots) {
    if (r->ws_slots >= n_slots && r->ws_slots > 0) {
        r->ws.n_slots = n_slots;
        return AKB_OK;
    }
    r->free_ws();
    r->free_graphs(); // they hold the old workspace's pointers
    Workspace &w = r->ws;
    const size_t n = (size_t)std::max(n_slots, 32);
    akb_status st;
#define WS(field, type)                                                                                                \
    if ((st = ws_alloc<type>(r, n, &w.field)) != AKB_OK) return st;
    WS(ray_o, float4) WS(ray_d, float4) WS(hit, float4) WS(beta, float4) WS(L, float4) WS(rng, unsigned long long)
    WS(sh_o, float4) WS(sh_d, float4) WS(sh_c, float4) WS(sh_b, float4) WS(acc, float4)
    for (int b = 0; b < 2; b++) {
        WS(q_ext[b], uint32_t) WS(q_sh[b], uint32_t)
        for (int k = 0; k < kNumKeys; k++) { WS(q_shade[b][k], uint32_t) }
    }
#undef WS
    w.ctr = r->d_ctr;
    w.slot_base = 0;
    w.n_slots = n_slots;
    r->ws_slots = n_slots;
    return AKB_OK;
}

akb_status ensure_film(Replica *r, size_t n_px) {
    if (r->d_film_px >= n_px) return AKB_OK;
    if (r->d_film) cudaFree(r->d_film);
    r->d_film = nullptr, r->d_film_px = 0;
    CU(cudaMalloc((void **)&r->d_film, std::max<size_t>(n_px, 1) * sizeof(float4)));
    r->d_film_px = n_px;
    return AKB_OK;
}

// Rings of the streaming wavefront: twice the next power of two of the slot count (a path has one entry in flight, so
// half a lap always separates a producer from the entries still being read).
akb_status ensure_rings(Replica *r, int n_slots) {
    int shift = 19; // at least 2^20 entries: a lap must take much longer than any reader sits on a ticket (kernels.cu)
    while ((1 << shift) < n_slots) shift++;
    shift++;
    if (!r->d_stream) {
        CU(cudaMalloc((void **)&r->d_stream, sizeof(StreamCtl)));
        CU(cudaMallocHost((void **)&r->h_stream, sizeof(StreamCtl)));
    }
    if (r->ring_shift == shift && r->r_ring) return AKB_OK;
    r->free_rings();
    CU(cudaMalloc((void **)&r->r_ring, sizeof(uint32_t) << shift));
    CU(cudaMalloc((void **)&r->q_ring, sizeof(uint4) << shift));
    r->ring_shift = shift;
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
    if (g_devs.empty())
        return set_err(AKB_NO_DEVICE, "%s: akb_init() has not succeeded (no CUDA device bound; there is no CPU fallback)", who);
    return AKB_OK;
}

// Binds one device: architecture check, arithmetic self-test, the kernels' constant-memory switches.
akb_status bind_device(int index, Device *out) {
    CU(cudaSetDevice(index));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, index));
    if (prop.major < 10)
        return set_err(AKB_NO_DEVICE, "akb_init: device %d is sm_%d%d; this library carries sm_100a code only", index, prop.major, prop.minor);
    // arithmetic self-test: the parity contract needs un-contracted fp32 (see vecmath.h)
    int *d_res = nullptr, h_res = 0;
    CU(cudaMalloc(&d_res, sizeof(int)));
    cudaError_t e = launch_selftest(d_res, 0);
    if (e == cudaSuccess) e = cudaMemcpy(&h_res, d_res, sizeof(int), cudaMemcpyDeviceToHost);
    cudaFree(d_res);
    if (e != cudaSuccess) return set_err(AKB_CUDA_ERROR, "akb_init: self-test launch failed on device %d: %s", index, cudaGetErrorString(e));
    if (h_res != 1) return set_err(AKB_CUDA_ERROR, "akb_init: kernels were built with FMA contraction; rebuild with -fmad=false");
    out->index = index, out->n_sms = prop.multiProcessorCount, out->peers_ok = false;
    return AKB_OK;
}

// Pushes the kernel-side switches (constant memory) of the current options to every bound device.
akb_status push_device_options() {
    const Options &o = options();
    for (const Device &d : g_devs) {
        CU(cudaSetDevice(d.index));
        CU(set_refill_below(o.refill_below));
        CU(set_no_park(o.no_park));
        CU(set_shade_partition(o.shade_partition));
        CU(set_tail_share(o.tail_share, o.tail_share_min));
        CU(set_fetch_chunk(o.fetch_chunk));
        CU(set_stream_options(o.stream_flex_mask, o.stream_watchdog_ms));
    }
    if (!g_devs.empty()) CU(cudaSetDevice(g_devs[0].index));
    return AKB_OK;
}

akb_status init_devices(const std::vector<int> &indices) {
    destroy_comms();
    g_devs.clear();
    std::vector<Device> devs(indices.size());
    for (size_t i = 0; i < indices.size(); i++) {
        akb_status st = bind_device(indices[i], &devs[i]);
        if (st != AKB_OK) return st;
    }
    // peer access between every pair (NVLink on an NVSwitch box): lets one GPU's kernels write another's film
    for (size_t i = 0; i < devs.size(); i++) {
        bool all = devs.size() > 1;
        for (size_t j = 0; j < devs.size(); j++) {
            if (i == j) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, devs[j].index, devs[i].index) != cudaSuccess || !can) {
                all = false;
                continue;
            }
            cudaSetDevice(devs[j].index);
            cudaError_t e = cudaDeviceEnablePeerAccess(devs[i].index, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) all = false, cudaGetLastError();
        }
        devs[i].peers_ok = all;
    }
    g_devs = devs;
    return push_device_options();
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

int akb_n_devices(void) { return (int)g_devs.size(); }

akb_status akb_init(int n_devices) {
    const int n = akb_device_count();
    if (n <= 0) return set_err(AKB_NO_DEVICE, "akb_init: no CUDA device visible (this backend has no CPU fallback)");
    if (n_devices < 0 || n_devices > kMaxDevices) return set_err(AKB_INVALID_ARG, "akb_init: n_devices %d out of range [0,%d]", n_devices, kMaxDevices);
    if (n_devices == 0) n_devices = std::min(n, kMaxDevices);
    if (n_devices > n) return set_err(AKB_NO_DEVICE, "akb_init: %d devices asked for, %d visible", n_devices, n);
    std::vector<int> idx(n_devices);
    for (int i = 0; i < n_devices; i++) idx[i] = i;
    return init_devices(idx);
}

akb_status akb_init_device(int device_index) {
    const int n = akb_device_count();
    if (n <= 0) return set_err(AKB_NO_DEVICE, "akb_init_device: no CUDA device visible (this backend has no CPU fallback)");
    if (device_index < 0 || device_index >= n) return set_err(AKB_INVALID_ARG, "akb_init_device: device %d out of range [0,%d)", device_index, n);
    return init_devices({device_index});
}

akb_status akb_shutdown(void) {
    for (const Device &d : g_devs) {
        cudaSetDevice(d.index);
        cudaDeviceSynchronize();
    }
    destroy_comms();
    g_devs.clear();
    return AKB_OK;
}

akb_status akb_set_option(const char *name, const char *value) {
    std::string err;
    if (!set_option(name, value, &err)) return set_err(AKB_INVALID_ARG, "%s", err.c_str());
    if (!g_devs.empty()) return push_device_options();
    return AKB_OK;
}

namespace {
// AKB_BVH_DEVICE: csrc/bvh_build.cu, with the host builder as the fallback for trees deeper than the stack and - when
// the device builder was the library's own choice - for any failure of the device build (e.g. no room for its scratch)
bool g_device_builder_forced = false;
bool device_bvh_builder(const float *lo, const float *hi, size_t n, float pad, int n_threads, Bvh8 *out, std::string *err) {
    const auto t0 = std::chrono::steady_clock::now();
    float node_cost, prim_cost;
    int max_leaf;
    bvh8_collapse_params(&node_cost, &prim_cost, &max_leaf);
    bool fallback = false;
    DeviceBuildTimes times;
    const cudaError_t e = build_bvh8_device(lo, hi, n, pad, node_cost, prim_cost, max_leaf, out, &fallback, &times);
    if (e != cudaSuccess) {
        cudaGetLastError();
        if (g_device_builder_forced) {
            *err = std::string(e == cudaErrorMemoryAllocation ? "out of device memory: " : "") + "device BVH build: " + cudaGetErrorString(e);
            return false;
        }
        fallback = true;
    }
    if (fallback) {
        build_bvh8(lo, hi, n, pad, n_threads, out);
        return true;
    }
    out->built_by = AKB_BVH_DEVICE;
    out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (options().bvh_verbose)
        fprintf(stderr, "[akb] device BVH build: %zu triangles -> %zu nodes, depth %d (%s, %d rounds); upload %.2f sort %.2f tree %.2f fit %.2f emit %.2f download %.2f ms\n",
                n, out->nodes.size(), out->depth, times.ploc_rounds ? "PLOC" : "radix tree", times.ploc_rounds, times.ms_upload, times.ms_sort, times.ms_tree, times.ms_fit, times.ms_emit, times.ms_download);
    return true;
}

// Uploads the flattened scene to one GPU and creates that GPU's streams / events / counters.
akb_status make_replica(const HostScene &hs, const Device &dev, Replica *r) {
    CU(cudaSetDevice(dev.index));
    r->dev = dev;
    SceneDev &S = r->S;
    memset(&S, 0, sizeof S);
    akb_status st = AKB_OK;
    { // BVH nodes and leaf triangles share one allocation (the traversal working set: 56 MB for 1 M triangles)
        const size_t node_f4 = hs.bvh.nodes.size() * 5, node_pad = (node_f4 + 15) & ~(size_t)15, tri_f4 = hs.tris.size();
        float4 *buf = nullptr;
        const size_t bytes = std::max<size_t>((node_pad + tri_f4) * sizeof(float4), 16);
        cudaError_t e = cudaMalloc((void **)&buf, bytes);
        if (e == cudaSuccess) r->allocs.push_back(buf);
        if (e == cudaSuccess && node_f4) e = cudaMemcpy(buf, hs.bvh.nodes.data(), node_f4 * sizeof(float4), cudaMemcpyHostToDevice);
        if (e == cudaSuccess && tri_f4) e = cudaMemcpy(buf + node_pad, hs.tris.data(), tri_f4 * sizeof(float4), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) st = set_err(e == cudaErrorMemoryAllocation ? AKB_OOM : AKB_CUDA_ERROR, "akb_scene_create: %s", cudaGetErrorString(e));
        S.nodes = buf, S.tris = buf + node_pad;
    }
    if (st == AKB_OK) st = upload(r, (const float4 *)hs.shading.data(), hs.shading.size(), &S.shading);
    if (st == AKB_OK) st = upload(r, hs.mats.data(), hs.mats.size(), &S.mats);
    if (st == AKB_OK) st = upload(r, (const float4 *)hs.lights.data(), hs.lights.size(), &S.lights);
    if (st == AKB_OK) st = upload(r, hs.light_cdf.data(), hs.light_cdf.size(), &S.light_cdf);
    if (st == AKB_OK) st = upload(r, hs.light_func.data(), hs.light_func.size(), &S.light_func);
    if (st == AKB_OK && !hs.tex_info.empty()) { // image textures: texels, per-texture info, per-triangle / per-light texcoords
        st = upload(r, (const float4 *)hs.texels.data(), hs.texels.size(), &S.texels);
        if (st == AKB_OK) st = upload(r, (const int4 *)hs.tex_info.data(), hs.tex_info.size() / 4, &S.tex_info);
        if (st == AKB_OK) st = upload(r, (const float4 *)hs.tri_tc.data(), hs.tri_tc.size(), &S.tri_tc);
        if (st == AKB_OK) st = upload(r, (const float4 *)hs.light_tc.data(), hs.light_tc.size(), &S.light_tc);
    }
    if (st != AKB_OK) return st;
    const bool timeline = options().timeline != 0;
    CU(cudaMalloc((void **)&r->d_ctr, kMaxPipes * sizeof(Counters)));
    CU(cudaMallocHost((void **)&r->h_ctr, kMaxPipes * sizeof(Counters)));
    CU(cudaStreamCreateWithFlags(&r->main_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&r->ev_fork, cudaEventDisableTiming));
    for (int k = 0; k + 1 < kMaxPipes; k++) {
        CU(cudaStreamCreateWithFlags(&r->aux_stream[k], cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&r->ev_join[k], cudaEventDisableTiming));
    }
    CU(cudaStreamCreateWithFlags(&r->cap_stream, cudaStreamNonBlocking));
    for (int k = 0; k < kMaxPipes; k++) CU(cudaEventCreateWithFlags(&r->ev_poll[k], timeline ? cudaEventDefault : cudaEventDisableTiming));
    CU(cudaEventCreate(&r->ev0));
    CU(cudaEventCreate(&r->ev1));
    S.light_func_int = hs.light_func_int;
    S.n_lights = (int)hs.light_func.size();
    S.n_mats = (int)hs.mats.size();
    // if-if traversal for real trees, fused step for tiny ones (option trav_ifif = 0/1 forces either; tuning aid)
    {
        const int f = options().trav_ifif;
        const bool ifif = f >= 0 ? f == 1 : hs.bvh.nodes.size() >= 64;
        S.trav_post_limit = ifif ? std::max(0, kStackSize - hs.bvh.depth - 1) : -1;
    }
    S.n_textures = (int)(hs.tex_info.size() / 4);
    S.n_nodes = (uint32_t)hs.bvh.nodes.size(), S.n_tris = (uint32_t)(hs.tris.size() / 3);
    S.cam = hs.cam;
    return AKB_OK;
}
} // namespace

akb_status akb_scene_create(const akb_scene_desc *desc, akb_scene **out_scene) {
    if (!out_scene) return set_err(AKB_INVALID_ARG, "akb_scene_create: out_scene == NULL");
    *out_scene = nullptr;
    akb_status st = check_ready("akb_scene_create");
    if (st != AKB_OK) return st;
    NvtxRange nvtx_range("akb_scene_create");
    CU(cudaSetDevice(g_devs[0].index));
    HostScene hs;
    std::string err;
    int build = desc ? (int)desc->bvh_build : 0;
    g_device_builder_forced = build == AKB_BVH_DEVICE;
    if (build == AKB_BVH_DEFAULT && options().bvh_build != AKB_BVH_DEFAULT) build = options().bvh_build, g_device_builder_forced = build == AKB_BVH_DEVICE;
    if (build == AKB_BVH_DEFAULT) {
        // host SAH for scenes it builds in about a second, the device build above that (measured on B200: C3's 1 M
        // triangles 0.43 s vs 0.03 s to build but 1 % slower to render (C5: 2 %); C4's 10 M: 4.4 s vs 0.08 s, 2.4 %
        // slower; a 1 M-triangle clustered soup: 0.9 s vs 0.03 s, 1-8 % FASTER to traverse)
        uint64_t nt = 0;
        if (desc && desc->meshes)
            for (uint32_t g = 0; g < desc->n_meshes; g++) nt += desc->meshes[g].n_triangles;
        build = nt >= (4u << 20) ? AKB_BVH_DEVICE : AKB_BVH_HOST_SAH;
    }
    if (build != AKB_BVH_HOST_SAH && build != AKB_BVH_DEVICE) return set_err(AKB_INVALID_ARG, "akb_scene_create: bad bvh_build %d", build);
    const auto t_create = std::chrono::steady_clock::now();
    // ONE flatten + BVH build (host threads, or device 0), then the same arrays go to every bound GPU
    st = flatten_scene(desc, 0, &hs, &err, build == AKB_BVH_DEVICE ? device_bvh_builder : nullptr);
    if (st != AKB_OK) return set_err(err.rfind("out of device memory", 0) == 0 ? AKB_OOM : st, "%s", err.c_str());
    const auto t_flat = std::chrono::steady_clock::now();
    std::unique_ptr<akb_scene> sc(new akb_scene());
    const size_t nd = g_devs.size();
    sc->rep.resize(nd);
    for (size_t i = 0; i < nd; i++) sc->rep[i].reset(new Replica());
    std::vector<akb_status> rst(nd, AKB_OK);
    std::vector<std::string> rerr(nd);
    auto up = [&](size_t i) {
        rst[i] = make_replica(hs, g_devs[i], sc->rep[i].get());
        if (rst[i] != AKB_OK) rerr[i] = g_err;
    };
    { // every GPU has its own PCIe link: upload the replicas side by side
        std::vector<std::thread> th;
        for (size_t i = 1; i < nd; i++) th.emplace_back(up, i);
        up(0);
        for (auto &t : th) t.join();
    }
    cudaSetDevice(g_devs[0].index);
    for (size_t i = 0; i < nd; i++)
        if (rst[i] != AKB_OK) return set_err(rst[i], "%s (device %d)", rerr[i].c_str(), g_devs[i].index);
    const auto t_up = std::chrono::steady_clock::now();
    if (options().bvh_verbose)
        fprintf(stderr, "[akb] scene create: flatten + BVH %.3f s (BVH %.3f s), upload to %zu GPU(s) %.3f s\n", std::chrono::duration<double>(t_flat - t_create).count(),
                hs.bvh.build_seconds, nd, std::chrono::duration<double>(t_up - t_flat).count());
    sc->n_tris = hs.n_tris, sc->n_nodes = hs.bvh.nodes.size(), sc->bvh_bytes = hs.bvh.nodes.size() * sizeof(Node8);
    sc->n_lights = hs.light_func.size(), sc->tri_bytes = hs.tris.size() * sizeof(f4);
    sc->bvh_depth = (uint64_t)hs.bvh.depth, sc->build_us = (uint64_t)(hs.bvh.build_seconds * 1e6);
    sc->built_by = (uint64_t)hs.bvh.built_by;
    sc->upload_us = (uint64_t)(std::chrono::duration<double>(t_up - t_flat).count() * 1e6);
    sc->width = hs.cam.width, sc->height = hs.cam.height;
    *out_scene = sc.release();
    return AKB_OK;
}

akb_status akb_scene_destroy(akb_scene *scene) {
    if (!scene) return AKB_OK;
    for (auto &r : scene->rep)
        if (r && r->dev.index >= 0) {
            cudaSetDevice(r->dev.index);
            cudaDeviceSynchronize();
        }
    delete scene;
    if (!g_devs.empty()) cudaSetDevice(g_devs[0].index);
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
    case 8: return sc->rep.size();
    case 9: return sc->upload_us;
    }
    return 0;
}

} // extern "C"

namespace {

akb_status validate_params(const char *who, const akb_render_params *p, int *tile_out) {
    if (p->integrator != AKB_INTEGRATOR_PATH && p->integrator != AKB_INTEGRATOR_AO) return set_err(AKB_INVALID_ARG, "%s: unknown integrator %d", who, p->integrator);
    if (p->integrator == AKB_INTEGRATOR_AO && !(p->ao_occlude > 0.0f)) return set_err(AKB_INVALID_ARG, "%s: AO.occlude must be positive", who);
    if (p->spp < 0 || p->max_depth < 0) return set_err(AKB_INVALID_ARG, "%s: negative spp/max_depth", who);
    if (p->sampler != AKB_SAMPLER_LCG && p->sampler != AKB_SAMPLER_PCG) return set_err(AKB_INVALID_ARG, "%s: bad sampler", who);
    const int world = p->world_size > 0 ? p->world_size : 1;
    if (p->rank < 0 || p->rank >= world) return set_err(AKB_INVALID_ARG, "%s: rank %d outside world_size %d", who, p->rank, world);
    const int tile = p->tile_size > 0 ? p->tile_size : 32;
    if ((tile & (tile - 1)) != 0 || tile < 8 || tile > 1024) return set_err(AKB_INVALID_ARG, "%s: tile_size must be a power of two in [8,1024]", who);
    if (p->split != AKB_SPLIT_TILE && p->split != AKB_SPLIT_SAMPLE) return set_err(AKB_INVALID_ARG, "%s: bad split %d", who, p->split);
    if (p->n_gpus < 0) return set_err(AKB_INVALID_ARG, "%s: negative n_gpus", who);
    *tile_out = tile;
    return AKB_OK;
}

// One GPU's share of a frame (the whole frame for rank 0 of 1): enqueues the wavefront bounce loop (or the megakernel)
// on `stream` of the replica's device, waits for it and fills `stats`. `film` may live on ANOTHER device (peer
// memory): with film_is_shared the share's pixels are written straight into it and nothing else of it is touched.
akb_status render_on_replica(Replica *r, int width, int height, const akb_render_params *p, int rank, int world, float4 *film,
                             bool film_is_shared, cudaStream_t stream, akb_render_stats *stats) {
    const Options &opt = options();
    CU(cudaSetDevice(r->dev.index));
    CU(set_debug_bounds(r->S.n_nodes, r->S.n_tris));
    const int n_sms = r->dev.n_sms;
    akb_render_params q = *p;
    q.rank = rank, q.world_size = world;
    RenderConfig rc;
    memset(&rc, 0, sizeof rc); // compared bytewise with the cached graphs' copy
    const Shard shard = make_shard(width, height, &q);
    rc.spp = shard.spp, rc.max_depth = p->max_depth, rc.rr_start = p->rr_start_depth, rc.sampler = p->sampler;
    rc.sort_materials = p->sort_materials, rc.ray_clamp = p->ray_clamp;
    rc.seed_offset = shard.seed_offset;
    rc.map = shard.map;
    rc.ao = p->integrator == AKB_INTEGRATOR_AO, rc.ao_occlude = p->ao_occlude;
    if (rc.ao) rc.sort_materials = 0, rc.ray_clamp = 0.0f; // AO: no materials, samples are exact 0/1
    const size_t n_px = (size_t)width * height;
    uint64_t launches = 0;
    bool streamed = false;
    std::vector<cudaEvent_t> evs;
    const bool profile_stages = opt.profile_stages != 0;
    if (!film_is_shared) CU(cudaMemsetAsync(film, 0, n_px * sizeof(float4), stream));
    CU(cudaMemsetAsync(r->d_ctr, 0, kMaxPipes * sizeof(Counters), stream));
    CU(cudaEventRecord(r->ev0, stream));
    if (p->integrator == AKB_INTEGRATOR_AO && p->mode == AKB_MODE_MEGAKERNEL) {
        CU(launch_ao(r->S, rc, p->ao_occlude, film, r->d_ctr, stream));
        launches = 1;
    } else if (p->mode == AKB_MODE_MEGAKERNEL) {
        CU(launch_megakernel(r->S, rc, film, r->d_ctr, stream));
        launches = 1;
    } else if (opt.stream && !rc.ao && !rc.sort_materials && !profile_stages && rc.map.n_slots() <= kStreamMaxSlots && rc.spp > 0 &&
               rc.map.n_slots() > 0) {
        // Streaming wavefront: one persistent launch, no barrier between bounces (kernels.cu)
        const int n_total = rc.map.n_slots();
        akb_status st = ensure_workspace(r, n_total);
        if (st == AKB_OK) st = ensure_rings(r, n_total);
        if (st != AKB_OK) return st;
        Workspace w = workspace_view(r->ws, 0, n_total, r->d_ctr);
        w.r_ring = r->r_ring, w.q_ring = r->q_ring, w.ring_shift = r->ring_shift, w.sc = r->d_stream;
        CU(cudaMemsetAsync(r->r_ring, 0xFF, sizeof(uint32_t) << r->ring_shift, stream)); // lap stamp 127: nothing there yet
        CU(cudaMemsetAsync(r->q_ring, 0xFF, sizeof(uint4) << r->ring_shift, stream));
        CU(cudaMemsetAsync(r->d_stream, 0, sizeof(StreamCtl), stream));
        CU(launch_wf_init(r->S, rc, w, n_sms, stream));
        CU(launch_wf_stream(r->S, rc, w, n_sms, stream));
        CU(launch_wf_resolve(rc, w, film, stream));
        launches = 4;
        streamed = true;
    } else {
        const int n_total = rc.map.n_slots();
        akb_status st = ensure_workspace(r, n_total);
        if (st != AKB_OK) return st;
        // Up to four pipelines on as many streams (default two), each a contiguous range of the owned tiles with its own
        // queues and counters: while one pipeline's traversal launch drains (persistent warps running their last,
        // longest rays at a few lanes each) the others' kernels fill the machine. Every slot keeps its pixel's seed and
        // accumulation order: the film is unchanged. Measured on C3 (profiles/r02_notes.md section 10): 1 / 2 / 3 / 4
        // pipelines 113.6 / 111.2 / 114.2 / 109.9 ms; dealing the tiles out in turn (option pipeline_interleave), so that
        // all pipelines hold the same mix of sky and ground pixels, makes two pipelines as slow as one (113.5 ms): the
        // pipelines then move in step, and what two pipelines win comes from their being out of step.
        const int tt = rc.map.tile * rc.map.tile;
        int n_pipes = (!profile_stages && n_total >= opt.pipeline_min_slots) ? std::min(opt.pipelines, rc.map.n_owned_tiles) : 1;
        n_pipes = std::max(1, std::min(n_pipes, kMaxPipes));
        rc.map.groups = opt.pipeline_interleave ? n_pipes : 1;
        struct Pipe {
            Workspace ws;
            cudaStream_t st;
            bool live;
        } pipe[kMaxPipes];
        for (int k = 0, base = 0; k < n_pipes; k++) {
            // (contiguous ranges: the same sizes, so that both settings cut the workspace alike)
            const int n_k = ((rc.map.n_owned_tiles - k + n_pipes - 1) / n_pipes) * tt;
            pipe[k] = Pipe{workspace_view(r->ws, base, n_k, r->d_ctr + k), k == 0 ? stream : r->aux_stream[k - 1], n_k > 0 && rc.spp > 0};
            base += n_k;
        }
        for (int k = n_pipes; k < kMaxPipes; k++) pipe[k].live = false;
        if (n_pipes > 1) {
            CU(cudaEventRecord(r->ev_fork, stream));
            for (int k = 1; k < n_pipes; k++) CU(cudaStreamWaitEvent(pipe[k].st, r->ev_fork, 0));
        }
        for (int k = 0; k < n_pipes; k++) {
            CU(launch_wf_init(r->S, rc, pipe[k].ws, n_sms, pipe[k].st));
            launches++;
        }
        const long max_it = (long)rc.spp * (rc.max_depth + 2) + 2;
        const int check_every = opt.check_every; // even: a snapshot taken after an odd bounce is read after an odd bounce
        bool poll_pending[kMaxPipes] = {};
        auto mark = [&]() {
            if (!profile_stages) return cudaSuccess;
            cudaEvent_t e;
            cudaError_t rr = cudaEventCreate(&e);
            if (rr != cudaSuccess) return rr;
            evs.push_back(e);
            return cudaEventRecord(e, stream);
        };
        CU(mark());
        const bool use_graphs = opt.graphs && !profile_stages;
        const int finish_div = (rc.ao || profile_stages) ? 0 : opt.finish_div; // the AO integrator keeps to its queues
        if (use_graphs)
            for (int k = 0; k < n_pipes; k++) {
                Replica::BounceGraph &g = r->bounce_graph[k];
                if (g.exec && g.n_bounces == check_every && !memcmp(&g.rc, &rc, sizeof rc) && !memcmp(&g.ws, &pipe[k].ws, sizeof(Workspace))) continue;
                if (g.exec) cudaGraphExecDestroy(g.exec);
                g.exec = nullptr;
                CU(cudaStreamBeginCapture(r->cap_stream, cudaStreamCaptureModeThreadLocal));
                cudaError_t ce = cudaSuccess;
                for (int j = 0; j < check_every && ce == cudaSuccess; j++) {
                    ce = launch_wf_trace(r->S, rc, pipe[k].ws, j & 1, n_sms, r->cap_stream);
                    if (ce == cudaSuccess) ce = launch_wf_shade(r->S, rc, pipe[k].ws, j & 1, n_sms, r->cap_stream);
                }
                cudaGraph_t graph = nullptr;
                const cudaError_t ee = cudaStreamEndCapture(r->cap_stream, &graph);
                if (ce == cudaSuccess) ce = ee;
                if (ce == cudaSuccess) ce = cudaGraphInstantiate(&g.exec, graph, 0);
                if (graph) cudaGraphDestroy(graph);
                CU(ce);
                g.rc = rc, g.ws = pipe[k].ws, g.n_bounces = check_every;
            }
        auto any_live = [&]() {
            for (int k = 0; k < n_pipes; k++)
                if (pipe[k].live) return true;
            return false;
        };
        for (long it = 0; it < max_it && any_live(); it++) {
            const int b = (int)(it & 1);
            if (use_graphs) {
                if ((it % check_every) == 0)
                    for (int k = 0; k < n_pipes; k++)
                        if (pipe[k].live) {
                            CU(cudaGraphLaunch(r->bounce_graph[k].exec, pipe[k].st));
                            launches += 2 * check_every;
                        }
            } else {
                for (int k = 0; k < n_pipes; k++)
                    if (pipe[k].live) {
                        CU(launch_wf_trace(r->S, rc, pipe[k].ws, b, n_sms, pipe[k].st));
                        launches++;
                    }
                CU(mark());
                for (int k = 0; k < n_pipes; k++)
                    if (pipe[k].live) {
                        CU(launch_wf_shade(r->S, rc, pipe[k].ws, b, n_sms, pipe[k].st));
                        launches++;
                    }
                CU(mark());
            }
            if ((it % check_every) == check_every - 1) {
                // The queue lengths are read one poll late: the copy issued here is looked at only after the next
                // check_every bounces have been enqueued, so the GPU never idles while the host waits for it. A
                // finished pipeline runs at most check_every more bounces on empty queues (microseconds).
                for (int k = 0; k < n_pipes; k++) {
                    if (!pipe[k].live) continue;
                    if (poll_pending[k] || !opt.poll_lag) {
                        if (!opt.poll_lag) {
                            CU(cudaMemcpyAsync(r->h_ctr + k, r->d_ctr + k, sizeof(Counters), cudaMemcpyDeviceToHost, pipe[k].st));
                            CU(cudaStreamSynchronize(pipe[k].st));
                        } else {
                            CU(cudaEventSynchronize(r->ev_poll[k]));
                        }
                        poll_pending[k] = false;
                        if (opt.timeline) {
                            float ms = 0.f;
                            if (cudaEventElapsedTime(&ms, r->ev0, r->ev_poll[k]) != cudaSuccess) // (option set after the scene was created: no timing events)
                                cudaGetLastError(), ms = -1.f;
                            fprintf(stderr, "akb timeline: gpu %d pipe %d after bounce %ld: %.3f ms, n_ext %u n_sh %u\n", r->dev.index, k, it - check_every,
                                    ms, r->h_ctr[k].n_ext[b ^ 1], r->h_ctr[k].n_sh[b ^ 1]);
                        }
                        const uint32_t left = r->h_ctr[k].n_ext[b ^ 1];
                        if (left == 0 && r->h_ctr[k].n_sh[b ^ 1] == 0) {
                            pipe[k].live = false;
                        } else if (finish_div > 0 && left <= (uint32_t)std::min(pipe[k].ws.n_slots / finish_div, opt.finish_max)) {
                            // few paths left (the count only falls from here on): one thread per path runs them out.
                            // Everything enqueued so far ends on a shade step of parity b, so the live queues are b ^ 1.
                            CU(launch_wf_finish(r->S, rc, pipe[k].ws, b ^ 1, left, pipe[k].st));
                            launches++;
                            pipe[k].live = false;
                        }
                    }
                    if (pipe[k].live && opt.poll_lag) {
                        CU(cudaMemcpyAsync(r->h_ctr + k, r->d_ctr + k, sizeof(Counters), cudaMemcpyDeviceToHost, pipe[k].st));
                        CU(cudaEventRecord(r->ev_poll[k], pipe[k].st));
                        poll_pending[k] = true;
                    }
                }
            }
        }
        for (int k = 1; k < n_pipes; k++) {
            CU(cudaEventRecord(r->ev_join[k - 1], pipe[k].st));
            CU(cudaStreamWaitEvent(stream, r->ev_join[k - 1], 0));
        }
        for (int k = 0; k < n_pipes; k++) {
            // "Film accumulate": slot accumulators -> the film; with a shared film this store goes over NVLink into
            // device 0's memory, which is the whole tile-split "collective" of the p2p combine
            CU(launch_wf_resolve(rc, pipe[k].ws, film, stream));
            launches++;
        }
    }
    CU(cudaEventRecord(r->ev1, stream));
    CU(cudaMemcpyAsync(r->h_ctr, r->d_ctr, kMaxPipes * sizeof(Counters), cudaMemcpyDeviceToHost, stream));
    if (streamed) CU(cudaMemcpyAsync(r->h_stream, r->d_stream, sizeof(StreamCtl), cudaMemcpyDeviceToHost, stream));
    CU(cudaStreamSynchronize(stream));
    if (streamed && (r->h_stream->abort || !r->h_stream->done))
        return set_err(AKB_CUDA_ERROR, "akb_render: the streaming wavefront stopped on its watchdog (%u of %u slots finished)",
                       r->h_stream->n_finished, r->h_stream->n_started);
    if (streamed && opt.timeline)
        fprintf(stderr, "akb stream: gpu %d: %llu shade batches, %llu rays traced by shade-first warps, %llu idle polls\n", r->dev.index, r->h_stream->stat_shade_batches,
                r->h_stream->stat_flex_rays, r->h_stream->stat_idle_polls);
#ifdef AKB_TAIL_PROF
    {
        unsigned long long tp[6];
        if (tail_prof_read(tp, true) == cudaSuccess)
            fprintf(stderr, "akb tail prof: queue live %llu steps at %.1f lanes | drained %llu steps at %.1f lanes | tail_share %llu steps at %.1f lanes\n", tp[0],
                    tp[1] / (double)std::max<unsigned long long>(tp[0], 1), tp[2], tp[3] / (double)std::max<unsigned long long>(tp[2], 1), tp[4],
                    tp[5] / (double)std::max<unsigned long long>(tp[4], 1));
    }
#endif
#ifdef AKB_STREAM_PROF
    if (streamed)
        for (int g = 0; g < 2; g++) {
            const unsigned long long *pr = r->h_stream->prof + 10 * g;
            const double all = (double)std::max<unsigned long long>(pr[0], 1);
            fprintf(stderr, "akb stream prof (%s warps): shade %.1f%% steps %.1f%% refill %.1f%% idle %.1f%% msg %.1f%% | %llu refills, %llu steps at %.1f lanes, %llu batches (%.0f cycles each)\n",
                    g ? "shade-first" : "tracing", 100 * pr[1] / all, 100 * pr[2] / all, 100 * pr[3] / all, 100 * pr[4] / all, 100 * pr[5] / all, pr[6], pr[7],
                    pr[8] / (double)std::max<unsigned long long>(pr[7], 1), pr[9], pr[1] / (double)std::max<unsigned long long>(pr[9], 1));
        }
#endif
    if (stats) {
        memset(stats, 0, sizeof *stats);
        for (int k = 0; k < kMaxPipes && !streamed; k++) {
            stats->n_samples += r->h_ctr[k].stat_samples;
            stats->n_closest_rays += r->h_ctr[k].stat_closest - r->h_ctr[k].stat_skipped;
            stats->n_shadow_rays += r->h_ctr[k].stat_shadow;
        }
        if (streamed) {
            stats->n_samples = r->h_stream->stat_samples;
            stats->n_closest_rays = r->h_stream->stat_closest;
            stats->n_shadow_rays = r->h_stream->stat_shadow;
        }
        stats->n_kernel_launches = launches;
        stats->n_gpus_used = 1;
        CU(cudaEventElapsedTime(&stats->ms_total, r->ev0, r->ev1));
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

// Page-locked host memory (cudaHostAlloc / cudaHostRegister by the caller): the DMA engines reach it directly, so the
// host-buffer entry points copy straight from / into it instead of through the library's own pinned staging.
bool host_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

// Device film -> caller's host buffer through a pinned staging buffer: the copy is cut into chunks, and while chunk k
// crosses PCIe a few host threads move the chunks already landed into the caller's (pageable) memory.
akb_status download_film(Replica *r, float *out, size_t n_px, cudaStream_t stream) {
    const size_t bytes = n_px * sizeof(float4);
    if (bytes == 0) return AKB_OK;
    if (host_pinned(out)) { // the caller's buffer is page-locked: one DMA, no staging
        CU(cudaMemcpyAsync(out, r->d_film, bytes, cudaMemcpyDeviceToHost, stream));
        CU(cudaStreamSynchronize(stream));
        return AKB_OK;
    }
    if (r->h_film_bytes < bytes) {
        if (r->h_film) cudaFreeHost(r->h_film);
        r->h_film = nullptr, r->h_film_bytes = 0;
        CU(cudaMallocHost((void **)&r->h_film, bytes));
        r->h_film_bytes = bytes;
    }
    const size_t chunk = (size_t)4 << 20;
    const size_t n_chunks = (bytes + chunk - 1) / chunk;
    while (r->ev_chunk.size() < n_chunks) {
        cudaEvent_t e;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        r->ev_chunk.push_back(e);
    }
    for (size_t k = 0; k < n_chunks; k++) {
        const size_t off = k * chunk, m = std::min(chunk, bytes - off);
        CU(cudaMemcpyAsync(r->h_film + off, (const char *)r->d_film + off, m, cudaMemcpyDeviceToHost, stream));
        CU(cudaEventRecord(r->ev_chunk[k], stream));
    }
    const int n_threads = (int)std::min<size_t>(4, n_chunks);
    std::vector<cudaError_t> terr(n_threads, cudaSuccess);
    auto work = [&](int t) {
        cudaSetDevice(r->dev.index);
        for (size_t k = t; k < n_chunks; k += n_threads) {
            cudaError_t e = cudaEventSynchronize(r->ev_chunk[k]);
            if (e != cudaSuccess) {
                terr[t] = e;
                return;
            }
            const size_t off = k * chunk, m = std::min(chunk, bytes - off);
            memcpy((char *)out + off, r->h_film + off, m);
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < n_threads; t++) th.emplace_back(work, t);
    work(0);
    for (auto &t : th) t.join();
    for (cudaError_t e : terr) CU(e);
    return AKB_OK;
}

// The frame split over the first n GPUs of this process: one host thread + stream per GPU renders that GPU's share
// (rank i of n: interleaved tiles or a share of the samples), then ONE collective leaves the whole film on device 0.
//   combine = nccl: every GPU renders into its own film (pixels it does not own stay 0) and one ncclReduce(sum) to
//             device 0 adds them: x + 0 = x, so the tile split stays bit-identical to one GPU.
//   combine = p2p : tile split - every GPU's resolve kernel stores its pixels straight into device 0's film over
//             NVLink (disjoint pixels: no reduction at all); sample split - device 0 sums the peers' films in rank
//             order with one kernel reading peer memory.
akb_status render_multi(akb_scene *sc, const akb_render_params *p, int n, akb_render_stats *stats, float *ms_combine) {
    const size_t n_px = (size_t)sc->width * sc->height;
    const bool want_p2p = options().combine == 1;
    bool p2p = want_p2p;
    if (p2p)
        for (int i = 0; i < n; i++) p2p = p2p && sc->rep[i]->dev.peers_ok;
    if (want_p2p && !p2p) return set_err(AKB_CUDA_ERROR, "akb_render: combine=p2p needs peer access between the %d GPUs", n);
    if (!p2p) {
        if (!nccl().ok()) return set_err(AKB_NCCL_ERROR, "akb_render: n_gpus = %d needs NCCL: %s", n, nccl().why.c_str());
        if ((int)g_comms.size() != n) { // one communicator per GPU of the frame, kept until a frame uses another count
            destroy_comms();
            std::vector<int> idx;
            for (int i = 0; i < n; i++) idx.push_back(g_devs[i].index);
            g_comms.assign(n, nullptr);
            const int rc = nccl().CommInitAll(g_comms.data(), n, idx.data());
            if (rc != 0) {
                g_comms.clear();
                return set_err(AKB_NCCL_ERROR, "akb_render: ncclCommInitAll over %d devices: %s", n, nccl().GetErrorString(rc));
            }
        }
    }
    const bool shared_film = p2p && p->split == AKB_SPLIT_TILE;
    for (int i = 0; i < n; i++) {
        CU(cudaSetDevice(sc->rep[i]->dev.index));
        if (i == 0 || !shared_film) {
            akb_status st = ensure_film(sc->rep[i].get(), n_px);
            if (st != AKB_OK) return st;
        }
    }
    std::vector<akb_status> rst(n, AKB_OK);
    std::vector<std::string> rerr(n);
    std::vector<akb_render_stats> rstats(n);
    auto work = [&](int i) {
        Replica *r = sc->rep[i].get();
        NvtxRange range("akb_render: one GPU's share");
        rst[i] = render_on_replica(r, sc->width, sc->height, p, i, n, shared_film ? sc->rep[0]->d_film : r->d_film, shared_film,
                                   r->main_stream, &rstats[i]);
        if (rst[i] != AKB_OK) rerr[i] = g_err;
    };
    {
        std::vector<std::thread> th;
        for (int i = 1; i < n; i++) th.emplace_back(work, i);
        work(0);
        for (auto &t : th) t.join();
    }
    for (int i = 0; i < n; i++)
        if (rst[i] != AKB_OK) return set_err(rst[i], "%s (device %d)", rerr[i].c_str(), sc->rep[i]->dev.index);
    // ---- the one collective of the frame (every share is complete: the threads synchronised their streams)
    const auto t0 = std::chrono::steady_clock::now();
    Replica *r0 = sc->rep[0].get();
    CU(cudaSetDevice(r0->dev.index));
    if (!p2p) {
        NvtxRange range("akb_render: ncclReduce film");
        int rc = nccl().GroupStart();
        for (int i = 0; i < n && rc == 0; i++) {
            Replica *r = sc->rep[i].get();
            rc = nccl().Reduce(r->d_film, r->d_film, n_px * 4, kNcclFloat32, kNcclSum, 0, g_comms[i], r->main_stream);
        }
        const int rc2 = nccl().GroupEnd();
        if (rc == 0) rc = rc2;
        if (rc != 0) return set_err(AKB_NCCL_ERROR, "akb_render: ncclReduce of the film: %s", nccl().GetErrorString(rc));
        CU(cudaSetDevice(r0->dev.index));
        CU(cudaStreamSynchronize(r0->main_stream));
    } else if (!shared_film) {
        NvtxRange range("akb_render: peer-memory film sum");
        PeerFilms pf;
        pf.n = n - 1;
        for (int i = 1; i < n; i++) pf.film[i - 1] = sc->rep[i]->d_film;
        CU(launch_film_sum_peers(r0->d_film, pf, n_px, r0->dev.n_sms, r0->main_stream));
        CU(cudaStreamSynchronize(r0->main_stream));
    }
    *ms_combine = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (stats) {
        memset(stats, 0, sizeof *stats);
        for (int i = 0; i < n; i++) {
            stats->n_samples += rstats[i].n_samples, stats->n_closest_rays += rstats[i].n_closest_rays;
            stats->n_shadow_rays += rstats[i].n_shadow_rays, stats->n_kernel_launches += rstats[i].n_kernel_launches;
            stats->ms_total = std::max(stats->ms_total, rstats[i].ms_total);
        }
        stats->n_kernel_launches += (p2p && !shared_film) ? 1 : 0;
        stats->n_gpus_used = n;
    }
    return AKB_OK;
}

} // namespace

extern "C" {

akb_status akb_render_device(akb_scene *sc, const akb_render_params *p, float *d_out_rgbw, void *cuda_stream,
                             akb_render_stats *stats) {
    akb_status st = check_ready("akb_render_device");
    if (st != AKB_OK) return st;
    if (!sc || !p || !d_out_rgbw) return set_err(AKB_INVALID_ARG, "akb_render_device: null argument");
    int tile = 0;
    st = validate_params("akb_render_device", p, &tile);
    if (st != AKB_OK) return st;
    if (p->n_gpus > 1) return set_err(AKB_INVALID_ARG, "akb_render_device: n_gpus = %d: the multi-GPU frame returns a host film (akb_render)", p->n_gpus);
    NvtxRange range("akb_render_device");
    const int world = p->world_size > 0 ? p->world_size : 1;
    return render_on_replica(sc->rep[0].get(), sc->width, sc->height, p, p->rank, world, (float4 *)d_out_rgbw, false,
                             (cudaStream_t)cuda_stream, stats);
}

akb_status akb_render(akb_scene *sc, const akb_render_params *p, float *out_rgbw, akb_render_stats *stats) {
    akb_status st = check_ready("akb_render");
    if (st != AKB_OK) return st;
    if (!sc || !p || !out_rgbw) return set_err(AKB_INVALID_ARG, "akb_render: null argument");
    int tile = 0;
    st = validate_params("akb_render", p, &tile);
    if (st != AKB_OK) return st;
    NvtxRange range("akb_render");
    const auto t0 = std::chrono::steady_clock::now();
    const size_t n_px = (size_t)sc->width * sc->height;
    const int n = p->n_gpus > 1 ? p->n_gpus : 1;
    Replica *r0 = sc->rep[0].get();
    float ms_combine = 0.f;
    if (n > 1) {
        if (n > (int)sc->rep.size())
            return set_err(AKB_INVALID_ARG, "akb_render: n_gpus = %d but the scene lives on %zu GPU(s) (akb_init(n_devices) before akb_scene_create)", n, sc->rep.size());
        if (p->world_size > 1 || p->rank != 0)
            return set_err(AKB_INVALID_ARG, "akb_render: n_gpus = %d together with rank/world_size = %d/%d (pick one arrangement)", n, p->rank, p->world_size);
        st = render_multi(sc, p, n, stats, &ms_combine);
        if (st != AKB_OK) return st;
    } else {
        CU(cudaSetDevice(r0->dev.index));
        st = ensure_film(r0, n_px);
        if (st != AKB_OK) return st;
        const int world = p->world_size > 0 ? p->world_size : 1;
        st = render_on_replica(r0, sc->width, sc->height, p, p->rank, world, r0->d_film, false, r0->main_stream, stats);
        if (st != AKB_OK) return st;
    }
    CU(cudaSetDevice(r0->dev.index));
    st = download_film(r0, out_rgbw, n_px, r0->main_stream);
    if (st != AKB_OK) return st;
    if (stats) {
        stats->ms_combine = ms_combine;
        stats->ms_wall = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return AKB_OK;
}

akb_status akb_trace_device(akb_scene *sc, const float *d_org, const float *d_dir, size_t n, int any_hit, akb_hit *d_hits,
                            int32_t *d_occ, void *cuda_stream, float *ms_kernel) {
    akb_status st = check_ready("akb_trace_device");
    if (st != AKB_OK) return st;
    if (!sc || (n && (!d_org || !d_dir)) || (n && !any_hit && !d_hits) || (n && any_hit && !d_occ))
        return set_err(AKB_INVALID_ARG, "akb_trace_device: null argument");
    NvtxRange range("akb_trace_device");
    Replica *r = sc->rep[0].get();
    CU(cudaSetDevice(r->dev.index));
    CU(set_debug_bounds(r->S.n_nodes, r->S.n_tris));
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    if (ms_kernel) CU(cudaEventRecord(r->ev0, stream));
    for (size_t off = 0; off < n; off += kMaxBatch) { // the kernel indexes rays with 32 bits
        const size_t m = std::min(kMaxBatch, n - off);
        CU(launch_trace_batch(r->S, (const float4 *)d_org + off, (const float4 *)d_dir + off, m, any_hit,
                              d_hits ? d_hits + off : nullptr, d_occ ? d_occ + off : nullptr, &r->d_ctr->fetch_trace, r->dev.n_sms, stream));
    }
    if (ms_kernel) CU(cudaEventRecord(r->ev1, stream));
    CU(cudaStreamSynchronize(stream));
    if (ms_kernel) CU(cudaEventElapsedTime(ms_kernel, r->ev0, r->ev1));
    return AKB_OK;
}

akb_status akb_trace(akb_scene *sc, const float *org, const float *dir, size_t n, int any_hit, akb_hit *out_hits,
                     int32_t *out_occ) {
    akb_status st = check_ready("akb_trace");
    if (st != AKB_OK) return st;
    if (!sc || (n && (!org || !dir)) || (n && !any_hit && !out_hits) || (n && any_hit && !out_occ))
        return set_err(AKB_INVALID_ARG, "akb_trace: null argument");
    if (n == 0) return AKB_OK;
    NvtxRange range("akb_trace");
    Replica *r = sc->rep[0].get();
    CU(cudaSetDevice(r->dev.index));
    CU(set_debug_bounds(r->S.n_nodes, r->S.n_tris));
    const size_t out_elt = any_hit ? sizeof(int32_t) : sizeof(akb_hit);
    const size_t chunk = std::min<size_t>(n, (size_t)options().trace_chunk);
    cudaError_t e = cudaSuccess;
    if (r->trace_cap < chunk) { // (re)size the staging slots; kept with the scene: pinning memory is slow
        r->free_trace();
        for (Replica::TraceSlot &t : r->trace_slot) {
            if (e == cudaSuccess) e = cudaMallocHost((void **)&t.h_in, chunk * 2 * sizeof(float4));
            if (e == cudaSuccess) e = cudaMallocHost((void **)&t.h_out, chunk * sizeof(akb_hit));
            if (e == cudaSuccess) e = cudaMalloc((void **)&t.d_o, chunk * sizeof(float4));
            if (e == cudaSuccess) e = cudaMalloc((void **)&t.d_d, chunk * sizeof(float4));
            if (e == cudaSuccess) e = cudaMalloc(&t.d_out, chunk * sizeof(akb_hit));
            if (e == cudaSuccess) e = cudaMalloc((void **)&t.cursor, sizeof(uint32_t));
            if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&t.st, cudaStreamNonBlocking);
        }
        if (e == cudaSuccess) r->trace_cap = chunk;
    }
    char *out = any_hit ? (char *)out_occ : (char *)out_hits;
    const size_t n_chunks = (n + chunk - 1) / chunk;
    if (e == cudaSuccess && host_pinned(org) && host_pinned(dir) && host_pinned(out)) {
        // Page-locked caller buffers: every chunk is H2D, kernel, D2H on its slot's stream, straight from / into the
        // caller's memory; the two slots' streams overlap both PCIe directions with the traversal, and stream order is
        // all the synchronisation there is until the end.
        for (size_t k = 0; k < n_chunks && e == cudaSuccess; k++) {
            Replica::TraceSlot &t = r->trace_slot[k & 1];
            const size_t off = k * chunk, m = std::min(chunk, n - off);
            e = cudaMemcpyAsync(t.d_o, org + 4 * off, m * sizeof(float4), cudaMemcpyHostToDevice, t.st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(t.d_d, dir + 4 * off, m * sizeof(float4), cudaMemcpyHostToDevice, t.st);
            if (e == cudaSuccess) e = launch_trace_batch(r->S, t.d_o, t.d_d, m, any_hit, (akb_hit *)t.d_out, (int32_t *)t.d_out, t.cursor, r->dev.n_sms, t.st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(out + off * out_elt, t.d_out, m * out_elt, cudaMemcpyDeviceToHost, t.st);
        }
        for (Replica::TraceSlot &t : r->trace_slot) {
            const cudaError_t rr = cudaStreamSynchronize(t.st);
            if (e == cudaSuccess) e = rr;
        }
        if (e != cudaSuccess) return set_err(AKB_CUDA_ERROR, "akb_trace: %s", cudaGetErrorString(e));
        return AKB_OK;
    }
    auto drain = [&](size_t k) { // chunk k's results: wait for its slot, then pinned -> caller
        Replica::TraceSlot &t = r->trace_slot[k & 1];
        cudaError_t rr = cudaStreamSynchronize(t.st);
        if (rr != cudaSuccess) return rr;
        const size_t off = k * chunk, m = std::min(chunk, n - off);
        host_copy(out + off * out_elt, t.h_out, m * out_elt);
        return cudaSuccess;
    };
    for (size_t k = 0; k < n_chunks && e == cudaSuccess; k++) {
        Replica::TraceSlot &t = r->trace_slot[k & 1];
        if (k >= 2) e = drain(k - 2);
        if (e != cudaSuccess) break;
        const size_t off = k * chunk, m = std::min(chunk, n - off);
        host_copy(t.h_in, org + 4 * off, m * sizeof(float4));
        host_copy(t.h_in + chunk * sizeof(float4), dir + 4 * off, m * sizeof(float4));
        e = cudaMemcpyAsync(t.d_o, t.h_in, m * sizeof(float4), cudaMemcpyHostToDevice, t.st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(t.d_d, t.h_in + chunk * sizeof(float4), m * sizeof(float4), cudaMemcpyHostToDevice, t.st);
        if (e == cudaSuccess) e = launch_trace_batch(r->S, t.d_o, t.d_d, m, any_hit, (akb_hit *)t.d_out, (int32_t *)t.d_out, t.cursor, r->dev.n_sms, t.st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(t.h_out, t.d_out, m * out_elt, cudaMemcpyDeviceToHost, t.st);
    }
    for (size_t k = n_chunks >= 2 ? n_chunks - 2 : 0; k < n_chunks && e == cudaSuccess; k++) e = drain(k);
    if (e != cudaSuccess) {
        cudaStreamSynchronize(r->trace_slot[0].st), cudaStreamSynchronize(r->trace_slot[1].st);
        if (e == cudaErrorMemoryAllocation) r->free_trace();
        return set_err(e == cudaErrorMemoryAllocation ? AKB_OOM : AKB_CUDA_ERROR, "akb_trace: %s", cudaGetErrorString(e));
    }
    return AKB_OK;
}

akb_status akb_generate_rays(akb_scene *sc, const int32_t *xy, const float *u2, size_t n, float *org, float *dir) {
    akb_status st = check_ready("akb_generate_rays");
    if (st != AKB_OK) return st;
    if (!sc || (n && (!xy || !u2 || !org || !dir))) return set_err(AKB_INVALID_ARG, "akb_generate_rays: null argument");
    if (n == 0) return AKB_OK;
    Replica *r = sc->rep[0].get();
    CU(cudaSetDevice(r->dev.index));
    int32_t *d_xy = nullptr;
    float *d_u = nullptr;
    float4 *d_o = nullptr, *d_d = nullptr;
    cudaError_t e = cudaMalloc((void **)&d_xy, n * 8);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_u, n * 8);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_o, n * 16);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_d, n * 16);
    if (e == cudaSuccess) e = cudaMemcpy(d_xy, xy, n * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_u, u2, n * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = launch_generate_rays(r->S, d_xy, d_u, n, d_o, d_d, 0);
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
    CU(cudaSetDevice(g_devs[0].index));
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

// ---------------------------------------------------------------------------------------------------------------------
// akb_film_to_srgb8_noalloc: the byte stage of akb_film_to_srgb8 for callers that may not allocate (CUDA graph capture
// around them, serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference. The device staging of
// akb_film_to_srgb8 (film in, RGB8 out: two cudaMalloc + cudaFree per call) lives in ONE file-scope device array
// instead: [0, 16 * kNoAllocMaxPixels) the float4 film, [16 * kNoAllocMaxPixels, 19 * kNoAllocMaxPixels) the bytes.
// Every GPU that loads this library holds the 38 MiB until the process exits. All calls share the array (it keeps the
// last call's contents; nothing in it has to start at zero: each call overwrites every byte it later reads), so calls
// that overlap on the GPU race.
constexpr size_t kNoAllocMaxPixels = AKB_NOALLOC_MAX_PIXELS;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocMaxPixels * 19];

namespace {
std::mutex g_noalloc_mu;
unsigned char *g_noalloc_ptr[kMaxDevices] = {}; // device address of g_noalloc_scratch per device ordinal

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: resolve it once per device, with
// that device current.
cudaError_t noalloc_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = (unsigned char *)p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}
} // namespace

extern "C" akb_status akb_film_to_srgb8_noalloc(const float *rgbw, int32_t width, int32_t height, uint8_t *out_rgb8) {
    akb_status st = check_ready("akb_film_to_srgb8_noalloc");
    if (st != AKB_OK) return st;
    if (!rgbw || !out_rgb8 || width <= 0 || height <= 0)
        return set_err(AKB_INVALID_ARG, "akb_film_to_srgb8_noalloc: bad argument");
    const size_t n = (size_t)width * (size_t)height;
    if (n > kNoAllocMaxPixels)
        return set_err(AKB_INVALID_ARG,
                       "akb_film_to_srgb8_noalloc: %d x %d = %zu pixels exceed the static scratch of %zu pixels "
                       "(AKB_NOALLOC_MAX_PIXELS); use akb_film_to_srgb8",
                       width, height, n, kNoAllocMaxPixels);
    const int device = g_devs[0].index;
    CU(cudaSetDevice(device));
    unsigned char *scratch = nullptr;
    CU(noalloc_scratch(device, &scratch));
    float4 *d_in = (float4 *)scratch;
    uint8_t *d_out = scratch + kNoAllocMaxPixels * sizeof(float4);
    CU(cudaMemcpy(d_in, rgbw, n * sizeof(float4), cudaMemcpyHostToDevice));
    CU(launch_film_to_srgb8(d_in, n, d_out, 0));
    CU(cudaMemcpy(out_rgb8, d_out, n * 3, cudaMemcpyDeviceToHost));
    return AKB_OK;
}
