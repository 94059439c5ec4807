// tests/hostcheck/hostcheck.cpp — TEST INFRASTRUCTURE ONLY. Not part of the product, never loaded by it.
//
// Compiles the product's host-side scene flattening + BVH8 build (csrc/scene_host.cpp, csrc/bvh8.cpp) together
// with the SAME AKB_HD traversal / shading headers the CUDA kernels use (csrc/traverse.h, csrc/shade.h) for the
// CPU, so that the `-m "not gpu"` tests can check the host logic (BVH build, quantisation conservativeness,
// slot map, flattening) and the control flow of the device functions against the oracle without a GPU.
// libakari_b200.so exports none of this: behind the C ABI there is only the CUDA path.
#define AKB_TRAV_STATS 1
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "../../akarirender-1_b200/csrc/scene_host.h"

using namespace akb;

namespace {
struct HC {
    HostScene hs;
    SceneDev dev;
};
template <class F> void parallel_rows(int n, int threads, F f) {
    std::atomic<int> next{0};
    auto work = [&]() {
        for (;;) {
            int i = next.fetch_add(1);
            if (i >= n) break;
            f(i);
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(work);
    work();
    for (auto &t : th) t.join();
}

template <int SAMP> void render_rows(HC *h, const akb_render_params *p, float *out, int threads) {
    const SceneDev &S = h->dev;
    const int W = S.cam.width, H = S.cam.height;
    const Shard shard = make_shard(W, H, p);
    const SlotMap sm = shard.map;
    const int tt = sm.tile * sm.tile;
    parallel_rows(sm.n_owned_tiles, threads, [&](int lt) {
        for (int in = 0; in < tt; in++) {
            int x, y;
            if (!slot_to_pixel(sm, lt * tt + in, x, y)) continue;
            Rng<SAMP> rng;
            rng.set_sample_index((uint64_t)(x + y * W) + shard.seed_offset);
            v3 sum = mk3(0, 0, 0);
            float weight = 0;
            for (int s = 0; s < shard.spp; s++) {
                if (p->integrator == AKB_INTEGRATOR_AO) { // the body of ao_kernel (csrc/kernels.cu)
                    float u1x, u1y, u2x, u2y;
                    rng.next2d(u2x, u2y);
                    rng.next2d(u1x, u1y);
                    v3 o, d;
                    camera_generate_ray(S.cam, u2x, u2y, x, y, o, d);
                    RayHit hit, dummy;
                    float L = 0.0f;
                    if (traverse8<false>(S.nodes, S.tris, o, d, AKB_EPS, INFINITY, hit)) {
                        const RaySeg r = ao_ray<SAMP>(S, rng, hit.u, hit.v, hit.tri, p->ao_occlude);
                        L = traverse8<true>(S.nodes, S.tris, r.o, r.d, r.tmin, r.tmax, dummy) ? 0.0f : 1.0f;
                    }
                    weight += 1.0f;
                    sum = add3(sum, mk3(L, L, L));
                    continue;
                }
                PathState<SAMP> ps;
                ps.L = mk3(0, 0, 0), ps.beta = mk3(1, 1, 1), ps.depth = 0, ps.rng = rng;
                float u1x, u1y, u2x, u2y;
                ps.rng.next2d(u2x, u2y);
                ps.rng.next2d(u1x, u1y);
                RaySeg ray;
                camera_generate_ray(S.cam, u2x, u2y, x, y, ray.o, ray.d);
                ray.tmin = AKB_EPS, ray.tmax = INFINITY;
                for (;;) {
                    RayHit hit;
                    if (!traverse8<false>(S.nodes, S.tris, ray.o, ray.d, ray.tmin, ray.tmax, hit)) break;
                    RaySeg next, sh;
                    v3 sc;
                    int fl = path_vertex<SAMP>(S, ps, ray.d, hit.u, hit.v, hit.tri, p->max_depth, p->rr_start_depth, next, sh, sc);
                    if (fl & VTX_SHADOW) {
                        RayHit dummy;
                        if (!traverse8<true>(S.nodes, S.tris, sh.o, sh.d, sh.tmin, sh.tmax, dummy)) ps.L = add3(ps.L, sc);
                    }
                    if (!(fl & VTX_CONTINUE)) break;
                    ray = next;
                }
                rng = ps.rng;
                weight += 1.0f;
                sum = add3(sum, clamp_sample(ps.L, p->ray_clamp));
            }
            float *o = out + 4 * ((size_t)y * W + x);
            o[0] = sum.x, o[1] = sum.y, o[2] = sum.z, o[3] = weight;
        }
    });
}
} // namespace

extern "C" {
void *hc_scene_create(const akb_scene_desc *d, int threads) {
    HC *h = new HC();
    std::string err;
    if (flatten_scene(d, threads, &h->hs, &err) != AKB_OK) {
        fprintf(stderr, "hostcheck: %s\n", err.c_str());
        delete h;
        return nullptr;
    }
    HostScene &hs = h->hs;
    SceneDev &S = h->dev;
    S.nodes = (const float4 *)hs.bvh.nodes.data();
    S.tris = (const float4 *)hs.tris.data();
    S.shading = (const float4 *)hs.shading.data();
    S.mats = hs.mats.data();
    S.lights = (const float4 *)hs.lights.data();
    S.light_cdf = hs.light_cdf.data(), S.light_func = hs.light_func.data();
    S.light_func_int = hs.light_func_int, S.n_lights = (int)hs.light_func.size(), S.n_mats = (int)hs.mats.size();
    S.n_textures = (int)(hs.tex_info.size() / 4);
    S.texels = (const float4 *)hs.texels.data(), S.tex_info = (const int4 *)hs.tex_info.data();
    S.tri_tc = (const float4 *)hs.tri_tc.data(), S.light_tc = (const float4 *)hs.light_tc.data();
    S.cam = hs.cam;
    return h;
}
void hc_scene_destroy(void *h) { delete (HC *)h; }
uint64_t hc_n_lights(void *h) { return ((HC *)h)->hs.light_func.size(); }
// 0 n_triangles, 1 n_bvh_nodes, 2 bvh_bytes, 3 n_lights, 4 triangle_bytes, 5 bvh depth
uint64_t hc_info(void *hv, int what) {
    HostScene &hs = ((HC *)hv)->hs;
    switch (what) {
    case 0: return hs.n_tris;
    case 1: return hs.bvh.nodes.size();
    case 2: return hs.bvh.nodes.size() * sizeof(Node8);
    case 3: return hs.light_func.size();
    case 4: return hs.tris.size() * 16;
    case 5: return (uint64_t)hs.bvh.depth;
    case 6: return (uint64_t)(hs.bvh.build_seconds * 1e6);
    }
    return 0;
}
int hc_trace(void *hv, const float *org, const float *dir, size_t n, int any_hit, akb_hit *hits, int32_t *occ, int threads) {
    HC *h = (HC *)hv;
    const SceneDev &S = h->dev;
    const int chunk = 4096;
    parallel_rows((int)((n + chunk - 1) / chunk), threads, [&](int c) {
        size_t e = std::min(n, (size_t)(c + 1) * chunk);
        for (size_t i = (size_t)c * chunk; i < e; i++) {
            v3 o = mk3(org[4 * i], org[4 * i + 1], org[4 * i + 2]), d = mk3(dir[4 * i], dir[4 * i + 1], dir[4 * i + 2]);
            RayHit hit;
            if (any_hit) {
                occ[i] = traverse8<true>(S.nodes, S.tris, o, d, org[4 * i + 3], dir[4 * i + 3], hit) ? 1 : 0;
            } else {
                akb_hit r;
                if (traverse8<false>(S.nodes, S.tris, o, d, org[4 * i + 3], dir[4 * i + 3], hit)) {
                    const float4 *sr = S.shading + 6 * (size_t)hit.tri;
                    r.t = hit.t, r.u = hit.u, r.v = hit.v;
                    r.geom_id = (int)akb_f2u(sr[4].w), r.prim_id = (int)akb_f2u(sr[5].w);
                } else {
                    r.t = INFINITY, r.u = r.v = 0, r.prim_id = r.geom_id = -1;
                }
                hits[i] = r;
            }
        }
    });
    return 0;
}
// the same, through the if-if step (trav_step_ii) the persistent warps use on real trees; post_limit < 0 = the value
// api.cu derives (stack size - tree depth - 1); small values force the "finish the group in place" path
int hc_trace_ii(void *hv, const float *org, const float *dir, size_t n, int any_hit, int post_limit, akb_hit *hits, int32_t *occ) {
    HC *h = (HC *)hv;
    const SceneDev &S = h->dev;
    if (post_limit < 0) post_limit = std::max(0, kStackSize - h->hs.bvh.depth - 1);
    for (size_t i = 0; i < n; i++) {
        Trav T;
        uint2 stack[kStackSize];
        trav_init(T, mk3(org[4 * i], org[4 * i + 1], org[4 * i + 2]), mk3(dir[4 * i], dir[4 * i + 1], dir[4 * i + 2]), org[4 * i + 3],
                  dir[4 * i + 3], any_hit != 0);
        while (!trav_step_ii(T, S.nodes, S.tris, stack, true, post_limit)) {
        }
        if (any_hit) {
            occ[i] = T.hit.tri >= 0 ? 1 : 0;
        } else {
            akb_hit r;
            if (T.hit.tri >= 0) {
                const float4 *sr = S.shading + 6 * (size_t)T.hit.tri;
                r.t = T.hit.t, r.u = T.hit.u, r.v = T.hit.v;
                r.geom_id = (int)akb_f2u(sr[4].w), r.prim_id = (int)akb_f2u(sr[5].w);
            } else {
                r.t = INFINITY, r.u = r.v = 0, r.prim_id = r.geom_id = -1;
            }
            hits[i] = r;
        }
    }
    return 0;
}
int hc_generate_rays(void *hv, const int32_t *xy, const float *u2, size_t n, float *org, float *dir) {
    const SceneDev &S = ((HC *)hv)->dev;
    for (size_t i = 0; i < n; i++) {
        v3 o, d;
        camera_generate_ray(S.cam, u2[2 * i], u2[2 * i + 1], xy[2 * i], xy[2 * i + 1], o, d);
        org[4 * i] = o.x, org[4 * i + 1] = o.y, org[4 * i + 2] = o.z, org[4 * i + 3] = AKB_EPS;
        dir[4 * i] = d.x, dir[4 * i + 1] = d.y, dir[4 * i + 2] = d.z, dir[4 * i + 3] = INFINITY;
    }
    return 0;
}
int hc_render(void *hv, const akb_render_params *p, float *out, int threads) {
    if (p->sampler == AKB_SAMPLER_PCG)
        render_rows<AKB_SAMPLER_PCG>((HC *)hv, p, out, threads);
    else
        render_rows<AKB_SAMPLER_LCG>((HC *)hv, p, out, threads);
    return 0;
}
}

// glibc sinf/cosf restatement (csrc/shade.h) against the host libm: returns the number of mismatches over n arguments
extern "C" long hc_trig_mismatches(long n) {
    long bad = 0;
    unsigned s = 1;
    for (long i = 0; i < n; i++) {
        s = s * 1103515245u + 12345u;
        float x = (float)s / 4294967296.0f * 8.0f - 1.0f;
        if (i % 7 == 0) x *= 1e-3f;
        if (i % 11 == 0) x *= 12.f;
        if (sinf(x) != akb::sincosf_glibc(x, 0)) bad++;
        if (cosf(x) != akb::sincosf_glibc(x, 1)) bad++;
    }
    return bad;
}

// traversal statistics over a ray batch: out[0] node visits, out[1] triangle tests, out[2] inner-child hits
extern "C" int hc_trace_stats(void *hv, const float *org, const float *dir, size_t n, int any_hit, double *out) {
    const SceneDev &S = ((HC *)hv)->dev;
    double a = 0, b = 0, c = 0;
    for (size_t i = 0; i < n; i++) {
        Trav T;
        uint2 stack[kStackSize];
        trav_init(T, mk3(org[4 * i], org[4 * i + 1], org[4 * i + 2]), mk3(dir[4 * i], dir[4 * i + 1], dir[4 * i + 2]), org[4 * i + 3],
                  dir[4 * i + 3], any_hit != 0);
        T.n_nodes = T.n_tris = T.n_child_hits = 0;
        while (!trav_step(T, S.nodes, S.tris, stack)) {
        }
        a += T.n_nodes, b += T.n_tris, c += T.n_child_hits;
    }
    out[0] = a / n, out[1] = b / n, out[2] = c / n;
    return 0;
}

// Warp-level cost simulation of traversal scheduling policies (development aid): 32 lanes, rays fetched in order,
// refill when fewer than `refill` lanes are busy. Costs in issued warp-instructions: node block cn, triangle round ct,
// refill cf. policy 0: node step then all its triangles; 1: node block (lanes without pending triangles) + one triangle
// round per iteration; 2: triangle round only when >= th lanes wait (or nobody can do node work); 3: as 2 but lanes
// with other work postpone their triangle group onto the stack instead of waiting.
extern "C" double hc_simulate(void *hv, const float *org, const float *dir, size_t n, int any_hit, int policy, int th,
                              int refill, double cn, double ct, double cf, double *useful_out) {
    const SceneDev &S = ((HC *)hv)->dev;
    struct Lane {
        Trav T;
        uint2 stack[kStackSize];
        bool active = false;
    };
    std::vector<Lane> L(32);
    size_t next = 0;
    double cost = 0, useful = 0;
    for (;;) {
        int n_act = 0;
        for (auto &l : L) n_act += l.active;
        if (n_act < refill && next < n) {
            for (auto &l : L)
                if (!l.active && next < n) {
                    size_t i = next++;
                    trav_init(l.T, mk3(org[4 * i], org[4 * i + 1], org[4 * i + 2]), mk3(dir[4 * i], dir[4 * i + 1], dir[4 * i + 2]),
                              org[4 * i + 3], dir[4 * i + 3], any_hit != 0);
                    l.active = true;
                }
            cost += cf;
            n_act = 0;
            for (auto &l : L) n_act += l.active;
        }
        if (n_act == 0) break;
        // node block
        int n_node = 0;
        for (auto &l : L) {
            if (!l.active) continue;
            if (policy != 0 && l.T.tmask) continue; // waiting for a triangle round
            if (!trav_node(l.T, S.nodes, l.stack)) {
                l.active = false;
                continue;
            }
            n_node++;
        }
        if (n_node) cost += cn, useful += cn * n_node / 32.0;
        // triangle rounds
        auto tri_round = [&](bool only_forced) {
            int k = 0;
            for (auto &l : L) {
                if (!l.active || !l.T.tmask) continue;
                if (only_forced && !l.T.tri_popped && (l.T.ng.y > 0x00FFFFFFu || l.T.sp > 0) && l.T.sp < kStackSize) continue;
                k++;
                if (trav_tri(l.T, S.tris)) l.active = false, l.T.tmask = 0;
            }
            if (k) cost += ct, useful += ct * k / 32.0;
            return k;
        };
        auto n_wait = [&]() {
            int k = 0;
            for (auto &l : L) k += (l.active && l.T.tmask != 0);
            return k;
        };
        if (policy == 0) {
            while (n_wait()) tri_round(false);
        } else if (policy == 1) {
            tri_round(false);
        } else if (policy == 2) {
            if (n_wait() >= th || (n_node == 0 && n_wait() > 0)) tri_round(false);
        } else {
            if (n_wait() >= th) {
                while (n_wait() >= th) tri_round(false);
            }
            // postpone or force
            bool forced = false;
            for (auto &l : L) {
                if (!l.active || !l.T.tmask) continue;
                if (!l.T.tri_popped && (l.T.ng.y > 0x00FFFFFFu || l.T.sp > 0) && l.T.sp < kStackSize) {
                    uint2 e;
                    e.x = l.T.tri_base, e.y = l.T.tmask;
                    // keep node groups above the triangle group: insert below the top? simple: push (processed when popped)
                    l.stack[l.T.sp++] = e;
                    l.T.tmask = 0;
                } else
                    forced = true;
            }
            if (forced) tri_round(true);
        }
        for (auto &l : L)
            if (l.active && trav_idle(l.T)) l.active = false;
    }
    if (useful_out) *useful_out = useful / cost;
    return cost / n;
}
