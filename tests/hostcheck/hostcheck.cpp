// tests/hostcheck/hostcheck.cpp — TEST INFRASTRUCTURE ONLY. Not part of the product, never loaded by it.
//
// Compiles the product's host-side scene flattening + BVH8 build (csrc/scene_host.cpp, csrc/bvh8.cpp) together
// with the SAME AKB_HD traversal / shading headers the CUDA kernels use (csrc/traverse.h, csrc/shade.h) for the
// CPU, so that the `-m "not gpu"` tests can check the host logic (BVH build, quantisation conservativeness,
// slot map, flattening) and the control flow of the device functions against the oracle without a GPU.
// libakari_b200.so exports none of this: behind the C ABI there is only the CUDA path.
#define AKB_TRAV_STATS 1
#include <atomic>
#include <functional>
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
// Debug aid: one pixel of the Path integrator (LCG sampler), every ray also traced by brute force over all
// triangles with the same triangle test and tie rule; prints the rays whose results differ. Returns their number.
int hc_debug_pixel(void *hv, const akb_render_params *p, int x, int y) {
    HC *h = (HC *)hv;
    const SceneDev &S = h->dev;
    const int W = S.cam.width;
    const size_t nt = h->hs.n_tris;
    auto brute = [&](const RaySeg &r, bool any, RayHit &best) {
        best.tri = -1, best.t = r.tmax;
        float tb = r.tmax;
        for (size_t i = 0; i < nt; i++) {
            const float4 *tp = S.tris + 3 * i;
            float t, u, v;
            if (!mt_test(r.o, r.d, r.tmin, any ? r.tmax : INFINITY, tp[0], tp[1], tp[2], t, u, v)) continue;
            if (any) {
                best.tri = (int)akb_f2u(tp[0].w), best.t = t, best.u = u, best.v = v;
                return true;
            }
            const int id = (int)akb_f2u(tp[0].w);
            if (t < tb || (t == tb && best.tri >= 0 && id < best.tri)) tb = t, best.t = t, best.u = u, best.v = v, best.tri = id;
        }
        return best.tri >= 0;
    };
    int bad = 0;
    Rng<AKB_SAMPLER_LCG> rng;
    rng.set_sample_index((uint64_t)(x + y * W));
    for (int s = 0; s < p->spp; s++) {
        PathState<AKB_SAMPLER_LCG> ps;
        ps.L = mk3(0, 0, 0), ps.beta = mk3(1, 1, 1), ps.depth = 0, ps.rng = rng;
        float u1x, u1y, u2x, u2y;
        ps.rng.next2d(u2x, u2y);
        ps.rng.next2d(u1x, u1y);
        RaySeg ray;
        camera_generate_ray(S.cam, u2x, u2y, x, y, ray.o, ray.d);
        ray.tmin = AKB_EPS, ray.tmax = INFINITY;
        for (;;) {
            RayHit hit, bh;
            const bool got = traverse8<false>(S.nodes, S.tris, ray.o, ray.d, ray.tmin, ray.tmax, hit);
            const bool bgot = brute(ray, false, bh);
            if (got != bgot || (got && (hit.tri != bh.tri || hit.t != bh.t))) {
                bad++;
                fprintf(stderr, "sample %d depth %d closest: tree %d t=%.9g (u %.9g v %.9g)  brute %d t=%.9g (u %.9g v %.9g)\n  o=(%.9g %.9g %.9g) d=(%.9g %.9g %.9g) tmin=%.9g\n",
                        s, ps.depth, got ? hit.tri : -1, hit.t, hit.u, hit.v, bgot ? bh.tri : -1, bh.t, bh.u, bh.v, ray.o.x, ray.o.y, ray.o.z,
                        ray.d.x, ray.d.y, ray.d.z, ray.tmin);
            }
            if (!got) break;
            RaySeg next, sh;
            v3 sc;
            int fl = path_vertex<AKB_SAMPLER_LCG>(S, ps, ray.d, hit.u, hit.v, hit.tri, p->max_depth, p->rr_start_depth, next, sh, sc);
            if (fl & VTX_SHADOW) {
                RayHit dummy, bd;
                const bool occ = traverse8<true>(S.nodes, S.tris, sh.o, sh.d, sh.tmin, sh.tmax, dummy);
                const bool bocc = brute(sh, true, bd);
                if (occ != bocc) {
                    bad++;
                    fprintf(stderr, "  brute occluder: flat tri %d t=%.9g u=%.9g v=%.9g\n", bd.tri, bd.t, bd.u, bd.v);
                    fprintf(stderr, "sample %d depth %d shadow: tree %d brute %d\n  o=(%.9g %.9g %.9g) d=(%.9g %.9g %.9g) tmin=%.9g tmax=%.9g\n", s, ps.depth,
                            (int)occ, (int)bocc, sh.o.x, sh.o.y, sh.o.z, sh.d.x, sh.d.y, sh.d.z, sh.tmin, sh.tmax);
                }
                if (!occ) ps.L = add3(ps.L, sc);
            }
            if (!(fl & VTX_CONTINUE)) break;
            ray = next;
        }
        rng = ps.rng;
    }
    return bad;
}
// Debug aid: walk from the root to the leaf slot holding flat triangle `flat_tri` and print, per ancestor, the
// traversal's slab arithmetic for the child on the path next to the exact (double) interval of the dequantised box.
int hc_debug_path(void *hv, const float *o, const float *d, float tmin, float tmax, int flat_tri) {
    HC *h = (HC *)hv;
    const Bvh8 &B = h->hs.bvh;
    // position of the triangle in leaf order
    long pos = -1;
    for (size_t i = 0; i < B.tri_order.size(); i++)
        if ((int)B.tri_order[i] == flat_tri) pos = (long)i;
    if (pos < 0) return -1;
    // find the path by DFS
    std::vector<std::pair<uint32_t, int>> path; // (node, slot)
    std::function<bool(uint32_t)> dfs = [&](uint32_t ni) {
        const Node8 &nd = B.nodes[ni];
        uint32_t ci = nd.child_base;
        for (int s = 0; s < 8; s++) {
            if (nd.meta[s] == 0) continue;
            if (nd.imask & (1u << s)) {
                path.push_back({ni, s});
                if (dfs(ci)) return true;
                path.pop_back();
                ci++;
            } else {
                const int off = nd.meta[s] & 31, cnt = __builtin_popcount(nd.meta[s] >> 5);
                if (pos >= (long)nd.tri_base + off && pos < (long)nd.tri_base + off + cnt) {
                    path.push_back({ni, s});
                    return true;
                }
            }
        }
        return false;
    };
    if (!dfs(0)) return -2;
    Trav T;
    trav_init(T, mk3(o[0], o[1], o[2]), mk3(d[0], d[1], d[2]), tmin, tmax, true);
    const float idv[3] = {T.idx, T.idy, T.idz}, ov[3] = {o[0], o[1], o[2]};
    for (auto &ps : path) {
        const Node8 &nd = B.nodes[ps.first];
        const int s = ps.second;
        const uint8_t *ql[3] = {nd.qlox, nd.qloy, nd.qloz}, *qh[3] = {nd.qhix, nd.qhiy, nd.qhiz};
        float cmin = tmin, cmax = tmax;
        double emin = tmin, emax = tmax;
        fprintf(stderr, "node %u slot %d meta %02x:\n", ps.first, s, nd.meta[s]);
        for (int a = 0; a < 3; a++) {
            const float adj = akb_u2f((uint32_t)nd.e[a] << 23) * idv[a];
            const float org = fmaf(-32768.0f, adj, (nd.p[a] - ov[a]) * idv[a]);
            const uint8_t qn = idv[a] < 0 ? qh[a][s] : ql[a][s], qf = idv[a] < 0 ? ql[a][s] : qh[a][s];
            const float tl = fmaf(32768.0f + (float)qn, adj, org), th = fmaf(32768.0f + (float)qf, adj, org);
            const double sc = ldexp(1.0, (int)nd.e[a] - 127);
            const double blo = (double)nd.p[a] + ql[a][s] * sc, bhi = (double)nd.p[a] + qh[a][s] * sc;
            const double e0 = (blo - (double)ov[a]) / (double)d[a], e1 = (bhi - (double)ov[a]) / (double)d[a];
            fprintf(stderr, "   axis %d: box [%.9g, %.9g] q [%d,%d]  traversal tl %.9g th %.9g   exact %.12g %.12g\n", a, blo, bhi, ql[a][s], qh[a][s], tl, th,
                    std::min(e0, e1), std::max(e0, e1));
            cmin = fmaxf(cmin, tl), cmax = fminf(cmax, th);
            emin = std::max(emin, std::min(e0, e1)), emax = std::min(emax, std::max(e0, e1));
        }
        fprintf(stderr, "   traversal [%.9g, %.9g] %s   exact [%.12g, %.12g] %s\n", cmin, cmax, cmin <= cmax ? "HIT" : "MISS", emin, emax,
                emin <= emax ? "hit" : "miss");
    }
    return (int)path.size();
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

// traversal statistics over a ray batch: out[0] node visits, out[1] triangle tests, out[2] inner-child hits,
// out[3] node visits in which no child was hit
extern "C" int hc_trace_stats(void *hv, const float *org, const float *dir, size_t n, int any_hit, double *out) {
    const SceneDev &S = ((HC *)hv)->dev;
    double a = 0, b = 0, c = 0, e = 0;
    for (size_t i = 0; i < n; i++) {
        Trav T;
        uint2 stack[kStackSize];
        trav_init(T, mk3(org[4 * i], org[4 * i + 1], org[4 * i + 2]), mk3(dir[4 * i], dir[4 * i + 1], dir[4 * i + 2]), org[4 * i + 3],
                  dir[4 * i + 3], any_hit != 0);
        T.n_nodes = T.n_tris = T.n_child_hits = T.n_empty = 0;
        while (!trav_step(T, S.nodes, S.tris, stack)) {
        }
        a += T.n_nodes, b += T.n_tris, c += T.n_child_hits, e += T.n_empty;
    }
    out[0] = a / n, out[1] = b / n, out[2] = c / n, out[3] = e / n;
    return 0;
}

// per-ray work: node visits and triangle tests of every ray of a batch (the tail of this distribution is what a
// traversal launch with few rays per thread waits for)
extern "C" int hc_trace_counts(void *hv, const float *org, const float *dir, size_t n, int any_hit, uint32_t *nodes_out, uint32_t *tris_out) {
    const SceneDev &S = ((HC *)hv)->dev;
    for (size_t i = 0; i < n; i++) {
        Trav T;
        uint2 stack[kStackSize];
        trav_init(T, mk3(org[4 * i], org[4 * i + 1], org[4 * i + 2]), mk3(dir[4 * i], dir[4 * i + 1], dir[4 * i + 2]), org[4 * i + 3],
                  dir[4 * i + 3], any_hit != 0);
        T.n_nodes = T.n_tris = T.n_child_hits = T.n_empty = 0;
        while (!trav_step(T, S.nodes, S.tris, stack)) {
        }
        nodes_out[i] = T.n_nodes, tris_out[i] = T.n_tris;
    }
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
