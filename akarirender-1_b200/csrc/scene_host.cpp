// scene_host.cpp — see scene_host.h. Host arithmetic here must round exactly like the reference's host build:
// compiled with -ffp-contract=off, fp32 expressions in the reference's order, glibc atanf.
#include "scene_host.h"
#include "options.h"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>

namespace akb {
namespace {

struct m44 {
    float m[4][4];
};
m44 m44_mul(const m44 &a, const m44 &b) { // common/math.h:97-105: m(i,j) = dot(row(i), rhs.col(j))
    m44 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float s = a.m[i][0] * b.m[0][j];
            for (int k = 1; k < 4; k++) s += a.m[i][k] * b.m[k][j];
            r.m[i][j] = s;
        }
    return r;
}
m44 m44_scale(float x, float y, float z) {
    m44 r;
    memset(&r, 0, sizeof r);
    r.m[0][0] = x, r.m[1][1] = y, r.m[2][2] = z, r.m[3][3] = 1;
    return r;
}
m44 m44_translate(float x, float y, float z) {
    m44 r = m44_scale(1, 1, 1);
    r.m[0][3] = x, r.m[1][3] = y, r.m[2][3] = z;
    return r;
}

// PerspectiveCamera ctor + preprocess(), kernel/camera.h:36-59 (note s = atan(fov/2), not tan)
void camera_init(CameraDev *c, const akb_camera *d) {
    memcpy(c->c2w, d->c2w, sizeof(float) * 16);
    c->width = d->width, c->height = d->height;
    m44 m = m44_scale(1, 1, 1);
    m = m44_mul(m44_scale(1.0f / (float)c->width, 1.0f / (float)c->height, 1), m);
    m = m44_mul(m44_scale(2, 2, 1), m);
    m = m44_mul(m44_translate(-1, -1, 0), m);
    m = m44_mul(m44_scale(1, -1, 1), m);
    float s = atanf(d->fov / 2);
    if (c->width > c->height)
        m = m44_mul(m44_scale(s, s * (float)c->height / (float)c->width, 1), m);
    else
        m = m44_mul(m44_scale(s * (float)c->width / (float)c->height, s, 1), m);
    memcpy(c->r2c, m.m, sizeof(float) * 16);
    v3 o = apply_point(c->c2w, mk3(0, 0, 0)); // generate_ray's ray origin, kernel/camera.h:83
    c->ox = o.x, c->oy = o.y, c->oz = o.z;
}

inline v3 ld3(const float *p) { return mk3(p[0], p[1], p[2]); }
inline f4 mkf4(v3 a, float w) { return f4{a.x, a.y, a.z, w}; }
inline float i2f(int32_t i) {
    float f;
    memcpy(&f, &i, 4);
    return f;
}

} // namespace

akb_status flatten_scene(const akb_scene_desc *d, int n_threads, HostScene *hs, std::string *err, BvhBuilder builder) {
    auto fail = [&](const char *msg) {
        if (err) *err = msg;
        return AKB_INVALID_ARG;
    };
    const auto t_f0 = std::chrono::steady_clock::now();
    if (!d) return fail("akb_scene_create: null description");
    if (d->abi_version != AKB_ABI_VERSION) return fail("akb_scene_create: abi_version mismatch");
    if (d->n_meshes && !d->meshes) return fail("akb_scene_create: meshes == NULL");
    if (d->n_materials && !d->materials) return fail("akb_scene_create: materials == NULL");
    if (d->camera.width <= 0 || d->camera.height <= 0) return fail("akb_scene_create: bad camera resolution");
    if (d->camera.lens_radius != 0.0f)
        return fail("akb_scene_create: thin-lens camera not supported (no reference node sets lens_radius)");
    hs->mats.assign(d->materials, d->materials + d->n_materials);
    for (uint32_t i = 0; i < d->n_materials; i++) {
        const akb_material &m = hs->mats[i];
        if (m.type < AKB_MAT_DIFFUSE || m.type > AKB_MAT_DIELECTRIC) return fail("akb_scene_create: bad material type");
        if (m.type == AKB_MAT_MIX && (m.child_a < 0 || m.child_b < 0 || (uint32_t)m.child_a >= d->n_materials ||
                                      (uint32_t)m.child_b >= d->n_materials))
            return fail("akb_scene_create: Mix child out of range");
    }
    if (d->n_textures && !d->textures) return fail("akb_scene_create: textures == NULL");
    for (uint32_t i = 0; i < d->n_materials; i++) {
        const akb_material &m = hs->mats[i];
        for (int32_t t : {m.color_tex, m.roughness_tex, m.fraction_tex})
            if (t < -1 || (t >= 0 && (uint32_t)t >= d->n_textures)) return fail("akb_scene_create: texture index out of range");
    }
    std::vector<float> tex_integral(d->n_textures, 0.f);
    for (uint32_t i = 0; i < d->n_textures; i++) {
        const akb_texture &t = d->textures[i];
        if (t.width <= 0 || t.height <= 0 || !t.rgba) return fail("akb_scene_create: bad texture");
        const size_t n = (size_t)t.width * t.height, first = hs->texels.size();
        if (first + n >= (1ull << 31)) return fail("akb_scene_create: more than 2^31 texels");
        hs->tex_info.insert(hs->tex_info.end(), {(int32_t)first, t.width, t.height, 0});
        hs->texels.resize(first + n);
        memcpy(&hs->texels[first], t.rgba, n * sizeof(f4));
        float I = 0; // ImageTexture::integral, kernel/texture.h:50-56: sequential float sum of luminances / (w*h)
        for (size_t k = 0; k < n; k++)
            I += dot3(mk3(t.rgba[4 * k], t.rgba[4 * k + 1], t.rgba[4 * k + 2]), mk3((float)0.2126, (float)0.7152, (float)0.0722));
        tex_integral[i] = I / (float)(t.width * t.height);
    }
    if (hs->mats.empty()) hs->mats.push_back(akb_material{AKB_MAT_DIFFUSE, {0, 0, 0}, 0, 0, -1, -1, 0, 0, -1, -1, -1});
    hs->geom_offset.resize(d->n_meshes + 1);
    uint64_t nt = 0;
    for (uint32_t g = 0; g < d->n_meshes; g++) {
        const akb_mesh &m = d->meshes[g];
        if (m.n_triangles && (!m.vertices || !m.normals || !m.indices || !m.material_indices))
            return fail("akb_scene_create: mesh with null arrays");
        hs->geom_offset[g] = nt;
        nt += m.n_triangles;
    }
    hs->geom_offset[d->n_meshes] = nt;
    if (nt >= (1ull << 31)) return fail("akb_scene_create: more than 2^31 triangles");
    hs->n_tris = nt;
    hs->shading.resize(6 * nt);
    const bool textured = d->n_textures > 0;
    if (textured) hs->tri_tc.resize(2 * nt);
    FloatVec blo(3 * nt), bhi(3 * nt);
    F4Vec leaf_rec(3 * nt); // flat order first; permuted after the build
    float max_abs = 0;
    struct LightSrc {
        uint64_t tri;
        float uv[3][2];
    };
    std::vector<LightSrc> light_src;
    if (n_threads <= 0) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    // per-triangle records, in chunks of one mesh's triangles over the host threads; emissive triangles are collected
    // per chunk and concatenated in (mesh, prim) order, which is the reference's light order
    struct Chunk {
        uint32_t g;
        uint64_t p0, p1;
        float max_abs = 0;
        const char *error = nullptr;
        std::vector<LightSrc> lights;
    };
    std::vector<Chunk> chunks;
    for (uint32_t g = 0; g < d->n_meshes; g++)
        for (uint64_t p0 = 0; p0 < d->meshes[g].n_triangles; p0 += 65536) {
            Chunk c;
            c.g = g, c.p0 = p0, c.p1 = std::min<uint64_t>(p0 + 65536, d->meshes[g].n_triangles);
            chunks.push_back(std::move(c));
        }
    auto do_chunk = [&](Chunk &ck) {
        auto fail = [&](const char *msg) { ck.error = msg; }; // shadows the outer one: a worker records, the caller reports
        float max_abs = 0;
        std::vector<LightSrc> &light_src = ck.lights;
        const uint32_t g = ck.g;
        const akb_mesh &m = d->meshes[g];
        for (uint64_t p = ck.p0; p < ck.p1; p++) {
            const uint64_t t = hs->geom_offset[g] + p;
            v3 v[3], n[3];
            for (int i = 0; i < 3; i++) { // get_triangle, kernel/instance.h:83-97
                int32_t idx = m.indices[3 * p + i];
                if (idx < 0 || (uint64_t)idx >= m.n_vertices) return fail("akb_scene_create: vertex index out of range");
                v[i] = ld3(m.vertices + 3 * (size_t)idx);
                n[i] = ld3(m.normals + 3 * (3 * p + i));
            }
            int32_t mi = m.material_indices[p];
            int32_t mat = (mi >= 0 && mi < m.n_mesh_materials) ? m.materials[mi] : -1;
            if (mat >= (int32_t)d->n_materials) return fail("akb_scene_create: material index out of range");
            const v3 e1 = sub3(v[1], v[0]), e2 = sub3(v[2], v[0]);
            const v3 ng = normalize3(cross3(e1, e2)); // Triangle::ng(), kernel/shape.h
            f4 *sr = &hs->shading[6 * t];
            sr[0] = mkf4(v[0], i2f(mat)), sr[1] = mkf4(v[1], ng.x), sr[2] = mkf4(v[2], ng.y);
            sr[3] = mkf4(n[0], ng.z), sr[4] = mkf4(n[1], i2f((int32_t)g)), sr[5] = mkf4(n[2], i2f((int32_t)p));
            leaf_rec[3 * t] = mkf4(v[0], i2f((int32_t)t));
            leaf_rec[3 * t + 1] = mkf4(e1, i2f(mat));
            leaf_rec[3 * t + 2] = mkf4(e2, 0.f);
            for (int a = 0; a < 3; a++) {
                float x0 = (&v[0].x)[a], x1 = (&v[1].x)[a], x2 = (&v[2].x)[a];
                float lo = fminf(x0, fminf(x1, x2)), hi = fmaxf(x0, fmaxf(x1, x2));
                blo[3 * t + a] = lo, bhi[3 * t + a] = hi;
                if (isfinite(lo) && isfinite(hi)) max_abs = fmaxf(max_abs, fmaxf(fabsf(lo), fabsf(hi)));
            }
            if (textured) {
                const float *tc = m.texcoords ? m.texcoords + 6 * p : nullptr;
                hs->tri_tc[2 * t] = tc ? f4{tc[0], tc[1], tc[2], tc[3]} : f4{0, 0, 0, 0};
                hs->tri_tc[2 * t + 1] = tc ? f4{tc[4], tc[5], 0, 0} : f4{0, 0, 0, 0};
            }
            if (mat >= 0 && hs->mats[mat].type == AKB_MAT_EMISSIVE) {
                LightSrc ls;
                ls.tri = t;
                for (int i = 0; i < 3; i++) {
                    ls.uv[i][0] = m.texcoords ? m.texcoords[2 * (3 * p + i)] : 0.f;
                    ls.uv[i][1] = m.texcoords ? m.texcoords[2 * (3 * p + i) + 1] : 0.f;
                }
                light_src.push_back(ls);
            }
        }
        ck.max_abs = max_abs;
    };
    {
        std::atomic<size_t> next{0};
        auto work = [&]() {
            for (size_t i; (i = next.fetch_add(1)) < chunks.size();) do_chunk(chunks[i]);
        };
        std::vector<std::thread> th;
        for (int i = 1; i < std::min<int>(n_threads, (int)chunks.size()); i++) th.emplace_back(work);
        work();
        for (auto &t : th) t.join();
    }
    for (Chunk &ck : chunks) {
        if (ck.error) return fail(ck.error);
        max_abs = fmaxf(max_abs, ck.max_abs);
        light_src.insert(light_src.end(), ck.lights.begin(), ck.lights.end());
    }
    // area lights in (mesh, prim) order + power distribution, core/nodes/scene.cpp:58-92
    const size_t nl = light_src.size();
    hs->lights.resize(4 * nl);
    if (textured) hs->light_tc.resize(2 * nl);
    hs->light_func.resize(nl);
    hs->light_cdf.assign(nl + 1, 0.f);
    for (size_t k = 0; k < nl; k++) {
        const f4 *sr = &hs->shading[6 * light_src[k].tri];
        const v3 v0 = mk3(sr[0].x, sr[0].y, sr[0].z), v1 = mk3(sr[1].x, sr[1].y, sr[1].z), v2 = mk3(sr[2].x, sr[2].y, sr[2].z);
        int32_t mat;
        memcpy(&mat, &sr[0].w, 4);
        const float *c = hs->mats[mat].color;
        const v3 color = mk3(c[0], c[1], c[2]);
        // AreaLight::sample's per-light constants (kernel/light.h:58-67), hoisted: ng and area
        const v3 cr = cross3(sub3(v1, v0), sub3(v2, v0));
        const v3 lng = normalize3(cr);
        const float area_half = sqrtf(dot3(cr, cr)) * 0.5f;
        hs->lights[4 * k] = mkf4(v0, area_half);
        hs->lights[4 * k + 1] = mkf4(v1, color.x);
        hs->lights[4 * k + 2] = mkf4(v2, color.y);
        hs->lights[4 * k + 3] = mkf4(lng, color.z);
        // power = |e1 x e2| * texcoord area * integral(texture) (core/nodes/scene.cpp:76-88);
        // ConstantTexture::integral = luminance (kernel/texture.h:36, common/color.h:52-56)
        const int32_t ctex = hs->mats[mat].color_tex;
        const float I = ctex >= 0 ? tex_integral[ctex] : dot3(color, mk3((float)0.2126, (float)0.7152, (float)0.0722));
        if (textured) {
            const float(*uv)[2] = light_src[k].uv;
            hs->light_tc[2 * k] = f4{uv[0][0], uv[0][1], uv[1][0], uv[1][1]};
            hs->light_tc[2 * k + 1] = f4{uv[2][0], uv[2][1], i2f(ctex), 0};
        }
        v3 tc[3];
        for (int j = 0; j < 3; j++) tc[j] = mk3(light_src[k].uv[j][0], light_src[k].uv[j][1], 0.0f);
        const v3 tcc = cross3(sub3(tc[1], tc[0]), sub3(tc[2], tc[0]));
        const double tc_area = (double)sqrtf(dot3(tcc, tcc)) * 0.5;
        const float area = sqrtf(dot3(cr, cr));
        hs->light_func[k] = (float)((double)area * tc_area * (double)I);
    }
    // Distribution1D ctor, common/distribution.h:52-67
    for (size_t i = 0; i < nl; i++) hs->light_cdf[i + 1] = hs->light_cdf[i] + hs->light_func[i] / (float)nl;
    hs->light_func_int = nl ? hs->light_cdf[nl] : 0.f;
    if (hs->light_func_int == 0) {
        for (size_t i = 1; i < nl + 1; i++) hs->light_cdf[i] = (float)i / (float)nl;
    } else {
        for (size_t i = 1; i < nl + 1; i++) hs->light_cdf[i] /= hs->light_func_int;
    }
    camera_init(&hs->cam, &d->camera);

    // conservative box padding: covers the rounding of the quantised slab test for ray origins up to ~12x the
    // scene's coordinate magnitude away (error <= |p - o| * 2^-21.6 per plane)
    const float pad = fmaxf(max_abs, 1e-30f) * 3.814697265625e-06f; // 2^-18
    const auto t_b0 = std::chrono::steady_clock::now();
    if (builder) {
        std::string why;
        if (!builder(blo.data(), bhi.data(), (size_t)nt, pad, n_threads, &hs->bvh, &why)) return fail(("akb_scene_create: " + why).c_str());
    } else {
        build_bvh8(blo.data(), bhi.data(), (size_t)nt, pad, n_threads, &hs->bvh);
    }
    if (hs->bvh.depth > kMaxBvh8Depth) return fail("akb_scene_create: BVH deeper than the traversal stack");
    const auto t_b1 = std::chrono::steady_clock::now();
    hs->tris.resize(3 * nt);
    {
        const uint64_t step = 1 << 16;
        std::atomic<uint64_t> next{0};
        auto work = [&]() {
            for (uint64_t b; (b = next.fetch_add(step)) < nt;)
                for (uint64_t i = b; i < std::min(nt, b + step); i++) {
                    const uint32_t t = hs->bvh.tri_order[i];
                    hs->tris[3 * i] = leaf_rec[3 * (size_t)t];
                    hs->tris[3 * i + 1] = leaf_rec[3 * (size_t)t + 1];
                    hs->tris[3 * i + 2] = leaf_rec[3 * (size_t)t + 2];
                }
        };
        std::vector<std::thread> th;
        for (int i = 1; i < std::min<uint64_t>((uint64_t)n_threads, (nt + step - 1) / step); i++) th.emplace_back(work);
        work();
        for (auto &t : th) t.join();
    }
    if (options().bvh_verbose) {
        const auto t_b2 = std::chrono::steady_clock::now();
        fprintf(stderr, "[akb] flatten: records %.3f s, BVH %.3f s, leaf order %.3f s\n", std::chrono::duration<double>(t_b0 - t_f0).count(),
                std::chrono::duration<double>(t_b1 - t_b0).count(), std::chrono::duration<double>(t_b2 - t_b1).count());
    }
    return AKB_OK;
}

Options &options() {
    static Options o;
    return o;
}

bool set_option(const char *name, const char *value, std::string *err) {
    if (!name || !value) {
        *err = "akb_set_option: null name or value";
        return false;
    }
    Options &o = options();
    const std::string n = name, v = value;
    char *end = nullptr;
    const double d = strtod(value, &end);
    const bool numeric = end != value && *end == 0;
    const long long i = (long long)d;
    auto need = [&](bool ok) {
        if (!ok) *err = "akb_set_option: bad value '" + v + "' for " + n;
        return ok;
    };
#define AKB_INT_OPT(field, lo, hi)                                                                                     \
    if (n == #field) {                                                                                                 \
        if (!need(numeric && i >= (lo) && i <= (hi))) return false;                                                    \
        o.field = (decltype(o.field))i;                                                                                \
        return true;                                                                                                   \
    }
    AKB_INT_OPT(pipelines, 1, 4)
    AKB_INT_OPT(pipeline_interleave, 0, 1)
    AKB_INT_OPT(pipeline_min_slots, 0, 1 << 30)
    AKB_INT_OPT(graphs, 0, 1)
    AKB_INT_OPT(poll_lag, 0, 1)
    AKB_INT_OPT(finish_div, 0, 1 << 30)
    AKB_INT_OPT(finish_max, 1, 1 << 30)
    AKB_INT_OPT(no_park, 0, 1)
    AKB_INT_OPT(shade_partition, 0, 1)
    AKB_INT_OPT(tail_share, 0, 2)
    AKB_INT_OPT(tail_share_min, 0, 1 << 20)
    AKB_INT_OPT(refill_below, 0, 32)
    AKB_INT_OPT(fetch_chunk, 0, 4096)
    AKB_INT_OPT(stream, 0, 1)
    AKB_INT_OPT(stream_flex_mask, 0, 15)
    AKB_INT_OPT(stream_watchdog_ms, 1, 600000)
    AKB_INT_OPT(profile_stages, 0, 1)
    AKB_INT_OPT(timeline, 0, 1)
    AKB_INT_OPT(trace_chunk, 1024, 1ll << 30)
    AKB_INT_OPT(trav_ifif, -1, 1)
    AKB_INT_OPT(bvh_verbose, 0, 1)
    AKB_INT_OPT(bvh_max_leaf, 1, 3)
    AKB_INT_OPT(ploc_radius, 1, 32)
    AKB_INT_OPT(nvtx, 0, 1)
#undef AKB_INT_OPT
    if (n == "check_every") {
        if (!need(numeric && i >= 2 && i <= 64)) return false;
        o.check_every = (int)i & ~1;
        return true;
    }
    if (n == "bvh_trav_cost" || n == "bvh_prim_cost") {
        if (!need(numeric && d > 0)) return false;
        (n == "bvh_trav_cost" ? o.bvh_trav_cost : o.bvh_prim_cost) = (float)d;
        return true;
    }
    if (n == "bvh_build") {
        if (!need(v == "default" || v == "host" || v == "device")) return false;
        o.bvh_build = v == "host" ? AKB_BVH_HOST_SAH : (v == "device" ? AKB_BVH_DEVICE : AKB_BVH_DEFAULT);
        return true;
    }
    if (n == "bvh_device_algo") {
        if (!need(v == "ploc" || v == "lbvh")) return false;
        o.bvh_device_lbvh = v == "lbvh";
        return true;
    }
    if (n == "combine") {
        if (!need(v == "nccl" || v == "p2p")) return false;
        o.combine = v == "p2p";
        return true;
    }
    *err = "akb_set_option: unknown option '" + n + "'";
    return false;
}

SlotMap make_slot_map(int width, int height, int tile, int rank, int world) {
    SlotMap m;
    m.width = width, m.height = height, m.tile = tile;
    m.tiles_x = (width + tile - 1) / tile, m.tiles_y = (height + tile - 1) / tile;
    m.rank = rank, m.world = world, m.groups = 1;
    int nt = m.tiles_x * m.tiles_y;
    m.n_owned_tiles = nt > rank ? (nt - rank + world - 1) / world : 0;
    return m;
}

Shard make_shard(int width, int height, const akb_render_params *p) {
    const int world = p->world_size > 0 ? p->world_size : 1;
    const int tile = p->tile_size > 0 ? p->tile_size : 32;
    Shard s;
    if (p->split == AKB_SPLIT_SAMPLE) {
        s.map = make_slot_map(width, height, tile, 0, 1);
        s.spp = p->spp / world + (p->rank < p->spp % world ? 1 : 0);
        s.seed_offset = (uint64_t)p->rank * (uint64_t)width * (uint64_t)height;
    } else {
        s.map = make_slot_map(width, height, tile, p->rank, world);
        s.spp = p->spp;
        s.seed_offset = 0;
    }
    return s;
}

} // namespace akb
