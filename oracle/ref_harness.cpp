// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// Thin C-ABI driver around the UNMODIFIED reference kernel headers, compiled in place from
// /root/reference/src (see oracle/Makefile). It takes the same flattened scene description the product
// library takes (include/akari_b200.h) and runs the reference's own code on it:
//   * Scene<C> + BVHAccelerator<C> build/intersect/occlude   (kernel/scene.{h,cpp}, kernel/bvh-accelerator.h)
//   * PerspectiveCamera<C>::generate_ray                      (kernel/camera.h:67-86)
//   * GenericPathTracer<C>::run_megakernel                    (kernel/pathtracer.h:136-163)
// The only logic restated here (because it lives in the node layer, which needs the SDL parser and image IO to
// compile) is the 30-line light-list / power-distribution loop of SceneNode::compile (core/nodes/scene.cpp:51-92)
// and the per-pixel sampling loop of cpu::PathTracer::render (kernel/integrators/cpu/integrator.cpp:115-135) —
// the latter because Film keeps its float buffers private and can only write 8-bit PNG (core/film.h:76-113).
// Built into oracle/_ref/libakari_ref.so (git-ignored). Used by tests/ (as the checker), tests/golden/make_golden.py
// and bench.py's `--impl reference` / cpu_baseline legs only.
//
// Compiled a second time with -DAKREF_GPU by nvcc (oracle/Makefile target `ref_gpu`) together with
// the reference's own CUDA wavefront (kernel/integrators/gpu/cuda/integrator.cpp, patched in transit by five sed
// rules): then the scene lives in the reference's managed memory (set_device_gpu(); default_resource() = managed,
// core/device.h:109) and akref_gpu_render() calls gpu::PathTracer<C>::render. Every allocation below therefore goes
// through default_resource(), which is the host heap in the CPU build.
#include <akari/kernel/scene.h>
#include <akari/kernel/bvh-accelerator.h>
#include <akari/kernel/pathtracer.h>
#include <akari/core/device.h>
#include <akari/core/film.h>
#ifdef AKREF_GPU
#include <akari/kernel/integrators/gpu/integrator.h>
#include <chrono>
#endif
#include <akari/core/parallel.h>

#include <atomic>
#include <ctime>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

#include "../include/akari_b200.h"

// Everything below sits inside namespace akari (the extern "C" entry points keep their plain C names): under nvcc the
// reference's int2 / float2 / float3 would otherwise collide with CUDA's built-in vector types.
namespace akari {
using C = Config<float, Color<float, 3>>;

namespace {
    template <class T> using RVec = astd::pmr::vector<T>; // lives in default_resource(): host heap, or managed memory
    template <class T> RVec<T> rvec(size_t n = 0) { return RVec<T>(n, TAllocator<T>(default_resource())); }

    struct RefScene {
        AKR_IMPORT_TYPES()
        std::vector<RVec<float>> P, N, UV;
        std::vector<RVec<int>> I, MI;
        std::vector<RVec<const Material<C> *>> mesh_mats;
        std::vector<Texture<C> *> textures;
        std::vector<Material<C> *> materials; // scene table
        RVec<MeshInstance<C>> instances = rvec<MeshInstance<C>>();
        RVec<AreaLight<C>> lights = rvec<AreaLight<C>>();
        Box<Distribution1D<C>> light_distribution;
        Box<BVHAccelerator<C>> accel;
        Scene<C> scene;
        int width = 0, height = 0;
        ~RefScene() {
            for (auto *m : materials)
                default_resource()->deallocate(m, sizeof(Material<C>), alignof(Material<C>));
            for (auto *t : textures)
                default_resource()->deallocate(t, sizeof(Texture<C>), alignof(Texture<C>));
        }
        std::vector<RVec<RGBA>> images; // texels of the description's image textures, as RGBAImage holds them
        std::vector<akari::int2> image_res;
        template <class... A> Texture<C> *new_tex(A &&...a) {
            auto *t = (Texture<C> *)default_resource()->allocate(sizeof(Texture<C>), alignof(Texture<C>));
            new (t) Texture<C>(std::forward<A>(a)...);
            textures.push_back(t);
            return t;
        }
        Texture<C> *tex(float r, float g, float b) { return new_tex(ConstantTexture<C>(Spectrum(r, g, b))); }
        // constant unless the description names an image texture: then the reference's ImageTexture over those texels
        Texture<C> *tex(int image, float r, float g, float b) {
            if (image < 0)
                return tex(r, g, b);
            RGBAImage::View view{images[image].data(), image_res[image]};
            return new_tex(ImageTexture<C>(view));
        }
    };

    thread_local std::string g_err;

    // core/nodes/scene.cpp:51-92 restated with the reference's own types.
    void build_lights(RefScene &rs) {
        AKR_IMPORT_TYPES()
        auto &scene = rs.scene;
        std::vector<Float> power;
        for (uint32_t mesh_id = 0; mesh_id < scene.meshes.size(); mesh_id++) {
            MeshInstance<C> &mesh = scene.meshes[mesh_id];
            for (uint32_t prim_id = 0; prim_id < mesh.indices.size() / 3; prim_id++) {
                auto triangle = scene.get_triangle(mesh_id, prim_id);
                auto material = triangle.material;
                if (!material)
                    continue;
                if (material->template isa<EmissiveMaterial<C>>())
                    rs.lights.emplace_back(triangle);
            }
        }
        for (size_t i = 0; i < rs.lights.size(); i++) {
            auto &light = rs.lights[i];
            // Texture::integral() (kernel/texture.h:63-65) does not compile under g++ 13 (dependent-base lookup);
            // every texture here is constant, so call ConstantTexture::integral (kernel/texture.h:36) directly.
            Float I = light.color->template isa<ImageTexture<C>>()
                          ? light.color->template get<ImageTexture<C>>()->integral()
                          : light.color->template get<ConstantTexture<C>>()->integral();
            auto &triangle = light.triangle;
            Float3 tc[3];
            for (int j = 0; j < 3; j++)
                tc[j] = Float3(triangle.texcoords[j].x, triangle.texcoords[j].y, 0.0);
            auto tc_area = length(cross(tc[1] - tc[0], tc[2] - tc[0])) * 0.5;
            auto area =
                length(cross(triangle.vertices[1] - triangle.vertices[0], triangle.vertices[2] - triangle.vertices[0]));
            power.emplace_back(area * tc_area * I);
        }
        rs.light_distribution = Box<Distribution1D<C>>::make(default_resource(), power.data(), power.size());
        scene.light_distribution = rs.light_distribution.get();
        scene.area_lights = BufferView<AreaLight<C>>(rs.lights.data(), rs.lights.size());
    }
} // namespace

extern "C" {

const char *akref_last_error(void) { return g_err.c_str(); }

void *akref_scene_create(const akb_scene_desc *d) {
    AKR_IMPORT_TYPES()
    try {
        if (!d || d->abi_version != AKB_ABI_VERSION) {
            g_err = "bad desc";
            return nullptr;
        }
#ifdef AKREF_GPU
        set_device_gpu(); // cmd/akari.cpp:41-76 (`--gpu`) does this before anything allocates (default_resource())
#else
        set_device_cpu();
#endif
        auto rs = std::make_unique<RefScene>();
        for (uint32_t i = 0; i < d->n_textures; i++) {
            const akb_texture &t = d->textures[i];
            RVec<RGBA> px = rvec<RGBA>((size_t)t.width * t.height);
            for (size_t k = 0; k < px.size(); k++)
                px[k] = RGBA(akari::float3(t.rgba[4 * k], t.rgba[4 * k + 1], t.rgba[4 * k + 2]), t.rgba[4 * k + 3]);
            rs->images.push_back(std::move(px));
            rs->image_res.push_back(akari::int2(t.width, t.height));
        }
        // materials: allocate raw storage first so Mix children can point at any entry
        rs->materials.resize(d->n_materials, nullptr);
        for (uint32_t i = 0; i < d->n_materials; i++)
            rs->materials[i] = (Material<C> *)default_resource()->allocate(sizeof(Material<C>), alignof(Material<C>));
        for (uint32_t i = 0; i < d->n_materials; i++) {
            const akb_material &m = d->materials[i];
            switch (m.type) {
            case AKB_MAT_DIFFUSE:
                new (rs->materials[i]) Material<C>(DiffuseMaterial<C>(rs->tex(m.color_tex, m.color[0], m.color[1], m.color[2])));
                break;
            case AKB_MAT_GLOSSY:
                new (rs->materials[i]) Material<C>(GlossyMaterial<C>(rs->tex(m.color_tex, m.color[0], m.color[1], m.color[2]),
                                                                     rs->tex(m.roughness_tex, m.roughness, m.roughness, m.roughness)));
                break;
            case AKB_MAT_EMISSIVE:
                new (rs->materials[i])
                    Material<C>(EmissiveMaterial<C>(rs->tex(m.color_tex, m.color[0], m.color[1], m.color[2]), m.double_sided != 0));
                break;
            case AKB_MAT_MIX:
                if (m.child_a < 0 || m.child_b < 0 || (uint32_t)m.child_a >= d->n_materials ||
                    (uint32_t)m.child_b >= d->n_materials) {
                    g_err = "mix child out of range";
                    return nullptr;
                }
                new (rs->materials[i]) Material<C>(MixMaterial<C>(rs->tex(m.fraction_tex, m.fraction, m.fraction, m.fraction),
                                                                  rs->materials[m.child_a], rs->materials[m.child_b]));
                break;
            default:
                g_err = "material type not representable in the reference (extension)";
                return nullptr;
            }
        }
        size_t nm = d->n_meshes;
        for (size_t k = 0; k < nm; k++) {
            rs->P.push_back(rvec<float>()), rs->N.push_back(rvec<float>()), rs->UV.push_back(rvec<float>());
            rs->I.push_back(rvec<int>()), rs->MI.push_back(rvec<int>());
            rs->mesh_mats.push_back(rvec<const Material<C> *>());
        }
        rs->instances.resize(nm);
        for (size_t k = 0; k < nm; k++) {
            const akb_mesh &m = d->meshes[k];
            auto fill = [](auto &v, const auto *src, size_t n) { // astd::pmr::vector has no range assign
                v.resize(n);
                if (n) std::memcpy(v.data(), src, n * sizeof(*src));
            };
            fill(rs->P[k], m.vertices, 3 * m.n_vertices);
            fill(rs->N[k], m.normals, 9 * m.n_triangles);
            fill(rs->UV[k], m.texcoords, 6 * m.n_triangles);
            fill(rs->I[k], m.indices, 3 * m.n_triangles);
            fill(rs->MI[k], m.material_indices, m.n_triangles);
            for (int j = 0; j < m.n_mesh_materials; j++)
                rs->mesh_mats[k].push_back(rs->materials[m.materials[j]]);
            MeshInstance<C> &inst = rs->instances[k];
            inst.vertices = BufferView<float>(rs->P[k].data(), rs->P[k].size());
            inst.normals = BufferView<float>(rs->N[k].data(), rs->N[k].size());
            inst.texcoords = BufferView<float>(rs->UV[k].data(), rs->UV[k].size());
            inst.indices = BufferView<int>(rs->I[k].data(), rs->I[k].size());
            inst.material_indices = BufferView<int>(rs->MI[k].data(), rs->MI[k].size());
            inst.materials = BufferView<const Material<C> *>(rs->mesh_mats[k].data(), rs->mesh_mats[k].size());
        }
        Scene<C> &scene = rs->scene;
        scene.meshes = BufferView<MeshInstance<C>>(rs->instances.data(), rs->instances.size());
        float cm[16];
        std::memcpy(cm, d->camera.c2w, sizeof(cm));
        Transform3f c2w{Matrix4f(cm)};
        rs->width = d->camera.width, rs->height = d->camera.height;
        scene.camera = PerspectiveCamera<C>(akari::int2(d->camera.width, d->camera.height), c2w, d->camera.fov);
        build_lights(*rs);
        scene.sampler = LCGSampler<C>(); // core/nodes/scene.cpp:126
        rs->accel = Box<BVHAccelerator<C>>::make();
        scene.accel = rs->accel.get(); // core/nodes/scene.cpp:129-130
        scene.commit();                // kernel/scene.cpp:64-81
        return rs.release();
    } catch (std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}

void akref_scene_destroy(void *h) { delete (RefScene *)h; }

uint64_t akref_n_lights(void *h) { return ((RefScene *)h)->lights.size(); }

// Scene::intersect / Scene::occlude over a batch (kernel/scene.h:58-67), n_threads host threads.
int akref_trace(void *h, const float *org_tmin, const float *dir_tmax, size_t n, int any_hit, akb_hit *out_hits,
                int32_t *out_occ, int n_threads) {
    AKR_IMPORT_TYPES()
    auto *rs = (RefScene *)h;
    const Scene<C> &scene = rs->scene;
    if (n_threads <= 0)
        n_threads = (int)std::thread::hardware_concurrency();
    std::atomic<size_t> next{0};
    const size_t chunk = 4096;
    auto work = [&]() {
        for (;;) {
            size_t b = next.fetch_add(chunk);
            if (b >= n)
                break;
            size_t e = std::min(n, b + chunk);
            for (size_t i = b; i < e; i++) {
                Ray3f ray(Float3(org_tmin[4 * i], org_tmin[4 * i + 1], org_tmin[4 * i + 2]),
                          Float3(dir_tmax[4 * i], dir_tmax[4 * i + 1], dir_tmax[4 * i + 2]), org_tmin[4 * i + 3],
                          dir_tmax[4 * i + 3]);
                if (any_hit) {
                    out_occ[i] = scene.occlude(ray) ? 1 : 0;
                } else {
                    Intersection<C> isct;
                    akb_hit o;
                    if (scene.intersect(ray, &isct)) {
                        o.t = isct.t, o.u = isct.uv.x, o.v = isct.uv.y, o.prim_id = isct.prim_id, o.geom_id = isct.geom_id;
                    } else {
                        o.t = Constants<Float>::Inf(), o.u = 0, o.v = 0, o.prim_id = -1, o.geom_id = -1;
                    }
                    out_hits[i] = o;
                }
            }
        }
    };
    std::vector<std::thread> ts;
    for (int t = 1; t < n_threads; t++)
        ts.emplace_back(work);
    work();
    for (auto &t : ts)
        t.join();
    return 0;
}

// Camera::generate_ray(u1,u2,raster) for a batch (kernel/camera.h:67-86). u1 is the lens sample (pinhole: unused).
int akref_generate_rays(void *h, const int32_t *xy, const float *u2, size_t n, float *org_tmin, float *dir_tmax) {
    AKR_IMPORT_TYPES()
    auto *rs = (RefScene *)h;
    for (size_t i = 0; i < n; i++) {
        auto s = rs->scene.camera.generate_ray(akari::float2(0.5f, 0.5f), akari::float2(u2[2 * i], u2[2 * i + 1]),
                                               akari::int2(xy[2 * i], xy[2 * i + 1]));
        org_tmin[4 * i] = s.ray.o.x, org_tmin[4 * i + 1] = s.ray.o.y, org_tmin[4 * i + 2] = s.ray.o.z;
        org_tmin[4 * i + 3] = s.ray.tmin;
        dir_tmax[4 * i] = s.ray.d.x, dir_tmax[4 * i + 1] = s.ray.d.y, dir_tmax[4 * i + 2] = s.ray.d.z;
        dir_tmax[4 * i + 3] = s.ray.tmax;
    }
    return 0;
}

// Sampler known-answer helper: n draws of next1d() after set_sample_index(seed) (kernel/sampler.h:48-66),
// and the (x,y) of the first next2d() on a fresh copy (pins the compiler-defined argument evaluation order).
int akref_sampler_kat(int sampler_kind, uint64_t seed, int n, float *out, float *out_next2d) {
    AKR_IMPORT_TYPES()
    Sampler<C> s = sampler_kind == AKB_SAMPLER_PCG ? Sampler<C>(PCGSampler<C>()) : Sampler<C>(LCGSampler<C>());
    s.set_sample_index(seed);
    Sampler<C> s2 = s;
    for (int i = 0; i < n; i++)
        out[i] = s.next1d();
    auto v = s2.next2d();
    out_next2d[0] = v.x, out_next2d[1] = v.y;
    return 0;
}

// The sampling loop of cpu::PathTracer::render (kernel/integrators/cpu/integrator.cpp:115-135) over a list of pixel
// rectangles [x0,x1) x [y0,y1) (rects = 4 ints each), writing (sum L, sum weight) per pixel as float4 into
// out_rgbw[(y*W+x)*4]. Pixels outside the rectangles are left untouched.
// Scheduling is the reference's: the rectangles are cut into 16x16-pixel tiles (tile_size, cpu/integrator.cpp:92,
// clipped to their rectangle) and ALL tiles of ALL rectangles form one work list that n_threads workers drain through
// one shared counter (core/parallel.cpp:119-129, parallel_for_2d over the tile grid) - so a call with a few thin bands
// still keeps every host thread busy. busy (may be NULL): [0] = CPU seconds summed over the workers
// (CLOCK_THREAD_CPUTIME_ID), [1] = wall seconds of the parallel region; [0]/[1] = threads actually busy.
int akref_render_rects(void *h, const akb_render_params *p, const int32_t *rects, int n_rects, float *out_rgbw,
                       int n_threads, double *busy) {
    AKR_IMPORT_TYPES()
    auto *rs = (RefScene *)h;
    const bool ao = p->integrator == AKB_INTEGRATOR_AO;
    if ((!ao && p->integrator != AKB_INTEGRATOR_PATH) || p->rr_start_depth != 0 || p->ray_clamp != 0.0f) {
        g_err = "reference CPU path has no RR/clamp";
        return 1;
    }
    const Float occlude = p->ao_occlude;
    // The body of the `Li` lambda of cpu::AmbientOcclusion::render (kernel/integrators/cpu/integrator.cpp:42-60), which
    // is local to that function (and that file needs Film/thread-pool code this harness does not link): the same calls
    // on the reference's own Scene, Triangle, Frame3f and sampling<C>, in the same order.
    auto ao_li = [&](Ray3f ray, Sampler<C> &sampler) -> Spectrum {
        Intersection<C> intersection;
        if (rs->scene.intersect(ray, &intersection)) {
            auto trig = rs->scene.get_triangle(intersection.geom_id, intersection.prim_id);
            Frame3f frame(trig.ng());
            auto w = sampling<C>::cosine_hemisphere_sampling(sampler.next2d());
            w = frame.local_to_world(w);
            ray = Ray3f(trig.p(intersection.uv), w);
            intersection = Intersection<C>();
            if (rs->scene.intersect(ray, &intersection) && intersection.t < occlude)
                return Spectrum(0);
            return Spectrum(1);
        }
        return Spectrum(0);
    };
    Scene<C> &scene = rs->scene;
    if (p->sampler == AKB_SAMPLER_PCG)
        scene.sampler = PCGSampler<C>();
    else
        scene.sampler = LCGSampler<C>();
    const int W = rs->width;
    if (n_threads <= 0)
        n_threads = (int)std::thread::hardware_concurrency();
    const int spp = p->spp, max_depth = p->max_depth;
    struct TileJob {
        int x0, y0, x1, y1;
    };
    std::vector<TileJob> tiles;
    const int T = 16;
    for (int r = 0; r < n_rects; r++) {
        const int rx0 = rects[4 * r], ry0 = rects[4 * r + 1], rx1 = rects[4 * r + 2], ry1 = rects[4 * r + 3];
        for (int ty = ry0; ty < ry1; ty += T)
            for (int tx = rx0; tx < rx1; tx += T)
                tiles.push_back(TileJob{tx, ty, std::min(tx + T, rx1), std::min(ty + T, ry1)});
    }
    std::atomic<size_t> next_tile{0};
    std::atomic<long long> cpu_ns{0};
    auto thread_ns = []() {
        timespec ts;
        clock_gettime(CLOCK_THREAD_CPUTIME_ID, &ts);
        return (long long)ts.tv_sec * 1000000000ll + ts.tv_nsec;
    };
    auto work = [&]() {
        const long long t0 = thread_ns();
        auto &camera = scene.camera;
        auto sampler = scene.sampler;
        for (;;) {
            const size_t k = next_tile.fetch_add(1);
            if (k >= tiles.size())
                break;
            const TileJob tj = tiles[k];
            for (int y = tj.y0; y < tj.y1; y++)
                for (int x = tj.x0; x < tj.x1; x++) {
                    sampler.set_sample_index(x + y * W);
                    Spectrum sum(0);
                    Float weight = 0;
                    for (int s = 0; s < spp; s++) {
                        sampler.start_next_sample();
                        if (ao) { // cpu/integrator.cpp:72-80
                            CameraSample<C> sample = camera.generate_ray(sampler.next2d(), sampler.next2d(), akari::int2(x, y));
                            auto L = ao_li(sample.ray, sampler);
                            weight += 1.0f;
                            sum += L;
                            continue;
                        }
                        GenericPathTracer<C> pt;
                        pt.depth = 0;
                        pt.max_depth = max_depth;
                        pt.sampler = sampler;
                        pt.run_megakernel(scene, camera, akari::int2(x, y));
                        sampler = pt.sampler;
                        weight += 1.0f; // Tile::add_sample, core/film.h:66-70
                        sum += pt.L;
                    }
                    float *o = out_rgbw + 4 * ((size_t)y * W + x);
                    o[0] = sum[0], o[1] = sum[1], o[2] = sum[2], o[3] = weight;
                }
        }
        cpu_ns += thread_ns() - t0;
    };
    timespec w0, w1;
    clock_gettime(CLOCK_MONOTONIC, &w0);
    std::vector<std::thread> ts;
    for (int t = 1; t < n_threads; t++)
        ts.emplace_back(work);
    work();
    for (auto &t : ts)
        t.join();
    clock_gettime(CLOCK_MONOTONIC, &w1);
    if (busy) {
        busy[0] = (double)cpu_ns.load() * 1e-9;
        busy[1] = (double)(w1.tv_sec - w0.tv_sec) + 1e-9 * (double)(w1.tv_nsec - w0.tv_nsec);
    }
    return 0;
}

int akref_render(void *h, const akb_render_params *p, int x0, int y0, int x1, int y1, float *out_rgbw,
                 int n_threads) {
    const int32_t rect[4] = {x0, y0, x1, y1};
    return akref_render_rects(h, p, rect, 1, out_rgbw, n_threads, nullptr);
}

// Film::write_image (core/film.h:97-113) -> GammaCorrection::process + DefaultImageWriter::write
// (core/image.cpp:38-71), the reference's own byte stage, run on a film given as (sum R, sum G, sum B, sum weight)
// per pixel: the sums enter a Film<C> through Tile::add_sample + Film::merge_tile (core/film.h:66-95), exactly as a
// render leaves them. `path` ends in .png or .jpg.
int akref_write_image(const float *rgbw, int w, int h, const char *path) {
    AKR_IMPORT_TYPES()
    try {
        set_device_cpu();
        Film<C> film(akari::int2(w, h));
        auto tile = film.tile(Bounds2i{akari::int2(0), akari::int2(w, h)});
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                const float *p = rgbw + 4 * ((size_t)y * w + x);
                tile.add_sample(akari::float2(x, y), Spectrum(p[0], p[1], p[2]), p[3]);
            }
        film.merge_tile(tile);
        film.write_image(fs::path(path));
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

// DefaultImageReader::read (core/image.cpp:78-125): stbi_loadf for .hdr, stbi_load (byte / 255.0f) otherwise.
// Returns RGBA float texels, top row first, as RGBAImage holds them. out may be NULL to query the size only.
int akref_read_image(const char *path, int *w, int *h, float *out_rgba, size_t out_floats) {
    try {
        set_device_cpu();
        auto img = default_image_reader()->read(fs::path(path));
        if (!img)
            return 1;
        *w = img->resolution().x, *h = img->resolution().y;
        if (!out_rgba)
            return 0;
        const size_t n = (size_t)*w * *h;
        if (out_floats < 4 * n)
            return 2;
        for (int y = 0; y < *h; y++)
            for (int x = 0; x < *w; x++) {
                const auto &t = (*img)(x, y);
                float *o = out_rgba + 4 * ((size_t)y * *w + x);
                o[0] = t.rgb[0], o[1] = t.rgb[1], o[2] = t.rgb[2], o[3] = t.alpha;
            }
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

#ifdef AKREF_GPU
// The reference's own CUDA wavefront / megakernel path tracer (gpu::PathTracer<C>::render,
// kernel/integrators/gpu/cuda/integrator.cpp:206-424) on the committed scene; film sums as in akref_render.
// ms_out (may be NULL): wall time of render() + device sync, as the reference's "render done took" measures it.
int akref_gpu_render(void *h, const akb_render_params *p, int tile_size, int wavefront, float *out_rgbw, double *ms_out) {
    AKR_IMPORT_TYPES()
    try {
        auto *rs = (RefScene *)h;
        Scene<C> &scene = rs->scene;
        scene.sampler = LCGSampler<C>();
        Film<C> film(akari::int2(rs->width, rs->height));
        gpu::PathTracer<C> pt(p->spp, p->max_depth, tile_size, p->ray_clamp, wavefront != 0);
        const auto t0 = std::chrono::steady_clock::now();
        pt.render(scene, &film);
        active_device()->sync();
        if (ms_out)
            *ms_out = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        // Film keeps its float buffers private and only writes 8-bit images (core/film.h:72-113): read them through a
        // struct of the same members in the same order.
        struct FilmView {
            TImage<Spectrum> radiance;
            TImage<Float> weight;
            Float splatScale;
        };
        static_assert(sizeof(FilmView) == sizeof(Film<C>), "Film layout changed");
        const FilmView &fv = *reinterpret_cast<const FilmView *>(&film);
        for (int y = 0; y < rs->height; y++)
            for (int x = 0; x < rs->width; x++) {
                float *o = out_rgbw + 4 * ((size_t)y * rs->width + x);
                const auto L = fv.radiance(x, y);
                o[0] = L[0], o[1] = L[1], o[2] = L[2], o[3] = fv.weight(x, y);
            }
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}
#endif

} // extern "C"
} // namespace akari
