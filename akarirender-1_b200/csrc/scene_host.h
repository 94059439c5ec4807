// scene_host.h — host-side flattening of an akb_scene_desc into the arrays the kernels read.
//
// Replaces SceneNode::compile (core/nodes/scene.cpp:43-95: mesh views, AreaLight list, power Distribution1D),
// PerspectiveCamera::preprocess (kernel/camera.h:45-59) and `scene.accel = BVHAccelerator; scene.commit()`
// (core/nodes/scene.cpp:127-135, kernel/scene.cpp:64-81). Pure C++ (no CUDA): api.cu uploads the result to HBM.
#pragma once
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "../../include/akari_b200.h"
#include "bvh8.h"
#include "shade.h"

namespace akb {

struct f4 {
    float x, y, z, w;
};

// std::allocator whose construct() default-initialises, so resize() of a vector that is about to be overwritten
// (hundreds of MB of per-triangle records) does not first write zeros through one thread.
template <class T> struct NoInitAlloc : std::allocator<T> {
    template <class U> struct rebind {
        using other = NoInitAlloc<U>;
    };
    template <class U, class... A> void construct(U *p, A &&...a) {
        if constexpr (sizeof...(A) == 0)
            ::new ((void *)p) U;
        else
            ::new ((void *)p) U(std::forward<A>(a)...);
    }
};
using F4Vec = std::vector<f4, NoInitAlloc<f4>>;
using FloatVec = std::vector<float, NoInitAlloc<float>>;

struct HostScene {
    uint64_t n_tris = 0;
    std::vector<uint64_t> geom_offset; // n_meshes + 1
    Bvh8 bvh;
    F4Vec tris;    // 3 per triangle, leaf order
    F4Vec shading; // 6 per triangle, flat order
    std::vector<akb_material> mats;
    std::vector<f4> lights; // 4 per light
    // image textures (only filled when the description has any): texels of all textures back to back, per texture
    // (first texel, width, height, 0), per triangle (flat order) and per light the three corner texcoords
    std::vector<f4> texels;
    std::vector<int32_t> tex_info; // 4 per texture
    F4Vec tri_tc;        // 2 per triangle: (u0,v0,u1,v1) (u2,v2,-,-)
    std::vector<f4> light_tc;      // 2 per light:    (u0,v0,u1,v1) (u2,v2,texture id bits,-)
    std::vector<float> light_func, light_cdf;
    float light_func_int = 0;
    CameraDev cam;
};

// Returns AKB_OK or AKB_INVALID_ARG (+ message).
// `builder` (optional): builds the BVH from the unpadded triangle boxes instead of build_bvh8 (api.cu passes the
// device builder for AKB_BVH_DEVICE); it returns false (+ message) on failure.
typedef bool (*BvhBuilder)(const float *lo, const float *hi, size_t n, float pad, int n_threads, Bvh8 *out, std::string *err);
akb_status flatten_scene(const akb_scene_desc *desc, int n_threads, HostScene *out, std::string *err, BvhBuilder builder = nullptr);

// Interleaved-tile pixel ownership: slot -> pixel. Slots are grouped per owned tile, Morton order
// inside the tile so that one warp covers an 8x4 pixel block.
struct SlotMap {
    int width, height, tile, tiles_x, tiles_y, rank, world;
    int n_owned_tiles;
    // The wavefront's pipelines (api.cu) are contiguous slot ranges; `groups` > 1 deals the owned tiles out to them in
    // turn (tile i of this rank -> group i % groups), so that every pipeline sees the whole picture - sky and ground
    // alike - and they run out of paths together. Slot order only: which pixel a slot is changes, never a pixel's result.
    int groups;
    AKB_HD int n_slots() const { return n_owned_tiles * tile * tile; }
    AKB_HD int group_tiles(int g) const { return (n_owned_tiles - g + groups - 1) / groups; } // owned tiles i with i % groups == g
};
SlotMap make_slot_map(int width, int height, int tile, int rank, int world);

// This rank's share of a frame (akb_split, include/akari_b200.h): which pixels, how many samples of each, and the
// offset added to the per-pixel sampler seed x + y*W.
struct Shard {
    SlotMap map;
    int spp;
    uint64_t seed_offset;
};
Shard make_shard(int width, int height, const akb_render_params *p);

AKB_HD uint32_t morton_compact(uint32_t v) { // even bits of v -> low half
    v &= 0x55555555u;
    v = (v | (v >> 1)) & 0x33333333u;
    v = (v | (v >> 2)) & 0x0F0F0F0Fu;
    v = (v | (v >> 4)) & 0x00FF00FFu;
    v = (v | (v >> 8)) & 0x0000FFFFu;
    return v;
}
// false when the slot falls outside the image (edge tiles)
AKB_HD bool slot_to_pixel(const SlotMap &m, int slot, int &x, int &y) {
    const int tt = m.tile * m.tile;
    int lt = slot / tt;
    const int in = slot - lt * tt;
    if (m.groups > 1) { // position in slot order -> (group, index in group) -> owned tile
        int g = 0;
        for (int c = m.group_tiles(0); g + 1 < m.groups && lt >= c; c = m.group_tiles(g)) lt -= c, g++;
        lt = lt * m.groups + g;
    }
    const int t = lt * m.world + m.rank;
    const int ty = t / m.tiles_x, tx = t - ty * m.tiles_x;
    x = tx * m.tile + (int)morton_compact((uint32_t)in);
    y = ty * m.tile + (int)morton_compact((uint32_t)in >> 1);
    return x < m.width && y < m.height;
}

} // namespace akb
