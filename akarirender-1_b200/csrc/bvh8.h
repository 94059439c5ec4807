// bvh8.h — compressed 8-wide BVH (host build, device traversal in traverse.cuh).
//
// Replaces the reference's two-level binary SBVH (kernel/bvh-accelerator.h:34-57 `BVHNode`, AoS 64 B, pointer
// stack; :125-475 recursiveBuild; :550-683 TLAS over per-mesh BLAS) with ONE flat 8-wide tree over all meshes'
// triangles: 80-byte nodes whose eight child boxes are quantised to 8 bits per plane relative to the node's
// origin and a per-axis power-of-two scale (layout after Ylitie, Karras & Laine, "Efficient Incoherent Ray
// Traversal on GPUs Through Compressed Wide BVHs", HPG 2017). A node is five aligned 16-byte words, so a visit is
// five 128-bit loads; leaves reference 1..3 triangles stored contiguously in leaf order as 3 x float4.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <vector>

namespace akb {

struct Node8 {
    float p[3];         // origin of the quantisation grid = node box minimum
    uint8_t e[3];       // biased (+127) power-of-two exponent of the grid step, per axis
    uint8_t imask;      // bit s set: slot s holds an internal child
    uint32_t child_base; // index of the first internal child (internal children are contiguous, in slot order)
    uint32_t tri_base;   // index (in leaf order) of the first triangle referenced by this node's leaf slots
    uint8_t meta[8];    // per slot: 0 empty; internal: 0b001_11sss (sss = slot); leaf: unary count<<5 | tri offset
    uint8_t qlox[8], qloy[8], qloz[8], qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(Node8) == 80, "Node8 must be five 16-byte words");

struct Bvh8 {
    std::vector<Node8> nodes;
    std::vector<uint32_t> tri_order; // leaf-order position -> flat triangle id
    int depth = 0;                   // wide-tree depth (bounds the traversal stack)
    float pad = 0;                   // conservative padding added to every triangle box
    double sah_cost = 0;
    double build_seconds = 0;
    int built_by = 1; // akb_bvh_build that produced the tree (1 host SAH, 2 device PLOC)
};

constexpr int kMaxBvh8Depth = 30; // traverse.cuh keeps a 32-entry stack

// Cost-model constants of the 8-wide collapse (node visit, primitive test, triangles per leaf slot), shared with the
// device builder (bvh_build.cu).
void bvh8_collapse_params(float *node_cost, float *prim_cost, int *max_leaf);

// lo/hi: n x 3 triangle bounds. Boxes are padded by `pad` on every side before the build.
void build_bvh8(const float *lo, const float *hi, size_t n, float pad, int n_threads, Bvh8 *out);

} // namespace akb
