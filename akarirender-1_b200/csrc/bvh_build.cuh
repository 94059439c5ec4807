// bvh_build.cuh — device build of the compressed 8-wide BVH.
//
// Replaces, for large scenes, the host build of bvh8.cpp (itself the replacement of the reference's
// TBVHAccelerator constructor, kernel/bvh-accelerator.h:72-87,125-475: 3.4 s per 1 M triangles on 8 cores): Morton
// order + binary topology by locally-ordered clustering (PLOC, Meister & Bittner 2018; or Karras' binary radix tree
// with AKB_BVH_DEVICE_ALGO=lbvh) + bottom-up box fit fused with the optimal 8-wide collapse DP + level-by-level
// emission of the same 80-byte nodes (bvh8.h). The result is a different tree from the host build's, which is
// allowed by the traversal's result contract (traverse.h: nearest t over all triangles, lowest flat id among equal t
// - independent of the tree), so films and hits stay bit-identical up to the ill-conditioned grazing tests described
// in traverse.h (2 of 2 M pixels on C3 at 16 spp); only traversal speed depends on tree quality.
#pragma once
#include <cuda_runtime.h>

#include "bvh8.h"

namespace akb {

struct DeviceBuildTimes {
    float ms_upload = 0, ms_sort = 0, ms_tree = 0, ms_fit = 0, ms_emit = 0, ms_download = 0;
    int ploc_rounds = 0; // 0: radix tree
};

// lo/hi: n x 3 host triangle boxes (unpadded; `pad` is added on the device as build_bvh8 does). Fills out->nodes,
// out->tri_order, out->depth. `*too_deep` is set when the wide tree is deeper than kMaxBvh8Depth (the caller then
// falls back to the host build, whose median fallback bounds the depth).
cudaError_t build_bvh8_device(const float *lo, const float *hi, size_t n, float pad, float node_cost, float prim_cost,
                              int max_leaf, Bvh8 *out, bool *too_deep, DeviceBuildTimes *times);

} // namespace akb
