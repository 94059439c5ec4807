// bvh_build.cu — see bvh_build.cuh. sm_100a only. None of this is on the per-frame path; it is HBM-bound setup
// work: one radix sort of 12 B per triangle, three passes over ~50 B per binary node, one pass per wide-tree level.
//
// Node numbering of the binary tree: internal nodes 0 .. n-2 (radix tree: 0 = root; PLOC: in merge order, the last one
// is the root), leaf k (k-th triangle in Morton order) = node n-1+k. Binary nodes are never stored as such after the build: only their boxes, ranges and the
// collapse DP tables live in HBM until the wide nodes are emitted.
#include "bvh_build.cuh"
#include "options.h"

#include <math.h>
#include <stdint.h>
#include <string.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace akb {
namespace {

constexpr int kThreads = 256;
inline int blocks_for(size_t n) { return (int)((n + kThreads - 1) / kThreads); }

// ---- 1. centroid bounds, Morton keys -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t float_to_ordered(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o);
}
__device__ __forceinline__ float safe_centre(float lo, float hi) {
    const float c = 0.5f * lo + 0.5f * hi;
    return isfinite(c) ? c : 0.0f; // a triangle with a non-finite corner can never be hit; park it anywhere
}

__global__ void centroid_bounds_kernel(const float *__restrict__ lo, const float *__restrict__ hi, uint32_t n, uint32_t *cb /*[6]*/) {
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int a = 0; a < 3; a++) {
            const float c = safe_centre(lo[3 * (size_t)i + a], hi[3 * (size_t)i + a]);
            mn[a] = fminf(mn[a], c), mx[a] = fmaxf(mx[a], c);
        }
    for (int a = 0; a < 3; a++) {
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xFFFFFFFFu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xFFFFFFFFu, mx[a], o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&cb[a], float_to_ordered(mn[a]));
            atomicMax(&cb[3 + a], float_to_ordered(mx[a]));
        }
    }
}

__device__ __forceinline__ uint64_t spread21(uint64_t v) { // bit i of v -> bit 3i
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x001F00000000FFFFull;
    v = (v | (v << 16)) & 0x001F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}

__global__ void morton_kernel(const float *__restrict__ lo, const float *__restrict__ hi, uint32_t n, const uint32_t *__restrict__ cb,
                              uint64_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t key = 0;
    // one cubic grid over the largest extent (not per-axis normalisation): a flat scene then spends no high key bits
    // on its thin axis, which would slice it into overlapping slabs
    float ext = 0.0f;
    for (int a = 0; a < 3; a++) ext = fmaxf(ext, ordered_to_float(cb[3 + a]) - ordered_to_float(cb[a]));
    for (int a = 0; a < 3; a++) {
        const float mn = ordered_to_float(cb[a]);
        const float c = safe_centre(lo[3 * (size_t)i + a], hi[3 * (size_t)i + a]);
        float u = ext > 0.0f ? (c - mn) / ext : 0.0f;
        u = fminf(fmaxf(u, 0.0f), 1.0f);
        const uint32_t q = min((uint32_t)(u * 2097152.0f), 2097151u);
        key |= spread21(q) << (2 - a); // x is the most significant axis of every triplet
    }
    keys[i] = key;
    vals[i] = i;
}

// ---- 2. binary radix tree (Karras 2012, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees")
__device__ __forceinline__ int common_prefix(const uint64_t *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    return a == b ? 64 + __clz(i ^ j) : __clzll((long long)(a ^ b)); // equal keys: the index breaks the tie
}

// left/right: child node ids (see the numbering above); first/count: the node's range in Morton order
__global__ void hierarchy_kernel(const uint64_t *__restrict__ keys, int n, int32_t *__restrict__ left, int32_t *__restrict__ right,
                                 int32_t *__restrict__ parent, int32_t *__restrict__ first, int32_t *__restrict__ count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = common_prefix(keys, n, i, i + 1) > common_prefix(keys, n, i, i - 1) ? 1 : -1;
    const int dmin = common_prefix(keys, n, i, i - d);
    int lmax = 2;
    while (common_prefix(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (common_prefix(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = common_prefix(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (common_prefix(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo_i = min(i, j), hi_i = max(i, j);
    const int32_t L = lo_i == gamma ? (n - 1) + gamma : gamma;
    const int32_t R = hi_i == gamma + 1 ? (n - 1) + gamma + 1 : gamma + 1;
    left[i] = L, right[i] = R;
    parent[L] = i, parent[R] = i;
    first[i] = lo_i, count[i] = hi_i - lo_i + 1;
    if (i == 0) parent[0] = -1;
}

struct TreeArrays {
    int n;
    int root; // binary node the wide root is rooted at: 0 (radix tree), the last merge (PLOC), n-1 when n == 1
    int32_t *left, *right, *parent, *first, *count; // internal nodes (parent: all 2n-1 nodes)
    float4 *nlo, *nhi;                              // all 2n-1 nodes
    float *cost;                                    // 7 per internal node
    uint32_t *dp;                                   // per internal node
    uint32_t *visit;                                // per internal node, zeroed
    const uint32_t *vals;                           // Morton position -> flat triangle id
};

// ---- 2b. PLOC (Meister & Bittner 2018, "Parallel Locally-Ordered Clustering for BVH Construction") -------------------
// Alternative to the radix tree with near-SAH quality: the clusters (initially the triangles in Morton order) each look
// `radius` positions left and right for the neighbour whose union box has the smallest area; mutual nearest neighbours
// merge into a new binary node; the array is compacted and the search repeats until one cluster is left. Ties are
// broken by position, so the globally best pair is always mutual and every round makes progress.
__global__ void leaf_box_kernel(TreeArrays T, const float *__restrict__ lo, const float *__restrict__ hi, float pad, uint32_t *__restrict__ cluster) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T.n) return;
    const size_t t = T.vals[k];
    float4 blo = make_float4(lo[3 * t] - pad, lo[3 * t + 1] - pad, lo[3 * t + 2] - pad, 0.f);
    float4 bhi = make_float4(hi[3 * t] + pad, hi[3 * t + 1] + pad, hi[3 * t + 2] + pad, 0.f);
    if (!(isfinite(blo.x) && isfinite(blo.y) && isfinite(blo.z) && isfinite(bhi.x) && isfinite(bhi.y) && isfinite(bhi.z))) {
        blo = make_float4(INFINITY, INFINITY, INFINITY, 0.f);
        bhi = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
    }
    T.nlo[T.n - 1 + k] = blo, T.nhi[T.n - 1 + k] = bhi;
    cluster[k] = (uint32_t)(T.n - 1 + k);
}

constexpr int kPlocMaxRadius = 32;

__global__ void __launch_bounds__(kThreads) ploc_nearest_kernel(TreeArrays T, const uint32_t *__restrict__ cluster, int c, int radius, int32_t *__restrict__ nn) {
    __shared__ float4 slo[kThreads + 2 * kPlocMaxRadius], shi[kThreads + 2 * kPlocMaxRadius];
    const int b0 = blockIdx.x * kThreads;
    for (int k = threadIdx.x; k < kThreads + 2 * radius; k += kThreads) {
        const int g = b0 - radius + k;
        if (g >= 0 && g < c) {
            const uint32_t id = cluster[g];
            slo[k] = T.nlo[id], shi[k] = T.nhi[id];
        }
    }
    __syncthreads();
    const int i = b0 + threadIdx.x;
    if (i >= c) return;
    const float4 alo = slo[threadIdx.x + radius], ahi = shi[threadIdx.x + radius];
    float best = INFINITY;
    int best_j = -1;
    unsigned best_key = 0;
    for (int dj = -radius; dj <= radius; dj++) {
        const int j = i + dj;
        if (dj == 0 || j < 0 || j >= c) continue;
        const float4 blo = slo[threadIdx.x + radius + dj], bhi = shi[threadIdx.x + radius + dj];
        const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y),
                    dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
        float a = dx * dy + dy * dz + dz * dx;
        if (!(a == a)) a = INFINITY; // two empty boxes
        // ties: a total order on pairs that both ends evaluate alike - nearer positions first, then pairs starting at
        // an even position (so that runs of identical boxes pair up (0,1)(2,3).. and halve every round)
        const unsigned key = ((unsigned)abs(dj) << 26) | ((unsigned)(min(i, j) & 1) << 25);
        if (best_j < 0 || a < best || (a == best && key < best_key)) best = a, best_j = j, best_key = key;
    }
    nn[i] = best_j;
}

// packs (keeps its slot, creates a node) per cluster for one scan: low 32 bits count kept slots, high 32 bits new nodes
__global__ void ploc_flag_kernel(const int32_t *__restrict__ nn, int c, unsigned long long *__restrict__ flags) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == i;
    const unsigned long long keep = (!mutual || i < j) ? 1ull : 0ull, make = (mutual && i < j) ? 1ull : 0ull;
    flags[i] = keep | (make << 32);
}

__global__ void ploc_merge_kernel(TreeArrays T, const uint32_t *__restrict__ cluster, const int32_t *__restrict__ nn, int c,
                                  const unsigned long long *__restrict__ flags, const unsigned long long *__restrict__ offs, int next_node,
                                  uint32_t *__restrict__ cluster_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    const unsigned long long f = flags[i], o = offs[i];
    if (!(f & 1ull)) return; // merged into its partner's slot
    const uint32_t a = cluster[i];
    uint32_t id = a;
    if (f >> 32) {
        const uint32_t b = cluster[nn[i]];
        const int p = next_node + (int)(o >> 32);
        const float4 alo = T.nlo[a], ahi = T.nhi[a], blo = T.nlo[b], bhi = T.nhi[b];
        T.nlo[p] = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
        T.nhi[p] = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
        T.left[p] = (int32_t)a, T.right[p] = (int32_t)b;
        T.parent[a] = p, T.parent[b] = p;
        id = (uint32_t)p;
    }
    cluster_out[(uint32_t)o] = id;
}

// ---- 3. bottom-up: boxes + the optimal 8-wide collapse tables (bvh8.cpp Collapser::solve, same recurrences) ----------
// dp word per internal node: bits 0-2 split8 (k of C_distribute(n,8)), bit 3 kind (1 = wide node, 0 = leaf slot),
// bits 4i..4i+2 for i = 1..6: k of the best distribution for i+1 roots (0 = use one root fewer).
__device__ __forceinline__ float half_area4(float4 lo, float4 hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}


__global__ void fit_kernel(TreeArrays T, const float *__restrict__ lo, const float *__restrict__ hi, float pad, float node_cost,
                           float prim_cost, int max_leaf) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = T.n;
    if (k >= n) return;
    const size_t t = T.vals[k];
    float4 blo = make_float4(lo[3 * t] - pad, lo[3 * t + 1] - pad, lo[3 * t + 2] - pad, 0.f);
    float4 bhi = make_float4(hi[3 * t] + pad, hi[3 * t + 1] + pad, hi[3 * t + 2] + pad, 0.f);
    if (!(isfinite(blo.x) && isfinite(blo.y) && isfinite(blo.z) && isfinite(bhi.x) && isfinite(bhi.y) && isfinite(bhi.z))) {
        blo = make_float4(INFINITY, INFINITY, INFINITY, 0.f); // cannot be hit: an empty box that unions ignore
        bhi = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
    }
    const int leaf = n - 1 + k;
    T.nlo[leaf] = blo, T.nhi[leaf] = bhi;
    int cur = T.parent[leaf];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&T.visit[cur], 1u) == 0u) return; // the sibling subtree is not done: its thread carries on
        __threadfence();
        const int L = T.left[cur], R = T.right[cur];
        const float4 llo = __ldcg(&T.nlo[L]), lhi = __ldcg(&T.nhi[L]), rlo = __ldcg(&T.nlo[R]), rhi = __ldcg(&T.nhi[R]);
        const float4 nlo = make_float4(fminf(llo.x, rlo.x), fminf(llo.y, rlo.y), fminf(llo.z, rlo.z), 0.f);
        const float4 nhi = make_float4(fmaxf(lhi.x, rhi.x), fmaxf(lhi.y, rhi.y), fmaxf(lhi.z, rhi.z), 0.f);
        T.nlo[cur] = nlo, T.nhi[cur] = nhi;
        float cl[7], cr[7];
        if (L >= n - 1) {
            const float c = half_area4(llo, lhi) * prim_cost;
            for (int i = 0; i < 7; i++) cl[i] = c;
        } else {
            for (int i = 0; i < 7; i++) cl[i] = __ldcg(&T.cost[7 * (size_t)L + i]);
        }
        if (R >= n - 1) {
            const float c = half_area4(rlo, rhi) * prim_cost;
            for (int i = 0; i < 7; i++) cr[i] = c;
        } else {
            for (int i = 0; i < 7; i++) cr[i] = __ldcg(&T.cost[7 * (size_t)R + i]);
        }
        const float area = half_area4(nlo, nhi);
        const int np = (L >= n - 1 ? 1 : __ldcg(&T.count[L])) + (R >= n - 1 ? 1 : __ldcg(&T.count[R]));
        T.count[cur] = np;
        const float c_leaf = np <= max_leaf ? area * (float)np * prim_cost : INFINITY;
        float c[7];
        uint32_t word = 0;
        {
            float best = INFINITY;
            uint32_t bk = 1;
            for (int kk = 1; kk < 8; kk++) {
                const float v = cl[kk - 1] + cr[8 - kk - 1];
                if (v < best) best = v, bk = (uint32_t)kk;
            }
            const float c_internal = best + area * node_cost;
            word |= bk;
            if (c_leaf <= c_internal)
                c[0] = c_leaf;
            else
                c[0] = c_internal, word |= 8u;
        }
        for (int i = 2; i <= 7; i++) {
            float best = INFINITY;
            uint32_t bk = 1;
            for (int kk = 1; kk < i; kk++) {
                const float v = cl[kk - 1] + cr[i - kk - 1];
                if (v < best) best = v, bk = (uint32_t)kk;
            }
            if (best < c[i - 2])
                c[i - 1] = best, word |= bk << (4 * (i - 1));
            else
                c[i - 1] = c[i - 2];
        }
        for (int i = 0; i < 7; i++) T.cost[7 * (size_t)cur + i] = c[i];
        T.dp[cur] = word;
        cur = T.parent[cur];
    }
}

// ---- 4. wide-node emission, one launch pair per level ----------------------------------------------------------------
__device__ __forceinline__ bool is_tri_node(const TreeArrays &T, int32_t b) { return b >= T.n - 1; }
__device__ __forceinline__ bool is_leaf_child(const TreeArrays &T, int32_t b) { return is_tri_node(T, b) || !(T.dp[b] & 8u); }
__device__ __forceinline__ int prims_of(const TreeArrays &T, int32_t b) { return is_tri_node(T, b) ? 1 : T.count[b]; }
// flat ids of the (at most 3) triangles below binary node b, left to right
__device__ int collect_tris(const TreeArrays &T, int32_t b, uint32_t *out) {
    int k = 0, sp = 0;
    int32_t st[4];
    st[sp++] = b;
    while (sp > 0 && k < 3) {
        const int32_t x = st[--sp];
        if (is_tri_node(T, x)) {
            out[k++] = T.vals[x - (T.n - 1)];
        } else {
            st[sp++] = T.right[x];
            st[sp++] = T.left[x];
        }
    }
    return k;
}

// the up-to-8 binary nodes that become the children of the wide node rooted at binary node b (Collapser::gather)
__device__ int gather_children(const TreeArrays &T, int32_t b, bool root, int32_t *child) {
    int k = 0;
    if (is_tri_node(T, b) || (root && !(T.dp[b] & 8u))) { // the whole scene fits one leaf slot
        child[k++] = b;
        return k;
    }
    int32_t stack_n[16];
    int stack_i[16], sp = 0;
    const int k8 = (int)(T.dp[b] & 7u);
    stack_n[sp] = T.right[b], stack_i[sp++] = 8 - k8;
    stack_n[sp] = T.left[b], stack_i[sp++] = k8;
    while (sp > 0) {
        const int32_t nd = stack_n[--sp];
        int i = stack_i[sp];
        if (is_tri_node(T, nd)) {
            child[k++] = nd;
            continue;
        }
        const uint32_t w = T.dp[nd];
        while (i > 1 && ((w >> (4 * (i - 1))) & 7u) == 0u) i--;
        if (i == 1) {
            child[k++] = nd;
            continue;
        }
        const int kl = (int)((w >> (4 * (i - 1))) & 7u);
        stack_n[sp] = T.right[nd], stack_i[sp++] = i - kl;
        stack_n[sp] = T.left[nd], stack_i[sp++] = kl;
    }
    return k;
}

// counts of one level: internal children and leaf triangles per item
__global__ void level_count_kernel(TreeArrays T, const int32_t *__restrict__ item_bnode, int begin, int end, uint32_t *__restrict__ n_int,
                                   uint32_t *__restrict__ n_tri) {
    const int w = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= end) return;
    int32_t child[8];
    const int k = gather_children(T, item_bnode[w], w == 0, child);
    uint32_t ni = 0, nt = 0;
    for (int i = 0; i < k; i++) {
        if (k == 1 || is_leaf_child(T, child[i]))
            nt += (uint32_t)prims_of(T, child[i]);
        else
            ni++;
    }
    n_int[w - begin] = ni, n_tri[w - begin] = nt;
}

// 8-bit quantisation of one child box on the node's grid: the same rule as bvh8.cpp Collapser::quantise (margin of
// 2^-8 grid steps beyond the padded child box; see the comment there and traverse.h for why).
__device__ void quantise_axis(double clo, double chi, double origin, double scale, uint8_t &qlo, uint8_t &qhi) {
    const double margin = scale * (1.0 / 256.0);
    double q0 = floor((clo - origin - margin) / scale);
    double q1 = ceil((chi - origin + margin) / scale);
    while (q0 > 0 && origin + q0 * scale > clo - margin) q0 -= 1;
    while (q1 < 255 && origin + q1 * scale < chi + margin) q1 += 1;
    q0 = q0 < 0 ? 0 : (q0 > 255 ? 255 : q0);
    q1 = q1 < 0 ? 0 : (q1 > 255 ? 255 : q1);
    qlo = (uint8_t)(q0 == q0 ? q0 : 255.0); // NaN (never produced by finite boxes) -> empty
    qhi = (uint8_t)(q1 == q1 ? q1 : 0.0);
}

__global__ void level_emit_kernel(TreeArrays T, int32_t *__restrict__ item_bnode, int begin, int end, const uint32_t *__restrict__ off_int,
                                  const uint32_t *__restrict__ off_tri, uint32_t node_base, uint32_t tri_base_level, Node8 *__restrict__ nodes,
                                  uint32_t *__restrict__ tri_order) {
    const int w = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= end) return;
    const int32_t b = item_bnode[w];
    int32_t child[8];
    const int k = gather_children(T, b, w == 0, child);
    const float4 blo = T.nlo[b], bhi = T.nhi[b];
    const float bl[3] = {blo.x, blo.y, blo.z}, bh[3] = {bhi.x, bhi.y, bhi.z};
    float clo[8][3], chi[8][3];
    for (int i = 0; i < k; i++) {
        const float4 l = T.nlo[child[i]], h = T.nhi[child[i]];
        clo[i][0] = l.x, clo[i][1] = l.y, clo[i][2] = l.z, chi[i][0] = h.x, chi[i][1] = h.y, chi[i][2] = h.z;
    }
    // slot s (bit0 = x, bit1 = y, bit2 = z) gets the child lying furthest towards corner s, so that `slot ^ octant`
    // enumerates children front to back (greedy assignment as in bvh8.cpp)
    float cst[8][8];
    for (int i = 0; i < k; i++) {
        float d[3];
        for (int a = 0; a < 3; a++) {
            const float c = 0.5f * (clo[i][a] + chi[i][a]) - 0.5f * (bl[a] + bh[a]);
            d[a] = isfinite(c) ? c : 0.0f;
        }
        for (int s = 0; s < 8; s++) cst[i][s] = ((s & 1) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 4) ? d[2] : -d[2]);
    }
    int slot_of[8], child_in_slot[8];
    uint32_t cdone = 0, sdone = 0;
    for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
    for (int r = 0; r < k; r++) {
        int bi = -1, bs = -1;
        float bc = -INFINITY;
        for (int i = 0; i < k; i++) {
            if (cdone & (1u << i)) continue;
            for (int s = 0; s < 8; s++)
                if (!(sdone & (1u << s)) && (bi < 0 || cst[i][s] > bc)) bc = cst[i][s], bi = i, bs = s;
        }
        cdone |= 1u << bi, sdone |= 1u << bs;
        slot_of[bi] = bs, child_in_slot[bs] = bi;
    }
    Node8 nd;
    memset(&nd, 0, sizeof nd);
    double scale[3], origin[3];
    for (int a = 0; a < 3; a++) {
        double ext = (double)bh[a] - (double)bl[a];
        if (!(ext > 0)) ext = 1e-30;
        float pf = (float)((double)bl[a] - ext * (1.0 / 128.0));
        if ((double)pf > (double)bl[a] - ext * (1.0 / 256.0)) pf = nextafterf(pf, -INFINITY);
        nd.p[a] = pf;
        origin[a] = (double)pf;
        const double span = ((double)bh[a] + ext * (1.0 / 128.0)) - origin[a];
        int ex;
        const double m = frexp(span / 255.0, &ex);
        int e = (m == 0.5) ? ex - 1 : ex;
        while (ldexp(255.0, e) < span) e++;
        e = e < -126 ? -126 : (e > 127 ? 127 : e);
        nd.e[a] = (uint8_t)(e + 127);
        scale[a] = ldexp(1.0, e);
    }
    for (int s = 0; s < 8; s++) nd.qlox[s] = nd.qloy[s] = nd.qloz[s] = 255, nd.qhix[s] = nd.qhiy[s] = nd.qhiz[s] = 0;
    for (int i = 0; i < k; i++) {
        const int s = slot_of[i];
        quantise_axis(clo[i][0], chi[i][0], origin[0], scale[0], nd.qlox[s], nd.qhix[s]);
        quantise_axis(clo[i][1], chi[i][1], origin[1], scale[1], nd.qloy[s], nd.qhiy[s]);
        quantise_axis(clo[i][2], chi[i][2], origin[2], scale[2], nd.qloz[s], nd.qhiz[s]);
    }
    nd.child_base = node_base + off_int[w - begin];
    nd.tri_base = tri_base_level + off_tri[w - begin];
    uint32_t ci = nd.child_base, tri_off = 0;
    for (int s = 0; s < 8; s++) {
        const int i = child_in_slot[s];
        if (i < 0) continue;
        const int32_t c = child[i];
        if (k == 1 || is_leaf_child(T, c)) {
            uint32_t ids[3];
            const int np = collect_tris(T, c, ids);
            nd.meta[s] = (uint8_t)((((1u << np) - 1u) << 5) | tri_off);
            for (int j = 0; j < np; j++) tri_order[nd.tri_base + tri_off + j] = ids[j];
            tri_off += (uint32_t)np;
        } else {
            nd.imask |= (uint8_t)(1u << s);
            nd.meta[s] = (uint8_t)((1u << 5) | (24u + (unsigned)s));
            item_bnode[ci++] = c;
        }
    }
    nodes[w] = nd;
}

#define CB(call)                                                                                                       \
    do {                                                                                                               \
        cudaError_t e_ = (call);                                                                                       \
        if (e_ != cudaSuccess) {                                                                                       \
            bufs.release();                                                                                            \
            return e_;                                                                                                 \
        }                                                                                                              \
    } while (0)

struct Buffers { // device scratch + timing events of one build; release() frees both (also on every error path)
    std::vector<void *> ptrs;
    std::vector<cudaEvent_t> events;
    cudaError_t event(cudaEvent_t *e) {
        cudaError_t r = cudaEventCreate(e);
        if (r == cudaSuccess) events.push_back(*e);
        return r;
    }
    template <class P> cudaError_t get(P **p, size_t count) {
        void *v = nullptr;
        cudaError_t e = cudaMalloc(&v, std::max<size_t>(count, 1) * sizeof(P));
        if (e == cudaSuccess) ptrs.push_back(v);
        *p = (P *)v;
        return e;
    }
    void release() {
        for (void *p : ptrs) cudaFree(p);
        ptrs.clear();
        for (cudaEvent_t e : events) cudaEventDestroy(e);
        events.clear();
    }
};

} // namespace

cudaError_t build_bvh8_device(const float *lo, const float *hi, size_t n_tris, float pad, float node_cost, float prim_cost,
                              int max_leaf, Bvh8 *out, bool *fallback, DeviceBuildTimes *times) {
    *fallback = false;
    if (n_tris == 0 || n_tris > 0x3FFFFFFFull) { // empty scenes and >1 G triangles stay with the host builder
        *fallback = true;
        return cudaSuccess;
    }
    const int n = (int)n_tris;
    Buffers bufs;
    cudaEvent_t ev[7];
    for (auto &e : ev) CB(bufs.event(&e));
    auto done = [&]() { bufs.release(); };
    float *d_lo, *d_hi;
    uint32_t *d_cb, *vals_in, *vals, *visit, *dp, *n_int, *n_tri, *off_int, *off_tri, *tri_order;
    uint64_t *keys_in, *keys;
    TreeArrays T;
    int32_t *item_bnode;
    Node8 *nodes;
    CB(bufs.get(&d_lo, 3 * n_tris));
    CB(bufs.get(&d_hi, 3 * n_tris));
    CB(bufs.get(&d_cb, 6));
    CB(bufs.get(&keys_in, n_tris));
    CB(bufs.get(&keys, n_tris));
    CB(bufs.get(&vals_in, n_tris));
    CB(bufs.get(&vals, n_tris));
    CB(bufs.get(&T.left, n_tris));
    CB(bufs.get(&T.right, n_tris));
    CB(bufs.get(&T.parent, 2 * n_tris));
    CB(bufs.get(&T.first, n_tris));
    CB(bufs.get(&T.count, n_tris));
    CB(bufs.get(&T.nlo, 2 * n_tris));
    CB(bufs.get(&T.nhi, 2 * n_tris));
    CB(bufs.get(&T.cost, 7 * n_tris));
    CB(bufs.get(&dp, n_tris));
    CB(bufs.get(&visit, n_tris));
    CB(bufs.get(&n_int, n_tris + 1));
    CB(bufs.get(&n_tri, n_tris + 1));
    CB(bufs.get(&off_int, n_tris + 1));
    CB(bufs.get(&off_tri, n_tris + 1));
    CB(bufs.get(&item_bnode, n_tris + 1));
    CB(bufs.get(&nodes, n_tris + 1));
    CB(bufs.get(&tri_order, n_tris));
    T.n = n, T.dp = dp, T.visit = visit, T.vals = vals;
    // binary topology: PLOC by default (near-SAH quality), akb_set_option("bvh_device_algo", "lbvh") for the plain radix tree
    const bool use_ploc = !options().bvh_device_lbvh;
    int ploc_radius = 3, ploc_rounds = 0; // measured 1..32 on C3 and a clustered soup: 3-4 is best (profiles/r01c_notes.md)
    ploc_radius = std::min(kPlocMaxRadius, std::max(1, options().ploc_radius));
    uint32_t *cluster_a = nullptr, *cluster_b = nullptr;
    int32_t *nn = nullptr;
    unsigned long long *ploc_offs = nullptr;
    if (use_ploc) {
        CB(bufs.get(&cluster_a, n_tris));
        CB(bufs.get(&cluster_b, n_tris));
        CB(bufs.get(&nn, n_tris));
        CB(bufs.get(&ploc_offs, n_tris));
    }
    size_t tmp_sort = 0, tmp_scan = 0, tmp_scan64 = 0;
    CB(cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, keys_in, keys, vals_in, vals, n, 0, 63));
    CB(cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, n_int, off_int, n));
    CB(cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan64, (unsigned long long *)nullptr, (unsigned long long *)nullptr, n));
    const size_t tmp_bytes = std::max(std::max(tmp_sort, tmp_scan), tmp_scan64);
    uint8_t *tmp;
    CB(bufs.get(&tmp, tmp_bytes));

    CB(cudaEventRecord(ev[0]));
    CB(cudaMemcpyAsync(d_lo, lo, 3 * n_tris * sizeof(float), cudaMemcpyHostToDevice));
    CB(cudaMemcpyAsync(d_hi, hi, 3 * n_tris * sizeof(float), cudaMemcpyHostToDevice));
    CB(cudaEventRecord(ev[1]));
    CB(cudaMemsetAsync(d_cb, 0xFF, 3 * sizeof(uint32_t)));
    CB(cudaMemsetAsync(d_cb + 3, 0x00, 3 * sizeof(uint32_t)));
    centroid_bounds_kernel<<<std::min(blocks_for(n_tris), 148 * 8), kThreads>>>(d_lo, d_hi, (uint32_t)n, d_cb);
    morton_kernel<<<blocks_for(n_tris), kThreads>>>(d_lo, d_hi, (uint32_t)n, d_cb, keys_in, vals_in);
    size_t tb = tmp_bytes;
    CB(cub::DeviceRadixSort::SortPairs(tmp, tb, keys_in, keys, vals_in, vals, n, 0, 63));
    CB(cudaEventRecord(ev[2]));
    CB(cudaMemsetAsync(T.parent, 0xFF, 2 * n_tris * sizeof(int32_t)));
    CB(cudaMemsetAsync(visit, 0, n_tris * sizeof(uint32_t)));
    T.root = 0; // also right for n == 1, where node 0 is the only triangle
    if (n > 1 && !use_ploc) hierarchy_kernel<<<blocks_for(n_tris - 1), kThreads>>>(keys, n, T.left, T.right, T.parent, T.first, T.count);
    if (n > 1 && use_ploc) {
        uint32_t *ca = cluster_a, *cb2 = cluster_b;
        leaf_box_kernel<<<blocks_for(n_tris), kThreads>>>(T, d_lo, d_hi, pad, ca);
        unsigned long long *flags = (unsigned long long *)keys_in; // the unsorted keys are dead by now
        int c = n, next_node = 0;
        while (c > 1) {
            ploc_nearest_kernel<<<blocks_for((size_t)c), kThreads>>>(T, ca, c, ploc_radius, nn);
            ploc_flag_kernel<<<blocks_for((size_t)c), kThreads>>>(nn, c, flags);
            tb = tmp_bytes;
            CB(cub::DeviceScan::ExclusiveSum(tmp, tb, flags, ploc_offs, c));
            unsigned long long last[2];
            CB(cudaMemcpyAsync(&last[0], flags + (c - 1), 8, cudaMemcpyDeviceToHost));
            CB(cudaMemcpyAsync(&last[1], ploc_offs + (c - 1), 8, cudaMemcpyDeviceToHost));
            CB(cudaStreamSynchronize(0));
            const unsigned long long tot = last[0] + last[1];
            const int kept = (int)(tot & 0xFFFFFFFFull), made = (int)(tot >> 32);
            if (made <= 0 || kept != c - made || next_node + made > n - 1) { // cannot happen: the best pair is always mutual
                done();
                return cudaErrorUnknown;
            }
            ploc_merge_kernel<<<blocks_for((size_t)c), kThreads>>>(T, ca, nn, c, flags, ploc_offs, next_node, cb2);
            next_node += made, c = kept;
            std::swap(ca, cb2);
            ploc_rounds++;
        }
        T.root = next_node - 1;
    }
    CB(cudaEventRecord(ev[3]));
    fit_kernel<<<blocks_for(n_tris), kThreads>>>(T, d_lo, d_hi, pad, node_cost, prim_cost, max_leaf);
    CB(cudaEventRecord(ev[4]));
    CB(cudaGetLastError());
    float4 root_box[2];
    CB(cudaMemcpy(&root_box[0], T.nlo + T.root, sizeof(float4), cudaMemcpyDeviceToHost));
    CB(cudaMemcpy(&root_box[1], T.nhi + T.root, sizeof(float4), cudaMemcpyDeviceToHost));
    if (!(isfinite(root_box[0].x) && isfinite(root_box[0].y) && isfinite(root_box[0].z) && isfinite(root_box[1].x) &&
          isfinite(root_box[1].y) && isfinite(root_box[1].z))) {
        *fallback = true;
        done();
        return cudaSuccess;
    }
    // wide nodes, level by level; item w = wide node w, item_bnode[w] = the binary node it is rooted at
    const int32_t root_item = T.root;
    CB(cudaMemcpyAsync(item_bnode, &root_item, sizeof(int32_t), cudaMemcpyHostToDevice));
    uint32_t begin = 0, end = 1, tri_total = 0;
    int depth = 0;
    while (begin < end) {
        if (++depth > kMaxBvh8Depth) {
            *fallback = true;
            done();
            return cudaSuccess;
        }
        const uint32_t m = end - begin;
        level_count_kernel<<<blocks_for(m), kThreads>>>(T, item_bnode, (int)begin, (int)end, n_int, n_tri);
        tb = tmp_bytes;
        CB(cub::DeviceScan::ExclusiveSum(tmp, tb, n_int, off_int, (int)m));
        tb = tmp_bytes;
        CB(cub::DeviceScan::ExclusiveSum(tmp, tb, n_tri, off_tri, (int)m));
        uint32_t last[4];
        CB(cudaMemcpyAsync(&last[0], n_int + (m - 1), 4, cudaMemcpyDeviceToHost));
        CB(cudaMemcpyAsync(&last[1], off_int + (m - 1), 4, cudaMemcpyDeviceToHost));
        CB(cudaMemcpyAsync(&last[2], n_tri + (m - 1), 4, cudaMemcpyDeviceToHost));
        CB(cudaMemcpyAsync(&last[3], off_tri + (m - 1), 4, cudaMemcpyDeviceToHost));
        CB(cudaStreamSynchronize(0));
        const uint32_t tot_int = last[0] + last[1], tot_tri = last[2] + last[3];
        if ((size_t)end + tot_int > n_tris + 1 || (size_t)tri_total + tot_tri > n_tris) { // cannot happen; never overrun
            done();
            return cudaErrorUnknown;
        }
        level_emit_kernel<<<blocks_for(m), kThreads>>>(T, item_bnode, (int)begin, (int)end, off_int, off_tri, end, tri_total, nodes, tri_order);
        begin = end, end += tot_int, tri_total += tot_tri;
    }
    CB(cudaEventRecord(ev[5]));
    CB(cudaGetLastError());
    out->nodes.resize(end);
    out->tri_order.resize(n_tris);
    CB(cudaMemcpy(out->nodes.data(), nodes, (size_t)end * sizeof(Node8), cudaMemcpyDeviceToHost));
    CB(cudaMemcpy(out->tri_order.data(), tri_order, n_tris * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    CB(cudaEventRecord(ev[6]));
    CB(cudaEventSynchronize(ev[6]));
    if (tri_total != (uint32_t)n) {
        done();
        return cudaErrorUnknown;
    }
    out->depth = depth;
    out->pad = pad;
    out->sah_cost = 0;
    if (times) {
        times->ploc_rounds = ploc_rounds;
        cudaEventElapsedTime(&times->ms_upload, ev[0], ev[1]);
        cudaEventElapsedTime(&times->ms_sort, ev[1], ev[2]);
        cudaEventElapsedTime(&times->ms_tree, ev[2], ev[3]);
        cudaEventElapsedTime(&times->ms_fit, ev[3], ev[4]);
        cudaEventElapsedTime(&times->ms_emit, ev[4], ev[5]);
        cudaEventElapsedTime(&times->ms_download, ev[5], ev[6]);
    }
    done();
    return cudaSuccess;
}

} // namespace akb
