// bvh8.cpp — host build of the compressed 8-wide BVH: binned-SAH binary tree (multi-threaded below the top
// levels) -> greedy surface-area collapse to 8 children per node -> octant-ordered slot assignment -> 8-bit
// quantisation. See bvh8.h. Replaces kernel/bvh-accelerator.h:72-87,125-475 (TBVHAccelerator build) of the
// reference; the algorithm is different on purpose (no spatial splits, one level, wide nodes).
#include "bvh8.h"
#include "options.h"

#include <stdlib.h>

#include <math.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>

namespace akb {
namespace {

struct BNode {
    float lo[3], hi[3];
    int32_t left, right; // -1 for leaves
    int32_t first, count; // count > 0 only for binary leaves
    int32_t n_prims;      // triangles below this node: idx[first .. first + n_prims)
};

inline float half_area(const float *lo, const float *hi) {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

constexpr int kBins = 16;
// Cost-model constants: akb_set_option("bvh_trav_cost" / "bvh_prim_cost" / "bvh_max_leaf") for tuning runs; see
// profiles/ for the sweep. A CWBVH leaf slot addresses at most 3 triangles (unary-coded count).
#define kTraversalCost (options().bvh_trav_cost)
#define kMaxLeaf (options().bvh_max_leaf)

struct Builder {
    const float *lo, *hi; // padded triangle boxes
    size_t n;
    std::vector<float> cen;
    std::vector<uint32_t> idx;
    std::vector<BNode> nodes;
    std::atomic<int32_t> n_nodes{0};
    bool median_only = false;

    int32_t alloc() { return n_nodes.fetch_add(1); }

    void bounds(int32_t first, int32_t count, float *blo, float *bhi, float *clo, float *chi) const {
        for (int a = 0; a < 3; a++) blo[a] = clo[a] = INFINITY, bhi[a] = chi[a] = -INFINITY;
        for (int32_t i = first; i < first + count; i++) {
            uint32_t t = idx[i];
            for (int a = 0; a < 3; a++) {
                blo[a] = std::min(blo[a], lo[3 * t + a]);
                bhi[a] = std::max(bhi[a], hi[3 * t + a]);
                clo[a] = std::min(clo[a], cen[3 * t + a]);
                chi[a] = std::max(chi[a], cen[3 * t + a]);
            }
        }
    }

    // returns the split position (first < mid < first+count) or -1 to make a leaf
    int32_t split(int32_t first, int32_t count, const float *blo, const float *bhi, const float *clo,
                  const float *chi) {
        int widest = 0;
        for (int a = 1; a < 3; a++)
            if (chi[a] - clo[a] > chi[widest] - clo[widest]) widest = a;
        if (!(chi[widest] - clo[widest] > 0)) { // all centroids coincide
            return count > kMaxLeaf ? first + count / 2 : -1;
        }
        int best_axis = -1, best_bin = -1;
        float best_cost = INFINITY;
        if (!median_only) {
            for (int a = 0; a < 3; a++) {
                float ext = chi[a] - clo[a];
                if (!(ext > 0)) continue;
                float k = kBins * (1.0f - 1e-6f) / ext;
                int cnt[kBins] = {0};
                float bl[kBins][3], bh[kBins][3];
                for (int b = 0; b < kBins; b++)
                    for (int c = 0; c < 3; c++) bl[b][c] = INFINITY, bh[b][c] = -INFINITY;
                for (int32_t i = first; i < first + count; i++) {
                    uint32_t t = idx[i];
                    int b = (int)(k * (cen[3 * t + a] - clo[a]));
                    b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
                    cnt[b]++;
                    for (int c = 0; c < 3; c++) {
                        bl[b][c] = std::min(bl[b][c], lo[3 * t + c]);
                        bh[b][c] = std::max(bh[b][c], hi[3 * t + c]);
                    }
                }
                float ra[kBins];
                int rc[kBins];
                float l[3] = {INFINITY, INFINITY, INFINITY}, h[3] = {-INFINITY, -INFINITY, -INFINITY};
                int c = 0;
                for (int b = kBins - 1; b > 0; b--) {
                    for (int d = 0; d < 3; d++) l[d] = std::min(l[d], bl[b][d]), h[d] = std::max(h[d], bh[b][d]);
                    c += cnt[b];
                    ra[b] = c ? half_area(l, h) : 0.f;
                    rc[b] = c;
                }
                for (int d = 0; d < 3; d++) l[d] = INFINITY, h[d] = -INFINITY;
                c = 0;
                for (int b = 0; b < kBins - 1; b++) {
                    for (int d = 0; d < 3; d++) l[d] = std::min(l[d], bl[b][d]), h[d] = std::max(h[d], bh[b][d]);
                    c += cnt[b];
                    if (c == 0 || rc[b + 1] == 0) continue;
                    float cost = half_area(l, h) * (float)c + ra[b + 1] * (float)rc[b + 1];
                    if (cost < best_cost) best_cost = cost, best_axis = a, best_bin = b;
                }
            }
        }
        if (best_axis >= 0) {
            float pa = half_area(blo, bhi);
            float split_cost = kTraversalCost * pa + best_cost;
            if (count <= kMaxLeaf && !(split_cost < (float)count * pa)) return -1;
            float ext = chi[best_axis] - clo[best_axis];
            float k = kBins * (1.0f - 1e-6f) / ext;
            float c0 = clo[best_axis];
            int a = best_axis, bb = best_bin;
            auto mid = std::partition(idx.begin() + first, idx.begin() + first + count, [&](uint32_t t) {
                int b = (int)(k * (cen[3 * t + a] - c0));
                b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
                return b <= bb;
            });
            int32_t m = (int32_t)(mid - idx.begin());
            if (m > first && m < first + count) return m;
        }
        if (count <= kMaxLeaf) return -1;
        // fall back: median split on the widest centroid axis
        int32_t m = first + count / 2;
        std::nth_element(idx.begin() + first, idx.begin() + m, idx.begin() + first + count,
                         [&](uint32_t x, uint32_t y) { return cen[3 * x + widest] < cen[3 * y + widest]; });
        return m;
    }

    void init_node(int32_t id, int32_t first, int32_t count, float *clo, float *chi) {
        BNode &nd = nodes[id];
        bounds(first, count, nd.lo, nd.hi, clo, chi);
        nd.left = nd.right = -1;
        nd.first = first, nd.count = count, nd.n_prims = count;
    }

    void build_subtree(int32_t id) { // node `id` initialised as a leaf over [first, first+count)
        std::vector<int32_t> stack{id};
        while (!stack.empty()) {
            int32_t cur = stack.back();
            stack.pop_back();
            int32_t first = nodes[cur].first, count = nodes[cur].count;
            float blo[3], bhi[3], clo[3], chi[3];
            bounds(first, count, blo, bhi, clo, chi);
            int32_t m = split(first, count, blo, bhi, clo, chi);
            if (m < 0) continue;
            int32_t l = alloc(), r = alloc();
            float t0[3], t1[3];
            init_node(l, first, m - first, t0, t1);
            init_node(r, m, first + count - m, t0, t1);
            nodes[cur].left = l, nodes[cur].right = r, nodes[cur].count = 0;
            stack.push_back(r);
            stack.push_back(l);
        }
    }

    void run(int n_threads) {
        nodes.resize(2 * n + 2);
        idx.resize(n);
        for (size_t i = 0; i < n; i++) idx[i] = (uint32_t)i;
        cen.resize(3 * n);
        for (size_t i = 0; i < 3 * n; i++) cen[i] = 0.5f * (lo[i] + hi[i]);
        n_nodes = 0;
        int32_t root = alloc();
        float t0[3], t1[3];
        init_node(root, 0, (int32_t)n, t0, t1);
        if (n_threads <= 1 || n < 65536) {
            build_subtree(root);
            return;
        }
        // phase 1: split the top of the tree sequentially until pieces are small enough to hand out
        size_t piece = std::max<size_t>(4096, n / (size_t)(8 * n_threads));
        std::vector<int32_t> open{root}, tasks;
        while (!open.empty()) {
            int32_t cur = open.back();
            open.pop_back();
            int32_t first = nodes[cur].first, count = nodes[cur].count;
            if ((size_t)count <= piece) {
                tasks.push_back(cur);
                continue;
            }
            float blo[3], bhi[3], clo[3], chi[3];
            bounds(first, count, blo, bhi, clo, chi);
            int32_t m = split(first, count, blo, bhi, clo, chi);
            if (m < 0) continue;
            int32_t l = alloc(), r = alloc();
            init_node(l, first, m - first, t0, t1);
            init_node(r, m, first + count - m, t0, t1);
            nodes[cur].left = l, nodes[cur].right = r, nodes[cur].count = 0;
            open.push_back(r);
            open.push_back(l);
        }
        // phase 2: subtrees in parallel (disjoint index ranges; node slots from the atomic counter)
        std::atomic<size_t> next{0};
        auto work = [&]() {
            for (;;) {
                size_t k = next.fetch_add(1);
                if (k >= tasks.size()) break;
                build_subtree(tasks[k]);
            }
        };
        std::vector<std::thread> th;
        for (int t = 1; t < n_threads; t++) th.emplace_back(work);
        work();
        for (auto &t : th) t.join();
    }
};

inline uint8_t unary(int count) { return (uint8_t)((1u << count) - 1u); }

struct Collapser {
    const Builder &b;
    Bvh8 &out;
    Collapser(const Builder &b_, Bvh8 &o) : b(b_), out(o) {}

    // 8-bit quantisation of the child boxes on a per-node grid. The traversal evaluates plane distances as
    // fma(32768 + q, 2^e/d, (p - o)/d - 32768 * 2^e/d) (traverse.h), whose constant term is rounded with an error of
    // up to 2^-9 grid steps; every quantised plane therefore keeps a margin of 2^-8 steps beyond the (already padded)
    // child box, and the grid itself starts 2^-7 extents below the node box so that the margin never hits the clamp.
    void quantise(Node8 &nd, const BNode &box, const int32_t *child, const int *slot_of, int k) {
        double scale[3], origin[3];
        for (int a = 0; a < 3; a++) {
            double ext = (double)box.hi[a] - (double)box.lo[a];
            if (!(ext > 0)) ext = 1e-30;
            float pf = (float)((double)box.lo[a] - ext * (1.0 / 128.0));
            if ((double)pf > (double)box.lo[a] - ext * (1.0 / 256.0)) pf = nextafterf(pf, -INFINITY);
            nd.p[a] = pf;
            origin[a] = (double)pf;
            double span = ((double)box.hi[a] + ext * (1.0 / 128.0)) - origin[a];
            int ex;
            double m = frexp(span / 255.0, &ex); // span/255 = m * 2^ex, m in [0.5,1)
            int e = (m == 0.5) ? ex - 1 : ex;
            while (ldexp(255.0, e) < span) e++;
            if (e < -126) e = -126;
            if (e > 127) e = 127;
            nd.e[a] = (uint8_t)(e + 127);
            scale[a] = ldexp(1.0, e);
        }
        uint8_t *ql[3] = {nd.qlox, nd.qloy, nd.qloz}, *qh[3] = {nd.qhix, nd.qhiy, nd.qhiz};
        for (int s = 0; s < 8; s++)
            for (int a = 0; a < 3; a++) ql[a][s] = 255, qh[a][s] = 0;
        for (int i = 0; i < k; i++) {
            const BNode &c = b.nodes[child[i]];
            int s = slot_of[i];
            for (int a = 0; a < 3; a++) {
                const double p = origin[a], margin = scale[a] * (1.0 / 256.0);
                double q0 = floor(((double)c.lo[a] - p - margin) / scale[a]);
                double q1 = ceil(((double)c.hi[a] - p + margin) / scale[a]);
                while (q0 > 0 && p + q0 * scale[a] > (double)c.lo[a] - margin) q0 -= 1;
                while (q1 < 255 && p + q1 * scale[a] < (double)c.hi[a] + margin) q1 += 1;
                q0 = q0 < 0 ? 0 : (q0 > 255 ? 255 : q0);
                q1 = q1 < 0 ? 0 : (q1 > 255 ? 255 : q1);
                ql[a][s] = (uint8_t)q0;
                qh[a][s] = (uint8_t)q1;
            }
        }
    }

    // Optimal collapse of the binary tree into 8-wide nodes (dynamic programme of Ylitie et al. 2017, §3.1):
    //   C(n,1) = min(C_leaf(n), C_internal(n));  C(n,i) = min(C_distribute(n,i), C(n,i-1))
    //   C_leaf(n) = A_n * P_n * c_prim (P_n <= 3);  C_internal(n) = C_distribute(n,8) + A_n * c_node
    //   C_distribute(n,j) = min_{0<k<j} C(left,k) + C(right,j-k)
    // cost[n][i-1] = C(n,i); split[n][i-1] = k of the best distribution for i roots (0: use C(n,i-1));
    // kind[n]: the single-root choice (leaf or wide node); split8[n]: k of C_distribute(n,8).
    const float kNodeCost = 1.0f, kPrimCost = options().bvh_prim_cost;
    std::vector<float> cost;
    std::vector<uint8_t> split, split8, kind;

    void solve() {
        const int32_t nn = b.n_nodes.load();
        cost.assign((size_t)nn * 7, INFINITY);
        split.assign((size_t)nn * 7, 0);
        split8.assign(nn, 0);
        kind.assign(nn, 0);
        for (int32_t n = nn - 1; n >= 0; n--) { // children always have larger indices than their parent
            const BNode &nd = b.nodes[n];
            const float area = half_area(nd.lo, nd.hi);
            float *c = &cost[(size_t)n * 7];
            const float c_leaf = nd.n_prims <= kMaxLeaf ? area * (float)nd.n_prims * kPrimCost : INFINITY;
            if (nd.left < 0) {
                for (int i = 0; i < 7; i++) c[i] = c_leaf;
                kind[n] = 0;
                continue;
            }
            const float *cl = &cost[(size_t)nd.left * 7], *cr = &cost[(size_t)nd.right * 7];
            auto distribute = [&](int j, uint8_t &best_k) { // j roots in total, 1..j-1 on the left (capped at 7 each)
                float best = INFINITY;
                for (int k = 1; k < j; k++) {
                    if (k > 7 || j - k > 7) continue;
                    float v = cl[k - 1] + cr[j - k - 1];
                    if (v < best) best = v, best_k = (uint8_t)k;
                }
                return best;
            };
            uint8_t k8 = 1;
            const float c_internal = distribute(8, k8) + area * kNodeCost;
            split8[n] = k8;
            if (c_leaf <= c_internal)
                c[0] = c_leaf, kind[n] = 0;
            else
                c[0] = c_internal, kind[n] = 1;
            for (int i = 2; i <= 7; i++) {
                uint8_t k = 1;
                float d = distribute(i, k);
                if (d < c[i - 2])
                    c[i - 1] = d, split[(size_t)n * 7 + i - 1] = k;
                else
                    c[i - 1] = c[i - 2], split[(size_t)n * 7 + i - 1] = 0;
            }
        }
    }
    void gather(int32_t n, int i, int32_t *child, int &k) const {
        while (i > 1 && split[(size_t)n * 7 + i - 1] == 0) i--;
        if (i == 1) {
            child[k++] = n;
            return;
        }
        const int kl = split[(size_t)n * 7 + i - 1];
        gather(b.nodes[n].left, kl, child, k);
        gather(b.nodes[n].right, i - kl, child, k);
    }
    bool is_leaf_child(int32_t n) const { return kind[n] == 0; }

    void run() {
        solve();
        out.nodes.clear();
        out.tri_order.clear();
        out.tri_order.reserve(b.n);
        out.depth = 0;
        out.sah_cost = 0;
        struct Item {
            int32_t bnode;
            uint32_t wnode;
            int depth;
        };
        std::vector<Item> queue;
        out.nodes.emplace_back();
        queue.push_back({0, 0, 1});
        const float root_area = std::max(half_area(b.nodes[0].lo, b.nodes[0].hi), 1e-30f);
        for (size_t qi = 0; qi < queue.size(); qi++) {
            Item it = queue[qi];
            out.depth = std::max(out.depth, it.depth);
            const BNode &bn = b.nodes[it.bnode];
            int32_t child[8];
            int k = 0;
            if (bn.left < 0 || (it.bnode == 0 && kind[0] == 0)) {
                child[k++] = it.bnode; // the whole scene fits one leaf
            } else {
                const int kl = split8[it.bnode];
                gather(bn.left, kl, child, k);
                gather(bn.right, 8 - kl, child, k);
            }
            // slot assignment: slot s (bit0=x, bit1=y, bit2=z) should hold the child lying furthest towards
            // the (+/-,+/-,+/-) corner s, so that `slot ^ ray_octant` enumerates children front to back.
            float cst[8][8];
            float nc[3];
            for (int a = 0; a < 3; a++) nc[a] = 0.5f * (bn.lo[a] + bn.hi[a]);
            for (int i = 0; i < k; i++) {
                const BNode &c = b.nodes[child[i]];
                float d[3];
                for (int a = 0; a < 3; a++) d[a] = 0.5f * (c.lo[a] + c.hi[a]) - nc[a];
                for (int s = 0; s < 8; s++)
                    cst[i][s] = ((s & 1) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 4) ? d[2] : -d[2]);
            }
            int slot_of[8];
            bool cdone[8] = {false}, sdone[8] = {false};
            for (int r = 0; r < k; r++) {
                int bi = -1, bs = -1;
                float bc = -INFINITY;
                for (int i = 0; i < k; i++) {
                    if (cdone[i]) continue;
                    for (int s = 0; s < 8; s++)
                        if (!sdone[s] && cst[i][s] > bc) bc = cst[i][s], bi = i, bs = s;
                }
                cdone[bi] = sdone[bs] = true;
                slot_of[bi] = bs;
            }
            int child_in_slot[8];
            for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
            for (int i = 0; i < k; i++) child_in_slot[slot_of[i]] = i;

            Node8 nd;
            memset(&nd, 0, sizeof nd);
            quantise(nd, bn, child, slot_of, k);
            nd.child_base = (uint32_t)out.nodes.size();
            nd.tri_base = (uint32_t)out.tri_order.size();
            int n_internal = 0, tri_off = 0;
            for (int s = 0; s < 8; s++) {
                int i = child_in_slot[s];
                if (i < 0) continue;
                const BNode &c = b.nodes[child[i]];
                const bool leaf = is_leaf_child(child[i]) || (it.bnode == child[i]);
                if (!leaf) {
                    nd.imask |= (uint8_t)(1u << s);
                    nd.meta[s] = (uint8_t)((1u << 5) | (24u + (unsigned)s));
                    n_internal++;
                } else {
                    nd.meta[s] = (uint8_t)((unary(c.n_prims) << 5) | (unsigned)tri_off);
                    for (int32_t t = c.first; t < c.first + c.n_prims; t++) out.tri_order.push_back(b.idx[t]);
                    tri_off += c.n_prims;
                    out.sah_cost += (double)c.n_prims * kPrimCost * half_area(c.lo, c.hi) / root_area;
                }
            }
            out.sah_cost += (double)kNodeCost * half_area(bn.lo, bn.hi) / root_area;
            out.nodes.resize(out.nodes.size() + (size_t)n_internal);
            uint32_t ci = nd.child_base;
            for (int s = 0; s < 8; s++) {
                int i = child_in_slot[s];
                if (i < 0) continue;
                if (is_leaf_child(child[i]) || it.bnode == child[i]) continue;
                queue.push_back({child[i], ci++, it.depth + 1});
            }
            out.nodes[it.wnode] = nd;
        }
    }
};

} // namespace

void bvh8_collapse_params(float *node_cost, float *prim_cost, int *max_leaf) {
    *node_cost = 1.0f, *prim_cost = options().bvh_prim_cost, *max_leaf = kMaxLeaf;
}

void build_bvh8(const float *lo, const float *hi, size_t n, float pad, int n_threads, Bvh8 *out) {
    out->built_by = 1;
    auto t0 = std::chrono::steady_clock::now();
    out->pad = pad;
    if (n == 0) {
        out->nodes.assign(1, Node8());
        memset(&out->nodes[0], 0, sizeof(Node8));
        out->nodes[0].e[0] = out->nodes[0].e[1] = out->nodes[0].e[2] = 127;
        out->tri_order.clear();
        out->depth = 1;
        return;
    }
    std::vector<float> plo(3 * n), phi(3 * n);
    for (size_t i = 0; i < 3 * n; i++) plo[i] = lo[i] - pad, phi[i] = hi[i] + pad;
    for (int attempt = 0; attempt < 2; attempt++) {
        Builder b;
        b.lo = plo.data(), b.hi = phi.data(), b.n = n;
        b.median_only = attempt == 1; // balanced fallback if the SAH tree came out too deep for the stack
        b.run(n_threads);
        Collapser c(b, *out);
        c.run();
        if (out->depth <= kMaxBvh8Depth) break;
    }
    out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

} // namespace akb
