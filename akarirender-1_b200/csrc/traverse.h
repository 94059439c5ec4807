// traverse.h — closest-hit / any-hit traversal of the compressed 8-wide BVH (bvh8.h) for ONE ray.
//
// Replaces TBVHAccelerator::intersect / occlude (kernel/bvh-accelerator.h:488-547), intersectAABB (:89-103),
// intersect_leaf (:476-486), BVHIntersector (:651-664) and MeshInstance::intersect (kernel/instance.h:42-80).
//
// Result contract (what makes it a drop-in): the nearest t over ALL triangles that pass the reference's
// Moeller-Trumbore test, evaluated with the reference's exact fp32 operation order (no FMA, IEEE divide), with
// the reference's constants and comparisons (|a| < 1e-6 reject, inclusive u in [0,1], u+v <= 1, exclusive
// tmin < t < tmax, strict t < best). Among candidates with bit-identical t the lowest flat triangle id wins
// (the reference resolves such exact ties by its own BVH's visit order, which no other tree can reproduce).
// Box culling is conservative: quantised child boxes are supersets of padded triangle boxes, NaNs from 0*inf
// drop out of fminf/fmaxf, so the box test only ever skips work, never a geometric hit.
// One class of results no tree can pin: for a grazing ray whose determinant a is barely above the 1e-6 reject, the
// fp32 test can accept a triangle the ray geometrically misses, at a computed t whose point lies outside the
// triangle's (padded) box - e.g. a = 1.2e-6, computed t 42.748 where the exact value is 42.773. Brute force over all
// triangles reports that "hit"; a BVH (the reference's own included) reports it only if its boxes happen to lead
// there. Measured on C3: about 2.5e-8 of rays (2 of 2 M pixels at 16 spp differ between two different trees).
//
// The code is AKB_HD so that tests/hostcheck can run the very same traversal on the CPU against the oracle
// (test infrastructure only; the library exports no CPU path).
#pragma once
#include "vecmath.h"

namespace akb {

#ifdef __CUDA_ARCH__
#define AKB_LD4(p) __ldg(p)
__device__ __forceinline__ uint32_t akb_f2u(float f) { return __float_as_uint(f); }
__device__ __forceinline__ float akb_u2f(uint32_t u) { return __uint_as_float(u); }
__device__ __forceinline__ int akb_bfind(uint32_t x) { return 31 - __clz((int)x); }
__device__ __forceinline__ int akb_popc(uint32_t x) { return __popc(x); }
__device__ __forceinline__ uint32_t akb_sext_s8x4(uint32_t x) { // per byte: 0xFF if bit 7 set else 0x00
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(0u), "r"(0xba98u));
    return r;
}
#else
#define AKB_LD4(p) (*(p))
#ifndef __CUDACC__
struct float4 {
    float x, y, z, w;
};
struct uint2 {
    uint32_t x, y;
};
struct int4 {
    int x, y, z, w;
};
#endif
inline uint32_t akb_f2u(float f) {
    uint32_t u;
    __builtin_memcpy(&u, &f, 4);
    return u;
}
inline float akb_u2f(uint32_t u) {
    float f;
    __builtin_memcpy(&f, &u, 4);
    return f;
}
inline int akb_bfind(uint32_t x) { return 31 - __builtin_clz(x); }
inline int akb_popc(uint32_t x) { return __builtin_popcount(x); }
inline uint32_t akb_sext_s8x4(uint32_t x) {
    uint32_t r = 0;
    for (int i = 0; i < 4; i++)
        if (x & (0x80u << (8 * i))) r |= 0xFFu << (8 * i);
    return r;
}
#endif

struct RayHit {
    float t, u, v;
    int tri; // flat triangle id, -1 = miss
    int mat; // scene material index of the hit triangle (from the leaf record)
};

// MeshInstance::intersect, kernel/instance.h:42-80, on a leaf-order triangle record
// (r0 = v0|flat id, r1 = e1|material, r2 = e2|-). e1 = v1 - v0 and e2 = v2 - v0 are the reference's own first two
// subtractions, done once at scene build in the same fp32 arithmetic (bit-identical to doing them per ray).
// Returns t or a negative sentinel; u,v written on success.
AKB_HD bool mt_test(v3 o, v3 d, float tmin, float tmax, const float4 &r0, const float4 &r1, const float4 &r2,
                    float &t_out, float &u_out, float &v_out) {
    v3 v0 = mk3(r0.x, r0.y, r0.z), e1 = mk3(r1.x, r1.y, r1.z), e2 = mk3(r2.x, r2.y, r2.z);
    v3 h = cross3(d, e2);
    float a = dot3(e1, h);
    if (a > -1e-6f && a < 1e-6f) return false;
    float f = 1.0f / a;
    v3 s = sub3(o, v0);
    float u = f * dot3(s, h);
    if (u < 0.0f || u > 1.0f) return false;
    v3 q = cross3(s, e1);
    float v = f * dot3(d, q);
    if (v < 0.0f || u + v > 1.0f) return false;
    float t = f * dot3(e2, q);
    if (t > tmin && t < tmax) {
        t_out = t, u_out = u, v_out = v;
        return true;
    }
    return false;
}

constexpr int kStackSize = 32;
#ifdef __CUDACC__
static __constant__ uint32_t c_k47 = 0x47000000u; // opaque to ptxas on purpose, see AKB_Q2F
#endif

// 32768 + byte j of w, built with one byte permute: the byte lands in mantissa bits 15..8 of 2^15 (0x47000000), so
// no integer->float conversion (a quarter-rate XU instruction, 48 per node visit) is needed.
// (The 2^15 pattern is kept in a register the compiler cannot see through, so that the selector is the PRMT's
// immediate; otherwise ptxas encodes the pattern as the immediate and re-materialises a selector register per plane.)
#ifdef __CUDA_ARCH__
#define AKB_Q2F(w, j) __uint_as_float(__byte_perm((w), k47, 0x7404u | ((j) << 4)))
#define AKB_Q2F_SETUP() const uint32_t k47 = c_k47
#else
#define AKB_Q2F(w, j) akb_u2f(0x47000000u | ((((w) >> (8 * (j))) & 0xFFu) << 8))
#define AKB_Q2F_SETUP()
#endif

// Traversal state of one ray. `trav_step` performs one node visit (pop if needed, five 128-bit loads, eight
// quantised slab tests, then the Moeller-Trumbore tests of the leaf triangles that were hit), so a kernel can
// interleave fetching new rays for finished lanes (persistent warps with ray replacement).
struct Trav {
    v3 o, d;
    float idx, idy, idz, tmin, tbest;
    uint32_t oct_inv4;
    uint2 ng;
    uint32_t tri_base, tmask; // leaf triangles of the last node visit still to be tested
    int sp;
    bool any, tri_popped;
    RayHit hit;
#ifdef AKB_TRAV_STATS
    uint32_t n_nodes, n_tris, n_child_hits, n_empty;
#endif
};

AKB_HD void trav_init(Trav &T, v3 o, v3 d, float tmin, float tmax, bool any_hit) {
    const float ooeps = 8.271806125530277e-25f; // 2^-80: keeps 1/d finite; far below any geometric scale
    T.o = o, T.d = d;
    T.idx = 1.0f / (fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
    T.idy = 1.0f / (fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
    T.idz = 1.0f / (fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
    T.oct_inv4 = (7u - ((T.idx < 0 ? 1u : 0u) | (T.idy < 0 ? 2u : 0u) | (T.idz < 0 ? 4u : 0u))) * 0x01010101u;
    T.tmin = tmin, T.tbest = tmax;
    T.ng.x = 0, T.ng.y = 0x80000000u;
    T.tri_base = 0, T.tmask = 0;
    T.tri_popped = false;
    T.sp = 0;
    T.any = any_hit;
    T.hit.tri = -1, T.hit.t = INFINITY, T.hit.u = 0.f, T.hit.v = 0.f, T.hit.mat = -1;
}

// A traversal stack that would overflow silently drops the entry (release builds rely on the builders' depth bound,
// kMaxBvh8Depth, which leaves room); -DAKB_DEBUG_TRAVERSAL traps instead.
#if defined(AKB_DEBUG_TRAVERSAL) && defined(__CUDA_ARCH__)
#define AKB_STACK_OVERFLOW() __trap()
#elif defined(AKB_DEBUG_TRAVERSAL)
#define AKB_STACK_OVERFLOW() __builtin_trap()
#else
#define AKB_STACK_OVERFLOW() ((void)0)
#endif

// -DAKB_DEBUG_CHECKS: bounds checks of our own on every node / triangle / queue / slot index (compute-sanitizer is not
// available on the GPU pool): a violated check traps, which the host sees as a CUDA error from the next API call.
#if defined(AKB_DEBUG_CHECKS) && defined(__CUDA_ARCH__)
#define AKB_CHECK(cond)                                                                                                \
    do {                                                                                                               \
        if (!(cond)) __trap();                                                                                         \
    } while (0)
#else
#define AKB_CHECK(cond) ((void)0)
#endif
#if defined(AKB_DEBUG_CHECKS) && defined(__CUDACC__)
static __constant__ uint32_t c_dbg_bounds[2]; // node count, leaf-triangle count of the scene being rendered (kernels.cu sets them)
#endif

// THE node visit (the one copy all traversal steps share): takes the nearest pending child out of the node group `ng`
// (pushing what is left of the group), loads that child node (five 128-bit loads), intersects its eight quantised
// child boxes with the ray and returns the hit mask: bits 31..24 inner children in front-to-back order, bits 23..0
// leaf triangles. On return `ng` is the visited node's own group (child base | inner hits | imask) and `tri_base` its
// first leaf triangle.
AKB_HD uint32_t visit_node(const Trav &T, uint2 &ng, uint2 *stack, int &sp, const float4 *__restrict__ nodes, uint32_t &tri_base) {
    const uint32_t oct_inv = T.oct_inv4 & 7u;
    const uint32_t imask = ng.y & 0xFFu;
    const int bit = akb_bfind(ng.y);
    ng.y &= ~(1u << bit);
    if (ng.y > 0x00FFFFFFu) {
        if (sp < kStackSize)
            stack[sp++] = ng;
        else
            AKB_STACK_OVERFLOW();
    }
    const uint32_t slot = (uint32_t)(bit - 24) ^ oct_inv;
    const uint32_t rel = (uint32_t)akb_popc(imask & ~(0xFFFFFFFFu << slot));
    AKB_CHECK(ng.x + rel < c_dbg_bounds[0]);
    const float4 *np = nodes + 5 * (size_t)(ng.x + rel);
    const float4 n0 = AKB_LD4(np), n1 = AKB_LD4(np + 1), n2 = AKB_LD4(np + 2), n3 = AKB_LD4(np + 3), n4 = AKB_LD4(np + 4);
    const uint32_t ew = akb_f2u(n0.w);
    const float idx = T.idx, idy = T.idy, idz = T.idz;
    const float adjx = akb_u2f((ew & 0xFFu) << 23) * idx;
    const float adjy = akb_u2f(((ew >> 8) & 0xFFu) << 23) * idy;
    const float adjz = akb_u2f(((ew >> 16) & 0xFFu) << 23) * idz;
    // plane distance = q*adj + (p - o)*idir with q carried as 32768 + q: fold the offset into the constant term
    // (its rounding error, <= 2^-9 of a grid step, is covered by the builder's quantisation margin)
    const float orgx = fmaf(-32768.0f, adjx, (n0.x - T.o.x) * idx);
    const float orgy = fmaf(-32768.0f, adjy, (n0.y - T.o.y) * idy);
    const float orgz = fmaf(-32768.0f, adjz, (n0.z - T.o.z) * idz);
    ng.x = akb_f2u(n1.x);
    const float tmin = T.tmin, tbest = T.tbest;
    uint32_t hitmask = 0;
    AKB_Q2F_SETUP();
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const uint32_t meta4 = akb_f2u(half ? n1.w : n1.z);
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = akb_sext_s8x4(is_inner4 << 3);
        const uint32_t bit_index4 = (meta4 ^ (T.oct_inv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = akb_f2u(half ? n2.y : n2.x), qloy = akb_f2u(half ? n2.w : n2.z);
        const uint32_t qloz = akb_f2u(half ? n3.y : n3.x), qhix = akb_f2u(half ? n3.w : n3.z);
        const uint32_t qhiy = akb_f2u(half ? n4.y : n4.x), qhiz = akb_f2u(half ? n4.w : n4.z);
        const uint32_t nx = idx < 0 ? qhix : qlox, fx = idx < 0 ? qlox : qhix;
        const uint32_t ny = idy < 0 ? qhiy : qloy, fy = idy < 0 ? qloy : qhiy;
        const uint32_t nz = idz < 0 ? qhiz : qloz, fz = idz < 0 ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float tlx = fmaf(AKB_Q2F(nx, j), adjx, orgx);
            const float tly = fmaf(AKB_Q2F(ny, j), adjy, orgy);
            const float tlz = fmaf(AKB_Q2F(nz, j), adjz, orgz);
            const float thx = fmaf(AKB_Q2F(fx, j), adjx, orgx);
            const float thy = fmaf(AKB_Q2F(fy, j), adjy, orgy);
            const float thz = fmaf(AKB_Q2F(fz, j), adjz, orgz);
            const float cmin = fmaxf(fmaxf(tlx, tly), fmaxf(tlz, tmin));
            const float cmax = fminf(fminf(thx, thy), fminf(thz, tbest));
            if (cmin <= cmax) hitmask |= ((child_bits4 >> (8 * j)) & 0xFFu) << ((bit_index4 >> (8 * j)) & 0xFFu);
        }
    }
    ng.y = (hitmask & 0xFF000000u) | (ew >> 24);
    tri_base = akb_f2u(n1.y);
    return hitmask;
}

// THE leaf-triangle test (shared by all traversal steps): leaf triangle `tri_base + j` against the ray, result folded
// into the ray's best hit under the contract's tie rule. Returns true when an any-hit ray found its occluder.
AKB_HD bool test_triangle(Trav &T, const float4 *__restrict__ tris, uint32_t tri_base, int j) {
    AKB_CHECK(tri_base + (uint32_t)j < c_dbg_bounds[1]);
    const float4 *tp = tris + 3 * (size_t)(tri_base + (uint32_t)j);
    const float4 r0 = AKB_LD4(tp), r1 = AKB_LD4(tp + 1), r2 = AKB_LD4(tp + 2);
    float t, u, v;
    if (mt_test(T.o, T.d, T.tmin, T.any ? T.tbest : INFINITY, r0, r1, r2, t, u, v)) {
        if (T.any) {
            T.hit.tri = 0;
            return true;
        }
        const int id = (int)akb_f2u(r0.w);
        if (t < T.tbest || (t == T.tbest && T.hit.tri >= 0 && id < T.hit.tri)) {
            T.tbest = t;
            T.hit.t = t, T.hit.u = u, T.hit.v = v, T.hit.tri = id, T.hit.mat = (int)akb_f2u(r1.w);
        }
    }
    return false;
}

// One node visit as a step of its own (tests/hostcheck's scheduling simulations): pops a stack entry if the current
// node group is empty and leaves the inner-child hits in T.ng and the leaf-triangle hits in (T.tri_base, T.tmask). A
// popped entry may also be a postponed triangle group (y <= 0x00FFFFFF), which just becomes the current one. Returns
// false when there was nothing left to do.
AKB_HD bool trav_node(Trav &T, const float4 *__restrict__ nodes, uint2 *stack) {
    uint2 ng = T.ng;
    if (ng.y <= 0x00FFFFFFu) {
        if (T.sp == 0) return false;
        ng = stack[--T.sp];
        if (ng.y <= 0x00FFFFFFu) { // postponed triangle group
            T.tri_base = ng.x, T.tmask = ng.y;
            T.ng.y = 0;
            T.tri_popped = true; // must be tested now, not postponed again
            return true;
        }
    }
    const uint32_t hitmask = visit_node(T, ng, stack, T.sp, nodes, T.tri_base);
    T.ng = ng;
    T.tmask = hitmask & 0x00FFFFFFu;
    T.tri_popped = false;
#ifdef AKB_TRAV_STATS
    T.n_nodes++, T.n_empty += hitmask == 0;
    T.n_tris += (uint32_t)akb_popc(T.tmask), T.n_child_hits += (uint32_t)akb_popc(hitmask >> 24);
#endif
    return true;
}

// Tests ONE pending leaf triangle (T.tmask != 0). Returns true when an any-hit ray found its occluder.
AKB_HD bool trav_tri(Trav &T, const float4 *__restrict__ tris) {
    const int j = akb_bfind(T.tmask);
    T.tmask &= ~(1u << j);
    return test_triangle(T, tris, T.tri_base, j);
}

AKB_HD bool trav_idle(const Trav &T) { return T.ng.y <= 0x00FFFFFFu && T.sp == 0 && T.tmask == 0; }

// One node visit followed by all of its triangle tests, fused (the form for tiny trees and for one ray per thread: the
// current node group and the leaf hit mask live only in registers for the duration of the step). Returns true when
// the ray is finished (closest: tree exhausted; any-hit: first accepted triangle, hit.tri = 0). Postponed triangle
// groups are never on the stack here.
AKB_HD bool trav_step(Trav &T, const float4 *__restrict__ nodes, const float4 *__restrict__ tris, uint2 *stack) {
    uint2 ng = T.ng;
    if (ng.y <= 0x00FFFFFFu) {
        if (T.sp == 0) return true;
        ng = stack[--T.sp];
    }
    uint32_t tri_base;
    const uint32_t hitmask = visit_node(T, ng, stack, T.sp, nodes, tri_base);
    uint32_t tmask = hitmask & 0x00FFFFFFu;
#ifdef AKB_TRAV_STATS
    T.n_nodes++, T.n_empty += hitmask == 0;
    T.n_tris += (uint32_t)akb_popc(tmask), T.n_child_hits += (uint32_t)akb_popc(hitmask >> 24);
#endif
    while (tmask) {
        const int j = akb_bfind(tmask);
        tmask &= ~(1u << j);
        if (test_triangle(T, tris, tri_base, j)) return true;
    }
    T.ng = ng;
    return ng.y <= 0x00FFFFFFu && T.sp == 0;
}

// "if-if" step for warps of many rays (persistent_trace): per call every lane does at most ONE node visit and then at
// most ONE triangle test, and the two phases are separated by a warp barrier, so the Moeller-Trumbore block is issued
// once per iteration for all lanes that have a triangle pending (ncu on trav_step: it ran 2.8 times per node step at 3.9
// of 32 lanes = 41 % of the kernel's issued instructions). A node visit does not wait for the lane's pending triangles:
// a leftover triangle group is parked on the traversal stack (entries with y <= 0x00FFFFFF), which is sound because
// the result is the minimum over all accepted triangles in any order and culling with a not-yet-shrunk tbest only
// visits more. `post_limit` bounds how high the stack may be for parking (stack size - tree depth - 1), so inner
// entries always fit. Must be called by all 32 lanes; lanes with live == false do nothing. Returns true when the
// lane's ray is finished.
#ifdef __CUDA_ARCH__
#define AKB_WARP_SYNC() __syncwarp()
#else
#define AKB_WARP_SYNC()
#endif
AKB_HD bool trav_step_ii(Trav &T, const float4 *__restrict__ nodes, const float4 *__restrict__ tris, uint2 *stack, bool live,
                         int post_limit) {
    bool visit = live && (T.ng.y > 0x00FFFFFFu || T.sp > 0) && (T.tmask == 0 || T.sp < post_limit);
    if (visit) {
        uint2 ng = T.ng;
        if (ng.y <= 0x00FFFFFFu) {
            ng = stack[--T.sp];
            if (ng.y <= 0x00FFFFFFu) { // a parked triangle group
                if (T.tmask == 0)
                    T.tri_base = ng.x, T.tmask = ng.y;
                else
                    stack[T.sp++] = ng; // still busy with another group: leave it parked
                visit = false;
            }
        }
        if (visit) {
            uint32_t tri_base;
            const uint32_t hitmask = visit_node(T, ng, stack, T.sp, nodes, tri_base);
            T.ng = ng;
            const uint32_t tm = hitmask & 0x00FFFFFFu;
            if (tm) {
                if (T.tmask) stack[T.sp++] = uint2{T.tri_base, T.tmask}; // park the leftover group (room: post_limit)
                T.tri_base = tri_base, T.tmask = tm;
            }
        }
    }
#if defined(__CUDA_ARCH__) && defined(AKB_PREFETCH_NODES)
    // Experiment (-DAKB_PREFETCH_NODES=1 L1, =2 L2), measured SLOWER and therefore not built by default: the node the
    // NEXT visit will load is known here - the nearest pending child of the node group in hand, else of the group on
    // top of the stack - and a triangle test lies between here and that visit, so its two cache lines can be asked
    // for early (CCTL.E.PF1). C3 extend stage 24.2 ms against 22.0, 2 M incoherent rays 4129 against 4491 Mrays/s: the
    // 15 instructions it takes to work the address out cost more than the shorter scoreboard waits win - the kernel
    // pays for every instruction it issues (+8 % instructions per step, +9.7 % time).
    if (live) {
        uint2 g = T.ng;
        if (g.y <= 0x00FFFFFFu && T.sp > 0) g = stack[T.sp - 1];
        if (g.y > 0x00FFFFFFu) {
            const uint32_t slot = (uint32_t)(akb_bfind(g.y) - 24) ^ (T.oct_inv4 & 7u);
            const float4 *np = nodes + 5 * (size_t)(g.x + (uint32_t)akb_popc((g.y & 0xFFu) & ~(0xFFFFFFFFu << slot)));
#if AKB_PREFETCH_NODES == 2
            asm volatile("prefetch.global.L2 [%0];" ::"l"(np));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(np + 4));
#else
            asm volatile("prefetch.global.L1 [%0];" ::"l"(np));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(np + 4));
#endif
        }
    }
#endif
    AKB_WARP_SYNC();
    bool done = false;
    if (live && T.tmask) {
        const int j = akb_bfind(T.tmask);
        T.tmask &= ~(1u << j);
        done = test_triangle(T, tris, T.tri_base, j);
    }
    AKB_WARP_SYNC();
    return live && (done || (T.tmask == 0 && T.ng.y <= 0x00FFFFFFu && T.sp == 0));
}

// Run one ray to completion (megakernel, tests/hostcheck).
template <bool ANY>
AKB_HD bool traverse8(const float4 *__restrict__ nodes, const float4 *__restrict__ tris, v3 o, v3 d, float tmin,
                      float tmax, RayHit &hit) {
    Trav T;
    uint2 stack[kStackSize];
    trav_init(T, o, d, tmin, tmax, ANY);
    while (!trav_step(T, nodes, tris, stack)) {
    }
    hit = T.hit;
    return T.hit.tri >= 0;
}

} // namespace akb
