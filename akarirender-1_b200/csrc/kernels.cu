// kernels.cu — hand-written sm_100a kernels of the Path hot path (no Triton/OptiX, no tensor cores: this is
// latency/issue/L2-bound pointer-light traversal + shading, not a dense contraction).
//
// Wavefront organisation (replaces render_tile_wavefront, kernel/integrators/gpu/cuda/integrator.cpp:206-406, whose
// 11 full-tile launches per bounce use state-predicated early-outs, no compaction and AoS managed memory):
//   wf_init    : one slot per owned pixel; first camera ray of every pixel        ("Set Path State" + "Ray Generation")
//   wf_trace   : ONE launch per bounce over the compacted extension-ray queue (closest hit, CWBVH traversal) AND the
//                compacted shadow-ray queue of the previous bounce (any hit, adds the carried radiance to L);
//                closest hits are binned by BSDF kind into material-keyed shading queues with warp-aggregated atomics
//                ("Extension Ray" + "Shadow Ray" + "Hit Nothing")
//   wf_shade   : ONE launch per bounce over the key-sorted shading queues (each key padded to a warp, so warps are
//                material-homogeneous): BSDF build + sample, light select + sample, emits the next extension ray and
//                the shadow ray; finished paths add to the film accumulator and regenerate the pixel's next sample
//                in place ("Evaluate Material" x3 + "Splat" + "Ray Generation")
//   wf_resolve : slot accumulators -> pixel-indexed film (Film::merge_tile)
// All queue lengths live in HBM (Counters); kernels run a fixed persistent grid and read them, so a bounce costs two
// launches and no host round trip.
#include "kernels.cuh"

#include <algorithm>

namespace akb {
namespace {

#ifndef AKB_BLOCK
#define AKB_BLOCK 128 // threads per block of the traversal / shade kernels (the *_MIN_BLOCKS below count blocks of this size)
#endif
constexpr int kBlock = AKB_BLOCK;
#ifndef AKB_TRACE_MIN_BLOCKS
#define AKB_TRACE_MIN_BLOCKS 8 // traversal kernels: 64 registers, no spills; measured 4 / 6 / 7 / 8 / 9 / 10 blocks: 29.3 / 23.6 / 22.7 / 21.9 - 22.2 / 22.8 / 24.7 ms
#endif
#ifndef AKB_SHADE_MIN_BLOCKS
#define AKB_SHADE_MIN_BLOCKS 7 // shade kernel: 72 registers; measured 4 / 5 / 6 / 7 / 8 blocks: C3 shade stage 8.58 / 7.69 / 7.32 / 7.15 / 7.90 ms, frame 28.83 / 28.18 / 28.12 / 27.59 / 28.13 ms
#endif
#ifndef AKB_MEGA_MIN_BLOCKS
#define AKB_MEGA_MIN_BLOCKS 5 // megakernel: 96 registers; measured 4 / 5 / 6 / 8 blocks: C3 at 16 spp 35.5 / 33.1 / 33.8 / 34.1 ms, a 1/8 tile-split share 5.80 / 5.35 / 5.72 / 5.97 ms
#endif
#ifndef AKB_STREAM_MIN_BLOCKS
#define AKB_STREAM_MIN_BLOCKS 8 // the streaming kernel holds both jobs
#endif
constexpr uint32_t kDeadRay = 0x80000000u; // queue-entry flag: path ends at the next shade step, nothing to trace
constexpr uint32_t kHasShadow = 0x40000000u; // queue-entry flag: the path also has an entry in the shadow queue (wf_finish_kernel)
constexpr uint32_t kSlotMask = ~(kDeadRay | kHasShadow);

__device__ __forceinline__ v3 xyz(const float4 &a) { return mk3(a.x, a.y, a.z); }
__device__ __forceinline__ float4 mk4(v3 a, float w) { return make_float4(a.x, a.y, a.z, w); }
__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Warp-aggregated queue append; must be reached by all 32 lanes. One atomic per warp.
__device__ __forceinline__ void queue_push(bool pred, uint32_t value, uint32_t *queue, uint32_t *counter, uint32_t cap) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if ((int)(threadIdx.x & 31) == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    AKB_CHECK(!pred || base + __popc(m & lanemask_lt()) < cap);
    if (pred) queue[base + __popc(m & lanemask_lt())] = value;
}

// Two appends at once (the shade step's extension and shadow queues): lanes 0 and 1 issue the two atomics in ONE
// instruction, so the warp waits for one round trip to L2 instead of two.
__device__ __forceinline__ void queue_push2(bool p0, uint32_t v0, uint32_t *q0, uint32_t *c0, bool p1, uint32_t v1, uint32_t *q1,
                                            uint32_t *c1, uint32_t cap) {
    const unsigned m0 = __ballot_sync(0xffffffffu, p0), m1 = __ballot_sync(0xffffffffu, p1);
    if ((m0 | m1) == 0) return;
    const unsigned lane = threadIdx.x & 31;
    uint32_t *ctr = lane == 0 ? c0 : c1;
    const uint32_t cnt = (uint32_t)__popc(lane == 0 ? m0 : m1);
    uint32_t base = 0;
    if (lane < 2 && cnt) base = atomicAdd(ctr, cnt);
    const uint32_t b0 = __shfl_sync(0xffffffffu, base, 0), b1 = __shfl_sync(0xffffffffu, base, 1);
    AKB_CHECK(!p0 || b0 + __popc(m0 & lanemask_lt()) < cap);
    AKB_CHECK(!p1 || b1 + __popc(m1 & lanemask_lt()) < cap);
    if (p0) q0[b0 + __popc(m0 & lanemask_lt())] = v0;
    if (p1) q1[b1 + __popc(m1 & lanemask_lt())] = v1;
}

// ---------------------------------------------------------------------------------------------------------------------
__global__ void selftest_kernel(int *result, float a, float b, float c) {
    // (1+2^-12)^2 = 1 + 2^-11 + 2^-24 rounds to 1 + 2^-11, so a*b + c == 0 unless the compiler fused it (2^-24).
    float r = a * b + c;
    *result = (r == 0.0f) ? 1 : 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// Persistent traversal warps with ray replacement. Rays of one warp finish after very different numbers of node
// visits (ncu on the first version: 7.9 of 32 lanes active per issued instruction), so a warp does not wait for its
// slowest ray: whenever fewer than kRefillBelow lanes are still traversing, the finished lanes write their results
// and fetch new rays from the queue (one atomic per refill), and the warp continues with all lanes busy.
//   Src::load(i, T, tag)  -> true: T initialised for item i; false: the item needs no traversal (result ready)
//   Src::complete(have, i, tag, T): called by all 32 lanes together; lanes with have == true consume their result
__constant__ int c_no_park = 1; // 0 (AKB_NO_PARK=0): a path that ends with a shadow ray pending always waits one step for it
#define g_no_park c_no_park
__constant__ int c_fetch_chunk = 0; // 0: reserve exactly the idle-lane count per refill; >0: warp-private chunks (AKB_FETCH_CHUNK)
__constant__ int c_shade_partition = 0; // 1 (option shade_partition): a block regroups its 128 entries into hits | misses before shading (measured slower)
__constant__ int c_tail_share = 1; // option tail_share: 0 never, 1 in launches of >= tail_share_min rays per 8 threads, 2 always: a drained warp's idle lanes help its busy ones
__constant__ int c_tail_share_min = 16; // option tail_share_min, in rays per 8 threads of the launch (16 = two rays per thread)
__constant__ int c_refill_below = 0; // > 0: AKB_REFILL_BELOW override; else 24 with the if-if step, 20 with the fused one (measured best on B200)

#ifdef AKB_TAIL_PROF
__device__ unsigned long long g_tail_prof[6];
#endif
// The tail of a traversal launch. Once the queue is drained a warp cannot refill its idle lanes, and the launch lasts
// as long as its slowest ray: one dependent chain of node visits (C3: mean 11, maximum 83 visits; 32 rays alone take
// 70 us, 2 M rays 560 us). Here the lanes that have run out of rays take pending subtrees off the traversal stacks of
// the lanes still busy and traverse them FOR THE SAME RAY: a helper lane copies the ray and the owner's best hit so far,
// works through one stack entry (a node group or a parked triangle group) and hands its best hit back, which the owner
// folds in under the contract's rule (nearest t, lowest triangle id among equal t) - the result is the minimum over
// all accepted triangles whatever the order, so nothing changes but the latency. An any-hit ray ends for everybody
// working on it as soon as one of them finds an occluder. Helpers hand out work themselves (root = the ray's owner).
// Must be entered by all 32 lanes; returns when every ray of the warp has been completed through src.complete().
template <class Src>
__device__ __forceinline__ void tail_share(const float4 *__restrict__ nodes, const float4 *__restrict__ tris, Src &src, Trav &T, uint2 *stack,
                                           bool active, uint32_t item, uint32_t tag, int post_limit) {
    const unsigned full = 0xffffffffu;
    const int lane = (int)(threadIdx.x & 31);
    int root = -1;                         // >= 0: this lane is helping lane `root` with its ray
    bool own_done = false, h_done = false; // owner finished its own share and waits for helpers / helper has a result to hand back
    for (;;) {
#ifdef AKB_TAIL_PROF
        if (lane == 0) atomicAdd(&g_tail_prof[4], 1ull), atomicAdd(&g_tail_prof[5], (unsigned long long)__popc(__ballot_sync(full, active)));
        else __ballot_sync(full, active);
#endif
        if (trav_step_ii(T, nodes, tris, stack, active, post_limit)) {
            active = false;
            if (root >= 0)
                h_done = true;
            else
                own_done = true;
        }
        for (unsigned hd = __ballot_sync(full, h_done); hd; hd &= hd - 1) { // helpers hand their best hit to the owner
            const int l = __ffs(hd) - 1;
            const int own = __shfl_sync(full, root, l);
            if (own < 0) continue; // sent back to the idle pool a moment ago: another helper found the ray's occluder
            const float ht = __shfl_sync(full, T.hit.t, l), hu = __shfl_sync(full, T.hit.u, l), hv = __shfl_sync(full, T.hit.v, l);
            const int htri = __shfl_sync(full, T.hit.tri, l), hmat = __shfl_sync(full, T.hit.mat, l);
            const bool occluder = __shfl_sync(full, (int)T.any, l) != 0 && htri >= 0;
            if (lane == own) {
                if (occluder) {
                    T.hit.tri = 0, active = false, own_done = true;
                } else if (!T.any && htri >= 0 && (ht < T.tbest || (ht == T.tbest && T.hit.tri >= 0 && htri < T.hit.tri))) {
                    T.tbest = ht;
                    T.hit.t = ht, T.hit.u = hu, T.hit.v = hv, T.hit.tri = htri, T.hit.mat = hmat;
                }
            }
            if (lane == l || (occluder && root == own)) active = false, h_done = false, root = -1; // back to the idle pool
        }
        const unsigned grp = __match_any_sync(full, root >= 0 ? root : lane); // an owner and everybody helping it
        bool have = false;
        if (own_done && __popc(grp) == 1) own_done = false, have = true;
        if (__any_sync(full, have)) src.complete(have, item, tag, T);
        const unsigned idle = __ballot_sync(full, !active && !own_done), donors = __ballot_sync(full, active && T.sp > 0);
        if (idle && donors) {
            const int n_pair = min(__popc(idle), __popc(donors));
            const bool give = active && T.sp > 0 && __popc(donors & lanemask_lt()) < n_pair;
            const bool take = !active && !own_done && __popc(idle & lanemask_lt()) < n_pair;
            uint2 e = make_uint2(0u, 0u);
            if (give) e = stack[--T.sp]; // (handing out the bottom entry instead - the farthest subtree - measured the same)
            const int from = take ? (int)__fns(donors, 0, __popc(idle & lanemask_lt()) + 1) : lane;
            const float ox = __shfl_sync(full, T.o.x, from), oy = __shfl_sync(full, T.o.y, from), oz = __shfl_sync(full, T.o.z, from);
            const float dx = __shfl_sync(full, T.d.x, from), dy = __shfl_sync(full, T.d.y, from), dz = __shfl_sync(full, T.d.z, from);
            const float ix = __shfl_sync(full, T.idx, from), iy = __shfl_sync(full, T.idy, from), iz = __shfl_sync(full, T.idz, from);
            const float tmn = __shfl_sync(full, T.tmin, from), tbs = __shfl_sync(full, T.tbest, from);
            const uint32_t oct = __shfl_sync(full, T.oct_inv4, from);
            const int any = __shfl_sync(full, (int)T.any, from), r_root = __shfl_sync(full, root >= 0 ? root : lane, from);
            const float ht = __shfl_sync(full, T.hit.t, from), hu = __shfl_sync(full, T.hit.u, from), hv = __shfl_sync(full, T.hit.v, from);
            const int htri = __shfl_sync(full, T.hit.tri, from), hmat = __shfl_sync(full, T.hit.mat, from);
            const uint32_t ex = __shfl_sync(full, e.x, from), ey = __shfl_sync(full, e.y, from);
            if (take) {
                T.o = mk3(ox, oy, oz), T.d = mk3(dx, dy, dz), T.idx = ix, T.idy = iy, T.idz = iz, T.tmin = tmn, T.tbest = tbs;
                T.oct_inv4 = oct, T.any = any != 0;
                T.hit.t = ht, T.hit.u = hu, T.hit.v = hv, T.hit.tri = T.any ? -1 : htri, T.hit.mat = hmat;
                T.sp = 0, T.tri_popped = false;
                if (ey > 0x00FFFFFFu) // a node group: the remaining children of a visited node
                    T.ng = make_uint2(ex, ey), T.tri_base = 0, T.tmask = 0;
                else // a parked triangle group
                    T.ng = make_uint2(0u, 0u), T.tri_base = ex, T.tmask = ey;
                root = r_root, active = true;
            }
        }
        if (!__any_sync(full, active || own_done)) break;
    }
}

template <class Src>
__device__ __forceinline__ void persistent_trace(const float4 *__restrict__ nodes, const float4 *__restrict__ tris, Src &src,
                                                 uint32_t total, uint32_t *cursor, int post_limit) {
    const unsigned lane = threadIdx.x & 31;
    uint2 stack[kStackSize];
    Trav T;
    T.hit.tri = -1, T.hit.mat = -1, T.hit.u = T.hit.v = 0.f, T.hit.t = INFINITY, T.any = false;
    bool active = false, have = false;
    uint32_t item = 0, tag = 0;
    uint32_t w_next = 0, w_end = 0; // warp-private reserved range of the queue (one atomic per kFetchChunk items)
    bool drained = false;           // the global cursor has passed the end
    const int refill_below = c_refill_below > 0 ? c_refill_below : (post_limit >= 0 ? 24 : 20);
    const uint32_t fetch_chunk = (uint32_t)c_fetch_chunk;
    for (;;) {
        src.complete(have, item, tag, T);
        have = false;
        const unsigned idle = __ballot_sync(0xffffffffu, !active);
        if (idle && !(drained && w_next == w_end)) {
            if (w_next == w_end) {
                uint32_t base = 0;
                const uint32_t want = fetch_chunk ? fetch_chunk : (uint32_t)__popc(idle);
                if (lane == 0) base = atomicAdd(cursor, want);
                base = __shfl_sync(0xffffffffu, base, 0);
                w_next = min(base, total), w_end = min(base + want, total);
                if (w_end == total) drained = true;
            }
            const uint32_t avail = w_end - w_next;
            const uint32_t rank = (uint32_t)__popc(idle & lanemask_lt());
            if (!active && rank < avail) {
                item = w_next + rank;
                if (src.load(item, T, tag))
                    active = true;
                else
                    have = true;
            }
            w_next += min((uint32_t)__popc(idle), avail);
        }
        const bool no_more = drained && w_next == w_end;
        unsigned act = __ballot_sync(0xffffffffu, active);
        if (act == 0) {
            if (__any_sync(0xffffffffu, have)) continue;
            if (no_more) break;
            continue;
        }
        // (Batching the leaf-triangle tests across lanes - one Moeller-Trumbore round per iteration, or only once N
        // lanes wait - was measured on B200 and did not pay: the node block then runs with fewer lanes (20 instead of
        // 26 of 32) and total issued instructions stay within 5 %; profiles/r01c_notes.md.)
        // (only for launches with at least two rays per thread - option tail_share_min - or option tail_share = 2: a
        // launch with fewer is all tail, and there the helpers' extra node fetches cost more than the shorter chains
        // win. Measured, threshold in rays per thread 4 / 2 / 1 / 0.5: C3 frame 105.8 / 104.5 / 104.1 / 104.1 ms, C5
        // 59.9 / 59.2 / 59.1 / 59.0 ms, a 1/8 tile-split share of C3 29.3 / 29.2 / 34.6 / 34.1 ms; 2 M incoherent rays:
        // +5 % closest, +10 % any hit)
        if (no_more && post_limit >= 0 && (c_tail_share == 2 || (c_tail_share == 1 && 8ull * total >= (unsigned long long)c_tail_share_min * gridDim.x * blockDim.x))) {
            src.complete(have, item, tag, T); // entries the refill above found ready (parked paths)
            have = false;
            tail_share(nodes, tris, src, T, stack, active, item, tag, post_limit);
            break; // every lane's ray is finished and consumed
        }
        do {
#ifdef AKB_TAIL_PROF
            if (lane == 0) atomicAdd(&g_tail_prof[no_more ? 2 : 0], 1ull), atomicAdd(&g_tail_prof[no_more ? 3 : 1], (unsigned long long)__popc(act));
#endif
            // post_limit >= 0: the "if-if" step (one node visit + one triangle test per lane and iteration, traverse.h);
            // < 0: the fused step, which is better for tiny trees where a ray is a handful of steps
            if (post_limit >= 0 ? trav_step_ii(T, nodes, tris, stack, active, post_limit) : (active && trav_step(T, nodes, tris, stack)))
                active = false, have = true;
            act = __ballot_sync(0xffffffffu, active);
        } while (act != 0 && (no_more || __popc(act) >= refill_below));
    }
}

struct BatchSrc {
    const float4 *org, *dir, *shading;
    akb_hit *hits;
    int32_t *occ;
    bool any;
    __device__ __forceinline__ bool load(uint32_t i, Trav &T, uint32_t &) const {
        const float4 o = __ldcs(org + i), d = __ldcs(dir + i);
        trav_init(T, xyz(o), xyz(d), o.w, d.w, any);
        return true;
    }
    __device__ __forceinline__ void complete(bool have, uint32_t i, uint32_t, const Trav &T) const {
        if (!have) return;
        if (any) {
            __stcs(occ + i, T.hit.tri >= 0 ? 1 : 0);
        } else {
            akb_hit out;
            if (T.hit.tri >= 0) {
                const float4 *sr = shading + 6 * (size_t)T.hit.tri;
                out.t = T.hit.t, out.u = T.hit.u, out.v = T.hit.v;
                out.geom_id = (int)__float_as_uint(__ldg(sr + 4).w);
                out.prim_id = (int)__float_as_uint(__ldg(sr + 5).w);
            } else {
                out.t = INFINITY, out.u = 0.f, out.v = 0.f, out.prim_id = -1, out.geom_id = -1;
            }
            hits[i] = out;
        }
    }
};

__global__ void __launch_bounds__(kBlock, AKB_TRACE_MIN_BLOCKS)
    trace_batch_kernel(SceneDev S, const float4 *__restrict__ org, const float4 *__restrict__ dir, uint32_t n, int any_hit,
                       akb_hit *__restrict__ hits, int32_t *__restrict__ occ, uint32_t *cursor) {
    BatchSrc src{org, dir, S.shading, hits, occ, any_hit != 0};
    persistent_trace(S.nodes, S.tris, src, n, cursor, S.trav_post_limit);
}

__global__ void generate_rays_kernel(SceneDev S, const int32_t *xy, const float *u2, size_t n, float4 *org, float4 *dir) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    v3 o, d;
    camera_generate_ray(S.cam, u2[2 * i], u2[2 * i + 1], xy[2 * i], xy[2 * i + 1], o, d);
    org[i] = mk4(o, AKB_EPS); // Ray ctor defaults, common/math.h:192-194
    dir[i] = mk4(d, INFINITY);
}

// Film::write_image post-process (core/film.h:97-113): radiance/weight; linear_to_srgb in fp32 (common/color.h:57-61);
// DefaultImageWriter (core/image.cpp:48-53): clamp to [0,1], (int)round(x * 255.5) in double, clamp to [0,255].
// powf is evaluated in double and rounded once, like glibc's.
__global__ void film_to_srgb8_kernel(const float4 *rgbw, size_t n, uint8_t *out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p = rgbw[i];
    const float c[3] = {p.x, p.y, p.z};
    for (int k = 0; k < 3; k++) {
        float L = c[k];
        if (p.w != 0) L = L / p.w;
        float v = (L < 0.0031308f) ? (L * 12.92f) : (1.055f * (float)pow((double)L, (double)(1.0f / 2.4f)) - 0.055f);
        v = min_std(max_std(v, 0.0f), 1.0f);
        int q = (int)round((double)v * 255.5);
        out[3 * i + k] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Megakernel mode (Path { megakernel: true }): one thread per pixel runs GenericPathTracer::run_megakernel for all its
// samples, exactly the loop of cpu::PathTracer::render (kernel/integrators/cpu/integrator.cpp:121-135).
template <int SAMP>
__global__ void __launch_bounds__(kBlock, AKB_MEGA_MIN_BLOCKS) megakernel(SceneDev S, RenderConfig rc, float4 *__restrict__ film, Counters *ctr) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    int x = 0, y = 0;
    const bool valid = slot < rc.map.n_slots() && slot_to_pixel(rc.map, slot, x, y);
    unsigned long long n_closest = 0, n_shadow = 0, n_samples = 0;
    if (valid) {
        Rng<SAMP> rng;
        rng.set_sample_index((uint64_t)(x + y * rc.map.width) + rc.seed_offset);
        v3 sum = mk3(0, 0, 0);
        float weight = 0;
        for (int s = 0; s < rc.spp; s++) {
            PathState<SAMP> ps;
            ps.L = mk3(0, 0, 0), ps.beta = mk3(1, 1, 1), ps.depth = 0, ps.rng = rng;
            float u1x, u1y, u2x, u2y;
            ps.rng.next2d(u2x, u2y); // g++ evaluates generate_ray's arguments right to left: u2 first
            ps.rng.next2d(u1x, u1y);
            RaySeg ray;
            camera_generate_ray(S.cam, u2x, u2y, x, y, ray.o, ray.d);
            ray.tmin = AKB_EPS, ray.tmax = INFINITY;
            for (;;) {
                RayHit hit;
                n_closest++;
                if (!traverse8<false>(S.nodes, S.tris, ray.o, ray.d, ray.tmin, ray.tmax, hit)) break;
                RaySeg next, sh;
                v3 sc;
                const int fl = path_vertex<SAMP>(S, ps, ray.d, hit.u, hit.v, hit.tri, rc.max_depth, rc.rr_start, next, sh, sc);
                if (fl & VTX_SHADOW) {
                    RayHit dummy;
                    n_shadow++;
                    if (!traverse8<true>(S.nodes, S.tris, sh.o, sh.d, sh.tmin, sh.tmax, dummy)) ps.L = add3(ps.L, sc);
                }
                if (!(fl & VTX_CONTINUE)) break;
                ray = next;
            }
            rng = ps.rng;
            weight += 1.0f;
            sum = add3(sum, clamp_sample(ps.L, rc.ray_clamp));
            n_samples++;
        }
        film[(size_t)y * rc.map.width + x] = mk4(sum, weight);
    }
    for (int o = 16; o > 0; o >>= 1) {
        n_closest += __shfl_down_sync(0xffffffffu, n_closest, o);
        n_shadow += __shfl_down_sync(0xffffffffu, n_shadow, o);
        n_samples += __shfl_down_sync(0xffffffffu, n_samples, o);
    }
    if ((threadIdx.x & 31) == 0 && n_samples) {
        atomicAdd(&ctr->stat_closest, n_closest);
        atomicAdd(&ctr->stat_shadow, n_shadow);
        atomicAdd(&ctr->stat_samples, n_samples);
    }
}

// AO integrator (`AO { spp, occlude }`, core/nodes/integrator.cpp:26-49): one thread per owned pixel runs the sample loop
// of AmbientOcclusion::render (kernel/integrators/cpu/integrator.cpp:72-80): camera ray -> closest hit -> one cosine ray.
// Film semantics are the CPU integrator's (add_sample(L, 1) per sample), like the Path integrator's.
template <int SAMP>
__global__ void __launch_bounds__(kBlock) ao_kernel(SceneDev S, RenderConfig rc, float occlude, float4 *__restrict__ film,
                                                    Counters *ctr) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    int x = 0, y = 0;
    const bool valid = slot < rc.map.n_slots() && slot_to_pixel(rc.map, slot, x, y);
    unsigned long long n_closest = 0, n_shadow = 0, n_samples = 0;
    if (valid) {
        Rng<SAMP> rng;
        rng.set_sample_index((uint64_t)(x + y * rc.map.width) + rc.seed_offset);
        float sum = 0, weight = 0;
        for (int s = 0; s < rc.spp; s++) {
            float u1x, u1y, u2x, u2y;
            rng.next2d(u2x, u2y);
            rng.next2d(u1x, u1y);
            v3 o, d;
            camera_generate_ray(S.cam, u2x, u2y, x, y, o, d);
            RayHit hit, dummy;
            float L = 0.0f;
            n_closest++;
            if (traverse8<false>(S.nodes, S.tris, o, d, AKB_EPS, INFINITY, hit)) {
                const RaySeg r = ao_ray<SAMP>(S, rng, hit.u, hit.v, hit.tri, occlude);
                n_shadow++;
                L = traverse8<true>(S.nodes, S.tris, r.o, r.d, r.tmin, r.tmax, dummy) ? 0.0f : 1.0f;
            }
            weight += 1.0f;
            sum += L;
            n_samples++;
        }
        film[(size_t)y * rc.map.width + x] = make_float4(sum, sum, sum, weight);
    }
    for (int o = 16; o > 0; o >>= 1) {
        n_closest += __shfl_down_sync(0xffffffffu, n_closest, o);
        n_shadow += __shfl_down_sync(0xffffffffu, n_shadow, o);
        n_samples += __shfl_down_sync(0xffffffffu, n_samples, o);
    }
    if ((threadIdx.x & 31) == 0 && n_samples) {
        atomicAdd(&ctr->stat_closest, n_closest);
        atomicAdd(&ctr->stat_shadow, n_shadow);
        atomicAdd(&ctr->stat_samples, n_samples);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Wavefront kernels.

// Starts sample `done` of the pixel owning `slot`: camera ray + fresh path state ("Ray Generation",
// gpu/cuda/integrator.cpp:236-256; sampler draw order of kernel/pathtracer.h:61-64 under g++).
template <int SAMP>
__device__ __forceinline__ void start_sample(const SceneDev &S, const Workspace &ws, int slot, int x, int y, Rng<SAMP> rng,
                                             int done) {
    float u1x, u1y, u2x, u2y;
    rng.next2d(u2x, u2y);
    rng.next2d(u1x, u1y);
    v3 o, d;
    camera_generate_ray(S.cam, u2x, u2y, x, y, o, d);
    ws.ray_o[slot] = mk4(o, AKB_EPS);
    ws.ray_d[slot] = mk4(d, INFINITY);
    ws.beta[slot] = make_float4(1.f, 1.f, 1.f, __int_as_float(0));
    ws.L[slot] = make_float4(0.f, 0.f, 0.f, __int_as_float(done));
    ws.rng[slot] = rng.s;
}

template <int SAMP> __global__ void __launch_bounds__(256) wf_init_kernel(SceneDev S, RenderConfig rc, Workspace ws) {
    const int n = ws.n_slots;
    const int n_pad = (n + 31) & ~31;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n_pad; slot += gridDim.x * blockDim.x) {
        int x = 0, y = 0;
        const bool valid = slot < n && slot_to_pixel(rc.map, slot + ws.slot_base, x, y) && rc.spp > 0;
        if (slot < n) ws.acc[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid) {
            Rng<SAMP> rng;
            rng.set_sample_index((uint64_t)(x + y * rc.map.width) + rc.seed_offset);
            start_sample<SAMP>(S, ws, slot, x, y, rng, 0);
        }
        queue_push(valid, (uint32_t)slot, ws.q_ext[0], &ws.ctr->n_ext[0], (uint32_t)ws.n_slots);
    }
}

// One launch per bounce: items [0, n_ext) are extension rays (closest hit), [n_ext, n_ext + n_sh) the shadow rays the
// previous shade step emitted (any hit; unoccluded ones add their carried radiance to the path's L).
// Path state read by the traversal step is read once per launch: how those loads treat L1 (which the BVH nodes live in)
// is a build switch for A/B runs: 0 plain loads, 1 ld.global.cs (streaming), 2 ld.global.cg (past L1). Measured on C3
// (16 spp): 27.62 - 27.64 / 27.76 - 27.89 / 27.71 - 27.73 ms - nothing to win: the default stays plain loads.
#ifndef AKB_WF_STREAM_LOADS
#define AKB_WF_STREAM_LOADS 0
#endif
#if AKB_WF_STREAM_LOADS == 1
#define AKB_WF_LD(p) __ldcs(p)
#elif AKB_WF_STREAM_LOADS == 2
#define AKB_WF_LD(p) __ldcg(p)
#else
#define AKB_WF_LD(p) (*(p))
#endif
template <int SAMP> struct WfSrc {
    const SceneDev &S;
    const RenderConfig &rc;
    const Workspace &ws;
    int b;
    uint32_t n_ext;
    __device__ __forceinline__ bool load(uint32_t i, Trav &T, uint32_t &slot) const {
        if (i < n_ext) {
            const uint32_t e = AKB_WF_LD(ws.q_ext[b] + i);
            slot = e & kSlotMask;
            AKB_CHECK(slot < (uint32_t)ws.n_slots);
            if (e & kDeadRay) { // parked path: nothing to trace, the shade step just finishes it
                T.hit.tri = -1, T.hit.mat = -1, T.any = false;
                return false;
            }
            const float4 o = AKB_WF_LD(ws.ray_o + slot), d = AKB_WF_LD(ws.ray_d + slot);
            trav_init(T, xyz(o), xyz(d), o.w, d.w, false);
        } else {
            slot = AKB_WF_LD(ws.q_sh[b] + (i - n_ext));
            AKB_CHECK(slot < (uint32_t)ws.n_slots);
            const float4 o = AKB_WF_LD(ws.sh_o + slot), d = AKB_WF_LD(ws.sh_d + slot);
            trav_init(T, xyz(o), xyz(d), o.w, d.w, true);
        }
        return true;
    }
    __device__ __forceinline__ void complete(bool have, uint32_t, uint32_t slot, const Trav &T) const {
        const unsigned lane = threadIdx.x & 31;
        int key = 0xFF;
        if (have) {
            if (!T.any) {
                ws.hit[slot] = make_float4(T.hit.u, T.hit.v, __int_as_float(T.hit.tri), __int_as_float(T.hit.mat));
                if (rc.sort_materials) {
                    key = 0;
                    if (rc.sort_materials == 2) { // two queues only: hits | misses and parked entries
                        key = T.hit.tri >= 0 ? 1 : 0;
                    } else if (T.hit.tri >= 0 && T.hit.mat >= 0) {
                        // BSDF kind this vertex will evaluate: resolve Mix with the number the shade step will draw
                        // (second draw from the path's current sampler state = x of the next next2d()).
                        int type = S.mats[T.hit.mat].type;
                        if (type == AKB_MAT_MIX) {
                            Rng<SAMP> peek;
                            peek.s = ws.rng[slot];
                            float ux, uy, cp;
                            peek.next2d(ux, uy);
                            float tu = 0.f, tv = 0.f;
                            if (S.n_textures)
                                tri_texcoord(__ldg(S.tri_tc + 2 * (size_t)T.hit.tri), __ldg(S.tri_tc + 2 * (size_t)T.hit.tri + 1),
                                             T.hit.u, T.hit.v, tu, tv);
                            type = S.mats[select_material(S, T.hit.mat, ux, tu, tv, cp)].type;
                        }
                        key = type == AKB_MAT_DIFFUSE ? 1 : (type == AKB_MAT_GLOSSY ? 2 : (type == AKB_MAT_DIELECTRIC ? 3 : 0));
                    }
                }
            } else {
                const float4 c = AKB_WF_LD(ws.sh_c + slot);
                if (c.w != 0.f) {
                    // last shadow ray of a sample whose pixel is already on its next one: this is the film add the
                    // shade step would have done one bounce later ("Splat", core/film.h:66-70). The pixel's next
                    // sample cannot finish before the next shade step, so the sums keep their order.
                    const float4 a = AKB_WF_LD(ws.acc + slot);
                    const v3 s = add3(xyz(a), T.hit.tri < 0 ? xyz(c) : xyz(AKB_WF_LD(ws.sh_b + slot)));
                    ws.acc[slot] = make_float4(s.x, s.y, s.z, a.w + 1.0f);
                } else if (T.hit.tri < 0) { // unoccluded: L += color, kernel/pathtracer.h:156
                    float4 L = AKB_WF_LD(ws.L + slot);
                    const v3 s = add3(xyz(L), xyz(c));
                    L.x = s.x, L.y = s.y, L.z = s.z;
                    ws.L[slot] = L;
                }
            }
        }
        if (rc.sort_materials) { // bin closest hits by BSDF kind: one atomic per (warp, key)
            const unsigned grp = __match_any_sync(0xffffffffu, key);
            if (key != 0xFF) {
                const int leader = __ffs(grp) - 1;
                uint32_t qb = 0;
                if ((int)lane == leader) qb = atomicAdd(&ws.ctr->n_shade[b][key], (uint32_t)__popc(grp));
                qb = __shfl_sync(grp, qb, leader);
                AKB_CHECK(qb + __popc(grp & lanemask_lt()) < (uint32_t)ws.n_slots);
                ws.q_shade[b][key][qb + __popc(grp & lanemask_lt())] = slot;
            }
        }
    }
};

template <int SAMP>
__global__ void __launch_bounds__(kBlock, AKB_TRACE_MIN_BLOCKS) wf_trace_kernel(SceneDev S, RenderConfig rc, Workspace ws, int b) {
    Counters *ctr = ws.ctr;
    const uint32_t n_ext = ctr->n_ext[b], n_sh = ctr->n_sh[b];
    WfSrc<SAMP> src{S, rc, ws, b, n_ext};
    persistent_trace(S.nodes, S.tris, src, n_ext + n_sh, &ctr->fetch_trace, S.trav_post_limit);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        // queues the NEXT shade step fills must start empty; nobody else touches them during this launch
        ctr->n_ext[b ^ 1] = 0;
        ctr->n_sh[b ^ 1] = 0;
        ctr->stat_closest += n_ext; // includes parked entries; the shade step counts those in stat_skipped
        ctr->stat_shadow += n_sh;
    }
}

// AO = true: the AO integrator on the same queues (kernel/integrators/cpu/integrator.cpp:72-80): a hit vertex emits one
// cosine-weighted any-hit ray carrying (1,1,1) and the path parks for one step, so that the sample is 1 iff that ray
// is unoccluded; a miss finishes with 0. Sampler order and film sums are those of ao_kernel.
template <int SAMP, bool TEX, bool AO>
__global__ void __launch_bounds__(kBlock, AKB_SHADE_MIN_BLOCKS) wf_shade_kernel(SceneDev S, RenderConfig rc, Workspace ws, int b) {
    Counters *ctr = ws.ctr;
    // Work list: with material sorting the kNumKeys keyed queues, each padded to a warp so that warps are
    // material-homogeneous; without it the extension queue itself (every traced or parked path gets shaded).
    uint32_t seg[kNumKeys + 1];
    uint32_t cnt[kNumKeys];
    seg[0] = 0;
#pragma unroll
    for (int k = 0; k < kNumKeys; k++) {
        cnt[k] = rc.sort_materials ? ctr->n_shade[b][k] : (k == 0 ? ctr->n_ext[b] : 0u);
        seg[k + 1] = seg[k] + ((cnt[k] + 31u) & ~31u);
    }
    const uint32_t total = seg[kNumKeys];
    const uint32_t lane = threadIdx.x & 31;
    uint32_t n_done = 0, n_parked = 0;
    // Work is taken a block at a time: 128 consecutive queue entries, one warp per 32 of them, in queue order (which is
    // near slot order, so the gathers of path state touch few lines). Option shade_partition = 1 first deals a block's
    // entries out again so that warps are homogeneous - hits, which run path_vertex, then misses and parked entries,
    // which only add to the film and regenerate; a STABLE two-way partition - and was measured SLOWER on B200 (C3 shade
    // stage 8.84 vs 7.30 ms, C5 8.97 vs 6.47 ms; profiles/r02_notes.md): the three block barriers per round tie the
    // four warps of a block to its slowest one, and the regrouped warps' gathers span twice the lines. Likewise two
    // queues filled by the traversal step (hits | misses: sort_materials = 2) and the four BSDF-keyed queues
    // (sort_materials = 1) lose to plain queue order: this kernel waits on memory, not on divergent issue slots.
    __shared__ uint32_t s_slot[kBlock];
    __shared__ float4 s_hit[kBlock];
    __shared__ uint32_t s_warp_hits[kBlock / 32];
    const bool partition = !rc.sort_materials && c_shade_partition;
    for (uint32_t blk = blockIdx.x * kBlock; blk < total; blk += gridDim.x * kBlock) {
        const uint32_t base = blk + (threadIdx.x & ~31u);
        int k = 0;
#pragma unroll
        for (int q = 1; q < kNumKeys; q++)
            if (base >= seg[q]) k = q;
        const uint32_t j = base - seg[k] + lane;
        bool valid = j < cnt[k];
        bool push_ext = false, push_sh = false;
        uint32_t ext_entry = 0, slot = 0;
        float4 hrec = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid) {
            slot = (rc.sort_materials ? ws.q_shade[b][k][j] : ws.q_ext[b][j]) & kSlotMask;
            hrec = ws.hit[slot];
        }
        if (partition) {
            const bool is_hit = valid && __float_as_int(hrec.z) >= 0;
            const unsigned mh = __ballot_sync(0xffffffffu, is_hit), mv = __ballot_sync(0xffffffffu, valid);
            if (lane == 0) s_warp_hits[threadIdx.x >> 5] = (uint32_t)__popc(mh) | ((uint32_t)__popc(mv) << 16);
            __syncthreads();
            uint32_t hits_before = 0, hits_total = 0, valid_before = 0, valid_total = 0;
#pragma unroll
            for (int w = 0; w < kBlock / 32; w++) {
                const uint32_t c = s_warp_hits[w];
                if (w < (int)(threadIdx.x >> 5)) hits_before += c & 0xFFFFu, valid_before += c >> 16;
                hits_total += c & 0xFFFFu, valid_total += c >> 16;
            }
            if (valid) {
                const uint32_t r_hit = hits_before + (uint32_t)__popc(mh & lanemask_lt());
                const uint32_t r_all = valid_before + (uint32_t)__popc(mv & lanemask_lt());
                const uint32_t dst = is_hit ? r_hit : hits_total + (r_all - r_hit);
                s_slot[dst] = slot, s_hit[dst] = hrec;
            }
            __syncthreads();
            valid = threadIdx.x < valid_total;
            if (valid) slot = s_slot[threadIdx.x], hrec = s_hit[threadIdx.x];
            __syncthreads(); // the arrays are rewritten by the next round
        }
        if (valid) {
            const float hit_u = hrec.x, hit_v = hrec.y;
            const int tri = __float_as_int(hrec.z);
            bool slot_done = false; // (only the streaming kernel counts finished slots)
#define AKB_WS_LD(p) (*(p))
#include "shade_vertex.inc"
#undef AKB_WS_LD
            (void)slot_done;
        }
        if (push_sh) ext_entry |= kHasShadow; // a shadow ray always comes with an extension entry (continuing or parked)
        AKB_CHECK(!valid || slot < (uint32_t)ws.n_slots);
        queue_push2(push_ext, ext_entry, ws.q_ext[b ^ 1], &ctr->n_ext[b ^ 1], push_sh, slot, ws.q_sh[b ^ 1], &ctr->n_sh[b ^ 1], (uint32_t)ws.n_slots);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < kNumKeys; k++) ctr->n_shade[b ^ 1][k] = 0;
        ctr->fetch_trace = 0; // the next traversal launch starts fetching at item 0
    }
    for (int o = 16; o > 0; o >>= 1) {
        n_done += __shfl_down_sync(0xffffffffu, n_done, o);
        n_parked += __shfl_down_sync(0xffffffffu, n_parked, o);
    }
    if (lane == 0 && n_done) atomicAdd(&ctr->stat_samples, (unsigned long long)n_done);
    if (lane == 0 && n_parked) atomicAdd(&ctr->stat_skipped, (unsigned long long)n_parked);
}

// Tail of a frame: once few paths are left, every further bounce costs two launches that each last as long as their
// slowest ray, with the machine nearly empty (C3: the last 90 of 220 bounces took 12 % of the frame). This kernel is
// launched instead, after a shade step, when the extension queue has dropped below a fraction of the slots: one thread
// per queue entry takes over the path where the wavefront left it (pending shadow ray, parked or continuing path) and
// runs that pixel's remaining samples to completion, in the order and arithmetic of wf_trace/wf_shade (L += shadow
// colour before the next vertex, rays leaving a vertex at max_depth not traced, sequential film sums), so the film and
// the ray counts do not depend on when the switch happens.
template <int SAMP, bool TEX>
__global__ void __launch_bounds__(kBlock) wf_finish_kernel(SceneDev S, RenderConfig rc, Workspace ws, int b) {
    Counters *ctr = ws.ctr;
    const uint32_t n = ctr->n_ext[b];
    unsigned long long n_closest = 0, n_shadow = 0, n_samples = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t e = ws.q_ext[b][i];
        const uint32_t slot = e & kSlotMask;
        bool dead = (e & kDeadRay) != 0, pend = (e & kHasShadow) != 0;
        if (dead) n_closest++; // the shade step that parked it counted it in stat_skipped; the trace launch that would balance that never runs
        PathState<SAMP> ps;
        ps.rng.s = ws.rng[slot];
        ps.beta = mk3(1, 1, 1), ps.depth = 0, ps.L = mk3(0, 0, 0);
        const float4 Lr = ws.L[slot];
        v3 L = xyz(Lr);
        int done = __float_as_int(Lr.w);
        const float4 a = ws.acc[slot];
        v3 sum = xyz(a);
        float weight = a.w;
        RaySeg ray, sh;
        v3 sc = mk3(0, 0, 0);
        ray.o = ray.d = sh.o = sh.d = mk3(0, 0, 0), ray.tmin = ray.tmax = sh.tmin = sh.tmax = 0.f;
        if (!dead) {
            const float4 o = ws.ray_o[slot], d = ws.ray_d[slot], bd = ws.beta[slot];
            ray.o = xyz(o), ray.d = xyz(d), ray.tmin = o.w, ray.tmax = d.w;
            ps.beta = xyz(bd), ps.depth = __float_as_int(bd.w);
        }
        bool last_of_sample = false; // the pending shadow ray decides an already finished sample's film value
        v3 sb = mk3(0, 0, 0);
        if (pend) {
            const float4 o = ws.sh_o[slot], d = ws.sh_d[slot], c = ws.sh_c[slot];
            sh.o = xyz(o), sh.d = xyz(d), sh.tmin = o.w, sh.tmax = d.w;
            sc = xyz(c);
            if (c.w != 0.f) last_of_sample = true, sb = xyz(ws.sh_b[slot]);
        }
        int x = 0, y = 0;
        slot_to_pixel(rc.map, (int)slot + ws.slot_base, x, y);
        for (;;) {
            if (pend) { // wf_trace, any-hit item: unoccluded -> L += colour
                RayHit dummy;
                n_shadow++;
                const bool unoccluded = !traverse8<true>(S.nodes, S.tris, sh.o, sh.d, sh.tmin, sh.tmax, dummy);
                if (last_of_sample) // wf_trace's film add; the sample itself was counted by the shade step
                    sum = add3(sum, unoccluded ? sc : sb), weight += 1.0f, last_of_sample = false;
                else if (unoccluded)
                    L = add3(L, sc);
                pend = false;
            }
            bool finished = true;
            ps.L = mk3(0, 0, 0);
            if (!dead) { // wf_trace closest-hit item + wf_shade
                RayHit hit;
                n_closest++;
                if (traverse8<false>(S.nodes, S.tris, ray.o, ray.d, ray.tmin, ray.tmax, hit)) {
                    RaySeg next;
                    const int fl = path_vertex<SAMP>(S, ps, ray.d, hit.u, hit.v, hit.tri, rc.max_depth, rc.rr_start, next, sh, sc, TEX);
                    bool cont = (fl & VTX_CONTINUE) != 0;
                    pend = (fl & VTX_SHADOW) != 0;
                    if (cont && ps.depth >= rc.max_depth && ps.depth > 0) cont = false;
                    if (cont)
                        ray = next, finished = false;
                    else if (pend)
                        dead = true, finished = false; // the film add waits for this vertex's shadow ray
                }
            }
            if (finished) {
                sum = add3(sum, clamp_sample(add3(L, ps.L), rc.ray_clamp));
                weight += 1.0f;
                n_samples++;
                if (++done >= rc.spp) break;
                float u1x, u1y, u2x, u2y; // start_sample
                ps.rng.next2d(u2x, u2y);
                ps.rng.next2d(u1x, u1y);
                camera_generate_ray(S.cam, u2x, u2y, x, y, ray.o, ray.d);
                ray.tmin = AKB_EPS, ray.tmax = INFINITY;
                ps.beta = mk3(1, 1, 1), ps.depth = 0;
                L = mk3(0, 0, 0);
                dead = false;
            }
        }
        ws.acc[slot] = mk4(sum, weight);
    }
    for (int o = 16; o > 0; o >>= 1) {
        n_closest += __shfl_down_sync(0xffffffffu, n_closest, o);
        n_shadow += __shfl_down_sync(0xffffffffu, n_shadow, o);
        n_samples += __shfl_down_sync(0xffffffffu, n_samples, o);
    }
    if ((threadIdx.x & 31) == 0 && (n_samples | n_closest | n_shadow)) {
        atomicAdd(&ctr->stat_closest, n_closest);
        atomicAdd(&ctr->stat_shadow, n_shadow);
        atomicAdd(&ctr->stat_samples, n_samples);
    }
}

__global__ void wf_resolve_kernel(RenderConfig rc, Workspace ws, float4 *__restrict__ film) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    int x, y;
    if (slot < ws.n_slots && slot_to_pixel(rc.map, slot + ws.slot_base, x, y)) film[(size_t)y * rc.map.width + x] = ws.acc[slot];
}

// dst (this GPU) += the peers' films, read over NVLink; fixed rank order, so the sum is reproducible
__global__ void __launch_bounds__(256) film_sum_peers_kernel(float4 *__restrict__ dst, PeerFilms peers, size_t n_px) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_px; i += (size_t)gridDim.x * blockDim.x) {
        float4 a = dst[i];
        for (int k = 0; k < peers.n; k++) {
            const float4 b = __ldcs(peers.film[k] + i);
            a.x += b.x, a.y += b.y, a.z += b.z, a.w += b.w;
        }
        dst[i] = a;
    }
}

#ifndef AKB_NO_STREAM // (development switch: a build without the streaming kernel, for A/B runs)
// ---------------------------------------------------------------------------------------------------------------------
// Streaming wavefront (akb_set_option("stream", 1)): ONE persistent launch renders the whole share, with no barrier
// between bounces. The bounce-synchronous kernels above pay, on every one of a frame's spp x depth rounds, for the
// slowest ray of that round (the launch cannot end before it does) and, once the short pixels have used up their
// samples, for rounds that hold few paths: at 1080p the same samples cost 18 % more than at 8K, where rounds are 16
// times larger (tools/res_scaling.py). Here a path moves on as soon as ITS rays are traced:
//   ray ring  (shade -> trace): one 32-bit entry per path that has rays to trace - its shadow ray, if any, and then
//              its extension ray, traced back to back by the lane that took the entry (so a path has one message in
//              flight at a time and needs no counting of outstanding rays);
//   hit ring  (trace -> shade): one 16-byte entry per traced path: (slot | shadow result, triangle, u, v) - the whole
//              result travels in the message, written by ONE 16-byte store, so the traversal side needs no fence.
// Both rings are multi-producer / multi-consumer: a producer takes tickets with one atomic per warp, writes its
// entries, each stamped with the ticket's lap number (7 bits); a consumer takes tickets the same way and reads an entry
// once the stamp it expects shows up. Path state (rays, throughput, sampler, accumulator) stays in the slot-indexed
// arrays, is written only by the shade side, published by `fence.release.gpu` (MEMBAR without the L1 invalidation of
// __threadfence(): the traversal warps keep their BVH nodes in L1) before the ring entry, and read with ld.global.cg.
// Every per-slot operation happens in the order of the bounce-synchronous kernels (shadow result -> L or the film,
// then the vertex, samples in sequence), so the film is bit-identical to theirs (tests/test_stream.py).
//   Warps are not tied to a job. Warp 0 of every block is "shade-first" (the warp-0s of the blocks resident on an SM
// share one scheduler and its instruction cache): it shades whenever the hit ring holds anything and otherwise traces
// ONE batch of rays at a time, without replacement; the other warps trace with ray replacement and shade only when the
// ray ring has run dry and none of their rays is in flight. So nobody ever gives up a half-full warp to change jobs,
// and the split of the machine between the two jobs follows the two rings' lengths by itself.
__constant__ int c_stream_flex_mask = 1;      // bit w: warp w of a block prefers shading
__constant__ int c_stream_watchdog_ms = 4000; // a warp that finds nothing to do for this long aborts the frame (a lost message)

constexpr uint32_t kRShadow = 1u << 24, kRDead = 1u << 23, kRSlot = (1u << 23) - 1u;
constexpr uint32_t kQUnocc = 1u << 24, kQHadShadow = 1u << 23;

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t *p, uint32_t v) { asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint4 ld_relaxed4(const uint4 *p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed4(uint4 *p, uint4 v) {
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void fence_release() { asm volatile("fence.release.gpu;" ::: "memory"); }
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// wf_init_kernel's list of started slots -> the first entries of the ray ring (lap 0)
__global__ void stream_seed_kernel(Workspace ws) {
    const uint32_t n = ws.ctr->n_ext[0];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) ws.r_ring[i] = ws.q_ext[0][i] & kRSlot;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        ws.sc->r_tail = n;
        ws.sc->n_started = n;
        if (n == 0) ws.sc->done = 1;
    }
}

// Takes up to 32 entries off the hit ring (only if it holds at least min_backlog) and shades them; all 32 lanes enter.
// Returns false if there was nothing to take. [q_next, q_end) = tickets this warp holds and has not consumed yet.
template <int SAMP, bool TEX>
__device__ __forceinline__ bool stream_shade_batch(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, uint32_t min_backlog,
                                                   uint32_t n_started, uint32_t &q_next, uint32_t &q_end, uint32_t &n_done) {
    const unsigned full = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31;
    StreamCtl *sc = ws.sc;
    const uint32_t mask = (1u << ws.ring_shift) - 1u;
    if (q_next == q_end) {
        uint32_t n = 0, base = 0;
        if (lane == 0) {
            const int backlog = (int)(ld_relaxed(&sc->q_tail) - ld_relaxed(&sc->q_head));
            if (backlog >= (int)min_backlog) {
                n = (uint32_t)min(32, backlog);
                base = atomicAdd(&sc->q_head, n);
            }
        }
        n = __shfl_sync(full, n, 0), base = __shfl_sync(full, base, 0);
        if (n == 0) return false;
        q_next = base, q_end = base + n;
    }
    const bool cand = lane < q_end - q_next;
    const uint32_t tk = q_next + lane;
    uint4 e = make_uint4(0u, 0u, 0u, 0u);
    bool ready = false;
    unsigned waiting = 0;
    for (int spin = 0; spin < 32; spin++) { // entries whose producers hold the tickets are a store away; others may never come
        if (cand && !ready) {
            e = ld_relaxed4(ws.q_ring + (tk & mask));
            ready = (e.x >> 25) == ((tk >> ws.ring_shift) & 127u);
        }
        waiting = __ballot_sync(full, cand && !ready);
        if (waiting == 0) break;
        __nanosleep(64);
    }
    const uint32_t n_take = waiting ? (uint32_t)(__ffs(waiting) - 1) : q_end - q_next;
    if (n_take == 0) return false;
    q_next += n_take;
    const bool valid = lane < n_take;
    bool push_ext = false, push_sh = false, slot_done = false;
    uint32_t ext_entry = 0, n_parked = 0;
    const uint32_t slot = e.x & kRSlot;
    if (valid) {
        AKB_CHECK(slot < (uint32_t)ws.n_slots);
        if (e.x & kQHadShadow) { // what WfSrc::complete does for an any-hit item
            const float4 c = __ldcg(ws.sh_c + slot);
            const bool unoccluded = (e.x & kQUnocc) != 0;
            if (c.w != 0.f) {
                const float4 a = __ldcg(ws.acc + slot);
                const v3 s = add3(xyz(a), unoccluded ? xyz(c) : xyz(__ldcg(ws.sh_b + slot)));
                ws.acc[slot] = make_float4(s.x, s.y, s.z, a.w + 1.0f);
            } else if (unoccluded) {
                float4 L = __ldcg(ws.L + slot);
                const v3 s = add3(xyz(L), xyz(c));
                L.x = s.x, L.y = s.y, L.z = s.z;
                ws.L[slot] = L;
            }
        }
        const float hit_u = __uint_as_float(e.z), hit_v = __uint_as_float(e.w);
        const int tri = (int)e.y;
        constexpr bool AO = false;
#define AKB_WS_LD(p) __ldcg(p)
#include "shade_vertex.inc"
#undef AKB_WS_LD
    }
    fence_release(); // the rays and the path state written above reach L2 before the ring entries that announce them
    const unsigned pm = __ballot_sync(full, push_ext);
    if (pm) {
        const int leader = __ffs(pm) - 1;
        uint32_t base = 0;
        if ((int)lane == leader) base = atomicAdd(&sc->r_tail, (uint32_t)__popc(pm));
        base = __shfl_sync(full, base, leader);
        if (push_ext) {
            const uint32_t t = base + (uint32_t)__popc(pm & lanemask_lt());
            st_relaxed(ws.r_ring + (t & mask), (((t >> ws.ring_shift) & 127u) << 25) | (push_sh ? kRShadow : 0u) | ((ext_entry & kDeadRay) ? kRDead : 0u) | slot);
        }
    }
    const unsigned fm = __ballot_sync(full, slot_done);
    if (fm && lane == 0) {
        const uint32_t nf = (uint32_t)__popc(fm);
        if (atomicAdd(&sc->n_finished, nf) + nf == n_started) st_relaxed(&sc->done, 1u);
    }
    return true;
}

template <int SAMP, bool TEX>
__global__ void __launch_bounds__(kBlock, AKB_STREAM_MIN_BLOCKS) wf_stream_kernel(SceneDev S, RenderConfig rc, Workspace ws) {
    const unsigned full = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31;
    StreamCtl *sc = ws.sc;
    const bool flex = ((c_stream_flex_mask >> (threadIdx.x >> 5)) & 1) != 0;
    const uint32_t mask = (1u << ws.ring_shift) - 1u;
    const int post_limit = S.trav_post_limit;
    const int refill_below = c_refill_below > 0 ? c_refill_below : (post_limit >= 0 ? 24 : 20);
    const uint32_t n_started = sc->n_started; // written by stream_seed_kernel, a launch ago
    const float4 *__restrict__ nodes = S.nodes;
    const float4 *__restrict__ tris = S.tris;

    uint2 stack[kStackSize];
    Trav T;
    trav_init(T, mk3(0, 0, 0), mk3(0, 0, 1), 0.f, 0.f, false);
    bool active = false, have = false; // traversing / a traversal of this lane has just ended
    bool in_shadow = false, unocc = false;
    uint32_t ent = 0;                                       // the ray-ring entry this lane works on
    uint32_t r_next = 0, r_end = 0, q_next = 0, q_end = 0; // tickets held and not yet consumed (warp-uniform)
    bool starved = false;                                   // the last refill found no entry for an idle lane
    uint32_t n_closest = 0, n_shadow = 0, n_done = 0, n_batches = 0, n_flex_rays = 0, n_idle = 0;
    unsigned long long idle_since = 0;
#ifdef AKB_STREAM_PROF
    long long t_shade = 0, t_steps = 0, t_refill = 0, t_idle = 0, t_msg = 0, n_refill = 0, n_steps = 0, lanes_steps = 0;
    const long long t_begin = clock64();
#define PROF_T0() const long long prof_t0 = clock64()
#define PROF_ADD(var) var += clock64() - prof_t0
#else
#define PROF_T0() ((void)0)
#define PROF_ADD(var) ((void)0)
#endif

    // Readers take tickets only for entries the ring holds (head < tail when they looked); two readers racing for the
    // same entries leave the loser with tickets whose entries come later. Such tickets are never sat on: ray-ring
    // tickets are looked at on every pass of this loop (at most eight traversal steps or one shade batch apart), and a
    // warp whose hit-ring tickets have been served takes no new rays until it has shaded them - so an entry waits
    // microseconds for its reader, while a lap of a ring (>= 2^20 entries) takes the whole GPU hundreds of
    // microseconds to produce.
    for (;;) {
        // ---- traversals that ended: shadow ray -> the path's extension ray; extension ray (or none) -> message ----
        bool msg = false;
        if (have) {
            have = false;
            if (in_shadow) {
                in_shadow = false;
                unocc = T.hit.tri < 0;
                if (!(ent & kRDead)) {
                    const uint32_t slot = ent & kRSlot;
                    const float4 o = __ldcg(ws.ray_o + slot), d = __ldcg(ws.ray_d + slot);
                    trav_init(T, xyz(o), xyz(d), o.w, d.w, false);
                    active = true;
                    n_closest++;
                } else {
                    T.hit.tri = -1; // a parked path: nothing else to trace
                    msg = true;
                }
            } else {
                msg = true;
            }
        }
        const unsigned mm = __ballot_sync(full, msg);
        if (mm) {
            PROF_T0();
            const int leader = __ffs(mm) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(&sc->q_tail, (uint32_t)__popc(mm));
            base = __shfl_sync(full, base, leader);
            if (msg) {
                const uint32_t t = base + (uint32_t)__popc(mm & lanemask_lt());
                uint4 v;
                v.x = (((t >> ws.ring_shift) & 127u) << 25) | (unocc ? kQUnocc : 0u) | ((ent & kRShadow) ? kQHadShadow : 0u) | (ent & kRSlot);
                v.y = (uint32_t)T.hit.tri, v.z = __float_as_uint(T.hit.u), v.w = __float_as_uint(T.hit.v);
                st_relaxed4(ws.q_ring + (t & mask), v);
            }
            PROF_ADD(t_msg);
        }
        bool progress = mm != 0;
        unsigned act = __ballot_sync(full, active);

        // ---- no ray in flight in this warp: it is free to shade. Shade-first warps whenever the hit ring holds
        //      anything; the others once the ray ring has nothing for them; everybody the tickets already held ----
        if (act == 0) {
            bool did = false;
            PROF_T0();
            if (flex || starved || q_next != q_end) did = stream_shade_batch<SAMP, TEX>(S, rc, ws, 1u, n_started, q_next, q_end, n_done);
            PROF_ADD(t_shade);
            trav_init(T, mk3(0, 0, 0), mk3(0, 0, 1), 0.f, 0.f, false); // nothing of T lives across the shading code
            if (did) {
                n_batches++;
                starved = false;
                idle_since = 0;
                if (r_next == r_end) continue; // (else: look at the ray-ring tickets this warp holds first)
                progress = true;
            }
        }

        // ---- idle lanes take entries off the ray ring (those whose stamp has arrived, in ticket order). Shade-first
        //      warps take one batch at a time and only what is there; the others refill whenever lanes fall idle ----
        const unsigned idle = __ballot_sync(full, !active);
        bool hits_waiting = false; // this warp holds hit-ring tickets whose entries have arrived: no new rays, shade first
        if (q_next != q_end && idle) {
            uint32_t x = 0;
            if (lane == 0) x = ld_relaxed(&ws.q_ring[q_next & mask].x);
            x = __shfl_sync(full, x, 0);
            hits_waiting = (x >> 25) == ((q_next >> ws.ring_shift) & 127u);
        }
        if (idle && !hits_waiting && (!flex || act == 0)) {
            PROF_T0();
            if (r_next == r_end) {
                // only what the ring holds (a racing reader may still get there first: then the tickets are kept and
                // looked at on every pass until their entries arrive)
                uint32_t base = 0, want = (uint32_t)__popc(idle);
                if (lane == 0) {
                    want = (uint32_t)max(0, min((int)want, (int)(ld_relaxed(&sc->r_tail) - ld_relaxed(&sc->r_head))));
                    if (want) base = atomicAdd(&sc->r_head, want);
                }
                base = __shfl_sync(full, base, 0), want = __shfl_sync(full, want, 0);
                r_next = base, r_end = base + want;
            }
            const uint32_t avail = r_end - r_next;
            const uint32_t rank = (uint32_t)__popc(idle & lanemask_lt());
            const bool cand = !active && rank < avail;
            uint32_t e = 0;
            bool ready = false;
            if (cand) {
                const uint32_t t = r_next + rank;
                e = ld_relaxed(ws.r_ring + (t & mask));
                ready = (e >> 25) == ((t >> ws.ring_shift) & 127u);
            }
            const unsigned late = __ballot_sync(full, cand && !ready);
            const uint32_t n_want = min((uint32_t)__popc(idle), avail);
            const uint32_t n_take = late ? (uint32_t)__popc(idle & ((1u << (__ffs(late) - 1)) - 1u)) : n_want;
            if (cand && rank < n_take) {
                ent = e;
                const uint32_t slot = e & kRSlot;
                AKB_CHECK(slot < (uint32_t)ws.n_slots);
                unocc = false;
                if (flex) n_flex_rays++;
                if (e & kRShadow) {
                    const float4 o = __ldcg(ws.sh_o + slot), d = __ldcg(ws.sh_d + slot);
                    trav_init(T, xyz(o), xyz(d), o.w, d.w, true);
                    in_shadow = true, active = true;
                    n_shadow++;
                } else if (!(e & kRDead)) {
                    const float4 o = __ldcg(ws.ray_o + slot), d = __ldcg(ws.ray_d + slot);
                    trav_init(T, xyz(o), xyz(d), o.w, d.w, false);
                    in_shadow = false, active = true;
                    n_closest++;
                } else { // (a parked entry always carries a shadow ray; kept for completeness)
                    T.hit.tri = -1, in_shadow = false, have = true;
                }
            }
            r_next += n_take;
            starved = n_take < (uint32_t)__popc(idle);
            if (n_take) progress = true;
#ifdef AKB_STREAM_PROF
            n_refill++;
            if (cand && rank < n_take && active) { // wait for the ray: the refill's whole latency chain
                volatile float sink = T.o.x + T.d.x;
                (void)sink;
            }
#endif
            PROF_ADD(t_refill);
        }
        act = __ballot_sync(full, active);
        const unsigned hv0 = __ballot_sync(full, have);
        if (act == 0) {
            if (hv0 || progress) {
                idle_since = 0;
                continue;
            }
            // nothing to trace, nothing to shade: the frame is ending (or has just begun)
            uint32_t flags = 0;
            if (lane == 0) flags = ld_relaxed(&sc->done) | ld_relaxed(&sc->abort);
            if (__shfl_sync(full, flags, 0)) break;
            n_idle++;
            if (lane == 0) {
                const unsigned long long now = global_ns();
                if (idle_since == 0) idle_since = now;
                if (now - idle_since > (unsigned long long)c_stream_watchdog_ms * 1000000ull) st_relaxed(&sc->abort, 1u);
            }
            starved = true;
            {
                PROF_T0();
                __nanosleep(400);
                PROF_ADD(t_idle);
            }
            continue;
        }
        idle_since = 0;
        // ---- traversal steps: until enough lanes have finished to make a refill worth its atomics; a warp that cannot
        //      refill (shade-first: one batch at a time; ray ring empty) sends its results a few at a time ----
        int steps = 0;
        unsigned hv = hv0;
        PROF_T0();
        do {
#ifdef AKB_STREAM_PROF
            n_steps++, lanes_steps += __popc(act);
#endif
            if (post_limit >= 0 ? trav_step_ii(T, nodes, tris, stack, active, post_limit) : (active && trav_step(T, nodes, tris, stack)))
                active = false, have = true;
            act = __ballot_sync(full, active);
            hv = __ballot_sync(full, have);
            steps++;
        } while (act != 0 && (flex ? __popc(hv) < 8 : (__popc(act) >= refill_below || (starved && hv == 0 && steps < 8))));
        PROF_ADD(t_steps);
    }
#ifdef AKB_STREAM_PROF
    if (lane == 0) {
        const long long t_all = clock64() - t_begin;
        unsigned long long *pr = sc->prof + (flex ? 10 : 0);
        atomicAdd(pr + 0, (unsigned long long)t_all), atomicAdd(pr + 1, (unsigned long long)t_shade), atomicAdd(pr + 2, (unsigned long long)t_steps);
        atomicAdd(pr + 3, (unsigned long long)t_refill), atomicAdd(pr + 4, (unsigned long long)t_idle), atomicAdd(pr + 5, (unsigned long long)t_msg);
        atomicAdd(pr + 6, (unsigned long long)n_refill), atomicAdd(pr + 7, (unsigned long long)n_steps), atomicAdd(pr + 8, (unsigned long long)lanes_steps);
        atomicAdd(pr + 9, (unsigned long long)n_batches);
    }
#endif
    // ---- statistics ----
    for (int o = 16; o > 0; o >>= 1) {
        n_closest += __shfl_down_sync(full, n_closest, o);
        n_shadow += __shfl_down_sync(full, n_shadow, o);
        n_done += __shfl_down_sync(full, n_done, o);
        n_flex_rays += __shfl_down_sync(full, n_flex_rays, o);
    }
    if (lane == 0) {
        if (n_closest) atomicAdd(&sc->stat_closest, (unsigned long long)n_closest);
        if (n_shadow) atomicAdd(&sc->stat_shadow, (unsigned long long)n_shadow);
        if (n_done) atomicAdd(&sc->stat_samples, (unsigned long long)n_done);
        if (n_batches) atomicAdd(&sc->stat_shade_batches, (unsigned long long)n_batches);
        if (n_flex_rays) atomicAdd(&sc->stat_flex_rays, (unsigned long long)n_flex_rays);
        if (n_idle) atomicAdd(&sc->stat_idle_polls, (unsigned long long)n_idle);
    }
}

#endif // AKB_NO_STREAM

template <class K> int blocks_per_sm(K kernel, int block) {
    int n = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, block, 0) != cudaSuccess || n < 1) n = 1;
    return n;
}

} // namespace

cudaError_t set_debug_bounds(uint32_t n_nodes, uint32_t n_tris) {
#ifdef AKB_DEBUG_CHECKS
    const uint32_t v[2] = {n_nodes, n_tris};
    return cudaMemcpyToSymbol(c_dbg_bounds, v, sizeof v);
#else
    (void)n_nodes, (void)n_tris;
    return cudaSuccess;
#endif
}
#ifdef AKB_TAIL_PROF
cudaError_t tail_prof_read(unsigned long long *out6, bool reset) {
    cudaError_t e = cudaMemcpyFromSymbol(out6, g_tail_prof, 6 * sizeof(unsigned long long));
    if (e == cudaSuccess && reset) {
        const unsigned long long z[6] = {};
        e = cudaMemcpyToSymbol(g_tail_prof, z, sizeof z);
    }
    return e;
}
#endif
cudaError_t set_refill_below(int v) { return cudaMemcpyToSymbol(c_refill_below, &v, sizeof(int)); }
cudaError_t set_no_park(int v) { return cudaMemcpyToSymbol(c_no_park, &v, sizeof(int)); }
cudaError_t set_tail_share(int v, int min_rays_per_8_threads) {
    cudaError_t e = cudaMemcpyToSymbol(c_tail_share, &v, sizeof(int));
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_tail_share_min, &min_rays_per_8_threads, sizeof(int));
    return e;
}
cudaError_t set_shade_partition(int v) { return cudaMemcpyToSymbol(c_shade_partition, &v, sizeof(int)); }
cudaError_t set_fetch_chunk(int v) { return cudaMemcpyToSymbol(c_fetch_chunk, &v, sizeof(int)); }

cudaError_t launch_selftest(int *d_result, cudaStream_t st) {
    const float a = 1.0f + 1.0f / 4096.0f;
    selftest_kernel<<<1, 1, 0, st>>>(d_result, a, a, -(1.0f + 1.0f / 2048.0f));
    return cudaGetLastError();
}

cudaError_t launch_trace_batch(const SceneDev &S, const float4 *org, const float4 *dir, size_t n, int any_hit,
                               akb_hit *hits, int32_t *occ, uint32_t *fetch_cursor, int n_sms, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (n > 0x7fffffffu) return cudaErrorInvalidValue; // callers chunk
    static int occ_b = blocks_per_sm(trace_batch_kernel, kBlock);
    cudaError_t e = cudaMemsetAsync(fetch_cursor, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) return e;
    size_t want = (n + kBlock - 1) / kBlock;
    int grid = (int)std::min<size_t>(want, (size_t)n_sms * occ_b);
    trace_batch_kernel<<<grid, kBlock, 0, st>>>(S, org, dir, (uint32_t)n, any_hit, hits, occ, fetch_cursor);
    return cudaGetLastError();
}

cudaError_t launch_generate_rays(const SceneDev &S, const int32_t *xy, const float *u2, size_t n, float4 *org,
                                 float4 *dir, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    generate_rays_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(S, xy, u2, n, org, dir);
    return cudaGetLastError();
}

cudaError_t launch_film_to_srgb8(const float4 *rgbw, size_t n, uint8_t *out, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    film_to_srgb8_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(rgbw, n, out);
    return cudaGetLastError();
}

cudaError_t launch_megakernel(const SceneDev &S, const RenderConfig &rc, float4 *film, Counters *ctr, cudaStream_t st) {
    const int n = rc.map.n_slots();
    if (n == 0) return cudaSuccess;
    const int grid = (n + kBlock - 1) / kBlock;
    if (rc.sampler == AKB_SAMPLER_PCG)
        megakernel<AKB_SAMPLER_PCG><<<grid, kBlock, 0, st>>>(S, rc, film, ctr);
    else
        megakernel<AKB_SAMPLER_LCG><<<grid, kBlock, 0, st>>>(S, rc, film, ctr);
    return cudaGetLastError();
}

cudaError_t launch_ao(const SceneDev &S, const RenderConfig &rc, float occlude, float4 *film, Counters *ctr, cudaStream_t st) {
    const int n = rc.map.n_slots();
    if (n == 0) return cudaSuccess;
    const int grid = (n + kBlock - 1) / kBlock;
    if (rc.sampler == AKB_SAMPLER_PCG)
        ao_kernel<AKB_SAMPLER_PCG><<<grid, kBlock, 0, st>>>(S, rc, occlude, film, ctr);
    else
        ao_kernel<AKB_SAMPLER_LCG><<<grid, kBlock, 0, st>>>(S, rc, occlude, film, ctr);
    return cudaGetLastError();
}

cudaError_t launch_wf_init(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int n_sms, cudaStream_t st) {
    const int grid = std::max(1, std::min((ws.n_slots + 255) / 256, n_sms * 8));
    if (rc.sampler == AKB_SAMPLER_PCG)
        wf_init_kernel<AKB_SAMPLER_PCG><<<grid, 256, 0, st>>>(S, rc, ws);
    else
        wf_init_kernel<AKB_SAMPLER_LCG><<<grid, 256, 0, st>>>(S, rc, ws);
    return cudaGetLastError();
}

cudaError_t launch_wf_trace(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int parity, int n_sms,
                            cudaStream_t st) {
    static int occ_l = blocks_per_sm(wf_trace_kernel<AKB_SAMPLER_LCG>, kBlock),
               occ_p = blocks_per_sm(wf_trace_kernel<AKB_SAMPLER_PCG>, kBlock);
    if (rc.sampler == AKB_SAMPLER_PCG)
        wf_trace_kernel<AKB_SAMPLER_PCG><<<n_sms * occ_p, kBlock, 0, st>>>(S, rc, ws, parity);
    else
        wf_trace_kernel<AKB_SAMPLER_LCG><<<n_sms * occ_l, kBlock, 0, st>>>(S, rc, ws, parity);
    return cudaGetLastError();
}

cudaError_t launch_wf_shade(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int parity, int n_sms,
                            cudaStream_t st) {
    // six instantiations: sampler x (scene has image textures), and the AO integrator per sampler
    static int occ[2][3] = {{blocks_per_sm(wf_shade_kernel<AKB_SAMPLER_LCG, false, false>, kBlock), blocks_per_sm(wf_shade_kernel<AKB_SAMPLER_LCG, true, false>, kBlock),
                             blocks_per_sm(wf_shade_kernel<AKB_SAMPLER_LCG, false, true>, kBlock)},
                            {blocks_per_sm(wf_shade_kernel<AKB_SAMPLER_PCG, false, false>, kBlock), blocks_per_sm(wf_shade_kernel<AKB_SAMPLER_PCG, true, false>, kBlock),
                             blocks_per_sm(wf_shade_kernel<AKB_SAMPLER_PCG, false, true>, kBlock)}};
    const bool pcg = rc.sampler == AKB_SAMPLER_PCG, tex = S.n_textures > 0;
    const int which = rc.ao ? 2 : (tex ? 1 : 0);
    const int grid = n_sms * occ[pcg][which];
    if (pcg && which == 2)
        wf_shade_kernel<AKB_SAMPLER_PCG, false, true><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else if (which == 2)
        wf_shade_kernel<AKB_SAMPLER_LCG, false, true><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else if (pcg && tex)
        wf_shade_kernel<AKB_SAMPLER_PCG, true, false><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else if (pcg)
        wf_shade_kernel<AKB_SAMPLER_PCG, false, false><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else if (tex)
        wf_shade_kernel<AKB_SAMPLER_LCG, true, false><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else
        wf_shade_kernel<AKB_SAMPLER_LCG, false, false><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    return cudaGetLastError();
}

cudaError_t launch_wf_finish(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int parity, uint32_t n_upper,
                             cudaStream_t st) {
    const unsigned grid = std::max(1u, (n_upper + kBlock - 1) / kBlock);
    const bool pcg = rc.sampler == AKB_SAMPLER_PCG, tex = S.n_textures > 0;
    if (pcg && tex)
        wf_finish_kernel<AKB_SAMPLER_PCG, true><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else if (pcg)
        wf_finish_kernel<AKB_SAMPLER_PCG, false><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else if (tex)
        wf_finish_kernel<AKB_SAMPLER_LCG, true><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    else
        wf_finish_kernel<AKB_SAMPLER_LCG, false><<<grid, kBlock, 0, st>>>(S, rc, ws, parity);
    return cudaGetLastError();
}

cudaError_t launch_film_sum_peers(float4 *dst, const PeerFilms &peers, size_t n_px, int n_sms, cudaStream_t st) {
    if (n_px == 0 || peers.n == 0) return cudaSuccess;
    const int grid = (int)std::min<size_t>((n_px + 255) / 256, (size_t)n_sms * 8);
    film_sum_peers_kernel<<<grid, 256, 0, st>>>(dst, peers, n_px);
    return cudaGetLastError();
}

#ifdef AKB_NO_STREAM
cudaError_t set_stream_options(int, int) { return cudaSuccess; }
cudaError_t launch_wf_stream(const SceneDev &, const RenderConfig &, const Workspace &, int, cudaStream_t) { return cudaErrorNotSupported; }
#else
cudaError_t set_stream_options(int flex_mask, int watchdog_ms) {
    cudaError_t e = cudaMemcpyToSymbol(c_stream_flex_mask, &flex_mask, sizeof(int));
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_stream_watchdog_ms, &watchdog_ms, sizeof(int));
    return e;
}

cudaError_t launch_wf_stream(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int n_sms, cudaStream_t st) {
    static int occ[2][2] = {{blocks_per_sm(wf_stream_kernel<AKB_SAMPLER_LCG, false>, kBlock), blocks_per_sm(wf_stream_kernel<AKB_SAMPLER_LCG, true>, kBlock)},
                            {blocks_per_sm(wf_stream_kernel<AKB_SAMPLER_PCG, false>, kBlock), blocks_per_sm(wf_stream_kernel<AKB_SAMPLER_PCG, true>, kBlock)}};
    if (ws.n_slots == 0) return cudaSuccess;
    if (ws.n_slots > kStreamMaxSlots || ws.ring_shift > 25 || ws.n_slots > (1 << ws.ring_shift)) return cudaErrorInvalidValue;
    stream_seed_kernel<<<std::max(1, std::min((ws.n_slots + 255) / 256, n_sms * 8)), 256, 0, st>>>(ws);
    const bool pcg = rc.sampler == AKB_SAMPLER_PCG, tex = S.n_textures > 0;
    const int grid = n_sms * occ[pcg][tex];
    if (pcg && tex)
        wf_stream_kernel<AKB_SAMPLER_PCG, true><<<grid, kBlock, 0, st>>>(S, rc, ws);
    else if (pcg)
        wf_stream_kernel<AKB_SAMPLER_PCG, false><<<grid, kBlock, 0, st>>>(S, rc, ws);
    else if (tex)
        wf_stream_kernel<AKB_SAMPLER_LCG, true><<<grid, kBlock, 0, st>>>(S, rc, ws);
    else
        wf_stream_kernel<AKB_SAMPLER_LCG, false><<<grid, kBlock, 0, st>>>(S, rc, ws);
    return cudaGetLastError();
}
#endif // AKB_NO_STREAM

cudaError_t launch_wf_resolve(const RenderConfig &rc, const Workspace &ws, float4 *film, cudaStream_t st) {
    if (ws.n_slots == 0) return cudaSuccess;
    wf_resolve_kernel<<<(ws.n_slots + 255) / 256, 256, 0, st>>>(rc, ws, film);
    return cudaGetLastError();
}

} // namespace akb
