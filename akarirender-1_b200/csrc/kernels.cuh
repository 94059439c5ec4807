// kernels.cuh — launch interface of the sm_100a kernels (kernels.cu), used by the C ABI (api.cu).
#pragma once
#include <cuda_runtime.h>

#include "scene_host.h"

namespace akb {

constexpr int kNumKeys = 4; // shading-queue keys: 0 terminate/other, 1 diffuse, 2 GGX, 3 dielectric

// Device-resident counters of the wavefront (one struct in HBM). Index [2] = iteration parity.
struct Counters {
    uint32_t n_ext[2];             // extension-ray queue length
    uint32_t n_sh[2];              // shadow-ray queue length
    uint32_t n_shade[2][kNumKeys]; // material-keyed shading queues
    uint32_t fetch_trace;          // work-fetch cursor of the persistent traversal warps (reset by the previous kernel)
    uint32_t pad[3];
    unsigned long long stat_closest, stat_shadow, stat_samples, stat_skipped;
    unsigned long long prof[6]; // -DAKB_TAIL_PROF builds: traversal steps / active lanes summed, [0,1] while the queue has rays, [2,3] after it is drained, [4,5] inside tail_share
};

// Control block of the streaming wavefront (option stream = 1; one struct in HBM, zeroed before every frame). Tickets
// are free-running 32-bit counters (wrap-around arithmetic); the hot ones sit on cache lines of their own.
struct StreamCtl {
    uint32_t r_tail, pad0[31]; // ray ring (shade -> trace): tickets handed to producers ...
    uint32_t r_head, pad1[31]; // ... and to consumers (may run ahead of r_tail: consumers then wait for their entries)
    uint32_t q_tail, pad2[31]; // hit ring (trace -> shade)
    uint32_t q_head, pad3[31];
    uint32_t n_started;        // slots that render at least one sample (set by stream_seed_kernel)
    uint32_t n_finished;       // slots whose last sample is in the accumulator
    uint32_t done;             // n_finished == n_started: every warp leaves
    uint32_t abort;            // a warp waited longer than the watchdog allows: every warp leaves, the frame fails
    uint32_t pad4[28];
    unsigned long long stat_closest, stat_shadow, stat_samples, stat_shade_batches, stat_flex_rays, stat_idle_polls;
    unsigned long long prof[20]; // -DAKB_STREAM_PROF builds: cycles per activity, [0..9] tracing warps, [10..19] shade-first warps
};

// Path / ray state, vectorised SoA in HBM, indexed by slot (one slot per owned pixel; replaces the reference's
// SOA<PathState>, RayQueue, ShadowRayQueue, MaterialQueue: kernel/integrators/gpu/workitem.h:27-107).
struct Workspace {
    int n_slots;
    int slot_base;              // this view's first slot in the frame's slot map (two half-frame pipelines, api.cu)
    float4 *ray_o, *ray_d;      // extension ray: (o|tmin) (d|tmax)
    float4 *hit;                // (u, v, tri id bits, material bits), tri < 0 = miss
    float4 *beta;               // (beta | depth bits)
    float4 *L;                  // (L | samples-done bits)
    unsigned long long *rng;    // sampler state
    float4 *sh_o, *sh_d, *sh_c; // shadow ray + the radiance it carries (w = 0), or, for the last shadow ray of a sample
                                // whose pixel has already started its next one (w = 1): the sample's film value if it is
    float4 *sh_b;               // unoccluded (sh_c) and if it is occluded (sh_b)
    float4 *acc;                // per-slot film accumulator (sum L | sum weight): Tile::add_sample
    uint32_t *q_ext[2], *q_sh[2], *q_shade[2][kNumKeys];
    Counters *ctr;
    // streaming wavefront only: the two rings (1 << ring_shift entries each) and their control block
    uint32_t *r_ring; // [31:25] tag = (ticket >> ring_shift) & 127, [24] shadow ray pending, [23] no extension ray, [22:0] slot
    uint4 *q_ring;    // x: [31:25] tag, [24] the shadow ray was unoccluded, [23] there was a shadow ray, [22:0] slot; y: triangle (< 0: miss); z, w: u, v
    int ring_shift;
    StreamCtl *sc;
};

struct RenderConfig {
    int spp, max_depth, rr_start, sampler, sort_materials;
    int ao;            // AKB_INTEGRATOR_AO on the wavefront queues
    float ao_occlude;
    float ray_clamp;
    unsigned long long seed_offset; // added to the per-pixel sampler seed x + y*W (AKB_SPLIT_SAMPLE)
    SlotMap map;
};

struct LaunchDims {
    int n_sms, trace_blocks, shade_blocks;
};

cudaError_t launch_selftest(int *d_result, cudaStream_t st);
cudaError_t set_refill_below(int v);
#ifdef AKB_TAIL_PROF
cudaError_t tail_prof_read(unsigned long long *out6, bool reset); // development: where the traversal steps of a frame go
#endif
cudaError_t set_debug_bounds(uint32_t n_nodes, uint32_t n_tris); // -DAKB_DEBUG_CHECKS builds only; a no-op otherwise
cudaError_t set_fetch_chunk(int v);
cudaError_t set_no_park(int v);
cudaError_t set_tail_share(int v, int min_rays_per_8_threads);
cudaError_t set_shade_partition(int v);
cudaError_t launch_trace_batch(const SceneDev &S, const float4 *org, const float4 *dir, size_t n, int any_hit,
                               akb_hit *hits, int32_t *occ, uint32_t *fetch_cursor, int n_sms, cudaStream_t st);
cudaError_t launch_generate_rays(const SceneDev &S, const int32_t *xy, const float *u2, size_t n, float4 *org,
                                 float4 *dir, cudaStream_t st);
cudaError_t launch_film_to_srgb8(const float4 *rgbw, size_t n, uint8_t *out, cudaStream_t st);
cudaError_t launch_megakernel(const SceneDev &S, const RenderConfig &rc, float4 *film, Counters *ctr, cudaStream_t st);

cudaError_t launch_ao(const SceneDev &S, const RenderConfig &rc, float occlude, float4 *film, Counters *ctr, cudaStream_t st);

cudaError_t launch_wf_init(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int n_sms, cudaStream_t st);
cudaError_t launch_wf_trace(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int parity, int n_sms,
                            cudaStream_t st);
cudaError_t launch_wf_shade(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int parity, int n_sms,
                            cudaStream_t st);
// tail of a frame: one thread per entry of q_ext[parity] runs its pixel's remaining samples to completion
cudaError_t launch_wf_finish(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int parity, uint32_t n_upper,
                             cudaStream_t st);
// Streaming wavefront: one persistent launch renders the whole share (after launch_wf_init filled q_ext[0]).
constexpr int kStreamMaxSlots = 1 << 23;
cudaError_t set_stream_options(int flex_mask, int watchdog_ms);
cudaError_t launch_wf_stream(const SceneDev &S, const RenderConfig &rc, const Workspace &ws, int n_sms, cudaStream_t st);
cudaError_t launch_wf_resolve(const RenderConfig &rc, const Workspace &ws, float4 *film, cudaStream_t st);

// Sample-split combine over peer memory (akb_set_option("combine", "p2p")): dst[i] = ((dst[i] + film[0][i]) + film[1][i]) ...
// in rank order, the peers' films read over NVLink by the GPU that owns dst.
struct PeerFilms {
    const float4 *film[15];
    int n;
};
cudaError_t launch_film_sum_peers(float4 *dst, const PeerFilms &peers, size_t n_px, int n_sms, cudaStream_t st);

} // namespace akb
