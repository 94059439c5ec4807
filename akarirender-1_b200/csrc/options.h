// options.h — process-wide tuning switches of the library, set through akb_set_option(name, value) (include/akari_b200.h).
// The library never reads the environment: a drop-in must not change behaviour on variables its host never heard of.
// Defaults are the measured best on B200 (profiles/r01c_notes.md, profiles/r02_notes.md); none changes a result bit.
#pragma once
#include <string>

namespace akb {

struct Options {
    // wavefront scheduling (api.cu)
    int pipelines = 2;                   // pipelines of the wavefront: the owned tiles in 1..4 groups, each with its own queues and stream
    int pipeline_interleave = 0;         // 1: tiles are dealt out to the pipelines in turn (all see the same mix of sky and ground; measured
                                         // slower: pipelines then move in step and might as well be one); 0: contiguous ranges of tiles
    int pipeline_min_slots = 128 * 1024; // below this many paths one pipeline keeps the launches large
    int graphs = 1;                      // 0: launch every bounce's kernels one by one
    int check_every = 4;                 // bounces per graph launch / queue-length poll (even)
    int poll_lag = 1;                    // 0: wait for every queue-length poll
    int finish_div = 4;                  // a pipeline whose queue is below slots / this is finished by wf_finish_kernel (0: never)
    int finish_max = 1 << 16;            // ... and below this many paths
    int no_park = 1;                     // 0: a path that ends with a shadow ray pending always waits one bounce for it
    int shade_partition = 0;             // 1: the shade step regroups a block's 128 entries into hits | misses (measured slower)
    int tail_share = 1;                  // once the ray queue is drained, a warp's idle lanes take subtrees off its busy lanes' stacks:
                                         // 0 never, 1 in launches with >= tail_share_min rays per 8 threads, 2 always
    int tail_share_min = 16;             // (16 = two rays per thread: C3 104.5 ms against 105.8 at 32 and 104.1 at 4; a 1/8 tile-split
                                         //  share 29.2 ms at 16 and 32 but 34.6 at 8 - launches with under two rays per thread are all tail)
    int refill_below = 0;                // persistent traversal warps refill below this many live lanes (0: 24 / 20 by step)
    int fetch_chunk = 0;                 // >0: warp-private fetch chunks
    int stream = 0;                      // 1: the streaming wavefront (one persistent launch per share, no barrier between bounces; kernels.cu)
    int stream_flex_mask = 1;            // warps of a block (bit w = warp w) that prefer shading
    int stream_watchdog_ms = 4000;       // a streaming warp idle for this long fails the frame instead of hanging the GPU
    int profile_stages = 0;              // per-stage events, one pipeline, no graphs, no finish kernel
    int timeline = 0;                    // print every poll's device time and queue lengths
    long long trace_chunk = 1ll << 21;   // rays per staging slot of akb_trace (host batches)
    int trav_ifif = -1;                  // -1: if-if traversal step for trees of >= 64 nodes, fused below; 0 / 1 force
    // BVH build (bvh8.cpp, bvh_build.cu)
    int bvh_build = 0;        // akb_bvh_build used when the description says DEFAULT (0: by size)
    int bvh_verbose = 0;      // build timings on stderr
    float bvh_trav_cost = 0.5f, bvh_prim_cost = 0.3f;
    int bvh_max_leaf = 3;     // a CWBVH leaf slot addresses at most 3 triangles
    int bvh_device_lbvh = 0;  // 1: plain binary radix tree instead of PLOC
    int ploc_radius = 3;
    // multi-GPU frame (api.cu)
    int combine = 0;          // 0: ncclReduce of the partial films; 1: peer-memory kernels over NVLink (no NCCL)
    int nvtx = 1;             // NVTX ranges around render / trace / scene build
};

Options &options();
// false (+ message) for an unknown name or an out-of-range value
bool set_option(const char *name, const char *value, std::string *err);

} // namespace akb
