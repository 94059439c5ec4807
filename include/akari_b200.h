/*
 * akari_b200.h — C ABI of the B200-native path-tracing backend (libakari_b200.so).
 *
 * This is the drop-in boundary for ONE hot path of AkariRender: BVH ray/scene intersection +
 * the unidirectional `Path` integrator bounce loop + Film accumulation.
 * Every entry point below names the reference interface it replaces (paths relative to
 * /root/reference/src/akari/). Plain pointers and sizes only; no C++/torch types cross the ABI;
 * no exceptions and no abort(): every call returns an akb_status and leaves a message
 * retrievable with akb_last_error() (reference behaviour being replaced: CUDA_CHECK ->
 * fatal()+abort(), kernel/cuda/util.h:28-36).
 *
 * There is NO CPU fallback behind this ABI. Without a CUDA device akb_init() returns
 * AKB_NO_DEVICE and every compute call fails loudly.
 */
#ifndef AKARI_B200_H
#define AKARI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AKB_ABI_VERSION 4

typedef enum akb_status {
    AKB_OK = 0,
    AKB_INVALID_ARG = 1,
    AKB_CUDA_ERROR = 2,
    AKB_NCCL_ERROR = 3,
    AKB_OOM = 4,
    AKB_NO_DEVICE = 5
} akb_status;

/* ---- flattened Scene<C> (kernel/scene.h:50-91, kernel/instance.h:30-35) -------------------- */

/* Material variant tags, kernel/material.h:244-246 (Diffuse, Glossy, Emissive, Mix);
 * DIELECTRIC is an extension the reference does not have. */
typedef enum akb_material_type {
    AKB_MAT_DIFFUSE = 0,
    AKB_MAT_GLOSSY = 1,
    AKB_MAT_EMISSIVE = 2,
    AKB_MAT_MIX = 3,
    AKB_MAT_DIELECTRIC = 4
} akb_material_type;

/* An image texture (kernel/texture.h:39-57): RGBA float texels, row-major, top row first, values as the reference's
 * image reader stores them (core/image.cpp:105-123: byte / 255.0f, no transfer function; alpha unused).
 * Lookup: tc = fmod(texcoords, 1); tc.y = 1 - tc.y; texel (int)(tc * resolution), clamped (core/image.hpp:46-68). */
typedef struct akb_texture {
    int32_t width, height;
    const float *rgba; /* 4 * width * height */
} akb_texture;

/* One entry of the scene-wide material table. Every texture slot is a constant (kernel/texture.h:30-37) unless its
 * *_tex index (into akb_scene_desc.textures) is >= 0, in which case the image texture replaces the constant. */
typedef struct akb_material {
    int32_t type;         /* akb_material_type */
    float color[3];       /* Diffuse/Glossy reflectance, Emissive radiance, Dielectric tint */
    float roughness;      /* Glossy: texture value r; alpha = r*r (kernel/material.h:225-226) */
    float fraction;       /* Mix: u < fraction selects child_b (kernel/material.h:255-268) */
    int32_t child_a;      /* Mix: `first`  -> material_A, index into the scene material table */
    int32_t child_b;      /* Mix: `second` -> material_B */
    int32_t double_sided; /* Emissive (core/nodes/material.cpp never sets it -> 0) */
    float ior;            /* Dielectric (extension) */
    int32_t color_tex;     /* image texture for `color`, or -1 */
    int32_t roughness_tex; /* Glossy: image texture for `roughness` (channel 0), or -1 */
    int32_t fraction_tex;  /* Mix: image texture for `fraction` (channel 0), or -1 */
} akb_material;

/* One MeshInstance (kernel/instance.h:30-35); array layout = the .mesh file's (core/mesh.cpp:48-85). */
typedef struct akb_mesh {
    uint64_t n_vertices;
    uint64_t n_triangles;
    const float *vertices;           /* 3*n_vertices */
    const float *normals;            /* 9*n_triangles, per corner */
    const float *texcoords;          /* 6*n_triangles, per corner */
    const int32_t *indices;          /* 3*n_triangles */
    const int32_t *material_indices; /* n_triangles, index into `materials` below; -1 = none */
    const int32_t *materials;        /* n_mesh_materials indices into akb_scene_desc.materials */
    int32_t n_mesh_materials;
    int32_t _pad;
} akb_mesh;

/* PerspectiveCamera(resolution, c2w, fov) (kernel/camera.h:36-66). c2w is row-major 4x4, built by the
 * caller as core/nodes/camera.cpp:32-39 does; r2c is derived inside the library exactly as
 * kernel/camera.h:45-59 (including s = atan(fov/2)). */
typedef struct akb_camera {
    float c2w[16];
    float fov; /* radians */
    int32_t width;
    int32_t height;
    float lens_radius;    /* no reference node sets these; must be 0 */
    float focal_distance;
} akb_camera;

/* Who builds the 8-wide BVH at akb_scene_create (the reference builds its binary SBVH on the host inside
 * scene.commit(), kernel/bvh-accelerator.h:72-87,671-678). Any choice gives the same hits and films (traverse.h's
 * result contract does not depend on the tree) except where the reference's triangle test itself is ill-conditioned
 * (a grazing ray with |a| barely above the 1e-6 reject can "hit" a triangle at a computed point outside the
 * triangle's box: whether any tree - the reference's own included - reaches that test depends on its boxes; measured
 * about 2.5e-8 of rays on C3). The choices differ in build time and traversal speed.
 *   DEFAULT: AKB_BVH_BUILD=host|device in the environment, else HOST_SAH below 4 Mi triangles and DEVICE from
 *            there on (where the host build takes seconds).
 *   HOST_SAH: binned-SAH binary build on the host threads + optimal 8-wide collapse (csrc/bvh8.cpp).
 *   DEVICE: Morton order + locally-ordered clustering (PLOC; AKB_BVH_DEVICE_ALGO=lbvh for the plain binary radix
 *           tree) + the same collapse, all on the GPU (csrc/bvh_build.cu); falls back to HOST_SAH when the tree
 *           would be deeper than the traversal stack. */
typedef enum akb_bvh_build { AKB_BVH_DEFAULT = 0, AKB_BVH_HOST_SAH = 1, AKB_BVH_DEVICE = 2 } akb_bvh_build;

typedef struct akb_scene_desc {
    uint32_t abi_version; /* AKB_ABI_VERSION */
    uint32_t n_meshes;
    const akb_mesh *meshes;
    uint32_t n_materials;
    uint32_t _pad;
    const akb_material *materials;
    akb_camera camera;
    uint32_t n_textures;
    uint32_t bvh_build; /* akb_bvh_build; 0 in descriptions written before the field had a meaning */
    const akb_texture *textures;
} akb_scene_desc;

/* ---- integrator parameters (kernel/integrators/gpu/integrator.h:39-53, core/nodes/integrator.cpp:26-84) */

typedef enum akb_integrator { AKB_INTEGRATOR_PATH = 0, AKB_INTEGRATOR_AO = 1 } akb_integrator;
typedef enum akb_sampler { AKB_SAMPLER_LCG = 0, AKB_SAMPLER_PCG = 1 } akb_sampler; /* kernel/sampler.h:28-66 */
typedef enum akb_mode { AKB_MODE_WAVEFRONT = 0, AKB_MODE_MEGAKERNEL = 1 } akb_mode; /* Path{wavefront|megakernel} */
/* How `world_size` processes (one per GPU) share a frame. The reference shards tiles over CPU threads
 * (kernel/integrators/cpu/integrator.cpp:115-141) and has no multi-GPU path. */
typedef enum akb_split {
    AKB_SPLIT_TILE = 0,  /* interleaved tile_size^2 tiles, owner = tile_id % world_size; pixels this rank does not own
                            stay zero; the sum over ranks is bit-identical to the single-process film */
    AKB_SPLIT_SAMPLE = 1 /* every rank renders every pixel with spp/world_size samples (the first spp % world_size
                            ranks take one more) from per-rank sampler seeds x + y*W + rank*W*H; the sum over ranks
                            is a `spp`-sample estimate of the same image, not bit-identical to the single-process
                            one (the reference's per-pixel sampler stream is sequential and cannot be skipped ahead).
                            Rank 0's share is bit-identical to a single-process render at that sample count. */
} akb_split;

typedef struct akb_render_params {
    int32_t integrator;     /* akb_integrator */
    int32_t spp;            /* Path.spp / AO.spp */
    int32_t max_depth;      /* Path.max_depth (CPU oracle hard-codes 5, cpu/integrator.cpp:91) */
    float ray_clamp;        /* >0: per-sample clamp to [0, ray_clamp] + NaN->0 as gpu/cuda/integrator.cpp:196-198;
                               0: off (CPU-oracle semantics) */
    int32_t sampler;        /* akb_sampler; the reference hard-wires LCG (core/nodes/scene.cpp:126) */
    int32_t rr_start_depth; /* 0 = off (reference); >0: Russian roulette from this depth (extension) */
    int32_t mode;           /* akb_mode */
    int32_t sort_materials; /* wavefront: 0 = shade in queue order (fastest measured); 1 = shading queues keyed by BSDF kind;
                               2 = two queues, hits | misses. All give the same film bits. */
    float ao_occlude;       /* AO.occlude distance */
    /* image-space sharding: interleaved tile_size^2 tiles, owner = tile_id % world_size.
       Pixels this rank does not own are left at zero in the film. */
    int32_t rank;
    int32_t world_size;
    int32_t tile_size;
    int32_t split;          /* akb_split */
    /* One process driving several GPUs (akb_init(n_devices >= 2)): akb_render splits the frame over the first n_gpus
       initialised devices (`split`, `tile_size` as above; one host thread + stream per GPU inside the library), combines
       the partial films with ONE collective on device 0 and returns the whole film. 0 or 1: one GPU. rank/world_size
       must then be 0/0 or 0/1 (they describe the other arrangement: one process per GPU). */
    int32_t n_gpus;
    int32_t _reserved[3];
} akb_render_params;

/* Per-render statistics (the reference prints per-kernel cudaEvent stats, kernel/cuda/launch.cpp:92-118). */
typedef struct akb_render_stats {
    uint64_t n_samples;       /* camera paths started */
    uint64_t n_closest_rays;  /* Scene::intersect queries */
    uint64_t n_shadow_rays;   /* Scene::occlude queries */
    uint64_t n_kernel_launches;
    float ms_total;           /* device time of the whole render (CUDA events) */
    float ms_raygen, ms_extend, ms_shade, ms_shadow, ms_film, ms_sort; /* per-stage device time when profiled */
    /* multi-GPU akb_render: counts above are sums over the GPUs, ms_total the slowest GPU's device time */
    int32_t n_gpus_used;
    float ms_combine;         /* host time of the film collective + its synchronisation */
    float ms_wall;            /* host time of the whole call (render + combine + device-to-host copy) */
    float _reserved;
} akb_render_stats;

/* Result of Scene::intersect (kernel/scene.h:40-49): prim_id/geom_id = -1 and t = +inf on a miss. */
typedef struct akb_hit {
    float t, u, v;
    int32_t prim_id;
    int32_t geom_id;
} akb_hit;

typedef struct akb_scene akb_scene; /* opaque */

/* ---- lifecycle --------------------------------------------------------------------------------------- */

/* Replaces Device selection (cmd/akari.cpp:41-76 `--gpu`, core/device.cpp:142-183): binds this process to the CUDA
 * devices 0 .. n_devices-1 (0 = every visible device). Scenes created afterwards are replicated on all of them;
 * akb_render_params.n_gpus chooses how many a frame uses. AKB_NO_DEVICE when fewer are usable. Calling it again
 * re-binds (existing scenes must have been destroyed). */
akb_status akb_init(int n_devices);
/* The other arrangement, one process per GPU (torchrun / MPI: LOCAL_RANK): binds this process to the single device
 * `device_index`; the caller passes rank/world_size in akb_render_params and combines the films itself. */
akb_status akb_init_device(int device_index);
akb_status akb_shutdown(void);
/* Number of devices the last akb_init / akb_init_device bound (0 before). */
int akb_n_devices(void);
/* Tuning switches (csrc/options.h lists them; defaults are the measured best, none changes a result bit). The
 * library never reads the environment. AKB_INVALID_ARG for an unknown name or a value out of range. Takes effect for
 * scenes created / frames rendered afterwards. */
akb_status akb_set_option(const char *name, const char *value);
const char *akb_last_error(void);
/* 0 when no CUDA device is usable. */
int akb_device_count(void);

/* Replaces SceneNode::compile + `scene.accel = BVHAccelerator; scene.commit()` (core/nodes/scene.cpp:43-95,
 * 119-135; kernel/scene.cpp:64-81; kernel/bvh-accelerator.h:671-678): deep-copies the description, derives the
 * area-light list and power distribution, builds the compressed 8-wide BVH and uploads everything to HBM. */
akb_status akb_scene_create(const akb_scene_desc *desc, akb_scene **out_scene);
akb_status akb_scene_destroy(akb_scene *scene);

/* Scene facts (for tests/benches): 0 n_triangles, 1 n_bvh_nodes, 2 bvh_bytes, 3 n_lights, 4 triangle_bytes,
 * 5 bvh_depth, 6 build_us (BVH build wall time), 7 the akb_bvh_build that produced the tree, 8 the number of GPUs
 * holding a replica, 9 upload_us (host-to-device copies of all replicas, wall time). */
uint64_t akb_scene_info(const akb_scene *scene, int what);

/* ---- the hot path ------------------------------------------------------------------------------------ */

/* Replaces gpu::Integrator<C>::render(const Scene<C>&, Film<C>*) (kernel/integrators/gpu/integrator.h:54-59,
 * called from SceneNode::render core/nodes/scene.cpp:141-162). Blocking. `out_rgbw` is a HOST buffer of
 * width*height*4 floats = (sum R, sum G, sum B, sum weight) per pixel, row-major: the contents of
 * Film::radiance / Film::weight after merge_tile (core/film.h:66-95). `stats` may be NULL.
 * params->n_gpus >= 2: the frame is split over that many GPUs and combined on device 0 with one ncclReduce
 * (AKB_NCCL_ERROR if NCCL cannot be loaded or fails; akb_set_option("combine", "p2p") uses the library's own
 * peer-memory kernels over NVLink instead). With AKB_SPLIT_TILE the film is bit-identical to the one-GPU film.
 * A page-locked `out_rgbw` (cudaHostAlloc / cudaHostRegister) receives the film by one DMA; pageable memory is filled
 * through the library's own pinned staging buffer. */
akb_status akb_render(akb_scene *scene, const akb_render_params *params, float *out_rgbw, akb_render_stats *stats);

/* Same, but the film stays in HBM: `d_out_rgbw` is a DEVICE pointer (width*height*4 floats) the caller owns
 * (e.g. a buffer it will hand to NCCL). Work is enqueued on `cuda_stream` (a cudaStream_t, 0 = default) and
 * the call returns after the stream has been synchronised. */
akb_status akb_render_device(akb_scene *scene, const akb_render_params *params, float *d_out_rgbw,
                             void *cuda_stream, akb_render_stats *stats);

/* Replaces Scene::intersect / Scene::occlude (kernel/scene.h:58-67) over a batch of rays.
 * Rays are two float4 streams (vectorised SoA): org_tmin[i] = (ox,oy,oz,tmin), dir_tmax[i] = (dx,dy,dz,tmax).
 * any_hit = 0: closest hit -> out_hits[i]; any_hit = 1: out_occluded[i] = 0/1. HOST pointers. Rays go to the device in
 * chunks (option trace_chunk) on two streams, so both PCIe directions overlap the traversal; page-locked buffers are
 * copied from / into directly, pageable ones through the library's pinned staging slots. */
akb_status akb_trace(akb_scene *scene, const float *org_tmin, const float *dir_tmax, size_t n_rays, int any_hit,
                     akb_hit *out_hits, int32_t *out_occluded);
/* DEVICE-pointer variant of akb_trace; `ms_kernel` (may be NULL) receives the traversal kernel's device time. */
akb_status akb_trace_device(akb_scene *scene, const float *d_org_tmin, const float *d_dir_tmax, size_t n_rays,
                            int any_hit, akb_hit *d_out_hits, int32_t *d_out_occluded, void *cuda_stream,
                            float *ms_kernel);

/* Replaces Camera::generate_ray (kernel/camera.h:67-86) for a batch: pixel (x,y) with lens sample u1 (unused,
 * pinhole) and film sample u2. HOST pointers; writes the two float4 ray streams. Test/KAT helper. */
akb_status akb_generate_rays(akb_scene *scene, const int32_t *xy, const float *u2, size_t n, float *org_tmin,
                             float *dir_tmax);

/* Film post-process of Film::write_image (core/film.h:97-113, core/image.cpp:38-61): radiance/weight ->
 * sRGB OETF -> round(255.5*x) clamp -> RGB8. HOST pointers; runs on the device. */
akb_status akb_film_to_srgb8(const float *rgbw, int32_t width, int32_t height, uint8_t *out_rgb8);

/* akb_film_to_srgb8 for callers that may not allocate device memory (code under CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc). This entry point has NO counterpart in the reference.
 * Same arguments (HOST pointers) and the same bytes out as akb_film_to_srgb8, but
 * the device staging is one static device array of AKB_NOALLOC_MAX_PIXELS pixels (19 bytes per pixel = 38 MiB, held
 * by every GPU that loads the library until the process exits) instead of two cudaMalloc / cudaFree per call: the
 * call allocates and frees nothing. width * height > AKB_NOALLOC_MAX_PIXELS -> AKB_INVALID_ARG. All calls share the
 * array (a process has one copy per GPU): calls of one process that overlap on the GPU race; serialise them. */
#define AKB_NOALLOC_MAX_PIXELS ((size_t)1 << 21) /* 2 097 152: a 1920x1080 film fits */
akb_status akb_film_to_srgb8_noalloc(const float *rgbw, int32_t width, int32_t height, uint8_t *out_rgb8);

/* ---- scene files (front-end) ------------------------------------------------------------------------ */

/* Replaces SceneGraphParser<C>::parse_file + the scene-graph nodes' object_field/compile for the Path hot path
 * (core/parser.cpp:150-363, core/nodes/{scene,camera,mesh,material,integrator}.cpp, core/mesh.cpp:48-85;
 * cmd/akari.cpp:78-92 looks up the export `scene`): reads a `.akari` file with its imports and `.mesh` files and
 * flattens it. Host-only (no device needed). The returned description and everything it points to belong to the
 * handle and stay valid until akb_scene_file_destroy; hand it to akb_scene_create, which deep-copies. */
typedef struct akb_scene_file akb_scene_file; /* opaque */
akb_status akb_scene_file_load(const char *path, akb_scene_file **out_file);
const akb_scene_desc *akb_scene_file_desc(const akb_scene_file *file);
/* Integrator node -> parameters (PathIntegratorNode / AOIntegratorNode defaults, core/nodes/integrator.cpp:26-84;
 * rank 0 of 1). */
akb_status akb_scene_file_params(const akb_scene_file *file, akb_render_params *out_params);
/* Scene.output (core/nodes/scene.cpp:100-101), default "out.png". */
const char *akb_scene_file_output(const akb_scene_file *file);
akb_status akb_scene_file_destroy(akb_scene_file *file);

/* Replaces the image write at the end of Film::write_image (core/film.h:97-113 -> core/image.cpp:38-61): 8-bit RGB
 * rows, top to bottom, as PNG. Host-only. */
akb_status akb_write_png(const char *path, const uint8_t *rgb8, int32_t width, int32_t height);
/* Float output the reference lacks: radiance/weight per pixel as a little-endian RGB .pfm. Host-only. */
akb_status akb_write_pfm(const char *path, const float *rgbw, int32_t width, int32_t height);

#ifdef __cplusplus
}
#endif
#endif /* AKARI_B200_H */
