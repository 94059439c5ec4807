"""ctypes view of include/akari_b200.h — the C ABI of libakari_b200.so.

Struct layouts mirror the header field by field. `load()` fails loudly when the CUDA library has not been
built: there is no CPU fallback on the product path.
"""
import ctypes as C
import os

ABI_VERSION = 4

OK, INVALID_ARG, CUDA_ERROR, NCCL_ERROR, OOM, NO_DEVICE = range(6)
MAT_DIFFUSE, MAT_GLOSSY, MAT_EMISSIVE, MAT_MIX, MAT_DIELECTRIC = range(5)
INTEGRATOR_PATH, INTEGRATOR_AO = 0, 1
SAMPLER_LCG, SAMPLER_PCG = 0, 1
MODE_WAVEFRONT, MODE_MEGAKERNEL = 0, 1
SPLIT_TILE, SPLIT_SAMPLE = 0, 1
BVH_DEFAULT, BVH_HOST_SAH, BVH_DEVICE = 0, 1, 2


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("color", C.c_float * 3), ("roughness", C.c_float), ("fraction", C.c_float),
                ("child_a", C.c_int32), ("child_b", C.c_int32), ("double_sided", C.c_int32), ("ior", C.c_float),
                ("color_tex", C.c_int32), ("roughness_tex", C.c_int32), ("fraction_tex", C.c_int32)]


class Texture(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgba", C.POINTER(C.c_float))]


class Mesh(C.Structure):
    _fields_ = [("n_vertices", C.c_uint64), ("n_triangles", C.c_uint64),
                ("vertices", C.POINTER(C.c_float)), ("normals", C.POINTER(C.c_float)),
                ("texcoords", C.POINTER(C.c_float)), ("indices", C.POINTER(C.c_int32)),
                ("material_indices", C.POINTER(C.c_int32)), ("materials", C.POINTER(C.c_int32)),
                ("n_mesh_materials", C.c_int32), ("_pad", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("c2w", C.c_float * 16), ("fov", C.c_float), ("width", C.c_int32), ("height", C.c_int32),
                ("lens_radius", C.c_float), ("focal_distance", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("n_meshes", C.c_uint32), ("meshes", C.POINTER(Mesh)),
                ("n_materials", C.c_uint32), ("_pad", C.c_uint32), ("materials", C.POINTER(Material)),
                ("camera", Camera), ("n_textures", C.c_uint32), ("bvh_build", C.c_uint32), ("textures", C.POINTER(Texture))]


class RenderParams(C.Structure):
    _fields_ = [("integrator", C.c_int32), ("spp", C.c_int32), ("max_depth", C.c_int32), ("ray_clamp", C.c_float),
                ("sampler", C.c_int32), ("rr_start_depth", C.c_int32), ("mode", C.c_int32),
                ("sort_materials", C.c_int32), ("ao_occlude", C.c_float), ("rank", C.c_int32),
                ("world_size", C.c_int32), ("tile_size", C.c_int32), ("split", C.c_int32), ("n_gpus", C.c_int32),
                ("_reserved", C.c_int32 * 3)]


class RenderStats(C.Structure):
    _fields_ = [("n_samples", C.c_uint64), ("n_closest_rays", C.c_uint64), ("n_shadow_rays", C.c_uint64),
                ("n_kernel_launches", C.c_uint64), ("ms_total", C.c_float), ("ms_raygen", C.c_float),
                ("ms_extend", C.c_float), ("ms_shade", C.c_float), ("ms_shadow", C.c_float), ("ms_film", C.c_float),
                ("ms_sort", C.c_float), ("n_gpus_used", C.c_int32), ("ms_combine", C.c_float), ("ms_wall", C.c_float),
                ("_reserved", C.c_float)]


class Hit(C.Structure):
    _fields_ = [("t", C.c_float), ("u", C.c_float), ("v", C.c_float), ("prim_id", C.c_int32), ("geom_id", C.c_int32)]


LIB_NAME = "libakari_b200.so"
# every symbol include/akari_b200.h declares (tests check the built library exports all of them)
SYMBOLS = ["akb_init", "akb_init_device", "akb_n_devices", "akb_set_option", "akb_shutdown", "akb_last_error", "akb_device_count", "akb_scene_create", "akb_scene_destroy",
           "akb_scene_info", "akb_render", "akb_render_device", "akb_trace", "akb_trace_device", "akb_generate_rays",
           "akb_film_to_srgb8", "akb_film_to_srgb8_noalloc", "akb_scene_file_load", "akb_scene_file_desc", "akb_scene_file_params",
           "akb_scene_file_output", "akb_scene_file_destroy", "akb_write_png", "akb_write_pfm"]


def lib_path():
    # AKB_LIB: development override to A/B differently compiled builds of the same library
    return os.environ.get("AKB_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), LIB_NAME)


_lib = None


def load():
    """dlopen libakari_b200.so and declare prototypes. Raises if it was not built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise RuntimeError(f"{p} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    L = C.CDLL(p)
    vp, fp, ip = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32)
    L.akb_init.argtypes = [C.c_int]
    L.akb_init.restype = C.c_int
    L.akb_init_device.argtypes = [C.c_int]
    L.akb_init_device.restype = C.c_int
    L.akb_n_devices.argtypes = []
    L.akb_n_devices.restype = C.c_int
    L.akb_set_option.argtypes = [C.c_char_p, C.c_char_p]
    L.akb_set_option.restype = C.c_int
    L.akb_shutdown.argtypes = []
    L.akb_shutdown.restype = C.c_int
    L.akb_last_error.argtypes = []
    L.akb_last_error.restype = C.c_char_p
    L.akb_device_count.argtypes = []
    L.akb_device_count.restype = C.c_int
    L.akb_scene_create.argtypes = [C.POINTER(SceneDesc), C.POINTER(vp)]
    L.akb_scene_create.restype = C.c_int
    L.akb_scene_destroy.argtypes = [vp]
    L.akb_scene_destroy.restype = C.c_int
    L.akb_scene_info.argtypes = [vp, C.c_int]
    L.akb_scene_info.restype = C.c_uint64
    L.akb_render.argtypes = [vp, C.POINTER(RenderParams), vp, C.POINTER(RenderStats)]
    L.akb_render.restype = C.c_int
    L.akb_render_device.argtypes = [vp, C.POINTER(RenderParams), vp, vp, C.POINTER(RenderStats)]
    L.akb_render_device.restype = C.c_int
    L.akb_trace.argtypes = [vp, vp, vp, C.c_size_t, C.c_int, vp, vp]
    L.akb_trace.restype = C.c_int
    L.akb_trace_device.argtypes = [vp, vp, vp, C.c_size_t, C.c_int, vp, vp, vp, fp]
    L.akb_trace_device.restype = C.c_int
    L.akb_generate_rays.argtypes = [vp, vp, vp, C.c_size_t, vp, vp]
    L.akb_generate_rays.restype = C.c_int
    L.akb_film_to_srgb8.argtypes = [vp, C.c_int32, C.c_int32, vp]
    L.akb_film_to_srgb8.restype = C.c_int
    L.akb_film_to_srgb8_noalloc.argtypes = [vp, C.c_int32, C.c_int32, vp]
    L.akb_film_to_srgb8_noalloc.restype = C.c_int
    L.akb_scene_file_load.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.akb_scene_file_load.restype = C.c_int
    L.akb_scene_file_desc.argtypes = [vp]
    L.akb_scene_file_desc.restype = C.POINTER(SceneDesc)
    L.akb_scene_file_params.argtypes = [vp, C.POINTER(RenderParams)]
    L.akb_scene_file_params.restype = C.c_int
    L.akb_scene_file_output.argtypes = [vp]
    L.akb_scene_file_output.restype = C.c_char_p
    L.akb_scene_file_destroy.argtypes = [vp]
    L.akb_scene_file_destroy.restype = C.c_int
    L.akb_write_png.argtypes = [C.c_char_p, vp, C.c_int32, C.c_int32]
    L.akb_write_png.restype = C.c_int
    L.akb_write_pfm.argtypes = [C.c_char_p, vp, C.c_int32, C.c_int32]
    L.akb_write_pfm.restype = C.c_int
    _lib = L
    return L
