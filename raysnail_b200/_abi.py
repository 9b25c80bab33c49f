"""ctypes view of include/raysnail_b200.h and include/raysnail_b200_host.h.

Pure plumbing: struct layouts and function prototypes of the C ABI.  The library is loaded
from the package directory (built in-tree by `make` / `__graft_entry__.build()`); there is no
fallback implementation — if the shared object is missing, importing the symbols raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RAYSNAIL_B200_LIB selects an experimental build of the same library (kernel tuning variants)
LIB_PATH = os.environ.get("RAYSNAIL_B200_LIB") or os.path.join(_HERE, "libraysnail_b200.so")

RS_OK = 0
RS_FLAG_PIXEL_CENTRE = 1
RS_FLAG_KERNEL_TIMING = 2
RS_FLAG_MIS = 4
RS_REDO_REFERENCE, RS_REDO_NOISE, RS_REDO_NOISE_FIXED = 0, 1, 2

# every symbol include/raysnail_b200.h declares
CORE_SYMBOLS = [
    "rs_abi_version", "rs_abi_struct_size", "rs_set_option", "rs_create", "rs_destroy", "rs_last_error", "rs_upload_scene", "rs_render",
    "rs_render_accumulate", "rs_finalize", "rs_combine_passes", "rs_noise_map", "rs_quantize_rgb8",
    "rs_debug_primary", "rs_debug_scatter", "rs_get_stats",
    "rs_device_alloc", "rs_device_free", "rs_ipc_export", "rs_ipc_open", "rs_ipc_close", "rs_finalize_peers",
    "rs_render_passes", "rs_pass_seed", "rs_philox4x32_10", "rs_philox_uniform",
    "rs_create_multi", "rs_destroy_multi", "rs_multi_last_error", "rs_multi_device_count", "rs_multi_set_option",
    "rs_multi_upload_scene", "rs_multi_render", "rs_signal", "rs_wait_counter", "rs_wait_check", "rs_redo_map_static",
]
# every symbol include/raysnail_b200_host.h declares
HOST_SYMBOLS = [
    "rs_host_last_error", "rs_host_scene_from_sdl_file", "rs_host_scene_from_sdl_text",
    "rs_host_scene_builtin", "rs_host_scene_free", "rs_host_scene_desc", "rs_host_scene_camera",
    "rs_host_camera_build", "rs_host_scene_new", "rs_host_tex_solid", "rs_host_tex_checker",
    "rs_host_tex_image", "rs_host_tex_image_file", "rs_host_tex_perlin", "rs_host_mat_lambertian",
    "rs_host_mat_metal", "rs_host_mat_diffuse_metal", "rs_host_mat_dielectric",
    "rs_host_mat_diffuse_light", "rs_host_mat_isotropic", "rs_host_mat_blinn_phong",
    "rs_host_mat_mixed", "rs_host_mat_set_phong", "rs_host_obj_sphere", "rs_host_obj_box",
    "rs_host_obj_rect", "rs_host_obj_quadric", "rs_host_obj_raymarcher", "rs_host_obj_difference", "rs_host_obj_intersection", "rs_host_obj_bvh",
    "rs_host_obj_transform", "rs_host_obj_medium", "rs_host_world_add", "rs_host_world_add_light",
    "rs_host_world_add_mesh", "rs_host_world_add_obj_file", "rs_host_world_set_background",
    "rs_host_world_set_time_range", "rs_host_scene_set_camera", "rs_host_seed_random_u32",
    "rs_host_chacha12_block", "rs_host_write_png_rgb8",
]


class RsNode(C.Structure):
    _fields_ = [("type", C.c_int32), ("material", C.c_int32), ("child0", C.c_int32), ("child1", C.c_int32),
                ("first_xform", C.c_int32), ("n_xform", C.c_int32), ("rect_axis", C.c_int32), ("reserved", C.c_int32),
                ("p", C.c_double * 10)]


class RsXform(C.Structure):
    _fields_ = [("m", C.c_double * 16), ("inv", C.c_double * 16)]


class RsMaterial(C.Structure):
    _fields_ = [("type", C.c_int32), ("texture", C.c_int32), ("mat1", C.c_int32), ("mat2", C.c_int32),
                ("has_curve", C.c_int32), ("phong_exponent", C.c_int32), ("phong_factor", C.c_double),
                ("a", C.c_double), ("b", C.c_double), ("color", C.c_float * 4)]


class RsTexture(C.Structure):
    _fields_ = [("type", C.c_int32), ("odd", C.c_int32), ("even", C.c_int32), ("width", C.c_int32),
                ("height", C.c_int32), ("texel_offset", C.c_int32), ("perlin_kind", C.c_int32),
                ("perlin_depth", C.c_int32), ("perlin_smooth", C.c_int32), ("perlin_vector", C.c_int32),
                ("point_count", C.c_int32), ("perm_offset", C.c_int32), ("value_offset", C.c_int32),
                ("reserved", C.c_int32), ("scale", C.c_double), ("color", C.c_float * 4)]


class RsSceneDesc(C.Structure):
    _fields_ = [("nodes", C.POINTER(RsNode)), ("n_nodes", C.c_int32),
                ("roots", C.POINTER(C.c_int32)), ("root_bbox", C.POINTER(C.c_double)), ("n_roots", C.c_int32),
                ("lights", C.POINTER(C.c_int32)), ("n_lights", C.c_int32),
                ("xforms", C.POINTER(RsXform)), ("n_xforms", C.c_int32),
                ("tri_verts", C.POINTER(C.c_double)), ("tri_normals", C.POINTER(C.c_double)),
                ("tri_material", C.POINTER(C.c_int32)), ("n_tris", C.c_int32),
                ("materials", C.POINTER(RsMaterial)), ("n_materials", C.c_int32),
                ("textures", C.POINTER(RsTexture)), ("n_textures", C.c_int32),
                ("texels", C.POINTER(C.c_uint8)), ("n_texel_bytes", C.c_int64),
                ("perlin_perm", C.POINTER(C.c_int32)), ("n_perlin_perm", C.c_int32),
                ("perlin_values", C.POINTER(C.c_double)), ("n_perlin_values", C.c_int32),
                ("default_material", C.c_int32), ("bg0", C.c_float * 4), ("bg1", C.c_float * 4)]


class RsCamera(C.Structure):
    _fields_ = [("origin", C.c_double * 3), ("lb", C.c_double * 3), ("horizontal_full", C.c_double * 3),
                ("vertical_full", C.c_double * 3), ("horizontal_unit", C.c_double * 3),
                ("vertical_unit", C.c_double * 3), ("aperture", C.c_double), ("shutter_speed", C.c_double),
                ("width", C.c_int32), ("height", C.c_int32)]


class RsRenderParams(C.Structure):
    _fields_ = [("sqrt_spp", C.c_int32), ("depth", C.c_int32), ("gamma", C.c_int32),
                ("stratum_row_begin", C.c_int32), ("stratum_row_end", C.c_int32), ("max_wave_paths", C.c_int32),
                ("seed", C.c_uint64), ("flags", C.c_uint32), ("reserved", C.c_uint32)]


class RsStats(C.Structure):
    _fields_ = [("n_samples", C.c_uint64), ("n_segments", C.c_uint64), ("n_launches", C.c_uint64),
                ("ms_raygen", C.c_double), ("ms_extend", C.c_double), ("ms_shade", C.c_double),
                ("ms_accumulate", C.c_double), ("ms_total", C.c_double)]


_lib = None


def load() -> C.CDLL:
    """Load libraysnail_b200.so (once) and attach prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()) — there is no fallback path")
    lib = C.CDLL(LIB_PATH)
    vp, i32, f64, f32 = C.c_void_p, C.c_int32, C.c_double, C.c_float
    P = C.POINTER

    def proto(name, restype, *argtypes):
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = list(argtypes)

    proto("rs_abi_version", C.c_int)
    proto("rs_abi_struct_size", C.c_int, C.c_int)
    proto("rs_set_option", C.c_int, vp, C.c_int, C.c_int)
    proto("rs_create", C.c_int, P(vp), C.c_int)
    proto("rs_destroy", None, vp)
    proto("rs_last_error", C.c_char_p, vp)
    proto("rs_upload_scene", C.c_int, vp, P(RsSceneDesc))
    proto("rs_render", C.c_int, vp, P(RsCamera), P(RsRenderParams), vp, vp)
    proto("rs_render_accumulate", C.c_int, vp, P(RsCamera), P(RsRenderParams), vp, vp, vp)
    proto("rs_finalize", C.c_int, vp, vp, vp, vp, i32, i32, i32, i32, vp)
    proto("rs_combine_passes", C.c_int, vp, vp, vp, vp, C.c_int64, f32, vp)
    proto("rs_noise_map", C.c_int, vp, vp, vp, i32, i32, i32, vp)
    proto("rs_quantize_rgb8", C.c_int, vp, vp, vp, C.c_int64, vp)
    proto("rs_redo_map_static", C.c_int, vp, vp, vp, vp, i32, i32, i32, f32, vp)
    proto("rs_debug_primary", C.c_int, vp, P(RsCamera), P(RsRenderParams), i32, vp, vp, vp, vp, vp, vp, vp)
    proto("rs_debug_scatter", C.c_int, vp, P(RsCamera), P(RsRenderParams), i32, vp, vp, vp, vp)
    proto("rs_get_stats", C.c_int, vp, P(RsStats))
    proto("rs_device_alloc", C.c_int, vp, C.c_int64, P(vp))
    proto("rs_device_free", C.c_int, vp, vp)
    proto("rs_ipc_export", C.c_int, vp, vp, vp)
    proto("rs_ipc_open", C.c_int, vp, vp, P(vp))
    proto("rs_ipc_close", C.c_int, vp, vp)
    proto("rs_finalize_peers", C.c_int, vp, P(vp), i32, vp, vp, C.c_int64, C.c_int64, i32, i32, vp)
    proto("rs_render_passes", C.c_int, vp, P(RsCamera), P(RsRenderParams), i32, i32, f32, vp, vp, vp)
    proto("rs_pass_seed", C.c_uint64, C.c_uint64, i32)
    proto("rs_create_multi", C.c_int, P(vp), P(i32), i32)
    proto("rs_destroy_multi", None, vp)
    proto("rs_multi_last_error", C.c_char_p, vp)
    proto("rs_multi_device_count", i32, vp)
    proto("rs_multi_set_option", C.c_int, vp, C.c_int, C.c_int)
    proto("rs_multi_upload_scene", C.c_int, vp, P(RsSceneDesc))
    proto("rs_multi_render", C.c_int, vp, P(RsCamera), P(RsRenderParams), vp, vp, P(RsStats))
    proto("rs_signal", C.c_int, vp, vp, vp)
    proto("rs_wait_counter", C.c_int, vp, vp, C.c_uint32, vp)
    proto("rs_wait_check", C.c_int, vp, vp)
    proto("rs_philox4x32_10", C.c_int, vp, vp, vp)
    proto("rs_philox_uniform", C.c_int, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, vp)

    proto("rs_host_last_error", C.c_char_p)
    proto("rs_host_scene_from_sdl_file", C.c_int, C.c_char_p, P(vp))
    proto("rs_host_scene_from_sdl_text", C.c_int, C.c_char_p, P(vp))
    proto("rs_host_scene_builtin", C.c_int, C.c_char_p, C.c_uint64, i32, P(vp))
    proto("rs_host_scene_free", None, vp)
    proto("rs_host_scene_desc", P(RsSceneDesc), vp)
    proto("rs_host_scene_camera", C.c_int, vp, i32, i32, P(RsCamera))
    proto("rs_host_camera_build", C.c_int, P(f64), P(f64), P(f64), f64, f64, f64, f64, i32, i32, P(RsCamera))
    proto("rs_host_scene_new", vp)
    proto("rs_host_tex_solid", C.c_int, vp, f32, f32, f32, f32)
    proto("rs_host_tex_checker", C.c_int, vp, C.c_int, C.c_int, f64)
    proto("rs_host_tex_image", C.c_int, vp, i32, i32, vp)
    proto("rs_host_tex_image_file", C.c_int, vp, C.c_char_p)
    proto("rs_host_tex_perlin", C.c_int, vp, i32, i32, C.c_uint64, i32, i32, i32, f64)
    proto("rs_host_mat_lambertian", C.c_int, vp, C.c_int)
    proto("rs_host_mat_metal", C.c_int, vp, C.c_int)
    proto("rs_host_mat_diffuse_metal", C.c_int, vp, f64, C.c_int)
    proto("rs_host_mat_dielectric", C.c_int, vp, f32, f32, f32, f64, C.c_int)
    proto("rs_host_mat_diffuse_light", C.c_int, vp, C.c_int, f64)
    proto("rs_host_mat_isotropic", C.c_int, vp, f32, f32, f32)
    proto("rs_host_mat_blinn_phong", C.c_int, vp, f64, f64, C.c_int)
    proto("rs_host_mat_mixed", C.c_int, vp, C.c_int, C.c_int, f64)
    proto("rs_host_mat_set_phong", C.c_int, vp, C.c_int, f64, i32)
    proto("rs_host_obj_sphere", C.c_int, vp, P(f64), f64, P(f64), C.c_int)
    proto("rs_host_obj_box", C.c_int, vp, P(f64), P(f64), C.c_int)
    proto("rs_host_obj_rect", C.c_int, vp, i32, f64, f64, f64, f64, f64, C.c_int)
    proto("rs_host_obj_quadric", C.c_int, vp, P(f64), C.c_int)
    proto("rs_host_obj_raymarcher", C.c_int, vp, C.c_int)
    proto("rs_host_obj_difference", C.c_int, vp, C.c_int, C.c_int, C.c_int)
    proto("rs_host_obj_intersection", C.c_int, vp, C.c_int, C.c_int, C.c_int)
    proto("rs_host_obj_bvh", C.c_int, vp, C.c_int32, C.POINTER(C.c_int32))
    proto("rs_host_obj_transform", C.c_int, vp, C.c_int, i32, P(i32), P(f64))
    proto("rs_host_obj_medium", C.c_int, vp, C.c_int, f32, f32, f32, f64)
    proto("rs_host_world_add", C.c_int, vp, C.c_int)
    proto("rs_host_world_add_light", C.c_int, vp, C.c_int)
    proto("rs_host_world_add_mesh", C.c_int, vp, vp, C.c_int64, vp, C.c_int64, vp, C.c_int64, f64, P(f64), f64, i32, C.c_int)
    proto("rs_host_world_add_obj_file", C.c_int, vp, C.c_char_p, f64, P(f64), f64, i32, C.c_int)
    proto("rs_host_world_set_background", C.c_int, vp, P(f32), P(f32))
    proto("rs_host_world_set_time_range", C.c_int, vp, f64, f64)
    proto("rs_host_scene_set_camera", C.c_int, vp, P(f64), P(f64), P(f64), f64, f64, f64, f64)
    proto("rs_host_seed_random_u32", C.c_int, C.c_uint64, vp, i32)
    proto("rs_host_chacha12_block", C.c_int, vp, C.c_uint64, vp)
    proto("rs_host_write_png_rgb8", C.c_int, C.c_char_p, vp, i32, i32)
    _lib = lib
    return lib
