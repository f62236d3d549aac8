"""ctypes mirror of include/henjou_b200.h (the C ABI of libhenjou_b200.so). Plain structs only."""
import ctypes as C

HJ_OK = 0
INTEGRATOR_PT, INTEGRATOR_NEE, INTEGRATOR_MIS = 0, 1, 2
TONEMAP_NONE, TONEMAP_ACES, TONEMAP_UCHIMURA = 0, 1, 2
AOV_COLOR, AOV_ALBEDO, AOV_NORMAL, AOV_RGBA8, AOV_ACCUM, AOV_PATH_RADIANCE, AOV_PATH_IDS, AOV_OUTPUT, AOV_OUTPUT_RGBA8 = range(9)
POST_PASSTHROUGH, POST_DENOISE, POST_DENOISE_UPSCALE2X = 0, 1, 2
ACCEL_BUILD, ACCEL_REFIT = 0, 1

f32p = C.POINTER(C.c_float)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)


class hj_material(C.Structure):
    _fields_ = [
        ("basecolor", C.c_float * 3), ("basecolor_tex", C.c_int32),
        ("emission", C.c_float * 3), ("emission_tex", C.c_int32),
        ("is_light", C.c_int32), ("metallic", C.c_float), ("metallic_tex", C.c_int32),
        ("roughness", C.c_float), ("roughness_tex", C.c_int32),
        ("specular", C.c_float * 3), ("sheen", C.c_float), ("clearcoat", C.c_float),
        ("clearcoat_tex", C.c_int32), ("bump_tex", C.c_int32), ("normal_tex", C.c_int32),
        ("transmission", C.c_float), ("ior", C.c_float),
        ("ideal_specular", C.c_int32), ("is_thinfilm", C.c_int32),
    ]


class hj_texture(C.Structure):
    _fields_ = [("rgba8", u8p), ("width", C.c_int32), ("height", C.c_int32), ("srgb", C.c_int32)]


class hj_scene_desc(C.Structure):
    _fields_ = [
        ("vertices", f32p), ("normals", f32p), ("texcoords", f32p),
        ("indices", u32p), ("material_ids", u32p), ("prim_offsets", u32p),
        ("n_vertices", C.c_uint32), ("n_triangles", C.c_uint32), ("n_instances", C.c_uint32),
        ("materials", C.POINTER(hj_material)), ("n_materials", C.c_uint32),
        ("textures", C.POINTER(hj_texture)), ("n_textures", C.c_uint32),
        ("light_prim_ids", u32p), ("light_prim_emission", f32p), ("n_light_prims", C.c_uint32),
        ("ibl_rgba32f", f32p), ("ibl_width", C.c_int32), ("ibl_height", C.c_int32),
        ("lut_rgba8", u8p), ("lut_width", C.c_int32), ("lut_height", C.c_int32),
    ]


class hj_frame_params(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32), ("spp", C.c_uint32),
        ("sample_begin", C.c_uint32), ("sample_end", C.c_uint32), ("frame", C.c_uint32),
        ("camera_pos", C.c_float * 3), ("camera_dir", C.c_float * 3),
        ("camera_up", C.c_float * 3), ("camera_right", C.c_float * 3),
        ("camera_f", C.c_float), ("ibl_intensity", C.c_float),
        ("integrator", C.c_int32), ("max_depth", C.c_int32), ("tonemap", C.c_int32),
        ("pixel_jitter", C.c_int32), ("debug_paths", C.c_int32), ("sky_sampling", C.c_int32),
    ]


class hj_stats(C.Structure):
    _fields_ = [
        ("rays_extend", C.c_uint64), ("rays_shadow", C.c_uint64), ("paths", C.c_uint64),
        ("nodes_fetched", C.c_uint64), ("tris_fetched", C.c_uint64), ("kernel_launches", C.c_uint64),
        ("ms_render", C.c_float), ("ms_trace", C.c_float), ("ms_shade", C.c_float), ("ms_accel", C.c_float),
        ("bvh_nodes", C.c_uint32), ("bvh_node_bytes", C.c_uint32),
        ("bvh_depth", C.c_uint32), ("max_nodes_per_ray", C.c_uint32), ("max_tris_per_ray", C.c_uint32), ("ms_post", C.c_float), ("post_launches", C.c_uint32),
        ("passes", C.c_uint32),
    ]


# every symbol include/henjou_b200.h declares (tests check the built library exports all of them)
EXPORTS = [
    "hj_abi_version", "hj_create", "hj_destroy", "hj_last_error", "hj_set_option", "hj_scene_upload",
    "hj_accel_build", "hj_render", "hj_render_async", "hj_sync", "hj_readback", "hj_device_ptr",
    "hj_finalize", "hj_get_stats", "hj_comm_unique_id", "hj_comm_init", "hj_comm_reduce", "hj_trace",
    "hj_host_pin", "hj_host_unpin", "hj_post", "hj_libm_eval", "hj_accel_build_async", "hj_accel_wait", "hj_comm_scene_upload", "hj_trace_noalloc",
]

TRACE_NOALLOC_MAX_RAYS = 65536      # HJ_TRACE_NOALLOC_MAX_RAYS (include/henjou_b200.h)
