"""ctypes binding of the C ABI declared in include/yafgpu.h (libyafgpu.so).

This is plumbing only: every struct below mirrors the header field for field.  The library is loaded eagerly and the
import FAILS LOUDLY when the CUDA extension has not been built — there is no CPU fallback and no eager/PyTorch path.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("YAFGPU_LIB") or os.path.join(_HERE, "libyafgpu.so")  # YAFGPU_LIB: kernel-variant experiments (tools/)

ABI_VERSION = 5

# enums (include/yafgpu.h)
OK, ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA, ERR_STATE, ERR_NOMEM = range(6)
VIS_NORMAL, VIS_NO_SHADOWS, VIS_SHADOW_ONLY, VIS_INVISIBLE = range(4)
(MAT_SHINYDIFFUSE, MAT_GLOSSY, MAT_COATED_GLOSSY, MAT_GLASS, MAT_ROUGH_GLASS, MAT_BLEND, MAT_MIRROR, MAT_NULL,
 MAT_LIGHT) = range(9)
MAT_MASK = 11
LIGHT_POINT, LIGHT_SUN, LIGHT_AREA, LIGHT_MESH, LIGHT_BACKGROUND, LIGHT_DIRECTIONAL, LIGHT_SPOT, LIGHT_SPHERE, LIGHT_BGPORTAL, LIGHT_IES = range(10)
CAM_PERSPECTIVE, CAM_ARCHITECT, CAM_ORTHOGRAPHIC, CAM_ANGULAR, CAM_EQUIRECTANGULAR = range(5)
ANGULAR_PROJECTIONS = {"equidistant": 0, "orthographic": 1, "stereographic": 2, "equisolid_angle": 3, "rectilinear": 4}
BK = {"disk1": 0, "disk2": 1, "triangle": 3, "square": 4, "pentagon": 5, "hexagon": 6, "ring": 7}
BB = {"uniform": 0, "center": 1, "edge": 2}
BG_NONE, BG_CONSTANT, BG_TEXTURE, BG_GRADIENT = range(4)
FILTERS = {"box": 0, "mitchell": 1, "gauss": 2, "lanczos": 3}
INT_DIRECTLIGHT, INT_PATHTRACING = range(2)
TRACE_CLOSEST, TRACE_SHADOW, TRACE_SHADOW_TRANSP = range(3)
TRACE_STATIC_MAX_RAYS = 65536  # YAFGPU_TRACE_STATIC_MAX_RAYS
FLAG_COUNT_TRAVERSAL, FLAG_TIME_STAGES, FLAG_SHARE_DEVICE = 1, 2, 4
TEX_IMAGE, TEX_BLEND, TEX_RGB_CUBE, TEX_CLOUDS, TEX_MARBLE, TEX_WOOD, TEX_VORONOI, TEX_MUSGRAVE, TEX_DISTORTED_NOISE = range(9)
DIST_METRIC = {"squared": 1, "manhattan": 2, "chebychev": 3, "minkovsky_half": 4, "minkovsky_four": 5, "minkovsky": 6}
MUSGRAVE = {"multifractal": 1, "heteroterrain": 2, "hybridmf": 3, "ridgedmf": 4}
NOISE = {"stdperlin": 1, "blender": 2, "voronoi_f1": 3, "voronoi_f2": 4, "voronoi_f3": 5, "voronoi_f4": 6, "voronoi_f2f1": 7, "voronoi_crackle": 8, "cellnoise": 9}
NOISE_TABLES_BYTES = 13340
TEXBLEND = {"lin": 0, "quad": 1, "ease": 2, "diag": 3, "sphere": 4, "halo": 5, "quad_sphere": 5, "radial": 6}
TEXINTERP = {"none": 0, "bilinear": 1, "bicubic": 2}
TEXCLIP = {"extend": 0, "clip": 1, "clipcube": 2, "repeat": 3, "checker": 4}
CS = {"Raw_Manual_Gamma": 1, "LinearRGB": 2, "sRGB": 3, "XYZ": 4}
NODE_TEXTURE_MAPPER, NODE_VALUE, NODE_MIX, NODE_LAYER = range(4)
TEXCO = {"uv": 0, "global": 1, "orco": 2, "transformed": 3, "normal": 4, "reflect": 5, "window": 6, "stick": 7, "stress": 8, "tangent": 9}
TEXPROJ = {"plain": 0, "cube": 1, "tube": 2, "sphere": 3}
SLOTS = {"diffuse_shader": 0, "mirror_color_shader": 1, "bump_shader": 2, "mirror_shader": 3, "transparency_shader": 4,
         "translucency_shader": 5, "sigma_oren_shader": 6, "diffuse_refl_shader": 7, "IOR_shader": 8, "wireframe_shader": 9,
         "glossy_shader": 10, "glossy_reflect_shader": 11, "exponent_shader": 12, "filter_color_shader": 13,
         "roughness_shader": 14, "mask": 15}
N_SLOTS = 16

f32, i32, u32, u64 = C.c_float, C.c_int32, C.c_uint32, C.c_uint64
F3 = f32 * 3
pf, pi = C.POINTER(f32), C.POINTER(i32)


class Object(C.Structure):
    _fields_ = [("has_orco", i32), ("has_uv", i32), ("smooth", i32), ("light", i32), ("first_tri", i32),
                ("n_tris", i32), ("obj_id", i32), ("pass_index", i32)]


class Geometry(C.Structure):
    _fields_ = [("points", pf), ("n_points", C.c_size_t), ("normals", pf), ("n_normals", C.c_size_t),
                ("uv", pf), ("n_uv", C.c_size_t), ("tri_vidx", pi), ("tri_nidx", pi), ("tri_uvidx", pi),
                ("tri_mat", pi), ("tri_obj", pi), ("n_tris", C.c_size_t),
                ("objects", C.POINTER(Object)), ("n_objects", C.c_size_t)]


class MatShinyDiffuse(C.Structure):
    _fields_ = [("color", F3), ("mirror_color", F3), ("diffuse_reflect", f32), ("specular_reflect", f32),
                ("transparency", f32), ("translucency", f32), ("emit", f32), ("ior", f32), ("fresnel_effect", i32),
                ("transmit_filter", f32), ("oren_nayar", i32), ("sigma", f32)]


class MatGlossy(C.Structure):
    _fields_ = [("color", F3), ("diffuse_color", F3), ("mirror_color", F3), ("diffuse_reflect", f32),
                ("glossy_reflect", f32), ("specular_reflect", f32), ("exponent", f32), ("exp_u", f32), ("exp_v", f32),
                ("ior", f32), ("as_diffuse", i32), ("anisotropic", i32), ("oren_nayar", i32), ("sigma", f32)]


class MatGlass(C.Structure):
    _fields_ = [("ior", f32), ("filter_color", F3), ("transmit_filter", f32), ("mirror_color", F3),
                ("dispersion_power", f32), ("fake_shadows", i32), ("alpha", f32), ("has_absorption", i32),
                ("absorption", F3), ("absorption_dist", f32)]


class MatBlend(C.Structure):
    _fields_ = [("material1", i32), ("material2", i32), ("blend_value", f32)]


class MatSimple(C.Structure):
    _fields_ = [("color", F3), ("strength", f32), ("double_sided", i32)]


class MatUnion(C.Union):
    _fields_ = [("shinydiffuse", MatShinyDiffuse), ("glossy", MatGlossy), ("glass", MatGlass), ("blend", MatBlend),
                ("simple", MatSimple), ("pad_", f32 * 32)]


class Material(C.Structure):
    _fields_ = [("type", i32), ("visibility", i32), ("receive_shadows", i32), ("flat_material", i32),
                ("additional_depth", i32), ("transparent_bias_factor", f32),
                ("transparent_bias_multiply_raydepth", i32), ("sampling_factor", f32), ("mat_pass_index", i32),
                ("node_first", i32), ("node_count", i32), ("reserved", i32), ("shader", i32 * N_SLOTS), ("u", MatUnion)]


class Texture(C.Structure):
    _fields_ = [("type", i32), ("width", i32), ("height", i32), ("rgba", pf), ("interpolate", i32), ("clip_mode", i32),
                ("xrepeat", i32), ("yrepeat", i32), ("mirror_x", i32), ("mirror_y", i32), ("rot90", i32),
                ("cropmin_x", f32), ("cropmin_y", f32), ("cropmax_x", f32), ("cropmax_y", f32),
                ("checker_even", i32), ("checker_odd", i32), ("checker_dist", f32), ("normalmap", i32),
                ("color_space", i32), ("gamma", f32), ("adj_intensity", f32), ("adj_contrast", f32),
                ("adj_saturation", f32), ("adj_hue", f32), ("adj_mult_red", f32), ("adj_mult_green", f32),
                ("adj_mult_blue", f32), ("adj_clamp", i32), ("blend_type", i32), ("use_flip_axis", i32),
                ("noise_type", i32), ("noise_depth", i32), ("noise_size", f32), ("noise_hard", i32), ("color1", f32 * 3), ("color2", f32 * 3),
                ("clouds_bias", i32), ("turbulence", f32), ("sharpness", f32), ("wave_shape", i32), ("wood_rings", i32),
                ("vor_weight", f32 * 4), ("mk_exponent", f32), ("intensity_scale", f32), ("distance_metric", i32), ("color_type", i32),
                ("musgrave_type", i32), ("mus_h", f32), ("mus_lacunarity", f32), ("mus_octaves", f32), ("mus_offset", f32), ("mus_gain", f32),
                ("distort", f32), ("noise_type2", i32),
                ("ramp_items", i32), ("ramp_mode", i32), ("ramp_interpolation", i32), ("ramp_hue_interpolation", i32), ("ramp", C.POINTER(C.c_float))]


class ShaderNode(C.Structure):
    _fields_ = [("type", i32), ("input1", i32), ("input2", i32), ("factor", i32), ("mode", i32), ("texture", i32),
                ("texco", i32), ("mapping", i32), ("proj_x", i32), ("proj_y", i32), ("proj_z", i32), ("scale", F3),
                ("offset", F3), ("bump_strength", f32), ("do_scalar", i32), ("transform", f32 * 16),
                ("color", f32 * 4), ("color2", f32 * 4), ("scalar", f32), ("scalar2", f32), ("colfac", f32),
                ("valfac", f32), ("do_color", i32), ("layer_do_scalar", i32), ("color_input", i32), ("use_alpha", i32),
                ("no_rgb", i32), ("stencil", i32), ("negative", i32)]


class Light(C.Structure):
    _fields_ = [("type", i32), ("light_enabled", i32), ("cast_shadows", i32), ("photon_only", i32), ("samples", i32),
                ("object", i32), ("double_sided", i32), ("abs_intersect", i32), ("color", F3), ("power", f32),
                ("from_", F3), ("point1", F3), ("point2", F3), ("angle", f32), ("clamp_intersect", f32),
                ("shoot_caustic", i32), ("shoot_diffuse", i32), ("radius", f32), ("cone_angle", f32), ("blend", f32),
                ("shadow_fuzzy", f32), ("soft_shadows", i32), ("infinite", i32), ("ies_data", C.c_char_p), ("ies_size", C.c_size_t)]


class Camera(C.Structure):
    _fields_ = [("type", i32), ("resx", i32), ("resy", i32), ("from_", F3), ("to", F3), ("up", F3), ("focal", f32),
                ("aperture", f32), ("dof_distance", f32), ("aspect_ratio", f32), ("near_clip", f32), ("far_clip", f32),
                ("bokeh_type", i32), ("bokeh_bias", i32), ("bokeh_rotation", f32), ("scale", f32), ("angle", C.c_double),
                ("max_angle", C.c_double), ("circular", i32), ("mirrored", i32), ("projection", i32)]


class Background(C.Structure):
    _fields_ = [("type", i32), ("color", F3), ("power", f32), ("ibl", i32), ("with_caustic", i32),
                ("mapping_angular", i32), ("rotation", f32), ("tex_rgb", pf), ("tex_w", i32), ("tex_h", i32),
                ("tex_interp", i32), ("horizon_color", F3), ("zenith_color", F3), ("horizon_ground_color", F3),
                ("zenith_ground_color", F3)]


class FilmDesc(C.Structure):
    _fields_ = [("width", i32), ("height", i32), ("xstart", i32), ("ystart", i32), ("aa_pixelwidth", f32),
                ("filter_type", i32), ("tile_size", i32), ("premult", i32), ("aa_clamp_samples", f32),
                ("aa_clamp_indirect", f32), ("aa_passes", i32), ("aa_light_sample_multiplier", f32),
                ("aa_indirect_sample_multiplier", f32), ("base_sampling_offset", u32)]


class IntegratorDesc(C.Structure):
    _fields_ = [("type", i32), ("raydepth", i32), ("transp_shad", i32), ("shadow_depth", i32), ("path_samples", i32),
                ("bounces", i32), ("rr_min_bounces", i32), ("no_recursive", i32), ("trace_caustics", i32),
                ("bg_transp", i32), ("bg_transp_refract", i32), ("do_ao", i32), ("ao_samples", i32),
                ("ao_distance", f32), ("ao_color", F3), ("shadow_bias", f32), ("shadow_bias_auto", i32),
                ("ray_min_dist", f32), ("ray_min_dist_auto", i32)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", u32), ("geometry", Geometry), ("materials", C.POINTER(Material)),
                ("n_materials", C.c_size_t), ("lights", C.POINTER(Light)), ("n_lights", C.c_size_t),
                ("textures", C.POINTER(Texture)), ("n_textures", C.c_size_t),
                ("nodes", C.POINTER(ShaderNode)), ("n_nodes", C.c_size_t),
                ("camera", Camera), ("background", Background), ("film", FilmDesc), ("integrator", IntegratorDesc),
                ("cie_xyz", C.POINTER(C.c_float)), ("n_cie_rows", C.c_size_t), ("noise_tables", C.c_void_p), ("noise_tables_bytes", C.c_size_t)]


class PassDesc(C.Structure):
    _fields_ = [("n_samples", i32), ("offset", i32), ("adaptive", i32), ("aa_pass_number", i32),
                ("resample_mask", C.POINTER(C.c_uint8)), ("sample_begin", i32), ("sample_end", i32),
                ("part_index", i32), ("part_count", i32)]


class AaNoiseParams(C.Structure):
    _fields_ = [("threshold", f32), ("dark_detection_type", i32), ("dark_threshold_factor", f32), ("variance_edge_size", i32), ("variance_pixels", i32),
                ("detect_color_noise", i32)]


DARK_DETECTION = {"linear": 1, "curve": 2}


class Stats(C.Structure):
    _fields_ = [("rays_camera", u64), ("rays_extension", u64), ("rays_shadow", u64), ("rays_shadow_transp", u64),
                ("rays_light_probe", u64), ("camera_samples", u64), ("kernel_launches", u64), ("bvh_nodes", u64),
                ("bvh_bytes", u64), ("tri_bytes", u64), ("nodes_visited", u64), ("tris_tested", u64),
                ("ms_build", C.c_double), ("ms_trace", C.c_double), ("ms_shade", C.c_double), ("ms_film", C.c_double),
                ("ms_total", C.c_double), ("waves", u64), ("wave_retries", u64), ("ms_shadow", C.c_double),
                ("launches_trace", u64), ("launches_shadow", u64), ("launches_shade", u64), ("rays_in_timed_trace", u64),
                ("rays_in_timed_shadow", u64), ("last_wave_tasks", C.c_uint32 * 16), ("last_wave_shadow", C.c_uint32 * 16),
                ("bvh_node_bytes", C.c_uint32), ("bvh_width", C.c_uint32), ("bvh_levels", C.c_uint32), ("reserved0", C.c_uint32),
                ("nodes_visited_shadow", u64), ("tris_tested_shadow", u64), ("film_reduces", u64), ("ms_film_reduce", C.c_double), ("nonfinite_samples", u64)]


# every entry point include/yafgpu.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "yafgpu_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_uint]),
    "yafgpu_destroy": (None, [C.c_void_p]),
    "yafgpu_last_error": (C.c_char_p, [C.c_void_p]),
    "yafgpu_abi_version": (C.c_int, []),
    "yafgpu_upload_scene": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc)]),
    "yafgpu_scene_bound": (C.c_int, [C.c_void_p, pf]),
    "yafgpu_trace": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int]),
    "yafgpu_trace_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int]),
    "yafgpu_trace_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]),
    "yafgpu_film_clear": (C.c_int, [C.c_void_p]),
    "yafgpu_render_pass": (C.c_int, [C.c_void_p, C.POINTER(PassDesc)]),
    "yafgpu_sync": (C.c_int, [C.c_void_p]),
    "yafgpu_next_pass": (C.c_int, [C.c_void_p, C.POINTER(AaNoiseParams), C.POINTER(C.c_int64)]),
    "yafgpu_host_register": (C.c_int, [C.c_void_p, C.c_size_t]),
    "yafgpu_host_unregister": (C.c_int, [C.c_void_p]),
    "yafgpu_read_film": (C.c_int, [C.c_void_p, C.c_void_p]),
    "yafgpu_film_device_ptr": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "yafgpu_write_film": (C.c_int, [C.c_void_p, C.c_void_p]),
    "yafgpu_set_aa_multipliers": (C.c_int, [C.c_void_p, f32, f32]),
    "yafgpu_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_longlong]),
    "yafgpu_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "yafgpu_reset_stats": (C.c_int, [C.c_void_p]),
    "yafgpu_stream": (C.c_void_p, [C.c_void_p]),
    "yafgpu_group_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int, C.c_uint]),
    "yafgpu_group_destroy": (None, [C.c_void_p]),
    "yafgpu_group_last_error": (C.c_char_p, [C.c_void_p]),
    "yafgpu_group_size": (C.c_int, [C.c_void_p]),
    "yafgpu_group_ctx": (C.c_void_p, [C.c_void_p, C.c_int]),
    "yafgpu_group_upload_scene": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc)]),
    "yafgpu_group_film_clear": (C.c_int, [C.c_void_p]),
    "yafgpu_group_write_film": (C.c_int, [C.c_void_p, C.c_void_p]),
    "yafgpu_group_render_pass": (C.c_int, [C.c_void_p, C.POINTER(PassDesc)]),
    "yafgpu_group_read_film": (C.c_int, [C.c_void_p, C.c_void_p]),
    "yafgpu_group_sync": (C.c_int, [C.c_void_p]),
    "yafgpu_group_next_pass": (C.c_int, [C.c_void_p, C.POINTER(AaNoiseParams), C.POINTER(C.c_int64)]),
    "yafgpu_group_set_aa_multipliers": (C.c_int, [C.c_void_p, f32, f32]),
    "yafgpu_group_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_longlong]),
    "yafgpu_group_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
}


class YafGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("yafgpu error %d: %s" % (code, msg))
        self.code = code


def load():
    """dlopen libyafgpu.so and bind every declared symbol; raises when the extension is missing (no fallback)."""
    if not os.path.exists(LIB_PATH):
        raise ImportError("libyafgpu.so is not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                          "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError when a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.yafgpu_abi_version() != ABI_VERSION:
        raise ImportError("libyafgpu.so ABI version mismatch")
    return lib


class _Lib:
    """`capi.lib`: binds libyafgpu.so on first use.  The package import (libyafaray_b200/__init__.py) forces the load right away, so
    a missing library still fails loudly at import; only a process that declares itself host-only (LIBYAFARAY_B200_HOST_ONLY=1:
    scene description and XML export for the reference arm of bench.py, which must not map the product library) defers it — and
    even there the first device call loads the library or raises, there is no fallback."""
    _lib = None

    def load(self):
        if _Lib._lib is None:
            _Lib._lib = load()
        return _Lib._lib

    def __getattr__(self, name):
        return getattr(self.load(), name)


lib = _Lib()
