"""ctypes declarations for include/rrt.h and include/rrt_host.h.

The structures mirror the C headers field for field (which in turn mirror the reference's
Camera / RenderOverrides / HittableObject records, see the headers for file:line).  The two
libraries are loaded from this package directory; a missing library is a hard error — there is
no Python or CPU implementation of the render path to fall back to.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)
IMAGE_DIR = os.path.join(REPO_ROOT, "images")

c_int = C.c_int32
c_dbl = C.c_double
D3 = c_dbl * 3


class Hittable(C.Structure):
    _fields_ = [("kind", c_int), ("material", c_int), ("first_child", c_int), ("n_children", c_int),
                ("a", D3), ("b", D3), ("c", D3)]


class Material(C.Structure):
    _fields_ = [("kind", c_int), ("texture", c_int), ("albedo", D3), ("fuzz", c_dbl),
                ("refraction_index", c_dbl)]


class Texture(C.Structure):
    _fields_ = [("kind", c_int), ("even", c_int), ("odd", c_int), ("image", c_int), ("perlin", c_int),
                ("_pad", c_int), ("color", D3), ("scale", c_dbl)]


class Image(C.Structure):
    _fields_ = [("width", c_int), ("height", c_int), ("rgb", C.POINTER(C.c_uint8))]


class Perlin(C.Structure):
    _fields_ = [("randvec", (c_dbl * 3) * 256), ("perm_x", c_int * 256), ("perm_y", c_int * 256),
                ("perm_z", c_int * 256)]


class SceneDesc(C.Structure):
    _fields_ = [("hittables", C.POINTER(Hittable)), ("n_hittables", c_int),
                ("children", C.POINTER(c_int)), ("n_children", c_int),
                ("materials", C.POINTER(Material)), ("n_materials", c_int),
                ("textures", C.POINTER(Texture)), ("n_textures", c_int),
                ("images", C.POINTER(Image)), ("n_images", c_int),
                ("perlins", C.POINTER(Perlin)), ("n_perlins", c_int),
                ("world", c_int), ("lights", c_int)]


class Camera(C.Structure):
    _fields_ = [("aspect_ratio", c_dbl), ("image_width", c_int), ("samples_per_pixel", c_int),
                ("max_depth", c_int), ("_pad", c_int), ("background", D3), ("vfov", c_dbl),
                ("lookfrom", D3), ("lookat", D3), ("vup", D3), ("defocus_angle", c_dbl),
                ("focus_dist", c_dbl)]


class RenderParams(C.Structure):
    _fields_ = [("integrator", c_int), ("shard_mode", c_int), ("rank", c_int), ("world", c_int),
                ("seed", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("width", c_int), ("height", c_int), ("spp_effective", c_int), ("kernel_launches", c_int),
                ("paths", c_dbl), ("kernel_ms", c_dbl), ("total_ms", c_dbl),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("n_prims", c_int), ("n_bvh_nodes", c_int), ("n_media", c_int), ("n_lights", c_int),
                ("sm_count", c_int), ("kernel_features", c_int), ("n_devices", c_int), ("n_grid_cells", c_int),
                ("exchange_ms", c_dbl)]


class Overrides(C.Structure):
    _fields_ = [("has", C.c_uint32), ("image_width", c_int), ("samples_per_pixel", c_int), ("max_depth", c_int),
                ("aspect_ratio", c_dbl), ("vfov", c_dbl), ("lookfrom", D3), ("lookat", D3), ("vup", D3),
                ("defocus_angle", c_dbl), ("focus_dist", c_dbl), ("background", D3)]


# enum values (include/rrt.h, include/rrt_host.h)
H_SPHERE, H_QUAD, H_LIST, H_BVH, H_TRANSLATE, H_ROTATE_Y, H_CONSTANT_MEDIUM = range(7)
M_LAMBERTIAN, M_METAL, M_DIELECTRIC, M_DIFFUSE_LIGHT, M_ISOTROPIC, M_EMPTY = range(6)
T_SOLID, T_CHECKER, T_IMAGE, T_NOISE = range(4)
INT_BOOK1, INT_BOOK2, INT_BOOK3 = 1, 2, 3
SHARD_NONE, SHARD_TILE, SHARD_SAMPLE = 0, 1, 2
SF_MOTION, SF_QUADS, SF_MEDIA, SF_TEXTURES, SF_NOISE, SF_BOXES, SF_ALL, SF_GRID, SF_GRID_MOVING = 1, 2, 4, 8, 16, 32, 63, 64, 128  # rrt_stats.kernel_features bits
BOOK_1, BOOK_2, BOOK_3, BOOK_1_GPU = 1, 2, 3, 4
OK, ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_NO_DEVICE = 0, -1, -2, -3, -4
O_BITS = {"aspect_ratio": 1 << 0, "image_width": 1 << 1, "samples_per_pixel": 1 << 2, "max_depth": 1 << 3,
          "vfov": 1 << 4, "lookfrom": 1 << 5, "lookat": 1 << 6, "vup": 1 << 7, "defocus_angle": 1 << 8,
          "focus_dist": 1 << 9, "background": 1 << 10}

# every symbol the two headers declare: (name, restype, argtypes)
VP = C.c_void_p
PROBE_STATIC_MAX_PIXELS = 1048576  # RRT_PROBE_STATIC_MAX_PIXELS (include/rrt.h)
RRT_SYMBOLS = [
    ("rrt_init", c_int, [c_int, C.POINTER(VP)]),
    ("rrt_shutdown", None, [VP]),
    ("rrt_init_multi", c_int, [C.POINTER(c_int), c_int, C.POINTER(VP)]),
    ("rrt_comm_unique_id", c_int, [VP]),
    ("rrt_comm_init", c_int, [VP, VP, c_int, c_int]),
    ("rrt_render_collective", c_int, [VP, VP, C.POINTER(Camera), C.POINTER(RenderParams), VP, VP]),
    ("rrt_render_frame_collective", c_int, [VP, C.POINTER(SceneDesc), C.POINTER(Camera), C.POINTER(RenderParams), VP]),
    ("rrt_tile_scatter_host", c_int, [C.POINTER(Camera), c_int, C.POINTER(VP), VP]),
    ("rrt_set_stream", c_int, [VP, VP]),
    ("rrt_scene_create", c_int, [VP, C.POINTER(SceneDesc), C.POINTER(VP)]),
    ("rrt_scene_destroy", None, [VP]),
    ("rrt_image_height", c_int, [C.POINTER(Camera)]),
    ("rrt_render", c_int, [VP, VP, C.POINTER(Camera), C.POINTER(RenderParams), VP]),
    ("rrt_render_device", c_int, [VP, VP, C.POINTER(Camera), C.POINTER(RenderParams), VP]),
    ("rrt_render_frame", c_int, [VP, C.POINTER(SceneDesc), C.POINTER(Camera), C.POINTER(RenderParams), VP]),
    ("rrt_probe", c_int, [VP, VP, C.POINTER(Camera), c_int, VP, VP]),
    ("rrt_probe_static", c_int, [VP, VP, C.POINTER(Camera), c_int, VP, VP]),
    ("rrt_resolve_rgb8_device", c_int, [VP, VP, c_int, c_int, c_int, VP]),
    ("rrt_resolve_rgb8", c_int, [VP, VP, c_int, c_int, c_int, VP]),
    ("rrt_get_stats", c_int, [VP, C.POINTER(Stats)]),
    ("rrt_shard_plan", c_int, [C.POINTER(Camera), C.POINTER(RenderParams), C.POINTER(c_int), C.POINTER(c_int),
                               C.POINTER(c_int), C.POINTER(c_dbl), VP]),
    ("rrt_flatten_check", c_int, [C.POINTER(SceneDesc)] + [C.POINTER(c_int)] * 5),
    ("rrt_scene_features", c_int, [C.POINTER(SceneDesc), C.POINTER(c_int)]),
    ("rrt_flatten_rects", c_int, [C.POINTER(SceneDesc), C.POINTER(c_int * 3)]),
    ("rrt_flatten_grid", c_int, [C.POINTER(SceneDesc), c_int, C.POINTER(c_int * 3), C.POINTER(c_int), C.POINTER(c_int), C.POINTER(c_int)]),
    ("rrt_philox_kat", None, [C.c_uint32] * 6 + [C.POINTER(C.c_uint32)]),
    ("rrt_philox_kat_rounds", c_int, [c_int] + [C.c_uint32] * 6 + [C.POINTER(C.c_uint32)]),
    ("rrt_philox_rounds", c_int, []),
    ("rrt_kat_light", c_int, [VP, VP, c_int, VP, VP, VP, VP, VP, VP]),
    ("rrt_kat_rng", c_int, [VP, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, c_int, VP, VP]),
    ("rrt_last_error", C.c_char_p, []),
    ("rrt_version", C.c_char_p, []),
]
RRTH_SYMBOLS = [
    ("rrth_overrides_shipped", None, [C.POINTER(Overrides)]),
    ("rrth_overrides_none", None, [C.POINTER(Overrides)]),
    ("rrth_normalize_book", c_int, [C.c_char_p]),
    ("rrth_scene_build", c_int, [c_int, c_int, C.c_uint64, C.POINTER(Overrides), C.c_char_p, C.POINTER(VP)]),
    ("rrth_scene_desc", C.POINTER(SceneDesc), [VP]),
    ("rrth_scene_camera", C.POINTER(Camera), [VP]),
    ("rrth_scene_integrator", c_int, [VP]),
    ("rrth_scene_free", None, [VP]),
    ("rrth_write_ppm_from_accum", c_int, [C.c_char_p, c_int, c_int, VP, C.c_uint32]),
    ("rrth_write_ppm_rgb8", c_int, [C.c_char_p, c_int, c_int, VP]),
    ("rrth_load_image", c_int, [C.c_char_p, C.c_char_p, C.POINTER(c_int), C.POINTER(c_int), VP, C.c_uint64]),
    ("rrth_rng_kat_state1234", C.c_uint64, [c_int]),
    ("rrth_rng_kat_seeded_f64", c_dbl, [C.c_uint64, c_int]),
    ("rrth_last_error", C.c_char_p, []),
]


def _load(name, symbols):
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `make` (or __graft_entry__.build()). "
            "rustraytrace_b200 has no fallback implementation.")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    for sym, res, args in symbols:
        fn = getattr(lib, sym)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    return lib


_rrt = None
_rrth = None


def rrt():
    """librrt.so (CUDA kernels + C ABI)."""
    global _rrt
    if _rrt is None:
        # RRT_LIB: development hook to load an experimental build (scripts/build_variant.sh)
        _rrt = _load(os.environ.get("RRT_LIB", "librrt.so"), RRT_SYMBOLS)
    return _rrt


def rrth():
    """librrt_host.so (scene builders, overrides, PPM writer)."""
    global _rrth
    if _rrth is None:
        _rrth = _load("librrt_host.so", RRTH_SYMBOLS)
    return _rrth
