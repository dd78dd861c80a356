"""ctypes binding of oracle/liboracle.so — the CPU restatement used as the checker.

Test infrastructure: imported only from tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs.  The product package never imports this module.
"""
import ctypes as C
import os

import numpy as np

from rustraytrace_b200 import _ffi as F

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = os.path.join(_ROOT, "oracle", "liboracle.so")

COUNTER_FIELDS = ["aabb", "sph_test", "sph_hit", "quad_test", "quad_full", "inst", "medium", "segments", "paths",
                  "lambertian", "metal", "dielectric", "isotropic", "emitted", "tex_checker", "tex_image",
                  "tex_noise", "pdf_mix", "rng_draws"]


class Counters(C.Structure):
    _fields_ = [(k, C.c_double) for k in COUNTER_FIELDS]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            raise ImportError(f"{_LIB} missing: run `make oracle`")
        L = C.CDLL(_LIB)
        VP, DP = C.c_void_p, C.POINTER(C.c_double)
        L.orc_scene_create.restype = VP
        L.orc_scene_create.argtypes = [C.POINTER(F.SceneDesc)]
        L.orc_scene_destroy.argtypes = [VP]
        L.orc_last_error.restype = C.c_char_p
        L.orc_image_height.argtypes = [C.POINTER(F.Camera)]
        L.orc_camera_internals.argtypes = [C.POINTER(F.Camera), C.c_int, VP]
        L.orc_probe.argtypes = [VP, C.POINTER(F.Camera), VP, VP]
        L.orc_render.argtypes = [VP, C.POINTER(F.Camera), C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int,
                                 VP, C.POINTER(Counters)]
        L.orc_render_moments.argtypes = [VP, C.POINTER(F.Camera), C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int,
                                         VP, VP, C.POINTER(Counters)]
        L.orc_write_color.argtypes = [VP, C.c_int, C.c_int, VP]
        L.orc_resolve_accum.argtypes = [VP, C.c_int, C.c_uint32, VP]
        L.orc_kat_hit.argtypes = [VP, DP, DP, C.c_double, C.c_double, C.c_double, DP]
        L.orc_kat_reflectance.restype = C.c_double
        L.orc_kat_reflectance.argtypes = [C.c_double, C.c_double]
        L.orc_kat_texture.argtypes = [VP, C.c_int, C.c_double, C.c_double, DP, DP]
        L.orc_kat_light_pdf.restype = C.c_double
        L.orc_kat_light_pdf.argtypes = [VP, DP, DP]
        L.orc_kat_refract.argtypes = [DP, DP, C.c_double, DP]
        L.orc_kat_world_bbox.argtypes = [VP, DP]
        _lib = L
    return _lib


def _d3(v):
    return (C.c_double * 3)(*[float(x) for x in v])


class OracleScene:
    def __init__(self, desc):
        self._h = lib().orc_scene_create(C.byref(desc))
        if not self._h:
            raise RuntimeError(lib().orc_last_error().decode())

    def close(self):
        if self._h:
            lib().orc_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def probe(self, cam):
        H = lib().orc_image_height(C.byref(cam))
        ids = np.empty((H, cam.image_width), dtype=np.int32)
        t = np.empty((H, cam.image_width), dtype=np.float64)
        lib().orc_probe(self._h, C.byref(cam), ids.ctypes.data, t.ctypes.data)
        return ids, t

    def render(self, cam, integrator, seed=1, nthreads=None, rows=None):
        """Returns (rgb (H,W,3) f64 pixel colours as write_color receives them, counters dict).
        rows = (begin, end, step) renders a subset; other rows stay NaN."""
        H = lib().orc_image_height(C.byref(cam))
        rgb = np.full((H, cam.image_width, 3), np.nan, dtype=np.float64)
        cnt = Counters()
        b, e, s = rows if rows else (0, H, 1)
        if nthreads is None:
            nthreads = os.cpu_count() or 1
        st = lib().orc_render(self._h, C.byref(cam), integrator, seed, nthreads, b, e, s, rgb.ctypes.data, C.byref(cnt))
        if st != 0:
            raise RuntimeError(lib().orc_last_error().decode())
        return rgb, {k: getattr(cnt, k) for k in COUNTER_FIELDS}

    def render_moments(self, cam, integrator, seed=1, nthreads=None):
        """(mean (H,W,3), mean of squares (H,W,3)) of the per-sample radiance, f64."""
        H = lib().orc_image_height(C.byref(cam))
        rgb = np.full((H, cam.image_width, 3), np.nan, dtype=np.float64)
        sq = np.full((H, cam.image_width, 3), np.nan, dtype=np.float64)
        if nthreads is None:
            nthreads = os.cpu_count() or 1
        st = lib().orc_render_moments(self._h, C.byref(cam), integrator, seed, nthreads, 0, H, 1, rgb.ctypes.data,
                                      sq.ctypes.data, None)
        if st != 0:
            raise RuntimeError(lib().orc_last_error().decode())
        return rgb, sq

    def hit(self, origin, direction, time=0.0, tmin=0.001, tmax=float("inf")):
        out = (C.c_double * 12)()
        if not lib().orc_kat_hit(self._h, _d3(origin), _d3(direction), time, tmin, tmax, out):
            return None
        o = list(out)
        return dict(t=o[0], p=o[1:4], n=o[4:7], u=o[7], v=o[8], front=bool(o[9]), id=int(o[10]), mat=int(o[11]))

    def texture(self, tex, u, v, p):
        out = (C.c_double * 3)()
        lib().orc_kat_texture(self._h, tex, u, v, _d3(p), out)
        return list(out)

    def light_pdf(self, origin, direction):
        return lib().orc_kat_light_pdf(self._h, _d3(origin), _d3(direction))

    def world_bbox(self):
        out = (C.c_double * 6)()
        lib().orc_kat_world_bbox(self._h, out)
        return list(out)


def camera_internals(cam, integrator=F.INT_BOOK2):
    out = (C.c_double * 22)()
    lib().orc_camera_internals(C.byref(cam), integrator, out)
    o = list(out)
    return dict(center=o[0:3], pixel00_loc=o[3:6], pixel_delta_u=o[6:9], pixel_delta_v=o[9:12],
                defocus_disk_u=o[12:15], defocus_disk_v=o[15:18], image_height=int(o[18]), sqrt_spp=int(o[19]),
                pixel_samples_scale=o[20], recip_sqrt_spp=o[21])


def write_color(rgb, nan_scrub):
    a = np.ascontiguousarray(rgb, dtype=np.float64).reshape(-1, 3)
    out = np.empty((a.shape[0], 3), dtype=np.uint8)
    lib().orc_write_color(a.ctypes.data, a.shape[0], int(nan_scrub), out.ctypes.data)
    return out.reshape(np.asarray(rgb).shape)


def resolve_accum(accum, spp):
    a = np.ascontiguousarray(accum, dtype=np.float32).reshape(-1, 4)
    out = np.empty((a.shape[0], 3), dtype=np.uint8)
    lib().orc_resolve_accum(a.ctypes.data, a.shape[0], int(spp), out.ctypes.data)
    return out.reshape(np.asarray(accum).shape[:-1] + (3,))


def reflectance(cosine, ri):
    return lib().orc_kat_reflectance(cosine, ri)


def refract(uv, n, eta):
    out = (C.c_double * 3)()
    lib().orc_kat_refract(_d3(uv), _d3(n), eta, out)
    return list(out)


# unit costs of SURVEY.md §8(d) ("algorithmic flops per path")
def flops_per_path(c):
    f = (27 * c["aabb"] + 30 * c["sph_test"] + 35 * c["sph_hit"] + 16 * c["quad_test"] + 47 * c["quad_full"]
         + 30 * c["inst"] + 30 * c["medium"] + 40 * c["lambertian"] + 60 * c["metal"] + 80 * c["dielectric"]
         + 40 * c["isotropic"] + 10 * c["tex_checker"] + 12 * c["tex_image"] + 1000 * c["tex_noise"]
         + 200 * c["pdf_mix"])
    return f / max(1.0, c["paths"])
