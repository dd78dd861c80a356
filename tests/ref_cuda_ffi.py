"""ctypes binding of oracle/_ref/libref_cuda.so — the REFERENCE'S OWN CUDA kernel
(src/cuda/mod.rs:15-335, extracted and compiled by oracle/ref_cuda/build.sh) driven with the
reference's launch shape.  Test/bench infrastructure: used as a second, real-reference witness for
converged book-1 renders and as the timed "existing src/cuda path" beside ours.  It supports what
the reference's GPU path supports: spheres with Lambertian/Metal/Dielectric, sky or solid background.
"""
import ctypes as C
import os

import numpy as np

from rustraytrace_b200 import _ffi as F

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(_ROOT, "oracle", "_ref", "libref_cuda.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.refcuda_render.argtypes = [C.c_void_p, C.c_void_p, C.c_uint, C.c_void_p, C.c_uint, C.c_uint, C.c_uint,
                                     C.c_uint, C.c_uint, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.refcuda_render.restype = C.c_int
        _lib = L
    return _lib


def pods_from_scene(scene, seed=0x1234567, centre_pixels=True):
    """(CameraUniform 144 B, SphereGpu[n] 32 B, MaterialGpu[m] 32 B) as build_in_one_weekend_scene
    fills them (src/gpu/mod.rs:174-197, 276-298) from a BookScene of spheres."""
    cam, d = scene.camera, scene.desc
    W, H = cam.image_width, F.rrt().rrt_image_height(C.byref(cam))
    lookfrom, lookat, vup = (np.array(list(x), dtype=np.float64) for x in (cam.lookfrom, cam.lookat, cam.vup))
    h = np.tan(np.radians(cam.vfov) / 2.0)
    vh = 2.0 * h * cam.focus_dist
    vw = vh * (W / H)
    w = (lookfrom - lookat) / np.linalg.norm(lookfrom - lookat)
    u = np.cross(vup, w); u /= np.linalg.norm(u)
    v = np.cross(w, u)
    vu, vv = u * vw, v * -vh
    du, dv = vu / W, vv / H
    p00 = lookfrom - w * cam.focus_dist - vu / 2 - vv / 2 + (du + dv) * 0.5
    if centre_pixels:
        # the reference GPU kernels jitter with x + U[0,1) around pixel00 (src/cuda/mod.rs:319-320),
        # its CPU path with x + U[-.5,.5) (in_one_weekend/camera.rs:152-166): half a pixel apart.  The
        # CPU path is the parity target, so hand the GPU kernel a pixel00 moved back by half a pixel.
        p00 = p00 - (du + dv) * 0.5
    defocus_radius = cam.focus_dist * np.tan(np.radians(cam.defocus_angle / 2.0))
    mats, sph = [], []
    mat_index = {}
    for i in range(d.n_hittables):
        hh = d.hittables[i]
        if hh.kind != F.H_SPHERE:
            continue
        if hh.material not in mat_index:
            m = d.materials[hh.material]
            if m.kind == F.M_LAMBERTIAN:
                col = list(d.textures[m.texture].color); rec = (col + [0.0], 0, 1.0)
            elif m.kind == F.M_METAL:
                rec = (list(m.albedo) + [m.fuzz], 1, 1.0)
            elif m.kind == F.M_DIELECTRIC:
                rec = ([1.0, 1.0, 1.0, 0.0], 2, m.refraction_index)
            else:
                raise ValueError("the reference CUDA kernel has no such material")
            mat_index[hh.material] = len(mats); mats.append(rec)
        sph.append((list(hh.a) + [hh.c[0]], mat_index[hh.material]))
    S = np.zeros(len(sph), dtype=[("cr", "<f4", 4), ("mat", "<u4"), ("pad", "<u4", 3)])
    for k, (cr, mi) in enumerate(sph):
        S[k]["cr"] = cr; S[k]["mat"] = mi
    M = np.zeros(len(mats), dtype=[("af", "<f4", 4), ("kind", "<u4"), ("ri", "<f4"), ("pad", "<u4", 2)])
    for k, (af, kind, ri) in enumerate(mats):
        M[k]["af"] = af; M[k]["kind"] = kind; M[k]["ri"] = ri
    cu = np.zeros(1, dtype=[("f", "<f4", 32), ("u", "<u4", 4)])
    f = cu[0]["f"]
    for row, vec in enumerate((lookfrom, p00, du, dv, u, v, np.array(list(cam.background)))):
        f[4 * row:4 * row + 3] = vec
    f[28:32] = [defocus_radius, W, H, cam.samples_per_pixel]
    cu[0]["u"] = [cam.max_depth, seed, len(sph), 1 if scene.integrator != F.INT_BOOK1 else 0]
    assert cu.nbytes == 144 and S.itemsize == 32 and M.itemsize == 32
    return cu, S, M, W, H


def render(scene, seed=0x1234567, centre_pixels=True):
    """Runs the reference kernel; returns (accum (H, W, 4) f32 RGBA sums, kernel_ms, passes)."""
    cu, S, M, W, H = pods_from_scene(scene, seed, centre_pixels)
    acc = np.zeros((H, W, 4), dtype=np.float32)
    ms, passes = C.c_float(), C.c_int()
    rc = lib().refcuda_render(cu.ctypes.data, S.ctypes.data, len(S), M.ctypes.data, len(M), seed,
                              scene.camera.samples_per_pixel, W, H, acc.ctypes.data, C.byref(ms), C.byref(passes))
    if rc != 0:
        raise RuntimeError(f"reference CUDA kernel failed (rc {rc})")
    return acc, ms.value, passes.value
