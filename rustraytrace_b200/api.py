"""Python face of the two C libraries: thin, no arithmetic of its own.

    scene = BookScene("the_next_week", 1, overrides=dict(image_width=400))   # books::<book>::run up to cam.render
    ctx = Context(0)                                                          # CudaDevice::new(0)
    gs = ctx.scene(scene.desc)                                                # upload + BVH
    accum = ctx.render(gs, scene.camera, scene.integrator, seed=1)            # (H, W, 4) f32 sums
    rgb = ctx.resolve(accum, spp)                                             # render_io.rs arithmetic

Errors raise RrtError with the library's message (the reference returns Err(String),
src/cuda/mod.rs:342-441).  There is no CPU fallback anywhere in this module.
"""
import ctypes as C

import numpy as np

from . import _ffi as F


class RrtError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"rrt status {status}: {message}")
        self.status = status


def _check(st):
    if st != 0:
        raise RrtError(st, F.rrt().rrt_last_error().decode(errors="replace"))


def make_overrides(**kw):
    """RenderOverrides (src/config.rs:1-14): keyword = Some(value)."""
    o = F.Overrides()
    F.rrth().rrth_overrides_none(C.byref(o))
    for k, v in kw.items():
        if v is None:
            continue
        if k not in F.O_BITS:
            raise KeyError(f"unknown override {k!r}")
        o.has |= F.O_BITS[k]
        if k in ("lookfrom", "lookat", "vup", "background"):
            setattr(o, k, F.D3(*[float(x) for x in v]))
        elif k in ("image_width", "samples_per_pixel", "max_depth"):
            setattr(o, k, int(v))
        else:
            setattr(o, k, float(v))
    return o


def shipped_overrides():
    """const OVERRIDES as the reference ships it (src/config.rs:50-62)."""
    o = F.Overrides()
    F.rrth().rrth_overrides_shipped(C.byref(o))
    return o


def copy_camera(cam, **changes):
    c = F.Camera()
    C.memmove(C.byref(c), C.byref(cam), C.sizeof(F.Camera))
    for k, v in changes.items():
        if k in ("lookfrom", "lookat", "vup", "background"):
            setattr(c, k, F.D3(*[float(x) for x in v]))
        else:
            setattr(c, k, v)
    return c


class BookScene:
    """books::<book>::run(scene) up to the cam.render call (host mirror, librrt_host.so)."""

    def __init__(self, book, scene_no=None, seed=0x5EED1234, overrides=None, image_dir=F.IMAGE_DIR):
        h = F.rrth()
        self.book = h.rrth_normalize_book(book.encode()) if isinstance(book, str) else int(book)
        if self.book == 0:
            raise ValueError(f"unknown book {book!r} (books: in_one_weekend, the_next_week, the_rest_of_your_life)")
        if isinstance(overrides, dict):
            overrides = make_overrides(**overrides)
        self._handle = C.c_void_p()
        st = h.rrth_scene_build(self.book, -1 if scene_no is None else int(scene_no), seed,
                                C.byref(overrides) if overrides is not None else None,
                                image_dir.encode() if image_dir else None, C.byref(self._handle))
        if st != 0:
            raise RuntimeError(h.rrth_last_error().decode())
        self.desc = h.rrth_scene_desc(self._handle).contents
        self.camera = copy_camera(h.rrth_scene_camera(self._handle).contents)
        self.integrator = h.rrth_scene_integrator(self._handle)

    @property
    def height(self):
        return F.rrt().rrt_image_height(C.byref(self.camera))

    @property
    def width(self):
        return self.camera.image_width

    def close(self):
        if self._handle:
            F.rrth().rrth_scene_free(self._handle)
            self._handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DescBuilder:
    """Hand-built rrt_scene_desc for tests (plays the role of make_ref/make_mat/make_tex)."""

    def __init__(self):
        self.h, self.ch, self.m, self.t, self.im, self.pn = [], [], [], [], [], []
        self._keep = []

    def solid(self, c):
        self.t.append(F.Texture(F.T_SOLID, -1, -1, -1, -1, 0, F.D3(*c), 1.0))
        return len(self.t) - 1

    def checker(self, scale, even, odd):
        self.t.append(F.Texture(F.T_CHECKER, even, odd, -1, -1, 0, F.D3(0, 0, 0), scale))
        return len(self.t) - 1

    def image(self, rgb):
        """rgb: (H, W, 3) uint8 array or None (missing file)."""
        if rgb is None:
            self.im.append(F.Image(0, 0, None))
        else:
            a = np.ascontiguousarray(rgb, dtype=np.uint8)
            self._keep.append(a)
            self.im.append(F.Image(a.shape[1], a.shape[0], a.ctypes.data_as(C.POINTER(C.c_uint8))))
        self.t.append(F.Texture(F.T_IMAGE, -1, -1, len(self.im) - 1, -1, 0, F.D3(0, 0, 0), 1.0))
        return len(self.t) - 1

    def noise(self, scale, perlin):
        self.pn.append(perlin)
        self.t.append(F.Texture(F.T_NOISE, -1, -1, -1, len(self.pn) - 1, 0, F.D3(0, 0, 0), scale))
        return len(self.t) - 1

    def material(self, kind, tex=-1, albedo=(0, 0, 0), fuzz=0.0, ri=1.0):
        self.m.append(F.Material(kind, tex, F.D3(*albedo), fuzz, ri))
        return len(self.m) - 1

    def lambertian(self, c):
        return self.material(F.M_LAMBERTIAN, self.solid(c), c)

    def _hit(self, kind, mat, first, n, a=(0, 0, 0), b=(0, 0, 0), c=(0, 0, 0)):
        self.h.append(F.Hittable(kind, mat, first, n, F.D3(*a), F.D3(*b), F.D3(*c)))
        return len(self.h) - 1

    def sphere(self, center, radius, mat, center2=None):
        b = (0, 0, 0) if center2 is None else tuple(center2[i] - center[i] for i in range(3))
        return self._hit(F.H_SPHERE, mat, -1, 0, center, b, (radius, 0, 0))

    def quad(self, q, u, v, mat):
        return self._hit(F.H_QUAD, mat, -1, 0, q, u, v)

    def list(self, objs, kind=F.H_LIST):
        first = len(self.ch)
        self.ch.extend(objs)
        return self._hit(kind, -1, first, len(objs))

    def bvh(self, objs):
        return self.list(objs, F.H_BVH)

    def translate(self, obj, offset):
        return self._hit(F.H_TRANSLATE, -1, obj, 0, offset)

    def rotate_y(self, obj, angle):
        return self._hit(F.H_ROTATE_Y, -1, obj, 0, (angle, 0, 0))

    def medium(self, boundary, density, phase_mat):
        return self._hit(F.H_CONSTANT_MEDIUM, phase_mat, boundary, 0, (density, 0, 0))

    def desc(self, world, lights=-1):
        def arr(T, items):
            a = (T * max(1, len(items)))(*items)
            self._keep.append(a)
            return a
        d = F.SceneDesc()
        d.hittables = arr(F.Hittable, self.h); d.n_hittables = len(self.h)
        d.children = arr(F.c_int, self.ch); d.n_children = len(self.ch)
        d.materials = arr(F.Material, self.m); d.n_materials = len(self.m)
        d.textures = arr(F.Texture, self.t); d.n_textures = len(self.t)
        d.images = arr(F.Image, self.im); d.n_images = len(self.im)
        d.perlins = arr(F.Perlin, self.pn); d.n_perlins = len(self.pn)
        d.world = world; d.lights = lights
        return d


def make_camera(image_width=100, aspect_ratio=1.0, samples_per_pixel=10, max_depth=10, background=(0, 0, 0),
                vfov=90.0, lookfrom=(0, 0, 0), lookat=(0, 0, -1), vup=(0, 1, 0), defocus_angle=0.0,
                focus_dist=10.0):
    """Camera::default (the_rest_of_your_life/camera.rs:31-47) with keyword changes."""
    return F.Camera(aspect_ratio, image_width, samples_per_pixel, max_depth, 0, F.D3(*background), vfov,
                    F.D3(*lookfrom), F.D3(*lookat), F.D3(*vup), defocus_angle, focus_dist)


def flatten_check(desc):
    """Host-only: run the flattener + BVH build and its self-check. Returns a dict of counts."""
    v = [F.c_int() for _ in range(5)]
    _check(F.rrt().rrt_flatten_check(C.byref(desc), *[C.byref(x) for x in v]))
    return dict(zip(("n_prims", "n_nodes", "bvh_depth", "n_media", "n_lights"), [x.value for x in v]))


def flatten_rects(desc):
    """Host-only: (nx, ny, nz) axis-aligned rectangles of a brute-force root leaf."""
    n = (F.c_int * 3)()
    _check(F.rrt().rrt_flatten_rects(C.byref(desc), C.byref(n)))
    return tuple(n)


def flatten_grid(desc, kind=0):
    """Host-only: the uniform grid the flattener chose for the scene's field of spheres (kind 0) or boxes (kind 1);
    res (0, 0, 0): none."""
    res, n_items, n_refs, most = (F.c_int * 3)(), F.c_int(), F.c_int(), F.c_int()
    _check(F.rrt().rrt_flatten_grid(C.byref(desc), int(kind), C.byref(res), C.byref(n_items), C.byref(n_refs), C.byref(most)))
    return dict(res=tuple(res), n_items=n_items.value, n_refs=n_refs.value, max_per_cell=most.value)


def scene_features(desc):
    """Host-only: SF_* feature bits of the flattened scene (which render-kernel instance it runs on)."""
    v = F.c_int()
    _check(F.rrt().rrt_scene_features(C.byref(desc), C.byref(v)))
    return v.value


def shard_plan(camera, integrator, shard_mode, rank, world, mask=False):
    """rrt_shard_plan (host only): this rank's sample range / tiles / pixel count [/ owner mask]."""
    p = F.RenderParams(int(integrator), int(shard_mode), int(rank), int(world), 0)
    sb, se, nt, npx = F.c_int(), F.c_int(), F.c_int(), F.c_dbl()
    m = None
    if mask:
        H = F.rrt().rrt_image_height(C.byref(camera))
        m = np.zeros((H, camera.image_width), dtype=np.uint8)
    _check(F.rrt().rrt_shard_plan(C.byref(camera), C.byref(p), C.byref(sb), C.byref(se), C.byref(nt), C.byref(npx),
                                  m.ctypes.data if m is not None else None))
    out = dict(sample_begin=sb.value, sample_end=se.value, n_super_tiles=nt.value, n_pixels=npx.value)
    if mask:
        out["mask"] = m
    return out


def tile_scatter_host(camera, shares):
    """rrt_tile_scatter_host (host only): per-rank compact tile buffers (float32, n_owned * 1024 * 4) -> (H, W, 4) frame."""
    H = F.rrt().rrt_image_height(C.byref(camera))
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in shares]
    ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
    frame = np.zeros((H, camera.image_width, 4), dtype=np.float32)
    _check(F.rrt().rrt_tile_scatter_host(C.byref(camera), len(arrs), ptrs, frame.ctypes.data))
    return frame


class GpuScene:
    def __init__(self, ctx, handle):
        self.ctx, self._handle = ctx, handle

    def close(self):
        if self._handle:
            F.rrt().rrt_scene_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """One CUDA device (rrt_init).  Raises RrtError(ERR_NO_DEVICE) when there is no GPU."""

    def __init__(self, device=0, devices=None):
        """device: one GPU (rrt_init).  devices=[0, 1, ...]: a group over several GPUs of this box driven by
        this process (rrt_init_multi); every method then returns the complete frame, exchanged over NCCL."""
        self._h = C.c_void_p()
        if devices is not None:
            arr = (F.c_int * len(devices))(*[int(d) for d in devices])
            _check(F.rrt().rrt_init_multi(arr, len(devices), C.byref(self._h)))
            self.device, self.devices = int(devices[0]), [int(d) for d in devices]
        else:
            _check(F.rrt().rrt_init(int(device), C.byref(self._h)))
            self.device, self.devices = int(device), [int(device)]

    # ---- one process per GPU: an NCCL communicator inside the library (rrt_comm_*) -----------------
    @staticmethod
    def comm_unique_id():
        """rrt_comm_unique_id: 128 bytes made on rank 0, to be handed to every rank (bytes)."""
        buf = (C.c_uint8 * 128)()
        _check(F.rrt().rrt_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, unique_id, rank, world):
        """rrt_comm_init (collective): join the communicator as `rank` of `world`."""
        buf = (C.c_uint8 * 128)(*unique_id)
        _check(F.rrt().rrt_comm_init(self._h, buf, int(rank), int(world)))
        self.rank, self.world = int(rank), int(world)

    def render_collective(self, gscene, camera, integrator, seed=1, shard_mode=F.SHARD_SAMPLE, accum_ptr=None, out=None):
        """rrt_render_collective (collective): this rank's share + the NCCL exchange; on rank 0 the frame goes to
        device memory `accum_ptr` (None: library-owned) and, if `out` (host (H, W, 4) float32 array) is given, there."""
        p = self.params(integrator, seed, shard_mode, 0, 1)
        _check(F.rrt().rrt_render_collective(self._h, gscene._handle, C.byref(camera), C.byref(p),
                                             C.c_void_p(int(accum_ptr)) if accum_ptr else None,
                                             out.ctypes.data if out is not None else None))
        return out

    def render_frame_collective(self, desc, camera, integrator, seed=1, shard_mode=F.SHARD_SAMPLE, out=None):
        """rrt_render_frame_collective (collective): upload, this rank's share, NCCL exchange, free; the frame lands in `out` on rank 0."""
        p = self.params(integrator, seed, shard_mode, 0, 1)
        _check(F.rrt().rrt_render_frame_collective(self._h, C.byref(desc), C.byref(camera), C.byref(p),
                                                   out.ctypes.data if out is not None else None))
        return out

    def close(self):
        if self._h:
            F.rrt().rrt_shutdown(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        """rrt_set_stream: launch on the caller's stream (int handle; 0/None = the library's own)."""
        _check(F.rrt().rrt_set_stream(self._h, C.c_void_p(int(cuda_stream) if cuda_stream else None)))

    def scene(self, desc):
        h = C.c_void_p()
        _check(F.rrt().rrt_scene_create(self._h, C.byref(desc), C.byref(h)))
        g = GpuScene(self, h)
        g.n_lights = self.stats()["n_lights"]
        return g

    @staticmethod
    def params(integrator, seed=1, shard_mode=F.SHARD_NONE, rank=0, world=1):
        return F.RenderParams(int(integrator), int(shard_mode), int(rank), int(world), int(seed))

    def render(self, gscene, camera, integrator, seed=1, shard_mode=F.SHARD_NONE, rank=0, world=1, out=None):
        """rrt_render: returns (H, W, 4) float32 host array of RGBA sums (.w = samples)."""
        H = F.rrt().rrt_image_height(C.byref(camera))
        if out is None:
            out = np.empty((H, camera.image_width, 4), dtype=np.float32)
        p = self.params(integrator, seed, shard_mode, rank, world)
        _check(F.rrt().rrt_render(self._h, gscene._handle, C.byref(camera), C.byref(p), out.ctypes.data))
        return out

    def render_device(self, gscene, camera, integrator, accum_ptr, seed=1, shard_mode=F.SHARD_NONE, rank=0, world=1):
        """rrt_render_device: accum_ptr is a device pointer (e.g. torch_tensor.data_ptr())."""
        p = self.params(integrator, seed, shard_mode, rank, world)
        _check(F.rrt().rrt_render_device(self._h, gscene._handle, C.byref(camera), C.byref(p), accum_ptr))

    def render_frame(self, desc, camera, integrator, seed=1, out=None):
        """rrt_render_frame: upload + render + read back + free (twin of src/cuda/mod.rs:342-441)."""
        H = F.rrt().rrt_image_height(C.byref(camera))
        if out is None:
            out = np.empty((H, camera.image_width, 4), dtype=np.float32)
        p = self.params(integrator, seed)
        _check(F.rrt().rrt_render_frame(self._h, C.byref(desc), C.byref(camera), C.byref(p), out.ctypes.data))
        return out

    def probe(self, gscene, camera, precision=64):
        H = F.rrt().rrt_image_height(C.byref(camera))
        ids = np.empty((H, camera.image_width), dtype=np.int32)
        t = np.empty((H, camera.image_width), dtype=np.float64)
        _check(F.rrt().rrt_probe(self._h, gscene._handle, C.byref(camera), int(precision), ids.ctypes.data, t.ctypes.data))
        return ids, t

    def probe_static(self, gscene, camera, precision=64):
        """rrt_probe_static: the probe without any device allocation (static scratch, <= F.PROBE_STATIC_MAX_PIXELS pixels)."""
        H = F.rrt().rrt_image_height(C.byref(camera))
        ids = np.empty((H, camera.image_width), dtype=np.int32)
        t = np.empty((H, camera.image_width), dtype=np.float64)
        _check(F.rrt().rrt_probe_static(self._h, gscene._handle, C.byref(camera), int(precision), ids.ctypes.data, t.ctypes.data))
        return ids, t

    def resolve(self, accum, spp):
        a = np.ascontiguousarray(accum, dtype=np.float32)
        H, W = a.shape[0], a.shape[1]
        rgb = np.empty((H, W, 3), dtype=np.uint8)
        _check(F.rrt().rrt_resolve_rgb8(self._h, a.ctypes.data, W, H, int(spp), rgb.ctypes.data))
        return rgb

    def resolve_device(self, accum_ptr, width, height, spp, rgb_ptr):
        _check(F.rrt().rrt_resolve_rgb8_device(self._h, accum_ptr, width, height, int(spp), rgb_ptr))

    def kat_light(self, gscene, origin, direction, u):
        """rrt_kat_light: the device light-sampling code on n queries.  origin, direction (n, 3), u (n, 2).
        Returns (pdf_each (n, n_lights), pdf_list (n,), random_dir (n, n_lights, 3))."""
        o = np.ascontiguousarray(origin, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(direction, dtype=np.float64).reshape(-1, 3)
        uu = np.ascontiguousarray(u, dtype=np.float64).reshape(-1, 2)
        n = o.shape[0]
        nl = self.stats()["n_lights"] if gscene is None else gscene.n_lights
        each = np.zeros((n, nl), dtype=np.float64)
        lst = np.zeros(n, dtype=np.float64)
        rnd = np.zeros((n, nl, 3), dtype=np.float64)
        _check(F.rrt().rrt_kat_light(self._h, gscene._handle, n, o.ctypes.data, d.ctypes.data, uu.ctypes.data,
                                     each.ctypes.data, lst.ctypes.data, rnd.ctypes.data))
        return each, lst, rnd

    def kat_rng(self, seed, pixel, sample0, bounce, block, n):
        """rrt_kat_rng: (raw (n, 4) uint32 Philox words, uniforms (n, 5) float32) of the device RNG."""
        raw = np.zeros((n, 4), dtype=np.uint32)
        u5 = np.zeros((n, 5), dtype=np.float32)
        _check(F.rrt().rrt_kat_rng(self._h, int(seed), int(pixel), int(sample0), int(bounce), int(block), int(n),
                                   raw.ctypes.data, u5.ctypes.data))
        return raw, u5

    def stats(self):
        s = F.Stats()
        _check(F.rrt().rrt_get_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in F.Stats._fields_ if not k.startswith("_")}
