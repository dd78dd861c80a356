"""A second witness for the scene builders of librrt_host.so (rustraytrace_b200/host/books.cpp).

Every parity test hands the oracle and the GPU the same rrt_scene_desc, so an error in the C++ scene
builders would be invisible to them.  This module restates the scenes behind every BASELINE.json config in
pure Python, read from the reference's Rust sources and NOT from books.cpp:
    in_one_weekend::run              src/books/in_one_weekend/mod.rs:57-126         (configs C1, C2)
    the_next_week::bouncing_spheres  src/books/the_next_week/mod.rs:83-152          (C3a)
    the_next_week::earth, perlin_spheres                    :196-254                (C3b, C3c)
    the_next_week::final_scene       src/books/the_next_week/mod.rs:507-624         (C4)
    the_rest_of_your_life::run       src/books/the_rest_of_your_life/mod.rs:69-161  (C5, with its light list)
with rand 0.8.5's SmallRng (Xoshiro256++ seeded through SplitMix64) and its gen_range(low..high)
mapping for f64 restated from their published algorithms, in the reference's order of draws.
tests/test_scene_witness.py walks the C description into the same nested form and compares.

Test infrastructure only.  Nested form: ("sphere", c, dc, r, mat) ("quad", Q, u, v, mat) ("list" | "bvh", [..])
("translate", offset, child) ("rotate_y", degrees, child) ("medium", density, phase_mat, child);
mat = (kind, albedo, fuzz, ri, tex); tex = ("solid", rgb) ("checker", scale, even, odd) ("image",) ("noise", scale, perlin) or None.
"""
import math
import struct

M64 = (1 << 64) - 1


class SmallRng:
    """rand 0.8.5 SmallRng on 64-bit targets: xoshiro256++ (Blackman & Vigna), seed_from_u64 = SplitMix64."""

    def __init__(self, seed):
        self.s = []
        state = seed & M64
        for _ in range(4):
            state = (state + 0x9E3779B97F4A7C15) & M64
            z = state
            z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
            z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
            self.s.append(z ^ (z >> 31))

    @staticmethod
    def _rotl(x, k):
        return ((x << k) | (x >> (64 - k))) & M64

    def next_u64(self):
        s = self.s
        result = (self._rotl((s[0] + s[3]) & M64, 23) + s[0]) & M64
        t = (s[1] << 17) & M64
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]
        s[2] ^= t
        s[3] = self._rotl(s[3], 45)
        return result

    def gen_range(self, low, high):
        """UniformFloat<f64>::sample_single: a [1, 2) mantissa fill minus one, scaled; redrawn if it rounds to high."""
        scale = high - low
        while True:
            v = struct.unpack("<d", struct.pack("<Q", (self.next_u64() >> 12) | 0x3FF0000000000000))[0]
            res = (v - 1.0) * scale + low
            if res < high:
                return res

    # rtweekend.rs:17-30
    def random_double(self):
        return self.gen_range(0.0, 1.0)

    def random_double_range(self, lo, hi):
        return self.gen_range(lo, hi)

    def random_int(self, lo, hi):
        return int(self.random_double_range(float(lo), float(hi + 1)))

    # vec3.rs:43-53
    def random_vec(self):
        return (self.random_double(), self.random_double(), self.random_double())

    def random_vec_range(self, lo, hi):
        return (self.random_double_range(lo, hi), self.random_double_range(lo, hi), self.random_double_range(lo, hi))


LAMBERTIAN, METAL, DIELECTRIC, DIFFUSE_LIGHT, ISOTROPIC = "lambertian", "metal", "dielectric", "diffuse_light", "isotropic"


def lambertian(c):
    return (LAMBERTIAN, tuple(c), 0.0, 1.0, ("solid", tuple(c)))


def metal(c, fuzz):
    return (METAL, tuple(c), min(fuzz, 1.0), 1.0, None)       # material.rs: fuzz clamped to 1


def dielectric(ri):
    return (DIELECTRIC, None, 0.0, ri, None)


def sphere(c, r, mat, c2=None):
    dc = (0.0, 0.0, 0.0) if c2 is None else tuple(c2[i] - c[i] for i in range(3))
    return ("sphere", tuple(c), dc, max(r, 0.0), mat)         # sphere.rs: radius.max(0.0)


def length(v):
    return math.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])


def book1_final(seed):
    """in_one_weekend/mod.rs:57-126.  Returns (world, camera dict)."""
    g = SmallRng(seed)
    world = [sphere((0.0, -1000.0, 0.0), 1000.0, lambertian((0.5, 0.5, 0.5)))]
    for a in range(-11, 11):
        for b in range(-11, 11):
            choose_mat = g.random_double()
            center = (a + 0.9 * g.random_double(), 0.2, b + 0.9 * g.random_double())
            if length((center[0] - 4.0, center[1] - 0.2, center[2] - 0.0)) > 0.9:
                if choose_mat < 0.8:
                    c1, c2 = g.random_vec(), g.random_vec()
                    mat = lambertian(tuple(c1[i] * c2[i] for i in range(3)))
                elif choose_mat < 0.95:
                    albedo = g.random_vec_range(0.5, 1.0)
                    mat = metal(albedo, g.random_double() * 0.5)
                else:
                    mat = dielectric(1.5)
                world.append(sphere(center, 0.2, mat))
    world.append(sphere((0.0, 1.0, 0.0), 1.0, dielectric(1.5)))
    world.append(sphere((-4.0, 1.0, 0.0), 1.0, lambertian((0.4, 0.2, 0.1))))
    world.append(sphere((4.0, 1.0, 0.0), 1.0, metal((0.7, 0.6, 0.5), 0.0)))
    cam = dict(aspect_ratio=16.0 / 9.0, image_width=1200, samples_per_pixel=10, max_depth=20, vfov=20.0,
               lookfrom=(13.0, 2.0, 3.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.6, focus_dist=10.0)
    return ("bvh", world), cam


def make_box(a, b, mat):
    """quad.rs:90-112: six quads, fixed order, as a list."""
    mn = tuple(min(a[i], b[i]) for i in range(3))
    mx = tuple(max(a[i], b[i]) for i in range(3))
    dx, dy, dz = (mx[0] - mn[0], 0.0, 0.0), (0.0, mx[1] - mn[1], 0.0), (0.0, 0.0, mx[2] - mn[2])
    neg = lambda v: tuple(-x for x in v)
    return ("list", [
        ("quad", (mn[0], mn[1], mx[2]), dx, dy, mat),
        ("quad", (mx[0], mn[1], mx[2]), neg(dz), dy, mat),
        ("quad", (mx[0], mn[1], mn[2]), neg(dx), dy, mat),
        ("quad", (mn[0], mn[1], mn[2]), dz, dy, mat),
        ("quad", (mn[0], mx[1], mx[2]), dx, neg(dz), mat),
        ("quad", (mn[0], mn[1], mn[2]), dx, dz, mat)])


def perlin_new(g):
    """perlin.rs:11-23, 65-76: 256 unit vectors, then three permutations (Fisher-Yates from the top, random_int(0, i))."""
    randvec = []
    for _ in range(256):
        v = g.random_vec_range(-1.0, 1.0)
        k = 1.0 / length(v)                      # unit_vector = v / |v| = (1 / |v|) * v, vec3.rs:142-147, 168-170
        randvec.append((k * v[0], k * v[1], k * v[2]))
    perms = []
    for _ in range(3):
        p = list(range(256))
        for i in range(255, 0, -1):
            t = g.random_int(0, i)
            p[i], p[t] = p[t], p[i]
        perms.append(p)
    return (randvec, perms[0], perms[1], perms[2])


def book2_final(seed, image_width=800, samples_per_pixel=10000, max_depth=40):
    """the_next_week/mod.rs:507-624 (called as final_scene(800, 10000, 40), :78)."""
    g = SmallRng(seed)
    ground = lambertian((0.48, 0.83, 0.53))
    boxes1 = []
    for i in range(20):
        for j in range(20):
            w = 100.0
            x0, z0, y0 = -1000.0 + i * w, -1000.0 + j * w, 0.0
            x1, y1, z1 = x0 + w, g.random_double() * 100.0 + 1.0, z0 + w
            boxes1.append(make_box((x0, y0, z0), (x1, y1, z1), ground))
    world = [("bvh", boxes1)]
    light = (DIFFUSE_LIGHT, (7.0, 7.0, 7.0), 0.0, 1.0, ("solid", (7.0, 7.0, 7.0)))
    world.append(("quad", (123.0, 554.0, 147.0), (300.0, 0.0, 0.0), (0.0, 0.0, 265.0), light))
    c1 = (400.0, 400.0, 200.0)
    world.append(sphere(c1, 50.0, lambertian((0.7, 0.3, 0.1)), c2=(c1[0] + 30.0, c1[1], c1[2])))
    world.append(sphere((260.0, 150.0, 45.0), 50.0, dielectric(1.5)))
    world.append(sphere((0.0, 150.0, 145.0), 50.0, metal((0.8, 0.8, 0.9), 1.0)))
    boundary = sphere((360.0, 150.0, 145.0), 70.0, dielectric(1.5))
    world.append(boundary)
    iso = lambda c: (ISOTROPIC, tuple(c), 0.0, 1.0, ("solid", tuple(c)))
    world.append(("medium", 0.2, iso((0.2, 0.4, 0.9)), boundary))
    world.append(("medium", 0.0001, iso((1.0, 1.0, 1.0)), sphere((0.0, 0.0, 0.0), 5000.0, dielectric(1.5))))
    world.append(sphere((400.0, 200.0, 400.0), 100.0, (LAMBERTIAN, None, 0.0, 1.0, ("image",))))
    pertext = ("noise", 0.2, perlin_new(g))
    world.append(sphere((220.0, 280.0, 300.0), 80.0, (LAMBERTIAN, None, 0.0, 1.0, pertext)))
    white = lambertian((0.73, 0.73, 0.73))
    boxes2 = [sphere(g.random_vec_range(0.0, 165.0), 10.0, white) for _ in range(1000)]
    world.append(("translate", (-100.0, 270.0, 395.0), ("rotate_y", 15.0, ("bvh", boxes2))))
    cam = dict(aspect_ratio=1.0, image_width=image_width, samples_per_pixel=samples_per_pixel, max_depth=max_depth,
               background=(0.0, 0.0, 0.0), vfov=40.0, lookfrom=(478.0, 278.0, -600.0), lookat=(278.0, 278.0, 0.0),
               vup=(0.0, 1.0, 0.0), defocus_angle=0.0)
    return ("list", world), cam


def checker(scale, even, odd):
    return ("checker", scale, ("solid", tuple(even)), ("solid", tuple(odd)))


def book2_bouncing_spheres(seed):
    """the_next_week/mod.rs:83-152 (scene 1; config C3a)."""
    g = SmallRng(seed)
    ground = (LAMBERTIAN, None, 0.0, 1.0, checker(0.32, (0.2, 0.3, 0.1), (0.9, 0.9, 0.9)))
    world = [sphere((0.0, -1000.0, 0.0), 1000.0, ground)]
    for a in range(-11, 11):
        for b in range(-11, 11):
            choose_mat = g.random_double()
            center = (a + 0.9 * g.random_double(), 0.2, b + 0.9 * g.random_double())
            if length((center[0] - 4.0, center[1] - 0.2, center[2] - 0.0)) > 0.9:
                if choose_mat < 0.8:
                    c1, c2 = g.random_vec(), g.random_vec()
                    mat = lambertian(tuple(c1[i] * c2[i] for i in range(3)))
                    center2 = (center[0], center[1] + g.random_double() * 0.5, center[2])
                    world.append(sphere(center, 0.2, mat, c2=center2))
                elif choose_mat < 0.95:
                    albedo = g.random_vec_range(0.5, 1.0)
                    world.append(sphere(center, 0.2, metal(albedo, g.random_double() * 0.5)))
                else:
                    world.append(sphere(center, 0.2, dielectric(1.5)))
    world.append(sphere((0.0, 1.0, 0.0), 1.0, dielectric(1.5)))
    world.append(sphere((-4.0, 1.0, 0.0), 1.0, lambertian((0.4, 0.2, 0.1))))
    world.append(sphere((4.0, 1.0, 0.0), 1.0, metal((0.7, 0.6, 0.5), 0.0)))
    cam = dict(aspect_ratio=16.0 / 9.0, image_width=400, samples_per_pixel=100, max_depth=50, background=(0.70, 0.80, 1.00),
               vfov=20.0, lookfrom=(13.0, 2.0, 3.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.6, focus_dist=10.0)
    return ("bvh", world), cam


def book2_earth():
    """the_next_week/mod.rs:196-219 (scene 3; config C3b)."""
    world = [sphere((0.0, 0.0, 0.0), 2.0, (LAMBERTIAN, None, 0.0, 1.0, ("image",)))]
    cam = dict(aspect_ratio=16.0 / 9.0, image_width=400, samples_per_pixel=100, max_depth=50, background=(0.70, 0.80, 1.00),
               vfov=20.0, lookfrom=(0.0, 0.0, 12.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0)
    return ("list", world), cam


def book2_perlin_spheres(seed):
    """the_next_week/mod.rs:221-254 (scene 4; config C3c): one NoiseTexture shared by both spheres."""
    g = SmallRng(seed)
    pertext = ("noise", 4.0, perlin_new(g))
    world = [sphere((0.0, -1000.0, 0.0), 1000.0, (LAMBERTIAN, None, 0.0, 1.0, pertext)),
             sphere((0.0, 2.0, 0.0), 2.0, (LAMBERTIAN, None, 0.0, 1.0, pertext))]
    cam = dict(aspect_ratio=16.0 / 9.0, image_width=400, samples_per_pixel=100, max_depth=50, background=(0.70, 0.80, 1.00),
               vfov=20.0, lookfrom=(13.0, 2.0, 3.0), lookat=(0.0, 0.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0)
    return ("list", world), cam


def book3_cornell():
    """the_rest_of_your_life/mod.rs:69-161 (config C5).  Returns (world, lights, camera dict)."""
    red, white, green = lambertian((0.65, 0.05, 0.05)), lambertian((0.73, 0.73, 0.73)), lambertian((0.12, 0.45, 0.15))
    light = (DIFFUSE_LIGHT, (15.0, 15.0, 15.0), 0.0, 1.0, ("solid", (15.0, 15.0, 15.0)))
    world = [("quad", (555.0, 0.0, 0.0), (0.0, 0.0, 555.0), (0.0, 555.0, 0.0), green),
             ("quad", (0.0, 0.0, 555.0), (0.0, 0.0, -555.0), (0.0, 555.0, 0.0), red),
             ("quad", (0.0, 555.0, 0.0), (555.0, 0.0, 0.0), (0.0, 0.0, 555.0), white),
             ("quad", (0.0, 0.0, 555.0), (555.0, 0.0, 0.0), (0.0, 0.0, -555.0), white),
             ("quad", (555.0, 0.0, 555.0), (-555.0, 0.0, 0.0), (0.0, 555.0, 0.0), white),
             ("quad", (213.0, 554.0, 227.0), (130.0, 0.0, 0.0), (0.0, 0.0, 105.0), light),
             ("translate", (265.0, 0.0, 295.0), ("rotate_y", 15.0, make_box((0.0, 0.0, 0.0), (165.0, 330.0, 165.0), white))),
             sphere((190.0, 90.0, 190.0), 90.0, dielectric(1.5))]
    empty = ("empty", None, 0.0, 1.0, None)
    lights = ("list", [("quad", (343.0, 554.0, 332.0), (-130.0, 0.0, 0.0), (0.0, 0.0, -105.0), empty),
                       sphere((190.0, 90.0, 190.0), 90.0, empty)])
    cam = dict(aspect_ratio=1.0, image_width=600, samples_per_pixel=100, max_depth=50, background=(0.0, 0.0, 0.0), vfov=40.0,
               lookfrom=(278.0, 278.0, -800.0), lookat=(278.0, 278.0, 0.0), vup=(0.0, 1.0, 0.0), defocus_angle=0.0)
    return ("bvh", world), lights, cam
