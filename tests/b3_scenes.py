"""Hand-built scenes for the book-3 integrator tests (tests/test_book3_sampling.py).

furnace(): the white-furnace test of SURVEY.md §4.  A closed box of emitters of radiance E (six quads,
normals inward, so that book 3's front-face-only DiffuseLight, material.rs:155-160, emits into the box)
around energy-conserving objects: albedo-1 Lambertian spheres and quads (solid / checker / image
textures that all evaluate to 1), a metal of albedo 1, a dielectric, a moving sphere, a constant
medium with an albedo-1 Isotropic phase function, and a second emitter (a sphere of radiance E).
Every surface a path can end on has radiance E and every scatter conserves energy, so every camera
ray's expected radiance is E, whatever the sampling pdf — any normalisation error in the mixture pdf
(quad pdf, sphere pdf, list mean, cosine lobe, 1/4pi phase function: quad.rs:93-107, sphere.rs:102-122,
hittable_list.rs:60-75, pdf.rs:8-100, material.rs:53-71, 163-190, camera.rs:239-253) shows up as a
deviation from E.

random_b3(seed): random scenes with a lights list (quad + sphere proxies) for the GPU-vs-oracle
comparison under RRT_INT_BOOK3: isotropic media, metal, dielectric, checker / image / noise textures,
moving spheres, instanced boxes."""
import numpy as np

import rustraytrace_b200 as R

F = R._ffi
E = (0.9, 0.7, 0.5)


def make_perlin(rng):
    """Perlin::new (perlin.rs:11-24): 256 unit vectors of U[-1,1)^3 and three permutations."""
    p = F.Perlin()
    for i in range(256):
        v = rng.uniform(-1, 1, 3)
        v = v / np.linalg.norm(v)
        for k in range(3):
            p.randvec[i][k] = float(v[k])
    for arr in (p.perm_x, p.perm_y, p.perm_z):
        perm = rng.permutation(256)
        for i in range(256):
            arr[i] = int(perm[i])
    return p


def inward_box(b, h, mat):
    """Six quads of the cube [-h, h]^3 whose normals cross(u, v) point inward."""
    s = 2.0 * h
    return [
        b.quad((-h, -h, -h), (0, 0, s), (s, 0, 0), mat),   # floor,  n = +y
        b.quad((-h, h, -h), (s, 0, 0), (0, 0, s), mat),    # ceiling, n = -y
        b.quad((-h, -h, -h), (0, s, 0), (0, 0, s), mat),   # x = -h, n = +x
        b.quad((h, -h, -h), (0, 0, s), (0, s, 0), mat),    # x = +h, n = -x
        b.quad((-h, -h, -h), (s, 0, 0), (0, s, 0), mat),   # z = -h, n = +z
        b.quad((-h, -h, h), (0, s, 0), (s, 0, 0), mat),    # z = +h, n = -z
    ]


def furnace(book3, width=48, spp=256, depth=50):
    b = R.DescBuilder()
    light = b.material(F.M_DIFFUSE_LIGHT, b.solid(E))
    one = b.solid((1, 1, 1))
    white = b.material(F.M_LAMBERTIAN, one)
    white_checker = b.material(F.M_LAMBERTIAN, b.checker(0.7, one, b.solid((1, 1, 1))))
    white_image = b.material(F.M_LAMBERTIAN, b.image(np.full((4, 8, 3), 255, dtype=np.uint8)))
    mirror = b.material(F.M_METAL, albedo=(1, 1, 1), fuzz=0.3 if book3 else 0.0)   # b1/b2 metal absorbs below-surface rays
    glass = b.material(F.M_DIELECTRIC, ri=1.5)
    phase = b.material(F.M_ISOTROPIC, one)
    walls = inward_box(b, 10.0, light)
    lamp = b.sphere((0, 0, -4), 1.5, light)
    fog_boundary = b.sphere((0, -3.5, 1), 2.0, glass)   # the boundary is not in the world, only inside the medium
    objs = walls + [
        lamp,
        b.sphere((-3.5, 0, 0), 2.0, white),
        b.sphere((3.5, 0, 0), 2.0, mirror),
        b.sphere((0, 3.5, 0), 1.5, glass),
        b.sphere((-3.5, 3.5, 2), 1.0, white_checker, center2=(-3.5, 4.2, 2)),
        b.sphere((3.5, 3.5, 2), 1.0, white_image),
        b.quad((-2, -1, 3), (4, 0, 0), (0, 1.5, 1), white),
        b.medium(fog_boundary, 0.6, phase),
        b.translate(b.rotate_y(b.list(inward_box(b, 0.8, white)), 30.0), (3.5, -3.5, 2.0)),
    ]
    world = b.bvh(objs)
    lights = b.list([walls[1], lamp]) if book3 else -1
    cam = R.make_camera(image_width=width, aspect_ratio=1.0, samples_per_pixel=spp, max_depth=depth, background=E,
                        vfov=70.0, lookfrom=(0.5, 0.3, 9.0), lookat=(0, 0, 0))
    return b, b.desc(world, lights), cam


def random_b3(seed, width=48, spp=1024):
    rng = np.random.default_rng(5000 + seed)
    b = R.DescBuilder()
    perlin = make_perlin(rng)
    img = rng.integers(30, 255, (8, 16, 3)).astype(np.uint8)
    texs = [b.solid(tuple(rng.uniform(0.3, 0.9, 3))), b.checker(0.8, b.solid((0.2, 0.3, 0.1)), b.solid((0.9, 0.9, 0.9))),
            b.image(img), b.noise(2.0, perlin)]
    lamb = [b.material(F.M_LAMBERTIAN, t) for t in texs]
    metal = b.material(F.M_METAL, albedo=tuple(rng.uniform(0.6, 0.95, 3)), fuzz=float(rng.uniform(0, 0.4)))
    glass = b.material(F.M_DIELECTRIC, ri=1.5)
    lightm = b.material(F.M_DIFFUSE_LIGHT, b.solid((7, 7, 6)))
    phase = b.material(F.M_ISOTROPIC, b.solid(tuple(rng.uniform(0.5, 0.95, 3))))
    mats = lamb + [metal, glass]
    pick = lambda: mats[rng.integers(len(mats))]
    objs = [b.quad((-6, -2, -6), (12, 0, 0), (0, 0, 12), lamb[1]),                 # floor (checker)
            b.quad((-6, -2, -6), (12, 0, 0), (0, 9, 0), lamb[0])]                   # back wall
    lq = b.quad((-1.5, 5.5, -1.5), (3, 0, 0), (0, 0, 3), lightm)                    # n = -y: emits downward
    ls = b.sphere((float(rng.uniform(-4, -2)), float(rng.uniform(1, 3)), float(rng.uniform(-3, 0))), 0.6, lightm)
    objs += [lq, ls]
    lc = np.array(b.h[ls].a)

    def clear_of_light(c, r):
        """Nothing may reach into the sphere light: a scattering point inside it has cos_theta_max = NaN
        (sphere.rs:110), which the reference turns into a black pixel and the GPU into a dropped sample."""
        return np.linalg.norm(np.array(c) - lc) > r + 0.6 + 0.15

    for _ in range(int(rng.integers(3, 7))):
        while True:
            c = (float(rng.uniform(-4, 4)), float(rng.uniform(-1.2, 2.5)), float(rng.uniform(-4, 3)))
            r = float(rng.uniform(0.4, 1.1))
            if clear_of_light(c, r + 0.5):    # + the moving spheres' travel
                break
        if rng.integers(3) == 0:
            objs.append(b.sphere(c, r, pick(), center2=(c[0], c[1] + float(rng.uniform(0.1, 0.5)), c[2])))
        else:
            objs.append(b.sphere(c, r, pick()))
    lo = rng.uniform(-3, 1, 3)
    while not clear_of_light(lo, 1.3):
        lo = rng.uniform(-3, 1, 3)
    bx = inward_box(b, 0.7, pick())
    objs.append(b.translate(b.rotate_y(b.list(bx), float(rng.uniform(-40, 40))), tuple(lo)))
    mc = (float(rng.uniform(-2, 2)), 0.5, float(rng.uniform(-2, 2)))
    while not clear_of_light(mc, 1.3):
        mc = (float(rng.uniform(-2, 2)), 0.5, float(rng.uniform(-2, 2)))
    objs.append(b.medium(b.sphere(mc, 1.3, glass), float(rng.uniform(0.3, 1.5)), phase))
    world = b.bvh(objs) if rng.integers(2) else b.list(objs)
    empty = b.material(F.M_EMPTY)
    lights = b.list([b.quad((1.5, 5.5, 1.5), (-3, 0, 0), (0, 0, -3), empty), b.sphere(tuple(b.h[ls].a), 0.6, empty)])
    cam = R.make_camera(image_width=width, aspect_ratio=1.0, samples_per_pixel=spp, max_depth=30, background=(0.02, 0.03, 0.05),
                        vfov=55.0, lookfrom=(float(rng.uniform(-1, 1)), 2.0, 11.0), lookat=(0, 1, 0))
    return b, b.desc(world, lights), cam
