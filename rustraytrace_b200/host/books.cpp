// books.cpp — host mirror of the reference's scene builders and camera defaults:
//   src/books/in_one_weekend/mod.rs:57-126, src/books/the_next_week/mod.rs:68-624,
//   src/books/the_rest_of_your_life/mod.rs:69-161, src/gpu/mod.rs:124-301 (seeded GPU-path scene),
//   apply_overrides (in_one_weekend/mod.rs:23-55 etc.) and src/config.rs.
// Every builder stops where the Rust calls cam.render(&world[, lights]); that call is the
// rrt.h boundary.  Random draws are issued in the reference's order (Rust evaluates operands and
// arguments left to right; here each draw is its own statement).
#include <cstdio>
#include <cstdlib>
#include <string>

#include "../../include/rrt_host.h"
#include "scene_builder.hpp"

namespace rrt_host {

struct Scene {
  SceneBuilder b;
  rrt_scene_desc desc;
  rrt_camera cam;
  int integrator;
  explicit Scene(uint64_t seed) : b(seed), integrator(RRT_INT_BOOK2) {}
};

// Camera::default, the_rest_of_your_life/camera.rs:31-47
static rrt_camera camera_default() {
  rrt_camera c; std::memset(&c, 0, sizeof c);
  c.aspect_ratio = 1.0; c.image_width = 100; c.samples_per_pixel = 10; c.max_depth = 10;
  c.background[0] = c.background[1] = c.background[2] = 0.0;
  c.vfov = 90.0;
  c.lookfrom[0] = 0; c.lookfrom[1] = 0; c.lookfrom[2] = 0;
  c.lookat[0] = 0; c.lookat[1] = 0; c.lookat[2] = -1;
  c.vup[0] = 0; c.vup[1] = 1; c.vup[2] = 0;
  c.defocus_angle = 0.0; c.focus_dist = 10.0;
  return c;
}
static void set3(double* d, double x, double y, double z) { d[0] = x; d[1] = y; d[2] = z; }

// apply_overrides: the_rest_of_your_life/mod.rs:32-67; book 1's copy has no background (:23-55).
static void apply_overrides(rrt_camera& cam, const rrth_overrides* o, bool has_background_field) {
  if (!o) return;
  if (o->has & RRTH_O_ASPECT_RATIO) cam.aspect_ratio = o->aspect_ratio;
  if (o->has & RRTH_O_IMAGE_WIDTH) cam.image_width = o->image_width;
  if (o->has & RRTH_O_SAMPLES_PER_PIXEL) cam.samples_per_pixel = o->samples_per_pixel;
  if (o->has & RRTH_O_MAX_DEPTH) cam.max_depth = o->max_depth;
  if (o->has & RRTH_O_VFOV) cam.vfov = o->vfov;
  if (o->has & RRTH_O_LOOKFROM) set3(cam.lookfrom, o->lookfrom[0], o->lookfrom[1], o->lookfrom[2]);
  if (o->has & RRTH_O_LOOKAT) set3(cam.lookat, o->lookat[0], o->lookat[1], o->lookat[2]);
  if (o->has & RRTH_O_VUP) set3(cam.vup, o->vup[0], o->vup[1], o->vup[2]);
  if (o->has & RRTH_O_DEFOCUS_ANGLE) cam.defocus_angle = o->defocus_angle;
  if (o->has & RRTH_O_FOCUS_DIST) cam.focus_dist = o->focus_dist;
  if (has_background_field && (o->has & RRTH_O_BACKGROUND))
    set3(cam.background, o->background[0], o->background[1], o->background[2]);
}

// The random small-sphere grid shared by in_one_weekend/mod.rs:67-98 (moving=false) and
// bouncing_spheres, the_next_week/mod.rs:98-123 (moving=true).
static void random_sphere_grid(SceneBuilder& b, std::vector<HittableRef>& world, bool moving) {
  for (int a = -11; a < 11; ++a) {
    for (int bb = -11; bb < 11; ++bb) {
      double choose_mat = b.random_double();
      double cx = a + 0.9 * b.random_double();
      double cz = bb + 0.9 * b.random_double();
      V3 center(cx, 0.2, cz);
      if (length(center - V3(4.0, 0.2, 0.0)) > 0.9) {
        if (choose_mat < 0.8) {
          V3 c1 = b.random_vec();
          V3 c2 = b.random_vec();
          MaterialRef m = b.lambertian(c1 * c2);
          if (moving) {
            double dy = b.random_double() * 0.5;
            world.push_back(b.sphere_moving(center, center + V3(0.0, dy, 0.0), 0.2, m));
          } else {
            world.push_back(b.sphere(center, 0.2, m));
          }
        } else if (choose_mat < 0.95) {
          V3 albedo = b.random_vec_range(0.5, 1.0);
          double fuzz = b.random_double() * 0.5;
          world.push_back(b.sphere(center, 0.2, b.metal(albedo, fuzz)));
        } else {
          world.push_back(b.sphere(center, 0.2, b.dielectric(1.5)));
        }
      }
    }
  }
}
static void three_big_spheres(SceneBuilder& b, std::vector<HittableRef>& world) {
  world.push_back(b.sphere(V3(0.0, 1.0, 0.0), 1.0, b.dielectric(1.5)));
  world.push_back(b.sphere(V3(-4.0, 1.0, 0.0), 1.0, b.lambertian(V3(0.4, 0.2, 0.1))));
  world.push_back(b.sphere(V3(4.0, 1.0, 0.0), 1.0, b.metal(V3(0.7, 0.6, 0.5), 0.0)));
}

// ---------------------------------------------------------------- book 1
// in_one_weekend::run, mod.rs:57-126
static void book1(Scene& s, const rrth_overrides* ov) {
  SceneBuilder& b = s.b;
  std::vector<HittableRef> world;
  world.push_back(b.sphere(V3(0.0, -1000.0, 0.0), 1000.0, b.lambertian(V3(0.5, 0.5, 0.5))));
  random_sphere_grid(b, world, false);
  three_big_spheres(b, world);
  rrt_camera& cam = s.cam;
  cam = camera_default();
  cam.aspect_ratio = 16.0 / 9.0; cam.image_width = 1200; cam.samples_per_pixel = 10; cam.max_depth = 20;
  cam.vfov = 20.0; set3(cam.lookfrom, 13, 2, 3); set3(cam.lookat, 0, 0, 0); set3(cam.vup, 0, 1, 0);
  cam.defocus_angle = 0.6; cam.focus_dist = 10.0;
  apply_overrides(cam, ov, false);
  int root = b.bvh(world);  // BvhNode::new(world), :124
  b.fill_desc(s.desc, root, -1);
  s.integrator = RRT_INT_BOOK1;
}

// build_in_one_weekend_scene, src/gpu/mod.rs:124-301: the only seeded scene of the reference.
// f32 draws via gen::<f32>() / gen_range(0.5..1.0) (f32), f64 centres via gen::<f64>().  The POD
// it fills is f32; albedo/fuzz are therefore f32 values widened to the f64 IR.  A background
// override switches it to solid-colour mode (:169-172), i.e. the book-2 integrator.
static void book1_gpu(Scene& s, const rrth_overrides* ov) {
  SceneBuilder& b = s.b;  // b.rng was seeded by the caller (0x5EED1234 reproduces the reference)
  std::vector<HittableRef> world;
  world.push_back(b.sphere(V3(0.0, -1000.0, 0.0), 1000.0, b.lambertian(V3(0.5, 0.5, 0.5))));
  for (int a = -11; a < 11; ++a) {
    for (int bb = -11; bb < 11; ++bb) {
      float choose_mat = b.rng.gen_f32();
      double cx = a + 0.9 * b.rng.gen_f64();
      double cz = bb + 0.9 * b.rng.gen_f64();
      V3 center(cx, 0.2, cz);
      // the reference narrows the centre to f32 when it fills SphereGpu (:225)
      V3 cf((double)(float)center.x, (double)(float)center.y, (double)(float)center.z);
      if (length(center - V3(4.0, 0.2, 0.0)) > 0.9) {
        if (choose_mat < 0.8f) {
          float al[3];
          for (int k = 0; k < 3; ++k) { float p = b.rng.gen_f32(); float q = b.rng.gen_f32(); al[k] = p * q; }
          world.push_back(b.sphere(cf, (double)0.2f, b.lambertian(V3(al[0], al[1], al[2]))));
        } else if (choose_mat < 0.95f) {
          float al[3];
          for (int k = 0; k < 3; ++k) al[k] = b.rng.gen_range_f32(0.5f, 1.0f);
          float fuzz = b.rng.gen_f32() * 0.5f;
          world.push_back(b.sphere(cf, (double)0.2f, b.metal(V3(al[0], al[1], al[2]), fuzz)));
        } else {
          world.push_back(b.sphere(cf, (double)0.2f, b.dielectric(1.5)));
        }
      }
    }
  }
  world.push_back(b.sphere(V3(0.0, 1.0, 0.0), 1.0, b.dielectric(1.5)));
  world.push_back(b.sphere(V3(-4.0, 1.0, 0.0), 1.0, b.lambertian(V3((float)0.4f, (float)0.2f, (float)0.1f))));
  world.push_back(b.sphere(V3(4.0, 1.0, 0.0), 1.0, b.metal(V3((float)0.7f, (float)0.6f, (float)0.5f), 0.0)));
  rrt_camera& cam = s.cam;
  cam = camera_default();
  cam.aspect_ratio = 16.0 / 9.0; cam.image_width = 1200; cam.samples_per_pixel = 10; cam.max_depth = 20;
  cam.vfov = 20.0; set3(cam.lookfrom, 13, 2, 3); set3(cam.lookat, 0, 0, 0); set3(cam.vup, 0, 1, 0);
  cam.defocus_angle = 0.6; cam.focus_dist = 10.0;
  apply_overrides(cam, ov, true);
  int root = b.bvh(world);
  b.fill_desc(s.desc, root, -1);
  s.integrator = (ov && (ov->has & RRTH_O_BACKGROUND)) ? RRT_INT_BOOK2 : RRT_INT_BOOK1;
}

// ---------------------------------------------------------------- book 2
static void cam16x9_book2(rrt_camera& cam) {  // the block shared by scenes 1-4, mod.rs:137-150
  cam = camera_default();
  cam.aspect_ratio = 16.0 / 9.0; cam.image_width = 400; cam.samples_per_pixel = 100; cam.max_depth = 50;
  set3(cam.background, 0.70, 0.80, 1.00);
  cam.vfov = 20.0; set3(cam.lookfrom, 13, 2, 3); set3(cam.lookat, 0, 0, 0); set3(cam.vup, 0, 1, 0);
  cam.defocus_angle = 0.0;
}

static void bouncing_spheres(Scene& s, const rrth_overrides* ov) {  // mod.rs:83-155
  SceneBuilder& b = s.b;
  std::vector<HittableRef> world;
  TextureRef chk = b.checker_from_colors(0.32, V3(0.2, 0.3, 0.1), V3(0.9, 0.9, 0.9));
  world.push_back(b.sphere(V3(0.0, -1000.0, 0.0), 1000.0, b.lambertian_tex(chk)));
  random_sphere_grid(b, world, true);
  three_big_spheres(b, world);
  int root = b.bvh(world);  // :135
  cam16x9_book2(s.cam);
  s.cam.defocus_angle = 0.6; s.cam.focus_dist = 10.0;
  apply_overrides(s.cam, ov, true);
  b.fill_desc(s.desc, root, -1);
}

static void checkered_spheres(Scene& s, const rrth_overrides* ov) {  // mod.rs:157-194
  SceneBuilder& b = s.b;
  TextureRef chk = b.checker_from_colors(0.32, V3(0.2, 0.3, 0.1), V3(0.9, 0.9, 0.9));
  std::vector<HittableRef> world;
  world.push_back(b.sphere(V3(0.0, -10.0, 0.0), 10.0, b.lambertian_tex(chk)));
  world.push_back(b.sphere(V3(0.0, 10.0, 0.0), 10.0, b.lambertian_tex(chk)));
  cam16x9_book2(s.cam);
  apply_overrides(s.cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

static void earth(Scene& s, const rrth_overrides* ov) {  // mod.rs:196-220
  SceneBuilder& b = s.b;
  TextureRef tex = b.image_texture("earthmap.jpg");
  std::vector<HittableRef> world;
  world.push_back(b.sphere(V3(0, 0, 0), 2.0, b.lambertian_tex(tex)));
  cam16x9_book2(s.cam);
  set3(s.cam.lookfrom, 0, 0, 12);
  apply_overrides(s.cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

static void perlin_spheres(Scene& s, const rrth_overrides* ov) {  // mod.rs:222-254
  SceneBuilder& b = s.b;
  TextureRef pertext = b.noise_texture(4.0);
  std::vector<HittableRef> world;
  world.push_back(b.sphere(V3(0.0, -1000.0, 0.0), 1000.0, b.lambertian_tex(pertext)));
  world.push_back(b.sphere(V3(0.0, 2.0, 0.0), 2.0, b.lambertian_tex(pertext)));
  cam16x9_book2(s.cam);
  apply_overrides(s.cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

static void quads(Scene& s, const rrth_overrides* ov) {  // mod.rs:256-314
  SceneBuilder& b = s.b;
  MaterialRef left_red = b.lambertian(V3(1.0, 0.2, 0.2));
  MaterialRef back_green = b.lambertian(V3(0.2, 1.0, 0.2));
  MaterialRef right_blue = b.lambertian(V3(0.2, 0.2, 1.0));
  MaterialRef upper_orange = b.lambertian(V3(1.0, 0.5, 0.0));
  MaterialRef lower_teal = b.lambertian(V3(0.2, 0.8, 0.8));
  std::vector<HittableRef> world;
  world.push_back(b.quad(V3(-3, -2, 5), V3(0, 0, -4), V3(0, 4, 0), left_red));
  world.push_back(b.quad(V3(-2, -2, 0), V3(4, 0, 0), V3(0, 4, 0), back_green));
  world.push_back(b.quad(V3(3, -2, 1), V3(0, 0, 4), V3(0, 4, 0), right_blue));
  world.push_back(b.quad(V3(-2, 3, 1), V3(4, 0, 0), V3(0, 0, 4), upper_orange));
  world.push_back(b.quad(V3(-2, -3, 5), V3(4, 0, 0), V3(0, 0, -4), lower_teal));
  rrt_camera& cam = s.cam;
  cam = camera_default();
  cam.aspect_ratio = 1.0; cam.image_width = 400; cam.samples_per_pixel = 100; cam.max_depth = 50;
  set3(cam.background, 0.70, 0.80, 1.00);
  cam.vfov = 80.0; set3(cam.lookfrom, 0, 0, 9); set3(cam.lookat, 0, 0, 0); set3(cam.vup, 0, 1, 0);
  cam.defocus_angle = 0.0;
  apply_overrides(cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

static void simple_light(Scene& s, const rrth_overrides* ov) {  // mod.rs:316-361
  SceneBuilder& b = s.b;
  TextureRef pertext = b.noise_texture(4.0);
  std::vector<HittableRef> world;
  world.push_back(b.sphere(V3(0.0, -1000.0, 0.0), 1000.0, b.lambertian_tex(pertext)));
  world.push_back(b.sphere(V3(0.0, 2.0, 0.0), 2.0, b.lambertian_tex(pertext)));
  MaterialRef difflight = b.diffuse_light(V3(4, 4, 4));
  world.push_back(b.sphere(V3(0.0, 7.0, 0.0), 2.0, difflight));
  world.push_back(b.quad(V3(3, 1, -2), V3(2, 0, 0), V3(0, 2, 0), difflight));
  rrt_camera& cam = s.cam;
  cam = camera_default();
  cam.aspect_ratio = 16.0 / 9.0; cam.image_width = 400; cam.samples_per_pixel = 100; cam.max_depth = 50;
  set3(cam.background, 0, 0, 0);
  cam.vfov = 20.0; set3(cam.lookfrom, 26, 3, 6); set3(cam.lookat, 0, 2, 0); set3(cam.vup, 0, 1, 0);
  cam.defocus_angle = 0.0;
  apply_overrides(cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

static void cornell_camera(rrt_camera& cam, int width, int spp, int depth) {
  cam = camera_default();
  cam.aspect_ratio = 1.0; cam.image_width = width; cam.samples_per_pixel = spp; cam.max_depth = depth;
  set3(cam.background, 0, 0, 0);
  cam.vfov = 40.0; set3(cam.lookfrom, 278, 278, -800); set3(cam.lookat, 278, 278, 0); set3(cam.vup, 0, 1, 0);
  cam.defocus_angle = 0.0;
}

static void cornell_box(Scene& s, const rrth_overrides* ov) {  // mod.rs:363-433
  SceneBuilder& b = s.b;
  MaterialRef red = b.lambertian(V3(0.65, 0.05, 0.05));
  MaterialRef white = b.lambertian(V3(0.73, 0.73, 0.73));
  MaterialRef green = b.lambertian(V3(0.12, 0.45, 0.15));
  MaterialRef light = b.diffuse_light(V3(15, 15, 15));
  std::vector<HittableRef> world;
  world.push_back(b.quad(V3(555, 0, 0), V3(0, 555, 0), V3(0, 0, 555), green));
  world.push_back(b.quad(V3(0, 0, 0), V3(0, 555, 0), V3(0, 0, 555), red));
  world.push_back(b.quad(V3(343, 554, 332), V3(-130, 0, 0), V3(0, 0, -105), light));
  world.push_back(b.quad(V3(0, 0, 0), V3(555, 0, 0), V3(0, 0, 555), white));
  world.push_back(b.quad(V3(555, 555, 555), V3(-555, 0, 0), V3(0, 0, -555), white));
  world.push_back(b.quad(V3(0, 0, 555), V3(555, 0, 0), V3(0, 555, 0), white));
  HittableRef box1 = b.make_box(V3(0, 0, 0), V3(165, 330, 165), white);
  box1 = b.rotate_y(box1, 15.0);
  box1 = b.translate(box1, V3(265, 0, 295));
  world.push_back(box1);
  HittableRef box2 = b.make_box(V3(0, 0, 0), V3(165, 165, 165), white);
  box2 = b.rotate_y(box2, -18.0);
  box2 = b.translate(box2, V3(130, 0, 65));
  world.push_back(box2);
  cornell_camera(s.cam, 600, 200, 50);
  apply_overrides(s.cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

static void cornell_smoke(Scene& s, const rrth_overrides* ov) {  // mod.rs:435-505
  SceneBuilder& b = s.b;
  MaterialRef red = b.lambertian(V3(0.65, 0.05, 0.05));
  MaterialRef white = b.lambertian(V3(0.73, 0.73, 0.73));
  MaterialRef green = b.lambertian(V3(0.12, 0.45, 0.15));
  MaterialRef light = b.diffuse_light(V3(7, 7, 7));
  std::vector<HittableRef> world;
  world.push_back(b.quad(V3(555, 0, 0), V3(0, 555, 0), V3(0, 0, 555), green));
  world.push_back(b.quad(V3(0, 0, 0), V3(0, 555, 0), V3(0, 0, 555), red));
  world.push_back(b.quad(V3(113, 554, 127), V3(330, 0, 0), V3(0, 0, 305), light));
  world.push_back(b.quad(V3(0, 555, 0), V3(555, 0, 0), V3(0, 0, 555), white));
  world.push_back(b.quad(V3(0, 0, 0), V3(555, 0, 0), V3(0, 0, 555), white));
  world.push_back(b.quad(V3(0, 0, 555), V3(555, 0, 0), V3(0, 555, 0), white));
  HittableRef box1 = b.make_box(V3(0, 0, 0), V3(165, 330, 165), white);
  box1 = b.rotate_y(box1, 15.0);
  box1 = b.translate(box1, V3(265, 0, 295));
  HittableRef box2 = b.make_box(V3(0, 0, 0), V3(165, 165, 165), white);
  box2 = b.rotate_y(box2, -18.0);
  box2 = b.translate(box2, V3(130, 0, 65));
  world.push_back(b.constant_medium(box1, 0.01, V3(0, 0, 0)));
  world.push_back(b.constant_medium(box2, 0.01, V3(1, 1, 1)));
  cornell_camera(s.cam, 600, 200, 50);
  apply_overrides(s.cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

static void final_scene(Scene& s, const rrth_overrides* ov, int image_width, int spp, int max_depth) {
  // mod.rs:507-624
  SceneBuilder& b = s.b;
  std::vector<HittableRef> boxes1;
  MaterialRef ground = b.lambertian(V3(0.48, 0.83, 0.53));
  const int boxes_per_side = 20;
  for (int i = 0; i < boxes_per_side; ++i) {
    for (int j = 0; j < boxes_per_side; ++j) {
      double w = 100.0;
      double x0 = -1000.0 + i * w, z0 = -1000.0 + j * w, y0 = 0.0;
      double x1 = x0 + w;
      double y1 = b.random_double() * 100.0 + 1.0;
      double z1 = z0 + w;
      boxes1.push_back(b.make_box(V3(x0, y0, z0), V3(x1, y1, z1), ground));
    }
  }
  std::vector<HittableRef> world;
  world.push_back(b.bvh(boxes1));
  world.push_back(b.quad(V3(123, 554, 147), V3(300, 0, 0), V3(0, 0, 265), b.diffuse_light(V3(7, 7, 7))));
  V3 center1(400, 400, 200);
  V3 center2 = center1 + V3(30, 0, 0);
  world.push_back(b.sphere_moving(center1, center2, 50.0, b.lambertian(V3(0.7, 0.3, 0.1))));
  world.push_back(b.sphere(V3(260, 150, 45), 50.0, b.dielectric(1.5)));
  world.push_back(b.sphere(V3(0, 150, 145), 50.0, b.metal(V3(0.8, 0.8, 0.9), 1.0)));
  HittableRef boundary = b.sphere(V3(360, 150, 145), 70.0, b.dielectric(1.5));
  world.push_back(boundary);
  world.push_back(b.constant_medium(boundary, 0.2, V3(0.2, 0.4, 0.9)));
  HittableRef boundary2 = b.sphere(V3(0, 0, 0), 5000.0, b.dielectric(1.5));
  world.push_back(b.constant_medium(boundary2, 0.0001, V3(1, 1, 1)));
  MaterialRef emat = b.lambertian_tex(b.image_texture("earthmap.jpg"));
  world.push_back(b.sphere(V3(400, 200, 400), 100.0, emat));
  TextureRef pertext = b.noise_texture(0.2);
  world.push_back(b.sphere(V3(220, 280, 300), 80.0, b.lambertian_tex(pertext)));
  std::vector<HittableRef> boxes2;
  MaterialRef white = b.lambertian(V3(0.73, 0.73, 0.73));
  for (int j = 0; j < 1000; ++j) {
    V3 c = b.random_vec_range(0.0, 165.0);
    boxes2.push_back(b.sphere(c, 10.0, white));
  }
  HittableRef bx = b.bvh(boxes2);
  bx = b.rotate_y(bx, 15.0);
  bx = b.translate(bx, V3(-100, 270, 395));
  world.push_back(bx);
  rrt_camera& cam = s.cam;
  cam = camera_default();
  cam.aspect_ratio = 1.0; cam.image_width = image_width; cam.samples_per_pixel = spp; cam.max_depth = max_depth;
  set3(cam.background, 0, 0, 0);
  cam.vfov = 40.0; set3(cam.lookfrom, 478, 278, -600); set3(cam.lookat, 278, 278, 0); set3(cam.vup, 0, 1, 0);
  cam.defocus_angle = 0.0;
  apply_overrides(cam, ov, true);
  b.fill_desc(s.desc, b.list(world), -1);
}

// the_next_week::run, mod.rs:68-81
static void book2(Scene& s, int scene_no, const rrth_overrides* ov) {
  s.integrator = RRT_INT_BOOK2;
  switch (scene_no < 0 ? 0 : scene_no) {
    case 1: bouncing_spheres(s, ov); break;
    case 2: checkered_spheres(s, ov); break;
    case 3: earth(s, ov); break;
    case 4: perlin_spheres(s, ov); break;
    case 5: quads(s, ov); break;
    case 6: simple_light(s, ov); break;
    case 7: cornell_box(s, ov); break;
    case 8: cornell_smoke(s, ov); break;
    case 9: final_scene(s, ov, 800, 10000, 40); break;
    default: final_scene(s, ov, 400, 250, 4); break;
  }
}

// ---------------------------------------------------------------- book 3
// the_rest_of_your_life::run, mod.rs:69-161
static void book3(Scene& s, const rrth_overrides* ov) {
  SceneBuilder& b = s.b;
  MaterialRef red = b.lambertian(V3(0.65, 0.05, 0.05));
  MaterialRef white = b.lambertian(V3(0.73, 0.73, 0.73));
  MaterialRef green = b.lambertian(V3(0.12, 0.45, 0.15));
  MaterialRef light = b.diffuse_light(V3(15, 15, 15));
  std::vector<HittableRef> world;
  world.push_back(b.quad(V3(555, 0, 0), V3(0, 0, 555), V3(0, 555, 0), green));
  world.push_back(b.quad(V3(0, 0, 555), V3(0, 0, -555), V3(0, 555, 0), red));
  world.push_back(b.quad(V3(0, 555, 0), V3(555, 0, 0), V3(0, 0, 555), white));
  world.push_back(b.quad(V3(0, 0, 555), V3(555, 0, 0), V3(0, 0, -555), white));
  world.push_back(b.quad(V3(555, 0, 555), V3(-555, 0, 0), V3(0, 555, 0), white));
  world.push_back(b.quad(V3(213, 554, 227), V3(130, 0, 0), V3(0, 0, 105), light));
  HittableRef box1 = b.make_box(V3(0, 0, 0), V3(165, 330, 165), white);
  box1 = b.rotate_y(box1, 15.0);
  box1 = b.translate(box1, V3(265, 0, 295));
  world.push_back(box1);
  MaterialRef glass = b.dielectric(1.5);
  world.push_back(b.sphere(V3(190, 90, 190), 90.0, glass));
  MaterialRef empty = b.empty_material();
  std::vector<HittableRef> lights;
  lights.push_back(b.quad(V3(343, 554, 332), V3(-130, 0, 0), V3(0, 0, -105), empty));
  lights.push_back(b.sphere(V3(190, 90, 190), 90.0, empty));
  cornell_camera(s.cam, 600, 100, 50);
  apply_overrides(s.cam, ov, true);
  int root = b.bvh(world);  // :159
  int lroot = b.list(lights);
  b.fill_desc(s.desc, root, lroot);
  s.integrator = RRT_INT_BOOK3;
}

static thread_local std::string g_err;

}  // namespace rrt_host

using namespace rrt_host;

struct rrth_scene { Scene s; explicit rrth_scene(uint64_t seed) : s(seed) {} };

extern "C" {

void rrth_overrides_none(rrth_overrides* o) { std::memset(o, 0, sizeof *o); }

void rrth_overrides_shipped(rrth_overrides* o) {  // src/config.rs:50-62
  std::memset(o, 0, sizeof *o);
  o->has = RRTH_O_IMAGE_WIDTH | RRTH_O_SAMPLES_PER_PIXEL | RRTH_O_MAX_DEPTH;
  o->image_width = 2160; o->samples_per_pixel = 5000; o->max_depth = 100;
}

int rrth_normalize_book(const char* name) {  // src/main.rs:7-12 + :87-92
  std::string key;
  for (const char* p = name ? name : ""; *p; ++p) {
    unsigned char c = (unsigned char)*p;
    if (c < 128 && std::isalnum(c)) key.push_back((char)std::tolower(c));
  }
  if (key == "inoneweekend" || key == "oneweekend" || key == "weekend") return RRTH_BOOK_1;
  if (key == "thenextweek" || key == "nextweek" || key == "next") return RRTH_BOOK_2;
  if (key == "therestofyourlife" || key == "restofyourlife" || key == "rest" || key == "restoflife")
    return RRTH_BOOK_3;
  if (key == "inoneweekendgpu" || key == "weekendgpu") return RRTH_BOOK_1_GPU;
  return RRTH_BOOK_UNKNOWN;
}

int rrth_scene_build(int book, int scene_no, uint64_t seed, const rrth_overrides* ov,
                     const char* image_dir, rrth_scene** out) {
  if (!out) { g_err = "rrth_scene_build: out is null"; return -1; }
  *out = nullptr;
  rrth_scene* sc = new rrth_scene(seed);
  if (image_dir) sc->s.b.image_dir = image_dir;
  switch (book) {
    case RRTH_BOOK_1: book1(sc->s, ov); break;
    case RRTH_BOOK_2: book2(sc->s, scene_no, ov); break;
    case RRTH_BOOK_3: book3(sc->s, ov); break;
    case RRTH_BOOK_1_GPU: book1_gpu(sc->s, ov); break;
    default:
      delete sc;
      g_err = "unknown book (books: in_one_weekend, the_next_week, the_rest_of_your_life)";
      return -1;
  }
  *out = sc;
  return 0;
}
const rrt_scene_desc* rrth_scene_desc(const rrth_scene* s) { return &s->s.desc; }
const rrt_camera* rrth_scene_camera(const rrth_scene* s) { return &s->s.cam; }
int rrth_scene_integrator(const rrth_scene* s) { return s->s.integrator; }
void rrth_scene_free(rrth_scene* s) { delete s; }

uint64_t rrth_rng_kat_state1234(int n) {
  SmallRng r = SmallRng::from_state(1, 2, 3, 4);
  uint64_t v = 0;
  for (int i = 0; i <= n; ++i) v = r.next_u64();
  return v;
}
double rrth_rng_kat_seeded_f64(uint64_t seed, int n) {
  SmallRng r = SmallRng::seed_from_u64(seed);
  double v = 0;
  for (int i = 0; i <= n; ++i) v = r.random_double();
  return v;
}
const char* rrth_last_error(void) { return g_err.c_str(); }

}  // extern "C"
