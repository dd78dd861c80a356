// scene_builder.hpp — the host's stand-ins for make_ref / make_mat / make_tex and the object
// constructors of the reference (the_rest_of_your_life/{hittable,material,texture,quad,sphere}.rs).
// Each call appends one record to the flat arrays of an rrt_scene_desc, in construction order;
// the returned int plays the role of the Arc<…> the Rust code passes around.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/rrt.h"
#include "rng.hpp"

namespace rrt_host {

struct V3 {
  double x, y, z;
  V3() : x(0), y(0), z(0) {}
  V3(double a, double b, double c) : x(a), y(b), z(c) {}
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(double s, V3 a) { return V3(s * a.x, s * a.y, s * a.z); }
inline double length(V3 a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
inline V3 unit_vector(V3 a) { double l = length(a); return V3(a.x / l, a.y / l, a.z / l); }

typedef int HittableRef;
typedef int MaterialRef;
typedef int TextureRef;

// RtwImage::new's file search + load (rtw_image.rs:10-67).  Decoding: see rtw_image.cpp.
struct LoadedImage {
  int width = 0, height = 0;
  std::vector<uint8_t> rgb;
};
bool load_rtw_image(const std::string& filename, const char* extra_dir, LoadedImage& out);
bool decode_jpeg(const std::vector<uint8_t>& file, LoadedImage& out, std::string& err);  // jpeg_decode.cpp

struct SceneBuilder {
  std::vector<rrt_hittable> hittables;
  std::vector<int32_t> children;
  std::vector<rrt_material> materials;
  std::vector<rrt_texture> textures;
  std::vector<rrt_image> images;
  std::vector<std::unique_ptr<LoadedImage>> image_store;
  std::vector<rrt_perlin> perlins;
  SmallRng rng;
  std::string image_dir;

  explicit SceneBuilder(uint64_t seed) : rng(SmallRng::seed_from_u64(seed)) {}

  // ---- rtweekend.rs / vec3.rs random helpers (draw order matters for reproducibility) ----
  double random_double() { return rng.random_double(); }
  double random_double_range(double a, double b) { return rng.random_double_range(a, b); }
  V3 random_vec() { double a = random_double(), b = random_double(), c = random_double(); return V3(a, b, c); }
  V3 random_vec_range(double mn, double mx) {
    double a = random_double_range(mn, mx), b = random_double_range(mn, mx), c = random_double_range(mn, mx);
    return V3(a, b, c);
  }

  // ---- textures (texture.rs) ----
  TextureRef push_tex(const rrt_texture& t) { textures.push_back(t); return (int)textures.size() - 1; }
  static rrt_texture blank_tex(int kind) {
    rrt_texture t; std::memset(&t, 0, sizeof t);
    t.kind = kind; t.even = t.odd = t.image = t.perlin = -1; t.scale = 1.0;
    return t;
  }
  TextureRef solid_color(V3 c) {  // SolidColor::new, texture.rs:22-25
    rrt_texture t = blank_tex(RRT_T_SOLID);
    t.color[0] = c.x; t.color[1] = c.y; t.color[2] = c.z;
    return push_tex(t);
  }
  TextureRef checker(double scale, TextureRef even, TextureRef odd) {  // CheckerTexture::new :46-52
    rrt_texture t = blank_tex(RRT_T_CHECKER);
    t.scale = scale; t.even = even; t.odd = odd;
    return push_tex(t);
  }
  TextureRef checker_from_colors(double scale, V3 c1, V3 c2) {  // :54-60
    TextureRef e = solid_color(c1);
    TextureRef o = solid_color(c2);
    return checker(scale, e, o);
  }
  TextureRef image_texture(const char* filename) {  // ImageTexture::new :84-86
    std::unique_ptr<LoadedImage> img(new LoadedImage());
    load_rtw_image(filename, image_dir.empty() ? nullptr : image_dir.c_str(), *img);
    rrt_image ri; ri.width = img->width; ri.height = img->height;
    ri.rgb = img->rgb.empty() ? nullptr : img->rgb.data();
    image_store.push_back(std::move(img));
    images.push_back(ri);
    rrt_texture t = blank_tex(RRT_T_IMAGE);
    t.image = (int)images.size() - 1;
    return push_tex(t);
  }
  // Perlin::new, perlin.rs:12-23 + perlin_generate_perm/permute :66-77
  int perlin_new() {
    rrt_perlin p;
    for (int i = 0; i < 256; ++i) {
      V3 v = unit_vector(random_vec_range(-1.0, 1.0));
      p.randvec[i][0] = v.x; p.randvec[i][1] = v.y; p.randvec[i][2] = v.z;
    }
    int32_t* perms[3] = {p.perm_x, p.perm_y, p.perm_z};
    for (int k = 0; k < 3; ++k) {
      for (int i = 0; i < 256; ++i) perms[k][i] = i;
      for (int i = 255; i >= 1; --i) {
        int target = rng.random_int(0, i);
        std::swap(perms[k][i], perms[k][target]);
      }
    }
    perlins.push_back(p);
    return (int)perlins.size() - 1;
  }
  TextureRef noise_texture(double scale) {  // NoiseTexture::new, texture.rs:116-119
    rrt_texture t = blank_tex(RRT_T_NOISE);
    t.perlin = perlin_new();
    t.scale = scale;
    return push_tex(t);
  }

  // ---- materials (material.rs) ----
  MaterialRef push_mat(int kind, TextureRef tex, V3 albedo, double fuzz, double ri) {
    rrt_material m; std::memset(&m, 0, sizeof m);
    m.kind = kind; m.texture = tex;
    m.albedo[0] = albedo.x; m.albedo[1] = albedo.y; m.albedo[2] = albedo.z;
    m.fuzz = fuzz; m.refraction_index = ri;
    materials.push_back(m);
    return (int)materials.size() - 1;
  }
  MaterialRef lambertian(V3 albedo) { return push_mat(RRT_M_LAMBERTIAN, solid_color(albedo), albedo, 0, 1); }
  MaterialRef lambertian_tex(TextureRef t) { return push_mat(RRT_M_LAMBERTIAN, t, V3(), 0, 1); }
  MaterialRef metal(V3 albedo, double fuzz) {  // Metal::new :79-81
    return push_mat(RRT_M_METAL, -1, albedo, fuzz < 1.0 ? fuzz : 1.0, 1);
  }
  MaterialRef dielectric(double ri) { return push_mat(RRT_M_DIELECTRIC, -1, V3(1, 1, 1), 0, ri); }
  MaterialRef diffuse_light(V3 emit) { return push_mat(RRT_M_DIFFUSE_LIGHT, solid_color(emit), emit, 0, 1); }
  MaterialRef isotropic(V3 albedo) { return push_mat(RRT_M_ISOTROPIC, solid_color(albedo), albedo, 0, 1); }
  MaterialRef isotropic_tex(TextureRef t) { return push_mat(RRT_M_ISOTROPIC, t, V3(), 0, 1); }
  MaterialRef empty_material() { return push_mat(RRT_M_EMPTY, -1, V3(), 0, 1); }

  // ---- hittables ----
  HittableRef push_hit(int kind, int mat, int first, int n, V3 a, V3 b, V3 c) {
    rrt_hittable h; std::memset(&h, 0, sizeof h);
    h.kind = kind; h.material = mat; h.first_child = first; h.n_children = n;
    h.a[0] = a.x; h.a[1] = a.y; h.a[2] = a.z;
    h.b[0] = b.x; h.b[1] = b.y; h.b[2] = b.z;
    h.c[0] = c.x; h.c[1] = c.y; h.c[2] = c.z;
    hittables.push_back(h);
    return (int)hittables.size() - 1;
  }
  HittableRef sphere(V3 center, double radius, MaterialRef m) {  // Sphere::new, sphere.rs:18-27
    return push_hit(RRT_H_SPHERE, m, -1, 0, center, V3(), V3(radius, 0, 0));
  }
  HittableRef sphere_moving(V3 c1, V3 c2, double radius, MaterialRef m) {  // new_moving :29-45
    return push_hit(RRT_H_SPHERE, m, -1, 0, c1, c2 - c1, V3(radius, 0, 0));
  }
  HittableRef quad(V3 q, V3 u, V3 v, MaterialRef m) {  // Quad::new, quad.rs:24-48
    return push_hit(RRT_H_QUAD, m, -1, 0, q, u, v);
  }
  HittableRef list(const std::vector<HittableRef>& objs, int kind = RRT_H_LIST) {
    int first = (int)children.size();
    for (HittableRef o : objs) children.push_back(o);
    return push_hit(kind, -1, first, (int)objs.size(), V3(), V3(), V3());
  }
  HittableRef bvh(const std::vector<HittableRef>& objs) { return list(objs, RRT_H_BVH); }  // BvhNode::new
  HittableRef translate(HittableRef o, V3 offset) { return push_hit(RRT_H_TRANSLATE, -1, o, 0, offset, V3(), V3()); }
  HittableRef rotate_y(HittableRef o, double angle) { return push_hit(RRT_H_ROTATE_Y, -1, o, 0, V3(angle, 0, 0), V3(), V3()); }
  HittableRef constant_medium(HittableRef boundary, double density, V3 albedo) {  // from_color :31-37
    MaterialRef phase = isotropic(albedo);
    return push_hit(RRT_H_CONSTANT_MEDIUM, phase, boundary, 0, V3(density, 0, 0), V3(), V3());
  }
  // make_box, quad.rs:112-134: six quads, fixed order front,right,back,left,top,bottom.
  HittableRef make_box(V3 a, V3 b, MaterialRef mat) {
    V3 mn(std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z));
    V3 mx(std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z));
    V3 dx(mx.x - mn.x, 0, 0), dy(0, mx.y - mn.y, 0), dz(0, 0, mx.z - mn.z);
    std::vector<HittableRef> sides;
    sides.push_back(quad(V3(mn.x, mn.y, mx.z), dx, dy, mat));
    sides.push_back(quad(V3(mx.x, mn.y, mx.z), -dz, dy, mat));
    sides.push_back(quad(V3(mx.x, mn.y, mn.z), -dx, dy, mat));
    sides.push_back(quad(V3(mn.x, mn.y, mn.z), dz, dy, mat));
    sides.push_back(quad(V3(mn.x, mx.y, mx.z), dx, -dz, mat));
    sides.push_back(quad(V3(mn.x, mn.y, mn.z), dx, dz, mat));
    return list(sides);
  }

  void fill_desc(rrt_scene_desc& d, int world, int lights) const {
    std::memset(&d, 0, sizeof d);
    d.hittables = hittables.data(); d.n_hittables = (int)hittables.size();
    d.children = children.data();   d.n_children = (int)children.size();
    d.materials = materials.data(); d.n_materials = (int)materials.size();
    d.textures = textures.data();   d.n_textures = (int)textures.size();
    d.images = images.data();       d.n_images = (int)images.size();
    d.perlins = perlins.data();     d.n_perlins = (int)perlins.size();
    d.world = world; d.lights = lights;
  }
};

}  // namespace rrt_host
