// orc_scene.hpp — ORACLE (test infrastructure only).  PARITY UNPINNED (see orc_math.hpp).
//
// CPU f64 restatement of the reference's scene graph, following the files line by line and
// keeping the recursive pointer tree (so it is structurally independent of the GPU's flat BVH):
//   Aabb            the_rest_of_your_life/aabb.rs:8-127
//   HitRecord/Hittable/Translate/RotateY   the_rest_of_your_life/hittable.rs:15-178
//   HittableList    the_rest_of_your_life/hittable_list.rs:8-76
//   BvhNode         the_rest_of_your_life/bvh.rs:10-177
//   Sphere          the_rest_of_your_life/sphere.rs:11-123
//   Quad            the_rest_of_your_life/quad.rs:11-107
//   ConstantMedium  the_rest_of_your_life/constant_medium.rs:10-84
//   Texture*        the_rest_of_your_life/texture.rs:8-168, perlin.rs:4-99, rtw_image.rs:46-78
// Departure #2: every leaf primitive carries `id` = its index in rrt_scene_desc.hittables.
#pragma once
#include <algorithm>
#include <memory>
#include <string>
#include <vector>

#include "../include/rrt.h"
#include "orc_math.hpp"

namespace orc {

// ---------------------------------------------------------------- Aabb
struct Aabb {
  Interval x, y, z;
  Aabb() {}  // EMPTY
  Aabb(Interval a, Interval b, Interval c) : x(a), y(b), z(c) { pad_to_minimums(); }  // :24-28
  static Aabb from_points(Point3 a, Point3 b) {  // :30-35
    Interval x = a[0] <= b[0] ? Interval(a[0], b[0]) : Interval(b[0], a[0]);
    Interval y = a[1] <= b[1] ? Interval(a[1], b[1]) : Interval(b[1], a[1]);
    Interval z = a[2] <= b[2] ? Interval(a[2], b[2]) : Interval(b[2], a[2]);
    return Aabb(x, y, z);
  }
  static Aabb from_boxes(const Aabb& b0, const Aabb& b1) {  // :37-43
    return Aabb(Interval::from_intervals(b0.x, b1.x), Interval::from_intervals(b0.y, b1.y),
                Interval::from_intervals(b0.z, b1.z));
  }
  const Interval& axis_interval(int n) const { return n == 1 ? y : (n == 2 ? z : x); }  // :45-50
  bool hit(const Ray& r, Interval ray_t, Ctx& c) const {  // :52-84
    c.n.aabb += 1;
    for (int axis = 0; axis < 3; ++axis) {
      const Interval& ax = axis_interval(axis);
      double adinv = 1.0 / r.dir[axis];
      double t0 = (ax.min - r.orig[axis]) * adinv;
      double t1 = (ax.max - r.orig[axis]) * adinv;
      if (t0 < t1) {
        if (t0 > ray_t.min) ray_t.min = t0;
        if (t1 < ray_t.max) ray_t.max = t1;
      } else {
        if (t1 > ray_t.min) ray_t.min = t1;
        if (t0 < ray_t.max) ray_t.max = t0;
      }
      if (ray_t.max <= ray_t.min) return false;
    }
    return true;
  }
  int longest_axis() const {  // :86-98
    if (x.size() > y.size()) return x.size() > z.size() ? 0 : 2;
    return y.size() > z.size() ? 1 : 2;
  }
  double surface_area() const {  // :100-105
    double a = x.size(), b = y.size(), c = z.size();
    return 2.0 * (a * b + a * c + b * c);
  }
  void pad_to_minimums() {  // :107-118
    double delta = 0.0001;
    if (x.size() < delta) x = x.expand(delta);
    if (y.size() < delta) y = y.expand(delta);
    if (z.size() < delta) z = z.expand(delta);
  }
  Aabb shifted(Vec3 o) const { return Aabb(x.shifted(o.x()), y.shifted(o.y()), z.shifted(o.z())); }  // :121-127
};

// ---------------------------------------------------------------- textures
struct Perlin {  // perlin.rs:4-9 (tables come from the scene description)
  Vec3 randvec[256];
  int perm_x[256], perm_y[256], perm_z[256];
  double noise(Point3 p) const {  // :26-49
    double u = p.x() - std::floor(p.x()), v = p.y() - std::floor(p.y()), w = p.z() - std::floor(p.z());
    int i = (int)std::floor(p.x()), j = (int)std::floor(p.y()), k = (int)std::floor(p.z());
    Vec3 c[2][2][2];
    for (int di = 0; di < 2; ++di)
      for (int dj = 0; dj < 2; ++dj)
        for (int dk = 0; dk < 2; ++dk)
          c[di][dj][dk] = randvec[perm_x[(i + di) & 255] ^ perm_y[(j + dj) & 255] ^ perm_z[(k + dk) & 255]];
    return interp(c, u, v, w);
  }
  double turb(Point3 p, int depth) const {  // :51-63
    double accum = 0.0, weight = 1.0;
    Point3 temp_p = p;
    for (int i = 0; i < depth; ++i) {
      accum += weight * noise(temp_p);
      weight *= 0.5;
      temp_p = temp_p * 2.0;
    }
    return std::fabs(accum);
  }
  static double interp(const Vec3 c[2][2][2], double u, double v, double w) {  // :79-98
    double uu = u * u * (3.0 - 2.0 * u), vv = v * v * (3.0 - 2.0 * v), ww = w * w * (3.0 - 2.0 * w);
    double accum = 0.0;
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 2; ++j)
        for (int k = 0; k < 2; ++k) {
          Vec3 weight_v(u - i, v - j, w - k);
          accum += (i * uu + (1.0 - i) * (1.0 - uu)) * (j * vv + (1.0 - j) * (1.0 - vv)) *
                   (k * ww + (1.0 - k) * (1.0 - ww)) * dot(c[i][j][k], weight_v);
        }
    return accum;
  }
};

struct Scene;  // fwd

struct Texture {
  int kind = RRT_T_SOLID;
  Color albedo;
  double inv_scale = 1.0, scale = 1.0;
  int even = -1, odd = -1;
  int img_w = 0, img_h = 0;
  const uint8_t* img = nullptr;
  const Perlin* perlin = nullptr;
};

// ---------------------------------------------------------------- materials (data only; behaviour in orc_shade.hpp)
struct Material {
  int kind = RRT_M_EMPTY;
  int tex = -1;
  Color albedo;
  double fuzz = 0, ri = 1;
};

// ---------------------------------------------------------------- hittables
struct HitRecord {  // hittable.rs:15-23 (+ id)
  Point3 p;
  Vec3 normal;
  int mat = -1;
  double t = 0, u = 0, v = 0;
  bool front_face = false;
  int id = -1;
  void set_face_normal(const Ray& r, Vec3 outward_normal) {  // :49-52
    front_face = dot(r.dir, outward_normal) < 0.0;
    normal = front_face ? outward_normal : -outward_normal;
  }
};

struct Hittable {
  virtual ~Hittable() {}
  virtual bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const = 0;
  virtual Aabb bounding_box() const = 0;
  virtual double pdf_value(Point3, Vec3, Ctx&) const { return 0.0; }           // hittable.rs:58-60
  virtual Vec3 random(Point3, Ctx&) const { return Vec3(1.0, 0.0, 0.0); }      // :62-64
  virtual bool is_medium() const { return false; }
};
typedef std::shared_ptr<Hittable> HittableRef;

struct Onb {  // onb.rs:3-35
  Vec3 axis[3];
  explicit Onb(Vec3 n) {
    Vec3 w = unit_vector(n);
    Vec3 a = std::fabs(w.x()) > 0.9 ? Vec3(0.0, 1.0, 0.0) : Vec3(1.0, 0.0, 0.0);
    Vec3 v = unit_vector(cross(w, a));
    Vec3 u = cross(w, v);
    axis[0] = u; axis[1] = v; axis[2] = w;
  }
  Vec3 w() const { return axis[2]; }
  Vec3 transform(Vec3 v) const { return v[0] * axis[0] + v[1] * axis[1] + v[2] * axis[2]; }
};

struct Sphere : Hittable {
  Ray center;
  double radius;
  int mat, id;
  Aabb bbox;
  Sphere(Point3 c0, Vec3 dc, double r, int m, int id_) : center(c0, dc), radius(std::fmax(r, 0.0)), mat(m), id(id_) {
    Vec3 rvec(r, r, r);  // sphere.rs:19-20, 35-38: the unclamped radius builds the box
    Aabb b1 = Aabb::from_points(c0 - rvec, c0 + rvec);
    if (dc.x() == 0.0 && dc.y() == 0.0 && dc.z() == 0.0) {
      bbox = b1;
    } else {
      Point3 c1 = c0 + dc;
      bbox = Aabb::from_boxes(b1, Aabb::from_points(c1 - rvec, c1 + rvec));
    }
  }
  static void get_sphere_uv(Point3 p, double& u, double& v) {  // :47-53
    double theta = std::acos(-p.y());
    double phi = std::atan2(-p.z(), p.x()) + PI;
    u = phi / (2.0 * PI);
    v = theta / PI;
  }
  bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const override {  // :69-96
    c.n.sph_test += 1;
    Point3 current_center = center.at(r.tm);
    Vec3 oc = current_center - r.orig;
    double a = r.dir.length_squared();
    double h = dot(r.dir, oc);
    double cc = oc.length_squared() - radius * radius;
    double discriminant = h * h - a * cc;
    if (discriminant < 0.0) return false;
    double sqrtd = std::sqrt(discriminant);
    double root = (h - sqrtd) / a;
    if (!ray_t.surrounds(root)) {
      root = (h + sqrtd) / a;
      if (!ray_t.surrounds(root)) return false;
    }
    c.n.sph_hit += 1;
    rec.t = root;
    rec.p = r.at(root);
    Vec3 outward_normal = (rec.p - current_center) / radius;
    get_sphere_uv(outward_normal, rec.u, rec.v);
    rec.mat = mat; rec.id = id;
    rec.set_face_normal(r, outward_normal);
    return true;
  }
  Aabb bounding_box() const override { return bbox; }
  double pdf_value(Point3 origin, Vec3 direction, Ctx& c) const override {  // :102-115
    HitRecord rec;
    if (!hit(Ray(origin, direction), Interval(0.001, INF), rec, c)) return 0.0;
    double dist_squared = (center.at(0.0) - origin).length_squared();
    double cos_theta_max = std::sqrt(1.0 - radius * radius / dist_squared);
    double solid_angle = 2.0 * PI * (1.0 - cos_theta_max);
    return 1.0 / solid_angle;
  }
  static Vec3 random_to_sphere(double radius, double distance_squared, Ctx& c) {  // :55-65
    double r1 = c.random_double(), r2 = c.random_double();
    double z = 1.0 + r2 * (std::sqrt(1.0 - radius * radius / distance_squared) - 1.0);
    double phi = 2.0 * PI * r1;
    double x = std::cos(phi) * std::sqrt(1.0 - z * z);
    double y = std::sin(phi) * std::sqrt(1.0 - z * z);
    return Vec3(x, y, z);
  }
  Vec3 random(Point3 origin, Ctx& c) const override {  // :117-122
    Vec3 direction = center.at(0.0) - origin;
    double distance_squared = direction.length_squared();
    Onb uvw(direction);
    return uvw.transform(random_to_sphere(radius, distance_squared, c));
  }
};

struct Quad : Hittable {
  Point3 q;
  Vec3 u, v, w, normal;
  double d, area;
  int mat, id;
  Aabb bbox;
  Quad(Point3 q_, Vec3 u_, Vec3 v_, int m, int id_) : q(q_), u(u_), v(v_), mat(m), id(id_) {  // quad.rs:24-48
    Vec3 n = cross(u, v);
    normal = unit_vector(n);
    d = dot(normal, q);
    w = n / dot(n, n);
    area = n.length();
    Aabb d1 = Aabb::from_points(q, q + u + v);  // :50-54
    Aabb d2 = Aabb::from_points(q + u, q + v);
    bbox = Aabb::from_boxes(d1, d2);
  }
  bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const override {  // :64-91
    c.n.quad_test += 1;
    double denom = dot(normal, r.dir);
    if (std::fabs(denom) < 1e-8) return false;
    double t = (d - dot(normal, r.orig)) / denom;
    if (!ray_t.contains(t)) return false;
    c.n.quad_full += 1;
    Point3 intersection = r.at(t);
    Vec3 planar_hitpt_vector = intersection - q;
    double alpha = dot(w, cross(planar_hitpt_vector, v));
    double beta = dot(w, cross(u, planar_hitpt_vector));
    Interval unit_interval(0.0, 1.0);  // is_interior :56-65
    if (!unit_interval.contains(alpha) || !unit_interval.contains(beta)) return false;
    rec.u = alpha; rec.v = beta;
    rec.t = t; rec.p = intersection; rec.mat = mat; rec.id = id;
    rec.set_face_normal(r, normal);
    return true;
  }
  Aabb bounding_box() const override { return bbox; }
  double pdf_value(Point3 origin, Vec3 direction, Ctx& c) const override {  // :93-102
    HitRecord rec;
    if (!hit(Ray(origin, direction), Interval(0.001, INF), rec, c)) return 0.0;
    double distance_squared = rec.t * rec.t * direction.length_squared();
    double cosine = std::fabs(dot(direction, rec.normal)) / direction.length();
    return distance_squared / (cosine * area);
  }
  Vec3 random(Point3 origin, Ctx& c) const override {  // :104-107
    double a = c.random_double(), b = c.random_double();
    Point3 p = q + (a * u) + (b * v);
    return p - origin;
  }
};

struct HittableList : Hittable {
  std::vector<HittableRef> objects;
  Aabb bbox;
  void add(HittableRef o) {  // hittable_list.rs:25-33
    objects.push_back(o);
    if (objects.size() == 1) bbox = o->bounding_box();
    else bbox = Aabb::from_boxes(bbox, o->bounding_box());
  }
  bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const override {  // :42-54
    bool hit_anything = false;
    double closest_so_far = ray_t.max;
    HitRecord tmp;
    for (const HittableRef& o : objects) {
      if (o->hit(r, Interval(ray_t.min, closest_so_far), tmp, c)) {
        closest_so_far = tmp.t;
        rec = tmp;
        hit_anything = true;
      }
    }
    return hit_anything;
  }
  Aabb bounding_box() const override { return bbox; }
  double pdf_value(Point3 origin, Vec3 direction, Ctx& c) const override {  // :60-69
    double weight = 1.0 / (double)objects.size();
    double sum = 0.0;
    for (const HittableRef& o : objects) sum += weight * o->pdf_value(origin, direction, c);
    return sum;
  }
  Vec3 random(Point3 origin, Ctx& c) const override {  // :71-75
    int int_size = (int)objects.size();
    int index = c.random_int(0, int_size - 1);
    return objects[index]->random(origin, c);
  }
};

struct Translate : Hittable {  // hittable.rs:67-98
  HittableRef object;
  Vec3 offset;
  Aabb bbox;
  Translate(HittableRef o, Vec3 off) : object(o), offset(off) { bbox = o->bounding_box().shifted(off); }
  bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const override {
    c.n.inst += 1;
    Ray offset_r(r.orig - offset, r.dir, r.tm);
    if (!object->hit(offset_r, ray_t, rec, c)) return false;
    rec.p = rec.p + offset;
    return true;
  }
  Aabb bounding_box() const override { return bbox; }
};

struct RotateY : Hittable {  // hittable.rs:100-178
  HittableRef object;
  double sin_theta, cos_theta;
  Aabb bbox;
  RotateY(HittableRef o, double angle) : object(o) {
    double radians = degrees_to_radians(angle);
    sin_theta = std::sin(radians);
    cos_theta = std::cos(radians);
    Aabb b = o->bounding_box();
    Point3 mn(INF, INF, INF), mx(-INF, -INF, -INF);
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 2; ++j)
        for (int k = 0; k < 2; ++k) {
          double x = i == 1 ? b.x.max : b.x.min, y = j == 1 ? b.y.max : b.y.min, z = k == 1 ? b.z.max : b.z.min;
          double newx = cos_theta * x + sin_theta * z;
          double newz = -sin_theta * x + cos_theta * z;
          Vec3 tester(newx, y, newz);
          for (int cc = 0; cc < 3; ++cc) { mn[cc] = std::fmin(mn[cc], tester[cc]); mx[cc] = std::fmax(mx[cc], tester[cc]); }
        }
    bbox = Aabb::from_points(mn, mx);
  }
  bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const override {  // :142-173
    c.n.inst += 1;
    Point3 origin((cos_theta * r.orig.x()) - (sin_theta * r.orig.z()), r.orig.y(),
                  (sin_theta * r.orig.x()) + (cos_theta * r.orig.z()));
    Vec3 direction((cos_theta * r.dir.x()) - (sin_theta * r.dir.z()), r.dir.y(),
                   (sin_theta * r.dir.x()) + (cos_theta * r.dir.z()));
    Ray rotated_r(origin, direction, r.tm);
    if (!object->hit(rotated_r, ray_t, rec, c)) return false;
    rec.p = Point3((cos_theta * rec.p.x()) + (sin_theta * rec.p.z()), rec.p.y(),
                   (-sin_theta * rec.p.x()) + (cos_theta * rec.p.z()));
    rec.normal = Vec3((cos_theta * rec.normal.x()) + (sin_theta * rec.normal.z()), rec.normal.y(),
                      (-sin_theta * rec.normal.x()) + (cos_theta * rec.normal.z()));
    return true;
  }
  Aabb bounding_box() const override { return bbox; }
};

struct ConstantMedium : Hittable {  // constant_medium.rs:10-84
  HittableRef boundary;
  double neg_inv_density;
  int phase_mat, id;
  bool probe_skip;  // departure #3: excluded in probe mode (set on the Ctx-free scene flag)
  const bool* probe_mode;
  ConstantMedium(HittableRef b, double density, int mat, int id_, const bool* pm)
      : boundary(b), neg_inv_density(-1.0 / density), phase_mat(mat), id(id_), probe_skip(false), probe_mode(pm) {}
  bool is_medium() const override { return true; }
  bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const override {
    if (*probe_mode) return false;
    c.n.medium += 1;
    HitRecord rec1, rec2;
    if (!boundary->hit(r, UNIVERSE, rec1, c)) return false;
    if (!boundary->hit(r, Interval(rec1.t + 0.0001, INF), rec2, c)) return false;
    if (rec1.t < ray_t.min) rec1.t = ray_t.min;
    if (rec2.t > ray_t.max) rec2.t = ray_t.max;
    if (rec1.t >= rec2.t) return false;
    if (rec1.t < 0.0) rec1.t = 0.0;
    double ray_length = r.dir.length();
    double distance_inside_boundary = (rec2.t - rec1.t) * ray_length;
    double hit_distance = neg_inv_density * std::log(c.random_double());
    if (hit_distance > distance_inside_boundary) return false;
    rec.t = rec1.t + hit_distance / ray_length;
    rec.p = r.at(rec.t);
    rec.normal = Vec3(1.0, 0.0, 0.0);
    rec.mat = phase_mat;
    rec.u = 0.0; rec.v = 0.0;
    rec.front_face = true;
    rec.id = id;
    return true;
  }
  Aabb bounding_box() const override { return boundary->bounding_box(); }
};

struct BvhNode : Hittable {  // bvh.rs
  HittableRef left, right;
  Aabb bbox;
  explicit BvhNode(std::vector<HittableRef>& objects) { build(objects.data(), objects.size()); }
  BvhNode(HittableRef* objects, size_t span) { build(objects, span); }
  static HittableRef make(HittableRef* objects, size_t span) { return HittableRef(new BvhNode(objects, span)); }
  void build(HittableRef* objects, size_t object_span) {  // :21-156
    bbox = Aabb();
    for (size_t i = 0; i < object_span; ++i) bbox = Aabb::from_boxes(bbox, objects[i]->bounding_box());
    if (object_span == 1) { left = objects[0]; right = objects[0]; return; }
    if (object_span == 2) { left = objects[0]; right = objects[1]; return; }
    const size_t NUM_BUCKETS = 12;
    int axis = bbox.longest_axis();
    auto comparator = [axis](const HittableRef& a, const HittableRef& b) {
      return a->bounding_box().axis_interval(axis).min < b->bounding_box().axis_interval(axis).min;
    };
    auto median_split = [&]() {
      std::stable_sort(objects, objects + object_span, comparator);  // slice::sort_by is stable
      size_t mid = object_span / 2;
      left = make(objects, mid);
      right = make(objects + mid, object_span - mid);
    };
    double centroid_min = INF, centroid_max = -INF;
    for (size_t i = 0; i < object_span; ++i) {
      Aabb b = objects[i]->bounding_box();
      double centroid = 0.5 * (b.axis_interval(axis).min + b.axis_interval(axis).max);
      if (centroid < centroid_min) centroid_min = centroid;
      if (centroid > centroid_max) centroid_max = centroid;
    }
    if (std::fabs(centroid_max - centroid_min) < 1e-12) { median_split(); return; }
    struct Bucket { size_t count = 0; Aabb bbox; };
    Bucket buckets[NUM_BUCKETS];
    auto bucket_of = [&](const Aabb& b) {
      double centroid = 0.5 * (b.axis_interval(axis).min + b.axis_interval(axis).max);
      size_t idx = (size_t)((centroid - centroid_min) / (centroid_max - centroid_min) * (double)NUM_BUCKETS);
      if (idx >= NUM_BUCKETS) idx = NUM_BUCKETS - 1;
      return idx;
    };
    for (size_t i = 0; i < object_span; ++i) {
      Aabb b = objects[i]->bounding_box();
      size_t idx = bucket_of(b);
      buckets[idx].count += 1;
      buckets[idx].bbox = Aabb::from_boxes(buckets[idx].bbox, b);
    }
    Aabb right_bbox[NUM_BUCKETS];
    size_t right_count[NUM_BUCKETS];
    Aabb accum_bbox;
    size_t accum_count = 0;
    for (size_t ii = NUM_BUCKETS; ii-- > 0;) {
      accum_count += buckets[ii].count;
      accum_bbox = Aabb::from_boxes(accum_bbox, buckets[ii].bbox);
      right_bbox[ii] = accum_bbox;
      right_count[ii] = accum_count;
    }
    Aabb left_bbox;
    size_t left_count = 0;
    double best_cost = INF;
    size_t best_split = 0;
    for (size_t i = 0; i < NUM_BUCKETS - 1; ++i) {
      left_count += buckets[i].count;
      left_bbox = Aabb::from_boxes(left_bbox, buckets[i].bbox);
      if (left_count == 0 || right_count[i + 1] == 0) continue;
      double cost = left_bbox.surface_area() * (double)left_count + right_bbox[i + 1].surface_area() * (double)right_count[i + 1];
      if (cost < best_cost) { best_cost = cost; best_split = i; }
    }
    if (!std::isfinite(best_cost)) { median_split(); return; }
    size_t mid = 0;
    for (size_t i = 0; i < object_span; ++i) {
      if (bucket_of(objects[i]->bounding_box()) <= best_split) { std::swap(objects[i], objects[mid]); mid += 1; }
    }
    if (mid == 0 || mid == object_span) { median_split(); return; }
    left = make(objects, mid);
    right = make(objects + mid, object_span - mid);
  }
  bool hit(const Ray& r, Interval ray_t, HitRecord& rec, Ctx& c) const override {  // :159-172
    if (!bbox.hit(r, ray_t, c)) return false;
    bool hit_left = left->hit(r, ray_t, rec, c);
    HitRecord rr;
    bool hit_right = right->hit(r, Interval(ray_t.min, hit_left ? rec.t : ray_t.max), rr, c);
    if (hit_right) rec = rr;
    return hit_left || hit_right;
  }
  Aabb bounding_box() const override { return bbox; }
};

// ---------------------------------------------------------------- the scene, built from the description
struct Scene {
  std::vector<Material> materials;
  std::vector<Texture> textures;
  std::vector<Perlin> perlins;
  std::vector<std::vector<uint8_t>> image_store;  // RtwImage owns its pixels (rtw_image.rs:4-8): copied, not borrowed
  HittableRef world, lights;
  bool probe_mode = false;
  std::string error;

  HittableRef build_node(const rrt_scene_desc* d, int idx, std::vector<HittableRef>& memo, int depth) {
    if (idx < 0 || idx >= d->n_hittables || depth > 64) { error = "bad hittable index"; return nullptr; }
    if (memo[idx]) return memo[idx];  // shared Arc (e.g. the_next_week/mod.rs:558-568 boundary.clone())
    const rrt_hittable& h = d->hittables[idx];
    HittableRef out;
    switch (h.kind) {
      case RRT_H_SPHERE:
        out.reset(new Sphere(Point3(h.a[0], h.a[1], h.a[2]), Vec3(h.b[0], h.b[1], h.b[2]), h.c[0], h.material, idx));
        break;
      case RRT_H_QUAD:
        out.reset(new Quad(Point3(h.a[0], h.a[1], h.a[2]), Vec3(h.b[0], h.b[1], h.b[2]), Vec3(h.c[0], h.c[1], h.c[2]), h.material, idx));
        break;
      case RRT_H_LIST: case RRT_H_BVH: {
        std::vector<HittableRef> objs;
        for (int k = 0; k < h.n_children; ++k) {
          HittableRef ch = build_node(d, d->children[h.first_child + k], memo, depth + 1);
          if (!ch) return nullptr;
          objs.push_back(ch);
        }
        if (h.kind == RRT_H_LIST) {
          HittableList* l = new HittableList();
          for (auto& o : objs) l->add(o);
          out.reset(l);
        } else {
          if (objs.empty()) { error = "BvhNode::new on an empty list"; return nullptr; }
          out.reset(new BvhNode(objs));
        }
        break;
      }
      case RRT_H_TRANSLATE: {
        HittableRef ch = build_node(d, h.first_child, memo, depth + 1);
        if (!ch) return nullptr;
        out.reset(new Translate(ch, Vec3(h.a[0], h.a[1], h.a[2])));
        break;
      }
      case RRT_H_ROTATE_Y: {
        HittableRef ch = build_node(d, h.first_child, memo, depth + 1);
        if (!ch) return nullptr;
        out.reset(new RotateY(ch, h.a[0]));
        break;
      }
      case RRT_H_CONSTANT_MEDIUM: {
        HittableRef ch = build_node(d, h.first_child, memo, depth + 1);
        if (!ch) return nullptr;
        out.reset(new ConstantMedium(ch, h.a[0], h.material, idx, &probe_mode));
        break;
      }
      default: error = "unknown hittable kind"; return nullptr;
    }
    memo[idx] = out;
    return out;
  }

  bool build(const rrt_scene_desc* d) {
    perlins.resize(d->n_perlins);
    for (int i = 0; i < d->n_perlins; ++i) {
      for (int k = 0; k < 256; ++k) {
        perlins[i].randvec[k] = Vec3(d->perlins[i].randvec[k][0], d->perlins[i].randvec[k][1], d->perlins[i].randvec[k][2]);
        perlins[i].perm_x[k] = d->perlins[i].perm_x[k];
        perlins[i].perm_y[k] = d->perlins[i].perm_y[k];
        perlins[i].perm_z[k] = d->perlins[i].perm_z[k];
      }
    }
    image_store.resize(d->n_images > 0 ? d->n_images : 0);
    for (int i = 0; i < d->n_images; ++i)
      if (d->images[i].rgb && d->images[i].width > 0 && d->images[i].height > 0)
        image_store[i].assign(d->images[i].rgb, d->images[i].rgb + (size_t)d->images[i].width * d->images[i].height * 3);
    for (int i = 0; i < d->n_textures; ++i) {
      const rrt_texture& t = d->textures[i];
      Texture x;
      x.kind = t.kind; x.albedo = Color(t.color[0], t.color[1], t.color[2]);
      x.scale = t.scale; x.inv_scale = 1.0 / t.scale;  // texture.rs:51
      x.even = t.even; x.odd = t.odd;
      if (t.kind == RRT_T_IMAGE && t.image >= 0 && t.image < d->n_images && d->images[t.image].rgb) {
        x.img_w = d->images[t.image].width; x.img_h = d->images[t.image].height; x.img = image_store[t.image].data();
      }
      if (t.kind == RRT_T_NOISE) x.perlin = &perlins[t.perlin];
      textures.push_back(x);
    }
    for (int i = 0; i < d->n_materials; ++i) {
      const rrt_material& m = d->materials[i];
      Material x;
      x.kind = m.kind; x.tex = m.texture; x.albedo = Color(m.albedo[0], m.albedo[1], m.albedo[2]);
      x.fuzz = m.fuzz; x.ri = m.refraction_index;
      materials.push_back(x);
    }
    std::vector<HittableRef> memo(d->n_hittables);
    world = build_node(d, d->world, memo, 0);
    if (!world) return false;
    if (d->lights >= 0) {
      lights = build_node(d, d->lights, memo, 0);
      if (!lights) return false;
    }
    return true;
  }

  // TextureObject::value, texture.rs:33-37, 63-77, 88-109, 121-126
  Color texture_value(int ti, double u, double v, Point3 p, Ctx& c) const {
    const Texture& t = textures[ti];
    switch (t.kind) {
      case RRT_T_SOLID: return t.albedo;
      case RRT_T_CHECKER: {
        c.n.tex_checker += 1;
        int xi = (int)std::floor(t.inv_scale * p.x());
        int yi = (int)std::floor(t.inv_scale * p.y());
        int zi = (int)std::floor(t.inv_scale * p.z());
        bool is_even = (xi + yi + zi) % 2 == 0;  // C++ % truncates like Rust's
        return texture_value(is_even ? t.even : t.odd, u, v, p, c);
      }
      case RRT_T_IMAGE: {
        c.n.tex_image += 1;
        if (t.img_h <= 0) return Color(0.0, 1.0, 1.0);
        double uu = Interval(0.0, 1.0).clamp(u);
        double vv = 1.0 - Interval(0.0, 1.0).clamp(v);
        int i = (int)(uu * (double)t.img_w);
        int j = (int)(vv * (double)t.img_h);
        auto clampi = [](int x, int lo, int hi) { return x < lo ? lo : (x < hi ? x : hi - 1); };  // rtw_image.rs:70-78
        i = clampi(i, 0, t.img_w); j = clampi(j, 0, t.img_h);
        const uint8_t* px = t.img + ((size_t)j * t.img_w + i) * 3;
        double s = 1.0 / 255.0;
        return Color(s * px[0], s * px[1], s * px[2]);
      }
      default: {
        c.n.tex_noise += 1;
        return Color(0.5, 0.5, 0.5) * (1.0 + std::sin(t.scale * p.z() + 10.0 * t.perlin->turb(p, 7)));
      }
    }
  }
};

}  // namespace orc
