// oracle.cpp — ORACLE (test infrastructure only; never linked into or imported by the product).
// PARITY UNPINNED: no reference golden vectors exist and the Rust reference cannot be built here
// (SURVEY.md §8c); pinned by hand-derived known-answer tests only.
//
// CPU f64 restatement of the three Camera implementations, materials, PDFs and colour output:
//   in_one_weekend/camera.rs:59-209, material.rs:28-102
//   the_next_week/camera.rs:61-201, material.rs:41-158
//   the_rest_of_your_life/camera.rs:63-254, material.rs:53-190, pdf.rs:8-151, color.rs:6-42
//   src/render_io.rs:3-30
// Documented departures: (1) seeded RNG, one stream per image row;
// (2) primitive ids; (3) probe entry point; (4) op counters; rows run on std::thread workers with
// dynamic scheduling (the reference: rayon into_par_iter over rows, camera.rs:70-72).
#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "orc_math.hpp"
#include "orc_scene.hpp"
#include "oracle.h"

namespace orc {

struct CameraData {  // Camera fields + CameraInternals (camera.rs:15-60)
  rrt_camera cam;
  int image_height, sqrt_spp;
  double pixel_samples_scale, recip_sqrt_spp;
  Point3 center, pixel00_loc;
  Vec3 pixel_delta_u, pixel_delta_v, defocus_disk_u, defocus_disk_v;
  Color background;
  int integrator;
};

// Camera::initialize, the_rest_of_your_life/camera.rs:108-154
static void initialize(const rrt_camera* cam, int integrator, CameraData& d) {
  d.cam = *cam; d.integrator = integrator;
  int image_height = (int)((double)cam->image_width / cam->aspect_ratio);
  if (image_height < 1) image_height = 1;
  d.image_height = image_height;
  d.sqrt_spp = (int)std::sqrt((double)cam->samples_per_pixel);
  if (integrator == RRT_INT_BOOK3) {
    d.pixel_samples_scale = 1.0 / (double)(d.sqrt_spp * d.sqrt_spp);
    d.recip_sqrt_spp = 1.0 / (double)d.sqrt_spp;
  } else {
    d.pixel_samples_scale = 1.0 / (double)cam->samples_per_pixel;
    d.recip_sqrt_spp = 0.0;
  }
  Point3 lookfrom(cam->lookfrom[0], cam->lookfrom[1], cam->lookfrom[2]);
  Point3 lookat(cam->lookat[0], cam->lookat[1], cam->lookat[2]);
  Vec3 vup(cam->vup[0], cam->vup[1], cam->vup[2]);
  d.center = lookfrom;
  double theta = degrees_to_radians(cam->vfov);
  double h = std::tan(theta / 2.0);
  double viewport_height = 2.0 * h * cam->focus_dist;
  double viewport_width = viewport_height * ((double)cam->image_width / (double)image_height);
  Vec3 w = unit_vector(lookfrom - lookat);
  Vec3 u = unit_vector(cross(vup, w));
  Vec3 v = cross(w, u);
  Vec3 viewport_u = viewport_width * u;
  Vec3 viewport_v = viewport_height * -v;
  d.pixel_delta_u = viewport_u / (double)cam->image_width;
  d.pixel_delta_v = viewport_v / (double)image_height;
  Point3 viewport_upper_left = d.center - (cam->focus_dist * w) - viewport_u / 2.0 - viewport_v / 2.0;
  d.pixel00_loc = viewport_upper_left + 0.5 * (d.pixel_delta_u + d.pixel_delta_v);
  double defocus_radius = cam->focus_dist * std::tan(degrees_to_radians(cam->defocus_angle / 2.0));
  d.defocus_disk_u = u * defocus_radius;
  d.defocus_disk_v = v * defocus_radius;
  d.background = Color(cam->background[0], cam->background[1], cam->background[2]);
}

// get_ray: book 1 camera.rs:152-180 (no time), book 2 :148-172, book 3 :156-182 (stratified)
static Ray get_ray(const CameraData& d, int i, int j, int s_i, int s_j, Ctx& c) {
  double ox, oy;
  if (d.integrator == RRT_INT_BOOK3) {
    ox = (((double)s_i + c.random_double()) * d.recip_sqrt_spp) - 0.5;
    oy = (((double)s_j + c.random_double()) * d.recip_sqrt_spp) - 0.5;
  } else {
    ox = c.random_double() - 0.5;
    oy = c.random_double() - 0.5;
  }
  Point3 pixel_sample = d.pixel00_loc + ((double)i + ox) * d.pixel_delta_u + ((double)j + oy) * d.pixel_delta_v;
  Point3 ray_origin = d.center;
  if (!(d.cam.defocus_angle <= 0.0)) {
    Vec3 p = random_in_unit_disk(c);
    ray_origin = d.center + (p[0] * d.defocus_disk_u) + (p[1] * d.defocus_disk_v);
  }
  Vec3 ray_direction = pixel_sample - ray_origin;
  if (d.integrator == RRT_INT_BOOK1) return Ray(ray_origin, ray_direction);
  double ray_time = c.random_double();
  return Ray(ray_origin, ray_direction, ray_time);
}

static double reflectance(double cosine, double refraction_index) {  // material.rs:107-111
  double r0 = (1.0 - refraction_index) / (1.0 + refraction_index);
  r0 *= r0;
  double m = 1.0 - cosine;
  return r0 + (1.0 - r0) * (m * m * m * m * m);  // powi(5)
}

static Vec3 dielectric_direction(const Material& m, const Ray& r_in, const HitRecord& rec, Ctx& c) {  // :113-137
  c.n.dielectric += 1;
  double ri = rec.front_face ? 1.0 / m.ri : m.ri;
  Vec3 unit_direction = unit_vector(r_in.dir);
  double cos_theta = std::fmin(-dot(unit_direction, rec.normal), 1.0);
  double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
  bool cannot_refract = ri * sin_theta > 1.0;
  if (cannot_refract || reflectance(cos_theta, ri) > c.random_double()) return reflect(unit_direction, rec.normal);
  return refract(unit_direction, rec.normal, ri);
}

// Material::scatter of books 1 and 2 (the_next_week/material.rs:41-158; book 1 has no textures
// but the same arithmetic, in_one_weekend/material.rs:28-102).  Returns false when absorbed.
static bool scatter_b12(const Scene& S, const Material& m, const Ray& r_in, const HitRecord& rec, Color& attenuation,
                        Ray& scattered, Ctx& c) {
  switch (m.kind) {
    case RRT_M_LAMBERTIAN: {
      c.n.lambertian += 1;
      Vec3 scatter_direction = rec.normal + random_unit_vector(c);
      if (scatter_direction.near_zero()) scatter_direction = rec.normal;
      scattered = Ray(rec.p, scatter_direction, r_in.tm);
      attenuation = S.texture_value(m.tex, rec.u, rec.v, rec.p, c);
      return true;
    }
    case RRT_M_METAL: {
      c.n.metal += 1;
      Vec3 reflected = reflect(r_in.dir, rec.normal);
      reflected = unit_vector(reflected) + m.fuzz * random_unit_vector(c);
      scattered = Ray(rec.p, reflected, r_in.tm);
      attenuation = m.albedo;
      return dot(scattered.dir, rec.normal) > 0.0;
    }
    case RRT_M_DIELECTRIC: {
      attenuation = Color(1.0, 1.0, 1.0);
      scattered = Ray(rec.p, dielectric_direction(m, r_in, rec, c), r_in.tm);
      return true;
    }
    case RRT_M_ISOTROPIC: {
      c.n.isotropic += 1;
      scattered = Ray(rec.p, random_unit_vector(c), r_in.tm);
      attenuation = S.texture_value(m.tex, rec.u, rec.v, rec.p, c);
      return true;
    }
    default: return false;  // DiffuseLight, EmptyMaterial
  }
}

static double rr_prob(Color a) {  // clamp(max channel, 0.05, 0.95)
  double p = std::fmax(std::fmax(a.x(), a.y()), a.z());
  return p < 0.05 ? 0.05 : (p > 0.95 ? 0.95 : p);
}

// in_one_weekend/camera.rs:182-209
static Color ray_color_b1(const CameraData& d, const Scene& S, const Ray& r, int depth, Ctx& c) {
  if (depth <= 0) return Color(0, 0, 0);
  HitRecord rec;
  c.n.segments += 1;
  if (S.world->hit(r, Interval(0.001, INF), rec, c)) {
    Color attenuation; Ray scattered;
    if (scatter_b12(S, S.materials[rec.mat], r, rec, attenuation, scattered, c)) {
      int bounces = d.cam.max_depth - depth;
      if (bounces >= 5) {
        double p = rr_prob(attenuation);
        if (c.random_double() > p) return Color(0, 0, 0);
        return attenuation * ray_color_b1(d, S, scattered, depth - 1, c) / p;
      }
      return attenuation * ray_color_b1(d, S, scattered, depth - 1, c);
    }
    return Color(0, 0, 0);
  }
  Vec3 unit_direction = unit_vector(r.dir);
  double a = 0.5 * (unit_direction.y() + 1.0);
  return (1.0 - a) * Color(1.0, 1.0, 1.0) + a * Color(0.5, 0.7, 1.0);
}

// the_next_week/camera.rs:174-201
static Color ray_color_b2(const CameraData& d, const Scene& S, const Ray& r, int depth, Ctx& c) {
  if (depth <= 0) return Color(0, 0, 0);
  HitRecord rec;
  c.n.segments += 1;
  if (!S.world->hit(r, Interval(0.001, INF), rec, c)) return d.background;
  const Material& m = S.materials[rec.mat];
  Color emitted(0, 0, 0);
  if (m.kind == RRT_M_DIFFUSE_LIGHT) { c.n.emitted += 1; emitted = S.texture_value(m.tex, rec.u, rec.v, rec.p, c); }  // material.rs:131-135
  Color attenuation; Ray scattered;
  if (scatter_b12(S, m, r, rec, attenuation, scattered, c)) {
    int bounces = d.cam.max_depth - depth;
    if (bounces >= 5) {
      double p = rr_prob(attenuation);
      if (c.random_double() > p) return emitted;
      return emitted + (attenuation * ray_color_b2(d, S, scattered, depth - 1, c) / p);
    }
    return emitted + attenuation * ray_color_b2(d, S, scattered, depth - 1, c);
  }
  return emitted;
}

// the_rest_of_your_life/camera.rs:184-254 with material.rs:53-190 and pdf.rs inlined by kind
static Color ray_color_b3(const CameraData& d, const Scene& S, const Ray& r, int depth, Ctx& c) {
  if (depth <= 0) return Color(0, 0, 0);
  HitRecord rec;
  c.n.segments += 1;
  if (!S.world->hit(r, Interval(0.001, INF), rec, c)) return d.background;
  const Material& m = S.materials[rec.mat];
  Color emitted(0, 0, 0);
  if (m.kind == RRT_M_DIFFUSE_LIGHT && rec.front_face) {  // material.rs:155-160
    c.n.emitted += 1;
    emitted = S.texture_value(m.tex, rec.u, rec.v, rec.p, c);
  }
  // scatter -> ScatterRecord
  bool skip_pdf = false;
  Color attenuation;
  Ray skip_pdf_ray;
  int pdf_kind = -1;  // 0 cosine, 1 sphere
  switch (m.kind) {
    case RRT_M_LAMBERTIAN:
      c.n.lambertian += 1;
      attenuation = S.texture_value(m.tex, rec.u, rec.v, rec.p, c);
      pdf_kind = 0;
      break;
    case RRT_M_METAL: {  // :84-96: never absorbs
      c.n.metal += 1;
      Vec3 reflected = reflect(r.dir, rec.normal);
      reflected = unit_vector(reflected) + m.fuzz * random_unit_vector(c);
      attenuation = m.albedo; skip_pdf = true;
      skip_pdf_ray = Ray(rec.p, reflected, r.tm);
      break;
    }
    case RRT_M_DIELECTRIC:
      attenuation = Color(1.0, 1.0, 1.0); skip_pdf = true;
      skip_pdf_ray = Ray(rec.p, dielectric_direction(m, r, rec, c), r.tm);
      break;
    case RRT_M_ISOTROPIC:
      c.n.isotropic += 1;
      attenuation = S.texture_value(m.tex, rec.u, rec.v, rec.p, c);
      pdf_kind = 1;
      break;
    default:
      return emitted;  // scatter() -> None
  }
  if (skip_pdf) {  // camera.rs:199-217
    int bounces = d.cam.max_depth - depth;
    if (bounces >= 5) {
      double p = rr_prob(attenuation);
      if (c.random_double() > p) return emitted;
      return emitted + (attenuation * ray_color_b3(d, S, skip_pdf_ray, depth - 1, c) / p);
    }
    return emitted + attenuation * ray_color_b3(d, S, skip_pdf_ray, depth - 1, c);
  }
  int bounces = d.cam.max_depth - depth;
  double rr = bounces >= 5 ? rr_prob(attenuation) : 1.0;
  if (rr < 1.0 && c.random_double() > rr) return emitted;
  // MixturePdf(HittablePdf(lights, rec.p), srec.pdf): generate (pdf.rs:93-99) then value (:89-91)
  c.n.pdf_mix += 1;
  Onb uvw(rec.normal);  // CosinePdf::new(rec.normal), material.rs:57
  Vec3 direction;
  if (c.random_double() < 0.5) {
    direction = S.lights->random(rec.p, c);
  } else if (pdf_kind == 0) {
    direction = uvw.transform(random_cosine_direction(c));  // pdf.rs:51-53
  } else {
    direction = random_unit_vector(c);  // pdf.rs:27-29
  }
  Ray scattered(rec.p, direction, r.tm);
  double p1;
  if (pdf_kind == 0) {  // CosinePdf::value, pdf.rs:42-49
    double cosine_theta = dot(unit_vector(direction), uvw.w());
    p1 = cosine_theta <= 0.0 ? 0.0 : cosine_theta / PI;
  } else {
    p1 = 1.0 / (4.0 * PI);
  }
  double pdf_value = 0.5 * S.lights->pdf_value(rec.p, direction, c) + 0.5 * p1;
  if (pdf_value <= 0.0) return emitted;
  double scattering_pdf;
  if (pdf_kind == 0) {  // material.rs:63-70
    double cos_theta = dot(rec.normal, unit_vector(scattered.dir));
    scattering_pdf = cos_theta < 0.0 ? 0.0 : cos_theta / PI;
  } else {
    scattering_pdf = 1.0 / (4.0 * PI);  // :186-188
  }
  Color sample_color = ray_color_b3(d, S, scattered, depth - 1, c);
  Color color_from_scatter = attenuation * scattering_pdf * sample_color / (pdf_value * rr);
  return emitted + color_from_scatter;
}

// color.rs:6-42
static void write_color(Color pc, bool nan_scrub, int out[3]) {
  double v[3] = {pc.x(), pc.y(), pc.z()};
  for (int k = 0; k < 3; ++k) {
    double x = v[k];
    if (nan_scrub && std::isnan(x)) x = 0.0;
    x = x > 0.0 ? std::sqrt(x) : 0.0;  // linear_to_gamma
    x = Interval(0.0, 0.999).clamp(x);
    out[k] = (int)(256.0 * x);
  }
}

static uint64_t row_seed(uint64_t seed, int row) {
  return seed * 0x9e3779b97f4a7c15ULL + 0x632be59bd9b4e019ULL * (uint64_t)(row + 1);
}

}  // namespace orc

using namespace orc;

struct orc_scene { Scene S; };
static thread_local std::string g_err;

extern "C" {

const char* orc_last_error(void) { return g_err.c_str(); }

orc_scene* orc_scene_create(const rrt_scene_desc* desc) {
  orc_scene* s = new orc_scene();
  if (!s->S.build(desc)) { g_err = s->S.error; delete s; return nullptr; }
  return s;
}
void orc_scene_destroy(orc_scene* s) { delete s; }

int orc_image_height(const rrt_camera* cam) {
  CameraData d; initialize(cam, RRT_INT_BOOK2, d);
  return d.image_height;
}

void orc_camera_internals(const rrt_camera* cam, int integrator, double* out) {
  CameraData d; initialize(cam, integrator, d);
  const Vec3* vs[6] = {&d.center, &d.pixel00_loc, &d.pixel_delta_u, &d.pixel_delta_v, &d.defocus_disk_u, &d.defocus_disk_v};
  for (int i = 0; i < 6; ++i) for (int k = 0; k < 3; ++k) out[3 * i + k] = (*vs[i])[k];
  out[18] = d.image_height; out[19] = d.sqrt_spp; out[20] = d.pixel_samples_scale; out[21] = d.recip_sqrt_spp;
}

// departure #3: pixel-centre ray from the camera centre, time 0, media skipped
int orc_probe(orc_scene* s, const rrt_camera* cam, int32_t* prim_id, double* t) {
  CameraData d; initialize(cam, RRT_INT_BOOK2, d);
  s->S.probe_mode = true;
  Ctx c(0);
  for (int j = 0; j < d.image_height; ++j)
    for (int i = 0; i < cam->image_width; ++i) {
      Point3 pixel_sample = d.pixel00_loc + ((double)i + 0.0) * d.pixel_delta_u + ((double)j + 0.0) * d.pixel_delta_v;
      Ray r(d.center, pixel_sample - d.center, 0.0);
      HitRecord rec;
      bool hit = s->S.world->hit(r, Interval(0.001, INF), rec, c);
      size_t k = (size_t)j * cam->image_width + i;
      prim_id[k] = hit ? rec.id : -1;
      t[k] = hit ? rec.t : -1.0;
    }
  s->S.probe_mode = false;
  return 0;
}

// Camera::render over rows [row_begin, row_end) step row_step.  rgb: W*H*3 doubles of
// pixel_samples_scale * sum (what write_color receives); rows not rendered are left untouched.
// rgb_sq (optional): per-pixel mean of the squared per-sample radiance, per channel — the second
// moment the golden fixtures store so that a test can form standard errors (not a reference output).
int orc_render_moments(orc_scene* s, const rrt_camera* cam, int integrator, uint64_t seed, int nthreads, int row_begin,
                       int row_end, int row_step, double* rgb, double* rgb_sq, orc_counters* counters) {
  if (integrator == RRT_INT_BOOK3 && !s->S.lights) { g_err = "book-3 integrator needs a lights list"; return -1; }
  CameraData d; initialize(cam, integrator, d);
  if (row_end > d.image_height) row_end = d.image_height;
  if (row_step < 1) row_step = 1;
  if (nthreads < 1) nthreads = 1;
  std::vector<int> rows;
  for (int j = row_begin; j < row_end; j += row_step) rows.push_back(j);
  std::atomic<size_t> next(0);
  std::vector<Counters> per(nthreads);
  const Scene& S = s->S;
  auto worker = [&](int tid) {
    for (;;) {
      size_t k = next.fetch_add(1);
      if (k >= rows.size()) break;
      int j = rows[k];
      Ctx c(row_seed(seed, j));
      for (int i = 0; i < cam->image_width; ++i) {
        Color pixel_color(0, 0, 0), pixel_sq(0, 0, 0);
        if (integrator == RRT_INT_BOOK3) {
          for (int s_j = 0; s_j < d.sqrt_spp; ++s_j)
            for (int s_i = 0; s_i < d.sqrt_spp; ++s_i) {
              Ray r = get_ray(d, i, j, s_i, s_j, c);
              c.n.paths += 1;
              Color L = ray_color_b3(d, S, r, cam->max_depth, c);
              pixel_color = pixel_color + L;
              if (rgb_sq) pixel_sq = pixel_sq + L * L;
            }
        } else {
          for (int sample = 0; sample < cam->samples_per_pixel; ++sample) {
            Ray r = get_ray(d, i, j, 0, 0, c);
            c.n.paths += 1;
            Color L = integrator == RRT_INT_BOOK1 ? ray_color_b1(d, S, r, cam->max_depth, c)
                                                  : ray_color_b2(d, S, r, cam->max_depth, c);
            pixel_color = pixel_color + L;
            if (rgb_sq) pixel_sq = pixel_sq + L * L;
          }
        }
        Color out = d.pixel_samples_scale * pixel_color;
        double* px = rgb + ((size_t)j * cam->image_width + i) * 3;
        px[0] = out.x(); px[1] = out.y(); px[2] = out.z();
        if (rgb_sq) {
          Color sq = d.pixel_samples_scale * pixel_sq;
          double* qx = rgb_sq + ((size_t)j * cam->image_width + i) * 3;
          qx[0] = sq.x(); qx[1] = sq.y(); qx[2] = sq.z();
        }
      }
      per[tid].add(c.n);
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < nthreads; ++t) th.emplace_back(worker, t);
  worker(0);
  for (auto& x : th) x.join();
  if (counters) {
    Counters tot;
    for (auto& p : per) tot.add(p);
    counters->aabb = tot.aabb; counters->sph_test = tot.sph_test; counters->sph_hit = tot.sph_hit;
    counters->quad_test = tot.quad_test; counters->quad_full = tot.quad_full; counters->inst = tot.inst;
    counters->medium = tot.medium; counters->segments = tot.segments; counters->paths = tot.paths;
    counters->lambertian = tot.lambertian; counters->metal = tot.metal; counters->dielectric = tot.dielectric;
    counters->isotropic = tot.isotropic; counters->emitted = tot.emitted; counters->tex_checker = tot.tex_checker;
    counters->tex_image = tot.tex_image; counters->tex_noise = tot.tex_noise; counters->pdf_mix = tot.pdf_mix;
    counters->rng_draws = tot.rng_draws;
  }
  return 0;
}

int orc_render(orc_scene* s, const rrt_camera* cam, int integrator, uint64_t seed, int nthreads, int row_begin,
               int row_end, int row_step, double* rgb, orc_counters* counters) {
  return orc_render_moments(s, cam, integrator, seed, nthreads, row_begin, row_end, row_step, rgb, nullptr, counters);
}

// color.rs:14-42 (book 3 scrubs NaN, books 1-2 do not)
void orc_write_color(const double* rgb, int n_pixels, int nan_scrub, uint8_t* out) {
  for (int i = 0; i < n_pixels; ++i) {
    int b[3];
    write_color(Color(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]), nan_scrub != 0, b);
    out[3 * i] = (uint8_t)b[0]; out[3 * i + 1] = (uint8_t)b[1]; out[3 * i + 2] = (uint8_t)b[2];
  }
}

// src/render_io.rs:8-26 in f32, as the reference computes it
void orc_resolve_accum(const float* accum, int n_pixels, uint32_t spp, uint8_t* out) {
  float scale = spp > 0 ? 1.0f / (float)spp : 0.0f;
  for (int i = 0; i < n_pixels; ++i)
    for (int k = 0; k < 3; ++k) {
      float r = accum[4 * i + k] * scale;
      if (!std::isfinite(r)) r = 0.0f;
      r = std::sqrt(std::fmax(r, 0.0f));
      r = r < 0.0f ? 0.0f : (r > 0.999f ? 0.999f : r);
      out[3 * i + k] = (uint8_t)(r * 256.0f);
    }
}

// ---- known-answer hooks -------------------------------------------------------------------------
int orc_kat_hit(orc_scene* s, const double* origin, const double* dir, double time, double tmin, double tmax,
                double* out /* t, p[3], n[3], u, v, front, id, mat */) {
  Ctx c(0);
  Ray r(Point3(origin[0], origin[1], origin[2]), Vec3(dir[0], dir[1], dir[2]), time);
  HitRecord rec;
  s->S.probe_mode = true;
  bool hit = s->S.world->hit(r, Interval(tmin, tmax), rec, c);
  s->S.probe_mode = false;
  if (!hit) return 0;
  out[0] = rec.t;
  for (int k = 0; k < 3; ++k) { out[1 + k] = rec.p[k]; out[4 + k] = rec.normal[k]; }
  out[7] = rec.u; out[8] = rec.v; out[9] = rec.front_face ? 1.0 : 0.0; out[10] = rec.id; out[11] = rec.mat;
  return 1;
}
double orc_kat_reflectance(double cosine, double ri) { return reflectance(cosine, ri); }
void orc_kat_texture(orc_scene* s, int tex, double u, double v, const double* p, double* out) {
  Ctx c(0);
  Color col = s->S.texture_value(tex, u, v, Point3(p[0], p[1], p[2]), c);
  out[0] = col.x(); out[1] = col.y(); out[2] = col.z();
}
double orc_kat_light_pdf(orc_scene* s, const double* origin, const double* dir) {
  Ctx c(0);
  if (!s->S.lights) return -1.0;
  return s->S.lights->pdf_value(Point3(origin[0], origin[1], origin[2]), Vec3(dir[0], dir[1], dir[2]), c);
}
void orc_kat_refract(const double* uv, const double* n, double eta, double* out) {
  Vec3 r = refract(Vec3(uv[0], uv[1], uv[2]), Vec3(n[0], n[1], n[2]), eta);
  out[0] = r.x(); out[1] = r.y(); out[2] = r.z();
}
// world bounding box of the reference-shaped tree (checks Aabb padding / RotateY boxes)
void orc_kat_world_bbox(orc_scene* s, double* out6) {
  Aabb b = s->S.world->bounding_box();
  out6[0] = b.x.min; out6[1] = b.x.max; out6[2] = b.y.min; out6[3] = b.y.max; out6[4] = b.z.min; out6[5] = b.z.max;
}

}  // extern "C"
