/* oracle.h — C entry points of the ORACLE (oracle/liboracle.so).
 *
 * TEST INFRASTRUCTURE ONLY: may be loaded by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs — never by the product path (rustraytrace_b200/).
 * PARITY UNPINNED: the reference has no tests or golden vectors and cannot be built here (no Rust
 * toolchain); the restatement is pinned only by hand-derived known-answer tests.
 * Consumes the same rrt_scene_desc as librrt.so, but keeps the reference's pointer-tree shape. */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#include "../include/rrt.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_scene orc_scene;

typedef struct orc_counters {
  double aabb, sph_test, sph_hit, quad_test, quad_full, inst, medium, segments, paths;
  double lambertian, metal, dielectric, isotropic, emitted, tex_checker, tex_image, tex_noise, pdf_mix,
      rng_draws;
} orc_counters;

orc_scene* orc_scene_create(const rrt_scene_desc* desc);
void orc_scene_destroy(orc_scene* s);
const char* orc_last_error(void);
int orc_image_height(const rrt_camera* cam);
void orc_camera_internals(const rrt_camera* cam, int integrator, double* out22);
int orc_probe(orc_scene* s, const rrt_camera* cam, int32_t* prim_id, double* t);
int orc_render(orc_scene* s, const rrt_camera* cam, int integrator, uint64_t seed, int nthreads,
               int row_begin, int row_end, int row_step, double* rgb, orc_counters* counters);
/* as orc_render, plus (rgb_sq, may be NULL) the per-pixel mean of squared per-sample radiance */
int orc_render_moments(orc_scene* s, const rrt_camera* cam, int integrator, uint64_t seed, int nthreads,
                       int row_begin, int row_end, int row_step, double* rgb, double* rgb_sq,
                       orc_counters* counters);
void orc_write_color(const double* rgb, int n_pixels, int nan_scrub, uint8_t* out);
void orc_resolve_accum(const float* accum, int n_pixels, uint32_t spp, uint8_t* out);
int orc_kat_hit(orc_scene* s, const double* origin, const double* dir, double time, double tmin,
                double tmax, double* out12);
double orc_kat_reflectance(double cosine, double ri);
void orc_kat_texture(orc_scene* s, int tex, double u, double v, const double* p, double* out3);
double orc_kat_light_pdf(orc_scene* s, const double* origin, const double* dir);
void orc_kat_refract(const double* uv, const double* n, double eta, double* out3);
void orc_kat_world_bbox(orc_scene* s, double* out6);

#ifdef __cplusplus
}
#endif
#endif
