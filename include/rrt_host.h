/* rrt_host.h — C exports of librrt_host.so, the host-side mirror of the reference's Rust surface
 * ABOVE the rrt.h boundary: config.rs overrides, the src/books scene builders, camera defaults,
 * image lookup and render_io.rs PPM output.  In the reference these are Rust (src/main.rs,
 * src/config.rs, src/books/<book>/mod.rs, src/render_io.rs); Rust is not available in this image, so they
 * are restated in C++ (rustraytrace_b200/host/) and exported here for the `rrt` CLI, the tests and
 * bench.py.  Nothing in this header touches the GPU.
 */
#ifndef RRT_HOST_H
#define RRT_HOST_H
#include <stdint.h>
#include "rrt.h"
#ifdef __cplusplus
extern "C" {
#endif

/* RenderOverrides, src/config.rs:1-14.  Option<T> -> bit i of `has` set + value. */
enum {
  RRTH_O_ASPECT_RATIO = 1 << 0,
  RRTH_O_IMAGE_WIDTH = 1 << 1,
  RRTH_O_SAMPLES_PER_PIXEL = 1 << 2,
  RRTH_O_MAX_DEPTH = 1 << 3,
  RRTH_O_VFOV = 1 << 4,
  RRTH_O_LOOKFROM = 1 << 5,
  RRTH_O_LOOKAT = 1 << 6,
  RRTH_O_VUP = 1 << 7,
  RRTH_O_DEFOCUS_ANGLE = 1 << 8,
  RRTH_O_FOCUS_DIST = 1 << 9,
  RRTH_O_BACKGROUND = 1 << 10
};
typedef struct rrth_overrides {
  uint32_t has;
  int32_t image_width;
  int32_t samples_per_pixel;
  int32_t max_depth;
  double aspect_ratio;
  double vfov;
  double lookfrom[3];
  double lookat[3];
  double vup[3];
  double defocus_angle;
  double focus_dist;
  double background[3];
} rrth_overrides;

/* const OVERRIDES as shipped, src/config.rs:50-62 (width 2160, 5000 spp, depth 100). */
void rrth_overrides_shipped(rrth_overrides* out);
/* RenderOverrides::none(), src/config.rs:17-32. */
void rrth_overrides_none(rrth_overrides* out);

/* Book keys after normalize_book_name + the alias match of src/main.rs:7-12, 87-92. */
enum { RRTH_BOOK_UNKNOWN = 0, RRTH_BOOK_1 = 1, RRTH_BOOK_2 = 2, RRTH_BOOK_3 = 3,
       RRTH_BOOK_1_GPU = 4 /* build_in_one_weekend_scene, src/gpu/mod.rs:124-301 */ };
int rrth_normalize_book(const char* name);

typedef struct rrth_scene rrth_scene;

/* books::<book>::run(scene) up to, but not including, cam.render(): builds the world (+lights),
 * the Camera with the scene's defaults, applies the overrides.
 *   scene_no: the_next_week scene number (src/books/the_next_week/mod.rs:68-81); <0 = None.
 *   seed:     seeds the scene RNG (the reference uses entropy; its GPU path uses 0x5EED1234).
 *   image_dir: extra directory searched first for image textures (like $RTW_IMAGES,
 *              rtw_image.rs:14-19); may be NULL. */
int rrth_scene_build(int book, int scene_no, uint64_t seed, const rrth_overrides* ov,
                     const char* image_dir, rrth_scene** out);
const rrt_scene_desc* rrth_scene_desc(const rrth_scene* s);
const rrt_camera* rrth_scene_camera(const rrth_scene* s);
int rrth_scene_integrator(const rrth_scene* s);
void rrth_scene_free(rrth_scene* s);

/* write_ppm_from_accum, src/render_io.rs:3-30: text "P3" from RGBA f32 sums.
 * path == NULL writes to stdout.  Returns 0 or -1. */
int rrth_write_ppm_from_accum(const char* path, int width, int height, const float* accum,
                              uint32_t samples_per_pixel);
/* Same text format from already-resolved RGB8 (the bytes rrt_resolve_rgb8 produces). */
int rrth_write_ppm_rgb8(const char* path, int width, int height, const uint8_t* rgb);

/* RtwImage::new + load (the_rest_of_your_life/rtw_image.rs:10-67): the same search order
 * ($RTW_IMAGES, cwd, images/ in up to six parents; extra_dir first) and a native decode (baseline JPEG,
 * binary PPM).  rgb may be NULL to query the size; returns 0, -1 (not found / undecodable), -2
 * (rgb_capacity too small). */
int rrth_load_image(const char* filename, const char* extra_dir, int32_t* width, int32_t* height,
                    uint8_t* rgb, uint64_t rgb_capacity);

/* Known-answer hooks for the scene RNG (tests only). */
uint64_t rrth_rng_kat_state1234(int n);          /* n-th output of xoshiro256++ from state {1,2,3,4} */
double rrth_rng_kat_seeded_f64(uint64_t seed, int n); /* n-th random_double() after seed_from_u64 */

const char* rrth_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
