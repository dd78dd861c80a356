/* rrt.h — C ABI of librrt.so: the B200 (sm_100a) path-tracing hot path of RustRayTrace.
 *
 * This is the drop-in boundary.  It generalises the only C-shaped seam
 * the reference has —
 *     fn render(camera: CameraUniform, spheres: &[SphereGpu], materials: &[MaterialGpu])
 *         -> Result<(), String>                       (src/cuda/mod.rs:342, src/gpu/mod.rs:308)
 *     write_ppm_from_accum(width, height, accum: &[f32], spp)      (src/render_io.rs:3)
 * — from "spheres + 3 materials" to the whole scene graph of the three books.
 *
 * The host (Rust in the reference; the C++ mirror in rustraytrace_b200/host here) walks its
 * HittableObject tree once and emits the arrays below (one rrt_hittable per Arc<HittableObject>,
 * in construction order).  The library deep-copies, bakes instances, builds its own BVH,
 * uploads SoA buffers and renders.  All host pointers are borrowed for the duration of the call
 * only (as htod_sync_copy borrows its slices, src/cuda/mod.rs:389-390).
 *
 * Errors: 0 = ok, negative rrt_status otherwise, text via rrt_last_error() (thread-local),
 * replacing Result<(), String> (src/cuda/mod.rs:342-441).  Never aborts; never falls back to CPU.
 * Threading: blocking calls from one host thread per context; one context = one CUDA device, or a
 * group of them (rrt_init_multi) driven by that one thread.
 */
#ifndef RRT_H
#define RRT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum rrt_status {
  RRT_OK = 0,
  RRT_ERR_INVALID = -1,     /* bad argument / malformed scene description */
  RRT_ERR_CUDA = -2,        /* CUDA runtime error ("cuda init failed", "kernel launch failed" …) */
  RRT_ERR_UNSUPPORTED = -3, /* scene uses a construct the flattener does not cover */
  RRT_ERR_NO_DEVICE = -4    /* no CUDA device: there is no CPU fallback */
} rrt_status;

/* ---- scene description (host side, f64 like every reference CPU quantity) ---------------- */

/* HittableObject variants, the_rest_of_your_life/hittable.rs:180-188 */
typedef enum rrt_hittable_kind {
  RRT_H_SPHERE = 0,          /* sphere.rs:11-45   a=center0, b=center1-center0, c[0]=radius       */
  RRT_H_QUAD = 1,            /* quad.rs:11-48     a=Q, b=u, c=v                                   */
  RRT_H_LIST = 2,            /* hittable_list.rs  children[first_child .. +n_children)           */
  RRT_H_BVH = 3,             /* bvh.rs:16-19      BvhNode::new(list): children = the list        */
  RRT_H_TRANSLATE = 4,       /* hittable.rs:73-84 child=first_child, a=offset                     */
  RRT_H_ROTATE_Y = 5,        /* hittable.rs:100-140 child=first_child, a[0]=angle in degrees      */
  RRT_H_CONSTANT_MEDIUM = 6  /* constant_medium.rs:10-37 boundary=first_child, a[0]=density,
                                material = the Isotropic phase function                          */
} rrt_hittable_kind;

typedef struct rrt_hittable {
  int32_t kind;
  int32_t material;    /* index into materials[]; -1 for containers */
  int32_t first_child; /* LIST/BVH: offset into children[]; TRANSLATE/ROTATE_Y/MEDIUM: hittable index */
  int32_t n_children;  /* LIST/BVH only */
  double a[3];
  double b[3];
  double c[3];
} rrt_hittable;

/* MaterialObject variants, the_rest_of_your_life/material.rs:195-202 */
typedef enum rrt_material_kind {
  RRT_M_LAMBERTIAN = 0,   /* texture */
  RRT_M_METAL = 1,        /* albedo, fuzz (already min(fuzz,1), material.rs:79-81) */
  RRT_M_DIELECTRIC = 2,   /* refraction_index */
  RRT_M_DIFFUSE_LIGHT = 3,/* texture */
  RRT_M_ISOTROPIC = 4,    /* texture */
  RRT_M_EMPTY = 5         /* EmptyMaterial (light proxies), material.rs:192-193 */
} rrt_material_kind;

typedef struct rrt_material {
  int32_t kind;
  int32_t texture; /* index into textures[] or -1 */
  double albedo[3];
  double fuzz;
  double refraction_index;
} rrt_material;

/* TextureObject variants, the_rest_of_your_life/texture.rs:128-133 */
typedef enum rrt_texture_kind {
  RRT_T_SOLID = 0,   /* color */
  RRT_T_CHECKER = 1, /* scale (inv_scale = 1/scale, texture.rs:51), even, odd = texture indices */
  RRT_T_IMAGE = 2,   /* image = index into images[]; missing image => height 0 => cyan (texture.rs:91-93) */
  RRT_T_NOISE = 3    /* scale, perlin = index into perlins[] */
} rrt_texture_kind;

typedef struct rrt_texture {
  int32_t kind;
  int32_t even;   /* checker */
  int32_t odd;    /* checker */
  int32_t image;  /* image  */
  int32_t perlin; /* noise  */
  int32_t _pad;
  double color[3];
  double scale;
} rrt_texture;

/* RtwImage after image::open(..).to_rgb8(), rtw_image.rs:57-67: row-major RGB8 */
typedef struct rrt_image {
  int32_t width, height;
  const uint8_t* rgb; /* width*height*3 bytes, NULL/0x0 when the file was not found */
} rrt_image;

/* Perlin tables, perlin.rs:4-23 */
typedef struct rrt_perlin {
  double randvec[256][3];
  int32_t perm_x[256], perm_y[256], perm_z[256];
} rrt_perlin;

typedef struct rrt_scene_desc {
  const rrt_hittable* hittables; int32_t n_hittables;
  const int32_t* children;       int32_t n_children;   /* hittable indices, for LIST/BVH */
  const rrt_material* materials; int32_t n_materials;
  const rrt_texture* textures;   int32_t n_textures;
  const rrt_image* images;       int32_t n_images;
  const rrt_perlin* perlins;     int32_t n_perlins;
  int32_t world;  /* hittable index of the object passed to Camera::render as `world` */
  int32_t lights; /* hittable index of book 3's `lights` list (camera.rs:63), or -1 */
} rrt_scene_desc;

/* Camera pub fields, the_rest_of_your_life/camera.rs:15-29 (book 1 lacks `background`). */
typedef struct rrt_camera {
  double aspect_ratio;
  int32_t image_width;
  int32_t samples_per_pixel;
  int32_t max_depth;
  int32_t _pad;
  double background[3];
  double vfov;
  double lookfrom[3];
  double lookat[3];
  double vup[3];
  double defocus_angle;
  double focus_dist;
} rrt_camera;

/* Which book's Camera::ray_color / get_ray to run. */
typedef enum rrt_integrator {
  RRT_INT_BOOK1 = 1, /* in_one_weekend/camera.rs:152-209: sky gradient, no emission, no ray time */
  RRT_INT_BOOK2 = 2, /* the_next_week/camera.rs:148-201: background colour, emitted, ray time   */
  RRT_INT_BOOK3 = 3  /* the_rest_of_your_life/camera.rs:156-254: stratified, mixture-PDF lights */
} rrt_integrator;

/* How one rank's share of the frame is chosen.  Every (pixel, sample) draws its
 * randomness from Philox keyed on (seed; pixel, sample, bounce), so any partition gives the same
 * per-sample radiance. */
typedef enum rrt_shard_mode {
  RRT_SHARD_NONE = 0,
  RRT_SHARD_TILE = 1,  /* interleaved 32x32-pixel tiles: tile k belongs to rank k % world; tile row sy is numbered from
                        * column sy on (wrapping), k = sy * tiles_per_row + (sx - sy) mod tiles_per_row: diagonals, not columns */
  RRT_SHARD_SAMPLE = 2 /* samples [rank*n/world, (rank+1)*n/world) of every pixel          */
} rrt_shard_mode;

typedef struct rrt_render_params {
  int32_t integrator; /* rrt_integrator */
  int32_t shard_mode; /* rrt_shard_mode */
  int32_t rank;
  int32_t world;
  uint64_t seed;
} rrt_render_params;

/* Filled by every render call. */
typedef struct rrt_stats {
  int32_t width, height;
  int32_t spp_effective;   /* spp, or floor(sqrt(spp))^2 for book 3 (camera.rs:114-116) */
  int32_t kernel_launches; /* CUDA kernels of this library launched by the last call */
  double paths;            /* camera paths traced by this rank in the last call */
  double kernel_ms;        /* device time of the path-tracing kernel (CUDA events, its own stream) */
  double total_ms;         /* device time of the whole call's stream work */
  uint64_t h2d_bytes, d2h_bytes;
  int32_t n_prims, n_bvh_nodes, n_media, n_lights;
  int32_t sm_count;
  int32_t kernel_features; /* feature set of the render-kernel instance the last call ran: bit 0 moving
                            * spheres, 1 quads, 2 constant media, 3 non-solid textures, 4 Perlin noise,
                            * 5 box leaves (63 = the general instance), 6 the scene is walked through grids instead of a tree,
                            * 7 and the members of a grid move;
                            * a scene runs on the smallest instance covering it.
                            * RRT_KERNEL=full in the environment at rrt_init forces the general instance (debugging). */
  int32_t n_devices;       /* GPUs behind the last call: 1, the size of an rrt_init_multi group, or the communicator's world */
  int32_t n_grid_cells;    /* cells of the uniform grids the scene's even fields of spheres / boxes are walked through, or 0:
                            * the BVH (RRT_NO_GRID=1 at rrt_scene_create keeps the BVH for every scene) */
  double exchange_ms;      /* device time of the frame exchange after the kernels (NCCL reduce, or gather of owned tiles
                            * + scatter into the frame) on the rank that receives it; 0 on one GPU.  With several GPUs
                            * kernel_ms is the slowest member's kernel and paths counts every member's. */
} rrt_stats;

typedef struct rrt_ctx rrt_ctx;
typedef struct rrt_scene rrt_scene;

/* CudaDevice::new(0), src/cuda/mod.rs:343.  One context per process-and-GPU.
 * Debugging switches read from the environment (results are unchanged up to float summation order):
 *   RRT_KERNEL=full        (rrt_init) always run the general render-kernel instance
 *   RRT_NO_STEAL=1         (rrt_init) lanes never take over a neighbour pixel's samples
 *   RRT_DETERMINISTIC=1    (rrt_init) no stealing and one work item per tile: every pixel is one lane's
 *                          sum in sample order, one float4 store — frames repeat bit for bit (the default
 *                          schedule adds per-chunk sums with float atomics in whatever order they finish)
 *   RRT_NO_BOX_LEAVES=1    (rrt_scene_create) axis-aligned boxes stay six separate quad leaves
 *   RRT_NO_FRONT=1         (rrt_scene_create) outsized primitives (a ground sphere) stay in the BVH instead of
 *                          being tested ahead of the traversal
 *   RRT_NO_RECT_QUADS=1    (rrt_scene_create) axis-aligned rectangles of a root leaf are tested like any other quad
 *   RRT_NO_GRID=1          (rrt_scene_create) an even field of spheres is walked through its BVH, not its grid
 *   RRT_FORCE_GRID=1       (rrt_scene_create) a grid for every sphere-only tree, however uneven */
int rrt_init(int device, rrt_ctx** out);
void rrt_shutdown(rrt_ctx* ctx);

/* ---- several GPUs of one box -------------------------------------------------------
 * The reference has one caller, one blocking call (src/main.rs:58-66 -> fn render, src/cuda/mod.rs:342).
 * Both multi-GPU forms keep that shape: the frame is partitioned inside the library — sample ranges
 * (params->shard_mode = RRT_SHARD_SAMPLE, the default for a group when RRT_SHARD_NONE is passed) or
 * interleaved 32x32 tiles (RRT_SHARD_TILE) — the shares are rendered concurrently, and one NCCL exchange
 * over NVLink brings the frame to the first device: a SUM-reduce of the RGBA-sum frames for sample
 * ranges, a gather of each rank's OWN tiles (packed back to back; 1/world of the frame per rank, not a
 * zero-padded full frame) followed by a scatter kernel for tiles.  params->rank / world are ignored.
 *
 * (A) one process, n GPUs — what a host binding this header uses:
 *   rrt_init_multi: a group context over devices[0..n); devices[0] receives the frame.  Every entry
 *   point takes the group like a plain context: rrt_scene_create uploads to every member,
 *   rrt_render / rrt_render_frame / rrt_render_device return the complete frame (host, or device memory
 *   of devices[0]); rrt_probe, rrt_resolve_* and the rrt_kat_* hooks run on devices[0].
 * (B) one process per GPU (torchrun, MPI): every rank creates a plain context with rrt_init, rank 0 calls
 *   rrt_comm_unique_id and hands the 128 bytes to the others by whatever means the launcher offers, every
 *   rank calls rrt_comm_init (collective, as ncclCommInitRank), and then
 *   rrt_render_collective (collective): renders this rank's share and exchanges.  On rank 0 the frame
 *   lands in accum_dev_root (device, W*H*4 floats; NULL = library-owned) and, when accum_host_root is not
 *   NULL, is copied there too; both are ignored on the other ranks. */
int rrt_init_multi(const int* devices, int n_devices, rrt_ctx** out);
int rrt_comm_unique_id(uint8_t* id128);
int rrt_comm_init(rrt_ctx* ctx, const uint8_t* id128, int rank, int world);
int rrt_render_collective(rrt_ctx* ctx, const rrt_scene* scene, const rrt_camera* cam,
                          const rrt_render_params* params, float* accum_dev_root, float* accum_host_root);
/* rrt_render_frame for (B): every rank passes the same description; upload, this rank's share, exchange, free — one
 * call, as the reference's one blocking render (src/cuda/mod.rs:342).  The frame arrives in accum_host_root on rank 0. */
int rrt_render_frame_collective(rrt_ctx* ctx, const rrt_scene_desc* desc, const rrt_camera* cam,
                                const rrt_render_params* params, float* accum_host_root);
/* Host-only mirror of the tile exchange's last step (tests): shares[r] = rank r's compact tiles
 * (owned 32x32 super tiles back to back, 1024 RGBA pixels each, row-major inside a tile; super tile k
 * belongs to rank k % world and is its (k / world)-th) -> row-major RGBA frame. */
int rrt_tile_scatter_host(const rrt_camera* cam, int world, const float* const* shares, float* frame);

/* Launch on the caller's CUDA stream (a cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream)
 * instead of the context's own; NULL restores the library's stream.  The reference has one implicit
 * stream per CudaDevice (cudarc, src/cuda/mod.rs:343); this makes the same choice explicit so the
 * render can be ordered with the caller's NCCL reduce and timed with the caller's events. */
int rrt_set_stream(rrt_ctx* ctx, void* cuda_stream);

/* htod of the scene, src/cuda/mod.rs:389-390 — plus instance baking and the BVH build that
 * replace BvhNode::new (bvh.rs:21-156). */
int rrt_scene_create(rrt_ctx* ctx, const rrt_scene_desc* desc, rrt_scene** out);
void rrt_scene_destroy(rrt_scene* scene);

/* Image height as Camera::initialize computes it (camera.rs:109-112). */
int rrt_image_height(const rrt_camera* cam);

/* The render loop of src/cuda/mod.rs:403-436 / Camera::render (camera.rs:63-106): traces this
 * rank's share and writes row-major RGBA f32 sums, .w = samples accumulated
 * (layout of src/render_io.rs:11, src/cuda/mod.rs:330-332).
 *   rrt_render:        accum_host is caller-allocated W*H*4 floats; includes the D2H copy.
 *   rrt_render_device: accum_dev is a device pointer (W*H*4 floats) on this context's GPU; the call
 *                      returns after the kernels completed; nothing leaves HBM.
 *
 * Non-finite samples (the one deliberate difference from the reference's accumulation): Camera::render
 * adds every sample to the pixel (camera.rs:79-83), so a single NaN — a scattering point inside a sphere
 * light, where sphere.rs:110 takes sqrt of a negative number — makes the whole pixel NaN and write_color
 * (color.rs:19-27) prints it black.  Here a path ends where its pdf stops being a positive number, a
 * sample whose radiance is not finite adds nothing to rgb, and .w still counts it; none of the book
 * scenes produces such a sample (tests/test_book3_sampling.py pins the behaviour on one that does).
 * In f32 the same rule keeps one overflowing weight among 6e10 paths from blacking out a pixel.
 *
 * Reproducibility: per-sample radiance is a pure function of (seed, pixel, sample); frames of the default
 * schedule differ from run to run in the last bits (float atomics in completion order, see
 * RRT_DETERMINISTIC above). */
int rrt_render(rrt_ctx* ctx, const rrt_scene* scene, const rrt_camera* cam,
               const rrt_render_params* params, float* accum_host);
int rrt_render_device(rrt_ctx* ctx, const rrt_scene* scene, const rrt_camera* cam,
                      const rrt_render_params* params, float* accum_dev);

/* One-shot twin of cuda::imp::render (src/cuda/mod.rs:342-441): upload, render, read back, free. */
int rrt_render_frame(rrt_ctx* ctx, const rrt_scene_desc* desc, const rrt_camera* cam,
                     const rrt_render_params* params, float* accum_host);

/* Deterministic probe: one pixel-centre ray per pixel from the camera centre (no jitter, no
 * defocus, time 0), constant media excluded.  prim_id = index into desc->hittables of the closest
 * leaf (or -1), t = ray parameter in units of |direction| (ray.rs:31-33).  Host outputs, W*H each.
 * precision: 64 = f64 kernel (parity gate), 32 = the f32 intersection code of the render kernel. */
int rrt_probe(rrt_ctx* ctx, const rrt_scene* scene, const rrt_camera* cam, int precision,
              int32_t* prim_id, double* t);

/* rrt_probe without any allocation (csrc/noalloc.cu).  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments, same results as rrt_probe; the device
 * staging of the per-pixel outputs (12 bytes a pixel) is a static device array of
 * RRT_PROBE_STATIC_MAX_PIXELS pixels = 12 MiB, which every GPU that loads the library holds until exit.
 * A frame with more pixels is refused with RRT_ERR_UNSUPPORTED (use rrt_probe).  The call makes no
 * cudaMalloc / cudaFree; the first call on a device looks the array's address up (cudaGetSymbolAddress).
 * All calls on a device share the array and it keeps the last call's contents (every pixel of the frame
 * is written before it is read back, so nothing needs clearing): calls that overlap on one GPU race —
 * serialise them. */
#define RRT_PROBE_STATIC_MAX_PIXELS 1048576
int rrt_probe_static(rrt_ctx* ctx, const rrt_scene* scene, const rrt_camera* cam, int precision,
                     int32_t* prim_id, double* t);

/* write_ppm_from_accum's per-pixel arithmetic (src/render_io.rs:8-26) on the device:
 * accum (device, RGBA f32) -> rgb8 (device, W*H*3).  spp<=0 uses accum.w per pixel. */
int rrt_resolve_rgb8_device(rrt_ctx* ctx, const float* accum_dev, int width, int height, int spp,
                            uint8_t* rgb_dev);
/* Same, host in / host out (copies inside). */
int rrt_resolve_rgb8(rrt_ctx* ctx, const float* accum_host, int width, int height, int spp,
                     uint8_t* rgb_host);

int rrt_get_stats(const rrt_ctx* ctx, rrt_stats* out);

/* Host-only: the share of the frame rrt_render gives to params->rank (what rayon's row split,
 * camera.rs:70-72, and the CUDA path's pass loop, src/cuda/mod.rs:403-432, become across GPUs).
 * Outputs may be NULL; owner_mask is W*H bytes, 1 where this rank writes the pixel. */
int rrt_shard_plan(const rrt_camera* cam, const rrt_render_params* params, int32_t* sample_begin,
                   int32_t* sample_end, int32_t* n_super_tiles, double* n_pixels, uint8_t* owner_mask);

/* Host-only hooks (no GPU needed; used by the CPU test tier):
 *   rrt_flatten_check: runs the flattener + BVH build on the host and verifies the result (every
 *   primitive in exactly one leaf, every child box inside its parent's).  Outputs may be NULL.
 *   rrt_scene_features: the feature bits the flattened scene needs (rrt_stats.kernel_features lists
 *   them); the render calls pick the smallest kernel instance that covers them.
 *   rrt_philox_kat:    one Philox4x32-10 block, for known-answer tests of the sampling RNG's round
 *   function (the render kernel runs the same function for 7 rounds). */
int rrt_flatten_check(const rrt_scene_desc* desc, int32_t* n_prims, int32_t* n_nodes, int32_t* bvh_depth,
                      int32_t* n_media, int32_t* n_lights);
int rrt_scene_features(const rrt_scene_desc* desc, int32_t* features);
/*   rrt_flatten_grid: the uniform grid the flattener lays over an even field of equal spheres (kind 0) or
 *   axis-aligned boxes (kind 1) (rrt_stats.n_grid_cells): cells per axis (0, 0, 0: no such grid — without any grid
 *   the scene is walked through its BVH), items in it, references in all cells, most in one cell. */
/*   rrt_flatten_rects: how many quads of a brute-force root leaf (scenes of at most 16 primitives: the Cornell
 *   boxes) are axis-aligned rectangles with their normal along x, y, z — they get a test of their own. */
int rrt_flatten_rects(const rrt_scene_desc* desc, int32_t n3[3]);
int rrt_flatten_grid(const rrt_scene_desc* desc, int kind, int32_t res3[3], int32_t* n_items, int32_t* n_refs,
                     int32_t* max_per_cell);
void rrt_philox_kat(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                    uint32_t* out4);
/*   rrt_philox_kat_rounds: the same block at 7 or 10 rounds (Random123 publishes known answers for both);
 *   rrt_philox_rounds: the round count the render kernel was compiled with (7). */
int rrt_philox_kat_rounds(int rounds, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                          uint32_t k1, uint32_t* out4);
int rrt_philox_rounds(void);

/* Device-side known-answer hooks (tests; need a GPU).  They run the kernels' own code, not a copy:
 *   rrt_kat_light: for n (origin, direction) pairs, HittableObject::pdf_value of every entry of the
 *     scene's light list (quad.rs:93-102, sphere.rs:102-115) -> pdf_each[n * n_lights], the list's mean
 *     (hittable_list.rs:60-69) -> pdf_list[n], and HittableObject::random(origin) drawn with the two
 *     uniforms u[2i], u[2i+1] (quad.rs:104-107, sphere.rs:117-122) -> random_dir[n * n_lights * 3].
 *   rrt_kat_rng: the sampling RNG's blocks for (seed; pixel, sample0 + i, bounce, block), i < n:
 *     raw4[4n] Philox words and/or uniform5[5n] = the five uniforms (x, y, z, w, v) the kernel draws. */
int rrt_kat_light(rrt_ctx* ctx, const rrt_scene* scene, int n, const double* origin, const double* dir,
                  const double* u, double* pdf_each, double* pdf_list, double* random_dir);
int rrt_kat_rng(rrt_ctx* ctx, uint64_t seed, uint32_t pixel, uint32_t sample0, uint32_t bounce,
                uint32_t block, int n, uint32_t* raw4, float* uniform5);
const char* rrt_last_error(void);
const char* rrt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RRT_H */
