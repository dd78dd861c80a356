// kernels.h — kernel argument blocks and host launchers (kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include "device_types.h"

#define RRT_BLOCK 128

namespace rrt {

struct RenderArgs {
  DevScene S;
  DevCamera cam;
  DevWork W;
  float4* accum;          // W*H float4, row-major: rgb sums, w = samples (src/render_io.rs:11)
  unsigned int* counter;  // work-item counter, zeroed before the launch
};

struct ProbeArgs {
  DevScene S;
  DevCamera cam;
  DevCamera64 cam64;
  int width, height;
  int* prim_id;
  double* t;
};

struct KatLightArgs {   // rrt_kat_light: n queries against the scene's light list
  DevScene S;
  int n;
  const float* origin;   // n x 3
  const float* dir;      // n x 3
  const float* u;        // n x 2 uniforms for random()
  float* pdf_each;       // n x n_lights
  float* pdf_list;       // n
  float* random_dir;     // n x n_lights x 3
};

// Resident CTAs per SM each render kernel instance is compiled for (its register budget).  Every
// instance without Perlin noise fits 64 registers spill-free: 8 CTAs, 32 warps per SM.  The
// turbulence loop (7 octaves of 8 gradient lookups) pushes an instance to 80 registers (6 CTAs).
// The noise instances that also carry media (the book-2 final scene) run 7 CTAs at 72 registers with
// the same 40 bytes of spills they have at 80: C4 +3.5 %; the plain noise instances lose 2 % there
// (C3c) and stay at 6; 64 registers spills 110-180 bytes, and a rolled corner loop in perlin_noise
// does not remove that and costs C3c 57 % (profiles/README.md, r01j).
#ifndef RRT_MIN_BLOCKS_LEAN
#define RRT_MIN_BLOCKS_LEAN 8
#endif
#ifndef RRT_MIN_BLOCKS
#define RRT_MIN_BLOCKS 6
#endif
#ifndef RRT_MIN_BLOCKS_NOISE_MEDIA
#define RRT_MIN_BLOCKS_NOISE_MEDIA 7
#endif
constexpr int render_min_blocks(int feat) {
  return (feat & SF_NOISE) ? ((feat & SF_MEDIA) ? RRT_MIN_BLOCKS_NOISE_MEDIA : RRT_MIN_BLOCKS) : RRT_MIN_BLOCKS_LEAN;
}

// One compiled instance of render_kernel<INTEGRATOR, FEAT> (kernels_book{1,2,3}.cu).
struct RenderVariant {
  int features;                 // SceneFeature bits the instance supports
  void (*kernel)(RenderArgs);      // __global__ entry, lanes take over neighbours' samples (STEAL)
  void (*kernel_own)(RenderArgs);  // every lane keeps to its own pixel
  void (*kernel_pool)(RenderArgs); // per-warp path pool with compacted phases (render_kernel_pool.cuh), or null
};
enum RenderSchedule { RS_OWN = 0, RS_STEAL = 1, RS_POOL = 2 };
const RenderVariant* render_variants_book1(int* n);
const RenderVariant* render_variants_book2(int* n);
const RenderVariant* render_variants_book3(int* n);

// Smallest instance of the integrator covering `features` (force_full: the SF_ALL instance).
const RenderVariant* select_render_variant(int integrator, int features, bool force_full);
cudaError_t launch_render(const RenderVariant* v, int schedule, const RenderArgs& A, int grid, cudaStream_t st);
int render_blocks_per_sm(const RenderVariant* v, int schedule);
cudaError_t launch_probe(const ProbeArgs& A, int precision, cudaStream_t st);
cudaError_t launch_probe64(const ProbeArgs& A, cudaStream_t st);  // probe64.cu (-fmad=false)
// Tile exchange, last step on the receiving rank: every rank's packed super tiles (rank-major in
// `staging`, rank r starting at float4 offset off[r]) -> the row-major frame.
struct TileScatterArgs {
  const float4* staging;
  float4* frame;
  int width, height, stiles_x, world;
  unsigned off[64];
};
cudaError_t launch_tile_scatter(const TileScatterArgs& A, cudaStream_t st);
cudaError_t launch_kat_light(const KatLightArgs& A, cudaStream_t st);
cudaError_t launch_kat_rng(unsigned k0, unsigned k1, unsigned pixel, unsigned sample0, unsigned bounce, unsigned block, int n,
                           uint32_t* raw4, float* u5, cudaStream_t st);
cudaError_t launch_resolve(const float* accum, int n, int spp, unsigned char* rgb, cudaStream_t st);

}  // namespace rrt
