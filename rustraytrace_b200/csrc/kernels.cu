// kernels.cu — probe / resolve kernels and the host launchers of librrt.so.
//
//   render_kernel<INTEGRATOR, FEAT>  render_kernel.cuh, instantiated in kernels_book{1,2,3}.cu
//   probe_kernel64/32          deterministic primary-hit probe (SURVEY.md §8d correctness gate).
//   resolve_kernel             write_ppm_from_accum's arithmetic (src/render_io.rs:8-26).
#include <cuda_runtime.h>
#include <math_constants.h>

#include "../../include/rrt.h"
#include "device_types.h"
#include "intersect.cuh"
#include "kernels.h"
#include "rng.cuh"
#include "shade.cuh"
#include "vecmath.cuh"

namespace rrt {

// ------------------------------------------------------------------------------------------------
// probe: one pixel-centre ray per pixel, origin = camera centre, time 0, media excluded.
__global__ void probe_kernel32(const __grid_constant__ ProbeArgs A) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= A.width || y >= A.height) return;
  const DevCamera& C = A.cam;
  Ray32 r;
  float fx = (float)x, fy = (float)y;
  r.o = V3f(C.cx, C.cy, C.cz);
  r.d = V3f(C.p00x + fx * C.dux + fy * C.dvx, C.p00y + fx * C.duy + fy * C.dvy, C.p00z + fx * C.duz + fy * C.dvz);
  r.tm = 0.0f;
  Hit32 h; h.t = CUDART_INF_F; h.prim = -1;
  if (A.S.n_grids > 0) world_hit32<SF_ALL | SF_GRID | SF_GRID_MOVING>(A.S, r, 0.001f, -1, h);
  else world_hit32<SF_ALL>(A.S, r, 0.001f, -1, h);
  int id = -1;
  if (h.prim >= 0) {
    int idx = h.prim & ~PRIM_QUAD_BIT;
    id = __float_as_int(ldg4(A.S.prims + 4 * idx + ((h.prim & PRIM_QUAD_BIT) ? 3 : 2)).y);
  }
  A.prim_id[(size_t)y * A.width + x] = id;
  A.t[(size_t)y * A.width + x] = h.prim >= 0 ? (double)h.t : -1.0;
}

// write_ppm_from_accum per-pixel arithmetic, src/render_io.rs:8-26
__global__ void resolve_kernel(const float4* __restrict__ accum, int n, int spp, unsigned char* __restrict__ rgb) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 a = accum[i];
  float scale = spp > 0 ? 1.0f / (float)spp : (a.w > 0.0f ? 1.0f / a.w : 0.0f);
  float c[3] = {a.x * scale, a.y * scale, a.z * scale};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float v = c[k];
    if (!isfinite(v)) v = 0.0f;
    v = sqrtf(fmaxf(v, 0.0f));
    v = fminf(fmaxf(v, 0.0f), 0.999f);
    rgb[3 * (size_t)i + k] = (unsigned char)(v * 256.0f);
  }
}

// Tile exchange: super tile k (device_types.h super_tile_index) belongs to rank k % world and is that rank's
// (k / world)-th; one float4 per pixel, coalesced on both sides (a warp reads one tile row).
__global__ void tile_scatter_kernel(const __grid_constant__ TileScatterArgs A) {
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (x >= A.width || y >= A.height) return;
  const int k = super_tile_index(x >> 5, y >> 5, A.stiles_x);
  const int r = k % A.world, j = k / A.world;
  A.frame[(size_t)y * A.width + x] = A.staging[(size_t)A.off[r] + (size_t)j * 1024 + (y & 31) * 32 + (x & 31)];
}

// ------------------------------------------------------------------------------------------------
// known-answer hooks: the DEVICE code of the book-3 light sampling (shade.cuh light_pdf / light_random,
// = quad.rs:93-107, sphere.rs:102-122, hittable_list.rs:60-75) and of the sampling RNG (rng.cuh), driven
// with caller-chosen arguments so that tests can hold them against closed forms and the oracle.
__global__ void kat_light_kernel(const __grid_constant__ KatLightArgs A) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n) return;
  const V3f o(A.origin[3 * i], A.origin[3 * i + 1], A.origin[3 * i + 2]);
  const V3f d(A.dir[3 * i], A.dir[3 * i + 1], A.dir[3 * i + 2]);
  const int nl = A.S.n_lights;
  float sum = 0.0f;
  for (int l = 0; l < nl; ++l) {
    const float p = light_pdf(A.S.lights[l], o, d);
    sum += p;
    A.pdf_each[(size_t)i * nl + l] = p;
    const V3f r = light_random(A.S.lights[l], o, A.u[2 * i], A.u[2 * i + 1]);
    float* dst = A.random_dir + ((size_t)i * nl + l) * 3;
    dst[0] = r.x; dst[1] = r.y; dst[2] = r.z;
  }
  A.pdf_list[i] = nl > 0 ? sum / (float)nl : 0.0f;   // hittable_list.rs:60-69
}

__global__ void kat_rng_kernel(unsigned k0, unsigned k1, unsigned pixel, unsigned sample0, unsigned bounce, unsigned block,
                               int n, uint32_t* raw4, float* u5) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  PathRng rng;
  uint32_t rk[2 * RRT_PHILOX_ROUNDS];
  philox_round_keys(k0, k1, rk, RRT_PHILOX_ROUNDS);
  rng.pixel = pixel; rng.sample = sample0 + (unsigned)i; rng.rk = rk;
  if (raw4) philox4x32<RRT_PHILOX_ROUNDS>(rng.pixel, rng.sample, bounce, block, k0, k1, raw4 + 4 * (size_t)i);
  if (u5) {
    const Rand4 r = rng.draw(bounce, block);
    float* dst = u5 + 5 * (size_t)i;
    dst[0] = r.x; dst[1] = r.y; dst[2] = r.z; dst[3] = r.w; dst[4] = r.v;
  }
}

// ------------------------------------------------------------------------------------------------
// launchers (host)
cudaError_t launch_tile_scatter(const TileScatterArgs& A, cudaStream_t st) {
  dim3 block(32, 8), grid((A.width + 31) / 32, (A.height + 7) / 8);
  tile_scatter_kernel<<<grid, block, 0, st>>>(A);
  return cudaGetLastError();
}
cudaError_t launch_kat_light(const KatLightArgs& A, cudaStream_t st) {
  if (A.n <= 0) return cudaSuccess;
  kat_light_kernel<<<(A.n + 127) / 128, 128, 0, st>>>(A);
  return cudaGetLastError();
}
cudaError_t launch_kat_rng(unsigned k0, unsigned k1, unsigned pixel, unsigned sample0, unsigned bounce, unsigned block, int n,
                           uint32_t* raw4, float* u5, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  kat_rng_kernel<<<(n + 255) / 256, 256, 0, st>>>(k0, k1, pixel, sample0, bounce, block, n, raw4, u5);
  return cudaGetLastError();
}

const RenderVariant* select_render_variant(int integrator, int features, bool force_full) {
  int n = 0;
  const RenderVariant* t = integrator == RRT_INT_BOOK1 ? render_variants_book1(&n)
                         : integrator == RRT_INT_BOOK2 ? render_variants_book2(&n) : render_variants_book3(&n);
  // the way the scene is walked (SF_GRID) must match; among those instances the first that covers the scene, or —
  // force_full — the general one, which is the last of its kind in every table
  // (SF_GRID_MOVING likewise; a table without such an instance — book 1 has no ray time — serves the scene with the
  // plain grid instances, which clip to the bounds of the whole shutter interval)
  const int WALK = SF_GRID | SF_GRID_MOVING;
  for (int pass = 0; pass < 2; ++pass) {
    const RenderVariant* general = nullptr;
    for (int i = 0; i < n; ++i) {
      if ((t[i].features & WALK) != (features & WALK)) continue;
      if (!force_full && (t[i].features & features) == features) return &t[i];
      general = &t[i];
    }
    if (general) return general;
    features &= ~SF_GRID_MOVING;
  }
  return &t[n - 1];
}

static const void* render_entry(const RenderVariant* v, int schedule) {
  return reinterpret_cast<const void*>(schedule == RS_POOL ? v->kernel_pool : schedule == RS_STEAL ? v->kernel : v->kernel_own);
}

cudaError_t launch_render(const RenderVariant* v, int schedule, const RenderArgs& A, int grid, cudaStream_t st) {
  void* args[] = {const_cast<RenderArgs*>(&A)};
  return cudaLaunchKernel(render_entry(v, schedule), dim3(grid), dim3(RRT_BLOCK), args, 0, st);
}

int render_blocks_per_sm(const RenderVariant* v, int schedule) {
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_entry(v, schedule), RRT_BLOCK, 0) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

cudaError_t launch_probe(const ProbeArgs& A, int precision, cudaStream_t st) {
  dim3 block(16, 8), grid((A.width + 15) / 16, (A.height + 7) / 8);
  if (precision == 64) return launch_probe64(A, st);
  probe_kernel32<<<grid, block, 0, st>>>(A);
  return cudaGetLastError();
}

cudaError_t launch_resolve(const float* accum, int n, int spp, unsigned char* rgb, cudaStream_t st) {
  resolve_kernel<<<(n + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(accum), n, spp, rgb);
  return cudaGetLastError();
}

}  // namespace rrt
