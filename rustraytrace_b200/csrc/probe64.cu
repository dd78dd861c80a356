// probe64.cu — the f64 probe kernel, in its own translation unit so it can be compiled with
// -fmad=false and without fast-math: every multiply and add then rounds exactly as the reference's
// f64 Rust arithmetic does (rustc never contracts a*b+c), which is what makes primary-hit ids
// bit-exact even for rays that graze an edge shared by two quads.
#include <cuda_runtime.h>
#include <math_constants.h>

#include "device_types.h"
#include "intersect.cuh"
#include "kernels.h"
#include "vecmath.cuh"

namespace rrt {

__global__ void probe_kernel64(const __grid_constant__ ProbeArgs A) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= A.width || y >= A.height) return;
  const DevCamera64& C = A.cam64;
  V3d o(C.c[0], C.c[1], C.c[2]);
  // pixel00_loc + (i + 0) * pixel_delta_u + (j + 0) * pixel_delta_v - center  (camera.rs:158-167)
  double fi = (double)x, fj = (double)y;
  V3d ps(C.p00[0] + fi * C.du[0] + fj * C.dv[0], C.p00[1] + fi * C.du[1] + fj * C.dv[1], C.p00[2] + fi * C.du[2] + fj * C.dv[2]);
  V3d d = ps - o;
  Hit64 h; h.t = CUDART_INF; h.prim = -1;
  world_hit64(A.S, o, d, 0.001, h);
  int id = -1;
  if (h.prim >= 0) {
    int idx = h.prim & ~PRIM_QUAD_BIT;
    id = __float_as_int(ldg4(A.S.prims + 4 * idx + ((h.prim & PRIM_QUAD_BIT) ? 3 : 2)).y);
  }
  A.prim_id[(size_t)y * A.width + x] = id;
  A.t[(size_t)y * A.width + x] = h.prim >= 0 ? h.t : -1.0;
}

cudaError_t launch_probe64(const ProbeArgs& A, cudaStream_t st) {
  dim3 block(16, 8), grid((A.width + 15) / 16, (A.height + 7) / 8);
  probe_kernel64<<<grid, block, 0, st>>>(A);
  return cudaGetLastError();
}

}  // namespace rrt
