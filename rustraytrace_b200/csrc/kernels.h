// kernels.h — kernel argument blocks and host launchers (kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include "device_types.h"

#define RRT_BLOCK 128
#ifndef RRT_MIN_BLOCKS
#define RRT_MIN_BLOCKS 6  // 80 registers/thread, no spills; 24 warps per SM
#endif

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

cudaError_t launch_render(const RenderArgs& A, int grid, cudaStream_t st);
int render_blocks_per_sm(int integrator);
cudaError_t launch_probe(const ProbeArgs& A, int precision, cudaStream_t st);
cudaError_t launch_probe64(const ProbeArgs& A, cudaStream_t st);  // probe64.cu (-fmad=false)
cudaError_t launch_resolve(const float* accum, int n, int spp, unsigned char* rgb, cudaStream_t st);

}  // namespace rrt
