// ctx.h — the two handle types behind include/rrt.h and the internals api.cu and multi.cu share.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/rrt.h"
#include "device_types.h"
#include "flatten.hpp"
#include "kernels.h"

namespace rrt {
int fail(int st, const std::string& m);  // records the thread-local text of rrt_last_error(), returns st
}
#define CK(call, what)                                                                   \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) return rrt::fail(RRT_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e_)); \
  } while (0)

#define RRT_MAX_WORLD 64  // ranks / devices one exchange can span (offsets travel in a kernel argument block)

struct rrt_scene;
struct rrt_ctx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;      // the stream every call launches on
  cudaStream_t own_stream = nullptr;  // created by rrt_init; `stream` points elsewhere after rrt_set_stream
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // call start, kernel start/end, D2H end, exchange start/end
  unsigned int* counter = nullptr;
  float* accum = nullptr;  // grow-only frame buffer for the host-output entry points
  size_t accum_floats = 0;
  rrt_stats stats;
  bool no_steal = false;              // RRT_NO_STEAL=1: lanes never take over neighbours' samples (debugging / A-B)
  bool force_steal = false;           // RRT_FORCE_STEAL=1: stealing also in the Perlin instances (A-B)
  bool deterministic = false;         // RRT_DETERMINISTIC=1: no stealing, one chunk per tile -> bit-reproducible frames
  bool use_pool = false;              // RRT_POOL=1: the per-warp path-pool schedule where an instance has it (render_kernel_pool.cuh)
  bool force_full_kernel = false;     // RRT_KERNEL=full: always run the SF_ALL render kernel instance (include/rrt.h)
  std::vector<rrt_scene*> scenes;  // live scenes of this context (released by rrt_shutdown if the caller did not)

  // ---- multi-GPU (multi.cu) ----------------------------------------------------------------------
  // A group context (rrt_init_multi) owns one member context per device and launches nothing itself;
  // a member, or a plain context after rrt_comm_init, is one rank of an NCCL communicator.
  std::vector<rrt_ctx*> members;
  rrt_ctx* group = nullptr;           // the group a member belongs to
  void* comm = nullptr;               // ncclComm_t
  int comm_rank = 0, comm_world = 1;
  float* share = nullptr;             // this rank's share before the exchange: a full frame (sample ranges) or its
  size_t share_floats = 0;            //   compact tiles (tile sharding)
  float* staging = nullptr;           // rank 0, tile sharding: every rank's compact tiles, rank-major
  size_t staging_floats = 0;
};

struct rrt_scene {
  rrt_ctx* ctx = nullptr;
  std::vector<void*> allocs;
  rrt::DevScene dev;
  int n_nodes = 0, n_prims = 0, n_media = 0, n_lights = 0, n_grid_cells = 0;
  uint64_t h2d_bytes = 0;
  std::vector<rrt_scene*> parts;      // scene of a group context: one uploaded copy per member device
};

namespace rrt {

// What one enqueued share of a frame amounts to (filled by render_enqueue, turned into rrt_stats after the sync).
struct ShareInfo {
  int width = 0, height = 0, spp_eff = 0, n_samples = 0, launches = 0, features = -1;
  double pixels = 0;
  int n_super = 0;  // 32x32 super tiles this rank owns (tile sharding)
};

int check_camera(const rrt_camera* cam, const rrt_render_params* p);
int scene_upload(rrt_ctx* c, const FlatScene& F, rrt_scene** out);
// rrt_scene_create; tree_optional: the scene will only be rendered (no BVH where grids do the walking)
int scene_create(rrt_ctx* c, const rrt_scene_desc* desc, rrt_scene** out, bool tree_optional);
void scene_release(rrt_scene* sc);  // frees device memory, orphans the handle (does not delete it)
// Enqueue this rank's share on c->stream: events ev[0] (start), ev[1]/ev[2] around the kernel; no host sync.
// compact: tile sharding writes the rank's super tiles back to back (1024 float4 each) instead of a
// zero-filled full frame, which is what the gather moves.
int render_enqueue(rrt_ctx* c, const rrt_scene* sc, const rrt_camera* cam, const rrt_render_params* p, float* dst_dev,
                   bool compact, ShareInfo* info);
void stats_from_share(rrt_ctx* c, const rrt_scene* sc, const ShareInfo& s);  // after the stream was synchronised
int ensure_floats(float** buf, size_t* have, size_t floats, const char* what);
int owned_super_tiles(int n_super, int rank, int world);

// multi.cu
int multi_scene_create(rrt_ctx* g, const FlatScene& F, rrt_scene** out);
int multi_render(rrt_ctx* g, const rrt_scene* sc, const rrt_camera* cam, const rrt_render_params* p, float* accum_dev_root,
                 float* accum_host);
int multi_comm_init_all(rrt_ctx* g, const int* devices);  // ncclCommInitAll over the members
void comm_release(rrt_ctx* c);

}  // namespace rrt
