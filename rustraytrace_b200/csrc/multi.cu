// multi.cu — several GPUs of one box behind the C ABI (include/rrt.h).
//
// The reference has no multi-device code: its only parallel split is rayon over rows
// (the_rest_of_your_life/camera.rs:70-72) and, on its GPU path, additive passes into one accum
// buffer whose .w counts samples (src/cuda/mod.rs:403-432).  Both carry over: a rank renders a sample
// range of every pixel, or every sample of its own interleaved 32x32 tiles, and — because every
// (pixel, sample) draws from Philox keyed on (seed; pixel, sample, bounce) — the union of the shares
// is the unsharded frame up to float summation order.  The path itself needs no collective; the one
// exchange is the frame at the end, over NCCL/NVLink:
//   sample ranges: ncclReduce(SUM) of the W*H RGBA-sum frames to rank 0 (in place there);
//   tiles:         every rank packs its OWN super tiles back to back (render_kernel's compact mode),
//                  rank 0 receives them rank-major (grouped ncclSend/ncclRecv: 1/world of the frame per
//                  rank instead of world zero-padded frames through a SUM) and scatters them into the
//                  row-major frame with one kernel.
// Two ways in, one implementation: a group context (rrt_init_multi, one host thread drives every
// device) and one process per GPU (rrt_comm_init + rrt_render_collective).
#include <cuda_runtime.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "ctx.h"

using namespace rrt;

#define NK(call, what)                                                                  \
  do {                                                                                  \
    ncclResult_t r_ = (call);                                                           \
    if (r_ != ncclSuccess) return rrt::fail(RRT_ERR_CUDA, std::string(what) + ": NCCL " + ncclGetErrorString(r_)); \
  } while (0)

namespace {

ncclComm_t comm_of(const rrt_ctx* c) { return reinterpret_cast<ncclComm_t>(c->comm); }

// NCCL prints its version banner (NCCL_DEBUG=VERSION/INFO) on stdout, which is where the CLI writes the
// PPM (render_io.rs:3-30 prints to stdout): send NCCL's text to stderr unless the user chose a file.
void nccl_text_to_stderr() { setenv("NCCL_DEBUG_FILE", "/dev/stderr", 0); }

int resolve_mode(const rrt_render_params* p) { return p->shard_mode == RRT_SHARD_TILE ? RRT_SHARD_TILE : RRT_SHARD_SAMPLE; }

struct FramePlan {
  int width, height, n_super, stiles_x;
  size_t frame_floats;
  size_t off[RRT_MAX_WORLD + 1];  // tile mode: float4 offset of rank r's tiles in the staging buffer
};

FramePlan plan_frame(const rrt_camera* cam, int world) {
  FramePlan f;
  f.width = cam->image_width;
  f.height = rrt_image_height(cam);
  f.stiles_x = (f.width + 31) / 32;
  f.n_super = f.stiles_x * ((f.height + 31) / 32);
  f.frame_floats = (size_t)f.width * f.height * 4;
  f.off[0] = 0;
  for (int r = 0; r < world; ++r) f.off[r + 1] = f.off[r] + (size_t)owned_super_tiles(f.n_super, r, world) * 1024;
  return f;
}

// Where rank `rank` renders its share, allocating what it needs.  Rank 0 renders straight into what
// it will hand out (sample ranges: the frame, reduced in place) or scatter from (tiles: the head of
// the staging buffer).
int share_target(rrt_ctx* c, int rank, int mode, const FramePlan& f, int world, float* frame_root, float** dst) {
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  int rc;
  if (mode == RRT_SHARD_SAMPLE) {
    if (rank == 0) { *dst = frame_root; return RRT_OK; }
    if ((rc = ensure_floats(&c->share, &c->share_floats, f.frame_floats, "cudaMalloc(share)"))) return rc;
    *dst = c->share;
    return RRT_OK;
  }
  if (rank == 0) {
    if ((rc = ensure_floats(&c->staging, &c->staging_floats, f.off[world] * 4, "cudaMalloc(staging)"))) return rc;
    *dst = c->staging;
    return RRT_OK;
  }
  if ((rc = ensure_floats(&c->share, &c->share_floats, (f.off[rank + 1] - f.off[rank]) * 4, "cudaMalloc(share)"))) return rc;
  *dst = c->share;
  return RRT_OK;
}

// This rank's part of the exchange, enqueued on its stream; the caller brackets the calls of all
// ranks it drives with ncclGroupStart/End.
int exchange_enqueue(rrt_ctx* c, int rank, int world, int mode, const FramePlan& f, float* src, float* frame_root) {
  if (mode == RRT_SHARD_SAMPLE) {
    NK(ncclReduce(src, rank == 0 ? frame_root : nullptr, f.frame_floats, ncclFloat, ncclSum, 0, comm_of(c), c->stream), "frame reduce");
    return RRT_OK;
  }
  if (rank == 0) {
    for (int r = 1; r < world; ++r) {
      const size_t n = (f.off[r + 1] - f.off[r]) * 4;
      if (n) NK(ncclRecv(c->staging + f.off[r] * 4, n, ncclFloat, r, comm_of(c), c->stream), "tile gather");
    }
  } else {
    const size_t n = (f.off[rank + 1] - f.off[rank]) * 4;
    if (n) NK(ncclSend(src, n, ncclFloat, 0, comm_of(c), c->stream), "tile gather");
  }
  return RRT_OK;
}

int scatter_enqueue(rrt_ctx* root, int world, const FramePlan& f, float* frame_root) {
  TileScatterArgs A;
  std::memset(&A, 0, sizeof A);
  A.staging = reinterpret_cast<const float4*>(root->staging);
  A.frame = reinterpret_cast<float4*>(frame_root);
  A.width = f.width; A.height = f.height; A.stiles_x = f.stiles_x; A.world = world;
  for (int r = 0; r < world; ++r) A.off[r] = (unsigned)f.off[r];
  CK(launch_tile_scatter(A, root->stream), "kernel launch failed");
  return RRT_OK;
}

}  // namespace

// ---- group context (one process, n devices) ------------------------------------------------------------
int rrt::multi_comm_init_all(rrt_ctx* g, const int* devices) {
  const int n = (int)g->members.size();
  std::vector<ncclComm_t> comms(n);
  nccl_text_to_stderr();
  NK(ncclCommInitAll(comms.data(), n, devices), "ncclCommInitAll");
  for (int i = 0; i < n; ++i) {
    g->members[i]->comm = comms[i];
    g->members[i]->comm_rank = i;
    g->members[i]->comm_world = n;
  }
  return RRT_OK;
}

void rrt::comm_release(rrt_ctx* c) {
  if (c->comm) { ncclCommDestroy(comm_of(c)); c->comm = nullptr; }
}

int rrt::multi_scene_create(rrt_ctx* g, const FlatScene& F, rrt_scene** out) {
  rrt_scene* sc = new rrt_scene();
  sc->ctx = g;
  std::memset(&sc->dev, 0, sizeof sc->dev);
  for (rrt_ctx* m : g->members) {
    rrt_scene* part = nullptr;
    int rc = scene_upload(m, F, &part);
    if (rc) { scene_release(sc); for (rrt_scene* p : sc->parts) delete p; delete sc; return rc; }
    sc->parts.push_back(part);
    sc->h2d_bytes += part->h2d_bytes;
  }
  sc->n_nodes = sc->parts[0]->n_nodes; sc->n_prims = sc->parts[0]->n_prims; sc->n_grid_cells = sc->parts[0]->n_grid_cells;
  sc->n_media = sc->parts[0]->n_media; sc->n_lights = sc->parts[0]->n_lights;
  g->scenes.push_back(sc);
  *out = sc;
  return RRT_OK;
}

int rrt::multi_render(rrt_ctx* g, const rrt_scene* sc, const rrt_camera* cam, const rrt_render_params* p, float* accum_dev_root,
                      float* accum_host) {
  const int world = (int)g->members.size();
  rrt_ctx* root = g->members[0];
  const int mode = resolve_mode(p);
  const FramePlan f = plan_frame(cam, world);
  int rc;
  CK(cudaSetDevice(root->device), "cudaSetDevice");
  float* frame = accum_dev_root;
  if (!frame) {
    if ((rc = ensure_floats(&root->accum, &root->accum_floats, f.frame_floats, "cudaMalloc(accum)"))) return rc;
    frame = root->accum;
  }
  // every member's share, launched back to back from this thread; nothing waits on the host in between
  std::vector<ShareInfo> info(world);
  std::vector<float*> dst(world, nullptr);
  for (int r = 0; r < world; ++r) {
    rrt_render_params q = *p;
    q.shard_mode = mode; q.rank = r; q.world = world;
    if ((rc = share_target(g->members[r], r, mode, f, world, frame, &dst[r]))) return rc;
    if ((rc = render_enqueue(g->members[r], sc->parts[r], cam, &q, dst[r], mode == RRT_SHARD_TILE, &info[r]))) return rc;
  }
  CK(cudaSetDevice(root->device), "cudaSetDevice");
  CK(cudaEventRecord(root->ev[4], root->stream), "cudaEventRecord");
  if (world > 1) {
    NK(ncclGroupStart(), "ncclGroupStart");
    for (int r = 0; r < world; ++r)
      if ((rc = exchange_enqueue(g->members[r], r, world, mode, f, dst[r], frame))) { ncclGroupEnd(); return rc; }
    NK(ncclGroupEnd(), "ncclGroupEnd");
  }
  CK(cudaSetDevice(root->device), "cudaSetDevice");
  if (mode == RRT_SHARD_TILE && (rc = scatter_enqueue(root, world, f, frame))) return rc;
  CK(cudaEventRecord(root->ev[5], root->stream), "cudaEventRecord");
  if (accum_host) CK(cudaMemcpyAsync(accum_host, frame, f.frame_floats * sizeof(float), cudaMemcpyDeviceToHost, root->stream), "accum read-back");
  CK(cudaEventRecord(root->ev[3], root->stream), "cudaEventRecord");
  for (int r = world - 1; r >= 0; --r) {
    CK(cudaSetDevice(g->members[r]->device), "cudaSetDevice");
    CK(cudaStreamSynchronize(g->members[r]->stream), "render kernel failed");
  }
  // stats: the slowest member's kernel, everyone's paths, the exchange as rank 0 saw it
  rrt_stats& st = g->stats;
  double paths = 0, kmax = 0;
  int launches = 0;
  for (int r = 0; r < world; ++r) {
    stats_from_share(g->members[r], sc->parts[r], info[r]);
    paths += g->members[r]->stats.paths;
    kmax = std::max(kmax, g->members[r]->stats.kernel_ms);
    launches += g->members[r]->stats.kernel_launches;
  }
  st = root->stats;
  float xms = 0, tms = 0;
  cudaEventElapsedTime(&xms, root->ev[4], root->ev[5]);
  cudaEventElapsedTime(&tms, root->ev[0], root->ev[3]);
  st.paths = paths; st.kernel_ms = kmax; st.kernel_launches = launches + (mode == RRT_SHARD_TILE ? 1 : 0);
  st.exchange_ms = xms; st.total_ms = tms;
  st.n_devices = world;
  st.d2h_bytes = accum_host ? f.frame_floats * sizeof(float) : 0;
  st.h2d_bytes = sc->h2d_bytes;
  return RRT_OK;
}

// ---- one process per GPU -------------------------------------------------------------------------------
extern "C" {

int rrt_comm_unique_id(uint8_t* id128) {
  if (!id128) return rrt::fail(RRT_ERR_INVALID, "rrt_comm_unique_id: id128 is null");
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  nccl_text_to_stderr();
  NK(ncclGetUniqueId(&id), "ncclGetUniqueId");
  std::memcpy(id128, &id, sizeof id);
  return RRT_OK;
}

int rrt_comm_init(rrt_ctx* c, const uint8_t* id128, int rank, int world) {
  if (!c || !id128) return rrt::fail(RRT_ERR_INVALID, "rrt_comm_init: null argument");
  if (!c->members.empty() || c->group) return rrt::fail(RRT_ERR_INVALID, "rrt_comm_init: a multi-device context has its communicator already");
  if (world < 1 || world > RRT_MAX_WORLD || rank < 0 || rank >= world) return rrt::fail(RRT_ERR_INVALID, "rrt_comm_init: rank/world out of range");
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  comm_release(c);
  nccl_text_to_stderr();
  ncclUniqueId id;
  std::memcpy(&id, id128, sizeof id);
  ncclComm_t comm;
  NK(ncclCommInitRank(&comm, world, id, rank), "ncclCommInitRank");
  c->comm = comm; c->comm_rank = rank; c->comm_world = world;
  return RRT_OK;
}

int rrt_render_frame_collective(rrt_ctx* c, const rrt_scene_desc* desc, const rrt_camera* cam, const rrt_render_params* p,
                                float* accum_host_root) {
  rrt_scene* sc = nullptr;
  int rc = rrt::scene_create(c, desc, &sc, true);   // rendered once and freed: no tree where grids do the walking
  if (rc) return rc;
  rc = rrt_render_collective(c, sc, cam, p, nullptr, accum_host_root);
  const uint64_t h2d = sc->h2d_bytes;
  rrt_scene_destroy(sc);
  if (rc == RRT_OK) c->stats.h2d_bytes = h2d;
  return rc;
}

int rrt_render_collective(rrt_ctx* c, const rrt_scene* sc, const rrt_camera* cam, const rrt_render_params* p,
                          float* accum_dev_root, float* accum_host_root) {
  if (!c || !sc || !p) return rrt::fail(RRT_ERR_INVALID, "rrt_render_collective: null argument");
  if (sc->ctx != c) return rrt::fail(RRT_ERR_INVALID, "scene belongs to another (or a shut down) context");
  int rc = check_camera(cam, p);
  if (rc) return rc;
  if (!c->members.empty()) return multi_render(c, sc, cam, p, accum_dev_root, accum_host_root);
  if (!c->comm) return rrt::fail(RRT_ERR_INVALID, "rrt_render_collective: call rrt_comm_init first");
  const int rank = c->comm_rank, world = c->comm_world;
  const int mode = resolve_mode(p);
  const FramePlan f = plan_frame(cam, world);
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  float* frame = nullptr;
  if (rank == 0) {
    frame = accum_dev_root;
    if (!frame) {
      if ((rc = ensure_floats(&c->accum, &c->accum_floats, f.frame_floats, "cudaMalloc(accum)"))) return rc;
      frame = c->accum;
    }
  }
  rrt_render_params q = *p;
  q.shard_mode = mode; q.rank = rank; q.world = world;
  float* dst = nullptr;
  ShareInfo info;
  if ((rc = share_target(c, rank, mode, f, world, frame, &dst))) return rc;
  if ((rc = render_enqueue(c, sc, cam, &q, dst, mode == RRT_SHARD_TILE, &info))) return rc;
  CK(cudaEventRecord(c->ev[4], c->stream), "cudaEventRecord");
  if (world > 1) {
    NK(ncclGroupStart(), "ncclGroupStart");
    if ((rc = exchange_enqueue(c, rank, world, mode, f, dst, frame))) { ncclGroupEnd(); return rc; }
    NK(ncclGroupEnd(), "ncclGroupEnd");
  }
  if (rank == 0 && mode == RRT_SHARD_TILE && (rc = scatter_enqueue(c, world, f, frame))) return rc;
  CK(cudaEventRecord(c->ev[5], c->stream), "cudaEventRecord");
  const bool d2h = rank == 0 && accum_host_root;
  if (d2h) CK(cudaMemcpyAsync(accum_host_root, frame, f.frame_floats * sizeof(float), cudaMemcpyDeviceToHost, c->stream), "accum read-back");
  CK(cudaEventRecord(c->ev[3], c->stream), "cudaEventRecord");
  CK(cudaStreamSynchronize(c->stream), "render kernel failed");
  stats_from_share(c, sc, info);
  float xms = 0, tms = 0;
  cudaEventElapsedTime(&xms, c->ev[4], c->ev[5]);
  cudaEventElapsedTime(&tms, c->ev[0], c->ev[3]);
  c->stats.exchange_ms = xms; c->stats.total_ms = tms;
  c->stats.n_devices = world;
  c->stats.kernel_launches = info.launches + (rank == 0 && mode == RRT_SHARD_TILE ? 1 : 0);
  c->stats.d2h_bytes = d2h ? f.frame_floats * sizeof(float) : 0;
  return RRT_OK;
}

int rrt_tile_scatter_host(const rrt_camera* cam, int world, const float* const* shares, float* frame) {
  if (!cam || !shares || !frame) return rrt::fail(RRT_ERR_INVALID, "rrt_tile_scatter_host: null argument");
  if (world < 1 || world > RRT_MAX_WORLD) return rrt::fail(RRT_ERR_INVALID, "rrt_tile_scatter_host: world out of range");
  int rc = check_camera(cam, nullptr);
  if (rc) return rc;
  const FramePlan f = plan_frame(cam, world);
  for (int y = 0; y < f.height; ++y)
    for (int x = 0; x < f.width; ++x) {
      const int k = rrt::super_tile_index(x >> 5, y >> 5, f.stiles_x);
      const int r = k % world, j = k / world;
      if (!shares[r]) return rrt::fail(RRT_ERR_INVALID, "rrt_tile_scatter_host: a share is null");
      std::memcpy(frame + ((size_t)y * f.width + x) * 4, shares[r] + ((size_t)j * 1024 + (y & 31) * 32 + (x & 31)) * 4, 4 * sizeof(float));
    }
  return RRT_OK;
}

}  // extern "C"
