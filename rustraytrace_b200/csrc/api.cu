// api.cu — the extern "C" surface of librrt.so (include/rrt.h).  Host orchestration only: the
// per-call flow mirrors cuda::imp::render (src/cuda/mod.rs:342-441) — device init, scene upload,
// one kernel over the frame, read-back — minus the NVRTC step (kernels are compiled ahead of time
// for sm_100a) and minus the per-pass relaunch/synchronise loop (:403-432): one persistent launch
// covers every sample.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rrt.h"
#include "device_types.h"
#include "flatten.hpp"
#include "ctx.h"
#include "kernels.h"
#include "rng.cuh"

using namespace rrt;

static thread_local std::string g_err;
int rrt::fail(int st, const std::string& m) { g_err = m; return st; }
using rrt::fail;

template <typename T>
static int upload(rrt_scene* sc, const std::vector<T>& v, const T** out) {
  *out = nullptr;
  if (v.empty()) return RRT_OK;
  void* d = nullptr;
  size_t bytes = v.size() * sizeof(T);
  CK(cudaMalloc(&d, bytes), "cudaMalloc(scene)");
  sc->allocs.push_back(d);
  CK(cudaMemcpyAsync(d, v.data(), bytes, cudaMemcpyHostToDevice, sc->ctx->stream), "scene upload");
  sc->h2d_bytes += bytes;
  *out = reinterpret_cast<const T*>(d);
  return RRT_OK;
}

// the device a single-device operation of a group context runs on, and its copy of the scene
static rrt_ctx* root_of(rrt_ctx* c) { return c && !c->members.empty() ? c->members[0] : c; }
static const rrt_scene* part0_of(const rrt_scene* sc) { return sc && !sc->parts.empty() ? sc->parts[0] : sc; }

// ---- internals shared with multi.cu (ctx.h) -------------------------------------------------------
void rrt::scene_release(rrt_scene* sc) {
  for (rrt_scene* part : sc->parts) {
    if (rrt_ctx* m = part->ctx) {
      cudaSetDevice(m->device);
      cudaStreamSynchronize(m->stream);
      for (size_t i = 0; i < m->scenes.size(); ++i)
        if (m->scenes[i] == part) { m->scenes.erase(m->scenes.begin() + i); break; }
    }
    scene_release(part);
  }
  for (void* p : sc->allocs) cudaFree(p);
  sc->allocs.clear();
  sc->ctx = nullptr;
}

int rrt::scene_upload(rrt_ctx* c, const FlatScene& F, rrt_scene** out) {
  *out = nullptr;
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  rrt_scene* sc = new rrt_scene();
  sc->ctx = c;
  std::memset(&sc->dev, 0, sizeof sc->dev);
  DevScene& D = sc->dev;
  const unsigned char* texels = nullptr;
  int rc = RRT_OK;
  if ((rc = upload(sc, F.prims, &D.prims)) || (rc = upload(sc, F.prims64, &D.prims64)) ||
      (rc = upload(sc, F.nodes, &D.nodes)) || (rc = upload(sc, F.bprims, &D.bprims)) ||
      (rc = upload(sc, F.boxes, &D.boxes)) || (rc = upload(sc, F.qpairs, &D.qpairs)) || (rc = upload(sc, F.rects, &D.rects)) || (rc = upload(sc, F.nodes4, &D.nodes4)) ||
      (rc = upload(sc, F.materials, &D.materials)) || (rc = upload(sc, F.textures, &D.textures)) ||
      (rc = upload(sc, F.images, &D.images)) || (rc = upload(sc, F.texels, &texels)) ||
      (rc = upload(sc, F.perlins, &D.perlins)) || (rc = upload(sc, F.media, &D.media)) ||
      (rc = upload(sc, F.lights, &D.lights)) || (rc = upload(sc, F.grid_cells, &D.grid_cells)) || (rc = upload(sc, F.grid_refs, &D.grid_refs))) {
    scene_release(sc); delete sc;
    return rc;
  }
  D.texels = texels;
  D.root_ref = F.root_ref;
  D.front_ref = F.front_ref;
  D.n_nodes = (int)(F.nodes.size() / 4);
  D.n_prims = F.n_prims;
  D.n_media = (int)F.media.size();
  D.n_lights = (int)F.lights.size();
  D.has_motion = F.has_motion;
  D.features = F.features;
  for (int a = 0; a < 3; ++a) D.n_rects[a] = F.n_rects[a];
  D.n_grids = 0;
  for (size_t g = 0; g < F.grids.size() && g < RRT_MAX_GRIDS; ++g) { D.grids[g] = F.grids[g]; D.n_grids += F.grids[g].n_items > 0; }
  sc->n_nodes = D.n_nodes; sc->n_prims = D.n_prims; sc->n_media = D.n_media; sc->n_lights = D.n_lights;
  sc->n_grid_cells = (int)F.grid_cells.size();
  cudaError_t e = cudaStreamSynchronize(c->stream);  // the host vectors die when the caller returns
  if (e != cudaSuccess) { scene_release(sc); delete sc; return fail(RRT_ERR_CUDA, std::string("scene upload: ") + cudaGetErrorString(e)); }
  c->scenes.push_back(sc);
  *out = sc;
  return RRT_OK;
}

int rrt::check_camera(const rrt_camera* cam, const rrt_render_params* p) {
  if (!cam) return fail(RRT_ERR_INVALID, "camera is null");
  if (cam->image_width < 1 || cam->image_width > 65536 || !(cam->aspect_ratio > 0.0))
    return fail(RRT_ERR_INVALID, "camera: image_width/aspect_ratio out of range");
  // pixel indices are 32-bit in the kernels (rng.pixel, py * width + px): cap the frame at 2^30 pixels
  // (a 16 GiB RGBA-sum frame) instead of overflowing
  if ((double)cam->image_width / cam->aspect_ratio * (double)cam->image_width > 1073741824.0)
    return fail(RRT_ERR_INVALID, "camera: frame larger than 2^30 pixels");
  if (p) {
    if (p->integrator < RRT_INT_BOOK1 || p->integrator > RRT_INT_BOOK3) return fail(RRT_ERR_INVALID, "params: unknown integrator");
    if (p->shard_mode < RRT_SHARD_NONE || p->shard_mode > RRT_SHARD_SAMPLE) return fail(RRT_ERR_INVALID, "params: unknown shard mode");
    if (p->shard_mode != RRT_SHARD_NONE && (p->world < 1 || p->rank < 0 || p->rank >= p->world))
      return fail(RRT_ERR_INVALID, "params: rank/world out of range");
    if (cam->samples_per_pixel < 0 || cam->max_depth < 0) return fail(RRT_ERR_INVALID, "camera: negative spp or depth");
  }
  return RRT_OK;
}
using rrt::check_camera;

int rrt::owned_super_tiles(int n_super, int rank, int world) { return rank < n_super ? (n_super - rank + world - 1) / world : 0; }

// This rank's share of the frame: which 32x32 super tiles and which sample range.
// Pure host arithmetic, shared by rrt_render_device and rrt_shard_plan.
static void plan_work(const rrt_camera* cam, const rrt_render_params* p, DevWork& W, DevCamera& dc, double& my_pixels) {
  DevCamera64 c64;
  const int H = camera_initialize(cam, dc, c64);
  const int Wd = cam->image_width;
  W.width = Wd; W.height = H;
  W.integrator = p->integrator;
  W.max_depth = cam->max_depth;
  W.shard_mode = p->shard_mode;
  W.rank = p->shard_mode == RRT_SHARD_NONE ? 0 : p->rank;
  W.world = p->shard_mode == RRT_SHARD_NONE ? 1 : p->world;
  W.seed_lo = (unsigned)(p->seed & 0xffffffffu); W.seed_hi = (unsigned)(p->seed >> 32);
  philox_round_keys(W.seed_lo, W.seed_hi, W.rkeys, 10);
  int spp_eff = cam->samples_per_pixel;
  if (p->integrator == RRT_INT_BOOK3) spp_eff = dc.sqrt_spp * dc.sqrt_spp;  // camera.rs:114-116
  W.spp_eff = spp_eff;
  W.s_begin = 0; W.s_end = spp_eff;
  if (p->shard_mode == RRT_SHARD_SAMPLE) {
    W.s_begin = (int)(((long long)spp_eff * W.rank) / W.world);
    W.s_end = (int)(((long long)spp_eff * (W.rank + 1)) / W.world);
  }
  W.stiles_x = (Wd + 31) / 32; W.stiles_y = (H + 31) / 32;
  const int n_super = W.stiles_x * W.stiles_y;
  const int tw = p->shard_mode == RRT_SHARD_TILE ? W.world : 1, tr = p->shard_mode == RRT_SHARD_TILE ? W.rank : 0;
  const int my_super = rrt::owned_super_tiles(n_super, tr, tw);
  W.n_tiles = my_super * 32;
  W.tiles_x = W.stiles_x * 4;
  my_pixels = 0;
  for (int k = tr; k < n_super; k += tw) {
    int sx, sy;
    rrt::super_tile_xy(k, W.stiles_x, sx, sy);
    my_pixels += (double)std::min(32, Wd - sx * 32) * (double)std::min(32, H - sy * 32);
  }
}

int rrt::render_enqueue(rrt_ctx* c, const rrt_scene* sc, const rrt_camera* cam, const rrt_render_params* p, float* dst_dev,
                        bool compact, ShareInfo* info) {
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  RenderArgs A;
  std::memset(&A, 0, sizeof A);
  A.S = sc->dev;
  double my_pixels = 0;
  plan_work(cam, p, A.W, A.cam, my_pixels);
  DevWork& W = A.W;
  const int Wd = W.width, H = W.height;
  const int n_samples = W.s_end - W.s_begin;
  W.compact = compact && p->shard_mode == RRT_SHARD_TILE ? 1 : 0;

  const RenderVariant* variant = select_render_variant(p->integrator, sc->dev.features, c->force_full_kernel);
  // Sample stealing (render_kernel.cuh) pays wherever path cost varies inside a warp (C1 +20 %, C2/C3a
  // +3.5 %, C5 +1-6 %).  Not for one- and two-primitive scenes (near-constant path cost; the shared
  // counters cost 9 % on C3b) nor for the Perlin instances, which sit at their register limit (C4 -4 %).
  const bool steal = sc->n_prims > 2 && (!(variant->features & SF_NOISE) || c->force_steal) && !c->no_steal && !c->deterministic;
  const int schedule = c->use_pool && variant->kernel_pool && !c->deterministic ? RS_POOL : steal ? RS_STEAL : RS_OWN;
  int bps = render_blocks_per_sm(variant, schedule);
  if (bps < 1) return fail(RRT_ERR_CUDA, "kernel launch failed: render kernel does not fit on an SM");
  const int grid = c->sm_count * bps;
  const long long resident_warps = (long long)grid * (RRT_BLOCK / 32);
  long long chunks = 1;
  if (W.n_tiles > 0 && n_samples > 0) {
    // work items per resident warp / fewest samples per lane and item (measured: profiles/README.md):
    // with stealing a short item has no intra-warp tail, so items can be smaller
    long long want = ((steal ? 96 : 48) * resident_warps + W.n_tiles - 1) / W.n_tiles;
    // at least 16 samples per lane and item: with 8 (tried for the stealing instances) a share of 62 spp — what one of
    // 8 GPUs gets of C2 — ran 2.3 % slower, 125 spp 4 % (profiles/README.md r02x)
    long long cap = std::max(1, n_samples / 16);
    chunks = std::max(1LL, std::min(want, cap));
    // ... unless that leaves fewer than six items a resident warp (a small frame: C1 has 2 850 tiles for 4 736 warps and
    // lost 4 % to the end of the launch with 6 chunks of its 100 spp instead of 12): then 8 samples
    if (steal && W.n_tiles * chunks < 6 * resident_warps) chunks = std::max(1LL, std::min(want, (long long)std::max(1, n_samples / 8)));
    if (const char* e = std::getenv("RRT_CHUNKS")) chunks = std::max(1L, std::min((long)n_samples, std::atol(e)));   // A/B of the work-item size
    if (c->deterministic) chunks = 1;   // one float4 store per pixel, samples summed in index order by one lane
  }
  W.chunks = (int)chunks;
  W.atomic_accum = chunks > 1 ? 1 : 0;
  A.accum = reinterpret_cast<float4*>(dst_dev);
  A.counter = c->counter;

  CK(cudaEventRecord(c->ev[0], c->stream), "cudaEventRecord");
  CK(cudaMemsetAsync(c->counter, 0, sizeof(unsigned int), c->stream), "cudaMemsetAsync(counter)");
  const size_t dst_floats = W.compact ? (size_t)(W.n_tiles / 32) * 1024 * 4 : (size_t)Wd * H * 4;
  if (W.atomic_accum || (p->shard_mode == RRT_SHARD_TILE && !W.compact) || n_samples <= 0)
    CK(cudaMemsetAsync(dst_dev, 0, dst_floats * sizeof(float), c->stream), "cudaMemsetAsync(accum)");
  int launches = 0;
  CK(cudaEventRecord(c->ev[1], c->stream), "cudaEventRecord");
  if (W.n_tiles > 0 && n_samples > 0) {
    CK(launch_render(variant, schedule, A, grid, c->stream), "kernel launch failed");
    launches = 1;
  }
  CK(cudaEventRecord(c->ev[2], c->stream), "cudaEventRecord");
  if (info) {
    info->width = Wd; info->height = H; info->spp_eff = W.spp_eff; info->n_samples = std::max(0, n_samples);
    info->launches = launches; info->features = launches ? variant->features : -1; info->pixels = my_pixels;
    info->n_super = W.n_tiles / 32;
  }
  return RRT_OK;
}

void rrt::stats_from_share(rrt_ctx* c, const rrt_scene* sc, const ShareInfo& s) {
  float kms = 0, tms = 0;
  cudaEventElapsedTime(&kms, c->ev[1], c->ev[2]);
  cudaEventElapsedTime(&tms, c->ev[0], c->ev[2]);
  rrt_stats& st = c->stats;
  st.width = s.width; st.height = s.height; st.spp_effective = s.spp_eff;
  st.kernel_launches = s.launches;
  st.paths = s.pixels * (double)s.n_samples;
  st.kernel_ms = kms; st.total_ms = tms;
  st.exchange_ms = 0; st.n_devices = 1;
  st.d2h_bytes = 0;
  st.n_prims = sc->n_prims; st.n_bvh_nodes = sc->n_nodes; st.n_media = sc->n_media; st.n_lights = sc->n_lights;
  st.n_grid_cells = sc->n_grid_cells;
  st.sm_count = c->sm_count;
  st.kernel_features = s.features;
}

int rrt::ensure_floats(float** buf, size_t* have, size_t floats, const char* what) {
  if (*have >= floats) return RRT_OK;
  if (*buf) { cudaFree(*buf); *buf = nullptr; *have = 0; }
  CK(cudaMalloc(buf, std::max<size_t>(floats, 4) * sizeof(float)), what);
  *have = floats;
  return RRT_OK;
}

int rrt::scene_create(rrt_ctx* c, const rrt_scene_desc* desc, rrt_scene** out, bool tree_optional) {
  if (!c || !out) return fail(RRT_ERR_INVALID, "rrt_scene_create: null argument");
  *out = nullptr;
  FlatScene F;
  std::string err;
  int st = flatten_scene(desc, F, err, RRT_SINGLE_LEAF_MAX, 1, tree_optional);
  if (st != RRT_OK) return fail(st, "scene flatten failed: " + err);
  if (F.bvh_depth >= RRT_STACK)   // the traversal stack holds one entry per level (intersect.cuh)
    return fail(RRT_ERR_UNSUPPORTED, "BVH depth " + std::to_string(F.bvh_depth) + " exceeds the traversal stack (" + std::to_string(RRT_STACK) + ")");
  if (F.media.size() > 5) return fail(RRT_ERR_UNSUPPORTED, "more than 5 constant media");
  rrt_scene* sc = nullptr;
  int rc = c->members.empty() ? rrt::scene_upload(c, F, &sc) : rrt::multi_scene_create(c, F, &sc);
  if (rc) return rc;
  const rrt_scene* p0 = part0_of(sc);
  c->stats.h2d_bytes = sc->h2d_bytes;
  c->stats.n_prims = p0->n_prims; c->stats.n_bvh_nodes = p0->n_nodes;
  c->stats.n_media = p0->n_media; c->stats.n_lights = p0->n_lights;
  c->stats.n_grid_cells = p0->n_grid_cells;
  *out = sc;
  return RRT_OK;
}


extern "C" {

const char* rrt_last_error(void) { return g_err.c_str(); }
const char* rrt_version(void) { return "rrt-b200 0.1 (sm_100a)"; }

static int init_device_ctx(int device, rrt_ctx** out) {
  CK(cudaSetDevice(device), "cuda init failed");
  rrt_ctx* c = new rrt_ctx();
  std::memset(&c->stats, 0, sizeof c->stats);
  c->device = device;
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) { delete c; return fail(RRT_ERR_CUDA, std::string("cudaGetDeviceProperties: ") + cudaGetErrorString(e)); }
  if (prop.major < 10) {
    delete c;
    return fail(RRT_ERR_NO_DEVICE, std::string("device '") + prop.name + "' is not sm_100-class; kernels are built for sm_100a only");
  }
  c->sm_count = prop.multiProcessorCount;
  const char* kv = std::getenv("RRT_KERNEL");
  c->force_full_kernel = kv && std::strcmp(kv, "full") == 0;
  c->no_steal = std::getenv("RRT_NO_STEAL") != nullptr;
  c->force_steal = std::getenv("RRT_FORCE_STEAL") != nullptr;
  c->deterministic = std::getenv("RRT_DETERMINISTIC") != nullptr;
  c->use_pool = std::getenv("RRT_POOL") != nullptr;
  e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
  c->stream = c->own_stream;
  for (auto& ev : c->ev) if (e == cudaSuccess) e = cudaEventCreate(&ev);
  if (e == cudaSuccess) e = cudaMalloc(&c->counter, sizeof(unsigned int));
  if (e != cudaSuccess) { rrt_shutdown(c); return fail(RRT_ERR_CUDA, std::string("cuda init failed: ") + cudaGetErrorString(e)); }
  c->stats.sm_count = c->sm_count;
  c->stats.n_devices = 1;
  *out = c;
  return RRT_OK;
}

static int device_count(int* n) {
  *n = 0;
  cudaError_t e = cudaGetDeviceCount(n);
  if (e != cudaSuccess || *n == 0) {
    cudaGetLastError();
    return fail(RRT_ERR_NO_DEVICE, std::string("cuda init failed: no CUDA device (") +
                                       (e != cudaSuccess ? cudaGetErrorString(e) : "count 0") +
                                       "); this library has no CPU fallback");
  }
  return RRT_OK;
}

int rrt_init(int device, rrt_ctx** out) {
  if (!out) return fail(RRT_ERR_INVALID, "rrt_init: out is null");
  *out = nullptr;
  int n = 0, rc = device_count(&n);
  if (rc) return rc;
  if (device < 0 || device >= n) return fail(RRT_ERR_INVALID, "rrt_init: device index out of range");
  return init_device_ctx(device, out);
}

int rrt_init_multi(const int* devices, int n_devices, rrt_ctx** out) {
  if (!out || !devices) return fail(RRT_ERR_INVALID, "rrt_init_multi: null argument");
  *out = nullptr;
  if (n_devices < 1 || n_devices > RRT_MAX_WORLD) return fail(RRT_ERR_INVALID, "rrt_init_multi: between 1 and 64 devices");
  int n = 0, rc = device_count(&n);
  if (rc) return rc;
  for (int i = 0; i < n_devices; ++i) {
    if (devices[i] < 0 || devices[i] >= n) return fail(RRT_ERR_INVALID, "rrt_init_multi: device index out of range");
    for (int j = 0; j < i; ++j) if (devices[j] == devices[i]) return fail(RRT_ERR_INVALID, "rrt_init_multi: device listed twice");
  }
  rrt_ctx* g = new rrt_ctx();
  std::memset(&g->stats, 0, sizeof g->stats);
  g->device = devices[0];
  for (int i = 0; i < n_devices; ++i) {
    rrt_ctx* m = nullptr;
    if ((rc = init_device_ctx(devices[i], &m))) { rrt_shutdown(g); return rc; }
    m->group = g;
    g->members.push_back(m);
  }
  if ((rc = rrt::multi_comm_init_all(g, devices))) { rrt_shutdown(g); return rc; }
  g->sm_count = g->members[0]->sm_count;
  g->stats.sm_count = g->sm_count;
  g->stats.n_devices = n_devices;
  *out = g;
  return RRT_OK;
}

void rrt_shutdown(rrt_ctx* c) {
  if (!c) return;
  if (!c->members.empty() || c->group == nullptr) {
    // scenes the caller still holds: free their device memory now and orphan the handles, so a late
    // rrt_scene_destroy (e.g. from a garbage collector) stays harmless
    for (rrt_scene* sc : c->scenes) rrt::scene_release(sc);
    c->scenes.clear();
  }
  if (!c->members.empty()) {  // a group: its members hold the devices
    for (rrt_ctx* m : c->members) { m->group = nullptr; rrt_shutdown(m); }
    c->members.clear();
    delete c;
    return;
  }
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  rrt::comm_release(c);
  for (rrt_scene* sc : c->scenes) rrt::scene_release(sc);
  c->scenes.clear();
  if (c->accum) cudaFree(c->accum);
  if (c->share) cudaFree(c->share);
  if (c->staging) cudaFree(c->staging);
  if (c->counter) cudaFree(c->counter);
  for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
}

int rrt_set_stream(rrt_ctx* c, void* stream) {
  if (!c) return fail(RRT_ERR_INVALID, "rrt_set_stream: ctx is null");
  if (!c->members.empty()) return fail(RRT_ERR_INVALID, "rrt_set_stream: a multi-device context launches on its own streams");
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  c->stream = stream ? reinterpret_cast<cudaStream_t>(stream) : c->own_stream;
  return RRT_OK;
}

void rrt_scene_destroy(rrt_scene* sc) {
  if (!sc) return;
  if (rrt_ctx* c = sc->ctx) {
    if (c->members.empty()) {
      cudaSetDevice(c->device);
      cudaStreamSynchronize(c->stream);
    }
    for (size_t i = 0; i < c->scenes.size(); ++i)
      if (c->scenes[i] == sc) { c->scenes.erase(c->scenes.begin() + i); break; }
  }
  rrt::scene_release(sc);
  for (rrt_scene* part : sc->parts) delete part;
  delete sc;
}

int rrt_scene_create(rrt_ctx* c, const rrt_scene_desc* desc, rrt_scene** out) { return rrt::scene_create(c, desc, out, false); }

int rrt_image_height(const rrt_camera* cam) {
  if (!cam || !(cam->aspect_ratio > 0.0)) return 0;
  int h = (int)((double)cam->image_width / cam->aspect_ratio);
  return h < 1 ? 1 : h;
}

int rrt_render_device(rrt_ctx* c, const rrt_scene* sc, const rrt_camera* cam, const rrt_render_params* p,
                      float* accum_dev) {
  if (!c || !sc || !p || !accum_dev) return fail(RRT_ERR_INVALID, "rrt_render_device: null argument");
  if (sc->ctx != c) return fail(RRT_ERR_INVALID, "scene belongs to another (or a shut down) context");
  int rc = check_camera(cam, p);
  if (rc) return rc;
  if (!c->members.empty()) return rrt::multi_render(c, sc, cam, p, accum_dev, nullptr);
  rrt::ShareInfo info;
  if ((rc = rrt::render_enqueue(c, sc, cam, p, accum_dev, false, &info))) return rc;
  CK(cudaStreamSynchronize(c->stream), "render kernel failed");
  rrt::stats_from_share(c, sc, info);
  return RRT_OK;
}

int rrt_render(rrt_ctx* c, const rrt_scene* sc, const rrt_camera* cam, const rrt_render_params* p, float* accum_host) {
  if (!c || !sc || !p || !accum_host) return fail(RRT_ERR_INVALID, "rrt_render: null argument");
  if (sc->ctx != c) return fail(RRT_ERR_INVALID, "scene belongs to another (or a shut down) context");
  int rc = check_camera(cam, p);
  if (rc) return rc;
  if (!c->members.empty()) return rrt::multi_render(c, sc, cam, p, nullptr, accum_host);
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  const int H = rrt_image_height(cam);
  const size_t floats = (size_t)cam->image_width * H * 4;
  if ((rc = rrt::ensure_floats(&c->accum, &c->accum_floats, floats, "cudaMalloc(accum)"))) return rc;
  if ((rc = rrt_render_device(c, sc, cam, p, c->accum))) return rc;
  CK(cudaMemcpyAsync(accum_host, c->accum, floats * sizeof(float), cudaMemcpyDeviceToHost, c->stream), "accum read-back");
  CK(cudaEventRecord(c->ev[3], c->stream), "cudaEventRecord");
  CK(cudaStreamSynchronize(c->stream), "accum read-back");
  float tms = 0;
  cudaEventElapsedTime(&tms, c->ev[0], c->ev[3]);
  c->stats.total_ms = tms;
  c->stats.d2h_bytes = floats * sizeof(float);
  return RRT_OK;
}

int rrt_render_frame(rrt_ctx* c, const rrt_scene_desc* desc, const rrt_camera* cam, const rrt_render_params* p,
                     float* accum_host) {
  rrt_scene* sc = nullptr;
  int rc = rrt::scene_create(c, desc, &sc, true);   // rendered once and freed: no tree where grids do the walking
  if (rc) return rc;
  rc = rrt_render(c, sc, cam, p, accum_host);
  uint64_t h2d = sc->h2d_bytes;
  rrt_scene_destroy(sc);
  if (rc == RRT_OK) c->stats.h2d_bytes = h2d;
  return rc;
}

int rrt_probe(rrt_ctx* c, const rrt_scene* sc, const rrt_camera* cam, int precision, int32_t* prim_id, double* t) {
  if (!c || !sc || !prim_id || !t) return fail(RRT_ERR_INVALID, "rrt_probe: null argument");
  if (precision != 32 && precision != 64) return fail(RRT_ERR_INVALID, "rrt_probe: precision must be 32 or 64");
  if (sc->ctx != c) return fail(RRT_ERR_INVALID, "scene belongs to another (or a shut down) context");
  if (!c->members.empty()) {   // one ray per pixel: the first device of a group does it alone
    int rc0 = rrt_probe(root_of(c), part0_of(sc), cam, precision, prim_id, t);
    c->stats.kernel_launches = 1;
    return rc0;
  }
  int rc = check_camera(cam, nullptr);
  if (rc) return rc;
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  ProbeArgs A;
  std::memset(&A, 0, sizeof A);
  A.S = sc->dev;
  A.height = camera_initialize(cam, A.cam, A.cam64);
  A.width = cam->image_width;
  size_t n = (size_t)A.width * A.height;
  CK(cudaMalloc(&A.prim_id, n * sizeof(int)), "cudaMalloc(probe)");
  cudaError_t e = cudaMalloc(&A.t, n * sizeof(double));
  if (e != cudaSuccess) { cudaFree(A.prim_id); return fail(RRT_ERR_CUDA, std::string("cudaMalloc(probe): ") + cudaGetErrorString(e)); }
  e = launch_probe(A, precision, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(prim_id, A.prim_id, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(t, A.t, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFree(A.prim_id); cudaFree(A.t);
  if (e != cudaSuccess) return fail(RRT_ERR_CUDA, std::string("probe kernel failed: ") + cudaGetErrorString(e));
  c->stats.kernel_launches = 1;
  c->stats.width = A.width; c->stats.height = A.height;
  return RRT_OK;
}

int rrt_resolve_rgb8_device(rrt_ctx* c, const float* accum_dev, int width, int height, int spp, uint8_t* rgb_dev) {
  if (!c || !accum_dev || !rgb_dev || width < 1 || height < 1) return fail(RRT_ERR_INVALID, "rrt_resolve_rgb8_device: bad argument");
  if (!c->members.empty()) return rrt_resolve_rgb8_device(root_of(c), accum_dev, width, height, spp, rgb_dev);
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  CK(launch_resolve(accum_dev, width * height, spp, rgb_dev, c->stream), "kernel launch failed");
  CK(cudaStreamSynchronize(c->stream), "resolve kernel failed");
  c->stats.kernel_launches = 1;
  return RRT_OK;
}

int rrt_resolve_rgb8(rrt_ctx* c, const float* accum_host, int width, int height, int spp, uint8_t* rgb_host) {
  if (!c || !accum_host || !rgb_host || width < 1 || height < 1) return fail(RRT_ERR_INVALID, "rrt_resolve_rgb8: bad argument");
  if (!c->members.empty()) return rrt_resolve_rgb8(root_of(c), accum_host, width, height, spp, rgb_host);
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  size_t n = (size_t)width * height;
  int rc = rrt::ensure_floats(&c->accum, &c->accum_floats, n * 4, "cudaMalloc(accum)");
  if (rc) return rc;
  unsigned char* rgb = nullptr;
  CK(cudaMalloc(&rgb, n * 3), "cudaMalloc(rgb)");
  cudaError_t e = cudaMemcpyAsync(c->accum, accum_host, n * 16, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = launch_resolve(c->accum, (int)n, spp, rgb, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(rgb_host, rgb, n * 3, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFree(rgb);
  if (e != cudaSuccess) return fail(RRT_ERR_CUDA, std::string("resolve failed: ") + cudaGetErrorString(e));
  c->stats.kernel_launches = 1;
  return RRT_OK;
}

int rrt_shard_plan(const rrt_camera* cam, const rrt_render_params* p, int32_t* sample_begin, int32_t* sample_end,
                   int32_t* n_super_tiles, double* n_pixels, uint8_t* owner_mask) {
  if (!p) return fail(RRT_ERR_INVALID, "rrt_shard_plan: params is null");
  int rc = check_camera(cam, p);
  if (rc) return rc;
  DevWork W; DevCamera dc; double px = 0;
  std::memset(&W, 0, sizeof W);
  plan_work(cam, p, W, dc, px);
  if (sample_begin) *sample_begin = W.s_begin;
  if (sample_end) *sample_end = W.s_end;
  if (n_super_tiles) *n_super_tiles = W.n_tiles / 32;
  if (n_pixels) *n_pixels = px;
  if (owner_mask) {  // W*H bytes: 1 where this rank writes the pixel
    const bool tile = W.shard_mode == RRT_SHARD_TILE;
    for (int y = 0; y < W.height; ++y)
      for (int x = 0; x < W.width; ++x) {
        int k = rrt::super_tile_index(x / 32, y / 32, W.stiles_x);
        owner_mask[(size_t)y * W.width + x] = (!tile || k % W.world == W.rank) ? 1 : 0;
      }
  }
  return RRT_OK;
}

int rrt_get_stats(const rrt_ctx* c, rrt_stats* out) {
  if (!c || !out) return fail(RRT_ERR_INVALID, "rrt_get_stats: null argument");
  *out = c->stats;
  return RRT_OK;
}

// Known-answer hook for the sampling RNG (tests): Philox4x32-10 block.
void rrt_philox_kat(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out4) {
  philox4x32_10(c0, c1, c2, c3, k0, k1, out4);
}

// The same block at any round count Random123 publishes known answers for (7 and 10); 0 = unsupported.
int rrt_philox_kat_rounds(int rounds, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out4) {
  if (!out4) return fail(RRT_ERR_INVALID, "rrt_philox_kat_rounds: out4 is null");
  if (rounds == 7) philox4x32<7>(c0, c1, c2, c3, k0, k1, out4);
  else if (rounds == 10) philox4x32<10>(c0, c1, c2, c3, k0, k1, out4);
  else return fail(RRT_ERR_INVALID, "rrt_philox_kat_rounds: rounds must be 7 or 10");
  return RRT_OK;
}
int rrt_philox_rounds(void) { return RRT_PHILOX_ROUNDS; }

// Device-side known-answer hooks (tests): run the kernels' own light-sampling and RNG code.
int rrt_kat_light(rrt_ctx* c, const rrt_scene* sc, int n, const double* origin, const double* dir, const double* u,
                  double* pdf_each, double* pdf_list, double* random_dir) {
  if (!c || !sc || n < 0 || !origin || !dir || !u || !pdf_each || !pdf_list || !random_dir)
    return fail(RRT_ERR_INVALID, "rrt_kat_light: null argument");
  if (sc->ctx != c) return fail(RRT_ERR_INVALID, "scene belongs to another (or a shut down) context");
  if (!c->members.empty()) return rrt_kat_light(root_of(c), part0_of(sc), n, origin, dir, u, pdf_each, pdf_list, random_dir);
  if (n == 0) return RRT_OK;
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  const int nl = sc->n_lights;
  const size_t n_in = (size_t)n * 8, n_out = (size_t)n * ((size_t)nl * 4 + 1);
  std::vector<float> hin(n_in), hout(n_out);
  for (int i = 0; i < n; ++i) {
    for (int k = 0; k < 3; ++k) { hin[3 * (size_t)i + k] = (float)origin[3 * (size_t)i + k]; hin[3 * (size_t)n + 3 * (size_t)i + k] = (float)dir[3 * (size_t)i + k]; }
    hin[6 * (size_t)n + 2 * (size_t)i] = (float)u[2 * (size_t)i]; hin[6 * (size_t)n + 2 * (size_t)i + 1] = (float)u[2 * (size_t)i + 1];
  }
  float* din = nullptr; float* dout = nullptr;
  CK(cudaMalloc(&din, n_in * sizeof(float)), "cudaMalloc(kat)");
  cudaError_t e = cudaMalloc(&dout, std::max<size_t>(1, n_out) * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpyAsync(din, hin.data(), n_in * sizeof(float), cudaMemcpyHostToDevice, c->stream);
  KatLightArgs A;
  std::memset(&A, 0, sizeof A);
  A.S = sc->dev; A.n = n;
  A.origin = din; A.dir = din + 3 * (size_t)n; A.u = din + 6 * (size_t)n;
  A.pdf_each = dout; A.pdf_list = dout + (size_t)n * nl; A.random_dir = dout + (size_t)n * nl + n;
  if (e == cudaSuccess) e = launch_kat_light(A, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(hout.data(), dout, n_out * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFree(din); if (dout) cudaFree(dout);
  if (e != cudaSuccess) return fail(RRT_ERR_CUDA, std::string("kat kernel failed: ") + cudaGetErrorString(e));
  for (size_t i = 0; i < (size_t)n * nl; ++i) pdf_each[i] = hout[i];
  for (size_t i = 0; i < (size_t)n; ++i) pdf_list[i] = hout[(size_t)n * nl + i];
  for (size_t i = 0; i < (size_t)n * nl * 3; ++i) random_dir[i] = hout[(size_t)n * nl + n + i];
  c->stats.kernel_launches = 1;
  return RRT_OK;
}

int rrt_kat_rng(rrt_ctx* c, uint64_t seed, uint32_t pixel, uint32_t sample0, uint32_t bounce, uint32_t block, int n,
                uint32_t* raw4, float* uniform5) {
  if (!c || n < 0 || (!raw4 && !uniform5)) return fail(RRT_ERR_INVALID, "rrt_kat_rng: bad argument");
  if (!c->members.empty()) return rrt_kat_rng(root_of(c), seed, pixel, sample0, bounce, block, n, raw4, uniform5);
  if (n == 0) return RRT_OK;
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  uint32_t* draw = nullptr; float* du = nullptr;
  cudaError_t e = cudaSuccess;
  if (raw4) e = cudaMalloc(&draw, (size_t)n * 16);
  if (e == cudaSuccess && uniform5) e = cudaMalloc(&du, (size_t)n * 20);
  if (e == cudaSuccess) e = launch_kat_rng((unsigned)(seed & 0xffffffffu), (unsigned)(seed >> 32), pixel, sample0, bounce, block, n, draw, du, c->stream);
  if (e == cudaSuccess && raw4) e = cudaMemcpyAsync(raw4, draw, (size_t)n * 16, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess && uniform5) e = cudaMemcpyAsync(uniform5, du, (size_t)n * 20, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (draw) cudaFree(draw);
  if (du) cudaFree(du);
  if (e != cudaSuccess) return fail(RRT_ERR_CUDA, std::string("kat kernel failed: ") + cudaGetErrorString(e));
  c->stats.kernel_launches = 1;
  return RRT_OK;
}

}  // extern "C"

// ---- host-only verification of the flattener (CPU test tier) --------------------------------
namespace {
struct Box { double lo[3], hi[3]; };
bool inside(const Box& c, const Box& p) {
  // boxes are padded and rounded independently, so allow the paddings to differ by a few ulps
  for (int a = 0; a < 3; ++a) {
    double tol = 2e-6 * std::max(1e-3, std::max(std::fabs(p.lo[a]), std::fabs(p.hi[a])));
    if (c.lo[a] < p.lo[a] - tol || c.hi[a] > p.hi[a] + tol) return false;
  }
  return true;
}
int as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
}  // namespace

extern "C" int rrt_scene_features(const rrt_scene_desc* desc, int32_t* features) {
  if (!features) return fail(RRT_ERR_INVALID, "rrt_scene_features: features is null");
  FlatScene F;
  std::string err;
  int st = flatten_scene(desc, F, err);
  if (st != RRT_OK) return fail(st, err);
  *features = F.features;
  return RRT_OK;
}

extern "C" int rrt_flatten_grid(const rrt_scene_desc* desc, int kind, int32_t res3[3], int32_t* n_items, int32_t* n_refs,
                                int32_t* max_per_cell) {
  if (kind < 0 || kind >= RRT_MAX_GRIDS) return fail(RRT_ERR_INVALID, "rrt_flatten_grid: kind must be 0 (spheres) or 1 (boxes)");
  FlatScene F;
  std::string err;
  int st = flatten_scene(desc, F, err);
  if (st != RRT_OK) return fail(st, "scene flatten failed: " + err);
  const bool on = (size_t)kind < F.grids.size() && F.grids[kind].n_items > 0;
  int most = 0, refs = 0;
  if (on) {
    const DevGrid& G = F.grids[kind];
    const size_t n_cells = (size_t)G.res[0] * G.res[1] * G.res[2];
    for (size_t c = 0; c < n_cells; ++c) { const int k = (int)(F.grid_cells[G.first_cell + c] & 255u); most = std::max(most, k); refs += k; }
  }
  if (res3) for (int a = 0; a < 3; ++a) res3[a] = on ? F.grids[kind].res[a] : 0;
  if (n_items) *n_items = on ? F.grids[kind].n_items : 0;
  if (n_refs) *n_refs = refs;
  if (max_per_cell) *max_per_cell = most;
  return RRT_OK;
}

extern "C" int rrt_flatten_rects(const rrt_scene_desc* desc, int32_t n3[3]) {
  FlatScene F;
  std::string err;
  int st = flatten_scene(desc, F, err);
  if (st != RRT_OK) return fail(st, "scene flatten failed: " + err);
  if (n3) for (int a = 0; a < 3; ++a) n3[a] = F.n_rects[a];
  return RRT_OK;
}

extern "C" int rrt_flatten_check(const rrt_scene_desc* desc, int32_t* n_prims, int32_t* n_nodes, int32_t* bvh_depth,
                                 int32_t* n_media, int32_t* n_lights) {
  FlatScene F;
  std::string err;
  int st = flatten_scene(desc, F, err);
  if (st != RRT_OK) return fail(st, "scene flatten failed: " + err);
  const int nn = (int)(F.nodes.size() / 4);
  std::vector<int> seen(F.n_prims, 0);
  std::vector<int> node_seen(nn, 0);
  auto mark_leaf = [&](int ref, bool is_root) -> bool {
    int leaf = ~ref, first = leaf >> 8, ns = (leaf >> 4) & 15, nq = leaf & 15;
    const bool is_box = (leaf & 0xFF) == LEAF_BOX_MARK;
    if (!is_root && !is_box && ns + nq != 1) return false;  // the traversal loop tests exactly one primitive per leaf
    if (is_box) {  // box leaf: six quads named by the box record
      if (first < 0 || 2 * first + 1 >= (int)F.boxes.size()) return false;
      first = as_int(F.boxes[2 * first].w); ns = 0; nq = 6;
    }
    if (first < 0 || first + ns + nq > F.n_prims) return false;
    for (int k = first; k < first + ns + nq; ++k) seen[k]++;
    return true;
  };
  if (!mark_leaf(F.front_ref, true)) return fail(RRT_ERR_INVALID, "flatten check: front run out of range");
  if (F.root_ref < 0) {
    if (!mark_leaf(F.root_ref, true)) return fail(RRT_ERR_INVALID, "flatten check: root leaf out of range");
  } else {
    struct Item { int node; Box box; bool has_box; };
    std::vector<Item> stack;
    Item root; root.node = 0; root.has_box = false;
    stack.push_back(root);
    while (!stack.empty()) {
      Item it = stack.back(); stack.pop_back();
      if (it.node < 0 || it.node >= nn) return fail(RRT_ERR_INVALID, "flatten check: node index out of range");
      if (node_seen[it.node]++) return fail(RRT_ERR_INVALID, "flatten check: node reached twice");
      const float4* n = &F.nodes[4 * it.node];
      const float* f = reinterpret_cast<const float*>(n);  // device_types.h: f[2a+k] centre, f[6+2a+k] half-extent
      Box c0, c1;
      for (int a = 0; a < 3; ++a) {
        c0.lo[a] = (double)f[2 * a] - f[6 + 2 * a];         c0.hi[a] = (double)f[2 * a] + f[6 + 2 * a];
        c1.lo[a] = (double)f[2 * a + 1] - f[6 + 2 * a + 1]; c1.hi[a] = (double)f[2 * a + 1] + f[6 + 2 * a + 1];
      }
      if (it.has_box && (!inside(c0, it.box) || !inside(c1, it.box)))
        return fail(RRT_ERR_INVALID, "flatten check: child box escapes its parent");
      int refs[2] = {as_int(n[3].x), as_int(n[3].y)};
      Box bx[2] = {c0, c1};
      for (int k = 0; k < 2; ++k) {
        if (refs[k] >= 0) { Item ch; ch.node = refs[k]; ch.box = bx[k]; ch.has_box = true; stack.push_back(ch); }
        else if (!mark_leaf(refs[k], false)) return fail(RRT_ERR_INVALID, "flatten check: leaf out of range or not a single primitive");
      }
    }
    for (int i = 0; i < nn; ++i) if (node_seen[i] != 1) return fail(RRT_ERR_INVALID, "flatten check: unreachable node");
  }
  for (int i = 0; i < F.n_prims; ++i)
    if (seen[i] != 1) return fail(RRT_ERR_INVALID, "flatten check: primitive not in exactly one leaf");
  // rectangles of a root leaf: its first quads, each record on the plane of its quad
  {
    const int nr = F.n_rects[0] + F.n_rects[1] + F.n_rects[2];
    if (nr > 0) {
      const int leaf = ~F.root_ref, first = (leaf >> 8) + ((leaf >> 4) & 15), nq = leaf & 15;
      if (F.root_ref >= 0 || nr > nq || (int)F.rects.size() != 2 * nr) return fail(RRT_ERR_INVALID, "flatten check: rectangle records do not match the root leaf");
      int k = 0;
      for (int a = 0; a < 3; ++a)
        for (int j = 0; j < F.n_rects[a]; ++j, ++k) {
          const float4 p0 = F.prims[4 * (first + k)];
          const float n[3] = {p0.x, p0.y, p0.z};
          const bool axis_ok = std::fabs(std::fabs(n[a]) - 1.0f) < 1e-6f && n[(a + 1) % 3] == 0.0f && n[(a + 2) % 3] == 0.0f;
          if (!axis_ok || std::fabs(n[a] * p0.w - F.rects[2 * k].x) > 1e-5f * (1.0f + std::fabs(p0.w)))
            return fail(RRT_ERR_INVALID, "flatten check: rectangle record off its quad's plane");
        }
    }
  }
  // grids: every reference in range, every item referenced from every cell its bounds overlap (so that a walk
  // through the cells cannot miss it), and — the render kernel walks no tree then — every primitive outside
  // the front run in exactly one grid
  if (!F.grids.empty()) {
    std::vector<int> in_grid(F.n_prims, 0);
    { const int lf = ~F.front_ref, first = lf >> 8, n = ((lf >> 4) & 15) + (lf & 15); for (int k = 0; k < n; ++k) in_grid[first + k] = -1000; }
    const int n_box_recs = (int)(F.boxes.size() / 2);
    for (size_t g = 0; g < F.grids.size(); ++g) {
      const DevGrid& G = F.grids[g];
      if (G.n_items <= 0) continue;
      const size_t n_cells = (size_t)G.res[0] * G.res[1] * G.res[2];
      if (G.res[0] < 1 || G.res[1] < 1 || G.res[2] < 1 || G.first_cell < 0 || (size_t)G.first_cell + n_cells > F.grid_cells.size())
        return fail(RRT_ERR_INVALID, "flatten check: grid resolution does not match its cell table");
      const int n_recs = G.kind == 1 ? n_box_recs : F.n_prims;
      std::vector<char> member(n_recs, 0);
      for (size_t c = 0; c < n_cells; ++c) {
        const unsigned first = F.grid_cells[G.first_cell + c] >> 8, cnt = F.grid_cells[G.first_cell + c] & 255u;
        if ((size_t)first + cnt > F.grid_refs.size()) return fail(RRT_ERR_INVALID, "flatten check: grid cell run out of range");
        for (unsigned k = 0; k < cnt; ++k) {
          const int p = F.grid_refs[first + k];
          if (p < 0 || p >= n_recs) return fail(RRT_ERR_INVALID, "flatten check: grid reference out of range");
          member[p] = 1;
        }
      }
      int n_members = 0;
      for (int p = 0; p < n_recs; ++p) {
        if (!member[p]) continue;
        ++n_members;
        float lo[3], hi[3];
        if (G.kind == 1) {   // a box record and its six quads
          const float4 b0 = F.boxes[2 * p], b1 = F.boxes[2 * p + 1];
          lo[0] = b0.x; lo[1] = b0.y; lo[2] = b0.z; hi[0] = b1.x; hi[1] = b1.y; hi[2] = b1.z;
          // the record is padded outwards and boxes that tile an axis own exactly their cell: test the box shrunk a little
          for (int a = 0; a < 3; ++a) { const float e = 1e-3f * (hi[a] - lo[a]); lo[a] += e; hi[a] -= e; }
          const int first = as_int(b0.w);
          if (first < 0 || first + 6 > F.n_prims) return fail(RRT_ERR_INVALID, "flatten check: box record out of range");
          for (int k = 0; k < 6; ++k) in_grid[first + k]++;
        } else {             // a sphere's swept box, the f32 way the kernel sees it
          const float4 p0 = F.prims[4 * p], p1 = F.prims[4 * p + 1];
          const float c[3] = {p0.x, p0.y, p0.z}, d[3] = {p1.x, p1.y, p1.z};
          for (int a = 0; a < 3; ++a) { lo[a] = std::min(c[a], c[a] + d[a]) - std::fabs(p0.w); hi[a] = std::max(c[a], c[a] + d[a]) + std::fabs(p0.w); }
          in_grid[p]++;
        }
        int c0[3], c1[3];
        for (int a = 0; a < 3; ++a) {
          if (G.kind == 0 && (lo[a] < G.lo[a] || hi[a] > G.hi[a])) return fail(RRT_ERR_INVALID, "flatten check: sphere outside its grid");
          c0[a] = std::max(0, std::min(G.res[a] - 1, (int)((lo[a] - G.lo[a]) * G.inv[a])));
          c1[a] = std::max(0, std::min(G.res[a] - 1, (int)((hi[a] - G.lo[a]) * G.inv[a])));
        }
        for (int z = c0[2]; z <= c1[2]; ++z) for (int y = c0[1]; y <= c1[1]; ++y) for (int x = c0[0]; x <= c1[0]; ++x) {
          if (G.kind == 0) {   // a sphere has to be listed only where the sphere itself reaches (here: 99.99 % of its radius)
            const float4 p0 = F.prims[4 * p], p1 = F.prims[4 * p + 1];
            const float c[3] = {p0.x, p0.y, p0.z}, dc[3] = {p1.x, p1.y, p1.z};
            const int ci[3] = {x, y, z};
            double d2 = 0;
            for (int a = 0; a < 3; ++a) {   // from the box of the centre's path, per axis
              const double lo2 = G.lo[a] + ci[a] * (double)G.cell[a], hi2 = lo2 + G.cell[a];
              const double ca = std::min(c[a], c[a] + dc[a]), cb = std::max(c[a], c[a] + dc[a]);
              const double d = cb < lo2 ? lo2 - cb : (ca > hi2 ? ca - hi2 : 0.0);
              d2 += d * d;
            }
            const double rr = 0.9999 * std::fabs(p0.w);
            if (d2 > rr * rr) continue;
          }
          const unsigned h = F.grid_cells[G.first_cell + ((size_t)z * G.res[1] + y) * G.res[0] + x];
          bool found = false;
          for (unsigned k = 0; k < (h & 255u); ++k) found |= F.grid_refs[(h >> 8) + k] == p;
          if (!found) return fail(RRT_ERR_INVALID, "flatten check: item missing from a grid cell it overlaps");
        }
      }
      if (n_members != G.n_items) return fail(RRT_ERR_INVALID, "flatten check: grid member count");
    }
    for (int p = 0; p < F.n_prims; ++p)
      if (in_grid[p] != 1 && in_grid[p] != -1000) return fail(RRT_ERR_INVALID, "flatten check: primitive neither in the front run nor in exactly one grid");
  }
  if (n_prims) *n_prims = F.n_prims;
  if (n_nodes) *n_nodes = nn;
  if (bvh_depth) *bvh_depth = F.bvh_depth;
  if (n_media) *n_media = (int)F.media.size();
  if (n_lights) *n_lights = (int)F.lights.size();
  return RRT_OK;
}
