// noalloc.cu — rrt_probe_static (include/rrt.h): rrt_probe for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
//
// rrt_probe (api.cu) stages the probe kernel's per-pixel outputs in two device buffers it allocates and
// frees on every call.  Here that scratch is a file-scope device array of fixed size: 12 bytes per pixel
// (f64 t + i32 id) for at most RRT_PROBE_STATIC_MAX_PIXELS pixels = 12 MiB, held by every GPU that loads
// librrt.so until the process exits.  No counterpart in the reference.
// Nothing in this file (and nothing it calls) allocates or frees device memory.
//
// The array's name in HOST code is a host-side stand-in, not a device address.  The device address is
// looked up once per device with cudaGetSymbolAddress, with that device current (scratch_of below), and
// that pointer is what the kernel gets.
//
// Sharing: every call on a device uses the same array, and the array keeps the last call's contents.
// The probe kernels write every pixel of the W x H frame before anything is read back and read nothing
// from the array, so nothing has to be cleared; pixels beyond W*H are stale and never copied out.  Two
// calls that overlap on one GPU (two host threads, two streams) race on the array: serialise them.
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>
#include <string>

#include "../../include/rrt.h"
#include "ctx.h"

using namespace rrt;

namespace {

constexpr size_t kMaxPixels = RRT_PROBE_STATIC_MAX_PIXELS;
// layout: kMaxPixels doubles (t), then kMaxPixels ints (prim_id)
static __device__ __align__(256) unsigned char g_probe_scratch[kMaxPixels * (sizeof(double) + sizeof(int))];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_of[kMaxDevices] = {nullptr};

// Device address of the scratch array on `device`; the lookup runs once per device, with it current.
int scratch_of(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxDevices) return fail(RRT_ERR_INVALID, "rrt_probe_static: device index out of range");
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_of[device]) {
    CK(cudaSetDevice(device), "cudaSetDevice");
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_probe_scratch), "cudaGetSymbolAddress(probe scratch)");
    if (!p) return fail(RRT_ERR_CUDA, "cudaGetSymbolAddress(probe scratch) returned null");
    g_scratch_of[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_of[device];
  return RRT_OK;
}

}  // namespace

extern "C" int rrt_probe_static(rrt_ctx* c, const rrt_scene* sc, const rrt_camera* cam, int precision, int32_t* prim_id, double* t) {
  if (!c || !sc || !prim_id || !t) return fail(RRT_ERR_INVALID, "rrt_probe_static: null argument");
  if (precision != 32 && precision != 64) return fail(RRT_ERR_INVALID, "rrt_probe_static: precision must be 32 or 64");
  if (sc->ctx != c) return fail(RRT_ERR_INVALID, "scene belongs to another (or a shut down) context");
  rrt_ctx* group = nullptr;
  if (!c->members.empty()) {   // a group: its first device does it alone, like rrt_probe
    if (sc->parts.empty()) return fail(RRT_ERR_INVALID, "rrt_probe_static: scene has no copy on the group's first device");
    group = c; c = c->members[0]; sc = sc->parts[0];
  }
  int rc = check_camera(cam, nullptr);
  if (rc) return rc;
  ProbeArgs A;
  std::memset(&A, 0, sizeof A);
  A.S = sc->dev;
  A.height = camera_initialize(cam, A.cam, A.cam64);
  A.width = cam->image_width;
  const size_t n = (size_t)A.width * (size_t)A.height;
  if (n > kMaxPixels)
    return fail(RRT_ERR_UNSUPPORTED, "rrt_probe_static: frame of " + std::to_string(n) + " pixels exceeds the static scratch of " +
                                         std::to_string(kMaxPixels) + " pixels (RRT_PROBE_STATIC_MAX_PIXELS); use rrt_probe");
  unsigned char* scratch = nullptr;
  rc = scratch_of(c->device, &scratch);
  if (rc) return rc;
  CK(cudaSetDevice(c->device), "cudaSetDevice");
  A.t = reinterpret_cast<double*>(scratch);
  A.prim_id = reinterpret_cast<int*>(scratch + kMaxPixels * sizeof(double));
  cudaError_t e = launch_probe(A, precision, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(prim_id, A.prim_id, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(t, A.t, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return fail(RRT_ERR_CUDA, std::string("probe kernel failed: ") + cudaGetErrorString(e));
  c->stats.kernel_launches = 1;
  c->stats.width = A.width; c->stats.height = A.height;
  if (group) { group->stats.kernel_launches = 1; group->stats.width = A.width; group->stats.height = A.height; }
  return RRT_OK;
}
