// render_kernel.cuh — the path-tracing megakernel template (instantiated by kernels_book{1,2,3}.cu).
//
//   render_kernel<INTEGRATOR, FEAT>  persistent-thread megakernel: camera/defocus ray generation,
//                              BVH traversal, sphere/quad/medium intersection, material scatter,
//                              book-3 mixture-PDF light sampling, Philox RNG, float4 accumulation.
//                              = Camera::render + get_ray + ray_color of the three books
//                              (in_one_weekend/camera.rs:59-209, the_next_week/camera.rs:61-201,
//                               the_rest_of_your_life/camera.rs:63-254).
//
// FEAT is the set of scene features (device_types.h SceneFeature) the instance supports: moving
// spheres, quads, constant media, non-solid textures.  A scene runs on the smallest instance that
// covers it, so the book-1 scenes (static spheres, solid colours) do not carry the registers and
// branches of the book-2 final scene: 60 registers and 8 CTAs/SM instead of 80 and 6 (+18 % on C2).
//
// Scheduling: a warp owns an 8x4-pixel tile and a chunk of the sample range; lane = pixel.  When a
// lane's path ends it regenerates the next sample of its own pixel immediately (no warp-wide wait
// for the longest path), radiance sums stay in registers and are written once per work item as one
// float4.  Work items are pulled from a global atomic counter, so the grid is exactly
// SMs x resident blocks.  The recursion `emitted + attenuation * ray_color(...)` of the reference
// is unrolled into a throughput/radiance loop.  BVH nodes and primitives are read through L1
// (ld.global.nc): a shared-memory copy of the node array measured 8-30 % slower.
#pragma once
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

__device__ inline V3f reflect(V3f v, V3f n) { return v - (2.0f * dot(v, n)) * n; }  // vec3.rs:201-203
__device__ inline V3f refract(V3f uv, V3f n, float eta) {                              // vec3.rs:205-210
  float cos_theta = fminf(-dot(uv, n), 1.0f);
  V3f perp = eta * (uv + cos_theta * n);
  V3f par = (-sqrtf(fabsf(1.0f - dot(perp, perp)))) * n;
  return perp + par;
}
__device__ inline float reflectance(float cosine, float ri) {  // material.rs:107-111 (Schlick)
  float r0 = (1.0f - ri) / (1.0f + ri);
  r0 *= r0;
  float m = 1.0f - cosine;
  float m2 = m * m;
  return r0 + (1.0f - r0) * m2 * m2 * m;
}
__device__ inline float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__device__ inline bool finite3(V3f v) { return isfinite(v.x + v.y + v.z); }  // radiance is >= 0: the sum is finite iff all three are

// tile index of this rank -> top-left pixel.  Super tiles are 32x32 pixels = 4x8 warp tiles;
// super tile k (device_types.h super_tile_xy) belongs to rank k % world (RRT_SHARD_TILE); otherwise world == 1.
__device__ inline void tile_origin(const DevWork& W, int tile, int& px0, int& py0) {
  int sj = tile >> 5, within = tile & 31;
  int world = W.shard_mode == RRT_SHARD_TILE ? W.world : 1;
  int rank = W.shard_mode == RRT_SHARD_TILE ? W.rank : 0;
  int k = rank + sj * world;
  int sx, sy;
  super_tile_xy(k, W.stiles_x, sx, sy);
  px0 = sx * 32 + (within & 3) * 8;
  py0 = sy * 32 + (within >> 2) * 4;
}

// Per-lane path state carried between the steps of the lane state machine.
struct Path {
  Ray32 ray;
  V3f T, L;    // throughput and radiance of the unrolled `emitted + attenuation * ray_color(...)`
  int bounce;  // max_depth - depth of the reference's recursion
  int skip;    // primitive the ray starts on
};

// ---- Camera::get_ray (camera.rs:156-182; b1 :152-180, b2 :148-172) --------------------------------
template <int INTEGRATOR, int FEAT>
__device__ __forceinline__ void generate_ray(const DevScene& S, const DevCamera& C, PathRng& rng, int px, int py, int s, Path& P) {
  rng.sample = (uint32_t)s;
  const Rand4 r0 = rng.draw(0xffffffffu, 0);
  float offx, offy;
  if (INTEGRATOR == RRT_INT_BOOK3) {  // sample_square_stratified, camera.rs:173-177
    int sj = s / C.sqrt_spp, si = s - sj * C.sqrt_spp;
    offx = ((float)si + r0.x) * C.recip_sqrt_spp - 0.5f;
    offy = ((float)sj + r0.y) * C.recip_sqrt_spp - 0.5f;
  } else {                            // sample_square
    offx = r0.x - 0.5f; offy = r0.y - 0.5f;
  }
  const float fx = (float)px + offx, fy = (float)py + offy;
  const V3f rel(C.p00x + fx * C.dux + fy * C.dvx, C.p00y + fx * C.duy + fy * C.dvy, C.p00z + fx * C.duz + fy * C.dvz);
  P.ray.o = V3f(C.cx, C.cy, C.cz);
  P.ray.d = rel;
  if (C.defocus) {  // defocus_disk_sample, camera.rs:179-182
    float lx, ly;
    uniform_disk(r0.z, r0.w, lx, ly);
    const V3f lens(lx * C.ddux + ly * C.ddvx, lx * C.dduy + ly * C.ddvy, lx * C.dduz + ly * C.ddvz);
    P.ray.o = P.ray.o + lens;
    P.ray.d = rel - lens;
  }
  P.ray.tm = 0.0f;
  if ((FEAT & SF_MOTION) && INTEGRATOR != RRT_INT_BOOK1 && S.has_motion) P.ray.tm = r0.v;  // random_double() for the ray time, camera.rs:169
  P.T = V3f(1.0f, 1.0f, 1.0f); P.L = V3f(0.0f, 0.0f, 0.0f);
  P.bounce = 0; P.skip = -1;
}

// ---- one call of ray_color after world.hit (camera.rs:184-254; b1 :182-209, b2 :174-201) ---------
// Returns true when the path continues with P.ray replaced, false when it ended (P.L is final).
template <int INTEGRATOR, int FEAT>
__device__ __forceinline__ bool shade_step(const DevScene& S, const DevCamera& C, const PathRng& rng, int max_depth,
                                           Hit32 h, Path& P, const DevPerlin* sh_perlin = nullptr) {
  Ray32& ray = P.ray;
  V3f& T = P.T;
  V3f& L = P.L;
  const int bounce = P.bounce;
  const Rand4 ra = rng.draw((uint32_t)bounce, 0);

  int medium = -1;
  if ((FEAT & SF_MEDIA) && S.n_media > 0) {  // ConstantMedium::hit, constant_medium.rs:39-79
    const float a = dot(ray.d, ray.d), inv_a = 1.0f / a;
    // free-flight draws: medium 0 takes ra.v, medium 1 takes ra.w where the bounce does not use it (books
    // 1-2: direction from ra.x, ra.y, roulette ra.z), so two media cost no second Philox block
    const int first_rb = INTEGRATOR == RRT_INT_BOOK3 ? 1 : 2;
    Rand4 rb;
    if (S.n_media > first_rb) rb = rng.draw((uint32_t)bounce, 1);
    for (int m = 0; m < S.n_media; ++m) {
      const DevMedium M = S.media[m];
      float t1, t2;
      const bool motion = (FEAT & SF_MOTION) && S.has_motion != 0;
      if (M.nsph == 1 && M.nquad == 0) {   // one sphere (both media of the book-2 final scene): one discriminant
        if (!boundary_sphere_span32(S.bprims, M, ray, a, inv_a, motion, t1, t2)) continue;
      } else {
        if (!boundary_hit32(S.bprims, M, ray, a, inv_a, -CUDART_INF_F, CUDART_INF_F, motion, t1)) continue;
        if (!boundary_hit32(S.bprims, M, ray, a, inv_a, t1 + 0.0001f, CUDART_INF_F, motion, t2)) continue;
      }
      if (t1 < 0.001f) t1 = 0.001f;
      if (t2 > h.t) t2 = h.t;
      if (t1 >= t2) continue;
      if (t1 < 0.0f) t1 = 0.0f;
      const int k = m - first_rb;  // index into the second block
      float xi = m == 0 ? ra.v : (k < 0 ? ra.w : (k == 0 ? rb.x : (k == 1 ? rb.y : (k == 2 ? rb.z : rb.w))));
      float len = sqrtf(a);
      float inside = (t2 - t1) * len;
      float hit_distance = M.neg_inv_density * __logf(xi);
      if (hit_distance > inside) continue;
      h.t = t1 + hit_distance / len;
      medium = m;
    }
  }

  if (h.prim < 0 && medium < 0) {  // miss
    if (INTEGRATOR == RRT_INT_BOOK1) {  // in_one_weekend/camera.rs:206-208
      V3f ud = normalize(ray.d);
      float a = 0.5f * (ud.y + 1.0f);
      L = L + T * V3f((1.0f - a) + a * 0.5f, (1.0f - a) + a * 0.7f, (1.0f - a) + a * 1.0f);
    } else {
      L = L + T * V3f(C.bgx, C.bgy, C.bgz);
    }
    return false;
  }

  Surface sf;
  if ((FEAT & SF_MEDIA) && medium >= 0) {  // constant_medium.rs:70-78
    sf.p = ray.o + h.t * ray.d;
    sf.n = V3f(1.0f, 0.0f, 0.0f); sf.n_obj = sf.n;
    sf.u = 0.0f; sf.v = 0.0f;
    sf.mat = S.media[medium].mat;
    sf.front = true; sf.is_sphere = false;
    P.skip = -1;
  } else {
    make_surface<FEAT>(S, ray, h, sf);
    P.skip = h.prim & ~PRIM_QUAD_BIT;
  }
  const DevMaterial M = S.materials[sf.mat];
  if ((FEAT & SF_TEXTURES) && M.albedo.w != 0.0f && sf.is_sphere) sphere_uv(sf.n_obj, sf.u, sf.v);
  V3f att(M.albedo.x, M.albedo.y, M.albedo.z);
  const bool textured = M.kind == RRT_M_LAMBERTIAN || M.kind == RRT_M_ISOTROPIC || M.kind == RRT_M_DIFFUSE_LIGHT;
  if ((FEAT & SF_TEXTURES) && textured && M.tex >= 0) att = texture_value<FEAT>(S, M.tex, sf.u, sf.v, sf.p, sh_perlin);

  // Single exit from here on: a `return` inside the material switch or the roulette test would leave
  // the lanes that skip it and the lanes that survive it in separate groups for the rest of the
  // function (ptxas only re-converges single-exit regions), and the mixture-PDF code below then ran
  // 1.7 times per call at 8.6 lanes on the book-3 scene (profiles/README.md, r01j).  `ok` carries the
  // reference's early returns instead.
  V3f dir(0.0f, 0.0f, 0.0f);
  bool pdf_branch = false;
  bool ok = true;
  switch (M.kind) {
    case RRT_M_LAMBERTIAN:
      if (INTEGRATOR == RRT_INT_BOOK3) {
        pdf_branch = true;
      } else {  // the_next_week/material.rs:41-53
        dir = sf.n + uniform_sphere(ra.x, ra.y);
        if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = sf.n;
      }
      break;
    case RRT_M_METAL: {  // material.rs:84-96 / the_next_week/material.rs:66-77
      dir = normalize(reflect(ray.d, sf.n)) + M.fuzz * uniform_sphere(ra.x, ra.y);
      if (INTEGRATOR != RRT_INT_BOOK3 && !(dot(dir, sf.n) > 0.0f)) ok = false;
      break;
    }
    case RRT_M_DIELECTRIC: {  // material.rs:113-137
      float ri = sf.front ? 1.0f / M.ior : M.ior;
      V3f ud = normalize(ray.d);
      float cos_theta = fminf(-dot(ud, sf.n), 1.0f);
      float sin_theta = sqrtf(fmaxf(0.0f, 1.0f - cos_theta * cos_theta));
      bool cannot_refract = ri * sin_theta > 1.0f;
      dir = (cannot_refract || reflectance(cos_theta, ri) > ra.x) ? reflect(ud, sf.n) : refract(ud, sf.n, ri);
      att = V3f(1.0f, 1.0f, 1.0f);
      break;
    }
    case RRT_M_ISOTROPIC:
      if (INTEGRATOR == RRT_INT_BOOK3) pdf_branch = true;
      else dir = uniform_sphere(ra.x, ra.y);  // the_next_week/material.rs:152-158
      break;
    case RRT_M_DIFFUSE_LIGHT:
      // emitted: book 2 both faces (material.rs:131-135), book 3 front face only (:155-160);
      // book 1 has no emission
      if (INTEGRATOR == RRT_INT_BOOK2 || (INTEGRATOR == RRT_INT_BOOK3 && sf.front)) L = L + T * att;
      ok = false;
      break;
    default:
      ok = false;  // EmptyMaterial
      break;
  }

  // Russian roulette after five bounces on the local attenuation (camera.rs:200-207, 223-237)
  const float rrc = clampf(max3(att), 0.05f, 0.95f);
  const bool deep = bounce >= 5;
  const float rr = deep ? rrc : 1.0f;
  if (deep && ra.z > rrc) ok = false;
  if (ok) {
    if (pdf_branch) {
      // MixturePdf(HittablePdf(lights, p), srec.pdf), camera.rs:239-253, pdf.rs:77-100
      // one Philox block per bounce: ra.x is the mixture coin, and its remaining mantissa bits
      // (2 ra.x mod 1, uniform and independent of the coin) pick the light; ra.y, ra.w give the
      // direction; ra.z is the roulette draw above; ra.v the medium's free-flight draw
      const bool lamb = M.kind == RRT_M_LAMBERTIAN;
      const int nl = S.n_lights;
      const bool pick_light = ra.x < 0.5f;
      const float u_pick = 2.0f * ra.x - (pick_light ? 0.0f : 1.0f);
      // The cosine lobe about the normal (pdf.rs:51-53, vec3.rs:212-222) and the cone towards a sphere
      // light (sphere.rs:55-65, 117-122) are the same code around different axes and heights z: one
      // basis, one sincos and one transform for both groups of lanes instead of a branch each.
      V3f axis = sf.n;
      float z = 0.0f;
      bool around_axis = true;
      if (nl > 0 && pick_light) {
        const int li = min((int)(u_pick * (float)nl), nl - 1);  // random_int(0, n-1), hittable_list.rs:71-75
        const DevLight& Lt = S.lights[li];
        if (Lt.kind == 0) {
          axis = V3f(Lt.a.x, Lt.a.y, Lt.a.z) - sf.p;
          z = 1.0f + ra.w * (sqrtf(fmaxf(0.0f, 1.0f - Lt.a.w * Lt.a.w / dot(axis, axis))) - 1.0f);
        } else {
          dir = light_random(Lt, sf.p, ra.y, ra.w);  // quad light, or the trait default
          around_axis = false;
        }
      } else if (lamb) {
        z = sqrtf(fmaxf(0.0f, 1.0f - ra.w));
      } else {
        dir = uniform_sphere(ra.y, ra.w);  // SpherePdf, pdf.rs:27-29
        around_axis = false;
      }
      if (around_axis) {
        float sn, cs;
        __sincosf(2.0f * RRT_PI * ra.y - RRT_PI, &sn, &cs);
        const float q = sqrtf(fmaxf(0.0f, 1.0f - z * z));
        dir = Onb(axis).transform(V3f(cs * q, sn * q, z));
      }
      V3f udir = normalize(dir);
      float cos_n = dot(sf.n, udir);
      float pdf_mat = lamb ? (cos_n > 0.0f ? cos_n * (1.0f / RRT_PI) : 0.0f) : (0.25f / RRT_PI);
      float pdf_value = pdf_mat;
      if (nl > 0) {
        float pl = 0.0f;
        for (int i = 0; i < nl; ++i) pl += light_pdf(S.lights[i], sf.p, dir);  // hittable_list.rs:60-69
        pdf_value = 0.5f * (pl / (float)nl) + 0.5f * pdf_mat;
      }
      ok = pdf_value > 0.0f;  // camera.rs:243-246
      float spdf = lamb ? (cos_n < 0.0f ? 0.0f : cos_n * (1.0f / RRT_PI)) : (0.25f / RRT_PI);  // material.rs:63-70, 186-188
      if (ok) T = T * (att * (spdf / (pdf_value * rr)));
    } else {
      T = T * (att * (1.0f / rr));
    }
  }
  if (ok) {
    ray.o = sf.p; ray.d = dir;
    P.bounce = bounce + 1;
    ok = P.bounce < max_depth;  // depth <= 0 -> black, camera.rs:185-187
  }
  return ok;
}

// ---- the megakernel ---------------------------------------------------------------------------------
template <int INTEGRATOR, int FEAT, bool STEAL>
__global__ void __launch_bounds__(RRT_BLOCK, render_min_blocks(FEAT))
render_kernel(const __grid_constant__ RenderArgs A) {
  const DevScene& S = A.S;
  const DevWork& W = A.W;
  const DevCamera& C = A.cam;
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int n_items = W.n_tiles * W.chunks;
  const int n_samples = W.s_end - W.s_begin;
  // Perlin instances: the first Perlin object's tables in shared memory (shade.cuh)
  __shared__ __align__(16) unsigned char sh_perlin_raw[(FEAT & SF_NOISE) ? sizeof(DevPerlin) : 16];
  const DevPerlin* sh_perlin = nullptr;
  if (FEAT & SF_NOISE) {
    if (S.perlins) {
      const uint4* src = reinterpret_cast<const uint4*>(S.perlins);
      uint4* dst = reinterpret_cast<uint4*>(sh_perlin_raw);
      for (int i = threadIdx.x; i < (int)(sizeof(DevPerlin) / 16); i += RRT_BLOCK) dst[i] = src[i];
      sh_perlin = reinterpret_cast<const DevPerlin*>(sh_perlin_raw);
    }
    __syncthreads();
  }

  for (;;) {
    int item = 0;
    if (lane == 0) item = (int)atomicAdd(A.counter, 1u);
    item = __shfl_sync(FULL, item, 0);
    if (item >= n_items) break;
    const int tile = item / W.chunks, chunk = item - tile * W.chunks;
    int px0, py0;
    tile_origin(W, tile, px0, py0);
    const int px = px0 + (lane & 7), py = py0 + (lane >> 3);
    const bool valid = px < W.width && py < W.height;
    const int s0 = W.s_begin + (int)(((long long)n_samples * chunk) / W.chunks);
    const int s1 = W.s_begin + (int)(((long long)n_samples * (chunk + 1)) / W.chunks);

    // where this pixel's sums go: the frame, or (tile sharding for a gather) this rank's packed super tiles
    const size_t out_idx = W.compact ? (size_t)(tile >> 5) * 1024 + (size_t)(((tile & 31) >> 2) * 4 + (lane >> 3)) * 32 + ((tile & 3) * 8 + (lane & 7))
                                     : (size_t)py * W.width + px;
    if (W.max_depth <= 0) {  // ray_color(r, depth <= 0) is black for every sample (camera.rs:185-187; b1 :183-185)
      if (valid) {
        float4* dst = A.accum + out_idx;
        if (W.atomic_accum) atomicAdd(&dst->w, (float)(s1 - s0));
        else *dst = make_float4(0.0f, 0.0f, 0.0f, (float)(s1 - s0));
      }
      continue;
    }
    PathRng rng;
    rng.pixel = (uint32_t)(py * W.width + px);
    rng.sample = 0;
    rng.rk = W.rkeys;

    V3f sum(0.0f, 0.0f, 0.0f);
    // Sample stealing inside the warp: a lane that has traced all samples of its own pixel goes on
    // with a neighbour's (shared-memory sample counters, one per pixel of the tile; foreign radiance
    // goes through shared-memory adds), so the warp stays full until the whole tile's samples are
    // handed out instead of waiting for its slowest pixel.  Every (pixel, sample) is still traced
    // exactly once with its own Philox key.
    __shared__ int sh_next[RRT_BLOCK];
    __shared__ float sh_sum[3][RRT_BLOCK];
    const int wbase = threadIdx.x & ~31;
    if (STEAL) {
      __syncwarp();
      sh_next[threadIdx.x] = valid ? s0 : s1;
      sh_sum[0][threadIdx.x] = 0.0f; sh_sum[1][threadIdx.x] = 0.0f; sh_sum[2][threadIdx.x] = 0.0f;
      __syncwarp();
    }
    int tgt = lane;
    int s = s0;
    bool alive = false;
    Path P;
    P.ray.o = V3f(0, 0, 0); P.ray.d = V3f(0, 0, 1); P.ray.tm = 0.0f;
    P.T = V3f(1.0f, 1.0f, 1.0f); P.L = V3f(0.0f, 0.0f, 0.0f);
    P.bounce = 0; P.skip = -1;

    for (;;) {
      if (!alive) {  // next sample of this lane's pixel, or of a neighbour's once its own are done
        int s_new = s1;
        if (STEAL) {
          int tries = 0;
          // (ptxas unrolls this four times, 199 instructions.  Do not roll it up: with `#pragma unroll 1` the lanes
          // leave the loop at different trips and no longer re-converge before the trace — C2 -24 %, C5 3.5x slower,
          // profiles/README.md r02o)
          for (; tries < 32; ++tries) {
            s_new = atomicAdd(&sh_next[wbase + tgt], 1);
            if (s_new < s1) break;
            tgt = (tgt + 1) & 31;
          }
          if (tries == 32) break;
        } else {
          if (!valid || s >= s1) break;
          s_new = s++;
        }
        const int tx = px0 + (tgt & 7), ty = py0 + (tgt >> 3);
        rng.pixel = (uint32_t)(ty * W.width + tx);
        generate_ray<INTEGRATOR, FEAT>(S, C, rng, tx, ty, s_new, P);
        alive = true;
      }
      Hit32 h;
      h.t = CUDART_INF_F; h.prim = -1;
      world_hit32<FEAT>(S, P.ray, 0.001f, P.skip, h);  // world.hit(r, Interval::new(0.001, INFINITY)), camera.rs:189
      if (!shade_step<INTEGRATOR, FEAT>(S, C, rng, W.max_depth, h, P, sh_perlin)) {
        if (finite3(P.L)) {  // a non-finite sample contributes nothing (color.rs:19-27)
          if (!STEAL || tgt == lane) sum = sum + P.L;
          else { atomicAdd(&sh_sum[0][wbase + tgt], P.L.x); atomicAdd(&sh_sum[1][wbase + tgt], P.L.y); atomicAdd(&sh_sum[2][wbase + tgt], P.L.z); }
        }
        alive = false;
      }
    }

    if (STEAL) {
      __syncwarp();
      sum = sum + V3f(sh_sum[0][threadIdx.x], sh_sum[1][threadIdx.x], sh_sum[2][threadIdx.x]);
    }
    if (valid) {
      float4* dst = A.accum + out_idx;
      float cnt = (float)(s1 - s0);
      if (W.atomic_accum) {
        atomicAdd(&dst->x, sum.x); atomicAdd(&dst->y, sum.y); atomicAdd(&dst->z, sum.z); atomicAdd(&dst->w, cnt);
      } else {
        *dst = make_float4(sum.x, sum.y, sum.z, cnt);
      }
    }
  }
}

}  // namespace rrt
