// render_kernel_pool.cuh — the cross-lane alternative to render_kernel's "lane = pixel" schedule
// (north_star: "a persistent-thread megakernel, or a wavefront kernel where ncu shows divergence wins").
//
// The megakernel ties a path to a lane from its camera ray to its end: a lane whose path ended
// regenerates alone (ray generation runs at 7-14 of 32 lanes), lanes that hit different materials
// run the material branches one after the other (the book-3 mixture-pdf code at 15 of 29 lanes), and a
// warp carries 21-28 live lanes through the traversal.  Here a warp owns a POOL of 64 path slots in
// shared memory (two per lane) and works through it in phases, each over a compacted list of slots:
//   regenerate  every dead slot, together, takes the next (pixel, sample) of the work item;
//   traverse    world.hit for every live slot, full warps;
//   shade       the slots sorted by what they hit (miss / material kind), so a pass of 32 lanes runs one
//               or two material branches instead of all of them; a path that ends adds its radiance to
//               its pixel's sum (shared-memory add) and frees its slot.
// Every (pixel, sample) keeps its own Philox key and every function below the schedule is the
// megakernel's (generate_ray, world_hit32, shade_step), so the two kernels render the same frame up to
// float summation order.  State traffic: 18 words per slot and phase through shared memory.
// Selected with RRT_POOL=1 (rrt_init); profiles/README.md r02 has the A/B.
#pragma once
#include "render_kernel.cuh"

namespace rrt {

#define RRT_POOL_SLOTS 64

// field f of slot s of this warp's pool: f * 64 + s (a warp reads 32 consecutive words: no bank conflicts)
enum PoolField { PF_OX, PF_OY, PF_OZ, PF_DX, PF_DY, PF_DZ, PF_TM, PF_TX, PF_TY, PF_TZ, PF_LX, PF_LY, PF_LZ, PF_BOUNCE, PF_SKIP, PF_PIXSMP,
                 PF_HT, PF_HPRIM, PF_COUNT };
enum PoolState { PS_DEAD = 0, PS_TRAVERSE = 1, PS_SHADE = 2 };   // PS_SHADE + (kind << 2)

struct WarpPool {
  float f[PF_COUNT][RRT_POOL_SLOTS];
  int state[RRT_POOL_SLOTS];
  int list[RRT_POOL_SLOTS];     // compacted slot lists (dead slots, then the shading order)
  int pixmap[32];               // k-th valid pixel of the tile
  float sum[3][32];             // per-pixel radiance of the work item
};

__device__ __forceinline__ void pool_store_path(WarpPool& Q, int s, const Path& P, int pixsmp) {
  Q.f[PF_OX][s] = P.ray.o.x; Q.f[PF_OY][s] = P.ray.o.y; Q.f[PF_OZ][s] = P.ray.o.z;
  Q.f[PF_DX][s] = P.ray.d.x; Q.f[PF_DY][s] = P.ray.d.y; Q.f[PF_DZ][s] = P.ray.d.z;
  Q.f[PF_TM][s] = P.ray.tm;
  Q.f[PF_TX][s] = P.T.x; Q.f[PF_TY][s] = P.T.y; Q.f[PF_TZ][s] = P.T.z;
  Q.f[PF_LX][s] = P.L.x; Q.f[PF_LY][s] = P.L.y; Q.f[PF_LZ][s] = P.L.z;
  Q.f[PF_BOUNCE][s] = __int_as_float(P.bounce); Q.f[PF_SKIP][s] = __int_as_float(P.skip);
  Q.f[PF_PIXSMP][s] = __int_as_float(pixsmp);
}
__device__ __forceinline__ void pool_load_ray(const WarpPool& Q, int s, Ray32& r, int& skip) {
  r.o = V3f(Q.f[PF_OX][s], Q.f[PF_OY][s], Q.f[PF_OZ][s]);
  r.d = V3f(Q.f[PF_DX][s], Q.f[PF_DY][s], Q.f[PF_DZ][s]);
  r.tm = Q.f[PF_TM][s];
  skip = __float_as_int(Q.f[PF_SKIP][s]);
}

template <int INTEGRATOR, int FEAT>
__global__ void __launch_bounds__(RRT_BLOCK, render_min_blocks(FEAT))
render_kernel_pool(const __grid_constant__ RenderArgs A) {
  const DevScene& S = A.S;
  const DevWork& W = A.W;
  const DevCamera& C = A.cam;
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  const int n_items = W.n_tiles * W.chunks;
  const int n_samples = W.s_end - W.s_begin;
  __shared__ WarpPool pools[RRT_BLOCK / 32];
  WarpPool& Q = pools[threadIdx.x >> 5];

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
    const size_t out_idx = W.compact ? (size_t)(tile >> 5) * 1024 + (size_t)(((tile & 31) >> 2) * 4 + (lane >> 3)) * 32 + ((tile & 3) * 8 + (lane & 7))
                                     : (size_t)py * W.width + px;
    if (W.max_depth <= 0) {
      if (valid) {
        float4* dst = A.accum + out_idx;
        if (W.atomic_accum) atomicAdd(&dst->w, (float)(s1 - s0));
        else *dst = make_float4(0.0f, 0.0f, 0.0f, (float)(s1 - s0));
      }
      continue;
    }
    // the tile's valid pixels, numbered: path q of the item is sample s0 + q / n_valid of pixel pixmap[q % n_valid]
    const unsigned vmask = __ballot_sync(FULL, valid);
    const int n_valid = __popc(vmask);
    __syncwarp();
    if (valid) Q.pixmap[__popc(vmask & lt)] = lane;
    Q.sum[0][lane] = 0.0f; Q.sum[1][lane] = 0.0f; Q.sum[2][lane] = 0.0f;
    Q.state[lane] = PS_DEAD; Q.state[lane + 32] = PS_DEAD;
    __syncwarp();
    const int n_total = n_valid * (s1 - s0);
    int issued = 0;
    PathRng rng;
    rng.rk = W.rkeys;

    for (;;) {
      // ---- regenerate: the dead slots, compacted ---------------------------------------------------------
      const bool deadA = Q.state[lane] == PS_DEAD, deadB = Q.state[lane + 32] == PS_DEAD;
      const unsigned mA = __ballot_sync(FULL, deadA), mB = __ballot_sync(FULL, deadB);
      const int nA = __popc(mA), ndead = nA + __popc(mB);
      const int nissue = min(ndead, n_total - issued);
      if (nissue == 0 && ndead == RRT_POOL_SLOTS) break;
      if (nissue > 0) {
        if (deadA) Q.list[__popc(mA & lt)] = lane;
        if (deadB) Q.list[nA + __popc(mB & lt)] = lane + 32;
        __syncwarp();
#pragma unroll 1
        for (int k = lane; k < nissue; k += 32) {
          const int slot = Q.list[k];
          const int q = issued + k;
          int pi, sm;
          if (n_valid == 32) { pi = q & 31; sm = q >> 5; } else { sm = q / n_valid; pi = q - sm * n_valid; }
          const int pl = Q.pixmap[pi];
          const int tx = px0 + (pl & 7), ty = py0 + (pl >> 3);
          rng.pixel = (uint32_t)(ty * W.width + tx);
          Path P;
          generate_ray<INTEGRATOR, FEAT>(S, C, rng, tx, ty, s0 + sm, P);
          pool_store_path(Q, slot, P, pl | ((s0 + sm) << 5));
          Q.state[slot] = PS_TRAVERSE;
        }
        issued += nissue;
        __syncwarp();
      }
      // ---- traverse: world.hit for every live slot; classify what was hit --------------------------------
#pragma unroll 1
      for (int slot = lane; slot < RRT_POOL_SLOTS; slot += 32) {
        if (Q.state[slot] != PS_TRAVERSE) continue;
        Ray32 r; int skip;
        pool_load_ray(Q, slot, r, skip);
        Hit32 h;
        h.t = CUDART_INF_F; h.prim = -1;
        world_hit32<FEAT>(S, r, 0.001f, skip, h);
        Q.f[PF_HT][slot] = h.t; Q.f[PF_HPRIM][slot] = __int_as_float(h.prim);
        int kind = 7;  // miss
        if (h.prim >= 0) {
          const int idx = h.prim & ~PRIM_QUAD_BIT;
          const int mat = __float_as_int(ldg4(S.prims + 4 * idx + ((h.prim & PRIM_QUAD_BIT) ? 3 : 2)).x);
          kind = S.materials[mat].kind;
        }
        Q.state[slot] = PS_SHADE | (kind << 2);
      }
      __syncwarp();
      // ---- shade: slots ordered by kind ---------------------------------------------------------------------
      const int kA = Q.state[lane], kB = Q.state[lane + 32];
      int base = 0;
#pragma unroll
      for (int kind = 0; kind < 8; ++kind) {
        const unsigned a = __ballot_sync(FULL, kA == (PS_SHADE | (kind << 2))), b = __ballot_sync(FULL, kB == (PS_SHADE | (kind << 2)));
        if (kA == (PS_SHADE | (kind << 2))) Q.list[base + __popc(a & lt)] = lane;
        if (kB == (PS_SHADE | (kind << 2))) Q.list[base + __popc(a) + __popc(b & lt)] = lane + 32;
        base += __popc(a) + __popc(b);
      }
      __syncwarp();
#pragma unroll 1
      for (int k = lane; k < base; k += 32) {
        const int slot = Q.list[k];
        Path P;
        pool_load_ray(Q, slot, P.ray, P.skip);
        P.T = V3f(Q.f[PF_TX][slot], Q.f[PF_TY][slot], Q.f[PF_TZ][slot]);
        P.L = V3f(Q.f[PF_LX][slot], Q.f[PF_LY][slot], Q.f[PF_LZ][slot]);
        P.bounce = __float_as_int(Q.f[PF_BOUNCE][slot]);
        const int pixsmp = __float_as_int(Q.f[PF_PIXSMP][slot]);
        const int pl = pixsmp & 31;
        rng.pixel = (uint32_t)((py0 + (pl >> 3)) * W.width + px0 + (pl & 7));
        rng.sample = (uint32_t)(pixsmp >> 5);
        Hit32 h;
        h.t = Q.f[PF_HT][slot]; h.prim = __float_as_int(Q.f[PF_HPRIM][slot]);
        if (shade_step<INTEGRATOR, FEAT>(S, C, rng, W.max_depth, h, P)) {
          pool_store_path(Q, slot, P, pixsmp);
          Q.state[slot] = PS_TRAVERSE;
        } else {
          if (finite3(P.L)) { atomicAdd(&Q.sum[0][pl], P.L.x); atomicAdd(&Q.sum[1][pl], P.L.y); atomicAdd(&Q.sum[2][pl], P.L.z); }
          Q.state[slot] = PS_DEAD;
        }
      }
      __syncwarp();
    }

    if (valid) {
      float4* dst = A.accum + out_idx;
      const float cnt = (float)(s1 - s0);
      const V3f sum(Q.sum[0][lane], Q.sum[1][lane], Q.sum[2][lane]);
      if (W.atomic_accum) {
        atomicAdd(&dst->x, sum.x); atomicAdd(&dst->y, sum.y); atomicAdd(&dst->z, sum.z); atomicAdd(&dst->w, cnt);
      } else {
        *dst = make_float4(sum.x, sum.y, sum.z, cnt);
      }
    }
  }
}

}  // namespace rrt
