// intersect.cuh — ray/primitive tests and the stack BVH traversal.
//
// What it restates (hot loops 3-4 of SURVEY.md §3.1):
//   Sphere::hit        the_rest_of_your_life/sphere.rs:69-96   (open interval: Interval::surrounds)
//   Quad::hit          the_rest_of_your_life/quad.rs:64-91     (closed interval: Interval::contains)
//   Aabb::hit          the_rest_of_your_life/aabb.rs:52-84     (slab test)
//   BvhNode::hit       the_rest_of_your_life/bvh.rs:159-172    (closest hit below the node)
//   HittableList::hit  the_rest_of_your_life/hittable_list.rs:42-54
// Translate/RotateY (hittable.rs:86-173) do not appear: instances are baked to world space on the
// host (flatten.cpp), which leaves t unchanged because the maps are rigid.
//
// Equal-t rule.  The reference's result at an exact tie is decided by list order plus the
// open/closed interval asymmetry: a later quad replaces an equal-t hit (contains), a later sphere
// does not (surrounds).  Traversal order here is different, so the rule is applied explicitly on
// the primitives' construction order: quad beats sphere; between quads the later wins; between
// spheres the earlier wins.
#pragma once
#include <math_constants.h>

#include "device_types.h"
#include "vecmath.cuh"

namespace rrt {

#define PRIM_QUAD_BIT 0x40000000

struct Ray32 {
  V3f o, d;
  float tm;
};

struct Hit32 {
  float t;   // closest so far (starts at +inf)
  int prim;  // flat prim index | PRIM_QUAD_BIT, or -1
};

__device__ inline int prim_order(const float4* prims, int prim) {
  int idx = prim & ~PRIM_QUAD_BIT;
  float4 m = ldg4(prims + 4 * idx + ((prim & PRIM_QUAD_BIT) ? 3 : 2));
  return __float_as_int(m.z);
}

// cand beats the current best at equal t?
__device__ inline bool tie_accept(const float4* prims, int cand, int best) {
  if (best < 0) return true;
  bool cq = cand & PRIM_QUAD_BIT, bq = best & PRIM_QUAD_BIT;
  if (cq != bq) return cq;
  int co = prim_order(prims, cand), bo = prim_order(prims, best);
  return cq ? (co > bo) : (co < bo);
}

// Sphere::hit in f32.  The discriminant is formed from the component of oc perpendicular to the
// ray (r^2 - |oc - (h/a) d|^2 = discriminant / a), which avoids the cancellation of
// |oc|^2 - r^2 for small far spheres; same roots as sphere.rs:71-89.
// skip: the sphere this ray starts on — only its far root is legal, and only when heading inward.
__device__ inline void sphere_hit32(const float4* __restrict__ prims, int i, const Ray32& r, float a,
                                    float inv_a, float tmin, int skip, bool motion, Hit32& h) {
  float4 p0 = ldg4(prims + 4 * i);
  V3f c(p0.x, p0.y, p0.z);
  if (motion) {
    float4 p1 = ldg4(prims + 4 * i + 1);
    c = c + r.tm * V3f(p1.x, p1.y, p1.z);  // center.at(time), sphere.rs:70
  }
  V3f oc = c - r.o;
  float hh = dot(r.d, oc);
  float k = hh * inv_a;
  V3f perp = oc - k * r.d;
  float disc = p0.w * p0.w - dot(perp, perp);
  if (disc < 0.0f) return;
  float sq = sqrtf(disc * a);
  float root;
  if (i == skip) {
    if (hh <= 0.0f) return;
    root = (hh + sq) * inv_a;
    if (!(root > tmin && root <= h.t)) return;
  } else {
    root = (hh - sq) * inv_a;
    if (!(root > tmin && root <= h.t)) {
      root = (hh + sq) * inv_a;
      if (!(root > tmin && root <= h.t)) return;
    }
  }
  if (root == h.t && !tie_accept(prims, i, h.prim)) return;
  h.t = root;
  h.prim = i;
}

// Quad::hit in f32.  alpha = w.(hp x v) = p.A + a0, beta = w.(u x hp) = p.B + b0 (quad.rs:76-79).
__device__ inline void quad_hit32(const float4* __restrict__ prims, int i, const Ray32& r, float tmin,
                                  int skip, Hit32& h) {
  if (i == skip) return;  // a ray leaving a plane cannot meet it again (reference: t ~ 1e-16 < 0.001)
  float4 p0 = ldg4(prims + 4 * i);
  V3f n(p0.x, p0.y, p0.z);
  float denom = dot(n, r.d);
  if (fabsf(denom) < 1e-8f) return;  // quad.rs:67
  float t = (p0.w - dot(n, r.o)) / denom;
  if (!(t >= tmin && t <= h.t)) return;  // contains, quad.rs:72
  V3f p = r.o + t * r.d;
  float4 p1 = ldg4(prims + 4 * i + 1);
  float alpha = dot(p, V3f(p1.x, p1.y, p1.z)) + p1.w;
  if (!(alpha >= 0.0f && alpha <= 1.0f)) return;  // is_interior, quad.rs:51-60
  float4 p2 = ldg4(prims + 4 * i + 2);
  float beta = dot(p, V3f(p2.x, p2.y, p2.z)) + p2.w;
  if (!(beta >= 0.0f && beta <= 1.0f)) return;
  if (t == h.t && !tie_accept(prims, i | PRIM_QUAD_BIT, h.prim)) return;
  h.t = t;
  h.prim = i | PRIM_QUAD_BIT;
}

__device__ inline float safe_rcp(float d) {
  // a zero direction component would turn the slab products into NaN (0 * inf); aabb.rs:58 has
  // the same hazard in f64 but never meets an exact zero.  Clamp instead.
  float ad = fabsf(d);
  return 1.0f / (ad > 1e-30f ? d : copysignf(1e-30f, d));
}

// Box leaf (device_types.h): the six quads of an axis-aligned box behind one slab test.  The slab
// only culls and names the entry face; the hit itself is always quad_hit32's, so t, the interior
// test and the tie rules are those of the six-quad list the reference holds (quad.rs box_()).
// Whenever the entry face does not settle it (origin inside, grazing an edge) all six are tested.
__device__ inline void box_leaf_hit32(const float4* __restrict__ prims, const float4* __restrict__ boxes, int box,
                                      const Ray32& r, V3f id, float tmin, int skip, Hit32& h) {
  const float4 b0 = ldg4(boxes + 2 * box), b1 = ldg4(boxes + 2 * box + 1);
  const int first = __float_as_int(b0.w);
  const unsigned from = (unsigned)(skip - first);
  if (from < 6u) {  // the ray starts on one of this box's faces
    const float dk = (from >> 1) == 0 ? r.d.x : ((from >> 1) == 1 ? r.d.y : r.d.z);
    if ((from & 1u) ? dk > 0.0f : dk < 0.0f) return;  // leaving outward: a convex box is not met again
  } else {
    const float x0 = (b0.x - r.o.x) * id.x, x1 = (b1.x - r.o.x) * id.x;
    const float y0 = (b0.y - r.o.y) * id.y, y1 = (b1.y - r.o.y) * id.y;
    const float z0 = (b0.z - r.o.z) * id.z, z1 = (b1.z - r.o.z) * id.z;
    const float nx = fminf(x0, x1), ny = fminf(y0, y1), nz = fminf(z0, z1);
    const float near = fmaxf(fmaxf(nx, ny), nz);
    const float far = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
    if (!(near <= far * 1.0000004f) || far < tmin || near > h.t * 1.0000004f) return;  // the box record is padded
    if (near >= tmin) {  // origin outside: the closest face is the one the ray enters through
      const int axis = near == nx ? 0 : (near == ny ? 1 : 2);
      const float da = axis == 0 ? r.d.x : (axis == 1 ? r.d.y : r.d.z);
      const int face = first + 2 * axis + (da > 0.0f ? 0 : 1);
      quad_hit32(prims, face, r, tmin, skip, h);
      if (h.prim == (face | PRIM_QUAD_BIT)) return;
    }
  }
#ifndef RRT_BOX_UNROLL
#pragma unroll 1   // the rare path (origin inside the box, grazing an edge): one copy of the quad test
#endif
  for (int f = 0; f < 6; ++f) quad_hit32(prims, first + f, r, tmin, skip, h);
}

// One leaf: a contiguous run, spheres first.  ~ref = (first << 8) | (nsph << 4) | nquad.
template <int FEAT>
__device__ inline void leaf_hit32(const float4* __restrict__ prims, const float4* __restrict__ boxes, int leaf,
                                  const Ray32& r, V3f id, float a, float inv_a, float tmin, int skip, bool motion, Hit32& h) {
  if ((FEAT & SF_BOXES) && (leaf & 0xFF) == LEAF_BOX_MARK) { box_leaf_hit32(prims, boxes, leaf >> 8, r, id, tmin, skip, h); return; }
  int first = leaf >> 8, ns = (leaf >> 4) & 15, nq = leaf & 15;
#ifdef RRT_FRONT_UNROLL1
#pragma unroll 1
#endif
  for (int k = 0; k < ns; ++k) sphere_hit32(prims, first + k, r, a, inv_a, tmin, skip, motion, h);
  first += ns;
  if (FEAT & SF_QUADS) {
#ifdef RRT_FRONT_UNROLL1
#pragma unroll 1
#endif
    for (int k = 0; k < nq; ++k) quad_hit32(prims, first + k, r, tmin, skip, h);
  }
}


// Packed FP32 FMA (Blackwell fma.rn.f32x2 -> SASS FFMA2): both halves of a 64-bit register pair in
// one issue slot; ptxas folds negation, |x| and a scalar broadcast into the operand modifiers.
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  unsigned long long ra, rb, rc, rd;
  ra = *reinterpret_cast<unsigned long long*>(&a);
  rb = *reinterpret_cast<unsigned long long*>(&b);
  rc = *reinterpret_cast<unsigned long long*>(&c);
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2*>(&rd);
}
__device__ __forceinline__ float2 bcast2(float v) { return make_float2(v, v); }
// sign flips as plain (non-.ftz) PTX so that ptxas can fold them into operand modifiers even under
// --use_fast_math, whose neg.ftz/abs.ftz must flush denormals and so cost an instruction each
__device__ __forceinline__ float neg_nf(float v) { float r; asm("neg.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float abs_nf(float v) { float r; asm("abs.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float2 neg2(float2 v) { return make_float2(neg_nf(v.x), neg_nf(v.y)); }

// Aabb::hit (aabb.rs:52-84) for both children of a node.  Boxes are (center, half-extent), so the
// two plane distances of an axis are tc -+ e*|1/d| with tc = c*(1/d) - o*(1/d); every pair of the
// node record is (child0, child1) (device_types.h), so that is 9 packed FMAs for both children,
// then one 3-input max/min per interval end — no per-axis min/max pairs.
__device__ __forceinline__ void slab2(float4 n0, float4 n1, float4 n2, float idx, float idy, float idz,
                                      float oodx, float oody, float oodz, float tmin, float tmax,
                                      float& near0, float& far0, float& near1, float& far1) {
  const float2 tx = fma2(make_float2(n0.x, n0.y), bcast2(idx), bcast2(neg_nf(oodx)));
  const float2 ty = fma2(make_float2(n0.z, n0.w), bcast2(idy), bcast2(neg_nf(oody)));
  const float2 tz = fma2(make_float2(n1.x, n1.y), bcast2(idz), bcast2(neg_nf(oodz)));
  const float2 ex = make_float2(n1.z, n1.w), ey = make_float2(n2.x, n2.y), ez = make_float2(n2.z, n2.w);
  const float2 ax = bcast2(abs_nf(idx)), ay = bcast2(abs_nf(idy)), az = bcast2(abs_nf(idz));
  const float2 fx = fma2(ex, ax, tx), fy = fma2(ey, ay, ty), fz = fma2(ez, az, tz);
  const float2 nx = fma2(neg2(ex), ax, tx), ny = fma2(neg2(ey), ay, ty), nz = fma2(neg2(ez), az, tz);
  near0 = fmaxf(fmaxf(nx.x, ny.x), fmaxf(nz.x, tmin));
  far0 = fminf(fminf(fx.x, fy.x), fminf(fz.x, tmax));
  near1 = fmaxf(fmaxf(nx.y, ny.y), fmaxf(nz.y, tmin));
  far1 = fminf(fminf(fx.y, fy.y), fminf(fz.y, tmax));
}

__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  unsigned long long ra, rb, rd;
  ra = *reinterpret_cast<unsigned long long*>(&a);
  rb = *reinterpret_cast<unsigned long long*>(&b);
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  return *reinterpret_cast<float2*>(&rd);
}

// The quads of a brute-force root leaf, two at a time (device_types.h quad pair record): Quad::hit
// (quad.rs:64-91) for both as 16 packed FMAs and two reciprocals, every lane running the same
// straight-line code; the reference's early-outs (parallel, t out of range, alpha / beta outside
// [0, 1]) become one predicate per quad.  Quads are accepted in leaf order with t <= h.t; the leaf order
// is the reference's visiting order, which makes that the reference's equal-t rule (see below).
__device__ __forceinline__ void quad_pairs_hit32(const float4* __restrict__ prims, const float4* __restrict__ qp, int first, int nq,
                                                 const Ray32& r, float tmin, int skip, Hit32& h) {
  const float2 ox = bcast2(r.o.x), oy = bcast2(r.o.y), oz = bcast2(r.o.z);
  const float2 dx = bcast2(r.d.x), dy = bcast2(r.d.y), dz = bcast2(r.d.z);
  for (int k = 0; k < nq; k += 2) {
    const float4* g = qp + 3 * k;
    const float4 q0 = ldg4(g), q1 = ldg4(g + 1), q2 = ldg4(g + 2), q3 = ldg4(g + 3), q4 = ldg4(g + 4), q5 = ldg4(g + 5);
    const float2 nx = make_float2(q0.x, q0.y), ny = make_float2(q0.z, q0.w), nz = make_float2(q1.x, q1.y);
    const float2 denom = fma2(nz, dz, fma2(ny, dy, mul2(nx, dx)));
    const float2 num = fma2(neg2(nz), oz, fma2(neg2(ny), oy, fma2(neg2(nx), ox, make_float2(q1.z, q1.w))));  // d - n.o
    const float2 t = make_float2(num.x / denom.x, num.y / denom.y);
    const float2 px = fma2(t, dx, ox), py = fma2(t, dy, oy), pz = fma2(t, dz, oz);
    const float2 al = fma2(pz, make_float2(q3.x, q3.y), fma2(py, make_float2(q2.z, q2.w), fma2(px, make_float2(q2.x, q2.y), make_float2(q3.z, q3.w))));
    const float2 be = fma2(pz, make_float2(q5.x, q5.y), fma2(py, make_float2(q4.z, q4.w), fma2(px, make_float2(q4.x, q4.y), make_float2(q5.z, q5.w))));
    const int i0 = first + k, i1 = first + k + 1;
    // exact-t ties: the leaf's quads are stored in the reference's visiting order (flatten.cpp) and come
    // after its spheres, so `t <= closest` alone gives "a later quad replaces an equal-t hit" (quad.rs:72)
    const bool ok0 = (fabsf(denom.x) >= 1e-8f) & (t.x >= tmin) & (t.x <= h.t) & (al.x >= 0.0f) & (al.x <= 1.0f) & (be.x >= 0.0f) & (be.x <= 1.0f) & (i0 != skip);
    if (ok0) { h.t = t.x; h.prim = i0 | PRIM_QUAD_BIT; }
    const bool ok1 = (fabsf(denom.y) >= 1e-8f) & (t.y >= tmin) & (t.y <= h.t) & (al.y >= 0.0f) & (al.y <= 1.0f) & (be.y >= 0.0f) & (be.y <= 1.0f) & (i1 != skip);
    if (ok1) { h.t = t.y; h.prim = i1 | PRIM_QUAD_BIT; }
  }
}

// Axis-aligned rectangles of a brute-force root leaf (device_types.h DevScene::rects), normal along axis A: Quad::hit
// (quad.rs:64-91) collapses to t = (plane - o_a) / d_a and two interval tests on the other two coordinates of the hit
// point — a dozen instructions against 35 for a quad of a pair.  |d_a| < 1e-8 is quad.rs:67's parallel test (the
// normal is a unit axis), alpha / beta in [0, 1] (quad.rs:51-60) is |p - centre| <= half extent.
template <int A>
__device__ __forceinline__ void rect_quads_hit32(const float4* __restrict__ rec, int n, int index, const Ray32& r, float tmin, int skip, Hit32& h) {
  const float oa = A == 0 ? r.o.x : (A == 1 ? r.o.y : r.o.z), da = A == 0 ? r.d.x : (A == 1 ? r.d.y : r.d.z);
  const float ob = A == 0 ? r.o.y : (A == 1 ? r.o.z : r.o.x), db = A == 0 ? r.d.y : (A == 1 ? r.d.z : r.d.x);
  const float oc = A == 0 ? r.o.z : (A == 1 ? r.o.x : r.o.y), dc = A == 0 ? r.d.z : (A == 1 ? r.d.x : r.d.y);
  const float ida = 1.0f / da;
  const bool steep = fabsf(da) >= 1e-8f;
  for (int k = 0; k < n; ++k, ++index) {
    const float4 r0 = ldg4(rec + 2 * k), r1 = ldg4(rec + 2 * k + 1);
    const float t = (r0.x - oa) * ida;
    const float pb = fmaf(t, db, ob), pc = fmaf(t, dc, oc);
    const bool ok = steep & (t >= tmin) & (t <= h.t) & (fabsf(pb - r0.y) <= r1.x) & (fabsf(pc - r0.z) <= r1.y) & (index != skip);
    if (ok) { h.t = t; h.prim = index | PRIM_QUAD_BIT; }
  }
}

// The brute-force root leaf (scenes of at most RRT_SINGLE_LEAF_MAX primitives): spheres one by one, axis-aligned
// rectangles by their own test, the other quads as pairs when the flattener supplied pair records.
template <int FEAT>
__device__ __forceinline__ void root_leaf_hit32(const DevScene& S, int leaf, const Ray32& r, float a, float inv_a, float tmin,
                                                int skip, bool motion, Hit32& h) {
  int first = leaf >> 8, ns = (leaf >> 4) & 15, nq = leaf & 15;
  for (int k = 0; k < ns; ++k) sphere_hit32(S.prims, first + k, r, a, inv_a, tmin, skip, motion, h);
  first += ns;
  if (FEAT & SF_QUADS) {
    // not in the instances with box leaves (the book-2 final scene): a scene with box leaves has a tree, and
    // those instances sit at their register limit
    if (!(FEAT & SF_BOXES)) {
      if (S.rects) {   // the leaf's first quads, grouped by the axis of their normal
        const int nx = S.n_rects[0], ny = S.n_rects[1], nz = S.n_rects[2];
        if (nx > 0) rect_quads_hit32<0>(S.rects, nx, first, r, tmin, skip, h);
        if (ny > 0) rect_quads_hit32<1>(S.rects + 2 * nx, ny, first + nx, r, tmin, skip, h);
        if (nz > 0) rect_quads_hit32<2>(S.rects + 2 * (nx + ny), nz, first + nx + ny, r, tmin, skip, h);
        first += nx + ny + nz; nq -= nx + ny + nz;
      }
      if (S.qpairs) { quad_pairs_hit32(S.prims, S.qpairs, first, nq, r, tmin, skip, h); nq = 0; }
    }
    for (int k = 0; k < nq; ++k) quad_hit32(S.prims, first + k, r, tmin, skip, h);
  }
}

// A leaf below the root holds exactly one primitive (or one box): the flattener builds the tree down
// to single primitives (flatten.hpp max_leaf = 1), so the traversal loop needs no per-leaf loops.
template <int FEAT>
__device__ __forceinline__ void leaf1_hit32(const float4* __restrict__ prims, const float4* __restrict__ boxes, int leaf,
                                            const Ray32& r, V3f id, float a, float inv_a, float tmin, int skip, bool motion, Hit32& h) {
  if ((FEAT & SF_BOXES) && (leaf & 0xFF) == LEAF_BOX_MARK) { box_leaf_hit32(prims, boxes, leaf >> 8, r, id, tmin, skip, h); return; }
  if (!(FEAT & SF_QUADS) || (leaf & 0x10)) sphere_hit32(prims, leaf >> 8, r, a, inv_a, tmin, skip, motion, h);
  else quad_hit32(prims, leaf >> 8, r, tmin, skip, h);
}

// One uniform grid of an even field (device_types.h DevGrid, flatten.cpp choose_grids / emit_grids): clip the ray
// to the grid, then walk the cells it crosses front to back (Amanatides & Woo) and test what each cell
// references — spheres (KIND 0) or boxes (KIND 1, the box-leaf test) — until the next cell begins behind the
// closest hit.  An item referenced by several cells is tested in each: sphere_hit32 / quad_hit32 do not
// depend on order or repetition, so the hit and the equal-t rule are those of the tree walk.
template <int FEAT, int KIND>
__device__ __forceinline__ void grid_hit32(const DevScene& S, const DevGrid& G, const Ray32& r, float idx, float idy, float idz,
                                           float a, float inv_a, float tmin, int skip, bool motion, Hit32& h) {
  // the field's bounds at the ray's time (device_types.h DevGrid::clo ..): moving spheres sweep a taller slab than they fill
#ifdef RRT_NO_TIME_CLIP
  float lx = G.lo[0], ly = G.lo[1], lz = G.lo[2], hx = G.hi[0], hy = G.hi[1], hz = G.hi[2];
  if (false) {
#else
  const bool tc = (FEAT & SF_GRID_MOVING) != 0;   // (instances without the bit clip to the bounds of the whole shutter interval)
  float lx = tc ? G.clo[0] : G.lo[0], ly = tc ? G.clo[1] : G.lo[1], lz = tc ? G.clo[2] : G.lo[2];
  float hx = tc ? G.chi[0] : G.hi[0], hy = tc ? G.chi[1] : G.hi[1], hz = tc ? G.chi[2] : G.hi[2];
  if ((FEAT & SF_GRID_MOVING) && motion && KIND == 0 && G.moving) {
#endif
    lx = fmaf(r.tm, G.dclo[0], lx); ly = fmaf(r.tm, G.dclo[1], ly); lz = fmaf(r.tm, G.dclo[2], lz);
    hx = fmaf(r.tm, G.dchi[0], hx); hy = fmaf(r.tm, G.dchi[1], hy); hz = fmaf(r.tm, G.dchi[2], hz);
  }
  const float x0 = (lx - r.o.x) * idx, x1 = (hx - r.o.x) * idx;
  const float y0 = (ly - r.o.y) * idy, y1 = (hy - r.o.y) * idy;
  const float z0 = (lz - r.o.z) * idz, z1 = (hz - r.o.z) * idz;
  const float t0 = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
  const float t1 = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), h.t));
  if (!(t0 <= t1)) return;
  const int rx = G.res[0], ry = G.res[1], rz = G.res[2];
  int ix = min(max((int)((fmaf(t0, r.d.x, r.o.x) - G.lo[0]) * G.inv[0]), 0), rx - 1);
  int iy = min(max((int)((fmaf(t0, r.d.y, r.o.y) - G.lo[1]) * G.inv[1]), 0), ry - 1);
  int iz = min(max((int)((fmaf(t0, r.d.z, r.o.z) - G.lo[2]) * G.inv[2]), 0), rz - 1);
  const bool px = r.d.x >= 0.0f, py = r.d.y >= 0.0f, pz = r.d.z >= 0.0f;
  // parameter at which the ray leaves the current cell along each axis, and its increment per cell
  float tmx = (fmaf((float)(ix + (px ? 1 : 0)), G.cell[0], G.lo[0]) - r.o.x) * idx;
  float tmy = (fmaf((float)(iy + (py ? 1 : 0)), G.cell[1], G.lo[1]) - r.o.y) * idy;
  float tmz = (fmaf((float)(iz + (pz ? 1 : 0)), G.cell[2], G.lo[2]) - r.o.z) * idz;
  const float tdx = G.cell[0] * fabsf(idx), tdy = G.cell[1] * fabsf(idy), tdz = G.cell[2] * fabsf(idz);
  const int sx = px ? 1 : -1, sy = py ? rx : -rx, sz = pz ? rx * ry : -(rx * ry);
  // cells still to cross along each axis before the ray leaves the grid
  int nx = px ? rx - 1 - ix : ix, ny = py ? ry - 1 - iy : iy, nz = pz ? rz - 1 - iz : iz;
  int cell = G.first_cell + (iz * ry + iy) * rx + ix;
  for (;;) {
    const unsigned hdr = __ldg(S.grid_cells + cell);
    const int* refs = S.grid_refs + (hdr >> 8);
    const int cnt = (int)(hdr & 255u);
#ifndef RRT_GRID_UNROLL
#pragma unroll 1   // few lanes are in the same cell loop together: copies of the test only cost instruction cache
#endif
    for (int k = 0; k < cnt; ++k) {
      if (KIND == 0) sphere_hit32(S.prims, __ldg(refs + k), r, a, inv_a, tmin, skip, motion, h);
      else box_leaf_hit32(S.prims, S.boxes, __ldg(refs + k), r, V3f(idx, idy, idz), tmin, skip, h);
    }
    const bool bx = tmx <= tmy && tmx <= tmz, by = !bx && tmy <= tmz;
    const float tn = bx ? tmx : (by ? tmy : tmz);
    int left;
    if (bx) { left = nx--; tmx += tdx; cell += sx; }
    else if (by) { left = ny--; tmy += tdy; cell += sy; }
    else { left = nz--; tmz += tdz; cell += sz; }
    if (left <= 0 || tn > fminf(t1, h.t)) break;
  }
}

// World::hit for the render kernel: closest hit in (tmin, +inf).
// Nodes and primitives come through L1 (ld.global.nc).
template <int FEAT>
__device__ inline void world_hit32(const DevScene& S, const Ray32& r, float tmin, int skip, Hit32& h) {
  const float a = dot(r.d, r.d);
  const float inv_a = 1.0f / a;
  const bool motion = (FEAT & SF_MOTION) && S.has_motion != 0;
  if (FEAT & SF_GRID) {   // even fields in grids, everything else in the front run: no tree to walk (device_types.h SF_GRID)
    if (S.front_ref != -1) leaf_hit32<FEAT>(S.prims, S.boxes, ~S.front_ref, r, V3f(0.0f, 0.0f, 0.0f), a, inv_a, tmin, skip, motion, h);
    const float gx = safe_rcp(r.d.x), gy = safe_rcp(r.d.y), gz = safe_rcp(r.d.z);
    if (S.grids[0].n_items > 0) grid_hit32<FEAT, 0>(S, S.grids[0], r, gx, gy, gz, a, inv_a, tmin, skip, motion, h);
    if ((FEAT & SF_BOXES) && S.grids[1].n_items > 0) grid_hit32<FEAT, 1>(S, S.grids[1], r, gx, gy, gz, a, inv_a, tmin, skip, motion, h);
    return;
  }
  int cur = S.root_ref;
  if (cur < 0) {  // the whole scene is one leaf: brute force, uniform control flow (a root leaf is never a box leaf)
    root_leaf_hit32<FEAT>(S, ~cur, r, a, inv_a, tmin, skip, motion, h);
    return;
  }
  // front run (flatten.cpp): primitives spanning most of the scene, tested by all lanes together
  if (S.front_ref != -1) leaf_hit32<FEAT>(S.prims, S.boxes, ~S.front_ref, r, V3f(0.0f, 0.0f, 0.0f), a, inv_a, tmin, skip, motion, h);
  const float idx = safe_rcp(r.d.x), idy = safe_rcp(r.d.y), idz = safe_rcp(r.d.z);
  float oodx = r.o.x * idx, oody = r.o.y * idy, oodz = r.o.z * idz;
  asm("" : "+f"(oodx), "+f"(oody), "+f"(oodz));  // opaque: keep in registers, do not rematerialise in the loop
  int stack[RRT_STACK];
  int sp = 0;
#ifdef RRT_BVH4
  // Experiment: 4-wide nodes (flatten.cpp, -DRRT_BVH4): one step tests four boxes (two packed slab tests),
  // orders the hits with a five-exchange network, descends into the nearest and pushes the rest.
  const char* const nodes4_base = reinterpret_cast<const char*>(S.nodes4);
  for (;;) {
    bool done = false;
    while (cur >= 0) {
      const float4* g = reinterpret_cast<const float4*>(nodes4_base + (size_t)(unsigned)cur * 128u);
      const float4 q0 = ldg4(g), q1 = ldg4(g + 1), q2 = ldg4(g + 2), q3 = ldg4(g + 3), q4 = ldg4(g + 4), q5 = ldg4(g + 5), q6 = ldg4(g + 6);
      float n0, f0, n1, f1, n2, f2, n3, f3;
      slab2(q0, q1, q2, idx, idy, idz, oodx, oody, oodz, tmin, h.t, n0, f0, n1, f1);
      slab2(q3, q4, q5, idx, idy, idz, oodx, oody, oodz, tmin, h.t, n2, f2, n3, f3);
      int r0 = __float_as_int(q6.x), r1 = __float_as_int(q6.y), r2 = __float_as_int(q6.z), r3 = __float_as_int(q6.w);
      if (!(n0 <= f0 * 1.0000004f)) n0 = CUDART_INF_F;
      if (!(n1 <= f1 * 1.0000004f)) n1 = CUDART_INF_F;
      if (!(n2 <= f2 * 1.0000004f)) n2 = CUDART_INF_F;
      if (!(n3 <= f3 * 1.0000004f)) n3 = CUDART_INF_F;
#define RRT_CSWAP(a, b, ra, rb) { const bool s_ = b < a; const float ta_ = s_ ? b : a, tb_ = s_ ? a : b; const int ia_ = s_ ? rb : ra, ib_ = s_ ? ra : rb; a = ta_; b = tb_; ra = ia_; rb = ib_; }
      RRT_CSWAP(n0, n1, r0, r1) RRT_CSWAP(n2, n3, r2, r3) RRT_CSWAP(n0, n2, r0, r2) RRT_CSWAP(n1, n3, r1, r3) RRT_CSWAP(n1, n2, r1, r2)
#undef RRT_CSWAP
      if (n0 < CUDART_INF_F) {   // nearest first; the others wait on the stack, farthest pushed first
        if (n3 < CUDART_INF_F && sp < RRT_STACK) stack[sp++] = r3;
        if (n2 < CUDART_INF_F && sp < RRT_STACK) stack[sp++] = r2;
        if (n1 < CUDART_INF_F && sp < RRT_STACK) stack[sp++] = r1;
        cur = r0;
      } else {
        if (sp == 0) { done = true; break; }
        cur = stack[--sp];
      }
    }
    if (done) break;
    if (!(FEAT & SF_QUADS)) {
      while (cur < 0) {
        leaf1_hit32<FEAT>(S.prims, S.boxes, ~cur, r, V3f(idx, idy, idz), a, inv_a, tmin, skip, motion, h);
        if (sp == 0) { done = true; break; }
        cur = stack[--sp];
      }
      if (done) break;
    } else if (cur < 0) {
      leaf1_hit32<FEAT>(S.prims, S.boxes, ~cur, r, V3f(idx, idy, idz), a, inv_a, tmin, skip, motion, h);
      if (sp == 0) break;
      cur = stack[--sp];
    }
  }
  return;
#endif
  const char* const nodes_base = reinterpret_cast<const char*>(S.nodes);
  // while-while traversal (Aila & Laine 2009): every lane first descends through inner nodes until
  // it holds a leaf, then the warp reconverges and processes leaves together; the single mixed
  // loop this replaces ran the inner-node and the leaf body serially whenever lanes disagreed
  // (warp execution efficiency 42 % in profiles/ncu_r01_*).  Measured alternatives that lost: postponing
  // a found leaf and descending speculatively to the next one (-25 %, wasted node visits), a lane
  // state machine voting on the next step every iteration (+35 % time, vote overhead), and leaving
  // this inner loop early once fewer than N lanes are still descending (+4 % for every N).
  for (;;) {
    bool done = false;
    while (cur >= 0) {
      float4 n0, n1, n2, n3;
      const float4* g = reinterpret_cast<const float4*>(nodes_base + (size_t)(unsigned)cur * 64u);
      n0 = ldg4(g); n1 = ldg4(g + 1); n2 = ldg4(g + 2); n3 = ldg4(g + 3);
      float near0, far0, near1, far1;
      slab2(n0, n1, n2, idx, idy, idz, oodx, oody, oodz, tmin, h.t, near0, far0, near1, far1);
      bool hit0 = near0 <= far0 * 1.0000004f;
      bool hit1 = near1 <= far1 * 1.0000004f;
      int r0 = __float_as_int(n3.x), r1 = __float_as_int(n3.y);
      if (hit0 && hit1) {
        if (near1 < near0) { int t = r0; r0 = r1; r1 = t; }
        if (sp < RRT_STACK) stack[sp++] = r1;
        cur = r0;
      } else if (hit0) {
        cur = r0;
      } else if (hit1) {
        cur = r1;
      } else {
        if (sp == 0) { done = true; break; }
        cur = stack[--sp];
      }
    }
    if (done) break;
    // Sphere-only instances run consecutive leaves of a lane (two sibling leaves, a popped leaf after a
    // leaf) in the same round instead of sending the lane through an empty descent between them
    // (C1 +3.5 %, C2 +1.3 %); with box leaves in the tree the longer rounds cost more than they save (C4 -3 %).
    if (!(FEAT & SF_QUADS)) {
      while (cur < 0) {
        leaf1_hit32<FEAT>(S.prims, S.boxes, ~cur, r, V3f(idx, idy, idz), a, inv_a, tmin, skip, motion, h);
        if (sp == 0) { done = true; break; }
        cur = stack[--sp];
      }
      if (done) break;
    } else if (cur < 0) {
      leaf1_hit32<FEAT>(S.prims, S.boxes, ~cur, r, V3f(idx, idy, idz), a, inv_a, tmin, skip, motion, h);
      if (sp == 0) break;
      cur = stack[--sp];
    }
  }
}
// Both boundary hits of ConstantMedium::hit (constant_medium.rs:41-42) when the boundary is one sphere:
// boundary.hit(r, UNIVERSE) returns the smaller root whatever its sign (sphere.rs:81-89), and the second
// call on (t1 + 0.0001, inf) can then only return the larger one — one discriminant instead of two.
__device__ inline bool boundary_sphere_span32(const float4* __restrict__ bprims, const DevMedium& m, const Ray32& r,
                                              float a, float inv_a, bool motion, float& t1, float& t2) {
  const float4 p0 = ldg4(bprims + 4 * m.first);
  V3f c(p0.x, p0.y, p0.z);
  if (motion) { float4 p1 = ldg4(bprims + 4 * m.first + 1); c = c + r.tm * V3f(p1.x, p1.y, p1.z); }
  const V3f oc = c - r.o;
  const float hh = dot(r.d, oc);
  const V3f perp = oc - (hh * inv_a) * r.d;
  const float disc = p0.w * p0.w - dot(perp, perp);
  if (disc < 0.0f) return false;
  const float sq = sqrtf(disc * a);
  t1 = (hh - sq) * inv_a;
  t2 = (hh + sq) * inv_a;
  return t2 > t1 + 0.0001f;
}

// Closest hit of a medium boundary run in [lo, hi] for quads / (lo, hi) for spheres
// (boundary.hit(r, interval), constant_medium.rs:41-42).  Returns t or -1 flagged by `found`.
__device__ inline bool boundary_hit32(const float4* __restrict__ bprims, const DevMedium& m, const Ray32& r,
                                      float a, float inv_a, float lo, float hi, bool motion, float& t_out) {
  bool found = false;
  float best = hi;
  int first = m.first;
  for (int k = 0; k < m.nsph; ++k) {
    int i = first + k;
    float4 p0 = ldg4(bprims + 4 * i);
    V3f c(p0.x, p0.y, p0.z);
    if (motion) { float4 p1 = ldg4(bprims + 4 * i + 1); c = c + r.tm * V3f(p1.x, p1.y, p1.z); }
    V3f oc = c - r.o;
    float hh = dot(r.d, oc);
    V3f perp = oc - (hh * inv_a) * r.d;
    float disc = p0.w * p0.w - dot(perp, perp);
    if (disc < 0.0f) continue;
    float sq = sqrtf(disc * a);
    float root = (hh - sq) * inv_a;
    if (!(root > lo && root < best)) {
      root = (hh + sq) * inv_a;
      if (!(root > lo && root < best)) continue;
    }
    best = root; found = true;
  }
  first += m.nsph;
  for (int k = 0; k < m.nquad; ++k) {
    int i = first + k;
    float4 p0 = ldg4(bprims + 4 * i);
    V3f n(p0.x, p0.y, p0.z);
    float denom = dot(n, r.d);
    if (fabsf(denom) < 1e-8f) continue;
    float t = (p0.w - dot(n, r.o)) / denom;
    if (!(t >= lo && t <= best)) continue;
    V3f p = r.o + t * r.d;
    float4 p1 = ldg4(bprims + 4 * i + 1);
    float alpha = dot(p, V3f(p1.x, p1.y, p1.z)) + p1.w;
    if (!(alpha >= 0.0f && alpha <= 1.0f)) continue;
    float4 p2 = ldg4(bprims + 4 * i + 2);
    float beta = dot(p, V3f(p2.x, p2.y, p2.z)) + p2.w;
    if (!(beta >= 0.0f && beta <= 1.0f)) continue;
    best = t; found = true;
  }
  t_out = best;
  return found;
}

// ------------------------------------------------------------------------------------------------
// f64 probe path: the reference's formulas verbatim on the f64 prim records.
struct Hit64 {
  double t;
  int prim;  // flat index | PRIM_QUAD_BIT, -1 none
};

__device__ inline void sphere_hit64(const double* __restrict__ q, int i, V3d o, V3d d, double tmin,
                                    const float4* prims, Hit64& h) {
  const double* s = q + 16 * i;
  V3d c(s[0], s[1], s[2]);  // probe time is 0: center.at(0) = center0
  V3d oc = c - o;
  double a = dot(d, d);
  double hh = dot(d, oc);
  double cc = dot(oc, oc) - s[3] * s[3];
  double disc = hh * hh - a * cc;  // sphere.rs:76
  if (disc < 0.0) return;
  double sq = sqrt(disc);
  double root = (hh - sq) / a;
  if (!(root > tmin && root <= h.t)) {
    root = (hh + sq) / a;
    if (!(root > tmin && root <= h.t)) return;
  }
  if (root == h.t && !tie_accept(prims, i, h.prim)) return;
  h.t = root; h.prim = i;
}

__device__ inline void quad_hit64(const double* __restrict__ q, int i, V3d o, V3d d, double tmin,
                                  const float4* prims, Hit64& h) {
  const double* s = q + 16 * i;
  V3d n(s[0], s[1], s[2]);
  double denom = dot(n, d);
  if (fabs(denom) < 1e-8) return;
  double t = (s[3] - dot(n, o)) / denom;
  if (!(t >= tmin && t <= h.t)) return;
  V3d p = o + t * d;
  V3d hp = p - V3d(s[4], s[5], s[6]);
  V3d u(s[7], s[8], s[9]), v(s[10], s[11], s[12]), w(s[13], s[14], s[15]);
  double alpha = dot(w, cross(hp, v));
  double beta = dot(w, cross(u, hp));
  if (!(alpha >= 0.0 && alpha <= 1.0 && beta >= 0.0 && beta <= 1.0)) return;
  if (t == h.t && !tie_accept(prims, i | PRIM_QUAD_BIT, h.prim)) return;
  h.t = t; h.prim = i | PRIM_QUAD_BIT;
}

__device__ inline void leaf_hit64(const DevScene& S, int leaf, V3d o, V3d d, double tmin, Hit64& h) {
  if ((leaf & 0xFF) == LEAF_BOX_MARK) {  // box leaf: the probe tests the six quads, no shortcut
    int first = __float_as_int(ldg4(S.boxes + 2 * (leaf >> 8)).w);
    for (int k = 0; k < 6; ++k) quad_hit64(S.prims64, first + k, o, d, tmin, S.prims, h);
    return;
  }
  int first = leaf >> 8, ns = (leaf >> 4) & 15, nq = leaf & 15;
  for (int k = 0; k < ns; ++k) sphere_hit64(S.prims64, first + k, o, d, tmin, S.prims, h);
  first += ns;
  for (int k = 0; k < nq; ++k) quad_hit64(S.prims64, first + k, o, d, tmin, S.prims, h);
}

__device__ inline void world_hit64(const DevScene& S, V3d o, V3d d, double tmin, Hit64& h) {
  int cur = S.root_ref;
  if (cur < 0) { leaf_hit64(S, ~cur, o, d, tmin, h); return; }
  if (S.front_ref != -1) leaf_hit64(S, ~S.front_ref, o, d, tmin, h);
  const double idx = 1.0 / d.x, idy = 1.0 / d.y, idz = 1.0 / d.z;
  int stack[RRT_STACK];
  int sp = 0;
  for (;;) {
    if (cur >= 0) {
      const float4* g = S.nodes + 4 * cur;
      float4 n0 = ldg4(g), n1 = ldg4(g + 1), n2 = ldg4(g + 2), n3 = ldg4(g + 3);
      // (center, half-extent) -> planes; the boxes are conservative supersets of the reference's
      const double cx0 = n0.x, cx1 = n0.y, cy0 = n0.z, cy1 = n0.w, cz0 = n1.x, cz1 = n1.y;
      const double ex0 = n1.z, ex1 = n1.w, ey0 = n2.x, ey1 = n2.y, ez0 = n2.z, ez1 = n2.w;
      double c0x0 = (cx0 - ex0 - o.x) * idx, c0x1 = (cx0 + ex0 - o.x) * idx;
      double c0y0 = (cy0 - ey0 - o.y) * idy, c0y1 = (cy0 + ey0 - o.y) * idy;
      double c0z0 = (cz0 - ez0 - o.z) * idz, c0z1 = (cz0 + ez0 - o.z) * idz;
      double c1x0 = (cx1 - ex1 - o.x) * idx, c1x1 = (cx1 + ex1 - o.x) * idx;
      double c1y0 = (cy1 - ey1 - o.y) * idy, c1y1 = (cy1 + ey1 - o.y) * idy;
      double c1z0 = (cz1 - ez1 - o.z) * idz, c1z1 = (cz1 + ez1 - o.z) * idz;
      double near0 = fmax(fmax(fmin(c0x0, c0x1), fmin(c0y0, c0y1)), fmax(fmin(c0z0, c0z1), tmin));
      double far0 = fmin(fmin(fmax(c0x0, c0x1), fmax(c0y0, c0y1)), fmin(fmax(c0z0, c0z1), h.t));
      double near1 = fmax(fmax(fmin(c1x0, c1x1), fmin(c1y0, c1y1)), fmax(fmin(c1z0, c1z1), tmin));
      double far1 = fmin(fmin(fmax(c1x0, c1x1), fmax(c1y0, c1y1)), fmin(fmax(c1z0, c1z1), h.t));
      bool hit0 = near0 <= far0, hit1 = near1 <= far1;
      int r0 = __float_as_int(n3.x), r1 = __float_as_int(n3.y);
      if (hit0 && hit1) {
        if (near1 < near0) { int t = r0; r0 = r1; r1 = t; }
        if (sp < RRT_STACK) stack[sp++] = r1;
        cur = r0;
        continue;
      }
      if (hit0) { cur = r0; continue; }
      if (hit1) { cur = r1; continue; }
    } else {
      leaf_hit64(S, ~cur, o, d, tmin, h);
    }
    if (sp == 0) break;
    cur = stack[--sp];
  }
}

}  // namespace rrt
