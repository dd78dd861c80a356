// shade.cuh — textures, hit-record reconstruction, light sampling (f32).
//
// Restates:
//   TextureObject::value   the_rest_of_your_life/texture.rs:33-37, 63-77, 88-109, 121-126
//   Perlin::noise/turb     the_rest_of_your_life/perlin.rs:26-64, 79-98
//   RtwImage::pixel_data   the_rest_of_your_life/rtw_image.rs:46-55, 70-78
//   get_sphere_uv          the_rest_of_your_life/sphere.rs:47-53
//   Onb                    the_rest_of_your_life/onb.rs:7-34
//   Quad/Sphere pdf_value, random   quad.rs:93-107, sphere.rs:55-65, 102-122
//   HittableList pdf_value/random   hittable_list.rs:60-75
#pragma once
#include "device_types.h"
#include "intersect.cuh"
#include "rng.cuh"
#include "vecmath.cuh"

namespace rrt {

#define RRT_PI 3.14159265358979323846f

// ---- Perlin -----------------------------------------------------------------------------------
// The tables of the scene's first Perlin object are staged in shared memory by the render kernel (4.9 KB:
// perlin.rs:4-9's 256 gradients + three permutations): 14 table reads per octave, 7 octaves per lookup,
// and a shared-memory read needs a 32-bit offset where a global one needs 64-bit address arithmetic —
// a quarter of the 140 instructions of an octave in the global version (profiles/README.md r02a, C3c:
// 68 % of the kernel is this function).  SHARED selects the address space at compile time.
template <bool SHARED>
__device__ __forceinline__ float4 perlin_grad(const DevPerlin* P, int i) {
  if (SHARED) {
    float4 g;
    const unsigned a = (unsigned)__cvta_generic_to_shared(&P->randvec[i]);
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w) : "r"(a));
    return g;
  }
  return ldg4(&P->randvec[i]);
}
template <bool SHARED>
__device__ __forceinline__ int perlin_perm(const DevPerlin* P, int axis, int i) {
  if (SHARED) {
    unsigned v;
    const unsigned a = (unsigned)__cvta_generic_to_shared(&P->perm[axis][i]);
    asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return (int)v;
  }
  return P->perm[axis][i];
}

template <bool SHARED>
__device__ inline float perlin_noise(const DevPerlin* __restrict__ P, V3f p) {
  float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
  float u = p.x - fx, v = p.y - fy, w = p.z - fz;
  int i = (int)fx, j = (int)fy, k = (int)fz;
  float uu = u * u * (3.0f - 2.0f * u);
  float vv = v * v * (3.0f - 2.0f * v);
  float ww = w * w * (3.0f - 2.0f * w);
  // perlin_interp (perlin.rs:78-96) is the trilinear blend of the eight corner products with weights uu, vv, ww; written
  // as nested lerps (along z, then y, then x: 7 x 2 operations) instead of eight triple products and sums (C3c +3 %)
  const int pz0 = perlin_perm<SHARED>(P, 2, k & 255), pz1 = perlin_perm<SHARED>(P, 2, (k + 1) & 255);
  float along_x[2];
#pragma unroll
  for (int di = 0; di < 2; ++di) {
    const int px = perlin_perm<SHARED>(P, 0, (i + di) & 255);
    float along_y[2];
#pragma unroll
    for (int dj = 0; dj < 2; ++dj) {
      const int pxy = px ^ perlin_perm<SHARED>(P, 1, (j + dj) & 255);
      const float4 g0 = perlin_grad<SHARED>(P, pxy ^ pz0), g1 = perlin_grad<SHARED>(P, pxy ^ pz1);
      const float d0 = g0.x * (u - di) + g0.y * (v - dj) + g0.z * w;
      const float d1 = g1.x * (u - di) + g1.y * (v - dj) + g1.z * (w - 1.0f);
      along_y[dj] = fmaf(ww, d1 - d0, d0);
    }
    along_x[di] = fmaf(vv, along_y[1] - along_y[0], along_y[0]);
  }
  return fmaf(uu, along_x[1] - along_x[0], along_x[0]);
}

template <bool SHARED>
__device__ inline float perlin_turb(const DevPerlin* __restrict__ P, V3f p, int depth) {
  float accum = 0.0f, weight = 1.0f;
  for (int i = 0; i < depth; ++i) {
    accum += weight * perlin_noise<SHARED>(P, p);
    weight *= 0.5f;
    p = 2.0f * p;
  }
  return fabsf(accum);
}

// ---- textures ---------------------------------------------------------------------------------
// sh_perlin: the shared-memory copy of S.perlins[0], or null (then every Perlin object is read from global memory)
template <int FEAT>
__device__ inline V3f texture_value(const DevScene& S, int tex, float u, float v, V3f p, const DevPerlin* sh_perlin) {
  for (int guard = 0; guard < 9; ++guard) {
    const DevTexture t = S.textures[tex];
    if (t.kind == RRT_T_CHECKER) {  // texture.rs:63-77
      float inv = t.v.w;
      int xi = (int)floorf(inv * p.x), yi = (int)floorf(inv * p.y), zi = (int)floorf(inv * p.z);
      tex = ((xi + yi + zi) & 1) == 0 ? t.a : t.b;  // Rust: % 2 == 0 (−1 % 2 = −1 → odd)
      continue;
    }
    if (t.kind == RRT_T_SOLID) return V3f(t.v.x, t.v.y, t.v.z);
    if (t.kind == RRT_T_IMAGE) {  // texture.rs:88-109
      const DevImage im = S.images[t.a];
      if (im.height <= 0) return V3f(0.0f, 1.0f, 1.0f);
      u = fminf(fmaxf(u, 0.0f), 1.0f);
      v = 1.0f - fminf(fmaxf(v, 0.0f), 1.0f);
      int i = (int)(u * (float)im.width), j = (int)(v * (float)im.height);
      i = i < 0 ? 0 : (i < im.width ? i : im.width - 1);   // rtw_image.rs:70-78
      j = j < 0 ? 0 : (j < im.height ? j : im.height - 1);
      uchar4 px = __ldg(reinterpret_cast<const uchar4*>(S.texels) + im.offset + j * im.width + i);
      const float s = 1.0f / 255.0f;
      return V3f(s * px.x, s * px.y, s * px.z);
    }
    // NoiseTexture, texture.rs:121-126
    if (!(FEAT & SF_NOISE)) return V3f(0.0f, 0.0f, 0.0f);  // unreachable: the flattener sets SF_NOISE for such scenes
    const float turb = (sh_perlin && t.a == 0) ? perlin_turb<true>(sh_perlin, p, 7) : perlin_turb<false>(S.perlins + t.a, p, 7);
    float c = 0.5f * (1.0f + sinf(t.v.w * p.z + 10.0f * turb));
    return V3f(c, c, c);
  }
  return V3f(0.0f, 0.0f, 0.0f);
}

// ---- hit record -------------------------------------------------------------------------------
struct Surface {
  V3f p;        // hit point (re-projected onto the primitive)
  V3f n;        // normal facing the ray (HitRecord::set_face_normal, hittable.rs:49-52)
  V3f n_obj;    // outward normal in object space (for sphere uv)
  float u, v;
  int mat;
  bool front;
  bool is_sphere;
};

// get_sphere_uv, sphere.rs:47-53
__device__ inline void sphere_uv(V3f n, float& u, float& v) {
  float theta = acosf(fminf(fmaxf(-n.y, -1.0f), 1.0f));
  float phi = atan2f(-n.z, n.x) + RRT_PI;
  u = phi * (1.0f / (2.0f * RRT_PI));
  v = theta * (1.0f / RRT_PI);
}

// Rebuild the HitRecord of the closest primitive (sphere.rs:91-95, quad.rs:76-88).
template <int FEAT>
__device__ inline void make_surface(const DevScene& S, const Ray32& r, const Hit32& h, Surface& s) {
  int idx = h.prim & ~PRIM_QUAD_BIT;
  const float4* rec = S.prims + 4 * idx;
  V3f p = r.o + h.t * r.d;
  V3f outward;
  if ((FEAT & SF_QUADS) && (h.prim & PRIM_QUAD_BIT)) {
    float4 p0 = ldg4(rec), p1 = ldg4(rec + 1), p2 = ldg4(rec + 2), m = ldg4(rec + 3);
    outward = V3f(p0.x, p0.y, p0.z);
    s.u = dot(p, V3f(p1.x, p1.y, p1.z)) + p1.w;
    s.v = dot(p, V3f(p2.x, p2.y, p2.z)) + p2.w;
    p = p - (dot(outward, p) - p0.w) * outward;  // back onto the plane
    s.mat = __float_as_int(m.x);
    s.is_sphere = false;
    s.n_obj = outward;
  } else {
    float4 p0 = ldg4(rec), m = ldg4(rec + 2);
    V3f c(p0.x, p0.y, p0.z);
    if ((FEAT & SF_MOTION) && S.has_motion) { float4 p1 = ldg4(rec + 1); c = c + r.tm * V3f(p1.x, p1.y, p1.z); }
    outward = normalize(p - c);                     // (p - center) / radius, sphere.rs:93, renormalised in f32
    p = c + p0.w * outward;                         // back onto the sphere
    s.mat = __float_as_int(m.x);
    s.is_sphere = true;
    s.n_obj = outward;
    if (__float_as_int(m.w) & PF_UV_ROTATED) {      // world -> object, hittable.rs:144-154
      float4 p3 = ldg4(rec + 3);
      s.n_obj = V3f(p3.x * outward.x - p3.y * outward.z, outward.y, p3.y * outward.x + p3.x * outward.z);
    }
    s.u = 0.0f; s.v = 0.0f;
  }
  s.p = p;
  s.front = dot(r.d, outward) < 0.0f;
  s.n = s.front ? outward : -outward;
}

// ---- sampling helpers ---------------------------------------------------------------------------
// random_unit_vector (vec3.rs:181-189) is uniform on the sphere; drawn directly, not by rejection.
__device__ inline V3f uniform_sphere(float u1, float u2) {
  float z = 1.0f - 2.0f * u1;
  float rr = sqrtf(fmaxf(0.0f, 1.0f - z * z));
  float s, c;
  __sincosf(2.0f * RRT_PI * u2 - RRT_PI, &s, &c);
  return V3f(rr * c, rr * s, z);
}
// random_in_unit_disk (vec3.rs:172-179): uniform on the disk
__device__ inline void uniform_disk(float u1, float u2, float& x, float& y) {
  float rr = sqrtf(u1);
  float s, c;
  __sincosf(2.0f * RRT_PI * u2 - RRT_PI, &s, &c);
  x = rr * c; y = rr * s;
}
// random_cosine_direction, vec3.rs:212-222
__device__ inline V3f cosine_direction(float r1, float r2) {
  float s, c;
  __sincosf(2.0f * RRT_PI * r1 - RRT_PI, &s, &c);
  float q = sqrtf(r2);
  return V3f(c * q, s * q, sqrtf(fmaxf(0.0f, 1.0f - r2)));
}

struct Onb {  // onb.rs:7-18
  V3f u, v, w;
  __device__ explicit Onb(V3f n) {
    w = normalize(n);
    V3f a = fabsf(w.x) > 0.9f ? V3f(0.0f, 1.0f, 0.0f) : V3f(1.0f, 0.0f, 0.0f);
    v = normalize(cross(w, a));
    u = cross(w, v);
  }
  __device__ V3f transform(V3f t) const { return t.x * u + t.y * v + t.z * w; }
};

// ---- lights (book 3) ----------------------------------------------------------------------------
// HittableObject::random(origin) of one light
__device__ inline V3f light_random(const DevLight& L, V3f origin, float r1, float r2) {
  if (L.kind == 1) {  // quad.rs:104-107
    V3f p = V3f(L.q.x, L.q.y, L.q.z) + r1 * V3f(L.u.x, L.u.y, L.u.z) + r2 * V3f(L.v.x, L.v.y, L.v.z);
    return p - origin;
  }
  if (L.kind == 0) {  // sphere.rs:55-65, 117-122
    V3f dir = V3f(L.a.x, L.a.y, L.a.z) - origin;
    float dist2 = dot(dir, dir);
    Onb uvw(dir);
    float z = 1.0f + r2 * (sqrtf(fmaxf(0.0f, 1.0f - L.a.w * L.a.w / dist2)) - 1.0f);
    float s, c;
    __sincosf(2.0f * RRT_PI * r1 - RRT_PI, &s, &c);
    float q = sqrtf(fmaxf(0.0f, 1.0f - z * z));
    return uvw.transform(V3f(c * q, s * q, z));
  }
  return V3f(1.0f, 0.0f, 0.0f);  // trait default, hittable.rs:62-64
}

// HittableObject::pdf_value(origin, direction) of one light
__device__ inline float light_pdf(const DevLight& L, V3f origin, V3f dir) {
  if (L.kind == 1) {  // quad.rs:93-102: hit test on (0.001, inf), then dist^2 / (cos * area)
    V3f n(L.a.x, L.a.y, L.a.z);
    float denom = dot(n, dir);
    if (fabsf(denom) < 1e-8f) return 0.0f;
    float t = (L.a.w - dot(n, origin)) / denom;
    if (!(t >= 0.001f)) return 0.0f;
    V3f p = origin + t * dir;
    float alpha = dot(p, V3f(L.b.x, L.b.y, L.b.z)) + L.b.w;
    float beta = dot(p, V3f(L.c.x, L.c.y, L.c.z)) + L.c.w;
    if (!(alpha >= 0.0f && alpha <= 1.0f && beta >= 0.0f && beta <= 1.0f)) return 0.0f;
    float len2 = dot(dir, dir);
    float dist2 = t * t * len2;
    float cosine = fabsf(denom) * rsqrtf(len2);
    return dist2 / (cosine * L.q.w);
  }
  if (L.kind == 0) {  // sphere.rs:102-115
    V3f oc = V3f(L.a.x, L.a.y, L.a.z) - origin;
    float a = dot(dir, dir), hh = dot(dir, oc);
    float inv_a = 1.0f / a;
    V3f perp = oc - (hh * inv_a) * dir;
    float disc = L.a.w * L.a.w - dot(perp, perp);
    if (disc < 0.0f) return 0.0f;
    float sq = sqrtf(disc * a);
    float root = (hh - sq) * inv_a;
    if (!(root > 0.001f)) {
      root = (hh + sq) * inv_a;
      if (!(root > 0.001f)) return 0.0f;
    }
    float dist2 = dot(oc, oc);
    float cos_theta_max = sqrtf(1.0f - L.a.w * L.a.w / dist2);  // NaN inside the sphere, as the reference
    float solid_angle = 2.0f * RRT_PI * (1.0f - cos_theta_max);
    return 1.0f / solid_angle;
  }
  return 0.0f;
}

}  // namespace rrt
