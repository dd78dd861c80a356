// vecmath.cuh — f32/f64 3-vectors for the kernels (Vec3 of the reference is f64 x 3,
// the_rest_of_your_life/vec3.rs:5-8; the render kernel computes in f32, the probe kernel in f64).
#pragma once
#include <cuda_runtime.h>

namespace rrt {

template <typename T>
struct V3 {
  T x, y, z;
  __host__ __device__ V3() {}
  __host__ __device__ V3(T a, T b, T c) : x(a), y(b), z(c) {}
};
typedef V3<float> V3f;
typedef V3<double> V3d;

template <typename T> __host__ __device__ inline V3<T> operator+(V3<T> a, V3<T> b) { return V3<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> __host__ __device__ inline V3<T> operator-(V3<T> a, V3<T> b) { return V3<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> __host__ __device__ inline V3<T> operator-(V3<T> a) { return V3<T>(-a.x, -a.y, -a.z); }
template <typename T> __host__ __device__ inline V3<T> operator*(V3<T> a, V3<T> b) { return V3<T>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename T> __host__ __device__ inline V3<T> operator*(T s, V3<T> a) { return V3<T>(s * a.x, s * a.y, s * a.z); }
template <typename T> __host__ __device__ inline V3<T> operator*(V3<T> a, T s) { return V3<T>(s * a.x, s * a.y, s * a.z); }
template <typename T> __host__ __device__ inline T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> __host__ __device__ inline V3<T> cross(V3<T> a, V3<T> b) {
  return V3<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ inline V3f normalize(V3f a) { float inv = rsqrtf(dot(a, a)); return inv * a; }
__device__ inline V3f xyz(float4 v) { return V3f(v.x, v.y, v.z); }
__device__ inline float max3(V3f a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }

__device__ inline float4 ldg4(const float4* p) { return __ldg(p); }

}  // namespace rrt
