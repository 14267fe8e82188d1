// vecmath.cuh -- fp32 3-vector helpers whose operation ORDER mirrors the reference's vec3
// (src/model/vector_and_matrix/vec3.h) so that, compiled with -fmad=false, results are
// bit-identical to the host replay.  Nothing here is copied: these are the obvious
// component-wise definitions, written to keep the same association order
// ((a0*b0 + a1*b1) + a2*b2 for dot, v / length(v) for normalisation, ...).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

struct V3 {
  float x, y, z;
};

__host__ __device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__host__ __device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__host__ __device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__host__ __device__ __forceinline__ V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
__host__ __device__ __forceinline__ V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
__host__ __device__ __forceinline__ V3 operator*(float t, V3 a) { return V3{t * a.x, t * a.y, t * a.z}; }
__host__ __device__ __forceinline__ V3 operator*(V3 a, float t) { return V3{t * a.x, t * a.y, t * a.z}; }
__host__ __device__ __forceinline__ V3 operator/(V3 a, float t) { return V3{a.x / t, a.y / t, a.z / t}; }
__host__ __device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return V3{a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x};
}
__host__ __device__ __forceinline__ float sqlen(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
__host__ __device__ __forceinline__ float length(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
__host__ __device__ __forceinline__ V3 unit(V3 a) { return a / length(a); }
// vec3::make_unit_vector: k = 1.0 / sqrt(.) (double division rounded to float == IEEE fp32
// division), then three multiplies (vec3.h:96-99)
__host__ __device__ __forceinline__ V3 make_unit(V3 a) {
  float k = 1.0f / sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
  return V3{a.x * k, a.y * k, a.z * k};
}
__host__ __device__ __forceinline__ float dist(V3 a, V3 b) { return length(a - b); }
__host__ __device__ __forceinline__ bool has_nan(V3 a) { return isnan(a.x) || isnan(a.y) || isnan(a.z); }
__host__ __device__ __forceinline__ bool has_inf(V3 a) { return isinf(a.x) || isinf(a.y) || isinf(a.z); }
__host__ __device__ __forceinline__ V3 de_nan(V3 a) {
  return V3{a.x == a.x ? a.x : 0.f, a.y == a.y ? a.y : 0.f, a.z == a.z ? a.z : 0.f};
}
__host__ __device__ __forceinline__ float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
