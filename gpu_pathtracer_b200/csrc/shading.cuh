// shading.cuh -- RNG state, local frames, samplers, pdfs and Phong / Schlick helpers.
//
// Restates, operation for operation (including the places where the reference's literals
// promote to double), the helpers of
//   src/util/vector_util.h:48-300, src/model/cartesian_system.h:20-80,
//   src/model/ray/ray_operations.h:16-35 and src/util/general.h:8-24.
// The XORWOW generator itself is the CUDA toolkit's own <curand_kernel.h> (the reference
// uses the same header), so per-pixel streams are identical by construction.
#pragma once
#include <curand_kernel.h>

#include "common.cuh"
#include "vecmath.cuh"

#define GPT_PI 3.14159265358979323846      /* M_PI (double, as in the reference) */
#define GPT_MAX_PHONG_N_S 1e6f             /* src/param.h:7 */

// ---- RNG: compact 24-byte XORWOW state (d, v[0..4]) ------------------------------------
struct Rng {
  curandStateXORWOW s;
};
__device__ __forceinline__ void rng_load(Rng &r, const uint32_t *__restrict__ st, size_t slot,
                                         size_t stride) {
  r.s.d = st[slot];
#pragma unroll
  for (int k = 0; k < 5; k++) r.s.v[k] = st[(k + 1) * stride + slot];
  r.s.boxmuller_flag = 0;
  r.s.boxmuller_flag_double = 0;
  r.s.boxmuller_extra = 0.f;
  r.s.boxmuller_extra_double = 0.0;
}
__device__ __forceinline__ void rng_store(const Rng &r, uint32_t *st, size_t slot,
                                          size_t stride) {
  st[slot] = r.s.d;
#pragma unroll
  for (int k = 0; k < 5; k++) st[(k + 1) * stride + slot] = r.s.v[k];
}
// packed variant: 8 words per slot (one 32-byte sector), for state that is visited in random
// slot order
template <bool AROUND_L1 = false>
__device__ __forceinline__ void rng_load_packed(Rng &r, const uint32_t *st, size_t slot) {
  const uint4 *p = reinterpret_cast<const uint4 *>(st + 8 * slot);
  uint4 a = AROUND_L1 ? __ldcg(p) : p[0], b = AROUND_L1 ? __ldcg(p + 1) : p[1];
  r.s.d = a.x; r.s.v[0] = a.y; r.s.v[1] = a.z; r.s.v[2] = a.w; r.s.v[3] = b.x; r.s.v[4] = b.y;
  r.s.boxmuller_flag = 0;
  r.s.boxmuller_flag_double = 0;
  r.s.boxmuller_extra = 0.f;
  r.s.boxmuller_extra_double = 0.0;
}
__device__ __forceinline__ void rng_store_packed(const Rng &r, uint32_t *st, size_t slot) {
  uint4 *p = reinterpret_cast<uint4 *>(st + 8 * slot);
  p[0] = make_uint4(r.s.d, r.s.v[0], r.s.v[1], r.s.v[2]);
  p[1] = make_uint4(r.s.v[3], r.s.v[4], 0u, 0u);
}
__device__ __forceinline__ float rng_uniform(Rng &r) { return curand_uniform(&r.s); }

// ---- clamp(lo, hi, v) (general.h:23-24) ------------------------------------------------
__device__ __forceinline__ float clamp_lohi(float lo, float hi, float v) {
  return fmaxf(lo, fminf(hi, v));
}

// ---- CartesianSystem (cartesian_system.h:37-66) -----------------------------------------
struct Frame {
  V3 x, y, z;
};
__device__ __forceinline__ Frame frame_from_z(V3 zin) {
  Frame f;
  f.z = unit(zin);
  if (fabsf(f.z.x) > fabsf(f.z.y))
    f.x = V3{f.z.z, 0.f, -f.z.x} / sqrtf(f.z.x * f.z.x + f.z.z * f.z.z);
  else
    f.x = V3{0.f, -f.z.z, f.z.y} / sqrtf(f.z.y * f.z.y + f.z.z * f.z.z);
  f.y = unit(cross(f.z, f.x));
  return f;
}
__device__ __forceinline__ Frame frame_from_zx(V3 zin, V3 xdraft) {
  Frame f;
  xdraft = unit(xdraft);
  f.z = unit(zin);
  f.y = unit(cross(f.z, xdraft));
  f.x = unit(cross(f.y, f.z));
  return f;
}
// to_world_system (cartesian_system.h:52-80), result re-normalised with make_unit_vector
__device__ __forceinline__ V3 frame_to_world(const Frame &f, V3 v) {
  V3 w = V3{v.x * f.x.x + v.y * f.y.x + v.z * f.z.x, v.x * f.x.y + v.y * f.y.y + v.z * f.z.y,
            v.x * f.x.z + v.y * f.y.z + v.z * f.z.z};
  return make_unit(w);
}

// ---- samplers (vector_util.h:52-54,248-300) ---------------------------------------------
__device__ __forceinline__ bool ns_too_large(float n) {
  return fabsf(n - GPT_MAX_PHONG_N_S) < GPT_SMALL || n >= GPT_MAX_PHONG_N_S;
}
__device__ __forceinline__ V3 sample_hemisphere_cos(Rng &rng) {
  float r1 = rng_uniform(rng), r2 = rng_uniform(rng);
  float z = sqrtf(1.f - r2);
  float phi = (float)(2 * GPT_PI * (double)r1);
  float x = cosf(phi) * sqrtf(r2);
  float y = sinf(phi) * sqrtf(r2);
  return make_unit(V3{x, y, z});
}
static __device__ __noinline__ V3 sample_phong(float n, Rng &rng) {
  if (ns_too_large(n)) return V3{0.f, 0.f, 1.f};  // no draws
  float r1 = rng_uniform(rng), r2 = rng_uniform(rng);
  float e2 = (float)(2.0 / (double)(n + 1.f)), e1 = (float)(1.0 / (double)(n + 1.f));
  // cos / sin take a double argument in the reference (2 * M_PI * r2)
  double ang = 2 * GPT_PI * (double)r2;
  float s = sqrtf(1.f - powf(r1, e2));
  float x = (float)((double)s * cos(ang));
  float y = (float)((double)s * sin(ang));
  float z = powf(r1, e1);
  return make_unit(V3{x, y, z});
}
__device__ __forceinline__ V3 sample_hemisphere(Rng &rng) {
  float st = rng_uniform(rng);
  float ct = sqrtf(1.f - st * st);
  float phi = (float)((double)(rng_uniform(rng) * 2.f) * GPT_PI);
  return make_unit(V3{st * cosf(phi), st * sinf(phi), ct});
}
// "disk" sampler: a point ON the unit circle (vector_util.h:290-300)
__device__ __forceinline__ V3 sample_disk(Rng &rng) {
  float st = 2.f * rng_uniform(rng) - 1.f;
  float ct = sqrtf(1.f - st * st);
  float rn = rng_uniform(rng);
  if (rn <= .5f) ct = -ct;
  return make_unit(V3{ct, st, 0.f});
}

// generate_ray (ray_operations.h:16-35) + Ray::Ray normalisation; returns the direction
static __device__ __noinline__ V3 generate_dir(V3 main_dir, V3 normal, int mode, float n, Rng &rng) {
  Frame f;
  V3 v;
  if (mode == 0) { f = frame_from_z(normal); v = sample_hemisphere_cos(rng); }
  else if (mode == 1) { f = frame_from_z(main_dir); v = sample_phong(n, rng); }
  else { f = frame_from_z(normal); v = sample_hemisphere(rng); }
  return unit(frame_to_world(f, v));
}

// ---- Phong filters, Schlick, reflect (vector_util.h:207-246) -----------------------------
static __device__ __noinline__ V3 phong_filter(V3 k, float n, V3 ideal, V3 dir) {
  if (ns_too_large(n)) return ((k * GPT_MAX_PHONG_N_S) * V3{1.f, 1.f, 1.f}) / 2.f;
  return ((k * (n + 2.f)) * powf(fmaxf(0.f, dot(ideal, dir)), n)) / 2.f;
}
__device__ __forceinline__ V3 phong_filter_2(V3 k, float n, V3 ideal, V3 dir) {
  if (ns_too_large(n)) return k;
  return k * powf(fmaxf(0.f, dot(ideal, dir)), n);
}
static __device__ __noinline__ float schlick(float cos_theta, float n1, float n2) {
  float r0 = powf((n1 - n2) / (n1 + n2), 2.f);
  return r0 + (1.f - r0) * powf(1.f - cos_theta, 5.f);
}
__device__ __forceinline__ V3 reflect_dir(V3 v, V3 n) { return v - (2.f * dot(v, n)) * n; }

// ---- pdfs (vector_util.h:75-205) ---------------------------------------------------------
__device__ __forceinline__ float scattering_pdf(V3 normal, V3 next, bool diffuse, V3 coming,
                                                bool refracted) {
  if (diffuse) return fmaxf(0.f, dot(normal, next));
  float d1 = dot(coming, normal), d2 = dot(next, normal);
  bool ok = (d1 >= 0 && d2 <= 0 && !refracted) || (d1 <= 0 && d2 >= 0 && !refracted) ||
            (d1 >= 0 && d2 >= 0 && refracted) || (d1 <= 0 && d2 <= 0 && refracted);
  return ok ? 1.f : 0.f;
}
__device__ __forceinline__ float diffuse_sampling_pdf(V3 normal, V3 dir) {
  return fmaxf(0.f, (float)((double)dot(normal, dir) / GPT_PI));
}
// compute_specular_sampling_pdf (the (n+1)-weighted variant used by _recompute_pdf)
static __device__ __noinline__ float specular_sampling_pdf(V3 in, V3 out, V3 normal, V3 perfect,
                                                       float n, bool refracted) {
  float d1 = dot(in, normal), d2 = dot(normal, out);
  bool side = refracted ? ((d1 >= 0 && d2 >= 0) || (d1 <= 0 && d2 <= 0))
                        : ((d1 >= 0 && d2 <= 0) || (d1 <= 0 && d2 >= 0));
  if (!side) return 0.f;
  if (ns_too_large(n)) return (float)((double)GPT_MAX_PHONG_N_S / (2 * GPT_PI));
  return (float)((double)((n + 1.f) * powf(fmaxf(0.f, dot(perfect, out)), n)) / (2 * GPT_PI));
}
