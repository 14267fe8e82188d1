// material.cuh -- texture fetch + material accessors (device).
// Behaviour follows Material::_get_texture and its wrappers
// (src/model/material/material.h:653-727): nearest texel, planar R/G/B atlas,
// iu = floor(frac(u) * (W-1)), iv = floor(frac(v) * (H-1)), idx = W*iv + iu.
#pragma once
#include "common.cuh"
#include "vecmath.cuh"

enum { TEX_DIFFUSE = 0, TEX_SPECULAR = 1, TEX_EMISSION = 2, TEX_NS = 3, TEX_BUMP = 4 };

struct TexAtlas {
  const float *r, *g, *b;
};

static __device__ __noinline__ V3 tex_fetch(const TexAtlas &a, const MaterialDev &m, int which,
                                        float u, float v, V3 filter) {
  int w = m.tex_w[which], h = m.tex_h[which];
  int iu = (int)floorf((u - floorf(u)) * (float)(w - 1));
  int iv = (int)floorf((v - floorf(v)) * (float)(h - 1));
  int idx = m.tex_off[which] + w * iv + iu;
  return V3{__ldg(a.r + idx), __ldg(a.g + idx), __ldg(a.b + idx)} * filter;
}

__device__ __forceinline__ V3 mat_emission(const TexAtlas &a, const MaterialDev &m, float u,
                                           float v) {
  return tex_fetch(a, m, TEX_EMISSION, u, v, V3{m.ke[0], m.ke[1], m.ke[2]});
}
__device__ __forceinline__ V3 mat_diffuse(const TexAtlas &a, const MaterialDev &m, float u,
                                          float v) {
  return tex_fetch(a, m, TEX_DIFFUSE, u, v, V3{m.kd[0], m.kd[1], m.kd[2]});
}
__device__ __forceinline__ V3 mat_specular(const TexAtlas &a, const MaterialDev &m, float u,
                                           float v) {
  return tex_fetch(a, m, TEX_SPECULAR, u, v, V3{m.ks[0], m.ks[1], m.ks[2]});
}
// get_texture_bump (material.h:690-703): zero unless the map is at least 2x2
__device__ __forceinline__ V3 mat_bump(const TexAtlas &a, const MaterialDev &m, float u,
                                       float v) {
  if (m.tex_h[TEX_BUMP] < 2 || m.tex_w[TEX_BUMP] < 2) return V3{0.f, 0.f, 0.f};
  V3 t = tex_fetch(a, m, TEX_BUMP, u, v, V3{1.f, 1.f, 1.f}) - V3{0.5f, 0.5f, 0.5f};
  return (2.f * m.bm) * t;  // `2 * bm * (...)`: (2*bm) first, then scale the vector
}
// _get_texture_n_s (material.h:713-727)
__device__ __forceinline__ float mat_ns(const TexAtlas &a, const MaterialDev &m, float u,
                                        float v) {
  if (isinf(m.n_s)) return m.n_s;
  float f = m.n_s / powf(3.f, .5f);
  return length(tex_fetch(a, m, TEX_NS, u, v, V3{f, f, f}));
}
