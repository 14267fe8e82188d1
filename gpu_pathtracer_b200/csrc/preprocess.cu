// preprocess.cu -- per-triangle derived data, the B200 replacement of create_world's
// `new Triangle(...)` (src/lib/world.h:194-248) and of the tangent kernels
// (src/model/geometry/triangle_operations.h:18-61).
//
// One thread per triangle, SoA in / SoA out, no device heap.  The arithmetic follows
// Triangle::Triangle (src/model/geometry/triangle.h:254-321) operation for operation so the
// tolerance, padded AABB, normal, area, flags and energy are bit-identical to the reference.
#include "common.cuh"
#include "material.cuh"
#include "vecmath.cuh"

namespace {

__device__ __forceinline__ V3 ldv(const float *x, const float *y, const float *z, int i) {
  return V3{__ldg(x + i), __ldg(y + i), __ldg(z + i)};
}

// Triangle::_compute_tangent (triangle.h:174-185): UV-space tangent of one triangle.
__device__ __forceinline__ void tri_tangent(V3 p1, V3 p2, V3 p3, float u1, float v1, float u2,
                                            float v2, float u3, float v3_, V3 &t, V3 &b) {
  V3 e1 = p2 - p1, e2 = p3 - p1;
  float x1 = u2 - u1, x2 = u3 - u1, y1 = v2 - v1, y2 = v3_ - v1;
  float r = 1.0f / (x1 * y2 - x2 * y1);
  t = (e1 * y2 - e2 * y1) * r;
  b = (e2 * x1 - e1 * x2) * r;
}

__global__ void __launch_bounds__(256) k_preprocess(SceneDev s, TexAtlas atlas) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.ntri) return;
  int a = __ldg(s.p1 + i), b = __ldg(s.p2 + i), c = __ldg(s.p3 + i);
  V3 p1 = ldv(s.x, s.y, s.z, a), p2 = ldv(s.x, s.y, s.z, b), p3 = ldv(s.x, s.y, s.z, c);
  const MaterialDev &m = s.mats[__ldg(s.mat_idx + i)];

  // _compute_tolerance (triangle.h:307-321)
  float d12 = dist(p1, p2), d13 = dist(p1, p3), d23 = dist(p2, p3);
  float tl = d12 < d13 ? d12 : d13;
  if (d23 < tl) tl = d23;
  float tol = fminf(GPT_SMALL, tl / 100.f);
  // _compute_triangle_area (triangle.h:443-452): Heron
  float sp = (d12 + d13 + d23) / 2.0f;
  float area = fabsf(sqrtf(sp * (sp - d12) * (sp - d13) * (sp - d23)));
  V3 nrm = unit(cross(p2 - p1, p3 - p1));

  V3 n1r = ldv(s.xn, s.yn, s.zn, __ldg(s.n1 + i));
  V3 n2r = ldv(s.xn, s.yn, s.zn, __ldg(s.n2 + i));
  V3 n3r = ldv(s.xn, s.yn, s.zn, __ldg(s.n3 + i));
  V3 n1 = unit(n1r), n2 = unit(n2r), n3 = unit(n3r);
  int ti1 = __ldg(s.t1 + i), ti2 = __ldg(s.t2 + i), ti3 = __ldg(s.t3 + i);
  float u1 = __ldg(s.xt + ti1), v1 = __ldg(s.yt + ti1);
  float u2 = __ldg(s.xt + ti2), v2 = __ldg(s.yt + ti2);
  float u3 = __ldg(s.xt + ti3), v3_ = __ldg(s.yt + ti3);

  // _check_if_light_emitting (triangle.h:123-135) / _compute_energy (:112-121)
  V3 em1 = mat_emission(atlas, m, u1, v1), em2 = mat_emission(atlas, m, u2, v2),
     em3 = mat_emission(atlas, m, u3, v3_);
  bool light = length(em1) > tol || length(em2) > tol || length(em3) > tol;
  bool transparent = m.t_r > 0.f;
  // vertex-normal fallback (triangle.h:292-300)
  if (length(n1r) < tol || length(n2r) < tol || length(n3r) < tol) n1 = n2 = n3 = nrm;

  // _compute_bounding_box (triangle.h:226-252)
  float *bb = s.aabb + 6 * (size_t)i;
  bb[0] = fminf(fminf(p1.x, p2.x), p3.x) - tol;
  bb[1] = fmaxf(fmaxf(p1.x, p2.x), p3.x) + tol;
  bb[2] = fminf(fminf(p1.y, p2.y), p3.y) - tol;
  bb[3] = fmaxf(fmaxf(p1.y, p2.y), p3.y) + tol;
  bb[4] = fminf(fminf(p1.z, p2.z), p3.z) - tol;
  bb[5] = fmaxf(fmaxf(p1.z, p2.z), p3.z) + tol;
  s.tolerance[i] = tol;
  s.normal[3 * (size_t)i] = nrm.x;
  s.normal[3 * (size_t)i + 1] = nrm.y;
  s.normal[3 * (size_t)i + 2] = nrm.z;

  V3 avg = (em1 + em2 + em3) / 3.0f;
  V3 en = (6.283185307179586f * avg) * area;  // (float)(2 * M_PI) * avg * area
  s.energy[3 * (size_t)i] = en.x;
  s.energy[3 * (size_t)i + 1] = en.y;
  s.energy[3 * (size_t)i + 2] = en.z;

  int flags = (light ? GPT_TRI_LIGHT : 0) | (transparent ? GPT_TRI_TRANSPARENT : 0) |
              (m.sss ? GPT_TRI_SSS : 0);
  float4 *o = s.shade + 7 * (size_t)i;
  o[0] = make_float4(n1.x, n1.y, n1.z, n2.x);
  o[1] = make_float4(n2.y, n2.z, n3.x, n3.y);
  o[2] = make_float4(n3.z, u1, v1, u2);
  o[3] = make_float4(v2, u3, v3_, area);
  // s4, s5 (corner tangents) are written by k_assign_tangent
  o[6] = make_float4(0.f, __int_as_float(__ldg(s.mat_idx + i)), __int_as_float(flags),
                     __int_as_float(__ldg(s.obj_idx + i)));
}

// (vertex, 3*tri+corner) pairs for the per-vertex tangent sums
__global__ void __launch_bounds__(256) k_tangent_pairs(SceneDev s, uint32_t *keys, int *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.ntri) return;
  keys[3 * (size_t)i] = (uint32_t)__ldg(s.p1 + i);
  keys[3 * (size_t)i + 1] = (uint32_t)__ldg(s.p2 + i);
  keys[3 * (size_t)i + 2] = (uint32_t)__ldg(s.p3 + i);
  vals[3 * (size_t)i] = 3 * i;
  vals[3 * (size_t)i + 1] = 3 * i + 1;
  vals[3 * (size_t)i + 2] = 3 * i + 2;
}

// sum_up_tangent_and_bitangent (triangle_operations.h:34-61) adds each triangle's tangent to
// its three vertices serially in triangle order.  fp32 addition is order-dependent, so each
// vertex sums ITS incident (triangle, corner) entries in the same increasing order: the
// pairs were stably sorted by vertex, one thread walks one vertex's segment.
__global__ void __launch_bounds__(256)
k_vertex_tangent(SceneDev s, const uint32_t *keys, const int *vals, int npairs) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= s.nv) return;
  int lo = 0, hi = npairs;  // first pair with key >= v
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (keys[mid] < (uint32_t)v) lo = mid + 1; else hi = mid;
  }
  V3 acc = V3{0.f, 0.f, 0.f};
  for (int k = lo; k < npairs && keys[k] == (uint32_t)v; k++) {
    int tri = vals[k] / 3;
    int a = __ldg(s.p1 + tri), b = __ldg(s.p2 + tri), c = __ldg(s.p3 + tri);
    const float4 *sh = s.shade + 7 * (size_t)tri;
    float4 s2 = sh[2], s3 = sh[3];
    V3 t, bt;
    tri_tangent(ldv(s.x, s.y, s.z, a), ldv(s.x, s.y, s.z, b), ldv(s.x, s.y, s.z, c), s2.y,
                s2.z, s2.w, s3.x, s3.y, s3.z, t, bt);
    if (!has_nan(t) && !has_nan(bt)) acc = acc + t;
  }
  s.vtangent[3 * (size_t)v] = acc.x;
  s.vtangent[3 * (size_t)v + 1] = acc.y;
  s.vtangent[3 * (size_t)v + 2] = acc.z;
}

// assign_tangent kernel + Triangle::assign_tangent (triangle_operations.h:18-32,
// triangle.h:160-172): Gram-Schmidt of the vertex tangent against each corner normal.
__global__ void __launch_bounds__(256) k_assign_tangent(SceneDev s) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.ntri) return;
  float4 *o = s.shade + 7 * (size_t)i;
  float4 s0 = o[0], s1 = o[1], s2 = o[2];
  V3 n[3] = {V3{s0.x, s0.y, s0.z}, V3{s0.w, s1.x, s1.y}, V3{s1.z, s1.w, s2.x}};
  int vi[3] = {__ldg(s.p1 + i), __ldg(s.p2 + i), __ldg(s.p3 + i)};
  V3 tg[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    V3 t = unit(ldv(s.vtangent, s.vtangent + 1, s.vtangent + 2, 3 * vi[k]));
    tg[k] = unit(t - dot(n[k], t) * n[k]);
  }
  o[4] = make_float4(tg[0].x, tg[0].y, tg[0].z, tg[1].x);
  o[5] = make_float4(tg[1].y, tg[1].z, tg[2].x, tg[2].y);
  float4 s6 = o[6];
  s6.x = tg[2].z;
  o[6] = s6;
}

}  // namespace

int gpt_preprocess(gpt_ctx *ctx) {
  SceneDev &s = ctx->scene;
  int n = s.ntri;
  GPT_TRY(gpt_alloc(ctx, &s.tolerance, (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &s.aabb, 6 * (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &s.normal, 3 * (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &s.energy, 3 * (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &s.shade, 7 * (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &s.vtangent, 3 * (size_t)s.nv));
  TexAtlas atlas{s.tex_r, s.tex_g, s.tex_b};
  int grid = (n + 255) / 256;
  k_preprocess<<<grid, 256, 0, ctx->stream>>>(s, atlas);
  GPT_LAUNCH_CHECK(ctx);

  // per-vertex tangent sums in the reference's order
  size_t np = 3 * (size_t)n;
  uint32_t *keys = nullptr, *keys_alt = nullptr, *hist = nullptr, *keys_out = nullptr;
  int *vals = nullptr, *vals_alt = nullptr, *vals_out = nullptr;
  GPT_TRY(gpt_alloc(ctx, &keys, np));
  GPT_TRY(gpt_alloc(ctx, &keys_alt, np));
  GPT_TRY(gpt_alloc(ctx, &vals, np));
  GPT_TRY(gpt_alloc(ctx, &vals_alt, np));
  GPT_TRY(gpt_alloc(ctx, &hist, (size_t)gpt_radix_hist_entries((int)np)));
  k_tangent_pairs<<<grid, 256, 0, ctx->stream>>>(s, keys, vals);
  GPT_LAUNCH_CHECK(ctx);
  int nbits = 1;
  while (nbits < 32 && (1ll << nbits) < (long long)s.nv) nbits++;
  GPT_TRY(gpt_radix_sort_pairs(ctx, keys, vals, keys_alt, vals_alt, hist, (int)np, nbits,
                               &keys_out, &vals_out));
  k_vertex_tangent<<<(s.nv + 255) / 256, 256, 0, ctx->stream>>>(s, keys_out, vals_out, (int)np);
  GPT_LAUNCH_CHECK(ctx);
  k_assign_tangent<<<grid, 256, 0, ctx->stream>>>(s);
  GPT_LAUNCH_CHECK(ctx);
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  gpt_free(ctx, keys); gpt_free(ctx, keys_alt); gpt_free(ctx, vals); gpt_free(ctx, vals_alt); gpt_free(ctx, hist);
  return GPT_OK;
}
