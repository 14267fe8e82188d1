// lighttree.cuh -- shading normal reconstruction, light importance sampling and MIS pdf.
//
// Restates Triangle::_get_normal / get_random_point_on_surface / get_hittable_pdf /
// compute_directed_energy (src/model/geometry/triangle.h:84-110,199-224,410-441),
// Node::compute_importance (src/model/bvh/bvh.h:60-96; only |E| / d^2 survives -- the cone and
// angle terms it computes are discarded by the reference itself),
// traverse_bvh_to_pick_a_target / get_node_pdf / traverse_bvh_target
// (src/model/bvh/bvh_traversal_target.h:48-184) and change_ref_ray / _recompute_pdf
// (src/model/material/material_operations.h:26-153).
#pragma once
#include "bsdf.cuh"
#include "trace_core.cuh"

struct TriShade {  // unpacked per-triangle shading record (common.cuh)
  V3 n1, n2, n3;
  float u1, v1, u2, v2, u3, v3;
  V3 g1, g2, g3;
  float area;
  int mat, flags, obj;
};
__device__ __forceinline__ TriShade load_tri_shade(const float4 *__restrict__ shade, int prim) {
  const float4 *o = shade + 7 * (size_t)prim;
  float4 s0 = __ldg(o), s1 = __ldg(o + 1), s2 = __ldg(o + 2), s3 = __ldg(o + 3),
         s4 = __ldg(o + 4), s5 = __ldg(o + 5), s6 = __ldg(o + 6);
  TriShade t;
  t.n1 = V3{s0.x, s0.y, s0.z}; t.n2 = V3{s0.w, s1.x, s1.y}; t.n3 = V3{s1.z, s1.w, s2.x};
  t.u1 = s2.y; t.v1 = s2.z; t.u2 = s2.w; t.v2 = s3.x; t.u3 = s3.y; t.v3 = s3.z;
  t.area = s3.w;
  t.g1 = V3{s4.x, s4.y, s4.z}; t.g2 = V3{s4.w, s5.x, s5.y}; t.g3 = V3{s5.z, s5.w, s6.x};
  t.mat = __float_as_int(s6.y);
  t.flags = __float_as_int(s6.z);
  t.obj = __float_as_int(s6.w);
  return t;
}

// Triangle::_get_normal (triangle.h:410-441): interpolated, bump-perturbed shading normal.
static __device__ __noinline__ V3 shading_normal(const TriShade &t, const MaterialDev &m,
                                             const TexAtlas &atlas, float w1, float w2,
                                             float w3) {
  V3 nn = unit(w1 * t.n1 + w2 * t.n2 + w3 * t.n3);
  if (m.tex_h[TEX_BUMP] < 2 || m.tex_w[TEX_BUMP] < 2) return unit(nn);  // bump == 0
  float u = w1 * t.u1 + w2 * t.u2 + w3 * t.u3;
  float v = w1 * t.v1 + w2 * t.v2 + w3 * t.v3;
  V3 bump = mat_bump(atlas, m, u, v);
  V3 tg = unit(w1 * t.g1 + w2 * t.g2 + w3 * t.g3);
  Frame f = has_nan(tg) ? frame_from_z(nn) : frame_from_zx(nn, tg);
  if (length(bump) > 0.f) nn = nn + ((bump.x * f.x) + (bump.y * f.y));
  return unit(nn);
}

// per light leaf (sorted order of the light BVH), 4 float4:
//   l0 = (box centre xyz, area) l1 = (centroid xyz, E.x) l2 = (avg normal xyz, E.y) l3 = (E.z,-,-,-)
// per light internal node, 1 float4: (box centre xyz, |E|); energies in a separate float3 array.
struct LightTree {
  const float4 *nodes;        // node64 of the light BVH
  const float4 *tris;         // 48-byte triangles, light-sorted order
  const int *parent_leaf;
  const float4 *leaf_info;
  const float4 *node_info;
  int n;
};

static __device__ __noinline__ float light_importance(const LightTree &lt, int ref, V3 point,
                                                  V3 pivot) {
  if (ref >= 0) {
    float4 ni = __ldg(lt.node_info + ref);
    return ni.w / sqlen(point - V3{ni.x, ni.y, ni.z});
  }
  const float4 *li = lt.leaf_info + 4 * (size_t)(~ref);
  float4 l0 = __ldg(li), l1 = __ldg(li + 1), l2 = __ldg(li + 2), l3 = __ldg(li + 3);
  // compute_directed_energy (triangle.h:84-93)
  V3 dir = unit(point - V3{l1.x, l1.y, l1.z});
  V3 avgn = V3{l2.x, l2.y, l2.z};
  float s = fmaxf(0.f, dot(avgn, dir) * dot(-dir, pivot));
  V3 e = V3{l1.w, l2.w, l3.x} * s;
  return length(e) / sqlen(point - V3{l0.x, l0.y, l0.z});
}

// traverse_bvh_to_pick_a_target: one uniform per level
__device__ __forceinline__ int light_pick(const LightTree &lt, V3 point, V3 pivot, Rng &rng) {
  int node = 0;
  while (node >= 0) {
    const int *q3 = reinterpret_cast<const int *>(lt.nodes + 4 * (size_t)node + 3);
    int l = __ldg(q3), r = __ldg(q3 + 1);
    float li = light_importance(lt, l, point, pivot), ri = light_importance(lt, r, point, pivot);
    float tot = li + ri;
    float f = tot < 1e-10f ? .5f : li / tot;
    node = rng_uniform(rng) < f ? l : r;
  }
  return ~node;
}

// get_node_pdf: walk to the root multiplying the branch probabilities
__device__ __forceinline__ float light_node_pdf(const LightTree &lt, int leaf, V3 point,
                                                V3 pivot) {
  float pdf = 1.0f;
  int ref = ~leaf;
  int pe = __ldg(lt.parent_leaf + leaf);
  while (pe >= 0) {
    int par = pe >> 1;
    const int *q3 = reinterpret_cast<const int *>(lt.nodes + 4 * (size_t)par + 3);
    int sib = (pe & 1) ? __ldg(q3) : __ldg(q3 + 1);
    float i1 = light_importance(lt, ref, point, pivot);
    float i2 = light_importance(lt, sib, point, pivot);
    float p = i1 / (i1 + i2);
    pdf *= (isnan(p) || isinf(p)) ? .5f : p;
    ref = par;
    pe = __ldg(q3 + 2);
  }
  return pdf;
}

// Triangle::get_random_point_on_surface (triangle.h:199-224): consumes 4 uniforms, only the
// point is used by the caller.
static __device__ __noinline__ V3 light_sample_point(const LightTree &lt, int leaf, Rng &rng) {
  const float4 *tp = lt.tris + 3 * (size_t)leaf;
  float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
  float r1 = rng_uniform(rng), r2 = rng_uniform(rng);
  float u = 1.f - powf(r1, .5f);
  float v = r2 * powf(r1, .5f);
  V3 p = (u * V3{t0.x, t0.y, t0.z} + v * V3{t0.w, t1.x, t1.y}) +
         (1.f - u - v) * V3{t1.z, t1.w, t2.x};
  (void)rng_uniform(rng);  // get_random_unit_vector_hemisphere: direction is discarded
  (void)rng_uniform(rng);
  return p;
}

// Triangle::get_hittable_pdf (triangle.h:95-110) for light leaf `leaf`
static __device__ __noinline__ float light_hittable_pdf(const LightTree &lt, const float4 *shade,
                                                    const MaterialDev *mats,
                                                    const TexAtlas &atlas, int leaf, V3 origin,
                                                    V3 dir) {
  dir = unit(dir);
  V3 rd = unit(dir);  // Ray::Ray normalises again
  const float4 *tp = lt.tris + 3 * (size_t)leaf;
  float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
  RaySetup r = make_ray(origin.x, origin.y, origin.z, rd.x, rd.y, rd.z);
  float t, b1, b2, b3;
  if (!tri_test(r, t0, t1, t2, t, b1, b2, b3)) return 0.f;
  float tol = t2.y;
  float inv_tol = 1.0f / tol;
  if (t > inv_tol || t < tol) return 0.f;
  int prim = __float_as_int(t2.z);
  TriShade ts = load_tri_shade(shade, prim);
  V3 n = shading_normal(ts, mats[ts.mat], atlas, b1, b2, b3);
  float cosv = fabsf(dot(dir, n));
  return (t * t) / (cosv * ts.area);
}

// traverse_bvh_target + the summation loop of _recompute_pdf, fused: leaves are consumed in
// the reference's visit order (no culling), capped at 399 candidates like the reference.
__device__ __forceinline__ float light_hittable_sum(const LightTree &lt, const float4 *shade,
                                                    const MaterialDev *mats,
                                                    const TexAtlas &atlas, V3 origin, V3 dir,
                                                    V3 hit_point, V3 pivot) {
  V3 rd = unit(dir);  // `Ray ray = Ray(origin, dir)` on the already unit dir
  RaySetup r = make_ray(origin.x, origin.y, origin.z, rd.x, rd.y, rd.z);
  int stack[64];
  int sp = 0, node = 0, count = 0;
  float sum = 0.f;
  while (true) {
    const float4 *q = lt.nodes + 4 * (size_t)node;
    float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
    int cref[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
    float bxn[2] = {q0.x, q1.z}, bxx[2] = {q0.y, q1.w};
    float byn[2] = {q0.z, q2.x}, byx[2] = {q0.w, q2.y};
    float bzn[2] = {q1.x, q2.z}, bzx[2] = {q1.y, q2.w};
    bool hit[2];
#pragma unroll
    for (int c = 0; c < 2; c++) {
      float a1 = (bxn[c] - r.ox) * r.ix, a2 = (bxx[c] - r.ox) * r.ix;
      float txn = a1 < a2 ? a1 : a2, txx = a1 < a2 ? a2 : a1;
      float c1 = (byn[c] - r.oy) * r.iy, c2 = (byx[c] - r.oy) * r.iy;
      float tyn = c1 < c2 ? c1 : c2, tyx = c1 < c2 ? c2 : c1;
      float d1 = (bzn[c] - r.oz) * r.iz, d2 = (bzx[c] - r.oz) * r.iz;
      float tzn = d1 < d2 ? d1 : d2, tzx = d1 < d2 ? d2 : d1;
      bool ok = (txn <= tyx) && (tyn <= txx);
      float txyn = fmaxf(txn, tyn), txyx = fminf(txx, tyx);
      hit[c] = ok && (txyn <= tzx) && (tzn <= txyx);
    }
#pragma unroll
    for (int c = 0; c < 2; c++) {
      if (hit[c] && cref[c] < 0) {
        int leaf = ~cref[c];
        float npdf = light_node_pdf(lt, leaf, origin, pivot);
        sum += npdf * light_hittable_pdf(lt, shade, mats, atlas, leaf, hit_point, dir);
        count++;
      }
    }
    bool tl = hit[0] && cref[0] >= 0, tr = hit[1] && cref[1] >= 0;
    if (!tl && !tr) {
      if (sp == 0) break;
      node = stack[--sp];
    } else {
      node = tl ? cref[0] : cref[1];
      if (tl && tr && sp < 64) stack[sp++] = cref[1];
    }
    if (count >= 399) break;
  }
  return sum;
}
