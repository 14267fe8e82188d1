// bsdf.cuh -- the material decision at a hit: nested-dielectric priority list, false hits,
// Schlick refract / reflect, diffuse vs Phong-lobe choice.
//
// Restates Material::check_next_path, _check_if_false_hit, find_highest_prioritised_materials,
// _get_false_hit_parameters, _refract (src/model/material/material.h:130-197,230-409,509-651)
// and add_new_material / remove_a_material (src/render/material_list_operations.h:9-41).
// The RNG draws happen in the same order as in the reference.
#pragma once
#include "material.cuh"
#include "shading.cuh"

#define MATLIST_MAX 12     /* the reference reserves 400 pointers; nesting this deep is a scene bug */
#define MAT_AIR 255        /* nullptr in the reference: priority 9999999, n_i = 1 */

struct MatList {
  unsigned char id[MATLIST_MAX];
  int len;
};

__device__ __forceinline__ int mat_priority(const MaterialDev *mats, int id) {
  return id == MAT_AIR ? 9999999 : mats[id].priority;
}
__device__ __forceinline__ float mat_ior(const MaterialDev *mats, int id) {
  return id == MAT_AIR ? 1.0f : mats[id].n_i;
}
__device__ __forceinline__ bool matlist_add(MatList &l, int id) {  // add_new_material
  for (int i = l.len - 1; i >= 0; i--)
    if (l.id[i] == id) return true;
  if (l.len >= MATLIST_MAX) return false;
  l.id[l.len++] = (unsigned char)id;
  return true;
}
__device__ __forceinline__ void matlist_remove(MatList &l, int id) {  // remove_a_material
  int sel = -1;
  for (int i = l.len - 1; i >= 0 && sel < 0; i--)
    if (l.id[i] == id) sel = i;
  if (sel < 0) return;
  for (int i = sel; i < l.len - 1; i++) l.id[i] = l.id[i + 1];
  l.len--;
}
// find_highest_prioritised_materials: scans from the END of the list, strict comparisons
__device__ __forceinline__ void matlist_top2(const MaterialDev *mats, const MatList &l, int &hi,
                                             int &second) {
  hi = MAT_AIR;
  second = MAT_AIR;
  for (int i = l.len - 1; i >= 0; i--)
    if (mat_priority(mats, l.id[i]) < mat_priority(mats, hi)) hi = l.id[i];
  for (int i = l.len - 1; i >= 0; i--) {
    int p = mat_priority(mats, l.id[i]);
    if (p < mat_priority(mats, second) && p > mat_priority(mats, hi)) second = l.id[i];
  }
}

struct RefRec {  // reflection_record (material.h:116-123)
  V3 dir;        // ref.ray.dir (origin is the hit point)
  V3 k, filter, perfect;
  V3 filter2;    // ref.filter_2 (compute_phong_filter_2): only filled by check_next_path<true>
  float n;
  bool diffuse, reflected, refracted, false_hit, entering, bent;
  int next_material;
};

// Returns with `ref` filled; `sss` as the reference sets it.  F2 = also fill ref.filter2 (the
// SPPM passes use it, the path tracer does not).
template <bool F2>
static __device__ __noinline__ void check_next_path(const MaterialDev *mats, const TexAtlas &atlas,
                                                int mat_id, V3 v_in, V3 normal, float u, float v,
                                                const MatList &list, RefRec &ref, bool &sss,
                                                Rng &rng, bool force_refract) {
  const MaterialDev &m = mats[mat_id];
  int hi, second;
  matlist_top2(mats, list, hi, second);
  ref.next_material = MAT_AIR;
  ref.false_hit = m.priority > mat_priority(mats, hi);
  if (ref.false_hit) {  // _get_false_hit_parameters
    sss = false;
    ref.reflected = false;
    ref.refracted = true;
    ref.dir = unit(v_in);
    ref.filter = V3{1.f, 1.f, 1.f};
    if (F2) ref.filter2 = V3{1.f, 1.f, 1.f};
    ref.diffuse = false;
    ref.bent = false;
    ref.entering = dot(v_in, normal) <= 0.f;
    ref.next_material = hi;
    return;
  }
  if (m.t_r > 0.f) {  // _refract
    float rn = rng_uniform(rng);
    float n_hi = mat_ior(mats, hi), n_2nd = mat_ior(mats, second);
    V3 k = V3{m.tf[0], m.tf[1], m.tf[2]} * m.t_r;
    float ns = mat_ns(atlas, m, u, v);
    ref.n = ns;
    sss = false;
    V3 v_out;
    if (dot(v_in, normal) <= 0.f) {
      float c1 = dot(v_in, -normal);
      float rp = schlick(c1, n_hi, m.n_i);
      if (rn >= rp || force_refract) {
        float s1 = powf(1.f - powf(c1, 2.f), .5f);
        V3 perp = (-c1) * normal;
        V3 par = v_in - perp;
        float s2 = n_hi / m.n_i * s1;
        float c2 = powf(1.f - powf(s2, 2.f), .5f);
        float t2 = s2 / c2;
        if (fabsf(t2) > GPT_SMALL) {
          V3 op = ((-1.f / t2) * length(par)) * normal;
          v_out = make_unit(par + op);
        } else {
          v_out = -normal;
        }
        ref.diffuse = false; ref.reflected = false; ref.refracted = true;
        ref.entering = true;
        ref.next_material = mat_id;
        ref.bent = fabsf(n_hi - m.n_i) > GPT_SMALL;
      } else {
        v_out = make_unit(reflect_dir(v_in, normal));
        ref.diffuse = false; ref.reflected = true; ref.refracted = false;
        ref.entering = false;
        ref.next_material = hi;
      }
    } else {
      float s1max = n_2nd / m.n_i;
      float c1 = dot(v_in, normal);
      float s1 = powf(1.f - powf(c1, 2.f), .5f);
      float rp = schlick(c1, m.n_i, n_2nd);
      if ((s1 >= s1max || rn <= rp) && !force_refract) {
        v_out = make_unit(reflect_dir(v_in, -normal));
        ref.diffuse = false; ref.reflected = true; ref.refracted = false;
        ref.entering = false;
        ref.next_material = mat_id;
      } else {
        V3 perp = c1 * normal;
        V3 par = v_in - perp;
        float s2 = m.n_i / n_2nd * s1;
        float c2 = powf(1.f - powf(s2, 2.f), .5f);
        float t2 = s2 / c2;
        if (fabsf(t2) > GPT_SMALL) {
          V3 op = ((1.f / t2) * length(par)) * normal;
          v_out = make_unit(par + op);
        } else {
          v_out = normal;
        }
        ref.diffuse = false; ref.reflected = false; ref.refracted = true;
        ref.entering = false;
        ref.next_material = second;
        ref.bent = fabsf(n_2nd - m.n_i) > GPT_SMALL;
      }
    }
    ref.perfect = v_out;
    ref.dir = generate_dir(v_out, normal, 1, ns, rng);
    ref.k = k;
    ref.filter = phong_filter(k, ns, v_out, ref.dir);
    if (F2) ref.filter2 = phong_filter_2(k, ns, v_out, ref.dir);
    if (ref.next_material != MAT_AIR && ref.next_material != mat_id) {
      const MaterialDev &nm = mats[ref.next_material];
      V3 k2 = V3{nm.tf[0], nm.tf[1], nm.tf[2]} * nm.t_r;
      float ns2 = mat_ns(atlas, nm, u, v);
      ref.k = ref.k * k2;
      ref.filter = ref.filter * phong_filter(k2, ns2, v_out, ref.dir);
      if (F2) ref.filter2 = ref.filter2 * phong_filter_2(k2, ns2, v_out, ref.dir);
    }
    return;
  }
  // opaque: diffuse vs Phong-lobe reflection (material.h:586-650)
  float rn = rng_uniform(rng);
  float kd_len = length(mat_diffuse(atlas, m, u, v));
  float ks_len = length(mat_specular(atlas, m, u, v));
  float factor = ks_len / (kd_len + ks_len);
  float ns = mat_ns(atlas, m, u, v);
  if (rn > factor) {
    ref.dir = generate_dir(V3{0.f, 0.f, 0.f}, normal, 0, 1.f, rng);
    ref.filter = mat_diffuse(atlas, m, u, v);
    if (F2) ref.filter2 = ref.filter;
    ref.diffuse = true; ref.reflected = false; ref.refracted = false;
    ref.k = ref.filter;
    ref.next_material = hi;
    ref.n = 1.f;
    sss = m.sss != 0;
    return;
  }
  V3 rdir = reflect_dir(v_in, normal);
  ref.dir = generate_dir(rdir, normal, 1, ns, rng);
  V3 k = mat_specular(atlas, m, u, v);
  ref.filter = phong_filter(k, ns, rdir, ref.dir);
  if (F2) ref.filter2 = phong_filter_2(k, ns, rdir, ref.dir);
  ref.diffuse = false; ref.reflected = true; ref.refracted = false;
  ref.perfect = rdir;
  ref.n = ns;
  ref.k = k;
  ref.next_material = hi;
  sss = false;
}
