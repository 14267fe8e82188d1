// integrator.cuh -- pieces shared by the path-tracing wavefront (render.cu) and the SPPM
// passes (sppm.cu): camera ray generation, light importance sampling + MIS (change_ref_ray),
// host-side camera derivation.
#pragma once
#include <cmath>

#include "lighttree.cuh"
#include "render_state.cuh"

// Camera::compute_ray (camera.h:68-83) + Ray::Ray normalisation.
__device__ __forceinline__ void camera_ray(const CameraDev &c, float i, float j, Rng &rng, V3 &o,
                                           V3 &d) {
  float w2 = (float)c.width / 2, h2 = (float)c.height / 2;
  float alpha = (float)(c.tan_half_fovx * (double)(j - w2) / (double)w2);
  float beta = (float)(c.tan_half_fovy * (double)(h2 - i) / (double)h2);
  V3 dir = unit(alpha * c.u + beta * c.v - c.w);
  V3 point = c.eye + c.focus_dist * dir;
  V3 rd = c.lens_radius * sample_disk(rng);
  V3 off = rd.x * c.u + rd.y * c.v;
  o = c.eye + off;
  d = unit(point - c.eye - off);
}

// change_ref_ray + _recompute_pdf (material_operations.h:26-153)
__device__ __forceinline__ float change_ref_ray(const LightTree &lt, const float4 *shade,
                                                const MaterialDev *mats, const TexAtlas &atlas,
                                                V3 point, V3 normal, V3 v_in, RefRec &ref,
                                                float weight, Rng &rng) {
  float rn = rng_uniform(rng);
  V3 pivot = ref.diffuse ? normal : ref.perfect;
  if (isinf(ref.n)) weight = 0.f;
  if (rn < weight) {
    int leaf = light_pick(lt, point, pivot, rng);
    V3 target = light_sample_point(lt, leaf, rng);
    ref.dir = unit(target - point);
    if (ref.reflected || ref.refracted) ref.filter = phong_filter(ref.k, ref.n, pivot, ref.dir);
  }
  V3 d1 = unit(ref.dir);
  float hsum = light_hittable_sum(lt, shade, mats, atlas, point, d1, point, pivot);
  float samp = ref.diffuse ? diffuse_sampling_pdf(normal, ref.dir)
                           : specular_sampling_pdf(v_in, ref.dir, normal, ref.perfect, ref.n,
                                                   ref.refracted);
  float pdf = weight * hsum + (1.f - weight) * samp;
  float spdf = scattering_pdf(normal, ref.dir, ref.diffuse, v_in, ref.refracted);
  return (float)((double)spdf / GPT_PI / (double)pdf);
}

// Camera::Camera (camera.h:30-52): same operations, evaluated on the host
static inline CameraDev derive_camera(const GptCamera &c) {
  CameraDev d;
  V3 eye = V3{c.eye[0], c.eye[1], c.eye[2]}, center = V3{c.center[0], c.center[1], c.center[2]};
  V3 up = V3{c.up[0], c.up[1], c.up[2]};
  float fovy = c.fovy;
  float fovx = (float)(2.0f * atan(((float)c.width / (float)c.height) *
                                   tan(M_PI * fovy / 180 / 2)) * 180 / M_PI);
  d.eye = eye;
  d.w = unit(eye - center);
  d.u = unit(cross(up, d.w));
  d.v = cross(d.w, d.u);
  d.tan_half_fovx = tan(fovx * M_PI / 180.0 / 2);
  d.tan_half_fovy = tan(fovy * M_PI / 180.0 / 2);
  d.focus_dist = c.focus_dist;
  d.lens_radius = (float)(c.aperture / 2.0);
  d.width = c.width;
  d.height = c.height;
  return d;
}

