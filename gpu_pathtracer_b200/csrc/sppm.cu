// sppm.cu -- stochastic progressive photon mapping on sm_100a.
//
// Replaces the render_mode == 2 branch of main() (src/main.cu:1282-1870):
//   ray_tracing_pass(init) + compute_average_radius + assign_radius_to_invalid_hit_points
//                                                         -> gpt_sppm_init
//   photon_pass + gather_recorded_photons                 -> k_photon_trace + stable 2-bit sort
//   photon LBVHs (bvh_building_photon.h, twice per pass)  -> gpt_lbvh_build_boxes (lbvh.cu) +
//                                                            k_pack_photons / k_sphere_refit
//   ray_tracing_pass                                      -> k_eye_pass (volume beam gather inside)
//   update_hit_point_parameters (traversal half)          -> k_gather      } gpt_sppm_pass
//   update_hit_point_parameters (update half)             -> k_update        gpt_sppm_update
//   get_ppm_image_output                                  -> k_resolve       gpt_sppm_resolve
//
// The per-pass state a multi-GPU run has to combine -- (flux.rgb, count) and the volume term
// per hit point -- lives in ONE buffer (gpt_sppm_flux_device) so the host can allreduce it
// between gpt_sppm_pass and gpt_sppm_update.
//
// Reference quirks, each REPLICATED unless stated:
//   * the eye pass never writes its RNG state back: pass p uses state(1984, pixel) advanced by
//     p draws (ray_tracing_pass.h:505-509)                                        replicated
//   * photon_pass draws its BSDF decisions from rand_state[0] -- ONE state raced on by every
//     photon thread (photon_pass.h:156-160).  DEVIATES: not reproducible on any parallel
//     machine.  Here each photon owns the state the reference's code would see if the kernel
//     were launched per photon with rand_state = &state[photon]: the BSDF stream G is
//     state[photon], advanced in place across passes; the local stream L is a copy of it
//     advanced by `pass` draws.  oracle/ref_replay.cpp runs the reference's own photon_pass
//     in exactly that way, so photon paths are compared one to one.
//   * max_bounce is overwritten to 64 in the eye pass (ray_tracing_pass.h:210)     replicated
//   * direct light clamp to 1.0 per channel (ray_tracing_pass.h:135-139)           replicated
//   * all radii = global mean after init (src/main.cu:1353), summed serially in pixel order
//     (fp32, ray_tracing_pass.h:421-431)                                           replicated
//   * the volume accumulation is prefix weighted (bvh_traversal_photon.h:62-78)    replicated
//   * the scattered photon's new origin uses the NEW free-flight distance
//     (photon_pass.h:194-196), prev_location is the previous ray's surface point   replicated
//   * a medium with scattering_prob == 0 and no enclosing surface loops forever in the
//     reference; DEVIATES: the walk is capped at 4096 scattering events.
#include <vector>

#include "integrator.cuh"
#include "trace_wide.cuh"

namespace {

constexpr int SP_BLOCK = 128;
constexpr int SP_STACK = TRACE_STACK;  // levels of the shared-memory traversal stack

// ---- scene view handed to every SPPM kernel ---------------------------------------------------
struct TreeRef {
  const float4 *nodes, *tris;
  const int *parent_leaf;
  int n;
};
struct SppmScene {
  TreeRef main_t, transp_t;
  SceneDev sc;
  TexAtlas atlas;
  LightTree lt;
  unsigned long long *counters;  // [0] V_int [1] V_leaf (profiling) [2] main-BVH rays, [3] stack overflow
  int *error_flag;               // nested-dielectric list overflow
  int count_visits;              // profiling: count V_int / V_leaf of the main-BVH traversals
};

// hit_record (primitive.h:111-119): persistent across traversals -- a miss only resets t.
struct SurfRec {
  float t;
  V3 point, normal;
  float u, v;
  int mat;     // material index of rec.object
  V3 in_dir;   // rec.coming_ray.dir
};
__device__ __forceinline__ void surf_init(SurfRec &r) {
  r.t = INFINITY;
  r.point = V3{0.f, 0.f, 0.f}; r.normal = V3{0.f, 0.f, 0.f};
  r.u = r.v = 0.f;
  r.mat = 0;
  r.in_dir = V3{1.f, 0.f, 0.f};  // Ray::Ray() default (ray.h:16-19)
}

struct Tracer {
  int *stack;       // column of the block's shared stack, stride SP_BLOCK
  unsigned nrays;   // closest-hit traversals of the MAIN BVH by this thread
};

// hit_record fill of Triangle::hit (triangle.h:399-405) from a closest-hit result; a miss only
// resets t (the record persists across traversals).
__device__ __forceinline__ bool fill_surf(const SppmScene &S, const TreeRef &tr, float t, int leaf,
                                          int prim, float b1, float b2, float b3, V3 dir,
                                          SurfRec &rec) {
  rec.t = INFINITY;
  if (leaf < 0) return false;
  const float4 *tp = tr.tris + 3 * (size_t)leaf;
  float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
  TriShade ts = load_tri_shade(S.sc.shade, prim);
  const MaterialDev &m = S.sc.mats[ts.mat];
  rec.t = t;
  rec.point = b1 * V3{t0.x, t0.y, t0.z} + b2 * V3{t0.w, t1.x, t1.y} + b3 * V3{t1.z, t1.w, t2.x};
  rec.u = b1 * ts.u1 + b2 * ts.u2 + b3 * ts.u3;
  rec.v = b1 * ts.v1 + b2 * ts.v2 + b3 * ts.v3;
  rec.normal = shading_normal(ts, m, S.atlas, b1, b2, b3);
  rec.mat = ts.mat;
  rec.in_dir = dir;
  return true;
}

// traverse_bvh (bvh_traversal.h:12-84) + hit_record fill of Triangle::hit (triangle.h:399-405).
// `dir` must already be the Ray's unit direction.
static __device__ __noinline__ bool trace_surf(const SppmScene &S, const TreeRef &tr, bool is_main,
                                               Tracer &T, V3 o, V3 dir, SurfRec &rec) {
  RaySetup r = make_ray(o.x, o.y, o.z, dir.x, dir.y, dir.z);
  int overflow = 0;
  HitResult h;
  if (S.count_visits && is_main)
    h = trace_ray<true>(tr.nodes, tr.tris, tr.parent_leaf, r, INFINITY, T.stack, SP_BLOCK,
                        S.counters, S.counters + 1, &overflow);
  else
    h = trace_ray<false>(tr.nodes, tr.tris, tr.parent_leaf, r, INFINITY, T.stack, SP_BLOCK,
                         nullptr, nullptr, &overflow);
  if (overflow) atomicExch(S.counters + 3, 1ull);
  if (is_main) T.nrays++;
  return fill_surf(S, tr, h.t, h.leaf, h.prim, h.b1, h.b2, h.b3, dir, rec);
}

// visibility rays only ask "how far is the first hit": the traversal without the hit-record fill
static __device__ __noinline__ bool trace_t(const SppmScene &S, const TreeRef &tr, Tracer &T, V3 o,
                                            V3 dir, float &t) {
  RaySetup r = make_ray(o.x, o.y, o.z, dir.x, dir.y, dir.z);
  int overflow = 0;
  HitResult h;
  if (S.count_visits)
    h = trace_ray<true>(tr.nodes, tr.tris, tr.parent_leaf, r, INFINITY, T.stack, SP_BLOCK,
                        S.counters, S.counters + 1, &overflow);
  else
    h = trace_ray<false>(tr.nodes, tr.tris, tr.parent_leaf, r, INFINITY, T.stack, SP_BLOCK,
                         nullptr, nullptr, &overflow);
  if (overflow) atomicExch(S.counters + 3, 1ull);
  T.nrays++;
  t = h.leaf < 0 ? INFINITY : h.t;
  return h.leaf >= 0;
}

// rearrange_material_list (material_list_operations.h:43-59)
__device__ __forceinline__ void rearrange_list(const SppmScene &S, MatList &l, int mat,
                                               const RefRec &ref) {
  if (ref.false_hit || ref.refracted) {
    if (ref.entering) {
      if (!matlist_add(l, mat)) *S.error_flag = 1;
    } else {
      matlist_remove(l, mat);
    }
  }
}

// check_if_entering_medium (render/ppm/common.h:8-24)
__device__ __forceinline__ bool entering_medium(const SppmScene &S, const RefRec &ref,
                                                bool in_medium, int &medium) {
  bool e = (!ref.false_hit && ref.next_material != MAT_AIR &&
            S.sc.mats[ref.next_material].extinction_coef > GPT_SMALL) ||
           (ref.false_hit && in_medium);
  if (e) medium = ref.next_material;
  return e;
}

// Material::get_transmittance (material.h:222-227); nullptr medium (AIR) has no extinction
__device__ __forceinline__ float transmittance(const SppmScene &S, int medium, float t) {
  if (medium == MAT_AIR) return 1.0f;
  float e = S.sc.mats[medium].extinction_coef;
  return e > GPT_SMALL ? expf(-t * e) : 1.0f;
}

// init_material_list (material_list_operations.h:61-92): walk back from 999 units along
// init_dir through the transparent-only BVH, forcing refraction, to find what encloses the point.
static __device__ __noinline__ void init_material_list(const SppmScene &S, Tracer &T, MatList &l,
                                                       V3 init_point, V3 init_dir, Rng &rng) {
  if (S.transp_t.n < 2) return;  // the reference cannot run without 2 transparent triangles
  float t = 999.f;
  V3 back = unit(-init_dir);
  V3 o = init_point + t * init_dir;
  SurfRec rec;
  surf_init(rec);
  bool hit = trace_surf(S, S.transp_t, false, T, o, back, rec);
  int guard = 0;
  while (hit && t - rec.t > GPT_SMALL && guard++ < 4096) {
    t -= rec.t;
    RefRec ref;
    bool sss;
    check_next_path<true>(S.sc.mats, S.atlas, rec.mat, rec.in_dir, rec.normal, rec.u, rec.v, l, ref,
                          sss, rng, true);
    rearrange_list(S, l, rec.mat, ref);
    hit = trace_surf(S, S.transp_t, false, T, rec.point, back, rec);
  }
}

// henyey_greenstein_cos_theta + Material::get_new_scattering_direction
// (vector_util.h:65-73, material.h:208-221)
static __device__ __noinline__ V3 new_scattering_direction(const MaterialDev &m, V3 cur, Rng &rng) {
  float g = m.g;
  float eta = rng_uniform(rng);
  float ct;
  if (fabsf(g) < GPT_SMALL) ct = 1.f - (2.f * eta);
  else ct = (-1.f / fabsf(2.f * g)) * (1.f + g * g - powf((1.f - g * g) / (1.f - g + 2.f * g * eta), 2.f));
  float st = powf(1.f - powf(ct, 2.f), .5f);
  float cot = ct / st;
  Frame f = frame_from_z(cur);
  V3 nd = sample_disk(rng);
  float z = cot * powf(nd.x * nd.x + nd.y * nd.y, .5f);
  nd = make_unit(V3{nd.x, nd.y, z});
  return frame_to_world(f, nd);
}

// henyey_greenstein_pdf (vector_util.h:56-63)
__device__ __forceinline__ float hg_pdf(float g, V3 d1, V3 d2) {
  float ct = dot(unit(d1), unit(d2));
  if (isnan(ct)) ct = 0.f;
  return (float)((1.0 / (4 * GPT_PI)) *
                 (double)((1.f - g * g) / powf(1.f + g * g - 2.f * g * ct, 1.5f)));
}

// ---- photon records ----------------------------------------------------------------------------
// raw, per photon of this rank's range: 3 float4 + kind
//   p0 = (loc.xyz, prev.x) p1 = (prev.yz, dir.xy) p2 = (dir.z, colour.rgb)
// kind: 0 surface, 1 volume, 2 not stored (location = INF in the reference)
struct PhotonOut {
  V3 loc, prev, dir, color;
  int kind;
};

struct PhotonParams {
  int pass, max_bounce;
  long long first;       // global index of this rank's first photon
  int count;             // photons of this rank
  const float *cdf;      // accumulated mean light energy, light-BVH order
  int nlight;
};

// pick_primitive_idx_for_sampling (photon_pass.h:77-95): first idx with !(rn2 > cdf[idx]),
// capped at n-1 -- the linear scan's answer, found by bisection (the cdf is non-decreasing)
__device__ __forceinline__ int pick_light(const float *__restrict__ cdf, int n, float rn) {
  float x = rn * __ldg(cdf + n - 1);
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (x > __ldg(cdf + mid)) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// photon_pass (photon_pass.h:97-237).  One thread per photon.
__global__ void __launch_bounds__(SP_BLOCK)
k_photon_trace(SppmScene S, PhotonParams pp, uint32_t *rng_state, float4 *raw, uint32_t *kind) {
  __shared__ int stack_mem[SP_STACK * SP_BLOCK];
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  Tracer T{stack_mem + threadIdx.x, 0u};
  if (i < pp.count) {
    PhotonOut out;
    out.kind = 2;
    out.loc = V3{INFINITY, INFINITY, INFINITY};
    out.prev = out.dir = out.color = V3{0.f, 0.f, 0.f};
    Rng G, L;
    rng_load(G, rng_state, i, pp.count);
    L = G;
    for (int k = 0; k < pp.pass; k++) (void)rng_uniform(L);
    const float max_energy = __ldg(pp.cdf + pp.nlight - 1);
    int li = pick_light(pp.cdf, pp.nlight, rng_uniform(L));
    // Triangle::get_random_point_on_surface (triangle.h:199-224)
    SurfRec rec;
    surf_init(rec);
    {
      const float4 *tp = S.lt.tris + 3 * (size_t)li;
      float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
      int prim = __float_as_int(t2.z);
      TriShade ts = load_tri_shade(S.sc.shade, prim);
      float r1 = rng_uniform(L), r2 = rng_uniform(L);
      float u = 1.f - powf(r1, .5f);
      float v = r2 * powf(r1, .5f);
      float w = 1.f - u - v;
      rec.point = (u * V3{t0.x, t0.y, t0.z} + v * V3{t0.w, t1.x, t1.y}) + w * V3{t1.z, t1.w, t2.x};
      rec.normal = unit((u * ts.n1 + v * ts.n2) + w * ts.n3);
      rec.u = (u * ts.u1 + v * ts.u2) + w * ts.u3;
      rec.v = (u * ts.v1 + v * ts.v2) + w * ts.v3;
      rec.mat = ts.mat;
      Frame f = frame_from_z(rec.normal);
      V3 rd = frame_to_world(f, sample_hemisphere(L));
      rec.in_dir = unit(-rd);
    }
    V3 color = mat_emission(S.atlas, S.sc.mats[rec.mat], rec.u, rec.v);
    color = color * (max_energy / ((color.x + color.y + color.z) / 3.f));
    const float mean_color = (color.x + color.y + color.z) / 3.f;
    (void)generate_dir(V3{0.f, 0.f, 0.f}, rec.normal, 2, 1.f, L);  // ray is overwritten below
    V3 prev_location = rec.point;
    MatList list;
#pragma unroll
    for (int k = 0; k < MATLIST_MAX; k++) list.id[k] = MAT_AIR;
    list.len = 0;
    SurfRec rec_medium;
    surf_init(rec_medium);
    RefRec ref;
    ref.reflected = ref.refracted = ref.diffuse = ref.false_hit = ref.entering = ref.bent = false;
    ref.n = 1.f; ref.k = V3{0.f, 0.f, 0.f}; ref.perfect = V3{1.f, 0.f, 0.f};
    ref.dir = V3{1.f, 0.f, 0.f}; ref.filter = ref.filter2 = V3{0.f, 0.f, 0.f};
    ref.next_material = MAT_AIR;
    int num_bounce = -1, medium = MAT_AIR;
    bool hit = false, sss = false, in_medium = false, scattered_in_medium = false;
    bool direct_check_surface = false;
    bool done = false;
    while (!done && ((hit && num_bounce < pp.max_bounce) || num_bounce < 0)) {
      bool scattered_now = false;
      bool have_trace = false;  // rec_medium already holds the trace of ref.dir from rec.point
      num_bounce++;
      if (num_bounce <= 0) init_material_list(S, T, list, rec.point, rec.normal, L);
      check_next_path<true>(S.sc.mats, S.atlas, rec.mat, rec.in_dir, rec.normal, rec.u, rec.v, list,
                            ref, sss, G, false);
      if (num_bounce > 0) rearrange_list(S, list, rec.mat, ref);
      in_medium = entering_medium(S, ref, in_medium, medium);
      if (!ref.false_hit) {
        if (in_medium && !direct_check_surface && medium != MAT_AIR) {
          const MaterialDev &md = S.sc.mats[medium];
          V3 ro = rec.point, rd = ref.dir;
          float d = -logf(rng_uniform(L)) / md.extinction_coef;
          hit = trace_surf(S, S.main_t, true, T, ro, rd, rec_medium);
          have_trace = true;
          int guard = 0;
          while (d - rec_medium.t < GPT_SMALL && guard++ < 4096) {
            scattered_in_medium = true;
            scattered_now = true;
            float rn = rng_uniform(L);
            if (rn < md.scattering_prob) {
              out.kind = 1;
              out.prev = prev_location;
              out.loc = ro + d * rd;
              out.color = color;
              out.dir = rec_medium.in_dir;
              done = true;
              break;
            }
            V3 nd = new_scattering_direction(md, rd, L);
            d = -logf(rng_uniform(L)) / md.extinction_coef;
            ro = ro + d * rd;
            rd = unit(nd);
            prev_location = rec_medium.point;
            hit = trace_surf(S, S.main_t, true, T, ro, rd, rec_medium);
          }
          if (done) break;
          direct_check_surface = true;
          if (scattered_now) rec = rec_medium;
        }
        if (!scattered_now && num_bounce > 0) {
          float rn = rng_uniform(L);
          float refl = fmaxf(fmaxf(ref.k.x, ref.k.y), ref.k.z);
          if (rn > refl) {
            if (ref.diffuse && (num_bounce > 1 || scattered_in_medium)) {
              out.kind = 0;
              out.prev = prev_location;
              out.loc = rec.point;
              out.color = color;
              out.dir = rec.in_dir;
            }
            done = true;
            break;
          }
          color = ref.k * color;
          float mean_tmp = (color.x + color.y + color.z) / 3.f;
          color = color * (mean_color / mean_tmp);
          direct_check_surface = false;
        }
      }
      if (!scattered_now) {
        prev_location = rec.point;
        if (have_trace) {  // the reference re-traces the identical ray (photon_pass.h:230-233)
          rec = rec_medium;
        } else {
          hit = trace_surf(S, S.main_t, true, T, rec.point, ref.dir, rec);
        }
      }
    }
    rng_store(G, rng_state, i, pp.count);
    raw[3 * (size_t)i] = make_float4(out.loc.x, out.loc.y, out.loc.z, out.prev.x);
    raw[3 * (size_t)i + 1] = make_float4(out.prev.y, out.prev.z, out.dir.x, out.dir.y);
    raw[3 * (size_t)i + 2] = make_float4(out.dir.z, out.color.x, out.color.y, out.color.z);
    kind[i] = (uint32_t)out.kind;
  }
  // warp-aggregated ray count
  unsigned tot = T.nrays;
  for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
  if ((threadIdx.x & 31) == 0 && tot) atomicAdd(S.counters + 2, (unsigned long long)tot);
}

// ---- photon pass as a wavefront ------------------------------------------------------------------
// The same photon_pass, cut at its closest-hit traversals of the main BVH: k_photon_step runs a
// photon's logic up to the next ray it needs (or to its end), parks the live state in HBM (16
// float4 = 8 whole sectors per photon) and queues the ray; k_ph_trace answers all queued rays with
// the persistent, ray-refilling traversal the path tracer uses; the next k_photon_step resumes
// every photon where it stopped.  One thread-per-photon megakernel (k_photon_trace above, kept as
// the bit-exact cross-check, GPT_SPPM_MEGAKERNEL=1) leaves most lanes idle: photons end after
// 1..max_bounce(+scatter events) rays, and its three traversal call sites serialise.
struct PhotonWave {
  float4 *state;   // 16 per photon
  float4 *rays;    // 2 per photon: {o, tmax} {unit d, -}
  float4 *hits;    // 2 per photon: {t, leaf, prim, -} {b1, b2, b3, -}
  int *queue[2];   // photons waiting for a main-BVH ray
  int *iqueue[2];  // photons waiting for a transparent-BVH ray (init_material_list)
  int *counts;     // [0] [1] queue lengths, [2] trace cursor; [3] [4] [5] same for iqueue
  uint32_t *key, *key_alt;  // coherence sort of the traced queue: (resume point, hit material)
  int *val, *val_alt;
};

struct PhotonRaySource {
  const int *queue;
  const int *qcount;
  const float4 *rays;
  float4 *hits;
  const float4 *shade;  // per-triangle shading records (material id in s6.y)
  uint32_t *sort_key;   // out: key of queue entry i (already offset for this queue)
  int *sort_val;        // out: its slot
  __device__ int count() const { return *qcount; }
  __device__ void load(int i, float4 &o, float4 &d) const {
    int s = queue[i];
    o = rays[2 * (size_t)s];
    d = rays[2 * (size_t)s + 1];
  }
  __device__ void store_warp(bool fin, int i, const TraceLane &l) const {
    if (fin) store(i, l);
  }
  __device__ void store(int i, const TraceLane &l) const {
    int s = queue[i];
    hits[2 * (size_t)s] = make_float4(l.t, __int_as_float(l.leaf), __int_as_float(l.prim), 0.f);
    hits[2 * (size_t)s + 1] = make_float4(l.b1, l.b2, l.b3, 0.f);
    // photons that will run the same code in k_photon_step become neighbours: same resume point
    // (stored by the step in the ray's spare word), same material at the hit
    uint32_t key = 63u;
    if (l.leaf >= 0) key = (uint32_t)__float_as_int(__ldg(&shade[7 * (size_t)l.prim + 6].y)) & 63u;
    key |= (uint32_t)__float_as_int(rays[2 * (size_t)s + 1].w) << 6;
    sort_key[i] = key;
    sort_val[i] = s;
  }
};

template <bool COUNT, bool COOP, bool DEFER = false>
__global__ void __launch_bounds__(TP_BLOCK, TP_MINBLOCKS)
k_ph_trace(PhotonRaySource src, const float4 *__restrict__ nodes, const float4 *__restrict__ tris,
           const int *__restrict__ parent_leaf, int *next_ray, unsigned long long *counters) {
  __shared__ int stack[TP_STACK_SMEM * TP_BLOCK];
  __shared__ int coop[COOP ? (TP_BLOCK / 32) * COOP_WARP_WORDS : 1];
  __shared__ int defer[DEFER ? 2 * DEFER_SLOTS * TP_BLOCK : 1];
  trace_persistent<COUNT, PhotonRaySource, COOP, DEFER>(nodes, tris, parent_leaf, src, next_ray,
                                                        stack, counters, coop, defer);
}

template <bool COUNT>
__global__ void __launch_bounds__(TP_BLOCK, W4_MINBLOCKS)
k_ph_trace_wide(PhotonRaySource src, WideTree tree, int *next_ray, unsigned long long *counters) {
  __shared__ int smem[W4_SMEM_WORDS];
  trace_wide<COUNT, PhotonRaySource>(tree, src, next_ray, smem, counters);
}

enum { PH_NEW = 0, PH_MEDIUM = 1, PH_SURFACE = 2, PH_INIT = 3 };  // where a photon resumes

struct PhotonLive {
  Rng G, L;
  SurfRec rec, recm;
  V3 refk, color, prev, ro, rd;
  float mean_color, d;
  int num_bounce, medium, guard, resume;
  bool hit, in_medium, scattered_in_medium, dcs, scattered_now, have_trace, ref_diffuse;
  MatList list;
};

__device__ __forceinline__ void rng_bare(Rng &r) {
  r.s.boxmuller_flag = 0;
  r.s.boxmuller_flag_double = 0;
  r.s.boxmuller_extra = 0.f;
  r.s.boxmuller_extra_double = 0.0;
}

__device__ __forceinline__ void photon_save(const PhotonLive &p, float4 *st) {
  auto F = [](uint32_t u) { return __uint_as_float(u); };
  auto I = [](int i) { return __int_as_float(i); };
  const Rng &G = p.G, &L = p.L;
  st[0] = make_float4(F(G.s.d), F(G.s.v[0]), F(G.s.v[1]), F(G.s.v[2]));
  st[1] = make_float4(F(G.s.v[3]), F(G.s.v[4]), F(L.s.d), F(L.s.v[0]));
  st[2] = make_float4(F(L.s.v[1]), F(L.s.v[2]), F(L.s.v[3]), F(L.s.v[4]));
  const SurfRec &a = p.rec, &b = p.recm;
  st[3] = make_float4(a.t, a.point.x, a.point.y, a.point.z);
  st[4] = make_float4(a.normal.x, a.normal.y, a.normal.z, a.u);
  st[5] = make_float4(a.v, I(a.mat), a.in_dir.x, a.in_dir.y);
  st[6] = make_float4(a.in_dir.z, b.t, b.point.x, b.point.y);
  st[7] = make_float4(b.point.z, b.normal.x, b.normal.y, b.normal.z);
  st[8] = make_float4(b.u, b.v, I(b.mat), b.in_dir.x);
  st[9] = make_float4(b.in_dir.y, b.in_dir.z, p.refk.x, p.refk.y);
  st[10] = make_float4(p.refk.z, p.color.x, p.color.y, p.color.z);
  st[11] = make_float4(p.mean_color, p.prev.x, p.prev.y, p.prev.z);
  st[12] = make_float4(p.ro.x, p.ro.y, p.ro.z, p.d);
  st[13] = make_float4(p.rd.x, p.rd.y, p.rd.z, I(p.num_bounce));
  int flags = (p.hit ? 1 : 0) | (p.in_medium ? 2 : 0) | (p.scattered_in_medium ? 4 : 0) |
              (p.dcs ? 8 : 0) | (p.scattered_now ? 16 : 0) | (p.have_trace ? 32 : 0) |
              (p.ref_diffuse ? 64 : 0) | (p.resume << 8);
  st[14] = make_float4(I(p.medium), I(p.guard), I(flags), I(p.list.len));
  uint32_t w[3];
#pragma unroll
  for (int k = 0; k < 3; k++)
    w[k] = (uint32_t)p.list.id[4 * k] | ((uint32_t)p.list.id[4 * k + 1] << 8) |
           ((uint32_t)p.list.id[4 * k + 2] << 16) | ((uint32_t)p.list.id[4 * k + 3] << 24);
  st[15] = make_float4(F(w[0]), F(w[1]), F(w[2]), 0.f);
}

__device__ __forceinline__ void photon_load(PhotonLive &p, const float4 *st) {
  auto U = [](float f) { return __float_as_uint(f); };
  auto I = [](float f) { return __float_as_int(f); };
  float4 q[16];
#pragma unroll
  for (int k = 0; k < 16; k++) q[k] = st[k];
  Rng &G = p.G, &L = p.L;
  G.s.d = U(q[0].x); G.s.v[0] = U(q[0].y); G.s.v[1] = U(q[0].z); G.s.v[2] = U(q[0].w);
  G.s.v[3] = U(q[1].x); G.s.v[4] = U(q[1].y); L.s.d = U(q[1].z); L.s.v[0] = U(q[1].w);
  L.s.v[1] = U(q[2].x); L.s.v[2] = U(q[2].y); L.s.v[3] = U(q[2].z); L.s.v[4] = U(q[2].w);
  rng_bare(G);
  rng_bare(L);
  SurfRec &a = p.rec, &b = p.recm;
  a.t = q[3].x; a.point = V3{q[3].y, q[3].z, q[3].w};
  a.normal = V3{q[4].x, q[4].y, q[4].z}; a.u = q[4].w;
  a.v = q[5].x; a.mat = I(q[5].y); a.in_dir = V3{q[5].z, q[5].w, q[6].x};
  b.t = q[6].y; b.point = V3{q[6].z, q[6].w, q[7].x};
  b.normal = V3{q[7].y, q[7].z, q[7].w};
  b.u = q[8].x; b.v = q[8].y; b.mat = I(q[8].z); b.in_dir = V3{q[8].w, q[9].x, q[9].y};
  p.refk = V3{q[9].z, q[9].w, q[10].x};
  p.color = V3{q[10].y, q[10].z, q[10].w};
  p.mean_color = q[11].x; p.prev = V3{q[11].y, q[11].z, q[11].w};
  p.ro = V3{q[12].x, q[12].y, q[12].z}; p.d = q[12].w;
  p.rd = V3{q[13].x, q[13].y, q[13].z}; p.num_bounce = I(q[13].w);
  p.medium = I(q[14].x); p.guard = I(q[14].y);
  int flags = I(q[14].z);
  p.list.len = I(q[14].w);
  p.hit = flags & 1; p.in_medium = flags & 2; p.scattered_in_medium = flags & 4;
  p.dcs = flags & 8; p.scattered_now = flags & 16; p.have_trace = flags & 32;
  p.ref_diffuse = flags & 64; p.resume = (flags >> 8) & 3;
  uint32_t w[3] = {U(q[15].x), U(q[15].y), U(q[15].z)};
#pragma unroll
  for (int k = 0; k < 12; k++) p.list.id[k] = (unsigned char)((w[k >> 2] >> (8 * (k & 3))) & 255u);
}

// One step of photon_pass (photon_pass.h:97-237) for photon slot s: from where the photon stopped
// to the next ray it needs -- returns 1: a main-BVH ray, 2: a transparent-BVH ray of
// init_material_list (material_list_operations.h:61-92); the ray is in W.rays and the state is
// parked -- or to its end (returns 0: record written).  The statements and their order are those
// of k_photon_trace.
template <bool FIRST>
__device__ __forceinline__ int photon_step_one(const SppmScene &S, const PhotonParams &pp,
                                               const PhotonWave &W, int s, uint32_t *rng_state,
                                               float4 *raw, uint32_t *kind) {
  int emit = 0;
  {
    PhotonLive p;
    PhotonOut out;
    out.kind = 2;
    out.loc = V3{INFINITY, INFINITY, INFINITY};
    out.prev = out.dir = out.color = V3{0.f, 0.f, 0.f};
    RefRec ref;
    bool sss = false, finished = false;
    int resume;
    V3 ray_o = V3{0.f, 0.f, 0.f}, ray_d = V3{1.f, 0.f, 0.f};
    if (FIRST) {
      resume = PH_NEW;
      rng_load(p.G, rng_state, s, pp.count);
      p.L = p.G;
      for (int k = 0; k < pp.pass; k++) (void)rng_uniform(p.L);
      const float max_energy = __ldg(pp.cdf + pp.nlight - 1);
      int li = pick_light(pp.cdf, pp.nlight, rng_uniform(p.L));
      surf_init(p.rec);
      {
        const float4 *tp = S.lt.tris + 3 * (size_t)li;
        float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
        int prim = __float_as_int(t2.z);
        TriShade ts = load_tri_shade(S.sc.shade, prim);
        float r1 = rng_uniform(p.L), r2 = rng_uniform(p.L);
        float u = 1.f - powf(r1, .5f);
        float v = r2 * powf(r1, .5f);
        float w = 1.f - u - v;
        p.rec.point = (u * V3{t0.x, t0.y, t0.z} + v * V3{t0.w, t1.x, t1.y}) + w * V3{t1.z, t1.w, t2.x};
        p.rec.normal = unit((u * ts.n1 + v * ts.n2) + w * ts.n3);
        p.rec.u = (u * ts.u1 + v * ts.u2) + w * ts.u3;
        p.rec.v = (u * ts.v1 + v * ts.v2) + w * ts.v3;
        p.rec.mat = ts.mat;
        Frame f = frame_from_z(p.rec.normal);
        V3 rd = frame_to_world(f, sample_hemisphere(p.L));
        p.rec.in_dir = unit(-rd);
      }
      p.color = mat_emission(S.atlas, S.sc.mats[p.rec.mat], p.rec.u, p.rec.v);
      p.color = p.color * (max_energy / ((p.color.x + p.color.y + p.color.z) / 3.f));
      p.mean_color = (p.color.x + p.color.y + p.color.z) / 3.f;
      (void)generate_dir(V3{0.f, 0.f, 0.f}, p.rec.normal, 2, 1.f, p.L);
      p.prev = p.rec.point;
#pragma unroll
      for (int k = 0; k < MATLIST_MAX; k++) p.list.id[k] = MAT_AIR;
      p.list.len = 0;
      surf_init(p.recm);
      p.refk = V3{0.f, 0.f, 0.f};
      p.ref_diffuse = false;
      p.ro = p.rd = V3{0.f, 0.f, 0.f};
      p.d = 0.f;
      p.num_bounce = -1;
      p.medium = MAT_AIR;
      p.guard = 0;
      p.hit = p.in_medium = p.scattered_in_medium = p.dcs = false;
      p.scattered_now = p.have_trace = false;
      if (S.transp_t.n >= 2) {  // init_material_list: walk back from 999 units along the normal
        p.d = 999.f;            // (t of the reference's loop, kept in the free-flight slot)
        ray_o = p.rec.point + p.d * p.rec.normal;
        ray_d = unit(-p.rec.normal);
        p.resume = PH_INIT;
        emit = 2;
      }
    } else {
      photon_load(p, W.state + 16 * (size_t)s);
      resume = p.resume;
      float4 h0 = W.hits[2 * (size_t)s], h1 = W.hits[2 * (size_t)s + 1];
      float4 rdv = W.rays[2 * (size_t)s + 1];
      SurfRec &target = resume == PH_SURFACE ? p.rec : p.recm;
      p.hit = fill_surf(S, resume == PH_INIT ? S.transp_t : S.main_t, h0.x, __float_as_int(h0.y),
                        __float_as_int(h0.z), h1.x, h1.y, h1.z, V3{rdv.x, rdv.y, rdv.z}, target);
      if (resume == PH_INIT) {
        if (p.hit && p.d - p.recm.t > GPT_SMALL && p.guard++ < 4096) {
          p.d -= p.recm.t;
          RefRec iref;
          bool isss;
          check_next_path<true>(S.sc.mats, S.atlas, p.recm.mat, p.recm.in_dir, p.recm.normal,
                                p.recm.u, p.recm.v, p.list, iref, isss, p.L, true);
          rearrange_list(S, p.list, p.recm.mat, iref);
          ray_o = p.recm.point;
          ray_d = unit(-p.rec.normal);
          emit = 2;
        } else {  // the walk is over: the loop proper starts with fresh records
          surf_init(p.recm);
          p.hit = false;
          p.guard = 0;
          p.d = 0.f;
        }
      }
    }
    ref.k = p.refk;
    ref.diffuse = p.ref_diffuse;
    ref.reflected = ref.refracted = ref.false_hit = ref.entering = ref.bent = false;
    ref.n = 1.f; ref.perfect = V3{1.f, 0.f, 0.f};
    ref.dir = V3{1.f, 0.f, 0.f}; ref.filter = ref.filter2 = V3{0.f, 0.f, 0.f};
    ref.next_material = MAT_AIR;
    bool in_medium_walk = resume == PH_MEDIUM;
    while (emit == 0) {
      if (!in_medium_walk) {
        // ---- loop head (photon_pass.h:141) ----
        if (!((p.hit && p.num_bounce < pp.max_bounce) || p.num_bounce < 0)) {
          finished = true;
          break;
        }
        p.scattered_now = false;
        p.have_trace = false;
        p.num_bounce++;
        // (init_material_list of bounce 0 has already run as the PH_INIT steps)
        check_next_path<true>(S.sc.mats, S.atlas, p.rec.mat, p.rec.in_dir, p.rec.normal, p.rec.u,
                              p.rec.v, p.list, ref, sss, p.G, false);
        if (p.num_bounce > 0) rearrange_list(S, p.list, p.rec.mat, ref);
        p.in_medium = entering_medium(S, ref, p.in_medium, p.medium);
      }
      if (in_medium_walk || !ref.false_hit) {
        if (in_medium_walk || (p.in_medium && !p.dcs && p.medium != MAT_AIR)) {
          const MaterialDev &md = S.sc.mats[p.medium];
          if (!in_medium_walk) {
            p.ro = p.rec.point;
            p.rd = ref.dir;
            p.d = -logf(rng_uniform(p.L)) / md.extinction_coef;
            p.guard = 0;
            ray_o = p.ro; ray_d = p.rd;
            p.resume = PH_MEDIUM;
            emit = 1;
            break;
          }
          in_medium_walk = false;
          p.have_trace = true;
          if (p.d - p.recm.t < GPT_SMALL && p.guard++ < 4096) {
            p.scattered_in_medium = true;
            p.scattered_now = true;
            float rn = rng_uniform(p.L);
            if (rn < md.scattering_prob) {
              out.kind = 1;
              out.prev = p.prev;
              out.loc = p.ro + p.d * p.rd;
              out.color = p.color;
              out.dir = p.recm.in_dir;
              finished = true;
              break;
            }
            V3 nd = new_scattering_direction(md, p.rd, p.L);
            p.d = -logf(rng_uniform(p.L)) / md.extinction_coef;
            p.ro = p.ro + p.d * p.rd;
            p.rd = unit(nd);
            p.prev = p.recm.point;
            ray_o = p.ro; ray_d = p.rd;
            p.resume = PH_MEDIUM;
            emit = 1;
            break;
          }
          p.dcs = true;
          if (p.scattered_now) p.rec = p.recm;
        }
        if (!p.scattered_now && p.num_bounce > 0) {
          float rn = rng_uniform(p.L);
          float refl = fmaxf(fmaxf(ref.k.x, ref.k.y), ref.k.z);
          if (rn > refl) {
            if (ref.diffuse && (p.num_bounce > 1 || p.scattered_in_medium)) {
              out.kind = 0;
              out.prev = p.prev;
              out.loc = p.rec.point;
              out.color = p.color;
              out.dir = p.rec.in_dir;
            }
            finished = true;
            break;
          }
          p.color = ref.k * p.color;
          float mean_tmp = (p.color.x + p.color.y + p.color.z) / 3.f;
          p.color = p.color * (p.mean_color / mean_tmp);
          p.dcs = false;
        }
      }
      if (!p.scattered_now) {
        p.prev = p.rec.point;
        if (p.have_trace) {  // the reference re-traces the identical ray (photon_pass.h:230-233)
          p.rec = p.recm;
        } else {
          ray_o = p.rec.point; ray_d = ref.dir;
          p.resume = PH_SURFACE;
          emit = 1;
          break;
        }
      }
    }
    if (emit) {
      p.refk = ref.k;
      p.ref_diffuse = ref.diffuse;
      W.rays[2 * (size_t)s] = make_float4(ray_o.x, ray_o.y, ray_o.z, INFINITY);
      W.rays[2 * (size_t)s + 1] = make_float4(ray_d.x, ray_d.y, ray_d.z, __int_as_float(p.resume));
      photon_save(p, W.state + 16 * (size_t)s);
    } else {
      (void)finished;
      rng_store(p.G, rng_state, s, pp.count);
      raw[3 * (size_t)s] = make_float4(out.loc.x, out.loc.y, out.loc.z, out.prev.x);
      raw[3 * (size_t)s + 1] = make_float4(out.prev.y, out.prev.z, out.dir.x, out.dir.y);
      raw[3 * (size_t)s + 2] = make_float4(out.dir.z, out.color.x, out.color.y, out.color.z);
      kind[s] = (uint32_t)out.kind;
    }
  }
  return emit;
}

// Persistent over the (sorted) slot list: a thread's local-memory frame is reused from photon to
// photon instead of one frame per photon going out to HBM.  n = n_init + n_main entries.
template <bool FIRST>
__global__ void __launch_bounds__(SP_BLOCK)
k_photon_step(SppmScene S, PhotonParams pp, PhotonWave W, int cur, const int *__restrict__ slots,
              int n, int n_main, uint32_t *rng_state, float4 *raw, uint32_t *kind) {
  if (!FIRST && blockIdx.x == 0 && threadIdx.x == 0)
    atomicAdd(S.counters + 2, (unsigned long long)n_main);  // main-BVH rays just traced
  const int lane = threadIdx.x & 31;
  const int nxt = FIRST ? 0 : (cur ^ 1);
  for (int base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
    const int a = base + threadIdx.x;
    int s = -1, emit = 0;
    if (a < n) {
      s = FIRST ? a : slots[a];
      emit = photon_step_one<FIRST>(S, pp, W, s, rng_state, raw, kind);
    }
    // warp-aggregated append to the next main / init queue
#pragma unroll
    for (int q = 1; q <= 2; q++) {
      unsigned m = __ballot_sync(0xffffffffu, emit == q);
      if (m) {
        int leader = __ffs(m) - 1, at = 0;
        int *cnt = W.counts + (q == 1 ? nxt : 3 + nxt);
        if (lane == leader) at = atomicAdd(cnt, __popc(m));
        at = __shfl_sync(0xffffffffu, at, leader);
        if (emit == q) (q == 1 ? W.queue[nxt] : W.iqueue[nxt])[at + __popc(m & ((1u << lane) - 1u))] = s;
      }
    }
  }
}

__global__ void __launch_bounds__(256) k_iota(int *v, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = i;
}

// sorted kinds -> [n_surface, n_volume]
__global__ void __launch_bounds__(256) k_kind_counts(const uint32_t *__restrict__ k, int n, int *counts) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t a = k[i], b = (i + 1 < n) ? k[i + 1] : 3u;
  if (a == 0u && b != 0u) counts[0] = i + 1;  // last surface photon
  if (a <= 1u && b >= 2u) counts[1] = i + 1;  // last stored photon (surface or volume)
}

// Point::assign_location boxes (point.h:58-66): location -/+ SMALL_DOUBLE, in list order
__global__ void __launch_bounds__(256)
k_photon_boxes(const float4 *__restrict__ raw, const int *__restrict__ list, int n, float *aabb) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  float4 p0 = __ldg(raw + 3 * (size_t)(list ? __ldg(list + j) : j));  // list == nullptr: raw is the list
  float *o = aabb + 6 * (size_t)j;
  o[0] = p0.x - GPT_SMALL; o[1] = p0.x + GPT_SMALL;
  o[2] = p0.y - GPT_SMALL; o[3] = p0.y + GPT_SMALL;
  o[4] = p0.z - GPT_SMALL; o[5] = p0.z + GPT_SMALL;
}

// photon records in LEAF order of their tree (48 B, same 3 float4 as raw)
__global__ void __launch_bounds__(256)
k_pack_photons(const float4 *__restrict__ raw, const int *__restrict__ list,
               const int *__restrict__ sorted_prim, int n, float4 *leaf) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  int at = sorted_prim ? __ldg(sorted_prim + j) : j;
  int src = list ? __ldg(list + at) : at;
  for (int k = 0; k < 3; k++) leaf[3 * (size_t)j + k] = __ldg(raw + 3 * (size_t)src + k);
}

// compute_bs_union (bounding_sphere.h:72-96)
__device__ __forceinline__ float4 sphere_union(float4 a, float4 b) {
  V3 c1 = V3{a.x, a.y, a.z}, c2 = V3{b.x, b.y, b.z};
  V3 cd = make_unit(c2 - c1);
  V3 p1 = c1 - a.w * cd, p2 = c1 + a.w * cd, p3 = c2 - b.w * cd, p4 = c2 + b.w * cd;
  float tmin = -fminf(dist(p1, c1), dist(p3, c1));
  float tmax = fmaxf(dist(p2, c1), dist(p4, c1));
  float r = (-tmin + tmax) / 2.f;
  V3 c = (((c1 + tmin * cd) + c1) + tmax * cd) / 2.f;
  return make_float4(c.x, c.y, c.z, r);
}

// compute_node_bounding_spheres (bvh_building.h:264-302): bottom-up, the second arrival unions
// (left, right) in that order.  Sphere node = 2 float4 (left child sphere, right child sphere);
// topology comes from the AABB tree's q3 words.
__global__ void __launch_bounds__(256)
k_sphere_refit(const float4 *__restrict__ leafrec, int n, const float4 *nodes,
               const int *__restrict__ parent_leaf, int *flags, float4 *spheres) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  float4 p0 = __ldg(leafrec + 3 * (size_t)j);
  float4 mine = make_float4(p0.x, p0.y, p0.z, GPT_SMALL);
  int pe = __ldg(parent_leaf + j);
  while (pe >= 0) {
    int par = pe >> 1, right = pe & 1;
    spheres[2 * (size_t)par + right] = mine;
    __threadfence();
    if (atomicAdd(flags + par, 1) == 0) return;
    __threadfence();
    float4 sib = __ldcg(spheres + 2 * (size_t)par + (right ^ 1));
    mine = right ? sphere_union(sib, mine) : sphere_union(mine, sib);
    pe = __ldcg(reinterpret_cast<const int *>(nodes + 4 * (size_t)par + 3) + 2);
  }
}

// ---- eye pass ------------------------------------------------------------------------------------
struct PhotonTree {
  const float4 *nodes;     // AABB nodes (topology + child boxes)
  const float4 *spheres;   // volume tree only
  const float4 *leaf;      // 3 float4 per photon, leaf order
  int n;
};

struct EyeParams {
  CameraDev cam;
  int pass;        // pass_iteration
  int init;
  float radius_multiplier;
  PhotonTree vol;
  // r02 sharded eye pass (multi-GPU, SURVEY 8(e) second variant): this rank traces the eye paths
  // of its own 32x32 tiles only (the path tracer's round-robin deal) and writes what the pass
  // produces -- location, normal, filter, this pass's direct light -- as 12 floats per pixel into
  // `xc`, MINUS zeros for the other ranks' pixels, so ONE sum all-reduce gives every rank every
  // hit point (x + (-0) + ... = x exactly; -0 and INFINITY included); k_eye_unpack applies it.
  int rank, nranks;
  float4 *xc;  // xstride float4 per pixel: (loc, nrm.x) (nrm.yz, filter.xy) (filter.z, direct.xyz) and,
               // in the init pass (xstride 4), (radius, -, -, -); nullptr = unsharded
  int xstride;
};

struct HitPoints {
  int P;
  float4 *loc;       // xyz, surface radius
  float4 *nrm;       // xyz, volume radius
  float4 *filter;
  float4 *direct;
  float4 *indirect;
  float4 *flux;      // [0,P): flux.rgb + count of this pass; [P,2P): tmp_accummulated_lm
};

// BoundingCylinder::is_point_inside (bounding_cylinder.h:59-76)
__device__ __forceinline__ bool cyl_inside(V3 p0, V3 axis, float l, float r, V3 loc, float buffer,
                                           float &dperp, float &dpar) {
  V3 sp = loc - p0;
  float d = length(sp);
  float ct = dot(axis, sp / d);
  float st = powf(1.f - ct * ct, .5f);
  dperp = d * st;
  if (dperp <= r + buffer) {
    dpar = d * ct;
    if (dpar <= l + buffer && dpar >= -buffer) return true;
  }
  return false;
}

// traverse_bvh_volume_photon + _compute_volume_photon_contribution
// (bvh_traversal_photon.h:15-106), beam (p0, axis, l, r = volume radius) vs the sphere tree
static __device__ __noinline__ V3 volume_gather(const SppmScene &S, const PhotonTree &vt, V3 p0,
                                                V3 axis, float l, float vr, int medium) {
  V3 contrib = V3{0.f, 0.f, 0.f}, tmp = V3{0.f, 0.f, 0.f};
  if (vt.n < 2 || medium == MAT_AIR) return contrib;
  const MaterialDev &md = S.sc.mats[medium];
  int stack[64];
  int sp = 0, node = 0;
  while (true) {
    const int *q3 = reinterpret_cast<const int *>(vt.nodes + 4 * (size_t)node + 3);
    int cref[2] = {__ldg(q3), __ldg(q3 + 1)};
    bool isect[2];
#pragma unroll
    for (int c = 0; c < 2; c++) {
      float4 s = __ldg(vt.spheres + 2 * (size_t)node + c);
      float dp, da;
      isect[c] = cyl_inside(p0, axis, l, vr, V3{s.x, s.y, s.z}, s.w, dp, da);
    }
#pragma unroll
    for (int c = 0; c < 2; c++) {
      if (isect[c] && cref[c] < 0) {
        const float4 *ph = vt.leaf + 3 * (size_t)(~cref[c]);
        float4 a = __ldg(ph), b = __ldg(ph + 1), d = __ldg(ph + 2);
        float dperp, dpar;
        V3 o = V3{0.f, 0.f, 0.f};
        if (cyl_inside(p0, axis, l, vr, V3{a.x, a.y, a.z}, 0.f, dperp, dpar)) {
          float x = dperp / vr;
          float silver = (float)((3 / GPT_PI) * (double)powf(1.f - x * x, 2.f));
          float kern = powf((float)(1.0 / (double)vr), 2.f) * silver;
          if (kern > GPT_SMALL) {
            float tr = md.extinction_coef > GPT_SMALL ? expf(-dpar * md.extinction_coef) : 1.0f;
            float phase = hg_pdf(md.g, axis, V3{b.z, b.w, d.x});
            o = (((kern * tr) * md.scattering_coef) * phase) * V3{d.y, d.z, d.w};
          }
        }
        tmp = tmp + de_nan(o);
        contrib = contrib + tmp;
      }
    }
    bool tl = isect[0] && cref[0] >= 0, tr = isect[1] && cref[1] >= 0;
    if (!tl && !tr) {
      if (sp == 0) break;
      node = stack[--sp];
    } else {
      node = tl ? cref[0] : cref[1];
      if (tl && tr && sp < 64) stack[sp++] = cref[1];
    }
  }
  return contrib;
}

// _compute_direct_radiance (ray_tracing_pass.h:19-141)
static __device__ __noinline__ V3 direct_radiance(const SppmScene &S, Tracer &T, V3 filter,
                                                  SurfRec rec, RefRec ref, bool in_medium,
                                                  int medium, MatList &list, Rng &rng) {
  V3 emittance = filter * mat_emission(S.atlas, S.sc.mats[rec.mat], rec.u, rec.v);
  V3 direct = V3{0.f, 0.f, 0.f};
  bool prev_in_medium = in_medium, flag = true;
  int prev_medium = medium;
  float factor = change_ref_ray(S.lt, S.sc.shade, S.sc.mats, S.atlas, rec.point, rec.normal,
                                rec.in_dir, ref, 1.f, rng);
  filter = filter * ref.filter;
  V3 prev_hit_point = rec.point;
  bool hit = trace_surf(S, S.main_t, true, T, rec.point, ref.dir, rec);
  int guard = 0;
  while (flag && hit && guard++ < 4096) {
    bool sss;
    check_next_path<true>(S.sc.mats, S.atlas, rec.mat, rec.in_dir, rec.normal, rec.u, rec.v, list,
                          ref, sss, rng, true);
    rearrange_list(S, list, rec.mat, ref);
    in_medium = entering_medium(S, ref, in_medium, medium);
    if (!ref.false_hit && prev_in_medium) {
      float l = length(rec.point - prev_hit_point);
      filter = filter * transmittance(S, prev_medium, l);
    }
    filter = filter * ref.filter2;
    if (!ref.false_hit && ref.diffuse)
      direct = direct + (filter * clamp_lohi(0.f, .999999f, factor)) *
                            mat_emission(S.atlas, S.sc.mats[rec.mat], rec.u, rec.v);
    if (!ref.refracted || ref.bent) flag = false;
    if (!ref.false_hit) {
      prev_hit_point = rec.point;
      prev_in_medium = in_medium;
      prev_medium = medium;
    }
    if (!flag) break;  // the reference traces one more ray whose result it never reads
    hit = trace_surf(S, S.main_t, true, T, rec.point, ref.dir, rec);
  }
  direct = direct + emittance;
  return V3{fminf(1.0f, direct.x), fminf(1.0f, direct.y), fminf(1.0f, direct.z)};
}

struct EyeResult {
  bool hit, diffuse;
  V3 point, normal;
};

// _get_hit_point_details (ray_tracing_pass.h:170-414).  `filter` and `direct` are in/out as in
// the reference (the init pass shares them across its five calls); lm accumulates the volume term.
static __device__ __noinline__ EyeResult hit_point_details(const SppmScene &S, const EyeParams &ep,
                                                           Tracer &T, int row, int col, float off_w,
                                                           float off_h, float vol_radius, Rng &rng,
                                                           V3 &filter, V3 &direct, V3 &lm) {
  EyeResult er;
  MatList list;
#pragma unroll
  for (int k = 0; k < MATLIST_MAX; k++) list.id[k] = MAT_AIR;
  list.len = 1;  // add_new_material(material_list, material_list_length, nullptr)
  direct = V3{0.f, 0.f, 0.f};
  V3 o, d;
  camera_ray(ep.cam, row + off_h, col + off_w, rng, o, d);
  filter = V3{1.f, 1.f, 1.f};
  V3 filter_lag = filter;
  init_material_list(S, T, list, o, V3{0.f, 0.f, 1.f}, rng);
  // _set_init_ref (ray_tracing_pass.h:143-168)
  RefRec ref;
  {
    int hi, second;
    matlist_top2(S.sc.mats, list, hi, second);
    ref.false_hit = false; ref.reflected = true; ref.refracted = false;
    ref.dir = d;
    ref.filter = ref.filter2 = V3{1.f, 1.f, 1.f};
    ref.diffuse = false; ref.bent = false; ref.entering = true;
    ref.next_material = hi;
    ref.k = V3{0.f, 0.f, 0.f}; ref.perfect = V3{1.f, 0.f, 0.f}; ref.n = 1.f;
  }
  int medium = MAT_AIR;
  bool in_medium = entering_medium(S, ref, false, medium);
  bool prev_in_medium = in_medium;
  int prev_medium = medium;
  V3 prev_hit_point = o;
  SurfRec rec;
  surf_init(rec);
  bool hit = trace_surf(S, S.main_t, true, T, o, d, rec);
  int num_bounce = 0;
  if (hit) {
    while (!ref.diffuse && hit && num_bounce < 64) {
      num_bounce++;
      bool sss;
      check_next_path<true>(S.sc.mats, S.atlas, rec.mat, rec.in_dir, rec.normal, rec.u, rec.v, list,
                            ref, sss, rng, false);
      rearrange_list(S, list, rec.mat, ref);
      in_medium = entering_medium(S, ref, in_medium, medium);
      if (!ref.false_hit && prev_in_medium && !ep.init) {
        V3 seg = rec.point - prev_hit_point;
        float l = length(seg);
        V3 axis = unit(seg);
        lm = lm + filter * volume_gather(S, ep.vol, prev_hit_point, axis, l, vol_radius, medium);
        filter = filter * transmittance(S, prev_medium, l);
      }
      if (!ref.false_hit) {
        filter_lag = filter;
        filter = filter * ref.filter2;
      }
      if (ref.diffuse) {
        direct = direct + direct_radiance(S, T, filter_lag, rec, ref, in_medium, medium, list, rng);
        break;
      }
      if (!ref.false_hit) {
        prev_in_medium = in_medium;
        prev_hit_point = rec.point;
        prev_medium = medium;
      }
      hit = trace_surf(S, S.main_t, true, T, rec.point, ref.dir, rec);
    }
  }
  er.hit = hit;
  er.diffuse = ref.diffuse;
  er.point = rec.point;
  er.normal = rec.normal;
  return er;
}

// ray_tracing_pass (ray_tracing_pass.h:484-592).  One thread per pixel.
__global__ void __launch_bounds__(SP_BLOCK)
k_eye_pass(SppmScene S, EyeParams ep, HitPoints hp, const uint32_t *__restrict__ rng_init) {
  __shared__ int stack_mem[SP_STACK * SP_BLOCK];
  int pix = blockIdx.x * blockDim.x + threadIdx.x;
  Tracer T{stack_mem + threadIdx.x, 0u};
  bool mine = pix < hp.P;
  if (mine && ep.xc) {
    const int row = pix / ep.cam.width, col = pix % ep.cam.width;
    const int tiles_x = (ep.cam.width + 31) / 32;
    mine = ((row / 32) * tiles_x + col / 32) % ep.nranks == ep.rank;
    if (!mine) {  // another rank's pixel: the identity of the sum into the exchange buffer and
                  // the volume term.  That identity is MINUS zero: x + (-0) == x for every x,
                  // -0 and INFINITY included, whereas (-0) + (+0) would turn into +0
      for (int k = 0; k < ep.xstride; k++)
        ep.xc[ep.xstride * (size_t)pix + k] = make_float4(-0.f, -0.f, -0.f, -0.f);
      if (ep.init) {
        hp.direct[pix] = make_float4(0.f, 0.f, 0.f, 0.f);
        hp.indirect[pix] = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        hp.flux[(size_t)hp.P + pix] = make_float4(-0.f, -0.f, -0.f, -0.f);
      }
    }
  }
  if (mine) {
    int row = pix / ep.cam.width, col = pix % ep.cam.width;
    Rng rng;
    rng_load(rng, rng_init, pix, hp.P);
    for (int k = 0; k < ep.pass; k++) (void)rng_uniform(rng);
    float off_h = .5f, off_w = .5f;
    float radius, vol_radius;
    if (ep.init) {
      radius = INFINITY;  // PPMHitPoint::PPMHitPoint (ppm_hit_point.h:173-188)
      vol_radius = INFINITY;
      hp.direct[pix] = make_float4(0.f, 0.f, 0.f, 0.f);
      hp.indirect[pix] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      off_h = rng_uniform(rng);
      off_w = rng_uniform(rng);
      radius = hp.loc[pix].w;
      vol_radius = hp.nrm[pix].w;
    }
    V3 filter = V3{1.f, 1.f, 1.f}, direct = V3{0.f, 0.f, 0.f}, lm = V3{0.f, 0.f, 0.f};
    EyeResult er = hit_point_details(S, ep, T, row, col, off_w, off_h, vol_radius, rng, filter,
                                     direct, lm);
    if (ep.init) {
      const float cw[4] = {0.f, 0.f, 1.f, 1.f}, ch[4] = {0.f, 1.f, 0.f, 1.f};
      V3 loc4[4];
      for (int k = 0; k < 4; k++) {
        V3 dummy_direct, dummy_lm = V3{0.f, 0.f, 0.f};
        EyeResult e2 = hit_point_details(S, ep, T, row, col, cw[k], ch[k], vol_radius, rng, filter,
                                         dummy_direct, dummy_lm);
        loc4[k] = (e2.hit && e2.diffuse) ? e2.point : V3{INFINITY, INFINITY, INFINITY};
      }
      for (int a = 0; a < 4; a++) {
        if (!has_inf(loc4[a]) && !has_inf(er.point) && er.hit && er.diffuse) {
          float rt = dist(loc4[a], er.point);
          if (radius > rt && rt > 0.f) radius = rt;
        }
        for (int b = a; b < 4; b++) {
          if (!has_inf(loc4[b]) && !has_inf(loc4[a])) {
            float rt = dist(loc4[a], loc4[b]);
            if (radius > rt && rt > 0.f) radius = rt;
          }
        }
      }
      radius *= ep.radius_multiplier;
    }
    if (ep.xc) {  // sharded: everything the pass produces goes through the exchange buffer
      V3 loc = V3{INFINITY, INFINITY, INFINITY}, nrm = V3{0.f, 0.f, 1.f}, dn = V3{0.f, 0.f, 0.f};
      if (er.hit && er.diffuse) { loc = er.point; nrm = er.normal; dn = de_nan(direct); }
      float4 *x = ep.xc + ep.xstride * (size_t)pix;
      x[0] = make_float4(loc.x, loc.y, loc.z, nrm.x);
      x[1] = make_float4(nrm.y, nrm.z, filter.x, filter.y);
      x[2] = make_float4(filter.z, dn.x, dn.y, dn.z);
      if (ep.init) x[3] = make_float4(radius, -0.f, -0.f, -0.f);
    } else {
    if (er.hit && er.diffuse) {
      hp.loc[pix] = make_float4(er.point.x, er.point.y, er.point.z, radius);
      hp.nrm[pix] = make_float4(er.normal.x, er.normal.y, er.normal.z, vol_radius);
      V3 dn = de_nan(direct);
      float4 dr = hp.direct[pix];
      hp.direct[pix] = make_float4(dr.x + dn.x, dr.y + dn.y, dr.z + dn.z, 0.f);
    } else {
      hp.loc[pix] = make_float4(INFINITY, INFINITY, INFINITY, radius);
      hp.nrm[pix] = make_float4(0.f, 0.f, 1.f, vol_radius);
    }
    hp.filter[pix] = make_float4(filter.x, filter.y, filter.z, 0.f);
    }
    if (!ep.init) hp.flux[(size_t)hp.P + pix] = make_float4(lm.x, lm.y, lm.z, 0.f);
  }
  unsigned tot = T.nrays;
  for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
  if ((threadIdx.x & 31) == 0 && tot) atomicAdd(S.counters + 2, (unsigned long long)tot);
}

// The sharded eye pass's results, after the all-reduce, into the hit points: exactly the stores of
// the unsharded kernel (radii stay, direct light accumulates; `0 + d` for a miss is what the
// unsharded pass's skipped add leaves, d being finite).
__global__ void __launch_bounds__(256)
k_eye_unpack(HitPoints hp, const float4 *__restrict__ xc, int xstride, int init) {
  int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= hp.P) return;
  const float4 *x = xc + xstride * (size_t)pix;
  const float4 a = x[0], b = x[1], c = x[2];
  // the init pass also sets the radii: surface = its own estimate, volume = INFINITY
  // (PPMHitPoint::PPMHitPoint, ppm_hit_point.h:173-188); later passes keep them
  const float radius = init ? x[3].x : hp.loc[pix].w, vradius = init ? INFINITY : hp.nrm[pix].w;
  hp.loc[pix] = make_float4(a.x, a.y, a.z, radius);
  hp.nrm[pix] = make_float4(a.w, b.x, b.y, vradius);
  hp.filter[pix] = make_float4(b.z, b.w, c.x, 0.f);
  if (!(a.x == INFINITY)) {  // er.hit && er.diffuse on the owner
    const float4 dr = hp.direct[pix];
    hp.direct[pix] = make_float4(dr.x + c.y, dr.y + c.z, dr.z + c.w, 0.f);
  }
}

// compute_average_radius (ray_tracing_pass.h:416-436): serial fp32 sum in pixel order, then
// assign_radius_to_invalid_hit_points gives EVERY hit point that mean (src/main.cu:1353).
// One warp streams the radii (coalesced), lane order = pixel order, lane 0 carries the sum.
// The sum is a chain of P dependent fp32 adds in pixel order; everything else is taken off that
// chain: 7 warps stage the next tile of radii into shared memory (coalesced) while thread 0 adds
// the current one.
constexpr int MR_TILE = 2048;
__global__ void __launch_bounds__(256) k_mean_radius_serial(const float4 *__restrict__ loc, int P, float *out) {
  __shared__ float tile[2][MR_TILE];
  const int t = threadIdx.x;
  float sum = 0.f;
  int valid = 0;
  auto stage = [&](int buf, int base) {
    if (t >= 32)
      for (int k = t - 32; k < MR_TILE; k += 224) {
        int i = base + k;
        tile[buf][k] = i < P ? loc[i].w : INFINITY;
      }
  };
  stage(0, 0);
  __syncthreads();
  int buf = 0;
  for (int base = 0; base < P; base += MR_TILE, buf ^= 1) {
    stage(buf ^ 1, base + MR_TILE);
    if (t == 0) {
#pragma unroll 16
      for (int k = 0; k < MR_TILE; k++) {
        float r = tile[buf][k];
        if (!isinf(r)) { valid++; sum += r; }
      }
    }
    __syncthreads();
  }
  if (t == 0) { out[0] = sum / (float)valid; out[1] = (float)valid; }
}
__global__ void __launch_bounds__(256)
k_assign_radius(HitPoints hp, const float *__restrict__ mean) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hp.P) return;
  float r = mean[0];
  hp.loc[i].w = r;           // PPMHitPoint::update_radius (ppm_hit_point.h:105-109)
  hp.nrm[i].w = 2.f * r;
}

// per-pass mean radius (printed only by the reference, src/main.cu:1787-1798): parallel sum
__global__ void __launch_bounds__(256)
k_mean_radius_stat(const float4 *__restrict__ loc, int P, double *acc) {
  double s = 0.0, c = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
    float r = loc[i].w;
    if (!isinf(r)) { s += (double)r; c += 1.0; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  if ((threadIdx.x & 31) == 0) { atomicAdd(acc, s); atomicAdd(acc + 1, c); }
}

// ---- gather --------------------------------------------------------------------------------------
// update_hit_point_parameters / _traverse_bvh_surface_photon (bvh_traversal_photon.h:108-225):
// AABB-vs-sphere descent of the surface-photon tree in the reference's left-first order (the
// fp32 flux sum is order dependent), exact distance test at the leaves, one visibility ray
// through the full geometry BVH per candidate.
__global__ void __launch_bounds__(SP_BLOCK)
k_gather(SppmScene S, PhotonTree st, HitPoints hp) {
  __shared__ int stack_mem[SP_STACK * SP_BLOCK];
  int pix = blockIdx.x * blockDim.x + threadIdx.x;
  Tracer T{stack_mem + threadIdx.x, 0u};
  if (pix < hp.P) {
    float4 hl = hp.loc[pix];
    V3 flux = V3{0.f, 0.f, 0.f};
    int count = 0;
    V3 c = V3{hl.x, hl.y, hl.z};
    if (!has_inf(c) && st.n >= 2) {
      float4 hn = hp.nrm[pix];
      V3 nrm = V3{hn.x, hn.y, hn.z};
      float r = hl.w;
      int stack[64];
      int sp = 0, node = 0;
      while (true) {
        const float4 *q = st.nodes + 4 * (size_t)node;
        float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
        int cref[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
        float bxn[2] = {q0.x, q1.z}, bxx[2] = {q0.y, q1.w};
        float byn[2] = {q0.z, q2.x}, byx[2] = {q0.w, q2.y};
        float bzn[2] = {q1.x, q2.z}, bzx[2] = {q1.y, q2.w};
        bool isect[2];
#pragma unroll
        for (int k = 0; k < 2; k++) {
          // BoundingBox::is_intersection(BoundingSphere*) (bounding_box.h:278-294); centre =
          // 0.5 * (min + max) (bounding_box.h:171-173)
          float cx = 0.5f * (bxn[k] + bxx[k]), cy = 0.5f * (byn[k] + byx[k]),
                cz = 0.5f * (bzn[k] + bzx[k]);
          isect[k] = fabsf(cx - c.x) <= (bxx[k] - cx) + r && fabsf(cy - c.y) <= (byx[k] - cy) + r &&
                     fabsf(cz - c.z) <= (bzx[k] - cz) + r;
        }
#pragma unroll
        for (int k = 0; k < 2; k++) {
          if (isect[k] && cref[k] < 0) {
            const float4 *ph = st.leaf + 3 * (size_t)(~cref[k]);
            float4 a = __ldg(ph);
            if (dist(c, V3{a.x, a.y, a.z}) <= r) {  // BoundingSphere::is_inside
              float4 b = __ldg(ph + 1), d = __ldg(ph + 2);
              V3 rd = V3{a.w, b.x, b.y} - c;  // prev_location - hit point
              float tv;
              bool gh = trace_t(S, S.main_t, T, c, unit(rd), tv);
              if (gh && tv > length(rd)) {
                count++;
                float f = fmaxf(0.0f, dot(nrm, -V3{b.z, b.w, d.x}));
                flux = flux + f * V3{d.y, d.z, d.w};
              }
            }
          }
        }
        bool tl = isect[0] && cref[0] >= 0, tr = isect[1] && cref[1] >= 0;
        if (!tl && !tr) {
          if (sp == 0) break;
          node = stack[--sp];
        } else {
          node = tl ? cref[0] : cref[1];
          if (tl && tr && sp < 64) stack[sp++] = cref[1];
        }
      }
    }
    hp.flux[pix] = make_float4(flux.x, flux.y, flux.z, (float)count);
  }
  unsigned tot = T.nrays;
  for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
  if ((threadIdx.x & 31) == 0 && tot) atomicAdd(S.counters + 2, (unsigned long long)tot);
}

// PPMHitPoint::update_accummulated_reflected_flux (ppm_hit_point.h:115-171)
__global__ void __launch_bounds__(256)
k_update(HitPoints hp, int iteration, float alpha, int emitted) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hp.P) return;
  float4 hl = hp.loc[i], hn = hp.nrm[i];
  float rs = hl.w, rv = hn.w, nrs = rs, nrv = rv;
  if (iteration >= 2) {
    float ratio = ((float)iteration + alpha) / (float)(iteration + 1);
    nrs = rs * powf(ratio, 0.5f);
    nrv = rv * powf(ratio, (float)(1.0 / 3));
  }
  float4 fl = hp.flux[i], lm = hp.flux[(size_t)hp.P + i], ft = hp.filter[i];
  float den = (float)((double)emitted * GPT_PI * (double)powf(rs, 2.f));
  V3 surf = de_nan((V3{ft.x, ft.y, ft.z} * V3{fl.x, fl.y, fl.z}) / den);
  V3 vol = de_nan(V3{lm.x, lm.y, lm.z} / (float)emitted);
  float4 in = hp.indirect[i];
  V3 add = surf + vol;
  hp.indirect[i] = make_float4(in.x + add.x, in.y + add.y, in.z + add.z, in.w + fl.w);
  hp.loc[i].w = nrs;
  hp.nrm[i].w = nrv;
}

// PPMHitPoint::compute_pixel_color + get_ppm_image_output (ppm_hit_point.h:78-103)
__global__ void __launch_bounds__(256)
k_resolve(HitPoints hp, int mode, int passes, float *out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hp.P) return;
  float4 d = hp.direct[i], in = hp.indirect[i];
  V3 dv = V3{d.x, d.y, d.z}, iv = V3{in.x, in.y, in.z};
  float np = (float)passes;
  V3 o = mode == 1 ? dv / np : (mode == 2 ? iv / np : (dv + iv) / np);
  o = de_nan(o);
  out[3 * (size_t)i] = o.x; out[3 * (size_t)i + 1] = o.y; out[3 * (size_t)i + 2] = o.z;
}

__global__ void __launch_bounds__(128)
k_sppm_rng_init(uint32_t *rng, long long first, int n) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  Rng r;
  curand_init(1984, (unsigned long long)(first + s), 0, &r.s);
  rng_store(r, rng, s, n);
}

// compute_accummulated_light_source_energy (photon_pass.h:60-75): serial fp32 prefix sum of
// de_nan(mean(energy)) in light-BVH order
__global__ void k_light_cdf(const float4 *__restrict__ lt_tris, const float *__restrict__ energy,
                            int n, float *cdf) {
  if (threadIdx.x || blockIdx.x) return;
  float acc = 0.f;
  for (int j = 0; j < n; j++) {
    int prim = __float_as_int(__ldg(&lt_tris[3 * (size_t)j + 2].z));
    const float *e = energy + 3 * (size_t)prim;
    float m = (e[0] + e[1] + e[2]) / 3.f;
    acc += (m == m ? m : 0.f);
    cdf[j] = acc;
  }
}

struct SppmCtx {
  int W = 0, H = 0, P = 0;
  HitPoints hp{};
  uint32_t *rng_pix = nullptr;
  // photons of this rank
  long long ph_first = -1;
  int ph_count = 0, ph_capacity = 0;
  uint32_t *rng_ph = nullptr;
  float4 *raw = nullptr;
  uint32_t *kind = nullptr, *kind_alt = nullptr, *hist = nullptr;
  int *idx = nullptr, *idx_alt = nullptr;
  float *aabb = nullptr;          // 6 * capacity, reused by both trees
  float4 *leaf_surf = nullptr, *leaf_vol = nullptr, *spheres = nullptr;
  BvhDev surf_bvh, vol_bvh;
  int n_surf = 0, n_vol = 0;
  PhotonWave wave{};               // wavefront photon pass: parked state, rays, hits, queues
  float *cdf = nullptr;
  int cdf_n = 0;
  int *d_counts = nullptr;        // [0] n_surf [1] n_surf + n_vol [2] error flag
  float *d_mean = nullptr;        // [0] mean radius [1] valid count
  double *d_stat = nullptr;
  float *d_out = nullptr;
  // light tree arrays (same as the path tracer's, built here when SPPM runs first)
  float4 *leaf_info = nullptr, *node_info = nullptr;
  float *node_energy = nullptr;
  int *lt_flags = nullptr;
  int light_n = 0;
  uint64_t light_epoch = 0;  // ctx->build_epoch the light-tree arrays were made for
  uint64_t rays_total = 0;
  // r02 sharded eye pass: exchange buffer, the volume photons of ALL ranks (compaction order of
  // the whole pass = rank order, ranks own consecutive photon ranges), their tree's buffers
  float4 *xc = nullptr;
  float4 *vol_send = nullptr, *vol_all = nullptr, *leaf_vol_all = nullptr, *spheres_all = nullptr;
  float *aabb_vol = nullptr;
  int *d_rank_counts = nullptr;
  int vol_send_cap = 0, vol_all_cap = 0;
  bool vol_is_global = false;   // this pass's volume tree holds every rank's photons
  int dbg_rank = 0, dbg_nranks = 1;  // gpt_sppm_debug_eye_shard: ownership without a communicator (tests)
};

// The eye pass is sharded when a communicator with > 1 ranks is bound (GPT_SPPM_SHARD_EYE=0 keeps
// the round-1 replicated form), or when a test asked for it on one device.
static bool eye_sharded(const gpt_ctx *ctx, const SppmCtx *sx, int *rank, int *nranks, bool *comm) {
  static int env = -1;
  if (env < 0) { const char *e = getenv("GPT_SPPM_SHARD_EYE"); env = e ? atoi(e) : 1; }
  *comm = false;
  if (ctx->nccl_comm && ctx->comm_nranks > 1 && env != 0) {
    *rank = ctx->comm_rank; *nranks = ctx->comm_nranks; *comm = true;
    return true;
  }
  if (sx->dbg_nranks > 1) { *rank = sx->dbg_rank; *nranks = sx->dbg_nranks; return true; }
  *rank = 0; *nranks = 1;
  return false;
}

SppmCtx *sx_of(gpt_ctx *ctx) {
  if (!ctx->sppm_state) ctx->sppm_state = new SppmCtx();
  return static_cast<SppmCtx *>(ctx->sppm_state);
}

}  // namespace

// light tree construction lives in render.cu
int gpt_build_light_tree_arrays(gpt_ctx *ctx, float4 **leaf_info, float4 **node_info,
                                float **node_energy, int **flags);

static int make_scene_view(gpt_ctx *ctx, SppmCtx *sx, SppmScene &S) {
  BvhDev &mb = ctx->bvh[GPT_BVH_MAIN], &tb = ctx->bvh[GPT_BVH_TRANSPARENT],
         &lb = ctx->bvh[GPT_BVH_LIGHT];
  if (lb.n < 2)
    return gpt_fail(ctx, GPT_E_INVALID, "SPPM needs >= 2 emissive triangles (src/main.cu:1124-1125)", __FILE__, __LINE__);
  if (sx->light_n != lb.n || sx->light_epoch != ctx->build_epoch) {
    GPT_TRY(gpt_build_light_tree_arrays(ctx, &sx->leaf_info, &sx->node_info, &sx->node_energy,
                                        &sx->lt_flags));
    sx->light_n = lb.n;
    sx->light_epoch = ctx->build_epoch;
  }
  S.main_t = TreeRef{mb.nodes, mb.tris, mb.parent_leaf, mb.n};
  S.transp_t = TreeRef{tb.nodes, tb.tris, tb.parent_leaf, tb.n};
  S.sc = ctx->scene;
  S.atlas = TexAtlas{ctx->scene.tex_r, ctx->scene.tex_g, ctx->scene.tex_b};
  S.lt = LightTree{lb.nodes, lb.tris, lb.parent_leaf, sx->leaf_info, sx->node_info, lb.n};
  S.counters = ctx->d_counters;
  S.error_flag = sx->d_counts + 2;
  S.count_visits = ctx->prof_count_visits ? 1 : 0;
  return GPT_OK;
}

static int sppm_reserve_pixels(gpt_ctx *ctx, SppmCtx *sx) {
  int W = ctx->cam.width, H = ctx->cam.height;
  if (W == sx->W && H == sx->H && sx->hp.loc) return GPT_OK;
  size_t P = (size_t)W * H;
  HitPoints &hp = sx->hp;
  GPT_TRY(gpt_alloc(ctx, &hp.loc, P));
  GPT_TRY(gpt_alloc(ctx, &hp.nrm, P));
  GPT_TRY(gpt_alloc(ctx, &hp.filter, P));
  GPT_TRY(gpt_alloc(ctx, &hp.direct, P));
  GPT_TRY(gpt_alloc(ctx, &hp.indirect, P));
  GPT_TRY(gpt_alloc(ctx, &hp.flux, 2 * P));
  GPT_TRY(gpt_alloc(ctx, &sx->rng_pix, 6 * P));
  GPT_TRY(gpt_alloc(ctx, &sx->d_out, 3 * P));
  if (!sx->d_counts) GPT_TRY(gpt_alloc(ctx, &sx->d_counts, (size_t)4));
  if (!sx->d_mean) GPT_TRY(gpt_alloc(ctx, &sx->d_mean, (size_t)2));
  if (!sx->d_stat) GPT_TRY(gpt_alloc(ctx, &sx->d_stat, (size_t)2));
  hp.P = (int)P;
  sx->W = W; sx->H = H; sx->P = (int)P;
  k_sppm_rng_init<<<((int)P + 127) / 128, 128, 0, ctx->stream>>>(sx->rng_pix, 0, (int)P);
  GPT_LAUNCH_CHECK(ctx);
  return GPT_OK;
}

static int sppm_reserve_photons(gpt_ctx *ctx, SppmCtx *sx, long long first, int count) {
  if (count > sx->ph_capacity) {
    size_t N = (size_t)count;
    GPT_TRY(gpt_alloc(ctx, &sx->rng_ph, 6 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->raw, 3 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->kind, N));
    GPT_TRY(gpt_alloc(ctx, &sx->kind_alt, N));
    GPT_TRY(gpt_alloc(ctx, &sx->idx, N));
    GPT_TRY(gpt_alloc(ctx, &sx->idx_alt, N));
    GPT_TRY(gpt_alloc(ctx, &sx->hist, (size_t)gpt_radix_hist_entries(count)));
    GPT_TRY(gpt_alloc(ctx, &sx->aabb, 6 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->leaf_surf, 3 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->leaf_vol, 3 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->spheres, 2 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.state, 16 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.rays, 2 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.hits, 2 * N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.queue[0], N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.queue[1], N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.iqueue[0], N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.iqueue[1], N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.key, N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.key_alt, N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.val, N));
    GPT_TRY(gpt_alloc(ctx, &sx->wave.val_alt, N));
    if (!sx->wave.counts) GPT_TRY(gpt_alloc(ctx, &sx->wave.counts, (size_t)8));
    sx->ph_capacity = count;
    sx->ph_first = -1;
  }
  if (sx->ph_first != first || sx->ph_count != count) {
    // init_curand_state<<<ppm_num_photon_per_pass, 1>>> (src/main.cu:1476): seq = photon index
    k_sppm_rng_init<<<(count + 127) / 128, 128, 0, ctx->stream>>>(sx->rng_ph, first, count);
    GPT_LAUNCH_CHECK(ctx);
    sx->ph_first = first;
    sx->ph_count = count;
  }
  return GPT_OK;
}

static int sppm_check_errors(gpt_ctx *ctx, SppmCtx *sx) {
  int h[4] = {0, 0, 0, 0};
  GPT_CUDA(ctx, cudaMemcpyAsync(h, sx->d_counts, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (h[2]) return gpt_fail(ctx, GPT_E_UNSUPPORTED, "nested-dielectric list deeper than 12", __FILE__, __LINE__);
  return GPT_OK;
}

static int sppm_collect_rays(gpt_ctx *ctx, SppmCtx *sx) {
  unsigned long long c[8];
  GPT_CUDA(ctx, cudaMemcpyAsync(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->stats.rays_traced = c[2];
  ctx->stats.internal_visits = c[0];
  ctx->stats.leaf_tests = c[1];
  ctx->stats.wide_node_visits = c[5];
  ctx->stats.wide_leaf_tests = c[7];
  sx->rays_total += c[2];
  if (c[3]) return gpt_fail(ctx, GPT_E_INVALID, "traversal stack overflow", __FILE__, __LINE__);
  return GPT_OK;
}

static int eye_pass(gpt_ctx *ctx, SppmCtx *sx, const SppmScene &S, int pass, bool init) {
  EyeParams ep;
  ep.cam = derive_camera(ctx->cam);
  ep.pass = pass;
  ep.init = init ? 1 : 0;
  ep.radius_multiplier = ctx->params.ppm_radius_scaling_factor;
  ep.vol = sx->vol_is_global
               ? PhotonTree{sx->vol_bvh.nodes, sx->spheres_all, sx->leaf_vol_all, init ? 0 : sx->n_vol}
               : PhotonTree{sx->vol_bvh.nodes, sx->spheres, sx->leaf_vol, init ? 0 : sx->n_vol};
  bool comm = false;
  const bool sharded = eye_sharded(ctx, sx, &ep.rank, &ep.nranks, &comm);  // the init pass too
  if (!sharded) { ep.rank = 0; ep.nranks = 1; }
  ep.xc = nullptr;
  ep.xstride = init ? 4 : 3;
  if (sharded) {
    if (!sx->xc) GPT_TRY(gpt_alloc(ctx, &sx->xc, 4 * (size_t)sx->P));
    ep.xc = sx->xc;
  }
  k_eye_pass<<<(sx->P + SP_BLOCK - 1) / SP_BLOCK, SP_BLOCK, 0, ctx->stream>>>(S, ep, sx->hp, sx->rng_pix);
  GPT_LAUNCH_CHECK(ctx);
  if (sharded) {
    if (comm) GPT_TRY(gpt_comm_allreduce_f32(ctx, sx->xc, 4 * (size_t)ep.xstride * sx->P));
    k_eye_unpack<<<(sx->P + 255) / 256, 256, 0, ctx->stream>>>(sx->hp, sx->xc, ep.xstride, ep.init);
    GPT_LAUNCH_CHECK(ctx);
  }
  return GPT_OK;
}

extern "C" {

void gpt_sppm_free(gpt_ctx *ctx) {
  if (!ctx || !ctx->sppm_state) return;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  void *p[] = {sx->hp.loc, sx->hp.nrm, sx->hp.filter, sx->hp.direct, sx->hp.indirect, sx->hp.flux,
               sx->rng_pix, sx->rng_ph, sx->raw, sx->kind, sx->kind_alt, sx->hist, sx->idx,
               sx->idx_alt, sx->aabb, sx->leaf_surf, sx->leaf_vol, sx->spheres, sx->cdf,
               sx->d_counts, sx->d_mean, sx->d_stat, sx->d_out, sx->leaf_info, sx->node_info,
               sx->node_energy, sx->lt_flags, sx->wave.state, sx->wave.rays, sx->wave.hits,
               sx->wave.queue[0], sx->wave.queue[1], sx->wave.iqueue[0], sx->wave.iqueue[1],
               sx->wave.counts, sx->wave.key,
               sx->wave.key_alt, sx->wave.val, sx->wave.val_alt, sx->xc, sx->vol_send, sx->vol_all,
               sx->leaf_vol_all, sx->spheres_all, sx->aabb_vol, sx->d_rank_counts};
  for (void *q : p) gpt_free(ctx, q);
  gpt_lbvh_free(ctx, sx->surf_bvh);
  gpt_lbvh_free(ctx, sx->vol_bvh);
  delete sx;
  ctx->sppm_state = nullptr;
}

// ray_tracing_pass(init = true) + compute_average_radius + assign_radius_to_invalid_hit_points
// + compute_accummulated_light_source_energy (src/main.cu:1311-1460)
int gpt_sppm_init(gpt_ctx *ctx) {
  if (!ctx) return GPT_E_INVALID;
  if (!ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  if (!ctx->have_cam) return gpt_fail(ctx, GPT_E_INVALID, "gpt_set_camera first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  SppmCtx *sx = sx_of(ctx);
  ctx->launches = 0;
  GPT_TRY(sppm_reserve_pixels(ctx, sx));
  SppmScene S;
  GPT_TRY(make_scene_view(ctx, sx, S));
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemsetAsync(sx->d_counts, 0, 4 * sizeof(int), st));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 5, 0, sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 7, 0, sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaMemsetAsync(sx->hp.flux, 0, sizeof(float4) * 2 * (size_t)sx->P, st));
  sx->rays_total = 0;
  sx->ph_first = -1;  // fresh photon streams
  sx->n_surf = sx->n_vol = 0;
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
  GPT_TRY(eye_pass(ctx, sx, S, 0, true));
  k_mean_radius_serial<<<1, 256, 0, st>>>(sx->hp.loc, sx->P, sx->d_mean);
  GPT_LAUNCH_CHECK(ctx);
  k_assign_radius<<<(sx->P + 255) / 256, 256, 0, st>>>(sx->hp, sx->d_mean);
  GPT_LAUNCH_CHECK(ctx);
  int nl = ctx->bvh[GPT_BVH_LIGHT].n;
  if (sx->cdf_n < nl) {
    GPT_TRY(gpt_alloc(ctx, &sx->cdf, (size_t)nl));
    sx->cdf_n = nl;
  }
  k_light_cdf<<<1, 32, 0, st>>>(ctx->bvh[GPT_BVH_LIGHT].tris, ctx->scene.energy, nl, sx->cdf);
  GPT_LAUNCH_CHECK(ctx);
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
  GPT_TRY(sppm_check_errors(ctx, sx));
  GPT_TRY(sppm_collect_rays(ctx, sx));
  float ms = 0.f;
  GPT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  ctx->stats.render_ms = ms;
  ctx->stats.kernel_launches = ctx->launches;
  return GPT_OK;
}

}  // extern "C"

// gather_recorded_photons (photon_pass.h:19-47) as a stable partition by kind (photon order is
// kept inside each kind), then the two photon trees (src/main.cu:1576-1752).
// photon_pass<<<ppm_num_photon_per_pass, 1>>> (src/main.cu:1496): the wavefront by default, the
// one-thread-per-photon megakernel when GPT_SPPM_MEGAKERNEL=1 (cross-check; identical records).
static int photon_stage(gpt_ctx *ctx, SppmCtx *sx, const SppmScene &S, const PhotonParams &pp) {
  cudaStream_t st = ctx->stream;
  const int count = pp.count;
  const char *mega_env = getenv("GPT_SPPM_MEGAKERNEL");  // read per call: tests flip it
  if (mega_env && atoi(mega_env) != 0) {
    k_photon_trace<<<(count + SP_BLOCK - 1) / SP_BLOCK, SP_BLOCK, 0, st>>>(S, pp, sx->rng_ph, sx->raw,
                                                                        sx->kind);
    GPT_LAUNCH_CHECK(ctx);
    return GPT_OK;
  }
  static int trace_grid = 0;
  if (!trace_grid) {
    int sms = 148, per_sm = 4;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ph_trace<false, true>, TP_BLOCK, 0);
    trace_grid = sms * (per_sm > 0 ? per_sm : 1);
  }
  static int step_grid[2] = {0, 0};
  if (!step_grid[0]) {
    int sms = 148, per_sm = 4;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_photon_step<false>, SP_BLOCK, 0);
    step_grid[0] = sms * (per_sm > 0 ? per_sm : 1);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_photon_step<true>, SP_BLOCK, 0);
    step_grid[1] = sms * (per_sm > 0 ? per_sm : 1);
    const char *e = getenv("GPT_PHOTON_WAVES");
    int waves = e ? atoi(e) : 1;
    if (waves <= 0) step_grid[0] = step_grid[1] = 0x7fffffff;
    else { step_grid[0] *= waves; step_grid[1] *= waves; }
  }
  PhotonWave &W = sx->wave;
  BvhDev &mb = ctx->bvh[GPT_BVH_MAIN], &tb = ctx->bvh[GPT_BVH_TRANSPARENT];
  GPT_CUDA(ctx, cudaMemsetAsync(W.counts, 0, 8 * sizeof(int), st));
  {
    int grid = (count + SP_BLOCK - 1) / SP_BLOCK;
    if (grid > step_grid[1]) grid = step_grid[1];
    k_photon_step<true><<<grid, SP_BLOCK, 0, st>>>(S, pp, W, 0, nullptr, count, 0, sx->rng_ph, sx->raw,
                                                   sx->kind);
    GPT_LAUNCH_CHECK(ctx);
  }
  int cur = 0;
  auto trace = [&](const BvhDev &tree, const int *queue, int *qcount, int *cursor, int n, int offset,
                   bool count_visits) -> int {
    PhotonRaySource src{queue, qcount, W.rays, W.hits, ctx->scene.shade, W.key + offset, W.val + offset};
    int grid = (n + TP_BLOCK - 1) / TP_BLOCK;
    if (grid > trace_grid) grid = trace_grid;
    if (gpt_trace_wide(ctx, tree)) {
      if (count_visits)
        k_ph_trace_wide<true><<<grid, TP_BLOCK, 0, st>>>(src, gpt_wide_tree(tree), cursor, ctx->d_counters);
      else
        k_ph_trace_wide<false><<<grid, TP_BLOCK, 0, st>>>(src, gpt_wide_tree(tree), cursor, ctx->d_counters);
      GPT_LAUNCH_CHECK(ctx);
      return GPT_OK;
    }
    const bool coop = gpt_trace_coop(ctx, tree);
    auto kern = count_visits ? (coop ? k_ph_trace<true, true> : k_ph_trace<true, false>)
                             : (coop ? k_ph_trace<false, true> : k_ph_trace<false, false>);
    if (gpt_trace_defer(ctx))
      kern = count_visits ? (coop ? k_ph_trace<true, true, true> : k_ph_trace<true, false, true>)
                          : (coop ? k_ph_trace<false, true, true> : k_ph_trace<false, false, true>);
    kern<<<grid, TP_BLOCK, 0, st>>>(src, tree.nodes, tree.tris, tree.parent_leaf, cursor,
                                    ctx->d_counters);
    GPT_LAUNCH_CHECK(ctx);
    return GPT_OK;
  };
  // every photon ends after at most max_bounce surface events + 4096 scatter events per bounce
  // (+ 4096 steps of the initial walk)
  for (long long it = 0; it < 4100LL * (pp.max_bounce + 3); it++) {
    int h[8];
    GPT_CUDA(ctx, cudaMemcpyAsync(h, W.counts, sizeof(h), cudaMemcpyDeviceToHost, st));
    GPT_CUDA(ctx, cudaStreamSynchronize(st));
    const int n_main = h[cur], n_init = h[3 + cur], n = n_main + n_init;
    if (n == 0) return GPT_OK;
    int zero_at[4] = {cur ^ 1, 2, 3 + (cur ^ 1), 5};
    for (int z : zero_at) GPT_CUDA(ctx, cudaMemsetAsync(W.counts + z, 0, sizeof(int), st));
    // keys / slots of the init queue land in [0, n_init), those of the main queue behind them
    if (n_init) GPT_TRY(trace(tb, W.iqueue[cur], W.counts + 3 + cur, W.counts + 5, n_init, 0, false));
    if (n_main) GPT_TRY(trace(mb, W.queue[cur], W.counts + cur, W.counts + 2, n_main, n_init, S.count_visits));
    // stable 8-bit radix sort of the traced photons by (resume point, hit material)
    uint32_t *sk = nullptr;
    int *sorted_slots = nullptr;
    GPT_TRY(gpt_radix_sort_pairs(ctx, W.key, W.val, W.key_alt, W.val_alt, sx->hist, n, 8, &sk,
                                 &sorted_slots, nullptr));
    int grid = (n + SP_BLOCK - 1) / SP_BLOCK;
    if (grid > step_grid[0]) grid = step_grid[0];
    k_photon_step<false><<<grid, SP_BLOCK, 0, st>>>(S, pp, W, cur, sorted_slots, n, n_main, sx->rng_ph,
                                                    sx->raw, sx->kind);
    GPT_LAUNCH_CHECK(ctx);
    cur ^= 1;
  }
  return gpt_fail(ctx, GPT_E_INVALID, "photon wavefront did not drain", __FILE__, __LINE__);
}

static int sppm_build_trees(gpt_ctx *ctx, SppmCtx *sx, int count) {
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemsetAsync(sx->d_counts, 0, 2 * sizeof(int), st));
  k_iota<<<(count + 255) / 256, 256, 0, st>>>(sx->idx, count);
  GPT_LAUNCH_CHECK(ctx);
  uint32_t *ko; int *vo;
  GPT_TRY(gpt_radix_sort_pairs(ctx, sx->kind, sx->idx, sx->kind_alt, sx->idx_alt, sx->hist, count,
                               2, &ko, &vo));
  k_kind_counts<<<(count + 255) / 256, 256, 0, st>>>(ko, count, sx->d_counts);
  GPT_LAUNCH_CHECK(ctx);
  int h[2];
  GPT_CUDA(ctx, cudaMemcpyAsync(h, sx->d_counts, sizeof(h), cudaMemcpyDeviceToHost, st));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  sx->n_surf = h[0];
  sx->n_vol = h[1] > h[0] ? h[1] - h[0] : 0;
  const int *surf_list = vo, *vol_list = vo + sx->n_surf;
  if (sx->n_surf >= 2) {
    int n = sx->n_surf;
    k_photon_boxes<<<(n + 255) / 256, 256, 0, st>>>(sx->raw, surf_list, n, sx->aabb);
    GPT_LAUNCH_CHECK(ctx);
    GPT_TRY(gpt_lbvh_build_boxes(ctx, sx->surf_bvh, sx->aabb, nullptr, n, false));
    k_pack_photons<<<(n + 255) / 256, 256, 0, st>>>(sx->raw, surf_list, sx->surf_bvh.sorted_prim, n,
                                                   sx->leaf_surf);
    GPT_LAUNCH_CHECK(ctx);
  }
  sx->vol_is_global = false;
  {
    int rank = 0, nranks = 1;
    bool comm = false;
    if (eye_sharded(ctx, sx, &rank, &nranks, &comm) && comm) {
      // every rank needs every volume photon: its eye paths gather against the WHOLE pass's
      // sphere tree, which also makes the (prefix weighted, order dependent) volume term the
      // single-GPU one.  Counts first (host), then a variable-length all-gather in rank order.
      const int mine = sx->n_vol;
      if (mine > sx->vol_send_cap) {
        GPT_TRY(gpt_alloc(ctx, &sx->vol_send, 3 * (size_t)mine));
        sx->vol_send_cap = mine;
      }
      if (mine > 0) {
        k_pack_photons<<<(mine + 255) / 256, 256, 0, st>>>(sx->raw, vol_list, nullptr, mine, sx->vol_send);
        GPT_LAUNCH_CHECK(ctx);
      }
      if (!sx->d_rank_counts) GPT_TRY(gpt_alloc(ctx, &sx->d_rank_counts, (size_t)nranks));
      std::vector<int> counts((size_t)nranks, 0);
      GPT_TRY(gpt_comm_allgather_count(ctx, mine, sx->d_rank_counts, counts.data()));
      long long total = 0;
      for (int c : counts) total += c;
      if (total > 0x7fffffff) return gpt_fail(ctx, GPT_E_INVALID, "more than 2^31 volume photons in a pass", __FILE__, __LINE__);
      if ((int)total > sx->vol_all_cap) {
        const size_t N = (size_t)total;
        GPT_TRY(gpt_alloc(ctx, &sx->vol_all, 3 * N));
        GPT_TRY(gpt_alloc(ctx, &sx->leaf_vol_all, 3 * N));
        GPT_TRY(gpt_alloc(ctx, &sx->spheres_all, 2 * N));
        GPT_TRY(gpt_alloc(ctx, &sx->aabb_vol, 6 * N));
        sx->vol_all_cap = (int)total;
      }
      GPT_TRY(gpt_comm_allgatherv(ctx, sx->vol_send, sx->vol_all, counts.data(), 3 * sizeof(float4)));
      sx->n_vol = (int)total;
      sx->vol_is_global = true;
      if (sx->n_vol >= 2) {
        const int n = sx->n_vol;
        k_photon_boxes<<<(n + 255) / 256, 256, 0, st>>>(sx->vol_all, nullptr, n, sx->aabb_vol);
        GPT_LAUNCH_CHECK(ctx);
        GPT_TRY(gpt_lbvh_build_boxes(ctx, sx->vol_bvh, sx->aabb_vol, nullptr, n, false));
        k_pack_photons<<<(n + 255) / 256, 256, 0, st>>>(sx->vol_all, nullptr, sx->vol_bvh.sorted_prim, n,
                                                       sx->leaf_vol_all);
        GPT_LAUNCH_CHECK(ctx);
        GPT_CUDA(ctx, cudaMemsetAsync(sx->vol_bvh.flags, 0, sizeof(int) * (size_t)n, st));
        k_sphere_refit<<<(n + 255) / 256, 256, 0, st>>>(sx->leaf_vol_all, n, sx->vol_bvh.nodes,
                                                       sx->vol_bvh.parent_leaf, sx->vol_bvh.flags,
                                                       sx->spheres_all);
        GPT_LAUNCH_CHECK(ctx);
      }
      return GPT_OK;
    }
  }
  if (sx->n_vol >= 2) {
    int n = sx->n_vol;
    k_photon_boxes<<<(n + 255) / 256, 256, 0, st>>>(sx->raw, vol_list, n, sx->aabb);
    GPT_LAUNCH_CHECK(ctx);
    GPT_TRY(gpt_lbvh_build_boxes(ctx, sx->vol_bvh, sx->aabb, nullptr, n, false));
    k_pack_photons<<<(n + 255) / 256, 256, 0, st>>>(sx->raw, vol_list, sx->vol_bvh.sorted_prim, n,
                                                   sx->leaf_vol);
    GPT_LAUNCH_CHECK(ctx);
    GPT_CUDA(ctx, cudaMemsetAsync(sx->vol_bvh.flags, 0, sizeof(int) * (size_t)n, st));
    k_sphere_refit<<<(n + 255) / 256, 256, 0, st>>>(sx->leaf_vol, n, sx->vol_bvh.nodes,
                                                   sx->vol_bvh.parent_leaf, sx->vol_bvh.flags,
                                                   sx->spheres);
    GPT_LAUNCH_CHECK(ctx);
  }
  return GPT_OK;
}

extern "C" {

// stage_mask: 1 = photon trace of the global range [photon_begin, photon_end) + compaction +
// both photon trees, 2 = eye pass (with the volume beam gather), 4 = surface gather.
int gpt_sppm_pass_stages(gpt_ctx *ctx, int pass, int64_t photon_begin, int64_t photon_end,
                         int stage_mask) {
  if (!ctx || pass < 0 || photon_end < photon_begin) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  if (photon_end - photon_begin > 0x7fffffff) return GPT_E_INVALID;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->launches = 0;
  int count = (int)(photon_end - photon_begin);
  SppmScene S;
  GPT_TRY(make_scene_view(ctx, sx, S));
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 5, 0, sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 7, 0, sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
  if (stage_mask & 1) {
    sx->n_surf = sx->n_vol = 0;
    if (count > 0) {
      GPT_TRY(sppm_reserve_photons(ctx, sx, photon_begin, count));
      PhotonParams pp;
      pp.pass = pass;
      pp.max_bounce = ctx->params.ppm_max_bounce;
      pp.first = photon_begin;
      pp.count = count;
      pp.cdf = sx->cdf;
      pp.nlight = ctx->bvh[GPT_BVH_LIGHT].n;
      GPT_TRY(photon_stage(ctx, sx, S, pp));
      GPT_TRY(sppm_build_trees(ctx, sx, count));
    }
  }
  if (stage_mask & 2) GPT_TRY(eye_pass(ctx, sx, S, pass, false));
  if (stage_mask & 4) {
    PhotonTree stree{sx->surf_bvh.nodes, nullptr, sx->leaf_surf, sx->n_surf};
    k_gather<<<(sx->P + SP_BLOCK - 1) / SP_BLOCK, SP_BLOCK, 0, st>>>(S, stree, sx->hp);
    GPT_LAUNCH_CHECK(ctx);
  }
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
  GPT_TRY(sppm_check_errors(ctx, sx));
  GPT_TRY(sppm_collect_rays(ctx, sx));
  float ms = 0.f;
  GPT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  ctx->stats.render_ms = ms;
  ctx->stats.kernel_launches = ctx->launches;
  return GPT_OK;
}

// Test hook: shard the eye pass as rank `rank` of `nranks` WITHOUT a communicator (no exchange:
// the other ranks' pixels come out zero), so one device can check that the shards tile the
// unsharded pass.  nranks <= 1 switches it off.
int gpt_sppm_debug_eye_shard(gpt_ctx *ctx, int rank, int nranks) {
  if (!ctx || rank < 0 || (nranks > 1 && rank >= nranks)) return GPT_E_INVALID;
  SppmCtx *sx = sx_of(ctx);
  sx->dbg_rank = nranks > 1 ? rank : 0;
  sx->dbg_nranks = nranks > 1 ? nranks : 1;
  return GPT_OK;
}

// One pass up to (not including) the hit-point update.
int gpt_sppm_pass(gpt_ctx *ctx, int pass, int64_t photon_begin, int64_t photon_end) {
  return gpt_sppm_pass_stages(ctx, pass, photon_begin, photon_end, 7);
}

int gpt_sppm_flux_device(gpt_ctx *ctx, void **d_flux, int64_t *num_floats) {
  if (!ctx || !d_flux || !num_floats) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.flux) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  *d_flux = sx->hp.flux;
  *num_floats = 8 * (int64_t)sx->P;
  return GPT_OK;
}

int gpt_sppm_update(gpt_ctx *ctx, int pass) {
  if (!ctx || pass < 0) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  k_update<<<(sx->P + 255) / 256, 256, 0, ctx->stream>>>(sx->hp, pass + 1, ctx->params.ppm_alpha,
                                                        ctx->params.ppm_num_photon_per_pass);
  GPT_LAUNCH_CHECK(ctx);
  return GPT_OK;
}

int gpt_sppm_resolve(gpt_ctx *ctx, int mode, int passes, float *out_rgb) {
  if (!ctx || !out_rgb || passes < 1 || mode < 0 || mode > 2) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  k_resolve<<<(sx->P + 255) / 256, 256, 0, ctx->stream>>>(sx->hp, mode, passes, sx->d_out);
  GPT_LAUNCH_CHECK(ctx);
  GPT_CUDA(ctx, cudaMemcpyAsync(out_rgb, sx->d_out, sizeof(float) * 3 * (size_t)sx->P,
                                cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return GPT_OK;
}

// ---- parity exports / imports -----------------------------------------------------------------
int gpt_sppm_export_hit_points(gpt_ctx *ctx, float *loc_radius /*P x 4*/, float *normal_vradius,
                               float *filter, float *direct, float *indirect, float *flux_lm /*2P x 4*/) {
  if (!ctx) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  size_t B = sizeof(float4) * (size_t)sx->P;
  auto dl = [&](float *h, const float4 *d, size_t bytes) -> cudaError_t {
    if (!h) return cudaSuccess;
    return cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream);
  };
  GPT_CUDA(ctx, dl(loc_radius, sx->hp.loc, B));
  GPT_CUDA(ctx, dl(normal_vradius, sx->hp.nrm, B));
  GPT_CUDA(ctx, dl(filter, sx->hp.filter, B));
  GPT_CUDA(ctx, dl(direct, sx->hp.direct, B));
  GPT_CUDA(ctx, dl(indirect, sx->hp.indirect, B));
  GPT_CUDA(ctx, dl(flux_lm, sx->hp.flux, 2 * B));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return GPT_OK;
}

// photons of the last pass in photon-index order: kind (0 surface, 1 volume, 2 none) and the
// 12 floats loc, prev, dir, colour
int gpt_sppm_export_photons(gpt_ctx *ctx, int32_t *kind, float *rec12, int64_t capacity,
                            int64_t *count) {
  if (!ctx || !count) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->raw) return gpt_fail(ctx, GPT_E_INVALID, "no photon pass yet", __FILE__, __LINE__);
  *count = sx->ph_count;
  if (capacity < sx->ph_count) return GPT_OK;
  // kinds were sorted in place: rebuild them from the raw records' stored flag
  std::vector<float> raw(12 * (size_t)sx->ph_count);
  GPT_CUDA(ctx, cudaMemcpy(raw.data(), sx->raw, sizeof(float) * raw.size(), cudaMemcpyDeviceToHost));
  std::vector<uint32_t> k(sx->ph_count);
  std::vector<int> order(sx->ph_count);
  // the final sorted (kind, idx) pairs live in kind/idx or their alt buffers (2 bits = 1 pass)
  GPT_CUDA(ctx, cudaMemcpy(k.data(), sx->kind_alt, 4 * (size_t)sx->ph_count, cudaMemcpyDeviceToHost));
  GPT_CUDA(ctx, cudaMemcpy(order.data(), sx->idx_alt, 4 * (size_t)sx->ph_count, cudaMemcpyDeviceToHost));
  for (int j = 0; j < sx->ph_count; j++) {
    int i = order[j];
    if (kind) kind[i] = (int32_t)k[j];
  }
  if (rec12) {
    for (int i = 0; i < sx->ph_count; i++) {
      const float *r = raw.data() + 12 * (size_t)i;
      float *o = rec12 + 12 * (size_t)i;
      o[0] = r[0]; o[1] = r[1]; o[2] = r[2];        // loc
      o[3] = r[3]; o[4] = r[4]; o[5] = r[5];        // prev
      o[6] = r[6]; o[7] = r[7]; o[8] = r[8];        // dir
      o[9] = r[9]; o[10] = r[10]; o[11] = r[11];    // colour
    }
  }
  return GPT_OK;
}

// Parity hook: replace the photons of the current pass by a caller-supplied list (host; kind
// 0 surface / 1 volume / 2 none, 12 floats each as exported) and build both photon trees from
// it, so the eye pass (stage 2) and the gather (stage 4) can be compared with the reference's
// on the reference's own photons.
int gpt_sppm_import_photons(gpt_ctx *ctx, const int32_t *kind, const float *rec12, int64_t n) {
  if (!ctx || !kind || !rec12 || n < 2 || n > 0x7fffffff) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->launches = 0;
  int count = (int)n;
  GPT_TRY(sppm_reserve_photons(ctx, sx, 0, count));
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemcpyAsync(sx->raw, rec12, sizeof(float) * 12 * (size_t)count,
                                cudaMemcpyHostToDevice, st));
  GPT_CUDA(ctx, cudaMemcpyAsync(sx->kind, kind, sizeof(int32_t) * (size_t)count,
                                cudaMemcpyHostToDevice, st));
  GPT_TRY(sppm_build_trees(ctx, sx, count));
  return GPT_OK;
}

// Parity hook: overwrite the hit-point geometry (location + surface radius, normal + volume
// radius, filter) with caller-supplied values, P x 4 host floats each (NULL = keep).
int gpt_sppm_import_hit_points(gpt_ctx *ctx, const float *loc_radius, const float *normal_vradius,
                               const float *filter) {
  if (!ctx) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  size_t B = sizeof(float4) * (size_t)sx->P;
  auto ul = [&](float4 *d, const float *h) -> cudaError_t {
    if (!h) return cudaSuccess;
    return cudaMemcpyAsync(d, h, B, cudaMemcpyHostToDevice, ctx->stream);
  };
  GPT_CUDA(ctx, ul(sx->hp.loc, loc_radius));
  GPT_CUDA(ctx, ul(sx->hp.nrm, normal_vradius));
  GPT_CUDA(ctx, ul(sx->hp.filter, filter));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return GPT_OK;
}

// create_point_image (point_operations.h:19-75): which 0 = hit points (colour = filter,
// src/main.cu:1362-1382), 1 = surface photons, 2 = volume photons of the last pass
// (src/main.cu:1517-1551).  out_rgb: P x 3 host floats.
int gpt_splat_device(gpt_ctx *ctx, const float *d_loc, int loc_stride, const float *d_col,
                     int col_stride, int n, const int *d_order, float *out_rgb_host);  // output.cu
int gpt_sppm_debug_image(gpt_ctx *ctx, int which, float *out_rgb) {
  if (!ctx || !out_rgb || which < 0 || which > 2) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  if (which == 0)
    return gpt_splat_device(ctx, reinterpret_cast<const float *>(sx->hp.loc), 4,
                            reinterpret_cast<const float *>(sx->hp.filter), 4, sx->P, nullptr, out_rgb);
  const float4 *leaf = which == 1 ? sx->leaf_surf : (sx->vol_is_global ? sx->leaf_vol_all : sx->leaf_vol);
  int n = which == 1 ? sx->n_surf : sx->n_vol;
  if (n < 2) n = 0;  // no tree was built: nothing recorded
  // the reference splats its photon lists BEFORE it sorts them for the trees (src/main.cu:1534,
  // 1545 against :1561-): array order = compaction order = sorted_prim of our leaf order
  const int *order = which == 1 ? sx->surf_bvh.sorted_prim : sx->vol_bvh.sorted_prim;
  return gpt_splat_device(ctx, reinterpret_cast<const float *>(leaf), 12,
                          reinterpret_cast<const float *>(leaf) + 9, 12, n, order, out_rgb);
}

int gpt_sppm_stats(gpt_ctx *ctx, int64_t *n_surface, int64_t *n_volume, double *mean_radius,
                   uint64_t *rays_total) {
  if (!ctx) return GPT_E_INVALID;
  SppmCtx *sx = static_cast<SppmCtx *>(ctx->sppm_state);
  if (!sx || !sx->hp.loc) return gpt_fail(ctx, GPT_E_INVALID, "gpt_sppm_init first", __FILE__, __LINE__);
  if (n_surface) *n_surface = sx->n_surf;
  if (n_volume) *n_volume = sx->n_vol;
  if (rays_total) *rays_total = sx->rays_total;
  if (mean_radius) {
    GPT_CUDA(ctx, cudaMemsetAsync(sx->d_stat, 0, 2 * sizeof(double), ctx->stream));
    k_mean_radius_stat<<<148 * 4, 256, 0, ctx->stream>>>(sx->hp.loc, sx->P, sx->d_stat);
    GPT_LAUNCH_CHECK(ctx);
    double h[2];
    GPT_CUDA(ctx, cudaMemcpyAsync(h, sx->d_stat, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *mean_radius = h[1] > 0 ? h[0] / h[1] : 0.0;
  }
  return GPT_OK;
}

}  // extern "C"
