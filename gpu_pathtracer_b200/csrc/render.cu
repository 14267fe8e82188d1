// render.cu -- wavefront path tracer: replaces init_curand_state + path_tracing_render
// (src/util/general.h:8-16, src/render/pathtracing.h:69-200,251-305, launched at
// src/main.cu:1208,1266) and the light-tree part of the build
// (compute_node_bounding_cones' energy sums, src/model/bvh/bvh_building.h:304-348).
//
// Stages per iteration, each a kernel over the queue of active pixel slots:
//   k_pt_trace  closest hit of every active path's ray against the main BVH
//   k_pt_shade  material decision, light sampling + MIS, throughput update, next ray or next
//               sample (ray-gen is folded into shade), warp-aggregated queue compaction
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#include "integrator.cuh"
#include "trace_wide.cuh"

namespace {

constexpr int PT_BLOCK = 128;
// frames of at most this many slots per rank use the streaming pipeline by default
#ifndef GPT_STREAM_AUTO_MAX_SLOTS
#define GPT_STREAM_AUTO_MAX_SLOTS 0
#endif

struct RenderCtx {
  PathState ps{};
  float4 *leaf_info = nullptr, *node_info = nullptr;
  float *node_energy = nullptr;
  int *lt_flags = nullptr;
  int light_n = 0;
  uint64_t light_epoch = 0;       // ctx->build_epoch the light-tree arrays were made for
  uint32_t *rng_init = nullptr;   // cached curand_init states for (width, height, rank set)
  int rng_w = 0, rng_h = 0, rng_rank = -1, rng_nranks = 0;
  int rng_lanes = 0, rng_conc = 0;  // rng_mode 1: lanes per pixel the cache holds / lanes per wave the pixel map holds
  size_t rng_init_cap = 0;        // words allocated for rng_init
  float4 *partial = nullptr;      // rng_mode 1: a slot's colour sum (one per slot of a wave)
  // Independent wavefront groups (slot ranges), each on its own stream with its own queues,
  // counters and read-back ring, so the tail of one group's trace overlaps the others' work.
  static constexpr int MAX_GROUPS = 8, RING = 4;
  // per group, 8 ints on the device: [0..2] queue lengths in rotation (read / append / being
  // zeroed), [3] the trace kernel's ray cursor, [4] error flag
  static constexpr int CNT = 8;
  int *h_counts = nullptr;        // pinned: MAX_GROUPS x RING x CNT ints
  cudaEvent_t ring_ev[MAX_GROUPS][RING] = {};
  cudaStream_t gstream[MAX_GROUPS] = {};
  cudaEvent_t fork_ev = nullptr, join_ev[MAX_GROUPS] = {};
  int hist_stride = 0;
  uint32_t *sort_key = nullptr, *sort_key_alt = nullptr, *sort_hist = nullptr;
  int *sort_val = nullptr, *sort_val_alt = nullptr;
  int *classq = nullptr;          // PT_NCLASS material-class lists per group (k_pt_trace -> k_pt_shade)
  int *ccount = nullptr;          // MAX_GROUPS x 3 (rotation) x PT_NCLASS list lengths
  float4 *sss_req = nullptr;      // deferred SSS lookups: 3 float4 per slot (a slot asks at most once a round)
  int *sss_req_cnt = nullptr;     // MAX_GROUPS x 2 request counters (alternating rounds)
  size_t sss_req_cap = 0;         // slots sss_req was allocated for
  int *sq_cells = nullptr, *sq_ctl = nullptr;  // streaming pipeline: 9 rings + control words
  int sq_capacity = 0;                         // cells per ring
  float *d_fb = nullptr;          // internal framebuffer for the host API
  size_t fb_capacity = 0;
  std::vector<cudaEvent_t> events;  // profiling only
};

struct RenderParams {
  CameraDev cam;
  int sample_size, level, dof_size;
  float pdf_weight;
  V3 sky;
  const float *bg_r, *bg_g, *bg_b;
  int bg_h, bg_w;
  float inv_samples;  // (float)(1.0 / sample_size / dof_sample_size)
  int mode;           // 0 = camera paths into the frame buffer, 1 = SSS first pass (paths start
                      //     at the SSS points, do_sss_first_pass pathtracing.h:202-249)
  int use_sss;        // mode 0: SSS point trees exist -> compute_color_sss at SSS hits
  // rng_mode 1 (fast_lanes > 0): a pixel's dof x spp samples are dealt to fast_lanes independent
  // streams (curand sequence pixel * fast_lanes + lane); a wave renders lanes lane0 .. of every
  // pixel, slot s = (lane - lane0) * npix + pixel slot, and leaves each slot's sum in `partial`
  int fast_lanes, lane0, npix, total_samples;
  float4 *partial;
  SssView sss;
  // r02 deferred SSS lookups (k_sss_lookup): a shade thread that hits an SSS material appends a
  // request of three float4 -- (point, path_length) (diffuse tint, slot) (object, -, -, -) --
  // instead of doing the lookup itself; nullptr = lookups stay in the shade thread
  float4 *sss_req;
  int *sss_req_count;
};

// ---- RNG init: curand_init(1984, pixel_index, 0) (general.h:15) ---------------------------
__global__ void __launch_bounds__(128) k_rng_init(uint32_t *rng, const int *pixel, int n) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  Rng r;
  curand_init(1984, pixel[s], 0, &r.s);
  rng_store_packed(r, rng, s);
}

// rng_mode 1: stream of (pixel, lane) = curand sequence pixel * lanes + lane; state index lane * npix + p
__global__ void __launch_bounds__(128)
k_rng_init_lanes(uint32_t *rng, const int *pixel, int npix, int lanes) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= (size_t)npix * lanes) return;
  const int lane = (int)(s / npix), pl = (int)(s % npix);
  Rng r;
  curand_init(1984, (unsigned long long)pixel[pl] * lanes + lane, 0, &r.s);
  rng_store_packed(r, rng, s);
}

// ---- light tree -----------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_light_leaves(SceneDev s, TexAtlas atlas, const float4 *lt_tris, int n, float4 *leaf_info) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const float4 *tp = lt_tris + 3 * (size_t)j;
  float4 t0 = tp[0], t1 = tp[1], t2 = tp[2];
  int prim = __float_as_int(t2.z);
  const float *bb = s.aabb + 6 * (size_t)prim;
  V3 bc = V3{0.5f * (bb[0] + bb[1]), 0.5f * (bb[2] + bb[3]), 0.5f * (bb[4] + bb[5])};
  V3 p1 = V3{t0.x, t0.y, t0.z}, p2 = V3{t0.w, t1.x, t1.y}, p3 = V3{t1.z, t1.w, t2.x};
  V3 cen = (p1 + p2 + p3) / 3.0f;
  TriShade ts = load_tri_shade(s.shade, prim);
  const float third = (float)(1.0 / 3);
  V3 an = shading_normal(ts, s.mats[ts.mat], atlas, third, third, third);
  const float *e = s.energy + 3 * (size_t)prim;
  float4 *o = leaf_info + 4 * (size_t)j;
  o[0] = make_float4(bc.x, bc.y, bc.z, ts.area);
  o[1] = make_float4(cen.x, cen.y, cen.z, e[0]);
  o[2] = make_float4(an.x, an.y, an.z, e[1]);
  o[3] = make_float4(e[2], 0.f, 0.f, 0.f);
}

// bottom-up energy sums: node.energy = left.energy + right.energy (bvh_building.h:338-339)
__global__ void __launch_bounds__(128)
k_light_nodes(const float4 *nodes, const int *parent_leaf, const float4 *leaf_info, int n,
              float *node_energy, float4 *node_info, int *flags) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  int pe = parent_leaf[j];
  while (pe >= 0) {
    int par = pe >> 1;
    __threadfence();
    if (atomicAdd(flags + par, 1) == 0) return;
    __threadfence();
    const float4 *q = nodes + 4 * (size_t)par;
    const int *q3 = reinterpret_cast<const int *>(q + 3);
    int cr[2] = {__ldcg(q3), __ldcg(q3 + 1)};
    V3 e[2];
    for (int c = 0; c < 2; c++) {
      if (cr[c] < 0) {
        const float4 *li = leaf_info + 4 * (size_t)(~cr[c]);
        e[c] = V3{__ldcg(&li[1].w), __ldcg(&li[2].w), __ldcg(&li[3].x)};
      } else {
        const float *ne = node_energy + 3 * (size_t)cr[c];
        e[c] = V3{__ldcg(ne), __ldcg(ne + 1), __ldcg(ne + 2)};
      }
    }
    V3 sum = e[0] + e[1];
    node_energy[3 * (size_t)par] = sum.x;
    node_energy[3 * (size_t)par + 1] = sum.y;
    node_energy[3 * (size_t)par + 2] = sum.z;
    const float *f = reinterpret_cast<const float *>(q);
    float c3[3];
    for (int k = 0; k < 3; k++) {
      float lo = fminf(__ldcg(f + 2 * k), __ldcg(f + 6 + 2 * k));
      float hi = fmaxf(__ldcg(f + 2 * k + 1), __ldcg(f + 6 + 2 * k + 1));
      c3[k] = 0.5f * (lo + hi);
    }
    node_info[par] = make_float4(c3[0], c3[1], c3[2], length(sum));
    pe = __ldcg(q3 + 2);
  }
}

}  // namespace

namespace {

// _get_sky_color (pathtracing.h:69-82)
__device__ __forceinline__ V3 sky_color(const RenderParams &p, V3 dir) {
  float u = (float)(.5 + (double)atan2f(dir.z, dir.x) / (2.0 * GPT_PI));
  float v = (float)(.5 - (double)asinf(dir.y) / GPT_PI);
  int iu = (int)floorf((u - floorf(u)) * (float)(p.bg_w - 1));
  int iv = (int)floorf((v - floorf(v)) * (float)(p.bg_h - 1));
  int idx = p.bg_w * iv + iu;
  return p.sky * V3{__ldg(p.bg_r + idx), __ldg(p.bg_g + idx), __ldg(p.bg_b + idx)};
}

__device__ __forceinline__ void start_sample(const PathState &ps, int slot, V3 o, V3 d) {
  ps.rays[2 * (size_t)slot] = make_float4(o.x, o.y, o.z, INFINITY);
  ps.rays[2 * (size_t)slot + 1] = make_float4(d.x, d.y, d.z, 0.f);
  MatList l;
#pragma unroll
  for (int k = 0; k < MATLIST_MAX; k++) l.id[k] = MAT_AIR;
  l.len = 1;  // add_new_material(material_list, length, nullptr)  (pathtracing.h:106)
  matlist_store(reinterpret_cast<uint4 *>(ps.cm + 2 * (size_t)slot + 1), l);
}

// SSS first pass: slot s = point s (tree order); its stream is rand_state_sss[s]
__global__ void __launch_bounds__(PT_BLOCK)
k_sss_start(PathState ps, RenderParams p, int n) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  Rng rng;
  rng_load(rng, p.sss.rng, s, n);
  float4 a = __ldg(p.sss.pts + 3 * (size_t)s), b = __ldg(p.sss.pts + 3 * (size_t)s + 1),
         c = __ldg(p.sss.pts + 3 * (size_t)s + 2);
  V3 o = V3{a.x, a.y, a.z}, nrm = V3{b.z, b.w, c.x};
  V3 d = generate_dir(V3{0.f, 0.f, 0.f}, nrm, 0, 1.f, rng);
  rng_store_packed(rng, ps.rng, s);
  ps.pixel[s] = s;
  ps.cam[2 * (size_t)s] = make_float4(o.x, o.y, o.z, 0.f);
  ps.cam[2 * (size_t)s + 1] = make_float4(nrm.x, nrm.y, nrm.z, 0.f);
  ps.ma[2 * (size_t)s] = make_float4(1.f, 1.f, 1.f, __int_as_float(0));
  ps.ma[2 * (size_t)s + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
  ps.cm[2 * (size_t)s] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
  start_sample(ps, s, o, d);
  ps.queue[0][s] = s;
}

// compute_color_sss + traverse_bvh_pts (pathtracing_sss.h:17-42, bvh_traversal_pts.h:14-104):
// inverse-distance weighted mean of the point colours within path_length of the hit, in the
// reference's left-first visit order (the fp32 sums are order dependent).
static __device__ __noinline__ V3 sss_lookup_ref_walk(const SssTree &t, V3 c, float r, bool &found) {
  V3 color = V3{0.f, 0.f, 0.f};
  float sum_w = 0.f;
  found = false;
  if (t.n < 2) return color;
  int stack[64];
  int sp = 0, node = 0;
  while (true) {
    const float4 *q = t.nodes + 4 * (size_t)node;
    float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
    int cref[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
    float bxn[2] = {q0.x, q1.z}, bxx[2] = {q0.y, q1.w};
    float byn[2] = {q0.z, q2.x}, byx[2] = {q0.w, q2.y};
    float bzn[2] = {q1.x, q2.z}, bzx[2] = {q1.y, q2.w};
    bool isect[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {  // BoundingBox::is_intersection(BoundingSphere) (bounding_box.h:296-312)
      float cx = 0.5f * (bxn[k] + bxx[k]), cy = 0.5f * (byn[k] + byx[k]), cz = 0.5f * (bzn[k] + bzx[k]);
      isect[k] = fabsf(cx - c.x) <= (bxx[k] - cx) + r && fabsf(cy - c.y) <= (byx[k] - cy) + r &&
                 fabsf(cz - c.z) <= (bzx[k] - cz) + r;
    }
#pragma unroll
    for (int k = 0; k < 2; k++) {
      if (isect[k] && cref[k] < 0) {
        const float4 *lf = t.leaf + 2 * (size_t)(~cref[k]);
        float4 a = __ldg(lf);
        float d = dist(c, V3{a.x, a.y, a.z});
        if (d <= r) {
          float4 col = __ldg(lf + 1);
          found = true;
          float w = 1.0f / dist(V3{a.x, a.y, a.z}, c);
          w = (float)fmin(9999.99, (double)w);
          sum_w += w;
          color = color + w * V3{col.x, col.y, col.z};
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
  if (found) {
    float k = 1.0f / sum_w;  // vec3::operator/=(float): k = 1.0 / t, then three multiplies
    color = V3{color.x * k, color.y * k, color.z * k};
  }
  return color;
}

// The same estimate, found fast.  The reference's point tree is keyed by Morton codes taken
// against the LAST point's box (sss.cu k_sss_morton), so its nodes span whole octants and a walk
// of it tests every point of every octant the lookup sphere touches (10^4..10^5 points for a
// few hundred accepted).  The accepted SET does not depend on the tree -- a point within r of c
// makes the sphere overlap every box that contains it -- only the ORDER of the fp32 sums does,
// and that order is the reference walk's visit order (a node's leaf children when the node is
// visited, nodes in left-first preorder).  So: find the set with a second LBVH over
// the same points keyed by their true Morton codes, mark each accepted point's position in
// that walk in a private bitmap, then add them up by scanning the bitmap in increasing order.  Same terms, same order, same result; ~100x fewer node visits.  The bitmap comes
// from a pool (probe + atomicCAS); with the pool exhausted the reference walk above runs.
__device__ int g_sss_debug_mode = 0;  // 0 fast, 1 reference walk, 2 skip (timing experiments)
// SssTree grid: the cell of a coordinate -- the same expression as grid_cell() in sss.cu
__device__ __forceinline__ int sss_grid_cell(float x, float o, float inv, int n) {
  float f = floorf((x - o) * inv);
  return f < 0.f ? 0 : (f >= (float)n ? n - 1 : (int)f);
}
// Candidates from the uniform grid (sss.cu build_candidate_grid): the cells the sphere's box
// touches -- at most 3 per axis, the cell edge is >= every lookup radius -- are runs of
// consecutive records along x, read with independent 16-byte loads instead of ~4000 DEPENDENT
// node fetches of the search tree.  The cell of a coordinate is the same fp32 expression as on
// the host and monotone in it, and the box is widened by more than the rounding of dist(), so
// no accepted point is missed.  (Its own function: inlined, it changed the register allocation
// of the whole shade kernel, 96 -> 72 with spills.)
static __device__ __noinline__ void sss_mark_grid(const SssTree &t, V3 c, float r, uint32_t *bm,
                                                  int &wlo_out, int &whi_out) {
  int wlo = wlo_out, whi = whi_out;
  const float g = 1e-5f * (fabsf(c.x) + fabsf(c.y) + fabsf(c.z) + r + 1.f);
  const float lo = r + g;
  // T = the largest float whose square root rounds to <= r
  float T = r * r;
  while (sqrtf(T) > r) T = __uint_as_float(__float_as_uint(T) - 1u);
  while (sqrtf(__uint_as_float(__float_as_uint(T) + 1u)) <= r) T = __uint_as_float(__float_as_uint(T) + 1u);
  int x0 = sss_grid_cell(c.x - lo, t.gox, t.ginv, t.gnx), x1 = sss_grid_cell(c.x + lo, t.gox, t.ginv, t.gnx);
  int y0 = sss_grid_cell(c.y - lo, t.goy, t.ginv, t.gny), y1 = sss_grid_cell(c.y + lo, t.goy, t.ginv, t.gny);
  int z0 = sss_grid_cell(c.z - lo, t.goz, t.ginv, t.gnz), z1 = sss_grid_cell(c.z + lo, t.goz, t.ginv, t.gnz);
  for (int z = z0; z <= z1; z++)
    for (int y = y0; y <= y1; y++) {
      const int row = t.gnx * (y + t.gny * z);
      const int jb = __ldg(t.gstart + row + x0), je = __ldg(t.gstart + row + x1 + 1);
      for (int j = jb; j < je; j++) {
        float4 a = __ldg(t.gpos + j);
        // dist(c, a) <= r without the square root: same d2 as length(c - a) forms, and sqrtf
        // (correctly rounded, monotone) is <= r exactly for d2 <= T
        const float dx = c.x - a.x, dy = c.y - a.y, dz = c.z - a.z;
        const float d2 = dx * dx + dy * dy + dz * dz;
        if (d2 <= T) {
          int i = __float_as_int(a.w);
          bm[i >> 5] |= 1u << (i & 31);
          wlo = min(wlo, i >> 5);
          whi = max(whi, i >> 5);
        }
      }
    }
  wlo_out = wlo;
  whi_out = whi;
}
struct BitmapPool {  // by value: a reference into the kernel parameters would be spilled
  uint32_t *bm_pool;
  int *bm_busy;
  int bm_n, bm_words;
};
static __device__ __noinline__ V3 sss_lookup(BitmapPool sv, const SssTree &t, V3 c, float r,
                                             bool &found) {
  found = false;
  if (t.n < 2) return V3{0.f, 0.f, 0.f};
  if (g_sss_debug_mode == 1) return sss_lookup_ref_walk(t, c, r, found);
  if (g_sss_debug_mode == 2) return V3{0.f, 0.f, 0.f};
  int slot = -1;
  {
    unsigned h = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u;
    for (int probe = 0; probe < 32 && slot < 0; probe++) {
      int i = (int)((h + (unsigned)probe * 40503u) % (unsigned)sv.bm_n);
      if (atomicCAS(sv.bm_busy + i, 0, 1) == 0) slot = i;
    }
  }
  if (slot < 0) return sss_lookup_ref_walk(t, c, r, found);
  uint32_t *bm = sv.bm_pool + (size_t)slot * sv.bm_words;
  int wlo = sv.bm_words, whi = -1;
  int stack[64];
  int sp = 0, node = 0;
  if (t.gpos) {
    sss_mark_grid(t, c, r, bm, wlo, whi);
  } else
  while (true) {
    const float4 *q = t.snodes + 4 * (size_t)node;
    float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
    int cref[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
    float bxn[2] = {q0.x, q1.z}, bxx[2] = {q0.y, q1.w};
    float byn[2] = {q0.z, q2.x}, byx[2] = {q0.w, q2.y};
    float bzn[2] = {q1.x, q2.z}, bzx[2] = {q1.y, q2.w};
    bool isect[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
      float cx = 0.5f * (bxn[k] + bxx[k]), cy = 0.5f * (byn[k] + byx[k]), cz = 0.5f * (bzn[k] + bzx[k]);
      // one ulp-scale guard band: this test only has to be conservative
      float g = 1e-5f * (fabsf(cx) + fabsf(cy) + fabsf(cz) + r + 1.f);
      isect[k] = fabsf(cx - c.x) <= (bxx[k] - cx) + r + g && fabsf(cy - c.y) <= (byx[k] - cy) + r + g &&
                 fabsf(cz - c.z) <= (bzx[k] - cz) + r + g;
    }
#pragma unroll
    for (int k = 0; k < 2; k++) {
      if (isect[k] && cref[k] < 0) {
        float4 a = __ldg(t.spos + (~cref[k]));
        float d = dist(c, V3{a.x, a.y, a.z});
        if (d <= r) {
          int i = __float_as_int(a.w);
          bm[i >> 5] |= 1u << (i & 31);
          wlo = min(wlo, i >> 5);
          whi = max(whi, i >> 5);
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
  V3 color = V3{0.f, 0.f, 0.f};
  float sum_w = 0.f;
  // (requesting the records of 8 marked points together, 16-byte bitmap loads around the L1 and
  //  fire-and-forget atomicOr marks were measured: 180 -> 218 ms on the C5 variant, so the
  //  plain form stays)
  for (int w = wlo; w <= whi; w++) {
    uint32_t bits = bm[w];
    if (!bits) continue;
    bm[w] = 0u;
    while (bits) {
      int b = __ffs(bits) - 1;
      bits &= bits - 1;
      const float4 *lf = t.vleaf + 2 * (size_t)(32 * w + b);  // visit order: no index load, neighbours share lines
      float4 a = lf[0];
      float4 col = lf[1];  // written after the first pass of this launch sequence: no __ldg
      found = true;
      float wgt = 1.0f / dist(V3{a.x, a.y, a.z}, c);
      // (float)fmin(9999.99, (double)wgt) without the fp64 round trip: (float)9999.99 is
      // 9999.990234375f > 9999.99, so every float below it is < 9999.99 and is kept, everything
      // else (NaN included, as fmin does) becomes it
      wgt = wgt < 9999.990234375f ? wgt : 9999.990234375f;
      sum_w += wgt;
      color = color + wgt * V3{col.x, col.y, col.z};
    }
  }
  __threadfence();
  atomicExch(sv.bm_busy + slot, 0);
  if (found) {
    float k = 1.0f / sum_w;
    color = V3{color.x * k, color.y * k, color.z * k};
  }
  return color;
}

// The same lookup by a WHOLE WARP (k_sss_lookup).  Why: ncu on the C5 frame (profiles/
// r02_summary.md) -- the thread-per-lookup form above runs its ordered sum at 5 of 32 lanes and
// its candidate scan at 13, every load a dependent round trip, and because the material sort
// puts all SSS hits into a few hundred warps those warps run ~8 ms while 90 % of the GPU idles.
// Here all 32 lanes call with the same (t, c, r):
//   mark:  the <= 9 candidate rows are read 32 records at a time (coalesced, independent loads),
//          accepted points set their bit with a fire-and-forget atomicOr;
//   sum:   32 bitmap words per step, one per lane; popc + warp prefix give every accepted point
//          its position in the reference's visit order; the 32 next terms (w, w r, w g, w b)
//          are computed side by side -- sqrt, IEEE divide, record fetch -- and parked in
//          shared memory, then lanes 0..3 add one component each IN ORDER.
// fp32 addition order, candidate test and weight are those of sss_lookup, so the result is
// bit-identical to it and to the reference walk (tests/test_gpu_sss.py).
static __device__ __noinline__ V3 sss_lookup_warp(BitmapPool sv, const SssTree &t, V3 c, float r,
                                                  bool &found, float4 *sterm) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  found = false;
  if (t.n < 2) return V3{0.f, 0.f, 0.f};
  int slot = -1;
  if (lane == 0) {
    unsigned h = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 2654435761u;
    for (int probe = 0; probe < 64 && slot < 0; probe++) {
      int i = (int)((h + (unsigned)probe * 40503u) % (unsigned)sv.bm_n);
      if (atomicCAS(sv.bm_busy + i, 0, 1) == 0) slot = i;
    }
  }
  slot = __shfl_sync(FULL, slot, 0);
  if (slot < 0 || !t.gpos) {  // no private bitmap / no candidate grid: one lane walks the reference's tree
    if (slot >= 0 && lane == 0) atomicExch(sv.bm_busy + slot, 0);
    V3 col = V3{0.f, 0.f, 0.f};
    bool f = false;
    if (lane == 0) col = sss_lookup_ref_walk(t, c, r, f);
    found = __shfl_sync(FULL, (int)f, 0) != 0;
    return V3{__shfl_sync(FULL, col.x, 0), __shfl_sync(FULL, col.y, 0), __shfl_sync(FULL, col.z, 0)};
  }
  uint32_t *bm = sv.bm_pool + (size_t)slot * sv.bm_words;
  // ---- mark (sss_mark_grid, 32 candidates at a time)
  int wlo = sv.bm_words, whi = -1;
  {
    const float g = 1e-5f * (fabsf(c.x) + fabsf(c.y) + fabsf(c.z) + r + 1.f);
    const float lo = r + g;
    float T = r * r;  // the largest float whose square root rounds to <= r
    while (sqrtf(T) > r) T = __uint_as_float(__float_as_uint(T) - 1u);
    while (sqrtf(__uint_as_float(__float_as_uint(T) + 1u)) <= r) T = __uint_as_float(__float_as_uint(T) + 1u);
    const int x0 = sss_grid_cell(c.x - lo, t.gox, t.ginv, t.gnx), x1 = sss_grid_cell(c.x + lo, t.gox, t.ginv, t.gnx);
    const int y0 = sss_grid_cell(c.y - lo, t.goy, t.ginv, t.gny), y1 = sss_grid_cell(c.y + lo, t.goy, t.ginv, t.gny);
    const int z0 = sss_grid_cell(c.z - lo, t.goz, t.ginv, t.gnz), z1 = sss_grid_cell(c.z + lo, t.goz, t.ginv, t.gnz);
    for (int z = z0; z <= z1; z++)
      for (int y = y0; y <= y1; y++) {
        const int row = t.gnx * (y + t.gny * z);
        const int jb = __ldg(t.gstart + row + x0), je = __ldg(t.gstart + row + x1 + 1);
#pragma unroll 4
        for (int j = jb + lane; j < je; j += 32) {
          const float4 a = __ldg(t.gpos + j);
          const float dx = c.x - a.x, dy = c.y - a.y, dz = c.z - a.z;
          const float d2 = dx * dx + dy * dy + dz * dz;
          if (d2 <= T) {
            const int i = __float_as_int(a.w);
            atomicOr(bm + (i >> 5), 1u << (i & 31));
            wlo = min(wlo, i >> 5);
            whi = max(whi, i >> 5);
          }
        }
      }
    wlo = __reduce_min_sync(FULL, wlo);
    whi = __reduce_max_sync(FULL, whi);
  }
  __threadfence();  // the marks, before any lane reads a word another lane marked
  __syncwarp();
  // ---- ordered sum: lane 0 owns sum_w, lanes 1..3 colour r, g, b
  // Software pipeline over "units" of <= 32 consecutive terms (ncu, first version: the stage
  // is latency bound -- long scoreboard 8 per issue -- by two dependent round trips per step,
  // bitmap word then point records): the next step's bitmap word is fetched one step ahead,
  // and a unit's record loads are issued BEFORE the four adding lanes run through the previous
  // unit's terms, so the adds cover the fetch.  Units are summed in order: same association.
  float accv = 0.f;
  int prev_nb = 0, buf = 0;   // terms parked in sterm[32 * (buf ^ 1) ..] that are not added yet
  auto add_parked = [&](int which, int nb) {
    if (lane < 4) {
      const float *f = reinterpret_cast<const float *>(sterm + 32 * which) + lane;
#pragma unroll 8
      for (int i = 0; i < nb; i++) accv = accv + f[4 * i];
    }
  };
  uint32_t bits_next = (wlo + lane <= whi) ? __ldcg(bm + wlo + lane) : 0u;
  for (int w0 = wlo; w0 <= whi; w0 += 32) {
    const int w = w0 + lane;
    const uint32_t bits = bits_next;
    bits_next = (w + 32 <= whi) ? __ldcg(bm + w + 32) : 0u;
    if (bits) bm[w] = 0u;  // leave the bitmap clean for its next user
    const int cnt = __popc(bits);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(FULL, incl, o);
      if (lane >= o) incl += y;
    }
    const int total = __shfl_sync(FULL, incl, 31), excl = incl - cnt;
    for (int base = 0; base < total; base += 32) {
      const int k = base + lane;
      int src = 0;  // the lane whose word holds the k-th accepted point of this step
#pragma unroll
      for (int step = 16; step >= 1; step >>= 1) {
        const int e = __shfl_sync(FULL, excl, (src + step) & 31);
        if (src + step < 32 && e <= k) src += step;
      }
      const uint32_t wb = __shfl_sync(FULL, bits, src);
      const int e0 = __shfl_sync(FULL, excl, src);
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), col = a;
      if (k < total) {  // issue this unit's record loads ...
        const int bbit = (int)__fns(wb, 0, k - e0 + 1);
        const float4 *lf = t.vleaf + 2 * (size_t)(32 * (w0 + src) + bbit);
        a = lf[0];
        col = lf[1];
      }
      add_parked(buf ^ 1, prev_nb);  // ... and add the previous unit's terms while they fly
      float4 term = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k < total) {
        float wgt = 1.0f / dist(V3{a.x, a.y, a.z}, c);
        wgt = wgt < 9999.990234375f ? wgt : 9999.990234375f;  // (float)fmin(9999.99, wgt), see sss_lookup
        term = make_float4(wgt, wgt * col.x, wgt * col.y, wgt * col.z);
      }
      sterm[32 * buf + lane] = term;
      __syncwarp();
      prev_nb = min(32, total - base);
      buf ^= 1;
      found = true;
    }
  }
  add_parked(buf ^ 1, prev_nb);
  __syncwarp();
  __threadfence();  // the cleared words, before the bitmap changes hands
  __syncwarp();
  if (lane == 0) atomicExch(sv.bm_busy + slot, 0);
  const float sum_w = __shfl_sync(FULL, accv, 0);
  V3 color = V3{__shfl_sync(FULL, accv, 1), __shfl_sync(FULL, accv, 2), __shfl_sync(FULL, accv, 3)};
  if (found) {
    const float k = 1.0f / sum_w;
    color = V3{color.x * k, color.y * k, color.z * k};
  }
  return color;
}

__global__ void __launch_bounds__(PT_BLOCK)
k_pt_start(PathState ps, RenderParams p, int n, float *fb) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  int pix = ps.pixel[s];
  int i = pix / p.cam.width, j = pix % p.cam.width;
  Rng rng;
  rng_load_packed(rng, ps.rng, s);
  V3 o, d;
  camera_ray(p.cam, i + .5f, j + .5f, rng, o, d);
  rng_store_packed(rng, ps.rng, s);
  ps.cam[2 * (size_t)s] = make_float4(o.x, o.y, o.z, 0.f);
  ps.cam[2 * (size_t)s + 1] = make_float4(d.x, d.y, d.z, 0.f);
  ps.ma[2 * (size_t)s] = make_float4(1.f, 1.f, 1.f, __int_as_float(0));
  ps.ma[2 * (size_t)s + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
  ps.cm[2 * (size_t)s] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
  start_sample(ps, s, o, d);
  ps.queue[0][s] = s;
  if (p.fast_lanes == 0) { fb[3 * (size_t)pix] = 0.f; fb[3 * (size_t)pix + 1] = 0.f; fb[3 * (size_t)pix + 2] = 0.f; }
}

// rng_mode 1, after a wave: fb[pixel] (+)= the sums of the wave's lanes in lane order; the last
// wave scales by 1 / samples.  The frame is the same whatever the number of lanes per wave.
__global__ void __launch_bounds__(PT_BLOCK)
k_pt_combine(const int *__restrict__ pixel, const float4 *__restrict__ partial, int npix, int lanes,
             int first, int last, float inv_samples, float *fb) {
  int pl = blockIdx.x * blockDim.x + threadIdx.x;
  if (pl >= npix) return;
  const size_t o = 3 * (size_t)pixel[pl];
  V3 acc = first ? V3{0.f, 0.f, 0.f} : V3{fb[o], fb[o + 1], fb[o + 2]};
  for (int c = 0; c < lanes; c++) {
    const float4 v = partial[(size_t)c * npix + pl];
    acc = acc + V3{v.x, v.y, v.z};
  }
  if (last) acc = acc * inv_samples;
  fb[o] = acc.x; fb[o + 1] = acc.y; fb[o + 2] = acc.z;
}

// trace stage: persistent warps pull active slots from the queue (trace_persistent.cuh)
//
// What the shade stage wants is its slots grouped by the code they will run (the material at the
// hit): that keeps its warps converged and its instruction cache warm.  Round 1 got there with a
// stable radix sort between the two stages -- three more launches in each of the spp x level
// dependent rounds of a frame (9 % of the GPU time at 1080p, and ~15 us of pure latency per
// round on a tile-sharded frame).  Now the trace kernel itself appends every
// finished slot to the list of its material class (PT_NCLASS lists per group, one warp-aggregated
// atomicAdd per class and warp) and the shade kernel reads the lists back to back: the same
// grouping, no sort, two launches per round.  A pixel's result does not depend on the order
// its slot is shaded in, so the frame is bit-identical (test).
constexpr int PT_NCLASS = 16;  // 15 material classes (material id mod 15) + misses
struct QueueSource {
  PathState ps;
  const int *queue;
  const int *qcount;
  const float4 *shade;   // per-triangle shading records (material id in s6.y)
  uint32_t *sort_key;    // sort mode -- out: shade-coherence key of queue entry i (material, 127 = miss)
  int *sort_val;         //              out: its slot
  int loc_bits, loc_shift;  // minor key: the hit leaf's top loc_bits bits (Morton order = place)
  int *classq;           // class mode (sort_key == nullptr): PT_NCLASS lists of cstride slots
  int cstride;
  int *ccount;           // their lengths
  __device__ int count() const { return *qcount; }
  __device__ void load(int i, float4 &o, float4 &d) const {
    int s = queue[i];
    o = ps.rays[2 * (size_t)s];
    d = ps.rays[2 * (size_t)s + 1];
  }
  __device__ int material_of(const TraceLane &l) const {
    return l.leaf >= 0 ? (__float_as_int(__ldg(&shade[7 * (size_t)l.prim + 6].y)) & 127) : 127;
  }
  __device__ int write_hit(int i, const TraceLane &l) const {
    int s = queue[i];
    ps.hits[2 * (size_t)s] = make_float4(l.t, __int_as_float(l.leaf), __int_as_float(l.prim), 0.f);
    ps.hits[2 * (size_t)s + 1] = make_float4(l.b1, l.b2, l.b3, 0.f);
    return s;
  }
  __device__ void store(int i, const TraceLane &l) const {  // one lane on its own
    const int s = write_hit(i, l);
    const int m = material_of(l);
    if (sort_key) {
      // Key for the material sort that precedes shade: paths that will run the same BSDF code
      // become neighbours ... and, within a material, paths that hit the same part of the
      // scene: their textures / shading records share cache lines, and the rays they spawn
      // start close together, which is what the NEXT trace wants.
      uint32_t key = (uint32_t)m << loc_bits;
      if (l.leaf >= 0) key |= (uint32_t)l.leaf >> loc_shift;
      sort_key[i] = key;
      sort_val[i] = s;
    } else {
      const int c = m == 127 ? PT_NCLASS - 1 : m % (PT_NCLASS - 1);
      classq[(size_t)c * cstride + atomicAdd(ccount + c, 1)] = s;
    }
  }
  // called by all 32 lanes; the lanes with fin set hand their ray's record over
  __device__ void store_warp(bool fin, int i, const TraceLane &l) const {
    if (sort_key) {
      if (fin) store(i, l);
      return;
    }
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int c = -1, s = 0;
    if (fin) {
      s = write_hit(i, l);
      const int m = material_of(l);
      c = m == 127 ? PT_NCLASS - 1 : m % (PT_NCLASS - 1);
    }
    const unsigned peers = __match_any_sync(FULL, c);
    if (fin) {
      const int leader = __ffs(peers) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(ccount + c, __popc(peers));
      base = __shfl_sync(peers, base, leader);
      classq[(size_t)c * cstride + base + __popc(peers & ((1u << lane) - 1u))] = s;
    }
  }
};

template <bool COUNT, bool COOP, bool DEFER = false>
__global__ void __launch_bounds__(TP_BLOCK, TP_MINBLOCKS)
k_pt_trace(QueueSource src, const float4 *__restrict__ nodes, const float4 *__restrict__ tris,
           const int *__restrict__ parent_leaf, int *next_ray, unsigned long long *counters) {
  __shared__ int stack[TP_STACK_SMEM * TP_BLOCK];
  __shared__ int coop[COOP ? (TP_BLOCK / 32) * COOP_WARP_WORDS : 1];
  __shared__ int defer[DEFER ? 2 * DEFER_SLOTS * TP_BLOCK : 1];
  trace_persistent<COUNT, QueueSource, COOP, DEFER>(nodes, tris, parent_leaf, src, next_ray,
                                                    stack, counters, coop, defer);
}

template <bool COUNT>
__global__ void __launch_bounds__(TP_BLOCK, W4_MINBLOCKS)
k_pt_trace_wide(QueueSource src, WideTree tree, int *next_ray, unsigned long long *counters) {
  __shared__ int smem[W4_SMEM_WORDS];
  trace_wide<COUNT, QueueSource>(tree, src, next_ray, smem, counters);
}

// One path vertex: the slot's hit is shaded and the slot either gets its next ray (returns
// true: it goes back to the trace stage) or its pixel is finished (false).  ASYNC = the
// streaming pipeline: a slot's state is written by whichever SM had it last while THIS kernel
// runs, so it is read around the (incoherent) L1.
#define PS_LD(ptr) (ASYNC ? __ldcg(ptr) : *(ptr))
// A path just ended with `result` (its radiance; pathtracing.h:251-305 adds it to the pixel and
// goes on with the pixel's next sample): accumulate, then either start the next sample of the
// slot (returns true: the slot goes back to the trace stage) or write the pixel / partial sum /
// SSS point colour.  Its own function since r02: k_sss_lookup ends the paths whose radiance is
// a deferred point-cloud lookup.
template <bool ASYNC>
__device__ __forceinline__ bool finish_sample(const PathState &ps, const RenderParams &p, int s,
                                              V3 result, Rng &rng, float *fb, int sidx) {
  bool alive = true;
  float4 cl = PS_LD(ps.cm + 2 * (size_t)s);
  V3 color = V3{cl.x, cl.y, cl.z} + de_nan(result);
  int k = __float_as_int(cl.w);
  sidx++;
  V3 o, d;
  bool finished = false;
  if (p.mode == 1) {
    // do_sss_first_pass: every sample draws a fresh cosine-hemisphere ray at the point
    if (sidx >= p.sample_size) {
      finished = true;
    } else {
      float4 co = PS_LD(ps.cam + 2 * (size_t)s), cn = PS_LD(ps.cam + 2 * (size_t)s + 1);
      o = V3{co.x, co.y, co.z};
      d = generate_dir(V3{0.f, 0.f, 0.f}, V3{cn.x, cn.y, cn.z}, 0, 1.f, rng);
    }
  } else if (p.fast_lanes) {
    // this slot's lane gets samples lane, lane + lanes, ... of the pixel's dof x spp (dof index =
    // sample / spp); as in the reference a new camera ray is drawn when the dof index changes
    // and reused otherwise, so one lane reproduces the reference's stream exactly
    const int lane = p.lane0 + s / p.npix;
    if (sidx >= (p.total_samples - lane + p.fast_lanes - 1) / p.fast_lanes) {
      finished = true;
    } else if ((lane + sidx * p.fast_lanes) / p.sample_size != (lane + (sidx - 1) * p.fast_lanes) / p.sample_size) {
      int pix = ps.pixel[s];
      camera_ray(p.cam, pix / p.cam.width + .5f, pix % p.cam.width + .5f, rng, o, d);
      ps.cam[2 * (size_t)s] = make_float4(o.x, o.y, o.z, 0.f);
      ps.cam[2 * (size_t)s + 1] = make_float4(d.x, d.y, d.z, 0.f);
    } else {
      float4 co = PS_LD(ps.cam + 2 * (size_t)s), cd = PS_LD(ps.cam + 2 * (size_t)s + 1);
      o = V3{co.x, co.y, co.z};
      d = V3{cd.x, cd.y, cd.z};
    }
  } else if (sidx >= p.sample_size) {
    sidx = 0;
    k++;
    if (k >= p.dof_size) {
      finished = true;
    } else {
      int pix = ps.pixel[s];
      camera_ray(p.cam, pix / p.cam.width + .5f, pix % p.cam.width + .5f, rng, o, d);
      ps.cam[2 * (size_t)s] = make_float4(o.x, o.y, o.z, 0.f);
      ps.cam[2 * (size_t)s + 1] = make_float4(d.x, d.y, d.z, 0.f);
    }
  } else {
    float4 co = PS_LD(ps.cam + 2 * (size_t)s), cd = PS_LD(ps.cam + 2 * (size_t)s + 1);
    o = V3{co.x, co.y, co.z};
    d = V3{cd.x, cd.y, cd.z};
  }
  if (finished && p.mode == 1) {
    // color *= (1.0 / sample_size); color *= filter; point -> assign_color(color)
    float4 pa = __ldg(p.sss.pts + 3 * (size_t)s), pb = __ldg(p.sss.pts + 3 * (size_t)s + 1);
    V3 out = (color * p.inv_samples) * V3{pa.w, pb.x, pb.y};
    p.sss.leaf[2 * (size_t)s + 1] = make_float4(out.x, out.y, out.z, 0.f);
    alive = false;
  } else if (finished && p.fast_lanes) {
    p.partial[s] = make_float4(color.x, color.y, color.z, 0.f);
    alive = false;
  } else if (finished) {
    int pix = ps.pixel[s];
    V3 out = color * p.inv_samples;
    fb[3 * (size_t)pix] = out.x; fb[3 * (size_t)pix + 1] = out.y; fb[3 * (size_t)pix + 2] = out.z;
    alive = false;
  } else {
    ps.cm[2 * (size_t)s] = make_float4(color.x, color.y, color.z, __int_as_float(k));
    ps.ma[2 * (size_t)s] = make_float4(1.f, 1.f, 1.f, __int_as_float(0));
    ps.ma[2 * (size_t)s + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(sidx));
    start_sample(ps, s, o, d);
  }
  return alive;
}

template <bool ASYNC>
__device__ __forceinline__ bool shade_slot(const PathState &ps, const RenderParams &p, int s,
                                           const SceneDev &sc, const TexAtlas &atlas,
                                           const float4 *__restrict__ main_tris,
                                           const LightTree &lt, float *fb, int *error_flag) {
  bool alive = true;
  float4 rdv = PS_LD(ps.rays + 2 * (size_t)s + 1);
  float4 h0 = PS_LD(ps.hits + 2 * (size_t)s), h1 = PS_LD(ps.hits + 2 * (size_t)s + 1);
  float4 mk = PS_LD(ps.ma + 2 * (size_t)s), ac = PS_LD(ps.ma + 2 * (size_t)s + 1);
  V3 dir = V3{rdv.x, rdv.y, rdv.z};
  V3 mask = V3{mk.x, mk.y, mk.z}, acc = V3{ac.x, ac.y, ac.z};
  int bounce = __float_as_int(mk.w), sidx = __float_as_int(ac.w);
  int leaf = __float_as_int(h0.y);
  Rng rng;
  rng_load_packed<ASYNC>(rng, ps.rng, s);
  bool done = false, deferred = false;
  V3 result = V3{0.f, 0.f, 0.f};
  if (leaf >= 0) {
    int prim = __float_as_int(h0.z);
    float b1 = h1.x, b2 = h1.y, b3 = h1.z;
    const float4 *tp = main_tris + 3 * (size_t)leaf;
    float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
    TriShade ts = load_tri_shade(sc.shade, prim);
    const MaterialDev &m = sc.mats[ts.mat];
    // hit_record (triangle.h:399-405)
    V3 point = b1 * V3{t0.x, t0.y, t0.z} + b2 * V3{t0.w, t1.x, t1.y} + b3 * V3{t1.z, t1.w, t2.x};
    float u = b1 * ts.u1 + b2 * ts.u2 + b3 * ts.u3;
    float v = b1 * ts.v1 + b2 * ts.v2 + b3 * ts.v3;
    V3 normal = shading_normal(ts, m, atlas, b1, b2, b3);
    MatList list;
    matlist_unpack(list, PS_LD(reinterpret_cast<const uint4 *>(ps.cm + 2 * (size_t)s + 1)));
    RefRec ref;
    ref.reflected = false; ref.refracted = false; ref.diffuse = false;
    ref.n = 1.f; ref.k = V3{0.f, 0.f, 0.f}; ref.perfect = V3{1.f, 0.f, 0.f}; ref.bent = false;
    bool sss = false;
    check_next_path<false>(sc.mats, atlas, ts.mat, dir, normal, u, v, list, ref, sss, rng, false);
    float factor = 1.f;
    const bool sss_hit = sss && p.use_sss;  // `sss && !sss_first_pass` (pathtracing.h:129-142)
    if (!ref.false_hit && !sss_hit)
      factor = change_ref_ray(lt, sc.shade, sc.mats, atlas, point, normal, dir, ref,
                              p.pdf_weight, rng);
    if (sss_hit) {  // return compute_color_sss(...): the sample IS the point-cloud estimate
      if (!ASYNC && p.sss_req) {  // deferred: a whole warp of k_sss_lookup does it
        const V3 tint = mat_diffuse(atlas, m, u, v);
        const int r = atomicAdd(p.sss_req_count, 1);
        p.sss_req[3 * (size_t)r] = make_float4(point.x, point.y, point.z, m.path_length);
        p.sss_req[3 * (size_t)r + 1] = make_float4(tint.x, tint.y, tint.z, __int_as_float(s));
        p.sss_req[3 * (size_t)r + 2] = make_float4(__int_as_float(ts.obj), 0.f, 0.f, 0.f);
        deferred = true;
      } else {
        bool found;
        V3 col = sss_lookup(BitmapPool{p.sss.bm_pool, p.sss.bm_busy, p.sss.bm_n, p.sss.bm_words},
                            p.sss.trees[ts.obj], point, m.path_length, found);
        result = found ? col * mat_diffuse(atlas, m, u, v) : V3{0.f, 0.f, 0.f};
      }
      done = true;
    }
    // material list bookkeeping (pathtracing.h:144-162)
    bool ok = true;
    if ((ref.false_hit || ref.refracted) && ref.entering) ok = matlist_add(list, ts.mat);
    else if ((ref.false_hit || ref.refracted) && !ref.entering) matlist_remove(list, ts.mat);
    if (!ok && !sss_hit) *error_flag = 1;
    matlist_store(reinterpret_cast<uint4 *>(ps.cm + 2 * (size_t)s + 1), list);
    if (!ref.false_hit && !sss_hit) {
      V3 add = de_nan(mask * mat_emission(atlas, m, u, v));
      acc = acc + add;
      if (factor > 0.f) {
        mask = mask * (ref.filter * clamp_lohi(0.f, .9999f, factor));
      } else {
        done = true;
        result = acc;
      }
    }
    if (!done) {
      bounce++;
      if (bounce >= p.level) {
        done = true;
        result = acc;
      } else {
        ps.rays[2 * (size_t)s] = make_float4(point.x, point.y, point.z, INFINITY);
        ps.rays[2 * (size_t)s + 1] = make_float4(ref.dir.x, ref.dir.y, ref.dir.z, 0.f);
        ps.ma[2 * (size_t)s] = make_float4(mask.x, mask.y, mask.z, __int_as_float(bounce));
        ps.ma[2 * (size_t)s + 1] = make_float4(acc.x, acc.y, acc.z, __int_as_float(sidx));
      }
    }
  } else {
    V3 sky = sky_color(p, dir);
    result = bounce < 1 ? sky : acc + mask * sky;
    done = true;
  }
  if (done && !deferred) alive = finish_sample<ASYNC>(ps, p, s, result, rng, fb, sidx);
  if (deferred) alive = false;  // k_sss_lookup finishes the sample and re-queues the slot
  rng_store_packed(rng, ps.rng, s);
  return alive;
}
#undef PS_LD

__global__ void __launch_bounds__(PT_BLOCK, 5)  // 5 blocks per SM = the 96-register form (ptxas picks 72 + spills or 114 on its own, both slower)
k_pt_shade(PathState ps, RenderParams p, const int *__restrict__ queue,
           const int *__restrict__ count, int *next_queue, int *next_count, SceneDev sc,
           TexAtlas atlas, const float4 *__restrict__ main_tris, LightTree lt, float *fb,
           int *error_flag, unsigned long long *ray_counter, int *zero_a, int *zero_b,
           const int *__restrict__ classq, int cstride, const int *__restrict__ ccount,
           int *czero) {
  int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a == 0) {
    atomicAdd(ray_counter, (unsigned long long)*count);  // rays traced this iteration
    // nobody reads these two until the next iteration of this group: the queue length of the
    // iteration after next, and the trace kernel's ray cursor (saves two memsets per iteration)
    *zero_a = 0;
    *zero_b = 0;
  }
  // class lengths of the round after next (read last by the previous round's shade)
  if (classq && a < PT_NCLASS) czero[a] = 0;
  bool alive = false;
  int s = -1;
  if (a < *count) {
    if (classq) {  // entry a of the PT_NCLASS class lists laid end to end
      int c = 0, before = 0;
#pragma unroll
      for (int k = 0; k < PT_NCLASS - 1; k++) {
        const int len = __ldg(ccount + k);
        if (c == k && a >= before + len) { c = k + 1; before += len; }
      }
      s = classq[(size_t)c * cstride + (a - before)];
    } else {
      s = queue[a];
    }
    alive = shade_slot<false>(ps, p, s, sc, atlas, main_tris, lt, fb, error_flag);
  }
  // warp-aggregated compaction of the surviving slots into the next queue
  unsigned ballot = __ballot_sync(0xffffffffu, alive);
  if (ballot) {
    int lane = threadIdx.x & 31;
    int leader = __ffs(ballot) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(next_count, __popc(ballot));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (alive) next_queue[base + __popc(ballot & ((1u << lane) - 1u))] = s;
  }
}


// ---- deferred SSS lookups (r02) -------------------------------------------------------------------
// compute_color_sss (pathtracing_sss.h:17-42) as a stage of its own: one WARP per request of the
// round's request list (sss_lookup_warp), then lane 0 ends the sample exactly as the shade thread
// would have -- result = found ? colour * tint : 0, finish_sample -- and puts the slot back into
// the next round's queue.  Runs after k_pt_shade of the round on the group's stream.  The two
// request counters of a group alternate between rounds; this launch zeroes the other one.
constexpr int LK_BLOCK = 128;
__global__ void __launch_bounds__(LK_BLOCK)
k_sss_lookup(PathState ps, RenderParams p, const int *__restrict__ req_count, int *req_zero,
             int *next_queue, int *next_count, float *fb) {
  __shared__ float4 sterm[LK_BLOCK / 32][64];  // two units of 32 terms per warp (double buffer)
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp = (blockIdx.x * LK_BLOCK + threadIdx.x) >> 5, nwarps = (gridDim.x * LK_BLOCK) >> 5;
  if (blockIdx.x == 0 && threadIdx.x == 0) *req_zero = 0;
  const int n = *req_count;
  const BitmapPool pool{p.sss.bm_pool, p.sss.bm_busy, p.sss.bm_n, p.sss.bm_words};
  for (int r = warp; r < n; r += nwarps) {
    const float4 q0 = p.sss_req[3 * (size_t)r], q1 = p.sss_req[3 * (size_t)r + 1], q2 = p.sss_req[3 * (size_t)r + 2];
    const int s = __float_as_int(q1.w), obj = __float_as_int(q2.x);
    bool found;
    const V3 col = sss_lookup_warp(pool, p.sss.trees[obj], V3{q0.x, q0.y, q0.z}, q0.w, found, sterm[wib]);
    if (lane == 0) {
      const V3 result = found ? col * V3{q1.x, q1.y, q1.z} : V3{0.f, 0.f, 0.f};
      Rng rng;
      rng_load_packed<false>(rng, ps.rng, s);
      const int sidx = __float_as_int(ps.ma[2 * (size_t)s + 1].w);
      const bool alive = finish_sample<false>(ps, p, s, result, rng, fb, sidx);
      rng_store_packed(rng, ps.rng, s);
      if (alive) next_queue[atomicAdd(next_count, 1)] = s;
    }
    __syncwarp();
  }
}

// ---- streaming pipeline (barrier-free wavefront) ------------------------------------------------
// The iteration loop above makes a frame a chain of dof x spp x level DEPENDENT rounds of
// launches, and each round lasts as long as its slowest ray + the latency of a sort + of one
// shade wave, however few rays it carries: fine for a whole frame on one GPU (groups overlap the
// rounds), but it is what bounds a tile-sharded frame (1/8 of 1080p: ~250 us a round, SMs 40 %
// idle).  Here the two stages are two PERSISTENT kernels running side by side, and a slot moves
// between them through rings in HBM the moment it is ready:
//   ray ring      shade -> trace   slots whose next ray is written
//   hit rings[8]  trace -> shade   slots whose hit record is written, one ring per material
//                                  class (ring 7 = miss), so a shade warp pops 32 slots that run
//                                  the same BSDF code -- the coherence sort, without the sort
// A pixel's chain is now the sum of ITS OWN ray and shade latencies, not of every round's
// maxima.  Pixels never interact and a pixel's samples stay sequential on its own RNG stream,
// so the frame is bit-identical to the iteration loop's (test).
// Ring protocol: a cell holds a slot index or -1.  Producers take a position with atomicAdd on
// the ring's tail, make their state writes visible (__threadfence) and store the slot into
// the cell; the consumer that owns the position spins until the cell is >= 0, empties it, and
// reads the slot's state around L1 (ld.cg).  Ray-ring positions are handed out unconditionally
// (a trace lane holds at most one and polls it between traversal steps); hit-ring positions
// only below the tail (CAS), because a shade warp must not commit to a material that may
// never come.  Rings hold >= 2 n cells for n slots, and a producer waits for its cell to be
// empty, so a wrap-around can never overwrite a slot.
constexpr int SQ_RINGS = 8;    // material classes (7 = miss)
constexpr int SQ_SUB = 16;     // sub-rings per class: producers and consumers of different SMs use
                               // different counters (ONE head per class was measured: 1184 shade
                               // warps serialised on it, 70 Mrays/s)
constexpr int SQ_NR = SQ_RINGS * SQ_SUB;
enum { SQ_RAY_HEAD = 0, SQ_RAY_TAIL = 8, SQ_DONE = 16, SQ_ABORT = 24, SQ_ERR = 32,
       SQ_RING0 = 40, SQ_WORDS = 40 + 24 * SQ_NR };
// per hit ring r: head (next position to hand to a consumer), tail (next position to hand to a
// producer), credits (published cells not yet promised to a consumer), one 32-byte sector each
__device__ __forceinline__ int sq_head(int r) { return SQ_RING0 + 24 * r; }
__device__ __forceinline__ int sq_tail(int r) { return SQ_RING0 + 24 * r + 8; }
__device__ __forceinline__ int sq_credits(int r) { return SQ_RING0 + 24 * r + 16; }
struct StreamQ {
  int *ray_ring;   // cells
  int *hit_ring;   // SQ_NR x hit_cells
  int *ctl;        // SQ_WORDS ints
  int mask;        // cells - 1
  int hit_mask;    // hit_cells - 1
  int n;           // slots = pixels to finish
};
__device__ __forceinline__ int ld_volatile(const int *p) { return *reinterpret_cast<const volatile int *>(p); }
__device__ __forceinline__ void st_volatile(int *p, int v) { *reinterpret_cast<volatile int *>(p) = v; }
// store slot s at an already reserved position of a ring (the caller fenced its state writes)
__device__ __forceinline__ void ring_put(int *cells, int mask, int pos, int s, int *ctl) {
  int *c = cells + (pos & mask);
  int spins = 0;
  while (ld_volatile(c) >= 0) {  // a full lap behind an unread cell: wait, never overwrite
    __nanosleep(100);
    if (++spins > (1 << 24)) { atomicExch(ctl + SQ_ABORT, 1); break; }
  }
  st_volatile(c, s);
}

struct StreamSource {
  PathState ps;
  StreamQ q;
  const float4 *shade;
  __device__ int reserve(int cnt) const { return atomicAdd(q.ctl + SQ_RAY_HEAD, cnt); }
  __device__ int poll(int pos) const {
    int *c = q.ray_ring + (pos & q.mask);
    int v = ld_volatile(c);
    if (v >= 0) {
      st_volatile(c, -1);
      __threadfence();
    }
    return v;
  }
  __device__ void load(int s, float4 &o, float4 &d) const {
    o = __ldcg(ps.rays + 2 * (size_t)s);
    d = __ldcg(ps.rays + 2 * (size_t)s + 1);
  }
  // a finished ray: the record goes out at once (fire and forget); the slot itself is handed to
  // the shade kernel later, by the whole warp (publish), and the return value says to which ring
  __device__ int store(int s, const TraceLane &l) const {
    __stcg(ps.hits + 2 * (size_t)s, make_float4(l.t, __int_as_float(l.leaf), __int_as_float(l.prim), 0.f));
    __stcg(ps.hits + 2 * (size_t)s + 1, make_float4(l.b1, l.b2, l.b3, 0.f));
    int m = SQ_RINGS - 1;  // miss
    if (l.leaf >= 0) m = (__float_as_int(__ldg(&shade[7 * (size_t)l.prim + 6].y)) & 127) % (SQ_RINGS - 1);
    return m * SQ_SUB + (int)(blockIdx.x % SQ_SUB);
  }
  // called by all 32 lanes; lanes with slot >= 0 hand it to ring r
  __device__ void publish(int slot, int r) const {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    __threadfence();  // hit records before the cells
    const unsigned has = __ballot_sync(FULL, slot >= 0);
    if (slot >= 0) {
      const unsigned peers = __match_any_sync(has, r);
      const int leader = __ffs(peers) - 1, cnt = __popc(peers);
      int base = 0;
      if (lane == leader) base = atomicAdd(q.ctl + sq_tail(r), cnt);
      base = __shfl_sync(peers, base, leader);
      ring_put(q.hit_ring + (size_t)r * (q.hit_mask + 1), q.hit_mask,
               base + __popc(peers & ((1u << lane) - 1u)), slot, q.ctl);
      __threadfence();  // cells before the credits
      __syncwarp(peers);
      if (lane == leader) atomicAdd(q.ctl + sq_credits(r), cnt);
    }
  }
  __device__ bool finished() const {
    return ld_volatile(q.ctl + SQ_DONE) >= q.n || ld_volatile(q.ctl + SQ_ABORT) != 0;
  }
  __device__ void abort() const { atomicExch(q.ctl + SQ_ABORT, 1); }
};

__global__ void __launch_bounds__(256) k_stream_init(StreamQ q) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t cells = (size_t)q.mask + 1, hit_total = (size_t)SQ_NR * ((size_t)q.hit_mask + 1);
  if (i < cells) q.ray_ring[i] = i < (size_t)q.n ? (int)i : -1;
  if (i < hit_total) q.hit_ring[i] = -1;
  if (i < SQ_WORDS) q.ctl[i] = i == SQ_RAY_TAIL ? q.n : 0;
}

template <bool COUNT>
__global__ void __launch_bounds__(TP_BLOCK, TP_MINBLOCKS)
k_pt_trace_stream(StreamSource src, const float4 *__restrict__ nodes,
                  const float4 *__restrict__ tris, const int *__restrict__ parent_leaf,
                  unsigned long long *counters) {
  __shared__ int stack[TP_STACK_SMEM * TP_BLOCK];
  trace_stream<COUNT, StreamSource>(nodes, tris, parent_leaf, src, stack, counters);
}

__global__ void __launch_bounds__(PT_BLOCK)
k_pt_shade_stream(PathState ps, RenderParams p, StreamQ q, SceneDev sc, TexAtlas atlas,
                  const float4 *__restrict__ main_tris, LightTree lt, float *fb,
                  unsigned long long *ray_counter) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const size_t hit_cells = (size_t)q.hit_mask + 1;
  // the four sub-rings this warp looks at first; lanes = 4 sub-rings x 8 classes
  int home = ((blockIdx.x * (PT_BLOCK / 32) + (threadIdx.x >> 5)) * 4) % SQ_SUB;
  unsigned shaded = 0;
  int misses = 0;
  long long idle_since = 0;
  while (true) {
    // ---- pick a ring: a full warp's worth wins, else the longest ------------------------------
    const int r_of_lane = (lane & 7) * SQ_SUB + (home + (lane >> 3)) % SQ_SUB;
    const int credits = ld_volatile(q.ctl + sq_credits(r_of_lane));
    const unsigned full = __ballot_sync(FULL, credits >= 32);
    int src_lane;
    if (full) {
      src_lane = __ffs(full) - 1;
    } else {
      int best = credits, bl = lane;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        int ob = __shfl_xor_sync(FULL, best, d), ol = __shfl_xor_sync(FULL, bl, d);
        if (ob > best || (ob == best && ol < bl)) { best = ob; bl = ol; }
      }
      src_lane = bl;
    }
    const int r = __shfl_sync(FULL, r_of_lane, src_lane);
    int take = __shfl_sync(FULL, credits, src_lane);
    take = take > 32 ? 32 : take;
    int h = -1;
    if (take > 0 && lane == 0) {
      // semaphore try-wait: credits are cells that ARE published, so positions promised against
      // them exist (head never passes tail); an overdraft is put back
      if (atomicAdd(q.ctl + sq_credits(r), -take) >= take) h = atomicAdd(q.ctl + sq_head(r), take);
      else atomicAdd(q.ctl + sq_credits(r), take);
    }
    h = __shfl_sync(FULL, h, 0);
    if (h < 0) {
      home = (home + 4) % SQ_SUB;
      if (++misses < SQ_SUB / 4) continue;  // look at the other sub-rings first
      misses = 0;
      if (ld_volatile(q.ctl + SQ_DONE) >= q.n || ld_volatile(q.ctl + SQ_ABORT) != 0) break;
      const long long now = clock64();
      if (idle_since == 0) idle_since = now;
      if (now - idle_since > TS_WATCHDOG_CYCLES) {
        atomicExch(q.ctl + SQ_ABORT, 1);
        break;
      }
      __nanosleep(100);
      continue;
    }
    idle_since = 0;
    misses = 0;
    // ---- the slots at positions h .. h + take - 1 ------------------------------------------------
    int s = -1;
    if (lane < take) {
      int *c = q.hit_ring + (size_t)r * hit_cells + ((h + lane) & q.hit_mask);
      int spins = 0;
      while ((s = ld_volatile(c)) < 0) {  // published out of order: its producer is a few instructions away
        __nanosleep(20);
        if (++spins > (1 << 24)) { atomicExch(q.ctl + SQ_ABORT, 1); break; }
      }
      if (s >= 0) st_volatile(c, -1);
    }
    __threadfence();
    bool alive = false;
    if (s >= 0) {
      alive = shade_slot<true>(ps, p, s, sc, atlas, main_tris, lt, fb, q.ctl + SQ_ERR);
      shaded++;
    }
    // ---- next ray -> ray ring; finished pixels -> the counter both kernels watch ----------------
    __threadfence();
    const unsigned am = __ballot_sync(FULL, alive), fm = __ballot_sync(FULL, s >= 0 && !alive);
    if (am) {
      const int leader = __ffs(am) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(q.ctl + SQ_RAY_TAIL, __popc(am));
      base = __shfl_sync(FULL, base, leader);
      if (alive) ring_put(q.ray_ring, q.mask, base + __popc(am & ((1u << lane) - 1u)), s, q.ctl);
    }
    if (fm && lane == __ffs(fm) - 1) atomicAdd(q.ctl + SQ_DONE, __popc(fm));
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) shaded += __shfl_xor_sync(FULL, shaded, d);
  if (lane == 0 && shaded) atomicAdd(ray_counter, (unsigned long long)shaded);
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------
static RenderCtx *rc_of(gpt_ctx *ctx) {
  if (!ctx->render_state) ctx->render_state = new RenderCtx();
  return static_cast<RenderCtx *>(ctx->render_state);
}

// Light-tree side arrays (leaf records, node centres + |E|, node energies) over the light BVH;
// shared by the path tracer and the SPPM passes (sppm.cu).
int gpt_build_light_tree_arrays(gpt_ctx *ctx, float4 **leaf_info, float4 **node_info,
                                float **node_energy, int **flags) {
  BvhDev &lb = ctx->bvh[GPT_BVH_LIGHT];
  int n = lb.n;
  if (n < 2)
    return gpt_fail(ctx, GPT_E_INVALID, "path tracing needs >= 2 emissive triangles (src/main.cu:1124-1125)", __FILE__, __LINE__);
  GPT_TRY(gpt_alloc(ctx, leaf_info, 4 * (size_t)n));
  GPT_TRY(gpt_alloc(ctx, node_info, (size_t)n));
  GPT_TRY(gpt_alloc(ctx, node_energy, 3 * (size_t)n));
  GPT_TRY(gpt_alloc(ctx, flags, (size_t)n));
  GPT_CUDA(ctx, cudaMemsetAsync(*flags, 0, sizeof(int) * (size_t)n, ctx->stream));
  TexAtlas atlas{ctx->scene.tex_r, ctx->scene.tex_g, ctx->scene.tex_b};
  int grid = (n + 127) / 128;
  k_light_leaves<<<grid, 128, 0, ctx->stream>>>(ctx->scene, atlas, lb.tris, n, *leaf_info);
  GPT_LAUNCH_CHECK(ctx);
  k_light_nodes<<<grid, 128, 0, ctx->stream>>>(lb.nodes, lb.parent_leaf, *leaf_info, n,
                                               *node_energy, *node_info, *flags);
  GPT_LAUNCH_CHECK(ctx);
  return GPT_OK;
}

static int build_light_tree(gpt_ctx *ctx, RenderCtx *rc) {
  GPT_TRY(gpt_build_light_tree_arrays(ctx, &rc->leaf_info, &rc->node_info, &rc->node_energy,
                                      &rc->lt_flags));
  rc->light_n = ctx->bvh[GPT_BVH_LIGHT].n;
  rc->light_epoch = ctx->build_epoch;
  return GPT_OK;
}

static int reserve_paths(gpt_ctx *ctx, RenderCtx *rc, int n) {
  PathState &ps = rc->ps;
  if (n <= ps.capacity) return GPT_OK;
  size_t N = (size_t)n;
  GPT_TRY(gpt_alloc(ctx, &ps.pixel, N));
  GPT_TRY(gpt_alloc(ctx, &ps.rng, 8 * N));
  GPT_TRY(gpt_alloc(ctx, &rc->rng_init, 8 * N));
  rc->rng_init_cap = 8 * N;
  GPT_TRY(gpt_alloc(ctx, &rc->partial, N));
  GPT_TRY(gpt_alloc(ctx, &ps.rays, 2 * N));
  GPT_TRY(gpt_alloc(ctx, &ps.hits, 2 * N));
  GPT_TRY(gpt_alloc(ctx, &ps.ma, 2 * N));
  GPT_TRY(gpt_alloc(ctx, &ps.cm, 2 * N));
  GPT_TRY(gpt_alloc(ctx, &ps.cam, 2 * N));
  GPT_TRY(gpt_alloc(ctx, &ps.queue[0], N));
  GPT_TRY(gpt_alloc(ctx, &ps.queue[1], N));
  GPT_TRY(gpt_alloc(ctx, &ps.counts, (size_t)RenderCtx::CNT * RenderCtx::MAX_GROUPS));
  GPT_TRY(gpt_alloc(ctx, &rc->sort_key, N));
  GPT_TRY(gpt_alloc(ctx, &rc->sort_key_alt, N));
  GPT_TRY(gpt_alloc(ctx, &rc->sort_val, N));
  GPT_TRY(gpt_alloc(ctx, &rc->sort_val_alt, N));
  GPT_TRY(gpt_alloc(ctx, &rc->classq, (size_t)PT_NCLASS * N));
  GPT_TRY(gpt_alloc(ctx, &rc->ccount, (size_t)RenderCtx::MAX_GROUPS * 3 * PT_NCLASS));
  rc->hist_stride = gpt_radix_hist_entries(n);
  GPT_TRY(gpt_alloc(ctx, &rc->sort_hist, (size_t)rc->hist_stride * RenderCtx::MAX_GROUPS));
  if (!rc->h_counts) {
    GPT_CUDA(ctx, cudaMallocHost((void **)&rc->h_counts,
                                 RenderCtx::MAX_GROUPS * RenderCtx::RING * RenderCtx::CNT * sizeof(int)));
    GPT_CUDA(ctx, cudaEventCreateWithFlags(&rc->fork_ev, cudaEventDisableTiming));
    for (int g = 0; g < RenderCtx::MAX_GROUPS; g++) {
      GPT_CUDA(ctx, cudaStreamCreateWithFlags(&rc->gstream[g], cudaStreamNonBlocking));
      GPT_CUDA(ctx, cudaEventCreateWithFlags(&rc->join_ev[g], cudaEventDisableTiming));
      for (int k = 0; k < RenderCtx::RING; k++)
        GPT_CUDA(ctx, cudaEventCreateWithFlags(&rc->ring_ev[g][k], cudaEventDisableTiming));
    }
  }
  ps.capacity = n;
  rc->rng_w = rc->rng_h = 0;  // cached states are gone
  return GPT_OK;
}

// slot -> pixel map: 32x32 tiles dealt round-robin to ranks, tile-major so that one warp
// covers one 32-pixel tile row (coherent primary rays)
static std::vector<int> tile_pixels(int W, int H, int rank, int nranks) {
  std::vector<int> px;
  int tx = (W + 31) / 32, ty = (H + 31) / 32;
  px.reserve((size_t)W * H / nranks + 1024);
  for (int t = 0; t < tx * ty; t++) {
    if (t % nranks != rank) continue;
    int x0 = (t % tx) * 32, y0 = (t / tx) * 32;
    for (int y = y0; y < y0 + 32 && y < H; y++)
      for (int x = x0; x < x0 + 32 && x < W; x++) px.push_back(y * W + x);
  }
  return px;
}

// rng_mode 1 ("fast", gpt.h GptParams.rng_mode).  The reference gives a pixel ONE XORWOW stream and
// runs its dof x spp samples one after the other on it (pathtracing.h:277-293), so a frame is a
// chain of dof x spp x level dependent rounds however few pixels a GPU renders: that chain, not
// throughput, is what an 8-way tile-sharded frame is made of.  Here the samples of
// a pixel are dealt to `lanes` independent streams (sequence pixel * lanes + lane) that run side
// by side, `conc` of them at a time: the chain is lanes times shorter, and a rank with an eighth
// of the pixels keeps as many paths in flight as a whole frame does.  The image is a function of
// the scene and of `lanes` only -- not of conc, the rank count or the run -- because lane sums are
// added in lane order; with lanes = 1 it IS the reference-mode image.  It is another Monte-Carlo
// estimate of the same integral, not the reference's bit pattern: compared statistically
// (tests/test_gpu_render.py::test_fast_rng_mode).
struct FastWave {
  int lanes, lane0, conc;
  bool first, last;
};
constexpr long long GPT_FAST_TARGET_SLOTS = 2200000;  // paths in flight a wave aims for (~ a 1080p frame)

static int render_device(gpt_ctx *ctx, float *d_fb, int rank, int nranks, bool sss_pass = false,
                         const FastWave *fw = nullptr) {
  if (!ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  if (!ctx->have_cam) return gpt_fail(ctx, GPT_E_INVALID, "gpt_set_camera first", __FILE__, __LINE__);
  if (nranks < 1 || rank < 0 || rank >= nranks) return GPT_E_INVALID;
  const GptParams &gp = ctx->params;
  if (gp.pathtracing_sample_size < 1 || gp.pathtracing_level < 1 || gp.dof_sample_size < 1)
    return gpt_fail(ctx, GPT_E_INVALID, "sample_size, level and dof_sample_size must be >= 1", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!sss_pass && !fw && gp.rng_mode == 1) {
    const long long total = (long long)gp.pathtracing_sample_size * gp.dof_sample_size;
    int lanes = ctx->fast_lanes > 0 ? ctx->fast_lanes : 8;
    if (lanes > total) lanes = (int)total;
    long long npix = 0;
    {
      const int W = ctx->cam.width, H = ctx->cam.height, tx = (W + 31) / 32, ty = (H + 31) / 32;
      for (int t = rank; t < tx * ty; t += nranks)
        npix += (long long)std::min(32, W - (t % tx) * 32) * std::min(32, H - (t / tx) * 32);
    }
    if (npix == 0) return GPT_OK;
    long long target = GPT_FAST_TARGET_SLOTS;
    if (const char *e = getenv("GPT_FAST_TARGET_SLOTS")) target = atoll(e);  // tests: force several waves
    int conc = (int)std::max(1LL, std::min((long long)lanes, target / npix));
    GptStats sum{};
    for (int lane0 = 0; lane0 < lanes; lane0 += conc) {
      FastWave w{lanes, lane0, conc, lane0 == 0, lane0 + conc >= lanes};
      GPT_TRY(render_device(ctx, d_fb, rank, nranks, false, &w));
      const GptStats &st = ctx->stats;
      sum.render_ms += st.render_ms; sum.rays_traced += st.rays_traced; sum.shade_events += st.shade_events;
      sum.kernel_launches += st.kernel_launches; sum.trace_launches += st.trace_launches;
      sum.trace_kernel_ms += st.trace_kernel_ms; sum.shade_kernel_ms += st.shade_kernel_ms;
      sum.internal_visits += st.internal_visits; sum.leaf_tests += st.leaf_tests;
    }
    ctx->stats.render_ms = sum.render_ms; ctx->stats.rays_traced = sum.rays_traced;
    ctx->stats.shade_events = sum.shade_events; ctx->stats.kernel_launches = sum.kernel_launches;
    ctx->stats.trace_launches = sum.trace_launches; ctx->stats.trace_kernel_ms = sum.trace_kernel_ms;
    ctx->stats.shade_kernel_ms = sum.shade_kernel_ms; ctx->stats.internal_visits = sum.internal_visits;
    ctx->stats.leaf_tests = sum.leaf_tests;
    return GPT_OK;
  }
  RenderCtx *rc = rc_of(ctx);
  SssView sv{};
  bool sss_done = true;
  const bool have_sss = gpt_sss_view(ctx, &sv, &sss_done);
  if (sss_pass && !have_sss) return GPT_OK;
  if (!sss_pass && have_sss && !sss_done) {
    // "Doing first pass for SSS objects" (src/main.cu:1213-1232), once per build
    GPT_TRY(render_device(ctx, d_fb, 0, 1, true));
    gpt_sss_mark_first_pass_done(ctx);
  }
  ctx->launches = 0;
  int W = ctx->cam.width, H = ctx->cam.height;
  bool new_map = !(rc->rng_w == W && rc->rng_h == H && rc->rng_rank == rank &&
                   rc->rng_nranks == nranks && rc->rng_lanes == (fw ? fw->lanes : 0) &&
                   rc->rng_conc == (fw ? fw->conc : 0));
  std::vector<int> px;
  int n, npix = 0;
  if (sss_pass) {
    n = sv.npoints;
    int cap_before = rc->ps.capacity;
    GPT_TRY(reserve_paths(ctx, rc, n));
    if (rc->ps.capacity != cap_before) rc->light_n = 0;
    new_map = false;
    rc->rng_w = rc->rng_h = 0;  // the slot -> pixel map is overwritten by this pass
  } else if (new_map) {
    px = tile_pixels(W, H, rank, nranks);
    npix = n = (int)px.size();
    if (n == 0) return GPT_OK;  // more ranks than tiles: nothing to render here
    if (fw) {  // the pixel list once per concurrent lane: slot (c, p) = c * npix + p
      if ((long long)npix * fw->conc > 0x7fffffffLL) return gpt_fail(ctx, GPT_E_INVALID, "too many slots", __FILE__, __LINE__);
      px.resize((size_t)npix * fw->conc);
      for (int c = 1; c < fw->conc; c++) std::copy(px.begin(), px.begin() + npix, px.begin() + (size_t)c * npix);
      n = npix * fw->conc;
    }
    GPT_TRY(reserve_paths(ctx, rc, n));
    GPT_CUDA(ctx, cudaMemcpyAsync(rc->ps.pixel, px.data(), sizeof(int) * (size_t)n,
                                  cudaMemcpyHostToDevice, ctx->stream));
    // init_curand_state (src/main.cu:1208): once per image layout, cached
    if (fw) {
      const size_t words = 8 * (size_t)npix * fw->lanes;
      if (words > rc->rng_init_cap) {
        GPT_TRY(gpt_alloc(ctx, &rc->rng_init, words));
        rc->rng_init_cap = words;
      }
      const size_t states = (size_t)npix * fw->lanes;
      k_rng_init_lanes<<<(unsigned)((states + 127) / 128), 128, 0, ctx->stream>>>(rc->rng_init, rc->ps.pixel, npix, fw->lanes);
    } else {
      k_rng_init<<<(n + 127) / 128, 128, 0, ctx->stream>>>(rc->rng_init, rc->ps.pixel, n);
    }
    GPT_LAUNCH_CHECK(ctx);
    GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rc->rng_w = W; rc->rng_h = H; rc->rng_rank = rank; rc->rng_nranks = nranks;
    rc->rng_lanes = fw ? fw->lanes : 0; rc->rng_conc = fw ? fw->conc : 0;
    rc->light_n = 0;
  } else {
    n = 0;
    int tx = (W + 31) / 32, ty = (H + 31) / 32;
    for (int t = rank; t < tx * ty; t += nranks) {
      int x0 = (t % tx) * 32, y0 = (t / tx) * 32;
      n += (std::min(32, W - x0)) * (std::min(32, H - y0));
    }
    npix = n;
  }
  if (fw) n = npix * std::min(fw->conc, fw->lanes - fw->lane0);  // slots of this wave
  if (n == 0) return GPT_OK;
  PathState ps = rc->ps;
  ps.capacity = n;  // stride of the RNG planes for this render
  // (a rebuilt scene with the same number of lights must not be sampled with the old records)
  if (rc->light_n != ctx->bvh[GPT_BVH_LIGHT].n || rc->light_epoch != ctx->build_epoch || new_map)
    GPT_TRY(build_light_tree(ctx, rc));

  RenderParams p;
  p.cam = derive_camera(ctx->cam);
  p.sample_size = gp.pathtracing_sample_size;
  p.level = gp.pathtracing_level;
  p.dof_size = gp.dof_sample_size;
  p.pdf_weight = gp.hittable_pdf_weight;
  p.sky = V3{gp.sky_emission[0], gp.sky_emission[1], gp.sky_emission[2]};
  p.bg_r = ctx->scene.bg_r; p.bg_g = ctx->scene.bg_g; p.bg_b = ctx->scene.bg_b;
  p.bg_h = ctx->scene.bg_h; p.bg_w = ctx->scene.bg_w;
  p.inv_samples = sss_pass ? (float)(1.0 / gp.pathtracing_sample_size)
                           : (float)(1.0 / gp.pathtracing_sample_size / gp.dof_sample_size);
  p.mode = sss_pass ? 1 : 0;
  p.fast_lanes = fw ? fw->lanes : 0;
  p.lane0 = fw ? fw->lane0 : 0;
  p.npix = npix;
  p.total_samples = gp.pathtracing_sample_size * gp.dof_sample_size;
  p.partial = rc->partial;
  p.use_sss = (!sss_pass && have_sss) ? 1 : 0;
  p.sss = sv;
  p.sss_req = nullptr;
  p.sss_req_count = nullptr;

  BvhDev &mb = ctx->bvh[GPT_BVH_MAIN], &lb = ctx->bvh[GPT_BVH_LIGHT];
  LightTree lt{lb.nodes, lb.tris, lb.parent_leaf, rc->leaf_info, rc->node_info, lb.n};
  TexAtlas atlas{ctx->scene.tex_r, ctx->scene.tex_g, ctx->scene.tex_b};
  cudaStream_t st = ctx->stream;

  static int lookup_grid = 0;  // persistent grid of k_sss_lookup: every resident warp takes requests in turn
  if (!lookup_grid) {
    int per_sm = 0, sms = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_sss_lookup, LK_BLOCK, 0);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    lookup_grid = (sms > 0 ? sms : 148) * (per_sm > 0 ? per_sm : 4);
  }
  bool defer_sss = false;  // SSS lookups as a stage of their own (k_sss_lookup), rounds pipeline only
  {
    const char *e = getenv("GPT_SSS_DEBUG_MODE");
    int mode = e ? atoi(e) : 0;
    GPT_CUDA(ctx, cudaMemcpyToSymbolAsync(g_sss_debug_mode, &mode, sizeof(int), 0, cudaMemcpyHostToDevice, st));
    // GPT_SSS_DEFER=0: the round-1 form, every lookup inside its shade thread (A/B, and what the
    // debug modes 1 = reference walk / 2 = skip apply to)
    const char *d = getenv("GPT_SSS_DEFER");  // read per render, like the debug mode: tests flip it
    const int env_defer = d ? atoi(d) : 1;
    defer_sss = p.use_sss && mode == 0 && env_defer != 0;
    if (defer_sss) {
      if (rc->sss_req_cap < (size_t)n) {
        gpt_free(ctx, rc->sss_req);
        rc->sss_req = nullptr;
        GPT_TRY(gpt_alloc(ctx, &rc->sss_req, 3 * (size_t)n));
        rc->sss_req_cap = (size_t)n;
      }
      if (!rc->sss_req_cnt) GPT_TRY(gpt_alloc(ctx, &rc->sss_req_cnt, (size_t)RenderCtx::MAX_GROUPS * 2));
      GPT_CUDA(ctx, cudaMemsetAsync(rc->sss_req_cnt, 0, sizeof(int) * RenderCtx::MAX_GROUPS * 2, st));
    }
  }
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
  if (!sss_pass)
    GPT_CUDA(ctx, cudaMemcpyAsync(ps.rng, rc->rng_init + (fw ? 8 * (size_t)npix * fw->lane0 : 0),
                                  sizeof(uint32_t) * 8 * (size_t)n, cudaMemcpyDeviceToDevice, st));
  if (sss_pass)
    k_sss_start<<<(n + PT_BLOCK - 1) / PT_BLOCK, PT_BLOCK, 0, st>>>(ps, p, n);
  else
    k_pt_start<<<(n + PT_BLOCK - 1) / PT_BLOCK, PT_BLOCK, 0, st>>>(ps, p, n, d_fb);
  GPT_LAUNCH_CHECK(ctx);
  const bool coop = gpt_trace_coop(ctx, mb);  // cooperative tail (trace_persistent.cuh)
  const bool defer_leaves = gpt_trace_defer(ctx);  // deferred leaf tests (trace_persistent.cuh)
  const bool wide_walk = gpt_trace_wide(ctx, mb);  // 4-wide tree (trace_wide.cuh)
  static int wide_min_rays = -1;
  if (wide_min_rays < 0) { const char *e = getenv("GPT_TRACE_WIDE_MIN_RAYS"); wide_min_rays = e ? atoi(e) : 190000; }
  // persistent grid = SMs x resident blocks of the kernel that runs: [0] binary walk, [1] wide walk
  static int trace_grid_of[2] = {0, 0};
  if (!trace_grid_of[0]) {
    int sms = 148, per_sm[2] = {4, 4};
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    // (the variant with the tail: 72 registers, 7 blocks per SM instead of 8 -- the grid size
    //  is flat between 4 and 8, profiles/r01d_summary.md)
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm[0], k_pt_trace<false, true>, TP_BLOCK, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm[1], k_pt_trace_wide<false>, TP_BLOCK, 0);
    // GPT_TRACE_BLOCKS_PER_SM < occupancy leaves room for another group's shade / sort blocks to
    // run beside a persistent trace kernel instead of after it (experiment knob)
    const char *e = getenv("GPT_TRACE_BLOCKS_PER_SM");
    int cap = e ? atoi(e) : 0;
    for (int k = 0; k < 2; k++) {
      if (cap > 0 && cap < per_sm[k]) per_sm[k] = cap;
      trace_grid_of[k] = sms * (per_sm[k] > 0 ? per_sm[k] : 1);
    }
  }
  uint64_t rays = 0, shades = 0, trace_launches = 0;
  const bool timing = ctx->prof_time_stages, counting = ctx->prof_count_visits;
  size_t nev = 0;
  // [0..2] visit counters (profiling), [3] traversal-stack overflow flag: a stale flag of an
  // earlier call must not fail this one, and an overflow in this one must not go unnoticed
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 5, 0, sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 7, 0, sizeof(unsigned long long), st));
  // GPT_TIMELINE=path (debug): the same three stamps per round, but on each group's own stream
  // with the groups left as they are, written as "group round t0 t1 t2" in ms since the frame
  // began -- where a round's 0.2 ms go on a 1/8-frame share
  static const char *timeline_path = getenv("GPT_TIMELINE");
  const bool timeline = timeline_path != nullptr && !timing;
  std::vector<int> stamp_group;
  cudaStream_t stamp_stream = st;
  int stamp_g = 0;
  auto stamp = [&]() -> cudaError_t {
    if (!timing && !timeline) return cudaSuccess;
    if (nev == rc->events.size()) {
      cudaEvent_t e;
      cudaError_t r = cudaEventCreate(&e);
      if (r != cudaSuccess) return r;
      rc->events.push_back(e);
    }
    if (timeline) stamp_group.push_back(stamp_g);
    return cudaEventRecord(rc->events[nev++], timeline ? stamp_stream : st);
  };
  // ---- which pipeline ------------------------------------------------------------------------
  int pipeline = ctx->pipeline;
  if (pipeline < 0) {
    static int env_pipe = -2;
    if (env_pipe == -2) {
      const char *e = getenv("GPT_PIPELINE");
      env_pipe = !e ? -1 : (!strcmp(e, "stream") || !strcmp(e, "1")) ? 1 : 0;
    }
    pipeline = env_pipe >= 0 ? env_pipe : (n <= GPT_STREAM_AUTO_MAX_SLOTS ? 1 : 0);
  }
  if (sss_pass || timing || fw) pipeline = 0;  // the SSS first pass, stage timing and rng_mode 1 use the rounds
  if (pipeline == 1) {
    // ---- streaming pipeline: two persistent kernels side by side (see k_pt_shade_stream) ------
    int cells = 1024;
    while (cells < 2 * n) cells <<= 1;
    const int hit_cells = cells / 4;  // per sub-ring: >= n / 2 slots of ONE class from ONE sixteenth of the trace blocks
    if (cells > rc->sq_capacity) {
      GPT_TRY(gpt_alloc(ctx, &rc->sq_cells, (size_t)cells + (size_t)SQ_NR * hit_cells));
      GPT_TRY(gpt_alloc(ctx, &rc->sq_ctl, (size_t)SQ_WORDS));
      rc->sq_capacity = cells;
    }
    StreamQ q{rc->sq_cells, rc->sq_cells + cells, rc->sq_ctl, cells - 1, hit_cells - 1, n};
    k_stream_init<<<(unsigned)(((size_t)SQ_NR * hit_cells + 255) / 256), 256, 0, st>>>(q);
    GPT_LAUNCH_CHECK(ctx);
    unsigned long long *ray_counter = ctx->d_counters + 6;
    GPT_CUDA(ctx, cudaMemsetAsync(ray_counter, 0, sizeof(unsigned long long), st));
    static int bt = 0, bs = 0, sms = 148;
    if (!bt) {
      // CUDA loads a kernel's code lazily at its first launch, and that load synchronises with
      // the kernels already running: the second kernel of the pair would wait for the first,
      // which waits for it.  Load all of them before the first pair goes out.
      cudaFuncAttributes fa;
      GPT_CUDA(ctx, cudaFuncGetAttributes(&fa, k_pt_shade_stream));
      GPT_CUDA(ctx, cudaFuncGetAttributes(&fa, k_pt_trace_stream<false>));
      GPT_CUDA(ctx, cudaFuncGetAttributes(&fa, k_pt_trace_stream<true>));
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
      const char *e1 = getenv("GPT_STREAM_TRACE_BLOCKS"), *e2 = getenv("GPT_STREAM_SHADE_BLOCKS");
      bs = e2 ? atoi(e2) : 2;
      bt = e1 ? atoi(e1) : 3;
      if (bt < 1) bt = 1;
      if (bs < 1) bs = 1;
      // Both kernels must be able to share an SM, and an SM's L1 / shared-memory split can only
      // change while the SM is empty: a shade block (no shared memory) that configures an SM
      // for "all L1" would lock the trace blocks (16 KB of traversal stacks each) out of it for
      // as long as it runs -- for ever (seen: the watchdog fired as soon as shade covered every
      // SM).  So all three ask for the same split: the smallest one that holds bt trace blocks.
      static const int cfg_kb[] = {8, 16, 32, 64, 100, 132, 164, 196, 228};
      int need_kb = bt * (TP_STACK_SMEM * TP_BLOCK * 4 / 1024 + 1), pick = 228;
      for (int c : cfg_kb)
        if (c >= need_kb) { pick = c; break; }
      const int carve = (pick * 100 + 227) / 228;
      GPT_CUDA(ctx, cudaFuncSetAttribute(k_pt_shade_stream, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
      GPT_CUDA(ctx, cudaFuncSetAttribute(k_pt_trace_stream<false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
      GPT_CUDA(ctx, cudaFuncSetAttribute(k_pt_trace_stream<true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    }
    const int need = (n + TP_BLOCK - 1) / TP_BLOCK;
    const int tgrid = std::min(sms * bt, need), sgrid = std::min(sms * bs, need);
    GPT_CUDA(ctx, cudaEventRecord(rc->fork_ev, st));
    GPT_CUDA(ctx, cudaStreamWaitEvent(rc->gstream[0], rc->fork_ev, 0));
    GPT_CUDA(ctx, cudaStreamWaitEvent(rc->gstream[1], rc->fork_ev, 0));
    StreamSource src{ps, q, ctx->scene.shade};
    // shade first: it has the larger local-memory frame, and a launch that makes the driver
    // grow the local-memory pool waits for every running kernel -- behind a persistent trace
    // kernel that itself waits for shade, that is a deadlock (seen: the watchdog fired)
    k_pt_shade_stream<<<sgrid, PT_BLOCK, 0, rc->gstream[1]>>>(ps, p, q, ctx->scene, atlas, mb.tris, lt, d_fb, ray_counter);
    GPT_LAUNCH_CHECK(ctx);
    if (counting)
      k_pt_trace_stream<true><<<tgrid, TP_BLOCK, 0, rc->gstream[0]>>>(src, mb.nodes, mb.tris, mb.parent_leaf, ctx->d_counters);
    else
      k_pt_trace_stream<false><<<tgrid, TP_BLOCK, 0, rc->gstream[0]>>>(src, mb.nodes, mb.tris, mb.parent_leaf, ctx->d_counters);
    GPT_LAUNCH_CHECK(ctx);
    for (int g = 0; g < 2; g++) {
      GPT_CUDA(ctx, cudaEventRecord(rc->join_ev[g], rc->gstream[g]));
      GPT_CUDA(ctx, cudaStreamWaitEvent(st, rc->join_ev[g], 0));
    }
    int h_ctl[SQ_WORDS];
    unsigned long long h_rays = 0;
    GPT_CUDA(ctx, cudaMemcpyAsync(h_ctl, q.ctl, sizeof(h_ctl), cudaMemcpyDeviceToHost, st));
    GPT_CUDA(ctx, cudaMemcpyAsync(&h_rays, ray_counter, sizeof(h_rays), cudaMemcpyDeviceToHost, st));
    GPT_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
    GPT_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    float ms = 0.f;
    GPT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->stats.render_ms = ms;
    ctx->stats.rays_traced = h_rays;
    ctx->stats.shade_events = h_rays;
    ctx->stats.kernel_launches = ctx->launches;
    ctx->stats.trace_launches = 1;
    if (counting) {
      unsigned long long c[8];
      GPT_CUDA(ctx, cudaMemcpy(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
      ctx->stats.internal_visits = c[0];
      ctx->stats.leaf_tests = c[1];
      ctx->stats.wide_node_visits = c[5];
      ctx->stats.wide_leaf_tests = c[7];
    }
    if (h_ctl[SQ_ERR])
      return gpt_fail(ctx, GPT_E_UNSUPPORTED, "nested-dielectric list deeper than 12", __FILE__, __LINE__);
    if (h_ctl[SQ_ABORT] || h_ctl[SQ_DONE] != n) {
      static char msg[512];
      int o = snprintf(msg, sizeof(msg), "streaming pipeline stalled: abort %d, done %d of %d, ray ring %d/%d, hit rings",
                       h_ctl[SQ_ABORT], h_ctl[SQ_DONE], n, h_ctl[SQ_RAY_HEAD], h_ctl[SQ_RAY_TAIL]);
      for (int m = 0; m < SQ_RINGS && o < (int)sizeof(msg) - 48; m++) {
        long long hd = 0, tl = 0, cr = 0;
        for (int k = 0; k < SQ_SUB; k++) {
          int r = m * SQ_SUB + k;
          hd += h_ctl[SQ_RING0 + 24 * r]; tl += h_ctl[SQ_RING0 + 24 * r + 8]; cr += h_ctl[SQ_RING0 + 24 * r + 16];
        }
        o += snprintf(msg + o, sizeof(msg) - o, " %lld/%lld(%lld)", hd, tl, cr);
      }
      snprintf(msg + o, sizeof(msg) - o, " (%.1f ms)", ms);
      return gpt_fail(ctx, GPT_E_INVALID, msg, __FILE__, __LINE__);
    }
    return GPT_OK;
  }
  // ---- wavefront groups -------------------------------------------------------------------
  // Slots are cut into G contiguous ranges that run their trace -> sort -> shade iterations
  // independently on their own streams (a pixel never interacts with another pixel).  Inside
  // a group the host runs LAG iterations ahead of the device: the queue length it sizes the
  // grids with is the exact length from LAG iterations earlier -- an upper bound, because
  // slots only ever leave a queue -- read back through a ring of pinned slots, so no
  // iteration waits for a device->host round trip.  Kernels take the exact length from device
  // memory.  Stage timing (profiling) needs one stream, so it forces G = 1.
  constexpr int RING = RenderCtx::RING;
  static int LAG = 0;  // rounds the host runs ahead of the device (1 .. RING - 1), env GPT_HOST_LAG
  static int host_prof = 0;
  if (!LAG) {
    const char *e = getenv("GPT_HOST_LAG");
    LAG = e ? atoi(e) : 2;
    if (LAG < 1) LAG = 1;
    if (LAG > RING - 1) LAG = RING - 1;
    host_prof = getenv("GPT_HOST_PROFILE") != nullptr;
  }
  double host_wait_s = 0.0;
  const auto host_t0 = std::chrono::steady_clock::now();
  int G = 1;
  {
    static int env_groups = -1;
    if (env_groups < 0) {
      const char *e = getenv("GPT_WAVEFRONT_GROUPS");
      env_groups = e ? atoi(e) : 0;
    }
    G = ctx->wavefront_groups > 0 ? ctx->wavefront_groups
        : env_groups > 0        ? env_groups
                                : (n >= (1 << 16) ? 4 : 1);
    if (G > RenderCtx::MAX_GROUPS) G = RenderCtx::MAX_GROUPS;
    if (timing || G < 1) G = 1;
  }
  // spatial minor key of the coherence sort: top bits of the hit leaf's Morton-sorted position
  int loc_bits = 0, loc_shift = 31;  // measured on C2: trace -7 % at 17 bits, but the extra sort passes cost more (default off)
  {
    static int env_loc = -1;
    if (env_loc < 0) {
      const char *e = getenv("GPT_SORT_LOC_BITS");
      env_loc = e ? atoi(e) : 0;
      if (env_loc < 0 || env_loc > 17) env_loc = 0;
    }
    loc_bits = env_loc;
    int leaf_bits = 1;
    while ((1 << leaf_bits) < mb.n) leaf_bits++;
    if (loc_bits > leaf_bits) loc_bits = leaf_bits;
    loc_shift = leaf_bits - loc_bits;
    if (loc_bits == 0) loc_shift = 31;
  }
  constexpr int CNT = RenderCtx::CNT;
  struct Group {
    int s0, n, active, cur;
    int c;  // which of counts[0..2] holds the length of the queue being traced
    int cur_of[RenderCtx::RING];
    long long it;
    bool done;
  };
  Group grp[RenderCtx::MAX_GROUPS];
  int h_init[RenderCtx::CNT * RenderCtx::MAX_GROUPS] = {0};
  for (int g = 0; g < G; g++) {
    // group boundaries on multiples of 1024 slots (= one 32x32 tile) so warps stay tile rows
    long long lo = (long long)n * g / G, hi = (long long)n * (g + 1) / G;
    if (g > 0) lo = (lo + 1023) / 1024 * 1024;
    if (g < G - 1) hi = (hi + 1023) / 1024 * 1024;
    if (lo > n) lo = n;
    if (hi > n) hi = n;
    grp[g] = Group{(int)lo, (int)(hi - lo), (int)(hi - lo), 0, 0, {0, 0, 0, 0}, 0, hi <= lo};
    h_init[CNT * g] = grp[g].n;
  }
  GPT_CUDA(ctx, cudaMemcpyAsync(ps.counts, h_init, sizeof(int) * CNT * G, cudaMemcpyHostToDevice, st));
  // how the shade stage gets its slots grouped by material: class lists written by the trace
  // kernel (default) or the round-1 radix sort between the stages (GPT_SHADE_SORT=1, A/B)
  // Measured on rank 0's share of C2 (tools/shard_time.py, r02): a full 1080p frame is faster with
  // the sort (the appends make the trace warps wait for their atomics: 27.6 vs 25.2 ms of trace
  // at 8 spp), a 1/8 share -- where a round is pure latency -- with the lists (17.4 vs 18.1 ms at
  // 16 spp).  GPT_SHADE_SORT=1 / 0 forces one; default: lists for groups of <= 100 000 slots.
  static int shade_sort = -2;
  if (shade_sort == -2) { const char *e = getenv("GPT_SHADE_SORT"); shade_sort = e ? atoi(e) : -1; }
  const bool class_lists = shade_sort >= 0 ? !shade_sort : (grp[0].n <= 100000);
  if (class_lists)
    GPT_CUDA(ctx, cudaMemsetAsync(rc->ccount, 0, sizeof(int) * RenderCtx::MAX_GROUPS * 3 * PT_NCLASS, st));

  unsigned long long *ray_counter = ctx->d_counters + 6;
  GPT_CUDA(ctx, cudaMemsetAsync(ray_counter, 0, sizeof(unsigned long long), st));
  GPT_CUDA(ctx, cudaEventRecord(rc->fork_ev, st));
  for (int g = 0; g < G; g++)
    if (G > 1) GPT_CUDA(ctx, cudaStreamWaitEvent(rc->gstream[g], rc->fork_ev, 0));
  const long long max_iters = (long long)p.sample_size * p.dof_size * p.level + 8 + LAG;
  bool list_overflow = false;
  std::vector<int> round_active;  // stage timing only: queue length the host sized each round with
  int remaining = 0;
  for (int g = 0; g < G; g++) remaining += grp[g].done ? 0 : 1;
  cudaStream_t const caller_stream = ctx->stream;
  int rc_err = GPT_OK;
  while (remaining > 0 && rc_err == GPT_OK) {
    for (int g = 0; g < G && rc_err == GPT_OK; g++) {
      Group &q = grp[g];
      if (q.done) continue;
      cudaStream_t gs = G > 1 ? rc->gstream[g] : st;
      int *h_ring = rc->h_counts + CNT * RING * g;
      int *d_counts = ps.counts + CNT * g;
      if (q.it >= LAG) {
        int slot = (int)((q.it - LAG) % RING);
        if (host_prof) {
          const auto w0 = std::chrono::steady_clock::now();
          GPT_CUDA(ctx, cudaEventSynchronize(rc->ring_ev[g][slot]));
          host_wait_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
        } else {
          GPT_CUDA(ctx, cudaEventSynchronize(rc->ring_ev[g][slot]));
        }
        const int *h = h_ring + CNT * slot;
        q.active = h[q.cur_of[slot]];
        if (h[4]) list_overflow = true;
      }
      if (q.active == 0 || list_overflow || q.it >= max_iters) {
        q.done = true;
        remaining--;
        continue;
      }
      int grid = (q.active + PT_BLOCK - 1) / PT_BLOCK;
      int *next_ray = d_counts + 3;
      const int c0 = q.c, c1 = (q.c + 1) % 3, c2 = (q.c + 2) % 3;
      uint32_t *hist = rc->sort_hist + (size_t)rc->hist_stride * g;
      stamp_stream = gs;
      stamp_g = g;
      GPT_CUDA(ctx, stamp());
      {
        QueueSource qs{ps, ps.queue[q.cur] + q.s0, d_counts + c0, ctx->scene.shade,
                       class_lists ? nullptr : rc->sort_key + q.s0, rc->sort_val + q.s0, loc_bits, loc_shift,
                       rc->classq + (size_t)PT_NCLASS * q.s0, q.n, rc->ccount + (3 * g + c0) * PT_NCLASS};
        const int tneed = (q.active + TP_BLOCK - 1) / TP_BLOCK;
        const bool wide_now = wide_walk && q.active >= wide_min_rays;
        const int trace_grid = trace_grid_of[wide_now ? 1 : 0];
        int tgrid = tneed < trace_grid ? tneed : trace_grid;
        // The wide walk wins on throughput (1080p frame at 16 spp: 76 vs 85 ms), the binary walk
        // with its cooperative tail on latency, i.e. when every ray of the launch is in flight at
        // once and the launch lasts as long as its slowest ray (a quarter of the frame: 26.5 vs
        // 29.2 ms, an eighth: 18.1 vs 20.7 ms).  GPT_TRACE_WIDE_MIN_RAYS moves the switch.
        if (wide_now) {
          if (counting)
            k_pt_trace_wide<true><<<tgrid, TP_BLOCK, 0, gs>>>(qs, gpt_wide_tree(mb), next_ray, ctx->d_counters);
          else
            k_pt_trace_wide<false><<<tgrid, TP_BLOCK, 0, gs>>>(qs, gpt_wide_tree(mb), next_ray, ctx->d_counters);
        } else {
        auto kern = counting ? (coop ? k_pt_trace<true, true> : k_pt_trace<true, false>)
                             : (coop ? k_pt_trace<false, true> : k_pt_trace<false, false>);
        if (defer_leaves)
          kern = counting ? (coop ? k_pt_trace<true, true, true> : k_pt_trace<true, false, true>)
                          : (coop ? k_pt_trace<false, true, true> : k_pt_trace<false, false, true>);
        kern<<<tgrid, TP_BLOCK, 0, gs>>>(qs, mb.nodes, mb.tris, mb.parent_leaf, next_ray,
                                         ctx->d_counters);
        }
      }
      GPT_LAUNCH_CHECK(ctx);
      trace_launches++;
      if (timing) round_active.push_back(q.active);
      GPT_CUDA(ctx, stamp());
      // stable 7-bit radix sort of the active slots by material (ties keep tile order)
      uint32_t *sk = nullptr;
      int *sorted_slots = nullptr;
      static int no_sort = -1;  // experiment knob: shade in queue order
      if (no_sort < 0) { const char *e = getenv("GPT_NO_SORT"); no_sort = e ? atoi(e) : 0; }
      ctx->stream = gs;  // the sort enqueues on ctx->stream
      if (class_lists) sorted_slots = nullptr;
      else if (no_sort) sorted_slots = rc->sort_val + q.s0;
      else rc_err = gpt_radix_sort_pairs(ctx, rc->sort_key + q.s0, rc->sort_val + q.s0,
                                    rc->sort_key_alt + q.s0, rc->sort_val_alt + q.s0,
                                    hist, q.active, 7 + loc_bits, &sk, &sorted_slots, d_counts + c0);
      ctx->stream = caller_stream;
      if (rc_err != GPT_OK) break;
      if (defer_sss) {
        p.sss_req = rc->sss_req + 3 * (size_t)q.s0;
        p.sss_req_count = rc->sss_req_cnt + 2 * g + (int)(q.it & 1);
      }
      k_pt_shade<<<grid, PT_BLOCK, 0, gs>>>(ps, p, sorted_slots, d_counts + c0,
                                           ps.queue[q.cur ^ 1] + q.s0, d_counts + c1,
                                           ctx->scene, atlas, mb.tris, lt, d_fb, d_counts + 4,
                                           ray_counter, d_counts + c2, d_counts + 3,
                                           class_lists ? rc->classq + (size_t)PT_NCLASS * q.s0 : nullptr, q.n,
                                           rc->ccount + (3 * g + c0) * PT_NCLASS,
                                           rc->ccount + (3 * g + c2) * PT_NCLASS);
      GPT_LAUNCH_CHECK(ctx);
      if (defer_sss) {  // the round's SSS requests: one warp per lookup, then the slots rejoin the queue
        int *cnt = rc->sss_req_cnt + 2 * g;
        const int par = (int)(q.it & 1);
        k_sss_lookup<<<lookup_grid, LK_BLOCK, 0, gs>>>(ps, p, cnt + par, cnt + (par ^ 1),
                                                     ps.queue[q.cur ^ 1] + q.s0, d_counts + c1, d_fb);
        GPT_LAUNCH_CHECK(ctx);
      }
      GPT_CUDA(ctx, stamp());
      q.cur ^= 1;
      q.c = c1;
      int slot = (int)(q.it % RING);
      q.cur_of[slot] = q.c;  // which of counts[0..2] the NEXT iteration's queue length is
      GPT_CUDA(ctx, cudaMemcpyAsync(h_ring + CNT * slot, d_counts, CNT * sizeof(int),
                                    cudaMemcpyDeviceToHost, gs));
      GPT_CUDA(ctx, cudaEventRecord(rc->ring_ev[g][slot], gs));
      q.it++;
    }
  }
  if (rc_err != GPT_OK) return rc_err;
  if (host_prof) {
    const double loop_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - host_t0).count();
    fprintf(stderr, "[gpt host] G=%d LAG=%d: launch loop %.3f ms, of which %.3f ms waiting for queue lengths, %llu trace launches\n",
            G, LAG, loop_s * 1e3, host_wait_s * 1e3, (unsigned long long)trace_launches);
  }
  int active = 0;
  {  // join the groups; final queue lengths, error flags, exact ray count
    for (int g = 0; g < G && G > 1; g++) {
      GPT_CUDA(ctx, cudaEventRecord(rc->join_ev[g], rc->gstream[g]));
      GPT_CUDA(ctx, cudaStreamWaitEvent(st, rc->join_ev[g], 0));
    }
    int h_fin[RenderCtx::CNT * RenderCtx::MAX_GROUPS];
    GPT_CUDA(ctx, cudaMemcpyAsync(h_fin, ps.counts, sizeof(int) * CNT * G, cudaMemcpyDeviceToHost, st));
    unsigned long long h_rays = 0, h_overflow = 0;
    GPT_CUDA(ctx, cudaMemcpyAsync(&h_rays, ray_counter, sizeof(h_rays), cudaMemcpyDeviceToHost, st));
    GPT_CUDA(ctx, cudaMemcpyAsync(&h_overflow, ctx->d_counters + 3, sizeof(h_overflow), cudaMemcpyDeviceToHost, st));
    GPT_CUDA(ctx, cudaStreamSynchronize(st));
    if (h_overflow)
      return gpt_fail(ctx, GPT_E_UNSUPPORTED, "traversal stack overflow (> 64 levels)", __FILE__, __LINE__);
    for (int g = 0; g < G; g++) {
      active += h_fin[CNT * g + grp[g].c];
      if (h_fin[CNT * g + 4]) list_overflow = true;
    }
    rays = shades = h_rays;
  }
  if (list_overflow)
    return gpt_fail(ctx, GPT_E_UNSUPPORTED, "nested-dielectric list deeper than 12", __FILE__, __LINE__);
  if (fw) {
    k_pt_combine<<<(npix + PT_BLOCK - 1) / PT_BLOCK, PT_BLOCK, 0, st>>>(
        ps.pixel, rc->partial, npix, n / npix, fw->first ? 1 : 0, fw->last ? 1 : 0, p.inv_samples, d_fb);
    GPT_LAUNCH_CHECK(ctx);
  }
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
  GPT_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
  float ms = 0.f;
  GPT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  ctx->stats.render_ms = ms;
  ctx->stats.rays_traced = rays;
  ctx->stats.shade_events = shades;
  ctx->stats.kernel_launches = ctx->launches;
  ctx->stats.trace_launches = trace_launches;
  if (timeline) {
    if (FILE *lf = fopen(timeline_path, "w")) {
      for (size_t k = 0; k + 2 < nev; k += 3) {
        float t[3] = {0.f, 0.f, 0.f};
        for (int j = 0; j < 3; j++) cudaEventElapsedTime(&t[j], ctx->ev0, rc->events[k + j]);
        fprintf(lf, "%d %zu %.4f %.4f %.4f\n", stamp_group[k], k / 3, t[0], t[1], t[2]);
      }
      fclose(lf);
    }
  }
  if (timing) {  // three stamps per iteration: before trace, between, after shade
    double tt = 0, ts = 0;
    // GPT_ROUND_LOG=path: one line per round of the stage-timed frame (round, queue length as
    // the host knew it, trace ms, sort + shade ms) -- the data behind DESIGN 4.5's chain analysis
    const char *log_path = getenv("GPT_ROUND_LOG");
    FILE *lf = log_path ? fopen(log_path, "w") : nullptr;
    for (size_t k = 0; k + 2 < nev; k += 3) {
      float a = 0.f, b = 0.f;
      cudaEventElapsedTime(&a, rc->events[k], rc->events[k + 1]);
      cudaEventElapsedTime(&b, rc->events[k + 1], rc->events[k + 2]);
      tt += a;
      ts += b;
      if (lf) fprintf(lf, "%zu %d %.4f %.4f\n", k / 3, k / 3 < round_active.size() ? round_active[k / 3] : -1, a, b);
    }
    if (lf) fclose(lf);
    ctx->stats.trace_kernel_ms = tt;
    ctx->stats.shade_kernel_ms = ts;
  }
  if (counting) {
    unsigned long long c[8];
    GPT_CUDA(ctx, cudaMemcpy(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
    ctx->stats.internal_visits = c[0];
    ctx->stats.leaf_tests = c[1];
    ctx->stats.wide_node_visits = c[5];
    ctx->stats.wide_leaf_tests = c[7];
  }
  if (active > 0) return gpt_fail(ctx, GPT_E_INVALID, "path queue did not drain", __FILE__, __LINE__);
  return GPT_OK;
}

extern "C" {

void gpt_render_free(gpt_ctx *ctx) {
  if (!ctx || !ctx->render_state) return;
  RenderCtx *rc = static_cast<RenderCtx *>(ctx->render_state);
  PathState &ps = rc->ps;
  void *p[] = {ps.pixel, ps.rng, ps.rays, ps.hits, ps.ma, ps.cm, ps.cam,
               ps.queue[0], ps.queue[1], ps.counts, rc->leaf_info, rc->node_info,
               rc->node_energy, rc->lt_flags, rc->rng_init, rc->d_fb, rc->sort_key,
               rc->sort_key_alt, rc->sort_hist, rc->sort_val, rc->sort_val_alt, rc->sq_cells, rc->sq_ctl,
               rc->classq, rc->ccount, rc->partial, rc->sss_req, rc->sss_req_cnt};
  for (void *q : p) gpt_free(ctx, q);
  if (rc->h_counts) cudaFreeHost(rc->h_counts);
  if (rc->fork_ev) cudaEventDestroy(rc->fork_ev);
  for (int g = 0; g < RenderCtx::MAX_GROUPS; g++) {
    if (rc->gstream[g]) cudaStreamDestroy(rc->gstream[g]);
    if (rc->join_ev[g]) cudaEventDestroy(rc->join_ev[g]);
    for (int k = 0; k < RenderCtx::RING; k++)
      if (rc->ring_ev[g][k]) cudaEventDestroy(rc->ring_ev[g][k]);
  }
  for (cudaEvent_t e : rc->events) cudaEventDestroy(e);
  delete rc;
  ctx->render_state = nullptr;
}

// do_sss_first_pass on demand (parity export); a render runs it automatically
int gpt_render_sss_first_pass(gpt_ctx *ctx) {
  if (!ctx) return GPT_E_INVALID;
  SssView sv{};
  bool done = true;
  if (!gpt_sss_view(ctx, &sv, &done) || done) return GPT_OK;
  GPT_TRY(render_device(ctx, nullptr, 0, 1, true));
  gpt_sss_mark_first_pass_done(ctx);
  return GPT_OK;
}

int gpt_render_pathtrace_device(gpt_ctx *ctx, void *d_out_rgb, int rank, int nranks) {
  if (!ctx || !d_out_rgb) return GPT_E_INVALID;
  return render_device(ctx, (float *)d_out_rgb, rank, nranks);
}

int gpt_render_pathtrace(gpt_ctx *ctx, float *out_rgb) {
  if (!ctx || !out_rgb) return GPT_E_INVALID;
  if (!ctx->have_cam) return gpt_fail(ctx, GPT_E_INVALID, "gpt_set_camera first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  RenderCtx *rc = rc_of(ctx);
  size_t nfl = 3 * (size_t)ctx->cam.width * ctx->cam.height;
  if (nfl > rc->fb_capacity) {
    GPT_TRY(gpt_alloc(ctx, &rc->d_fb, nfl));
    rc->fb_capacity = nfl;
  }
  GPT_TRY(render_device(ctx, rc->d_fb, 0, 1));
  GPT_CUDA(ctx, cudaMemcpyAsync(out_rgb, rc->d_fb, sizeof(float) * nfl, cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return GPT_OK;
}

}  // extern "C"
