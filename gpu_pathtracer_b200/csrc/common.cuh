// common.cuh -- shared declarations of the sm_100a hot path (context, device layouts).
//
// ARITHMETIC CONTRACT: every .cu of this library is compiled with -fmad=false, so a*b+c is
// evaluated as an IEEE fp32 multiply followed by an IEEE fp32 add, exactly like the
// reference's host replay (g++ -ffp-contract=off).  Where a fused multiply-add is wanted for
// speed in code that is NOT parity-critical it is written explicitly as fmaf().
// Division and sqrt are the IEEE-rounded defaults (-prec-div=true -prec-sqrt=true).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string.h>

#include <string>
#include <vector>

#include "../../include/gpt.h"

#define GPT_SMALL 1e-5f  // SMALL_DOUBLE, src/param.h:6

// ---- device layouts ------------------------------------------------------------------
// BVH node, 64 B, four 16-byte words fetched with ld.global.nc.v4:
//   q0 = (L.xmin, L.xmax, L.ymin, L.ymax)
//   q1 = (L.zmin, L.zmax, R.xmin, R.xmax)
//   q2 = (R.ymin, R.ymax, R.zmin, R.zmax)
//   q3 = (left_ref, right_ref, parent_enc, range_first)  as int bits
// child ref c >= 0: internal node c; c < 0: leaf ~c (sorted position).
// parent_enc = parent * 2 + is_right_child, -1 for the root.
// Internal node i lives at index i and the root is node 0, as in the reference
// (src/model/bvh/bvh_building.h:63,72,179-183).
//
// Trace triangle, 48 B, three 16-byte words, stored in SORTED (leaf) order:
//   t0 = (p1.x, p1.y, p1.z, p2.x)  t1 = (p2.y, p2.z, p3.x, p3.y)
//   t2 = (p3.z, tolerance, prim_index bits, 0)

struct BvhDev {
  int n = 0;             // leaves
  int capacity = 0;      // allocated leaves
  float4 *nodes = nullptr;      // 4 * (n-1) float4
  float4 *tris = nullptr;       // 3 * n float4 (triangle BVHs only)
  int *parent_leaf = nullptr;   // n, encoded like parent_enc
  uint32_t *morton = nullptr;   // n sorted keys
  int *sorted_prim = nullptr;   // n: index into the primitive list the BVH was built over
  int *prim_list = nullptr;     // optional subset list -> original triangle index (NULL = identity)
  uint32_t *world_enc = nullptr;  // 6 order-preserving encoded floats: min xyz, max xyz
  // build scratch
  uint32_t *keys_alt = nullptr;
  int *vals_alt = nullptr;
  uint32_t *block_hist = nullptr;
  int hist_capacity = 0;
  int *flags = nullptr;         // per internal node: last leaf of its range (k_emit -> k_node_boxes); arrival counters of the chain refit
  float *levels = nullptr;      // 4-ary hierarchy of partial box unions over the sorted leaves (lbvh.cu k_leaf_levels)
  // 4-wide quantised tree collapsed from the LBVH (wide.cu, trace_wide.cuh)
  uint4 *wq = nullptr;          // 3 x 16 bytes per wide node, node 0 = root
  float4 *wtris = nullptr;      // the 48-byte trace triangles once more, in wide-tree leaf order; t2.w = LBVH leaf
  float *wmeta = nullptr;       // [0] max |coordinate| of the world box, [1] 2^23
  int wide_capacity = 0;        // leaves the three buffers were allocated for
  int wide_nodes = 0;           // wide nodes of the current tree, 0 = none (the binary walk is used)
};

// Per-triangle shading record in ORIGINAL order (7 float4 = 112 B), filled by preprocess:
//   s0 = (n1.x n1.y n1.z n2.x) s1 = (n2.y n2.z n3.x n3.y) s2 = (n3.z uv1.u uv1.v uv2.u)
//   s3 = (uv2.v uv3.u uv3.v area) s4 = (tg1.x tg1.y tg1.z tg2.x) s5 = (tg2.y tg2.z tg3.x tg3.y)
//   s6 = (tg3.z, material_idx bits, flags bits, object_idx bits)

struct MaterialDev {  // one per material, constant-ish, read through the read-only path
  float ka[3], kd[3], ks[3], ke[3], tf[3];
  float path_length, t_r, n_s, n_i, bm;
  float scattering_coef, absorption_coef, extinction_coef, g, scattering_prob;
  int priority;
  int sss;
  int tex_h[5], tex_w[5], tex_off[5];  // diffuse, specular, emission, n_s, bump
  int pad;
};

struct SceneDev {
  // raw SoA as uploaded (src/main.cu:386-450)
  int nv = 0, nvn = 0, nvt = 0, ntri = 0, nobj = 0, nmat = 0;
  float *x = nullptr, *y = nullptr, *z = nullptr;
  float *xn = nullptr, *yn = nullptr, *zn = nullptr;
  float *xt = nullptr, *yt = nullptr;
  int *p1 = nullptr, *p2 = nullptr, *p3 = nullptr;
  int *n1 = nullptr, *n2 = nullptr, *n3 = nullptr;
  int *t1 = nullptr, *t2 = nullptr, *t3 = nullptr;
  int *mat_idx = nullptr, *obj_idx = nullptr;
  MaterialDev *mats = nullptr;
  float *tex_r = nullptr, *tex_g = nullptr, *tex_b = nullptr;
  int64_t tex_len = 0;
  float *bg_r = nullptr, *bg_g = nullptr, *bg_b = nullptr;
  int bg_h = 1, bg_w = 1;
  // derived per triangle (original order)
  float *tolerance = nullptr;  // N
  float *aabb = nullptr;       // 6N: xmin,xmax,ymin,ymax,zmin,zmax
  float *normal = nullptr;     // 3N
  float *energy = nullptr;     // 3N
  float4 *shade = nullptr;     // 7N
  float *vtangent = nullptr;   // 3 * nv accumulated per-vertex tangents
};

struct gpt_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  SceneDev scene;
  BvhDev bvh[3];
  GptCamera cam{};
  GptParams params{};
  GptStats stats{};
  bool have_mesh = false, have_mats = false, have_cam = false, built = false;
  // bumped by gpt_build: anything derived from the built scene and cached elsewhere (the
  // light-tree side arrays of render.cu / sppm.cu) is stale when its epoch differs
  uint64_t build_epoch = 0;
  cudaMemPool_t pool = nullptr;              // this context's own stream-ordered pool (gpt_alloc)
  unsigned long long *d_counters = nullptr;  // [0] V_int [1] V_leaf [2] rays [3] overflow [4] trace.cu cursor [5] wide-node share of [0] [6] render ray count [7] wide-walk share of [1]
  void *render_state = nullptr;              // owned by render.cu
  void *sppm_state = nullptr;                // owned by sppm.cu
  void *sss_state = nullptr;                 // owned by sss.cu
  std::vector<int> obj_num, obj_off;         // object_num_primitives / object_primitive_offset_idx
  uint64_t launches = 0;                     // kernels launched since last reset
  bool prof_time_stages = false, prof_count_visits = false;
  float max_path_length = 0.f;  // largest sub-surface lookup radius of the materials (sss.cu grid)
  int wavefront_groups = 0;  // 0 = automatic (render.cu)
  void *nccl_comm = nullptr;  // ncclComm_t of gpt_comm_init (comm.cu)
  int comm_rank = 0, comm_nranks = 1;
  int fast_lanes = 0;        // rng_mode 1: independent sample streams per pixel, 0 = default (8)
  int pipeline = -1;         // path tracer: 0 = rounds of launches, 1 = streaming (two persistent kernels), -1 = default (env GPT_PIPELINE, else by frame size)
  int trace_defer = -1;      // path tracer: batch the triangle tests of a warp (deferred leaf tests): -1 = default (on, env GPT_TRACE_DEFER), 0 off, 1 on
  int trace_wide = -1;       // walk the 4-wide tree (trace_wide.cuh): -1 = default (on, env GPT_TRACE_WIDE), 0 = binary LBVH walk, 1 on
  int trace_tail = -1;       // cooperative tail of the persistent trace: -1 = default (on, env GPT_TRACE_COOP), 0 off, 1 on
};

// ---- error plumbing --------------------------------------------------------------------
int gpt_fail(gpt_ctx *ctx, int code, const char *what, const char *file, int line);
#define GPT_CUDA(ctx, call)                                                         \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess)                                                         \
      return gpt_fail((ctx), GPT_E_CUDA, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)
#define GPT_TRY(expr)            \
  do {                           \
    int rc__ = (expr);           \
    if (rc__ != GPT_OK) return rc__; \
  } while (0)
#define GPT_LAUNCH_CHECK(ctx)    \
  do {                           \
    (ctx)->launches++;           \
    GPT_CUDA((ctx), cudaGetLastError()); \
  } while (0)

// Device memory comes from a stream-ordered pool OWNED by the context (release threshold = keep
// everything; destroyed, and so returned to the device, by gpt_destroy -- the device's default
// pool is process-global and is left alone): a scene upload + build is ~60 buffers, and with plain
// cudaMalloc / cudaFree the driver calls alone cost more than the kernels (C2: build 10.6 ms
// of which < 2 ms is GPU work).  Every allocation and free is ordered on ctx->stream; the
// wavefront-group streams fork from and join back into it, so that order covers them too.
static inline void gpt_free(gpt_ctx *ctx, void *p) {
  if (p) cudaFreeAsync(p, ctx->stream);
}
template <class T>
static inline int gpt_alloc(gpt_ctx *ctx, T **p, size_t count) {
  if (*p) { gpt_free(ctx, *p); *p = nullptr; }
  if (count == 0) count = 1;
  GPT_CUDA(ctx, cudaMallocFromPoolAsync((void **)p, count * sizeof(T), ctx->pool, ctx->stream));
  return GPT_OK;
}
template <class T>
static inline int gpt_upload(gpt_ctx *ctx, T **p, const T *host, size_t count) {
  GPT_TRY(gpt_alloc(ctx, p, count));
  if (count)
    GPT_CUDA(ctx, cudaMemcpyAsync(*p, host, count * sizeof(T), cudaMemcpyHostToDevice,
                                  ctx->stream));
  return GPT_OK;
}

// ---- order-preserving float <-> uint encoding for atomicMin/atomicMax -------------------
__host__ __device__ static inline uint32_t enc_f32(float f) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(f);
#else
  union { float f; uint32_t u; } c; c.f = f; uint32_t b = c.u;
#endif
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ static inline float dec_f32(uint32_t e) {
  uint32_t b = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  union { float f; uint32_t u; } c; c.u = b; return c.f;
#endif
}

// ---- stage entry points (implemented in the .cu named) ----------------------------------
int gpt_preprocess(gpt_ctx *ctx);                                    // preprocess.cu
int gpt_lbvh_build(gpt_ctx *ctx, int which, const int *d_prim_list, int n);  // lbvh.cu
int gpt_lbvh_export(gpt_ctx *ctx, int which, uint32_t *morton, int32_t *sorted_to_orig,
                    int32_t *left, int32_t *right, int32_t *parent_internal,
                    int32_t *parent_leaf, float *node_aabb, float *leaf_aabb,
                    float *world_aabb);
int gpt_lbvh_build_boxes(gpt_ctx *ctx, BvhDev &b, const float *d_aabb, const int *d_prim_list,
                         int n, bool with_tris,
                         const uint32_t *d_keys = nullptr);           // lbvh.cu
void gpt_lbvh_free(gpt_ctx *ctx, BvhDev &b);
int gpt_lbvh_build_wide(gpt_ctx *ctx, BvhDev &b);                     // lbvh.cu
int gpt_collect_subset(gpt_ctx *ctx, int flag_mask, int **d_list, int *count);  // lbvh.cu
int gpt_trace_launch(gpt_ctx *ctx, int which, const float4 *d_rays, int64_t n,
                     float4 *d_hits, bool normalize, bool count);     // trace.cu
// SSS (sss.cu): point clouds + per-object point trees at build time; the first pass (radiance
// at the points) runs through the path-tracing wavefront (render.cu) on first use
struct SssTree {           // per scene OBJECT; n == 0 for objects without sub-surface scattering
  const float4 *nodes;     // AABB node64 of the object's point LBVH
  const float4 *leaf;      // 2 float4 per point in leaf order: (location, -) (colour, -)
  int n;
  int pad;
  // search structure (ours): an LBVH of the same points keyed by their TRUE Morton codes;
  // spos[j] = (location, VISIT position of the point in the reference walk) of search leaf j;
  // visit[k] = leaf-order index of the k-th point the reference walk would test
  const float4 *snodes;
  const float4 *spos;
  const int *visit;
  // candidate grid (ours): the same (location, visit position) records sorted by the cell of a
  // uniform grid over the object's points, cell edge >= the largest lookup radius of the scene;
  // gstart[c] .. gstart[c + 1] = the records of cell c = x + gnx * (y + gny * z).
  // gpos == nullptr: no grid, the search tree finds the candidates.
  const float4 *gpos;
  const int *gstart;
  float gox, goy, goz, ginv;
  int gnx, gny, gnz;
  // the leaf records once more, in VISIT order (vleaf[2 k] = leaf[2 visit[k]]; filled after the
  // first pass wrote the colours): the ordered sum of a lookup reads them front to back
  const float4 *vleaf;
};
struct SssView {
  const SssTree *trees;    // device, one per object; nullptr when the scene has no SSS points
  const float4 *pts;       // 3 float4 per point, tree order: (loc, filter.r)(filter.gb, normal.xy)(normal.z, -, -, -)
  float4 *leaf;            // all trees' leaf records, tree order (colour written by the first pass)
  const uint32_t *rng;     // rand_state_sss, 6 planes of stride npoints
  int npoints;
  // pool of private candidate bitmaps for sss_lookup (one bit per point of an object)
  uint32_t *bm_pool;
  int *bm_busy;
  int bm_n, bm_words;
};
// collectives used inside gpt_sppm_pass when a communicator is bound               // comm.cu
int gpt_comm_allreduce_f32(gpt_ctx *ctx, void *d_buf, size_t n);
int gpt_comm_allgather_count(gpt_ctx *ctx, int mine, int *d_scratch, int *counts);
int gpt_comm_allgatherv(gpt_ctx *ctx, const void *d_send, void *d_recv, const int *counts,
                        size_t elem_bytes);
int gpt_sss_prepare(gpt_ctx *ctx);                                   // sss.cu
bool gpt_sss_view(gpt_ctx *ctx, SssView *out, bool *first_pass_done);  // sss.cu
void gpt_sss_mark_first_pass_done(gpt_ctx *ctx);                     // sss.cu
// stable LSD radix sort of (key, value) pairs on key bits [0, nbits)   // lbvh.cu
int gpt_radix_sort_pairs(gpt_ctx *ctx, uint32_t *keys, int *vals, uint32_t *keys_alt,
                         int *vals_alt, uint32_t *block_hist, int n, int nbits,
                         uint32_t **keys_out, int **vals_out,
                         const int *d_n = nullptr);  // d_n: exact count on the device (n = bound)
int gpt_radix_hist_entries(int n);
bool gpt_trace_coop(const gpt_ctx *ctx, const BvhDev &b);  // trace.cu
bool gpt_trace_defer(const gpt_ctx *ctx);                   // trace.cu
bool gpt_trace_wide(const gpt_ctx *ctx, const BvhDev &b);   // trace.cu
