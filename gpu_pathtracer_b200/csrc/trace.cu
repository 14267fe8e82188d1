// trace.cu -- stand-alone closest-hit kernel over a ray queue (gpt_trace*).
//
// Persistent warps with ray refill and a cooperative tail (trace_persistent.cuh), 128-thread CTAs, traversal stack in shared memory laid out
// [level][thread] (bank-conflict free), nodes / triangles fetched as 16-byte words through
// the read-only path.  Rays: 2 x float4 {ox,oy,oz,tmax}{dx,dy,dz,-}; hits: 2 x float4
// {t, leaf, prim, -}{b1, b2, b3, -}.
#include <cstdlib>

#include "trace_wide.cuh"

namespace {

template <bool NORMALIZE>
struct ArraySource {
  const float4 *rays;
  float4 *hits;
  int n;
  __device__ int count() const { return n; }
  __device__ void load(int i, float4 &o, float4 &d) const {
    o = __ldg(rays + 2 * (size_t)i);
    d = __ldg(rays + 2 * (size_t)i + 1);
    if (NORMALIZE) {  // Ray::Ray -> unit_vector(dir) = dir / dir.length()  (ray.h:21-24)
      float len = sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
      d.x = d.x / len; d.y = d.y / len; d.z = d.z / len;
    }
  }
  __device__ void store_warp(bool fin, int i, const TraceLane &l) const {
    if (fin) store(i, l);
  }
  __device__ void store(int i, const TraceLane &l) const {
    hits[2 * (size_t)i] = make_float4(l.t, __int_as_float(l.leaf), __int_as_float(l.prim), 0.f);
    hits[2 * (size_t)i + 1] = make_float4(l.b1, l.b2, l.b3, 0.f);
  }
};

template <bool NORMALIZE, bool COUNT, bool COOP, bool DEFER>
__global__ void __launch_bounds__(TP_BLOCK, TP_MINBLOCKS)
k_trace(const float4 *__restrict__ nodes, const float4 *__restrict__ tris,
        const int *__restrict__ parent_leaf, ArraySource<NORMALIZE> src, int *next_ray,
        unsigned long long *counters) {
  __shared__ int stack[TP_STACK_SMEM * TP_BLOCK];
  __shared__ int coop[COOP ? (TP_BLOCK / 32) * COOP_WARP_WORDS : 1];
  __shared__ int defer[DEFER ? 2 * DEFER_SLOTS * TP_BLOCK : 1];
  trace_persistent<COUNT, ArraySource<NORMALIZE>, COOP, DEFER>(nodes, tris, parent_leaf, src,
                                                               next_ray, stack, counters, coop,
                                                               defer);
}

// the same queue through the 4-wide tree (trace_wide.cuh)
template <bool NORMALIZE, bool COUNT>
__global__ void __launch_bounds__(TP_BLOCK, W4_MINBLOCKS)
k_trace_wide(WideTree tree, ArraySource<NORMALIZE> src, int *next_ray,
             unsigned long long *counters) {
  __shared__ int smem[W4_SMEM_WORDS];
  trace_wide<COUNT, ArraySource<NORMALIZE>>(tree, src, next_ray, smem, counters);
}

template <class K>
int persistent_grid(K kernel) {
  int dev = 0, sms = 148, per_sm = 4;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, TP_BLOCK, 0);
  return sms * (per_sm > 0 ? per_sm : 1);
}

}  // namespace

// Whether the persistent trace kernels finish their last rays cooperatively
// (trace_persistent.cuh): on unless gpt_set_trace_tail / GPT_TRACE_COOP=0 says otherwise, and
// only while a node index fits the work-list entry.
bool gpt_trace_coop(const gpt_ctx *ctx, const BvhDev &b) {
  static int env = -1;
  if (env < 0) {
    const char *e = getenv("GPT_TRACE_COOP");
    env = e ? (atoi(e) != 0) : 1;
  }
  int on = ctx->trace_tail >= 0 ? ctx->trace_tail : env;
  return on && b.n < (1 << COOP_NODE_BITS);
}

// Whether the persistent trace kernels batch a warp's triangle tests (deferred leaf tests,
// trace_persistent.cuh): on unless gpt_set_trace_leaf_batching / GPT_TRACE_DEFER=0 says otherwise.
bool gpt_trace_defer(const gpt_ctx *ctx) {
  static int env = -1;
  if (env < 0) {
    const char *e = getenv("GPT_TRACE_DEFER");
    env = e ? (atoi(e) != 0) : 1;
  }
  return ctx->trace_defer >= 0 ? ctx->trace_defer != 0 : env != 0;
}

// Whether the trace kernels walk the 4-wide tree (trace_wide.cuh) instead of the binary LBVH: on
// unless gpt_set_trace_bvh / GPT_TRACE_WIDE=0 says otherwise, and only for a tree that has one.
bool gpt_trace_wide(const gpt_ctx *ctx, const BvhDev &b) {
  static int env = -1;
  if (env < 0) {
    const char *e = getenv("GPT_TRACE_WIDE");
    env = e ? (atoi(e) != 0) : 1;
  }
  int on = ctx->trace_wide >= 0 ? ctx->trace_wide : env;
  return on && b.wide_nodes > 0 && b.n >= 2;
}

int gpt_trace_launch(gpt_ctx *ctx, int which, const float4 *d_rays, int64_t n, float4 *d_hits,
                     bool normalize, bool count) {
  if (which < 0 || which > 2) return gpt_fail(ctx, GPT_E_INVALID, "bad BVH id", __FILE__, __LINE__);
  BvhDev &b = ctx->bvh[which];
  if (b.n < 2) return gpt_fail(ctx, GPT_E_INVALID, "BVH not built (needs >= 2 leaves)", __FILE__, __LINE__);
  if (n <= 0) return GPT_OK;
  if (n > 0x7fffffff) return gpt_fail(ctx, GPT_E_INVALID, "more than 2^31 rays in one call", __FILE__, __LINE__);
  unsigned long long *c = ctx->d_counters;
  int *next_ray = reinterpret_cast<int *>(c + 4);
  // [0..2] visit counters, [3] the stack-overflow flag: cleared for every call (it was sticky)
  GPT_CUDA(ctx, cudaMemsetAsync(c, 0, 4 * sizeof(unsigned long long), ctx->stream));
  GPT_CUDA(ctx, cudaMemsetAsync(next_ray, 0, sizeof(int), ctx->stream));
  cudaStream_t st = ctx->stream;
  if (gpt_trace_wide(ctx, b)) {
#define GPT_TRACE_LAUNCH_W(NORM, CNT)                                                      \
  do {                                                                                     \
    static int grid = persistent_grid(k_trace_wide<NORM, CNT>);                            \
    long long need = (n + TP_BLOCK - 1) / TP_BLOCK;                                        \
    ArraySource<NORM> src{d_rays, d_hits, (int)n};                                         \
    k_trace_wide<NORM, CNT><<<(unsigned)(need < grid ? need : grid), TP_BLOCK, 0, st>>>(   \
        gpt_wide_tree(b), src, next_ray, c);                                               \
  } while (0)
    if (normalize) { if (count) GPT_TRACE_LAUNCH_W(true, true); else GPT_TRACE_LAUNCH_W(true, false); }
    else { if (count) GPT_TRACE_LAUNCH_W(false, true); else GPT_TRACE_LAUNCH_W(false, false); }
#undef GPT_TRACE_LAUNCH_W
    GPT_LAUNCH_CHECK(ctx);
    return GPT_OK;
  }
  const bool coop = gpt_trace_coop(ctx, b);
  const bool defer = gpt_trace_defer(ctx);
#define GPT_TRACE_LAUNCH(NORM, CNT, COOP, DEFER)                                           \
  do {                                                                                     \
    static int grid = persistent_grid(k_trace<NORM, CNT, COOP, DEFER>);                    \
    long long need = (n + TP_BLOCK - 1) / TP_BLOCK;                                        \
    ArraySource<NORM> src{d_rays, d_hits, (int)n};                                         \
    k_trace<NORM, CNT, COOP, DEFER><<<(unsigned)(need < grid ? need : grid), TP_BLOCK, 0, st>>>( \
        b.nodes, b.tris, b.parent_leaf, src, next_ray, c);                                 \
  } while (0)
#define GPT_TRACE_LAUNCH2(NORM, CNT)                                                       \
  do {                                                                                     \
    if (coop) { if (defer) GPT_TRACE_LAUNCH(NORM, CNT, true, true); else GPT_TRACE_LAUNCH(NORM, CNT, true, false); } \
    else { if (defer) GPT_TRACE_LAUNCH(NORM, CNT, false, true); else GPT_TRACE_LAUNCH(NORM, CNT, false, false); } \
  } while (0)
  if (normalize) { if (count) GPT_TRACE_LAUNCH2(true, true); else GPT_TRACE_LAUNCH2(true, false); }
  else { if (count) GPT_TRACE_LAUNCH2(false, true); else GPT_TRACE_LAUNCH2(false, false); }
#undef GPT_TRACE_LAUNCH2
#undef GPT_TRACE_LAUNCH
  GPT_LAUNCH_CHECK(ctx);
  return GPT_OK;
}
