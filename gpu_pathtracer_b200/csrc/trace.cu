// trace.cu -- stand-alone closest-hit kernel over a ray queue (gpt_trace*).
//
// Persistent warps with ray refill and a cooperative tail (trace_persistent.cuh), 128-thread CTAs, traversal stack in shared memory laid out
// [level][thread] (bank-conflict free), nodes / triangles fetched as 16-byte words through
// the read-only path.  Rays: 2 x float4 {ox,oy,oz,tmax}{dx,dy,dz,-}; hits: 2 x float4
// {t, leaf, prim, -}{b1, b2, b3, -}.
#include <cstdio>
#include <cstdlib>
#include <mutex>

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
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 5, 0, sizeof(unsigned long long), ctx->stream));
  GPT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters + 7, 0, sizeof(unsigned long long), ctx->stream));
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

// ---- gpt_trace_static: closest hits for HOST ray / hit buffers without any allocation -----------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  gpt_trace stages its rays and hit records in buffers it takes from the context's pool on
// every call; here the staging AND the kernel's own scratch (the persistent kernel's ray cursor,
// the visit counters, the stack-overflow flag) live in one file-scope device array of a fixed,
// modest size that every GPU that loads the library holds until exit:
//   float4 [0, 16)                       256 bytes of control words, laid out like ctx->d_counters
//                                        (u64 [0] V_int [1] V_leaf [2] rays [3] overflow, int cursor at u64 [4])
//   float4 [16, 16 + 2 MAX)              rays, 2 float4 each
//   float4 [16 + 2 MAX, 16 + 4 MAX)      hit records, 2 float4 each
// = 256 + 64 * GPT_TRACE_STATIC_MAX_RAYS bytes = 8 MiB + 256 B.  The array keeps the last call's
// contents: the control words are cleared at the start of every call (the cursor MUST start at
// zero), rays and hits are fully overwritten for the n rays of the call.  All calls on a device
// share the array, whatever their context: two calls that overlap on the GPU race.  (This
// function returns only after its own work is complete, so calls from one host thread never do.)
constexpr int TS_MAX = GPT_TRACE_STATIC_MAX_RAYS;
constexpr int TS_CTL = 16;  // float4 words of control data
static __device__ __align__(256) float4 g_trace_static[TS_CTL + 4 * (size_t)TS_MAX];

// In host code `g_trace_static` is a host-side stand-in for the symbol, NOT a device address:
// the address a kernel can use comes from cudaGetSymbolAddress, once per device, with that
// device current.
static int trace_static_base(gpt_ctx *ctx, float4 **base) {
  static float4 *per_device[64] = {nullptr};
  static std::mutex mu;
  const int dev = ctx->device;
  if (dev < 0 || dev >= 64) return gpt_fail(ctx, GPT_E_INVALID, "device index beyond 63", __FILE__, __LINE__);
  std::lock_guard<std::mutex> lock(mu);
  if (!per_device[dev]) {
    GPT_CUDA(ctx, cudaSetDevice(dev));
    float4 *p = nullptr;
    GPT_CUDA(ctx, cudaGetSymbolAddress((void **)&p, g_trace_static));
    if (!p) return gpt_fail(ctx, GPT_E_CUDA, "cudaGetSymbolAddress returned a null address", __FILE__, __LINE__);
    per_device[dev] = p;
  }
  *base = per_device[dev];
  return GPT_OK;
}

extern "C" int gpt_trace_static(gpt_ctx *ctx, int which, const float *rays, int64_t n,
                                int32_t *leaf_id, int32_t *tri_id, float *t, float *bary) {
  if (!ctx || !ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  if (n < 0 || (n > 0 && !rays)) return GPT_E_INVALID;
  if (n > TS_MAX) {
    char msg[200];
    snprintf(msg, sizeof(msg), "gpt_trace_static: %lld rays exceed the static scratch (GPT_TRACE_STATIC_MAX_RAYS = %d); split the call or use gpt_trace",
             (long long)n, TS_MAX);
    return gpt_fail(ctx, GPT_E_INVALID, msg, __FILE__, __LINE__);
  }
  if (which < 0 || which > 2) return gpt_fail(ctx, GPT_E_INVALID, "bad BVH id", __FILE__, __LINE__);
  BvhDev &b = ctx->bvh[which];
  if (b.n < 2) return gpt_fail(ctx, GPT_E_INVALID, "BVH not built (needs >= 2 leaves)", __FILE__, __LINE__);
  if (n == 0) return GPT_OK;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  float4 *base = nullptr;
  GPT_TRY(trace_static_base(ctx, &base));
  unsigned long long *ctl = reinterpret_cast<unsigned long long *>(base);
  int *cursor = reinterpret_cast<int *>(ctl + 4);
  float4 *d_rays = base + TS_CTL, *d_hits = base + TS_CTL + 2 * (size_t)TS_MAX;
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemsetAsync(base, 0, TS_CTL * sizeof(float4), st));  // cursor, counters, overflow flag
  GPT_CUDA(ctx, cudaMemcpyAsync(d_rays, rays, 32 * (size_t)n, cudaMemcpyHostToDevice, st));
  ctx->launches = 0;
  const long long need = (n + TP_BLOCK - 1) / TP_BLOCK;
  ArraySource<true> src{d_rays, d_hits, (int)n};
  if (gpt_trace_wide(ctx, b)) {
    static int grid = persistent_grid(k_trace_wide<true, false>);
    k_trace_wide<true, false><<<(unsigned)(need < grid ? need : grid), TP_BLOCK, 0, st>>>(
        gpt_wide_tree(b), src, cursor, ctl);
  } else {
    static int grid = persistent_grid(k_trace<true, false, false, false>);
    k_trace<true, false, false, false><<<(unsigned)(need < grid ? need : grid), TP_BLOCK, 0, st>>>(
        b.nodes, b.tris, b.parent_leaf, src, cursor, ctl);
  }
  GPT_LAUNCH_CHECK(ctx);
  // hit record of ray i = float4 pair {t, leaf, prim, -}{b1, b2, b3, -}: strided copies scatter the
  // fields straight into the caller's arrays (no host staging buffer either)
  const char *h = reinterpret_cast<const char *>(d_hits);
  if (t) GPT_CUDA(ctx, cudaMemcpy2DAsync(t, 4, h, 32, 4, (size_t)n, cudaMemcpyDeviceToHost, st));
  if (leaf_id) GPT_CUDA(ctx, cudaMemcpy2DAsync(leaf_id, 4, h + 4, 32, 4, (size_t)n, cudaMemcpyDeviceToHost, st));
  if (tri_id) GPT_CUDA(ctx, cudaMemcpy2DAsync(tri_id, 4, h + 8, 32, 4, (size_t)n, cudaMemcpyDeviceToHost, st));
  if (bary) GPT_CUDA(ctx, cudaMemcpy2DAsync(bary, 12, h + 16, 32, 12, (size_t)n, cudaMemcpyDeviceToHost, st));
  unsigned long long overflow = 0;
  GPT_CUDA(ctx, cudaMemcpyAsync(&overflow, ctl + 3, sizeof(overflow), cudaMemcpyDeviceToHost, st));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  if (overflow) return gpt_fail(ctx, GPT_E_UNSUPPORTED, "traversal stack overflow (> 64 levels)", __FILE__, __LINE__);
  ctx->stats.rays_traced = (uint64_t)n;
  ctx->stats.kernel_launches = ctx->launches;
  return GPT_OK;
}
