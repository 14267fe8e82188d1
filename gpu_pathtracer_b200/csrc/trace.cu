// trace.cu -- stand-alone closest-hit kernel over a ray queue (gpt_trace*).
//
// One thread per ray, 128-thread CTAs, traversal stack in shared memory laid out
// [level][thread] (bank-conflict free), nodes / triangles fetched as 16-byte words through
// the read-only path.  Rays: 2 x float4 {ox,oy,oz,tmax}{dx,dy,dz,-}; hits: 2 x float4
// {t, leaf, prim, -}{b1, b2, b3, -}.
#include "trace_core.cuh"

namespace {

constexpr int TRACE_BLOCK = 128;

template <bool NORMALIZE, bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_trace(const float4 *__restrict__ nodes, const float4 *__restrict__ tris,
        const int *__restrict__ parent_leaf, const float4 *__restrict__ rays, float4 *hits,
        long long n, unsigned long long *counters) {
  __shared__ int stack[TRACE_STACK * TRACE_BLOCK];
  long long i = (long long)blockIdx.x * TRACE_BLOCK + threadIdx.x;
  if (i >= n) return;
  float4 ro = __ldg(rays + 2 * i), rd = __ldg(rays + 2 * i + 1);
  float dx = rd.x, dy = rd.y, dz = rd.z;
  if (NORMALIZE) {  // Ray::Ray -> unit_vector(dir) = dir / dir.length()  (ray.h:21-24)
    float len = sqrtf(dx * dx + dy * dy + dz * dz);
    dx = dx / len; dy = dy / len; dz = dz / len;
  }
  RaySetup r = make_ray(ro.x, ro.y, ro.z, dx, dy, dz);
  int overflow = 0;
  HitResult h = trace_ray<COUNT>(nodes, tris, parent_leaf, r, ro.w, stack + threadIdx.x,
                                 TRACE_BLOCK, counters, counters + 1, &overflow);
  if (overflow) atomicExch(counters + 3, 1ull);
  hits[2 * i] = make_float4(h.t, __int_as_float(h.leaf), __int_as_float(h.prim), 0.f);
  hits[2 * i + 1] = make_float4(h.b1, h.b2, h.b3, 0.f);
}

}  // namespace

int gpt_trace_launch(gpt_ctx *ctx, int which, const float4 *d_rays, int64_t n, float4 *d_hits,
                     bool normalize, bool count) {
  if (which < 0 || which > 2) return gpt_fail(ctx, GPT_E_INVALID, "bad BVH id", __FILE__, __LINE__);
  BvhDev &b = ctx->bvh[which];
  if (b.n < 2) return gpt_fail(ctx, GPT_E_INVALID, "BVH not built (needs >= 2 leaves)", __FILE__, __LINE__);
  if (n <= 0) return GPT_OK;
  unsigned grid = (unsigned)((n + TRACE_BLOCK - 1) / TRACE_BLOCK);
  unsigned long long *c = ctx->d_counters;
  if (count) GPT_CUDA(ctx, cudaMemsetAsync(c, 0, 3 * sizeof(unsigned long long), ctx->stream));
  if (normalize) {
    if (count) k_trace<true, true><<<grid, TRACE_BLOCK, 0, ctx->stream>>>(b.nodes, b.tris, b.parent_leaf, d_rays, d_hits, n, c);
    else k_trace<true, false><<<grid, TRACE_BLOCK, 0, ctx->stream>>>(b.nodes, b.tris, b.parent_leaf, d_rays, d_hits, n, c);
  } else {
    if (count) k_trace<false, true><<<grid, TRACE_BLOCK, 0, ctx->stream>>>(b.nodes, b.tris, b.parent_leaf, d_rays, d_hits, n, c);
    else k_trace<false, false><<<grid, TRACE_BLOCK, 0, ctx->stream>>>(b.nodes, b.tris, b.parent_leaf, d_rays, d_hits, n, c);
  }
  GPT_LAUNCH_CHECK(ctx);
  return GPT_OK;
}
