// wide.cu -- the 4-wide, 8-bit quantised tree the trace kernels walk (trace_wide.cuh).
//
// This is OUR acceleration structure, not the reference's: the reference's LBVH (lbvh.cu,
// src/model/bvh/bvh_building.h:44-262) is built first, bit for bit, and stays what
// gpt_export_bvh exports and what equal-t ties are resolved on.  The wide tree is collapsed
// from it so that a ray needs fewer and smaller node fetches:
//
//   ncu, r02b (profiles/r02_summary.md): both the binary walk and a 4-wide walk over 128-byte
//   fp32 nodes sit at 66-68 % of the L1 data-pipe wavefront rate with the issue slots half
//   idle, and adding k dummy 16-byte loads to a node visit adds k/7 of the whole kernel time.
//   A lane's 16-byte load of its own node is one L1 wavefront whatever else the warp does, so
//   the trace stage is bound by (16-byte loads per visit) x (visits per ray) and by nothing
//   else: 4 x 49 on the binary LBVH, 7 x 21 on the fp32 4-wide tree, 3 x 24 here.
//
// Node = 48 bytes = three 16-byte words:
//   w0 = (origin[0], origin[1], origin[2], meta)   origin = lower corner of the node's box (fp32)
//        meta = E0 | E1 << 8 | E2 << 16 | ni << 24 | nl << 27 | axis << 30
//        E* = biased exponent of the per-axis grid step 2^e (as in an IEEE float); ni / nl =
//        number of internal / leaf children; axis = the axis slot the children are sorted along
//        (ascending box centre), which is the order a ray visits them in
//   w1 = (qlo[0], qlo[1], qlo[2], qhi[0])           one byte per child slot: bound = origin + q * 2^e
//   w2 = (qhi[1], qhi[2], child_base, leaf_base)
// Child slots: the ni internal children first, then the nl leaves, each group sorted along
// `axis`; internal child c is node child_base + c, leaf child c is triangle leaf_base + c of
// wtris (a copy of the LBVH's 48-byte trace triangles in this order, carrying the LBVH leaf
// index).  Unused slots hold the inverted box (qlo = 255, qhi = 0).
// The tree is stored THREE times, once per dominant ray axis kz, with the axis slots of every
// node permuted to (kz + 1, kz + 2, kz) mod 3: a ray reads the copy of its own kz, so the
// dominant axis -- the only one it may cull on (trace_core.cuh) -- is always slot 2 and the walk
// needs no per-box selects.  Child boxes are quantised OUTWARD (floor / ceil), so
// every wide box contains the exact LBVH box it stands for; trace_wide.cuh adds the margin
// that covers the rounding of its own arithmetic, and filters every candidate hit with the
// reference's exact predicate on the triangle's own box, which is why a looser tree cannot
// change the answer.
//
// Build: breadth-first from the root, one cooperative launch.  A work item is (binary node b,
// wide node w): the thread collapses b's subtree top down -- b's two children, then twice the
// internal child with the larger box area is replaced by its two children -- sorts the (up to)
// four children, reserves consecutive slots for the internal / leaf ones (one atomicAdd each),
// writes node w and appends the internal children to the next level's queue.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int WB_THREADS = 256;
constexpr int WB_MAX_LEVELS = 512;
enum { WC_NODES = 0, WC_LEAVES = 1, WC_ERROR = 2, WC_LEVEL0 = 4 };  // counters; WC_LEVEL0 + L = items of level L

__device__ __forceinline__ float half_area6(const float *b) {
  const float x = b[1] - b[0], y = b[3] - b[2], z = b[5] - b[4];
  return x * y + y * z + z * x;
}

struct WChild {
  float box[6];
  int ref;
};
// order: internal children (ref >= 0, class 0), then leaves (class 1), then unused slots (class
// 2, marked by an infinite box); inside a class by box centre along `axis`
__device__ __forceinline__ int wclass(const WChild &c) {
  return c.box[0] == INFINITY ? 2 : (c.ref >= 0 ? 0 : 1);
}
__device__ __forceinline__ void wswap(WChild &a, WChild &b, int axis) {
  const float ca = a.box[2 * axis] + a.box[2 * axis + 1], cb = b.box[2 * axis] + b.box[2 * axis + 1];
  const int ka = wclass(a), kb = wclass(b);
  if (kb < ka || (kb == ka && cb < ca)) {
    WChild t = a;
    a = b;
    b = t;
  }
}

__device__ void wide_emit(const float4 *__restrict__ nodes, const float4 *__restrict__ tris, int b,
                          int w, int level, int2 *qnext, int *counters, uint4 *wq, float4 *wtris,
                          int stride) {
  WChild ch[4];
  int nc = 2;
  {
    const float4 *q = nodes + 4 * (size_t)b;
    const float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
    ch[0].box[0] = q0.x; ch[0].box[1] = q0.y; ch[0].box[2] = q0.z; ch[0].box[3] = q0.w; ch[0].box[4] = q1.x; ch[0].box[5] = q1.y;
    ch[1].box[0] = q1.z; ch[1].box[1] = q1.w; ch[1].box[2] = q2.x; ch[1].box[3] = q2.y; ch[1].box[4] = q2.z; ch[1].box[5] = q2.w;
    ch[0].ref = __float_as_int(q3.x);
    ch[1].ref = __float_as_int(q3.y);
  }
#pragma unroll 1
  for (int it = 0; it < 2; it++) {
    int pick = -1, pickref = -1;
    float pa = -1.f;
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (k < nc && ch[k].ref >= 0) {
        const float a = half_area6(ch[k].box);
        if (a > pa || pick < 0) { pa = a; pick = k; pickref = ch[k].ref; }
      }
    if (pick < 0) break;
    const float4 *q = nodes + 4 * (size_t)pickref;
    const float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
#pragma unroll
    for (int k = 0; k < 4; k++) {  // compile-time k: the arrays stay in registers
      if (k == pick) {
        ch[k].box[0] = q0.x; ch[k].box[1] = q0.y; ch[k].box[2] = q0.z; ch[k].box[3] = q0.w; ch[k].box[4] = q1.x; ch[k].box[5] = q1.y;
        ch[k].ref = __float_as_int(q3.x);
      }
      if (k == nc) {
        ch[k].box[0] = q1.z; ch[k].box[1] = q1.w; ch[k].box[2] = q2.x; ch[k].box[3] = q2.y; ch[k].box[4] = q2.z; ch[k].box[5] = q2.w;
        ch[k].ref = __float_as_int(q3.y);
      }
    }
    nc++;
  }
  // node box, and the axis along which the children's centres spread most
  float lo[3], hi[3], spread[3];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float cmin = INFINITY, cmax = -INFINITY;
    lo[a] = INFINITY; hi[a] = -INFINITY;
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (k < nc) {
        lo[a] = fminf(lo[a], ch[k].box[2 * a]);
        hi[a] = fmaxf(hi[a], ch[k].box[2 * a + 1]);
        const float c = ch[k].box[2 * a] + ch[k].box[2 * a + 1];
        cmin = fminf(cmin, c); cmax = fmaxf(cmax, c);
      }
    spread[a] = cmax - cmin;
  }
  const int axis = (spread[0] >= spread[1] && spread[0] >= spread[2]) ? 0 : (spread[1] >= spread[2] ? 1 : 2);
  // unused slots sort last: give them an infinite centre
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (k >= nc) {
#pragma unroll
      for (int a = 0; a < 6; a++) ch[k].box[a] = INFINITY;
      ch[k].ref = 0;
    }
  // 4-element sorting network on the centre along `axis` (switch keeps the indices static)
#define WSORT(AX)                                                                   \
  do {                                                                              \
    wswap(ch[0], ch[1], AX); wswap(ch[2], ch[3], AX); wswap(ch[0], ch[2], AX);      \
    wswap(ch[1], ch[3], AX); wswap(ch[1], ch[2], AX);                               \
  } while (0)
  if (axis == 0) WSORT(0); else if (axis == 1) WSORT(1); else WSORT(2);
#undef WSORT
  // per-axis grid step 2^e >= extent / 254, as a biased exponent byte
  unsigned E[3], qlo[3] = {0, 0, 0}, qhi[3] = {0, 0, 0};
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const float ext = hi[a] - lo[a];  // >= 0, finite
    int e = 0;
    frexpf(ext * (1.0f / 254.0f), &e);  // ext / 254 = m * 2^e, m in [0.5, 1)  =>  2^e > ext / 254
    int Eb = e + 127;
    Eb = Eb < 1 ? 1 : (Eb > 253 ? 253 : Eb);
    const float inv_step = __uint_as_float((unsigned)(254 - Eb) << 23);  // 2^-(Eb - 127), exact
    E[a] = (unsigned)Eb;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      unsigned l = 255u, h = 0u;  // unused slot: the inverted box
      if (k < nc) {
        // (bound - origin) carries one rounding (<= 2^-24 relative, i.e. <= 2e-5 grid steps);
        // scaling by a power of two and floor / ceil are exact.  The runtime margin of
        // trace_wide.cuh (> 1 step) covers that rounding.
        const float fl = floorf((ch[k].box[2 * a] - lo[a]) * inv_step);
        const float fh = ceilf((ch[k].box[2 * a + 1] - lo[a]) * inv_step);
        l = (unsigned)fminf(fmaxf(fl, 0.f), 255.f);
        h = (unsigned)fminf(fmaxf(fh, 0.f), 255.f);
      }
      qlo[a] |= l << (8 * k);
      qhi[a] |= h << (8 * k);
    }
  }
  int ni = 0;
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (k < nc && ch[k].ref >= 0) ni++;  // sorted: these are slots 0 .. ni - 1
  const int nl = nc - ni;
  const int cbase = ni ? atomicAdd(counters + WC_NODES, ni) : 0;
  const int lbase = nl ? atomicAdd(counters + WC_LEAVES, nl) : 0;
  const int qpos = ni ? atomicAdd(counters + WC_LEVEL0 + level + 1, ni) : 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    if (k >= nc) continue;
    if (k < ni) {
      qnext[qpos + k] = make_int2(ch[k].ref, cbase + k);
    } else {
      const int leaf = ~ch[k].ref;
      const float4 *tp = tris + 3 * (size_t)leaf;
      float4 t2 = __ldg(tp + 2);
      t2.w = __int_as_float(leaf);
      float4 *o = wtris + 3 * (size_t)(lbase + (k - ni));
      o[0] = __ldg(tp);
      o[1] = __ldg(tp + 1);
      o[2] = t2;
    }
  }
  // three copies, axis slots (kz + 1, kz + 2, kz) mod 3
#pragma unroll
  for (int kz = 0; kz < 3; kz++) {
    const int a0 = (kz + 1) % 3, a1 = (kz + 2) % 3, a2 = kz;
    const int axis_slot = (axis - a0 + 3) % 3;
    const unsigned meta = E[a0] | E[a1] << 8 | E[a2] << 16 | (unsigned)ni << 24 | (unsigned)nl << 27 |
                          (unsigned)axis_slot << 30;
    uint4 *o = wq + 3 * ((size_t)kz * stride + (size_t)w);
    o[0] = make_uint4(__float_as_uint(lo[a0]), __float_as_uint(lo[a1]), __float_as_uint(lo[a2]), meta);
    o[1] = make_uint4(qlo[a0], qlo[a1], qlo[a2], qhi[a0]);
    o[2] = make_uint4(qhi[a1], qhi[a2], (unsigned)cbase, (unsigned)lbase);
  }
}

__global__ void __launch_bounds__(WB_THREADS)
k_wide_build(const float4 *__restrict__ nodes, const float4 *__restrict__ tris,
             const uint32_t *__restrict__ world_enc, int2 *q0, int2 *q1, int *counters, uint4 *wq,
             float4 *wtris, float *wmeta, int stride) {
  cg::grid_group grid = cg::this_grid();
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
  if (gtid == 0) {
    float wmax = 0.f;
    for (int a = 0; a < 6; a++) wmax = fmaxf(wmax, fabsf(dec_f32(world_enc[a])));
    wmeta[0] = wmax;
    wmeta[1] = 8388608.f;  // 2^23 = 0x4b000000: trace_wide.cuh's byte -> float constant
  }
  int2 *qa = q0, *qb = q1;
  for (int level = 0;; level++) {
    const int cnt = *reinterpret_cast<volatile int *>(counters + WC_LEVEL0 + level);
    if (cnt == 0) break;
    if (level + 1 >= WB_MAX_LEVELS) {  // absurdly deep: report, the caller keeps the binary walk
      if (gtid == 0) counters[WC_ERROR] = 1;
      break;
    }
    for (int i = gtid; i < cnt; i += gsize) {
      const int2 it = qa[i];
      wide_emit(nodes, tris, it.x, it.y, level, qb, counters, wq, wtris, stride);
    }
    grid.sync();
    int2 *t = qa; qa = qb; qb = t;
  }
}

}  // namespace

// The wide tree over an already built triangle LBVH.  One cooperative launch; the node count is
// read back (the caller synchronises anyway to time the build).
int gpt_lbvh_build_wide(gpt_ctx *ctx, BvhDev &b) {
  b.wide_nodes = 0;
  if (b.n < 2 || !b.tris) return GPT_OK;
  const size_t N = (size_t)b.n;
  if (b.wide_capacity < b.n) {
    GPT_TRY(gpt_alloc(ctx, &b.wq, 9 * (size_t)b.capacity));  // 3 copies x 3 words x (at most) capacity nodes
    GPT_TRY(gpt_alloc(ctx, &b.wtris, 3 * (size_t)b.capacity));
    GPT_TRY(gpt_alloc(ctx, &b.wmeta, (size_t)4));
    b.wide_capacity = b.capacity;
  }
  int2 *queues = nullptr;
  int *counters = nullptr;
  GPT_TRY(gpt_alloc(ctx, &queues, 2 * N));
  GPT_TRY(gpt_alloc(ctx, &counters, (size_t)(WC_LEVEL0 + WB_MAX_LEVELS + 1)));
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemsetAsync(counters, 0, sizeof(int) * (WC_LEVEL0 + WB_MAX_LEVELS + 1), st));
  {  // root: work item (binary node 0 -> wide node 0), one wide node and one queue entry so far
    const int2 root = make_int2(0, 0);
    const int one = 1;
    GPT_CUDA(ctx, cudaMemcpyAsync(queues, &root, sizeof(root), cudaMemcpyHostToDevice, st));
    GPT_CUDA(ctx, cudaMemcpyAsync(counters + WC_NODES, &one, sizeof(int), cudaMemcpyHostToDevice, st));
    GPT_CUDA(ctx, cudaMemcpyAsync(counters + WC_LEVEL0, &one, sizeof(int), cudaMemcpyHostToDevice, st));
  }
  static int grid_cache[64] = {0};
  int &grid = grid_cache[ctx->device & 63];
  if (!grid) {
    int sms = 148, per_sm = 1;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_wide_build, WB_THREADS, 0);
    grid = sms * (per_sm > 0 ? per_sm : 1);
  }
  const float4 *nodes = b.nodes, *tris = b.tris;
  const uint32_t *we = b.world_enc;
  int2 *q0 = queues, *q1 = queues + N;
  uint4 *wq = b.wq;
  float4 *wtris = b.wtris;
  float *wmeta = b.wmeta;
  int stride = b.wide_capacity;
  void *args[] = {&nodes, &tris, &we, &q0, &q1, &counters, &wq, &wtris, &wmeta, &stride};
  GPT_CUDA(ctx, cudaLaunchCooperativeKernel((const void *)k_wide_build, dim3(grid), dim3(WB_THREADS), args, 0, st));
  ctx->launches++;
  int h[4] = {0, 0, 0, 0};
  GPT_CUDA(ctx, cudaMemcpyAsync(h, counters, sizeof(h), cudaMemcpyDeviceToHost, st));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  gpt_free(ctx, queues);
  gpt_free(ctx, counters);
  // a stack / queue entry of trace_wide.cuh holds a 23-bit node base and a 24-bit leaf base
  if (h[WC_ERROR] == 0 && h[WC_LEAVES] == b.n && h[WC_NODES] < (1 << 23) && b.n < (1 << 24))
    b.wide_nodes = h[WC_NODES];
  return GPT_OK;
}
