// lbvh.cu -- Karras LBVH construction for sm_100a.
//
// Replaces src/main.cu:826-918 (and the transparent / light repeats :930-1197):
//   compute_world_bounding_box  (1 thread, serial)      -> k_world_reduce (grid reduction)
//   compute_morton_code_batch   (1 thread / block)      -> k_morton
//   thrust::stable_sort w/ pointer comparator           -> 8-bit LSD radix sort, warp-match
//                                                          ranking (stable), 4 passes
//   set_node_relationship       (linear scans)          -> k_emit (binary searches)
//   compute_node_bounding_boxes (flags, no fences)      -> k_refit (atomic counters + fence),
//                                                          fused with the 48-byte triangle pack
// Results (codes, permutation, topology, boxes) are bit-identical to the reference: the
// Morton arithmetic is the same fp32 sequence, the sort is stable, delta() carries the
// same index tie-break, and min/max unions are order-independent.
#include "common.cuh"
#include <cuda/atomic>

#include "vecmath.cuh"

namespace {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;                         // keys per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;   // 4096 keys per block
// Short inputs (the path tracer's per-group queues on a tile-sharded frame, 10^4..10^5 slots)
// are latency: with 4096-key tiles 65 536 keys are 16 blocks whose threads walk 16 keys each,
// 10 + 12 us for histogram + scatter.  Up to SORT_SMALL_N keys the tile is 1024 keys.
constexpr int SORT_ITEMS_SMALL = 4;
constexpr int SORT_SMALL_N = 1 << 21;
constexpr int SORT_WARPS = SORT_THREADS / 32;

// ---- world box ---------------------------------------------------------------------
// compute_world_bounding_box (bounding_box_operations.h:12-45): min/max over the leaf boxes.
__global__ void __launch_bounds__(256)
k_world_reduce(const float *__restrict__ aabb, const int *__restrict__ prim_list, int n,
               uint32_t *world_enc) {
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    int p = prim_list ? __ldg(prim_list + i) : i;
    const float2 *b = reinterpret_cast<const float2 *>(aabb + 6 * (size_t)p);
    float2 bx = __ldg(b), by = __ldg(b + 1), bz = __ldg(b + 2);
    mn[0] = fminf(mn[0], bx.x); mx[0] = fmaxf(mx[0], bx.y);
    mn[1] = fminf(mn[1], by.x); mx[1] = fmaxf(mx[1], by.y);
    mn[2] = fminf(mn[2], bz.x); mx[2] = fmaxf(mx[2], bz.y);
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  }
  __shared__ float smn[3][8], smx[3][8];
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0)
    for (int a = 0; a < 3; a++) { smn[a][w] = mn[a]; smx[a][w] = mx[a]; }
  __syncthreads();
  if (threadIdx.x < 3) {
    int a = threadIdx.x;
    float lo = smn[a][0], hi = smx[a][0];
    for (int k = 1; k < 8; k++) { lo = fminf(lo, smn[a][k]); hi = fmaxf(hi, smx[a][k]); }
    atomicMin(world_enc + a, enc_f32(lo));
    atomicMax(world_enc + 3 + a, enc_f32(hi));
  }
}

__global__ void k_world_init(uint32_t *world_enc) {
  if (threadIdx.x < 3) world_enc[threadIdx.x] = 0xffffffffu;
  else if (threadIdx.x < 6) world_enc[threadIdx.x] = 0u;
}

// ---- Morton codes --------------------------------------------------------------------
__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {  // bvh_util.h:4-10
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ uint32_t quant10(float x) {  // bvh_util.h:13-18
  return (uint32_t)fminf(fmaxf(x * 1024.0f, 0.0f), 1023.0f);
}

// compute_morton_code_batch + compute_normalized_center + compute_bb_morton_3d
// (bvh_building.h:44-55, bounding_box.h:126-141,171-173)
__global__ void __launch_bounds__(256)
k_morton(const float *__restrict__ aabb, const int *__restrict__ prim_list, int n,
         const uint32_t *__restrict__ world_enc, uint32_t *keys, int *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // world box = reduced min/max -/+ SMALL_DOUBLE, length = max - min
  // (bounding_box_operations.h:34-43, bounding_box.h:175-177)
  float wmin[3], wlen[3];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float lo = dec_f32(world_enc[a]) - GPT_SMALL, hi = dec_f32(world_enc[3 + a]) + GPT_SMALL;
    wmin[a] = lo;
    wlen[a] = hi - lo;
  }
  int p = prim_list ? __ldg(prim_list + i) : i;
  const float2 *b = reinterpret_cast<const float2 *>(aabb + 6 * (size_t)p);
  float2 bx = __ldg(b), by = __ldg(b + 1), bz = __ldg(b + 2);
  float cx = 0.5f * (bx.x + bx.y), cy = 0.5f * (by.x + by.y), cz = 0.5f * (bz.x + bz.y);
  float nx = (cx - wmin[0]) / wlen[0], ny = (cy - wmin[1]) / wlen[1],
        nz = (cz - wmin[2]) / wlen[2];
  uint32_t code = expand_bits10(quant10(nx)) * 4u + expand_bits10(quant10(ny)) * 2u +
                  expand_bits10(quant10(nz));
  keys[i] = code;
  vals[i] = i;
}

// ---- stable LSD radix sort -------------------------------------------------------------
// Pass = k_sort_hist (per-block digit counts, digit-major) -> k_sort_scan_rows (exclusive scan per digit)
//        -> k_sort_scatter (stable ranks via __match_any_sync + per-warp running counters).
template <int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS)
k_sort_hist(const uint32_t *__restrict__ keys, int n, int shift, uint32_t mask,
            uint32_t *__restrict__ block_hist, int nblocks, const int *__restrict__ d_n) {
  __shared__ uint32_t h[256];
  if (d_n) n = min(n, __ldg(d_n));  // n is only an upper bound: the exact count is on the device
  h[threadIdx.x] = 0;
  __syncthreads();
  int base = blockIdx.x * (SORT_THREADS * ITEMS);
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    int i = base + k * SORT_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[(__ldg(keys + i) >> shift) & mask], 1u);
  }
  __syncthreads();
  block_hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// block-wide exclusive scan of one value per thread (SORT_THREADS threads); returns the
// exclusive prefix, *total = the sum over the block
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *wsum, uint32_t *total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) wsum[w] = x;
  __syncthreads();
  uint32_t base = 0, tot = 0;
#pragma unroll
  for (int k = 0; k < SORT_WARPS; k++) {
    const uint32_t c = wsum[k];
    if (k < w) base += c;
    tot += c;
  }
  __syncthreads();
  *total = tot;
  return base + x - v;
}

// One block per digit: exclusive scan of that digit's row of per-block counts (digit-major
// layout) in place, the row's sum into totals[digit].  The bases of the digits themselves -- a
// 256-entry scan -- are recomputed by every scatter block (r02: the single-block scan over all
// digits x blocks counters was 18 % of the 1 M-triangle build, 3 x 21 us).
__global__ void __launch_bounds__(SORT_THREADS)
k_sort_scan_rows(uint32_t *block_hist, int nblocks, uint32_t *totals) {
  __shared__ uint32_t wsum[SORT_WARPS];
  uint32_t *row = block_hist + (size_t)blockIdx.x * nblocks;
  uint32_t carry = 0;
  for (int base = 0; base < nblocks; base += SORT_THREADS) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < nblocks ? row[i] : 0u;
    uint32_t tot;
    const uint32_t ex = block_exclusive_scan(v, wsum, &tot);
    if (i < nblocks) row[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

template <int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS)
k_sort_scatter(const uint32_t *__restrict__ keys, const int *__restrict__ vals, int n,
               int shift, uint32_t mask, const uint32_t *__restrict__ block_hist, int nblocks,
               uint32_t *__restrict__ keys_out, int *__restrict__ vals_out,
               const int *__restrict__ d_n, const uint32_t *__restrict__ totals) {
  __shared__ uint32_t cnt[SORT_WARPS][256];  // running / total count per warp and digit
  __shared__ uint32_t wsum[SORT_WARPS];
  if (d_n) n = min(n, __ldg(d_n));
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  for (int d = threadIdx.x; d < SORT_WARPS * 256; d += SORT_THREADS) (&cnt[0][0])[d] = 0;
  __syncthreads();
  // each warp owns a contiguous chunk of the tile, read 32 keys at a time in order
  int wbase = blockIdx.x * (SORT_THREADS * ITEMS) + w * (32 * ITEMS);
  uint32_t key[ITEMS];
  uint16_t rank[ITEMS];
  uint32_t lt = (1u << l) - 1u;
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    int i = wbase + k * 32 + l;
    bool ok = i < n;
    key[k] = ok ? __ldg(keys + i) : 0xffffffffu;
    uint32_t d = ok ? ((key[k] >> shift) & mask) : 256u;  // 256 = "no element"
    uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t before = 0;
    if (ok) before = cnt[w][d];
    __syncwarp();
    rank[k] = (uint16_t)(before + __popc(peers & lt));
    if (ok && (peers & lt) == 0) cnt[w][d] = before + __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  {  // exclusive scan over warps for digit d = threadIdx.x, plus the global base = keys with
     // a smaller digit (scan of the digit totals) + keys of this digit in earlier blocks
    int d = threadIdx.x;
    uint32_t tot_all;
    const uint32_t digit_base = block_exclusive_scan((uint32_t)d <= mask ? __ldg(totals + d) : 0u, wsum, &tot_all);
    uint32_t run = digit_base + ((uint32_t)d <= mask ? __ldg(block_hist + (size_t)d * nblocks + blockIdx.x) : 0u);
#pragma unroll
    for (int k = 0; k < SORT_WARPS; k++) {
      uint32_t c = cnt[k][d];
      cnt[k][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    int i = wbase + k * 32 + l;
    if (i < n) {
      uint32_t d = (key[k] >> shift) & mask;
      uint32_t dst = cnt[w][d] + rank[k];
      keys_out[dst] = key[k];
      vals_out[dst] = __ldg(vals + i);
    }
  }
}

// ---- Karras hierarchy ---------------------------------------------------------------------
// length_longest_common_prefix (bvh_util.h:22-31): clz(k_i ^ k_j), plus clz(i ^ j) when the
// keys are equal, -1 out of range.
__device__ __forceinline__ int delta(const uint32_t *__restrict__ keys, int n, int i,
                                     uint32_t ki, int j) {
  if (j < 0 || j >= n) return -1;
  uint32_t kj = __ldg(keys + j);
  int extra = (ki == kj) ? __clz(i ^ j) : 0;
  return __clz(ki ^ kj) + extra;
}

// set_node_relationship (bvh_building.h:88-185) with the linear walks replaced by the
// doubling + binary searches of Karras 2012; same direction rule (d1 < d2 -> -1), same
// delta_min, same unique split, same child / parent assignment.
__global__ void __launch_bounds__(256)
k_emit(const uint32_t *__restrict__ keys, int n, float4 *nodes, int *parent_leaf, int *range_last) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  uint32_t ki = __ldg(keys + i);
  int d = (delta(keys, n, i, ki, i + 1) < delta(keys, n, i, ki, i - 1)) ? -1 : 1;
  int dmin = delta(keys, n, i, ki, i - d);
  int lmax = 2;
  while (delta(keys, n, i, ki, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (delta(keys, n, i, ki, i + (l + t) * d) > dmin) l += t;
  int j = i + l * d;
  int dnode = delta(keys, n, i, ki, j);
  int s = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (delta(keys, n, i, ki, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  int gamma = i + s * d + min(d, 0);
  int first = min(i, j), last = max(i, j);
  int left = (first == gamma) ? ~gamma : gamma;
  int right = (last == gamma + 1) ? ~(gamma + 1) : gamma + 1;
  int *q3 = reinterpret_cast<int *>(nodes + 4 * (size_t)i + 3);
  q3[0] = left;
  q3[1] = right;
  q3[3] = first;
  range_last[i] = last;  // for k_node_boxes
  if (i == 0) q3[2] = -1;
  if (left >= 0) reinterpret_cast<int *>(nodes + 4 * (size_t)left + 3)[2] = 2 * i;
  else parent_leaf[~left] = 2 * i;
  if (right >= 0) reinterpret_cast<int *>(nodes + 4 * (size_t)right + 3)[2] = 2 * i + 1;
  else parent_leaf[~right] = 2 * i + 1;
}

// compute_node_bounding_boxes (bvh_building.h:221-262): bottom-up union.  Each child writes
// its box into its slot of the parent's 64-byte node; the second arrival (atomic counter,
// fenced) reads the sibling slot and climbs.  Fused with the pack of the 48-byte trace
// triangle for triangle BVHs (TRIS) -- the leaf thread already owns the primitive.
template <bool TRIS>
__global__ void __launch_bounds__(256)
k_refit(const float *__restrict__ aabb, const int *__restrict__ prim_list,
        const int *__restrict__ sorted_prim, int n, float4 *nodes,
        const int *__restrict__ parent_leaf, int *flags, SceneDev s, float4 *tris) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  int sp = __ldg(sorted_prim + j);
  int p = prim_list ? __ldg(prim_list + sp) : sp;
  const float2 *b = reinterpret_cast<const float2 *>(aabb + 6 * (size_t)p);
  float2 bx = __ldg(b), by = __ldg(b + 1), bz = __ldg(b + 2);
  if (TRIS) {
    int a = __ldg(s.p1 + p), bb = __ldg(s.p2 + p), c = __ldg(s.p3 + p);
    float4 *t = tris + 3 * (size_t)j;
    t[0] = make_float4(__ldg(s.x + a), __ldg(s.y + a), __ldg(s.z + a), __ldg(s.x + bb));
    t[1] = make_float4(__ldg(s.y + bb), __ldg(s.z + bb), __ldg(s.x + c), __ldg(s.y + c));
    t[2] = make_float4(__ldg(s.z + c), __ldg(s.tolerance + p), __int_as_float(p), 0.f);
  }
  int pe = __ldg(parent_leaf + j);
  while (pe >= 0) {
    int par = pe >> 1, right = pe & 1;
    float *nf = reinterpret_cast<float *>(nodes + 4 * (size_t)par);
    float *slot = nf + (right ? 6 : 0);
    // own slot: (xmin,xmax,ymin,ymax,zmin,zmax)
    reinterpret_cast<float2 *>(slot)[0] = bx;
    reinterpret_cast<float2 *>(slot)[1] = by;
    reinterpret_cast<float2 *>(slot)[2] = bz;
    // one acq_rel RMW instead of fence.sc + atom + fence.sc (r02: the refit is a chain of
    // ~40 such hand-overs for the top of the tree): release publishes this child's slot,
    // acquire makes the sibling's slot visible to the second arrival
    cuda::atomic_ref<int, cuda::thread_scope_device> arrived(flags[par]);
    if (arrived.fetch_add(1, cuda::std::memory_order_acq_rel) == 0) return;  // first arrival: the sibling will climb
    const float2 *sib = reinterpret_cast<const float2 *>(nf + (right ? 0 : 6));
    float2 sx = __ldcg(sib), sy = __ldcg(sib + 1), sz = __ldcg(sib + 2);
    bx = make_float2(fminf(bx.x, sx.x), fmaxf(bx.y, sx.y));
    by = make_float2(fminf(by.x, sy.x), fmaxf(by.y, sy.y));
    bz = make_float2(fminf(bz.x, sz.x), fmaxf(bz.y, sz.y));
    pe = __ldcg(reinterpret_cast<const int *>(nodes + 4 * (size_t)par + 3) + 2);
  }
}


// ---- node boxes without the bottom-up chain (r02) ---------------------------------------------
// k_refit above is a chain: the box of a node needs both children, so the top ~40 levels of a
// 1 M-leaf tree are ~40 dependent (fence, atomic, load) hand-overs, 155 us of a 0.37 ms build.
// An LBVH node covers a contiguous range [first, last] of the SORTED leaves, and a box is a
// min / max -- associative, commutative, exact -- so the box of a range can be taken from ANY
// set of partial unions that tile it.  k_leaf_levels builds a 4-ary hierarchy of partial unions
// over the sorted leaves (level k = unions of 4^k consecutive leaves; levels 0..5 inside one
// 1024-leaf block with shuffles, k_top_levels adds the few entries above), and k_node_boxes
// gives every internal node its two child boxes with the usual bottom-up range query (at most
// 3 + 3 entries per level): every node independent, no atomics, no fences.  Same boxes bit for
// bit (tests/test_gpu_build_trace.py::test_bvh_bit_exact compares them with the reference's).
constexpr int LV_BLOCK = 1024;  // leaves per block of k_leaf_levels = 4^5
constexpr int LV_MAX = 17;      // 4^16 > 2^31 leaves
struct Levels {
  float *base;           // all levels back to back, 6 floats per entry: xmin xmax ymin ymax zmin zmax
  long long off[LV_MAX]; // first float of level k
  int count[LV_MAX];     // entries of level k = ceil(n / 4^k)
  int nlevels;           // levels 0 .. nlevels - 1; the last one has a single entry
};
struct Box6 { float v[6]; };
__device__ __forceinline__ Box6 box_empty() {
  Box6 b;
#pragma unroll
  for (int k = 0; k < 3; k++) { b.v[2 * k] = INFINITY; b.v[2 * k + 1] = -INFINITY; }
  return b;
}
__device__ __forceinline__ void box_add(Box6 &a, const float *p) {
  const float2 x = *reinterpret_cast<const float2 *>(p), y = *reinterpret_cast<const float2 *>(p + 2),
               z = *reinterpret_cast<const float2 *>(p + 4);
  a.v[0] = fminf(a.v[0], x.x); a.v[1] = fmaxf(a.v[1], x.y);
  a.v[2] = fminf(a.v[2], y.x); a.v[3] = fmaxf(a.v[3], y.y);
  a.v[4] = fminf(a.v[4], z.x); a.v[5] = fmaxf(a.v[5], z.y);
}
__device__ __forceinline__ void box_store(float *p, const Box6 &a) {
  reinterpret_cast<float2 *>(p)[0] = make_float2(a.v[0], a.v[1]);
  reinterpret_cast<float2 *>(p)[1] = make_float2(a.v[2], a.v[3]);
  reinterpret_cast<float2 *>(p)[2] = make_float2(a.v[4], a.v[5]);
}
__device__ __forceinline__ void box_shfl_xor(Box6 &a, int m) {
#pragma unroll
  for (int k = 0; k < 3; k++) {
    a.v[2 * k] = fminf(a.v[2 * k], __shfl_xor_sync(0xffffffffu, a.v[2 * k], m));
    a.v[2 * k + 1] = fmaxf(a.v[2 * k + 1], __shfl_xor_sync(0xffffffffu, a.v[2 * k + 1], m));
  }
}

// One thread per sorted leaf: its box into level 0, the unions of 4 / 16 / 64 / 256 / 1024
// leaves into levels 1..5; fused with the pack of the 48-byte trace triangle (TRIS).
template <bool TRIS>
__global__ void __launch_bounds__(LV_BLOCK)
k_leaf_levels(const float *__restrict__ aabb, const int *__restrict__ prim_list,
              const int *__restrict__ sorted_prim, int n, Levels lv, SceneDev s, float4 *tris) {
  __shared__ float wbox[LV_BLOCK / 32][6];
  const int j = blockIdx.x * LV_BLOCK + threadIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Box6 b = box_empty();
  if (j < n) {
    const int sp = __ldg(sorted_prim + j);
    const int p = prim_list ? __ldg(prim_list + sp) : sp;
    box_add(b, aabb + 6 * (size_t)p);
    box_store(lv.base + lv.off[0] + 6 * (size_t)j, b);
    if (TRIS) {
      const int a = __ldg(s.p1 + p), bb = __ldg(s.p2 + p), c = __ldg(s.p3 + p);
      float4 *t = tris + 3 * (size_t)j;
      t[0] = make_float4(__ldg(s.x + a), __ldg(s.y + a), __ldg(s.z + a), __ldg(s.x + bb));
      t[1] = make_float4(__ldg(s.y + bb), __ldg(s.z + bb), __ldg(s.x + c), __ldg(s.y + c));
      t[2] = make_float4(__ldg(s.z + c), __ldg(s.tolerance + p), __int_as_float(p), 0.f);
    }
  }
  box_shfl_xor(b, 1); box_shfl_xor(b, 2);
  if ((lane & 3) == 0 && j < n && lv.nlevels > 1) box_store(lv.base + lv.off[1] + 6 * (size_t)(j >> 2), b);
  box_shfl_xor(b, 4); box_shfl_xor(b, 8);
  if ((lane & 15) == 0 && j < n && lv.nlevels > 2) box_store(lv.base + lv.off[2] + 6 * (size_t)(j >> 4), b);
  box_shfl_xor(b, 16);
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < 6; k++) wbox[w][k] = b.v[k];
  }
  __syncthreads();
  // levels 3 (2 warps), 4 (8 warps), 5 (the block): a handful of entries, thread t takes one
  const int t = threadIdx.x;
  const int base_leaf = blockIdx.x * LV_BLOCK;
  if (t < 16 + 4 + 1) {
    const int level = t < 16 ? 3 : (t < 20 ? 4 : 5);
    const int e = t < 16 ? t : (t < 20 ? t - 16 : 0);           // entry inside the block
    const int span = level == 3 ? 2 : (level == 4 ? 8 : 32);    // warps per entry
    const int first_leaf = base_leaf + e * span * 32;
    if (first_leaf < n && level < lv.nlevels) {
      Box6 u = box_empty();
      for (int k = 0; k < span; k++) box_add(u, wbox[e * span + k]);
      box_store(lv.base + lv.off[level] + 6 * (size_t)(first_leaf >> (2 * level)), u);
    }
  }
}

// levels 6 and up: a few hundred entries at most, one block, level by level
__global__ void __launch_bounds__(1024) k_top_levels(Levels lv) {
  for (int level = 6; level < lv.nlevels; level++) {
    const float *below = lv.base + lv.off[level - 1];
    float *mine = lv.base + lv.off[level];
    const int nb = lv.count[level - 1];
    for (int e = threadIdx.x; e < lv.count[level]; e += blockDim.x) {
      Box6 u = box_empty();
      for (int k = 0; k < 4; k++)
        if (4 * e + k < nb) box_add(u, below + 6 * (size_t)(4 * e + k));
      box_store(mine + 6 * (size_t)e, u);
    }
    __syncthreads();
  }
}

// union of the sorted leaves [a, b]: bottom-up walk of the 4-ary hierarchy
__device__ __forceinline__ Box6 range_box(const Levels &lv, int a, int b) {
  Box6 u = box_empty();
  int lo = a, hi = b + 1, level = 0;
  while (lo < hi) {
    const float *L = lv.base + lv.off[level];
    while (lo < hi && (lo & 3)) { box_add(u, L + 6 * (size_t)lo); lo++; }
    while (lo < hi && (hi & 3)) { hi--; box_add(u, L + 6 * (size_t)hi); }
    lo >>= 2; hi >>= 2; level++;
  }
  return u;
}

// compute_node_bounding_boxes (bvh_building.h:221-262) for every internal node at once: its
// children cover [first, gamma] and [gamma + 1, last]
__global__ void __launch_bounds__(256)
k_node_boxes(float4 *nodes, const int *__restrict__ range_last, int n, Levels lv) {
  // one thread per CHILD (2 i = left of node i, 2 i + 1 = right): half the work per thread and
  // twice the threads of a thread per node; a child's box is 6 floats = 3 float2 of the node
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = c >> 1, right = c & 1;
  if (i >= n - 1) return;
  const int4 q3 = *reinterpret_cast<const int4 *>(nodes + 4 * (size_t)i + 3);
  const int gamma = q3.x >= 0 ? q3.x : ~q3.x;
  const int a = right ? gamma + 1 : q3.w, b = right ? __ldg(range_last + i) : gamma;
  const Box6 u = range_box(lv, a, b);
  float2 *slot = reinterpret_cast<float2 *>(nodes + 4 * (size_t)i) + (right ? 3 : 0);
  slot[0] = make_float2(u.v[0], u.v[1]);
  slot[1] = make_float2(u.v[2], u.v[3]);
  slot[2] = make_float2(u.v[4], u.v[5]);
}

// flagged-subset compaction in index order (collect_*_geom, 1 thread serial in the reference:
// transparent_geom_operations.h:28-40, pathtracing_target_geom_operations.h:36-48).
// NOTE the reference collects from the Morton-SORTED main list (src/main.cu:952,1083).
__global__ void __launch_bounds__(256)
k_subset_flags(const float4 *__restrict__ shade, const int *__restrict__ sorted_main, int n,
               int mask, uint32_t *keys, int *vals) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  int p = __ldg(sorted_main + j);
  int flags = __float_as_int(__ldg(&shade[7 * (size_t)p + 6].z));
  keys[j] = (flags & mask) ? 0u : 1u;  // stable partition: flagged first, order kept
  vals[j] = p;
}
__global__ void k_count_zero_keys(const uint32_t *keys, int n, int *count) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  // keys are sorted: the boundary 0 -> 1 gives the count
  if (keys[j] == 0u && (j == n - 1 || keys[j + 1] != 0u)) *count = j + 1;
}

__global__ void __launch_bounds__(256)
k_export(const float4 *__restrict__ nodes, const int *__restrict__ parent_leaf,
         const int *__restrict__ sorted_prim, const int *__restrict__ prim_list,
         const float *__restrict__ aabb, int n, int *sorted_to_orig, int *left, int *right,
         int *parent_internal, int *parent_leaf_out, float *node_aabb, float *leaf_aabb) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    int sp = sorted_prim[i];
    int p = prim_list ? prim_list[sp] : sp;
    sorted_to_orig[i] = p;
    parent_leaf_out[i] = parent_leaf[i] >> 1;
    for (int k = 0; k < 6; k++) leaf_aabb[6 * (size_t)i + k] = aabb[6 * (size_t)p + k];
  }
  if (i < n - 1) {
    const int *q3 = reinterpret_cast<const int *>(nodes + 4 * (size_t)i + 3);
    left[i] = q3[0];
    right[i] = q3[1];
    parent_internal[i] = q3[2] < 0 ? -1 : (q3[2] >> 1);
    const float *f = reinterpret_cast<const float *>(nodes + 4 * (size_t)i);
    for (int k = 0; k < 3; k++) {  // node box = union of its two child slots
      node_aabb[6 * (size_t)i + 2 * k] = fminf(f[2 * k], f[6 + 2 * k]);
      node_aabb[6 * (size_t)i + 2 * k + 1] = fmaxf(f[2 * k + 1], f[6 + 2 * k + 1]);
    }
  }
}

}  // namespace

static int sort_tile_of(int n) { return SORT_THREADS * (n <= SORT_SMALL_N ? SORT_ITEMS_SMALL : SORT_ITEMS); }

// counters a sort of up to n keys needs -- and of ANY shorter input (the path tracer sorts its
// shrinking queues with the buffer of the first round), whichever tile that one uses
int gpt_radix_hist_entries(int n) {
  int small_n = n < SORT_SMALL_N ? n : SORT_SMALL_N;
  int a = (small_n + SORT_THREADS * SORT_ITEMS_SMALL - 1) / (SORT_THREADS * SORT_ITEMS_SMALL);
  int b = (n + SORT_TILE - 1) / SORT_TILE;
  int nblocks = a > b ? a : b;
  return 256 * (nblocks > 0 ? nblocks : 1) + 256;  // + the 256 digit totals behind the rows
}

int gpt_radix_sort_pairs(gpt_ctx *ctx, uint32_t *keys, int *vals, uint32_t *keys_alt,
                         int *vals_alt, uint32_t *block_hist, int n, int nbits,
                         uint32_t **keys_out, int **vals_out, const int *d_n) {
  const int tile = sort_tile_of(n);
  const bool small = tile != SORT_TILE;
  int nblocks = (n + tile - 1) / tile;
  if (nblocks < 1) nblocks = 1;
  uint32_t *ka = keys, *kb = keys_alt;
  int *va = vals, *vb = vals_alt;
  for (int shift = 0; shift < nbits; shift += 8) {
    int bits = nbits - shift < 8 ? nbits - shift : 8;
    uint32_t mask = (1u << bits) - 1u;
    if (small) k_sort_hist<SORT_ITEMS_SMALL><<<nblocks, SORT_THREADS, 0, ctx->stream>>>(ka, n, shift, mask, block_hist, nblocks, d_n);
    else k_sort_hist<SORT_ITEMS><<<nblocks, SORT_THREADS, 0, ctx->stream>>>(ka, n, shift, mask, block_hist, nblocks, d_n);
    GPT_LAUNCH_CHECK(ctx);
    // digit-major layout: one block per digit this pass can produce scans its row; the digit
    // totals live behind the 256 rows (gpt_radix_hist_entries reserves them)
    uint32_t *totals = block_hist + (size_t)256 * nblocks;
    k_sort_scan_rows<<<(int)(mask + 1), SORT_THREADS, 0, ctx->stream>>>(block_hist, nblocks, totals);
    GPT_LAUNCH_CHECK(ctx);
    if (small) k_sort_scatter<SORT_ITEMS_SMALL><<<nblocks, SORT_THREADS, 0, ctx->stream>>>(ka, va, n, shift, mask, block_hist, nblocks, kb, vb, d_n, totals);
    else k_sort_scatter<SORT_ITEMS><<<nblocks, SORT_THREADS, 0, ctx->stream>>>(ka, va, n, shift, mask, block_hist, nblocks, kb, vb, d_n, totals);
    GPT_LAUNCH_CHECK(ctx);
    uint32_t *tk = ka; ka = kb; kb = tk;
    int *tv = va; va = vb; vb = tv;
  }
  *keys_out = ka;
  *vals_out = va;
  return GPT_OK;
}

void gpt_lbvh_free(gpt_ctx *ctx, BvhDev &b) {
  gpt_free(ctx, b.nodes); gpt_free(ctx, b.tris); gpt_free(ctx, b.parent_leaf); gpt_free(ctx, b.morton);
  gpt_free(ctx, b.sorted_prim); gpt_free(ctx, b.prim_list); gpt_free(ctx, b.world_enc);
  gpt_free(ctx, b.keys_alt); gpt_free(ctx, b.vals_alt); gpt_free(ctx, b.block_hist); gpt_free(ctx, b.flags); gpt_free(ctx, b.levels);
  gpt_free(ctx, b.wq); gpt_free(ctx, b.wtris); gpt_free(ctx, b.wmeta);
  b = BvhDev();
}

static int lbvh_reserve(gpt_ctx *ctx, BvhDev &b, int n, bool with_tris) {
  if (n <= b.capacity && (!with_tris || b.tris)) return GPT_OK;
  if (n < b.capacity) n = b.capacity;
  int *keep = b.prim_list;
  b.prim_list = nullptr;
  gpt_lbvh_free(ctx, b);
  b.prim_list = keep;
  size_t N = (size_t)n;
  GPT_TRY(gpt_alloc(ctx, &b.nodes, 4 * N));
  if (with_tris) GPT_TRY(gpt_alloc(ctx, &b.tris, 3 * N));
  GPT_TRY(gpt_alloc(ctx, &b.parent_leaf, N));
  GPT_TRY(gpt_alloc(ctx, &b.morton, N));
  GPT_TRY(gpt_alloc(ctx, &b.sorted_prim, N));
  GPT_TRY(gpt_alloc(ctx, &b.keys_alt, N));
  GPT_TRY(gpt_alloc(ctx, &b.vals_alt, N));
  GPT_TRY(gpt_alloc(ctx, &b.block_hist, (size_t)gpt_radix_hist_entries(n)));
  GPT_TRY(gpt_alloc(ctx, &b.flags, N));
  GPT_TRY(gpt_alloc(ctx, &b.levels, 8 * N + 6 * (size_t)LV_MAX));  // 6 N (1 + 1/4 + 1/16 + ...) floats, rounded up per level
  GPT_TRY(gpt_alloc(ctx, &b.world_enc, (size_t)8));
  b.capacity = n;
  return GPT_OK;
}

// Build the LBVH over `n` boxes (aabb: 6 floats each, indexed through d_prim_list when not
// NULL).  with_tris also packs the 48-byte trace triangles (triangle BVHs); point BVHs (SPPM
// photons, SSS samples) pass false.  All work is enqueued on ctx->stream; nothing
// synchronises with the host.
__global__ void __launch_bounds__(256)
k_copy_keys(const uint32_t *__restrict__ src, int n, uint32_t *keys, int *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = src[i];
  vals[i] = i;
}

int gpt_lbvh_build_boxes(gpt_ctx *ctx, BvhDev &b, const float *aabb, const int *d_prim_list,
                         int n, bool with_tris, const uint32_t *d_keys) {
  if (n < 2) {  // the reference cannot build or traverse such a tree (note 9)
    b.n = n;
    return GPT_OK;
  }
  GPT_TRY(lbvh_reserve(ctx, b, n, with_tris));
  b.n = n;
  const SceneDev &s = ctx->scene;
  cudaStream_t st = ctx->stream;
  int grid = (n + 255) / 256;
  k_world_init<<<1, 32, 0, st>>>(b.world_enc);
  GPT_LAUNCH_CHECK(ctx);
  int rgrid = grid < 148 * 8 ? grid : 148 * 8;
  k_world_reduce<<<rgrid, 256, 0, st>>>(aabb, d_prim_list, n, b.world_enc);
  GPT_LAUNCH_CHECK(ctx);
  // keys/vals ping-pong between (morton, sorted_prim) and (keys_alt, vals_alt); 30 bits = 4
  // passes, an even number, so the result lands back in (morton, sorted_prim).
  if (d_keys)  // caller-computed codes (the SSS point clouds normalise against their own box)
    k_copy_keys<<<grid, 256, 0, st>>>(d_keys, n, b.morton, b.sorted_prim);
  else
    k_morton<<<grid, 256, 0, st>>>(aabb, d_prim_list, n, b.world_enc, b.morton, b.sorted_prim);
  GPT_LAUNCH_CHECK(ctx);
  uint32_t *ko; int *vo;
  GPT_TRY(gpt_radix_sort_pairs(ctx, b.morton, b.sorted_prim, b.keys_alt, b.vals_alt,
                               b.block_hist, n, 30, &ko, &vo));
  if (ko != b.morton) return gpt_fail(ctx, GPT_E_INVALID, "sort parity", __FILE__, __LINE__);
  k_emit<<<grid, 256, 0, st>>>(b.morton, n, b.nodes, b.parent_leaf, b.flags);
  GPT_LAUNCH_CHECK(ctx);
  static int chain_refit = -1;  // GPT_LBVH_CHAIN_REFIT=1: the round-1 bottom-up refit (A/B)
  if (chain_refit < 0) { const char *e = getenv("GPT_LBVH_CHAIN_REFIT"); chain_refit = e ? atoi(e) : 0; }
  if (chain_refit) {
    GPT_CUDA(ctx, cudaMemsetAsync(b.flags, 0, sizeof(int) * (size_t)n, st));
    if (with_tris)
      k_refit<true><<<grid, 256, 0, st>>>(aabb, d_prim_list, b.sorted_prim, n, b.nodes,
                                          b.parent_leaf, b.flags, s, b.tris);
    else
      k_refit<false><<<grid, 256, 0, st>>>(aabb, d_prim_list, b.sorted_prim, n, b.nodes,
                                           b.parent_leaf, b.flags, s, nullptr);
    GPT_LAUNCH_CHECK(ctx);
    return GPT_OK;
  }
  Levels lv;
  lv.base = b.levels;
  long long off = 0;
  int cnt = n, nl = 0;
  for (;;) {
    lv.off[nl] = off;
    lv.count[nl] = cnt;
    off += 6LL * cnt;
    nl++;
    if (cnt == 1 || nl == LV_MAX) break;
    cnt = (cnt + 3) / 4;
  }
  lv.nlevels = nl;
  for (int k = nl; k < LV_MAX; k++) { lv.off[k] = 0; lv.count[k] = 0; }
  const int lgrid = (n + LV_BLOCK - 1) / LV_BLOCK;
  if (with_tris) k_leaf_levels<true><<<lgrid, LV_BLOCK, 0, st>>>(aabb, d_prim_list, b.sorted_prim, n, lv, s, b.tris);
  else k_leaf_levels<false><<<lgrid, LV_BLOCK, 0, st>>>(aabb, d_prim_list, b.sorted_prim, n, lv, s, nullptr);
  GPT_LAUNCH_CHECK(ctx);
  if (nl > 6) {
    k_top_levels<<<1, 1024, 0, st>>>(lv);
    GPT_LAUNCH_CHECK(ctx);
  }
  k_node_boxes<<<(int)((2LL * n + 255) / 256), 256, 0, st>>>(b.nodes, b.flags, n, lv);
  GPT_LAUNCH_CHECK(ctx);
  return GPT_OK;
}

int gpt_lbvh_build(gpt_ctx *ctx, int which, const int *d_prim_list, int n) {
  GPT_TRY(gpt_lbvh_build_boxes(ctx, ctx->bvh[which], ctx->scene.aabb, d_prim_list, n, true, nullptr));
  // the light tree is only ever walked by the light sampler (lighttree.cuh), never traced
  if (which != GPT_BVH_LIGHT) GPT_TRY(gpt_lbvh_build_wide(ctx, ctx->bvh[which]));
  return GPT_OK;
}

// Ordered list of the triangles whose flags & mask != 0, taken in MAIN-BVH sorted order.
int gpt_collect_subset(gpt_ctx *ctx, int flag_mask, int **d_list, int *count) {
  BvhDev &m = ctx->bvh[GPT_BVH_MAIN];
  int n = m.n;
  uint32_t *keys = nullptr, *keys_alt = nullptr, *hist = nullptr, *ko = nullptr;
  int *vals = nullptr, *vals_alt = nullptr, *vo = nullptr, *d_count = nullptr;
  GPT_TRY(gpt_alloc(ctx, &keys, (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &keys_alt, (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &vals, (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &vals_alt, (size_t)n));
  GPT_TRY(gpt_alloc(ctx, &hist, (size_t)gpt_radix_hist_entries(n)));
  GPT_TRY(gpt_alloc(ctx, &d_count, (size_t)1));
  GPT_CUDA(ctx, cudaMemsetAsync(d_count, 0, sizeof(int), ctx->stream));
  int grid = (n + 255) / 256;
  k_subset_flags<<<grid, 256, 0, ctx->stream>>>(ctx->scene.shade, m.sorted_prim, n, flag_mask,
                                               keys, vals);
  GPT_LAUNCH_CHECK(ctx);
  GPT_TRY(gpt_radix_sort_pairs(ctx, keys, vals, keys_alt, vals_alt, hist, n, 1, &ko, &vo));
  k_count_zero_keys<<<grid, 256, 0, ctx->stream>>>(ko, n, d_count);
  GPT_LAUNCH_CHECK(ctx);
  int c = 0;
  GPT_CUDA(ctx, cudaMemcpyAsync(&c, d_count, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *count = c;
  if (*d_list) { gpt_free(ctx, *d_list); *d_list = nullptr; }
  GPT_TRY(gpt_alloc(ctx, d_list, (size_t)(c > 0 ? c : 1)));
  if (c > 0)
    GPT_CUDA(ctx, cudaMemcpyAsync(*d_list, vo, sizeof(int) * (size_t)c,
                                  cudaMemcpyDeviceToDevice, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  gpt_free(ctx, keys); gpt_free(ctx, keys_alt); gpt_free(ctx, vals); gpt_free(ctx, vals_alt); gpt_free(ctx, hist);
  gpt_free(ctx, d_count);
  return GPT_OK;
}

int gpt_lbvh_export(gpt_ctx *ctx, int which, uint32_t *morton, int32_t *sorted_to_orig,
                    int32_t *left, int32_t *right, int32_t *parent_internal,
                    int32_t *parent_leaf, float *node_aabb, float *leaf_aabb,
                    float *world_aabb) {
  BvhDev &b = ctx->bvh[which];
  int n = b.n;
  if (n < 2) return gpt_fail(ctx, GPT_E_INVALID, "BVH has fewer than 2 leaves", __FILE__, __LINE__);
  size_t N = (size_t)n;
  int *d_i = nullptr; float *d_f = nullptr;
  GPT_TRY(gpt_alloc(ctx, &d_i, 5 * N));
  GPT_TRY(gpt_alloc(ctx, &d_f, 12 * N));
  int *d_s2o = d_i, *d_l = d_i + N, *d_r = d_i + 2 * N, *d_pi = d_i + 3 * N, *d_pl = d_i + 4 * N;
  float *d_na = d_f, *d_la = d_f + 6 * N;
  k_export<<<(n + 255) / 256, 256, 0, ctx->stream>>>(b.nodes, b.parent_leaf, b.sorted_prim,
                                                    b.prim_list, ctx->scene.aabb, n, d_s2o, d_l,
                                                    d_r, d_pi, d_pl, d_na, d_la);
  GPT_LAUNCH_CHECK(ctx);
  auto dl = [&](void *h, const void *d, size_t bytes) -> cudaError_t {
    if (!h) return cudaSuccess;
    return cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream);
  };
  GPT_CUDA(ctx, dl(morton, b.morton, 4 * N));
  GPT_CUDA(ctx, dl(sorted_to_orig, d_s2o, 4 * N));
  GPT_CUDA(ctx, dl(left, d_l, 4 * (N - 1)));
  GPT_CUDA(ctx, dl(right, d_r, 4 * (N - 1)));
  GPT_CUDA(ctx, dl(parent_internal, d_pi, 4 * (N - 1)));
  GPT_CUDA(ctx, dl(parent_leaf, d_pl, 4 * N));
  GPT_CUDA(ctx, dl(node_aabb, d_na, 24 * (N - 1)));
  GPT_CUDA(ctx, dl(leaf_aabb, d_la, 24 * N));
  uint32_t we[6];
  GPT_CUDA(ctx, cudaMemcpyAsync(we, b.world_enc, sizeof(we), cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (world_aabb) {
    for (int a = 0; a < 3; a++) {  // (xmin,xmax,ymin,ymax,zmin,zmax) -/+ SMALL_DOUBLE
      world_aabb[2 * a] = dec_f32(we[a]) - GPT_SMALL;
      world_aabb[2 * a + 1] = dec_f32(we[3 + a]) + GPT_SMALL;
    }
  }
  gpt_free(ctx, d_i); gpt_free(ctx, d_f);
  return GPT_OK;
}
