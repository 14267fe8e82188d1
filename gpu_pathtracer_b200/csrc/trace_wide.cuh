// trace_wide.cuh -- closest hit through the quantised 4-wide tree of wide.cu.
//
// WHY.  ncu (profiles/r02_summary.md): the persistent walk of the binary LBVH
// (trace_persistent.cuh) runs at 66-68 % of the L1 data-pipe wavefront rate with the issue
// slots half idle -- every lane's 16-byte load of its own node is one wavefront -- and a probe
// that adds k dummy 16-byte loads per node visit adds k/7 of the kernel time.  The trace stage
// is bound by (16-byte loads per visit) x (visits per ray): 4 x 49 on the binary tree.  Here a
// visit is THREE 16-byte loads (a 48-byte node: origin, per-axis step exponents, 4 x 6 bytes of
// quantised child bounds, two base indices) and a bounce ray needs ~24 visits.
//
// PARITY ARGUMENT.  For a ray whose three direction components are non-zero
// and whose reciprocals are finite and moderate ("generic" below), the reference's slab
// predicate (bounding_box.h:217-276) is MONOTONE in the box: enlarging a box can only lower
// the near values and raise the far values, in floating point exactly as in the reals,
// because fp32 subtract and multiply are monotone and no NaN can arise.  Every ancestor box of
// a leaf is an exact min/max union that contains the leaf's own padded box, so the reference
// reaches leaf j  <=>  its predicate passes on j's OWN box A_j.  The reference's answer is
// therefore  argmin t over { j : pred(A_j) and Triangle::hit accepts }, ties to the later
// leaf in its visit order -- a statement about leaves only, not about the tree that is
// walked.  This kernel
//   1. walks the wide tree with a CONSERVATIVE slab test: the quantised boxes contain the
//      exact ones, and the near / far constants carry a margin that dominates every rounding
//      of this kernel's arithmetic and of the reference's (w4_node_consts), so it reaches every
//      leaf the reference reaches;
//   2. culls a subtree only by the dominant-axis slab against t_best (the argument of
//      trace_core.cuh, with the margin taken against a per-ray bound of |t|);
//   3. before a triangle may become the closest hit, evaluates the reference's exact
//      predicate on A_j, recomputed from the vertices with the operations of
//      Triangle::_compute_bounding_box (triangle.h:226-252), and drops it if that fails;
//   4. resolves equal t by the reference's visit rank on the BINARY tree (leaf_visit_rank).
// Rays that are not generic (a zero / denormal / non-finite direction component: their slab
// values can be inf or NaN, where the monotonicity argument does not hold) and rays whose
// stack would overflow take the ordered walk of the binary tree (trace_ray) instead.
#pragma once
#include "trace_persistent.cuh"

#ifndef W4_MINBLOCKS
#define W4_MINBLOCKS 7
#endif
#ifndef W4_PASS_LANES
#define W4_PASS_LANES 33   // a test pass also runs once this many lanes have a queued leaf (33 = never)
#endif
#ifndef W4_DRAIN_ALL
#define W4_DRAIN_ALL 0     // at the refill point: 1 = empty every queue, 0 = only until the lanes whose walk is over are done
#endif
#ifndef W4_QPASS_V
#define W4_QPASS_V 4
#endif
constexpr int W4_QPASS = W4_QPASS_V;     // a test pass runs when a lane has more than this many leaves queued
constexpr int W4_QSLOTS = W4_QPASS + 2;  // queue entries per lane (<= one per queued leaf before a step, one more after it)
constexpr float W4_EPS = 1.1e-5f;    // relative margin of the dominant-axis cull (trace_core.cuh: 1e-5)
constexpr float W4_MARGIN = 9.5367431640625e-07f;  // 2^-20 = 16 ulp: see w4_node_consts
constexpr float W4_STEPS = 1.25f;    // margin in grid steps: see w4_node_consts
constexpr float W4_IMAX = 1.0995116e12f;           // 2^40: |1/d| above this is "not generic"

// 2-bit lookup tables indexed by a 4-bit mask m (entry at bits 2m, 2m+1)
constexpr unsigned w4_lut(bool low) {
  unsigned v = 0;
  for (unsigned m = 1; m < 16; m++) {
    unsigned b = 0;
    if (low) { while (!((m >> b) & 1u)) b++; }
    else { b = 3; while (!((m >> b) & 1u)) b--; }
    v |= b << (2 * m);
  }
  return v;
}
constexpr unsigned W4_LOWBIT = w4_lut(true), W4_HIGHBIT = w4_lut(false);

constexpr int W4_STACK = 24;         // stack levels per lane (a deeper walk falls back to the binary tree)
// shared memory of a block, in 4-byte words, all laid out [slot][thread]: the stack, the leaf
// queue, the hit record (t, b1, b2, b3, leaf, prim) and the Woop shear (sx, sy, sz).  The last two
// are only touched by a triangle test / at the end of a ray: keeping them out of the register
// file is what lets 7 blocks of 128 threads share an SM (ncu r02d: the walk is latency bound at
// 5 blocks, issue slots 54 % busy, L1 wavefronts 40 %).
constexpr int W4_SMEM_WORDS = (W4_STACK + W4_QSLOTS + 6 + 3) * TP_BLOCK;
enum { W4_HAVE_HIT = 32 };

struct WLane {
  // axis SLOTS (0, 1, 2) = axes (kz + 1, kz + 2, kz) mod 3: slot 2 is the ray's dominant axis,
  // as in the copy of the tree this ray reads (wide.cu)
  float o[3];      // origin
  float i[3];      // 1/d (IEEE division, as the reference)
  float g;         // spatial margin 2^-20 * (|world|max + |o|max)
  float best;      // t_max, then the closest accepted t
  float bcull;     // best + eps * T: a box whose dominant-axis near value exceeds it holds no acceptable hit
  float bneg;      // -eps * T: a box whose dominant-axis far value is below it is behind the ray
  int node;        // wide node to visit next, -1 = walk over (queued leaves may remain)
  int sp;
  int ray;         // index in the source queue, -1 = idle
  int kzneg;       // kz | neg << 2 | W4_HAVE_HIT: dominant axis; bit s of neg set when 1/d of axis slot s is < 0
};

__device__ __forceinline__ float pick3(int k, float x, float y, float z) {
  return k == 0 ? x : (k == 1 ? y : z);
}

// false: the ray is not generic and must take the binary walk
__device__ __forceinline__ bool make_wlane(WLane &L, float4 ro, float4 rd, float wmax, float *shear) {
  const RaySetup r = make_ray(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z);
  const int kz = r.kz, kx = kz == 2 ? 0 : kz + 1, ky = kx == 2 ? 0 : kx + 1;
  const int ax[3] = {kx, ky, kz};
  bool generic = true;
  int neg = 0;
  float omax = 0.f;
#pragma unroll
  for (int s = 0; s < 3; s++) {
    const float o = pick3(ax[s], r.ox, r.oy, r.oz), inv = pick3(ax[s], r.ix, r.iy, r.iz);
    L.o[s] = o;
    L.i[s] = inv;
    // |inv| in (0, 2^40) and not NaN; o finite and moderate
    generic = generic && fabsf(inv) > 0.f && fabsf(inv) < W4_IMAX && fabsf(o) < 1e18f;
    omax = fmaxf(omax, fabsf(o));
    neg |= (inv < 0.f ? 1 : 0) << s;
  }
  L.g = W4_MARGIN * (wmax + omax);
  L.kzneg = kz | neg << 2;
  shear[0] = r.sx; shear[TP_BLOCK] = r.sy; shear[2 * TP_BLOCK] = r.sz;  // Woop shear (triangle.h:358-360)
  // every box lies inside the world box, so on the dominant axis |t| <= T = (|world|max + |o|) |1/d|
  // (+ this kernel's own margins, a 1e-6 relative excess that the 1.1 in W4_EPS absorbs)
  const float T = (wmax + fabsf(L.o[2])) * fabsf(L.i[2]);
  L.best = ro.w;
  L.bneg = -W4_EPS * T;
  L.bcull = ro.w - L.bneg;
  L.node = 0; L.sp = 0;
  return generic;
}

// a triangle's vertices in the lane's axis slots
struct WTri {
  float p[3][3];  // [slot][vertex]
};
__device__ __forceinline__ WTri w4_permute_tri(int kz, float4 t0, float4 t1, float4 t2) {
  WTri v;
  permute3(kz, t0.x, t0.y, t0.z, v.p[0][0], v.p[1][0], v.p[2][0]);
  permute3(kz, t0.w, t1.x, t1.y, v.p[0][1], v.p[1][1], v.p[2][1]);
  permute3(kz, t1.z, t1.w, t2.x, v.p[0][2], v.p[1][2], v.p[2][2]);
  return v;
}

// Triangle::hit (triangle.h:343-408) on a lane whose origin is stored permuted: the vertex is
// permuted first and the origin subtracted after -- the same values as tri_test's
// subtract-then-permute.
__device__ __forceinline__ bool w4_tri_test(const WLane &L, const WTri &v, float sx, float sy,
                                            float sz, float &t, float &b1, float &b2, float &b3) {
  float ax = v.p[0][0] - L.o[0], ay = v.p[1][0] - L.o[1], az = v.p[2][0] - L.o[2];
  float bx = v.p[0][1] - L.o[0], by = v.p[1][1] - L.o[1], bz = v.p[2][1] - L.o[2];
  float cx = v.p[0][2] - L.o[0], cy = v.p[1][2] - L.o[1], cz = v.p[2][2] - L.o[2];
  ax = ax + sx * az; ay = ay + sy * az;
  bx = bx + sx * bz; by = by + sy * bz;
  cx = cx + sx * cz; cy = cy + sy * cz;
  float e1 = bx * cy - by * cx;
  float e2 = cx * ay - cy * ax;
  float e3 = ax * by - ay * bx;
  if (e1 == 0.f || e2 == 0.f || e3 == 0.f) {  // fp64 fallback (triangle.h:369-379)
    e1 = (float)((double)bx * (double)cy - (double)by * (double)cx);
    e2 = (float)((double)cx * (double)ay - (double)cy * (double)ax);
    e3 = (float)((double)ax * (double)by - (double)ay * (double)bx);
  }
  if ((e1 < 0.f || e2 < 0.f || e3 < 0.f) && (e1 > 0.f || e2 > 0.f || e3 > 0.f)) return false;
  float det = e1 + e2 + e3;
  if (det == 0.f) return false;
  az = az * sz; bz = bz * sz; cz = cz * sz;
  float t_scaled = e1 * az + e2 * bz + e3 * cz;
  float inv_det = 1.0f / det;
  b1 = e1 * inv_det; b2 = e2 * inv_det; b3 = e3 * inv_det;
  t = t_scaled * inv_det;
  return true;
}

// The reference's slab predicate on the leaf's own box, which is recomputed from the vertices
// exactly as preprocess.cu / Triangle::_compute_bounding_box do: min/max of the three
// coordinates (exact), -+ tolerance (one rounding).  Near / far by the sign of 1/d is the
// reference's `t1 < t2 ? ...` selection for a generic ray (trace_persistent.cuh header).
// _are_intersecting twice (bounding_box.h:61-69,217-276),
//   (xn <= yx && yn <= xx) && (max(xn, yn) <= zx && zn <= min(xx, yx)),
// is, for values that are not NaN, "every near <= every other axis' far": symmetric under the
// relabelling of the axes this lane uses.
__device__ __forceinline__ bool w4_leaf_box_pred(const WLane &L, const WTri &v, float tol) {
  float tn[3], tf[3];
#pragma unroll
  for (int s = 0; s < 3; s++) {
    const float lo = fminf(fminf(v.p[s][0], v.p[s][1]), v.p[s][2]) - tol;
    const float hi = fmaxf(fmaxf(v.p[s][0], v.p[s][1]), v.p[s][2]) + tol;
    const float t1v = (lo - L.o[s]) * L.i[s], t2v = (hi - L.o[s]) * L.i[s];
    tn[s] = t1v < t2v ? t1v : t2v;
    tf[s] = t1v < t2v ? t2v : t1v;
  }
  return tn[0] <= tf[1] && tn[1] <= tf[0] && tn[0] <= tf[2] && tn[1] <= tf[2] && tn[2] <= tf[0] &&
         tn[2] <= tf[1];
}

struct W4Queue {   // per-lane LIFO of leaf groups waiting for their triangle tests, [slot][thread]
  unsigned *q;     // entry = first leaf slot's triangle index (24 bits) | pending leaf slots << 24
  int cnt = 0;     // entries
  int pend = 0;    // leaves in them
  __device__ __forceinline__ W4Queue(int *smem) : q(reinterpret_cast<unsigned *>(smem) + threadIdx.x) {}
};

// the triangle test of the next queued leaf of this lane; rec = this lane's hit record and shear
template <bool COUNT>
__device__ __forceinline__ void w4_leaf_test(const float4 *__restrict__ bnodes,
                                             const float4 *__restrict__ wtris,
                                             const int *__restrict__ parent_leaf, WLane &L,
                                             W4Queue &Q, float *rec, unsigned &vl) {
  const unsigned e = Q.q[(Q.cnt - 1) * TP_BLOCK];
  const unsigned lh = e >> 24;
  const unsigned c = (W4_LOWBIT >> (2 * lh)) & 3u;
  const unsigned tri = (e & 0xffffffu) + c;
  const unsigned lh2 = lh & (lh - 1u);
  if (lh2) Q.q[(Q.cnt - 1) * TP_BLOCK] = (e & 0xffffffu) | lh2 << 24;
  else Q.cnt--;
  Q.pend--;
  const float4 *tp = wtris + 3 * (size_t)tri;
  const float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
  const float sx = rec[6 * TP_BLOCK], sy = rec[7 * TP_BLOCK], sz = rec[8 * TP_BLOCK];
  if (COUNT) vl++;
  const WTri v = w4_permute_tri(L.kzneg & 3, t0, t1, t2);
  float t, b1, b2, b3;
  if (!w4_tri_test(L, v, sx, sy, sz, t, b1, b2, b3)) return;
  if (t > L.best || t < t2.y) return;         // reference: reject if (t > t_max || t < tolerance)
  if (!w4_leaf_box_pred(L, v, t2.y)) return;  // the reference never reaches this leaf
  const int lf = __float_as_int(t2.w);        // LBVH leaf (sorted position)
  bool take = true;
  if ((L.kzneg & W4_HAVE_HIT) && t == L.best)  // equal t: the later one in reference order wins
    take = leaf_visit_rank(bnodes, parent_leaf, lf) >
           leaf_visit_rank(bnodes, parent_leaf, __float_as_int(rec[4 * TP_BLOCK]));
  if (take) {
    rec[0] = t; rec[TP_BLOCK] = b1; rec[2 * TP_BLOCK] = b2; rec[3 * TP_BLOCK] = b3;
    rec[4 * TP_BLOCK] = t2.w; rec[5 * TP_BLOCK] = t2.z;  // leaf, prim (as bits)
    L.kzneg |= W4_HAVE_HIT;
    L.best = t;
    L.bcull = t - L.bneg;
  }
}

// Near / far constants of one axis of one node.  A child bound is origin + q * step (q a byte).
// With Q = 2^23 + q (built by placing the byte in the mantissa of 2^23, exact),
//   t = fl(Q * S + C),   S = step / d (exact scaling of 1/d by a power of two),
//                        C = fl(fl(fl(fl(origin - o) * (1/d)) - 2^23 * S) -+ m)
// Rounding: C carries <= 1 ulp of 2^23 |S| = 1 grid step in t (the two roundings next to the
// 2^23 S term), the rest <= 2^-23 (3 |origin - o| + |t|) |1/d|; the reference's own value
// fl(fl(b - o) * (1/d)) is within 2^-23 |b - o| |1/d| of the real one; wide.cu's floor / ceil
// may be off by 2e-5 steps.  m = (1.25 step + 2^-20 (|world|max + |o|max)) |1/d| covers the
// sum, so near <= the reference's near and far >= the reference's far for every box below.
__device__ __forceinline__ void w4_node_consts(float org, unsigned ebits, float inv, float o,
                                               float g, float &S, float &Cn, float &Cf) {
  const float step = __uint_as_float(ebits);
  S = step * inv;
  const float C1 = fmaf(-8388608.f, S, (org - o) * inv);
  const float m = fmaf(W4_STEPS, step, g) * fabsf(inv);
  Cn = C1 - m;
  Cf = C1 + m;
}

// byte c of w as the float 2^23 + byte (prmt with an IMMEDIATE selector: written as a*b+c in C
// the compiler keeps the 0x4b000000 as the immediate and spends a move per selector)
template <int C>
__device__ __forceinline__ float w4_q(unsigned w, unsigned magic) {
  unsigned r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(magic), "n"(0x7650 + C));
  return __uint_as_float(r);
}

template <int C>
__device__ __forceinline__ unsigned w4_child(const unsigned (&nw)[3], const unsigned (&fw)[3],
                                             const float (&S)[3], const float (&Cn)[3],
                                             const float (&Cf)[3], unsigned magic, float bcull,
                                             float bneg) {
  const float tn0 = fmaf(w4_q<C>(nw[0], magic), S[0], Cn[0]), tf0 = fmaf(w4_q<C>(fw[0], magic), S[0], Cf[0]);
  const float tn1 = fmaf(w4_q<C>(nw[1], magic), S[1], Cn[1]), tf1 = fmaf(w4_q<C>(fw[1], magic), S[1], Cf[1]);
  const float tn2 = fmaf(w4_q<C>(nw[2], magic), S[2], Cn[2]), tf2 = fmaf(w4_q<C>(fw[2], magic), S[2], Cf[2]);
  const float tnear = fmaxf(fmaxf(tn0, tn1), tn2), tfar = fminf(fminf(tf0, tf1), tf2);
  // slot 2 is the dominant axis: the cull of trace_core.cuh, with the margin eps * T of make_wlane
  return (tnear <= tfar && tn2 <= bcull && tf2 >= bneg) ? (1u << C) : 0u;
}

// One visit of wide node L.node: four conservative slab tests, the leaf children that pass
// queued as one entry, the internal children that pass visited in the node's sort order (one
// stack entry holds the ones not yet visited).  The caller guarantees a free stack level.
//   stack entry = child_base (23 bits) | not-yet-visited internal slots << 23 | ascending << 27
template <bool COUNT>
__device__ __forceinline__ void w4_step(const uint4 *__restrict__ wq, WLane &L, unsigned *stack,
                                        W4Queue &Q, unsigned &vi, unsigned magic) {
  const uint4 *n = wq + 3 * (size_t)L.node;
  const uint4 w0 = __ldg(n), w1 = __ldg(n + 1), w2 = __ldg(n + 2);
  const unsigned top = L.sp > 0 ? stack[(L.sp - 1) * TP_BLOCK] : 0u;  // wanted if nothing below is hit
  if (COUNT) vi++;
  const unsigned meta = w0.w;
  const int neg = L.kzneg >> 2;
  float S[3], Cn[3], Cf[3];
  w4_node_consts(__uint_as_float(w0.x), (meta << 23) & 0x7f800000u, L.i[0], L.o[0], L.g, S[0], Cn[0], Cf[0]);
  w4_node_consts(__uint_as_float(w0.y), (meta << 15) & 0x7f800000u, L.i[1], L.o[1], L.g, S[1], Cn[1], Cf[1]);
  w4_node_consts(__uint_as_float(w0.z), (meta << 7) & 0x7f800000u, L.i[2], L.o[2], L.g, S[2], Cn[2], Cf[2]);
  // near / far bound bytes per axis slot by the sign of 1/d
  const unsigned nw[3] = {(neg & 1) ? w1.w : w1.x, (neg & 2) ? w2.x : w1.y, (neg & 4) ? w2.y : w1.z};
  const unsigned fw[3] = {(neg & 1) ? w1.x : w1.w, (neg & 2) ? w1.y : w2.x, (neg & 4) ? w1.z : w2.y};
  const unsigned hm = w4_child<0>(nw, fw, S, Cn, Cf, magic, L.bcull, L.bneg) |
                      w4_child<1>(nw, fw, S, Cn, Cf, magic, L.bcull, L.bneg) |
                      w4_child<2>(nw, fw, S, Cn, Cf, magic, L.bcull, L.bneg) |
                      w4_child<3>(nw, fw, S, Cn, Cf, magic, L.bcull, L.bneg);
  // slots: ni internal children, then nl leaves, then unused ones (their bits are dropped here)
  const unsigned ni = (meta >> 24) & 7u, nl = (meta >> 27) & 7u;
  const unsigned ih = hm & ~(~0u << ni), lh = (hm >> ni) & ~(~0u << nl);
  if (lh) {
    Q.q[Q.cnt * TP_BLOCK] = w2.w | lh << 24;
    Q.cnt++;
    Q.pend += __popc(lh);
  }
  // children are sorted by box centre along the node's axis: ascending for a ray going up it
  const unsigned asc = (((unsigned)neg >> (meta >> 30)) & 1u) ^ 1u;
  const unsigned cand = ih ? (w2.z | ih << 23 | asc << 27) : top;
  if (cand == 0u) {
    L.node = -1;
    return;
  }
  const unsigned rem = (cand >> 23) & 15u;
  const unsigned c = ((((cand >> 27) & 1u) ? W4_LOWBIT : W4_HIGHBIT) >> (2 * rem)) & 3u;
  L.node = (int)((cand & 0x7fffffu) + c);
  const unsigned rem2 = rem & ~(1u << c);
  const unsigned e2 = (cand & ~(15u << 23)) | rem2 << 23;
  if (ih) {
    if (rem2) { stack[L.sp * TP_BLOCK] = e2; L.sp++; }
  } else {
    if (rem2) stack[(L.sp - 1) * TP_BLOCK] = e2;
    else L.sp--;
  }
}

// a ray the wide walk cannot take: the ordered walk of the binary tree, on a local stack.
// (Takes the ray and returns the record by value: a reference to the caller's Source here would
// force that whole struct, kernel parameters included, into local memory.)
template <bool COUNT>
static __device__ __noinline__ HitResult w4_fallback(const float4 *__restrict__ bnodes,
                                                     const float4 *__restrict__ tris,
                                                     const int *__restrict__ parent_leaf, float4 ro,
                                                     float4 rd, unsigned long long *counters) {
  const RaySetup r = make_ray(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z);
  int stack[TRACE_STACK];
  int overflow = 0;
  const HitResult h = trace_ray<COUNT>(bnodes, tris, parent_leaf, r, ro.w, stack, 1, counters,
                                       counters + 1, &overflow);
  if (overflow) atomicExch(counters + 3, 1ull);
  return h;
}

struct WideTree {        // what the kernels need of a BvhDev
  const uint4 *wq;       // wide nodes (wide.cu): three copies of `stride` nodes, one per dominant axis
  int stride;
  const float4 *wtris;   // triangles in wide-leaf order
  const float *wmeta;    // [0] max |coordinate| of the world box, [1] the float 2^23
  const float4 *nodes;   // the binary LBVH: tie-break ranks and the fallback walk
  const float4 *tris;
  const int *parent_leaf;
};

// Same Source concept as trace_persistent plus
//   void store_warp(bool fin, int i, const TraceLane &l);  called by all 32 lanes, stores for those with fin
// and the same persistent-warp schedule (ray refill at
// TP_REFILL idle lanes, triangle tests batched per warp); the step is the wide one.
template <bool COUNT, class Source>
__device__ __forceinline__ void trace_wide(const WideTree tr, Source src, int *next_ray,
                                           int *smem /* W4_SMEM_WORDS */,
                                           unsigned long long *counters) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int n = src.count();
  unsigned *stack = reinterpret_cast<unsigned *>(smem) + threadIdx.x;  // [level][thread], stride TP_BLOCK
  float *rec = reinterpret_cast<float *>(smem) + (W4_STACK + W4_QSLOTS) * TP_BLOCK + threadIdx.x;
  const float wmax = __ldg(tr.wmeta);
  // 0x4b000000 (the float 2^23), read from memory so that it stays in a register: as a literal
  // it becomes PRMT's one immediate operand and every byte selector costs a move instead
  const unsigned magic = __float_as_uint(__ldg(tr.wmeta + 1));
  WLane L;
  L.ray = -1;
  L.node = -1;
  W4Queue Q(smem + W4_STACK * TP_BLOCK);
  unsigned vi = 0, vl = 0;
  bool exhausted = false;
  // called by all 32 lanes: the lanes whose walk is over and whose queue is empty hand their
  // record to the source (store_warp lets a source aggregate what the lanes of a warp append)
  auto finish = [&]() {
    const bool fin = L.ray >= 0 && L.node < 0 && Q.pend == 0;
    if (__ballot_sync(FULL, fin) == 0) return;
    TraceLane R;
    R.t = INFINITY; R.leaf = -1; R.prim = -1; R.b1 = R.b2 = R.b3 = 0.f;
    if (fin && (L.kzneg & W4_HAVE_HIT)) {
      R.t = rec[0]; R.b1 = rec[TP_BLOCK]; R.b2 = rec[2 * TP_BLOCK]; R.b3 = rec[3 * TP_BLOCK];
      R.leaf = __float_as_int(rec[4 * TP_BLOCK]); R.prim = __float_as_int(rec[5 * TP_BLOCK]);
    }
    src.store_warp(fin, L.ray, R);
    if (fin) L.ray = -1;
  };
  auto fallback = [&](int ray) {
    float4 ro, rd;
    src.load(ray, ro, rd);
    const HitResult h = w4_fallback<COUNT>(tr.nodes, tr.tris, tr.parent_leaf, ro, rd, counters);
    TraceLane R;
    R.t = h.t; R.leaf = h.leaf; R.prim = h.prim; R.b1 = h.b1; R.b2 = h.b2; R.b3 = h.b3;
    src.store(ray, R);
  };
  while (true) {
    // ---- refill idle lanes -----------------------------------------------------------
    unsigned idle = __ballot_sync(FULL, L.ray < 0);
    if (idle && !exhausted) {
      int cnt = __popc(idle), leader = __ffs(idle) - 1, base = 0;
      if (lane == leader) base = atomicAdd(next_ray, cnt);
      base = __shfl_sync(FULL, base, leader);
      if (base + cnt >= n) exhausted = true;
      if (L.ray < 0) {
        int idx = base + __popc(idle & ((1u << lane) - 1u));
        if (idx < n) {
          float4 ro, rd;
          src.load(idx, ro, rd);
          if (make_wlane(L, ro, rd, wmax, rec + 6 * TP_BLOCK)) L.ray = idx;
          else fallback(idx);
        }
      }
    }
    if (__ballot_sync(FULL, L.ray >= 0) == 0) {
      if (exhausted) break;
      continue;  // every ray of this refill took the fallback: fetch more
    }
    // ---- traverse until enough lanes have gone idle ------------------------------------
    while (true) {
      if (L.ray >= 0 && L.node >= 0) {
        if (L.sp < W4_STACK) {
          w4_step<COUNT>(tr.wq + 3 * (size_t)(L.kzneg & 3) * tr.stride, L, stack, Q, vi, magic);
        } else {  // no free stack level: redo this ray on the binary tree (local stack)
          fallback(L.ray);
          L.ray = -1; L.node = -1; Q.cnt = 0; Q.pend = 0;
        }
      }
      bool pass = __ballot_sync(FULL, Q.pend > W4_QPASS) != 0;
      if (W4_PASS_LANES <= 32 && !pass) pass = __popc(__ballot_sync(FULL, Q.pend > 0)) >= W4_PASS_LANES;
      while (pass) {
        if (Q.pend > 0) w4_leaf_test<COUNT>(tr.nodes, tr.wtris, tr.parent_leaf, L, Q, rec, vl);
        pass = __ballot_sync(FULL, Q.pend > W4_QPASS) != 0;
      }
      finish();
      const int nwalk = __popc(__ballot_sync(FULL, L.ray >= 0 && L.node >= 0));
      if (nwalk == 0 || (!exhausted && 32 - nwalk >= TP_REFILL)) break;
    }
    // back at the refill point: the lanes whose walk is over need their queues empty before
    // their record can be stored (every lane with a queued leaf takes part in those passes)
    while (__ballot_sync(FULL, W4_DRAIN_ALL ? Q.pend > 0 : (Q.pend > 0 && L.ray >= 0 && L.node < 0))) {
      if (Q.pend > 0) w4_leaf_test<COUNT>(tr.nodes, tr.wtris, tr.parent_leaf, L, Q, rec, vl);
    }
    finish();
  }
  if (COUNT) {
    atomicAdd(counters, (unsigned long long)vi);
    atomicAdd(counters + 1, (unsigned long long)vl);
    // the wide walk's own share of the two totals (48-byte nodes: the roofline numerator
    // prices a wide visit differently from a 64-byte binary node)
    atomicAdd(counters + 5, (unsigned long long)vi);
    atomicAdd(counters + 7, (unsigned long long)vl);
  }
}

static inline WideTree gpt_wide_tree(const BvhDev &b) {
  return WideTree{b.wq, b.wide_capacity, b.wtris, b.wmeta, b.nodes, b.tris, b.parent_leaf};
}
