// trace_persistent.cuh -- persistent-warp closest-hit traversal with ray refill.
//
// Same per-ray semantics as trace_ray (trace_core.cuh, see the parity argument there); what
// changes is the execution model.  With one ray per thread a warp runs until its SLOWEST ray
// ends (ncu r01: 13.6 of 32 lanes active).  Here a fixed grid of warps pulls rays from a global
// counter; whenever >= REFILL lanes of a warp are idle they are re-armed with the next rays of
// the queue (warp-aggregated atomicAdd), so lanes stay busy and neighbouring lanes keep
// neighbouring rays.  Slab ordering is decided once per ray from the sign of 1/d instead of per
// box (exactly equivalent to the reference's `t1 < t2 ? ...` selection: for 1/d >= 0 the
// products are already ordered, for 1/d < 0 they are reversed, and any NaN slab value makes
// the predicate false in either order -- bounding_box.h:61-69,217-276).
#pragma once
#include "trace_core.cuh"

constexpr int TP_BLOCK = 128;
#ifndef TP_STACK_SMEM_V
#define TP_STACK_SMEM_V 32
#endif
#ifndef TP_MINBLOCKS
#define TP_MINBLOCKS 1
#endif
#ifndef TP_PREFETCH
#define TP_PREFETCH 0
#endif
constexpr int TP_STACK_SMEM = TP_STACK_SMEM_V;   // levels kept in shared memory; deeper levels spill to local
#ifndef TP_REFILL_V
#define TP_REFILL_V 20
#endif
constexpr int TP_REFILL = TP_REFILL_V;  // idle lanes that trigger a refill

struct TraceLane {
  float ox, oy, oz, ix, iy, iz;
  float sx, sy, sz;
  float best;
  float t, b1, b2, b3;
  int leaf, prim;
  int node, sp;
  int ray;         // index in the source queue, -1 = idle
  int kz;
  unsigned neg;    // bit a set: 1/d[a] < 0 -> the "near" bound of that axis is the box max
};

// Source concept:
//   int count() const;                       number of rays in the queue
//   void load(int i, float4 &o, float4 &d);  ray i: {o, tmax}{unit d, -}
//   void store(int i, const TraceLane &l);   hit record of ray i
template <bool COUNT, class Source>
__device__ __forceinline__ void trace_persistent(const float4 *__restrict__ nodes,
                                                 const float4 *__restrict__ tris,
                                                 const int *__restrict__ parent_leaf,
                                                 Source src, int *next_ray, int *stack_smem,
                                                 unsigned long long *counters) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int n = src.count();
  int *stack = stack_smem + threadIdx.x;  // [level][thread], stride TP_BLOCK
  int spill[TRACE_STACK - TP_STACK_SMEM];
  TraceLane L;
  L.ray = -1;
  unsigned vi = 0, vl = 0;
  bool exhausted = false;
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
          RaySetup r = make_ray(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z);
          L.ox = r.ox; L.oy = r.oy; L.oz = r.oz; L.ix = r.ix; L.iy = r.iy; L.iz = r.iz;
          L.sx = r.sx; L.sy = r.sy; L.sz = r.sz; L.kz = r.kz;
          L.neg = (r.ix < 0.f ? 1u : 0u) | (r.iy < 0.f ? 2u : 0u) | (r.iz < 0.f ? 4u : 0u);
          L.best = ro.w;
          L.t = INFINITY; L.leaf = -1; L.prim = -1; L.b1 = L.b2 = L.b3 = 0.f;
          L.node = 0; L.sp = 0;
          L.ray = idx;
        }
      }
    }
    if (__ballot_sync(FULL, L.ray >= 0) == 0) break;
    // ---- traverse until enough lanes have gone idle ------------------------------------
    // (postponing leaf tests into separate warp-wide "intersect rounds" was measured and was
    //  10-20 % SLOWER on C2 -- the extra rounds and votes cost more than the divergence they
    //  remove with one triangle per leaf -- so leaves are tested where they are found.)
    while (true) {
      if (L.ray >= 0) {
        const float4 *q = nodes + 4 * (size_t)L.node;
        float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
        if (COUNT) vi++;
        int cref[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
        // near / far bound per axis from the sign of 1/d
        const bool nx = L.neg & 1u, ny = L.neg & 2u, nz = L.neg & 4u;
        float xn[2] = {nx ? q0.y : q0.x, nx ? q1.w : q1.z}, xf[2] = {nx ? q0.x : q0.y, nx ? q1.z : q1.w};
        float yn[2] = {ny ? q0.w : q0.z, ny ? q2.y : q2.x}, yf[2] = {ny ? q0.z : q0.w, ny ? q2.x : q2.y};
        float zn[2] = {nz ? q1.y : q1.x, nz ? q2.w : q2.z}, zf[2] = {nz ? q1.x : q1.y, nz ? q2.z : q2.w};
        bool go[2];
        float tnear[2], tcull[2];
#pragma unroll
        for (int c = 0; c < 2; c++) {
          float txn = (xn[c] - L.ox) * L.ix, txx = (xf[c] - L.ox) * L.ix;
          float tyn = (yn[c] - L.oy) * L.iy, tyx = (yf[c] - L.oy) * L.iy;
          float tzn = (zn[c] - L.oz) * L.iz, tzx = (zf[c] - L.oz) * L.iz;
          bool ok = (txn <= tyx) && (tyn <= txx);
          float txyn = fmaxf(txn, tyn), txyx = fminf(txx, tyx);
          ok = ok && (txyn <= tzx) && (tzn <= txyx);
          float tn = L.kz == 0 ? txn : (L.kz == 1 ? tyn : tzn);
          float tf = L.kz == 0 ? txx : (L.kz == 1 ? tyx : tzx);
          float margin = 1e-5f * fmaxf(fabsf(tn), fabsf(tf));
          tnear[c] = tn;
          tcull[c] = tn - margin;
          go[c] = ok && !(tf + margin < 0.f);
        }
        // Leaves: the two children's tests share ONE copy of the intersection code (a leaf on
        // either side of any lane runs in the same round), instead of two unrolled copies that
        // each execute with ~2 of 32 lanes (ncu r01b: 37 % of issued instructions).
        int lf0 = (go[0] && cref[0] < 0) ? ~cref[0] : -1, lf1 = (go[1] && cref[1] < 0) ? ~cref[1] : -1;
        float lc0 = tcull[0], lc1 = tcull[1];
        if (lf0 < 0) { lf0 = lf1; lc0 = lc1; lf1 = -1; }
#pragma unroll 1
        while (lf0 >= 0) {
          if (!(lc0 > L.best)) {
            const float4 *tp = tris + 3 * (size_t)lf0;
            float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
            if (COUNT) vl++;
            RaySetup r;
            r.ox = L.ox; r.oy = L.oy; r.oz = L.oz; r.sx = L.sx; r.sy = L.sy; r.sz = L.sz; r.kz = L.kz;
            float t, b1, b2, b3;
            if (tri_test(r, t0, t1, t2, t, b1, b2, b3)) {
              if (!(t > L.best) && !(t < t2.y)) {
                bool take = true;
                if (t == L.t && L.leaf >= 0)
                  take = leaf_visit_rank(nodes, parent_leaf, lf0) >
                         leaf_visit_rank(nodes, parent_leaf, L.leaf);
                if (take) {
                  L.t = t; L.leaf = lf0; L.b1 = b1; L.b2 = b2; L.b3 = b3;
                  L.prim = __float_as_int(t2.z);
                  L.best = t;
                }
              }
            }
          }
          lf0 = lf1; lc0 = lc1; lf1 = -1;
        }
        // Internal children, written to compile into predicated code rather than a three-way
        // divergent branch (push / step / pop): near child first, far child pushed.
        const bool gl = go[0] && cref[0] >= 0 && !(tcull[0] > L.best);
        const bool gr = go[1] && cref[1] >= 0 && !(tcull[1] > L.best);
        const bool both = gl && gr, any = gl || gr;
        const bool lfirst = tnear[0] <= tnear[1];
        const int nearc = (gl && (!gr || lfirst)) ? cref[0] : cref[1];
        const int farc = lfirst ? cref[1] : cref[0];
        if (both && L.sp < TP_STACK_SMEM) stack[L.sp * TP_BLOCK] = farc;
#if TP_PREFETCH
        if (both && farc >= 0) {
          const float4 *fq = nodes + 4 * (size_t)farc;
          asm volatile("prefetch.global.L2 [%0];" ::"l"(fq));
        }
#endif
        if (both && L.sp >= TP_STACK_SMEM) {  // rare: deeper than the shared-memory levels
          if (L.sp < TRACE_STACK) spill[L.sp - TP_STACK_SMEM] = farc;
          else atomicExch(counters + 3, 1ull);
        }
        if (both && L.sp < TRACE_STACK) L.sp++;
        const bool finished = !any && L.sp == 0;
        if (any) {
          L.node = nearc;
        } else if (L.sp > 0) {
          L.sp--;
          L.node = L.sp < TP_STACK_SMEM ? stack[L.sp * TP_BLOCK] : spill[L.sp - TP_STACK_SMEM];
        }
        if (finished) {
          src.store(L.ray, L);
          L.ray = -1;
        }
      }
      unsigned nidle = __popc(__ballot_sync(FULL, L.ray < 0));
      if (nidle == 32 || (!exhausted && nidle >= TP_REFILL)) break;
    }
  }
  if (COUNT) {
    atomicAdd(counters, (unsigned long long)vi);
    atomicAdd(counters + 1, (unsigned long long)vl);
  }
}
