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

#ifndef TP_BLOCK_V
#define TP_BLOCK_V 128
#endif
constexpr int TP_BLOCK = TP_BLOCK_V;  // threads per persistent trace block (a block keeps its SM resources until its slowest warp ends)
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
#define TP_REFILL_V 12
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


// ---- cooperative tail -----------------------------------------------------------------
// A launch ends when its slowest ray ends, and a ray is a chain of dependent node steps of
// ~0.5-0.7 us each.  On a 1/8-frame shard every ray of the queue is in flight at once, so the
// launch time IS that one chain (ncu / shard_time: ~250 us of a 310 us iteration), and the
// 320 dependent iterations of a 64-spp frame are what bounds strong scaling.  So, once the
// queue is exhausted and a warp is down to <= COOP_SLOTS live rays, the warp stops walking
// them one lane each: the pending nodes of ALL its live rays (current node + stack entries)
// go into one warp-wide work list in shared memory, and every round each lane pops one entry,
// tests that node's two children against ITS ray, tests leaves, and pushes the children that
// pass.  Closest hit with the visit-rank tie-break does not depend on the visit order (the
// parity argument in trace_core.cuh never uses it), so the result is bit-identical; culling
// sees t_best of the round's start only, so a few more nodes are visited than in the ordered
// walk, which the tail has idle lanes to pay for.
#ifndef COOP_SLOTS_V
#define COOP_SLOTS_V 16
#endif
constexpr int COOP_SLOTS = COOP_SLOTS_V;        // live rays a warp may take into the tail (<= 32)
constexpr int COOP_LIST = 32 * TP_STACK_SMEM;   // work-list entries per warp (its share of the stack array)
constexpr int COOP_RAY_WORDS = 18;
constexpr int COOP_WARP_WORDS = COOP_RAY_WORDS * COOP_SLOTS;  // ray table; the list reuses the stack array
constexpr int COOP_NODE_BITS = 27;              // entry = slot << 27 | node
enum { CR_OX, CR_OY, CR_OZ, CR_IX, CR_IY, CR_IZ, CR_SX, CR_SY, CR_SZ, CR_KZNEG, CR_BEST, CR_T,
       CR_LEAF, CR_B1, CR_B2, CR_B3, CR_PRIM, CR_RAY };

template <bool COUNT, class Source>
__device__ __forceinline__ void coop_finish(const float4 *__restrict__ nodes,
                                         const float4 *__restrict__ tris,
                                         const int *__restrict__ parent_leaf, Source &src,
                                         const TraceLane &L, unsigned live, const int *stack,
                                         int *sm, unsigned &vi, unsigned &vl) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  // The work list lives in the warp's own part of the traversal-stack array: 32 levels x 32
  // lanes, entry i at level i / 32, column i % 32.  (A separate list costs shared memory that
  // the unified L1 then lacks: 37 KB instead of 16 KB per block made the FULL-frame trace 12 %
  // slower, with the tail code never even reached.)
  int *const wbase = const_cast<int *>(stack) - lane;
#define WL(i) wbase[((i) >> 5) * TP_BLOCK + ((i) & 31)]
  float *rt = reinterpret_cast<float *>(sm);  // ray table [word][slot]
  int *rti = sm;
  const bool mine = (live >> lane) & 1u;
  const int nslots = __popc(live);
  const int slot0 = __popc(live & ((1u << lane) - 1u));
  // ---- the live lanes' stacks, compacted in place level by level: the entries of level k
  //      land at list positions < 32 (k + 1), i.e. in levels that were already read ----------
  int cnt = 0;
  {
    int maxsp = mine ? L.sp : 0;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) maxsp = max(maxsp, __shfl_xor_sync(FULL, maxsp, d));
    const int tag = slot0 << COOP_NODE_BITS;
    for (int k = 0; k <= maxsp; k++) {
      // level sp of a lane = its current node (the most urgent entry, so the last one)
      const bool has = mine && k <= L.sp;
      int v = 0;
      if (has) v = tag | (k < L.sp ? stack[k * TP_BLOCK] : L.node);
      const unsigned hm = __ballot_sync(FULL, has);
      __syncwarp();
      if (has) WL(cnt + __popc(hm & ((1u << lane) - 1u))) = v;
      cnt += __popc(hm);
      __syncwarp();
    }
  }
  if (mine) {
    const int slot = slot0;
    rt[CR_OX * COOP_SLOTS + slot] = L.ox; rt[CR_OY * COOP_SLOTS + slot] = L.oy;
    rt[CR_OZ * COOP_SLOTS + slot] = L.oz; rt[CR_IX * COOP_SLOTS + slot] = L.ix;
    rt[CR_IY * COOP_SLOTS + slot] = L.iy; rt[CR_IZ * COOP_SLOTS + slot] = L.iz;
    rt[CR_SX * COOP_SLOTS + slot] = L.sx; rt[CR_SY * COOP_SLOTS + slot] = L.sy;
    rt[CR_SZ * COOP_SLOTS + slot] = L.sz;
    rti[CR_KZNEG * COOP_SLOTS + slot] = L.kz | (int)(L.neg << 2);
    rt[CR_BEST * COOP_SLOTS + slot] = L.best; rt[CR_T * COOP_SLOTS + slot] = L.t;
    rti[CR_LEAF * COOP_SLOTS + slot] = L.leaf; rt[CR_B1 * COOP_SLOTS + slot] = L.b1;
    rt[CR_B2 * COOP_SLOTS + slot] = L.b2; rt[CR_B3 * COOP_SLOTS + slot] = L.b3;
    rti[CR_PRIM * COOP_SLOTS + slot] = L.prim; rti[CR_RAY * COOP_SLOTS + slot] = L.ray;
  }
  __syncwarp();
  // ---- rounds: every lane pops one (ray, node) entry ------------------------------------
  while (cnt > 0) {
    int m = cnt < 32 ? cnt : 32;
    if (cnt > COOP_LIST - 128) m = 1;  // nearly full: depth-first with one lane, which cannot grow it by more than the tree depth
    const bool has_e = lane < m;
    const int e = has_e ? WL(cnt - 1 - lane) : 0;
    cnt -= m;
    __syncwarp();
    int npush = 0, push_far = 0, push_near = 0, slot = 0;
    bool cand = false;
    float ct = 0.f, cb1 = 0.f, cb2 = 0.f, cb3 = 0.f, ctol = 0.f;
    int cleaf = -1, cprim = -1;
    if (has_e) {
      slot = (int)((unsigned)e >> COOP_NODE_BITS);
      const int node = e & ((1 << COOP_NODE_BITS) - 1);
      RaySetup r;
      r.ox = rt[CR_OX * COOP_SLOTS + slot]; r.oy = rt[CR_OY * COOP_SLOTS + slot];
      r.oz = rt[CR_OZ * COOP_SLOTS + slot]; r.ix = rt[CR_IX * COOP_SLOTS + slot];
      r.iy = rt[CR_IY * COOP_SLOTS + slot]; r.iz = rt[CR_IZ * COOP_SLOTS + slot];
      r.sx = rt[CR_SX * COOP_SLOTS + slot]; r.sy = rt[CR_SY * COOP_SLOTS + slot];
      r.sz = rt[CR_SZ * COOP_SLOTS + slot];
      const int kzneg = rti[CR_KZNEG * COOP_SLOTS + slot];
      r.kz = kzneg & 3;
      const unsigned neg = (unsigned)kzneg >> 2;
      float best = rt[CR_BEST * COOP_SLOTS + slot];
      const float4 *q = nodes + 4 * (size_t)node;
      float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
      if (COUNT) vi++;
      int cref[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
      const bool nx = neg & 1u, ny = neg & 2u, nz = neg & 4u;
      float xn[2] = {nx ? q0.y : q0.x, nx ? q1.w : q1.z}, xf[2] = {nx ? q0.x : q0.y, nx ? q1.z : q1.w};
      float yn[2] = {ny ? q0.w : q0.z, ny ? q2.y : q2.x}, yf[2] = {ny ? q0.z : q0.w, ny ? q2.x : q2.y};
      float zn[2] = {nz ? q1.y : q1.x, nz ? q2.w : q2.z}, zf[2] = {nz ? q1.x : q1.y, nz ? q2.z : q2.w};
      bool go[2];
      float tnear[2], tcull[2];
#pragma unroll
      for (int c = 0; c < 2; c++) {  // the same predicate, operation for operation, as the ordered walk
        float txn = (xn[c] - r.ox) * r.ix, txx = (xf[c] - r.ox) * r.ix;
        float tyn = (yn[c] - r.oy) * r.iy, tyx = (yf[c] - r.oy) * r.iy;
        float tzn = (zn[c] - r.oz) * r.iz, tzx = (zf[c] - r.oz) * r.iz;
        bool ok = (txn <= tyx) && (tyn <= txx);
        float txyn = fmaxf(txn, tyn), txyx = fminf(txx, tyx);
        ok = ok && (txyn <= tzx) && (tzn <= txyx);
        float tn = r.kz == 0 ? txn : (r.kz == 1 ? tyn : tzn);
        float tf = r.kz == 0 ? txx : (r.kz == 1 ? tyx : tzx);
        float margin = 1e-5f * fmaxf(fabsf(tn), fabsf(tf));
        tnear[c] = tn;
        tcull[c] = tn - margin;
        go[c] = ok && !(tf + margin < 0.f);
      }
      int lf0 = (go[0] && cref[0] < 0) ? ~cref[0] : -1, lf1 = (go[1] && cref[1] < 0) ? ~cref[1] : -1;
      float lc0 = tcull[0], lc1 = tcull[1];
      if (lf0 < 0) { lf0 = lf1; lc0 = lc1; lf1 = -1; }
#pragma unroll 1
      while (lf0 >= 0) {
        if (!(lc0 > best)) {
          const float4 *tp = tris + 3 * (size_t)lf0;
          float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
          if (COUNT) vl++;
          float t, b1, b2, b3;
          if (tri_test(r, t0, t1, t2, t, b1, b2, b3)) {
            if (!(t > best) && !(t < t2.y)) {
              bool take = true;
              if (cand && t == ct)  // the node's two leaves tie: the later one in reference order
                take = leaf_visit_rank(nodes, parent_leaf, lf0) >
                       leaf_visit_rank(nodes, parent_leaf, cleaf);
              if (take) {
                cand = true; ct = t; cb1 = b1; cb2 = b2; cb3 = b3; cleaf = lf0;
                cprim = __float_as_int(t2.z); ctol = t2.y;
                best = t;
              }
            }
          }
        }
        lf0 = lf1; lc0 = lc1; lf1 = -1;
      }
      const bool gl = go[0] && cref[0] >= 0 && !(tcull[0] > best);
      const bool gr = go[1] && cref[1] >= 0 && !(tcull[1] > best);
      const bool lfirst = tnear[0] <= tnear[1];
      npush = (gl ? 1 : 0) + (gr ? 1 : 0);
      const int tag = slot << COOP_NODE_BITS;
      if (gl && gr) { push_near = tag | (lfirst ? cref[0] : cref[1]); push_far = tag | (lfirst ? cref[1] : cref[0]); }
      else if (npush) push_near = tag | (gl ? cref[0] : cref[1]);
    }
    // ---- append the children that passed (far below near: near is popped first) -----------
    const unsigned m1 = __ballot_sync(FULL, npush >= 1), m2 = __ballot_sync(FULL, npush == 2);
    const unsigned lt = (1u << lane) - 1u;
    const int off = cnt + __popc(m1 & lt) + __popc(m2 & lt);
    if (npush == 2) { WL(off) = push_far; WL(off + 1) = push_near; }
    else if (npush == 1) WL(off) = push_near;
    cnt += __popc(m1) + __popc(m2);
    // ---- hits: merged one lane at a time into the ray's record (few per round) ------------
    unsigned cm = __ballot_sync(FULL, cand);
    while (cm) {
      const int j = __ffs(cm) - 1;
      cm &= cm - 1;
      if (lane == j) {
        const float best = rt[CR_BEST * COOP_SLOTS + slot];
        if (!(ct > best) && !(ct < ctol)) {
          bool take = true;
          const int hleaf = rti[CR_LEAF * COOP_SLOTS + slot];
          if (ct == rt[CR_T * COOP_SLOTS + slot] && hleaf >= 0)
            take = leaf_visit_rank(nodes, parent_leaf, cleaf) >
                   leaf_visit_rank(nodes, parent_leaf, hleaf);
          if (take) {
            rt[CR_T * COOP_SLOTS + slot] = ct; rti[CR_LEAF * COOP_SLOTS + slot] = cleaf;
            rt[CR_B1 * COOP_SLOTS + slot] = cb1; rt[CR_B2 * COOP_SLOTS + slot] = cb2;
            rt[CR_B3 * COOP_SLOTS + slot] = cb3; rti[CR_PRIM * COOP_SLOTS + slot] = cprim;
            rt[CR_BEST * COOP_SLOTS + slot] = ct;
          }
        }
      }
      __syncwarp();
    }
    __syncwarp();
  }
  // ---- hand the records back ---------------------------------------------------------------
  if (lane < nslots) {
    TraceLane R = L;
    R.t = rt[CR_T * COOP_SLOTS + lane]; R.leaf = rti[CR_LEAF * COOP_SLOTS + lane];
    R.b1 = rt[CR_B1 * COOP_SLOTS + lane]; R.b2 = rt[CR_B2 * COOP_SLOTS + lane];
    R.b3 = rt[CR_B3 * COOP_SLOTS + lane]; R.prim = rti[CR_PRIM * COOP_SLOTS + lane];
    src.store(rti[CR_RAY * COOP_SLOTS + lane], R);
  }
#undef WL
}
// One step of the ordered walk for a lane that holds a ray: test the two children of L.node,
// test leaves where they are found, descend / push / pop.  Returns true when the walk is over
// (the caller stores the record).
template <bool COUNT>
__device__ __forceinline__ bool tp_step(const float4 *__restrict__ nodes,
                                        const float4 *__restrict__ tris,
                                        const int *__restrict__ parent_leaf, TraceLane &L,
                                        int *stack, int *spill, unsigned long long *counters,
                                        unsigned &vi, unsigned &vl) {
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
    return finished;
}

// ---- deferred leaf tests (DEFER) ----------------------------------------------------------
// ncu (profiles/r01g_k_pt_trace, source page): 82 % of the node steps are followed by a pass of
// the ~95-instruction triangle test with 2.9 of 32 lanes on -- 43 % of the loop's issue slots.
// Here a step only QUEUES the leaves it finds, the walk goes on with the cull bound it has,
// and the warp runs a test pass -- every lane with a queued leaf tests its newest one -- only
// when some lane holds more than DEFER_MAX, and empties the queues when it returns to the
// refill point.  Closest hit with the visit-rank tie-break does not depend on the order of
// the tests (trace_core.cuh) and a stale bound only culls less (V_int 47.8 -> 49.2, V_leaf
// 4.98 -> 5.13 on C2), so the hits are bit-identical; a lane whose walk is over waits
// (L.node = -1) for its queue to drain.  C2: trace stage 32.8 -> 29.1 ms at 8 spp.
#ifndef DEFER_MAX_V
#define DEFER_MAX_V 4
#endif
#ifndef DEFER_LANES_V
#define DEFER_LANES_V 33
#endif
constexpr int DEFER_MAX = DEFER_MAX_V;      // a lane may keep this many leaves queued between passes (2 more arrive per step at most)
constexpr int DEFER_LANES = DEFER_LANES_V;  // ... and a pass also runs once this many lanes have one (33 = never)
// The queue is a per-lane LIFO in shared memory ([slot][thread], like the stack): a push is two
// predicated stores.  (A front-packed queue in registers was measured first: its select chains
// cost every step what the batching saved.)
constexpr int DEFER_SLOTS = DEFER_MAX_V + 2;
struct LeafQueue {
  int *qi;     // [slot * TP_BLOCK]: leaf
  float *qc;   // [slot * TP_BLOCK]: cull bound
  int cnt = 0;
  __device__ __forceinline__ LeafQueue(int *smem) : qi(smem + threadIdx.x), qc(reinterpret_cast<float *>(smem + DEFER_SLOTS * TP_BLOCK + threadIdx.x)) {}
  __device__ __forceinline__ void push(int lf, float lc) {
    if (lf >= 0) {
      qi[cnt * TP_BLOCK] = lf;
      qc[cnt * TP_BLOCK] = lc;
      cnt++;
    }
  }
  __device__ __forceinline__ void pop(int &lf, float &lc) {
    cnt--;
    lf = qi[cnt * TP_BLOCK];
    lc = qc[cnt * TP_BLOCK];
  }
};

// the triangle test of one queued leaf (the body of tp_step's leaf loop, same operations)
template <bool COUNT>
__device__ __forceinline__ void tp_leaf_test(const float4 *__restrict__ nodes,
                                             const float4 *__restrict__ tris,
                                             const int *__restrict__ parent_leaf, TraceLane &L,
                                             int lf0, float lc0, unsigned &vl) {
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
}

// tp_step without its leaf tests: the leaves found go to (lfA, lcA), (lfB, lcB)
template <bool COUNT>
__device__ __forceinline__ bool tp_node_step(const float4 *__restrict__ nodes, TraceLane &L,
                                             int *stack, int *spill,
                                             unsigned long long *counters, unsigned &vi,
                                             int &lfA, float &lcA, int &lfB, float &lcB) {
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
    lfA = (go[0] && cref[0] < 0) ? ~cref[0] : -1;
    lfB = (go[1] && cref[1] < 0) ? ~cref[1] : -1;
    lcA = tcull[0]; lcB = tcull[1];
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
    return finished;
}

// Source concept:
//   int count() const;                       number of rays in the queue
//   void load(int i, float4 &o, float4 &d);  ray i: {o, tmax}{unit d, -}
//   void store(int i, const TraceLane &l);   hit record of ray i
template <bool COUNT, class Source, bool COOP = false, bool DEFER = false>
__device__ __forceinline__ void trace_persistent(const float4 *__restrict__ nodes,
                                                 const float4 *__restrict__ tris,
                                                 const int *__restrict__ parent_leaf,
                                                 Source src, int *next_ray, int *stack_smem,
                                                 unsigned long long *counters,
                                                 int *coop_smem = nullptr,
                                                 int *defer_smem = nullptr) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int n = src.count();
  int *stack = stack_smem + threadIdx.x;  // [level][thread], stride TP_BLOCK
  int spill[TRACE_STACK - TP_STACK_SMEM];
  TraceLane L;
  L.ray = -1;
  unsigned vi = 0, vl = 0;
  bool exhausted = false;
  unsigned coop_live = 0;
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
    if (COOP && exhausted) {
      // the queue is empty and only a few long rays are left in this warp: leave the ordered
      // walk and finish them with all 32 lanes (coop_finish, after this loop -- NOT called from
      // inside it: a call here costs the hot loop 26 registers and 18 % of its speed)
      coop_live = __ballot_sync(FULL, L.ray >= 0);
      if (__popc(coop_live) <= COOP_SLOTS) {
        // (a lane deeper than the shared-memory levels counts as too much)
        int tot = L.ray >= 0 ? (L.sp < TP_STACK_SMEM ? 1 + L.sp : COOP_LIST) : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) tot += __shfl_xor_sync(FULL, tot, d);
        if (tot <= COOP_LIST - 128) break;  // else: very deep stacks, stay with the ordered walk
      }
      coop_live = 0;
    }
    // ---- traverse until enough lanes have gone idle ------------------------------------
    // (postponing leaf tests into separate warp-wide "intersect rounds" was measured and was
    //  10-20 % SLOWER on C2 -- the extra rounds and votes cost more than the divergence they
    //  remove with one triangle per leaf -- so leaves are tested where they are found.)
    if constexpr (!DEFER) {
      while (true) {
        if (L.ray >= 0) {
          if (tp_step<COUNT>(nodes, tris, parent_leaf, L, stack, spill, counters, vi, vl)) {
            src.store(L.ray, L);
            L.ray = -1;
          }
        }
        unsigned nidle = __popc(__ballot_sync(FULL, L.ray < 0));
        if (nidle == 32 || (!exhausted && nidle >= TP_REFILL)) break;
        if (COOP && exhausted && nidle >= 32 - COOP_SLOTS) break;
      }
    } else {
      LeafQueue Q(defer_smem);
      while (true) {
        if (L.ray >= 0 && L.node >= 0) {
          int lfA, lfB;
          float lcA, lcB;
          if (tp_node_step<COUNT>(nodes, L, stack, spill, counters, vi, lfA, lcA, lfB, lcB))
            L.node = -1;  // walk over; the record is stored once the queue is empty
          Q.push(lfA, lcA);
          Q.push(lfB, lcB);
        }
        bool pass = __ballot_sync(FULL, Q.cnt > DEFER_MAX) != 0;
        if (DEFER_LANES <= 32 && !pass) pass = __popc(__ballot_sync(FULL, Q.cnt > 0)) >= DEFER_LANES;
        while (pass) {
          if (Q.cnt > 0) {
            int lf;
            float lc;
            Q.pop(lf, lc);
            tp_leaf_test<COUNT>(nodes, tris, parent_leaf, L, lf, lc, vl);
          }
          pass = __ballot_sync(FULL, Q.cnt > DEFER_MAX) != 0;
        }
        if (L.ray >= 0 && L.node < 0 && Q.cnt == 0) {
          src.store(L.ray, L);
          L.ray = -1;
        }
        const int nwalk = __popc(__ballot_sync(FULL, L.ray >= 0 && L.node >= 0));
        if (nwalk == 0 || (!exhausted && 32 - nwalk >= TP_REFILL)) break;
        if (COOP && exhausted && nwalk <= COOP_SLOTS) break;
      }
      while (__ballot_sync(FULL, Q.cnt > 0)) {  // back at the refill point with empty queues
        if (Q.cnt > 0) {
          int lf;
          float lc;
          Q.pop(lf, lc);
          tp_leaf_test<COUNT>(nodes, tris, parent_leaf, L, lf, lc, vl);
        }
      }
      if (L.ray >= 0 && L.node < 0) {
        src.store(L.ray, L);
        L.ray = -1;
      }
    }
  }
  if (COOP && coop_live)
    coop_finish<COUNT>(nodes, tris, parent_leaf, src, L, coop_live, stack,
                       coop_smem + (threadIdx.x >> 5) * COOP_WARP_WORDS, vi, vl);
  if (COUNT) {
    atomicAdd(counters, (unsigned long long)vi);
    atomicAdd(counters + 1, (unsigned long long)vl);
  }
}

// ---- streaming trace (the path tracer's barrier-free pipeline, render.cu) -----------------
// The same ordered walk, fed from a RING that the shade kernel fills while this kernel runs:
// there is no "end of the queue", so there is no tail to wait for.  An idle lane reserves the
// next ring position (warp-aggregated atomicAdd on the ring's head) and polls that cell between
// traversal steps until a slot appears in it; a finished ray is handed to the shade kernel by
// the source's store().  The launch ends when the source says every pixel is finished.
// Source concept:
//   int  reserve(int cnt);                    first of cnt consecutive ring positions
//   int  poll(int pos);                       slot published at that position, or -1 (not yet)
//   void load(int slot, float4 &o, float4 &d);
//   int  store(int slot, const TraceLane &l); writes the hit record, returns the ring it goes to
//   void publish(int slot, int ring);         warp-wide: lanes with slot >= 0 hand it over
//   bool finished();                          every pixel done (or the pipeline was aborted)
//   void abort();
// Handing a slot over costs a fence and two atomics, i.e. several round trips to L2 during
// which the whole warp stands still; so a finished lane keeps its slot until the warp comes
// back to the refill point (>= TS_REFILL lanes without a ray, or a poll of reserved cells is
// due) and the warp publishes all of them together, one atomic per ring.
constexpr int TS_POLL_STEPS = 8;                          // steps between polls of reserved cells
#ifndef TS_REFILL_V
#define TS_REFILL_V 12
#endif
constexpr int TS_REFILL = TS_REFILL_V;                    // lanes without a ray that end a traversal stretch
constexpr long long TS_WATCHDOG_CYCLES = 20000000000ll;   // ~10 s without any ray: give up loudly
template <bool COUNT, class Source>
__device__ __forceinline__ void trace_stream(const float4 *__restrict__ nodes,
                                             const float4 *__restrict__ tris,
                                             const int *__restrict__ parent_leaf, Source src,
                                             int *stack_smem, unsigned long long *counters) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  int *stack = stack_smem + threadIdx.x;
  int spill[TRACE_STACK - TP_STACK_SMEM];
  TraceLane L;
  L.ray = -1;
  int pend = -1;      // ring position this lane has reserved and is waiting on
  int fin_slot = -1;  // finished, not yet handed to the shade kernel
  int fin_ring = 0;
  unsigned vi = 0, vl = 0;
  long long idle_since = 0;
  while (true) {
    if (__ballot_sync(FULL, fin_slot >= 0)) {
      src.publish(fin_slot, fin_ring);
      fin_slot = -1;
    }
    const unsigned want = __ballot_sync(FULL, L.ray < 0 && pend < 0);
    const unsigned busy = __ballot_sync(FULL, L.ray >= 0);
    if (want && (__popc(want) >= TS_REFILL || busy == 0)) {
      const int cnt = __popc(want), leader = __ffs(want) - 1;
      int base = 0;
      if (lane == leader) base = src.reserve(cnt);
      base = __shfl_sync(FULL, base, leader);
      if (L.ray < 0 && pend < 0) pend = base + __popc(want & ((1u << lane) - 1u));
    }
    if (pend >= 0) {
      const int slot = src.poll(pend);
      if (slot >= 0) {
        float4 ro, rd;
        src.load(slot, ro, rd);
        RaySetup r = make_ray(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z);
        L.ox = r.ox; L.oy = r.oy; L.oz = r.oz; L.ix = r.ix; L.iy = r.iy; L.iz = r.iz;
        L.sx = r.sx; L.sy = r.sy; L.sz = r.sz; L.kz = r.kz;
        L.neg = (r.ix < 0.f ? 1u : 0u) | (r.iy < 0.f ? 2u : 0u) | (r.iz < 0.f ? 4u : 0u);
        L.best = ro.w;
        L.t = INFINITY; L.leaf = -1; L.prim = -1; L.b1 = L.b2 = L.b3 = 0.f;
        L.node = 0; L.sp = 0;
        L.ray = slot;
        pend = -1;
      }
    }
    if (__ballot_sync(FULL, L.ray >= 0) == 0) {
      // nothing to walk: either the frame is over or the shade kernel has not produced yet
      if (src.finished()) break;
      const long long now = clock64();
      if (idle_since == 0) idle_since = now;
      if (now - idle_since > TS_WATCHDOG_CYCLES) {
        src.abort();
        break;
      }
      __nanosleep(100);
      continue;
    }
    idle_since = 0;
    int steps = __ballot_sync(FULL, pend >= 0) ? TS_POLL_STEPS : (1 << 30);
    while (true) {
      if (L.ray >= 0) {
        if (tp_step<COUNT>(nodes, tris, parent_leaf, L, stack, spill, counters, vi, vl)) {
          fin_ring = src.store(L.ray, L);
          fin_slot = L.ray;
          L.ray = -1;
        }
      }
      const int nwant = __popc(__ballot_sync(FULL, L.ray < 0 && pend < 0));
      if (nwant >= TS_REFILL || --steps <= 0) break;
    }
  }
  if (COUNT) {
    atomicAdd(counters, (unsigned long long)vi);
    atomicAdd(counters + 1, (unsigned long long)vl);
  }
}
