// trace_core.cuh -- closest-hit BVH traversal + watertight ray/triangle test (device).
//
// Replaces traverse_bvh (src/model/bvh/bvh_traversal.h:12-84), BoundingBox::is_intersection
// (src/model/grid/bounding_box.h:217-276) and Triangle::hit
// (src/model/geometry/triangle.h:343-408).
//
// PARITY ARGUMENT (DESIGN.md "Trace").  The reference visits every leaf whose own box and
// whose ancestors' boxes pass its no-cull slab predicate, in a fixed left-first order, and
// keeps the LAST visited among hits of equal t (t > t_max rejects, so equality replaces).
// This kernel returns the same triangle while visiting far fewer nodes:
//   1. a child is entered only if the reference's exact predicate passes (same fp32 ops);
//   2. it is additionally culled only by the DOMINANT-AXIS slab [tn, tf] of its box:
//      the accepted t of any triangle below is a convex combination of (p_i[kz]-o[kz])*sz,
//      each of which lies inside that slab by monotonicity of fp32 sub/mul, so
//      tn - margin > t_best or tf + margin < 0 can never discard an acceptable hit;
//   3. equal-t hits are resolved by the reference's visit rank, computed on demand from
//      the parent links (rank = 2 * preorder(parent) + is_right).
#pragma once
#include "common.cuh"
#include "vecmath.cuh"

struct RaySetup {
  float ox, oy, oz;     // origin
  float dx, dy, dz;     // unit direction
  float ix, iy, iz;     // 1.0 / d (IEEE division, bounding_box.h:220)
  float sx, sy, sz;     // Woop shear (triangle.h:358-360)
  int kz;               // dominant axis (max_dimension, vec3.h:200-208)
};

struct HitResult {
  float t;
  int leaf;          // sorted position, -1 = miss
  float b1, b2, b3;  // e_i * inv_det (triangle.h:392-393)
  int prim;          // original triangle index
};

__device__ __forceinline__ RaySetup make_ray(float ox, float oy, float oz, float dx, float dy,
                                             float dz) {
  RaySetup r;
  r.ox = ox; r.oy = oy; r.oz = oz;
  r.dx = dx; r.dy = dy; r.dz = dz;
  r.ix = 1.0f / dx; r.iy = 1.0f / dy; r.iz = 1.0f / dz;
  float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
  int kz = (ax >= ay && ax >= az) ? 0 : ((ay >= ax && ay >= az) ? 1 : 2);
  r.kz = kz;
  // (kx, ky, kz) is a rotation of (0, 1, 2): no winding swap (triangle.h:349-351)
  float pdx, pdy, pdz;
  if (kz == 0) { pdx = dy; pdy = dz; pdz = dx; }
  else if (kz == 1) { pdx = dz; pdy = dx; pdz = dy; }
  else { pdx = dx; pdy = dy; pdz = dz; }
  r.sx = -pdx / pdz;
  r.sy = -pdy / pdz;
  r.sz = 1.0f / pdz;
  return r;
}

__device__ __forceinline__ void permute3(int kz, float x, float y, float z, float &a, float &b,
                                         float &c) {
  if (kz == 0) { a = y; b = z; c = x; }
  else if (kz == 1) { a = z; b = x; c = y; }
  else { a = x; b = y; c = z; }
}

// Triangle::hit up to the t computation.  Returns false on the sign / det rejections.
__device__ __forceinline__ bool tri_test(const RaySetup &r, float4 t0, float4 t1, float4 t2,
                                         float &t, float &b1, float &b2, float &b3) {
  float ax, ay, az, bx, by, bz, cx, cy, cz;
  // p_it = permute(point_i - ray.p0): subtract first, then permute (same values)
  permute3(r.kz, t0.x - r.ox, t0.y - r.oy, t0.z - r.oz, ax, ay, az);
  permute3(r.kz, t0.w - r.ox, t1.x - r.oy, t1.y - r.oz, bx, by, bz);
  permute3(r.kz, t1.z - r.ox, t1.w - r.oy, t2.x - r.oz, cx, cy, cz);
  ax = ax + r.sx * az; ay = ay + r.sy * az;
  bx = bx + r.sx * bz; by = by + r.sy * bz;
  cx = cx + r.sx * cz; cy = cy + r.sy * cz;
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
  az = az * r.sz; bz = bz * r.sz; cz = cz * r.sz;
  float t_scaled = e1 * az + e2 * bz + e3 * cz;
  float inv_det = 1.0f / det;
  b1 = e1 * inv_det; b2 = e2 * inv_det; b3 = e3 * inv_det;
  t = t_scaled * inv_det;
  return true;
}

// Reference visit rank of a leaf: leaf children are tested when their parent is processed
// (left before right), internal nodes in left-first preorder.  preorder(X) = range_first(X)
// + number of left-child edges on the root..X path.
static __device__ __noinline__ int leaf_visit_rank(const float4 *__restrict__ nodes,
                                            const int *__restrict__ parent_leaf, int leaf) {
  int pe = __ldg(parent_leaf + leaf);
  int par = pe >> 1;
  const int *q3 = reinterpret_cast<const int *>(nodes + 4 * (size_t)par + 3);
  int pre = __ldg(q3 + 3);
  int up = __ldg(q3 + 2);
  while (up >= 0) {
    if ((up & 1) == 0) pre++;
    up = __ldg(reinterpret_cast<const int *>(nodes + 4 * (size_t)(up >> 1) + 3) + 2);
  }
  return 2 * pre + (pe & 1);
}

constexpr int TRACE_STACK = 64;  // >= 30 + ceil(log2 N) levels

// One ray, closest hit.  stack: per-thread column of a shared-memory array, stride = blockDim.
template <bool COUNT>
__device__ __forceinline__ HitResult trace_ray(const float4 *__restrict__ nodes,
                                               const float4 *__restrict__ tris,
                                               const int *__restrict__ parent_leaf,
                                               const RaySetup &r, float tmax, int *stack,
                                               int stride, unsigned long long *cnt_int,
                                               unsigned long long *cnt_leaf, int *overflow) {
  HitResult h;
  h.t = INFINITY; h.leaf = -1; h.b1 = h.b2 = h.b3 = 0.f; h.prim = -1;
  float best = tmax;  // cull bound; equals h.t once something was hit
  int sp = 0;
  int node = 0;
  // per-ray selection of the dominant-axis slab
  const int kz = r.kz;
  unsigned vi = 0, vl = 0;
  while (true) {
    const float4 *q = nodes + 4 * (size_t)node;
    float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
    if (COUNT) vi++;
    int cref[2] = {__float_as_int(q3.x), __float_as_int(q3.y)};
    // child boxes: L = (q0.x q0.y | q0.z q0.w | q1.x q1.y), R = (q1.z q1.w | q2.x q2.y | q2.z q2.w)
    float bxn[2] = {q0.x, q1.z}, bxx[2] = {q0.y, q1.w};
    float byn[2] = {q0.z, q2.x}, byx[2] = {q0.w, q2.y};
    float bzn[2] = {q1.x, q2.z}, bzx[2] = {q1.y, q2.w};
    bool go[2];
    float tnear[2], tcull[2];
#pragma unroll
    for (int c = 0; c < 2; c++) {
      // _compute_t_{x,y,z}_range + _are_intersecting (bounding_box.h:61-69,217-276)
      float a1 = (bxn[c] - r.ox) * r.ix, a2 = (bxx[c] - r.ox) * r.ix;
      float txn = a1 < a2 ? a1 : a2, txx = a1 < a2 ? a2 : a1;
      float c1 = (byn[c] - r.oy) * r.iy, c2 = (byx[c] - r.oy) * r.iy;
      float tyn = c1 < c2 ? c1 : c2, tyx = c1 < c2 ? c2 : c1;
      float d1 = (bzn[c] - r.oz) * r.iz, d2 = (bzx[c] - r.oz) * r.iz;
      float tzn = d1 < d2 ? d1 : d2, tzx = d1 < d2 ? d2 : d1;
      bool ok = (txn <= tyx) && (tyn <= txx);
      float txyn = fmaxf(txn, tyn), txyx = fminf(txx, tyx);
      ok = ok && (txyn <= tzx) && (tzn <= txyx);
      // dominant-axis slab [tn, tf] for the conservative cull (header comment, point 2)
      float tn = kz == 0 ? txn : (kz == 1 ? tyn : tzn);
      float tf = kz == 0 ? txx : (kz == 1 ? tyx : tzx);
      float margin = 1e-5f * fmaxf(fabsf(tn), fabsf(tf));
      tnear[c] = tn;
      tcull[c] = tn - margin;
      go[c] = ok && !(tf + margin < 0.f);
    }
    // leaves first (left, then right), as the reference does at each node
#pragma unroll
    for (int c = 0; c < 2; c++) {
      if (go[c] && cref[c] < 0) {
        go[c] = false;
        if (tcull[c] > best) continue;
        int leaf = ~cref[c];
        const float4 *tp = tris + 3 * (size_t)leaf;
        float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
        if (COUNT) vl++;
        float t, b1, b2, b3;
        if (tri_test(r, t0, t1, t2, t, b1, b2, b3)) {
          float tol = t2.y;
          // reference: reject if (t > t_max || t < tolerance)
          if (!(t > best) && !(t < tol)) {
            bool take = true;
            if (t == h.t && h.leaf >= 0)  // equal t: the later one in reference order wins
              take = leaf_visit_rank(nodes, parent_leaf, leaf) >
                     leaf_visit_rank(nodes, parent_leaf, h.leaf);
            if (take) {
              h.t = t; h.leaf = leaf; h.b1 = b1; h.b2 = b2; h.b3 = b3;
              h.prim = __float_as_int(t2.z);
              best = t;
            }
          }
        }
      }
    }
    // internal children: far cull against the updated best, nearer one first
    bool gl = go[0] && !(tcull[0] > best);
    bool gr = go[1] && !(tcull[1] > best);
    if (gl && gr) {
      bool lfirst = tnear[0] <= tnear[1];
      int nearc = lfirst ? cref[0] : cref[1], farc = lfirst ? cref[1] : cref[0];
      if (sp < TRACE_STACK) { stack[sp * stride] = farc; sp++; } else *overflow = 1;
      node = nearc;
    } else if (gl || gr) {
      node = gl ? cref[0] : cref[1];
    } else {
      if (sp == 0) break;
      sp--;
      node = stack[sp * stride];
    }
  }
  if (COUNT) {
    atomicAdd(cnt_int, (unsigned long long)vi);
    atomicAdd(cnt_leaf, (unsigned long long)vl);
  }
  return h;
}
