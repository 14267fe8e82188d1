/*
 * oracle.c -- TEST INFRASTRUCTURE ONLY: plain-C CPU restatement of the reference's hot path.
 *
 * This file is the checker, never the product: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (gpu_pathtracer_b200/) has no CPU fallback and never links or imports anything here.
 *
 * Parity status: PINNED.  The reference has no tests or golden vectors of its own
 * (SURVEY.md section 4), so every function below is checked bit-for-bit against the
 * reference's own headers replayed on the host (oracle/_ref/libref_replay.so, built from
 * /root/reference/src by oracle/Makefile) in tests/test_oracle_vs_reference.py, and against
 * the golden fixtures under tests/golden/ that were generated from that replay.
 *
 * Arithmetic: IEEE fp32, no FMA contraction (-ffp-contract=off), operations in the same
 * order as the reference.  Each function cites the reference file:line it restates.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/gpt.h"

#define SMALL 1e-5f /* SMALL_DOUBLE, src/param.h:6 */

typedef struct { float x, y, z; } v3;

static v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 vscale(float t, v3 a) { return V(t * a.x, t * a.y, t * a.z); }
static v3 vdiv(v3 a, float t) { return V(a.x / t, a.y / t, a.z / t); }
static v3 vmul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
/* vec3::length (vec3.h:35): sqrt(e0*e0 + e1*e1 + e2*e2), left to right */
static float vlen(v3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
static float vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* cross (vec3.h:128-132) */
static v3 vcross(v3 a, v3 b) {
  return V(a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x);
}
static v3 vunit(v3 a) { return vdiv(a, vlen(a)); } /* unit_vector, vec3.h:224-226 */
static int vnan(v3 a) { return isnan(a.x) || isnan(a.y) || isnan(a.z); }

typedef struct {
  int n;                 /* leaves */
  int *prim;             /* sorted position -> original triangle index */
  uint32_t *morton;      /* sorted codes */
  int *left, *right;     /* child refs: >= 0 internal, < 0 leaf ~c */
  int *parent_internal, *parent_leaf;
  float *node_aabb;      /* 6 (n-1) */
  float world[6];
} OrcBvh;

typedef struct OrcCtx {
  /* inputs (owned copies) */
  int nv, nvn, nvt, nt, nmat;
  float *x, *y, *z, *xn, *yn, *zn, *xt, *yt;
  int *p1, *p2, *p3, *n1, *n2, *n3, *t1, *t2, *t3, *mat_idx;
  float *mat_f[23];
  int *mat_i[16];
  float *tex[3];
  int64_t tex_len;
  /* derived per triangle (original order) */
  float *tolerance, *aabb, *normal, *area, *energy, *tangents;
  v3 *cn; /* 3 corner normals per triangle after the fallback */
  int *flags;
  OrcBvh bvh[3];
  double last_ms;
} OrcCtx;

static void *dup_mem(const void *src, size_t bytes) {
  void *p = malloc(bytes ? bytes : 4);
  if (bytes) memcpy(p, src, bytes);
  else memset(p, 0, 4);
  return p;
}

int orc_create(OrcCtx **out) {
  *out = (OrcCtx *)calloc(1, sizeof(OrcCtx));
  return *out ? GPT_OK : GPT_E_INVALID;
}

static void free_bvh(OrcBvh *b) {
  free(b->prim); free(b->morton); free(b->left); free(b->right);
  free(b->parent_internal); free(b->parent_leaf); free(b->node_aabb);
  memset(b, 0, sizeof(*b));
}

void orc_destroy(OrcCtx *c) {
  if (!c) return;
  float **fp[] = {&c->x, &c->y, &c->z, &c->xn, &c->yn, &c->zn, &c->xt, &c->yt, &c->tolerance,
                  &c->aabb, &c->normal, &c->area, &c->energy, &c->tangents};
  for (size_t i = 0; i < sizeof(fp) / sizeof(fp[0]); i++) free(*fp[i]);
  int **ip[] = {&c->p1, &c->p2, &c->p3, &c->n1, &c->n2, &c->n3, &c->t1, &c->t2, &c->t3,
                &c->mat_idx, &c->flags};
  for (size_t i = 0; i < sizeof(ip) / sizeof(ip[0]); i++) free(*ip[i]);
  for (int i = 0; i < 23; i++) free(c->mat_f[i]);
  for (int i = 0; i < 16; i++) free(c->mat_i[i]);
  for (int i = 0; i < 3; i++) { free(c->tex[i]); free_bvh(&c->bvh[i]); }
  free(c->cn);
  free(c);
}

int orc_set_mesh(OrcCtx *c, const GptMesh *m) {
  c->nv = m->num_vertices; c->nvn = m->num_normals; c->nvt = m->num_texcoords;
  c->nt = m->num_triangles;
  size_t f = sizeof(float), i = sizeof(int);
  c->x = dup_mem(m->x, f * c->nv); c->y = dup_mem(m->y, f * c->nv); c->z = dup_mem(m->z, f * c->nv);
  c->xn = dup_mem(m->x_norm, f * c->nvn); c->yn = dup_mem(m->y_norm, f * c->nvn);
  c->zn = dup_mem(m->z_norm, f * c->nvn);
  c->xt = dup_mem(m->x_tex, f * c->nvt); c->yt = dup_mem(m->y_tex, f * c->nvt);
  c->p1 = dup_mem(m->point_1_idx, i * c->nt); c->p2 = dup_mem(m->point_2_idx, i * c->nt);
  c->p3 = dup_mem(m->point_3_idx, i * c->nt);
  c->n1 = dup_mem(m->norm_1_idx, i * c->nt); c->n2 = dup_mem(m->norm_2_idx, i * c->nt);
  c->n3 = dup_mem(m->norm_3_idx, i * c->nt);
  c->t1 = dup_mem(m->tex_1_idx, i * c->nt); c->t2 = dup_mem(m->tex_2_idx, i * c->nt);
  c->t3 = dup_mem(m->tex_3_idx, i * c->nt);
  c->mat_idx = dup_mem(m->material_idx, i * c->nt);
  return GPT_OK;
}

/* material float / int columns in GptMaterials declaration order */
enum { KA_X, KA_Y, KA_Z, KD_X, KD_Y, KD_Z, KS_X, KS_Y, KS_Z, KE_X, KE_Y, KE_Z, TF_X, TF_Y,
       TF_Z, PATH_LENGTH, T_R, N_S, N_I, BM, SCAT, ABSORB, G_ };
enum { PRIORITY, H_DIFF, W_DIFF, O_DIFF, H_SPEC, W_SPEC, O_SPEC, H_EMIS, W_EMIS, O_EMIS,
       H_NS, W_NS, O_NS, H_BUMP, W_BUMP, O_BUMP };

int orc_set_materials(OrcCtx *c, const GptMaterials *m) {
  const float *const *fsrc = (const float *const *)&m->ka_x;
  const int32_t *const *isrc = (const int32_t *const *)&m->material_priority;
  c->nmat = m->num_materials;
  for (int k = 0; k < 23; k++) c->mat_f[k] = dup_mem(fsrc[k], sizeof(float) * c->nmat);
  for (int k = 0; k < 16; k++) c->mat_i[k] = dup_mem(isrc[k], sizeof(int) * c->nmat);
  return GPT_OK;
}

int orc_set_textures(OrcCtx *c, const float *r, const float *g, const float *b, int64_t n) {
  c->tex[0] = dup_mem(r, sizeof(float) * n);
  c->tex[1] = dup_mem(g, sizeof(float) * n);
  c->tex[2] = dup_mem(b, sizeof(float) * n);
  c->tex_len = n;
  return GPT_OK;
}

/* ---- texture fetch: Material::_get_texture (src/model/material/material.h:653-673) ------ */
static v3 tex_fetch(const OrcCtx *c, int mat, int h_col, float u, float v, v3 filter) {
  int h = c->mat_i[h_col][mat], w = c->mat_i[h_col + 1][mat], off = c->mat_i[h_col + 2][mat];
  int iu = (int)floorf((u - floorf(u)) * (float)(w - 1));
  int iv = (int)floorf((v - floorf(v)) * (float)(h - 1));
  int idx = off + w * iv + iu;
  return vmul(V(c->tex[0][idx], c->tex[1][idx], c->tex[2][idx]), filter);
}
static v3 mat_emission(const OrcCtx *c, int m, float u, float v) { /* material.h:675-681 */
  return tex_fetch(c, m, H_EMIS, u, v, V(c->mat_f[KE_X][m], c->mat_f[KE_Y][m], c->mat_f[KE_Z][m]));
}

/* ---- per-triangle derived data: Triangle::Triangle (src/model/geometry/triangle.h:254-305),
 *      _compute_tolerance (:307-321), _compute_triangle_area (:443-452),
 *      _compute_bounding_box (:226-252), _check_if_light_emitting (:123-135),
 *      _compute_energy (:112-121) ---------------------------------------------------------- */
static void tri_tangent(v3 p1, v3 p2, v3 p3, float u1, float v1, float u2, float v2, float u3,
                        float v3_, v3 *t, v3 *b) { /* _compute_tangent, triangle.h:174-185 */
  v3 e1 = vsub(p2, p1), e2 = vsub(p3, p1);
  float x1 = u2 - u1, x2 = u3 - u1, y1 = v2 - v1, y2 = v3_ - v1;
  float r = 1.0f / (x1 * y2 - x2 * y1);
  *t = vscale(r, vsub(vscale(y2, e1), vscale(y1, e2)));
  *b = vscale(r, vsub(vscale(x1, e2), vscale(x2, e1)));
}

static void preprocess(OrcCtx *c) {
  int n = c->nt;
  c->tolerance = malloc(sizeof(float) * n); c->aabb = malloc(sizeof(float) * 6 * n);
  c->normal = malloc(sizeof(float) * 3 * n); c->area = malloc(sizeof(float) * n);
  c->energy = malloc(sizeof(float) * 3 * n); c->tangents = malloc(sizeof(float) * 9 * n);
  c->flags = malloc(sizeof(int) * n); c->cn = malloc(sizeof(v3) * 3 * n);
  v3 *vt = calloc(c->nv > 0 ? c->nv : 1, sizeof(v3));
  for (int i = 0; i < n; i++) {
    int a = c->p1[i], b = c->p2[i], d = c->p3[i], m = c->mat_idx[i];
    v3 p1 = V(c->x[a], c->y[a], c->z[a]), p2 = V(c->x[b], c->y[b], c->z[b]),
       p3 = V(c->x[d], c->y[d], c->z[d]);
    float d12 = vlen(vsub(p1, p2)), d13 = vlen(vsub(p1, p3)), d23 = vlen(vsub(p2, p3));
    float tl = d12 < d13 ? d12 : d13;
    if (d23 < tl) tl = d23;
    float tol = fminf(SMALL, tl / 100.f);
    float s = (d12 + d13 + d23) / 2.0f;
    float area = fabsf(sqrtf(s * (s - d12) * (s - d13) * (s - d23)));
    v3 nrm = vunit(vcross(vsub(p2, p1), vsub(p3, p1)));
    v3 n1r = V(c->xn[c->n1[i]], c->yn[c->n1[i]], c->zn[c->n1[i]]);
    v3 n2r = V(c->xn[c->n2[i]], c->yn[c->n2[i]], c->zn[c->n2[i]]);
    v3 n3r = V(c->xn[c->n3[i]], c->yn[c->n3[i]], c->zn[c->n3[i]]);
    v3 n1 = vunit(n1r), n2 = vunit(n2r), n3 = vunit(n3r);
    float u1 = c->xt[c->t1[i]], v1 = c->yt[c->t1[i]], u2 = c->xt[c->t2[i]], v2 = c->yt[c->t2[i]],
          u3 = c->xt[c->t3[i]], v3_ = c->yt[c->t3[i]];
    v3 em1 = mat_emission(c, m, u1, v1), em2 = mat_emission(c, m, u2, v2),
       em3 = mat_emission(c, m, u3, v3_);
    int light = vlen(em1) > tol || vlen(em2) > tol || vlen(em3) > tol;
    int transparent = c->mat_f[T_R][m] > 0.f;
    int sss = c->mat_f[PATH_LENGTH][m] > 0.f;
    if (vlen(n1r) < tol || vlen(n2r) < tol || vlen(n3r) < tol) n1 = n2 = n3 = nrm;
    float *bb = c->aabb + 6 * i;
    bb[0] = fminf(fminf(p1.x, p2.x), p3.x) - tol; bb[1] = fmaxf(fmaxf(p1.x, p2.x), p3.x) + tol;
    bb[2] = fminf(fminf(p1.y, p2.y), p3.y) - tol; bb[3] = fmaxf(fmaxf(p1.y, p2.y), p3.y) + tol;
    bb[4] = fminf(fminf(p1.z, p2.z), p3.z) - tol; bb[5] = fmaxf(fmaxf(p1.z, p2.z), p3.z) + tol;
    c->tolerance[i] = tol; c->area[i] = area;
    c->normal[3 * i] = nrm.x; c->normal[3 * i + 1] = nrm.y; c->normal[3 * i + 2] = nrm.z;
    v3 avg = vdiv(vadd(vadd(em1, em2), em3), 3.0f);
    v3 en = vscale(area, vscale(6.283185307179586f, avg)); /* (float)(2 * M_PI) * avg * area */
    c->energy[3 * i] = en.x; c->energy[3 * i + 1] = en.y; c->energy[3 * i + 2] = en.z;
    c->flags[i] = (light ? GPT_TRI_LIGHT : 0) | (transparent ? GPT_TRI_TRANSPARENT : 0) |
                  (sss ? GPT_TRI_SSS : 0);
    c->cn[3 * i] = n1; c->cn[3 * i + 1] = n2; c->cn[3 * i + 2] = n3;
    /* sum_up_tangent_and_bitangent (triangle_operations.h:34-61): serial, in triangle order */
    v3 t, bt;
    tri_tangent(p1, p2, p3, u1, v1, u2, v2, u3, v3_, &t, &bt);
    if (!vnan(t) && !vnan(bt)) {
      vt[a] = vadd(vt[a], t); vt[b] = vadd(vt[b], t); vt[d] = vadd(vt[d], t);
    }
  }
  /* assign_tangent (triangle_operations.h:18-32, triangle.h:160-172) */
  for (int i = 0; i < n; i++) {
    int vi[3] = {c->p1[i], c->p2[i], c->p3[i]};
    for (int k = 0; k < 3; k++) {
      v3 t = vunit(vt[vi[k]]), nk = c->cn[3 * i + k];
      v3 g = vunit(vsub(t, vscale(vdot(nk, t), nk)));
      c->tangents[9 * i + 3 * k] = g.x; c->tangents[9 * i + 3 * k + 1] = g.y;
      c->tangents[9 * i + 3 * k + 2] = g.z;
    }
  }
  free(vt);
}

/* ---- LBVH: src/main.cu:826-918 ------------------------------------------------------------ */
static uint32_t expand_bits(uint32_t v) { /* src/util/bvh_util.h:4-10 */
  v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
static uint32_t morton3(float x, float y, float z) { /* src/util/bvh_util.h:12-20 */
  x = fminf(fmaxf(x * 1024.0f, 0.0f), 1023.0f); y = fminf(fmaxf(y * 1024.0f, 0.0f), 1023.0f);
  z = fminf(fmaxf(z * 1024.0f, 0.0f), 1023.0f);
  return expand_bits((uint32_t)x) * 4 + expand_bits((uint32_t)y) * 2 + expand_bits((uint32_t)z);
}
static int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
static int lcp(const uint32_t *k, int i, int j, int n) { /* src/util/bvh_util.h:22-31 */
  if (i < 0 || j < 0 || i >= n || j >= n) return -1;
  int extra = 0;
  if (k[i] == k[j]) extra = clz32((uint32_t)i ^ (uint32_t)j);
  return clz32(k[i] ^ k[j]) + extra;
}

/* stable merge sort of (key, value) by key == thrust::stable_sort (src/main.cu:846-855) */
static void stable_sort_pairs(uint32_t *k, int *v, int n) {
  uint32_t *k2 = malloc(sizeof(uint32_t) * (n ? n : 1));
  int *v2 = malloc(sizeof(int) * (n ? n : 1));
  for (int w = 1; w < n; w *= 2) {
    for (int lo = 0; lo < n; lo += 2 * w) {
      int mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
      int a = lo, b = mid, o = lo;
      while (a < mid && b < hi) {
        if (k[b] < k[a]) { k2[o] = k[b]; v2[o++] = v[b++]; }
        else { k2[o] = k[a]; v2[o++] = v[a++]; }
      }
      while (a < mid) { k2[o] = k[a]; v2[o++] = v[a++]; }
      while (b < hi) { k2[o] = k[b]; v2[o++] = v[b++]; }
    }
    memcpy(k, k2, sizeof(uint32_t) * n); memcpy(v, v2, sizeof(int) * n);
  }
  free(k2); free(v2);
}

static void bb_union(const float *a, const float *b, float *o) { /* bounding_box.h:329-339 */
  o[0] = fminf(a[0], b[0]); o[1] = fmaxf(a[1], b[1]); o[2] = fminf(a[2], b[2]);
  o[3] = fmaxf(a[3], b[3]); o[4] = fminf(a[4], b[4]); o[5] = fmaxf(a[5], b[5]);
}
static const float *child_box(const OrcCtx *c, const OrcBvh *b, int ref) {
  return ref >= 0 ? b->node_aabb + 6 * ref : c->aabb + 6 * b->prim[~ref];
}
/* compute_node_bounding_boxes (bvh_building.h:221-262) as a post-order recursion-free walk:
 * the union is order independent, so any bottom-up order gives the reference's boxes */
static void refit(const OrcCtx *c, OrcBvh *b) {
  int n = b->n;
  int *pending = calloc(n, sizeof(int));
  for (int j = 0; j < n; j++) {
    int p = b->parent_leaf[j];
    while (p >= 0 && ++pending[p] == 2) {
      bb_union(child_box(c, b, b->left[p]), child_box(c, b, b->right[p]), b->node_aabb + 6 * p);
      p = b->parent_internal[p];
    }
  }
  free(pending);
}

/* One LBVH over `list[0..n)` (original triangle indices). */
static void build_bvh(OrcCtx *c, OrcBvh *b, const int *list, int n) {
  free_bvh(b);
  b->n = n;
  if (n < 2) return;
  b->prim = malloc(sizeof(int) * n); b->morton = malloc(sizeof(uint32_t) * n);
  b->left = malloc(sizeof(int) * n); b->right = malloc(sizeof(int) * n);
  b->parent_internal = malloc(sizeof(int) * n); b->parent_leaf = malloc(sizeof(int) * n);
  b->node_aabb = malloc(sizeof(float) * 6 * n);
  /* compute_world_bounding_box (bounding_box_operations.h:12-45) */
  float w[6] = {INFINITY, -INFINITY, INFINITY, -INFINITY, INFINITY, -INFINITY};
  for (int i = 0; i < n; i++) {
    const float *bb = c->aabb + 6 * list[i];
    w[0] = fminf(w[0], bb[0]); w[1] = fmaxf(w[1], bb[1]); w[2] = fminf(w[2], bb[2]);
    w[3] = fmaxf(w[3], bb[3]); w[4] = fminf(w[4], bb[4]); w[5] = fmaxf(w[5], bb[5]);
  }
  w[0] -= SMALL; w[1] += SMALL; w[2] -= SMALL; w[3] += SMALL; w[4] -= SMALL; w[5] += SMALL;
  memcpy(b->world, w, sizeof(w));
  float lx = w[1] - w[0], ly = w[3] - w[2], lz = w[5] - w[4];
  /* compute_normalized_center + compute_bb_morton_3d (bounding_box.h:126-141,171-173) */
  for (int i = 0; i < n; i++) {
    const float *bb = c->aabb + 6 * list[i];
    float cx = 0.5f * (bb[0] + bb[1]), cy = 0.5f * (bb[2] + bb[3]), cz = 0.5f * (bb[4] + bb[5]);
    b->morton[i] = morton3((cx - w[0]) / lx, (cy - w[2]) / ly, (cz - w[4]) / lz);
    b->prim[i] = list[i];
  }
  stable_sort_pairs(b->morton, b->prim, n);
  /* set_node_relationship (bvh_building.h:88-185), with the reference's own linear walks */
  const uint32_t *k = b->morton;
  b->parent_internal[0] = -1;
  for (int idx = 0; idx < n - 1; idx++) {
    int d = lcp(k, idx, idx + 1, n) < lcp(k, idx, idx - 1, n) ? -1 : 1;
    int start = idx, end = start + d, min_delta = lcp(k, idx, idx - d, n), flag = 1;
    while (end >= 0 && end <= n - 1 && flag) {
      if (lcp(k, start, end, n) > min_delta) end += d;
      else { end -= d; flag = 0; }
    }
    if (end < 0) end = 0;
    if (end > n - 1) end = n - 1;
    int split = start + d, i1 = start, i2 = start + d, min2 = lcp(k, i1, i2, n);
    while ((d > 0 && i2 <= end) || (d < 0 && i2 >= end)) {
      int cur = lcp(k, i1, i2, n);
      if (cur < min2) { split = i2; min2 = cur; }
      i1 += d; i2 += d;
    }
    int left, right;
    if (d > 0) {
      left = (split - start > 1) ? split - 1 : ~start;
      right = (end - split > 0) ? split : ~end;
    } else {
      right = (start - split > 1) ? split + 1 : ~start;
      left = (split - end > 0) ? split : ~end;
    }
    b->left[idx] = left; b->right[idx] = right;
    if (left >= 0) b->parent_internal[left] = idx; else b->parent_leaf[~left] = idx;
    if (right >= 0) b->parent_internal[right] = idx; else b->parent_leaf[~right] = idx;
  }
  refit(c, b);
}

int orc_build(OrcCtx *c) {
  if (!c->tex[0]) { float one = 1.f; orc_set_textures(c, &one, &one, &one, 1); }
  preprocess(c);
  int n = c->nt;
  int *all = malloc(sizeof(int) * n), *sub = malloc(sizeof(int) * n);
  for (int i = 0; i < n; i++) all[i] = i;
  build_bvh(c, &c->bvh[GPT_BVH_MAIN], all, n);
  /* the subsets are collected from the SORTED main list (src/main.cu:952,1083) */
  const int masks[2] = {GPT_TRI_TRANSPARENT, GPT_TRI_LIGHT};
  const int which[2] = {GPT_BVH_TRANSPARENT, GPT_BVH_LIGHT};
  for (int s = 0; s < 2; s++) {
    int cnt = 0;
    for (int j = 0; j < n; j++) {
      int p = c->bvh[GPT_BVH_MAIN].prim[j];
      if (c->flags[p] & masks[s]) sub[cnt++] = p;
    }
    build_bvh(c, &c->bvh[which[s]], sub, cnt);
  }
  free(all); free(sub);
  return GPT_OK;
}

int orc_export_triangles(OrcCtx *c, float *tolerance, float *aabb, float *normal, float *area,
                         int32_t *flags, float *energy, float *tangents) {
  size_t n = (size_t)c->nt;
  if (tolerance) memcpy(tolerance, c->tolerance, 4 * n);
  if (aabb) memcpy(aabb, c->aabb, 24 * n);
  if (normal) memcpy(normal, c->normal, 12 * n);
  if (area) memcpy(area, c->area, 4 * n);
  if (flags) memcpy(flags, c->flags, 4 * n);
  if (energy) memcpy(energy, c->energy, 12 * n);
  if (tangents) memcpy(tangents, c->tangents, 36 * n);
  return GPT_OK;
}

int orc_bvh_num_leaves(OrcCtx *c, int which, int32_t *n) { *n = c->bvh[which].n; return GPT_OK; }

int orc_export_bvh(OrcCtx *c, int which, uint32_t *morton, int32_t *sorted_to_orig, int32_t *left,
                   int32_t *right, int32_t *parent_internal, int32_t *parent_leaf,
                   float *node_aabb, float *leaf_aabb, float *world_aabb) {
  OrcBvh *b = &c->bvh[which];
  size_t n = (size_t)b->n;
  if (n < 2) return GPT_E_INVALID;
  if (morton) memcpy(morton, b->morton, 4 * n);
  if (sorted_to_orig) memcpy(sorted_to_orig, b->prim, 4 * n);
  if (left) memcpy(left, b->left, 4 * (n - 1));
  if (right) memcpy(right, b->right, 4 * (n - 1));
  if (parent_internal) memcpy(parent_internal, b->parent_internal, 4 * (n - 1));
  if (parent_leaf) memcpy(parent_leaf, b->parent_leaf, 4 * n);
  if (node_aabb) memcpy(node_aabb, b->node_aabb, 24 * (n - 1));
  if (leaf_aabb)
    for (size_t j = 0; j < n; j++) memcpy(leaf_aabb + 6 * j, c->aabb + 6 * b->prim[j], 24);
  if (world_aabb) memcpy(world_aabb, b->world, 24);
  return GPT_OK;
}

/* ---- BoundingBox::is_intersection (src/model/grid/bounding_box.h:61-69,217-276) ---------- */
static void slab(float lo, float hi, float o, float d, float *tmin, float *tmax) {
  float inv = 1.0f / d; /* 1.0 / dir as double, rounded to float == IEEE fp32 division */
  float t1 = (lo - o) * inv, t2 = (hi - o) * inv;
  if (t1 < t2) { *tmin = t1; *tmax = t2; } else { *tmin = t2; *tmax = t1; }
}
static int box_hit(const float *bb, v3 o, v3 d) {
  float xn, xx, yn, yx, zn, zx;
  slab(bb[0], bb[1], o.x, d.x, &xn, &xx);
  slab(bb[2], bb[3], o.y, d.y, &yn, &yx);
  slab(bb[4], bb[5], o.z, d.z, &zn, &zx);
  if (xn <= yx && yn <= xx) {
    float xyn = fmaxf(xn, yn), xyx = fminf(xx, yx);
    if (xyn <= zx && zn <= xyx) return 1;
  }
  return 0;
}

/* ---- Triangle::hit (src/model/geometry/triangle.h:343-408) -------------------------------- */
static float pick(v3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }
static int tri_hit(const OrcCtx *c, int prim, v3 o, v3 d, float t_max, float *t_out, float *bary) {
  v3 P1 = V(c->x[c->p1[prim]], c->y[c->p1[prim]], c->z[c->p1[prim]]);
  v3 P2 = V(c->x[c->p2[prim]], c->y[c->p2[prim]], c->z[c->p2[prim]]);
  v3 P3 = V(c->x[c->p3[prim]], c->y[c->p3[prim]], c->z[c->p3[prim]]);
  v3 a = vsub(P1, o), b = vsub(P2, o), e = vsub(P3, o);
  float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  int kz = (ax >= ay && ax >= az) ? 0 : ((ay >= ax && ay >= az) ? 1 : 2); /* vec3.h:200-208 */
  int kx = kz + 1; if (kx == 3) kx = 0;
  int ky = kx + 1; if (ky == 3) ky = 0;
  float dx = pick(d, kx), dy = pick(d, ky), dz = pick(d, kz);
  float p1x = pick(a, kx), p1y = pick(a, ky), p1z = pick(a, kz);
  float p2x = pick(b, kx), p2y = pick(b, ky), p2z = pick(b, kz);
  float p3x = pick(e, kx), p3y = pick(e, ky), p3z = pick(e, kz);
  float sx = -dx / dz, sy = -dy / dz, sz = 1.0f / dz;
  p1x = p1x + sx * p1z; p1y = p1y + sy * p1z;
  p2x = p2x + sx * p2z; p2y = p2y + sy * p2z;
  p3x = p3x + sx * p3z; p3y = p3y + sy * p3z;
  float e1 = p2x * p3y - p2y * p3x, e2 = p3x * p1y - p3y * p1x, e3 = p1x * p2y - p1y * p2x;
  if (e1 == 0 || e2 == 0 || e3 == 0) {
    e1 = (float)((double)p2x * (double)p3y - (double)p2y * (double)p3x);
    e2 = (float)((double)p3x * (double)p1y - (double)p3y * (double)p1x);
    e3 = (float)((double)p1x * (double)p2y - (double)p1y * (double)p2x);
  }
  if ((e1 < 0 || e2 < 0 || e3 < 0) && (e1 > 0 || e2 > 0 || e3 > 0)) return 0;
  float det = e1 + e2 + e3;
  if (det == 0) return 0;
  p1z = p1z * sz; p2z = p2z * sz; p3z = p3z * sz;
  float t_scaled = e1 * p1z + e2 * p2z + e3 * p3z;
  float inv_det = 1.0f / det;
  float t = t_scaled * inv_det;
  if (t > t_max || t < c->tolerance[prim]) return 0;
  *t_out = t;
  bary[0] = e1 * inv_det; bary[1] = e2 * inv_det; bary[2] = e3 * inv_det;
  return 1;
}

/* ---- traverse_bvh (src/model/bvh/bvh_traversal.h:12-84): reference order, no culling,
 *      equal t replaces (t > t_max rejects) --------------------------------------------------- */
static int traverse(const OrcCtx *c, const OrcBvh *b, v3 o, v3 d, int *leaf, float *t,
                    float *bary) {
  int stack[400], sp = 0, found = 0, node = 0;
  float best = INFINITY;
  stack[sp++] = -1;
  do {
    int cl = b->left[node], cr = b->right[node];
    int il = box_hit(child_box(c, b, cl), o, d), ir = box_hit(child_box(c, b, cr), o, d);
    const int ch[2] = {cl, cr}, ih[2] = {il, ir};
    for (int s = 0; s < 2; s++) {
      if (ih[s] && ch[s] < 0) {
        float tt, bb[3];
        if (tri_hit(c, b->prim[~ch[s]], o, d, best, &tt, bb)) {
          best = tt; *leaf = ~ch[s]; *t = tt; memcpy(bary, bb, sizeof(bb)); found = 1;
        }
      }
    }
    int tl = il && cl >= 0, tr = ir && cr >= 0;
    if (!tl && !tr) {
      node = stack[--sp];
    } else {
      node = tl ? cl : cr;
      if (tl && tr) stack[sp++] = cr;
    }
  } while (sp > 0 && sp < 400 && node >= 0);
  return found;
}

/* rays: n x 8 {o, tmax, d, 0}; the direction is normalised as Ray::Ray does (ray.h:21-24) */
int orc_trace(OrcCtx *c, int which, const float *rays, int64_t n, int32_t *leaf_id,
              int32_t *tri_id, float *t_out, float *bary) {
  const OrcBvh *b = &c->bvh[which];
  if (b->n < 2) return GPT_E_INVALID;
  for (int64_t r = 0; r < n; r++) {
    const float *q = rays + 8 * r;
    v3 o = V(q[0], q[1], q[2]), d = vunit(V(q[4], q[5], q[6]));
    int leaf = -1;
    float t = INFINITY, bb[3] = {0, 0, 0};
    int hit = traverse(c, b, o, d, &leaf, &t, bb);
    if (hit && t > q[3]) hit = 0;
    if (!hit) { leaf = -1; t = INFINITY; bb[0] = bb[1] = bb[2] = 0; }
    if (leaf_id) leaf_id[r] = leaf;
    if (tri_id) tri_id[r] = hit ? b->prim[leaf] : -1;
    if (t_out) t_out[r] = t;
    if (bary) { bary[3 * r] = bb[0]; bary[3 * r + 1] = bb[1]; bary[3 * r + 2] = bb[2]; }
  }
  return GPT_OK;
}

/* ---- visit counts of the ORDERED, CULLING walk the CUDA path's binary-tree kernels do
 *      (gpu_pathtracer_b200/csrc/trace_core.cuh trace_ray, trace_persistent.cuh tp_step): the
 *      reference's predicate per child (above), the extra dominant-axis cull against t_best
 *      with the 1e-5 relative margin, leaves tested where they are found (left, then right),
 *      nearer internal child first.  SURVEY.md 8(d) takes the algorithmic bytes of the trace
 *      stage from these counts (gpt_stats.internal_visits / leaf_tests); this restatement pins
 *      them (tests/test_gpu_counters.py).  It is the checker of OUR culling rule -- the
 *      reference itself never culls (traverse() above). -------------------------------------- */
static void slab_nf(float lo, float hi, float o, float inv, float *tn, float *tf) {
  float a1 = (lo - o) * inv, a2 = (hi - o) * inv;
  if (a1 < a2) { *tn = a1; *tf = a2; } else { *tn = a2; *tf = a1; }
}
int orc_trace_count(OrcCtx *c, int which, const float *rays, int64_t n, uint64_t *v_int,
                    uint64_t *v_leaf, int32_t *leaf_id) {
  const OrcBvh *b = &c->bvh[which];
  if (b->n < 2) return GPT_E_INVALID;
  uint64_t vi = 0, vl = 0;
  for (int64_t r = 0; r < n; r++) {
    const float *q = rays + 8 * r;
    v3 o = V(q[0], q[1], q[2]), d = vunit(V(q[4], q[5], q[6]));
    float ix = 1.0f / d.x, iy = 1.0f / d.y, iz = 1.0f / d.z;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz = (ax >= ay && ax >= az) ? 0 : ((ay >= ax && ay >= az) ? 1 : 2);
    float best = q[3];
    int hit_leaf = -1, stack[128], sp = 0, node = 0;
    for (;;) {
      int cr[2] = {b->left[node], b->right[node]}, go[2];
      float tnear[2], tcull[2];
      vi++;
      for (int k = 0; k < 2; k++) {
        const float *bb = child_box(c, b, cr[k]);
        float txn, txx, tyn, tyx, tzn, tzx;
        slab_nf(bb[0], bb[1], o.x, ix, &txn, &txx);
        slab_nf(bb[2], bb[3], o.y, iy, &tyn, &tyx);
        slab_nf(bb[4], bb[5], o.z, iz, &tzn, &tzx);
        int ok = (txn <= tyx) && (tyn <= txx);
        float txyn = fmaxf(txn, tyn), txyx = fminf(txx, tyx);
        ok = ok && (txyn <= tzx) && (tzn <= txyx);
        float tn = kz == 0 ? txn : (kz == 1 ? tyn : tzn), tf = kz == 0 ? txx : (kz == 1 ? tyx : tzx);
        float margin = 1e-5f * fmaxf(fabsf(tn), fabsf(tf));
        tnear[k] = tn;
        tcull[k] = tn - margin;
        go[k] = ok && !(tf + margin < 0.f);
      }
      for (int k = 0; k < 2; k++) {
        if (go[k] && cr[k] < 0) {
          go[k] = 0;
          if (tcull[k] > best) continue;
          float tt, bary[3];
          vl++;
          /* tri_hit rejects t > t_max; equal t replaces only for a later visit rank, which does
           * not change the counts */
          if (tri_hit(c, b->prim[~cr[k]], o, d, best, &tt, bary)) { best = tt; hit_leaf = ~cr[k]; }
        }
      }
      int gl = go[0] && !(tcull[0] > best), gr = go[1] && !(tcull[1] > best);
      if (gl && gr) {
        int lfirst = tnear[0] <= tnear[1];
        stack[sp++] = lfirst ? cr[1] : cr[0];
        node = lfirst ? cr[0] : cr[1];
      } else if (gl || gr) {
        node = gl ? cr[0] : cr[1];
      } else {
        if (sp == 0) break;
        node = stack[--sp];
      }
    }
    if (leaf_id) leaf_id[r] = hit_leaf;
  }
  *v_int = vi;
  *v_leaf = vl;
  return GPT_OK;
}
