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
