// sss.cu -- sub-surface-scattering pre-pass set-up.
//
// Replaces src/main.cu:555-824: create_objects' area CDFs, compute_num_sss_objects,
// create_sss_pts (object_operations.h:120-141), compute_object_boundaries_batch,
// compute_pts_morton_code_batch, the per-object stable sort and the per-object point LBVHs
// (bvh_building_pts.h).  The radiance at the points (do_sss_first_pass, pathtracing.h:202-249)
// is computed by the path-tracing wavefront in render.cu; the lookup (compute_color_sss +
// traverse_bvh_pts) lives in the shade kernel.
//
// Reference quirks, REPLICATED unless stated:
//   * Object::compute_boundaries never updates its running min/max (object.h:118-131), so an
//     object's "bounds" are the box of its LAST point; Morton codes are normalised against
//     that 2e-5 wide box and clamp to 0/1023 on every axis -- the trees are built from these
//     codes (ties broken by index, as in the reference).
//   * create_sss_pts draws the primitive from a COPY of state[idx] and the surface point from
//     the state array itself; the reference passes the array base, so every thread races on
//     state[0] (object_operations.h:135-136).  DEVIATES: each point uses state[idx] in place
//     -- the reference's kernel launched per point (oracle/ref_replay.cpp does exactly that).
//     State idx is reused by every SSS object in turn, as in the reference.
#include <vector>

#include "lighttree.cuh"

namespace {

struct SssObj {
  int obj = 0, tri_off = 0, tri_num = 0, pt_off = 0;
  BvhDev bvh;      // the reference's tree (degenerate keys, see k_sss_morton)
  BvhDev search;   // same points, true Morton keys: finds the candidates of a lookup fast
  float4 *gpos = nullptr;  // candidate grid (SssTree::gpos / gstart)
  int *gstart = nullptr;
};

struct SssCtx {
  int pts_per_obj = 0, npoints = 0;
  std::vector<SssObj> objs;
  float *cdf = nullptr;        // N: per-object accumulated triangle area (object-local prefix)
  uint32_t *rng = nullptr;     // 6 planes x npoints
  float4 *pts_created = nullptr, *pts = nullptr, *leaf = nullptr;
  float *aabb = nullptr;       // 6 x npoints, creation order
  uint32_t *keys = nullptr;
  int *obj_flag = nullptr;
  SssTree *trees = nullptr;    // device, one per scene object
  float4 *spos = nullptr;      // npoints: (location, leaf-order index) per search leaf
  int *rank = nullptr;         // pts scratch: creation index -> leaf-order index
  int *leafpos = nullptr;      // pts scratch: leaf-order index -> position in the reference walk
  int *visit = nullptr;        // npoints: position in the walk -> leaf-order index, per object
  int *visit_global = nullptr; // the same with the object's point offset added (index into leaf)
  float4 *vleaf = nullptr;     // 2 x npoints: leaf records in visit order (after the first pass)
  uint32_t *vkeys = nullptr, *vkeys_alt = nullptr, *vhist = nullptr;  // pts scratch (rank sort)
  int *vvals = nullptr, *vvals_alt = nullptr;
  uint32_t *bm_pool = nullptr;
  int *bm_busy = nullptr;
  int bm_n = 0, bm_words = 0;
  bool first_pass_done = false;
};

__global__ void __launch_bounds__(256)
k_obj_flags(const float4 *__restrict__ shade, int n, int *flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 s6 = __ldg(&shade[7 * (size_t)i + 6]);
  if (__float_as_int(s6.z) & GPT_TRI_SSS) flag[__float_as_int(s6.w)] = 1;
}

// Object::compute_accummulated_triangle_area (object.h:104-112): serial fp32 prefix per object.
// Only the SSS objects' tables are ever read (create_sss_pts), so only those are computed: one
// block per SSS object; the block stages a tile of areas, thread 0 runs the dependent adds over
// shared memory, the block writes the tile back.  (One thread walking a 10 M-triangle terrain
// through global memory was 3.9 s of C5's build.)
constexpr int CDF_TILE = 2048;
__global__ void __launch_bounds__(256)
k_area_cdf(const float4 *__restrict__ shade, const int *__restrict__ off,
           const int *__restrict__ num, float *cdf) {
  __shared__ float tile[CDF_TILE];
  __shared__ float carry;
  const int first = off[blockIdx.x], count = num[blockIdx.x];
  if (threadIdx.x == 0) carry = 0.f;
  for (int base = 0; base < count; base += CDF_TILE) {
    const int m = min(CDF_TILE, count - base);
    for (int k = threadIdx.x; k < m; k += blockDim.x) {
      float a = __ldg(&shade[7 * (size_t)(first + base + k) + 3].w);
      tile[k] = (a == a ? a : 0.f);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      float acc = carry;
      for (int k = 0; k < m; k++) {
        acc += tile[k];
        tile[k] = acc;
      }
      carry = acc;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < m; k += blockDim.x) cdf[first + base + k] = tile[k];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(128) k_sss_rng_init(uint32_t *rng, int n) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  Rng r;
  curand_init(1984, (unsigned long long)s, 0, &r.s);
  rng_store(r, rng, s, n);
}

// create_sss_pts for every SSS object in turn (the host loop at src/main.cu:690-698), one
// thread per point slot idx; Object::pick_primitive_idx_for_sampling by bisection.
__global__ void __launch_bounds__(128)
k_sss_points(SceneDev sc, TexAtlas atlas, const float *__restrict__ cdf,
             const int *__restrict__ tri_off, const int *__restrict__ tri_num,
             const int *__restrict__ pt_off, int nsss, int pts, int npoints, uint32_t *rng,
             float4 *out, float *aabb) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= pts) return;
  Rng G;
  rng_load(G, rng, idx, npoints);
  for (int k = 0; k < nsss; k++) {
    Rng L = G;
    int off = tri_off[k], num = tri_num[k];
    float x = rng_uniform(L) * __ldg(cdf + off + num - 1);
    int lo = 0, hi = num - 1;
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (x > __ldg(cdf + off + mid)) lo = mid + 1; else hi = mid;
    }
    int prim = off + lo;
    // Triangle::get_random_point_on_surface (triangle.h:199-224)
    int a = __ldg(sc.p1 + prim), b = __ldg(sc.p2 + prim), c = __ldg(sc.p3 + prim);
    V3 p1 = V3{__ldg(sc.x + a), __ldg(sc.y + a), __ldg(sc.z + a)};
    V3 p2 = V3{__ldg(sc.x + b), __ldg(sc.y + b), __ldg(sc.z + b)};
    V3 p3 = V3{__ldg(sc.x + c), __ldg(sc.y + c), __ldg(sc.z + c)};
    TriShade ts = load_tri_shade(sc.shade, prim);
    float r1 = rng_uniform(G), r2 = rng_uniform(G);
    float u = 1.f - powf(r1, .5f);
    float v = r2 * powf(r1, .5f);
    float w = 1.f - u - v;
    V3 loc = (u * p1 + v * p2) + w * p3;
    V3 nrm = unit((u * ts.n1 + v * ts.n2) + w * ts.n3);
    float tu = (u * ts.u1 + v * ts.u2) + w * ts.u3;
    float tv = (u * ts.v1 + v * ts.v2) + w * ts.v3;
    (void)rng_uniform(G);  // get_random_unit_vector_hemisphere: the direction is not used
    (void)rng_uniform(G);
    V3 filter = mat_diffuse(atlas, sc.mats[ts.mat], tu, tv);
    size_t q = (size_t)pt_off[k] + idx;
    out[3 * q] = make_float4(loc.x, loc.y, loc.z, filter.x);
    out[3 * q + 1] = make_float4(filter.y, filter.z, nrm.x, nrm.y);
    out[3 * q + 2] = make_float4(nrm.z, 0.f, 0.f, 0.f);
    float *bx = aabb + 6 * q;  // Point::_create_bounding_box (point.h:74-80)
    bx[0] = loc.x - GPT_SMALL; bx[1] = loc.x + GPT_SMALL;
    bx[2] = loc.y - GPT_SMALL; bx[3] = loc.y + GPT_SMALL;
    bx[4] = loc.z - GPT_SMALL; bx[5] = loc.z + GPT_SMALL;
  }
  rng_store(G, rng, idx, npoints);
}

__device__ __forceinline__ uint32_t sss_expand_bits10(uint32_t v) {  // bvh_util.h:4-10
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ uint32_t sss_quant10(float x) {  // bvh_util.h:13-18
  return (uint32_t)fminf(fmaxf(x * 1024.0f, 0.0f), 1023.0f);
}

// compute_pts_morton_code_batch (object_operations.h:48-65) against the object's "bounds" =
// the box of its LAST created point (Object::compute_boundaries, see the header comment)
__global__ void __launch_bounds__(256)
k_sss_morton(const float *__restrict__ aabb, const int *__restrict__ pt_off, int nsss, int pts,
             uint32_t *keys) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nsss * pts) return;
  int k = i / pts;
  const float *ob = aabb + 6 * ((size_t)pt_off[k] + pts - 1);
  const float *b = aabb + 6 * (size_t)i;
  float cx = 0.5f * (b[0] + b[1]), cy = 0.5f * (b[2] + b[3]), cz = 0.5f * (b[4] + b[5]);
  float nx = (cx - ob[0]) / (ob[1] - ob[0]);
  float ny = (cy - ob[2]) / (ob[3] - ob[2]);
  float nz = (cz - ob[4]) / (ob[5] - ob[4]);
  keys[i] = sss_expand_bits10(sss_quant10(nx)) * 4u + sss_expand_bits10(sss_quant10(ny)) * 2u +
            sss_expand_bits10(sss_quant10(nz));
}

// creation order -> tree (sorted) order
__global__ void __launch_bounds__(256)
k_sss_pack(const float4 *__restrict__ created, const int *__restrict__ sorted_prim, int n,
           float4 *pts, float4 *leaf) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  int src = __ldg(sorted_prim + j);
  float4 a = __ldg(created + 3 * (size_t)src);
  pts[3 * (size_t)j] = a;
  pts[3 * (size_t)j + 1] = __ldg(created + 3 * (size_t)src + 1);
  pts[3 * (size_t)j + 2] = __ldg(created + 3 * (size_t)src + 2);
  leaf[2 * (size_t)j] = make_float4(a.x, a.y, a.z, 0.f);
  leaf[2 * (size_t)j + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
}

SssCtx *sss_of(gpt_ctx *ctx) { return static_cast<SssCtx *>(ctx->sss_state); }

}  // namespace

// The cell a coordinate falls in; host and device (render.cu sss_lookup) must agree bit for bit,
// so it is ONE expression of non-contracted fp32 operations: floorf((x - origin) * inv), clamped.
static inline int grid_cell(float x, float o, float inv, int n) {
  float f = floorf((x - o) * inv);
  int c = f < 0.f ? 0 : (f >= (float)n ? n - 1 : (int)f);
  return c;
}

// Candidate grid of one SSS object (SssTree::gpos / gstart): built on the host from the search
// records -- 10^5 points, once per build -- by a counting sort over the cells.
static int build_candidate_grid(gpt_ctx *ctx, SssObj &so, const float4 *d_spos, int pts, SssTree &tr) {
  static int use_grid = -1;
  if (use_grid < 0) {
    const char *e = getenv("GPT_SSS_GRID");
    use_grid = e ? atoi(e) : 1;
  }
  if (!use_grid || pts < 2 || !(ctx->max_path_length > 0.f)) return GPT_OK;
  std::vector<float4> h((size_t)pts);
  GPT_CUDA(ctx, cudaMemcpyAsync(h.data(), d_spos, sizeof(float4) * (size_t)pts, cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  float lo[3] = {h[0].x, h[0].y, h[0].z}, hi[3] = {h[0].x, h[0].y, h[0].z};
  for (const float4 &q : h) {
    const float v[3] = {q.x, q.y, q.z};
    for (int a = 0; a < 3; a++) {
      if (!(v[a] == v[a]) || fabsf(v[a]) > 1e30f) return GPT_OK;  // not a finite cloud: keep the tree
      lo[a] = fminf(lo[a], v[a]);
      hi[a] = fmaxf(hi[a], v[a]);
    }
  }
  float cell = ctx->max_path_length;
  for (int a = 0; a < 3; a++) cell = fmaxf(cell, (hi[a] - lo[a]) / 200.f);  // <= ~200 cells per axis
  if (!(cell > 0.f)) return GPT_OK;
  const float inv = 1.0f / cell;
  int n[3];
  for (int a = 0; a < 3; a++) {
    n[a] = (int)floorf((hi[a] - lo[a]) * inv) + 1;
    if (n[a] < 1) n[a] = 1;
    if (n[a] > 256) n[a] = 256;
  }
  const size_t ncell = (size_t)n[0] * n[1] * n[2];
  std::vector<int> start(ncell + 1, 0), cid((size_t)pts);
  for (int j = 0; j < pts; j++) {
    int cx = grid_cell(h[j].x, lo[0], inv, n[0]), cy = grid_cell(h[j].y, lo[1], inv, n[1]),
        cz = grid_cell(h[j].z, lo[2], inv, n[2]);
    cid[j] = cx + n[0] * (cy + n[1] * cz);
    start[(size_t)cid[j] + 1]++;
  }
  for (size_t c = 0; c < ncell; c++) start[c + 1] += start[c];
  std::vector<int> fill(start.begin(), start.end() - 1);
  std::vector<float4> sorted((size_t)pts);
  for (int j = 0; j < pts; j++) sorted[(size_t)fill[cid[j]]++] = h[j];
  GPT_TRY(gpt_upload(ctx, &so.gpos, sorted.data(), sorted.size()));
  GPT_TRY(gpt_upload(ctx, &so.gstart, start.data(), start.size()));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the uploads read the vectors above
  tr.gpos = so.gpos;
  tr.gstart = so.gstart;
  tr.gox = lo[0]; tr.goy = lo[1]; tr.goz = lo[2];
  tr.ginv = inv;
  tr.gnx = n[0]; tr.gny = n[1]; tr.gnz = n[2];
  return GPT_OK;
}

extern "C" void gpt_sss_free(gpt_ctx *ctx) {
  if (!ctx || !ctx->sss_state) return;
  SssCtx *sx = sss_of(ctx);
  void *p[] = {sx->cdf, sx->rng, sx->pts_created, sx->pts, sx->leaf, sx->aabb, sx->keys,
               sx->obj_flag, sx->trees, sx->spos, sx->rank, sx->bm_pool, sx->bm_busy,
               sx->leafpos, sx->visit, sx->visit_global, sx->vleaf, sx->vkeys, sx->vkeys_alt, sx->vhist, sx->vvals, sx->vvals_alt};
  for (void *q : p) gpt_free(ctx, q);
  for (SssObj &o : sx->objs) {
    gpt_lbvh_free(ctx, o.bvh);
    gpt_lbvh_free(ctx, o.search);
    gpt_free(ctx, o.gpos);
    gpt_free(ctx, o.gstart);
  }
  delete sx;
  ctx->sss_state = nullptr;
}

bool gpt_sss_view(gpt_ctx *ctx, SssView *out, bool *first_pass_done) {
  SssCtx *sx = sss_of(ctx);
  if (!sx || sx->npoints == 0) return false;
  out->trees = sx->trees;
  out->pts = sx->pts;
  out->leaf = sx->leaf;
  out->rng = sx->rng;
  out->npoints = sx->npoints;
  out->bm_pool = sx->bm_pool;
  out->bm_busy = sx->bm_busy;
  out->bm_n = sx->bm_n;
  out->bm_words = sx->bm_words;
  if (first_pass_done) *first_pass_done = sx->first_pass_done;
  return true;
}

__global__ void __launch_bounds__(256)
k_sss_add_offset(const int *__restrict__ in, int n, int off, int *out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = in[k] + off;
}

// vleaf[2 k .. 2 k + 1] = leaf[2 visit[k] .. ]: the records in the order the lookups add them up
__global__ void __launch_bounds__(256)
k_sss_vleaf(const float4 *__restrict__ leaf, const int *__restrict__ visit, int n, float4 *vleaf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  int j = visit[k];
  vleaf[2 * (size_t)k] = leaf[2 * (size_t)j];
  vleaf[2 * (size_t)k + 1] = leaf[2 * (size_t)j + 1];
}

void gpt_sss_mark_first_pass_done(gpt_ctx *ctx) {
  SssCtx *sx = sss_of(ctx);
  if (!sx) return;
  sx->first_pass_done = true;
  if (sx->vleaf && sx->npoints > 0)  // the colours exist now: the visit-order copy of the records
    k_sss_vleaf<<<(sx->npoints + 255) / 256, 256, 0, ctx->stream>>>(sx->leaf, sx->visit_global,
                                                                  sx->npoints, sx->vleaf);
}

// The reference walk (bvh_traversal_pts.h) tests a node's leaf children when it visits the node,
// left before right, and visits nodes in left-first preorder: leaf_visit_rank (trace_core.cuh)
// is that order.  Sorting the leaves by it gives every point its position in the walk.
__global__ void __launch_bounds__(256)
k_sss_visit_keys(const float4 *__restrict__ nodes, const int *__restrict__ parent_leaf, int n,
                 uint32_t *keys, int *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = (uint32_t)leaf_visit_rank(nodes, parent_leaf, i);
  vals[i] = i;
}
__global__ void __launch_bounds__(256)
k_sss_invert(const int *__restrict__ order, int n, int *pos) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) pos[order[k]] = k;
}
// leaf-order index of every point (inverse of the reference tree's sorted_prim) ...
__global__ void __launch_bounds__(256)
k_sss_rank(const int *__restrict__ sorted_prim, int n, int *rank) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rank[sorted_prim[i]] = i;
}
// ... and the search tree's leaf records: location + where the point sits in the reference order
__global__ void __launch_bounds__(256)
k_sss_spos(const float4 *__restrict__ created, const int *__restrict__ search_sorted,
           const int *__restrict__ rank, const int *__restrict__ leafpos, int n, float4 *spos) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  int p = search_sorted[j];
  float4 a = created[3 * (size_t)p];
  spos[j] = make_float4(a.x, a.y, a.z, __int_as_float(leafpos[rank[p]]));
}

int gpt_sss_prepare(gpt_ctx *ctx) {
  gpt_sss_free(ctx);
  int pts = ctx->params.sss_pts_per_object;
  int nobj = (int)ctx->obj_num.size();
  const SceneDev &sc = ctx->scene;
  if (pts <= 0 || nobj == 0) return GPT_OK;
  cudaStream_t st = ctx->stream;
  SssCtx *sx = new SssCtx();
  ctx->sss_state = sx;
  sx->pts_per_obj = pts;
  GPT_TRY(gpt_alloc(ctx, &sx->obj_flag, (size_t)nobj));
  GPT_CUDA(ctx, cudaMemsetAsync(sx->obj_flag, 0, sizeof(int) * (size_t)nobj, st));
  k_obj_flags<<<(sc.ntri + 255) / 256, 256, 0, st>>>(sc.shade, sc.ntri, sx->obj_flag);
  GPT_LAUNCH_CHECK(ctx);
  std::vector<int> flag(nobj);
  GPT_CUDA(ctx, cudaMemcpyAsync(flag.data(), sx->obj_flag, sizeof(int) * (size_t)nobj,
                                cudaMemcpyDeviceToHost, st));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  int off = 0;
  for (int o = 0; o < nobj; o++)
    if (flag[o] && ctx->obj_num[o] > 0) {
      SssObj so;
      so.obj = o;
      so.tri_off = ctx->obj_off[o];
      so.tri_num = ctx->obj_num[o];
      so.pt_off = off;
      off += pts;
      sx->objs.push_back(so);
    }
  int nsss = (int)sx->objs.size();
  if (nsss == 0) { gpt_sss_free(ctx); return GPT_OK; }
  int np = nsss * pts;
  sx->npoints = np;
  // device copies of the per-object tables (all objects for the CDF, SSS objects for sampling)
  int *d_tab = nullptr;
  std::vector<int> tab(2 * (size_t)nobj + 3 * (size_t)nsss);
  for (int o = 0; o < nobj; o++) { tab[o] = ctx->obj_off[o]; tab[nobj + o] = ctx->obj_num[o]; }
  for (int k = 0; k < nsss; k++) {
    tab[2 * nobj + k] = sx->objs[k].tri_off;
    tab[2 * nobj + nsss + k] = sx->objs[k].tri_num;
    tab[2 * nobj + 2 * nsss + k] = sx->objs[k].pt_off;
  }
  GPT_TRY(gpt_upload(ctx, &d_tab, tab.data(), tab.size()));
  GPT_TRY(gpt_alloc(ctx, &sx->cdf, (size_t)sc.ntri));
  k_area_cdf<<<nsss, 256, 0, st>>>(sc.shade, d_tab + 2 * nobj, d_tab + 2 * nobj + nsss, sx->cdf);
  GPT_LAUNCH_CHECK(ctx);
  GPT_TRY(gpt_alloc(ctx, &sx->rng, 6 * (size_t)np));
  k_sss_rng_init<<<(np + 127) / 128, 128, 0, st>>>(sx->rng, np);
  GPT_LAUNCH_CHECK(ctx);
  GPT_TRY(gpt_alloc(ctx, &sx->pts_created, 3 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->pts, 3 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->leaf, 2 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->aabb, 6 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->keys, (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->spos, (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->rank, (size_t)pts));
  GPT_TRY(gpt_alloc(ctx, &sx->leafpos, (size_t)pts));
  GPT_TRY(gpt_alloc(ctx, &sx->visit, (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->visit_global, (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->vleaf, 2 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->vkeys, (size_t)pts));
  GPT_TRY(gpt_alloc(ctx, &sx->vkeys_alt, (size_t)pts));
  GPT_TRY(gpt_alloc(ctx, &sx->vvals, (size_t)pts));
  GPT_TRY(gpt_alloc(ctx, &sx->vvals_alt, (size_t)pts));
  GPT_TRY(gpt_alloc(ctx, &sx->vhist, (size_t)gpt_radix_hist_entries(pts)));
  TexAtlas atlas{sc.tex_r, sc.tex_g, sc.tex_b};
  const int *d_soff = d_tab + 2 * nobj, *d_snum = d_soff + nsss, *d_poff = d_snum + nsss;
  k_sss_points<<<(pts + 127) / 128, 128, 0, st>>>(sc, atlas, sx->cdf, d_soff, d_snum, d_poff, nsss,
                                                 pts, np, sx->rng, sx->pts_created, sx->aabb);
  GPT_LAUNCH_CHECK(ctx);
  k_sss_morton<<<(np + 255) / 256, 256, 0, st>>>(sx->aabb, d_poff, nsss, pts, sx->keys);
  GPT_LAUNCH_CHECK(ctx);
  std::vector<SssTree> trees(nobj, SssTree{nullptr, nullptr, 0, 0});
  for (int k = 0; k < nsss; k++) {
    SssObj &so = sx->objs[k];
    if (pts < 2) continue;  // the reference cannot build or traverse a one-point tree
    GPT_TRY(gpt_lbvh_build_boxes(ctx, so.bvh, sx->aabb + 6 * (size_t)so.pt_off, nullptr, pts, false,
                                 sx->keys + so.pt_off));
    k_sss_pack<<<(pts + 255) / 256, 256, 0, st>>>(sx->pts_created + 3 * (size_t)so.pt_off,
                                                 so.bvh.sorted_prim, pts,
                                                 sx->pts + 3 * (size_t)so.pt_off,
                                                 sx->leaf + 2 * (size_t)so.pt_off);
    GPT_LAUNCH_CHECK(ctx);
    // the search structure: same boxes, the builder's own (true) Morton codes
    GPT_TRY(gpt_lbvh_build_boxes(ctx, so.search, sx->aabb + 6 * (size_t)so.pt_off, nullptr, pts, false,
                                 nullptr));
    k_sss_rank<<<(pts + 255) / 256, 256, 0, st>>>(so.bvh.sorted_prim, pts, sx->rank);
    GPT_LAUNCH_CHECK(ctx);
    // positions in the reference walk: sort the leaves by visit rank (< 2 * pts)
    k_sss_visit_keys<<<(pts + 255) / 256, 256, 0, st>>>(so.bvh.nodes, so.bvh.parent_leaf, pts, sx->vkeys,
                                                       sx->vvals);
    GPT_LAUNCH_CHECK(ctx);
    int vbits = 1;
    while ((1ll << vbits) < 2ll * pts) vbits++;
    uint32_t *ko = nullptr;
    int *order = nullptr;
    GPT_TRY(gpt_radix_sort_pairs(ctx, sx->vkeys, sx->vvals, sx->vkeys_alt, sx->vvals_alt, sx->vhist, pts,
                                 vbits, &ko, &order, nullptr));
    GPT_CUDA(ctx, cudaMemcpyAsync(sx->visit + so.pt_off, order, sizeof(int) * (size_t)pts,
                                  cudaMemcpyDeviceToDevice, st));
    k_sss_invert<<<(pts + 255) / 256, 256, 0, st>>>(sx->visit + so.pt_off, pts, sx->leafpos);
    GPT_LAUNCH_CHECK(ctx);
    k_sss_add_offset<<<(pts + 255) / 256, 256, 0, st>>>(sx->visit + so.pt_off, pts, so.pt_off,
                                                       sx->visit_global + so.pt_off);
    GPT_LAUNCH_CHECK(ctx);
    k_sss_spos<<<(pts + 255) / 256, 256, 0, st>>>(sx->pts_created + 3 * (size_t)so.pt_off,
                                                 so.search.sorted_prim, sx->rank, sx->leafpos, pts,
                                                 sx->spos + so.pt_off);
    GPT_LAUNCH_CHECK(ctx);
    SssTree tr{so.bvh.nodes, sx->leaf + 2 * (size_t)so.pt_off, pts, 0,
               so.search.nodes, sx->spos + so.pt_off, sx->visit + so.pt_off,
               nullptr, nullptr, 0.f, 0.f, 0.f, 0.f, 0, 0, 0, sx->vleaf + 2 * (size_t)so.pt_off};
    GPT_TRY(build_candidate_grid(ctx, so, sx->spos + so.pt_off, pts, tr));
    trees[so.obj] = tr;
  }
  // candidate bitmaps: one bit per point; enough of them for every resident shade thread to hold
  // one (148 SMs x 640 threads) at 1.4x for the probing, within 2 GB
  sx->bm_words = (((pts + 31) / 32) + 3) & ~3;  // whole 16-byte groups
  sx->bm_n = (int)std::min<size_t>(131072, ((size_t)2 << 30) / (4 * (size_t)sx->bm_words));
  if (sx->bm_n < 1) sx->bm_n = 1;
  GPT_TRY(gpt_alloc(ctx, &sx->bm_pool, (size_t)sx->bm_n * sx->bm_words));
  GPT_TRY(gpt_alloc(ctx, &sx->bm_busy, (size_t)sx->bm_n));
  GPT_CUDA(ctx, cudaMemsetAsync(sx->bm_pool, 0, sizeof(uint32_t) * (size_t)sx->bm_n * sx->bm_words, st));
  GPT_CUDA(ctx, cudaMemsetAsync(sx->bm_busy, 0, sizeof(int) * (size_t)sx->bm_n, st));
  GPT_TRY(gpt_upload(ctx, &sx->trees, trees.data(), trees.size()));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  gpt_free(ctx, d_tab);
  return GPT_OK;
}

// parity export: points in tree order -- object index + 12 floats (location, filter, normal,
// colour); colour is valid after the first pass (a render, or run_first_pass != 0)
extern "C" int gpt_render_sss_first_pass(gpt_ctx *ctx);  // render.cu
extern "C" int gpt_sss_export_points(gpt_ctx *ctx, int run_first_pass, int32_t *object_idx,
                                     float *rec12, int64_t capacity, int64_t *count) {
  if (!ctx || !count) return GPT_E_INVALID;
  SssCtx *sx = sss_of(ctx);
  *count = sx ? sx->npoints : 0;
  if (!sx || capacity < sx->npoints) return GPT_OK;
  if (run_first_pass) GPT_TRY(gpt_render_sss_first_pass(ctx));
  int np = sx->npoints;
  std::vector<float> pts(12 * (size_t)np), leaf(8 * (size_t)np);
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  GPT_CUDA(ctx, cudaMemcpy(pts.data(), sx->pts, sizeof(float) * pts.size(), cudaMemcpyDeviceToHost));
  GPT_CUDA(ctx, cudaMemcpy(leaf.data(), sx->leaf, sizeof(float) * leaf.size(), cudaMemcpyDeviceToHost));
  for (const SssObj &so : sx->objs)
    for (int j = 0; j < sx->pts_per_obj; j++) {
      size_t q = (size_t)so.pt_off + j;
      if (object_idx) object_idx[q] = so.obj;
      if (rec12) {
        const float *p = pts.data() + 12 * q, *l = leaf.data() + 8 * q;
        float *o = rec12 + 12 * q;
        o[0] = p[0]; o[1] = p[1]; o[2] = p[2];      // location
        o[3] = p[3]; o[4] = p[4]; o[5] = p[5];      // filter
        o[6] = p[6]; o[7] = p[7]; o[8] = p[8];      // normal
        o[9] = l[4]; o[10] = l[5]; o[11] = l[6];    // colour
      }
    }
  return GPT_OK;
}

// "Creating point image" of the SSS points (src/main.cu:1244-1262): colour = point colour
extern "C" int gpt_splat_device(gpt_ctx *ctx, const float *d_loc, int loc_stride,
                                const float *d_col, int col_stride, int n, const int *d_order,
                                float *out_rgb_host);  // output.cu
extern "C" int gpt_sss_debug_image(gpt_ctx *ctx, float *out_rgb) {
  if (!ctx || !out_rgb) return GPT_E_INVALID;
  SssCtx *sx = sss_of(ctx);
  const float *leaf = sx ? reinterpret_cast<const float *>(sx->leaf) : nullptr;
  return gpt_splat_device(ctx, leaf, 8, leaf ? leaf + 4 : nullptr, 8, sx ? sx->npoints : 0, nullptr, out_rgb);
}
