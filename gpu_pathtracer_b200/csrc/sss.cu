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
  BvhDev bvh;
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
  bool first_pass_done = false;
};

__global__ void __launch_bounds__(256)
k_obj_flags(const float4 *__restrict__ shade, int n, int *flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 s6 = __ldg(&shade[7 * (size_t)i + 6]);
  if (__float_as_int(s6.z) & GPT_TRI_SSS) flag[__float_as_int(s6.w)] = 1;
}

// Object::compute_accummulated_triangle_area (object.h:104-112): serial fp32 prefix per object
__global__ void k_area_cdf(const float4 *__restrict__ shade, const int *__restrict__ off,
                           const int *__restrict__ num, int nobj, float *cdf) {
  int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= nobj) return;
  float acc = 0.f;
  for (int i = off[o]; i < off[o] + num[o]; i++) {
    float a = __ldg(&shade[7 * (size_t)i + 3].w);
    acc += (a == a ? a : 0.f);
    cdf[i] = acc;
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

extern "C" void gpt_sss_free(gpt_ctx *ctx) {
  if (!ctx || !ctx->sss_state) return;
  SssCtx *sx = sss_of(ctx);
  void *p[] = {sx->cdf, sx->rng, sx->pts_created, sx->pts, sx->leaf, sx->aabb, sx->keys,
               sx->obj_flag, sx->trees};
  for (void *q : p) cudaFree(q);
  for (SssObj &o : sx->objs) gpt_lbvh_free(o.bvh);
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
  if (first_pass_done) *first_pass_done = sx->first_pass_done;
  return true;
}

void gpt_sss_mark_first_pass_done(gpt_ctx *ctx) {
  if (sss_of(ctx)) sss_of(ctx)->first_pass_done = true;
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
  k_area_cdf<<<(nobj + 63) / 64, 64, 0, st>>>(sc.shade, d_tab, d_tab + nobj, nobj, sx->cdf);
  GPT_LAUNCH_CHECK(ctx);
  GPT_TRY(gpt_alloc(ctx, &sx->rng, 6 * (size_t)np));
  k_sss_rng_init<<<(np + 127) / 128, 128, 0, st>>>(sx->rng, np);
  GPT_LAUNCH_CHECK(ctx);
  GPT_TRY(gpt_alloc(ctx, &sx->pts_created, 3 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->pts, 3 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->leaf, 2 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->aabb, 6 * (size_t)np));
  GPT_TRY(gpt_alloc(ctx, &sx->keys, (size_t)np));
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
    trees[so.obj] = SssTree{so.bvh.nodes, sx->leaf + 2 * (size_t)so.pt_off, pts, 0};
  }
  GPT_TRY(gpt_upload(ctx, &sx->trees, trees.data(), trees.size()));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  cudaFree(d_tab);
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
                                const float *d_col, int col_stride, int n, float *out_rgb_host);  // output.cu
extern "C" int gpt_sss_debug_image(gpt_ctx *ctx, float *out_rgb) {
  if (!ctx || !out_rgb) return GPT_E_INVALID;
  SssCtx *sx = sss_of(ctx);
  const float *leaf = sx ? reinterpret_cast<const float *>(sx->leaf) : nullptr;
  return gpt_splat_device(ctx, leaf, 8, leaf ? leaf + 4 : nullptr, 8, sx ? sx->npoints : 0, out_rgb);
}
