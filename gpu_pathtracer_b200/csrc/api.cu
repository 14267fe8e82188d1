// api.cu -- the C-ABI of include/gpt.h: context lifetime, scene upload, build, parity
// exports and the trace entry points.  Rendering entry points live in render.cu / sppm.cu.
#include <cstdio>
#include <cstring>
#include <vector>

#include "common.cuh"

int gpt_fail(gpt_ctx *ctx, int code, const char *what, const char *file, int line) {
  if (ctx) {
    char buf[512];
    snprintf(buf, sizeof(buf), "%s (%s:%d)", what, file, line);
    ctx->err = buf;
  }
  return code;
}

extern "C" {

const char *gpt_version(void) { return "gpu_pathtracer_b200 0.1 (sm_100a)"; }

int gpt_create(int device, gpt_ctx **out) {
  if (!out) return GPT_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return GPT_E_NODEVICE;
  if (device < 0 || device >= ndev) return GPT_E_INVALID;
  if (cudaSetDevice(device) != cudaSuccess) return GPT_E_CUDA;
  gpt_ctx *ctx = new gpt_ctx();
  ctx->device = device;
  if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess ||
      cudaMalloc((void **)&ctx->d_counters, 8 * sizeof(unsigned long long)) != cudaSuccess) {
    delete ctx;
    return GPT_E_CUDA;
  }
  cudaMemset(ctx->d_counters, 0, 8 * sizeof(unsigned long long));
  {  // gpt_alloc (common.cuh): a private pool that keeps freed blocks instead of returning them
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    if (cudaMemPoolCreate(&ctx->pool, &props) != cudaSuccess) {
      cudaFree(ctx->d_counters);
      delete ctx;
      return GPT_E_CUDA;
    }
    uint64_t keep = UINT64_MAX;
    cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  ctx->stream = ctx->own_stream;
  ctx->params.render_mode = 1;
  ctx->params.pathtracing_sample_size = 1;
  ctx->params.pathtracing_level = 1;
  ctx->params.dof_sample_size = 1;
  *out = ctx;
  return GPT_OK;
}

void gpt_render_free(gpt_ctx *ctx);  // render.cu
void gpt_sppm_free(gpt_ctx *ctx);    // sppm.cu
void gpt_sss_free(gpt_ctx *ctx);     // sss.cu

static void free_scene(gpt_ctx *ctx, SceneDev &s) {
  void *ptrs[] = {s.x, s.y, s.z, s.xn, s.yn, s.zn, s.xt, s.yt, s.p1, s.p2, s.p3, s.n1, s.n2,
                  s.n3, s.t1, s.t2, s.t3, s.mat_idx, s.obj_idx, s.tolerance, s.aabb, s.normal,
                  s.energy, s.shade, s.vtangent};
  for (void *p : ptrs) gpt_free(ctx, p);
  MaterialDev *mats = s.mats;
  float *tr = s.tex_r, *tg = s.tex_g, *tb = s.tex_b, *br = s.bg_r, *bg = s.bg_g, *bb = s.bg_b;
  int nmat = s.nmat, bh = s.bg_h, bw = s.bg_w;
  int64_t tl = s.tex_len;
  s = SceneDev();
  s.mats = mats; s.nmat = nmat;
  s.tex_r = tr; s.tex_g = tg; s.tex_b = tb; s.tex_len = tl;
  s.bg_r = br; s.bg_g = bg; s.bg_b = bb; s.bg_h = bh; s.bg_w = bw;
}

int gpt_comm_destroy(gpt_ctx *ctx);  // comm.cu

void gpt_destroy(gpt_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  gpt_comm_destroy(ctx);
  gpt_render_free(ctx);
  gpt_sppm_free(ctx);
  gpt_sss_free(ctx);
  for (int k = 0; k < 3; k++) gpt_lbvh_free(ctx, ctx->bvh[k]);
  SceneDev &s = ctx->scene;
  free_scene(ctx, s);
  gpt_free(ctx, s.mats); gpt_free(ctx, s.tex_r); gpt_free(ctx, s.tex_g); gpt_free(ctx, s.tex_b);
  gpt_free(ctx, s.bg_r); gpt_free(ctx, s.bg_g); gpt_free(ctx, s.bg_b);
  cudaFree(ctx->d_counters);  // plain cudaMalloc in gpt_create
  cudaStreamSynchronize(ctx->stream);
  cudaStreamSynchronize(ctx->own_stream);
  if (ctx->pool) cudaMemPoolDestroy(ctx->pool);  // everything gpt_alloc ever handed out goes back to the device
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

const char *gpt_last_error(const gpt_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int gpt_set_stream(gpt_ctx *ctx, uint64_t cuda_stream) {
  if (!ctx) return GPT_E_INVALID;
  cudaStream_t next = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  // allocations and frees are ordered on ctx->stream: drain the old one before switching
  if (next != ctx->stream) GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->stream = next;
  return GPT_OK;
}

int gpt_synchronize(gpt_ctx *ctx) {
  if (!ctx) return GPT_E_INVALID;
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  // the asynchronous entry points (gpt_trace_device) cannot report a traversal-stack overflow
  // themselves: it surfaces here
  unsigned long long overflow = 0;
  GPT_CUDA(ctx, cudaMemcpy(&overflow, ctx->d_counters + 3, sizeof(overflow), cudaMemcpyDeviceToHost));
  if (overflow) {
    cudaMemset(ctx->d_counters + 3, 0, sizeof(overflow));
    return gpt_fail(ctx, GPT_E_UNSUPPORTED, "traversal stack overflow (> 64 levels) in an earlier asynchronous call", __FILE__, __LINE__);
  }
  return GPT_OK;
}

int gpt_set_mesh(gpt_ctx *ctx, const GptMesh *m) {
  if (!ctx || !m || m->num_triangles < 0) return GPT_E_INVALID;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  SceneDev &s = ctx->scene;
  free_scene(ctx, s);
  ctx->built = false;
  s.nv = m->num_vertices; s.nvn = m->num_normals; s.nvt = m->num_texcoords;
  s.ntri = m->num_triangles; s.nobj = m->num_objects;
  size_t nt = (size_t)s.ntri;
  GPT_TRY(gpt_upload(ctx, &s.x, m->x, (size_t)s.nv));
  GPT_TRY(gpt_upload(ctx, &s.y, m->y, (size_t)s.nv));
  GPT_TRY(gpt_upload(ctx, &s.z, m->z, (size_t)s.nv));
  GPT_TRY(gpt_upload(ctx, &s.xn, m->x_norm, (size_t)s.nvn));
  GPT_TRY(gpt_upload(ctx, &s.yn, m->y_norm, (size_t)s.nvn));
  GPT_TRY(gpt_upload(ctx, &s.zn, m->z_norm, (size_t)s.nvn));
  GPT_TRY(gpt_upload(ctx, &s.xt, m->x_tex, (size_t)s.nvt));
  GPT_TRY(gpt_upload(ctx, &s.yt, m->y_tex, (size_t)s.nvt));
  GPT_TRY(gpt_upload(ctx, &s.p1, m->point_1_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.p2, m->point_2_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.p3, m->point_3_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.n1, m->norm_1_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.n2, m->norm_2_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.n3, m->norm_3_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.t1, m->tex_1_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.t2, m->tex_2_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.t3, m->tex_3_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.mat_idx, m->material_idx, nt));
  GPT_TRY(gpt_upload(ctx, &s.obj_idx, m->triangle_object_idx, nt));
  ctx->obj_num.clear();
  ctx->obj_off.clear();
  if (m->num_objects > 0 && m->object_num_primitives && m->object_primitive_offset_idx) {
    ctx->obj_num.assign(m->object_num_primitives, m->object_num_primitives + m->num_objects);
    ctx->obj_off.assign(m->object_primitive_offset_idx,
                        m->object_primitive_offset_idx + m->num_objects);
  }
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // host arrays may be freed after return
  ctx->have_mesh = true;
  return GPT_OK;
}

int gpt_set_materials(gpt_ctx *ctx, const GptMaterials *m) {
  if (!ctx || !m || m->num_materials < 1 || m->num_materials > 100)
    return gpt_fail(ctx, GPT_E_INVALID, "1..100 materials (src/main.cu:130)", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  int n = m->num_materials;
  std::vector<MaterialDev> h((size_t)n);
  for (int i = 0; i < n; i++) {
    MaterialDev &d = h[i];
    memset(&d, 0, sizeof(d));
    d.ka[0] = m->ka_x[i]; d.ka[1] = m->ka_y[i]; d.ka[2] = m->ka_z[i];
    d.kd[0] = m->kd_x[i]; d.kd[1] = m->kd_y[i]; d.kd[2] = m->kd_z[i];
    d.ks[0] = m->ks_x[i]; d.ks[1] = m->ks_y[i]; d.ks[2] = m->ks_z[i];
    d.ke[0] = m->ke_x[i]; d.ke[1] = m->ke_y[i]; d.ke[2] = m->ke_z[i];
    d.tf[0] = m->tf_x[i]; d.tf[1] = m->tf_y[i]; d.tf[2] = m->tf_z[i];
    d.path_length = m->path_length[i];
    if (i == 0) ctx->max_path_length = 0.f;
    if (d.path_length > ctx->max_path_length) ctx->max_path_length = d.path_length;
    d.t_r = m->t_r[i];
    d.n_i = m->n_i[i];
    d.bm = m->bm[i];
    // Material::Material (material.h:411-507)
    d.n_s = (m->n_s[i] >= 1e6f && d.t_r > 0.f) ? INFINITY : m->n_s[i];
    d.scattering_coef = m->scattering_coef[i];
    d.absorption_coef = m->absorption_coef[i];
    d.extinction_coef = m->scattering_coef[i] + m->absorption_coef[i];
    d.g = m->g[i];
    d.scattering_prob = d.scattering_coef / d.extinction_coef;
    d.priority = m->material_priority[i];
    d.sss = d.path_length > 0.f;
    const int32_t *hh[5] = {m->height_diffuse, m->height_specular, m->height_emission,
                            m->height_n_s, m->height_bump};
    const int32_t *ww[5] = {m->width_diffuse, m->width_specular, m->width_emission,
                            m->width_n_s, m->width_bump};
    const int32_t *oo[5] = {m->offset_diffuse, m->offset_specular, m->offset_emission,
                            m->offset_n_s, m->offset_bump};
    for (int k = 0; k < 5; k++) {
      d.tex_h[k] = hh[k][i]; d.tex_w[k] = ww[k][i]; d.tex_off[k] = oo[k][i];
    }
  }
  GPT_TRY(gpt_upload(ctx, &ctx->scene.mats, h.data(), (size_t)n));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->scene.nmat = n;
  ctx->have_mats = true;
  ctx->built = false;
  return GPT_OK;
}

int gpt_set_textures(gpt_ctx *ctx, const float *r, const float *g, const float *b,
                     int64_t n) {
  if (!ctx || n < 1 || !r || !g || !b) return GPT_E_INVALID;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  SceneDev &s = ctx->scene;
  GPT_TRY(gpt_upload(ctx, &s.tex_r, r, (size_t)n));
  GPT_TRY(gpt_upload(ctx, &s.tex_g, g, (size_t)n));
  GPT_TRY(gpt_upload(ctx, &s.tex_b, b, (size_t)n));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  s.tex_len = n;
  ctx->built = false;
  return GPT_OK;
}

int gpt_set_background(gpt_ctx *ctx, const float *r, const float *g, const float *b, int32_t h,
                       int32_t w) {
  if (!ctx || h < 1 || w < 1 || !r || !g || !b) return GPT_E_INVALID;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  SceneDev &s = ctx->scene;
  size_t n = (size_t)h * w;
  GPT_TRY(gpt_upload(ctx, &s.bg_r, r, n));
  GPT_TRY(gpt_upload(ctx, &s.bg_g, g, n));
  GPT_TRY(gpt_upload(ctx, &s.bg_b, b, n));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  s.bg_h = h; s.bg_w = w;
  return GPT_OK;
}

int gpt_set_camera(gpt_ctx *ctx, const GptCamera *cam) {
  if (!ctx || !cam || cam->width < 1 || cam->height < 1) return GPT_E_INVALID;
  ctx->cam = *cam;
  ctx->have_cam = true;
  return GPT_OK;
}

int gpt_set_params(gpt_ctx *ctx, const GptParams *p) {
  if (!ctx || !p) return GPT_E_INVALID;
  if (p->rng_mode != 0 && p->rng_mode != 1)
    return gpt_fail(ctx, GPT_E_INVALID, "rng_mode must be 0 (reference streams) or 1 (per-sample streams)", __FILE__, __LINE__);
  ctx->params = *p;
  return GPT_OK;
}

}  // extern "C"

// ---- build / exports / trace -----------------------------------------------------------
static int build_main_timed(gpt_ctx *ctx) {
  // build_ms = the LBVH itself (the reference's structure: Morton codes, sort, emit, refit);
  // build_wide_ms = our 4-wide tree collapsed from it (trace_wide.cuh), timed separately
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  GPT_TRY(gpt_lbvh_build_boxes(ctx, ctx->bvh[GPT_BVH_MAIN], ctx->scene.aabb, nullptr,
                               ctx->scene.ntri, true, nullptr));
  GPT_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  GPT_TRY(gpt_lbvh_build_wide(ctx, ctx->bvh[GPT_BVH_MAIN]));
  cudaEvent_t ev2;
  GPT_CUDA(ctx, cudaEventCreate(&ev2));
  GPT_CUDA(ctx, cudaEventRecord(ev2, ctx->stream));
  GPT_CUDA(ctx, cudaEventSynchronize(ev2));
  float ms = 0.f, ms_w = 0.f;
  GPT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  GPT_CUDA(ctx, cudaEventElapsedTime(&ms_w, ctx->ev1, ev2));
  cudaEventDestroy(ev2);
  ctx->stats.build_ms = ms;
  ctx->stats.build_wide_ms = ms_w;
  return GPT_OK;
}

extern "C" {

int gpt_build(gpt_ctx *ctx) {
  if (!ctx) return GPT_E_INVALID;
  if (!ctx->have_mesh || !ctx->have_mats)
    return gpt_fail(ctx, GPT_E_INVALID, "gpt_set_mesh and gpt_set_materials first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  SceneDev &s = ctx->scene;
  if (s.ntri < 2)
    return gpt_fail(ctx, GPT_E_INVALID, "need >= 2 triangles (the reference's root is an internal node)", __FILE__, __LINE__);
  if (!s.tex_r) {  // default 1x1 white atlas (read_image_util.h:135-144)
    float one = 1.f;
    GPT_TRY(gpt_set_textures(ctx, &one, &one, &one, 1));
  }
  if (!s.bg_r) {
    float one = 1.f;
    GPT_TRY(gpt_set_background(ctx, &one, &one, &one, 1, 1));
  }
  ctx->launches = 0;
  cudaEvent_t a, b;
  GPT_CUDA(ctx, cudaEventCreate(&a));
  GPT_CUDA(ctx, cudaEventCreate(&b));
  GPT_CUDA(ctx, cudaEventRecord(a, ctx->stream));
  GPT_TRY(gpt_preprocess(ctx));
  GPT_TRY(build_main_timed(ctx));
  // transparent and light subsets, collected from the SORTED main list (src/main.cu:952,1083)
  const int which[2] = {GPT_BVH_TRANSPARENT, GPT_BVH_LIGHT};
  const int mask[2] = {GPT_TRI_TRANSPARENT, GPT_TRI_LIGHT};
  for (int k = 0; k < 2; k++) {
    BvhDev &bv = ctx->bvh[which[k]];
    int cnt = 0;
    GPT_TRY(gpt_collect_subset(ctx, mask[k], &bv.prim_list, &cnt));
    GPT_TRY(gpt_lbvh_build(ctx, which[k], bv.prim_list, cnt));
  }
  // SSS objects: surface point clouds + per-object point LBVHs (src/main.cu:642-824)
  GPT_TRY(gpt_sss_prepare(ctx));
  GPT_CUDA(ctx, cudaEventRecord(b, ctx->stream));
  GPT_CUDA(ctx, cudaEventSynchronize(b));
  float ms = 0.f;
  GPT_CUDA(ctx, cudaEventElapsedTime(&ms, a, b));
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  ctx->stats.build_all_ms = ms;
  ctx->stats.kernel_launches = ctx->launches;
  ctx->built = true;
  ctx->build_epoch++;
  return GPT_OK;
}

int gpt_rebuild_main_bvh(gpt_ctx *ctx) {
  if (!ctx || !ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->launches = 0;
  GPT_TRY(build_main_timed(ctx));
  ctx->stats.kernel_launches = ctx->launches;
  return GPT_OK;
}

int gpt_export_triangles(gpt_ctx *ctx, float *tolerance, float *aabb, float *normal, float *area,
                         int32_t *flags, float *energy, float *tangents) {
  if (!ctx || !ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  SceneDev &s = ctx->scene;
  size_t n = (size_t)s.ntri;
  auto dl = [&](void *h, const void *d, size_t bytes) -> cudaError_t {
    return h ? cudaMemcpy(h, d, bytes, cudaMemcpyDeviceToHost) : cudaSuccess;
  };
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  GPT_CUDA(ctx, dl(tolerance, s.tolerance, 4 * n));
  GPT_CUDA(ctx, dl(aabb, s.aabb, 24 * n));
  GPT_CUDA(ctx, dl(normal, s.normal, 12 * n));
  GPT_CUDA(ctx, dl(energy, s.energy, 12 * n));
  if (area || flags || tangents) {
    std::vector<float4> sh(7 * n);
    GPT_CUDA(ctx, cudaMemcpy(sh.data(), s.shade, sizeof(float4) * 7 * n, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) {
      const float4 *o = &sh[7 * i];
      if (area) area[i] = o[3].w;
      if (flags) memcpy(&flags[i], &o[6].z, 4);
      if (tangents) {
        float *t = tangents + 9 * i;
        t[0] = o[4].x; t[1] = o[4].y; t[2] = o[4].z; t[3] = o[4].w; t[4] = o[5].x;
        t[5] = o[5].y; t[6] = o[5].z; t[7] = o[5].w; t[8] = o[6].x;
      }
    }
  }
  return GPT_OK;
}

int gpt_bvh_num_leaves(gpt_ctx *ctx, int which, int32_t *n) {
  if (!ctx || !n || which < 0 || which > 2) return GPT_E_INVALID;
  *n = ctx->bvh[which].n;
  return GPT_OK;
}

int gpt_export_bvh(gpt_ctx *ctx, int which, uint32_t *morton, int32_t *sorted_to_orig,
                   int32_t *left, int32_t *right, int32_t *parent_internal,
                   int32_t *parent_leaf, float *node_aabb, float *leaf_aabb, float *world_aabb) {
  if (!ctx || !ctx->built || which < 0 || which > 2)
    return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first / bad BVH id", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  return gpt_lbvh_export(ctx, which, morton, sorted_to_orig, left, right, parent_internal,
                         parent_leaf, node_aabb, leaf_aabb, world_aabb);
}

static int check_overflow(gpt_ctx *ctx) {
  unsigned long long c[4];
  GPT_CUDA(ctx, cudaMemcpyAsync(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (c[3]) return gpt_fail(ctx, GPT_E_UNSUPPORTED, "traversal stack overflow (> 64 levels)", __FILE__, __LINE__);
  return GPT_OK;
}

int gpt_trace_device(gpt_ctx *ctx, int which, const void *d_rays, int64_t n, void *d_hits) {
  if (!ctx || !ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  ctx->launches = 0;
  GPT_TRY(gpt_trace_launch(ctx, which, (const float4 *)d_rays, n, (float4 *)d_hits, true, false));
  ctx->stats.kernel_launches = ctx->launches;
  ctx->stats.rays_traced = (uint64_t)n;
  return GPT_OK;
}

int gpt_trace_count_device(gpt_ctx *ctx, int which, const void *d_rays, int64_t n) {
  if (!ctx || !ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  float4 *hits = nullptr;
  GPT_TRY(gpt_alloc(ctx, &hits, 2 * (size_t)n));
  int rc = gpt_trace_launch(ctx, which, (const float4 *)d_rays, n, hits, true, true);
  unsigned long long c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (rc == GPT_OK) {
    cudaMemcpyAsync(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
  }
  gpt_free(ctx, hits);
  GPT_TRY(rc);
  ctx->stats.internal_visits = c[0];
  ctx->stats.leaf_tests = c[1];
  ctx->stats.wide_node_visits = c[5];
  ctx->stats.wide_leaf_tests = c[7];
  ctx->stats.rays_traced = (uint64_t)n;
  return GPT_OK;
}

int gpt_trace(gpt_ctx *ctx, int which, const float *rays, int64_t n, int32_t *leaf_id,
              int32_t *tri_id, float *t, float *bary) {
  if (!ctx || !ctx->built) return gpt_fail(ctx, GPT_E_INVALID, "gpt_build first", __FILE__, __LINE__);
  if (n < 0 || (n > 0 && !rays)) return GPT_E_INVALID;
  if (n == 0) return GPT_OK;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  float4 *d_rays = nullptr, *d_hits = nullptr;
  GPT_TRY(gpt_alloc(ctx, &d_rays, 2 * (size_t)n));
  int rc = gpt_alloc(ctx, &d_hits, 2 * (size_t)n);
  if (rc != GPT_OK) {
    gpt_free(ctx, d_rays);
    return rc;
  }
  std::vector<float4> h(2 * (size_t)n);
  do {
    if (cudaMemcpyAsync(d_rays, rays, 32 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { rc = GPT_E_CUDA; break; }
    cudaEventRecord(ctx->ev0, ctx->stream);
    ctx->launches = 0;
    rc = gpt_trace_launch(ctx, which, d_rays, n, d_hits, true, false);
    if (rc != GPT_OK) break;
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (cudaMemcpyAsync(h.data(), d_hits, 32 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) { rc = GPT_E_CUDA; break; }
    rc = check_overflow(ctx);
  } while (0);
  gpt_free(ctx, d_rays);
  gpt_free(ctx, d_hits);
  GPT_TRY(rc);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  ctx->stats.trace_ms = ms;
  ctx->stats.rays_traced = (uint64_t)n;
  ctx->stats.kernel_launches = ctx->launches;
  for (int64_t i = 0; i < n; i++) {
    const float4 &a = h[2 * i], &b = h[2 * i + 1];
    if (t) t[i] = a.x;
    if (leaf_id) memcpy(&leaf_id[i], &a.y, 4);
    if (tri_id) memcpy(&tri_id[i], &a.z, 4);
    if (bary) { bary[3 * i] = b.x; bary[3 * i + 1] = b.y; bary[3 * i + 2] = b.z; }
  }
  return GPT_OK;
}

int gpt_set_profiling(gpt_ctx *ctx, int time_stages, int count_visits) {
  if (!ctx) return GPT_E_INVALID;
  ctx->prof_time_stages = time_stages != 0;
  ctx->prof_count_visits = count_visits != 0;
  return GPT_OK;
}

int gpt_set_wavefront_groups(gpt_ctx *ctx, int groups) {
  if (!ctx || groups < 0 || groups > 8) return GPT_E_INVALID;
  ctx->wavefront_groups = groups;
  return GPT_OK;
}

int gpt_set_trace_tail(gpt_ctx *ctx, int mode) {
  if (!ctx || mode < -1 || mode > 1) return GPT_E_INVALID;
  ctx->trace_tail = mode;
  return GPT_OK;
}

int gpt_set_fast_rng_lanes(gpt_ctx *ctx, int lanes) {
  if (!ctx || lanes < 0 || lanes > 64) return GPT_E_INVALID;
  ctx->fast_lanes = lanes;
  return GPT_OK;
}

int gpt_set_trace_bvh(gpt_ctx *ctx, int mode) {
  if (!ctx || mode < -1 || mode > 1) return GPT_E_INVALID;
  ctx->trace_wide = mode;
  return GPT_OK;
}

int gpt_set_pipeline(gpt_ctx *ctx, int mode) {
  if (!ctx || mode < -1 || mode > 1) return GPT_E_INVALID;
  ctx->pipeline = mode;
  return GPT_OK;
}

int gpt_set_trace_leaf_batching(gpt_ctx *ctx, int mode) {
  if (!ctx || mode < -1 || mode > 1) return GPT_E_INVALID;
  ctx->trace_defer = mode;
  return GPT_OK;
}

int gpt_stats(gpt_ctx *ctx, GptStats *out) {
  if (!ctx || !out) return GPT_E_INVALID;
  *out = ctx->stats;
  return GPT_OK;
}

}  // extern "C"
