/*
 * ref_replay.cpp -- TEST INFRASTRUCTURE ONLY.  Host replay of the reference's OWN code.
 *
 * Builds oracle/_ref/libref_replay.so from the headers under /root/reference/src (included
 * in place, nothing copied) through oracle/ref_shim.h.  Every function here only marshals
 * the C-ABI arrays of include/gpt.h into the reference's classes and then calls the
 * reference's own kernels/functions by looping blockIdx/threadIdx exactly the way
 * src/main.cu launches them.  It is the ground truth the CUDA path and the C oracle
 * (oracle/oracle.c) are pinned against, and the "reference" CPU baseline of bench.py.
 *
 * It is never imported by the product path (gpu_pathtracer_b200/), only by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 */
#include "ref_shim.h"

// same include order as src/main.cu:10-45 (minus the jpeg / input / world.h host loaders)
#include "model/geometry/triangle.h"
#include "model/bvh/bvh.h"
#include "model/bvh/bvh_traversal.h"

// Ray counter for the Mrays/s baseline: every later CALL of traverse_bvh (the integrators in
// pathtracing.h, photon_pass.h, ray_tracing_pass.h, bvh_traversal_photon.h) goes through this
// wrapper, which counts closest-hit traversals of the MAIN tree and forwards to the
// reference's own function.  The reference source is untouched.
static Node *g_counted_root = nullptr;
static thread_local unsigned long long tl_ray_count = 0;
static inline bool counted_traverse_bvh(Node *root, Ray ray, hit_record &rec) {
  if (root == g_counted_root) tl_ray_count++;
  return (traverse_bvh)(root, ray, rec);
}
#define traverse_bvh(...) counted_traverse_bvh(__VA_ARGS__)

#include "model/bvh/bvh_building.h"
#include "model/bvh/bvh_building_photon.h"
#include "model/bvh/bvh_building_pts.h"
#include "model/bvh/bvh_traversal_photon.h"
#include "model/camera.h"
#include "model/geometry/primitive.h"
#include "model/geometry/triangle_operations.h"
#include "model/grid/bounding_box.h"
#include "model/grid/bounding_box_operations.h"
#include "model/material/material.h"
#include "model/object/object.h"
#include "model/object/object_operations.h"
#include "model/point/point.h"
#include "model/point/ppm_hit_point.h"
#include "model/point/point_operations.h"
#include "model/ray/ray.h"
#include "model/vector_and_matrix/vec3.h"
#include "render/pathtracing.h"
#include "render/pathtracing_target_geom_operations.h"
#include "render/ppm/image_output.h"
#include "render/ppm/photon_pass.h"
#include "render/ppm/ray_tracing_pass.h"
#include "render/transparent_geom_operations.h"
#include "util/general.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../include/gpt.h"

// The reference never initialises Node::parent (src/model/bvh/bvh.h:42-54), yet
// compute_node_bounding_boxes stops climbing on `parent != nullptr`
// (src/model/bvh/bvh_building.h:238-243): it silently relies on the CUDA device heap handing
// out zero-filled memory.  The host replay reproduces that by zero-filling every `new`.
#include <new>
void *operator new(std::size_t n) {
  void *p = calloc(1, n ? n : 1);
  if (!p) throw std::bad_alloc();
  return p;
}
void *operator new[](std::size_t n) {
  void *p = calloc(1, n ? n : 1);
  if (!p) throw std::bad_alloc();
  return p;
}
void operator delete(void *p) noexcept { free(p); }
void operator delete[](void *p) noexcept { free(p); }
void operator delete(void *p, std::size_t) noexcept { free(p); }
void operator delete[](void *p, std::size_t) noexcept { free(p); }

namespace {

// Emulates `kernel<<<grid, block>>>(...)` with 1-D grid/block: one host call per thread.
template <class F>
void launch_1d(unsigned grid, unsigned block, F f) {
  shim_gridDim = dim3(grid, 1, 1);
  shim_blockDim = dim3(block, 1, 1);
  for (unsigned b = 0; b < grid; b++) {
    shim_blockIdx = {b, 0, 0};
    for (unsigned t = 0; t < block; t++) {
      shim_threadIdx = {t, 0, 0};
      f();
    }
  }
  shim_blockIdx = {0, 0, 0};
  shim_threadIdx = {0, 0, 0};
}

struct Bvh {
  int n = 0;
  Primitive **geom = nullptr;  // sorted in place like the reference
  Node **node_list = nullptr, **leaf_list = nullptr;
  BoundingBox **world = nullptr;
  std::vector<unsigned int> morton;  // sorted codes (extract_morton_code_list)
};

struct RefCtx {
  // inputs
  std::vector<float> x, y, z, xn, yn, zn, xt, yt;
  std::vector<int> p1, p2, p3, n1, n2, n3, t1, t2, t3, mat_idx, obj_idx;
  std::vector<int> obj_num, obj_off;
  int num_triangles = 0, num_vertices = 0, num_objects = 0;
  std::vector<float> tex_r, tex_g, tex_b, bg_r, bg_g, bg_b;
  int bg_h = 1, bg_w = 1;
  GptMaterials mats{};
  std::vector<std::vector<float>> mat_f;  // owned copies
  std::vector<std::vector<int>> mat_i;
  GptCamera cam{};
  GptParams params{};
  bool have_mesh = false, have_mats = false, have_cam = false;

  // reference objects
  Material **my_material = nullptr;
  Primitive **my_geom = nullptr;  // ORIGINAL order copy (my_geom_sorted is the sorted one)
  Camera **my_camera = nullptr;
  std::unordered_map<Primitive *, int> orig_index;
  Bvh main_bvh, transparent_bvh, light_bvh;
  std::string err;
  double last_ms = 0;
  unsigned long long last_rays = 0;
  struct Sppm *sppm = nullptr;
  struct Sss *sss = nullptr;
};

// SSS state, the locals of src/main.cu:555-824
struct Sss {
  Object **my_objects = nullptr;
  float *triangle_area = nullptr, *accumulated_triangle_area = nullptr;
  int num_sss_objects = 0, num_sss_points = 0, pts_per_object = 0;
  int *pt_offset_array = nullptr, *num_pt_array = nullptr;
  Point **sss_pts = nullptr;
  curandState *rand_state_sss = nullptr;
  Node **node_list = nullptr, **leaf_list = nullptr;
  unsigned int *morton = nullptr;
  bool first_pass_done = false;
};

// SPPM state, the locals of main()'s render_mode == 2 branch (src/main.cu:1282-1494)
struct Sppm {
  int P = 0, nphot = 0;
  PPMHitPoint **hit_point_list = nullptr;
  Point **photon_list = nullptr, **surface_photon_list = nullptr, **volume_photon_list = nullptr;
  Node **surface_photon_node_list = nullptr, **surface_photon_leaf_list = nullptr;
  Node **volume_photon_node_list = nullptr, **volume_photon_leaf_list = nullptr;
  float *acc_energy = nullptr;
  curandState *rand_state_ppm = nullptr, *rand_state_image = nullptr;
  unsigned int *surface_morton = nullptr, *volume_morton = nullptr;
  int num_surface = 0, num_volume = 0;
  float average_radius = 0.f;
  unsigned long long rays = 0;
};

void build_bvh(Bvh &b, bool cones) {
  // src/main.cu:826-918 (main), :959-1063 (transparent), :1090-1197 (light)
  int n = b.n;
  b.world = new BoundingBox *[1];
  launch_1d(1, 1, [&] { compute_world_bounding_box(b.world, b.geom, n); });
  launch_1d(n + 1, 1, [&] { compute_morton_code_batch(b.geom, b.world, n); });
  // thrust::stable_sort(thrust::device, geom, geom + n, sort_geom)  src/main.cu:846-855
  std::stable_sort(b.geom, b.geom + n, [](Primitive *o1, Primitive *o2) {
    return o1->get_bounding_box()->morton_code < o2->get_bounding_box()->morton_code;
  });
  b.node_list = new Node *[std::max(1, n - 1)];
  b.leaf_list = new Node *[std::max(1, n)];
  launch_1d(std::max(1, n), 1, [&] { build_leaf_list(b.leaf_list, b.geom, n, cones); });
  launch_1d(std::max(1, n), 1, [&] { build_node_list(b.node_list, n, cones); });
  b.morton.assign(std::max(1, n), 0u);
  launch_1d(std::max(1, n), 1,
            [&] { extract_morton_code_list(b.geom, b.morton.data(), n); });
  launch_1d(std::max(1, n), 1, [&] {
    set_node_relationship(b.node_list, b.leaf_list, b.morton.data(), n);
  });
  launch_1d(std::max(1, n), 1,
            [&] { compute_node_bounding_boxes(b.leaf_list, b.node_list, n); });
  if (cones) {
    launch_1d(std::max(1, n), 1,
              [&] { compute_node_bounding_cones(b.leaf_list, b.node_list, n); });
  }
}

int child_ref(Node *c) { return c->is_leaf ? ~c->idx : c->idx; }

Bvh *pick(RefCtx *c, int which) {
  if (which == GPT_BVH_MAIN) return &c->main_bvh;
  if (which == GPT_BVH_TRANSPARENT) return &c->transparent_bvh;
  if (which == GPT_BVH_LIGHT) return &c->light_bvh;
  return nullptr;
}

}  // namespace

// SSS pre-pass set-up, src/main.cu:555-824: objects, area CDFs, surface point samples, the
// per-object point LBVHs.  create_sss_pts is launched once PER POINT with rand_state =
// &state[idx] (the reference passes the global array to get_random_point_on_surface, so all
// threads race on rand_state[0], object_operations.h:135-136; the per-point launch makes
// rand_state[0] that point's own state -- deterministic, same convention as the photons).
static void sss_prepare(RefCtx *c) {
  int pts = c->params.sss_pts_per_object, nt = c->num_triangles, nobj = c->num_objects;
  if (pts <= 0 || nobj <= 0) return;
  Sss *s = new Sss();
  s->pts_per_object = pts;
  s->triangle_area = new float[nt];
  s->accumulated_triangle_area = new float[nt];
  s->my_objects = new Object *[nobj];
  for (int i = 0; i < nobj; i++)  // create_objects (src/lib/world.h:177-192): marshalling only
    s->my_objects[i] = new Object(c->obj_off[i], c->obj_num[i], s->triangle_area + c->obj_off[i],
                                  s->accumulated_triangle_area + c->obj_off[i]);
  bool any = false;
  for (int i = 0; i < nt; i++) {  // the tail of create_world (src/lib/world.h:236-246)
    s->triangle_area[i] = c->my_geom[i]->get_area();
    if (c->my_geom[i]->is_sub_surface_scattering()) {
      s->my_objects[c->obj_idx[i]]->set_as_sub_surface_scattering(pts);
      any = true;
    }
  }
  if (!any) { delete s; return; }
  s->pt_offset_array = new int[nobj];
  s->num_pt_array = new int[nobj];
  launch_1d(1, 1, [&] {
    compute_num_sss_objects(&s->num_sss_objects, s->my_objects, s->pt_offset_array,
                            s->num_pt_array, nobj, pts);
  });
  int n = pts * s->num_sss_objects;
  s->num_sss_points = n;
  s->sss_pts = new Point *[std::max(1, n)];
  launch_1d(nobj, 1, [&] { allocate_pts_sss(s->my_objects, s->sss_pts, s->pt_offset_array, nobj); });
  s->rand_state_sss = new curandState[std::max(1, n)];
  launch_1d(std::max(1, n), 1, [&] { init_curand_state(n, s->rand_state_sss); });
  for (int i = 0; i < nobj; i++) {
    if (!s->my_objects[i]->sub_surface_scattering) continue;
    for (int idx = 0; idx < pts; idx++) {
      // one launch per point: the kernel's thread 0 sees rand_state[0] = state[idx] and
      // writes point_array[pt_offset[i] + 0]
      launch_1d(1, 1, [&] {
        create_sss_pts(s->my_objects, c->my_geom, s->sss_pts + idx, s->pt_offset_array,
                       s->rand_state_sss + idx, i, pts);
      });
    }
  }
  launch_1d(nobj, 1, [&] { compute_object_boundaries_batch(s->my_objects, nobj); });
  launch_1d(nt + 1, 1, [&] { compute_pts_morton_code_batch(s->my_objects, s->sss_pts, n); });
  for (int i = 0; i < nobj; i++)
    std::stable_sort(s->sss_pts + s->pt_offset_array[i],
                     s->sss_pts + s->pt_offset_array[i] + s->num_pt_array[i],
                     [](Point *a, Point *b) {
                       return a->bounding_box->morton_code < b->bounding_box->morton_code;
                     });
  launch_1d(1, 1, [&] { compute_sss_pts_offset(s->my_objects, nobj); });
  s->node_list = new Node *[std::max(1, n - s->num_sss_objects)];
  s->leaf_list = new Node *[std::max(1, n)];
  for (int i = 0; i < nobj; i++)
    launch_1d(std::max(1, n), 1, [&] {
      build_sss_pts_leaf_list(s->leaf_list, s->sss_pts, s->my_objects, i, s->pt_offset_array);
    });
  for (int i = 0; i < nobj; i++)
    launch_1d(std::max(1, n), 1,
              [&] { build_sss_pts_node_list(s->node_list, s->my_objects, i); });
  s->morton = new unsigned int[std::max(1, n)];
  launch_1d(std::max(1, n), 1, [&] { extract_sss_morton_code_list(s->sss_pts, s->morton, n); });
  for (int i = 0; i < nobj; i++)
    launch_1d(std::max(1, n), 1, [&] {
      set_pts_sss_node_relationship(s->node_list, s->leaf_list, s->morton, s->my_objects, i);
    });
  launch_1d(std::max(1, n), 1,
            [&] { compute_node_bounding_boxes(s->leaf_list, s->node_list, n); });
  c->sss = s;
}

// do_sss_first_pass (src/main.cu:1213-1232, pathtracing.h:202-249) over all points, threaded
static void sss_first_pass(RefCtx *c, int nthreads) {
  Sss *s = c->sss;
  if (!s || s->first_pass_done) return;
  const GptParams &p = c->params;
  vec3 sky(p.sky_emission[0], p.sky_emission[1], p.sky_emission[2]);
  int n = s->num_sss_points;
  std::atomic<int> next(0);
  auto work = [&] {
    shim_blockDim = dim3(1, 1, 1);
    shim_threadIdx = {0, 0, 0};
    for (;;) {
      int k = next.fetch_add(64);
      if (k >= n) break;
      for (int q = k; q < std::min(n, k + 64); q++) {
        shim_blockIdx = {(unsigned)q, 0, 0};
        do_sss_first_pass(s->sss_pts, n, p.pathtracing_sample_size, p.pathtracing_level, sky,
                          c->bg_h, c->bg_w, c->bg_r.data(), c->bg_g.data(), c->bg_b.data(),
                          c->main_bvh.node_list, s->rand_state_sss, c->light_bvh.geom,
                          c->light_bvh.node_list, c->light_bvh.leaf_list, c->light_bvh.n,
                          p.hittable_pdf_weight);
      }
    }
    shim_blockIdx = {0, 0, 0};
  };
  if (nthreads <= 1) work();
  else {
    std::vector<std::thread> th;
    for (int k = 0; k < nthreads; k++) th.emplace_back(work);
    for (auto &x : th) x.join();
  }
  s->first_pass_done = true;
}

extern "C" {

int ref_create(RefCtx **out) {
  *out = new RefCtx();
  return GPT_OK;
}

void ref_destroy(RefCtx *c) { delete c; /* reference objects are leaked like main.cu does */ }

int ref_set_mesh(RefCtx *c, const GptMesh *m) {
  int nt = m->num_triangles;
  auto cpf = [](std::vector<float> &d, const float *s, int n) {
    d.assign(s, s + std::max(0, n));
    if (d.empty()) d.push_back(0.f);  // max(1, n) allocations, src/main.cu:412-429
  };
  cpf(c->x, m->x, m->num_vertices);
  cpf(c->y, m->y, m->num_vertices);
  cpf(c->z, m->z, m->num_vertices);
  cpf(c->xn, m->x_norm, m->num_normals);
  cpf(c->yn, m->y_norm, m->num_normals);
  cpf(c->zn, m->z_norm, m->num_normals);
  cpf(c->xt, m->x_tex, m->num_texcoords);
  cpf(c->yt, m->y_tex, m->num_texcoords);
  auto cpi = [&](std::vector<int> &d, const int *s) { d.assign(s, s + nt); };
  cpi(c->p1, m->point_1_idx);
  cpi(c->p2, m->point_2_idx);
  cpi(c->p3, m->point_3_idx);
  cpi(c->n1, m->norm_1_idx);
  cpi(c->n2, m->norm_2_idx);
  cpi(c->n3, m->norm_3_idx);
  cpi(c->t1, m->tex_1_idx);
  cpi(c->t2, m->tex_2_idx);
  cpi(c->t3, m->tex_3_idx);
  cpi(c->mat_idx, m->material_idx);
  cpi(c->obj_idx, m->triangle_object_idx);
  c->num_objects = m->num_objects;
  c->obj_num.assign(m->object_num_primitives, m->object_num_primitives + m->num_objects);
  c->obj_off.assign(m->object_primitive_offset_idx,
                    m->object_primitive_offset_idx + m->num_objects);
  c->num_triangles = nt;
  c->num_vertices = m->num_vertices;
  c->have_mesh = true;
  return GPT_OK;
}

int ref_set_materials(RefCtx *c, const GptMaterials *m) {
  int n = m->num_materials;
  c->mat_f.clear();
  c->mat_i.clear();
  c->mat_f.reserve(64);
  c->mat_i.reserve(64);
  GptMaterials o = *m;
  auto F = [&](const float *&p) {
    c->mat_f.emplace_back(p, p + n);
    p = c->mat_f.back().data();
  };
  auto I = [&](const int *&p) {
    c->mat_i.emplace_back(p, p + n);
    p = c->mat_i.back().data();
  };
  F(o.ka_x); F(o.ka_y); F(o.ka_z); F(o.kd_x); F(o.kd_y); F(o.kd_z);
  F(o.ks_x); F(o.ks_y); F(o.ks_z); F(o.ke_x); F(o.ke_y); F(o.ke_z);
  F(o.tf_x); F(o.tf_y); F(o.tf_z);
  F(o.path_length); F(o.t_r); F(o.n_s); F(o.n_i); F(o.bm);
  F(o.scattering_coef); F(o.absorption_coef); F(o.g);
  I(o.material_priority);
  I(o.height_diffuse); I(o.width_diffuse); I(o.offset_diffuse);
  I(o.height_specular); I(o.width_specular); I(o.offset_specular);
  I(o.height_emission); I(o.width_emission); I(o.offset_emission);
  I(o.height_n_s); I(o.width_n_s); I(o.offset_n_s);
  I(o.height_bump); I(o.width_bump); I(o.offset_bump);
  c->mats = o;
  c->have_mats = true;
  return GPT_OK;
}

int ref_set_textures(RefCtx *c, const float *r, const float *g, const float *b,
                     int64_t n) {
  c->tex_r.assign(r, r + n);
  c->tex_g.assign(g, g + n);
  c->tex_b.assign(b, b + n);
  return GPT_OK;
}

int ref_set_background(RefCtx *c, const float *r, const float *g, const float *b, int h,
                       int w) {
  c->bg_r.assign(r, r + (size_t)h * w);
  c->bg_g.assign(g, g + (size_t)h * w);
  c->bg_b.assign(b, b + (size_t)h * w);
  c->bg_h = h;
  c->bg_w = w;
  return GPT_OK;
}

int ref_set_camera(RefCtx *c, const GptCamera *cam) {
  c->cam = *cam;
  c->have_cam = true;
  return GPT_OK;
}

int ref_set_params(RefCtx *c, const GptParams *p) {
  c->params = *p;
  return GPT_OK;
}

int ref_build(RefCtx *c) {
  if (!c->have_mesh || !c->have_mats) return GPT_E_INVALID;
  const GptMaterials &m = c->mats;
  int nm = m.num_materials, nt = c->num_triangles;
  if (c->tex_r.empty()) {  // the 1x1 white default texture, read_image_util.h:135-144
    c->tex_r = {1.f}; c->tex_g = {1.f}; c->tex_b = {1.f};
  }
  float *ir = c->tex_r.data(), *ig = c->tex_g.data(), *ib = c->tex_b.data();

  // create_material<<<1, num_materials>>>  src/main.cu:497, src/lib/world.h:101-175
  c->my_material = new Material *[100];
  for (int i = 0; i < nm; i++) {
    c->my_material[i] = new Material(
        vec3(m.ka_x[i], m.ka_y[i], m.ka_z[i]), vec3(m.kd_x[i], m.kd_y[i], m.kd_z[i]),
        vec3(m.ks_x[i], m.ks_y[i], m.ks_z[i]), vec3(m.ke_x[i], m.ke_y[i], m.ke_z[i]),
        vec3(m.tf_x[i], m.tf_y[i], m.tf_z[i]), m.path_length[i], m.t_r[i], m.n_s[i],
        m.n_i[i], m.bm[i], m.scattering_coef[i], m.absorption_coef[i], m.g[i],
        m.material_priority[i],
        m.height_diffuse[i], m.width_diffuse[i], ir + m.offset_diffuse[i],
        ig + m.offset_diffuse[i], ib + m.offset_diffuse[i],
        m.height_specular[i], m.width_specular[i], ir + m.offset_specular[i],
        ig + m.offset_specular[i], ib + m.offset_specular[i],
        m.height_emission[i], m.width_emission[i], ir + m.offset_emission[i],
        ig + m.offset_emission[i], ib + m.offset_emission[i],
        m.height_n_s[i], m.width_n_s[i], ir + m.offset_n_s[i], ig + m.offset_n_s[i],
        ib + m.offset_n_s[i],
        m.height_bump[i], m.width_bump[i], ir + m.offset_bump[i], ig + m.offset_bump[i],
        ib + m.offset_bump[i]);
  }

  // create_world<<<N+1, 1>>>  src/main.cu:582, src/lib/world.h:194-248
  Primitive **geom = new Primitive *[nt];
  c->my_geom = new Primitive *[nt];
  const float *x = c->x.data(), *y = c->y.data(), *z = c->z.data();
  const float *xn = c->xn.data(), *yn = c->yn.data(), *zn = c->zn.data();
  const float *xt = c->xt.data(), *yt = c->yt.data();
  for (int i = 0; i < nt; i++) {
    int a = c->p1[i], b = c->p2[i], d = c->p3[i];
    geom[i] = new Triangle(
        vec3(x[a], y[a], z[a]), vec3(x[b], y[b], z[b]), vec3(x[d], y[d], z[d]), a, b, d,
        c->my_material[c->mat_idx[i]], c->obj_idx[i],
        vec3(xn[c->n1[i]], yn[c->n1[i]], zn[c->n1[i]]),
        vec3(xn[c->n2[i]], yn[c->n2[i]], zn[c->n2[i]]),
        vec3(xn[c->n3[i]], yn[c->n3[i]], zn[c->n3[i]]),
        vec3(xt[c->t1[i]], yt[c->t1[i]], 0), vec3(xt[c->t2[i]], yt[c->t2[i]], 0),
        vec3(xt[c->t3[i]], yt[c->t3[i]], 0));
    c->my_geom[i] = geom[i];
    c->orig_index[geom[i]] = i;
  }

  // tangents  src/main.cu:602-625
  std::vector<vec3> tangent(std::max(1, c->num_vertices)), bitangent(tangent.size());
  launch_1d(1, 1, [&] {
    sum_up_tangent_and_bitangent(tangent.data(), bitangent.data(), geom, nt);
  });
  launch_1d(nt, 1, [&] { assign_tangent(tangent.data(), geom, nt); });

  // SSS objects / points / per-object point trees  src/main.cu:555-824 (before the sort)
  sss_prepare(c);

  // main BVH  src/main.cu:826-918
  c->main_bvh.n = nt;
  c->main_bvh.geom = geom;
  build_bvh(c->main_bvh, false);

  // transparent subset, collected from the SORTED main list  src/main.cu:930-1063
  {
    int cnt = 0;
    launch_1d(1, 1, [&] { compute_num_transparent_geom(geom, nt, &cnt); });
    Bvh &b = c->transparent_bvh;
    b.n = cnt;
    b.geom = new Primitive *[std::max(1, cnt)];
    launch_1d(1, 1, [&] { collect_transparent_geom(geom, nt, b.geom); });
    if (cnt >= 2) build_bvh(b, false);
  }
  // light ("target") subset  src/main.cu:1065-1197
  {
    int cnt = 0;
    launch_1d(1, 1, [&] { compute_num_target_geom(geom, nt, &cnt); });
    Bvh &b = c->light_bvh;
    b.n = cnt;
    b.geom = new Primitive *[std::max(1, cnt)];
    launch_1d(1, 1, [&] { collect_target_geom(geom, nt, b.geom); });
    if (cnt >= 2) build_bvh(b, true);
  }

  if (c->have_cam) {
    const GptCamera &k = c->cam;
    c->my_camera = new Camera *[1];
    c->my_camera[0] = new Camera(vec3(k.eye[0], k.eye[1], k.eye[2]),
                                 vec3(k.center[0], k.center[1], k.center[2]),
                                 vec3(k.up[0], k.up[1], k.up[2]), k.fovy, k.width,
                                 k.height, k.aperture, k.focus_dist);
  }
  return GPT_OK;
}

// Times only the main-BVH stage (src/main.cu:826-918) on the already constructed triangles;
// used by bench.py's cpu_baseline leg.  The geometry list is restored to OBJ order first.
int ref_time_main_bvh_build(RefCtx *c, double *ms) {
  int nt = c->num_triangles;
  Primitive **geom = new Primitive *[nt];
  memcpy(geom, c->my_geom, sizeof(Primitive *) * nt);
  Bvh b;
  b.n = nt;
  b.geom = geom;
  auto t0 = std::chrono::steady_clock::now();
  build_bvh(b, false);
  auto t1 = std::chrono::steady_clock::now();
  *ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
  // the triangles' boxes now carry the main-BVH codes again; keep this tree as the main one
  c->main_bvh = b;
  return GPT_OK;
}

int ref_export_triangles(RefCtx *c, float *tolerance, float *aabb, float *normal,
                         float *area, int32_t *flags, float *energy, float *tangents) {
  for (int i = 0; i < c->num_triangles; i++) {
    Triangle *t = static_cast<Triangle *>(c->my_geom[i]);
    const Triangle *p = t;
    BoundingBox *bb = t->get_bounding_box();
    if (tolerance) tolerance[i] = p->tolerance;
    if (aabb) {
      float *a = aabb + 6 * i;
      a[0] = bb->x_min; a[1] = bb->x_max; a[2] = bb->y_min;
      a[3] = bb->y_max; a[4] = bb->z_min; a[5] = bb->z_max;
    }
    if (normal) {
      vec3 n = t->get_fixed_normal();
      normal[3 * i] = n.x(); normal[3 * i + 1] = n.y(); normal[3 * i + 2] = n.z();
    }
    if (area) area[i] = t->get_area();
    if (flags)
      flags[i] = (t->is_light_source() ? GPT_TRI_LIGHT : 0) |
                 (t->is_transparent_geom() ? GPT_TRI_TRANSPARENT : 0) |
                 (t->is_sub_surface_scattering() ? GPT_TRI_SSS : 0);
    if (energy) {
      vec3 e = t->get_energy();
      energy[3 * i] = e.x(); energy[3 * i + 1] = e.y(); energy[3 * i + 2] = e.z();
    }
    if (tangents) {
      const vec3 *tg[3] = {&p->tangent_1, &p->tangent_2, &p->tangent_3};
      for (int k = 0; k < 3; k++)
        for (int a = 0; a < 3; a++) tangents[9 * i + 3 * k + a] = (*tg[k])[a];
    }
  }
  return GPT_OK;
}

int ref_bvh_num_leaves(RefCtx *c, int which, int32_t *n) {
  Bvh *b = pick(c, which);
  if (!b) return GPT_E_INVALID;
  *n = b->n;
  return GPT_OK;
}

int ref_export_bvh(RefCtx *c, int which, uint32_t *morton, int32_t *sorted_to_orig,
                   int32_t *left, int32_t *right, int32_t *parent_internal,
                   int32_t *parent_leaf, float *node_aabb, float *leaf_aabb,
                   float *world_aabb) {
  Bvh *b = pick(c, which);
  if (!b || b->n < 2) return GPT_E_INVALID;
  int n = b->n;
  auto put = [](float *a, BoundingBox *bb) {
    a[0] = bb->x_min; a[1] = bb->x_max; a[2] = bb->y_min;
    a[3] = bb->y_max; a[4] = bb->z_min; a[5] = bb->z_max;
  };
  for (int j = 0; j < n; j++) {
    if (morton) morton[j] = b->morton[j];
    if (sorted_to_orig) sorted_to_orig[j] = c->orig_index[b->geom[j]];
    if (parent_leaf) parent_leaf[j] = b->leaf_list[j]->parent->idx;
    if (leaf_aabb) put(leaf_aabb + 6 * j, b->leaf_list[j]->bounding_box);
  }
  for (int i = 0; i < n - 1; i++) {
    Node *nd = b->node_list[i];
    if (left) left[i] = child_ref(nd->left);
    if (right) right[i] = child_ref(nd->right);
    if (parent_internal) parent_internal[i] = (i == 0) ? -1 : nd->parent->idx;
    if (node_aabb) put(node_aabb + 6 * i, nd->bounding_box);
  }
  if (world_aabb) put(world_aabb, b->world[0]);
  return GPT_OK;
}

// Light-tree extras (cones + energies), for the light-sampling parity tests.
int ref_export_light_tree(RefCtx *c, float *node_cone /*5(n-1): axis,theta_0,theta_e*/,
                          float *node_energy /*3(n-1)*/, float *leaf_energy /*3n*/) {
  Bvh *b = &c->light_bvh;
  if (b->n < 2) return GPT_E_INVALID;
  for (int i = 0; i < b->n - 1; i++) {
    Node *nd = b->node_list[i];
    if (node_cone) {
      float *o = node_cone + 5 * i;
      o[0] = nd->bounding_cone->axis.x(); o[1] = nd->bounding_cone->axis.y();
      o[2] = nd->bounding_cone->axis.z(); o[3] = nd->bounding_cone->theta_0;
      o[4] = nd->bounding_cone->theta_e;
    }
    if (node_energy)
      for (int a = 0; a < 3; a++) node_energy[3 * i + a] = nd->energy[a];
  }
  if (leaf_energy)
    for (int j = 0; j < b->n; j++)
      for (int a = 0; a < 3; a++) leaf_energy[3 * j + a] = b->leaf_list[j]->energy[a];
  return GPT_OK;
}

// traverse_bvh + Triangle::hit on a ray list (src/model/bvh/bvh_traversal.h:12-84).
// nthreads <= 1 is the single-core replay BASELINE.md 3.2 asks for.
int ref_trace(RefCtx *c, int which, const float *rays, int64_t n, int32_t *leaf_id,
              int32_t *tri_id, float *t_out, float *bary, int nthreads) {
  Bvh *b = pick(c, which);
  if (!b || b->n < 2) return GPT_E_INVALID;
  std::unordered_map<Primitive *, int> leaf_of;
  leaf_of.reserve(b->n * 2);
  for (int j = 0; j < b->n; j++) leaf_of[b->geom[j]] = j;
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int64_t lo, int64_t hi) {
    for (int64_t r = lo; r < hi; r++) {
      const float *q = rays + 8 * r;
      Ray ray(vec3(q[0], q[1], q[2]), vec3(q[4], q[5], q[6]));
      hit_record rec;
      bool hit = traverse_bvh(b->node_list[0], ray, rec);
      // rec.t starts at INFINITY inside traverse_bvh; an optional finite tmax only filters
      if (hit && rec.t > q[3]) hit = false;
      int lj = -1, oi = -1;
      float tt = INFINITY, b1 = 0, b2 = 0, b3 = 0;
      if (hit) {
        lj = leaf_of[rec.object];
        oi = c->orig_index[rec.object];
        tt = rec.t;
      }
      if (leaf_id) leaf_id[r] = lj;
      if (tri_id) tri_id[r] = oi;
      if (t_out) t_out[r] = tt;
      if (bary) { bary[3 * r] = b1; bary[3 * r + 1] = b2; bary[3 * r + 2] = b3; }
      if (hit && bary) {
        // hit_record has no barycentrics; rec.point = b1*p1 + b2*p2 + b3*p3 is exported
        // instead so the CUDA path's reconstruction can be compared bitwise.
        bary[3 * r] = rec.point.x(); bary[3 * r + 1] = rec.point.y();
        bary[3 * r + 2] = rec.point.z();
      }
    }
  };
  if (nthreads <= 1) {
    work(0, n);
  } else {
    std::vector<std::thread> th;
    int64_t chunk = (n + nthreads - 1) / nthreads;
    for (int k = 0; k < nthreads; k++) {
      int64_t lo = k * chunk, hi = std::min<int64_t>(n, lo + chunk);
      if (lo < hi) th.emplace_back(work, lo, hi);
    }
    for (auto &x : th) x.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  c->last_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
  return GPT_OK;
}

double ref_last_ms(RefCtx *c) { return c->last_ms; }

// Camera::compute_ray for every pixel centre with the reference's per-pixel XORWOW stream
// (src/model/camera.h:68-83, src/util/general.h:8-16).  rays: P x 8 as gpt_trace takes.
int ref_primary_rays(RefCtx *c, float *rays) {
  if (!c->my_camera) return GPT_E_INVALID;
  int W = c->cam.width, H = c->cam.height;
  for (int i = 0; i < H; i++)
    for (int j = 0; j < W; j++) {
      int pix = i * W + j;
      curandState st;
      curand_init(1984, pix, 0, &st);
      Ray r = c->my_camera[0]->compute_ray(i + .5, j + .5, &st);
      float *q = rays + 8 * (size_t)pix;
      q[0] = r.p0.x(); q[1] = r.p0.y(); q[2] = r.p0.z(); q[3] = INFINITY;
      q[4] = r.dir.x(); q[5] = r.dir.y(); q[6] = r.dir.z(); q[7] = 0;
    }
  return GPT_OK;
}

// init_curand_state + path_tracing_render (src/main.cu:1208,1266) over a pixel window
// [row0,row1) x [col0,col1); pixels outside are left untouched in out_rgb (P x 3).
int ref_render_pathtrace_window(RefCtx *c, float *out_rgb, int row0, int row1, int col0,
                                int col1, int nthreads) {
  if (!c->my_camera || c->light_bvh.n < 2) return GPT_E_INVALID;
  int W = c->cam.width, H = c->cam.height;
  const GptParams &p = c->params;
  vec3 sky(p.sky_emission[0], p.sky_emission[1], p.sky_emission[2]);
  if (c->bg_r.empty()) { c->bg_r = {1.f}; c->bg_g = {1.f}; c->bg_b = {1.f}; }
  std::vector<vec3> fb((size_t)W * H);
  std::vector<curandState> rs((size_t)W * H);
  row1 = std::min(row1, H);
  col1 = std::min(col1, W);
  std::atomic<int> next_row(row0);
  std::atomic<unsigned long long> total_rays(0);
  g_counted_root = nullptr;
  sss_first_pass(c, nthreads);  // src/main.cu:1213-1232 (not part of the timed "Rendering")
  Object **objs = c->sss ? c->sss->my_objects : nullptr;
  Node **sss_nodes = c->sss ? c->sss->node_list : nullptr;
  g_counted_root = c->main_bvh.node_list[0];
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&] {
    shim_blockDim = dim3(1, 1, 1);
    tl_ray_count = 0;
    for (;;) {
      int i = next_row.fetch_add(1);
      if (i >= row1) { total_rays += tl_ray_count; break; }
      for (int j = col0; j < col1; j++) {
        shim_blockIdx = {(unsigned)j, (unsigned)i, 0};
        shim_threadIdx = {0, 0, 0};
        curand_init(1984, i * W + j, 0, &rs[(size_t)i * W + j]);
        path_tracing_render(fb.data(), c->my_camera, rs.data(), p.pathtracing_sample_size,
                            p.pathtracing_level, p.dof_sample_size, sky, c->bg_h, c->bg_w,
                            c->bg_r.data(), c->bg_g.data(), c->bg_b.data(),
                            c->main_bvh.node_list, objs, sss_nodes,
                            c->light_bvh.node_list, c->light_bvh.leaf_list,
                            c->light_bvh.geom, c->light_bvh.n, p.hittable_pdf_weight);
        float *o = out_rgb + 3 * ((size_t)i * W + j);
        const vec3 &v = fb[(size_t)i * W + j];
        o[0] = v.x(); o[1] = v.y(); o[2] = v.z();
      }
    }
  };
  if (nthreads <= 1) {
    work();
  } else {
    std::vector<std::thread> th;
    for (int k = 0; k < nthreads; k++) th.emplace_back(work);
    for (auto &x : th) x.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  c->last_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
  c->last_rays = total_rays.load();
  return GPT_OK;
}

unsigned long long ref_last_rays(RefCtx *c) { return c->last_rays; }

// SSS points in tree (sorted) order: object index + 12 floats (location, filter, normal,
// colour after the first pass).  *count = number of points; arrays may be NULL.
int ref_sss_export_points(RefCtx *c, int run_first_pass, int32_t *object_idx, float *rec12,
                          int64_t *count, int nthreads) {
  if (!c->sss) { *count = 0; return GPT_OK; }
  Sss *s = c->sss;
  if (run_first_pass) sss_first_pass(c, nthreads);
  *count = s->num_sss_points;
  for (int i = 0; i < s->num_sss_points; i++) {
    Point *p = s->sss_pts[i];
    if (object_idx) object_idx[i] = p->object_idx;
    if (rec12) {
      const vec3 *v[4] = {&p->location, &p->filter, &p->normal, &p->color};
      for (int a = 0; a < 4; a++)
        for (int b = 0; b < 3; b++) rec12[12 * (size_t)i + 3 * a + b] = (*v[a])[b];
    }
  }
  return GPT_OK;
}

}  // extern "C"

// ---- SPPM: the reference's own kernels, launched the way src/main.cu:1282-1870 does ---------

// rows [0,H) of a 2-D pixel kernel, spread over host threads (pixels are independent)
template <class F>
static void launch_pixels(int W, int H, int nthreads, std::atomic<unsigned long long> &rays, F f) {
  std::atomic<int> next_row(0);
  auto work = [&] {
    shim_blockDim = dim3(1, 1, 1);
    tl_ray_count = 0;
    for (;;) {
      int i = next_row.fetch_add(1);
      if (i >= H) break;
      for (int j = 0; j < W; j++) {
        shim_blockIdx = {(unsigned)j, (unsigned)i, 0};
        shim_threadIdx = {0, 0, 0};
        f();
      }
    }
    rays += tl_ray_count;
    shim_blockIdx = {0, 0, 0};
  };
  if (nthreads <= 1) { work(); return; }
  std::vector<std::thread> th;
  for (int k = 0; k < nthreads; k++) th.emplace_back(work);
  for (auto &x : th) x.join();
}

extern "C" {

int ref_sppm_init(RefCtx *c, int nthreads) {
  if (!c->my_camera || c->light_bvh.n < 2 || c->transparent_bvh.n < 2) return GPT_E_INVALID;
  const GptParams &p = c->params;
  int W = c->cam.width, H = c->cam.height, P = W * H, nph = p.ppm_num_photon_per_pass;
  Sppm *s = new Sppm();
  c->sppm = s;
  s->P = P;
  s->nphot = nph;
  s->rand_state_image = new curandState[P];
  launch_1d(P / 8 + 1, 8, [&] { init_curand_state(P, s->rand_state_image); });
  s->hit_point_list = new PPMHitPoint *[P];
  s->photon_list = new Point *[nph];
  s->surface_photon_list = new Point *[nph];
  s->volume_photon_list = new Point *[nph];
  g_counted_root = c->main_bvh.node_list[0];
  std::atomic<unsigned long long> rays(0);
  auto t0 = std::chrono::steady_clock::now();
  launch_pixels(W, H, nthreads, rays, [&] {
    ray_tracing_pass(s->hit_point_list, c->my_camera, s->rand_state_image, nullptr,
                     c->main_bvh.node_list, true, p.ppm_max_bounce, p.ppm_alpha, 0,
                     c->light_bvh.n, c->light_bvh.geom, c->light_bvh.node_list,
                     c->light_bvh.leaf_list, c->transparent_bvh.node_list,
                     p.pathtracing_sample_size, p.ppm_radius_scaling_factor);
  });
  launch_1d(1, 1, [&] { compute_average_radius(s->hit_point_list, P, &s->average_radius); });
  float avg = s->average_radius;
  launch_1d(P, 1, [&] { assign_radius_to_invalid_hit_points(s->hit_point_list, P, avg); });
  launch_1d(nph, 1, [&] { create_photon_list(s->photon_list, nph); });
  s->surface_photon_node_list = new Node *[std::max(1, nph - 1)];
  s->surface_photon_leaf_list = new Node *[nph];
  s->volume_photon_node_list = new Node *[std::max(1, nph - 1)];
  s->volume_photon_leaf_list = new Node *[nph];
  s->acc_energy = new float[c->light_bvh.n];
  launch_1d(nph, 1, [&] { init_photon_leaves(s->surface_photon_leaf_list, nph); });
  launch_1d(nph, 1, [&] { init_photon_nodes(s->surface_photon_node_list, nph); });
  launch_1d(nph, 1, [&] { init_photon_leaves(s->volume_photon_leaf_list, nph); });
  launch_1d(nph, 1, [&] { init_photon_nodes(s->volume_photon_node_list, nph); });
  launch_1d(1, 1, [&] {
    compute_accummulated_light_source_energy(c->light_bvh.geom, c->light_bvh.n, s->acc_energy);
  });
  s->rand_state_ppm = new curandState[nph];
  launch_1d(nph, 1, [&] { init_curand_state(nph, s->rand_state_ppm); });
  s->surface_morton = new unsigned int[std::max(1, nph)];
  s->volume_morton = new unsigned int[std::max(1, nph)];
  auto t1 = std::chrono::steady_clock::now();
  c->last_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
  s->rays = rays.load();
  c->last_rays = s->rays;
  return GPT_OK;
}

}  // extern "C"

// The photon tree of one kind (src/main.cu:1576-1752); spheres = the volume variant.
static void build_photon_tree(Point **list, int n, int nph, Node **node_list, Node **leaf_list,
                              unsigned int *morton, bool spheres) {
  BoundingBox **world = new BoundingBox *[1];
  launch_1d(1, 1, [&] { compute_world_bounding_box(world, list, n); });
  if (n <= 0) return;
  launch_1d(1, 1, [&] { modify_world_bounding_box(world, list, n); });
  launch_1d(n, 1, [&] { compute_photon_morton_code_batch(list, n, world); });
  std::stable_sort(list, list + n, [](Point *a, Point *b) {
    return a->bounding_box->morton_code < b->bounding_box->morton_code;
  });
  launch_1d(nph, 1, [&] { reset_photon_nodes(node_list, nph); });
  launch_1d(std::max(1, n), 1, [&] { assign_photons(leaf_list, list, n); });
  launch_1d(std::max(1, n), 1, [&] { extract_sss_morton_code_list(list, morton, n); });
  launch_1d(std::max(1, n), 1,
            [&] { set_photon_node_relationship(node_list, leaf_list, morton, n); });
  if (spheres)
    launch_1d(std::max(1, n), 1, [&] { compute_node_bounding_spheres(leaf_list, node_list, n); });
  else
    launch_1d(std::max(1, n), 1, [&] { compute_node_bounding_boxes(leaf_list, node_list, n); });
}

extern "C" {

// One pass of the loop at src/main.cu:1496-1868.
//   photon_rng 0: photon_pass is launched once PER PHOTON with rand_state = &state[photon],
//                 photon_list = &list[photon], num_photons = 1 -- the reference's own kernel,
//                 with rand_state[0] (its BSDF stream, photon_pass.h:156-160) being that
//                 photon's state instead of a state all threads race on.  Deterministic.
//   photon_rng 1: one launch over all photons, executed sequentially: every photon draws its
//                 BSDF decisions from rand_state[0] (one possible serialisation of the race).
//   stage_mask: 1 photon trace + trees, 2 eye pass, 4 gather + update
int ref_sppm_pass(RefCtx *c, int pass, int photon_rng, int stage_mask, int nthreads) {
  Sppm *s = c->sppm;
  if (!s) return GPT_E_INVALID;
  const GptParams &p = c->params;
  int W = c->cam.width, H = c->cam.height, P = s->P, nph = s->nphot;
  g_counted_root = c->main_bvh.node_list[0];
  std::atomic<unsigned long long> rays(0);
  auto t0 = std::chrono::steady_clock::now();
  if (stage_mask & 1) {
    if (photon_rng == 0) {
      std::atomic<int> next(0);
      auto work = [&] {
        shim_blockDim = dim3(1, 1, 1);
        shim_blockIdx = {0, 0, 0};
        shim_threadIdx = {0, 0, 0};
        tl_ray_count = 0;
        for (;;) {
          int k = next.fetch_add(64);
          if (k >= nph) break;
          for (int q = k; q < std::min(nph, k + 64); q++)
            photon_pass(c->light_bvh.geom, c->main_bvh.node_list, c->transparent_bvh.node_list,
                        s->photon_list + q, c->light_bvh.n, s->acc_energy, 1, p.ppm_max_bounce,
                        pass, s->rand_state_ppm + q);
        }
        rays += tl_ray_count;
      };
      if (nthreads <= 1) work();
      else {
        std::vector<std::thread> th;
        for (int k = 0; k < nthreads; k++) th.emplace_back(work);
        for (auto &x : th) x.join();
      }
    } else {
      tl_ray_count = 0;
      launch_1d(nph, 1, [&] {
        photon_pass(c->light_bvh.geom, c->main_bvh.node_list, c->transparent_bvh.node_list,
                    s->photon_list, c->light_bvh.n, s->acc_energy, nph, p.ppm_max_bounce, pass,
                    s->rand_state_ppm);
      });
      rays += tl_ray_count;
    }
    launch_1d(1, 1, [&] {
      gather_recorded_photons(s->photon_list, s->surface_photon_list, s->volume_photon_list, nph,
                              &s->num_surface, &s->num_volume);
    });
    build_photon_tree(s->surface_photon_list, s->num_surface, nph, s->surface_photon_node_list,
                      s->surface_photon_leaf_list, s->surface_morton, false);
    build_photon_tree(s->volume_photon_list, s->num_volume, nph, s->volume_photon_node_list,
                      s->volume_photon_leaf_list, s->volume_morton, true);
  }
  if (stage_mask & 2) {
    launch_pixels(W, H, nthreads, rays, [&] {
      ray_tracing_pass(s->hit_point_list, c->my_camera, s->rand_state_image,
                       s->volume_photon_node_list, c->main_bvh.node_list, false,
                       p.ppm_max_bounce, p.ppm_alpha, pass, c->light_bvh.n, c->light_bvh.geom,
                       c->light_bvh.node_list, c->light_bvh.leaf_list,
                       c->transparent_bvh.node_list, p.pathtracing_sample_size,
                       p.ppm_radius_scaling_factor);
    });
  }
  if (stage_mask & 4) {
    std::atomic<int> next(0);
    auto work = [&] {
      shim_blockDim = dim3(1, 1, 1);
      shim_threadIdx = {0, 0, 0};
      tl_ray_count = 0;
      for (;;) {
        int k = next.fetch_add(256);
        if (k >= P) break;
        for (int q = k; q < std::min(P, k + 256); q++) {
          shim_blockIdx = {(unsigned)q, 0, 0};
          update_hit_point_parameters(pass + 1, s->surface_photon_node_list,
                                      c->main_bvh.node_list, s->hit_point_list, P, nph);
        }
      }
      rays += tl_ray_count;
      shim_blockIdx = {0, 0, 0};
    };
    if (nthreads <= 1) work();
    else {
      std::vector<std::thread> th;
      for (int k = 0; k < nthreads; k++) th.emplace_back(work);
      for (auto &x : th) x.join();
    }
    launch_1d(1, 1, [&] { compute_average_radius(s->hit_point_list, P, &s->average_radius); });
  }
  auto t1 = std::chrono::steady_clock::now();
  c->last_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
  c->last_rays = rays.load();
  s->rays += c->last_rays;
  return GPT_OK;
}

int ref_sppm_counts(RefCtx *c, int64_t *n_surface, int64_t *n_volume, double *mean_radius) {
  Sppm *s = c->sppm;
  if (!s) return GPT_E_INVALID;
  *n_surface = s->num_surface;
  *n_volume = s->num_volume;
  *mean_radius = s->average_radius;
  return GPT_OK;
}

// hit-point SoA in the layout of gpt_sppm_export_hit_points (P x 4 floats each); lm = the
// tmp_accummulated_lm of the last eye pass
int ref_sppm_export_hit_points(RefCtx *c, float *loc_radius, float *normal_vradius, float *filter,
                               float *direct, float *indirect, float *lm) {
  Sppm *s = c->sppm;
  if (!s) return GPT_E_INVALID;
  for (int i = 0; i < s->P; i++) {
    PPMHitPoint *h = s->hit_point_list[i];
    auto put = [&](float *o, const vec3 &v, float w) {
      if (!o) return;
      o[4 * i] = v.x(); o[4 * i + 1] = v.y(); o[4 * i + 2] = v.z(); o[4 * i + 3] = w;
    };
    put(loc_radius, h->location, h->surface_radius);
    put(normal_vradius, h->normal, h->volume_radius);
    put(filter, h->filter, 0.f);
    put(direct, h->direct_radiance, 0.f);
    put(indirect, h->accummulated_indirect_radiance, 0.f);
    put(lm, h->tmp_accummulated_lm, 0.f);
  }
  return GPT_OK;
}

// photons of the last photon pass in photon-index order (layout of gpt_sppm_export_photons)
int ref_sppm_export_photons(RefCtx *c, int32_t *kind, float *rec12) {
  Sppm *s = c->sppm;
  if (!s) return GPT_E_INVALID;
  for (int i = 0; i < s->nphot; i++) {
    Point *p = s->photon_list[i];
    int k = p->location.vector_is_inf() ? 2 : (p->on_surface ? 0 : 1);
    if (kind) kind[i] = k;
    if (rec12) {
      float *o = rec12 + 12 * (size_t)i;
      const vec3 *v[4] = {&p->location, &p->prev_location, &p->direction, &p->color};
      for (int a = 0; a < 4; a++)
        for (int b = 0; b < 3; b++) o[3 * a + b] = (*v[a])[b];
    }
  }
  return GPT_OK;
}

// The traversal half of update_hit_point_parameters (bvh_traversal_photon.h:108-225) for every
// hit point, WITHOUT the update: flux (rgb) + count per pixel, P x 4 floats.
int ref_sppm_gather_only(RefCtx *c, float *flux_count, int nthreads) {
  Sppm *s = c->sppm;
  if (!s) return GPT_E_INVALID;
  std::atomic<int> next(0);
  auto work = [&] {
    for (;;) {
      int k = next.fetch_add(256);
      if (k >= s->P) break;
      for (int q = k; q < std::min(s->P, k + 256); q++) {
        vec3 flux(0, 0, 0);
        int cnt = 0;
        if (!s->hit_point_list[q]->location.vector_is_inf())
          _traverse_bvh_surface_photon(s->surface_photon_node_list[0], c->main_bvh.node_list[0],
                                       s->hit_point_list[q], flux, cnt);
        flux_count[4 * q] = flux.x(); flux_count[4 * q + 1] = flux.y();
        flux_count[4 * q + 2] = flux.z(); flux_count[4 * q + 3] = (float)cnt;
      }
    }
  };
  if (nthreads <= 1) work();
  else {
    std::vector<std::thread> th;
    for (int k = 0; k < nthreads; k++) th.emplace_back(work);
    for (auto &x : th) x.join();
  }
  return GPT_OK;
}

// get_ppm_image_output (image_output.h:8-31); mode 0 direct, 1 indirect, 2 global
int ref_sppm_resolve(RefCtx *c, int image_mode, int passes, float *out_rgb) {
  Sppm *s = c->sppm;
  if (!s) return GPT_E_INVALID;
  int W = c->cam.width, H = c->cam.height;
  std::vector<vec3> fb((size_t)s->P);
  std::atomic<unsigned long long> dummy(0);
  launch_pixels(W, H, 1, dummy, [&] {
    get_ppm_image_output(passes, fb.data(), s->hit_point_list, c->my_camera, image_mode);
  });
  for (int i = 0; i < s->P; i++) {
    out_rgb[3 * i] = fb[i].x(); out_rgb[3 * i + 1] = fb[i].y(); out_rgb[3 * i + 2] = fb[i].z();
  }
  return GPT_OK;
}

// create_point_image (point_operations.h:19-75) launched as src/main.cu:1248, 1370, 1534, 1545
// do, threads in index order (so the last writer of a pixel is the highest index).  which:
// 0 = hit points (colour = filter), 1 / 2 = surface / volume photons of the last photon pass in
// the order gather_recorded_photons left them (the reference splats them BEFORE the tree build
// sorts the lists), 3 = SSS points (sorted per object by then).  out_rgb: P x 3, zero = no point.
int ref_point_image(RefCtx *c, int which, float *out_rgb) {
  const int W = c->cam.width, H = c->cam.height, P = W * H;
  std::vector<vec3> fb((size_t)P, vec3(0, 0, 0));
  const unsigned tx = 8;
  if (which == 0) {
    if (!c->sppm) return GPT_E_INVALID;
    PPMHitPoint **hp = c->sppm->hit_point_list;
    launch_1d(P / tx + 1, tx, [&] { create_point_image(fb.data(), c->my_camera, hp, P); });
  } else if (which == 1 || which == 2) {
    Sppm *s = c->sppm;
    if (!s) return GPT_E_INVALID;
    std::vector<Point *> surf((size_t)std::max(1, s->nphot)), vol((size_t)std::max(1, s->nphot));
    int ns = 0, nv = 0;
    launch_1d(1, 1, [&] { gather_recorded_photons(s->photon_list, surf.data(), vol.data(), s->nphot, &ns, &nv); });
    Point **list = which == 1 ? surf.data() : vol.data();
    const int n = which == 1 ? ns : nv;
    launch_1d(n / tx + 1, tx, [&] { create_point_image(fb.data(), c->my_camera, list, n); });
  } else if (which == 3) {
    Sss *s = c->sss;
    if (!s) return GPT_E_INVALID;
    launch_1d(s->num_sss_points / tx + 1, tx,
              [&] { create_point_image(fb.data(), c->my_camera, s->sss_pts, s->num_sss_points); });
  } else {
    return GPT_E_INVALID;
  }
  for (int i = 0; i < P; i++) {
    out_rgb[3 * i] = fb[i].x(); out_rgb[3 * i + 1] = fb[i].y(); out_rgb[3 * i + 2] = fb[i].z();
  }
  return GPT_OK;
}

}  // extern "C"
