/*
 * ref_loader.cpp -- TEST INFRASTRUCTURE ONLY.  The reference's OWN host loaders, run as they are.
 *
 * Builds oracle/_ref/libref_loader.so from the reference's headers: input/input_param.h,
 * input/read_file_util.h, input/read_image_util.h (and what they include).  read_file_util.h
 * includes external/libjpeg_cpp/jpeg.h by a relative path, and that wrapper needs <jpeglib.h>,
 * which this image does not have; so, exactly like the `refcuda` target, the Makefile compiles
 * against a scratch copy of the reference tree under /tmp in which ONLY that one header is
 * replaced by oracle/jpeg_stub/jpeg.h (raw "GPTIMG" images instead of JPEG).  Nothing of the
 * reference enters the repository.
 *
 * ref_scene_files_parse() performs the call sequence of src/main.cu:165-471 -- same functions,
 * same order, same array sizes -- and hands the arrays back through the GptSceneFilesView of
 * include/gpt.h, so tests/test_loader.py can compare them field by field with what
 * gpt_scene_files_parse (gpu_pathtracer_b200/csrc/loader.cu) produces from the same files.
 */
#include <algorithm>
#include <fstream>
#include <iostream>
#include <iterator>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "ref_shim.h"

#include "input/input_param.h"
#include "input/read_file_util.h"
#include "input/read_image_util.h"

#include <chrono>
#include <cstring>

#include "../include/gpt.h"

namespace {

struct RefScene {
  InputParam ip;
  std::vector<std::string> material_file, material_name, texture_name;
  std::vector<int> texture_offset, texture_height, texture_width;
  std::vector<float> tex[3], bg[3];
  int bg_h = 0, bg_w = 0;
  std::vector<float> mf[23];
  std::vector<int> mi[16];
  int num_materials = 0;
  std::vector<float> x, y, z, xn, yn, zn, xt, yt;
  std::vector<int> p1, p2, p3, n1, n2, n3, t1, t2, t3, mat, obj, obj_num, obj_off;
  int num_triangles = 0;
  double parse_ms = 0, obj_ms = 0;
};

}  // namespace

extern "C" {

int ref_scene_files_parse(const char *master, void **out, char *err, int err_len) {
  RefScene *s = new RefScene();
  try {
    auto t0 = std::chrono::steady_clock::now();
    s->ip.aperture = 0;  // the class leaves its scalars uninitialised (input_param.h:11-33)
    s->ip.dof_sample_size = 1;
    s->ip.extract_parameters(master);
    const std::string folder = s->ip.input_folder_path, objf = s->ip.obj_filename;
    // main.cu:165-186 background
    extract_single_image_requirement(folder, s->ip.texture_bg_path, s->bg_h, s->bg_w);
    for (auto &c : s->bg) c.assign((size_t)s->bg_h * s->bg_w, 0.f);
    int next_idx = 0;
    extract_single_image(folder, s->ip.texture_bg_path, s->bg[0].data(), s->bg[1].data(),
                         s->bg[2].data(), next_idx);
    auto t1 = std::chrono::steady_clock::now();
    // main.cu:192-199
    extract_material_file_names(folder, objf, s->material_file);
    // main.cu:201-243 texture atlas
    long int texture_length = 0;
    extract_image_resource_requirement(folder, s->material_file, s->texture_name, s->texture_offset,
                                       s->texture_height, s->texture_width, texture_length);
    for (auto &c : s->tex) c.assign((size_t)texture_length, 0.f);
    extract_textures(folder, s->texture_name, s->tex[0].data(), s->tex[1].data(), s->tex[2].data());
    // main.cu:245-254
    int num_objects, num_vertices, num_vt, num_vn, num_faces;
    extract_num_elements(folder, objf, num_objects, num_vertices, num_vt, num_vn, num_faces);
    // main.cu:256-383 materials (max_num_materials = 100, main.cu:130)
    for (auto &v : s->mf) v.assign(100, 0.f);
    for (auto &v : s->mi) v.assign(100, 0);
    s->num_materials = 0;
    float **f = nullptr;
    (void)f;
    auto F = [&](int i) { return s->mf[i].data(); };
    auto I = [&](int i) { return s->mi[i].data(); };
    // argument order of extract_material_data: ka kd ks ke tf path_length t_r n_s n_i bm
    // scattering absorption g priority, then (height, width, offset) of diffuse, specular,
    // emission, n_s, bump
    extract_material_data(folder, s->material_file, s->texture_name, s->texture_offset,
                          s->texture_height, s->texture_width,
                          F(0), F(1), F(2), F(3), F(4), F(5), F(6), F(7), F(8), F(9), F(10), F(11),
                          F(12), F(13), F(14), F(15), F(16), F(17), F(18), F(19), F(20), F(21), F(22),
                          I(0), I(1), I(2), I(3), I(4), I(5), I(6), I(7), I(8), I(9), I(10), I(11),
                          I(12), I(13), I(14), I(15), &s->num_materials, s->material_name);
    // main.cu:386-471 mesh (array sizes as allocated there)
    auto fz = [](std::vector<float> &v, int n) { v.assign((size_t)std::max(1, n), 0.f); };
    fz(s->x, num_vertices); fz(s->y, num_vertices); fz(s->z, num_vertices);
    // +1: extract_triangle_data writes x_tex[norm_idx] when the file has no vt
    // (read_file_util.h:646-650), one past the reference's own allocation when vn exist
    fz(s->xn, num_vn); fz(s->yn, num_vn); fz(s->zn, num_vn);
    s->xt.assign((size_t)std::max(1, num_vt) + (size_t)std::max(1, num_vn) + 1, 0.f);
    s->yt = s->xt;
    for (auto *v : {&s->p1, &s->p2, &s->p3, &s->n1, &s->n2, &s->n3, &s->t1, &s->t2, &s->t3, &s->mat, &s->obj})
      v->assign((size_t)num_faces, 0);
    s->obj_num.assign((size_t)num_objects, 0);
    s->obj_off.assign((size_t)num_objects, 0);
    extract_triangle_data(folder, objf, s->x.data(), s->y.data(), s->z.data(), s->xn.data(),
                          s->yn.data(), s->zn.data(), s->xt.data(), s->yt.data(), s->p1.data(),
                          s->p2.data(), s->p3.data(), s->n1.data(), s->n2.data(), s->n3.data(),
                          s->t1.data(), s->t2.data(), s->t3.data(), s->material_name, s->mat.data(),
                          &s->num_triangles, &s->num_materials, s->obj.data(), s->obj_num.data(),
                          s->obj_off.data());
    s->xt.resize((size_t)std::max(1, num_vt));
    s->yt.resize(s->xt.size());
    auto t2 = std::chrono::steady_clock::now();
    s->parse_ms = std::chrono::duration<double, std::milli>(t2 - t0).count();
    s->obj_ms = std::chrono::duration<double, std::milli>(t2 - t1).count();
  } catch (std::exception &e) {
    if (err && err_len > 0) snprintf(err, (size_t)err_len, "%s", e.what());
    delete s;
    return -1;
  }
  *out = s;
  return 0;
}

void ref_scene_files_free(void *p) { delete static_cast<RefScene *>(p); }

int ref_scene_files_view(void *p, GptSceneFilesView *v) {
  RefScene *s = static_cast<RefScene *>(p);
  memset(v, 0, sizeof *v);
  GptMesh &m = v->mesh;
  m.num_vertices = (int)s->x.size();
  m.num_normals = (int)s->xn.size();
  m.num_texcoords = (int)s->xt.size();
  m.num_triangles = s->num_triangles;
  m.num_objects = (int)s->obj_num.size();
  m.x = s->x.data(); m.y = s->y.data(); m.z = s->z.data();
  m.x_norm = s->xn.data(); m.y_norm = s->yn.data(); m.z_norm = s->zn.data();
  m.x_tex = s->xt.data(); m.y_tex = s->yt.data();
  m.point_1_idx = s->p1.data(); m.point_2_idx = s->p2.data(); m.point_3_idx = s->p3.data();
  m.norm_1_idx = s->n1.data(); m.norm_2_idx = s->n2.data(); m.norm_3_idx = s->n3.data();
  m.tex_1_idx = s->t1.data(); m.tex_2_idx = s->t2.data(); m.tex_3_idx = s->t3.data();
  m.material_idx = s->mat.data();
  m.triangle_object_idx = s->obj.data();
  m.object_num_primitives = s->obj_num.data();
  m.object_primitive_offset_idx = s->obj_off.data();
  GptMaterials &q = v->materials;
  q.num_materials = s->num_materials;
  const float **fp = &q.ka_x;
  for (int i = 0; i < 23; i++) fp[i] = s->mf[i].data();
  const int32_t **ipp = &q.material_priority;
  for (int i = 0; i < 16; i++) ipp[i] = s->mi[i].data();
  const InputParam &ip = s->ip;
  GptCamera &c = v->camera;
  c.eye[0] = ip.eye_x; c.eye[1] = ip.eye_y; c.eye[2] = ip.eye_z;
  c.center[0] = ip.center_x; c.center[1] = ip.center_y; c.center[2] = ip.center_z;
  c.up[0] = ip.up_x; c.up[1] = ip.up_y; c.up[2] = ip.up_z;
  c.fovy = ip.fovy; c.width = ip.image_width; c.height = ip.image_height;
  c.aperture = ip.aperture; c.focus_dist = ip.focus_dist;
  GptParams &r = v->params;
  r.render_mode = ip.render_mode;
  r.pathtracing_sample_size = ip.pathtracing_sample_size;
  r.pathtracing_level = ip.pathtracing_level;
  r.dof_sample_size = ip.dof_sample_size;
  r.sss_pts_per_object = ip.sss_pts_per_object;
  r.hittable_pdf_weight = ip.hittable_pdf_weight;
  r.sky_emission[0] = ip.sky_emission_r; r.sky_emission[1] = ip.sky_emission_g; r.sky_emission[2] = ip.sky_emission_b;
  r.ppm_num_photon_per_pass = ip.ppm_num_photon_per_pass;
  r.ppm_num_pass = ip.ppm_num_pass;
  r.ppm_max_bounce = ip.ppm_max_bounce;
  r.ppm_alpha = ip.ppm_alpha;
  r.ppm_radius_scaling_factor = ip.ppm_radius_scaling_factor;
  v->tex_r = s->tex[0].data(); v->tex_g = s->tex[1].data(); v->tex_b = s->tex[2].data();
  v->texture_length = (int64_t)s->tex[0].size();
  v->bg_r = s->bg[0].data(); v->bg_g = s->bg[1].data(); v->bg_b = s->bg[2].data();
  v->bg_height = s->bg_h; v->bg_width = s->bg_w;
  v->num_textures = (int)s->texture_name.size();
  v->texture_offset = s->texture_offset.data();
  v->texture_height = s->texture_height.data();
  v->texture_width = s->texture_width.data();
  v->input_folder_path = ip.input_folder_path.c_str();
  v->obj_filename = ip.obj_filename.c_str();
  v->texture_bg_path = ip.texture_bg_path.c_str();
  v->image_output_path = ip.image_output_path.c_str();
  v->ppm_image_output_iteration = (int)ip.ppm_image_output_iteration;
  v->parse_ms = s->parse_ms;
  v->obj_parse_ms = s->obj_ms;
  v->threads = 1;
  return 0;
}

const char *ref_scene_files_material_name(void *p, int i) {
  RefScene *s = static_cast<RefScene *>(p);
  return (i >= 0 && (size_t)i < s->material_name.size()) ? s->material_name[(size_t)i].c_str() : nullptr;
}

const char *ref_scene_files_texture_name(void *p, int i) {
  RefScene *s = static_cast<RefScene *>(p);
  return (i >= 0 && (size_t)i < s->texture_name.size()) ? s->texture_name[(size_t)i].c_str() : nullptr;
}

}  // extern "C"
