/*
 * examples/main.cpp -- the reference's main() with its device half replaced by the C-ABI.
 *
 * This is the host program INTEGRATION.md describes, complete and compiled: it keeps the
 * reference's OWN host code -- InputParam (src/input/input_param.h), the OBJ / MTL / image
 * loaders (src/input/read_file_util.h, read_image_util.h) and save_image
 * (src/util/image_util.h:50-72) -- by #including those headers unchanged, runs the loader calls
 * of src/main.cu:165-471 in their order, and then, where src/main.cu:474-1870 launches ~60
 * kernels, calls include/gpt.h.  Output files have the reference's names (src/main.cu:1236-1262,
 * 1830-1875).  Nothing of the reference is copied into this repository: examples/Makefile
 * compiles against the tree at $(REF) (through a scratch copy in which only the libjpeg wrapper
 * is replaced by oracle/jpeg_stub/jpeg.h, because this image has no <jpeglib.h>).
 *
 *   gpt_main MASTER_FILE OUTPUT_PREFIX [--rank R --nranks N --id-file PATH] [--device D]
 *
 * One process per GPU.  With --nranks > 1 rank 0 writes the NCCL unique id to --id-file and the
 * other ranks read it (any launcher that can share a file works; MPI_Bcast would do the same),
 * gpt_comm_init binds the library's communicator, the frame is tile-sharded
 * (gpt_render_pathtrace_device(rank, nranks) + gpt_gather_frame) or, in SPPM, the photons are
 * split by index range with gpt_sppm_allreduce between the two halves of
 * update_hit_point_parameters; rank 0 writes the images.
 */
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <iterator>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "input/input_param.h"
#include "input/read_file_util.h"
#include "input/read_image_util.h"
#include "util/image_util.h"

#include "gpt.h"

namespace {

struct Args {
  std::string master, out;
  int rank = 0, nranks = 1, device = -1;
  std::string id_file;
};

bool parse_args(int argc, char **argv, Args &a) {
  std::vector<std::string> pos;
  for (int i = 1; i < argc; i++) {
    std::string s = argv[i];
    auto next = [&]() -> const char * { return i + 1 < argc ? argv[++i] : ""; };
    if (s == "--rank") a.rank = atoi(next());
    else if (s == "--nranks") a.nranks = atoi(next());
    else if (s == "--id-file") a.id_file = next();
    else if (s == "--device") a.device = atoi(next());
    else pos.push_back(s);
  }
  if (pos.size() != 2 || a.nranks < 1 || a.rank < 0 || a.rank >= a.nranks) return false;
  if (a.nranks > 1 && a.id_file.empty()) return false;
  a.master = pos[0];
  a.out = pos[1];
  if (a.device < 0) a.device = a.rank;
  return true;
}

int fail(gpt_ctx *ctx, const char *what) {
  fprintf(stderr, "gpt_main: %s failed: %s\n", what, ctx ? gpt_last_error(ctx) : "(no context)");
  return 99;  // the reference's exit code for every CUDA error (src/main.cu:47-59)
}
#define CALL(expr) do { if ((expr) != GPT_OK) return fail(ctx, #expr); } while (0)

// 128 bytes of ncclUniqueId from rank 0 to everybody through a file (written under a
// temporary name and renamed, so a reader never sees half of it)
bool share_unique_id(const Args &a, unsigned char id[128]) {
  if (a.rank == 0) {
    if (gpt_comm_unique_id(id) != GPT_OK) return false;
    const std::string tmp = a.id_file + ".tmp";
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f || fwrite(id, 1, 128, f) != 128) return false;
    fclose(f);
    return rename(tmp.c_str(), a.id_file.c_str()) == 0;
  }
  for (int tries = 0; tries < 6000; tries++) {  // up to 60 s
    FILE *f = fopen(a.id_file.c_str(), "rb");
    if (f) {
      const size_t n = fread(id, 1, 128, f);
      fclose(f);
      if (n == 128) return true;
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(10));
  }
  return false;
}

struct HostScene {  // what src/main.cu:139-156,386-450 keeps in cudaMallocManaged arrays
  std::vector<float> bg[3], tex[3], mf[23];
  std::vector<int> mi[16];
  std::vector<float> x, y, z, xn, yn, zn, xt, yt;
  std::vector<int> idx[11], obj_num, obj_off;
  int bg_h = 0, bg_w = 0, n_materials = 0, n_triangles = 0;
  int n_objects = 0, n_v = 0, n_vt = 0, n_vn = 0, n_faces = 0;
  long int texture_length = 0;
};

// src/main.cu:165-471, the reference's own functions in the reference's order
void load_with_the_reference_loaders(InputParam &in, HostScene &s) {
  const std::string folder = in.input_folder_path, obj = in.obj_filename;
  std::vector<std::string> mtl_files, texture_names, material_names;
  std::vector<int> tex_off, tex_h, tex_w;

  extract_single_image_requirement(folder, in.texture_bg_path, s.bg_h, s.bg_w);
  for (auto &c : s.bg) c.assign((size_t)s.bg_h * s.bg_w, 0.f);
  int cursor = 0;
  extract_single_image(folder, in.texture_bg_path, s.bg[0].data(), s.bg[1].data(), s.bg[2].data(), cursor);

  extract_material_file_names(folder, obj, mtl_files);
  extract_image_resource_requirement(folder, mtl_files, texture_names, tex_off, tex_h, tex_w,
                                     s.texture_length);
  for (auto &c : s.tex) c.assign((size_t)std::max<long int>(1, s.texture_length), 0.f);
  extract_textures(folder, texture_names, s.tex[0].data(), s.tex[1].data(), s.tex[2].data());
  extract_num_elements(folder, obj, s.n_objects, s.n_v, s.n_vt, s.n_vn, s.n_faces);

  const int max_materials = 100;  // src/main.cu:130
  for (auto &v : s.mf) v.assign(max_materials, 0.f);
  for (auto &v : s.mi) v.assign(max_materials, 0);
  float *f[23];
  int *m[16];
  for (int i = 0; i < 23; i++) f[i] = s.mf[i].data();
  for (int i = 0; i < 16; i++) m[i] = s.mi[i].data();
  extract_material_data(folder, mtl_files, texture_names, tex_off, tex_h, tex_w,
                        f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8], f[9], f[10], f[11],
                        f[12], f[13], f[14], f[15], f[16], f[17], f[18], f[19], f[20], f[21], f[22],
                        m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10], m[11],
                        m[12], m[13], m[14], m[15], &s.n_materials, material_names);

  auto floats = [](std::vector<float> &v, int n) { v.assign((size_t)std::max(1, n), 0.f); };
  floats(s.x, s.n_v); floats(s.y, s.n_v); floats(s.z, s.n_v);
  floats(s.xn, s.n_vn); floats(s.yn, s.n_vn); floats(s.zn, s.n_vn);
  // extract_triangle_data writes x_tex[norm_idx] for a file without vt (read_file_util.h:646-650):
  // room for that stray write, as the managed allocation of the reference happens to have
  s.xt.assign((size_t)std::max(1, s.n_vt) + (size_t)std::max(1, s.n_vn) + 1, 0.f);
  s.yt = s.xt;
  for (auto &v : s.idx) v.assign((size_t)std::max(1, s.n_faces), 0);
  s.obj_num.assign((size_t)std::max(1, s.n_objects), 0);
  s.obj_off.assign((size_t)std::max(1, s.n_objects), 0);
  extract_triangle_data(folder, obj, s.x.data(), s.y.data(), s.z.data(), s.xn.data(), s.yn.data(),
                        s.zn.data(), s.xt.data(), s.yt.data(), s.idx[0].data(), s.idx[1].data(),
                        s.idx[2].data(), s.idx[3].data(), s.idx[4].data(), s.idx[5].data(),
                        s.idx[6].data(), s.idx[7].data(), s.idx[8].data(), material_names,
                        s.idx[9].data(), &s.n_triangles, &s.n_materials, s.idx[10].data(),
                        s.obj_num.data(), s.obj_off.data());
}

int hand_over(gpt_ctx *ctx, const InputParam &in, const HostScene &s) {
  GptMesh mesh;
  memset(&mesh, 0, sizeof mesh);
  mesh.num_vertices = std::max(1, s.n_v);
  mesh.num_normals = std::max(1, s.n_vn);
  mesh.num_texcoords = std::max(1, s.n_vt);
  mesh.num_triangles = s.n_triangles;
  mesh.num_objects = s.n_objects;
  mesh.x = s.x.data(); mesh.y = s.y.data(); mesh.z = s.z.data();
  mesh.x_norm = s.xn.data(); mesh.y_norm = s.yn.data(); mesh.z_norm = s.zn.data();
  mesh.x_tex = s.xt.data(); mesh.y_tex = s.yt.data();
  const int32_t **ix = &mesh.point_1_idx;  // point_1..3, norm_1..3, tex_1..3, material, object
  for (int i = 0; i < 11; i++) ix[i] = s.idx[i].data();
  mesh.object_num_primitives = s.obj_num.data();
  mesh.object_primitive_offset_idx = s.obj_off.data();

  GptMaterials mats;
  memset(&mats, 0, sizeof mats);
  mats.num_materials = s.n_materials;
  const float **fp = &mats.ka_x;  // 23 float arrays in extract_material_data's order
  for (int i = 0; i < 23; i++) fp[i] = s.mf[i].data();
  const int32_t **ip = &mats.material_priority;  // priority, then 5 x (height, width, offset)
  for (int i = 0; i < 16; i++) ip[i] = s.mi[i].data();

  GptCamera cam = {{in.eye_x, in.eye_y, in.eye_z}, {in.center_x, in.center_y, in.center_z},
                   {in.up_x, in.up_y, in.up_z}, in.fovy, in.image_width, in.image_height,
                   in.aperture, in.focus_dist};
  GptParams par;
  memset(&par, 0, sizeof par);
  par.render_mode = in.render_mode;
  par.pathtracing_sample_size = in.pathtracing_sample_size;
  par.pathtracing_level = in.pathtracing_level;
  par.dof_sample_size = in.dof_sample_size;
  par.sss_pts_per_object = in.sss_pts_per_object;
  par.hittable_pdf_weight = in.hittable_pdf_weight;
  par.sky_emission[0] = in.sky_emission_r;
  par.sky_emission[1] = in.sky_emission_g;
  par.sky_emission[2] = in.sky_emission_b;
  par.ppm_num_photon_per_pass = in.ppm_num_photon_per_pass;
  par.ppm_num_pass = in.ppm_num_pass;
  par.ppm_max_bounce = in.ppm_max_bounce;
  par.ppm_alpha = in.ppm_alpha;
  par.ppm_radius_scaling_factor = in.ppm_radius_scaling_factor;
  par.rng_mode = 0;  // the reference's streams

  CALL(gpt_set_mesh(ctx, &mesh));
  CALL(gpt_set_materials(ctx, &mats));
  CALL(gpt_set_textures(ctx, s.tex[0].data(), s.tex[1].data(), s.tex[2].data(), s.texture_length));
  CALL(gpt_set_background(ctx, s.bg[0].data(), s.bg[1].data(), s.bg[2].data(), s.bg_h, s.bg_w));
  CALL(gpt_set_camera(ctx, &cam));
  CALL(gpt_set_params(ctx, &par));
  return GPT_OK;
}

}  // namespace

int main(int argc, char **argv) {
  Args a;
  if (!parse_args(argc, argv, a)) {
    fprintf(stderr, "usage: %s MASTER_FILE OUTPUT_PREFIX [--rank R --nranks N --id-file PATH] [--device D]\n", argv[0]);
    return 2;
  }
  InputParam in;
  in.aperture = 0;  // InputParam leaves its scalars uninitialised (input_param.h:11-33)
  in.dof_sample_size = 1;
  HostScene scene;
  try {
    in.extract_parameters(a.master);
    load_with_the_reference_loaders(in, scene);
  } catch (std::exception &e) {
    fprintf(stderr, "gpt_main: loading %s: %s\n", a.master.c_str(), e.what());
    return 99;
  }
  const int W = in.image_width, H = in.image_height;
  const std::string out = a.out;

  gpt_ctx *ctx = nullptr;
  if (gpt_create(a.device, &ctx) != GPT_OK) return fail(ctx, "gpt_create");
  if (a.nranks > 1) {
    unsigned char id[128];
    if (!share_unique_id(a, id)) { fprintf(stderr, "gpt_main: could not share the NCCL id\n"); return 99; }
    CALL(gpt_comm_init(ctx, id, a.rank, a.nranks));
  }
  if (hand_over(ctx, in, scene) != GPT_OK) return 99;
  CALL(gpt_build(ctx));

  std::vector<vec3> image((size_t)W * H);  // vec3 = three floats (src/model/vector_and_matrix/vec3.h)
  float *rgb = reinterpret_cast<float *>(image.data());
  const bool writer = a.rank == 0;

  if (in.render_mode == 1) {  // src/main.cu:1211-1280
    if (in.sss_pts_per_object > 0 && writer) {
      CALL(gpt_sss_debug_image(ctx, rgb));
      save_image(image.data(), W, H, out + "_pts.ppm");
    }
    if (a.nranks == 1) {
      CALL(gpt_render_pathtrace(ctx, rgb));
    } else {
      void *d_frame = nullptr;
      const size_t bytes = (size_t)W * H * 3 * sizeof(float);
      cudaSetDevice(a.device);
      if (cudaMalloc(&d_frame, bytes) != cudaSuccess || cudaMemset(d_frame, 0, bytes) != cudaSuccess) {
        fprintf(stderr, "gpt_main: cudaMalloc of the frame failed\n");
        return 99;
      }
      CALL(gpt_render_pathtrace_device(ctx, d_frame, a.rank, a.nranks));  // this rank's 32x32 tiles
      CALL(gpt_gather_frame(ctx, d_frame, 0));                            // tiles are disjoint: a sum
      CALL(gpt_synchronize(ctx));
      if (writer) cudaMemcpy(rgb, d_frame, bytes, cudaMemcpyDeviceToHost);
      cudaFree(d_frame);
    }
  } else if (in.render_mode == 2) {  // src/main.cu:1282-1870
    CALL(gpt_sppm_init(ctx));
    const int64_t photons = in.ppm_num_photon_per_pass;
    const int64_t begin = photons * a.rank / a.nranks, end = photons * (a.rank + 1) / a.nranks;
    for (int i = 0; i < in.ppm_num_pass; i++) {
      CALL(gpt_sppm_pass(ctx, i, begin, end));
      CALL(gpt_sppm_allreduce(ctx));  // a no-op without a communicator
      CALL(gpt_sppm_update(ctx, i));
      if (writer && i % std::max(1, (int)in.ppm_image_output_iteration) == 0) {  // src/main.cu:1830-1867
        CALL(gpt_sppm_debug_image(ctx, 1, rgb));  // surface photons of this pass
        save_image(image.data(), W, H, out + "_surface_photon.ppm");
        CALL(gpt_sppm_debug_image(ctx, 2, rgb));  // volume photons
        save_image(image.data(), W, H, out + "_volume_photon.ppm");
        CALL(gpt_sppm_resolve(ctx, 1, i + 1, rgb));
        save_image(image.data(), W, H, out + "_direct.ppm");
        CALL(gpt_sppm_resolve(ctx, 2, i + 1, rgb));
        save_image(image.data(), W, H, out + "_indirect.ppm");
        CALL(gpt_sppm_resolve(ctx, 0, i + 1, rgb));
        save_image(image.data(), W, H, out + "_global.ppm");
      }
    }
    CALL(gpt_sppm_resolve(ctx, 0, in.ppm_num_pass, rgb));
  } else {
    fprintf(stderr, "gpt_main: render_mode %d is neither 1 (path tracing) nor 2 (SPPM)\n", in.render_mode);
    return 99;
  }
  if (writer) save_image(image.data(), W, H, out + ".ppm");  // src/main.cu:1872-1875

  GptStats st;
  if (gpt_stats(ctx, &st) == GPT_OK && writer)
    printf("gpt_main: %d triangles, build %.3f ms (all %.3f ms), render %.3f ms, %llu rays\n",
           scene.n_triangles, st.build_ms, st.build_all_ms, st.render_ms,
           (unsigned long long)st.rays_traced);
  gpt_destroy(ctx);
  return 0;
}
