/*
 * gpt.h -- C-ABI of the B200-native hot path of gpu_pathtracer.
 *
 * The reference (aloysius-aryaputera/gpu_pathtracer) has no plugin or FFI seam: its host
 * loaders fill flat SoA arrays (src/main.cu:165-471) and everything after that is device
 * code launched from main() (src/main.cu:474-1870).  This header IS that seam: each entry
 * point below replaces the launch sequence it cites, takes the same flat arrays with the
 * same names and meaning, and returns an int status instead of exit(99)
 * (src/main.cu:47-59).
 *
 * Conventions
 *   - every pointer in an input struct is a caller-owned HOST pointer unless the function
 *     name ends in _device; the library copies what it needs into its own HBM buffers.
 *   - return value: GPT_OK (0) or a negative GPT_E_* code; gpt_last_error() gives text.
 *   - no exceptions, no exit(), no torch types cross this boundary.
 *   - all floating point is IEEE fp32, parity-critical stages are evaluated without FMA
 *     contraction so they are bit-identical to the reference's host replay.
 */
#ifndef GPT_H
#define GPT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPT_OK 0
#define GPT_E_INVALID (-1)   /* bad argument / call order                                  */
#define GPT_E_CUDA (-2)      /* a CUDA runtime call failed (text in gpt_last_error)        */
#define GPT_E_NODEVICE (-3)  /* no CUDA device: there is NO CPU fallback                   */
#define GPT_E_UNSUPPORTED (-4)
#define GPT_E_NCCL (-5)      /* an NCCL call failed, or libnccl.so.2 could not be loaded         */

/* which acceleration structure (the reference builds three: src/main.cu:826-1197) */
#define GPT_BVH_MAIN 0         /* all triangles          src/main.cu:826-918   */
#define GPT_BVH_TRANSPARENT 1  /* t_r > 0 triangles      src/main.cu:930-1063  */
#define GPT_BVH_LIGHT 2        /* emissive triangles     src/main.cu:1065-1197 */

/* per-triangle flag bits exported by gpt_export_triangles */
#define GPT_TRI_LIGHT 1        /* Triangle::_check_if_light_emitting  triangle.h:123-135 */
#define GPT_TRI_TRANSPARENT 2  /* Material::is_transparent            material.h:532-534 */
#define GPT_TRI_SSS 4          /* material path_length > 0            material.h:497     */

typedef struct gpt_ctx gpt_ctx;

/* Geometry SoA exactly as produced by extract_triangle_data (src/main.cu:386-471,
 * src/input/read_file_util.h:96-101).  Indices are 0-based; 0 where the OBJ had none. */
typedef struct GptMesh {
  int32_t num_vertices, num_normals, num_texcoords, num_triangles, num_objects;
  const float *x, *y, *z;                        /* [num_vertices]  */
  const float *x_norm, *y_norm, *z_norm;         /* [num_normals]   */
  const float *x_tex, *y_tex;                    /* [num_texcoords] */
  const int32_t *point_1_idx, *point_2_idx, *point_3_idx; /* [num_triangles] */
  const int32_t *norm_1_idx, *norm_2_idx, *norm_3_idx;
  const int32_t *tex_1_idx, *tex_2_idx, *tex_3_idx;
  const int32_t *material_idx;                   /* [num_triangles] */
  const int32_t *triangle_object_idx;            /* [num_triangles] */
  const int32_t *object_num_primitives;          /* [num_objects]   */
  const int32_t *object_primitive_offset_idx;    /* [num_objects]   */
} GptMesh;

/* Material SoA exactly as produced by extract_material_data (src/main.cu:139-156,355-383)
 * and consumed by create_material (src/lib/world.h:101-175).  Texture (height, width,
 * offset) triples index the planar atlas given to gpt_set_textures. */
typedef struct GptMaterials {
  int32_t num_materials;                         /* <= 100 (src/main.cu:130) */
  const float *ka_x, *ka_y, *ka_z;
  const float *kd_x, *kd_y, *kd_z;
  const float *ks_x, *ks_y, *ks_z;
  const float *ke_x, *ke_y, *ke_z;
  const float *tf_x, *tf_y, *tf_z;
  const float *path_length, *t_r, *n_s, *n_i, *bm;
  const float *scattering_coef, *absorption_coef, *g;
  const int32_t *material_priority;
  const int32_t *height_diffuse, *width_diffuse, *offset_diffuse;
  const int32_t *height_specular, *width_specular, *offset_specular;
  const int32_t *height_emission, *width_emission, *offset_emission;
  const int32_t *height_n_s, *width_n_s, *offset_n_s;
  const int32_t *height_bump, *width_bump, *offset_bump;
} GptMaterials;

/* Camera constructor arguments (src/lib/world.h:85-99, src/model/camera.h:30-52). */
typedef struct GptCamera {
  float eye[3], center[3], up[3];
  float fovy;
  int32_t width, height;
  float aperture, focus_dist;
} GptCamera;

/* Mirror of InputParam's render scalars (src/input/input_param.h:11-33). */
typedef struct GptParams {
  int32_t render_mode;              /* 1 = path tracing, 2 = SPPM */
  int32_t pathtracing_sample_size;
  int32_t pathtracing_level;
  int32_t dof_sample_size;
  int32_t sss_pts_per_object;
  float hittable_pdf_weight;
  float sky_emission[3];
  int32_t ppm_num_photon_per_pass;
  int32_t ppm_num_pass;
  int32_t ppm_max_bounce;
  float ppm_alpha;
  float ppm_radius_scaling_factor;
  int32_t rng_mode;                 /* 0 = reference streams (seed 1984, seq = pixel): a pixel's
                                       samples run one after the other on its one stream
                                       (src/render/pathtracing.h:277-293, src/util/general.h:8-16);
                                       1 = fast: the samples of a pixel are dealt to L independent
                                       streams (seq = pixel * L + lane, gpt_set_fast_rng_lanes)
                                       that run side by side -- same estimator, other random
                                       numbers, a chain of dependent rounds L times shorter */
} GptParams;

typedef struct GptStats {
  double build_ms;        /* last gpt_build, main BVH only (src/main.cu:826-918)           */
  double build_all_ms;    /* preprocessing + all three BVHs                                */
  double trace_ms;        /* last gpt_trace* kernel time (CUDA events, device only)        */
  double render_ms;       /* last gpt_render_pathtrace, device only                        */
  uint64_t rays_traced;   /* closest-hit traversals of the MAIN BVH in the last call       */
  uint64_t internal_visits; /* V_int: internal nodes whose two child boxes were tested     */
  uint64_t leaf_tests;      /* V_leaf: triangles intersected                               */
  uint64_t kernel_launches; /* kernels of this library launched by the last call           */
  uint64_t shade_events;    /* path vertices shaded in the last render                     */
  double trace_kernel_ms;   /* sum of the trace-stage kernels of the last render (profiling) */
  double shade_kernel_ms;   /* sum of the shade-stage kernels of the last render (profiling) */
  uint64_t trace_launches;  /* trace-stage launches in the last render                       */
  double build_wide_ms;     /* last gpt_build: the 4-wide tree collapsed from the main LBVH  */
  uint64_t wide_node_visits; /* the share of internal_visits that were 48-byte nodes of the 4-wide tree */
  uint64_t wide_leaf_tests;  /* the share of leaf_tests made by the wide walk                            */
} GptStats;

/* ---- lifetime ---------------------------------------------------------------------- */
int gpt_create(int device, gpt_ctx **out);
void gpt_destroy(gpt_ctx *ctx);
const char *gpt_last_error(const gpt_ctx *ctx);
const char *gpt_version(void);
/* Launch on the caller's CUDA stream (a cudaStream_t passed as an integer); 0 = the
 * library's own stream.  Lets a host time the kernels with events on that stream. */
int gpt_set_stream(gpt_ctx *ctx, uint64_t cuda_stream);

/* ---- scene upload: replaces src/main.cu:474-640 (create_camera, create_material,
 *      create_objects, create_world, tangent accumulation) ------------------------------ */
int gpt_set_mesh(gpt_ctx *ctx, const GptMesh *mesh);
int gpt_set_materials(gpt_ctx *ctx, const GptMaterials *mats);
/* planar atlas, offset 0 = the 1x1 white default (src/input/read_image_util.h:135-144) */
int gpt_set_textures(gpt_ctx *ctx, const float *r, const float *g, const float *b,
                     int64_t texture_length);
int gpt_set_background(gpt_ctx *ctx, const float *r, const float *g, const float *b,
                       int32_t bg_height, int32_t bg_width);
int gpt_set_camera(gpt_ctx *ctx, const GptCamera *cam);
int gpt_set_params(gpt_ctx *ctx, const GptParams *params);

/* ---- scene packer: replaces the host loaders of src/main.cu:165-471 --------------------------
 *      (InputParam::extract_parameters input_param.h:36-122, extract_material_file_names /
 *      extract_material_data / extract_num_elements / extract_triangle_data
 *      read_file_util.h:125-655, extract_image_resource_requirement / extract_textures
 *      read_image_util.h:47-176).  Host work: the OBJ file is mapped once and parsed by
 *      `threads` host threads (0 = all cores) in two passes (count, scan, fill); the SoA it
 *      produces is value-identical to the reference loader's (deviations: loader.cu header). */

/* Decodes `path` into 8-bit RGB scan lines, top to bottom, in a malloc()ed buffer the library
 * frees; returns 0 on success.  This is where the host's libjpeg goes (the reference reads JPEG
 * through external/libjpeg_cpp/jpeg.h); with reader == NULL the built-in readers accept the raw
 * "GPTIMG <w> <h>\n<rgb bytes>" container and binary PPM (P6). */
typedef int (*gpt_image_reader)(void *user, const char *path, int32_t *height, int32_t *width,
                                uint8_t **rgb);

typedef struct gpt_scene_files gpt_scene_files;

/* Borrowed pointers into a parsed scene; valid until gpt_scene_files_free. */
typedef struct GptSceneFilesView {
  GptMesh mesh;
  GptMaterials materials;
  GptCamera camera;
  GptParams params;
  const float *tex_r, *tex_g, *tex_b; /* planar atlas, texel 0 = the white default */
  int64_t texture_length;
  const float *bg_r, *bg_g, *bg_b;
  int32_t bg_height, bg_width;
  int32_t num_textures;               /* atlas entries incl. the default */
  const int32_t *texture_offset, *texture_height, *texture_width;
  const char *input_folder_path, *obj_filename, *texture_bg_path, *image_output_path;
  int32_t ppm_image_output_iteration;
  double parse_ms;                    /* whole gpt_scene_files_parse                      */
  double obj_parse_ms;                /* OBJ + MTL + textures                             */
  int64_t obj_bytes;
  int32_t threads;                    /* host threads the OBJ passes actually used        */
} GptSceneFilesView;

/* Needs no GPU.  On failure returns GPT_E_INVALID and writes the reason into err. */
int gpt_scene_files_parse(const char *master_file_path, int threads, gpt_image_reader reader,
                          void *reader_user, gpt_scene_files **out, char *err, int err_len);
int gpt_scene_files_view(const gpt_scene_files *s, GptSceneFilesView *view);
const char *gpt_scene_files_material_name(const gpt_scene_files *s, int32_t i);
const char *gpt_scene_files_texture_name(const gpt_scene_files *s, int32_t i);
void gpt_scene_files_free(gpt_scene_files *s);
/* gpt_set_mesh + _materials + _textures + _background + _camera + _params from a parsed scene */
int gpt_load_scene_files(gpt_ctx *ctx, const gpt_scene_files *s);

/* ---- build: replaces create_world's derived data (triangle.h:254-321), tangents
 *      (triangle_operations.h:18-61) and the three LBVH builds (src/main.cu:826-1197) --- */
int gpt_build(gpt_ctx *ctx);
/* Rebuild only the main BVH from the resident SoA (the timed "LBVH build ms" stage). */
int gpt_rebuild_main_bvh(gpt_ctx *ctx);

/* ---- parity exports ------------------------------------------------------------------ */
/* Per-triangle derived data in ORIGINAL (OBJ face) order.  Any pointer may be NULL. */
int gpt_export_triangles(gpt_ctx *ctx, float *tolerance /*N*/, float *aabb /*6N: xmin,xmax,
                         ymin,ymax,zmin,zmax*/, float *normal /*3N*/, float *area /*N*/,
                         int32_t *flags /*N*/, float *energy /*3N*/,
                         float *tangents /*9N: corner 1,2,3*/);
int gpt_bvh_num_leaves(gpt_ctx *ctx, int which, int32_t *n);
/* Child references: c >= 0 is internal node c, c < 0 is leaf ~c (sorted position).
 * Internal node i lives at index i, the root is 0 (bvh_building.h:63,72,179-183). */
int gpt_export_bvh(gpt_ctx *ctx, int which, uint32_t *morton /*n, sorted*/,
                   int32_t *sorted_to_orig /*n: original triangle index of leaf j*/,
                   int32_t *left /*n-1*/, int32_t *right /*n-1*/,
                   int32_t *parent_internal /*n-1, root = -1*/, int32_t *parent_leaf /*n*/,
                   float *node_aabb /*6(n-1)*/, float *leaf_aabb /*6n*/,
                   float *world_aabb /*6*/);

/* ---- trace: replaces traverse_bvh + Triangle::hit (bvh_traversal.h:12-84,
 *      triangle.h:343-408).  rays: n x 8 floats {ox,oy,oz,tmax, dx,dy,dz,0}; the direction
 *      is normalised exactly as Ray::Ray does (ray.h:21-24); tmax = INFINITY reproduces
 *      rec.t = INFINITY.  Outputs (any may be NULL): leaf_id = sorted position or -1,
 *      tri_id = original triangle index or -1, t, bary = {b1,b2,b3}. ------------------- */
int gpt_trace(gpt_ctx *ctx, int which, const float *rays, int64_t n, int32_t *leaf_id,
              int32_t *tri_id, float *t, float *bary);
/* gpt_trace without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments, same results bit for bit as
 * gpt_trace, but the device staging of rays and hit records and the trace kernel's scratch (ray
 * cursor, counters, overflow flag) live in one file-scope `static __device__` array of
 * 256 + 64 * GPT_TRACE_STATIC_MAX_RAYS bytes (8 MiB) that every GPU loading the library holds
 * until exit; the call makes no cudaMalloc / cudaFree / pool allocation and needs no host
 * staging buffer.  n > GPT_TRACE_STATIC_MAX_RAYS returns GPT_E_INVALID with a message (split the
 * call, or use gpt_trace).  The array is shared by every context on the device and keeps the last
 * call's contents (what must start at zero is cleared per call): calls that overlap on the GPU
 * race -- the function returns only once its own work is complete, so serialise callers per
 * device. */
#define GPT_TRACE_STATIC_MAX_RAYS 131072
int gpt_trace_static(gpt_ctx *ctx, int which, const float *rays, int64_t n, int32_t *leaf_id,
                     int32_t *tri_id, float *t, float *bary);
/* Same, all pointers are DEVICE pointers; hits: n x 4 x 4 bytes {t, leaf_id, b1, b2}
 * with leaf_id stored as int32 bits.  Asynchronous on the context's stream. */
int gpt_trace_device(gpt_ctx *ctx, int which, const void *d_rays, int64_t n, void *d_hits);
/* Count V_int / V_leaf of the same traversal (not timed) into gpt_stats. */
int gpt_trace_count_device(gpt_ctx *ctx, int which, const void *d_rays, int64_t n);

/* ---- path tracing: replaces init_curand_state + path_tracing_render
 *      (src/main.cu:1208,1266; src/render/pathtracing.h:251-305).  out_rgb: P x 3 linear
 *      fp32, row-major, top row first. ------------------------------------------------- */
int gpt_render_pathtrace(gpt_ctx *ctx, float *out_rgb);
/* Render only pixels whose 32x32 tile index t satisfies t % nranks == rank; other pixels
 * are left untouched in d_out_rgb (DEVICE pointer, P x 3).  Asynchronous. */
int gpt_render_pathtrace_device(gpt_ctx *ctx, void *d_out_rgb, int rank, int nranks);

/* ---- multi-GPU: one process (or thread) per GPU, each with its own context and
 *      the whole scene; rank r renders the 32x32 tiles t with t % nranks == r and traces the
 *      photons of its index range.  The reference has no counterpart (single GPU).
 *      NCCL is loaded at run time (libnccl.so.2); all calls are enqueued on the context's stream.
 *
 *      path tracing:  gpt_render_pathtrace_device(ctx, d_rgb, rank, nranks)   -- d_rgb zeroed first
 *                     gpt_gather_frame(ctx, d_rgb, 0)                          -- root has the frame
 *      SPPM, per pass: gpt_sppm_pass(ctx, pass, begin_r, end_r);
 *                     gpt_sppm_allreduce(ctx);   <- the ONE collective of the data path: between the
 *                                                   traversal and the update half of
 *                                                   update_hit_point_parameters
 *                                                   (bvh_traversal_photon.h:207-225, main.cu:1780)
 *                     gpt_sppm_update(ctx, pass);
 * ----------------------------------------------------------------------------------------- */
#define GPT_NCCL_UNIQUE_ID_BYTES 128
/* ncclGetUniqueId: call on one rank, hand the 128 bytes to the others by any means. */
int gpt_comm_unique_id(void *out_id);
/* ncclCommInitRank on the context's device; collective over the nranks contexts. */
int gpt_comm_init(gpt_ctx *ctx, const void *nccl_unique_id, int rank, int nranks);
/* rank / nranks of the communicator (0 / 1 without one) and the NCCL version in use. */
int gpt_comm_info(gpt_ctx *ctx, int *rank, int *nranks, int *nccl_version);
int gpt_comm_destroy(gpt_ctx *ctx);
/* ncclAllReduce(sum) of the 8 P floats of gpt_sppm_flux_device, in place.  No-op with one rank. */
int gpt_sppm_allreduce(gpt_ctx *ctx);
/* ncclReduce(sum) of a P x 3 fp32 DEVICE frame to `root`, in place: frames that are zero outside
 * their own tiles add up to the image exactly.  No-op with one rank. */
int gpt_gather_frame(gpt_ctx *ctx, void *d_rgb, int root);

/* ---- sub-surface scattering: replaces src/main.cu:642-824 (point clouds per
 *      SSS object + per-object point LBVHs, built by gpt_build when sss_pts_per_object > 0),
 *      do_sss_first_pass (src/main.cu:1213-1232, run automatically by the first render) and
 *      compute_color_sss / traverse_bvh_pts inside the bounce loop.
 *      Parity export, points in tree order: object index + 12 floats (location, filter, normal,
 *      colour); run_first_pass != 0 computes the colours first.  *count = number of points;
 *      nothing is written when capacity < *count. -------------------------------------------- */
int gpt_sss_export_points(gpt_ctx *ctx, int run_first_pass, int32_t *object_idx, float *rec12,
                          int64_t capacity, int64_t *count);

/* ---- SPPM: replaces the render_mode == 2 branch, src/main.cu:1282-1870 ------------------
 * Call order of one frame (single GPU):
 *     gpt_sppm_init(ctx);
 *     for (pass = 0; pass < ppm_num_pass; pass++) {
 *         gpt_sppm_pass(ctx, pass, 0, ppm_num_photon_per_pass);
 *         gpt_sppm_update(ctx, pass);
 *     }
 *     gpt_sppm_resolve(ctx, 0, ppm_num_pass, out_rgb);
 * Multi-GPU: every rank runs the same sequence with its own photon range and all-reduces
 * (sum) the buffer of gpt_sppm_flux_device between gpt_sppm_pass and gpt_sppm_update. */
/* ray_tracing_pass(init=true) + compute_average_radius + assign_radius_to_invalid_hit_points
 * + compute_accummulated_light_source_energy (src/main.cu:1311-1460). */
int gpt_sppm_init(gpt_ctx *ctx);
/* photon_pass + gather_recorded_photons + both photon LBVHs + ray_tracing_pass + the
 * traversal half of update_hit_point_parameters (src/main.cu:1496-1779), for the global
 * photon indices [photon_begin, photon_end).  Leaves this pass' per-hit-point accumulators
 * in the buffer of gpt_sppm_flux_device. */
int gpt_sppm_pass(gpt_ctx *ctx, int pass, int64_t photon_begin, int64_t photon_end);
/* DEVICE buffer of 8 P floats: [0, 4P) = (flux.r, flux.g, flux.b, photon count) per pixel,
 * [4P, 8P) = (tmp_accummulated_lm.rgb, 0) per pixel. */
int gpt_sppm_flux_device(gpt_ctx *ctx, void **d_flux, int64_t *num_floats);
/* PPMHitPoint::update_accummulated_reflected_flux with iteration = pass + 1
 * (ppm_hit_point.h:115-171). */
int gpt_sppm_update(gpt_ctx *ctx, int pass);
/* get_ppm_image_output (image_output.h:8-31): mode 0 = global (direct + indirect),
 * 1 = direct, 2 = indirect; divided by `passes`.  out_rgb: P x 3 host floats. */
int gpt_sppm_resolve(gpt_ctx *ctx, int mode, int passes, float *out_rgb);
/* last pass: stored surface / volume photons, mean surface radius (compute_average_radius,
 * src/main.cu:1787-1798), main-BVH rays since gpt_sppm_init.  Any pointer may be NULL. */
int gpt_sppm_stats(gpt_ctx *ctx, int64_t *n_surface, int64_t *n_volume, double *mean_radius,
                   uint64_t *rays_total);
/* parity exports (host pointers, any may be NULL): hit points as P x 4 floats each --
 * (location, surface radius) (normal, volume radius) (filter, 0) (direct radiance, 0)
 * (indirect radiance, photons) and the 2P x 4 flux buffer described above. */
int gpt_sppm_export_hit_points(gpt_ctx *ctx, float *loc_radius, float *normal_vradius,
                               float *filter, float *direct, float *indirect, float *flux_lm);
/* photons of the last pass in photon-index order: kind (0 surface, 1 volume, 2 not stored),
 * 12 floats each (location, prev_location, direction, colour).  *count = photons traced;
 * nothing is written when capacity < *count. */
int gpt_sppm_export_photons(gpt_ctx *ctx, int32_t *kind, float *rec12, int64_t capacity,
                            int64_t *count);
/* one pass split into stages (bit mask): 1 = photon trace + compaction + both photon trees,
 * 2 = eye pass with the volume beam gather, 4 = surface gather.  gpt_sppm_pass = mask 7. */
int gpt_sppm_pass_stages(gpt_ctx *ctx, int pass, int64_t photon_begin, int64_t photon_end,
                         int stage_mask);
/* parity import: replace the photons of the current pass by n caller-supplied ones (kind +
 * 12 floats each, as exported) and build both photon trees from them. */
int gpt_sppm_import_photons(gpt_ctx *ctx, const int32_t *kind, const float *rec12, int64_t n);
/* Multi-GPU SPPM, r02 (second variant).  With a communicator of > 1 ranks bound
 * (gpt_comm_init), gpt_sppm_pass shards the eye pass too: the volume photons of all ranks are
 * all-gathered in photon order (so every rank builds the WHOLE pass's sphere tree and the
 * reference's prefix-weighted volume term, bvh_traversal_photon.h:62-78, is the single-GPU one),
 * each rank traces the eye paths of its own 32x32 tiles only (ray_tracing_pass.h:490-594) and
 * one sum all-reduce of 12 floats per pixel hands every rank every hit point; the surface
 * gather and gpt_sppm_allreduce / gpt_sppm_update are unchanged.  Environment
 * GPT_SPPM_SHARD_EYE=0 keeps the replicated eye pass.  This test hook applies the tile
 * ownership WITHOUT a communicator (other ranks' pixels come out zero); nranks <= 1 = off. */
int gpt_sppm_debug_eye_shard(gpt_ctx *ctx, int rank, int nranks);
/* parity import: overwrite hit-point (location, surface radius) / (normal, volume radius) /
 * filter with caller-supplied P x 4 host floats (NULL = keep). */
int gpt_sppm_import_hit_points(gpt_ctx *ctx, const float *loc_radius,
                               const float *normal_vradius, const float *filter);

/* ---- output stage: replaces save_image (src/util/image_util.h:50-72) ----
 * display transform (int)(255 * clamp(0, 1, sqrt(c))) of n_values host floats */
int gpt_tonemap_u8(gpt_ctx *ctx, const float *rgb, int64_t n_values, uint8_t *out);
/* ASCII PPM exactly as save_image writes it: "P3\nW H\n255\n", then "r g b " per pixel */
int gpt_write_ppm_u8(const char *path, const uint8_t *rgb_u8, int32_t width, int32_t height);
int gpt_save_image(gpt_ctx *ctx, const float *rgb, int32_t width, int32_t height, const char *path);
/* ---- debug splats (8 f-4): create_point_image (point_operations.h:19-75).  which: 0 = SPPM
 * hit points (colour = filter), 1 = surface photons, 2 = volume photons of the last pass. */
int gpt_sppm_debug_image(gpt_ctx *ctx, int which, float *out_rgb);
int gpt_sss_debug_image(gpt_ctx *ctx, float *out_rgb);

/* Profiling switches for the NEXT renders (never on in a timed headline run):
 *   time_stages  - bracket every trace / shade launch with CUDA events (fills
 *                  trace_kernel_ms / shade_kernel_ms)
 *   count_visits - count V_int / V_leaf of the main-BVH traversals (fills internal_visits /
 *                  leaf_tests) */
int gpt_set_profiling(gpt_ctx *ctx, int time_stages, int count_visits);
/* Number of independent wavefront groups (1..8) the path tracer cuts a rank's pixels into, each
 * on its own stream; 0 = automatic (4 for frames of >= 65536 pixels per rank, else 1; the
 * environment variable GPT_WAVEFRONT_GROUPS overrides the automatic choice).  The image does
 * not depend on it: a pixel's samples never interact with another pixel's. */
int gpt_set_wavefront_groups(gpt_ctx *ctx, int groups);
/* How the persistent trace kernels end a launch.  1: once the ray queue is empty and a warp is
 * down to <= 16 live rays, the warp finishes them cooperatively -- their pending nodes go into
 * one shared work list that all 32 lanes pop from -- instead of walking them one lane each
 * (the chain of dependent node steps of the slowest ray is what a launch's tail, and with it
 * strong scaling, is made of).  0: every ray stays on its lane.  -1 (default): on, unless the
 * environment variable GPT_TRACE_COOP=0.  Hits are bit-identical either way: closest hit with
 * the reference's visit-rank tie-break does not depend on the order nodes are visited in. */
int gpt_set_trace_tail(gpt_ctx *ctx, int mode);
/* Path tracer's trace stage (traverse_bvh + Triangle::hit, src/model/bvh/bvh_traversal.h:12-84,
 * src/model/geometry/triangle.h:343-408).  1: a traversal step only queues the leaves it finds
 * (per-lane LIFO in shared memory) and the warp runs the triangle test for all its lanes
 * together once a lane holds more than 4 -- the test is ~95 instructions and, done where the
 * leaf is found, runs with ~3 of 32 lanes on.  0: every leaf is tested where it is found.
 * -1 (default): on, unless the environment variable GPT_TRACE_DEFER=0.  Hits are bit-identical
 * either way (closest hit with the reference's visit-rank tie-break does not depend on the
 * order of the tests; a later test only means a looser cull bound in between). */
int gpt_set_trace_leaf_batching(gpt_ctx *ctx, int mode);
/* rng_mode 1: independent sample streams per pixel, 1..64 (0 = default, 8); clamped to
 * dof_sample_size * pathtracing_sample_size.  The image depends on this number (and the scene)
 * only, not on the rank count; with 1 lane it is the rng_mode 0 image bit for bit. */
int gpt_set_fast_rng_lanes(gpt_ctx *ctx, int lanes);
/* Which tree the trace kernels walk (traverse_bvh, src/model/bvh/bvh_traversal.h:12-84).
 *   1: the 4-wide tree gpt_build collapses from the LBVH (one 128-byte node per visit holds four
 *      boxes; ~23 dependent node fetches per bounce ray on the 1 M-triangle config instead of
 *      ~49).  The walk is conservative and every candidate hit is filtered by the reference's
 *      exact slab predicate on the triangle's own box, so the hit is the one the reference's
 *      walk of its binary tree returns, bit for bit; rays with a zero / denormal / non-finite
 *      direction component take the binary walk.
 *   0: the ordered walk of the binary LBVH itself.
 *  -1 (default): 1, unless the environment variable GPT_TRACE_WIDE=0.
 * gpt_export_bvh always exports the binary LBVH (the reference's tree). */
int gpt_set_trace_bvh(gpt_ctx *ctx, int mode);
/* How gpt_render_pathtrace* runs path_tracing_render's dof x spp x level loop
 * (src/render/pathtracing.h:251-305).
 *   0 "rounds":  one round of launches (trace, material sort, shade) per bounce of every pixel,
 *                in 1..8 independent groups (gpt_set_wavefront_groups).
 *   1 "stream":  two persistent kernels, trace and shade, run side by side for the whole frame
 *                and hand pixel slots to each other through rings in device memory (one ring
 *                per material class towards shade), so no pixel waits for another pixel's ray.
 *  -1 (default): environment variable GPT_PIPELINE ("rounds" / "stream"), else chosen by the
 *                number of pixels this rank renders.
 * The frame and the ray count are bit-identical either way: pixels never interact, and a
 * pixel's samples stay sequential on its own RNG stream. */
int gpt_set_pipeline(gpt_ctx *ctx, int mode);
int gpt_stats(gpt_ctx *ctx, GptStats *out);
int gpt_synchronize(gpt_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* GPT_H */
