// loader.cu -- the scene packer: master file + OBJ + MTL + textures ->
// the flat SoA of include/gpt.h, then gpt_set_* (the tangent accumulation and every other
// derived quantity is device work in preprocess.cu).  HOST code only; it lives in libgpt.so so
// that the reference's main() can replace src/main.cu:165-471 with two calls.
//
// What it reproduces, line by line, is the reference's loaders:
//   InputParam::extract_parameters          src/input/input_param.h:36-122
//   extract_material_file_names             src/input/read_file_util.h:167-198
//   extract_image_resource_requirement      src/input/read_image_util.h:100-176
//   extract_textures / extract_single_image src/input/read_image_util.h:47-98
//   extract_material_data                   src/input/read_file_util.h:200-466
//   extract_num_elements                    src/input/read_file_util.h:125-165
//   extract_triangle_data                   src/input/read_file_util.h:523-655
// with a different machine underneath: the OBJ file is mapped once and cut into one slice per
// host thread; pass 1 counts elements per slice, an exclusive scan turns the counts into
// output offsets, pass 2 parses floats with std::from_chars (correctly rounded, like the
// strtof behind std::stof, so the floats are bit-identical) straight into the SoA.  The
// reference instead reads the file three times with getline + split + stof on one core.
//
// Reference behaviours kept on purpose (tests/test_loader.py pins them against the reference's
// own headers compiled in oracle/ref_loader.cpp):
//   - material 0 is the built-in "Default_123", texture 0 the 1x1 white "Default_texture";
//   - `clamp(stof(v), 0, 1)` in the MTL reader is called with the arguments in the wrong order
//     for clamp(lo, hi, v) (src/util/general.h:23): it evaluates to max(v, min(0, 1)), i.e. a
//     LOWER bound only.  So d / Tr / Ns are only clamped at 0, Ni only at 1, and priority only
//     from below at -num_materials_so_far;
//   - a new material starts with t_r = 1 (fully transmissive) until `d` or `Tr` says otherwise;
//   - a property line before the first newmtl of a file edits the last material of the
//     previous file;
//   - `f` polygons are fanned (v1, v[i+1], v[i+2]); whether a face has vt / vn is decided from
//     its FIRST corner only; absent vt / vn index 0;
//   - texture rows are flipped (atlas row 0 = last scan line).
// Deviations (all are undefined behaviour in the reference; here they are defined):
//   - x_tex[0] / y_tex[0] are set to 0 when the file has no `vt` (the reference writes the
//     zero at index norm_idx, read_file_util.h:646-650, leaving slot 0 uninitialised);
//   - a face before the first `usemtl`, an unknown usemtl / texture name, or a malformed
//     number is an error (GPT_E_INVALID + text) instead of an out-of-bounds write or a throw;
//   - JPEG decoding belongs to the host's libjpeg (absent from this image): the caller passes
//     a gpt_image_reader; built in are the raw "GPTIMG" container of the test scenes and
//     binary PPM (P6).
#include <fcntl.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <charconv>
#include <chrono>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/gpt.h"

namespace {

struct LoadError {
  std::string text;
};

[[noreturn]] void fail(const char *fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  throw LoadError{buf};
}

// ---- the reference's string helpers (src/util/string_util.h), restated on ranges ---------
inline bool is_blank(char c) { return c == ' ' || c == '\t'; }

// clean_string_end (string_util.h:68-81): ONE trailing byte outside 33..126 is dropped when
// the string is longer than one byte.
inline void clean_end(const char *b, const char *&e) {
  if (e - b > 1) {
    int c = (int)(signed char)e[-1];
    if (c < 33 || c > 126) --e;
  }
}

// reduce + clean_string_end + split(' ') (string_util.h:20-66): tokens separated by runs of
// blanks; a trailing \r (or other control byte) is not part of the last token.
struct Tokens {
  static constexpr int MAX = 64;
  const char *b[MAX], *e[MAX];
  int n = 0;
  std::vector<std::pair<const char *, const char *>> more;  // polygons with > 62 corners
  int size() const { return n + (int)more.size(); }
  const char *begin(int i) const { return i < MAX ? b[i] : more[i - MAX].first; }
  const char *end(int i) const { return i < MAX ? e[i] : more[i - MAX].second; }
  std::string str(int i) const { return std::string(begin(i), end(i)); }
  bool is(int i, const char *lit) const {
    size_t l = strlen(lit);
    return (size_t)(end(i) - begin(i)) == l && memcmp(begin(i), lit, l) == 0;
  }
};

inline void tokenize(const char *b, const char *e, Tokens &t) {
  t.n = 0;
  t.more.clear();
  while (b < e && is_blank(*b)) ++b;
  while (e > b && is_blank(e[-1])) --e;
  clean_end(b, e);
  while (b < e) {
    while (b < e && is_blank(*b)) ++b;
    if (b >= e) break;
    const char *s = b;
    while (b < e && !is_blank(*b)) ++b;
    if (t.n < Tokens::MAX) {
      t.b[t.n] = s;
      t.e[t.n] = b;
      t.n++;
    } else {
      t.more.emplace_back(s, b);
    }
  }
}

// std::stof == strtof on the token (libstdc++ basic_string.h __stoa): longest valid prefix,
// correctly rounded.  from_chars is also correctly rounded; it does not take a leading '+',
// blanks, hex or inf/nan spellings, so anything it does not consume goes to strtof.
inline float parse_float(const char *b, const char *e, const char *what) {
  const char *p = b;
  if (p < e && *p == '+' && p + 1 < e && (p[1] == '.' || (p[1] >= '0' && p[1] <= '9'))) ++p;
  float v;
  auto r = std::from_chars(p, e, v, std::chars_format::general);
  if (r.ec == std::errc() && r.ptr == e) return v;
  std::string s(b, e);
  char *endp = nullptr;
  v = strtof(s.c_str(), &endp);
  if (endp == s.c_str()) fail("%s: '%s' is not a number", what, s.c_str());
  return v;
}

// std::stoi: optional blanks and sign, decimal digits, longest prefix
inline int parse_int(const char *b, const char *e, const char *what) {
  const char *p = b;
  while (p < e && (*p == ' ' || (*p >= 9 && *p <= 13))) ++p;
  bool neg = false;
  if (p < e && (*p == '-' || *p == '+')) neg = *p++ == '-';
  if (p >= e || *p < '0' || *p > '9') fail("%s: '%s' is not an integer", what, std::string(b, e).c_str());
  long long v = 0;
  while (p < e && *p >= '0' && *p <= '9') {
    v = v * 10 + (*p++ - '0');
    if (v > 2147483648LL) fail("%s: integer out of range", what);
  }
  v = neg ? -v : v;
  if (v > 2147483647LL || v < -2147483648LL) fail("%s: integer out of range", what);
  return (int)v;
}

// a whole file, mapped read-only (or empty when it cannot be opened: the reference's
// `if (myfile.is_open())` silently skips missing files)
struct MappedFile {
  const char *data = nullptr;
  size_t size = 0;
  bool ok = false;
  explicit MappedFile(const std::string &path) {
    int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) return;
    struct stat st;
    if (fstat(fd, &st) == 0 && S_ISREG(st.st_mode)) {
      ok = true;
      size = (size_t)st.st_size;
      if (size) {
        void *p = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (p == MAP_FAILED) {
          ok = false;
          size = 0;
        } else {
          data = (const char *)p;
          madvise(p, size, MADV_SEQUENTIAL);
        }
      }
    }
    close(fd);
  }
  ~MappedFile() {
    if (data) munmap((void *)data, size);
  }
  MappedFile(const MappedFile &) = delete;
  MappedFile &operator=(const MappedFile &) = delete;
};

template <class F>
void for_each_line(const char *b, const char *e, F &&f) {
  while (b < e) {
    const char *nl = (const char *)memchr(b, '\n', (size_t)(e - b));
    const char *le = nl ? nl : e;
    if (le > b) f(b, le);
    b = nl ? nl + 1 : e;
  }
}

// ---- images -----------------------------------------------------------------------------
struct Image {
  int h = 0, w = 0;
  std::vector<uint8_t> rgb;  // scan lines top to bottom
};

bool read_builtin_image(const std::string &path, Image &im) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) fail("cannot open image %s", path.c_str());
  char magic[8] = {0};
  unsigned long w = 0, h = 0, maxv = 0;
  bool ok = false;
  if (fread(magic, 1, 2, f) == 2) {
    if (magic[0] == 'G' && magic[1] == 'P') {  // "GPTIMG <w> <h>\n" + RGB bytes
      rewind(f);
      ok = fscanf(f, "GPTIMG %lu %lu", &w, &h) == 2 && fgetc(f) == '\n';
    } else if (magic[0] == 'P' && magic[1] == '6') {  // binary PPM, 8 bit
      ok = fscanf(f, " %lu %lu %lu", &w, &h, &maxv) == 3 && maxv == 255 && fgetc(f) != EOF;
    }
  }
  if (ok && w > 0 && h > 0 && w * h <= (1ul << 30)) {
    im.w = (int)w;
    im.h = (int)h;
    im.rgb.resize(3 * w * h);
    ok = fread(im.rgb.data(), 1, im.rgb.size(), f) == im.rgb.size();
  } else {
    ok = false;
  }
  fclose(f);
  return ok;
}

struct ImageSource {
  gpt_image_reader reader;
  void *user;
  std::unordered_map<std::string, Image> cache;  // every file is decoded once
  const Image &get(const std::string &path) {
    auto it = cache.find(path);
    if (it != cache.end()) return it->second;
    Image im;
    if (reader) {
      int32_t h = 0, w = 0;
      uint8_t *px = nullptr;
      if (reader(user, path.c_str(), &h, &w, &px) != 0 || !px || h <= 0 || w <= 0)
        fail("the image reader failed on %s", path.c_str());
      im.h = h;
      im.w = w;
      im.rgb.assign(px, px + 3 * (size_t)h * (size_t)w);
      free(px);
    } else if (!read_builtin_image(path, im)) {
      fail("%s: not a GPTIMG / binary PPM image; JPEG needs a gpt_image_reader from the host "
           "(libjpeg is not part of this library)", path.c_str());
    }
    return cache.emplace(path, std::move(im)).first->second;
  }
};

// extract_single_image (read_image_util.h:47-72): rows bottom-up, value = 1.0 * c / 255.0
void unpack_image(const Image &im, float *r, float *g, float *b) {
  size_t k = 0;
  for (int y = im.h - 1; y >= 0; --y) {
    const uint8_t *row = im.rgb.data() + 3 * (size_t)y * (size_t)im.w;
    for (int x = 0; x < im.w; ++x, ++k) {
      r[k] = (float)(1.0 * (double)row[3 * x + 0] / 255.0);
      g[k] = (float)(1.0 * (double)row[3 * x + 1] / 255.0);
      b[k] = (float)(1.0 * (double)row[3 * x + 2] / 255.0);
    }
  }
}

}  // namespace

// ---- the parsed scene ---------------------------------------------------------------------
struct gpt_scene_files {
  // InputParam (input_param.h:11-33)
  std::string input_folder_path, obj_filename, texture_bg_path, image_output_path;
  GptCamera cam{};
  GptParams par{};
  int ppm_image_output_iteration = 0;
  // textures
  std::vector<std::string> texture_name;
  std::vector<int32_t> texture_offset, texture_height, texture_width;
  std::vector<float> tex_r, tex_g, tex_b, bg_r, bg_g, bg_b;
  int bg_h = 0, bg_w = 0;
  // materials
  std::vector<std::string> material_file, material_name;
  std::vector<float> mf[23];    // in GptMaterials order: ka_x .. g
  std::vector<int32_t> mi[16];  // material_priority, then (height, width, offset) x 5
  // mesh
  std::vector<float> x, y, z, xn, yn, zn, xt, yt;
  std::vector<int32_t> p1, p2, p3, n1, n2, n3, t1, t2, t3, mat, obj, obj_num, obj_off;
  // bookkeeping
  double parse_ms = 0, obj_ms = 0;
  int64_t obj_bytes = 0;
  int threads = 1;
};

namespace {

enum { KA = 0, KD = 3, KS = 6, KE = 9, TF = 12, PATH_LENGTH = 15, T_R, N_S, N_I, BM, SCAT, ABSO, GG };
enum { PRIORITY = 0, TEX_DIFFUSE = 1, TEX_SPECULAR = 4, TEX_EMISSION = 7, TEX_NS = 10, TEX_BUMP = 13 };

// InputParam::extract_parameters (input_param.h:36-122): lines are split on TABs only; a key
// seen twice keeps the last value; dof_sample_size is forced to 1 unless an aperture > 0 was
// read EARLIER in the file (:100-105).
void parse_master(const std::string &path, gpt_scene_files &s) {
  MappedFile mf(path);
  if (!mf.ok) fail("cannot open master file %s", path.c_str());
  s.par.dof_sample_size = 1;
  for_each_line(mf.data, mf.data + mf.size, [&](const char *b, const char *e) {
    clean_end(b, e);
    const char *tab = (const char *)memchr(b, '\t', (size_t)(e - b));
    if (!tab) return;
    std::string key(b, tab);
    const char *vb = tab + 1;
    const char *ve = (const char *)memchr(vb, '\t', (size_t)(e - vb));
    if (!ve) ve = e;
    auto F = [&]() { return parse_float(vb, ve, key.c_str()); };
    auto I = [&]() { return parse_int(vb, ve, key.c_str()); };
    if (key == "input_folder_path") s.input_folder_path.assign(vb, ve);
    else if (key == "obj_filename") s.obj_filename.assign(vb, ve);
    else if (key == "texture_bg_path") s.texture_bg_path.assign(vb, ve);
    else if (key == "image_output_path") s.image_output_path.assign(vb, ve);
    else if (key == "image_width") s.cam.width = I();
    else if (key == "image_height") s.cam.height = I();
    else if (key == "render_mode") s.par.render_mode = I();
    else if (key == "pathtracing_sample_size") s.par.pathtracing_sample_size = I();
    else if (key == "pathtracing_level") s.par.pathtracing_level = I();
    else if (key == "ppm_num_photon_per_pass") s.par.ppm_num_photon_per_pass = I();
    else if (key == "ppm_num_pass") s.par.ppm_num_pass = I();
    else if (key == "ppm_max_bounce") s.par.ppm_max_bounce = I();
    else if (key == "ppm_alpha") s.par.ppm_alpha = F();
    else if (key == "ppm_radius_scaling_factor") s.par.ppm_radius_scaling_factor = F();
    else if (key == "ppm_image_output_iteration") s.ppm_image_output_iteration = I();
    else if (key == "eye_x") s.cam.eye[0] = F();
    else if (key == "eye_y") s.cam.eye[1] = F();
    else if (key == "eye_z") s.cam.eye[2] = F();
    else if (key == "center_x") s.cam.center[0] = F();
    else if (key == "center_y") s.cam.center[1] = F();
    else if (key == "center_z") s.cam.center[2] = F();
    else if (key == "up_x") s.cam.up[0] = F();
    else if (key == "up_y") s.cam.up[1] = F();
    else if (key == "up_z") s.cam.up[2] = F();
    else if (key == "fovy") s.cam.fovy = F();
    else if (key == "aperture") s.cam.aperture = F();
    else if (key == "dof_sample_size") s.par.dof_sample_size = s.cam.aperture > 0 ? I() : 1;
    else if (key == "focus_dist") s.cam.focus_dist = F();
    else if (key == "sky_emission_r") s.par.sky_emission[0] = F();
    else if (key == "sky_emission_g") s.par.sky_emission[1] = F();
    else if (key == "sky_emission_b") s.par.sky_emission[2] = F();
    else if (key == "sss_pts_per_object") s.par.sss_pts_per_object = I();
    else if (key == "hittable_pdf_weight") s.par.hittable_pdf_weight = F();
  });
  if (s.obj_filename.empty()) fail("%s: no obj_filename", path.c_str());
  if (s.cam.width <= 0 || s.cam.height <= 0) fail("%s: image_width / image_height missing", path.c_str());
}

int find_first(const std::vector<std::string> &v, const std::string &name) {
  for (size_t i = 0; i < v.size(); i++)
    if (v[i] == name) return (int)i;
  return -1;
}

inline bool is_map_key(const Tokens &t) {
  return t.is(0, "map_Kd") || t.is(0, "map_Ks") || t.is(0, "map_Ns") || t.is(0, "map_Ke") ||
         t.is(0, "map_Bump");
}

// extract_image_resource_requirement + extract_textures (read_image_util.h:74-176): the atlas
// is [Default_texture 1x1 white] + every distinct file named by the LAST token of a map_*
// line, in order of first appearance over the material files in mtllib order.
void build_atlas(gpt_scene_files &s, ImageSource &img) {
  s.texture_name.assign(1, "Default_texture");
  s.texture_offset.assign(1, 0);
  s.texture_height.assign(1, 1);
  s.texture_width.assign(1, 1);
  long next = 1;
  Tokens t;
  for (const std::string &mfile : s.material_file) {
    MappedFile mf(s.input_folder_path + mfile);
    if (!mf.ok) continue;
    for_each_line(mf.data, mf.data + mf.size, [&](const char *b, const char *e) {
      tokenize(b, e, t);
      if (t.size() < 2 || !is_map_key(t)) return;
      std::string name = t.str(t.size() - 1);
      if (find_first(s.texture_name, name) >= 0) return;
      const Image &im = img.get(s.input_folder_path + name);
      s.texture_name.push_back(name);
      s.texture_offset.push_back((int32_t)next);
      s.texture_height.push_back(im.h);
      s.texture_width.push_back(im.w);
      next += (long)im.h * im.w;
      if (next > 0x7fffffffL) fail("texture atlas exceeds 2^31 texels");
    });
  }
  s.tex_r.assign((size_t)next, 0.f);
  s.tex_g.assign((size_t)next, 0.f);
  s.tex_b.assign((size_t)next, 0.f);
  s.tex_r[0] = s.tex_g[0] = s.tex_b[0] = 1.f;
  for (size_t i = 1; i < s.texture_name.size(); i++) {
    size_t o = (size_t)s.texture_offset[i];
    unpack_image(img.get(s.input_folder_path + s.texture_name[i]), &s.tex_r[o], &s.tex_g[o], &s.tex_b[o]);
  }
}

void push_material(gpt_scene_files &s, const std::string &name, bool builtin_default) {
  s.material_name.push_back(name);
  const float zero3[3] = {0, 0, 0};
  auto set3 = [&](int base, float a, float b, float c) {
    s.mf[base].push_back(a);
    s.mf[base + 1].push_back(b);
    s.mf[base + 2].push_back(c);
  };
  (void)zero3;
  set3(KA, 0, 0, 0);
  set3(KD, .9f, .9f, .9f);
  set3(KS, 0, 0, 0);
  set3(KE, 0, 0, 0);
  set3(TF, 1, 1, 1);
  s.mf[PATH_LENGTH].push_back(0);
  s.mf[T_R].push_back(builtin_default ? 0.f : 1.f);  // read_file_util.h:262 vs :319
  s.mf[N_S].push_back(0);
  s.mf[N_I].push_back(0);
  s.mf[BM].push_back(1);
  s.mf[SCAT].push_back(builtin_default ? -1.f : 0.f);  // :265-266 vs :322-323
  s.mf[ABSO].push_back(builtin_default ? -1.f : 0.f);
  s.mf[GG].push_back(0);
  s.mi[PRIORITY].push_back(-1);
  for (int k = 0; k < 5; k++) {
    s.mi[1 + 3 * k].push_back(s.texture_height[0]);
    s.mi[2 + 3 * k].push_back(s.texture_width[0]);
    s.mi[3 + 3 * k].push_back(s.texture_offset[0]);
  }
}

// extract_material_data (read_file_util.h:200-466).  `lower(v, lo)` is what the reference's
// clamp(v, lo, hi) call evaluates to under clamp(lo, hi, v) = fmaxf(lo, fminf(hi, v)).
void parse_materials(gpt_scene_files &s) {
  for (auto &v : s.mf) v.clear();
  for (auto &v : s.mi) v.clear();
  s.material_name.clear();
  int num_so_far = 0;  // num_materials[0]: updated at the END of every file (:463)
  Tokens t;
  for (const std::string &mfile : s.material_file) {
    if (s.material_name.empty()) push_material(s, "Default_123", true);
    int idx = (int)s.material_name.size() - 1;
    MappedFile mf(s.input_folder_path + mfile);
    if (mf.ok) {
      for_each_line(mf.data, mf.data + mf.size, [&](const char *b, const char *e) {
        tokenize(b, e, t);
        if (t.size() == 0) return;
        const std::string key = t.str(0);
        auto need = [&](int n) {
          if (t.size() < n) fail("%s: '%s' needs %d values", mfile.c_str(), key.c_str(), n - 1);
        };
        auto F = [&](int i) { return parse_float(t.begin(i), t.end(i), key.c_str()); };
        auto set3 = [&](int base) {
          need(4);
          s.mf[base][idx] = F(1);
          s.mf[base + 1][idx] = F(2);
          s.mf[base + 2][idx] = F(3);
        };
        auto set_tex = [&](int base, const std::string &name) {
          int k = find_first(s.texture_name, name);
          if (k < 0) fail("%s: %s names '%s', which is not the last token of the line", mfile.c_str(), key.c_str(), name.c_str());
          s.mi[base][idx] = s.texture_height[k];
          s.mi[base + 1][idx] = s.texture_width[k];
          s.mi[base + 2][idx] = s.texture_offset[k];
        };
        if (key == "newmtl") {
          need(2);
          push_material(s, t.str(1), false);
          idx++;
        } else if (key == "Ka") set3(KA);
        else if (key == "Kd") set3(KD);
        else if (key == "Ks") set3(KS);
        else if (key == "Ke") set3(KE);
        else if (key == "Tf") set3(TF);
        else if (key == "path_length") { need(2); s.mf[PATH_LENGTH][idx] = F(1); }
        else if (key == "d") { need(2); s.mf[T_R][idx] = (float)(1.0 - (double)fmaxf(F(1), fminf(0.f, 1.f))); }
        else if (key == "Tr") { need(2); s.mf[T_R][idx] = fmaxf(F(1), fminf(0.f, 1.f)); }
        else if (key == "Ns") { need(2); s.mf[N_S][idx] = fmaxf(F(1), fminf(0.f, 1e6f)); }
        else if (key == "Ni") { need(2); s.mf[N_I][idx] = fmaxf(F(1), fminf(1.f, 1000.f)); }
        else if (key == "scattering_coef") { need(2); s.mf[SCAT][idx] = F(1); }
        else if (key == "absorption_coef") { need(2); s.mf[ABSO][idx] = F(1); }
        else if (key == "g") { need(2); s.mf[GG][idx] = F(1); }
        else if (key == "priority") {
          need(2);
          float p = (float)parse_int(t.begin(1), t.end(1), "priority");
          s.mi[PRIORITY][idx] = (int32_t)fmaxf(p, fminf((float)-num_so_far, (float)num_so_far));
        } else if (key == "map_Kd") { need(2); set_tex(TEX_DIFFUSE, t.str(1)); }
        else if (key == "map_Ks") { need(2); set_tex(TEX_SPECULAR, t.str(1)); }
        else if (key == "map_Ke") { need(2); set_tex(TEX_EMISSION, t.str(1)); }
        else if (key == "map_Ns") { need(2); set_tex(TEX_NS, t.str(1)); }
        else if (key == "map_Bump") {
          need(2);
          for (int i = 1; i + 1 < t.size(); i += 2)
            if (t.is(i, "-bm")) s.mf[BM][idx] = F(i + 1);
          set_tex(TEX_BUMP, t.str(t.size() - 1));
        }
      });
    }
    num_so_far = (int)s.material_name.size();
  }
  if (s.material_name.size() > 100)
    fail("%d materials: the reference holds at most 100 (src/main.cu:130)", (int)s.material_name.size());
}

}  // namespace

// ---- OBJ: two parallel passes over one mapping ------------------------------------------------
namespace {

struct SliceCount {
  size_t v = 0, vn = 0, vt = 0, tri = 0, usemtl = 0;
  std::vector<std::string> usemtl_name;                   // in order, this slice
  std::vector<std::pair<size_t, std::string>> mtllib;     // (byte position, file name)
  int last_material = -2;                                 // resolved after the MTL files are read
  std::string error;
};

inline int line_kind(const char *b, const char *e) {  // 0 other, 1 v, 2 vn, 3 vt, 4 f, 5 usemtl, 6 mtllib
  while (b < e && is_blank(*b)) ++b;
  size_t n = (size_t)(e - b);
  if (n < 2) return 0;
  if (b[0] == 'v') {
    if (is_blank(b[1])) return 1;
    if (n >= 3 && b[1] == 'n' && is_blank(b[2])) return 2;
    if (n >= 3 && b[1] == 't' && is_blank(b[2])) return 3;
    return 0;
  }
  if (b[0] == 'f') return is_blank(b[1]) ? 4 : 0;
  if (n >= 7 && memcmp(b, "usemtl", 6) == 0 && is_blank(b[6])) return 5;
  if (n >= 7 && memcmp(b, "mtllib", 6) == 0 && is_blank(b[6])) return 6;
  return 0;
}

// split(chunk, '/') (string_util.h:20-30): istringstream + getline drops ONE trailing empty
// field, so "7/" has 1 part, "7//" 2, "7//3" 3.
struct Corner {
  const char *b[3], *e[3];
  int parts = 0;
};

inline void split_corner(const char *b, const char *e, Corner &c) {
  c.parts = 0;
  const char *s = b;
  for (const char *p = b;; ++p) {
    if (p == e || *p == '/') {
      if (p == e && s == e) break;  // nothing after the last '/'
      if (c.parts < 3) {
        c.b[c.parts] = s;
        c.e[c.parts] = p;
      }
      c.parts++;
      if (p == e) break;
      s = p + 1;
    }
  }
}

template <class F>
void run_slices(int T, F &&f) {
  if (T == 1) {
    f(0);
    return;
  }
  std::vector<std::thread> th;
  th.reserve(T);
  for (int i = 0; i < T; i++) th.emplace_back([&f, i] { f(i); });
  for (auto &t : th) t.join();
}

void parse_obj(gpt_scene_files &s, int threads, ImageSource &img) {
  const std::string path = s.input_folder_path + s.obj_filename;
  MappedFile mf(path);
  if (!mf.ok) fail("cannot open OBJ file %s", path.c_str());
  auto t0 = std::chrono::steady_clock::now();
  s.obj_bytes = (int64_t)mf.size;
  const char *B = mf.data, *E = mf.data + mf.size;
  int T = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
  if (T < 1) T = 1;
  if (T > 256) T = 256;
  if (mf.size < (size_t)T * 65536) T = (int)std::max<size_t>(1, mf.size / 65536);
  s.threads = T;
  std::vector<const char *> cut(T + 1);  // slice boundaries on line starts
  cut[0] = B;
  cut[T] = E;
  for (int i = 1; i < T; i++) {
    const char *p = B + mf.size * (size_t)i / (size_t)T;
    if (p < cut[i - 1]) p = cut[i - 1];
    const char *nl = p < E ? (const char *)memchr(p, '\n', (size_t)(E - p)) : nullptr;
    cut[i] = nl ? nl + 1 : E;
  }
  // pass 1: counts (extract_num_elements, read_file_util.h:125-165) + the names that decide
  // material and object indices (extract_material_file_names, :167-198)
  std::vector<SliceCount> cnt(T);
  run_slices(T, [&](int i) {
    SliceCount &c = cnt[i];
    Tokens t;
    for_each_line(cut[i], cut[i + 1], [&](const char *b, const char *e) {
      switch (line_kind(b, e)) {
        case 1: c.v++; break;
        case 2: c.vn++; break;
        case 3: c.vt++; break;
        case 4:
          tokenize(b, e, t);
          if (t.size() >= 4) c.tri += (size_t)t.size() - 3;
          break;
        case 5:
          tokenize(b, e, t);
          c.usemtl++;
          c.usemtl_name.push_back(t.size() > 1 ? t.str(1) : std::string());
          break;
        case 6:
          tokenize(b, e, t);
          for (int k = 1; k < t.size(); k++) c.mtllib.emplace_back((size_t)(b - B), t.str(k));
          break;
        default: break;
      }
    });
  });
  // material files in order of first appearance, after the built-in default (:177)
  s.material_file.assign(1, "file_default_123.mtl");
  for (auto &c : cnt)
    for (auto &m : c.mtllib)
      if (find_first(s.material_file, m.second) < 0) s.material_file.push_back(m.second);
  build_atlas(s, img);
  parse_materials(s);
  std::unordered_map<std::string, int> mat_of;  // find_in_vector: FIRST material of that name
  for (size_t i = 0; i < s.material_name.size(); i++) mat_of.emplace(s.material_name[i], (int)i);

  // exclusive scan of the slice counts
  std::vector<size_t> ov(T + 1, 0), on(T + 1, 0), ot(T + 1, 0), of(T + 1, 0), ou(T + 1, 0);
  for (int i = 0; i < T; i++) {
    ov[i + 1] = ov[i] + cnt[i].v;
    on[i + 1] = on[i] + cnt[i].vn;
    ot[i + 1] = ot[i] + cnt[i].vt;
    of[i + 1] = of[i] + cnt[i].tri;
    ou[i + 1] = ou[i] + cnt[i].usemtl;
  }
  const size_t NV = ov[T], NN = on[T], NT = ot[T], NF = of[T], NO = ou[T];
  if (NF > 0x7fffffffu || NV > 0x7fffffffu) fail("more than 2^31 elements");
  // the material a slice starts in = the last usemtl of the slices before it
  const bool named = s.material_name.size() > 1;  // :551: with no MTL every usemtl maps to 0
  auto resolve = [&](const std::string &name) {
    if (!named) return 0;
    auto it = mat_of.find(name);
    if (it == mat_of.end()) fail("usemtl %s: no such material in the MTL files", name.c_str());
    return it->second;
  };
  std::vector<int> start_mat(T, 0);
  {
    int cur = 0;
    for (int i = 0; i < T; i++) {
      start_mat[i] = cur;
      if (!cnt[i].usemtl_name.empty()) cur = resolve(cnt[i].usemtl_name.back());
    }
  }
  s.x.resize(std::max<size_t>(1, NV)); s.y.resize(s.x.size()); s.z.resize(s.x.size());
  s.xn.assign(std::max<size_t>(1, NN), 0.f); s.yn = s.xn; s.zn = s.xn;
  s.xt.assign(std::max<size_t>(1, NT), 0.f); s.yt = s.xt;
  for (auto *v : {&s.p1, &s.p2, &s.p3, &s.n1, &s.n2, &s.n3, &s.t1, &s.t2, &s.t3, &s.mat, &s.obj}) v->resize(NF);
  s.obj_num.assign(NO, 0);
  s.obj_off.assign(NO, 0);

  // pass 2: fill (extract_triangle_data, read_file_util.h:523-655)
  run_slices(T, [&](int i) {
    SliceCount &c = cnt[i];
    try {
      Tokens t;
      Corner c1, c2, c3;
      size_t iv = ov[i], in = on[i], it = ot[i], jf = of[i], ju = ou[i];
      size_t k_use = 0;
      int cur_mat = start_mat[i];
      long cur_obj = (long)ju - 1;
      for_each_line(cut[i], cut[i + 1], [&](const char *b, const char *e) {
        int kind = line_kind(b, e);
        if (!kind || kind == 6) return;
        tokenize(b, e, t);
        auto F = [&](int k, const char *what) { return parse_float(t.begin(k), t.end(k), what); };
        switch (kind) {
          case 1:
            if (t.size() < 4) fail("OBJ byte %zu: v needs 3 numbers", (size_t)(b - B));
            s.x[iv] = F(1, "v"); s.y[iv] = F(2, "v"); s.z[iv] = F(3, "v");
            iv++;
            break;
          case 2:
            if (t.size() < 4) fail("OBJ byte %zu: vn needs 3 numbers", (size_t)(b - B));
            s.xn[in] = F(1, "vn"); s.yn[in] = F(2, "vn"); s.zn[in] = F(3, "vn");
            in++;
            break;
          case 3:
            if (t.size() < 3) fail("OBJ byte %zu: vt needs 2 numbers", (size_t)(b - B));
            s.xt[it] = F(1, "vt"); s.yt[it] = F(2, "vt");
            it++;
            break;
          case 5:
            cur_mat = resolve(c.usemtl_name[k_use++]);
            cur_obj++;
            s.obj_off[(size_t)cur_obj] = (int32_t)jf;
            ju++;
            break;
          case 4: {
            if (t.size() < 4) return;
            if (cur_obj < 0) fail("OBJ byte %zu: face before the first usemtl", (size_t)(b - B));
            split_corner(t.begin(1), t.end(1), c1);
            const bool has_t = c1.parts > 1 && c1.e[1] > c1.b[1];
            const bool has_n = c1.parts > 2;
            auto idx = [&](const Corner &c, int part) {
              if (c.parts <= part) fail("OBJ byte %zu: face corner lacks field %d", (size_t)(b - B), part + 1);
              return parse_int(c.b[part], c.e[part], "f") - 1;
            };
            const int p_1 = idx(c1, 0);
            const int t_1 = has_t ? idx(c1, 1) : 0, n_1 = has_n ? idx(c1, 2) : 0;
            for (int k = 0; k + 3 < t.size(); k++, jf++) {
              split_corner(t.begin(2 + k), t.end(2 + k), c2);
              split_corner(t.begin(3 + k), t.end(3 + k), c3);
              s.obj[jf] = (int32_t)cur_obj;
              s.p1[jf] = p_1; s.p2[jf] = idx(c2, 0); s.p3[jf] = idx(c3, 0);
              s.t1[jf] = t_1; s.t2[jf] = has_t ? idx(c2, 1) : 0; s.t3[jf] = has_t ? idx(c3, 1) : 0;
              s.n1[jf] = n_1; s.n2[jf] = has_n ? idx(c2, 2) : 0; s.n3[jf] = has_n ? idx(c3, 2) : 0;
              s.mat[jf] = cur_mat;
            }
            break;
          }
        }
      });
    } catch (LoadError &err) {
      c.error = err.text;
    }
  });
  for (auto &c : cnt)
    if (!c.error.empty()) throw LoadError{c.error};
  // object_num_primitives (:627): every face belongs to the latest usemtl
  for (size_t o = 0; o < NO; o++)
    s.obj_num[o] = (int32_t)((o + 1 < NO ? (size_t)s.obj_off[o + 1] : NF) - (size_t)s.obj_off[o]);
  // index range: the reference would read out of bounds
  const int nv = (int)s.x.size(), nn = (int)s.xn.size(), nt = (int)s.xt.size();
  for (size_t f = 0; f < NF; f++) {
    auto bad = [](int i, int n) { return i < 0 || i >= n; };
    if (bad(s.p1[f], nv) || bad(s.p2[f], nv) || bad(s.p3[f], nv) || bad(s.n1[f], nn) ||
        bad(s.n2[f], nn) || bad(s.n3[f], nn) || bad(s.t1[f], nt) || bad(s.t2[f], nt) || bad(s.t3[f], nt))
      fail("face %zu indexes an element the OBJ file does not have (relative indices are not "
           "supported by the reference loader either)", f);
  }
  s.obj_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

// ---- C-ABI ------------------------------------------------------------------------------------
extern "C" {

int gpt_scene_files_parse(const char *master_file_path, int threads, gpt_image_reader reader,
                          void *reader_user, gpt_scene_files **out, char *err, int err_len) {
  auto say = [&](const std::string &m) {
    if (err && err_len > 0) snprintf(err, (size_t)err_len, "%s", m.c_str());
  };
  if (out) *out = nullptr;
  if (!master_file_path || !out) {
    say("gpt_scene_files_parse: null argument");
    return GPT_E_INVALID;
  }
  gpt_scene_files *s = nullptr;
  try {
    s = new gpt_scene_files();
    auto t0 = std::chrono::steady_clock::now();
    parse_master(master_file_path, *s);
    ImageSource img{reader, reader_user, {}};
    // background first (src/main.cu:165-186), then materials + textures, then the mesh
    {
      const Image &bg = img.get(s->input_folder_path + s->texture_bg_path);
      s->bg_h = bg.h;
      s->bg_w = bg.w;
      size_t n = (size_t)bg.h * bg.w;
      s->bg_r.resize(n); s->bg_g.resize(n); s->bg_b.resize(n);
      unpack_image(bg, s->bg_r.data(), s->bg_g.data(), s->bg_b.data());
    }
    parse_obj(*s, threads, img);
    s->par.rng_mode = 0;
    s->parse_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  } catch (LoadError &e) {
    say(e.text);
    delete s;
    return GPT_E_INVALID;
  } catch (std::exception &e) {
    say(std::string("scene loader: ") + e.what());
    delete s;
    return GPT_E_INVALID;
  }
  *out = s;
  return GPT_OK;
}

void gpt_scene_files_free(gpt_scene_files *s) { delete s; }

static void fill_mesh(const gpt_scene_files *s, GptMesh *m) {
  m->num_vertices = (int32_t)s->x.size();
  m->num_normals = (int32_t)s->xn.size();
  m->num_texcoords = (int32_t)s->xt.size();
  m->num_triangles = (int32_t)s->p1.size();
  m->num_objects = (int32_t)s->obj_num.size();
  m->x = s->x.data(); m->y = s->y.data(); m->z = s->z.data();
  m->x_norm = s->xn.data(); m->y_norm = s->yn.data(); m->z_norm = s->zn.data();
  m->x_tex = s->xt.data(); m->y_tex = s->yt.data();
  m->point_1_idx = s->p1.data(); m->point_2_idx = s->p2.data(); m->point_3_idx = s->p3.data();
  m->norm_1_idx = s->n1.data(); m->norm_2_idx = s->n2.data(); m->norm_3_idx = s->n3.data();
  m->tex_1_idx = s->t1.data(); m->tex_2_idx = s->t2.data(); m->tex_3_idx = s->t3.data();
  m->material_idx = s->mat.data();
  m->triangle_object_idx = s->obj.data();
  m->object_num_primitives = s->obj_num.data();
  m->object_primitive_offset_idx = s->obj_off.data();
}

static void fill_materials(const gpt_scene_files *s, GptMaterials *m) {
  m->num_materials = (int32_t)s->material_name.size();
  const float **fp = &m->ka_x;  // 23 float pointers, then 16 int pointers, in declaration order
  for (int i = 0; i < 23; i++) fp[i] = s->mf[i].data();
  const int32_t **ip = &m->material_priority;
  for (int i = 0; i < 16; i++) ip[i] = s->mi[i].data();
}

int gpt_scene_files_view(const gpt_scene_files *s, GptSceneFilesView *v) {
  if (!s || !v) return GPT_E_INVALID;
  memset(v, 0, sizeof *v);
  fill_mesh(s, &v->mesh);
  fill_materials(s, &v->materials);
  v->camera = s->cam;
  v->params = s->par;
  v->tex_r = s->tex_r.data(); v->tex_g = s->tex_g.data(); v->tex_b = s->tex_b.data();
  v->texture_length = (int64_t)s->tex_r.size();
  v->bg_r = s->bg_r.data(); v->bg_g = s->bg_g.data(); v->bg_b = s->bg_b.data();
  v->bg_height = s->bg_h;
  v->bg_width = s->bg_w;
  v->num_textures = (int32_t)s->texture_name.size();
  v->texture_offset = s->texture_offset.data();
  v->texture_height = s->texture_height.data();
  v->texture_width = s->texture_width.data();
  v->input_folder_path = s->input_folder_path.c_str();
  v->obj_filename = s->obj_filename.c_str();
  v->texture_bg_path = s->texture_bg_path.c_str();
  v->image_output_path = s->image_output_path.c_str();
  v->ppm_image_output_iteration = s->ppm_image_output_iteration;
  v->parse_ms = s->parse_ms;
  v->obj_parse_ms = s->obj_ms;
  v->obj_bytes = s->obj_bytes;
  v->threads = s->threads;
  return GPT_OK;
}

const char *gpt_scene_files_material_name(const gpt_scene_files *s, int32_t i) {
  return (s && i >= 0 && (size_t)i < s->material_name.size()) ? s->material_name[(size_t)i].c_str() : nullptr;
}

const char *gpt_scene_files_texture_name(const gpt_scene_files *s, int32_t i) {
  return (s && i >= 0 && (size_t)i < s->texture_name.size()) ? s->texture_name[(size_t)i].c_str() : nullptr;
}

// main.cu:474-640 in one call: everything the parsed files hold goes through the gpt_set_* entry
// points (which copy to HBM); gpt_build is the caller's next step.
int gpt_load_scene_files(gpt_ctx *ctx, const gpt_scene_files *s) {
  if (!ctx || !s) return GPT_E_INVALID;
  GptSceneFilesView v;
  gpt_scene_files_view(s, &v);
  int rc;
  if ((rc = gpt_set_mesh(ctx, &v.mesh)) != GPT_OK) return rc;
  if ((rc = gpt_set_materials(ctx, &v.materials)) != GPT_OK) return rc;
  if ((rc = gpt_set_textures(ctx, v.tex_r, v.tex_g, v.tex_b, v.texture_length)) != GPT_OK) return rc;
  if ((rc = gpt_set_background(ctx, v.bg_r, v.bg_g, v.bg_b, v.bg_height, v.bg_width)) != GPT_OK) return rc;
  if ((rc = gpt_set_camera(ctx, &v.camera)) != GPT_OK) return rc;
  return gpt_set_params(ctx, &v.params);
}

}  // extern "C"
