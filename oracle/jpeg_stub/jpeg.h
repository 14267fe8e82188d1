// jpeg.h -- TEST INFRASTRUCTURE: stand-in for the reference's external/libjpeg_cpp/jpeg.h
// (a libjpeg wrapper; jpeglib.h exists neither here nor on the GPU box).  It provides only
// the surface the reference's loader uses (src/input/read_image_util.h:47-66,178-187):
// marengo::jpeg::Image{ctor(path), getHeight, getWidth, getLuminance(x, y, channel)}, plus the
// TRUE / FALSE macros the reference otherwise gets from jpeglib.h
// (src/model/bvh/bvh_building.h:117,126).  Images are read from a trivial raw format written
// by gpu_pathtracer_b200/scene.py::write_scene_files: "GPTIMG <width> <height>\n" + RGB bytes.
#ifndef GPT_JPEG_STUB_H
#define GPT_JPEG_STUB_H
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>
#ifndef TRUE
#define TRUE 1
#endif
#ifndef FALSE
#define FALSE 0
#endif
namespace marengo {
namespace jpeg {
class Image {
 public:
  explicit Image(const std::string &fileName) {
    FILE *f = fopen(fileName.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open image " + fileName);
    unsigned long w = 0, h = 0;
    if (fscanf(f, "GPTIMG %lu %lu", &w, &h) != 2) { fclose(f); throw std::runtime_error("bad image " + fileName); }
    fgetc(f);  // the newline
    m_width = w; m_height = h;
    m_data.resize(3 * w * h);
    size_t got = fread(m_data.data(), 1, m_data.size(), f);
    fclose(f);
    if (got != m_data.size()) throw std::runtime_error("short image " + fileName);
  }
  size_t getHeight() const { return m_height; }
  size_t getWidth() const { return m_width; }
  float getLuminance(size_t x, size_t y, int channel) const {
    return (float)m_data[3 * (y * m_width + x) + channel];
  }
 private:
  size_t m_width = 0, m_height = 0;
  std::vector<uint8_t> m_data;
};
}  // namespace jpeg
}  // namespace marengo
#endif
