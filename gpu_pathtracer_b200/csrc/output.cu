// output.cu -- the stage immediately downstream of the hot path:
// display transform + ASCII PPM writer byte-identical to save_image (src/util/image_util.h:
// 50-72), and the debug point splats of create_point_image (src/model/point/point_operations.h:
// 19-75) for hit points, photons and SSS points.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "integrator.cuh"

namespace {

// (int)(255 * clamp(0, 1, sqrt(c))) per channel (image_util.h:64-68); sqrt of a float is the
// IEEE fp32 square root on both sides, NaN and negatives clamp to 1 through fminf/fmaxf
__global__ void __launch_bounds__(256)
k_tonemap(const float *__restrict__ rgb, size_t n, unsigned char *out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // nvcc folds fmaxf(0, fminf(1, x)) into a .sat operation, which maps NaN to 0; the host
  // code of save_image gets 1 from fminf(1, NaN), so NaN is handled explicitly
  float r = sqrtf(rgb[i]);
  float c = (r != r) ? 1.f : fmaxf(0.f, fminf(1.f, r));
  out[i] = (unsigned char)(int)(255.f * c);
}

// Camera::compute_i / compute_j (camera.h:54-66) of the direction eye -> point.  Several
// points can land on one pixel; the reference lets the last writer win (a race), here the
// point with the HIGHEST index IN THE REFERENCE'S ARRAY ORDER wins -- what the reference's
// kernel does when its threads run in index order, i.e. the host replay's answer.  `order`
// maps a point of our array to its index in the reference's array (nullptr: the same order);
// winner words are (order + 1) << 32 | our index, 0 = no point.
__global__ void __launch_bounds__(256)
k_splat_pick(CameraDev cam, const float *__restrict__ loc, int loc_stride, int n,
             const int *__restrict__ order, unsigned long long *winner) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  V3 p = V3{loc[(size_t)idx * loc_stride], loc[(size_t)idx * loc_stride + 1],
            loc[(size_t)idx * loc_stride + 2]};
  if (has_inf(p)) return;
  V3 dir = unit(p - cam.eye);
  float w2 = (float)cam.width / 2, h2 = (float)cam.height / 2;
  int j = (int)floorf((float)((double)w2 + (double)w2 / cam.tan_half_fovx * (double)dot(dir + cam.w, cam.u)));
  int i = (int)floorf((float)((double)h2 - (double)h2 / cam.tan_half_fovy * (double)dot(dir + cam.w, cam.v)));
  int pix = i * cam.width + j;
  if (pix < cam.width * cam.height && i >= 0 && j >= 0) {
    const unsigned long long key = (unsigned long long)((order ? order[idx] : idx) + 1);
    atomicMax(winner + pix, (key << 32) | (unsigned)idx);
  }
}
__global__ void __launch_bounds__(256)
k_splat_write(const unsigned long long *__restrict__ winner, const float *__restrict__ col,
              int col_stride, int npix, float *out) {
  int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const unsigned long long w = winner[pix];
  const size_t idx = (size_t)(w & 0xffffffffull);
  for (int k = 0; k < 3; k++) out[3 * (size_t)pix + k] = w == 0 ? 0.f : col[idx * col_stride + k];
}

}  // namespace

// device-side splat used by the SPPM / SSS debug images
extern "C" int gpt_splat_device(gpt_ctx *ctx, const float *d_loc, int loc_stride,
                                const float *d_col, int col_stride, int n, const int *d_order,
                                float *out_rgb_host) {
  if (!ctx->have_cam) return gpt_fail(ctx, GPT_E_INVALID, "gpt_set_camera first", __FILE__, __LINE__);
  int W = ctx->cam.width, H = ctx->cam.height, P = W * H;
  unsigned long long *d_win = nullptr;
  float *d_out = nullptr;
  GPT_TRY(gpt_alloc(ctx, &d_win, (size_t)P));
  GPT_TRY(gpt_alloc(ctx, &d_out, 3 * (size_t)P));
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemsetAsync(d_win, 0, sizeof(unsigned long long) * (size_t)P, st));
  if (n > 0) {
    k_splat_pick<<<(n + 255) / 256, 256, 0, st>>>(derive_camera(ctx->cam), d_loc, loc_stride, n, d_order, d_win);
    GPT_LAUNCH_CHECK(ctx);
  }
  k_splat_write<<<(P + 255) / 256, 256, 0, st>>>(d_win, d_col, col_stride, P, d_out);
  GPT_LAUNCH_CHECK(ctx);
  GPT_CUDA(ctx, cudaMemcpyAsync(out_rgb_host, d_out, sizeof(float) * 3 * (size_t)P,
                                cudaMemcpyDeviceToHost, st));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  gpt_free(ctx, d_win);
  gpt_free(ctx, d_out);
  return GPT_OK;
}

extern "C" {

// host -> device -> host display transform of n_values floats (3 per pixel)
int gpt_tonemap_u8(gpt_ctx *ctx, const float *rgb, int64_t n_values, uint8_t *out) {
  if (!ctx || !rgb || !out || n_values < 0) return GPT_E_INVALID;
  if (n_values == 0) return GPT_OK;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  float *d_in = nullptr;
  unsigned char *d_out = nullptr;
  GPT_TRY(gpt_alloc(ctx, &d_in, (size_t)n_values));
  GPT_TRY(gpt_alloc(ctx, &d_out, (size_t)n_values));
  cudaStream_t st = ctx->stream;
  GPT_CUDA(ctx, cudaMemcpyAsync(d_in, rgb, sizeof(float) * (size_t)n_values, cudaMemcpyHostToDevice, st));
  k_tonemap<<<(unsigned)((n_values + 255) / 256), 256, 0, st>>>(d_in, (size_t)n_values, d_out);
  GPT_LAUNCH_CHECK(ctx);
  GPT_CUDA(ctx, cudaMemcpyAsync(out, d_out, (size_t)n_values, cudaMemcpyDeviceToHost, st));
  GPT_CUDA(ctx, cudaStreamSynchronize(st));
  gpt_free(ctx, d_in);
  gpt_free(ctx, d_out);
  return GPT_OK;
}

// "P3\nW H\n255\n" then "r g b " per pixel, no line breaks: the bytes save_image writes
int gpt_write_ppm_u8(const char *path, const uint8_t *rgb_u8, int32_t width, int32_t height) {
  if (!path || !rgb_u8 || width <= 0 || height <= 0) return GPT_E_INVALID;
  FILE *f = fopen(path, "wb");
  if (!f) return GPT_E_INVALID;
  fprintf(f, "P3\n%d %d\n255\n", width, height);
  static const char digits[] = "0123456789";
  std::vector<char> buf;
  buf.reserve(1 << 20);
  size_t n = 3 * (size_t)width * height;
  for (size_t i = 0; i < n; i++) {
    unsigned v = rgb_u8[i];
    if (v >= 100) buf.push_back(digits[v / 100]);
    if (v >= 10) buf.push_back(digits[(v / 10) % 10]);
    buf.push_back(digits[v % 10]);
    buf.push_back(' ');
    if (buf.size() > (1 << 20) - 8) { fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
  }
  fwrite(buf.data(), 1, buf.size(), f);
  return fclose(f) == 0 ? GPT_OK : GPT_E_INVALID;
}

// save_image(image, width, height, filename) (image_util.h:50-72)
int gpt_save_image(gpt_ctx *ctx, const float *rgb, int32_t width, int32_t height, const char *path) {
  if (!ctx || !rgb || !path || width <= 0 || height <= 0) return GPT_E_INVALID;
  std::vector<uint8_t> u8(3 * (size_t)width * height);
  GPT_TRY(gpt_tonemap_u8(ctx, rgb, (int64_t)u8.size(), u8.data()));
  if (gpt_write_ppm_u8(path, u8.data(), width, height) != GPT_OK)
    return gpt_fail(ctx, GPT_E_INVALID, "cannot write the PPM file", __FILE__, __LINE__);
  return GPT_OK;
}

}  // extern "C"
