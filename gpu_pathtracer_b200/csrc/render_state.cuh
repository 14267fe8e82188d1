// render_state.cuh -- wavefront path-tracer state in HBM (SoA, one slot per pixel).
//
// The reference's megakernel keeps one thread per pixel alive for dof x spp x level bounces
// (src/render/pathtracing.h:251-305).  Here a pixel is a SLOT whose state lives in HBM between
// the stages of the wavefront (ray-gen/shade -> trace -> shade ...).  A pixel's samples are
// processed sequentially on its own XORWOW stream, so RNG consumption order -- and therefore
// every sample -- is identical to the reference; parallelism is across pixels.
#pragma once
#include "bsdf.cuh"

struct CameraDev {  // Camera (src/model/camera.h:30-52), derived on the host
  V3 eye, u, v, w;
  double tan_half_fovx, tan_half_fovy;  // tan(fov * M_PI / 180.0 / 2), double as in compute_ray
  float focus_dist, lens_radius;
  int width, height;
};

struct PathState {
  int capacity;       // slots (pixels owned by this rank)
  int *pixel;         // slot -> global pixel index (row * width + col)
  // Slots are visited in MATERIAL-sorted order by the shade stage, i.e. at random: everything a
  // slot touches together is packed into whole 32-byte sectors (ncu r01b: the previous
  // plane-per-field layout moved 730 B per path vertex for ~290 B of useful state).
  uint32_t *rng;      // 8 x capacity, packed per slot: (d, v0..v4, -, -)            1 sector
  float4 *rays;       // 2 x capacity: {o, tmax}{d, 0}   (trace input)               1 sector
  float4 *hits;       // 2 x capacity: {t, leaf, prim, -}{b1, b2, b3, -}             1 sector
  float4 *ma;         // 2 x capacity: {mask.rgb, bounce}{acc_color.rgb, sample idx} 1 sector
  float4 *cm;         // 2 x capacity: {pixel colour sum, dof idx}{12 material ids + length}
  float4 *cam;        // 2 x capacity: camera ray {o}{d} of the current dof sample
  int *queue[2];      // active slot lists (ping-pong)
  int *counts;        // [0],[1] queue lengths; [2] matlist overflow flag
};

__device__ __forceinline__ void matlist_store(uint4 *dst, const MatList &l) {
  uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
  for (int i = 0; i < MATLIST_MAX; i++) w[i >> 2] |= (uint32_t)l.id[i] << (8 * (i & 3));
  w[3] = (uint32_t)l.len;
  *dst = make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ void matlist_unpack(MatList &l, uint4 q) {
  uint32_t w[3] = {q.x, q.y, q.z};
#pragma unroll
  for (int i = 0; i < MATLIST_MAX; i++) l.id[i] = (unsigned char)((w[i >> 2] >> (8 * (i & 3))) & 0xff);
  l.len = (int)q.w;
}
__device__ __forceinline__ void matlist_load(MatList &l, const uint4 *src) { matlist_unpack(l, *src); }
