/*
 * ref_shim.h -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * Lets g++ compile the reference's header-only CUDA sources (/root/reference/src) as plain
 * host C++ so that the reference's OWN code can be replayed on the CPU as the ground truth
 * for the parity tests.  Nothing of the reference is copied: the headers are #included from
 * where they lie; this file only supplies the handful of CUDA built-ins they use.
 *
 *   - __device__/__host__/__global__ come from <cuda_runtime.h> (ignored attributes on g++)
 *   - curandState / curand_init / curand_uniform come from the CUDA toolkit's own
 *     <curand_kernel.h>, whose host path is bit-identical to the device path (XORWOW)
 *   - threadIdx / blockIdx / blockDim become settable thread_local variables
 *   - __clz, min/max get the CUDA device semantics (fminf/fmaxf for floats)
 */
#ifndef REF_SHIM_H
#define REF_SHIM_H

#include <cuda_runtime.h>
#include <curand_kernel.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <fstream>   // before "#define private public": util/image_util.h pulls these in later
#include <sstream>

static thread_local uint3 shim_threadIdx = {0, 0, 0}, shim_blockIdx = {0, 0, 0};
static thread_local dim3 shim_blockDim(1, 1, 1), shim_gridDim(1, 1, 1);
#define threadIdx shim_threadIdx
#define blockIdx shim_blockIdx
#define blockDim shim_blockDim
#define gridDim shim_gridDim

/* the reference gets TRUE/FALSE from jpeglib.h (bvh_building.h:117,126) */
#ifndef TRUE
#define TRUE 1
#endif
#ifndef FALSE
#define FALSE 0
#endif

static inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
static inline float min(float a, float b) { return fminf(a, b); }
static inline float max(float a, float b) { return fmaxf(a, b); }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }

/* The reference printf()s progress and diagnostics from inside its kernels
 * (pathtracing.h:298-304 ...).  Silence them unless REF_VERBOSE is set. */
static inline int shim_printf(const char *fmt, ...) {
  static const bool verbose = getenv("REF_VERBOSE") != nullptr;
  if (!verbose) return 0;
  va_list ap;
  va_start(ap, fmt);
  int r = vprintf(fmt, ap);
  va_end(ap);
  return r;
}
#define printf(...) shim_printf(__VA_ARGS__)

/* the parity exports need Triangle's private derived data (tolerance, tangents) */
#define private public

#endif
