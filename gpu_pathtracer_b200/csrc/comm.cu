// comm.cu -- the collectives of the path, behind the C-ABI (include/gpt.h "multi-GPU").
//
// north_star: scene replicated, image tile-partitioned, SPPM photons split per GPU; per-hit-point
// photon flux and photon counts are combined each pass with ONE NCCL all-reduce over NVLink --
// between the traversal half and the update half of update_hit_point_parameters
// (src/model/bvh/bvh_traversal_photon.h:207-225, launched at src/main.cu:1780) -- and that is the
// only collective of the data path.  The end-of-frame tile gather (every rank's frame buffer is
// zero outside its own tiles, so a SUM reduce to the root assembles the image exactly) is output
// plumbing after the path.  The reference has no multi-device code at all.
//
// NCCL is bound at run time (dlopen "libnccl.so.2"), not at link time: inside a process that
// already carries an NCCL (PyTorch bundles its own) the same library instance is used, a plain
// C++ host gets the system one, and a single-GPU host needs none.  Every call is enqueued on the
// context's stream: no host synchronisation between a pass, its all-reduce and the update.
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

#include "common.cuh"

extern "C" int gpt_sppm_flux_device(gpt_ctx *ctx, void **d_flux, int64_t *num_floats);  // sppm.cu

namespace {

struct NcclApi {
  void *lib = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclReduce) Reduce = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclBroadcast) Broadcast = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
  bool ok = false;
};

NcclApi &nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
      api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (!api.lib) return;
#define GPT_SYM(f) api.f = reinterpret_cast<decltype(api.f)>(dlsym(api.lib, "nccl" #f))
    GPT_SYM(GetUniqueId); GPT_SYM(CommInitRank); GPT_SYM(CommDestroy); GPT_SYM(AllReduce);
    GPT_SYM(Reduce); GPT_SYM(GetErrorString); GPT_SYM(GetVersion);
    GPT_SYM(AllGather); GPT_SYM(Broadcast); GPT_SYM(GroupStart); GPT_SYM(GroupEnd);
#undef GPT_SYM
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.Reduce &&
             api.AllGather && api.Broadcast && api.GroupStart && api.GroupEnd;
  });
  return api;
}

int nccl_fail(gpt_ctx *ctx, ncclResult_t r, const char *file, int line) {
  NcclApi &a = nccl();
  return gpt_fail(ctx, GPT_E_NCCL, a.GetErrorString ? a.GetErrorString(r) : "NCCL error", file, line);
}
#define GPT_NCCL(ctx, call)                                              \
  do {                                                                   \
    ncclResult_t r__ = (call);                                           \
    if (r__ != ncclSuccess) return nccl_fail((ctx), r__, __FILE__, __LINE__); \
  } while (0)

}  // namespace

// ---- internal collectives of the sharded SPPM eye pass (sppm.cu; declared in common.cuh) -------
// In-place SUM all-reduce of n floats on the context's stream.
int gpt_comm_allreduce_f32(gpt_ctx *ctx, void *d_buf, size_t n) {
  if (!ctx->nccl_comm || ctx->comm_nranks == 1) return GPT_OK;
  GPT_NCCL(ctx, nccl().AllReduce(d_buf, d_buf, n, ncclFloat, ncclSum,
                                 static_cast<ncclComm_t>(ctx->nccl_comm), ctx->stream));
  return GPT_OK;
}

// Every rank's `mine` into counts[0 .. nranks) on the HOST (synchronises the stream).
int gpt_comm_allgather_count(gpt_ctx *ctx, int mine, int *d_scratch /* nranks ints */, int *counts) {
  const int R = ctx->comm_nranks, r = ctx->comm_rank;
  GPT_CUDA(ctx, cudaMemcpyAsync(d_scratch + r, &mine, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
  GPT_NCCL(ctx, nccl().AllGather(d_scratch + r, d_scratch, 1, ncclInt,
                                 static_cast<ncclComm_t>(ctx->nccl_comm), ctx->stream));
  GPT_CUDA(ctx, cudaMemcpyAsync(counts, d_scratch, sizeof(int) * R, cudaMemcpyDeviceToHost, ctx->stream));
  GPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return GPT_OK;
}

// Variable-length all-gather in rank order: rank r's counts[r] elements of elem_bytes bytes land
// behind those of the ranks before it in d_recv on every rank (one broadcast per rank, grouped).
int gpt_comm_allgatherv(gpt_ctx *ctx, const void *d_send, void *d_recv, const int *counts,
                        size_t elem_bytes) {
  const int R = ctx->comm_nranks, me = ctx->comm_rank;
  size_t off = 0, my_off = 0;
  for (int r = 0; r < me; r++) my_off += (size_t)counts[r] * elem_bytes;
  if (counts[me] > 0)
    GPT_CUDA(ctx, cudaMemcpyAsync(static_cast<char *>(d_recv) + my_off, d_send,
                                  (size_t)counts[me] * elem_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
  GPT_NCCL(ctx, nccl().GroupStart());
  for (int r = 0; r < R; r++) {
    const size_t bytes = (size_t)counts[r] * elem_bytes;
    if (bytes) {
      char *at = static_cast<char *>(d_recv) + off;
      ncclResult_t rc = nccl().Broadcast(at, at, bytes, ncclChar, r,
                                         static_cast<ncclComm_t>(ctx->nccl_comm), ctx->stream);
      if (rc != ncclSuccess) { nccl().GroupEnd(); return nccl_fail(ctx, rc, __FILE__, __LINE__); }
    }
    off += bytes;
  }
  GPT_NCCL(ctx, nccl().GroupEnd());
  return GPT_OK;
}

extern "C" {

int gpt_comm_unique_id(void *out_id) {
  if (!out_id) return GPT_E_INVALID;
  static_assert(sizeof(ncclUniqueId) == GPT_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size");
  NcclApi &a = nccl();
  if (!a.ok) return GPT_E_NCCL;
  ncclUniqueId id;
  if (a.GetUniqueId(&id) != ncclSuccess) return GPT_E_NCCL;
  memcpy(out_id, &id, sizeof(id));
  return GPT_OK;
}

int gpt_comm_init(gpt_ctx *ctx, const void *nccl_unique_id, int rank, int nranks) {
  if (!ctx || !nccl_unique_id || nranks < 1 || rank < 0 || rank >= nranks) return GPT_E_INVALID;
  NcclApi &a = nccl();
  if (!a.ok) return gpt_fail(ctx, GPT_E_NCCL, "libnccl.so.2 not found (dlopen)", __FILE__, __LINE__);
  if (ctx->nccl_comm) return gpt_fail(ctx, GPT_E_INVALID, "communicator already initialised", __FILE__, __LINE__);
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  ncclUniqueId id;
  memcpy(&id, nccl_unique_id, sizeof(id));
  ncclComm_t comm = nullptr;
  GPT_NCCL(ctx, a.CommInitRank(&comm, nranks, id, rank));
  ctx->nccl_comm = comm;
  ctx->comm_rank = rank;
  ctx->comm_nranks = nranks;
  return GPT_OK;
}

int gpt_comm_info(gpt_ctx *ctx, int *rank, int *nranks, int *nccl_version) {
  if (!ctx) return GPT_E_INVALID;
  if (rank) *rank = ctx->nccl_comm ? ctx->comm_rank : 0;
  if (nranks) *nranks = ctx->nccl_comm ? ctx->comm_nranks : 1;
  if (nccl_version) {
    *nccl_version = 0;
    NcclApi &a = nccl();
    if (a.ok && a.GetVersion) a.GetVersion(nccl_version);
  }
  return GPT_OK;
}

int gpt_comm_destroy(gpt_ctx *ctx) {
  if (!ctx) return GPT_E_INVALID;
  if (ctx->nccl_comm) {
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    nccl().CommDestroy(static_cast<ncclComm_t>(ctx->nccl_comm));
    ctx->nccl_comm = nullptr;
    ctx->comm_rank = 0;
    ctx->comm_nranks = 1;
  }
  return GPT_OK;
}

int gpt_sppm_allreduce(gpt_ctx *ctx) {
  if (!ctx) return GPT_E_INVALID;
  if (!ctx->nccl_comm || ctx->comm_nranks == 1) return GPT_OK;  // one rank: the sum is the buffer
  void *flux = nullptr;
  int64_t n = 0;
  GPT_TRY(gpt_sppm_flux_device(ctx, &flux, &n));
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  GPT_NCCL(ctx, nccl().AllReduce(flux, flux, (size_t)n, ncclFloat, ncclSum,
                                 static_cast<ncclComm_t>(ctx->nccl_comm), ctx->stream));
  return GPT_OK;
}

int gpt_gather_frame(gpt_ctx *ctx, void *d_rgb, int root) {
  if (!ctx || !d_rgb) return GPT_E_INVALID;
  if (!ctx->have_cam) return gpt_fail(ctx, GPT_E_INVALID, "gpt_set_camera first", __FILE__, __LINE__);
  if (!ctx->nccl_comm || ctx->comm_nranks == 1) return GPT_OK;
  if (root < 0 || root >= ctx->comm_nranks) return GPT_E_INVALID;
  GPT_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t n = 3 * (size_t)ctx->cam.width * ctx->cam.height;
  GPT_NCCL(ctx, nccl().Reduce(d_rgb, d_rgb, n, ncclFloat, ncclSum, root,
                              static_cast<ncclComm_t>(ctx->nccl_comm), ctx->stream));
  return GPT_OK;
}

}  // extern "C"
