// comm.cu -- multi-GPU reduction of the per-shard step results.
//
// Cells shard across GPUs by row bands (one process per GPU); the only exchange on the path is
// the argmax / safe-count reduction.  Each rank contributes one 56-byte record
// (ucb value+index, width value+index, safe / clamped / low-sigma^2 counts); a single ncclAllGather moves
// the records over NVLink and a one-warp kernel folds them with the library's total order
// (value desc, index asc).  fp64 values do not fit a packed 64-bit max-key, hence gather+fold
// rather than ncclMax on a key.
//
// NCCL is bound at run time (dlopen of libnccl.so.2 -- the copy PyTorch already loaded when the
// host is a torch.distributed process), so libsbo.so has no link-time dependency on it.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace sbo {
namespace {

typedef struct { char internal[128]; } NcclUniqueId;
typedef int (*GetUniqueIdFn)(NcclUniqueId*);
typedef int (*CommInitRankFn)(void**, int, NcclUniqueId, int);
typedef int (*AllGatherFn)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*BroadcastFn)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*CommDestroyFn)(void*);
typedef const char* (*GetErrorStringFn)(int);

struct NcclApi {
  void* handle = nullptr;
  GetUniqueIdFn get_unique_id = nullptr;
  CommInitRankFn comm_init_rank = nullptr;
  AllGatherFn all_gather = nullptr;
  BroadcastFn broadcast = nullptr;
  CommDestroyFn comm_destroy = nullptr;
  GetErrorStringFn get_error_string = nullptr;
  std::string err;
};

NcclApi& nccl() {
  static NcclApi api;
  if (api.handle || !api.err.empty()) return api;
  const char* env = std::getenv("SBO_NCCL_LIB");
  const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    if (!nm) continue;
    api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (api.handle) break;
  }
  if (!api.handle) {
    const char* e = dlerror();   // (one call: dlerror clears the state it returns)
    api.err = std::string("cannot dlopen libnccl.so.2 (set SBO_NCCL_LIB): ") + (e ? e : "");
    return api;
  }
  api.get_unique_id = (GetUniqueIdFn)dlsym(api.handle, "ncclGetUniqueId");
  api.comm_init_rank = (CommInitRankFn)dlsym(api.handle, "ncclCommInitRank");
  api.all_gather = (AllGatherFn)dlsym(api.handle, "ncclAllGather");
  api.broadcast = (BroadcastFn)dlsym(api.handle, "ncclBroadcast");
  api.comm_destroy = (CommDestroyFn)dlsym(api.handle, "ncclCommDestroy");
  api.get_error_string = (GetErrorStringFn)dlsym(api.handle, "ncclGetErrorString");
  if (!api.get_unique_id || !api.comm_init_rank || !api.all_gather || !api.broadcast || !api.comm_destroy) {
    api.err = "libnccl is missing a required symbol";
    api.handle = nullptr;
  }
  return api;
}

__global__ void fold_records_kernel(const SetsResult* __restrict__ parts, int count, SetsResult* __restrict__ out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  SetsResult r = parts[0];
  for (int i = 1; i < count; ++i) {
    const SetsResult o = parts[i];
    if (better(o.ucb_val, o.ucb_idx, r.ucb_val, r.ucb_idx)) { r.ucb_val = o.ucb_val; r.ucb_idx = o.ucb_idx; }
    if (better(o.width_val, o.width_idx, r.width_val, r.width_idx)) { r.width_val = o.width_val; r.width_idx = o.width_idx; }
    r.safe_count += o.safe_count;
    r.clamped += o.clamped;
    r.low_sigma2 += o.low_sigma2;
  }
  *out = r;
}

}  // namespace

int32_t comm_gather_reduce(sbo_ctx* ctx) {
  if (!ctx->nccl_comm || ctx->world <= 1) return SBO_OK;
  NcclApi& api = nccl();
  const int rc = api.all_gather(ctx->d_result, ctx->d_gather, sizeof(SetsResult), /*ncclInt8*/ 0, ctx->nccl_comm,
                                ctx->stream);
  if (rc != 0)
    return fail(ctx, SBO_ERR_COMM, std::string("ncclAllGather: ") + (api.get_error_string ? api.get_error_string(rc) : "?"));
  ctx->launches++;  // NCCL's kernel
  fold_records_kernel<<<1, 32, 0, ctx->stream>>>(ctx->d_gather, ctx->world, ctx->d_result);
  SBO_LAUNCH_CHECK(ctx, "fold_records_kernel");
  return SBO_OK;
}

// The two collectives of the shared fit (factor_dist.cuh): panel k from its owner to every rank, and the block columns of
// L^-1 from every rank to every rank.  Byte counts (ncclInt8), in stream order like every NCCL call.
int32_t comm_broadcast(sbo_ctx* ctx, void* buf, size_t bytes, int root, cudaStream_t stream) {
  NcclApi& api = nccl();
  if (!ctx->nccl_comm || !api.handle) return fail(ctx, SBO_ERR_COMM, "comm_broadcast: no communicator");
  const int rc = api.broadcast(buf, buf, bytes, /*ncclInt8*/ 0, root, ctx->nccl_comm, stream);
  if (rc != 0)
    return fail(ctx, SBO_ERR_COMM, std::string("ncclBroadcast: ") + (api.get_error_string ? api.get_error_string(rc) : "?"));
  ctx->launches++;
  return SBO_OK;
}

int32_t comm_all_gather(sbo_ctx* ctx, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t stream) {
  NcclApi& api = nccl();
  if (!ctx->nccl_comm || !api.handle) return fail(ctx, SBO_ERR_COMM, "comm_all_gather: no communicator");
  const int rc = api.all_gather(send, recv, bytes_per_rank, /*ncclInt8*/ 0, ctx->nccl_comm, stream);
  if (rc != 0)
    return fail(ctx, SBO_ERR_COMM, std::string("ncclAllGather: ") + (api.get_error_string ? api.get_error_string(rc) : "?"));
  ctx->launches++;
  return SBO_OK;
}

}  // namespace sbo

extern "C" {

int32_t sbo_comm_unique_id(uint8_t id_out[128]) {
  if (!id_out) return SBO_ERR_INVALID_ARG;
  sbo::NcclApi& api = sbo::nccl();
  if (!api.handle) return SBO_ERR_COMM;
  sbo::NcclUniqueId id;
  if (api.get_unique_id(&id) != 0) return SBO_ERR_COMM;
  std::memcpy(id_out, id.internal, 128);
  return SBO_OK;
}

int32_t sbo_comm_init(sbo_ctx* ctx, const uint8_t id[128], int32_t rank, int32_t world) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (!id || world < 1 || rank < 0 || rank >= world) return sbo::fail(ctx, SBO_ERR_INVALID_ARG, "sbo_comm_init: bad rank/world/id");
  sbo::NcclApi& api = sbo::nccl();
  if (!api.handle) return sbo::fail(ctx, SBO_ERR_COMM, api.err);
  if (ctx->nccl_comm) sbo_comm_destroy(ctx);
  SBO_CUDA(ctx, cudaSetDevice(ctx->device));
  sbo::NcclUniqueId uid;
  std::memcpy(uid.internal, id, 128);
  void* comm = nullptr;
  const int rc = api.comm_init_rank(&comm, world, uid, rank);
  if (rc != 0)
    return sbo::fail(ctx, SBO_ERR_COMM, std::string("ncclCommInitRank: ") + (api.get_error_string ? api.get_error_string(rc) : "?"));
  ctx->nccl_comm = comm;
  ctx->rank = rank;
  ctx->world = world;
  SBO_CUDA(ctx, cudaMalloc(&ctx->d_gather, sizeof(sbo::SetsResult) * world));
  return SBO_OK;
}

int32_t sbo_comm_destroy(sbo_ctx* ctx) {
  if (!ctx) return SBO_ERR_INVALID_ARG;
  if (ctx->nccl_comm) {
    sbo::nccl().comm_destroy(ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
  }
  if (ctx->d_gather) { cudaFree(ctx->d_gather); ctx->d_gather = nullptr; }
  ctx->rank = 0;
  ctx->world = 1;
  return SBO_OK;
}

}  // extern "C"
