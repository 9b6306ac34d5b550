// ctx.cu — context lifetime, error text, timing, pinned host memory and the optional NCCL
// communicator of libigpu_rollup.so.  See include/igpu_rollup.h for the contract of each export.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

static thread_local std::string g_last_global_error;

int32_t igpu_fail(igpu_ctx* ctx, int32_t status, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  else g_last_global_error = buf;
  return status;
}

extern "C" {

int32_t igpu_version(void) { return (IGPU_VERSION_MAJOR << 16) | (IGPU_VERSION_MINOR << 8) | IGPU_VERSION_PATCH; }

const char* igpu_strerror(int32_t status) {
  switch (status) {
    case IGPU_OK: return "ok";
    case IGPU_ERR_ARG: return "invalid argument";
    case IGPU_ERR_CUDA: return "CUDA runtime error";
    case IGPU_ERR_NCCL: return "NCCL unavailable or collective failed";
    case IGPU_ERR_OOM: return "out of memory";
    case IGPU_ERR_NODEVICE: return "no usable CUDA device (sm_100 required); there is no CPU fallback";
    case IGPU_ERR_RANGE: return "size or id beyond a documented limit";
    case IGPU_ERR_STATE: return "call order violated";
    default: return "unknown status";
  }
}

const char* igpu_last_error(const igpu_ctx* ctx) { return ctx ? ctx->err.c_str() : g_last_global_error.c_str(); }

int32_t igpu_ctx_create(int32_t device, void* stream, igpu_ctx** out) {
  if (!out) return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_ctx_create: out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return igpu_fail(nullptr, IGPU_ERR_NODEVICE, "no CUDA device: %s", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
  if (device < 0 || device >= n) return igpu_fail(nullptr, IGPU_ERR_ARG, "device %d out of range [0,%d)", device, n);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major < 10)
    return igpu_fail(nullptr, IGPU_ERR_NODEVICE, "device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
  igpu_ctx* ctx = new igpu_ctx();
  igpu_read_switches(ctx->sw);
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  DeviceGuard g(device);
  if (stream) {
    ctx->stream = static_cast<cudaStream_t>(stream);
  } else {
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
      delete ctx;
      return igpu_fail(nullptr, IGPU_ERR_CUDA, "cudaStreamCreate failed");
    }
    ctx->own_stream = true;
  }
  for (auto& ev : ctx->ev) cudaEventCreate(&ev);
  for (auto& ev : ctx->kev) cudaEventCreate(&ev);
  bool ok = cudaMalloc(&ctx->d_totals, IGPU_TOTALS_LEN * sizeof(int64_t)) == cudaSuccess &&
            cudaMalloc(&ctx->d_fleet, 2 * IGPU_FLEET_LEN * sizeof(double)) == cudaSuccess &&
            cudaMallocHost(&ctx->h_totals, IGPU_TOTALS_LEN * sizeof(int64_t)) == cudaSuccess &&
            cudaMallocHost(&ctx->h_fleet, 2 * IGPU_FLEET_LEN * sizeof(double)) == cudaSuccess;
  if (!ok) {
    igpu_ctx_destroy(ctx);
    return igpu_fail(nullptr, IGPU_ERR_OOM, "context scratch allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  *out = ctx;
  return IGPU_OK;
}

int32_t igpu_ctx_destroy(igpu_ctx* ctx) {
  if (!ctx) return IGPU_OK;
  DeviceGuard g(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->cached_fleet) igpu_fleet_destroy(ctx, ctx->cached_fleet);
  igpu_comm_destroy(ctx);
  for (auto& ev : ctx->ev)
    if (ev) cudaEventDestroy(ev);
  for (auto& ev : ctx->kev)
    if (ev) cudaEventDestroy(ev);
  if (ctx->d_totals) cudaFree(ctx->d_totals);
  if (ctx->d_fleet) cudaFree(ctx->d_fleet);
  if (ctx->d_pj) cudaFree(ctx->d_pj);
  if (ctx->h_totals) cudaFreeHost(ctx->h_totals);
  if (ctx->h_fleet) cudaFreeHost(ctx->h_fleet);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return IGPU_OK;
}

int32_t igpu_ctx_sync(igpu_ctx* ctx) {
  if (!ctx) return IGPU_ERR_ARG;
  DeviceGuard g(ctx->device);
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return IGPU_OK;
}

int32_t igpu_ctx_reload_env(igpu_ctx* ctx) {
  if (!ctx) return IGPU_ERR_ARG;
  igpu_read_switches(ctx->sw);
  return IGPU_OK;
}

int32_t igpu_get_timing(const igpu_ctx* cctx, igpu_timing* out) {
  if (!cctx || !out) return IGPU_ERR_ARG;
  igpu_ctx* ctx = const_cast<igpu_ctx*>(cctx);
  DeviceGuard g(ctx->device);
  auto span = [&](int a, int b) -> float {
    if (!ctx->ev_set[a] || !ctx->ev_set[b]) return 0.f;
    if (cudaEventSynchronize(ctx->ev[b]) != cudaSuccess) return 0.f;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->ev[a], ctx->ev[b]) != cudaSuccess) return 0.f;
    return ms;
  };
  ctx->timing.h2d_ms = span(0, 1);
  ctx->timing.kernel_ms = span(2, 3);
  ctx->timing.d2h_ms = span(4, 5);
  ctx->timing.total_ms = span(6, 7);
  for (int k = 0; k < 12; ++k) {
    float ms = 0.f;
    if (ctx->kev_set[k] && cudaEventSynchronize(ctx->kev[2 * k + 1]) == cudaSuccess &&
        cudaEventElapsedTime(&ms, ctx->kev[2 * k], ctx->kev[2 * k + 1]) == cudaSuccess)
      ctx->timing.kernel_detail_ms[k] = ms;
  }
  cudaGetLastError();
  *out = ctx->timing;
  return IGPU_OK;
}

int32_t igpu_host_alloc(uint64_t bytes, void** out) {
  if (!out) return IGPU_ERR_ARG;
  *out = nullptr;
  cudaError_t e = cudaMallocHost(out, bytes ? bytes : 1);
  if (e != cudaSuccess) return igpu_fail(nullptr, IGPU_ERR_OOM, "cudaMallocHost(%llu) failed: %s", (unsigned long long)bytes, cudaGetErrorString(e));
  return IGPU_OK;
}

int32_t igpu_host_free(void* p) {
  if (!p) return IGPU_OK;
  return cudaFreeHost(p) == cudaSuccess ? IGPU_OK : IGPU_ERR_CUDA;
}

// ---------------------------------------------------------------------------------------------
// NCCL through dlopen: the library has no link-time dependency on NCCL.
// ---------------------------------------------------------------------------------------------
namespace {
struct NcclUid {
  char internal[128];
};
struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(NcclUid*) = nullptr;
  int (*CommInitRank)(void**, int, NcclUid, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GroupStart)() = nullptr;  // optional: fuses the small all-reduces of one call into one launch
  int (*GroupEnd)() = nullptr;
  bool ok() const { return handle && GetUniqueId && CommInitRank && CommDestroy && AllReduce; }
};
NcclApi g_nccl;

bool nccl_load(std::string* why) {
  if (g_nccl.ok()) return true;
  const char* env = getenv("IGPU_NCCL_LIB");
  void* h = nullptr;
  if (env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // already in the process (e.g. torch's)
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    if (why) *why = std::string("dlopen(libnccl.so.2) failed: ") + (dlerror() ? dlerror() : "?");
    return false;
  }
  g_nccl.handle = h;
  g_nccl.GetUniqueId = reinterpret_cast<int (*)(NcclUid*)>(dlsym(h, "ncclGetUniqueId"));
  g_nccl.CommInitRank = reinterpret_cast<int (*)(void**, int, NcclUid, int)>(dlsym(h, "ncclCommInitRank"));
  g_nccl.CommDestroy = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclCommDestroy"));
  g_nccl.AllReduce =
      reinterpret_cast<int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t)>(dlsym(h, "ncclAllReduce"));
  g_nccl.GetErrorString = reinterpret_cast<const char* (*)(int)>(dlsym(h, "ncclGetErrorString"));
  g_nccl.GroupStart = reinterpret_cast<int (*)()>(dlsym(h, "ncclGroupStart"));
  g_nccl.GroupEnd = reinterpret_cast<int (*)()>(dlsym(h, "ncclGroupEnd"));
  if (!g_nccl.ok()) {
    if (why) *why = "libnccl is missing one of ncclGetUniqueId/ncclCommInitRank/ncclCommDestroy/ncclAllReduce";
    return false;
  }
  return true;
}
const char* nccl_err(int r) { return g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "nccl error"; }
}  // namespace

int32_t igpu_comm_unique_id(void* out128) {
  if (!out128) return IGPU_ERR_ARG;
  std::string why;
  if (!nccl_load(&why)) return igpu_fail(nullptr, IGPU_ERR_NCCL, "%s", why.c_str());
  NcclUid id;
  int r = g_nccl.GetUniqueId(&id);
  if (r != 0) return igpu_fail(nullptr, IGPU_ERR_NCCL, "ncclGetUniqueId: %s", nccl_err(r));
  memcpy(out128, id.internal, 128);
  return IGPU_OK;
}

int32_t igpu_comm_init(igpu_ctx* ctx, const void* id128, int32_t rank, int32_t nranks) {
  if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_comm_init: bad arguments");
  std::string why;
  if (!nccl_load(&why)) return igpu_fail(ctx, IGPU_ERR_NCCL, "%s", why.c_str());
  DeviceGuard g(ctx->device);
  igpu_comm_destroy(ctx);
  NcclUid id;
  memcpy(id.internal, id128, 128);
  void* comm = nullptr;
  int r = g_nccl.CommInitRank(&comm, nranks, id, rank);
  if (r != 0) return igpu_fail(ctx, IGPU_ERR_NCCL, "ncclCommInitRank: %s", nccl_err(r));
  ctx->comm = comm;
  ctx->rank = rank;
  ctx->nranks = nranks;
  return IGPU_OK;
}

int32_t igpu_comm_destroy(igpu_ctx* ctx) {
  if (!ctx) return IGPU_ERR_ARG;
  if (ctx->comm && g_nccl.ok()) {
    DeviceGuard g(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    g_nccl.CommDestroy(ctx->comm);
  }
  ctx->comm = nullptr;
  ctx->rank = 0;
  ctx->nranks = 1;
  return IGPU_OK;
}

}  // extern "C"

int32_t igpu_allreduce_i64_sum(igpu_ctx* ctx, int64_t* d_buf, size_t count) {
  if (!ctx->comm || ctx->nranks <= 1) return IGPU_OK;
  int r = g_nccl.AllReduce(d_buf, d_buf, count, /*ncclInt64*/ 4, /*ncclSum*/ 0, ctx->comm, ctx->stream);
  if (r != 0) return igpu_fail(ctx, IGPU_ERR_NCCL, "ncclAllReduce(int64,sum): %s", nccl_err(r));
  return IGPU_OK;
}

// Brackets several all-reduces of one call so NCCL issues them as one launch (no-ops without a communicator).
int32_t igpu_comm_group_begin(igpu_ctx* ctx) {
  if (!ctx->comm || ctx->nranks <= 1 || !g_nccl.GroupStart || !g_nccl.GroupEnd) return IGPU_OK;
  const int r = g_nccl.GroupStart();
  return r == 0 ? IGPU_OK : igpu_fail(ctx, IGPU_ERR_NCCL, "ncclGroupStart: %s", nccl_err(r));
}
int32_t igpu_comm_group_end(igpu_ctx* ctx) {
  if (!ctx->comm || ctx->nranks <= 1 || !g_nccl.GroupStart || !g_nccl.GroupEnd) return IGPU_OK;
  const int r = g_nccl.GroupEnd();
  return r == 0 ? IGPU_OK : igpu_fail(ctx, IGPU_ERR_NCCL, "ncclGroupEnd: %s", nccl_err(r));
}

int32_t igpu_allreduce_f64(igpu_ctx* ctx, double* d_buf, size_t count, int op) {
  if (!ctx->comm || ctx->nranks <= 1) return IGPU_OK;
  int r = g_nccl.AllReduce(d_buf, d_buf, count, /*ncclFloat64*/ 8, op, ctx->comm, ctx->stream);
  if (r != 0) return igpu_fail(ctx, IGPU_ERR_NCCL, "ncclAllReduce(f64,op %d): %s", op, nccl_err(r));
  return IGPU_OK;
}
