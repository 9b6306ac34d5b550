// common.cuh — context object, error plumbing and small device helpers shared by every kernel file
// of libigpu_rollup.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "../../include/igpu_rollup.h"

// Debugging / comparison / sweep switches (environment variables IGPU_*).  They are read ONCE, when the context is created
// (and again on igpu_ctx_reload_env), never on the call paths.  0 / unset = the tuned default.
struct IgpuSwitches {
  bool k1_legacy = false, k1_pod_parallel = false;          // IGPU_K1_LEGACY, IGPU_K1_POD_PARALLEL
  bool q_force_fallback = false;                            // IGPU_Q_FORCE_FALLBACK
  bool no_fused = false, no_wide = false, wide_no_l2pf = false;  // IGPU_NO_FUSED, IGPU_NO_WIDE, IGPU_WIDE_NO_L2PF
  bool rate_force_generic = false, ws_force_generic = false;     // IGPU_RATE_FORCE_GENERIC, IGPU_WS_FORCE_GENERIC
  int wide_ctas = 0, wide_stages = 0, wide_slots = 0;       // IGPU_WIDE_CTAS / _STAGES / _SLOTS
  int fused_ctas = 0, fused_stages = 0;                     // IGPU_FUSED_CTAS / _STAGES
  int rate_seg = 0, rate_grid = 0;                          // IGPU_RATE_SEG, IGPU_RATE_GRID
  bool fleet_quantile_exact = false;                        // IGPU_FLEET_QUANTILE_EXACT: skip the sampled bracket, three exact passes
  bool window_rollup_from_slab = false;                     // IGPU_WINDOW_ROLLUP_FROM_SLAB: never take the node planes from kept window planes
  int fleet_min_sample = 0;                                 // IGPU_FLEET_QUANTILE_MIN_SAMPLE: sampled values from which the bracket is used (tests: 1)
  int rate_tier = 0;                                        // IGPU_RATE_TIER: 0 by the slab's gap statistic, 1 streaming kernel, 2 gap kernel
};

struct igpu_ctx {
  IgpuSwitches sw;
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  // event pairs: [0]/[1] h2d, [2]/[3] kernels, [4]/[5] d2h, [6]/[7] whole one-shot call
  cudaEvent_t ev[8] = {};
  bool ev_set[8] = {};
  cudaEvent_t kev[24] = {};  // begin/end pair per IGPU_K_* kernel id
  bool kev_set[12] = {};
  igpu_timing timing = {};
  std::string err;
  igpu_fleet* cached_fleet = nullptr;  // grow-only staging for the one-shot call
  // communicator (NCCL through dlopen, see comm.cpp)
  void* comm = nullptr;
  int rank = 0, nranks = 1;
  // small device scratch: totals for reductions / collectives
  int64_t* d_totals = nullptr;   // IGPU_TOTALS_LEN
  double* d_fleet = nullptr;     // 2 * IGPU_FLEET_LEN (sum-part, min/max-part)
  int64_t* h_totals = nullptr;   // pinned mirrors
  double* h_fleet = nullptr;
  char* d_pj = nullptr;          // grow-only device arena of the instant power join (power_join.cu)
  size_t pj_bytes = 0;
};

int32_t igpu_fail(igpu_ctx* ctx, int32_t status, const char* fmt, ...);

#define IGPU_CUDA(ctx, call)                                                                       \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess)                                                                        \
      return igpu_fail((ctx), e__ == cudaErrorMemoryAllocation ? IGPU_ERR_OOM : IGPU_ERR_CUDA,     \
                       "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

#define IGPU_REQUIRE(ctx, cond, status, ...) \
  do {                                       \
    if (!(cond)) return igpu_fail((ctx), (status), __VA_ARGS__); \
  } while (0)

// Binds the calling thread to the context's device for the duration of a call.  `status` is what cudaSetDevice
// answered: a caller that takes the ordinal from outside (not from a context, whose device was checked when it was
// created) must look at it before it does anything on "the" device.
struct DeviceGuard {
  int prev = -1;
  cudaError_t status = cudaSuccess;
  explicit DeviceGuard(int dev) {
    status = cudaGetDevice(&prev);
    if (status != cudaSuccess) prev = -1;
    else if (prev != dev) {
      status = cudaSetDevice(dev);
      if (status != cudaSuccess) prev = -1;  // nothing was changed, nothing to restore
    } else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

inline void igpu_mark(igpu_ctx* ctx, int i) {
  cudaEventRecord(ctx->ev[i], ctx->stream);
  ctx->ev_set[i] = true;
}
inline void igpu_kbegin(igpu_ctx* ctx, int k) { cudaEventRecord(ctx->kev[2 * k], ctx->stream); }
inline void igpu_kend(igpu_ctx* ctx, int k) {
  cudaEventRecord(ctx->kev[2 * k + 1], ctx->stream);
  ctx->kev_set[k] = true;
}
inline void igpu_count_launch(igpu_ctx* ctx, unsigned n = 1) { ctx->timing.kernel_launches += n; }

// Collective on the context's stream over its communicator (comm.cpp); no-ops when nranks == 1.
int32_t igpu_allreduce_i64_sum(igpu_ctx* ctx, int64_t* d_buf, size_t count);
int32_t igpu_allreduce_f64(igpu_ctx* ctx, double* d_buf, size_t count, int op /*0 sum, 2 max, 3 min*/);
int32_t igpu_comm_group_begin(igpu_ctx* ctx);  // several all-reduces between begin/end go out as one NCCL launch
int32_t igpu_comm_group_end(igpu_ctx* ctx);

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
// debugging / comparison switches: set and neither empty nor "0"
static inline bool igpu_env_on(const char* name) {
  const char* v = getenv(name);
  return v && *v && !(v[0] == '0' && !v[1]);
}
static inline int igpu_env_int(const char* name) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : 0;
}
static inline void igpu_read_switches(IgpuSwitches& sw) {
  sw.k1_legacy = igpu_env_on("IGPU_K1_LEGACY"), sw.k1_pod_parallel = igpu_env_on("IGPU_K1_POD_PARALLEL");
  sw.q_force_fallback = igpu_env_on("IGPU_Q_FORCE_FALLBACK");
  sw.no_fused = igpu_env_on("IGPU_NO_FUSED"), sw.no_wide = igpu_env_on("IGPU_NO_WIDE"), sw.wide_no_l2pf = igpu_env_on("IGPU_WIDE_NO_L2PF");
  sw.rate_force_generic = igpu_env_on("IGPU_RATE_FORCE_GENERIC"), sw.ws_force_generic = igpu_env_on("IGPU_WS_FORCE_GENERIC");
  sw.wide_ctas = igpu_env_int("IGPU_WIDE_CTAS"), sw.wide_stages = igpu_env_int("IGPU_WIDE_STAGES"), sw.wide_slots = igpu_env_int("IGPU_WIDE_SLOTS");
  sw.fused_ctas = igpu_env_int("IGPU_FUSED_CTAS"), sw.fused_stages = igpu_env_int("IGPU_FUSED_STAGES");
  sw.rate_seg = igpu_env_int("IGPU_RATE_SEG"), sw.rate_grid = igpu_env_int("IGPU_RATE_GRID");
  sw.rate_tier = igpu_env_int("IGPU_RATE_TIER");
  sw.window_rollup_from_slab = igpu_env_on("IGPU_WINDOW_ROLLUP_FROM_SLAB");
  sw.fleet_quantile_exact = igpu_env_on("IGPU_FLEET_QUANTILE_EXACT");
  sw.fleet_min_sample = igpu_env_int("IGPU_FLEET_QUANTILE_MIN_SAMPLE");
}

#ifdef __CUDACC__
// ---- device helpers -------------------------------------------------------------------------

// Math.round on the device: nearest integer, ties toward +Infinity (IEEE-exact double ops).
__device__ __forceinline__ double js_round_dev(double x) {
  double f = floor(x);
  return (x - f) >= 0.5 ? f + 1.0 : f;
}

__device__ __forceinline__ long long warp_sum_i64(long long v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Fire-and-forget global adds, spelled in PTX: through atomicAdd() nvcc wraps every call in its warp-aggregation
// sequence (match / ballot / shuffle-reduce, ~13 instructions) in case lanes share an address — ncu showed 14 % of K1's
// instructions there, for adds that are already aggregated by hand or go to distinct nodes.
__device__ __forceinline__ void red_add_i64(int64_t* p, long long v) {
  asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// the same for a shared-memory histogram bin (p must point into shared memory)
__device__ __forceinline__ void smem_red_add_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(p))), "r"(v) : "memory");
}
__device__ __forceinline__ void smem_red_or_b32(uint32_t* p, uint32_t v) {
  asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(p))), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// mbarrier + bulk-copy (TMA 1-D) wrappers, sm_90+ PTX.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy; completes `bytes` transaction bytes on `bar`. 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// asks the memory system to bring [src, src+bytes) into L2 (no shared memory, no completion to wait for)
__device__ __forceinline__ void bulk_prefetch_l2(const void* src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}
// streaming 128-bit load that does not allocate in L1
__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
// order-preserving map float32 -> uint32 (total order on non-NaN values, -0 < +0)
__device__ __forceinline__ uint32_t f32_key(float f) {
  uint32_t u = __float_as_uint(f);
  return u ^ (static_cast<uint32_t>(static_cast<int32_t>(u) >> 31) | 0x80000000u);
}
__device__ __forceinline__ float key_f32(uint32_t k) {
  uint32_t u = (k & 0x80000000u) ? (k ^ 0x80000000u) : ~k;
  return __uint_as_float(u);
}
#endif
