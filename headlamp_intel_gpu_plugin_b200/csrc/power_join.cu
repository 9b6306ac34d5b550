// power_join.cu — K2: the instant-vector join of fetchGpuMetrics and the fleet power sums.
//
// Reference arithmetic replaced:
//   three Map builds + per-chip assemble      src/api/metrics.ts:118-149
//   chips.reduce sums + PowerBar percentage   src/components/MetricsPage.tsx:330-332, :50-53
// chip / instance / node-name strings arrive as dictionary ids; value columns hold
// parseFloat(value[1]) (metrics.ts:130,137), both done by the host marshal layer (metrics.py).
//
// Map.set semantics — the LAST row carrying a key wins (metrics.ts:123,130,137; the chip maps are
// keyed by the chip label only) — become an atomicMax on (row index + 1) per dictionary id; the
// assemble step is then a pure gather.  Fleet sums: per-CTA partials in float64, combined by one
// CTA in a fixed order (deterministic; the reference adds left to right, the difference is below
// 1e-15 relative and the contract is 1e-5).  `?? 0` only replaces null, so a NaN sample poisons
// the sum exactly as it does in the reference.
#include <algorithm>
#include <cmath>
#include <mutex>

#include "common.cuh"

namespace {

constexpr int PJ_BLOCKS = 148;

__global__ void __launch_bounds__(256) pj_last_writer_kernel(const uint32_t* __restrict__ key, uint32_t n, uint32_t n_ids,
                                                             uint32_t* __restrict__ row_plus1) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t k = key[i];
    if (k != IGPU_NODE_NONE && k < n_ids) atomicMax(row_plus1 + k, i + 1);
  }
}

struct PjParams {
  uint32_t n_chips, n_chip_ids, n_instance_ids;
  const uint32_t* __restrict__ chip_chip;
  const uint32_t* __restrict__ chip_instance;
  const double* __restrict__ energy_value;
  const double* __restrict__ power_max_value;
  const uint32_t* __restrict__ uname_node;
  const uint32_t* __restrict__ pw_row;
  const uint32_t* __restrict__ mx_row;
  const uint32_t* __restrict__ inst_row;
  uint8_t* __restrict__ chip_flags;
  double* __restrict__ chip_watts;
  double* __restrict__ chip_max;
  uint32_t* __restrict__ chip_node;
  double* __restrict__ partial;  // [PJ_BLOCKS][2]
};

__global__ void __launch_bounds__(256) pj_assemble_kernel(const PjParams p) {
  __shared__ double s_w[8], s_m[8];
  double w_sum = 0, m_sum = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < p.n_chips; i += gridDim.x * blockDim.x) {
    const uint32_t chip = p.chip_chip[i], inst = p.chip_instance[i];
    uint32_t fl = 0, node = IGPU_NODE_NONE;
    double w = 0, m = 0;
    if (chip != IGPU_NODE_NONE && chip < p.n_chip_ids) {
      const uint32_t a = p.pw_row[chip], b = p.mx_row[chip];
      if (a) fl |= IGPU_POWER_HAS_WATTS, w = p.energy_value[a - 1];      // chipToPower.has(chip), metrics.ts:145
      if (b) fl |= IGPU_POWER_HAS_MAX, m = p.power_max_value[b - 1];     // metrics.ts:146
    }
    if (inst != IGPU_NODE_NONE && inst < p.n_instance_ids) {
      const uint32_t c = p.inst_row[inst];
      if (c) fl |= IGPU_POWER_HAS_NODE, node = p.uname_node[c - 1];      // metrics.ts:144
    }
    p.chip_flags[i] = (uint8_t)fl;
    p.chip_watts[i] = w;
    p.chip_max[i] = m;
    p.chip_node[i] = node;
    w_sum += w;  // absent => 0 (`?? 0`), NaN stays NaN
    m_sum += m;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    w_sum += __shfl_xor_sync(0xffffffffu, w_sum, o);
    m_sum += __shfl_xor_sync(0xffffffffu, m_sum, o);
  }
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = w_sum, s_m[threadIdx.x >> 5] = m_sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < 8; ++i) a += s_w[i], b += s_m[i];
    p.partial[2 * blockIdx.x] = a;
    p.partial[2 * blockIdx.x + 1] = b;
  }
}

__global__ void pj_final_kernel(const double* __restrict__ partial, int n_blocks, uint32_t n_chips, double* __restrict__ fleet) {
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < n_blocks; ++i) a += partial[2 * i], b += partial[2 * i + 1];
    fleet[0] = a;
    fleet[1] = b;
    fleet[2] = (double)n_chips;
    fleet[3] = 0;
  }
}

template <class T>
T* carve(char*& cur, size_t count) {
  size_t off = align_up(reinterpret_cast<size_t>(cur), 256);
  cur = reinterpret_cast<char*>(off + count * sizeof(T));
  return reinterpret_cast<T*>(off);
}

}  // namespace

extern "C" int32_t igpu_power_join_sum(igpu_ctx* ctx, const igpu_power_cols* c, const igpu_power_out* o) {
  if (!ctx || !c || !o) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_power_join_sum: NULL argument");
  IGPU_REQUIRE(ctx, (c->n_chips == 0 || (c->chip_chip && c->chip_instance)) && (c->n_energy == 0 || (c->energy_chip && c->energy_value)) &&
                        (c->n_power_max == 0 || (c->power_max_chip && c->power_max_value)) &&
                        (c->n_uname == 0 || (c->uname_instance && c->uname_node)),
               IGPU_ERR_ARG, "a required power column is NULL");
  DeviceGuard g(ctx->device);
  const size_t nc = c->n_chips, ne = c->n_energy, nm = c->n_power_max, nu = c->n_uname, ci = c->n_chip_ids, ii = c->n_instance_ids;
  const size_t need = 256 * 20 + nc * (4 + 4 + 1 + 8 + 8 + 4) + ne * 12 + nm * 12 + nu * 8 + (2 * ci + ii) * 4 + PJ_BLOCKS * 16 + 64;
  if (need > ctx->pj_bytes) {
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->d_pj) cudaFree(ctx->d_pj);
    ctx->d_pj = nullptr;
    ctx->pj_bytes = 0;
    IGPU_CUDA(ctx, cudaMalloc(&ctx->d_pj, need * 2));
    ctx->pj_bytes = need * 2;
  }
  char* cur = ctx->d_pj;
  uint32_t* d_chip_chip = carve<uint32_t>(cur, nc);
  uint32_t* d_chip_inst = carve<uint32_t>(cur, nc);
  uint32_t* d_en_chip = carve<uint32_t>(cur, ne);
  double* d_en_val = carve<double>(cur, ne);
  uint32_t* d_mx_chip = carve<uint32_t>(cur, nm);
  double* d_mx_val = carve<double>(cur, nm);
  uint32_t* d_un_inst = carve<uint32_t>(cur, nu);
  uint32_t* d_un_node = carve<uint32_t>(cur, nu);
  char* tables_begin = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(cur), 256));
  uint32_t* d_pw_row = carve<uint32_t>(cur, ci);
  uint32_t* d_mx_row = carve<uint32_t>(cur, ci);
  uint32_t* d_inst_row = carve<uint32_t>(cur, ii);
  char* tables_end = cur;
  uint8_t* d_flags = carve<uint8_t>(cur, nc);
  double* d_watts = carve<double>(cur, nc);
  double* d_max = carve<double>(cur, nc);
  uint32_t* d_node = carve<uint32_t>(cur, nc);
  double* d_partial = carve<double>(cur, PJ_BLOCKS * 2);
  double* d_fleet = carve<double>(cur, 8);

  igpu_mark(ctx, 6);
  igpu_mark(ctx, 0);
  uint64_t up_bytes = 0;
  auto up = [&](void* dst, const void* src, size_t n) -> cudaError_t {
    if (!n) return cudaSuccess;
    up_bytes += n;
    return cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, ctx->stream);
  };
  IGPU_CUDA(ctx, up(d_chip_chip, c->chip_chip, nc * 4));
  IGPU_CUDA(ctx, up(d_chip_inst, c->chip_instance, nc * 4));
  IGPU_CUDA(ctx, up(d_en_chip, c->energy_chip, ne * 4));
  IGPU_CUDA(ctx, up(d_en_val, c->energy_value, ne * 8));
  IGPU_CUDA(ctx, up(d_mx_chip, c->power_max_chip, nm * 4));
  IGPU_CUDA(ctx, up(d_mx_val, c->power_max_value, nm * 8));
  IGPU_CUDA(ctx, up(d_un_inst, c->uname_instance, nu * 4));
  IGPU_CUDA(ctx, up(d_un_node, c->uname_node, nu * 4));
  igpu_mark(ctx, 1);
  ctx->timing.h2d_bytes = up_bytes;

  igpu_mark(ctx, 2);
  igpu_kbegin(ctx, IGPU_K_POWER_JOIN);
  IGPU_CUDA(ctx, cudaMemsetAsync(tables_begin, 0, tables_end - tables_begin, ctx->stream));
  auto blocks = [&](size_t n) { return (unsigned)std::max<size_t>(1, std::min<size_t>((n + 255) / 256, (size_t)ctx->sm_count * 4)); };
  unsigned launches = 0;
  if (ne && ci) pj_last_writer_kernel<<<blocks(ne), 256, 0, ctx->stream>>>(d_en_chip, (uint32_t)ne, (uint32_t)ci, d_pw_row), ++launches;
  if (nm && ci) pj_last_writer_kernel<<<blocks(nm), 256, 0, ctx->stream>>>(d_mx_chip, (uint32_t)nm, (uint32_t)ci, d_mx_row), ++launches;
  if (nu && ii) pj_last_writer_kernel<<<blocks(nu), 256, 0, ctx->stream>>>(d_un_inst, (uint32_t)nu, (uint32_t)ii, d_inst_row), ++launches;
  PjParams p;
  p.n_chips = (uint32_t)nc;
  p.n_chip_ids = (uint32_t)ci;
  p.n_instance_ids = (uint32_t)ii;
  p.chip_chip = d_chip_chip;
  p.chip_instance = d_chip_inst;
  p.energy_value = d_en_val;
  p.power_max_value = d_mx_val;
  p.uname_node = d_un_node;
  p.pw_row = d_pw_row;
  p.mx_row = d_mx_row;
  p.inst_row = d_inst_row;
  p.chip_flags = d_flags;
  p.chip_watts = d_watts;
  p.chip_max = d_max;
  p.chip_node = d_node;
  p.partial = d_partial;
  pj_assemble_kernel<<<PJ_BLOCKS, 256, 0, ctx->stream>>>(p);
  pj_final_kernel<<<1, 32, 0, ctx->stream>>>(d_partial, PJ_BLOCKS, (uint32_t)nc, d_fleet);
  igpu_kend(ctx, IGPU_K_POWER_JOIN);
  IGPU_CUDA(ctx, cudaGetLastError());
  igpu_count_launch(ctx, launches + 2);
  int32_t st = igpu_allreduce_f64(ctx, d_fleet, 3, /*sum*/ 0);
  if (st != IGPU_OK) return st;
  igpu_mark(ctx, 3);

  igpu_mark(ctx, 4);
  uint64_t down_bytes = 0;
  auto down = [&](void* dst, const void* src, size_t n) -> cudaError_t {
    if (!dst || !n) return cudaSuccess;
    down_bytes += n;
    return cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, ctx->stream);
  };
  IGPU_CUDA(ctx, down(o->chip_flags, d_flags, nc));
  IGPU_CUDA(ctx, down(o->chip_watts, d_watts, nc * 8));
  IGPU_CUDA(ctx, down(o->chip_max_watts, d_max, nc * 8));
  IGPU_CUDA(ctx, down(o->chip_node, d_node, nc * 4));
  IGPU_CUDA(ctx, down(ctx->h_fleet, d_fleet, 4 * 8));
  igpu_mark(ctx, 5);
  igpu_mark(ctx, 7);
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->timing.d2h_bytes = down_bytes;
  if (o->fleet) {
    const double total = ctx->h_fleet[0], max_total = ctx->h_fleet[1];
    o->fleet[0] = total;
    o->fleet[1] = max_total;
    // PowerBar (MetricsPage.tsx:50-53): null unless maxTotal > 0; Math.min(100, Math.round(x)) keeps NaN
    double pct = NAN;
    if (max_total > 0) {
      const double x = total / max_total * 100.0;
      if (!std::isnan(x)) {
        const double f = std::floor(x);
        pct = std::fmin(100.0, (x - f) >= 0.5 ? f + 1.0 : f);
      }
    }
    o->fleet[2] = pct;
    o->fleet[3] = ctx->h_fleet[2];
  }
  return IGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// igpu_power_join_sum_static — the same join for callers that must not allocate.
//
// Who that is: a host program that has handed the whole device to its own memory manager (an
// inference server that reserves all of HBM at start-up, so a later cudaMalloc simply fails); an
// exporter that runs this join inside a periodic scrape loop next to other streams, where the
// device-wide synchronisation hidden in cudaMalloc / cudaFree would stall its neighbours; or any
// program whose rule is "only our allocator touches device memory" and which therefore passes this
// library device pointers.  For them the intermediate tables of the join (one "last row that wrote
// this id" word per chip id for power, per chip id for TDP and per instance id, plus the per-CTA
// partial sums) live in a fixed array inside the library image instead of a cudaMalloc'ed arena.
// ---------------------------------------------------------------------------------------------
namespace {

constexpr size_t PJ_STATIC_TABLE_BYTES = (size_t)3 * IGPU_STATIC_JOIN_MAX_IDS * sizeof(uint32_t);
constexpr size_t PJ_STATIC_PARTIAL_OFF = PJ_STATIC_TABLE_BYTES;  // multiple of 256
constexpr size_t PJ_STATIC_BYTES = PJ_STATIC_PARTIAL_OFF + (size_t)PJ_BLOCKS * 2 * sizeof(double);

// Lives in every GPU that loads this library, until the process exits (3 MB + 2.3 KB).
static __device__ __align__(256) unsigned char g_pj_static_scratch[PJ_STATIC_BYTES];

// In host code `g_pj_static_scratch` is only a handle for the symbol, not an address a kernel could
// use: the device address is asked for once per device and cached here.
constexpr int PJ_MAX_DEVICES = 64;
void* g_pj_static_addr[PJ_MAX_DEVICES] = {};
std::mutex g_pj_static_mu;

// [sum watts, sum tdp, n chips, 0]  ->  [sum watts, sum tdp, PowerBar pct (NaN = none), n chips]
__global__ void pj_static_pct_kernel(double* __restrict__ fleet) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const double total = fleet[0], max_total = fleet[1], n = fleet[2];
    double pct = __longlong_as_double(0x7ff8000000000000LL);
    if (max_total > 0) {  // MetricsPage.tsx:51,331
      const double x = total / max_total * 100.0;
      if (x == x) {
        const double f = floor(x);
        pct = fmin(100.0, (x - f) >= 0.5 ? f + 1.0 : f);
      }
    }
    fleet[2] = pct;
    fleet[3] = n;
  }
}

}  // namespace

extern "C" int32_t igpu_power_join_sum_static(int32_t device, void* stream, const igpu_power_cols* c, const igpu_power_out* o) {
  if (!c || !o) return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_power_join_sum_static: NULL argument");
  if (device < 0 || device >= PJ_MAX_DEVICES) return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_power_join_sum_static: device %d out of range", device);
  if (c->n_chip_ids > IGPU_STATIC_JOIN_MAX_IDS || c->n_instance_ids > IGPU_STATIC_JOIN_MAX_IDS)
    return igpu_fail(nullptr, IGPU_ERR_RANGE,
                     "igpu_power_join_sum_static: %u chip ids / %u instance ids exceed the built-in scratch limit of %u ids each; "
                     "use igpu_power_join_sum (which allocates) for larger fleets",
                     c->n_chip_ids, c->n_instance_ids, (unsigned)IGPU_STATIC_JOIN_MAX_IDS);
  if (!o->chip_flags || !o->chip_watts || !o->chip_max_watts || !o->chip_node || !o->fleet)
    return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_power_join_sum_static: every output pointer must be a device buffer");
  if ((c->n_chips && (!c->chip_chip || !c->chip_instance)) || (c->n_energy && (!c->energy_chip || !c->energy_value)) ||
      (c->n_power_max && (!c->power_max_chip || !c->power_max_value)) || (c->n_uname && (!c->uname_instance || !c->uname_node)))
    return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_power_join_sum_static: a required input column is NULL");
  int n_devices = 0;
  if (cudaGetDeviceCount(&n_devices) != cudaSuccess || n_devices <= 0) {
    (void)cudaGetLastError();
    return igpu_fail(nullptr, IGPU_ERR_NODEVICE, "igpu_power_join_sum_static: no CUDA device");
  }
  if (device >= n_devices)
    return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_power_join_sum_static: device %d, but this machine has %d", device, n_devices);
  DeviceGuard g(device);
  if (g.status != cudaSuccess) {  // nothing may be looked up, cached or launched on whatever device is current instead
    (void)cudaGetLastError();
    return igpu_fail(nullptr, IGPU_ERR_CUDA, "igpu_power_join_sum_static: cannot select device %d: %s", device, cudaGetErrorString(g.status));
  }
  unsigned char* base = nullptr;
  {  // one-time set-up per device, with that device current
    std::lock_guard<std::mutex> lk(g_pj_static_mu);
    if (!g_pj_static_addr[device]) {
      void* ptr = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&ptr, g_pj_static_scratch);
      if (e != cudaSuccess || !ptr)
        return igpu_fail(nullptr, IGPU_ERR_CUDA, "cudaGetSymbolAddress(static join scratch) on device %d failed: %s", device,
                         cudaGetErrorString(e));
      g_pj_static_addr[device] = ptr;
    }
    base = static_cast<unsigned char*>(g_pj_static_addr[device]);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const uint32_t nc = c->n_chips, ne = c->n_energy, nm = c->n_power_max, nu = c->n_uname, ci = c->n_chip_ids, ii = c->n_instance_ids;
  uint32_t* pw_row = reinterpret_cast<uint32_t*>(base);
  uint32_t* mx_row = pw_row + IGPU_STATIC_JOIN_MAX_IDS;
  uint32_t* inst_row = mx_row + IGPU_STATIC_JOIN_MAX_IDS;
  double* partial = reinterpret_cast<double*>(base + PJ_STATIC_PARTIAL_OFF);
  // The array still holds the previous call's tables: the words this call will read must start at zero.
  // (The partial sums need no clearing: every one of the PJ_BLOCKS CTAs overwrites its two slots.)
  cudaError_t e = cudaSuccess;
  if (ci) e = cudaMemsetAsync(pw_row, 0, (size_t)ci * 4, st);
  if (e == cudaSuccess && ci) e = cudaMemsetAsync(mx_row, 0, (size_t)ci * 4, st);
  if (e == cudaSuccess && ii) e = cudaMemsetAsync(inst_row, 0, (size_t)ii * 4, st);
  if (e != cudaSuccess) return igpu_fail(nullptr, IGPU_ERR_CUDA, "cudaMemsetAsync(static join scratch) failed: %s", cudaGetErrorString(e));
  auto blocks = [&](uint32_t n) { return (unsigned)std::max<uint32_t>(1, std::min<uint32_t>((n + 255) / 256, 148u * 4)); };
  if (ne && ci) pj_last_writer_kernel<<<blocks(ne), 256, 0, st>>>(c->energy_chip, ne, ci, pw_row);
  if (nm && ci) pj_last_writer_kernel<<<blocks(nm), 256, 0, st>>>(c->power_max_chip, nm, ci, mx_row);
  if (nu && ii) pj_last_writer_kernel<<<blocks(nu), 256, 0, st>>>(c->uname_instance, nu, ii, inst_row);
  PjParams p;
  p.n_chips = nc;
  p.n_chip_ids = ci;
  p.n_instance_ids = ii;
  p.chip_chip = c->chip_chip;
  p.chip_instance = c->chip_instance;
  p.energy_value = c->energy_value;
  p.power_max_value = c->power_max_value;
  p.uname_node = c->uname_node;
  p.pw_row = pw_row;
  p.mx_row = mx_row;
  p.inst_row = inst_row;
  p.chip_flags = o->chip_flags;
  p.chip_watts = o->chip_watts;
  p.chip_max = o->chip_max_watts;
  p.chip_node = o->chip_node;
  p.partial = partial;
  pj_assemble_kernel<<<PJ_BLOCKS, 256, 0, st>>>(p);
  pj_final_kernel<<<1, 32, 0, st>>>(partial, PJ_BLOCKS, nc, o->fleet);
  pj_static_pct_kernel<<<1, 32, 0, st>>>(o->fleet);
  e = cudaGetLastError();
  if (e != cudaSuccess) return igpu_fail(nullptr, IGPU_ERR_CUDA, "static power join launch failed: %s", cudaGetErrorString(e));
  return IGPU_OK;
}
