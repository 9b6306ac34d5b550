// fleet_series.cu — fleet-level views of the power slab (north star: "per-node and per-fleet windowed
// min/mean/max/p95"): the quantile over EVERY valid sample of the fleet, and window statistics per node (over
// the node's chips) and per fleet.  The reference has no such code (only the mock-up
// docs/screenshots/03-metrics.svg:26-37,85-89 shows "Peak Power Draw" / "Avg Power (cluster)" over a time range);
// the definitions are SURVEY §8(c)'s PromQL ones (quantile_over_time, min / max / avg over the joined series), which
// the test suite's CPU restatement follows as well.  Parity unpinned by the reference.
//
// Multi-GPU: every rank holds the series of its node range.  Node planes are local.  The fleet quantile is an exact
// radix select over the order-preserving 32-bit keys of the samples — three histogram passes of 11 + 11 + 10 bits,
// each pass one read of the rank's slab into CTA-private shared-memory histograms, each histogram all-reduced over
// NCCL (ncclAllReduce(int64[2 x 2048], sum), SURVEY §8(e) "Optional") before the host picks the bin that holds the
// wanted rank.  The fleet window planes are all-reduced the same way (sum / count: sum; min: min; max: max).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "series.cuh"

namespace {

constexpr int FH_BINS = 2048, FH_THREADS = 512;
constexpr uint32_t FQ_EVERY = 64;            // the sample: one 128-byte line (32 samples) in 64
constexpr double FQ_CLUSTER = 32.0;          // a line's 32 samples are neighbours in time: counted as ONE independent draw
constexpr double FQ_SIGMAS = 5.0;            // half-width of the bracket in standard deviations of the sample rank
constexpr uint64_t FQ_MIN_SAMPLE = 1u << 16; // below that many sampled values the three exact passes are cheap anyway

struct FhParams {
  const float* x;
  uint64_t n;        // samples in the slab
  uint32_t every;    // 1: all of them; else one 128-byte line in `every` (whole lines only)
  uint32_t shift;    // bin = (key >> shift) & bin_mask
  uint32_t bin_mask;
  uint32_t pmask;    // a sample takes part in histogram t when (key & pmask) == prefix[t]
  uint32_t prefix[2];
  uint32_t two;      // the two wanted ranks fell into different bins of an earlier pass
  unsigned long long* hist;  // [2][FH_BINS]
};

__device__ __forceinline__ uint32_t fq_mix(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352dU;
  x ^= x >> 15;
  x *= 0x846ca68bU;
  return x ^ (x >> 16);
}

__device__ __forceinline__ void fh_one(const FhParams& p, uint32_t (*h)[FH_BINS], float v) {
  if (v != v) return;  // NaN = no sample
  const uint32_t key = f32_key(v);
  const uint32_t bin = (key >> p.shift) & p.bin_mask;
  // (atomicAdd, not a hand-written RED: nvcc's warp aggregation pays off here — power samples share their leading bits,
  // so in the first pass most lanes of a warp hit the same bin)
  if ((key & p.pmask) == p.prefix[0]) atomicAdd(&h[0][bin], 1u);
  if (p.two && (key & p.pmask) == p.prefix[1]) atomicAdd(&h[1][bin], 1u);
}

__global__ void __launch_bounds__(FH_THREADS) fleet_hist_kernel(const FhParams p) {
  __shared__ uint32_t h[2][FH_BINS];
  for (int i = threadIdx.x; i < 2 * FH_BINS; i += FH_THREADS) (&h[0][0])[i] = 0;
  __syncthreads();
  const float4* x4 = reinterpret_cast<const float4*>(p.x);  // the slab is 256-byte aligned
  const uint64_t stride = (uint64_t)gridDim.x * FH_THREADS;
  uint64_t i = (uint64_t)blockIdx.x * FH_THREADS + threadIdx.x;
  if (p.every > 1) {
    // the sample: line `pick * every + hash(pick) % every` of every group of `every` lines, eight float4s each
    const uint64_t n_item = (p.n / 32 / p.every) * 8;
    for (; i < n_item; i += stride) {
      const uint64_t pick = i >> 3;
      const uint64_t line = pick * p.every + fq_mix((uint32_t)pick) % p.every;
      const float4 v = ldg_stream_f4(x4 + line * 8 + (i & 7));
      fh_one(p, h, v.x), fh_one(p, h, v.y), fh_one(p, h, v.z), fh_one(p, h, v.w);
    }
  } else {
    const uint64_t n4 = p.n >> 2;
    for (; i + 3 * stride < n4; i += 4 * stride) {  // four 128-bit loads in flight per thread
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = ldg_stream_f4(x4 + i + j * stride);
#pragma unroll
      for (int j = 0; j < 4; ++j) fh_one(p, h, v[j].x), fh_one(p, h, v[j].y), fh_one(p, h, v[j].z), fh_one(p, h, v[j].w);
    }
    for (; i < n4; i += stride) {
      const float4 v = ldg_stream_f4(x4 + i);
      fh_one(p, h, v.x), fh_one(p, h, v.y), fh_one(p, h, v.z), fh_one(p, h, v.w);
    }
    if (blockIdx.x == 0 && threadIdx.x < (p.n & 3)) fh_one(p, h, p.x[(n4 << 2) + threadIdx.x]);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < 2 * FH_BINS; b += FH_THREADS) {
    const uint32_t c = (&h[0][0])[b];
    if (c) atomicAdd(p.hist + b, (unsigned long long)c);
  }
}

// The one full read of the bracketed select: how many valid samples lie below the bracket [lo_key, hi_key], how many there
// are at all, and the samples inside it, appended to `cand` (one global atomic per warp and 128 samples that hold a
// candidate at all — the bracket holds about a thousandth of the samples).
struct FbParams {
  const float* x;
  uint64_t n;
  uint32_t lo_key, hi_key;
  float* cand;
  uint32_t cap;
  unsigned long long* counters;  // [0] below, [1] inside (may exceed cap), [2] valid, [3] overflowed
};

__global__ void __launch_bounds__(FH_THREADS) fleet_bracket_kernel(const FbParams p) {
  const uint32_t lane = threadIdx.x & 31, span = p.hi_key - p.lo_key;
  const uint64_t n4 = p.n >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(p.x);
  const uint64_t stride = (uint64_t)gridDim.x * FH_THREADS;
  uint32_t below = 0, valid = 0;
  auto take = [&](float4 v, bool live) {  // called by whole warps
    const float f[4] = {v.x, v.y, v.z, v.w};
    uint32_t in = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const bool ok = live && f[c] == f[c];
      const uint32_t key = f32_key(f[c]);
      valid += ok;
      below += ok && key < p.lo_key;
      in |= (uint32_t)(ok && key - p.lo_key <= span) << c;
    }
    if (!__any_sync(0xffffffffu, in)) return;
    // append: lanes in order, a lane's components in order
    const uint32_t mine = __popc(in);
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= (uint32_t)o) incl += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(p.counters + 1, (unsigned long long)total);
    base = __shfl_sync(0xffffffffu, base, 0);
    unsigned long long at = base + incl - mine;
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (in & (1u << c)) {
        if (at < p.cap) p.cand[at] = f[c];
        ++at;
      }
  };
  // warp-uniform trip count: every lane of a warp runs the same iterations, lanes past the end carry dead values
  const uint64_t warp0 = (uint64_t)blockIdx.x * FH_THREADS + (threadIdx.x & ~31u);
  for (uint64_t w = warp0; w < n4; w += 4 * stride) {
    float4 v[4];
    bool live[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint64_t i = w + j * stride + lane;
      live[j] = i < n4;
      v[j] = live[j] ? ldg_stream_f4(x4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (w + j * stride < n4) take(v[j], live[j]);  // uniform
  }
  if (blockIdx.x == 0 && threadIdx.x < 32) {  // the up to three samples after the last float4
    const bool live = lane < (p.n & 3);
    const float v = live ? p.x[(n4 << 2) + lane] : 0.f;
    take(make_float4(v, __int_as_float(0x7fc00000), __int_as_float(0x7fc00000), __int_as_float(0x7fc00000)), live);
  }
  below = __reduce_add_sync(0xffffffffu, below);
  valid = __reduce_add_sync(0xffffffffu, valid);
  if (lane == 0) {
    if (below) atomicAdd(p.counters, (unsigned long long)below);
    if (valid) atomicAdd(p.counters + 2, (unsigned long long)valid);
  }
}

}  // namespace

// grow-only device scratch of the fleet-level calls, owned by the series handle
static int32_t fleet_scratch(igpu_ctx* ctx, igpu_series* s, size_t bytes, char** out) {
  if (s->fs_bytes < bytes) {
    if (s->d_fs) cudaFree(s->d_fs);
    s->d_fs = nullptr, s->fs_bytes = 0;
    IGPU_CUDA(ctx, cudaMalloc(&s->d_fs, bytes));
    s->fs_bytes = bytes;
  }
  *out = s->d_fs;
  return IGPU_OK;
}

static inline float host_key_f32(uint32_t k) {
  const uint32_t u = (k & 0x80000000u) ? (k ^ 0x80000000u) : ~k;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// Exact radix select over the valid samples of x[0..n) on all ranks: three histogram passes of 11 + 11 + 10 key bits,
// each all-reduced (the exchange step: every rank then takes the same decision).  `every` > 1 selects among a sparse
// sample of the slab instead.  The two wanted ranks are decided after the first pass, when the number of valid samples
// is known: want(n_valid, rank) (false: nothing to select, e.g. no valid sample).  keys[t] = the key at rank[t].
template <class Want>
static int32_t fleet_radix_select(igpu_ctx* ctx, unsigned long long* d_hist, const float* x, uint64_t n, uint32_t every, Want want,
                                  uint64_t* n_valid_out, uint32_t keys[2], bool* selected) {
  std::vector<unsigned long long> h(2 * FH_BINS);
  FhParams p;
  p.x = x;
  p.n = n;
  p.every = every;
  p.hist = d_hist;
  p.pmask = 0, p.prefix[0] = p.prefix[1] = 0, p.two = 0;
  const uint32_t shifts[3] = {21, 10, 0}, bits[3] = {11, 11, 10};
  uint64_t rank[2] = {0, 0}, n_valid = 0;
  const unsigned grid = (unsigned)ctx->sm_count * 4u;
  *selected = false;
  for (int pass = 0; pass < 3; ++pass) {
    p.shift = shifts[pass];
    p.bin_mask = (1u << bits[pass]) - 1u;
    IGPU_CUDA(ctx, cudaMemsetAsync(d_hist, 0, 2 * FH_BINS * sizeof(unsigned long long), ctx->stream));
    fleet_hist_kernel<<<grid, FH_THREADS, 0, ctx->stream>>>(p);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
    const int32_t st = igpu_allreduce_i64_sum(ctx, reinterpret_cast<int64_t*>(d_hist), 2 * FH_BINS);
    if (st != IGPU_OK) return st;
    IGPU_CUDA(ctx, cudaMemcpyAsync(h.data(), d_hist, 2 * FH_BINS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (pass == 0) {
      for (int b = 0; b < FH_BINS; ++b) n_valid += h[b];
      *n_valid_out = n_valid;
      if (!want(n_valid, rank)) return IGPU_OK;
    }
    // the bin that holds each wanted rank; the rank becomes relative to that bin
    for (int t = 0; t < 2; ++t) {
      const unsigned long long* ht = h.data() + ((p.two && t == 1) ? FH_BINS : 0);
      uint64_t below = 0;
      uint32_t b = 0;
      for (; b < (uint32_t)FH_BINS; ++b) {
        if (rank[t] < below + ht[b]) break;
        below += ht[b];
      }
      IGPU_REQUIRE(ctx, b < (uint32_t)FH_BINS, IGPU_ERR_STATE, "fleet quantile: rank outside the histogram (the slab changed during the call?)");
      rank[t] -= below;
      p.prefix[t] |= b << p.shift;
    }
    p.pmask |= p.bin_mask << p.shift;
    p.two = p.prefix[0] != p.prefix[1];
  }
  keys[0] = p.prefix[0], keys[1] = p.prefix[1];
  *selected = true;
  return IGPU_OK;
}

// the two ranks an interpolated (mode 0, PromQL quantile_over_time) or nearest-rank (mode 1) quantile of n values needs
static void fleet_wanted_ranks(double q, int32_t mode, uint64_t n, uint64_t rank[2], double* weight) {
  if (mode == 1) {  // x[ceil(q n) - 1]
    const double r = std::ceil(q * (double)n) - 1.0;
    uint64_t k = r < 0 ? 0 : (uint64_t)r;
    if (k >= n) k = n - 1;
    rank[0] = rank[1] = k;
    *weight = 0.0;
  } else {
    const double pos = q * (double)(n - 1);
    rank[0] = (uint64_t)std::floor(pos);
    rank[1] = std::min<uint64_t>(n - 1, rank[0] + 1);
    *weight = pos - (double)rank[0];
  }
}

extern "C" int32_t igpu_series_quantile_fleet(igpu_ctx* ctx, igpu_series* s, double q, int32_t mode, double* host_quantile,
                                              uint64_t* host_count) {
  if (!ctx || !s || !host_quantile) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_quantile_fleet: NULL argument");
  IGPU_REQUIRE(ctx, q >= 0.0 && q <= 1.0, IGPU_ERR_ARG, "quantile %g outside [0, 1]", q);
  IGPU_REQUIRE(ctx, mode == 0 || mode == 1, IGPU_ERR_ARG, "mode %d (0 = interpolate, 1 = nearest rank)", mode);
  DeviceGuard g(ctx->device);
  const uint64_t n = (uint64_t)s->n_series * s->n_time;
  // scratch: histograms | counters | candidates of the bracketed pass (twice the expected number and a floor)
  const uint32_t cap = (uint32_t)std::min<uint64_t>(n / 256 + (1u << 20), 0x7fffffffu);
  const size_t off_cnt = 2 * FH_BINS * sizeof(unsigned long long), off_cand = off_cnt + 256;
  char* ws = nullptr;
  int32_t st = fleet_scratch(ctx, s, off_cand + (size_t)cap * sizeof(float), &ws);
  if (st != IGPU_OK) return st;
  unsigned long long* d_hist = reinterpret_cast<unsigned long long*>(ws);
  unsigned long long* d_cnt = reinterpret_cast<unsigned long long*>(ws + off_cnt);
  float* d_cand = reinterpret_cast<float*>(ws + off_cand);
  double weight = 0.0;
  uint64_t n_valid = 0;
  uint32_t keys[2] = {0, 0};
  bool selected = false;
  igpu_mark(ctx, 2);
  s->fleet_quantile_reads = 0;

  // ---- 1. a bracket [lo, hi] of keys around the wanted ranks, from a sparse sample of every rank's slab ----
  uint64_t m = 0;
  uint32_t bracket[2] = {0u, 0xFFFFFFFFu};
  bool have_bracket = false;
  if (!ctx->sw.fleet_quantile_exact) {
    st = fleet_radix_select(
        ctx, d_hist, s->d_x, n, FQ_EVERY,
        [&](uint64_t mv, uint64_t rank[2]) {
          if (mv < (ctx->sw.fleet_min_sample > 0 ? (uint64_t)ctx->sw.fleet_min_sample : FQ_MIN_SAMPLE)) return false;
          const double pos = q * (double)(mv - 1);
          const double sigma = std::sqrt(std::max(q * (1.0 - q), 0.0) * (double)mv * FQ_CLUSTER);
          const double d = std::ceil(FQ_SIGMAS * sigma) + 2.0;
          rank[0] = pos - d < 0.0 ? 0 : (uint64_t)(pos - d);
          rank[1] = pos + d + 1.0 >= (double)mv ? mv - 1 : (uint64_t)(pos + d + 1.0);
          return true;
        },
        &m, bracket, &have_bracket);
    if (st != IGPU_OK) return st;
  }
  if (have_bracket) {
    // the sample's extremes stand for "no bound": the true extremes lie outside them
    const double pos = q * (double)(m - 1), sigma = std::sqrt(std::max(q * (1.0 - q), 0.0) * (double)m * FQ_CLUSTER);
    const double d = std::ceil(FQ_SIGMAS * sigma) + 2.0;
    if (pos - d < 0.0) bracket[0] = 0u;
    if (pos + d + 1.0 >= (double)m) bracket[1] = 0xFFFFFFFFu;
    // ---- 2. ONE read of the slab: valid samples, samples below the bracket, the samples inside it ----
    IGPU_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, 4 * sizeof(unsigned long long), ctx->stream));
    FbParams bp{s->d_x, n, bracket[0], bracket[1], d_cand, cap, d_cnt};
    fleet_bracket_kernel<<<(unsigned)ctx->sm_count * 4u, FH_THREADS, 0, ctx->stream>>>(bp);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
    s->fleet_quantile_reads = 1;
    unsigned long long local[4];
    IGPU_CUDA(ctx, cudaMemcpyAsync(local, d_cnt, sizeof local, cudaMemcpyDeviceToHost, ctx->stream));
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint64_t n_mid_local = local[1];
    local[3] = n_mid_local > cap ? 1 : 0;  // a rank whose candidates did not fit sends every rank to the exact passes
    IGPU_CUDA(ctx, cudaMemcpyAsync(d_cnt, local, sizeof local, cudaMemcpyHostToDevice, ctx->stream));
    st = igpu_allreduce_i64_sum(ctx, reinterpret_cast<int64_t*>(d_cnt), 4);
    if (st != IGPU_OK) return st;
    unsigned long long tot[4];
    IGPU_CUDA(ctx, cudaMemcpyAsync(tot, d_cnt, sizeof tot, cudaMemcpyDeviceToHost, ctx->stream));
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    n_valid = tot[2];
    if (host_count) *host_count = n_valid;
    if (!n_valid) {
      *host_quantile = std::nan("");
      igpu_mark(ctx, 3);
      return IGPU_OK;
    }
    uint64_t rank[2];
    fleet_wanted_ranks(q, mode, n_valid, rank, &weight);
    if (!tot[3] && rank[0] >= tot[0] && rank[1] < tot[0] + tot[1]) {
      // ---- 3. the exact ranks among the candidates (a thousandth of the slab) ----
      const uint64_t r0 = rank[0] - tot[0], r1 = rank[1] - tot[0];
      uint64_t n_mid = 0;
      st = fleet_radix_select(
          ctx, d_hist, d_cand, n_mid_local, 1,
          [&](uint64_t, uint64_t rk[2]) {
            rk[0] = r0, rk[1] = r1;
            return true;
          },
          &n_mid, keys, &selected);
      if (st != IGPU_OK) return st;
      IGPU_REQUIRE(ctx, selected && n_mid == tot[1], IGPU_ERR_STATE, "fleet quantile: candidate count changed between passes");
    }
  }
  if (!selected) {
    // ---- the bracket missed, overflowed or there was no sample to speak of: three exact passes over the slab ----
    st = fleet_radix_select(
        ctx, d_hist, s->d_x, n, 1,
        [&](uint64_t nv, uint64_t rank[2]) {
          if (!nv) return false;
          fleet_wanted_ranks(q, mode, nv, rank, &weight);
          return true;
        },
        &n_valid, keys, &selected);
    if (st != IGPU_OK) return st;
    s->fleet_quantile_reads += 3;
    if (host_count) *host_count = n_valid;
    if (!selected) {
      *host_quantile = std::nan("");
      igpu_mark(ctx, 3);
      return IGPU_OK;
    }
  }
  igpu_mark(ctx, 3);
  const double xlo = (double)host_key_f32(keys[0]), xhi = (double)host_key_f32(keys[1]);
  *host_quantile = mode == 1 ? xlo : xlo * (1.0 - weight) + xhi * weight;
  return IGPU_OK;
}

extern "C" uint32_t igpu_series_fleet_quantile_reads(const igpu_series* s) { return s ? s->fleet_quantile_reads : 0; }

// ---- window statistics per node and per fleet ------------------------------------------------------------------
namespace {

// One warp per (node, window): the lanes walk the node's rows of that window (128-bit loads when the rows allow it) and
// fold min / max / sum (float64) / count; the node planes are written by lane 0, together with the float64 sum the
// fleet reduction needs.
template <bool VEC>
__global__ void __launch_bounds__(256) node_window_kernel(const float* __restrict__ x, const uint32_t* __restrict__ node_off,
                                                          uint32_t n_nodes, uint32_t n_time, uint32_t window, uint32_t n_win,
                                                          float* __restrict__ nmin, float* __restrict__ nmean, float* __restrict__ nmax,
                                                          uint32_t* __restrict__ ncount, double* __restrict__ nsum) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint64_t total = (uint64_t)n_nodes * n_win;
  for (uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; gw < total; gw += warps) {
    const uint32_t node = (uint32_t)(gw / n_win), w = (uint32_t)(gw - (uint64_t)node * n_win);
    const uint32_t t0 = w * window, t1 = min(n_time, t0 + window);
    const uint32_t s0 = __ldg(node_off + node), s1 = __ldg(node_off + node + 1);
    Acc a{INFINITY, -INFINITY, 0.0, 0u};
    if (VEC) {
      const uint32_t nvec = (t1 - t0) >> 2;  // W % 4 == 0 and T % 4 == 0
      const uint32_t items = (s1 - s0) * nvec;
      const float inv = 1.0f / (float)nvec;
      for (uint32_t i = lane; i < items; i += 32) {
        uint32_t r = (uint32_t)(((float)i + 0.5f) * inv);  // i / nvec for the small numbers at hand, fixed up below
        r -= (r * nvec > i);
        r += ((r + 1) * nvec <= i);
        const float4 v = ldg_stream_f4(reinterpret_cast<const float4*>(x + (size_t)(s0 + r) * n_time + t0) + (i - r * nvec));
        acc_sample(a, v.x), acc_sample(a, v.y), acc_sample(a, v.z), acc_sample(a, v.w);
      }
    } else {
      for (uint32_t srow = s0; srow < s1; ++srow)
        for (uint32_t t = t0 + lane; t < t1; t += 32) acc_sample(a, __ldg(x + (size_t)srow * n_time + t));
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      a.mn = fminf(a.mn, __shfl_xor_sync(0xffffffffu, a.mn, o));
      a.mx = fmaxf(a.mx, __shfl_xor_sync(0xffffffffu, a.mx, o));
      a.sum += __shfl_xor_sync(0xffffffffu, a.sum, o);
      a.cnt += __shfl_xor_sync(0xffffffffu, a.cnt, o);
    }
    if (lane == 0) {
      const float nanv = __int_as_float(0x7fc00000);
      nmin[gw] = a.cnt ? a.mn : nanv;
      nmax[gw] = a.cnt ? a.mx : nanv;
      nmean[gw] = a.cnt ? (float)(a.sum / (double)a.cnt) : nanv;
      ncount[gw] = a.cnt;
      nsum[gw] = a.sum;
    }
  }
}

// The same node planes WITHOUT the slab, from the per-series window planes the single-read rollup left behind (min, mean,
// max as float32 and — igpu_series_keep_window_counts — the number of valid samples): min of mins, max of maxes, count =
// sum of counts, sum = sum of mean x count in float64.  The mean differs from the mean over the samples only by the
// float32 rounding of the per-series means (relative 6e-8, far inside the 1e-5 of the north star); min, max and count are
// exact.  Thread per (node, window), consecutive threads consecutive windows: every plane is read coalesced.
__global__ void __launch_bounds__(256) node_planes_kernel(const float* __restrict__ wmin, const float* __restrict__ wmean,
                                                          const float* __restrict__ wmax, const uint16_t* __restrict__ wcnt,
                                                          const uint32_t* __restrict__ node_off, uint32_t n_nodes, uint32_t n_win,
                                                          float* __restrict__ nmin, float* __restrict__ nmean, float* __restrict__ nmax,
                                                          uint32_t* __restrict__ ncount, double* __restrict__ nsum) {
  const uint64_t total = (uint64_t)n_nodes * n_win;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t node = (uint32_t)(i / n_win), w = (uint32_t)(i - (uint64_t)node * n_win);
    const uint32_t s0 = __ldg(node_off + node), s1 = __ldg(node_off + node + 1);
    float mn = INFINITY, mx = -INFINITY;
    double sum = 0.0;
    uint32_t cnt = 0;
    for (uint32_t r = s0; r < s1; ++r) {
      const size_t o = (size_t)r * n_win + w;
      const uint32_t c = wcnt[o];
      if (c) {
        mn = fminf(mn, wmin[o]);
        mx = fmaxf(mx, wmax[o]);
        sum += (double)wmean[o] * (double)c;
        cnt += c;
      }
    }
    const float nanv = __int_as_float(0x7fc00000);
    nmin[i] = cnt ? mn : nanv;
    nmax[i] = cnt ? mx : nanv;
    nmean[i] = cnt ? (float)(sum / (double)cnt) : nanv;
    ncount[i] = cnt;
    nsum[i] = sum;
  }
}

// Fleet planes: thread (g, j) of a CTA adds up nodes g, g + 8, ... for window 32 * blockIdx + j (coalesced across j), the
// eight partial results of a window are combined in a fixed order: the float64 sums are deterministic.
// out: [4][n_win] doubles — sum, count, min, max (min / max +-inf where no node has a sample)
__global__ void __launch_bounds__(256) fleet_window_kernel(const float* __restrict__ nmin, const float* __restrict__ nmax,
                                                           const uint32_t* __restrict__ ncount, const double* __restrict__ nsum,
                                                           uint32_t n_nodes, uint32_t n_win, double* __restrict__ out) {
  __shared__ double s_sum[8][32], s_cnt[8][32], s_min[8][32], s_max[8][32];
  const uint32_t j = threadIdx.x & 31, gsel = threadIdx.x >> 5, w = blockIdx.x * 32 + j;
  double sum = 0, cnt = 0, mn = INFINITY, mx = -INFINITY;
  if (w < n_win)
    for (uint32_t n = gsel; n < n_nodes; n += 8) {
      const size_t o = (size_t)n * n_win + w;
      const uint32_t c = ncount[o];
      sum += nsum[o];
      cnt += (double)c;
      if (c) mn = fmin(mn, (double)nmin[o]), mx = fmax(mx, (double)nmax[o]);
    }
  s_sum[gsel][j] = sum, s_cnt[gsel][j] = cnt, s_min[gsel][j] = mn, s_max[gsel][j] = mx;
  __syncthreads();
  if (gsel == 0 && w < n_win) {
    for (int k = 1; k < 8; ++k) sum += s_sum[k][j], cnt += s_cnt[k][j], mn = fmin(mn, s_min[k][j]), mx = fmax(mx, s_max[k][j]);
    out[w] = sum, out[n_win + w] = cnt, out[2 * (size_t)n_win + w] = mn, out[3 * (size_t)n_win + w] = mx;
  }
}

}  // namespace

extern "C" int32_t igpu_series_window_rollup(igpu_ctx* ctx, igpu_series* s, uint32_t window, const igpu_window_rollup_out* o) {
  if (!ctx || !s || !o) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_window_rollup: NULL argument");
  IGPU_REQUIRE(ctx, window >= 1, IGPU_ERR_ARG, "window must be >= 1");
  DeviceGuard g(ctx->device);
  const uint32_t N = s->n_nodes, T = s->n_time;  // T: the row pitch (padding is NaN)
  const uint32_t n_win = s->n_time_user ? (s->n_time_user + window - 1) / window : 0;
  const size_t NW = (size_t)N * n_win;
  if (!n_win) return IGPU_OK;
  // scratch: nsum f64 [NW] | fleet f64 [4][n_win] | nmin, nmean, nmax f32 [NW] | ncount u32 [NW]
  const size_t off_fleet = align_up(NW * 8, 256), off_min = off_fleet + align_up((size_t)4 * n_win * 8, 256);
  const size_t plane = align_up(NW * 4, 256);
  char* ws = nullptr;
  int32_t st = fleet_scratch(ctx, s, off_min + 4 * plane, &ws);
  if (st != IGPU_OK) return st;
  double* d_nsum = reinterpret_cast<double*>(ws);
  double* d_fleet = reinterpret_cast<double*>(ws + off_fleet);
  float* d_nmin = reinterpret_cast<float*>(ws + off_min);
  float* d_nmean = reinterpret_cast<float*>(ws + off_min + plane);
  float* d_nmax = reinterpret_cast<float*>(ws + off_min + 2 * plane);
  uint32_t* d_ncount = reinterpret_cast<uint32_t*>(ws + off_min + 3 * plane);
  igpu_mark(ctx, 2);
  // No read of the slab when the single-read rollup of this window kept its per-series counts and the slab is unchanged
  const bool from_planes = s->wcnt_valid && s->window == window && s->n_win == n_win && !ctx->sw.window_rollup_from_slab;
  if (NW) {
    if (from_planes) {
      const unsigned grid = (unsigned)std::min<uint64_t>((NW + 255) / 256, (uint64_t)ctx->sm_count * 16);
      node_planes_kernel<<<grid, 256, 0, ctx->stream>>>(s->d_wmin, s->d_wmean, s->d_wmax, s->d_wcnt, s->d_node_off, N, n_win, d_nmin, d_nmean,
                                                        d_nmax, d_ncount, d_nsum);
    } else {
      const bool vec = window % 4 == 0 && T % 4 == 0;
      const unsigned grid = (unsigned)std::min<uint64_t>((NW + 7) / 8, (uint64_t)ctx->sm_count * 16);
      if (vec)
        node_window_kernel<true><<<grid, 256, 0, ctx->stream>>>(s->d_x, s->d_node_off, N, T, window, n_win, d_nmin, d_nmean, d_nmax, d_ncount, d_nsum);
      else
        node_window_kernel<false><<<grid, 256, 0, ctx->stream>>>(s->d_x, s->d_node_off, N, T, window, n_win, d_nmin, d_nmean, d_nmax, d_ncount, d_nsum);
    }
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
  }
  if (o->slab_reads) *o->slab_reads = from_planes ? 0u : 1u;
  fleet_window_kernel<<<(n_win + 31) / 32, 256, 0, ctx->stream>>>(d_nmin, d_nmax, d_ncount, d_nsum, N, n_win, d_fleet);
  IGPU_CUDA(ctx, cudaGetLastError());
  igpu_count_launch(ctx);
  // the exchange step: the fleet planes over the node-range shards
  st = igpu_comm_group_begin(ctx);
  if (st == IGPU_OK) st = igpu_allreduce_f64(ctx, d_fleet, (size_t)2 * n_win, /*sum*/ 0);
  if (st == IGPU_OK) st = igpu_allreduce_f64(ctx, d_fleet + (size_t)2 * n_win, n_win, /*min*/ 3);
  if (st == IGPU_OK) st = igpu_allreduce_f64(ctx, d_fleet + (size_t)3 * n_win, n_win, /*max*/ 2);
  const int32_t st_end = igpu_comm_group_end(ctx);
  if (st == IGPU_OK) st = st_end;
  if (st != IGPU_OK) return st;
  igpu_mark(ctx, 3);
  std::vector<double> hf((size_t)4 * n_win);
  IGPU_CUDA(ctx, cudaMemcpyAsync(hf.data(), d_fleet, hf.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (NW) {
    if (o->node_min) IGPU_CUDA(ctx, cudaMemcpyAsync(o->node_min, d_nmin, NW * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (o->node_mean) IGPU_CUDA(ctx, cudaMemcpyAsync(o->node_mean, d_nmean, NW * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (o->node_max) IGPU_CUDA(ctx, cudaMemcpyAsync(o->node_max, d_nmax, NW * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (o->node_count) IGPU_CUDA(ctx, cudaMemcpyAsync(o->node_count, d_ncount, NW * 4, cudaMemcpyDeviceToHost, ctx->stream));
  }
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const float nanv = std::nanf("");
  for (uint32_t w = 0; w < n_win; ++w) {
    const double sum = hf[w], cnt = hf[n_win + w];
    if (o->fleet_min) o->fleet_min[w] = cnt > 0 ? (float)hf[2 * (size_t)n_win + w] : nanv;
    if (o->fleet_max) o->fleet_max[w] = cnt > 0 ? (float)hf[3 * (size_t)n_win + w] : nanv;
    if (o->fleet_mean) o->fleet_mean[w] = cnt > 0 ? (float)(sum / cnt) : nanv;
    if (o->fleet_count) o->fleet_count[w] = (uint64_t)cnt;
  }
  return IGPU_OK;
}
