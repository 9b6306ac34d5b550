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

struct FhParams {
  const float* x;
  uint64_t n;        // samples in the slab
  uint32_t shift;    // bin = (key >> shift) & bin_mask
  uint32_t bin_mask;
  uint32_t pmask;    // a sample takes part in histogram t when (key & pmask) == prefix[t]
  uint32_t prefix[2];
  uint32_t two;      // the two wanted ranks fell into different bins of an earlier pass
  unsigned long long* hist;  // [2][FH_BINS]
};

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
  const uint64_t n4 = p.n >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(p.x);  // the slab is 256-byte aligned
  const uint64_t stride = (uint64_t)gridDim.x * FH_THREADS;
  uint64_t i = (uint64_t)blockIdx.x * FH_THREADS + threadIdx.x;
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
  __syncthreads();
  for (int b = threadIdx.x; b < 2 * FH_BINS; b += FH_THREADS) {
    const uint32_t c = (&h[0][0])[b];
    if (c) atomicAdd(p.hist + b, (unsigned long long)c);
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

extern "C" int32_t igpu_series_quantile_fleet(igpu_ctx* ctx, igpu_series* s, double q, int32_t mode, double* host_quantile,
                                              uint64_t* host_count) {
  if (!ctx || !s || !host_quantile) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_quantile_fleet: NULL argument");
  IGPU_REQUIRE(ctx, q >= 0.0 && q <= 1.0, IGPU_ERR_ARG, "quantile %g outside [0, 1]", q);
  IGPU_REQUIRE(ctx, mode == 0 || mode == 1, IGPU_ERR_ARG, "mode %d (0 = interpolate, 1 = nearest rank)", mode);
  DeviceGuard g(ctx->device);
  char* ws = nullptr;
  int32_t st = fleet_scratch(ctx, s, 2 * FH_BINS * sizeof(unsigned long long), &ws);
  if (st != IGPU_OK) return st;
  unsigned long long* d_hist = reinterpret_cast<unsigned long long*>(ws);
  std::vector<unsigned long long> h(2 * FH_BINS);
  FhParams p;
  p.x = s->d_x;
  p.n = (uint64_t)s->n_series * s->n_time;
  p.hist = d_hist;
  p.pmask = 0, p.prefix[0] = p.prefix[1] = 0, p.two = 0;
  const uint32_t shifts[3] = {21, 10, 0}, bits[3] = {11, 11, 10};
  uint64_t rank[2] = {0, 0}, n_valid = 0;
  double weight = 0.0;  // interpolation weight of the upper rank
  const unsigned grid = (unsigned)ctx->sm_count * 4u;
  igpu_mark(ctx, 2);
  for (int pass = 0; pass < 3; ++pass) {
    p.shift = shifts[pass];
    p.bin_mask = (1u << bits[pass]) - 1u;
    IGPU_CUDA(ctx, cudaMemsetAsync(d_hist, 0, 2 * FH_BINS * sizeof(unsigned long long), ctx->stream));
    fleet_hist_kernel<<<grid, FH_THREADS, 0, ctx->stream>>>(p);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
    // the exchange step: the histogram of the whole fleet (every rank then takes the same decision)
    st = igpu_allreduce_i64_sum(ctx, reinterpret_cast<int64_t*>(d_hist), 2 * FH_BINS);
    if (st != IGPU_OK) return st;
    IGPU_CUDA(ctx, cudaMemcpyAsync(h.data(), d_hist, 2 * FH_BINS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (pass == 0) {
      for (int b = 0; b < FH_BINS; ++b) n_valid += h[b];
      if (host_count) *host_count = n_valid;
      if (!n_valid) {
        *host_quantile = std::nan("");
        igpu_mark(ctx, 3);
        return IGPU_OK;
      }
      if (mode == 1) {  // nearest rank: x[ceil(q n) - 1]
        const double r = std::ceil(q * (double)n_valid) - 1.0;
        uint64_t k = r < 0 ? 0 : (uint64_t)r;
        if (k >= n_valid) k = n_valid - 1;
        rank[0] = rank[1] = k;
      } else {  // linear interpolation between the closest ranks (PromQL quantile_over_time)
        const double pos = q * (double)(n_valid - 1);
        rank[0] = (uint64_t)std::floor(pos);
        rank[1] = std::min<uint64_t>(n_valid - 1, rank[0] + 1);
        weight = pos - (double)rank[0];
      }
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
  igpu_mark(ctx, 3);
  const double xlo = (double)host_key_f32(p.prefix[0]), xhi = (double)host_key_f32(p.prefix[1]);
  *host_quantile = mode == 1 ? xlo : xlo * (1.0 - weight) + xhi * weight;
  return IGPU_OK;
}

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
  const uint32_t N = s->n_nodes, T = s->n_time;
  const uint32_t n_win = T ? (T + window - 1) / window : 0;
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
