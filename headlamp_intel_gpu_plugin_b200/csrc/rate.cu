// rate.cu — N3: PromQL rate()/increase()/delta() over cumulative energy counters, evaluated on
// the device so the plugin can fetch the raw `node_hwmon_energy_joule_total` matrix once instead of asking the
// server for `rate(...[5m])` (metrics.ts:106-109) per refresh.  Arithmetic: Prometheus extrapolatedRate
// (promql/functions.go, v3.x; the test suite holds a CPU restatement of it); every float64 operation
// below is an explicitly rounded intrinsic so nvcc cannot contract a*b+c and drift from the Go/C++ evaluation.
//
// Layout: float64 [n_series][n_time] counters on a regular grid, NaN = no sample.  8 B read + 4 B written per
// evaluation point: an HBM-streaming kernel.  A warp walks one series 32 grid samples at a time; lane l owns grid
// sample 32k+l AND the evaluation that ends there.  Which samples exist is a 64-bit ballot mask (previous block,
// current block), so "first / last / count of samples in range" are bit operations, the values travel by shuffles,
// and the counter-reset terms are added in time order (the same order as the Go loop), so the float64 result is
// bit-identical to the serial evaluation.
#include <cstdlib>
#include <cstring>

#include "series.cuh"

struct igpu_counters {
  uint32_t n_series = 0, n_time = 0;
  double* d_v = nullptr;  // [n_series][n_time]
};

namespace {

constexpr int RATE_THREADS = 256;
constexpr int RATE_WARPS = RATE_THREADS / 32;
constexpr int RATE_SEG_BLOCKS = 64;  // a warp's work item: 64 blocks of 32 grid samples of one series
constexpr int RATE_UNROLL = 4;       // blocks in flight per warp
constexpr uint32_t RATE_MAX_KEFF = 33;

__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// Synthetic energy counters: integer microjoules and one rounding, so a CPU twin can reproduce them bit for bit.
__global__ void __launch_bounds__(256) counters_gen_kernel(double* __restrict__ out, uint32_t n_series, uint32_t n_time, uint64_t seed,
                                                           uint64_t global_first, uint32_t step_ms, uint32_t reset_period,
                                                           uint32_t missing_per_million) {
  const uint64_t total = (uint64_t)n_series * n_time;
  const uint64_t S_CPARAM = 0x21ull << 56, S_CNOISE = 0x22ull << 56, S_CMISS = 0x23ull << 56;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t s = (uint32_t)(i / n_time), t = (uint32_t)(i - (uint64_t)s * n_time);
    const uint64_t gs = global_first + s;
    const uint64_t rp = splitmix64(seed ^ S_CPARAM ^ gs);
    const int64_t mw = 50000 + (int64_t)(rp % 250000u);
    const int64_t inc = mw * (int64_t)step_ms;
    const uint32_t phase = reset_period ? (uint32_t)((rp >> 32) % reset_period) : 0;
    const int64_t boot = (int64_t)((rp >> 20) & 0xFFFFF);
    const uint64_t idx = gs * (uint64_t)n_time + t;
    const int64_t noise = (int64_t)(splitmix64(seed ^ S_CNOISE ^ idx) % (uint64_t)(inc / 4 + 1));
    const int64_t since = reset_period ? (int64_t)((t + phase) % reset_period) : (int64_t)t + boot;
    const int64_t uj = inc * since + noise;
    const bool miss = missing_per_million && (uint32_t)(splitmix64(seed ^ S_CMISS ^ idx) % 1000000u) < missing_per_million;
    out[i] = miss ? __longlong_as_double(0x7ff8000000000000ll) : __dmul_rn(__ll2double_rn(uj), 1e-6);
  }
}

struct RateParams {
  const double* __restrict__ v;
  float* __restrict__ out;
  double* __restrict__ out64;  // optional
  uint32_t n_series, n_time, n_out;
  uint32_t K, Keff;  // range in grid steps; samples a full range holds (K, or K+1 closed-left)
  uint32_t first_eval, stride, step_ms;
  uint32_t kb_first, segs_per_row;  // first 32-sample block that holds an evaluation; work items per series
  uint64_t n_items;
  double range_seconds;
  uint32_t is_counter, is_rate;
};

__device__ __forceinline__ double seconds_of(uint32_t steps, uint32_t step_ms) {
  return __ddiv_rn(__ll2double_rn((long long)steps * (long long)step_ms), 1000.0);
}

// Everything after "result = last - first (+ resets)"; `dts` / `dte` are the raw distances to the range boundaries.
__device__ __forceinline__ double rate_finish(double res, double first_v, double si, double dts, double dte, uint32_t n,
                                              bool is_counter, bool is_rate, double range_seconds) {
  const double avg = __ddiv_rn(si, __uint2double_rn(n - 1));
  const double thr = __dmul_rn(avg, 1.1);
  const double half = __ddiv_rn(avg, 2.0);
  if (dts >= thr) dts = half;
  if (is_counter && res > 0.0 && first_v >= 0.0) {
    const double dtz = __dmul_rn(si, __ddiv_rn(first_v, res));  // where the counter would have been zero
    if (dtz < dts) dts = dtz;
  }
  double eti = __dadd_rn(si, dts);
  if (dte >= thr) dte = half;
  eti = __dadd_rn(eti, dte);
  double factor = __ddiv_rn(eti, si);
  if (is_rate) factor = __ddiv_rn(factor, range_seconds);
  return __dmul_rn(res, factor);
}

__device__ __forceinline__ double ldg_stream_f64(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}

template <bool STRIDE1>
__global__ void __launch_bounds__(RATE_THREADS) counter_rate_kernel(const RateParams p) {
  __shared__ double sec_tab[RATE_MAX_KEFF + 3];  // seconds spanned by k grid steps
  if (threadIdx.x < RATE_MAX_KEFF + 3) sec_tab[threadIdx.x] = seconds_of(threadIdx.x, p.step_ms);
  __syncthreads();

  const uint32_t lane = threadIdx.x & 31;
  const uint32_t B = 32 + lane;  // bit of this lane's sample in the (previous, current) mask
  const uint32_t Keff = p.Keff, K = p.K;
  const bool is_counter = p.is_counter, is_rate = p.is_rate;
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  const uint64_t win_mask = ((1ull << Keff) - 1ull) << (B - Keff + 1);

  // constants of a range in which every grid step has a sample
  const double si_full = sec_tab[Keff - 1];
  double dts_full = sec_tab[K - Keff + 1], dte_full = sec_tab[0], factor_full = 0.0, zc_full = 0.0;
  if (Keff >= 2) {
    const double avg = __ddiv_rn(si_full, __uint2double_rn(Keff - 1));
    const double thr = __dmul_rn(avg, 1.1);
    if (dts_full >= thr) dts_full = __ddiv_rn(avg, 2.0);
    if (dte_full >= thr) dte_full = __ddiv_rn(avg, 2.0);
    factor_full = __ddiv_rn(__dadd_rn(__dadd_rn(si_full, dts_full), dte_full), si_full);
    if (is_rate) factor_full = __ddiv_rn(factor_full, p.range_seconds);
    // the zero-point clamp can only fire when first / result < dts / si; twice that is a rounding-proof skip test
    zc_full = __dmul_rn(2.0, __ddiv_rn(dts_full, si_full));
  }

  const uint32_t n_blocks = (p.n_time + 31) / 32;
  const uint64_t warp0 = (uint64_t)blockIdx.x * RATE_WARPS + (threadIdx.x >> 5), n_warps = (uint64_t)gridDim.x * RATE_WARPS;

  for (uint64_t item = warp0; item < p.n_items; item += n_warps) {
    const uint32_t row = (uint32_t)(item / p.segs_per_row), seg = (uint32_t)(item - (uint64_t)row * p.segs_per_row);
    const double* __restrict__ x = p.v + (size_t)row * p.n_time;
    const int kb0 = (int)(p.kb_first + seg * RATE_SEG_BLOCKS);
    const int kb1 = min((int)n_blocks, kb0 + RATE_SEG_BLOCKS);
    auto load_block = [&](int k) -> double {
      const long long i = (long long)k * 32 + lane;
      return (k >= 0 && i < (long long)p.n_time) ? ldg_stream_f64(x + i) : qnan;
    };
    // blocks kb0-2 (history of the warm-up block) and kb0-1 (warm-up: gives the reset terms of the first real block's history)
    double vp = load_block(kb0 - 2);
    uint32_t pm_prev = __ballot_sync(0xffffffffu, vp == vp);
    double c_prev = 0.0;
    uint32_t rm_prev = 0;
    double cur[RATE_UNROLL], nxt[RATE_UNROLL];
#pragma unroll
    for (int u = 0; u < RATE_UNROLL; ++u) cur[u] = load_block(kb0 - 1 + u);

    for (int kg = kb0 - 1; kg < kb1; kg += RATE_UNROLL) {
#pragma unroll
      for (int u = 0; u < RATE_UNROLL; ++u) nxt[u] = (kg + RATE_UNROLL + u < kb1) ? load_block(kg + RATE_UNROLL + u) : qnan;
#pragma unroll
      for (int u = 0; u < RATE_UNROLL; ++u) {
        const int k = kg + u;
        if (k >= kb1) break;  // warp-uniform
        const double vc = cur[u];
        const bool pres = vc == vc;
        const uint32_t pm_cur = __ballot_sync(0xffffffffu, pres);
        const uint64_t M = ((uint64_t)pm_cur << 32) | pm_prev;
        const bool full = M == ~0ull;  // warp-uniform: both blocks complete
        // ---- reset term of this lane's sample: the value it dropped from ----
        double c_cur = 0.0;
        if (full) {
          const double up = __shfl_up_sync(0xffffffffu, vc, 1), w = __shfl_sync(0xffffffffu, vp, 31);
          const double pv = lane ? up : w;
          if (is_counter && vc < pv) c_cur = pv;
        } else {
          const uint64_t below = M & ((1ull << B) - 1ull);
          const int pp = 63 - __clzll((long long)below);  // -1 when nothing earlier is in reach
          const double a = __shfl_sync(0xffffffffu, vc, pp & 31), b = __shfl_sync(0xffffffffu, vp, pp & 31);
          const double pv = pp >= 32 ? a : b;
          if (is_counter && pres && below && vc < pv) c_cur = pv;
        }
        const uint32_t rm_cur = __ballot_sync(0xffffffffu, c_cur != 0.0);
        const uint64_t Rm = ((uint64_t)rm_cur << 32) | rm_prev;

        if (k >= kb0) {  // warp-uniform: the warm-up block produces no output
          const uint32_t e = (uint32_t)k * 32 + lane;
          double r = qnan;
          if (full) {
            const uint32_t src = (lane - (Keff - 1)) & 31;
            const double a = __shfl_sync(0xffffffffu, vc, src), b = __shfl_sync(0xffffffffu, vp, src);
            const double vf = lane >= Keff - 1 ? a : b;
            double res = __dsub_rn(vc, vf);
            if (Rm) {  // warp-uniform, rare: add the drops in time order, as the serial loop does
              uint64_t rr = Rm;
              const uint32_t f = B - Keff + 1;
              while (rr) {
                const int bpos = __ffsll((long long)rr) - 1;
                rr &= rr - 1;
                const double cb = bpos >= 32 ? __shfl_sync(0xffffffffu, c_cur, bpos - 32) : __shfl_sync(0xffffffffu, c_prev, bpos);
                if (((win_mask >> bpos) & 1ull) && (uint32_t)bpos != f) res = __dadd_rn(res, cb);
              }
            }
            if (Keff >= 2) {
              if (is_counter && res > 0.0 && vf >= 0.0 && vf < __dmul_rn(res, zc_full)) {
                r = rate_finish(res, vf, si_full, sec_tab[K - Keff + 1], sec_tab[0], Keff, is_counter, is_rate, p.range_seconds);
              } else {
                r = __dmul_rn(res, factor_full);
              }
            }
          } else {
            const uint64_t PW = M & win_mask;
            const uint32_t n = __popcll(PW);
            const int f = __ffsll((long long)PW) - 1, li = 63 - __clzll((long long)PW);
            const double fa = __shfl_sync(0xffffffffu, vc, f & 31), fb = __shfl_sync(0xffffffffu, vp, f & 31);
            const double la = __shfl_sync(0xffffffffu, vc, li & 31), lb = __shfl_sync(0xffffffffu, vp, li & 31);
            const double vf = f >= 32 ? fa : fb, vl = li >= 32 ? la : lb;
            double res = __dsub_rn(vl, vf);
            if (Rm) {
              uint64_t rr = Rm;
              while (rr) {
                const int bpos = __ffsll((long long)rr) - 1;
                rr &= rr - 1;
                const double cb = bpos >= 32 ? __shfl_sync(0xffffffffu, c_cur, bpos - 32) : __shfl_sync(0xffffffffu, c_prev, bpos);
                if (((PW >> bpos) & 1ull) && bpos != f) res = __dadd_rn(res, cb);
              }
            }
            if (n >= 2)
              r = rate_finish(res, vf, sec_tab[li - f], sec_tab[f + K - B], sec_tab[B - li], n, is_counter, is_rate, p.range_seconds);
          }
          if (e >= p.first_eval && e < p.n_time) {
            uint32_t j = e - p.first_eval;
            bool emit = true;
            if (!STRIDE1) {
              const uint32_t q = j / p.stride;
              emit = q * p.stride == j;
              j = q;
            }
            if (emit) {
              const size_t o = (size_t)row * p.n_out + j;
              p.out[o] = __double2float_rn(r);
              if (p.out64) p.out64[o] = r;
            }
          }
        }
        vp = vc;
        pm_prev = pm_cur;
        c_prev = c_cur;
        rm_prev = rm_cur;
      }
#pragma unroll
      for (int u = 0; u < RATE_UNROLL; ++u) cur[u] = nxt[u];
    }
  }
}

// Any range length: one thread per evaluation, two serial passes over its range (first/last/count, then the drops).
__global__ void __launch_bounds__(256) counter_rate_generic_kernel(const RateParams p) {
  const uint64_t total = (uint64_t)p.n_series * p.n_out;
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  const bool is_counter = p.is_counter, is_rate = p.is_rate;
  for (uint64_t o = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; o < total; o += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t row = (uint32_t)(o / p.n_out), j = (uint32_t)(o - (uint64_t)row * p.n_out);
    const double* __restrict__ x = p.v + (size_t)row * p.n_time;
    const long long e = (long long)p.first_eval + (long long)j * p.stride;
    long long lo = e - (long long)p.Keff + 1;
    if (lo < 0) lo = 0;
    long long f = -1, l = -1;
    uint32_t n = 0;
    for (long long i = lo; i <= e; ++i) {
      const double v = x[i];
      if (v == v) {
        if (f < 0) f = i;
        l = i;
        ++n;
      }
    }
    double r = qnan;
    if (n >= 2) {
      const double vf = x[f];
      double res = __dsub_rn(x[l], vf);
      if (is_counter) {
        double prev = vf;
        for (long long i = f + 1; i <= l; ++i) {
          const double v = x[i];
          if (v == v) {
            if (v < prev) res = __dadd_rn(res, prev);
            prev = v;
          }
        }
      }
      r = rate_finish(res, vf, seconds_of((uint32_t)(l - f), p.step_ms), seconds_of((uint32_t)(f - (e - (long long)p.K)), p.step_ms),
                      seconds_of((uint32_t)(e - l), p.step_ms), n, is_counter, is_rate, p.range_seconds);
    }
    p.out[o] = __double2float_rn(r);
    if (p.out64) p.out64[o] = r;
  }
}

}  // namespace

extern "C" {

int32_t igpu_counters_create(igpu_ctx* ctx, uint32_t n_series, uint32_t n_time, igpu_counters** out) {
  if (!ctx || !out) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_counters_create: NULL argument");
  *out = nullptr;
  IGPU_REQUIRE(ctx, n_time > 0, IGPU_ERR_ARG, "n_time must be positive");
  DeviceGuard g(ctx->device);
  igpu_counters* c = new igpu_counters();
  c->n_series = n_series;
  c->n_time = n_time;
  const size_t bytes = std::max<size_t>((size_t)n_series * n_time * sizeof(double), 16);
  if (cudaMalloc(&c->d_v, bytes) != cudaSuccess) {
    cudaGetLastError();
    delete c;
    return igpu_fail(ctx, IGPU_ERR_OOM, "device allocation of %zu counter bytes failed", bytes);
  }
  *out = c;
  return IGPU_OK;
}

int32_t igpu_counters_destroy(igpu_ctx* ctx, igpu_counters* c) {
  if (!ctx || !c) return IGPU_OK;
  DeviceGuard g(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (c->d_v) cudaFree(c->d_v);
  delete c;
  return IGPU_OK;
}

int32_t igpu_counters_upload(igpu_ctx* ctx, igpu_counters* c, uint32_t first, uint32_t count, const double* host) {
  if (!ctx || !c || (!host && count)) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_counters_upload: NULL argument");
  IGPU_REQUIRE(ctx, (uint64_t)first + count <= c->n_series, IGPU_ERR_ARG, "rows [%u,%u) outside the slab of %u series", first,
               first + count, c->n_series);
  DeviceGuard g(ctx->device);
  const size_t bytes = (size_t)count * c->n_time * sizeof(double);
  igpu_mark(ctx, 0);
  if (bytes) IGPU_CUDA(ctx, cudaMemcpyAsync(c->d_v + (size_t)first * c->n_time, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
  igpu_mark(ctx, 1);
  ctx->timing.h2d_bytes = bytes;
  return IGPU_OK;
}

int32_t igpu_counters_generate(igpu_ctx* ctx, igpu_counters* c, uint64_t seed, uint64_t global_first, uint32_t step_ms,
                               uint32_t reset_period, uint32_t missing_per_million) {
  if (!ctx || !c) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_counters_generate: NULL argument");
  IGPU_REQUIRE(ctx, step_ms > 0, IGPU_ERR_ARG, "step_ms must be positive");
  DeviceGuard g(ctx->device);
  if (c->n_series) {
    counters_gen_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(c->d_v, c->n_series, c->n_time, seed, global_first, step_ms,
                                                                   reset_period, missing_per_million);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
  }
  return IGPU_OK;
}

uint32_t igpu_counters_rate_len(uint32_t n_time, const igpu_rate_params* p) {
  if (!p || !p->eval_stride || p->first_eval >= n_time) return 0;
  return (n_time - 1 - p->first_eval) / p->eval_stride + 1;
}

int32_t igpu_counters_rate(igpu_ctx* ctx, const igpu_counters* c, const igpu_rate_params* rp, igpu_series* dst, double* host_rate64) {
  if (!ctx || !c || !rp || !dst) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_counters_rate: NULL argument");
  IGPU_REQUIRE(ctx, rp->step_ms > 0 && rp->range_samples > 0 && rp->eval_stride > 0, IGPU_ERR_ARG,
               "step_ms, range_samples and eval_stride must be positive");
  IGPU_REQUIRE(ctx, (rp->options & ~(IGPU_RATE_CLOSED_LEFT | IGPU_RATE_INCREASE | IGPU_RATE_GAUGE)) == 0, IGPU_ERR_ARG,
               "unknown option bits 0x%x", rp->options);
  IGPU_REQUIRE(ctx, (uint64_t)rp->range_samples * rp->step_ms < (1ull << 40), IGPU_ERR_RANGE, "range of %u steps of %u ms is too long",
               rp->range_samples, rp->step_ms);
  const uint32_t n_out = igpu_counters_rate_len(c->n_time, rp);
  IGPU_REQUIRE(ctx, dst->n_series == c->n_series && dst->n_time == n_out, IGPU_ERR_ARG,
               "destination slab is %u x %u, the rate of %u x %u counters is %u x %u", dst->n_series, dst->n_time, c->n_series, c->n_time,
               c->n_series, n_out);
  DeviceGuard g(ctx->device);
  RateParams p;
  memset(&p, 0, sizeof p);
  p.v = c->d_v;
  p.out = dst->d_x;
  p.n_series = c->n_series;
  p.n_time = c->n_time;
  p.n_out = n_out;
  p.K = rp->range_samples;
  p.Keff = rp->range_samples + ((rp->options & IGPU_RATE_CLOSED_LEFT) ? 1u : 0u);
  p.first_eval = rp->first_eval;
  p.stride = rp->eval_stride;
  p.step_ms = rp->step_ms;
  p.range_seconds = (double)((int64_t)rp->range_samples * (int64_t)rp->step_ms) / 1000.0;
  p.is_counter = !(rp->options & IGPU_RATE_GAUGE);
  p.is_rate = !(rp->options & IGPU_RATE_INCREASE);
  double* d_out64 = nullptr;
  const size_t n_res = (size_t)c->n_series * n_out;
  if (host_rate64 && n_res) IGPU_CUDA(ctx, cudaMalloc(&d_out64, n_res * sizeof(double)));
  p.out64 = d_out64;
  igpu_mark(ctx, 2);
  if (n_res) {
    igpu_kbegin(ctx, IGPU_K_COUNTER_RATE);
    const bool warp_path = p.Keff <= RATE_MAX_KEFF && !getenv("IGPU_RATE_FORCE_GENERIC");
    if (warp_path) {
      const uint32_t n_blocks = (c->n_time + 31) / 32;
      p.kb_first = rp->first_eval / 32;
      p.segs_per_row = (n_blocks - p.kb_first + RATE_SEG_BLOCKS - 1) / RATE_SEG_BLOCKS;
      p.n_items = (uint64_t)c->n_series * p.segs_per_row;
      const uint64_t want = (p.n_items + RATE_WARPS - 1) / RATE_WARPS;
      const unsigned grid = (unsigned)std::min<uint64_t>(want, (uint64_t)ctx->sm_count * 8);
      if (p.stride == 1)
        counter_rate_kernel<true><<<grid, RATE_THREADS, 0, ctx->stream>>>(p);
      else
        counter_rate_kernel<false><<<grid, RATE_THREADS, 0, ctx->stream>>>(p);
    } else {
      const unsigned grid = (unsigned)std::min<uint64_t>((n_res + 255) / 256, (uint64_t)ctx->sm_count * 16);
      counter_rate_generic_kernel<<<grid, 256, 0, ctx->stream>>>(p);
    }
    cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) {
      if (d_out64) cudaFree(d_out64);
      return igpu_fail(ctx, IGPU_ERR_CUDA, "rate kernel launch failed: %s", cudaGetErrorString(le));
    }
    igpu_count_launch(ctx);
    igpu_kend(ctx, IGPU_K_COUNTER_RATE);
  }
  igpu_mark(ctx, 3);
  // the slab changed: window statistics of an earlier call no longer describe it
  dst->window = 0;
  dst->n_win = 0;
  if (d_out64) {
    cudaError_t ce = cudaMemcpyAsync(host_rate64, d_out64, n_res * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_out64);
    if (ce != cudaSuccess) return igpu_fail(ctx, IGPU_ERR_CUDA, "copy of the float64 rates failed: %s", cudaGetErrorString(ce));
  }
  return IGPU_OK;
}

}  // extern "C"
