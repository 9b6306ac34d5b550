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
  double* d_v = nullptr;    // [n_series][n_time]
  double* d_ftab = nullptr; // extrapolation factors of the last rate call, see rate_table_kernel
  unsigned long long* d_gapstat = nullptr;  // [2] missing / looked-at samples of a sparse sample of the slab, see counters_gapstat_kernel
  uint32_t* d_dirty = nullptr;              // long-range rate: bitmap of the blocks left to the second kernel (allocated on first use)
};

namespace {

constexpr int RATE_THREADS = 256;
constexpr int RATE_WARPS = RATE_THREADS / 32;
constexpr int RATE_UNROLL = 4;       // blocks per chunk (one bulk copy, 1 KB)
static_assert(RATE_UNROLL == 4, "the gap kernel's chunk test names four blocks");
constexpr int RATE_STAGES = 4;       // chunks in a warp's shared-memory ring
constexpr int RATE_CHUNK_ELEMS = RATE_UNROLL * 32 + 2;  // +2: the copy may start one element early to be 16-byte aligned
constexpr uint32_t RATE_CHUNK_BYTES = RATE_CHUNK_ELEMS * 8;
constexpr uint32_t RATE_MAX_KEFF = 33;
constexpr uint32_t RATE_TAB_DIM = RATE_MAX_KEFF + 1;  // indices 0..33 for steps and sample counts
constexpr uint32_t RATE_FTAB_N = RATE_TAB_DIM * RATE_TAB_DIM * RATE_TAB_DIM;  // extrapolation factors
constexpr uint32_t RATE_ZTAB_N = RATE_TAB_DIM * RATE_TAB_DIM;                 // clamp thresholds, stored behind the factors
constexpr unsigned long long RATE_GAP_PPM = 3000;  // missing samples per million from which the gap kernel is the faster one

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
  uint32_t out_pitch;  // row pitch of `out` (the destination slab pads its rows to a multiple of 4 samples); out64 is dense
  uint32_t K, Keff;  // range in grid steps; samples a full range holds (K, or K+1 closed-left)
  uint32_t first_eval, stride, step_ms;
  uint32_t kb_first, segs_per_row, seg_blocks;  // first 32-sample block that holds an evaluation; work items per series and their length
  uint64_t n_items;
  double range_seconds;
  uint32_t is_counter, is_rate;
  const double* ftab;
  uint32_t* dirty;                    // long-range path: one bit per 32-sample block the streaming kernel left to the second kernel
  const unsigned long long* gapstat;  // the slab's gap statistic
  uint32_t tier;                      // 0: the kernel that suits the statistic runs, the other returns at once; 1 / 2: forced
};

__device__ __forceinline__ double seconds_of(uint32_t steps, uint32_t step_ms) {
  return __ddiv_rn(__ll2double_rn((long long)steps * (long long)step_ms), 1000.0);
}

// Everything after "result = last - first (+ resets)"; `dts` / `dte` are the raw distances to the range boundaries.
__device__ __forceinline__ double rate_finish(double res, double first_v, double si, double dts, double dte, uint32_t n,
                                              bool is_counter, bool is_rate, double range_seconds) {
  const double avg = __ddiv_rn(si, __uint2double_rn(n - 1));
  const double thr = __dmul_rn(avg, 1.1);
  const double half = __dmul_rn(avg, 0.5);  // avg / 2 exactly (a power of two): the same bits as the division, a tenth of the instructions
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

// position of the highest set bit, -1 for 0 (one instruction; volatile so that it is never hidden behind a branch)
__device__ __forceinline__ int bfind_u32(uint32_t x) {
  int r;
  asm volatile("bfind.u32 %0, %1;" : "=r"(r) : "r"(x));
  return r;
}

__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}

__device__ __forceinline__ double ldg_stream_f64(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}

// How gappy is the slab?  Looks at one 128-byte line (16 samples) out of every `every` of the n samples at v and adds
// (missing, looked at) to stat[0..1].  Runs behind every upload / generate; the rate call picks its kernel by it.
__global__ void __launch_bounds__(256) counters_gapstat_kernel(const double* __restrict__ v, uint64_t n, uint32_t every,
                                                               unsigned long long* __restrict__ stat) {
  const uint64_t n_pick = n / 16 / every;
  uint32_t miss = 0, seen = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_pick * 16; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t pick = i >> 4;
    const uint64_t line = pick * every + splitmix64(pick) % every;
    const double x = ldg_stream_f64(v + line * 16 + (i & 15));
    miss += x != x;
    ++seen;
  }
  miss = __reduce_add_sync(0xffffffffu, miss);
  seen = __reduce_add_sync(0xffffffffu, seen);
  if ((threadIdx.x & 31) == 0 && seen) {
    atomicAdd(stat, (unsigned long long)miss);
    atomicAdd(stat + 1, (unsigned long long)seen);
  }
}

// Extrapolation factor of a range that is NOT clamped at the counter's zero point: with the grid fixed it depends only
// on (steps between first and last sample, steps from the range start to the first sample, samples in range) — the steps
// to the range end follow from the range length.  One launch tabulates every combination with the same rounded
// operations as rate_finish, so ranges with gaps need no division at evaluation time (they were 2-3 divisions each).
__global__ void rate_table_kernel(double* __restrict__ tab, uint32_t K, uint32_t step_ms, double range_seconds, uint32_t is_rate,
                                  uint32_t is_counter) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= RATE_FTAB_N + RATE_ZTAB_N) return;
  if (i >= RATE_FTAB_N) {
    // Second table, by (sampled steps, steps to start): the zero-point clamp can fire only when first / result is below
    // (time to the range start) / (sampled time); twice that is a rounding-proof skip test.  Gauges never clamp.
    const uint32_t j = i - RATE_FTAB_N, ds = j % RATE_TAB_DIM, ss = j / RATE_TAB_DIM;
    tab[i] = (is_counter && ss >= 1 && ss + ds <= K) ? __dmul_rn(2.0, __ddiv_rn(seconds_of(ds, step_ms), seconds_of(ss, step_ms))) : 0.0;
    return;
  }
  const uint32_t n = i % RATE_TAB_DIM, ds = (i / RATE_TAB_DIM) % RATE_TAB_DIM, ss = i / (RATE_TAB_DIM * RATE_TAB_DIM);
  // combinations no range has are NaN: the gap kernel sends ranges with fewer than two samples to entries 0 and 1 of row 0
  // and gets its "no result" from the multiplication
  double f = __longlong_as_double(0x7ff8000000000000ll);
  if (ss >= 1 && n >= 2 && ss + ds <= K) {
    const double si = seconds_of(ss, step_ms);
    double dts = seconds_of(ds, step_ms), dte = seconds_of(K - ss - ds, step_ms);
    const double avg = __ddiv_rn(si, __uint2double_rn(n - 1));
    const double thr = __dmul_rn(avg, 1.1);
    const double half = __dmul_rn(avg, 0.5);  // avg / 2 exactly (a power of two): the same bits as the division, a tenth of the instructions
    if (dts >= thr) dts = half;
    double eti = __dadd_rn(si, dts);
    if (dte >= thr) dte = half;
    eti = __dadd_rn(eti, dte);
    f = __ddiv_rn(eti, si);
    if (is_rate) f = __ddiv_rn(f, range_seconds);
  }
  tab[i] = f;
}

// Per-CTA constants in shared memory: the out-of-line slow paths read them through one pointer, the streaming loop
// keeps the few it needs in registers.
struct RateShared {
  double sec_tab[RATE_MAX_KEFF + 3];  // seconds spanned by k grid steps
  double factor_full;                 // extrapolation factor (and 1/range for rate()) when every grid step has a sample
  double zc_full;                     // skip test of the zero-point clamp, see counter_rate_kernel
  double range_seconds;
  const double* ftab;                 // [RATE_TAB_DIM]^3 extrapolation factors by (sampled steps, steps to start, samples)
  uint32_t K, Keff, is_counter, is_rate;
};

// State a warp carries from one 32-sample block of a series to the next.
struct RateCarry {
  double vp;          // previous block's samples (lane l = sample l)
  uint32_t pm_prev;   // which of them exist
  uint32_t rm_prev;   // which of them are counter resets (smaller than the sample before them)
};

__device__ __forceinline__ uint64_t rate_win_mask(uint32_t lane, uint32_t Keff) { return ((1ull << Keff) - 1ull) << (33 + lane - Keff); }

struct RateBlockOut {
  double r;     // this lane's rate (NaN = fewer than two samples in range)
  uint32_t rm;  // reset flags of the block's 32 samples
};

// One block of any shape, for everything the whole-chunk fast path declines (a counter reset in reach, a zero-point
// clamp candidate, gaps, a series' first and last blocks).  Called by the whole warp from warp-uniform branches and
// kept out of line so the streaming loop stays small.  Two tiers: a block whose 64 samples in reach all exist keeps
// the fixed shuffle pattern and only adds the reset terms / the clamp; otherwise positions come from the bit masks.
__device__ __noinline__ RateBlockOut rate_block_any(double vc, double vp, uint32_t pm_cur, uint32_t pm_prev, uint32_t rm_prev,
                                                    const RateShared* sh) {
  const uint32_t lane = threadIdx.x & 31, B = 32 + lane, Keff = sh->Keff, K = sh->K;
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  RateBlockOut o;
  if ((pm_cur & pm_prev) == 0xffffffffu) {
    const uint32_t back = Keff - 1;
    const uint64_t win = rate_win_mask(lane, Keff);
    const double pv = __shfl_sync(0xffffffffu, lane == 31 ? vp : vc, (lane + 31) & 31);
    o.rm = sh->is_counter ? __ballot_sync(0xffffffffu, vc < pv) : 0u;
    const double vf = __shfl_sync(0xffffffffu, (lane + back >= 32 && back) ? vp : vc, (lane - back) & 31);
    double res = __dsub_rn(vc, vf);
    uint64_t rr = ((uint64_t)o.rm << 32) | rm_prev;
    while (rr) {  // warp-uniform: add what the counter dropped from (the step before each reset), in time order
      const int bpos = __ffsll((long long)rr) - 1;
      rr &= rr - 1;
      const int pp = bpos ? bpos - 1 : 0;
      const double cb = pp >= 32 ? __shfl_sync(0xffffffffu, vc, pp - 32) : __shfl_sync(0xffffffffu, vp, pp);
      if (((win >> bpos) & 1ull) && (uint32_t)bpos != B - back) res = __dadd_rn(res, cb);
    }
    if (Keff < 2)
      o.r = qnan;
    else if (vf >= 0.0 && vf < __dmul_rn(res, sh->zc_full))  // right after a reset: the start-side extrapolation is clamped
      o.r = rate_finish(res, vf, sh->sec_tab[back], sh->sec_tab[K - back], sh->sec_tab[0], Keff, sh->is_counter, sh->is_rate,
                        sh->range_seconds);
    else
      o.r = __dmul_rn(res, sh->factor_full);
    return o;
  }
  // ---- gaps: positions come from the bit masks.  Bit b of prev:cur is grid sample b of the 64 in reach; this lane's
  //      own sample is bit B and its range reaches back Keff-1 <= 32 samples, so the earlier part of the range fits one
  //      32-bit word E (bit i = sample s+i, s = B-Keff+1) and popc / ffs / clz are single instructions ----
  const bool have = vc == vc;
  const uint32_t s = 33 + lane - Keff;  // 0..62
  const uint32_t emask = Keff >= 33 ? 0xffffffffu : (1u << (Keff - 1)) - 1u;
  const uint32_t E = __funnelshift_r(s < 32 ? pm_prev : pm_cur, s < 32 ? pm_cur : 0u, s) & emask;
  // the nearest earlier sample: in this block when one exists below this lane, else the previous block's last
  const uint32_t c = pm_cur & ((1u << lane) - 1u);
  const double pvc = __shfl_sync(0xffffffffu, vc, (31 - __clz(c)) & 31);
  const double pvp = __shfl_sync(0xffffffffu, vp, (31 - __clz(pm_prev)) & 31);
  const double pv = c ? pvc : pvp;
  o.rm = __ballot_sync(0xffffffffu, sh->is_counter && have && (c | pm_prev) && vc < pv);
  const uint32_t n = __popc(E) + (have ? 1u : 0u);
  const uint32_t fi = E ? (uint32_t)__ffs(E) - 1u : Keff - 1;       // first sample of the range, relative to s
  const uint32_t li = have ? Keff - 1 : (uint32_t)(31 - __clz(E));  // last one (meaningless when the range is empty: n < 2)
  const uint32_t f = s + fi;
  const double fa = __shfl_sync(0xffffffffu, vc, f & 31), fb = __shfl_sync(0xffffffffu, vp, f & 31);
  const double vf = f >= 32 ? fa : fb;
  // without a sample of its own the last of the range is the nearest earlier sample (were that outside the range, n = 0)
  double res = __dsub_rn(have ? vc : pv, vf);
  uint64_t rr = ((uint64_t)o.rm << 32) | rm_prev;
  if (rr) {  // warp-uniform and rare: a counter reset in reach
    const uint64_t M = ((uint64_t)pm_cur << 32) | pm_prev, PW = M & rate_win_mask(lane, Keff);
    while (rr) {
      const int bpos = __ffsll((long long)rr) - 1;
      rr &= rr - 1;
      const uint64_t below = M & ((1ull << bpos) - 1ull);
      const int pp = below ? 63 - __clzll((long long)below) : 0;
      const double cb = pp >= 32 ? __shfl_sync(0xffffffffu, vc, pp - 32) : __shfl_sync(0xffffffffu, vp, pp);
      if (((PW >> bpos) & 1ull) && (uint32_t)bpos != f) res = __dadd_rn(res, cb);
    }
  }
  const uint32_t span = li - fi, lead = fi + K + 1 - Keff;  // grid steps first..last, range start..first
  if (n < 2)
    o.r = qnan;
  else if (E == emask && have && !(vf >= 0.0 && vf < __dmul_rn(res, sh->zc_full)))
    // this lane's own range has every sample (the gap lies outside it): constant factor, no divisions
    o.r = __dmul_rn(res, sh->factor_full);
  else if (sh->is_counter && res > 0.0 && vf >= 0.0 &&
           __dmul_rn(vf, sh->sec_tab[span]) < __dmul_rn(__dmul_rn(res, 2.0), sh->sec_tab[lead]))
    // could be clamped at the zero point (that needs first / result < time to the range start / sampled time; twice
    // that is a rounding-proof bound): the full formula
    o.r = rate_finish(res, vf, sh->sec_tab[span], sh->sec_tab[lead], sh->sec_tab[Keff - 1 - li], n, sh->is_counter, sh->is_rate,
                      sh->range_seconds);
  else
    o.r = __dmul_rn(res, __ldg(sh->ftab + (span * RATE_TAB_DIM + lead) * RATE_TAB_DIM + n));
  return o;
}

// One block through rate_block_any, with the carry update.
__device__ __forceinline__ double rate_block(double vc, uint32_t pm_cur, RateCarry& st, const RateShared* sh) {
  const RateBlockOut o = rate_block_any(vc, st.vp, pm_cur, st.pm_prev, st.rm_prev, sh);
  st.vp = vc;
  st.pm_prev = pm_cur;
  st.rm_prev = o.rm;
  return o.r;
}

// Two instantiations per output shape.  GAPS = false streams: its whole-chunk fast path assumes every sample is there
// and sends everything else block by block through rate_block_any.  GAPS = true evaluates every block from the bit
// masks, four side by side, and calls out only for counter resets and clamp candidates; it costs about twice the
// instructions per block, whatever the data.  With tier 0 both are launched and the one that does not suit the
// slab's gap statistic returns at once (a few microseconds; no host round trip to read the statistic).
template <bool STRIDE1, bool WANT64, bool GAPS>
__global__ void __launch_bounds__(RATE_THREADS, 3) counter_rate_kernel(const RateParams p) {
  if (p.tier == 0) {
    const bool gappy = p.gapstat[0] * 1000000ull > p.gapstat[1] * RATE_GAP_PPM;
    if (gappy != GAPS) return;
  }
  __shared__ RateShared sh;
  __shared__ __align__(16) double ring_all[RATE_WARPS][RATE_STAGES * RATE_CHUNK_ELEMS];
  __shared__ uint64_t bar_all[RATE_WARPS][RATE_STAGES];
  if (threadIdx.x < RATE_MAX_KEFF + 3) sh.sec_tab[threadIdx.x] = seconds_of(threadIdx.x, p.step_ms);
  const uint32_t lane = threadIdx.x & 31;
  double* ring = ring_all[threadIdx.x >> 5];
  uint64_t* bar = bar_all[threadIdx.x >> 5];
  if (lane == 0) {
    for (int i = 0; i < RATE_STAGES; ++i) mbar_init(&bar[i], 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    sh.K = p.K;
    sh.Keff = p.Keff;
    sh.is_counter = p.is_counter;
    sh.is_rate = p.is_rate;
    sh.range_seconds = p.range_seconds;
    sh.ftab = p.ftab;
    double factor = 0.0, zc = 0.0;
    if (p.Keff >= 2) {  // a range in which every grid step has a sample: n = Keff, first at the range's first step, last at its end
      const double si = sh.sec_tab[p.Keff - 1];
      double dts = sh.sec_tab[p.K - p.Keff + 1], dte = sh.sec_tab[0];
      const double avg = __ddiv_rn(si, __uint2double_rn(p.Keff - 1));
      const double thr = __dmul_rn(avg, 1.1);
      if (dts >= thr) dts = __ddiv_rn(avg, 2.0);
      if (dte >= thr) dte = __ddiv_rn(avg, 2.0);
      factor = __ddiv_rn(__dadd_rn(__dadd_rn(si, dts), dte), si);
      if (p.is_rate) factor = __ddiv_rn(factor, p.range_seconds);
      // The zero-point clamp fires only when first / result < dts / si; twice that is a rounding-proof skip test
      // (it also implies result > 0).  Gauges never clamp: a zero threshold makes the test always false.
      zc = p.is_counter ? __dmul_rn(2.0, __ddiv_rn(dts, si)) : 0.0;
    }
    sh.factor_full = factor;
    sh.zc_full = zc;
  }
  __syncthreads();
  const double factor_full = sh.factor_full, zc_full = sh.zc_full;
  const uint32_t back = p.Keff - 1;  // steps from an evaluation back to the first sample of a full range
  const uint32_t src_prev = (lane + 31) & 31, src_first = (lane - back) & 31;
  const bool from_prev = lane + back >= 32 && back > 0;  // this lane's first-of-range lies in the previous block
  // gap kernel: the part of this lane's range before its own sample, as bits gs.. of prev:cur (see rate_block_any)
  const uint32_t gs = 33 + lane - p.Keff, emask = p.Keff >= 33 ? 0xffffffffu : (1u << back) - 1u, lane_lt = (1u << lane) - 1u;
  const uint32_t lead0 = p.K + 1 - p.Keff, lane_bit = 1u << lane, counter_mask = p.is_counter ? 0xffffffffu : 0u;
  const int lane_back = (int)lane - (int)back;
  const uint32_t ring_a = (uint32_t)__cvta_generic_to_shared(ring);
  // chunks in flight per warp besides the one in work.  The gap kernel keeps one fewer: the stage consumed last stays
  // intact for a round, so "the block before" is always readable in the ring and positions become plain shared-memory reads
  constexpr int AHEAD = GAPS ? RATE_STAGES - 2 : RATE_STAGES - 1;
  const double* __restrict__ ztab = p.ftab + RATE_FTAB_N;
  uint32_t cc = 0;  // chunks this warp has consumed so far: stage = cc % RATE_STAGES, phase parity = (cc / RATE_STAGES) & 1

  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  const uint32_t n_blocks = (p.n_time + 31) / 32;
  // blocks [k_lo, k_hi) need no bounds checks: all their samples exist in the slab and all their evaluations are emitted
  const uint32_t k_lo = (p.first_eval + 31) / 32, k_hi = p.n_time / 32;
  const uint32_t q32 = 32 / p.stride, r32 = 32 % p.stride;  // how an output index moves when e advances by a block
  const uint64_t warp0 = (uint64_t)blockIdx.x * RATE_WARPS + (threadIdx.x >> 5), n_warps = (uint64_t)gridDim.x * RATE_WARPS;

  for (uint64_t item = warp0; item < p.n_items; item += n_warps) {
    const uint32_t row = (uint32_t)(item / p.segs_per_row), seg = (uint32_t)(item - (uint64_t)row * p.segs_per_row);
    const double* __restrict__ x = p.v + (size_t)row * p.n_time;
    float* __restrict__ orow = p.out + (size_t)row * p.out_pitch;
    double* __restrict__ orow64 = WANT64 ? p.out64 + (size_t)row * p.n_out : nullptr;
    const int kb0 = (int)(p.kb_first + seg * p.seg_blocks);
    const int kb1 = min((int)n_blocks, kb0 + (int)p.seg_blocks);
    auto load_checked = [&](int k) -> double {
      const long long i = (long long)k * 32 + lane;
      return (k >= 0 && i < (long long)p.n_time) ? ldg_stream_f64(x + i) : qnan;
    };
    auto emit_checked = [&](int k, double r) {
      const uint32_t e = (uint32_t)k * 32 + lane;
      if (e >= p.first_eval && e < p.n_time) {
        uint32_t j = e - p.first_eval;
        if (!STRIDE1) {
          const uint32_t q = j / p.stride;
          if (q * p.stride != j) return;
          j = q;
        }
        orow[j] = __double2float_rn(r);
        if (WANT64) orow64[j] = r;
      }
    };
    // Two blocks of history: kb0-2 seeds the carry, kb0-1 runs without output so its reset flags exist.
    RateCarry st;
    st.vp = load_checked(kb0 - 2);
    st.pm_prev = __ballot_sync(0xffffffffu, st.vp == st.vp);
    st.rm_prev = 0;
    {
      const double w = load_checked(kb0 - 1);
      (void)rate_block(w, __ballot_sync(0xffffffffu, w == w), st, &sh);
    }

    int k = kb0;
    while (k < kb1) {
      const int fast_end = min(kb1, (int)k_hi);
      if (k >= (int)k_lo && k + RATE_UNROLL <= fast_end) {
        // ---- streaming part: chunks of RATE_UNROLL blocks arrive in this warp's shared-memory ring by bulk copy;
        //      RATE_STAGES-1 chunks are in flight while one is being evaluated ----
        const int n_chunks = (fast_end - k) / RATE_UNROLL;
        const uint64_t g0 = (uint64_t)row * p.n_time + (uint64_t)k * 32;  // first element of the first chunk
        const uint32_t off = (uint32_t)(g0 & 1);                          // bulk copies start 16-byte aligned
        const double* src = p.v + (g0 - off);
        uint32_t j = (uint32_t)k * 32 + lane - p.first_eval, ph = 0;
        if (!STRIDE1) {
          const uint32_t q = j / p.stride;
          ph = j - q * p.stride;
          j = q;
        }
        float* op = orow + j;  // gap kernel, stride 1: this lane's next output
        double* op64 = WANT64 ? orow64 + j : nullptr;
        if constexpr (GAPS) {
          // the block before the first chunk lives in registers: park it where the ring would have it (the last block
          // of the stage before), which no copy touches until the second chunk is in work
          __syncwarp();
          ring[((cc + RATE_STAGES - 1) % RATE_STAGES) * RATE_CHUNK_ELEMS + off + 32 * (RATE_UNROLL - 1) + lane] = st.vp;
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
        }
        if (lane == 0) {
          for (int ci = 0; ci < AHEAD && ci < n_chunks; ++ci) {
            const uint32_t sg = (cc + ci) % RATE_STAGES;
            mbar_expect_tx(&bar[sg], RATE_CHUNK_BYTES);
            bulk_g2s(ring + sg * RATE_CHUNK_ELEMS, src + (size_t)ci * (32 * RATE_UNROLL), RATE_CHUNK_BYTES, &bar[sg]);
          }
        }
        for (int ci = 0; ci < n_chunks; ++ci, ++cc) {
          const uint32_t sg = cc % RATE_STAGES;
          mbar_wait(&bar[sg], (cc / RATE_STAGES) & 1);
          double cur[RATE_UNROLL];
#pragma unroll
          for (int u = 0; u < RATE_UNROLL; ++u) cur[u] = ring[sg * RATE_CHUNK_ELEMS + off + 32 * u + lane];
          __syncwarp();  // every lane has read the stage consumed last round before it is refilled
          if (lane == 0 && ci + AHEAD < n_chunks) {
            const uint32_t ng = (cc + AHEAD) % RATE_STAGES;
            mbar_expect_tx(&bar[ng], RATE_CHUNK_BYTES);
            bulk_g2s(ring + ng * RATE_CHUNK_ELEMS, src + (size_t)(ci + AHEAD) * (32 * RATE_UNROLL), RATE_CHUNK_BYTES, &bar[ng]);
          }
          auto emit = [&](double r) {
            if (STRIDE1) {
              orow[j] = __double2float_rn(r);
              if (WANT64) orow64[j] = r;
              j += 32;
            } else {
              // predicated store (no divergent region around it): only every stride-th evaluation is an output
              asm volatile("{\n\t.reg .pred q;\n\tsetp.eq.u32 q, %2, 0;\n\t@q st.global.f32 [%0], %1;\n\t}" ::"l"(orow + j),
                           "f"(__double2float_rn(r)), "r"(ph)
                           : "memory");
              if (WANT64 && ph == 0) orow64[j] = r;
              ph += r32;
              j += q32;
              if (ph >= p.stride) ph -= p.stride, ++j;
            }
          };
          if constexpr (GAPS) {
            // Shared-memory byte addresses and selects on integers throughout: nothing here may become a branch (a divergent
            // warp would put a convergence check in front of every vote).
            uint32_t pm[RATE_UNROLL], rm[RATE_UNROLL], bw[RATE_UNROLL];
            double r[RATE_UNROLL];
            const uint32_t a_stage0 = ring_a + (sg * RATE_CHUNK_ELEMS + off) * 8u;  // sample 0 of block 0
            // sample 0 of the block before block 0: the last block of the stage consumed last round
            const uint32_t a_before0 = ring_a + (((cc + RATE_STAGES - 1) % RATE_STAGES) * RATE_CHUNK_ELEMS + off + 32 * (RATE_UNROLL - 1)) * 8u;
#pragma unroll
            for (int u = 0; u < RATE_UNROLL; ++u) pm[u] = __ballot_sync(0xffffffffu, cur[u] == cur[u]);
#pragma unroll
            for (int u = 0; u < RATE_UNROLL; ++u) {
              // a_cb + 8 i is sample i of this block; negative i reaches into the block before, which for u > 0 lies right
              // in front of it in the stage
              const uint32_t a_cb = a_stage0 + 256u * u;
              const double vc = cur[u];
              const uint32_t pmc = pm[u], pmb = u ? pm[u - 1] : st.pm_prev;
              const bool have = pmc & lane_bit;
              const uint32_t E = (gs < 32 ? __funnelshift_r(pmb, pmc, gs) : pmc >> (gs & 31)) & emask;
              // the nearest earlier sample: in this block when there is one below this lane, else the last of the block before
              const uint32_t c = pmc & lane_lt;
              // (pmb | 1, E | bit 31 below: the same positions whenever there is a sample to find, and an address inside the
              // ring when there is none — the value is not used then)
              const int bc = bfind_u32(c), bb = bfind_u32(pmb | 1u);
              const uint32_t a_pv = u ? a_cb + 8 * (c ? bc : bb - 32) : (c ? a_cb : a_before0) + 8 * (c ? bc : bb);
              const double pv = lds_f64(a_pv);
              // (a lane without a sample compares NaN: false)
              rm[u] = __ballot_sync(0xffffffffu, (c | pmb) && vc < pv) & counter_mask;
              const uint32_t n = __popc(E) + (have ? 1u : 0u);
              // first and last sample of the range in steps from its start; with fewer than two samples they are not used
              const int fi = __ffs(E | 0x80000000u) - 1, hb = bfind_u32(E), li = have ? (int)back : hb;
              const int rel = lane_back + fi;  // the first one as a sample index of this block
              const uint32_t a_vf = (u == 0 && rel < 0 ? a_before0 + 256u : a_cb) + 8 * rel;
              const double vf = lds_f64(a_vf);
              const double res = __dsub_rn(have ? vc : pv, vf);
              // table row (sampled steps, steps to start); row 0 (and its entries n = 0, 1) holds NaN factors and a zero threshold
              const uint32_t t = n >= 2 ? (uint32_t)(li * (int)RATE_TAB_DIM - fi * (int)(RATE_TAB_DIM - 1)) + lead0 : 0u;
              r[u] = __dmul_rn(res, __ldg(p.ftab + (t * RATE_TAB_DIM + n)));
              // out of line: a reset in reach (its drop has to be added) or a lane that may be clamped at the zero point
              const bool cand = vf >= 0.0 && vf < __dmul_rn(res, __ldg(ztab + t));
              bw[u] = __ballot_sync(0xffffffffu, cand) | rm[u] | (u ? rm[u - 1] : st.rm_prev);
            }
            auto put = [&](int u, double v) {
              if constexpr (STRIDE1) {
                op[32 * u] = __double2float_rn(v);
                if (WANT64) op64[32 * u] = v;
              } else {
                emit(v);
              }
            };
            if ((bw[0] | bw[1] | bw[2] | bw[3]) == 0) {
#pragma unroll
              for (int u = 0; u < RATE_UNROLL; ++u) put(u, r[u]);
              st.vp = cur[RATE_UNROLL - 1];
              st.pm_prev = pm[RATE_UNROLL - 1];
              st.rm_prev = 0;
            } else {
#pragma unroll
              for (int u = 0; u < RATE_UNROLL; ++u) {
                if (bw[u]) {
                  put(u, rate_block(cur[u], pm[u], st, &sh));
                } else {
                  put(u, r[u]);
                  st.vp = cur[u];
                  st.pm_prev = pm[u];
                  st.rm_prev = 0;
                }
              }
            }
            op += 32 * RATE_UNROLL;
            if (WANT64) op64 += 32 * RATE_UNROLL;
          } else {
            // Whole-chunk fast path: the four blocks are evaluated side by side (independent shuffles and votes).  Three
            // warp-uniform tests guard it: every grid step of the five blocks in reach has a sample; no counter reset in
            // reach; no lane is a zero-point clamp candidate.  Anything else goes block by block through rate_block_any.
            uint32_t pm[RATE_UNROLL], pm_and = st.pm_prev;
#pragma unroll
            for (int u = 0; u < RATE_UNROLL; ++u) {
              pm[u] = __ballot_sync(0xffffffffu, cur[u] == cur[u]);
              pm_and &= pm[u];
            }
            // (the shuffles below run on whatever the blocks hold; blocks with gaps are re-done by rate_block_any)
            double res[RATE_UNROLL];
            uint32_t rm[RATE_UNROLL], rm_or = st.rm_prev, clamp_bits = 0;
#pragma unroll
            for (int u = 0; u < RATE_UNROLL; ++u) {
              const double before = u ? cur[u - 1] : st.vp;
              // the sample before this lane's: one rotate, with the previous block's last sample parked in lane 31
              const double pv = __shfl_sync(0xffffffffu, lane == 31 ? before : cur[u], src_prev);
              rm[u] = p.is_counter ? __ballot_sync(0xffffffffu, cur[u] < pv) : 0u;
              rm_or |= rm[u];
              // the first sample of this lane's range, `back` steps earlier
              const double vf = __shfl_sync(0xffffffffu, from_prev ? before : cur[u], src_first);
              res[u] = __dsub_rn(cur[u], vf);
              clamp_bits |= (uint32_t)(vf >= 0.0 && vf < __dmul_rn(res[u], zc_full)) << u;
            }
            if (pm_and == 0xffffffffu && rm_or == 0 && back > 0 && !__any_sync(0xffffffffu, clamp_bits)) {
#pragma unroll
              for (int u = 0; u < RATE_UNROLL; ++u) emit(__dmul_rn(res[u], factor_full));
              st.vp = cur[RATE_UNROLL - 1];
              st.pm_prev = 0xffffffffu;
              st.rm_prev = 0;
            } else {
              // block by block: a block keeps the inline result unless a gap, a reset or a clamp candidate is in ITS reach
#pragma unroll
              for (int u = 0; u < RATE_UNROLL; ++u) {
                const bool plain = (pm[u] & st.pm_prev) == 0xffffffffu && (rm[u] | st.rm_prev) == 0 && back > 0 &&
                                   !__any_sync(0xffffffffu, clamp_bits & (1u << u));
                if (plain) {
                  emit(__dmul_rn(res[u], factor_full));
                  st.vp = cur[u];
                  st.pm_prev = pm[u];
                  st.rm_prev = 0;
                } else {
                  emit(rate_block(cur[u], pm[u], st, &sh));
                }
              }
            }
                  }
        }
        k += n_chunks * RATE_UNROLL;
      } else {
        const double w = load_checked(k);
        emit_checked(k, rate_block(w, __ballot_sync(0xffffffffu, w == w), st, &sh));
        ++k;
      }
    }
  }
}

// One evaluation, serially: two passes over its range (first / last / count, then the drops).  Any range length, any
// gaps; the reference evaluation every other path has to match bit for bit.
struct RateEvalArgs {  // by value: a reference to the kernel's parameter struct would force a copy of it onto the stack
  uint32_t K, Keff, step_ms, is_counter, is_rate;
  double range_seconds;
};
__device__ __forceinline__ RateEvalArgs rate_eval_args(const RateParams& p) {
  return RateEvalArgs{p.K, p.Keff, p.step_ms, p.is_counter, p.is_rate, p.range_seconds};
}
__device__ __noinline__ double rate_eval_serial(const double* __restrict__ x, long long e, const RateEvalArgs p) {
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
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
  if (n < 2) return qnan;
  const double vf = x[f];
  double res = __dsub_rn(x[l], vf);
  if (p.is_counter) {
    double prev = vf;
    for (long long i = f + 1; i <= l; ++i) {
      const double v = x[i];
      if (v == v) {
        if (v < prev) res = __dadd_rn(res, prev);
        prev = v;
      }
    }
  }
  return rate_finish(res, vf, seconds_of((uint32_t)(l - f), p.step_ms), seconds_of((uint32_t)(f - (e - (long long)p.K)), p.step_ms),
                     seconds_of((uint32_t)(e - l), p.step_ms), n, p.is_counter, p.is_rate, p.range_seconds);
}

// Any range length: one thread per evaluation.
__global__ void __launch_bounds__(256) counter_rate_generic_kernel(const RateParams p) {
  const uint64_t total = (uint64_t)p.n_series * p.n_out;
  for (uint64_t o = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; o < total; o += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t row = (uint32_t)(o / p.n_out), j = (uint32_t)(o - (uint64_t)row * p.n_out);
    const double r = rate_eval_serial(p.v + (size_t)row * p.n_time, (long long)p.first_eval + (long long)j * p.stride, rate_eval_args(p));
    p.out[(size_t)row * p.out_pitch + j] = __double2float_rn(r);
    if (p.out64) p.out64[o] = r;
  }
}

// Ranges of more than 33 samples (rate(x[5m]) at a 5 s step is 60, [1h] at 15 s is 240), streaming.  In a stretch of
// the series without gaps and counter resets an evaluation needs two samples — its own and the one Keff-1 steps earlier —
// whatever the range length, so the kernel streams like the short-range one: a warp walks a series in chunks of four
// 32-sample blocks through a bulk-copy ring, here five stages deep with two chunks in flight, which leaves the two chunks
// before the one in work intact: 256 samples of history, read straight from shared memory.  A warp-uniform counter knows
// how many consecutive blocks before the current one were clean (every sample there, none smaller than its predecessor);
// a block whose reach is clean, and in which no lane is a zero-point clamp candidate, costs one subtraction and one
// multiplication per evaluation.  Every other block — a gap or a reset within reach, a clamp candidate, a range cut by the
// start of the series — only gets its bit set in a bitmap; counter_rate_dirty_kernel evaluates those afterwards.
constexpr uint32_t RATE_LONG_MAX = 256;
constexpr int RATE_LSTAGES = 5, RATE_LAHEAD = 2;

// One 32-evaluation block of any shape for ranges of up to RATE_LONG_MAX samples, by the whole warp and without state: the
// (need + 1) blocks in reach are loaded from global memory (they were streamed a moment ago: cache hits), their presence
// and reset masks rebuilt by ballots (a reset = a sample smaller than the nearest earlier sample, carried across blocks),
// then every lane finds count / first / last of its own range with bit operations on those masks, fetches the two values,
// adds the drops of the resets in its range in time order (a warp-uniform loop: resets are rare) and finishes with the
// full formula.  ~300 instructions per block, against Keff x 16 per LANE for the serial evaluation it replaces.
// A reset whose predecessor lies before the blocks in reach is not flagged: it could only be the first sample of a range
// that holds it, and the first sample's drop never counts.
constexpr int RATE_REACH = RATE_LONG_MAX / 32 + 1;  // blocks a range can touch

__device__ __forceinline__ double rate_block_coop(const double* __restrict__ x, int kb, uint32_t n_time, const RateEvalArgs a) {
  const uint32_t lane = threadIdx.x & 31, back = a.Keff - 1, lane_lt = (1u << lane) - 1u;
  const int need = (int)((back + 31) / 32);
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  uint32_t pm[RATE_REACH], rm[RATE_REACH];  // [j]: the block j blocks before kb
  double vj[RATE_REACH];
#pragma unroll
  for (int j = 0; j < RATE_REACH; ++j) {  // all the loads first: one memory latency, not one per block
    const int i = (kb - j) * 32 + (int)lane;  // (n_time < 2^31: positions fit an int)
    vj[j] = (j <= need && i >= 0 && i < (int)n_time) ? x[i] : qnan;
  }
  double carry = 0.0;  // the last sample seen so far, walking the reach oldest block first
  bool have_carry = false;
#pragma unroll
  for (int j = RATE_REACH - 1; j >= 0; --j) {
    pm[j] = rm[j] = 0u;
    if (j <= need) {  // warp-uniform
      const double v = vj[j];
      const uint32_t m = __ballot_sync(0xffffffffu, v == v);
      const uint32_t lower = m & lane_lt;
      const double pv_in = __shfl_sync(0xffffffffu, v, lower ? 31 - __clz(lower) : 0);
      const double pv = lower ? pv_in : carry;
      pm[j] = m;
      rm[j] = a.is_counter ? __ballot_sync(0xffffffffu, v == v && (lower || have_carry) && v < pv) : 0u;
      if (m) {  // warp-uniform
        carry = __shfl_sync(0xffffffffu, v, 31 - __clz(m));
        have_carry = true;
      }
    }
  }
  // this lane's range: samples e - back .. e
  const int e = kb * 32 + (int)lane, ws = e - (int)back;
  uint32_t n = 0;
  int f_pos = -1, l_pos = -1;
#pragma unroll
  for (int j = 0; j < RATE_REACH; ++j) {  // newest block first
    if (j <= need) {
      const int b0 = (kb - j) * 32;
      const int lo = ws - b0, hi = e - b0;  // bits lo .. hi of this block lie in the range (hi >= 0)
      uint32_t mask = 0u;
      if (lo <= 31) mask = (0xffffffffu >> (31 - (uint32_t)(hi < 31 ? hi : 31))) & (0xffffffffu << (uint32_t)(lo > 0 ? lo : 0));
      const uint32_t w = pm[j] & mask;
      n += __popc(w);
      if (w) {
        if (l_pos < 0) l_pos = b0 + 31 - __clz(w);
        f_pos = b0 + __ffs(w) - 1;
      }
    }
  }
  const bool ok = n >= 2;
  const double vf = ok ? x[f_pos] : 0.0;
  double res = __dsub_rn(ok ? x[l_pos] : 0.0, vf);
  // the drops: every reset of the reach, oldest first; a lane adds the sample before the reset when the reset lies in its
  // range and is not its first sample
#pragma unroll
  for (int j = RATE_REACH - 1; j >= 0; --j) {
    if (j <= need) {
      uint32_t rr = rm[j];
      while (rr) {  // warp-uniform
        const int bit = __ffs(rr) - 1;
        rr &= rr - 1;
        const int rpos = (kb - j) * 32 + bit;
        int pp = -1;  // the nearest earlier sample: below `bit` in this block, else the last of an older block
        const uint32_t lowm = pm[j] & ((1u << bit) - 1u);
        if (lowm) pp = (kb - j) * 32 + 31 - __clz(lowm);
#pragma unroll
        for (int jj = j + 1; jj < RATE_REACH; ++jj)
          if (jj <= need && pp < 0 && pm[jj]) pp = (kb - jj) * 32 + 31 - __clz(pm[jj]);
        if (pp >= 0) {
          const double cb = x[pp];
          if (ok && rpos > f_pos && rpos <= e) res = __dadd_rn(res, cb);
        }
      }
    }
  }
  if (!ok) return qnan;
  return rate_finish(res, vf, seconds_of((uint32_t)(l_pos - f_pos), a.step_ms), seconds_of((uint32_t)(f_pos - (e - (int)a.K)), a.step_ms),
                     seconds_of((uint32_t)(e - l_pos), a.step_ms), n, a.is_counter, a.is_rate, a.range_seconds);
}

// Where a row's results go: everything rate_chunk_coop needs to store evaluations, by value.
struct RateEmit {
  float* orow;
  double* orow64;  // or NULL
  uint32_t first_eval, n_time, stride;
};
__device__ __forceinline__ void rate_emit(const RateEmit& o, int k, double r) {  // the evaluation at grid step 32 k + lane, if it is one
  const uint32_t e = (uint32_t)k * 32 + (threadIdx.x & 31);
  if (e >= o.first_eval && e < o.n_time) {
    uint32_t j = e - o.first_eval;
    if (o.stride != 1) {
      const uint32_t q = j / o.stride;
      if (q * o.stride != j) return;
      j = q;
    }
    o.orow[j] = __double2float_rn(r);
    if (o.orow64) o.orow64[j] = r;
  }
}
// Second kernel of the long-range path: the blocks the streaming kernel declined (it only sets their bit in `dirty`, one
// 32-bit word per 32 blocks of a row; any call out of its loop cost the clean path a third of its speed), each through
// rate_block_coop.  A warp takes a word at a time.
__global__ void __launch_bounds__(256) counter_rate_dirty_kernel(const RateParams p) {
  const uint32_t wpr = (((p.n_time + 31) / 32) + 31) / 32;  // bitmap words per row
  const uint64_t n_words = (uint64_t)p.n_series * wpr;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const RateEvalArgs a = rate_eval_args(p);
  for (uint64_t w = warp0; w < n_words; w += n_warps) {
    uint32_t m = p.dirty[w];  // warp-uniform
    if (!m) continue;
    const uint32_t row = (uint32_t)(w / wpr), word = (uint32_t)(w - (uint64_t)row * wpr);
    const double* __restrict__ x = p.v + (size_t)row * p.n_time;
    const RateEmit out{p.out + (size_t)row * p.out_pitch, p.out64 ? p.out64 + (size_t)row * p.n_out : nullptr, p.first_eval, p.n_time, p.stride};
    while (m) {
      const int kb = (int)(word * 32 + (uint32_t)__ffs(m) - 1);
      m &= m - 1;
      if (p.Keff <= RATE_LONG_MAX) {  // warp-uniform
        rate_emit(out, kb, rate_block_coop(x, kb, p.n_time, a));
      } else {  // beyond the cooperative routine's reach: every lane walks its own range
        const uint32_t e = (uint32_t)kb * 32 + (threadIdx.x & 31);
        double r = 0.0;
        if (e >= p.first_eval && e < p.n_time && (p.stride == 1 || (e - p.first_eval) % p.stride == 0)) r = rate_eval_serial(x, (long long)e, a);
        rate_emit(out, kb, r);
      }
    }
  }
}

template <bool STRIDE1, bool FAR>  // FAR: ranges longer than the ring remembers (more than RATE_LONG_MAX samples)
__global__ void __launch_bounds__(RATE_THREADS, 3) counter_rate_long_kernel(const RateParams p) {
  __shared__ __align__(16) double ring_all[RATE_WARPS][RATE_LSTAGES * RATE_CHUNK_ELEMS];
  __shared__ uint64_t bar_all[RATE_WARPS][RATE_LSTAGES];
  const uint32_t lane = threadIdx.x & 31;
  double* ring = ring_all[threadIdx.x >> 5];
  uint64_t* bar = bar_all[threadIdx.x >> 5];
  if (lane == 0) {
    for (int i = 0; i < RATE_LSTAGES; ++i) mbar_init(&bar[i], 1);
    mbar_fence_init();
  }
  __syncthreads();
  // a range in which every grid step has a sample: n = Keff, first at the range's first step, last at its end
  double factor_full, zc_full;
  {
    const double si = seconds_of(p.Keff - 1, p.step_ms);
    double dts = seconds_of(p.K - p.Keff + 1, p.step_ms), dte = seconds_of(0, p.step_ms);
    const double avg = __ddiv_rn(si, __uint2double_rn(p.Keff - 1));
    const double thr = __dmul_rn(avg, 1.1);
    if (dts >= thr) dts = __ddiv_rn(avg, 2.0);
    if (dte >= thr) dte = __ddiv_rn(avg, 2.0);
    factor_full = __ddiv_rn(__dadd_rn(__dadd_rn(si, dts), dte), si);
    if (p.is_rate) factor_full = __ddiv_rn(factor_full, p.range_seconds);
    zc_full = p.is_counter ? __dmul_rn(2.0, __ddiv_rn(dts, si)) : 0.0;  // see counter_rate_kernel
  }
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  const uint32_t back = p.Keff - 1;
  const int need = (int)((back + 31) / 32);  // clean blocks wanted in front of a block for its fast path
  const bool head_ok = p.first_eval >= back;  // no evaluation's range is cut by the start of the series
  const uint32_t wpr = (((p.n_time + 31) / 32) + 31) / 32;  // words per row of the dirty-block bitmap
  const uint32_t src_prev = (lane + 31) & 31;
  const uint32_t k_lo = (p.first_eval + 31) / 32, k_hi = p.n_time / 32, n_blocks = (p.n_time + 31) / 32;
  const uint64_t warp0 = (uint64_t)blockIdx.x * RATE_WARPS + (threadIdx.x >> 5), n_warps = (uint64_t)gridDim.x * RATE_WARPS;
  uint32_t cc = 0;  // chunks consumed so far by this warp: stage = cc % RATE_LSTAGES, parity = (cc / RATE_LSTAGES) & 1

  for (uint64_t item = warp0; item < p.n_items; item += n_warps) {
    const uint32_t row = (uint32_t)(item / p.segs_per_row), seg = (uint32_t)(item - (uint64_t)row * p.segs_per_row);
    const double* __restrict__ x = p.v + (size_t)row * p.n_time;
    float* __restrict__ orow = p.out + (size_t)row * p.out_pitch;
    double* __restrict__ orow64 = p.out64 ? p.out64 + (size_t)row * p.n_out : nullptr;
    const int kb0 = (int)(p.kb_first + seg * p.seg_blocks);
    const int kb1 = min((int)n_blocks, kb0 + (int)p.seg_blocks);
    const RateEmit out{orow, orow64, p.first_eval, p.n_time, STRIDE1 ? 1u : p.stride};
    uint32_t* drow = p.dirty + (size_t)row * wpr;
    // the stream starts `need` blocks (rounded up to whole chunks) before the first block with evaluations, so that the
    // history is in the ring and its cleanliness is known; those warm-up blocks are not evaluated
    const int warm = min(kb0, (need + RATE_UNROLL - 1) / RATE_UNROLL * RATE_UNROLL);
    int k = kb0 - warm;
    const int stream_end = min(kb1, (int)k_hi);
    const int n_chunks = stream_end > k ? (stream_end - k) / RATE_UNROLL : 0;
    // consecutive clean blocks right before block k.  Nothing precedes the series — no gap and no reset there — so a stream
    // that starts at block 0 starts with all the clean history it needs (ranges that begin before sample 0 are kept off the
    // fast path by `whole` below)
    int clean_run = k == 0 ? need : 0;
    double last_prev = qnan;  // the sample right before block k
    if (k > 0) last_prev = ldg_stream_f64(x + (size_t)k * 32 - 1);
    if (n_chunks > 0) {
      const uint64_t g0 = (uint64_t)row * p.n_time + (uint64_t)k * 32;
      const uint32_t off = (uint32_t)(g0 & 1);  // bulk copies start 16-byte aligned
      const double* src = p.v + (g0 - off);
      __syncwarp();
      if (lane == 0) {
        for (int ci = 0; ci < RATE_LAHEAD && ci < n_chunks; ++ci) {
          const uint32_t sg = (cc + ci) % RATE_LSTAGES;
          mbar_expect_tx(&bar[sg], RATE_CHUNK_BYTES);
          bulk_g2s(ring + sg * RATE_CHUNK_ELEMS, src + (size_t)ci * (32 * RATE_UNROLL), RATE_CHUNK_BYTES, &bar[sg]);
        }
      }
      for (int ci = 0; ci < n_chunks; ++ci, ++cc) {
        const uint32_t sg = cc % RATE_LSTAGES;
        mbar_wait(&bar[sg], (cc / RATE_LSTAGES) & 1);
        double cur[RATE_UNROLL];
#pragma unroll
        for (int u = 0; u < RATE_UNROLL; ++u) cur[u] = ring[sg * RATE_CHUNK_ELEMS + off + 32 * u + lane];
        __syncwarp();  // every lane is done with the chunk consumed three rounds ago: its stage may be refilled
        if (lane == 0 && ci + RATE_LAHEAD < n_chunks) {
          const uint32_t ng = (cc + RATE_LAHEAD) % RATE_LSTAGES;
          mbar_expect_tx(&bar[ng], RATE_CHUNK_BYTES);
          bulk_g2s(ring + ng * RATE_CHUNK_ELEMS, src + (size_t)(ci + RATE_LAHEAD) * (32 * RATE_UNROLL), RATE_CHUNK_BYTES, &bar[ng]);
        }
        // the four blocks side by side (independent shuffles, votes and shared-memory reads); a chunk that is clean as a
        // whole, behind a clean history, is finished with four multiplications and four stores
        const int kc = k + ci * RATE_UNROLL;  // first block of the chunk
        uint32_t pm[RATE_UNROLL], rm[RATE_UNROLL], cand = 0;
        double res[RATE_UNROLL];
#pragma unroll
        for (int u = 0; u < RATE_UNROLL; ++u) {
          const double before = u ? cur[u - 1] : last_prev;  // only lane 31's matters
          const double pv = __shfl_sync(0xffffffffu, lane == 31 ? before : cur[u], src_prev);
          pm[u] = __ballot_sync(0xffffffffu, cur[u] == cur[u]);
          // (a sample smaller than its predecessor; NaN compares false.  Gauges have no resets.)
          rm[u] = p.is_counter ? __ballot_sync(0xffffffffu, cur[u] < pv) : 0u;
          // the first sample of the range: `back` samples earlier, in this chunk or one of the two before it.  (Where the
          // history is not in the ring yet the index is meaningless but inside the ring; such blocks are not fast.)
          const int rel = ci * (32 * RATE_UNROLL) + 32 * u + (int)lane - (int)back;  // relative to the run's first sample
          // its stage: 0, 1 or 2 chunks before this one (rel >= 128 ci - 255, also where it is meaningless)
          const int dch = ci - (rel >> 7);
          const uint32_t fsg = sg >= (uint32_t)dch ? sg - (uint32_t)dch : sg + RATE_LSTAGES - (uint32_t)dch;
          double vf;
          if (!FAR) {
            vf = ring[fsg * RATE_CHUNK_ELEMS + off + (rel & 127)];
          } else {  // ranges longer than the ring remembers: the same sample from global memory (streamed not long ago: L2)
            const long long ef = ((long long)kc + u) * 32 + lane - back;
            vf = ef >= 0 ? ldg_stream_f64(x + ef) : 0.0;
          }
          res[u] = __dsub_rn(cur[u], vf);
          cand |= (uint32_t)(vf >= 0.0 && vf < __dmul_rn(res[u], zc_full)) << u;
        }
        // which blocks hold a clamp candidate (warp-uniform)
        const uint32_t cand_bits = __reduce_or_sync(0xffffffffu, cand);
        const bool chunk_clean = (pm[0] & pm[1] & pm[2] & pm[3]) == 0xffffffffu && (rm[0] | rm[1] | rm[2] | rm[3]) == 0u;
        if (STRIDE1 && chunk_clean && !cand_bits && clean_run >= need && kc >= (int)k_lo && kc >= kb0 && (head_ok || (uint32_t)kc * 32u >= back)) {
          float* op = orow + ((size_t)kc * 32 + lane - p.first_eval);
#pragma unroll
          for (int u = 0; u < RATE_UNROLL; ++u) op[32 * u] = __double2float_rn(__dmul_rn(res[u], factor_full));
          if (orow64) {
            double* op64 = orow64 + ((size_t)kc * 32 + lane - p.first_eval);
#pragma unroll
            for (int u = 0; u < RATE_UNROLL; ++u) op64[32 * u] = __dmul_rn(res[u], factor_full);
          }
          clean_run += RATE_UNROLL;
        } else {
          // anything else, block by block: a clean block behind enough clean history still takes the two-sample form (this
          // is also where the first chunk of a series ends up: it starts before the first evaluation); the rest is left to
          // the second kernel
#pragma unroll
          for (int u = 0; u < RATE_UNROLL; ++u) {
            const int kb = kc + u;
            const bool clean = pm[u] == 0xffffffffu && rm[u] == 0u;
            if (kb >= kb0) {
              const bool whole = head_ok || (uint32_t)kb * 32u >= back;  // no emitted range is cut by the start of the series
              if (clean && clean_run >= need && whole && !(cand_bits & (1u << u))) rate_emit(out, kb, __dmul_rn(res[u], factor_full));
              else if (lane == 0) atomicOr(drow + (kb >> 5), 1u << (kb & 31));
            }
            clean_run = clean ? clean_run + 1 : 0;
          }
        }
        last_prev = __shfl_sync(0xffffffffu, cur[RATE_UNROLL - 1], 31);
      }
      k += n_chunks * RATE_UNROLL;
    }
    // fewer than four whole blocks left, and the cut block at the series' end: the same two tiers, samples from global memory
    // (blocks before kb0 that the stream did not reach — a row or segment shorter than its warm-up — are only looked at)
    for (; k < kb1; ++k) {
      const long long i = (long long)k * 32 + lane;
      const bool in = i < (long long)p.n_time;
      const double c = in ? ldg_stream_f64(x + i) : qnan;
      const double pv = __shfl_sync(0xffffffffu, lane == 31 ? last_prev : c, src_prev);
      const uint32_t live = __ballot_sync(0xffffffffu, in);  // lanes of the block that are samples of the series (a prefix)
      const uint32_t pm = __ballot_sync(0xffffffffu, c == c);
      const uint32_t rm = p.is_counter ? __ballot_sync(0xffffffffu, c < pv) : 0u;
      const bool clean = pm == live && rm == 0u;  // what exists of the block is clean (nothing after it is ever evaluated)
      bool fast = clean && clean_run >= need && (head_ok || (uint32_t)k * 32u >= back) && i >= (long long)back;
      double res = 0.0;
      fast = __all_sync(0xffffffffu, fast || !in);
      if (fast) {
        const double vf = in ? ldg_stream_f64(x + i - back) : 0.0;
        res = __dsub_rn(c, vf);
        fast = !__any_sync(0xffffffffu, in && vf >= 0.0 && vf < __dmul_rn(res, zc_full));
      }
      if (k >= kb0) {
        if (fast) rate_emit(out, k, __dmul_rn(res, factor_full));
        else if (lane == 0) atomicOr(drow + (k >> 5), 1u << (k & 31));
      }
      clean_run = (clean && live == 0xffffffffu) ? clean_run + 1 : 0;
      last_prev = __shfl_sync(0xffffffffu, c, 31);
    }
  }
}

// Adds rows [first, first+count) to the slab's gap statistic (small slabs are looked at whole).
cudaError_t launch_gapstat(igpu_ctx* ctx, igpu_counters* c, uint32_t first, uint32_t count) {
  const uint64_t n = (uint64_t)count * c->n_time;
  if (n < 16) return cudaSuccess;
  const uint32_t every = n / 16 >= (1u << 16) ? 64u : 1u;
  const uint64_t threads = n / 16 / every * 16;
  const unsigned grid = (unsigned)std::min<uint64_t>((threads + 255) / 256, (uint64_t)ctx->sm_count * 8);
  counters_gapstat_kernel<<<grid, 256, 0, ctx->stream>>>(c->d_v + (size_t)first * c->n_time, n, every, c->d_gapstat);
  igpu_count_launch(ctx);
  return cudaGetLastError();
}

}  // namespace

extern "C" {

int32_t igpu_counters_create(igpu_ctx* ctx, uint32_t n_series, uint32_t n_time, igpu_counters** out) {
  if (!ctx || !out) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_counters_create: NULL argument");
  *out = nullptr;
  IGPU_REQUIRE(ctx, n_time > 0, IGPU_ERR_ARG, "n_time must be positive");
  IGPU_REQUIRE(ctx, n_time < 0x7FFFFFE0u, IGPU_ERR_RANGE, "rows of %u samples are too long (sample positions are 31-bit)", n_time);
  DeviceGuard g(ctx->device);
  igpu_counters* c = new igpu_counters();
  c->n_series = n_series;
  c->n_time = n_time;
  // +64: the rate kernel's bulk copies may read up to two elements past the last sample (alignment slack)
  const size_t bytes = (size_t)n_series * n_time * sizeof(double) + 64;
  if (cudaMalloc(&c->d_v, bytes) != cudaSuccess ||
      cudaMalloc(&c->d_ftab, (size_t)(RATE_FTAB_N + RATE_ZTAB_N) * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&c->d_gapstat, 2 * sizeof(unsigned long long)) != cudaSuccess) {
    if (c->d_v) cudaFree(c->d_v);
    if (c->d_ftab) cudaFree(c->d_ftab);
    cudaGetLastError();
    delete c;
    return igpu_fail(ctx, IGPU_ERR_OOM, "device allocation of %zu counter bytes failed", bytes);
  }
  cudaMemset(c->d_gapstat, 0, 2 * sizeof(unsigned long long));
  *out = c;
  return IGPU_OK;
}

int32_t igpu_counters_destroy(igpu_ctx* ctx, igpu_counters* c) {
  if (!ctx || !c) return IGPU_OK;
  DeviceGuard g(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (c->d_v) cudaFree(c->d_v);
  if (c->d_ftab) cudaFree(c->d_ftab);
  if (c->d_gapstat) cudaFree(c->d_gapstat);
  if (c->d_dirty) cudaFree(c->d_dirty);
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
  if (bytes) IGPU_CUDA(ctx, launch_gapstat(ctx, c, first, count));
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
    IGPU_CUDA(ctx, cudaMemsetAsync(c->d_gapstat, 0, 2 * sizeof(unsigned long long), ctx->stream));
    IGPU_CUDA(ctx, launch_gapstat(ctx, c, 0, c->n_series));
  }
  return IGPU_OK;
}

int32_t igpu_counters_gap_stat(igpu_ctx* ctx, const igpu_counters* c, uint64_t* missing, uint64_t* looked_at) {
  if (!ctx || !c || !missing || !looked_at) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_counters_gap_stat: NULL argument");
  DeviceGuard g(ctx->device);
  unsigned long long st[2] = {0, 0};
  IGPU_CUDA(ctx, cudaMemcpyAsync(st, c->d_gapstat, sizeof st, cudaMemcpyDeviceToHost, ctx->stream));
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *missing = st[0];
  *looked_at = st[1];
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
  IGPU_REQUIRE(ctx, dst->n_series == c->n_series && dst->n_time_user == n_out, IGPU_ERR_ARG,
               "destination slab is %u x %u, the rate of %u x %u counters is %u x %u", dst->n_series, dst->n_time_user, c->n_series, c->n_time,
               c->n_series, n_out);
  DeviceGuard g(ctx->device);
  RateParams p;
  memset(&p, 0, sizeof p);
  p.v = c->d_v;
  p.out = dst->d_x;
  p.n_series = c->n_series;
  p.n_time = c->n_time;
  p.n_out = n_out;
  p.out_pitch = dst->n_time;
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
    const bool warp_path = p.Keff <= RATE_MAX_KEFF && !ctx->sw.rate_force_generic;
    if (warp_path) {  // factor table of this (grid step, range, rate/increase) — 39,304 entries, a few microseconds
      igpu_counters* cm = const_cast<igpu_counters*>(c);
      const uint32_t tab_n = RATE_FTAB_N + RATE_ZTAB_N;
      rate_table_kernel<<<(tab_n + 255) / 256, 256, 0, ctx->stream>>>(cm->d_ftab, p.K, p.step_ms, p.range_seconds, p.is_rate, p.is_counter);
      IGPU_CUDA(ctx, cudaGetLastError());
      igpu_count_launch(ctx);
      p.ftab = cm->d_ftab;
    }
    const bool long_path = !warp_path && !ctx->sw.rate_force_generic;  // any longer range (the generic kernel stays as the cross-check)
    igpu_kbegin(ctx, IGPU_K_COUNTER_RATE);
    if (warp_path || long_path) {
      const uint32_t n_blocks = (c->n_time + 31) / 32;
      p.kb_first = rp->first_eval / 32;
      // A work item is a run of blocks of one series; every item re-reads two blocks of history and restarts the
      // copy pipeline, so items are whole series when there are enough series to balance the warps (16 per warp),
      // and series are cut (never below 16 blocks) only when there are few of them.
      const uint32_t row_blocks = n_blocks - p.kb_first;
      const uint64_t warp_slots = (uint64_t)ctx->sm_count * 3 * RATE_WARPS;
      uint64_t segs = c->n_series ? (16 * warp_slots + c->n_series - 1) / c->n_series : 1;
      segs = std::max<uint64_t>(1, std::min<uint64_t>(segs, (row_blocks + 15) / 16));
      p.seg_blocks = (uint32_t)((row_blocks + segs - 1) / segs);
      if (ctx->sw.rate_seg) p.seg_blocks = (uint32_t)std::max(4, ctx->sw.rate_seg);
      p.seg_blocks = (p.seg_blocks + RATE_UNROLL - 1) / RATE_UNROLL * RATE_UNROLL;
      p.segs_per_row = (row_blocks + p.seg_blocks - 1) / p.seg_blocks;
      p.n_items = (uint64_t)c->n_series * p.segs_per_row;
      const uint64_t want = (p.n_items + RATE_WARPS - 1) / RATE_WARPS;
      const int grid_mult = ctx->sw.rate_grid ? std::max(1, ctx->sw.rate_grid) : 3;
      const unsigned grid = (unsigned)std::min<uint64_t>(want, (uint64_t)ctx->sm_count * grid_mult);
      p.gapstat = c->d_gapstat;
      p.tier = (uint32_t)std::min(2, std::max(0, ctx->sw.rate_tier));
      if (long_path) {  // 34 .. RATE_LONG_MAX samples per range: the streaming kernel with its history in the ring,
                        // then the blocks it declined
        igpu_counters* cm = const_cast<igpu_counters*>(c);
        const size_t dirty_bytes = (size_t)c->n_series * ((n_blocks + 31) / 32) * sizeof(uint32_t);
        if (!cm->d_dirty) {
          cudaError_t ae = cudaMalloc(&cm->d_dirty, std::max<size_t>(dirty_bytes, 4));
          if (ae != cudaSuccess) {
            cudaGetLastError();
            if (d_out64) cudaFree(d_out64);
            return igpu_fail(ctx, IGPU_ERR_OOM, "device allocation of %zu bitmap bytes failed", dirty_bytes);
          }
        }
        cudaMemsetAsync(cm->d_dirty, 0, dirty_bytes, ctx->stream);
        p.dirty = cm->d_dirty;
        const bool far = p.Keff > RATE_LONG_MAX;
        void (*lk)(const RateParams) = p.stride == 1 ? (far ? counter_rate_long_kernel<true, true> : counter_rate_long_kernel<true, false>)
                                                     : (far ? counter_rate_long_kernel<false, true> : counter_rate_long_kernel<false, false>);
        lk<<<grid, RATE_THREADS, 0, ctx->stream>>>(p);
        igpu_count_launch(ctx);
        counter_rate_dirty_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(p);
        igpu_count_launch(ctx);
      }
      for (int gaps = 0; warp_path && gaps < 2; ++gaps) {
        if (p.tier && (int)p.tier != gaps + 1) continue;
        void (*kern)(const RateParams);
        if (p.stride == 1 && !d_out64)
          kern = gaps ? counter_rate_kernel<true, false, true> : counter_rate_kernel<true, false, false>;
        else if (p.stride == 1)
          kern = gaps ? counter_rate_kernel<true, true, true> : counter_rate_kernel<true, true, false>;
        else if (!d_out64)
          kern = gaps ? counter_rate_kernel<false, false, true> : counter_rate_kernel<false, false, false>;
        else
          kern = gaps ? counter_rate_kernel<false, true, true> : counter_rate_kernel<false, true, false>;
        kern<<<grid, RATE_THREADS, 0, ctx->stream>>>(p);
        igpu_count_launch(ctx);
      }
    } else {
      const unsigned grid = (unsigned)std::min<uint64_t>((n_res + 255) / 256, (uint64_t)ctx->sm_count * 16);
      counter_rate_generic_kernel<<<grid, 256, 0, ctx->stream>>>(p);
      igpu_count_launch(ctx);
    }
    cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) {
      if (d_out64) cudaFree(d_out64);
      return igpu_fail(ctx, IGPU_ERR_CUDA, "rate kernel launch failed: %s", cudaGetErrorString(le));
    }
    igpu_kend(ctx, IGPU_K_COUNTER_RATE);
  }
  igpu_mark(ctx, 3);
  // the slab changed: window statistics of an earlier call no longer describe it
  dst->window = 0;
  dst->n_win = 0;
  dst->wcnt_valid = false;
  if (d_out64) {
    cudaError_t ce = cudaMemcpyAsync(host_rate64, d_out64, n_res * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_out64);
    if (ce != cudaSuccess) return igpu_fail(ctx, IGPU_ERR_CUDA, "copy of the float64 rates failed: %s", cudaGetErrorString(ce));
  }
  return IGPU_OK;
}

}  // extern "C"
