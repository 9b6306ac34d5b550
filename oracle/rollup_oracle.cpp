// rollup_oracle.cpp — CPU restatement of the reference's fleet rollup over the columnar layout.
//
// TEST INFRASTRUCTURE ONLY.  Not shipped, not the product: only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference leg may load this library.  The product path
// (libigpu_rollup.so) never links or calls it.
//
// Parity status (see DESIGN.md "Oracle"):
//  * allocation / classification: the arithmetic below follows the cited reference lines with JS
//    `number` semantics (every accumulator is an IEEE double, exactly as in the reference).  It is
//    pinned indirectly: tests/test_oracle_golden.py pins oracle/ref_objects.py (the object-level
//    restatement) against the reference's own known-answer tests (src/api/k8s.test.ts), and
//    tests/test_oracle_columnar.py proves this file == ref_objects.py on generated and adversarial
//    fleets.  Page loops (OverviewPage/NodesPage/NodeDetailSection) and all of metrics.ts have no
//    numeric test in the reference: those rows are "parity unpinned by the reference".
//  * window statistics / quantiles: the reference has no such code (metrics.ts:67-75,101-116 are
//    instant queries); the definitions are SURVEY.md §8(c)'s PromQL-style ones.  PARITY UNPINNED.
//
// Build: see oracle/Makefile (g++ -O2 -fopenmp -ffp-contract=off).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/igpu_rollup.h"

namespace {

// Math.round: nearest integer, ties toward +Infinity (ECMA-262 21.3.2.28).
inline double js_round(double x) {
  if (std::isnan(x) || std::isinf(x)) return x;
  double f = std::floor(x);
  return (x - f) >= 0.5 ? f + 1.0 : f;
}

inline int64_t to_i64(double x) {
  if (std::isnan(x)) return 0;
  if (x >= 9.2e18) return std::numeric_limits<int64_t>::max();
  if (x <= -9.2e18) return std::numeric_limits<int64_t>::min();
  return (int64_t)x;
}

inline uint8_t level_90_70(double pct) { return pct >= 90 ? 2 : (pct >= 70 ? 1 : 0); }

// -- synthetic series generator (DESIGN.md "generator"); exact f32 ops only, no libm, no FMA ----
inline uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
constexpr uint64_t STREAM_SERIES_PARAM = 0x11ull << 56;
constexpr uint64_t STREAM_SAMPLE = 0x12ull << 56;
constexpr uint64_t STREAM_NAN = 0x13ull << 56;

struct SeriesParam {
  float base, tdp;
  uint32_t phase;
};
inline SeriesParam series_param(uint64_t seed, uint64_t gs) {
  uint64_t r = splitmix64(seed ^ STREAM_SERIES_PARAM ^ gs);
  SeriesParam p;
  p.base = 40.0f + (float)(r & 0xFFFF) * (180.0f / 65536.0f);
  uint32_t k = (uint32_t)((r >> 16) & 0xFF) % 3u;
  p.tdp = k == 0 ? 150.0f : (k == 1 ? 250.0f : 300.0f);
  p.phase = (uint32_t)((r >> 24) % 86400u);
  return p;
}
inline float series_sample(uint64_t seed, uint64_t gs, uint32_t n_time, uint32_t t, const SeriesParam& p,
                           uint32_t nan_per_million) {
  uint64_t idx = gs * (uint64_t)n_time + t;
  uint64_t r = splitmix64(seed ^ STREAM_SAMPLE ^ idx);
  if (nan_per_million) {
    uint64_t rn = splitmix64(seed ^ STREAM_NAN ^ idx);
    if ((uint32_t)(rn % 1000000u) < nan_per_million) return std::numeric_limits<float>::quiet_NaN();
  }
  float xi = (float)(uint32_t)(r >> 40) * (1.0f / 8388608.0f) - 1.0f;  // U(-1,1), 24 bits
  uint32_t ph = (t + p.phase) % 86400u;
  float x = (float)ph / 86400.0f;  // day fraction in [0,1)
  float u = 1.0f - x;
  float y = x * u;
  float wave = 8.0f * y - 1.0f;    // parabolic diurnal wave in [-1,1]
  float a = 0.25f * p.base;
  float b = 0.05f * p.base;
  float v = p.base + a * wave;
  v = v + b * xi;
  v = v < 0.0f ? 0.0f : v;
  v = v > p.tdp ? p.tdp : v;
  return v;
}

}  // namespace

extern "C" {

int32_t oracle_version(void) { return (IGPU_VERSION_MAJOR << 16) | (IGPU_VERSION_MINOR << 8) | IGPU_VERSION_PATCH; }

int32_t oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// ---------------------------------------------------------------------------------------------
// Allocation rollup + classification.  Single-threaded when threads <= 1, else OpenMP over pods
// with per-thread node accumulators (merged in thread order; integer-valued doubles, so exact).
// ---------------------------------------------------------------------------------------------
int32_t oracle_alloc_rollup(const igpu_fleet_cols* c, const igpu_fleet_out* o, int32_t threads) {
  if (!c || !o) return IGPU_ERR_ARG;
  const uint32_t N = c->n_nodes, P = c->n_pods, R = c->n_resources;
  const bool want_lim = (c->options & IGPU_OPT_POD_LIMITS) != 0;
  if (R < 4 || R > IGPU_MAX_RESOURCES) return IGPU_ERR_RANGE;
  if (threads < 1) threads = 1;

  std::vector<double> totals(IGPU_TOTALS_LEN, 0.0);
  std::vector<double> node_cnt(N, 0.0), node_inuse(N, 0.0);
  std::vector<double> node_req((size_t)R * N, 0.0), node_lim(want_lim ? (size_t)R * N : 0, 0.0);

#ifdef _OPENMP
  const int T = threads;
#else
  const int T = 1;
#endif
  std::vector<std::vector<double>> t_tot(T, std::vector<double>(IGPU_TOTALS_LEN, 0.0));
  std::vector<std::vector<double>> t_cnt(T), t_inuse(T), t_req(T), t_lim(T);
  for (int t = 0; t < T; ++t) {
    t_cnt[t].assign(N, 0.0);
    t_inuse[t].assign(N, 0.0);
    t_req[t].assign((size_t)R * N, 0.0);
    if (want_lim) t_lim[t].assign((size_t)R * N, 0.0);
  }

#pragma omp parallel num_threads(T)
  {
#ifdef _OPENMP
    const int tid = omp_get_thread_num();
#else
    const int tid = 0;
#endif
    std::vector<double>& tot = t_tot[tid];
    double req[IGPU_MAX_RESOURCES], lim[IGPU_MAX_RESOURCES];
#pragma omp for schedule(static)
    for (int64_t p = 0; p < (int64_t)P; ++p) {
      const uint32_t e0 = c->pod_entry_off[p], e1 = c->pod_entry_off[p + 1];
      for (uint32_t r = 0; r < R; ++r) req[r] = lim[r] = 0.0;
      uint32_t req_mask = 0, lim_mask = 0;
      double inuse_main = 0.0;
      uint64_t seen_main[4] = {0, 0, 0, 0};  // which spec.containers indices own a prefixed key (PodsPage.tsx:50-54)
      for (uint32_t e = e0; e < e1; ++e) {
        const igpu_entry& en = c->entries[e];
        if (!(en.flags & IGPU_ENTRY_INIT)) seen_main[en.container >> 6] |= 1ull << (en.container & 63);
        const double v = (double)en.value;  // parseInt(value, 10) || 0, done by the marshal layer
        if (!(en.flags & IGPU_ENTRY_LIMIT)) {
          if (en.flags & IGPU_ENTRY_TRUTHY) {  // k8s.ts:295 `&& value`
            req[en.res] += v;                  // k8s.ts:296
            req_mask |= 1u << en.res;
          }
          // NodeDetailSection.tsx:78-86: main containers only, i915/xe, no truthiness test
          // (parseInt('') || 0 is 0 anyway, which is what the marshal layer stored).
          if (!(en.flags & IGPU_ENTRY_INIT) && (en.res == IGPU_RES_I915 || en.res == IGPU_RES_XE)) inuse_main += v;
        } else if (en.flags & IGPU_ENTRY_TRUTHY) {
          lim[en.res] += v;
          lim_mask |= 1u << en.res;
        }
      }
      const bool is_gpu = e1 > e0;  // k8s.ts:257-263: any prefixed key, requests or limits
      if (o->pod_gpu_containers) {  // gpuContainers.length, PodDetailSection.tsx:48-52,106-108
        int n = 0;
        for (int w = 0; w < 4; ++w) n += __builtin_popcountll(seen_main[w]);
        o->pod_gpu_containers[p] = (uint8_t)n;
      }
      if (o->pod_is_gpu) o->pod_is_gpu[p] = is_gpu ? 1 : 0;
      if (o->pod_req_mask) o->pod_req_mask[p] = req_mask;
      if (o->pod_req)
        for (uint32_t r = 0; r < R; ++r) o->pod_req[(size_t)r * P + p] = to_i64(req[r]);
      if (want_lim) {
        if (o->pod_lim_mask) o->pod_lim_mask[p] = lim_mask;
        if (o->pod_lim)
          for (uint32_t r = 0; r < R; ++r) o->pod_lim[(size_t)r * P + p] = to_i64(lim[r]);
      }
      if (!is_gpu) continue;  // everything below iterates gpuPods (IntelGpuDataContext.tsx:204-208)
      tot[IGPU_T_GPU_PODS] += 1;
      const uint8_t ph = c->pod_phase[p];
      if (ph <= IGPU_PHASE_OTHER) tot[IGPU_T_PHASE0 + ph] += 1;  // OverviewPage.tsx:122-130
      const bool running = ph == IGPU_PHASE_RUNNING;
      if (running) {
        // OverviewPage.tsx:105-113: String(total) -> parseInt -> || 0 is the identity on these integers
        tot[IGPU_T_ALLOCATED] += req[IGPU_RES_I915] + req[IGPU_RES_XE];
        for (uint32_t r = 0; r < R; ++r) {
          tot[IGPU_T_REQ0 + r] += req[r];
          if (want_lim) tot[IGPU_T_LIM0 + r] += lim[r];
        }
      }
      const uint32_t nd = c->pod_node[p];
      if (nd == IGPU_NODE_NONE) continue;  // NodesPage.tsx:155
      tot[IGPU_T_SCHEDULED_PODS] += 1;
      t_cnt[tid][nd] += 1;  // NodesPage.tsx:156-158, any phase
      if (running) {
        t_inuse[tid][nd] += inuse_main;  // NodeDetailSection.tsx:77
        for (uint32_t r = 0; r < R; ++r) {
          t_req[tid][(size_t)r * N + nd] += req[r];
          if (want_lim) t_lim[tid][(size_t)r * N + nd] += lim[r];
        }
      }
    }
  }
  for (int t = 0; t < T; ++t) {
    for (int i = 0; i < IGPU_TOTALS_LEN; ++i) totals[i] += t_tot[t][i];
    for (uint32_t n = 0; n < N; ++n) {
      node_cnt[n] += t_cnt[t][n];
      node_inuse[n] += t_inuse[t][n];
    }
    for (size_t i = 0; i < (size_t)R * N; ++i) {
      node_req[i] += t_req[t][i];
      if (want_lim) node_lim[i] += t_lim[t][i];
    }
  }

  // node pass
  for (uint32_t n = 0; n < N; ++n) {
    const uint8_t f = c->node_flags[n];
    const bool is_gpu = (f & (IGPU_NODE_DISCRETE_ROLE | IGPU_NODE_IGPU_ROLE | IGPU_NODE_NFD_LABEL |
                              IGPU_NODE_HAS_GPU_CAPACITY)) != 0;                       // k8s.ts:134-149
    const uint8_t type = (f & IGPU_NODE_DISCRETE_ROLE) ? IGPU_TYPE_DISCRETE
                         : (f & IGPU_NODE_IGPU_ROLE)   ? IGPU_TYPE_INTEGRATED
                                                       : IGPU_TYPE_UNKNOWN;            // k8s.ts:187-191
    const double gpu_count = (double)c->node_cap_i915[n] + (double)c->node_cap_xe[n];  // k8s.ts:174-178
    const bool cap_nan = f & IGPU_NODE_CAP_NAN, alloc_nan = f & IGPU_NODE_ALLOC_NAN;
    const double nanv = std::numeric_limits<double>::quiet_NaN();
    const double total_cap = cap_nan ? nanv : gpu_count;                               // NodesPage.tsx:173
    const double total_alloc =
        alloc_nan ? nanv : (double)c->node_alloc_i915[n] + (double)c->node_alloc_xe[n];  // NodesPage.tsx:174
    // NodesPage.tsx:270-273 + :35-38
    const double bar_alloc = (std::isnan(total_alloc) || total_alloc == 0) ? gpu_count : total_alloc;
    uint8_t oflags = (is_gpu ? IGPU_NODEOUT_IS_GPU : 0) | ((f & IGPU_NODE_READY) ? IGPU_NODEOUT_READY : 0) |
                     (cap_nan ? IGPU_NODEOUT_CAP_NAN : 0) | (alloc_nan ? IGPU_NODEOUT_ALLOC_NAN : 0);
    double bar_pct = 0;
    uint8_t bar_level = 0;
    if (bar_alloc == 0) {
      oflags |= IGPU_NODEOUT_BAR_DASH;
    } else {
      bar_pct = std::min(100.0, js_round(node_cnt[n] / bar_alloc * 100));
      bar_level = level_90_70(bar_pct);
    }
    // NodeDetailSection.tsx:72-76, :89-91
    const double own = (double)c->node_alloc_own[n];
    const double util = own > 0 ? js_round(node_inuse[n] / own * 100) : 0.0;

    if (o->node_out_flags) o->node_out_flags[n] = oflags;
    if (o->node_type) o->node_type[n] = type;
    if (o->node_gpu_count) o->node_gpu_count[n] = to_i64(gpu_count);
    if (o->node_cap_total) o->node_cap_total[n] = to_i64(total_cap);
    if (o->node_alloc_total) o->node_alloc_total[n] = to_i64(total_alloc);
    if (o->node_pod_count) o->node_pod_count[n] = (uint32_t)node_cnt[n];
    if (o->node_bar_pct) o->node_bar_pct[n] = to_i64(bar_pct);
    if (o->node_bar_level) o->node_bar_level[n] = bar_level;
    if (o->node_inuse) o->node_inuse[n] = to_i64(node_inuse[n]);
    if (o->node_util_pct) o->node_util_pct[n] = to_i64(util);
    if (o->node_util_level) o->node_util_level[n] = level_90_70(util);
    for (uint32_t r = 0; r < R; ++r) {
      if (o->node_req) o->node_req[(size_t)r * N + n] = to_i64(node_req[(size_t)r * N + n]);
      if (want_lim && o->node_lim) o->node_lim[(size_t)r * N + n] = to_i64(node_lim[(size_t)r * N + n]);
    }
    if (!is_gpu) continue;  // OverviewPage iterates gpuNodes only
    totals[IGPU_T_GPU_NODES] += 1;
    totals[IGPU_T_DISCRETE + type] += 1;        // OverviewPage.tsx:79-82
    totals[IGPU_T_GPU_DEVICES] += gpu_count;    // :84
    if (f & IGPU_NODE_READY) totals[IGPU_T_READY_NODES] += 1;  // :85
    if (cap_nan) totals[IGPU_T_CAP_NAN] += 1; else totals[IGPU_T_CAPACITY] += gpu_count;         // :98
    if (alloc_nan) totals[IGPU_T_ALLOC_NAN] += 1; else totals[IGPU_T_ALLOCATABLE] += total_alloc;  // :99
  }
  // OverviewPage.tsx:115-116: NaN > 0 is false, so a poisoned capacity gives 0
  {
    const double cap = totals[IGPU_T_CAP_NAN] != 0 ? std::numeric_limits<double>::quiet_NaN() : totals[IGPU_T_CAPACITY];
    totals[IGPU_T_UTIL_PCT] = cap > 0 ? js_round(totals[IGPU_T_ALLOCATED] / cap * 100) : 0.0;
  }
  if (o->fleet_totals)
    for (int i = 0; i < IGPU_TOTALS_LEN; ++i) o->fleet_totals[i] = to_i64(totals[i]);
  return IGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Instant power join + fleet sums: metrics.ts:118-149, MetricsPage.tsx:330-332, :50-53.
// ---------------------------------------------------------------------------------------------
int32_t oracle_power_join_sum(const igpu_power_cols* c, const igpu_power_out* o) {
  if (!c || !o) return IGPU_ERR_ARG;
  const uint32_t NONE = IGPU_NODE_NONE;
  std::vector<int64_t> inst_row(c->n_instance_ids, -1), pw_row(c->n_chip_ids, -1), mx_row(c->n_chip_ids, -1);
  for (uint32_t i = 0; i < c->n_uname; ++i)  // Map.set: last write wins (metrics.ts:123)
    if (c->uname_instance[i] != NONE) inst_row[c->uname_instance[i]] = i;
  for (uint32_t i = 0; i < c->n_energy; ++i)
    if (c->energy_chip[i] != NONE) pw_row[c->energy_chip[i]] = i;  // metrics.ts:130
  for (uint32_t i = 0; i < c->n_power_max; ++i)
    if (c->power_max_chip[i] != NONE) mx_row[c->power_max_chip[i]] = i;  // metrics.ts:137
  double total = 0, max_total = 0;
  for (uint32_t i = 0; i < c->n_chips; ++i) {  // metrics.ts:141-149
    uint8_t fl = 0;
    double w = 0, m = 0;
    uint32_t node = NONE;
    const uint32_t chip = c->chip_chip[i], inst = c->chip_instance[i];
    if (chip != NONE && pw_row[chip] >= 0) { fl |= IGPU_POWER_HAS_WATTS; w = c->energy_value[pw_row[chip]]; }
    if (chip != NONE && mx_row[chip] >= 0) { fl |= IGPU_POWER_HAS_MAX; m = c->power_max_value[mx_row[chip]]; }
    if (inst != NONE && inst_row[inst] >= 0) { fl |= IGPU_POWER_HAS_NODE; node = c->uname_node[inst_row[inst]]; }
    if (o->chip_flags) o->chip_flags[i] = fl;
    if (o->chip_watts) o->chip_watts[i] = w;
    if (o->chip_max_watts) o->chip_max_watts[i] = m;
    if (o->chip_node) o->chip_node[i] = node;
    total += (fl & IGPU_POWER_HAS_WATTS) ? w : 0.0;      // MetricsPage.tsx:330 (`?? 0`: NaN stays NaN)
    max_total += (fl & IGPU_POWER_HAS_MAX) ? m : 0.0;    // MetricsPage.tsx:331
  }
  if (o->fleet) {
    o->fleet[0] = total;
    o->fleet[1] = max_total;
    // PowerBar: maxWatts = maxTotal > 0 ? maxTotal : null; pct = min(100, round(w / max * 100)) or null
    o->fleet[2] = (max_total > 0) ? std::fmin(100.0, js_round(total / max_total * 100))
                                  : std::numeric_limits<double>::quiet_NaN();
    if (max_total > 0 && std::isnan(total)) o->fleet[2] = std::numeric_limits<double>::quiet_NaN();
    o->fleet[3] = (double)c->n_chips;
  }
  return IGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Window statistics (PromQL min/avg/max_over_time over non-overlapping windows).  UNPINNED.
// ---------------------------------------------------------------------------------------------
int32_t oracle_window_stats(const float* x, uint32_t n_series, uint32_t n_time, uint32_t window, float* out_min,
                            float* out_mean, float* out_max, int32_t threads) {
  if (!x || window == 0) return IGPU_ERR_ARG;
  const uint32_t nw = (n_time + window - 1) / window;
  if (threads < 1) threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
  for (int64_t s = 0; s < (int64_t)n_series; ++s) {
    const float* row = x + (size_t)s * n_time;
    for (uint32_t w = 0; w < nw; ++w) {
      const uint32_t t0 = w * window, t1 = std::min(n_time, t0 + window);
      float mn = std::numeric_limits<float>::infinity(), mx = -std::numeric_limits<float>::infinity();
      double sum = 0;
      uint32_t cnt = 0;
      for (uint32_t t = t0; t < t1; ++t) {
        const float v = row[t];
        if (v != v) continue;  // NaN = missing sample
        mn = v < mn ? v : mn;
        mx = v > mx ? v : mx;
        sum += (double)v;
        ++cnt;
      }
      const size_t o = (size_t)s * nw + w;
      const float nanf_ = std::numeric_limits<float>::quiet_NaN();
      if (out_min) out_min[o] = cnt ? mn : nanf_;
      if (out_max) out_max[o] = cnt ? mx : nanf_;
      if (out_mean) out_mean[o] = cnt ? (float)(sum / (double)cnt) : nanf_;
    }
  }
  return IGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Per-node quantile (PromQL quantile_over_time) + node stats + fleet vector.  UNPINNED.
// ---------------------------------------------------------------------------------------------
int32_t oracle_quantile_by_node(const float* x, uint32_t n_series, uint32_t n_time, uint32_t n_nodes,
                                const uint32_t* node_series_off, double q, int32_t mode, double* out_q,
                                igpu_node_series_stats* out_stats, double* out_fleet, int32_t threads) {
  if (!x || !node_series_off || !(q >= 0.0 && q <= 1.0)) return IGPU_ERR_ARG;
  if (node_series_off[n_nodes] > n_series) return IGPU_ERR_ARG;
  if (threads < 1) threads = 1;
  std::vector<igpu_node_series_stats> st(n_nodes);
#pragma omp parallel num_threads(threads)
  {
    std::vector<float> buf;
#pragma omp for schedule(dynamic, 1)
    for (int64_t n = 0; n < (int64_t)n_nodes; ++n) {
      const size_t b = (size_t)node_series_off[n] * n_time, e = (size_t)node_series_off[n + 1] * n_time;
      buf.clear();
      buf.reserve(e - b);
      double sum = 0;
      float mn = std::numeric_limits<float>::infinity(), mx = -std::numeric_limits<float>::infinity();
      for (size_t i = b; i < e; ++i) {
        const float v = x[i];
        if (v != v) continue;
        buf.push_back(v);
        sum += (double)v;
        mn = v < mn ? v : mn;
        mx = v > mx ? v : mx;
      }
      const size_t cnt = buf.size();
      double res = std::numeric_limits<double>::quiet_NaN();
      if (cnt) {
        if (mode == 1) {  // nearest rank: x[ceil(q*n) - 1]
          double r = std::ceil(q * (double)cnt) - 1.0;
          size_t k = r < 0 ? 0 : (size_t)r;
          if (k >= cnt) k = cnt - 1;
          std::nth_element(buf.begin(), buf.begin() + k, buf.end());
          res = (double)buf[k];
        } else {  // linear interpolation between closest ranks
          const double rank = q * (double)(cnt - 1);
          const size_t lo = (size_t)std::floor(rank);
          const size_t hi = std::min(cnt - 1, lo + 1);
          const double w = rank - (double)lo;
          std::nth_element(buf.begin(), buf.begin() + lo, buf.end());
          const double xlo = (double)buf[lo];
          double xhi = xlo;
          if (hi != lo) xhi = (double)*std::min_element(buf.begin() + lo + 1, buf.end());
          res = xlo * (1.0 - w) + xhi * w;
        }
      }
      if (out_q) out_q[n] = res;
      st[n].sum = sum;
      st[n].count = cnt;
      st[n].min = cnt ? mn : std::numeric_limits<float>::quiet_NaN();
      st[n].max = cnt ? mx : std::numeric_limits<float>::quiet_NaN();
    }
  }
  double fsum = 0, fcnt = 0, fmin = std::numeric_limits<double>::infinity(), fmax = -fmin, fn = 0;
  for (uint32_t n = 0; n < n_nodes; ++n) {
    if (out_stats) out_stats[n] = st[n];
    fsum += st[n].sum;
    fcnt += (double)st[n].count;
    if (st[n].count) {
      fmin = std::fmin(fmin, (double)st[n].min);
      fmax = std::fmax(fmax, (double)st[n].max);
      fn += 1;
    }
  }
  if (out_fleet) {
    for (int i = 0; i < IGPU_FLEET_LEN; ++i) out_fleet[i] = 0;
    out_fleet[IGPU_F_SUM] = fsum;
    out_fleet[IGPU_F_COUNT] = fcnt;
    out_fleet[IGPU_F_MIN] = fn ? fmin : std::numeric_limits<double>::quiet_NaN();
    out_fleet[IGPU_F_MAX] = fn ? fmax : std::numeric_limits<double>::quiet_NaN();
    out_fleet[IGPU_F_NODES] = fn;
  }
  return IGPU_OK;
}

// Synthetic series rows [0, n_series) of the slab whose first row is fleet-wide row global_first.
int32_t oracle_generate_series(float* out, uint32_t n_series, uint32_t n_time, uint64_t seed, uint64_t global_first,
                               uint32_t nan_per_million, int32_t threads) {
  if (!out) return IGPU_ERR_ARG;
  if (threads < 1) threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
  for (int64_t s = 0; s < (int64_t)n_series; ++s) {
    const uint64_t gs = global_first + (uint64_t)s;
    const SeriesParam p = series_param(seed, gs);
    float* row = out + (size_t)s * n_time;
    for (uint32_t t = 0; t < n_time; ++t) row[t] = series_sample(seed, gs, n_time, t, p, nan_per_million);
  }
  return IGPU_OK;
}

}  // extern "C"

// ---- N3: Prometheus extrapolatedRate on a regular grid (test infrastructure) ---------------------------
// Restates promql/functions.go extrapolatedRate (Prometheus v3.x; prometheus/prometheus is a server the
// reference queries at metrics.ts:106-109, it is not vendored in /root/reference) with explicit integer
// millisecond timestamps, one evaluation at a time, the way the engine does.  NaN in the slab = no sample.
struct RateSample {
  int64_t t_ms;
  double v;
};

static double extrapolated_rate(const RateSample* smp, size_t n, int64_t range_start_ms, int64_t range_end_ms, double range_seconds,
                                bool is_counter, bool is_rate) {
  if (n < 2) return std::numeric_limits<double>::quiet_NaN();  // "no sense in trying to compute a rate without at least two points"
  const size_t n_minus_one = n - 1;
  const int64_t first_t = smp[0].t_ms, last_t = smp[n_minus_one].t_ms;
  double result = smp[n_minus_one].v - smp[0].v;
  if (is_counter) {
    double prev = smp[0].v;
    for (size_t i = 1; i < n; ++i) {
      if (smp[i].v < prev) result += prev;  // counter reset: the counter restarted from zero
      prev = smp[i].v;
    }
  }
  double duration_to_start = (double)(first_t - range_start_ms) / 1000.0;
  double duration_to_end = (double)(range_end_ms - last_t) / 1000.0;
  const double sampled_interval = (double)(last_t - first_t) / 1000.0;
  const double avg_between = sampled_interval / (double)n_minus_one;
  const double threshold = avg_between * 1.1;
  double extrapolate_to = sampled_interval;
  if (duration_to_start >= threshold) duration_to_start = avg_between / 2.0;
  if (is_counter && result > 0 && smp[0].v >= 0) {
    const double duration_to_zero = sampled_interval * (smp[0].v / result);
    if (duration_to_zero < duration_to_start) duration_to_start = duration_to_zero;
  }
  extrapolate_to += duration_to_start;
  if (duration_to_end >= threshold) duration_to_end = avg_between / 2.0;
  extrapolate_to += duration_to_end;
  double factor = extrapolate_to / sampled_interval;
  if (is_rate) factor /= range_seconds;
  return result * factor;
}

extern "C" {

uint32_t oracle_counters_rate_len(uint32_t n_time, const igpu_rate_params* p) {
  if (!p || !p->eval_stride || p->first_eval >= n_time) return 0;
  return (n_time - 1 - p->first_eval) / p->eval_stride + 1;
}

int32_t oracle_counters_rate(const double* v, uint32_t n_series, uint32_t n_time, const igpu_rate_params* p, float* out32,
                             double* out64, int32_t threads) {
  if (!v || !p || !p->step_ms || !p->range_samples || !p->eval_stride) return IGPU_ERR_ARG;
  if (threads < 1) threads = 1;
  const uint32_t n_out = oracle_counters_rate_len(n_time, p);
  const bool closed = p->options & IGPU_RATE_CLOSED_LEFT, is_rate = !(p->options & IGPU_RATE_INCREASE),
             is_counter = !(p->options & IGPU_RATE_GAUGE);
  const int64_t step = p->step_ms, range_ms = (int64_t)p->range_samples * step;
  const double range_seconds = (double)range_ms / 1000.0;
#pragma omp parallel for schedule(static) num_threads(threads)
  for (int64_t s = 0; s < (int64_t)n_series; ++s) {
    const double* row = v + (size_t)s * n_time;
    std::vector<RateSample> smp;
    for (uint32_t j = 0; j < n_out; ++j) {
      const int64_t e = (int64_t)p->first_eval + (int64_t)j * p->eval_stride;
      const int64_t end_ms = e * step, start_ms = end_ms - range_ms;
      smp.clear();
      for (int64_t i = e - (int64_t)p->range_samples; i <= e; ++i) {
        if (i < 0 || i >= (int64_t)n_time) continue;
        const int64_t t = i * step;
        if (closed ? t < start_ms : t <= start_ms) continue;  // (start, end] in v3, [start, end] in v2
        if (row[i] != row[i]) continue;
        smp.push_back({t, row[i]});
      }
      const double r = extrapolated_rate(smp.data(), smp.size(), start_ms, end_ms, range_seconds, is_counter, is_rate);
      if (out64) out64[(size_t)s * n_out + j] = r;
      if (out32) out32[(size_t)s * n_out + j] = (float)r;
    }
  }
  return IGPU_OK;
}

// Synthetic energy counters; mirrors counters_gen_kernel (csrc/rate.cu) bit for bit: integer microjoules, one rounding.
int32_t oracle_counters_generate(double* out, uint32_t n_series, uint32_t n_time, uint64_t seed, uint64_t global_first, uint32_t step_ms,
                                 uint32_t reset_period, uint32_t missing_per_million, int32_t threads) {
  if (!out || !step_ms) return IGPU_ERR_ARG;
  if (threads < 1) threads = 1;
  const uint64_t S_CPARAM = 0x21ull << 56, S_CNOISE = 0x22ull << 56, S_CMISS = 0x23ull << 56;
#pragma omp parallel for schedule(static) num_threads(threads)
  for (int64_t s = 0; s < (int64_t)n_series; ++s) {
    const uint64_t gs = global_first + (uint64_t)s;
    const uint64_t rp = splitmix64(seed ^ S_CPARAM ^ gs);
    const int64_t mw = 50000 + (int64_t)(rp % 250000u);       // milliwatts
    const int64_t inc = mw * (int64_t)step_ms;                // microjoules per grid step
    const uint32_t phase = reset_period ? (uint32_t)((rp >> 32) % reset_period) : 0;
    const int64_t boot = (int64_t)((rp >> 20) & 0xFFFFF);     // grid steps the counter ran before index 0 when it never resets
    double* row = out + (size_t)s * n_time;
    for (uint32_t t = 0; t < n_time; ++t) {
      const uint64_t idx = gs * (uint64_t)n_time + t;
      const int64_t noise = (int64_t)(splitmix64(seed ^ S_CNOISE ^ idx) % (uint64_t)(inc / 4 + 1));
      const int64_t since = reset_period ? (int64_t)((t + phase) % reset_period) : (int64_t)t + boot;
      const int64_t uj = inc * since + noise;
      const bool miss = missing_per_million && (uint32_t)(splitmix64(seed ^ S_CMISS ^ idx) % 1000000u) < missing_per_million;
      row[t] = miss ? std::numeric_limits<double>::quiet_NaN() : (double)uj * 1e-6;
    }
  }
  return IGPU_OK;
}


// ---------------------------------------------------------------------------------------------
// Fleet-wide quantile (north star: "per-fleet ... p95"): the quantile_over_time definition of
// oracle_quantile_by_node applied to every valid sample of the slab at once.  UNPINNED.
// ---------------------------------------------------------------------------------------------
int32_t oracle_quantile_fleet(const float* x, uint64_t n_samples, double q, int32_t mode, double* out_q, uint64_t* out_count) {
  if (!x || !out_q || !(q >= 0.0 && q <= 1.0)) return IGPU_ERR_ARG;
  std::vector<float> buf;
  buf.reserve(n_samples);
  for (uint64_t i = 0; i < n_samples; ++i)
    if (x[i] == x[i]) buf.push_back(x[i]);
  const size_t cnt = buf.size();
  double res = std::numeric_limits<double>::quiet_NaN();
  if (cnt) {
    if (mode == 1) {
      double r = std::ceil(q * (double)cnt) - 1.0;
      size_t k = r < 0 ? 0 : (size_t)r;
      if (k >= cnt) k = cnt - 1;
      std::nth_element(buf.begin(), buf.begin() + k, buf.end());
      res = (double)buf[k];
    } else {
      const double rank = q * (double)(cnt - 1);
      const size_t lo = (size_t)std::floor(rank);
      const size_t hi = std::min(cnt - 1, lo + 1);
      const double w = rank - (double)lo;
      std::nth_element(buf.begin(), buf.begin() + lo, buf.end());
      const double xlo = (double)buf[lo];
      double xhi = xlo;
      if (hi != lo) xhi = (double)*std::min_element(buf.begin() + lo + 1, buf.end());
      res = xlo * (1.0 - w) + xhi * w;
    }
  }
  *out_q = res;
  if (out_count) *out_count = cnt;
  return IGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Window statistics per NODE (over the node's chips) and per FLEET (over all series): min of the
// samples, max of the samples, mean = sum of the samples / number of valid samples (PromQL
// min / max / avg over the joined series), for the mocked-up range view
// (docs/screenshots/03-metrics.svg:26-37,85-89).  Planes are [n_nodes][n_win] and [n_win]; NULL = not wanted.
// ---------------------------------------------------------------------------------------------
int32_t oracle_window_rollup(const float* x, uint32_t n_series, uint32_t n_time, uint32_t n_nodes, const uint32_t* node_series_off,
                             uint32_t window, float* node_min, float* node_mean, float* node_max, uint32_t* node_count,
                             float* fleet_min, float* fleet_mean, float* fleet_max, uint64_t* fleet_count) {
  if (!x || !node_series_off || !window || node_series_off[n_nodes] > n_series) return IGPU_ERR_ARG;
  const uint32_t nw = (n_time + window - 1) / window;
  const float inf = std::numeric_limits<float>::infinity(), nanf_ = std::numeric_limits<float>::quiet_NaN();
  std::vector<double> fsum(nw, 0.0);
  std::vector<uint64_t> fcnt(nw, 0);
  std::vector<float> fmn(nw, inf), fmx(nw, -inf);
  for (uint32_t n = 0; n < n_nodes; ++n)
    for (uint32_t w = 0; w < nw; ++w) {
      const uint32_t t0 = w * window, t1 = std::min(n_time, t0 + window);
      float mn = inf, mx = -inf;
      double sum = 0;
      uint32_t cnt = 0;
      for (uint32_t s = node_series_off[n]; s < node_series_off[n + 1]; ++s)
        for (uint32_t t = t0; t < t1; ++t) {
          const float v = x[(size_t)s * n_time + t];
          if (v != v) continue;
          mn = v < mn ? v : mn, mx = v > mx ? v : mx, sum += (double)v, ++cnt;
        }
      const size_t o = (size_t)n * nw + w;
      if (node_min) node_min[o] = cnt ? mn : nanf_;
      if (node_max) node_max[o] = cnt ? mx : nanf_;
      if (node_mean) node_mean[o] = cnt ? (float)(sum / (double)cnt) : nanf_;
      if (node_count) node_count[o] = cnt;
      fsum[w] += sum, fcnt[w] += cnt;
      if (cnt) fmn[w] = std::min(fmn[w], mn), fmx[w] = std::max(fmx[w], mx);
    }
  for (uint32_t w = 0; w < nw; ++w) {
    if (fleet_min) fleet_min[w] = fcnt[w] ? fmn[w] : nanf_;
    if (fleet_max) fleet_max[w] = fcnt[w] ? fmx[w] : nanf_;
    if (fleet_mean) fleet_mean[w] = fcnt[w] ? (float)(fsum[w] / (double)fcnt[w]) : nanf_;
    if (fleet_count) fleet_count[w] = fcnt[w];
  }
  return IGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Per-container key sets (PodsPage.tsx:49-62, PodDetailSection.tsx:48-65): one row per main container
// that owns a prefixed key, pods ascending, containers ascending.  Does not rely on the entries of
// a container being adjacent.  Returns the number of rows through *n_rows; writes at most `capacity`.
// ---------------------------------------------------------------------------------------------
int32_t oracle_container_rows(const igpu_fleet_cols* c, igpu_container_row* rows, uint64_t capacity, uint64_t* n_rows) {
  if (!c || !n_rows) return IGPU_ERR_ARG;
  uint64_t n = 0;
  for (uint32_t p = 0; p < c->n_pods; ++p) {
    const uint32_t e0 = c->pod_entry_off[p], e1 = c->pod_entry_off[p + 1];
    for (uint32_t ci = 0; ci < IGPU_MAX_CONTAINERS; ++ci) {
      igpu_container_row r = {p, ci, 0, 0, 0, 0, 0, 0};
      for (uint32_t e = e0; e < e1; ++e) {
        const igpu_entry& en = c->entries[e];
        if ((en.flags & IGPU_ENTRY_INIT) || en.container != ci) continue;
        const uint32_t bit = 1u << en.res;
        ++r.n_entries;
        if (en.flags & IGPU_ENTRY_LIMIT) {
          r.lim_keys |= bit;
          if (en.flags & IGPU_ENTRY_TRUTHY) r.lim_truthy |= bit;
        } else {
          r.req_keys |= bit;
          if (en.flags & IGPU_ENTRY_TRUTHY) r.req_truthy |= bit;
        }
        if (en.flags & IGPU_ENTRY_SAME) r.same |= bit;
      }
      if (!r.n_entries) continue;
      if (rows && n < capacity) rows[n] = r;
      ++n;
    }
  }
  *n_rows = n;
  return IGPU_OK;
}

}  // extern "C"
