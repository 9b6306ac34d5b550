// quantile.cu — K4: per-node quantile (p95) of the power series + per-node sum / count / min / max
// and the fleet vector, in ~1.1 reads of the data.
//
// No reference counterpart (see series.cu header); semantics = PromQL quantile_over_time, fixed in
// NaN = missing; mode 0: rank = q(n-1), v = x[lo](1-w) + x[hi]w in float64 over the
// node's sorted valid samples; mode 1: x[ceil(qn)-1].
//
// Why not a plain multi-pass radix select: each radix pass re-reads the whole segment, so a 3-pass
// select can reach at most a third of the one-read HBM roofline (40 MB node segments do not sit in
// L2 while 148 SMs stream other nodes).  Instead (per node, all exact):
//   A  sample     one CTA reads ~n/256 (1024..4096) jittered samples, and with a block-radix select in
//                 shared memory picks two splitter keys lo <= hi that bracket the wanted rank with
//                 ~4.5 sigma margin on each side;
//   B  scan       the one full pass: every sample is turned into an order-preserving uint32 key and
//                 classified against (lo, hi): counted below, equal to lo, equal to hi, or — the ~4 %
//                 that fall strictly between — appended to the node's candidate list (scratch);
//                 the same pass folds the node's min / max / sum / NaN count;
//   C  finish     one CTA per node locates the wanted ranks inside
//                 [below | ==lo | candidates | ==hi | above] and, when they are candidates, runs the
//                 block-radix select (shared-memory histogram refinement) on that short list;
//   D  fallback   nodes whose bracket missed (or whose candidate list overflowed) are solved exactly by
//                 the same block-radix select over the whole segment — rare, slow, correct.
// Shared-memory atomics are not the limit here: profiles/microbench shows a CTA-private histogram
// absorbing one atomic per sample at the full HBM stream rate.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "series.cuh"

namespace {

constexpr uint32_t ITEM = 8192;        // samples per scan work item (32 KB)
constexpr int SCAN_THREADS = 256;
constexpr int SEL_BINS = 2048;
constexpr uint32_t SAMPLE_ROUND = 4096;  // samples taken per round of the sample kernel (16 per thread)
constexpr uint32_t SAMPLE_MAX = 16384;   // most samples per node (64 KB of keys in shared memory)
constexpr uint32_t SAMPLE_MIN = 1024;
constexpr float MARGIN_SIGMA = 4.5f;
constexpr uint32_t KEY_NAN = 0xFFFFFFFFu;  // canonical key of a missing sample; no finite value maps to it

struct NodeState {
  float lo_f, hi_f;         // splitters as floats (-inf / +inf when the bracket reaches the end of the sample)
  uint32_t lo_key, hi_key;  // the same as order-preserving keys
  uint32_t ties;            // 1: a splitter value repeats in the sample -> count ties exactly (exclusive candidates)
  uint32_t n_mid;           // candidates appended (may exceed capacity -> overflow)
  uint32_t ge_lo;           // valid samples >= lo
  uint32_t gt_lo, gt_hi;    // ties mode only: > lo, > hi
  uint32_t n_nan;
  uint32_t min_key, max_key;
  uint32_t overflow;
  uint32_t pad[3];
};
static_assert(sizeof(NodeState) == 64, "NodeState is one 64-byte record");

struct QuantWs {
  uint32_t n_nodes = 0;
  bool ready = false;
  uint64_t n_items = 0, cand_total = 0;
  uint32_t* d_item_off = nullptr;   // [n_nodes + 1]
  uint64_t* d_cand_off = nullptr;   // [n_nodes + 1]
  uint32_t* d_cand = nullptr;       // candidate keys (scratch)
  // node sums: each scan CTA owns a contiguous item range, i.e. a few whole nodes plus at most one
  // partial run at each end; partial runs are combined in CTA order by the finish kernel => deterministic
  double* d_node_sum = nullptr;     // [n_nodes] sum of a node that lies inside one CTA's range
  double* d_run_sum = nullptr;      // [max_ctas][2] partial sums of a CTA's first / last node
  uint32_t* d_cta_nodes = nullptr;  // [max_ctas][2] first / last node of each CTA's range
  uint32_t max_ctas = 0;
  NodeState* d_state = nullptr;     // [n_nodes]
  uint32_t* d_fail = nullptr;       // [1 + n_nodes] count, then node ids
  double* d_q = nullptr;            // [n_nodes]
  igpu_node_series_stats* d_stats = nullptr;
};

// ---- block-wide helpers -------------------------------------------------------------------------

struct SelSmem {
  uint32_t hist[SEL_BINS];
  uint32_t warp_tot[32];
  uint32_t res[2];
};

// smallest shift that maps a key range of `width` keys onto SEL_BINS bins
__device__ __forceinline__ int sel_shift(uint64_t width) {
  const uint64_t wm1 = width - 1;
  return wm1 >= (uint64_t)SEL_BINS ? (64 - __clzll((long long)wm1)) - 11 : 0;
}
static_assert(SEL_BINS == 2048, "sel_shift keeps 11 bits");

// Finds the bin where the running count crosses r; res[0] = bin, res[1] = r - count before the bin.
__device__ void block_find_bin(SelSmem& s, uint32_t r) {
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const int bpt = SEL_BINS / nt;  // nt in {256, 512}: 8 or 4 bins per thread
  uint32_t local = 0;
  for (int j = 0; j < bpt; ++j) local += s.hist[tid * bpt + j];
  uint32_t incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s.warp_tot[warp] = incl;
  __syncthreads();
  uint32_t before = 0;
  for (int w = 0; w < warp; ++w) before += s.warp_tot[w];
  const uint32_t excl = before + incl - local;
  if (r >= excl && r < excl + local) {
    uint32_t c = excl;
    for (int j = 0; j < bpt; ++j) {
      const uint32_t h = s.hist[tid * bpt + j];
      if (r < c + h) {
        s.res[0] = tid * bpt + j;
        s.res[1] = r - c;
        break;
      }
      c += h;
    }
  }
  __syncthreads();
}

// r-th smallest (0-based) of the keys that lie in [lo, hi]; keys outside the range are ignored.
// Histogram refinement: each round narrows the key range by up to 2048x. All threads must call.
template <class KeyAt>
__device__ uint32_t block_select(KeyAt key_at, uint32_t n, uint32_t lo, uint32_t hi, uint32_t r, SelSmem& s) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (;;) {
    const uint64_t width = (uint64_t)hi - lo + 1;
    const int shift = sel_shift(width);  // largest bin index (width-1)>>shift must be < SEL_BINS
    for (int i = tid; i < SEL_BINS; i += nt) s.hist[i] = 0;
    __syncthreads();
#pragma unroll 4
    for (uint32_t i = tid; i < n; i += nt) {
      const uint32_t k = key_at(i);
      if (k >= lo && k <= hi) smem_red_add_u32(&s.hist[(k - lo) >> shift], 1u);
    }
    __syncthreads();
    block_find_bin(s, r);
    const uint32_t b = s.res[0];
    r = s.res[1];
    __syncthreads();
    if (shift == 0) return lo + b;
    const uint64_t nlo = (uint64_t)lo + ((uint64_t)b << shift);
    const uint64_t cand_hi = nlo + ((1ull << shift) - 1);
    const uint64_t nhi = cand_hi < (uint64_t)hi ? cand_hi : (uint64_t)hi;
    lo = (uint32_t)nlo;
    hi = (uint32_t)nhi;
  }
}

template <class KeyAt>
__device__ void block_count_le_min_gt(KeyAt key_at, uint32_t n, uint32_t lo, uint32_t hi, uint32_t K, SelSmem& s, uint32_t& c_le,
                                      uint32_t& min_gt);

// Calls f(key) for each of the n keys at `keys` (global memory, 16-byte aligned): 128-bit loads, four of
// them in flight per thread before any key is used.
template <int DEPTH = 8, class F>  // DEPTH: 128-bit loads in flight per thread
__device__ __forceinline__ void for_each_key_vec(const uint32_t* __restrict__ keys, uint32_t n, F f) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint4* k4 = reinterpret_cast<const uint4*>(keys);
  const uint32_t n4 = n >> 2;
  for (uint32_t i0 = 0; i0 < n4; i0 += DEPTH * nt) {
    uint4 v[DEPTH];
    bool ok[DEPTH];
#pragma unroll
    for (int j = 0; j < DEPTH; ++j) {
      const uint32_t i = i0 + j * nt + tid;
      ok[j] = i < n4;
      v[j] = ok[j] ? __ldg(k4 + i) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int j = 0; j < DEPTH; ++j)
      if (ok[j]) f(v[j].x), f(v[j].y), f(v[j].z), f(v[j].w);
  }
  for (uint32_t i = (n4 << 2) + tid; i < n; i += nt) f(__ldg(keys + i));
}

// Two ranks r and r+1 (or just r when `both` is false) among the keys in [lo, hi], in two passes over
// the keys: (1) a 2048-bin histogram over the range locates the bin of rank r; (2) the keys of that
// bin are collected into shared memory while the smallest key above the bin is tracked; the answer
// is then found inside the (small) collected list.  Falls back to block_select when the bin holds
// more keys than the list (heavy ties inside the bracket).  All threads must call.
constexpr uint32_t PAIR_LIST = 1024;
constexpr int PAIR_DEPTH = 8;  // 4 with 6 CTAs per SM measured the same (0.367 vs 0.365 ms): not occupancy-bound
struct PairSmem {
  uint32_t list[PAIR_LIST];
  uint32_t n, min_above;
  uint32_t out[2];
};

__device__ void block_select_pair(const uint32_t* __restrict__ keys, uint32_t n, uint32_t lo, uint32_t hi, uint32_t r, bool both,
                                  SelSmem& s, PairSmem& ps, uint32_t& k_r, uint32_t& k_r1) {
  const int tid = threadIdx.x, nt = blockDim.x;
  auto key_at = [&](uint32_t i) { return __ldg(keys + i); };
  const uint64_t width = (uint64_t)hi - lo + 1;
  const int shift = sel_shift(width);  // largest bin index (width-1)>>shift must be < SEL_BINS
  for (int i = tid; i < SEL_BINS; i += nt) s.hist[i] = 0;
  if (tid == 0) ps.n = 0, ps.min_above = KEY_NAN;
  __syncthreads();
  // every key handed to this function lies in [lo, hi] (the candidates of a bracket are exactly the keys inside it), so
  // the histogram needs no range test; the clamp only keeps a key that broke that promise from writing outside the bins
  for_each_key_vec<PAIR_DEPTH>(keys, n, [&](uint32_t k) { smem_red_add_u32(&s.hist[min((k - lo) >> shift, (uint32_t)SEL_BINS - 1u)], 1u); });
  __syncthreads();
  block_find_bin(s, r);
  const uint32_t b = s.res[0], r_in = s.res[1], in_bin = s.hist[b];
  __syncthreads();
  const uint64_t blo64 = (uint64_t)lo + ((uint64_t)b << shift), bhi64 = blo64 + ((1ull << shift) - 1);
  const uint32_t blo = (uint32_t)blo64, bhi = (uint32_t)(bhi64 < hi ? bhi64 : hi);
  if (in_bin > PAIR_LIST) {  // uniform: too many equal-ish keys, refine the classic way
    k_r = block_select(key_at, n, blo, bhi, r_in, s);
    k_r1 = k_r;
    if (both) {
      uint32_t c_le, min_gt;
      block_count_le_min_gt(key_at, n, lo, hi, k_r, s, c_le, min_gt);
      k_r1 = (c_le >= r + 2) ? k_r : min_gt;
    }
    return;
  }
  uint32_t my_min = KEY_NAN;
  const uint32_t bspan = bhi - blo;
  for_each_key_vec<PAIR_DEPTH>(keys, n, [&](uint32_t k) {
    if (k - blo <= bspan) ps.list[atomicAdd(&ps.n, 1u)] = k;  // in the bin (unsigned: also false below it)
    my_min = min(my_min, k > bhi ? k : KEY_NAN);                // smallest key above the bin (none exceeds hi)
  });
  my_min = __reduce_min_sync(0xffffffffu, my_min);
  if ((tid & 31) == 0 && my_min != KEY_NAN) atomicMin(&ps.min_above, my_min);
  __syncthreads();
  // rank inside the list by counting: element e has rank (#smaller) + (#equal with lower index)
  const uint32_t m = ps.n;
  for (uint32_t e = tid; e < m; e += nt) {
    const uint32_t ke = ps.list[e];
    uint32_t rank = 0;
    for (uint32_t j = 0; j < m; ++j) {
      const uint32_t kj = ps.list[j];
      rank += (kj < ke) | ((kj == ke) & (j < e));
    }
    if (rank == r_in) ps.out[0] = ke;
    if (rank == r_in + 1) ps.out[1] = ke;
  }
  __syncthreads();
  k_r = ps.out[0];
  k_r1 = both ? (r_in + 1 < m ? ps.out[1] : ps.min_above) : k_r;
  __syncthreads();
}

// Among keys in [lo, hi]: how many are <= K, and the smallest one > K (KEY_NAN if none).
template <class KeyAt>
__device__ void block_count_le_min_gt(KeyAt key_at, uint32_t n, uint32_t lo, uint32_t hi, uint32_t K, SelSmem& s, uint32_t& c_le,
                                      uint32_t& min_gt) {
  const int tid = threadIdx.x, nt = blockDim.x;
  if (tid == 0) s.res[0] = 0, s.res[1] = KEY_NAN;
  __syncthreads();
  uint32_t c = 0, m = KEY_NAN;
  for (uint32_t i = tid; i < n; i += nt) {
    const uint32_t k = key_at(i);
    if (k >= lo && k <= hi) {
      c += k <= K;
      if (k > K) m = min(m, k);
    }
  }
  c = __reduce_add_sync(0xffffffffu, c);
  m = __reduce_min_sync(0xffffffffu, m);
  if ((tid & 31) == 0) {
    if (c) atomicAdd(&s.res[0], c);
    atomicMin(&s.res[1], m);
  }
  __syncthreads();
  c_le = s.res[0];
  min_gt = s.res[1];
  __syncthreads();
}

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352dU;
  x ^= x >> 15;
  x *= 0x846ca68bU;
  x ^= x >> 16;
  return x;
}

// Ranks ra <= rb among n keys held in shared memory, in two passes: one 2048-bin histogram over
// [kmin, kmax] locates both bins, then the keys of those bins are collected and ranked by counting.
// A bin that holds more than its list (heavy ties) is refined with block_select instead.
constexpr uint32_t TWO_LIST = 512;
struct TwoSmem {
  uint32_t list[2][TWO_LIST];
  uint32_t n[2];
  uint32_t out[2];
  uint32_t dup[2];  // how many keys equal out[w] (counted in the list: every key of the bin is in it)
};

// dup_known: ts.dup[] is valid on return (false when a bin was too full for its list).
__device__ void block_two_ranks(const uint32_t* keys, uint32_t n, uint32_t kmin, uint32_t kmax, uint32_t ra, uint32_t rb,
                                SelSmem& s, TwoSmem& ts, uint32_t& ka, uint32_t& kb, bool& dup_known) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint64_t width = (uint64_t)kmax - kmin + 1;
  const int shift = sel_shift(width);  // largest bin index (width-1)>>shift must be < SEL_BINS
  for (int i = tid; i < SEL_BINS; i += nt) s.hist[i] = 0;
  if (tid == 0) ts.n[0] = ts.n[1] = 0, ts.dup[0] = ts.dup[1] = 0;
  __syncthreads();
  for (uint32_t i = tid; i < n; i += nt) smem_red_add_u32(&s.hist[(keys[i] - kmin) >> shift], 1u);
  __syncthreads();
  block_find_bin(s, ra);
  const uint32_t b0 = s.res[0], r0 = s.res[1];
  __syncthreads();
  block_find_bin(s, rb);
  const uint32_t b1 = s.res[0], r1 = s.res[1];
  const uint32_t c0 = s.hist[b0], c1 = s.hist[b1];
  __syncthreads();
  auto bin_lo = [&](uint32_t b) { return (uint32_t)((uint64_t)kmin + ((uint64_t)b << shift)); };
  auto bin_hi = [&](uint32_t b) {
    const uint64_t h = (uint64_t)kmin + ((uint64_t)b << shift) + ((1ull << shift) - 1);
    return (uint32_t)(h < kmax ? h : kmax);
  };
  const bool big0 = c0 > TWO_LIST, big1 = c1 > TWO_LIST;
  for (uint32_t i = tid; i < n; i += nt) {
    const uint32_t k = keys[i], b = (k - kmin) >> shift;
    if (b == b0 && !big0) ts.list[0][atomicAdd(&ts.n[0], 1u)] = k;
    if (b == b1 && !big1) ts.list[1][atomicAdd(&ts.n[1], 1u)] = k;
  }
  __syncthreads();
  for (int w = 0; w < 2; ++w) {
    const uint32_t m = ts.n[w], want = w ? r1 : r0;
    for (uint32_t e = tid; e < m; e += nt) {
      const uint32_t ke = ts.list[w][e];
      uint32_t rank = 0;
      for (uint32_t j = 0; j < m; ++j) {
        const uint32_t kj = ts.list[w][j];
        rank += (kj < ke) | ((kj == ke) & (j < e));
      }
      if (rank == want) ts.out[w] = ke;
    }
  }
  __syncthreads();
  ka = ts.out[0];
  kb = ts.out[1];
  dup_known = !big0 && !big1;
  if (dup_known) {
    for (int w = 0; w < 2; ++w) {
      const uint32_t m = ts.n[w], kw = w ? kb : ka;
      uint32_t c = 0;
      for (uint32_t e = tid; e < m; e += nt) c += ts.list[w][e] == kw;
      if (c) atomicAdd(&ts.dup[w], c);
    }
  }
  __syncthreads();
  auto at = [&](uint32_t i) { return keys[i]; };
  if (big0) ka = block_select(at, n, bin_lo(b0), bin_hi(b0), r0, s);  // uniform branches
  if (big1) kb = block_select(at, n, bin_lo(b1), bin_hi(b1), r1, s);
}

// ---- A: sample + splitters ------------------------------------------------------------------------
// -0.0 is folded into +0.0 (v + 0.0f) wherever a value becomes a key, so that key order and float
// comparison order coincide exactly.
__device__ __forceinline__ uint32_t canon_key(float v) { return f32_key(v + 0.0f); }

__global__ void __launch_bounds__(256) quantile_sample_kernel(const float* __restrict__ x, const uint32_t* __restrict__ node_off,
                                                              uint32_t n_time, double q, NodeState* __restrict__ state,
                                                              uint32_t force_miss, uint32_t key_cap) {
  extern __shared__ uint32_t keys[];  // [SAMPLE_MAX]
  __shared__ SelSmem sel;
  __shared__ TwoSmem two;
  __shared__ uint32_t s_m, s_dup[2], s_rng[2];
  const uint32_t node = blockIdx.x;
  const int tid = threadIdx.x;
  const float* seg = x + (size_t)node_off[node] * n_time;
  const uint32_t n = (node_off[node + 1] - node_off[node]) * n_time;
  if (tid == 0) s_m = 0, s_dup[0] = 0, s_dup[1] = 0, s_rng[0] = KEY_NAN, s_rng[1] = 0;
  __syncthreads();
  // Sampling plan.  Small segments are read whole.  Otherwise the segment is cut into `lines` equal
  // buckets and one jittered 128-byte line (32 floats) is touched per bucket — DRAM moves whole lines, so
  // a lone 4-byte sample costs as much as the line (ncu: 128 B per scattered sample) — and 4 floats
  // spaced 8 apart are taken from it.  The margin below treats such a quad as 2 independent draws.
  // How many lines: sampling costs 128 B per line, the bracket it buys costs 2 x 4 B per candidate and narrows
  // like 1/sqrt(lines); the sum is least near lines = (0.043 n)^(2/3) — about 1,000 for a 0.7 M-sample node,
  // 4,096 (the cap) for the 10 M-sample nodes of the C5 shape.
  const bool whole = n <= SAMPLE_ROUND;
  const uint32_t best = whole ? 0u : (uint32_t)__powf(0.0433f * (float)n, 0.6667f);
  const uint32_t lines = whole ? 0u : min(key_cap / 4u, max(SAMPLE_MIN / 4u, min(best, n / 512u)));  // key_cap: shared memory given
  const uint32_t m = whole ? n : lines * 4u;
  const uint32_t bucket = whole ? 1u : n / lines;  // >= 512 floats
  uint32_t kmn = KEY_NAN, kmx = 0;  // key range of this thread's samples
  for (uint32_t round0 = 0; round0 < m; round0 += SAMPLE_ROUND) {  // block-uniform
    // all of a thread's 16 scattered loads of the round are issued before any is consumed.  (Four neighbouring threads
    // share a line and each computes its two hashes; a thread-per-line mapping that computes them once was measured
    // SLOWER — 0.31 -> 0.36 ms at C4: its loads touch 32 lines per instruction instead of 8.)
    float v[SAMPLE_ROUND / 256];
#pragma unroll
    for (int r = 0; r < (int)(SAMPLE_ROUND / 256); ++r) {
      const uint32_t j = round0 + r * 256 + tid;
      uint32_t idx = j;
      if (!whole) {
        const uint32_t line = j >> 2, sub = j & 3u;
        // hash -> [0, bucket - 31) by multiply-high (no division)
        const uint32_t start = line * bucket + __umulhi(mix32(line * 2654435761u + node), bucket - 31u);
        idx = (start & ~31u) + sub * 8u + (mix32(line + 0x9e3779b9u * node) & 7u);
        if (idx >= n) idx = n - 1;
      }
      v[r] = j < m ? __ldg(seg + idx) : __int_as_float(0x7fc00000);
    }
    // compaction of the valid samples: one shared-memory atomic per warp and round, not one per sample
#pragma unroll
    for (int r = 0; r < (int)(SAMPLE_ROUND / 256); ++r) {
      const bool ok = v[r] == v[r];
      const uint32_t mask = __ballot_sync(0xffffffffu, ok);
      uint32_t base = 0;
      if ((tid & 31) == 0 && mask) base = atomicAdd(&s_m, (uint32_t)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (ok) {
        const uint32_t k = canon_key(v[r]);
        keys[base + __popc(mask & ((1u << (tid & 31)) - 1u))] = k;
        kmn = min(kmn, k), kmx = max(kmx, k);
      }
    }
  }
  kmn = __reduce_min_sync(0xffffffffu, kmn);
  kmx = __reduce_max_sync(0xffffffffu, kmx);
  if ((tid & 31) == 0 && kmn != KEY_NAN) atomicMin(&s_rng[0], kmn), atomicMax(&s_rng[1], kmx);
  __syncthreads();
  const uint32_t mv = s_m;
  const uint32_t KEY_NEG_INF = 0x007FFFFFu, KEY_POS_INF = 0xFF800000u;
  uint32_t lo = KEY_NEG_INF, hi = KEY_POS_INF;
  if (mv > 0) {
    const double pos = q * (double)(mv - 1);
    // rank uncertainty of the sample quantile, in sample ranks: sqrt(q(1-q) m) for m independent draws;
    // line-clustered quads count as 2 draws each, which widens it by sqrt(2)
    const double sigma = sqrt(fmax(q * (1.0 - q), 0.0) * (double)mv * (whole ? 1.0 : 2.0));
    const int d = (int)ceil((double)MARGIN_SIGMA * sigma) + 2;
    const long long lo_i = (long long)floor(pos) - d, hi_i = (long long)ceil(pos) + d;
    // both splitter ranks in two passes over the keys (their range was folded while they were collected)
    const bool want_lo = lo_i > 0, want_hi = hi_i < (long long)mv - 1;
    bool dup_known = false;
    if (want_lo || want_hi) {
      const uint32_t ra = want_lo ? (uint32_t)lo_i : 0u, rb = want_hi ? (uint32_t)hi_i : mv - 1;
      uint32_t ka, kb;
      block_two_ranks(keys, mv, s_rng[0], s_rng[1], ra, rb, sel, two, ka, kb, dup_known);
      if (want_lo) lo = ka;
      if (want_hi) hi = kb;
      dup_known = dup_known && want_lo && want_hi;
    }
    // how often do the splitter values repeat in the sample?  Known from the select's lists when both splitters came from
    // it; otherwise (a splitter at infinity, a bin too full for its list) one pass over the keys
    if (dup_known) {
      if (tid == 0) s_dup[0] = two.dup[0], s_dup[1] = two.dup[1];
    } else {
      uint32_t c0 = 0, c1 = 0;
      for (uint32_t i = tid; i < mv; i += blockDim.x) c0 += keys[i] == lo, c1 += keys[i] == hi;
      c0 = __reduce_add_sync(0xffffffffu, c0);
      c1 = __reduce_add_sync(0xffffffffu, c1);
      if ((tid & 31) == 0) atomicAdd(&s_dup[0], c0), atomicAdd(&s_dup[1], c1);
    }
    __syncthreads();
  }
  if (tid == 0) {
    NodeState st;
    if (force_miss) lo = hi = 0x00800000u;  // test hook: a bracket far below every sample -> every node goes to the exact fallback
    st.lo_key = lo;
    st.hi_key = hi;
    st.lo_f = key_f32(lo);
    st.hi_f = key_f32(hi);
    st.ties = (lo == hi || s_dup[0] >= 3 || s_dup[1] >= 3) ? 1u : 0u;
    st.n_mid = st.ge_lo = st.gt_lo = st.gt_hi = st.n_nan = 0;
    st.min_key = KEY_NAN;
    st.max_key = 0;
    st.overflow = 0;
    st.pad[0] = st.pad[1] = st.pad[2] = 0;
    state[node] = st;
  }
}

// ---- B: the full pass -------------------------------------------------------------------------------
// Work item = ITEM consecutive samples of one node.  CTA b owns the contiguous items
// [b*ipc, (b+1)*ipc): it finds its first node once and then only advances, keeps the node's counters
// in registers for as long as the node lasts, and flushes candidates from the thread-private slots
// only when some thread's slots are half full (one __syncthreads_or per item).
struct ScanParams {
  const float* __restrict__ x;
  const uint32_t* __restrict__ node_off;
  const uint32_t* __restrict__ item_off;
  const uint64_t* __restrict__ cand_off;
  uint32_t* __restrict__ cand;
  double* __restrict__ node_sum;
  double* __restrict__ run_sum;
  uint32_t* __restrict__ cta_nodes;
  NodeState* state;
  uint32_t n_nodes, n_time;
  uint32_t n_items, items_per_cta;
};

constexpr int SPT = ITEM / SCAN_THREADS;   // samples per thread per item (32)
constexpr int SLOT_CAP = 2 * SPT;          // private candidate slots per thread (64 -> 64 KB per CTA)

struct ScanAcc {
  uint32_t ge_lo, gt_lo, gt_hi, n_nan;
  uint32_t slot_addr;  // shared-memory address of this thread's next private slot (advances by 1 KB)
  float mn, mx;
  double sum;
};

// One sample against the splitters, as a fixed predicated sequence (no branches, no selects):
//   inclusive mode (TIES = false): candidates are lo <= v <= hi; 5 instructions
//       setp.ge p1 | @p1 add ge | setp.le.and p2 | @p2 st.shared | @p2 add addr
//   ties mode: candidates are lo < v < hi; > lo and > hi are counted so both tie masses are exact; 9 instructions.
// NaN compares false everywhere, so missing samples are neither counted nor kept.
template <bool TIES, int STRIDE_BYTES = SCAN_THREADS * 4>
__device__ __forceinline__ void classify(float v, float lo, float hi, ScanAcc& a) {
  if (TIES) {
    asm volatile(
        "{\n\t"
        ".reg .pred p1, p2, p3, p4;\n\t"
        "setp.ge.f32 p1, %5, %6;\n\t"
        "@p1 add.u32 %0, %0, 1;\n\t"
        "setp.gt.f32 p2, %5, %6;\n\t"
        "@p2 add.u32 %1, %1, 1;\n\t"
        "setp.gt.f32 p3, %5, %7;\n\t"
        "@p3 add.u32 %2, %2, 1;\n\t"
        "setp.lt.and.f32 p4, %5, %7, p2;\n\t"
        "@p4 st.shared.f32 [%3], %5;\n\t"
        "@p4 add.u32 %3, %3, %4;\n\t"
        "}"
        : "+r"(a.ge_lo), "+r"(a.gt_lo), "+r"(a.gt_hi), "+r"(a.slot_addr)
        : "n"(STRIDE_BYTES), "f"(v), "f"(lo), "f"(hi)
        : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p1, p2;\n\t"
        "setp.ge.f32 p1, %3, %4;\n\t"
        "@p1 add.u32 %0, %0, 1;\n\t"
        "setp.le.and.f32 p2, %3, %5, p1;\n\t"
        "@p2 st.shared.f32 [%1], %3;\n\t"
        "@p2 add.u32 %1, %1, %2;\n\t"
        "}"
        : "+r"(a.ge_lo), "+r"(a.slot_addr)
        : "n"(STRIDE_BYTES), "f"(v), "f"(lo), "f"(hi)
        : "memory");
  }
}

template <bool TIES>
__device__ __forceinline__ void scan_item(const float* seg, uint32_t cnt, float lo, float hi, ScanAcc& a) {
  const int tid = threadIdx.x;
  if (cnt == ITEM && ((reinterpret_cast<uintptr_t>(seg) & 15u) == 0)) {
    const float4* seg4 = reinterpret_cast<const float4*>(seg);
    float4 v[SPT / 4];
#pragma unroll
    for (int r = 0; r < SPT / 4; ++r) v[r] = ldg_stream_f4(seg4 + r * SCAN_THREADS + tid);
    double sum = 0.0;
    float mx = -INFINITY;
#pragma unroll
    for (int r = 0; r < SPT / 4; ++r) {
      sum += (double)((v[r].x + v[r].y) + (v[r].z + v[r].w));  // a NaN poisons `sum`; repaired below
      a.mn = fminf(fminf(a.mn, v[r].x), fminf(v[r].y, fminf(v[r].z, v[r].w)));
      mx = fmaxf(fmaxf(mx, v[r].x), fmaxf(v[r].y, fmaxf(v[r].z, v[r].w)));
    }
    a.mx = fmaxf(a.mx, mx);
    // warp-level early-out: nothing to count or keep when no lane holds a sample >= lo (see wide_group)
    if (__any_sync(0xffffffffu, mx >= lo)) {
#pragma unroll
      for (int r = 0; r < SPT / 4; ++r) {
        classify<TIES>(v[r].x, lo, hi, a);
        classify<TIES>(v[r].y, lo, hi, a);
        classify<TIES>(v[r].z, lo, hi, a);
        classify<TIES>(v[r].w, lo, hi, a);
      }
    }
    if (sum != sum) {  // some NaN (or +inf with -inf) among this thread's 32 samples: exact recount from registers
      sum = 0.0;
#pragma unroll
      for (int r = 0; r < SPT / 4; ++r) {
        const float c[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool ok = c[j] == c[j];
          a.n_nan += !ok;
          sum += ok ? (double)c[j] : 0.0;
        }
      }
    }
    a.sum += sum;
  } else {
    for (uint32_t i = tid; i < cnt; i += SCAN_THREADS) {
      const float v = __ldg(seg + i);
      if (v == v) a.sum += (double)v; else ++a.n_nan;
      a.mn = fminf(a.mn, v);
      a.mx = fmaxf(a.mx, v);
      classify<TIES>(v, lo, hi, a);
    }
  }
}

struct SmallShared {
  uint32_t warp_cnt[SCAN_THREADS / 32];
  uint32_t base;
  uint32_t u[4];   // ge_lo, gt_lo, gt_hi, n_nan
  uint32_t mm[2];  // min key, max key
  double sum[SCAN_THREADS / 32];
};

struct ScanShared {
  float priv[SLOT_CAP * SCAN_THREADS];  // thread t owns priv[slot * 256 + t]: no atomics, no bank conflicts
  SmallShared sm;
};

// where a node's results go in global memory (shared by the scan kernel and the fused kernel)
struct NodeSink {
  NodeState* state;
  const uint64_t* __restrict__ cand_off;
  uint32_t* __restrict__ cand;
};

// Moves every thread's private candidates of `node` to the node's global candidate list.
__device__ __forceinline__ void flush_candidates(const NodeSink& p, SmallShared& sh, const float* priv, uint32_t node, ScanAcc& acc,
                                                 uint32_t slot0) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n_local = (acc.slot_addr - slot0) / (SCAN_THREADS * 4);
  uint32_t incl = n_local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) sh.warp_cnt[warp] = incl;
  __syncthreads();
  uint32_t total = 0, before = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    const uint32_t c = sh.warp_cnt[w];
    if (w < warp) before += c;
    total += c;
  }
  if (total) {  // uniform
    if (tid == 0) sh.base = atomicAdd(&p.state[node].n_mid, total);
    __syncthreads();
    const uint64_t c0 = __ldg(p.cand_off + node), cap = __ldg(p.cand_off + node + 1) - c0;
    const uint64_t my0 = (uint64_t)sh.base + before + (incl - n_local);
    if ((uint64_t)sh.base + total > cap && tid == 0) atomicExch(&p.state[node].overflow, 1u);
    for (uint32_t k = 0; k < n_local; ++k)
      if (my0 + k < cap) p.cand[c0 + my0 + k] = canon_key(priv[k * SCAN_THREADS + tid]);
  }
  __syncthreads();  // warp_cnt / base / the slots are reused
  acc.slot_addr = slot0;
}

// Adds the node-run's counters to the node's state and stores the run's partial sum.
__device__ __forceinline__ void flush_counters(const NodeSink& p, SmallShared& sh, uint32_t node, ScanAcc& acc, bool ties,
                                               double* sum_dst) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    sh.u[0] = sh.u[1] = sh.u[2] = sh.u[3] = 0;
    sh.mm[0] = KEY_NAN;
    sh.mm[1] = 0;
  }
  const uint32_t r0 = __reduce_add_sync(0xffffffffu, acc.ge_lo), r3 = __reduce_add_sync(0xffffffffu, acc.n_nan);
  const uint32_t r1 = ties ? __reduce_add_sync(0xffffffffu, acc.gt_lo) : 0u, r2 = ties ? __reduce_add_sync(0xffffffffu, acc.gt_hi) : 0u;
  const uint32_t kmn = __reduce_min_sync(0xffffffffu, canon_key(acc.mn)), kmx = __reduce_max_sync(0xffffffffu, canon_key(acc.mx));
  double sum = acc.sum;
#pragma unroll
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  __syncthreads();
  if (lane == 0) {
    if (r0) atomicAdd(&sh.u[0], r0);
    if (r1) atomicAdd(&sh.u[1], r1);
    if (r2) atomicAdd(&sh.u[2], r2);
    if (r3) atomicAdd(&sh.u[3], r3);
    atomicMin(&sh.mm[0], kmn);
    atomicMax(&sh.mm[1], kmx);
    sh.sum[warp] = sum;
  }
  __syncthreads();
  if (tid == 0) {
    NodeState* sw = p.state + node;
    if (sh.u[0]) atomicAdd(&sw->ge_lo, sh.u[0]);
    if (sh.u[1]) atomicAdd(&sw->gt_lo, sh.u[1]);
    if (sh.u[2]) atomicAdd(&sw->gt_hi, sh.u[2]);
    if (sh.u[3]) atomicAdd(&sw->n_nan, sh.u[3]);
    // the +inf / -inf placeholders of threads that saw no valid sample cannot beat real values
    atomicMin(&sw->min_key, sh.mm[0]);
    atomicMax(&sw->max_key, sh.mm[1]);
    double t = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) t += sh.sum[w];
    *sum_dst = t;
  }
  __syncthreads();
  acc.ge_lo = acc.gt_lo = acc.gt_hi = acc.n_nan = 0;
  acc.mn = INFINITY;
  acc.mx = -INFINITY;
  acc.sum = 0.0;
}

__global__ void __launch_bounds__(SCAN_THREADS) quantile_scan_kernel(const ScanParams p) {
  extern __shared__ __align__(16) unsigned char scan_smem_raw[];
  ScanShared& sh = *reinterpret_cast<ScanShared*>(scan_smem_raw);
  const NodeSink sink{p.state, p.cand_off, p.cand};
  const int tid = threadIdx.x;
  const uint32_t i0 = min(p.n_items, blockIdx.x * p.items_per_cta), i1 = min(p.n_items, i0 + p.items_per_cta);
  if (i0 >= i1) {
    if (tid == 0) p.cta_nodes[2 * blockIdx.x] = p.cta_nodes[2 * blockIdx.x + 1] = IGPU_NODE_NONE;
    return;
  }
  // owner node of the first item: last node with item_off[node] <= i0 (done once per CTA)
  uint32_t na = 0, nb = p.n_nodes;
  while (nb - na > 1) {
    const uint32_t mid = (na + nb) >> 1;
    if (__ldg(p.item_off + mid) <= i0) na = mid; else nb = mid;
  }
  uint32_t node = na;
  const uint32_t first_node = node;
  uint32_t node_item0 = __ldg(p.item_off + node), node_item1 = __ldg(p.item_off + node + 1);
  const uint32_t slot0 = smem_u32(sh.priv + tid);
  ScanAcc acc{0, 0, 0, 0, slot0, INFINITY, -INFINITY, 0.0};
  float lo = p.state[node].lo_f, hi = p.state[node].hi_f;
  bool ties = p.state[node].ties != 0;
  size_t seg_base = (size_t)__ldg(p.node_off + node) * p.n_time;
  uint32_t n = (__ldg(p.node_off + node + 1) - __ldg(p.node_off + node)) * p.n_time;
  for (uint32_t item = i0; item < i1; ++item) {
    if (item >= node_item1) {  // uniform: the node is finished, move to the next one that has items
      flush_candidates(sink, sh.sm, sh.priv, node, acc, slot0);
      flush_counters(sink, sh.sm, node, acc, ties, node == first_node ? p.run_sum + 2 * blockIdx.x : p.node_sum + node);
      do {
        ++node;
        node_item0 = node_item1;
        node_item1 = __ldg(p.item_off + node + 1);
      } while (item >= node_item1);
      lo = p.state[node].lo_f;
      hi = p.state[node].hi_f;
      ties = p.state[node].ties != 0;
      seg_base = (size_t)__ldg(p.node_off + node) * p.n_time;
      n = (__ldg(p.node_off + node + 1) - __ldg(p.node_off + node)) * p.n_time;
    }
    const uint32_t base = (item - node_item0) * ITEM;
    const uint32_t cnt = min(ITEM, n - base);
    const float* seg = p.x + seg_base + base;
    // the item after this one (same node, contiguous in memory): ask for it in L2 now, so that its loads
    // find it there; items that start a new node are simply not prefetched
    if (tid == 0 && cnt == ITEM && base + 2 * ITEM <= n && ((reinterpret_cast<uintptr_t>(seg) & 15u) == 0))
      bulk_prefetch_l2(seg + ITEM, ITEM * 4u);
    if (ties) scan_item<true>(seg, cnt, lo, hi, acc);
    else scan_item<false>(seg, cnt, lo, hi, acc);
    // the next item can add up to SPT candidates per thread
    if (__syncthreads_or((acc.slot_addr - slot0) > (uint32_t)(SLOT_CAP - SPT) * SCAN_THREADS * 4))
      flush_candidates(sink, sh.sm, sh.priv, node, acc, slot0);
  }
  flush_candidates(sink, sh.sm, sh.priv, node, acc, slot0);
  // the last run goes to slot 1 unless it is also the first run of this CTA (then slot 0)
  flush_counters(sink, sh.sm, node, acc, ties, p.run_sum + 2 * blockIdx.x + (node == first_node ? 0 : 1));
  if (tid == 0) {
    p.cta_nodes[2 * blockIdx.x] = first_node;
    p.cta_nodes[2 * blockIdx.x + 1] = node;
  }
}

// ---- B': fused single read — window statistics AND quantile classification ------------------------
// The K3 structure (bulk-copy ring of row tiles in shared memory, thread t folds chunk t of L samples,
// PW lanes combine a window) with the K4 scan folded into the same pass over the staged samples: each
// sample is also classified against the node's splitters (same 5-instruction predicated sequence, same
// thread-private candidate slots), the node's min / max / sum / NaN count come for free from the chunk
// results, and — as in the scan kernel — a CTA owns a contiguous range of tiles, so the node is found
// once and then only advanced.  One read of the slab instead of two.
constexpr int FUSED_MAX_STAGES = 4;

struct FusedParams {
  const float* __restrict__ x;
  float* __restrict__ omin;
  float* __restrict__ omean;
  float* __restrict__ omax;
  uint32_t n_series, n_time, n_win;
  uint32_t PW, log2PW, K, tiles_per_row, n_tiles, tiles_per_cta;
  uint32_t slot_cap;  // private candidate slots per thread (> K*L)
  uint32_t stages;    // depth of the bulk-copy ring (2..4)
  double inv_window;  // 1.0 / window: full windows multiply instead of dividing
  const uint32_t* __restrict__ node_off;
  NodeSink sink;
  double* __restrict__ node_sum;
  double* __restrict__ run_sum;
  uint32_t* __restrict__ cta_nodes;
  uint32_t n_nodes;
};

template <int L, bool TIES>
__device__ __forceinline__ void fused_chunk(const float* src, uint32_t sample0, uint32_t n_time, float lo, float hi, ScanAcc& acc,
                                            Acc& a) {
  a = Acc{INFINITY, -INFINITY, 0.0, (uint32_t)L};
  uint32_t span = L;  // samples of this chunk that exist
  if (sample0 + L <= n_time) {
    float v[L];
#pragma unroll
    for (int j = 0; j < L; ++j) v[j] = src[j];
    // min / max as balanced trees (three-input FMNMX3), sum in float64, then the splitter classification
    float mn = v[0], mx = v[0];
#pragma unroll
    for (int j = 1; j + 1 < L; j += 2) {
      mn = fminf(mn, fminf(v[j], v[j + 1]));
      mx = fmaxf(mx, fmaxf(v[j], v[j + 1]));
    }
    if ((L & 1) == 0) mn = fminf(mn, v[L - 1]), mx = fmaxf(mx, v[L - 1]);
    double sum = 0.0;
#pragma unroll
    for (int j = 0; j < L; ++j) sum += (double)v[j];
#pragma unroll
    for (int j = 0; j < L; ++j) classify<TIES>(v[j], lo, hi, acc);
    if (sum == sum) {
      a.mn = mn;  // no NaN among the L samples, so fminf never had to skip one
      a.mx = mx;
      a.sum = sum;
    } else {  // a NaN (or +inf with -inf): redo the window fold exactly; classification is NaN-safe already.  (Taking only the
              // sum and the count again, as the thread-per-window kernel does, measured slower here: 7.5 -> 8.5 ms at W = 30.)
      a = Acc{INFINITY, -INFINITY, 0.0, 0u};
      for (int j = 0; j < L; ++j) acc_sample(a, src[j]);
    }
  } else {  // the chunk is cut by the end of the row
    span = sample0 < n_time ? n_time - sample0 : 0u;
    a = Acc{INFINITY, -INFINITY, 0.0, 0u};
    for (uint32_t j = 0; j < span; ++j) {
      const float v = src[j];
      acc_sample(a, v);
      classify<TIES>(v, lo, hi, acc);
    }
  }
  acc.mn = fminf(acc.mn, a.mn);
  acc.mx = fmaxf(acc.mx, a.mx);
  acc.sum += a.sum;
  acc.n_nan += span - a.cnt;
}

template <int L>
__global__ void __launch_bounds__(WS_THREADS) series_rollup_kernel(const FusedParams p) {
  extern __shared__ __align__(128) unsigned char fused_smem[];
  const uint32_t tid = threadIdx.x;
  const uint32_t tile_chunks = WS_THREADS * p.K, tile_floats = tile_chunks * L;
  float* ring = reinterpret_cast<float*>(fused_smem);
  const uint32_t STAGES = p.stages;
  float* priv = ring + (size_t)STAGES * tile_floats;
  SmallShared& sm = *reinterpret_cast<SmallShared*>(priv + (size_t)p.slot_cap * WS_THREADS);
  uint64_t* full = reinterpret_cast<uint64_t*>(&sm + 1);
  const uint32_t t0 = min(p.n_tiles, blockIdx.x * p.tiles_per_cta), t1 = min(p.n_tiles, t0 + p.tiles_per_cta);
  if (t0 >= t1) {
    if (tid == 0) p.cta_nodes[2 * blockIdx.x] = p.cta_nodes[2 * blockIdx.x + 1] = IGPU_NODE_NONE;
    return;
  }
  if (tid == 0) {
    for (uint32_t s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  const uint32_t count = t1 - t0, tpr = p.tiles_per_row;
  TileCursor prod{t0 / tpr, t0 % tpr};
  TileCursor cons = prod;
  auto issue = [&](uint32_t i) {  // thread 0 only
    const uint32_t start = prod.tr * tile_floats;
    const uint32_t len = min(tile_floats, p.n_time - start);
    const uint32_t stage = i % STAGES;
    mbar_expect_tx(&full[stage], len * 4u);
    bulk_g2s(ring + (size_t)stage * tile_floats, p.x + (size_t)prod.row * p.n_time + start, len * 4u, &full[stage]);
    prod.advance(0, 1, tpr);
  };
  if (tid == 0)
    for (uint32_t i = 0; i < count && i < STAGES; ++i) issue(i);

  // node of the first row: last node with node_off[node] <= row (found once)
  uint32_t na = 0, nb = p.n_nodes;
  while (nb - na > 1) {
    const uint32_t mid = (na + nb) >> 1;
    if (__ldg(p.node_off + mid) <= cons.row) na = mid; else nb = mid;
  }
  uint32_t node = na;
  const uint32_t first_node = node;
  uint32_t node_row_end = __ldg(p.node_off + node + 1);
  float lo = p.sink.state[node].lo_f, hi = p.sink.state[node].hi_f;
  bool ties = p.sink.state[node].ties != 0;
  const uint32_t slot0 = smem_u32(priv + tid);
  ScanAcc acc{0, 0, 0, 0, slot0, INFINITY, -INFINITY, 0.0};
  const uint32_t flush_above = (p.slot_cap - p.K * L) * WS_THREADS * 4;  // a tile adds at most K*L candidates per thread
  const uint32_t PW = p.PW;

  uint32_t stage = 0, parity = 0;
  for (uint32_t i = 0; i < count; ++i) {
    if (cons.row >= node_row_end) {  // uniform: the node is finished, move on to the owner of this row
      flush_candidates(p.sink, sm, priv, node, acc, slot0);
      flush_counters(p.sink, sm, node, acc, ties, node == first_node ? p.run_sum + 2 * blockIdx.x : p.node_sum + node);
      do {
        ++node;
        node_row_end = __ldg(p.node_off + node + 1);
      } while (cons.row >= node_row_end);
      lo = p.sink.state[node].lo_f;
      hi = p.sink.state[node].hi_f;
      ties = p.sink.state[node].ties != 0;
    }
    mbar_wait(&full[stage], parity);
    const float* sbuf = ring + (size_t)stage * tile_floats;
    const uint32_t tile_sample0 = cons.tr * tile_floats;
    float* const row_min = p.omin + (size_t)cons.row * p.n_win;
    float* const row_mean = p.omean + (size_t)cons.row * p.n_win;
    float* const row_max = p.omax + (size_t)cons.row * p.n_win;
    for (uint32_t k = 0; k < p.K; ++k) {
      const uint32_t chunk_in_tile = k * WS_THREADS + tid;
      const uint32_t sample0 = tile_sample0 + chunk_in_tile * L;
      const float* src = sbuf + chunk_in_tile * L;
      Acc a;
      if (ties) fused_chunk<L, true>(src, sample0, p.n_time, lo, hi, acc, a);
      else fused_chunk<L, false>(src, sample0, p.n_time, lo, hi, acc, a);
#pragma unroll
      for (uint32_t o = 1; o < 32; o <<= 1) {  // the PW chunks of a window sit in adjacent lanes
        if (o < PW) {                          // uniform
          a.mn = fminf(a.mn, __shfl_xor_sync(0xffffffffu, a.mn, o));
          a.mx = fmaxf(a.mx, __shfl_xor_sync(0xffffffffu, a.mx, o));
          a.sum += __shfl_xor_sync(0xffffffffu, a.sum, o);
          a.cnt += __shfl_xor_sync(0xffffffffu, a.cnt, o);
        }
      }
      const uint32_t chunk = cons.tr * tile_chunks + chunk_in_tile;
      const uint32_t w = chunk >> p.log2PW;
      if ((chunk & (PW - 1)) == 0 && w < p.n_win) {
        const float nanv = __int_as_float(0x7fc00000);
        // a full window multiplies by 1/W (exactly rounded reciprocal); others divide
        const double mean = a.cnt == PW * L ? a.sum * p.inv_window : a.sum / (double)a.cnt;
        row_min[w] = a.cnt ? a.mn : nanv;
        row_max[w] = a.cnt ? a.mx : nanv;
        row_mean[w] = a.cnt ? (float)mean : nanv;
      }
    }
    cons.advance(0, 1, tpr);
    // one barrier per tile: releases the ring stage and decides whether the private slots must be emptied
    if (__syncthreads_or((acc.slot_addr - slot0) > flush_above)) flush_candidates(p.sink, sm, priv, node, acc, slot0);
    if (tid == 0 && i + STAGES < count) issue(i + STAGES);
    if (++stage == STAGES) stage = 0, parity ^= 1u;
  }
  flush_candidates(p.sink, sm, priv, node, acc, slot0);
  flush_counters(p.sink, sm, node, acc, ties, p.run_sum + 2 * blockIdx.x + (node == first_node ? 0 : 1));
  if (tid == 0) {
    p.cta_nodes[2 * blockIdx.x] = first_node;
    p.cta_nodes[2 * blockIdx.x + 1] = node;
  }
}

// ---- B'': the same fused pass, one thread per WHOLE window ---------------------------------------------
// For windows of W = 4*NV samples with NV odd (W = 60 -> NV = 15): thread t reads its window as NV
// 128-bit shared loads (stride W*4 bytes between threads is conflict-free for LDS.128 exactly when NV is
// odd), so there are no cross-lane shuffles, the three outputs of consecutive threads are consecutive
// (coalesced), and the fixed per-tile work (barrier, cursor, node check) is spread over W samples per
// thread instead of L.  ncu on the chunk-per-thread kernel above showed that fixed work costing as much
// as the 15-sample body; this mapping roughly halves the instructions per sample.
// Small CTAs (64 threads, a tile = 64 windows) keep the shared-memory footprint per CTA low; the private
// candidate slots are checked for fullness after every group of 16 samples.
constexpr int WIDE_THREADS = 64;
// float4s per group = samples / 4 between two fullness checks of the private candidate slots.  A divisor of NV, so that all
// groups of a window are alike and the loop over them stays ROLLED: with four unrolled groups x {ties, no ties} the kernel's
// code outgrew the instruction cache (ncu at q = 0.5, where the classification actually runs: "no instruction" was the top
// stall, 2.9 per issue).
constexpr int wide_group_of(int nv) { return nv % 5 == 0 ? 5 : nv % 3 == 0 ? 3 : nv % 4 == 0 ? 4 : nv; }

struct WideParams {
  const float* __restrict__ x;
  float* __restrict__ omin;
  float* __restrict__ omean;
  float* __restrict__ omax;
  uint16_t* __restrict__ ocnt;  // valid samples per (series, window), or NULL (igpu_series_keep_window_counts)
  uint32_t n_series, n_time, n_win, window;
  uint32_t tiles_per_row, n_tiles, tiles_per_cta;
  uint32_t slot_cap, stages, l2_prefetch;
  uint32_t k;       // sub-windows (of 4*NV samples, one per thread) that make up one window; 1 = the thread owns the window
  uint32_t active;  // threads of a tile that own a sub-window: the largest multiple of k that fits the CTA
  double inv_window;
  const uint32_t* __restrict__ node_off;
  NodeSink sink;
  double* __restrict__ node_sum;
  double* __restrict__ run_sum;
  uint32_t* __restrict__ cta_nodes;
  uint32_t n_nodes;
};

// a thread's sub-window result on its way to the thread that combines a window (k > 1)
struct WidePart {
  float mn, mx;
  double sum;
  uint32_t cnt, pad;
};

struct SmallSharedWide {  // the flush helpers are written for SCAN_THREADS; this kernel has its own small copies
  uint32_t warp_cnt[WIDE_THREADS / 32];
  uint32_t base;
  uint32_t u[4];
  uint32_t mm[2];
  uint32_t pad;
  double sum[WIDE_THREADS / 32];
};

__device__ __forceinline__ void wide_flush_candidates(const NodeSink& p, SmallSharedWide& sh, const float* priv, uint32_t node,
                                                      uint32_t& slot_addr, uint32_t slot0) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n_local = (slot_addr - slot0) / (WIDE_THREADS * 4);
  uint32_t incl = n_local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) sh.warp_cnt[warp] = incl;
  __syncthreads();
  const uint32_t c0w = sh.warp_cnt[0], total = c0w + sh.warp_cnt[1], before = warp ? c0w : 0u;
  if (total) {  // uniform
    if (tid == 0) sh.base = atomicAdd(&p.state[node].n_mid, total);
    __syncthreads();
    const uint64_t c0 = __ldg(p.cand_off + node), cap = __ldg(p.cand_off + node + 1) - c0;
    const uint64_t my0 = (uint64_t)sh.base + before + (incl - n_local);
    if ((uint64_t)sh.base + total > cap && tid == 0) atomicExch(&p.state[node].overflow, 1u);
    for (uint32_t k = 0; k < n_local; ++k)
      if (my0 + k < cap) p.cand[c0 + my0 + k] = canon_key(priv[k * WIDE_THREADS + tid]);
  }
  __syncthreads();
  slot_addr = slot0;
}

__device__ __forceinline__ void wide_flush_counters(const NodeSink& p, SmallSharedWide& sh, uint32_t node, ScanAcc& acc, bool ties,
                                                    double* sum_dst) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t r0 = __reduce_add_sync(0xffffffffu, acc.ge_lo), r3 = __reduce_add_sync(0xffffffffu, acc.n_nan);
  const uint32_t r1 = ties ? __reduce_add_sync(0xffffffffu, acc.gt_lo) : 0u, r2 = ties ? __reduce_add_sync(0xffffffffu, acc.gt_hi) : 0u;
  const uint32_t kmn = __reduce_min_sync(0xffffffffu, canon_key(acc.mn)), kmx = __reduce_max_sync(0xffffffffu, canon_key(acc.mx));
  double sum = acc.sum;
#pragma unroll
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  __syncthreads();
  if (lane == 0 && warp == 1) {
    sh.u[0] = r0, sh.u[1] = r1, sh.u[2] = r2, sh.u[3] = r3;
    sh.mm[0] = kmn, sh.mm[1] = kmx;
    sh.sum[1] = sum;
  }
  __syncthreads();
  if (tid == 0) {
    NodeState* sw = p.state + node;
    const uint32_t a0 = r0 + sh.u[0], a1 = r1 + sh.u[1], a2 = r2 + sh.u[2], a3 = r3 + sh.u[3];
    if (a0) atomicAdd(&sw->ge_lo, a0);
    if (a1) atomicAdd(&sw->gt_lo, a1);
    if (a2) atomicAdd(&sw->gt_hi, a2);
    if (a3) atomicAdd(&sw->n_nan, a3);
    atomicMin(&sw->min_key, min(kmn, sh.mm[0]));
    atomicMax(&sw->max_key, max(kmx, sh.mm[1]));
    *sum_dst = sum + sh.sum[1];
  }
  __syncthreads();
  acc.ge_lo = acc.gt_lo = acc.gt_hi = acc.n_nan = 0;
  acc.mn = INFINITY;
  acc.mx = -INFINITY;
  acc.sum = 0.0;
}

// one group of G float4s (4G samples) of the thread's (sub-)window, which starts at base4; the group's float4s are
// j0 .. j0 + G - 1, or — ROTN != 0, a window of ROTN float4s with ROTN a power of two — those positions rotated by `rot`
// (the fold does not care about order, and thread t starting at float4 t keeps LDS.128 conflict-free where the plain
// stride of an even number of float4s would put every lane of a quarter-warp on the same banks).
// `valid` = how many of the group's samples exist (always all of them when rot != 0).
template <int G, bool TIES, int ROTN>
__device__ __forceinline__ void wide_group(const float4* base4, uint32_t j0, uint32_t rot, uint32_t valid, float lo, float hi, ScanAcc& acc,
                                           Acc& a) {
  // `valid == 4*G` is NOT warp-uniform at a row end, so the warp vote below sits outside that branch.
  float4 q[G];
  float mx = -INFINITY;
  const bool whole = valid == 4 * G;
  auto at = [&](uint32_t i) -> uint32_t { return ROTN ? ((j0 + i + rot) & (uint32_t)(ROTN - 1)) : j0 + i; };
  if (whole) {
#pragma unroll
    for (int i = 0; i < G; ++i) q[i] = base4[at(i)];
    float mn = q[0].x;
    double sum = 0.0;
#pragma unroll
    for (int i = 0; i < G; ++i) {
      mn = fminf(fminf(mn, q[i].x), fminf(q[i].y, fminf(q[i].z, q[i].w)));
      mx = fmaxf(fmaxf(mx, q[i].x), fmaxf(q[i].y, fmaxf(q[i].z, q[i].w)));
      sum += (double)q[i].x;
      sum += (double)q[i].y;
      sum += (double)q[i].z;
      sum += (double)q[i].w;
    }
    if (sum == sum) {
      a.mn = fminf(a.mn, mn);
      a.mx = fmaxf(a.mx, mx);
      a.sum += sum;
      a.cnt += 4 * G;
    } else {
      // a NaN (or +inf with -inf) in the group.  min and max above are right as they are — fminf / fmaxf return the operand
      // that is not NaN, a group without any sample leaves NaN / -inf, which the folds below ignore — so only the sum and
      // the count are taken again, one 128-bit read and four predicated additions per float4.  (Slabs with 0.1 % of the
      // samples missing hit this in half of all warp-groups: the sample-by-sample fold it replaces cost 7.3 ms at C4.)
      a.mn = fminf(a.mn, mn);
      a.mx = fmaxf(a.mx, mx);
#pragma unroll 1  // (unrolled from the registers: clean data 4.52 -> 4.69 ms, W = 300 4.84 -> 5.63: code size)
      for (int i = 0; i < G; ++i) {
        const float4 v = base4[at(i)];
        a.sum += v.x == v.x ? (double)v.x : 0.0;
        a.sum += v.y == v.y ? (double)v.y : 0.0;
        a.sum += v.z == v.z ? (double)v.z : 0.0;
        a.sum += v.w == v.w ? (double)v.w : 0.0;
        a.cnt += (uint32_t)(v.x == v.x) + (uint32_t)(v.y == v.y) + (uint32_t)(v.z == v.z) + (uint32_t)(v.w == v.w);
      }
    }
  } else {  // the window is cut by the end of the row (rot == 0 there): fold and classify sample by sample
    const float* s1 = reinterpret_cast<const float*>(base4 + j0);
#pragma unroll 1
    for (uint32_t j = 0; j < valid; ++j) {
      const float v = s1[j];
      acc_sample(a, v);
      classify<TIES, WIDE_THREADS * 4>(v, lo, hi, acc);
    }
  }
  // Early-out for the whole warp: when no lane's group maximum reaches the lower splitter, none of these
  // 32 x 4G samples is counted or kept, so the 5-instructions-per-sample classification is skipped.  Power
  // is temporally clustered (idle chips, off-peak hours), which makes this the common case; for data
  // without such structure it costs one compare and one vote per group.
  if (__any_sync(0xffffffffu, whole && mx >= lo)) {
    if (whole) {
#pragma unroll
      for (int i = 0; i < G; ++i) {
        classify<TIES, WIDE_THREADS * 4>(q[i].x, lo, hi, acc);
        classify<TIES, WIDE_THREADS * 4>(q[i].y, lo, hi, acc);
        classify<TIES, WIDE_THREADS * 4>(q[i].z, lo, hi, acc);
        classify<TIES, WIDE_THREADS * 4>(q[i].w, lo, hi, acc);
      }
    }
  }
}

// SUB = false: the thread owns the whole window (k == 1; tile size and thread count are compile-time constants, which is
// worth 10 % on the W = 60 headline shape); SUB = true: k threads share a window.
template <int NV, bool SUB>
__global__ void __launch_bounds__(WIDE_THREADS) series_rollup_wide_kernel(const WideParams p) {
  extern __shared__ __align__(128) unsigned char wide_smem[];
  constexpr uint32_t W = 4 * NV;  // samples per thread: the whole window, or one of its k sub-windows
  constexpr int G = wide_group_of(NV);
  constexpr uint32_t NG = NV / G;  // groups per (sub-)window
  const uint32_t tid = threadIdx.x;
  const uint32_t n_active = SUB ? p.active : (uint32_t)WIDE_THREADS, n_sub = SUB ? p.k : 1u;
  const uint32_t tile_floats = n_active * W;
  const uint32_t STAGES = p.stages;
  float* ring = reinterpret_cast<float*>(wide_smem);
  float* priv = ring + (size_t)STAGES * tile_floats;
  SmallSharedWide& sm = *reinterpret_cast<SmallSharedWide*>(priv + (size_t)p.slot_cap * WIDE_THREADS);
  uint64_t* full = reinterpret_cast<uint64_t*>(&sm + 1);
  WidePart* parts = reinterpret_cast<WidePart*>(full + FUSED_MAX_STAGES);  // [2][WIDE_THREADS], used when p.k > 1
  const uint32_t t0 = min(p.n_tiles, blockIdx.x * p.tiles_per_cta), t1 = min(p.n_tiles, t0 + p.tiles_per_cta);
  if (t0 >= t1) {
    if (tid == 0) p.cta_nodes[2 * blockIdx.x] = p.cta_nodes[2 * blockIdx.x + 1] = IGPU_NODE_NONE;
    return;
  }
  if (tid == 0) {
    for (uint32_t s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  const uint32_t count = t1 - t0, tpr = p.tiles_per_row;
  TileCursor prod{t0 / tpr, t0 % tpr};
  TileCursor cons = prod;
  uint32_t issued = 0;
  auto issue = [&](uint32_t stage) {  // thread 0 only
    const uint32_t start = prod.tr * tile_floats;
    const uint32_t len = min(tile_floats, p.n_time - start);
    mbar_expect_tx(&full[stage], len * 4u);
    bulk_g2s(ring + (size_t)stage * tile_floats, p.x + (size_t)prod.row * p.n_time + start, len * 4u, &full[stage]);
    prod.advance(0, 1, tpr);
    if (p.l2_prefetch && ++issued < count) {  // the tile after this one: into L2 now, into shared memory later
      const uint32_t nstart = prod.tr * tile_floats;
      bulk_prefetch_l2(p.x + (size_t)prod.row * p.n_time + nstart, min(tile_floats, p.n_time - nstart) * 4u);
    }
  };
  if (tid == 0)
    for (uint32_t i = 0; i < count && i < STAGES; ++i) issue(i);

  uint32_t na = 0, nb = p.n_nodes;
  while (nb - na > 1) {
    const uint32_t mid = (na + nb) >> 1;
    if (__ldg(p.node_off + mid) <= cons.row) na = mid; else nb = mid;
  }
  uint32_t node = na;
  const uint32_t first_node = node;
  uint32_t node_row_end = __ldg(p.node_off + node + 1);
  float lo = p.sink.state[node].lo_f, hi = p.sink.state[node].hi_f;
  bool ties = p.sink.state[node].ties != 0;
  const uint32_t slot0 = smem_u32(priv + tid);
  ScanAcc acc{0, 0, 0, 0, slot0, INFINITY, -INFINITY, 0.0};
  const uint32_t flush_above = (p.slot_cap - 4 * G) * WIDE_THREADS * 4;  // a group adds at most 4 G per thread

  uint32_t stage = 0, parity = 0;
  for (uint32_t i = 0; i < count; ++i) {
    if (cons.row >= node_row_end) {  // uniform
      wide_flush_candidates(p.sink, sm, priv, node, acc.slot_addr, slot0);
      wide_flush_counters(p.sink, sm, node, acc, ties, node == first_node ? p.run_sum + 2 * blockIdx.x : p.node_sum + node);
      do {
        ++node;
        node_row_end = __ldg(p.node_off + node + 1);
      } while (cons.row >= node_row_end);
      lo = p.sink.state[node].lo_f;
      hi = p.sink.state[node].hi_f;
      ties = p.sink.state[node].ties != 0;
    }
    mbar_wait(&full[stage], parity);
    const bool owner = !SUB || tid < n_active;  // (threads beyond the last whole window of the tile idle through it)
    const float4* src4 = reinterpret_cast<const float4*>(ring + (size_t)stage * tile_floats) + (owner ? tid : 0u) * NV;
    const uint32_t w = cons.tr * n_active + tid;                   // (sub-)window index within the row
    const uint32_t sample0 = w * W;
    const uint32_t wlen = owner && sample0 < p.n_time ? min(W, p.n_time - sample0) : 0u;  // samples of it that exist
    Acc a{INFINITY, -INFINITY, 0.0, 0u};
    // an even number of float4s per thread: every thread starts its fold at its own float4 (a cut window reads in order)
    constexpr int ROTN = (NV % 2 == 0) ? NV : 0;
    static_assert(ROTN == 0 || (NV & (NV - 1)) == 0, "rotation needs a power-of-two window");
    const uint32_t rot = (ROTN && wlen == W) ? tid : 0u;
#pragma unroll 1
    for (uint32_t g = 0; g < NG; ++g) {
      constexpr uint32_t GS = 4 * G;
      const uint32_t gvalid = wlen > g * GS ? min(wlen - g * GS, GS) : 0u;
      if (ties) wide_group<G, true, ROTN>(src4, g * G, rot, gvalid, lo, hi, acc, a);
      else wide_group<G, false, ROTN>(src4, g * G, rot, gvalid, lo, hi, acc, a);
      if (SUB && g + 1 == NG) parts[(i & 1u) * WIDE_THREADS + tid] = WidePart{a.mn, a.mx, a.sum, a.cnt, 0u};  // before the barrier
      // uniform: every thread runs all NG groups; the last barrier of the tile also releases the ring stage
      if (__syncthreads_or((acc.slot_addr - slot0) > flush_above)) wide_flush_candidates(p.sink, sm, priv, node, acc.slot_addr, slot0);
    }
    if (SUB) {
      // the first thread of each window combines its k sub-windows (the buffer of tile i is written again at tile i + 2,
      // two barriers later at the least)
      if (owner && tid % n_sub == 0 && (uint64_t)(w / n_sub) * p.window < p.n_time) {
        const WidePart* pt = parts + (i & 1u) * WIDE_THREADS + tid;
        Acc t{INFINITY, -INFINITY, 0.0, 0u};
        for (uint32_t j = 0; j < n_sub; ++j) t.mn = fminf(t.mn, pt[j].mn), t.mx = fmaxf(t.mx, pt[j].mx), t.sum += pt[j].sum, t.cnt += pt[j].cnt;
        const size_t o = (size_t)cons.row * p.n_win + w / n_sub;
        const float nanv = __int_as_float(0x7fc00000);
        const double mean = t.cnt == p.window ? t.sum * p.inv_window : t.sum / (double)t.cnt;
        p.omin[o] = t.cnt ? t.mn : nanv;
        p.omax[o] = t.cnt ? t.mx : nanv;
        p.omean[o] = t.cnt ? (float)mean : nanv;
        if (p.ocnt) p.ocnt[o] = (uint16_t)t.cnt;
      }
    }
    if (wlen) {
      if (!SUB) {
        const size_t o = (size_t)cons.row * p.n_win + w;
        const float nanv = __int_as_float(0x7fc00000);
        const double mean = a.cnt == W ? a.sum * p.inv_window : a.sum / (double)a.cnt;
        p.omin[o] = a.cnt ? a.mn : nanv;
        p.omax[o] = a.cnt ? a.mx : nanv;
        p.omean[o] = a.cnt ? (float)mean : nanv;
        if (p.ocnt) p.ocnt[o] = (uint16_t)a.cnt;
      }
      acc.mn = fminf(acc.mn, a.mn);
      acc.mx = fmaxf(acc.mx, a.mx);
      acc.sum += a.sum;
      acc.n_nan += wlen - a.cnt;
    }
    cons.advance(0, 1, tpr);
    if (tid == 0 && i + STAGES < count) issue(stage);
    if (++stage == STAGES) stage = 0, parity ^= 1u;
  }
  wide_flush_candidates(p.sink, sm, priv, node, acc.slot_addr, slot0);
  wide_flush_counters(p.sink, sm, node, acc, ties, p.run_sum + 2 * blockIdx.x + (node == first_node ? 0 : 1));
  if (tid == 0) {
    p.cta_nodes[2 * blockIdx.x] = first_node;
    p.cta_nodes[2 * blockIdx.x + 1] = node;
  }
}

using WideKernel = void (*)(const WideParams);
static WideKernel wide_kernel_for(uint32_t nv, bool sub) {
  switch (nv) {
    case 1: return sub ? nullptr : series_rollup_wide_kernel<1, false>;
    case 3: return sub ? series_rollup_wide_kernel<3, true> : series_rollup_wide_kernel<3, false>;
    case 5: return sub ? series_rollup_wide_kernel<5, true> : series_rollup_wide_kernel<5, false>;
    case 7: return sub ? series_rollup_wide_kernel<7, true> : series_rollup_wide_kernel<7, false>;
    case 9: return sub ? series_rollup_wide_kernel<9, true> : series_rollup_wide_kernel<9, false>;
    case 11: return sub ? series_rollup_wide_kernel<11, true> : series_rollup_wide_kernel<11, false>;
    case 13: return sub ? series_rollup_wide_kernel<13, true> : series_rollup_wide_kernel<13, false>;
    case 15: return sub ? series_rollup_wide_kernel<15, true> : series_rollup_wide_kernel<15, false>;
    case 16: return sub ? series_rollup_wide_kernel<16, true> : series_rollup_wide_kernel<16, false>;  // 64, 256, 1024: rotated reads
    default: return nullptr;
  }
}

using FusedKernel = void (*)(const FusedParams);
static FusedKernel fused_kernel_for(uint32_t L) {
  switch (L) {
    case 1: return series_rollup_kernel<1>;
    case 3: return series_rollup_kernel<3>;
    case 5: return series_rollup_kernel<5>;
    case 7: return series_rollup_kernel<7>;
    case 9: return series_rollup_kernel<9>;
    case 11: return series_rollup_kernel<11>;
    case 13: return series_rollup_kernel<13>;
    case 15: return series_rollup_kernel<15>;
    default: return nullptr;
  }
}

// ---- C / D: resolve ranks -----------------------------------------------------------------------------
struct FinishParams {
  const float* __restrict__ x;
  const uint32_t* __restrict__ node_off;
  const uint32_t* __restrict__ item_off;
  const uint64_t* __restrict__ cand_off;
  const uint32_t* __restrict__ cand;
  const double* __restrict__ node_sum;
  const double* __restrict__ run_sum;
  const uint32_t* __restrict__ cta_nodes;
  uint32_t items_per_cta, n_ctas;
  uint32_t unit_tiles_per_row;  // 0: CTA ranges are in scan work items; else in fused-kernel tiles (rows * tiles_per_row)
  NodeState* state;
  uint32_t* fail;  // [0] count, [1..] node ids
  double* out_q;
  igpu_node_series_stats* out_stats;
  uint32_t n_nodes, n_time;
  double q;
  int mode;
};

__device__ __forceinline__ void wanted_ranks(double q, int mode, uint64_t nv, uint64_t& r_lo, uint64_t& r_hi, double& w) {
  if (mode == 1) {
    double r = ceil(q * (double)nv) - 1.0;
    uint64_t k = r < 0 ? 0 : (uint64_t)r;
    if (k >= nv) k = nv - 1;
    r_lo = r_hi = k;
    w = 0.0;
  } else {
    const double rank = q * (double)(nv - 1);
    r_lo = (uint64_t)floor(rank);
    r_hi = (r_lo + 1 < nv) ? r_lo + 1 : nv - 1;
    w = rank - (double)r_lo;
  }
}

__device__ __forceinline__ double interpolate(uint32_t k_lo, uint32_t k_hi, double w) {
  const double xlo = (double)key_f32(k_lo), xhi = (double)key_f32(k_hi);
  return __dadd_rn(__dmul_rn(xlo, 1.0 - w), __dmul_rn(xhi, w));
}

__global__ void __launch_bounds__(256) quantile_finish_kernel(const FinishParams p) {
  __shared__ SelSmem sel;
  __shared__ PairSmem pair;
  __shared__ double s_part[1];
  const uint32_t node = blockIdx.x;
  const int tid = threadIdx.x;
  const NodeState st = p.state[node];
  const uint32_t n = (p.node_off[node + 1] - p.node_off[node]) * p.n_time;
  const uint64_t nv = (uint64_t)n - st.n_nan;
  // deterministic node sum: the scan CTAs that touched this node, in CTA order.  A CTA holds the node as
  // its first run (slot 0), its last run (slot 1), or — when the node lies strictly inside its range —
  // wrote node_sum[node] directly.
  if (tid == 0) {
    double sum = 0;
    const uint32_t it0 = p.unit_tiles_per_row ? p.node_off[node] * p.unit_tiles_per_row : p.item_off[node];
    const uint32_t it1 = p.unit_tiles_per_row ? p.node_off[node + 1] * p.unit_tiles_per_row : p.item_off[node + 1];
    if (it1 > it0) {
      const uint32_t ca = it0 / p.items_per_cta, cb = (it1 - 1) / p.items_per_cta;
      for (uint32_t c = ca; c <= cb; ++c) {
        if (p.cta_nodes[2 * c] == node) sum += p.run_sum[2 * c];
        else if (p.cta_nodes[2 * c + 1] == node) sum += p.run_sum[2 * c + 1];
        else sum += p.node_sum[node];
      }
    }
    s_part[0] = sum;
  }
  __syncthreads();
  if (tid == 0) {
    igpu_node_series_stats o;
    o.sum = s_part[0];
    o.count = nv;
    o.min = nv ? key_f32(st.min_key) : __int_as_float(0x7fc00000);
    o.max = nv ? key_f32(st.max_key) : __int_as_float(0x7fc00000);
    p.out_stats[node] = o;
  }
  if (nv == 0) {
    if (tid == 0) p.out_q[node] = __longlong_as_double(0x7ff8000000000000LL);
    return;
  }
  uint64_t r_lo, r_hi;
  double w;
  wanted_ranks(p.q, p.mode, nv, r_lo, r_hi, w);
  const uint64_t c0 = p.cand_off[node];
  const uint32_t* cand = p.cand + c0;
  // zones in sorted order: [below | == lo | candidates | == hi | above]; inclusive mode has no tie zones
  const uint64_t below = nv - st.ge_lo;
  uint64_t eq_lo = 0, eq_hi = 0;
  uint32_t mid_lo = st.lo_key, mid_hi = st.hi_key;
  if (st.ties) {
    eq_lo = st.ge_lo - st.gt_lo;
    // gt_lo = candidates + (== hi) + (> hi)
    eq_hi = st.hi_key != st.lo_key ? (uint64_t)st.gt_lo - st.n_mid - st.gt_hi : 0;
    mid_lo = st.lo_key + 1;
    mid_hi = st.hi_key - 1;
  }
  const uint64_t z1 = below, z2 = z1 + eq_lo, z3 = z2 + st.n_mid, z4 = z3 + eq_hi;
  bool fail = st.overflow != 0 || r_lo < z1 || r_hi >= z4;
  uint32_t k_lo = 0, k_hi = 0;
  if (!fail) {
    // zone of each rank: 1 == lo, 2 candidates, 3 == hi
    const int zone_lo = r_lo < z2 ? 1 : (r_lo < z3 ? 2 : 3);
    const int zone_hi = r_hi < z2 ? 1 : (r_hi < z3 ? 2 : 3);
    if (zone_lo == 1) k_lo = st.lo_key;
    if (zone_lo == 3) k_lo = st.hi_key;
    if (zone_hi == 1) k_hi = st.lo_key;
    if (zone_hi == 3) k_hi = st.hi_key;
    if (zone_lo == 2) {
      const uint32_t r2 = (uint32_t)(r_lo - z2);
      const bool both = zone_hi == 2 && r_hi != r_lo;
      uint32_t a_lo, a_hi;
      block_select_pair(cand, st.n_mid, mid_lo, mid_hi, r2, both, sel, pair, a_lo, a_hi);
      k_lo = a_lo;
      if (zone_hi == 2) k_hi = a_hi;
    } else if (zone_hi == 2) {  // r_lo is the last "== lo" sample, r_hi the smallest candidate
      uint32_t a_lo, a_hi;
      block_select_pair(cand, st.n_mid, mid_lo, mid_hi, (uint32_t)(r_hi - z2), false, sel, pair, a_lo, a_hi);
      k_hi = a_lo;
    }
  }
  if (tid == 0) {
    if (fail) {
      const uint32_t slot = atomicAdd(p.fail, 1u);
      p.fail[1 + slot] = node;
    } else {
      p.out_q[node] = interpolate(k_lo, k_hi, w);
    }
  }
}

// D: exact select over the whole segment for the nodes listed in fail[]
__global__ void __launch_bounds__(512) quantile_fallback_kernel(const FinishParams p) {
  __shared__ SelSmem sel;
  const uint32_t n_fail = p.fail[0];
  for (uint32_t f = blockIdx.x; f < n_fail; f += gridDim.x) {
    const uint32_t node = p.fail[1 + f];
    const NodeState st = p.state[node];
    const float* seg = p.x + (size_t)p.node_off[node] * p.n_time;
    const uint32_t n = (p.node_off[node + 1] - p.node_off[node]) * p.n_time;
    const uint64_t nv = (uint64_t)n - st.n_nan;
    uint64_t r_lo, r_hi;
    double w;
    wanted_ranks(p.q, p.mode, nv, r_lo, r_hi, w);
    auto at = [&](uint32_t i) {
      const float v = __ldg(seg + i);
      return v == v ? canon_key(v) : KEY_NAN;
    };
    const uint32_t k_lo = block_select(at, n, st.min_key, st.max_key, (uint32_t)r_lo, sel);
    uint32_t k_hi = k_lo;
    if (r_hi != r_lo) {
      uint32_t c_le, min_gt;
      block_count_le_min_gt(at, n, st.min_key, st.max_key, k_lo, sel, c_le, min_gt);
      k_hi = ((uint64_t)c_le >= r_lo + 2) ? k_lo : min_gt;
    }
    if (threadIdx.x == 0) p.out_q[node] = interpolate(k_lo, k_hi, w);
    __syncthreads();
  }
}

// fleet vector: deterministic single-CTA reduction over the per-node stats
__global__ void __launch_bounds__(1024) fleet_reduce_kernel(const igpu_node_series_stats* __restrict__ st, uint32_t n_nodes,
                                                            double* __restrict__ fleet /* [2*IGPU_FLEET_LEN] */) {
  __shared__ double s_sum[1024], s_cnt[1024], s_min[1024], s_max[1024], s_n[1024];
  const int tid = threadIdx.x;
  double sum = 0, cnt = 0, mn = INFINITY, mx = -INFINITY, nn = 0;
  for (uint32_t i = tid; i < n_nodes; i += 1024) {
    const igpu_node_series_stats s = st[i];
    sum += s.sum;
    cnt += (double)s.count;
    if (s.count) {
      mn = fmin(mn, (double)s.min);
      mx = fmax(mx, (double)s.max);
      nn += 1;
    }
  }
  s_sum[tid] = sum, s_cnt[tid] = cnt, s_min[tid] = mn, s_max[tid] = mx, s_n[tid] = nn;
  __syncthreads();
  for (int o = 512; o; o >>= 1) {
    if (tid < o) {
      s_sum[tid] += s_sum[tid + o];
      s_cnt[tid] += s_cnt[tid + o];
      s_n[tid] += s_n[tid + o];
      s_min[tid] = fmin(s_min[tid], s_min[tid + o]);
      s_max[tid] = fmax(s_max[tid], s_max[tid + o]);
    }
    __syncthreads();
  }
  if (tid == 0) {
    for (int i = 0; i < 2 * IGPU_FLEET_LEN; ++i) fleet[i] = 0;
    fleet[IGPU_F_SUM] = s_sum[0];
    fleet[IGPU_F_COUNT] = s_cnt[0];
    fleet[IGPU_F_NODES] = s_n[0];
    fleet[IGPU_FLEET_LEN + 0] = s_min[0];  // reduced with min across ranks
    fleet[IGPU_FLEET_LEN + 1] = s_max[0];  // reduced with max across ranks
  }
}

int32_t ensure_ws(igpu_ctx* ctx, igpu_series* s, QuantWs** out) {
  if (s->q_ws && static_cast<QuantWs*>(s->q_ws)->ready) {
    *out = static_cast<QuantWs*>(s->q_ws);
    return IGPU_OK;
  }
  igpu_quantile_ws_free(s);
  QuantWs* ws = new QuantWs();
  s->q_ws = ws;  // from here igpu_quantile_ws_free releases whatever was allocated
  const uint32_t N = s->n_nodes;
  ws->n_nodes = N;
  std::vector<uint32_t> item_off(N + 1, 0);
  std::vector<uint64_t> cand_off(N + 1, 0);
  for (uint32_t i = 0; i < N; ++i) {
    const uint64_t n = (uint64_t)(s->h_node_off[i + 1] - s->h_node_off[i]) * s->n_time;
    const uint64_t items = (n + ITEM - 1) / ITEM;
    IGPU_REQUIRE(ctx, item_off[i] + items < 0xFFFFFFFFull, IGPU_ERR_RANGE, "series slab too large for 32-bit work-item ids");
    item_off[i + 1] = item_off[i] + (uint32_t)items;
    const uint64_t cap = std::min<uint64_t>(n, std::max<uint64_t>(2048, n / 8 + 1024));
    cand_off[i + 1] = cand_off[i] + ((cap + 3) & ~3ull);  // keeps every list 16-byte aligned for 128-bit loads
  }
  ws->n_items = item_off[N];
  ws->cand_total = cand_off[N];
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_item_off, (N + 1) * sizeof(uint32_t)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_cand_off, (N + 1) * sizeof(uint64_t)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_cand, std::max<uint64_t>(ws->cand_total, 1) * sizeof(uint32_t)));
  ws->max_ctas = (uint32_t)ctx->sm_count * 32;
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_node_sum, std::max<uint32_t>(N, 1) * sizeof(double)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_run_sum, (size_t)ws->max_ctas * 2 * sizeof(double)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_cta_nodes, (size_t)ws->max_ctas * 2 * sizeof(uint32_t)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_state, std::max<uint32_t>(N, 1) * sizeof(NodeState)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_fail, ((size_t)N + 1) * sizeof(uint32_t)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_q, std::max<uint32_t>(N, 1) * sizeof(double)));
  IGPU_CUDA(ctx, cudaMalloc(&ws->d_stats, std::max<uint32_t>(N, 1) * sizeof(igpu_node_series_stats)));
  IGPU_CUDA(ctx, cudaMemcpyAsync(ws->d_item_off, item_off.data(), (N + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
  IGPU_CUDA(ctx, cudaMemcpyAsync(ws->d_cand_off, cand_off.data(), (N + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the host vectors go out of scope
  ws->ready = true;
  *out = ws;
  return IGPU_OK;
}

}  // namespace

void igpu_quantile_ws_free(igpu_series* s) {
  if (!s || !s->q_ws) return;
  QuantWs* ws = static_cast<QuantWs*>(s->q_ws);
  cudaFree(ws->d_item_off);
  cudaFree(ws->d_cand_off);
  cudaFree(ws->d_cand);
  cudaFree(ws->d_node_sum);
  cudaFree(ws->d_run_sum);
  cudaFree(ws->d_cta_nodes);
  cudaFree(ws->d_state);
  cudaFree(ws->d_fail);
  cudaFree(ws->d_q);
  cudaFree(ws->d_stats);
  delete ws;
  s->q_ws = nullptr;
}

namespace {

// finish + fallback + fleet reduction + the exchange step + copies back.  `ipc` / `n_ctas` describe the CTA
// ranges of whichever full-pass kernel ran; tiles_per_row != 0 means the fused kernel (ranges in tiles).
int32_t finish_and_download(igpu_ctx* ctx, igpu_series* s, QuantWs* ws, double q, int32_t mode, uint32_t ipc, uint32_t n_ctas,
                            uint32_t tiles_per_row, double* host_quantile, igpu_node_series_stats* host_stats, double* host_fleet) {
  const uint32_t N = s->n_nodes;
  if (N) {
    FinishParams fp;
    fp.x = s->d_x;
    fp.node_off = s->d_node_off;
    fp.item_off = ws->d_item_off;
    fp.cand_off = ws->d_cand_off;
    fp.cand = ws->d_cand;
    fp.node_sum = ws->d_node_sum;
    fp.run_sum = ws->d_run_sum;
    fp.cta_nodes = ws->d_cta_nodes;
    fp.items_per_cta = ipc;
    fp.n_ctas = n_ctas;
    fp.unit_tiles_per_row = tiles_per_row;
    fp.state = ws->d_state;
    fp.fail = ws->d_fail;
    fp.out_q = ws->d_q;
    fp.out_stats = ws->d_stats;
    fp.n_nodes = N;
    fp.n_time = s->n_time;
    fp.q = q;
    fp.mode = mode;
    igpu_kbegin(ctx, IGPU_K_QUANTILE_FINISH);
    quantile_finish_kernel<<<N, 256, 0, ctx->stream>>>(fp);
    igpu_kend(ctx, IGPU_K_QUANTILE_FINISH);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_kbegin(ctx, IGPU_K_QUANTILE_TAIL);
    quantile_fallback_kernel<<<std::min<unsigned>(N, (unsigned)ctx->sm_count * 2), 512, 0, ctx->stream>>>(fp);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx, 2);
  }
  if (!N) igpu_kbegin(ctx, IGPU_K_QUANTILE_TAIL);
  fleet_reduce_kernel<<<1, 1024, 0, ctx->stream>>>(ws->d_stats, N, ctx->d_fleet);
  igpu_kend(ctx, IGPU_K_QUANTILE_TAIL);
  IGPU_CUDA(ctx, cudaGetLastError());
  igpu_count_launch(ctx);
  // the exchange step: fleet totals across the node-range shards
  int32_t st = igpu_comm_group_begin(ctx);  // sum / min / max: three tiny all-reduces, one NCCL launch
  if (st == IGPU_OK) st = igpu_allreduce_f64(ctx, ctx->d_fleet, IGPU_FLEET_LEN, /*sum*/ 0);
  if (st == IGPU_OK) st = igpu_allreduce_f64(ctx, ctx->d_fleet + IGPU_FLEET_LEN, 1, /*min*/ 3);
  if (st == IGPU_OK) st = igpu_allreduce_f64(ctx, ctx->d_fleet + IGPU_FLEET_LEN + 1, 1, /*max*/ 2);
  const int32_t st_end = igpu_comm_group_end(ctx);
  if (st == IGPU_OK) st = st_end;
  if (st != IGPU_OK) return st;
  igpu_mark(ctx, 3);
  igpu_mark(ctx, 4);
  uint64_t moved = 0;
  if (host_quantile && N) {
    IGPU_CUDA(ctx, cudaMemcpyAsync(host_quantile, ws->d_q, N * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    moved += N * sizeof(double);
  }
  if (host_stats && N) {
    IGPU_CUDA(ctx, cudaMemcpyAsync(host_stats, ws->d_stats, N * sizeof(igpu_node_series_stats), cudaMemcpyDeviceToHost, ctx->stream));
    moved += N * sizeof(igpu_node_series_stats);
  }
  IGPU_CUDA(ctx, cudaMemcpyAsync(ctx->h_fleet, ctx->d_fleet, 2 * IGPU_FLEET_LEN * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  moved += 2 * IGPU_FLEET_LEN * sizeof(double);
  igpu_mark(ctx, 5);
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->timing.d2h_bytes = moved;
  if (host_fleet) {
    const double* f = ctx->h_fleet;
    for (int i = 0; i < IGPU_FLEET_LEN; ++i) host_fleet[i] = 0;
    host_fleet[IGPU_F_SUM] = f[IGPU_F_SUM];
    host_fleet[IGPU_F_COUNT] = f[IGPU_F_COUNT];
    host_fleet[IGPU_F_NODES] = f[IGPU_F_NODES];
    const bool any = f[IGPU_F_NODES] > 0;
    host_fleet[IGPU_F_MIN] = any ? f[IGPU_FLEET_LEN + 0] : NAN;
    host_fleet[IGPU_F_MAX] = any ? f[IGPU_FLEET_LEN + 1] : NAN;
  }
  return IGPU_OK;
}

int32_t launch_sample(igpu_ctx* ctx, igpu_series* s, QuantWs* ws, double q) {
  IGPU_CUDA(ctx, cudaMemsetAsync(ws->d_fail, 0, sizeof(uint32_t), ctx->stream));
  igpu_kbegin(ctx, IGPU_K_QUANTILE_SAMPLE);
  // shared memory for the keys: as much as the largest node needs (4 K keys suffice below 2.6 M samples)
  uint32_t max_node = 0;
  for (uint32_t i = 0; i < s->n_nodes; ++i) max_node = std::max(max_node, (s->h_node_off[i + 1] - s->h_node_off[i]));
  const uint64_t max_n = (uint64_t)max_node * s->n_time;
  const double best = std::pow(0.0433 * (double)max_n, 0.6667) * 1.01 + 2.0;  // the kernel's plan, rounded up
  const uint32_t key_cap = (uint32_t)std::min<double>(SAMPLE_MAX, std::max<double>(SAMPLE_ROUND, 4.0 * best));
  const size_t smem = (size_t)key_cap * sizeof(uint32_t);
  IGPU_CUDA(ctx, cudaFuncSetAttribute(quantile_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SAMPLE_MAX * sizeof(uint32_t))));
  quantile_sample_kernel<<<s->n_nodes, 256, smem, ctx->stream>>>(s->d_x, s->d_node_off, s->n_time, q, ws->d_state,
                                                                 ctx->sw.q_force_fallback ? 1u : 0u, key_cap);
  igpu_kend(ctx, IGPU_K_QUANTILE_SAMPLE);
  IGPU_CUDA(ctx, cudaGetLastError());
  igpu_count_launch(ctx);
  return IGPU_OK;
}

}  // namespace

extern "C" int32_t igpu_series_quantile_by_node(igpu_ctx* ctx, igpu_series* s, double q, int32_t mode, double* host_quantile,
                                                igpu_node_series_stats* host_stats, double* host_fleet) {
  if (!ctx || !s) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_quantile_by_node: NULL argument");
  IGPU_REQUIRE(ctx, q >= 0.0 && q <= 1.0, IGPU_ERR_ARG, "q = %g outside [0,1]", q);
  IGPU_REQUIRE(ctx, mode == 0 || mode == 1, IGPU_ERR_ARG, "mode %d is not 0 (linear) or 1 (nearest rank)", mode);
  DeviceGuard g(ctx->device);
  QuantWs* ws = nullptr;
  int32_t st = ensure_ws(ctx, s, &ws);
  if (st != IGPU_OK) return st;
  const uint32_t N = s->n_nodes;
  igpu_mark(ctx, 2);
  uint32_t n_ctas = 0, ipc = 1;
  if (N) {
    st = launch_sample(ctx, s, ws, q);
    if (st != IGPU_OK) return st;
    if (ws->n_items) {
      ScanParams sp;
      sp.x = s->d_x;
      sp.node_off = s->d_node_off;
      sp.item_off = ws->d_item_off;
      sp.cand_off = ws->d_cand_off;
      sp.cand = ws->d_cand;
      sp.node_sum = ws->d_node_sum;
      sp.run_sum = ws->d_run_sum;
      sp.cta_nodes = ws->d_cta_nodes;
      sp.state = ws->d_state;
      sp.n_nodes = N;
      sp.n_time = s->n_time;
      sp.n_items = (uint32_t)ws->n_items;
      const size_t smem = sizeof(ScanShared);
      IGPU_CUDA(ctx, cudaFuncSetAttribute(quantile_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int per_sm = 3;
      IGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, quantile_scan_kernel, SCAN_THREADS, smem));
      const uint64_t want = (uint64_t)ctx->sm_count * std::max(per_sm, 1);
      n_ctas = (uint32_t)std::min<uint64_t>(std::min<uint64_t>(ws->n_items, want), ws->max_ctas);
      ipc = (uint32_t)((ws->n_items + n_ctas - 1) / n_ctas);
      n_ctas = (uint32_t)((ws->n_items + ipc - 1) / ipc);
      sp.items_per_cta = ipc;
      igpu_kbegin(ctx, IGPU_K_QUANTILE_SCAN);
      quantile_scan_kernel<<<n_ctas, SCAN_THREADS, smem, ctx->stream>>>(sp);
      igpu_kend(ctx, IGPU_K_QUANTILE_SCAN);
      IGPU_CUDA(ctx, cudaGetLastError());
      igpu_count_launch(ctx);
    }
  }
  return finish_and_download(ctx, s, ws, q, mode, ipc, n_ctas, 0, host_quantile, host_stats, host_fleet);
}

// One read of the slab for both results (see series_rollup_kernel).  Shapes the fused kernel does not
// cover (window not of the form PW*L with L odd <= 15, rows not 16-byte aligned, no nodes) run the two
// separate passes instead; the results are the same either way.
extern "C" int32_t igpu_series_rollup(igpu_ctx* ctx, igpu_series* s, uint32_t window, double q, int32_t mode, float* host_min,
                                      float* host_mean, float* host_max, double* host_quantile, igpu_node_series_stats* host_stats,
                                      double* host_fleet) {
  if (!ctx || !s) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_rollup: NULL argument");
  IGPU_REQUIRE(ctx, window > 0, IGPU_ERR_ARG, "window must be positive");
  IGPU_REQUIRE(ctx, q >= 0.0 && q <= 1.0, IGPU_ERR_ARG, "q = %g outside [0,1]", q);
  IGPU_REQUIRE(ctx, mode == 0 || mode == 1, IGPU_ERR_ARG, "mode %d is not 0 (linear) or 1 (nearest rank)", mode);
  const WsPlan plan = igpu_ws_plan(window, s->n_time, s->n_series);
  // The thread-per-window kernel takes window = 4 * NV * k: every thread folds a sub-window of NV float4s (NV odd <= 15,
  // conflict-free LDS.128; or NV = 16 with every thread starting at its own float4), k neighbouring threads make up a
  // window (k = 1: 20, 60, 64; k = 5: 300; k = 24: 1440; k = 4 with NV = 16: 256; ...).
  uint32_t wide_nv = 0, wide_k = 0;
  if (window % 4 == 0 && s->n_time % 4 == 0 && !ctx->sw.no_wide) {
    uint32_t best = 0;
    for (int nvi = 16; nvi >= 1; nvi -= (nvi == 16 ? 1 : 2)) {  // 16 (power-of-two windows, rotated reads), then the odd ones
      const uint32_t nv = (uint32_t)nvi;
      if ((window / 4) % nv) continue;
      const uint32_t k = window / 4 / nv;
      if (k > WIDE_THREADS || (nv < 3 && k > 1)) continue;  // (4-sample sub-windows would be all overhead)
      const uint32_t active = WIDE_THREADS / k * k;
      const uint32_t score = std::min(active, 60u) * 32 + nv;  // enough busy threads first, then the longer sub-window
      if (score > best) best = score, wide_nv = nv, wide_k = k;
    }
  }
  const bool nodes_ok = s->n_nodes > 0 && s->n_series > 0 && s->h_node_off[s->n_nodes] == s->n_series;
  const bool fused = nodes_ok && !ctx->sw.no_fused && (wide_nv || (plan.fast && !plan.vec4 && plan.L <= 15));
  if (!fused) {
    int32_t st = igpu_series_window_stats(ctx, s, window, host_min, host_mean, host_max);
    if (st != IGPU_OK) return st;
    return igpu_series_quantile_by_node(ctx, s, q, mode, host_quantile, host_stats, host_fleet);
  }
  DeviceGuard g(ctx->device);
  QuantWs* ws = nullptr;
  int32_t st = ensure_ws(ctx, s, &ws);
  if (st != IGPU_OK) return st;
  st = igpu_series_ensure_window_buffers(ctx, s, window);
  if (st != IGPU_OK) return st;
  igpu_mark(ctx, 2);
  st = launch_sample(ctx, s, ws, q);
  if (st != IGPU_OK) return st;
  if (wide_nv) {
    const uint32_t NV = wide_nv;
    WideParams wp;
    wp.k = wide_k;
    wp.active = WIDE_THREADS / wide_k * wide_k;
    wp.x = s->d_x;
    wp.omin = s->d_wmin;
    wp.omean = s->d_wmean;
    wp.omax = s->d_wmax;
    wp.ocnt = (s->keep_counts && window <= 0xFFFFu) ? s->d_wcnt : nullptr;
    wp.n_series = s->n_series;
    wp.n_time = s->n_time;
    wp.n_win = s->n_win;  // windows of the caller's row length (the plan counts those of the padded pitch)
    wp.window = window;
    const uint32_t tile_samples = wp.active * 4 * NV;
    wp.tiles_per_row = (s->n_time + tile_samples - 1) / tile_samples;
    const uint64_t n_tiles = (uint64_t)wp.tiles_per_row * s->n_series;
    wp.inv_window = 1.0 / (double)window;
    wp.node_off = s->d_node_off;
    wp.sink = NodeSink{ws->d_state, ws->d_cand_off, ws->d_cand};
    wp.node_sum = ws->d_node_sum;
    wp.run_sum = ws->d_run_sum;
    wp.cta_nodes = ws->d_cta_nodes;
    wp.n_nodes = s->n_nodes;
    const int e_ctas = ctx->sw.wide_ctas, e_stages = ctx->sw.wide_stages, e_slots = ctx->sw.wide_slots;
    wp.l2_prefetch = ctx->sw.wide_no_l2pf ? 0u : 1u;
    wp.stages = e_stages ? (uint32_t)e_stages : 1u;  // measured: occupancy (10 CTAs/SM) beats ring depth here
    const uint32_t group_samples = 4u * (uint32_t)wide_group_of((int)NV);
    wp.slot_cap = e_slots ? (uint32_t)e_slots : group_samples + 8;
    if (wp.stages < 1 || wp.stages > FUSED_MAX_STAGES) wp.stages = 1;
    if (wp.slot_cap < group_samples + 1) wp.slot_cap = group_samples + 8;
    const size_t tile_bytes = (size_t)tile_samples * sizeof(float);
    const size_t smem = wp.stages * tile_bytes + (size_t)wp.slot_cap * WIDE_THREADS * 4 + sizeof(SmallSharedWide) +
                        FUSED_MAX_STAGES * sizeof(uint64_t) + (wide_k > 1 ? 2 * WIDE_THREADS * sizeof(WidePart) : 0) + 128;
    WideKernel kern = wide_kernel_for(NV, wide_k > 1);
    if (n_tiles < 0xFFFFFFFFull && kern && smem <= 200 * 1024) {
      wp.n_tiles = (uint32_t)n_tiles;
      IGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int per_sm = 1;
      IGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WIDE_THREADS, smem));
      // measured at C4 (W = 60): 8 CTAs per SM 4.50 ms, 9: 4.58, 10: 4.60 — past eight the extra warps only add L2 pressure
      per_sm = std::min(per_sm, e_ctas ? e_ctas : 8);
      per_sm = std::max(per_sm, 1);
      uint32_t n_ctas = (uint32_t)std::min<uint64_t>(std::min<uint64_t>(n_tiles, (uint64_t)ctx->sm_count * per_sm), ws->max_ctas);
      const uint32_t tpc = (uint32_t)((n_tiles + n_ctas - 1) / n_ctas);
      n_ctas = (uint32_t)((n_tiles + tpc - 1) / tpc);
      wp.tiles_per_cta = tpc;
      igpu_kbegin(ctx, IGPU_K_SERIES_FUSED);
      kern<<<n_ctas, WIDE_THREADS, smem, ctx->stream>>>(wp);
      igpu_kend(ctx, IGPU_K_SERIES_FUSED);
      IGPU_CUDA(ctx, cudaGetLastError());
      igpu_count_launch(ctx);
      s->wcnt_valid = wp.ocnt != nullptr;
      st = finish_and_download(ctx, s, ws, q, mode, tpc, n_ctas, wp.tiles_per_row, host_quantile, host_stats, host_fleet);
      if (st != IGPU_OK) return st;
      if (host_min || host_mean || host_max) return igpu_series_read_window_stats(ctx, s, 0, s->n_series, host_min, host_mean, host_max);
      return IGPU_OK;
    }
  }
  if (!(plan.fast && !plan.vec4 && plan.L <= 15)) {  // the thread-per-window launch did not fit and there is no other fused shape
    st = igpu_series_window_stats(ctx, s, window, host_min, host_mean, host_max);
    if (st != IGPU_OK) return st;
    return igpu_series_quantile_by_node(ctx, s, q, mode, host_quantile, host_stats, host_fleet);
  }
  FusedParams fp;
  fp.x = s->d_x;
  fp.omin = s->d_wmin;
  fp.omean = s->d_wmean;
  fp.omax = s->d_wmax;
  fp.n_series = s->n_series;
  fp.n_time = s->n_time;
  fp.n_win = s->n_win;
  fp.PW = plan.PW;
  fp.log2PW = plan.log2PW;
  fp.K = plan.K;
  fp.tiles_per_row = plan.tiles_per_row;
  fp.n_tiles = (uint32_t)plan.n_tiles;
  fp.node_off = s->d_node_off;
  fp.sink = NodeSink{ws->d_state, ws->d_cand_off, ws->d_cand};
  fp.node_sum = ws->d_node_sum;
  fp.run_sum = ws->d_run_sum;
  fp.cta_nodes = ws->d_cta_nodes;
  fp.n_nodes = s->n_nodes;
  // shared memory = ring (stages x tile) + private slots + bookkeeping.  Occupancy matters more than ring
  // depth here (the kernel is issue-bound, not latency-bound): try 4 CTAs/SM with a 2-deep ring, then 3 with 3.
  const uint32_t KL = plan.K * plan.L;
  const size_t tile_bytes = (size_t)WS_THREADS * KL * sizeof(float);
  const size_t fixed = sizeof(SmallShared) + FUSED_MAX_STAGES * sizeof(uint64_t) + 128;
  FusedKernel kern = fused_kernel_for(plan.L);
  size_t smem = 0;
  unsigned per_sm = 0;
  const int e_ctas = ctx->sw.fused_ctas, e_stages = ctx->sw.fused_stages;
  const unsigned try_ctas[4] = {e_ctas ? (unsigned)e_ctas : 4u, 3u, 2u, 1u};
  for (unsigned t = 0; t < 4 && !per_sm; ++t) {
    const unsigned want = try_ctas[t];
    if (want < 1 || want > 8) continue;
    const unsigned stages = e_stages ? (unsigned)e_stages : (want >= 4 ? 2u : 3u);
    if (stages < 2 || stages > FUSED_MAX_STAGES) continue;
    const size_t budget = (size_t)227 * 1024 / want - 1024;
    const size_t ring = stages * tile_bytes;
    const size_t min_slots = KL + 8;
    if (budget < ring + fixed + min_slots * WS_THREADS * 4) continue;
    uint32_t slots = (uint32_t)((budget - ring - fixed) / (WS_THREADS * 4));
    slots = std::min<uint32_t>(slots, 4 * KL + 32);
    fp.slot_cap = slots;
    fp.stages = stages;
    smem = ring + (size_t)slots * WS_THREADS * 4 + fixed;
    per_sm = want;
  }
  fp.inv_window = 1.0 / (double)window;
  IGPU_REQUIRE(ctx, per_sm > 0 && kern, IGPU_ERR_RANGE, "window shape does not fit the fused kernel's shared memory");
  IGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  uint32_t n_ctas = (uint32_t)std::min<uint64_t>(std::min<uint64_t>(plan.n_tiles, (uint64_t)ctx->sm_count * per_sm), ws->max_ctas);
  const uint32_t tpc = (uint32_t)((plan.n_tiles + n_ctas - 1) / n_ctas);
  n_ctas = (uint32_t)((plan.n_tiles + tpc - 1) / tpc);
  fp.tiles_per_cta = tpc;
  igpu_kbegin(ctx, IGPU_K_SERIES_FUSED);
  kern<<<n_ctas, WS_THREADS, smem, ctx->stream>>>(fp);
  igpu_kend(ctx, IGPU_K_SERIES_FUSED);
  IGPU_CUDA(ctx, cudaGetLastError());
  igpu_count_launch(ctx);
  st = finish_and_download(ctx, s, ws, q, mode, tpc, n_ctas, plan.tiles_per_row, host_quantile, host_stats, host_fleet);
  if (st != IGPU_OK) return st;
  if (host_min || host_mean || host_max) return igpu_series_read_window_stats(ctx, s, 0, s->n_series, host_min, host_mean, host_max);
  return IGPU_OK;
}
