// alloc_rollup.cu — K1: per-pod request/limit sums, per-node accumulation, classification and the
// fleet totals, for a whole fleet in two launches (pods, then nodes).
//
// Reference arithmetic replaced (all per-object, single-threaded TypeScript):
//   getPodGpuRequests                 src/api/k8s.ts:289-301
//   isGpuRequestingPod                src/api/k8s.ts:250-264
//   isIntelGpuNode / getNodeGpuType / getNodeGpuCount / isNodeReady   k8s.ts:125-152,171-192,329-331
//   Overview loops                    src/components/OverviewPage.tsx:71-130
//   Nodes page loops + bar            src/components/NodesPage.tsx:35-38,152-189,270-273
//   Node detail loops                 src/components/NodeDetailSection.tsx:61-91
//
// Data layout in HBM: one input arena (node columns, pod columns, the 16-byte entry stream), one
// accumulator arena that is zeroed per rollup (per-node counters + the totals vector) and one
// output arena (per-pod planes [R][P], per-node planes [R][N], small node columns).
//
// pod kernel: persistent grid, one thread per pod per iteration.  A pod's entries are contiguous
// (pod_entry_off), so a thread streams its 16-byte entries with 128-bit loads and keeps the R
// per-resource sums for requests and limits in its own shared-memory column (bank-conflict free:
// column index == threadIdx).  Per-pod results leave as coalesced plane stores; per-node
// contributions are fire-and-forget RED.ADD on the node accumulators; fleet totals are reduced
// warp -> block (shared) -> one global RED per touched slot per CTA.
// Integer adds are associative, so every output is bit-exact regardless of scheduling.
#include <algorithm>
#include <cstring>

#include "common.cuh"

struct igpu_fleet {
  uint32_t max_nodes = 0, max_pods = 0, R = 0, options = 0;
  uint64_t max_entries = 0;
  uint32_t n_nodes = 0, n_pods = 0;
  uint64_t n_entries = 0;
  bool uploaded = false, rolled = false;
  char *d_in = nullptr, *d_acc = nullptr, *d_out = nullptr;
  size_t in_bytes = 0, acc_bytes = 0, out_bytes = 0;
  size_t in_span = 0, out_span = 0;  // bytes of the input / output arena the current fleet occupies
  // Small fleets (the plugin's real scale): columns are packed into one pinned staging buffer per direction, so a call
  // costs two or three copies instead of thirty (each small copy is ~4 us of driver time).
  char *h_in = nullptr, *h_out = nullptr;
  size_t h_in_cap = 0, h_out_cap = 0;
  bool stage_busy = false;  // an upload out of h_in may still be in flight
  // inputs
  uint8_t* node_flags;
  int32_t *cap_i915, *cap_xe, *alloc_i915, *alloc_xe, *alloc_own;
  uint32_t* pod_node;
  uint8_t* pod_phase;
  uint32_t* pod_off;
  igpu_entry* entries;
  // accumulators (zeroed per rollup); node_req/node_lim/totals double as outputs
  uint32_t* node_pod_count;
  int64_t *node_inuse, *node_req, *node_lim, *totals;
  // outputs
  uint8_t* pod_is_gpu;
  uint8_t* pod_gpu_containers;
  uint32_t *pod_req_mask, *pod_lim_mask;
  int64_t *pod_req, *pod_lim;
  uint8_t *node_out_flags, *node_type, *node_bar_level, *node_util_level;
  int64_t *node_gpu_count, *node_cap_total, *node_alloc_total, *node_bar_pct, *node_util_pct;
  // on-request scratch of igpu_fleet_container_rows (grow-only)
  uint32_t* d_cont_block = nullptr;
  size_t cont_block_cap = 0;
  igpu_container_row* d_cont_rows = nullptr;
  size_t cont_rows_cap = 0;
};

namespace {

struct PodParams {
  uint32_t P, N, R, want_lim;
  uint32_t E;  // entries in the stream: offsets beyond it are rejected, not followed
  const uint32_t* __restrict__ pod_node;
  const uint8_t* __restrict__ pod_phase;
  const uint32_t* __restrict__ pod_off;
  const int4* __restrict__ entries;
  uint8_t* __restrict__ pod_is_gpu;
  uint8_t* __restrict__ pod_gpu_containers;
  uint32_t* __restrict__ pod_req_mask;
  uint32_t* __restrict__ pod_lim_mask;
  int64_t* __restrict__ pod_req;
  int64_t* __restrict__ pod_lim;
  uint32_t* node_pod_count;
  int64_t* node_inuse;
  int64_t* node_req;
  int64_t* node_lim;
  int64_t* totals;
};

// RT > 0: the number of resources is a compile-time constant (every per-resource loop unrolls and the
// plane addresses strength-reduce); RT == 0: generic, R read from the parameters.
template <int RT>
__global__ void __launch_bounds__(256) pod_rollup_kernel(const PodParams p) {
  extern __shared__ long long sm[];
  __shared__ long long btot[IGPU_TOTALS_LEN];
  const int tid = threadIdx.x, nt = RT ? 256 : blockDim.x;
  const uint32_t R = RT ? (uint32_t)RT : p.R;
  long long* cur_req = sm + tid;                 // [r * nt]
  long long* cur_lim = sm + (size_t)R * nt + tid;
  long long* tot_req = sm + (size_t)2 * R * nt + tid;
  long long* tot_lim = sm + (size_t)3 * R * nt + tid;
  for (uint32_t r = 0; r < R; ++r) tot_req[r * nt] = 0, tot_lim[r * nt] = 0;
  for (int i = tid; i < IGPU_TOTALS_LEN; i += nt) btot[i] = 0;

  unsigned n_gpu = 0, n_sched = 0, ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0, bad = 0;
  long long allocated = 0;

  // One entry folded into the thread's shared-memory columns.
  uint32_t n_cont = 0, prev_cont = 0xFFFFFFFFu;  // GPU containers of the current pod (PodDetailSection.tsx:48-52)
  uint32_t cur_pod = 0;
  auto apply = [&](const int4 raw, uint32_t& req_mask, uint32_t& lim_mask, long long& inuse_main) {
    const uint32_t res = (uint32_t)raw.y & 0xFFFFu;
    const uint32_t fl = ((uint32_t)raw.y >> 16) & 0xFFu;
    if ((uint32_t)raw.x != cur_pod) {  // entries[i].pod must agree with pod_entry_off: counted, not followed
      ++bad;
      return;
    }
    {  // a main container counts once: its entries are adjacent (igpu_rollup.h, igpu_entry)
      const uint32_t cont = (uint32_t)raw.y >> 24;
      const bool is_main = !(fl & IGPU_ENTRY_INIT);
      n_cont += (is_main && cont != prev_cont) ? 1u : 0u;
      prev_cont = is_main ? cont : prev_cont;
    }
    const long long v = (long long)(((unsigned long long)(uint32_t)raw.w << 32) | (uint32_t)raw.z);
    if (res >= R) {  // resource id outside the planes: counted (IGPU_T_BAD_INPUT), not followed
      ++bad;
      return;
    }
    if (!(fl & IGPU_ENTRY_LIMIT)) {
      if (fl & IGPU_ENTRY_TRUTHY) {  // k8s.ts:295-296
        cur_req[res * nt] += v;
        req_mask |= 1u << res;
      }
      // NodeDetailSection.tsx:78-86 — main containers, i915 / xe
      if (!(fl & IGPU_ENTRY_INIT) && res <= IGPU_RES_XE) inuse_main += v;
    } else if (fl & IGPU_ENTRY_TRUTHY) {
      cur_lim[res * nt] += v;
      lim_mask |= 1u << res;
    }
  };

  // The chain offsets -> entries -> atomics is what bounds this kernel (ncu: issue-active 39 %, stalls on
  // the first use of loaded entries), so the loop is software-pipelined: the NEXT pod's offsets are
  // requested before the current pod is processed, and the current pod's first four entries, phase
  // and node are all requested together before any of them is used.
  const uint32_t stride = gridDim.x * nt;
  uint32_t pod = blockIdx.x * nt + tid;
  uint32_t e0 = 0, e1 = 0;
  if (pod < p.P) e0 = __ldg(p.pod_off + pod), e1 = __ldg(p.pod_off + pod + 1);
  for (; pod < p.P; pod += stride) {
    const uint32_t nxt = pod + stride;
    uint32_t ne0 = 0, ne1 = 0;
    if (nxt < p.P) ne0 = __ldg(p.pod_off + nxt), ne1 = __ldg(p.pod_off + nxt + 1);
    if (e1 < e0 || e1 > p.E) {  // offsets not ascending or past the stream: the pod is counted as bad and has no entries
      ++bad;
      e0 = e1 = 0;
    }
    const uint32_t n_ent = e1 - e0;
    int4 first[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) first[j] = (uint32_t)j < n_ent ? __ldg(p.entries + e0 + j) : make_int4(0, 0, 0, 0);
    const uint32_t ph = __ldg(p.pod_phase + pod);
    uint32_t node = __ldg(p.pod_node + pod);
    if (node != IGPU_NODE_NONE && node >= p.N) node = IGPU_NODE_NONE, ++bad;
#pragma unroll
    for (uint32_t r = 0; r < R; ++r) cur_req[r * nt] = 0, cur_lim[r * nt] = 0;
    uint32_t req_mask = 0, lim_mask = 0;
    long long inuse_main = 0;
    n_cont = 0, prev_cont = 0xFFFFFFFFu, cur_pod = pod;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if ((uint32_t)j < n_ent) apply(first[j], req_mask, lim_mask, inuse_main);
    for (uint32_t e = e0 + 4; e < e1; ++e) apply(__ldg(p.entries + e), req_mask, lim_mask, inuse_main);
    const bool is_gpu = e1 > e0;  // k8s.ts:257-263
    p.pod_is_gpu[pod] = is_gpu ? 1 : 0;
    p.pod_gpu_containers[pod] = (uint8_t)min(n_cont, 255u);
    p.pod_req_mask[pod] = req_mask;
#pragma unroll
    for (uint32_t r = 0; r < R; ++r) p.pod_req[(size_t)r * p.P + pod] = cur_req[r * nt];
    if (p.want_lim) {
      p.pod_lim_mask[pod] = lim_mask;
#pragma unroll
      for (uint32_t r = 0; r < R; ++r) p.pod_lim[(size_t)r * p.P + pod] = cur_lim[r * nt];
    }
    if (!is_gpu) {
      e0 = ne0, e1 = ne1;
      continue;
    }
    ++n_gpu;
    ph0 += ph == IGPU_PHASE_RUNNING;
    ph1 += ph == IGPU_PHASE_PENDING;
    ph2 += ph == IGPU_PHASE_SUCCEEDED;
    ph3 += ph == IGPU_PHASE_FAILED;
    ph4 += ph == IGPU_PHASE_OTHER;
    const bool running = ph == IGPU_PHASE_RUNNING;
    const bool sched = node != IGPU_NODE_NONE;
    if (sched) {
      ++n_sched;
      red_add_u32(p.node_pod_count + node, 1u);  // NodesPage.tsx:156-158 (any phase)
    }
    if (running) {
      allocated += cur_req[IGPU_RES_I915 * nt] + cur_req[IGPU_RES_XE * nt];  // OverviewPage.tsx:105-113
      if (sched && inuse_main != 0) red_add_i64(p.node_inuse + node, inuse_main);
#pragma unroll
      for (uint32_t r = 0; r < R; ++r) {
        const long long q = cur_req[r * nt];
        if (q != 0) {
          tot_req[r * nt] += q;
          if (sched) red_add_i64(p.node_req + (size_t)r * p.N + node, q);
        }
        if (p.want_lim) {
          const long long l = cur_lim[r * nt];
          if (l != 0) {
            tot_lim[r * nt] += l;
            if (sched) red_add_i64(p.node_lim + (size_t)r * p.N + node, l);
          }
        }
      }
    }
    e0 = ne0, e1 = ne1;
  }

  // ---- block reduction of the fleet totals -----------------------------------------------------
  __syncthreads();
  const int lane = tid & 31, warp = tid >> 5, nwarps = nt >> 5;
  {
    const unsigned g = __reduce_add_sync(0xffffffffu, n_gpu), s = __reduce_add_sync(0xffffffffu, n_sched);
    const unsigned a0 = __reduce_add_sync(0xffffffffu, ph0), a1 = __reduce_add_sync(0xffffffffu, ph1);
    const unsigned a2 = __reduce_add_sync(0xffffffffu, ph2), a3 = __reduce_add_sync(0xffffffffu, ph3);
    const unsigned a4 = __reduce_add_sync(0xffffffffu, ph4);
    const long long al = warp_sum_i64(allocated);
    if (lane == 0) {
      auto add = [&](int slot, long long v) {
        if (v) atomicAdd(reinterpret_cast<unsigned long long*>(&btot[slot]), (unsigned long long)v);
      };
      add(IGPU_T_GPU_PODS, g);
      add(IGPU_T_SCHEDULED_PODS, s);
      add(IGPU_T_PHASE0 + 0, a0);
      add(IGPU_T_PHASE0 + 1, a1);
      add(IGPU_T_PHASE0 + 2, a2);
      add(IGPU_T_PHASE0 + 3, a3);
      add(IGPU_T_PHASE0 + 4, a4);
      add(IGPU_T_ALLOCATED, al);
    }
    const unsigned nb = __reduce_add_sync(0xffffffffu, bad);
    if (lane == 0 && nb) atomicAdd(reinterpret_cast<unsigned long long*>(&btot[IGPU_T_BAD_INPUT]), (unsigned long long)nb);
  }
  for (uint32_t row = warp; row < 2 * R; row += nwarps) {  // rows 0..R-1 requests, R..2R-1 limits
    const long long* src = sm + (size_t)(2 * R + row) * nt;
    long long v = 0;
    for (int i = lane; i < nt; i += 32) v += src[i];
    v = warp_sum_i64(v);
    if (lane == 0 && v) btot[(row < R ? IGPU_T_REQ0 + row : IGPU_T_LIM0 + (row - R))] = v;  // one writer per row
  }
  __syncthreads();
  for (int i = tid; i < IGPU_TOTALS_LEN; i += nt)
    if (btot[i]) red_add_i64(p.totals + i, btot[i]);
}

// ---- K1, tiled: the same rollup with the dependent chain offsets -> entries taken off the threads -----------
// The per-pod loop above waits twice on HBM for every pod (its offsets, then its entries; ncu: 34 % issue-active,
// stalls on the first use of a loaded entry).  Here a CTA walks tiles of 256 consecutive pods.  The tile's CSR slice
// is contiguous — entries [off[t0], off[t0+256]) — so one thread fetches the two bounding offsets (one tile ahead)
// and starts four bulk copies (entries, offsets, node ids, phases) into a two-stage shared-memory ring; the 256
// threads then fold their pods out of shared memory while the next tile is in flight.  Fleet totals per resource
// live in registers (the loops over r unroll), only the per-pod columns that an entry indexes dynamically stay in
// shared memory.  A tile with more than K1_EC entries reads the excess from global memory (rare).
constexpr int K1_TP = 128;  // pods per tile = threads per CTA.  Measured at C2 (entry-parallel kernel, CTAs per SM scaled to keep 768
                            // threads resident): 256 -> 0.0483 ms, 128 -> 0.0457, 64 -> 0.0667; 10 M pods: 0.380 / 0.363 / 0.558
constexpr int K1_EC = 960;
struct K1Stage {
  int4 ent[K1_EC];
  uint32_t off[K1_TP + 4];
  uint32_t node[K1_TP];
  uint8_t phase[K1_TP];
};

// Sum of a 64-bit integer over the warp (mod 2^64, so signed values work) with REDUX on 22-bit limbs: the upper limbs
// are almost always zero (small non-negative counts), and then cost one vote instead of two reductions.
__device__ __forceinline__ long long warp_sum_i64_limbs(long long v) {
  const unsigned long long u = (unsigned long long)v;
  unsigned long long s = __reduce_add_sync(0xffffffffu, (unsigned)(u & 0x3FFFFFu));
  if (__any_sync(0xffffffffu, (u >> 22) != 0)) {
    s += (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)((u >> 22) & 0x3FFFFFu)) << 22;
    s += (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)(u >> 44)) << 44;
  }
  return (long long)s;
}

template <int RT>
__global__ void __launch_bounds__(K1_TP, 8) pod_rollup_tiled_kernel(const PodParams p) {
  extern __shared__ __align__(128) unsigned char k1_smem[];
  __shared__ uint64_t full[2];
  __shared__ long long btot[IGPU_TOTALS_LEN];
  K1Stage* stg = reinterpret_cast<K1Stage*>(k1_smem);
  long long* cols = reinterpret_cast<long long*>(k1_smem + 2 * sizeof(K1Stage));
  const int tid = threadIdx.x;
  constexpr int nt = K1_TP;
  long long* cur_req = cols + tid;              // [r * nt]
  long long* cur_lim = cols + RT * nt + tid;
  for (int i = tid; i < IGPU_TOTALS_LEN; i += nt) btot[i] = 0;
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
  }
  __syncthreads();

  const uint32_t n_tiles = (p.P + K1_TP - 1) / K1_TP;
  const uint32_t first = blockIdx.x, step = gridDim.x;
  const uint32_t my_count = n_tiles > first ? (n_tiles - first + step - 1) / step : 0;
  // thread 0: bounding offsets of the tile to issue next, fetched one tile ahead of their use
  uint32_t nb0 = 0, nb1 = 0;
  auto fetch_bounds = [&](uint32_t i) {
    const uint32_t t0 = (first + i * step) * K1_TP, cnt = min((uint32_t)K1_TP, p.P - t0);
    nb0 = __ldg(p.pod_off + t0);
    nb1 = __ldg(p.pod_off + t0 + cnt);
  };
  auto issue = [&](uint32_t i) {  // uses nb0 / nb1 fetched for tile i
    const uint32_t t0 = (first + i * step) * K1_TP, cnt = min((uint32_t)K1_TP, p.P - t0);
    K1Stage& st = stg[i & 1];
    // offsets come from the caller: never follow them past the entry stream (a broken tile stages nothing and its
    // pods are rejected one by one below)
    const uint32_t lo = min(nb0, p.E), hi = min(max(nb1, lo), p.E);
    const uint32_t n_ent = min(hi - lo, (uint32_t)K1_EC);
    const uint32_t b_ent = n_ent * 16u, b_off = ((cnt + 1 + 3) & ~3u) * 4u, b_node = ((cnt + 3) & ~3u) * 4u, b_ph = (cnt + 15) & ~15u;
    mbar_expect_tx(&full[i & 1], b_ent + b_off + b_node + b_ph);
    if (b_ent) bulk_g2s(st.ent, p.entries + lo, b_ent, &full[i & 1]);
    bulk_g2s(st.off, p.pod_off + t0, b_off, &full[i & 1]);
    bulk_g2s(st.node, p.pod_node + t0, b_node, &full[i & 1]);
    bulk_g2s(st.phase, p.pod_phase + t0, b_ph, &full[i & 1]);
  };
  if (tid == 0 && my_count) {
    fetch_bounds(0);
    issue(0);
    if (my_count > 1) {
      fetch_bounds(1);
      issue(1);
    }
    if (my_count > 2) fetch_bounds(2);
  }

  unsigned n_gpu = 0, n_sched = 0, ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0, bad = 0;
  long long allocated = 0;
  long long tot_req[RT], tot_lim[RT];
#pragma unroll
  for (int r = 0; r < RT; ++r) tot_req[r] = 0, tot_lim[r] = 0;

  for (uint32_t i = 0; i < my_count; ++i) {
    const uint32_t t0 = (first + i * step) * K1_TP, cnt = min((uint32_t)K1_TP, p.P - t0);
    K1Stage& st = stg[i & 1];
    mbar_wait(&full[i & 1], (i >> 1) & 1u);
    uint32_t meta = 0, my_node = IGPU_NODE_NONE;
    long long my_inuse = 0;
    if ((uint32_t)tid < cnt) {
      const uint32_t pod = t0 + tid;
      const uint32_t base = st.off[0];
      const uint32_t raw0 = st.off[tid], raw1 = st.off[tid + 1];
      // column invariants are checked, not trusted: offsets ascending and inside the stream, node id inside the node
      // columns, resource ids inside the planes.  A violation is counted (IGPU_T_BAD_INPUT) and skipped.
      const bool off_ok = raw0 >= base && raw1 >= raw0 && raw1 <= p.E;
      bad += !off_ok;
      const uint32_t e0 = off_ok ? raw0 - base : 0u, e1 = off_ok ? raw1 - base : 0u;
      const uint32_t ph = st.phase[tid];
      uint32_t node = st.node[tid];
      if (node != IGPU_NODE_NONE && node >= p.N) node = IGPU_NODE_NONE, ++bad;
#pragma unroll
      for (int r = 0; r < RT; ++r) cur_req[r * nt] = 0, cur_lim[r * nt] = 0;
      uint32_t req_mask = 0, lim_mask = 0;
      long long inuse_main = 0;
      uint32_t n_cont = 0, prev_cont = 0xFFFFFFFFu;  // GPU containers of this pod (PodDetailSection.tsx:48-52)
      for (uint32_t e = e0; e < e1; ++e) {
        const int4 raw = e < (uint32_t)K1_EC ? st.ent[e] : __ldg(p.entries + base + e);
        const uint32_t res = (uint32_t)raw.y & 0xFFFFu;
        const uint32_t fl = ((uint32_t)raw.y >> 16) & 0xFFu;
        const long long v = (long long)(((unsigned long long)(uint32_t)raw.w << 32) | (uint32_t)raw.z);
        if ((uint32_t)raw.x != pod) {  // entries[i].pod must agree with pod_entry_off: counted, not followed
          ++bad;
          continue;
        }
        {  // a main container counts once: its entries are adjacent (igpu_rollup.h, igpu_entry)
          const uint32_t cont = (uint32_t)raw.y >> 24;
          const bool is_main = !(fl & IGPU_ENTRY_INIT);
          n_cont += (is_main && cont != prev_cont) ? 1u : 0u;
          prev_cont = is_main ? cont : prev_cont;
        }
        if (res >= (uint32_t)RT) {
          ++bad;
          continue;
        }
        // straight-line fold (the lanes of a warp already diverge on their entry counts; no branches inside):
        // a truthy request adds to its column and sets its mask bit (k8s.ts:295-296), a truthy limit likewise;
        // requests of main containers for i915 / xe also count as in use, truthy or not (NodeDetailSection.tsx:78-86)
        const bool is_lim = fl & IGPU_ENTRY_LIMIT, truthy = fl & IGPU_ENTRY_TRUTHY;
        long long* dst = (is_lim ? cur_lim : cur_req) + res * nt;
        *dst += truthy ? v : 0;
        const uint32_t bit = truthy ? 1u << res : 0u;
        req_mask |= is_lim ? 0u : bit;
        lim_mask |= is_lim ? bit : 0u;
        inuse_main += (!is_lim && !(fl & IGPU_ENTRY_INIT) && res <= IGPU_RES_XE) ? v : 0;
      }
      const bool is_gpu = e1 > e0;  // k8s.ts:257-263
      p.pod_is_gpu[pod] = is_gpu ? 1 : 0;
      p.pod_gpu_containers[pod] = (uint8_t)min(n_cont, 255u);
      p.pod_req_mask[pod] = req_mask;
#pragma unroll
      for (int r = 0; r < RT; ++r) p.pod_req[(size_t)r * p.P + pod] = cur_req[r * nt];
      if (p.want_lim) {
        p.pod_lim_mask[pod] = lim_mask;
#pragma unroll
        for (int r = 0; r < RT; ++r) p.pod_lim[(size_t)r * p.P + pod] = cur_lim[r * nt];
      }
      if (is_gpu) {
        ++n_gpu;
        ph0 += ph == IGPU_PHASE_RUNNING;
        ph1 += ph == IGPU_PHASE_PENDING;
        ph2 += ph == IGPU_PHASE_SUCCEEDED;
        ph3 += ph == IGPU_PHASE_FAILED;
        ph4 += ph == IGPU_PHASE_OTHER;
        const bool running = ph == IGPU_PHASE_RUNNING;
        const bool sched = node != IGPU_NODE_NONE;
        n_sched += sched;
        meta = (sched ? 1u : 0u) | (running && sched ? 2u : 0u);
        if (running) {
          allocated += cur_req[IGPU_RES_I915 * nt] + cur_req[IGPU_RES_XE * nt];  // OverviewPage.tsx:105-113
#pragma unroll
          for (int r = 0; r < RT; ++r) {
            tot_req[r] += cur_req[r * nt];
            if (p.want_lim) tot_lim[r] += cur_lim[r * nt];
          }
        }
      }
      my_inuse = (meta & 2u) ? inuse_main : 0;
      my_node = node;
    }
    // ---- per-node accumulation, warp by warp ---------------------------------------------------------------------
    // Pods of one node that sit next to each other (the host marshal can group them: FleetColumns.grouped_by_node)
    // let a warp add its 32 pods up first and issue ONE set of REDs per node instead of one per pod: up to two nodes
    // per warp are reduced (REDUX on 22-bit limbs of the 64-bit values), anything beyond that — API-order pods, a
    // different node in every lane — keeps the per-pod REDs.  The sums are the same integers either way, so the
    // choice never shows in the results.
    {
      const int lane = tid & 31;
      const uint32_t prev_node = __shfl_up_sync(0xffffffffu, my_node, 1);
      const uint32_t changes = __ballot_sync(0xffffffffu, lane > 0 && meta != 0 && my_node != prev_node);
      uint32_t remaining = __ballot_sync(0xffffffffu, meta != 0);
      if (__popc(changes) <= 2) {  // warp-uniform: at most three runs in this warp
        for (int g = 0; g < 2 && remaining; ++g) {
          const int leader = __ffs(remaining) - 1;
          const uint32_t nd = __shfl_sync(0xffffffffu, my_node, leader);
          const bool in = meta != 0 && my_node == nd;
          remaining &= ~__ballot_sync(0xffffffffu, in);
          const bool sums = in && (meta & 2u);
          const unsigned v_cnt = __reduce_add_sync(0xffffffffu, in ? (meta & 1u) : 0u);
          const long long v_inuse = warp_sum_i64_limbs(sums ? my_inuse : 0);
          if (lane == leader) {
            if (v_cnt) red_add_u32(p.node_pod_count + nd, v_cnt);  // NodesPage.tsx:156-158 (any phase)
            if (v_inuse) red_add_i64(p.node_inuse + nd, v_inuse);
          }
#pragma unroll
          for (int r = 0; r < RT; ++r) {
            const long long q = warp_sum_i64_limbs(sums ? cur_req[r * nt] : 0);
            if (lane == leader && q) red_add_i64(p.node_req + (size_t)r * p.N + nd, q);
            if (p.want_lim) {
              const long long l = warp_sum_i64_limbs(sums ? cur_lim[r * nt] : 0);
              if (lane == leader && l) red_add_i64(p.node_lim + (size_t)r * p.N + nd, l);
            }
          }
        }
      }
      if ((remaining >> lane) & 1u) {  // not covered above: this pod adds itself
        if (meta & 1u) red_add_u32(p.node_pod_count + my_node, 1u);
        if (meta & 2u) {
          if (my_inuse != 0) red_add_i64(p.node_inuse + my_node, my_inuse);
#pragma unroll
          for (int r = 0; r < RT; ++r) {
            const long long q = cur_req[r * nt];
            if (q != 0) red_add_i64(p.node_req + (size_t)r * p.N + my_node, q);
            if (p.want_lim) {
              const long long l = cur_lim[r * nt];
              if (l != 0) red_add_i64(p.node_lim + (size_t)r * p.N + my_node, l);
            }
          }
        }
      }
    }
    __syncthreads();  // every thread is done with this stage
    if (tid == 0 && i + 2 < my_count) {
      issue(i + 2);  // bounds were fetched an iteration ago
      if (i + 3 < my_count) fetch_bounds(i + 3);
    }
  }

  // ---- block reduction of the fleet totals ---------------------------------------------------------------
  const int lane = tid & 31;
  auto add = [&](int slot, long long v) {
    v = warp_sum_i64(v);
    if (lane == 0 && v) atomicAdd(reinterpret_cast<unsigned long long*>(&btot[slot]), (unsigned long long)v);
  };
  add(IGPU_T_GPU_PODS, n_gpu);
  add(IGPU_T_SCHEDULED_PODS, n_sched);
  add(IGPU_T_PHASE0 + 0, ph0);
  add(IGPU_T_PHASE0 + 1, ph1);
  add(IGPU_T_PHASE0 + 2, ph2);
  add(IGPU_T_PHASE0 + 3, ph3);
  add(IGPU_T_PHASE0 + 4, ph4);
  add(IGPU_T_ALLOCATED, allocated);
  add(IGPU_T_BAD_INPUT, bad);
#pragma unroll
  for (int r = 0; r < RT; ++r) add(IGPU_T_REQ0 + r, tot_req[r]), add(IGPU_T_LIM0 + r, tot_lim[r]);
  __syncthreads();
  for (int i = tid; i < IGPU_TOTALS_LEN; i += nt)
    if (btot[i]) red_add_i64(p.totals + i, btot[i]);
}


// ---- K1, entry-parallel: one lane per ENTRY, then one thread per pod ------------------------------------------------
// The tiled kernel above lets thread t fold pod t's entries: a warp runs as many iterations as its longest pod (ncu:
// 9.7 for an average of 3.6 entries, ~52 instructions each, 64-bit read-modify-writes of shared memory), and reduces
// eleven 64-bit quantities per node run with limb REDUX — 1,340 warp instructions per 32 pods, issue-bound at 0.38 of the
// HBM roofline.  Here the tile's entries are spread over the 256 threads (3.6 converged iterations per 32 pods).  A lane
// adds its value into the pod's row of 32-bit accumulators with a native shared-memory RED (row stride 2R+3 words, odd,
// so the entries of a pod and of neighbouring pods fall on different banks), ORs its mask bit, counts its container when
// it opens one — and, when the pod is Running on a node, adds the same value into a small per-tile table of node sums
// (pods grouped by node put a tile's nodes in one short ascending range; a pod on a node outside the table is added by
// its own thread afterwards, one RED per quantity, which is what API-order input costs).  A thread per pod then reads its row back and writes the planes; the
// node table is flushed with one global RED per touched word; pods per node are counted by the first lane of each run.
// 32 bits are exact because a fast tile holds at most K1E_EC entries and every value is checked to lie in [0, 2^20): any
// sum inside the tile stays below 2^31.  A tile that breaks the condition (a negative parseInt, memory in bytes, more
// entries than the staging buffer) is detected by the lanes that read it and is folded by the 64-bit pod-per-thread loop
// instead; integer sums are the same either way, so the choice never shows in the results.
constexpr uint32_t K1_FAST_MAX_VALUE = 1u << 20;
constexpr int K1E_EC = 640;     // staged entries per tile (mean 3.6 x 128 = 461, sigma ~32)
constexpr int K1E_SLOTS = 16;   // nodes per tile with shared-memory sums
struct K1EStage {
  int4 ent[K1E_EC];
  uint32_t off[K1_TP + 4];
  uint32_t node[K1_TP];
  uint8_t phase[K1_TP];
};

__device__ __forceinline__ void sred_add(uint32_t saddr, uint32_t v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory"); }
__device__ __forceinline__ void sred_or(uint32_t saddr, uint32_t v) { asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory"); }

template <int RT>
__global__ void __launch_bounds__(K1_TP, 6) pod_rollup_ep_kernel(const PodParams p) {
  constexpr int S = 2 * RT + 3;  // words per pod row: req[RT] lim[RT] inuse containers masks(req | lim << 16)
  constexpr int W_INUSE = 2 * RT, W_CONT = 2 * RT + 1, W_MASK = 2 * RT + 2;
  constexpr int NW = 2 * RT + 1;  // words per node slot: req[RT] lim[RT] inuse
  constexpr int NSLOT_TOT = 2 * RT + 9;  // per-warp partial totals: 8 counters, allocated, then req[RT], lim[RT]
  static_assert(RT <= 16, "the two masks share one word");
  extern __shared__ __align__(128) unsigned char k1_smem[];
  __shared__ uint64_t full[2];
  __shared__ uint32_t tile_slow[2];
  __shared__ uint32_t ntab[K1E_SLOTS * NW];
  __shared__ long long wtot[K1_TP / 32][NSLOT_TOT];
  K1EStage* stg = reinterpret_cast<K1EStage*>(k1_smem);
  uint32_t* acc = reinterpret_cast<uint32_t*>(k1_smem + 2 * sizeof(K1EStage));    // [K1_TP][S], zero between tiles
  long long* cols = reinterpret_cast<long long*>(k1_smem + 2 * sizeof(K1EStage));  // [2 RT][K1_TP]: the 64-bit path, same bytes
  const uint32_t acc_s = smem_u32(acc), ntab_s = smem_u32(ntab);
  const int tid = threadIdx.x, lane = tid & 31;
  constexpr int nt = K1_TP;
  for (int i = tid; i < K1_TP * S; i += nt) acc[i] = 0;
  for (int i = tid; i < K1E_SLOTS * NW; i += nt) ntab[i] = 0;
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
    tile_slow[0] = tile_slow[1] = 0;
  }
  __syncthreads();

  const uint32_t n_tiles = (p.P + K1_TP - 1) / K1_TP;
  const uint32_t first = blockIdx.x, step = gridDim.x;
  const uint32_t my_count = n_tiles > first ? (n_tiles - first + step - 1) / step : 0;
  uint32_t nb0 = 0, nb1 = 0;  // thread 0: bounding offsets of the tile to issue next, fetched one tile ahead
  auto fetch_bounds = [&](uint32_t i) {
    const uint32_t t0 = (first + i * step) * K1_TP, cnt = min((uint32_t)K1_TP, p.P - t0);
    nb0 = __ldg(p.pod_off + t0);
    nb1 = __ldg(p.pod_off + t0 + cnt);
  };
  auto issue = [&](uint32_t i) {
    const uint32_t t0 = (first + i * step) * K1_TP, cnt = min((uint32_t)K1_TP, p.P - t0);
    K1EStage& st = stg[i & 1];
    const uint32_t lo = min(nb0, p.E), hi = min(max(nb1, lo), p.E);  // never past the entry stream
    const uint32_t n_ent = min(hi - lo, (uint32_t)K1E_EC);
    const uint32_t b_ent = n_ent * 16u, b_off = ((cnt + 1 + 3) & ~3u) * 4u, b_node = ((cnt + 3) & ~3u) * 4u, b_ph = (cnt + 15) & ~15u;
    mbar_expect_tx(&full[i & 1], b_ent + b_off + b_node + b_ph);
    if (b_ent) bulk_g2s(st.ent, p.entries + lo, b_ent, &full[i & 1]);
    bulk_g2s(st.off, p.pod_off + t0, b_off, &full[i & 1]);
    bulk_g2s(st.node, p.pod_node + t0, b_node, &full[i & 1]);
    bulk_g2s(st.phase, p.pod_phase + t0, b_ph, &full[i & 1]);
  };
  if (tid == 0 && my_count) {
    fetch_bounds(0);
    issue(0);
    if (my_count > 1) {
      fetch_bounds(1);
      issue(1);
    }
    if (my_count > 2) fetch_bounds(2);
  }

  unsigned n_gpu = 0, n_sched = 0, ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0, bad = 0;
  long long allocated = 0;
  long long tot_req[RT], tot_lim[RT];
#pragma unroll
  for (int r = 0; r < RT; ++r) tot_req[r] = 0, tot_lim[r] = 0;

  for (uint32_t i = 0; i < my_count; ++i) {
    const uint32_t t0 = (first + i * step) * K1_TP, cnt = min((uint32_t)K1_TP, p.P - t0);
    K1EStage& st = stg[i & 1];
    mbar_wait(&full[i & 1], (i >> 1) & 1u);
    const uint32_t base = st.off[0];
    const uint32_t lo = min(base, p.E), n_tile = min(max(st.off[cnt], lo), p.E) - lo;
    const uint32_t node0 = st.node[0];
    const bool use_tab = node0 < p.N;  // (IGPU_NODE_NONE is not)

    // ---- entries: one lane per entry ------------------------------------------------------------------------------
    uint32_t slow_seen = n_tile > (uint32_t)K1E_EC ? 1u : 0u;
    const uint32_t n_staged = min(n_tile, (uint32_t)K1E_EC);  // (a longer tile is folded by the 64-bit path below)
    for (uint32_t e = tid; e < n_staged; e += nt) {
      const int4 raw = st.ent[e];
      const uint32_t py = (uint32_t)st.ent[e ? e - 1 : 0].y;  // the entry before: does this one open a container?
      const uint32_t pl = (uint32_t)raw.x - t0;               // the pod of this tile the entry says it belongs to
      if (pl >= cnt) {
        ++bad;
        continue;
      }
      // column invariants are checked, not trusted: the entry must lie inside the slice its pod's offsets give
      const uint32_t r0 = st.off[pl], r1 = st.off[pl + 1], ge = lo + e;
      if (r1 < r0 || ge - r0 >= r1 - r0) {
        ++bad;
        continue;
      }
      const uint32_t res = (uint32_t)raw.y & 0xFFFFu, fl = ((uint32_t)raw.y >> 16) & 0xFFu, cont = (uint32_t)raw.y >> 24;
      const uint32_t row = acc_s + pl * (S * 4);
      const bool is_main = !(fl & IGPU_ENTRY_INIT);
      // a main container is counted by its first entry (its entries are adjacent; igpu_rollup.h, igpu_entry)
      if (is_main && (ge == r0 || (py >> 24) != cont || ((py >> 16) & IGPU_ENTRY_INIT))) sred_add(row + W_CONT * 4, 1u);
      if (res >= (uint32_t)RT) {
        ++bad;
        continue;
      }
      const uint32_t v = (uint32_t)raw.z;
      slow_seen |= (raw.w != 0 || v >= K1_FAST_MAX_VALUE) ? 1u : 0u;
      const bool is_lim = fl & IGPU_ENTRY_LIMIT, truthy = fl & IGPU_ENTRY_TRUTHY;
      // requests of main containers for i915 / xe count as in use, truthy or not (NodeDetailSection.tsx:78-86)
      const bool inuse = !is_lim && is_main && res <= IGPU_RES_XE;
      const uint32_t word = (is_lim ? RT : 0) + res;
      if (truthy) {  // k8s.ts:295-296; the limits rollup likewise
        sred_add(row + word * 4, v);
        sred_or(row + W_MASK * 4, 1u << (res + (is_lim ? 16u : 0u)));
      }
      if (inuse) sred_add(row + W_INUSE * 4, v);
      // per-node sums take Running pods that sit on a known node (OverviewPage.tsx:105-113, NodeDetailSection.tsx:77-87);
      // a node outside the tile's table is served by its pod's thread below (nothing irrevocable happens in this loop:
      // the tile may still turn out to need the 64-bit path)
      const uint32_t slot = st.node[pl] - node0;
      if (use_tab && slot < (uint32_t)K1E_SLOTS && st.phase[pl] == IGPU_PHASE_RUNNING) {
        const uint32_t nrow = ntab_s + slot * (NW * 4);
        if (truthy) sred_add(nrow + word * 4, v);
        if (inuse) sred_add(nrow + W_INUSE * 4, v);
      }
    }
    if (slow_seen) tile_slow[i & 1] = 1u;
    __syncthreads();
    const bool slow = tile_slow[i & 1] != 0;  // CTA-uniform
    if (tid == 0) tile_slow[(i + 1) & 1] = 0;  // its last readers left a tile ago, its next writers start after this tile's last barrier

    // ---- pods: one thread per pod -----------------------------------------------------------------------------------
    const bool have = (uint32_t)tid < cnt;
    uint32_t meta = 0, my_node = IGPU_NODE_NONE;
    bool is_gpu = false;
    uint32_t ph = IGPU_PHASE_OTHER;
    uint32_t e0 = 0, e1 = 0;
    if (have) {
      const uint32_t raw0 = st.off[tid], raw1 = st.off[tid + 1];
      const bool off_ok = raw0 >= base && raw1 >= raw0 && raw1 <= p.E;
      bad += !off_ok;
      e0 = off_ok ? raw0 - base : 0u, e1 = off_ok ? raw1 - base : 0u;
      is_gpu = e1 > e0;  // k8s.ts:257-263
      ph = st.phase[tid];
      uint32_t node = st.node[tid];
      if (node != IGPU_NODE_NONE && node >= p.N) node = IGPU_NODE_NONE, ++bad;
      my_node = node;
      if (is_gpu) {
        ++n_gpu;
        ph0 += ph == IGPU_PHASE_RUNNING;
        ph1 += ph == IGPU_PHASE_PENDING;
        ph2 += ph == IGPU_PHASE_SUCCEEDED;
        ph3 += ph == IGPU_PHASE_FAILED;
        ph4 += ph == IGPU_PHASE_OTHER;
        const bool sched = node != IGPU_NODE_NONE;
        n_sched += sched;
        meta = (sched ? 1u : 0u) | (ph == IGPU_PHASE_RUNNING && sched ? 2u : 0u);
      }
    }
    const uint32_t pod = t0 + tid;
    if (!slow) {
      if (have) {
        uint32_t* row = acc + tid * S;
        uint32_t q[RT], l[RT];
#pragma unroll
        for (int r = 0; r < RT; ++r) q[r] = row[r], l[r] = row[RT + r];
        uint32_t n_cont = row[W_CONT], m = row[W_MASK];
        const uint32_t my_inuse = row[W_INUSE];
#pragma unroll
        for (int w = 0; w < S; ++w) row[w] = 0;  // ready for the next tile
        if (!is_gpu) {  // a pod whose offsets were rejected owns nothing, whatever the lanes added for it
#pragma unroll
          for (int r = 0; r < RT; ++r) q[r] = 0, l[r] = 0;
          n_cont = 0, m = 0;
        }
        if ((meta & 2u) && !(use_tab && my_node - node0 < (uint32_t)K1E_SLOTS)) {  // Running on a node outside the table
          if (my_inuse) red_add_i64(p.node_inuse + my_node, (long long)my_inuse);
#pragma unroll
          for (int r = 0; r < RT; ++r) {
            if (q[r]) red_add_i64(p.node_req + (size_t)r * p.N + my_node, (long long)q[r]);
            if (p.want_lim && l[r]) red_add_i64(p.node_lim + (size_t)r * p.N + my_node, (long long)l[r]);
          }
        }
        p.pod_is_gpu[pod] = is_gpu ? 1 : 0;
        p.pod_gpu_containers[pod] = (uint8_t)min(n_cont, 255u);
        p.pod_req_mask[pod] = m & 0xFFFFu;
#pragma unroll
        for (int r = 0; r < RT; ++r) p.pod_req[(size_t)r * p.P + pod] = (long long)q[r];
        if (p.want_lim) {
          p.pod_lim_mask[pod] = m >> 16;
#pragma unroll
          for (int r = 0; r < RT; ++r) p.pod_lim[(size_t)r * p.P + pod] = (long long)l[r];
        }
        if (is_gpu && ph == IGPU_PHASE_RUNNING) {
          allocated += (long long)(q[IGPU_RES_I915] + q[IGPU_RES_XE]);  // OverviewPage.tsx:105-113 (each < 2^31)
#pragma unroll
          for (int r = 0; r < RT; ++r) {
            tot_req[r] += q[r];
            if (p.want_lim) tot_lim[r] += l[r];
          }
        }
      }
      // node sums of the tile: one global RED per touched word of the table
      for (uint32_t t = tid; t < (uint32_t)(K1E_SLOTS * NW); t += nt) {  // (more words than threads when R = 8)
        const uint32_t v = ntab[t];
        if (v) {
          ntab[t] = 0;
          const uint32_t slot = t / NW, w = t % NW, node = node0 + slot;
          int64_t* dst = w < (uint32_t)RT ? p.node_req + (size_t)w * p.N + node
                                          : (w < 2u * RT ? p.node_lim + (size_t)(w - RT) * p.N + node : p.node_inuse + node);
          // (a pod_node beyond the node columns was counted as bad input by its pod's thread; it must not be followed here)
          if (node < p.N && (p.want_lim || w < (uint32_t)RT || w == (uint32_t)W_INUSE)) red_add_i64(dst, (long long)v);
        }
      }
      // pods per node (NodesPage.tsx:156-158, any phase): the first lane of each run of equal nodes adds the run's count
      {
        const uint32_t counted = __ballot_sync(0xffffffffu, (meta & 1u) != 0);
        const uint32_t prev_node = __shfl_up_sync(0xffffffffu, my_node, 1);
        const uint32_t heads = __ballot_sync(0xffffffffu, lane == 0 || my_node != prev_node);
        if (counted && ((heads >> lane) & 1u) && my_node != IGPU_NODE_NONE) {
          // a run is a maximal stretch of lanes with EQUAL my_node, so every counted lane of the run shares the head's node
          const uint32_t above = lane == 31 ? 0u : (heads >> (lane + 1)) << (lane + 1);  // heads after this lane
          const uint32_t run = (above ? ((1u << (__ffs(above) - 1)) - 1u) : 0xffffffffu) & ~((1u << lane) - 1u);
          const uint32_t c = (uint32_t)__popc(counted & run);
          if (c) red_add_u32(p.node_pod_count + my_node, c);
        }
      }
    } else {
      // 64-bit path: thread t folds pod t into its own shared-memory column (the accumulator bytes, re-used); per-node
      // contributions by warp reduction on 22-bit limbs.  The 32-bit partial sums of this tile are discarded.
      long long* cur_req = cols + tid;
      long long* cur_lim = cols + RT * nt + tid;
      long long req[RT], lim[RT], my_inuse = 0;
      uint32_t req_mask = 0, lim_mask = 0, n_cont = 0;
#pragma unroll
      for (int r = 0; r < RT; ++r) req[r] = 0, lim[r] = 0;
      for (int t = tid; t < K1E_SLOTS * NW; t += nt) ntab[t] = 0;
      if (have) {
#pragma unroll
        for (int r = 0; r < RT; ++r) cur_req[r * nt] = 0, cur_lim[r * nt] = 0;
        uint32_t prev_cont = 0xFFFFFFFFu;
        for (uint32_t e = e0; e < e1; ++e) {
          const int4 raw = e < (uint32_t)K1E_EC ? st.ent[e] : __ldg(p.entries + base + e);
          if ((uint32_t)raw.x != pod) {  // counted above when it was staged; beyond the staging buffer it is counted here
            bad += e >= (uint32_t)K1E_EC;
            continue;
          }
          const uint32_t res = (uint32_t)raw.y & 0xFFFFu, fl = ((uint32_t)raw.y >> 16) & 0xFFu, cont = (uint32_t)raw.y >> 24;
          const long long v = (long long)(((unsigned long long)(uint32_t)raw.w << 32) | (uint32_t)raw.z);
          const bool is_main = !(fl & IGPU_ENTRY_INIT);
          n_cont += (is_main && cont != prev_cont) ? 1u : 0u;
          prev_cont = is_main ? cont : prev_cont;
          if (res >= (uint32_t)RT) {
            bad += e >= (uint32_t)K1E_EC;
            continue;
          }
          const bool is_lim = fl & IGPU_ENTRY_LIMIT, truthy = fl & IGPU_ENTRY_TRUTHY;
          long long* dst = (is_lim ? cur_lim : cur_req) + res * nt;
          *dst += truthy ? v : 0;
          const uint32_t bit = truthy ? 1u << res : 0u;
          req_mask |= is_lim ? 0u : bit;
          lim_mask |= is_lim ? bit : 0u;
          my_inuse += (!is_lim && is_main && res <= IGPU_RES_XE) ? v : 0;
        }
#pragma unroll
        for (int r = 0; r < RT; ++r) req[r] = cur_req[r * nt], lim[r] = cur_lim[r * nt];
        p.pod_is_gpu[pod] = is_gpu ? 1 : 0;
        p.pod_gpu_containers[pod] = (uint8_t)min(n_cont, 255u);
        p.pod_req_mask[pod] = req_mask;
#pragma unroll
        for (int r = 0; r < RT; ++r) p.pod_req[(size_t)r * p.P + pod] = req[r];
        if (p.want_lim) {
          p.pod_lim_mask[pod] = lim_mask;
#pragma unroll
          for (int r = 0; r < RT; ++r) p.pod_lim[(size_t)r * p.P + pod] = lim[r];
        }
        if (is_gpu && ph == IGPU_PHASE_RUNNING) {
          allocated += req[IGPU_RES_I915] + req[IGPU_RES_XE];
#pragma unroll
          for (int r = 0; r < RT; ++r) {
            tot_req[r] += req[r];
            if (p.want_lim) tot_lim[r] += lim[r];
          }
        }
        if (!(meta & 2u)) my_inuse = 0;
      }
      __syncthreads();  // the columns are read; the bytes become accumulator rows again
      for (int w = tid; w < K1_TP * S; w += nt) acc[w] = 0;
      {
        const uint32_t prev_node = __shfl_up_sync(0xffffffffu, my_node, 1);
        const uint32_t changes = __ballot_sync(0xffffffffu, lane > 0 && meta != 0 && my_node != prev_node);
        uint32_t remaining = __ballot_sync(0xffffffffu, meta != 0);
        if (__popc(changes) <= 2) {
          for (int g = 0; g < 2 && remaining; ++g) {
            const int leader = __ffs(remaining) - 1;
            const uint32_t nd = __shfl_sync(0xffffffffu, my_node, leader);
            const bool in = meta != 0 && my_node == nd;
            remaining &= ~__ballot_sync(0xffffffffu, in);
            const bool sums = in && (meta & 2u);
            const unsigned v_cnt = __reduce_add_sync(0xffffffffu, in ? (meta & 1u) : 0u);
            const long long v_inuse = warp_sum_i64_limbs(sums ? my_inuse : 0);
            if (lane == leader) {
              if (v_cnt) red_add_u32(p.node_pod_count + nd, v_cnt);
              if (v_inuse) red_add_i64(p.node_inuse + nd, v_inuse);
            }
#pragma unroll
            for (int r = 0; r < RT; ++r) {
              const long long qs = warp_sum_i64_limbs(sums ? req[r] : 0);
              if (lane == leader && qs) red_add_i64(p.node_req + (size_t)r * p.N + nd, qs);
              if (p.want_lim) {
                const long long ls = warp_sum_i64_limbs(sums ? lim[r] : 0);
                if (lane == leader && ls) red_add_i64(p.node_lim + (size_t)r * p.N + nd, ls);
              }
            }
          }
        }
        if ((remaining >> lane) & 1u) {  // not covered above: this pod adds itself
          if (meta & 1u) red_add_u32(p.node_pod_count + my_node, 1u);
          if (meta & 2u) {
            if (my_inuse != 0) red_add_i64(p.node_inuse + my_node, my_inuse);
#pragma unroll
            for (int r = 0; r < RT; ++r) {
              if (req[r] != 0) red_add_i64(p.node_req + (size_t)r * p.N + my_node, req[r]);
              if (p.want_lim && lim[r] != 0) red_add_i64(p.node_lim + (size_t)r * p.N + my_node, lim[r]);
            }
          }
        }
      }
    }
    __syncthreads();  // every thread is done with this stage, the accumulator rows and the node table
    if (tid == 0 && i + 2 < my_count) {
      issue(i + 2);  // bounds were fetched an iteration ago
      if (i + 3 < my_count) fetch_bounds(i + 3);
    }
  }

  // ---- fleet totals: warp sums to shared memory, then one thread per slot adds the eight warps and issues one RED ------
  {
    const int warp = tid >> 5;
    long long* mine = wtot[warp];
    auto put32 = [&](int k, unsigned v) {
      v = __reduce_add_sync(0xffffffffu, v);
      if (lane == 0) mine[k] = (long long)v;
    };
    auto put64 = [&](int k, long long v) {
      v = warp_sum_i64_limbs(v);
      if (lane == 0) mine[k] = v;
    };
    put32(0, n_gpu), put32(1, n_sched), put32(2, ph0), put32(3, ph1), put32(4, ph2), put32(5, ph3), put32(6, ph4), put32(7, bad);
    put64(8, allocated);
#pragma unroll
    for (int r = 0; r < RT; ++r) put64(9 + r, tot_req[r]), put64(9 + RT + r, tot_lim[r]);
    __syncthreads();
    if (tid < NSLOT_TOT) {
      long long v = 0;
#pragma unroll
      for (int w = 0; w < K1_TP / 32; ++w) v += wtot[w][tid];
      const int slot = tid == 0   ? IGPU_T_GPU_PODS
                       : tid == 1 ? IGPU_T_SCHEDULED_PODS
                       : tid < 7  ? IGPU_T_PHASE0 + (tid - 2)
                       : tid == 7 ? IGPU_T_BAD_INPUT
                       : tid == 8 ? IGPU_T_ALLOCATED
                       : (tid < 9 + RT ? IGPU_T_REQ0 + (tid - 9) : IGPU_T_LIM0 + (tid - 9 - RT));
      if (v) red_add_i64(p.totals + slot, v);
    }
  }
}

struct NodeParams {
  uint32_t N, R;
  const uint8_t* __restrict__ node_flags;
  const int32_t* __restrict__ cap_i915;
  const int32_t* __restrict__ cap_xe;
  const int32_t* __restrict__ alloc_i915;
  const int32_t* __restrict__ alloc_xe;
  const int32_t* __restrict__ alloc_own;
  const uint32_t* __restrict__ node_pod_count;
  const int64_t* __restrict__ node_inuse;
  uint8_t* __restrict__ out_flags;
  uint8_t* __restrict__ type;
  int64_t* __restrict__ gpu_count;
  int64_t* __restrict__ cap_total;
  int64_t* __restrict__ alloc_total;
  int64_t* __restrict__ bar_pct;
  uint8_t* __restrict__ bar_level;
  int64_t* __restrict__ util_pct;
  uint8_t* __restrict__ util_level;
  int64_t* totals;
};

__device__ __forceinline__ long long sat_i64(double x) {
  if (x >= 9.2e18) return 0x7fffffffffffffffLL;
  if (x <= -9.2e18) return (long long)0x8000000000000000ULL;
  return (long long)x;
}

__global__ void __launch_bounds__(256) node_rollup_kernel(const NodeParams p) {
  __shared__ long long btot[16];
  if (threadIdx.x < 16) btot[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  long long t[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // slots IGPU_T_GPU_NODES .. IGPU_T_ALLOC_NAN
  if (n < p.N) {
    const uint32_t f = p.node_flags[n];
    const bool is_gpu = (f & (IGPU_NODE_DISCRETE_ROLE | IGPU_NODE_IGPU_ROLE | IGPU_NODE_NFD_LABEL | IGPU_NODE_HAS_GPU_CAPACITY)) != 0;
    const uint32_t type = (f & IGPU_NODE_DISCRETE_ROLE) ? IGPU_TYPE_DISCRETE : (f & IGPU_NODE_IGPU_ROLE) ? IGPU_TYPE_INTEGRATED : IGPU_TYPE_UNKNOWN;
    const long long gpu_count = (long long)p.cap_i915[n] + (long long)p.cap_xe[n];      // k8s.ts:174-178
    const bool cap_nan = f & IGPU_NODE_CAP_NAN, alloc_nan = f & IGPU_NODE_ALLOC_NAN;
    const long long total_alloc = alloc_nan ? 0 : (long long)p.alloc_i915[n] + (long long)p.alloc_xe[n];
    // `totalAllocatable || gpuCount`, NodesPage.tsx:272 — NaN and 0 are falsy
    const long long bar_alloc = (alloc_nan || total_alloc == 0) ? gpu_count : total_alloc;
    uint32_t of = (is_gpu ? IGPU_NODEOUT_IS_GPU : 0) | ((f & IGPU_NODE_READY) ? IGPU_NODEOUT_READY : 0) |
                  (cap_nan ? IGPU_NODEOUT_CAP_NAN : 0) | (alloc_nan ? IGPU_NODEOUT_ALLOC_NAN : 0);
    long long bar_pct = 0;
    uint32_t bar_level = 0;
    if (bar_alloc == 0) {
      of |= IGPU_NODEOUT_BAR_DASH;  // NodesPage.tsx:36
    } else {
      const double pct = fmin(100.0, js_round_dev((double)p.node_pod_count[n] / (double)bar_alloc * 100.0));  // :37
      bar_pct = sat_i64(pct);
      bar_level = pct >= 90 ? 2 : (pct >= 70 ? 1 : 0);  // :38
    }
    const long long own = p.alloc_own[n];
    const double util = own > 0 ? js_round_dev((double)p.node_inuse[n] / (double)own * 100.0) : 0.0;  // NodeDetailSection.tsx:89
    p.out_flags[n] = (uint8_t)of;
    p.type[n] = (uint8_t)type;
    p.gpu_count[n] = gpu_count;
    p.cap_total[n] = cap_nan ? 0 : gpu_count;
    p.alloc_total[n] = total_alloc;
    p.bar_pct[n] = bar_pct;
    p.bar_level[n] = (uint8_t)bar_level;
    p.util_pct[n] = sat_i64(util);
    p.util_level[n] = util >= 90 ? 2 : (util >= 70 ? 1 : 0);  // :90-91
    if (is_gpu) {  // OverviewPage.tsx:78-102 iterates gpuNodes
      t[IGPU_T_GPU_NODES] = 1;
      t[IGPU_T_DISCRETE + type] = 1;
      t[IGPU_T_GPU_DEVICES] = gpu_count;
      t[IGPU_T_READY_NODES] = (f & IGPU_NODE_READY) ? 1 : 0;
      if (cap_nan) t[IGPU_T_CAP_NAN] = 1; else t[IGPU_T_CAPACITY] = gpu_count;
      if (alloc_nan) t[IGPU_T_ALLOC_NAN] = 1; else t[IGPU_T_ALLOCATABLE] = total_alloc;
    }
  }
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const long long v = warp_sum_i64(t[i]);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(reinterpret_cast<unsigned long long*>(&btot[i]), (unsigned long long)v);
  }
  __syncthreads();
  if (threadIdx.x < 10 && btot[threadIdx.x]) red_add_i64(p.totals + threadIdx.x, btot[threadIdx.x]);
}


// ---- per-container key rows (igpu_fleet_container_rows): PodsPage.tsx:49-62, PodDetailSection.tsx:48-65 ------------------
// Row positions are an exclusive scan of pod_gpu_containers (written by the pod kernel): block sums, a one-CTA scan of
// the block sums, then every block scans its own pods again and each thread writes the rows of its pods.  A thread owns
// CR_PER consecutive pods, so rows come out in (pod, container) order.
constexpr int CR_THREADS = 256, CR_PER = 8, CR_BLOCK = CR_THREADS * CR_PER;

__device__ __forceinline__ uint32_t cr_block_exclusive(uint32_t v, uint32_t* warp_sums, uint32_t& block_total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  uint32_t base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < CR_THREADS / 32; ++w) {
    const uint32_t ws = warp_sums[w];
    base += w < warp ? ws : 0u;
    total += ws;
  }
  __syncthreads();
  block_total = total;
  return base + inc - v;
}

__global__ void __launch_bounds__(CR_THREADS) cont_block_sums_kernel(const uint8_t* __restrict__ cnt, uint32_t P, uint32_t* __restrict__ block_sum) {
  __shared__ uint32_t ws[CR_THREADS / 32];
  const uint32_t p0 = blockIdx.x * CR_BLOCK + threadIdx.x * CR_PER;
  uint32_t v = 0;
#pragma unroll
  for (int j = 0; j < CR_PER; ++j)
    if (p0 + j < P) v += cnt[p0 + j];
  uint32_t total;
  (void)cr_block_exclusive(v, ws, total);
  if (threadIdx.x == 0) block_sum[blockIdx.x] = total;
}

// exclusive scan of block_sum[0..nb) in place; block_sum[nb] = total
__global__ void __launch_bounds__(CR_THREADS) cont_scan_blocks_kernel(uint32_t* __restrict__ block_sum, uint32_t nb) {
  __shared__ uint32_t ws[CR_THREADS / 32];
  uint32_t carry = 0;
  for (uint32_t i0 = 0; i0 < nb; i0 += CR_THREADS) {
    const uint32_t i = i0 + threadIdx.x;
    const uint32_t v = i < nb ? block_sum[i] : 0u;
    uint32_t total;
    const uint32_t ex = cr_block_exclusive(v, ws, total);
    if (i < nb) block_sum[i] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) block_sum[nb] = carry;
}

__global__ void __launch_bounds__(CR_THREADS) cont_rows_kernel(const uint8_t* __restrict__ cnt, const uint32_t* __restrict__ pod_off,
                                                                const int4* __restrict__ entries, uint32_t P, uint32_t E, uint32_t R,
                                                                const uint32_t* __restrict__ block_base, igpu_container_row* __restrict__ rows,
                                                                uint32_t cap) {
  __shared__ uint32_t ws[CR_THREADS / 32];
  const uint32_t p0 = blockIdx.x * CR_BLOCK + threadIdx.x * CR_PER;
  uint32_t c[CR_PER], v = 0;
#pragma unroll
  for (int j = 0; j < CR_PER; ++j) {
    c[j] = p0 + j < P ? cnt[p0 + j] : 0u;
    v += c[j];
  }
  uint32_t total;
  uint32_t at = block_base[blockIdx.x] + cr_block_exclusive(v, ws, total);
  if (!v) return;
#pragma unroll 1
  for (int j = 0; j < CR_PER; ++j) {
    if (!c[j]) continue;
    const uint32_t pod = p0 + j;
    const uint32_t e0 = pod_off[pod], e1 = pod_off[pod + 1];
    if (e1 < e0 || e1 > E) continue;  // broken offsets were counted by the pod kernel; nothing is followed
    const uint32_t end = at + c[j];  // never write past this pod's share, whatever the entries say
    igpu_container_row r = {pod, 0xFFFFFFFFu, 0, 0, 0, 0, 0, 0};
    for (uint32_t e = e0; e < e1; ++e) {
      const int4 raw = __ldg(entries + e);
      const uint32_t res = (uint32_t)raw.y & 0xFFFFu, fl = ((uint32_t)raw.y >> 16) & 0xFFu, cont = (uint32_t)raw.y >> 24;
      if (fl & IGPU_ENTRY_INIT) continue;
      if (cont != r.container) {
        if (r.container != 0xFFFFFFFFu && at < end) {
          if (at < cap) rows[at] = r;
          ++at;
        }
        r = igpu_container_row{pod, cont, 0, 0, 0, 0, 0, 0};
      }
      ++r.n_entries;
      if (res >= R) continue;
      const uint32_t bit = 1u << res;
      const bool truthy = fl & IGPU_ENTRY_TRUTHY;
      if (fl & IGPU_ENTRY_LIMIT) r.lim_keys |= bit, r.lim_truthy |= truthy ? bit : 0u;
      else r.req_keys |= bit, r.req_truthy |= truthy ? bit : 0u;
      r.same |= (fl & IGPU_ENTRY_SAME) ? bit : 0u;
    }
    if (r.container != 0xFFFFFFFFu && at < end) {
      if (at < cap) rows[at] = r;
      ++at;
    }
    at = end;
  }
}

constexpr size_t FLEET_PACK_MAX = 1u << 20;  // arenas up to this size travel as one packed copy per direction

size_t carve(char*& cur, size_t bytes) {
  size_t off = reinterpret_cast<size_t>(cur);
  off = align_up(off, 256);
  cur = reinterpret_cast<char*>(off + bytes);
  return off;
}
template <class T>
T* take(char*& cur, size_t count) {
  return reinterpret_cast<T*>(carve(cur, count * sizeof(T)));
}

// Lays the typed pointers over the three arenas for the current (n_nodes, n_pods, n_entries).
void bind_pointers(igpu_fleet* f) {
  const size_t N = f->n_nodes, P = f->n_pods, E = f->n_entries, R = f->R;
  char* c = f->d_in;
  f->node_flags = take<uint8_t>(c, N);
  f->cap_i915 = take<int32_t>(c, N);
  f->cap_xe = take<int32_t>(c, N);
  f->alloc_i915 = take<int32_t>(c, N);
  f->alloc_xe = take<int32_t>(c, N);
  f->alloc_own = take<int32_t>(c, N);
  f->pod_node = take<uint32_t>(c, P);
  f->pod_phase = take<uint8_t>(c, P);
  f->pod_off = take<uint32_t>(c, P + 1);
  f->entries = take<igpu_entry>(c, E);
  f->in_span = (size_t)(c - f->d_in);
  c = f->d_acc;
  f->node_pod_count = take<uint32_t>(c, N);
  f->node_inuse = take<int64_t>(c, N);
  f->node_req = take<int64_t>(c, R * N);
  f->node_lim = take<int64_t>(c, R * N);
  f->totals = take<int64_t>(c, IGPU_TOTALS_LEN);
  f->acc_bytes = c - f->d_acc;
  c = f->d_out;
  f->pod_is_gpu = take<uint8_t>(c, P);
  f->pod_gpu_containers = take<uint8_t>(c, P);
  f->pod_req_mask = take<uint32_t>(c, P);
  f->pod_lim_mask = take<uint32_t>(c, P);
  f->pod_req = take<int64_t>(c, R * P);
  f->pod_lim = take<int64_t>(c, R * P);
  f->node_out_flags = take<uint8_t>(c, N);
  f->node_type = take<uint8_t>(c, N);
  f->node_bar_level = take<uint8_t>(c, N);
  f->node_util_level = take<uint8_t>(c, N);
  f->node_gpu_count = take<int64_t>(c, N);
  f->node_cap_total = take<int64_t>(c, N);
  f->node_alloc_total = take<int64_t>(c, N);
  f->node_bar_pct = take<int64_t>(c, N);
  f->node_util_pct = take<int64_t>(c, N);
  f->out_span = (size_t)(c - f->d_out);
}

size_t arena_bytes(size_t per_item_bytes_sum, size_t n_columns) { return per_item_bytes_sum + 256 * (n_columns + 1); }

}  // namespace

extern "C" {

int32_t igpu_fleet_create(igpu_ctx* ctx, uint32_t max_nodes, uint32_t max_pods, uint64_t max_entries, uint32_t n_resources,
                          uint32_t options, igpu_fleet** out) {
  if (!ctx || !out) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_fleet_create: NULL argument");
  *out = nullptr;
  IGPU_REQUIRE(ctx, n_resources >= 4 && n_resources <= IGPU_MAX_RESOURCES, IGPU_ERR_RANGE, "n_resources %u outside [4,%d]",
               n_resources, IGPU_MAX_RESOURCES);
  IGPU_REQUIRE(ctx, max_entries < 0xFFFFFFFFull, IGPU_ERR_RANGE, "max_entries %llu does not fit the uint32 offsets",
               (unsigned long long)max_entries);
  DeviceGuard g(ctx->device);
  igpu_fleet* f = new igpu_fleet();
  f->max_nodes = max_nodes;
  f->max_pods = max_pods;
  f->max_entries = max_entries;
  f->R = n_resources;
  f->options = options;
  const size_t N = max_nodes, P = max_pods, E = max_entries, R = n_resources;
  f->in_bytes = arena_bytes(N * 21 + P * 5 + (P + 1) * 4 + E * 16, 10);
  const size_t acc = arena_bytes(N * 12 + 2 * R * N * 8 + IGPU_TOTALS_LEN * 8, 5);
  f->out_bytes = arena_bytes(P * 10 + 2 * R * P * 8 + N * 4 + N * 40, 15);
  cudaError_t e1 = cudaMalloc(&f->d_in, f->in_bytes), e2 = cudaMalloc(&f->d_acc, acc), e3 = cudaMalloc(&f->d_out, f->out_bytes);
  if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) {
    cudaGetLastError();
    if (f->d_in) cudaFree(f->d_in);
    if (f->d_acc) cudaFree(f->d_acc);
    if (f->d_out) cudaFree(f->d_out);
    const size_t ib = f->in_bytes, ob = f->out_bytes;
    delete f;
    return igpu_fail(ctx, IGPU_ERR_OOM, "device allocation for fleet (%zu + %zu + %zu bytes) failed", ib, acc, ob);
  }
  f->n_nodes = max_nodes;
  f->n_pods = max_pods;
  f->n_entries = max_entries;
  bind_pointers(f);
  *out = f;
  return IGPU_OK;
}

int32_t igpu_fleet_destroy(igpu_ctx* ctx, igpu_fleet* f) {
  if (!ctx || !f) return IGPU_OK;
  DeviceGuard g(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaFree(f->d_in);
  cudaFree(f->d_acc);
  cudaFree(f->d_out);
  cudaFree(f->d_cont_block);
  cudaFree(f->d_cont_rows);
  if (f->h_in) cudaFreeHost(f->h_in);
  if (f->h_out) cudaFreeHost(f->h_out);
  if (ctx->cached_fleet == f) ctx->cached_fleet = nullptr;
  delete f;
  return IGPU_OK;
}

int32_t igpu_fleet_upload(igpu_ctx* ctx, igpu_fleet* f, const igpu_fleet_cols* c) {
  if (!ctx || !f || !c) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_fleet_upload: NULL argument");
  IGPU_REQUIRE(ctx, c->n_nodes <= f->max_nodes && c->n_pods <= f->max_pods && c->n_entries <= f->max_entries, IGPU_ERR_RANGE,
               "fleet (%u nodes, %u pods, %llu entries) exceeds the handle's capacity (%u, %u, %llu)", c->n_nodes, c->n_pods,
               (unsigned long long)c->n_entries, f->max_nodes, f->max_pods, (unsigned long long)f->max_entries);
  IGPU_REQUIRE(ctx, c->n_resources == f->R, IGPU_ERR_ARG, "n_resources %u differs from the handle's %u", c->n_resources, f->R);
  IGPU_REQUIRE(ctx, (c->n_nodes == 0 || (c->node_flags && c->node_cap_i915 && c->node_cap_xe && c->node_alloc_i915 &&
                                         c->node_alloc_xe && c->node_alloc_own)) &&
                        c->pod_entry_off && (c->n_pods == 0 || (c->pod_node && c->pod_phase)) && (c->n_entries == 0 || c->entries),
               IGPU_ERR_ARG, "a required input column is NULL");
  IGPU_REQUIRE(ctx, c->pod_entry_off[0] == 0 && c->pod_entry_off[c->n_pods] == c->n_entries, IGPU_ERR_ARG,
               "pod_entry_off must start at 0 and end at n_entries");
  DeviceGuard g(ctx->device);
  f->n_nodes = c->n_nodes;
  f->n_pods = c->n_pods;
  f->n_entries = c->n_entries;
  f->options = c->options;
  bind_pointers(f);
  igpu_mark(ctx, 0);
  const size_t N = c->n_nodes, P = c->n_pods, E = c->n_entries;
  uint64_t bytes = 0;
  auto up = [&](void* dst, const void* src, size_t n) -> cudaError_t {
    if (!n) return cudaSuccess;
    bytes += n;
    return cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, ctx->stream);
  };
  if (f->in_span <= FLEET_PACK_MAX) {
    // small fleet: pack the columns at their arena offsets in pinned staging and send them as one copy
    if (f->h_in_cap < f->in_span) {
      if (f->h_in) cudaFreeHost(f->h_in);
      f->h_in = nullptr;
      f->h_in_cap = 0;
      IGPU_CUDA(ctx, cudaMallocHost(&f->h_in, FLEET_PACK_MAX));
      f->h_in_cap = FLEET_PACK_MAX;
    }
    if (f->stage_busy) IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the previous packed upload still owns h_in
    auto put = [&](const void* ddst, const void* src, size_t n) {
      if (n) memcpy(f->h_in + (reinterpret_cast<const char*>(ddst) - f->d_in), src, n), bytes += n;
    };
    put(f->node_flags, c->node_flags, N);
    put(f->cap_i915, c->node_cap_i915, N * 4);
    put(f->cap_xe, c->node_cap_xe, N * 4);
    put(f->alloc_i915, c->node_alloc_i915, N * 4);
    put(f->alloc_xe, c->node_alloc_xe, N * 4);
    put(f->alloc_own, c->node_alloc_own, N * 4);
    put(f->pod_node, c->pod_node, P * 4);
    put(f->pod_phase, c->pod_phase, P);
    put(f->pod_off, c->pod_entry_off, (P + 1) * 4);
    put(f->entries, c->entries, E * 16);
    IGPU_CUDA(ctx, cudaMemcpyAsync(f->d_in, f->h_in, f->in_span, cudaMemcpyHostToDevice, ctx->stream));
    f->stage_busy = true;
    igpu_mark(ctx, 1);
    ctx->timing.h2d_bytes = bytes;
    f->uploaded = true;
    f->rolled = false;
    return IGPU_OK;
  }
  IGPU_CUDA(ctx, up(f->node_flags, c->node_flags, N));
  IGPU_CUDA(ctx, up(f->cap_i915, c->node_cap_i915, N * 4));
  IGPU_CUDA(ctx, up(f->cap_xe, c->node_cap_xe, N * 4));
  IGPU_CUDA(ctx, up(f->alloc_i915, c->node_alloc_i915, N * 4));
  IGPU_CUDA(ctx, up(f->alloc_xe, c->node_alloc_xe, N * 4));
  IGPU_CUDA(ctx, up(f->alloc_own, c->node_alloc_own, N * 4));
  IGPU_CUDA(ctx, up(f->pod_node, c->pod_node, P * 4));
  IGPU_CUDA(ctx, up(f->pod_phase, c->pod_phase, P));
  IGPU_CUDA(ctx, up(f->pod_off, c->pod_entry_off, (P + 1) * 4));
  IGPU_CUDA(ctx, up(f->entries, c->entries, E * 16));
  igpu_mark(ctx, 1);
  ctx->timing.h2d_bytes = bytes;
  f->uploaded = true;
  f->rolled = false;
  return IGPU_OK;
}

int32_t igpu_fleet_rollup(igpu_ctx* ctx, igpu_fleet* f) {
  if (!ctx || !f) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_fleet_rollup: NULL argument");
  IGPU_REQUIRE(ctx, f->uploaded, IGPU_ERR_STATE, "igpu_fleet_rollup before igpu_fleet_upload");
  DeviceGuard g(ctx->device);
  const uint32_t R = f->R;
  igpu_mark(ctx, 2);
  IGPU_CUDA(ctx, cudaMemsetAsync(f->d_acc, 0, f->acc_bytes, ctx->stream));
  if (f->n_pods) {
    const int nt = R <= 8 ? 256 : 64;
    const size_t smem = (size_t)4 * R * nt * sizeof(long long);
    using PodKernel = void (*)(const PodParams);
    PodKernel kern = pod_rollup_kernel<0>;
    switch (R) {
      case 4: kern = pod_rollup_kernel<4>; break;
      case 5: kern = pod_rollup_kernel<5>; break;
      case 6: kern = pod_rollup_kernel<6>; break;
      case 7: kern = pod_rollup_kernel<7>; break;
      case 8: kern = pod_rollup_kernel<8>; break;
      default: break;
    }
    IGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    PodKernel tiled = nullptr;
    switch (R) {
      case 4: tiled = pod_rollup_tiled_kernel<4>; break;
      case 5: tiled = pod_rollup_tiled_kernel<5>; break;
      case 6: tiled = pod_rollup_tiled_kernel<6>; break;
      case 7: tiled = pod_rollup_tiled_kernel<7>; break;
      case 8: tiled = pod_rollup_tiled_kernel<8>; break;
      default: break;
    }
    PodKernel ep = nullptr;
    switch (R) {
      case 4: ep = pod_rollup_ep_kernel<4>; break;
      case 5: ep = pod_rollup_ep_kernel<5>; break;
      case 6: ep = pod_rollup_ep_kernel<6>; break;
      case 7: ep = pod_rollup_ep_kernel<7>; break;
      case 8: ep = pod_rollup_ep_kernel<8>; break;
      default: break;
    }
    if (ctx->sw.k1_legacy) tiled = nullptr, ep = nullptr;
    if (ctx->sw.k1_pod_parallel) ep = nullptr;  // the previous tiled kernel (thread per pod all the way), kept for comparison
    PodParams p;
    p.P = f->n_pods;
    p.N = f->n_nodes;
    p.R = R;
    p.E = (uint32_t)f->n_entries;
    p.want_lim = (f->options & IGPU_OPT_POD_LIMITS) ? 1 : 0;
    p.pod_node = f->pod_node;
    p.pod_phase = f->pod_phase;
    p.pod_off = f->pod_off;
    p.entries = reinterpret_cast<const int4*>(f->entries);
    p.pod_is_gpu = f->pod_is_gpu;
    p.pod_gpu_containers = f->pod_gpu_containers;
    p.pod_req_mask = f->pod_req_mask;
    p.pod_lim_mask = f->pod_lim_mask;
    p.pod_req = f->pod_req;
    p.pod_lim = f->pod_lim;
    p.node_pod_count = f->node_pod_count;
    p.node_inuse = f->node_inuse;
    p.node_req = f->node_req;
    p.node_lim = f->node_lim;
    p.totals = f->totals;
    const unsigned need = (f->n_pods + nt - 1) / nt;
    const unsigned grid = std::min<unsigned>(need, (unsigned)ctx->sm_count * 4u);
    igpu_kbegin(ctx, IGPU_K_POD_ROLLUP);
    if (ep) {
      const size_t region = std::max((size_t)K1_TP * (2 * R + 3) * sizeof(uint32_t), (size_t)2 * R * K1_TP * sizeof(long long));
      const size_t tsmem = 2 * sizeof(K1EStage) + region;
      IGPU_CUDA(ctx, cudaFuncSetAttribute(ep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem));
      const unsigned tiles = (f->n_pods + K1_TP - 1) / K1_TP;
      ep<<<std::min<unsigned>(tiles, (unsigned)ctx->sm_count * 6u), K1_TP, tsmem, ctx->stream>>>(p);
    } else if (tiled) {
      const size_t tsmem = 2 * sizeof(K1Stage) + (size_t)2 * R * K1_TP * sizeof(long long);
      IGPU_CUDA(ctx, cudaFuncSetAttribute(tiled, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem));
      const unsigned tiles = (f->n_pods + K1_TP - 1) / K1_TP;
      tiled<<<std::min<unsigned>(tiles, (unsigned)ctx->sm_count * 8u), K1_TP, tsmem, ctx->stream>>>(p);
    } else {
      kern<<<grid, nt, smem, ctx->stream>>>(p);
    }
    igpu_kend(ctx, IGPU_K_POD_ROLLUP);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
  }
  if (f->n_nodes) {
    NodeParams q;
    q.N = f->n_nodes;
    q.R = R;
    q.node_flags = f->node_flags;
    q.cap_i915 = f->cap_i915;
    q.cap_xe = f->cap_xe;
    q.alloc_i915 = f->alloc_i915;
    q.alloc_xe = f->alloc_xe;
    q.alloc_own = f->alloc_own;
    q.node_pod_count = f->node_pod_count;
    q.node_inuse = f->node_inuse;
    q.out_flags = f->node_out_flags;
    q.type = f->node_type;
    q.gpu_count = f->node_gpu_count;
    q.cap_total = f->node_cap_total;
    q.alloc_total = f->node_alloc_total;
    q.bar_pct = f->node_bar_pct;
    q.bar_level = f->node_bar_level;
    q.util_pct = f->node_util_pct;
    q.util_level = f->node_util_level;
    q.totals = f->totals;
    igpu_kbegin(ctx, IGPU_K_NODE_ROLLUP);
    node_rollup_kernel<<<(f->n_nodes + 255) / 256, 256, 0, ctx->stream>>>(q);
    igpu_kend(ctx, IGPU_K_NODE_ROLLUP);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
  }
  // the one exchange step of the path: fleet totals over the node-range shards
  int32_t st = igpu_allreduce_i64_sum(ctx, f->totals, IGPU_TOTALS_LEN);
  if (st != IGPU_OK) return st;
  igpu_mark(ctx, 3);
  f->rolled = true;
  return IGPU_OK;
}

int32_t igpu_fleet_download(igpu_ctx* ctx, igpu_fleet* f, const igpu_fleet_out* o) {
  if (!ctx || !f || !o) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_fleet_download: NULL argument");
  IGPU_REQUIRE(ctx, f->rolled, IGPU_ERR_STATE, "igpu_fleet_download before igpu_fleet_rollup");
  DeviceGuard g(ctx->device);
  const size_t N = f->n_nodes, P = f->n_pods, R = f->R;
  const bool lim = f->options & IGPU_OPT_POD_LIMITS;
  uint64_t bytes = 0;
  auto down = [&](void* dst, const void* src, size_t n) -> cudaError_t {
    if (!dst || !n) return cudaSuccess;
    bytes += n;
    return cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, ctx->stream);
  };
  igpu_mark(ctx, 4);
  const size_t acc_at = align_up(f->out_span, 256);
  const bool packed = acc_at + f->acc_bytes <= FLEET_PACK_MAX;
  if (packed) {
    // small fleet: the output arena and the accumulator arena come back as two copies; the columns are handed out
    // from pinned staging after the one synchronisation
    if (!f->h_out) {
      IGPU_CUDA(ctx, cudaMallocHost(&f->h_out, FLEET_PACK_MAX));
      f->h_out_cap = FLEET_PACK_MAX;
    }
    if (f->out_span) IGPU_CUDA(ctx, cudaMemcpyAsync(f->h_out, f->d_out, f->out_span, cudaMemcpyDeviceToHost, ctx->stream));
    IGPU_CUDA(ctx, cudaMemcpyAsync(f->h_out + acc_at, f->d_acc, f->acc_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    igpu_mark(ctx, 5);
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    f->stage_busy = false;
    auto take_out = [&](void* dst, const void* dsrc, size_t n) {
      if (dst && n) memcpy(dst, f->h_out + (reinterpret_cast<const char*>(dsrc) - f->d_out), n), bytes += n;
    };
    auto take_acc = [&](void* dst, const void* dsrc, size_t n) {
      if (dst && n) memcpy(dst, f->h_out + acc_at + (reinterpret_cast<const char*>(dsrc) - f->d_acc), n), bytes += n;
    };
    take_out(o->pod_is_gpu, f->pod_is_gpu, P);
    take_out(o->pod_gpu_containers, f->pod_gpu_containers, P);
    take_out(o->pod_req_mask, f->pod_req_mask, P * 4);
    take_out(o->pod_req, f->pod_req, R * P * 8);
    if (lim) {
      take_out(o->pod_lim_mask, f->pod_lim_mask, P * 4);
      take_out(o->pod_lim, f->pod_lim, R * P * 8);
    }
    take_out(o->node_out_flags, f->node_out_flags, N);
    take_out(o->node_type, f->node_type, N);
    take_out(o->node_gpu_count, f->node_gpu_count, N * 8);
    take_out(o->node_cap_total, f->node_cap_total, N * 8);
    take_out(o->node_alloc_total, f->node_alloc_total, N * 8);
    take_acc(o->node_pod_count, f->node_pod_count, N * 4);
    take_out(o->node_bar_pct, f->node_bar_pct, N * 8);
    take_out(o->node_bar_level, f->node_bar_level, N);
    take_acc(o->node_inuse, f->node_inuse, N * 8);
    take_out(o->node_util_pct, f->node_util_pct, N * 8);
    take_out(o->node_util_level, f->node_util_level, N);
    take_acc(o->node_req, f->node_req, R * N * 8);
    if (lim) take_acc(o->node_lim, f->node_lim, R * N * 8);
    take_acc(ctx->h_totals, f->totals, IGPU_TOTALS_LEN * 8);
  } else {
  IGPU_CUDA(ctx, down(o->pod_is_gpu, f->pod_is_gpu, P));
  IGPU_CUDA(ctx, down(o->pod_gpu_containers, f->pod_gpu_containers, P));
  IGPU_CUDA(ctx, down(o->pod_req_mask, f->pod_req_mask, P * 4));
  IGPU_CUDA(ctx, down(o->pod_req, f->pod_req, R * P * 8));
  if (lim) {
    IGPU_CUDA(ctx, down(o->pod_lim_mask, f->pod_lim_mask, P * 4));
    IGPU_CUDA(ctx, down(o->pod_lim, f->pod_lim, R * P * 8));
  }
  IGPU_CUDA(ctx, down(o->node_out_flags, f->node_out_flags, N));
  IGPU_CUDA(ctx, down(o->node_type, f->node_type, N));
  IGPU_CUDA(ctx, down(o->node_gpu_count, f->node_gpu_count, N * 8));
  IGPU_CUDA(ctx, down(o->node_cap_total, f->node_cap_total, N * 8));
  IGPU_CUDA(ctx, down(o->node_alloc_total, f->node_alloc_total, N * 8));
  IGPU_CUDA(ctx, down(o->node_pod_count, f->node_pod_count, N * 4));
  IGPU_CUDA(ctx, down(o->node_bar_pct, f->node_bar_pct, N * 8));
  IGPU_CUDA(ctx, down(o->node_bar_level, f->node_bar_level, N));
  IGPU_CUDA(ctx, down(o->node_inuse, f->node_inuse, N * 8));
  IGPU_CUDA(ctx, down(o->node_util_pct, f->node_util_pct, N * 8));
  IGPU_CUDA(ctx, down(o->node_util_level, f->node_util_level, N));
  IGPU_CUDA(ctx, down(o->node_req, f->node_req, R * N * 8));
  if (lim) IGPU_CUDA(ctx, down(o->node_lim, f->node_lim, R * N * 8));
  IGPU_CUDA(ctx, cudaMemcpyAsync(ctx->h_totals, f->totals, IGPU_TOTALS_LEN * 8, cudaMemcpyDeviceToHost, ctx->stream));
  bytes += IGPU_TOTALS_LEN * 8;
  igpu_mark(ctx, 5);
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  f->stage_busy = false;
  }
  ctx->timing.d2h_bytes = bytes;
  // OverviewPage.tsx:115-116 on the reduced totals: Math.round(allocated / capacity * 100), 0 when the
  // capacity is not > 0 (which includes a NaN-poisoned capacity).
  int64_t* t = ctx->h_totals;
  const bool cap_ok = t[IGPU_T_CAP_NAN] == 0 && t[IGPU_T_CAPACITY] > 0;
  if (cap_ok) {
    const double x = (double)t[IGPU_T_ALLOCATED] / (double)t[IGPU_T_CAPACITY] * 100.0;
    const double fl = floor(x);
    const double r = (x - fl) >= 0.5 ? fl + 1.0 : fl;
    t[IGPU_T_UTIL_PCT] = r >= 9.2e18 ? INT64_MAX : (r <= -9.2e18 ? INT64_MIN : (int64_t)r);
  } else {
    t[IGPU_T_UTIL_PCT] = 0;
  }
  if (o->fleet_totals) memcpy(o->fleet_totals, t, IGPU_TOTALS_LEN * 8);
  if (t[IGPU_T_BAD_INPUT] != 0)
    return igpu_fail(ctx, IGPU_ERR_ARG,
                     "%lld pods / entries break the column invariants (pod_entry_off ascending and <= n_entries, entry.res < "
                     "n_resources, pod_node < n_nodes): they were skipped, the outputs are not meaningful",
                     (long long)t[IGPU_T_BAD_INPUT]);
  return IGPU_OK;
}

int32_t igpu_alloc_rollup(igpu_ctx* ctx, const igpu_fleet_cols* c, const igpu_fleet_out* o) {
  if (!ctx || !c || !o) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_alloc_rollup: NULL argument");
  IGPU_REQUIRE(ctx, c->n_resources >= 4 && c->n_resources <= IGPU_MAX_RESOURCES, IGPU_ERR_RANGE, "n_resources %u outside [4,%d]",
               c->n_resources, IGPU_MAX_RESOURCES);
  DeviceGuard g(ctx->device);
  igpu_fleet* f = ctx->cached_fleet;
  if (!f || f->max_nodes < c->n_nodes || f->max_pods < c->n_pods || f->max_entries < c->n_entries || f->R != c->n_resources) {
    if (f) igpu_fleet_destroy(ctx, f);
    ctx->cached_fleet = nullptr;
    const uint32_t gn = std::max<uint32_t>(c->n_nodes, 64), gp = std::max<uint32_t>(c->n_pods, 256);
    const uint64_t ge = std::max<uint64_t>(c->n_entries, 1024);
    int32_t st = igpu_fleet_create(ctx, gn, gp, ge, c->n_resources, c->options, &f);
    if (st != IGPU_OK) return st;
    ctx->cached_fleet = f;
  }
  igpu_mark(ctx, 6);
  int32_t st = igpu_fleet_upload(ctx, f, c);
  if (st == IGPU_OK) st = igpu_fleet_rollup(ctx, f);
  if (st == IGPU_OK) st = igpu_fleet_download(ctx, f, o);
  igpu_mark(ctx, 7);
  if (st == IGPU_OK) IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return st;
}

int32_t igpu_fleet_container_rows(igpu_ctx* ctx, igpu_fleet* f, igpu_container_row* host_rows, uint64_t capacity, uint64_t* n_rows) {
  if (!ctx || !f || !n_rows || (capacity && !host_rows)) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_fleet_container_rows: NULL argument");
  IGPU_REQUIRE(ctx, f->rolled, IGPU_ERR_STATE, "igpu_fleet_container_rows before igpu_fleet_rollup");
  *n_rows = 0;
  const uint32_t P = f->n_pods;
  if (!P) return IGPU_OK;
  DeviceGuard g(ctx->device);
  const uint32_t nb = (P + CR_BLOCK - 1) / CR_BLOCK;
  if (f->cont_block_cap < (size_t)nb + 1) {
    cudaFree(f->d_cont_block);
    f->d_cont_block = nullptr, f->cont_block_cap = 0;
    IGPU_CUDA(ctx, cudaMalloc(&f->d_cont_block, ((size_t)nb + 1) * sizeof(uint32_t)));
    f->cont_block_cap = (size_t)nb + 1;
  }
  cont_block_sums_kernel<<<nb, CR_THREADS, 0, ctx->stream>>>(f->pod_gpu_containers, P, f->d_cont_block);
  cont_scan_blocks_kernel<<<1, CR_THREADS, 0, ctx->stream>>>(f->d_cont_block, nb);
  IGPU_CUDA(ctx, cudaGetLastError());
  igpu_count_launch(ctx, 2);
  uint32_t total = 0;
  IGPU_CUDA(ctx, cudaMemcpyAsync(&total, f->d_cont_block + nb, sizeof total, cudaMemcpyDeviceToHost, ctx->stream));
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *n_rows = total;
  const uint64_t want = std::min<uint64_t>(total, capacity);
  if (!want) return IGPU_OK;
  if (f->cont_rows_cap < total) {
    cudaFree(f->d_cont_rows);
    f->d_cont_rows = nullptr, f->cont_rows_cap = 0;
    IGPU_CUDA(ctx, cudaMalloc(&f->d_cont_rows, (size_t)total * sizeof(igpu_container_row)));
    f->cont_rows_cap = total;
  }
  cont_rows_kernel<<<nb, CR_THREADS, 0, ctx->stream>>>(f->pod_gpu_containers, f->pod_off, reinterpret_cast<const int4*>(f->entries), P,
                                                        (uint32_t)f->n_entries, f->R, f->d_cont_block, f->d_cont_rows, total);
  IGPU_CUDA(ctx, cudaGetLastError());
  igpu_count_launch(ctx);
  IGPU_CUDA(ctx, cudaMemcpyAsync(host_rows, f->d_cont_rows, (size_t)want * sizeof(igpu_container_row), cudaMemcpyDeviceToHost, ctx->stream));
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return IGPU_OK;
}

}  // extern "C"
