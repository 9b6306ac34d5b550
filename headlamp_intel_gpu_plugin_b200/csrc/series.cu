// series.cu — device-resident power series slab, synthetic generator, and K3: windowed
// min / mean / max downsampling.
//
// No reference counterpart: src/api/metrics.ts only issues instant queries (:67-75, :101-116) and
// the one windowed quantity, rate(...[5m]) (:106), is evaluated inside Prometheus.  Semantics are
// PromQL's min/avg/max_over_time over non-overlapping windows; NaN = missing.
//
// HBM layout: x[series][time] float32, series grouped by node; outputs are three planes
// [series][ceil(T/W)] float32.  Algorithmic traffic: 4 B per sample in, 12 B per window out.
//
// Fast path (window = PW * L with PW a power of two <= 32 and L odd <= 97 — W itself may be odd —, T % 4 == 0): a
// persistent CTA streams tiles of 256*K "chunks" (L consecutive samples each) of one series row
// into a 4-stage shared-memory ring with 1-D bulk copies (cp.async.bulk, mbarrier completion).
// Thread t owns chunk t of the tile: it reads shared memory at stride L words (odd => no bank
// conflicts), folds min / max / float64 sum / count, and the PW threads of a window combine with
// log2(PW) shuffle steps.  Every other shape takes the generic warp-per-window kernel.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

#include "series.cuh"

namespace {

// ---- synthetic generator: integer hash + individually rounded f32 ops (matches synth.py) ------
__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256) series_gen_kernel(float* __restrict__ out, uint32_t n_series, uint32_t n_time, uint32_t pitch,
                                                         uint64_t seed, uint64_t global_first, uint32_t nan_per_million) {
  const uint64_t total = (uint64_t)n_series * n_time;
  const uint64_t S_PARAM = 0x11ull << 56, S_SAMPLE = 0x12ull << 56, S_NAN = 0x13ull << 56;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t s = (uint32_t)(i / n_time), t = (uint32_t)(i - (uint64_t)s * n_time);
    const uint64_t gs = global_first + s;
    const uint64_t rp = splitmix64(seed ^ S_PARAM ^ gs);
    const float base = __fadd_rn(40.0f, __fmul_rn(__uint2float_rn((uint32_t)(rp & 0xFFFF)), 180.0f / 65536.0f));
    const uint32_t k = (uint32_t)((rp >> 16) & 0xFF) % 3u;
    const float tdp = k == 0 ? 150.0f : (k == 1 ? 250.0f : 300.0f);
    const uint32_t phase = (uint32_t)((rp >> 24) % 86400u);
    const uint64_t idx = gs * (uint64_t)n_time + t;
    const uint64_t r = splitmix64(seed ^ S_SAMPLE ^ idx);
    float v;
    bool knocked = false;
    if (nan_per_million) {
      const uint64_t rn = splitmix64(seed ^ S_NAN ^ idx);
      knocked = (uint32_t)(rn % 1000000u) < nan_per_million;
    }
    const float xi = __fsub_rn(__fmul_rn(__uint2float_rn((uint32_t)(r >> 40)), 1.0f / 8388608.0f), 1.0f);
    const uint32_t ph = (t + phase) % 86400u;
    const float x = __fdiv_rn(__uint2float_rn(ph), 86400.0f);
    const float u = __fsub_rn(1.0f, x);
    const float y = __fmul_rn(x, u);
    const float wave = __fsub_rn(__fmul_rn(8.0f, y), 1.0f);
    const float a = __fmul_rn(0.25f, base), b = __fmul_rn(0.05f, base);
    v = __fadd_rn(base, __fmul_rn(a, wave));
    v = __fadd_rn(v, __fmul_rn(b, xi));
    v = v < 0.0f ? 0.0f : v;
    v = v > tdp ? tdp : v;
    out[(size_t)s * pitch + t] = knocked ? __int_as_float(0x7fc00000) : v;
  }
}

// ---- K3 fast path --------------------------------------------------------------------------------
constexpr int WS_STAGES = 4;

struct WsParams {
  const float* __restrict__ x;
  float* __restrict__ omin;
  float* __restrict__ omean;
  float* __restrict__ omax;
  uint32_t n_series, n_time, window, n_win;
  uint32_t PW, log2PW, K;      // window = PW * L; a tile has K rounds of 256 chunks
  uint32_t L, stages;          // L: the template parameter, or the run-time chunk length of the <0> instantiation; ring depth <= WS_STAGES
  uint32_t tiles_per_row;
  uint32_t n_tiles;            // < 2^32 (host checks)
};

// LT > 0: chunk length known at compile time (fully unrolled, L <= 31).  LT == 0: chunk length p.L at run time
// (odd, up to 97: W = 300 = 4 x 75, 1440 = 32 x 45, ...), same tile / ring / shuffle structure with a shallower ring.
// V4: the unit is a float4 instead of a float — window = PW * L float4s, each thread folds L consecutive float4s with
// 128-bit shared-memory loads.  Takes the windows whose float count has no usable odd factor (64, 128, 256, ...):
// L = 1 is conflict-free, L = 2 / 4 / 8 cost 2 / 4 / 8 shared-memory wavefronts per load, still far above the HBM rate.
template <int LT, bool V4>
__global__ void __launch_bounds__(WS_THREADS) window_stats_staged_kernel(const WsParams p) {
  constexpr uint32_t ELEM = V4 ? 4u : 1u;  // floats per unit
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full[WS_STAGES];
  float* buf = reinterpret_cast<float*>(smem_raw);
  const uint32_t tid = threadIdx.x;
  const uint32_t L = LT ? (uint32_t)LT : p.L;
  const uint32_t NS = LT ? (uint32_t)WS_STAGES : p.stages;
  const uint32_t tile_chunks = WS_THREADS * p.K;
  const uint32_t LF = L * ELEM;  // floats per chunk
  const uint32_t tile_floats = tile_chunks * LF;
  if (tid == 0) {
    for (int s = 0; s < WS_STAGES; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  const uint32_t first = blockIdx.x, step = gridDim.x, tpr = p.tiles_per_row;
  const uint32_t my_count = p.n_tiles > first ? (p.n_tiles - first + step - 1) / step : 0;
  const uint32_t d_row = step / tpr, d_tr = step % tpr;
  TileCursor prod{first / tpr, first % tpr};  // next tile to issue (thread 0)
  TileCursor cons = prod;                      // tile being consumed

  auto issue = [&](uint32_t i) {  // thread 0 only: start the bulk copy of this CTA's i-th tile
    const uint32_t start = prod.tr * tile_floats;
    const uint32_t len = min(tile_floats, p.n_time - start);
    const uint32_t stage = i % NS;
    mbar_expect_tx(&full[stage], len * 4u);
    bulk_g2s(buf + (size_t)stage * tile_floats, p.x + (size_t)prod.row * p.n_time + start, len * 4u, &full[stage]);
    prod.advance(d_row, d_tr, tpr);
  };
  if (tid == 0)
    for (uint32_t i = 0; i < my_count && i < NS; ++i) issue(i);

  const uint32_t PW = p.PW;
  for (uint32_t i = 0; i < my_count; ++i) {
    const uint32_t stage = i % NS;
    mbar_wait(&full[stage], (i / NS) & 1u);
    const float* sbuf = buf + (size_t)stage * tile_floats;
    const uint32_t tile_sample0 = cons.tr * tile_floats;  // first sample of the tile within its row
    for (uint32_t k = 0; k < p.K; ++k) {
      const uint32_t chunk_in_tile = k * WS_THREADS + tid;
      const uint32_t sample0 = tile_sample0 + chunk_in_tile * LF;
      const float* src = sbuf + chunk_in_tile * LF;
      Acc a{INFINITY, -INFINITY, 0.0, LF};
      bool slow = sample0 + LF > p.n_time;  // chunk is cut by the end of the row
      if (!slow) {
        // NaN-free fast path: 1 LDS + min + max + cvt + add per sample, no per-sample predicate.  A NaN
        // sample (or +inf with -inf) makes the sum NaN, which sends just this chunk to the exact path.
        if (V4) {
          const float4* src4 = reinterpret_cast<const float4*>(src);
          // power-of-two chunk lengths put the lanes of a quarter-warp on the same banks; each lane starts its
          // (order-free) fold at a different float4 so that the eight 16-byte accesses of a wavefront spread out
          const bool pow2 = (L & (L - 1)) == 0;
          const uint32_t rot = pow2 ? (L >= 8 ? tid : (tid * L) >> 3) : 0u;
#pragma unroll 4
          for (uint32_t j = 0; j < L; ++j) {
            const float4 v = src4[pow2 ? ((j + rot) & (L - 1)) : j];
            a.mn = fminf(fminf(a.mn, v.x), fminf(v.y, fminf(v.z, v.w)));
            a.mx = fmaxf(fmaxf(a.mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
            a.sum += (double)v.x;
            a.sum += (double)v.y;
            a.sum += (double)v.z;
            a.sum += (double)v.w;
          }
        } else if (LT) {
#pragma unroll
          for (int j = 0; j < LT; ++j) {
            const float v = src[j];
            a.mn = fminf(a.mn, v);
            a.mx = fmaxf(a.mx, v);
            a.sum += (double)v;
          }
        } else {
#pragma unroll 5
          for (uint32_t j = 0; j < L; ++j) {
            const float v = src[j];
            a.mn = fminf(a.mn, v);
            a.mx = fmaxf(a.mx, v);
            a.sum += (double)v;
          }
        }
        if (a.sum != a.sum) {
          // a NaN (or +inf with -inf) in the chunk.  fminf / fmaxf returned the operand that is not NaN, so min and max are
          // right as they are; only the sum and the count are taken again (half the work of a full NaN-aware fold: slabs
          // with 0.1 % of the samples missing send most warps here)
          a.sum = 0.0;
          a.cnt = 0u;
#pragma unroll 4
          for (uint32_t j = 0; j < LF; ++j) {
            const float v = src[j];
            a.sum += v == v ? (double)v : 0.0;
            a.cnt += v == v;
          }
        }
      }
      if (slow) {  // the chunk is cut by the end of the row
        const int valid = sample0 < p.n_time ? (int)min(LF, p.n_time - sample0) : 0;
        a = Acc{INFINITY, -INFINITY, 0.0, 0u};
        for (int j = 0; j < valid; ++j) acc_sample(a, src[j]);
      }
      for (uint32_t o = 1; o < PW; o <<= 1) {  // the PW chunks of a window sit in adjacent lanes
        a.mn = fminf(a.mn, __shfl_xor_sync(0xffffffffu, a.mn, o));
        a.mx = fmaxf(a.mx, __shfl_xor_sync(0xffffffffu, a.mx, o));
        a.sum += __shfl_xor_sync(0xffffffffu, a.sum, o);
        a.cnt += __shfl_xor_sync(0xffffffffu, a.cnt, o);
      }
      const uint32_t chunk = cons.tr * tile_chunks + chunk_in_tile;  // chunk index within the row
      const uint32_t w = chunk >> p.log2PW;
      if ((chunk & (PW - 1)) == 0 && w < p.n_win) {
        const size_t o = (size_t)cons.row * p.n_win + w;
        const float nanv = __int_as_float(0x7fc00000);
        p.omin[o] = a.cnt ? a.mn : nanv;
        p.omax[o] = a.cnt ? a.mx : nanv;
        p.omean[o] = a.cnt ? (float)(a.sum / (double)a.cnt) : nanv;
      }
    }
    cons.advance(d_row, d_tr, tpr);
    __syncthreads();  // every thread is done with this stage: it may be refilled
    if (tid == 0 && i + NS < my_count) issue(i + NS);
  }
}

using WsKernel = void (*)(const WsParams);
template <int L>
WsKernel ws_pick() {
  return window_stats_staged_kernel<L, false>;
}
static WsKernel ws_kernel_for(uint32_t L) {
  switch (L) {
    case 1: return ws_pick<1>();
    case 3: return ws_pick<3>();
    case 5: return ws_pick<5>();
    case 7: return ws_pick<7>();
    case 9: return ws_pick<9>();
    case 11: return ws_pick<11>();
    case 13: return ws_pick<13>();
    case 15: return ws_pick<15>();
    case 17: return ws_pick<17>();
    case 19: return ws_pick<19>();
    case 21: return ws_pick<21>();
    case 23: return ws_pick<23>();
    case 25: return ws_pick<25>();
    case 27: return ws_pick<27>();
    case 29: return ws_pick<29>();
    case 31: return ws_pick<31>();
    default: return ws_pick<0>();  // run-time chunk length
  }
}

// ---- K3 generic path: one warp per window, lanes stride through it ------------------------------
// VEC = true when windows start 16-byte aligned (W % 4 == 0 and T % 4 == 0): 128-bit loads, up to four
// in flight per lane; else scalar loads (still coalesced), four in flight.
template <bool VEC>
__global__ void __launch_bounds__(256) window_stats_generic_kernel(const float* __restrict__ x, float* __restrict__ omin,
                                                                   float* __restrict__ omean, float* __restrict__ omax,
                                                                   uint32_t n_series, uint32_t n_time, uint32_t window, uint32_t n_win) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint64_t total = (uint64_t)n_series * n_win;
  for (uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; gw < total; gw += warps) {
    const uint64_t row = gw / n_win;
    const uint32_t w = (uint32_t)(gw - row * n_win);
    const uint32_t t0 = w * window, t1 = min(n_time, t0 + window);
    const float* src = x + row * n_time;
    Acc a{INFINITY, -INFINITY, 0.0, 0u};
    if (VEC) {
      const float4* src4 = reinterpret_cast<const float4*>(src + t0);
      const uint32_t nvec = (t1 - t0) >> 2;  // (t1 - t0) % 4 == 0 because W % 4 == 0 and T % 4 == 0
      for (uint32_t i0 = 0; i0 < nvec; i0 += 128) {
        float4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t i = i0 + j * 32 + lane;
          v[j] = i < nvec ? __ldg(src4 + i) : make_float4(NAN, NAN, NAN, NAN);  // NaN = ignored by the fold
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) acc_sample(a, v[j].x), acc_sample(a, v[j].y), acc_sample(a, v[j].z), acc_sample(a, v[j].w);
      }
    } else {
      for (uint32_t tb = t0; tb < t1; tb += 128) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t t = tb + j * 32 + lane;
          v[j] = t < t1 ? __ldg(src + t) : NAN;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) acc_sample(a, v[j]);
      }
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
      omin[gw] = a.cnt ? a.mn : nanv;
      omax[gw] = a.cnt ? a.mx : nanv;
      omean[gw] = a.cnt ? (float)(a.sum / (double)a.cnt) : nanv;
    }
  }
}

}  // namespace

WsPlan igpu_ws_plan(uint32_t window, uint32_t n_time, uint32_t n_series) {
  WsPlan p{};
  p.n_win = (n_time + window - 1) / window;
  p.PW = 1;
  p.log2PW = 0;
  while (p.PW < 32 && (window % (p.PW * 2)) == 0) p.PW *= 2, ++p.log2PW;
  p.L = window / p.PW;
  p.vec4 = false;
  if (window % 4 == 0 && (!(p.L & 1) || p.L > WS_MAX_L)) {
    // no usable odd chunk in floats: try float4 units, window = PW * L float4s with L <= WS_MAX_L4
    const uint32_t v = window / 4;
    uint32_t pw = 1, lg = 0;
    while (pw < 32 && (v % (pw * 2)) == 0) pw *= 2, ++lg;
    // a pure power of two: keep 16 float4s per thread (no shuffles up to W = 64, few beyond) rather than PW = 32
    if ((v & (v - 1)) == 0)
      while (pw > 1 && v / pw < 16) pw /= 2, --lg;
    if (v / pw <= WS_MAX_L4) p.PW = pw, p.log2PW = lg, p.L = v / pw, p.vec4 = true;
  }
  p.K = std::max<uint32_t>(1, 16 / (p.L * (p.vec4 ? 4 : 1)));
  const uint32_t chunks_per_row = p.n_win * p.PW;
  p.tiles_per_row = (chunks_per_row + WS_THREADS * p.K - 1) / (WS_THREADS * p.K);
  p.n_tiles = (uint64_t)p.tiles_per_row * n_series;
  // (rows must start 16-byte aligned for the bulk copies; the window itself need not be a multiple of 4 samples: a tile is
  // 256 K chunks, whatever the chunk length — W = 30 is 2 x 15, W = 10 is 2 x 5, W = 15 is 1 x 15)
  p.fast = (n_time % 4 == 0) && (p.vec4 || ((p.L & 1) && p.L <= WS_MAX_L)) && p.n_tiles < 0xFFFFFFFFull;
  return p;
}

int32_t igpu_series_ensure_window_buffers(igpu_ctx* ctx, igpu_series* s, uint32_t window) {
  const uint32_t n_win = (s->n_time_user + window - 1) / window;
  const size_t need = std::max<size_t>((size_t)s->n_series * n_win, 4);
  if (need > s->w_capacity) {
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (s->d_wcnt) cudaFree(s->d_wcnt);
  if (s->d_wmin) cudaFree(s->d_wmin);
    if (s->d_wmean) cudaFree(s->d_wmean);
    if (s->d_wmax) cudaFree(s->d_wmax);
    s->d_wmin = s->d_wmean = s->d_wmax = nullptr;
    s->w_capacity = 0;
    IGPU_CUDA(ctx, cudaMalloc(&s->d_wmin, need * sizeof(float)));
    IGPU_CUDA(ctx, cudaMalloc(&s->d_wmean, need * sizeof(float)));
    IGPU_CUDA(ctx, cudaMalloc(&s->d_wmax, need * sizeof(float)));
    s->w_capacity = need;
  }
  if (s->keep_counts && need > s->wcnt_capacity) {
    IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (s->d_wcnt) cudaFree(s->d_wcnt);
    s->d_wcnt = nullptr;
    s->wcnt_capacity = 0;
    IGPU_CUDA(ctx, cudaMalloc(&s->d_wcnt, need * sizeof(uint16_t)));
    s->wcnt_capacity = need;
  }
  s->window = window;
  s->n_win = n_win;
  s->wcnt_valid = false;  // the planes are about to be rewritten; the kernel that also writes counts says so afterwards
  return IGPU_OK;
}

extern "C" {

int32_t igpu_series_create(igpu_ctx* ctx, uint32_t n_series, uint32_t n_time, uint32_t n_nodes, const uint32_t* node_series_off,
                           igpu_series** out) {
  if (!ctx || !out || !node_series_off) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_create: NULL argument");
  *out = nullptr;
  IGPU_REQUIRE(ctx, n_time > 0, IGPU_ERR_ARG, "n_time must be positive");
  IGPU_REQUIRE(ctx, n_time < 0x7FFFFFFCu, IGPU_ERR_RANGE, "rows of %u samples are too long (the pitch is rounded up to a multiple of 4)", n_time);
  IGPU_REQUIRE(ctx, node_series_off[0] == 0 && node_series_off[n_nodes] == n_series, IGPU_ERR_ARG,
               "node_series_off must start at 0 and end at n_series");
  for (uint32_t i = 0; i < n_nodes; ++i) {
    IGPU_REQUIRE(ctx, node_series_off[i] <= node_series_off[i + 1], IGPU_ERR_ARG, "node_series_off must be non-decreasing");
    IGPU_REQUIRE(ctx, (uint64_t)(node_series_off[i + 1] - node_series_off[i]) * ((n_time + 3ull) & ~3ull) < 0xFFFFFFFFull, IGPU_ERR_RANGE,
                 "node %u holds 2^32 samples or more", i);
  }
  DeviceGuard g(ctx->device);
  igpu_series* s = new igpu_series();
  s->n_series = n_series;
  s->n_time_user = n_time;
  s->n_time = (n_time + 3u) & ~3u;  // row pitch
  s->n_nodes = n_nodes;
  s->h_node_off = new uint32_t[n_nodes + 1];
  memcpy(s->h_node_off, node_series_off, (n_nodes + 1) * sizeof(uint32_t));
  const size_t bytes = std::max<size_t>((size_t)n_series * s->n_time * sizeof(float), 16);
  cudaError_t e1 = cudaMalloc(&s->d_x, bytes), e2 = cudaMalloc(&s->d_node_off, (n_nodes + 1) * sizeof(uint32_t));
  if (e1 != cudaSuccess || e2 != cudaSuccess) {
    cudaGetLastError();
    igpu_series_destroy(ctx, s);
    return igpu_fail(ctx, IGPU_ERR_OOM, "device allocation of %zu series bytes failed", bytes);
  }
  IGPU_CUDA(ctx, cudaMemcpyAsync(s->d_node_off, s->h_node_off, (n_nodes + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
  if (s->n_time != n_time && n_series)  // the padding: all-ones words are NaNs, and nothing ever writes there again
    IGPU_CUDA(ctx, cudaMemset2DAsync(s->d_x + n_time, (size_t)s->n_time * sizeof(float), 0xFF, (s->n_time - n_time) * sizeof(float), n_series,
                                     ctx->stream));
  *out = s;
  return IGPU_OK;
}

int32_t igpu_series_destroy(igpu_ctx* ctx, igpu_series* s) {
  if (!ctx || !s) return IGPU_OK;
  DeviceGuard g(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  igpu_quantile_ws_free(s);
  if (s->d_x) cudaFree(s->d_x);
  if (s->d_node_off) cudaFree(s->d_node_off);
  if (s->d_wmin) cudaFree(s->d_wmin);
  if (s->d_wmean) cudaFree(s->d_wmean);
  if (s->d_wmax) cudaFree(s->d_wmax);
  if (s->d_wcnt) cudaFree(s->d_wcnt);
  if (s->d_fs) cudaFree(s->d_fs);
  delete[] s->h_node_off;
  delete s;
  return IGPU_OK;
}

int32_t igpu_series_upload(igpu_ctx* ctx, igpu_series* s, uint32_t first, uint32_t count, const float* host) {
  if (!ctx || !s || (!host && count)) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_upload: NULL argument");
  IGPU_REQUIRE(ctx, (uint64_t)first + count <= s->n_series, IGPU_ERR_ARG, "rows [%u,%u) outside the slab of %u series", first,
               first + count, s->n_series);
  DeviceGuard g(ctx->device);
  const size_t row = (size_t)s->n_time_user * sizeof(float), bytes = (size_t)count * row;
  igpu_mark(ctx, 0);
  if (bytes) {
    if (s->n_time == s->n_time_user)
      IGPU_CUDA(ctx, cudaMemcpyAsync(s->d_x + (size_t)first * s->n_time, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    else
      IGPU_CUDA(ctx, cudaMemcpy2DAsync(s->d_x + (size_t)first * s->n_time, (size_t)s->n_time * sizeof(float), host, row, row, count,
                                       cudaMemcpyHostToDevice, ctx->stream));
  }
  igpu_mark(ctx, 1);
  ctx->timing.h2d_bytes = bytes;
  s->wcnt_valid = false;  // the slab changed
  return IGPU_OK;
}

int32_t igpu_series_download(igpu_ctx* ctx, igpu_series* s, uint32_t first, uint32_t count, float* host) {
  if (!ctx || !s || (!host && count)) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_download: NULL argument");
  IGPU_REQUIRE(ctx, (uint64_t)first + count <= s->n_series, IGPU_ERR_ARG, "rows [%u,%u) outside the slab of %u series", first,
               first + count, s->n_series);
  DeviceGuard g(ctx->device);
  const size_t row = (size_t)s->n_time_user * sizeof(float), bytes = (size_t)count * row;
  if (bytes) {
    if (s->n_time == s->n_time_user)
      IGPU_CUDA(ctx, cudaMemcpyAsync(host, s->d_x + (size_t)first * s->n_time, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    else
      IGPU_CUDA(ctx, cudaMemcpy2DAsync(host, row, s->d_x + (size_t)first * s->n_time, (size_t)s->n_time * sizeof(float), row, count,
                                       cudaMemcpyDeviceToHost, ctx->stream));
  }
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return IGPU_OK;
}

int32_t igpu_series_generate(igpu_ctx* ctx, igpu_series* s, uint64_t seed, uint64_t global_first, uint32_t nan_per_million) {
  if (!ctx || !s) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_generate: NULL argument");
  DeviceGuard g(ctx->device);
  if (s->n_series) {
    series_gen_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(s->d_x, s->n_series, s->n_time_user, s->n_time, seed, global_first,
                                                                  nan_per_million);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
  }
  s->wcnt_valid = false;  // the slab changed
  return IGPU_OK;
}

int32_t igpu_series_keep_window_counts(igpu_ctx* ctx, igpu_series* s, int32_t on) {
  if (!ctx || !s) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_keep_window_counts: NULL argument");
  s->keep_counts = on != 0;
  if (!s->keep_counts) s->wcnt_valid = false;
  return IGPU_OK;
}

int32_t igpu_series_window_stats(igpu_ctx* ctx, igpu_series* s, uint32_t window, float* host_min, float* host_mean, float* host_max) {
  if (!ctx || !s) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_window_stats: NULL argument");
  IGPU_REQUIRE(ctx, window > 0, IGPU_ERR_ARG, "window must be positive");
  DeviceGuard g(ctx->device);
  int32_t bst = igpu_series_ensure_window_buffers(ctx, s, window);
  if (bst != IGPU_OK) return bst;
  const uint32_t n_win = s->n_win;
  igpu_mark(ctx, 2);
  if (s->n_series) {
    igpu_kbegin(ctx, IGPU_K_WINDOW_STATS);
    const WsPlan plan = igpu_ws_plan(window, s->n_time, s->n_series);
    const uint32_t PW = plan.PW, log2PW = plan.log2PW, L = plan.L, K = plan.K, tiles_per_row = plan.tiles_per_row;
    const uint64_t n_tiles = plan.n_tiles;
    const bool fast = plan.fast && !ctx->sw.ws_force_generic;
    if (fast) {
      WsParams p;
      p.x = s->d_x;
      p.omin = s->d_wmin;
      p.omean = s->d_wmean;
      p.omax = s->d_wmax;
      p.n_series = s->n_series;
      p.n_time = s->n_time;
      p.window = window;
      p.n_win = n_win;
      p.PW = PW;
      p.log2PW = log2PW;
      p.K = K;
      p.tiles_per_row = tiles_per_row;
      p.n_tiles = (uint32_t)n_tiles;
      const uint32_t tile_floats = WS_THREADS * K * L * (plan.vec4 ? 4u : 1u);
      // chunk lengths above 31 floats use a run-time-L instantiation with as deep a ring as 200 KB allows (>= 2 stages)
      p.L = L;
      p.stages = (!plan.vec4 && L <= 31) ? WS_STAGES : std::min<uint32_t>(WS_STAGES, (200u * 1024u) / (tile_floats * 4u));
      const size_t smem = (size_t)p.stages * tile_floats * sizeof(float);
      WsKernel kern = plan.vec4 ? window_stats_staged_kernel<0, true> : ws_kernel_for(L);
      IGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      const unsigned per_sm = std::max<unsigned>(1, std::min<unsigned>(4, (unsigned)((220 * 1024) / (smem + 1024))));
      const unsigned grid = (unsigned)std::min<uint64_t>(n_tiles, (uint64_t)ctx->sm_count * per_sm);
      kern<<<grid, WS_THREADS, smem, ctx->stream>>>(p);
    } else {
      const uint64_t total = (uint64_t)s->n_series * n_win;
      const unsigned grid = (unsigned)std::min<uint64_t>((total + 7) / 8, (uint64_t)ctx->sm_count * 8);
      if (window % 4 == 0 && s->n_time % 4 == 0)
        window_stats_generic_kernel<true><<<grid, 256, 0, ctx->stream>>>(s->d_x, s->d_wmin, s->d_wmean, s->d_wmax, s->n_series,
                                                                         s->n_time, window, n_win);
      else
        window_stats_generic_kernel<false><<<grid, 256, 0, ctx->stream>>>(s->d_x, s->d_wmin, s->d_wmean, s->d_wmax, s->n_series,
                                                                          s->n_time, window, n_win);
    }
    igpu_kend(ctx, IGPU_K_WINDOW_STATS);
    IGPU_CUDA(ctx, cudaGetLastError());
    igpu_count_launch(ctx);
  }
  igpu_mark(ctx, 3);
  if (host_min || host_mean || host_max) return igpu_series_read_window_stats(ctx, s, 0, s->n_series, host_min, host_mean, host_max);
  return IGPU_OK;
}

int32_t igpu_series_read_window_stats(igpu_ctx* ctx, igpu_series* s, uint32_t first, uint32_t count, float* host_min,
                                      float* host_mean, float* host_max) {
  if (!ctx || !s) return igpu_fail(ctx, IGPU_ERR_ARG, "igpu_series_read_window_stats: NULL argument");
  IGPU_REQUIRE(ctx, s->window != 0, IGPU_ERR_STATE, "no window statistics computed yet");
  IGPU_REQUIRE(ctx, (uint64_t)first + count <= s->n_series, IGPU_ERR_ARG, "rows [%u,%u) outside the slab of %u series", first,
               first + count, s->n_series);
  DeviceGuard g(ctx->device);
  const size_t off = (size_t)first * s->n_win, bytes = (size_t)count * s->n_win * sizeof(float);
  uint64_t moved = 0;
  igpu_mark(ctx, 4);
  if (bytes) {
    if (host_min) {
      IGPU_CUDA(ctx, cudaMemcpyAsync(host_min, s->d_wmin + off, bytes, cudaMemcpyDeviceToHost, ctx->stream));
      moved += bytes;
    }
    if (host_mean) {
      IGPU_CUDA(ctx, cudaMemcpyAsync(host_mean, s->d_wmean + off, bytes, cudaMemcpyDeviceToHost, ctx->stream));
      moved += bytes;
    }
    if (host_max) {
      IGPU_CUDA(ctx, cudaMemcpyAsync(host_max, s->d_wmax + off, bytes, cudaMemcpyDeviceToHost, ctx->stream));
      moved += bytes;
    }
  }
  igpu_mark(ctx, 5);
  IGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->timing.d2h_bytes = moved;
  return IGPU_OK;
}

}  // extern "C"
