// series.cuh — device-resident power time-series slab shared by series.cu (K3) and quantile.cu (K4).
#pragma once
#include "common.cuh"

struct igpu_series {
  // n_time is the row PITCH every kernel sees: the caller's row length (n_time_user) rounded up to a multiple of 4 samples,
  // the samples past n_time_user being NaN for good — "no sample" to every kernel.  Rows then start 16-byte aligned whatever
  // the caller's length (a query_range of N steps returns N + 1 points), so the bulk-copy kernels take every slab.  Window
  // counts (n_win) and everything the API reports are about n_time_user.
  uint32_t n_series = 0, n_time = 0, n_time_user = 0, n_nodes = 0;
  float* d_x = nullptr;            // [n_series][n_time]
  uint32_t* d_node_off = nullptr;  // [n_nodes + 1] first series of each node
  uint32_t* h_node_off = nullptr;  // host copy
  // window statistics of the last igpu_series_window_stats call
  uint32_t window = 0, n_win = 0;
  float *d_wmin = nullptr, *d_wmean = nullptr, *d_wmax = nullptr;
  size_t w_capacity = 0;  // floats per plane
  // valid samples per (series, window), kept by the single-read rollup when asked to (igpu_series_keep_window_counts):
  // with them the node / fleet window planes follow from the per-series planes without another read of the slab
  uint16_t* d_wcnt = nullptr;
  size_t wcnt_capacity = 0;
  bool keep_counts = false;
  bool wcnt_valid = false;  // d_wcnt and the three planes describe `window` of the slab as it is now
  // quantile workspace (allocated on first use, see quantile.cu)
  void* q_ws = nullptr;
  // scratch of the fleet-level calls (fleet_series.cu), grow-only
  char* d_fs = nullptr;
  size_t fs_bytes = 0;
  uint32_t fleet_quantile_reads = 0;  // full passes over the slab the last igpu_series_quantile_fleet made (1, 3 or 4)
};

void igpu_quantile_ws_free(igpu_series* s);

// ---- pieces of K3 shared with the fused single-read kernel (quantile.cu) -------------------------
constexpr int WS_THREADS = 256;
constexpr uint32_t WS_MAX_L = 97;  // longest odd chunk the staged window kernel takes (two 97 KB stages)
constexpr uint32_t WS_MAX_L4 = 24; // longest chunk in float4 units (two 96 KB stages)

struct WsPlan {
  uint32_t n_win, PW, log2PW, L, K, tiles_per_row;
  uint64_t n_tiles;
  bool fast;  // window = PW * L, PW power of two <= 32, L odd <= WS_MAX_L, rows 16-byte aligned
  bool vec4;  // L counts float4s instead of floats (windows without a usable odd factor: 64, 128, 256, ...)
};
WsPlan igpu_ws_plan(uint32_t window, uint32_t n_time, uint32_t n_series);
// (Re)allocates the three window output planes for `window`; sets s->window / s->n_win.
int32_t igpu_series_ensure_window_buffers(igpu_ctx* ctx, igpu_series* s, uint32_t window);

#ifdef __CUDACC__
struct Acc {
  float mn, mx;
  double sum;
  uint32_t cnt;
};

__device__ __forceinline__ void acc_sample(Acc& a, float v) {
  const bool ok = v == v;
  a.mn = fminf(a.mn, v);  // fminf / fmaxf return the non-NaN operand
  a.mx = fmaxf(a.mx, v);
  a.sum += ok ? (double)v : 0.0;
  a.cnt += ok;
}

// (row, tile-in-row) of a CTA's tile stream, advanced without a division
struct TileCursor {
  uint32_t row, tr;
  __device__ __forceinline__ void advance(uint32_t d_row, uint32_t d_tr, uint32_t tpr) {
    row += d_row;
    tr += d_tr;
    if (tr >= tpr) tr -= tpr, ++row;
  }
};
#endif
