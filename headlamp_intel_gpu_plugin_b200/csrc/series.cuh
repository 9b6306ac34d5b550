// series.cuh — device-resident power time-series slab shared by series.cu (K3) and quantile.cu (K4).
#pragma once
#include "common.cuh"

struct igpu_series {
  uint32_t n_series = 0, n_time = 0, n_nodes = 0;
  float* d_x = nullptr;            // [n_series][n_time]
  uint32_t* d_node_off = nullptr;  // [n_nodes + 1] first series of each node
  uint32_t* h_node_off = nullptr;  // host copy
  // window statistics of the last igpu_series_window_stats call
  uint32_t window = 0, n_win = 0;
  float *d_wmin = nullptr, *d_wmean = nullptr, *d_wmax = nullptr;
  size_t w_capacity = 0;  // floats per plane
  // quantile workspace (allocated on first use, see quantile.cu)
  void* q_ws = nullptr;
};

void igpu_quantile_ws_free(igpu_series* s);
