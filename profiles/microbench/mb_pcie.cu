// mb_pcie.cu — host<->device copy rates on this box: default pinned vs write-combined pinned memory, one direction and
// both at once (two streams).  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mb_pcie mb_pcie.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
int main() {
  const size_t bytes = 2ull << 30;
  void *d0, *d1, *h_plain, *h_wc, *h_back;
  CK(cudaMalloc(&d0, bytes));
  CK(cudaMalloc(&d1, bytes));
  CK(cudaHostAlloc(&h_plain, bytes, cudaHostAllocDefault));
  CK(cudaHostAlloc(&h_wc, bytes, cudaHostAllocWriteCombined));
  CK(cudaHostAlloc(&h_back, bytes, cudaHostAllocDefault));
  memset(h_plain, 1, bytes);
  memset(h_wc, 1, bytes);
  cudaStream_t s0, s1;
  CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  auto time = [&](const char* name, auto fn, double gb) {
    fn();
    cudaDeviceSynchronize();
    float best = 1e9f;
    for (int r = 0; r < 3; ++r) {
      cudaEventRecord(a, s0);
      fn();
      cudaStreamSynchronize(s1);
      cudaEventRecord(b, s0);
      cudaEventSynchronize(b);
      float ms;
      cudaEventElapsedTime(&ms, a, b);
      if (ms < best) best = ms;
    }
    printf("%-44s %7.2f ms  %6.1f GB/s\n", name, best, gb / best * 1e3);
  };
  const double gb = bytes / 1e9;
  time("H2D pinned", [&] { cudaMemcpyAsync(d0, h_plain, bytes, cudaMemcpyHostToDevice, s0); }, gb);
  time("H2D write-combined", [&] { cudaMemcpyAsync(d0, h_wc, bytes, cudaMemcpyHostToDevice, s0); }, gb);
  time("D2H pinned", [&] { cudaMemcpyAsync(h_back, d1, bytes, cudaMemcpyDeviceToHost, s0); }, gb);
  time("H2D pinned + D2H pinned (two streams)", [&] {
    cudaMemcpyAsync(d0, h_plain, bytes, cudaMemcpyHostToDevice, s0);
    cudaMemcpyAsync(h_back, d1, bytes, cudaMemcpyDeviceToHost, s1);
  }, 2 * gb);
  time("H2D write-combined + D2H pinned", [&] {
    cudaMemcpyAsync(d0, h_wc, bytes, cudaMemcpyHostToDevice, s0);
    cudaMemcpyAsync(h_back, d1, bytes, cudaMemcpyDeviceToHost, s1);
  }, 2 * gb);
  time("H2D pinned, 8 x 256 MB pieces", [&] {
    for (int i = 0; i < 8; ++i) cudaMemcpyAsync((char*)d0 + i * (bytes / 8), (char*)h_plain + i * (bytes / 8), bytes / 8, cudaMemcpyHostToDevice, s0);
  }, gb);
  time("H2D pinned, two streams x half", [&] {
    cudaMemcpyAsync(d0, h_plain, bytes / 2, cudaMemcpyHostToDevice, s0);
    cudaMemcpyAsync((char*)d0 + bytes / 2, (char*)h_plain + bytes / 2, bytes / 2, cudaMemcpyHostToDevice, s1);
  }, gb);
  return 0;
}
