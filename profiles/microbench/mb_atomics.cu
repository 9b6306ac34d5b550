// Micro-benchmarks that decide two design points of the series / allocation kernels:
//  (1) how fast can a CTA-private shared-memory histogram absorb one atomic per f32 sample
//      (spread bins vs. clustered bins) compared with the HBM stream rate;
//  (2) how fast are global RED.ADD.64 updates onto ~1e4 per-node accumulators;
//  (3) plain streaming read rate of a float4 grid-stride loop and a compare+count loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o mb_atomics mb_atomics.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t fkey(float f) {
  uint32_t u = __float_as_uint(f);
  return u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);
}

__global__ void fill(float* x, size_t n, int mode) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint64_t z = i * 0x9E3779B97F4A7C15ull + 0x1234567ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    float u = (float)(z >> 40) * (1.0f / 16777216.0f);
    x[i] = mode == 0 ? 40.0f + 200.0f * u                     // spread
                     : 150.0f + 0.5f * u;                     // clustered (smooth series)
  }
}

template <int BINS>
__global__ void __launch_bounds__(512) hist_smem(const float4* __restrict__ x, size_t n4, float lo, float inv, unsigned* out) {
  __shared__ unsigned h[BINS];
  for (int i = threadIdx.x; i < BINS; i += blockDim.x) h[i] = 0;
  __syncthreads();
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    float4 v = x[i];
    int b0 = min(BINS - 1, max(0, (int)((v.x - lo) * inv)));
    int b1 = min(BINS - 1, max(0, (int)((v.y - lo) * inv)));
    int b2 = min(BINS - 1, max(0, (int)((v.z - lo) * inv)));
    int b3 = min(BINS - 1, max(0, (int)((v.w - lo) * inv)));
    atomicAdd(&h[b0], 1u); atomicAdd(&h[b1], 1u); atomicAdd(&h[b2], 1u); atomicAdd(&h[b3], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < BINS; i += blockDim.x) if (h[i]) atomicAdd(&out[i], h[i]);
}

__global__ void __launch_bounds__(512) count_below(const float4* __restrict__ x, size_t n4, float p0, float p1, unsigned long long* out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  unsigned c0 = 0, c1 = 0; float mn = 3e38f, mx = -3e38f; double s = 0;
  for (; i < n4; i += stride) {
    float4 v = x[i];
    c0 += (v.x < p0) + (v.y < p0) + (v.z < p0) + (v.w < p0);
    c1 += (v.x <= p1) + (v.y <= p1) + (v.z <= p1) + (v.w <= p1);
    mn = fminf(fminf(mn, v.x), fminf(v.y, fminf(v.z, v.w)));
    mx = fmaxf(fmaxf(mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
    s += (double)((v.x + v.y) + (v.z + v.w));
  }
  for (int o = 16; o; o >>= 1) { c0 += __shfl_xor_sync(~0u, c0, o); c1 += __shfl_xor_sync(~0u, c1, o); s += __shfl_xor_sync(~0u, s, o); }
  if ((threadIdx.x & 31) == 0) { atomicAdd(&out[0], c0); atomicAdd(&out[1], c1); if (mn > mx || s == 1.2345) out[2] = 1; }
}

__global__ void __launch_bounds__(256) red_nodes(const uint32_t* __restrict__ node, size_t n, int nacc, unsigned long long* acc, int nnodes) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint32_t nd = node[i];
    for (int a = 0; a < nacc; ++a) atomicAdd(&acc[(size_t)a * nnodes + nd], 1ull + a);
  }
}

__global__ void fill_nodes(uint32_t* node, size_t n, int nnodes, int sorted) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint64_t z = i * 0x9E3779B97F4A7C15ull + 99;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    node[i] = sorted ? (uint32_t)((i * (uint64_t)nnodes) / n) : (uint32_t)(z % nnodes);
  }
}

template <class F> float timeit(F f, int reps) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); for (int i = 0; i < reps; ++i) f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / reps;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  printf("device %s SMs %d smem/blk optin %zu L2 %d MB\n", p.name, p.multiProcessorCount, p.sharedMemPerBlockOptin, p.l2CacheSize >> 20);
  size_t n = (size_t)1 << 30;  // 4 GiB of f32
  float* x; CK(cudaMalloc(&x, n * 4));
  unsigned* hout; CK(cudaMalloc(&hout, 65536 * 4)); CK(cudaMemset(hout, 0, 65536 * 4));
  unsigned long long* cout_; CK(cudaMalloc(&cout_, 64)); CK(cudaMemset(cout_, 0, 64));
  int sms = p.multiProcessorCount;
  for (int mode = 0; mode < 2; ++mode) {
    fill<<<sms * 8, 512>>>(x, n, mode); CK(cudaDeviceSynchronize());
    float lo = mode == 0 ? 40.f : 150.f, span = mode == 0 ? 200.f : 0.5f;
    float ms;
    ms = timeit([&] { count_below<<<sms * 4, 512>>>((const float4*)x, n / 4, lo + 0.9f * span, lo + 0.95f * span, cout_); }, 5);
    printf("mode %d count_below+minmax+sum : %.3f ms  %.1f GB/s\n", mode, ms, n * 4 / ms / 1e6);
    ms = timeit([&] { hist_smem<2048><<<sms * 4, 512>>>((const float4*)x, n / 4, lo, 2048.f / span, hout); }, 3);
    printf("mode %d hist_smem<2048> 4cta/sm : %.3f ms  %.1f GB/s  %.1f Gatom/s\n", mode, ms, n * 4 / ms / 1e6, n / ms / 1e6);
    ms = timeit([&] { hist_smem<4096><<<sms * 2, 512>>>((const float4*)x, n / 4, lo, 4096.f / span, hout); }, 3);
    printf("mode %d hist_smem<4096> 2cta/sm : %.3f ms  %.1f GB/s  %.1f Gatom/s\n", mode, ms, n * 4 / ms / 1e6, n / ms / 1e6);
    ms = timeit([&] { hist_smem<256><<<sms * 4, 512>>>((const float4*)x, n / 4, lo, 256.f / span, hout); }, 3);
    printf("mode %d hist_smem<256>  4cta/sm : %.3f ms  %.1f GB/s  %.1f Gatom/s\n", mode, ms, n * 4 / ms / 1e6, n / ms / 1e6);
  }
  // global RED onto per-node accumulators
  size_t np = 4u << 20; int nnodes = 10000;
  uint32_t* node; CK(cudaMalloc(&node, np * 4));
  unsigned long long* acc; CK(cudaMalloc(&acc, (size_t)8 * nnodes * 8)); CK(cudaMemset(acc, 0, (size_t)8 * nnodes * 8));
  for (int sorted = 0; sorted < 2; ++sorted) {
    fill_nodes<<<sms * 4, 256>>>(node, np, nnodes, sorted); CK(cudaDeviceSynchronize());
    for (int nacc = 1; nacc <= 8; nacc *= 2) {
      float ms = timeit([&] { red_nodes<<<sms * 8, 256>>>(node, np, nacc, acc, nnodes); }, 5);
      printf("red sorted=%d nacc=%d : %.1f us  %.2f Gred/s\n", sorted, nacc, ms * 1e3, np * (double)nacc / ms / 1e6);
    }
  }
  // pinned H2D bandwidth
  size_t hb = (size_t)1 << 30; void* h; CK(cudaMallocHost(&h, hb));
  float ms = timeit([&] { cudaMemcpyAsync(x, h, hb, cudaMemcpyHostToDevice, 0); }, 3);
  printf("H2D pinned 1GiB: %.2f ms %.1f GB/s\n", ms, hb / ms / 1e6);
  ms = timeit([&] { cudaMemcpyAsync(h, x, hb, cudaMemcpyDeviceToHost, 0); }, 3);
  printf("D2H pinned 1GiB: %.2f ms %.1f GB/s\n", ms, hb / ms / 1e6);
  CK(cudaDeviceSynchronize());
  return 0;
}
