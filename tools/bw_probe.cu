// HBM bandwidth by read:write mix (the early conv layers and the preprocess kernel are write-heavy):
// grid-stride uint4 kernels over buffers larger than L2.  build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>

// each thread: R loads and W stores of 16 bytes per iteration (different buffers)
template <int R, int W>
__global__ void __launch_bounds__(256) mix(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint4 v = __ldcs(src + i + (size_t)r * n);
      acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
    }
#pragma unroll
    for (int w = 0; w < W; ++w) __stcs(dst + i + (size_t)w * n, acc);
    if (W == 0 && acc.x == 0x12345u) dst[0] = acc;
  }
}

template <int R, int W>
void run(uint4* a, uint4* b, size_t n, int ctas) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int k = 0; k < 6; ++k) {
    cudaEventRecord(e0);
    mix<R, W><<<ctas, 256>>>(a, b, n);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (k && ms < best) best = ms;
  }
  const double bytes = (double)(R + W) * n * 16;
  printf("read:write %d:%d  %4d CTAs  %.1f MB  %.3f ms  %.0f GB/s total (%.0f read, %.0f write)  %s\n", R, W, ctas, bytes / 1e6,
         best, bytes / best / 1e6, R * n * 16.0 / best / 1e6, W * n * 16.0 / best / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const size_t n = (size_t)16 << 20;  // 16 Mi uint4 = 256 MiB per unit stream
  uint4 *a, *b;
  cudaMalloc(&a, n * 16 * 4);
  cudaMalloc(&b, n * 16 * 4);
  cudaMemset(a, 1, n * 16 * 4);
  cudaMemset(b, 2, n * 16 * 4);
  for (int ctas : {148 * 8, 148 * 16}) {
    run<1, 0>(a, b, n, ctas);
    run<4, 0>(a, b, n, ctas);
    run<0, 1>(a, b, n, ctas);
    run<0, 4>(a, b, n, ctas);
    run<1, 1>(a, b, n, ctas);
    run<2, 2>(a, b, n, ctas);
    run<1, 3>(a, b, n, ctas);
    run<1, 2>(a, b, n, ctas);
    run<2, 1>(a, b, n, ctas);
    run<3, 1>(a, b, n, ctas);
  }
  return 0;
}
