// MUFU throughput probe: clocks per warp-instruction per SM sub-partition for tanh.approx / ex2.approx / rcp.approx
// (the conv epilogue evaluates one MUFU per output element; this tells how many clocks that costs).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mufu_probe tools/mufu_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__device__ __forceinline__ float mufu(float x) {
  float y;
  if (OP == 0) asm volatile("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  else if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  else if (OP == 2) asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  else {  // tanh.approx.f16x2: two elements per MUFU
    unsigned u = __float_as_uint(x), r;
    asm volatile("tanh.approx.f16x2 %0, %1;" : "=r"(r) : "r"(u));
    y = __uint_as_float(r);
  }
  return y;
}

template <int OP>
__global__ void probe(float* out, long long* clk, int iters) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = 0.001f * (threadIdx.x + i);
  __syncthreads();
  const long long t0 = clock64();
  for (int k = 0; k < iters; ++k) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = mufu<OP>(v[i]);
  }
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int threads) {
  float* out; long long* clk;
  cudaMalloc(&out, 148 * 1024 * sizeof(float));
  cudaMalloc(&clk, 148 * sizeof(long long));
  const int iters = 2000;
  probe<OP><<<148, threads>>>(out, clk, iters);
  probe<OP><<<148, threads>>>(out, clk, iters);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  const double warps_per_smsp = threads / 32 / 4.0;
  const double per_instr = (double)h[0] / (iters * 16.0 * warps_per_smsp);
  printf("%-18s %4d threads/SM: %8lld clk for %d x 16 MUFU per warp -> %.2f clk per warp-instruction per SMSP (%s)\n", name,
         threads, h[0], iters, per_instr, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(clk);
}

int main() {
  for (int threads : {128, 512, 1024}) {
    run<0>("tanh.approx.f32", threads);
    run<1>("ex2.approx.f32", threads);
    run<2>("rcp.approx.f32", threads);
    run<3>("tanh.approx.f16x2", threads);
  }
  return 0;
}
