// Hardware probes that decide kernel design choices on sm_100a (run on the GPU box; results are recorded
// in DESIGN.md):
//   T1  UMMA shared-memory descriptors on shifted / strided views of one swizzled patch
//       (start address offset by whole 128-byte rows, SBO that is not a multiple of 1024, base_offset field)
//   T3  TMA tensor maps whose pixel stride is smaller than the box inner extent (overlapping rows)
//   T4  TMA load throughput per SM as a function of the box inner row size (32 / 64 / 128 bytes)
//   T5  accuracy of SiLU variants (ex2+rcp vs tanh.approx)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/umma_probe tools/umma_probe.cu
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../aquaculture_b200/csrc/ptx.cuh"

using namespace aq::ptx;

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) {                                                                    \
      printf("CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                                  \
    }                                                                                           \
  } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode;

__device__ __forceinline__ uint64_t make_desc_ex(uint32_t addr, uint32_t sbo_bytes, uint32_t base_off) {
  return (uint64_t)((addr & 0x3FFFF) >> 4) | (1ull << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46) |
         ((uint64_t)(base_off & 7) << 49) | (2ull << 61);
}

// ------------------------------------------------------------------------------------------------ T1
constexpr int kPN = 32;  // UMMA N of the probe
__global__ void __launch_bounds__(128) umma_view_kernel(const __half* Apatch, const __half* B, float* D, int off_rows,
                                                        int sbo_bytes, int bo, int R) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tslot;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  for (int i = threadIdx.x; i < R * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    *reinterpret_cast<uint4*>(sm + r * 128 + ((j ^ (r & 7)) << 4)) = reinterpret_cast<const uint4*>(Apatch)[i];
  }
  const uint32_t boff = ((R * 128 + 1023) / 1024) * 1024;
  for (int i = threadIdx.x; i < kPN * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    *reinterpret_cast<uint4*>(sm + boff + r * 128 + ((j ^ (r & 7)) << 4)) = reinterpret_cast<const uint4*>(B)[i];
  }
  fence_proxy_async_smem();
  if (warp == 0) {
    tmem_alloc(smem_u32(&tslot), 32);
    tmem_relinquish();
  }
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tslot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_f16(128, kPN);
    const uint32_t a_addr = base + off_rows * 128;
    for (int k = 0; k < 4; ++k)
      umma_f16(tmem, make_desc_ex(a_addr + k * 32, sbo_bytes, bo), make_desc_ex(base + boff + k * 32, 1024, 0), idesc,
               k > 0);
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  for (int c0 = 0; c0 < kPN; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tmem + ((warp * 32u) << 16) + c0, v);
    tmem_ld_wait();
    for (int i = 0; i < 16; ++i) D[(warp * 32 + lane) * kPN + c0 + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 32);
}

static void run_t1() {
  const int R = 320;
  std::vector<__half> A(R * 64), B(kPN * 64);
  srand(1);
  for (auto& v : A) v = __float2half((float)(rand() % 9 - 4));
  for (auto& v : B) v = __float2half((float)(rand() % 7 - 3));
  __half *dA, *dB;
  float* dD;
  CK(cudaMalloc(&dA, A.size() * 2));
  CK(cudaMalloc(&dB, B.size() * 2));
  CK(cudaMalloc(&dD, 128 * kPN * 4));
  CK(cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice));
  const size_t smem = R * 128 + kPN * 128 + 2048;
  CK(cudaFuncSetAttribute(umma_view_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // expected rows: full[r][n] = sum_k A[r][k] B[n][k]
  std::vector<float> full(R * kPN);
  for (int r = 0; r < R; ++r)
    for (int n = 0; n < kPN; ++n) {
      float s = 0;
      for (int k = 0; k < 64; ++k) s += __half2float(A[r * 64 + k]) * __half2float(B[n * 64 + k]);
      full[r * kPN + n] = s;
    }
  struct V { int off, sbo, bo; };
  const V vars[] = {{0, 1024, 0},  {8, 1024, 0},  {16, 1024, 0}, {1, 1024, 0},  {1, 1024, 1},  {3, 1024, 0},
                    {3, 1024, 3},  {4, 1024, 0},  {4, 1024, 4},  {0, 1280, 0},  {11, 1280, 0}, {11, 1280, 3},
                    {0, 2304, 0},  {19, 2304, 0}, {19, 2304, 3}, {0, 512, 0},   {2, 512, 0},   {0, 2048, 0},
                    {9, 2048, 0},  {9, 2048, 1}};
  for (const V& v : vars) {
    CK(cudaMemset(dD, 0, 128 * kPN * 4));
    umma_view_kernel<<<1, 128, smem>>>(dA, dB, dD, v.off, v.sbo, v.bo, R);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("T1 off=%d sbo=%d bo=%d: CUDA error %s\n", v.off, v.sbo, v.bo, cudaGetErrorString(e));
      exit(3);
    }
    std::vector<float> D(128 * kPN);
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    int ok_rows = 0, matched_other = 0, unmatched = 0;
    char map[256] = {0};
    int mp = 0;
    for (int m = 0; m < 128; ++m) {
      const int want = v.off + (m / 8) * (v.sbo / 128) + (m % 8);
      int found = -1;
      for (int r = 0; r < R && found < 0; ++r)
        if (!memcmp(&D[m * kPN], &full[r * kPN], kPN * 4)) found = r;
      if (found == want) ++ok_rows;
      else if (found >= 0) ++matched_other;
      else ++unmatched;
      if (m < 20 && mp < 200) mp += snprintf(map + mp, sizeof(map) - mp, "%d>%d ", want, found);
    }
    printf("T1 off=%2d sbo=%4d bo=%d : rows ok %3d / other-row %3d / garbage %3d | want>got: %s\n", v.off, v.sbo, v.bo,
           ok_rows, matched_other, unmatched, map);
  }
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
}

// ------------------------------------------------------------------------------------------------ T3
__global__ void __launch_bounds__(128) tma_overlap_kernel(const __grid_constant__ CUtensorMap tm, __half* out, int rank,
                                                          int c0, int c1, int c2, int c3, int rows) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(smem_u32(&bar), rows * 128);
    if (rank == 2) tma_load_2d(base, &tm, smem_u32(&bar), c0, c1);
    else tma_load_4d(base, &tm, smem_u32(&bar), c0, c1, c2, c3);
  }
  mbar_wait(smem_u32(&bar), 0);
  for (int i = threadIdx.x; i < rows * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    reinterpret_cast<uint4*>(out)[i] = *reinterpret_cast<const uint4*>(sm + r * 128 + ((j ^ (r & 7)) << 4));
  }
}

static void run_t3() {
  const int W = 40, H = 10, N = 2, C = 16;  // 16 channels per pixel (32 B), rows of 64 elements (4 pixels)
  const size_t n_el = (size_t)N * H * W * C + 256;
  std::vector<__half> src(n_el);
  for (size_t i = 0; i < n_el; ++i) src[i] = __float2half((float)(i % 2039));
  __half *d, *dout;
  CK(cudaMalloc(&d, n_el * 2));
  CK(cudaMalloc(&dout, 64 * 64 * 2));
  CK(cudaMemcpy(d, src.data(), n_el * 2, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(tma_overlap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
  {
    CUtensorMap tm;
    cuuint64_t gd[2] = {64, (cuuint64_t)N * H * W};
    cuuint64_t gs[1] = {(cuuint64_t)C * 2};
    cuuint32_t bx[2] = {64, 8}, es[2] = {1, 1};
    CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("T3 2-D overlapped map (dim0 64 el, pixel stride 16 el): encode result %d\n", (int)r);
    if (r == CUDA_SUCCESS) {
      tma_overlap_kernel<<<1, 128, 16384>>>(tm, dout, 2, 0, 5, 0, 0, 8);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("T3 2-D: CUDA error %s\n", cudaGetErrorString(e)); exit(3); }
      std::vector<__half> o(8 * 64);
      CK(cudaMemcpy(o.data(), dout, o.size() * 2, cudaMemcpyDeviceToHost));
      int bad = 0;
      for (int p = 0; p < 8; ++p)
        for (int k = 0; k < 64; ++k)
          if (__half2float(o[p * 64 + k]) != __half2float(src[(5 + p) * C + k])) ++bad;
      printf("T3 2-D overlapped load: %d / 512 elements wrong\n", bad);
    }
  }
  {
    CUtensorMap tm;
    cuuint64_t gd[4] = {64, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t gs[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t bx[4] = {64, 8, 4, 1}, es[4] = {1, 1, 1, 1};
    CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("T3 4-D overlapped map: encode result %d\n", (int)r);
    if (r == CUDA_SUCCESS) {
      tma_overlap_kernel<<<1, 128, 16384>>>(tm, dout, 4, 0, -1, -1, 1, 32);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("T3 4-D: CUDA error %s\n", cudaGetErrorString(e)); exit(3); }
      std::vector<__half> o(32 * 64);
      CK(cudaMemcpy(o.data(), dout, o.size() * 2, cudaMemcpyDeviceToHost));
      int bad = 0;
      for (int yy = 0; yy < 4; ++yy)
        for (int xx = 0; xx < 8; ++xx)
          for (int k = 0; k < 64; ++k) {
            const int y = yy - 1, x = xx - 1;
            float want = 0.f;
            if (y >= 0 && x >= 0) want = __half2float(src[(((size_t)1 * H + y) * W + x) * C + k]);
            if (__half2float(o[(yy * 8 + xx) * 64 + k]) != want) ++bad;
          }
      printf("T3 4-D overlapped load with negative start (OOB rows -> 0): %d / 2048 elements wrong\n", bad);
    }
  }
  cudaFree(d); cudaFree(dout);
}

// ------------------------------------------------------------------------------------------------ T4
constexpr int kS = 6;
__global__ void __launch_bounds__(96) tma_rate_kernel(const __grid_constant__ CUtensorMap tm, int iters, int loads,
                                                      int box_rows, int row_bytes, long rows_total, int nprod,
                                                      unsigned long long* sink) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bars[2 * kS];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t stage_bytes = loads * box_rows * row_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kS; ++s) {
      mbar_init(smem_u32(&bars[s]), nprod);
      mbar_init(smem_u32(&bars[kS + s]), 1);
    }
    fence_barrier_init();
  }
  __syncthreads();
  if (warp < (uint32_t)nprod && lane == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(smem_u32(&bars[kS + stage]), phase ^ 1);
      const int my_loads = loads / nprod;
      mbar_arrive_expect_tx(smem_u32(&bars[stage]), my_loads * box_rows * row_bytes);
      for (int l = 0; l < my_loads; ++l) {
        const int li = warp * my_loads + l;
        const long row = (((long)blockIdx.x * iters + it) * loads + li) * box_rows % rows_total;
        tma_load_2d(base + stage * stage_bytes + li * box_rows * row_bytes, &tm, smem_u32(&bars[stage]), 0, (int)row);
      }
      if (++stage == kS) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 2 && lane == 0) {
    int stage = 0;
    uint32_t phase = 0;
    unsigned long long acc = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(smem_u32(&bars[stage]), phase);
      acc += *reinterpret_cast<volatile uint32_t*>(raw + 1024 + stage * stage_bytes);
      mbar_arrive(smem_u32(&bars[kS + stage]));
      if (++stage == kS) { stage = 0; phase ^= 1; }
    }
    if (acc == 0x1234567) *sink = acc;
  }
}

static void run_t4() {
  const size_t big = (size_t)6 << 30;  // 6 GiB source for the HBM runs
  uint8_t* d;
  unsigned long long* sink;
  CK(cudaMalloc(&d, big));
  CK(cudaMemset(d, 1, big));
  CK(cudaMalloc(&sink, 8));
  CK(cudaFuncSetAttribute(tma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  struct V { int cb, box_rows, loads, nprod; bool l2; };
  const V vars[] = {{16, 128, 4, 1, true}, {32, 128, 4, 1, true}, {64, 128, 1, 1, true},  {64, 128, 1, 1, false},
                    {32, 128, 2, 1, false}, {16, 128, 4, 1, false}, {64, 128, 2, 2, true}, {16, 128, 4, 2, true},
                    {64, 256, 1, 1, true}, {64, 32, 4, 1, true},   {64, 8, 16, 1, true},  {16, 256, 2, 1, true}};
  for (const V& v : vars) {
    const int row_bytes = v.cb * 2;
    const long rows_total = v.l2 ? (long)(32 << 20) / row_bytes : (long)(big / row_bytes);
    CUtensorMap tm;
    cuuint64_t gd[2] = {(cuuint64_t)v.cb, (cuuint64_t)rows_total};
    cuuint64_t gs[1] = {(cuuint64_t)row_bytes};
    cuuint32_t bx[2] = {(cuuint32_t)v.cb, (cuuint32_t)v.box_rows}, es[2] = {1, 1};
    CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                           : (row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B),
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("T4 encode failed %d\n", (int)r); continue; }
    const int iters = v.l2 ? 4000 : 1500;
    const size_t smem = (size_t)kS * v.loads * v.box_rows * row_bytes + 2048;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      tma_rate_kernel<<<148, 96, smem>>>(tm, iters, v.loads, v.box_rows, row_bytes, rows_total - 4096, v.nprod, sink);
      cudaEventRecord(e1);
      cudaError_t e = cudaEventSynchronize(e1);
      if (e != cudaSuccess) { printf("T4: CUDA error %s\n", cudaGetErrorString(e)); exit(3); }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep == 1) {
        const double rows = 148.0 * iters * v.loads * v.box_rows;
        const double bytes = rows * row_bytes;
        printf("T4 row %3d B, box %3d rows x %2d loads/stage, %d producer(s), %s: %.3f ms  %.2f TB/s  %.1f ns/row/SM  (%.2f Grows/s chip)\n",
               row_bytes, v.box_rows, v.loads, v.nprod, v.l2 ? "L2-resident" : "HBM", ms, bytes / ms / 1e9,
               ms * 1e6 / (rows / 148.0), rows / ms / 1e6);
      }
    }
  }
  cudaFree(d); cudaFree(sink);
}

// ------------------------------------------------------------------------------------------------ T5
__global__ void silu_kernel(const float* x, float* a, float* b, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = x[i];
  a[i] = __fdividef(v, 1.0f + __expf(-v));
  const float h = 0.5f * v;
  b[i] = fmaf(h, tanh_approx(h), h);
}
static void run_t5() {
  const int n = 1 << 20;
  std::vector<float> x(n), a(n), b(n);
  for (int i = 0; i < n; ++i) x[i] = -16.f + 32.f * (float)i / n;
  float *dx, *da, *db;
  CK(cudaMalloc(&dx, n * 4)); CK(cudaMalloc(&da, n * 4)); CK(cudaMalloc(&db, n * 4));
  CK(cudaMemcpy(dx, x.data(), n * 4, cudaMemcpyHostToDevice));
  silu_kernel<<<n / 256, 256>>>(dx, da, db, n);
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(a.data(), da, n * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(b.data(), db, n * 4, cudaMemcpyDeviceToHost));
  double ea = 0, eb = 0, ra = 0, rb = 0, ha = 0, hb = 0;
  for (int i = 0; i < n; ++i) {
    const double v = x[i], ref = v / (1.0 + exp(-v));
    const double da_ = fabs(a[i] - ref), db_ = fabs(b[i] - ref);
    ea = fmax(ea, da_); eb = fmax(eb, db_);
    if (fabs(ref) > 1e-3) { ra = fmax(ra, da_ / fabs(ref)); rb = fmax(rb, db_ / fabs(ref)); }
    // error measured in fp16 ulps of the result (what matters after the fp16 store)
    const double ulp = fmax(ldexp(1.0, (int)floor(log2(fmax(fabs(ref), 6.1e-5))) - 10), 5.96e-8);
    ha = fmax(ha, da_ / ulp); hb = fmax(hb, db_ / ulp);
  }
  printf("T5 SiLU ex2+rcp : max abs %.3e  max rel %.3e  max err in fp16 ulps %.3f\n", ea, ra, ha);
  printf("T5 SiLU tanh    : max abs %.3e  max rel %.3e  max err in fp16 ulps %.3f\n", eb, rb, hb);
  cudaFree(dx); cudaFree(da); cudaFree(db);
}

// ------------------------------------------------------------------------------------------------ T6
// TMA *store* of a box whose start coordinate is negative (rows before 0 must simply be dropped)
__global__ void __launch_bounds__(128) tma_store_neg_kernel(const __grid_constant__ CUtensorMap tm, int y0, int use_reduce) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  // smem box [16 rows y][8 px][64 ch] fp16, value = y * 8 + x + 1, swizzled
  for (int i = threadIdx.x; i < 128 * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    __half v[8];
    for (int k = 0; k < 8; ++k) v[k] = __float2half((float)(r + 1));
    *reinterpret_cast<uint4*>(sm + r * 128 + ((j ^ (r & 7)) << 4)) = *reinterpret_cast<uint4*>(v);
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (use_reduce) tma_reduce_add_4d(&tm, base, 0, 0, y0, 0);
    else tma_store_4d(&tm, base, 0, 0, y0, 0);
    tma_store_commit();
    tma_store_wait0();
  }
}
static void run_t6() {
  const int W = 8, H = 20, C = 64;
  __half* d;
  CK(cudaMalloc(&d, (size_t)H * W * C * 2));
  CK(cudaFuncSetAttribute(tma_store_neg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
  CUtensorMap tm;
  cuuint64_t gd[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, 1};
  cuuint64_t gs[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t bx[4] = {64, 8, 16, 1}, es[4] = {1, 1, 1, 1};
  CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("T6 encode failed %d\n", (int)r); return; }
  const int tests[][2] = {{3, 0}, {10, 0}, {-3, 0}, {-3, 1}, {-15, 0}};
  for (auto& t : tests) {
    CK(cudaMemset(d, 0, (size_t)H * W * C * 2));
    tma_store_neg_kernel<<<1, 128, 32768>>>(tm, t[0], t[1]);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("T6 y0=%d reduce=%d: CUDA error %s\n", t[0], t[1], cudaGetErrorString(e)); exit(3); }
    std::vector<__half> o((size_t)H * W * C);
    CK(cudaMemcpy(o.data(), d, o.size() * 2, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x) {
        const int by = y - t[0];
        const float want = (by >= 0 && by < 16) ? (float)(by * 8 + x + 1) : 0.f;
        for (int c = 0; c < C; ++c) bad += __half2float(o[((size_t)y * W + x) * C + c]) != want;
      }
    printf("T6 store box rows [%d, %d) into H=%d (%s): %d wrong elements\n", t[0], t[0] + 16, H, t[1] ? "reduce-add" : "store", bad);
  }
  cudaFree(d);
}


// ------------------------------------------------------------------------------------------------ T7
// tcgen05.mma issue / execution rate of ONE thread: clk per K=16 MMA as a function of N, of the number of
// accumulators the thread cycles through (dependent chains), of tcgen05.commit frequency and of a second
// issuing thread.  Operands are zero-filled shared memory (no loads in flight).
__global__ void __launch_bounds__(128) mma_rate_kernel(int N, int n_acc, int commit_every, int n_issuers, int iters,
                                                       long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bars[8];
  __shared__ uint32_t tslot;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < (16384 + 32768) / 16; i += 128)
    *reinterpret_cast<uint4*>(raw + (base - smem_u32(raw)) + i * 16) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (warp == 0) {
    tmem_alloc(smem_u32(&tslot), 512);
    tmem_relinquish();
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bars[i]), 1);
    fence_barrier_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tslot;
  if (warp < (uint32_t)n_issuers) {
    const uint32_t idesc = make_idesc_f16(128, N);
    const uint32_t a_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    const uint32_t a_lo = ((base & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t b_lo = (((base + 16384) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t pitch = (N + 31) / 32 * 32;
    int acc = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tmem + (warp * n_acc + acc) * pitch;
      if (elect_one()) {
        umma_f16_split(d, a_lo, a_hi, b_lo, a_hi, idesc, 1u);
        umma_f16_split(d, a_lo + 2, a_hi, b_lo + 2, a_hi, idesc, 1u);
        umma_f16_split(d, a_lo + 4, a_hi, b_lo + 4, a_hi, idesc, 1u);
        umma_f16_split(d, a_lo + 6, a_hi, b_lo + 6, a_hi, idesc, 1u);
        if (commit_every && (i % commit_every) == commit_every - 1) umma_commit(smem_u32(&bars[2 + warp]));
      }
      __syncwarp();
      if (++acc == n_acc) acc = 0;
    }
    const long long t1 = clock64();
    if (elect_one()) umma_commit(smem_u32(&bars[warp]));
    __syncwarp();
    mbar_wait(smem_u32(&bars[warp]), 0);
    const long long t2 = clock64();
    if (lane == 0 && blockIdx.x == 0) {
      out[warp * 2] = t1 - t0;
      out[warp * 2 + 1] = t2 - t0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

static void run_t7() {
  long long* d;
  CK(cudaMalloc(&d, 64));
  CK(cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  struct V { int N, n_acc, commit_every, n_issuers; };
  std::vector<V> vars;
  for (int N : {48, 96, 192, 256}) {
    vars.push_back({N, 1, 0, 1});
    vars.push_back({N, 1, 1, 1});
    vars.push_back({N, 1, 9, 1});
    if (2 * ((N + 31) / 32 * 32) <= 512) {
      vars.push_back({N, 2, 0, 1});
      vars.push_back({N, 1, 0, 2});
      vars.push_back({N, 1, 1, 2});
    }
    if (4 * ((N + 31) / 32 * 32) <= 512) vars.push_back({N, 2, 0, 2});
  }
  const int iters = 2000;
  for (const V& v : vars) {
    long long h[4] = {0, 0, 0, 0};
    for (int rep = 0; rep < 2; ++rep) {
      mma_rate_kernel<<<148, 128, 56 * 1024>>>(v.N, v.n_acc, v.commit_every, v.n_issuers, iters, d);
      CK(cudaDeviceSynchronize());
    }
    CK(cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost));
    printf("T7 N=%3d acc/thread=%d commit_every=%d issuers=%d : issue %.1f clk/MMA, complete %.1f clk/MMA per thread (ideal %d); "
           "combined %.1f clk/MMA\n",
           v.N, v.n_acc, v.commit_every, v.n_issuers, (double)h[0] / (iters * 4), (double)h[1] / (iters * 4), v.N / 2,
           (double)h[1] / (iters * 4) / v.n_issuers);
  }
  cudaFree(d);
}

// ------------------------------------------------------------------------------------------------ T8
// TMA load latency and throughput per SM vs boxes in flight: thread 0 issues `depth` boxes (128 rows x 128 B)
// on one mbarrier, waits for all, repeats.  L2-resident source (32 MiB window).
__global__ void __launch_bounds__(64) tma_lat_kernel(const __grid_constant__ CUtensorMap tm, int depth, int box_rows,
                                                     int iters, long rows_total, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long tot = 0, tissue = 0;
    for (int it = 0; it < iters; ++it) {
      const long long t0 = clock64();
      mbar_arrive_expect_tx(smem_u32(&bar), depth * box_rows * 128);
      for (int l = 0; l < depth; ++l) {
        const long row = (((long)blockIdx.x * iters + it) * depth + l) * box_rows % rows_total;
        tma_load_2d(base + l * box_rows * 128, &tm, smem_u32(&bar), 0, (int)row);
      }
      const long long t1 = clock64();
      mbar_wait(smem_u32(&bar), it & 1);
      tot += clock64() - t0;
      tissue += t1 - t0;
    }
    if (blockIdx.x == 0) {
      out[0] = tot / iters;
      out[1] = tissue / iters;
    }
  }
}

static void run_t8() {
  const size_t big = (size_t)1 << 30;
  uint8_t* d;
  long long* o;
  CK(cudaMalloc(&d, big));
  CK(cudaMemset(d, 1, big));
  CK(cudaMalloc(&o, 64));
  CK(cudaFuncSetAttribute(tma_lat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  for (int l2 = 1; l2 >= 0; --l2) {
    for (int box_rows : {128, 256, 32}) {
      const long rows_total = l2 ? (long)(32 << 20) / 128 : (long)(big / 128);
      CUtensorMap tm;
      cuuint64_t gd[2] = {64, (cuuint64_t)rows_total};
      cuuint64_t gs[1] = {128};
      cuuint32_t bx[2] = {64, (cuuint32_t)box_rows}, es[2] = {1, 1};
      CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("T8 encode failed %d\n", (int)r); continue; }
      for (int grid : {1, 148}) {
        for (int depth : {1, 2, 4, 8}) {
          if ((size_t)depth * box_rows * 128 > 190 * 1024) continue;
          long long h[2];
          for (int rep = 0; rep < 2; ++rep) {
            tma_lat_kernel<<<grid, 64, depth * box_rows * 128 + 2048>>>(tm, depth, box_rows, l2 ? 400 : 100, rows_total - 4096, o);
            CK(cudaDeviceSynchronize());
          }
          CK(cudaMemcpy(h, o, 16, cudaMemcpyDeviceToHost));
          printf("T8 %s box %3d rows x128B, %d in flight, grid %3d: round trip %6lld clk (issue %5lld) -> %.1f B/clk/SM\n",
                 l2 ? "L2 " : "HBM", box_rows, depth, grid, h[0], h[1], (double)depth * box_rows * 128 / h[0]);
        }
      }
    }
  }
  cudaFree(d); cudaFree(o);
}

// ------------------------------------------------------------------------------------------------ T9
// Halo-patch box loads: one 4-D TMA box (64 ch, pw, ph, 1 image) from an NHWC fp16 map, `depth` boxes in
// flight per SM, coordinates precomputed (no div/mod in the issue loop).
__global__ void __launch_bounds__(64) tma_patch_kernel(const __grid_constant__ CUtensorMap tm, int depth, int pw, int ph,
                                                       int iters, int W, int H, int N, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t slot = ((uint32_t)(pw * ph * 128) + 1023u) & ~1023u;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long tot = 0, tissue = 0;
    int x = (blockIdx.x * 8) % W, y = (blockIdx.x * 16) % H, n = blockIdx.x % N;
    for (int it = 0; it < iters; ++it) {
      const long long t0 = clock64();
      mbar_arrive_expect_tx(smem_u32(&bar), depth * pw * ph * 128);
      for (int l = 0; l < depth; ++l) {
        tma_load_4d(base + l * slot, &tm, smem_u32(&bar), 0, x - 1, y - 1, n);
        x += pw - 2;
        if (x >= W) { x = 0; y += ph - 2; if (y >= H) { y = 0; if (++n == N) n = 0; } }
      }
      const long long t1 = clock64();
      mbar_wait(smem_u32(&bar), it & 1);
      tot += clock64() - t0;
      tissue += t1 - t0;
    }
    if (blockIdx.x == 0) {
      out[0] = tot / iters;
      out[1] = tissue / iters;
    }
  }
}

static void run_t9() {
  const int W = 160, H = 160, N = 64, C = 64;
  const size_t bytes = (size_t)N * H * W * C * 2;
  uint8_t* d;
  long long* o;
  CK(cudaMalloc(&d, bytes));
  CK(cudaMemset(d, 1, bytes));
  CK(cudaMalloc(&o, 64));
  CK(cudaFuncSetAttribute(tma_patch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  const int shapes[4][2] = {{10, 18}, {18, 18}, {18, 34}, {34, 18}};
  for (auto& sh : shapes) {
    const int pw = sh[0], ph = sh[1];
    CUtensorMap tm;
    cuuint64_t gd[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t gs[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t bx[4] = {64, (cuuint32_t)pw, (cuuint32_t)ph, 1}, es[4] = {1, 1, 1, 1};
    CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("T9 encode failed %d\n", (int)r); continue; }
    const size_t slot = ((size_t)pw * ph * 128 + 1023) & ~(size_t)1023;
    for (int grid : {1, 148}) {
      for (int depth : {1, 2, 3}) {
        if (depth * slot > 200 * 1024) continue;
        long long h[2];
        for (int rep = 0; rep < 2; ++rep) {
          tma_patch_kernel<<<grid, 64, depth * slot + 2048>>>(tm, depth, pw, ph, 300, W, H, N, o);
          CK(cudaDeviceSynchronize());
        }
        CK(cudaMemcpy(h, o, 16, cudaMemcpyDeviceToHost));
        printf("T9 patch box 64ch x %2d x %2d (%5.1f KB), %d in flight, grid %3d: round trip %6lld clk (issue %5lld) -> %.1f B/clk/SM\n",
               pw, ph, pw * ph * 128 / 1024.0, depth, grid, h[0], h[1], (double)depth * pw * ph * 128 / h[0]);
      }
    }
  }
  cudaFree(d); cudaFree(o);
}

// ------------------------------------------------------------------------------------------------ T10
// Does a SWIZZLE_128B TMA load honour the absolute-address swizzle when its shared-memory destination is only
// 128-byte aligned (not 1024)?  Needed to land halo-patch segments at arbitrary patch rows.
__global__ void __launch_bounds__(128) tma_unaligned_kernel(const __grid_constant__ CUtensorMap tm, int off_bytes, int pw,
                                                            int ph, int x0, int y0, int* bad) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0xFFFFFFFFu;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(smem_u32(&bar), pw * ph * 128);
    tma_load_4d(base + off_bytes, &tm, smem_u32(&bar), 0, x0, y0, 0);
  }
  mbar_wait(smem_u32(&bar), 0);
  int nbad = 0;
  for (int i = threadIdx.x; i < pw * ph * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    const int py = r / pw, px = r - py * pw;
    const uint32_t row_addr = base + off_bytes + r * 128;
    const uint32_t a = row_addr + ((j ^ ((row_addr >> 7) & 7)) << 4);
    const __half got = *reinterpret_cast<const __half*>(sm + (a - base));
    const float want = (float)(((y0 + py) * 32 + (x0 + px)) * 8 + j);
    if (__half2float(got) != want) ++nbad;
  }
  atomicAdd(bad, nbad);
}

static void run_t10() {
  const int W = 32, H = 8, C = 64;
  std::vector<__half> h((size_t)H * W * C);
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x)
      for (int c = 0; c < C; ++c) h[((size_t)y * W + x) * C + c] = __float2half((float)((y * 32 + x) * 8 + (c >> 3)));
  __half* d;
  int* bad;
  CK(cudaMalloc(&d, h.size() * 2));
  CK(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&bad, 4));
  CK(cudaFuncSetAttribute(tma_unaligned_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  const int pw = 18, ph = 2;
  CUtensorMap tm;
  cuuint64_t gd[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, 1};
  cuuint64_t gs[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t bx[4] = {64, (cuuint32_t)pw, (cuuint32_t)ph, 1}, es[4] = {1, 1, 1, 1};
  CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("T10 encode failed %d\n", (int)r); return; }
  for (int off : {0, 1024, 128, 256, 384, 896, 2304, 3 * 2304, 5 * 2304 + 128}) {
    CK(cudaMemset(bad, 0, 4));
    tma_unaligned_kernel<<<1, 128, 52 * 1024>>>(tm, off, pw, ph, 3, 1, bad);
    cudaError_t e = cudaDeviceSynchronize();
    int hb = -1;
    if (e == cudaSuccess) CK(cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost));
    printf("T10 SW128 box 64x%dx%d to smem base+%5d: %s, %d / %d chunks wrong\n", pw, ph, off,
           e == cudaSuccess ? "ok" : cudaGetErrorString(e), hb, pw * ph * 8);
    if (e != cudaSuccess) break;
  }
}

// ------------------------------------------------------------------------------------------------ T11
// Nearest 2x upsample folded into a TMA load: 5-D map (c, dx, xs, dy, ys) over the LOW-resolution tensor whose
// dx / dy dimensions (extent 2) have byte stride 0.  Does cuTensorMapEncodeTiled accept stride 0, and does the box
// {64, 2, tw/2, 2, th/2} land as a row-major (2*ys+dy, 2*xs+dx) tile of replicated pixels?
__global__ void __launch_bounds__(128) tma_upsample_kernel(const __grid_constant__ CUtensorMap tm, int xs0, int ys0,
                                                           int tw, int th, int Ws, int* bad) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  for (int i = threadIdx.x; i < 32 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0xFFFFFFFFu;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(smem_u32(&bar), tw * th * 128);
    tma_load_5d(base, &tm, smem_u32(&bar), 0, 0, xs0, 0, ys0);
  }
  mbar_wait(smem_u32(&bar), 0);
  int nbad = 0;
  for (int i = threadIdx.x; i < tw * th * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    const int y = r / tw, x = r - y * tw;  // tile-local high-resolution pixel
    const uint32_t row_addr = base + r * 128;
    const uint32_t a = row_addr + ((j ^ ((row_addr >> 7) & 7)) << 4);
    const __half got = *reinterpret_cast<const __half*>(sm + (a - base));
    const int ys = ys0 + y / 2, xs = xs0 + x / 2;
    const float want = (ys < 0 || xs < 0 || xs >= Ws || ys >= 8) ? 0.f : (float)((ys * Ws + xs) * 8 + j);
    if (__half2float(got) != want) ++nbad;
  }
  atomicAdd(bad, nbad);
}

static void run_t11() {
  const int Ws = 16, Hs = 8, C = 64;
  std::vector<__half> h((size_t)Hs * Ws * C);
  for (int y = 0; y < Hs; ++y)
    for (int x = 0; x < Ws; ++x)
      for (int c = 0; c < C; ++c) h[((size_t)y * Ws + x) * C + c] = __float2half((float)((y * Ws + x) * 8 + (c >> 3)));
  __half* d;
  int* bad;
  CK(cudaMalloc(&d, h.size() * 2));
  CK(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&bad, 4));
  CK(cudaFuncSetAttribute(tma_upsample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  for (int variant = 0; variant < 2; ++variant) {
    const int tw = variant ? 8 : 16, th = variant ? 16 : 8;
    CUtensorMap tm;
    cuuint64_t gd[5] = {(cuuint64_t)C, 2, (cuuint64_t)Ws, 2, (cuuint64_t)Hs};
    cuuint64_t gs[4] = {0, (cuuint64_t)C * 2, 0, (cuuint64_t)Ws * C * 2};
    cuuint32_t bx[5] = {64, 2, (cuuint32_t)tw / 2, 2, (cuuint32_t)th / 2}, es[5] = {1, 1, 1, 1, 1};
    CUresult r = g_encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, d, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("T11 encode 5-D map with stride-0 dx/dy dims, box {64,2,%d,2,%d}: %s (%d)\n", tw / 2, th / 2,
           r == CUDA_SUCCESS ? "accepted" : "REJECTED", (int)r);
    if (r != CUDA_SUCCESS) continue;
    const int coords[3][2] = {{0, 0}, {4, 2}, {12, 6}};  // the last one runs past Ws / Hs: OOB zero fill
    for (auto& c : coords) {
      CK(cudaMemset(bad, 0, 4));
      tma_upsample_kernel<<<1, 128, 40 * 1024>>>(tm, c[0], c[1], tw, th, Ws, bad);
      cudaError_t e = cudaDeviceSynchronize();
      int hb = -1;
      if (e == cudaSuccess) CK(cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost));
      printf("T11 upsampled tile %dx%d at low-res (%d,%d): %s, %d / %d chunks wrong\n", tw, th, c[0], c[1],
             e == cudaSuccess ? "ok" : cudaGetErrorString(e), hb, tw * th * 8);
      if (e != cudaSuccess) return;
    }
  }
}

int main(int argc, char** argv) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  g_encode = (EncodeTiledFn)fn;
  const char* which = argc > 1 ? argv[1] : "1345";
  if (strchr(which, '1')) run_t1();
  if (strchr(which, '3')) run_t3();
  if (strchr(which, '5')) run_t5();
  if (strchr(which, '4')) run_t4();
  if (strchr(which, '6')) run_t6();
  if (strchr(which, '7')) run_t7();
  if (strchr(which, '8')) run_t8();
  if (strchr(which, '9')) run_t9();
  if (strchr(which, 'a')) run_t10();
  if (strchr(which, 'b')) run_t11();
  return 0;
}
