// See conv_igemm.cuh for the design.  sm_100a only (tcgen05 / TMEM / TMA).
#include "conv_igemm.cuh"
#include "ptx.cuh"

namespace aq {

using namespace ptx;

namespace {

struct TileCoord {
  int m_tile, n_tile;
  int n, y0, x0;  // spatial modes
  int m0;         // flat mode
};

__device__ __forceinline__ TileCoord decode_tile(const ConvKParams& p, int tile) {
  TileCoord t;
  t.m_tile = tile / p.n_tiles_n;
  t.n_tile = tile - t.m_tile * p.n_tiles_n;
  t.m0 = t.m_tile * kTileM;
  if (p.mode != CONV_MODE_FLAT) {
    const int per_img = p.tiles_x * p.tiles_y;
    t.n = t.m_tile / per_img;
    const int r = t.m_tile - t.n * per_img;
    const int ty = r / p.tiles_x;
    t.y0 = ty * p.th;
    t.x0 = (r - ty * p.tiles_x) * p.tw;
  } else {
    t.n = t.y0 = t.x0 = 0;
  }
  return t;
}

// Byte offset of 16-byte chunk (row r, chunk j) inside a swizzled [128 x ob] fp16 staging group
// (TMA swizzle modes 32B/64B/128B = XOR of address bits [4,4+B) with bits [7,7+B)).
__device__ __forceinline__ uint32_t swz_off(uint32_t r, uint32_t j, uint32_t pitch, uint32_t mask) {
  uint32_t a = r * pitch + (j << 4);
  return a ^ (((a >> 7) & mask) << 4);
}

__device__ __forceinline__ float silu(float v) { return __fdividef(v, 1.0f + __expf(-v)); }

template <int EPI>
__global__ void __launch_bounds__(kConvThreads, 1) conv_igemm_kernel(const __grid_constant__ ConvKParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * kMaxStages + 6];
  __shared__ uint32_t tmem_slot;

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t c_base = smem_base + p.num_stages * p.stage_bytes;

  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * kMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * kMaxStages + 2 + a); };
  const uint32_t res_full = bar0 + 8u * (2 * kMaxStages + 4);
  const uint32_t res_empty = bar0 + 8u * (2 * kMaxStages + 5);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmA);
    prefetch_tmap(&p.tmB);
    if (EPI == EPI_STORE) {
      prefetch_tmap(&p.tmOut);
      if (p.has_res) prefetch_tmap(&p.tmRes);
    }
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < p.num_stages; ++s) {
        mbar_init(full_bar(s), 1);
        mbar_init(empty_bar(s), 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(tfull_bar(a), 1);
        mbar_init(tempty_bar(a), 4);
      }
      mbar_init(res_full, 1);
      mbar_init(res_empty, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(smem_u32(&tmem_slot), p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  const int total_tiles = p.m_tiles * p.n_tiles_n;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const TileCoord t = decode_tile(p, tile);
        const int n0 = t.n_tile * p.tile_n;
        if (EPI == EPI_STORE && p.has_res) {
          // residual tile -> output staging buffer, once the previous tile's store has drained it
          mbar_wait(res_empty, (it & 1) ^ 1);
          mbar_arrive_expect_tx(res_full, (p.a_box_bytes / (p.cb * 2)) * p.tile_n * 2);
          const uint32_t grp_bytes = kTileM * p.ob * 2;
          for (int g = 0; g < p.tile_n / p.ob; ++g) {
            if (p.mode == CONV_MODE_FLAT)
              tma_load_2d(c_base + g * grp_bytes, &p.tmRes, res_full, n0 + g * p.ob, t.m0);
            else
              tma_load_4d(c_base + g * grp_bytes, &p.tmRes, res_full, n0 + g * p.ob, t.x0, t.y0, t.n);
          }
        }
        int chunk = 0;
        for (int ks = 0; ks < p.kstages; ++ks) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const int nch = min(p.cps, p.total_chunks - chunk);
          mbar_arrive_expect_tx(full_bar(stage), nch * (p.a_box_bytes + p.b_box_bytes));
          const uint32_t sa = smem_base + stage * p.stage_bytes;
          const uint32_t sb = sa + p.cps * p.a_sub_bytes;
          for (int j = 0; j < nch; ++j, ++chunk) {
            const int tap = chunk / p.cpt;
            const int c0 = (chunk - tap * p.cpt) * p.cb;
            const int kh = tap / p.ksize;
            const int kw = tap - kh * p.ksize;
            const uint32_t dst = sa + j * p.a_sub_bytes;
            if (p.mode == CONV_MODE_FLAT) {
              tma_load_2d(dst, &p.tmA, full_bar(stage), c0, t.m0);
            } else if (p.mode == CONV_MODE_S1) {
              tma_load_4d(dst, &p.tmA, full_bar(stage), c0, t.x0 + kw - p.pad, t.y0 + kh - p.pad, t.n);
            } else {
              // input row 2*Y + kh - 1: kh=0 -> (parity 1, Y-1), kh=1 -> (0, Y), kh=2 -> (1, Y)
              const int yp = (kh + 1) & 1, dy = (kh == 0) ? -1 : 0;
              const int xp = (kw + 1) & 1, dx = (kw == 0) ? -1 : 0;
              tma_load_5d(dst, &p.tmA, full_bar(stage), xp * p.src_ct + p.src_c_off + c0, t.x0 + dx, yp,
                          t.y0 + dy, t.n);
            }
            tma_load_2d(sb + j * p.b_sub_bytes, &p.tmB, full_bar(stage), tap * p.cpt * p.cb + c0, n0);
          }
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer (one thread) ===========================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_f16(kTileM, p.tile_n);
      const uint32_t swz = p.cb * 2;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * p.acc_stride;
        int chunk = 0;
        uint32_t accumulate = 0;
        for (int ks = 0; ks < p.kstages; ++ks) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const int nch = min(p.cps, p.total_chunks - chunk);
          const uint32_t sa = smem_base + stage * p.stage_bytes;
          const uint32_t sb = sa + p.cps * p.a_sub_bytes;
          for (int j = 0; j < nch; ++j) {
            for (uint32_t kk = 0; kk < swz; kk += 32) {  // UMMA K = 16 fp16 = 32 bytes
              umma_f16(d_tmem, make_kmajor_desc(sa + j * p.a_sub_bytes + kk, swz),
                       make_kmajor_desc(sb + j * p.b_sub_bytes + kk, swz), idesc, accumulate);
              accumulate = 1;
            }
          }
          chunk += nch;
          umma_commit(empty_bar(stage));  // smem slot reusable once these MMAs retire
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(tfull_bar(as));  // accumulator ready for the epilogue
      }
    }
  } else {
    // =========================== epilogue warps 2..5 ===========================
    const uint32_t q = warp & 3;        // TMEM lane quarter this warp may access
    const uint32_t row = q * 32 + lane;  // accumulator row == TMEM lane
    const bool store_thread = (warp == 2 && lane == 0);
    const uint32_t pitch = p.ob * 2;
    const uint32_t mask = (pitch >> 4) - 1;
    const uint32_t grp_bytes = kTileM * pitch;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const TileCoord t = decode_tile(p, tile);
      const int n0 = t.n_tile * p.tile_n;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((q * 32u) << 16) + as * p.acc_stride;

      if (EPI == EPI_STORE) {
        if (p.has_res) {
          mbar_wait(res_full, it & 1);
        } else if (it > 0) {
          if (store_thread) tma_store_wait_read0();  // previous tile's store has finished reading smem
          named_bar_sync(1, 128);
        }
        for (int c0 = 0; c0 < p.tile_n; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(t_addr + c0, v);
          tmem_ld_wait();
          const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + c0);
          float f[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 b = __ldg(bp + i);
            f[4 * i + 0] = __uint_as_float(v[4 * i + 0]) + b.x;
            f[4 * i + 1] = __uint_as_float(v[4 * i + 1]) + b.y;
            f[4 * i + 2] = __uint_as_float(v[4 * i + 2]) + b.z;
            f[4 * i + 3] = __uint_as_float(v[4 * i + 3]) + b.w;
          }
          if (p.act) {
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = silu(f[i]);
          }
          const int g = c0 / p.ob;
          const uint32_t j = (uint32_t)(c0 - g * p.ob) >> 3;
          const uint32_t a0 = c_base + g * grp_bytes + swz_off(row, j, pitch, mask);
          const uint32_t a1 = c_base + g * grp_bytes + swz_off(row, j + 1, pitch, mask);
          if (p.has_res) {
            uint32_t r[8];
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a0));
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(a1));
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float2 rr = __half22float2(*reinterpret_cast<const __half2*>(&r[i]));
              f[2 * i] += rr.x;
              f[2 * i + 1] += rr.y;
            }
          }
          uint32_t h[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const __half2 hh = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
            h[i] = *reinterpret_cast<const uint32_t*>(&hh);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a0), "r"(h[0]), "r"(h[1]), "r"(h[2]), "r"(h[3])
                       : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a1), "r"(h[4]), "r"(h[5]), "r"(h[6]), "r"(h[7])
                       : "memory");
        }
        // accumulator stage drained -> MMA warp may overwrite it
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(as));
        // make the staging writes visible to the async proxy, then one thread stores the tile
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (store_thread) {
          for (int g = 0; g < p.tile_n / p.ob; ++g) {
            if (p.mode == CONV_MODE_FLAT)
              tma_store_2d(&p.tmOut, c_base + g * grp_bytes, n0 + g * p.ob, t.m0);
            else
              tma_store_4d(&p.tmOut, c_base + g * grp_bytes, n0 + g * p.ob, t.x0, t.y0, t.n);
          }
          tma_store_commit();
          if (p.has_res) {
            tma_store_wait_read0();
            mbar_arrive(res_empty);
          }
        }
      } else {
        // ---- Detect decode: row = one grid cell, columns = 3 anchors x (5 + nc)
        const int m = t.m0 + (int)row;
        const int cells = p.det_ny * p.det_nx;
        const int b = m / cells;
        const int pos = m - b * cells;
        const int gy = pos / p.det_nx;
        const int gx = pos - gy * p.det_nx;
        const bool valid = m < p.det_m_total;
        const int ncols = 3 * p.det_no;
        for (int c0 = 0; c0 < p.tile_n; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(t_addr + c0, v);
          tmem_ld_wait();
          if (valid) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int col = c0 + i;
              if (col < ncols) {
                const int a = col / p.det_no;
                const int o = col - a * p.det_no;
                const float z = __uint_as_float(v[i]) + __ldg(p.bias + col);
                const float s = __fdividef(1.0f, 1.0f + __expf(-z));
                float out;
                if (o == 0) out = (s * 2.0f + ((float)gx - 0.5f)) * p.det_stride;
                else if (o == 1) out = (s * 2.0f + ((float)gy - 0.5f)) * p.det_stride;
                else if (o == 2) out = (s * 2.0f) * (s * 2.0f) * p.det_anchor[2 * a];
                else if (o == 3) out = (s * 2.0f) * (s * 2.0f) * p.det_anchor[2 * a + 1];
                else out = s;
                const size_t r = (size_t)b * p.det_rows + p.det_level_off + (size_t)a * cells + pos;
                p.pred[r * p.det_no + o] = out;
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(as));
      }
    }
    if (EPI == EPI_STORE && store_thread) tma_store_wait0();
  }

  // teardown: everyone done with TMEM before the allocating warp frees it
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

}  // namespace

size_t conv_smem_bytes(const ConvKParams& p) {
  return (size_t)p.num_stages * p.stage_bytes + p.c_bytes + 1024;
}

cudaError_t conv_kernels_init() {
  // 227 KB per CTA includes the kernel's static shared memory (barriers): leave 1 KB for it
  cudaError_t e = cudaFuncSetAttribute(conv_igemm_kernel<EPI_STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kConvMaxDynSmem);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(conv_igemm_kernel<EPI_DETECT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              kConvMaxDynSmem);
}

cudaError_t launch_conv(const ConvKParams& p, int epi, int num_sms, cudaStream_t stream) {
  const int total = p.m_tiles * p.n_tiles_n;
  const int grid = total < num_sms ? total : num_sms;
  const size_t smem = conv_smem_bytes(p);
  if (epi == EPI_STORE)
    conv_igemm_kernel<EPI_STORE><<<grid, kConvThreads, smem, stream>>>(p);
  else
    conv_igemm_kernel<EPI_DETECT><<<grid, kConvThreads, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace aq
