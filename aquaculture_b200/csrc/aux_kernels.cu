// Helper kernels (HBM-bound).  Upstream functions replaced [UPSTREAM-RECALLED; reached from
// `yolov5/detect.py`, REF README.md:77]:
//   preprocess : utils/augmentations.py::letterbox (cv2.resize INTER_LINEAR uint8 + copyMakeBorder(114)),
//                utils/dataloaders.py::LoadImages.__next__ (HWC->CHW, BGR->RGB), detect.py::run (.float()/255)
//   sppf_pool  : models/common.py::SPPF's three chained nn.MaxPool2d(5, 1, 2)
//   upsample2x : nn.Upsample(scale_factor=2, mode='nearest') feeding models/common.py::Concat
#include "aux_kernels.cuh"

namespace aq {
namespace {

// One letterboxed uint8 sample (net coordinates y, x; rgb channel c), bit-exact with cv2:
//   horizontal pass  S = p[x0]*ax0 + p[x1]*ax1              (int32, weights sum to 2048)
//   vertical pass    out = (((ay0*(S0>>4))>>16) + ((ay1*(S1>>4))>>16) + 2) >> 2
__device__ __forceinline__ int letterbox_sample(const PreParams& p, const uint8_t* img, int y, int x, int c) {
  const int ry = y - p.top, rx = x - p.left;
  if (ry < 0 || ry >= p.new_h || rx < 0 || rx >= p.new_w) return 114;
  const int sc = p.bgr ? 2 - c : c;
  if (!p.resize) return img[((size_t)ry * p.in_w + rx) * 3 + sc];
  const int xa = __ldg(p.x0 + rx), xb = __ldg(p.x1 + rx);
  const int wa = __ldg(p.ax0 + rx), wb = __ldg(p.ax1 + rx);
  const int ya = __ldg(p.y0 + ry), yb = __ldg(p.y1 + ry);
  const int va = __ldg(p.ay0 + ry), vb = __ldg(p.ay1 + ry);
  const uint8_t* r0 = img + (size_t)ya * p.in_w * 3;
  const uint8_t* r1 = img + (size_t)yb * p.in_w * 3;
  const int s0 = __ldg(r0 + xa * 3 + sc) * wa + __ldg(r0 + xb * 3 + sc) * wb;
  const int s1 = __ldg(r1 + xa * 3 + sc) * wa + __ldg(r1 + xb * 3 + sc) * wb;
  return (((va * (s0 >> 4)) >> 16) + ((vb * (s1 >> 4)) >> 16) + 2) >> 2;
}

// One thread per space-to-depth output pixel: 2x2 source pixels x 3 channels -> 16 fp16 (two 16-byte stores).
// u / 255 -> fp16 comes from a 256-entry shared-memory table built with the same IEEE division + rounding upstream's
// `.half() / 255` path produces (bit-identical, 12 divisions per thread saved).  Tiles that need no resize and
// no padding (the reference's 640x640 case) take a vectorised path: per source row the thread's two pixels are
// six consecutive bytes, read as three aligned 16-bit loads (a warp reads 192 contiguous bytes per row).
__global__ void __launch_bounds__(256) preprocess_kernel(const PreParams p) {
  __shared__ __half lut[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = __float2half_rn(__fdiv_rn((float)i, 255.0f));
  __syncthreads();
  const int W2 = p.net_w >> 1, H2 = p.net_h >> 1;
  const size_t total = (size_t)p.n * H2 * W2;
  const bool direct = !p.resize && p.top == 0 && p.left == 0 && p.new_h == p.net_h && p.new_w == p.net_w && !(p.in_w & 1);
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int X = (int)(idx % W2);
    const size_t t = idx / W2;
    const int Y = (int)(t % H2);
    const int n = (int)(t / H2);
    const uint8_t* img = p.src + (size_t)n * p.in_h * p.in_w * 3;
    __align__(16) __half v[16];
    if (direct) {
#pragma unroll
      for (int py = 0; py < 2; ++py) {
        const uint16_t* row = reinterpret_cast<const uint16_t*>(img + ((size_t)(2 * Y + py) * p.in_w + 2 * X) * 3);
        const uint32_t w0 = __ldg(row), w1 = __ldg(row + 1), w2 = __ldg(row + 2);
        // pixel 0 = bytes 0..2, pixel 1 = bytes 3..5 of the row segment; bgr swaps the outer bytes of each pixel
        const uint32_t a0 = w0 & 255u, a1 = w0 >> 8, a2 = w1 & 255u, a3 = w1 >> 8, a4 = w2 & 255u, a5 = w2 >> 8;
        v[py * 6 + 0] = lut[p.bgr ? a2 : a0];
        v[py * 6 + 1] = lut[a1];
        v[py * 6 + 2] = lut[p.bgr ? a0 : a2];
        v[py * 6 + 3] = lut[p.bgr ? a5 : a3];
        v[py * 6 + 4] = lut[a4];
        v[py * 6 + 5] = lut[p.bgr ? a3 : a5];
      }
    } else {
#pragma unroll
      for (int py = 0; py < 2; ++py)
#pragma unroll
        for (int px = 0; px < 2; ++px)
#pragma unroll
          for (int c = 0; c < 3; ++c) v[(py * 2 + px) * 3 + c] = lut[letterbox_sample(p, img, 2 * Y + py, 2 * X + px, c)];
    }
#pragma unroll
    for (int c = 12; c < 16; ++c) v[c] = __float2half_rn(0.f);
    uint4* dst = reinterpret_cast<uint4*>(p.out_s2d + idx * 16);
    dst[0] = reinterpret_cast<const uint4*>(v)[0];
    dst[1] = reinterpret_cast<const uint4*>(v)[1];
  }
}

__global__ void __launch_bounds__(256) preprocess_u8_kernel(const PreParams p) {
  const size_t total = (size_t)p.n * p.net_h * p.net_w * 3;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % 3);
    size_t t = idx / 3;
    const int x = (int)(t % p.net_w);
    t /= p.net_w;
    const int y = (int)(t % p.net_h);
    const int n = (int)(t / p.net_h);
    p.out_u8[idx] = (uint8_t)letterbox_sample(p, p.src + (size_t)n * p.in_h * p.in_w * 3, y, x, c);
  }
}

// ---------------------------------------------------------------- SPPF pooling
// grid (c / CG, n); the CTA keeps an (h*w) x CG-channel slab in shared memory and runs three separable
// 5x5 max pools (horizontal then vertical), writing each result to its concat slice.
// kPoolV = 16-byte vectors per pixel per CTA: 4 (64 contiguous bytes = two full 32-byte sectors) while the slab fits
// shared memory (20x20 maps of the 640 configs), 1 for larger maps (40x40 at --img 1280, up to 4266 pixels).

__device__ __forceinline__ uint4 hmax8(uint4 a, uint4 b) {
  uint4 r;
  const __half2* pa = reinterpret_cast<const __half2*>(&a);
  const __half2* pb = reinterpret_cast<const __half2*>(&b);
  __half2* pr = reinterpret_cast<__half2*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) pr[i] = __hmax2(pa[i], pb[i]);
  return r;
}

template <int kPoolV>
__global__ void __launch_bounds__(256) sppf_pool_kernel(__half* buf, int h, int w, int ct, int c_off, int c) {
  constexpr int kPoolCG = 8 * kPoolV;  // channels per CTA
  extern __shared__ uint4 s_pool[];
  const int hw = h * w, nv = hw * kPoolV;  // element i = pixel (i / kPoolV), vector (i % kPoolV)
  uint4* s0 = s_pool;
  uint4* s1 = s_pool + nv;
  uint4* tmp = s_pool + 2 * nv;
  const int cg = blockIdx.x;
  const int n = blockIdx.y;
  __half* base = buf + (size_t)n * hw * ct + c_off + cg * kPoolCG;
  for (int i = threadIdx.x; i < nv; i += blockDim.x)
    s0[i] = *reinterpret_cast<const uint4*>(base + (size_t)(i / kPoolV) * ct + (i % kPoolV) * 8);
  __syncthreads();
  uint4* src = s0;
  uint4* dst = s1;
  for (int pass = 1; pass <= 3; ++pass) {
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
      const int pix = i / kPoolV;
      const int y = pix / w, x = pix - y * w;
      uint4 m = src[i];
      for (int d = 1; d <= 2; ++d) {
        if (x - d >= 0) m = hmax8(m, src[i - d * kPoolV]);
        if (x + d < w) m = hmax8(m, src[i + d * kPoolV]);
      }
      tmp[i] = m;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
      const int pix = i / kPoolV;
      const int y = pix / w;
      uint4 m = tmp[i];
      for (int d = 1; d <= 2; ++d) {
        if (y - d >= 0) m = hmax8(m, tmp[i - d * w * kPoolV]);
        if (y + d < h) m = hmax8(m, tmp[i + d * w * kPoolV]);
      }
      dst[i] = m;
      *reinterpret_cast<uint4*>(base + (size_t)pix * ct + (i % kPoolV) * 8 + (size_t)pass * c) = m;
    }
    __syncthreads();
    uint4* t = src; src = dst; dst = t;
  }
}

__global__ void __launch_bounds__(256) upsample2x_kernel(const __half* __restrict__ src, int n, int h, int w,
                                                         int src_ct, int src_c_off, int c, __half* __restrict__ dst,
                                                         int dst_ct, int dst_c_off) {
  const int cv = c >> 3;  // 16-byte vectors per pixel
  const int ho = 2 * h, wo = 2 * w;
  const size_t total = (size_t)n * ho * wo * cv;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int v = (int)(idx % cv);
    size_t t = idx / cv;
    const int x = (int)(t % wo);
    t /= wo;
    const int y = (int)(t % ho);
    const int b = (int)(t / ho);
    const uint4 val = __ldg(reinterpret_cast<const uint4*>(
        src + (((size_t)b * h + (y >> 1)) * w + (x >> 1)) * src_ct + src_c_off + v * 8));
    *reinterpret_cast<uint4*>(dst + (((size_t)b * ho + y) * wo + x) * dst_ct + dst_c_off + v * 8) = val;
  }
}

inline int grid_for(size_t total, int block) {
  size_t g = (total + block - 1) / block;
  const size_t cap = 148 * 16;
  return (int)(g < cap ? (g ? g : 1) : cap);
}

}  // namespace

cudaError_t launch_preprocess(const PreParams& p, cudaStream_t stream) {
  const size_t total = (size_t)p.n * (p.net_h / 2) * (p.net_w / 2);
  preprocess_kernel<<<grid_for(total, 256), 256, 0, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_preprocess_u8(const PreParams& p, cudaStream_t stream) {
  const size_t total = (size_t)p.n * p.net_h * p.net_w * 3;
  preprocess_u8_kernel<<<grid_for(total, 256), 256, 0, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t sppf_init() {
  cudaError_t e = cudaFuncSetAttribute(sppf_pool_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(sppf_pool_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

cudaError_t launch_sppf_pool(__half* buf, int n, int h, int w, int ct, int c_off, int c, cudaStream_t stream) {
  const size_t smem1 = (size_t)3 * h * w * sizeof(uint4);  // one 8-channel vector per pixel
  if (4 * smem1 <= 200 * 1024 && (c % 32) == 0) {
    sppf_pool_kernel<4><<<dim3(c / 32, n), 256, 4 * smem1, stream>>>(buf, h, w, ct, c_off, c);
  } else {
    if (smem1 > 200 * 1024 || (c % 8) != 0) return cudaErrorInvalidValue;  // maps beyond 4266 pixels: not supported
    sppf_pool_kernel<1><<<dim3(c / 8, n), 256, smem1, stream>>>(buf, h, w, ct, c_off, c);
  }
  return cudaGetLastError();
}

cudaError_t launch_upsample2x(const __half* src, int n, int h, int w, int src_ct, int src_c_off, int c, __half* dst,
                              int dst_ct, int dst_c_off, cudaStream_t stream) {
  const size_t total = (size_t)n * 4 * h * w * (c / 8);
  upsample2x_kernel<<<grid_for(total, 256), 256, 0, stream>>>(src, n, h, w, src_ct, src_c_off, c, dst, dst_ct,
                                                              dst_c_off);
  return cudaGetLastError();
}

}  // namespace aq
