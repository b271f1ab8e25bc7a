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
__device__ __forceinline__ const uint8_t* tile_base(const PreParams& p, int n) {
  if (!p.windowed) return p.src + (size_t)n * p.in_h * p.row_pitch;
  const int w = p.win_list ? __ldg(p.win_list + n) : p.win0 + n, ix = w / p.wins_y, iy = w - ix * p.wins_y;
  return p.src + (size_t)iy * p.in_h * p.row_pitch + (size_t)ix * p.in_w * 3;
}

__device__ __forceinline__ int letterbox_sample(const PreParams& p, const uint8_t* img, int y, int x, int c) {
  const int ry = y - p.top, rx = x - p.left;
  if (ry < 0 || ry >= p.new_h || rx < 0 || rx >= p.new_w) return 114;
  const int sc = p.bgr ? 2 - c : c;
  if (!p.resize) return img[(size_t)ry * p.row_pitch + rx * 3 + sc];
  const int xa = __ldg(p.x0 + rx), xb = __ldg(p.x1 + rx);
  const int wa = __ldg(p.ax0 + rx), wb = __ldg(p.ax1 + rx);
  const int ya = __ldg(p.y0 + ry), yb = __ldg(p.y1 + ry);
  const int va = __ldg(p.ay0 + ry), vb = __ldg(p.ay1 + ry);
  const uint8_t* r0 = img + (size_t)ya * p.row_pitch;
  const uint8_t* r1 = img + (size_t)yb * p.row_pitch;
  const int s0 = __ldg(r0 + xa * 3 + sc) * wa + __ldg(r0 + xb * 3 + sc) * wb;
  const int s1 = __ldg(r1 + xa * 3 + sc) * wa + __ldg(r1 + xb * 3 + sc) * wb;
  return (((va * (s0 >> 4)) >> 16) + ((vb * (s1 >> 4)) >> 16) + 2) >> 2;
}

// One thread per space-to-depth output pixel: 2x2 source pixels x 3 channels -> 16 fp16 (two 16-byte stores).
// u / 255 -> fp16 is evaluated as half_rn(float(u) * (1.0f / 255.0f)): for every u in 0..255 this equals both
// half_rn(u / 255.0f) and upstream's `.half() / 255` (checked exhaustively, tests/test_oracle.py), so neither a
// division nor a table lookup is needed.  Tiles that need no resize and no padding (the reference's 640x640 case)
// take a vectorised path: per source row the thread's two pixels are six consecutive bytes, read as three aligned
// 16-bit loads (a warp reads 192 contiguous bytes per row); two pixels per thread are in flight at once.
__device__ __forceinline__ uint32_t norm2(uint32_t lo, uint32_t hi) {
  const float k = 1.0f / 255.0f;
  const __half2 h = __floats2half2_rn(__fmul_rn((float)lo, k), __fmul_rn((float)hi, k));
  return *reinterpret_cast<const uint32_t*>(&h);
}

struct RowSeg { uint32_t w0, w1, w2; };

__device__ __forceinline__ void emit_direct(__half* out, size_t idx, const RowSeg (&r)[2], int bgr) {
  uint32_t b[2][6];
#pragma unroll
  for (int py = 0; py < 2; ++py) {
    // pixel 0 = bytes 0..2, pixel 1 = bytes 3..5 of the row segment; bgr swaps the outer bytes of each pixel
    const uint32_t a0 = r[py].w0 & 255u, a1 = r[py].w0 >> 8, a2 = r[py].w1 & 255u, a3 = r[py].w1 >> 8,
                   a4 = r[py].w2 & 255u, a5 = r[py].w2 >> 8;
    b[py][0] = bgr ? a2 : a0; b[py][1] = a1; b[py][2] = bgr ? a0 : a2;
    b[py][3] = bgr ? a5 : a3; b[py][4] = a4; b[py][5] = bgr ? a3 : a5;
  }
  uint4 lo, hi;
  lo.x = norm2(b[0][0], b[0][1]); lo.y = norm2(b[0][2], b[0][3]); lo.z = norm2(b[0][4], b[0][5]);
  lo.w = norm2(b[1][0], b[1][1]); hi.x = norm2(b[1][2], b[1][3]); hi.y = norm2(b[1][4], b[1][5]);
  hi.z = 0u; hi.w = 0u;
  uint4* dst = reinterpret_cast<uint4*>(out + idx * 16);
  __stcs(dst, lo);   // streamed: the stem conv reads it once
  __stcs(dst + 1, hi);
}

__global__ void __launch_bounds__(256) preprocess_kernel(const PreParams p) {
  const int W2 = p.net_w >> 1, H2 = p.net_h >> 1;
  const size_t total = (size_t)p.n * H2 * W2;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const bool direct = !p.windowed && !p.resize && p.top == 0 && p.left == 0 && p.new_h == p.net_h && p.new_w == p.net_w && !(p.in_w & 1);
  if (direct) {
    // in_h == net_h, in_w == net_w: s2d pixel idx = (n, Y, X) reads bytes [6 * idx' ..) of rows 2Y and 2Y+1
    const size_t row_bytes = (size_t)p.in_w * 3;
    auto load = [&](size_t idx, RowSeg (&r)[2]) {
      const int X = (int)(idx % W2);
      const size_t t = idx / W2;  // = n * H2 + Y; image n starts at row n * in_h = 2 * n * H2
      const uint16_t* row = reinterpret_cast<const uint16_t*>(p.src + 2 * t * row_bytes + (size_t)X * 6);
#pragma unroll
      for (int py = 0; py < 2; ++py) {
        const uint16_t* q = reinterpret_cast<const uint16_t*>(reinterpret_cast<const uint8_t*>(row) + py * row_bytes);
        r[py].w0 = __ldg(q); r[py].w1 = __ldg(q + 1); r[py].w2 = __ldg(q + 2);
      }
    };
    size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; idx + stride < total; idx += 2 * stride) {
      RowSeg ra[2], rb[2];
      load(idx, ra);
      load(idx + stride, rb);
      emit_direct(p.out_s2d, idx, ra, p.bgr);
      emit_direct(p.out_s2d, idx + stride, rb, p.bgr);
    }
    if (idx < total) {
      RowSeg ra[2];
      load(idx, ra);
      emit_direct(p.out_s2d, idx, ra, p.bgr);
    }
    return;
  }
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const int X = (int)(idx % W2);
    const size_t t = idx / W2;
    const int Y = (int)(t % H2);
    const int n = (int)(t / H2);
    const uint8_t* img = tile_base(p, n);
    uint32_t b[12];
#pragma unroll
    for (int py = 0; py < 2; ++py)
#pragma unroll
      for (int px = 0; px < 2; ++px)
#pragma unroll
        for (int c = 0; c < 3; ++c) b[(py * 2 + px) * 3 + c] = (uint32_t)letterbox_sample(p, img, 2 * Y + py, 2 * X + px, c);
    uint4 lo, hi;
    lo.x = norm2(b[0], b[1]); lo.y = norm2(b[2], b[3]); lo.z = norm2(b[4], b[5]); lo.w = norm2(b[6], b[7]);
    hi.x = norm2(b[8], b[9]); hi.y = norm2(b[10], b[11]); hi.z = 0u; hi.w = 0u;
    uint4* dst = reinterpret_cast<uint4*>(p.out_s2d + idx * 16);
    dst[0] = lo;
    dst[1] = hi;
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
    p.out_u8[idx] = (uint8_t)letterbox_sample(p, tile_base(p, n), y, x, c);
  }
}

// ---------------------------------------------------------------- blank-tile statistics
// grid (ceil(in_h / kStatRows), n), 256 threads.  A CTA walks kStatRows rows of one tile: thread t owns columns
// t, t + 256, ... (a warp reads 96 contiguous bytes per row), keeps their partial column sums in registers, and the
// CTA reduces each row's sum (blank-row test) and the luminance extrema.  Column sums meet in `col_sums` (one
// atomicAdd per column per CTA); tile_stats_finish counts the blank columns.
constexpr int kStatRows = 16;
constexpr int kStatColSlots = 8;  // columns per thread: tiles up to 2048 pixels wide

__global__ void __launch_bounds__(256) tile_stats_kernel(const PreParams p, unsigned int* col_sums, TileStats* stats) {
  __shared__ unsigned int s_row[kStatRows];
  __shared__ int s_min, s_max;
  const int n = blockIdx.y, y0 = blockIdx.x * kStatRows;
  const uint8_t* img = tile_base(p, n);
  if (threadIdx.x < kStatRows) s_row[threadIdx.x] = 0;
  if (threadIdx.x == 0) { s_min = 255; s_max = 0; }
  __syncthreads();
  unsigned int col[kStatColSlots];
#pragma unroll
  for (int k = 0; k < kStatColSlots; ++k) col[k] = 0;
  int lmin = 255, lmax = 0;
  const int rows = min(kStatRows, p.in_h - y0);
  for (int r = 0; r < rows; ++r) {
    const uint8_t* row = img + (size_t)(y0 + r) * p.row_pitch;
    unsigned int rs = 0;
#pragma unroll
    for (int k = 0; k < kStatColSlots; ++k) {
      const int x = threadIdx.x + 256 * k;
      if (x < p.in_w) {
        const unsigned int c0 = __ldg(row + 3 * x), c1 = __ldg(row + 3 * x + 1), c2 = __ldg(row + 3 * x + 2);
        const unsigned int R = p.bgr ? c2 : c0, B = p.bgr ? c0 : c2;
        const int L = (int)((R * 19595u + c1 * 38470u + B * 7471u + 0x8000u) >> 16);  // PIL ImageConvert rgb2l
        lmin = min(lmin, L);
        lmax = max(lmax, L);
        const unsigned int s3 = c0 + c1 + c2;
        col[k] += s3;
        rs += s3;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rs += __shfl_xor_sync(0xffffffffu, rs, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_row[r], rs);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lmin = min(lmin, __shfl_xor_sync(0xffffffffu, lmin, o));
    lmax = max(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicMin(&s_min, lmin); atomicMax(&s_max, lmax); }
#pragma unroll
  for (int k = 0; k < kStatColSlots; ++k) {
    const int x = threadIdx.x + 256 * k;
    if (x < p.in_w) atomicAdd(&col_sums[(size_t)n * p.in_w + x], col[k]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int blank = 0;
    for (int r = 0; r < rows; ++r) blank += s_row[r] >= 750u * (unsigned int)p.in_w;  // mean over (W, 3) >= 250
    if (blank) atomicAdd(&stats[n].blank_rows, blank);
    atomicMin(&stats[n].l_min, s_min);
    atomicMax(&stats[n].l_max, s_max);
  }
}

// Vectorised variant for tiles whose rows are 16-byte aligned (in_w % 16 == 0, pitch % 16 == 0): an item = 16 pixels =
// 48 bytes = three 16-byte loads; the CTA walks kStatBand rows, items strided over its threads; column sums meet in
// shared memory (conflict-free: the lanes of a warp hold neighbouring pixel groups), one global atomicAdd per column
// per CTA.  grid (ceil(in_h / kStatBand), n).
constexpr int kStatBand = 32;

// REGCOL (256 % G == 0, e.g. 1024-wide tiles): a thread meets the same pixel group in every row it visits, so its 16
// column sums stay in registers until the end of the band.
template <bool REGCOL>
__global__ void __launch_bounds__(256) tile_stats_vec_kernel(const PreParams p, unsigned int* col_sums, TileStats* stats) {
  extern __shared__ unsigned int s_dyn[];  // [in_w] column sums, then [kStatBand] row sums
  unsigned int* s_col = s_dyn;
  unsigned int* s_rowv = s_dyn + p.in_w;
  __shared__ int s_min, s_max;
  const int n = blockIdx.y, y0 = blockIdx.x * kStatBand;
  const uint8_t* img = tile_base(p, n);
  for (int i = threadIdx.x; i < p.in_w + kStatBand; i += blockDim.x) s_dyn[i] = 0;
  if (threadIdx.x == 0) { s_min = 255; s_max = 0; }
  __syncthreads();
  const int G = p.in_w >> 4;
  const int rows = min(kStatBand, p.in_h - y0);
  const int items = rows * G;
  int lmin = 255, lmax = 0;
  unsigned int col[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) col[k] = 0;
  for (int it = threadIdx.x; it - (int)(threadIdx.x & 31) < items; it += blockDim.x) {  // whole warps stay in the loop
    const bool live = it < items;
    const int r = live ? it / G : -1;
    const int g = live ? it - r * G : 0;
    unsigned int rs = 0;
    if (live) {
      const uint4* q = reinterpret_cast<const uint4*>(img + (size_t)(y0 + r) * p.row_pitch + (size_t)g * 48);
      const uint4 v0 = __ldg(q), v1 = __ldg(q + 1), v2 = __ldg(q + 2);
      const unsigned int w[12] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};
#pragma unroll
      for (int px = 0; px < 16; ++px) {
        const int b0 = 3 * px, b1 = b0 + 1, b2 = b0 + 2;
        const unsigned int c0 = (w[b0 >> 2] >> (8 * (b0 & 3))) & 255u, c1 = (w[b1 >> 2] >> (8 * (b1 & 3))) & 255u,
                           c2 = (w[b2 >> 2] >> (8 * (b2 & 3))) & 255u;
        const unsigned int R = p.bgr ? c2 : c0, B = p.bgr ? c0 : c2;
        const int L = (int)((R * 19595u + c1 * 38470u + B * 7471u + 0x8000u) >> 16);
        lmin = min(lmin, L);
        lmax = max(lmax, L);
        const unsigned int s3 = c0 + c1 + c2;
        rs += s3;
        if (REGCOL) col[px] += s3;
        else atomicAdd(&s_col[px * G + g], s3);  // [px][g]: the lanes of a warp (consecutive g) hit consecutive banks
      }
    }
    // row sum: one shared atomic per warp when its lanes sit in one row, else one per lane
    const int r0 = __shfl_sync(0xffffffffu, r, 0);
    if (__all_sync(0xffffffffu, r == r0)) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) rs += __shfl_xor_sync(0xffffffffu, rs, o);
      if ((threadIdx.x & 31) == 0 && r0 >= 0) atomicAdd(&s_rowv[r0], rs);
    } else if (live) {
      atomicAdd(&s_rowv[r], rs);
    }
  }
  if (REGCOL) {
    const int g = (int)threadIdx.x % G;
#pragma unroll
    for (int k = 0; k < 16; ++k) atomicAdd(&s_col[k * G + g], col[k]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lmin = min(lmin, __shfl_xor_sync(0xffffffffu, lmin, o));
    lmax = max(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicMin(&s_min, lmin); atomicMax(&s_max, lmax); }
  __syncthreads();
  for (int x = threadIdx.x; x < p.in_w; x += blockDim.x) atomicAdd(&col_sums[(size_t)n * p.in_w + x], s_col[(x & 15) * G + (x >> 4)]);
  if (threadIdx.x == 0) {
    int blank = 0;
    for (int r = 0; r < rows; ++r) blank += s_rowv[r] >= 750u * (unsigned int)p.in_w;
    if (blank) atomicAdd(&stats[n].blank_rows, blank);
    atomicMin(&stats[n].l_min, s_min);
    atomicMax(&stats[n].l_max, s_max);
  }
}

__global__ void __launch_bounds__(256) tile_stats_init(TileStats* stats, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) stats[i] = TileStats{255, 0, 0, 0};
}

__global__ void __launch_bounds__(256) tile_stats_finish(const unsigned int* col_sums, TileStats* stats, int in_h, int in_w) {
  const int n = blockIdx.x;
  int blank = 0;
  for (int x = threadIdx.x; x < in_w; x += blockDim.x) blank += col_sums[(size_t)n * in_w + x] >= 750u * (unsigned int)in_h;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) blank += __shfl_xor_sync(0xffffffffu, blank, o);
  if ((threadIdx.x & 31) == 0 && blank) atomicAdd(&stats[n].blank_cols, blank);
}

// ---------------------------------------------------------------- SPPF pooling
// grid (c / CG, n); the CTA keeps an (h*w) x CG-channel slab in shared memory and runs three separable
// 5x5 max pools (horizontal then vertical), writing each result to its concat slice.
// kPoolV = 16-byte vectors per pixel per CTA: 4 (64 contiguous bytes = two full 32-byte sectors) for the 20x20 maps of the 640
// configs, 2 for the 40x40 maps of --img 1280, 1 for anything larger or for channel counts that are no multiple of 16.

__device__ __forceinline__ uint4 hmax8(uint4 a, uint4 b) {
  uint4 r;
  const __half2* pa = reinterpret_cast<const __half2*>(&a);
  const __half2* pb = reinterpret_cast<const __half2*>(&b);
  __half2* pr = reinterpret_cast<__half2*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) pr[i] = __hmax2(pa[i], pb[i]);
  return r;
}

// One 5-wide clipped max along the lines of a map held in shared memory: out[i] = max(in[i - 2 .. i + 2] inside the line).  A thread
// owns (segment of a line, 8-channel vector) and slides a five-value window kept in registers along it: one shared-memory read
// and one write per element (the first version re-read every neighbour from shared memory and was bound by that bandwidth).
// kGlobal: the result also goes to the tensor (the pool's output slice); `elem_of(line, i)` = pixel index of element i of a line.
template <int kPoolV, bool kRows, bool kGlobal>
__device__ __forceinline__ void sppf_line_pass(const uint4* in, uint4* out, int h, int w, int pitch, __half* gbase, int ct) {
  const int lines = kRows ? h : w, len = kRows ? w : h;
  const int stride_line = kRows ? pitch : kPoolV, stride_elem = kRows ? kPoolV : pitch;
  const int per = lines * kPoolV;
  const int segs = (int)blockDim.x / per > 0 ? (int)blockDim.x / per : 1;
  const int seg = (len + segs - 1) / segs;
  const uint4 neg = make_uint4(0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u);  // -inf: the identity of max (clipped windows)
  for (int t = threadIdx.x; t < per * segs; t += blockDim.x) {
    const int v = t % kPoolV, line = (t / kPoolV) % lines, sg = t / per;
    const int i0 = sg * seg, i1 = i0 + seg < len ? i0 + seg : len;
    if (i0 >= i1) continue;
    const uint4* src = in + line * stride_line + v;
    uint4* dst = out + line * stride_line + v;
    auto at = [&](int i) { return i >= 0 && i < len ? src[i * stride_elem] : neg; };
    uint4 a = at(i0 - 2), b = at(i0 - 1), c = at(i0), d = at(i0 + 1);
    for (int i = i0; i < i1; ++i) {
      const uint4 e = at(i + 2);
      const uint4 m = hmax8(hmax8(hmax8(a, b), hmax8(c, d)), e);
      dst[i * stride_elem] = m;
      if (kGlobal) {
        const int pix = kRows ? line * w + i : i * w + line;
        *reinterpret_cast<uint4*>(gbase + (size_t)pix * ct + v * 8) = m;
      }
      a = b; b = c; c = d; d = e;
    }
  }
}

// SPPF's three chained 5x5 / stride 1 / pad 2 max pools of one (image, 8 kPoolV channels) slab, the map in shared memory (rows
// padded by one vector group: conflict-free row walks): per pool a row pass and a column pass, the column pass writing the slice.
template <int kPoolV>
__global__ void __launch_bounds__(256) sppf_pool_kernel(__half* buf, int h, int w, int ct, int c_off, int c) {
  constexpr int kPoolCG = 8 * kPoolV;  // channels per CTA
  extern __shared__ uint4 s_pool[];
  const int hw = h * w, nv = hw * kPoolV;  // element i = pixel (i / kPoolV), vector (i % kPoolV)
  const int pitch = (w + 1) * kPoolV;
  uint4* s0 = s_pool;
  uint4* s1 = s_pool + h * pitch;
  const int cg = blockIdx.x;
  const int n = blockIdx.y;
  __half* base = buf + (size_t)n * hw * ct + c_off + cg * kPoolCG;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    const int pix = i / kPoolV, y = pix / w, x = pix - y * w;
    s0[y * pitch + x * kPoolV + (i % kPoolV)] = *reinterpret_cast<const uint4*>(base + (size_t)pix * ct + (i % kPoolV) * 8);
  }
  __syncthreads();
  for (int pass = 1; pass <= 3; ++pass) {
    sppf_line_pass<kPoolV, true, false>(s0, s1, h, w, pitch, nullptr, 0);
    __syncthreads();
    sppf_line_pass<kPoolV, false, true>(s1, s0, h, w, pitch, base + (size_t)pass * c, ct);
    __syncthreads();
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

cudaError_t launch_tile_stats(const PreParams& p, unsigned int* col_sums, TileStats* stats, cudaStream_t stream) {
  if (p.in_w > 256 * kStatColSlots || p.in_h > 4096) return cudaErrorInvalidValue;  // 750 * 4096 * ... stays inside uint32
  cudaError_t e = cudaMemsetAsync(col_sums, 0, (size_t)p.n * p.in_w * sizeof(unsigned int), stream);
  if (e != cudaSuccess) return e;
  tile_stats_init<<<(p.n + 255) / 256, 256, 0, stream>>>(stats, p.n);
  const bool vec = (p.in_w % 16) == 0 && (p.row_pitch % 16) == 0 && (reinterpret_cast<uintptr_t>(p.src) % 16) == 0 &&
                   (!p.windowed || ((size_t)p.in_w * 3) % 16 == 0);
  if (vec) {
    const dim3 grid((p.in_h + kStatBand - 1) / kStatBand, p.n);
    const size_t smem = (p.in_w + kStatBand) * sizeof(unsigned int);
    if (256 % (p.in_w / 16) == 0) tile_stats_vec_kernel<true><<<grid, 256, smem, stream>>>(p, col_sums, stats);
    else tile_stats_vec_kernel<false><<<grid, 256, smem, stream>>>(p, col_sums, stats);
  } else
    tile_stats_kernel<<<dim3((p.in_h + kStatRows - 1) / kStatRows, p.n), 256, 0, stream>>>(p, col_sums, stats);
  tile_stats_finish<<<p.n, 256, 0, stream>>>(col_sums, stats, p.in_h, p.in_w);
  return cudaGetLastError();
}

// Scratch of aq_tile_stats_static (see aux_kernels.cuh): shared by every call on this device, keeps the last call's contents
// (launch_tile_stats clears the part it uses), so calls that overlap on the GPU race.
static __device__ __align__(256) unsigned int g_tile_stats_cols[kTileStatsStaticCols];

cudaError_t tile_stats_static_scratch(unsigned int** ptr) {
  constexpr int kMaxDev = 64;
  static unsigned int* cache[kMaxDev] = {};  // device address per device; written once per device at context set-up
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
  if (!cache[dev]) {
    unsigned int* a = nullptr;
    e = cudaGetSymbolAddress((void**)&a, g_tile_stats_cols);  // NOT &g_tile_stats_cols: that is a host address
    if (e != cudaSuccess) return e;
    cache[dev] = a;
  }
  *ptr = cache[dev];
  return cudaSuccess;
}

cudaError_t sppf_init() {
  cudaError_t e = cudaFuncSetAttribute(sppf_pool_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(sppf_pool_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(sppf_pool_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

cudaError_t launch_sppf_pool(__half* buf, int n, int h, int w, int ct, int c_off, int c, cudaStream_t stream) {
  const size_t smem1 = (size_t)2 * h * (w + 1) * sizeof(uint4);  // two maps, one 8-channel vector per pixel, rows padded by one
  // 32 channels x 256 threads per CTA measured best at 20x20 (29.0 us per 64 images; 16 / 8 channels or 64 - 128 threads: 31 - 46 us),
  // 16 channels at 40x40 (39 us at batch 16)
  if (4 * smem1 <= 100 * 1024 && (c % 32) == 0) {  // two CTAs per SM
    sppf_pool_kernel<4><<<dim3(c / 32, n), 256, 4 * smem1, stream>>>(buf, h, w, ct, c_off, c);
  } else if (2 * smem1 <= 110 * 1024 && (c % 16) == 0) {
    sppf_pool_kernel<2><<<dim3(c / 16, n), 256, 2 * smem1, stream>>>(buf, h, w, ct, c_off, c);
  } else {
    if (smem1 > 200 * 1024 || (c % 8) != 0) return cudaErrorInvalidValue;  // maps beyond ~6000 pixels: not supported
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
