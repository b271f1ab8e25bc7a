// Device half of the hybrid JPEG decoder: dequantisation + 8x8 inverse DCT into component planes, then
// chroma upsampling + YCbCr -> RGB into the uint8 HWC tile the preprocess kernel reads.  Bit-identical to what
// cv2.imread returns inside upstream utils/dataloaders.py::LoadImages [UPSTREAM-RECALLED], i.e. to libjpeg's default
// decompression settings: the accurate integer IDCT (Loeffler-Ligtenberg-Moschytz, 13-bit constants, 2 extra bits
// kept between the passes), the 3:1 triangle-filter chroma upsampling and the 16-bit fixed-point colour matrix of
// ITU-R BT.601 as JFIF prescribes.  Both kernels are HBM bound (integer / byte work; no tensor cores).
//   idct:   reads 2 B per coefficient (+ the 384 B table), writes 1 B per sample
//   colour: reads 1 B per sample, writes 3 B per pixel
#include "jpeg.cuh"

namespace aq {
namespace {

constexpr int CONST_BITS = 13, PASS1_BITS = 2;
constexpr int F_0_298 = 2446, F_0_390 = 3196, F_0_541 = 4433, F_0_765 = 6270, F_0_899 = 7373, F_1_175 = 9633,
              F_1_501 = 12299, F_1_847 = 15137, F_1_961 = 16069, F_2_053 = 16819, F_2_562 = 20995, F_3_072 = 25172;

// One 1-D pass of the 8-point inverse DCT (even part: rotator sqrt(2) c(-6); odd part: 4-rotator form).
// In: s[0..7]; out: the eight outputs before the descale.
__device__ __forceinline__ void idct8(const int (&s)[8], int (&o)[8]) {
  int z2 = s[2], z3 = s[6];
  int z1 = (z2 + z3) * F_0_541;
  int tmp2 = z1 + z3 * (-F_1_847);
  int tmp3 = z1 + z2 * F_0_765;
  int tmp0 = (s[0] + s[4]) << CONST_BITS;
  int tmp1 = (s[0] - s[4]) << CONST_BITS;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = s[7]; tmp1 = s[5]; tmp2 = s[3]; tmp3 = s[1];
  z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
  int z4 = tmp1 + tmp3;
  const int z5 = (z3 + z4) * F_1_175;
  tmp0 *= F_0_298; tmp1 *= F_2_053; tmp2 *= F_3_072; tmp3 *= F_1_501;
  z1 *= -F_0_899; z2 *= -F_2_562; z3 *= -F_1_961; z4 *= -F_0_390;
  z3 += z5; z4 += z5;
  tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
  o[0] = tmp10 + tmp3; o[7] = tmp10 - tmp3;
  o[1] = tmp11 + tmp2; o[6] = tmp11 - tmp2;
  o[2] = tmp12 + tmp1; o[5] = tmp12 - tmp1;
  o[3] = tmp13 + tmp0; o[4] = tmp13 - tmp0;
}

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// The decoder's output clamp: a 1024-entry wrap-around table in libjpeg; identical for every in-range value.
__device__ __forceinline__ unsigned range_limit(int v) {
  v &= 1023;
  return v < 128 ? 128 + v : v < 512 ? 255 : v < 896 ? 0 : v - 896;
}

// Eight lanes per 8x8 block, 32 blocks per 256-thread CTA.  Lane t of a block loads coefficient row t (one 16-byte
// load: a warp reads 512 contiguous bytes), the block is transposed through shared memory for the column pass, the
// workspace goes back through shared memory for the row pass, and lane t stores output row t (8 bytes).
constexpr int IDCT_THREADS = 256, IDCT_BLOCKS = IDCT_THREADS / 8, WS_PITCH = 9, WS_BLOCK = 8 * WS_PITCH + 1;

template <bool PACKED>
__global__ void __launch_bounds__(IDCT_THREADS) jpeg_idct_kernel(JpegParams p) {
  __shared__ int ws[IDCT_BLOCKS * WS_BLOCK];
  __shared__ unsigned short qs[192];
  const int img = blockIdx.y;
  if (threadIdx.x < 192) qs[threadIdx.x] = p.quant[(size_t)img * 192 + threadIdx.x];
  __syncthreads();
  const int g = threadIdx.x >> 3, t = threadIdx.x & 7;
  const long long blk = (long long)blockIdx.x * IDCT_BLOCKS + g;
  const bool live = blk < p.blocks_per_image;
  int comp = 0, bw = p.luma_bw, pw = p.luma_pw;
  long long b = blk, plane_off = 0;
  if (live && blk >= p.luma_blocks) {
    b = blk - p.luma_blocks;
    comp = 1;
    plane_off = p.luma_plane;
    if (b >= p.chroma_blocks) { b -= p.chroma_blocks; comp = 2; plane_off += p.chroma_plane; }
    bw = p.chroma_bw;
    pw = p.chroma_pw;
  }
  int* w = ws + g * WS_BLOCK;
  if (PACKED) {
    // the block arrives as a short list of non-zero coefficients: clear the workspace, scatter the list
    if (live) {
#pragma unroll
      for (int i = 0; i < 8; ++i) w[t * WS_PITCH + i] = 0;
    }
    __syncwarp();
    if (live) {
      const unsigned desc = p.block_desc[(size_t)img * p.blocks_per_image + blk];
      const unsigned* e = p.entries + p.tile_entry_base[img] + (desc >> 7);
      const int cnt = desc & 127;
      for (int i = t; i < cnt; i += 8) {
        const unsigned v = e[i];
        const int pos = (v >> 16) & 63;
        w[(pos >> 3) * WS_PITCH + (pos & 7)] = (int)(short)(v & 0xFFFF) * (int)qs[comp * 64 + pos];
      }
    }
  } else if (live) {
    const int4 raw = *reinterpret_cast<const int4*>(p.coef + ((size_t)img * p.blocks_per_image + blk) * 64 + t * 8);
    const short* c = reinterpret_cast<const short*>(&raw);
    const unsigned short* q = qs + comp * 64 + t * 8;
#pragma unroll
    for (int i = 0; i < 8; ++i) w[t * WS_PITCH + i] = (int)c[i] * (int)q[i];
    if (p.dc && t == 0) {
      // the block's place in decode order: MCU (my, mx), slot inside the MCU (luma blocks row-major, then Cb, Cr)
      const int by = (int)(b / bw), bx = (int)(b - (long long)by * bw);
      const int v = comp == 0 ? p.hs : 1;
      const int my = by / v, mx = bx / v;
      const int slot = comp == 0 ? (by - my * v) * v + (bx - mx * v) : p.hs * p.hs + comp - 1;
      const long long g = ((long long)my * p.dc_mcus_x + mx) * p.dc_bpm + slot;
      w[0] = p.dc[(size_t)img * p.blocks_per_image + g] * (int)q[0];
    }
  }
  __syncwarp();
  int s[8], o[8];
  if (live) {
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = w[i * WS_PITCH + t];  // column t
    idct8(s, o);
  }
  __syncwarp();
  if (live) {
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i * WS_PITCH + t] = descale(o[i], CONST_BITS - PASS1_BITS);
  }
  __syncwarp();
  if (live) {
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = w[t * WS_PITCH + i];  // row t
    idct8(s, o);
    unsigned lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      lo |= range_limit(descale(o[i], CONST_BITS + PASS1_BITS + 3)) << (8 * i);
      hi |= range_limit(descale(o[i + 4], CONST_BITS + PASS1_BITS + 3)) << (8 * i);
    }
    const int by = (int)(b / bw), bx = (int)(b - (long long)by * bw);
    uint8_t* dst = p.planes + (size_t)img * p.planes_per_image + plane_off + (size_t)(by * 8 + t) * pw + bx * 8;
    *reinterpret_cast<uint2*>(dst) = make_uint2(lo, hi);
  }
}

__device__ __forceinline__ unsigned clamp255(int v) { return (unsigned)min(max(v, 0), 255); }

// Eight horizontally adjacent pixels per thread: one 8-byte luma load, chroma through 32-bit loads (4:2:0: the four
// samples under the pixels + one neighbour each side, for two rows; 4:4:4: 8 bytes per plane), 24 output bytes as three
// 8-byte streaming stores when rows are 8-byte aligned.  Threads at the right edge (or of narrow images) take the
// byte-wise path with clamped indices.
__global__ void __launch_bounds__(256) jpeg_colour_kernel(JpegParams p) {
  const int octs = (p.width + 7) >> 3;   // groups per row; consecutive threads walk a row, then the next one
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int y = (int)(gid / octs), img = blockIdx.y;
  if (y >= p.height) return;
  const int x0 = (int)(gid - (long long)y * octs) * 8;
  const uint8_t* planes = p.planes + (size_t)img * p.planes_per_image;
  const uint2 yy8 = *reinterpret_cast<const uint2*>(planes + (size_t)y * p.luma_pw + x0);  // planes are padded to whole blocks
  int cb[8], cr[8];
  if (p.ncomp == 1) {
#pragma unroll
    for (int i = 0; i < 8; ++i) cb[i] = cr[i] = 128;
  } else if (p.hs == 1) {
    const uint2 b8 = *reinterpret_cast<const uint2*>(planes + p.luma_plane + (size_t)y * p.chroma_pw + x0);
    const uint2 r8 = *reinterpret_cast<const uint2*>(planes + p.luma_plane + p.chroma_plane + (size_t)y * p.chroma_pw + x0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      cb[i] = (b8.x >> (8 * i)) & 255; cb[i + 4] = (b8.y >> (8 * i)) & 255;
      cr[i] = (r8.x >> (8 * i)) & 255; cr[i + 4] = (r8.y >> (8 * i)) & 255;
    }
  } else {
    // triangle filter: 3/4 nearer + 1/4 further sample in each axis; rows / columns beyond the component's real
    // extent (ceil(size / 2)) are replaced by the edge sample
    const int ch = (p.height + 1) >> 1, cw = (p.width + 1) >> 1;
    const int cy = y >> 1;
    const int fy = min(max((y & 1) ? cy + 1 : cy - 1, 0), ch - 1);
    const int ci = x0 >> 1;    // multiple of 4
    const bool box = cw <= 2;  // libjpeg filters only components more than two samples wide; narrower ones are replicated
    const bool interior = ci >= 1 && ci + 4 <= cw - 1;
#pragma unroll
    for (int pl = 0; pl < 2; ++pl) {
      const uint8_t* base = planes + p.luma_plane + (size_t)pl * p.chroma_plane;
      const uint8_t* near = base + (size_t)cy * p.chroma_pw;
      const uint8_t* far = base + (size_t)fy * p.chroma_pw;
      int sum[6];  // columns ci - 1 .. ci + 4
      if (interior) {
        const unsigned n4 = *reinterpret_cast<const unsigned*>(near + ci), f4 = *reinterpret_cast<const unsigned*>(far + ci);
        sum[0] = 3 * (int)near[ci - 1] + (int)far[ci - 1];
        sum[5] = 3 * (int)near[ci + 4] + (int)far[ci + 4];
#pragma unroll
        for (int j = 0; j < 4; ++j) sum[1 + j] = 3 * (int)((n4 >> (8 * j)) & 255) + (int)((f4 >> (8 * j)) & 255);
      } else {
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          const int c = min(max(ci - 1 + j, 0), cw - 1);
          sum[j] = box ? 4 * (int)near[c] : 3 * (int)near[c] + (int)far[c];
        }
      }
      int* o = pl ? cr : cb;
#pragma unroll
      for (int j = 0; j < 4; ++j) {  // chroma column ci + j -> pixels 2j, 2j + 1
        o[2 * j] = (3 * sum[1 + j] + (box ? sum[1 + j] : sum[j]) + 8) >> 4;
        o[2 * j + 1] = (3 * sum[1 + j] + (box ? sum[1 + j] : sum[2 + j]) + 7) >> 4;
      }
    }
  }
  unsigned char px[24];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int yy = ((i < 4 ? yy8.x : yy8.y) >> (8 * (i & 3))) & 255;
    const int u = cb[i] - 128, v = cr[i] - 128;
    const unsigned r = clamp255(yy + ((91881 * v + 32768) >> 16));
    const unsigned g = clamp255(yy + ((-22554 * u + (-46802 * v + 32768)) >> 16));
    const unsigned bl = clamp255(yy + ((116130 * u + 32768) >> 16));
    px[3 * i + 0] = p.bgr ? bl : r;
    px[3 * i + 1] = g;
    px[3 * i + 2] = p.bgr ? r : bl;
  }
  uint8_t* dst = p.out + ((size_t)img * p.height + y) * (size_t)p.width * 3 + (size_t)x0 * 3;
  if ((p.width & 7) == 0) {  // whole group inside the row, rows 8-byte aligned
    unsigned w32[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) w32[i] = px[4 * i] | (px[4 * i + 1] << 8) | (px[4 * i + 2] << 16) | ((unsigned)px[4 * i + 3] << 24);
    uint2* d64 = reinterpret_cast<uint2*>(dst);
#pragma unroll
    for (int i = 0; i < 3; ++i) __stcs(d64 + i, make_uint2(w32[2 * i], w32[2 * i + 1]));
  } else {
    const int nb = min(8, p.width - x0) * 3;
    for (int i = 0; i < nb; ++i) dst[i] = px[i];
  }
}

}  // namespace

void jpeg_geometry(JpegParams& p) {
  const int mcus_x = (p.width + 8 * p.hs - 1) / (8 * p.hs), mcus_y = (p.height + 8 * p.hs - 1) / (8 * p.hs);
  p.luma_bw = mcus_x * p.hs;
  p.luma_pw = p.luma_bw * 8;
  p.luma_blocks = (long long)p.luma_bw * mcus_y * p.hs;
  p.luma_plane = p.luma_blocks * 64;
  p.chroma_bw = mcus_x;
  p.chroma_pw = mcus_x * 8;
  p.chroma_blocks = p.ncomp == 3 ? (long long)mcus_x * mcus_y : 0;
  p.chroma_plane = p.chroma_blocks * 64;
  p.blocks_per_image = p.luma_blocks + 2 * p.chroma_blocks;
  p.planes_per_image = p.blocks_per_image * 64;
}

cudaError_t launch_jpeg_reconstruct(const JpegParams& p, cudaStream_t stream) {
  if (p.n <= 0 || p.n > 65535 || p.height > 65535) return cudaErrorInvalidValue;
  const unsigned gx = (unsigned)((p.blocks_per_image + IDCT_BLOCKS - 1) / IDCT_BLOCKS);
  if (p.coef) jpeg_idct_kernel<false><<<dim3(gx, p.n), IDCT_THREADS, 0, stream>>>(p);
  else jpeg_idct_kernel<true><<<dim3(gx, p.n), IDCT_THREADS, 0, stream>>>(p);
  const long long groups = (long long)((p.width + 7) / 8) * p.height;
  jpeg_colour_kernel<<<dim3((unsigned)((groups + 255) / 256), p.n), 256, 0, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace aq
