// Implicit-GEMM convolution on tcgen05 tensor cores (sm_100a), operands fed by TMA.
//
// Replaces upstream models/common.py::Conv.forward_fuse (+ Bottleneck residual) and, with EPI_DETECT,
// models/yolo.py::Detect.forward [UPSTREAM-RECALLED; the reference reaches them through
// `yolov5/detect.py`, REF README.md:77].
//
//   D[M = pixels, N = c_out] = sum over taps (kh,kw) and channel chunks of  A_tap[M, cb] * W_tap[N, cb]^T
//
// * activations NHWC fp16; one M tile = 128 output pixels: 128 consecutive flat pixels for 1x1 convs, or a
//   tw x th spatial patch of one image for 3x3 convs.  Each (tap, channel-chunk) A operand is ONE TMA box
//   load of the shifted patch; conv zero padding is TMA out-of-bounds fill.  Stride-2 convs use a 5-D
//   view (x-parity*C, W/2, y-parity, H/2, N) of the same buffer so the strided gather is still a dense box.
// * weights [c_out][k*k*c_in] fp16, K-major; accumulators fp32 in TMEM (double-buffered so the epilogue of
//   tile i overlaps the MMAs of tile i+1).
// * warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
//   warps 2..5 = epilogue (TMEM -> regs -> bias + SiLU (+ residual) -> fp16 -> swizzled smem -> TMA store
//   into a channel slice of the destination buffer, which is how concat disappears).
// * persistent CTAs (grid = min(tiles, SMs)), static round-robin tile order with N tiles adjacent so
//   CTAs sharing an A patch run concurrently and hit L2.
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace aq {

enum { CONV_MODE_FLAT = 0, CONV_MODE_S1 = 1, CONV_MODE_S2 = 2 };
enum { EPI_STORE = 0, EPI_DETECT = 1 };

constexpr int kConvThreads = 192;
constexpr int kMaxStages = 8;
constexpr int kTileM = 128;
constexpr int kConvMaxDynSmem = 232448 - 1024;  // 227 KB opt-in limit minus static smem

struct alignas(64) ConvKParams {
  CUtensorMap tmA;    // activation loads
  CUtensorMap tmB;    // weight loads
  CUtensorMap tmOut;  // output stores        (EPI_STORE)
  CUtensorMap tmRes;  // residual tile loads  (EPI_STORE, has_res)
  const float* bias;  // [n_tiles_n * tile_n], zero padded
  int mode, ksize, pad;
  int cb, cpt;        // channels per chunk (16/32/64), chunks per tap (= c_in / cb)
  int total_chunks, cps, kstages;  // chunks per tile, chunks per pipeline stage, ceil(total/cps)
  int num_stages;
  int tile_n, n_tiles_n, ob;       // UMMA N, number of N tiles, channels per output store group
  int m_tiles, tw, th, tiles_x, tiles_y;
  int src_ct, src_c_off;           // mode S2: channels of the source buffer / first channel
  int act, has_res;
  uint32_t a_sub_bytes, b_sub_bytes, stage_bytes;  // smem strides (1024-aligned)
  uint32_t a_box_bytes, b_box_bytes;               // bytes one TMA box actually delivers
  uint32_t c_bytes;                                // output staging bytes
  uint32_t tmem_cols, acc_stride;
  // EPI_DETECT
  float* pred;
  int det_no, det_rows, det_level_off, det_ny, det_nx, det_m_total;
  float det_stride;
  float det_anchor[6];
};

size_t conv_smem_bytes(const ConvKParams& p);
cudaError_t launch_conv(const ConvKParams& p, int epi, int num_sms, cudaStream_t stream);
cudaError_t conv_kernels_init();  // sets max dynamic smem attributes

}  // namespace aq
