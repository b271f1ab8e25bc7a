// Bandwidth-bound helper kernels of the tile path: letterbox/normalise, SPPF pooling, nearest upsample.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace aq {

struct PreParams {
  const uint8_t* src;  // (n, in_h, in_w, 3) uint8
  int n, in_h, in_w, bgr;
  int net_h, net_w;    // letterboxed size (multiples of 32)
  int new_h, new_w;    // resized (unpadded) size
  int top, left;       // padding offsets
  int resize;          // 0: copy, 1: bilinear (cv2 INTER_LINEAR fixed point)
  // per destination column / row (resize == 1): source indices and 11-bit weights
  const int* x0; const int* x1; const short* ax0; const short* ax1;
  const int* y0; const int* y1; const short* ay0; const short* ay1;
  // Source addressing.  Plain batch: tile n starts at src + n * in_h * row_pitch, row_pitch = in_w * 3.
  // Raster windows (tiling fused into inference, replaces src/load_data/tile_tifs.py::split_all_tiffs): src is one
  // (raster_h, raster_w, 3) image, tile n is window win0 + n of the regular grid of in_h x in_w windows (column-major
  // like the reference's loop: index = ix * wins_y + iy, origin (ix * in_w, iy * in_h)), row_pitch = raster_w * 3.
  size_t row_pitch;
  int windowed, wins_y, win0;
  const int* win_list;  // optional (device, n entries): tile n is window win_list[n] instead of win0 + n (blank windows left out)
  __half* out_s2d;     // (n, net_h/2, net_w/2, 16) fp16, channel (py*2+px)*3 + rgb, 12..15 = 0
  uint8_t* out_u8;     // (n, net_h, net_w, 3) RGB, optional debug/parity output
};

cudaError_t launch_preprocess(const PreParams& p, cudaStream_t stream);
cudaError_t launch_preprocess_u8(const PreParams& p, cudaStream_t stream);

// Blank-tile statistics (replaces the reference's src/utils.py::is_blank / is_partly_blank [REF src/utils.py:325-369]):
// per tile the PIL "L" luminance extrema and the number of rows / columns whose mean over (pixels, channels) is >= 250.
struct TileStats {
  int l_min, l_max;              // extrema of L = (19595 R + 38470 G + 7471 B + 0x8000) >> 16
  int blank_rows, blank_cols;    // rows / columns with sum >= 250 * 3 * width (height)
};
// `col_sums` is scratch of n * in_w uint32 (zeroed by the call); stats is a device array of n entries.
cudaError_t launch_tile_stats(const PreParams& p, unsigned int* col_sums, TileStats* stats, cudaStream_t stream);

// Allocation-free blank-tile statistics (aq_tile_stats_static): the column sums live in a file-scope __device__ array of
// kTileStatsStaticCols uint32 (512 KB, held by every GPU that loads the library until the process exits) instead of
// cudaMalloc'd scratch.  tile_stats_static_scratch resolves the array's DEVICE address for the CURRENT device
// (cudaGetSymbolAddress; the array's name in host code is only a host-side stand-in) -- once per device, cached.
constexpr int kTileStatsStaticCols = 131072;  // n * in_w of one call, e.g. 64 tiles of 2048 columns
cudaError_t tile_stats_static_scratch(unsigned int** ptr);

// SPPF: three chained 5x5/s1/p2 max pools of channels [c_off, c_off+c) written to the next three slices.
cudaError_t launch_sppf_pool(__half* buf, int n, int h, int w, int ct, int c_off, int c, cudaStream_t stream);
cudaError_t sppf_init();
// nearest 2x upsample of a channel slice into a channel slice of the destination
cudaError_t launch_upsample2x(const __half* src, int n, int h, int w, int src_ct, int src_c_off, int c, __half* dst,
                              int dst_ct, int dst_c_off, cudaStream_t stream);

}  // namespace aq
