// Device half of the hybrid JPEG decoder (jpeg.cu): parameters shared with the C ABI (api.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace aq {

struct JpegParams {
  const int16_t* coef;     // n x blocks_per_image x 64, planes Y | Cb | Cr, blocks in raster order, natural order inside
  const uint16_t* quant;   // n x 3 x 64 quantisation steps, natural order
  // packed input (coef == nullptr): per block (first entry << 7 | count), entries (position << 16 | value) and the
  // index of each tile's first entry
  const uint32_t* block_desc;
  const uint32_t* entries;
  const unsigned long long* tile_entry_base;
  // optional (all-device decoder): DC coefficients kept apart in decode order, dc[tile][mcu * bpm + slot]; when set they
  // replace coefficient 0 of every dense block
  const int* dc;
  int dc_bpm, dc_mcus_x;
  uint8_t* planes;         // scratch: n x planes_per_image reconstructed samples (block-padded planes)
  uint8_t* out;            // (n, height, width, 3) uint8
  int n, width, height, ncomp, hs, bgr;
  // derived by jpeg_geometry()
  int luma_bw, luma_pw, chroma_bw, chroma_pw;
  long long luma_blocks, chroma_blocks, luma_plane, chroma_plane, blocks_per_image, planes_per_image;
};

void jpeg_geometry(JpegParams& p);
cudaError_t launch_jpeg_reconstruct(const JpegParams& p, cudaStream_t stream);

}  // namespace aq
