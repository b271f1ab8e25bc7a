// Device-side Huffman decoding of baseline JPEG scans (jpeg_huff.cu): parameters shared with the C ABI (api.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "jpeg_huff_core.h"

namespace aq {

struct HuffParams {
  const uint8_t* scans;                 // prepared entropy-coded segments (byte stuffing removed, zero padded), device
  const unsigned long long* scan_off;   // per tile: byte offset of its segment in `scans` (multiple of 4)
  const uint32_t* scan_bytes;           // per tile: segment length in bytes
  const uint8_t* tables;                // per tile: kTablesPerTile x DevHuff
  int16_t* coef;                        // out: n x total_blocks x 64 dense coefficients (planes Y | Cb | Cr), DC excluded
  int* dc;                              // out: n x total_blocks DC coefficients in decode order (mcu * bpm + slot)
  aqj::Geo geo;
  int sub_bits;                         // subsequence length in bits (multiple of 32)
  // set by launch_jpeg_huffman from the state scratch
  int nsub_max;
  unsigned long long* start;            // [tile][sub] state a subsequence was last decoded from
  unsigned long long* end;              // [tile][sub] state that decoding left
  long long* base;                      // [tile][sub] blocks completed before the subsequence
  int* nblk;                            // [tile][sub] blocks completed inside it
  int* status;                          // [tile] 0 or AQ_ERR_IO
};

size_t jpeg_huff_state_bytes(int n, int nsub_max);
const int* jpeg_huff_status_ptr(void* state_scratch, int n, int nsub_max);
cudaError_t launch_jpeg_huffman(HuffParams p, void* state_scratch, int n, int nsub_max, cudaStream_t stream);

}  // namespace aq
