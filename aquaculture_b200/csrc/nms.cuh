// Batched NMS (see nms.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace aq {

struct NmsParams {
  const float* pred;  // (batch, rows, no) fp32
  int rows, no;
  float conf_thres;
  float iou_thr_f32;  // largest fp32 <= iou_thres (host computes it in double)
  int max_det, max_nms, agnostic;
  float max_wh;
  int n_classes;
  int classes[32];
  // candidate bitmask, (batch, mask_words = ceil(rows / 32)): bit r % 32 of word r / 32 set iff row r passes
  // obj > conf and max_c(obj * cls_c) > conf (and the class filter, if any).  Written by nms_mask_kernel (all SMs, one
  // pass over pred) or, in the planned path, by the Detect epilogues themselves (conv_igemm.cu, EPI_DETECT).
  const uint32_t* cand_mask;
  // scratch, per image `rows` entries (keys/vals: 2*rows for ping-pong)
  float4* cand_box;
  float* cand_score;
  float* cand_cls;
  int* cand_row;
  uint32_t* keys;
  uint32_t* vals;
  // outputs
  float* dets;     // (batch, max_det, 6)
  int* counts;     // (batch)
  int* keep_rows;  // (batch, max_det) or nullptr
  long long* dbg_clk;  // optional (AQUA_NMS_CLK): 16 clock64 stamps of image 0's CTA at the phase boundaries
  int dbg_mode;        // AQUA_NMS_CLK=2 / 3: timing experiments that skip parts of the suppression (wrong results)
};

size_t nms_scratch_bytes(int batch, int rows);
inline int nms_mask_words(int rows) { return (rows + 31) / 32; }
// Builds p.cand_mask (the pointer is written to, despite the const of the consumer's view) from p.pred on every SM.
cudaError_t launch_nms_mask(const NmsParams& p, int batch, cudaStream_t stream);
cudaError_t launch_nms(const NmsParams& p, int batch, cudaStream_t stream);
cudaError_t nms_init();

}  // namespace aq
