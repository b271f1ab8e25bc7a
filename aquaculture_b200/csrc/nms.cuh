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
};

size_t nms_scratch_bytes(int batch, int rows);
cudaError_t launch_nms(const NmsParams& p, int batch, cudaStream_t stream);
cudaError_t nms_init();

}  // namespace aq
