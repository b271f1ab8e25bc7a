// Host-side label epilogue: box rescale + '%g' text + file append.  No GPU code here.
//
// Replaces upstream utils/general.py::{scale_boxes, clip_boxes, xyxy2xywh} and the --save-txt block of
// detect.py::run [UPSTREAM-RECALLED]; the grammar is what the consumer parses with np.loadtxt and
// bbox[0..5] [REF src/process_yolo/geocode_results.py:140-172].  fp32 arithmetic in torch's operation
// order (compile with -ffp-contract=off): (x - pad) / gain, clamp, round-half-even, then
// ((x1+x2)/2)/w0 etc.; '%g' of the value widened to double, exactly what Python's '%g' prints.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "../../include/aqua_b200.h"

namespace {

inline float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

// Appends the text of one image's detections (ascending confidence) to `out`.
void format_one(const float* det, int count, int net_h, int net_w, int h0, int w0, int save_conf, std::string& out) {
  // gain/pad are Python floats (double) in upstream, then applied to fp32 tensors as fp32 scalars
  const double gain_d = fmin((double)net_h / h0, (double)net_w / w0);
  const float gain = (float)gain_d;
  const float pad_x = (float)(((double)net_w - w0 * gain_d) / 2.0);
  const float pad_y = (float)(((double)net_h - h0 * gain_d) / 2.0);
  const float fw = (float)w0, fh = (float)h0;
  char line[256];
  for (int i = count - 1; i >= 0; --i) {
    const float* d = det + (size_t)i * 6;
    float x1 = (d[0] - pad_x) / gain, y1 = (d[1] - pad_y) / gain;
    float x2 = (d[2] - pad_x) / gain, y2 = (d[3] - pad_y) / gain;
    x1 = nearbyintf(clampf(x1, 0.f, fw));
    y1 = nearbyintf(clampf(y1, 0.f, fh));
    x2 = nearbyintf(clampf(x2, 0.f, fw));
    y2 = nearbyintf(clampf(y2, 0.f, fh));
    const float cx = ((x1 + x2) / 2.0f) / fw;
    const float cy = ((y1 + y2) / 2.0f) / fh;
    const float bw = (x2 - x1) / fw;
    const float bh = (y2 - y1) / fh;
    int len;
    if (save_conf)
      len = snprintf(line, sizeof line, "%g %g %g %g %g %g\n", (double)d[5], (double)cx, (double)cy, (double)bw,
                     (double)bh, (double)d[4]);
    else
      len = snprintf(line, sizeof line, "%g %g %g %g %g\n", (double)d[5], (double)cx, (double)cy, (double)bw,
                     (double)bh);
    out.append(line, (size_t)len);
  }
}

}  // namespace

extern "C" int aq_format_label_text(const float* dets, int count, int net_h, int net_w, int h0, int w0, int save_conf,
                                    char* out, size_t out_cap) {
  if (!dets || !out || count < 0 || h0 <= 0 || w0 <= 0) return AQ_ERR_ARG;
  std::string s;
  format_one(dets, count, net_h, net_w, h0, w0, save_conf, s);
  if (s.size() + 1 > out_cap) return AQ_ERR_NOMEM;
  memcpy(out, s.data(), s.size());
  out[s.size()] = 0;
  return (int)s.size();
}

extern "C" int aq_format_labels(const float* dets, const int* counts, int n_images, int det_stride, int net_h,
                                int net_w, const int* h0, const int* w0, const char* const* paths, int save_conf,
                                int n_threads) {
  if (!dets || !counts || !h0 || !w0 || !paths || n_images < 0) return AQ_ERR_ARG;
  std::atomic<int> next(0), written(0), failed(0);
  auto work = [&]() {
    std::string s;
    for (;;) {
      const int i = next.fetch_add(1);
      if (i >= n_images) break;
      if (counts[i] <= 0) continue;  // zero detections -> NO file (an empty file breaks the consumer)
      s.clear();
      format_one(dets + (size_t)i * det_stride * 6, counts[i], net_h, net_w, h0[i], w0[i], save_conf, s);
      FILE* f = fopen(paths[i], "a");  // append, as upstream does
      if (!f) {
        failed.fetch_add(1);
        continue;
      }
      const size_t w = fwrite(s.data(), 1, s.size(), f);
      if (fclose(f) != 0 || w != s.size()) failed.fetch_add(1);
      else written.fetch_add(1);
    }
  };
  if (n_threads <= 1 || n_images < 8) {
    work();
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; ++t) th.emplace_back(work);
    for (auto& t : th) t.join();
  }
  if (failed.load()) return AQ_ERR_IO;
  return written.load();
}
