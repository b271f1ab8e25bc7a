// Host-side label epilogue: box rescale + '%g' text + file append.  No GPU code here.
//
// Replaces upstream utils/general.py::{scale_boxes, clip_boxes, xyxy2xywh} and the --save-txt block of
// detect.py::run [UPSTREAM-RECALLED]; the grammar is what the consumer parses with np.loadtxt and
// bbox[0..5] [REF src/process_yolo/geocode_results.py:140-172].  fp32 arithmetic in torch's operation
// order (compile with -ffp-contract=off): (x - pad) / gain, clamp, round-half-even, then
// ((x1+x2)/2)/w0 etc.; '%g' of the value widened to double, exactly what Python's '%g' prints.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
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

// ---------------------------------------------------------------------------------------------------------
// the arithmetic of the reference's post-processing, host C as upstream's is host Python.
//
// aq_geocode_label_text = the per-row body of geocode_all_detections [REF src/process_yolo/geocode_results.py:158-170]
// + convert_pix_to_m_bboxes [REF src/process_yolo/geocode_results.py:71-101]: label rows as np.loadtxt parses them
// (float64) -> integer pixel box (Python int() truncation) -> EPSG:3857 metres by the affine map of the large tif's
// bounds.  Every product / quotient is evaluated in the reference's order in double (-ffp-contract=off).
// The EPSG:3857 -> EPSG:3035 reprojection upstream delegates to pyproj is not restated here (pyproj is absent from the
// build container, so it could not be pinned).
extern "C" int aq_geocode_label_text(const char* text, int x_offset, int y_offset, const double bounds_3857[4],
                                     int large_tif_size, int im_width, int im_height, aq_geo_row* out, int cap) {
  if (!text || !bounds_3857 || !out || cap < 0 || large_tif_size <= 0) return AQ_ERR_ARG;
  const double xmin_b = bounds_3857[0], ymin_b = bounds_3857[1], xmax_b = bounds_3857[2], ymax_b = bounds_3857[3];
  const double mx = (xmax_b - xmin_b) / large_tif_size, my = (ymax_b - ymin_b) / large_tif_size;  // m per pixel
  int n = 0;
  const char* p = text;
  while (*p) {
    double b[6];
    int got = 0;
    char* end = nullptr;
    while (got < 6) {
      const double v = strtod(p, &end);
      if (end == p) break;
      b[got++] = v;
      p = end;
    }
    while (*p && *p != '\n') ++p;  // rest of the line
    if (*p == '\n') ++p;
    if (got == 0) continue;            // blank line
    if (got != 6) return AQ_ERR_ARG;   // the consumer reads bbox[5]: --save-conf is mandatory [REF geocode_results.py:165]
    if (n >= cap) return AQ_ERR_NOMEM;
    aq_geo_row& r = out[n++];
    r.cls = (int)b[0];
    r.xmin = (int)(im_width * (b[1] - b[3] / 2));
    r.ymin = (int)(im_height * (b[2] - b[4] / 2));
    r.xmax = (int)(im_width * (b[1] + b[3] / 2));
    r.ymax = (int)(im_height * (b[2] + b[4] / 2));
    r.conf = b[5];
    // (xmin, ymin) -> (xmin_m, ymax_m); (xmax, ymax) -> (xmax_m, ymin_m): image y grows downwards
    r.xmin_m = (r.xmin + x_offset) * mx + xmin_b;
    r.ymax_m = ymax_b - (r.ymin + y_offset) * my;
    r.xmax_m = (r.xmax + x_offset) * mx + xmin_b;
    r.ymin_m = ymax_b - (r.ymax + y_offset) * my;
  }
  return n;
}

// ---------------------------------------------------------------------------------------------------------
// EPSG:3857 -> EPSG:3035, the reprojection geocode_all_detections delegates to pyproj
// [REF src/process_yolo/geocode_results.py:31,175-177; CRS_DICT src/utils.py].  Closed forms of the two published
// methods (EPSG Guidance Note 7-2): inverse of Popular Visualisation Pseudo-Mercator (method 1024: the SPHERICAL
// formulas on the WGS 84 semi-major axis, giving geodetic longitude / latitude directly), the null WGS 84 -> ETRS89
// transformation (EPSG:1149, what PROJ selects for this pair), and the ellipsoidal Lambert Azimuthal Equal Area
// projection (method 9820) on GRS 1980 with latitude / longitude of natural origin 52 N / 10 E, false easting /
// northing 4 321 000 / 3 210 000 m.  PARITY UNPINNED against PROJ (pyproj is absent here): pinned instead by the
// published EPSG test point (50 N, 5 E -> 3 962 799.45, 2 999 718.85) in tests/test_lib_abi.py; PROJ evaluates the
// authalic latitude through a series, so agreement is at the sub-micrometre level, not bit for bit.
namespace {
struct Laea3035 {
  double a, e, e2, qp, beta0, sb0, cb0, rq, d, lam0;
  static double qfun(double sphi, double e, double e2) {
    return (1.0 - e2) * (sphi / (1.0 - e2 * sphi * sphi) - (1.0 / (2.0 * e)) * log((1.0 - e * sphi) / (1.0 + e * sphi)));
  }
  Laea3035() {
    const double pi = 3.141592653589793238;
    a = 6378137.0;
    const double f = 1.0 / 298.257222101;  // GRS 1980
    e2 = 2.0 * f - f * f;
    e = sqrt(e2);
    const double phi0 = 52.0 * pi / 180.0;
    lam0 = 10.0 * pi / 180.0;
    qp = qfun(1.0, e, e2);
    const double q0 = qfun(sin(phi0), e, e2);
    beta0 = asin(q0 / qp);
    sb0 = sin(beta0);
    cb0 = cos(beta0);
    rq = a * sqrt(qp / 2.0);
    d = a * (cos(phi0) / sqrt(1.0 - e2 * sin(phi0) * sin(phi0))) / (rq * cb0);
  }
  void forward(double lam, double phi, double* E, double* N) const {
    const double q = qfun(sin(phi), e, e2);
    const double beta = asin(q / qp);
    const double sb = sin(beta), cb = cos(beta), dl = lam - lam0;
    const double B = rq * sqrt(2.0 / (1.0 + sb0 * sb + cb0 * cb * cos(dl)));
    *E = 4321000.0 + B * d * cb * sin(dl);
    *N = 3210000.0 + (B / d) * (cb0 * sb - sb0 * cb * cos(dl));
  }
};
const Laea3035& laea() {
  static const Laea3035 p;
  return p;
}
void merc_to_laea(double x, double y, double* E, double* N) {
  const double R = 6378137.0, half_pi = 1.5707963267948966;
  laea().forward(x / R, half_pi - 2.0 * atan(exp(-y / R)), E, N);
}
}  // namespace

extern "C" int aq_transform_3857_to_3035(const double* xy_in, int n, double* xy_out) {
  if (!xy_in || !xy_out || n < 0) return AQ_ERR_ARG;
  for (int i = 0; i < n; ++i) merc_to_laea(xy_in[2 * i], xy_in[2 * i + 1], &xy_out[2 * i], &xy_out[2 * i + 1]);
  return AQ_OK;
}

extern "C" int aq_lonlat_to_3035(const double* lonlat_deg, int n, double* xy_out) {
  if (!lonlat_deg || !xy_out || n < 0) return AQ_ERR_ARG;
  const double rad = 3.141592653589793238 / 180.0;
  for (int i = 0; i < n; ++i) laea().forward(lonlat_deg[2 * i] * rad, lonlat_deg[2 * i + 1] * rad, &xy_out[2 * i], &xy_out[2 * i + 1]);
  return AQ_OK;
}

// REVERSE_CLASS_MAPPING [REF src/process_yolo/geocode_results.py:24-30]
extern "C" const char* aq_geo_class_name(int cls) {
  static const char* const names[5] = {"circle_farm", "square_farm", "triangle_farm", "other_farm", "rectangle_farm"};
  return cls >= 0 && cls < 5 ? names[cls] : nullptr;
}

// aq_geocode_dets = geocode_all_detections' loop over label files [REF src/process_yolo/geocode_results.py:139-193] fed by the
// detector's output instead of the files: no np.loadtxt per file, no cv.imread per file (upstream opens every image only to see
// that it exists, :148-151).  The values are the ones the text hop would have produced: each detection goes through the label
// formatter's arithmetic and a '%g' print / strtod round trip IN MEMORY (np.loadtxt reads what '%g' wrote: six significant
// digits), then through the per-row body (:158-177).  Rows come out per image in label-file order (ascending confidence).
extern "C" int aq_geocode_dets(const float* dets, const int* counts, int n_images, int det_stride, int net_h, int net_w,
                               const int* h0, const int* w0, const char* const* stems, const double* download_bounds_3857,
                               int n_boxes, int large_tif_size, int im_width, int im_height, aq_geo_det* out, int cap) {
  if (!dets || !counts || !h0 || !w0 || !stems || !download_bounds_3857 || !out || n_images < 0 || cap < 0 || large_tif_size <= 0)
    return AQ_ERR_ARG;
  int n = 0;
  std::string text;
  for (int i = 0; i < n_images; ++i) {
    if (counts[i] <= 0) continue;  // no label file -> nothing for the consumer to glob
    // `_, bbox_ind, x_offset, y_offset = basename.replace('.txt', '').split('_')` (:89): exactly four fields;
    // `year = int(im_name.split('_')[0][-4:])` (:143)
    const char* st = stems[i];
    if (!st) return AQ_ERR_ARG;
    const char* f[4] = {st, nullptr, nullptr, nullptr};
    int nf = 1;
    for (const char* q = st; *q; ++q)
      if (*q == '_') {
        if (nf == 4) return AQ_ERR_ARG;
        f[nf++] = q + 1;
      }
    if (nf != 4) return AQ_ERR_ARG;
    const long len0 = (long)(f[1] - 1 - f[0]);
    if (len0 < 1) return AQ_ERR_ARG;
    char* end = nullptr;
    char ybuf[8] = {0};
    const long yl = len0 < 4 ? len0 : 4;
    memcpy(ybuf, f[0] + len0 - yl, (size_t)yl);
    const long year = strtol(ybuf, &end, 10);
    if (*end) return AQ_ERR_ARG;
    const long ind = strtol(f[1], &end, 10);
    if (end == f[1] || *end != '_') return AQ_ERR_ARG;
    const long xo = strtol(f[2], &end, 10);
    if (end == f[2] || *end != '_') return AQ_ERR_ARG;
    const long yo = strtol(f[3], &end, 10);
    if (end == f[3] || *end) return AQ_ERR_ARG;
    if (ind < 0 || ind >= n_boxes) return AQ_ERR_ARG;  // wanted_bboxes.loc[int(bbox_ind)] would raise KeyError
    text.clear();
    format_one(dets + (size_t)i * det_stride * 6, counts[i], net_h, net_w, h0[i], w0[i], 1, text);
    if (n + counts[i] > cap) return AQ_ERR_NOMEM;
    aq_geo_row rows[64];
    // the per-row body lives in aq_geocode_label_text: same arithmetic, same order
    const char* p = text.c_str();
    while (*p) {
      // hand the formatter's text over in pieces of <= 64 rows
      const char* q = p;
      int lines = 0;
      while (*q && lines < 64) {
        if (*q == '\n') ++lines;
        ++q;
      }
      const std::string piece(p, q);
      const int got = aq_geocode_label_text(piece.c_str(), (int)xo, (int)yo, download_bounds_3857 + 4 * ind, large_tif_size,
                                            im_width, im_height, rows, 64);
      if (got < 0) return got;
      for (int k = 0; k < got; ++k) {
        aq_geo_det& o = out[n++];
        o.image = i;
        o.cls = rows[k].cls;
        o.year = (int)year;
        o.xmin = rows[k].xmin; o.xmax = rows[k].xmax; o.ymin = rows[k].ymin; o.ymax = rows[k].ymax;
        o.conf = rows[k].conf;
        o.xmin_m = rows[k].xmin_m; o.ymin_m = rows[k].ymin_m; o.xmax_m = rows[k].xmax_m; o.ymax_m = rows[k].ymax_m;
        // `xmin_a, ymax_a = transform(xmin_m, ymax_m); xmax_a, ymin_a = transform(xmax_m, ymin_m)` (:175-177)
        merc_to_laea(o.xmin_m, o.ymax_m, &o.xmin_a, &o.ymax_a);
        merc_to_laea(o.xmax_m, o.ymin_m, &o.xmax_a, &o.ymin_a);
      }
      p = q;
    }
  }
  return n;
}

// get_circle_area_from_bbox / get_square_area_from_bbox [REF src/process_yolo/calc_net_areas.py:24-82]:
// out = (estimate, variance, min area, max area) in the units of width * height.
extern "C" int aq_cage_areas(int square, double w, double h, int x_border, int y_border, double out[4]) {
  if (!out) return AQ_ERR_ARG;
  const double pi = 3.141592653589793;
  double lo, hi;
  if (square) {
    lo = w * h * (1.0 / 2);
    hi = w * h;
  } else if (x_border || y_border) {
    lo = h * w / 2;  // approximating triangle
    if (x_border && y_border) hi = pi * h * w / 4;       // quarter ellipse
    else if (x_border) hi = pi * (h / 2) * w / 2;         // half ellipse
    else hi = pi * h * (w / 2) / 2;
  } else {
    const double a = w / 2, b = h / 2;
    out[0] = pi * a * b; out[1] = 0; out[2] = pi * a * b; out[3] = pi * a * b;
    return AQ_OK;
  }
  out[0] = (lo + hi) / 2;
  out[1] = ((hi - lo) * (hi - lo)) / 12;
  out[2] = lo;
  out[3] = hi;
  return AQ_OK;
}

// Facility clustering = get_clusters [REF src/cluster_facilities.py:11-15]: sklearn.cluster.DBSCAN(eps, min_samples)
// .fit(center_points).labels_ on 2-D points (cage centres in metres).  Semantics restated from the published algorithm
// (Ester et al. 1996) in scikit-learn's conventions: a point's neighbourhood holds every point within eps (inclusive, the
// point itself included; squared distances dx*dx + dy*dy compared with eps*eps in float64 as the KD-tree query does),
// core = at least min_samples neighbours, clusters are numbered in order of their lowest-index core point, each cluster
// is grown completely before the next one starts (so a border point two clusters reach keeps the lower label), noise = -1.
// Neighbour search: points sorted by the cell of a uniform grid slightly wider than eps; 3 x 3 cells per query.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

extern "C" int aq_dbscan(const double* xy, int n, double eps, int min_samples, int* labels) {
  if (n < 0 || (n > 0 && (!xy || !labels)) || !(eps > 0) || min_samples < 1) return AQ_ERR_ARG;
  if (n == 0) return AQ_OK;
  double minx = xy[0], miny = xy[1], maxx = xy[0], maxy = xy[1];
  for (int i = 0; i < n; ++i) {
    if (!std::isfinite(xy[2 * i]) || !std::isfinite(xy[2 * i + 1])) return AQ_ERR_ARG;
    minx = std::min(minx, xy[2 * i]); maxx = std::max(maxx, xy[2 * i]);
    miny = std::min(miny, xy[2 * i + 1]); maxy = std::max(maxy, xy[2 * i + 1]);
  }
  const double cell = eps * (1.0 + 1e-9);  // two points within eps are then at most one cell apart despite rounding
  if ((maxx - minx) / cell > 2e9 || (maxy - miny) / cell > 2e9) return AQ_ERR_ARG;
  auto key_of = [](int64_t ix, int64_t iy) { return (uint64_t)(ix + 1) << 32 | (uint64_t)(iy + 1); };
  std::vector<int64_t> cx(n), cy(n);
  std::vector<std::pair<uint64_t, int>> order(n);
  for (int i = 0; i < n; ++i) {
    cx[i] = (int64_t)std::floor((xy[2 * i] - minx) / cell);
    cy[i] = (int64_t)std::floor((xy[2 * i + 1] - miny) / cell);
    order[i] = {key_of(cx[i], cy[i]), i};
  }
  std::sort(order.begin(), order.end());
  const double eps2 = eps * eps;
  // neighbourhoods in CSR form
  std::vector<size_t> first(n + 1, 0);
  std::vector<int> nbr;
  nbr.reserve((size_t)n * 8);
  for (int i = 0; i < n; ++i) {
    first[i] = nbr.size();
    const double x = xy[2 * i], y = xy[2 * i + 1];
    for (int64_t dx = -1; dx <= 1; ++dx)
      for (int64_t dy = -1; dy <= 1; ++dy) {
        const int64_t ix = cx[i] + dx, iy = cy[i] + dy;
        if (ix < 0 || iy < 0) continue;
        const uint64_t k = key_of(ix, iy);
        auto it = std::lower_bound(order.begin(), order.end(), std::make_pair(k, -1));
        for (; it != order.end() && it->first == k; ++it) {
          const double ddx = x - xy[2 * it->second], ddy = y - xy[2 * it->second + 1];
          if (ddx * ddx + ddy * ddy <= eps2) nbr.push_back(it->second);
        }
      }
  }
  first[n] = nbr.size();
  std::vector<char> core(n);
  for (int i = 0; i < n; ++i) {
    core[i] = (first[i + 1] - first[i]) >= (size_t)min_samples;
    labels[i] = -1;
  }
  int label = 0;
  std::vector<int> stack;
  for (int s = 0; s < n; ++s) {
    if (labels[s] != -1 || !core[s]) continue;
    int i = s;
    for (;;) {
      if (labels[i] == -1) {
        labels[i] = label;
        if (core[i])
          for (size_t k = first[i]; k < first[i + 1]; ++k)
            if (labels[nbr[k]] == -1) stack.push_back(nbr[k]);
      }
      if (stack.empty()) break;
      i = stack.back();
      stack.pop_back();
    }
    ++label;
  }
  return label;  // number of clusters
}
