"""TEST INFRASTRUCTURE (oracle) -- CPU restatement of the reference's geocoding / area arithmetic.  Not imported by the
product.

  pixel_box(row)          the four int() lines of geocode_all_detections [REF src/process_yolo/geocode_results.py:158-161]
  pix_to_m(x, y, ...)     convert_pix_to_m_bboxes [REF src/process_yolo/geocode_results.py:71-101]
  circle_area / square_area   get_circle_area_from_bbox / get_square_area_from_bbox
                          [REF src/process_yolo/calc_net_areas.py:24-82]
tests/golden/geo_ref.npz holds outputs of the reference's OWN source of convert_pix_to_m_bboxes and of the two area
functions (extracted with ast by tools/make_golden_geo.py) and pins this file.  The EPSG:3857 -> 3035 step upstream
delegates to pyproj is not restated (pyproj absent from the build container: it could not be pinned).
"""
import math

IM_WIDTH = IM_HEIGHT = 1024          # [REF src/utils.py:18-19]
LARGE_TIF_SIZE = 1024 * 6            # [REF src/utils.py:17]


def pixel_box(bbox):
    xmin = int(IM_WIDTH * (bbox[1] - bbox[3] / 2))
    ymin = int(IM_HEIGHT * (bbox[2] - bbox[4] / 2))
    xmax = int(IM_WIDTH * (bbox[1] + bbox[3] / 2))
    ymax = int(IM_HEIGHT * (bbox[2] + bbox[4] / 2))
    return xmin, ymin, xmax, ymax


def pix_to_m(x, y, x_offset, y_offset, bounds, large_tif_size=LARGE_TIF_SIZE):
    xmin_m, ymin_m, xmax_m, ymax_m = bounds
    x_loc = x + int(x_offset)
    y_loc = y + int(y_offset)
    x_m = x_loc * ((xmax_m - xmin_m) / large_tif_size) + xmin_m
    y_m = ymax_m - y_loc * ((ymax_m - ymin_m) / large_tif_size)
    return x_m, y_m


def geocode_rows(rows, x_offset, y_offset, bounds):
    """rows: iterable of (cls, cx, cy, w, h, conf) float64 as np.loadtxt returns them."""
    out = []
    for b in rows:
        xmin, ymin, xmax, ymax = pixel_box(b)
        xmin_m, ymax_m = pix_to_m(xmin, ymin, x_offset, y_offset, bounds)
        xmax_m, ymin_m = pix_to_m(xmax, ymax, x_offset, y_offset, bounds)
        out.append(dict(cls=int(b[0]), xmin=xmin, xmax=xmax, ymin=ymin, ymax=ymax, conf=float(b[5]), xmin_m=xmin_m,
                        ymin_m=ymin_m, xmax_m=xmax_m, ymax_m=ymax_m))
    return out


def circle_area(w, h, x_border=False, y_border=False):
    if x_border or y_border:
        lo = h * w / 2
        if x_border and y_border:
            hi = math.pi * h * w / 4
        elif x_border:
            hi = math.pi * (h / 2) * w / 2
        else:
            hi = math.pi * h * (w / 2) / 2
        return (lo + hi) / 2, ((hi - lo) ** 2) / 12, lo, hi
    a, b = w / 2, h / 2
    return math.pi * a * b, 0, math.pi * a * b, math.pi * a * b


def square_area(w, h):
    lo = w * h * (1 / 2)
    hi = w * h
    return (lo + hi) / 2, ((hi - lo) ** 2) / 12, lo, hi
