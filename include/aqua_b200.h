/*
 * aqua_b200.h — C ABI of the B200-native YOLOv5m tile-detection path.
 *
 * This library replaces the compute behind the reference's one GPU-shaped stage,
 *     python3 yolov5/detect.py --weights ...exp2.pt --source data/jpegs --nosave --save-txt --save-conf
 *                                                                    [REF README.md:77]
 * The reference exposes NO plugin / FFI interface for this path (it is a CLI on one side and a directory
 * of label files on the other, consumed by [REF src/process_yolo/geocode_results.py:140-172]); the entry
 * points below are therefore the boundary a maintainer binds with ctypes from a detect.py replacement
 * (INTEGRATION.md).  Each entry cites the upstream function whose arithmetic it replaces
 * ([UPSTREAM-RECALLED] = ultralytics/yolov5, the absent submodule of [REF .SUBMODULES.json:8-16]).
 *
 * Conventions: plain pointers and sizes only; every function returns AQ_OK (0) or a negative aq_status
 * and never throws; aq_last_error() gives the text.  The caller owns every buffer it passes; the context
 * owns its workspace, weights and activation buffers.  One context per (process, device); not thread
 * safe.  `stream` is a cudaStream_t passed as void* (NULL = the context's own stream).  There is no CPU
 * fallback: without a CUDA device aq_create fails.
 */
#ifndef AQUA_B200_H
#define AQUA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AQ_ABI_VERSION 5

typedef enum aq_status {
  AQ_OK = 0,
  AQ_ERR_ARG = -1,          /* bad argument */
  AQ_ERR_CUDA = -2,         /* CUDA runtime / driver error (text in aq_last_error) */
  AQ_ERR_STATE = -3,        /* call out of order (e.g. forward before plan_end) */
  AQ_ERR_NOMEM = -4,
  AQ_ERR_UNSUPPORTED = -5,  /* shape or mode this build has no kernel for; never a silent fallback */
  AQ_ERR_IO = -6
} aq_status;

typedef struct aq_ctx aq_ctx;

/* Activation flags of a conv op. */
#define AQ_ACT_NONE 0
#define AQ_ACT_SILU 1       /* x * sigmoid(x) with ex2 + rcp (~1e-6 absolute) */
#define AQ_ACT_SILU_FAST 2  /* h + h * tanh.approx(h), h = x / 2: one MUFU per element, <= 1.1e-5 absolute
                               (profiles/r1_hw_probe_v1.txt T5), below the fp16 rounding of the stored value */

/*
 * One fused convolution = upstream models/common.py::Conv.forward_fuse (conv + folded BN bias + SiLU)
 * optionally followed by models/common.py::Bottleneck's residual add, executed as an implicit GEMM on
 * tcgen05 tensor cores.  Activations are NHWC fp16; an op reads channels [src_c_off, src_c_off+c_in) of
 * buffer `src` and writes channels [dst_c_off, dst_c_off+c_out) of buffer `dst`, which is how
 * torch.cat / models/common.py::Concat disappear (producers write into slices of the concat buffer).
 */
typedef struct aq_conv_desc {
  int src, src_c_off, c_in;     /* input buffer id, first channel, channels (multiple of 16) */
  int dst, dst_c_off, c_out;    /* output buffer id, first channel, channels */
  int k;                        /* 1 or 3 (the 6x6/s2 stem is given as 3x3/s1 over the space-to-depth input) */
  int stride;                   /* 1 or 2 */
  int act;                      /* AQ_ACT_* */
  int res, res_c_off;           /* residual buffer id (or -1) and first channel; added AFTER the activation.
                                   Must name the destination slice itself (Bottleneck runs in place in the C3
                                   concat buffer): the add is a TMA reduce-add into dst, it costs no load. */
  const uint16_t* weight;       /* host, fp16 bits, [c_out][k][k][c_in] (KRSC), BN already folded */
  const float* bias;            /* host, fp32 [c_out] */
  /* tuning knobs; 0 = let the library choose */
  int tile_n;                   /* UMMA N per tile */
  int mblk;                     /* 128-pixel M blocks per CTA tile (1 or 2) */
  int mode;                     /* 3x3/s1 only: 1 = halo patch in shared memory, 2 = one box per tap */
  int cluster;                  /* CTAs per thread-block cluster sharing each weight tile by TMA multicast (1, 2, 4) */
  /* nn.Upsample(2,'nearest') + Concat folded into this conv's loads (1x1 convs only; up_c = 0: off): input channels
   * [0, up_c) are the nearest-2x upsampling of channels [up_c_off, up_c_off+up_c) of the half-resolution buffer
   * up_src, read through a TMA tensor map whose replication dimensions have stride 0; channels [up_c, c_in) are
   * src[src_c_off, src_c_off + c_in - up_c).  Neither the upsampled tensor nor the concat exists in memory. */
  int up_src, up_c_off, up_c;
} aq_conv_desc;

/*
 * Detect head level = upstream models/yolo.py::Detect.forward (eval branch): 1x1 conv with bias, then
 * sigmoid, xy = (s*2 + grid - 0.5) * stride, wh = (s*2)^2 * anchor_px, conf = s, written as fp32 rows
 * [level_row_off + a*ny*nx + y*nx + x][5+nc] of the prediction tensor (B, rows, 5+nc).
 */
typedef struct aq_detect_desc {
  int src, src_c_off, c_in;
  int nc;                       /* classes; no = nc + 5; conv N = 3*no */
  int level_row_off;            /* first row of this level inside one image's prediction */
  float stride;                 /* 8, 16, 32 */
  float anchors_px[6];          /* 3 anchors (w,h) in pixels */
  const uint16_t* weight;       /* host fp16 [3*no][c_in] */
  const float* bias;            /* host fp32 [3*no] */
} aq_detect_desc;

/* ------------------------------------------------------------------ lifetime */
int aq_abi_version(void);
/* Fails with AQ_ERR_CUDA when no sm_100 device is usable: there is no CPU path. */
int aq_create(int device, aq_ctx** out);
void aq_destroy(aq_ctx* ctx);
const char* aq_last_error(const aq_ctx* ctx); /* ctx may be NULL: error of the last failed aq_create */

/* Tiling table: one "signature choice" line per conv shape (what AQUA_TUNE_CACHE=<file> collects on a B200).  Shapes the
 * table lists are planned with the listed tiling -- the same in every process and on every rank, so label text does not
 * depend on plan-time timing noise; a shape it does not list is timed on the device at plan time and reported on stderr
 * (AQUA_TUNE_LOG=0 silences it; AQUA_AUTOTUNE=0 uses the cost model instead, also deterministic).  Returns the number of
 * entries added.  aq_tune_misses: convs of this context that had to be timed because no entry matched. */
int aq_tune_load(aq_ctx* ctx, const char* path);
int aq_tune_misses(aq_ctx* ctx);

/* ------------------------------------------------------------------ plan (replaces DetectionModel.__init__/fuse/_forward_once wiring) */
/* Begin a plan for `batch` tiles whose letterboxed network input is net_h x net_w (multiples of 32). */
int aq_plan_begin(aq_ctx* ctx, int batch, int net_h, int net_w, int nc);
/* Declare an NHWC fp16 activation buffer (h, w, c); returns its id >= 0.  Buffer 0 must be the network
 * input in space-to-depth form: (net_h/2, net_w/2, 16) with channel (py*2+px)*3 + rgb, 12..15 zero. */
int aq_plan_buffer(aq_ctx* ctx, int h, int w, int c);
/* Depth-first execution of a run of layers (no upstream counterpart: upstream runs every layer over the whole batch).
 * The ops planned between _begin and _end form a stage that aq_forward executes sub-batch by sub-batch (`sub_batch`
 * images at a time, all ops of the stage per sub-batch); buffers declared inside the stage are TRANSIENT: they hold
 * one sub-batch per lane, are visible to the ops of this stage only and are overwritten by the next sub-batch, so
 * the activations between the stage's layers stay in the 126 MB L2 instead of travelling to HBM and back.  Buffers
 * declared outside are entered at image s * sub_batch by sub-batch s.  `lanes` (1..4) sub-batches run as independent
 * chains on parallel branches of the captured graph, filling each other's launch / drain bubbles.  Results are
 * identical to the unstaged plan (same kernels on the same values).  sub_batch >= batch, or not a divisor of it:
 * the ops run as ordinary full-batch ops.  Stages do not nest. */
int aq_plan_stage_begin(aq_ctx* ctx, int sub_batch, int lanes);
int aq_plan_stage_end(aq_ctx* ctx);
int aq_plan_conv(aq_ctx* ctx, const aq_conv_desc* d);
/* Two convs in ONE kernel: `first` as aq_plan_conv, then `second` -- a 1x1 / stride-1 conv (+ bias + activation, no residual) whose
 * input is the first second->c_in channels of first's output slice -- evaluated on the 128 x 64 fp16 groups first's epilogue
 * stages in shared memory, before they leave for HBM (a second tcgen05.mma on the staged tiles; conv_igemm.cuh EPI_CHAIN).
 * upstream: models/common.py::C3.forward's cv1 followed by m[0].cv1, or any Conv feeding a pointwise Conv.  Values are those of
 * two separate aq_plan_conv calls (the second reads the same fp16-rounded activations); store_first = 0 keeps first's output
 * out of memory altogether.  first: no residual, <= 256 output channels, FLAT / TAP feeds; AQ_ERR_UNSUPPORTED otherwise (plan
 * the two convs separately then). */
int aq_plan_conv_chain(aq_ctx* ctx, const aq_conv_desc* first, const aq_conv_desc* second, int store_first);
/* models/common.py::SPPF's three chained MaxPool2d(5,1,2): reads channels [c_off, c_off+c) of `buf` and
 * writes the pooled maps to [c_off+c, c_off+2c), [c_off+2c, ..), [c_off+3c, ..) of the same buffer. */
int aq_plan_sppf_pool(aq_ctx* ctx, int buf, int c_off, int c);
/* nn.Upsample(scale_factor=2, mode='nearest') into a channel slice of the consumer's concat buffer. */
int aq_plan_upsample2x(aq_ctx* ctx, int src, int src_c_off, int c, int dst, int dst_c_off);
int aq_plan_detect(aq_ctx* ctx, const aq_detect_desc* d);
/* Allocate, upload weights, encode TMA tensor maps, capture the forward as a CUDA graph. */
int aq_plan_end(aq_ctx* ctx);
/* Introspection for tests: device pointer and geometry of a planned buffer. */
int aq_buffer_info(aq_ctx* ctx, int buf, void** dev_ptr, int* n, int* h, int* w, int* c);
/* Device pointer of the decoded predictions (batch, rows, 5+nc) fp32 written by aq_forward. */
int aq_pred_info(aq_ctx* ctx, float** dev_ptr, int* rows, int* no);
int aq_plan_num_launches(aq_ctx* ctx); /* kernels launched by one aq_forward */
/* Measurement aids: time every planned op alone (CUDA events on the context stream, `iters` launches each,
 * after one warm-up) into ms_per_op[0..n_ops); returns n_ops.  aq_op_describe writes a one-line description
 * (kind, GEMM shape, tile configuration, algorithmic FLOPs and bytes) of op `i`. */
int aq_profile_ops(aq_ctx* ctx, int iters, float* ms_per_op, int cap);
/* Same table measured in situ: `iters` whole forward passes with an event after every op (real cache state). */
int aq_profile_ops_insitu(aq_ctx* ctx, int iters, float* ms_per_op, int cap);
int aq_op_describe(aq_ctx* ctx, int i, char* buf, size_t cap);
/* Launch-overhead probe: conv op `i` cut down to its first waves x SMs tiles, `iters` launches back to back; *ms_out =
 * time per launch, returns the tiles launched.  Slope over `waves` = one wave, intercept = fixed drain + fill cost of a
 * launch.  Leaves the op's output partly unwritten: measurement only. */
int aq_profile_op_waves(aq_ctx* ctx, int i, int waves, int iters, float* ms_out);
/* Per-role timeline of planned conv op `i` (instrumented build of the same kernel, launched once after one
 * warm-up): out[grid][6 roles][48 tiles][4] clock64 values (layout in conv_igemm.cuh); returns the grid size
 * (CTAs) or a negative status; cap_vals = capacity of `out` in 64-bit values. */
int aq_trace_op(aq_ctx* ctx, int i, long long* out, size_t cap_vals);
/* Host-only planner probe (no device needed): the tiling the library would choose for one conv of the given
 * shape, written as the same one-line description aq_op_describe gives.  mblk / mode = 0 lets it choose. */
int aq_debug_plan_conv(int batch, int h, int w, int c_in, int c_out, int k, int stride, int num_sms, int mblk, int mode,
                       char* buf, size_t cap);

/* ------------------------------------------------------------------ execution */
/*
 * Letterbox + to-tensor: upstream utils/augmentations.py::letterbox (cv2.resize INTER_LINEAR in cv2's
 * 11-bit fixed point + copyMakeBorder 114) fused with LoadImages' HWC->CHW/BGR->RGB and detect.py's
 * float()/255 [UPSTREAM-RECALLED].  `tiles` = n uint8 HWC images of in_h x in_w x 3 (device pointer, or
 * pinned/pageable host pointer when src_is_device == 0, in which case the H2D copy is issued here).
 * Result: buffer 0 of the plan.  `bgr` != 0 means the source is BGR (cv2.imread order).
 */
int aq_preprocess(aq_ctx* ctx, const uint8_t* tiles, int n, int in_h, int in_w, int bgr,
                  int src_is_device, void* stream);
/* Letterboxed uint8 image (n, net_h, net_w, 3) in RGB order, for bit-exact checks of the resize/pad. */
int aq_preprocess_u8(aq_ctx* ctx, const uint8_t* tiles_dev, int n, int in_h, int in_w, int bgr,
                     uint8_t* out_dev, void* stream);
/* Conv stack + Detect decode (one CUDA graph launch). */
int aq_forward(aq_ctx* ctx, void* stream);
/*
 * Batched NMS = upstream utils/general.py::non_max_suppression (+ xywh2xyxy, torchvision.ops.nms)
 * [UPSTREAM-RECALLED].  pred: device fp32 (batch, rows, no) or NULL to use the plan's own predictions.
 * dets_out: device fp32 (batch, max_det, 6) rows [x1,y1,x2,y2,conf,cls] in descending confidence;
 * counts_out: device int32 (batch); keep_rows_out: device int32 (batch, max_det) source row of each kept
 * detection (may be NULL).  classes: host int array or NULL.  Keep indices are bit-exact w.r.t. the
 * reference semantics: stable descending sort, fp32 IoU inter/(a_i+a_j-inter), suppress iff
 * (double)iou > iou_thres.
 */
int aq_nms(aq_ctx* ctx, const float* pred, int batch, int rows, int no, float conf_thres, double iou_thres,
           int max_det, int agnostic, const int* classes, int n_classes, float max_wh, int max_nms,
           float* dets_out, int* counts_out, int* keep_rows_out, void* stream);
/*
 * Whole path for one batch: preprocess -> forward -> NMS -> copy detections to host.
 * dets_host (batch, max_det, 6) fp32 and counts_host (batch) int32 should be pinned for overlap.
 * Synchronises `stream` before returning.
 */
int aq_detect(aq_ctx* ctx, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device,
              float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes,
              int n_classes, float* dets_host, int* counts_host, void* stream);
/* Asynchronous halves of aq_detect for double-buffered pipelines.  aq_detect_submit enqueues the whole batch
 * (kernels and the D2H copies on `stream`; for host tiles the H2D copy goes through one of two staging buffers on
 * the library's own copy stream, so the copy of batch i+1 overlaps the kernels of batch i) and returns a ticket
 * (0 or 1, alternating; negative = aq_status error).  aq_detect_wait blocks until the batch of that ticket has
 * its detections in dets_host / counts_host.  At most two batches may be in flight; each needs its own
 * dets_host / counts_host, and the host tiles of a batch must stay valid until its ticket has been waited for. */
int aq_detect_submit(aq_ctx* ctx, const uint8_t* tiles, int n, int in_h, int in_w, int bgr,
                     int src_is_device, float conf_thres, double iou_thres, int max_det, int agnostic,
                     const int* classes, int n_classes, float* dets_host, int* counts_host, void* stream);
int aq_detect_wait(aq_ctx* ctx, int ticket);
/* Device time of the three stages upstream's detect.py reports in its "Speed: %.1fms pre-process, %.1fms inference,
 * %.1fms NMS per image" line, for the batch behind `ticket` (after aq_detect_wait): ms[0] letterbox + normalise kernel,
 * ms[1] conv stack + Detect decode, ms[2] NMS -- CUDA events on the compute stream, whole batch (divide by n). */
int aq_detect_stage_ms(aq_ctx* ctx, int ticket, float ms[3]);

/* ------------------------------------------------------------------ host-side label epilogue (no GPU) */
/*
 * upstream utils/general.py::{scale_boxes, clip_boxes, xyxy2xywh} + detect.py::run's --save-txt block
 * [UPSTREAM-RECALLED]; grammar pinned by the consumer [REF src/process_yolo/geocode_results.py:140-172].
 * For image i: dets + i*det_stride*6 holds counts[i] rows; boxes are rescaled from the letterboxed
 * (net_h, net_w) to the original (h0[i], w0[i]), clipped, rounded, normalised; lines are emitted in
 * ASCENDING confidence as "cls cx cy w h[ conf]\n" with "%g"; zero detections produce NO file; files
 * are opened in append mode as upstream does.  paths[i] = full path of the label file.
 * Returns the number of files written (>= 0) or a negative aq_status.
 */
int aq_format_labels(const float* dets, const int* counts, int n_images, int det_stride, int net_h, int net_w,
                     const int* h0, const int* w0, const char* const* paths, int save_conf, int n_threads);
/* Same text into a caller buffer (for tests): returns bytes written or negative status. */
int aq_format_label_text(const float* dets, int count, int net_h, int net_w, int h0, int w0, int save_conf,
                         char* out, size_t out_cap);

/* ------------------------------------------------------------------ post-processing arithmetic */
typedef struct aq_geo_row {
  int cls;                      /* bbox[0] */
  int xmin, xmax, ymin, ymax;   /* pixel box: int(IM_WIDTH * (cx -+ w/2)) ... [REF src/process_yolo/geocode_results.py:158-161] */
  double conf;                  /* bbox[5] */
  double xmin_m, ymin_m, xmax_m, ymax_m; /* EPSG:3857 metres [REF src/process_yolo/geocode_results.py:96-99,167-168] */
} aq_geo_row;
/*
 * Geocode the rows of one label file (its text, as written by aq_format_labels with save_conf = 1).  x_offset, y_offset:
 * the tile's origin inside its large tif (fields 2, 3 of the file stem), bounds_3857 = (xmin, ymin, xmax, ymax) of that
 * tif's download box, large_tif_size = 6144, im_width = im_height = 1024 in the reference [REF src/utils.py:17-19].
 * Returns the number of rows written to out (<= cap) or a negative aq_status.  Host only; needs no context.
 */
int aq_geocode_label_text(const char* text, int x_offset, int y_offset, const double bounds_3857[4], int large_tif_size,
                          int im_width, int im_height, aq_geo_row* out, int cap);
/* One geocoded detection of a batch: what geocode_all_detections appends to its frame per label row
 * [REF src/process_yolo/geocode_results.py:158-193]. */
typedef struct aq_geo_det {
  int image;                    /* index of the tile inside the batch handed in */
  int cls;                      /* bbox[0]; aq_geo_class_name(cls) = REVERSE_CLASS_MAPPING[cls] (:24-30,172) */
  int year;                     /* int(stem.split('_')[0][-4:]) (:143) */
  int xmin, xmax, ymin, ymax;   /* pixel box (:158-161) */
  double conf;                  /* bbox[5] */
  double xmin_m, ymin_m, xmax_m, ymax_m; /* EPSG:3857 metres (:167-168) -- the frame's `geometry` box */
  double xmin_a, ymin_a, xmax_a, ymax_a; /* EPSG:3035 metres (:175-177) -- the frame's xmin_m .. ymax_m columns */
} aq_geo_det;
/*
 * geocode_all_detections fed by the detector's output instead of the label directory: dets / counts / det_stride / net_h /
 * net_w / h0 / w0 exactly as given to aq_format_labels (which still writes the .txt files), stems[i] = the tile's file stem
 * "<layer><year>_<download box index>_<x offset>_<y offset>" [REF src/load_data/tile_tifs.py:33-41], download_bounds_3857 =
 * (xmin, ymin, xmax, ymax) of every download box, indexed as wanted_bboxes.loc[ind].  One call per batch (or per run); no
 * text files are read and no images are opened.  Values are identical to what the text hop yields (the '%g' print /
 * parse round trip is done in memory).  Returns the number of rows written (<= cap) or a negative aq_status; a stem that
 * upstream's split / int() would raise on returns AQ_ERR_ARG.  Host only; needs no context.
 */
int aq_geocode_dets(const float* dets, const int* counts, int n_images, int det_stride, int net_h, int net_w, const int* h0,
                    const int* w0, const char* const* stems, const double* download_bounds_3857, int n_boxes,
                    int large_tif_size, int im_width, int im_height, aq_geo_det* out, int cap);
const char* aq_geo_class_name(int cls); /* NULL where REVERSE_CLASS_MAPPING has no key */
/* EPSG:3857 -> EPSG:3035 (Transformer.from_crs(CRS_DICT['mapping'], CRS_DICT['area']), :31): n (x, y) pairs.  Closed-form
 * EPSG methods 1024 (inverse) + 9820; pinned by the published EPSG test point, not bit-pinned against PROJ. */
int aq_transform_3857_to_3035(const double* xy_in, int n, double* xy_out);
int aq_lonlat_to_3035(const double* lonlat_deg, int n, double* xy_out); /* the LAEA leg alone (degrees on GRS 1980) */
/* Cage area estimate from a box in metres: square = 0 -> get_circle_area_from_bbox, 1 -> get_square_area_from_bbox
 * [REF src/process_yolo/calc_net_areas.py:24-82]; out = (estimate, variance, min, max). */
int aq_cage_areas(int square, double w, double h, int x_border, int y_border, double out[4]);

/* Facility clustering = get_clusters [REF src/cluster_facilities.py:11-15], i.e. sklearn.cluster.DBSCAN(eps = distance
 * threshold in metres, min_samples).fit(center_points).labels_: xy = n points (x, y) float64; labels receives n int32
 * (cluster number in sklearn's numbering, -1 = noise).  Returns the number of clusters or a negative aq_status.
 * Host only; needs no context. */
int aq_dbscan(const double* xy, int n, double eps, int min_samples, int* labels);

/* ------------------------------------------------------------------ JPEG tiles */
/*
 * Decoder for the tiles gdal.Translate -of JPEG writes [REF src/load_data/tile_tifs.py:74] and upstream
 * utils/dataloaders.py::LoadImages reads with cv2.imread [UPSTREAM-RECALLED].  Two ways in: the all-device decoder
 * (aq_jpeg_scan_prepare + aq_jpeg_decode_device, further down: the host only walks the markers, the GPU undoes the Huffman
 * coding too) and the hybrid one declared first (the host undoes the Huffman coding, the GPU does the arithmetic:
 * dequantisation, the 8x8 inverse DCT, chroma upsampling and YCbCr -> RGB).  Results are bit-identical to cv2.imread /
 * libjpeg's defaults (accurate integer IDCT "islow", triangle-filter "fancy" upsampling, 16-bit fixed-point colour
 * conversion).  Supported: baseline / 8-bit
 * extended sequential Huffman files, one interleaved scan, grayscale, YCbCr 4:4:4 and 4:2:0, restart intervals.
 * Anything else (progressive, arithmetic, CMYK, 4:2:2 ...) returns AQ_ERR_UNSUPPORTED -- the caller decides, nothing is
 * decoded silently some other way.
 */
typedef struct aq_jpeg_info {
  int width, height;
  int ncomp;            /* 1 or 3 */
  int hs, vs;           /* luma sampling factors: (1,1) or (2,2) */
  int mcus_x, mcus_y;   /* MCU grid: ceil(width / (8 hs)) x ceil(height / (8 vs)) */
  int restart_interval;
  size_t coef_count;    /* int16 values aq_jpeg_entropy_decode writes */
} aq_jpeg_info;
/* Host only; needs no context. */
int aq_jpeg_probe(const uint8_t* data, size_t len, aq_jpeg_info* info);
/*
 * Host only.  coef: component planes one after another (Y, then Cb, Cr), each a raster of 8x8 blocks
 * [block row][block column][64 coefficients in natural (row-major) order], luma plane (mcus_y vs) x (mcus_x hs) blocks,
 * chroma planes mcus_y x mcus_x blocks; quantised values exactly as coded (DC prediction undone).
 * quant: 3 x 64 uint16 quantisation steps in natural order (table of component 0, 1, 2).
 */
int aq_jpeg_entropy_decode(const uint8_t* data, size_t len, int16_t* coef, size_t coef_cap, uint16_t* quant);
/*
 * Same decoding into the PACKED form the pipelined path ships to the GPU (most quantised coefficients are zero: a
 * 1024 x 1024 tile at quality 75 is ~0.7 MB packed against 3.1 MB dense).  block_desc[b] for block b (same plane /
 * raster order as above, coef_count / 64 blocks) = (index of the block's first entry << 7) | number of entries;
 * entries[i] = (natural-order position << 16) | (uint16_t) value, in decoding order.  entries_cap >= coef_count is
 * always enough.  *n_entries receives the number of entries written.  Host only.
 */
int aq_jpeg_entropy_decode_packed(const uint8_t* data, size_t len, uint32_t* block_desc, size_t block_cap, uint32_t* entries,
                                  size_t entries_cap, size_t* n_entries, uint16_t* quant);
/*
 * Device half: n images of identical geometry (as tiles are).  coef = n x coef_count int16 and quant = n x 192 uint16 as
 * written by aq_jpeg_entropy_decode (device pointers, or host pointers when src_is_device == 0: the H2D copies are issued
 * here).  out_dev: (n, height, width, 3) uint8 in BGR order (bgr_out != 0, what cv2.imread returns) or RGB; grayscale
 * files give three equal channels as cv2.IMREAD_COLOR does.  Feed out_dev to aq_preprocess / aq_detect with
 * src_is_device = 1.  Does not synchronise.
 */
int aq_jpeg_reconstruct(aq_ctx* ctx, const int16_t* coef, const uint16_t* quant, int n, int width, int height, int ncomp,
                        int hs, int bgr_out, int src_is_device, uint8_t* out_dev, void* stream);
/*
 * Device half for packed input, pipelined: block_desc = n x (coef_count / 64) words, entries = the n tiles' entry lists
 * one after another, tile_entry_base[i] = index in `entries` where tile i's list starts, total_entries = their sum;
 * all HOST pointers (pinned for overlap).  The H2D copies are issued on the library's copy stream into one of two
 * alternating device staging areas, so the copies of batch i+1 overlap the kernels of batch i; the kernels run on
 * `stream`.  The host buffers of a call may be reused once the batch that consumed out_dev has been waited for
 * (aq_detect_wait) or after aq_sync.  Does not synchronise.
 */
int aq_jpeg_reconstruct_packed(aq_ctx* ctx, const uint32_t* block_desc, const uint32_t* entries, const uint64_t* tile_entry_base,
                               size_t total_entries, const uint16_t* quant, int n, int width, int height, int ncomp, int hs,
                               int bgr_out, uint8_t* out_dev, void* stream);
/*
 * Entropy decoding ON THE DEVICE (self-synchronising parallel Huffman decoding; csrc/jpeg_huff_core.h): the host only walks
 * the markers.  aq_jpeg_scan_prepare (host only): geometry, the entropy-coded segment with its byte stuffing removed
 * (scan_out: 4-byte aligned, scan_cap >= file length + 16; *scan_bytes = its length, zero padded to whole words + 8),
 * the tile's six Huffman tables in device form (tables_out: AQ_JPEG_TABLE_BYTES, 4-byte aligned) and the quantisation
 * tables (quant: 192 uint16).  Files with restart intervals return AQ_ERR_UNSUPPORTED here (use the host decoder).
 */
#define AQ_JPEG_TABLE_BYTES 8640
int aq_jpeg_scan_prepare(const uint8_t* data, size_t len, aq_jpeg_info* info, uint8_t* scan_out, size_t scan_cap,
                         size_t* scan_bytes, uint8_t* tables_out, uint16_t* quant);
/* The same for a batch in one call, written straight into the staging areas aq_jpeg_decode_device takes and spread over
 * n_threads host threads: tile i's segment goes to scans + scan_off[i] (offsets are chosen here from the file lengths;
 * scans_cap >= sum over i of (lens[i] + 16 rounded up to 4) always suffices), *total_bytes = the extent used; infos /
 * status receive n entries (status[i] = that tile's aq_status).  Returns AQ_OK or the first failing tile's status. */
int aq_jpeg_scan_prepare_batch(const uint8_t* const* datas, const size_t* lens, int n, uint8_t* scans, size_t scans_cap,
                               uint64_t* scan_off, uint32_t* scan_bytes, size_t* total_bytes, uint8_t* tables, uint16_t* quant,
                               aq_jpeg_info* infos, int* status, int n_threads);
/*
 * n tiles of identical geometry: scans = their prepared segments one after another, tile i at byte scan_off[i] (a multiple
 * of 4) with scan_bytes[i] bytes, total_scan_bytes = extent of `scans` (padding included); tables = n x
 * AQ_JPEG_TABLE_BYTES; quant = n x 192; all HOST pointers (pinned for overlap; copied on the library's copy stream into
 * one of two alternating staging areas).  Kernels on `stream`: parallel Huffman decoding -> dense coefficients ->
 * aq_jpeg_reconstruct's kernels -> out_dev as there.  status_host (n int32, pinned; may be NULL) receives, asynchronously
 * on `stream`, 0 per tile or AQ_ERR_IO when a tile's scan did not decode to the expected number of blocks (read it after
 * the batch has been waited for).  Does not synchronise.
 */
int aq_jpeg_decode_device(aq_ctx* ctx, const uint8_t* scans, const uint64_t* scan_off, const uint32_t* scan_bytes,
                          size_t total_scan_bytes, const uint8_t* tables, const uint16_t* quant, int n, int width, int height,
                          int ncomp, int hs, int bgr_out, uint8_t* out_dev, int* status_host, void* stream);
/* CPU rehearsal of the device schedule with the same decoding core (tests without a GPU; not a product path): decodes one
 * prepared scan with subsequences of sub_bits bits (multiple of 32) into dense coefficients laid out as
 * aq_jpeg_entropy_decode writes them; *rounds_out = relaxation rounds needed. */
int aq_jpeg_debug_parallel_decode(const uint8_t* scan, size_t scan_bytes, const uint8_t* tables, const aq_jpeg_info* info,
                                  int sub_bits, int16_t* coef, size_t coef_cap, int* rounds_out);

/* ------------------------------------------------------------------ misc */
/*
 * Tiling fused into inference: detect on windows of one device-resident raster instead of tiles cut
 * by src/load_data/tile_tifs.py::split_all_tiffs [REF src/load_data/tile_tifs.py:33-47].  raster_dev: (raster_h,
 * raster_w, 3) uint8.  Window k of the grid of FULL win_h x win_w windows is enumerated like the reference's loop
 * (x outer, y inner): k = ix * wins_y + iy, origin (x, y) = (ix * win_w, iy * win_h), i.e. the tile the reference
 * names `<tif>_<x>_<y>`.  Windows [win0, win0 + n) are letterboxed to the plan's network size, run through the conv
 * stack and NMS; results as aq_detect.  Ragged right / bottom strips are not covered (cut them on the host and use
 * aq_detect).  Synchronises `stream`.  Returns a ticket >= 0 (for aq_detect_stage_ms) or a negative aq_status.
 */
int aq_detect_windows(aq_ctx* ctx, const uint8_t* raster_dev, int raster_h, int raster_w, int win_h, int win_w, int win0,
                      int n, int bgr, float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes,
                      int n_classes, float* dets_host, int* counts_host, void* stream);

/*
 * The same for an arbitrary list of windows: tile i of the batch is window win_idx[i] (host array, n <= plan batch) of the
 * grid described above.  This is how --skip-blank works: aq_tile_stats measures every window first and the windows
 * src/utils.py::is_blank calls blank [REF src/utils.py:325-350] never reach the network -- the reference does not even
 * download such images [REF src/load_data/download_french_data.py:58-60].
 */
int aq_detect_window_list(aq_ctx* ctx, const uint8_t* raster_dev, int raster_h, int raster_w, int win_h, int win_w,
                          const int* win_idx, int n, int bgr, float conf_thres, double iou_thres, int max_det, int agnostic,
                          const int* classes, int n_classes, float* dets_host, int* counts_host, void* stream);
/* Letterboxed uint8 windows (n, net_h, net_w, 3) RGB, for bit-exact checks against cropping + aq_preprocess_u8. */
int aq_preprocess_windows_u8(aq_ctx* ctx, const uint8_t* raster_dev, int raster_h, int raster_w, int win_h, int win_w,
                             int win0, int n, int bgr, uint8_t* out_dev, void* stream);
/*
 * Blank-tile statistics = the measurements behind src/utils.py::is_blank / is_partly_blank
 * [REF src/utils.py:325-369].  src_dev: n tiles (n, in_h, in_w, 3) uint8, or -- when raster_h > 0 -- a raster of which
 * windows [win0, win0 + n) of size in_h x in_w are taken as above.  stats_host receives n x 4 int32:
 *   [0] min, [1] max of PIL's "L" luminance (19595 R + 38470 G + 7471 B + 0x8000) >> 16   (is_blank's getextrema())
 *   [2] rows, [3] columns whose mean over pixels and channels is >= 250                    (is_partly_blank)
 * Tiles up to 2048 wide and 4096 tall.  Needs no plan.  Synchronises `stream` (stats_host == NULL: kernels only, no
 * copy, no synchronisation -- for timing).
 */
int aq_tile_stats(aq_ctx* ctx, const uint8_t* src_dev, int n, int in_h, int in_w, int bgr, int raster_h, int raster_w,
                  int win0, int* stats_host, void* stream);

/*
 * aq_tile_stats for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same arguments
 * and results as aq_tile_stats, except that
 *   - the column-sum scratch is a file-scope static __device__ array inside the library (512 KB, held by every GPU that
 *     loads the library until the process exits); its device address is resolved once per device at aq_create;
 *   - n * in_w must not exceed AQ_TILE_STATS_STATIC_MAX_COLS (e.g. 64 tiles 2048 wide, 128 tiles 1024 wide); larger
 *     requests return AQ_ERR_ARG with a message naming the limit -- use aq_tile_stats for those;
 *   - stats_dev is a DEVICE buffer of n x 4 int32 owned by the caller; nothing is copied to the host and `stream` is
 *     not synchronised;
 *   - the call performs no cudaMalloc / cudaFree and calls no allocating library routine: one cudaMemsetAsync (the
 *     part of the shared array it uses must start at zero) and four kernel launches on `stream`.
 * Every call on a device shares the one array: calls that overlap on the GPU (different streams, two contexts of one
 * device) race on it; order them on one stream or with events.
 */
#define AQ_TILE_STATS_STATIC_MAX_COLS 131072
int aq_tile_stats_static(aq_ctx* ctx, const uint8_t* src_dev, int n, int in_h, int in_w, int bgr, int raster_h,
                         int raster_w, int win0, int* stats_dev, void* stream);
/* cudaMemcpyAsync wrapper so hosts without a CUDA binding can move test data: kind 0=H2D 1=D2H 2=D2D. */
int aq_memcpy(aq_ctx* ctx, void* dst, const void* src, size_t bytes, int kind, void* stream);
int aq_sync(aq_ctx* ctx, void* stream);
/* Elapsed GPU milliseconds of the last aq_forward's conv graph per op class is not exposed; use events. */

#ifdef __cplusplus
}
#endif
#endif /* AQUA_B200_H */
