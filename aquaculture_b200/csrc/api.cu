// C ABI of the tile-detection path (include/aqua_b200.h): context, plan builder, TMA tensor-map encoding,
// CUDA-graph capture of the conv stack, and the batch entry points.
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/aqua_b200.h"
#include "aux_kernels.cuh"
#include "conv_igemm.cuh"
#include "nms.cuh"

using namespace aq;

namespace {

thread_local std::string g_create_error;

struct Buffer {
  int h, w, c;
  __half* ptr = nullptr;
};

enum OpKind { OP_CONV, OP_DETECT, OP_POOL, OP_UPSAMPLE };

struct Op {
  OpKind kind;
  // conv / detect
  aq_conv_desc conv{};
  aq_detect_desc det{};
  std::vector<uint16_t> w_host;
  std::vector<float> b_host;
  __half* w_dev = nullptr;
  float* b_dev = nullptr;
  ConvKParams kp{};
  // pool / upsample
  int buf = 0, c_off = 0, c = 0, dst = 0, dst_c_off = 0;
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

}  // namespace

struct aq_ctx {
  int device = 0;
  int num_sms = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  EncodeTiledFn encode = nullptr;
  // plan
  bool planning = false, planned = false;
  int batch = 0, net_h = 0, net_w = 0, nc = 0, no = 0, rows = 0;
  std::vector<Buffer> bufs;
  std::vector<Op> ops;
  float* pred = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  int n_launches = 0;
  // preprocess state
  uint8_t* stage_u8 = nullptr;
  size_t stage_bytes = 0;
  int tab_in_h = -1, tab_in_w = -1;
  void* tab_dev = nullptr;
  PreParams pre{};
  // nms scratch
  void* nms_scratch = nullptr;
  size_t nms_scratch_bytes = 0;
  float* nms_dets = nullptr;
  int* nms_counts = nullptr;
  size_t nms_out_cap = 0;
};

namespace {

int fail(aq_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  else g_create_error = buf;
  return code;
}

#define CU_CHECK(c, expr)                                                                              \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess) return fail((c), AQ_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                       __FILE__, __LINE__);                                            \
  } while (0)

int round_up(int v, int m) { return (v + m - 1) / m * m; }

CUtensorMapSwizzle swizzle_for(int inner_bytes) {
  return inner_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                            : inner_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
}

int encode_map(aq_ctx* c, CUtensorMap* m, void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
               const uint32_t* box, CUtensorMapSwizzle swz) {
  cuuint64_t gd[5] = {1, 1, 1, 1, 1}, gs[4] = {0, 0, 0, 0};
  cuuint32_t bx[5] = {1, 1, 1, 1, 1}, es[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i < rank - 1) gs[i] = strides_bytes[i];
  }
  CUresult r = c->encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank, base, gd, gs, bx, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(c, AQ_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu %llu %llu %llu %llu] box [%u %u %u %u %u]",
                (int)r, rank, (unsigned long long)gd[0], (unsigned long long)(rank > 1 ? gd[1] : 0),
                (unsigned long long)(rank > 2 ? gd[2] : 0), (unsigned long long)(rank > 3 ? gd[3] : 0),
                (unsigned long long)(rank > 4 ? gd[4] : 0), bx[0], rank > 1 ? bx[1] : 0, rank > 2 ? bx[2] : 0,
                rank > 3 ? bx[3] : 0, rank > 4 ? bx[4] : 0);
  return AQ_OK;
}

// Pick the spatial M tile (tw x th <= 128 pixels) that wastes the fewest MMA rows on an (h, w) map.
void pick_spatial_tile(int h, int w, int* tw_out, int* th_out) {
  double best = -1;
  int btw = 16, bth = 8;
  for (int tw = 1; tw <= 128 && tw <= 256; ++tw) {
    if (tw > w && tw != 1) break;
    int th = 128 / tw;
    if (th > h) th = h;
    if (th < 1) continue;
    const long tiles = (long)((w + tw - 1) / tw) * ((h + th - 1) / th);
    double eff = (double)h * w / (tiles * 128.0);
    // prefer squarer patches on ties (more halo reuse in L2)
    const double sq = 1e-6 * (tw < th * 2 ? (double)tw / (th * 2) : (double)(th * 2) / tw);
    eff += sq;
    if (eff > best) {
      best = eff;
      btw = tw;
      bth = th;
    }
  }
  *tw_out = btw;
  *th_out = bth;
}

int largest_divisor_tile_n(int c_out) {
  for (int n = 256; n >= 16; n -= 16)
    if (c_out % n == 0) return n;
  return 0;
}

int pow2_at_least(int v) {
  int p = 32;
  while (p < v) p <<= 1;
  return p;
}

// Fill everything of ConvKParams except device pointers / tensor maps.
int plan_conv_geometry(aq_ctx* c, Op& op) {
  const bool det = op.kind == OP_DETECT;
  ConvKParams& k = op.kp;
  memset(&k, 0, sizeof k);
  const int src = det ? op.det.src : op.conv.src;
  const int c_in = det ? op.det.c_in : op.conv.c_in;
  const int ksize = det ? 1 : op.conv.k;
  const int stride = det ? 1 : op.conv.stride;
  const Buffer& sb = c->bufs[src];
  const int ho = sb.h / stride, wo = sb.w / stride;
  k.ksize = ksize;
  k.pad = ksize / 2;
  k.cb = c_in % 64 == 0 ? 64 : (c_in % 32 == 0 ? 32 : 16);
  if (c_in % 16) return fail(c, AQ_ERR_UNSUPPORTED, "conv c_in %d is not a multiple of 16", c_in);
  k.cpt = c_in / k.cb;
  k.total_chunks = ksize * ksize * k.cpt;
  if (det) {
    k.tile_n = round_up(3 * (op.det.nc + 5), 16);
    if (k.tile_n > 256) return fail(c, AQ_ERR_UNSUPPORTED, "detect head with nc=%d needs N=%d > 256", op.det.nc, k.tile_n);
    k.n_tiles_n = 1;
  } else {
    k.tile_n = op.conv.tile_n > 0 ? op.conv.tile_n : largest_divisor_tile_n(op.conv.c_out);
    if (k.tile_n <= 0 || k.tile_n > 256 || k.tile_n % 16 || op.conv.c_out % k.tile_n)
      return fail(c, AQ_ERR_UNSUPPORTED, "conv c_out %d has no valid UMMA N tile (got %d)", op.conv.c_out, k.tile_n);
    k.n_tiles_n = op.conv.c_out / k.tile_n;
  }
  k.ob = k.tile_n % 64 == 0 ? 64 : (k.tile_n % 32 == 0 ? 32 : 16);
  int rows_per_tile;
  if (ksize == 1 && stride == 1) {
    k.mode = CONV_MODE_FLAT;
    k.m_tiles = (c->batch * ho * wo + kTileM - 1) / kTileM;
    k.tw = kTileM;
    k.th = 1;
    k.tiles_x = k.tiles_y = 1;
    rows_per_tile = kTileM;
  } else {
    if (ksize != 3 || (stride != 1 && stride != 2))
      return fail(c, AQ_ERR_UNSUPPORTED, "conv k=%d s=%d unsupported", ksize, stride);
    if (stride == 2 && ((sb.h | sb.w) & 1)) return fail(c, AQ_ERR_UNSUPPORTED, "stride-2 conv on odd map");
    k.mode = stride == 1 ? CONV_MODE_S1 : CONV_MODE_S2;
    pick_spatial_tile(ho, wo, &k.tw, &k.th);
    k.tiles_x = (wo + k.tw - 1) / k.tw;
    k.tiles_y = (ho + k.th - 1) / k.th;
    k.m_tiles = c->batch * k.tiles_x * k.tiles_y;
    rows_per_tile = k.tw * k.th;
  }
  k.src_ct = sb.c;
  k.src_c_off = det ? op.det.src_c_off : op.conv.src_c_off;
  k.act = det ? 0 : op.conv.act;
  k.has_res = (!det && op.conv.res >= 0) ? 1 : 0;
  k.a_sub_bytes = kTileM * k.cb * 2;
  k.b_sub_bytes = round_up(k.tile_n * k.cb * 2, 1024);
  k.a_box_bytes = rows_per_tile * k.cb * 2;
  k.b_box_bytes = k.tile_n * k.cb * 2;
  k.cps = (!det && op.conv.chunks_per_stage > 0) ? op.conv.chunks_per_stage : (64 / k.cb);
  if (k.cps > k.total_chunks) k.cps = k.total_chunks;
  k.kstages = (k.total_chunks + k.cps - 1) / k.cps;
  k.stage_bytes = k.cps * (k.a_sub_bytes + k.b_sub_bytes);
  k.c_bytes = det ? 0 : kTileM * k.tile_n * 2;
  const int budget = kConvMaxDynSmem - (int)k.c_bytes - 1024;
  int st = budget / (int)k.stage_bytes;
  if (st > kMaxStages) st = kMaxStages;
  if (st < 2) return fail(c, AQ_ERR_UNSUPPORTED, "conv stage of %u bytes does not fit twice in shared memory", k.stage_bytes);
  k.num_stages = st;
  k.tmem_cols = pow2_at_least(2 * k.tile_n);
  k.acc_stride = k.tmem_cols / 2;
  if (k.tmem_cols > 512) return fail(c, AQ_ERR_UNSUPPORTED, "tile_n %d needs more than 512 TMEM columns", k.tile_n);
  return AQ_OK;
}

// Tensor maps for one planned conv (needs device pointers).
int plan_conv_maps(aq_ctx* c, Op& op) {
  const bool det = op.kind == OP_DETECT;
  ConvKParams& k = op.kp;
  const int src = det ? op.det.src : op.conv.src;
  const int c_in = det ? op.det.c_in : op.conv.c_in;
  const Buffer& sb = c->bufs[src];
  const uint64_t N = c->batch;
  int rc;
  // ---- A
  if (k.mode == CONV_MODE_FLAT) {
    uint64_t dims[2] = {(uint64_t)c_in, N * sb.h * sb.w};
    uint64_t str[1] = {(uint64_t)sb.c * 2};
    uint32_t box[2] = {(uint32_t)k.cb, (uint32_t)kTileM};
    rc = encode_map(c, &k.tmA, sb.ptr + k.src_c_off, 2, dims, str, box, swizzle_for(k.cb * 2));
  } else if (k.mode == CONV_MODE_S1) {
    uint64_t dims[4] = {(uint64_t)c_in, (uint64_t)sb.w, (uint64_t)sb.h, N};
    uint64_t str[3] = {(uint64_t)sb.c * 2, (uint64_t)sb.w * sb.c * 2, (uint64_t)sb.h * sb.w * sb.c * 2};
    uint32_t box[4] = {(uint32_t)k.cb, (uint32_t)k.tw, (uint32_t)k.th, 1};
    rc = encode_map(c, &k.tmA, sb.ptr + k.src_c_off, 4, dims, str, box, swizzle_for(k.cb * 2));
  } else {
    // (x-parity*Ct + channel, W/2, y-parity, H/2, N): a stride-2 gather becomes a dense box
    uint64_t dims[5] = {(uint64_t)2 * sb.c, (uint64_t)sb.w / 2, 2, (uint64_t)sb.h / 2, N};
    uint64_t str[4] = {(uint64_t)2 * sb.c * 2, (uint64_t)sb.w * sb.c * 2, (uint64_t)2 * sb.w * sb.c * 2,
                       (uint64_t)sb.h * sb.w * sb.c * 2};
    uint32_t box[5] = {(uint32_t)k.cb, (uint32_t)k.tw, 1, (uint32_t)k.th, 1};
    rc = encode_map(c, &k.tmA, sb.ptr, 5, dims, str, box, swizzle_for(k.cb * 2));
  }
  if (rc) return rc;
  // ---- B: weights [rows = c_out][K = k*k*c_in]
  {
    const int rows = det ? 3 * (op.det.nc + 5) : op.conv.c_out;
    const uint64_t K = (uint64_t)k.ksize * k.ksize * c_in;
    uint64_t dims[2] = {K, (uint64_t)rows};
    uint64_t str[1] = {K * 2};
    uint32_t box[2] = {(uint32_t)k.cb, (uint32_t)k.tile_n};
    rc = encode_map(c, &k.tmB, op.w_dev, 2, dims, str, box, swizzle_for(k.cb * 2));
    if (rc) return rc;
  }
  k.bias = op.b_dev;
  if (det) {
    k.pred = c->pred;
    k.det_no = op.det.nc + 5;
    k.det_rows = c->rows;
    k.det_level_off = op.det.level_row_off;
    k.det_ny = sb.h;
    k.det_nx = sb.w;
    k.det_m_total = c->batch * sb.h * sb.w;
    k.det_stride = op.det.stride;
    for (int i = 0; i < 6; ++i) k.det_anchor[i] = op.det.anchors_px[i];
    return AQ_OK;
  }
  // ---- output (and residual) maps over the destination channel slice
  auto out_map = [&](CUtensorMap* m, int buf, int c_off) -> int {
    const Buffer& db = c->bufs[buf];
    if (k.mode == CONV_MODE_FLAT) {
      uint64_t dims[2] = {(uint64_t)op.conv.c_out, N * db.h * db.w};
      uint64_t str[1] = {(uint64_t)db.c * 2};
      uint32_t box[2] = {(uint32_t)k.ob, (uint32_t)kTileM};
      return encode_map(c, m, db.ptr + c_off, 2, dims, str, box, swizzle_for(k.ob * 2));
    }
    uint64_t dims[4] = {(uint64_t)op.conv.c_out, (uint64_t)db.w, (uint64_t)db.h, N};
    uint64_t str[3] = {(uint64_t)db.c * 2, (uint64_t)db.w * db.c * 2, (uint64_t)db.h * db.w * db.c * 2};
    uint32_t box[4] = {(uint32_t)k.ob, (uint32_t)k.tw, (uint32_t)k.th, 1};
    return encode_map(c, m, db.ptr + c_off, 4, dims, str, box, swizzle_for(k.ob * 2));
  };
  rc = out_map(&k.tmOut, op.conv.dst, op.conv.dst_c_off);
  if (rc) return rc;
  if (k.has_res) rc = out_map(&k.tmRes, op.conv.res, op.conv.res_c_off);
  return rc;
}

int check_buf(aq_ctx* c, int id, int c_off, int ch, const char* what) {
  if (id < 0 || id >= (int)c->bufs.size()) return fail(c, AQ_ERR_ARG, "%s: buffer id %d out of range", what, id);
  const Buffer& b = c->bufs[id];
  if (c_off < 0 || ch <= 0 || c_off + ch > b.c) return fail(c, AQ_ERR_ARG, "%s: channels [%d,%d) outside buffer %d (c=%d)", what, c_off, c_off + ch, id, b.c);
  if (c_off % 8) return fail(c, AQ_ERR_ARG, "%s: channel offset %d is not 16-byte aligned", what, c_off);
  return AQ_OK;
}

void free_plan(aq_ctx* c) {
  if (c->graph_exec) cudaGraphExecDestroy(c->graph_exec);
  if (c->graph) cudaGraphDestroy(c->graph);
  c->graph_exec = nullptr;
  c->graph = nullptr;
  for (auto& b : c->bufs)
    if (b.ptr) cudaFree(b.ptr);
  for (auto& o : c->ops) {
    if (o.w_dev) cudaFree(o.w_dev);
    if (o.b_dev) cudaFree(o.b_dev);
  }
  if (c->pred) cudaFree(c->pred);
  c->pred = nullptr;
  c->bufs.clear();
  c->ops.clear();
  c->planned = c->planning = false;
}

int run_ops(aq_ctx* c, cudaStream_t s) {
  for (auto& op : c->ops) {
    cudaError_t e = cudaSuccess;
    switch (op.kind) {
      case OP_CONV: e = launch_conv(op.kp, EPI_STORE, c->num_sms, s); break;
      case OP_DETECT: e = launch_conv(op.kp, EPI_DETECT, c->num_sms, s); break;
      case OP_POOL: {
        const Buffer& b = c->bufs[op.buf];
        e = launch_sppf_pool(b.ptr, c->batch, b.h, b.w, b.c, op.c_off, op.c, s);
        break;
      }
      case OP_UPSAMPLE: {
        const Buffer& sb = c->bufs[op.buf];
        const Buffer& db = c->bufs[op.dst];
        e = launch_upsample2x(sb.ptr, c->batch, sb.h, sb.w, sb.c, op.c_off, op.c, db.ptr, db.c, op.dst_c_off, s);
        break;
      }
    }
    if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "launch of op kind %d failed: %s", (int)op.kind, cudaGetErrorString(e));
  }
  return AQ_OK;
}

// cv2 INTER_LINEAR uint8 coefficient tables (11-bit fixed point), computed exactly as cv2 does on the host.
// Horizontally cv2 zeroes the fraction where the 2-tap window leaves the image; vertically it keeps the weights
// and clips the row indices instead.
void resize_tables(int dst, int src, bool vertical, std::vector<int>& i0, std::vector<int>& i1,
                   std::vector<short>& a0, std::vector<short>& a1) {
  const double inv_scale = (double)dst / src;
  const double scale = 1.0 / inv_scale;
  i0.resize(dst); i1.resize(dst); a0.resize(dst); a1.resize(dst);
  for (int d = 0; d < dst; ++d) {
    float fx = (float)((d + 0.5) * scale - 0.5);
    int sx = (int)floorf(fx);
    fx -= sx;
    if (!vertical) {
      if (sx < 0) { fx = 0; sx = 0; }
      if (sx >= src - 1) { fx = 0; sx = src - 1; }
    }
    auto clip = [src](int v) { return v < 0 ? 0 : (v > src - 1 ? src - 1 : v); };
    i0[d] = clip(sx);
    i1[d] = clip(sx + 1);
    a0[d] = (short)lrintf((1.f - fx) * 2048.f);
    a1[d] = (short)lrintf(fx * 2048.f);
  }
}

int setup_preprocess(aq_ctx* c, int in_h, int in_w) {
  if (c->tab_in_h == in_h && c->tab_in_w == in_w) return AQ_OK;
  // upstream letterbox(im, new_shape=(net_h, net_w), auto=False here: the plan fixes the padded shape)
  PreParams& p = c->pre;
  memset(&p, 0, sizeof p);
  const double r = fmin((double)c->net_h / in_h, (double)c->net_w / in_w);
  p.new_w = (int)nearbyint(in_w * r);
  p.new_h = (int)nearbyint(in_h * r);
  const double dw = (c->net_w - p.new_w) / 2.0, dh = (c->net_h - p.new_h) / 2.0;
  if (dw < 0 || dh < 0) return fail(c, AQ_ERR_ARG, "letterbox: resized %dx%d exceeds network input %dx%d", p.new_w, p.new_h, c->net_w, c->net_h);
  p.top = (int)nearbyint(dh - 0.1);
  p.left = (int)nearbyint(dw - 0.1);
  p.resize = (p.new_w != in_w || p.new_h != in_h) ? 1 : 0;
  p.in_h = in_h; p.in_w = in_w; p.net_h = c->net_h; p.net_w = c->net_w;
  if (c->tab_dev) { cudaFree(c->tab_dev); c->tab_dev = nullptr; }
  if (p.resize) {
    std::vector<int> x0, x1, y0, y1;
    std::vector<short> ax0, ax1, ay0, ay1;
    resize_tables(p.new_w, in_w, false, x0, x1, ax0, ax1);
    resize_tables(p.new_h, in_h, true, y0, y1, ay0, ay1);
    std::vector<uint8_t> host;
    auto put = [&](const void* src, size_t n) {
      const size_t o = host.size();
      host.resize(o + ((n + 15) & ~size_t(15)));
      memcpy(host.data() + o, src, n);
      return o;
    };
    const size_t nx = p.new_w, ny = p.new_h;
    const size_t ox0 = put(x0.data(), nx * 4), ox1 = put(x1.data(), nx * 4);
    const size_t oy0 = put(y0.data(), ny * 4), oy1 = put(y1.data(), ny * 4);
    const size_t oax0 = put(ax0.data(), nx * 2), oax1 = put(ax1.data(), nx * 2);
    const size_t oay0 = put(ay0.data(), ny * 2), oay1 = put(ay1.data(), ny * 2);
    CU_CHECK(c, cudaMalloc(&c->tab_dev, host.size()));
    CU_CHECK(c, cudaMemcpy(c->tab_dev, host.data(), host.size(), cudaMemcpyHostToDevice));
    const uint8_t* d = (const uint8_t*)c->tab_dev;
    p.x0 = (const int*)(d + ox0); p.x1 = (const int*)(d + ox1);
    p.y0 = (const int*)(d + oy0); p.y1 = (const int*)(d + oy1);
    p.ax0 = (const short*)(d + oax0); p.ax1 = (const short*)(d + oax1);
    p.ay0 = (const short*)(d + oay0); p.ay1 = (const short*)(d + oay1);
  }
  c->tab_in_h = in_h;
  c->tab_in_w = in_w;
  return AQ_OK;
}

int stage_tiles(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int src_is_device, cudaStream_t s,
                const uint8_t** dev_out) {
  if (src_is_device) {
    *dev_out = tiles;
    return AQ_OK;
  }
  const size_t bytes = (size_t)n * in_h * in_w * 3;
  if (bytes > c->stage_bytes) {
    if (c->stage_u8) cudaFree(c->stage_u8);
    c->stage_u8 = nullptr;
    CU_CHECK(c, cudaMalloc(&c->stage_u8, bytes));
    c->stage_bytes = bytes;
  }
  CU_CHECK(c, cudaMemcpyAsync(c->stage_u8, tiles, bytes, cudaMemcpyHostToDevice, s));
  *dev_out = c->stage_u8;
  return AQ_OK;
}

}  // namespace

// =====================================================================================================
extern "C" {

int aq_abi_version(void) { return AQ_ABI_VERSION; }

const char* aq_last_error(const aq_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int aq_create(int device, aq_ctx** out) {
  if (!out) return fail(nullptr, AQ_ERR_ARG, "aq_create: out is NULL");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(nullptr, AQ_ERR_CUDA, "no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
  if (device < 0 || device >= count) return fail(nullptr, AQ_ERR_ARG, "device %d out of range (%d devices)", device, count);
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail(nullptr, AQ_ERR_CUDA, "%s", cudaGetErrorString(e));
  if (prop.major != 10)
    return fail(nullptr, AQ_ERR_UNSUPPORTED, "device %d is sm_%d%d; kernels are built for sm_100a only", device, prop.major, prop.minor);
  if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, AQ_ERR_CUDA, "%s", cudaGetErrorString(e));
  aq_ctx* c = new aq_ctx();
  c->device = device;
  c->num_sms = prop.multiProcessorCount;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || !fn) {
    delete c;
    return fail(nullptr, AQ_ERR_CUDA, "cuTensorMapEncodeTiled unavailable: %s", cudaGetErrorString(e));
  }
  c->encode = (EncodeTiledFn)fn;
  const char* what = "stream";
  if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (what = "conv smem attr", e = conv_kernels_init()) != cudaSuccess ||
      (what = "nms smem attr", e = nms_init()) != cudaSuccess ||
      (what = "sppf smem attr", e = sppf_init()) != cudaSuccess) {
    delete c;
    return fail(nullptr, AQ_ERR_CUDA, "init (%s): %s", what, cudaGetErrorString(e));
  }
  *out = c;
  return AQ_OK;
}

void aq_destroy(aq_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  free_plan(c);
  if (c->stage_u8) cudaFree(c->stage_u8);
  if (c->tab_dev) cudaFree(c->tab_dev);
  if (c->nms_scratch) cudaFree(c->nms_scratch);
  if (c->nms_dets) cudaFree(c->nms_dets);
  if (c->nms_counts) cudaFree(c->nms_counts);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

int aq_plan_begin(aq_ctx* c, int batch, int net_h, int net_w, int nc) {
  if (!c) return AQ_ERR_ARG;
  if (batch <= 0 || net_h <= 0 || net_w <= 0 || net_h % 32 || net_w % 32 || nc <= 0)
    return fail(c, AQ_ERR_ARG, "plan_begin: batch %d, input %dx%d (multiples of 32), nc %d", batch, net_h, net_w, nc);
  cudaSetDevice(c->device);
  free_plan(c);
  c->batch = batch; c->net_h = net_h; c->net_w = net_w; c->nc = nc; c->no = nc + 5;
  c->rows = 3 * ((net_h / 8) * (net_w / 8) + (net_h / 16) * (net_w / 16) + (net_h / 32) * (net_w / 32));
  c->planning = true;
  c->tab_in_h = c->tab_in_w = -1;
  return AQ_OK;
}

int aq_plan_buffer(aq_ctx* c, int h, int w, int ch) {
  if (!c || !c->planning) return c ? fail(c, AQ_ERR_STATE, "plan_buffer outside plan_begin/plan_end") : AQ_ERR_ARG;
  if (h <= 0 || w <= 0 || ch <= 0 || ch % 8) return fail(c, AQ_ERR_ARG, "plan_buffer: %dx%dx%d (c must be a multiple of 8)", h, w, ch);
  Buffer b;
  b.h = h; b.w = w; b.c = ch;
  c->bufs.push_back(b);
  return (int)c->bufs.size() - 1;
}

int aq_plan_conv(aq_ctx* c, const aq_conv_desc* d) {
  if (!c || !d) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_conv outside plan_begin/plan_end");
  int rc;
  if ((rc = check_buf(c, d->src, d->src_c_off, d->c_in, "conv src"))) return rc;
  if ((rc = check_buf(c, d->dst, d->dst_c_off, d->c_out, "conv dst"))) return rc;
  if (d->res >= 0 && (rc = check_buf(c, d->res, d->res_c_off, d->c_out, "conv residual"))) return rc;
  if (!d->weight || !d->bias) return fail(c, AQ_ERR_ARG, "conv: weight/bias NULL");
  const Buffer& sb = c->bufs[d->src];
  const Buffer& db = c->bufs[d->dst];
  if (d->stride != 1 && d->stride != 2) return fail(c, AQ_ERR_UNSUPPORTED, "conv stride %d", d->stride);
  if (sb.h / d->stride != db.h || sb.w / d->stride != db.w)
    return fail(c, AQ_ERR_ARG, "conv: src %dx%d / stride %d != dst %dx%d", sb.h, sb.w, d->stride, db.h, db.w);
  if (d->res >= 0 && (c->bufs[d->res].h != db.h || c->bufs[d->res].w != db.w)) return fail(c, AQ_ERR_ARG, "conv: residual geometry mismatch");
  Op op;
  op.kind = OP_CONV;
  op.conv = *d;
  const size_t nw = (size_t)d->c_out * d->k * d->k * d->c_in;
  op.w_host.assign(d->weight, d->weight + nw);
  op.b_host.assign(d->bias, d->bias + d->c_out);
  op.conv.weight = nullptr;
  op.conv.bias = nullptr;
  if ((rc = plan_conv_geometry(c, op))) return rc;
  c->ops.push_back(std::move(op));
  return AQ_OK;
}

int aq_plan_detect(aq_ctx* c, const aq_detect_desc* d) {
  if (!c || !d) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_detect outside plan_begin/plan_end");
  int rc;
  if ((rc = check_buf(c, d->src, d->src_c_off, d->c_in, "detect src"))) return rc;
  if (d->nc != c->nc) return fail(c, AQ_ERR_ARG, "detect nc %d != plan nc %d", d->nc, c->nc);
  const Buffer& sb = c->bufs[d->src];
  if (d->level_row_off < 0 || d->level_row_off + 3 * sb.h * sb.w > c->rows) return fail(c, AQ_ERR_ARG, "detect level rows out of range");
  Op op;
  op.kind = OP_DETECT;
  op.det = *d;
  const int n_out = 3 * (d->nc + 5);
  op.w_host.assign(d->weight, d->weight + (size_t)n_out * d->c_in);
  op.b_host.assign(d->bias, d->bias + n_out);
  op.det.weight = nullptr;
  op.det.bias = nullptr;
  if ((rc = plan_conv_geometry(c, op))) return rc;
  c->ops.push_back(std::move(op));
  return AQ_OK;
}

int aq_plan_sppf_pool(aq_ctx* c, int buf, int c_off, int ch) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_sppf_pool outside plan_begin/plan_end");
  int rc;
  if ((rc = check_buf(c, buf, c_off, 4 * ch, "sppf pool"))) return rc;
  Op op;
  op.kind = OP_POOL;
  op.buf = buf; op.c_off = c_off; op.c = ch;
  c->ops.push_back(std::move(op));
  return AQ_OK;
}

int aq_plan_upsample2x(aq_ctx* c, int src, int src_c_off, int ch, int dst, int dst_c_off) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_upsample2x outside plan_begin/plan_end");
  int rc;
  if ((rc = check_buf(c, src, src_c_off, ch, "upsample src"))) return rc;
  if ((rc = check_buf(c, dst, dst_c_off, ch, "upsample dst"))) return rc;
  if (c->bufs[dst].h != 2 * c->bufs[src].h || c->bufs[dst].w != 2 * c->bufs[src].w) return fail(c, AQ_ERR_ARG, "upsample: dst is not 2x src");
  Op op;
  op.kind = OP_UPSAMPLE;
  op.buf = src; op.c_off = src_c_off; op.c = ch; op.dst = dst; op.dst_c_off = dst_c_off;
  c->ops.push_back(std::move(op));
  return AQ_OK;
}

int aq_plan_end(aq_ctx* c) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_end without plan_begin");
  cudaSetDevice(c->device);
  for (auto& b : c->bufs) {
    const size_t bytes = (size_t)c->batch * b.h * b.w * b.c * sizeof(__half);
    CU_CHECK(c, cudaMalloc(&b.ptr, bytes));
    CU_CHECK(c, cudaMemsetAsync(b.ptr, 0, bytes, c->stream));
  }
  {
    const size_t bytes = (size_t)c->batch * c->rows * c->no * sizeof(float);
    CU_CHECK(c, cudaMalloc(&c->pred, bytes));
    CU_CHECK(c, cudaMemsetAsync(c->pred, 0, bytes, c->stream));
  }
  int launches = 0;
  for (auto& op : c->ops) {
    ++launches;
    if (op.kind != OP_CONV && op.kind != OP_DETECT) continue;
    const ConvKParams& k = op.kp;
    // weights: rows padded up to n_tiles_n * tile_n are handled by TMA OOB fill; bias is padded here
    CU_CHECK(c, cudaMalloc(&op.w_dev, op.w_host.size() * sizeof(uint16_t)));
    CU_CHECK(c, cudaMemcpyAsync(op.w_dev, op.w_host.data(), op.w_host.size() * sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream));
    std::vector<float> bias((size_t)k.n_tiles_n * k.tile_n, 0.f);
    memcpy(bias.data(), op.b_host.data(), op.b_host.size() * sizeof(float));
    CU_CHECK(c, cudaMalloc(&op.b_dev, bias.size() * sizeof(float)));
    CU_CHECK(c, cudaMemcpy(op.b_dev, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
    CU_CHECK(c, cudaStreamSynchronize(c->stream));
    int rc = plan_conv_maps(c, op);
    if (rc) return rc;
    std::vector<uint16_t>().swap(op.w_host);
  }
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  c->n_launches = launches;
  // capture the whole forward as one graph
  if (!c->ops.empty()) {
    CU_CHECK(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    int rc = run_ops(c, c->stream);
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture(c->stream, &g);
    if (rc) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "graph capture: %s", cudaGetErrorString(e));
    c->graph = g;
    CU_CHECK(c, cudaGraphInstantiate(&c->graph_exec, c->graph, 0));
  }
  c->planning = false;
  c->planned = true;
  return AQ_OK;
}

int aq_buffer_info(aq_ctx* c, int buf, void** dev_ptr, int* n, int* h, int* w, int* ch) {
  if (!c) return AQ_ERR_ARG;
  if (buf < 0 || buf >= (int)c->bufs.size()) return fail(c, AQ_ERR_ARG, "buffer id %d out of range", buf);
  const Buffer& b = c->bufs[buf];
  if (dev_ptr) *dev_ptr = b.ptr;
  if (n) *n = c->batch;
  if (h) *h = b.h;
  if (w) *w = b.w;
  if (ch) *ch = b.c;
  return AQ_OK;
}

int aq_pred_info(aq_ctx* c, float** dev_ptr, int* rows, int* no) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "no plan");
  if (dev_ptr) *dev_ptr = c->pred;
  if (rows) *rows = c->rows;
  if (no) *no = c->no;
  return AQ_OK;
}

int aq_plan_num_launches(aq_ctx* c) { return c ? c->n_launches : AQ_ERR_ARG; }

int aq_profile_ops(aq_ctx* c, int iters, float* ms_per_op, int cap) {
  if (!c || !ms_per_op || iters <= 0) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "profile before plan_end");
  cudaSetDevice(c->device);
  const int n = (int)c->ops.size();
  if (cap < n) return fail(c, AQ_ERR_ARG, "profile: need room for %d ops", n);
  cudaEvent_t e0, e1;
  CU_CHECK(c, cudaEventCreate(&e0));
  CU_CHECK(c, cudaEventCreate(&e1));
  std::vector<Op> all;
  all.swap(c->ops);
  int rc = AQ_OK;
  for (int i = 0; i < n && rc == AQ_OK; ++i) {
    c->ops.clear();
    c->ops.push_back(all[i]);
    rc = run_ops(c, c->stream);  // warm-up
    if (rc) break;
    cudaEventRecord(e0, c->stream);
    for (int k = 0; k < iters && rc == AQ_OK; ++k) rc = run_ops(c, c->stream);
    cudaEventRecord(e1, c->stream);
    if (cudaEventSynchronize(e1) != cudaSuccess) rc = fail(c, AQ_ERR_CUDA, "profile: op %d failed: %s", i, cudaGetErrorString(cudaGetLastError()));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    ms_per_op[i] = ms / iters;
  }
  c->ops.swap(all);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return rc ? rc : n;
}

int aq_op_describe(aq_ctx* c, int i, char* buf, size_t cap) {
  if (!c || !buf || i < 0 || i >= (int)c->ops.size()) return AQ_ERR_ARG;
  const Op& op = c->ops[i];
  if (op.kind == OP_CONV || op.kind == OP_DETECT) {
    const bool det = op.kind == OP_DETECT;
    const ConvKParams& k = op.kp;
    const Buffer& sb = c->bufs[det ? op.det.src : op.conv.src];
    const int stride = det ? 1 : op.conv.stride;
    const int c_in = det ? op.det.c_in : op.conv.c_in;
    const int c_out = det ? 3 * (op.det.nc + 5) : op.conv.c_out;
    const double M = (double)c->batch * (sb.h / stride) * (sb.w / stride);
    const double flops = 2.0 * M * c_out * c_in * k.ksize * k.ksize;
    const double bytes = 2.0 * ((double)c->batch * sb.h * sb.w * c_in + M * c_out * (k.has_res ? 2 : 1)) +
                         2.0 * c_out * c_in * k.ksize * k.ksize;
    snprintf(buf, cap, "%s k%d s%d %dx%d cin=%d cout=%d res=%d | M=%.0f tileN=%d cb=%d cps=%d stages=%d tile=%dx%d mt=%d nt=%d | flop=%.4g bytes=%.4g",
             det ? "detect" : "conv", k.ksize, stride, sb.h, sb.w, c_in, c_out, k.has_res, M, k.tile_n, k.cb, k.cps,
             k.num_stages, k.tw, k.th, k.m_tiles, k.n_tiles_n, flops, bytes);
  } else if (op.kind == OP_POOL) {
    const Buffer& b = c->bufs[op.buf];
    snprintf(buf, cap, "sppf_pool %dx%d c=%d | flop=0 bytes=%.4g", b.h, b.w, op.c, 2.0 * c->batch * b.h * b.w * op.c * 4.0);
  } else {
    const Buffer& b = c->bufs[op.buf];
    snprintf(buf, cap, "upsample2x %dx%d c=%d | flop=0 bytes=%.4g", b.h, b.w, op.c, 2.0 * c->batch * b.h * b.w * op.c * 5.0);
  }
  return AQ_OK;
}

int aq_preprocess(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device, void* stream) {
  if (!c || !tiles) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "preprocess before plan_end");
  if (n <= 0 || n > c->batch) return fail(c, AQ_ERR_ARG, "preprocess: n %d outside plan batch %d", n, c->batch);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  cudaSetDevice(c->device);
  const Buffer& b0 = c->bufs[0];
  if (b0.h != c->net_h / 2 || b0.w != c->net_w / 2 || b0.c != 16)
    return fail(c, AQ_ERR_STATE, "buffer 0 must be the (net_h/2, net_w/2, 16) space-to-depth input");
  int rc = setup_preprocess(c, in_h, in_w);
  if (rc) return rc;
  const uint8_t* dev = nullptr;
  if ((rc = stage_tiles(c, tiles, n, in_h, in_w, src_is_device, s, &dev))) return rc;
  PreParams p = c->pre;
  p.src = dev; p.n = n; p.bgr = bgr; p.out_s2d = b0.ptr; p.out_u8 = nullptr;
  CU_CHECK(c, launch_preprocess(p, s));
  return AQ_OK;
}

int aq_preprocess_u8(aq_ctx* c, const uint8_t* tiles_dev, int n, int in_h, int in_w, int bgr, uint8_t* out_dev, void* stream) {
  if (!c || !tiles_dev || !out_dev) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "preprocess before plan_end");
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  cudaSetDevice(c->device);
  int rc = setup_preprocess(c, in_h, in_w);
  if (rc) return rc;
  PreParams p = c->pre;
  p.src = tiles_dev; p.n = n; p.bgr = bgr; p.out_s2d = nullptr; p.out_u8 = out_dev;
  CU_CHECK(c, launch_preprocess_u8(p, s));
  return AQ_OK;
}

int aq_forward(aq_ctx* c, void* stream) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "forward before plan_end");
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  cudaSetDevice(c->device);
  if (c->graph_exec) CU_CHECK(c, cudaGraphLaunch(c->graph_exec, s));
  return AQ_OK;
}

int aq_nms(aq_ctx* c, const float* pred, int batch, int rows, int no, float conf_thres, double iou_thres, int max_det,
           int agnostic, const int* classes, int n_classes, float max_wh, int max_nms, float* dets_out,
           int* counts_out, int* keep_rows_out, void* stream) {
  if (!c) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  if (!pred) {
    if (!c->planned) return fail(c, AQ_ERR_STATE, "nms: no plan and no predictions given");
    pred = c->pred;
    batch = batch > 0 ? batch : c->batch;
    rows = c->rows;
    no = c->no;
  }
  if (batch <= 0 || rows <= 0 || no < 6 || max_det <= 0 || max_det > 8192 || !dets_out || !counts_out)
    return fail(c, AQ_ERR_ARG, "nms: batch %d rows %d no %d max_det %d", batch, rows, no, max_det);
  if (n_classes < 0 || n_classes > 32 || (n_classes > 0 && !classes)) return fail(c, AQ_ERR_ARG, "nms: at most 32 filter classes");
  if (!(conf_thres >= 0.f && conf_thres <= 1.f) || !(iou_thres >= 0.0 && iou_thres <= 1.0))
    return fail(c, AQ_ERR_ARG, "nms: thresholds must lie in [0,1]");
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  const size_t need = nms_scratch_bytes(batch, rows);
  if (need > c->nms_scratch_bytes) {
    if (c->nms_scratch) cudaFree(c->nms_scratch);
    c->nms_scratch = nullptr;
    CU_CHECK(c, cudaMalloc(&c->nms_scratch, need));
    c->nms_scratch_bytes = need;
  }
  NmsParams p{};
  p.pred = pred; p.rows = rows; p.no = no; p.conf_thres = conf_thres;
  // largest fp32 <= iou_thres: (double)ovr > iou_thres  <=>  ovr > thr_f
  float tf = (float)iou_thres;
  if ((double)tf > iou_thres) tf = nextafterf(tf, -INFINITY);
  p.iou_thr_f32 = tf;
  p.max_det = max_det; p.max_nms = max_nms > 0 ? max_nms : 30000; p.agnostic = agnostic; p.max_wh = max_wh;
  p.n_classes = n_classes;
  for (int i = 0; i < n_classes; ++i) p.classes[i] = classes[i];
  uint8_t* sc = (uint8_t*)c->nms_scratch;
  const size_t br = (size_t)batch * rows;
  p.cand_box = (float4*)sc; sc += br * sizeof(float4);
  p.keys = (uint32_t*)sc; sc += br * 2 * 4;
  p.vals = (uint32_t*)sc; sc += br * 2 * 4;
  p.cand_score = (float*)sc; sc += br * 4;
  p.cand_cls = (float*)sc; sc += br * 4;
  p.cand_row = (int*)sc;
  p.dets = dets_out; p.counts = counts_out; p.keep_rows = keep_rows_out;
  CU_CHECK(c, launch_nms(p, batch, s));
  return AQ_OK;
}

int aq_detect_submit(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device,
                     float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes, int n_classes,
                     float* dets_host, int* counts_host, void* stream) {
  if (!c || !dets_host || !counts_host) return AQ_ERR_ARG;
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  int rc = aq_preprocess(c, tiles, n, in_h, in_w, bgr, src_is_device, s);
  if (rc) return rc;
  if ((rc = aq_forward(c, s))) return rc;
  const size_t cap = (size_t)c->batch * max_det;
  if (cap > c->nms_out_cap) {
    if (c->nms_dets) cudaFree(c->nms_dets);
    if (c->nms_counts) cudaFree(c->nms_counts);
    c->nms_dets = nullptr; c->nms_counts = nullptr;
    CU_CHECK(c, cudaMalloc(&c->nms_dets, cap * 6 * sizeof(float)));
    CU_CHECK(c, cudaMalloc(&c->nms_counts, (size_t)c->batch * sizeof(int)));
    c->nms_out_cap = cap;
  }
  if ((rc = aq_nms(c, nullptr, n, 0, 0, conf_thres, iou_thres, max_det, agnostic, classes, n_classes, 7680.f, 30000,
                   c->nms_dets, c->nms_counts, nullptr, s)))
    return rc;
  CU_CHECK(c, cudaMemcpyAsync(counts_host, c->nms_counts, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s));
  CU_CHECK(c, cudaMemcpyAsync(dets_host, c->nms_dets, (size_t)n * max_det * 6 * sizeof(float), cudaMemcpyDeviceToHost, s));
  return AQ_OK;
}

int aq_detect(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device, float conf_thres,
              double iou_thres, int max_det, int agnostic, const int* classes, int n_classes, float* dets_host,
              int* counts_host, void* stream) {
  int rc = aq_detect_submit(c, tiles, n, in_h, in_w, bgr, src_is_device, conf_thres, iou_thres, max_det, agnostic,
                            classes, n_classes, dets_host, counts_host, stream);
  if (rc) return rc;
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  CU_CHECK(c, cudaStreamSynchronize(s));
  return AQ_OK;
}

int aq_memcpy(aq_ctx* c, void* dst, const void* src, size_t bytes, int kind, void* stream) {
  if (!c || !dst || !src) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  const cudaMemcpyKind k = kind == 0 ? cudaMemcpyHostToDevice : kind == 1 ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
  CU_CHECK(c, cudaMemcpyAsync(dst, src, bytes, k, s));
  return AQ_OK;
}

int aq_sync(aq_ctx* c, void* stream) {
  if (!c) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  CU_CHECK(c, cudaStreamSynchronize(stream ? (cudaStream_t)stream : c->stream));
  return AQ_OK;
}

}  // extern "C"
