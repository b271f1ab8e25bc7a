// C ABI of the tile-detection path (include/aqua_b200.h): context, plan builder, TMA tensor-map encoding,
// CUDA-graph capture of the conv stack, and the batch entry points.
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/aqua_b200.h"
#include "aux_kernels.cuh"
#include "conv_igemm.cuh"
#include "jpeg.cuh"
#include "jpeg_huff.cuh"
#include "nms.cuh"

using namespace aq;

namespace {

thread_local std::string g_create_error;

struct Buffer {
  int h, w, c;
  __half* ptr = nullptr;
  int stage = -1;  // >= 0: transient buffer of that stage (holds sub * lanes images and is reused by every sub-batch)
};

// A stage = a run of consecutive ops executed depth-first over sub-batches of `sub` images (aq_plan_stage_begin):
// for each sub-batch all ops of the stage run before the next sub-batch starts, so the activations between them
// (the stage's transient buffers, sized for one sub-batch) are produced and consumed while they still sit in the
// 126 MB L2 and are overwritten there by the next sub-batch -- they never travel to HBM.  `lanes` > 1 runs that many
// sub-batches as independent chains on parallel branches of the graph (one set of transient buffers each), so one
// chain's launch / drain bubbles are filled by the other's CTAs.
struct Stage {
  int sub = 0, lanes = 1;
  int op_begin = 0, op_end = 0;
};
constexpr int kMaxLanes = 4;

enum OpKind { OP_CONV, OP_DETECT, OP_POOL, OP_UPSAMPLE };

struct Op {
  OpKind kind;
  // conv / detect
  aq_conv_desc conv{};
  aq_detect_desc det{};
  std::vector<uint16_t> w_host;
  std::vector<float> b_host;
  __half* w_dev = nullptr;
  __half* w_dev_dense = nullptr;  // 3x3 convs with c_in <= 16: [c_out][3 * 64], K = tap * 16 + channel (see Cand::b_dense)
  float* b_dev = nullptr;
  // EPI_CHAIN (aq_plan_conv_chain): a 1x1 conv evaluated behind this conv inside the same kernel
  bool chained = false;
  int chain_store1 = 1;
  aq_conv_desc conv2{};
  std::vector<uint16_t> w2_host;
  std::vector<float> b2_host;
  __half* w2_dev = nullptr;
  float* b2_dev = nullptr;
  ConvKParams kp{};               // geometry + the tensor maps of sub-batch 0
  std::vector<ConvKParams> kps;   // staged ops: one parameter block per sub-batch (same geometry, shifted tensor maps)
  int stage = -1;
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
  std::vector<Stage> stages;
  int cur_stage = -1;  // stage being declared (between aq_plan_stage_begin / _end)
  cudaStream_t lane_stream[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};  // [0] unused: lane 0 runs on the caller's stream
  cudaEvent_t ev_fork = nullptr, ev_join[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr};
  float* pred = nullptr;
  // early NMS candidate filter: the Detect epilogues set one bit per passing prediction row (nc = 3 plans)
  uint32_t* det_mask = nullptr;
  float* det_conf_dev = nullptr;
  float det_conf_now = -1.f;    // value *det_conf_dev holds
  float det_conf_stage = 0.f;   // host source of the pending update
  float mask_conf = -1.f;       // conf_thres the mask of the latest forward pass was built with (-1: no usable mask)
  int use_det_mask = 1;         // AQUA_NMS_MASK=kernel: always rebuild the mask from pred with nms_mask_kernel
  // Alternate the tile order from op to op (AQUA_SERPENTINE=0: every op ascending).  Measured with interleaved rounds
  // (tools/ab_env.py: the first spec timed on a cool chip looks 2-4 % faster, which hid it at first): 4.524 vs 4.615 ms
  // per batch-64 conv stack = -2.0 % (profiles/r2_ab_chain_serpentine.txt).
  int serpentine = 1;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  int n_launches = 0;
  int pdl = 1;  // programmatic dependent launch between conv kernels (AQUA_PDL=0 disables)
  // Tiling choice per conv: 2 (default) = the table loaded with aq_tune_load decides, a shape it does not list is
  // timed on the device at plan time (logged); 1 = AQUA_AUTOTUNE=1, the same but meant for building a table
  // (AQUA_TUNE_CACHE=<file> collects the choices); 0 = AQUA_AUTOTUNE=0, the cost model alone (no timing launches).
  int autotune = 2;
  int tune_log = 1, tune_misses = 0;
  int tune_iters = 3;  // timed launches per candidate tiling (AQUA_TUNE_ITERS; table builds use more)
  // AQUA_TUNE_CACHE=<file>: tilings chosen by an earlier run are replayed instead of re-timed (one "signature choice"
  // line per conv); new measurements are appended.  Lets a profiler pass see the step, not the tuning launches.
  std::string tune_path;
  std::map<std::string, int> tune_cache;
  // preprocess state: host tiles are staged through two device buffers filled on a dedicated copy stream, so the
  // H2D copy of batch i+1 overlaps the kernels of batch i (aq_detect_submit / aq_detect_wait)
  uint8_t* stage_u8[2] = {nullptr, nullptr};
  size_t stage_bytes[2] = {0, 0};
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_h2d[2] = {nullptr, nullptr};    // staging buffer filled (copy stream)
  cudaEvent_t ev_pre[2] = {nullptr, nullptr};    // staging buffer consumed by the preprocess kernel (compute stream)
  cudaEvent_t ev_done[2] = {nullptr, nullptr};   // detections of the batch are in host memory (compute stream)
  bool pre_recorded[2] = {false, false};
  int pipe_slot = 0;
  // stage boundaries of a submitted batch on the compute stream (timing events): start, preprocess done, conv stack +
  // Detect decode done, NMS done -- upstream detect.py's three "Speed:" stages, measured on the device
  cudaEvent_t ev_stage[2][4] = {{nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr}};
  bool stage_timed[2] = {false, false};
  int tab_in_h = -1, tab_in_w = -1;
  void* tab_dev = nullptr;
  PreParams pre{};
  void* stats_scratch = nullptr;  // aq_tile_stats: column sums + TileStats
  int* win_list_dev = nullptr;    // aq_detect_window_list: window indices of one batch
  int win_list_cap = 0;
  size_t stats_bytes = 0;
  unsigned int* stats_static = nullptr;        // aq_tile_stats_static: device address of the library's static column-sum array
  cudaError_t stats_static_err = cudaSuccess;  // ... or why it could not be resolved at aq_create
  void* jpeg_scratch = nullptr;   // aq_jpeg_reconstruct: component planes (+ staged coefficients / tables for host input)
  size_t jpeg_bytes = 0;
  // aq_jpeg_reconstruct_packed: two device staging areas filled on the copy stream
  void* jpk_stage[2] = {nullptr, nullptr};
  size_t jpk_bytes[2] = {0, 0};
  cudaEvent_t ev_jpk_h2d[2] = {nullptr, nullptr};   // staging area filled (copy stream)
  cudaEvent_t ev_jpk_used[2] = {nullptr, nullptr};  // staging area consumed by the IDCT kernel (compute stream)
  bool jpk_recorded[2] = {false, false};
  int jpk_slot = 0;
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

// Images one launch of `op` processes, and the number of launches (sub-batches) per forward pass.
int op_batch(const aq_ctx* c, const Op& op) { return op.stage >= 0 ? c->stages[op.stage].sub : c->batch; }
int op_subs(const aq_ctx* c, const Op& op) { return op.stage >= 0 ? c->batch / c->stages[op.stage].sub : 1; }
int op_epi(const Op& op) { return op.kind == OP_DETECT ? EPI_DETECT : (op.chained ? EPI_CHAIN : EPI_STORE); }
// Images a buffer holds.
int buf_images(const aq_ctx* c, const Buffer& b) { return b.stage >= 0 ? c->stages[b.stage].sub * c->stages[b.stage].lanes : c->batch; }
// First element of buffer `id` as sub-batch `s` of `op` sees it: a plan-wide buffer is entered at image s * sub, a
// transient buffer of the op's stage at the copy of the lane that runs sub-batch s.
__half* buf_base(const aq_ctx* c, int id, const Op& op, int s) {
  const Buffer& b = c->bufs[id];
  if (op.stage < 0) return b.ptr;
  const Stage& st = c->stages[op.stage];
  const size_t img = (size_t)b.h * b.w * b.c;
  return b.ptr + (size_t)(b.stage >= 0 ? s % st.lanes : s) * st.sub * img;
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

// Pick the spatial M block (tw x th <= 128 pixels) that wastes the fewest MMA rows on an (h, w) map.
void pick_spatial_tile(int h, int w, int* tw_out, int* th_out, bool even = false) {
  double best = -1;
  int btw = 16, bth = 8;
  for (int tw = 1; tw <= 128; ++tw) {
    if (tw > w && tw != 1) break;
    if (even && (tw & 1)) continue;
    int th = 128 / tw;
    if (th > h) th = h;
    if (even) th &= ~1;
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

// One candidate tiling of a conv; fill_candidate() completes the shared-memory / TMEM plan and the cost.
struct Cand {
  int mode = 0, mblk = 1, mx = 1, my = 1, tw = 0, th = 0;
  int cluster = 1;  // CTAs per thread-block cluster sharing each weight tile by multicast (1, 2, 4)
  int pair = 0;     // 1: cluster of two CTAs running tcgen05.mma.cta_group::2 (each keeps half of every weight tile)
  int epi_sets = 1; // epilogue organisation: 1 = sixteen warps in lockstep, 2 = two sets of eight on alternate groups
  int pref = 0;     // shared-memory split: 0 = staging / activation rings first, 1 = weight ring first (streamed weights)
  int b_dense = 0;  // HALO with c_in <= 16: four taps share one 64-wide weight chunk (resident weights = 3 slots, not 9)
  // derived
  int tiles_x = 1, tiles_y = 1, m_tiles = 0, pw = 0, ph = 0;
  int na = 0, nb = 0, b_res = 0, n_stg = 0, acc_stages = 0;
  int acc2_stages = 0;  // chained conv: TMEM stages of its accumulator
  uint32_t a_slot = 0, a_tx = 0;
  double cycles = 1e30;
  bool ok = false;
};

// Rough per-SM cycle model used only to rank candidates: each tile costs the slowest of its MMA time
// (128 x N x 16 per N/2 cycles), its L2 -> SM operand traffic (~40 B/clk/SM measured), the TMA issue rate of
// one producer thread (~600 clk per box, profiles/r1_hw_probe_v1.txt T4) and the epilogue.
void fill_candidate(Cand& cd, int batch, int ho, int wo, int tile_n, int n_tiles_n, int n_steps, int cpt, int num_sms,
                    int extra_smem = 0, int extra_tmem = 0, int chain_kg = 0) {  // what a chained conv adds (resident weights, accumulator columns, K groups)
  const int groups = (tile_n + 63) / 64;
  const uint32_t b_slot = (uint32_t)round_up((cd.pair ? tile_n / 2 : tile_n) * 128, 1024);
  (void)groups;
  const uint32_t stg = kTileM * 128;  // staging ring: 16 KB buffers, one 64-channel group of an M block each
  int a_instr_per_group = 1, n_groups = n_steps;
  if (cd.mode == CONV_MODE_FLAT) {
    const int m_blocks = (batch * ho * wo + kTileM - 1) / kTileM;
    cd.m_tiles = (m_blocks + cd.mblk - 1) / cd.mblk;
    cd.a_slot = cd.a_tx = (uint32_t)cd.mblk * kTileM * 128;
  } else if (cd.mode == CONV_MODE_HALO) {
    cd.mblk = cd.mx * cd.my;
    cd.pw = 8 * cd.mx + 2;
    cd.ph = 16 * cd.my + 2;
    cd.tiles_x = (wo + 8 * cd.mx - 1) / (8 * cd.mx);
    cd.tiles_y = (batch * (ho + 1) + 16 * cd.my - 1) / (16 * cd.my);  // tall rows: images + one zero row each
    cd.m_tiles = cd.tiles_x * cd.tiles_y;
    cd.a_tx = (uint32_t)cd.pw * cd.ph * 128;
    cd.a_slot = (uint32_t)round_up((int)cd.a_tx, 1024);
    n_groups = cpt;
  } else if (cd.mode == CONV_MODE_HALO2) {
    // stride 2: four parity planes (input y-parity, x-parity) of (8mx+1) x (16my+1) half-resolution pixels each,
    // one ring slot per plane; plane order (1,1) (1,0) (0,1) (0,0) carries 4 / 2 / 2 / 1 taps
    cd.mblk = cd.mx * cd.my;
    cd.pw = 8 * cd.mx + 1;
    cd.ph = 16 * cd.my + 1;
    cd.tiles_x = (wo + 8 * cd.mx - 1) / (8 * cd.mx);
    cd.tiles_y = (batch * (ho + 1) + 16 * cd.my - 1) / (16 * cd.my);
    cd.m_tiles = cd.tiles_x * cd.tiles_y;
    cd.a_tx = (uint32_t)cd.pw * cd.ph * 128;
    cd.a_slot = (uint32_t)round_up((int)cd.a_tx, 1024);
    n_groups = cpt * 4;
  } else {
    cd.tiles_x = (wo + cd.tw - 1) / cd.tw;
    cd.tiles_y = (ho + cd.th - 1) / cd.th;
    const int m_blocks = batch * cd.tiles_x * cd.tiles_y;
    cd.m_tiles = (m_blocks + cd.mblk - 1) / cd.mblk;
    cd.a_tx = (uint32_t)cd.mblk * cd.tw * cd.th * 128;
    cd.a_slot = (uint32_t)cd.mblk * kTileM * 128;
    a_instr_per_group = cd.mblk;
  }
  const int acc_pitch = round_up(tile_n, 32);  // each accumulator starts on a 32-column boundary
  // extra_tmem = columns of ONE chained accumulator stage: prefer (2, 2) stages, then one chained stage, then one own stage
  if (cd.mblk * acc_pitch + extra_tmem > 512) return;
  if (2 * cd.mblk * acc_pitch + 2 * extra_tmem <= 512) { cd.acc_stages = 2; cd.acc2_stages = 2; }
  else if (2 * cd.mblk * acc_pitch + extra_tmem <= 512) { cd.acc_stages = 2; cd.acc2_stages = 1; }
  else if (cd.mblk * acc_pitch + 2 * extra_tmem <= 512) { cd.acc_stages = 1; cd.acc2_stages = 2; }
  else { cd.acc_stages = 1; cd.acc2_stages = 1; }
  // ---- shared memory: minimal plan first, then grow the rings
  const long budget = kConvMaxDynSmem - 1024 - extra_smem;
  if (cd.b_dense && (cd.mode != CONV_MODE_HALO || cpt != 1 || n_tiles_n != 1)) cd.b_dense = 0;
  const int res_slots = cd.b_dense ? (n_steps + 3) / 4 : n_steps;  // weight slots that hold the whole (single) N tile
  cd.b_res = (n_tiles_n == 1 && res_slots <= kMaxB && (long)res_slots * b_slot <= 112 * 1024) ? 1 : 0;
  cd.na = 2;
  cd.n_stg = 2;
  cd.nb = cd.b_res ? res_slots : 3;
  auto used = [&]() { return (long)cd.na * cd.a_slot + (long)cd.nb * b_slot + (long)cd.n_stg * stg; };
  if (used() > budget) {
    if (!cd.b_res) return;
    cd.b_res = 0;
    cd.b_dense = 0;
    cd.nb = 3;
    if (used() > budget) return;
  }
  // HALO: one patch per 64-channel chunk; with a single chunk per tile (c_in <= 64) the A ring depth is the number of
  // TILES in flight, and a patch load takes 2-3k clk under load -- such layers want a deep ring, the others 3 slots
  const int halo_na_cap = cd.mode == CONV_MODE_HALO ? (cpt == 1 ? kMaxA : 3) : kMaxA;
  auto grow = [&](int& v, int cap, long unit) {
    while (v < cap && used() + unit <= budget) ++v;
  };
  // Growth order: a third staging buffer first (a TMA store needs ~1-2k clk before its shared-memory source may
  // be overwritten, so two buffers serialise the epilogue on it), then deeper operand rings.
  if (cd.pref == 1) {
    // Streamed weights are latency bound through a shallow ring: a slot is busy from its TMA issue (~1-1.5k clk to
    // land) until the MMAs that read it retire, so nb slots carry nb - 1 tiles in flight; layers whose operand
    // traffic L2 -> SM is the limit (3x3 convs with few M blocks per tile) want the bytes in the weight ring
    // rather than in staging buffers.  The autotuner times both splits.
    if (cd.b_res) return;
    if (cd.mode == CONV_MODE_HALO2) grow(cd.na, 4, cd.a_slot);
    grow(cd.nb, 6, b_slot);
    grow(cd.n_stg, 3, stg);
    if (cd.mode == CONV_MODE_HALO) grow(cd.na, halo_na_cap, cd.a_slot);
    grow(cd.nb, 12, b_slot);
    grow(cd.n_stg, 4, stg);
    grow(cd.na, halo_na_cap, cd.a_slot);
  } else {
    grow(cd.n_stg, 3, stg);
    if (cd.mode == CONV_MODE_HALO) grow(cd.na, halo_na_cap, cd.a_slot);
    if (cd.mode == CONV_MODE_HALO2) grow(cd.na, 4, cd.a_slot);
    if (!cd.b_res) grow(cd.nb, 5, b_slot);
    grow(cd.n_stg, 4, stg);
    grow(cd.na, halo_na_cap, cd.a_slot);
    if (!cd.b_res) grow(cd.nb, 12, b_slot);
  }
  // ---- cost
  const double mma = (double)n_steps * cd.mblk * 4.0 * (tile_n / 2.0) + 40.0 * n_steps;
  const double a_bytes = (double)n_groups * cd.a_tx;
  if (cd.cluster > 1 && !cd.pair && (cd.b_res || (tile_n / cd.cluster) % 8 || num_sms / cd.cluster < 1)) return;  // nothing to share / swizzle atom split
  if (cd.pair && (cd.cluster != 2 || (tile_n / 2) % 8 || cd.b_dense)) return;
  const double b_bytes = cd.b_res ? 0.0 : (double)n_steps * tile_n * 128.0 / cd.cluster;  // per SM: a multicast cluster still delivers every byte to every SM (the model is optimistic there), a pair really halves it
  const double o_bytes = (double)cd.mblk * kTileM * tile_n * 2.0;
  const double l2 = (a_bytes + b_bytes + o_bytes) / 40.0;
  // TMA: one producer thread per operand, ~600 clk per box and ~4500 clk latency per ring slot; the HALO
  // patch is gathered by 128 cp.async threads (~1.5 clk per 128-byte row)
  const double lat = 4500.0;
  double issue = cd.b_res ? 0.0 : fmax(600.0 * n_steps, lat * n_steps / cd.nb);
  if (cd.mode == CONV_MODE_HALO || cd.mode == CONV_MODE_HALO2) issue = fmax(issue, 1.5 * cd.pw * cd.ph * n_groups);
  else issue = fmax(issue, fmax(600.0 * n_groups * a_instr_per_group, lat * n_groups / cd.na));
  const double epi = (double)cd.mblk * tile_n * 8.0 + 400.0;
  double tile_cyc = fmax(fmax(mma, l2), fmax(issue, epi)) + 300.0;
  if (cd.acc_stages == 1) tile_cyc += epi;
  const long total_tiles = (long)((cd.m_tiles + cd.cluster - 1) / cd.cluster) * n_tiles_n;  // cluster-tiles
  const long n_clusters = num_sms / cd.cluster;
  const long per_sm = (total_tiles + n_clusters - 1) / n_clusters;
  cd.cycles = per_sm * tile_cyc + 4000.0 + (cd.cluster > 1 ? 500.0 : 0.0);
  cd.ok = true;
}

// Fill everything of ConvKParams except device pointers / tensor maps.  The feasible tilings (UMMA N x A-feed
// mode x M blocks per tile) are enumerated; `choice` < 0 takes the one the cost model ranks first, otherwise the
// choice-th one (plan-time autotuning times them all on the device).  *n_choices receives their number.
int plan_conv_geometry(aq_ctx* c, Op& op, int choice = -1, int* n_choices = nullptr) {
  const bool det = op.kind == OP_DETECT;
  ConvKParams& k = op.kp;
  memset(&k, 0, sizeof k);
  const int src = det ? op.det.src : op.conv.src;
  const int c_in = det ? op.det.c_in : op.conv.c_in;
  const int c_out = det ? 3 * (op.det.nc + 5) : op.conv.c_out;
  const int ksize = det ? 1 : op.conv.k;
  const int stride = det ? 1 : op.conv.stride;
  const Buffer& sb = c->bufs[src];
  const int ho = sb.h / stride, wo = sb.w / stride;
  if (c_in % 8) return fail(c, AQ_ERR_UNSUPPORTED, "conv c_in %d is not a multiple of 8", c_in);
  if (!((ksize == 1 && stride == 1) || (ksize == 3 && (stride == 1 || stride == 2))))
    return fail(c, AQ_ERR_UNSUPPORTED, "conv k=%d s=%d unsupported", ksize, stride);
  if (stride == 2 && ((sb.h | sb.w) & 1)) return fail(c, AQ_ERR_UNSUPPORTED, "stride-2 conv on odd map");
  k.stride = stride;
  k.taps = ksize * ksize;
  k.c_in = c_in;
  k.cpt = (c_in + 63) / 64;
  k.n_steps = k.taps * k.cpt;
  const bool chain = !det && op.chained;
  const int c_in2 = chain ? op.conv2.c_in : 0, tile_n2 = chain ? round_up(op.conv2.c_out, 16) : 0;
  const int chain_kg = (c_in2 + 63) / 64, b2_slot = round_up(tile_n2 * 128, 1024), acc2_pitch = round_up(tile_n2, 32);
  // ---- UMMA N options
  std::vector<int> tile_ns;
  if (det) {
    tile_ns.push_back(round_up(c_out, 16));
  } else if (op.conv.tile_n > 0 && !chain) {
    tile_ns.push_back(op.conv.tile_n);
  } else if (c_out <= 256) {
    tile_ns.push_back(round_up(c_out, 16));
  } else {
    for (int n = 256; n >= 128; n -= 16)
      if (c_out % n == 0) tile_ns.push_back(n);
    if (tile_ns.empty()) tile_ns.push_back(largest_divisor_tile_n(c_out));
  }
  // ---- A-feed candidates
  std::vector<Cand> shapes;
  const int want_mblk = det ? 0 : op.conv.mblk;
  const int want_mode = det ? 0 : op.conv.mode;  // 0 auto, 1 halo, 2 tap (3x3/s1 only)
  const int want_cluster = det ? 1 : op.conv.cluster;  // 0 auto, else 1 / 2 / 4
  const int up_c = det ? 0 : op.conv.up_c;
  if (ksize == 1 && up_c > 0) {
    // folded upsample: spatial (even x even) M blocks so the replicated 2x2 pixel groups stay inside a box
    for (int mb = 1; mb <= 2; ++mb) {
      Cand cd;
      cd.mode = CONV_MODE_TAP;
      cd.mblk = mb;
      pick_spatial_tile(ho, wo, &cd.tw, &cd.th, true);
      shapes.push_back(cd);
    }
  } else if (ksize == 1) {
    for (int mb = 1; mb <= 2; ++mb) {
      Cand cd;
      cd.mode = CONV_MODE_FLAT;
      cd.mblk = mb;
      shapes.push_back(cd);
    }
  } else {
    if (stride == 1 && want_mode != 2) {
      const int arr[5][2] = {{1, 1}, {2, 1}, {1, 2}, {2, 2}, {4, 1}};
      for (auto& a : arr) {
        Cand cd;
        cd.mode = CONV_MODE_HALO;
        cd.mx = a[0];
        cd.my = a[1];
        cd.mblk = a[0] * a[1];
        cd.b_dense = (c_in <= 16 && !getenv("AQUA_NO_DENSE")) ? 1 : 0;
        shapes.push_back(cd);
      }
    }
    if (stride == 2 && want_mode != 2) {
      const int arr[5][2] = {{1, 1}, {2, 1}, {1, 2}, {2, 2}, {4, 1}};
      for (auto& a : arr) {
        Cand cd;
        cd.mode = CONV_MODE_HALO2;
        cd.mx = a[0];
        cd.my = a[1];
        cd.mblk = a[0] * a[1];
        shapes.push_back(cd);
      }
    }
    if ((stride == 2 && want_mode != 1) || (stride == 1 && want_mode != 1)) {
      for (int mb = 1; mb <= 2; ++mb) {
        Cand cd;
        cd.mode = CONV_MODE_TAP;
        cd.mblk = mb;
        pick_spatial_tile(ho, wo, &cd.tw, &cd.th);
        shapes.push_back(cd);
      }
    }
  }
  struct Choice { int tile_n, n_tiles_n; Cand cd; };
  std::vector<Choice> choices;
  for (int tn : tile_ns) {
    if (tn <= 0 || tn > 256 || tn % 16) continue;
    const int ntn = (c_out + tn - 1) / tn;
    if (ntn * tn > kMaxCout) continue;
    for (Cand cd0 : shapes) {
      if (want_mblk > 0 && cd0.mblk != want_mblk) continue;
      // want_cluster: 0 auto, 1 / 2 / 4 multicast clusters, 12 = CTA pair (cta_group::2).  Auto = single CTAs, plus pairs as
      // candidates for the autotuner unless AQUA_PAIR=0
      const bool pair_auto = !(getenv("AQUA_PAIR") && atoi(getenv("AQUA_PAIR")) == 0);
      for (int ci = 0; ci < 4; ++ci) {
        const int cl = ci == 3 ? 2 : (1 << ci);
        const int pr = ci == 3 ? 1 : 0;
        // measured (profiles/r1_cluster_multicast_sweep.txt): sharing weight tiles by multicast does not speed any
        // layer up (cluster 2: +-1 %, cluster 4: slower), so multicast clusters are only planned on request
        if (want_cluster == 12) { if (!pr) continue; }
        else if (want_cluster > 0) { if (pr || cl != want_cluster) continue; }
        else if (pr ? (!pair_auto || chain) : cl != 1) continue;
        // AQUA_EPI_SETS=1|2 pins the epilogue organisation (tests); default: both are candidates
        const char* fs = getenv("AQUA_EPI_SETS");
        const int min_sets = (det || chain) ? 1 : (fs ? (atoi(fs) == 2 ? 2 : 1) : 1);
        const int max_sets = (det || chain) ? 1 : (fs ? min_sets : 2);  // the chained epilogue is the lockstep one
        for (int sets = min_sets; sets <= max_sets; ++sets)
          for (int pref = 0; pref < 2; ++pref) {
            Cand cd = cd0;
            cd.cluster = cl;
            cd.pair = pr;
            cd.pref = pref;
            cd.epi_sets = sets;
            fill_candidate(cd, op_batch(c, op), ho, wo, tn, ntn, k.n_steps, k.cpt, c->num_sms, chain ? chain_kg * b2_slot : 0,
                           chain ? acc2_pitch : 0, chain ? chain_kg : 0);
            if (!cd.ok) continue;
            if (chain && (ntn != 1 || cd.cluster != 1)) continue;
            if (pref == 1 && !choices.empty()) {  // same split as the default order: not a new choice
              const Cand& o = choices.back().cd;
              if (o.mode == cd.mode && o.mblk == cd.mblk && o.mx == cd.mx && o.na == cd.na && o.nb == cd.nb && o.n_stg == cd.n_stg && o.pair == cd.pair) continue;
            }
            choices.push_back({tn, ntn, cd});
          }
      }
    }
  }
  if (n_choices) *n_choices = (int)choices.size();
  if (choices.empty())
    return fail(c, AQ_ERR_UNSUPPORTED, "conv %dx%d c_in %d c_out %d: no tiling fits shared memory / TMEM (UMMA N must be a multiple of 16 <= 256)", ksize, ksize, c_in, c_out);
  int pick = choice;
  if (pick < 0) {
    pick = 0;
    for (int i = 1; i < (int)choices.size(); ++i)
      if (choices[i].cd.cycles < choices[pick].cd.cycles) pick = i;
  }
  if (pick >= (int)choices.size()) return fail(c, AQ_ERR_ARG, "conv tiling choice %d out of range", pick);
  const Cand best = choices[pick].cd;
  k.tile_n = choices[pick].tile_n;
  k.n_tiles_n = choices[pick].n_tiles_n;
  k.groups = (k.tile_n + 63) / 64;
  k.mode = best.mode;
  k.mblk = best.mblk;
  k.mx = best.mx;
  k.my = best.my;
  k.pw = best.pw;
  k.ph = best.ph;
  k.pw_inv = best.pw > 0 ? (65536 + best.pw - 1) / best.pw : 0;
  const bool patch = best.mode == CONV_MODE_HALO || best.mode == CONV_MODE_HALO2;
  k.tw = patch ? 8 : (best.mode == CONV_MODE_FLAT ? kTileM : best.tw);
  k.th = patch ? 16 : (best.mode == CONV_MODE_FLAT ? 1 : best.th);
  k.tiles_x = best.tiles_x;
  k.tiles_y = best.tiles_y;
  k.m_tiles = best.m_tiles;
  k.up_chunks = up_c / 64;
  k.up_hs = sb.h / 2;
  k.cshift = best.cluster == 4 ? 2 : (best.cluster == 2 ? 1 : 0);
  k.pair = best.pair;
  k.total_tiles = (best.m_tiles + best.cluster - 1) / best.cluster * k.n_tiles_n;
  k.a_group = best.mode == CONV_MODE_HALO ? 9 : 1;
  k.na = best.na;
  k.nb = best.nb;
  k.b_res = best.b_res;
  k.b_dense = best.b_res ? best.b_dense : 0;
  k.epi_sets = det ? 1 : best.epi_sets;
  k.n_stg = best.n_stg;
  k.acc_stages = best.acc_stages;
  k.n_mma = best.mblk > 1 ? 2 : 1;
  k.a_slot_bytes = best.a_slot;
  k.a_tx_bytes = best.a_tx;
  k.b_slot_bytes = (uint32_t)round_up((k.pair ? k.tile_n / 2 : k.tile_n) * 128, 1024);
  k.b_tx_bytes = (uint32_t)(k.pair ? k.tile_n / 2 : k.tile_n) * 128;  // per CTA
  k.stg_bytes = det ? 0 : (uint32_t)kTileM * 128;
  if (det) k.n_stg = 1;
  k.act = det ? ACT_NONE : op.conv.act;
  k.has_res = (!det && op.conv.res >= 0) ? 1 : 0;
  { const char* v = getenv("AQUA_DBG"); k.dbg = v ? atoi(v) : 0; }
  k.acc_pitch = round_up(k.tile_n, 32);
  k.acc_stride = k.mblk * k.acc_pitch;
  k.tmem_cols = pow2_at_least(k.acc_stages * k.acc_stride);
  if (chain) {
    k.chain = 1;
    k.chain_kg = chain_kg;
    k.chain_kk_last = (c_in2 - 64 * (chain_kg - 1) + 15) / 16;
    k.chain_store1 = op.chain_store1;
    k.tile_n2 = tile_n2;
    k.groups2 = (tile_n2 + 63) / 64;
    k.act2 = op.conv2.act;
    k.b2_slot_bytes = (uint32_t)b2_slot;
    k.acc2_pitch = (uint32_t)acc2_pitch;
    k.acc2_base = k.acc_stages * k.acc_stride;
    k.acc2_stages = best.acc2_stages;
    k.tmem_cols = pow2_at_least((int)(k.acc2_base + k.acc2_stages * k.acc2_pitch));
    k.epi_sets = 1;
  }
  return AQ_OK;
}

// Tensor maps of one planned conv for sub-batch `s` (needs device pointers), written into `k` (geometry already set).
int build_conv_maps(aq_ctx* c, Op& op, ConvKParams& k, int s) {
  const bool det = op.kind == OP_DETECT;
  const int src = det ? op.det.src : op.conv.src;
  const int c_in = det ? op.det.c_in : op.conv.c_in;
  const int src_c_off = det ? op.det.src_c_off : op.conv.src_c_off;
  const Buffer& sb = c->bufs[src];
  __half* const sb_ptr = buf_base(c, src, op, s);
  const int nb_img = op_batch(c, op);
  const uint64_t N = nb_img;
  const CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_128B;
  int rc;
  // ---- A: dim 0 is exactly the conv's input channels, so a 64-wide box past them is zero-filled
  if (k.mode == CONV_MODE_FLAT) {
    uint64_t dims[2] = {(uint64_t)c_in, N * sb.h * sb.w};
    uint64_t str[1] = {(uint64_t)sb.c * 2};
    uint32_t box[2] = {64, (uint32_t)(kTileM * k.mblk)};
    rc = encode_map(c, &k.tmA, sb_ptr + src_c_off, 2, dims, str, box, sw);
  } else if (k.mode == CONV_MODE_HALO) {
    // patch rows are loaded as boxes of 32 / 16 / 8 / 4 / 2 / 1 image rows (per-image segments of the tall patch)
    k.a_src = sb_ptr + src_c_off;
    k.a_n = nb_img; k.a_h = sb.h; k.a_w = sb.w; k.a_ct = sb.c; k.a_cin = c_in;
    uint64_t dims[4] = {(uint64_t)c_in, (uint64_t)sb.w, (uint64_t)sb.h, N};
    uint64_t str[3] = {(uint64_t)sb.c * 2, (uint64_t)sb.w * sb.c * 2, (uint64_t)sb.h * sb.w * sb.c * 2};
    rc = AQ_OK;
    for (int hb = 0; hb < 6 && !rc; ++hb) {
      uint32_t box[4] = {64, (uint32_t)k.pw, (uint32_t)(32 >> hb), 1};
      rc = encode_map(c, &k.tmH[hb], sb_ptr + src_c_off, 4, dims, str, box, sw);
    }
  } else if (k.mode == CONV_MODE_HALO2) {
    // per x-parity: (channel, W/2, y-parity, H/2, N); a plane segment is a box (64 ch, pw, 1, h rows, 1)
    k.a_src = sb_ptr + src_c_off;
    k.a_n = nb_img; k.a_h = sb.h / 2; k.a_w = sb.w / 2; k.a_ct = sb.c; k.a_cin = c_in;
    uint64_t dims[5] = {(uint64_t)c_in, (uint64_t)sb.w / 2, 2, (uint64_t)sb.h / 2, N};
    uint64_t str[4] = {(uint64_t)2 * sb.c * 2, (uint64_t)sb.w * sb.c * 2, (uint64_t)2 * sb.w * sb.c * 2,
                       (uint64_t)sb.h * sb.w * sb.c * 2};
    rc = AQ_OK;
    for (int hb = 0; hb < 6 && !rc; ++hb) {
      uint32_t box[5] = {64, (uint32_t)k.pw, 1, (uint32_t)(32 >> hb), 1};
      rc = encode_map(c, &k.tmH[hb], sb_ptr + src_c_off, 5, dims, str, box, sw);
      if (!rc) rc = encode_map(c, &k.tmH2[hb], sb_ptr + src_c_off + sb.c, 5, dims, str, box, sw);
    }
  } else if (k.stride == 1) {
    const int up_c = det ? 0 : op.conv.up_c;
    uint64_t dims[4] = {(uint64_t)(c_in - up_c), (uint64_t)sb.w, (uint64_t)sb.h, N};
    uint64_t str[3] = {(uint64_t)sb.c * 2, (uint64_t)sb.w * sb.c * 2, (uint64_t)sb.h * sb.w * sb.c * 2};
    uint32_t box[4] = {64, (uint32_t)k.tw, (uint32_t)k.th, 1};
    rc = encode_map(c, &k.tmA, sb_ptr + src_c_off, 4, dims, str, box, sw);
    if (!rc && up_c > 0) {
      // (channel, dx, W/2, dy, H/2 * N) over the half-resolution tensor; dx / dy replicate with byte stride 0
      // (accepted by cuTensorMapEncodeTiled and exact on sm_100a: probe T11, profiles/r1_hw_probe_t11.txt)
      const Buffer& ub = c->bufs[op.conv.up_src];
      __half* const ub_ptr = buf_base(c, op.conv.up_src, op, s);
      uint64_t ud[5] = {(uint64_t)up_c, 2, (uint64_t)ub.w, 2, (uint64_t)ub.h * N};
      uint64_t us[4] = {0, (uint64_t)ub.c * 2, 0, (uint64_t)ub.w * ub.c * 2};
      uint32_t ubox[5] = {64, 2, (uint32_t)k.tw / 2, 2, (uint32_t)k.th / 2};
      rc = encode_map(c, &k.tmUp, ub_ptr + op.conv.up_c_off, 5, ud, us, ubox, sw);
    }
  } else {
    // per x-parity: (channel, W/2, y-parity, H/2, N) -- a stride-2 gather becomes a dense box
    uint64_t dims[5] = {(uint64_t)c_in, (uint64_t)sb.w / 2, 2, (uint64_t)sb.h / 2, N};
    uint64_t str[4] = {(uint64_t)2 * sb.c * 2, (uint64_t)sb.w * sb.c * 2, (uint64_t)2 * sb.w * sb.c * 2,
                       (uint64_t)sb.h * sb.w * sb.c * 2};
    uint32_t box[5] = {64, (uint32_t)k.tw, 1, (uint32_t)k.th, 1};
    rc = encode_map(c, &k.tmA, sb_ptr + src_c_off, 5, dims, str, box, sw);
    if (rc) return rc;
    rc = encode_map(c, &k.tmA2, sb_ptr + src_c_off + sb.c, 5, dims, str, box, sw);
  }
  if (rc) return rc;
  // ---- B: weights [rows = c_out][K = taps * cpt * 64] (zero padded per tap); rows past c_out are zero-filled
  {
    const int rows = det ? 3 * (op.det.nc + 5) : op.conv.c_out;
    // b_dense: the second copy [c_out][ceil(taps / 4) * 64] with K index = tap * 16 + channel
    const uint64_t K = k.b_dense ? (uint64_t)(k.taps + 3) / 4 * 64 : (uint64_t)k.taps * k.cpt * 64;
    uint64_t dims[2] = {K, (uint64_t)rows};
    uint64_t str[1] = {K * 2};
    uint32_t box[2] = {64, (uint32_t)(k.tile_n >> k.cshift)};  // cluster: each CTA loads its share of the rows
    rc = encode_map(c, &k.tmB, k.b_dense ? op.w_dev_dense : op.w_dev, 2, dims, str, box, sw);
    if (rc) return rc;
  }
  k.bias = op.b_dev;
  if (!det && op.chained) {
    const uint64_t K2 = (uint64_t)k.chain_kg * 64;
    uint64_t dims[2] = {K2, (uint64_t)op.conv2.c_out};
    uint64_t str[1] = {K2 * 2};
    uint32_t box[2] = {64, (uint32_t)k.tile_n2};
    if ((rc = encode_map(c, &k.tmB2, op.w2_dev, 2, dims, str, box, sw))) return rc;
    k.bias2 = op.b2_dev;
    const Buffer& d2 = c->bufs[op.conv2.dst];
    __half* const d2_ptr = buf_base(c, op.conv2.dst, op, s);
    if (k.mode == CONV_MODE_FLAT) {
      uint64_t od[2] = {(uint64_t)op.conv2.c_out, N * d2.h * d2.w};
      uint64_t os[1] = {(uint64_t)d2.c * 2};
      uint32_t ob[2] = {64, (uint32_t)kTileM};
      if ((rc = encode_map(c, &k.tmOut2, d2_ptr + op.conv2.dst_c_off, 2, od, os, ob, sw))) return rc;
    } else {
      uint64_t od[4] = {(uint64_t)op.conv2.c_out, (uint64_t)d2.w, (uint64_t)d2.h, N};
      uint64_t os[3] = {(uint64_t)d2.c * 2, (uint64_t)d2.w * d2.c * 2, (uint64_t)d2.h * d2.w * d2.c * 2};
      uint32_t ob[4] = {64, (uint32_t)k.tw, (uint32_t)k.th, 1};
      if ((rc = encode_map(c, &k.tmOut2, d2_ptr + op.conv2.dst_c_off, 4, od, os, ob, sw))) return rc;
      for (int hb = 0; hb < 4; ++hb) {
        uint32_t bx[4] = {64, 8, (uint32_t)(8 >> hb), 1};
        if ((rc = encode_map(c, &k.tmOut2R[hb], d2_ptr + op.conv2.dst_c_off, 4, od, os, bx, sw))) return rc;
      }
    }
  }
  if (det) {
    k.pred = c->pred + (size_t)s * nb_img * c->rows * c->no;
    k.det_mask = (c->det_mask && op.det.nc == 3) ? c->det_mask + (size_t)s * nb_img * nms_mask_words(c->rows) : nullptr;
    k.det_conf = c->det_conf_dev;
    k.det_mask_words = nms_mask_words(c->rows);
    k.det_no = op.det.nc + 5;
    k.det_rows = c->rows;
    k.det_level_off = op.det.level_row_off;
    k.det_ny = sb.h;
    k.det_nx = sb.w;
    k.det_m_total = nb_img * sb.h * sb.w;
    k.det_stride = op.det.stride;
    for (int i = 0; i < 6; ++i) k.det_anchor[i] = op.det.anchors_px[i];
    return AQ_OK;
  }
  // ---- output map over the destination channel slice (dim 0 = c_out: the 64-wide store box is clipped)
  const Buffer& db = c->bufs[op.conv.dst];
  __half* const db_ptr = buf_base(c, op.conv.dst, op, s);
  if (k.mode == CONV_MODE_FLAT) {
    uint64_t dims[2] = {(uint64_t)op.conv.c_out, N * db.h * db.w};
    uint64_t str[1] = {(uint64_t)db.c * 2};
    uint32_t box[2] = {64, (uint32_t)kTileM};
    return encode_map(c, &k.tmOut, db_ptr + op.conv.dst_c_off, 2, dims, str, box, sw);
  }
  uint64_t dims[4] = {(uint64_t)op.conv.c_out, (uint64_t)db.w, (uint64_t)db.h, N};
  uint64_t str[3] = {(uint64_t)db.c * 2, (uint64_t)db.w * db.c * 2, (uint64_t)db.h * db.w * db.c * 2};
  uint32_t box[4] = {64, (uint32_t)k.tw, (uint32_t)k.th, 1};
  rc = encode_map(c, &k.tmOut, db_ptr + op.conv.dst_c_off, 4, dims, str, box, sw);
  if (rc || (k.mode != CONV_MODE_HALO && k.mode != CONV_MODE_HALO2)) return rc;
  for (int hb = 0; hb < 4 && !rc; ++hb) {
    uint32_t bx[4] = {64, 8, (uint32_t)(8 >> hb), 1};
    rc = encode_map(c, &k.tmOutR[hb], db_ptr + op.conv.dst_c_off, 4, dims, str, bx, sw);
  }
  return rc;
}

// Tensor maps of a planned conv: op.kp for sub-batch 0 and, for an op of a stage, one block per sub-batch.
int plan_conv_maps(aq_ctx* c, Op& op) {
  int rc = build_conv_maps(c, op, op.kp, 0);
  if (rc || op.stage < 0) return rc;
  const int n = op_subs(c, op);
  op.kps.assign((size_t)n, op.kp);
  for (int s = 1; s < n && !rc; ++s) rc = build_conv_maps(c, op, op.kps[s], s);
  return rc;
}

// Plan-time autotuning: launch every feasible tiling of this conv on its real buffers (1 warm-up + 3 timed
// launches, CUDA events on the context stream) and keep the fastest.  The cost model only orders the search.
std::string tune_signature(const aq_ctx* c, const Op& op, int n) {
  char b[256];
  if (op.kind == OP_DETECT) {
    const Buffer& sb = c->bufs[op.det.src];
    snprintf(b, sizeof b, "det:b%d:%dx%d:ct%d:ci%d:nc%d:n%d", op_batch(c, op), sb.h, sb.w, sb.c, op.det.c_in, op.det.nc, n);
  } else {
    const Buffer& sb = c->bufs[op.conv.src];
    const Buffer& db = c->bufs[op.conv.dst];
    snprintf(b, sizeof b, "conv:b%d:%dx%d:ct%d:ci%d:co%d:dt%d:k%d:s%d:a%d:r%d:tn%d:mb%d:md%d:n%d", op_batch(c, op), sb.h, sb.w, sb.c,
             op.conv.c_in, op.conv.c_out, db.c, op.conv.k, op.conv.stride, op.conv.act, op.conv.res >= 0 ? 1 : 0,
             op.conv.tile_n, op.conv.mblk, op.conv.mode * 8 + op.conv.cluster + 64 * op.conv.up_c, n);
    if (op.chained) snprintf(b + strlen(b), sizeof b - strlen(b), ":ch%dx%d:s%d", op.conv2.c_in, op.conv2.c_out, op.chain_store1);
  }
  return b;
}

int autotune_conv(aq_ctx* c, Op& op) {
  int n = 0;
  int rc = plan_conv_geometry(c, op, -1, &n);
  if (rc) return rc;
  if (n <= 1) return plan_conv_maps(c, op);
  const std::string sig = tune_signature(c, op, n);
  {
    auto it = c->tune_cache.find(sig);
    if (it != c->tune_cache.end() && it->second >= 0 && it->second < n && !plan_conv_geometry(c, op, it->second))
      return plan_conv_maps(c, op);
  }
  if (c->autotune == 2) {
    // the shipped table has no entry for this shape: time its tilings now -- and say so, because two processes
    // may then keep different tilings, whose fp16 results differ in the last bits
    if (c->tune_log) fprintf(stderr, "aqua_b200: no tuned tiling for %s; timing %d candidates on the device\n", sig.c_str(), n);
    ++c->tune_misses;
  }
  const int epi = op_epi(op);
  cudaEvent_t e0, e1;
  CU_CHECK(c, cudaEventCreate(&e0));
  CU_CHECK(c, cudaEventCreate(&e1));
  float best_ms = 1e30f;
  int best = -1;
  for (int i = 0; i < n; ++i) {
    if (plan_conv_geometry(c, op, i) || plan_conv_maps(c, op)) continue;
    cudaError_t e = launch_conv(op.kp, epi, c->num_sms, 0, c->stream);
    cudaEventRecord(e0, c->stream);
    for (int r = 0; r < c->tune_iters && e == cudaSuccess; ++r) e = launch_conv(op.kp, epi, c->num_sms, 0, c->stream);
    cudaEventRecord(e1, c->stream);
    if (e == cudaSuccess) e = cudaEventSynchronize(e1);
    if (e != cudaSuccess) {
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
      return fail(c, AQ_ERR_CUDA, "autotune: tiling %d of a conv failed: %s", i, cudaGetErrorString(e));
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (c->tune_log >= 2)
      fprintf(stderr, "aqua_b200: tune %s #%d tile_n %d mode %d mblk %d sets %d %s na %d nb %d stg %d res %d: %.4f ms\n", sig.c_str(), i, op.kp.tile_n,
              op.kp.mode, op.kp.mblk, op.kp.epi_sets, op.kp.pair ? "pair" : "single", op.kp.na, op.kp.nb, op.kp.n_stg, op.kp.b_res, ms / c->tune_iters);
    if (ms < best_ms) {
      best_ms = ms;
      best = i;
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (best < 0) return fail(c, AQ_ERR_UNSUPPORTED, "autotune: no tiling could be planned");
  c->tune_cache[sig] = best;
  if (!c->tune_path.empty())
    if (FILE* f = fopen(c->tune_path.c_str(), "a")) {
      fprintf(f, "%s %d\n", sig.c_str(), best);
      fclose(f);
    }
  if ((rc = plan_conv_geometry(c, op, best))) return rc;
  return plan_conv_maps(c, op);
}

// [c_out][taps][c_in] fp16 -> [c_out][taps][cpt * 64] with zero padding (so a 64-wide K chunk never straddles taps)
void repack_weights(const uint16_t* w, int c_out, int taps, int c_in, std::vector<uint16_t>& out) {
  const int cp = (c_in + 63) / 64 * 64;
  out.assign((size_t)c_out * taps * cp, 0);
  for (int o = 0; o < c_out; ++o)
    for (int t = 0; t < taps; ++t)
      memcpy(&out[((size_t)o * taps + t) * cp], &w[((size_t)o * taps + t) * c_in], (size_t)c_in * 2);
}

int check_buf(aq_ctx* c, int id, int c_off, int ch, const char* what) {
  if (id < 0 || id >= (int)c->bufs.size()) return fail(c, AQ_ERR_ARG, "%s: buffer id %d out of range", what, id);
  const Buffer& b = c->bufs[id];
  if (c_off < 0 || ch <= 0 || c_off + ch > b.c) return fail(c, AQ_ERR_ARG, "%s: channels [%d,%d) outside buffer %d (c=%d)", what, c_off, c_off + ch, id, b.c);
  if (c_off % 8) return fail(c, AQ_ERR_ARG, "%s: channel offset %d is not 16-byte aligned", what, c_off);
  return AQ_OK;
}

// Tag a new op with the stage being declared and check that every buffer it names is visible to it: a transient
// buffer exists only inside its own stage.
int tag_op(aq_ctx* c, Op& op, const int* ids, int n_ids) {
  op.stage = c->cur_stage >= 0 ? c->cur_stage : -1;
  for (int i = 0; i < n_ids; ++i) {
    if (ids[i] < 0) continue;
    const int bs = c->bufs[ids[i]].stage;
    if (bs >= 0 && bs != op.stage)
      return fail(c, AQ_ERR_ARG, "buffer %d is transient to stage %d and cannot be used by an op %s", ids[i], bs,
                  op.stage < 0 ? "outside any stage" : "of another stage");
  }
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
    if (o.w_dev_dense) cudaFree(o.w_dev_dense);
    if (o.b_dev) cudaFree(o.b_dev);
    if (o.w2_dev) cudaFree(o.w2_dev);
    if (o.b2_dev) cudaFree(o.b2_dev);
  }
  if (c->pred) cudaFree(c->pred);
  c->pred = nullptr;
  if (c->det_mask) cudaFree(c->det_mask);
  if (c->det_conf_dev) cudaFree(c->det_conf_dev);
  c->det_mask = nullptr;
  c->det_conf_dev = nullptr;
  c->det_conf_now = c->mask_conf = -1.f;
  c->bufs.clear();
  c->ops.clear();
  c->stages.clear();
  c->cur_stage = -1;
  c->planned = c->planning = false;
}

// One launch: sub-batch `sub` of op `op` on stream `s`.
int launch_op(aq_ctx* c, Op& op, int sub, cudaStream_t s) {
  cudaError_t e = cudaSuccess;
  switch (op.kind) {
    case OP_CONV: e = launch_conv(op.kps.empty() ? op.kp : op.kps[sub], op_epi(op), c->num_sms, c->pdl, s); break;
    case OP_DETECT: e = launch_conv(op.kps.empty() ? op.kp : op.kps[sub], EPI_DETECT, c->num_sms, c->pdl, s); break;
    case OP_POOL: {
      const Buffer& b = c->bufs[op.buf];
      e = launch_sppf_pool(buf_base(c, op.buf, op, sub), op_batch(c, op), b.h, b.w, b.c, op.c_off, op.c, s);
      break;
    }
    case OP_UPSAMPLE: {
      const Buffer& sb = c->bufs[op.buf];
      const Buffer& db = c->bufs[op.dst];
      e = launch_upsample2x(buf_base(c, op.buf, op, sub), op_batch(c, op), sb.h, sb.w, sb.c, op.c_off, op.c,
                            buf_base(c, op.dst, op, sub), db.c, op.dst_c_off, s);
      break;
    }
  }
  if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "launch of op kind %d failed: %s", (int)op.kind, cudaGetErrorString(e));
  return AQ_OK;
}

// The forward pass: ops in plan order; the ops of a stage run depth-first over its sub-batches, lanes > 1 as that
// many independent chains forked from / joined back into `s` with events (captured into the graph as branches).
int run_ops(aq_ctx* c, cudaStream_t s) {
  int rc = AQ_OK;
  if (c->det_mask)  // the Detect epilogues OR their candidates' bits into it
    CU_CHECK(c, cudaMemsetAsync(c->det_mask, 0, (size_t)c->batch * nms_mask_words(c->rows) * sizeof(uint32_t), s));
  size_t i = 0;
  while (i < c->ops.size() && !rc) {
    const int sid = c->ops[i].stage;
    if (sid < 0) {
      rc = launch_op(c, c->ops[i], 0, s);
      ++i;
      continue;
    }
    const Stage& st = c->stages[sid];
    const int n_sub = c->batch / st.sub;
    const int lanes = st.lanes < n_sub ? st.lanes : n_sub;
    if (lanes > 1) {
      CU_CHECK(c, cudaEventRecord(c->ev_fork, s));
      for (int l = 1; l < lanes; ++l) CU_CHECK(c, cudaStreamWaitEvent(c->lane_stream[l], c->ev_fork, 0));
    }
    // sub-batches are issued round-robin over the lanes so the branches of the graph advance together
    for (int sub = 0; sub < n_sub && !rc; ++sub) {
      cudaStream_t ls = (sub % lanes) ? c->lane_stream[sub % lanes] : s;
      for (int j = st.op_begin; j < st.op_end && !rc; ++j) rc = launch_op(c, c->ops[j], sub, ls);
    }
    if (lanes > 1 && !rc) {
      for (int l = 1; l < lanes; ++l) {
        CU_CHECK(c, cudaEventRecord(c->ev_join[l], c->lane_stream[l]));
        CU_CHECK(c, cudaStreamWaitEvent(s, c->ev_join[l], 0));
      }
    }
    i = (size_t)st.op_end;
  }
  return rc;
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
  p.row_pitch = (size_t)in_w * 3; p.windowed = 0; p.wins_y = 1; p.win0 = 0;
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

// Host tiles -> staging buffer `slot` on the copy stream; stream `s` (the compute stream) is made to wait for the
// copy.  The copy itself waits until the preprocess kernel that last read this staging buffer has finished.
cudaError_t create_pipe_events(aq_ctx* c) {
  cudaError_t e = cudaSuccess;
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&c->ev_h2d[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_pre[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming);
  }
  for (int i = 0; i < 2; ++i)
    for (int k = 0; k < 4 && e == cudaSuccess; ++k) e = cudaEventCreate(&c->ev_stage[i][k]);
  // stage lanes: branches of the forward graph
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
  for (int l = 1; l < kMaxLanes && e == cudaSuccess; ++l) {
    e = cudaStreamCreateWithFlags(&c->lane_stream[l], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_join[l], cudaEventDisableTiming);
  }
  return e;
}

int stage_tiles(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int src_is_device, cudaStream_t s,
                int slot, const uint8_t** dev_out) {
  if (src_is_device) {
    *dev_out = tiles;
    return AQ_OK;
  }
  const size_t bytes = (size_t)n * in_h * in_w * 3;
  if (bytes > c->stage_bytes[slot]) {
    if (c->stage_u8[slot]) {
      CU_CHECK(c, cudaDeviceSynchronize());
      cudaFree(c->stage_u8[slot]);
    }
    c->stage_u8[slot] = nullptr;
    CU_CHECK(c, cudaMalloc(&c->stage_u8[slot], bytes));
    c->stage_bytes[slot] = bytes;
  }
  if (c->pre_recorded[slot]) CU_CHECK(c, cudaStreamWaitEvent(c->copy_stream, c->ev_pre[slot], 0));
  CU_CHECK(c, cudaMemcpyAsync(c->stage_u8[slot], tiles, bytes, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaEventRecord(c->ev_h2d[slot], c->copy_stream));
  CU_CHECK(c, cudaStreamWaitEvent(s, c->ev_h2d[slot], 0));
  *dev_out = c->stage_u8[slot];
  return AQ_OK;
}

int preprocess_slot(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device, int slot,
                    cudaStream_t s);
int forward_nms_d2h(aq_ctx* c, int n, float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes,
                    int n_classes, float* dets_host, int* counts_host, cudaStream_t s, int slot);

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
  if (const char* v = getenv("AQUA_PDL")) c->pdl = atoi(v) != 0;
  if (const char* v = getenv("AQUA_AUTOTUNE")) c->autotune = atoi(v) != 0;
  if (const char* v = getenv("AQUA_TUNE_LOG")) c->tune_log = atoi(v);  // 2: one line per timed candidate
  if (const char* v = getenv("AQUA_NMS_MASK")) c->use_det_mask = strcmp(v, "kernel") != 0;
  if (const char* v = getenv("AQUA_SERPENTINE")) c->serpentine = atoi(v) != 0;
  if (const char* v = getenv("AQUA_TUNE_ITERS")) c->tune_iters = atoi(v) > 0 ? atoi(v) : 3;
  if (const char* v = getenv("AQUA_TUNE_CACHE")) {
    c->tune_path = v;
    if (FILE* f = fopen(v, "r")) {
      char sig[256];
      int choice;
      while (fscanf(f, "%255s %d", sig, &choice) == 2) c->tune_cache[sig] = choice;
      fclose(f);
    }
  }
  const char* what = "stream";
  if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (what = "copy stream", e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (what = "events", e = create_pipe_events(c)) != cudaSuccess ||
      (what = "conv smem attr", e = conv_kernels_init()) != cudaSuccess ||
      (what = "nms smem attr", e = nms_init()) != cudaSuccess ||
      (what = "sppf smem attr", e = sppf_init()) != cudaSuccess) {
    delete c;
    return fail(nullptr, AQ_ERR_CUDA, "init (%s): %s", what, cudaGetErrorString(e));
  }
  // set-up for aq_tile_stats_static (device `device` is current): the static scratch array's address on this device
  c->stats_static_err = tile_stats_static_scratch(&c->stats_static);
  *out = c;
  return AQ_OK;
}

void aq_destroy(aq_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  free_plan(c);
  for (int i = 0; i < 2; ++i) {
    if (c->stage_u8[i]) cudaFree(c->stage_u8[i]);
    if (c->ev_h2d[i]) cudaEventDestroy(c->ev_h2d[i]);
    if (c->ev_pre[i]) cudaEventDestroy(c->ev_pre[i]);
    if (c->ev_done[i]) cudaEventDestroy(c->ev_done[i]);
    for (int k = 0; k < 4; ++k)
      if (c->ev_stage[i][k]) cudaEventDestroy(c->ev_stage[i][k]);
  }
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  for (int l = 1; l < kMaxLanes; ++l) {
    if (c->lane_stream[l]) cudaStreamDestroy(c->lane_stream[l]);
    if (c->ev_join[l]) cudaEventDestroy(c->ev_join[l]);
  }
  if (c->tab_dev) cudaFree(c->tab_dev);
  if (c->nms_scratch) cudaFree(c->nms_scratch);
  if (c->stats_scratch) cudaFree(c->stats_scratch);
  if (c->win_list_dev) cudaFree(c->win_list_dev);
  if (c->jpeg_scratch) cudaFree(c->jpeg_scratch);
  for (int i = 0; i < 2; ++i) {
    if (c->jpk_stage[i]) cudaFree(c->jpk_stage[i]);
    if (c->ev_jpk_h2d[i]) cudaEventDestroy(c->ev_jpk_h2d[i]);
    if (c->ev_jpk_used[i]) cudaEventDestroy(c->ev_jpk_used[i]);
  }
  if (c->nms_dets) cudaFree(c->nms_dets);
  if (c->nms_counts) cudaFree(c->nms_counts);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

int aq_tune_load(aq_ctx* c, const char* path) {
  if (!c || !path) return AQ_ERR_ARG;
  FILE* f = fopen(path, "r");
  if (!f) return fail(c, AQ_ERR_IO, "tune_load: cannot open %s", path);
  char line[512], sig[256];
  int choice, n = 0;
  while (fgets(line, sizeof line, f)) {
    if (line[0] == '#' || sscanf(line, "%255s %d", sig, &choice) != 2) continue;
    // entries of the explicitly given AQUA_TUNE_CACHE (read at aq_create) win over the table's
    if (c->tune_cache.emplace(sig, choice).second) ++n;
  }
  fclose(f);
  return n;
}

int aq_tune_misses(aq_ctx* c) { return c ? c->tune_misses : AQ_ERR_ARG; }

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
  b.stage = c->cur_stage >= 0 ? c->cur_stage : -1;  // declared inside a stage: transient, one sub-batch per lane
  c->bufs.push_back(b);
  return (int)c->bufs.size() - 1;
}

int aq_plan_stage_begin(aq_ctx* c, int sub_batch, int lanes) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_stage_begin outside plan_begin/plan_end");
  if (c->cur_stage >= 0) return fail(c, AQ_ERR_STATE, "plan_stage_begin: stages do not nest");
  if (sub_batch <= 0 || lanes < 1 || lanes > kMaxLanes) return fail(c, AQ_ERR_ARG, "plan_stage_begin: sub_batch %d, lanes %d (1..%d)", sub_batch, lanes, kMaxLanes);
  if (sub_batch >= c->batch || c->batch % sub_batch) {
    // nothing to split (or the batch is not a whole number of sub-batches): the ops run as ordinary full-batch ops
    c->cur_stage = -2;
    return AQ_OK;
  }
  Stage st;
  st.sub = sub_batch;
  st.lanes = lanes < c->batch / sub_batch ? lanes : c->batch / sub_batch;
  st.op_begin = st.op_end = (int)c->ops.size();
  c->stages.push_back(st);
  c->cur_stage = (int)c->stages.size() - 1;
  return AQ_OK;
}

int aq_plan_stage_end(aq_ctx* c) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planning || c->cur_stage == -1) return fail(c, AQ_ERR_STATE, "plan_stage_end without plan_stage_begin");
  if (c->cur_stage >= 0) c->stages[c->cur_stage].op_end = (int)c->ops.size();
  c->cur_stage = -1;
  return AQ_OK;
}

int aq_plan_conv(aq_ctx* c, const aq_conv_desc* d) {
  if (!c || !d) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_conv outside plan_begin/plan_end");
  int rc;
  if (d->up_c < 0 || d->up_c >= d->c_in) return fail(c, AQ_ERR_ARG, "conv: up_c %d outside [0, c_in %d)", d->up_c, d->c_in);
  if ((rc = check_buf(c, d->src, d->src_c_off, d->c_in - d->up_c, "conv src"))) return rc;
  if ((rc = check_buf(c, d->dst, d->dst_c_off, d->c_out, "conv dst"))) return rc;
  if (d->up_c > 0) {
    if ((rc = check_buf(c, d->up_src, d->up_c_off, d->up_c, "conv upsample src"))) return rc;
    const Buffer& ub = c->bufs[d->up_src];
    const Buffer& s0 = c->bufs[d->src];
    if (d->k != 1 || d->stride != 1 || d->up_c % 64 || 2 * ub.h != s0.h || 2 * ub.w != s0.w)
      return fail(c, AQ_ERR_UNSUPPORTED, "conv: a folded upsample needs a 1x1/s1 conv, up_c %% 64 == 0 and a half-resolution source "
                                         "(got k %d s %d up_c %d, %dx%d vs %dx%d)", d->k, d->stride, d->up_c, ub.h, ub.w, s0.h, s0.w);
  }
  if (d->res >= 0 && (rc = check_buf(c, d->res, d->res_c_off, d->c_out, "conv residual"))) return rc;
  if (!d->weight || !d->bias) return fail(c, AQ_ERR_ARG, "conv: weight/bias NULL");
  const Buffer& sb = c->bufs[d->src];
  const Buffer& db = c->bufs[d->dst];
  if (d->stride != 1 && d->stride != 2) return fail(c, AQ_ERR_UNSUPPORTED, "conv stride %d", d->stride);
  if (sb.h / d->stride != db.h || sb.w / d->stride != db.w)
    return fail(c, AQ_ERR_ARG, "conv: src %dx%d / stride %d != dst %dx%d", sb.h, sb.w, d->stride, db.h, db.w);
  if (d->res >= 0 && (d->res != d->dst || d->res_c_off != d->dst_c_off))
    return fail(c, AQ_ERR_UNSUPPORTED, "conv: the residual operand must be the destination slice itself (dst += f(conv)); "
                                       "got res (%d,%d) vs dst (%d,%d)", d->res, d->res_c_off, d->dst, d->dst_c_off);
  if (d->act != AQ_ACT_NONE && d->act != AQ_ACT_SILU && d->act != AQ_ACT_SILU_FAST) return fail(c, AQ_ERR_ARG, "conv: unknown activation %d", d->act);
  Op op;
  op.kind = OP_CONV;
  op.conv = *d;
  const size_t nw = (size_t)d->c_out * d->k * d->k * d->c_in;
  (void)nw;
  repack_weights(d->weight, d->c_out, d->k * d->k, d->c_in, op.w_host);
  op.b_host.assign(d->bias, d->bias + d->c_out);
  op.conv.weight = nullptr;
  op.conv.bias = nullptr;
  {
    const int ids[3] = {d->src, d->dst, d->up_c > 0 ? d->up_src : -1};
    if ((rc = tag_op(c, op, ids, 3))) return rc;
  }
  if ((rc = plan_conv_geometry(c, op))) return rc;
  c->ops.push_back(std::move(op));
  return AQ_OK;
}

int aq_plan_conv_chain(aq_ctx* c, const aq_conv_desc* d1, const aq_conv_desc* d2, int store_first) {
  if (!c || !d1 || !d2) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_conv_chain outside plan_begin/plan_end");
  int rc;
  // the second conv: 1x1 / stride 1 over the FIRST c_in channels of the first conv's output, no residual, no upsample
  if (d2->k != 1 || d2->stride != 1 || d2->res >= 0 || d2->up_c != 0)
    return fail(c, AQ_ERR_UNSUPPORTED, "conv chain: the chained conv must be 1x1 / stride 1 without residual or folded upsample");
  if (d2->src != d1->dst || d2->src_c_off != d1->dst_c_off || d2->c_in > d1->c_out || d2->c_in % 16)
    return fail(c, AQ_ERR_UNSUPPORTED, "conv chain: the chained conv must read the first channels of the first conv's output slice "
                                       "(got src (%d,%d) c_in %d vs dst (%d,%d) c_out %d)", d2->src, d2->src_c_off, d2->c_in, d1->dst, d1->dst_c_off, d1->c_out);
  if (d1->res >= 0 || d1->c_out > 256 || d2->c_out > kMaxCout2)
    return fail(c, AQ_ERR_UNSUPPORTED, "conv chain: first conv without residual and <= 256 output channels, chained conv <= %d", kMaxCout2);
  if (d2->dst < 0 || d2->dst >= (int)c->bufs.size()) return fail(c, AQ_ERR_ARG, "chained conv dst: buffer id %d out of range", d2->dst);
  if ((rc = check_buf(c, d2->dst, d2->dst_c_off, d2->c_out, "chained conv dst"))) return rc;
  if (!d2->weight || !d2->bias) return fail(c, AQ_ERR_ARG, "conv chain: weight/bias NULL");
  if (d2->act != AQ_ACT_NONE && d2->act != AQ_ACT_SILU && d2->act != AQ_ACT_SILU_FAST) return fail(c, AQ_ERR_ARG, "conv chain: unknown activation %d", d2->act);
  if ((rc = aq_plan_conv(c, d1))) return rc;  // validates and appends the first conv
  Op& op = c->ops.back();
  const Buffer& db = c->bufs[d1->dst];
  const Buffer& d2b = c->bufs[d2->dst];
  const int want_stage = op.stage;
  if (db.h != d2b.h || db.w != d2b.w || (d2b.stage >= 0 && d2b.stage != want_stage)) {
    c->ops.pop_back();
    return fail(c, AQ_ERR_ARG, "conv chain: chained dst %dx%d (stage %d) does not match %dx%d (stage %d)", d2b.h, d2b.w, d2b.stage, db.h, db.w, want_stage);
  }
  op.chained = true;
  op.chain_store1 = store_first ? 1 : 0;
  op.conv2 = *d2;
  repack_weights(d2->weight, d2->c_out, 1, d2->c_in, op.w2_host);
  op.b2_host.assign(d2->bias, d2->bias + d2->c_out);
  op.conv2.weight = nullptr;
  op.conv2.bias = nullptr;
  if ((rc = plan_conv_geometry(c, op))) {  // no tiling leaves room for the chained conv's weights / accumulators
    c->ops.pop_back();
    return rc;
  }
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
  repack_weights(d->weight, n_out, 1, d->c_in, op.w_host);
  op.b_host.assign(d->bias, d->bias + n_out);
  op.det.weight = nullptr;
  op.det.bias = nullptr;
  if ((rc = tag_op(c, op, &d->src, 1))) return rc;
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
  if ((rc = tag_op(c, op, &buf, 1))) return rc;
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
  {
    const int ids[2] = {src, dst};
    if ((rc = tag_op(c, op, ids, 2))) return rc;
  }
  c->ops.push_back(std::move(op));
  return AQ_OK;
}

int aq_plan_end(aq_ctx* c) {
  if (!c) return AQ_ERR_ARG;
  if (!c->planning) return fail(c, AQ_ERR_STATE, "plan_end without plan_begin");
  if (c->cur_stage != -1) return fail(c, AQ_ERR_STATE, "plan_end inside an open stage");
  cudaSetDevice(c->device);
  for (auto& b : c->bufs) {
    const size_t bytes = (size_t)buf_images(c, b) * b.h * b.w * b.c * sizeof(__half);
    CU_CHECK(c, cudaMalloc(&b.ptr, bytes));
    CU_CHECK(c, cudaMemsetAsync(b.ptr, 0, bytes, c->stream));
  }
  {
    const size_t bytes = (size_t)c->batch * c->rows * c->no * sizeof(float);
    CU_CHECK(c, cudaMalloc(&c->pred, bytes));
    CU_CHECK(c, cudaMemsetAsync(c->pred, 0, bytes, c->stream));
    if (c->nc == 3 && c->use_det_mask) {  // the Detect epilogue's static 8-column path evaluates the candidate test
      const size_t mb = (size_t)c->batch * nms_mask_words(c->rows) * sizeof(uint32_t);
      CU_CHECK(c, cudaMalloc(&c->det_mask, mb));
      CU_CHECK(c, cudaMemsetAsync(c->det_mask, 0, mb, c->stream));
      CU_CHECK(c, cudaMalloc(&c->det_conf_dev, sizeof(float)));
      c->det_conf_stage = 0.25f;  // detect.py's default; aq_detect* update it when another threshold is asked for
      CU_CHECK(c, cudaMemcpyAsync(c->det_conf_dev, &c->det_conf_stage, sizeof(float), cudaMemcpyHostToDevice, c->stream));
      c->det_conf_now = 0.25f;
    }
  }
  int launches = 0;
  for (auto& op : c->ops) {
    launches += op_subs(c, op);
    if (op.kind != OP_CONV && op.kind != OP_DETECT) continue;
    const ConvKParams& k = op.kp;
    // weights: rows padded up to n_tiles_n * tile_n are handled by TMA OOB fill; bias is padded here
    CU_CHECK(c, cudaMalloc(&op.w_dev, op.w_host.size() * sizeof(uint16_t)));
    CU_CHECK(c, cudaMemcpyAsync(op.w_dev, op.w_host.data(), op.w_host.size() * sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream));
    if (op.kind == OP_CONV && op.conv.k == 3 && op.conv.c_in <= 16) {
      // dense-tap copy for HALO tilings: K index = tap * 16 + channel, three 64-wide chunks instead of nine
      const int taps = 9, co = op.conv.c_out;
      std::vector<uint16_t> dense((size_t)co * 192, 0);
      for (int o = 0; o < co; ++o)
        for (int t = 0; t < taps; ++t)
          memcpy(&dense[(size_t)o * 192 + t * 16], &op.w_host[((size_t)o * taps + t) * 64], 16 * sizeof(uint16_t));
      CU_CHECK(c, cudaMalloc(&op.w_dev_dense, dense.size() * sizeof(uint16_t)));
      CU_CHECK(c, cudaMemcpy(op.w_dev_dense, dense.data(), dense.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
    }
    std::vector<float> bias((size_t)kMaxCout, 0.f);  // padded to the largest N any tiling may read
    memcpy(bias.data(), op.b_host.data(), op.b_host.size() * sizeof(float));
    if (k.act == ACT_SILU_FAST)  // the kernel evaluates silu(x) = h + h * tanh(h) with h = acc / 2 + bias / 2
      for (auto& b : bias) b *= 0.5f;
    CU_CHECK(c, cudaMalloc(&op.b_dev, bias.size() * sizeof(float)));
    CU_CHECK(c, cudaMemcpy(op.b_dev, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
    if (op.chained) {
      CU_CHECK(c, cudaMalloc(&op.w2_dev, op.w2_host.size() * sizeof(uint16_t)));
      CU_CHECK(c, cudaMemcpy(op.w2_dev, op.w2_host.data(), op.w2_host.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
      std::vector<float> b2((size_t)kMaxCout2, 0.f);
      memcpy(b2.data(), op.b2_host.data(), op.b2_host.size() * sizeof(float));
      if (op.conv2.act == ACT_SILU_FAST)
        for (auto& b : b2) b *= 0.5f;
      CU_CHECK(c, cudaMalloc(&op.b2_dev, b2.size() * sizeof(float)));
      CU_CHECK(c, cudaMemcpy(op.b2_dev, b2.data(), b2.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    CU_CHECK(c, cudaStreamSynchronize(c->stream));
    int rc = c->autotune ? autotune_conv(c, op) : plan_conv_maps(c, op);
    if (rc) return rc;
    // serpentine tile order: every other op walks its tiles from the end, so it starts on what its producer wrote last
    // (still in L2) -- see ConvKParams::reverse.  Results do not depend on the order.
    op.kp.reverse = (c->serpentine && ((&op - c->ops.data()) & 1)) ? 1 : 0;
    for (auto& k : op.kps) k.reverse = op.kp.reverse;
    std::vector<uint16_t>().swap(op.w_host);
  }
  if (c->autotune) {  // the timing launches wrote into the activation buffers
    for (auto& b : c->bufs) CU_CHECK(c, cudaMemsetAsync(b.ptr, 0, (size_t)buf_images(c, b) * b.h * b.w * b.c * sizeof(__half), c->stream));
    CU_CHECK(c, cudaMemsetAsync(c->pred, 0, (size_t)c->batch * c->rows * c->no * sizeof(float), c->stream));
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
  if (n) *n = buf_images(c, b);  // a stage's transient buffer: sub-batch images per lane x lanes
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
  int rc = AQ_OK;
  // an op of a stage is timed over all of its sub-batches, back to back on one stream
  auto run_all = [&](Op& op) {
    int r = AQ_OK;
    for (int sub = 0; sub < op_subs(c, op) && !r; ++sub) r = launch_op(c, op, sub, c->stream);
    return r;
  };
  for (int i = 0; i < n && rc == AQ_OK; ++i) {
    Op& op = c->ops[i];
    rc = run_all(op);  // warm-up
    if (rc) break;
    cudaEventRecord(e0, c->stream);
    for (int k = 0; k < iters && rc == AQ_OK; ++k) rc = run_all(op);
    cudaEventRecord(e1, c->stream);
    if (cudaEventSynchronize(e1) != cudaSuccess) rc = fail(c, AQ_ERR_CUDA, "profile: op %d failed: %s", i, cudaGetErrorString(cudaGetLastError()));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    ms_per_op[i] = ms / iters;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return rc ? rc : n;
}

// In-situ variant: whole forward passes with an event after every launch, so each op sees the cache state and the
// neighbours it has in the real step (the isolated table above re-runs one op on warm buffers).  Launches follow the
// real order (a stage's ops depth-first over its sub-batches) on ONE stream: lanes are not overlapped here.
int aq_profile_ops_insitu(aq_ctx* c, int iters, float* ms_per_op, int cap) {
  if (!c || !ms_per_op || iters <= 0) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "profile before plan_end");
  cudaSetDevice(c->device);
  const int n = (int)c->ops.size();
  if (cap < n) return fail(c, AQ_ERR_ARG, "profile: need room for %d ops", n);
  std::vector<std::pair<int, int>> order;  // (op, sub-batch)
  for (int i = 0; i < n;) {
    const int sid = c->ops[i].stage;
    if (sid < 0) {
      order.push_back({i, 0});
      ++i;
      continue;
    }
    const Stage& st = c->stages[sid];
    for (int sub = 0; sub < c->batch / st.sub; ++sub)
      for (int j = st.op_begin; j < st.op_end; ++j) order.push_back({j, sub});
    i = st.op_end;
  }
  const int L = (int)order.size();
  std::vector<cudaEvent_t> ev(L + 1);
  for (auto& e : ev) CU_CHECK(c, cudaEventCreate(&e));
  std::vector<double> acc(n, 0.0);
  int rc = AQ_OK;
  for (int k = 0; k <= iters && rc == AQ_OK; ++k) {  // pass 0 = warm-up
    cudaEventRecord(ev[0], c->stream);
    for (int i = 0; i < L && rc == AQ_OK; ++i) {
      rc = launch_op(c, c->ops[order[i].first], order[i].second, c->stream);
      cudaEventRecord(ev[i + 1], c->stream);
    }
    if (rc == AQ_OK && cudaEventSynchronize(ev[L]) != cudaSuccess)
      rc = fail(c, AQ_ERR_CUDA, "profile: %s", cudaGetErrorString(cudaGetLastError()));
    for (int i = 0; i < L && rc == AQ_OK && k > 0; ++i) {
      float ms = 0;
      cudaEventElapsedTime(&ms, ev[i], ev[i + 1]);
      acc[order[i].first] += ms;
    }
  }
  for (int i = 0; i < n; ++i) ms_per_op[i] = (float)(acc[i] / iters);
  for (auto& e : ev) cudaEventDestroy(e);
  return rc ? rc : n;
}

// Launch-overhead probe: conv op `i` restricted to its first `waves` x SMs tiles, `iters` launches back to back (with
// PDL, as in the graph).  Time per launch over waves = 1, 2, 4 .. gives the cost of one wave (slope) and the fixed
// drain + fill cost of a launch (intercept).  Results are garbage (part of the output is not written).
int aq_profile_op_waves(aq_ctx* c, int i, int waves, int iters, float* ms_out) {
  if (!c || !ms_out || iters <= 0 || waves <= 0) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "profile before plan_end");
  if (i < 0 || i >= (int)c->ops.size()) return fail(c, AQ_ERR_ARG, "op %d out of range", i);
  Op& op = c->ops[i];
  if (op.kind != OP_CONV && op.kind != OP_DETECT) return fail(c, AQ_ERR_ARG, "op %d is not a conv", i);
  cudaSetDevice(c->device);
  ConvKParams kp = op.kp;
  const int cl = 1 << kp.cshift;
  const int cap = waves * (c->num_sms / cl);
  if (kp.total_tiles > cap) kp.total_tiles = cap;
  const int epi = op_epi(op);
  cudaEvent_t e0, e1;
  CU_CHECK(c, cudaEventCreate(&e0));
  CU_CHECK(c, cudaEventCreate(&e1));
  cudaError_t e = launch_conv(kp, epi, c->num_sms, c->pdl, c->stream);
  cudaEventRecord(e0, c->stream);
  for (int k = 0; k < iters && e == cudaSuccess; ++k) e = launch_conv(kp, epi, c->num_sms, c->pdl, c->stream);
  cudaEventRecord(e1, c->stream);
  if (e == cudaSuccess) e = cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "profile_op_waves: %s", cudaGetErrorString(e));
  *ms_out = ms / iters;
  return kp.total_tiles;
}

int aq_trace_op(aq_ctx* c, int i, long long* out, size_t cap_vals) {
  if (!c || !out) return AQ_ERR_ARG;
  if (!c->planned) return fail(c, AQ_ERR_STATE, "trace before plan_end");
  if (i < 0 || i >= (int)c->ops.size()) return fail(c, AQ_ERR_ARG, "trace: op %d out of range", i);
  Op& op = c->ops[i];
  if ((op.kind != OP_CONV && op.kind != OP_DETECT) || op.chained) return fail(c, AQ_ERR_ARG, "trace: op %d is not a plain conv", i);
  cudaSetDevice(c->device);
  const int total = op.kp.m_tiles * op.kp.n_tiles_n;
  const int grid = total < c->num_sms ? total : c->num_sms;
  const size_t vals = (size_t)grid * kTraceRoles * kTraceTiles * kTraceVals;
  if (cap_vals < vals) return fail(c, AQ_ERR_ARG, "trace: need room for %zu values", vals);
  const int epi = op.kind == OP_CONV ? EPI_STORE : EPI_DETECT;
  unsigned long long* dev = nullptr;
  CU_CHECK(c, cudaMalloc(&dev, vals * 8));
  CU_CHECK(c, cudaMemsetAsync(dev, 0, vals * 8, c->stream));
  ConvKParams kp = op.kp;
  cudaError_t e = launch_conv(kp, epi, c->num_sms, 0, c->stream);  // warm-up, untraced
  kp.trace = dev;
  if (e == cudaSuccess) e = launch_conv(kp, epi, c->num_sms, 0, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, dev, vals * 8, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFree(dev);
  if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "trace: %s", cudaGetErrorString(e));
  return grid;
}

int aq_debug_plan_conv(int batch, int h, int w, int c_in, int c_out, int k, int stride, int num_sms, int mblk, int mode,
                       char* buf, size_t cap) {
  if (!buf || batch <= 0 || h <= 0 || w <= 0) return AQ_ERR_ARG;
  aq_ctx tmp;  // host-only: no CUDA call is made on this path
  tmp.batch = batch;
  tmp.num_sms = num_sms > 0 ? num_sms : 148;
  Buffer s{}, d{};
  s.h = h; s.w = w; s.c = c_in;
  d.h = h / stride; d.w = w / stride; d.c = c_out;
  tmp.bufs.push_back(s);
  tmp.bufs.push_back(d);
  Op op;
  op.kind = OP_CONV;
  op.conv.src = 0; op.conv.c_in = c_in; op.conv.dst = 1; op.conv.c_out = c_out; op.conv.k = k; op.conv.stride = stride;
  op.conv.act = AQ_ACT_SILU; op.conv.res = -1; op.conv.mblk = mblk; op.conv.mode = mode;
  int rc = plan_conv_geometry(&tmp, op, getenv("AQUA_DEBUG_CHOICE") ? atoi(getenv("AQUA_DEBUG_CHOICE")) : -1);
  if (rc) {
    snprintf(buf, cap, "%s", tmp.err.c_str());
    return rc;
  }
  tmp.ops.push_back(op);
  return aq_op_describe(&tmp, 0, buf, cap);
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
    double flops = 2.0 * M * c_out * c_in * k.taps;
    double bytes = 2.0 * ((double)c->batch * sb.h * sb.w * c_in + M * c_out * (k.has_res ? 2 : 1)) +
                   2.0 * c_out * c_in * k.taps;
    if (op.chained) {  // + the chained 1x1 (its input never leaves the SM)
      flops += 2.0 * M * op.conv2.c_out * op.conv2.c_in;
      bytes += 2.0 * M * op.conv2.c_out + 2.0 * op.conv2.c_out * op.conv2.c_in - (op.chain_store1 ? 0.0 : 2.0 * M * c_out);
    }
    const int ksz = k.taps == 9 ? 3 : 1;
    static const char* mode_name[] = {"flat", "halo", "tap", "halo2"};
    char stg[96] = "";
    if (op.stage >= 0) snprintf(stg, sizeof stg, " sub=%dx%d lanes=%d", op_batch(c, op), op_subs(c, op), c->stages[op.stage].lanes);
    if (k.reverse) strncat(stg, " rev", sizeof stg - strlen(stg) - 1);
    if (op.chained) snprintf(stg + strlen(stg), sizeof stg - strlen(stg), " +1x1 %d->%d", op.conv2.c_in, op.conv2.c_out);
    snprintf(buf, cap, "%s k%d s%d %dx%d cin=%d cout=%d res=%d | M=%.0f %s mblk=%d(%dx%d) tileN=%d steps=%d na=%d nb=%d%s stg=%d acc=%d es=%d tile=%dx%d mt=%d nt=%d cl=%d%s | flop=%.4g bytes=%.4g",
             det ? "detect" : "conv", ksz, stride, sb.h, sb.w, c_in, c_out, k.has_res, M, mode_name[k.mode], k.mblk,
             k.mx, k.my, k.tile_n, k.n_steps, k.na, k.nb, k.b_res ? "(res)" : "", k.n_stg, k.acc_stages, k.epi_sets, k.tw, k.th,
             k.m_tiles, k.n_tiles_n, k.pair ? 12 : (1 << k.cshift), stg, flops, bytes);
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
  return preprocess_slot(c, tiles, n, in_h, in_w, bgr, src_is_device, 0, stream ? (cudaStream_t)stream : c->stream);
}

}  // extern "C"

namespace {
int preprocess_slot(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device, int slot,
                    cudaStream_t s) {
  if (!c->planned) return fail(c, AQ_ERR_STATE, "preprocess before plan_end");
  if (n <= 0 || n > c->batch) return fail(c, AQ_ERR_ARG, "preprocess: n %d outside plan batch %d", n, c->batch);
  cudaSetDevice(c->device);
  const Buffer& b0 = c->bufs[0];
  if (b0.h != c->net_h / 2 || b0.w != c->net_w / 2 || b0.c != 16)
    return fail(c, AQ_ERR_STATE, "buffer 0 must be the (net_h/2, net_w/2, 16) space-to-depth input");
  int rc = setup_preprocess(c, in_h, in_w);
  if (rc) return rc;
  const uint8_t* dev = nullptr;
  if ((rc = stage_tiles(c, tiles, n, in_h, in_w, src_is_device, s, slot, &dev))) return rc;
  PreParams p = c->pre;
  p.src = dev; p.n = n; p.bgr = bgr; p.out_s2d = b0.ptr; p.out_u8 = nullptr;
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][0], s));  // after the wait for the H2D copy: the kernel alone is timed
  c->stage_timed[slot] = false;
  CU_CHECK(c, launch_preprocess(p, s));
  if (!src_is_device) {
    CU_CHECK(c, cudaEventRecord(c->ev_pre[slot], s));
    c->pre_recorded[slot] = true;
  }
  return AQ_OK;
}
}  // namespace

extern "C" {

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
  c->mask_conf = c->det_mask ? c->det_conf_now : -1.f;
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
  p.cand_row = (int*)sc; sc += br * 4;
  if (pred == c->pred && c->det_mask && n_classes == 0 && conf_thres == c->mask_conf && batch <= c->batch) {
    p.cand_mask = c->det_mask;  // written by the Detect epilogues of the forward pass that produced `pred`
  } else {
    p.cand_mask = (uint32_t*)sc;
    CU_CHECK(c, launch_nms_mask(p, batch, s));
  }
  p.dets = dets_out; p.counts = counts_out; p.keep_rows = keep_rows_out;
  static const bool nms_clk = getenv("AQUA_NMS_CLK") != nullptr;
  p.dbg_clk = nullptr;
  p.dbg_mode = 0;
  if (nms_clk) {
    p.dbg_mode = atoi(getenv("AQUA_NMS_CLK"));  // phase timeline of image 0's CTA (debug): printed after a synchronising copy
    static long long* d_clk = nullptr;
    if (!d_clk) cudaMalloc(&d_clk, 16 * sizeof(long long));
    p.dbg_clk = d_clk;
    cudaMemsetAsync(d_clk, 0, 16 * sizeof(long long), s);
    CU_CHECK(c, launch_nms(p, batch, s));
    long long h[16] = {0};
    cudaMemcpyAsync(h, d_clk, sizeof h, cudaMemcpyDeviceToHost, s);
    cudaStreamSynchronize(s);
    fprintf(stderr, "nms phases (clk): count %lld, compaction %lld, sort %lld, suppression %lld (%lld clk to the end of the first 1024 candidates, %lld more to the end of the second block's shared catch-up), total %lld\n",
            h[1] - h[0], h[2] - h[1], h[3] - h[2], h[4] - h[3], h[5] - h[3], h[6] ? h[6] - h[5] : 0, h[4] - h[0]);
    if (h[8])
      fprintf(stderr, "nms sort, first radix pass (clk): zero + histogram walk %lld, scan over (digit, warp) %lld, digit scan %lld, scatter walk %lld\n", h[8] - h[2], h[9] - h[8],
              h[10] - h[9], h[11] - h[10]);
    return AQ_OK;
  }
  CU_CHECK(c, launch_nms(p, batch, s));
  return AQ_OK;
}

int aq_detect_submit(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device,
                     float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes, int n_classes,
                     float* dets_host, int* counts_host, void* stream) {
  if (!c || !tiles || !dets_host || !counts_host) return AQ_ERR_ARG;
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  const int slot = c->pipe_slot;
  c->pipe_slot ^= 1;
  int rc = preprocess_slot(c, tiles, n, in_h, in_w, bgr, src_is_device, slot, s);
  if (rc) return rc;
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][1], s));
  return forward_nms_d2h(c, n, conf_thres, iou_thres, max_det, agnostic, classes, n_classes, dets_host, counts_host, s, slot);
}

}  // extern "C"

namespace {
// conv stack + Detect decode + NMS + async copy of the detections to the host; records ev_done[slot]
int forward_nms_d2h(aq_ctx* c, int n, float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes,
                    int n_classes, float* dets_host, int* counts_host, cudaStream_t s, int slot) {
  int rc;
  if (c->det_mask && conf_thres != c->det_conf_now) {
    // the threshold the Detect epilogues filter with lives in device memory (the graph is captured once)
    CU_CHECK(c, cudaStreamSynchronize(s));  // det_conf_stage may still be the source of an earlier pending copy
    c->det_conf_stage = conf_thres;
    CU_CHECK(c, cudaMemcpyAsync(c->det_conf_dev, &c->det_conf_stage, sizeof(float), cudaMemcpyHostToDevice, s));
    c->det_conf_now = conf_thres;
  }
  if ((rc = aq_forward(c, s))) return rc;
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][2], s));
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
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][3], s));
  c->stage_timed[slot] = true;
  CU_CHECK(c, cudaMemcpyAsync(counts_host, c->nms_counts, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s));
  CU_CHECK(c, cudaMemcpyAsync(dets_host, c->nms_dets, (size_t)n * max_det * 6 * sizeof(float), cudaMemcpyDeviceToHost, s));
  CU_CHECK(c, cudaEventRecord(c->ev_done[slot], s));
  return slot;  // ticket for aq_detect_wait
}

// PreParams for window `win0 ..` of a device-resident raster (checks the grid; see aux_kernels.cuh)
int window_params(aq_ctx* c, const uint8_t* raster_dev, int raster_h, int raster_w, int win_h, int win_w, int win0, int n,
                  int bgr, PreParams* out) {
  if (!c->planned) return fail(c, AQ_ERR_STATE, "windows before plan_end");
  if (!raster_dev || win_h <= 0 || win_w <= 0 || raster_h < win_h || raster_w < win_w)
    return fail(c, AQ_ERR_ARG, "windows: raster %dx%d, window %dx%d", raster_h, raster_w, win_h, win_w);
  const int wins_x = raster_w / win_w, wins_y = raster_h / win_h;  // full windows only: the ragged right / bottom
  if (n <= 0 || n > c->batch || win0 < 0 || win0 + n > wins_x * wins_y)  // strips go through the plain tile path
    return fail(c, AQ_ERR_ARG, "windows: [%d, %d) outside the %d x %d grid of full windows (plan batch %d)", win0, win0 + n, wins_x, wins_y, c->batch);
  int rc = setup_preprocess(c, win_h, win_w);
  if (rc) return rc;
  PreParams p = c->pre;
  p.src = raster_dev; p.n = n; p.bgr = bgr;
  p.row_pitch = (size_t)raster_w * 3; p.windowed = 1; p.wins_y = wins_y; p.win0 = win0;
  p.out_s2d = nullptr; p.out_u8 = nullptr; p.win_list = nullptr;
  *out = p;
  return AQ_OK;
}
}  // namespace

extern "C" {

int aq_detect_windows(aq_ctx* c, const uint8_t* raster_dev, int raster_h, int raster_w, int win_h, int win_w, int win0, int n,
                      int bgr, float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes,
                      int n_classes, float* dets_host, int* counts_host, void* stream) {
  if (!c || !dets_host || !counts_host) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  PreParams p;
  int rc = window_params(c, raster_dev, raster_h, raster_w, win_h, win_w, win0, n, bgr, &p);
  if (rc) return rc;
  p.out_s2d = c->bufs[0].ptr;
  const int slot = c->pipe_slot;
  c->pipe_slot ^= 1;
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][0], s));
  c->stage_timed[slot] = false;
  CU_CHECK(c, launch_preprocess(p, s));
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][1], s));
  rc = forward_nms_d2h(c, n, conf_thres, iou_thres, max_det, agnostic, classes, n_classes, dets_host, counts_host, s, slot);
  if (rc < 0) return rc;
  CU_CHECK(c, cudaStreamSynchronize(s));
  return slot;  // >= 0: the ticket aq_detect_stage_ms takes
}

int aq_detect_window_list(aq_ctx* c, const uint8_t* raster_dev, int raster_h, int raster_w, int win_h, int win_w, const int* win_idx,
                          int n, int bgr, float conf_thres, double iou_thres, int max_det, int agnostic, const int* classes,
                          int n_classes, float* dets_host, int* counts_host, void* stream) {
  if (!c || !dets_host || !counts_host || !win_idx) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  PreParams p;
  int rc = window_params(c, raster_dev, raster_h, raster_w, win_h, win_w, 0, 1, bgr, &p);
  if (rc) return rc;
  const int n_wins = (raster_w / win_w) * (raster_h / win_h);
  if (n <= 0 || n > c->batch) return fail(c, AQ_ERR_ARG, "window list: %d windows (plan batch %d)", n, c->batch);
  for (int i = 0; i < n; ++i)
    if (win_idx[i] < 0 || win_idx[i] >= n_wins) return fail(c, AQ_ERR_ARG, "window list: index %d outside the %d full windows", win_idx[i], n_wins);
  if (c->win_list_cap < n) {
    if (c->win_list_dev) cudaFree(c->win_list_dev);
    c->win_list_dev = nullptr;
    c->win_list_cap = 0;
    CU_CHECK(c, cudaMalloc(&c->win_list_dev, (size_t)c->batch * sizeof(int)));
    c->win_list_cap = c->batch;
  }
  CU_CHECK(c, cudaMemcpyAsync(c->win_list_dev, win_idx, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
  p.n = n;
  p.win_list = c->win_list_dev;
  p.out_s2d = c->bufs[0].ptr;
  const int slot = c->pipe_slot;
  c->pipe_slot ^= 1;
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][0], s));
  c->stage_timed[slot] = false;
  CU_CHECK(c, launch_preprocess(p, s));
  CU_CHECK(c, cudaEventRecord(c->ev_stage[slot][1], s));
  rc = forward_nms_d2h(c, n, conf_thres, iou_thres, max_det, agnostic, classes, n_classes, dets_host, counts_host, s, slot);
  if (rc < 0) return rc;
  CU_CHECK(c, cudaStreamSynchronize(s));
  return slot;
}

int aq_preprocess_windows_u8(aq_ctx* c, const uint8_t* raster_dev, int raster_h, int raster_w, int win_h, int win_w, int win0,
                             int n, int bgr, uint8_t* out_dev, void* stream) {
  if (!c || !out_dev) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  PreParams p;
  int rc = window_params(c, raster_dev, raster_h, raster_w, win_h, win_w, win0, n, bgr, &p);
  if (rc) return rc;
  p.out_u8 = out_dev;
  CU_CHECK(c, launch_preprocess_u8(p, s));
  return AQ_OK;
}

int aq_tile_stats(aq_ctx* c, const uint8_t* src_dev, int n, int in_h, int in_w, int bgr, int raster_h, int raster_w, int win0,
                  int* stats_host, void* stream) {
  if (!c || !src_dev || n <= 0 || in_h <= 0 || in_w <= 0) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  PreParams p{};
  p.src = src_dev; p.n = n; p.bgr = bgr; p.in_h = in_h; p.in_w = in_w;
  p.row_pitch = (size_t)in_w * 3; p.wins_y = 1;
  if (raster_h > 0) {  // windows of a raster
    const int wins_x = raster_w / in_w, wins_y = raster_h / in_h;
    if (raster_h < in_h || raster_w < in_w || win0 < 0 || win0 + n > wins_x * wins_y)
      return fail(c, AQ_ERR_ARG, "tile_stats: windows [%d, %d) outside the %d x %d grid", win0, win0 + n, wins_x, wins_y);
    p.row_pitch = (size_t)raster_w * 3; p.windowed = 1; p.wins_y = wins_y; p.win0 = win0;
  }
  const size_t need = (size_t)n * in_w * sizeof(unsigned int) + (size_t)n * sizeof(TileStats);
  if (need > c->stats_bytes) {
    if (c->stats_scratch) cudaFree(c->stats_scratch);
    c->stats_scratch = nullptr;
    CU_CHECK(c, cudaMalloc(&c->stats_scratch, need));
    c->stats_bytes = need;
  }
  unsigned int* col = (unsigned int*)c->stats_scratch;
  TileStats* st = (TileStats*)(col + (size_t)n * in_w);
  cudaError_t e = launch_tile_stats(p, col, st, s);
  if (e != cudaSuccess) return fail(c, e == cudaErrorInvalidValue ? AQ_ERR_UNSUPPORTED : AQ_ERR_CUDA, "tile_stats: %s (tiles up to 2048 x 4096)", cudaGetErrorString(e));
  if (!stats_host) return AQ_OK;  // kernels only (timing): nothing is copied, the stream is not synchronised
  CU_CHECK(c, cudaMemcpyAsync(stats_host, st, (size_t)n * sizeof(TileStats), cudaMemcpyDeviceToHost, s));
  CU_CHECK(c, cudaStreamSynchronize(s));
  return AQ_OK;
}

// aq_tile_stats without any allocation: column sums in the library's static __device__ array, results written to the
// caller's DEVICE buffer, nothing copied, nothing synchronised.  No counterpart in the reference (added
// for callers that cannot call cudaMalloc: graph capture, tight serving loops).
int aq_tile_stats_static(aq_ctx* c, const uint8_t* src_dev, int n, int in_h, int in_w, int bgr, int raster_h, int raster_w, int win0,
                         int* stats_dev, void* stream) {
  if (!c || !src_dev || !stats_dev || n <= 0 || in_h <= 0 || in_w <= 0) return AQ_ERR_ARG;
  if ((long long)n * in_w > (long long)kTileStatsStaticCols)
    return fail(c, AQ_ERR_ARG, "tile_stats_static: n * in_w = %d * %d = %lld columns exceed the static scratch of %d (AQ_TILE_STATS_STATIC_MAX_COLS); use aq_tile_stats",
                n, in_w, (long long)n * in_w, kTileStatsStaticCols);
  if (c->stats_static_err != cudaSuccess || !c->stats_static)
    return fail(c, AQ_ERR_CUDA, "tile_stats_static: the static scratch could not be resolved at aq_create: %s", cudaGetErrorString(c->stats_static_err));
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  PreParams p{};
  p.src = src_dev; p.n = n; p.bgr = bgr; p.in_h = in_h; p.in_w = in_w;
  p.row_pitch = (size_t)in_w * 3; p.wins_y = 1;
  if (raster_h > 0) {  // windows of a raster
    const int wins_x = raster_w / in_w, wins_y = raster_h / in_h;
    if (raster_h < in_h || raster_w < in_w || win0 < 0 || win0 + n > wins_x * wins_y)
      return fail(c, AQ_ERR_ARG, "tile_stats_static: windows [%d, %d) outside the %d x %d grid", win0, win0 + n, wins_x, wins_y);
    p.row_pitch = (size_t)raster_w * 3; p.windowed = 1; p.wins_y = wins_y; p.win0 = win0;
  }
  // launch_tile_stats clears the n * in_w column sums it uses (cudaMemsetAsync on `stream`) and initialises stats_dev
  cudaError_t e = launch_tile_stats(p, c->stats_static, reinterpret_cast<TileStats*>(stats_dev), s);
  if (e != cudaSuccess) return fail(c, e == cudaErrorInvalidValue ? AQ_ERR_UNSUPPORTED : AQ_ERR_CUDA, "tile_stats_static: %s (tiles up to 2048 x 4096)", cudaGetErrorString(e));
  return AQ_OK;
}

int aq_jpeg_reconstruct(aq_ctx* c, const int16_t* coef, const uint16_t* quant, int n, int width, int height, int ncomp, int hs,
                        int bgr_out, int src_is_device, uint8_t* out_dev, void* stream) {
  if (!c || !coef || !quant || !out_dev || n <= 0 || width <= 0 || height <= 0) return AQ_ERR_ARG;
  if ((ncomp != 1 && ncomp != 3) || (hs != 1 && hs != 2) || (ncomp == 1 && hs != 1))
    return fail(c, AQ_ERR_UNSUPPORTED, "jpeg: %d components with %dx%d luma sampling (grayscale, 4:4:4 and 4:2:0 only)", ncomp, hs, hs);
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  JpegParams p{};
  p.n = n; p.width = width; p.height = height; p.ncomp = ncomp; p.hs = hs; p.bgr = bgr_out; p.out = out_dev;
  jpeg_geometry(p);
  const size_t planes = (size_t)n * p.planes_per_image;
  const size_t coef_bytes = (size_t)n * p.blocks_per_image * 64 * sizeof(int16_t), quant_bytes = (size_t)n * 192 * sizeof(uint16_t);
  const size_t planes_al = (planes + 255) / 256 * 256, coef_al = (coef_bytes + 255) / 256 * 256;
  const size_t need = planes_al + (src_is_device ? 0 : coef_al + quant_bytes);
  if (need > c->jpeg_bytes) {
    if (c->jpeg_scratch) { cudaStreamSynchronize(s); cudaFree(c->jpeg_scratch); }
    c->jpeg_scratch = nullptr;
    c->jpeg_bytes = 0;
    CU_CHECK(c, cudaMalloc(&c->jpeg_scratch, need));
    c->jpeg_bytes = need;
  }
  p.planes = (uint8_t*)c->jpeg_scratch;
  if (src_is_device) {
    p.coef = coef;
    p.quant = quant;
  } else {
    int16_t* cd = (int16_t*)(p.planes + planes_al);
    uint16_t* qd = (uint16_t*)(p.planes + planes_al + coef_al);
    CU_CHECK(c, cudaMemcpyAsync(cd, coef, coef_bytes, cudaMemcpyHostToDevice, s));
    CU_CHECK(c, cudaMemcpyAsync(qd, quant, quant_bytes, cudaMemcpyHostToDevice, s));
    p.coef = cd;
    p.quant = qd;
  }
  cudaError_t e = launch_jpeg_reconstruct(p, s);
  if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "jpeg_reconstruct: %s", cudaGetErrorString(e));
  return AQ_OK;
}

int aq_jpeg_reconstruct_packed(aq_ctx* c, const uint32_t* block_desc, const uint32_t* entries, const uint64_t* tile_entry_base,
                               size_t total_entries, const uint16_t* quant, int n, int width, int height, int ncomp, int hs,
                               int bgr_out, uint8_t* out_dev, void* stream) {
  if (!c || !block_desc || !entries || !tile_entry_base || !quant || !out_dev || n <= 0 || width <= 0 || height <= 0) return AQ_ERR_ARG;
  if ((ncomp != 1 && ncomp != 3) || (hs != 1 && hs != 2) || (ncomp == 1 && hs != 1))
    return fail(c, AQ_ERR_UNSUPPORTED, "jpeg: %d components with %dx%d luma sampling (grayscale, 4:4:4 and 4:2:0 only)", ncomp, hs, hs);
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  JpegParams p{};
  p.n = n; p.width = width; p.height = height; p.ncomp = ncomp; p.hs = hs; p.bgr = bgr_out; p.out = out_dev;
  jpeg_geometry(p);
  auto al = [](size_t b) { return (b + 255) / 256 * 256; };
  const size_t desc_b = (size_t)n * p.blocks_per_image * sizeof(uint32_t), ent_b = total_entries * sizeof(uint32_t);
  const size_t base_b = (size_t)n * sizeof(uint64_t), quant_b = (size_t)n * 192 * sizeof(uint16_t);
  const size_t need = al(desc_b) + al(ent_b) + al(base_b) + al(quant_b);
  const size_t planes = (size_t)n * p.planes_per_image;
  if (planes > c->jpeg_bytes) {
    if (c->jpeg_scratch) { CU_CHECK(c, cudaDeviceSynchronize()); cudaFree(c->jpeg_scratch); }
    c->jpeg_scratch = nullptr;
    c->jpeg_bytes = 0;
    CU_CHECK(c, cudaMalloc(&c->jpeg_scratch, planes));
    c->jpeg_bytes = planes;
  }
  const int slot = c->jpk_slot;
  c->jpk_slot ^= 1;
  if (!c->ev_jpk_h2d[slot]) {
    CU_CHECK(c, cudaEventCreateWithFlags(&c->ev_jpk_h2d[slot], cudaEventDisableTiming));
    CU_CHECK(c, cudaEventCreateWithFlags(&c->ev_jpk_used[slot], cudaEventDisableTiming));
  }
  if (need > c->jpk_bytes[slot]) {
    if (c->jpk_stage[slot]) { CU_CHECK(c, cudaDeviceSynchronize()); cudaFree(c->jpk_stage[slot]); }
    c->jpk_stage[slot] = nullptr;
    c->jpk_bytes[slot] = 0;
    const size_t grow = need + need / 4;  // the packed size varies from batch to batch: leave headroom
    CU_CHECK(c, cudaMalloc(&c->jpk_stage[slot], grow));
    c->jpk_bytes[slot] = grow;
    c->jpk_recorded[slot] = false;
  }
  uint8_t* st = (uint8_t*)c->jpk_stage[slot];
  uint32_t* d_desc = (uint32_t*)st;
  uint32_t* d_ent = (uint32_t*)(st + al(desc_b));
  unsigned long long* d_base = (unsigned long long*)(st + al(desc_b) + al(ent_b));
  uint16_t* d_quant = (uint16_t*)(st + al(desc_b) + al(ent_b) + al(base_b));
  if (c->jpk_recorded[slot]) CU_CHECK(c, cudaStreamWaitEvent(c->copy_stream, c->ev_jpk_used[slot], 0));
  CU_CHECK(c, cudaMemcpyAsync(d_desc, block_desc, desc_b, cudaMemcpyHostToDevice, c->copy_stream));
  if (ent_b) CU_CHECK(c, cudaMemcpyAsync(d_ent, entries, ent_b, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaMemcpyAsync(d_base, tile_entry_base, base_b, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaMemcpyAsync(d_quant, quant, quant_b, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaEventRecord(c->ev_jpk_h2d[slot], c->copy_stream));
  CU_CHECK(c, cudaStreamWaitEvent(s, c->ev_jpk_h2d[slot], 0));
  p.planes = (uint8_t*)c->jpeg_scratch;
  p.coef = nullptr;
  p.block_desc = d_desc; p.entries = d_ent; p.tile_entry_base = d_base; p.quant = d_quant;
  cudaError_t e = launch_jpeg_reconstruct(p, s);
  if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "jpeg_reconstruct_packed: %s", cudaGetErrorString(e));
  CU_CHECK(c, cudaEventRecord(c->ev_jpk_used[slot], s));
  c->jpk_recorded[slot] = true;
  return AQ_OK;
}

int aq_jpeg_decode_device(aq_ctx* c, const uint8_t* scans, const uint64_t* scan_off, const uint32_t* scan_bytes, size_t total_scan_bytes,
                          const uint8_t* tables, const uint16_t* quant, int n, int width, int height, int ncomp, int hs, int bgr_out,
                          uint8_t* out_dev, int* status_host, void* stream) {
  if (!c || !scans || !scan_off || !scan_bytes || !tables || !quant || !out_dev || n <= 0 || width <= 0 || height <= 0 ||
      (total_scan_bytes & 3))
    return AQ_ERR_ARG;
  if ((ncomp != 1 && ncomp != 3) || (hs != 1 && hs != 2) || (ncomp == 1 && hs != 1))
    return fail(c, AQ_ERR_UNSUPPORTED, "jpeg: %d components with %dx%d luma sampling (grayscale, 4:4:4 and 4:2:0 only)", ncomp, hs, hs);
  cudaSetDevice(c->device);
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  static const int sub_bits = [] { const char* e = getenv("AQUA_JPEG_SUB_BITS"); const int v = e ? atoi(e) : 1024; return v >= 32 && v % 32 == 0 ? v : 1024; }();
  JpegParams p{};
  p.n = n; p.width = width; p.height = height; p.ncomp = ncomp; p.hs = hs; p.bgr = bgr_out; p.out = out_dev;
  jpeg_geometry(p);
  HuffParams h{};
  h.geo.hs = hs; h.geo.vs = hs; h.geo.ncomp = ncomp; h.geo.bpm = ncomp == 1 ? 1 : hs * hs + 2;
  h.geo.mcus_x = p.chroma_bw; h.geo.mcus_y = (height + 8 * hs - 1) / (8 * hs);
  h.geo.luma_bw = p.luma_bw; h.geo.luma_blocks = p.luma_blocks; h.geo.chroma_blocks = p.chroma_blocks;
  h.geo.total_blocks = p.blocks_per_image;
  h.sub_bits = sub_bits;
  int nsub_max = 1;
  for (int i = 0; i < n; ++i) {
    if ((scan_off[i] & 3) || scan_off[i] + ((size_t)scan_bytes[i] + 3) / 4 * 4 + 8 > total_scan_bytes || scan_bytes[i] >= (1u << 28))
      return fail(c, AQ_ERR_ARG, "jpeg_decode_device: tile %d: segment [%llu, +%u) not inside the %zu staged bytes (whole words + 8 bytes of zero padding are read)",
                  i, (unsigned long long)scan_off[i], scan_bytes[i], total_scan_bytes);
    const int ns = (int)(((size_t)scan_bytes[i] * 8 + sub_bits - 1) / sub_bits);
    nsub_max = ns > nsub_max ? ns : nsub_max;
  }
  auto al = [](size_t b) { return (b + 255) / 256 * 256; };
  // scratch on the compute side: planes | dense coefficients | decoder states
  const size_t planes_b = al((size_t)n * p.planes_per_image), coef_b = al((size_t)n * p.blocks_per_image * 64 * sizeof(int16_t));
  const size_t dc_b = al((size_t)n * p.blocks_per_image * sizeof(int));
  const size_t state_b = al(jpeg_huff_state_bytes(n, nsub_max));
  if (planes_b + coef_b + dc_b + state_b > c->jpeg_bytes) {
    if (c->jpeg_scratch) { CU_CHECK(c, cudaDeviceSynchronize()); cudaFree(c->jpeg_scratch); }
    c->jpeg_scratch = nullptr;
    c->jpeg_bytes = 0;
    const size_t grow = planes_b + coef_b + dc_b + state_b + state_b / 2;
    CU_CHECK(c, cudaMalloc(&c->jpeg_scratch, grow));
    c->jpeg_bytes = grow;
  }
  // staging on the copy side
  const size_t off_b = (size_t)n * sizeof(uint64_t), len_b = (size_t)n * sizeof(uint32_t), tab_b = (size_t)n * AQ_JPEG_TABLE_BYTES;
  const size_t quant_b = (size_t)n * 192 * sizeof(uint16_t);
  const size_t need = al(total_scan_bytes) + al(off_b) + al(len_b) + al(tab_b) + al(quant_b);
  const int slot = c->jpk_slot;
  c->jpk_slot ^= 1;
  if (!c->ev_jpk_h2d[slot]) {
    CU_CHECK(c, cudaEventCreateWithFlags(&c->ev_jpk_h2d[slot], cudaEventDisableTiming));
    CU_CHECK(c, cudaEventCreateWithFlags(&c->ev_jpk_used[slot], cudaEventDisableTiming));
  }
  if (need > c->jpk_bytes[slot]) {
    if (c->jpk_stage[slot]) { CU_CHECK(c, cudaDeviceSynchronize()); cudaFree(c->jpk_stage[slot]); }
    c->jpk_stage[slot] = nullptr;
    c->jpk_bytes[slot] = 0;
    const size_t grow = need + need / 4;
    CU_CHECK(c, cudaMalloc(&c->jpk_stage[slot], grow));
    c->jpk_bytes[slot] = grow;
    c->jpk_recorded[slot] = false;
  }
  uint8_t* st = (uint8_t*)c->jpk_stage[slot];
  uint8_t* d_scans = st;
  unsigned long long* d_off = (unsigned long long*)(st + al(total_scan_bytes));
  uint32_t* d_len = (uint32_t*)((uint8_t*)d_off + al(off_b));
  uint8_t* d_tab = (uint8_t*)d_len + al(len_b);
  uint16_t* d_quant = (uint16_t*)(d_tab + al(tab_b));
  if (c->jpk_recorded[slot]) CU_CHECK(c, cudaStreamWaitEvent(c->copy_stream, c->ev_jpk_used[slot], 0));
  CU_CHECK(c, cudaMemcpyAsync(d_scans, scans, total_scan_bytes, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaMemcpyAsync(d_off, scan_off, off_b, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaMemcpyAsync(d_len, scan_bytes, len_b, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaMemcpyAsync(d_tab, tables, tab_b, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaMemcpyAsync(d_quant, quant, quant_b, cudaMemcpyHostToDevice, c->copy_stream));
  CU_CHECK(c, cudaEventRecord(c->ev_jpk_h2d[slot], c->copy_stream));
  CU_CHECK(c, cudaStreamWaitEvent(s, c->ev_jpk_h2d[slot], 0));
  uint8_t* scratch = (uint8_t*)c->jpeg_scratch;
  h.scans = d_scans; h.scan_off = d_off; h.scan_bytes = d_len; h.tables = d_tab;
  h.coef = (int16_t*)(scratch + planes_b);
  h.dc = (int*)(scratch + planes_b + coef_b);
  void* state = scratch + planes_b + coef_b + dc_b;
  cudaError_t e = launch_jpeg_huffman(h, state, n, nsub_max, s);
  if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "jpeg_decode_device (entropy decoding): %s", cudaGetErrorString(e));
  p.planes = scratch;
  p.coef = h.coef;
  p.dc = h.dc; p.dc_bpm = h.geo.bpm; p.dc_mcus_x = h.geo.mcus_x;
  p.quant = d_quant;
  e = launch_jpeg_reconstruct(p, s);
  if (e != cudaSuccess) return fail(c, AQ_ERR_CUDA, "jpeg_decode_device (reconstruction): %s", cudaGetErrorString(e));
  if (status_host)
    CU_CHECK(c, cudaMemcpyAsync(status_host, jpeg_huff_status_ptr(state, n, nsub_max), (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s));
  CU_CHECK(c, cudaEventRecord(c->ev_jpk_used[slot], s));
  c->jpk_recorded[slot] = true;
  return AQ_OK;
}

int aq_detect_stage_ms(aq_ctx* c, int ticket, float ms[3]) {
  if (!c || !ms || ticket < 0 || ticket > 1) return AQ_ERR_ARG;
  if (!c->stage_timed[ticket]) return fail(c, AQ_ERR_STATE, "detect_stage_ms: slot %d holds no batch submitted through aq_detect_submit", ticket);
  cudaSetDevice(c->device);
  CU_CHECK(c, cudaEventSynchronize(c->ev_stage[ticket][3]));
  for (int k = 0; k < 3; ++k) CU_CHECK(c, cudaEventElapsedTime(&ms[k], c->ev_stage[ticket][k], c->ev_stage[ticket][k + 1]));
  return AQ_OK;
}

int aq_detect_wait(aq_ctx* c, int ticket) {
  if (!c || ticket < 0 || ticket > 1) return AQ_ERR_ARG;
  cudaSetDevice(c->device);
  CU_CHECK(c, cudaEventSynchronize(c->ev_done[ticket]));
  return AQ_OK;
}

int aq_detect(aq_ctx* c, const uint8_t* tiles, int n, int in_h, int in_w, int bgr, int src_is_device, float conf_thres,
              double iou_thres, int max_det, int agnostic, const int* classes, int n_classes, float* dets_host,
              int* counts_host, void* stream) {
  int rc = aq_detect_submit(c, tiles, n, in_h, in_w, bgr, src_is_device, conf_thres, iou_thres, max_det, agnostic,
                            classes, n_classes, dets_host, counts_host, stream);
  if (rc < 0) return rc;
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
