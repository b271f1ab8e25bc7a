// Implicit-GEMM convolution on tcgen05 tensor cores (sm_100a), operands fed by TMA.
//
// Replaces upstream models/common.py::Conv.forward_fuse (+ Bottleneck residual) and, with EPI_DETECT,
// models/yolo.py::Detect.forward [UPSTREAM-RECALLED; the reference reaches them through
// `yolov5/detect.py`, REF README.md:77].
//
//   D[M = pixels, N = c_out] = sum over taps (kh,kw) and 64-channel chunks of  A_tap[M, 64] * W_tap[N, 64]^T
//
// * activations NHWC fp16.  Every TMA row is one pixel x 64 channels = 128 bytes (SWIZZLE_128B); channel
//   counts that are not multiples of 64 are completed by TMA out-of-bounds zero fill (loads) / clipping
//   (stores), so no buffer is padded in HBM.  Weights are stored [c_out][tap][chunk][64] with zero padding.
// * an "M block" is 128 output pixels = one UMMA M=128 accumulator; a CTA tile is 1 or 2 M blocks that share
//   every weight tile (halves the weight traffic L2 -> SM per output).  Three ways to feed A:
//     FLAT  1x1 convs: 128 consecutive pixels per M block, one box load per 64-channel chunk.
//     HALO  3x3/s1 convs: the CTA loads the (8*mx+2) x (16*my+2) pixel halo patch ONCE per chunk; the nine
//           taps are nine UMMA descriptors into that patch (start address shifted by whole 128-byte rows,
//           stride-byte-offset = patch row pitch; the swizzle is a function of the absolute smem address, so
//           any 128-byte-multiple offset is legal -- measured, profiles/r1_hw_probe_v1.txt T1).
//     HALO2 3x3/s2 convs: the input is viewed as four parity planes (y-parity, x-parity) at half resolution; the
//           CTA loads each plane's (8*mx+1) x (16*my+1) patch once per chunk (one ring slot per plane) and the
//           nine taps are shifted UMMA views of the planes (4 / 2 / 2 / 1 taps per plane) -- 2.1 patches' worth
//           of bytes per chunk instead of nine boxes.  Same tall output coordinates and stores as HALO.
//     TAP   3x3/s2 convs and small maps: one shifted box per (tap, chunk); stride 2 uses a parity view
//           (c, W/2, y-parity, H/2, N) per x-parity so the strided gather is still a dense box.
// * warp roles (512 threads): warps 0..3 = A producer (HALO: 128 threads of 16-byte cp.async with zero fill --
//   a multi-line TMA box costs ~0.2-0.4 us per line; FLAT/TAP: one thread issuing TMA boxes), warp 4 = weight
//   producer (TMA), warp 5 = TMEM allocator + single-thread MMA issuer, warp 6 = TMA store issuer,
//   warps 8..15 = epilogue (TMEM -> regs -> bias + SiLU -> fp16 -> swizzled smem staging).  Separate warps
//   because one thread sustains only one cp.async.bulk.tensor per ~0.2-0.4 us (measured, T4): loads, weight
//   loads and stores must not queue behind each other.
// * HALO tiles live in a "tall" coordinate system: all images stacked with one shared zero row between them,
//   so small maps (20x20, 40x40) fill 16-row M blocks without per-image waste; stores are clipped per image.
// * the Bottleneck residual (dst += f(conv)) is a TMA reduce-add of the staged tile into the destination
//   (the residual operand is the destination slice itself), so it costs no load.
// * weights of small layers stay resident in shared memory for the whole (persistent) kernel.
// * accumulators fp32 in TMEM, double-buffered when 2 * mblk * N <= 512 columns.
// * persistent CTAs (grid = min(tiles, SMs)), static round-robin tile order with N tiles adjacent; launched
//   with programmatic dependent launch so the prologue (barrier init, TMEM alloc, weight prefetch) overlaps
//   the previous kernel's tail.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace aq {

enum { CONV_MODE_FLAT = 0, CONV_MODE_HALO = 1, CONV_MODE_TAP = 2, CONV_MODE_HALO2 = 3 };
enum { EPI_STORE = 0, EPI_DETECT = 1, EPI_CHAIN = 2 };
enum { ACT_NONE = 0, ACT_SILU = 1, ACT_SILU_FAST = 2 };

constexpr int kConvThreads = 672;  // 4 role warps + 16 epilogue warps + 1 store warp
constexpr int kConvThreadsChain = 704;  // EPI_CHAIN: + 1 warp issuing the chained 1x1 conv's MMAs
constexpr int kMaxCout2 = 256;      // EPI_CHAIN: output channels of the chained conv (one UMMA N)
constexpr int kEpiWarps = 16;
constexpr int kMaxA = 6;    // A ring slots
constexpr int kMaxB = 20;   // weight ring slots (resident layers: one per K step)
constexpr int kTileM = 128;
constexpr int kMaxCout = 768;
constexpr int kTraceRoles = 6, kTraceTiles = 48, kTraceVals = 4;  // roles: A producer, B producer, MMA0, epilogue warp 0, its parts, store warp
constexpr int kConvMaxDynSmem = 232448 - 6144;  // 227 KB opt-in limit minus static smem (barriers, bias)

struct alignas(64) ConvKParams {
  CUtensorMap tmA;    // activation loads (TAP stride 2: x-parity 0)
  CUtensorMap tmA2;   // TAP stride 2: x-parity 1
  CUtensorMap tmB;    // weight loads
  CUtensorMap tmOut;  // output stores / reduce-adds (EPI_STORE)
  CUtensorMap tmH[6];  // HALO patch loads: boxes (64 ch, pw, h rows) with h = 32 / 16 / 8 / 4 / 2 / 1
  CUtensorMap tmH2[6];  // HALO2 (3x3/s2 parity planes): tmH = x-parity 0, tmH2 = x-parity 1, 5-D (c, W/2, y-parity, H/2, N)
  CUtensorMap tmUp;     // TAP 1x1 with a folded nearest-2x upsample: (c, dx, W/2, dy, H/2 * N), dx / dy strides 0
  CUtensorMap tmOutR[4];  // HALO: the same with boxes of 8 / 4 / 2 / 1 rows (blocks that start on a zero row)
  const float* bias;  // [n_tiles_n * tile_n], zero padded (pre-multiplied by 0.5 for ACT_SILU_FAST)
  int mode, stride;
  int c_in;                // input channels of the conv (valid K columns per tap)
  int cpt, taps, n_steps;  // 64-channel chunks per tap, taps, K steps per tile (= taps * cpt)
  int a_group;             // K steps that share one A slot (HALO: 9, else 1)
  int mblk, mx, my;        // M blocks per tile; HALO arrangement (mx side by side, my stacked)
  int pw, ph, pw_inv;      // HALO patch size in pixels, ceil(65536 / pw)
  const __half* a_src;     // HALO: source slice (first channel), geometry of the source buffer
  int a_n, a_h, a_w, a_ct, a_cin;
  int na, nb, b_res, n_stg, acc_stages;
  int b_dense;             // HALO, c_in <= 16, resident: weight slot = tap / 4, K offset = (tap % 4) * 16 (see api.cu Cand)
  int n_mma;               // MMA-issuing threads (2 when the tile has more than one M block)
  int tile_n, n_tiles_n, groups;  // UMMA N, N tiles, 64-channel store groups per tile
  int m_tiles;                    // tiles along M
  // Thread-block clusters of 1 << cshift CTAs (1, 2 or 4): the CTAs of a cluster work on consecutive M tiles of the
  // same N tile in lockstep and share every weight tile -- each loads 1 / cluster of its rows and multicasts them to
  // all, which divides the weight traffic L2 -> SM by the cluster size.  total_tiles = ceil(m_tiles / cluster) *
  // n_tiles_n cluster-tiles; M tiles past m_tiles are computed on zero-filled operands and clipped by the stores.
  int cshift, total_tiles;
  int pair;  // 1: clusters of two CTAs running tcgen05.mma.cta_group::2 (M = 256 per pair; each CTA keeps half of every weight tile)
  // Tile order: 0 = ascending, 1 = descending (tile t of the persistent loop works on tile total - 1 - t).  The plan
  // alternates it from op to op ("serpentine"): a layer then starts on the tiles its producer wrote LAST, which still
  // sit in the 126 MB L2, instead of on the ones written first, which have long been evicted to HBM.
  int reverse;
  int up_chunks, up_hs;  // leading 64-channel chunks that come from tmUp; rows of the half-resolution source
  int tw, th, tiles_x, tiles_y;   // HALO: tile origin grid (tile = 8mx x 16my); TAP: M block = tw x th
  int act, has_res;
  int epi_sets;  // 1: sixteen epilogue warps in lockstep; 2: two sets of eight taking alternate groups (conv_igemm.cu)
  int dbg;  // AQUA_DBG attribution experiments (results wrong): 1 = no TMA stores issued, 2 = epilogue skips math + staging, 4 = one store per tile only
  uint32_t a_slot_bytes, b_slot_bytes, stg_bytes;
  uint32_t a_tx_bytes, b_tx_bytes;  // bytes one slot fill delivers
  uint32_t tmem_cols, acc_stride, acc_pitch;  // TMEM columns: allocation, per accumulator stage, per M block
  // EPI_DETECT
  float* pred;
  int det_no, det_rows, det_level_off, det_ny, det_nx, det_m_total;
  float det_stride;
  float det_anchor[6];
  // Early candidate filter for NMS (nc = 3 only): one bit per prediction row of the image, set iff obj > *det_conf and
  // max_c(obj * cls_c) > *det_conf -- upstream non_max_suppression's candidate test, evaluated on the very fp32 values
  // this epilogue stores.  det_mask = (images, det_mask_words) zeroed before the forward, or NULL.
  uint32_t* det_mask;
  const float* det_conf;
  int det_mask_words;
  // EPI_CHAIN: a 1x1 / stride-1 conv consuming this conv's output, evaluated in the same kernel.  The 128 x 64 fp16
  // groups the epilogue stages for the TMA store are valid SWIZZLE_128B K-major A operands, so a second tcgen05.mma
  // reads them out of shared memory before (while) the store leaves: D2[128, tile_n2] = sum over the first chain_kg groups
  // of stage[g] * W2[g]^T, accumulated in its own TMEM columns, then bias2 + activation -> staging -> tmOut2.  The
  // consumer's values are exactly those of a separate launch (it reads the same fp16-rounded activations); what
  // disappears is one launch boundary and one read of the tensor from HBM / L2.
  CUtensorMap tmB2;        // chained weights [c_out2][chain_kg * 64] (zero padded), box {64, tile_n2}
  CUtensorMap tmOut2;      // chained output stores (same geometry as tmOut)
  CUtensorMap tmOut2R[4];  // HALO: boxes of 8 / 4 / 2 / 1 rows
  const float* bias2;      // [tile_n2] (pre-multiplied by 0.5 for ACT_SILU_FAST)
  int chain;               // 0 = off
  int chain_kg;            // 64-channel groups of THIS conv's output that feed the chained conv (its K = c_in2)
  int chain_kk_last;       // K = 16 MMAs that carry data in the last of them
  int chain_store1;        // 1: this conv's own output is stored too (0: only the chained conv's output leaves)
  int tile_n2, groups2, act2;
  uint32_t b2_slot_bytes;  // one resident weight slot of the chained conv (tile_n2 rows x 128 B, 1024-aligned)
  uint32_t acc2_base, acc2_pitch;  // TMEM columns: first column of the chained accumulators (acc2_stages stages, acc2_pitch apart)
  int acc2_stages;                 // 1 or 2 (block b uses stage b % acc2_stages)
  unsigned long long* trace;  // device buffer [grid][kTraceRoles][kTraceTiles][kTraceVals] or NULL
};

size_t conv_smem_bytes(const ConvKParams& p);
cudaError_t launch_conv(const ConvKParams& p, int epi, int num_sms, int pdl, cudaStream_t stream);
cudaError_t conv_kernels_init();  // sets max dynamic smem attributes

}  // namespace aq
