// Batched non-max suppression: candidate test on all SMs, then one CTA per image.
//
// Replaces upstream utils/general.py::non_max_suppression (+ xywh2xyxy) and the torchvision.ops.nms core it
// calls [UPSTREAM-RECALLED; reached from `yolov5/detect.py`, REF README.md:77].  Semantics restated in
// SURVEY.md §8 a-8 / App. C1 and pinned by torchvision's CPU kernel in tests/:
//   1. keep rows with obj > conf;  cls_conf *= obj;  xyxy = (cx -/+ w/2, cy -/+ h/2)
//   2. conf, j = max over classes (first max on ties); keep conf > conf; optional class filter
//   3. stable sort by conf descending (ties: lower source row first); keep first max_nms
//   4. boxes + j*max_wh (fp32) unless agnostic; greedy: suppress iff (double)iou > iou_thres with
//      iou = inter / (area_i + area_j - inter) in fp32, inter = max(0,dx)*max(0,dy)
//   5. first max_det keeps, rows [x1,y1,x2,y2,conf,cls]
// All arithmetic that decides keep indices uses explicit round-to-nearest intrinsics so nvcc cannot
// contract it into FMAs: the keep list is bit-exact w.r.t. the CPU reference.
//
// The candidate test (steps 1-2) runs before this kernel, over all SMs: either nms_mask_kernel (one pass over pred) or
// the Detect epilogues of the conv stack themselves, which set one bit per passing row (conv_igemm.cu) -- this kernel
// then reads 32 rows per mask word and touches only the rows that pass (~0.2 % of 25 200 on sparse tiles).
// Phases per CTA (1024 threads): A ordered compaction of the mask's set bits (block scan over the mask words, then one thread
// per row: coalesced) -> B stable sort by score: a rank sort in shared memory for <= 1024 candidates, else an LSD radix sort of
// (score key, position), 8 bits per pass, each warp walking its own contiguous segment (three block barriers per pass), passes
// with a single occupied bin skipped -> C greedy suppression in blocks of 1024 candidates against the <= max_det kept boxes
// held in shared memory: warps catch up on the boxes kept since they last looked (four tests side by side, the exact division
// only inside a 2e-6 band around the threshold), each warp in turn resolves its own 32 candidates with bit arithmetic on a
// 32 x 32 suppression relation, warps that cannot be reached before max_det wait, and a later block's first warps share their
// long catch-up with the whole CTA.  BASELINE.json configs[3] (~20 k candidates per image): 1.09 -> 0.35 ms per 64 images;
// sparse tiles 45 -> 19 us (profiles/r2_nms_stress.txt).
#include <cuda_runtime.h>
#include <stdint.h>

#include "nms.cuh"

namespace aq {
namespace {

constexpr int kT = 1024;
constexpr int kWarps = kT / 32;

// Lanes holding the same 8-bit digit (valid lanes only): eight ballots instead of match.any, which is several times slower
__device__ __forceinline__ unsigned match_digit(uint32_t d, bool valid) {
  unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    const bool bit = (d >> b) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

// Sort key, ascending = score descending: 0x7F800000 - bits for non-negative scores (conf 0.25 .. 1 share the top byte, whose
// pass is then skipped), the raw bits for negative ones (after every positive, more negative later)
__device__ __forceinline__ uint32_t score_key(float v) {
  const uint32_t b = __float_as_uint(v);
  return b <= 0x7F800000u ? 0x7F800000u - b : b;
}

__host__ __device__ constexpr size_t nms_kept_region_bytes(int max_det) {
  const size_t kept = ((size_t)max_det * (sizeof(float4) + sizeof(float)) + 15) / 16 * 16;
  return kept > (size_t)kWarps * 256 * 4 ? kept : (size_t)kWarps * 256 * 4;
}

struct Box {
  float x1, y1, x2, y2;
};

// IoU test of the greedy scan: RN(inter / uni) > thr with thr = the largest fp32 <= iou_thres, i.e. (double)ovr > iou_thres.
// RN is monotone and within 2^-24 relative of the true quotient, so a pair with inter < (thr * 0.999998) * uni (two more
// roundings of 2^-24 each) is certainly not above the threshold: almost every pair leaves here, without the division.
// The argument needs normal numbers: a tiny or non-positive union or a tiny threshold take the exact path, and so do NaNs
// (every comparison with them is false).  `thr_lo` = thr * 0.999998f, or a negative number to switch the shortcut off.
__device__ __forceinline__ bool iou_gt(const Box& a, float area_a, const Box& b, float area_b, float thr, float thr_lo) {
  const float w = fmaxf(0.0f, __fsub_rn(fminf(a.x2, b.x2), fmaxf(a.x1, b.x1)));
  const float h = fmaxf(0.0f, __fsub_rn(fminf(a.y2, b.y2), fmaxf(a.y1, b.y1)));
  const float inter = __fmul_rn(w, h);
  const float uni = __fsub_rn(__fadd_rn(area_a, area_b), inter);
  if (inter < __fmul_rn(thr_lo, uni) && uni > 1.0e-30f) return false;
  return __fdiv_rn(inter, uni) > thr;
}

// Does any kept box k0 <= k < k1 suppress the candidate?  Four boxes per round: their intersections and unions are computed
// side by side (one test is a chain of ~10 dependent instructions) and one branch leaves when all four fail the quick test of
// iou_gt; otherwise the four exact quotients decide.
__device__ __forceinline__ bool any_kept_suppresses(const float4* s_kbox, const float* s_karea, int k0, int k1, const Box& ob, float area,
                                                    float thr, float thr_lo) {
  int k = k0;
  for (; k + 4 <= k1; k += 4) {
    float inter[4], uni[4];
    bool quick = true;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float4 kb = s_kbox[k + u];
      const float w = fmaxf(0.0f, __fsub_rn(fminf(kb.z, ob.x2), fmaxf(kb.x, ob.x1)));
      const float h = fmaxf(0.0f, __fsub_rn(fminf(kb.w, ob.y2), fmaxf(kb.y, ob.y1)));
      inter[u] = __fmul_rn(w, h);
      uni[u] = __fsub_rn(__fadd_rn(s_karea[k + u], area), inter[u]);
      quick &= (inter[u] < __fmul_rn(thr_lo, uni[u])) & (uni[u] > 1.0e-30f);
    }
    if (quick) continue;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (__fdiv_rn(inter[u], uni[u]) > thr) return true;
  }
  for (; k < k1; ++k) {
    const float4 kb = s_kbox[k];
    const Box kbx = {kb.x, kb.y, kb.z, kb.w};
    if (iou_gt(kbx, s_karea[k], ob, area, thr, thr_lo)) return true;
  }
  return false;
}

// Candidate test of every row, 256 rows per CTA over the whole grid: warp w of CTA (x, img) writes mask word 8 x + w.
__global__ void __launch_bounds__(256) nms_mask_kernel(NmsParams p, uint32_t* mask_out) {
  const int img = blockIdx.y;
  const int r = blockIdx.x * 256 + threadIdx.x;
  const int nc = p.no - 5;
  bool pass = false;
  if (r < p.rows) {
    const float* row = p.pred + ((size_t)img * p.rows + r) * p.no;
    const float obj = row[4];
    if (obj > p.conf_thres) {
      int best = 0;
      float bestv = __fmul_rn(row[5], obj);
      for (int c = 1; c < nc; ++c) {
        const float v = __fmul_rn(row[5 + c], obj);
        if (v > bestv) {
          bestv = v;
          best = c;
        }
      }
      pass = bestv > p.conf_thres;
      if (pass && p.n_classes > 0) {
        pass = false;
        for (int k = 0; k < p.n_classes; ++k) pass |= (p.classes[k] == best);
      }
    }
  }
  const unsigned bal = __ballot_sync(0xffffffffu, pass);
  const int words = (p.rows + 31) >> 5;
  const int w = r >> 5;
  if ((threadIdx.x & 31) == 0 && w < words) mask_out[(size_t)img * words + w] = bal;
}

__global__ void __launch_bounds__(kT, 1) nms_kernel(NmsParams p) {
  const int img = blockIdx.x;
  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;

  const float* pred = p.pred + (size_t)img * p.rows * p.no;
  const size_t sbase = (size_t)img * p.rows;
  float4* cbox = p.cand_box + sbase;
  float* cscore = p.cand_score + sbase;
  float* ccls = p.cand_cls + sbase;
  int* crow = p.cand_row + sbase;
  uint32_t* keyA = p.keys + 2 * sbase;
  uint32_t* keyB = keyA + p.rows;
  uint32_t* valA = p.vals + 2 * sbase;
  uint32_t* valB = valA + p.rows;

  __shared__ int s_warp_cnt[kWarps];
  __shared__ int s_total;
  // small images (<= kT candidates, the usual case: ~40 per 640^2 tile at conf 0.25): candidates, sort and suppression stay in
  // shared memory -- no global scratch round trips, a rank sort instead of four radix passes
  // (carved out of the dynamic shared memory behind the kept boxes: s_cbox, s_cscore, s_ccls, s_crow, s_order)
  __shared__ uint32_t s_bins[256];
  __shared__ int s_skip;
  __shared__ uint32_t s_rmask[32];
  __shared__ float4 s_rbox[32];
  __shared__ float s_rarea[32];
  __shared__ int s_kept;
  // dynamic shared memory: [kept boxes (class-offset) [max_det] | their areas [max_det]] padded to 32 KB, then the small path's
  // candidate arrays.  The kept region is idle until phase C, so the radix sort's per-warp digit counters (and phase A's word
  // positions) live in it: s_whist[kWarps][256].
  extern __shared__ float4 s_dyn[];
  float4* s_kbox = s_dyn;
  float* s_karea = reinterpret_cast<float*>(s_kbox + p.max_det);
  uint32_t (*s_whist)[256] = reinterpret_cast<uint32_t (*)[256]>(s_dyn);
  float4* s_cbox = s_dyn + nms_kept_region_bytes(p.max_det) / sizeof(float4);
  float* s_cscore = reinterpret_cast<float*>(s_cbox + kT);
  float* s_ccls = s_cscore + kT;
  int* s_crow = reinterpret_cast<int*>(s_ccls + kT);
  uint32_t* s_order = reinterpret_cast<uint32_t*>(s_crow + kT);

#define NMS_STAMP(i) do { if (p.dbg_clk && img == 0 && tid == 0) p.dbg_clk[i] = clock64(); } while (0)
  NMS_STAMP(0);
  // ------------------------------------------------------------------ count first: which path?
  const int nc = p.no - 5;
  const int words = (p.rows + 31) >> 5;
  const uint32_t* mask = p.cand_mask + (size_t)img * words;
  {
    int cnt = 0;
    for (int w = tid; w < words; w += kT) cnt += __popc(mask[w]);
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) s_warp_cnt[warp] = cnt;
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int k = 0; k < kWarps; ++k) t += s_warp_cnt[k];
      s_total = t;
    }
    __syncthreads();
  }
  const bool small = s_total <= kT;
  __syncthreads();
  if (small) {
    cbox = s_cbox;
    cscore = s_cscore;
    ccls = s_ccls;
    crow = s_crow;
  }
  NMS_STAMP(1);
  // ------------------------------------------------------------------ A: ordered compaction of the passing rows
  // thread t takes mask word base + t (rows 32 (base + t) ..): exclusive block scan of the popcounts gives the
  // position of its first candidate, so candidates stay in source-row order (upstream's stable sort breaks score
  // ties by it)
  // pass 1: exclusive scan of the mask words' popcounts -> position of each word's first candidate (s_wpos, carved
  // from the radix sort's per-warp histograms, which are idle until phase B; images with more than 8192 words --
  // 262 144 rows -- keep the positions in the sort's global scratch)
  uint32_t* s_wpos = words <= kWarps * 256 ? &s_whist[0][0] : keyB;
  if (tid == 0) s_total = 0;
  __syncthreads();
  for (int base = 0; base < words; base += kT) {
    const int w = base + tid;
    const int cnt = w < words ? __popc(mask[w]) : 0;
    int inc = cnt;
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += y;
    }
    if (lane == 31) s_warp_cnt[warp] = inc;
    __syncthreads();
    int pos = s_total + inc - cnt;
    for (int k = 0; k < warp; ++k) pos += s_warp_cnt[k];
    if (w < words) s_wpos[w] = (uint32_t)pos;
    __syncthreads();
    if (tid == 0) {
      int t = s_total;
      for (int k = 0; k < kWarps; ++k) t += s_warp_cnt[k];
      s_total = t;
    }
    __syncthreads();
  }
  // pass 2: one thread per row (a warp reads 32 consecutive rows: coalesced), position = word base + set bits below
  for (int r0 = warp * 32; r0 < p.rows; r0 += kT) {
    const uint32_t m = mask[r0 >> 5];
    if (!((m >> lane) & 1u)) continue;
    const int r = r0 + lane;
    const int pos = (int)s_wpos[r0 >> 5] + __popc(m & ((1u << lane) - 1u));
    const float* row = pred + (size_t)r * p.no;
    const float obj = row[4];
    int best = 0;
    float bestv = __fmul_rn(row[5], obj);
    for (int c = 1; c < nc; ++c) {
      const float v = __fmul_rn(row[5 + c], obj);
      if (v > bestv) {
        bestv = v;
        best = c;
      }
    }
    const float cx = row[0], cy = row[1];
    const float hw = __fdiv_rn(row[2], 2.0f), hh = __fdiv_rn(row[3], 2.0f);
    cbox[pos] = make_float4(__fsub_rn(cx, hw), __fsub_rn(cy, hh), __fadd_rn(cx, hw), __fadd_rn(cy, hh));
    cscore[pos] = bestv;
    ccls[pos] = (float)best;
    crow[pos] = r;
    if (!small) {
      keyA[pos] = score_key(bestv);
      valA[pos] = (uint32_t)pos;
    }
  }
  __syncthreads();
  NMS_STAMP(2);
  const int n_all = s_total;
  if (n_all == 0) {
    if (tid == 0) p.counts[img] = 0;
    return;
  }

  // ------------------------------------------------------------------ B: stable sort by descending score
  uint32_t* ksrc = keyA;
  uint32_t* kdst = keyB;
  uint32_t* vsrc = valA;
  uint32_t* vdst = valB;
  if (small) {
    // rank sort: candidate i goes to position #{j : score_j > score_i, or equal and j < i} (stable, as upstream's argsort is made)
    __syncthreads();
    if (tid < n_all) {
      const float si = s_cscore[tid];
      int rank = 0;
      for (int j = 0; j < n_all; ++j) {
        const float sj = s_cscore[j];
        rank += (sj > si || (sj == si && j < tid)) ? 1 : 0;
      }
      s_order[rank] = (uint32_t)tid;
    }
    __syncthreads();
    vsrc = s_order;
  } else {
    // LSD radix sort, 8 bits per pass.  Warp w owns the contiguous segment [w * seg, (w + 1) * seg) of the keys and walks it in
    // order, 32 keys per step: per-warp digit histograms (pass 1), one block scan over (digit, warp) giving every warp its
    // first output position per digit, then the same walk scatters (pass 2) -- stable, and three block barriers per digit
    // instead of four per 1024 keys.
    const int seg = (n_all + kT - 1) / kT * 32;  // keys per warp
    const int w0 = warp * seg;
    const int w1 = w0 + seg < n_all ? w0 + seg : n_all;
    for (int shift = 0; shift < 32; shift += 8) {
      for (int i = tid; i < kWarps * 256 / 4; i += kT) reinterpret_cast<uint4*>(&s_whist[0][0])[i] = make_uint4(0, 0, 0, 0);
      if (tid == 0) s_skip = 0;
      __syncthreads();
      // (four steps' keys are loaded together: the walk is a chain of L2 round trips otherwise)
      for (int i0 = w0; i0 < w1; i0 += 128) {
        uint32_t kk[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) kk[u] = i0 + u * 32 + lane < w1 ? ksrc[i0 + u * 32 + lane] : 0u;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const bool valid = i0 + u * 32 + lane < w1;
          const uint32_t d = (kk[u] >> shift) & 255u;
          const unsigned peers = match_digit(d, valid);
          if (valid && (peers & ((1u << lane) - 1u)) == 0) s_whist[warp][d] += (uint32_t)__popc(peers);  // one lane per digit
          __syncwarp();
        }
      }
      __syncthreads();
      if (shift == 0) NMS_STAMP(8);
      if (tid < 256) {
        uint32_t run = 0;
        for (int w = 0; w < kWarps; ++w) {
          const uint32_t c = s_whist[w][tid];
          s_whist[w][tid] = run;  // keys of this digit in earlier warps' segments
          run += c;
        }
        s_bins[tid] = run;
        if (run == (uint32_t)n_all) s_skip = 1;
      }
      __syncthreads();
      if (shift == 0) NMS_STAMP(9);
      if (s_skip) {  // every key has the same digit: the pass is the identity
        __syncthreads();  // (s_skip is reset at the top of the next pass)
        continue;
      }
      if (warp == 0) {
        // exclusive scan of the 256 digit totals, 8 per lane -> first output position of each digit
        uint32_t loc[8], sum = 0;
        for (int k = 0; k < 8; ++k) {
          loc[k] = s_bins[lane * 8 + k];
          sum += loc[k];
        }
        uint32_t inc = sum;
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += y;
        }
        uint32_t run = inc - sum;
        for (int k = 0; k < 8; ++k) {
          s_bins[lane * 8 + k] = run;
          run += loc[k];
        }
      }
      __syncthreads();
      if (shift == 0) NMS_STAMP(10);
      for (int i0 = w0; i0 < w1; i0 += 128) {
        uint32_t kk[4], vv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const bool valid = i0 + u * 32 + lane < w1;
          kk[u] = valid ? ksrc[i0 + u * 32 + lane] : 0u;
          vv[u] = valid ? vsrc[i0 + u * 32 + lane] : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const bool valid = i0 + u * 32 + lane < w1;
          const uint32_t d = (kk[u] >> shift) & 255u;
          const unsigned peers = match_digit(d, valid);
          const int rank = __popc(peers & ((1u << lane) - 1u));
          if (valid) {
            const uint32_t dst = s_bins[d] + s_whist[warp][d] + (uint32_t)rank;
            kdst[dst] = kk[u];
            vdst[dst] = vv[u];
          }
          __syncwarp();
          if (valid && rank == 0) s_whist[warp][d] += (uint32_t)__popc(peers);
          __syncwarp();
        }
      }
      __syncthreads();  // the scattered keys are read back by other warps in the next pass
      if (shift == 0) NMS_STAMP(11);
      uint32_t* t = ksrc; ksrc = kdst; kdst = t;
      t = vsrc; vsrc = vdst; vdst = t;
    }
  }
  const int n = n_all < p.max_nms ? n_all : p.max_nms;

  NMS_STAMP(3);
  // ------------------------------------------------------------------ C: greedy suppression
  if (tid == 0) s_kept = 0;
  __syncthreads();
  float* dets = p.dets + (size_t)img * p.max_det * 6;
  int* keep_rows = p.keep_rows ? p.keep_rows + (size_t)img * p.max_det : nullptr;
  const float thr = p.iou_thr_f32;
  const float thr_lo = thr >= 0.01f ? __fmul_rn(thr, 0.999998f) : -1.0f;
  const int dbg_mode = p.dbg_mode;  // timing experiments (AQUA_NMS_CLK=2: no catch-up tests, 3: no tests inside the warp): wrong results
  for (int base = 0; base < n; base += kT) {
    if (s_kept >= p.max_det) break;  // uniform: read after a barrier
    const int i = base + tid;
    bool alive = i < n;
    Box ob = {0, 0, 0, 0}, raw = {0, 0, 0, 0};
    float area = 0.f, score = 0.f, cls = 0.f;
    int src_row = 0;
    if (alive) {
      const uint32_t pos = vsrc[i];
      const float4 bb = cbox[pos];
      raw = {bb.x, bb.y, bb.z, bb.w};
      score = cscore[pos];
      cls = ccls[pos];
      src_row = crow[pos];
      const float off = p.agnostic ? 0.0f : __fmul_rn(cls, p.max_wh);
      ob = {__fadd_rn(raw.x1, off), __fadd_rn(raw.y1, off), __fadd_rn(raw.x2, off), __fadd_rn(raw.y2, off)};
      area = __fmul_rn(__fsub_rn(ob.x2, ob.x1), __fsub_rn(ob.y2, ob.y1));
    }
    int checked = 0;
    if (base > 0) {
      // A later block starts with a long kept list, and (see the window below) only its first warps are likely to be reached:
      // all 1024 threads share those warps' tests against the list -- candidate t % C against slice t / C of it -- instead
      // of leaving each of the few warps a serial walk over the whole list.  (The small path's arrays park the candidates.)
      const int kc0 = s_kept;
      const int need0 = (p.max_det - kc0 + 31) >> 5;
      const int wc = need0 * 2 > 4 ? need0 * 2 : 4;  // = the window of step 0
      if (wc < kWarps / 2 && kc0 > 0) {
        const int C = wc * 32, S = kT / C;
        const int L = (kc0 + S - 1) / S;
        int* s_flag = reinterpret_cast<int*>(s_ccls);
        if (tid < C) {
          s_cbox[tid] = make_float4(ob.x1, ob.y1, ob.x2, ob.y2);
          s_cscore[tid] = area;
          s_flag[tid] = 0;
        }
        __syncthreads();
        const int cnd = tid % C, sl = tid / C;
        if (sl < S && base + cnd < n) {
          const float4 cb = s_cbox[cnd];
          const Box cbx = {cb.x, cb.y, cb.z, cb.w};
          const int k0 = sl * L, k1 = k0 + L < kc0 ? k0 + L : kc0;
          if (any_kept_suppresses(s_kbox, s_karea, k0, k1, cbx, s_cscore[cnd], thr, thr_lo)) s_flag[cnd] = 1;
        }
        __syncthreads();
        if (tid < C) {
          alive = alive && s_flag[tid] == 0;
          checked = kc0;
        }
      }
      if (base == kT) NMS_STAMP(6);
    }
    for (int w = 0; w < kWarps; ++w) {
      __syncthreads();
      const int kc = s_kept;
      if (kc >= p.max_det) break;  // uniform
      if (base + w * 32 >= n) break;  // uniform: no candidates left in this block
      // Warps ahead of the resolving one test their candidates against the boxes kept since they last looked -- but only
      // the warps that can still be reached: with r keeps missing at least ceil(r / 32) more warps resolve, so twice that
      // many (>= 4) stay caught up and the rest wait (the block in which max_det is reached is mostly never looked at)
      const int need = (p.max_det - kc + 31) >> 5;
      const int window = need * 2 > 4 ? need * 2 : 4;
      if (warp >= w && warp < w + window) {
        // catch up on boxes kept since this thread last looked
        if (dbg_mode != 2) alive = alive && !any_kept_suppresses(s_kbox, s_karea, checked, kc, ob, area, thr, thr_lo);
        checked = kc;
      }
      if (warp == w) {
        // Greedy scan inside the warp's 32 candidates (score order = lane order).  Every lane first finds the set of
        // earlier lanes whose box would suppress its own -- tests against the 32 boxes parked in shared memory, four side by
        // side, no shuffles or votes in the loop -- then the scan itself is bit arithmetic on those 32 words: a candidate is
        // kept iff it is alive and no kept earlier lane is in its set.
        const unsigned alive0 = __ballot_sync(0xffffffffu, alive);
        s_rbox[lane] = make_float4(ob.x1, ob.y1, ob.x2, ob.y2);
        s_rarea[lane] = area;
        __syncwarp();
        unsigned sup_by = 0;
        if (alive && dbg_mode != 3) {
          for (int l = 0; l < lane; l += 4) {
            float inter[4], uni[4];
            bool quick = true;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float4 kb = s_rbox[l + u];
              const float w = fmaxf(0.0f, __fsub_rn(fminf(kb.z, ob.x2), fmaxf(kb.x, ob.x1)));
              const float h = fmaxf(0.0f, __fsub_rn(fminf(kb.w, ob.y2), fmaxf(kb.y, ob.y1)));
              inter[u] = __fmul_rn(w, h);
              uni[u] = __fsub_rn(__fadd_rn(s_rarea[l + u], area), inter[u]);
              // (the lane itself and later lanes do not matter: they never reach the exact test)
              quick &= ((inter[u] < __fmul_rn(thr_lo, uni[u])) & (uni[u] > 1.0e-30f)) | (l + u >= lane);
            }
            if (quick) continue;
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (l + u < lane && __fdiv_rn(inter[u], uni[u]) > thr) sup_by |= 1u << (l + u);
          }
        }
        s_rmask[lane] = sup_by;
        __syncwarp();
        unsigned keptmask = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (((alive0 >> j) & 1u) && !(s_rmask[j] & keptmask)) keptmask |= 1u << j;
        __syncwarp();
        if ((keptmask >> lane) & 1u) {
          const int slot = kc + __popc(keptmask & ((1u << lane) - 1u));
          if (slot < p.max_det) {
            s_kbox[slot] = make_float4(ob.x1, ob.y1, ob.x2, ob.y2);
            s_karea[slot] = area;
            float* o = dets + (size_t)slot * 6;
            o[0] = raw.x1; o[1] = raw.y1; o[2] = raw.x2; o[3] = raw.y2; o[4] = score; o[5] = cls;
            if (keep_rows) keep_rows[slot] = src_row;
          }
        }
        if (lane == 0) {
          const int nk = kc + __popc(keptmask);
          s_kept = nk < p.max_det ? nk : p.max_det;
        }
      }
    }
    __syncthreads();
    if (base == 0) NMS_STAMP(5);
  }
  __syncthreads();
  if (tid == 0) p.counts[img] = s_kept;
  NMS_STAMP(4);
}

}  // namespace

size_t nms_scratch_bytes(int batch, int rows) {
  return (size_t)batch * rows * (sizeof(float4) + 3 * 4 + 4 * 4) + (size_t)batch * nms_mask_words(rows) * 4;
}

cudaError_t nms_init() {
  return cudaFuncSetAttribute(nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

cudaError_t launch_nms_mask(const NmsParams& p, int batch, cudaStream_t stream) {
  nms_mask_kernel<<<dim3((p.rows + 255) / 256, batch), 256, 0, stream>>>(p, const_cast<uint32_t*>(p.cand_mask));
  return cudaGetLastError();
}

cudaError_t launch_nms(const NmsParams& p, int batch, cudaStream_t stream) {
  const size_t dyn = nms_kept_region_bytes(p.max_det) + (size_t)kT * (sizeof(float4) + 4 * sizeof(float));
  nms_kernel<<<batch, kT, dyn, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace aq
