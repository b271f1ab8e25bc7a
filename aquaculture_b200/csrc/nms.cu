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
// Phases per CTA (1024 threads): A ordered compaction of the mask's set bits (block scan) -> B stable LSD radix sort of
// (~score bits, position), 8 bits per pass, passes with a single occupied bin skipped -> C greedy
// suppression in blocks of 1024 candidates against the <= max_det kept boxes held in shared memory,
// warp-ballot resolution inside each 32-candidate group, early exit at max_det keeps.
#include <cuda_runtime.h>
#include <stdint.h>

#include "nms.cuh"

namespace aq {
namespace {

constexpr int kT = 1024;
constexpr int kWarps = kT / 32;

struct Box {
  float x1, y1, x2, y2;
};

__device__ __forceinline__ bool iou_gt(const Box& a, float area_a, const Box& b, float area_b, float thr) {
  const float xx1 = fmaxf(a.x1, b.x1);
  const float yy1 = fmaxf(a.y1, b.y1);
  const float xx2 = fminf(a.x2, b.x2);
  const float yy2 = fminf(a.y2, b.y2);
  const float w = fmaxf(0.0f, __fsub_rn(xx2, xx1));
  const float h = fmaxf(0.0f, __fsub_rn(yy2, yy1));
  const float inter = __fmul_rn(w, h);
  const float ovr = __fdiv_rn(inter, __fsub_rn(__fadd_rn(area_a, area_b), inter));
  return ovr > thr;  // thr = largest fp32 <= iou_thres  <=>  (double)ovr > iou_thres
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
  __shared__ __align__(16) uint16_t s_whist[kWarps][256];
  __shared__ uint32_t s_chunk_base[256];
  __shared__ int s_skip;
  __shared__ int s_kept;
  extern __shared__ float4 s_dyn[];  // kept boxes (class-offset) [max_det] then areas [max_det]
  float4* s_kbox = s_dyn;
  float4* s_cbox = s_dyn + p.max_det;
  float* s_karea = reinterpret_cast<float*>(s_cbox + kT);
  float* s_cscore = s_karea + p.max_det;
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
  if (tid == 0) s_total = 0;
  __syncthreads();
  for (int base = 0; base < words; base += kT) {
    const int w = base + tid;
    uint32_t m = w < words ? mask[w] : 0u;
    const int cnt = __popc(m);
    int inc = cnt;
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += y;
    }
    if (lane == 31) s_warp_cnt[warp] = inc;
    __syncthreads();
    int pos = s_total + inc - cnt;
    for (int k = 0; k < warp; ++k) pos += s_warp_cnt[k];
    while (m) {
      const int bit = __ffs(m) - 1;
      m &= m - 1;
      const int r = (w << 5) + bit;
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
        keyA[pos] = ~__float_as_uint(bestv);
        valA[pos] = (uint32_t)pos;
      }
      ++pos;
    }
    __syncthreads();
    if (tid == 0) {
      int t = s_total;
      for (int k = 0; k < kWarps; ++k) t += s_warp_cnt[k];
      s_total = t;
    }
    __syncthreads();
  }
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
  } else
  for (int shift = 0; shift < 32; shift += 8) {
    if (tid < 256) s_bins[tid] = 0;
    if (tid == 0) s_skip = 0;
    __syncthreads();
    for (int i = tid; i < n_all; i += kT) atomicAdd(&s_bins[(ksrc[i] >> shift) & 255u], 1u);
    __syncthreads();
    if (tid < 256 && s_bins[tid] == (uint32_t)n_all) s_skip = 1;
    __syncthreads();
    if (s_skip) continue;  // every key has the same digit: the pass is the identity
    if (warp == 0) {
      // exclusive scan of the 256 bins, 8 per lane -> running output position of each digit
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
    for (int base = 0; base < n_all; base += kT) {
      // zero the per-warp digit counts (32 x 256 x 2 B = 16 KB = one 16-byte store per thread)
      reinterpret_cast<uint4*>(&s_whist[0][0])[tid] = make_uint4(0, 0, 0, 0);
      __syncthreads();
      const int i = base + tid;
      const bool valid = i < n_all;
      uint32_t key = 0, val = 0, d = 0;
      if (valid) {
        key = ksrc[i];
        val = vsrc[i];
        d = (key >> shift) & 255u;
      }
      const uint32_t tag = valid ? d : (256u + (uint32_t)lane);
      const unsigned peers = __match_any_sync(0xffffffffu, tag);
      const int rank = __popc(peers & ((1u << lane) - 1u));
      if (valid && rank == 0) s_whist[warp][d] = (uint16_t)__popc(peers);
      __syncthreads();
      if (tid < 256) {
        const uint32_t b0 = s_bins[tid];
        uint32_t run = 0;
        for (int w = 0; w < kWarps; ++w) {
          const uint32_t c = s_whist[w][tid];
          s_whist[w][tid] = (uint16_t)run;  // elements of this digit in earlier warps of the chunk
          run += c;
        }
        s_chunk_base[tid] = b0;
        s_bins[tid] = b0 + run;
      }
      __syncthreads();
      if (valid) {
        const uint32_t dst = s_chunk_base[d] + s_whist[warp][d] + (uint32_t)rank;
        kdst[dst] = key;
        vdst[dst] = val;
      }
      __syncthreads();
    }
    uint32_t* t = ksrc; ksrc = kdst; kdst = t;
    t = vsrc; vsrc = vdst; vdst = t;
  }
  const int n = n_all < p.max_nms ? n_all : p.max_nms;

  NMS_STAMP(3);
  // ------------------------------------------------------------------ C: greedy suppression
  if (tid == 0) s_kept = 0;
  __syncthreads();
  float* dets = p.dets + (size_t)img * p.max_det * 6;
  int* keep_rows = p.keep_rows ? p.keep_rows + (size_t)img * p.max_det : nullptr;
  const float thr = p.iou_thr_f32;
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
    for (int w = 0; w < kWarps; ++w) {
      __syncthreads();
      const int kc = s_kept;
      if (kc >= p.max_det) break;  // uniform
      if (base + w * 32 >= n) break;  // uniform: no candidates left in this block
      if (warp >= w) {
        // catch up on boxes kept since this thread last looked
        for (int k = checked; k < kc && alive; ++k) {
          const float4 kb = s_kbox[k];
          const Box kbx = {kb.x, kb.y, kb.z, kb.w};
          if (iou_gt(kbx, s_karea[k], ob, area, thr)) alive = false;
        }
        checked = kc;
      }
      if (warp == w) {
        unsigned rem = __ballot_sync(0xffffffffu, alive);
        unsigned keptmask = 0;
        while (rem) {
          const int l = __ffs(rem) - 1;
          keptmask |= 1u << l;
          Box lb;
          lb.x1 = __shfl_sync(0xffffffffu, ob.x1, l);
          lb.y1 = __shfl_sync(0xffffffffu, ob.y1, l);
          lb.x2 = __shfl_sync(0xffffffffu, ob.x2, l);
          lb.y2 = __shfl_sync(0xffffffffu, ob.y2, l);
          const float la = __shfl_sync(0xffffffffu, area, l);
          if (alive && lane > l && iou_gt(lb, la, ob, area, thr)) alive = false;
          rem = __ballot_sync(0xffffffffu, alive && lane > l);
        }
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
  const size_t dyn = (size_t)p.max_det * (sizeof(float4) + sizeof(float)) + (size_t)kT * (sizeof(float4) + 4 * sizeof(float));
  nms_kernel<<<batch, kT, dyn, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace aq
