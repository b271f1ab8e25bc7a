// See conv_igemm.cuh for the design.  sm_100a only (tcgen05 / TMEM / TMA).
#include "conv_igemm.cuh"
#include "ptx.cuh"

namespace aq {

using namespace ptx;

namespace {

// This file is compiled twice (see the Makefile): AQ_CLUSTER=0 (default) builds the kernels every planned layer
// uses; AQ_CLUSTER=1 builds the thread-block-cluster variants (weight tiles shared by TMA multicast) as a separate
// set of kernels.  A runtime `if (p.cshift)` in the MMA issue loop cost the streamed-weight 3x3 layers 5-8 %
// (A/B on one box, profiles/r1_ab_cluster_branch.txt), hence compile-time.
#ifndef AQ_CLUSTER
#define AQ_CLUSTER 0
#endif
// AQ_PAIR=1 (third build of this file): clusters of exactly two CTAs executing tcgen05.mma.cta_group::2 -- one M = 256 MMA per
// pair, each CTA holding its own 128 rows of A and HALF of every weight tile, so an SM ingests half the weight bytes per output
// (what bounds the streamed-weight layers: L2 -> SM operand delivery).  The leader (cluster rank 0) issues the
// MMAs; operand arrivals of both CTAs are counted on the leader's mbarriers, completions are multicast to both.
#ifndef AQ_PAIR
#define AQ_PAIR 0
#endif
#if AQ_PAIR
#undef AQ_CLUSTER
#define AQ_CLUSTER 1
#endif
#if AQ_CLUSTER
#define CSHIFT(p) ((p).cshift)
#else
#define CSHIFT(p) 0
#endif

// Operand loads / barrier traffic of a CTA, single or paired.  `bar` is always the LOCAL address of the barrier.
__device__ __forceinline__ uint32_t full_bar(uint32_t bar) { return AQ_PAIR ? mapa_rank(bar, 0) : bar; }
__device__ __forceinline__ void arm_full(uint32_t bar, uint32_t bytes) {  // expect the bytes of one slot fill (of both CTAs of a pair)
  if (AQ_PAIR) {
    if (cluster_ctarank() == 0) mbar_arrive_expect_tx(bar, 2u * bytes);
  } else {
    mbar_arrive_expect_tx(bar, bytes);
  }
}
__device__ __forceinline__ void ld2d(uint32_t dst, const void* tm, uint32_t bar, int c0, int c1) {
  if (AQ_PAIR) tma2_load_2d(dst, tm, full_bar(bar), c0, c1);
  else tma_load_2d(dst, tm, bar, c0, c1);
}
__device__ __forceinline__ void ld4d(uint32_t dst, const void* tm, uint32_t bar, int c0, int c1, int c2, int c3) {
  if (AQ_PAIR) tma2_load_4d(dst, tm, full_bar(bar), c0, c1, c2, c3);
  else tma_load_4d(dst, tm, bar, c0, c1, c2, c3);
}
__device__ __forceinline__ void ld5d(uint32_t dst, const void* tm, uint32_t bar, int c0, int c1, int c2, int c3, int c4) {
  if (AQ_PAIR) tma2_load_5d(dst, tm, full_bar(bar), c0, c1, c2, c3, c4);
  else tma_load_5d(dst, tm, bar, c0, c1, c2, c3, c4);
}
// MMA completion -> mbarrier (pair: the barrier at this offset in BOTH CTAs)
__device__ __forceinline__ void commit_to(uint32_t bar) {
  if (AQ_PAIR) umma2_commit_mc(bar, 3);
  else umma_commit(bar);
}
// an epilogue warp has drained its part of an accumulator stage (pair: the leader's barrier counts both CTAs' warps)
__device__ __forceinline__ void arrive_tempty(uint32_t bar) {
  if (AQ_PAIR && cluster_ctarank() != 0) mbar_arrive_cluster(mapa_rank(bar, 0));
  else mbar_arrive(bar);
}

struct MBlock {
  int n, y0, x0;  // spatial modes: image, first output row / column
  int m0;         // flat mode: first flat pixel
};

// Output coordinates of M block `mb` of M tile `m_tile`.
__device__ __forceinline__ MBlock mblock_coords(const ConvKParams& p, int m_tile, int mb) {
  MBlock r;
  r.n = r.y0 = r.x0 = r.m0 = 0;
  if (p.mode == CONV_MODE_FLAT) {
    r.m0 = (m_tile * p.mblk + mb) * kTileM;
  } else if (p.mode == CONV_MODE_HALO || p.mode == CONV_MODE_HALO2) {
    // y0 is a row of the "tall" image: all images stacked with one shared zero row between them,
    // tall row R = n * (H + 1) + y + 1 (rows n * (H + 1) are the zero rows)
    const int ty = m_tile / p.tiles_x;
    const int tx = m_tile - ty * p.tiles_x;
    const int jx = mb % p.mx, jy = mb / p.mx;
    r.x0 = (tx * p.mx + jx) * 8;
    r.y0 = (ty * p.my + jy) * 16;
  } else {
    const int idx = m_tile * p.mblk + mb;
    const int per_img = p.tiles_x * p.tiles_y;
    r.n = idx / per_img;  // >= batch for the padding block of an odd count: TMA zero-fills / clips it
    const int rem = idx - r.n * per_img;
    const int ty = rem / p.tiles_x;
    r.y0 = ty * p.th;
    r.x0 = (rem - ty * p.tiles_x) * p.tw;
  }
  return r;
}

// Weight chunk index (64 K elements each) of K step s: weights are [c_out][tap][chunk][64].
__device__ __forceinline__ int weight_chunk(const ConvKParams& p, int s) {
  if (p.mode != CONV_MODE_HALO && p.mode != CONV_MODE_HALO2) return s;  // FLAT: s = chunk; TAP: s = tap * cpt + chunk
  const int c = s / 9;                     // HALO: s = chunk * 9 + tap
  int tap = s - c * 9;
  // HALO2: taps in plane order (1,1): 0 2 6 8 | (1,0): 1 7 | (0,1): 3 5 | (0,0): 4
  if (p.mode == CONV_MODE_HALO2) tap = (int)((0x453718620ull >> (4 * tap)) & 15ull);
  return tap * p.cpt + c;
}

// Ring position advanced with compare/reset: `x % n` and `x / n` by a runtime n cost a ~100-clk dependent
// instruction chain each, which made every role's inner loop scalar-latency bound (measured, trace_s7b).
__device__ __forceinline__ void split_tile(const ConvKParams& p, int tile, int& m_tile, int& n_idx) {
  if (p.reverse) tile = p.total_tiles - 1 - tile;
  if (p.n_tiles_n == 1) {
    m_tile = tile;
    n_idx = 0;
  } else {
    m_tile = tile / p.n_tiles_n;
    n_idx = tile - m_tile * p.n_tiles_n;
  }
  if (CSHIFT(p)) m_tile = (m_tile << CSHIFT(p)) + (int)cluster_ctarank();  // CTA r of a cluster: M tile r of the group
}

// Persistent tile loop of a CTA: cluster-tiles tile_first(), + tile_step(), ... < p.total_tiles
__device__ __forceinline__ int tile_first(const ConvKParams& p) { return (int)(blockIdx.x >> CSHIFT(p)); }
__device__ __forceinline__ int tile_step(const ConvKParams& p) { return (int)(gridDim.x >> CSHIFT(p)); }

struct Ring {
  uint32_t slot = 0, phase = 0;
  __device__ __forceinline__ void advance(uint32_t n) {
    if (++slot == n) {
      slot = 0;
      phase ^= 1u;
    }
  }
};

constexpr uint32_t kWarpA = 0, kWarpB = 1, kWarpMma = 2, kWarpMma1 = 3, kWarpEpi0 = 4, kWarpStore = 20;
// named barriers: 1..4 = staging buffer b filled (16 epilogue warps arrive, the store warp waits),
// 5..8 = staging buffer b read by its TMA store (the store warp arrives, the epilogue warps wait); with two epilogue
// sets (epi_store_role2) set 1 waits on 9..12 instead -- see kBarStgEmptySet1 below
constexpr uint32_t kBarStgFull = 1, kBarStgEmpty = 5, kBarStgThreads = kEpiWarps * 32 + 32;
// A hardware named barrier counts arrivals, not who arrives.  With an ODD number of staging buffers a buffer is filled
// by the two sets alternately, so set 0 (waiting for group k's turn on buffer b) and set 1 (already waiting for group
// k + 3's turn on the same buffer) could sit on ONE barrier id at the same time; the store warp's single arrival then
// released whichever eight warps got there first -- warps of the wrong set overwrote a buffer its TMA store had not read
// yet, while the right set's late warps filled theirs after the store had gone (found by tools/repro_probe.py: ~1 % of
// forward passes, last tiles of the 768 -> 768 1x1 conv with three buffers; profiles/r2_race_hunt.txt).  Each set
// therefore owns its own "buffer read" barrier ids.
constexpr uint32_t kBarStgEmptySet1 = 9;

__device__ __forceinline__ float silu_exact(float v) { return __fdividef(v, 1.0f + __expf(-v)); }

// Per-role timeline (TRACE builds only): role r of CTA b writes kTraceVals clock64() values per tile at
// trace[((b * kTraceRoles + r) * kTraceTiles + it) * kTraceVals]; see tools/trace_op.py.
#define TR_NOW() (TRACE ? clock64() : 0ll)
#define TR_WAIT(acc, ...)                  \
  do {                                     \
    if (TRACE) {                           \
      const long long t_ = clock64();      \
      __VA_ARGS__;                         \
      acc += clock64() - t_;               \
    } else {                               \
      __VA_ARGS__;                         \
    }                                      \
  } while (0)

template <bool TRACE>
__device__ __forceinline__ void trace_put(const ConvKParams& p, int role, int it, long long v0, long long v1,
                                          long long v2, long long v3) {
  if (TRACE) {
    if (it < kTraceTiles) {
      long long* d = reinterpret_cast<long long*>(p.trace) +
                     (((size_t)blockIdx.x * kTraceRoles + role) * kTraceTiles + it) * kTraceVals;
      d[0] = v0; d[1] = v1; d[2] = v2; d[3] = v3;
    }
  }
}


// TMA store (or reduce-add, for the fused Bottleneck residual) of one staged 64-channel group
// [128 rows][128 B, swizzled] of the M block at `t`; `ch0` = first output channel of the group.
__device__ __forceinline__ void store_group_to(const ConvKParams& p, const CUtensorMap* tm_out, const CUtensorMap* tm_out_r, bool has_res,
                                               const MBlock& t, uint32_t src, int ch0);
__device__ __forceinline__ void store_group(const ConvKParams& p, const MBlock& t, uint32_t src, int ch0) {
  store_group_to(p, &p.tmOut, p.tmOutR, p.has_res != 0, t, src, ch0);
}
// The same through explicit tensor maps (EPI_CHAIN stores the chained conv's output through tmOut2 / tmOut2R).
__device__ __forceinline__ void store_group_to(const ConvKParams& p, const CUtensorMap* tm_out, const CUtensorMap* tm_out_r, bool has_res,
                                               const MBlock& t, uint32_t src, int ch0) {
  if (p.mode == CONV_MODE_FLAT) {
    if (has_res) tma_reduce_add_2d(tm_out, src, ch0, t.m0);
    else tma_store_2d(tm_out, src, ch0, t.m0);
  } else if (p.mode == CONV_MODE_HALO || p.mode == CONV_MODE_HALO2) {
    // the 16 tall rows of the block touch one or more images: one clipped box per image (rows outside
    // [0, H) -- the shared zero row and the other image's rows -- are dropped by TMA)
    const int hp1 = p.a_h + 1;
    const int n_first = t.y0 / hp1;
    int n_last = n_first;  // usually one or two images; maps shorter than 15 rows put more into a block
    while ((n_last + 1) * hp1 <= t.y0 + 15) ++n_last;
    if (n_last > p.a_n - 1) n_last = p.a_n - 1;
    for (int n = n_first; n <= n_last; ++n) {
      const int ys = t.y0 - (n * hp1 + 1);  // image row of staging row 0
      if (ys >= 0) {
        if (has_res) tma_reduce_add_4d(tm_out, src, ch0, t.x0, ys, n);
        else tma_store_4d(tm_out, src, ch0, t.x0, ys, n);
      } else {
        // a TMA store may not start at a negative coordinate (illegal instruction, probe T6):
        // write staging rows [-ys, 16) to image rows [0, 16 + ys) as boxes of 8 / 4 / 2 / 1 rows
        int s0 = -ys, y = 0;
        int len = 16 - s0;
        if (len > p.a_h) len = p.a_h;
#pragma unroll
        for (int hb = 0; hb < 4; ++hb) {
          const int h = 8 >> hb;
          if (len & h) {
            const uint32_t ssrc = src + (uint32_t)s0 * 1024u;
            if (has_res) tma_reduce_add_4d(&tm_out_r[hb], ssrc, ch0, t.x0, y, n);
            else tma_store_4d(&tm_out_r[hb], ssrc, ch0, t.x0, y, n);
            s0 += h;
            y += h;
          }
        }
      }
    }
  } else {
    if (has_res) tma_reduce_add_4d(tm_out, src, ch0, t.x0, t.y0, t.n);
    else tma_store_4d(tm_out, src, ch0, t.x0, t.y0, t.n);
  }
}

// Per-image segments of a halo patch in "tall" coordinates (images stacked with one zero row above each):
// rows R0 .. R0 + ph - 1 cut where an image ends.  Worked out ONCE per tile (one runtime division per segment;
// a scalar warp runs ~8 clk per dependent instruction) and reused by every chunk / parity plane of the tile.
constexpr int kMaxSeg = 4;
struct PatchSegs {
  int n[kMaxSeg], y[kMaxSeg], len[kMaxSeg];
  int count;  // > kMaxSeg: maps so short that the patch spans more images -- callers take the general loop
};
__device__ __forceinline__ PatchSegs patch_segments(int R0, int ph, int a_h) {
  PatchSegs sg;
  const int hp1 = a_h + 1;
  int R = R0, rem = ph, k = 0;
#pragma unroll
  for (int i = 0; i < kMaxSeg; ++i) {
    sg.n[i] = sg.y[i] = sg.len[i] = 0;
    if (rem > 0) {
      const int n = R / hp1;
      const int y = R - n * hp1 - 1;  // -1 = the zero row above image n
      int len = a_h - y;              // rows y .. H-1
      if (len > rem) len = rem;
      sg.n[i] = n; sg.y[i] = y; sg.len[i] = len;
      R += len;
      rem -= len;
      ++k;
    }
  }
  sg.count = rem > 0 ? kMaxSeg + 1 : k;
  return sg;
}
// Issue the boxes of one patch: ld(hb, dst, y, n) loads 32 >> hb rows of image n starting at row y.
template <class F>
__device__ __forceinline__ void patch_issue(const PatchSegs& sg, int R0, int ph, int a_h, uint32_t dst, uint32_t row_bytes, F&& ld) {
  if (sg.count <= kMaxSeg) {
#pragma unroll
    for (int i = 0; i < kMaxSeg; ++i) {
      if (i < sg.count) {
        int y = sg.y[i];
#pragma unroll
        for (int hb = 0; hb < 6; ++hb) {
          const int h = 32 >> hb;
          if (sg.len[i] & h) {
            ld(hb, dst, y, sg.n[i]);
            dst += (uint32_t)h * row_bytes;
            y += h;
          }
        }
      }
    }
    return;
  }
  const int hp1 = a_h + 1;
  int R = R0, rem = ph;
  while (rem > 0) {
    const int n = R / hp1;
    int y = R - n * hp1 - 1;
    int len = a_h - y;
    if (len > rem) len = rem;
    R += len;
    rem -= len;
#pragma unroll
    for (int hb = 0; hb < 6; ++hb) {
      const int h = 32 >> hb;
      if (len & h) {
        ld(hb, dst, y, n);
        dst += (uint32_t)h * row_bytes;
        y += h;
      }
    }
  }
}

// Barrier addresses inside the static `bars` array (see the kernel for the layout).
struct MmaCtx {
  uint32_t bar0, a_base, b_base, tmem_base;
  int mi, total_tiles;
};

// MMA issuer role, specialised at compile time so that the per-K-step path is a handful of instructions:
// the generic runtime-parameter loop needed ~130 dependent SASS instructions (~640 clk) per step, several
// times the 4 x 46..128 clk the tensor core needs (profiles/r1_trace_*.txt, r1_hw_probe T7).
//   HALO : the nine taps of a 3x3/s1 conv are shifted views of one patch slot (else: one A slot per step)
//   BRES : every weight tile is resident in shared memory (loaded once), no ring
//   NMY  : M blocks issued by this thread (block indices mi, mi + n_mma)
template <int FEED, bool BRES, int NMY, bool TRACE>
__device__ __forceinline__ void mma_role(const ConvKParams& p, const MmaCtx& mc) {
  constexpr bool HALO = FEED != 0;  // FEED: 0 = one A slot per K step, 1 = HALO (3x3/s1 patch), 2 = HALO2 (3x3/s2 planes)
  const uint32_t bar0 = mc.bar0;
  auto a_full = [&](uint32_t s) { return bar0 + 8u * s; };
  auto a_empty = [&](uint32_t s) { return bar0 + 8u * (kMaxA + s); };
  auto b_full = [&](uint32_t s) { return bar0 + 8u * (2 * kMaxA + s); };
  auto b_empty = [&](uint32_t s) { return bar0 + 8u * (2 * kMaxA + kMaxB + s); };
  auto tfull = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + a); };
  auto tempty = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + 2 + a); };
  // loop invariants in registers
  if (AQ_PAIR && cluster_ctarank() != 0) return;  // the leader issues for the pair
  const uint32_t idesc = make_idesc_f16(AQ_PAIR ? 2 * kTileM : kTileM, p.tile_n);
  const uint32_t pw8 = p.pw * 8u;
  // descriptor halves: lo = start>>4 | LBO(1)<<16 ; hi = SBO>>4 | version(1)<<14 | SWIZZLE_128B(2)<<29
  const uint32_t a_hi = ((HALO ? p.pw * 128u : 1024u) >> 4) | (1u << 14) | (2u << 29);
  const uint32_t b_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  const uint32_t lo_flags = 1u << 16;
  const uint32_t a_lo0 = ((mc.a_base & 0x3FFFFu) >> 4) | lo_flags, a_slot16 = p.a_slot_bytes >> 4;
  const uint32_t b_lo0 = ((mc.b_base & 0x3FFFFu) >> 4) | lo_flags, b_slot16 = p.b_slot_bytes >> 4;
  const uint32_t na = p.na, nb = p.nb, acc_stages = p.acc_stages, acc_stride = p.acc_stride;
  const int cpt = p.cpt, n_steps = p.n_steps;
  const bool b_dense = p.b_dense != 0;
  // K = 16 MMAs that carry data in the last 64-channel chunk (the rest of the chunk is zero padding)
  const int c_in = HALO ? p.a_cin : p.c_in;
  const int kk_last = (c_in - 64 * (cpt - 1) + 15) >> 4;
  uint32_t mb_a[NMY], mb_d[NMY];
#pragma unroll
  for (int q = 0; q < NMY; ++q) {
    const int mb = mc.mi + q * p.n_mma;
    const int jx = mb % p.mx, jy = mb / p.mx;
    mb_a[q] = HALO ? (uint32_t)(jy * 16 * p.pw + jx * 8) * 8u : (uint32_t)mb * (kTileM * 128 / 16);
    mb_d[q] = mb * p.acc_pitch;
  }
  const bool leader = elect_one();
  const uint16_t cmask = (uint16_t)((1u << (1u << CSHIFT(p))) - 1u);  // all CTAs of the cluster
  // a weight slot is refilled by multicast into every CTA of the cluster, so its release goes to all of them
  auto release_b = [&](uint32_t bs) {
    if (AQ_PAIR) commit_to(b_empty(bs));
    else if (CSHIFT(p)) umma_commit_mc(b_empty(bs), cmask);
    else umma_commit(b_empty(bs));
  };
  Ring ra, rb, rt;
  int it = 0;
  for (int tile = tile_first(p); tile < mc.total_tiles; tile += tile_step(p), ++it, rt.advance(acc_stages)) {
    const uint32_t as = rt.slot;
    const long long tr0 = TR_NOW();
    long long trwa = 0, trwb = 0;
    mbar_wait(tempty(as), rt.phase ^ 1u);
    const long long tr1 = TR_NOW();
    tc_fence_after();
    const uint32_t d_tmem = mc.tmem_base + as * acc_stride;
    uint32_t acc = 0;
    // one K step: `kk` MMAs of K = 16 per owned M block on A view `al`, weight slot `bs`
    auto issue = [&](uint32_t al, uint32_t bl, int kk) {
#pragma unroll
      for (int q = 0; q < NMY; ++q) {
        const uint32_t d = d_tmem + mb_d[q];
        const uint32_t aq = al + mb_a[q];
#if AQ_PAIR
#define AQ_UMMA umma2_f16_split
#else
#define AQ_UMMA umma_f16_split
#endif
        AQ_UMMA(d, aq, a_hi, bl, b_hi, idesc, acc);
        if (kk > 1) AQ_UMMA(d, aq + 2, a_hi, bl + 2, b_hi, idesc, 1u);
        if (kk > 2) AQ_UMMA(d, aq + 4, a_hi, bl + 4, b_hi, idesc, 1u);
        if (kk > 3) AQ_UMMA(d, aq + 6, a_hi, bl + 6, b_hi, idesc, 1u);
      }
    };
    if (FEED == 2) {
      // four parity planes per chunk, one A slot each; tap (kh, kw) of plane (yp, xp) is the plane shifted by
      // (dy, dx) = (kh > 0, kw > 0) pixels: plane 0 = (1,1): 4 taps, 1 = (1,0): 2, 2 = (0,1): 2, 3 = (0,0): 1
      int s = 0;
      for (int c = 0; c < cpt; ++c) {
        const int kk = c == cpt - 1 ? kk_last : 4;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          TR_WAIT(trwa, mbar_wait(a_full(ra.slot), ra.phase));
          const uint32_t a_lo = a_lo0 + ra.slot * a_slot16;
          constexpr int kTaps[4] = {4, 2, 2, 1};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (j >= (q == 0 ? 4 : (q == 3 ? 1 : 2))) continue;
            // (dy, dx) of tap j in plane q
            const int dy = q == 0 ? (j >> 1) : (q == 1 ? j : 1);
            const int dx = q == 0 ? (j & 1) : (q == 2 ? j : 1);
            const uint32_t tap_off = dy ? pw8 + (dx ? 8u : 0u) : (dx ? 8u : 0u);
            const uint32_t bs = BRES ? (uint32_t)s : rb.slot;
            if (BRES) {
              if (it == 0) TR_WAIT(trwb, mbar_wait(b_full(bs), 0));
            } else {
              TR_WAIT(trwb, mbar_wait(b_full(bs), rb.phase));
            }
            tc_fence_after();
            if (leader) {
              issue(a_lo + tap_off, b_lo0 + bs * b_slot16, kk);
              if (!BRES) release_b(bs);
              if (j == kTaps[q] - 1) {
                commit_to(a_empty(ra.slot));
                if (q == 3 && c == cpt - 1) commit_to(tfull(as));
              }
            }
            __syncwarp();
            acc = 1u;
            if (!BRES) rb.advance(nb);
            ++s;
          }
          ra.advance(na);
        }
      }
    } else if (HALO) {
      int s = 0;
      for (int c = 0; c < cpt; ++c) {
        TR_WAIT(trwa, mbar_wait(a_full(ra.slot), ra.phase));
        const uint32_t a_lo = a_lo0 + ra.slot * a_slot16;
        const int kk = c == cpt - 1 ? kk_last : 4;
        uint32_t tap_off = 0;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap, ++s) {
          // dense weights (c_in <= 16, resident): four taps per 64-wide chunk, K offset 16 elements = 32 bytes
          const bool dense = BRES && b_dense;
          const uint32_t bs = BRES ? (dense ? (uint32_t)(tap >> 2) : (uint32_t)s) : rb.slot;
          if (BRES) {
            if (it == 0) TR_WAIT(trwb, mbar_wait(b_full(bs), 0));
          } else {
            TR_WAIT(trwb, mbar_wait(b_full(bs), rb.phase));
          }
          tc_fence_after();
          if (leader) {
            issue(a_lo + tap_off, b_lo0 + bs * b_slot16 + (dense ? 2u * (uint32_t)(tap & 3) : 0u), kk);
            if (!BRES) release_b(bs);  // weight slot reusable once these MMAs retire
            if (tap == 8) {
              commit_to(a_empty(ra.slot));
              if (c == cpt - 1) commit_to(tfull(as));  // accumulators ready for the epilogue
            }
          }
          __syncwarp();
          acc = 1u;
          if (!BRES) rb.advance(nb);
          tap_off += (tap % 3 == 2) ? pw8 - 16u : 8u;
        }
        ra.advance(na);
      }
    } else {
      int chunk = 0;
      for (int s = 0; s < n_steps; ++s) {
        TR_WAIT(trwa, mbar_wait(a_full(ra.slot), ra.phase));
        const uint32_t bs = BRES ? (uint32_t)s : rb.slot;
        if (BRES) {
          if (it == 0) TR_WAIT(trwb, mbar_wait(b_full(bs), 0));
        } else {
          TR_WAIT(trwb, mbar_wait(b_full(bs), rb.phase));
        }
        tc_fence_after();
        const int kk = chunk == cpt - 1 ? kk_last : 4;
        if (leader) {
          issue(a_lo0 + ra.slot * a_slot16, b_lo0 + bs * b_slot16, kk);
          if (!BRES) release_b(bs);
          commit_to(a_empty(ra.slot));
          if (s == n_steps - 1) commit_to(tfull(as));
        }
        __syncwarp();
        acc = 1u;
        if (!BRES) rb.advance(nb);
        ra.advance(na);
        if (++chunk == cpt) chunk = 0;
      }
    }
    if (mc.mi == 0 && leader) trace_put<TRACE>(p, 2, it, tr0, tr1 - tr0, (trwa << 32) | trwb, TR_NOW());
  }
}


struct EpiCtx {
  uint32_t bar0, stg_base, tmem_base, bias_smem;
  int total_tiles;
  uint32_t e, lane;  // epilogue warp index 0..15, lane
};

// Epilogue role of the storing kernel: TMEM -> registers -> bias + activation -> fp16 -> swizzled staging -> TMA
// store / reduce-add.  Warp e reads TMEM lanes 32 * (e % 4) .. +31 (rows of the M block) and the 16-column chunk
// e / 4 of every 64-channel group, so its staging offsets and TMEM column are loop invariants.  Output leaves in
// 64-channel groups through a ring of 16 KB staging buffers: all sixteen warps fill a group and ARRIVE (without
// waiting) on the buffer's "full" named barrier; the store warp waits there, issues the TMA store, and arrives on
// the buffer's "empty" barrier once the store has read shared memory.  An epilogue warp only blocks when it comes
// back to a buffer whose store is still in flight (issuing a store costs the issuing thread ~500 clk, measured).
// Every warp runs at the latency of its own dependent instruction stream (~8 clk per instruction measured,
// profiles/r1_trace_*.txt), so the per-group path is kept to the arithmetic: all parameters live in registers.
template <int ACT, bool TRACE>
__device__ __forceinline__ void epi_store_role(const ConvKParams& p, const EpiCtx& ec) {
  const uint32_t bar0 = ec.bar0;
  auto tfull = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + a); };
  auto tempty = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + 2 + a); };
  const uint32_t e = ec.e, lane = ec.lane;
  const uint32_t q = e & 3, sub = e >> 2;
  const uint32_t row = q * 32 + lane;  // accumulator row == TMEM lane == M row of the block
  // staging offsets of this thread's two 16-byte pieces inside a group buffer [128 rows][128 B]:
  // 16-byte chunk j of row r sits at j ^ (r & 7)
  const uint32_t st0 = row * 128u + (((2 * sub) ^ (row & 7u)) << 4);
  const uint32_t st1 = row * 128u + (((2 * sub + 1) ^ (row & 7u)) << 4);
  const int n_chunks = p.tile_n >> 4, n_groups = p.groups, mblk = p.mblk, tile_n = p.tile_n;
  const uint32_t n_stg = p.n_stg, acc_stages = p.acc_stages, acc_stride = p.acc_stride, acc_pitch = p.acc_pitch;
  const uint32_t t_lane = ec.tmem_base + ((q * 32u) << 16) + sub * 16u;
  const uint32_t stg_base = ec.stg_base;
  const int total_tiles = ec.total_tiles;
  uint32_t v[16];
  bool preloaded = false;
  uint32_t gcount = 0;  // groups staged so far by this CTA
  Ring rs, rt;
  int it = 0;
  for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), ++it, rt.advance(acc_stages)) {
    int m_tile, n_idx;
    split_tile(p, tile, m_tile, n_idx);
    const int n0 = n_idx * tile_n;
    const uint32_t as = rt.slot;
    const long long tr0 = TR_NOW();
    long long trs = 0, d0 = 0, d1 = 0, d2 = 0, d3 = 0, d4 = 0, d5 = 0, d6 = 0, d7 = 0;
    long long tp = TR_NOW();
#define TR_PT(acc) do { if (TRACE) { const long long n_ = clock64(); acc += n_ - tp; tp = n_; } } while (0)
    mbar_wait(tfull(as), rt.phase);
    const long long tr1 = TR_NOW();
    tc_fence_after();
    uint32_t ta = t_lane + as * acc_stride;                      // TMEM address of (mb, g = 0, this chunk)
    const uint32_t bias0 = ec.bias_smem + (uint32_t)(n0 + (int)sub * 16) * 4u;
    for (int mb = 0; mb < mblk; ++mb, ta += acc_pitch) {
      for (int g = 0; g < n_groups; ++g, ++gcount) {
        const uint32_t buf = rs.slot;
        const uint32_t stg = stg_base + buf * (kTileM * 128);
        rs.advance(n_stg);
        const bool valid = 4 * g + (int)sub < n_chunks;  // warp-uniform
        TR_PT(d0);
        if (!preloaded && valid && !(p.dbg & 8)) tmem_ld16(ta + g * 64, v);
        tmem_ld_wait();
        TR_PT(d1);
        const bool last_of_tile = mb == mblk - 1 && g == n_groups - 1;
        if (last_of_tile) {
          // every TMEM read of this tile is done -> the MMA warps may overwrite the accumulators
          tc_fence_before();
          __syncwarp();
          if (lane == 0) arrive_tempty(tempty(as));
        }
        TR_PT(d2);
        // the TMA store that last used this staging buffer must have read it
        if (gcount >= n_stg && !(p.dbg & 32)) TR_WAIT(trs, named_bar_sync(kBarStgEmpty + buf, kBarStgThreads));
        if (valid && !(p.dbg & 2)) {
          uint32_t b4[16];
          const uint32_t ba = bias0 + (uint32_t)g * 256u;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(b4[4 * i]), "=r"(b4[4 * i + 1]), "=r"(b4[4 * i + 2]), "=r"(b4[4 * i + 3])
                         : "r"(ba + 16u * i));
          float f[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (ACT == ACT_SILU_FAST) {
              // silu(x) = h + h * tanh(h), h = x / 2; s_bias holds bias / 2
              const float h = fmaf(__uint_as_float(v[i]), 0.5f, __uint_as_float(b4[i]));
              f[i] = fmaf(h, tanh_approx(h), h);
            } else {
              const float x = __uint_as_float(v[i]) + __uint_as_float(b4[i]);
              f[i] = ACT == ACT_SILU ? silu_exact(x) : x;
            }
          }
          uint32_t h[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) h[i] = pack_half2(f[2 * i], f[2 * i + 1]);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(stg + st0), "r"(h[0]), "r"(h[1]), "r"(h[2]), "r"(h[3])
                       : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(stg + st1), "r"(h[4]), "r"(h[5]), "r"(h[6]), "r"(h[7])
                       : "memory");
        }
        TR_PT(d3);
        // the next group's accumulators travel TMEM -> registers while this group is stored
        preloaded = !last_of_tile;
        if (preloaded) {
          const bool same = g + 1 < n_groups;
          const int ng = same ? g + 1 : 0;
          if (4 * ng + (int)sub < n_chunks && !(p.dbg & 8)) tmem_ld16((same ? ta : ta + acc_pitch) + ng * 64, v);
        }
        TR_PT(d4);
        if (!(p.dbg & 16)) fence_proxy_async_smem();  // staging writes -> visible to the async proxy (TMA)
        TR_PT(d5);
        if (!(p.dbg & 32)) named_bar_arrive(kBarStgFull + buf, kBarStgThreads);  // hand the group to the store warp
        TR_PT(d6);
      }
    }
    if (e == 0 && lane == 0) {
      trace_put<TRACE>(p, 3, it, tr0, tr1 - tr0, trs, TR_NOW());
      trace_put<TRACE>(p, 4, it, d0 | (d1 << 32), d2 | (d3 << 32), d4 | (d5 << 32), d6 | (d7 << 32));
    }
#undef TR_PT
  }
}


// Two-set variant of the epilogue (p.epi_sets == 2).  The sixteen-warp epilogue above runs every phase of a group in
// lockstep -- TMEM load, the MUFU burst of the activation (8 clk per warp instruction per SM sub-partition, all
// warps at once), staging, proxy fence, barrier hand-off -- so a group costs the SUM of its phases (~1.2-1.4k clk,
// measured with AQUA_DBG: profiles/r1_attribution_stem_v7.txt), which bounds every layer with few K steps (the stem,
// the 1x1 / 3x3 convs of the 160x160 stage).  Here the warps form two sets of eight (TMEM lane quarter x two
// 16-column chunks each); set s takes the groups with global index = s (mod 2), so one set's MUFU burst overlaps the
// other's loads / fences / barrier latency.  Staging ring, store warp and the order of the stores are unchanged;
// the named barriers count 8 warps + the store warp.
constexpr uint32_t kBarStgThreads2 = (kEpiWarps / 2) * 32 + 32;

template <int ACT>
__device__ __forceinline__ void epi_math_store(const uint32_t (&v)[16], uint32_t bias_addr, uint32_t dst0, uint32_t dst1) {
  uint32_t b4[16];
#pragma unroll
  for (int i = 0; i < 4; ++i)
    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(b4[4 * i]), "=r"(b4[4 * i + 1]), "=r"(b4[4 * i + 2]), "=r"(b4[4 * i + 3])
                 : "r"(bias_addr + 16u * i));
  float f[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    if (ACT == ACT_SILU_FAST) {
      const float h = fmaf(__uint_as_float(v[i]), 0.5f, __uint_as_float(b4[i]));  // s_bias holds bias / 2
      f[i] = fmaf(h, tanh_approx(h), h);
    } else {
      const float x = __uint_as_float(v[i]) + __uint_as_float(b4[i]);
      f[i] = ACT == ACT_SILU ? silu_exact(x) : x;
    }
  }
  uint32_t h[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) h[i] = pack_half2(f[2 * i], f[2 * i + 1]);
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst0), "r"(h[0]), "r"(h[1]), "r"(h[2]), "r"(h[3]) : "memory");
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst1), "r"(h[4]), "r"(h[5]), "r"(h[6]), "r"(h[7]) : "memory");
}

template <int ACT>
__device__ __forceinline__ void epi_store_role2(const ConvKParams& p, const EpiCtx& ec) {
  const uint32_t bar0 = ec.bar0;
  auto tfull = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + a); };
  auto tempty = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + 2 + a); };
  const uint32_t e = ec.e, lane = ec.lane;
  const uint32_t q = e & 3, sub = e >> 2;
  const uint32_t set = sub >> 1, half = sub & 1;  // this warp's chunks of a 64-channel group: half and half + 2
  const uint32_t row = q * 32 + lane;
  const uint32_t sw = row & 7u, rb = row * 128u;
  // staging offsets of the four 16-byte pieces (chunk c = pieces 2c, 2c + 1; piece j of row r sits at j ^ (r & 7))
  const uint32_t sa0 = rb + (((2 * half) ^ sw) << 4), sa1 = rb + (((2 * half + 1) ^ sw) << 4);
  const uint32_t sb0 = rb + (((2 * half + 4) ^ sw) << 4), sb1 = rb + (((2 * half + 5) ^ sw) << 4);
  const int n_chunks = p.tile_n >> 4, n_groups = p.groups, tile_n = p.tile_n;
  const int G = p.mblk * n_groups;  // groups per tile
  const uint32_t n_stg = p.n_stg, acc_stages = p.acc_stages, acc_stride = p.acc_stride, acc_pitch = p.acc_pitch;
  const uint32_t t_lane = ec.tmem_base + ((q * 32u) << 16);
  const int total_tiles = ec.total_tiles;
  uint32_t va[16], vb[16];
  uint32_t k0 = 0;    // global index of the tile's first group
  uint32_t buf0 = 0;  // k0 % n_stg
  Ring rt;
  for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), rt.advance(acc_stages)) {
    int m_tile, n_idx;
    split_tile(p, tile, m_tile, n_idx);
    const int n0 = n_idx * tile_n;
    const uint32_t as = rt.slot;
    mbar_wait(tfull(as), rt.phase);
    tc_fence_after();
    const uint32_t ta0 = t_lane + as * acc_stride;
    // own groups of this tile: j = j0, j0 + 2, ... < G with (k0 + j) % 2 == set
    int j = ((k0 & 1u) == set) ? 0 : 1;
    int mb = 0, g = j;
    while (g >= n_groups) { g -= n_groups; ++mb; }
    uint32_t buf = buf0 + (uint32_t)j;
    if (buf >= n_stg) buf -= n_stg;
    bool loaded = false;
    if (j >= G) {  // nothing of this tile is ours: release the accumulators straight away
      tc_fence_before();
      __syncwarp();
      if (lane == 0) arrive_tempty(tempty(as));
    }
    for (; j < G; j += 2) {
      const uint32_t ta = ta0 + (uint32_t)mb * acc_pitch + (uint32_t)g * 64u;
      const bool v0 = 4 * g + (int)half < n_chunks, v1 = 4 * g + (int)half + 2 < n_chunks;  // warp-uniform
      if (!loaded) {
        if (v0) tmem_ld16(ta + half * 16u, va);
        if (v1) tmem_ld16(ta + (half + 2u) * 16u, vb);
      }
      tmem_ld_wait();
      const bool last_own = j + 2 >= G;
      if (last_own) {  // every TMEM read of this tile by this warp is done
        tc_fence_before();
        __syncwarp();
        if (lane == 0) arrive_tempty(tempty(as));
      }
      // the TMA store that last used this staging buffer (group k - n_stg) must have read it
      if (k0 + (uint32_t)j >= n_stg) named_bar_sync((set ? kBarStgEmptySet1 : kBarStgEmpty) + buf, kBarStgThreads2);
      const uint32_t stg = ec.stg_base + buf * (kTileM * 128);
      const uint32_t bias_g = ec.bias_smem + (uint32_t)(n0 + g * 64) * 4u;
      if (v0) epi_math_store<ACT>(va, bias_g + half * 64u, stg + sa0, stg + sa1);
      if (v1) epi_math_store<ACT>(vb, bias_g + (half + 2u) * 64u, stg + sb0, stg + sb1);
      // next own group (same tile): its accumulators travel TMEM -> registers while this group is stored
      int ng = g + 2, nmb = mb;
      while (ng >= n_groups) { ng -= n_groups; ++nmb; }
      loaded = !last_own;
      if (loaded) {
        const uint32_t tn = ta0 + (uint32_t)nmb * acc_pitch + (uint32_t)ng * 64u;
        if (4 * ng + (int)half < n_chunks) tmem_ld16(tn + half * 16u, va);
        if (4 * ng + (int)half + 2 < n_chunks) tmem_ld16(tn + (half + 2u) * 16u, vb);
      }
      fence_proxy_async_smem();
      named_bar_arrive(kBarStgFull + buf, kBarStgThreads2);
      g = ng;
      mb = nmb;
      buf += 2;
      while (buf >= n_stg) buf -= n_stg;
    }
    k0 += (uint32_t)G;
    buf0 += (uint32_t)G;
    while (buf0 >= n_stg) buf0 -= n_stg;
  }
}

// Store warp: waits for each staged group, issues its TMA store / reduce-add, frees staging buffers.
template <bool TRACE>
__device__ __forceinline__ void store_role(const ConvKParams& p, uint32_t stg_base, int total_tiles, uint32_t lane) {
  const int n_groups = p.groups, mblk = p.mblk, tile_n = p.tile_n;
  const uint32_t n_stg = p.n_stg;
  const int my_tiles = (total_tiles - tile_first(p) + tile_step(p) - 1) / tile_step(p);
  const uint32_t total_groups = (uint32_t)(my_tiles * mblk * n_groups);
  Ring rs;
  uint32_t k = 0;  // groups stored so far
  int it = 0;
  const bool two_sets = p.epi_sets == 2 && !TRACE;  // traced builds keep the lockstep epilogue
  const uint32_t bar_threads = two_sets ? kBarStgThreads2 : kBarStgThreads;
  const bool halo = (p.mode == CONV_MODE_HALO || p.mode == CONV_MODE_HALO2) && !(p.dbg & 64);
  const int hp1 = p.a_h + 1;
  for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), ++it) {
    int m_tile, n_idx;
    split_tile(p, tile, m_tile, n_idx);
    const int n0 = n_idx * tile_n;
    const MBlock t0 = mblock_coords(p, m_tile, 0);
    const long long tr0 = TR_NOW();
    long long trw = 0, trr = 0;
    // HALO: lane L works out where M block L falls in the stack of images (one division) while the epilogue is
    // still filling the first group, so the issuing lane's path after the barrier is the TMA instruction alone
    // (one scalar warp runs ~8 clk per dependent instruction: the division + image loop cost ~1k clk per store)
    int pre_n = 0, pre_ys = 0;
    if (halo) {
      const int mbl = (int)lane < mblk ? (int)lane : 0;
      const int y0l = t0.y0 + (mbl >> (p.mx >> 1)) * 16;
      pre_n = y0l / hp1;
      pre_ys = y0l - (pre_n * hp1 + 1);
    }
    for (int mb = 0; mb < mblk; ++mb) {
      MBlock t = t0;
      if (p.mode == CONV_MODE_TAP) {
        t = mblock_coords(p, m_tile, mb);
      } else {  // derive from the tile origin without divisions (mx is 1, 2 or 4)
        const int jx = mb & (p.mx - 1), jy = mb >> (p.mx >> 1);
        t.m0 += mb * kTileM;
        t.x0 += jx * 8;
        t.y0 += jy * 16;
      }
      const int img = __shfl_sync(0xffffffffu, pre_n, mb), ys = __shfl_sync(0xffffffffu, pre_ys, mb);
      // the 16 rows lie inside one image (plus, at most, the zero row below it, which TMA clips)
      const bool one_image = halo && ys >= 0 && ys + 15 <= p.a_h;
      for (int g = 0; g < n_groups; ++g, ++k) {
        const uint32_t buf = rs.slot;
        rs.advance(n_stg);
        if (p.dbg & 32) continue;
        TR_WAIT(trw, named_bar_sync(kBarStgFull + buf, bar_threads));
        if (lane == 0) {
          if (p.dbg & 1) {
          } else if (one_image) {
            if (img < p.a_n) {
              if (p.has_res) tma_reduce_add_4d(&p.tmOut, stg_base + buf * (kTileM * 128), n0 + g * 64, t.x0, ys, img);
              else tma_store_4d(&p.tmOut, stg_base + buf * (kTileM * 128), n0 + g * 64, t.x0, ys, img);
            }
          } else {
            store_group(p, t, stg_base + buf * (kTileM * 128), n0 + g * 64);
          }
          tma_store_commit();
          // all but the latest n_stg - 2 stores have read their buffers -> the buffer of store k - (n_stg - 2),
          // which group k + 2 will fill, is free
          const long long tq = TR_NOW();
          if (n_stg == 2) tma_store_wait_read0();
          else if (n_stg == 3) tma_store_wait_read1();
          else tma_store_wait_read2();  // n_stg == 4
          trr += TR_NOW() - tq;
        }
        __syncwarp();
        if (k + 2 >= n_stg && k + 2 < total_groups) {
          uint32_t fb = buf + 2;  // (k + 2) % n_stg
          if (fb >= n_stg) fb -= n_stg;
          // two sets: group k + 2 belongs to set (k + 2) % 2 = k % 2 (groups per tile are counted globally in both roles)
          named_bar_arrive((two_sets && (k & 1u) ? kBarStgEmptySet1 : kBarStgEmpty) + fb, bar_threads);
        }
      }
    }
    if (lane == 0) trace_put<TRACE>(p, 5, it, tr0, trw, trr, TR_NOW());
  }
  // the CTA may exit (and its shared memory be handed to the next kernel's CTA) once every store has READ its staging
  // buffer; the writes themselves are complete and visible when the grid completes, which is what a dependent kernel's
  // griddepcontrol.wait / the stream order waits for (AQUA build flag AQ_STORE_WAIT_FULL restores the full wait)
#ifdef AQ_STORE_WAIT_FULL
  if (lane == 0) tma_store_wait0();
#else
  if (lane == 0) tma_store_wait_read0();
#endif
}


// ===================================================================================================================
// EPI_CHAIN: a 1x1 / stride-1 conv chained behind this conv inside the same kernel (ConvKParams::chain, conv_igemm.cuh).
//
// Per CTA the staged groups follow ONE sequence that the epilogue warps, the store warp and the chained-MMA warp walk in
// step (ring of n_stg staging buffers, group k -> buffer k % n_stg):
//     block 0: G1 groups of this conv | block 1: G1 groups | G2 groups of block 0's chained conv | block 2: G1 | G2 of block 1 ...
// i.e. the chained conv's epilogue lags one M block behind, so its MMAs (issued as soon as the block's K groups are staged)
// have a block's worth of epilogue time to complete.  G1 = every 64-channel group of this conv when its own output is stored
// too, else only the chain_kg groups the chained conv consumes.  Barriers beside the usual ones (mbarriers, explicit phases):
//   c_full[buf]  the 16 epilogue warps -> chained-MMA warp: K group staged in `buf` (after the proxy fence)
//   c_done[buf]  tcgen05.commit -> store warp: the chained MMAs have read `buf` (it may be refilled once its store has too)
//   tfull2[s] / tempty2[s]   chained accumulator stage s (block index & 1) ready / drained
constexpr int kBarChain = 2 * kMaxA + 2 * kMaxB + 8;  // first chain barrier inside `bars`
constexpr uint32_t kWarpMma2 = 21;

struct ChainBars {
  uint32_t bar0;
  __device__ __forceinline__ uint32_t c_full(uint32_t b) const { return bar0 + 8u * (kBarChain + b); }
  __device__ __forceinline__ uint32_t c_done(uint32_t b) const { return bar0 + 8u * (kBarChain + 4 + b); }
  __device__ __forceinline__ uint32_t b2_full() const { return bar0 + 8u * (kBarChain + 8); }
  __device__ __forceinline__ uint32_t tfull2(uint32_t s) const { return bar0 + 8u * (kBarChain + 9 + s); }
  __device__ __forceinline__ uint32_t tempty2(uint32_t s) const { return bar0 + 8u * (kBarChain + 11 + s); }
};

template <int ACT>
__device__ __forceinline__ void epi_chunk_math_store(const uint32_t (&v)[16], uint32_t bias_addr, uint32_t dst0, uint32_t dst1) {
  epi_math_store<ACT>(v, bias_addr, dst0, dst1);
}

__device__ __forceinline__ void epi_chunk_dispatch(int act, const uint32_t (&v)[16], uint32_t bias_addr, uint32_t dst0, uint32_t dst1) {
  if (act == ACT_SILU_FAST) epi_chunk_math_store<ACT_SILU_FAST>(v, bias_addr, dst0, dst1);
  else if (act == ACT_SILU) epi_chunk_math_store<ACT_SILU>(v, bias_addr, dst0, dst1);
  else epi_chunk_math_store<ACT_NONE>(v, bias_addr, dst0, dst1);
}

// Epilogue warps (sixteen, lockstep): this conv's groups, then the previous block's chained groups.
__device__ __forceinline__ void epi_chain_role(const ConvKParams& p, const EpiCtx& ec, uint32_t bias2_smem) {
  const uint32_t bar0 = ec.bar0;
  const ChainBars cb{bar0};
  auto tfull = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + a); };
  auto tempty = [&](uint32_t a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + 2 + a); };
  const uint32_t e = ec.e, lane = ec.lane;
  const uint32_t q = e & 3, sub = e >> 2;
  const uint32_t row = q * 32 + lane;
  const uint32_t st0 = row * 128u + (((2 * sub) ^ (row & 7u)) << 4);
  const uint32_t st1 = row * 128u + (((2 * sub + 1) ^ (row & 7u)) << 4);
  const int n_chunks1 = p.tile_n >> 4, n_chunks2 = p.tile_n2 >> 4, mblk = p.mblk;
  const int G1 = p.chain_store1 ? p.groups : p.chain_kg, KG = p.chain_kg, G2 = p.groups2;
  const uint32_t n_stg = p.n_stg, acc_stages = p.acc_stages;
  const uint32_t t_lane = ec.tmem_base + ((q * 32u) << 16) + sub * 16u;
  uint32_t v[16];
  uint32_t gcount = 0;
  Ring rs, rt;
  uint32_t blk = 0;  // blocks whose own groups have been staged
  // With two chained accumulator stages the chained epilogue lags ONE block behind (its MMAs get a block's worth of time);
  // with one stage it follows its own block at once: the chained MMAs of block i wait for the chained epilogue of block
  // i - 1, which needs staging buffers -- and in ring order those are held by block i's not yet consumed K groups.
  const uint32_t lag = p.acc2_stages == 2 ? 1u : 0u;
  // one staged group: accumulator chunk `ta` (this warp's 16 columns) -> bias + activation -> staging buffer
  auto stage = [&](uint32_t ta, bool valid, uint32_t bias_addr, int act, bool k_group, uint32_t release_bar) {
    const uint32_t buf = rs.slot;
    rs.advance(n_stg);
    if (valid) tmem_ld16(ta, v);
    tmem_ld_wait();
    if (release_bar) {  // every TMEM read of the accumulator stage is done
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(release_bar);
    }
    if (gcount >= n_stg) named_bar_sync(kBarStgEmpty + buf, kBarStgThreads);
    const uint32_t stg = ec.stg_base + buf * (kTileM * 128);
    if (valid) epi_chunk_dispatch(act, v, bias_addr, stg + st0, stg + st1);
    fence_proxy_async_smem();
    if (k_group) {  // the chained MMAs read this buffer through the async proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(cb.c_full(buf));
    }
    named_bar_arrive(kBarStgFull + buf, kBarStgThreads);
    ++gcount;
  };
  auto chained_groups = [&](uint32_t b) {  // epilogue of block b's chained conv
    const uint32_t s2 = p.acc2_stages == 2 ? (b & 1u) : 0u;
    mbar_wait(cb.tfull2(s2), (p.acc2_stages == 2 ? (b >> 1) : b) & 1u);
    tc_fence_after();
    const uint32_t ta2 = t_lane + p.acc2_base + s2 * p.acc2_pitch;
    for (int g = 0; g < G2; ++g)
      stage(ta2 + g * 64, 4 * g + (int)sub < n_chunks2, bias2_smem + (uint32_t)(g * 64 + (int)sub * 16) * 4u, p.act2, false,
            g == G2 - 1 ? cb.tempty2(s2) : 0u);
  };
  for (int tile = tile_first(p); tile < ec.total_tiles; tile += tile_step(p), rt.advance(acc_stages)) {
    const uint32_t as = rt.slot;
    mbar_wait(tfull(as), rt.phase);
    tc_fence_after();
    for (int mb = 0; mb < mblk; ++mb, ++blk) {
      const uint32_t ta = t_lane + as * p.acc_stride + (uint32_t)mb * p.acc_pitch;
      for (int g = 0; g < G1; ++g)
        stage(ta + g * 64, 4 * g + (int)sub < n_chunks1, ec.bias_smem + (uint32_t)(g * 64 + (int)sub * 16) * 4u, p.act, g < KG,
              (mb == mblk - 1 && g == G1 - 1) ? tempty(as) : 0u);
      if (blk >= lag) chained_groups(blk - lag);
    }
  }
  if (lag && blk > 0) chained_groups(blk - 1);
}

// Chained-MMA warp: D2[128, tile_n2] (+)= stage[g][128, 64] * W2[g][tile_n2, 64]^T over the block's K groups.
__device__ __forceinline__ void mma2_role(const ConvKParams& p, uint32_t bar0, uint32_t stg_base, uint32_t b2_base, uint32_t tmem_base,
                                          int total_tiles) {
  const ChainBars cb{bar0};
  const uint32_t idesc2 = make_idesc_f16(kTileM, p.tile_n2);
  const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29), lo_flags = 1u << 16;
  const int G1 = p.chain_store1 ? p.groups : p.chain_kg, KG = p.chain_kg, G2 = p.groups2, mblk = p.mblk;
  const uint32_t n_stg = p.n_stg;
  const bool leader = elect_one();
  const uint32_t lag = p.acc2_stages == 2 ? 1u : 0u;
  mbar_wait(cb.b2_full(), 0);
  uint32_t cf_phase = 0;  // bit b: parity of c_full[b]'s next completion
  Ring rs;
  uint32_t blk = 0;
  for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p)) {
    for (int mb = 0; mb < mblk; ++mb, ++blk) {
      const uint32_t s2 = p.acc2_stages == 2 ? (blk & 1u) : 0u;
      mbar_wait(cb.tempty2(s2), ((p.acc2_stages == 2 ? (blk >> 1) : blk) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d = tmem_base + p.acc2_base + s2 * p.acc2_pitch;
      for (int g = 0; g < G1; ++g) {
        const uint32_t buf = rs.slot;
        rs.advance(n_stg);
        if (g >= KG) continue;
        mbar_wait(cb.c_full(buf), (cf_phase >> buf) & 1u);
        cf_phase ^= 1u << buf;
        tc_fence_after();
        if (leader) {
          const uint32_t a_lo = (((stg_base + buf * (kTileM * 128)) & 0x3FFFFu) >> 4) | lo_flags;
          const uint32_t b_lo = (((b2_base + (uint32_t)g * p.b2_slot_bytes) & 0x3FFFFu) >> 4) | lo_flags;
          const int kk = g == KG - 1 ? p.chain_kk_last : 4;
          for (int j = 0; j < kk; ++j) umma_f16_split(d, a_lo + 2u * j, hi, b_lo + 2u * j, hi, idesc2, (g > 0 || j > 0) ? 1u : 0u);
          umma_commit(cb.c_done(buf));                      // the buffer has been read
          if (g == KG - 1) umma_commit(cb.tfull2(s2));      // the chained accumulators are complete
        }
        __syncwarp();
      }
      if (blk >= lag)
        for (int g = 0; g < G2; ++g) rs.advance(n_stg);  // block (blk - lag)'s chained groups pass through the ring here
    }
  }
}

// Store warp of the chained kernel: same sequence; a buffer is handed back once its TMA store has read it AND -- if it held
// a K group -- the chained MMAs have.
__device__ __forceinline__ void store_chain_role(const ConvKParams& p, uint32_t bar0, uint32_t stg_base, int total_tiles, uint32_t lane) {
  const ChainBars cb{bar0};
  const int G1 = p.chain_store1 ? p.groups : p.chain_kg, KG = p.chain_kg, G2 = p.groups2, mblk = p.mblk;
  const uint32_t n_stg = p.n_stg;
  const int my_tiles = (total_tiles - tile_first(p) + tile_step(p) - 1) / tile_step(p);
  const uint32_t total_groups = (uint32_t)(my_tiles * mblk * (G1 + G2));
  Ring rs;
  uint32_t k = 0;          // groups handled so far
  uint32_t kmask = 0;      // bit (j % 32): group j was a K group (n_stg <= 4 groups back is all that is ever looked at)
  uint32_t cd_phase = 0;   // bit b: parity of c_done[b]'s next completion
  uint32_t blk = 0;
  const uint32_t lag = p.acc2_stages == 2 ? 1u : 0u;
  MBlock prev = {};
  auto one = [&](bool chained, bool k_group, const MBlock& t, int ch0) {
    const uint32_t buf = rs.slot;
    rs.advance(n_stg);
    kmask = (kmask & ~(1u << (k & 31u))) | ((k_group ? 1u : 0u) << (k & 31u));
    named_bar_sync(kBarStgFull + buf, kBarStgThreads);
    if (lane == 0) {
      const uint32_t src = stg_base + buf * (kTileM * 128);
      if (chained) store_group_to(p, &p.tmOut2, p.tmOut2R, false, t, src, ch0);
      else if (p.chain_store1) store_group_to(p, &p.tmOut, p.tmOutR, p.has_res != 0, t, src, ch0);
      tma_store_commit();  // an empty group when nothing was stored: the wait below stays uniform
      if (n_stg == 2) tma_store_wait_read0();
      else if (n_stg == 3) tma_store_wait_read1();
      else tma_store_wait_read2();
    }
    __syncwarp();
    if (k + 2 >= n_stg && k + 2 < total_groups) {
      uint32_t fb = buf + 2;  // (k + 2) % n_stg: last used by group k + 2 - n_stg
      if (fb >= n_stg) fb -= n_stg;
      const uint32_t j = k + 2 - n_stg;
      if ((kmask >> (j & 31u)) & 1u) {
        mbar_wait(cb.c_done(fb), (cd_phase >> fb) & 1u);
        cd_phase ^= 1u << fb;
      }
      named_bar_arrive(kBarStgEmpty + fb, kBarStgThreads);
    }
    ++k;
  };
  for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p)) {
    int m_tile, n_idx;
    split_tile(p, tile, m_tile, n_idx);
    for (int mb = 0; mb < mblk; ++mb, ++blk) {
      const MBlock t = mblock_coords(p, m_tile, mb);
      for (int g = 0; g < G1; ++g) one(false, g < KG, t, g * 64);
      if (blk >= lag)
        for (int g = 0; g < G2; ++g) one(true, false, lag ? prev : t, g * 64);
      prev = t;
    }
  }
  if (lag && blk > 0)
    for (int g = 0; g < G2; ++g) one(true, false, prev, g * 64);
  if (lane == 0) tma_store_wait_read0();
}

template <int EPI, bool TRACE>
__global__ void __launch_bounds__(EPI == EPI_CHAIN ? kConvThreadsChain : kConvThreads, 1) conv_igemm_kernel(const __grid_constant__ ConvKParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * kMaxA + 2 * kMaxB + 8 + (EPI == EPI_CHAIN ? 16 : 0)];
  __shared__ __align__(16) float s_bias[kMaxCout];
  __shared__ __align__(16) float s_bias2[EPI == EPI_CHAIN ? kMaxCout2 : 4];
  __shared__ uint32_t tmem_slot;
  __shared__ uint8_t s_det_a[EPI == EPI_DETECT ? 256 : 1], s_det_o[EPI == EPI_DETECT ? 256 : 1];  // column -> anchor, output

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t a_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_base = a_base + p.na * p.a_slot_bytes;
  const uint32_t stg_base = b_base + p.nb * p.b_slot_bytes;
  const uint32_t b2_base = stg_base + p.n_stg * p.stg_bytes;  // EPI_CHAIN: resident weights of the chained conv

  const uint32_t bar0 = smem_u32(bars);
  auto a_full = [&](int s) { return bar0 + 8u * s; };
  auto a_empty = [&](int s) { return bar0 + 8u * (kMaxA + s); };
  auto b_full = [&](int s) { return bar0 + 8u * (2 * kMaxA + s); };
  auto b_empty = [&](int s) { return bar0 + 8u * (2 * kMaxA + kMaxB + s); };
  auto tfull = [&](int a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + a); };
  auto tempty = [&](int a) { return bar0 + 8u * (2 * kMaxA + 2 * kMaxB + 2 + a); };

  if (warp == 0 && lane == 0 && (p.mode == CONV_MODE_HALO || p.mode == CONV_MODE_HALO2)) {
    for (int hb = 0; hb < 6; ++hb) prefetch_tmap(&p.tmH[hb]);
    if (p.mode == CONV_MODE_HALO2)
      for (int hb = 0; hb < 6; ++hb) prefetch_tmap(&p.tmH2[hb]);
  }
  if (warp == 0 && lane == 0 && (p.mode == CONV_MODE_FLAT || p.mode == CONV_MODE_TAP)) {
    prefetch_tmap(&p.tmA);
    if (p.mode == CONV_MODE_TAP && p.stride == 2) prefetch_tmap(&p.tmA2);
    if (p.up_chunks) prefetch_tmap(&p.tmUp);
  }
  if (warp == kWarpStore && lane == 0 && EPI != EPI_DETECT) prefetch_tmap(&p.tmOut);
  if (warp == kWarpStore && lane == 0 && EPI == EPI_CHAIN) {
    prefetch_tmap(&p.tmOut2);
    prefetch_tmap(&p.tmB2);
  }
  if (warp == kWarpB && lane == 0) {
    prefetch_tmap(&p.tmB);
    for (int s = 0; s < p.na; ++s) {
      mbar_init(a_full(s), 1);
      mbar_init(a_empty(s), p.n_mma);
    }
    for (int s = 0; s < p.nb; ++s) {
      mbar_init(b_full(s), 1);
      mbar_init(b_empty(s), AQ_PAIR ? p.n_mma : (p.n_mma << CSHIFT(p)));  // the MMA issuers of every CTA of a multicast cluster / of the pair's leader
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull(a), p.n_mma);
      mbar_init(tempty(a), AQ_PAIR ? 2 * kEpiWarps : kEpiWarps);  // pair: the leader's counts both CTAs' epilogue warps
    }
    if (EPI == EPI_CHAIN) {
      const ChainBars cb{bar0};
      for (uint32_t b = 0; b < 4; ++b) {
        mbar_init(cb.c_full(b), kEpiWarps);
        mbar_init(cb.c_done(b), 1);
      }
      mbar_init(cb.b2_full(), 1);
      for (uint32_t a = 0; a < 2; ++a) {
        mbar_init(cb.tfull2(a), 1);
        mbar_init(cb.tempty2(a), kEpiWarps);
      }
    }
    fence_barrier_init();
  }
  if (warp == kWarpMma) {
    if (AQ_PAIR) {
      tmem_alloc2(smem_u32(&tmem_slot), p.tmem_cols);
      tmem_relinquish2();
    } else {
      tmem_alloc(smem_u32(&tmem_slot), p.tmem_cols);
      tmem_relinquish();
    }
  }
  for (int i = threadIdx.x; i < p.n_tiles_n * p.tile_n; i += blockDim.x) s_bias[i] = __ldg(p.bias + i);
  if (EPI == EPI_CHAIN)
    for (int i = threadIdx.x; i < p.tile_n2; i += blockDim.x) s_bias2[i] = __ldg(p.bias2 + i);
  if (EPI == EPI_DETECT) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
      const int a = i / p.det_no;
      s_det_a[i] = (uint8_t)a;
      s_det_o[i] = (uint8_t)(i - a * p.det_no);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (CSHIFT(p)) cluster_sync_all();  // peers' barriers are initialised before anything is multicast to them
  const uint32_t tmem_base = tmem_slot;
  griddep_launch_dependents();

  const int total_tiles = p.total_tiles;

  if (warp == kWarpA) {
    // =========================== A producer ===========================
    if (p.mode == CONV_MODE_HALO) {
      // One thread loads the halo patch with TMA boxes (64 ch, pw pixels, h rows).  The patch lives in "tall"
      // coordinates (images stacked with one shared zero row): it is cut into per-image segments, each a few
      // boxes of 32/16/8/4/2/1 rows; rows outside an image (the shared zero rows, x = -1 / W, channels past
      // c_in, images past the batch) come from TMA out-of-bounds zero fill.  Box destinations are only
      // 128-byte aligned, which SWIZZLE_128B honours (absolute-address swizzle, probe T10); a patch costs
      // 2-5 TMA instructions instead of ~2600 cp.async (probe T9: 18x18 patch round trip ~0.8k clk).
      {
        griddep_wait();  // activations are written by the previous kernel

        const uint32_t row_bytes = (uint32_t)p.pw * 128u;
        Ring ra;
        int it = 0;
        for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), ++it) {
          int m_tile, n_idx;
          split_tile(p, tile, m_tile, n_idx);
          const MBlock t0 = mblock_coords(p, m_tile, 0);
          const long long tr0 = TR_NOW();
          long long trw = 0;
          const int xl = t0.x0 - 1;
          // R = -1 (above the first zero row) resolves to n = 0, y = -2
          const PatchSegs sg = patch_segments(t0.y0 - 1, p.ph, p.a_h);
          for (int c = 0; c < p.cpt; ++c, ra.advance(p.na)) {
            const uint32_t slot = ra.slot;
            TR_WAIT(trw, mbar_wait(a_empty(slot), ra.phase ^ 1u));
            if (!elect_one()) continue;
            arm_full(a_full(slot), p.a_tx_bytes);
            patch_issue(sg, t0.y0 - 1, p.ph, p.a_h, a_base + slot * p.a_slot_bytes, row_bytes,
                        [&](int hb, uint32_t dst, int y, int n) { ld4d(dst, &p.tmH[hb], a_full(slot), c * 64, xl, y, n); });
          }
          if (lane == 0) trace_put<TRACE>(p, 0, it, tr0, trw, 0, TR_NOW());
        }
      }
    } else if (p.mode == CONV_MODE_HALO2) {
      // 3x3/s2: per chunk the four parity planes (y-parity, x-parity) = (1,1) (1,0) (0,1) (0,0), one ring slot each.
      // A plane holds half-resolution pixels xh = x0-1 .. x0+8mx-1, rows T = y0-1 .. y0+16my-1 of the tall plane
      // (per image: one zero row = yh -1, then yh 0 .. H/2-1), cut into per-image segments exactly like HALO.
      griddep_wait();

      const uint32_t row_bytes = (uint32_t)p.pw * 128u;
      Ring ra;
      int it = 0;
      for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), ++it) {
        int m_tile, n_idx;
        split_tile(p, tile, m_tile, n_idx);
        const MBlock t0 = mblock_coords(p, m_tile, 0);
        const long long tr0 = TR_NOW();
        long long trw = 0;
        const int xl = t0.x0 - 1;
        const PatchSegs sg = patch_segments(t0.y0 - 1, p.ph, p.a_h);
        for (int cq = 0; cq < p.cpt * 4; ++cq, ra.advance(p.na)) {
          const int c = cq >> 2, q = cq & 3;
          const int yp = q < 2 ? 1 : 0, xp = (q & 1) ? 0 : 1;
          const uint32_t slot = ra.slot;
          TR_WAIT(trw, mbar_wait(a_empty(slot), ra.phase ^ 1u));
          if (!elect_one()) continue;
          arm_full(a_full(slot), p.a_tx_bytes);
          patch_issue(sg, t0.y0 - 1, p.ph, p.a_h, a_base + slot * p.a_slot_bytes, row_bytes, [&](int hb, uint32_t dst, int y, int n) {
            ld5d(dst, xp ? &p.tmH2[hb] : &p.tmH[hb], a_full(slot), c * 64, xl, yp, y, n);
          });
        }
        if (lane == 0) trace_put<TRACE>(p, 0, it, tr0, trw, 0, TR_NOW());
      }
    } else {
      griddep_wait();  // activations are written by the previous kernel
      Ring ra;
      int it = 0;
      for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), ++it) {
        int m_tile, n_idx;
        split_tile(p, tile, m_tile, n_idx);
        const long long tr0 = TR_NOW();
        long long trw = 0;
        MBlock tb0 = {}, tb1 = {};
        if (p.mode == CONV_MODE_TAP) {
          tb0 = mblock_coords(p, m_tile, 0);
          tb1 = mblock_coords(p, m_tile, p.mblk > 1 ? 1 : 0);
        }
        int tap = 0, chunk = 0;  // TAP: grp = tap * cpt + chunk
        for (int grp = 0; grp < p.n_steps; ++grp, ra.advance(p.na)) {
          const uint32_t slot = ra.slot;
          TR_WAIT(trw, mbar_wait(a_empty(slot), ra.phase ^ 1u));
          const uint32_t dst = a_base + slot * p.a_slot_bytes;
          const int c0 = chunk * 64;
          const int kh = tap / 3, kw = tap - kh * 3;
          if (++chunk == p.cpt) {
            chunk = 0;
            ++tap;
          }
          if (!elect_one()) continue;
          arm_full(a_full(slot), p.a_tx_bytes);
          if (p.mode == CONV_MODE_FLAT) {
            ld2d(dst, &p.tmA, a_full(slot), grp * 64, m_tile * p.mblk * kTileM);
          } else {
            for (int mb = 0; mb < p.mblk; ++mb) {
              const MBlock t = mb ? tb1 : tb0;
              const uint32_t d = dst + mb * (kTileM * 128);
              if (p.taps == 1) {
                // 1x1 conv over [upsample2x(low-res tensor) | src]: the first up_chunks chunks are read from the
                // half-resolution tensor through tmUp = (c, dx, W/2, dy, H/2 * N) whose dx / dy dimensions have
                // stride 0, so the box {64, 2, tw/2, 2, th/2} lands as the tw x th tile of replicated pixels
                if (grp < p.up_chunks) ld5d(d, &p.tmUp, a_full(slot), c0, 0, t.x0 >> 1, 0, t.n * p.up_hs + (t.y0 >> 1));
                else ld4d(d, &p.tmA, a_full(slot), c0 - p.up_chunks * 64, t.x0, t.y0, t.n);
              } else if (p.stride == 1) {
                ld4d(d, &p.tmA, a_full(slot), c0, t.x0 + kw - 1, t.y0 + kh - 1, t.n);
              } else {
                // input row 2*Y + kh - 1: kh=0 -> (parity 1, Y-1), kh=1 -> (0, Y), kh=2 -> (1, Y); same in x
                const int yp = (kh + 1) & 1, dy = (kh == 0) ? -1 : 0;
                const int xp = (kw + 1) & 1, dx = (kw == 0) ? -1 : 0;
                ld5d(d, xp ? &p.tmA2 : &p.tmA, a_full(slot), c0, t.x0 + dx, yp, t.y0 + dy, t.n);
              }
            }
          }
        }
        if (lane == 0) trace_put<TRACE>(p, 0, it, tr0, trw, 0, TR_NOW());
      }
    }
  } else if (warp == kWarpB) {
    // =========================== weight producer ===========================
    if (EPI == EPI_CHAIN && elect_one()) {  // the chained conv's weights: resident, one barrier for all K groups
      const ChainBars cb{bar0};
      mbar_arrive_expect_tx(cb.b2_full(), (uint32_t)p.chain_kg * (uint32_t)p.tile_n2 * 128u);
      for (int g = 0; g < p.chain_kg; ++g) tma_load_2d(b2_base + (uint32_t)g * p.b2_slot_bytes, &p.tmB2, cb.b2_full(), g * 64, 0);
    }
    __syncwarp();
    if (p.b_res) {
      // all K steps of the (single) N tile stay resident: load once, never released
      const int n_res = p.b_dense ? p.nb : p.n_steps;
      for (int s = 0; s < n_res; ++s) {
        if (elect_one()) {
          arm_full(b_full(s), p.b_tx_bytes);
          // pair: this CTA keeps rows [rank * tile_n / 2, +tile_n / 2) of the tile (tmB's box is tile_n >> cshift rows)
          ld2d(b_base + s * p.b_slot_bytes, &p.tmB, b_full(s), (p.b_dense ? s : weight_chunk(p, s)) * 64,
               AQ_PAIR ? (int)(cluster_ctarank() * ((uint32_t)p.tile_n >> 1)) : 0);
        }
      }
    } else {
      Ring rb;
      int it = 0;
      for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), ++it) {
        int m_tile, n_idx;
        split_tile(p, tile, m_tile, n_idx);
        const int n0 = n_idx * p.tile_n;
        const long long tr0 = TR_NOW();
        long long trw = 0;
        for (int s = 0; s < p.n_steps; ++s, rb.advance(p.nb)) {
          const uint32_t slot = rb.slot;
          TR_WAIT(trw, mbar_wait(b_empty(slot), rb.phase ^ 1u));
          if (elect_one()) {
            arm_full(b_full(slot), p.b_tx_bytes);
            if (AQ_PAIR) {
              // this CTA's half of the tile's rows into ITS shared memory; the MMA reads both halves
              ld2d(b_base + slot * p.b_slot_bytes, &p.tmB, b_full(slot), weight_chunk(p, s) * 64,
                   n0 + (int)(cluster_ctarank() * ((uint32_t)p.tile_n >> 1)));
            } else if (CSHIFT(p)) {
              // this CTA's share of the tile's rows, multicast to every CTA of the cluster (tmB's box is
              // tile_n >> cshift rows); the other shares arrive from the peers on the same barrier
              const uint32_t r = cluster_ctarank();
              const uint32_t rows = (uint32_t)p.tile_n >> CSHIFT(p);
              tma_load_2d_mc(b_base + slot * p.b_slot_bytes + r * rows * 128u, &p.tmB, b_full(slot),
                             weight_chunk(p, s) * 64, n0 + (int)(r * rows), (uint16_t)((1u << (1u << CSHIFT(p))) - 1u));
            } else {
              tma_load_2d(b_base + slot * p.b_slot_bytes, &p.tmB, b_full(slot), weight_chunk(p, s) * 64, n0);
            }
          }
        }
        if (lane == 0) trace_put<TRACE>(p, 1, it, tr0, trw, 0, TR_NOW());
      }
    }
  } else if (warp == kWarpMma || warp == kWarpMma1) {
    // =========================== MMA issuers (one elected lane each) ===========================
    const int mi = warp == kWarpMma ? 0 : 1;
    if (mi < p.n_mma) {
      const MmaCtx mc{bar0, a_base, b_base, tmem_base, mi, total_tiles};
      const int n_my = p.mblk > p.n_mma ? 2 : 1;  // M blocks this issuer owns: mi, mi + n_mma
      if (p.mode == CONV_MODE_HALO) {
        if (p.b_res) { if (n_my == 2) mma_role<1, true, 2, TRACE>(p, mc); else mma_role<1, true, 1, TRACE>(p, mc); }
        else         { if (n_my == 2) mma_role<1, false, 2, TRACE>(p, mc); else mma_role<1, false, 1, TRACE>(p, mc); }
      } else if (p.mode == CONV_MODE_HALO2) {
        if (p.b_res) { if (n_my == 2) mma_role<2, true, 2, TRACE>(p, mc); else mma_role<2, true, 1, TRACE>(p, mc); }
        else         { if (n_my == 2) mma_role<2, false, 2, TRACE>(p, mc); else mma_role<2, false, 1, TRACE>(p, mc); }
      } else {
        if (p.b_res) mma_role<0, true, 1, TRACE>(p, mc);
        else mma_role<0, false, 1, TRACE>(p, mc);
      }
    }
  } else if (warp == kWarpStore) {
    if (EPI == EPI_STORE) store_role<TRACE>(p, stg_base, total_tiles, lane);
    if (EPI == EPI_CHAIN) store_chain_role(p, bar0, stg_base, total_tiles, lane);
  } else if (EPI == EPI_CHAIN && warp == kWarpMma2) {
    mma2_role(p, bar0, stg_base, b2_base, tmem_base, total_tiles);
  } else if (EPI == EPI_CHAIN && warp >= kWarpEpi0) {
    const EpiCtx ec{bar0, stg_base, tmem_base, smem_u32(s_bias), total_tiles, warp - kWarpEpi0, lane};
    epi_chain_role(p, ec, smem_u32(s_bias2));
  } else if (warp >= kWarpEpi0 && EPI == EPI_STORE) {
    // =========================== epilogue warps 4..19: bias + SiLU + TMA store ===========================
    const EpiCtx ec{bar0, stg_base, tmem_base, smem_u32(s_bias), total_tiles, warp - kWarpEpi0, lane};
    if (p.epi_sets == 2 && !TRACE) {
      if (p.act == ACT_SILU_FAST) epi_store_role2<ACT_SILU_FAST>(p, ec);
      else if (p.act == ACT_SILU) epi_store_role2<ACT_SILU>(p, ec);
      else epi_store_role2<ACT_NONE>(p, ec);
    } else if (p.act == ACT_SILU_FAST) epi_store_role<ACT_SILU_FAST, TRACE>(p, ec);
    else if (p.act == ACT_SILU) epi_store_role<ACT_SILU, TRACE>(p, ec);
    else epi_store_role<ACT_NONE, TRACE>(p, ec);
  } else if (warp >= kWarpEpi0) {
    // =========================== epilogue warps 4..19: Detect decode ===========================
    const uint32_t e = warp - kWarpEpi0;
    const uint32_t q = e & 3;            // TMEM lane quarter this warp may access (warp % 4)
    const uint32_t sub = e >> 2;         // which 16-column chunk of every 64-channel group (0..3)
    const uint32_t row = q * 32 + lane;  // accumulator row == TMEM lane == M row of the block
    const int n_chunks = p.tile_n >> 4;
    const float conf_thr = p.det_mask ? __ldg(p.det_conf) : 0.f;
    Ring rt;
    int it = 0;
    for (int tile = tile_first(p); tile < total_tiles; tile += tile_step(p), ++it, rt.advance(p.acc_stages)) {
      int m_tile, n_idx;
      split_tile(p, tile, m_tile, n_idx);
      const int as = rt.slot;
      const long long tr0 = TR_NOW();
      mbar_wait(tfull(as), rt.phase);
      const long long tr1 = TR_NOW();
      tc_fence_after();
      for (int mb = 0; mb < p.mblk; ++mb) {
        const uint32_t t_addr = tmem_base + ((q * 32u) << 16) + as * p.acc_stride + mb * p.acc_pitch;
        {
          // ---- Detect decode: row = one grid cell, columns = 3 anchors x (5 + nc)
          const MBlock t = mblock_coords(p, m_tile, mb);
          const int m = t.m0 + (int)row;
          const int cells = p.det_ny * p.det_nx;
          const int b = m / cells;
          const int pos = m - b * cells;
          const int gy = pos / p.det_nx;
          const int gx = pos - gy * p.det_nx;
          const bool valid = m < p.det_m_total;
          const int ncols = 3 * p.det_no;
          if (p.det_no == 8) {
            // nc = 3 (the reference's model): a 16-column chunk is exactly two anchors x 8 outputs, so the column ->
            // (anchor, output) mapping is static: no table lookups, no per-element branches.  With two M blocks per
            // tile and two chunks (24 columns) every one of the 16 warps takes one (M block, chunk) pair.
            const bool split = p.mblk == 2 && n_chunks == 2;
            if (split && (int)(sub >> 1) != mb) goto det_block_done;
            for (int ch = split ? (int)(sub & 1) : (int)sub; ch < n_chunks; ch += split ? 2 : 4) {
              const int c0 = ch << 4;
              uint32_t v[16];
              tmem_ld16(t_addr + c0, v);
              tmem_ld_wait();
              const float fx = ((float)gx - 0.5f), fy = ((float)gy - 0.5f);
#pragma unroll
              for (int aa = 0; aa < 2; ++aa) {
                const int a = (c0 >> 3) + aa;
                if (a >= 3) continue;
                float sg[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float z = __uint_as_float(v[8 * aa + j]) + s_bias[c0 + 8 * aa + j];
                  sg[j] = __fdividef(1.0f, 1.0f + __expf(-z));
                }
                const float aw = p.det_anchor[2 * a], ah = p.det_anchor[2 * a + 1];
                const float4 lo = make_float4((sg[0] * 2.0f + fx) * p.det_stride, (sg[1] * 2.0f + fy) * p.det_stride,
                                              (sg[2] * 2.0f) * (sg[2] * 2.0f) * aw, (sg[3] * 2.0f) * (sg[3] * 2.0f) * ah);
                if (valid) {
                  const int ri = p.det_level_off + a * cells + pos;  // row inside the image's prediction
                  const size_t r = (size_t)b * p.det_rows + ri;
                  float4* dst = reinterpret_cast<float4*>(p.pred + r * 8);
                  dst[0] = lo;
                  dst[1] = make_float4(sg[4], sg[5], sg[6], sg[7]);
                  if (p.det_mask) {
                    const float best = fmaxf(fmaxf(__fmul_rn(sg[5], sg[4]), __fmul_rn(sg[6], sg[4])), __fmul_rn(sg[7], sg[4]));
                    if (sg[4] > conf_thr && best > conf_thr)  // rare (~0.2 % of the rows on sparse tiles)
                      atomicOr(p.det_mask + (size_t)b * p.det_mask_words + (ri >> 5), 1u << (ri & 31));
                  }
                }
              }
            }
          } else
          for (int ch = sub; ch < n_chunks; ch += 4) {
            const int c0 = ch << 4;
            uint32_t v[16];
            tmem_ld16(t_addr + c0, v);
            tmem_ld_wait();
            float o[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int col = c0 + i;
              const int a = s_det_a[col];
              const int oi = s_det_o[col];
              const float z = __uint_as_float(v[i]) + s_bias[col < kMaxCout ? col : 0];
              const float s = __fdividef(1.0f, 1.0f + __expf(-z));
              float out;
              if (oi == 0) out = (s * 2.0f + ((float)gx - 0.5f)) * p.det_stride;
              else if (oi == 1) out = (s * 2.0f + ((float)gy - 0.5f)) * p.det_stride;
              else if (oi == 2) out = (s * 2.0f) * (s * 2.0f) * p.det_anchor[a < 3 ? 2 * a : 0];
              else if (oi == 3) out = (s * 2.0f) * (s * 2.0f) * p.det_anchor[a < 3 ? 2 * a + 1 : 0];
              else out = s;
              o[i] = out;
            }
            if (valid) {
              if (p.det_no == 8) {
                // two anchors per 16-column chunk, 8 floats = 32 contiguous bytes per (anchor, cell)
#pragma unroll
                for (int aa = 0; aa < 2; ++aa) {
                  const int a = (c0 >> 3) + aa;
                  if (a < 3) {
                    const size_t r = (size_t)b * p.det_rows + p.det_level_off + (size_t)a * cells + pos;
                    float4* dst = reinterpret_cast<float4*>(p.pred + r * 8);
                    dst[0] = make_float4(o[8 * aa + 0], o[8 * aa + 1], o[8 * aa + 2], o[8 * aa + 3]);
                    dst[1] = make_float4(o[8 * aa + 4], o[8 * aa + 5], o[8 * aa + 6], o[8 * aa + 7]);
                  }
                }
              } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                  const int col = c0 + i;
                  if (col < ncols) {
                    const int a = s_det_a[col];
                    const int oi = s_det_o[col];
                    const size_t r = (size_t)b * p.det_rows + p.det_level_off + (size_t)a * cells + pos;
                    p.pred[r * p.det_no + oi] = o[i];
                  }
                }
              }
            }
          }
        det_block_done:
          if (mb == p.mblk - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) arrive_tempty(tempty(as));
          }
        }
      }
      if (e == 0 && lane == 0) trace_put<TRACE>(p, 3, it, tr0, tr1 - tr0, 0, TR_NOW());
    }
  }

  // teardown: everyone done with TMEM before the allocating warp frees it; in a cluster no CTA may exit while a
  // peer can still multicast into its shared memory or arrive on its barriers
  tc_fence_before();
  __syncthreads();
  if (CSHIFT(p)) cluster_sync_all();
  if (warp == kWarpMma) {
    tc_fence_after();
    if (AQ_PAIR) tmem_dealloc2(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

}  // namespace

#if AQ_PAIR
#define AQ_VARIANT(x) x##_pair
#elif AQ_CLUSTER
#define AQ_VARIANT(x) x##_cl
#else
#define AQ_VARIANT(x) x##_nocl
#endif
cudaError_t conv_kernels_init_cl();
cudaError_t conv_kernels_init_nocl();
cudaError_t conv_kernels_init_pair();
cudaError_t launch_conv_cl(const ConvKParams& p, int epi, int num_sms, int pdl, cudaStream_t stream);
cudaError_t launch_conv_nocl(const ConvKParams& p, int epi, int num_sms, int pdl, cudaStream_t stream);
cudaError_t launch_conv_pair(const ConvKParams& p, int epi, int num_sms, int pdl, cudaStream_t stream);

#if !AQ_CLUSTER && !AQ_PAIR
size_t conv_smem_bytes(const ConvKParams& p) {
  return (size_t)p.na * p.a_slot_bytes + (size_t)p.nb * p.b_slot_bytes + (size_t)p.n_stg * p.stg_bytes + 1024 +
         (p.chain ? (size_t)p.chain_kg * p.b2_slot_bytes : 0);
}
cudaError_t conv_kernels_init() {
  cudaError_t e = conv_kernels_init_nocl();
  if (e == cudaSuccess) e = conv_kernels_init_cl();
  return e == cudaSuccess ? conv_kernels_init_pair() : e;
}
cudaError_t launch_conv(const ConvKParams& p, int epi, int num_sms, int pdl, cudaStream_t stream) {
  if (p.pair) return launch_conv_pair(p, epi, num_sms, pdl, stream);
  return p.cshift ? launch_conv_cl(p, epi, num_sms, pdl, stream) : launch_conv_nocl(p, epi, num_sms, pdl, stream);
}
#endif

cudaError_t AQ_VARIANT(conv_kernels_init)() {
  cudaError_t e = cudaSuccess;
  const void* fns[5] = {(const void*)conv_igemm_kernel<EPI_STORE, false>, (const void*)conv_igemm_kernel<EPI_DETECT, false>,
                        (const void*)conv_igemm_kernel<EPI_STORE, true>, (const void*)conv_igemm_kernel<EPI_DETECT, true>,
                        (const void*)conv_igemm_kernel<EPI_CHAIN, false>};
  for (int i = 0; i < 5 && e == cudaSuccess; ++i)
    e = cudaFuncSetAttribute(fns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, kConvMaxDynSmem);
  return e;
}

cudaError_t AQ_VARIANT(launch_conv)(const ConvKParams& p, int epi, int num_sms, int pdl, cudaStream_t stream) {
  const int cluster = 1 << p.cshift;
  const int max_clusters = num_sms / cluster;
  const int n_clusters = p.total_tiles < max_clusters ? p.total_tiles : max_clusters;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_clusters * cluster);
  cfg.blockDim = dim3(epi == EPI_CHAIN ? kConvThreadsChain : kConvThreads);
  cfg.dynamicSmemBytes = conv_smem_bytes(p);
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cluster > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cluster;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  if (p.trace) {
    if (epi == EPI_STORE) return cudaLaunchKernelEx(&cfg, conv_igemm_kernel<EPI_STORE, true>, p);
    return cudaLaunchKernelEx(&cfg, conv_igemm_kernel<EPI_DETECT, true>, p);
  }
  if (epi == EPI_STORE) return cudaLaunchKernelEx(&cfg, conv_igemm_kernel<EPI_STORE, false>, p);
  if (epi == EPI_CHAIN) return cudaLaunchKernelEx(&cfg, conv_igemm_kernel<EPI_CHAIN, false>, p);
  return cudaLaunchKernelEx(&cfg, conv_igemm_kernel<EPI_DETECT, false>, p);
}

}  // namespace aq
