// Core of the parallel (self-synchronising) Huffman decoder for baseline JPEG scans -- compiled for the device
// (jpeg_huff.cu) and, unchanged, for the host (jpeg_host.cpp: aq_jpeg_debug_parallel_decode, the CPU rehearsal of the
// same schedule used by the CPU tests).  SURVEY.md 8f-1.
//
// A scan is one long bit string of variable-length codes; a decoder is the state (bit position p, block slot s inside
// the MCU, zigzag index z; z == 0 means "the next symbol is a DC size").  The string is cut into subsequences of
// `sub_bits` bits.  A decoder started at a subsequence boundary in a guessed state produces garbage at first but, because
// Huffman codes are self-synchronising, usually falls into step with the true decoder within a few dozen symbols.  The
// schedule (Klein & Wiseman 2003; Weissenberger & Schmidt 2018 for GPUs):
//   1. every subsequence is decoded from the guess (p = its first bit, s = 0, z = 0); its end state is recorded;
//   2. relaxation: subsequence i is re-decoded from the end state of i-1 whenever that differs from the state it was
//      last started from, until nothing changes (the first subsequence starts from the true state, so the fixed point
//      is the sequential decoder's trajectory);
//   3. an exclusive scan over the per-subsequence counts of completed blocks gives every subsequence its first block;
//   4. every subsequence is decoded once more, now writing coefficients (DC still as differences);
//   5. a scan per component over the DC differences undoes the prediction.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define AQ_HD __host__ __device__ __forceinline__
#else
#define AQ_HD inline
#endif

namespace aqj {

// One Huffman table in the form both decoders read: 9-bit first-level lookup ((length << 8) | symbol, 0 = longer code)
// and the canonical-code arrays of T.81 Annex F.2.2.3 for the rest.
struct DevHuff {
  uint16_t look[512];
  int32_t maxcode[18];
  int32_t valoff[17];
  uint8_t vals[256];
  uint8_t pad[20];
};
static_assert(sizeof(DevHuff) == 1440, "DevHuff layout is part of the C ABI (AQ_JPEG_TABLE_BYTES)");
constexpr int kTablesPerTile = 6;  // component c: [2c] = DC table, [2c + 1] = AC table

struct State {  // packed into 64 bits for single-store updates
  uint32_t p;
  int s, z;
};
AQ_HD uint64_t pack_state(const State& st) { return (uint64_t)st.p | ((uint64_t)st.s << 32) | ((uint64_t)st.z << 40); }
AQ_HD State unpack_state(uint64_t v) { return State{(uint32_t)v, (int)((v >> 32) & 255), (int)((v >> 40) & 255)}; }

AQ_HD uint32_t be32(uint32_t w) {
#if defined(__CUDA_ARCH__)
  return __byte_perm(w, 0, 0x0123);
#else
  return __builtin_bswap32(w);
#endif
}

// Sliding 64-bit view of the stream (the buffer is padded with zero words past the end): one 32-bit load per 32 bits
// consumed instead of two per symbol.  Positions only move forward.
struct BitWindow {
  const uint32_t* words;
  uint32_t wi;   // acc holds words[wi] (high half) and words[wi + 1] (low half)
  uint64_t acc;
  AQ_HD void init(const uint32_t* w, uint32_t p) {
    words = w;
    wi = p >> 5;
    acc = ((uint64_t)be32(w[wi]) << 32) | be32(w[wi + 1]);
  }
  // 32 bits of the stream starting at bit p
  AQ_HD uint32_t get(uint32_t p) {
    uint32_t rel = p - (wi << 5);
    while (rel >= 32) {
      ++wi;
      acc = (acc << 32) | be32(words[wi + 1]);
      rel -= 32;
    }
    return (uint32_t)((acc << rel) >> 32);
  }
};

AQ_HD int extend(uint32_t v, int size) { return v < (1u << (size - 1)) ? (int)v - (1 << size) + 1 : (int)v; }

// Decodes the symbol at st.p and advances the state.  emit(zigzag position, value) is called for a coded coefficient
// (position 0 = DC difference); returns true when the symbol completed a block (st.s then names the next slot).
// slot_comp: component of MCU slot s in nibble s.  Written with selects rather than branches: the lanes of a warp sit in
// different places of their blocks (DC size, coefficient, end-of-block, zero run), and one common instruction stream
// costs less than the four diverged ones; only the rare long-code search (> 9 bits) branches.
template <class Emit>
AQ_HD bool decode_one(BitWindow& bits, const DevHuff* tabs, uint32_t slot_comp, int bpm, State& st, Emit&& emit) {
  const uint32_t w = bits.get(st.p);
  const bool is_dc = st.z == 0;
  const DevHuff& t = tabs[((slot_comp >> (4 * st.s)) & 15) * 2 + (is_dc ? 0 : 1)];
  const uint32_t lk = t.look[w >> 23];
  int len = (int)(lk >> 8), sym = (int)(lk & 255);
  if (lk == 0) {
    const int c16 = (int)(w >> 16);
    int code = 0;
    for (len = 10; len <= 16; ++len) {
      code = c16 >> (16 - len);
      if (code <= t.maxcode[len]) break;
    }
    if (len > 16) {  // no code matches (only possible while out of step): any deterministic rule that advances will do
      len = 16;
      sym = 0;
    } else {
      sym = t.vals[(code + t.valoff[len]) & 255];
    }
  }
  const int run = is_dc ? 0 : sym >> 4;
  const int size = is_dc ? (sym > 16 ? 16 : sym) : (sym & 15);
  const bool no_coef = !is_dc && size == 0;             // end of block, or a run of 16 zeros
  const int pos = st.z + run;
  if (size != 0 && pos < 64) emit(pos, extend((w << len) >> (32 - size), size));
  st.z = no_coef ? (run == 15 ? st.z + 16 : 64) : pos + 1;
  st.p += (uint32_t)(len + size);
  const bool done = st.z >= 64;
  st.z = done ? 0 : st.z;
  st.s = done ? (st.s + 1 == bpm ? 0 : st.s + 1) : st.s;
  return done;
}

// Decodes every symbol that STARTS before bit `limit`; returns the number of blocks completed.
AQ_HD int decode_range(const uint32_t* words, const DevHuff* tabs, uint32_t slot_comp, int bpm, State& st, uint32_t limit) {
  int nb = 0;
  BitWindow bits;
  bits.init(words, st.p);
  while (st.p < limit) nb += decode_one(bits, tabs, slot_comp, bpm, st, [](int, int) {}) ? 1 : 0;
  return nb;
}

// Component of every MCU slot, one nibble per slot: luma blocks first, then Cb, Cr.
AQ_HD uint32_t pack_slot_comp(int ncomp, int hs, int vs) {
  if (ncomp == 1) return 0;
  uint32_t v = 0;
  const int nl = hs * vs;
  for (int s = 0; s < nl + 2 && s < 8; ++s) v |= (uint32_t)(s < nl ? 0 : s - nl + 1) << (4 * s);
  return v;
}

// Geometry shared by the writers: where block g (decode order) of a tile lives in the dense coefficient planes.
struct Geo {
  int hs, vs, ncomp, bpm, mcus_x, mcus_y, luma_bw;
  long long luma_blocks, chroma_blocks, total_blocks;
};
AQ_HD long long block_address(const Geo& g, long long blk) {
  const long long mcu = blk / g.bpm;
  const int slot = (int)(blk - mcu * g.bpm);
  const int my = (int)(mcu / g.mcus_x), mx = (int)(mcu - (long long)my * g.mcus_x);
  const int nl = g.hs * g.vs;
  if (slot < nl) return (long long)(my * g.vs + slot / g.hs) * g.luma_bw + mx * g.hs + slot % g.hs;
  return g.luma_blocks + (long long)(slot - nl) * g.chroma_blocks + (long long)my * g.mcus_x + mx;
}

}  // namespace aqj
