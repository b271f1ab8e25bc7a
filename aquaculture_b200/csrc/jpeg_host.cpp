// Host half of the hybrid JPEG decoder: marker parsing + Huffman entropy decoding of baseline
// (SOF0 / 8-bit SOF1) JPEG files into quantised DCT coefficients.  The arithmetic half (dequantisation, inverse DCT,
// chroma upsampling, YCbCr -> RGB) runs on the GPU (jpeg.cu).  Replaces the entropy-decoding part of the cv2.imread
// call inside upstream utils/dataloaders.py::LoadImages.__next__ [UPSTREAM-RECALLED]; the files are the tiles written
// by gdal.Translate -of JPEG [REF src/load_data/tile_tifs.py:74].  Algorithm: ITU-T T.81 Annex F.2 (sequential
// Huffman decoding), table construction as Annex C; no code from libjpeg is used.
#include <stdint.h>
#include <string.h>

#include <atomic>
#include <thread>
#include <vector>

#include "../../include/aqua_b200.h"
#include "jpeg_huff_core.h"

namespace {

const uint8_t kZigzagToNatural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                      41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                      30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

constexpr int kLookBits = 9;

struct HuffTable {
  bool present = false;
  uint8_t bits[17];
  uint8_t vals[256];
  // Annex F.2.2.3 decoding tables: for code length l, codes run mincode[l] .. maxcode[l]
  int32_t maxcode[18];
  int32_t valoff[17];
  uint16_t look[1 << kLookBits];  // (length << 8) | symbol for codes of <= kLookBits bits, 0 = longer code
  // AC tables only: code and magnitude bits resolved by ONE lookup when both fit in kLookBits bits:
  // (value << 16) | (run << 8) | total bits, 0 = take the general path
  int32_t fast_ac[1 << kLookBits];
};

bool build_table(HuffTable& t) {
  int n = 0;
  for (int l = 1; l <= 16; ++l) n += t.bits[l];
  if (n > 256) return false;
  uint32_t code = 0;
  int p = 0;
  memset(t.look, 0, sizeof t.look);
  t.maxcode[0] = -1;
  t.valoff[0] = 0;
  memset(t.vals + n, 0, sizeof t.vals - (size_t)n);
  for (int l = 1; l <= 16; ++l) {
    if (t.bits[l]) {
      // over-subscribed lengths are rejected BEFORE the lookup table is written: `first + f` below would run far
      // past `look` for a crafted DHT (e.g. bits[1] = 255)
      if (code + t.bits[l] > (1u << l)) return false;
      t.valoff[l] = p - (int32_t)code;
      for (int i = 0; i < t.bits[l]; ++i, ++p, ++code) {
        if (l <= kLookBits) {
          const uint32_t first = code << (kLookBits - l);
          for (uint32_t f = 0; f < (1u << (kLookBits - l)); ++f) t.look[first + f] = (uint16_t)((l << 8) | t.vals[p]);
        }
      }
      t.maxcode[l] = (int32_t)code - 1;
      if (code > (1u << l)) return false;  // over-subscribed
    } else {
      t.maxcode[l] = -1;
      t.valoff[l] = 0;
    }
    code <<= 1;
  }
  t.maxcode[17] = 0x7fffffff;
  for (int i = 0; i < (1 << kLookBits); ++i) {
    t.fast_ac[i] = 0;
    const uint32_t lk = t.look[i];
    if (!lk) continue;
    const int len = lk >> 8, rs = lk & 0xFF, run = rs >> 4, mag = rs & 15;
    if (mag == 0 || len + mag > kLookBits) continue;
    int v = (i >> (kLookBits - len - mag)) & ((1 << mag) - 1);
    if (v < (1 << (mag - 1))) v -= (1 << mag) - 1;  // Annex F.2.2.1 EXTEND
    t.fast_ac[i] = (int32_t)((uint32_t)v << 16) | (run << 8) | (len + mag);
  }
  t.present = true;
  return true;
}

struct Component {
  int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
};

struct Frame {
  int width = 0, height = 0, ncomp = 0;
  Component comp[3];
  uint16_t quant[4][64];  // natural order
  bool quant_present[4] = {false, false, false, false};
  HuffTable dc[4], ac[4];
  int restart_interval = 0;
  bool jfif = false, adobe = false;
  int adobe_transform = 0;
  size_t scan_off = 0;  // first entropy-coded byte
  bool progressive = false, have_sof = false, have_sos = false;
  int sof_marker = 0;
};

inline int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

// Walks the marker segments up to (and including) the first SOS header.
int parse_headers(const uint8_t* d, size_t len, Frame& f) {
  if (len < 4 || d[0] != 0xFF || d[1] != 0xD8) return AQ_ERR_ARG;
  size_t p = 2;
  while (p + 4 <= len) {
    if (d[p] != 0xFF) return AQ_ERR_ARG;
    while (p < len && d[p] == 0xFF) ++p;  // fill bytes
    if (p >= len) return AQ_ERR_ARG;
    const int m = d[p++];
    if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;  // standalone markers
    if (m == 0xD9) return AQ_ERR_ARG;                                  // EOI before any scan
    if (p + 2 > len) return AQ_ERR_ARG;
    const int seg = be16(d + p);
    if (seg < 2 || p + seg > len) return AQ_ERR_ARG;
    const uint8_t* s = d + p + 2;
    const int n = seg - 2;
    if (m == 0xDB) {  // DQT
      int q = 0;
      while (q < n) {
        const int pq = s[q] >> 4, tq = s[q] & 15;
        ++q;
        if (tq > 3 || q + (pq ? 128 : 64) > n) return AQ_ERR_ARG;
        for (int i = 0; i < 64; ++i) {
          f.quant[tq][kZigzagToNatural[i]] = pq ? (uint16_t)be16(s + q + 2 * i) : s[q + i];
        }
        q += pq ? 128 : 64;
        f.quant_present[tq] = true;
      }
    } else if (m == 0xC4) {  // DHT
      int q = 0;
      while (q < n) {
        if (q + 17 > n) return AQ_ERR_ARG;
        const int tc = s[q] >> 4, th = s[q] & 15;
        if (tc > 1 || th > 3) return AQ_ERR_ARG;
        HuffTable& t = tc ? f.ac[th] : f.dc[th];
        int cnt = 0;
        t.bits[0] = 0;
        for (int l = 1; l <= 16; ++l) cnt += (t.bits[l] = s[q + l]);
        q += 17;
        if (cnt > 256 || q + cnt > n) return AQ_ERR_ARG;
        memcpy(t.vals, s + q, cnt);
        q += cnt;
        if (!build_table(t)) return AQ_ERR_ARG;
      }
    } else if (m == 0xDD) {  // DRI
      if (n < 2) return AQ_ERR_ARG;
      f.restart_interval = be16(s);
    } else if (m == 0xE0) {
      if (n >= 5 && !memcmp(s, "JFIF", 5)) f.jfif = true;
    } else if (m == 0xEE) {
      if (n >= 12 && !memcmp(s, "Adobe", 5)) {
        f.adobe = true;
        f.adobe_transform = s[11];
      }
    } else if (m >= 0xC0 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {  // SOFn
      f.sof_marker = m;
      if (m != 0xC0 && m != 0xC1) {
        f.progressive = true;  // progressive / lossless / arithmetic: no kernel path for these
        return AQ_ERR_UNSUPPORTED;
      }
      if (n < 6) return AQ_ERR_ARG;
      if (s[0] != 8) return AQ_ERR_UNSUPPORTED;  // 12-bit samples
      f.height = be16(s + 1);
      f.width = be16(s + 3);
      f.ncomp = s[5];
      if (f.width <= 0 || f.height <= 0) return AQ_ERR_UNSUPPORTED;  // DNL-defined height
      if (f.ncomp != 1 && f.ncomp != 3) return AQ_ERR_UNSUPPORTED;   // CMYK / YCCK
      if (n < 6 + 3 * f.ncomp) return AQ_ERR_ARG;
      for (int i = 0; i < f.ncomp; ++i) {
        f.comp[i].id = s[6 + 3 * i];
        f.comp[i].h = s[7 + 3 * i] >> 4;
        f.comp[i].v = s[7 + 3 * i] & 15;
        f.comp[i].tq = s[8 + 3 * i];
        if (f.comp[i].tq > 3) return AQ_ERR_ARG;
      }
      f.have_sof = true;
    } else if (m == 0xDA) {  // SOS
      if (!f.have_sof) return AQ_ERR_ARG;
      if (n < 1) return AQ_ERR_ARG;
      const int ns = s[0];
      if (ns != f.ncomp) return AQ_ERR_UNSUPPORTED;  // one scan per component (non-interleaved colour)
      if (n < 1 + 2 * ns + 3) return AQ_ERR_ARG;
      for (int i = 0; i < ns; ++i) {
        if (s[1 + 2 * i] != f.comp[i].id) return AQ_ERR_UNSUPPORTED;
        f.comp[i].td = s[2 + 2 * i] >> 4;
        f.comp[i].ta = s[2 + 2 * i] & 15;
        if (f.comp[i].td > 3 || f.comp[i].ta > 3) return AQ_ERR_ARG;
      }
      f.scan_off = p + seg;
      f.have_sos = true;
      return AQ_OK;
    }
    p += seg;
  }
  return AQ_ERR_ARG;
}

int validate(const Frame& f, aq_jpeg_info* info) {
  if (!f.have_sof || !f.have_sos) return AQ_ERR_ARG;
  int hs = f.comp[0].h, vs = f.comp[0].v;
  if (f.ncomp == 1) {
    hs = vs = 1;  // a single-component scan is never interleaved: one block per MCU whatever the factors say
  } else {
    if (f.comp[1].h != 1 || f.comp[1].v != 1 || f.comp[2].h != 1 || f.comp[2].v != 1) return AQ_ERR_UNSUPPORTED;
    if (!((hs == 1 && vs == 1) || (hs == 2 && vs == 2))) return AQ_ERR_UNSUPPORTED;  // 4:4:4 and 4:2:0 only
    // colour space as libjpeg infers it: JFIF -> YCbCr; Adobe transform 0 -> RGB; ids 'R','G','B' -> RGB
    if (!f.jfif && f.adobe && f.adobe_transform != 1) return AQ_ERR_UNSUPPORTED;
    if (!f.jfif && !f.adobe && f.comp[0].id == 'R' && f.comp[1].id == 'G' && f.comp[2].id == 'B') return AQ_ERR_UNSUPPORTED;
  }
  for (int i = 0; i < f.ncomp; ++i) {
    if (!f.quant_present[f.comp[i].tq]) return AQ_ERR_ARG;
    if (!f.dc[f.comp[i].td].present || !f.ac[f.comp[i].ta].present) return AQ_ERR_ARG;
  }
  if (info) {
    info->width = f.width;
    info->height = f.height;
    info->ncomp = f.ncomp;
    info->hs = hs;
    info->vs = vs;
    info->mcus_x = (f.width + 8 * hs - 1) / (8 * hs);
    info->mcus_y = (f.height + 8 * vs - 1) / (8 * vs);
    info->restart_interval = f.restart_interval;
    const size_t luma = (size_t)info->mcus_x * hs * info->mcus_y * vs;
    const size_t chroma = f.ncomp == 3 ? 2 * (size_t)info->mcus_x * info->mcus_y : 0;
    info->coef_count = (luma + chroma) * 64;
  }
  return AQ_OK;
}

// Entropy-coded segment reader: byte stuffing removed, markers stop the stream (zeros are fed after one).
struct BitReader {
  const uint8_t* d;
  size_t len, pos;
  uint64_t acc = 0;
  int nbits = 0;
  bool hit_marker = false;
  inline void fill() {
    if (!hit_marker && pos + 8 <= len) {
      uint64_t v;
      memcpy(&v, d + pos, 8);
      const uint64_t x = ~v;  // a 0xFF byte becomes 0x00
      if (!((x - 0x0101010101010101ULL) & ~x & 0x8080808080808080ULL)) {
        // no 0xFF among the next eight bytes: append as many whole bytes as fit.  The bits of a partly fitting byte
        // that land below `nbits` are the stream's own next bits, so OR-ing the same byte again later is harmless.
        const int take = (64 - nbits) >> 3;
        acc |= __builtin_bswap64(v) >> nbits;
        pos += take;
        nbits += take * 8;
        return;
      }
    }
    while (nbits <= 56) {
      uint32_t b = 0;
      if (!hit_marker && pos < len) {
        b = d[pos];
        if (b == 0xFF) {
          if (pos + 1 < len && d[pos + 1] == 0x00) {
            pos += 2;
          } else {
            hit_marker = true;  // leave pos on the 0xFF
            b = 0;
          }
        } else {
          ++pos;
        }
      } else {
        hit_marker = true;
      }
      acc |= (uint64_t)b << (56 - nbits);
      nbits += 8;
    }
  }
  inline uint32_t peek(int n) { return (uint32_t)(acc >> (64 - n)); }
  inline void skip(int n) {
    acc <<= n;
    nbits -= n;
  }
  inline void reset() {
    acc = 0;
    nbits = 0;
    hit_marker = false;
  }
};

inline int decode_symbol(BitReader& br, const HuffTable& t) {
  if (br.nbits < 16) br.fill();
  const uint32_t lk = t.look[br.peek(kLookBits)];
  if (lk) {
    br.skip(lk >> 8);
    return lk & 0xFF;
  }
  const uint32_t w = br.peek(16);
  for (int l = kLookBits + 1; l <= 16; ++l) {
    const int32_t code = (int32_t)(w >> (16 - l));
    if (code <= t.maxcode[l]) {
      br.skip(l);
      return t.vals[(code + t.valoff[l]) & 0xFF];
    }
  }
  br.skip(16);
  return -1;  // corrupt
}

// Annex F.2.2.1 EXTEND
inline int receive_extend(BitReader& br, int s) {
  if (br.nbits < s) br.fill();
  const int v = (int)br.peek(s);
  br.skip(s);
  return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v;
}

// Sinks: where a decoded coefficient goes.  Dense = the block's 64-entry natural-order array (pre-zeroed);
// packed = a list of (position << 16 | value) words, the form shipped to the GPU (most coefficients are zero).
struct DenseSink {
  int16_t* blk;
  inline void put(int natural_pos, int v) { blk[natural_pos] = (int16_t)v; }
};
struct PackedSink {
  uint32_t* out;
  uint32_t n = 0;
  inline void put(int natural_pos, int v) { out[n++] = ((uint32_t)natural_pos << 16) | (uint16_t)(int16_t)v; }
};

template <class Sink>
inline bool decode_block(BitReader& br, const HuffTable& dc, const HuffTable& ac, int& pred, Sink& sink) {
  int s = decode_symbol(br, dc);
  if (s < 0 || s > 16) return false;
  if (s) pred += receive_extend(br, s);
  if (pred) sink.put(0, pred);
  for (int k = 1; k < 64;) {
    if (br.nbits < 16) br.fill();
    const int32_t fa = ac.fast_ac[br.peek(kLookBits)];
    if (fa) {  // run, size and magnitude bits in one lookup
      k += (fa >> 8) & 15;
      if (k > 63) return false;
      br.skip(fa & 0xFF);
      sink.put(kZigzagToNatural[k], fa >> 16);
      ++k;
      continue;
    }
    const int rs = decode_symbol(br, ac);
    if (rs < 0) return false;
    const int r = rs >> 4;
    s = rs & 15;
    if (s) {
      k += r;
      if (k > 63) return false;
      sink.put(kZigzagToNatural[k], receive_extend(br, s));
      ++k;
    } else {
      if (r != 15) break;  // EOB
      k += 16;
    }
  }
  return true;
}

// Walks the MCUs of the one interleaved scan; `visit(plane, block index in the plane)` returns the sink of that block.
template <class Visit>
int decode_scan(const uint8_t* data, size_t len, const Frame& f, const aq_jpeg_info& info, Visit&& visit) {
  const int hs = info.hs, vs = info.vs;
  const size_t luma_bw = (size_t)info.mcus_x * hs;
  BitReader br{data, len, f.scan_off};
  int pred[3] = {0, 0, 0};
  int until_restart = f.restart_interval, next_rst = 0;
  const HuffTable* dct[3];
  const HuffTable* act[3];
  for (int c = 0; c < f.ncomp; ++c) {
    dct[c] = &f.dc[f.comp[c].td];
    act[c] = &f.ac[f.comp[c].ta];
  }
  for (int my = 0; my < info.mcus_y; ++my) {
    for (int mx = 0; mx < info.mcus_x; ++mx) {
      if (f.restart_interval && until_restart == 0) {
        // byte-align, expect RSTn; DC predictions restart at 0 (T.81 F.2.1.3.1)
        br.reset();
        size_t p = br.pos;
        while (p + 1 < len && !(data[p] == 0xFF && data[p + 1] >= 0xD0 && data[p + 1] <= 0xD7)) ++p;
        if (p + 1 >= len || data[p + 1] != 0xD0 + next_rst) return AQ_ERR_IO;
        br.pos = p + 2;
        next_rst = (next_rst + 1) & 7;
        pred[0] = pred[1] = pred[2] = 0;
        until_restart = f.restart_interval;
      }
      for (int by = 0; by < vs; ++by)
        for (int bx = 0; bx < hs; ++bx)
          if (!visit(0, ((size_t)my * vs + by) * luma_bw + (size_t)mx * hs + bx, br, *dct[0], *act[0], pred[0])) return AQ_ERR_IO;
      for (int c = 1; c < f.ncomp; ++c)
        if (!visit(c, (size_t)my * info.mcus_x + mx, br, *dct[c], *act[c], pred[c])) return AQ_ERR_IO;
      --until_restart;
    }
  }
  return AQ_OK;
}

int open_file(const uint8_t* data, size_t len, Frame& f, aq_jpeg_info& info, uint16_t* quant) {
  int rc = parse_headers(data, len, f);
  if (rc != AQ_OK) return rc;
  rc = validate(f, &info);
  if (rc != AQ_OK) return rc;
  for (int c = 0; c < 3; ++c) memcpy(quant + 64 * c, f.quant[f.comp[c < f.ncomp ? c : 0].tq], 64 * sizeof(uint16_t));
  return AQ_OK;
}

}  // namespace

extern "C" int aq_jpeg_probe(const uint8_t* data, size_t len, aq_jpeg_info* info) {
  if (!data || !info) return AQ_ERR_ARG;
  Frame f;
  const int rc = parse_headers(data, len, f);
  if (rc != AQ_OK) return rc;
  return validate(f, info);
}

extern "C" int aq_jpeg_entropy_decode(const uint8_t* data, size_t len, int16_t* coef, size_t coef_cap, uint16_t* quant) {
  if (!data || !coef || !quant) return AQ_ERR_ARG;
  Frame f;
  aq_jpeg_info info;
  const int rc = open_file(data, len, f, info, quant);
  if (rc != AQ_OK) return rc;
  if (coef_cap < info.coef_count) return AQ_ERR_ARG;
  memset(coef, 0, info.coef_count * sizeof(int16_t));
  const size_t luma_blocks = (size_t)info.mcus_x * info.hs * info.mcus_y * info.vs;
  const size_t chroma_blocks = (size_t)info.mcus_x * info.mcus_y;
  int16_t* plane[3] = {coef, coef + luma_blocks * 64, coef + (luma_blocks + chroma_blocks) * 64};
  return decode_scan(data, len, f, info, [&](int c, size_t b, BitReader& br, const HuffTable& dc, const HuffTable& ac, int& pred) {
    DenseSink sink{plane[c] + b * 64};
    return decode_block(br, dc, ac, pred, sink);
  });
}

extern "C" int aq_jpeg_entropy_decode_packed(const uint8_t* data, size_t len, uint32_t* block_desc, size_t block_cap,
                                             uint32_t* entries, size_t entries_cap, size_t* n_entries, uint16_t* quant) {
  if (!data || !block_desc || !entries || !n_entries || !quant) return AQ_ERR_ARG;
  Frame f;
  aq_jpeg_info info;
  const int rc = open_file(data, len, f, info, quant);
  if (rc != AQ_OK) return rc;
  if (block_cap < info.coef_count / 64 || entries_cap < info.coef_count || info.coef_count >= ((size_t)1 << 25)) return AQ_ERR_ARG;
  const size_t luma_blocks = (size_t)info.mcus_x * info.hs * info.mcus_y * info.vs;
  const size_t chroma_blocks = (size_t)info.mcus_x * info.mcus_y;
  const size_t plane_first[3] = {0, luma_blocks, luma_blocks + chroma_blocks};
  uint32_t total = 0;
  const int rc2 = decode_scan(data, len, f, info, [&](int c, size_t b, BitReader& br, const HuffTable& dc, const HuffTable& ac, int& pred) {
    PackedSink sink{entries + total};
    const bool ok = decode_block(br, dc, ac, pred, sink);
    block_desc[plane_first[c] + b] = (total << 7) | sink.n;  // first entry (25 bits) | entries in the block (<= 64)
    total += sink.n;
    return ok;
  });
  *n_entries = total;
  return rc2;
}

// ------------------------------------------------------------------ device-side entropy decoding: host preparation
namespace {

void export_table(const HuffTable& t, aqj::DevHuff& o) {
  memset(&o, 0, sizeof o);
  memcpy(o.look, t.look, sizeof o.look);
  memcpy(o.maxcode, t.maxcode, sizeof o.maxcode);
  memcpy(o.valoff, t.valoff, sizeof o.valoff);
  memcpy(o.vals, t.vals, sizeof o.vals);
}

aqj::Geo make_geo(const aq_jpeg_info& info) {
  aqj::Geo g;
  g.hs = info.hs; g.vs = info.vs; g.ncomp = info.ncomp;
  g.bpm = info.ncomp == 1 ? 1 : info.hs * info.vs + 2;
  g.mcus_x = info.mcus_x; g.mcus_y = info.mcus_y;
  g.luma_bw = info.mcus_x * info.hs;
  g.luma_blocks = (long long)g.luma_bw * info.mcus_y * info.vs;
  g.chroma_blocks = info.ncomp == 3 ? (long long)info.mcus_x * info.mcus_y : 0;
  g.total_blocks = g.luma_blocks + 2 * g.chroma_blocks;
  return g;
}

}  // namespace

extern "C" int aq_jpeg_scan_prepare(const uint8_t* data, size_t len, aq_jpeg_info* info, uint8_t* scan_out, size_t scan_cap,
                                    size_t* scan_bytes, uint8_t* tables_out, uint16_t* quant) {
  if (!data || !info || !scan_out || !scan_bytes || !tables_out || !quant) return AQ_ERR_ARG;
  static_assert(AQ_JPEG_TABLE_BYTES == aqj::kTablesPerTile * sizeof(aqj::DevHuff), "AQ_JPEG_TABLE_BYTES");
  Frame f;
  const int rc = open_file(data, len, f, *info, quant);
  if (rc != AQ_OK) return rc;
  if (f.restart_interval) return AQ_ERR_UNSUPPORTED;  // restart markers re-align the stream: use the host decoder for such files
  aqj::DevHuff* tabs = reinterpret_cast<aqj::DevHuff*>(tables_out);
  for (int c = 0; c < 3; ++c) {
    const Component& k = f.comp[c < f.ncomp ? c : 0];
    export_table(f.dc[k.td], tabs[2 * c]);
    export_table(f.ac[k.ta], tabs[2 * c + 1]);
  }
  // entropy-coded segment with the byte stuffing removed (0xFF 0x00 -> 0xFF); any other marker ends it
  size_t n = 0;
  const uint8_t* p = data + f.scan_off;
  const uint8_t* end = data + len;
  while (p < end) {
    const uint8_t* ff = (const uint8_t*)memchr(p, 0xFF, (size_t)(end - p));
    const size_t run = (size_t)((ff ? ff : end) - p);
    if (n + run + 1 + 8 > scan_cap) return AQ_ERR_NOMEM;
    memcpy(scan_out + n, p, run);
    n += run;
    if (!ff || ff + 1 >= end) break;
    if (ff[1] != 0x00) break;  // a marker (EOI normally)
    scan_out[n++] = 0xFF;
    p = ff + 2;
  }
  *scan_bytes = n;
  const size_t padded = (n + 3) / 4 * 4 + 8;  // whole words + two zero words the 32-bit window may touch
  if (padded > scan_cap) return AQ_ERR_NOMEM;
  memset(scan_out + n, 0, padded - n);
  return AQ_OK;
}

// A whole batch in one call, written straight into the staging areas aq_jpeg_decode_device reads; tiles are spread over
// n_threads host threads (no interpreter in the loop).
extern "C" int aq_jpeg_scan_prepare_batch(const uint8_t* const* datas, const size_t* lens, int n, uint8_t* scans, size_t scans_cap,
                                          uint64_t* scan_off, uint32_t* scan_bytes, size_t* total_bytes, uint8_t* tables,
                                          uint16_t* quant, aq_jpeg_info* infos, int* status, int n_threads) {
  if (!datas || !lens || n <= 0 || !scans || !scan_off || !scan_bytes || !total_bytes || !tables || !quant || !infos || !status)
    return AQ_ERR_ARG;
  size_t pos = 0;
  for (int i = 0; i < n; ++i) {  // a scan is never longer than its file
    scan_off[i] = pos;
    pos += (lens[i] + 16 + 3) / 4 * 4;
  }
  if (pos > scans_cap) return AQ_ERR_NOMEM;
  std::atomic<int> next{0};
  auto work = [&]() {
    for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) {
      size_t nb = 0;
      const size_t cap = (lens[i] + 16 + 3) / 4 * 4;
      status[i] = datas[i] ? aq_jpeg_scan_prepare(datas[i], lens[i], &infos[i], scans + scan_off[i], cap, &nb,
                                                  tables + (size_t)i * AQ_JPEG_TABLE_BYTES, quant + (size_t)i * 192)
                           : AQ_ERR_ARG;
      scan_bytes[i] = (uint32_t)nb;
    }
  };
  const int nt = n_threads < 1 ? 1 : (n_threads > n ? n : n_threads);
  std::vector<std::thread> pool;
  for (int t = 1; t < nt; ++t) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  *total_bytes = pos;
  for (int i = 0; i < n; ++i)
    if (status[i] != AQ_OK) return status[i];
  return AQ_OK;
}

// CPU rehearsal of the device schedule (same core, subsequences visited one after another): for tests without a GPU.
extern "C" int aq_jpeg_debug_parallel_decode(const uint8_t* scan, size_t scan_bytes, const uint8_t* tables, const aq_jpeg_info* info,
                                             int sub_bits, int16_t* coef, size_t coef_cap, int* rounds_out) {
  if (!scan || !tables || !info || !coef || sub_bits < 32 || (sub_bits & 31) || ((uintptr_t)scan & 3)) return AQ_ERR_ARG;
  if (coef_cap < info->coef_count) return AQ_ERR_ARG;
  const aqj::Geo g = make_geo(*info);
  const aqj::DevHuff* tabs = reinterpret_cast<const aqj::DevHuff*>(tables);
  const uint32_t* words = reinterpret_cast<const uint32_t*>(scan);
  const uint32_t slot_comp = aqj::pack_slot_comp(g.ncomp, g.hs, g.vs);
  const uint32_t total_bits = (uint32_t)(scan_bytes * 8);
  const int nsub = (int)((total_bits + sub_bits - 1) / sub_bits);
  if (nsub <= 0) return AQ_ERR_ARG;
  std::vector<uint64_t> start(nsub), endst(nsub);
  std::vector<int> nblk(nsub);
  auto limit = [&](int i) { const uint64_t l = (uint64_t)(i + 1) * sub_bits; return (uint32_t)(l < total_bits ? l : total_bits); };
  for (int i = 0; i < nsub; ++i) {  // step 1
    aqj::State st{(uint32_t)((uint64_t)i * sub_bits), 0, 0};
    start[i] = aqj::pack_state(st);
    nblk[i] = aqj::decode_range(words, tabs, slot_comp, g.bpm, st, limit(i));
    endst[i] = aqj::pack_state(st);
  }
  int rounds = 0;
  for (bool changed = true; changed; ++rounds) {  // step 2 (Jacobi order, as the device runs it)
    changed = false;
    std::vector<uint64_t> prev_end(endst);
    for (int i = 1; i < nsub; ++i) {
      if (prev_end[i - 1] == start[i]) continue;
      start[i] = prev_end[i - 1];
      aqj::State st = aqj::unpack_state(start[i]);
      nblk[i] = aqj::decode_range(words, tabs, slot_comp, g.bpm, st, limit(i));
      endst[i] = aqj::pack_state(st);
      changed = true;
    }
    if (rounds > nsub + 1) return AQ_ERR_IO;
  }
  if (rounds_out) *rounds_out = rounds;
  memset(coef, 0, info->coef_count * sizeof(int16_t));
  long long base = 0;
  for (int i = 0; i < nsub; ++i) {  // steps 3 + 4
    aqj::State st = aqj::unpack_state(start[i]);
    long long blk = base;
    const uint32_t lim = limit(i);
    aqj::BitWindow bits;
    bits.init(words, st.p);
    while (st.p < lim && blk < g.total_blocks) {
      int16_t* dst = coef + aqj::block_address(g, blk) * 64;
      if (aqj::decode_one(bits, tabs, slot_comp, g.bpm, st, [&](int z, int v) { dst[kZigzagToNatural[z]] = (int16_t)v; })) ++blk;
    }
    base += nblk[i];
  }
  if (base < g.total_blocks) return AQ_ERR_IO;
  for (int c = 0; c < g.ncomp; ++c) {  // step 5
    int pred = 0;
    const int per_mcu = c == 0 ? (g.ncomp == 1 ? 1 : g.hs * g.vs) : 1, first = c == 0 ? 0 : g.hs * g.vs + c - 1;
    for (long long m = 0; m < (long long)g.mcus_x * g.mcus_y; ++m)
      for (int k = 0; k < per_mcu; ++k) {
        int16_t* dc = coef + aqj::block_address(g, m * g.bpm + first + k) * 64;
        pred += *dc;
        *dc = (int16_t)pred;
      }
  }
  return AQ_OK;
}
