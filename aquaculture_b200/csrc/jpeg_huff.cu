// Entropy decoding of baseline JPEG scans on the GPU: the schedule described in jpeg_huff_core.h.
// Replaces the Huffman stage of the cv2.imread call inside upstream utils/dataloaders.py::LoadImages [UPSTREAM-RECALLED]
// for the tiles gdal.Translate -of JPEG writes [REF src/load_data/tile_tifs.py:74]; the host only walks the markers.
// Latency / integer bound (a chain of dependent table lookups per symbol), not bandwidth bound: parallelism comes from
// the subsequences (1024 bits each: ~1.4k per 1024^2 tile, ~90k per 64-tile batch).
#include "jpeg_huff.cuh"

namespace aq {
namespace {

using namespace aqj;

#ifndef AQ_HUFF_THREADS
#define AQ_HUFF_THREADS 1024
#endif
constexpr int HT = AQ_HUFF_THREADS;  // threads per CTA = subsequences per CTA

__constant__ uint8_t c_zigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                     41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                     30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct Shared {
  DevHuff tabs[kTablesPerTile];
};

__device__ __forceinline__ void load_tables(Shared& sh, const HuffParams& p, int tile) {
  const uint32_t* src = reinterpret_cast<const uint32_t*>(p.tables + (size_t)tile * kTablesPerTile * sizeof(DevHuff));
  uint32_t* dst = reinterpret_cast<uint32_t*>(sh.tabs);
  for (int i = threadIdx.x; i < (int)(kTablesPerTile * sizeof(DevHuff) / 4); i += blockDim.x) dst[i] = src[i];
  __syncthreads();
}

__device__ __forceinline__ uint32_t sub_limit(const HuffParams& p, uint32_t total_bits, int i) {
  const unsigned long long l = (unsigned long long)(i + 1) * p.sub_bits;
  return (uint32_t)(l < total_bits ? l : total_bits);
}

// Steps 1 + 2.  CTA (x, tile) owns the REAL = HT - GHOST subsequences [x * REAL, (x + 1) * REAL) of the tile; its first
// GHOST threads re-decode the GHOST subsequences in front of that range as a warm-up (first launch only, results not
// stored), so that the CTA's first own subsequence usually starts from a state that has already fallen into step.
// first != 0: decode from the guessed states, then relax inside the CTA until nothing changes; first == 0: relax again
// from the boundary state the previous CTA left in global memory -- a CTA whose boundary did not move returns at
// once.  Launching it (CTAs per tile) times reaches the fixed point whatever the data: launch k finds CTA k-1 final.
constexpr int GHOST = 8, REAL = HT - GHOST;

__global__ void __launch_bounds__(HT) jpeg_huff_sync_kernel(HuffParams p, int first) {
  __shared__ Shared sh;
  __shared__ unsigned long long s_end[HT];
  const int tile = blockIdx.y, t = threadIdx.x;
  const int i = (int)blockIdx.x * REAL - GHOST + t;
  const uint32_t total_bits = p.scan_bytes[tile] * 8u;
  const int nsub = (int)((total_bits + p.sub_bits - 1) / p.sub_bits);
  if ((int)blockIdx.x * REAL >= nsub) return;
  const bool real = t >= GHOST;
  const bool live = i >= 0 && i < nsub && (real || first);
  const size_t slot = (size_t)tile * p.nsub_max + (i >= 0 ? i : 0);
  unsigned long long boundary = 0;
  if (!first) {
    const bool at_boundary = t == GHOST && live && i > 0;
    if (at_boundary) boundary = p.end[slot - 1];
    if (!__syncthreads_or(at_boundary && boundary != p.start[slot])) return;  // the previous CTA's last state did not move
  }
  load_tables(sh, p, tile);
  const uint32_t* words = reinterpret_cast<const uint32_t*>(p.scans + p.scan_off[tile]);
  const uint32_t slot_comp = pack_slot_comp(p.geo.ncomp, p.geo.hs, p.geo.vs);
  unsigned long long my_start = 0, my_end = 0;
  int my_nblk = 0;
  if (live) {
    if (first) {
      State st{(uint32_t)((unsigned long long)i * p.sub_bits), 0, 0};
      my_start = pack_state(st);
      my_nblk = decode_range(words, sh.tabs, slot_comp, p.geo.bpm, st, sub_limit(p, total_bits, i));
      my_end = pack_state(st);
    } else {
      my_start = p.start[slot];
      my_end = p.end[slot];
      my_nblk = p.nblk[slot];
    }
  }
  // where a thread's predecessor state comes from: the neighbouring thread (a warm-up thread included, first launch),
  // or -- first own subsequence in later launches -- the previous CTA's final state
  const bool from_boundary = !first && t == GHOST && live && i > 0;
  const bool from_neighbour = live && i > 0 && t > 0 && !from_boundary;
  s_end[t] = my_end;
  __syncthreads();
  for (;;) {
    bool changed = false;
    if (from_boundary || from_neighbour) {
      const unsigned long long prev = from_boundary ? boundary : s_end[t - 1];
      if (prev != my_start) {
        my_start = prev;
        State st = unpack_state(prev);
        my_nblk = decode_range(words, sh.tabs, slot_comp, p.geo.bpm, st, sub_limit(p, total_bits, i));
        my_end = pack_state(st);
        changed = true;
      }
    }
    __syncthreads();  // every s_end[t - 1] has been read
    if (changed) s_end[t] = my_end;
    if (!__syncthreads_or(changed)) break;
  }
  if (live && real) {
    p.start[slot] = my_start;
    p.end[slot] = my_end;
    p.nblk[slot] = my_nblk;
  }
}

// Step 3 + verification: one CTA per tile.  base[i] = blocks completed before subsequence i; status[tile] = 0 when the
// chain of states is consistent and covers the expected number of blocks.
__global__ void __launch_bounds__(HT) jpeg_huff_scan_kernel(HuffParams p) {
  __shared__ long long s_part[HT];
  __shared__ int s_bad;
  const int tile = blockIdx.x;
  const uint32_t total_bits = p.scan_bytes[tile] * 8u;
  const int nsub = (int)((total_bits + p.sub_bits - 1) / p.sub_bits);
  const size_t off = (size_t)tile * p.nsub_max;
  if (threadIdx.x == 0) s_bad = 0;
  const int per = (nsub + HT - 1) / HT, lo = min(threadIdx.x * per, nsub), hi = min(lo + per, nsub);
  long long sum = 0;
  bool bad = false;
  for (int i = lo; i < hi; ++i) {
    sum += p.nblk[off + i];
    if (i > 0 && p.start[off + i] != p.end[off + i - 1]) bad = true;
  }
  s_part[threadIdx.x] = sum;
  __syncthreads();
  if (bad) s_bad = 1;
  for (int d = 1; d < HT; d <<= 1) {  // inclusive scan of the per-thread sums
    const long long v = threadIdx.x >= d ? s_part[threadIdx.x - d] : 0;
    __syncthreads();
    s_part[threadIdx.x] += v;
    __syncthreads();
  }
  long long run = s_part[threadIdx.x] - sum;
  for (int i = lo; i < hi; ++i) {
    p.base[off + i] = run;
    run += p.nblk[off + i];
  }
  if (threadIdx.x == HT - 1) p.status[tile] = (s_bad || s_part[HT - 1] < p.geo.total_blocks || nsub <= 0) ? -6 : 0;
}

// Step 4: every subsequence is decoded from its final start state, writing the coefficients it codes.
__global__ void __launch_bounds__(HT) jpeg_huff_write_kernel(HuffParams p) {
  __shared__ Shared sh;
  const int tile = blockIdx.y, i = blockIdx.x * HT + threadIdx.x;
  const uint32_t total_bits = p.scan_bytes[tile] * 8u;
  const int nsub = (int)((total_bits + p.sub_bits - 1) / p.sub_bits);
  if (blockIdx.x * HT >= nsub) return;
  load_tables(sh, p, tile);
  if (i >= nsub) return;
  const uint32_t* words = reinterpret_cast<const uint32_t*>(p.scans + p.scan_off[tile]);
  const uint32_t slot_comp = pack_slot_comp(p.geo.ncomp, p.geo.hs, p.geo.vs);
  const size_t slot = (size_t)tile * p.nsub_max + i;
  State st = unpack_state(p.start[slot]);
  long long blk = p.base[slot];
  const uint32_t lim = sub_limit(p, total_bits, i);
  int16_t* tile_coef = p.coef + (size_t)tile * p.geo.total_blocks * 64;
  int* tile_dc = p.dc + (size_t)tile * p.geo.total_blocks;
  BitWindow bits;
  bits.init(words, st.p);
  while (st.p < lim && blk < p.geo.total_blocks) {
    int16_t* dst = tile_coef + block_address(p.geo, blk) * 64;
    int* dcd = tile_dc + blk;  // DC differences stay in decode order: the prediction scan below runs over a compact array
    if (decode_one(bits, sh.tabs, slot_comp, p.geo.bpm, st, [&](int z, int v) {
          if (z == 0) *dcd = v;
          else dst[c_zigzag[z]] = (int16_t)v;
        }))
      ++blk;
  }
}

// Step 5: DC prediction undone by a scan over each component's blocks in decode order (compact int32 array; the IDCT
// kernel picks the values up from there).  CTA (component, tile).
__global__ void __launch_bounds__(HT) jpeg_huff_dc_kernel(HuffParams p) {
  __shared__ int s_part[HT];
  const int c = blockIdx.x, tile = blockIdx.y;
  if (c >= p.geo.ncomp) return;
  const Geo& g = p.geo;
  const int per_mcu = c == 0 ? (g.ncomp == 1 ? 1 : g.hs * g.vs) : 1, first = c == 0 ? 0 : g.hs * g.vs + c - 1;
  const long long n = (long long)g.mcus_x * g.mcus_y * per_mcu;
  const long long per = (n + HT - 1) / HT, lo = min((long long)threadIdx.x * per, n), hi = min(lo + per, n);
  int* tile_dc = p.dc + (size_t)tile * g.total_blocks;
  auto addr = [&](long long j) { return tile_dc + (j / per_mcu) * g.bpm + first + (int)(j % per_mcu); };
  int sum = 0;
  for (long long j = lo; j < hi; ++j) sum += *addr(j);
  s_part[threadIdx.x] = sum;
  __syncthreads();
  for (int d = 1; d < HT; d <<= 1) {
    const int v = threadIdx.x >= d ? s_part[threadIdx.x - d] : 0;
    __syncthreads();
    s_part[threadIdx.x] += v;
    __syncthreads();
  }
  int run = s_part[threadIdx.x] - sum;
  for (long long j = lo; j < hi; ++j) {
    int* a = addr(j);
    run += *a;
    *a = (int)(short)run;  // coefficients are 16-bit in the format
  }
}

}  // namespace

size_t jpeg_huff_state_bytes(int n, int nsub_max) { return (size_t)n * nsub_max * (8 + 8 + 4 + 8) + (size_t)n * 4 + 256; }

cudaError_t launch_jpeg_huffman(HuffParams p, void* state_scratch, int n, int nsub_max, cudaStream_t stream) {
  if (n <= 0 || n > 65535 || nsub_max <= 0) return cudaErrorInvalidValue;
  uint8_t* s = (uint8_t*)state_scratch;
  const size_t cells = (size_t)n * nsub_max;
  p.start = (unsigned long long*)s;
  p.end = p.start + cells;
  p.base = (long long*)(p.end + cells);
  p.nblk = (int*)(p.base + cells);
  p.status = p.nblk + cells;
  p.nsub_max = nsub_max;
  const int ctas = (nsub_max + HT - 1) / HT, sync_ctas = (nsub_max + REAL - 1) / REAL;
  cudaError_t e = cudaMemsetAsync(p.coef, 0, (size_t)n * p.geo.total_blocks * 64 * sizeof(int16_t), stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(p.dc, 0, (size_t)n * p.geo.total_blocks * sizeof(int), stream);
  if (e != cudaSuccess) return e;
  for (int k = 0; k < sync_ctas; ++k) jpeg_huff_sync_kernel<<<dim3(sync_ctas, n), HT, 0, stream>>>(p, k == 0);
  jpeg_huff_scan_kernel<<<n, HT, 0, stream>>>(p);
  jpeg_huff_write_kernel<<<dim3(ctas, n), HT, 0, stream>>>(p);
  jpeg_huff_dc_kernel<<<dim3(3, n), HT, 0, stream>>>(p);
  return cudaGetLastError();
}

const int* jpeg_huff_status_ptr(void* state_scratch, int n, int nsub_max) {
  const size_t cells = (size_t)n * nsub_max;
  return (const int*)((uint8_t*)state_scratch + cells * (8 + 8 + 8)) + cells;
}

}  // namespace aq
