// ia_jpeg_kernels.cuh — device side of the JPEG decoder (see ia_jpeg.cu for the overview).
//
//   U1 unstuff_count_kernel   per 8 KB tile of the entropy-coded bytes: how many bytes
//                             survive (no FF00 stuffing, no RSTn markers), how many restart
//                             markers, where the scan ends (first other marker)
//   U2 unstuff_scan_kernel    per image: exclusive scan of the tile counts, totals
//   U3 unstuff_scatter_kernel per tile: write the clean stream, the region starts and the code
//                             of every marker inside the data
//   R  resync_walk_kernel     damaged restart markers only: which region of the stream is which
//                             restart segment, the way libjpeg's resynchronisation decides it
//   H  huffman_kernel         one CTA per chunk of the clean stream (256 or 512 subsequences of
//                             128 or 256 bytes, one per thread: 32 to 128 KB), self-synchronising
//                             decode inside the CTA, chunks chained through a ticketed look-back,
//                             libjpeg's concealment of damaged segments in the write pass
//   I  idct_kernel            one thread per 8x8 block, jidctint.c "islow" as the SIMD code evaluates it
//   C  colour_kernel          eight pixels per thread: fancy upsampling, YCbCr->RGB, orientation
#pragma once

#include "ia_common.cuh"

namespace ia {
namespace jpeg {

// Bytes of the clean stream per thread of huffman_kernel.  Two instantiations: 128 (more CTAs per
// image, lowest latency for a few frames) and 256 (fewer rounds of the self-synchronising decode
// and less repeated work, best throughput once the batch fills the GPU; 64, 384 and 512 are
// slower on both counts).  The workspace is sized for the smaller one.
constexpr int kSubseqBytes = 128;
constexpr int kSubseqBytesLarge = 256;
// Cross-overs measured on B200 (profiles/README.md).  256-byte subsequences win once their CTAs
// alone fill most of the GPU (256 x VGA: 1.93 -> 1.59 ms; 128 x VGA, 256 CTAs: 1.08 -> 1.26 ms),
// and for large files — many chunks chained per image — already from a few megabytes (8 x 1080p
// 1.03 -> 0.95 ms, one frame 0.54 -> 0.66 ms).
constexpr unsigned long long kLargeSubseqCtasFrom = 384;
constexpr unsigned long long kLargeSubseqBytesFrom = 6ull << 20;
constexpr unsigned long long kLargeFileBytes = 256ull << 10;
constexpr int kSubseqBits = kSubseqBytes * 8;
// huffman_kernel runs as 256 threads x 4 CTAs per SM (64 registers, nothing spilled).  512
// threads x 3 CTAs (42 registers, the first shape of this kernel) is still instantiated for
// comparison (IA_JPEG_THREADS=512): 4 % ahead at exactly 64 x 1080p, behind everywhere else
// measured.  The constants here are the smallest configuration: they size the workspace.
constexpr int kHuffThreads = 256;
constexpr int kHuffThreadsLarge = 512;
constexpr uint32_t kChunkBits = (uint32_t)kHuffThreads * kSubseqBits;
constexpr uint32_t kChunkBytes = kChunkBits / 8;
constexpr int kLutBits = 10;
constexpr int kUnstuffThreads = 512;
constexpr int kUnstuffPerThread = 16;
constexpr uint32_t kTileBytes = kUnstuffThreads * kUnstuffPerThread;
constexpr int kCleanPad = 64;   // readable zero bytes after the clean stream
constexpr int kBroadcastEvery = 16;  // huffman_kernel: rounds before the first phase broadcast (then 32, 64 ...)
constexpr int kChainWords = 16;  // per chunk: tentative and final hand-over records (huffman_kernel)
constexpr int kTileWords = 4;   // per tile: kept bytes, markers, end position, pad

__host__ __device__ inline uint32_t scan_bytes(const ia_jpeg_job& j) { return j.file_bytes - j.scan_offset; }
__host__ __device__ inline uint32_t tile_capacity(const ia_jpeg_job& j) {
  return (scan_bytes(j) + kTileBytes - 1) / kTileBytes + 1;
}
// ia_jpeg_job::segment_dev: [0] clean bytes, [1] markers recorded, [2] flags (1 = no marker behind
// the data, 2 = the region table below holds segments), [3, 3 + cap) clean byte offset at which
// the region behind marker r starts, [3 + cap, 3 + 2 cap) that marker's code, later the restart
// segment of the region.  A damaged stream can hold more markers than segments; beyond cap they
// are not told apart any more (the image is flagged long before).
__host__ __device__ inline uint32_t region_capacity(const ia_jpeg_job& j) { return 2u * (uint32_t)j.num_segments + 8u; }
__host__ __device__ inline uint32_t chunk_capacity(const ia_jpeg_job& j) {
  return (scan_bytes(j) + kChunkBytes - 1) / kChunkBytes + 1;
}

// ================================================================================ unstuff
// Classify the 16 bytes [r0, r0+16) of the scan: which survive, which end a restart marker,
// and the first position that starts any other marker (the end of the scan).
//
// Threads are laid out over the 16-byte aligned stream that contains the scan (v0 = this
// thread's offset from the aligned address at or below the scan start), so the 16 bytes are one
// 128-bit load; the byte before and the byte after come from the neighbouring lanes.  r0 (out)
// is the scan-relative position of the thread's first byte and may be negative for the very
// first thread; bytes outside [0, n) are never kept (before the scan they read as 0, after it
// as an EOI marker).
//
// rst_restarts: the frame has a restart interval.  Without one libjpeg never consumes an RSTn
// marker: like any other marker it ends the data.
__device__ __forceinline__ void classify(const uint8_t* __restrict__ src, uint32_t n, uint32_t v0, bool rst_restarts,
                                         uint8_t (&b)[18], uint32_t& keep_mask, uint32_t& bnd_mask, uint32_t& my_end,
                                         long long& r_first) {
  const uint32_t a = (uint32_t)((uintptr_t)src & 15);
  const uint8_t* __restrict__ base = src - a;
  const uint32_t nv = n + a;  // length of the aligned stream that still belongs to the scan
  auto fix = [&](uint32_t v, uint32_t raw) -> uint8_t {  // byte at aligned offset v
    return v < a ? (uint8_t)0 : (v < nv ? (uint8_t)raw : (uint8_t)0xD9);
  };
  uint4 q = make_uint4(0, 0, 0, 0);
  if (v0 < nv) q = __ldg(reinterpret_cast<const uint4*>(base + v0));
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int k = 0; k < kUnstuffPerThread; ++k) b[k + 1] = fix(v0 + k, (w[k >> 2] >> (8 * (k & 3))) & 255u);
  const int lane = threadIdx.x & 31;
  uint32_t prev = __shfl_up_sync(0xffffffffu, (uint32_t)b[kUnstuffPerThread], 1);
  uint32_t next = __shfl_down_sync(0xffffffffu, (uint32_t)b[1], 1);
  if (lane == 0) prev = v0 == 0 ? 0u : fix(v0 - 1, (v0 - 1 < nv) ? __ldg(base + v0 - 1) : 0u);
  if (lane == 31) next = fix(v0 + 16, (v0 + 16 < nv) ? __ldg(base + v0 + 16) : 0u);
  b[0] = (uint8_t)prev, b[kUnstuffPerThread + 1] = (uint8_t)next;
  r_first = (long long)v0 - (long long)a;
  const uint32_t r0 = v0 - a;  // wraps for bytes before the scan: those fail r < n below
  keep_mask = 0, bnd_mask = 0, my_end = 0xffffffffu;
#pragma unroll
  for (int k = 0; k < kUnstuffPerThread; ++k) {
    const uint32_t r = r0 + k;
    const uint8_t prev = b[k], cur = b[k + 1], next = b[k + 2];
    if (r < n) {
      // With a restart interval two kinds of marker do not end the data for good: RSTn, and codes
      // below C0, which libjpeg's resynchronisation skips as garbage at the next restart
      // (jpeg_resync_to_restart).  Both cut the stream into regions; resync_walk_kernel decides
      // which region is which restart segment.
      if (cur == 0xFF) {
        if (next == 0x00) keep_mask |= 1u << k;
        else if (next != 0xFF && !(rst_restarts && ((next >= 0xD0 && next <= 0xD7) || next < 0xC0)))
          my_end = min(my_end, r);
      } else if (prev == 0xFF) {
        if (rst_restarts && ((cur >= 0xD0 && cur <= 0xD7) || (cur != 0 && cur < 0xC0)))
          bnd_mask |= 1u << k;  // cur == 0 is stuffing: dropped
      } else {
        keep_mask |= 1u << k;
      }
    }
  }
}

__device__ __forceinline__ void cut_at(uint32_t end, long long r_first, uint32_t& keep_mask, uint32_t& bnd_mask) {
  if ((long long)end < r_first + kUnstuffPerThread) {
    const long long valid = (long long)end > r_first ? (long long)end - r_first : 0;  // < 16
    const uint32_t m = (1u << (int)valid) - 1u;
    keep_mask &= m, bnd_mask &= m;
  }
}

// CTA-wide inclusive scan of a packed (markers << 16 | kept) count; returns the exclusive
// value of this thread and, through `total`, the CTA total.
__device__ __forceinline__ uint32_t cta_scan_packed(uint32_t mine, uint32_t* s_warp, uint32_t& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < kUnstuffThreads / 32 ? s_warp[lane] : 0;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += o;
    }
    if (lane < kUnstuffThreads / 32) s_warp[lane] = w;
  }
  __syncthreads();
  total = s_warp[kUnstuffThreads / 32 - 1];
  return incl - mine + (warp ? s_warp[warp - 1] : 0);
}

__global__ void __launch_bounds__(kUnstuffThreads) unstuff_count_kernel(const ia_jpeg_job* __restrict__ jobs) {
  const ia_jpeg_job& J = jobs[blockIdx.y];
  const uint32_t n = scan_bytes(J);
  const uint32_t tile = blockIdx.x;
  if (tile >= tile_capacity(J)) return;
  __shared__ uint32_t s_end, s_warp[kUnstuffThreads / 32];
  if (threadIdx.x == 0) s_end = 0xffffffffu;
  __syncthreads();
  const uint32_t r0 = tile * kTileBytes + threadIdx.x * kUnstuffPerThread;
  uint8_t b[18];
  uint32_t keep_mask, bnd_mask, my_end;
  long long r_first;
  classify(J.file_dev + J.scan_offset, n, r0, J.restart_interval != 0, b, keep_mask, bnd_mask, my_end, r_first);
  if (my_end != 0xffffffffu) atomicMin(&s_end, my_end);
  __syncthreads();
  const uint32_t end = s_end;
  cut_at(end, r_first, keep_mask, bnd_mask);
  uint32_t total;
  cta_scan_packed((uint32_t)__popc(keep_mask) | ((uint32_t)__popc(bnd_mask) << 16), s_warp, total);
  if (threadIdx.x == 0) {
    uint32_t* t = J.tile_dev + (size_t)tile * kTileWords;
    t[0] = total & 0xffffu, t[1] = total >> 16, t[2] = end;
  }
}

// One warp per image walks the tiles: counts -> exclusive bases; tiles after the end of the
// scan are marked dead (end = 0).  Also writes the totals and the zero padding.
__global__ void __launch_bounds__(32) unstuff_scan_kernel(const ia_jpeg_job* __restrict__ jobs, int32_t* __restrict__ status) {
  const ia_jpeg_job& J = jobs[blockIdx.x];
  const uint32_t tiles = tile_capacity(J);
  const int lane = threadIdx.x;
  uint32_t base_keep = 0, base_bnd = 0, scan_end = 0xffffffffu;
  for (uint32_t t0 = 0; t0 < tiles; t0 += 32) {
    const uint32_t t = t0 + lane;
    uint32_t* rec = J.tile_dev + (size_t)t * kTileWords;
    uint32_t keep = 0, bnd = 0, end = 0xffffffffu;
    if (t < tiles) keep = rec[0], bnd = rec[1], end = rec[2];
    // the first tile (in order) that saw an end marker closes the scan
    const uint32_t ended = __ballot_sync(0xffffffffu, end != 0xffffffffu);
    const int first = ended ? __ffs(ended) - 1 : 32;
    const bool dead = scan_end != 0xffffffffu || lane > first;
    if (dead) keep = 0, bnd = 0;
    uint32_t ik = keep, ib = bnd;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t ok = __shfl_up_sync(0xffffffffu, ik, d), ob = __shfl_up_sync(0xffffffffu, ib, d);
      if (lane >= d) ik += ok, ib += ob;
    }
    if (t < tiles) {
      rec[0] = base_keep + ik - keep;
      rec[1] = base_bnd + ib - bnd;
      rec[2] = dead ? 0u : end;  // dead tiles keep nothing; live ones cut at their own end
    }
    base_keep += __shfl_sync(0xffffffffu, ik, 31);
    base_bnd += __shfl_sync(0xffffffffu, ib, 31);
    if (scan_end == 0xffffffffu && ended) scan_end = __shfl_sync(0xffffffffu, end, first);
  }
  for (int i = lane; i < kCleanPad; i += 32) J.clean_dev[base_keep + i] = 0;
  if (lane == 0) {
    J.segment_dev[0] = base_keep;
    J.segment_dev[1] = min(base_bnd, region_capacity(J));
    // no marker behind the entropy-coded bytes (or a lone FF as the last byte): libjpeg would
    // run out of input if the blocks are not all there
    J.segment_dev[2] = scan_end == 0xffffffffu || scan_end + 1 >= scan_bytes(J);
    // the Huffman kernel ORs its findings in; a restart marker count other than the interval
    // implies is damage already (without an interval libjpeg stops reading at the first one)
    status[blockIdx.x] = base_bnd + 1 == (uint32_t)J.num_segments ? 0 : 1;
  }
}

__global__ void __launch_bounds__(kUnstuffThreads) unstuff_scatter_kernel(const ia_jpeg_job* __restrict__ jobs,
                                                                          int32_t* __restrict__ status) {
  const ia_jpeg_job& J = jobs[blockIdx.y];
  const uint32_t n = scan_bytes(J);
  const uint32_t tile = blockIdx.x;
  if (tile >= tile_capacity(J)) return;
  const uint32_t* rec = J.tile_dev + (size_t)tile * kTileWords;
  const uint32_t end = rec[2];
  if (end == 0) return;  // after the end of the scan
  __shared__ uint32_t s_warp[kUnstuffThreads / 32];
  const uint32_t r0 = tile * kTileBytes + threadIdx.x * kUnstuffPerThread;
  uint8_t b[18];
  uint32_t keep_mask, bnd_mask, my_end;
  long long r_first;
  classify(J.file_dev + J.scan_offset, n, r0, J.restart_interval != 0, b, keep_mask, bnd_mask, my_end, r_first);
  cut_at(end, r_first, keep_mask, bnd_mask);
  uint32_t total;
  const uint32_t excl = cta_scan_packed((uint32_t)__popc(keep_mask) | ((uint32_t)__popc(bnd_mask) << 16), s_warp, total);
  // compact the tile in shared memory first, then write it out with consecutive lanes on
  // consecutive bytes (a thread's own run of ~16 bytes would cost a sector per lane and store)
  __shared__ uint8_t s_tile[kTileBytes];
  const uint32_t tile_base = rec[0];
  uint32_t local = excl & 0xffffu;
  uint32_t bpos = rec[1] + (excl >> 16);
  uint32_t* __restrict__ seg = J.segment_dev;
  const uint32_t seg_cap = region_capacity(J);
#pragma unroll
  for (int k = 0; k < kUnstuffPerThread; ++k) {
    if (keep_mask >> k & 1u) s_tile[local++] = b[k + 1];
    if (bnd_mask >> k & 1u) {
      // RSTn markers count modulo 8 and libjpeg resynchronises by these numbers.  Keep the code of
      // every marker (8 = a garbage code below C0); one out of sequence flags the image and makes
      // resync_walk_kernel work out which region belongs to which segment.
      const uint32_t code = (b[k + 1] >= 0xD0 && b[k + 1] <= 0xD7) ? (b[k + 1] & 7u) : 8u;
      if (bpos < seg_cap) {
        seg[3 + bpos] = tile_base + local;  // the next region starts at this clean byte
        seg[3 + seg_cap + bpos] = code;
      }
      if (code != (bpos & 7u)) atomicOr(status + blockIdx.y, 1);
      ++bpos;
    }
  }
  __syncthreads();
  uint8_t* __restrict__ clean = J.clean_dev + tile_base;
  const uint32_t kept = total & 0xffffu;
  for (uint32_t i = threadIdx.x; i < kept; i += kUnstuffThreads) clean[i] = s_tile[i];
}

// Damaged restart markers only (flagged images; intact ones leave at once).  libjpeg restarts by
// MCU count and takes whatever marker comes next in the stream (jdmarker.c read_restart_marker /
// jpeg_resync_to_restart): the expected number is consumed; one or two ahead is left pending
// (that interval gets no data and the marker is looked at again at the next restart); one or
// two behind and garbage codes are skipped together with the data that follows them; anything
// else is taken as the expected one.  One thread walks the marker list and turns each code
// (second table of segment_dev) into the restart segment that the data after that marker belongs
// to, or -1 for data libjpeg never decodes.  The last non-RST marker (>= C0) stays pending for good.
__global__ void __launch_bounds__(32) resync_walk_kernel(const ia_jpeg_job* __restrict__ jobs,
                                                          const int32_t* __restrict__ status) {
  const ia_jpeg_job& J = jobs[blockIdx.x];
  if (threadIdx.x != 0) return;
  uint32_t* flags = J.segment_dev + 2;
  if (status[blockIdx.x] == 0 || J.restart_interval == 0) return;
  const int m = (int)J.segment_dev[1], segments = J.num_segments;
  int32_t* code = reinterpret_cast<int32_t*>(J.segment_dev + 3 + region_capacity(J));
  int c = 0;  // next unread marker
  for (int k = 1; k < segments; ++k) {
    const int desired = (k - 1) & 7;
    while (c < m) {
      const int v = code[c];
      const int ahead = (v - desired) & 7;
      if (v == desired || (v < 8 && ahead != 1 && ahead != 2 && ahead != 7 && ahead != 6)) {
        code[c++] = k;  // consumed: the data behind it is segment k
        break;
      }
      if (v < 8 && (ahead == 1 || ahead == 2)) break;  // left pending: segment k gets no data
      code[c++] = -1;  // skipped, with the data behind it
    }
    if (c >= m) break;  // the marker that ends the scan stays pending: no more data for anyone
  }
  for (; c < m; ++c) code[c] = -1;
  *flags |= 2u;  // the Huffman kernel maps regions through the table
}

// ================================================================================ Huffman
constexpr int kHuffSlots = 6;  // look-ups a scan can need: one DC and one AC per component

// Packed look-up entry, one byte per field so that each is one PRMT/LOP away:
//   byte 0  bits the symbol occupies (code length + magnitude bits, <= 31)
//   byte 1  code length
//   byte 2  32 - magnitude bits (the right shift that extracts them)
//   byte 3  how far the zig-zag index moves: 1 for a DC symbol, run + 1 for a coefficient,
//           16 for ZRL, 64 for end-of-block (anything that reaches 64 ends the block)
// 0 = the code is longer than the look-up covers (huff_slow builds the entry on the fly).
struct HuffTables {
  uint32_t lut[kHuffSlots][1 << kLutBits];
  int32_t maxcode[kHuffSlots][18];  // exclusive upper bound of the codes of each length, left-aligned to 16 bits
  int32_t valoff[kHuffSlots][18];   // index of the first value of that length minus its first code
  uint8_t vals[kHuffSlots][256];
  uint32_t blk_info[16];  // per block of an MCU: component | DC slot << 8 | AC slot << 16
  uint8_t natural[64];    // zig-zag index -> natural (row-major) index
};

__device__ __forceinline__ uint32_t pack_entry(int slot, int len, int sym) {
  const int s = sym & 15, r = sym >> 4;
  const int zinc = slot < 3 ? 1 : (s ? r + 1 : (r == 15 ? 16 : 64));
  return (uint32_t)(len + s) | ((uint32_t)len << 8) | ((uint32_t)(32 - s) << 16) | ((uint32_t)zinc << 24);
}

struct Agg {  // what a run of the stream contributes: completed blocks, DC sums, "predictors were reset"
  int nb, flag, d0, d1, d2;
};

__device__ __forceinline__ Agg combine(const Agg& a, const Agg& b) {
  Agg r;
  r.nb = b.flag ? b.nb : a.nb + b.nb;  // blocks since the last restart boundary
  if (b.flag) r.flag = 1, r.d0 = b.d0, r.d1 = b.d1, r.d2 = b.d2;
  else r.flag = a.flag, r.d0 = a.d0 + b.d0, r.d1 = a.d1 + b.d1, r.d2 = a.d2 + b.d2;
  return r;
}

__device__ __forceinline__ Agg shfl_up_agg(const Agg& a, int d) {
  Agg r;
  r.nb = __shfl_up_sync(0xffffffffu, a.nb, d);
  r.flag = __shfl_up_sync(0xffffffffu, a.flag, d);
  r.d0 = __shfl_up_sync(0xffffffffu, a.d0, d);
  r.d1 = __shfl_up_sync(0xffffffffu, a.d1, d);
  r.d2 = __shfl_up_sync(0xffffffffu, a.d2, d);
  return r;
}

// entry for a code longer than the look-up covers
__device__ __forceinline__ uint32_t huff_slow(const HuffTables& T, int slot, uint32_t bits) {
  const int32_t code = (int32_t)(bits >> 16);
  for (int l = kLutBits + 1; l <= 16; ++l) {
    if (code < T.maxcode[slot][l]) return pack_entry(slot, l, T.vals[slot][(T.valoff[slot][l] + (code >> (16 - l))) & 255]);
  }
  // an invalid code yields symbol 0 after 17 bits: libjpeg's search loop (jpeg_huff_decode) is only
  // stopped by the sentinel maxcode[17]
  return pack_entry(slot, 17, 0);
}

struct Stream {
  const uint32_t* words;   // clean stream
  uint32_t total_bits;
  const uint32_t* bounds;  // restart boundaries (clean byte offsets), ascending
  int num_bounds;
  int bpm;
  int seg_blocks;    // blocks per restart segment (0 without restart intervals)
  int segments;      // restart segments of the frame
  uint32_t first_end_bit;  // where the first subsequence of the stream ends
  int32_t* damaged;  // status word of the image: the write pass ORs 1 into it
  const int32_t* seg_of;  // damaged markers only (resync_walk_kernel): segment of the region behind marker r, or -1

  // restart segment that region r (the data between marker r - 1 and marker r) belongs to
  __device__ __forceinline__ int segment_of(int r) const { return r == 0 ? 0 : (seg_of ? seg_of[r - 1] : r); }
};

__device__ __forceinline__ uint32_t load_be(const uint32_t* __restrict__ words, uint32_t i) {
  return __byte_perm(__ldg(words + i), 0, 0x0123);
}

// What libjpeg does when a restart segment (or the whole scan) stops short of its blocks at a
// marker (jdhuff.c decode_mcu, insufficient_data): it supplies zero bits, finishes the MCU in
// hand from them — a whole one if the data ended exactly between two — and leaves the rest
// zero.  q: bit position to go on from; end_pos: where the real bits stop; limit_blk: first
// block that is not this segment's.  Cold path of the write pass.
struct TailState {
  int b, z, blk, dc0, dc1, dc2;
};

__device__ __noinline__ void conceal(const Stream& S, const HuffTables& T, TailState& s, uint32_t q, uint32_t end_pos,
                                     int limit_blk, int j, int16_t* __restrict__ coef) {
  bool crossed = false;
  uint32_t info = T.blk_info[s.b];
  while (s.blk < limit_blk && !(crossed && s.b == 0 && s.z == 0)) {
    // libjpeg restarts by MCU count: the predictors are reset even when the marker is missing
    // (and what is left of the last byte, padding, is dropped).  j segments were entered at
    // their markers; a restart that is due beyond those is one without a marker.
    if (s.z == 0 && s.b == 0 && S.seg_blocks && s.blk % S.seg_blocks == 0 && s.blk / S.seg_blocks > j)
      s.dc0 = s.dc1 = s.dc2 = 0, q = max(q, end_pos);
    uint32_t bits = 0;
    if (q < end_pos) {
      bits = __funnelshift_l(__byte_perm(__ldg(S.words + (q >> 5) + 1), 0, 0x0123),
                             __byte_perm(__ldg(S.words + (q >> 5)), 0, 0x0123), q);
      if (end_pos - q < 32u) bits &= ~(0xffffffffu >> (end_pos - q));
    }
    const bool dc_mode = s.z == 0;
    const uint32_t slot = dc_mode ? (info >> 8) & 255u : info >> 16;
    uint32_t e = T.lut[slot][bits >> (32 - kLutBits)];
    if (e == 0) e = huff_slow(T, (int)slot, bits);
    const uint32_t len = (e >> 8) & 255u, shift = (e >> 16) & 255u;
    q += e & 255u;
    crossed |= q > end_pos;
    const uint32_t tail = bits << len;
    const int raw = (int)__funnelshift_rc(tail, 0u, shift);
    const int value = raw + (~((int)tail >> 31) & (1 - (int)(1u << (32u - shift))));
    const uint32_t comp = info & 255u;
    int& dc = comp == 0 ? s.dc0 : (comp == 1 ? s.dc1 : s.dc2);
    if (dc_mode) dc += value;
    const int zn = s.z + (int)(e >> 24);
    if (dc_mode || shift != 32u) coef[(size_t)s.blk * 64 + T.natural[min(zn - 1, 63)]] = (int16_t)(dc_mode ? dc : value);
    s.z = zn;
    if (zn >= 64) {
      s.z = 0, ++s.blk;
      s.b = s.b + 1 == S.bpm ? 0 : s.b + 1;
      info = T.blk_info[s.b];
    }
  }
}

// Write pass, cold: the decoder leaves region r (the marker that ends it sits at bit end_pos) in
// state (b, z) with `blk` the next block it would write.  Unless the region held exactly its
// segment's blocks and the next region is the next segment, the image is damaged and libjpeg's
// concealment applies: a segment that ran dry is finished from zero bits; after one that did not,
// a following segment without data (its marker missing, or left pending by the resynchronisation)
// still decodes its first MCU from zero bits before libjpeg notices — the rest stays zero.
__device__ __noinline__ void leave_region(const Stream& S, const HuffTables& T, int r, int b, int z, int blk, int dc0,
                                          int dc1, int dc2, uint32_t p, uint32_t end_pos, int total_blocks,
                                          int16_t* __restrict__ coef) {
  const int s = S.segment_of(r);
  if (s < 0) return;  // data libjpeg never decodes
  const int seg_end = min(total_blocks, (s + 1) * S.seg_blocks);
  int next = r + 1;  // the next region that is decoded at all
  while (next <= S.num_bounds && S.segment_of(next) < 0) ++next;
  const int next_seg = next <= S.num_bounds ? S.segment_of(next) : S.segments;
  if (z == 0 && b == 0 && blk == seg_end && next_seg == s + 1 && next == r + 1) return;  // intact
  atomicOr(S.damaged, 1);
  if (blk < seg_end) {
    TailState ts = {b, z, blk, dc0, dc1, dc2};
    conceal(S, T, ts, p, end_pos, seg_end, s, coef);
  } else if (next_seg > s + 1 && seg_end < total_blocks) {
    TailState ts = {0, 0, seg_end, 0, 0, 0};
    conceal(S, T, ts, end_pos, end_pos, min(total_blocks, seg_end + S.seg_blocks), s + 1, coef);
  }
}

// Decode from state (p, b, z) while p < end_bit.  kWrite: store coefficients with the DC
// predictors in agg.d*, `since` = blocks completed since the last restart boundary before p
// (a segment's blocks start at segment index x blocks per interval, as libjpeg restarts by MCU
// count); otherwise only count (agg.nb = blocks since the last boundary, agg.d* the diffs).
template <bool kWrite>
__device__ __forceinline__ void decode_range(const Stream& S, const HuffTables& T, uint32_t& p_io, int& b_io,
                                             int& z_io, uint32_t end_bit, Agg& agg, int since,
                                             int16_t* __restrict__ coef, int total_blocks) {
  uint32_t p = p_io;
  int b = b_io, z = z_io;
  // first boundary strictly after p; sitting exactly on one means "segment start"
  int lo = 0, hi = S.num_bounds;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(S.bounds + mid) * 8u > p) hi = mid;
    else lo = mid + 1;
  }
  int j = lo;
  int dc0 = agg.d0, dc1 = agg.d1, dc2 = agg.d2;
  if (j > 0 && __ldg(S.bounds + j - 1) * 8u == p) {
    if (kWrite && S.seg_blocks && (!(since == 0 && b == 0 && z == 0) || p == 0)) {
      // the predecessor stopped exactly here without stepping over the boundary (it would have
      // handed over a count of 0) — or there is none, the scan begins with a marker: the region
      // that ends here is looked after from this side
      int j0 = j - 1;  // several markers in a row: the regions between them are empty
      while (j0 > 0 && __ldg(S.bounds + j0 - 1) * 8u == p) --j0;
      leave_region(S, T, j0, b, z, S.segment_of(j0) * S.seg_blocks + since, dc0, dc1, dc2, p, p, total_blocks, coef);
      for (int e = j0 + 1; e < j; ++e)
        leave_region(S, T, e, 0, 0, S.segment_of(e) * S.seg_blocks, 0, 0, 0, p, p, total_blocks, coef);
    }
    b = 0, z = 0, agg.flag = 1;
    dc0 = dc1 = dc2 = 0;
    since = 0;
  }
  uint32_t next_bound = j < S.num_bounds ? __ldg(S.bounds + j) * 8u : 0xffffffffu;
  const uint32_t limit = min(end_bit, S.total_bits);
  uint32_t stop = min(S.total_bits, next_bound);  // no symbol may cross this bit
  uint32_t info = T.blk_info[b];                  // component | DC slot << 8 | AC slot << 16
  // Region j is restart segment s (s = j unless markers are damaged; -1: data libjpeg never
  // decodes, nothing is written).  Segment s owns blocks [s * seg_blocks, seg_end); what a
  // damaged segment decodes beyond them is dropped.
  int s = kWrite ? S.segment_of(j) : 0;
  int blk = s < 0 ? 0 : s * S.seg_blocks + since;
  int seg_end = s < 0 ? 0 : (S.seg_blocks ? min(total_blocks, (s + 1) * S.seg_blocks) : total_blocks);
  int16_t* blk_ptr = kWrite ? coef + (size_t)blk * 64 : nullptr;
  int nb = 0;
  uint32_t tail_from = 0xffffffffu;  // write pass: where the data ran out inside a symbol

  // three big-endian words of the stream: (w0:w1) holds bit p, w2 is fetched ahead
  uint32_t w0 = 0, w1 = 0, w2 = 0, widx = 0;
  auto reload = [&]() {
    widx = p >> 5;
    w0 = load_be(S.words, widx), w1 = load_be(S.words, widx + 1), w2 = load_be(S.words, widx + 2);
  };
  if (p < limit) reload();
  // The body has no data-dependent branches apart from the rare cases (a code longer than the
  // look-up, a restart boundary, the end of the stream, the end of a block): the 32 lanes of a
  // warp sit at unrelated places of their blocks, and every divergent path is paid by all.
  // (Stepping over a restart boundary happens outside the symbol loop: it may call the cold
  // concealment code, and a call inside the loop costs the loop its registers.)
  for (;;) {
  bool crossing = false;
  while (p < limit) {
    const uint32_t bits = __funnelshift_l(w1, w0, p);
    const bool dc_mode = z == 0;
    const uint32_t slot = dc_mode ? (info >> 8) & 255u : info >> 16;
    uint32_t e = T.lut[slot][bits >> (32 - kLutBits)];
    if (e == 0) e = huff_slow(T, (int)slot, bits);
    const uint32_t used = e & 255u, len = (e >> 8) & 255u, shift = (e >> 16) & 255u;
    const uint32_t p_next = p + used;
    if (p_next > stop) {
      // (a symbol that starts in the padding before a marker can reach past a short last segment
      // as well: the boundary comes first)
      if (next_bound > S.total_bits) {  // the padding at the very end: nothing more to decode
        if (kWrite) tail_from = p;  // ... unless blocks are missing: see below
        p = S.total_bits;
        break;
      }
      crossing = true;  // ran into the padding before a restart marker: new segment
      break;
    }
    // EXTEND (T.81 F.2.2.1): magnitude bits with a leading 0 mean value - (2^s - 1); s = 0 gives 0
    const uint32_t tail = bits << len;
    const int raw = (int)__funnelshift_rc(tail, 0u, shift);
    const int value = raw + (~((int)tail >> 31) & (1 - (int)(1u << (32u - shift))));
    if ((p_next ^ p) >> 5) {  // crossed into the next word
      w0 = w1, w1 = w2;
      ++widx;
      w2 = load_be(S.words, widx + 2);
    }
    p = p_next;
    const uint32_t comp = info & 255u;
    const int add = dc_mode ? value : 0;
    dc0 += comp == 0 ? add : 0;
    dc1 += comp == 1 ? add : 0;
    dc2 += comp == 2 ? add : 0;
    const int zn = z + (int)(e >> 24);
    if (kWrite) {
      // every stored symbol lands at zig-zag index zn - 1 (the DC at 0); ZRL / EOB store nothing.
      // A run past the end of the block (corrupt data) lands on the last coefficient:
      // jpeg_natural_order has 16 extra entries, all 63.
      if ((dc_mode || shift != 32u) && blk < seg_end)
        blk_ptr[T.natural[min(zn - 1, 63)]] = (int16_t)(dc_mode ? (comp == 0 ? dc0 : (comp == 1 ? dc1 : dc2)) : value);
    }
    z = zn;
    if (zn >= 64) {  // block complete
      z = 0;
      ++nb;
      if (kWrite) ++blk, blk_ptr += 64;
      b = b + 1 == S.bpm ? 0 : b + 1;
      info = T.blk_info[b];
    }
  }
  if (!crossing) break;
    {  // step over the boundary into the next region
      if (kWrite) {
        // too few or too many blocks, or markers that are not what they should be: concealment
        if (z != 0 || b != 0 || blk != seg_end || S.seg_of)
          leave_region(S, T, j, b, z, blk, dc0, dc1, dc2, p, next_bound, total_blocks, coef);
        s = S.segment_of(j + 1);
        blk = s < 0 ? 0 : s * S.seg_blocks;
        seg_end = s < 0 ? 0 : min(total_blocks, blk + S.seg_blocks);
        blk_ptr = coef + (size_t)blk * 64;
      }
      nb = 0;  // blocks are counted from the last boundary
      p = next_bound, b = 0, z = 0, agg.flag = 1;
      dc0 = dc1 = dc2 = 0;
      ++j;
      next_bound = j < S.num_bounds ? __ldg(S.bounds + j) * 8u : 0xffffffffu;
      while (kWrite && next_bound == p) {  // two markers in a row: an empty region, left at once
        leave_region(S, T, j, 0, 0, blk, 0, 0, 0, p, p, total_blocks, coef);
        ++j;
        s = S.segment_of(j);
        blk = s < 0 ? 0 : s * S.seg_blocks;
        seg_end = s < 0 ? 0 : min(total_blocks, blk + S.seg_blocks);
        blk_ptr = coef + (size_t)blk * 64;
        next_bound = j < S.num_bounds ? __ldg(S.bounds + j) * 8u : 0xffffffffu;
      }
      stop = min(S.total_bits, next_bound);
      info = T.blk_info[0];
      if (p < limit) reload();
    }
  }
  if (kWrite) {
    // Blocks are missing and the data is over (a file cut short and closed by a marker; intact
    // restart segments before it): libjpeg supplies zero bits, finishes the MCU in hand from
    // them — one whole MCU if the data ended exactly between two — and leaves the rest zero
    // (jdhuff.c decode_mcu, insufficient_data).  The counting passes do not see these blocks,
    // so the image is still reported as damaged.
    // one thread owns the end: the one that decoded up to it (the first one if there is no data)
    const bool owns_end = p_io < S.total_bits || (S.total_bits == 0 && end_bit == S.first_end_bit);
    if (tail_from == 0xffffffffu && p == S.total_bits && end_bit >= S.total_bits) tail_from = p;
    if (owns_end && s >= 0 && tail_from != 0xffffffffu && stop == S.total_bits && min(blk, seg_end) < total_blocks) {
      // (beyond this segment's blocks too when the data ended at or after its last MCU and the
      // restart that is due has no marker: one more MCU from zeros, predictors reset)
      TailState ts = {b, z, blk, dc0, dc1, dc2};
      uint32_t q = tail_from;
      if (blk >= seg_end) ts = {0, 0, seg_end, 0, 0, 0}, q = S.total_bits;  // what a damaged segment had beyond its blocks is skipped
      if (ts.blk < total_blocks) conceal(S, T, ts, q, S.total_bits, total_blocks, s, coef);
      p = S.total_bits;
    }
  }
  agg.nb += nb;
  if (p >= S.total_bits) p = S.total_bits, b = 0, z = 0;  // one canonical state past the end
  agg.d0 = dc0, agg.d1 = dc1, agg.d2 = dc2;
  p_io = p, b_io = b, z_io = z;
}

// Slots 0..2: the DC table of component 0..2; slots 3..5: the AC tables.  (Components sharing
// a table get identical copies: simpler than a mapping, and the copies are built in parallel.)
__device__ __forceinline__ void build_tables(const ia_jpeg_job& J, HuffTables& T) {
  const int tid = threadIdx.x;
  for (int i = tid; i < kHuffSlots * (1 << kLutBits); i += (int)blockDim.x) (&T.lut[0][0])[i] = 0;
  for (int i = tid; i < kHuffSlots * 256; i += (int)blockDim.x) {
    const int slot = i >> 8, k = i & 255, c = slot % 3;
    uint8_t v = 0;
    if (c < J.ncomp) v = slot < 3 ? (k < 16 ? J.dc_vals[J.comp_dc[c]][k] : 0) : J.ac_vals[J.comp_ac[c]][k];
    T.vals[slot][k] = v;
  }
  __syncthreads();
  if (tid < kHuffSlots) {
    const int slot = tid, c = slot % 3;
    const uint8_t* counts = slot < 3 ? J.dc_counts[J.comp_dc[c < J.ncomp ? c : 0]] : J.ac_counts[J.comp_ac[c < J.ncomp ? c : 0]];
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
      const int cnt = c < J.ncomp ? counts[l - 1] : 0;
      T.valoff[slot][l] = k - code;
      for (int n = 0; n < cnt; ++n, ++k, ++code) {
        if (l <= kLutBits) {
          const uint32_t e = pack_entry(slot, l, T.vals[slot][k & 255]);
          const int first = code << (kLutBits - l);
          for (int f = 0; f < (1 << (kLutBits - l)); ++f) T.lut[slot][(first + f) & ((1 << kLutBits) - 1)] = e;
        }
      }
      // codes of this length are < maxcode (left-aligned to 16 bits); a length without codes
      // inherits the previous bound so that it never matches
      T.maxcode[slot][l] = cnt ? (code << (16 - l)) : (l > 1 ? T.maxcode[slot][l - 1] : 0);
      code <<= 1;
    }
  } else if (tid >= 32 && tid < 48) {
    const int i = tid - 32;
    const int k = i < J.blocks_per_mcu ? i : 0;
    const int c = J.block_comp[k];
    T.blk_info[i] = (uint32_t)c | ((uint32_t)c << 8) | ((uint32_t)(3 + c) << 16);
  } else if (tid >= 64 && tid < 128) {
    constexpr uint8_t kNatural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                      12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                      35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                      58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    T.natural[tid - 64] = kNatural[tid - 64];
  }
  __syncthreads();
}

// Chunks of one image are chained: a chunk needs the exact decoder state and the running
// block count / DC predictors at its first bit, which only its predecessor knows.  Each CTA
// first converges on guesses (no waiting), then takes a ticket-ordered look at the
// predecessor's published record, repairs its first subsequences, publishes its own record and
// only then writes coefficients.  Tickets are handed out in launch order, so the CTA a chunk
// waits for is always already running.
template <int kThreads, int kMinBlocks, int kSub>
__global__ void __launch_bounds__(kThreads, kMinBlocks) huffman_kernel(const ia_jpeg_job* __restrict__ jobs, int max_chunks,
                                                                    uint32_t* __restrict__ ticket_counter,
                                                                    int32_t* __restrict__ status) {
  constexpr int kHuffThreads = kThreads;  // this instantiation's, not the namespace's
  constexpr int kSubseqBits = kSub * 8;
  constexpr uint32_t kChunkBits = (uint32_t)kHuffThreads * kSubseqBits;
  __shared__ HuffTables T;
  // per subsequence: assumed entry state, resulting exit state and contribution
  __shared__ uint32_t s_entry_p[kHuffThreads], s_entry_bz[kHuffThreads];
  __shared__ uint32_t s_exit_p[kHuffThreads], s_exit_bz[kHuffThreads];
  __shared__ int s_nb[kHuffThreads], s_d0[kHuffThreads], s_d1[kHuffThreads], s_d2[kHuffThreads];
  __shared__ uint8_t s_flag[kHuffThreads];
  __shared__ uint16_t s_task[kHuffThreads];
  __shared__ uint32_t s_votes[kHuffThreads / 32];
  constexpr int kHistory = 128;  // fronts of the last rounds, indexed by subsequence & 127
  __shared__ uint32_t s_hist_p[kHistory], s_hist_bz[kHistory];
  __shared__ int s_hist_t[kHistory];
  __shared__ Agg s_warp_agg[kHuffThreads / 32];
  __shared__ Agg s_carry;
  __shared__ uint32_t s_in_p, s_in_bz, s_ticket;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < kHistory) s_hist_t[tid] = -1;
  if (tid == 0) s_ticket = atomicAdd(ticket_counter, 1u);
  __syncthreads();
  const int image = (int)(s_ticket / (uint32_t)max_chunks), chunk = (int)(s_ticket % (uint32_t)max_chunks);
  const ia_jpeg_job& J = jobs[image];

  Stream S;
  S.words = reinterpret_cast<const uint32_t*>(J.clean_dev);
  S.total_bits = J.segment_dev[0] * 8u;
  S.bounds = J.segment_dev + 3;
  S.num_bounds = (int)J.segment_dev[1];
  S.bpm = J.blocks_per_mcu;
  S.seg_blocks = J.restart_interval * J.blocks_per_mcu;
  S.segments = J.num_segments;
  S.first_end_bit = (uint32_t)kSubseqBits;
  S.damaged = status + image;
  S.seg_of = (J.segment_dev[2] & 2u)
                 ? reinterpret_cast<const int32_t*>(J.segment_dev + 3 + region_capacity(J))
                 : nullptr;
  const int num_chunks = S.total_bits ? (int)((S.total_bits + kChunkBits - 1) / kChunkBits) : 1;
  if (chunk >= num_chunks) return;  // nothing waits on the spare CTAs of shorter images
  build_tables(J, T);

  const uint32_t chunk_begin = (uint32_t)chunk * kChunkBits;
  // decode subsequence t from its assumed entry state; results go to shared memory
  auto run = [&](int t) {
    uint32_t p = s_entry_p[t];
    const uint32_t bz = s_entry_bz[t];
    int b = (int)(bz >> 8), z = (int)(bz & 255u);
    Agg a;
    a.nb = 0, a.flag = 0, a.d0 = a.d1 = a.d2 = 0;
    decode_range<false>(S, T, p, b, z, chunk_begin + (uint32_t)(t + 1) * kSubseqBits, a, 0, nullptr, 0);
    s_exit_p[t] = p;
    s_exit_bz[t] = ((uint32_t)b << 8) | (uint32_t)z;
    s_nb[t] = a.nb, s_d0[t] = a.d0, s_d1[t] = a.d1, s_d2[t] = a.d2;
    s_flag[t] = (uint8_t)a.flag;
  };
  // Repeat: every subsequence whose predecessor ended in another state than assumed is decoded
  // again from that state.  The repeats thin out quickly (hundreds, tens, a few ...), so they
  // are compacted into a task list and run by the first threads of the CTA: a round with
  // three tasks costs one warp, not sixteen.
  int rounds = 0, tasks = 0;  // diagnostics, published with the chain record
  auto converge = [&](bool first_known) {
    int local_rounds = 0;  // of this call: the broadcasts below come after 16, 32, 64 ... of them
    for (;;) {
      __syncthreads();  // exits of the previous round are visible
      bool redo = false;
      uint32_t np = 0, nbz = 0;
      if (tid > 0) np = s_exit_p[tid - 1], nbz = s_exit_bz[tid - 1];
      else if (first_known) np = s_in_p, nbz = s_in_bz;
      if (tid > 0 || first_known) redo = np != s_entry_p[tid] || nbz != s_entry_bz[tid];
      const uint32_t votes = __ballot_sync(0xffffffffu, redo);
      if (lane == 0) s_votes[warp] = votes;
      __syncthreads();
      int base = 0, total = 0, front = kHuffThreads;  // front: the first subsequence that repeats
#pragma unroll
      for (int w = 0; w < kHuffThreads / 32; ++w) {
        const uint32_t v = s_votes[w];
        const int c = __popc(v);
        base += w < warp ? c : 0;
        total += c;
        if (v && front == kHuffThreads) front = w * 32 + __ffs(v) - 1;
      }
      if (total == 0) break;
      ++rounds, ++local_rounds, tasks += total;
      if (local_rounds >= kBroadcastEvery && (local_rounds & (local_rounds - 1)) == 0) {
        // Still repeating after this many rounds (ordinary streams settle in a dozen): the
        // stream does not resynchronise by itself — flat or periodic content, where every start
        // phase parses self-consistently — and the exact state would crawl forward one
        // subsequence per round.  In such a stream the state relative to the subsequence start
        // repeats (period dividing the subsequence length) or cycles through a few values; the
        // fronts of the last rounds, one subsequence apart and exact, show that cycle.  Give every
        // later subsequence the state of the recorded front congruent to it and decode them all
        // once; if the guess is wrong, the rounds simply go on from there.
        const uint32_t my_begin = chunk_begin + (uint32_t)tid * kSubseqBits;
        if (tid == front) {
          s_hist_p[front & (kHistory - 1)] = np - my_begin, s_hist_bz[front & (kHistory - 1)] = nbz;
          s_hist_t[front & (kHistory - 1)] = front;
        }
        __syncthreads();
        // smallest q < kHistory / 2 for which the last q fronts repeat the q before them (a cycle
        // longer than the rounds spent so far cannot be seen yet; the next broadcast will)
        int cycle = 1;
        for (int q = 1; q < kHistory / 2; ++q) {
          bool ok = true;
          for (int i = front; i > front - q && ok; --i) {
            const int a = i & (kHistory - 1), b = (i - q) & (kHistory - 1);
            ok = i - q >= 0 && s_hist_t[a] == i && s_hist_t[b] == i - q && s_hist_p[a] == s_hist_p[b] &&
                 s_hist_bz[a] == s_hist_bz[b];
          }
          if (ok) {
            cycle = q;
            break;
          }
        }
        // The guess must not cost a stream that IS converging (slowly: very high qualities need
        // dozens of rounds) what it has reached: keep the old results in registers, and keep
        // the guess only if it leaves more subsequences consistent with their predecessors.
        const uint32_t old_ep = s_entry_p[tid], old_ebz = s_entry_bz[tid];
        const uint32_t old_xp = s_exit_p[tid], old_xbz = s_exit_bz[tid];
        const int old_nb = s_nb[tid], old_d0 = s_d0[tid], old_d1 = s_d1[tid], old_d2 = s_d2[tid];
        const uint8_t old_flag = s_flag[tid];
        const int agreed_before = __syncthreads_count(tid > front && !redo);
        if (tid >= front) {
          const int j = (tid - front) % cycle;  // state of subsequence front - cycle + j (j > 0) or front
          const int from = (j ? front - cycle + j : front) & (kHistory - 1);
          // only the front's predecessor is known to be exact
          s_entry_p[tid] = tid == front ? np : min(my_begin + s_hist_p[from], S.total_bits);
          s_entry_bz[tid] = tid == front ? nbz : s_hist_bz[from];
          run(tid);
        }
        __syncthreads();
        const bool agrees = tid > front && s_entry_p[tid] == s_exit_p[tid - 1] && s_entry_bz[tid] == s_exit_bz[tid - 1];
        const int agreed_after = __syncthreads_count(agrees);
        // (a tie goes to the guess: a self-consistent but wrong parse of a periodic stream also
        // "agrees" everywhere, until the front reaches it)
        if (agreed_after < agreed_before && tid > front) {  // worse: put the old results back
          s_entry_p[tid] = old_ep, s_entry_bz[tid] = old_ebz;
          s_exit_p[tid] = old_xp, s_exit_bz[tid] = old_xbz;
          s_nb[tid] = old_nb, s_d0[tid] = old_d0, s_d1[tid] = old_d1, s_d2[tid] = old_d2;
          s_flag[tid] = old_flag;
        }
        continue;
      }
      if (tid == front) {  // remember the front of every round (see the broadcast above)
        s_hist_p[front & (kHistory - 1)] = np - (chunk_begin + (uint32_t)tid * kSubseqBits);
        s_hist_bz[front & (kHistory - 1)] = nbz;
        s_hist_t[front & (kHistory - 1)] = front;
      }
      if (redo) {
        s_entry_p[tid] = np, s_entry_bz[tid] = nbz;
        s_task[base + __popc(votes & ((1u << lane) - 1u))] = (uint16_t)tid;
      }
      __syncthreads();
      if (tid < total) run(s_task[tid]);
    }
  };
  s_entry_p[tid] = min(chunk_begin + (uint32_t)tid * kSubseqBits, S.total_bits);  // a guess (exact for
  s_entry_bz[tid] = 0;                                                             // chunk 0, thread 0)
  auto now_us = []() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return (uint32_t)(t >> 10);  // ~microseconds
  };
  uint32_t t_mark[6];  // diagnostics (thread 0): phase boundaries
  t_mark[0] = now_us();
  run(tid);
  __syncthreads();
  t_mark[1] = now_us();
  converge(false);
  t_mark[2] = now_us();

  // Hand-over between the chunks of one image.  Record of a chunk (kChainWords words):
  //   [0] tentative flag, [1..2] exit state after converging on guesses — right with near
  //       certainty (the last subsequence is 64 KB downstream of the guess), so the successor
  //       repairs its head from it at once, without waiting for this chunk's own repair;
  //   [3] final flag, [4..5] exit state after the repair, [6..9] blocks and DC predictors up
  //       to the end of this chunk, [10] diagnostics.
  // The final records form the only serial chain, and a link of it is a compare and five adds.
  uint32_t* my_rec = J.chain_dev + (size_t)chunk * kChainWords;
  const volatile uint32_t* prev_rec = J.chain_dev + (size_t)(chunk > 0 ? chunk - 1 : 0) * kChainWords;
  __shared__ int s_used_final, s_mismatch;
  if (tid == 0) {
    my_rec[1] = s_exit_p[kHuffThreads - 1], my_rec[2] = s_exit_bz[kHuffThreads - 1];
    __threadfence();
    *reinterpret_cast<volatile uint32_t*>(my_rec) = 1u;
    if (chunk == 0) {
      s_in_p = 0, s_in_bz = 0, s_used_final = 1;
    } else {
      for (;;) {
        if (prev_rec[3]) {
          __threadfence();
          s_in_p = prev_rec[4], s_in_bz = prev_rec[5], s_used_final = 1;
          break;
        }
        if (prev_rec[0]) {
          __threadfence();
          s_in_p = prev_rec[1], s_in_bz = prev_rec[2], s_used_final = 0;
          break;
        }
        __nanosleep(32);
      }
    }
  }
  converge(true);  // its first barrier publishes s_in_*

  // CTA scan of the contributions (without the carry): per-thread inclusive value, warp totals
  Agg incl;
  auto local_scan = [&]() {
    incl.nb = s_nb[tid], incl.flag = s_flag[tid], incl.d0 = s_d0[tid], incl.d1 = s_d1[tid], incl.d2 = s_d2[tid];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const Agg o = shfl_up_agg(incl, d);
      if (lane >= d) incl = combine(o, incl);
    }
    if (lane == 31) s_warp_agg[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      Agg w;
      if (lane < kHuffThreads / 32) w = s_warp_agg[lane];
      else w.nb = w.flag = w.d0 = w.d1 = w.d2 = 0;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const Agg o = shfl_up_agg(w, d);
        if (lane >= d) w = combine(o, w);
      }
      if (lane < kHuffThreads / 32) s_warp_agg[lane] = w;
    }
    __syncthreads();
  };
  local_scan();
  t_mark[3] = now_us();

  // the predecessor's final record: running totals, and a check of the state this chunk assumed
  if (tid == 0) {
    s_mismatch = 0;
    s_carry.nb = 0, s_carry.flag = 0, s_carry.d0 = s_carry.d1 = s_carry.d2 = 0;
    if (chunk > 0) {
      while (prev_rec[3] == 0) __nanosleep(32);
      __threadfence();
      const uint32_t fp = prev_rec[4], fbz = prev_rec[5];
      s_carry.nb = (int)prev_rec[6];
      s_carry.d0 = (int)prev_rec[7], s_carry.d1 = (int)prev_rec[8], s_carry.d2 = (int)prev_rec[9];
      if (!s_used_final && (fp != s_in_p || fbz != s_in_bz)) s_in_p = fp, s_in_bz = fbz, s_mismatch = 1;
    }
  }
  __syncthreads();
  if (s_mismatch) {  // the tentative state was wrong after all (uniform branch): repair again
    converge(true);
    local_scan();
  }
  t_mark[4] = now_us();
  Agg before = s_carry;  // everything before this thread: carry (+) earlier warps (+) earlier lanes
  if (warp > 0) before = combine(before, s_warp_agg[warp - 1]);
  {
    const Agg prev_lane = shfl_up_agg(incl, 1);
    if (lane > 0) before = combine(before, prev_lane);
  }
  if (tid == 0) {  // publish before writing coefficients: the successor does not need them
    const Agg out = combine(s_carry, s_warp_agg[kHuffThreads / 32 - 1]);
    my_rec[4] = s_exit_p[kHuffThreads - 1], my_rec[5] = s_exit_bz[kHuffThreads - 1];
    my_rec[6] = (uint32_t)out.nb, my_rec[7] = (uint32_t)out.d0, my_rec[8] = (uint32_t)out.d1;
    my_rec[9] = (uint32_t)out.d2;
    my_rec[10] = (uint32_t)rounds | ((uint32_t)min(tasks, 0xffff) << 16);
    __threadfence();
    *reinterpret_cast<volatile uint32_t*>(my_rec + 3) = 1u;
    if (chunk == num_chunks - 1) {
      // 1: the stream held fewer or more blocks than the frame (damaged: libjpeg conceals, pixels
      // of the damaged part differ); 2: and it simply ends — no marker behind it — before the
      // last block: libjpeg runs out of input there and OpenCV reports a failed decode
      // (out.nb counts from the last restart boundary; the segments before it are checked where
      // the write pass steps over their boundaries)
      const int want_nb = J.total_blocks - S.num_bounds * S.seg_blocks;
      int st = out.nb == want_nb ? 0 : 1;
      if (out.nb < want_nb && (J.segment_dev[2] & 1u)) st |= 2;
      if (st) atomicOr(status + image, st);
    }
  }

  // ---- write pass from the exact entry state
  Agg w;
  w.nb = 0, w.flag = 0, w.d0 = before.d0, w.d1 = before.d1, w.d2 = before.d2;
  uint32_t p = s_entry_p[tid];
  int b = (int)(s_entry_bz[tid] >> 8), z = (int)(s_entry_bz[tid] & 255u);
  decode_range<true>(S, T, p, b, z, chunk_begin + (uint32_t)(tid + 1) * kSubseqBits, w, before.nb, J.coef_dev,
                     J.total_blocks);
  __syncthreads();
  if (tid == 0) {  // phase durations for tools: guesses, repeats, repair, wait for the chain, write
    t_mark[5] = now_us();
    for (int k = 0; k < 5; ++k) my_rec[11 + k] = t_mark[k + 1] - t_mark[k];
  }
}

// =================================================================================== IDCT
constexpr int kIdctThreads = 128;

// One 8-point pass of jidctint.c (Loeffler-Ligtenberg-Moschytz, CONST_BITS 13) evaluated the way
// libjpeg-turbo's SIMD code does (jidctint-sse2 / -avx2), which is what OpenCV runs: the same
// integers as the C code for every block an encoder can produce, and for out-of-range
// coefficients (corrupt streams, broken tables) the same wrap-arounds — 16-bit lanes in the
// sums in0 +- in4, in7 + in3, in5 + in1; rotations as sums of two 16x16 products with the
// constants pre-combined (pmaddwd); 32-bit sums that wrap; results descaled and packed to
// 16 bits with saturation.  Inputs hold 16-bit values.
__device__ __forceinline__ int wrap16(int x) { return (int)(int16_t)(uint16_t)(uint32_t)x; }

// `left16` collects (as non-zero bits) every place where a 16-bit lane wrapped or a result was
// saturated: while it stays 0 the SIMD evaluation is plain jidctint.c integer arithmetic.
template <int kShift>
__device__ __forceinline__ void idct8(const int d[8], int o[8], uint32_t& left16) {
  typedef uint32_t u;
  const u tmp3 = (u)(d[2] * (4433 + 6270)) + (u)(d[6] * 4433);
  const u tmp2 = (u)(d[2] * 4433) + (u)(d[6] * (4433 - 15137));
  const int s04 = d[0] + d[4], d04 = d[0] - d[4], s73 = d[7] + d[3], s51 = d[5] + d[1];
  const u tmp0 = (u)(wrap16(s04) * 8192), tmp1 = (u)(wrap16(d04) * 8192);
  const u tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  const int z3 = wrap16(s73), z4 = wrap16(s51);
  left16 |= (u)(s04 ^ wrap16(s04)) | (u)(d04 ^ wrap16(d04)) | (u)(s73 ^ z3) | (u)(s51 ^ z4);
  const u z3n = (u)(z3 * (9633 - 16069)) + (u)(z4 * 9633);
  const u z4n = (u)(z3 * 9633) + (u)(z4 * (9633 - 3196));
  const u t0 = (u)(d[7] * (2446 - 7373)) + (u)(d[1] * -7373) + z3n;
  const u t3 = (u)(d[7] * -7373) + (u)(d[1] * (12299 - 7373)) + z4n;
  const u t1 = (u)(d[5] * (16819 - 20995)) + (u)(d[3] * -20995) + z4n;
  const u t2 = (u)(d[5] * -20995) + (u)(d[3] * (25172 - 20995)) + z3n;
  auto fin = [&left16](u x) -> int {
    const int v = (int)(x + (1u << (kShift - 1))) >> kShift;
    const int c = max(-32768, min(32767, v));
    left16 |= (u)(v ^ c);
    return c;
  };
  o[0] = fin(tmp10 + t3), o[7] = fin(tmp10 - t3), o[1] = fin(tmp11 + t2), o[6] = fin(tmp11 - t2);
  o[2] = fin(tmp12 + t1), o[5] = fin(tmp12 - t1), o[3] = fin(tmp13 + t0), o[4] = fin(tmp13 - t0);
}

__global__ void __launch_bounds__(kIdctThreads) idct_kernel(const ia_jpeg_job* __restrict__ jobs,
                                                            int32_t* __restrict__ status) {
  const ia_jpeg_job& J = jobs[blockIdx.y];
  __shared__ uint16_t s_quant[IA_JPEG_MAX_COMPONENTS][64];
  for (int i = threadIdx.x; i < IA_JPEG_MAX_COMPONENTS * 64; i += kIdctThreads)
    (&s_quant[0][0])[i] = (&J.quant[0][0])[i];
  __syncthreads();
  const int g = blockIdx.x * kIdctThreads + threadIdx.x;
  if (g >= J.total_blocks) return;
  const int bpm = J.blocks_per_mcu;
  const int mcu = g / bpm, k = g - mcu * bpm;
  const int my = mcu / J.mcux, mx = mcu - my * J.mcux;
  const int c = J.block_comp[k];
  const int bx = mx * J.h[c] + J.block_bx[k], by = my * J.v[c] + J.block_by[k];
  const int pitch = J.mcux * J.h[c] * 8;
  uint8_t* __restrict__ dst = J.plane_dev[c] + (size_t)by * 8 * pitch + (size_t)bx * 8;

  const uint4* __restrict__ src = reinterpret_cast<const uint4*>(J.coef_dev + (size_t)g * 64);
  const uint16_t* q = s_quant[c];
  int ws[64];
  uint32_t left16 = 0;  // a 16-bit lane wrapped / saturated somewhere in this block
  uint32_t below = 0;   // any coefficient in rows 1..7
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const uint4 v = __ldg(src + r);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    if (r) below |= v.x | v.y | v.z | v.w;
#pragma unroll
    for (int i = 0; i < 4; ++i) {  // dequantise in a 16-bit lane (pmullw)
      const int lo = (int)(int16_t)(w[i] & 0xffffu) * (int)q[r * 8 + 2 * i];
      const int hi = (int)(int16_t)(w[i] >> 16) * (int)q[r * 8 + 2 * i + 1];
      ws[r * 8 + 2 * i] = wrap16(lo);
      ws[r * 8 + 2 * i + 1] = wrap16(hi);
      left16 |= (uint32_t)(lo ^ wrap16(lo)) | (uint32_t)(hi ^ wrap16(hi));
    }
  }
  if (below == 0) {
    // rows 1..7 empty: every column is its row-0 term << PASS1_BITS, in a 16-bit lane
#pragma unroll
    for (int col = 0; col < 8; ++col) {
      const int v = wrap16(ws[col] << 2);
      left16 |= (uint32_t)((ws[col] << 2) ^ v);
#pragma unroll
      for (int r = 0; r < 8; ++r) ws[r * 8 + col] = v;
    }
  } else {
    // pass 1: columns, descale by CONST_BITS - PASS1_BITS = 11
#pragma unroll
    for (int col = 0; col < 8; ++col) {
      int d[8], o[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) d[r] = ws[r * 8 + col];
      idct8<11>(d, o, left16);
#pragma unroll
      for (int r = 0; r < 8; ++r) ws[r * 8 + col] = o[r];
    }
  }
  // pass 2: rows, descale by CONST_BITS + PASS1_BITS + 3 = 18, clamp to [-128, 127], + 128, store
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    int d[8], o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) d[i] = ws[r * 8 + i];
    idct8<18>(d, o, left16);
    uint32_t px[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) px[i] = (uint32_t)(max(-128, min(127, o[i])) + 128);
    uint2 out;
    out.x = px[0] | (px[1] << 8) | (px[2] << 16) | (px[3] << 24);
    out.y = px[4] | (px[5] << 8) | (px[6] << 16) | (px[7] << 24);
    *reinterpret_cast<uint2*>(dst + (size_t)r * pitch) = out;
  }
  // Coefficients no encoder produces (damaged data) pushed a 16-bit lane of the SIMD evaluation over
  // its range.  The wrap-arounds above follow libjpeg-turbo's code, but only the in-range case is
  // plain jidctint.c arithmetic and therefore certain to be cv2's pixels: such a frame is reported
  // as damaged (status bit 0), never as clean.
  if (left16) atomicOr(status + blockIdx.y, 1);
}

// ====================================================================== upsample + colour
__device__ __forceinline__ int chroma_sample(const uint8_t* __restrict__ plane, int pitch, int dw, int dh, int hexp,
                                             int vexp, int x, int y) {
  // jdsample.c: fancy upsampling for factors of two (h2v1 / h2v2 only when the subsampled
  // width exceeds 2), plain replication otherwise
  if (hexp == 1 && vexp == 1) return plane[(size_t)y * pitch + x];
  if (hexp == 2 && vexp == 1 && dw > 2) {
    const int xc = x >> 1;
    const uint8_t* row = plane + (size_t)y * pitch;
    const int a = row[xc];
    if (x == 0 || x == 2 * dw - 1) return a;
    return (x & 1) ? (3 * a + row[xc + 1] + 2) >> 2 : (3 * a + row[xc - 1] + 1) >> 2;
  }
  if (hexp == 2 && vexp == 2 && dw > 2) {
    const int xc = x >> 1, yc = y >> 1;
    int yo = (y & 1) ? yc + 1 : yc - 1;
    yo = yo < 0 ? 0 : (yo > dh - 1 ? dh - 1 : yo);
    const uint8_t* r0 = plane + (size_t)yc * pitch;
    const uint8_t* r1 = plane + (size_t)yo * pitch;
    const int cur = 3 * r0[xc] + r1[xc];
    const int xo = (x & 1) ? xc + 1 : xc - 1;
    const int bias = (x & 1) ? 7 : 8;
    if (xo < 0 || xo > dw - 1) return (4 * cur + bias) >> 4;
    return (3 * cur + 3 * r0[xo] + r1[xo] + bias) >> 4;
  }
  if (hexp == 1 && vexp == 2) {
    const int yc = y >> 1;
    int yo = (y & 1) ? yc + 1 : yc - 1;
    yo = yo < 0 ? 0 : (yo > dh - 1 ? dh - 1 : yo);
    return (3 * plane[(size_t)yc * pitch + x] + plane[(size_t)yo * pitch + x] + ((y & 1) ? 2 : 1)) >> 2;
  }
  return plane[(size_t)(y / vexp) * pitch + x / hexp];
}

constexpr int kColourPixels = 8;  // consecutive pixels of a row per thread

__global__ void __launch_bounds__(256) colour_kernel(const ia_jpeg_job* __restrict__ jobs) {
  const ia_jpeg_job& J = jobs[blockIdx.z];
  const int x0 = (blockIdx.x * 32 + (threadIdx.x & 31)) * kColourPixels, y = blockIdx.y * 8 + (threadIdx.x >> 5);
  const int W = J.width, H = J.height;
  if (x0 >= W || y >= H) return;
  const uint8_t* __restrict__ yrow = J.plane_dev[0] + (size_t)y * J.comp_pitch[0];
  const int ncomp = J.ncomp, orientation = J.orientation;
  const int count = min(kColourPixels, W - x0);
  // chroma of the eight pixels.  Interior groups of the common layouts (4:2:0, 4:2:2) read one
  // aligned word + two bytes per chroma row; everything else (image borders, other factors,
  // tiny planes) goes pixel by pixel through chroma_sample.
  int chroma[2][kColourPixels];
  if (ncomp == 3) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int hexp = J.hexp[k + 1], vexp = J.vexp[k + 1], dw = J.comp_w[k + 1], dh = J.comp_h[k + 1];
      const int pitch = J.comp_pitch[k + 1];
      const uint8_t* __restrict__ plane = J.plane_dev[k + 1];
      const int xc = x0 >> 1;  // a multiple of 4: the four samples xc..xc+3 are one aligned word
      if (hexp == 2 && vexp <= 2 && count == kColourPixels && xc + 3 <= dw - 1 && dw > 2) {
        // the first / last group of a row has no left / right neighbour sample: jdsample.c's
        // edge formulas for those two pixels, everything else as in the interior
        const bool has_left = xc >= 1, has_right = xc + 4 <= dw - 1;
        const int yc = vexp == 2 ? y >> 1 : y;
        int yo = yc;
        if (vexp == 2) {
          yo = (y & 1) ? yc + 1 : yc - 1;
          yo = yo < 0 ? 0 : (yo > dh - 1 ? dh - 1 : yo);
        }
        const uint8_t* r0 = plane + (size_t)yc * pitch + xc;
        const uint8_t* r1 = plane + (size_t)yo * pitch + xc;
        const uint32_t w0 = *reinterpret_cast<const uint32_t*>(r0), w1 = *reinterpret_cast<const uint32_t*>(r1);
        int near_[6], far_[6];  // samples xc-1 .. xc+4 of the near and the far chroma row
        near_[0] = r0[has_left ? -1 : 0], near_[5] = r0[has_right ? 4 : 3];
        far_[0] = r1[has_left ? -1 : 0], far_[5] = r1[has_right ? 4 : 3];
#pragma unroll
        for (int i = 0; i < 4; ++i) near_[i + 1] = (w0 >> (8 * i)) & 255u, far_[i + 1] = (w1 >> (8 * i)) & 255u;
        int cs[6];  // h2v2: 3 * near row + far row ("colsum"); h2v1: the sample itself
#pragma unroll
        for (int i = 0; i < 6; ++i) cs[i] = vexp == 2 ? 3 * near_[i] + far_[i] : near_[i];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (vexp == 2) {
            chroma[k][2 * i] = (3 * cs[i + 1] + cs[i] + 8) >> 4;
            chroma[k][2 * i + 1] = (3 * cs[i + 1] + cs[i + 2] + 7) >> 4;
          } else {
            chroma[k][2 * i] = (3 * cs[i + 1] + cs[i] + 1) >> 2;
            chroma[k][2 * i + 1] = (3 * cs[i + 1] + cs[i + 2] + 2) >> 2;
          }
        }
        if (!has_left) chroma[k][0] = vexp == 2 ? (4 * cs[1] + 8) >> 4 : cs[1];
        if (!has_right) chroma[k][7] = vexp == 2 ? (4 * cs[4] + 7) >> 4 : cs[4];
      } else if (hexp == 1 && vexp == 1) {
#pragma unroll
        for (int i = 0; i < kColourPixels; ++i) chroma[k][i] = plane[(size_t)y * pitch + min(x0 + i, W - 1)];
      } else {
#pragma unroll
        for (int i = 0; i < kColourPixels; ++i)
          chroma[k][i] = chroma_sample(plane, pitch, dw, dh, hexp, vexp, min(x0 + i, W - 1), y);
      }
    }
  }
  uint8_t rgb[kColourPixels * 3];
  int luma[kColourPixels];
  if (count == kColourPixels) {  // x0 is a multiple of 8 and so is the pitch: one aligned 64-bit load
    const uint2 yv = *reinterpret_cast<const uint2*>(yrow + x0);
#pragma unroll
    for (int i = 0; i < 4; ++i) luma[i] = (yv.x >> (8 * i)) & 255u, luma[i + 4] = (yv.y >> (8 * i)) & 255u;
  } else {
#pragma unroll
    for (int i = 0; i < kColourPixels; ++i) luma[i] = yrow[min(x0 + i, W - 1)];
  }
#pragma unroll
  for (int i = 0; i < kColourPixels; ++i) {
    const int yy = luma[i];
    int r, g, b;
    if (ncomp == 1) {
      r = g = b = yy;
    } else {
      const int cb = chroma[0][i] - 128, cr = chroma[1][i] - 128;
      // jdcolor.c build_ycc_rgb_table: FIX(1.40200)=91881, FIX(1.77200)=116130,
      // FIX(0.71414)=46802, FIX(0.34414)=22554, ONE_HALF=32768, arithmetic shifts
      r = yy + ((91881 * cr + 32768) >> 16);
      g = yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
      b = yy + ((116130 * cb + 32768) >> 16);
      r = r < 0 ? 0 : (r > 255 ? 255 : r);
      g = g < 0 ? 0 : (g > 255 ? 255 : g);
      b = b < 0 ? 0 : (b > 255 ? 255 : b);
    }
    rgb[3 * i] = (uint8_t)r, rgb[3 * i + 1] = (uint8_t)g, rgb[3 * i + 2] = (uint8_t)b;
  }
  if (orientation == 1) {
    uint8_t* o = J.out_dev + ((size_t)y * W + x0) * 3;
    if (count == kColourPixels && ((uintptr_t)o & 3) == 0) {
      uint32_t* o32 = reinterpret_cast<uint32_t*>(o);
#pragma unroll
      for (int k = 0; k < kColourPixels * 3 / 4; ++k)
        o32[k] = rgb[4 * k] | (rgb[4 * k + 1] << 8) | (rgb[4 * k + 2] << 16) | ((uint32_t)rgb[4 * k + 3] << 24);
    } else {
#pragma unroll
      for (int k = 0; k < kColourPixels * 3; ++k)  // constant indices: rgb[] stays in registers
        if (k < count * 3) o[k] = rgb[k];
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < kColourPixels; ++i) {  // OpenCV's ExifTransform
    if (i >= count) break;
    const int x = x0 + i;
    int ox = x, oy = y;
    switch (orientation) {
      case 2: ox = W - 1 - x; break;
      case 3: ox = W - 1 - x, oy = H - 1 - y; break;
      case 4: oy = H - 1 - y; break;
      case 5: ox = y, oy = x; break;
      case 6: ox = H - 1 - y, oy = x; break;
      case 7: ox = H - 1 - y, oy = W - 1 - x; break;
      case 8: ox = y, oy = W - 1 - x; break;
      default: break;
    }
    uint8_t* o = J.out_dev + ((size_t)oy * J.out_width + ox) * 3;
    o[0] = rgb[3 * i], o[1] = rgb[3 * i + 1], o[2] = rgb[3 * i + 2];
  }
}

}  // namespace jpeg
}  // namespace ia
