// ia_jpeg.cu — SURVEY §8f row N3: baseline JPEG decode on the device, bit-identical to what
// the reference's load_image_from_bytes gets from OpenCV/libjpeg-turbo
// (src/shared/processing/transforms.py:77-110).  oracle/jpeg.py restates the algorithm.
//
// Four kernels per batch, one launch each:
//   J1 unstuff_kernel   one CTA per image: drops the FF00 stuffing and the RSTn markers from
//                       the entropy-coded bytes (CTA-wide scan + scatter), records where each
//                       restart segment begins, finds the end of the scan.
//   J2 huffman_kernel   one CTA per image, one thread per 128-byte subsequence of the clean
//                       stream.  Huffman codes cannot be cut at arbitrary bits, so every
//                       thread first decodes its subsequence from a guessed state (bit 0 of
//                       the subsequence, start of a block); a thread whose predecessor ended
//                       in a different state (bit position, block within the MCU, zig-zag
//                       index) decodes again from that state, until nothing changes — JPEG
//                       codes resynchronise within a few symbols, so this takes a handful of
//                       rounds instead of one serial walk (self-synchronising decode,
//                       Weissenberger & Schmidt 2018/2021).  A CTA scan of per-thread block
//                       counts and DC sums then gives every thread its output block and DC
//                       predictor, and one more pass writes the coefficients.
//   J3 idct_kernel      one thread per 8x8 block: dequantise + libjpeg's jidctint.c "islow"
//                       integer IDCT -> component planes.
//   J4 colour_kernel    one thread per pixel: libjpeg's fancy (triangle) chroma upsampling,
//                       jdcolor.c fixed-point YCbCr->RGB, EXIF orientation, RGB store.
#include <string.h>

#include "ia_common.cuh"

namespace ia {
namespace jpeg {

constexpr int kSubseqBytes = 128;
constexpr int kSubseqBits = kSubseqBytes * 8;
constexpr int kHuffThreads = 512;
constexpr int kLutBits = 10;
constexpr int kUnstuffThreads = 512;
constexpr int kUnstuffPerThread = 16;
constexpr int kCleanPad = 64;  // readable zero bytes after the clean stream

__constant__ uint8_t c_natural[80] = {
    0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13,
    6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31,
    39, 46, 53, 60, 61, 54, 47, 55, 62, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

static const uint8_t h_natural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                      12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                      35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                      58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// =========================================================================== host: headers
static inline uint32_t be16(const uint8_t* p) { return ((uint32_t)p[0] << 8) | p[1]; }

// Orientation tag (0x0112) of an APP1 "Exif" payload; 1 when absent or malformed.
static int exif_orientation(const uint8_t* seg, size_t len) {
  if (len < 14 || memcmp(seg, "Exif\0\0", 6) != 0) return 1;
  const uint8_t* t = seg + 6;
  const size_t n = len - 6;
  bool le;
  if (t[0] == 'I' && t[1] == 'I') le = true;
  else if (t[0] == 'M' && t[1] == 'M') le = false;
  else return 1;
  auto rd16 = [&](size_t o) -> uint32_t { return le ? (t[o] | (t[o + 1] << 8)) : ((t[o] << 8) | t[o + 1]); };
  auto rd32 = [&](size_t o) -> uint32_t {
    return le ? (t[o] | (t[o + 1] << 8) | (t[o + 2] << 16) | ((uint32_t)t[o + 3] << 24))
              : (((uint32_t)t[o] << 24) | (t[o + 1] << 16) | (t[o + 2] << 8) | t[o + 3]);
  };
  if (n < 8 || rd16(2) != 42) return 1;
  const size_t ifd = rd32(4);
  if (ifd + 2 > n) return 1;
  const uint32_t entries = rd16(ifd);
  for (uint32_t e = 0; e < entries; ++e) {
    const size_t o = ifd + 2 + 12 * (size_t)e;
    if (o + 12 > n) break;
    if (rd16(o) == 0x0112) {
      const uint32_t v = rd16(o + 8);
      return (v >= 1 && v <= 8) ? (int)v : 1;
    }
  }
  return 1;
}

static int parse(const uint8_t* d, size_t n, ia_jpeg_job* job) {
  memset(job, 0, sizeof(*job));
  IA_RETURN_IF(n < 4 || d[0] != 0xFF || d[1] != 0xD8, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(n >= 0xfffffff0ull, IA_ERR_UNSUPPORTED);
  uint16_t qt[4][64];
  bool have_qt[4] = {false, false, false, false}, have_dc[4] = {false, false, false, false},
       have_ac[4] = {false, false, false, false};
  int tq[IA_JPEG_MAX_COMPONENTS] = {0, 0, 0}, comp_id[IA_JPEG_MAX_COMPONENTS] = {0, 0, 0};
  bool have_sof = false, saw_exif = false;
  int adobe_transform = -1;
  job->orientation = 1;
  size_t i = 2;
  for (;;) {
    while (i < n && d[i] != 0xFF) ++i;
    while (i < n && d[i] == 0xFF) ++i;
    IA_RETURN_IF(i >= n, IA_ERR_INVALID_ARGUMENT);
    const uint8_t m = d[i++];
    if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;
    IA_RETURN_IF(m == 0xD9, IA_ERR_INVALID_ARGUMENT);
    IA_RETURN_IF(i + 2 > n, IA_ERR_INVALID_ARGUMENT);
    const size_t L = be16(d + i);
    IA_RETURN_IF(L < 2 || i + L > n, IA_ERR_INVALID_ARGUMENT);
    const uint8_t* seg = d + i + 2;
    const size_t len = L - 2;
    if (m == 0xDB) {
      size_t k = 0;
      while (k < len) {
        const int pq = seg[k] >> 4, t = seg[k] & 15;
        ++k;
        IA_RETURN_IF(t > 3 || pq > 1 || k + (pq ? 128 : 64) > len, IA_ERR_INVALID_ARGUMENT);
        for (int j = 0; j < 64; ++j) {
          qt[t][h_natural[j]] = pq ? (uint16_t)be16(seg + k + 2 * j) : seg[k + j];
        }
        k += pq ? 128 : 64;
        have_qt[t] = true;
      }
    } else if (m == 0xC0 || m == 0xC1) {
      IA_RETURN_IF(len < 6, IA_ERR_INVALID_ARGUMENT);
      IA_RETURN_IF(seg[0] != 8, IA_ERR_UNSUPPORTED);
      job->height = (int32_t)be16(seg + 1);
      job->width = (int32_t)be16(seg + 3);
      job->ncomp = seg[5];
      IA_RETURN_IF(job->ncomp != 1 && job->ncomp != 3, IA_ERR_UNSUPPORTED);
      IA_RETURN_IF(len < 6 + 3 * (size_t)job->ncomp, IA_ERR_INVALID_ARGUMENT);
      for (int c = 0; c < job->ncomp; ++c) {
        comp_id[c] = seg[6 + 3 * c];
        job->h[c] = seg[7 + 3 * c] >> 4;
        job->v[c] = seg[7 + 3 * c] & 15;
        tq[c] = seg[8 + 3 * c];
        IA_RETURN_IF(job->h[c] < 1 || job->h[c] > 4 || job->v[c] < 1 || job->v[c] > 4 || tq[c] > 3,
                     IA_ERR_INVALID_ARGUMENT);
      }
      have_sof = true;
    } else if (m == 0xC2 || m == 0xC3 || (m >= 0xC5 && m <= 0xC7) || (m >= 0xC9 && m <= 0xCB) ||
               (m >= 0xCD && m <= 0xCF)) {
      return IA_ERR_UNSUPPORTED;  // progressive, lossless, arithmetic
    } else if (m == 0xC4) {
      size_t k = 0;
      while (k < len) {
        IA_RETURN_IF(k + 17 > len, IA_ERR_INVALID_ARGUMENT);
        const int tc = seg[k] >> 4, th = seg[k] & 15;
        IA_RETURN_IF(tc > 1 || th > 3, IA_ERR_INVALID_ARGUMENT);
        int total = 0;
        for (int j = 0; j < 16; ++j) total += seg[k + 1 + j];
        IA_RETURN_IF(total > (tc ? 256 : 16) || k + 17 + total > len, IA_ERR_INVALID_ARGUMENT);
        uint8_t* counts = tc ? job->ac_counts[th] : job->dc_counts[th];
        uint8_t* vals = tc ? job->ac_vals[th] : job->dc_vals[th];
        memcpy(counts, seg + k + 1, 16);
        memset(vals, 0, tc ? 256 : 16);
        memcpy(vals, seg + k + 17, total);
        (tc ? have_ac : have_dc)[th] = true;
        k += 17 + total;
      }
    } else if (m == 0xDD) {
      IA_RETURN_IF(len < 2, IA_ERR_INVALID_ARGUMENT);
      job->restart_interval = (int32_t)be16(seg);
    } else if (m == 0xE1 && !saw_exif && len >= 6 && memcmp(seg, "Exif\0\0", 6) == 0) {
      saw_exif = true;
      job->orientation = exif_orientation(seg, len);
    } else if (m == 0xEE && len >= 12 && memcmp(seg, "Adobe", 5) == 0) {
      adobe_transform = seg[11];
    } else if (m == 0xDA) {
      IA_RETURN_IF(!have_sof || len < 1, IA_ERR_INVALID_ARGUMENT);
      const int ns = seg[0];
      IA_RETURN_IF(ns != job->ncomp, IA_ERR_UNSUPPORTED);  // non-interleaved / multi-scan
      IA_RETURN_IF(len < 1 + 2 * (size_t)ns + 3, IA_ERR_INVALID_ARGUMENT);
      for (int s = 0; s < ns; ++s) {
        IA_RETURN_IF(seg[1 + 2 * s] != comp_id[s], IA_ERR_UNSUPPORTED);
        job->comp_dc[s] = seg[2 + 2 * s] >> 4;
        job->comp_ac[s] = seg[2 + 2 * s] & 15;
        IA_RETURN_IF(job->comp_dc[s] > 3 || job->comp_ac[s] > 3, IA_ERR_INVALID_ARGUMENT);
        IA_RETURN_IF(!have_dc[job->comp_dc[s]] || !have_ac[job->comp_ac[s]], IA_ERR_INVALID_ARGUMENT);
      }
      job->scan_offset = (uint32_t)(i + L);
      break;
    }
    i += L;
  }
  IA_RETURN_IF(job->width < 1 || job->height < 1, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(job->ncomp == 3 && adobe_transform == 0, IA_ERR_UNSUPPORTED);  // stored as RGB
  if (job->ncomp == 1) job->h[0] = job->v[0] = 1;  // sampling factors mean nothing for one component
  job->hmax = job->vmax = 1;
  int bpm = 0;
  for (int c = 0; c < job->ncomp; ++c) {
    IA_RETURN_IF(!have_qt[tq[c]], IA_ERR_INVALID_ARGUMENT);
    memcpy(job->quant[c], qt[tq[c]], sizeof(qt[0]));
    job->hmax = job->h[c] > job->hmax ? job->h[c] : job->hmax;
    job->vmax = job->v[c] > job->vmax ? job->v[c] : job->vmax;
  }
  for (int c = 0; c < job->ncomp; ++c) {
    IA_RETURN_IF(job->hmax % job->h[c] || job->vmax % job->v[c], IA_ERR_UNSUPPORTED);
    for (int by = 0; by < job->v[c]; ++by) {
      for (int bx = 0; bx < job->h[c]; ++bx) {
        IA_RETURN_IF(bpm >= IA_JPEG_MAX_BLOCKS_PER_MCU, IA_ERR_INVALID_ARGUMENT);
        job->block_comp[bpm] = (uint8_t)c;
        job->block_bx[bpm] = (uint8_t)bx;
        job->block_by[bpm] = (uint8_t)by;
        ++bpm;
      }
    }
  }
  job->blocks_per_mcu = bpm;
  job->mcux = (job->width + 8 * job->hmax - 1) / (8 * job->hmax);
  job->mcuy = (job->height + 8 * job->vmax - 1) / (8 * job->vmax);
  const long long mcus = (long long)job->mcux * job->mcuy;
  IA_RETURN_IF(mcus * bpm > 0x3fffffffLL, IA_ERR_UNSUPPORTED);
  job->total_blocks = (int32_t)(mcus * bpm);
  job->num_segments = job->restart_interval
                          ? (int32_t)((mcus + job->restart_interval - 1) / job->restart_interval)
                          : 1;
  job->file_bytes = (uint32_t)n;
  const bool swap = job->orientation >= 5;
  job->out_width = swap ? job->height : job->width;
  job->out_height = swap ? job->width : job->height;
  return IA_OK;
}

// ------------------------------------------------------------------- host: workspace carve
struct Carve {
  size_t clean, segment, coef, plane[IA_JPEG_MAX_COMPONENTS];
};

static inline size_t round256(size_t v) { return (v + 255) & ~(size_t)255; }

static Carve sizes(const ia_jpeg_job& j) {
  Carve c{};
  c.clean = round256((size_t)(j.file_bytes - j.scan_offset) + kCleanPad);
  c.segment = round256(((size_t)j.num_segments + 4) * sizeof(uint32_t));
  c.coef = round256((size_t)j.total_blocks * 64 * sizeof(int16_t));
  for (int k = 0; k < j.ncomp; ++k)
    c.plane[k] = round256((size_t)j.mcux * j.h[k] * 8 * (size_t)j.mcuy * j.v[k] * 8);
  return c;
}

// =============================================================================== J1 unstuff
__global__ void __launch_bounds__(kUnstuffThreads) unstuff_kernel(const ia_jpeg_job* __restrict__ jobs) {
  const ia_jpeg_job& J = jobs[blockIdx.x];
  const uint8_t* __restrict__ src = J.file_dev + J.scan_offset;
  const uint32_t n = J.file_bytes - J.scan_offset;
  uint8_t* __restrict__ clean = J.clean_dev;
  uint32_t* __restrict__ seg = J.segment_dev;
  const uint32_t seg_cap = (uint32_t)J.num_segments;

  __shared__ uint32_t s_end, s_base_keep, s_base_bnd;
  __shared__ uint32_t s_warp[kUnstuffThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_end = n, s_base_keep = 0, s_base_bnd = 0;
  __syncthreads();

  for (uint32_t tile = 0; tile < n; tile += kUnstuffThreads * kUnstuffPerThread) {
    const uint32_t r0 = tile + (uint32_t)tid * kUnstuffPerThread;
    uint8_t b[kUnstuffPerThread + 2];  // b[k] = byte r0 - 1 + k
#pragma unroll
    for (int k = 0; k < kUnstuffPerThread + 2; ++k) {
      const uint32_t idx = r0 + k - 1;  // wraps for the byte before the scan
      b[k] = (r0 + k == 0) ? 0 : (idx < n ? __ldg(src + idx) : 0xD9);
    }
    uint32_t keep_mask = 0, bnd_mask = 0, my_end = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < kUnstuffPerThread; ++k) {
      const uint32_t r = r0 + k;
      const uint8_t prev = b[k], cur = b[k + 1], next = b[k + 2];
      if (r < n) {
        if (cur == 0xFF) {
          if (next == 0x00) keep_mask |= 1u << k;
          else if (next != 0xFF && !(next >= 0xD0 && next <= 0xD7)) my_end = min(my_end, r);
        } else if (prev == 0xFF) {
          if (cur >= 0xD0 && cur <= 0xD7) bnd_mask |= 1u << k;  // cur == 0 is stuffing: dropped
        } else {
          keep_mask |= 1u << k;
        }
      }
    }
    if (my_end != 0xffffffffu) atomicMin(&s_end, my_end);
    __syncthreads();
    const uint32_t end = s_end;
    if (end < r0 + kUnstuffPerThread) {  // cut everything at and after the end marker
      const uint32_t valid = end > r0 ? end - r0 : 0;
      const uint32_t m = valid >= 32 ? 0xffffffffu : ((1u << valid) - 1u);
      keep_mask &= m, bnd_mask &= m;
    }
    // exclusive CTA scan of (boundaries << 16 | kept bytes)
    const uint32_t mine = (uint32_t)__popc(keep_mask) | ((uint32_t)__popc(bnd_mask) << 16);
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
    const uint32_t excl = incl - mine + (warp ? s_warp[warp - 1] : 0);
    uint32_t pos = s_base_keep + (excl & 0xffffu);
    uint32_t bpos = s_base_bnd + (excl >> 16);
    const uint32_t tile_total = s_warp[kUnstuffThreads / 32 - 1];
#pragma unroll
    for (int k = 0; k < kUnstuffPerThread; ++k) {
      if (keep_mask >> k & 1u) clean[pos++] = b[k + 1];
      if (bnd_mask >> k & 1u) {
        if (bpos < seg_cap) seg[2 + bpos] = pos;  // the next segment starts at this clean byte
        ++bpos;
      }
    }
    __syncthreads();
    if (tid == 0) s_base_keep += tile_total & 0xffffu, s_base_bnd += tile_total >> 16;
    __syncthreads();
    if (end < n) break;  // uniform: s_end is shared
  }
  const uint32_t total = s_base_keep;
  if (tid < kCleanPad) clean[total + tid] = 0;
  if (tid == 0) {
    seg[0] = total;
    seg[1] = min(s_base_bnd, seg_cap);
  }
}

// =============================================================================== J2 Huffman
struct HuffTables {
  uint16_t lut[8][1 << kLutBits];  // (length << 8) | symbol for codes of <= kLutBits bits, else 0
  int32_t maxcode[8][18];          // exclusive upper bound of the codes of each length, left-aligned to 16 bits
  int32_t valoff[8][18];           // index of the first value of that length minus its first code
  uint8_t vals[8][256];
  uint8_t blk_comp[16], blk_dc[16], blk_ac[16];  // per block of an MCU: component, table rows
};

struct Agg {  // what a subsequence contributes: completed blocks, DC sums, "predictors were reset"
  int nb, flag, d0, d1, d2;
};

__device__ __forceinline__ Agg combine(const Agg& a, const Agg& b) {
  Agg r;
  r.nb = a.nb + b.nb;
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

__device__ __forceinline__ uint32_t peek32(const uint32_t* __restrict__ words, uint32_t p) {
  const uint32_t i = p >> 5;
  const uint32_t hi = __byte_perm(__ldg(words + i), 0, 0x0123);
  const uint32_t lo = __byte_perm(__ldg(words + i + 1), 0, 0x0123);
  return __funnelshift_l(lo, hi, p & 31);
}

// symbol and code length for the code at the top of `bits`
__device__ __forceinline__ void huff_lookup(const HuffTables& T, int t, uint32_t bits, int& sym, int& len) {
  const uint32_t e = T.lut[t][bits >> (32 - kLutBits)];
  if (e) {
    sym = (int)(e & 255u), len = (int)(e >> 8);
    return;
  }
  const int32_t code = (int32_t)(bits >> 16);
  sym = 0, len = 16;  // an invalid code eats 16 bits and yields 0 (libjpeg warns and does the same)
  for (int l = kLutBits + 1; l <= 16; ++l) {
    if (code < T.maxcode[t][l]) {
      sym = T.vals[t][(T.valoff[t][l] + (code >> (16 - l))) & 255];
      len = l;
      return;
    }
  }
}

__device__ __forceinline__ int extend(uint32_t v, int s) {  // T.81 F.2.2.1 EXTEND
  return (int)v < (1 << (s - 1)) ? (int)v - (1 << s) + 1 : (int)v;
}

struct Stream {
  const uint32_t* words;   // clean stream
  uint32_t total_bits;
  const uint32_t* bounds;  // restart boundaries (clean byte offsets), ascending
  int num_bounds;
  const ia_jpeg_job* job;
  int bpm;
};

// Decode from state (p, b, z) while p < end_bit.  kWrite: store coefficients of block `blk`
// onwards with the DC predictors in `dc`; otherwise only count (dc accumulates the diffs).
template <bool kWrite>
__device__ __forceinline__ void decode_range(const Stream& S, const HuffTables& T, uint32_t& p, int& b, int& z,
                                             uint32_t end_bit, Agg& agg, int blk, int16_t* __restrict__ coef,
                                             int total_blocks) {
  // first boundary strictly after p; sitting exactly on one means "segment start"
  int lo = 0, hi = S.num_bounds;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(S.bounds + mid) * 8u > p) hi = mid;
    else lo = mid + 1;
  }
  int j = lo;
  int dc[3] = {agg.d0, agg.d1, agg.d2};
  if (j > 0 && __ldg(S.bounds + j - 1) * 8u == p) {
    b = 0, z = 0, agg.flag = 1;
    dc[0] = dc[1] = dc[2] = 0;
  }
  uint32_t next_bound = j < S.num_bounds ? __ldg(S.bounds + j) * 8u : 0xffffffffu;
  const uint32_t limit = min(end_bit, S.total_bits);
  while (p < limit) {
    const uint32_t bits = peek32(S.words, p);
    const int comp = T.blk_comp[b];
    int sym, len;
    huff_lookup(T, z == 0 ? T.blk_dc[b] : T.blk_ac[b], bits, sym, len);
    const int s = sym & 15;  // DC: the category itself (<= 11 for 8-bit data); AC: low nibble
    const uint32_t used = (uint32_t)(len + s);
    if (p + used > S.total_bits) {  // the padding at the very end: nothing more to decode
      p = S.total_bits;
      break;
    }
    if (p + used > next_bound) {  // ran into the padding before a restart marker: new segment
      p = next_bound, b = 0, z = 0, agg.flag = 1;
      dc[0] = dc[1] = dc[2] = 0;
      ++j;
      next_bound = j < S.num_bounds ? __ldg(S.bounds + j) * 8u : 0xffffffffu;
      continue;
    }
    const int value = s ? extend((bits << len) >> (32 - s), s) : 0;
    p += used;
    if (z == 0) {
      dc[comp] += value;
      if (kWrite && blk < total_blocks) coef[(size_t)blk * 64] = (int16_t)dc[comp];
      z = 1;
    } else {
      const int r = sym >> 4;
      if (s == 0) {
        z = r == 15 ? z + 16 : 64;
      } else {
        z += r;
        if (kWrite && z < 64 && blk < total_blocks) coef[(size_t)blk * 64 + c_natural[z]] = (int16_t)value;
        ++z;
      }
    }
    if (z >= 64) {
      z = 0;
      ++agg.nb;
      if (kWrite) ++blk;
      b = b + 1 == S.bpm ? 0 : b + 1;
    }
  }
  if (p >= S.total_bits) p = S.total_bits, b = 0, z = 0;  // one canonical state past the end
  agg.d0 = dc[0], agg.d1 = dc[1], agg.d2 = dc[2];
}

__global__ void __launch_bounds__(kHuffThreads) huffman_kernel(const ia_jpeg_job* __restrict__ jobs,
                                                               int32_t* __restrict__ status) {
  const ia_jpeg_job& J = jobs[blockIdx.x];
  extern __shared__ __align__(16) unsigned char huff_smem[];
  HuffTables& T = *reinterpret_cast<HuffTables*>(huff_smem);
  __shared__ uint32_t s_exit_p[kHuffThreads], s_exit_bz[kHuffThreads];
  __shared__ Agg s_warp_agg[kHuffThreads / 32];
  __shared__ Agg s_carry;
  __shared__ uint32_t s_carry_p, s_carry_bz;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- decoding tables: zero the look-ups, then one thread per table walks its codes
  for (int i = tid; i < 8 * (1 << kLutBits); i += kHuffThreads) (&T.lut[0][0])[i] = 0;
  for (int i = tid; i < 8 * 256; i += kHuffThreads) {
    const int t = i >> 8, k = i & 255;
    T.vals[t][k] = t < 4 ? (k < 16 ? J.dc_vals[t][k] : 0) : J.ac_vals[t - 4][k];
  }
  __syncthreads();
  if (tid < 8) {
    const uint8_t* counts = tid < 4 ? J.dc_counts[tid] : J.ac_counts[tid - 4];
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
      const int cnt = counts[l - 1];
      T.valoff[tid][l] = k - code;
      for (int c = 0; c < cnt; ++c, ++k, ++code) {
        if (l <= kLutBits) {
          const uint16_t e = (uint16_t)((l << 8) | T.vals[tid][k]);
          const int first = code << (kLutBits - l);
          for (int f = 0; f < (1 << (kLutBits - l)); ++f) T.lut[tid][(first + f) & ((1 << kLutBits) - 1)] = e;
        }
      }
      T.maxcode[tid][l] = cnt ? (code << (16 - l)) : 0;  // codes of this length are < maxcode (left-aligned)
      code <<= 1;
    }
    // lengths without codes must never match: give them the bound of the previous length
    for (int l = 1; l <= 16; ++l)
      if (!counts[l - 1]) T.maxcode[tid][l] = l > 1 ? T.maxcode[tid][l - 1] : 0;
  }
  if (tid < 16) {
    const int k = tid < J.blocks_per_mcu ? tid : 0;
    const int c = J.block_comp[k];
    T.blk_comp[tid] = (uint8_t)c;
    T.blk_dc[tid] = J.comp_dc[c];
    T.blk_ac[tid] = (uint8_t)(4 + J.comp_ac[c]);
  }
  if (tid == 0) {
    s_carry.nb = 0, s_carry.flag = 0, s_carry.d0 = s_carry.d1 = s_carry.d2 = 0;
    s_carry_p = 0, s_carry_bz = 0;
  }
  __syncthreads();

  Stream S;
  S.words = reinterpret_cast<const uint32_t*>(J.clean_dev);
  S.total_bits = J.segment_dev[0] * 8u;
  S.bounds = J.segment_dev + 2;
  S.num_bounds = (int)J.segment_dev[1];
  S.job = &J;
  S.bpm = J.blocks_per_mcu;

  for (uint32_t chunk_bit = 0; chunk_bit < S.total_bits; chunk_bit += kHuffThreads * kSubseqBits) {
    const uint32_t begin = chunk_bit + (uint32_t)tid * kSubseqBits, end = begin + kSubseqBits;
    // entry state: exact for the first thread (carried over), a guess for the others
    uint32_t ep = tid == 0 ? s_carry_p : min(begin, S.total_bits);
    uint32_t ebz = tid == 0 ? s_carry_bz : 0;
    Agg agg;
    uint32_t p;
    int b, z;
    auto run = [&]() {
      p = ep, b = (int)(ebz >> 8), z = (int)(ebz & 255u);
      agg.nb = 0, agg.flag = 0, agg.d0 = agg.d1 = agg.d2 = 0;
      decode_range<false>(S, T, p, b, z, end, agg, 0, nullptr, 0);
      s_exit_p[tid] = p;
      s_exit_bz[tid] = ((uint32_t)b << 8) | (uint32_t)z;
    };
    run();
    for (;;) {
      __syncthreads();
      bool redo = false;
      uint32_t np = 0, nbz = 0;
      if (tid > 0) {
        np = s_exit_p[tid - 1], nbz = s_exit_bz[tid - 1];
        redo = np != ep || nbz != ebz;
      }
      const int any = __syncthreads_or(redo);  // also orders the reads above before the writes below
      if (!any) break;
      if (redo) {
        ep = np, ebz = nbz;
        run();
      }
    }

    // ---- CTA scan of the contributions: output block and DC predictors at each thread's entry
    Agg incl = agg;
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
    Agg before = s_carry;  // everything before this thread: carry (+) earlier warps (+) earlier lanes
    if (warp > 0) before = combine(before, s_warp_agg[warp - 1]);
    {
      const Agg prev_lane = shfl_up_agg(incl, 1);
      if (lane > 0) before = combine(before, prev_lane);
    }
    const Agg chunk_total = s_warp_agg[kHuffThreads / 32 - 1];
    const uint32_t last_p = s_exit_p[kHuffThreads - 1], last_bz = s_exit_bz[kHuffThreads - 1];

    // ---- write pass from the now exact entry state
    {
      Agg w;
      w.nb = 0, w.flag = 0, w.d0 = before.d0, w.d1 = before.d1, w.d2 = before.d2;
      p = ep, b = (int)(ebz >> 8), z = (int)(ebz & 255u);
      decode_range<true>(S, T, p, b, z, end, w, before.nb, J.coef_dev, J.total_blocks);
    }
    __syncthreads();
    if (tid == 0) {
      s_carry = combine(s_carry, chunk_total);
      s_carry.flag = 0;
      s_carry_p = last_p, s_carry_bz = last_bz;
    }
    __syncthreads();
  }
  if (tid == 0) status[blockIdx.x] = s_carry.nb == J.total_blocks ? 0 : 1;
}

// ================================================================================== J3 IDCT
constexpr int kIdctThreads = 128;

#define IA_FIX_0_298631336 2446
#define IA_FIX_0_390180644 3196
#define IA_FIX_0_541196100 4433
#define IA_FIX_0_765366865 6270
#define IA_FIX_0_899976223 7373
#define IA_FIX_1_175875602 9633
#define IA_FIX_1_501321110 12299
#define IA_FIX_1_847759065 15137
#define IA_FIX_1_961570560 16069
#define IA_FIX_2_053119869 16819
#define IA_FIX_2_562915447 20995
#define IA_FIX_3_072711026 25172

// one 8-point pass of jidctint.c; outputs are not yet descaled
__device__ __forceinline__ void idct8(const int d[8], int o[8]) {
  int z1 = (d[2] + d[6]) * IA_FIX_0_541196100;
  const int tmp2 = z1 + d[6] * (-IA_FIX_1_847759065);
  const int tmp3 = z1 + d[2] * IA_FIX_0_765366865;
  const int tmp0 = (d[0] + d[4]) << 13, tmp1 = (d[0] - d[4]) << 13;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  int t0 = d[7], t1 = d[5], t2 = d[3], t3 = d[1];
  z1 = t0 + t3;
  int z2 = t1 + t2, z3 = t0 + t2, z4 = t1 + t3;
  const int z5 = (z3 + z4) * IA_FIX_1_175875602;
  t0 *= IA_FIX_0_298631336, t1 *= IA_FIX_2_053119869, t2 *= IA_FIX_3_072711026, t3 *= IA_FIX_1_501321110;
  z1 *= -IA_FIX_0_899976223, z2 *= -IA_FIX_2_562915447, z3 *= -IA_FIX_1_961570560, z4 *= -IA_FIX_0_390180644;
  z3 += z5, z4 += z5;
  t0 += z1 + z3, t1 += z2 + z4, t2 += z2 + z3, t3 += z1 + z4;
  o[0] = tmp10 + t3, o[7] = tmp10 - t3, o[1] = tmp11 + t2, o[6] = tmp11 - t2;
  o[2] = tmp12 + t1, o[5] = tmp12 - t1, o[3] = tmp13 + t0, o[4] = tmp13 - t0;
}

// libjpeg's range_limit[x & RANGE_MASK] with the table centred on 128
__device__ __forceinline__ uint32_t range_limit(int x) {
  const int i = x & 1023;
  return i < 128 ? i + 128 : (i < 512 ? 255 : (i < 896 ? 0 : i - 896));
}

__global__ void __launch_bounds__(kIdctThreads) idct_kernel(const ia_jpeg_job* __restrict__ jobs) {
  const ia_jpeg_job& J = jobs[blockIdx.y];
  __shared__ uint16_t s_quant[IA_JPEG_MAX_COMPONENTS][64];
  for (int i = threadIdx.x; i < IA_JPEG_MAX_COMPONENTS * 64; i += kIdctThreads) (&s_quant[0][0])[i] = (&J.quant[0][0])[i];
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
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const uint4 v = __ldg(src + r);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      ws[r * 8 + 2 * i] = (int)(int16_t)(w[i] & 0xffffu) * (int)q[r * 8 + 2 * i];
      ws[r * 8 + 2 * i + 1] = (int)(int16_t)(w[i] >> 16) * (int)q[r * 8 + 2 * i + 1];
    }
  }
  // pass 1: columns, descale by CONST_BITS - PASS1_BITS = 11
#pragma unroll
  for (int col = 0; col < 8; ++col) {
    int d[8], o[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) d[r] = ws[r * 8 + col];
    idct8(d, o);
#pragma unroll
    for (int r = 0; r < 8; ++r) ws[r * 8 + col] = (o[r] + (1 << 10)) >> 11;
  }
  // pass 2: rows, descale by CONST_BITS + PASS1_BITS + 3 = 18, range-limit, store
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    int d[8], o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) d[i] = ws[r * 8 + i];
    idct8(d, o);
    uint32_t px[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) px[i] = range_limit((o[i] + (1 << 17)) >> 18);
    uint2 out;
    out.x = px[0] | (px[1] << 8) | (px[2] << 16) | (px[3] << 24);
    out.y = px[4] | (px[5] << 8) | (px[6] << 16) | (px[7] << 24);
    *reinterpret_cast<uint2*>(dst + (size_t)r * pitch) = out;
  }
}

// ================================================================= J4 upsample + colour
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

__global__ void __launch_bounds__(256) colour_kernel(const ia_jpeg_job* __restrict__ jobs) {
  const ia_jpeg_job& J = jobs[blockIdx.z];
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  const int W = J.width, H = J.height;
  if (x >= W || y >= H) return;
  const int pitch_y = J.mcux * J.h[0] * 8;
  const int yy = J.plane_dev[0][(size_t)y * pitch_y + x];
  int r, g, b;
  if (J.ncomp == 1) {
    r = g = b = yy;
  } else {
    int cc[2];
#pragma unroll
    for (int k = 1; k <= 2; ++k) {
      const int hexp = J.hmax / J.h[k], vexp = J.vmax / J.v[k];
      const int dw = (W * J.h[k] + J.hmax - 1) / J.hmax, dh = (H * J.v[k] + J.vmax - 1) / J.vmax;
      cc[k - 1] = chroma_sample(J.plane_dev[k], J.mcux * J.h[k] * 8, dw, dh, hexp, vexp, x, y);
    }
    const int cb = cc[0] - 128, cr = cc[1] - 128;
    // jdcolor.c build_ycc_rgb_table: FIX(1.40200)=91881, FIX(1.77200)=116130,
    // FIX(0.71414)=46802, FIX(0.34414)=22554, ONE_HALF=32768, arithmetic shifts
    r = yy + ((91881 * cr + 32768) >> 16);
    g = yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
    b = yy + ((116130 * cb + 32768) >> 16);
    r = r < 0 ? 0 : (r > 255 ? 255 : r);
    g = g < 0 ? 0 : (g > 255 ? 255 : g);
    b = b < 0 ? 0 : (b > 255 ? 255 : b);
  }
  int ox = x, oy = y;  // OpenCV's ExifTransform
  switch (J.orientation) {
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
  o[0] = (uint8_t)r, o[1] = (uint8_t)g, o[2] = (uint8_t)b;
}

}  // namespace jpeg
}  // namespace ia

// ===================================================================================== C ABI
extern "C" size_t ia_jpeg_job_size(void) { return sizeof(ia_jpeg_job); }

extern "C" int ia_jpeg_parse(const uint8_t* file_host, size_t file_bytes, ia_jpeg_job* job) {
  IA_RETURN_IF(!file_host || !job, IA_ERR_INVALID_ARGUMENT);
  return ia::jpeg::parse(file_host, file_bytes, job);
}

extern "C" size_t ia_jpeg_workspace_bytes(const ia_jpeg_job* jobs_host, int n) {
  if (!jobs_host || n <= 0) return 0;
  size_t total = 0;
  for (int i = 0; i < n; ++i) {
    const ia::jpeg::Carve c = ia::jpeg::sizes(jobs_host[i]);
    total += c.clean + c.segment + c.coef + c.plane[0] + c.plane[1] + c.plane[2];
  }
  return total + 256;
}

extern "C" int ia_jpeg_plan(ia_jpeg_job* jobs_host, int n, const uint8_t* const* file_dev,
                            uint8_t* const* out_dev, void* workspace_dev, size_t workspace_bytes) {
  using namespace ia;
  IA_RETURN_IF(n < 0, IA_ERR_INVALID_ARGUMENT);
  if (n == 0) return IA_OK;
  IA_RETURN_IF(!jobs_host || !file_dev || !out_dev || !workspace_dev, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(((uintptr_t)workspace_dev & 255) != 0, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(workspace_bytes < ia_jpeg_workspace_bytes(jobs_host, n), IA_ERR_WORKSPACE_TOO_SMALL);
  // coefficients of all images first and contiguous: one memset clears them
  uint8_t* base = static_cast<uint8_t*>(workspace_dev);
  size_t off = 0;
  for (int i = 0; i < n; ++i) {
    jobs_host[i].coef_dev = reinterpret_cast<int16_t*>(base + off);
    off += jpeg::sizes(jobs_host[i]).coef;
  }
  for (int i = 0; i < n; ++i) {
    IA_RETURN_IF(!file_dev[i] || !out_dev[i], IA_ERR_INVALID_ARGUMENT);
    const jpeg::Carve c = jpeg::sizes(jobs_host[i]);
    ia_jpeg_job& j = jobs_host[i];
    j.file_dev = file_dev[i];
    j.out_dev = out_dev[i];
    j.clean_dev = base + off, off += c.clean;
    j.segment_dev = reinterpret_cast<uint32_t*>(base + off), off += c.segment;
    for (int k = 0; k < IA_JPEG_MAX_COMPONENTS; ++k) {
      j.plane_dev[k] = k < j.ncomp ? base + off : nullptr;
      off += c.plane[k];
    }
  }
  return IA_OK;
}

extern "C" int ia_jpeg_decode(const ia_jpeg_job* jobs_host, const ia_jpeg_job* jobs_dev, int n,
                              int32_t* status_dev, void* stream_) {
  using namespace ia;
  using namespace ia::jpeg;
  cudaStream_t stream = (cudaStream_t)stream_;
  IA_RETURN_IF(n < 0, IA_ERR_INVALID_ARGUMENT);
  if (n == 0) return IA_OK;
  IA_RETURN_IF(!jobs_host || !jobs_dev || !status_dev, IA_ERR_INVALID_ARGUMENT);
  size_t coef_bytes = 0;
  int max_blocks = 0, max_w = 0, max_h = 0;
  for (int i = 0; i < n; ++i) {
    const ia_jpeg_job& j = jobs_host[i];
    IA_RETURN_IF(!j.coef_dev || !j.clean_dev || !j.out_dev || !j.file_dev, IA_ERR_INVALID_ARGUMENT);
    coef_bytes += sizes(j).coef;
    max_blocks = j.total_blocks > max_blocks ? j.total_blocks : max_blocks;
    max_w = j.width > max_w ? j.width : max_w;
    max_h = j.height > max_h ? j.height : max_h;
  }
  IA_RETURN_IF(n > 65535, IA_ERR_UNSUPPORTED);
  cudaError_t e = cudaMemsetAsync(jobs_host[0].coef_dev, 0, coef_bytes, stream);
  if (e != cudaSuccess) return (int)e;
  unstuff_kernel<<<n, kUnstuffThreads, 0, stream>>>(jobs_dev);
  int rc = launch_status();
  if (rc != IA_OK) return rc;
  static_assert(sizeof(HuffTables) + 8 * 1024 <= 48 * 1024, "huffman_kernel shared memory");
  huffman_kernel<<<n, kHuffThreads, sizeof(HuffTables), stream>>>(jobs_dev, status_dev);
  rc = launch_status();
  if (rc != IA_OK) return rc;
  idct_kernel<<<dim3((max_blocks + kIdctThreads - 1) / kIdctThreads, n), kIdctThreads, 0, stream>>>(jobs_dev);
  rc = launch_status();
  if (rc != IA_OK) return rc;
  colour_kernel<<<dim3((max_w + 31) / 32, (max_h + 7) / 8, n), 256, 0, stream>>>(jobs_dev);
  return launch_status();
}
