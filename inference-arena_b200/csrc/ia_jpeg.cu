// ia_jpeg.cu — SURVEY §8f row N3: baseline JPEG decode on the device, bit-identical to what
// the reference's load_image_from_bytes gets from OpenCV/libjpeg-turbo
// (src/shared/processing/transforms.py:77-110).  oracle/jpeg.py restates the algorithm.
//
// Host side here: the header parser (markers, tables, geometry, EXIF orientation), the
// workspace carve and the launches.  Kernels: ia_jpeg_kernels.cuh —
//   unstuff (count / scan / scatter): the entropy-coded bytes without FF00 stuffing and RSTn
//       markers, plus where each restart segment begins and where the scan ends;
//   resync walk: for streams with damaged restart markers, libjpeg's resynchronisation as a
//       table (which stretch of data belongs to which restart segment);
//   huffman: one thread per 128- or 256-byte subsequence.  Huffman codes cannot be cut at arbitrary
//       bits, so every thread first decodes from a guessed state (bit 0 of its subsequence,
//       start of a block); a thread whose predecessor ended in a different state (bit
//       position, block within the MCU, zig-zag index) decodes again from that state, until
//       nothing changes — JPEG codes resynchronise within a few symbols, so this takes a
//       handful of rounds instead of one serial walk (self-synchronising decode, Weissenberger
//       & Schmidt 2018/2021).  A CTA scan of block counts and DC sums gives every thread its
//       output block and DC predictors; one more pass writes the coefficients (and conceals
//       damaged segments the way jdhuff.c does).  The chunks of one image (32 to 128 KB) run in
//       different CTAs, chained by a ticket-ordered look-back;
//   idct: dequantise + libjpeg's jidctint.c "islow" integer IDCT, one thread per block;
//   colour: libjpeg's fancy (triangle) chroma upsampling, jdcolor.c fixed-point YCbCr->RGB,
//       EXIF orientation, RGB store.
#include <stdlib.h>
#include <string.h>

#include "ia_jpeg_kernels.cuh"

namespace ia {
namespace jpeg {

static const uint8_t h_natural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                      12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                      35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                      58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// =========================================================================== host: headers
// ITU T.81 Annex K.3-K.6: the tables libjpeg installs for ids 0 (luminance) and 1 (chrominance)
// when a file defines none.  DC symbols are 0..11 in order for both.
static const uint8_t kStdDcCounts[2][16] = {{0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0},
                                            {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0}};
static const uint8_t kStdAcCounts[2][16] = {{0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 125},
                                            {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 119}};
static const uint8_t kStdAcVals[2][162] = {
    {
     0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07,
     0x22, 0x71, 0x14, 0x32, 0x81, 0x91, 0xA1, 0x08, 0x23, 0x42, 0xB1, 0xC1, 0x15, 0x52, 0xD1, 0xF0,
     0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0A, 0x16, 0x17, 0x18, 0x19, 0x1A, 0x25, 0x26, 0x27, 0x28,
     0x29, 0x2A, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3A, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49,
     0x4A, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5A, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69,
     0x6A, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7A, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89,
     0x8A, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9A, 0xA2, 0xA3, 0xA4, 0xA5, 0xA6, 0xA7,
     0xA8, 0xA9, 0xAA, 0xB2, 0xB3, 0xB4, 0xB5, 0xB6, 0xB7, 0xB8, 0xB9, 0xBA, 0xC2, 0xC3, 0xC4, 0xC5,
     0xC6, 0xC7, 0xC8, 0xC9, 0xCA, 0xD2, 0xD3, 0xD4, 0xD5, 0xD6, 0xD7, 0xD8, 0xD9, 0xDA, 0xE1, 0xE2,
     0xE3, 0xE4, 0xE5, 0xE6, 0xE7, 0xE8, 0xE9, 0xEA, 0xF1, 0xF2, 0xF3, 0xF4, 0xF5, 0xF6, 0xF7, 0xF8,
     0xF9, 0xFA,
    },
    {
     0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71,
     0x13, 0x22, 0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xA1, 0xB1, 0xC1, 0x09, 0x23, 0x33, 0x52, 0xF0,
     0x15, 0x62, 0x72, 0xD1, 0x0A, 0x16, 0x24, 0x34, 0xE1, 0x25, 0xF1, 0x17, 0x18, 0x19, 0x1A, 0x26,
     0x27, 0x28, 0x29, 0x2A, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3A, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48,
     0x49, 0x4A, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5A, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68,
     0x69, 0x6A, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7A, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87,
     0x88, 0x89, 0x8A, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9A, 0xA2, 0xA3, 0xA4, 0xA5,
     0xA6, 0xA7, 0xA8, 0xA9, 0xAA, 0xB2, 0xB3, 0xB4, 0xB5, 0xB6, 0xB7, 0xB8, 0xB9, 0xBA, 0xC2, 0xC3,
     0xC4, 0xC5, 0xC6, 0xC7, 0xC8, 0xC9, 0xCA, 0xD2, 0xD3, 0xD4, 0xD5, 0xD6, 0xD7, 0xD8, 0xD9, 0xDA,
     0xE2, 0xE3, 0xE4, 0xE5, 0xE6, 0xE7, 0xE8, 0xE9, 0xEA, 0xF2, 0xF3, 0xF4, 0xF5, 0xF6, 0xF7, 0xF8,
     0xF9, 0xFA,
    }};

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

// libjpeg's fatal checks on a table a scan uses (jdhuff.c jpeg_make_d_derived_tbl): the codes of
// a length may not reach the all-ones code of that length; DC symbols are categories 0..15.
static bool huffman_table_ok(const uint8_t* counts, const uint8_t* vals, bool dc) {
  int code = 0, total = 0;
  for (int l = 1; l <= 16; ++l) {
    code += counts[l - 1];
    total += counts[l - 1];
    if (code >= (1 << l)) return false;
    code <<= 1;
  }
  if (dc) {
    for (int k = 0; k < total; ++k)
      if (vals[k] > 15) return false;
  }
  return true;
}

static int parse(const uint8_t* d, size_t n, ia_jpeg_job* job) {
  memset(job, 0, sizeof(*job));
  // What makes libjpeg give up (ERREXIT: cv2.imdecode returns None and the reference raises) is
  // IA_ERR_INVALID_ARGUMENT here; jdmarker.c read_markers / get_sof / get_sos / get_dht / get_dqt /
  // get_dri, jdinput.c initial_setup / per_scan_setup / latch_quant_tables.  OpenCV picks its
  // decoder by the three-byte signature FF D8 FF.
  IA_RETURN_IF(n < 4 || d[0] != 0xFF || d[1] != 0xD8 || d[2] != 0xFF, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(n >= 0xfffffff0ull, IA_ERR_UNSUPPORTED);
  uint16_t qt[4][64];
  bool have_qt[4] = {false, false, false, false}, have_dc[4] = {false, false, false, false},
       have_ac[4] = {false, false, false, false}, long_dc[4] = {false, false, false, false};
  int tq[IA_JPEG_MAX_COMPONENTS] = {0, 0, 0}, comp_id[IA_JPEG_MAX_COMPONENTS] = {0, 0, 0};
  bool have_sof = false, saw_exif = false, saw_jfif = false;
  int adobe_transform = -1;
  job->orientation = 1;
  size_t i = 2;
  for (;;) {
    // next_marker: skip to an FF, skip fill FFs; FF00 is stuffing, not a marker
    for (;;) {
      while (i < n && d[i] != 0xFF) ++i;
      while (i < n && d[i] == 0xFF) ++i;
      IA_RETURN_IF(i >= n, IA_ERR_INVALID_ARGUMENT);
      if (d[i] != 0) break;
      ++i;
    }
    const uint8_t m = d[i++];
    if ((m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;  // parameterless
    IA_RETURN_IF(m == 0xD8 || m == 0xD9, IA_ERR_INVALID_ARGUMENT);  // second SOI, EOI before the scan
    const bool known = (m >= 0xC0 && m <= 0xCF) || m == 0xDA || m == 0xDB || m == 0xDC || m == 0xDD ||
                       m == 0xFE || (m >= 0xE0 && m <= 0xEF);
    IA_RETURN_IF(!known, IA_ERR_INVALID_ARGUMENT);
    IA_RETURN_IF(i + 2 > n, IA_ERR_INVALID_ARGUMENT);
    size_t L = be16(d + i);
    // skip_variable / get_interesting_appn skip nothing for a length below 2
    if (L < 2 && ((m >= 0xE0 && m <= 0xEF) || m == 0xFE || m == 0xDC)) L = 2;
    IA_RETURN_IF(L < 2 || i + L > n, IA_ERR_INVALID_ARGUMENT);
    const uint8_t* seg = d + i + 2;
    const size_t len = L - 2;
    if (m == 0xCC) {  // DAC: pointless in a Huffman file, but get_dac still checks it
      IA_RETURN_IF(len % 2, IA_ERR_INVALID_ARGUMENT);
      for (size_t k = 0; k < len; k += 2) {
        const int index = seg[k], val = seg[k + 1];
        IA_RETURN_IF(index >= 32 || (index < 16 && (val & 15) > (val >> 4)), IA_ERR_INVALID_ARGUMENT);
      }
    }
    if (m == 0xDB) {
      size_t k = 0;
      while (k < len) {
        const int pq = seg[k] >> 4, t = seg[k] & 15;
        ++k;
        // any non-zero precision nibble means 16-bit entries (get_dqt)
        IA_RETURN_IF(t > 3 || k + (pq ? 128 : 64) > len, IA_ERR_INVALID_ARGUMENT);
        for (int j = 0; j < 64; ++j) {
          qt[t][h_natural[j]] = pq ? (uint16_t)be16(seg + k + 2 * j) : seg[k + j];
        }
        k += pq ? 128 : 64;
        have_qt[t] = true;
      }
    } else if (m == 0xC0 || m == 0xC1) {
      IA_RETURN_IF(have_sof || len < 6, IA_ERR_INVALID_ARGUMENT);
      job->height = (int32_t)be16(seg + 1);
      job->width = (int32_t)be16(seg + 3);
      job->ncomp = seg[5];
      IA_RETURN_IF(job->height == 0 || job->width == 0 || job->ncomp == 0, IA_ERR_INVALID_ARGUMENT);
      IA_RETURN_IF(len != 6 + 3 * (size_t)job->ncomp, IA_ERR_INVALID_ARGUMENT);
      IA_RETURN_IF(seg[0] != 8, IA_ERR_UNSUPPORTED);
      IA_RETURN_IF(job->ncomp != 1 && job->ncomp != 3, IA_ERR_UNSUPPORTED);
      for (int c = 0; c < job->ncomp; ++c) {
        comp_id[c] = seg[6 + 3 * c];
        job->h[c] = seg[7 + 3 * c] >> 4;
        job->v[c] = seg[7 + 3 * c] & 15;
        tq[c] = seg[8 + 3 * c];
      }
      have_sof = true;
    } else if (m >= 0xC0 && m <= 0xCF && m != 0xC4 && m != 0xCC) {
      return IA_ERR_UNSUPPORTED;  // progressive, lossless, arithmetic
    } else if (m == 0xC4) {
      size_t k = 0;
      while (k < len) {
        IA_RETURN_IF(k + 17 > len, IA_ERR_INVALID_ARGUMENT);
        const int tc = seg[k] >> 4, th = seg[k] & 15;
        IA_RETURN_IF(tc > 1 || th > 3, IA_ERR_INVALID_ARGUMENT);
        int total = 0;
        for (int j = 0; j < 16; ++j) total += seg[k + 1 + j];
        IA_RETURN_IF(total > 256 || k + 17 + total > len, IA_ERR_INVALID_ARGUMENT);
        uint8_t* counts = tc ? job->ac_counts[th] : job->dc_counts[th];
        uint8_t* vals = tc ? job->ac_vals[th] : job->dc_vals[th];
        memcpy(counts, seg + k + 1, 16);
        memset(vals, 0, tc ? 256 : 16);
        // a DC table of more than 16 symbols repeats one or holds a category above 15: refused when used
        if (!tc) long_dc[th] = total > 16;
        memcpy(vals, seg + k + 17, !tc && total > 16 ? 16 : total);
        (tc ? have_ac : have_dc)[th] = true;
        k += 17 + total;
      }
    } else if (m == 0xDD) {
      IA_RETURN_IF(len != 2, IA_ERR_INVALID_ARGUMENT);
      job->restart_interval = (int32_t)be16(seg);
    } else if (m == 0xE1 && !saw_exif && len >= 6 && memcmp(seg, "Exif\0\0", 6) == 0) {
      saw_exif = true;
      job->orientation = exif_orientation(seg, len);
    } else if (m == 0xEE && len >= 12 && memcmp(seg, "Adobe", 5) == 0) {
      adobe_transform = seg[11];
    } else if (m == 0xE0 && len >= 14 && memcmp(seg, "JFIF\0", 5) == 0) {
      saw_jfif = true;
    } else if (m == 0xDA) {
      IA_RETURN_IF(!have_sof, IA_ERR_INVALID_ARGUMENT);
      const int ns = len ? seg[0] : 0;
      IA_RETURN_IF(ns < 1 || ns > 4 || len != 2 * (size_t)ns + 4, IA_ERR_INVALID_ARGUMENT);
      for (int s = 0; s < ns; ++s) {
        bool found = false;
        for (int c = 0; c < job->ncomp; ++c) found |= seg[1 + 2 * s] == comp_id[c];
        IA_RETURN_IF(!found, IA_ERR_INVALID_ARGUMENT);
      }
      IA_RETURN_IF(ns != job->ncomp, IA_ERR_UNSUPPORTED);  // non-interleaved / multi-scan
      for (int s = 0; s < ns; ++s) {
        IA_RETURN_IF(seg[1 + 2 * s] != comp_id[s], IA_ERR_UNSUPPORTED);
        job->comp_dc[s] = seg[2 + 2 * s] >> 4;
        job->comp_ac[s] = seg[2 + 2 * s] & 15;
      }
      job->scan_offset = (uint32_t)(i + L);
      break;
    }
    i += L;
  }
  IA_RETURN_IF(job->width > 65500 || job->height > 65500, IA_ERR_INVALID_ARGUMENT);
  for (int c = 0; c < job->ncomp; ++c)
    IA_RETURN_IF(job->h[c] < 1 || job->h[c] > 4 || job->v[c] < 1 || job->v[c] > 4, IA_ERR_INVALID_ARGUMENT);
  // libjpeg's guess of the colour space (jdapimin.c default_decompress_parms): JFIF means YCbCr;
  // else an Adobe marker decides; else the component ids 'R','G','B' mean RGB.  RGB is not decoded here.
  if (job->ncomp == 3 && !saw_jfif) {
    const bool ids_rgb = comp_id[0] == 'R' && comp_id[1] == 'G' && comp_id[2] == 'B';
    IA_RETURN_IF(adobe_transform == 0 || (adobe_transform < 0 && ids_rgb), IA_ERR_UNSUPPORTED);
  }
  if (job->ncomp == 1) job->h[0] = job->v[0] = 1;  // sampling factors mean nothing for one component
  job->hmax = job->vmax = 1;
  int bpm = 0;
  // Table ids 0 and 1 fall back to the standard tables of T.81 Annex K when the file defines
  // none (jdhuff.c jinit_huff_decoder -> std_huff_tables: Motion-JPEG frames leave them out).
  for (int t = 0; t < 2; ++t) {
    if (!have_dc[t]) {
      memcpy(job->dc_counts[t], kStdDcCounts[t], 16);
      memset(job->dc_vals[t], 0, 16);
      for (int k = 0; k < 12; ++k) job->dc_vals[t][k] = (uint8_t)k;
      have_dc[t] = true;
    }
    if (!have_ac[t]) {
      memcpy(job->ac_counts[t], kStdAcCounts[t], 16);
      memset(job->ac_vals[t], 0, 256);
      memcpy(job->ac_vals[t], kStdAcVals[t], 162);
      have_ac[t] = true;
    }
  }
  for (int c = 0; c < job->ncomp; ++c) {
    const int td = job->comp_dc[c], ta = job->comp_ac[c];
    IA_RETURN_IF(tq[c] > 3 || !have_qt[tq[c]], IA_ERR_INVALID_ARGUMENT);
    IA_RETURN_IF(td > 3 || ta > 3 || !have_dc[td] || !have_ac[ta] || long_dc[td], IA_ERR_INVALID_ARGUMENT);
    IA_RETURN_IF(!huffman_table_ok(job->dc_counts[td], job->dc_vals[td], true) ||
                     !huffman_table_ok(job->ac_counts[ta], job->ac_vals[ta], false),
                 IA_ERR_INVALID_ARGUMENT);
    memcpy(job->quant[c], qt[tq[c]], sizeof(qt[0]));
    job->hmax = job->h[c] > job->hmax ? job->h[c] : job->hmax;
    job->vmax = job->v[c] > job->vmax ? job->v[c] : job->vmax;
  }
  // the colour stage reads the first component at full resolution (every encoder's layout)
  IA_RETURN_IF(job->h[0] != job->hmax || job->v[0] != job->vmax, IA_ERR_UNSUPPORTED);
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
  for (int c = 0; c < job->ncomp; ++c) {
    job->hexp[c] = job->hmax / job->h[c];
    job->vexp[c] = job->vmax / job->v[c];
    job->comp_w[c] = (job->width * job->h[c] + job->hmax - 1) / job->hmax;
    job->comp_h[c] = (job->height * job->v[c] + job->vmax - 1) / job->vmax;
    job->comp_pitch[c] = job->mcux * job->h[c] * 8;
  }
  const long long mcus = (long long)job->mcux * job->mcuy;
  IA_RETURN_IF(mcus * bpm > 0x3fffffffLL, IA_ERR_UNSUPPORTED);
  job->total_blocks = (int32_t)(mcus * bpm);
  job->num_segments = job->restart_interval
                          ? (int32_t)((mcus + job->restart_interval - 1) / job->restart_interval)
                          : 1;
  // bit positions inside the entropy-coded data are 32-bit on the device: 2^29 bytes = 2^32 bits
  IA_RETURN_IF(n - job->scan_offset >= (1ull << 29), IA_ERR_UNSUPPORTED);
  job->file_bytes = (uint32_t)n;
  const bool swap = job->orientation >= 5;
  job->out_width = swap ? job->height : job->width;
  job->out_height = swap ? job->width : job->height;
  return IA_OK;
}

// ------------------------------------------------------------------- host: workspace carve
struct Carve {
  size_t clean, segment, tile, coef, chain, plane[IA_JPEG_MAX_COMPONENTS];
};

static inline size_t round256(size_t v) { return (v + 255) & ~(size_t)255; }

static Carve sizes(const ia_jpeg_job& j) {
  Carve c{};
  c.clean = round256((size_t)scan_bytes(j) + kCleanPad);
  c.segment = round256((3 + 2 * (size_t)region_capacity(j)) * sizeof(uint32_t));  // header, offsets, marker codes
  c.tile = round256((size_t)tile_capacity(j) * kTileWords * sizeof(uint32_t));
  c.coef = round256((size_t)j.total_blocks * 64 * sizeof(int16_t));
  c.chain = round256((size_t)chunk_capacity(j) * kChainWords * sizeof(uint32_t));
  for (int k = 0; k < j.ncomp; ++k)
    c.plane[k] = round256((size_t)j.mcux * j.h[k] * 8 * (size_t)j.mcuy * j.v[k] * 8);
  return c;
}

// bytes at the front of the workspace that every call clears: coefficients, chain records, ticket
static size_t zeroed_bytes(const ia_jpeg_job* jobs, int n) {
  size_t total = 256;  // the ticket counter
  for (int i = 0; i < n; ++i) {
    const Carve c = sizes(jobs[i]);
    total += c.coef + c.chain;
  }
  return total;
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
  size_t total = ia::jpeg::zeroed_bytes(jobs_host, n);
  for (int i = 0; i < n; ++i) {
    const ia::jpeg::Carve c = ia::jpeg::sizes(jobs_host[i]);
    total += c.clean + c.segment + c.tile + c.plane[0] + c.plane[1] + c.plane[2];
  }
  return total;
}

extern "C" int ia_jpeg_plan(ia_jpeg_job* jobs_host, int n, const uint8_t* const* file_dev,
                            uint8_t* const* out_dev, void* workspace_dev, size_t workspace_bytes) {
  using namespace ia;
  IA_RETURN_IF(n < 0, IA_ERR_INVALID_ARGUMENT);
  if (n == 0) return IA_OK;
  IA_RETURN_IF(!jobs_host || !file_dev || !out_dev || !workspace_dev, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(((uintptr_t)workspace_dev & 255) != 0, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(workspace_bytes < ia_jpeg_workspace_bytes(jobs_host, n), IA_ERR_WORKSPACE_TOO_SMALL);
  // everything a call must clear sits at the front, contiguous: one memset
  uint8_t* base = static_cast<uint8_t*>(workspace_dev);
  size_t off = 0;
  uint32_t* ticket = reinterpret_cast<uint32_t*>(base);
  off += 256;
  for (int i = 0; i < n; ++i) {
    const jpeg::Carve c = jpeg::sizes(jobs_host[i]);
    jobs_host[i].ticket_dev = ticket;
    jobs_host[i].coef_dev = reinterpret_cast<int16_t*>(base + off), off += c.coef;
    jobs_host[i].chain_dev = reinterpret_cast<uint32_t*>(base + off), off += c.chain;
  }
  for (int i = 0; i < n; ++i) {
    IA_RETURN_IF(!file_dev[i] || !out_dev[i], IA_ERR_INVALID_ARGUMENT);
    const jpeg::Carve c = jpeg::sizes(jobs_host[i]);
    ia_jpeg_job& j = jobs_host[i];
    j.file_dev = file_dev[i];
    j.out_dev = out_dev[i];
    j.clean_dev = base + off, off += c.clean;
    j.segment_dev = reinterpret_cast<uint32_t*>(base + off), off += c.segment;
    j.tile_dev = reinterpret_cast<uint32_t*>(base + off), off += c.tile;
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
  IA_RETURN_IF(n > 65535, IA_ERR_UNSUPPORTED);
  uint32_t max_tiles = 1, max_chunks = 1;
  int max_blocks = 0, max_w = 0, max_h = 0;
  bool any_restart = false;
  for (int i = 0; i < n; ++i) {
    const ia_jpeg_job& j = jobs_host[i];
    any_restart |= j.restart_interval != 0;
    IA_RETURN_IF(!j.coef_dev || !j.clean_dev || !j.out_dev || !j.file_dev || !j.ticket_dev, IA_ERR_INVALID_ARGUMENT);
    max_tiles = tile_capacity(j) > max_tiles ? tile_capacity(j) : max_tiles;
    max_chunks = chunk_capacity(j) > max_chunks ? chunk_capacity(j) : max_chunks;
    max_blocks = j.total_blocks > max_blocks ? j.total_blocks : max_blocks;
    max_w = j.width > max_w ? j.width : max_w;
    max_h = j.height > max_h ? j.height : max_h;
  }
  IA_RETURN_IF((unsigned long long)max_chunks * (unsigned)n > 0x7fffffffull, IA_ERR_UNSUPPORTED);
  cudaError_t e = cudaMemsetAsync(jobs_host[0].ticket_dev, 0, zeroed_bytes(jobs_host, n), stream);
  if (e != cudaSuccess) return (int)e;
  unstuff_count_kernel<<<dim3(max_tiles, n), kUnstuffThreads, 0, stream>>>(jobs_dev);
  int rc = launch_status();
  if (rc != IA_OK) return rc;
  unstuff_scan_kernel<<<n, 32, 0, stream>>>(jobs_dev, status_dev);
  rc = launch_status();
  if (rc != IA_OK) return rc;
  unstuff_scatter_kernel<<<dim3(max_tiles, n), kUnstuffThreads, 0, stream>>>(jobs_dev, status_dev);
  rc = launch_status();
  if (rc != IA_OK) return rc;
  if (any_restart) {  // does nothing for images whose restart markers are in sequence
    resync_walk_kernel<<<n, 32, 0, stream>>>(jobs_dev, status_dev);
    rc = launch_status();
    if (rc != IA_OK) return rc;
  }
  // Configuration of the Huffman kernel (measured cross-overs, see ia_jpeg_kernels.cuh):
  // 256 threads x 128-byte subsequences (32 KB chunks) for a few frames, 256-byte subsequences
  // once the batch holds enough chunks.  The 512-thread instantiations are kept for
  // IA_JPEG_THREADS=512 (IA_JPEG_SUBSEQ=128|256 forces the other choice): they win by 4 % at
  // exactly 64 x 1080p and lose at 32, at 128 and on 4K frames, so they are not chosen.
  unsigned long long total_bytes = 0, ctas_large_sub = 0;
  uint32_t max_scan = 0;
  for (int i = 0; i < n; ++i) {
    const uint32_t bytes = scan_bytes(jobs_host[i]);
    total_bytes += bytes;
    ctas_large_sub += (bytes + kHuffThreads * kSubseqBytesLarge - 1) / (kHuffThreads * kSubseqBytesLarge);
    max_scan = bytes > max_scan ? bytes : max_scan;
  }
  static const int forced_sub = [] { const char* e = getenv("IA_JPEG_SUBSEQ"); return e ? atoi(e) : 0; }();
  static const int forced_threads = [] { const char* e = getenv("IA_JPEG_THREADS"); return e ? atoi(e) : 0; }();
  const unsigned long long avg_bytes = total_bytes / (unsigned)n;
  const bool large_files = avg_bytes >= kLargeFileBytes;
  int threads = kHuffThreads;
  int sub = ctas_large_sub >= kLargeSubseqCtasFrom || (large_files && total_bytes >= kLargeSubseqBytesFrom)
                ? kSubseqBytesLarge
                : kSubseqBytes;
  if (forced_threads == kHuffThreads || forced_threads == kHuffThreadsLarge) threads = forced_threads;
  if (forced_sub == kSubseqBytes || forced_sub == kSubseqBytesLarge) sub = forced_sub;
  const uint32_t chunk_bytes = (uint32_t)threads * (uint32_t)sub;
  const uint32_t chunks = (max_scan + chunk_bytes - 1) / chunk_bytes + 1;
  const unsigned grid = chunks * (unsigned)n;
  uint32_t* ticket = jobs_host[0].ticket_dev;
  if (threads == kHuffThreadsLarge && sub == kSubseqBytesLarge)
    huffman_kernel<kHuffThreadsLarge, 3, kSubseqBytesLarge><<<grid, threads, 0, stream>>>(jobs_dev, (int)chunks, ticket, status_dev);
  else if (threads == kHuffThreadsLarge)
    huffman_kernel<kHuffThreadsLarge, 3, kSubseqBytes><<<grid, threads, 0, stream>>>(jobs_dev, (int)chunks, ticket, status_dev);
  else if (sub == kSubseqBytesLarge)
    huffman_kernel<kHuffThreads, 4, kSubseqBytesLarge><<<grid, threads, 0, stream>>>(jobs_dev, (int)chunks, ticket, status_dev);
  else
    huffman_kernel<kHuffThreads, 4, kSubseqBytes><<<grid, threads, 0, stream>>>(jobs_dev, (int)chunks, ticket, status_dev);
  rc = launch_status();
  if (rc != IA_OK) return rc;
  idct_kernel<<<dim3((max_blocks + kIdctThreads - 1) / kIdctThreads, n), kIdctThreads, 0, stream>>>(jobs_dev, status_dev);
  rc = launch_status();
  if (rc != IA_OK) return rc;
  const int px_per_cta = 32 * kColourPixels;
  colour_kernel<<<dim3((max_w + px_per_cta - 1) / px_per_cta, (max_h + 7) / 8, n), 256, 0, stream>>>(jobs_dev);
  return launch_status();
}
