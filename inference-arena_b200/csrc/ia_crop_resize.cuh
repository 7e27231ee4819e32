// ia_crop_resize.cuh — K4 stage B: ROI resize + normalise + CHW.
// Warp-specialised, double-buffered, persistent.
//
// Work is a stream of steps; a step is up to 7 output rows (half of a 14-row band, fewer when
// the crop rows are too wide for one staging buffer) of one ROI.  CTAs are persistent
// (grid = SMs x resident CTAs) and stride over the (ROI, band) items; the ROI total is read
// from device memory, so nothing returns to the host between NMS and here.  A CTA is
// 7 consumer warps + 1 producer warp, coupled only by mbarriers (full / empty per staging
// buffer) — no CTA-wide barrier in the loop:
//   producer: lane (row j, tap) evaluates OpenCV's yofs/beta entry, duplicates are removed
//           with __match_any_sync and the copy geometry is worked out — all before waiting
//           for the consumers to release the buffer; then the row plan is published and each
//           distinct crop row segment is pulled into the staging buffer with one 1-D bulk
//           async copy (TMA engine, complete_tx on full[buf]) of its 16-byte-aligned superset,
//           clamped to the frame buffer (bytes a clamp cuts off are copied by the lane).
//   consumers: one output row per warp, lanes over consecutive pixels.  The 6 source bytes of
//           a pixel's two taps come from three aligned 32-bit shared loads + a funnel shift; a
//           PRMT pairs (p[c], p[c+3]) and one IDP.2A does p0*a0 + p1*a1 against the packed
//           coefficient pair — OpenCV's 11-bit fixed point, bit-identical to cv2.resize
//           INTER_LINEAR (copy and the exact-2x box average use the same arithmetic with
//           tables a0=b0=2048 resp. a0=a1=b0=b1=1024, which reduce to the identity and to
//           (p00+p01+p10+p11+2)>>2).  Then the ((v/255 - mean)/std) LUT, replicated 4x across
//           banks, and one 128-byte streaming store per warp and colour plane.
// The global side sees only wide, fully used requests; byte gathers hit shared memory.
#pragma once

#include "ia_common.cuh"

namespace ia {

// 7 consumer warps (one output row each per 7-row step; 224 = 7 x 32 pixels per row at
// out_size 224, so no lane idles) + 1 producer warp = 256 threads.
constexpr int kCropConsumers = 224;
constexpr int kCropThreads = kCropConsumers + 32;
constexpr int kBandRows = 14;  // two 7-row steps per band: one output row per consumer warp
constexpr int kLutRep = 4;  // copies of each LUT entry, one per (lane & 3): fewer bank conflicts

struct CropParams {
  const uint8_t* frames;
  size_t frames_bytes;  // n * src_h * src_w * 3: staging never reads outside [frames, frames+bytes)
  int src_h, src_w;
  const ia_frame_desc* descs;  // ragged batches: per-image pointer/height/width (device) or null
  const int* rois;
  const int* total_dev;  // number of valid ROIs (device) or null
  int total_host;        // used when total_dev is null
  int roi_slots;         // ROIs the output buffer can hold
  int out_size, bands;
  int stage_bytes;  // bytes of ONE staging buffer (two follow each other in dynamic smem)
  float mean[3], stdv[3];
  float* out;
};

struct YPlan {
  int off0, off1;  // byte offsets (within a staging buffer) of crop byte 0 of the two rows
  int b0, b1;
};

// Position in the step stream; producer and consumers advance their own copy identically.
struct Cursor {
  int item, sb, nsub, sub_rows;
  int r, y_band, rows;  // ROI index, first output row of the band, rows in the band
  int empty, cw, ch, mode;
  int slot_bytes;
  const uint8_t* crop;     // first byte of the crop rectangle in the frame
  uint32_t row_bytes;      // pitch of the frame the ROI lives in
  uintptr_t buf_lo, buf_hi;  // that frame's allocation: staging never reads outside it
};

__device__ __forceinline__ bool load_item(Cursor& c, const CropParams& p, int item, int items) {
  c.item = item;
  c.sb = 0;
  if (item >= items) return false;
  const int S = p.out_size;
  c.r = item / p.bands;
  c.y_band = (item - c.r * p.bands) * kBandRows;
  c.rows = min(kBandRows, S - c.y_band);
  const int* roi = p.rois + (size_t)c.r * 8;
  c.empty = roi[5];
  c.cw = c.empty ? 1 : roi[3] - roi[1];
  c.ch = c.empty ? 1 : roi[4] - roi[2];
  c.mode = resize_mode(c.ch, c.cw, S, S);
  if (p.descs != nullptr) {
    const ia_frame_desc d = p.descs[roi[0]];
    c.row_bytes = (uint32_t)d.w * 3u;
    c.buf_lo = (uintptr_t)d.data_dev;
    c.buf_hi = c.buf_lo + (size_t)d.h * c.row_bytes;
    c.crop = d.data_dev + (size_t)roi[2] * c.row_bytes + (size_t)roi[1] * 3;
  } else {
    c.row_bytes = (uint32_t)p.src_w * 3u;
    c.buf_lo = (uintptr_t)p.frames;
    c.buf_hi = c.buf_lo + p.frames_bytes;
    c.crop = p.frames + ((size_t)roi[0] * p.src_h + roi[2]) * c.row_bytes + (size_t)roi[1] * 3;
  }
  c.slot_bytes = (c.cw * 3 + 30 + 15) & ~15;  // aligned superset of one crop row segment
  const int max_slots = p.stage_bytes / c.slot_bytes;
  const int cap = max(1, min(kBandRows, max_slots >> 1));  // two source rows per output row
  c.nsub = (c.rows + cap - 1) / cap;
  c.sub_rows = (c.rows + c.nsub - 1) / c.nsub;  // balanced split
  return true;
}

__device__ __forceinline__ bool advance(Cursor& c, const CropParams& p, int items) {
  if (++c.sb < c.nsub) return true;
  return load_item(c, p, c.item + (int)gridDim.x, items);
}

// Producer warp: plan + stage the crop rows of step `c`; arrives on `full`.
// Everything that does not touch shared memory (coefficients, duplicate removal, copy
// geometry) is computed BEFORE waiting for the consumers to release the buffer.
__device__ __forceinline__ void produce_step(const Cursor& c, const CropParams& p, double sc_y,
                                             uint8_t* stage, YPlan* plan, uint64_t* full,
                                             uint64_t* empty, int empty_parity, double* scx_slot,
                                             double sc_x, bool new_roi, int lane) {
  if (c.empty) {  // nothing to stage; keep the barrier protocol uniform
    if (empty_parity >= 0) mbar_wait(empty, (uint32_t)empty_parity);
    if (lane == 0) mbar_expect_tx(full, 0);
    return;
  }
  const int S = p.out_size;
  const size_t row_bytes = c.row_bytes;
  const uintptr_t buf_lo = c.buf_lo, buf_hi = c.buf_hi;
  const int y0 = c.y_band + c.sb * c.sub_rows;
  const int nr = min(c.sub_rows, c.rows - c.sb * c.sub_rows);
  const uint32_t seg = (uint32_t)c.cw * 3u;
  const int j = lane >> 1, tap = lane & 1;
  int row = -1, b0 = kCoefOne, b1 = 0;
  if (j < nr) {
    const int y = y0 + j;
    if (c.mode == kResizeCopy) {
      row = tap ? -1 : y;
    } else if (c.mode == kResizeBox2) {
      row = 2 * y + tap;
      b0 = b1 = kCoefOne / 2;
    } else {
      const AxisCoef k = axis_coef(y, sc_y, c.ch, false);
      b0 = k.w0;
      b1 = k.w1;
      row = tap ? (k.w1 != 0 ? clampi(k.s0 + 1, 0, c.ch - 1) : -1) : clampi(k.s0, 0, c.ch - 1);
    }
  }
  const unsigned same = __match_any_sync(0xffffffffu, row);
  const int leader = __ffs(same) - 1;
  const bool copy = leader == lane && row >= 0;
  const unsigned leaders = __ballot_sync(0xffffffffu, copy);
  const int slot = __popc(leaders & ((1u << leader) - 1u));
  const uintptr_t g_lo = (uintptr_t)(c.crop + (size_t)max(row, 0) * row_bytes);
  const int off = slot * c.slot_bytes + (int)(g_lo & 15);
  uintptr_t a_lo = 0, a_hi = 0;
  uint32_t bulk = 0;
  uint8_t* s = stage + off;  // where crop byte g_lo lands
  if (copy) {
    a_lo = max(g_lo & ~(uintptr_t)15, (buf_lo + 15) & ~(uintptr_t)15);
    a_hi = min((g_lo + seg + 15) & ~(uintptr_t)15, buf_hi & ~(uintptr_t)15);
    if (a_hi > a_lo) bulk = (uint32_t)(a_hi - a_lo);
  }
  uint32_t tx = bulk;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) tx += __shfl_xor_sync(0xffffffffu, tx, o);

  if (empty_parity >= 0) mbar_wait(empty, (uint32_t)empty_parity);  // consumers released the buffer
  if (new_roi && lane == 0) *scx_slot = sc_x;  // OpenCV scale_x for the consumers' x table
  if (copy) {
    if (bulk) {
      for (uintptr_t q = g_lo; q < a_lo; ++q) s[q - g_lo] = *reinterpret_cast<const uint8_t*>(q);
      for (uintptr_t q = max(a_hi, g_lo); q < g_lo + seg; ++q) s[q - g_lo] = *reinterpret_cast<const uint8_t*>(q);
    } else {
      for (uint32_t q = 0; q < seg; ++q) s[q] = reinterpret_cast<const uint8_t*>(g_lo)[q];
    }
  }
  if (j < nr) {
    if (tap == 0) {
      plan[j].off0 = off;
      plan[j].b0 = b0;
      plan[j].b1 = b1;
    } else {
      plan[j].off1 = row >= 0 ? off : 0;  // unused tap: any valid offset, its weight is 0
    }
  }
  __syncwarp();  // plan and lane-copied bytes are ordered before the releasing arrive
  if (lane == 0) mbar_expect_tx(full, tx);
  __syncwarp();
  if (bulk) bulk_g2s(s + (ptrdiff_t)(a_lo - g_lo), reinterpret_cast<const void*>(a_lo), bulk, full);
}

__global__ void __launch_bounds__(kCropThreads, 4) crop_resize_kernel(const CropParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ float lut[3][256 * kLutRep];
  __shared__ YPlan yplan[2][kBandRows];
  __shared__ double scx[2];  // OpenCV scale_x of the step's ROI, from the producer
  __shared__ __align__(8) uint64_t full[2], empty[2];

  const int tid = threadIdx.x, lane = tid & 31;
  const int S = p.out_size;
  const int xg = (S + 3) >> 2;
  int* xoff = reinterpret_cast<int*>(smem + 2 * (size_t)p.stage_bytes);  // [2][4*xg] int2 (offset, coefs)
  const int total = p.total_dev ? min(*p.total_dev, p.roi_slots) : p.total_host;
  const int items = total * p.bands;
  const size_t plane = (size_t)S * S;

  for (int i = tid; i < 256 * kLutRep; i += kCropThreads) {
    const float u = __fdiv_rn((float)(i / kLutRep), 255.0f);  // transforms.py:263
#pragma unroll
    for (int c = 0; c < 3; ++c) lut[c][i] = __fdiv_rn(__fsub_rn(u, p.mean[c]), p.stdv[c]);  // :270
  }
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_init(&empty[0], kCropConsumers / 32);
    mbar_init(&empty[1], kCropConsumers / 32);
    mbar_fence_init();
  }
  __syncthreads();

  Cursor cur;
  bool have = load_item(cur, p, blockIdx.x, items);

  if (tid >= kCropConsumers) {
    // =============================== producer warp ====================================
    double sc_y = 0.0;
    for (int s = 0; have; ++s) {
      const int buf = s & 1;
      const int empty_parity = s >= 2 ? (((s >> 1) - 1) & 1) : -1;
      double sc_x = 0.0;
      const bool new_roi = cur.sb == 0 && !cur.empty;
      if (new_roi) {
        sc_y = axis_scale(S, cur.ch);
        sc_x = axis_scale(S, cur.cw);
      }
      produce_step(cur, p, sc_y, smem + (size_t)buf * p.stage_bytes, yplan[buf], &full[buf], &empty[buf],
                   empty_parity, &scx[buf], sc_x, new_roi, lane);
      have = advance(cur, p, items);
    }
    return;
  }

  // ================================= consumer warps ====================================
  const int rep = lane & (kLutRep - 1);
  int n_item = 0;
  for (int s = 0; have; ++s) {
    const int buf = s & 1;
    mbar_wait(&full[buf], (s >> 1) & 1);
    if (cur.sb == 0) {
      if (!cur.empty) {
        // x table of this ROI, double-buffered by the parity of the NON-EMPTY items (an empty
        // ROI has no table and no barrier, so it must not flip the parity), then a
        // consumers-only barrier: a warp can only overwrite a table two barriers after the
        // slowest warp last read it
        ++n_item;
        int2* xt_w = reinterpret_cast<int2*>(xoff + (n_item & 1) * 8 * xg);
        const double sc_x = scx[buf];
        for (int x = tid; x < S; x += kCropConsumers) {
          int xo, a0, a1;
          if (cur.mode == kResizeCopy) {
            xo = x * 3, a0 = kCoefOne, a1 = 0;
          } else if (cur.mode == kResizeBox2) {
            xo = x * 6, a0 = a1 = kCoefOne / 2;
          } else {
            const AxisCoef k = axis_coef(x, sc_x, cur.cw, true);
            xo = k.s0 * 3, a0 = k.w0, a1 = k.w1;
          }
          xt_w[x] = make_int2(xo, a0 | (a1 << 16));
        }
        named_barrier_sync(1, kCropConsumers);
      }
    }

    const uint32_t* stage32 = reinterpret_cast<const uint32_t*>(smem + (size_t)buf * p.stage_bytes);
    const int2* xt = reinterpret_cast<const int2*>(xoff + (n_item & 1) * 8 * xg);  // (byte offset, a0 | a1 << 16)
    const int y0 = cur.y_band + cur.sb * cur.sub_rows;
    const int nr = min(cur.sub_rows, cur.rows - cur.sb * cur.sub_rows);
    const int warp = tid >> 5;

    // one output row per warp, lanes over consecutive pixels: neighbouring lanes read
    // neighbouring (often the same) staging words and write one 128-byte line per plane
    for (int j = warp; j < nr; j += kCropConsumers / 32) {
      float* orow = p.out + ((size_t)cur.r * 3) * plane + (size_t)(y0 + j) * S;
      if (cur.empty) {
        for (int x = lane; x < S; x += 32) {  // mobilenet_preprocess.py:266-267: a 1x1 black crop
#pragma unroll
          for (int c = 0; c < 3; ++c) st_stream_f1(orow + c * plane + x, lut[c][rep]);
        }
        continue;
      }
      const YPlan yp = yplan[buf][j];
      // one pointer per colour plane, advanced by a constant: under the unroll the stores use
      // immediate offsets instead of 64-bit address arithmetic per element
      float* o0 = orow + lane;
      float* o1 = o0 + plane;
      float* o2 = o1 + plane;
      auto pixel = [&](const int2 e, float* d0, float* d1, float* d2) {
        uint32_t lo0, hi0, lo1, hi1;
        load6(stage32, yp.off0 + e.x, lo0, hi0);
        load6(stage32, yp.off1 + e.x, lo1, hi1);
        float* const dst[3] = {d0, d1, d2};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          // h = p[c]*a0 + p[c+3]*a1 in one IDP.2A; a zero weight multiplies whatever byte sits
          // there (always inside the staging buffer) by 0
          const int h0 = (int)__dp2a_lo((uint32_t)e.y, tap_pair(lo0, hi0, c), 0u);
          const int h1 = (int)__dp2a_lo((uint32_t)e.y, tap_pair(lo1, hi1, c), 0u);
          const int v = vresize(h0, h1, yp.b0, yp.b1);
          st_stream_f1(dst[c], lut[c][v * kLutRep + rep]);
        }
      };
      if (S == 7 * 32) {  // MobileNet's 224: a fixed trip count, so every offset is an immediate
        const int2* xl = xt + lane;
#pragma unroll
        for (int k = 0; k < 7; ++k) pixel(xl[32 * k], o0 + 32 * k, o1 + 32 * k, o2 + 32 * k);
      } else {
#pragma unroll 4
        for (int x = lane; x < S; x += 32, o0 += 32, o1 += 32, o2 += 32) pixel(xt[x], o0, o1, o2);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[buf]);  // this warp is done with the staging buffer + plan
    have = advance(cur, p, items);
  }
}

static int launch_crops(const uint8_t* frames, size_t frames_bytes, int src_h, int src_w,
                        const ia_frame_desc* descs, const int* rois, const int* total_dev, int total_host, int roi_slots,
                        int out_size, const float mean[3], const float stdv[3], float* out,
                        cudaStream_t stream) {
  if (roi_slots == 0) return IA_OK;
  CropParams c;
  c.frames = frames;
  c.frames_bytes = frames_bytes;
  c.src_h = src_h;
  c.src_w = src_w;
  c.descs = descs;
  c.rois = rois;
  c.total_dev = total_dev;
  c.total_host = total_host;
  c.roi_slots = roi_slots;
  c.out_size = out_size;
  c.bands = (out_size + kBandRows - 1) / kBandRows;
  for (int i = 0; i < 3; ++i) {
    c.mean[i] = mean[i];
    c.stdv[i] = stdv[i];
  }
  c.out = out;
  const long long items = (long long)c.bands * roi_slots;
  IA_RETURN_IF(items > 0x7fffffffLL, IA_ERR_UNSUPPORTED);
  // two staging buffers; each holds at least two full-width rows, normally 19 KB
  const size_t widest = ((size_t)src_w * 3 + 30 + 15) & ~(size_t)15;
  size_t stage = 19 * 1024;  // with the 4x LUT and 64 registers: 4 CTAs (28 consumer warps) per SM
  if (stage < 2 * widest) stage = 2 * widest;
  const size_t xg = ((size_t)out_size + 3) / 4;
  const size_t smem = 2 * stage + 4 * xg * sizeof(int4);
  IA_RETURN_IF(smem > 180 * 1024 || stage >= (1u << 20), IA_ERR_UNSUPPORTED);
  c.stage_bytes = (int)stage;
  int dev = 0, sms = 148, per_sm = 1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(crop_resize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e == cudaSuccess)
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, crop_resize_kernel, kCropThreads, smem);
  if (e != cudaSuccess) return (int)e;
  if (per_sm < 1) per_sm = 1;
  const long long resident = (long long)sms * per_sm;
  const long long grid = items < resident ? items : resident;
  crop_resize_kernel<<<(unsigned)grid, kCropThreads, smem, stream>>>(c);
  return launch_status();
}

}  // namespace ia
