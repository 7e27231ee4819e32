// ia_crop_resize.cuh — K4 stage B: ROI resize + normalise + CHW, double-buffered row staging.
//
// Work is a stream of steps; a step is a sub-band (<= 16 output rows, sized to the staging
// buffer) of one ROI.  CTAs are persistent (grid = SMs x resident CTAs) and stride over the
// (ROI, band) items; the ROI total is read from device memory, so nothing returns to the host
// between NMS and here.  Per step:
//   issue   (warp 0, one step AHEAD of the math): lane (row j, tap) evaluates OpenCV's
//           yofs/beta entry, duplicates are removed with __match_any_sync, and each distinct
//           crop row segment is pulled into the idle staging buffer with one 1-D bulk async
//           copy (TMA engine, mbarrier completion) of its 16-byte-aligned superset, clamped
//           to the frame buffer (bytes a clamp cuts off are copied by the lane);
//   compute (all warps, warp per output row): 4 consecutive pixels per lane — branch-free
//           11-bit fixed-point bilinear on bytes from shared memory (bit-identical to
//           cv2.resize INTER_LINEAR; copy and the exact-2x box average use the same arithmetic
//           with tables a0=b0=2048 resp. a0=a1=b0=b1=1024, which reduce to the identity and to
//           (p00+p01+p10+p11+2)>>2), the ((v/255 - mean)/std) LUT replicated 8x across banks,
//           three 128-bit streaming plane stores.
// The global side sees only wide, fully used requests; byte gathers hit shared memory.
#pragma once

#include "ia_common.cuh"

namespace ia {

constexpr int kCropThreads = 256;
constexpr int kCropWarps = kCropThreads / 32;
constexpr int kBandRows = 16;
constexpr int kLutRep = 8;  // copies of each LUT entry, one per (lane & 7): fewer bank conflicts

struct CropParams {
  const uint8_t* frames;
  size_t frames_bytes;  // n * src_h * src_w * 3: staging never reads outside [frames, frames+bytes)
  int src_h, src_w;
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
  int off0, off1;  // shared-memory byte offsets (within a staging buffer) of the two crop rows
  int b0, b1;
};

// Position in the step stream; advanced identically by every thread of the CTA.
struct Cursor {
  int item, sb, nsub, sub_rows;
  int r, y_band, rows;  // ROI index, first output row of the band, rows in the band
  int empty, cw, ch, mode;
  int slot_bytes;
  const uint8_t* crop;  // first byte of the crop rectangle in the frame
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
  c.crop = p.frames + ((size_t)roi[0] * p.src_h + roi[2]) * ((size_t)p.src_w * 3) + (size_t)roi[1] * 3;
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

// Source rows / weights of output row y of the crop (OpenCV's vertical table).
__device__ __forceinline__ void y_entry(const Cursor& c, int S, int y, int tap, int& row, int& b0, int& b1) {
  b0 = kCoefOne;
  b1 = 0;
  if (c.mode == kResizeCopy) {
    row = tap ? -1 : y;
  } else if (c.mode == kResizeBox2) {
    row = 2 * y + tap;
    b0 = b1 = kCoefOne / 2;
  } else {
    const AxisCoef k = axis_coef(y, axis_scale(S, c.ch), c.ch, false);
    b0 = k.w0;
    b1 = k.w1;
    row = tap ? (k.w1 != 0 ? clampi(k.s0 + 1, 0, c.ch - 1) : -1) : clampi(k.s0, 0, c.ch - 1);
  }
}

// Warp 0: stage the crop rows of step `c` into `stage`, publish the row plan.
__device__ __forceinline__ void issue_step(const Cursor& c, const CropParams& p, uint8_t* stage,
                                           YPlan* plan, uint64_t* bar, int lane) {
  const int S = p.out_size;
  const size_t row_bytes = (size_t)p.src_w * 3;
  const uintptr_t buf_lo = (uintptr_t)p.frames, buf_hi = buf_lo + p.frames_bytes;
  const int y0 = c.y_band + c.sb * c.sub_rows;
  const int nr = min(c.sub_rows, c.rows - c.sb * c.sub_rows);
  const uint32_t seg = (uint32_t)c.cw * 3u;
  const int j = lane >> 1, tap = lane & 1;
  int row = -1, b0 = kCoefOne, b1 = 0;
  if (j < nr) y_entry(c, S, y0 + j, tap, row, b0, b1);
  const unsigned same = __match_any_sync(0xffffffffu, row);
  const int leader = __ffs(same) - 1;
  const bool copy = leader == lane && row >= 0;
  const unsigned leaders = __ballot_sync(0xffffffffu, copy);
  const int slot = __popc(leaders & ((1u << leader) - 1u));
  const uintptr_t g_lo = (uintptr_t)(c.crop + (size_t)max(row, 0) * row_bytes);
  const int off = slot * c.slot_bytes + (int)(g_lo & 15);
  uintptr_t a_lo = 0, a_hi = 0;
  uint32_t bulk = 0;
  if (copy) {
    a_lo = max(g_lo & ~(uintptr_t)15, (buf_lo + 15) & ~(uintptr_t)15);
    a_hi = min((g_lo + seg + 15) & ~(uintptr_t)15, buf_hi & ~(uintptr_t)15);
    if (a_hi > a_lo) bulk = (uint32_t)(a_hi - a_lo);
  }
  uint32_t tx = bulk;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) tx += __shfl_xor_sync(0xffffffffu, tx, o);
  if (lane == 0) mbar_expect_tx(bar, tx);
  __syncwarp();
  if (copy) {
    uint8_t* s = stage + off;  // where crop byte g_lo lands
    if (bulk) {
      bulk_g2s(s + (ptrdiff_t)(a_lo - g_lo), reinterpret_cast<const void*>(a_lo), bulk, bar);
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
}

__global__ void __launch_bounds__(kCropThreads) crop_resize_kernel(const CropParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ float lut[3][256 * kLutRep];
  __shared__ YPlan yplan[2][kBandRows];
  __shared__ __align__(8) uint64_t bar[2];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int S = p.out_size;
  const int xg = (S + 3) >> 2;
  int4* xoff4 = reinterpret_cast<int4*>(smem + 2 * (size_t)p.stage_bytes);  // [xg] byte offsets
  int4* coef4 = xoff4 + xg;                                               // [xg] a0 | a1 << 16
  const int total = p.total_dev ? min(*p.total_dev, p.roi_slots) : p.total_host;
  const int items = total * p.bands;
  const bool vec_ok = (S & 3) == 0;
  const size_t plane = (size_t)S * S;

  for (int i = tid; i < 256 * kLutRep; i += kCropThreads) {
    const float u = __fdiv_rn((float)(i / kLutRep), 255.0f);  // transforms.py:263
#pragma unroll
    for (int c = 0; c < 3; ++c) lut[c][i] = __fdiv_rn(__fsub_rn(u, p.mean[c]), p.stdv[c]);  // :270
  }
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    mbar_fence_init();
  }
  __syncthreads();

  Cursor cur;
  bool have = load_item(cur, p, blockIdx.x, items);
  if (have && !cur.empty && tid < 32) issue_step(cur, p, smem, yplan[0], &bar[0], lane);
  uint32_t phase0 = 0, phase1 = 0;

  for (int s = 0; have; ++s) {
    const int buf = s & 1;
    Cursor nxt = cur;
    const bool have_next = advance(nxt, p, items);
    // prefetch the next step into the other buffer (its readers finished one iteration ago)
    if (have_next && !nxt.empty && tid < 32)
      issue_step(nxt, p, smem + (size_t)(buf ^ 1) * p.stage_bytes, yplan[buf ^ 1], &bar[buf ^ 1], lane);

    // x table of this ROI, once per item (previous item's readers are past the last barrier)
    if (cur.sb == 0 && !cur.empty) {
      for (int x = tid; x < 4 * xg; x += kCropThreads) {
        const int xc = min(x, S - 1);
        int xo, a0, a1;
        if (cur.mode == kResizeCopy) {
          xo = xc * 3, a0 = kCoefOne, a1 = 0;
        } else if (cur.mode == kResizeBox2) {
          xo = xc * 6, a0 = a1 = kCoefOne / 2;
        } else {
          const AxisCoef k = axis_coef(xc, axis_scale(S, cur.cw), cur.cw, true);
          xo = k.s0 * 3, a0 = k.w0, a1 = k.w1;
        }
        reinterpret_cast<int*>(xoff4)[x] = xo;
        reinterpret_cast<int*>(coef4)[x] = a0 | (a1 << 16);
      }
    }
    __syncthreads();  // x table, row plan and lane-copied bytes of this step are visible
    if (!cur.empty) {
      if (buf == 0) {
        mbar_wait(&bar[0], phase0);
        phase0 ^= 1;
      } else {
        mbar_wait(&bar[1], phase1);
        phase1 ^= 1;
      }
    }

    // ---- compute: warp per output row, 4 pixels x 3 channels per lane --------------------
    const uint8_t* stage = smem + (size_t)buf * p.stage_bytes;
    const int y0 = cur.y_band + cur.sb * cur.sub_rows;
    const int nr = min(cur.sub_rows, cur.rows - cur.sb * cur.sub_rows);
    const int rep = lane & (kLutRep - 1);
    for (int j = warp; j < nr; j += kCropWarps) {
      const YPlan yp = yplan[buf][j];
      const uint8_t* s0 = stage + yp.off0;
      const uint8_t* s1 = stage + yp.off1;
      float* orow = p.out + ((size_t)cur.r * 3) * plane + (size_t)(y0 + j) * S;
      for (int g = lane; g < xg; g += 32) {
        uint32_t px[4][3];
        if (cur.empty) {
#pragma unroll
          for (int k = 0; k < 4; ++k) px[k][0] = px[k][1] = px[k][2] = 0;  // mobilenet_preprocess.py:266-267
        } else {
          const int4 xo4 = xoff4[g];
          const int4 cf4 = coef4[g];
          const int xo[4] = {xo4.x, xo4.y, xo4.z, xo4.w};
          const int cf[4] = {cf4.x, cf4.y, cf4.z, cf4.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int a0 = cf[k] & 0xffff, a1 = cf[k] >> 16;
            const uint8_t* a = s0 + xo[k];
            const uint8_t* b = s1 + xo[k];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              // branch-free: a zero weight multiplies whatever byte sits there (always inside
              // the staging buffer) by 0, exactly as if the tap had been skipped
              const int h0 = a[c] * a0 + a[c + 3] * a1;
              const int h1 = b[c] * a0 + b[c + 3] * a1;
              px[k][c] = (uint32_t)vresize(h0, h1, yp.b0, yp.b1);
            }
          }
        }
        float* o = orow + (g << 2);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float* l = lut[c] + rep;
          if (vec_ok) {
            st_stream_f4(o + c * plane, l[px[0][c] * kLutRep], l[px[1][c] * kLutRep],
                         l[px[2][c] * kLutRep], l[px[3][c] * kLutRep]);
          } else {
            for (int k = 0; k < 4 && (g << 2) + k < S; ++k)
              st_stream_f1(o + c * plane + k, l[px[k][c] * kLutRep]);
          }
        }
      }
    }
    __syncthreads();  // everyone is done with this staging buffer, its plan and the x table
    cur = nxt;
    have = have_next;
  }
}

static int launch_crops(const uint8_t* frames, size_t frames_bytes, int src_h, int src_w,
                        const int* rois, const int* total_dev, int total_host, int roi_slots,
                        int out_size, const float mean[3], const float stdv[3], float* out,
                        cudaStream_t stream) {
  if (roi_slots == 0) return IA_OK;
  CropParams c;
  c.frames = frames;
  c.frames_bytes = frames_bytes;
  c.src_h = src_h;
  c.src_w = src_w;
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
  // two staging buffers; each holds at least two full-width rows, normally 24 KB
  const size_t widest = ((size_t)src_w * 3 + 30 + 15) & ~(size_t)15;
  size_t stage = 24 * 1024;
  if (stage < 2 * widest) stage = 2 * widest;
  const size_t xg = ((size_t)out_size + 3) / 4;
  const size_t smem = 2 * stage + 2 * xg * sizeof(int4);
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
