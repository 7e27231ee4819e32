// ia_crop_resize.cuh — K4 stage B: ROI resize + normalise + CHW.
// Warp-specialised, double-buffered, persistent; separable like cv::resize itself.
//
// One CTA per SM holds up to four independent TEAMS of 7 consumer warps + 1 producer warp and
// one shared table of ((v/255 - mean)/std), replicated once per lane (bank) so that a warp's
// 32 look-ups of 32 unrelated pixel values are one shared-memory wavefront.  A team strides
// over work items of 28 output rows of one ROI (the ROI total is read from device memory:
// nothing returns to the host between NMS and here) and walks an item in steps of up to 14
// rows, coupled to its producer only by mbarriers (full / empty per staging buffer):
//   producer: lane (row j, tap) evaluates OpenCV's yofs/beta entry, duplicates are removed
//           with __match_any_sync and the copy geometry is worked out — all before waiting
//           for the consumers to release the buffer; then the row plan is published and each
//           distinct crop row segment is pulled into the staging buffer with one 1-D bulk
//           async copy (TMA engine, complete_tx on full[buf]) of its 16-byte-aligned superset,
//           clamped to the frame buffer (bytes a clamp cuts off are copied by the lane).
//   consumers: a lane owns ONE output column for the whole item and walks down the rows; a
//           warp is 32 neighbouring columns, so every plane store is one 128-byte line.  The
//           lane's x coefficients live in registers (no table in shared memory).  The
//           horizontal pass of a source row — three aligned 32-bit shared loads + a funnel
//           shift for the 6 bytes of the two taps, a PRMT pairing (p[c], p[c+3]) and one
//           IDP.2A against the packed coefficient pair, >> 4 — is evaluated once per source
//           row and kept in registers for as long as output rows use it (two rows are live:
//           an up-scaled crop reuses each ~2.5 times), which is the order cv::resize works in.
//           The vertical step is two IMAD.HI against coefficients pre-shifted by 16 — OpenCV's
//           11-bit fixed point, bit-identical to cv2.resize INTER_LINEAR (copy and the exact-2x
//           box average use the same arithmetic with tables a0=b0=2048 resp.
//           a0=a1=b0=b1=1024, which reduce to the identity and to (p00+p01+p10+p11+2)>>2).
// The global side sees only wide, fully used requests; byte gathers hit shared memory.
#pragma once

#include "ia_common.cuh"

namespace ia {

// 7 consumer warps (224 = 7 x 32 columns at out_size 224, so no lane idles) + 1 producer warp.
constexpr int kTeamConsumers = 224;
constexpr int kTeamThreads = kTeamConsumers + 32;
constexpr int kMaxTeams = 4;
constexpr int kItemRows = 28;  // output rows of one work item (8 items per 224-row ROI)
constexpr int kStepRows = 14;  // rows per staging step: producer lane = (row, tap)
constexpr int kStages = 2;     // staging buffers per team
constexpr int kLutRep = 32;    // one copy of the 3x256 table per bank
constexpr int kLutFloats = 3 * 256 * kLutRep;

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
  int teams;        // teams per CTA (blockDim.x = teams * kTeamThreads)
  int stage_bytes;  // bytes of ONE staging buffer (each team has kStages)
  // Work distribution.  With `tickets` (two zeroed ints: next item, teams finished) a team draws
  // its items one at a time, so ROIs of very different cost still finish together; the last
  // team to finish zeroes both again.  Without (the ROI table is read-only) teams stride.
  int* tickets;
  float mean[3], stdv[3];
  float* out;
};

// One output row of a step.  A consumer lane keeps the horizontal passes of two source rows in
// registers X and Y (a source row lives in the register of its parity, so the two taps of an
// output row never collide); the producer, which knows the whole row sequence, tells it
// which of them to refill — the shared-memory address of crop byte 0 of the staged row, or
// -1 to keep what is there — and the vertical weight of each, shifted left by 16 for IMAD.HI
// (0 for a tap that is not used).
struct __align__(16) YPlan {
  int src_x, src_y;
  uint32_t w_x, w_y;
};

enum : int { kStepFirst = 1, kStepEmpty = 2, kStepModeShift = 2 };

// What the consumers need to know about a step; they never read the ROI table themselves.
struct __align__(16) StepInfo {
  int item;    // -1: no more work
  int roi;     // ROI slot in the output
  int y0, nr;  // first output row and number of rows of the step
  int flags;   // kStepFirst: first step of an item; kStepEmpty: empty ROI; mode << kStepModeShift
  int cw;      // crop width
  double scx;  // OpenCV scale_x of the ROI
};

// Per-team control block in shared memory.
struct __align__(16) TeamCtl {
  YPlan plan[kStages][kStepRows];
  StepInfo info[kStages];
  uint64_t full[kStages], empty[kStages];
};

// The ROI record of a work item (and its frame, for ragged batches), read ahead of its use.
struct ItemRecord {
  int item;
  int4 a, b;  // image x1 y1 x2 | y2 empty - -
  ia_frame_desc d;
};

__device__ __forceinline__ ItemRecord fetch_item(const CropParams& p, int item) {
  ItemRecord r;
  r.item = item;
  r.a = make_int4(0, 0, 0, 0);
  r.b = make_int4(0, 1, 0, 0);
  r.d.data_dev = nullptr;
  r.d.h = r.d.w = 0;
  if (item >= 0) {
    const int4* roi = reinterpret_cast<const int4*>(p.rois + (size_t)(item / p.bands) * 8);
    r.a = roi[0];
    r.b = roi[1];
    if (p.descs != nullptr) r.d = p.descs[r.a.x];
  }
  return r;
}

// A work item as the producer sees it.
struct Cursor {
  int item, sb, nsub, sub_rows;
  int r, y_band, rows;  // ROI index, first output row of the item, rows in the item
  int empty, cw, ch, mode;
  int slot_bytes;
  const uint8_t* crop;     // first byte of the crop rectangle in the frame
  uint32_t row_bytes;      // pitch of the frame the ROI lives in
  uintptr_t buf_lo, buf_hi;  // that frame's allocation: staging never reads outside it
};

__device__ __forceinline__ void load_item(Cursor& c, const CropParams& p, const ItemRecord& rec) {
  c.item = rec.item;
  c.sb = 0;
  const int S = p.out_size;
  c.r = rec.item / p.bands;
  c.y_band = (rec.item - c.r * p.bands) * kItemRows;
  c.rows = min(kItemRows, S - c.y_band);
  c.empty = rec.b.y;
  c.cw = c.empty ? 1 : rec.a.w - rec.a.y;
  c.ch = c.empty ? 1 : rec.b.x - rec.a.z;
  c.mode = resize_mode(c.ch, c.cw, S, S);
  if (p.descs != nullptr) {
    c.row_bytes = (uint32_t)rec.d.w * 3u;
    c.buf_lo = (uintptr_t)rec.d.data_dev;
    c.buf_hi = c.buf_lo + (size_t)rec.d.h * c.row_bytes;
    c.crop = rec.d.data_dev + (size_t)rec.a.z * c.row_bytes + (size_t)rec.a.y * 3;
  } else {
    c.row_bytes = (uint32_t)p.src_w * 3u;
    c.buf_lo = (uintptr_t)p.frames;
    c.buf_hi = c.buf_lo + p.frames_bytes;
    c.crop = p.frames + ((size_t)rec.a.x * p.src_h + rec.a.z) * c.row_bytes + (size_t)rec.a.y * 3;
  }
  c.slot_bytes = (c.cw * 3 + 30 + 15) & ~15;  // aligned superset of one crop row segment
  const int max_slots = p.stage_bytes / c.slot_bytes;
  const int cap = max(1, min(kStepRows, max_slots >> 1));  // two source rows per output row
  c.nsub = (c.rows + cap - 1) / cap;
  c.sub_rows = (c.rows + c.nsub - 1) / c.nsub;  // balanced split
}

// Producer warp: plan + stage the crop rows of step `c`; arrives on `full`.
// Everything that does not touch shared memory (coefficients, duplicate removal, copy
// geometry, which register each source row goes to) is computed BEFORE waiting for the
// consumers to release the buffer.  held_x / held_y: the crop rows the consumers' registers hold
// on entry (uniform across the warp), updated to what they hold after the step.
__device__ __forceinline__ void produce_step(const Cursor& c, double sc_y, double sc_x, uint8_t* stage,
                                             TeamCtl& ctl, int buf, int empty_parity, int& held_x,
                                             int& held_y, int lane) {
  uint64_t* full = &ctl.full[buf];
  uint64_t* empty = &ctl.empty[buf];
  const int y0 = c.y_band + c.sb * c.sub_rows;
  const int nr = min(c.sub_rows, c.rows - c.sb * c.sub_rows);
  StepInfo info;
  info.item = c.item;
  info.roi = c.r;
  info.y0 = y0;
  info.nr = nr;
  info.flags = (c.sb == 0 ? kStepFirst : 0) | (c.empty ? kStepEmpty : 0) | (c.mode << kStepModeShift);
  info.cw = c.cw;
  info.scx = sc_x;
  if (c.empty) {  // nothing to stage; keep the barrier protocol uniform
    if (empty_parity >= 0) mbar_wait(empty, (uint32_t)empty_parity);
    if (lane == 0) {
      ctl.info[buf] = info;
      mbar_expect_tx(full, 0);
    }
    return;
  }
  YPlan* plan = ctl.plan[buf];
  const size_t row_bytes = c.row_bytes;
  const uintptr_t buf_lo = c.buf_lo, buf_hi = c.buf_hi;
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
  // ---- which register: a row goes to the register of its parity; a second tap clamped onto the
  // first tap's row (bottom border) and an unused tap take the other one
  const int row_first = __shfl_sync(0xffffffffu, row, lane & ~1);  // tap 0's row of this output row
  const bool valid = row >= 0;
  const int reg = tap == 0 ? (row & 1) : ((valid && row != row_first) ? (row & 1) : ((row_first & 1) ^ 1));
  // what the register held before this output row: the closest earlier valid lane with the same
  // register, else what it held on entry
  const unsigned same_reg = __ballot_sync(0xffffffffu, reg == 1);
  const unsigned valid_mask = __ballot_sync(0xffffffffu, valid);
  const unsigned mine = (reg ? same_reg : ~same_reg) & valid_mask;
  const unsigned before = mine & ((1u << lane) - 1u);
  const int prev_lane = 31 - __clz(before);  // -1 when none
  const int prev_row_l = __shfl_sync(0xffffffffu, row, prev_lane & 31);
  const int prev_row = before ? prev_row_l : (reg ? held_y : held_x);
  const bool refill = valid && row != prev_row;
  {  // registers after the step
    const unsigned mx = ~same_reg & valid_mask, my = same_reg & valid_mask;
    const int lx = __shfl_sync(0xffffffffu, row, (31 - __clz(mx)) & 31);
    const int ly = __shfl_sync(0xffffffffu, row, (31 - __clz(my)) & 31);
    if (mx) held_x = lx;
    if (my) held_y = ly;
  }
  // ---- staging: one slot per distinct row that some lane refills from
  const unsigned same = __match_any_sync(0xffffffffu, row);
  const unsigned refills = __ballot_sync(0xffffffffu, refill);
  const int leader = __ffs(same) - 1;
  const bool copy = leader == lane && valid && (same & refills) != 0;
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
  if (lane == 0) ctl.info[buf] = info;
  if (copy) {
    if (bulk) {
      for (uintptr_t q = g_lo; q < a_lo; ++q) s[q - g_lo] = *reinterpret_cast<const uint8_t*>(q);
      for (uintptr_t q = max(a_hi, g_lo); q < g_lo + seg; ++q) s[q - g_lo] = *reinterpret_cast<const uint8_t*>(q);
    } else {
      for (uint32_t q = 0; q < seg; ++q) s[q] = reinterpret_cast<const uint8_t*>(g_lo)[q];
    }
  }
  if (j < nr) {
    const int src = refill ? (int)smem_u32(s) : -1;
    const uint32_t w = valid ? (uint32_t)(tap ? b1 : b0) << 16 : 0u;
    if (reg == 0) {
      plan[j].src_x = src;
      plan[j].w_x = w;
    } else {
      plan[j].src_y = src;
      plan[j].w_y = w;
    }
  }
  __syncwarp();  // plan and lane-copied bytes are ordered before the releasing arrive
  if (lane == 0) mbar_expect_tx(full, tx);
  __syncwarp();
  if (bulk) bulk_g2s(s + (ptrdiff_t)(a_lo - g_lo), reinterpret_cast<const void*>(a_lo), bulk, full);
}

// Horizontal pass of one staged source row at this lane's column: (p[x0]*a0 + p[x1]*a1) >> 4 per
// channel (HResizeLinear + the >> 4 VResizeLinear applies to its int32 rows).  `a` is the
// shared-memory address of the first tap's R byte.
__device__ __forceinline__ void hpass(uint32_t a, uint32_t coef, uint32_t& h0, uint32_t& h1, uint32_t& h2) {
  const uint32_t w = a & ~3u;
  uint32_t w0, w1, w2;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(w));
  asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(w1) : "r"(w));
  asm volatile("ld.shared.u32 %0, [%1+8];" : "=r"(w2) : "r"(w));
  const uint32_t sh = a << 3;  // the funnel shift takes it modulo 32: (a & 3) * 8
  const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
  // a zero weight multiplies whatever byte sits there (always inside the staging buffer) by 0
  h0 = __dp2a_lo(coef, tap_pair(lo, hi, 0), 0u) >> 4;
  h1 = __dp2a_lo(coef, tap_pair(lo, hi, 1), 0u) >> 4;
  h2 = __dp2a_lo(coef, tap_pair(lo, hi, 2), 0u) >> 4;
}

__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// kFixed224: out_size == 224 (MobileNet's): one column per consumer lane, every stride a constant.
template <bool kFixed224>
__global__ void __launch_bounds__(kMaxTeams * kTeamThreads, 1) crop_resize_kernel(const CropParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  float* lut = reinterpret_cast<float*>(smem);  // [3][256][32]
  uint8_t* stage_all = smem + kLutFloats * sizeof(float);
  TeamCtl* ctl_all = reinterpret_cast<TeamCtl*>(stage_all + (size_t)p.teams * kStages * p.stage_bytes);

  const int tid = threadIdx.x, lane = tid & 31;
  const int team = tid / kTeamThreads, ttid = tid - team * kTeamThreads;
  const int S = kFixed224 ? 224 : p.out_size;
  const int total = p.total_dev ? min(*p.total_dev, p.roi_slots) : p.total_host;
  const int items = total * p.bands;
  if (items <= 0) return;  // (the ticket words are only valid when ROI 0 exists)
  const size_t plane = (size_t)S * S;
  TeamCtl& ctl = ctl_all[team];
  uint8_t* stage = stage_all + (size_t)team * kStages * p.stage_bytes;

  // the table: 768 values, each written as one 128-byte row (a copy per bank)
  for (int i = tid; i < 3 * 256; i += blockDim.x) {
    const int c = i >> 8;
    const float u = __fdiv_rn((float)(i & 255), 255.0f);                    // transforms.py:263
    const float mean = c == 0 ? p.mean[0] : (c == 1 ? p.mean[1] : p.mean[2]);
    const float stdv = c == 0 ? p.stdv[0] : (c == 1 ? p.stdv[1] : p.stdv[2]);
    const float f = __fdiv_rn(__fsub_rn(u, mean), stdv);                     // :270
    float4* row = reinterpret_cast<float4*>(lut + (size_t)i * kLutRep);
#pragma unroll
    for (int k = 0; k < kLutRep / 4; ++k) row[k] = make_float4(f, f, f, f);
  }
  if (ttid == 0) {
#pragma unroll
    for (int b = 0; b < kStages; ++b) {
      mbar_init(&ctl.full[b], 1);
      mbar_init(&ctl.empty[b], kTeamConsumers / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (ttid >= kTeamConsumers) {
    // =============================== producer warp ====================================
    const int n_teams = (int)gridDim.x * p.teams;
    auto draw = [&](int current) -> int {  // the next work item, or -1
      int it;
      if (p.tickets != nullptr) {
        it = lane == 0 ? atomicAdd(p.tickets, 1) : 0;
        it = __shfl_sync(0xffffffffu, it, 0);
      } else {
        it = current < 0 ? (int)blockIdx.x * p.teams + team : current + n_teams;
      }
      return it < items ? it : -1;
    };
    Cursor cur;
    int s = 0;
    ItemRecord rec = fetch_item(p, draw(-1));
    while (rec.item >= 0) {
      load_item(cur, p, rec);
      // the next ticket and its ROI record are fetched while this item is staged
      const int next = draw(rec.item);
      double sc_y = 0.0, sc_x = 0.0;
      if (!cur.empty) {  // both scales at once: lane parity picks the axis
        const double sc = axis_scale(S, (lane & 1) ? cur.cw : cur.ch);
        sc_y = __shfl_sync(0xffffffffu, sc, 0);
        sc_x = __shfl_sync(0xffffffffu, sc, 1);
      }
      int held_x = -2, held_y = -2;  // a new item starts with nothing in the consumers' registers
      for (; cur.sb < cur.nsub; ++cur.sb, ++s) {
        const int buf = s % kStages, use = s / kStages;
        // more than 224 columns: a lane walks the step once per column chunk, so nothing carries over
        if (!kFixed224 && S > kTeamConsumers) held_x = held_y = -2;
        produce_step(cur, sc_y, sc_x, stage + (size_t)buf * p.stage_bytes, ctl, buf, use ? ((use - 1) & 1) : -1,
                     held_x, held_y, lane);
        if (cur.sb == 0) rec = fetch_item(p, next);
      }
    }
    {  // the end marker travels like a step
      const int buf = s % kStages, use = s / kStages;
      if (use) mbar_wait(&ctl.empty[buf], (uint32_t)((use - 1) & 1));
      if (lane == 0) {
        ctl.info[buf].item = -1;
        mbar_expect_tx(&ctl.full[buf], 0);
        if (p.tickets != nullptr && atomicAdd(p.tickets + 1, 1) == n_teams - 1) {
          p.tickets[0] = 0;  // every team has drawn its last ticket: leave the words as they were
          p.tickets[1] = 0;
        }
      }
    }
    return;
  }

  // ================================= consumer warps ====================================
  const uint32_t lut_lane = smem_u32(lut) + ((uint32_t)lane << 2);
  const int col0 = ttid;  // this lane's column within a 224-column chunk
  uint32_t hx0 = 0, hx1 = 0, hx2 = 0, hy0 = 0, hy1 = 0, hy2 = 0;  // the two live horizontal passes
  int xo = 0;
  uint32_t coef = 0;
  for (int s = 0;; ++s) {
    const int buf = s % kStages;
    mbar_wait(&ctl.full[buf], (uint32_t)((s / kStages) & 1));
    const int4 ia = *reinterpret_cast<const int4*>(&ctl.info[buf]);  // item roi y0 nr
    if (ia.x < 0) break;
    const int2 ib = *reinterpret_cast<const int2*>(&ctl.info[buf].flags);  // flags cw
    const int nr = ia.w;
    float* const obase = p.out + ((size_t)ia.y * 3) * plane + (size_t)ia.z * S;

    if (ib.x & kStepEmpty) {
      // mobilenet_preprocess.py:266-267: a 1x1 black crop
      const float e0 = lds_f32(lut_lane), e1 = lds_f32(lut_lane + 256 * kLutRep * 4),
                  e2 = lds_f32(lut_lane + 512 * kLutRep * 4);
      for (int x = col0; x < S; x += kTeamConsumers) {
        float* o = obase + x;
        for (int j = 0; j < nr; ++j, o += S) {
          st_stream_f1(o, e0);
          st_stream_f1(o + plane, e1);
          st_stream_f1(o + 2 * plane, e2);
        }
      }
    } else {
      const int chunks = kFixed224 ? 1 : (S + kTeamConsumers - 1) / kTeamConsumers;
      for (int xb = 0; xb < chunks; ++xb) {
        const int x = xb * kTeamConsumers + col0;
        const bool live = kFixed224 || x < S;
        if ((ib.x & kStepFirst) || chunks > 1) {
          // this lane's x coefficients (xofs / alpha entry), kept in registers for the whole item
          const int mode = ib.x >> kStepModeShift;
          xo = 0, coef = 0;
          if (live) {
            if (mode == kResizeCopy) {
              xo = x * 3, coef = kCoefOne;
            } else if (mode == kResizeBox2) {
              xo = x * 6, coef = (kCoefOne / 2) | ((kCoefOne / 2) << 16);
            } else {
              const AxisCoef k = axis_coef(x, ctl.info[buf].scx, ib.y, true);
              xo = k.s0 * 3, coef = (uint32_t)k.w0 | ((uint32_t)k.w1 << 16);
            }
          }
        }
        float* o = obase + x;
        const YPlan* plan = ctl.plan[buf];
        for (int j = 0; j < nr; ++j, o += S) {
          const int4 pl = *reinterpret_cast<const int4*>(&plan[j]);  // src_x src_y w_x w_y, warp-uniform
          if (pl.x >= 0) hpass((uint32_t)(pl.x + xo), coef, hx0, hx1, hx2);
          if (pl.y >= 0) hpass((uint32_t)(pl.y + xo), coef, hy0, hy1, hy2);
          // VResizeLinear: (((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) + 2) >> 2; an unused tap has
          // weight 0 and multiplies whatever its register holds by 0
          const uint32_t t0 = __umulhi((uint32_t)pl.z, hx0) + __umulhi((uint32_t)pl.w, hy0) + 2u;
          const uint32_t t1 = __umulhi((uint32_t)pl.z, hx1) + __umulhi((uint32_t)pl.w, hy1) + 2u;
          const uint32_t t2 = __umulhi((uint32_t)pl.z, hx2) + __umulhi((uint32_t)pl.w, hy2) + 2u;
          if (live) {
            // table row (t >> 2) is 128 bytes; this lane's copy sits at lane * 4 inside it
            st_stream_f1(o, lds_f32(lut_lane + ((t0 & ~3u) << 5)));
            st_stream_f1(o + plane, lds_f32(lut_lane + 256 * kLutRep * 4 + ((t1 & ~3u) << 5)));
            st_stream_f1(o + 2 * plane, lds_f32(lut_lane + 512 * kLutRep * 4 + ((t2 & ~3u) << 5)));
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&ctl.empty[buf]);  // this warp is done with the staging buffer + plan
  }
}

static int launch_crops(const uint8_t* frames, size_t frames_bytes, int src_h, int src_w,
                        const ia_frame_desc* descs, const int* rois, const int* total_dev, int total_host, int roi_slots,
                        int out_size, const float mean[3], const float stdv[3], float* out, int* tickets,
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
  c.bands = (out_size + kItemRows - 1) / kItemRows;
  for (int i = 0; i < 3; ++i) {
    c.mean[i] = mean[i];
    c.stdv[i] = stdv[i];
  }
  c.out = out;
  c.tickets = tickets;
  const long long items = (long long)c.bands * roi_slots;
  IA_RETURN_IF(items > 0x7fffffffLL, IA_ERR_UNSUPPORTED);
  int dev = 0, sms = 148, optin = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return (int)e;
  // shared memory of the one CTA per SM: the table, then per team two staging buffers (each at
  // least two full-width rows) and a control block.  Narrow frames get four teams of ~16 KB
  // buffers; wide ones fewer teams with larger buffers.
  const size_t widest = ((size_t)src_w * 3 + 30 + 15) & ~(size_t)15;
  const size_t fixed = kLutFloats * sizeof(float) + kMaxTeams * sizeof(TeamCtl);
  IA_RETURN_IF((size_t)optin < fixed + 2 * kStages * widest, IA_ERR_UNSUPPORTED);
  const size_t avail = (size_t)optin - fixed;
  int teams = (int)(avail / (2 * kStages * widest));
  if (teams > kMaxTeams) teams = kMaxTeams;
  const size_t stage = (avail / ((size_t)teams * kStages)) & ~(size_t)127;
  IA_RETURN_IF(stage >= (1u << 20), IA_ERR_UNSUPPORTED);
  c.teams = teams;
  c.stage_bytes = (int)stage;
  const size_t smem = kLutFloats * sizeof(float) + (size_t)teams * kStages * stage + (size_t)teams * sizeof(TeamCtl);
  void (*kernel)(CropParams) = out_size == 224 ? crop_resize_kernel<true> : crop_resize_kernel<false>;
  e = allow_max_dynamic_smem(reinterpret_cast<const void*>(kernel));
  if (e != cudaSuccess) return (int)e;
  const long long want = (items + teams - 1) / teams;
  const long long grid = want < sms ? want : sms;
  kernel<<<(unsigned)grid, teams * kTeamThreads, smem, stream>>>(c);
  return launch_status();
}

}  // namespace ia
