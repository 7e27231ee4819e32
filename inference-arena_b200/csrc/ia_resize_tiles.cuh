// ia_resize_tiles.cuh — the separable 8-bit bilinear resize + table + CHW kernel behind
//   K4 stage B (ROI crop -> 224x224 -> ImageNet normalise, mobilenet_preprocess.py:114-160) and
//   K1's general linear mode (letterbox -> /255, yolo_preprocess.py:128-174, when the resize is
//   neither a copy, an exact 2x box nor a pure gather).
// Warp-specialised, multi-buffered, persistent; separable like cv::resize itself.
//
// Work is cut into TILES of the output: up to 28 rows x up to 224 columns of one resized image
// (an ROI's 224x224 output is 8 tiles; a letterboxed 640x480 area is 18 x 3), plus FILL tiles
// for what is constant (letterbox padding, empty ROIs).  One CTA per SM holds up to four
// independent TEAMS of 7 consumer warps + 1 producer warp and one shared table of the output
// values — ((v/255 - mean)/std) or v/divisor — replicated once per lane (bank) so that a warp's
// 32 look-ups of 32 unrelated pixel values are one shared-memory wavefront.  A team takes tiles
// one at a time — from a ticket counter in device memory when the caller has one, else from a
// counter the teams of a CTA share over the CTA's own tiles: tiles of very different cost still
// finish together — and walks a tile in steps, coupled to its producer only by mbarriers
// (full / empty per staging buffer):
//   producer: plans a whole tile at once, one lane per output row: OpenCV's yofs/beta entry,
//           which of the consumers' two registers each source row goes to (its parity, so the
//           taps of one output row never collide and every source row is staged and filled once
//           per register), and the copy of each row (16-byte-aligned superset of the needed
//           segment, clamped to the frame's allocation; bytes a clamp cuts off are copied by the
//           lane).  A step is the longest run of rows whose fills fit a staging buffer (counted
//           with ballots).  Per step: wait for the consumers to release the buffer, post the
//           lanes' expect_tx, issue one 1-D bulk async copy per row (TMA engine, complete_tx on
//           full[buf]), publish a 16-byte plan per output row and a 48-byte step record, arrive.
//           The next ticket and its ROI record are fetched while the current tile is staged.
//   consumers: a lane owns ONE output column for the whole tile and walks down the rows; a
//           warp is 32 neighbouring columns, so every plane store is one 128-byte line.  The
//           lane's x coefficients live in registers.  The horizontal pass of a source row —
//           three aligned 32-bit shared loads + a funnel shift for the 6 bytes of the two taps,
//           a PRMT pairing (p[c], p[c+3]) and one IDP.2A against the packed coefficient pair,
//           >> 4 — is evaluated once per source row and kept in registers for as long as output
//           rows use it (two rows are live: an up-scaled image reuses each ~2.5 times), which is
//           the order cv::resize works in.  The vertical step is two IMAD.HI against
//           coefficients pre-shifted by 16 — OpenCV's 11-bit fixed point, bit-identical to
//           cv2.resize INTER_LINEAR (copy and the exact-2x box average use the same arithmetic
//           with tables a0=b0=2048 resp. a0=a1=b0=b1=1024, which reduce to the identity and to
//           (p00+p01+p10+p11+2)>>2).  Consumers never read global memory.
// The global side sees only wide, fully used requests; byte gathers hit shared memory.
#pragma once

#include "ia_common.cuh"

namespace ia {

// 7 consumer warps (224 = 7 x 32 columns: an ROI's whole width) + 1 producer warp.
constexpr int kTeamConsumers = 224;
constexpr int kTeamThreads = kTeamConsumers + 32;
constexpr int kMaxTeams = 4;
constexpr int kTileRows = 28;  // most output rows of one tile: one producer lane per row
constexpr int kStages = 2;     // staging buffers per team
constexpr int kLutRep = 32;    // one copy of the 3x256 table per bank
constexpr int kLutFloats = 3 * 256 * kLutRep;

struct TileParams {
  // ---- source frames
  const uint8_t* frames;
  size_t frames_bytes;  // n * src_h * src_w * 3: staging never reads outside [frames, frames+bytes)
  int src_h, src_w;
  const ia_frame_desc* descs;  // ragged batches: per-image pointer/height/width (device) or null
  // ---- what is resized: rectangles from a table (K4) or, when rois is null, whole frames (K1)
  const int* rois;
  const int* total_dev;  // number of valid images / ROIs (device) or null
  int total_host;        // used when total_dev is null
  int slots;             // images / ROIs the output buffer can hold
  // ---- the resized image (out_w x out_h) and the canvas it sits in
  int out_w, out_h;
  int canvas, canvas_x0, canvas_y0;  // canvas edge (S or T), position of the resized image in it
  int tiles_x, tiles_y;              // image tiles per image
  int tile_rows;                     // output rows per tile (<= kTileRows)
  int fill_top, fill_bottom;         // fill tiles (bands of tile_rows rows) above / below
  int fill_left, fill_right;         // 1 when there is padding left / right of the image
  int tiles_per_image;
  uint32_t fill_rgb;  // table indices of the fill colour, r | g << 8 | b << 16
  // ---- execution
  int teams;        // teams per CTA (blockDim.x = teams * kTeamThreads)
  int stage_bytes;  // bytes of ONE staging buffer (each team has kStages)
  // Work distribution.  With `tickets` (two zeroed ints: next tile, teams finished) a team draws
  // its tiles one at a time; the last team to finish zeroes both again.  Without, CTAs stride over
  // the tiles and the teams of a CTA share a counter in shared memory.
  int* tickets;
  // ---- whole frames (rois == null): one resize geometry for the launch, so OpenCV's scales come
  // from the host and the x coefficients of every column are tabulated once per CTA
  double sc_x, sc_y;
  int xtab_bytes;  // out_w * 8 (16-byte multiple) or 0
  // ---- output values: ((v/255) - mean) / std per channel, or v / divisor when divisor != 0
  float divisor;
  float mean[3], stdv[3];
  float* out;
};

// One output row of a step.  A consumer lane keeps the horizontal passes of two source rows in
// registers X and Y (a source row lives in the register of its parity, so the two taps of an
// output row never collide); the producer, which knows the whole row sequence, tells it
// which of them to refill — the shared-memory address of the staged segment's first byte, or
// -1 to keep what is there — and the vertical weight of each, shifted left by 16 for IMAD.HI
// (0 for a tap that is not used).
struct __align__(16) YPlan {
  int src_x, src_y;
  uint32_t w_x, w_y;
};

enum : int { kStepFirst = 1, kStepFill = 2, kStepModeShift = 2 };

// What the consumers need to know about a step; they never read global memory themselves.
struct __align__(16) StepInfo {
  int tile;     // -1: no more work
  int nr;       // rows of the step
  int flags;    // kStepFirst: first step of a tile; kStepFill: constant tile; mode << kStepModeShift
  int cols;     // columns of the tile
  long long out_off;  // element offset of (first row of the step, first column of the tile), plane 0
  double scx;   // OpenCV scale_x of the resize
  int x0;       // first column of the tile in the resized image
  int seg_x0;   // first source column of the staged segments
  int cw;       // source width
  uint32_t fill_rgb;
};

// Per-team control block in shared memory.
struct __align__(16) TeamCtl {
  YPlan plan[kStages][kTileRows];
  StepInfo info[kStages];
  uint64_t full[kStages], empty[kStages];
};

// The ROI record of a tile (and its frame, for ragged batches), read ahead of its use.
struct TileRecord {
  int tile;
  int4 a, b;  // image x1 y1 x2 | y2 empty - -
  ia_frame_desc d;
};

__device__ __forceinline__ TileRecord fetch_tile(const TileParams& p, int tile) {
  TileRecord r;
  r.tile = tile;
  r.a = make_int4(0, 0, 0, 0);
  r.b = make_int4(0, 1, 0, 0);
  r.d.data_dev = nullptr;
  r.d.h = r.d.w = 0;
  if (tile >= 0 && p.rois != nullptr) {
    const int4* roi = reinterpret_cast<const int4*>(p.rois + (size_t)(tile / p.tiles_per_image) * 8);
    r.a = roi[0];
    r.b = roi[1];
    if (p.descs != nullptr) r.d = p.descs[r.a.x];
  }
  return r;
}

// A tile as the producer sees it.
struct Cursor {
  int tile;
  int y_first, rows;   // first output row (in the resized image) and rows of the tile
  int x0, cols;        // first output column (in the resized image) and columns
  int fill, cw, ch, mode;
  uint32_t fill_rgb;
  long long out_off;   // element offset of the tile's first output element (plane 0)
  int seg_x0, seg_cols;  // source columns staged per row
  int slot_bytes;
  const uint8_t* crop;     // first byte of the source rectangle in the frame
  uint32_t row_bytes;      // pitch of the frame it lives in
  uintptr_t buf_lo, buf_hi;  // that frame's allocation: staging never reads outside it
};

// Tile index -> geometry.  Tiles of one image: tiles_y x tiles_x image tiles, then the fill
// bands above, below, and the fill columns left and right of every image band.
__device__ __forceinline__ void load_tile(Cursor& c, const TileParams& p, const TileRecord& rec) {
  c.tile = rec.tile;
  const int image = rec.tile / p.tiles_per_image;
  int local = rec.tile - image * p.tiles_per_image;
  const int T = p.canvas;
  const long long image_off = (long long)image * 3 * T * T;
  const int n_image_tiles = p.tiles_x * p.tiles_y;
  c.fill = 0;
  c.fill_rgb = p.fill_rgb;
  c.mode = kResizeCopy;
  c.cw = c.ch = 1;
  c.seg_x0 = 0;
  c.seg_cols = 1;
  c.crop = nullptr;
  c.row_bytes = 0;
  c.buf_lo = c.buf_hi = 0;
  int cy, cx;  // canvas position of the tile
  if (local < n_image_tiles) {
    const int ty = local / p.tiles_x, tx = local - ty * p.tiles_x;
    c.y_first = ty * p.tile_rows;
    c.rows = min(p.tile_rows, p.out_h - c.y_first);
    c.x0 = tx * kTeamConsumers;
    c.cols = min(kTeamConsumers, p.out_w - c.x0);
    cy = p.canvas_y0 + c.y_first;
    cx = p.canvas_x0 + c.x0;
    if (p.rois != nullptr) {  // a rectangle of a frame
      c.fill = rec.b.y;       // empty ROI: a 1x1 black crop (mobilenet_preprocess.py:266-267)
      c.fill_rgb = 0;
      c.cw = c.fill ? 1 : rec.a.w - rec.a.y;
      c.ch = c.fill ? 1 : rec.b.x - rec.a.z;
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
    } else {  // the whole frame `image`
      c.cw = p.src_w;
      c.ch = p.src_h;
      c.row_bytes = (uint32_t)p.src_w * 3u;
      c.buf_lo = (uintptr_t)p.frames;
      c.buf_hi = c.buf_lo + p.frames_bytes;
      c.crop = p.frames + (size_t)image * p.src_h * c.row_bytes;
    }
    c.mode = resize_mode(c.ch, c.cw, p.out_h, p.out_w);
  } else {
    c.fill = 1;
    local -= n_image_tiles;
    c.y_first = 0;
    c.x0 = 0;
    if (local < p.fill_top + p.fill_bottom) {  // full-width bands above / below the image
      const bool below = local >= p.fill_top;
      const int k = below ? local - p.fill_top : local;
      const int lo = below ? p.canvas_y0 + p.out_h : 0, hi = below ? T : p.canvas_y0;
      cy = lo + k * p.tile_rows;
      c.rows = min(p.tile_rows, hi - cy);
      cx = 0;
      c.cols = T;
    } else {  // columns left / right of one band of image rows
      local -= p.fill_top + p.fill_bottom;
      const bool right = p.fill_left ? local >= p.tiles_y : true;
      const int ty = (p.fill_left && right) ? local - p.tiles_y : local;
      cy = p.canvas_y0 + ty * p.tile_rows;
      c.rows = min(p.tile_rows, p.out_h - ty * p.tile_rows);
      cx = right ? p.canvas_x0 + p.out_w : 0;
      c.cols = right ? T - cx : p.canvas_x0;
    }
  }
  c.out_off = image_off + (long long)cy * T + cx;
  c.slot_bytes = 16;
}

// The source columns a tile's horizontal taps touch.  first/last: xofs of the tile's first and
// last column.
__device__ __forceinline__ void size_tile(Cursor& c, int s0_first, int s0_last) {
  c.seg_x0 = s0_first;
  c.seg_cols = min(c.cw - 1, s0_last + 1) - s0_first + 1;
  c.slot_bytes = (c.seg_cols * 3 + 30 + 15) & ~15;  // aligned superset of one row segment
}

// The vertical plan of a whole tile, one producer lane per output row: the source row each of
// the consumers' two registers must hold during that output row (-1: not used), its weight, and
// whether it has to be (re)filled there.  A source row lives in the register of its parity, so
// the two taps of an output row never collide and — the rows only move forward — every source
// row is staged and filled once per register; a second tap clamped onto the first tap's row
// (bottom border) and an unused tap take the other register.
struct LanePlan {
  int row_x, row_y;
  uint32_t w_x, w_y;  // << 16; 0 for an unused tap
  bool fill_x, fill_y;
};

__device__ __forceinline__ LanePlan plan_tile(const Cursor& c, double sc_y, int lane, unsigned& fills_x,
                                              unsigned& fills_y) {
  int r0 = -1, r1 = -1, b0 = kCoefOne, b1 = 0;
  if (lane < c.rows) {
    const int y = c.y_first + lane;
    if (c.mode == kResizeCopy) {
      r0 = y;
    } else if (c.mode == kResizeBox2) {
      r0 = 2 * y;
      r1 = r0 + 1;
      b0 = b1 = kCoefOne / 2;
    } else {
      const AxisCoef k = axis_coef(y, sc_y, c.ch, false);
      b0 = k.w0;
      b1 = k.w1;
      r0 = clampi(k.s0, 0, c.ch - 1);
      if (k.w1 != 0) r1 = clampi(k.s0 + 1, 0, c.ch - 1);
    }
  }
  if (r1 < 0) b1 = 0;
  const bool swap = (r0 & 1) != 0;  // tap 0 goes to the register of its row's parity
  LanePlan l;
  l.row_x = swap ? r1 : r0;
  l.row_y = swap ? r0 : r1;
  l.w_x = (uint32_t)(swap ? b1 : b0) << 16;
  l.w_y = (uint32_t)(swap ? b0 : b1) << 16;
  if (lane >= c.rows) l.row_x = l.row_y = -1, l.w_x = l.w_y = 0;
  // a register is filled when the closest earlier row that used it held another source row
  const unsigned lt = (1u << lane) - 1u;
  const unsigned used_x = __ballot_sync(0xffffffffu, l.row_x >= 0) & lt;
  const unsigned used_y = __ballot_sync(0xffffffffu, l.row_y >= 0) & lt;
  const int prev_x = __shfl_sync(0xffffffffu, l.row_x, (31 - __clz(used_x)) & 31);
  const int prev_y = __shfl_sync(0xffffffffu, l.row_y, (31 - __clz(used_y)) & 31);
  l.fill_x = l.row_x >= 0 && (used_x == 0 || prev_x != l.row_x);
  l.fill_y = l.row_y >= 0 && (used_y == 0 || prev_y != l.row_y);
  fills_x = __ballot_sync(0xffffffffu, l.fill_x);
  fills_y = __ballot_sync(0xffffffffu, l.fill_y);
  return l;
}

// Where the needed segment of one source row comes from: one bulk async copy of the
// 16-byte-aligned superset, clamped to the frame's allocation (bytes a clamp cuts off are copied
// by the lane).  A source row is staged once per tile and register, so this is worked out once
// per tile; a step only adds the slot it lands in.
struct RowCopy {
  uintptr_t g_lo;  // first needed byte in global memory
  uintptr_t a_lo;  // start of the bulk copy (16-byte aligned)
  uint32_t bulk;   // its length (16-byte multiple; 0: everything by the lane)
};

__device__ __forceinline__ RowCopy plan_copy(const Cursor& c, int row) {
  RowCopy r;
  r.g_lo = (uintptr_t)(c.crop + (size_t)max(row, 0) * c.row_bytes + (size_t)c.seg_x0 * 3);
  const uint32_t seg = (uint32_t)c.seg_cols * 3u;
  r.a_lo = max(r.g_lo & ~(uintptr_t)15, (c.buf_lo + 15) & ~(uintptr_t)15);
  const uintptr_t a_hi = min((r.g_lo + seg + 15) & ~(uintptr_t)15, c.buf_hi & ~(uintptr_t)15);
  r.bulk = (row >= 0 && a_hi > r.a_lo) ? (uint32_t)(a_hi - r.a_lo) : 0u;
  return r;
}

__device__ __forceinline__ void sts_u8(uint32_t addr, uint8_t v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"((uint32_t)v) : "memory");
}

// dst: shared-memory address where byte g_lo lands.
__device__ __forceinline__ void issue_copy(const RowCopy& r, uint32_t seg, uint32_t dst, uint64_t* full) {
  const uintptr_t a_hi = r.a_lo + r.bulk;
  if (r.bulk == 0 || r.a_lo > r.g_lo || a_hi < r.g_lo + seg) {
    // first / last bytes of the frame's allocation: the lane fetches what the clamp cut off
    for (uintptr_t q = r.g_lo; q < r.g_lo + seg; ++q)
      if (r.bulk == 0 || q < r.a_lo || q >= a_hi) sts_u8(dst + (uint32_t)(q - r.g_lo), *reinterpret_cast<const uint8_t*>(q));
  }
  if (r.bulk)
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     dst + (uint32_t)(r.a_lo - r.g_lo)),
                 "l"(r.a_lo), "r"(r.bulk), "r"(smem_u32(full))
                 : "memory");
}

// Producer warp, one step: rows [j0, j1) of the tile.  Publishes the step record and the row
// plans, stages the rows those output rows fill their registers from, arrives on `full`.
// Everything that does not touch shared memory is worked out BEFORE waiting for the consumers
// to release the buffer: between their release and the copies' issue lie only the stores of
// the plan and the copy instructions themselves.
__device__ __forceinline__ void produce_step(const Cursor& c, const LanePlan& l, const RowCopy& cx, const RowCopy& cy,
                                             unsigned fills_x, unsigned fills_y, int j0, int j1, int canvas,
                                             double sc_x, uint32_t stage_addr, TeamCtl& ctl, int buf,
                                             int empty_parity, int lane) {
  uint64_t* full = &ctl.full[buf];
  StepInfo info;
  info.tile = c.tile;
  info.nr = j1 - j0;
  info.flags = (j0 == 0 ? kStepFirst : 0) | (c.fill ? kStepFill : 0) | (c.mode << kStepModeShift);
  info.cols = c.cols;
  info.out_off = c.out_off + (long long)j0 * canvas;
  info.scx = sc_x;
  info.x0 = c.x0;
  info.seg_x0 = c.seg_x0;
  info.cw = c.cw;
  info.fill_rgb = c.fill_rgb;
  // slots of this lane's rows: fills of earlier rows of the step come first, X before Y
  const unsigned step_lt = ((1u << lane) - 1u) & ~((1u << j0) - 1u);
  const int slot_x = __popc(fills_x & step_lt) + __popc(fills_y & step_lt);
  const bool mine = lane >= j0 && lane < j1 && !c.fill;
  const bool do_x = mine && l.fill_x, do_y = mine && l.fill_y;
  const uint32_t dst_x = stage_addr + (uint32_t)(slot_x * c.slot_bytes) + (uint32_t)(cx.g_lo & 15);
  const uint32_t dst_y = stage_addr + (uint32_t)((slot_x + (l.fill_x ? 1 : 0)) * c.slot_bytes) + (uint32_t)(cy.g_lo & 15);
  const uint32_t seg = (uint32_t)c.seg_cols * 3u;
  YPlan yp;
  yp.src_x = do_x ? (int)dst_x : -1;  // -1: keep
  yp.src_y = do_y ? (int)dst_y : -1;
  yp.w_x = l.w_x;
  yp.w_y = l.w_y;
  const uint32_t tx = (do_x ? cx.bulk : 0u) + (do_y ? cy.bulk : 0u);

  if (empty_parity >= 0) mbar_wait(&ctl.empty[buf], (uint32_t)empty_parity);  // consumers released the buffer
  if (tx) mbar_add_tx(full, tx);
  if (do_x) issue_copy(cx, seg, dst_x, full);
  if (do_y) issue_copy(cy, seg, dst_y, full);
  if (mine) ctl.plan[buf][lane - j0] = yp;
  if (lane == 0) ctl.info[buf] = info;
  __syncwarp();  // plans, lane-copied bytes and every lane's expect_tx are ordered before the arrive
  if (lane == 0) mbar_arrive(full);
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

// xofs of output column x (HResize's first tap) for the three modes.
__device__ __forceinline__ AxisCoef x_entry(int mode, int x, double sc_x, int cw) {
  AxisCoef k;
  if (mode == kResizeCopy) {
    k.s0 = x, k.w0 = kCoefOne, k.w1 = 0;
  } else if (mode == kResizeBox2) {
    k.s0 = 2 * x, k.w0 = k.w1 = kCoefOne / 2;
  } else {
    k = axis_coef(x, sc_x, cw, true);
  }
  return k;
}

// kCanvas: canvas edge known at compile time (224: MobileNet's, 640: YOLO's; 0: from the
// parameters), which makes every output stride a constant.
template <int kCanvas>
__global__ void __launch_bounds__(kMaxTeams * kTeamThreads, 1) resize_tiles_kernel(const TileParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int cta_ticket;  // next of this CTA's tiles (used when there are no device-memory tickets)
  float* lut = reinterpret_cast<float*>(smem);  // [3][256][32]
  uint8_t* stage_all = smem + kLutFloats * sizeof(float);
  TeamCtl* ctl_all = reinterpret_cast<TeamCtl*>(stage_all + (size_t)p.teams * kStages * p.stage_bytes);
  int2* xtab = reinterpret_cast<int2*>(ctl_all + p.teams);  // [out_w] (xofs, a0 | a1 << 16) when xtab_bytes

  const int tid = threadIdx.x, lane = tid & 31;
  // Warps 0 .. 7*teams-1 are the consumers (seven per team), the last `teams` warps the producers:
  // warp w sits on scheduler w % 4, so every scheduler gets one producer next to seven consumers
  // instead of one scheduler getting all four.
  const int warp = tid >> 5;
  const bool producer = warp >= 7 * p.teams;
  const int team = producer ? warp - 7 * p.teams : warp / 7;
  const int ttid = producer ? kTeamConsumers + lane : tid - team * kTeamConsumers;
  const int T = kCanvas ? kCanvas : p.canvas;
  const int total = p.total_dev ? min(*p.total_dev, p.slots) : p.total_host;
  const long long tiles_ll = (long long)total * p.tiles_per_image;
  const int tiles = (int)min(tiles_ll, (long long)0x7fffffff);
  if (tiles <= 0) return;  // (the ticket words are only valid when ROI 0 exists)
  const size_t plane = (size_t)T * T;
  TeamCtl& ctl = ctl_all[team];
  uint8_t* stage = stage_all + (size_t)team * kStages * p.stage_bytes;

  // the table: 768 values, each written as one 128-byte row (a copy per bank)
  for (int i = tid; i < 3 * 256; i += blockDim.x) {
    const int c = i >> 8;
    float f;
    if (p.divisor != 0.f) {
      f = __fdiv_rn((float)(i & 255), p.divisor);  // yolo_preprocess.py:158
    } else {
      const float u = __fdiv_rn((float)(i & 255), 255.0f);  // transforms.py:263
      const float mean = c == 0 ? p.mean[0] : (c == 1 ? p.mean[1] : p.mean[2]);
      const float stdv = c == 0 ? p.stdv[0] : (c == 1 ? p.stdv[1] : p.stdv[2]);
      f = __fdiv_rn(__fsub_rn(u, mean), stdv);  // :270
    }
    float4* row = reinterpret_cast<float4*>(lut + (size_t)i * kLutRep);
#pragma unroll
    for (int k = 0; k < kLutRep / 4; ++k) row[k] = make_float4(f, f, f, f);
  }
  if (p.xtab_bytes) {
    const int mode = resize_mode(p.src_h, p.src_w, p.out_h, p.out_w);
    for (int x = tid; x < p.out_w; x += blockDim.x) {
      const AxisCoef k = x_entry(mode, x, p.sc_x, p.src_w);
      xtab[x] = make_int2(k.s0, k.w0 | (k.w1 << 16));
    }
  }
  if (tid == 0) cta_ticket = 0;
  if (producer && lane == 0) {
#pragma unroll
    for (int b = 0; b < kStages; ++b) {
      mbar_init(&ctl.full[b], 1);
      mbar_init(&ctl.empty[b], kTeamConsumers / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (producer) {
    // =============================== producer warp ====================================
    const int n_teams = (int)gridDim.x * p.teams;
    auto draw = [&](int current) -> int {  // the next tile, or -1
      int it;
      if (p.tickets != nullptr) {
        it = lane == 0 ? atomicAdd(p.tickets, 1) : 0;
        it = __shfl_sync(0xffffffffu, it, 0);
      } else {
        // no counter in device memory: the CTA owns tiles blockIdx.x, blockIdx.x + gridDim.x, ... and
        // its teams draw from those through a counter in shared memory (a team that gets cheap
        // tiles simply takes more of them)
        it = lane == 0 ? atomicAdd(&cta_ticket, 1) : 0;
        it = __shfl_sync(0xffffffffu, it, 0);
        it = (long long)it * gridDim.x + blockIdx.x < 0x7fffffffLL ? it * (int)gridDim.x + (int)blockIdx.x : -1;
      }
      return (it >= 0 && it < tiles) ? it : -1;
    };
    Cursor cur;
    int s = 0;
    TileRecord rec = fetch_tile(p, draw(-1));
    while (rec.tile >= 0) {
      load_tile(cur, p, rec);
      // the next ticket and its ROI record are fetched while this tile is staged
      const int next = draw(rec.tile);
      double sc_x = 0.0;
      LanePlan l;
      l.row_x = l.row_y = -1, l.w_x = l.w_y = 0, l.fill_x = l.fill_y = false;
      unsigned fills_x = 0, fills_y = 0;
      if (!cur.fill) {
        // both scales at once (lane parity picks the axis), then the tile's first and last xofs
        double sc_y;
        if (p.xtab_bytes) {
          sc_y = p.sc_y;
          sc_x = p.sc_x;
        } else {
          const double sc = axis_scale((lane & 1) ? p.out_w : p.out_h, (lane & 1) ? cur.cw : cur.ch);
          sc_y = __shfl_sync(0xffffffffu, sc, 0);
          sc_x = __shfl_sync(0xffffffffu, sc, 1);
        }
        if (p.tiles_x == 1) {  // the tile spans the whole width: stage whole source rows
          size_tile(cur, 0, cur.cw - 1);
        } else if (p.xtab_bytes) {
          size_tile(cur, xtab[cur.x0].x, xtab[cur.x0 + cur.cols - 1].x);
        } else {
          const int s0 = x_entry(cur.mode, (lane & 1) ? cur.x0 + cur.cols - 1 : cur.x0, sc_x, cur.cw).s0;
          size_tile(cur, __shfl_sync(0xffffffffu, s0, 0), __shfl_sync(0xffffffffu, s0, 1));
        }
        l = plan_tile(cur, sc_y, lane, fills_x, fills_y);
      }
      RowCopy cx, cy;
      cx.g_lo = cx.a_lo = cy.g_lo = cy.a_lo = 0, cx.bulk = cy.bulk = 0;
      if (!cur.fill) {
        cx = plan_copy(cur, l.fill_x ? l.row_x : -1);
        cy = plan_copy(cur, l.fill_y ? l.row_y : -1);
      }
      const int max_slots = p.stage_bytes / cur.slot_bytes;
      const unsigned le = (2u << lane) - 1u;  // lanes 0 .. lane
      for (int j0 = 0; j0 < cur.rows; ++s) {
        // the longest run of rows whose fills fit one staging buffer (a row needs at most two
        // slots): every lane counts the fills of rows j0 .. lane, the first lane over the limit ends it
        const unsigned from = ~((1u << j0) - 1u);
        const bool fits = __popc(fills_x & from & le) + __popc(fills_y & from & le) <= max_slots;
        const unsigned over = ~__ballot_sync(0xffffffffu, fits) & from;
        const int j1 = max(j0 + 1, min(cur.rows, over ? __ffs(over) - 1 : 32));
        const int buf = s % kStages, use = s / kStages;
        produce_step(cur, l, cx, cy, fills_x, fills_y, j0, j1, T, sc_x, smem_u32(stage + (size_t)buf * p.stage_bytes),
                     ctl, buf, use ? ((use - 1) & 1) : -1, lane);
        if (j0 == 0) rec = fetch_tile(p, next);
        j0 = j1;
      }
    }
    {  // the end marker travels like a step
      const int buf = s % kStages, use = s / kStages;
      if (use) mbar_wait(&ctl.empty[buf], (uint32_t)((use - 1) & 1));
      if (lane == 0) {
        ctl.info[buf].tile = -1;
        mbar_arrive(&ctl.full[buf]);
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
  const int col = ttid;  // this lane's column within the tile
  uint32_t hx0 = 0, hx1 = 0, hx2 = 0, hy0 = 0, hy1 = 0, hy2 = 0;  // the two live horizontal passes
  int xo = 0;
  uint32_t coef = 0;
  for (int s = 0;; ++s) {
    const int buf = s % kStages;
    mbar_wait(&ctl.full[buf], (uint32_t)((s / kStages) & 1));
    const StepInfo& info = ctl.info[buf];
    const int4 ia = *reinterpret_cast<const int4*>(&info);  // tile nr flags cols
    if (ia.x < 0) break;
    const int nr = ia.y;
    float* const obase = p.out + info.out_off;

    if (ia.z & kStepFill) {
      const uint32_t rgb = info.fill_rgb;
      const float e0 = lds_f32(lut_lane + (rgb & 0xff) * (kLutRep * 4)),
                  e1 = lds_f32(lut_lane + (256 + ((rgb >> 8) & 0xff)) * (kLutRep * 4)),
                  e2 = lds_f32(lut_lane + (512 + ((rgb >> 16) & 0xff)) * (kLutRep * 4));
      if (((ia.w | T) & 3) == 0 && (reinterpret_cast<uintptr_t>(obase) & 15) == 0) {
        // whole rows of padding: 128-bit stores, a warp writes 512 contiguous bytes per plane
        for (int x = col * 4; x < ia.w; x += kTeamConsumers * 4) {
          float* o = obase + x;
          for (int j = 0; j < nr; ++j, o += T) {
            st_stream_f4(o, e0, e0, e0, e0);
            st_stream_f4(o + plane, e1, e1, e1, e1);
            st_stream_f4(o + 2 * plane, e2, e2, e2, e2);
          }
        }
      } else {
        for (int x = col; x < ia.w; x += kTeamConsumers) {
          float* o = obase + x;
          for (int j = 0; j < nr; ++j, o += T) {
            st_stream_f1(o, e0);
            st_stream_f1(o + plane, e1);
            st_stream_f1(o + 2 * plane, e2);
          }
        }
      }
    } else {
      const bool live = col < ia.w;
      if (ia.z & kStepFirst) {
        // this lane's x coefficients (xofs / alpha entry), kept in registers for the whole tile
        xo = 0, coef = 0;
        if (live && p.xtab_bytes) {
          const int2 e = xtab[info.x0 + col];
          xo = (e.x - info.seg_x0) * 3;
          coef = (uint32_t)e.y;
        } else if (live) {
          const AxisCoef k = x_entry(ia.z >> kStepModeShift, info.x0 + col, info.scx, info.cw);
          xo = (k.s0 - info.seg_x0) * 3;
          coef = (uint32_t)k.w0 | ((uint32_t)k.w1 << 16);
        }
      }
      float* o = obase + col;
      const YPlan* plan = ctl.plan[buf];
      for (int j = 0; j < nr; ++j, o += T) {
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
    __syncwarp();
    if (lane == 0) mbar_arrive(&ctl.empty[buf]);  // this warp is done with the staging buffer + plan
  }
}

// Teams per CTA, staging size and launch.  `widest`: the most source columns a staged row
// segment can have.
static int launch_tiles(TileParams& c, int widest_cols, cudaStream_t stream) {
  const long long tiles = (long long)c.tiles_per_image * c.slots;
  IA_RETURN_IF(tiles > 0x7fffffffLL, IA_ERR_UNSUPPORTED);
  int dev = 0, sms = 148, optin = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return (int)e;
  optin -= 128;  // the kernel's few bytes of static shared memory count against the same limit
  // shared memory of the one CTA per SM: the table, then per team kStages staging buffers (each
  // at least two of the widest row segments) and a control block.  Narrow sources get four teams
  // of ~16 KB buffers; wide ones fewer teams with larger buffers.
  const size_t widest = ((size_t)widest_cols * 3 + 30 + 15) & ~(size_t)15;
  const size_t fixed = kLutFloats * sizeof(float) + kMaxTeams * sizeof(TeamCtl) + (size_t)c.xtab_bytes;
  IA_RETURN_IF((size_t)optin < fixed + 2 * kStages * widest, IA_ERR_UNSUPPORTED);
  const size_t avail = (size_t)optin - fixed;
  int teams = (int)(avail / (2 * kStages * widest));
  if (teams > kMaxTeams) teams = kMaxTeams;
  const size_t stage = (avail / ((size_t)teams * kStages)) & ~(size_t)127;
  IA_RETURN_IF(stage >= (1u << 20), IA_ERR_UNSUPPORTED);
  c.teams = teams;
  c.stage_bytes = (int)stage;
  const size_t smem = kLutFloats * sizeof(float) + (size_t)teams * kStages * stage + (size_t)teams * sizeof(TeamCtl) +
                      (size_t)c.xtab_bytes;
  void (*kernel)(TileParams) = c.canvas == 224   ? resize_tiles_kernel<224>
                               : c.canvas == 640 ? resize_tiles_kernel<640>
                                                 : resize_tiles_kernel<0>;
  e = allow_max_dynamic_smem(reinterpret_cast<const void*>(kernel));
  if (e != cudaSuccess) return (int)e;
  const long long want = (tiles + teams - 1) / teams;
  const long long grid = want < sms ? want : sms;
  kernel<<<(unsigned)grid, teams * kTeamThreads, smem, stream>>>(c);
  return launch_status();
}

// K4 stage B: every ROI of the table -> out_size x out_size, ImageNet statistics.
static int launch_crops(const uint8_t* frames, size_t frames_bytes, int src_h, int src_w,
                        const ia_frame_desc* descs, const int* rois, const int* total_dev, int total_host, int roi_slots,
                        int out_size, const float mean[3], const float stdv[3], float* out, int* tickets,
                        cudaStream_t stream) {
  if (roi_slots == 0) return IA_OK;
  TileParams c;
  c.frames = frames;
  c.frames_bytes = frames_bytes;
  c.src_h = src_h;
  c.src_w = src_w;
  c.descs = descs;
  c.rois = rois;
  c.total_dev = total_dev;
  c.total_host = total_host;
  c.slots = roi_slots;
  c.out_w = c.out_h = c.canvas = out_size;
  c.canvas_x0 = c.canvas_y0 = 0;
  c.tiles_x = (out_size + kTeamConsumers - 1) / kTeamConsumers;
  // a handful of ROIs (a single request) is latency-bound: shorter tiles spread it over more teams
  c.tile_rows = kTileRows;
  while (c.tile_rows > 7 && (long long)roi_slots * c.tiles_x * ((out_size + c.tile_rows - 1) / c.tile_rows) < 2LL * kMaxTeams * 148)
    c.tile_rows /= 2;
  c.tiles_y = (out_size + c.tile_rows - 1) / c.tile_rows;
  c.fill_top = c.fill_bottom = c.fill_left = c.fill_right = 0;
  c.tiles_per_image = c.tiles_x * c.tiles_y;
  c.fill_rgb = 0;
  c.sc_x = c.sc_y = 0.0;
  c.xtab_bytes = 0;
  c.divisor = 0.f;
  for (int i = 0; i < 3; ++i) {
    c.mean[i] = mean[i];
    c.stdv[i] = stdv[i];
  }
  c.out = out;
  c.tickets = tickets;
  return launch_tiles(c, src_w, stream);
}

// K1, general linear mode: n equally sized frames -> letterboxed target x target, v / divisor.
int launch_letterbox_tiles(const uint8_t* frames, int n, int src_h, int src_w, int target, int new_h, int new_w,
                           int pad_h, int pad_w, float divisor, uint32_t pad_rgb, float* out, cudaStream_t stream) {
  if (n == 0) return IA_OK;
  IA_RETURN_IF(divisor == 0.f, IA_ERR_INVALID_ARGUMENT);
  TileParams c;
  c.frames = frames;
  c.frames_bytes = (size_t)n * src_h * src_w * 3;
  c.src_h = src_h;
  c.src_w = src_w;
  c.descs = nullptr;
  c.rois = nullptr;
  c.total_dev = nullptr;
  c.total_host = n;
  c.slots = n;
  c.out_w = new_w;
  c.out_h = new_h;
  c.canvas = target;
  c.canvas_x0 = pad_w;
  c.canvas_y0 = pad_h;
  c.tiles_x = (new_w + kTeamConsumers - 1) / kTeamConsumers;
  // Tiles are dealt round-robin (no scratch memory for tickets here): a small batch takes shorter
  // tiles so that every team still gets a dozen of them
  int sms = 148;
  {
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  c.tile_rows = kTileRows;
  while (c.tile_rows > 7 && (long long)n * c.tiles_x * ((target + c.tile_rows - 1) / c.tile_rows) < 12LL * kMaxTeams * sms)
    c.tile_rows /= 2;
  c.tiles_y = (new_h + c.tile_rows - 1) / c.tile_rows;
  c.fill_top = (pad_h + c.tile_rows - 1) / c.tile_rows;
  c.fill_bottom = (target - pad_h - new_h + c.tile_rows - 1) / c.tile_rows;
  c.fill_left = pad_w > 0;
  c.fill_right = target - pad_w - new_w > 0;
  c.tiles_per_image = c.tiles_x * c.tiles_y + c.fill_top + c.fill_bottom + (c.fill_left + c.fill_right) * c.tiles_y;
  c.sc_x = axis_scale(new_w, src_w);
  c.sc_y = axis_scale(new_h, src_h);
  c.xtab_bytes = (new_w * 8 + 15) & ~15;
  c.fill_rgb = pad_rgb;
  c.divisor = divisor;
  for (int i = 0; i < 3; ++i) c.mean[i] = 0.f, c.stdv[i] = 1.f;
  c.out = out;
  c.tickets = nullptr;
  // a tile's staged segment spans the source columns of 224 output columns
  const long long seg = ((long long)kTeamConsumers * src_w + new_w - 1) / new_w + 3;
  return launch_tiles(c, (int)(seg < src_w ? seg : src_w), stream);
}

}  // namespace ia
