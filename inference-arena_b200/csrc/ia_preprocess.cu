// ia_preprocess.cu — K1: fused letterbox + normalise + HWC->CHW (sm_100a).
//
// Replaces YOLOPreprocessor.preprocess (reference src/shared/processing/
// yolo_preprocess.py:128-174) and letterbox (transforms.py:118-180) for a batch of
// equally-sized uint8 HWC frames.  One CTA produces a group of output rows of one image:
//   1. warp 0 works out which source rows those output rows need (OpenCV's yofs/beta
//      table), drops duplicates and zero-weight rows, and stages each distinct row into
//      shared memory with a 1-D bulk async copy (cp.async.bulk -> TMA engine, completion on
//      an mbarrier); the <= 15 unaligned bytes at either end of a row are copied by the lane;
//   2. meanwhile the other warps build the u8 -> float LUT (exact float32 division, same
//      as numpy's `astype(float32) / 255.0`; 8 copies across banks) and the x-coefficient
//      table;
//   3. every thread then produces 4 consecutive output pixels: 8-bit fixed-point bilinear
//      (bit-identical to cv2.resize INTER_LINEAR), LUT, three coalesced 128-bit streaming
//      stores — one per colour plane.
// The host evaluates the same coefficient tables to size the staging buffer exactly (a 3x
// downscale of 1080p needs ONE source row per output row), which sets the occupancy.
// HBM traffic = source rows actually needed + the output, each once.
#include <math.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "ia_common.cuh"

namespace ia {

constexpr int kPreThreads = 256;
constexpr int kPreMaxRows = 8;
constexpr int kPreLutRep = 8;

struct PreParams {
  const uint8_t* frames;
  void* out;
  int n, src_h, src_w, target;
  int new_h, new_w, pad_h, pad_w;
  int rows_per_cta, groups;
  int slot_bytes;  // shared-memory bytes reserved per staged source row (16-byte multiple)
  int slots_cap;   // staged rows a CTA has room for (computed exactly by the host)
  float divisor;
  double sc_x, sc_y;
  uint32_t pad_rgb;  // r | g << 8 | b << 16
};

struct RowPlan {
  int image_row;   // 0: padding row
  int off0, off1;  // shared-memory byte offsets of the two staged source rows
  int b0, b1;      // vertical coefficients
};

// Host twin of axis_coef (ia_common.cuh): plain IEEE double/float operations, one rounding
// per step (volatile keeps the host compiler from contracting or reassociating).
static AxisCoef axis_coef_host(int i, double sc, int src, bool zero_frac_at_border) {
  volatile double t = (double)i + 0.5;
  t = t * sc;
  t = t - 0.5;
  volatile float f = (float)t;
  int s0 = (int)floorf(f);
  f = f - (float)s0;
  if (zero_frac_at_border) {
    if (s0 < 0) s0 = 0, f = 0.f;
    if (s0 >= src - 1) s0 = src - 1, f = 0.f;
  }
  volatile float w1 = f * (float)kCoefOne;
  volatile float w0 = 1.f - f;
  w0 = w0 * (float)kCoefOne;
  AxisCoef c;
  c.s0 = s0;
  c.w1 = (int)lrintf(w1);
  c.w0 = (int)lrintf(w0);
  return c;
}

// Source rows (r0, r1; -1 = not needed) and weights of letterboxed output row y.
__host__ __device__ inline bool row_sources(int mode, int dy, int new_h, int src_h, double sc_y,
                                            int& r0, int& r1, int& b0, int& b1) {
  r0 = r1 = -1;
  b0 = kCoefOne;
  b1 = 0;
  if (dy < 0 || dy >= new_h) return false;
  if (mode == kResizeCopy) {
    r0 = dy;
  } else if (mode == kResizeBox2) {
    r0 = 2 * dy;
    r1 = 2 * dy + 1;
    b0 = b1 = kCoefOne / 2;
  } else {
#ifdef __CUDA_ARCH__
    const AxisCoef c = axis_coef(dy, sc_y, src_h, false);
#else
    const AxisCoef c = axis_coef_host(dy, sc_y, src_h, false);
#endif
    b0 = c.w0;
    b1 = c.w1;
    r0 = c.s0 < 0 ? 0 : (c.s0 > src_h - 1 ? src_h - 1 : c.s0);
    if (c.w1 != 0) r1 = c.s0 + 1 < 0 ? 0 : (c.s0 + 1 > src_h - 1 ? src_h - 1 : c.s0 + 1);
  }
  return true;
}

// Ragged batches (frames of different sizes): images of one shape are launched together, and
// their base pointers / output slots travel in the kernel parameters (no device scratch).
constexpr int kRaggedMax = 256;
template <bool kRagged>
struct PreRagged {};
template <>
struct PreRagged<true> {
  const uint8_t* frame[kRaggedMax];
  int out_index[kRaggedMax];
};

// The work of one CTA: `rows_per_cta` output rows of one image, starting at y_begin.  The
// shared arrays belong to the calling kernel (one set, whatever modes it instantiates).
template <bool kOutU8, int kMode>
__device__ __forceinline__ void letterbox_body(const PreParams& p, const uint8_t* frame, int out_img, int y_begin,
                                               uint8_t* smem, float* lut, RowPlan* plan, int* seg_row,
                                               uint64_t& bar) {
  const int tid = threadIdx.x, lane = tid & 31;
  const int rows = min(p.rows_per_cta, p.target - y_begin);
  const uint32_t row_bytes = (uint32_t)p.src_w * 3u;
  int2* xtab = reinterpret_cast<int2*>(smem + (size_t)p.slots_cap * p.slot_bytes);

  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  // ---- per-row plan: which source rows, which vertical weights -------------------
  if (tid < rows) {
    int r0, r1, b0, b1;
    RowPlan rp;
    rp.image_row = row_sources(kMode, y_begin + tid - p.pad_h, p.new_h, p.src_h, p.sc_y, r0, r1, b0, b1) ? 1 : 0;
    rp.off0 = rp.off1 = 0;
    rp.b0 = b0;
    rp.b1 = b1;
    plan[tid] = rp;
    seg_row[2 * tid] = r0;
    seg_row[2 * tid + 1] = r1;
  }
  __syncthreads();

  // ---- stage the distinct source rows: bulk async copy of the 16-byte aligned body -------
  if (tid < 32) {
    const int j = lane >> 1, tap = lane & 1;
    const int row = (j < rows) ? seg_row[lane] : -1;
    const unsigned same = __match_any_sync(0xffffffffu, row);
    const int leader = __ffs(same) - 1;
    const bool copy = (leader == lane) && row >= 0;
    const unsigned leaders = __ballot_sync(0xffffffffu, copy);
    const int slot_raw = __popc(leaders & ((1u << leader) - 1u));
    // The host sized the staging buffer from the same tables (IEEE twins of this code), so a row
    // beyond it is not expected; if the two ever disagreed the row would share the last slot —
    // wrong pixels in this row group, never an access outside the buffer and never a trap that
    // would take the caller's CUDA context down.
    const int slot = min(slot_raw, p.slots_cap - 1);
    const uint8_t* g = frame + (size_t)max(row, 0) * row_bytes;
    const uint32_t mis = (uint32_t)((uintptr_t)g & 15);
    const int off = slot * p.slot_bytes + (int)mis;
    uint32_t head = 0, body = 0, tail = 0;
    if (copy) {
      head = min(row_bytes, (16u - mis) & 15u);
      body = (row_bytes - head) & ~15u;
      tail = row_bytes - head - body;
    }
    uint32_t total = body;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    if (lane == 0) mbar_expect_tx(&bar, total);
    __syncwarp();
    if (copy) {
      uint8_t* s = smem + off;
      if (body) bulk_g2s(s + head, g + head, body, &bar);
      for (uint32_t i = 0; i < head; ++i) s[i] = g[i];
      for (uint32_t i = 0; i < tail; ++i) s[head + body + i] = g[head + body + i];
    }
    if (j < rows && row >= 0) {
      if (tap == 0) plan[j].off0 = off;
      else plan[j].off1 = off;
    }
  }

  // ---- tables ---------------------------------------------------------------------
  if (!kOutU8) {  // one exact division per pixel value, then replicate it across the banks
    static_assert(kPreThreads == 256 && kPreLutRep == 8, "one LUT value per thread, two float4 stores");
    const float q = __fdiv_rn((float)tid, p.divisor);
    float4* dst = reinterpret_cast<float4*>(lut + tid * kPreLutRep);
    dst[0] = make_float4(q, q, q, q);
    dst[1] = make_float4(q, q, q, q);
  }
  if (kMode == kResizeLinear || kMode == kResizeGather) {
    for (int dx = tid; dx < p.new_w; dx += kPreThreads) {
      const AxisCoef c = axis_coef(dx, p.sc_x, p.src_w, true);
      xtab[dx] = make_int2(c.s0 * 3, c.w0 | (c.w1 << 16));
    }
  }
  __syncthreads();
  mbar_wait(&bar, 0);

  // ---- compute: 4 pixels per item --------------------------------------------------
  const int T = p.target;
  const int xg = (T + 3) >> 2;
  const uint32_t pr = p.pad_rgb & 0xff, pg = (p.pad_rgb >> 8) & 0xff, pb = (p.pad_rgb >> 16) & 0xff;
  const bool vec_ok = (T & 3) == 0;
  const float* l = lut + (kOutU8 ? 0 : (lane & (kPreLutRep - 1)));

  // (row j, pixel group gx) of item `tid`, then advanced by kPreThreads items per iteration
  int j = tid / xg, gx = tid - j * xg;
  const int dj = kPreThreads / xg, dg = kPreThreads - dj * xg;
  for (int it = tid; it < rows * xg; it += kPreThreads, j += dj, gx += dg) {
    if (gx >= xg) gx -= xg, ++j;
    const int x = gx << 2;
    const int y = y_begin + j;
    const RowPlan rp = plan[j];
    const uint8_t* p0 = smem + rp.off0;
    const uint8_t* p1 = smem + rp.off1;
    (void)p1;  // (only the modes with a second tap row read it)
    uint32_t px[4][3];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int dx = x + k - p.pad_w;
      if (!rp.image_row || dx < 0 || dx >= p.new_w) {
        px[k][0] = pr;
        px[k][1] = pg;
        px[k][2] = pb;
      } else if (kMode == kResizeCopy) {
        const uint8_t* s = p0 + dx * 3;
        px[k][0] = s[0];
        px[k][1] = s[1];
        px[k][2] = s[2];
      } else if (kMode == kResizeBox2) {
        uint32_t lo0, hi0, lo1, hi1;  // two adjacent pixels of each of the two rows
        load6(reinterpret_cast<const uint32_t*>(smem), rp.off0 + dx * 6, lo0, hi0);
        load6(reinterpret_cast<const uint32_t*>(smem), rp.off1 + dx * 6, lo1, hi1);
#pragma unroll
        for (int c = 0; c < 3; ++c)  // IDP.2A with weights (1, 1): p[c] + p[c+3]
          px[k][c] = (__dp2a_lo(0x00010001u, tap_pair(lo0, hi0, c), 2u) +
                      __dp2a_lo(0x00010001u, tap_pair(lo1, hi1, c), 0u)) >> 2;
      } else if (kMode == kResizeGather) {
        const uint8_t* s = p0 + xtab[dx].x;
        px[k][0] = s[0];
        px[k][1] = s[1];
        px[k][2] = s[2];
      } else {
        // six bytes per row through three aligned word loads + a funnel shift, PRMT pairs the
        // two taps of a channel, one IDP.2A does p0*w0 + p1*w1 (branch-free: a zero weight
        // multiplies a byte that is always inside the staging buffer — every slot has 32
        // spare bytes — by 0)
        const int2 e = xtab[dx];
        uint32_t lo0, hi0, lo1, hi1;
        load6(reinterpret_cast<const uint32_t*>(smem), rp.off0 + e.x, lo0, hi0);
        load6(reinterpret_cast<const uint32_t*>(smem), rp.off1 + e.x, lo1, hi1);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const int h0 = (int)__dp2a_lo((uint32_t)e.y, tap_pair(lo0, hi0, c), 0u);
          const int h1 = (int)__dp2a_lo((uint32_t)e.y, tap_pair(lo1, hi1, c), 0u);
          px[k][c] = (uint32_t)vresize(h0, h1, rp.b0, rp.b1);
        }
      }
    }
    if (kOutU8) {
      uint8_t* o = reinterpret_cast<uint8_t*>(p.out) + (((size_t)out_img * T + y) * T + x) * 3;
      if (vec_ok) {
        uint32_t* o32 = reinterpret_cast<uint32_t*>(o);
        o32[0] = px[0][0] | px[0][1] << 8 | px[0][2] << 16 | px[1][0] << 24;
        o32[1] = px[1][1] | px[1][2] << 8 | px[2][0] << 16 | px[2][1] << 24;
        o32[2] = px[2][2] | px[3][0] << 8 | px[3][1] << 16 | px[3][2] << 24;
      } else {
        for (int k = 0; k < 4 && x + k < T; ++k)
          for (int c = 0; c < 3; ++c) o[k * 3 + c] = (uint8_t)px[k][c];
      }
    } else {
      float* o = reinterpret_cast<float*>(p.out) + (((size_t)out_img * 3) * T + y) * T + x;
      const size_t plane = (size_t)T * T;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        if (vec_ok) {
          st_stream_f4(o + c * plane, l[px[0][c] * kPreLutRep], l[px[1][c] * kPreLutRep],
                       l[px[2][c] * kPreLutRep], l[px[3][c] * kPreLutRep]);
        } else {
          for (int k = 0; k < 4 && x + k < T; ++k) st_stream_f1(o + c * plane + k, l[px[k][c] * kPreLutRep]);
        }
      }
    }
  }
}

template <bool kOutU8, int kMode, bool kRagged>
__global__ void __launch_bounds__(kPreThreads) letterbox_kernel(const PreParams p,
                                                                const __grid_constant__ PreRagged<kRagged> rg) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(16) float lut[kOutU8 ? 4 : 256 * kPreLutRep];
  __shared__ RowPlan plan[kPreMaxRows];
  __shared__ int seg_row[2 * kPreMaxRows];  // source row per (output row, tap), -1 = none
  __shared__ __align__(8) uint64_t bar;
  const int img = blockIdx.x / p.groups;
  const int y_begin = (blockIdx.x % p.groups) * p.rows_per_cta;
  const uint8_t* frame;
  int out_img;
  if constexpr (kRagged) {
    frame = rg.frame[img];
    out_img = rg.out_index[img];
  } else {
    frame = p.frames + (size_t)img * p.src_h * p.src_w * 3u;
    out_img = img;
  }
  letterbox_body<kOutU8, kMode>(p, frame, out_img, y_begin, smem, lut, plan, seg_row, bar);
}

// A ragged batch with several frame shapes in ONE launch: every shape keeps its own geometry
// and resize mode (a CTA looks its shape up from its block index), so a batch of real camera
// or dataset frames costs one launch instead of one per shape.
constexpr int kMixedMaxShapes = 32;
struct MixedParams {
  int num_shapes;
  int first_cta[kMixedMaxShapes + 1];  // CTAs of shape s: [first_cta[s], first_cta[s + 1])
  int first_img[kMixedMaxShapes];      // its frames: frame[first_img[s] ...]
  int mode[kMixedMaxShapes];
  PreParams p[kMixedMaxShapes];
  const uint8_t* frame[kRaggedMax];
  int out_index[kRaggedMax];
};

__global__ void __launch_bounds__(kPreThreads) letterbox_mixed_kernel(const __grid_constant__ MixedParams mp) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(16) float lut[256 * kPreLutRep];
  __shared__ RowPlan plan[kPreMaxRows];
  __shared__ int seg_row[2 * kPreMaxRows];
  __shared__ __align__(8) uint64_t bar;
  int s = 0;
  while (s + 1 < mp.num_shapes && (int)blockIdx.x >= mp.first_cta[s + 1]) ++s;
  const PreParams& p = mp.p[s];
  const int local = (int)blockIdx.x - mp.first_cta[s];
  const int img = mp.first_img[s] + local / p.groups;
  const int y_begin = (local % p.groups) * p.rows_per_cta;
  const uint8_t* frame = mp.frame[img];
  const int out_img = mp.out_index[img];
  switch (mp.mode[s]) {  // uniform over the CTA
    case kResizeCopy: letterbox_body<false, kResizeCopy>(p, frame, out_img, y_begin, smem, lut, plan, seg_row, bar); break;
    case kResizeBox2: letterbox_body<false, kResizeBox2>(p, frame, out_img, y_begin, smem, lut, plan, seg_row, bar); break;
    case kResizeGather: letterbox_body<false, kResizeGather>(p, frame, out_img, y_begin, smem, lut, plan, seg_row, bar); break;
    default: letterbox_body<false, kResizeLinear>(p, frame, out_img, y_begin, smem, lut, plan, seg_row, bar); break;
  }
}

// Distinct source rows the busiest CTA has to stage when each CTA owns `rows` output rows.
static int max_distinct_rows(int mode, int rows, const ia_letterbox_geom& g, double sc_y) {
  int worst = 1;
  for (int y0 = 0; y0 < g.target; y0 += rows) {
    int seen[2 * kPreMaxRows];
    int n = 0;
    for (int j = 0; j < rows && y0 + j < g.target; ++j) {
      int r[2], b0, b1;
      row_sources(mode, y0 + j - g.pad_h, g.new_h, g.src_h, sc_y, r[0], r[1], b0, b1);
      for (int t = 0; t < 2; ++t) {
        if (r[t] < 0) continue;
        bool dup = false;
        for (int k = 0; k < n; ++k) dup |= seen[k] == r[t];
        if (!dup) seen[n++] = r[t];
      }
    }
    if (n > worst) worst = n;
  }
  return worst;
}

// Geometry, mode, rows per CTA and staging size for one frame shape (cached: the table walk
// costs tens of microseconds on the host).
static int plan_letterbox(int src_h, int src_w, int target, PreParams& p, int& mode, size_t& smem,
                          bool linear_in_tile_kernel = false) {
  ia_letterbox_geom g;
  int rc = ia_letterbox_geometry(src_h, src_w, target, &g);
  if (rc != IA_OK) return rc;
  p.src_h = src_h;
  p.src_w = src_w;
  p.target = target;
  p.new_h = g.new_h;
  p.new_w = g.new_w;
  p.pad_h = g.pad_h;
  p.pad_w = g.pad_w;
  p.sc_x = axis_scale(g.new_w, src_w);
  p.sc_y = axis_scale(g.new_h, src_h);
  p.slot_bytes = (((src_w * 3 + 15) & ~15) + 32);  // row + misalignment + branch-free over-read
  mode = resize_mode(src_h, src_w, g.new_h, g.new_w);
  const size_t xtab_bytes = mode == kResizeLinear ? (size_t)g.new_w * sizeof(int2) : 0;
  // rows per CTA: the largest group whose staging buffer leaves room for >= 6 CTAs per SM
  int rows = kPreMaxRows, slots = 0;
  {
    static std::mutex mu;
    static std::map<std::tuple<int, int, int>, std::tuple<int, int, int>> cache;
    std::lock_guard<std::mutex> lock(mu);
    const auto key = std::make_tuple(src_h, src_w, target);
    auto it = cache.find(key);
    if (it == cache.end()) {
      if (mode == kResizeLinear) {  // all second coefficients zero on both axes -> pure gather
        bool gather = true;
        for (int x = 0; x < g.new_w && gather; ++x) gather = axis_coef_host(x, p.sc_x, src_w, true).w1 == 0;
        for (int y = 0; y < g.new_h && gather; ++y) gather = axis_coef_host(y, p.sc_y, src_h, false).w1 == 0;
        if (gather) mode = kResizeGather;
      }
      for (;; rows >>= 1) {
        slots = max_distinct_rows(mode, rows, g, p.sc_y);
        if (rows == 1 || (size_t)slots * p.slot_bytes + xtab_bytes <= 28 * 1024) break;
      }
      if (cache.size() > 256) cache.clear();
      cache[key] = std::make_tuple(rows, slots, mode);
    } else {
      rows = std::get<0>(it->second);
      slots = std::get<1>(it->second);
      mode = std::get<2>(it->second);
    }
  }
  smem = (size_t)slots * p.slot_bytes + xtab_bytes;
  if (linear_in_tile_kernel && mode == kResizeLinear) return IA_OK;  // geometry and mode only
  IA_RETURN_IF(smem > 200 * 1024, IA_ERR_UNSUPPORTED);
  IA_RETURN_IF((size_t)slots * src_w * 3 >= (1u << 20), IA_ERR_UNSUPPORTED);  // mbarrier tx range
  p.rows_per_cta = rows;
  p.slots_cap = slots;
  p.groups = (target + rows - 1) / rows;
  return IA_OK;
}

template <bool kOutU8, bool kRagged>
static int launch_planned(PreParams& p, int mode, size_t smem, int n, const PreRagged<kRagged>& rg,
                          cudaStream_t stream) {
  p.n = n;
  const long long blocks = (long long)p.groups * n;
  IA_RETURN_IF(blocks > 0x7fffffffLL, IA_ERR_UNSUPPORTED);
  void (*kernel)(PreParams, PreRagged<kRagged>) = nullptr;
  switch (mode) {
    case kResizeCopy: kernel = letterbox_kernel<kOutU8, kResizeCopy, kRagged>; break;
    case kResizeBox2: kernel = letterbox_kernel<kOutU8, kResizeBox2, kRagged>; break;
    case kResizeGather: kernel = letterbox_kernel<kOutU8, kResizeGather, kRagged>; break;
    default: kernel = letterbox_kernel<kOutU8, kResizeLinear, kRagged>; break;
  }
  cudaError_t e = allow_max_dynamic_smem(reinterpret_cast<const void*>(kernel));
  if (e != cudaSuccess) return (int)e;
  kernel<<<(unsigned)blocks, kPreThreads, smem, stream>>>(p, rg);
  return launch_status();
}

template <bool kOutU8>
static int launch_letterbox(const uint8_t* frames, int n, int src_h, int src_w, int target,
                            float divisor, uint32_t pad_rgb, void* out, cudaStream_t stream) {
  IA_RETURN_IF(n < 0 || src_h < 1 || src_w < 1 || target < 1, IA_ERR_INVALID_ARGUMENT);
  if (n == 0) return IA_OK;
  IA_RETURN_IF(!frames || !out, IA_ERR_INVALID_ARGUMENT);
  PreParams p;
  int mode;
  size_t smem;
  int rc = plan_letterbox(src_h, src_w, target, p, mode, smem, !kOutU8);
  if (rc != IA_OK) return rc;
  if (!kOutU8 && mode == kResizeLinear)  // the separable tile kernel (ia_resize_tiles.cuh)
    return launch_letterbox_tiles(frames, n, src_h, src_w, target, p.new_h, p.new_w, p.pad_h, p.pad_w, divisor,
                                  pad_rgb, static_cast<float*>(out), stream);
  p.frames = frames;
  p.out = out;
  p.divisor = divisor;
  p.pad_rgb = pad_rgb;
  return launch_planned<kOutU8, false>(p, mode, smem, n, PreRagged<false>{}, stream);
}

// Frames of different sizes: one launch per (shape, chunk of kRaggedMax images).
static int launch_letterbox_ragged(const ia_frame_desc* frames, int n, int target, float divisor,
                                   uint32_t pad_rgb, float* out, cudaStream_t stream) {
  IA_RETURN_IF(n < 0 || target < 1, IA_ERR_INVALID_ARGUMENT);
  if (n == 0) return IA_OK;
  IA_RETURN_IF(!frames || !out, IA_ERR_INVALID_ARGUMENT);
  std::map<std::pair<int, int>, std::vector<int>> by_shape;
  for (int i = 0; i < n; ++i) {
    IA_RETURN_IF(!frames[i].data_dev || frames[i].h < 1 || frames[i].w < 1, IA_ERR_INVALID_ARGUMENT);
    by_shape[std::make_pair(frames[i].h, frames[i].w)].push_back(i);
  }
  if (by_shape.size() > 1 && by_shape.size() <= (size_t)kMixedMaxShapes && n <= kRaggedMax) {
    // several shapes, one launch: per-shape geometry and frame pointers in the kernel parameters
    MixedParams mp;
    mp.num_shapes = 0;
    size_t smem_max = 0;
    long long ctas = 0;
    int images = 0;
    for (const auto& kv : by_shape) {
      const int s = mp.num_shapes++;
      PreParams& p = mp.p[s];
      size_t smem;
      int rc = plan_letterbox(kv.first.first, kv.first.second, target, p, mp.mode[s], smem);
      if (rc != IA_OK) return rc;
      p.frames = nullptr;
      p.out = out;
      p.divisor = divisor;
      p.pad_rgb = pad_rgb;
      p.n = (int)kv.second.size();
      smem_max = smem > smem_max ? smem : smem_max;
      mp.first_cta[s] = (int)ctas;
      mp.first_img[s] = images;
      for (int i : kv.second) {
        mp.frame[images] = frames[i].data_dev;
        mp.out_index[images] = i;
        ++images;
      }
      ctas += (long long)p.groups * p.n;
      IA_RETURN_IF(ctas > 0x7fffffffLL, IA_ERR_UNSUPPORTED);
    }
    mp.first_cta[mp.num_shapes] = (int)ctas;
    cudaError_t e = allow_max_dynamic_smem(reinterpret_cast<const void*>(letterbox_mixed_kernel));
    if (e != cudaSuccess) return (int)e;
    letterbox_mixed_kernel<<<(unsigned)ctas, kPreThreads, smem_max, stream>>>(mp);
    return launch_status();
  }
  for (const auto& kv : by_shape) {
    PreParams p;
    int mode;
    size_t smem;
    int rc = plan_letterbox(kv.first.first, kv.first.second, target, p, mode, smem);
    if (rc != IA_OK) return rc;
    p.frames = nullptr;
    p.out = out;
    p.divisor = divisor;
    p.pad_rgb = pad_rgb;
    const std::vector<int>& idx = kv.second;
    for (size_t lo = 0; lo < idx.size(); lo += kRaggedMax) {
      const int m = (int)std::min(idx.size() - lo, (size_t)kRaggedMax);
      PreRagged<true> rg;
      for (int k = 0; k < m; ++k) {
        rg.frame[k] = frames[idx[lo + k]].data_dev;
        rg.out_index[k] = idx[lo + k];
      }
      rc = launch_planned<false, true>(p, mode, smem, m, rg, stream);
      if (rc != IA_OK) return rc;
    }
  }
  return IA_OK;
}

}  // namespace ia

extern "C" int ia_letterbox_geometry(int src_h, int src_w, int target, ia_letterbox_geom* out) {
  if (!out || src_h < 1 || src_w < 1 || target < 1) return IA_ERR_INVALID_ARGUMENT;
  // transforms.py:153-157,174-175 — Python float (binary64) arithmetic, int() truncation.
  const double sh = (double)target / (double)src_h;
  const double sw = (double)target / (double)src_w;
  const double scale = sh < sw ? sh : sw;
  out->src_h = src_h;
  out->src_w = src_w;
  out->target = target;
  out->scale = scale;
  out->new_w = (int)((double)src_w * scale);
  out->new_h = (int)((double)src_h * scale);
  // cv2.resize rejects an empty destination; mirror it as an error.
  if (out->new_w < 1 || out->new_h < 1) return IA_ERR_UNSUPPORTED;
  out->pad_w = (target - out->new_w) / 2;
  out->pad_h = (target - out->new_h) / 2;
  return IA_OK;
}

extern "C" int ia_yolo_preprocess(const uint8_t* frames_dev, int n, int src_h, int src_w,
                                  int target, float divisor, float* out_dev, void* stream) {
  return ia::launch_letterbox<false>(frames_dev, n, src_h, src_w, target, divisor,
                                     114u | 114u << 8 | 114u << 16, out_dev, (cudaStream_t)stream);
}

extern "C" int ia_letterbox_u8(const uint8_t* frames_dev, int n, int src_h, int src_w, int target,
                               const uint8_t pad_rgb[3], uint8_t* out_dev, void* stream) {
  if (!pad_rgb) return IA_ERR_INVALID_ARGUMENT;
  const uint32_t rgb = (uint32_t)pad_rgb[0] | (uint32_t)pad_rgb[1] << 8 | (uint32_t)pad_rgb[2] << 16;
  return ia::launch_letterbox<true>(frames_dev, n, src_h, src_w, target, 255.0f, rgb, out_dev,
                                    (cudaStream_t)stream);
}

extern "C" int ia_yolo_preprocess_ragged(const ia_frame_desc* frames_host, int n, int target,
                                         float divisor, float* out_dev, void* stream) {
  return ia::launch_letterbox_ragged(frames_host, n, target, divisor, 114u | 114u << 8 | 114u << 16,
                                     out_dev, (cudaStream_t)stream);
}
